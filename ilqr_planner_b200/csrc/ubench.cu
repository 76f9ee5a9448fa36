// ubench.cu — FP64 peak rates of the device measured in place (ilqrb_fp64_peaks): the denominators of the FP64 rooflines.
// MEASURED_PEAKS.json carries HBM and bf16 figures only; the solve path is FP64 (DFMA + DMMA m8n8k4, the warp-level FP64 tensor
// shape — tcgen05 has no f64 kind), so its peaks are measured here with CUDA events: every SM full (one 1024-thread block per
// SM and pass, 8 independent chains per thread), best of several passes.
#include <cuda_runtime.h>

#include "handle.hpp"
#include "wmat.cuh"

namespace ilqrb {
namespace {
constexpr int UB_ITERS = 2048, UB_ILP = 8;

template <int OP>
__global__ void __launch_bounds__(1024) k_fp64_peak(double* out, double seed) {
    double v[UB_ILP], w[UB_ILP];
#pragma unroll
    for (int i = 0; i < UB_ILP; i++) {
        v[i] = seed + threadIdx.x * 1e-3 + i;
        w[i] = 1.0;
    }
    for (int it = 0; it < UB_ITERS; it++) {
#pragma unroll
        for (int i = 0; i < UB_ILP; i++) {
            if (OP == 0) v[i] = fma(v[i], 1.0000001, 1e-9);
            else dmma884(v[i], w[i], v[i], w[i]);
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < UB_ILP; i++) s += v[i] + w[i];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP>
int peak(int sms, double* out, cudaStream_t st, double* tflops) {
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    const int waves = 4;
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
        CU(cudaEventRecord(e0, st));
        k_fp64_peak<OP><<<sms * waves, 1024, 0, st>>>(out, 1.0);
        CU(cudaEventRecord(e1, st));
        CU(cudaEventSynchronize(e1));
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    CU(cudaGetLastError());
    // DFMA: 2 flop per thread instruction; DMMA m8n8k4: 2 * 8 * 8 * 4 = 512 flop per warp instruction
    const double per_thread = OP == 0 ? 2.0 : 512.0 / 32.0;
    *tflops = (double)sms * waves * 1024.0 * UB_ILP * UB_ITERS * per_thread / (best * 1e-3) / 1e12;
    return 0;
}
}  // namespace
}  // namespace ilqrb

extern "C" int ilqrb_fp64_peaks(int device, double* dfma_tflops, double* dmma_tflops) {
    using namespace ilqrb;
    CU(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    const int sms = prop.multiProcessorCount;
    double* out = nullptr;
    CU(cudaMalloc(&out, (size_t)sms * 4 * 1024 * sizeof(double)));
    cudaStream_t st;
    CU(cudaStreamCreate(&st));
    int rc = 0;
    double a = 0, b = 0;
    if (!(rc = peak<0>(sms, out, st, &a))) rc = peak<1>(sms, out, st, &b);
    cudaStreamDestroy(st);
    cudaFree(out);
    if (rc) return rc;
    if (dfma_tflops) *dfma_tflops = a;
    if (dmma_tflops) *dmma_tflops = b;
    return 0;
}
