// Micro-benchmarks for the FP64 paths the iLQR kernels lean on (B200, sm_100a): throughput (full SM, 8 independent
// chains per thread) and latency (one warp, one dependent chain) of DFMA, DMUL, SHFL.64, MUFU.RCP64H, DMMA m8n8k4, LDS.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_ubench fp64_ubench.cu && ./fp64_ubench
#include <cstdio>
#include <cuda_runtime.h>
constexpr int ITERS = 4096;
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
template <int OP, int ILP>
__global__ void k(double* out, long long* cyc, double seed) {
    double v[ILP], w[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) { v[i] = seed + threadIdx.x * 1e-3 + i; w[i] = 1.0; }
    __shared__ double sm[1024];
    sm[threadIdx.x & 1023] = seed;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) {
            if (OP == 0) v[i] = fma(v[i], 1.0000001, 1e-9);
            if (OP == 1) v[i] = v[i] * 1.0000001;
            if (OP == 2) v[i] = __shfl_sync(0xffffffffu, v[i], (threadIdx.x + 1) & 31);
            if (OP == 3) { double y; asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(v[i])); v[i] = y; }
            if (OP == 4) dmma(v[i], w[i], v[i], w[i]);
            if (OP == 5) v[i] = sm[(__double2loint(v[i]) + threadIdx.x) & 1023];
            if (OP == 6) { int x = __shfl_sync(0xffffffffu, __double2loint(v[i]), (threadIdx.x + 1) & 31); v[i] = __hiloint2double(__double2hiint(v[i]), x); }
        }
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += v[i] + w[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int OP>
void run(const char* name, double* out, long long* cyc) {
    long long h;
    k<OP, 1><<<1, 32>>>(out, cyc, 1.0);
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    double lat = (double)h / ITERS;
    // throughput: 148 blocks x 1024 threads (8 warps per scheduler), 8 independent chains each
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<OP, 8><<<148, 1024>>>(out, cyc, 1.0);
    cudaEventRecord(e0);
    k<OP, 8><<<148, 1024>>>(out, cyc, 1.0);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    double warp_insts_per_sm = 32.0 * 8 * ITERS;  // 32 warps x ILP x ITERS
    printf("%-14s latency %6.1f cyc   throughput %6.3f warp-inst/cyc/SM (%5.2f cyc per warp-inst per scheduler)  [%s]\n", name, lat, warp_insts_per_sm / h, 4.0 * h / warp_insts_per_sm,
           cudaGetErrorString(cudaGetLastError()));
}
int main() {
    double* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 8); cudaMalloc(&cyc, 8);
    run<0>("DFMA", out, cyc);
    run<1>("DMUL", out, cyc);
    run<2>("SHFL.64", out, cyc);
    run<6>("SHFL.32", out, cyc);
    run<3>("MUFU.RCP64H", out, cyc);
    run<4>("DMMA.884", out, cyc);
    run<5>("LDS.64", out, cyc);
    return 0;
}
