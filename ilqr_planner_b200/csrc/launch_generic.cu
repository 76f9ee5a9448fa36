// launch_generic.cu — launchers of the generic Recursive / AL kernels (recursive.cuh): initial rollout, keypoint derivatives,
// all-in-shared-memory backward pass, the two line searches, accept, f(X), and the kernel-level test hooks.
#include <cstdlib>

#include "handle.hpp"
#include "recursive.cuh"

namespace ilqrb {

bool generic_forced() {
    static const bool forced = getenv("ILQRB_GENERIC") != nullptr;
    return forced;
}

namespace {

// ---- dispatch over the compiled size classes -------------------------------------------------------------
int size_class(int n, int m) {  // thread kernels <N, M>
    if (n <= 8 && m <= 8) return 0;
    if (n <= 16 && m <= 8) return 1;
    if (n <= 16 && m <= 16) return 2;
    return 3;
}
#define DISPATCH_NM(cls, CALL)          \
    switch (cls) {                      \
        case 0: { CALL(8, 8); } break;  \
        case 1: { CALL(16, 8); } break; \
        case 2: { CALL(16, 16); } break;\
        default: { CALL(32, 16); } break;\
    }

int tile_class(int n, int m) {  // backward kernel <NXT, NUT>: 8*NXT >= n+1, 8*NUT >= m
    if (n + 1 <= 8 && m <= 8) return 0;
    if (n + 1 <= 16 && m <= 8) return 1;
    if (n + 1 <= 16 && m <= 16) return 2;
    return 3;
}

template <int NXT, int NUT, bool TIME, bool AL, int WARPS, int MINB>
int launch_backward_t(ilqrb_handle h, double penalty_frozen) {
    using S = BwdSmem<NXT, NUT>;
    const size_t smem = ((TIME ? WARPS : 1) * S::DYN + (size_t)WARPS * S::PER_WARP) * sizeof(double);
    auto kern = k_backward<NXT, NUT, TIME, AL, WARPS, MINB>;
    static DynSmem attr;
    int rc = attr.ensure(kern, smem, h->device);
    if (rc) return rc;
    const int B = h->hp.B;
    kern<<<(B + WARPS - 1) / WARPS, 32 * WARPS, smem, h->stream>>>(h->dp, h->W, penalty_frozen);
    return 0;
}
template <bool TIME, bool AL>
int launch_backward_ta(ilqrb_handle h, double pf) {
    switch (tile_class(h->hp.n, h->hp.m)) {
        case 0: return launch_backward_t<1, 1, TIME, AL, 4, 7>(h, pf);
        case 1: return launch_backward_t<2, 1, TIME, AL, 4, 3>(h, pf);
        case 2: return launch_backward_t<2, 2, TIME, AL, 4, 2>(h, pf);
        default: return launch_backward_t<4, 2, TIME, AL, 2, 1>(h, pf);
    }
}

// affine line search (k_linesearch_affine): linear dynamics
template <int N, int M>
int launch_linesearch_affine_t(ilqrb_handle h, int line_search) {
    const DevProblem& P = h->hp;
    const size_t smem = 4 * (size_t)ls_affine_smem_doubles(P.H, h->W.ns, h->W.ms) * sizeof(double);
    auto kern = k_linesearch_affine<N, M>;
    static DynSmem attr;
    int rc = attr.ensure(kern, smem, h->device);
    if (rc) return rc;
    kern<<<(P.B + 3) / 4, 128, smem, h->stream>>>(h->dp, h->W, line_search);
    return 0;
}

// one thread per (problem, timestep): stacked task Jacobian (test hook)
__global__ void k_rollout_jac(const DevProblem* __restrict__ Pp, Work W, double* __restrict__ Jout) {
    const DevProblem& P = *Pp;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P.B * P.H) return;
    const int b = idx / P.H, k = idx % P.H, n = P.n, dof = P.dof;
    const double* x = W.X + ((size_t)b * P.H + k) * W.ns;
    double xs[MAXN], dqs[MAXM];
    for (int i = 0; i < n; i++) xs[i] = x[i];
    for (int i = 0; i < dof; i++) dqs[i] = (P.nb_deriv == 2) ? xs[dof + i] : 0.0;
    double* out = Jout + (size_t)idx * P.r * n;
    for (int t = 0; t < P.n_tasks; t++) {
        const DevTask& tk = P.tasks[t];
        EEState e;
        task_ee(P, tk, xs, dqs, e);
        task_jacobian(P, tk, e, out + (size_t)tk.r_off * n, n);
    }
}
__global__ void k_kinematics(const DevProblem* __restrict__ Pp, int robot, int nq, const double* __restrict__ q, const double* __restrict__ dq, double* x, double* quat,
                             double* J, double* dx, double* w) {
    const DevProblem& P = *Pp;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    const DevRobot& rb = P.robots[robot];
    double qs[RDOF], dqs[RDOF];
    for (int c = 0; c < RDOF; c++) {
        qs[c] = c < rb.dof ? q[(size_t)i * rb.dof + c] : 0.0;
        dqs[c] = c < rb.dof ? dq[(size_t)i * rb.dof + c] : 0.0;
    }
    EEState e;
    if (rb.kind == ILQRB_ROBOT_2D) robot2d_update_kinematics(rb.lengths, rb.dof, qs, dqs, e);
    else kdl_update_kinematics(rb.segs, rb.n_segments, rb.dof, qs, dqs, e);
    for (int a = 0; a < 3; a++) {
        x[(size_t)i * 3 + a] = e.x[a];
        dx[(size_t)i * 3 + a] = e.dx[a];
        w[(size_t)i * 3 + a] = e.w[a];
    }
    for (int a = 0; a < 4; a++) quat[(size_t)i * 4 + a] = e.quat[a];
    for (int a = 0; a < 6; a++)
        for (int c = 0; c < rb.dof; c++) J[((size_t)i * 6 + a) * rb.dof + c] = a < 3 ? e.Jv[c][a] : e.Jw[c][a - 3];
}

}  // namespace

int launch_backward_generic(ilqrb_handle h, bool al, double pf) {
    h->begin(PS_BACKWARD);
    int rc;
    if (h->hp.time) rc = al ? launch_backward_ta<true, true>(h, pf) : launch_backward_ta<true, false>(h, pf);
    else rc = al ? launch_backward_ta<false, true>(h, pf) : launch_backward_ta<false, false>(h, pf);
    h->end();
    if (rc) return rc;
    CU(cudaGetLastError());
    return 0;
}

int launch_init(ilqrb_handle h, int b0, int nB, cudaStream_t st) {
    if (nB < 0) nB = h->hp.B;
    if (!st) st = h->stream;
    h->begin(PS_ROLLOUT_INIT);
    const int tot = nB * h->hp.H;
    {  // threads per problem: joints + time state
        const int need = h->hp.dof + (h->hp.time ? 1 : 0);
        if (need <= 8) k_rollout_x<8><<<(nB * 8 + 127) / 128, 128, 0, st>>>(h->dp, h->W, b0, nB);
        else if (need <= 16) k_rollout_x<16><<<(nB * 16 + 127) / 128, 128, 0, st>>>(h->dp, h->W, b0, nB);
        else k_rollout_x<32><<<(nB * 32 + 127) / 128, 128, 0, st>>>(h->dp, h->W, b0, nB);
    }
#define CALL(N, M) k_stage_costs<N, M><<<(tot + 127) / 128, 128, 0, st>>>(h->dp, h->W, b0, nB)
    DISPATCH_NM(size_class(h->hp.n, h->hp.m), CALL)
#undef CALL
    k_sum_costs<<<(nB + 31) / 32, 32, 0, st>>>(h->dp, h->W, b0, nB);
    h->end();
    CU(cudaGetLastError());
    return 0;
}
int launch_derivs(ilqrb_handle h, int b0, int nB, cudaStream_t st) {
    if (nB < 0) nB = h->hp.B;
    if (!st) st = h->stream;
    const int tot = nB * h->hp.n_kp_steps;
    h->begin(PS_KP_DERIVS);
    k_kp_derivs<<<(tot + 63) / 64, 64, 0, st>>>(h->dp, h->W, b0, nB);
    h->end();
    CU(cudaGetLastError());
    return 0;
}
int launch_linesearch(ilqrb_handle h, int line_search, bool al) {
    const int tot = h->hp.B * CAND_LANES;
    h->begin(PS_LINESEARCH);
    const size_t need = 4 * (size_t)ls_affine_smem_doubles(h->hp.H, h->W.ns, h->W.ms) * sizeof(double);
    if (!generic_forced() && !h->hp.time && h->hp.m <= 8 && need <= 200 * 1024) {  // AL line searches on the plain cost too
        int rc = 0;
#define CALL(N, M) rc = launch_linesearch_affine_t<N, M>(h, line_search)
        DISPATCH_NM(size_class(h->hp.n, h->hp.m), CALL)
#undef CALL
        h->end();
        if (rc) return rc;
        CU(cudaGetLastError());
        return 0;
    }
#define CALL(N, M) k_linesearch<N, M><<<(tot + 127) / 128, 128, 0, h->stream>>>(h->dp, h->W, line_search)
    DISPATCH_NM(size_class(h->hp.n, h->hp.m), CALL)
#undef CALL
    h->end();
    CU(cudaGetLastError());
    return 0;
}
int launch_accept(ilqrb_handle h, int it, int nb_iter, int line_search, int early_stop, int al, int lag, double pen_next) {
    const int tot = h->hp.B * 32;
    h->begin(PS_ACCEPT);
#define CALL(N, M) k_accept<N, M><<<(tot + 127) / 128, 128, 0, h->stream>>>(h->dp, h->W, it, nb_iter, line_search, early_stop, al, lag, pen_next)
    DISPATCH_NM(size_class(h->hp.n, h->hp.m), CALL)
#undef CALL
    h->end();
    CU(cudaGetLastError());
    return 0;
}
int launch_al_init(ilqrb_handle h) {
    h->begin(PS_MISC);
    k_al_init<<<(h->hp.B * 32 + 127) / 128, 128, 0, h->stream>>>(h->dp, h->W);
    h->end();
    CU(cudaGetLastError());
    return 0;
}
int launch_final_fx(ilqrb_handle h, int b0, int nB, cudaStream_t st) {
    const int tot = nB * h->hp.H;
    k_final_fx<<<(tot + 127) / 128, 128, 0, st>>>(h->dp, h->W, h->d_fX, b0, nB);
    CU(cudaGetLastError());
    return 0;
}
int launch_rollout_jac(ilqrb_handle h, double* Jout) {
    const int tot = h->hp.B * h->hp.H;
    k_rollout_jac<<<(tot + 127) / 128, 128, 0, h->stream>>>(h->dp, h->W, Jout);
    CU(cudaGetLastError());
    return 0;
}
int launch_kinematics(ilqrb_handle h, int robot, int nq, const double* q, const double* dq, double* x, double* quat, double* J, double* dx, double* w) {
    h->begin(PS_MISC);
    k_kinematics<<<(nq + 127) / 128, 128, 0, h->stream>>>(h->dp, robot, nq, q, dq, x, quat, J, dx, w);
    h->end();
    CU(cudaGetLastError());
    return 0;
}


// ---- kinematics hook without any allocation -------------------------------------------------------------------------
// ilqrb_kinematics stages its inputs and outputs in a cudaMalloc'ed buffer per call.  Callers that cannot allocate (code under
// CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc) get the same kernel with its staging area in a
// file-scope device array: every GPU that loads the library holds these ILQRB_KIN_STATIC_MAX_NQ * 69 doubles (2.26 MB) until
// exit.  The array is shared by all handles and calls of the process on that device and keeps the last call's contents: the
// velocity slots are cleared when dq is NULL, everything else is overwritten before it is read; two calls that overlap on the
// GPU (different streams or handles, no synchronisation between them) race on it.
static __device__ __align__(256) double g_kin_scratch[(size_t)ILQRB_KIN_STATIC_MAX_NQ * (2 * RDOF + 3 + 4 + 6 * RDOF + 3 + 3)];

namespace {
// In host code `g_kin_scratch` is a host-side stand-in, not a device address: the address a kernel may be given comes from
// cudaGetSymbolAddress, asked once per device with that device current.
int kin_scratch_ptr(int device, double** out) {
    static double* ptr[MAX_DEVICES];
    static std::atomic<bool> have[MAX_DEVICES];
    if (device < 0 || device >= MAX_DEVICES) return fail(ILQRB_EINVAL, "device index beyond the per-device table");
    if (!have[device].load(std::memory_order_acquire)) {
        double* p = nullptr;
        CU(cudaGetSymbolAddress((void**)&p, g_kin_scratch));
        ptr[device] = p;
        have[device].store(true, std::memory_order_release);
    }
    *out = ptr[device];
    return 0;
}
}  // namespace

}  // namespace ilqrb

extern "C" int ilqrb_kinematics_static(ilqrb_handle h, int robot, int nq, const double* q, const double* dq, double* x, double* quat, double* J, double* dx, double* w) {
    using namespace ilqrb;
    if (!h) return fail(ILQRB_EINVAL, "null handle");
    if (robot < 0 || robot >= h->hp.n_robots) return fail(ILQRB_EINVAL, "robot out of range");
    if (nq < 1 || !q) return fail(ILQRB_EINVAL, "no configurations");
    if (nq > ILQRB_KIN_STATIC_MAX_NQ)
        return fail(ILQRB_EINVAL, "ilqrb_kinematics_static: nq = " + std::to_string(nq) + " exceeds the static scratch area (ILQRB_KIN_STATIC_MAX_NQ = " +
                                      std::to_string(ILQRB_KIN_STATIC_MAX_NQ) + " configurations); split the call or use ilqrb_kinematics");
    CU(cudaSetDevice(h->device));
    double* buf = nullptr;
    int rc = kin_scratch_ptr(h->device, &buf);
    if (rc) return rc;
    const int dof = h->hp.robots[robot].dof;
    const size_t in = (size_t)nq * dof;
    double *dq_d = buf + in, *ox = buf + 2 * in, *oq = ox + (size_t)nq * 3, *oJ = oq + (size_t)nq * 4, *odx = oJ + (size_t)nq * 6 * dof, *ow = odx + (size_t)nq * 3;
    CU(cudaMemcpyAsync(buf, q, in * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    if (dq) CU(cudaMemcpyAsync(dq_d, dq, in * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    else CU(cudaMemsetAsync(dq_d, 0, in * sizeof(double), h->stream));  // the array keeps the last call's contents
    if ((rc = launch_kinematics(h, robot, nq, buf, dq_d, ox, oq, oJ, odx, ow))) return rc;
    if (x) CU(cudaMemcpyAsync(x, ox, (size_t)nq * 3 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    if (quat) CU(cudaMemcpyAsync(quat, oq, (size_t)nq * 4 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    if (J) CU(cudaMemcpyAsync(J, oJ, (size_t)nq * 6 * dof * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    if (dx) CU(cudaMemcpyAsync(dx, odx, (size_t)nq * 3 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    if (w) CU(cudaMemcpyAsync(w, ow, (size_t)nq * 3 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return 0;
}
