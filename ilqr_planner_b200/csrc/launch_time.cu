// launch_time.cu — launchers of the time-system fast path (fused_time.cuh): k_backward_time / k_forward_time.
#include <cstdlib>

#include "fused_time.cuh"
#include "handle.hpp"

namespace ilqrb {

// PosOrnTimePlannerSys with nb_deriv = 1 and end-effector tasks on robots of at most 7 joints (n = m = dof + 1 <= 8)
bool fast_time_applies(ilqrb_handle h, bool al) {
    const DevProblem& P = h->hp;
    static const bool off_switch = getenv("ILQRB_NO_FAST_TIME") != nullptr;  // A/B switch for tests and profiling
    if (off_switch || generic_forced() || al || !P.time || P.nb_deriv != 1 || P.n != P.m || P.n != P.dof + 1 || P.n > 8) return false;
    for (int t = 0; t < P.n_tasks; t++)
        if (P.tasks[t].kind != ILQRB_TASK_POSE) return false;  // joint-space tasks: generic kernels
    return true;
}

namespace {

template <int DOF, int WARPS, int MINB>
int launch_backward_time_t(ilqrb_handle h, int b0, int nB, cudaStream_t st) {
    const size_t smem = (size_t)WARPS * BwdTimeSmem::PER_WARP * sizeof(double);
    auto kern = k_backward_time<DOF, WARPS, MINB>;
    static DynSmem attr;
    int rc = attr.ensure(kern, smem, h->device);
    if (rc) return rc;
    kern<<<(nB + WARPS - 1) / WARPS, 32 * WARPS, smem, st>>>(h->dp, h->W, b0, nB);
    return 0;
}

template <int DOF, int RING, int SPS, int WARPS, int MINB>
int launch_forward_time_t(ilqrb_handle h, int b0, int nB, cudaStream_t st, int it, int nb_iter, int line_search, int early_stop) {
    const int n = h->hp.n, ns = h->W.ns;
    const size_t smem = 128 + (size_t)WARPS * fwd_time_warp_bytes<RING, SPS>(n, ns, h->hp.n_kp_steps);
    if (smem > 220 * 1024) return fail(ILQRB_EUNSUPPORTED, "time-system forward pass: too many keypoint steps for the shared-memory candidate slots");
    auto kern = k_forward_time<DOF, RING, SPS, WARPS, MINB>;
    static DynSmem attr;
    int rc = attr.ensure(kern, smem, h->device);
    if (rc) return rc;
    kern<<<(nB + WARPS - 1) / WARPS, 32 * WARPS, smem, st>>>(h->dp, h->W, b0, nB, it, nb_iter, line_search, early_stop);
    return 0;
}

}  // namespace

int launch_derivs_time(ilqrb_handle h, int b0, int nB, cudaStream_t st) {
    h->begin(PS_KP_DERIVS);
    k_kp_derivs_time<4><<<(nB + 3) / 4, 128, 0, st>>>(h->dp, h->W, b0, nB);
    h->end();
    CU(cudaGetLastError());
    return 0;
}

int launch_backward_time(ilqrb_handle h, int b0, int nB, cudaStream_t st) {
    h->begin(PS_BACKWARD);
    int rc = h->hp.dof == 7 ? launch_backward_time_t<7, 4, 4>(h, b0, nB, st) : launch_backward_time_t<0, 4, 4>(h, b0, nB, st);
    h->end();
    if (rc) return rc;
    CU(cudaGetLastError());
    return 0;
}

int launch_forward_time(ilqrb_handle h, int b0, int nB, cudaStream_t st, int it, int nb_iter, int line_search, int early_stop) {
    h->begin(PS_LINESEARCH);
    // ring shape <RING, SPS>: stages in flight x steps per stage (4 bulk copies per stage); ILQRB_TIME_FWD picks another for sweeps
    static const int variant = getenv("ILQRB_TIME_FWD") ? atoi(getenv("ILQRB_TIME_FWD")) : 0;
    int rc;
#define FWD(RING, SPS, WARPS, MINB)                                                                                             \
    (h->hp.dof == 7 ? launch_forward_time_t<7, RING, SPS, WARPS, MINB>(h, b0, nB, st, it, nb_iter, line_search, early_stop) \
                    : launch_forward_time_t<0, RING, SPS, WARPS, MINB>(h, b0, nB, st, it, nb_iter, line_search, early_stop))
    switch (variant) {
        case 1: rc = FWD(3, 2, 4, 4); break;
        case 2: rc = FWD(2, 8, 4, 4); break;
        case 3: rc = FWD(3, 4, 4, 4); break;
        case 4: rc = FWD(2, 4, 4, 5); break;
        case 5: rc = FWD(2, 2, 4, 6); break;
        case 6: rc = FWD(2, 4, 2, 10); break;
        default: rc = FWD(2, 4, 4, 4); break;
    }
#undef FWD
    h->end();
    if (rc) return rc;
    CU(cudaGetLastError());
    return 0;
}

}  // namespace ilqrb
