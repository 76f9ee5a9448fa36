// launch_fused2.cu — launcher of the double-integrator fast backward pass (fused2.cuh: k_backward_n2, with and without AL terms).
#include "fused2.cuh"
#include "handle.hpp"

namespace ilqrb {

// nb_deriv = 2, no time state, dof <= 7, at most 32 AL rows
bool fast_n2_applies(ilqrb_handle h, bool al) {
    const DevProblem& P = h->hp;
    (void)al;
    return !generic_forced() && !P.time && P.nb_deriv == 2 && P.dof <= 7 && P.n == 2 * P.dof && P.m == P.dof && h->W.nc <= 32;
}

namespace {
template <int DOF, bool AL, int WARPS, int MINB>
int launch_backward_n2_t(ilqrb_handle h, double pf) {
    const size_t smem = (size_t)WARPS * BwdN2Smem::PER_WARP * sizeof(double);
    auto kern = k_backward_n2<DOF, AL, WARPS, MINB>;
    static DynSmem attr;
    int rc = attr.ensure(kern, smem, h->device);
    if (rc) return rc;
    const int B = h->hp.B;
    kern<<<(B + WARPS - 1) / WARPS, 32 * WARPS, smem, h->stream>>>(h->dp, h->W, 0, B, pf);
    return 0;
}
}  // namespace

int launch_backward_n2(ilqrb_handle h, bool al, double pf) {
    h->begin(PS_BACKWARD);
    int rc;
    if (al) rc = h->hp.dof == 7 ? launch_backward_n2_t<7, true, 4, 4>(h, pf) : launch_backward_n2_t<0, true, 4, 4>(h, pf);
    else rc = h->hp.dof == 7 ? launch_backward_n2_t<7, false, 4, 4>(h, pf) : launch_backward_n2_t<0, false, 4, 4>(h, pf);
    h->end();
    if (rc) return rc;
    CU(cudaGetLastError());
    return 0;
}

}  // namespace ilqrb
