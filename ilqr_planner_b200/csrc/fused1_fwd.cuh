// fused1_fwd.cuh — forward pass of the single-integrator fast path as TWO kernels (the single k_forward_n1 of fused1.cuh stays
// selectable with ILQRB_FWD_V1=1 for A/B runs):
//
//   k_fwd_rollout_n1   phase A: the alpha = 1 closed-loop deviation rollout  du_k = K_k dx_k + d_k, dx_{k+1} = dx_k + dt du_k
//                      (ILQRRecursive.cpp:124-136 minus the nominal; linear dynamics make every candidate affine in alpha).
//                      One problem per warp.  The gains are the only large stream of the forward pass (K: 38.8 KB per problem and
//                      iteration at C2): they are staged by the bulk-copy engine — `cp.async.bulk.shared::cluster.global` (1-D TMA,
//                      SASS UBLKCP) into a shared-memory ring of RING stages of SPS steps each, completion on one mbarrier per stage
//                      (`mbarrier::complete_tx::bytes`) — so that RING * SPS steps of gains are in flight per warp and no global load
//                      sits on the serial 99-step chain.  K_k is the DMMA A operand: with the contraction index of slice s, lane t
//                      mapped to column 2t + s, the two A fragments of a step are ONE 16-byte shared-memory load per lane, and the
//                      whole warp reads the 512 contiguous bytes of K_k without bank conflicts.
//                      du goes to a global scratch (L2-resident until k_fwd_select_n1 reads it), dx at the keypoint steps too.
//
//   k_fwd_select_n1    phases B-F of forward_n1_body<.., SPLIT = true>: candidate costs (lane = (candidate, keypoint step), lean FK),
//                      the reference's sequential choice, accept (u += alpha du, x re-integrated in the reference's operation
//                      order), l_x / l_xx of the accepted trajectory.  The FK sweeps want ~100 registers: with its own launch bound
//                      this kernel holds them without the local-memory spills the fused kernel has at its 72-register cap
//                      (ncu: 4.8 M + 4.1 M local-memory sectors per launch).
#pragma once
#include "fused1.cuh"

namespace ilqrb {

// ---- mbarrier / bulk-copy PTX (sm_90+; sm_100a here) ---------------------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// global -> shared bulk copy (size and both addresses multiples of 16 bytes), completion counted on `bar`
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes),
                 "r"(smem_u32(bar))
                 : "memory");
}

// shared memory of one warp of k_fwd_rollout_n1: RING stages of SPS steps each: K [SPS][m][ns], d [SPS][ms], X [SPS][ns] (doubles)
template <int SPS>
__host__ __device__ inline int fwd_ring_stage_doubles(int m, int ns) { return SPS * (m * ns + 2 * ns); }
template <int RING, int SPS>
__host__ __device__ inline size_t fwd_ring_smem_bytes(int m, int ns, int warps) { return (size_t)warps * (RING * fwd_ring_stage_doubles<SPS>(m, ns) * 8 + RING * 8); }

// phase A for problem b by one warp; ring: RING stages of this warp, bars: its RING mbarriers
template <int NT, int DOF, int RING, int SPS>
__device__ __forceinline__ void fwd_rollout_ring(const DevProblem& P, const Work& W, int b, int lane, double* ring, unsigned long long* bars) {
    constexpr unsigned FULL = 0xffffffffu;
    const int g = lane >> 2, t = lane & 3;
    const int n = DOF > 0 ? DOF : P.n, m = n, H = P.H, S = P.n_kp_steps;
    const int ns = (n + 1) & ~1, ms = ns;
    const double dt = P.dt;
    const int stage_d = fwd_ring_stage_doubles<SPS>(m, ns);
    const double* Xb = W.X + (size_t)b * H * ns;
    const double* Kb = W.K + (size_t)b * (H - 1) * m * ns;
    const double* db = W.d + (size_t)b * (H - 1) * ms;
    double* DUb = W.DU + (size_t)b * (H - 1) * ms;
    double* DXb = W.DXk + (size_t)b * S * ns;

    // joint limits of the first limit-carrying task in registers (weight 0 -> +-inf), as in k_backward_n1
    int l0 = -1, l1 = -1;
    if (P.any_limits)
        for (int ti = P.n_tasks - 1; ti >= 0; ti--)
            if (P.tasks[ti].limits_set) {
                l1 = l0;
                l0 = ti;
            }
    double lmx[NT], lmn[NT];
#pragma unroll
    for (int rt = 0; rt < NT; rt++) {
        const int i = 8 * rt + g;
        lmx[rt] = INFINITY;
        lmn[rt] = -INFINITY;
        if (l0 >= 0 && i < n && P.tasks[l0].limit_weight[i] != 0) {
            lmx[rt] = P.tasks[l0].state_max[i];
            lmn[rt] = P.tasks[l0].state_min[i];
        }
    }
    const bool want_x = l0 >= 0;  // warp-uniform: the state rows travel with the gains only when a limit can be violated

    // ---- the ring: stage q holds steps [q SPS, (q+1) SPS) of the horizon, in slot q mod RING -------------------------------
    const int n_stages = (H - 1 + SPS - 1) / SPS;
    auto issue = [&](int q) {  // lane 0 only
        const int slot = q % RING, k0 = q * SPS;
        const int steps = (H - 1 - k0) < SPS ? (H - 1 - k0) : SPS;
        double* sK = ring + (size_t)slot * stage_d;
        double* sd = sK + SPS * m * ns;
        double* sx = sd + SPS * ms;
        const unsigned bK = (unsigned)(steps * m * ns * 8), bd = (unsigned)(steps * ms * 8), bx = want_x ? (unsigned)(steps * ns * 8) : 0u;
        mbar_expect_tx(&bars[slot], bK + bd + bx);
        bulk_g2s(sK, Kb + (size_t)k0 * m * ns, bK, &bars[slot]);
        bulk_g2s(sd, db + (size_t)k0 * ms, bd, &bars[slot]);
        if (want_x) bulk_g2s(sx, Xb + (size_t)(k0 + 1) * ns, bx, &bars[slot]);
    };
    if (lane == 0) {
        for (int s = 0; s < RING; s++) mbar_init(&bars[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    if (lane == 0)
        for (int q = 0; q < RING && q < n_stages; q++) issue(q);

    bool viol = (l1 >= 0);  // several limit-carrying sub-systems: the select kernel always takes its exact path
    if (l0 >= 0) {          // the initial state is common to all candidates but still pays its penalty
#pragma unroll
        for (int rt = 0; rt < NT; rt++) {
            const int i = 8 * rt + g;
            const double x00 = (i < n) ? Xb[i] : 0.0;
            viol = viol || (x00 > lmx[rt]) || (x00 < lmn[rt]);
        }
    }
    bool vrow[NT], vcol[NT];  // validity of this lane's fragment rows / of its column pair (2t, 2t+1) per k-tile
#pragma unroll
    for (int rt = 0; rt < NT; rt++) {
        vrow[rt] = 8 * rt + g < m;
        vcol[rt] = 8 * rt + 2 * t < ns;  // the pad column of K holds zeros (the backward pass writes them)
    }
    double dxr[NT], bx[NT][2];
#pragma unroll
    for (int rt = 0; rt < NT; rt++) {
        dxr[rt] = 0.0;
        bx[rt][0] = bx[rt][1] = 0.0;
    }
    int skp = 0;
    int next_kp = S > 0 ? P.kp_steps[0] : -1;
    if (next_kp == 0) {  // a keypoint on the initial state: dx = 0
        if (lane < ns) DXb[lane] = 0.0;
        skp++;
        next_kp = skp < S ? P.kp_steps[skp] : -1;
    }
    for (int q = 0; q < n_stages; q++) {
        const int slot = q % RING, k0 = q * SPS;
        const double* sK = ring + (size_t)slot * stage_d;
        const double* sd = sK + SPS * m * ns;
        const double* sx = sd + SPS * ms;
        mbar_wait(&bars[slot], (unsigned)((q / RING) & 1));
#pragma unroll
        for (int j = 0; j < SPS; j++) {
            const int k = k0 + j;
            if (k < H - 1) {  // warp-uniform
#pragma unroll
                for (int rt = 0; rt < NT; rt++) {
                    double c0 = 0.0, c1 = 0.0;
#pragma unroll
                    for (int kt = 0; kt < NT; kt++) {
                        // A fragments of slices 0 and 1: K_k[8rt+g][8kt + 2t], [.. + 2t + 1] in one 16-byte load
                        double2 a = make_double2(0.0, 0.0);
                        if (vrow[rt] && vcol[kt]) a = *reinterpret_cast<const double2*>(sK + (size_t)j * m * ns + (8 * rt + g) * ns + 8 * kt + 2 * t);
                        dmma884(c0, c1, a.x, bx[kt][0]);
                        dmma884(c0, c1, a.y, bx[kt][1]);
                    }
                    const double dk = vrow[rt] ? sd[j * ms + 8 * rt + g] : 0.0;
                    const double du = c0 + dk;
                    dxr[rt] = dxr[rt] + dt * du;
                    if (t == 0 && 8 * rt + g < ms) DUb[(size_t)k * ms + 8 * rt + g] = du;
                    if (want_x) {
                        const double x0 = vrow[rt] ? sx[j * ns + 8 * rt + g] : 0.0, x1 = x0 + dxr[rt];  // state k+1 (state 0 is fixed and shared)
                        viol = viol || (x0 > lmx[rt]) || (x0 < lmn[rt]) || (x1 > lmx[rt]) || (x1 < lmn[rt]);
                    }
                }
                if (k + 1 == next_kp) {
#pragma unroll
                    for (int rt = 0; rt < NT; rt++)
                        if (t == 0 && 8 * rt + g < ns) DXb[skp * ns + 8 * rt + g] = dxr[rt];
                    skp++;
                    next_kp = skp < S ? P.kp_steps[skp] : -1;
                }
                // B fragments of the next step: slice s of lane t carries dx[8kt + 2t + s], held by the lanes of row group 2t + s
#pragma unroll
                for (int kt = 0; kt < NT; kt++)
#pragma unroll
                    for (int s = 0; s < 2; s++) bx[kt][s] = __shfl_sync(FULL, dxr[kt], 4 * (2 * t + s));
            }
        }
        __syncwarp();  // every lane is done with the slot before the copy engine refills it
        if (lane == 0 && q + RING < n_stages) issue(q + RING);
    }
    viol = __any_sync(FULL, viol);
    if (lane == 0) W.viol[b] = viol ? 1 : 0;
}

template <int NT, int DOF, int RING, int SPS, int WARPS, int MINB>
__global__ void __launch_bounds__(32 * WARPS, MINB) k_fwd_rollout_n1(const DevProblem* __restrict__ Pp, Work W, int b0, int nB) {
    extern __shared__ __align__(128) unsigned char fwd_smem[];
    const DevProblem& P = *Pp;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int bl = blockIdx.x * WARPS + wid;
    if (bl >= nB) return;
    const int b = b0 + bl;
    if (W.done[b]) return;
    const int n = DOF > 0 ? DOF : P.n, ns = (n + 1) & ~1;
    const int stage_d = fwd_ring_stage_doubles<SPS>(n, ns);
    double* ring = reinterpret_cast<double*>(fwd_smem) + (size_t)wid * RING * stage_d;
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(reinterpret_cast<double*>(fwd_smem) + (size_t)WARPS * RING * stage_d) + wid * RING;
    fwd_rollout_ring<NT, DOF, RING, SPS>(P, W, b, lane, ring, bars);
}

// The whole forward pass in ONE kernel again (default): phase A through the bulk-copy ring, then phases B-F on the deviation it left
// in global memory (L2-resident: written and re-read by the same warp), warps of different problems in different phases sharing an SM —
// the FK chains of phases C and F are latency-bound (ncu on k_fwd_select_n1 alone: 24 % issue-active, 8 warps per issue on the long
// scoreboard) and need the rollout warps next to them to fill the issue slots.  Per warp: the ring (phase A), reused for the Jacobian
// columns of phase F, + [NCAND][S] candidate costs.
template <int RING, int SPS>
__host__ __device__ inline size_t fwd_fused_warp_bytes(int m, int ns, int S) {
    size_t ringb = (size_t)RING * fwd_ring_stage_doubles<SPS>(m, ns) * 8;
    if (ringb < 32 * 7 * 8) ringb = 32 * 7 * 8;
    return (ringb + (size_t)NCAND * S * 8 + RING * 8 + 127) / 128 * 128;  // bulk copies need 16-byte aligned destinations
}
template <int NT, int DOF, int RING, int SPS, int WARPS, int MINB, int RR = 6>
__global__ void __launch_bounds__(32 * WARPS, MINB) k_forward_ring_n1(const DevProblem* __restrict__ Pp, Work W, int b0, int nB, int it, int nb_iter, int line_search,
                                                                      int early_stop) {
    extern __shared__ __align__(128) unsigned char fwd_smem[];
    const DevProblem& P = *Pp;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int bl = blockIdx.x * WARPS + wid;
    if (bl >= nB) return;
    const int b = b0 + bl;
    if (W.done[b]) return;
    const int n = DOF > 0 ? DOF : P.n, ns = (n + 1) & ~1, S = P.n_kp_steps;
    unsigned char* mine = fwd_smem + (size_t)wid * fwd_fused_warp_bytes<RING, SPS>(n, ns, S);
    double* ring = reinterpret_cast<double*>(mine);
    size_t ringb = (size_t)RING * fwd_ring_stage_doubles<SPS>(n, ns) * 8;
    if (ringb < 32 * 7 * 8) ringb = 32 * 7 * 8;
    double* s_cc = reinterpret_cast<double*>(mine + ringb);
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(mine + ringb + (size_t)NCAND * S * 8);
    fwd_rollout_ring<NT, DOF, RING, SPS>(P, W, b, lane, ring, bars);
    __syncwarp();  // the deviation and the limit flag written by some lanes are read by others
    forward_n1_body<NT, DOF, true, RR>(P, W, b, lane, ring, it, nb_iter, line_search, early_stop, s_cc);
}

// phases B-F: one warp per block (the FK sweeps of phases C and F get the registers they need)
__host__ __device__ inline int fwd_select_smem_doubles(int S) { return 32 * 7 + NCAND * S; }

template <int NT, int DOF, int MINB>
__global__ void __launch_bounds__(32, MINB) k_fwd_select_n1(const DevProblem* __restrict__ Pp, Work W, int b0, int it, int nb_iter, int line_search, int early_stop) {
    extern __shared__ double sel_smem[];
    const int b = b0 + blockIdx.x;
    if (W.done[b]) return;
    forward_n1_body<NT, DOF, true>(*Pp, W, b, threadIdx.x, sel_smem, it, nb_iter, line_search, early_stop);
}

}  // namespace ilqrb
