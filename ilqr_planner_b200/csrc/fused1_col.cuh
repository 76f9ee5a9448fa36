// fused1_col.cuh — backward Riccati pass of the single-integrator fast path with FOUR problems per warp (n = m <= 7, and per 7-joint
// block of composite robots).
//
// k_backward_n1 (fused1.cuh) gives a whole warp to one problem: the 8x8 augmented matrices sit in the DMMA accumulator layout, two
// elements per lane, and the in-register Gauss-Jordan spends ~37 instructions per stage on lane-role bookkeeping for 4 useful FP64
// operations per lane (ncu: 397 instructions per problem and step, 100 of them FP64, issue-bound).  Here 8 lanes own one problem and
// lane j owns COLUMN j of every augmented matrix ([P|p], [Qux|Qu], [Qxx|Qx], [K|d]; column 7 is the vector), eight rows in
// registers with compile-time indices:
//   * the elimination of [Quu + 1e-6 I | -Qux -Qu] needs one broadcast per stage — the multiplier column l_ik / l_kk, 7 doubles
//     shuffled from the pivot column's owner — and then updates its two columns with 14 independent FMAs; it is a solve (no
//     explicit inverse, no product with it);
//   * products C = A B take column k of A as a shared-memory broadcast and B[k][j] from the lane's own registers: 8 FMAs per 16-byte
//     pair of loads.  Quu [K|d] is formed as dt^2 (P [K|d]) + R o [K|d], so one product P [K|d] serves both the bracket
//     M = Quu [K|d] + [Qux|Qu] and Qxu [K|d] = dt P [K|d]; the second product is K^T M.            ILQRRecursive.cpp:82-94
// ~175 instructions per problem and step (ncu), the same FP64 work.  Plain FMA chains in the reference's summation order; no tensor pipe
// (DMMA and DFMA peak at the same rate on B200: 37.0 against 36.1 TFLOP/s measured, bench.py `peaks_measured_here_tflops`).
// A problem whose elimination meets a non-positive pivot is redone by the warp-per-problem pass with the pivoted inverse
// (backward_n1_redo), like there.
#pragma once
#include "fused1.cuh"

namespace ilqrb {

struct BwdColSmem {
    // per problem: [K|d] row-major (odd problems shifted by 32 bytes so that the 64-byte row stores of two problems fall on disjoint
    // banks), P column-major with a column stride of 10 doubles (the 16-byte column stores of 8 lanes hit 8 different bank groups),
    // the limit residuals.  The block stride keeps the problems of a half-warp on distinct 16-byte bank groups for the broadcast
    // loads (164 * 8 B = 32 mod 128).
    static constexpr int KS = 0, PC = 68, PCS = 10, LQ = 148, BLK = 164;
    static constexpr int PER_WARP = 4 * BLK > BwdN1Smem<1>::PER_WARP ? 4 * BLK : BwdN1Smem<1>::PER_WARP;
};

__device__ __forceinline__ void ld8(const double* p, double (&o)[8]) {
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const double2 v = *reinterpret_cast<const double2*>(p + 2 * q);
        o[2 * q] = v.x;
        o[2 * q + 1] = v.y;
    }
}
__device__ __forceinline__ void st8(double* p, const double (&o)[8]) {
#pragma unroll
    for (int q = 0; q < 4; q++) *reinterpret_cast<double2*>(p + 2 * q) = make_double2(o[2 * q], o[2 * q + 1]);
}

// returns the ballot of lanes whose problem met a non-positive pivot.  The 8 lanes of a group work on problem b (idle groups shadow
// a live problem and store nothing).  BLK = false: the problem's state is the 7 joints (row stride 8 doubles: 16-byte accesses).
// BLK = true: the group handles the diagonal block [off, off + 7) of a problem whose tasks touch disjoint joint ranges (composite
// robots, C5; see backward_n1_body): strides are those of the full state.
// N < 7 (BLK = false): the same lane groups of 8 with lanes N + 1 .. 7 idle (their columns stay zero); global rows are then narrower than
// 8 doubles and take the scalar accesses of the block form.
template <int N, bool BLK>
__device__ __forceinline__ unsigned backward_n1_col(const DevProblem& P, const Work& W, int b, bool live, int off, int lane, double* wsmem) {
    static_assert(N >= 1 && N <= 7, "8 x 8 augmented matrices");
    constexpr unsigned FULL = 0xffffffffu;
    constexpr bool VEC = !BLK && N == 7;  // 16-byte accesses to the rows of U and d
    const int nf = BLK ? P.n : N, ns = (nf + 1) & ~1, ms = ns;
    const int sub = lane >> 3, j = lane & 7;
    const int H = P.H;
    const double dt = 1.0 * P.dt;  // B = Identity * dt
    double* sm = wsmem + sub * BwdColSmem::BLK;
    double* sK = sm + BwdColSmem::KS + 4 * (sub & 1);
    const int gsrc = lane & ~7;  // first lane of this problem's group
    double* sP = sm + BwdColSmem::PC;
    double* sQ = sm + BwdColSmem::LQ;

    // R, and per-lane constants that stand in for selects on the lane's column index: the diagonal terms of Quu + 1e-6 I are added
    // as x + 0 everywhere else (exact), R u enters the vector column only
    double R[N], Rd[N], Ed[N], Rv[N];
#pragma unroll
    for (int i = 0; i < N; i++) {
        R[i] = P.r_diag[off + i];
        Rd[i] = (i == j) ? R[i] : 0.0;
        Ed[i] = (i == j) ? 1e-6 * 1.0 : 0.0;
        Rv[i] = (j == N) ? R[i] : 0.0;
    }
    const double* Xb = W.X + (size_t)b * H * ns + off;
    const double* Ub = W.U + (size_t)b * (H - 1) * ms + off;
    const int S_kp = P.n_kp_steps;

    // joint limits: lane j checks state off + j against the one sub-system that limits it (the launcher sends systems where two
    // sub-systems limit the same state to the warp-per-problem kernel)   System.cpp:121-142, SequentialSystem.cpp:119-165
    int l0 = -1;
    double lmx = INFINITY, lmn = -INFINITY;
    if (P.any_limits)
        for (int ti = 0; ti < P.n_tasks; ti++)
            if (P.tasks[ti].limits_set) {
                l0 = ti;
                if (j < N && P.tasks[ti].limit_weight[off + j] != 0) {
                    lmx = P.tasks[ti].state_max[off + j];
                    lmn = P.tasks[ti].state_min[off + j];
                }
            }
    // e += [L^T L | -L^T q] with L = diag(violated), q = limit - x                      System.cpp:263-269,299-305
    auto add_limits = [&](double (&e)[8], double x) {
        double q = 0.0;
        bool act = false;
        if (x > lmx) {
            q = lmx - x;
            act = true;
        } else if (x < lmn) {
            q = lmn - x;
            act = true;
        }
        sQ[j] = act ? -1.0 * q : 0.0;
        __syncwarp();
        double qv[8];
        ld8(sQ, qv);
#pragma unroll
        for (int i = 0; i < N; i++) {
            if (i == j && act) e[i] += 1.0 * 1.0;
            if (j == N) e[i] += qv[i];
        }
        __syncwarp();
    };
    // e += [l_xx | l_x] of keypoint step index s                                         System.cpp:252-260,290-296
    auto add_kp = [&](double (&e)[8], int s) {
        const double* lx = W.LX + ((size_t)b * S_kp + s) * nf + off;
        const double* lxx = W.LXX + ((size_t)b * S_kp + s) * nf * nf + (size_t)off * nf + off;
#pragma unroll
        for (int i = 0; i < N; i++) e[i] += (j < N) ? lxx[i * nf + j] : (j == N ? lx[i] : 0.0);
    };

    // terminal: P = cost_F_xx, p = cost_F_x   (ILQRRecursive.cpp:69-71)
    double Pc[8];
#pragma unroll
    for (int i = 0; i < 8; i++) Pc[i] = 0.0;
    int sp = S_kp - 1;
    int next_kp = sp >= 0 ? P.kp_steps[sp] : -1;
    if (next_kp == H - 1) {
        add_kp(Pc, sp);
        sp--;
        next_kp = sp >= 0 ? P.kp_steps[sp] : -1;
    }
    if (l0 >= 0) add_limits(Pc, j < N ? Xb[(size_t)(H - 1) * ns + j] : 0.0);
    st8(sP + j * BwdColSmem::PCS, Pc);
    auto load_u = [&](int k, double (&o)[8]) {  // the controls enter the vector column only
        if constexpr (!VEC) {
#pragma unroll
            for (int i = 0; i < N; i++) o[i] = (j == N) ? Ub[(size_t)k * ms + i] : 0.0;
        } else ld8(Ub + (size_t)k * ms, o);
    };
    double un[8] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0}, xn = 0.0;
    load_u(H - 2, un);
    if (l0 >= 0 && j < N) xn = Xb[(size_t)(H - 2) * ns + j];
    bool ok = true;
    double* Kg = W.K + ((size_t)b * (H - 1) + (H - 2)) * nf * ns + (size_t)off * ns + off;  // walks backwards with k
    double* dg = W.d + ((size_t)b * (H - 1) + (H - 2)) * ms + off;
    __syncwarp();

    for (int k = H - 2; k >= 0; k--, Kg -= nf * ns, dg -= ms) {
        double u[8];
#pragma unroll
        for (int i = 0; i < 8; i++) u[i] = un[i];
        const double x = xn;
        if (k > 0) {  // prefetch the next step's controls / state
            load_u(k - 1, un);
            if (l0 >= 0 && j < N) xn = Xb[(size_t)(k - 1) * ns + j];
        }
        // [Qux|Qu] = [0 | R u] + (B^T [P|p]) A_aug ;  Quu = R + (B^T P) B ;  [Qxx|Qx] = [l_xx|l_x] + (A^T [P|p]) A_aug   ILQRRecursive.cpp:82-87
        double Qa[8], L[8], Rh[8], Qx[8];
#pragma unroll
        for (int i = 0; i < N; i++) {
            const double bp = dt * Pc[i];
            Qa[i] = fma(Rv[i], u[i], bp);
            L[i] = ((bp * dt) + Rd[i]) + Ed[i];  // Quu + 1e-6 I (ILQRRecursive.cpp:89); lane 7 carries the vector column: its L is never a pivot column
            Rh[i] = -Qa[i];
            Qx[i] = Pc[i];
        }
        if (k == next_kp) {
            add_kp(Qx, sp);
            sp--;
            next_kp = sp >= 0 ? P.kp_steps[sp] : -1;
        }
        if (l0 >= 0) add_limits(Qx, x);

        // [K|d] = -(Quu + 1e-6 I)^-1 [Qux|Qu]: Gauss-Jordan without pivoting on the columns; per stage every lane takes column s
        // from its owner by shuffle (raw, so the shuffles do not wait for the reciprocal) and scales its own pivot-row entries
#pragma unroll
        for (int s = 0; s < N; s++) {
            double f[N];
#pragma unroll
            for (int i = 0; i < N; i++) f[i] = __shfl_sync(FULL, L[i], gsrc + s);
            const double inv = fast_rcp(f[s]);
            const int hi = __double2hiint(inv);  // 1 / pivot: positive and finite iff the pivot was
            ok = ok && (hi > 0) && (hi < 0x7ff00000);
            const double lk = L[s] * inv, rk = Rh[s] * inv;
#pragma unroll
            for (int i = 0; i < N; i++) {
                if (i == s) {
                    L[i] = lk;
                    Rh[i] = rk;
                } else {
                    L[i] = fma(-f[i], lk, L[i]);
                    Rh[i] = fma(-f[i], rk, Rh[i]);
                }
            }
        }
        // Rh = column j of [K|d].  Publish the rows for K^T M, stream K and d to global (time-forward index)
#pragma unroll
        for (int i = 0; i < N; i++) sK[i * 8 + j] = Rh[i];
        if (live) {
#pragma unroll
            for (int i = 0; i < N; i++)
                if (j < N) Kg[i * ns + j] = Rh[i];  // the pad column stays as the solve's memset left it
            if (j == N) {
                if constexpr (!VEC) {
#pragma unroll
                    for (int i = 0; i < N; i++) dg[i] = Rh[i];
                } else {
                    double dv[8];
#pragma unroll
                    for (int i = 0; i < 8; i++) dv[i] = i < N ? Rh[i] : 0.0;
                    st8(dg, dv);
                }
            }
        }
        __syncwarp();
        // PK = P [K|d]
        double PK[8];
#pragma unroll
        for (int i = 0; i < 8; i++) PK[i] = 0.0;
#pragma unroll
        for (int c = 0; c < N; c++) {
            double pc[8];
            ld8(sP + c * BwdColSmem::PCS, pc);
#pragma unroll
            for (int i = 0; i < N; i++) PK[i] = fma(pc[i], Rh[c], PK[i]);
        }
        // M = Quu [K|d] + [Qux|Qu], un-regularised Quu = R + dt^2 P                        ILQRRecursive.cpp:93
        double M[8];
#pragma unroll
        for (int i = 0; i < N; i++) M[i] = fma(R[i], Rh[i], (dt * PK[i]) * dt) + Qa[i];
        // [P|p] = [Qxx|Qx] + K^T M + Qxu [K|d]                                             ILQRRecursive.cpp:93-94
        double acc[8];
#pragma unroll
        for (int i = 0; i < 8; i++) acc[i] = 0.0;
#pragma unroll
        for (int c = 0; c < N; c++) {
            double kr[8];
            ld8(sK + c * 8, kr);
#pragma unroll
            for (int i = 0; i < N; i++) acc[i] = fma(kr[i], M[c], acc[i]);
        }
#pragma unroll
        for (int i = 0; i < N; i++) Pc[i] = (Qx[i] + acc[i]) + dt * PK[i];
        __syncwarp();  // every lane has read sP and sK
        st8(sP + j * BwdColSmem::PCS, Pc);
        __syncwarp();
    }
    return __ballot_sync(FULL, !ok);
}

template <int N, int WARPS>
__global__ void __maxnreg__(200) k_backward_n1_col(const DevProblem* __restrict__ Pp, Work W, int b0, int nB) {
    extern __shared__ __align__(16) double smem_col[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int bfirst = b0 + 4 * (blockIdx.x * WARPS + wid), bend = b0 + nB;
    if (bfirst >= bend) return;
    double* ws = smem_col + (size_t)wid * BwdColSmem::PER_WARP;
    int shadow = -1;
    for (int s = 3; s >= 0; s--)
        if (bfirst + s < bend && !W.done[bfirst + s]) shadow = bfirst + s;
    if (shadow < 0) return;
    const int bq = bfirst + (lane >> 3);
    const bool live = bq < bend && !W.done[bq < bend ? bq : shadow];
    const unsigned bad = backward_n1_col<N, false>(*Pp, W, live ? bq : shadow, live, 0, lane, ws);
    if (bad == 0) return;
    for (int s = 0; s < 4; s++) {  // rare: the reference's pivoted inverse for the problems that need it, a warp each
        const int b = bfirst + s;
        if (((bad >> (8 * s)) & 0xffu) == 0 || b >= bend || W.done[b]) continue;
        if (lane == 0) atomicOr(W.status + b, ILQRB_ST_QUU_NOT_SPD);
        __syncwarp();
        backward_n1_redo<1, (N == 7 ? 7 : 0)>(*Pp, W, b, lane, ws, 0, 0);
    }
}

// composite robots: one group of 8 lanes per (problem, diagonal block of 7 joints)
template <int N, int WARPS>
__global__ void __maxnreg__(200) k_backward_n1_col_blocks(const DevProblem* __restrict__ Pp, Work W, int b0, int nB, StateBlocks sb) {
    extern __shared__ __align__(16) double smem_col[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int ufirst = 4 * (blockIdx.x * WARPS + wid), uend = nB * sb.nblk;
    if (ufirst >= uend) return;
    double* ws = smem_col + (size_t)wid * BwdColSmem::PER_WARP;
    int shadow = -1;
    for (int s = 3; s >= 0; s--)
        if (ufirst + s < uend && !W.done[b0 + (ufirst + s) / sb.nblk]) shadow = ufirst + s;
    if (shadow < 0) return;
    const int uq = ufirst + (lane >> 3);
    const bool live = uq < uend && !W.done[b0 + (uq < uend ? uq : shadow) / sb.nblk];
    const int u = live ? uq : shadow;
    const int bl = u / sb.nblk, blk = u - bl * sb.nblk;
    int off = 0;
    for (int q = 0; q < 4; q++)
        if (q == blk) off = sb.off[q];
    const unsigned bad = backward_n1_col<N, true>(*Pp, W, b0 + bl, live, off, lane, ws);
    if (bad == 0) return;
    for (int s = 0; s < 4; s++) {
        const int us = ufirst + s;
        if (((bad >> (8 * s)) & 0xffu) == 0 || us >= uend) continue;
        const int b = b0 + us / sb.nblk, bk = us % sb.nblk;
        if (W.done[b]) continue;
        if (lane == 0) atomicOr(W.status + b, ILQRB_ST_QUU_NOT_SPD);
        __syncwarp();
        int o = 0;
        for (int q = 0; q < 4; q++)
            if (q == bk) o = sb.off[q];
        backward_n1_redo<1, N>(*Pp, W, b, lane, ws, o, N);
    }
}

}  // namespace ilqrb
