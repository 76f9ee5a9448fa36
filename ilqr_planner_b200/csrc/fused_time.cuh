// fused_time.cuh — fast path of the Recursive iLQR solve for first-order systems WITH the time state
// (PosOrnTimePlannerSys nb_deriv = 1: x = [q, t], u = [dq, s], dt = s^2; PosOrnTimePlannerSys.cpp:149-177):
//   A = I,   B = [ dt I_d   2 s dq ]      n = m = dof + 1 <= 8
//                [  0        2 s   ]
// Two kernels per iteration (plus k_kp_derivs for l_x / l_xx), one problem per warp:
//
//   k_backward_time   Riccati recursion (ILQRRecursive.cpp:69-97).  [P | p] lives in registers as an 8 x 16 pair of DMMA accumulator
//                     tiles (the vector rides in column 8).  B is a scaled identity plus one column, so its products are scalings
//                     and one reduction each: B^T [P|p] = dt [P|p] on the joint rows and the column sums c^T [P|p] on the time row
//                     (3 shuffle rounds), (B^T P) B and P B = a scaling of the columns and one row sum (2 shuffle rounds).  What is
//                     left per step are the dense products of the reference's update — K = -(Quu + 1e-6 I)^-1 Qux, M = Quu K + Qux,
//                     K^T M, Qxu K — 16 mma.sync m8n8k4.f64, operands re-laid-out through shared memory, the inverse by the
//                     in-register Gauss-Jordan of fused1.cuh.
//
//   k_forward_time    line search + accept (ILQRRecursive.cpp:101-176).  dt = s^2 makes the rollout bilinear in the controls, so
//                     the candidates are not affine in alpha (fused1.cuh's trick does not apply): lane c rolls candidate c out in
//                     registers (x' [8], one 8 x 8 gain product per step), all lanes reading the same K_k, d_k, X_k, U_k from a
//                     shared-memory ring the bulk-copy engine fills (cp.async.bulk + mbarrier, as in fused1_fwd.cuh); kinematics
//                     only at the keypoint steps (and where a joint limit is violated).  The reference's sequential choice, then the
//                     chosen candidate is rolled out once more — same code, same alpha, same numbers — with lane 0 writing X and U.
//                     Replaces k_linesearch + k_accept, which wrote all 11 candidate control trajectories to global memory
//                     (285 MB per iteration at B = 4096, T = 100).
#pragma once
#include "fused1_fwd.cuh"

namespace ilqrb {

struct BwdTimeSmem {
    static constexpr int LD = 20, SZ = 8 * LD;
    // [Qux|Qu], [K|d] (pivot rows of the elimination before that), (Quu + 1e-6 I)^-1 then M = Quu [K|d] + [Qux|Qu], [Quu | Qxu]
    static constexpr int PER_WARP = 4 * SZ;
};

// store an 8 x 16 accumulator pair: lane (g, t) owns (g, 8 ct + 2t .. + 1)
__device__ __forceinline__ void st_c2(double* M, const double (&e)[2][2], int g, int t) {
    constexpr int LD = BwdTimeSmem::LD;
#pragma unroll
    for (int ct = 0; ct < 2; ct++) *reinterpret_cast<double2*>(M + g * LD + 8 * ct + 2 * t) = make_double2(e[ct][0], e[ct][1]);
}

// DOF > 0: dof at compile time (n = m = DOF + 1); DOF == 0: read from the problem
// PIVOT: see backward_n1_body (fused1.cuh)
template <int DOF, bool PIVOT = false>
__device__ __forceinline__ bool backward_time_body(const DevProblem& P, const Work& W, int b, int lane, double* wsmem) {
    constexpr unsigned FULL = 0xffffffffu;
    constexpr int LD = BwdTimeSmem::LD, SZ = BwdTimeSmem::SZ;
    const int g = lane >> 2, t = lane & 3;
    const int d = DOF > 0 ? DOF : P.dof, n = d + 1, m = n, H = P.H;
    const int ns = (n + 1) & ~1, ms = ns;
    double* sQux = wsmem;
    double* sKd = sQux + SZ;
    double* sM = sKd + SZ;
    double* sQuu = sM + SZ;  // columns 0..7: Quu, columns 8..15: Qxu

    const double Rrow = (g < m) ? P.r_diag[g] : 0.0;
    const double* Xb = W.X + (size_t)b * H * ns;
    const double* Ub = W.U + (size_t)b * (H - 1) * ms;
    const int S_kp = P.n_kp_steps;
    const int j0 = 2 * t, j1 = 2 * t + 1;  // this lane's columns inside a tile

    // joint limits of the first limit-carrying task in registers (weight 0 -> +-inf); further tasks: slow loop
    int l0 = -1, l1 = -1;
    if (P.any_limits)
        for (int ti = P.n_tasks - 1; ti >= 0; ti--)
            if (P.tasks[ti].limits_set) {
                l1 = l0;
                l0 = ti;
            }
    double lmx = INFINITY, lmn = -INFINITY;
    if (l0 >= 0 && g < n && P.tasks[l0].limit_weight[g] != 0) {
        lmx = P.tasks[l0].state_max[g];
        lmn = P.tasks[l0].state_min[g];
    }
    // e += [l_xx | l_x] of keypoint step index s                                        System.cpp:252-260,290-296
    auto add_kp = [&](double (&e)[2][2], int s) {
        const double* lx = W.LX + ((size_t)b * S_kp + s) * n;
        const double* lxx = W.LXX + ((size_t)b * S_kp + s) * n * n;
        if (g < n) {
            if (j0 < n) e[0][0] += lxx[g * n + j0];
            if (j1 < n) e[0][1] += lxx[g * n + j1];
            if (t == 0) e[1][0] += lx[g];
        }
    };
    // e += [L^T L | -L^T q], L = diag(violated), q = limit - x                          System.cpp:121-142,263-269,299-305
    auto add_lq = [&](double (&e)[2][2], double q) {
        if (j0 == g) e[0][0] += 1.0 * 1.0;
        if (j1 == g) e[0][1] += 1.0 * 1.0;
        if (t == 0) e[1][0] += -1.0 * q;
    };
    auto add_limits = [&](double (&e)[2][2], double xr) {
        if (xr > lmx) add_lq(e, lmx - xr);
        else if (xr < lmn) add_lq(e, lmn - xr);
        if (l1 >= 0)
            for (int ti = l1; ti < P.n_tasks; ti++) {
                const DevTask& tk = P.tasks[ti];
                if (!tk.limits_set || g >= n) continue;
                bool act;
                const double q = limit_q(tk, g, xr, act);
                if (act) add_lq(e, q);
            }
    };

    // terminal: P = cost_F_xx, p = cost_F_x   (ILQRRecursive.cpp:69-71)
    double Pa[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
    int sp = S_kp - 1;
    int next_kp = sp >= 0 ? P.kp_steps[sp] : -1;
    if (next_kp == H - 1) {
        add_kp(Pa, sp);
        sp--;
        next_kp = sp >= 0 ? P.kp_steps[sp] : -1;
    }
    if (l0 >= 0) add_limits(Pa, g < n ? Xb[(size_t)(H - 1) * ns + g] : 0.0);

    auto load_u = [&](int k, double& uR, double& uC0, double& uC1, double& sq, double& xr) {
        const double* Uk = Ub + (size_t)k * ms;
        uR = (g < m) ? Uk[g] : 0.0;
        uC0 = (j0 < m) ? Uk[j0] : 0.0;
        uC1 = (j1 < m) ? Uk[j1] : 0.0;
        sq = Uk[d];
        xr = (l0 >= 0 && g < n) ? Xb[(size_t)k * ns + g] : 0.0;
    };
    double uRn, uC0n, uC1n, sqn, xrn;
    load_u(H - 2, uRn, uC0n, uC1n, sqn, xrn);
    bool spd = true;
    double* Kg = W.K + ((size_t)b * (H - 1) + (H - 2)) * m * ns;  // walks backwards with k
    double* dg = W.d + ((size_t)b * (H - 1) + (H - 2)) * ms;

    for (int k = H - 2; k >= 0; k--, Kg -= m * ns, dg -= ms) {
        const double uR = uRn, uC0 = uC0n, uC1 = uC1n, sq = sqn, xr = xrn;
        if (k > 0) load_u(k - 1, uRn, uC0n, uC1n, sqn, xrn);  // prefetch the next step's rows
        // dynamics of this step (PosOrnTimePlannerSys.cpp:149-176): dt = s^2, last control column c = [2 s dq ; 2 s]
        const double dt = sq * sq, s2 = 2 * sq;
        const double cR = (g < d) ? s2 * uR : (g == d ? s2 : 0.0);
        const double cC0 = (j0 < d) ? s2 * uC0 : (j0 == d ? s2 : 0.0), cC1 = (j1 < d) ? s2 * uC1 : (j1 == d ? s2 : 0.0);

        // BP = B^T [P|p]: joint rows scaled by dt, the time row = c^T [P|p]                ILQRRecursive.cpp:82-83,86
        double BP[2][2];
        {
            double w00 = cR * Pa[0][0], w01 = cR * Pa[0][1], w10 = cR * Pa[1][0];
#pragma unroll
            for (int o = 4; o < 32; o <<= 1) {
                w00 += __shfl_xor_sync(FULL, w00, o);
                w01 += __shfl_xor_sync(FULL, w01, o);
                w10 += __shfl_xor_sync(FULL, w10, o);
            }
            BP[0][0] = (g < d) ? dt * Pa[0][0] : (g == d ? w00 : 0.0);
            BP[0][1] = (g < d) ? dt * Pa[0][1] : (g == d ? w01 : 0.0);
            BP[1][0] = (g < d) ? dt * Pa[1][0] : (g == d ? w10 : 0.0);
            BP[1][1] = 0.0;
        }
        // [Qux|Qu] = [0 | R u] + BP A_aug (A = I)
        double Qa[2][2] = {{BP[0][0], BP[0][1]}, {BP[1][0], 0.0}};
        if (t == 0) Qa[1][0] = Rrow * uR + BP[1][0];
        // Quu = R + BP B ;  Qxu = (A^T P) B = P B: columns scaled by dt, the s column = row sums against c
        double Quu[2], Qxu[2], Inv[1][1][2];
        {
            double ru = BP[0][0] * cC0 + BP[0][1] * cC1, rx = Pa[0][0] * cC0 + Pa[0][1] * cC1;
            ru += __shfl_xor_sync(FULL, ru, 1);
            rx += __shfl_xor_sync(FULL, rx, 1);
            ru += __shfl_xor_sync(FULL, ru, 2);
            rx += __shfl_xor_sync(FULL, rx, 2);
            double q0 = (j0 < d) ? BP[0][0] * dt : (j0 == d ? ru : 0.0), q1 = (j1 < d) ? BP[0][1] * dt : (j1 == d ? ru : 0.0);
            if (g == j0) q0 = Rrow + q0;
            if (g == j1) q1 = Rrow + q1;
            Quu[0] = (g < m) ? q0 : 0.0;
            Quu[1] = (g < m) ? q1 : 0.0;
            Qxu[0] = (j0 < d) ? Pa[0][0] * dt : (j0 == d ? rx : 0.0);
            Qxu[1] = (j1 < d) ? Pa[0][1] * dt : (j1 == d ? rx : 0.0);
            // Quu + 1e-6 I, identity outside the leading block                            ILQRRecursive.cpp:89
            Inv[0][0][0] = (g == j0) ? ((g < m) ? Quu[0] + 1e-6 * 1.0 : 1.0) : Quu[0];
            Inv[0][0][1] = (g == j1) ? ((g < m) ? Quu[1] + 1e-6 * 1.0 : 1.0) : Quu[1];
        }
        // [Qxx|Qx] = [l_xx|l_x] + (A^T [P|p]) A_aug
        double Qx[2][2] = {{Pa[0][0], Pa[0][1]}, {Pa[1][0], 0.0}};
        if (k == next_kp) {
            add_kp(Qx, sp);
            sp--;
            next_kp = sp >= 0 ? P.kp_steps[sp] : -1;
        }
        if (l0 >= 0) add_limits(Qx, xr);

        st_c2(sQux, Qa, g, t);
        {
            const double QQ[2][2] = {{Quu[0], Quu[1]}, {Qxu[0], Qxu[1]}};
            st_c2(sQuu, QQ, g, t);
        }
        // [K|d] = -(Quu + 1e-6 I)^-1 [Qux|Qu]                                            ILQRRecursive.cpp:89-91
        bool okk = true;
        if constexpr (PIVOT) {
            okk = __all_sync(FULL, gj_inverse_regs<1, (DOF > 0 ? DOF + 1 : 0), LD>(Inv, sKd, m, g, t));
            if (!okk) {  // a pivot was not positive: Quu + 1e-6 I again, for the reference's pivoted LU inverse
                spd = false;
                Inv[0][0][0] = (g == j0) ? ((g < m) ? Quu[0] + 1e-6 * 1.0 : 1.0) : Quu[0];
                Inv[0][0][1] = (g == j1) ? ((g < m) ? Quu[1] + 1e-6 * 1.0 : 1.0) : Quu[1];
            }
        } else {
            spd = gj_inverse_regs<1, (DOF > 0 ? DOF + 1 : 0), LD>(Inv, sKd, m, g, t) && spd;
        }
        *reinterpret_cast<double2*>(sM + g * LD + 2 * t) = make_double2(Inv[0][0][0], Inv[0][0][1]);
        if (PIVOT && !okk) wm_inverse_pivoted(sM, LD, m, sKd, lane);
        __syncwarp();
        double Kd[2][2];
#pragma unroll
        for (int ct = 0; ct < 2; ct++) {
            double c0 = 0.0, c1 = 0.0;
#pragma unroll
            for (int s = 0; s < 2; s++) dmma884(c0, c1, ld_a<LD>(sM, 0, 0, s, g, t), ld_b<LD>(sQux, 0, s, ct, g, t));
            Kd[ct][0] = -c0;
            Kd[ct][1] = -c1;
        }
        st_c2(sKd, Kd, g, t);
        // stream K, d to global (time-forward index); pad columns of K carry zeros
        if (g < m) {
            if (j0 < ns) *reinterpret_cast<double2*>(Kg + g * ns + j0) = make_double2(Kd[0][0], Kd[0][1]);
            if (t == 0) dg[g] = Kd[1][0];
        }
        __syncwarp();
        // [P|p] = [Qxx|Qx] + K^T (Quu [K|d] + [Qux|Qu]) + Qxu [K|d], the bracket with the un-regularised Quu (see fused1.cuh on
        // why the reference's form is kept rather than its collapsed equivalent)       ILQRRecursive.cpp:93-94
        double kb[2][2], mb[2][2];
#pragma unroll
        for (int s = 0; s < 2; s++)
#pragma unroll
            for (int ct = 0; ct < 2; ct++) kb[s][ct] = ld_b<LD>(sKd, 0, s, ct, g, t);  // B fragment of [K|d]; kb[s][0] is also the A fragment of K^T
        {
            double Mm[2][2];
#pragma unroll
            for (int ct = 0; ct < 2; ct++) {
                double c0 = Qa[ct][0], c1 = Qa[ct][1];
#pragma unroll
                for (int s = 0; s < 2; s++) dmma884(c0, c1, ld_a<LD>(sQuu, 0, 0, s, g, t), kb[s][ct]);
                Mm[ct][0] = c0;
                Mm[ct][1] = c1;
            }
            __syncwarp();  // every lane has read the inverse in sM
            st_c2(sM, Mm, g, t);
            __syncwarp();
        }
#pragma unroll
        for (int s = 0; s < 2; s++)
#pragma unroll
            for (int ct = 0; ct < 2; ct++) mb[s][ct] = ld_b<LD>(sM, 0, s, ct, g, t);
#pragma unroll
        for (int ct = 0; ct < 2; ct++) {
            double c0 = Qx[ct][0], c1 = Qx[ct][1];
#pragma unroll
            for (int s = 0; s < 2; s++) dmma884(c0, c1, kb[s][0], mb[s][ct]);
#pragma unroll
            for (int s = 0; s < 2; s++) dmma884(c0, c1, ld_a<LD>(sQuu, 0, 1, s, g, t), kb[s][ct]);
            const bool v0 = g < n && (ct == 0 ? j0 < n : t == 0), v1 = g < n && ct == 0 && j1 < n;
            Pa[ct][0] = v0 ? c0 : 0.0;
            Pa[ct][1] = v1 ? c1 : 0.0;
        }
        __syncwarp();  // shared operands are rewritten by the next step
    }
    return __all_sync(FULL, spd);
}
template <int DOF>
__device__ __noinline__ void backward_time_redo(const DevProblem& P, const Work& W, int b, int lane, double* wsmem) {
    backward_time_body<DOF, true>(P, W, b, lane, wsmem);
}

template <int DOF, int WARPS, int MINB>
__global__ void __launch_bounds__(32 * WARPS, MINB) k_backward_time(const DevProblem* __restrict__ Pp, Work W, int b0, int nB) {
    extern __shared__ double smem[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int bl = blockIdx.x * WARPS + wid;
    if (bl >= nB) return;
    const int b = b0 + bl;
    if (W.done[b]) return;
    double* ws = smem + (size_t)wid * BwdTimeSmem::PER_WARP;
    if (!backward_time_body<DOF, false>(*Pp, W, b, lane, ws)) {
        if (lane == 0) atomicOr(W.status + b, ILQRB_ST_QUU_NOT_SPD);
        backward_time_redo<DOF>(*Pp, W, b, lane, ws);
    }
}

// ------------------------------------------------------------------------------------------------------------------------------
// forward pass.  Ring stage = SPS steps of K [m][ns], d [ms], X [ns], U [ms] (doubles)
template <int SPS>
__host__ __device__ inline int fwd_time_stage_doubles(int m, int ns) { return SPS * (m * ns + 3 * ns); }

// System::cost at one step, out of line: the kinematics want ~100 registers and index their arguments at run time
static __device__ __noinline__ double time_stage_cost(const DevProblem& P, int b, const double* x, const double* u, int k) { return lean_stage_cost<8>(P, b, x, u, k); }

// per-warp shared memory of k_forward_time: the ring, its barriers, (x, u) of every candidate at the keypoint steps and at the final
// state ([S + 1][NCAND][16] doubles), and the 8 x 16 deviation matrix of the current step (row stride 20) + 16 limit-cost totals
constexpr int FWD_TIME_DX_DOUBLES = 8 * 20 + 16;
template <int RING, int SPS>
__host__ __device__ inline size_t fwd_time_warp_bytes(int m, int ns, int S) {
    return ((size_t)RING * fwd_time_stage_doubles<SPS>(m, ns) * 8 + RING * 8 + (size_t)(S + 1) * NCAND * 16 * 8 + FWD_TIME_DX_DOUBLES * 8 + 127) / 128 * 128;
}

// the bulk-copy ring of one warp: stage q = steps [q SPS, (q+1) SPS) of K, d, X, U; both passes of the forward kernel share it
template <int RING, int SPS>
struct TimeRing {
    double* ring;
    unsigned long long* bars;
    const double *Kb, *db, *Xb, *Ub;
    int m, ns, H, n_stages, stage_d;
    int issued, waited;  // stages issued / consumed since the barriers were initialised (issued: lane 0's copy is the truth)
    __device__ __forceinline__ void issue(int q) {  // lane 0 only
        const int slot = issued % RING, k0 = q * SPS;
        const int steps = (H - 1 - k0) < SPS ? (H - 1 - k0) : SPS;
        double* sK = ring + (size_t)slot * stage_d;
        double* sd = sK + SPS * m * ns;
        double* sx = sd + SPS * ns;
        double* su = sx + SPS * ns;
        issued++;
        if (steps <= 0) {  // the stage of the final cost alone: the phase completes with this arrival
            mbar_expect_tx(&bars[slot], 0u);
            return;
        }
        const unsigned bK = (unsigned)(steps * m * ns * 8), bv = (unsigned)(steps * ns * 8);
        mbar_expect_tx(&bars[slot], bK + 3 * bv);
        bulk_g2s(sK, Kb + (size_t)k0 * m * ns, bK, &bars[slot]);
        bulk_g2s(sd, db + (size_t)k0 * ns, bv, &bars[slot]);
        bulk_g2s(sx, Xb + (size_t)k0 * ns, bv, &bars[slot]);
        bulk_g2s(su, Ub + (size_t)k0 * ns, bv, &bars[slot]);
    }
    __device__ __forceinline__ const double* wait_next() {  // all lanes: the next stage in order
        const int slot = waited % RING;
        mbar_wait(&bars[slot], (unsigned)((waited / RING) & 1));
        waited++;
        return ring + (size_t)slot * stage_d;
    }
};

// One closed-loop rollout per lane with that lane's step length (ILQRRecursive.cpp:124-154).
//   MODE 0  candidates, fast: nothing is called inside the step loop (x, u, dx live in registers); (x, u) at the keypoint steps and the
//           final state go to `slots`, the costs are evaluated after the loop in step order.  Exact as long as no candidate violates a
//           joint limit anywhere (then every other step adds an exact zero to the reference's running sum); `viol` reports otherwise.
//   MODE 1  candidates, exact: System::cost inside the loop at every step that needs it, in the reference's order (slow; only for
//           problems whose fast pass reported a violation, or with several limit-carrying sub-systems).
//   MODE 2  accept: no cost; the lane stores its controls and states to Us / Xs.
template <int DOF, int RING, int SPS, int MODE>
__device__ __forceinline__ void time_rollout(const DevProblem& P, int b, int lane, TimeRing<RING, SPS>& R, const double alpha, const bool active, const double* s_lim,
                                             const bool slow_limits, double* slots, double* Us, double* Xs, double& cost_out, double& dun_out, bool& viol_out) {
    constexpr unsigned FULL = 0xffffffffu;
    constexpr int N = 8;
    const int d = DOF > 0 ? DOF : P.dof, n = d + 1, m = n, H = P.H, S = P.n_kp_steps;
    const int ns = (n + 1) & ~1, ms = ns;
    double x[N], u[N];
    load_x0<N>(P, b, x);
#pragma unroll
    for (int i = 0; i < N; i++) u[i] = 0.0;
    double cost = 0.0, dun = 0.0;
    bool viol = false;
    int skp = 0;
    int next_kp = S > 0 ? P.kp_steps[0] : -1;
    if (lane == 0)
        for (int q = 0; q < RING && q < R.n_stages; q++) R.issue(q);
    const double* sK = R.ring;
#pragma unroll 1
    for (int k = 0; k < H; k++) {
        const int j = k % SPS;
        if (j == 0) sK = R.wait_next();  // first step of a stage: its copies have landed
        const bool last = k == H - 1;
        if (!last && active) {
            const double* Kk = sK + (size_t)j * m * ns;
            const double* sd = sK + SPS * m * ns;
            const double* sx = sd + SPS * ms;
            const double* su = sx + SPS * ns;
            double dx[N];
#pragma unroll
            for (int c = 0; c < N; c++) dx[c] = (c < n) ? __dsub_rn(x[c], sx[j * ns + c]) : 0.0;
            double nrm2 = 0.0;
#pragma unroll
            for (int i = 0; i < N; i++) {
                if (i < m) {
                    double acc = 0.0;
#pragma unroll
                    for (int c = 0; c < N; c += 2) {  // ns is even and the pad column of K holds zeros
                        if (c < ns) {
                            const double2 kk = *reinterpret_cast<const double2*>(Kk + i * ns + c);
                            acc += kk.x * dx[c];
                            acc += kk.y * dx[c + 1];
                        }
                    }
                    const double du = acc + alpha * sd[j * ms + i];
                    nrm2 += du * du;
                    u[i] = su[j * ms + i] + du;
                }
            }
            dun += sqrt(nrm2);
        }
        if (MODE != 2 && active) {
            bool out = false;  // this state is beyond a joint limit
            if (s_lim) {
#pragma unroll
                for (int c = 0; c < N; c++)
                    if (c < n) out = out || (x[c] > s_lim[c]) || (x[c] < s_lim[8 + c]);
            }
            if (MODE == 0) {
                viol = viol || out;
                if (k == next_kp || last) {  // (x, u) of this candidate at a step that costs something
                    double* sl = slots + ((size_t)(last ? S : skp) * NCAND + lane) * 16;
#pragma unroll
                    for (int c = 0; c < N; c += 2) {
                        *reinterpret_cast<double2*>(sl + c) = make_double2(x[c], x[c + 1]);
                        *reinterpret_cast<double2*>(sl + 8 + c) = make_double2(u[c], u[c + 1]);
                    }
                }
            } else if (k == next_kp || last || out || slow_limits) {
                double xc[N], uc[N];  // the kinematics index their arguments at run time: they get copies
#pragma unroll
                for (int c = 0; c < N; c++) {
                    xc[c] = x[c];
                    uc[c] = u[c];
                }
                cost += time_stage_cost(P, b, xc, last ? nullptr : uc, k);
            }
        }
        if (!last && active) {
            if (MODE == 2) {
#pragma unroll
                for (int i = 0; i < N; i += 2)
                    if (i < ms) *reinterpret_cast<double2*>(Us + (size_t)k * ms + i) = make_double2(u[i], u[i + 1]);
            }
            // one integrator step with the reference's operation order and no FMA contraction (step_state, work.cuh):
            // dt = s^2 ; q += dt dq + dt^2/2 * 0 ; t += dt                               SimulationInterface.cpp:19-31
            double sq = 0.0;
#pragma unroll
            for (int c = 0; c < N; c++)
                if (c == d) sq = u[c];
            const double dts = __dmul_rn(sq, sq), hdt2 = __dmul_rn(dts, dts) / 2;
#pragma unroll
            for (int c = 0; c < N; c++) {
                if (c < d) x[c] = __dadd_rn(x[c], __dadd_rn(__dmul_rn(dts, u[c]), __dmul_rn(hdt2, 0.0)));
                else if (c == d) x[c] = __dadd_rn(x[c], dts);
            }
            if (MODE == 2) {
#pragma unroll
                for (int c = 0; c < N; c += 2)
                    if (c < ns) *reinterpret_cast<double2*>(Xs + (size_t)k * ns + c) = make_double2(x[c], x[c + 1]);
            }
        }
        if (k == next_kp) {
            skp++;
            next_kp = skp < S ? P.kp_steps[skp] : -1;
        }
        if (j == SPS - 1 || last) {  // last step of a stage: every lane is done with the slot before the copy engine refills it
            __syncwarp();
            const int q = k / SPS;
            if (lane == 0 && q + RING < R.n_stages) R.issue(q + RING);
        }
    }
    if (MODE == 0) {
        // costs of the stored states, in step order (a keypoint on the last step is part of the final cost, not a slot of its own)
        __syncwarp();
        const int nslots = (S > 0 && P.kp_steps[S - 1] == H - 1) ? S - 1 : S;
        for (int q = 0; q <= nslots; q++) {
            const bool fin = q == nslots;
            const double* sl = slots + ((size_t)(fin ? S : q) * NCAND + lane) * 16;
            if (active) cost += time_stage_cost(P, b, sl, fin ? nullptr : sl + 8, fin ? H - 1 : P.kp_steps[q]);
        }
    }
    cost_out = cost;
    dun_out = dun;
    viol_out = viol;
}

// The rollouts of ALL candidates as one matrix product per step on the FP64 tensor pipe: column c of the 8 x 16 deviation matrix DX is
// candidate c's x' - x (columns >= NCAND ride along with alpha = 0), and  K_k DX  = 4 mma.sync m8n8k4.f64 replaces 11 x 64 FMAs.
// Lane (g, t) owns row g of candidates 2t, 2t+1, 8+2t, 8+2t+1 (the accumulator layout): their u'_g, x'_g, the limits of state g.
//   dt of a candidate = (its u'_d)^2 comes from the lane holding row d of the same columns (one shuffle per column);
//   the next step's B fragments (DX re-laid-out, contraction index 2t+s so that both A fragments of K_k are one 16-byte load) go
//   through an 8 x 20 shared-memory matrix.
// MODE 0: candidates (alpha per column); (x, u) at the keypoint steps and at the end go to `slots`, joint-limit penalties of the other
//         steps are accumulated per lane; the costs are completed after the loop, lane c = candidate c (kinematics out of line).
//         Every candidate's controls go to the scratch Us = [H-1][NCAND][ms] (W.Uc), from which the accept step takes the chosen ones.
//         The sum runs over the same terms as the reference's but not in step order when a limit is violated (then it differs in
//         the last bits; without violations every skipped term is an exact zero).
// MODE 2: accept: every column carries the chosen alpha (the same arithmetic as that candidate's column in MODE 0, bit for bit);
//         the lanes of column 0 store U' and X', and sum |du| for the early-stop test.
template <int DOF, int RING, int SPS, int MODE>
__device__ __forceinline__ void time_rollout_mma(const DevProblem& P, int b, int lane, TimeRing<RING, SPS>& R, const double alpha_sel, const int line_search,
                                                 const double* s_lim, double* slots, double* sDX, double* Us, double* Xs, double& cost_out, double& dun_out) {
    constexpr unsigned FULL = 0xffffffffu;
    constexpr int LDX = 20;
    const int g = lane >> 2, t = lane & 3;
    const int d = DOF > 0 ? DOF : P.dof, n = d + 1, m = n, H = P.H, S = P.n_kp_steps;
    const int ns = (n + 1) & ~1, ms = ns;
    double* sLim = sDX + 8 * LDX;  // [16] limit-cost totals per column
    // this lane's four columns
    int col[4];
    double al[4], xs[4], u[4], limc[4];
    const double x0g = (g < d) ? P.q0[(size_t)b * d + g] : 0.0;  // x0 = [q0, 0]   PosOrnTimePlannerSys.cpp:55-62
#pragma unroll
    for (int e = 0; e < 4; e++) {
        col[e] = 8 * (e >> 1) + 2 * t + (e & 1);
        al[e] = MODE == 2 ? alpha_sel : ((col[e] < NCAND && (line_search || col[e] == 0)) ? alpha_of(col[e]) : 0.0);
        xs[e] = x0g;
        u[e] = 0.0;
        limc[e] = 0.0;
    }
    const double lmx = (s_lim && g < n) ? s_lim[g] : INFINITY, lmn = (s_lim && g < n) ? s_lim[8 + g] : -INFINITY;
    const bool vrow = g < m, vcolk = 2 * t < ns;
    double* up[4];  // where this lane's control of column e goes at the current step
#pragma unroll
    for (int e = 0; e < 4; e++) up[e] = Us + (size_t)col[e] * ms + g;
    double dun = 0.0;
    int skp = 0;
    int next_kp = S > 0 ? P.kp_steps[0] : -1;
    if (lane == 0)
        for (int q = 0; q < RING && q < R.n_stages; q++) R.issue(q);
    const double* sK = R.ring;
#pragma unroll 1
    for (int k = 0; k < H; k++) {
        const int j = k % SPS;
        if (j == 0) sK = R.wait_next();  // first step of a stage: its copies have landed
        const bool last = k == H - 1;
        const double* sd = sK + SPS * m * ns;
        const double* sx = sd + SPS * ms;
        const double* su = sx + SPS * ns;
        if (!last) {
            // DX of this step, re-laid-out for the B operand
            const double xk = (g < n) ? sx[j * ns + g] : 0.0;
            double dx[4];
#pragma unroll
            for (int e = 0; e < 4; e++) dx[e] = (g < n) ? __dsub_rn(xs[e], xk) : 0.0;
            *reinterpret_cast<double2*>(sDX + g * LDX + 2 * t) = make_double2(dx[0], dx[1]);
            *reinterpret_cast<double2*>(sDX + g * LDX + 8 + 2 * t) = make_double2(dx[2], dx[3]);
            __syncwarp();
            double2 a = make_double2(0.0, 0.0);
            if (vrow && vcolk) a = *reinterpret_cast<const double2*>(sK + (size_t)j * m * ns + g * ns + 2 * t);  // K_k[g][2t], [2t+1]
            double du[4];
#pragma unroll
            for (int ct = 0; ct < 2; ct++) {
                double c0 = 0.0, c1 = 0.0;
                dmma884(c0, c1, a.x, sDX[(2 * t) * LDX + 8 * ct + g]);      // contraction index 2t + 0
                dmma884(c0, c1, a.y, sDX[(2 * t + 1) * LDX + 8 * ct + g]);  // contraction index 2t + 1
                du[2 * ct] = c0;
                du[2 * ct + 1] = c1;
            }
            __syncwarp();  // sDX is rewritten by the next step
            const double dk = vrow ? sd[j * ms + g] : 0.0, uk = vrow ? su[j * ms + g] : 0.0;
#pragma unroll
            for (int e = 0; e < 4; e++) {
                du[e] = du[e] + al[e] * dk;  // du = K dx + alpha d ; u' = u + du        ILQRRecursive.cpp:128-130
                u[e] = vrow ? uk + du[e] : 0.0;
            }
            if (MODE == 2) {  // |du| of column 0 (rows over g)
                double n2 = du[0] * du[0];
                n2 += __shfl_xor_sync(FULL, n2, 4);
                n2 += __shfl_xor_sync(FULL, n2, 8);
                n2 += __shfl_xor_sync(FULL, n2, 16);
                dun += sqrt(n2);
            }
        }
        if (MODE == 0) {
            if (k == next_kp || last) {  // (x, u) of the candidates at a step whose cost needs the kinematics
                const int q = last ? S : skp;
#pragma unroll
                for (int e = 0; e < 4; e++)
                    if (col[e] < NCAND) {
                        double* sl = slots + ((size_t)q * NCAND + col[e]) * 16;
                        sl[g] = xs[e];
                        sl[8 + g] = last ? 0.0 : u[e];
                    }
            } else {  // System::inspectJointLimit terms of a step without a keypoint (System.cpp:121-142, 226-232)
                bool anyv = false;
#pragma unroll
                for (int e = 0; e < 4; e++) anyv = anyv || (xs[e] > lmx) || (xs[e] < lmn);
                if (anyv) {  // rare: one branch per step instead of one per column
#pragma unroll
                    for (int e = 0; e < 4; e++) {
                        if (xs[e] > lmx) {
                            const double q = lmx - xs[e];
                            limc[e] += (q * 1.0) * q;
                        } else if (xs[e] < lmn) {
                            const double q = lmn - xs[e];
                            limc[e] += (q * 1.0) * q;
                        }
                    }
                }
            }
        }
        if (!last) {
            if (MODE == 2 && t == 0 && g < ms) Us[(size_t)k * ms + g] = u[0];
            if (MODE == 0 && g < ms) {  // the controls of every candidate: the accept step reads the chosen one's back (L2)
#pragma unroll
                for (int e = 0; e < 4; e++)
                    if (col[e] < NCAND) *up[e] = u[e];
            }
            if (MODE == 0) {
#pragma unroll
                for (int e = 0; e < 4; e++) up[e] += NCAND * ms;
            }
            // one integrator step with the reference's operation order and no FMA contraction (step_state, work.cuh):
            // dt = s^2 ; q += dt dq + dt^2/2 * 0 ; t += dt                               SimulationInterface.cpp:19-31
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const double sq = __shfl_sync(FULL, u[e], 4 * d + t);  // row d of the same columns
                const double dts = __dmul_rn(sq, sq), hdt2 = __dmul_rn(dts, dts) / 2;
                if (g < d) xs[e] = __dadd_rn(xs[e], __dadd_rn(__dmul_rn(dts, u[e]), __dmul_rn(hdt2, 0.0)));
                else if (g == d) xs[e] = __dadd_rn(xs[e], dts);
            }
            if (MODE == 2 && t == 0 && g < ns) Xs[(size_t)k * ns + g] = (g < n) ? xs[0] : 0.0;
        }
        if (k == next_kp) {
            skp++;
            next_kp = skp < S ? P.kp_steps[skp] : -1;
        }
        if (j == SPS - 1 || last) {  // last step of a stage: every lane is done with the slot before the copy engine refills it
            __syncwarp();
            const int q = k / SPS;
            if (lane == 0 && q + RING < R.n_stages) R.issue(q + RING);
        }
    }
    double cost = 0.0;
    if (MODE == 0) {
        // limit penalties of each column over its rows, then lane c = candidate c completes its cost with the stored states
#pragma unroll
        for (int e = 0; e < 4; e++) {
            limc[e] += __shfl_xor_sync(FULL, limc[e], 4);
            limc[e] += __shfl_xor_sync(FULL, limc[e], 8);
            limc[e] += __shfl_xor_sync(FULL, limc[e], 16);
            if (g == 0) sLim[col[e]] = limc[e];
        }
        __syncwarp();
        const bool active = lane < NCAND && (line_search || lane == 0);
        const int nslots = (S > 0 && P.kp_steps[S - 1] == H - 1) ? S - 1 : S;
        for (int q = 0; q <= nslots; q++) {  // in step order (a keypoint on the last step is part of the final cost)
            const bool fin = q == nslots;
            const double* sl = slots + ((size_t)(fin ? S : q) * NCAND + (lane < NCAND ? lane : 0)) * 16;
            if (active) cost += time_stage_cost(P, b, sl, fin ? nullptr : sl + 8, fin ? H - 1 : P.kp_steps[q]);
        }
        if (active) cost += sLim[lane];
    }
    cost_out = cost;
    dun_out = dun;
}

template <int DOF, int RING, int SPS>
__device__ __forceinline__ void forward_time_body(const DevProblem& P, const Work& W, int b, int lane, double* ring, unsigned long long* bars, double* slots,
                                                  double* sDX, const double* s_lim, const bool slow_limits, int it, int nb_iter, int line_search, int early_stop) {
    constexpr unsigned FULL = 0xffffffffu;
    const int d = DOF > 0 ? DOF : P.dof, n = d + 1, m = n, H = P.H;
    const int ns = (n + 1) & ~1, ms = ns;
    double* Xb = W.X + (size_t)b * H * ns;
    double* Ub = W.U + (size_t)b * (H - 1) * ms;
    // this problem's slice of the candidate-control scratch: the controls of all candidates [H-1][NCAND][ms] (tensor-pipe pass), or, on
    // the scalar pass, the accepted trajectory staged as U' [H-1][ms], X'_1.. [H-1][ns] (written in place it would race with the bulk
    // copies of the steps ahead, which must still see the old X and U)
    double* Us = W.Uc + (size_t)b * (H - 1) * NCAND * ms;
    double* Xs = Us + (size_t)(H - 1) * ms;

    TimeRing<RING, SPS> R;
    R.ring = ring;
    R.bars = bars;
    R.Kb = W.K + (size_t)b * (H - 1) * m * ns;
    R.db = W.d + (size_t)b * (H - 1) * ms;
    R.Xb = Xb;
    R.Ub = Ub;
    R.m = m, R.ns = ns, R.H = H;
    R.n_stages = (H + SPS - 1) / SPS;  // steps 0 .. H-1: step H-1 is the final cost only (no control, no copies)
    R.stage_d = fwd_time_stage_doubles<SPS>(m, ns);
    R.issued = R.waited = 0;
    if (lane == 0) {
        for (int s = 0; s < RING; s++) mbar_init(&bars[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();

    // ---- candidates: column c of the rollout = step length 2^-c ----------------------------------------------------------------
    const bool cand = lane < NCAND && (line_search || lane == 0);
    double cost, dun;
    if (!slow_limits) {
        time_rollout_mma<DOF, RING, SPS, 0>(P, b, lane, R, 0.0, line_search, s_lim, slots, sDX, Us, Xs, cost, dun);
    } else {  // several limit-carrying sub-systems: the scalar pass, System::cost at every step
        bool viol;
        time_rollout<DOF, RING, SPS, 1>(P, b, lane, R, alpha_of(lane < NCAND ? lane : 0), cand, s_lim, true, slots, Us, Xs, cost, dun, viol);
    }
    // the reference's do/while (ILQRRecursive.cpp:107-155): first alpha in 1, 1/2, .. whose cost is not (>= cost0 or NaN);
    // alpha = 2^-10 is taken unconditionally (Q2)
    const double cost0 = W.cost0[b];
    int sel = 0;
    if (line_search) {
        const unsigned ok = __ballot_sync(FULL, cand && !((cost >= cost0) || isnan(cost)));
        sel = ok ? __ffs(ok) - 1 : NCAND - 1;
    }
    const double newCost = __shfl_sync(FULL, cost, sel), alpha = alpha_of(sel);
    // ---- accept ---------------------------------------------------------------------------------------------------------------
    double dun_sel;
    if (!slow_limits) {
        // the chosen candidate's controls come back from the scratch; sum |u' - u| for the early-stop test (steps over lanes), then X is
        // re-integrated from them with the operations of the rollout, in order — the trajectory the candidate's cost was computed on
        const double* Usel = Us + (size_t)sel * ms;  // row k at + k NCAND ms
        double ds = 0.0;
        for (int k = lane; k < H - 1; k += 32) {
            double n2 = 0.0;
            for (int i = 0; i < m; i++) {
                const double dd = Usel[(size_t)k * NCAND * ms + i] - Ub[(size_t)k * ms + i];
                n2 += dd * dd;
            }
            ds += sqrt(n2);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ds += __shfl_xor_sync(FULL, ds, o);
        dun_sel = ds;
        __syncwarp();
        double xg = (lane < d) ? P.q0[(size_t)b * d + lane] : 0.0;  // lane g < n owns state g
        constexpr int PFK = 8;
        for (int k0 = 0; k0 < H - 1; k0 += PFK) {
            double ub[PFK];
#pragma unroll
            for (int q = 0; q < PFK; q++) ub[q] = (k0 + q < H - 1 && lane < ms) ? Usel[(size_t)(k0 + q) * NCAND * ms + lane] : 0.0;
#pragma unroll
            for (int q = 0; q < PFK; q++) {
                const int k = k0 + q;
                if (k < H - 1) {  // warp-uniform
                    const double sq = __shfl_sync(FULL, ub[q], d);
                    const double dts = __dmul_rn(sq, sq), hdt2 = __dmul_rn(dts, dts) / 2;
                    if (lane < d) xg = __dadd_rn(xg, __dadd_rn(__dmul_rn(dts, ub[q]), __dmul_rn(hdt2, 0.0)));
                    else if (lane == d) xg = __dadd_rn(xg, dts);
                    if (lane < ms) Ub[(size_t)k * ms + lane] = ub[q];
                    if (lane < ns) Xb[(size_t)(k + 1) * ns + lane] = (lane < n) ? xg : 0.0;
                }
            }
        }
    } else {
        double c2, d2;
        bool v2;
        time_rollout<DOF, RING, SPS, 2>(P, b, lane, R, alpha, lane == 0, s_lim, false, slots, Us, Xs, c2, d2, v2);
        dun_sel = __shfl_sync(FULL, d2, 0);
        __syncwarp();
        // the staged trajectory becomes the current one (X_0 is the fixed initial state)
        for (int e = lane; e < (H - 1) * ms; e += 32) Ub[e] = Us[e];
        for (int e = lane; e < (H - 1) * ns; e += 32) Xb[ns + e] = Xs[e];
    }
    if (lane == 0) {
        W.cost0[b] = newCost;
        W.cost_hist[(size_t)b * nb_iter + it] = newCost;
        W.alpha_hist[(size_t)b * nb_iter + it] = alpha;
        W.alpha_last[b] = alpha;
        W.iters[b] = it + 1;
        int st = W.status[b];
        if (isnan(newCost)) st |= ILQRB_ST_NAN_COST;
        const bool stop = early_stop && (alpha * sqrt(dun_sel) < 1e-3) && (newCost < 1e-3);
        if (stop) {
            st |= ILQRB_ST_EARLY_STOP;
            W.done[b] = 1;
        }
        W.status[b] = st;
    }
}

template <int DOF, int RING, int SPS, int WARPS, int MINB>
__global__ void __launch_bounds__(32 * WARPS, MINB) k_forward_time(const DevProblem* __restrict__ Pp, Work W, int b0, int nB, int it, int nb_iter, int line_search,
                                                                   int early_stop) {
    extern __shared__ __align__(128) unsigned char fwd_smem[];
    const DevProblem& P = *Pp;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int d = DOF > 0 ? DOF : P.dof, n = d + 1, ns = (n + 1) & ~1;
    // limits of the single limit-carrying sub-system, shared by the block: [0..7] max, [8..15] min (weight 0 -> +-inf)
    double* s_lim = reinterpret_cast<double*>(fwd_smem);
    int l0 = -1, nl = 0;
    if (P.any_limits)
        for (int ti = 0; ti < P.n_tasks; ti++)
            if (P.tasks[ti].limits_set) {
                if (l0 < 0) l0 = ti;
                nl++;
            }
    if (threadIdx.x < 16) {
        const int c = threadIdx.x & 7;
        double v = threadIdx.x < 8 ? INFINITY : -INFINITY;
        if (l0 >= 0 && c < n && P.tasks[l0].limit_weight[c] != 0) v = threadIdx.x < 8 ? P.tasks[l0].state_max[c] : P.tasks[l0].state_min[c];
        s_lim[threadIdx.x] = v;
    }
    __syncthreads();
    const int bl = blockIdx.x * WARPS + wid;
    if (bl >= nB) return;
    const int b = b0 + bl;
    if (W.done[b]) return;
    unsigned char* mine = fwd_smem + 128 + (size_t)wid * fwd_time_warp_bytes<RING, SPS>(n, ns, P.n_kp_steps);
    double* ring = reinterpret_cast<double*>(mine);
    const size_t ringb = (size_t)RING * fwd_time_stage_doubles<SPS>(n, ns) * 8;
    double* slots = reinterpret_cast<double*>(mine + ringb);
    double* sDX = slots + (size_t)(P.n_kp_steps + 1) * NCAND * 16;
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(sDX + FWD_TIME_DX_DOUBLES);
    forward_time_body<DOF, RING, SPS>(P, W, b, lane, ring, bars, slots, sDX, nl == 1 ? s_lim : nullptr, nl > 1, it, nb_iter, line_search, early_stop);
}

// ------------------------------------------------------------------------------------------------------------------------------
// l_x = sum_tasks ((-J^T) Q) diff, l_xx = sum_tasks (J^T Q) J at the keypoint steps of the current trajectory (System.cpp:252-260,
// 290-296) for the time systems of this file: one warp per problem, lane (sg, i) = (keypoint step s0 + sg, state column i).  Every
// lane runs the lean kinematics of its step and forms its own Jacobian column in closed form (the time state's column is the unit
// vector of the time row, PosOrnTimePlannerSys.cpp:104-110); the columns meet in shared memory.  Replaces k_kp_derivs (one thread
// per (problem, keypoint step): 8192 threads at B = 4096, 77 us) on this path.
template <int WARPS>
__global__ void __launch_bounds__(32 * WARPS) k_kp_derivs_time(const DevProblem* __restrict__ Pp, Work W, int b0, int nB) {
    constexpr unsigned FULL = 0xffffffffu;
    constexpr int NX = 8, RR = 7;  // state columns per keypoint step, residual rows (6 + time)
    __shared__ double s_Jall[WARPS][32 * RR];
    const DevProblem& P = *Pp;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int bl = blockIdx.x * WARPS + wid;
    if (bl >= nB) return;
    const int b = b0 + bl;
    if (W.done[b]) return;
    double* s_J = s_Jall[wid];
    const int d = P.dof, n = d + 1, H = P.H, S = P.n_kp_steps, ns = (n + 1) & ~1;
    const double* Xb = W.X + (size_t)b * H * ns;
    const int sg = lane / NX, i = lane % NX;
    for (int s0 = 0; s0 < S; s0 += 32 / NX) {
        const int s = s0 + sg;
        const bool act = (s < S) && (i < n);
        const int k = P.kp_steps[s < S ? s : 0];
        const double* xk = Xb + (size_t)k * ns;
        double lxi = 0.0, row[NX];
#pragma unroll
        for (int j = 0; j < NX; j++) row[j] = 0.0;
        for (int ti = 0; ti < P.n_tasks; ti++) {
            const DevTask& tk = P.tasks[ti];
            const int kp = task_kp_at(tk, k);
            const bool has = act && kp >= 0;
            if (!__any_sync(FULL, has)) continue;
            const int off = tk.joint_offset, r = tk.r_t;  // r = 7
            const int cj_ = i - off;
            const int cj = (cj_ >= 0 && cj_ < P.robots[tk.robot].dof && i < d) ? cj_ : -1;
            double fx[8], Jc[RR], res[RR], jq[RR];
#pragma unroll
            for (int a = 0; a < RR; a++) res[a] = 0.0;
            lean_task_eval(P, tk, [&](int j) { return xk[off + j]; }, cj, fx, Jc);
            Jc[6] = (i == d) ? 1.0 : 0.0;  // the time row of the task Jacobian
            fx[7] = xk[n - 1];
            if (kp >= 0) {
                const double* tg = tk.kp_target + ((size_t)b * tk.n_keypoints + kp) * ILQRB_KP_STRIDE;
                task_residual(P, tk, fx, tg, res, task_deadzone(tk, kp));
            }
            const double* Q = tk.kp_precision + ((size_t)(tk.precision_per_problem ? b : 0) * tk.n_keypoints + (kp >= 0 ? kp : 0)) * r * r;
            double gsum = 0.0;
#pragma unroll
            for (int c = 0; c < RR; c++) {
                double a = 0.0, na = 0.0;
                if (c < r) {
#pragma unroll
                    for (int a2 = 0; a2 < RR; a2++)
                        if (a2 < r) {
                            const double qv = Q[a2 * r + c];
                            a += Jc[a2] * qv;
                            na += (-1.0 * Jc[a2]) * qv;
                        }
                }
                jq[c] = a;
                gsum += na * res[c];
            }
#pragma unroll
            for (int a = 0; a < RR; a++) s_J[lane * RR + a] = has ? Jc[a] : 0.0;
            __syncwarp();
            if (has) {
                lxi += gsum;
#pragma unroll
                for (int j = 0; j < NX; j++) {
                    if (j < n) {
                        const double* Jj = s_J + (sg * NX + j) * RR;
                        double hh = 0.0;
#pragma unroll
                        for (int c = 0; c < RR; c++)
                            if (c < r) hh += jq[c] * Jj[c];
                        row[j] += hh;
                    }
                }
            }
            __syncwarp();
        }
        if (act) {
            W.LX[((size_t)b * S + s) * n + i] = lxi;
            double* out = W.LXX + (((size_t)b * S + s) * n + i) * n;
#pragma unroll
            for (int j = 0; j < NX; j++)
                if (j < n) out[j] = row[j];
        }
    }
}

}  // namespace ilqrb
