// fused1.cuh — fast path of the Recursive iLQR solve for single-integrator systems without a time state
// (PosOrnPlannerSys nb_deriv = 1, the POSITION system; any robot / task list):  n = m = dof,
//   A = I,  B = dt * I          (PosOrnPlannerSys.cpp:115-121)
//
// Two kernels per iteration, one problem per warp:
//
//   k_backward_n1   Riccati recursion (ILQRRecursive.cpp:69-97).  [P|p] lives in registers in the DMMA accumulator
//                   layout; the products with A and B are exact scalings (x*1 and +0 are exact in IEEE, so
//                   (B^T P) A == dt*P bit for bit), (Quu + 1e-6 I)^-1 is an in-register division-free Gauss-Jordan, and the
//                   four dense products that remain per step (K = -Inv Qux; M = Quu K + Qux, K^T M and Qxu K of the P update)
//                   run on the FP64 tensor pipe (mma.sync m8n8k4.f64), operands re-laid-out through shared memory.
//
//   k_forward_n1    line search + accept (ILQRRecursive.cpp:101-176).  With linear dynamics the candidate rollout is
//                   affine in alpha:  x'(alpha) = x + alpha*dx,  u'(alpha) = u + alpha*du  where (dx,du) is the
//                   alpha = 1 closed-loop rollout of  du_k = K_k dx_k + d_k,  dx_{k+1} = dx_k + dt du_k.  One serial
//                   sweep produces (dx,du); the 11 candidates then only differ in the cost terms (FK + log-map
//                   residual at the keypoint steps, joint-limit penalties), which the lanes evaluate in parallel:
//                   lane = (candidate, keypoint step).  Selection follows the reference's sequential rule (first
//                   alpha in 1, 1/2, .. that is not (>= cost0 or NaN); 2^-10 unconditionally, Q2), then U += alpha*du,
//                   X is re-integrated from U with the reference's operation order, and l_x / l_xx of the keypoint
//                   steps of the accepted trajectory are produced for the next backward pass.
#pragma once
#include "kin_lean.cuh"
#include "wmat.cuh"
#include "work.cuh"

namespace ilqrb {



// ---- DMMA fragment loads from a row-major shared-memory matrix with leading dimension LD -------------------
//   A (8x4 slice s of k-tile kt, row tile rt):  M[8rt+g][8kt+4s+t]
//   B (4x8 slice, column tile ct):              M[8kt+4s+t][8ct+g]
template <int LD>
__device__ __forceinline__ double ld_a(const double* M, int rt, int kt, int s, int g, int t) {
    return M[(8 * rt + g) * LD + 8 * kt + 4 * s + t];
}
template <int LD>
__device__ __forceinline__ double ld_b(const double* M, int kt, int s, int ct, int g, int t) {
    return M[(8 * kt + 4 * s + t) * LD + 8 * ct + g];
}
// store an accumulator-layout register matrix: lane (g,t) owns (8rt+g, 8ct+2t .. +1)
template <int NT, int LD>
__device__ __forceinline__ void st_c(double* M, const double (&e)[NT][NT][2], int g, int t) {
#pragma unroll
    for (int rt = 0; rt < NT; rt++)
#pragma unroll
        for (int ct = 0; ct < NT; ct++) *reinterpret_cast<double2*>(M + (8 * rt + g) * LD + 8 * ct + 2 * t) = make_double2(e[rt][ct][0], e[rt][ct][1]);
}

// 1/x to full double precision: MUFU.RCP64H seed (2^-23) + two Newton steps; no slow path (x is a pivot of an SPD
// matrix in normal operation; NaN / non-positive pivots are reported through the status word).
__device__ __forceinline__ double fast_rcp(double x) {
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double e = fma(-x, y, 1.0);
    y = fma(y, e, y);
    e = fma(-x, y, 1.0);
    y = fma(y, e, y);
    return y;
}

// exact power of two 2^(1022 - E) for x = f * 2^(E - 1023): x * pow2_unscale(x) lies in [0.5, 1)
__device__ __forceinline__ double pow2_unscale(double x) { return __hiloint2double(0x7fd00000 - (__double2hiint(x) & 0x7ff00000), 0); }

// (L)^-1 in registers, in place: division-free Gauss-Jordan on [L | I] without pivoting, (8NT)^2 in accumulator layout
// (any row scaling leaves the solution unchanged, so no reciprocal sits on the pivot-to-pivot dependency chain).
// L = Quu + 1e-6 I in its leading m x m block and identity outside; this is the explicit inverse the reference takes
// (ILQRRecursive.cpp:89).  In-place storage: after stage k, column k of L is dead and its slots hold column k of the
// right-hand block.  Stage k replaces row i by s (p_kk row_i - l_ik row_k), s an exact power of two near 1/p_kk, so every
// row carries a scale: d_i (its diagonal, tracked per lane) for pivoted rows, the warp-uniform S for the others; rows are
// divided by d_i at the end.  The shuffle/LSU data pipe bounds the backward kernels (ncu: l1tex__data_pipe_lsu_wavefronts
// 83 %), so the pivot row travels through shared memory as one predicated 16-byte store per owner lane and one broadcast
// 16-byte load per lane (1 wavefront each, against 4 for the shuffles), the pivot as a broadcast 8-byte load; only the
// pivot-column element (quad-local) is shuffled.  sRow: (8NT) x LD scratch, row k used by stage k only.
// Returns (per lane, for its own rows) false if a pivot was <= 0, infinite or NaN.
template <int NT, int M, int LD>
__device__ __forceinline__ bool gj_inverse_regs(double (&Wm)[NT][NT][2], double* sRow, int m_rt, int g, int t) {
    constexpr unsigned FULL = 0xffffffffu;
    const int m = M > 0 ? M : m_rt;
    double dg[NT], S = 1.0;
#pragma unroll
    for (int rt = 0; rt < NT; rt++) dg[rt] = 1.0;
#pragma unroll
    for (int k = 0; k < 8 * NT; k++) {
        if (k < m) {  // warp-uniform (compile-time when M > 0)
            const int kt = k >> 3, kg = k & 7, kh = k & 1;
            if (g == kg) {
#pragma unroll
                for (int ct = 0; ct < NT; ct++) *reinterpret_cast<double2*>(sRow + k * LD + 8 * ct + 2 * t) = make_double2(Wm[kt][ct][0], Wm[kt][ct][1]);
            }
            __syncwarp();
            double rk[NT][2], colk[NT];
#pragma unroll
            for (int ct = 0; ct < NT; ct++) {
                const double2 v = *reinterpret_cast<const double2*>(sRow + k * LD + 8 * ct + 2 * t);
                rk[ct][0] = v.x;
                rk[ct][1] = v.y;
            }
            const double pkk = sRow[k * LD + k];
#pragma unroll
            for (int rt = 0; rt < NT; rt++) colk[rt] = __shfl_sync(FULL, Wm[rt][kt][kh], 4 * g + (kg >> 1));
            const double sc = pow2_unscale(pkk);
            const double a = pkk * sc;  // every row, the pivot row included, is scaled by a
            const double Sn = S * a;
            const bool kcol = (t == (kg >> 1));
#pragma unroll
            for (int rt = 0; rt < NT; rt++) {
                const bool prow = (rt == kt) && (g == kg);  // this lane owns row k: scaled only
                const double c = prow ? 0.0 : colk[rt] * sc;
                dg[rt] = (prow ? pkk : dg[rt]) * a;
#pragma unroll
                for (int ct = 0; ct < NT; ct++)
#pragma unroll
                    for (int h = 0; h < 2; h++) Wm[rt][ct][h] = fma(-c, rk[ct][h], a * Wm[rt][ct][h]);
                if (kcol) Wm[rt][kt][kh] = prow ? Sn : -(c * S);  // column k of the right-hand block (row k: the scale of the unpivoted rows)
            }
            S = Sn;
        }
    }
    // every scaling factor is positive when the pivots are, and the last non-positive pivot leaves its own row's diagonal
    // non-positive (zero and NaN propagate): the pivot test of the elimination is the sign test of the final diagonals
    bool ok = true;
#pragma unroll
    for (int rt = 0; rt < NT; rt++) {
        const int hi = __double2hiint(dg[rt]);
        ok = ok && (hi > 0) && (hi < 0x7ff00000);
        const double inv = fast_rcp(dg[rt]);
#pragma unroll
        for (int ct = 0; ct < NT; ct++) {
            Wm[rt][ct][0] *= inv;
            Wm[rt][ct][1] *= inv;
        }
    }
    return ok;  // per lane: combine with a vote where it is consumed
}

// Solve (L) X = Rr in registers for the larger systems (NT = 2).  There the elimination of the augmented matrix [L | Rr]
// touches 14 elements per lane and stage and the kernel holds 2 warps per scheduler (255 registers), so it is bound by the
// latency of the pivot-to-pivot chain: one reciprocal per stage and pivot rows by shuffle (26 cycles) measured faster on
// B200 than the division-free form (a second multiply per element), than the shared-memory broadcast (store -> barrier
// -> load per stage) and than the explicit inverse (16 more DMMA): C5 backward pass 38.9 / 40.8 / 40.9 / 43.0 ms.
// Gauss-Jordan without pivoting and without row scaling (rows are divided by their diagonal once, at the end).
// Returns false if a pivot was <= 0, infinite or NaN.
template <int NT, int M>
__device__ __forceinline__ bool gj_solve_regs(double (&L)[NT][NT][2], double (&Rr)[NT][NT][2], int m_rt, int g, int t) {
    constexpr unsigned FULL = 0xffffffffu;
    const int m = M > 0 ? M : m_rt;
    bool ok = true;
#pragma unroll
    for (int k = 0; k < 8 * NT; k++) {
        if (k < m) {  // warp-uniform (compile-time when M > 0)
            const int kt = k >> 3, kg = k & 7, kh = k & 1;
            const double pkk = __shfl_sync(FULL, L[kt][kt][kh], 4 * kg + (kg >> 1));
            double rl[NT][2], rr[NT][2], colk[NT];
#pragma unroll
            for (int ct = 0; ct < NT; ct++) {
                if (ct >= kt) {  // columns left of the pivot tile are dead
                    rl[ct][0] = __shfl_sync(FULL, L[kt][ct][0], 4 * kg + t);
                    rl[ct][1] = __shfl_sync(FULL, L[kt][ct][1], 4 * kg + t);
                }
                rr[ct][0] = __shfl_sync(FULL, Rr[kt][ct][0], 4 * kg + t);
                rr[ct][1] = __shfl_sync(FULL, Rr[kt][ct][1], 4 * kg + t);
            }
#pragma unroll
            for (int rt = 0; rt < NT; rt++) colk[rt] = __shfl_sync(FULL, L[rt][kt][kh], 4 * g + (kg >> 1));
            const int hi = __double2hiint(pkk);
            ok = ok && (hi > 0) && (hi < 0x7ff00000);
            const double piv = fast_rcp(pkk);
#pragma unroll
            for (int rt = 0; rt < NT; rt++) {
                const bool prow = (rt == kt) && (g == kg);  // this lane owns row k: left as it is
                const double f = prow ? 0.0 : colk[rt] * piv;
#pragma unroll
                for (int ct = 0; ct < NT; ct++)
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        if (ct >= kt) L[rt][ct][h] = fma(-f, rl[ct][h], L[rt][ct][h]);
                        Rr[rt][ct][h] = fma(-f, rr[ct][h], Rr[rt][ct][h]);
                    }
            }
        }
    }
    // divide every row by its diagonal entry: (i, i) lives in lane (g, g>>1), slot g&1 of tile [rt][rt]
#pragma unroll
    for (int rt = 0; rt < NT; rt++) {
        const double mine = (g & 1) ? L[rt][rt][1] : L[rt][rt][0];
        const double dgi = __shfl_sync(FULL, mine, 4 * g + ((g >> 1) & 3));
        const double inv = fast_rcp(dgi);
#pragma unroll
        for (int ct = 0; ct < NT; ct++) {
            Rr[rt][ct][0] *= inv;
            Rr[rt][ct][1] *= inv;
        }
    }
    return ok;
}

template <int NT>
struct BwdN1Smem {
    static constexpr int NX = 8 * NT, LD = NX + 4, SZ = NX * LD;
    // [Qux|Qu], [K|d] (pivot rows of the elimination before that), (Quu + 1e-6 I)^-1 then M = Quu [K|d] + [Qux|Qu], Quu
    static constexpr int PER_WARP = 4 * SZ;
};

// ------------------------------------------------------------------------------------------------------------
// backward Riccati pass, A = I, B = dt I.  Augmented layout: column n of the state-indexed matrices carries the
// vector ([P|p], [Qux|Qu], [Qxx|Qx], [K|d]); rows >= n and columns > n are kept at exact zero.
// DOF > 0: n = m = DOF at compile time; DOF == 0: read from the problem.
// nsub > 0: the warp handles the diagonal block [off, off + nsub) of a problem whose tasks touch disjoint joint ranges
// (composite robots, C5): l_xx, R and the dynamics are block diagonal there, so P, Quu and K are — in the reference's dense
// arithmetic too, where the off-diagonal blocks are sums and products of exact zeros — and the pass splits into
// independent passes per block; strides stay those of the full state.
// PIVOT = false (the kernels' first pass): the unpivoted eliminations only; returns false (warp-uniform) if some pivot was not positive.
// PIVOT = true  (the out-of-line second pass of such a problem, rare): a step whose elimination meets a non-positive pivot is
// redone with the reference's partial-pivot LU inverse (wmat.cuh wm_inverse_pivoted).  The fallback code lives in the second pass
// only: inside the first it cost registers in the step loop (72-register cap: spills, 251 -> 276 us per launch at C2).
template <int NT, int DOF, bool PIVOT = false>
__device__ __forceinline__ bool backward_n1_body(const DevProblem& P, const Work& W, int b, int lane, double* wsmem, const int off = 0, const int nsub = 0) {
    using S = BwdN1Smem<NT>;
    constexpr int LD = S::LD, SZ = S::SZ;
    const int g = lane >> 2, t = lane & 3;
    const int nf = P.n;  // full state size: strides of LX, LXX, K
    const int n = DOF > 0 ? DOF : (nsub > 0 ? nsub : P.n), m = n, H = P.H;
    const int ns = (nf + 1) & ~1, ms = ns;
    const double dt = 1.0 * P.dt;  // B = Identity * dt
    double* sQux = wsmem;
    double* sKd = sQux + SZ;
    double* sM = sKd + SZ;
    double* sQuu = sM + SZ;

    double Rrow[NT];
#pragma unroll
    for (int rt = 0; rt < NT; rt++) Rrow[rt] = (8 * rt + g < m) ? P.r_diag[off + 8 * rt + g] : 0.0;

    const double* Xb = W.X + (size_t)b * H * ns + off;
    const double* Ub = W.U + (size_t)b * (H - 1) * ms + off;
    const int S_kp = P.n_kp_steps;

    // joint limits of the first limit-carrying task live in registers (weight 0 -> +-inf); further tasks: slow loop
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
        if (l0 >= 0 && i < n && P.tasks[l0].limit_weight[off + i] != 0) {
            lmx[rt] = P.tasks[l0].state_max[off + i];
            lmn[rt] = P.tasks[l0].state_min[off + i];
        }
    }

    // e += [l_xx | l_x] of keypoint step index s (k_kp_derivs / k_forward_n1 wrote them)    System.cpp:252-260,290-296
    auto add_kp = [&](double (&e)[NT][NT][2], int s) {
        const double* lx = W.LX + ((size_t)b * S_kp + s) * nf + off;
        const double* lxx = W.LXX + ((size_t)b * S_kp + s) * nf * nf + (size_t)off * nf + off;
#pragma unroll
        for (int rt = 0; rt < NT; rt++) {
            const int i = 8 * rt + g;
#pragma unroll
            for (int ct = 0; ct < NT; ct++)
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int j = 8 * ct + 2 * t + h;
                    if (i < n) {
                        if (j < n) e[rt][ct][h] += lxx[i * nf + j];
                        else if (j == n) e[rt][ct][h] += lx[i];
                    }
                }
        }
    };
    // e += [L^T L | -L^T q] with L = diag(violated), q = limit - x                      System.cpp:121-142,263-269,299-305
    auto add_lq = [&](double (&e)[NT][NT][2], int rt, double q) {
        const int i = 8 * rt + g;
#pragma unroll
        for (int ct = 0; ct < NT; ct++)
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int j = 8 * ct + 2 * t + h;
                if (j == i) e[rt][ct][h] += 1.0 * 1.0;
                else if (j == n) e[rt][ct][h] += -1.0 * q;
            }
    };
    auto add_limits = [&](double (&e)[NT][NT][2], const double (&xr)[NT]) {
#pragma unroll
        for (int rt = 0; rt < NT; rt++) {
            if (xr[rt] > lmx[rt]) add_lq(e, rt, lmx[rt] - xr[rt]);
            else if (xr[rt] < lmn[rt]) add_lq(e, rt, lmn[rt] - xr[rt]);
        }
        if (l1 >= 0) {  // more than one sub-system with limits (SequentialSystem.cpp:119-165 sums them)
            for (int ti = l1; ti < P.n_tasks; ti++) {
                const DevTask& tk = P.tasks[ti];
                if (!tk.limits_set) continue;
#pragma unroll
                for (int rt = 0; rt < NT; rt++) {
                    const int i = 8 * rt + g;
                    if (i < n) {
                        bool act;
                        const double q = limit_q(tk, off + i, xr[rt], act);
                        if (act) add_lq(e, rt, q);
                    }
                }
            }
        }
    };
    auto load_rows = [&](const double* base, int stride, int k, int lim, double (&o)[NT]) {
#pragma unroll
        for (int rt = 0; rt < NT; rt++) o[rt] = (8 * rt + g < lim) ? base[(size_t)k * stride + 8 * rt + g] : 0.0;
    };

    // terminal: P = cost_F_xx, p = cost_F_x   (ILQRRecursive.cpp:69-71)
    double Pa[NT][NT][2];
#pragma unroll
    for (int rt = 0; rt < NT; rt++)
#pragma unroll
        for (int ct = 0; ct < NT; ct++) Pa[rt][ct][0] = Pa[rt][ct][1] = 0.0;
    int sp = S_kp - 1;
    int next_kp = sp >= 0 ? P.kp_steps[sp] : -1;
    if (next_kp == H - 1) {
        add_kp(Pa, sp);
        sp--;
        next_kp = sp >= 0 ? P.kp_steps[sp] : -1;
    }
    if (l0 >= 0) {
        double xr[NT];
        load_rows(Xb, ns, H - 1, n, xr);
        add_limits(Pa, xr);
    }
    double un[NT], xn[NT];
    load_rows(Ub, ms, H - 2, m, un);
    load_rows(Xb, ns, H - 2, n, xn);
    bool spd = true;
    double* Kg = W.K + ((size_t)b * (H - 1) + (H - 2)) * nf * ns + (size_t)off * ns + off;  // walks backwards with k
    double* dg = W.d + ((size_t)b * (H - 1) + (H - 2)) * ms + off;

    for (int k = H - 2; k >= 0; k--, Kg -= nf * ns, dg -= ms) {
        double u[NT], x[NT];
#pragma unroll
        for (int rt = 0; rt < NT; rt++) {
            u[rt] = un[rt];
            x[rt] = xn[rt];
        }
        if (k > 0) {  // prefetch the next step's control / state rows
            load_rows(Ub, ms, k - 1, m, un);
            if (l0 >= 0) load_rows(Xb, ns, k - 1, n, xn);
        }
        // [Qux|Qu] = [0 | R u] + (B^T [P|p]) A_aug ;  Quu = R + (B^T P) B ;  [Qxx|Qx] = [l_xx|l_x] + (A^T [P|p]) A_aug
        // Qxu = (A^T P) B = P*dt has the same entries as Qux = dt*P                     ILQRRecursive.cpp:82-87
        double Qa[NT][NT][2], Qx[NT][NT][2], Inv[NT][NT][2], Quu[NT][NT][2];
#pragma unroll
        for (int rt = 0; rt < NT; rt++) {
            const int i = 8 * rt + g;
#pragma unroll
            for (int ct = 0; ct < NT; ct++)
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int j = 8 * ct + 2 * t + h;
                    const double bp = dt * Pa[rt][ct][h];
                    Qa[rt][ct][h] = (j == n) ? Rrow[rt] * u[rt] + bp : bp;
                    double quu = bp * dt;
                    if (i == j) quu = Rrow[rt] + quu;
                    quu = (j < m) ? quu : 0.0;  // rows >= m of [P|p] are zero already
                    // Quu + 1e-6 I, identity outside the leading block                    ILQRRecursive.cpp:89
                    Inv[rt][ct][h] = (i == j) ? ((i < m) ? quu + 1e-6 * 1.0 : 1.0) : quu;
                    Quu[rt][ct][h] = quu;  // the un-regularised Quu of the P update below
                    Qx[rt][ct][h] = Pa[rt][ct][h];
                }
        }
        if (k == next_kp) {
            add_kp(Qx, sp);
            sp--;
            next_kp = sp >= 0 ? P.kp_steps[sp] : -1;
        }
        if (l0 >= 0) add_limits(Qx, x);

        st_c<NT, LD>(sQux, Qa, g, t);
        st_c<NT, LD>(sQuu, Quu, g, t);
        // [K|d] = -(Quu + 1e-6 I)^-1 [Qux|Qu]                                            ILQRRecursive.cpp:89-91
        double Kd[NT][NT][2];
        if constexpr (NT == 1) {  // explicit inverse, then the product on DMMA
            if constexpr (PIVOT) {
                const bool okk = __all_sync(0xffffffffu, gj_inverse_regs<NT, DOF, LD>(Inv, sKd, m, g, t));
                if (!okk) {  // a pivot was not positive: Quu + 1e-6 I again, for the reference's pivoted LU inverse
                    spd = false;
                    const int j = 2 * t;
                    Inv[0][0][0] = Quu[0][0][0] + ((g == j) ? ((g < m) ? 1e-6 * 1.0 : 1.0) : 0.0);
                    Inv[0][0][1] = Quu[0][0][1] + ((g == j + 1) ? ((g < m) ? 1e-6 * 1.0 : 1.0) : 0.0);
                }
                st_c<NT, LD>(sM, Inv, g, t);
                if (!okk) wm_inverse_pivoted(sM, LD, m, sKd, lane);
            } else {
                spd = gj_inverse_regs<NT, DOF, LD>(Inv, sKd, m, g, t) && spd;
                st_c<NT, LD>(sM, Inv, g, t);
            }
            __syncwarp();
            double c0 = 0.0, c1 = 0.0;
#pragma unroll
            for (int s = 0; s < 2; s++) dmma884(c0, c1, ld_a<LD>(sM, 0, 0, s, g, t), ld_b<LD>(sQux, 0, s, 0, g, t));
            Kd[0][0][0] = -c0;
            Kd[0][0][1] = -c1;
        } else {  // elimination of the augmented matrix
#pragma unroll
            for (int rt = 0; rt < NT; rt++)
#pragma unroll
                for (int ct = 0; ct < NT; ct++) {
                    Kd[rt][ct][0] = -Qa[rt][ct][0];
                    Kd[rt][ct][1] = -Qa[rt][ct][1];
                }
            const bool oks = gj_solve_regs<NT, DOF>(Inv, Kd, m, g, t);
            if constexpr (!PIVOT) spd = oks && spd;
            if (PIVOT && !__all_sync(0xffffffffu, oks)) {
                // a pivot was not positive: the reference's pivoted LU inverse of Quu + 1e-6 I, then the product (rare, plain loops)
                spd = false;
                __syncwarp();
                for (int e = lane; e < 8 * NT * 8 * NT; e += 32) {
                    const int i = e / (8 * NT), j = e % (8 * NT);
                    double v = sQuu[i * LD + j];
                    if (i == j) v = (i < m) ? v + 1e-6 * 1.0 : 1.0;
                    sM[i * LD + j] = v;
                }
                wm_inverse_pivoted(sM, LD, m, sKd, lane);
#pragma unroll
                for (int rt = 0; rt < NT; rt++)
#pragma unroll
                    for (int ct = 0; ct < NT; ct++)
#pragma unroll
                        for (int h = 0; h < 2; h++) {
                            const int i = 8 * rt + g, j = 8 * ct + 2 * t + h;
                            double acc = 0.0;
                            for (int l = 0; l < m; l++) acc += sM[i * LD + l] * sQux[l * LD + j];
                            Kd[rt][ct][h] = (i < m) ? -acc : 0.0;
                        }
                __syncwarp();
            }
        }
        st_c<NT, LD>(sKd, Kd, g, t);
        // stream K, d to global (time-forward index)
#pragma unroll
        for (int rt = 0; rt < NT; rt++) {
            const int i = 8 * rt + g;
            if (i < m) {
#pragma unroll
                for (int ct = 0; ct < NT; ct++) {
                    const int j = 8 * ct + 2 * t;
                    if (nsub > 0) {  // a diagonal block: its own columns only (the offset may be odd: scalar stores)
                        if (j < n) Kg[i * ns + j] = Kd[rt][ct][0];
                        if (j + 1 < n) Kg[i * ns + j + 1] = Kd[rt][ct][1];
                    } else if (j < ns) *reinterpret_cast<double2*>(Kg + i * ns + j) = make_double2(Kd[rt][ct][0], j + 1 < n ? Kd[rt][ct][1] : 0.0);
                    if (j == n) dg[i] = Kd[rt][ct][0];
                    else if (j + 1 == n) dg[i] = Kd[rt][ct][1];
                }
            }
        }
        __syncwarp();

        // [P|p] = [Qxx|Qx] + K^T (Quu [K|d] + [Qux|Qu]) + Qxu [K|d]                       ILQRRecursive.cpp:93-94
        // The bracket M is formed as written, with the un-regularised Quu.  Analytically M = -1e-6 [K|d] (because
        // [K|d] = -(Quu + 1e-6 I)^-1 [Qux|Qu]), and round 1 used that identity to save two products — but the reference's form is
        // stationary in K: an error dK in the computed gains (cond(Quu) eps) enters P through (K^T Quu + Qxu) dK = M^T dK ~ 0,
        // whereas the collapsed form passes Qxu dK on at first order, step after step.  Measured against a long-double run of
        // the oracle (tests/_parity.py): gains 1.1e-9 from the truth with the collapsed form, 2e-11 with this one, which is
        // what the reference's own arithmetic gives.  (Row n = d^T [..] is produced too and masked out.)
        double kb[NT][2][NT], mb[NT][2][NT];
#pragma unroll
        for (int kt = 0; kt < NT; kt++)
#pragma unroll
            for (int s = 0; s < 2; s++)
#pragma unroll
                for (int ct = 0; ct < NT; ct++) kb[kt][s][ct] = ld_b<LD>(sKd, kt, s, ct, g, t);  // B fragment of [K|d] == A fragment of [K|d]^T
        {
            double Mm[NT][NT][2];
#pragma unroll
            for (int rt = 0; rt < NT; rt++)
#pragma unroll
                for (int ct = 0; ct < NT; ct++) {
                    double c0 = Qa[rt][ct][0], c1 = Qa[rt][ct][1];
#pragma unroll
                    for (int kt = 0; kt < NT; kt++)
#pragma unroll
                        for (int s = 0; s < 2; s++) dmma884(c0, c1, ld_a<LD>(sQuu, rt, kt, s, g, t), kb[kt][s][ct]);
                    Mm[rt][ct][0] = c0;
                    Mm[rt][ct][1] = c1;
                }
            __syncwarp();  // NT == 1: every lane has read the inverse in sM
            st_c<NT, LD>(sM, Mm, g, t);
            __syncwarp();
        }
#pragma unroll
        for (int kt = 0; kt < NT; kt++)
#pragma unroll
            for (int s = 0; s < 2; s++)
#pragma unroll
                for (int ct = 0; ct < NT; ct++) mb[kt][s][ct] = ld_b<LD>(sM, kt, s, ct, g, t);
#pragma unroll
        for (int rt = 0; rt < NT; rt++)
#pragma unroll
            for (int ct = 0; ct < NT; ct++) {
                double c0 = Qx[rt][ct][0], c1 = Qx[rt][ct][1];
#pragma unroll
                for (int kt = 0; kt < NT; kt++)
#pragma unroll
                    for (int s = 0; s < 2; s++) dmma884(c0, c1, kb[kt][s][rt], mb[kt][s][ct]);
#pragma unroll
                for (int kt = 0; kt < NT; kt++)
#pragma unroll
                    for (int s = 0; s < 2; s++) dmma884(c0, c1, ld_a<LD>(sQux, rt, kt, s, g, t), kb[kt][s][ct]);
                const int i = 8 * rt + g, j = 8 * ct + 2 * t;
                Pa[rt][ct][0] = (i < n && j <= n) ? c0 : 0.0;
                Pa[rt][ct][1] = (i < n && j + 1 <= n) ? c1 : 0.0;
            }
        __syncwarp();  // shared operands are rewritten by the next step
    }
    return __all_sync(0xffffffffu, spd);
}
// second pass of a problem whose first pass met a non-positive pivot: the same recursion with the pivoted fallback, out of line
template <int NT, int DOF>
__device__ __noinline__ void backward_n1_redo(const DevProblem& P, const Work& W, int b, int lane, double* wsmem, int off, int nsub) {
    backward_n1_body<NT, DOF, true>(P, W, b, lane, wsmem, off, nsub);
}
template <int NT, int DOF>
__device__ __forceinline__ void backward_n1_run(const DevProblem& P, const Work& W, int b, int lane, double* wsmem, const int off = 0, const int nsub = 0) {
    if (!backward_n1_body<NT, DOF, false>(P, W, b, lane, wsmem, off, nsub)) {
        if (lane == 0) atomicOr(W.status + b, ILQRB_ST_QUU_NOT_SPD);
        backward_n1_redo<NT, DOF>(P, W, b, lane, wsmem, off, nsub);
    }
}

// diagonal blocks of the state that no task couples (ilqrb_api.cu: find_state_blocks)
struct StateBlocks {
    int nblk, off[4], size[4];
};

// one warp per (problem, diagonal block): see backward_n1_body
template <int DOF, int WARPS, int MINB>
__global__ void __launch_bounds__(32 * WARPS, MINB) k_backward_n1_blocks(const DevProblem* __restrict__ Pp, Work W, int b0, int nB, StateBlocks sb) {
    extern __shared__ double smem[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int w = blockIdx.x * WARPS + wid;
    const int bl = w / sb.nblk, blk = w - bl * sb.nblk;
    if (bl >= nB) return;
    const int b = b0 + bl;
    if (W.done[b]) return;
    backward_n1_run<1, DOF>(*Pp, W, b, lane, smem + (size_t)wid * BwdN1Smem<1>::PER_WARP, sb.off[blk], sb.size[blk]);
}

template <int NT, int DOF, int WARPS, int MINB>
__global__ void __launch_bounds__(32 * WARPS, MINB) k_backward_n1(const DevProblem* __restrict__ Pp, Work W, int b0, int nB) {
    extern __shared__ double smem[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int bl = blockIdx.x * WARPS + wid;
    if (bl >= nB) return;
    const int b = b0 + bl;
    if (W.done[b]) return;
    backward_n1_run<NT, DOF>(*Pp, W, b, lane, smem + (size_t)wid * BwdN1Smem<NT>::PER_WARP);
}

// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// per-warp shared memory of k_forward_n1 (doubles): deviation controls of the whole horizon, deviation states at the
// keypoint steps, candidate x keypoint costs, one Jacobian column per lane
__host__ __device__ inline int fwd_n1_smem_doubles(int H, int ns, int S) {
    const int du = (H - 1) * ns > 32 * 7 ? (H - 1) * ns : 32 * 7;  // the Jacobian columns of phase F reuse the sDU area
    return du + S * ns + NCAND * S;
}

// forward pass: line search over all candidates, accept, derivatives of the accepted trajectory.  One warp per block.
// returns true when the early-stop rule fired for this problem
// SPLIT: phase A has run in its own kernel (k_fwd_rollout_n1, fused1_fwd.cuh) and left du, dx at the keypoint steps and the limit
// flag in global memory; this body then does phases B-F only, with `smem` = [32][6] Jacobian columns + [NCAND][S] costs.
// RR: residual rows of a task the FK phases provide for: 6 (end-effector tasks), 7 when the system has a joint-space task
// (JointSpacePlannerSys sub-systems, f(x) = x, J = I: 7 joints = 7 rows).
template <int NT, int DOF, bool SPLIT = false, int RR = 6>
__device__ __forceinline__ bool forward_n1_body(const DevProblem& P, const Work& W, int b, int lane, double* smem, int it, int nb_iter, int line_search, int early_stop,
                                                double* s_cc_split = nullptr) {
    constexpr int NX = 8 * NT;
    constexpr unsigned FULL = 0xffffffffu;
    constexpr int PF = (NT == 1) ? 4 : 3;  // register ring of gain fragments, in steps (L2 prefetch runs further ahead)
    const int g = lane >> 2, t = lane & 3;
    const int n = DOF > 0 ? DOF : P.n, m = n, dof = n, H = P.H, S = P.n_kp_steps;
    const int ns = (n + 1) & ~1, ms = ns;
    const double dt = P.dt;
    double* sDU = SPLIT ? W.DU + (size_t)b * (H - 1) * ms : smem;        // [H-1][ms]  alpha = 1 control deviation
    double* sDXk = SPLIT ? W.DXk + (size_t)b * S * ns : sDU + (H - 1) * ms;  // [S][ns]    alpha = 1 state deviation at the keypoint steps
    double* s_cc = SPLIT ? (s_cc_split ? s_cc_split : smem + 32 * 6) : sDXk + S * ns;  // [NCAND][S]
    double* s_J = SPLIT ? smem : sDU;                                      // [32][6]    phase F only (not SPLIT: when sDU is dead)
    double* Xb = W.X + (size_t)b * H * ns;
    double* Ub = W.U + (size_t)b * (H - 1) * ms;
    const double* Kb = W.K + (size_t)b * (H - 1) * m * ns;
    const double* db = W.d + (size_t)b * (H - 1) * ms;

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

    // ---- phase A: alpha = 1 deviation rollout on the FP64 tensor pipe.
    //      du = K (x' - x) + d ; x'+ = x' + dt u'     (ILQRRecursive.cpp:124-136 with alpha = 1, minus the nominal)
    //      K_k is the A operand straight from global memory (prefetched PF steps ahead), dx the B operand (all eight
    //      columns equal), so every lane of row group g ends up with du_g.  Along the way: is any joint limit violated
    //      at either end of the candidate segment [x, x + dx]?  (if not, no candidate pays a limit penalty)
    bool viol = (l1 >= 0);  // several limit-carrying sub-systems: always take the exact path below
    if constexpr (SPLIT) viol = viol || (W.viol[b] != 0);
    if (!SPLIT && l0 >= 0) {  // the initial state is common to all candidates but still pays its penalty
#pragma unroll
        for (int rt = 0; rt < NT; rt++) {
            const int i = 8 * rt + g;
            const double x00 = (i < n) ? Xb[i] : 0.0;
            viol = viol || (x00 > lmx[rt]) || (x00 < lmn[rt]);
        }
    }
    if constexpr (!SPLIT) {
        // L2 prefetch stream, PFD steps ahead of the register ring: lanes 0-3 cover the 4 cache lines a gain matrix can
        // straddle, lane 4 the feed-forward row, lane 5 the control row (phases C, E), lane 6 the state row
        constexpr int PFD = 24;
        const char* pf_ptr = nullptr;
        int pf_stride = 0;
        if (lane < 4) {
            pf_ptr = reinterpret_cast<const char*>(Kb) + lane * 128;
            pf_stride = m * ns * 8;
            if (lane * 128 >= pf_stride) pf_ptr = nullptr;
        } else if (lane == 4) {
            pf_ptr = reinterpret_cast<const char*>(db);
            pf_stride = ms * 8;
        } else if (lane == 5) {
            pf_ptr = reinterpret_cast<const char*>(Ub);
            pf_stride = ms * 8;
        } else if (lane == 6 && l0 >= 0) {
            pf_ptr = reinterpret_cast<const char*>(Xb + ns);
            pf_stride = ns * 8;
        }
        auto prefetch_step = [&](int k) {
            if (pf_ptr != nullptr && k < H - 1) asm volatile("prefetch.global.L2 [%0];" ::"l"(pf_ptr + (size_t)k * pf_stride));
        };
        for (int k = 0; k < PFD; k++) prefetch_step(k);

        // per-lane validity of the fragment entries and running pointers
        bool vrow[NT], vcol[NT][2];
#pragma unroll
        for (int rt = 0; rt < NT; rt++) {
            vrow[rt] = 8 * rt + g < m;
#pragma unroll
            for (int s = 0; s < 2; s++) vcol[rt][s] = 8 * rt + 4 * s + t < n;
        }
        const double* pK = Kb + g * ns + t;   // + rt*8*ns + kt*8 + s*4, advances m*ns per step
        const double* pD = db + g;            // + rt*8, advances ms per step
        const double* pX = Xb + ns + g;       // state k+1, + rt*8, advances ns per step
        double aK[PF][NT][NT][2], aD[PF][NT], aX[PF][NT];
        auto loadK = [&](int k, double (&a)[NT][NT][2], double (&dd)[NT], double (&xx)[NT]) {
            const bool in = k < H - 1;
            const double* qK = pK + (size_t)k * m * ns;
#pragma unroll
            for (int rt = 0; rt < NT; rt++) {
                dd[rt] = (in && vrow[rt]) ? __ldcs(pD + (size_t)k * ms + 8 * rt) : 0.0;
                xx[rt] = (in && vrow[rt] && l0 >= 0) ? pX[(size_t)k * ns + 8 * rt] : 0.0;
#pragma unroll
                for (int kt = 0; kt < NT; kt++)
#pragma unroll
                    for (int s = 0; s < 2; s++) a[rt][kt][s] = (in && vrow[rt] && vcol[kt][s]) ? __ldcs(qK + rt * 8 * ns + kt * 8 + s * 4) : 0.0;
            }
        };
#pragma unroll
        for (int pp = 0; pp < PF; pp++) loadK(pp, aK[pp], aD[pp], aX[pp]);
        double dxr[NT], bx[NT][2];
#pragma unroll
        for (int rt = 0; rt < NT; rt++) {
            dxr[rt] = 0.0;
            bx[rt][0] = bx[rt][1] = 0.0;
        }
        int skp = 0;
        int next_kp = S > 0 ? P.kp_steps[0] : -1;
        if (next_kp == 0) {  // a keypoint on the initial state: dx = 0
            if (lane < ns) sDXk[lane] = 0.0;
            skp++;
            next_kp = skp < S ? P.kp_steps[skp] : -1;
        }
        for (int k0 = 0; k0 < H - 1; k0 += PF) {
#pragma unroll
            for (int pp = 0; pp < PF; pp++) {
                const int k = k0 + pp;
                if (k < H - 1) {
                    prefetch_step(k + PFD);
#pragma unroll
                    for (int rt = 0; rt < NT; rt++) {
                        double c0 = 0.0, c1 = 0.0;
#pragma unroll
                        for (int kt = 0; kt < NT; kt++)
#pragma unroll
                            for (int s = 0; s < 2; s++) dmma884(c0, c1, aK[pp][rt][kt][s], bx[kt][s]);
                        const double du = c0 + aD[pp][rt];
                        dxr[rt] = dxr[rt] + dt * du;
                        if (t == 0 && 8 * rt + g < ms) sDU[k * ms + 8 * rt + g] = du;
                        const double x0 = aX[pp][rt], x1 = x0 + dxr[rt];  // state k+1 (state 0 is fixed and shared)
                        viol = viol || (x0 > lmx[rt]) || (x0 < lmn[rt]) || (x1 > lmx[rt]) || (x1 < lmn[rt]);
                    }
                    if (k + 1 == next_kp) {
#pragma unroll
                        for (int rt = 0; rt < NT; rt++)
                            if (t == 0 && 8 * rt + g < ns) sDXk[skp * ns + 8 * rt + g] = dxr[rt];
                        skp++;
                        next_kp = skp < S ? P.kp_steps[skp] : -1;
                    }
#pragma unroll
                    for (int kt = 0; kt < NT; kt++)
#pragma unroll
                        for (int s = 0; s < 2; s++) bx[kt][s] = __shfl_sync(FULL, dxr[kt], 4 * (4 * s + t));
                    loadK(k + PF, aK[pp], aD[pp], aX[pp]);
                }
            }
        }
    }
    __syncwarp();

    // ---- phase B: sum_k |du_k| at alpha = 1 (early-stop test, ILQRRecursive.cpp:129,174); exact limit penalties if needed
    double dun1 = 0.0;
    if (early_stop) {
        for (int k = lane; k < H - 1; k += 32) {
            double s2 = 0.0;
            for (int i = 0; i < m; i++) {
                const double v = sDU[k * ms + i];
                s2 += v * v;
            }
            dun1 += sqrt(s2);
        }
        dun1 = warp_sum(dun1);
    }
    const int ncand = line_search ? NCAND : 1;
    double lim_mine = 0.0;  // lane c < ncand: joint-limit cost of candidate c over the whole horizon (System.cpp:225-231)
    if (__any_sync(FULL, viol)) {
        // rare: lane i < n walks the horizon for state entry i, rebuilding dx_i from du_i, one accumulator per candidate
        double acc[NCAND];
#pragma unroll
        for (int c = 0; c < NCAND; c++) acc[c] = 0.0;
        if (lane < n) {
            double dxi = 0.0;
            for (int k = 0; k < H; k++) {
                const double xi = Xb[(size_t)k * ns + lane];
#pragma unroll
                for (int c = 0; c < NCAND; c++) {
                    const double xc = xi + alpha_of(c) * dxi;
                    for (int ti = 0; ti < P.n_tasks; ti++) {
                        const DevTask& tk = P.tasks[ti];
                        if (!tk.limits_set) continue;
                        bool act;
                        const double q = limit_q(tk, lane, xc, act);
                        if (act) acc[c] += (q * 1.0) * q;
                    }
                }
                if (k < H - 1) dxi = dxi + dt * sDU[k * ms + lane];
            }
        }
#pragma unroll
        for (int c = 0; c < NCAND; c++) {
            const double v = warp_sum(acc[c]);
            if (lane == c) lim_mine = v;
        }
    }

    // ---- phase C: task cost of every (candidate, keypoint step) pair, one pair per lane      System.cpp:213-222
    for (int base = 0; base < ncand * S; base += 32) {
        const int idx = base + lane;
        const bool act = idx < ncand * S;
        const int c = act ? idx / S : 0, s = act ? idx % S : 0, k = P.kp_steps[s];
        const double alpha = alpha_of(c);
        const double* xk = Xb + (size_t)k * ns;
        const double* dxk = sDXk + s * ns;
        double cost = 0.0;
        for (int ti = 0; ti < P.n_tasks; ti++) {
            const DevTask& tk = P.tasks[ti];
            const int kp = task_kp_at(tk, k);
            if (kp < 0) continue;
            const int off = tk.joint_offset;
            double fx[7], Jc[RR], res[RR];
            lean_task_eval<RR>(P, tk, [&](int j) { return xk[off + j] + alpha * dxk[off + j]; }, -1, fx, Jc);
            double cst = lean_task_cost(P, tk, b, kp, fx, res);
            if (k < H - 1) {
                double uRu = 0.0, uo[NX], ud[NX];  // compile-time bounds: the loads of a row are in flight together
#pragma unroll
                for (int j = 0; j < NX; j++) {
                    uo[j] = j < m ? Ub[(size_t)k * ms + j] : 0.0;
                    ud[j] = j < m ? sDU[k * ms + j] : 0.0;
                }
#pragma unroll
                for (int j = 0; j < NX; j++)
                    if (j < m) {
                        const double uj = uo[j] + alpha * ud[j];
                        uRu += (uj * tk.r_diag[j]) * uj;
                    }
                cst += uRu;
            }
            cost += cst;
        }
        if (act) s_cc[c * S + s] = cost;
    }
    __syncwarp();

    // ---- phase D: the reference's sequential choice (ILQRRecursive.cpp:107-155, Q2)
    const double cost0 = W.cost0[b];
    double tot = 0.0;
    if (lane < ncand) {
        for (int s = 0; s < S; s++) tot += s_cc[lane * S + s];
        tot += lim_mine;
    }
    const bool good = (lane < ncand) && !((tot >= cost0) || isnan(tot));
    const unsigned bal = __ballot_sync(FULL, good);
    int sel = 0;
    if (line_search) sel = bal ? (__ffs(bal) - 1) : (NCAND - 1);
    const double newCost = __shfl_sync(FULL, tot, sel);
    const double alpha = alpha_of(sel);

    // ---- phase E: accept.  u += alpha du ; x re-integrated from u with the reference's operation order
    //      (SimulationInterface.cpp:19-31).  E1, all lanes: the new controls of every (step, joint) pair, coalesced, and the increment
    //      dt u' + dt^2/2 * 0 left in the deviation's slot.  E2, lane i < dof: joint i walks the horizon, one load and one add per step.
    {
        const double hdt2 = __dmul_rn(dt, dt) / 2;
        const double zterm = __dmul_rn(hdt2, 0.0);
        const int tot = (H - 1) * ms;
        for (int e0 = 0; e0 < tot; e0 += 128) {
            double uu[4], dd[4];
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int idx = e0 + 32 * r + lane;
                uu[r] = idx < tot ? Ub[idx] : 0.0;
                dd[r] = idx < tot ? sDU[idx] : 0.0;
            }
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int idx = e0 + 32 * r + lane;
                if (idx < tot) {
                    const double un = uu[r] + alpha * dd[r];
                    Ub[idx] = un;
                    sDU[idx] = __dadd_rn(__dmul_rn(dt, un), zterm);
                }
            }
        }
        __syncwarp();
        if (lane < dof) {
            double xq = P.q0[(size_t)b * dof + lane];
            for (int k0 = 0; k0 < H - 1; k0 += 8) {
                double inc[8];
#pragma unroll
                for (int j = 0; j < 8; j++) inc[j] = (k0 + j < H - 1) ? sDU[(k0 + j) * ms + lane] : 0.0;
#pragma unroll
                for (int j = 0; j < 8; j++)
                    if (k0 + j < H - 1) {
                        xq = __dadd_rn(xq, inc[j]);
                        Xb[(size_t)(k0 + j + 1) * ns + lane] = xq;
                    }
            }
        }
    }
    __syncwarp();

    // ---- phase F: l_x = sum_tasks ((-J^T) Q) diff, l_xx = sum_tasks (J^T Q) J at the keypoint steps of the accepted
    //      trajectory (System.cpp:252-260,290-296).  Lane (sg, i): keypoint step s0 + sg, state column i; every lane
    //      forms its own Jacobian column and reads the others' through shared memory.
    if (it + 1 < nb_iter) {
        constexpr int GRP = 32 / NX;
        const int sg = lane / NX, i = lane % NX;
        for (int s0 = 0; s0 < S; s0 += GRP) {
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
                const int off = tk.joint_offset, r = tk.r_t;
                const int cj_ = i - off;
                const int cj = (cj_ >= 0 && cj_ < P.robots[tk.robot].dof) ? cj_ : -1;
                double fx[7], Jc[RR], res[RR], jq[RR];
#pragma unroll
                for (int a = 0; a < RR; a++) {
                    res[a] = 0.0;
                    Jc[a] = 0.0;
                }
                lean_task_eval<RR>(P, tk, [&](int j) { return xk[off + j]; }, cj, fx, Jc);
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
                            double h = 0.0;
#pragma unroll
                            for (int c = 0; c < RR; c++)
                                if (c < r) h += jq[c] * Jj[c];
                            row[j] += h;
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
    const bool stop = early_stop && (alpha * sqrt(alpha * dun1) < 1e-3) && (newCost < 1e-3);  // ILQRRecursive.cpp:174
    if (lane == 0) {
        W.cost0[b] = newCost;
        W.cost_hist[(size_t)b * nb_iter + it] = newCost;
        W.alpha_hist[(size_t)b * nb_iter + it] = alpha;
        W.alpha_last[b] = alpha;
        W.iters[b] = it + 1;
        int st = W.status[b];
        if (isnan(newCost)) st |= ILQRB_ST_NAN_COST;
        if (stop) {
            st |= ILQRB_ST_EARLY_STOP;
            W.done[b] = 1;
        }
        W.status[b] = st;
    }
    return stop;
}

template <int NT, int DOF, int MINB>
__global__ void __launch_bounds__(32, MINB) k_forward_n1(const DevProblem* __restrict__ Pp, Work W, int b0, int it, int nb_iter, int line_search, int early_stop) {
    extern __shared__ double smem[];
    const int b = b0 + blockIdx.x;
    if (W.done[b]) return;
    forward_n1_body<NT, DOF>(*Pp, W, b, threadIdx.x, smem, it, nb_iter, line_search, early_stop);
}

}  // namespace ilqrb
