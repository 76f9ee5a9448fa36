// fused1.cuh — fast path of the Recursive iLQR solve for single-integrator systems without a time state
// (PosOrnPlannerSys nb_deriv = 1, the POSITION system; any robot / task list):  n = m = dof,
//   A = I,  B = dt * I          (PosOrnPlannerSys.cpp:115-121)
//
// Two kernels per iteration, one problem per warp:
//
//   k_backward_n1   Riccati recursion (ILQRRecursive.cpp:69-97).  [P|p] lives in registers in the DMMA accumulator
//                   layout; the products with A and B are exact scalings (x*1 and +0 are exact in IEEE, so
//                   (B^T P) A == dt*P bit for bit), Quu^-1 is an in-register Gauss-Jordan driven by shuffles, and the
//                   five dense products that remain per step (K = -Inv Qux, K^T Quu, and the three terms of the P
//                   update) run on the FP64 tensor pipe (mma.sync m8n8k4.f64), operands re-laid-out through shared
//                   memory.
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
#include "recursive.cuh"

namespace ilqrb {

constexpr int MAX_KP_STEPS = ILQRB_MAX_TASKS * ILQRB_MAX_KEYPOINTS;

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

// In-register inverse of the leading m x m block of an (8NT)^2 matrix held in accumulator layout: Gauss-Jordan
// without pivoting (Quu + 1e-6 I is symmetric positive definite in normal operation; replaces the PartialPivLU
// inverse of ILQRRecursive.cpp:89).  Pivot row / column / element travel by warp shuffles; no shared memory, no
// barrier.  Entries outside the leading block are left untouched.  Returns false if a pivot was <= 0 or NaN.
template <int NT>
__device__ __forceinline__ bool gj_inverse_regs(double (&e)[NT][NT][2], int m, int g, int t) {
    constexpr unsigned FULL = 0xffffffffu;
    bool ok = true;
#pragma unroll
    for (int k = 0; k < 8 * NT; k++) {
        if (k < m) {  // warp-uniform
            const int kt = k >> 3, kg = k & 7, kh = k & 1;
            double rowk[NT][2], colk[NT];
#pragma unroll
            for (int ct = 0; ct < NT; ct++) {
                rowk[ct][0] = __shfl_sync(FULL, e[kt][ct][0], 4 * kg + t);
                rowk[ct][1] = __shfl_sync(FULL, e[kt][ct][1], 4 * kg + t);
            }
#pragma unroll
            for (int rt = 0; rt < NT; rt++) colk[rt] = __shfl_sync(FULL, e[rt][kt][kh], 4 * g + (kg >> 1));
            const double pkk = __shfl_sync(FULL, e[kt][kt][kh], 4 * kg + (kg >> 1));
            if (!(pkk > 0.0)) ok = false;
            const double piv = 1.0 / pkk;
#pragma unroll
            for (int rt = 0; rt < NT; rt++) {
                const int i = 8 * rt + g;
#pragma unroll
                for (int ct = 0; ct < NT; ct++)
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        const int j = 8 * ct + 2 * t + h;
                        const double rkj = rowk[ct][h] * piv;
                        double v;
                        if (i == k) v = (j == k) ? piv : rkj;
                        else if (j == k) v = -colk[rt] * piv;
                        else v = e[rt][ct][h] - colk[rt] * rkj;
                        if (i < m && j < m) e[rt][ct][h] = v;
                    }
            }
        }
    }
    return ok;
}

template <int NT>
struct BwdN1Smem {
    static constexpr int NX = 8 * NT, LD = NX + 4, SZ = NX * LD;
    static constexpr int PER_WARP = 5 * SZ;  // Inv, [Qux|Qu], Quu, [K|d], K^T Quu
};

// ------------------------------------------------------------------------------------------------------------
// backward Riccati pass, A = I, B = dt I.  Augmented layout: column n of the state-indexed matrices carries the
// vector ([P|p], [Qux|Qu], [Qxx|Qx], [K|d]); rows >= n and columns > n are kept at exact zero.
template <int NT, int WARPS, int MINB>
__global__ void __launch_bounds__(32 * WARPS, MINB) k_backward_n1(const DevProblem* __restrict__ Pp, Work W) {
    using S = BwdN1Smem<NT>;
    constexpr int LD = S::LD, SZ = S::SZ;
    extern __shared__ double smem[];
    const DevProblem& P = *Pp;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int b = blockIdx.x * WARPS + wid;
    if (b >= P.B) return;
    if (W.done[b]) return;
    const int g = lane >> 2, t = lane & 3;
    const int n = P.n, m = P.m, H = P.H, ns = W.ns, ms = W.ms;
    const double dt = 1.0 * P.dt;  // B = Identity * dt
    double* sInv = smem + (size_t)wid * S::PER_WARP;
    double* sQux = sInv + SZ;
    double* sQuu = sQux + SZ;
    double* sKd = sQuu + SZ;
    double* sKtQ = sKd + SZ;

    double Rrow[NT];
#pragma unroll
    for (int rt = 0; rt < NT; rt++) Rrow[rt] = (8 * rt + g < m) ? P.r_diag[8 * rt + g] : 0.0;

    const double* Xb = W.X + (size_t)b * H * ns;
    const double* Ub = W.U + (size_t)b * (H - 1) * ms;

    // e += [l_xx | l_x] of keypoint step index s (k_kp_derivs / k_forward_n1 wrote them)    System.cpp:252-260,290-296
    auto add_kp = [&](double (&e)[NT][NT][2], int s) {
        const double* lx = W.LX + ((size_t)b * P.n_kp_steps + s) * n;
        const double* lxx = W.LXX + ((size_t)b * P.n_kp_steps + s) * n * n;
#pragma unroll
        for (int rt = 0; rt < NT; rt++) {
            const int i = 8 * rt + g;
#pragma unroll
            for (int ct = 0; ct < NT; ct++)
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int j = 8 * ct + 2 * t + h;
                    if (i < n) {
                        if (j < n) e[rt][ct][h] += lxx[i * n + j];
                        else if (j == n) e[rt][ct][h] += lx[i];
                    }
                }
        }
    };
    // e += [L^T L | -L^T q] of the joint-limit penalty at state row values xr          System.cpp:263-269,299-305
    auto add_limits = [&](double (&e)[NT][NT][2], const double (&xr)[NT]) {
        for (int tk_i = 0; tk_i < P.n_tasks; tk_i++) {
            const DevTask& tk = P.tasks[tk_i];
            if (!tk.limits_set) continue;
#pragma unroll
            for (int rt = 0; rt < NT; rt++) {
                const int i = 8 * rt + g;
                if (i < n) {
                    bool act;
                    const double q = limit_q(tk, i, xr[rt], act);
                    if (act) {
#pragma unroll
                        for (int ct = 0; ct < NT; ct++)
#pragma unroll
                            for (int h = 0; h < 2; h++) {
                                const int j = 8 * ct + 2 * t + h;
                                if (j == i) e[rt][ct][h] += 1.0 * 1.0;
                                else if (j == n) e[rt][ct][h] += -1.0 * q;
                            }
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
    int sp = P.n_kp_steps - 1;
    if (sp >= 0 && P.kp_steps[sp] == H - 1) {
        add_kp(Pa, sp);
        sp--;
    }
    if (P.any_limits) {
        double xr[NT];
        load_rows(Xb, ns, H - 1, n, xr);
        add_limits(Pa, xr);
    }
    double un[NT], xn[NT];
    load_rows(Ub, ms, H - 2, m, un);
    load_rows(Xb, ns, H - 2, n, xn);
    bool spd = true;

    for (int k = H - 2; k >= 0; k--) {
        const size_t bk = (size_t)b * (H - 1) + k;
        double u[NT], x[NT];
#pragma unroll
        for (int rt = 0; rt < NT; rt++) {
            u[rt] = un[rt];
            x[rt] = xn[rt];
        }
        if (k > 0) {  // prefetch the next step's control / state rows
            load_rows(Ub, ms, k - 1, m, un);
            load_rows(Xb, ns, k - 1, n, xn);
        }
        // [Qux|Qu] = [0 | R u] + (B^T [P|p]) A_aug ;  Quu = R + (B^T P) B ;  [Qxx|Qx] = [l_xx|l_x] + (A^T [P|p]) A_aug
        // Qxu = (A^T P) B = P*dt has the same entries as Qux = dt*P                     ILQRRecursive.cpp:82-87
        double Qa[NT][NT][2], Qu2[NT][NT][2], Qx[NT][NT][2];
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
                    Qu2[rt][ct][h] = (i < m && j < m) ? quu : 0.0;
                    Qx[rt][ct][h] = Pa[rt][ct][h];
                }
        }
        if (sp >= 0 && P.kp_steps[sp] == k) {
            add_kp(Qx, sp);
            sp--;
        }
        if (P.any_limits) add_limits(Qx, x);

        // Inv = (Quu + 1e-6 I)^-1                                                        ILQRRecursive.cpp:89
        double Inv[NT][NT][2];
#pragma unroll
        for (int rt = 0; rt < NT; rt++)
#pragma unroll
            for (int ct = 0; ct < NT; ct++)
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int i = 8 * rt + g, j = 8 * ct + 2 * t + h;
                    double v = Qu2[rt][ct][h];
                    if (i == j) v = (i < m) ? v + 1e-6 * 1.0 : 1.0;
                    Inv[rt][ct][h] = v;
                }
        spd = gj_inverse_regs<NT>(Inv, m, g, t) && spd;
        st_c<NT, LD>(sInv, Inv, g, t);
        st_c<NT, LD>(sQux, Qa, g, t);
        st_c<NT, LD>(sQuu, Qu2, g, t);
        __syncwarp();

        // [K|d] = -(Inv [Qux|Qu])                                                        ILQRRecursive.cpp:90-91
        double bQ[NT][2][NT];
#pragma unroll
        for (int kt = 0; kt < NT; kt++)
#pragma unroll
            for (int s = 0; s < 2; s++)
#pragma unroll
                for (int ct = 0; ct < NT; ct++) bQ[kt][s][ct] = ld_b<LD>(sQux, kt, s, ct, g, t);
        double Kd[NT][NT][2];
#pragma unroll
        for (int rt = 0; rt < NT; rt++)
#pragma unroll
            for (int ct = 0; ct < NT; ct++) {
                double c0 = 0.0, c1 = 0.0;
#pragma unroll
                for (int kt = 0; kt < NT; kt++)
#pragma unroll
                    for (int s = 0; s < 2; s++) dmma884(c0, c1, ld_a<LD>(sInv, rt, kt, s, g, t), bQ[kt][s][ct]);
                const bool row = 8 * rt + g < m;
                Kd[rt][ct][0] = row ? -c0 : 0.0;
                Kd[rt][ct][1] = row ? -c1 : 0.0;
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
                    if (j < ns) {
                        double2 v = make_double2(j < n ? Kd[rt][ct][0] : 0.0, j + 1 < n ? Kd[rt][ct][1] : 0.0);
                        *reinterpret_cast<double2*>(W.K + (bk * m + i) * ns + j) = v;
                    }
                    if (j == n) W.d[bk * ms + i] = Kd[rt][ct][0];
                    else if (j + 1 == n) W.d[bk * ms + i] = Kd[rt][ct][1];
                }
            }
        }
        __syncwarp();

        // K^T Quu  (row n = d^T Quu is produced too and masked out of P below)
        double kb[NT][2][NT], bU[NT][2][NT];
#pragma unroll
        for (int kt = 0; kt < NT; kt++)
#pragma unroll
            for (int s = 0; s < 2; s++)
#pragma unroll
                for (int ct = 0; ct < NT; ct++) {
                    kb[kt][s][ct] = ld_b<LD>(sKd, kt, s, ct, g, t);  // B fragment of [K|d] == A fragment of [K|d]^T
                    bU[kt][s][ct] = ld_b<LD>(sQuu, kt, s, ct, g, t);
                }
        double KtQ[NT][NT][2];
#pragma unroll
        for (int rt = 0; rt < NT; rt++)
#pragma unroll
            for (int ct = 0; ct < NT; ct++) {
                double c0 = 0.0, c1 = 0.0;
#pragma unroll
                for (int kt = 0; kt < NT; kt++)
#pragma unroll
                    for (int s = 0; s < 2; s++) dmma884(c0, c1, kb[kt][s][rt], bU[kt][s][ct]);
                KtQ[rt][ct][0] = c0;
                KtQ[rt][ct][1] = c1;
            }
        st_c<NT, LD>(sKtQ, KtQ, g, t);
        __syncwarp();

        // [P|p] = [Qxx|Qx] + (K^T Quu)[K|d] + K^T [Qux|Qu] + Qxu [K|d]                   ILQRRecursive.cpp:93-94
#pragma unroll
        for (int rt = 0; rt < NT; rt++)
#pragma unroll
            for (int ct = 0; ct < NT; ct++) {
                double c0 = Qx[rt][ct][0], c1 = Qx[rt][ct][1];
#pragma unroll
                for (int kt = 0; kt < NT; kt++)
#pragma unroll
                    for (int s = 0; s < 2; s++) dmma884(c0, c1, ld_a<LD>(sKtQ, rt, kt, s, g, t), kb[kt][s][ct]);
#pragma unroll
                for (int kt = 0; kt < NT; kt++)
#pragma unroll
                    for (int s = 0; s < 2; s++) dmma884(c0, c1, kb[kt][s][rt], bQ[kt][s][ct]);
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
    if (!spd && lane == 0) W.status[b] |= ILQRB_ST_QUU_NOT_SPD;
}

// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// column j of the task Jacobian (r_t entries) — same content as task_jacobian(), one column at a time
__device__ __forceinline__ void task_jac_col(const DevProblem& P, const DevTask& tk, const EEState& e, int j, double* col) {
    const DevRobot& rb = P.robots[tk.robot];
    const int nd = P.nb_deriv, dof = P.dof;
    for (int a = 0; a < tk.r_t; a++) col[a] = 0.0;
    const int rows = (P.sys_kind == ILQRB_SYS_POSITION) ? tk.nt_t / nd : 6;
    if (j < nd * dof) {
        const int blk = j / dof, c = j - blk * dof - tk.joint_offset;
        if (c >= 0 && c < rb.dof)
            for (int a = 0; a < rows; a++) col[blk * rows + a] = (a < 3) ? e.Jv[c][a] : e.Jw[c][a - 3];
    }
    if (P.time && j == P.n - 1) col[tk.r_t - 1] = 1.0;
}

// row i of l_xx and entry i of l_x at keypoint-step index s of the current trajectory (System.cpp:252-260,290-296):
//   l_x = sum_tasks ((-J^T) Q) diff ,  l_xx = sum_tasks (J^T Q) J
__device__ __noinline__ void kp_deriv_row(const DevProblem& P, const Work& W, int b, int s, int i) {
    const int k = P.kp_steps[s], n = P.n, dof = P.dof;
    const double* x = W.X + ((size_t)b * P.H + k) * W.ns;
    double xs[MAXN], dqz[MAXM], row[MAXN];
    for (int a = 0; a < n; a++) {
        xs[a] = x[a];
        row[a] = 0.0;
    }
    for (int a = 0; a < dof; a++) dqz[a] = 0.0;
    double lxi = 0.0;
    for (int ti = 0; ti < P.n_tasks; ti++) {
        const DevTask& tk = P.tasks[ti];
        const int kp = task_kp_at(tk, k);
        if (kp < 0) continue;
        EEState e;
        task_ee(P, tk, xs, P.nb_deriv == 2 ? xs + dof : dqz, e);
        double fx[MAX_NTT], res[MAX_RT];
        task_fx(P, tk, e, P.time ? xs[n - 1] : 0.0, fx);
        const double* tg = tk.kp_target + ((size_t)b * tk.n_keypoints + kp) * ILQRB_KP_STRIDE;
        task_residual(P, tk, fx, tg, res);
        const int r = tk.r_t;
        const double* Q = tk.kp_precision + ((size_t)(tk.precision_per_problem ? b : 0) * tk.n_keypoints + kp) * r * r;
        double Ji[MAX_RT], jq[MAX_RT];
        task_jac_col(P, tk, e, i, Ji);
        double gsum = 0.0;
        for (int c = 0; c < r; c++) {
            double a = 0.0, na = 0.0;
            for (int a2 = 0; a2 < r; a2++) {
                a += Ji[a2] * Q[a2 * r + c];
                na += (-1.0 * Ji[a2]) * Q[a2 * r + c];
            }
            jq[c] = a;
            gsum += na * res[c];
        }
        lxi += gsum;
        for (int j = 0; j < n; j++) {
            double Jj[MAX_RT];
            task_jac_col(P, tk, e, j, Jj);
            double h = 0.0;
            for (int c = 0; c < r; c++) h += jq[c] * Jj[c];
            row[j] += h;
        }
    }
    W.LX[((size_t)b * P.n_kp_steps + s) * n + i] = lxi;
    double* out = W.LXX + (((size_t)b * P.n_kp_steps + s) * n + i) * n;
    for (int j = 0; j < n; j++) out[j] = row[j];
}

// joint-limit cost q^T L q of one state vector element (System.cpp:225-231), summed over the tasks that set limits
__device__ __forceinline__ double limit_cost_elem(const DevProblem& P, int i, double xi, bool& any) {
    double c = 0.0;
    for (int ti = 0; ti < P.n_tasks; ti++) {
        const DevTask& tk = P.tasks[ti];
        if (!tk.limits_set) continue;
        bool act;
        const double q = limit_q(tk, i, xi, act);
        if (act) {
            c += (q * 1.0) * q;
            any = true;
        }
    }
    return c;
}

// ------------------------------------------------------------------------------------------------------------
// forward pass: line search over all candidates, accept, derivatives of the accepted trajectory.
// Scratch (per problem, inside W.Uc): DX [H][ns] and DU [H-1][ms] = the alpha = 1 closed-loop deviation.
template <int NT, int WARPS>
__global__ void __launch_bounds__(32 * WARPS) k_forward_n1(const DevProblem* __restrict__ Pp, Work W, int it, int nb_iter, int line_search, int early_stop) {
    constexpr int NX = 8 * NT;
    constexpr unsigned FULL = 0xffffffffu;
    __shared__ double s_cc[WARPS][NCAND][MAX_KP_STEPS];
    const DevProblem& P = *Pp;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int b = blockIdx.x * WARPS + wid;
    if (b >= P.B) return;
    if (W.done[b]) return;
    const int H = P.H, n = P.n, m = P.m, ns = W.ns, ms = W.ms, dof = P.dof, S = P.n_kp_steps;
    const double dt = P.dt;
    double* DX = W.Uc + (size_t)b * (H - 1) * NCAND * ms;
    double* DU = DX + (size_t)H * ns;
    double* Xb = W.X + (size_t)b * H * ns;
    double* Ub = W.U + (size_t)b * (H - 1) * ms;
    const double* Kb = W.K + (size_t)b * (H - 1) * m * ns;
    const double* db = W.d + (size_t)b * (H - 1) * ms;

    // ---- phase A: alpha = 1 deviation rollout.  lane i < n owns dx_i and row i of K_k.
    //      du = K (x' - x) + d ; x'+ = x' + dt u'     (ILQRRecursive.cpp:124-136 with alpha = 1, minus the nominal)
    {
        double dx = 0.0;
        if (lane < ns) DX[lane] = 0.0;
        double kc[NX], dc = 0.0;
        auto loadK = [&](int k, double (&kr)[NX], double& dd) {
            if (lane < m) {
                const double2* p = reinterpret_cast<const double2*>(Kb + ((size_t)k * m + lane) * ns);
#pragma unroll
                for (int j2 = 0; j2 < NX / 2; j2++)
                    if (2 * j2 < ns) {
                        const double2 v = p[j2];
                        kr[2 * j2] = v.x;
                        kr[2 * j2 + 1] = v.y;
                    }
                dd = db[(size_t)k * ms + lane];
            }
        };
#pragma unroll
        for (int j = 0; j < NX; j++) kc[j] = 0.0;
        loadK(0, kc, dc);
        for (int k = 0; k < H - 1; k++) {
            double kn[NX], dn = 0.0;
#pragma unroll
            for (int j = 0; j < NX; j++) kn[j] = 0.0;
            if (k + 1 < H - 1) loadK(k + 1, kn, dn);
            double a0 = 0.0, a1 = 0.0;
#pragma unroll
            for (int j = 0; j < NX; j += 2) {
                if (j < n) a0 += kc[j] * __shfl_sync(FULL, dx, j);
                if (j + 1 < n) a1 += kc[j + 1] * __shfl_sync(FULL, dx, j + 1);
            }
            const double du = (lane < m) ? (a0 + a1) + dc : 0.0;
            if (lane < ms) DU[(size_t)k * ms + lane] = du;
            if (lane < dof) dx = dx + dt * du;
            if (lane < ns) DX[(size_t)(k + 1) * ns + lane] = (lane < n) ? dx : 0.0;
#pragma unroll
            for (int j = 0; j < NX; j++) kc[j] = kn[j];
            dc = dn;
        }
    }
    __syncwarp();

    // ---- phase B: terms that run over every step, lanes stride over k
    double dun1 = 0.0;  // sum_k |du_k| at alpha = 1 (early-stop test, ILQRRecursive.cpp:129,174)
    if (early_stop) {
        for (int k = lane; k < H - 1; k += 32) {
            double s2 = 0.0;
            for (int i = 0; i < m; i++) {
                const double v = DU[(size_t)k * ms + i];
                s2 += v * v;
            }
            dun1 += sqrt(s2);
        }
        dun1 = warp_sum(dun1);
    }
    const int ncand = line_search ? NCAND : 1;
    double lim_mine = 0.0;  // lane c < ncand: joint-limit cost of candidate c over the whole horizon
    if (P.any_limits) {
        // the candidates lie on the segment [x, x + dx]: if neither end violates a limit, none of them does
        bool viol = false;
        for (int k = lane; k < H; k += 32)
            for (int i = 0; i < n; i++) {
                const double x0 = Xb[(size_t)k * ns + i], x1 = x0 + DX[(size_t)k * ns + i];
                limit_cost_elem(P, i, x0, viol);
                limit_cost_elem(P, i, x1, viol);
            }
        if (__any_sync(FULL, viol)) {
            for (int c = 0; c < ncand; c++) {
                const double alpha = alpha_of(c);
                double acc = 0.0;
                bool dummy = false;
                for (int k = lane; k < H; k += 32)
                    for (int i = 0; i < n; i++) acc += limit_cost_elem(P, i, Xb[(size_t)k * ns + i] + alpha * DX[(size_t)k * ns + i], dummy);
                acc = warp_sum(acc);
                if (lane == c) lim_mine = acc;
            }
        }
    }

    // ---- phase C: task cost of every (candidate, keypoint step) pair, one pair per lane
    for (int idx = lane; idx < ncand * S; idx += 32) {
        const int c = idx / S, s = idx % S, k = P.kp_steps[s];
        const double alpha = alpha_of(c);
        double x[NX], u[NX];
#pragma unroll
        for (int i = 0; i < NX; i++) {
            x[i] = (i < n) ? Xb[(size_t)k * ns + i] + alpha * DX[(size_t)k * ns + i] : 0.0;
            u[i] = (i < m && k < H - 1) ? Ub[(size_t)k * ms + i] + alpha * DU[(size_t)k * ms + i] : 0.0;
        }
        s_cc[wid][c][s] = stage_cost<NX, false>(P, b, x, k < H - 1 ? u : nullptr, k);
    }
    __syncwarp();

    // ---- phase D: the reference's sequential choice (ILQRRecursive.cpp:107-155, Q2)
    const double cost0 = W.cost0[b];
    double tot = 0.0;
    if (lane < ncand) {
        for (int s = 0; s < S; s++) tot += s_cc[wid][lane][s];
        tot += lim_mine;
    }
    const bool good = (lane < ncand) && !((tot >= cost0) || isnan(tot));
    const unsigned bal = __ballot_sync(FULL, good);
    int sel = 0;
    if (line_search) sel = bal ? (__ffs(bal) - 1) : (NCAND - 1);
    const double newCost = __shfl_sync(FULL, tot, sel);
    const double alpha = alpha_of(sel);

    // ---- phase E: accept.  U += alpha du ; X re-integrated from U (SimulationInterface.cpp:19-31 operation order)
    for (int e = lane; e < (H - 1) * ms; e += 32) Ub[e] = Ub[e] + alpha * DU[e];
    __syncwarp();
    if (lane < dof) {
        double xq = P.q0[(size_t)b * dof + lane];
        const double hdt2 = __dmul_rn(dt, dt) / 2;
        for (int k = 0; k < H - 1; k++) {
            const double uu = Ub[(size_t)k * ms + lane];
            xq = __dadd_rn(xq, __dadd_rn(__dmul_rn(dt, uu), __dmul_rn(hdt2, 0.0)));
            Xb[(size_t)(k + 1) * ns + lane] = xq;
        }
    }
    __syncwarp();

    // ---- phase F: l_x, l_xx at the keypoint steps of the accepted trajectory; bookkeeping
    if (it + 1 < nb_iter)
        for (int idx = lane; idx < S * n; idx += 32) kp_deriv_row(P, W, b, idx / n, idx % n);
    if (lane == 0) {
        W.cost0[b] = newCost;
        W.cost_hist[(size_t)b * nb_iter + it] = newCost;
        W.alpha_hist[(size_t)b * nb_iter + it] = alpha;
        W.alpha_last[b] = alpha;
        W.iters[b] = it + 1;
        int st = W.status[b];
        if (isnan(newCost)) st |= ILQRB_ST_NAN_COST;
        if (early_stop && (alpha * sqrt(alpha * dun1) < 1e-3) && (newCost < 1e-3)) {
            st |= ILQRB_ST_EARLY_STOP;
            W.done[b] = 1;
        }
        W.status[b] = st;
    }
}

}  // namespace ilqrb
