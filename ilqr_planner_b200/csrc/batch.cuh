// batch.cuh — BatchILQR / BatchILQRCP path (BatchILQR.cpp:22-173, BatchILQRCP.cpp:21-174, System.cpp:111-211,321-361).
//
// Per iteration and problem.  The reference's `M = [A M, B]` recursion of buildSuJL (BatchILQR.cpp:62-98) makes the
// column block j of the row block of keypoint step t equal to A_{t-1} .. A_{j+1} B_j (B_j = dynamics B of control u_{j-1},
// B_0 = 0: the column shift Q3).  The integrator chains of the planners have closed-form products, so Su is written
// directly: single integrators A = I; double integrators A_l = [[I, dt_l I], [0, I]] => (prod A) B = B_q + tau B_v with
// tau = sum of the dt_l in between (batch_Su_entry).
// Keypoint slots: the reference indexes residuals / precisions by the time-sorted keypoint list WITH duplicates
// (truncateStates, diffBatch, getQMatrix(true)) but packs J, L and Su by DISTINCT timestep, leaving the trailing
// blocks zero (Q16).  BatchDev::vp / vds carry that mapping; without duplicates both indexings coincide.
//
//   k_batch_fp        fpBatch for the current controls AND for the 11 line-search candidates u + alpha du in one launch
//                     (thread = problem x candidate): rollout, task vector / residual at the keypoint rows, pre-step
//                     joint-limit terms (Q4), cost  e^T Q e + u^T R u + ql^T L ql ; candidate 0 (alpha = 0) also
//                     writes J, e, L, ql of the linearisation.
//   k_batch_assemble  W = J^T Q J + L, g = J^T Q e + L ql, the dense sensitivity matrix Su with the reference's column
//                     shift (Q3: row block of x_i is [0, B_1, .., B_{i-1}]), T = W Su, rhs = Su^T g - R u.
//   k_batch_syrk      lstq_A = Su^T T + R, dense N x N, FP64 tensor pipe (DMMA m8n8k4).          BatchILQR.cpp:130
//   k_batch_chol      in-place blocked Cholesky of lstq_A (one CTA per problem; diagonal block and panel in shared
//                     memory, trailing update on DMMA) — replaces `lstq_A.inverse()` (PartialPivLU) :132
//   k_batch_trsv      du = lstq_A^-1 rhs by the two triangular solves.
//   CP variant        lstq_A = (Su Psi)^T W (Su Psi) + Psi^T R Psi is nw x nw: one warp per problem factors and solves
//                     it in shared memory (k_batch_cp).                                          BatchILQRCP.cpp:129-133
//   k_batch_select    sequential line-search rule over the 11 candidate costs (accept if cost < cost0 or alpha < 1e-3,
//                     :135-157), u update, trace (the PRE-step cost is what the reference prints, :160), early stop.
#pragma once
#include <cuda.h>  // CUtensorMap (types only: cuTensorMapEncodeTiled is fetched through cudaGetDriverEntryPoint)

#include <algorithm>
#include <functional>
#include <string>
#include <vector>

#include "batch_types.hpp"
#include "handle.hpp"
#include "kin.cuh"
#include "wmat.cuh"
#include "work.cuh"

namespace ilqrb {

__device__ __forceinline__ double batch_alpha(int c) { return c == 0 ? 0.0 : 1.0 / (double)(1 << (c - 1)); }

// B block of the dynamics for the control u_j (row a, column c):  PosOrnPlannerSys.cpp:117-121 (dt I) and
// PosOrnTimePlannerSys.cpp:154-176 for nb_deriv = 1 (dt = s^2 ; last column [2 s u ; 2 s])
__device__ __forceinline__ double batch_B_entry(const DevProblem& P, const double* uj, int a, int c) {
    const int dof = P.dof, m = P.m, n = P.n;
    if (!P.time) return (a == c && a < dof) ? 1.0 * P.dt : 0.0;
    const double s = uj[m - 1];
    double v = 0.0;
    if (a < dof && c == a) v = 1.0 * (s * s);
    if (a < dof && c == m - 1) v = 2 * s * uj[a];
    if (a == n - 1 && c == m - 1) v = 2 * s;
    return v;
}

// (A_{ts-1} .. A_{j+1}) B_j, entry (a, c): the column block j (1 <= j <= ts-1) of the row block of keypoint step ts.
// u: controls of the problem, X: its states [H][n], cs: prefix sums of dt (cs[l] = dt_1 + .. + dt_l, dt_l = step of u_{l-1}).
// PosOrnPlannerSys.cpp:114-138, PosOrnTimePlannerSys.cpp:149-184 (Q15: the time column uses the POST-step velocity).
__device__ __forceinline__ double batch_Su_entry(const DevProblem& P, const double* u, const double* X, const double* cs, int ts, int j, int a, int c) {
    const int dof = P.dof, m = P.m, n = P.n;
    const double* uj = u + (size_t)(j - 1) * m;
    if (P.nb_deriv == 1) return batch_B_entry(P, uj, a, c);
    double dt = P.dt, sq = 0.0, tau = (double)(ts - 1 - j) * P.dt;
    if (P.time) {
        sq = uj[m - 1];
        dt = sq * sq;
        tau = cs[ts - 1] - cs[j];
        if (a == n - 1) return c == m - 1 ? 2 * sq : 0.0;
    }
    const int i = a < dof ? a : a - dof;
    double bv = (c == i) ? 1.0 * dt : 0.0;
    if (P.time && c == m - 1) bv = 2 * sq * uj[i];
    if (a >= dof) return bv;
    double bq = (c == i) ? 1.0 * dt * dt / 2 : 0.0;
    if (P.time && c == m - 1) bq = 2 * sq * X[(size_t)j * n + dof + i] + 2 * sq * sq * sq * uj[i];
    return bq + tau * bv;
}

// ------------------------------------------------------------------------------------------------------------
// fpBatch + cost for candidate c of problem b (System.cpp:181-211, BatchILQR.cpp:134,141-147)
__global__ void __launch_bounds__(64) k_batch_fp(const DevProblem* __restrict__ Pp, Work W, BatchDev BW, int b0, int nB, int first_iter) {
    const DevProblem& P = *Pp;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nB * BNC) return;
    const int bl = idx / BNC, c = idx % BNC, b = b0 + bl;
    if (W.done[b]) return;
    if (first_iter && c > 0) return;  // before the first solve there is no step to try
    const int H = P.H, n = P.n, m = P.m, dof = P.dof, S = BW.S, N = BW.N;
    const double alpha = batch_alpha(c);
    const double* u = BW.u + (size_t)b * N;
    const double* du = BW.du + (size_t)b * N;
    double x[MAXN], xprev[MAXN], uk[MAXM];
    for (int i = 0; i < MAXN; i++) x[i] = xprev[i] = 0.0;
    for (int i = 0; i < dof; i++) x[i] = P.q0[(size_t)b * dof + i];
    if (P.nb_deriv == 2)
        for (int i = 0; i < dof; i++) x[dof + i] = P.dq0[(size_t)b * dof + i];
    double uRu = 0.0, qLq = 0.0;
    double evec[MAX_KP_STEPS * MAX_RT];
    int s = 0;
    for (int row = 0; row < H; row++) {
        if (c == 0)
            for (int i = 0; i < n; i++) BW.X[((size_t)bl * H + row) * n + i] = x[i];
        if (s < S && P.kp_steps[s] == row) {
            // task vector, residual (and for candidate 0 the Jacobian) at the keypoint row
            double dqz[MAXM];
            for (int i = 0; i < dof; i++) dqz[i] = 0.0;
            for (int ti = 0; ti < P.n_tasks; ti++) {
                const DevTask& tk = P.tasks[ti];
                EEState ee;
                task_ee(P, tk, x, P.nb_deriv == 2 ? x + dof : dqz, ee);
                double fx[MAX_NTT], res[MAX_RT];
                task_fx(P, tk, ee, P.time ? x[n - 1] : 0.0, fx);
                const int kp = task_kp_at(tk, row);
                for (int a = 0; a < tk.r_t; a++) res[a] = 0.0;
                if (kp >= 0) {
                    const double* tg = tk.kp_target + ((size_t)b * tk.n_keypoints + kp) * ILQRB_KP_STRIDE;
                    task_residual(P, tk, fx, tg, res, task_deadzone(tk, kp));
                }
                for (int a = 0; a < tk.r_t; a++) evec[s * P.r + tk.r_off + a] = res[a];
                if (c == 0) {
                    double* Jo = BW.Jt + (((size_t)bl * S + s) * P.r + tk.r_off) * n;
                    task_jacobian(P, tk, ee, Jo, n);
                }
            }
            // joint-limit terms of the row come from the state BEFORE the step that produced it (System.cpp:158, Q4)
            // (a SequentialSystem has its own limits disabled, SequentialSystem.cpp:16-17, and fpBatch never asks its
            //  sub-systems: only a plain single system contributes here)
            for (int i = 0; i < n; i++) {
                double ld = 0.0, qv = 0.0;
                if (row > 0 && P.n_tasks == 1)
                    for (int ti = 0; ti < P.n_tasks; ti++) {
                        const DevTask& tk = P.tasks[ti];
                        if (!tk.limits_set) continue;
                        bool act;
                        const double q = limit_q(tk, i, xprev[i], act);
                        if (act) {
                            ld += 1.0;
                            qv = q;  // one limit-carrying system per state entry (Sequential systems disable their own limits)
                        }
                    }
                qLq += (qv * ld) * qv;
                if (c == 0) {
                    BW.Ld[((size_t)bl * S + s) * n + i] = ld;
                    BW.ql[((size_t)bl * S + s) * n + i] = qv;
                }
            }
            s++;
        }
        if (row == H - 1) break;
        for (int i = 0; i < m; i++) {
            uk[i] = u[(size_t)row * m + i] + alpha * du[(size_t)row * m + i];
            uRu += (uk[i] * P.r_diag[i]) * uk[i];
        }
        for (int i = 0; i < n; i++) xprev[i] = x[i];
        dynamics_step(P, x, uk);
    }
    // e^T Q e over the keypoint slots (duplicates counted as often as they appear, Q16)
    const int Sr = BW.Sr, rr = P.r;
    double eQe = 0.0;
    if (BW.Qfull) {
        for (int j = 0; j < Sr; j++) {
            double v = 0.0;
            for (int i = 0; i < Sr; i++) v += evec[BW.vds[i / rr] * rr + i % rr] * BW.Qfull[(size_t)i * Sr + j];
            eQe += v * evec[BW.vds[j / rr] * rr + j % rr];
        }
    } else {
        for (int ss = 0; ss < BW.nk; ss++)
            for (int ti = 0; ti < P.n_tasks; ti++) {
                const DevTask& tk = P.tasks[ti];
                const int kp = task_kp_at(tk, BW.vp[ss]);
                if (kp < 0) continue;
                const int r = tk.r_t;
                const double* Q = tk.kp_precision + ((size_t)(tk.precision_per_problem ? b : 0) * tk.n_keypoints + kp) * r * r;
                const double* ev = evec + BW.vds[ss] * rr + tk.r_off;
                for (int j = 0; j < r; j++) {
                    double v = 0.0;
                    for (int i = 0; i < r; i++) v += ev[i] * Q[i * r + j];
                    eQe += v * ev[j];
                }
            }
    }
    BW.cc[(size_t)b * BNC + c] = (eQe + uRu) + qLq;
    if (c == 0)
        for (int i = 0; i < Sr; i++) BW.e[(size_t)bl * Sr + i] = evec[BW.vds[i / rr] * rr + i % rr];
}

// ------------------------------------------------------------------------------------------------------------
// block-diagonal precision of the keypoint rows (System::getQMatrix(true), System.cpp:341-361 / SequentialSystem.cpp:229-272)
__device__ __forceinline__ double batch_q_elem(const DevProblem& P, const BatchDev& BW, int b, int i, int j) {
    if (BW.Qfull) return BW.Qfull[(size_t)i * BW.Sr + j];
    const int si = i / P.r, sj = j / P.r;
    if (si != sj) return 0.0;
    const int ci = i % P.r, cj = j % P.r;
    for (int ti = 0; ti < P.n_tasks; ti++) {
        const DevTask& tk = P.tasks[ti];
        if (ci >= tk.r_off && ci < tk.r_off + tk.r_t) {
            if (cj < tk.r_off || cj >= tk.r_off + tk.r_t) return 0.0;
            const int kp = task_kp_at(tk, BW.vp[si]);
            if (kp < 0) return 0.0;
            const double* Q = tk.kp_precision + ((size_t)(tk.precision_per_problem ? b : 0) * tk.n_keypoints + kp) * tk.r_t * tk.r_t;
            return Q[(ci - tk.r_off) * tk.r_t + (cj - tk.r_off)];
        }
    }
    return 0.0;
}

// W = J^T Q J + L, g = J^T Q e + L ql (BatchILQR.cpp:130-131); Su, T = W Su, rhs = Su^T g - R u.  One block per problem.
// Shared memory: JQ [Sn][Sr], W [Sn][Sn], g [Sn], cs [H] (prefix sums of dt).
__global__ void __launch_bounds__(256) k_batch_assemble(const DevProblem* __restrict__ Pp, Work W, BatchDev BW, int b0) {
    extern __shared__ double sm[];
    const DevProblem& P = *Pp;
    const int bl = blockIdx.x, b = b0 + bl;
    if (W.done[b]) return;
    const int n = P.n, m = P.m, r = P.r, S = BW.S, Sn = BW.Sn, Sr = BW.Sr, N = BW.N, Np = BW.Np;
    double* sJQ = sm;             // [Sn][Sr]
    double* sW = sJQ + Sn * Sr;   // [Sn][Sn]
    double* sg = sW + Sn * Sn;    // [Sn]
    double* scs = sg + Sn;        // [H]
    const double* Jt = BW.Jt + (size_t)bl * S * r * n;
    const double* e = BW.e + (size_t)bl * Sr;
    const double* Ld = BW.Ld + (size_t)bl * Sn;
    const double* ql = BW.ql + (size_t)bl * Sn;
    for (int idx = threadIdx.x; idx < Sn * Sr; idx += blockDim.x) {
        const int ia = idx / Sr, j = idx % Sr, s = ia / n, a = ia % n;
        double v = 0.0;
        for (int c = 0; c < r; c++) v += Jt[((size_t)s * r + c) * n + a] * batch_q_elem(P, BW, b, s * r + c, j);
        sJQ[idx] = v;
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < Sn * Sn; idx += blockDim.x) {
        const int ia = idx / Sn, jb = idx % Sn, s2 = jb / n, bb = jb % n;
        double v = 0.0;
        for (int c = 0; c < r; c++) v += sJQ[ia * Sr + s2 * r + c] * Jt[((size_t)s2 * r + c) * n + bb];
        if (ia == jb) v += Ld[ia];
        sW[idx] = v;
        BW.Wm[(size_t)bl * Sn * Sn + idx] = v;
    }
    for (int ia = threadIdx.x; ia < Sn; ia += blockDim.x) {
        double v = 0.0;
        for (int j = 0; j < Sr; j++) v += sJQ[ia * Sr + j] * e[j];
        v += Ld[ia] * ql[ia];
        sg[ia] = v;
        BW.g[(size_t)bl * Sn + ia] = v;
    }
    __syncthreads();
    // columns of Su / T / rhs.  Row block s of Su is [0, A..A B_1, .., B_{t_s - 1}, 0 ..]: column block j = control u_j holds
    // the B of control u_{j-1} (BatchILQR.cpp:71-94 with ABJLs[0].B = 0, Q3)
    const double* u = BW.u + (size_t)b * N;
    const double* X = BW.X + (size_t)bl * P.H * n;
    if (threadIdx.x == 0) {
        double acc = 0.0;
        scs[0] = 0.0;
        for (int l = 1; l < P.H; l++) {
            const double sq = P.time ? u[(size_t)(l - 1) * m + m - 1] : 0.0;
            acc += P.time ? sq * sq : P.dt;
            scs[l] = acc;
        }
    }
    __syncthreads();
    double* Su = BW.Su + (size_t)bl * BW.SnP * Np;
    double* T = BW.T + (size_t)bl * BW.SnP * Np;
    for (int col = threadIdx.x; col < Np; col += blockDim.x) {
        const int j = col / m, c = col % m;
        double rhs = 0.0;
        if (col < N) {
            for (int ia = 0; ia < Sn; ia++) {
                const int s = ia / n, a = ia % n;
                double v = 0.0;
                if (j >= 1 && j <= P.kp_steps[s] - 1) v = batch_Su_entry(P, u, X, scs, P.kp_steps[s], j, a, c);
                Su[(size_t)ia * Np + col] = v;
                rhs += v * sg[ia];
            }
            for (int ia = 0; ia < Sn; ia++) {
                double tv = 0.0;
                for (int kb = 0; kb < Sn; kb++) tv += sW[ia * Sn + kb] * Su[(size_t)kb * Np + col];  // this thread's own column
                T[(size_t)ia * Np + col] = tv;
            }
            rhs -= P.r_diag[c] * u[col];
        } else {
            for (int ia = 0; ia < Sn; ia++) {
                Su[(size_t)ia * Np + col] = 0.0;
                T[(size_t)ia * Np + col] = 0.0;
            }
        }
        BW.rhs[(size_t)bl * Np + col] = rhs;
    }
}

// ------------------------------------------------------------------------------------------------------------
// lstq_A = Su^T T + R  (BatchILQR.cpp:130), lower triangle incl. diagonal blocks, FP64 tensor pipe.
// A is row-major [Np][Np]; padded rows/columns (>= N) get the identity so that the factorisation passes through them.
// One warp per 8 x 64 strip of the output: the A fragments (Su^T rows) are reused across the eight column tiles.
__global__ void __launch_bounds__(128) k_batch_syrk(const DevProblem* __restrict__ Pp, Work W, BatchDev BW, int b0) {
    const DevProblem& P = *Pp;
    const int bl = blockIdx.y, b = b0 + bl;
    if (W.done[b]) return;
    const int Np = BW.Np, N = BW.N, SnP = BW.SnP, m = P.m;
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int strips = Np / 64, rtiles = Np / 8;
    const int wid = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (wid >= rtiles * strips) return;
    const int it = wid / strips, js = wid % strips;
    const int i0 = it * 8, j0 = js * 64;
    if (j0 > i0) return;  // strictly upper strip: never read
    const double* Su = BW.Su + (size_t)bl * SnP * Np;
    const double* T = BW.T + (size_t)bl * SnP * Np;
    double* A = BW.A + (size_t)bl * Np * Np;
    double c0[8], c1[8];
#pragma unroll
    for (int ct = 0; ct < 8; ct++) c0[ct] = c1[ct] = 0.0;
    for (int ks = 0; ks < SnP / 4; ks++) {
        const double a = Su[(size_t)(4 * ks + t) * Np + i0 + g];  // (Su^T)[i0+g][4ks+t]
#pragma unroll
        for (int ct = 0; ct < 8; ct++) {
            const double bb = T[(size_t)(4 * ks + t) * Np + j0 + 8 * ct + g];
            dmma884(c0[ct], c1[ct], a, bb);
        }
    }
    const int i = i0 + g;
#pragma unroll
    for (int ct = 0; ct < 8; ct++) {
        const int j = j0 + 8 * ct + 2 * t;
        double v0 = c0[ct], v1 = c1[ct];
        if (i == j) v0 += (i < N) ? P.r_diag[i % m] : 1.0;
        if (i == j + 1) v1 += (i < N) ? P.r_diag[i % m] : 1.0;
        *reinterpret_cast<double2*>(A + (size_t)i * Np + j) = make_double2(v0, v1);
    }
}

// ------------------------------------------------------------------------------------------------------------
// In-place blocked Cholesky A = L L^T (lower, row-major) of one problem per CTA: 256 threads, block size 64.
//   per panel k0:  D = A[k0:k0+64, k0:k0+64] factored in shared memory (right-looking, 64 steps);
//                  Linv = D^-1 (lower) by 64 independent forward substitutions;
//                  panel rows  L21 = A21 Linv^T            (DMMA, C = X Y^T form)
//                  trailing    A22 -= L21 L21^T  (lower)   (DMMA, same form, read-modify-write through L2)
// Replaces `lstq_A.inverse()` (Eigen PartialPivLU, BatchILQR.cpp:132): lstq_A = R + Su^T W Su is symmetric positive
// definite (R > 0).  A non-positive pivot sets ILQRB_ST_QUU_NOT_SPD and the factorisation carries on with NaNs.
constexpr int CH_NB = 64, CH_LD = 68;

// C(64x64 at Cg, leading dim ldc) = beta_old * C + sign * X Y^T with X, Y 64 x 64 in shared memory (leading dim CH_LD).
// 8 warps: warp w owns rows 8w..8w+7, all eight column tiles.  lower_only: skip tiles strictly above the diagonal.
template <bool ACCUM, int NCT = 8>
__device__ __forceinline__ void ch_xyT(double* Cg, int ldc, const double* sX, const double* sY, double sign, bool lower_only, int warp, int lane, int ct0 = 0) {
    const int g = lane >> 2, t = lane & 3;
    double a[16];
#pragma unroll
    for (int ks = 0; ks < 16; ks++) a[ks] = sX[(8 * warp + g) * CH_LD + 4 * ks + t];
#pragma unroll
    for (int cti = 0; cti < NCT; cti++) {
        const int ct = ct0 + cti;
        if (lower_only && ct > warp) continue;
        double c0 = 0.0, c1 = 0.0;
#pragma unroll
        for (int ks = 0; ks < 16; ks++) dmma884(c0, c1, a[ks], sY[(8 * ct + g) * CH_LD + 4 * ks + t]);
        double2* dst = reinterpret_cast<double2*>(Cg + (size_t)(8 * warp + g) * ldc + 8 * ct + 2 * t);
        if (ACCUM) {
            const double2 o = *dst;
            *dst = make_double2(o.x + sign * c0, o.y + sign * c1);
        } else {
            *dst = make_double2(sign * c0, sign * c1);
        }
    }
}
__device__ __forceinline__ void ch_load64(double* sdst, const double* Ag, int lda, int tid) {
    for (int e = tid; e < 64 * 32; e += 256) {
        const int r = e >> 5, c2 = e & 31;
        const double2 v = *reinterpret_cast<const double2*>(Ag + (size_t)r * lda + 2 * c2);
        sdst[r * CH_LD + 2 * c2] = v.x;
        sdst[r * CH_LD + 2 * c2 + 1] = v.y;
    }
}

__global__ void __launch_bounds__(256) k_batch_chol(Work W, BatchDev BW, int b0) {
    extern __shared__ double sm[];
    double* sD = sm;                    // diagonal block / its factor
    double* sX = sD + 64 * CH_LD;       // row block operand
    double* sY = sX + 64 * CH_LD;       // Linv, then column block operand
    const int bl = blockIdx.x, b = b0 + bl;
    if (W.done[b]) return;
    const int Np = BW.Np, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    double* A = BW.A + (size_t)bl * Np * Np;
    __shared__ int s_bad;
    if (tid == 0) s_bad = 0;
    for (int k0 = 0; k0 < Np; k0 += CH_NB) {
        __syncthreads();
        ch_load64(sD, A + (size_t)k0 * Np + k0, Np, tid);
        __syncthreads();
        // factor the diagonal block (lower part)
        for (int j = 0; j < 64; j++) {
            __syncthreads();  // rank-1 update of step j-1 complete
            const double djj = sD[j * CH_LD + j];
            __syncthreads();  // everybody holds the pivot before it is overwritten
            const double d = sqrt(djj);
            if (tid == 0) {
                if (!(djj > 0.0)) s_bad = 1;
                sD[j * CH_LD + j] = d;
            }
            if (tid > j && tid < 64) sD[tid * CH_LD + j] /= d;
            __syncthreads();
            // rank-1 update of the trailing lower triangle: elements (i, k) with j < k <= i < 64
            for (int e = tid; e < 64 * 64; e += 256) {
                const int i = e >> 6, k = e & 63;
                if (k > j && k <= i) sD[i * CH_LD + k] -= sD[i * CH_LD + j] * sD[k * CH_LD + j];
            }
        }
        __syncthreads();
        // write the factor back (upper part of the block zeroed) ; Linv into sY by forward substitution, one column per thread
        for (int e = tid; e < 64 * 64; e += 256) {
            const int i = e >> 6, k = e & 63;
            const double v = (k <= i) ? sD[i * CH_LD + k] : 0.0;
            sD[i * CH_LD + k] = v;
            A[(size_t)(k0 + i) * Np + k0 + k] = v;
        }
        __syncthreads();
        if (tid < 64) {
            const int j = tid;  // column j of the inverse: L x = e_j
            for (int i = 0; i < 64; i++) {
                double acc = (i == j) ? 1.0 : 0.0;
                for (int k = j; k < i; k++) acc -= sD[i * CH_LD + k] * sY[k * CH_LD + j];
                sY[i * CH_LD + j] = (i < j) ? 0.0 : acc / sD[i * CH_LD + i];
            }
        }
        __syncthreads();
        // panel: L21 = A21 Linv^T, 64 rows at a time
        for (int I = k0 + CH_NB; I < Np; I += 64) {
            ch_load64(sX, A + (size_t)I * Np + k0, Np, tid);
            __syncthreads();
            ch_xyT<false>(A + (size_t)I * Np + k0, Np, sX, sY, 1.0, false, warp, lane);
            __syncthreads();
        }
        // trailing update: A22 -= L21 L21^T (lower block triangle)
        for (int I = k0 + CH_NB; I < Np; I += 64) {
            ch_load64(sX, A + (size_t)I * Np + k0, Np, tid);
            for (int J = k0 + CH_NB; J <= I; J += 64) {
                __syncthreads();
                ch_load64(sY, A + (size_t)J * Np + k0, Np, tid);
                __syncthreads();
                ch_xyT<true>(A + (size_t)I * Np + J, Np, sX, sY, -1.0, J == I, warp, lane);
            }
            __syncthreads();
        }
    }
    __syncthreads();
    if (tid == 0 && s_bad) W.status[b] |= ILQRB_ST_QUU_NOT_SPD;
}

// Left-looking variant of the same factorisation (default).  The right-looking kernel above reads and rewrites every
// trailing 64x64 block once per panel (96 KB of traffic per 0.5 Mflop block product; at C4, N = 1592, that is 250 MB per
// problem and iteration with 6 GB of matrices in flight, i.e. HBM traffic, next to the DMMA work).  Here block (I, J)
// accumulates  A(I,J) - sum_k L(I,k) L(J,k)^T  in registers over all finished panels k < J (two 32 KB operand loads per
// block product, the row panel L(J,k) shared by all I of the column), is read and written once, then the column is
// finished: diagonal block factored in shared memory, its inverse kept there, L(I,J) = C Linv^T for the blocks below.
__global__ void __launch_bounds__(256, 2) k_batch_chol_ll(Work W, BatchDev BW, int b0) {
    extern __shared__ double sm[];
    double* sD = sm;                    // diagonal block, its factor, then the factor's inverse for the rest of the column
    double* sX = sD + 64 * CH_LD;       // L(I,k) ; then the updated block C
    double* sY = sX + 64 * CH_LD;       // L(J,k)
    const int bl = blockIdx.x, b = b0 + bl;
    if (W.done[b]) return;
    const int Np = BW.Np, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t = lane & 3;
    double* A = BW.A + (size_t)bl * Np * Np;
    __shared__ int s_bad;
    if (tid == 0) s_bad = 0;
    for (int J0 = 0; J0 < Np; J0 += CH_NB) {
        for (int I0 = J0; I0 < Np; I0 += 64) {
            const bool diag = (I0 == J0);
            double c[8][2];
#pragma unroll
            for (int ct = 0; ct < 8; ct++) c[ct][0] = c[ct][1] = 0.0;
            for (int k0 = 0; k0 < J0; k0 += CH_NB) {
                __syncthreads();  // the previous product is done with sX, sY
                ch_load64(sX, A + (size_t)I0 * Np + k0, Np, tid);
                if (!diag) ch_load64(sY, A + (size_t)J0 * Np + k0, Np, tid);
                __syncthreads();
                const double* sB = diag ? sX : sY;
                double a[16];
#pragma unroll
                for (int ks = 0; ks < 16; ks++) a[ks] = sX[(8 * warp + g) * CH_LD + 4 * ks + t];
#pragma unroll
                for (int ct = 0; ct < 8; ct++) {
                    if (diag && ct > warp) continue;  // lower triangle only
#pragma unroll
                    for (int ks = 0; ks < 16; ks++) dmma884(c[ct][0], c[ct][1], a[ks], sB[(8 * ct + g) * CH_LD + 4 * ks + t]);
                }
            }
            __syncthreads();
            // C = A(I,J) - sum, into shared memory (the diagonal block into sD, the others into sX)
            double* sC = diag ? sD : sX;
#pragma unroll
            for (int ct = 0; ct < 8; ct++) {
                const double2 o = *reinterpret_cast<const double2*>(A + (size_t)(I0 + 8 * warp + g) * Np + J0 + 8 * ct + 2 * t);
                sC[(8 * warp + g) * CH_LD + 8 * ct + 2 * t] = o.x - c[ct][0];
                sC[(8 * warp + g) * CH_LD + 8 * ct + 2 * t + 1] = o.y - c[ct][1];
            }
            __syncthreads();
            if (diag) {
                // factor the diagonal block (lower part)
                for (int j = 0; j < 64; j++) {
                    __syncthreads();  // rank-1 update of step j-1 complete
                    const double djj = sD[j * CH_LD + j];
                    __syncthreads();  // everybody holds the pivot before it is overwritten
                    const double d = sqrt(djj);
                    if (tid == 0) {
                        if (!(djj > 0.0)) s_bad = 1;
                        sD[j * CH_LD + j] = d;
                    }
                    if (tid > j && tid < 64) sD[tid * CH_LD + j] /= d;
                    __syncthreads();
                    for (int e = tid; e < 64 * 64; e += 256) {
                        const int i = e >> 6, k = e & 63;
                        if (k > j && k <= i) sD[i * CH_LD + k] -= sD[i * CH_LD + j] * sD[k * CH_LD + j];
                    }
                }
                __syncthreads();
                // write the factor back (upper part of the block zeroed) ; Linv into sY by forward substitution, one column per thread
                for (int e = tid; e < 64 * 64; e += 256) {
                    const int i = e >> 6, k = e & 63;
                    const double v = (k <= i) ? sD[i * CH_LD + k] : 0.0;
                    sD[i * CH_LD + k] = v;
                    A[(size_t)(J0 + i) * Np + J0 + k] = v;
                }
                __syncthreads();
                if (tid < 64) {
                    const int j = tid;  // column j of the inverse: L x = e_j
                    for (int i = 0; i < 64; i++) {
                        double acc = (i == j) ? 1.0 : 0.0;
                        for (int k = j; k < i; k++) acc -= sD[i * CH_LD + k] * sY[k * CH_LD + j];
                        sY[i * CH_LD + j] = (i < j) ? 0.0 : acc / sD[i * CH_LD + i];
                    }
                }
                __syncthreads();
                for (int e = tid; e < 64 * 64; e += 256) sD[(e >> 6) * CH_LD + (e & 63)] = sY[(e >> 6) * CH_LD + (e & 63)];
                __syncthreads();
            } else {
                // L(I,J) = C Linv^T
                ch_xyT<false>(A + (size_t)I0 * Np + J0, Np, sX, sD, 1.0, false, warp, lane);
            }
        }
    }
    __syncthreads();
    if (tid == 0 && s_bad) W.status[b] |= ILQRB_ST_QUU_NOT_SPD;
}

// ------------------------------------------------------------------------------------------------------------
// The same left-looking factorisation with the operand tiles moved by the TMA engine (default).  k_batch_chol_ll above stages every
// 64 x 64 operand tile global -> registers -> shared memory with two __syncthreads per block product (ncu: 3.8 warps per issue
// stalled at the barrier, DMMA pipe 56 % active).  Here L(I,k) and L(J,k) arrive by `cp.async.bulk.tensor.3d` (SASS UTMALDG) through
// a tensor map over the chunk's [problem][row][column] matrices into a two-stage shared-memory ring, completion on one mbarrier
// per stage: thread 0 issues the tiles of product k+1 before the CTA computes product k, one __syncthreads per product (the ring
// slot's readers are done before it is refilled).  Tiles land as four [64][16] boxes in the 128-byte swizzle, which makes the
// DMMA fragment reads (8 rows x 4 column pairs per warp) conflict-free without the 68-double row padding the register path needs.
// 198 KB of shared memory: one CTA of 8 warps per SM (the register-staged kernel runs two); the copies no longer wait for
// the warps and the warps no longer wait for the copies.
constexpr int CHT_TILE = 64 * 64;  // doubles per operand tile (4 boxes of 64 x 16)

__device__ __forceinline__ void tma_load_box(double* dst, const CUtensorMap* tm, int col, int row, int prob, unsigned long long* bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"((unsigned)__cvta_generic_to_shared(dst)),
                 "l"(tm), "r"(col), "r"(row), "r"(prob), "r"((unsigned)__cvta_generic_to_shared(bar))
                 : "memory");
}
__device__ __forceinline__ void cht_mbar_init(unsigned long long* bar) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)));
}
__device__ __forceinline__ void cht_expect(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
// bounded wait: a copy that never lands (bad tensor map) traps instead of hanging the device
__device__ __forceinline__ void cht_wait(unsigned long long* bar, unsigned parity) {
    unsigned done = 0;
    for (long long spin = 0; !done; spin++) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"((unsigned)__cvta_generic_to_shared(bar)), "r"(parity)
            : "memory");
        if (spin > (1ll << 26)) __trap();
    }
}

// NW = 8: warp w owns tile rows 8w..8w+7 and all eight column groups; NW = 16: warps w and w + 8 share the rows and split the
// column groups (twice the warps to cover the DMMA and shared-memory latencies at one CTA per SM; the A fragments are read twice).
template <int NW>
__global__ void __launch_bounds__(32 * NW, 1) k_batch_chol_tma(const __grid_constant__ CUtensorMap tmA, const DevProblem* __restrict__ Pp, Work W, BatchDev BW, int b0) {
    extern __shared__ __align__(1024) unsigned char cht_smem[];
    double* ring = reinterpret_cast<double*>(cht_smem);  // [2 stages][X tile, Y tile]
    double* sD = ring + 4 * CHT_TILE;                    // diagonal block, its factor, then the factor's inverse (row stride CH_LD)
    double* sC = sD + 64 * CH_LD;                        // the updated block C ; Linv while the diagonal block is inverted
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(sC + 64 * CH_LD);
    const int bl = blockIdx.x, b = b0 + bl;
    if (W.done[b]) return;
    constexpr int NT = 32 * NW, NCT = 64 / NW;  // threads; column groups per warp (8 or 4)
    const int Np = BW.Np, tid = threadIdx.x, lane = tid & 31;
    const int warp = (tid >> 5) & 7, ct0 = ((tid >> 5) >> 3) * NCT;  // row group ; first column group
    const int g = lane >> 2, t = lane & 3;
    double* A = BW.A + (size_t)bl * Np * Np;
    __shared__ int s_bad;
    if (tid == 0) {
        s_bad = 0;
        cht_mbar_init(&bars[0]);
        cht_mbar_init(&bars[1]);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    // element (row, col) of a tile: box col/16, 128-byte rows, 16-byte chunks XOR-ed with row % 8 (CU_TENSOR_MAP_SWIZZLE_128B).
    // A 64-bit shared-memory load is served per half-warp (lanes g = 0..3 and g = 4..7): with fragment row g = tile row g the four
    // rows of a half differ in two bits only and the XOR maps both column pairs of a lane quad onto the same four chunks (ncu: 46 %
    // of the wavefronts were conflicts).  Fragment row g therefore stands for tile row pg = 2 (g % 4) + g / 4 — rows 0, 2, 4, 6 in
    // one half, 1, 3, 5, 7 in the other — for the A rows, the B rows (= output columns) and the output alike.
    const int pg = 2 * (g & 3) + (g >> 2);
    int sw[4];
#pragma unroll
    for (int q = 0; q < 4; q++) sw[q] = (((2 * q + (t >> 1)) ^ pg) << 1) + (t & 1);
    const int arow = (8 * warp + pg) * 16;
    const int pc0 = 2 * ((2 * t) & 3) + ((2 * t) >> 2), pc1 = 2 * ((2 * t + 1) & 3) + ((2 * t + 1) >> 2);  // output columns of slots 0, 1
    auto issue = [&](int stage, int I0, int J0, int k0, bool diag) {  // thread 0
        double* sX = ring + (size_t)stage * 2 * CHT_TILE;
        double* sY = sX + CHT_TILE;
        cht_expect(&bars[stage], (diag ? 1u : 2u) * CHT_TILE * 8u);
#pragma unroll
        for (int bx = 0; bx < 4; bx++) tma_load_box(sX + bx * 1024, &tmA, k0 + 16 * bx, I0, bl, &bars[stage]);
        if (!diag) {
#pragma unroll
            for (int bx = 0; bx < 4; bx++) tma_load_box(sY + bx * 1024, &tmA, k0 + 16 * bx, J0, bl, &bars[stage]);
        }
    };
    unsigned cnt = 0;  // block products so far: stage = cnt % 2, phase parity = (cnt / 2) % 2
    bool prefetched = false;  // the first tiles of the coming block are already on their way
    for (int J0 = 0; J0 < Np; J0 += CH_NB) {
        // the blocks of column J0 - 64 were written with ordinary stores: order them before the copy engine's reads
        asm volatile("fence.proxy.async;" ::: "memory");
        __syncthreads();
        for (int I0 = J0; I0 < Np; I0 += 64) {
            const bool diag = (I0 == J0);
            double c[NCT][2];
#pragma unroll
            for (int ct = 0; ct < NCT; ct++) c[ct][0] = c[ct][1] = 0.0;
            const int nk = J0 / CH_NB;
            if (nk > 0 && tid == 0 && !prefetched) issue(cnt & 1, I0, J0, 0, diag);
            prefetched = false;
            for (int kk = 0; kk < nk; kk++, cnt++) {
                if (tid == 0 && kk + 1 < nk) issue((cnt + 1) & 1, I0, J0, (kk + 1) * CH_NB, diag);
                const double* sX = ring + (size_t)(cnt & 1) * 2 * CHT_TILE;
                const double* sB = diag ? sX : sX + CHT_TILE;
                cht_wait(&bars[cnt & 1], (cnt >> 1) & 1);
                double a[16];
#pragma unroll
                for (int ks = 0; ks < 16; ks++) a[ks] = sX[(ks >> 2) * 1024 + arow + sw[ks & 3]];
#pragma unroll
                for (int ct = 0; ct < NCT; ct++) {
                    if (diag && ct0 + ct > warp) continue;  // lower triangle only
#pragma unroll
                    for (int ks = 0; ks < 16; ks++) dmma884(c[ct][0], c[ct][1], a[ks], sB[(ks >> 2) * 1024 + (8 * (ct0 + ct) + pg) * 16 + sw[ks & 3]]);
                }
                __syncthreads();  // the slot's readers are done: it is refilled by the issue of the next iteration
            }
            // the first operand tiles of the next block of this column travel while this block is finished (its slot's readers passed
            // the barrier of product cnt - 2; a column's first block is not prefetched: it reads blocks this column is still writing)
            if (nk > 0 && I0 + 64 < Np) {
                if (tid == 0) issue(cnt & 1, I0 + 64, J0, 0, false);
                prefetched = true;
            }
            // C = A(I,J) - sum, into shared memory (the diagonal block into sD, the others into sC)
            // lstq_A(I,J) = Su(:,I)^T T(:,J) + R (BatchILQR.cpp:130; identity on the padding) is formed here, from the rank-SnP factors,
            // instead of being written by k_batch_syrk and read back (2 x 20 MB per problem and iteration at C4)
            double* sCC = diag ? sD : sC;
            {
                const int SnP = BW.SnP, N = BW.N, m = Pp->m;
                const double* Su = BW.Su + (size_t)bl * SnP * Np + I0 + 8 * warp + pg;
                const double* Tm = BW.T + (size_t)bl * SnP * Np + J0 + pg;
                double e[NCT][2];
#pragma unroll
                for (int ct = 0; ct < NCT; ct++) e[ct][0] = e[ct][1] = 0.0;
                for (int ks = 0; ks < SnP / 4; ks++) {
                    const double av = Su[(size_t)(4 * ks + t) * Np];
#pragma unroll
                    for (int ct = 0; ct < NCT; ct++) dmma884(e[ct][0], e[ct][1], av, Tm[(size_t)(4 * ks + t) * Np + 8 * (ct0 + ct)]);
                }
                const int gi = I0 + 8 * warp + pg;
                double* Crow = sCC + (8 * warp + pg) * CH_LD;
#pragma unroll
                for (int ct = 0; ct < NCT; ct++) {
                    const int j0 = 8 * (ct0 + ct) + pc0, j1 = 8 * (ct0 + ct) + pc1;
                    double v0 = e[ct][0], v1 = e[ct][1];
                    if (gi == J0 + j0) v0 += (gi < N) ? Pp->r_diag[gi % m] : 1.0;
                    if (gi == J0 + j1) v1 += (gi < N) ? Pp->r_diag[gi % m] : 1.0;
                    Crow[j0] = v0 - c[ct][0];
                    Crow[j1] = v1 - c[ct][1];
                }
            }
            __syncthreads();
            if (diag) {
                // factor the diagonal block (lower part)
                for (int j = 0; j < 64; j++) {
                    __syncthreads();  // rank-1 update of step j-1 complete
                    const double djj = sD[j * CH_LD + j];
                    __syncthreads();  // everybody holds the pivot before it is overwritten
                    const double d = sqrt(djj);
                    if (tid == 0) {
                        if (!(djj > 0.0)) s_bad = 1;
                        sD[j * CH_LD + j] = d;
                    }
                    if (tid > j && tid < 64) sD[tid * CH_LD + j] /= d;
                    __syncthreads();
                    for (int e = tid; e < 64 * 64; e += NT) {
                        const int i = e >> 6, k = e & 63;
                        if (k > j && k <= i) sD[i * CH_LD + k] -= sD[i * CH_LD + j] * sD[k * CH_LD + j];
                    }
                }
                __syncthreads();
                // write the factor back (upper part of the block zeroed) ; Linv into sC by forward substitution, one column per thread
                for (int e = tid; e < 64 * 64; e += NT) {
                    const int i = e >> 6, k = e & 63;
                    const double v = (k <= i) ? sD[i * CH_LD + k] : 0.0;
                    sD[i * CH_LD + k] = v;
                    A[(size_t)(J0 + i) * Np + J0 + k] = v;
                }
                __syncthreads();
                if (tid < 64) {
                    const int j = tid;  // column j of the inverse: L x = e_j
                    for (int i = 0; i < 64; i++) {
                        double acc = (i == j) ? 1.0 : 0.0;
                        for (int k = j; k < i; k++) acc -= sD[i * CH_LD + k] * sC[k * CH_LD + j];
                        sC[i * CH_LD + j] = (i < j) ? 0.0 : acc / sD[i * CH_LD + i];
                    }
                }
                __syncthreads();
                for (int e = tid; e < 64 * 64; e += NT) sD[(e >> 6) * CH_LD + (e & 63)] = sC[(e >> 6) * CH_LD + (e & 63)];
                __syncthreads();
            } else {
                // L(I,J) = C Linv^T
                ch_xyT<false, NCT>(A + (size_t)I0 * Np + J0, Np, sC, sD, 1.0, false, warp, lane, ct0);
                __syncthreads();  // sC is rewritten by the next block
            }
        }
    }
    __syncthreads();
    if (tid == 0 && s_bad) W.status[b] |= ILQRB_ST_QUU_NOT_SPD;
}

// tensor map over the chunk's matrices A [nB][Np][Np] (FP64): box = 16 columns x 64 rows x 1 problem, 128-byte swizzle.
// cuTensorMapEncodeTiled comes from the driver through the runtime (no link against libcuda); false = not available.
inline bool cht_make_tensor_map(CUtensorMap* out, const double* A, int nB, int Np) {
    typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_fn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess) fn = (encode_fn)p;
    }
    if (!fn) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)Np, (cuuint64_t)Np, (cuuint64_t)nB};
    const cuuint64_t strides[2] = {(cuuint64_t)Np * 8, (cuuint64_t)Np * Np * 8};
    const cuuint32_t box[3] = {16, 64, 1}, estr[3] = {1, 1, 1};
    return fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<double*>(A), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// ------------------------------------------------------------------------------------------------------------
// du = (L L^T)^-1 rhs : forward then backward substitution, one CTA (256 threads) per problem, 64-wide blocks.
__global__ void __launch_bounds__(256) k_batch_trsv(Work W, BatchDev BW, int b0) {
    extern __shared__ double sm[];
    double* sD = sm;              // [64][CH_LD]
    double* y = sD + 64 * CH_LD;  // [Np]
    const int bl = blockIdx.x, b = b0 + bl;
    if (W.done[b]) return;
    const int Np = BW.Np, N = BW.N, tid = threadIdx.x;
    const double* A = BW.A + (size_t)bl * Np * Np;
    for (int i = tid; i < Np; i += 256) y[i] = BW.rhs[(size_t)bl * Np + i];
    // L z = rhs
    for (int k0 = 0; k0 < Np; k0 += 64) {
        __syncthreads();
        ch_load64(sD, A + (size_t)k0 * Np + k0, Np, tid);
        __syncthreads();
        for (int j = 0; j < 64; j++) {
            if (tid == 0) y[k0 + j] /= sD[j * CH_LD + j];
            __syncthreads();
            if (tid > j && tid < 64) y[k0 + tid] -= sD[tid * CH_LD + j] * y[k0 + j];
            __syncthreads();
        }
        // rows below the block: 8 lanes per row, 64 contiguous bytes each (whole 128-byte lines per row), reduced by shuffles
        {
            const int l8 = tid & 7;
            double yk[8];
#pragma unroll
            for (int q = 0; q < 8; q++) yk[q] = y[k0 + 8 * l8 + q];
            for (int i0 = k0 + 64; i0 < Np; i0 += 32) {  // Np is a multiple of 64: all 32 rows of a pass exist
                const int i = i0 + (tid >> 3);
                const double2* row = reinterpret_cast<const double2*>(A + (size_t)i * Np + k0 + 8 * l8);
                const double2 v0 = row[0], v1 = row[1], v2 = row[2], v3 = row[3];
                double acc = v0.x * yk[0] + v0.y * yk[1] + v1.x * yk[2] + v1.y * yk[3] + v2.x * yk[4] + v2.y * yk[5] + v3.x * yk[6] + v3.y * yk[7];
                acc += __shfl_xor_sync(0xffffffffu, acc, 4);
                acc += __shfl_xor_sync(0xffffffffu, acc, 2);
                acc += __shfl_xor_sync(0xffffffffu, acc, 1);
                if (l8 == 0) y[i] -= acc;
            }
        }
    }
    // L^T x = z
    for (int k0 = Np - 64; k0 >= 0; k0 -= 64) {
        __syncthreads();
        ch_load64(sD, A + (size_t)k0 * Np + k0, Np, tid);
        __syncthreads();
        for (int j = 63; j >= 0; j--) {
            if (tid == 0) y[k0 + j] /= sD[j * CH_LD + j];
            __syncthreads();
            if (tid < j) y[k0 + tid] -= sD[j * CH_LD + tid] * y[k0 + j];
            __syncthreads();
        }
        for (int c = tid; c < k0; c += 256) {
            double acc = 0.0;
#pragma unroll 8
            for (int i = 0; i < 64; i++) acc += A[(size_t)(k0 + i) * Np + c] * y[k0 + i];
            y[c] -= acc;
        }
    }
    __syncthreads();
    for (int i = tid; i < N; i += 256) BW.du[(size_t)b * N + i] = y[i];
}

// ------------------------------------------------------------------------------------------------------------
// control primitives (BatchILQRCP.cpp:129-133): lstq_A = ((Psi^T Su^T) W) Su Psi + (Psi^T R) Psi,
// lstq_B = (Psi^T Su^T) g - (Psi^T R) u, dw = lstq_A^-1 lstq_B (partial-pivot elimination, as PartialPivLU), du = Psi dw.
// One block of 128 threads per problem.  Shared memory: SP [Sn][nw], WS [Sn][nw], Aw [nw][nw+1].
__global__ void __launch_bounds__(128) k_batch_cp(const DevProblem* __restrict__ Pp, Work W, BatchDev BW, int b0) {
    extern __shared__ double sm[];
    const DevProblem& P = *Pp;
    const int bl = blockIdx.x, b = b0 + bl;
    if (W.done[b]) return;
    const int Sn = BW.Sn, nw = BW.nw, N = BW.N, Np = BW.Np, m = P.m, tid = threadIdx.x;
    double* sSP = sm;                 // Su Psi
    double* sWS = sSP + Sn * nw;      // W (Su Psi)
    double* sA = sWS + Sn * nw;       // [nw][nw+1] augmented system
    const double* Su = BW.Su + (size_t)bl * BW.SnP * Np;
    const double* Wm = BW.Wm + (size_t)bl * Sn * Sn;
    const double* gv = BW.g + (size_t)bl * Sn;
    const double* u = BW.u + (size_t)b * N;
    for (int idx = tid; idx < Sn * nw; idx += blockDim.x) {
        const int k = idx / nw, w = idx % nw;
        double v = 0.0;
        for (int col = 0; col < N; col++) v += Su[(size_t)k * Np + col] * BW.psi[(size_t)col * nw + w];
        sSP[idx] = v;
    }
    __syncthreads();
    for (int idx = tid; idx < Sn * nw; idx += blockDim.x) {
        const int k = idx / nw, w = idx % nw;
        double v = 0.0;
        for (int k2 = 0; k2 < Sn; k2++) v += Wm[k * Sn + k2] * sSP[k2 * nw + w];
        sWS[idx] = v;
    }
    __syncthreads();
    for (int idx = tid; idx < nw * (nw + 1); idx += blockDim.x) {
        const int w = idx / (nw + 1), w2 = idx % (nw + 1);
        double v = 0.0;
        if (w2 < nw) {
            for (int k = 0; k < Sn; k++) v += sSP[k * nw + w] * sWS[k * nw + w2];
            v += BW.psiRpsi[w * nw + w2];
        } else {
            for (int k = 0; k < Sn; k++) v += sSP[k * nw + w] * gv[k];
            double ru = 0.0;
            for (int col = 0; col < N; col++) ru += (BW.psi[(size_t)col * nw + w] * P.r_diag[col % m]) * u[col];
            v -= ru;
        }
        sA[idx] = v;
    }
    __syncthreads();
    if (tid == 0) {  // nw is small (number of primitives x m): Gaussian elimination with partial pivoting
        const int ld = nw + 1;
        for (int k = 0; k < nw; k++) {
            int piv = k;
            double best = fabs(sA[k * ld + k]);
            for (int i = k + 1; i < nw; i++)
                if (fabs(sA[i * ld + k]) > best) {
                    best = fabs(sA[i * ld + k]);
                    piv = i;
                }
            if (piv != k)
                for (int j = 0; j < ld; j++) {
                    const double tmp = sA[k * ld + j];
                    sA[k * ld + j] = sA[piv * ld + j];
                    sA[piv * ld + j] = tmp;
                }
            const double d = sA[k * ld + k];
            for (int i = k + 1; i < nw; i++) {
                const double f = sA[i * ld + k] / d;
                for (int j = k; j < ld; j++) sA[i * ld + j] -= f * sA[k * ld + j];
            }
        }
        for (int k = nw - 1; k >= 0; k--) {
            double acc = sA[k * ld + nw];
            for (int j = k + 1; j < nw; j++) acc -= sA[k * ld + j] * sA[j * ld + nw];
            sA[k * ld + nw] = acc / sA[k * ld + k];
        }
    }
    __syncthreads();
    for (int col = tid; col < N; col += blockDim.x) {
        double v = 0.0;
        for (int w = 0; w < nw; w++) v += BW.psi[(size_t)col * nw + w] * sA[w * (nw + 1) + nw];
        BW.du[(size_t)b * N + col] = v;
    }
}

// ------------------------------------------------------------------------------------------------------------
// OPTIONAL fast mode of BatchILQR (ILQRB_BATCH_WOODBURY).  Not the reference's algorithm: lstq_A = R + Su^T W Su is a
// rank-Sn update (Sn = keypoint steps x n, 16 at C4) of the diagonal R, so  du = lstq_A^-1 rhs  follows from the Woodbury identity
//   du = R^-1 (rhs - Su^T y),   (I + W G) y = W z,   G = Su R^-1 Su^T,   z = Su R^-1 rhs
// with one Sn x Sn system instead of the N x N factorisation (N = 1592: 10^4 times fewer flops).  The answer differs from the dense
// path's in its rounding (cond(lstq_A) eps either way); it is parity-checked on its own (tests) and never the reported number.
__global__ void __launch_bounds__(128) k_batch_woodbury(const DevProblem* __restrict__ Pp, Work W, BatchDev BW, int b0) {
    extern __shared__ double sm[];
    const DevProblem& P = *Pp;
    const int bl = blockIdx.x, b = b0 + bl;
    if (W.done[b]) return;
    const int Sn = BW.Sn, N = BW.N, Np = BW.Np, m = P.m, tid = threadIdx.x, ld = Sn + 1;
    double* sG = sm;             // [Sn][Sn]   G, then W G
    double* sM = sG + Sn * Sn;   // [Sn][Sn+1] augmented system (I + W G | W z)
    double* sz = sM + Sn * ld;   // [Sn]
    const double* Su = BW.Su + (size_t)bl * BW.SnP * Np;
    const double* Wm = BW.Wm + (size_t)bl * Sn * Sn;
    const double* rhs = BW.rhs + (size_t)bl * Np;
    for (int idx = tid; idx < Sn * (Sn + 1); idx += blockDim.x) {
        const int k = idx / (Sn + 1), k2 = idx % (Sn + 1);
        double v = 0.0;
        if (k2 < Sn) {
            for (int col = 0; col < N; col++) v += (Su[(size_t)k * Np + col] / P.r_diag[col % m]) * Su[(size_t)k2 * Np + col];
            sG[k * Sn + k2] = v;
        } else {
            for (int col = 0; col < N; col++) v += (Su[(size_t)k * Np + col] / P.r_diag[col % m]) * rhs[col];
            sz[k] = v;
        }
    }
    __syncthreads();
    for (int idx = tid; idx < Sn * ld; idx += blockDim.x) {
        const int k = idx / ld, k2 = idx % ld;
        double v = 0.0;
        if (k2 < Sn) {
            for (int q = 0; q < Sn; q++) v += Wm[k * Sn + q] * sG[q * Sn + k2];
            if (k == k2) v += 1.0;
        } else {
            for (int q = 0; q < Sn; q++) v += Wm[k * Sn + q] * sz[q];
        }
        sM[idx] = v;
    }
    __syncthreads();
    if (tid == 0) {  // Sn is small: Gaussian elimination with partial pivoting
        for (int k = 0; k < Sn; k++) {
            int piv = k;
            double best = fabs(sM[k * ld + k]);
            for (int i = k + 1; i < Sn; i++)
                if (fabs(sM[i * ld + k]) > best) {
                    best = fabs(sM[i * ld + k]);
                    piv = i;
                }
            if (piv != k)
                for (int j = 0; j < ld; j++) {
                    const double tmp = sM[k * ld + j];
                    sM[k * ld + j] = sM[piv * ld + j];
                    sM[piv * ld + j] = tmp;
                }
            const double d = sM[k * ld + k];
            for (int i = k + 1; i < Sn; i++) {
                const double f = sM[i * ld + k] / d;
                for (int j = k; j < ld; j++) sM[i * ld + j] -= f * sM[k * ld + j];
            }
        }
        for (int k = Sn - 1; k >= 0; k--) {
            double acc = sM[k * ld + Sn];
            for (int j = k + 1; j < Sn; j++) acc -= sM[k * ld + j] * sM[j * ld + Sn];
            sM[k * ld + Sn] = acc / sM[k * ld + k];
        }
    }
    __syncthreads();
    for (int col = tid; col < N; col += blockDim.x) {
        double v = rhs[col];
        for (int k = 0; k < Sn; k++) v -= Su[(size_t)k * Np + col] * sM[k * ld + Sn];
        BW.du[(size_t)b * N + col] = v / P.r_diag[col % m];
    }
}

// Su-only variant of the assemble kernel is not needed: k_batch_assemble also serves the CP path (T and rhs unused there).

// ------------------------------------------------------------------------------------------------------------
// line-search decision, accept, trace, early stop (BatchILQR.cpp:135-170).  One thread per problem.
__global__ void k_batch_select(const DevProblem* __restrict__ Pp, Work W, BatchDev BW, int b0, int nB, int it, int nb_iter, int early_stop) {
    const int bl = blockIdx.x * blockDim.x + threadIdx.x;
    if (bl >= nB) return;
    const int b = b0 + bl;
    if (W.done[b]) return;
    const int N = BW.N;
    const double* cc = BW.cc + (size_t)b * BNC;
    const double cost0 = cc[0];
    int sel = NCAND;  // alpha = 2^-10 < 1e-3 is accepted whatever its cost
    for (int c = 1; c <= NCAND; c++)
        if (cc[c] < cost0) {
            sel = c;
            break;
        }
    const double alpha = batch_alpha(sel);
    double* u = BW.u + (size_t)b * N;
    const double* du = BW.du + (size_t)b * N;
    double nrm = 0.0;
    for (int i = 0; i < N; i++) {
        nrm += du[i] * du[i];
        u[i] = u[i] + alpha * du[i];
    }
    W.cost_hist[(size_t)b * nb_iter + it] = cost0;  // the reference prints the PRE-step cost (:160)
    W.alpha_hist[(size_t)b * nb_iter + it] = alpha;
    W.cost0[b] = cost0;
    W.iters[b] = it + 1;
    int st = W.status[b];
    if (isnan(cost0)) st |= ILQRB_ST_NAN_COST;
    if (early_stop && alpha * sqrt(nrm) < 1e-3) {
        st |= ILQRB_ST_EARLY_STOP;
        W.done[b] = 1;
    }
    W.status[b] = st;
}

// ------------------------------------------------------------------------------------------------------------
// host orchestration
__global__ void k_bfill(double* p, size_t n, double v) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}
__global__ void k_bpack_u(double* __restrict__ dst, const double* __restrict__ src, size_t rows, int w, int stride) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * w) return;
    dst[i] = src[(i / w) * stride + (i % w)];
}

#define BCU(call)                                                              \
    do {                                                                       \
        cudaError_t e_ = (call);                                               \
        if (e_ != cudaSuccess) {                                               \
            err = std::string(#call) + ": " + cudaGetErrorString(e_);          \
            return ILQRB_ECUDA;                                                \
        }                                                                      \
    } while (0)

template <class T>
inline int batch_alloc(BatchWork& BW, T** p, size_t count, std::string& err) {
    void* q = nullptr;
    cudaError_t e = cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T));
    if (e != cudaSuccess) {
        err = std::string("cudaMalloc (batch solver workspace): ") + cudaGetErrorString(e);
        return ILQRB_ENOMEM;
    }
    BW.allocs.push_back(q);
    *p = (T*)q;
    return 0;
}
inline void batch_free_impl(BatchWork& BW) {
    for (void* p : BW.allocs) cudaFree(p);
    BW = BatchWork();
}

// solver::BatchILQR::solve / BatchILQRCP::solve for the whole batch, chunk by chunk (the dense N x N workspaces of a
// chunk must fit the workspace budget, ILQRB_BATCH_WS_GB, default 24 GB).  `begin(slot)` / `end()` bracket every launch.
inline int batch_run(cudaStream_t st, int device, const DevProblem& P, const DevProblem* dp, Work& W, BatchWork& BW, const ilqrb_batch_desc* bd,
                     const ilqrb_solve_opts* opts, std::string& err, std::function<void(int)> begin, std::function<void()> end) {
    // keypoint slots (time-sorted, duplicates kept) and their distinct-step index
    std::vector<int> vp;
    for (int t = 0; t < P.n_tasks; t++)
        for (int k = 0; k < P.tasks[t].n_keypoints; k++) vp.push_back(P.tasks[t].kp_timestep[k]);
    std::stable_sort(vp.begin(), vp.end());
    const int nk = (int)vp.size();
    if (nk != P.n_kp_steps && P.n_tasks == 1 && P.any_limits) {
        err = "Batch solvers: a single system with joint limits and several keypoints at the same timestep is not built (SURVEY Q16)";
        return ILQRB_EUNSUPPORTED;
    }
    const int B = P.B, H = P.H, n = P.n, m = P.m, r = P.r, S = P.n_kp_steps;
    const int N = (H - 1) * m, Np = (N + 63) / 64 * 64, Sn = S * n, SnP = (Sn + 3) / 4 * 4, Sr = nk * r;
    const int nw = (bd && bd->psi) ? bd->nw : 0, cp = nw > 0;
    const bool woodbury = bd && (bd->flags & ILQRB_BATCH_WOODBURY) && !cp;
    if (cp && nw > 128) {
        err = "BatchILQRCP: more than 128 primitive weights";
        return ILQRB_EUNSUPPORTED;
    }
    if (S * r > MAX_KP_STEPS * MAX_RT || nk > MAX_KP_STEPS) {
        err = "Batch solvers: too many keypoint rows";
        return ILQRB_EUNSUPPORTED;
    }
    if (BW.allocated && (BW.N != N || BW.nw != nw || BW.S != S || BW.nk != nk)) batch_free_impl(BW);
    int rc;
    if (!BW.allocated) {
        BW.N = N; BW.Np = Np; BW.S = S; BW.Sn = Sn; BW.SnP = SnP; BW.Sr = Sr; BW.nw = nw; BW.cp = cp; BW.nk = nk;
        for (int i = 0; i < nk; i++) {
            BW.vp[i] = vp[i];
            BW.vds[i] = (int)(std::find(P.kp_steps, P.kp_steps + S, vp[i]) - P.kp_steps);
        }
        const char* ev = getenv("ILQRB_BATCH_WS_GB");
        const double budget = (ev ? atof(ev) : 24.0) * 1e9;
        const double per = 8.0 * ((cp ? 0.0 : (double)Np * Np) + 2.0 * SnP * Np + Np + (double)H * n + (double)S * r * n + Sr + 2.0 * Sn + (double)Sn * Sn + Sn + (double)Sn * std::max(nw, 1));
        BW.chunk = (int)std::max(1.0, std::min((double)B, budget / per));
        const size_t C = BW.chunk;
        if ((rc = batch_alloc(BW, &BW.u, (size_t)B * N, err)) || (rc = batch_alloc(BW, &BW.du, (size_t)B * N, err)) || (rc = batch_alloc(BW, &BW.cc, (size_t)B * BNC, err)) ||
            (rc = batch_alloc(BW, &BW.X, C * H * n, err)) || (rc = batch_alloc(BW, &BW.Jt, C * S * r * n, err)) || (rc = batch_alloc(BW, &BW.e, C * Sr, err)) || (rc = batch_alloc(BW, &BW.Ld, C * Sn, err)) ||
            (rc = batch_alloc(BW, &BW.ql, C * Sn, err)) || (rc = batch_alloc(BW, &BW.Wm, C * Sn * Sn, err)) || (rc = batch_alloc(BW, &BW.g, C * Sn, err)) ||
            (rc = batch_alloc(BW, &BW.Su, C * SnP * Np, err)) || (rc = batch_alloc(BW, &BW.T, C * SnP * Np, err)) || (rc = batch_alloc(BW, &BW.rhs, C * Np, err)))
            return rc;
        if (!cp && (rc = batch_alloc(BW, &BW.A, C * Np * Np, err))) return rc;
        if (cp && ((rc = batch_alloc(BW, &BW.psi, (size_t)N * nw, err)) || (rc = batch_alloc(BW, &BW.psiRpsi, (size_t)nw * nw, err)))) return rc;
        if ((rc = batch_alloc(BW, &BW.Qfull, (size_t)Sr * Sr, err))) return rc;
        BCU(cudaMemsetAsync(BW.Su, 0, C * SnP * Np * sizeof(double), st));
        BCU(cudaMemsetAsync(BW.T, 0, C * SnP * Np * sizeof(double), st));
        BW.allocated = true;
    }
    // solver-constructor inputs: Q override (BatchILQR.cpp:27-33), Psi (BatchILQRCP.cpp:21-32)
    BatchDev K = BW;  // kernel argument copy (Qfull pointer nulled when there is no override)
    if (bd && bd->Q) BCU(cudaMemcpyAsync(BW.Qfull, bd->Q, (size_t)Sr * Sr * sizeof(double), cudaMemcpyHostToDevice, st));
    else K.Qfull = nullptr;
    if (cp) {
        BCU(cudaMemcpyAsync(BW.psi, bd->psi, (size_t)N * nw * sizeof(double), cudaMemcpyHostToDevice, st));
        std::vector<double> prp((size_t)nw * nw, 0.0);
        for (int w = 0; w < nw; w++)
            for (int w2 = 0; w2 < nw; w2++) {
                double v = 0.0;
                for (int col = 0; col < N; col++) v += (bd->psi[(size_t)col * nw + w] * P.r_diag[col % m]) * bd->psi[(size_t)col * nw + w2];
                prp[(size_t)w * nw + w2] = v;
            }
        BCU(cudaMemcpyAsync(BW.psiRpsi, prp.data(), prp.size() * sizeof(double), cudaMemcpyHostToDevice, st));
        BCU(cudaStreamSynchronize(st));  // prp is a stack-lifetime buffer
    }
    // state: u <- W.U (padded rows -> compact), traces, flags
    const int nit = std::max(opts->nb_iter, 1);
    {
        const size_t tot = (size_t)B * N;
        k_bpack_u<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(BW.u, W.U, (size_t)B * (H - 1), m, W.ms);
        BCU(cudaMemsetAsync(BW.du, 0, tot * sizeof(double), st));
        const size_t nh = (size_t)B * nit;
        k_bfill<<<(unsigned)((nh + 255) / 256), 256, 0, st>>>(W.cost_hist, nh, NAN);
        k_bfill<<<(unsigned)((nh + 255) / 256), 256, 0, st>>>(W.alpha_hist, nh, NAN);
        BCU(cudaMemsetAsync(W.done, 0, B * sizeof(int), st));
        BCU(cudaMemsetAsync(W.iters, 0, B * sizeof(int), st));
        BCU(cudaMemsetAsync(W.status, 0, B * sizeof(int), st));
        BCU(cudaMemsetAsync(W.cost0, 0, B * sizeof(double), st));
    }
    const size_t smem_chol = 3 * 64 * CH_LD * sizeof(double), smem_trsv = (64 * CH_LD + Np) * sizeof(double);
    const size_t smem_tma = (4 * (size_t)CHT_TILE + 2 * 64 * CH_LD) * sizeof(double) + 16;
    const size_t smem_asm = ((size_t)Sn * Sr + (size_t)Sn * Sn + Sn + H) * sizeof(double), smem_cp = (2 * (size_t)Sn * std::max(nw, 1) + (size_t)nw * (nw + 1)) * sizeof(double);
    static const bool chol_right_looking = getenv("ILQRB_CHOL_RL") != nullptr;  // A/B switch: the first (right-looking) kernel
    static const bool chol_no_tma = getenv("ILQRB_CHOL_NO_TMA") != nullptr;     // A/B switch: the register-staged left-looking kernel
    static const int chol_tma_warps = getenv("ILQRB_CHOL_TMA_WARPS") ? atoi(getenv("ILQRB_CHOL_TMA_WARPS")) : 16;  // 8 | 16 warps per CTA
    CUtensorMap tmA;
    const bool use_tma = !cp && !woodbury && !chol_right_looking && !chol_no_tma && cht_make_tensor_map(&tmA, BW.A, BW.chunk, Np);
    const size_t smem_wb = ((size_t)Sn * Sn + (size_t)Sn * (Sn + 1) + Sn) * sizeof(double);
    if (woodbury && smem_wb > 200 * 1024) {
        err = "BatchILQR Woodbury mode: too many keypoint rows for the shared-memory reduced system";
        return ILQRB_EUNSUPPORTED;
    }
    {
        static DynSmem a_chol, a_chol_ll, a_trsv, a_asm, a_cp, a_tma, a_tma16, a_wb;  // per-device "raised to" values
        auto ens = [&](DynSmem& a, auto kern, size_t smem) -> int {
            if (smem <= 48 * 1024) return 0;
            const bool tracked = device >= 0 && device < MAX_DEVICES;
            if (!tracked || smem > a.raised[device].load(std::memory_order_relaxed)) {
                BCU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                if (tracked) a.raised[device].store(smem, std::memory_order_relaxed);
            }
            return 0;
        };
        if ((rc = ens(a_chol, k_batch_chol, smem_chol)) || (rc = ens(a_chol_ll, k_batch_chol_ll, smem_chol)) || (rc = ens(a_trsv, k_batch_trsv, smem_trsv)) ||
            (rc = ens(a_asm, k_batch_assemble, smem_asm)) || (rc = ens(a_cp, k_batch_cp, smem_cp)) || (use_tma && ((rc = ens(a_tma, k_batch_chol_tma<8>, smem_tma)) || (rc = ens(a_tma16, k_batch_chol_tma<16>, smem_tma)))) ||
            (woodbury && (rc = ens(a_wb, k_batch_woodbury, smem_wb))))
            return rc;
    }
    for (int b0 = 0; b0 < B; b0 += BW.chunk) {
        const int nB = std::min(BW.chunk, B - b0);
        const unsigned fp_grid = (unsigned)(((size_t)nB * BNC + 63) / 64);
        for (int it = 0; it < opts->nb_iter; it++) {
            begin(0);
            k_batch_fp<<<fp_grid, 64, 0, st>>>(dp, W, K, b0, nB, 1);
            end();
            begin(1);
            k_batch_assemble<<<nB, 256, smem_asm, st>>>(dp, W, K, b0);
            end();
            if (woodbury) {
                begin(3);
                k_batch_woodbury<<<nB, 128, smem_wb, st>>>(dp, W, K, b0);
                end();
            } else if (!cp) {
                if (!use_tma) {  // the TMA factorisation forms the blocks of lstq_A itself
                    begin(2);
                    k_batch_syrk<<<dim3((unsigned)((Np / 8) * (Np / 64) + 3) / 4, nB), 128, 0, st>>>(dp, W, K, b0);
                    end();
                }
                begin(3);
                if (use_tma && chol_tma_warps == 8) k_batch_chol_tma<8><<<nB, 256, smem_tma, st>>>(tmA, dp, W, K, b0);
                else if (use_tma) k_batch_chol_tma<16><<<nB, 512, smem_tma, st>>>(tmA, dp, W, K, b0);
                else if (chol_right_looking) k_batch_chol<<<nB, 256, smem_chol, st>>>(W, K, b0);
                else k_batch_chol_ll<<<nB, 256, smem_chol, st>>>(W, K, b0);
                end();
                begin(3);
                k_batch_trsv<<<nB, 256, smem_trsv, st>>>(W, K, b0);
                end();
            } else {
                begin(3);
                k_batch_cp<<<nB, 128, smem_cp, st>>>(dp, W, K, b0);
                end();
            }
            begin(0);
            k_batch_fp<<<fp_grid, 64, 0, st>>>(dp, W, K, b0, nB, 0);
            end();
            begin(4);
            k_batch_select<<<(nB + 127) / 128, 128, 0, st>>>(dp, W, K, b0, nB, it, nit, opts->early_stop);
            end();
            BCU(cudaGetLastError());
        }
    }
    return 0;
}

inline int batch_download(cudaStream_t st, const DevProblem& P, BatchWork& BW, ilqrb_result* out, const Work& W, int nit, std::string& err) {
    const size_t B = P.B;
    if (out->U) BCU(cudaMemcpyAsync(out->U, BW.u, B * BW.N * sizeof(double), cudaMemcpyDefault, st));
    if (out->cost) BCU(cudaMemcpyAsync(out->cost, W.cost0, B * sizeof(double), cudaMemcpyDefault, st));
    if (out->cost_hist) BCU(cudaMemcpyAsync(out->cost_hist, W.cost_hist, B * nit * sizeof(double), cudaMemcpyDefault, st));
    if (out->alpha_hist) BCU(cudaMemcpyAsync(out->alpha_hist, W.alpha_hist, B * nit * sizeof(double), cudaMemcpyDefault, st));
    if (out->iters) BCU(cudaMemcpyAsync(out->iters, W.iters, B * sizeof(int), cudaMemcpyDefault, st));
    if (out->status) BCU(cudaMemcpyAsync(out->status, W.status, B * sizeof(int), cudaMemcpyDefault, st));
    BCU(cudaStreamSynchronize(st));
    return 0;
}

}  // namespace ilqrb
