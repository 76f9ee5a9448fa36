// recursive.cuh — kernels of the Recursive / AL iLQR solve path (ILQRRecursive.cpp:21-181, AL-ILQR.cpp:50-232).
//
//   k_rollout_x, k_stage_costs, k_sum_costs   initial rollout + cost   ILQRRecursive.cpp:41-56
//   k_kp_derivs      l_x, l_xx of the task cost at keypoint steps    System.cpp:248-272,286-308
//   k_backward       Riccati recursion, one problem per warp (DMMA)  ILQRRecursive.cpp:69-97 (+ AL terms AL-ILQR.cpp:110-145)
//   k_linesearch     all 11 alpha candidates rolled out in parallel  ILQRRecursive.cpp:101-155
//   k_accept         sequential-rule selection, accept, early stop   ILQRRecursive.cpp:155-176 (+ AL-ILQR.cpp:189-208)
//   k_final_fx       f(X) of the returned trajectory                 ILQRRecursive.cpp:37,51 (f_X)
#pragma once
#include "kin_lean.cuh"
#include "wmat.cuh"
#include "work.cuh"

namespace ilqrb {

// ------------------------------------------------------------------------------------------------------------
// initial rollout + cost (ILQRRecursive.cpp:41-56) as three launches: the states by one thread per problem (the only serial part:
// step_state), the stage costs by one thread per (problem, timestep) into a scratch (the first B*H doubles of W.Uc, which the line
// search fills later), their sum by one thread per problem in timestep order — the same values added in the same order as a single
// serial pass, which took 370 us per solve at C2 (one thread per problem walking 99 x [stage cost, step] on 32 SMs).
// states: the integrator is independent per joint (SimulationInterface.cpp:19-31: q_i += dt dq_i + dt^2/2 ddq_i ; dq_i += dt ddq_i ;
// t += dt), so one thread per (problem, joint) — plus one for the time state — walks the horizon with step_state's operations;
// TP threads per problem (a power of two >= dof + 1), loads of the controls four steps ahead of the serial chain
template <int TP>
__global__ void __launch_bounds__(128) k_rollout_x(const DevProblem* __restrict__ Pp, Work W, int b0, int nB) {
    const DevProblem& P = *Pp;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int bl = idx / TP, c = idx % TP;
    if (bl >= nB) return;
    const int b = b0 + bl;
    const int H = P.H, dof = P.dof, n = P.n, ns = W.ns, ms = W.ms;
    const bool second = P.nb_deriv == 2, joint = c < dof, tim = P.time && c == dof;
    if (c == 0) {
        W.done[b] = 0;
        W.iters[b] = 0;
        W.alpha_last[b] = 1.0;
    }
    double q = joint ? P.q0[(size_t)b * dof + c] : 0.0;
    double v = (joint && second) ? P.dq0[(size_t)b * dof + c] : 0.0;
    double* Xb = W.X + (size_t)b * H * ns;
    const double* Ub = W.U + (size_t)b * (H - 1) * ms;
    auto put = [&](int k) {
        double* Xk = Xb + (size_t)k * ns;
        if (joint) {
            Xk[c] = q;
            if (second) Xk[dof + c] = v;
        } else if (tim) Xk[n - 1] = q;
        if (c == 0)
            for (int pad = n; pad < ns; pad++) Xk[pad] = 0.0;  // the padding entries of the row
    };
    put(0);
    constexpr int PF = 4;
    for (int k0 = 0; k0 < H - 1; k0 += PF) {
        double uu[PF], ss[PF];
#pragma unroll
        for (int p = 0; p < PF; p++) {
            const int k = k0 + p < H - 1 ? k0 + p : H - 2;
            uu[p] = joint ? Ub[(size_t)k * ms + c] : 0.0;
            ss[p] = P.time ? Ub[(size_t)k * ms + P.m - 1] : 0.0;
        }
#pragma unroll
        for (int p = 0; p < PF; p++) {
            if (k0 + p >= H - 1) break;
            double dt = P.dt;
            if (P.time) dt = __dmul_rn(ss[p], ss[p]);
            const double hdt2 = __dmul_rn(dt, dt) / 2;
            if (joint) {
                if (!second) q = __dadd_rn(q, __dadd_rn(__dmul_rn(dt, uu[p]), __dmul_rn(hdt2, 0.0)));
                else {
                    q = __dadd_rn(q, __dadd_rn(__dmul_rn(dt, v), __dmul_rn(hdt2, uu[p])));
                    v = __dadd_rn(v, __dmul_rn(dt, uu[p]));
                }
            } else if (tim) q = __dadd_rn(q, dt);
            put(k0 + p + 1);
        }
    }
}
template <int N, int M>
__global__ void __launch_bounds__(128) k_stage_costs(const DevProblem* __restrict__ Pp, Work W, int b0, int nB) {
    const DevProblem& P = *Pp;
    const int H = P.H;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nB * H) return;
    const int k = idx / nB, bl = idx - k * nB, b = b0 + bl;  // a warp holds one timestep of 32 problems: the keypoint branch is uniform
    double x[N], u[M];
    const double* Xk = W.X + ((size_t)b * H + k) * W.ns;
#pragma unroll
    for (int i = 0; i < N; i++) x[i] = (i < P.n) ? Xk[i] : 0.0;
    if (k < H - 1) {
        const double* Uk = W.U + ((size_t)b * (H - 1) + k) * W.ms;
#pragma unroll
        for (int i = 0; i < M; i++) u[i] = (i < P.m) ? Uk[i] : 0.0;
    }
    W.Uc[(size_t)b * H + k] = lean_stage_cost<N>(P, b, x, k < H - 1 ? u : nullptr, k);  // the FK sweep the line search's candidate costs use
}
__global__ void __launch_bounds__(128) k_sum_costs(const DevProblem* __restrict__ Pp, Work W, int b0, int nB) {
    const int bl = blockIdx.x * blockDim.x + threadIdx.x;
    if (bl >= nB) return;
    const int b = b0 + bl, H = Pp->H;
    const double* c = W.Uc + (size_t)b * H;
    double cost = 0.0;
    for (int k = 0; k < H; k++) cost += c[k];
    W.cost0[b] = cost;
    W.status[b] = isnan(cost) ? ILQRB_ST_NAN_COST : 0;
}

// ------------------------------------------------------------------------------------------------------------
// keypoint derivatives of the accepted trajectory: one thread per (problem, keypoint step)
//   l_x  = sum_tasks ((-J^T) Q) diff      System.cpp:252-260
//   l_xx = sum_tasks (J^T Q) J            System.cpp:290-296
__global__ void __launch_bounds__(64) k_kp_derivs(const DevProblem* __restrict__ Pp, Work W, int b0, int nB) {
    const DevProblem& P = *Pp;
    const int S = P.n_kp_steps;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nB * S) return;
    const int b = b0 + idx / S, s = idx % S;
    if (W.done[b]) return;
    const int k = P.kp_steps[s], n = P.n, dof = P.dof;
    const double* x = W.X + ((size_t)b * P.H + k) * W.ns;
    double* lx = W.LX + ((size_t)b * S + s) * n;
    double* lxx = W.LXX + ((size_t)b * S + s) * n * n;
    for (int i = 0; i < n; i++) lx[i] = 0.0;
    for (int i = 0; i < n * n; i++) lxx[i] = 0.0;
    double xs[MAXN], dqz[MAXM];
    for (int i = 0; i < n; i++) xs[i] = x[i];
    for (int i = 0; i < dof; i++) dqz[i] = 0.0;
    for (int t = 0; t < P.n_tasks; t++) {
        const DevTask& tk = P.tasks[t];
        const int kp = task_kp_at(tk, k);
        if (kp < 0) continue;
        EEState e;
        task_ee(P, tk, xs, P.nb_deriv == 2 ? xs + dof : dqz, e);
        double fx[MAX_NTT], res[MAX_RT];
        task_fx(P, tk, e, P.time ? xs[n - 1] : 0.0, fx);
        const double* tg = tk.kp_target + ((size_t)b * tk.n_keypoints + kp) * ILQRB_KP_STRIDE;
        task_residual(P, tk, fx, tg, res, task_deadzone(tk, kp));
        const int r = tk.r_t;
        const double* Q = tk.kp_precision + ((size_t)(tk.precision_per_problem ? b : 0) * tk.n_keypoints + kp) * r * r;
        double J[MAX_RT * MAXN];
        task_jacobian(P, tk, e, J, n);
        // column i of J (= row i of J^T):  jq = (J^T Q)[i][:]
        for (int i = 0; i < n; i++) {
            double jq[MAX_RT], njq[MAX_RT];
            for (int c = 0; c < r; c++) {
                double a = 0.0, na = 0.0;
                for (int a2 = 0; a2 < r; a2++) {
                    a += J[a2 * n + i] * Q[a2 * r + c];
                    na += (-1.0 * J[a2 * n + i]) * Q[a2 * r + c];
                }
                jq[c] = a;
                njq[c] = na;
            }
            double g = 0.0;
            for (int c = 0; c < r; c++) g += njq[c] * res[c];
            lx[i] += g;
            for (int j = 0; j < n; j++) {
                double h = 0.0;
                for (int c = 0; c < r; c++) h += jq[c] * J[c * n + j];
                lxx[i * n + j] += h;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------------------
// line search: all NCAND candidates of a problem rolled out by adjacent lanes
//   du = K (x' - x) + alpha d ; u' = u + du ; x'+ = f(x', u') ; cost += l(x', u')      ILQRRecursive.cpp:124-154
template <int N, int M>
__global__ void __launch_bounds__(128) k_linesearch(const DevProblem* __restrict__ Pp, Work W, int line_search) {
    const DevProblem& P = *Pp;
    const int gid = blockIdx.x * blockDim.x + threadIdx.x;
    const int b = gid / CAND_LANES, c = gid % CAND_LANES;
    if (b >= P.B || c >= NCAND) return;
    if (W.done[b]) return;
    if (!line_search && c > 0) return;
    const int H = P.H, n = P.n, m = P.m, ns = W.ns, ms = W.ms;
    const double alpha = alpha_of(c);
    double x[N], u[M];
    load_x0<N>(P, b, x);
    double cost = 0.0, dun = 0.0;
    for (int k = 0; k < H - 1; k++) {
        const size_t bk = (size_t)b * (H - 1) + k;
        const double* Xk = W.X + ((size_t)b * H + k) * ns;
        const double* Kk = W.K + bk * m * ns;
        const double* dk = W.d + bk * ms;
        const double* Uk = W.U + bk * ms;
        double dx[N];
#pragma unroll
        for (int j = 0; j < N; j++) dx[j] = (j < n) ? __dsub_rn(x[j], Xk[j]) : 0.0;
        double nrm2 = 0.0;
#pragma unroll
        for (int i = 0; i < M; i++) {
            u[i] = 0.0;
            if (i < m) {
                double acc = 0.0;
#pragma unroll
                for (int j = 0; j < N; j++)
                    if (j < n) acc += Kk[i * ns + j] * dx[j];
                const double du = acc + alpha * dk[i];
                nrm2 += du * du;
                u[i] = Uk[i] + du;
            }
        }
        dun += sqrt(nrm2);
        double* Uo = W.Uc + (bk * NCAND + c) * ms;
#pragma unroll
        for (int i = 0; i < M; i++)
            if (i < ms) Uo[i] = u[i];
        cost += stage_cost<N>(P, b, x, u, k);
        step_state<N>(P, x, u);
    }
    cost += stage_cost<N>(P, b, x, nullptr, H - 1);
    W.cost_c[(size_t)b * NCAND + c] = cost;
    W.dun_c[(size_t)b * NCAND + c] = dun;
}

// ------------------------------------------------------------------------------------------------------------
// line search for systems with LINEAR dynamics (no time state) and no AL terms: one warp per problem.
// With x+ = A x + B u the candidate of step length alpha is affine in alpha, x'(alpha) = x + alpha dx, u'(alpha) = u + alpha du,
// where (dx, du) is the alpha = 1 closed-loop deviation du_k = K_k dx_k + d_k, dx_{k+1} = A dx_k + B du_k.  The warp rolls
// that deviation out ONCE (lane i owns control row i and joint i), keeps it in shared memory, and the 11 candidates then
// only differ in cost terms: lane c evaluates candidate c at the keypoint steps (every step if some joint limit is violated
// at either end of a segment [x, x + dx]: box limits are convex, so otherwise no candidate violates one).
// Selection follows k_accept's rule; only the SELECTED candidate's controls are written (k_accept applies the same rule
// to the same costs and reads that slot).  Replaces k_linesearch's 11 full rollouts (ILQRRecursive.cpp:101-155).
__host__ __device__ inline int ls_affine_smem_doubles(int H, int ns, int ms) { return (H - 1) * ms + (MAX_KP_STEPS + 2) * ns + MAX_KP_STEPS * NCAND; }

template <int N, int M>
__global__ void __launch_bounds__(128, 3) k_linesearch_affine(const DevProblem* __restrict__ Pp, Work W, int line_search) {
    extern __shared__ double ls_smem[];
    constexpr unsigned FULL = 0xffffffffu;
    const DevProblem& P = *Pp;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int b = blockIdx.x * 4 + wid;
    if (b >= P.B) return;
    if (W.done[b]) return;
    const int H = P.H, n = P.n, m = P.m, ns = W.ns, ms = W.ms, dof = P.dof, S = P.n_kp_steps;
    double* sDU = ls_smem + (size_t)wid * ls_affine_smem_doubles(H, ns, ms);  // [H-1][ms]  du of every step
    double* sKX = sDU + (size_t)(H - 1) * ms;                                  // [S][ns]    dx at the keypoint steps
    double* sX2 = sKX + (size_t)MAX_KP_STEPS * ns;                             // [2][ns]    dx of the current step (double buffer)
    double* sCk = sX2 + 2 * ns;                                                // [S][NCAND] keypoint-step costs
    const double dt = P.dt, hdt2 = __dmul_rn(dt, dt) / 2;
    const double* Kb = W.K + (size_t)b * (H - 1) * m * ns;
    const double* db = W.d + (size_t)b * (H - 1) * ms;
    const double* Xb = W.X + (size_t)b * H * ns;

    // ---- phase A: deviation rollout; limits checked on the fly at both ends of every segment --------------------------
    // lane (g, t) = (control row g, quarter t of the state columns): 4 lanes form one dot product, lane (i, 0) owns joint i
    // (its deviation, its limits).  Gains, d and X rows run through 4-deep register rings so that no global load sits on
    // the serial chain.
    constexpr int PER = N / 4, PF = 4;
    const int g = lane >> 2, t = lane & 3;
    const bool rowl = g < m, owner = rowl && t == 0 && g < dof;
    for (int j = lane; j < 2 * ns; j += 32) sX2[j] = 0.0;
    double dq = 0.0, dv = 0.0;
    bool viol = false;
    // limits of the first limit-carrying task in registers (weight 0 -> +-inf); more than one such task: slow loop
    int l0 = -1, l1 = -1;
    if (P.any_limits)
        for (int ti = P.n_tasks - 1; ti >= 0; ti--)
            if (P.tasks[ti].limits_set) {
                l1 = l0;
                l0 = ti;
            }
    double qmx = INFINITY, qmn = -INFINITY, vmx = INFINITY, vmn = -INFINITY;
    if (l0 >= 0 && owner) {
        const DevTask& tk = P.tasks[l0];
        if (tk.limit_weight[g] != 0) qmx = tk.state_max[g], qmn = tk.state_min[g];
        if (P.nb_deriv == 2 && tk.limit_weight[dof + g] != 0) vmx = tk.state_max[dof + g], vmn = tk.state_min[dof + g];
    }
    auto check = [&](int k, double xq, double xv) {  // owner lanes: state entries g and dof + g of step k
        if (l0 < 0 || !owner) return;
        viol = viol || xq > qmx || xq < qmn || xq + dq > qmx || xq + dq < qmn;
        if (P.nb_deriv == 2) viol = viol || xv > vmx || xv < vmn || xv + dv > vmx || xv + dv < vmn;
        if (l1 >= 0)
            for (int ti = l1; ti < P.n_tasks; ti++) {
                const DevTask& tk = P.tasks[ti];
                if (!tk.limits_set) continue;
                bool a0, a1;
                limit_q(tk, g, xq, a0);
                limit_q(tk, g, xq + dq, a1);
                viol = viol || a0 || a1;
                if (P.nb_deriv == 2) {
                    limit_q(tk, dof + g, xv, a0);
                    limit_q(tk, dof + g, xv + dv, a1);
                    viol = viol || a0 || a1;
                }
            }
    };
    double kr[PF][PER], dr[PF], xqr[PF], xvr[PF];
    auto fetch = [&](int slot, int k) {  // operands of step k: gain row piece, d_g, X_{k+1} entries of the owner
#pragma unroll
        for (int j = 0; j < PER; j++) kr[slot][j] = (rowl && k < H - 1 && PER * t + j < n) ? __ldcs(Kb + ((size_t)k * m + g) * ns + PER * t + j) : 0.0;
        dr[slot] = (rowl && t == 0 && k < H - 1) ? db[(size_t)k * ms + g] : 0.0;
        xqr[slot] = (owner && l0 >= 0 && k < H - 1) ? Xb[(size_t)(k + 1) * ns + g] : 0.0;
        xvr[slot] = (owner && l0 >= 0 && k < H - 1 && P.nb_deriv == 2) ? Xb[(size_t)(k + 1) * ns + dof + g] : 0.0;
    };
#pragma unroll
    for (int r = 0; r < PF; r++) fetch(r, r);
    int sq = 0;  // next keypoint step index
    if (S > 0 && P.kp_steps[0] == 0) {
        for (int j = lane; j < ns; j += 32) sKX[j] = 0.0;
        sq = 1;
    }
    if (owner && l0 >= 0) check(0, Xb[g], P.nb_deriv == 2 ? Xb[dof + g] : 0.0);
    __syncwarp();
    for (int k0 = 0; k0 < H - 1; k0 += PF) {
#pragma unroll
        for (int r = 0; r < PF; r++) {
            const int k = k0 + r;
            if (k < H - 1) {  // warp-uniform
                const double* cur = sX2 + (k & 1) * ns;
                double* nxt = sX2 + ((k + 1) & 1) * ns;
                double acc = 0.0;
#pragma unroll
                for (int j = 0; j < PER; j++) acc += kr[r][j] * ((PER * t + j < n) ? cur[PER * t + j] : 0.0);
                acc += __shfl_xor_sync(FULL, acc, 1);
                acc += __shfl_xor_sync(FULL, acc, 2);
                const double du = acc + dr[r];
                const double xq = xqr[r], xv = xvr[r];
                if (k + 8 < H - 1 && rowl) asm volatile("prefetch.global.L2 [%0];" ::"l"(Kb + ((size_t)(k + 8) * m + g) * ns + PER * t));
                fetch(r, k + PF);
                if (t == 0 && g < ms) sDU[(size_t)k * ms + g] = rowl ? du : 0.0;
                if (owner) {
                    if (P.nb_deriv == 1) {
                        dq = __dadd_rn(dq, __dmul_rn(dt, du));
                        nxt[g] = dq;
                    } else {
                        dq = __dadd_rn(dq, __dadd_rn(__dmul_rn(dt, dv), __dmul_rn(hdt2, du)));
                        dv = __dadd_rn(dv, __dmul_rn(dt, du));
                        nxt[g] = dq;
                        nxt[dof + g] = dv;
                    }
                    check(k + 1, xq, xv);
                }
                __syncwarp();
                if (sq < S && P.kp_steps[sq] == k + 1) {
                    for (int j = lane; j < ns; j += 32) sKX[(size_t)sq * ns + j] = nxt[j];
                    sq++;
                }
            }
        }
    }
    // sum_k |du_k| of the alpha = 1 candidate (early-stop test of k_accept)
    double dsum = 0.0;
    for (int k = lane; k < H - 1; k += 32) {
        double nrm2 = 0.0;
        for (int i = 0; i < m; i++) nrm2 += sDU[(size_t)k * ms + i] * sDU[(size_t)k * ms + i];
        dsum += sqrt(nrm2);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dsum += __shfl_xor_sync(FULL, dsum, o);
    viol = __any_sync(FULL, viol);
    __syncwarp();
    // ---- phase B: candidate costs -----------------------------------------------------------------------------------------
    // no limit violated: only keypoint steps cost anything; lane = (keypoint step, candidate) pair, summed per candidate in
    // step order (the other steps add exact zeros in the reference's running sum).  Otherwise lane = candidate walks every
    // step, re-integrating its deviation from du.
    const bool mine = lane < NCAND && (line_search || lane == 0);
    auto cost_at = [&](int k, double alpha, const double* dx) -> double {
        const double* Xk = Xb + (size_t)k * ns;
        double x[N], u[M];
#pragma unroll
        for (int j = 0; j < N; j++) x[j] = (j < n) ? __dadd_rn(Xk[j], __dmul_rn(alpha, dx[j])) : 0.0;
        if (k < H - 1) {
            const double* Uk = W.U + ((size_t)b * (H - 1) + k) * ms;
#pragma unroll
            for (int i = 0; i < M; i++) u[i] = (i < m) ? __dadd_rn(Uk[i], __dmul_rn(alpha, sDU[(size_t)k * ms + i])) : 0.0;
        }
        return lean_stage_cost<N>(P, b, x, k < H - 1 ? u : nullptr, k);
    };
    double cost = 0.0;
    if (!viol) {
        const int ncand = line_search ? NCAND : 1;
        for (int p = lane; p < S * ncand; p += 32) {
            const int q = p / ncand, c = p % ncand;
            sCk[q * NCAND + c] = cost_at(P.kp_steps[q], alpha_of(c), sKX + (size_t)q * ns);
        }
        __syncwarp();
        if (mine)
            for (int q = 0; q < S; q++) cost += sCk[q * NCAND + lane];
    } else if (mine) {
        double dx[N];
#pragma unroll
        for (int j = 0; j < N; j++) dx[j] = 0.0;
        for (int k = 0; k < H; k++) {
            cost += cost_at(k, alpha_of(lane), dx);
            if (k < H - 1) {
#pragma unroll
                for (int i = 0; i < N / 2; i++)
                    if (i < dof) {
                        const double du = sDU[(size_t)k * ms + i];
                        if (P.nb_deriv == 1) {
                            dx[i] = __dadd_rn(dx[i], __dmul_rn(dt, du));
                        } else {
                            dx[i] = __dadd_rn(dx[i], __dadd_rn(__dmul_rn(dt, dx[dof + i]), __dmul_rn(hdt2, du)));
                            dx[dof + i] = __dadd_rn(dx[dof + i], __dmul_rn(dt, du));
                        }
                    }
            }
        }
    }
    if (mine) {
        W.cost_c[(size_t)b * NCAND + lane] = cost;
        W.dun_c[(size_t)b * NCAND + lane] = alpha_of(lane) * dsum;
    }
    // ---- selection (the rule of k_accept) and the selected candidate's controls -------------------------------------------
    const double cost0 = W.cost0[b];
    const bool good = mine && !((cost >= cost0) || isnan(cost));
    const unsigned okm = __ballot_sync(FULL, good);
    int sel = 0;
    if (line_search) sel = okm ? (__ffs(okm) - 1) : NCAND - 1;
    const double asel = alpha_of(sel);
    for (int e = lane; e < (H - 1) * ms; e += 32) {
        const int k = e / ms, i = e % ms;
        const size_t bk = (size_t)b * (H - 1) + k;
        W.Uc[(bk * NCAND + sel) * ms + i] = (i < m) ? __dadd_rn(W.U[bk * ms + i], __dmul_rn(asel, sDU[e])) : 0.0;
    }
}

// AL: constraint values and active set of the current trajectory (AL-ILQR.cpp:21-44,71-73,189-191) and, when `upd`,
// the multiplier update lambda = max(0, lambda + penalty * C) (AL-ILQR.cpp:202-208).  One warp per problem.
__device__ __forceinline__ void al_constraints(const DevProblem& P, const Work& W, int b, int lane, bool upd, double penalty_next) {
    const int H = P.H, n = P.n, m = P.m, ns = W.ns, ms = W.ms;
    const int nc = W.nc, nsu = n + m;
    // lanes 0-15 take the even step of a pair, lanes 16-31 the odd one (nc <= 16), four pairs in flight: the loads of a
    // group are issued before its first store, so the L2 latency is paid once per 8 steps instead of once per step
    constexpr int G = 4;
    const int half = lane >> 4, r = lane & 15;
    if (nc > 16) {  // wide constraint sets: one step at a time
        for (int k = 0; k < H - 1; k++) {
            const size_t bk = (size_t)b * (H - 1) + k;
            const double* Ac = W.alA + (size_t)(W.al_per_step ? k : 0) * nc * nsu;
            const double* bc = W.alb + (size_t)(W.al_per_step ? k : 0) * nc;
            const double* Xk = W.X + ((size_t)b * H + k) * ns;
            const double* Uk = W.U + bk * ms;
            bool act = false;
            if (lane < nc) {
                double g = 0.0;
                for (int j = 0; j < n; j++) g += Ac[lane * nsu + j] * Xk[j];
                for (int j = 0; j < m; j++) g += Ac[lane * nsu + n + j] * Uk[j];
                g = g - bc[lane];
                const double lam = W.lam[bk * nc + lane];
                act = !(g < 0 && lam == 0);
                W.Cs[bk * nc + lane] = g;
                if (upd) {
                    const double v = lam + penalty_next * g;
                    W.lam[bk * nc + lane] = (v < 0.0) ? 0.0 : v;  // cwiseMax(0)
                }
            }
            const unsigned bits = __ballot_sync(0xffffffffu, act);
            if (lane == 0) W.mask[bk] = bits;
        }
        return;
    }
    for (int k0 = 0; k0 < H - 1; k0 += 2 * G) {
        double g[G], lam[G];
        bool in[G];
#pragma unroll
        for (int q = 0; q < G; q++) {
            const int k = k0 + 2 * q + half;
            in[q] = (k < H - 1) && (r < nc);
            g[q] = 0.0;
            lam[q] = 0.0;
            if (in[q]) {
                const size_t bk = (size_t)b * (H - 1) + k;
                const double* Ac = W.alA + (size_t)(W.al_per_step ? k : 0) * nc * nsu + (size_t)r * nsu;
                const double* Xk = W.X + ((size_t)b * H + k) * ns;
                const double* Uk = W.U + bk * ms;
                double acc = 0.0;
                for (int j = 0; j < n; j++) acc += Ac[j] * Xk[j];
                for (int j = 0; j < m; j++) acc += Ac[n + j] * Uk[j];
                g[q] = acc - W.alb[(size_t)(W.al_per_step ? k : 0) * nc + r];
                lam[q] = W.lam[bk * nc + r];
            }
        }
#pragma unroll
        for (int q = 0; q < G; q++) {
            const int k = k0 + 2 * q + half;
            const size_t bk = (size_t)b * (H - 1) + k;
            const bool act = in[q] && !(g[q] < 0 && lam[q] == 0);
            if (in[q]) {
                W.Cs[bk * nc + r] = g[q];
                if (upd) {
                    const double v = lam[q] + penalty_next * g[q];
                    W.lam[bk * nc + r] = (v < 0.0) ? 0.0 : v;  // cwiseMax(0)
                }
            }
            const unsigned bits = __ballot_sync(0xffffffffu, act);
            if (r == 0 && k < H - 1) W.mask[bk] = (bits >> (16 * half)) & 0xffffu;
        }
    }
}
// the same for constraint sets shared by all steps with nc <= 16 (the C3 shape): lane (half, r) keeps row r of A in
// registers, the trip counts are compile-time bounds, so the loads of 8 steps are in flight together (the runtime-bounded
// loop above leaves 2-4: ncu shows 70 % of k_accept waiting on them).  Same summation order as al_constraints.
template <int N, int M>
__device__ __forceinline__ void al_constraints_t(const DevProblem& P, const Work& W, int b, int lane, bool upd, double penalty_next) {
    const int H = P.H, n = P.n, m = P.m, ns = W.ns, ms = W.ms;
    const int nc = W.nc, nsu = n + m;
    constexpr int G = 4;
    const int half = lane >> 4, r = lane & 15;
    const bool row = r < nc;
    double Ax[N], Au[M];
#pragma unroll
    for (int j = 0; j < N; j++) Ax[j] = (row && j < n) ? W.alA[(size_t)r * nsu + j] : 0.0;
#pragma unroll
    for (int j = 0; j < M; j++) Au[j] = (row && j < m) ? W.alA[(size_t)r * nsu + n + j] : 0.0;
    const double br = row ? W.alb[r] : 0.0;
    for (int k0 = 0; k0 < H - 1; k0 += 2 * G) {
        double g[G], lam[G];
        bool in[G];
#pragma unroll
        for (int q = 0; q < G; q++) {
            const int k = k0 + 2 * q + half;
            in[q] = (k < H - 1) && row;
            const int kc = k < H - 1 ? k : H - 2;  // clamped: the loads stay unconditional and independent
            const size_t bk = (size_t)b * (H - 1) + kc;
            const double* Xk = W.X + ((size_t)b * H + kc) * ns;
            const double* Uk = W.U + bk * ms;
            double acc = 0.0;
#pragma unroll
            for (int j = 0; j < N; j++)
                if (j < n) acc = fma(Ax[j], Xk[j], acc);
#pragma unroll
            for (int j = 0; j < M; j++)
                if (j < m) acc = fma(Au[j], Uk[j], acc);
            g[q] = acc - br;
            lam[q] = row ? W.lam[bk * nc + r] : 0.0;
        }
#pragma unroll
        for (int q = 0; q < G; q++) {
            const int k = k0 + 2 * q + half;
            const size_t bk = (size_t)b * (H - 1) + k;
            const bool act = in[q] && !(g[q] < 0 && lam[q] == 0);
            if (in[q]) {
                W.Cs[bk * nc + r] = g[q];
                if (upd) {
                    const double v = lam[q] + penalty_next * g[q];
                    W.lam[bk * nc + r] = (v < 0.0) ? 0.0 : v;  // cwiseMax(0)
                }
            }
            const unsigned bits = __ballot_sync(0xffffffffu, act);
            if (r == 0 && k < H - 1) W.mask[bk] = (bits >> (16 * half)) & 0xffffu;
        }
    }
}
__global__ void __launch_bounds__(128) k_al_init(const DevProblem* __restrict__ Pp, Work W) {
    const DevProblem& P = *Pp;
    const int lane = threadIdx.x & 31;
    const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (b >= P.B) return;
    al_constraints(P, W, b, lane, false, 0.0);
}

// ------------------------------------------------------------------------------------------------------------
// accept: one warp per problem.
//   candidate choice = the reference's do/while (ILQRRecursive.cpp:107-155): first alpha in 1, 1/2, .. whose cost is
//   not (>= cost0 or NaN); alpha = 2^-10 is taken unconditionally (Q2).  Then X,U <- candidate; early-stop test
//   (ILQRRecursive.cpp:174: alpha*sqrt(sum|du|) < 1e-3 && cost < 1e-3 ; AL-ILQR.cpp:225 without the cost test).
//   AL: Cs / active set of the accepted rollout (AL-ILQR.cpp:189-191) and multiplier update (AL-ILQR.cpp:202-208).
template <int N, int M>
__global__ void __launch_bounds__(128) k_accept(const DevProblem* __restrict__ Pp, Work W, int it, int nb_iter, int line_search, int early_stop, int al,
                                                int lag_update_step, double penalty_next) {
    const DevProblem& P = *Pp;
    const int lane = threadIdx.x & 31;
    const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (b >= P.B) return;
    if (W.done[b]) return;
    const int H = P.H, n = P.n, m = P.m, ns = W.ns, ms = W.ms, dof = P.dof;
    const double cost0 = W.cost0[b];
    int sel = 0;
    if (line_search) {
        sel = NCAND - 1;
        for (int c = 0; c < NCAND; c++) {
            const double cc = W.cost_c[(size_t)b * NCAND + c];
            if (!((cc >= cost0) || isnan(cc))) {
                sel = c;
                break;
            }
        }
    }
    const double alpha = alpha_of(sel);
    const double newCost = W.cost_c[(size_t)b * NCAND + sel];
    const double dun = W.dun_c[(size_t)b * NCAND + sel];
    // copy U and recompute X with the same recurrence as the candidate rollout
    double xq = 0.0, xv = 0.0, xt = 0.0;  // lane i < dof owns joint i ; lane 0 also owns t
    if (lane < dof) {
        xq = P.q0[(size_t)b * dof + lane];
        if (P.nb_deriv == 2) xv = P.dq0[(size_t)b * dof + lane];
    }
    constexpr int PFK = 8;  // the controls of 8 steps are loaded ahead of the serial state recurrence
    for (int k0 = 0; k0 < H - 1; k0 += PFK) {
        double ub[PFK];
#pragma unroll
        for (int q = 0; q < PFK; q++) {
            const int k = k0 + q;
            ub[q] = (k < H - 1 && lane < ms) ? W.Uc[(((size_t)b * (H - 1) + k) * NCAND + sel) * ms + lane] : 0.0;
        }
#pragma unroll
        for (int q = 0; q < PFK; q++) {
            const int k = k0 + q;
            if (k < H - 1) {
                const size_t bk = (size_t)b * (H - 1) + k;
                const double uu = ub[q];
                if (lane < ms) W.U[bk * ms + lane] = uu;
                double dt = P.dt;
                if (P.time) {
                    const double s = __shfl_sync(0xffffffffu, uu, m - 1);
                    dt = __dmul_rn(s, s);
                }
                const double hdt2 = __dmul_rn(dt, dt) / 2;
                if (lane < dof) {
                    if (P.nb_deriv == 1) {
                        xq = __dadd_rn(xq, __dadd_rn(__dmul_rn(dt, uu), __dmul_rn(hdt2, 0.0)));
                    } else {
                        xq = __dadd_rn(xq, __dadd_rn(__dmul_rn(dt, xv), __dmul_rn(hdt2, uu)));
                        xv = __dadd_rn(xv, __dmul_rn(dt, uu));
                    }
                }
                xt = __dadd_rn(xt, dt);
                double* Xn = W.X + ((size_t)b * H + k + 1) * ns;
                if (lane < dof) {
                    Xn[lane] = xq;
                    if (P.nb_deriv == 2) Xn[dof + lane] = xv;
                }
                if (P.time && lane == 0) Xn[n - 1] = xt;
            }
        }
    }
    __syncwarp();
    if (al) {
        const bool upd = ((it + 1) % lag_update_step) == 0;
        if (W.nc <= 16 && !W.al_per_step) al_constraints_t<N, M>(P, W, b, lane, upd, penalty_next);
        else al_constraints(P, W, b, lane, upd, penalty_next);
    }
    if (lane == 0) {
        W.cost0[b] = newCost;
        W.cost_hist[(size_t)b * nb_iter + it] = newCost;
        W.alpha_hist[(size_t)b * nb_iter + it] = alpha;
        W.alpha_last[b] = alpha;
        W.iters[b] = it + 1;
        int st = W.status[b];
        if (isnan(newCost)) st |= ILQRB_ST_NAN_COST;
        bool stop = early_stop && (alpha * sqrt(dun) < 1e-3) && (al || newCost < 1e-3);
        if (stop) {
            st |= ILQRB_ST_EARLY_STOP;
            W.done[b] = 1;
        }
        W.status[b] = st;
    }
}

// ------------------------------------------------------------------------------------------------------------
// f(X) for the returned trajectory: one thread per (problem, timestep)
__global__ void __launch_bounds__(128) k_final_fx(const DevProblem* __restrict__ Pp, Work W, double* __restrict__ fX, int b0, int nB) {
    const DevProblem& P = *Pp;
    const int idx0 = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx0 >= nB * P.H) return;
    const int b = b0 + idx0 / P.H, k = idx0 % P.H;
    const size_t idx = (size_t)b * P.H + k;
    const int n = P.n, dof = P.dof;
    const double* x = W.X + ((size_t)b * P.H + k) * W.ns;
    double xs[MAXN], dqs[MAXM];
    for (int i = 0; i < n; i++) xs[i] = x[i];
    // f_X[k] comes from forwardPass(k-1): the robot's dq is the last command (nd=1: u_{k-1}; k=0: dq0) — only the
    // velocity part of fx depends on it, which exists for nb_deriv == 2 where dq is part of the state.
    for (int i = 0; i < dof; i++) dqs[i] = (P.nb_deriv == 2) ? xs[dof + i] : 0.0;
    double* out = fX + (size_t)idx * P.nt;
    for (int t = 0; t < P.n_tasks; t++) {
        const DevTask& tk = P.tasks[t];
        if (P.nb_deriv == 1 && P.sys_kind != ILQRB_SYS_POSITION && tk.kind == ILQRB_TASK_POSE && P.robots[tk.robot].kind == ILQRB_ROBOT_KDL) {
            // first-order end-effector tasks: pose only (the sweep of the forward kernels' cost evaluation, no Jacobian)
            double fx[7], Jc[6];
            const double* qb = xs + tk.joint_offset;
            lean_task_eval(P, tk, [&](int j) { return qb[j]; }, -1, fx, Jc);
            double* o = out + tk.fx_off;
#pragma unroll
            for (int i = 0; i < 7; i++) o[i] = fx[i];
            if (P.time) o[7] = xs[n - 1];
            continue;
        }
        EEState e;
        task_ee(P, tk, xs, dqs, e);
        task_fx(P, tk, e, P.time ? xs[n - 1] : 0.0, out + tk.fx_off);
    }
}

// ------------------------------------------------------------------------------------------------------------
// backward Riccati pass: one problem per warp, all blocks in shared memory, products on the FP64 tensor pipe.
// NXT, NUT: 8-tiles of the (augmented) state and control dimensions: 8*NXT >= n+1, 8*NUT >= m.
// The vectors ride in column n of the state-indexed matrices:  Paug = [P | p], QuxQu = [Qux | Qu], Kd = [K | d].
template <int NXT, int NUT>
struct BwdSmem {
    static constexpr int NX = 8 * NXT, NU = 8 * NUT, LDX = NX + 4, LDU = NU + 4;
    static constexpr int SZ_XX = NX * LDX, SZ_UX = NU * LDX, SZ_XU = NX * LDU, SZ_UU = NU * LDU;
    // per warp: Paug, AtP, QxxQx (XX) ; BtP, QuxQu, Kd (UX) ; Qxu (XU) ; Quu (UU).
    // Inv reuses BtP (dead by then; UX >= UU because NX >= NU).
    static_assert(NXT >= NUT, "state tiles must cover control tiles");
    static constexpr int PER_WARP = 3 * SZ_XX + 3 * SZ_UX + 1 * SZ_XU + 1 * SZ_UU;
    static constexpr int DYN = SZ_XX + SZ_XU;  // A_aug, B
};

template <int NXT, int NUT, bool TIME, bool AL, int WARPS, int MINB>
__global__ void __launch_bounds__(32 * WARPS, MINB) k_backward(const DevProblem* __restrict__ Pp, Work W, double penalty_frozen) {
    using S = BwdSmem<NXT, NUT>;
    constexpr int NX = S::NX, NU = S::NU, LDX = S::LDX, LDU = S::LDU;
    extern __shared__ double smem[];
    const DevProblem& P = *Pp;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int b = blockIdx.x * WARPS + wid;
    const int n = P.n, m = P.m, H = P.H, ns = W.ns, ms = W.ms, dof = P.dof;
    // dynamics matrices: shared by the block when constant, per warp for time systems
    double* Aaug = smem + (TIME ? (size_t)wid * S::DYN : 0);
    double* Bm = Aaug + S::SZ_XX;
    double* base = smem + (TIME ? WARPS : 1) * S::DYN + (size_t)wid * S::PER_WARP;
    double* Paug = base;
    double* AtP = Paug + S::SZ_XX;
    double* QxxQx = AtP + S::SZ_XX;
    double* BtP = QxxQx + S::SZ_XX;
    double* QuxQu = BtP + S::SZ_UX;
    double* Kd = QuxQu + S::SZ_UX;
    double* Qxu = Kd + S::SZ_UX;
    double* Quu = Qxu + S::SZ_XU;
    double* Inv = BtP;    // alias (BtP is dead once Quu is formed)

    // ---- A_aug = blockdiag(A, 1), B  (PosOrnPlannerSys.cpp:115-128) ----
    if (!TIME) {
        for (int e = threadIdx.x; e < NX * LDX; e += 32 * WARPS) {
            const int i = e / LDX, j = e % LDX;
            double v = 0.0;
            if (i == j && i <= n) v = 1.0;
            if (P.nb_deriv == 2 && i < dof && j == dof + i) v = 1.0 * P.dt;
            Aaug[e] = v;
        }
        for (int e = threadIdx.x; e < NX * LDU; e += 32 * WARPS) {
            const int i = e / LDU, j = e % LDU;
            double v = 0.0;
            if (P.nb_deriv == 1) {
                if (i == j && i < dof) v = 1.0 * P.dt;
            } else {
                if (i < dof && j == i) v = 1.0 * P.dt * P.dt / 2;
                if (i >= dof && i < 2 * dof && j == i - dof) v = 1.0 * P.dt;
            }
            Bm[e] = v;
        }
        __syncthreads();
    }
    if (b >= P.B) return;
    if (W.done[b]) return;

    // l_xx | l_x of timestep k into dst (NX x LDX):  keypoint part from k_kp_derivs + joint-limit part (System.cpp:263-269,299-305)
    auto fill_L = [&](double* dst, int k) {
        int s = -1;
        for (int i = 0; i < P.n_kp_steps; i++)
            if (P.kp_steps[i] == k) s = i;
        const double* lx = s >= 0 ? W.LX + ((size_t)b * P.n_kp_steps + s) * n : nullptr;
        const double* lxx = s >= 0 ? W.LXX + ((size_t)b * P.n_kp_steps + s) * n * n : nullptr;
        for (int e = lane; e < NX * LDX; e += 32) {
            const int i = e / LDX, j = e % LDX;
            double v = 0.0;
            if (s >= 0 && i < n) {
                if (j < n) v = lxx[i * n + j];
                else if (j == n) v = lx[i];
            }
            dst[e] = v;
        }
        __syncwarp();
        if (P.any_limits) {
            const double* Xk = W.X + ((size_t)b * H + k) * ns;
            if (lane < n) {
                for (int t = 0; t < P.n_tasks; t++) {
                    const DevTask& tk = P.tasks[t];
                    if (!tk.limits_set) continue;
                    bool act;
                    const double q = limit_q(tk, lane, Xk[lane], act);
                    if (act) {
                        dst[lane * LDX + lane] += 1.0 * 1.0;      // L^T L
                        dst[lane * LDX + n] += -1.0 * q;          // -L^T q
                    }
                }
            }
            __syncwarp();
        }
    };

    // terminal: P = cost_F_xx, p = cost_F_x  (ILQRRecursive.cpp:69-71)
    fill_L(Paug, H - 1);
    bool spd = true;

    for (int k = H - 2; k >= 0; k--) {
        const size_t bk = (size_t)b * (H - 1) + k;
        const double* Uk = W.U + bk * ms;
        if (TIME) {
            // per-step dynamics (PosOrnTimePlannerSys.cpp:149-177), from the accepted trajectory (Q15: post-step velocity)
            const double sq = Uk[m - 1];
            const double dt = sq * sq;
            const double* Xn = W.X + ((size_t)b * H + k + 1) * ns;
            for (int e = lane; e < NX * LDX; e += 32) {
                const int i = e / LDX, j = e % LDX;
                double v = 0.0;
                if (i == j && i <= n) v = 1.0;
                if (P.nb_deriv == 2 && i < dof && j == dof + i) v = dt * 1.0;
                Aaug[e] = v;
            }
            for (int e = lane; e < NX * LDU; e += 32) {
                const int i = e / LDU, j = e % LDU;
                double v = 0.0;
                if (P.nb_deriv == 1) {
                    if (i < dof && j == i) v = 1.0 * dt;
                    if (i < dof && j == m - 1) v = 2 * sq * Uk[i];
                } else {
                    if (i < dof && j == i) v = 1.0 * dt * dt / 2;
                    if (i >= dof && i < 2 * dof && j == i - dof) v = 1.0 * dt;
                    if (i < dof && j == m - 1) v = 2 * sq * Xn[dof + i] + 2 * sq * sq * sq * Uk[i];
                    if (i >= dof && i < 2 * dof && j == m - 1) v = 2 * sq * Uk[i - dof];
                }
                if (i == n - 1 && j == m - 1) v = 2 * sq;
                Bm[e] = v;
            }
        }
        // QuxQu <- [cost_ux = 0 | cost_u = R u]  (System.cpp:274-280) ; QxxQx <- [l_xx | l_x]
        for (int e = lane; e < NU * LDX; e += 32) {
            const int i = e / LDX, j = e % LDX;
            QuxQu[e] = (j == n && i < m) ? P.r_diag[i] * Uk[i] : 0.0;
        }
        fill_L(QxxQx, k);  // ends with __syncwarp

        // BtP = B^T [P|p]
        wm_gemm<NUT, NXT, NXT, true, false, false, false>(BtP, LDX, Bm, LDU, Paug, LDX, lane);
        // [Qux|Qu] += BtP A_aug           ILQRRecursive.cpp:82,86
        wm_gemm<NUT, NXT, NXT, false, false, true, false>(QuxQu, LDX, BtP, LDX, Aaug, LDX, lane);
        // Quu = R + BtP B                 ILQRRecursive.cpp:83
        for (int e = lane; e < NU * LDU; e += 32) {
            const int i = e / LDU, j = e % LDU;
            Quu[e] = (i == j && i < m) ? P.r_diag[i] : 0.0;
        }
        __syncwarp();
        wm_gemm<NUT, NUT, NXT, false, false, true, false>(Quu, LDU, BtP, LDX, Bm, LDU, lane);
        // AtP = A^T [P|p] ; [Qxx|Qx] += AtP A_aug ; Qxu = AtP B      ILQRRecursive.cpp:84,85,87
        wm_gemm<NXT, NXT, NXT, true, false, false, false>(AtP, LDX, Aaug, LDX, Paug, LDX, lane);
        wm_gemm<NXT, NXT, NXT, false, false, true, false>(QxxQx, LDX, AtP, LDX, Aaug, LDX, lane);
        wm_gemm<NXT, NUT, NXT, false, false, false, false>(Qxu, LDU, AtP, LDX, Bm, LDU, lane);

        if (AL) {
            // AL terms (AL-ILQR.cpp:116-134): ckx = A[:, :n], cku = A[:, n:], Ik = penalty * diag(active), lic = lambda + Ik c
            const int nc = W.nc, nsu = n + m;
            const double* Ac = W.alA + (size_t)(W.al_per_step ? k : 0) * nc * nsu;
            const unsigned bits = W.mask[bk];
            const double* lam = W.lam + bk * nc;
            const double* cs = W.Cs + bk * nc;
            // element loops: every lane owns entries (i, j) of the four Q blocks
            for (int e = lane; e < m * (n + 1); e += 32) {  // Qux | Qu
                const int i = e / (n + 1), j = e % (n + 1);
                double acc = 0.0;
                for (int r = 0; r < nc; r++) {
                    const double ik = ((bits >> r) & 1u) ? penalty_frozen : 0.0;
                    const double cu = Ac[r * nsu + n + i];
                    if (j < n) acc += (cu * ik) * Ac[r * nsu + j];
                    else acc += cu * (lam[r] + ik * cs[r]);
                }
                QuxQu[i * LDX + j] += acc;
            }
            for (int e = lane; e < m * m; e += 32) {  // Quu
                const int i = e / m, j = e % m;
                double acc = 0.0;
                for (int r = 0; r < nc; r++) {
                    const double ik = ((bits >> r) & 1u) ? penalty_frozen : 0.0;
                    acc += (Ac[r * nsu + n + i] * ik) * Ac[r * nsu + n + j];
                }
                Quu[i * LDU + j] += acc;
            }
            for (int e = lane; e < n * (n + 1); e += 32) {  // Qxx | Qx
                const int i = e / (n + 1), j = e % (n + 1);
                double acc = 0.0;
                for (int r = 0; r < nc; r++) {
                    const double ik = ((bits >> r) & 1u) ? penalty_frozen : 0.0;
                    const double cx = Ac[r * nsu + i];
                    if (j < n) acc += (cx * ik) * Ac[r * nsu + j];
                    else acc += cx * (lam[r] + ik * cs[r]);
                }
                QxxQx[i * LDX + j] += acc;
            }
            for (int e = lane; e < n * m; e += 32) {  // Qxu
                const int i = e / m, j = e % m;
                double acc = 0.0;
                for (int r = 0; r < nc; r++) {
                    const double ik = ((bits >> r) & 1u) ? penalty_frozen : 0.0;
                    acc += (Ac[r * nsu + i] * ik) * Ac[r * nsu + n + j];
                }
                Qxu[i * LDU + j] += acc;
            }
            __syncwarp();
        }

        // Inv = (Quu + 1e-6 I)^-1        ILQRRecursive.cpp:89
        for (int e = lane; e < NU * LDU; e += 32) {
            const int i = e / LDU, j = e % LDU;
            double v = Quu[e];
            if (i == j) v = (i < m) ? v + 1e-6 * 1.0 : 1.0;
            Inv[e] = v;
        }
        __syncwarp();
        if (!wm_inverse_spd<NUT>(Inv, LDU, m, lane)) {  // a pivot was not positive: the reference's pivoted LU inverse (rare)
            spd = false;
            for (int e = lane; e < NU * LDU; e += 32) {
                const int i = e / LDU, j = e % LDU;
                double v = Quu[e];
                if (i == j) v = (i < m) ? v + 1e-6 * 1.0 : 1.0;
                Inv[e] = v;
            }
            wm_inverse_pivoted(Inv, LDU, m, Kd, lane);
        }
        // [K|d] = -(Inv [Qux|Qu])          ILQRRecursive.cpp:90-91
        wm_gemm<NUT, NXT, NUT, false, false, false, true>(Kd, LDX, Inv, LDU, QuxQu, LDX, lane);
        // stream K, d to global (time-forward index)
        {
            double* Kg = W.K + bk * m * ns;
            for (int e = lane; e < m * (ns / 2); e += 32) {
                const int i = e / (ns / 2), j2 = e % (ns / 2);
                double2 v = *reinterpret_cast<const double2*>(Kd + i * LDX + 2 * j2);
                if (2 * j2 + 1 >= n) v.y = 0.0;
                *reinterpret_cast<double2*>(Kg + i * ns + 2 * j2) = v;
            }
            if (lane < ms) W.d[bk * ms + lane] = (lane < m) ? Kd[lane * LDX + n] : 0.0;
        }
        __syncwarp();
        // [P|p] = [Qxx|Qx] + K^T (Quu [K|d] + [Qux|Qu]) + Qxu [K|d]      ILQRRecursive.cpp:93-94
        // the bracket M (un-regularised Quu) is formed as written, in place over [Qux|Qu]: stationary in K (see backward_n1_body)
        wm_gemm<NUT, NXT, NUT, false, false, true, false>(QuxQu, LDX, Quu, LDU, Kd, LDX, lane);
        {
            const int g = lane >> 2, t = lane & 3;
#pragma unroll
            for (int rt = 0; rt < NXT; rt++)
#pragma unroll
                for (int ct = 0; ct < NXT; ct++) {
                    double a0 = 0.0, a1 = 0.0, c0 = 0.0, c1 = 0.0;
                    wm_tile_acc<NUT, true, false>(a0, a1, Kd, LDX, QuxQu, LDX, rt, ct, g, t);
                    wm_tile_acc<NUT, false, false>(c0, c1, Qxu, LDU, Kd, LDX, rt, ct, g, t);
                    const int i = 8 * rt + g, j = 8 * ct + 2 * t;
                    const double2 q = *reinterpret_cast<const double2*>(QxxQx + i * LDX + j);
                    double v0 = (q.x + a0) + c0;
                    double v1 = (q.y + a1) + c1;
                    if (i >= n || j > n) v0 = 0.0;
                    if (i >= n || j + 1 > n) v1 = 0.0;
                    *reinterpret_cast<double2*>(Paug + i * LDX + j) = make_double2(v0, v1);
                }
            __syncwarp();
        }
    }
    if (!spd && lane == 0) W.status[b] |= ILQRB_ST_QUU_NOT_SPD;
}

}  // namespace ilqrb
