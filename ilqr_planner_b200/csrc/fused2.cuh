// fused2.cuh — fast path of the Recursive solver for DOUBLE-integrator systems without a time state
// (PosOrnPlannerSys nb_deriv = 2, PosOrnPlannerSys.cpp:122-128):  x = [q; v],  A = [[I, dt I], [0, I]],  B = [dt^2/2 I; dt I],
// n = 2 dof, m = dof <= 7.  One problem per warp, everything in registers in the DMMA accumulator layout.
//
// Padded state index: q_i -> i, v_i -> 8 + i, so that each dof x dof block of a state-indexed matrix is its own 8x8
// tile and the augmented column (the vector that rides along: p, Qx, Qu, d) is column 15 = local column 7 of the
// second column tile.  With this layout the products with A and B are tile-wise scalings and sums:
//   B^T [P|p] = [M1 M2],  M1 = c2 P11 + dt P21,  M2 = c2 P12 + dt P22          (c2 = dt^2/2)
//   [Qux|Qu]  = [M1, dt M1 + M2] + [0 | R u]        Quu = R + c2 M1 + dt M2
//   A^T [P|p] = N,  N1c = P1c,  N2c = dt P1c + P2c ;   [Qxx|Qx] = [l_xx|l_x] + [Nr1, dt Nr1 + Nr2] ;   Qxu = c2 Nr1 + dt Nr2
// Only [K|d] = -(Quu + 1e-6 I)^-1 [Qux|Qu] (in-register Gauss-Jordan inverse, 4 DMMA for the product) and
// M = Quu [K|d] + [Qux|Qu] (4 DMMA), P' = Qxx + K^T M + Qxu K (16 DMMA m8n8k4.f64) remain as dense work per step
// (ILQRRecursive.cpp:69-97).
#pragma once
#include "fused1.cuh"

namespace ilqrb {

struct BwdN2Smem {
    static constexpr int LDW = 20, LDQ = 12;                 // 8 x 16 matrices / the 8 x 8 Quu
    static constexpr int SZW = 8 * LDW, SZQ = 8 * LDQ;
    static constexpr int PER_WARP = 2 * SZW + 4 * SZQ;       // [K|d], [Qux|Qu] then M, Qxu (16 x 8), (Quu + 1e-6 I)^-1, Quu
};

// fragment loads of the 8 x 16 matrices (row-major, leading dimension LDW)
__device__ __forceinline__ double ld_bw(const double* M, int s, int c, int g, int t) { return M[(4 * s + t) * BwdN2Smem::LDW + 8 * c + g]; }
__device__ __forceinline__ void st_w(double* M, const double (&e)[2][2], int g, int t) {
#pragma unroll
    for (int c = 0; c < 2; c++) *reinterpret_cast<double2*>(M + g * BwdN2Smem::LDW + 8 * c + 2 * t) = make_double2(e[c][0], e[c][1]);
}

// PIVOT: see backward_n1_body (fused1.cuh) — false = first pass, unpivoted only, returns whether every pivot was positive;
// true = the out-of-line second pass with the pivoted fallback.
template <int DOF, bool AL, bool PIVOT = false>
__device__ __forceinline__ bool backward_n2_body(const DevProblem& P, const Work& W, int b, int lane, double* wsmem, double penalty_frozen) {
    using S = BwdN2Smem;
    const int g = lane >> 2, t = lane & 3;
    const int dof = DOF > 0 ? DOF : P.dof, n = 2 * dof, m = dof, H = P.H;
    const int ns = W.ns, ms = W.ms;
    const double dt = 1.0 * P.dt, c2 = 1.0 * P.dt * P.dt / 2;
    double* sKd = wsmem;
    double* sM = sKd + S::SZW;
    double* sQxu = sM + S::SZW;  // 16 x 8
    double* sInv = sQxu + 2 * S::SZQ;
    double* sQuu = sInv + S::SZQ;
    const bool rowv = g < dof;  // this lane's row is a real control / state row of its tile
    const double Rg = rowv ? P.r_diag[g] : 0.0;
    const double* Xb = W.X + (size_t)b * H * ns;
    const double* Ub = W.U + (size_t)b * (H - 1) * ms;
    const int S_kp = P.n_kp_steps;

    // e += [l_xx | l_x] of keypoint step index s ; natural index of (tile r, local i) is r dof + i      System.cpp:252-260,290-296
    auto add_kp = [&](double (&e)[2][2][2], int s) {
        const double* lx = W.LX + ((size_t)b * S_kp + s) * n;
        const double* lxx = W.LXX + ((size_t)b * S_kp + s) * n * n;
        if (!rowv) return;
#pragma unroll
        for (int r = 0; r < 2; r++)
#pragma unroll
            for (int c = 0; c < 2; c++)
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int jl = 2 * t + h;
                    if (jl < dof) e[r][c][h] += lxx[(r * dof + g) * n + c * dof + jl];
                    else if (c == 1 && jl == 7) e[r][c][h] += lx[r * dof + g];
                }
    };
    // joint limits (System.cpp:121-142,263-269,299-305): e += [L^T L | -L^T q], L = diag(violated), q = limit - x
    auto add_limits = [&](double (&e)[2][2][2], int k) {
        if (!P.any_limits || !rowv) return;
        const double* Xk = Xb + (size_t)k * ns;
        for (int ti = 0; ti < P.n_tasks; ti++) {
            const DevTask& tk = P.tasks[ti];
            if (!tk.limits_set) continue;
#pragma unroll
            for (int r = 0; r < 2; r++) {
                bool act;
                const double q = limit_q(tk, r * dof + g, Xk[r * dof + g], act);
                if (act) {
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        if (2 * t + h == g) e[r][r][h] += 1.0 * 1.0;
                        if (2 * t + h == 7) e[r][1][h] += -1.0 * q;
                    }
                }
            }
        }
    };

    // terminal: P = cost_F_xx, p = cost_F_x   (ILQRRecursive.cpp:69-71)
    double Pa[2][2][2];
#pragma unroll
    for (int r = 0; r < 2; r++)
#pragma unroll
        for (int c = 0; c < 2; c++) Pa[r][c][0] = Pa[r][c][1] = 0.0;
    int sp = S_kp - 1;
    int next_kp = sp >= 0 ? P.kp_steps[sp] : -1;
    if (next_kp == H - 1) {
        add_kp(Pa, sp);
        sp--;
        next_kp = sp >= 0 ? P.kp_steps[sp] : -1;
    }
    add_limits(Pa, H - 1);
    bool spd = true;
    double* Kg = W.K + ((size_t)b * (H - 1) + (H - 2)) * m * ns;
    double* dg = W.d + ((size_t)b * (H - 1) + (H - 2)) * ms;
    double un = rowv ? Ub[(size_t)(H - 2) * ms + g] : 0.0;

    for (int k = H - 2; k >= 0; k--, Kg -= m * ns, dg -= ms) {
        const double u = un;
        if (k > 0) un = rowv ? Ub[(size_t)(k - 1) * ms + g] : 0.0;
        double Qa[2][2], Quu[2], Lr[2], Qx[2][2][2], Qxu[2][2];
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int jl = 2 * t + h;
            // separate multiplies and adds (no FMA contraction): the same roundings as the reference's dense products,
            // whose other terms are exact zeros
            const double m1 = __dadd_rn(__dmul_rn(c2, Pa[0][0][h]), __dmul_rn(dt, Pa[1][0][h]));  // B^T P, columns of q
            const double m2 = __dadd_rn(__dmul_rn(c2, Pa[0][1][h]), __dmul_rn(dt, Pa[1][1][h]));  // B^T [P|p], columns of v and the vector
            Qa[0][h] = m1;
            const double qv = __dadd_rn(__dmul_rn(m1, dt), m2);
            Qa[1][h] = (jl == 7) ? __dadd_rn(__dmul_rn(Rg, u), qv) : qv;
            double quu = __dadd_rn(__dmul_rn(m1, c2), __dmul_rn(m2, dt));
            if (jl == g) quu = __dadd_rn(Rg, quu);
            quu = (jl < dof && rowv) ? quu : 0.0;
            Quu[h] = quu;
            Lr[h] = (jl == g) ? (rowv ? quu + 1e-6 * 1.0 : 1.0) : quu;  // Quu + 1e-6 I, identity outside          ILQRRecursive.cpp:89
            const double n20 = __dadd_rn(__dmul_rn(dt, Pa[0][0][h]), Pa[1][0][h]);  // A^T [P|p], second row block
            const double n21 = __dadd_rn(__dmul_rn(dt, Pa[0][1][h]), Pa[1][1][h]);
            // Qxu = (A^T P) B, from P itself: NOT Qux^T — the transposed shortcut leaves the antisymmetric rounding part of P
            // without the cross terms that damp it, E' = A^T E A + (BK)^T E (BK), and that mode grows along the horizon
            Qxu[0][h] = (jl < dof) ? __dadd_rn(__dmul_rn(Pa[0][0][h], c2), __dmul_rn(Pa[0][1][h], dt)) : 0.0;
            Qxu[1][h] = (jl < dof) ? __dadd_rn(__dmul_rn(n20, c2), __dmul_rn(n21, dt)) : 0.0;
            Qx[0][0][h] = Pa[0][0][h];
            Qx[1][0][h] = n20;
            Qx[0][1][h] = __dadd_rn(__dmul_rn(Pa[0][0][h], dt), Pa[0][1][h]);
            Qx[1][1][h] = __dadd_rn(__dmul_rn(n20, dt), n21);
        }
        if (AL) {
            // AL terms (AL-ILQR.cpp:116-134): rows r of the active set add Ik_r a_r a_r^T to the Hessian blocks and a_r (lambda_r + Ik_r c_r)
            // to the gradients, a_r = [ckx_r cku_r], Ik_r = penalty.  Inactive rows contribute exact zeros (lambda = 0 there), so the
            // sums run over the set bits only; each sum is formed on its own and then added, as the reference's products are.
            const size_t bk = (size_t)b * (H - 1) + k;
            const int nc = W.nc, nsu = n + m;
            const double* Ac = W.alA + (size_t)(W.al_per_step ? k : 0) * nc * nsu;
            unsigned bits = W.mask[bk];
            if (bits) {
                double Gxx[2][2][2] = {{{0, 0}, {0, 0}}, {{0, 0}, {0, 0}}}, Gux[2][2] = {{0, 0}, {0, 0}}, Guu[2] = {0, 0}, Gxu[2][2] = {{0, 0}, {0, 0}};
                while (bits) {
                    const int r = __ffs(bits) - 1;
                    bits &= bits - 1;
                    const double* a = Ac + (size_t)r * nsu;
                    const double lic = W.lam[bk * nc + r] + penalty_frozen * W.Cs[bk * nc + r];
                    const double aur = rowv ? a[n + g] : 0.0;
                    const double axr[2] = {rowv ? a[g] : 0.0, rowv ? a[dof + g] : 0.0};
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        const int jl = 2 * t + h;
                        const bool cv = jl < dof;
                        const double ac0 = cv ? a[jl] : 0.0, ac1 = cv ? a[dof + jl] : 0.0, auc = cv ? a[n + jl] : 0.0;
                        Gux[0][h] += (aur * penalty_frozen) * ac0;
                        Gux[1][h] += (jl == 7) ? aur * lic : (aur * penalty_frozen) * ac1;
                        Guu[h] += (aur * penalty_frozen) * auc;
#pragma unroll
                        for (int rr = 0; rr < 2; rr++) {
                            Gxx[rr][0][h] += (axr[rr] * penalty_frozen) * ac0;
                            Gxx[rr][1][h] += (jl == 7) ? axr[rr] * lic : (axr[rr] * penalty_frozen) * ac1;
                            Gxu[rr][h] += (axr[rr] * penalty_frozen) * auc;
                        }
                    }
                }
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    Qa[0][h] += Gux[0][h];
                    Qa[1][h] += Gux[1][h];
                    Quu[h] += Guu[h];
                    Lr[h] = (2 * t + h == g) ? (rowv ? Quu[h] + 1e-6 * 1.0 : 1.0) : Quu[h];
#pragma unroll
                    for (int rr = 0; rr < 2; rr++) {
                        Qx[rr][0][h] += Gxx[rr][0][h];
                        Qx[rr][1][h] += Gxx[rr][1][h];
                        Qxu[rr][h] += Gxu[rr][h];
                    }
                }
            }
        }
        if (k == next_kp) {
            add_kp(Qx, sp);
            sp--;
            next_kp = sp >= 0 ? P.kp_steps[sp] : -1;
        }
        add_limits(Qx, k);

#pragma unroll
        for (int r = 0; r < 2; r++) *reinterpret_cast<double2*>(sQxu + (8 * r + g) * S::LDQ + 2 * t) = make_double2(Qxu[r][0], Qxu[r][1]);
        // [K|d] = -(Quu + 1e-6 I)^-1 [Qux|Qu]: explicit inverse (gj_inverse_regs; pivot rows travel through the sKd
        // area), then the product on DMMA                                                ILQRRecursive.cpp:89-91
        st_w(sM, Qa, g, t);
        double Iv[1][1][2] = {{{Lr[0], Lr[1]}}};
        bool okk = true;
        if constexpr (PIVOT) {
            okk = __all_sync(0xffffffffu, gj_inverse_regs<1, DOF, S::LDW>(Iv, sKd, m, g, t));
            if (!okk) {  // a pivot was not positive: Quu + 1e-6 I again, for the reference's pivoted LU inverse
                spd = false;
                Iv[0][0][0] = Lr[0];
                Iv[0][0][1] = Lr[1];
            }
        } else {
            spd = gj_inverse_regs<1, DOF, S::LDW>(Iv, sKd, m, g, t) && spd;
        }
        *reinterpret_cast<double2*>(sInv + g * S::LDQ + 2 * t) = make_double2(Iv[0][0][0], Iv[0][0][1]);
        *reinterpret_cast<double2*>(sQuu + g * S::LDQ + 2 * t) = make_double2(Quu[0], Quu[1]);  // un-regularised, for the P update
        if (PIVOT && !okk) wm_inverse_pivoted(sInv, S::LDQ, m, sKd, lane);
        __syncwarp();
        double Kd[2][2];
        {
            const double ia0 = sInv[g * S::LDQ + t], ia1 = sInv[g * S::LDQ + 4 + t];
#pragma unroll
            for (int c = 0; c < 2; c++) {
                double c0 = 0.0, c1 = 0.0;
                dmma884(c0, c1, ia0, ld_bw(sM, 0, c, g, t));
                dmma884(c0, c1, ia1, ld_bw(sM, 1, c, g, t));
                Kd[c][0] = -c0;
                Kd[c][1] = -c1;
            }
        }
#pragma unroll
        for (int c = 0; c < 2; c++)
#pragma unroll
            for (int h = 0; h < 2; h++)
                if (!rowv) Kd[c][h] = 0.0;
        st_w(sKd, Kd, g, t);
        if (rowv) {  // stream K (m x n, natural column order) and d
#pragma unroll
            for (int c = 0; c < 2; c++)
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int jl = 2 * t + h;
                    if (jl < dof) Kg[g * ns + c * dof + jl] = Kd[c][h];
                    else if (c == 1 && jl == 7) dg[g] = Kd[c][h];
                }
        }
        __syncwarp();

        // [P|p] = [Qxx|Qx] + K^T (Quu [K|d] + [Qux|Qu]) + Qxu [K|d]   (ILQRRecursive.cpp:93-94).  The bracket M is formed as
        // written, with the un-regularised Quu: this form is stationary in K, so the rounding error of the computed gains does not
        // enter P at first order (see backward_n1_body; the collapsed form M = -1e-6 [K|d] of round 1 lost that)
        double kb[2][2], mb[2][2], xa[2][2];  // B fragments of [K|d] (= A fragments of [K|d]^T) and of M; A fragments of Qxu
#pragma unroll
        for (int s = 0; s < 2; s++)
#pragma unroll
            for (int c = 0; c < 2; c++) {
                kb[s][c] = ld_bw(sKd, s, c, g, t);
                xa[s][c] = sQxu[(8 * c + g) * S::LDQ + 4 * s + t];
            }
        {
            const double qa0 = sQuu[g * S::LDQ + t], qa1 = sQuu[g * S::LDQ + 4 + t];
            double Mm[2][2];
#pragma unroll
            for (int c = 0; c < 2; c++) {
                double c0 = Qa[c][0], c1 = Qa[c][1];
                dmma884(c0, c1, qa0, kb[0][c]);
                dmma884(c0, c1, qa1, kb[1][c]);
                Mm[c][0] = c0;
                Mm[c][1] = c1;
            }
            __syncwarp();  // every lane has read [Qux|Qu] in sM
            st_w(sM, Mm, g, t);
            __syncwarp();
        }
#pragma unroll
        for (int s = 0; s < 2; s++)
#pragma unroll
            for (int c = 0; c < 2; c++) mb[s][c] = ld_bw(sM, s, c, g, t);
#pragma unroll
        for (int r = 0; r < 2; r++)
#pragma unroll
            for (int c = 0; c < 2; c++) {
                double c0 = Qx[r][c][0], c1 = Qx[r][c][1];
#pragma unroll
                for (int s = 0; s < 2; s++) dmma884(c0, c1, kb[s][r], mb[s][c]);
#pragma unroll
                for (int s = 0; s < 2; s++) dmma884(c0, c1, xa[s][r], kb[s][c]);
                // keep rows >= dof (incl. the d row the transposed fragment drags in) and unused columns at exact zero
                const int j0 = 2 * t;
                Pa[r][c][0] = (rowv && (j0 < dof || (c == 1 && j0 == 7))) ? c0 : 0.0;
                Pa[r][c][1] = (rowv && (j0 + 1 < dof || (c == 1 && j0 + 1 == 7))) ? c1 : 0.0;
            }
        __syncwarp();
    }
    return __all_sync(0xffffffffu, spd);
}
template <int DOF, bool AL>
__device__ __noinline__ void backward_n2_redo(const DevProblem& P, const Work& W, int b, int lane, double* wsmem, double penalty_frozen) {
    backward_n2_body<DOF, AL, true>(P, W, b, lane, wsmem, penalty_frozen);
}

template <int DOF, bool AL, int WARPS, int MINB>
__global__ void __launch_bounds__(32 * WARPS, MINB) k_backward_n2(const DevProblem* __restrict__ Pp, Work W, int b0, int nB, double penalty_frozen) {
    extern __shared__ double smem[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int bl = blockIdx.x * WARPS + wid;
    if (bl >= nB) return;
    const int b = b0 + bl;
    if (W.done[b]) return;
    double* ws = smem + (size_t)wid * BwdN2Smem::PER_WARP;
    if (!backward_n2_body<DOF, AL, false>(*Pp, W, b, lane, ws, penalty_frozen)) {
        if (lane == 0) atomicOr(W.status + b, ILQRB_ST_QUU_NOT_SPD);
        backward_n2_redo<DOF, AL>(*Pp, W, b, lane, ws, penalty_frozen);
    }
}

}  // namespace ilqrb
