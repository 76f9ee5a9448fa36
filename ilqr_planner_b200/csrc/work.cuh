// work.cuh — device work buffers of one handle and the small helpers every kernel family of the Recursive / AL path shares
// (state initialisation and the integrator step in the reference's operation order).
#pragma once
#include "kin.cuh"

namespace ilqrb {

constexpr int NCAND = ILQRB_N_ALPHA;  // 11
constexpr int CAND_LANES = 16;        // lanes reserved per problem in k_linesearch
constexpr int MAX_NC = 32;            // AL constraint rows per step
constexpr int MAX_KP_STEPS = ILQRB_MAX_TASKS * ILQRB_MAX_KEYPOINTS;  // distinct keypoint timesteps of a system

// device work buffers of one handle (padded strides: ns = even(n), ms = even(m))
struct Work {
    int ns, ms;            // padded row strides of X / U,d ; K rows have stride ns
    double* X;             // [B][H][ns]
    double* U;             // [B][H-1][ms]
    double* K;             // [B][H-1][m][ns]   time-forward index
    double* d;             // [B][H-1][ms]
    double* Uc;            // [B][H-1][NCAND][ms]  candidate controls
    double* cost_c;        // [B][NCAND]
    double* dun_c;         // [B][NCAND]  sum_k |du_k|
    double* cost0;         // [B]
    double* LX;            // [B][S][n]        keypoint l_x
    double* LXX;           // [B][S][n][n]     keypoint l_xx
    double* cost_hist;     // [B][nb_iter]
    double* alpha_hist;    // [B][nb_iter]
    double* alpha_last;    // [B]
    double* DU;            // [B][H-1][ms]     alpha = 1 control deviation of the forward pass (fused1_fwd.cuh)
    double* DXk;           // [B][S][ns]       alpha = 1 state deviation at the keypoint steps
    int* viol;             // [B]              a joint limit is violated at an end of some candidate segment
    int* done;             // [B]
    int* iters;            // [B]
    int* status;           // [B]
    // AL
    int nc, al_per_step;
    const double* alA;     // [(H-1)|1][nc][n+m]
    const double* alb;     // [(H-1)|1][nc]
    double* lam;           // [B][H-1][nc]
    double* Cs;            // [B][H-1][nc]
    unsigned* mask;        // [B][H-1]  active-set bits (Ik diagonal without the penalty factor)
};

__device__ __forceinline__ double alpha_of(int c) { return 1.0 / (double)(1 << c); }

// x0 = [q0, (dq0), (0)]   (PosOrnPlannerSys.cpp:61-69, PosOrnTimePlannerSys.cpp:55-62)
template <int N>
__device__ __forceinline__ void load_x0(const DevProblem& P, int b, double* x) {
#pragma unroll
    for (int i = 0; i < N; i++) x[i] = 0.0;
#pragma unroll
    for (int i = 0; i < N; i++) {
        if (i < P.dof) x[i] = P.q0[(size_t)b * P.dof + i];
        else if (P.nb_deriv == 2 && i < 2 * P.dof) x[i] = P.dq0[(size_t)b * P.dof + i - P.dof];
    }
}

// one integrator step with the reference's operation order and no FMA contraction
// (SimulationInterface.cpp:19-31: q += dt*dq + dt*dt/2*ddq ; dq += dt*ddq ; t += dt)
template <int N>
__device__ __forceinline__ void step_state(const DevProblem& P, double* x, const double* u) {
    const int dof = P.dof;
    double dt = P.dt;
    if (P.time) {
        const double s = u[P.m - 1];
        dt = __dmul_rn(s, s);
    }
    const double hdt2 = __dmul_rn(dt, dt) / 2;
    if (P.nb_deriv == 1) {
#pragma unroll
        for (int i = 0; i < N; i++)
            if (i < dof) x[i] = __dadd_rn(x[i], __dadd_rn(__dmul_rn(dt, u[i]), __dmul_rn(hdt2, 0.0)));
    } else {
#pragma unroll
        for (int i = 0; i < N / 2; i++)
            if (i < dof) {
                // x[dof+i] with a compile-time-bounded scan keeps x in registers
                double v = 0.0;
#pragma unroll
                for (int j = 0; j < N; j++)
                    if (j == dof + i) v = x[j];
                x[i] = __dadd_rn(x[i], __dadd_rn(__dmul_rn(dt, v), __dmul_rn(hdt2, u[i])));
                const double vn = __dadd_rn(v, __dmul_rn(dt, u[i]));
#pragma unroll
                for (int j = 0; j < N; j++)
                    if (j == dof + i) x[j] = vn;
            }
    }
    if (P.time) {
#pragma unroll
        for (int j = 0; j < N; j++)
            if (j == P.n - 1) x[j] = __dadd_rn(x[j], dt);
    }
}

}  // namespace ilqrb
