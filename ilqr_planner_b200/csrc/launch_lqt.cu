// launch_lqt.cu — solver::LQT (lqt.h:24-88, lqt.cpp:16-126) for a batch of independent linear tracking problems:
// one warp per problem, every matrix of the recursion in that warp's shared memory, products on the FP64 tensor pipe (wmat.cuh).
// One backward sweep produces what the reference's two solvers produce:
//   * solveDP's P_t, d_t exactly as lqt.cpp:41-49 forms them, and the gain / feed-forward getCommand(timestep, state) builds from them
//     (lqt.cpp:105-124, including its use of mu_{t+1} in both places of  A mu - mu);
//   * solveLinAl's controls: the reference inverts the dense (T-1)m x (T-1)m matrix Su^T Q Su + R (lqt.cpp:95).  P_t above IS the
//     Riccati matrix of that objective, so the same minimiser comes from the exact linear term p_t (p = Q mu + (A - B K)^T p') and a
//     forward rollout from x_0 = mu_0 — O(T n^3) instead of O((T m)^3).  tests/test_gpu_lqt.py holds it to the oracle's dense form.
// Not on the iLQR solve path; stateless entry points that stage their operands per call.
#include <vector>

#include "handle.hpp"
#include "wmat.cuh"

namespace ilqrb {
namespace {

struct DevBytes {  // one cudaMalloc per call, freed on scope exit
    void* p = nullptr;
    ~DevBytes() {
        if (p) cudaFree(p);
    }
};

template <int NT>
struct LqtSmem {
    static constexpr int D = 8 * NT, LD = D + 4, MAT = D * LD;
    enum { mA, mB, mP, mPB, mG, mM1, mM2, mM3, mK, mT, NMAT };
    enum { vMuI, vMuN, vD, vP, vV, vW, vT1, vT2, NVEC };
    static constexpr int DOUBLES = NMAT * MAT + NVEC * D;
};

template <int NT>
__global__ void __launch_bounds__(32) k_lqt(int n, int m, int T, int ab_per, const double* __restrict__ Ag, const double* __restrict__ Bg, const double* __restrict__ Qs,
                                            const double* __restrict__ mu, double r, double* __restrict__ Ps, double* __restrict__ ds, double* Kg, double* __restrict__ fg,
                                            double* kg, double* __restrict__ u, double* __restrict__ xhat, int* __restrict__ status) {
    using S = LqtSmem<NT>;
    constexpr int D = S::D, LD = S::LD;
    extern __shared__ __align__(16) double lqt_sm[];
    const int q = blockIdx.x, lane = threadIdx.x;
    auto mat = [&](int i) { return lqt_sm + i * S::MAT; };
    auto vec = [&](int i) { return lqt_sm + S::NMAT * S::MAT + i * D; };
    double *A = mat(S::mA), *B = mat(S::mB), *P = mat(S::mP), *PB = mat(S::mPB), *G = mat(S::mG), *M1 = mat(S::mM1), *M2 = mat(S::mM2), *M3 = mat(S::mM3),
           *K = mat(S::mK), *Tm = mat(S::mT);
    double *muI = vec(S::vMuI), *muN = vec(S::vMuN), *d = vec(S::vD), *p = vec(S::vP), *v = vec(S::vV), *w = vec(S::vW), *t1 = vec(S::vT1), *t2 = vec(S::vT2);
    for (int e = lane; e < S::DOUBLES; e += 32) lqt_sm[e] = 0.0;
    __syncwarp();
    const double* Aq = Ag + (ab_per ? (size_t)q * n * n : 0);
    const double* Bq = Bg + (ab_per ? (size_t)q * n * m : 0);
    const double* Qq = Qs + (size_t)q * T * n * n;
    const double* muq = mu + (size_t)q * T * n;
    double* Pq = Ps + (size_t)q * T * n * n;
    double* dq = ds + (size_t)q * T * n;
    double* Kq = Kg + (size_t)q * (T - 1) * m * n;
    double* fq = fg + (size_t)q * (T - 1) * m;
    double* kq = kg + (size_t)q * (T - 1) * m;
    wm_foreach(n, n, lane, [&](int i, int j) {
        A[i * LD + j] = Aq[i * n + j];
        const double qv = Qq[(size_t)(T - 1) * n * n + i * n + j];
        P[i * LD + j] = qv;  // PT = Qs.back()   lqt.cpp:35
        Pq[(size_t)(T - 1) * n * n + i * n + j] = qv;
    });
    wm_foreach(n, m, lane, [&](int i, int j) { B[i * LD + j] = Bq[i * m + j]; });
    __syncwarp();
    if (lane < n) {
        double s = 0.0;
        for (int j = 0; j < n; j++) s += P[lane * LD + j] * muq[(size_t)(T - 1) * n + j];
        p[lane] = s;
        d[lane] = 0.0;  // dT = 0   lqt.cpp:36
        dq[(size_t)(T - 1) * n + lane] = 0.0;
    }
    __syncwarp();
    int st = 0;
    for (int i = T - 2; i >= 0; i--) {
        const double* Qi = Qq + (size_t)i * n * n;
        if (lane < n) {
            muI[lane] = muq[(size_t)i * n + lane];
            muN[lane] = muq[(size_t)(i + 1) * n + lane];
        }
        // (B^T P' B + Rt)^-1   lqt.cpp:46
        wm_gemm<NT, NT, NT, false, false, false, false>(PB, LD, P, LD, B, LD, lane);
        wm_gemm<NT, NT, NT, true, false, false, false>(G, LD, B, LD, PB, LD, lane);
        if (lane < D) G[lane * LD + lane] = (lane < m) ? G[lane * LD + lane] + r : 1.0;
        __syncwarp();
        wm_foreach(D, D, lane, [&](int a, int b) { Tm[a * LD + b] = G[a * LD + b]; });
        __syncwarp();
        if (!wm_inverse_spd<NT>(G, LD, m, lane)) {  // what the reference computes is PartialPivLU's inverse: take it when a pivot was not positive
            wm_foreach(D, D, lane, [&](int a, int b) { G[a * LD + b] = Tm[a * LD + b]; });
            wm_inverse_pivoted(G, LD, m, M1, lane);
            st |= ILQRB_ST_QUU_NOT_SPD;
        }
        // P' B G^-1 B^T P' - P'   lqt.cpp:46
        wm_gemm<NT, NT, NT, false, false, false, false>(M1, LD, PB, LD, G, LD, lane);
        wm_gemm<NT, NT, NT, false, true, false, false>(M2, LD, M1, LD, B, LD, lane);
        wm_gemm<NT, NT, NT, false, false, false, false>(M3, LD, M2, LD, P, LD, lane);
        wm_foreach(D, D, lane, [&](int a, int b) { M3[a * LD + b] -= P[a * LD + b]; });
        // v = P' (A mu_{i+1} - mu_{i+1}) + d'  (getCommand, lqt.cpp:119-120);  w = P' (A mu_i - mu_{i+1}) + d'  (solveDP, lqt.cpp:48)
        if (lane < n) {
            double a1 = 0.0, a2 = 0.0;
            for (int j = 0; j < n; j++) {
                a1 += A[lane * LD + j] * muN[j];
                a2 += A[lane * LD + j] * muI[j];
            }
            t1[lane] = a1 - muN[lane];
            t2[lane] = a2 - muN[lane];
        }
        __syncwarp();
        if (lane < n) {
            double s1 = 0.0, s2 = 0.0;
            for (int j = 0; j < n; j++) {
                s1 += P[lane * LD + j] * t1[j];
                s2 += P[lane * LD + j] * t2[j];
            }
            v[lane] = s1 + d[lane];
            w[lane] = s2 + d[lane];
        }
        __syncwarp();
        // K = G^-1 B^T P' A   lqt.cpp:117
        wm_gemm<NT, NT, NT, true, false, false, false>(Tm, LD, PB, LD, A, LD, lane);
        wm_gemm<NT, NT, NT, false, false, false, false>(K, LD, G, LD, Tm, LD, lane);
        wm_foreach(m, n, lane, [&](int a, int b) { Kq[(size_t)i * m * n + a * n + b] = K[a * LD + b]; });
        // f = -G^-1 B^T v (lqt.cpp:118-120);  k = G^-1 B^T p' (the exact feed-forward of the tracking objective)
        if (lane < m) {
            double s1 = 0.0, s2 = 0.0;
            for (int j = 0; j < n; j++) {
                s1 += B[j * LD + lane] * v[j];
                s2 += B[j * LD + lane] * p[j];
            }
            t1[lane] = s1;
            t2[lane] = s2;
        }
        __syncwarp();
        if (lane < m) {
            double s1 = 0.0, s2 = 0.0;
            for (int j = 0; j < m; j++) {
                s1 += G[lane * LD + j] * t1[j];
                s2 += G[lane * LD + j] * t2[j];
            }
            fq[(size_t)i * m + lane] = -s1;
            kq[(size_t)i * m + lane] = s2;
        }
        __syncwarp();
        // d = (A^T - A^T P' B G^-1 B^T) w = A^T (w - M2 w)   lqt.cpp:47-48
        if (lane < n) {
            double s = 0.0;
            for (int j = 0; j < n; j++) s += M2[lane * LD + j] * w[j];
            t1[lane] = w[lane] - s;
        }
        // A - B K
        wm_gemm<NT, NT, NT, false, false, false, true>(Tm, LD, B, LD, K, LD, lane);
        wm_foreach(n, n, lane, [&](int a, int b) { Tm[a * LD + b] += A[a * LD + b]; });
        __syncwarp();
        double dn = 0.0, pn = 0.0;
        if (lane < n) {
            for (int j = 0; j < n; j++) {
                dn += A[j * LD + lane] * t1[j];
                pn += Tm[j * LD + lane] * p[j];
            }
            double s = 0.0;
            for (int j = 0; j < n; j++) s += Qi[lane * n + j] * muI[j];
            pn = s + pn;
        }
        __syncwarp();
        if (lane < n) {
            d[lane] = dn;
            p[lane] = pn;
            dq[(size_t)i * n + lane] = dn;
        }
        // P = Q_i - A^T (...) A   lqt.cpp:46
        wm_gemm<NT, NT, NT, true, false, false, false>(M1, LD, A, LD, M3, LD, lane);
        wm_gemm<NT, NT, NT, false, false, false, true>(P, LD, M1, LD, A, LD, lane);
        wm_foreach(n, n, lane, [&](int a, int b) {
            const double pv = Qi[a * n + b] + P[a * LD + b];
            P[a * LD + b] = pv;
            Pq[(size_t)i * n * n + a * n + b] = pv;
        });
        __syncwarp();
    }
    // the minimiser of solveLinAl's objective and its states: x_0 = mu_0, u_t = k_t - K_t x_t
    double* x = v;
    if (lane < n) x[lane] = muq[lane];
    __syncwarp();
    for (int i = 0; i < T - 1; i++) {
        if (lane < n) xhat[(size_t)q * T * n + (size_t)i * n + lane] = x[lane];
        if (lane < m) {
            double s = 0.0;
            for (int j = 0; j < n; j++) s += Kq[(size_t)i * m * n + lane * n + j] * x[j];
            const double uu = kq[(size_t)i * m + lane] - s;
            w[lane] = uu;
            u[(size_t)q * (T - 1) * m + (size_t)i * m + lane] = uu;
        }
        __syncwarp();
        double xn = 0.0;
        if (lane < n) {
            double s1 = 0.0, s2 = 0.0;
            for (int j = 0; j < n; j++) s1 += A[lane * LD + j] * x[j];
            for (int j = 0; j < m; j++) s2 += B[lane * LD + j] * w[j];
            xn = s1 + s2;
        }
        __syncwarp();
        if (lane < n) x[lane] = xn;
        __syncwarp();
    }
    if (lane < n) xhat[(size_t)q * T * n + (size_t)(T - 1) * n + lane] = x[lane];
    if (lane == 0) status[q] = st;
}

// one thread per query: u = K (mu_{ts+1} - state) + f   lqt.cpp:121
__global__ void k_lqt_command(int n, int m, int T, int nc, const double* __restrict__ K, const double* __restrict__ f, const double* __restrict__ mu,
                              const int* __restrict__ prob, const int* __restrict__ ts, const double* __restrict__ state, double* __restrict__ u) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= nc) return;
    const int q = prob ? prob[c] : 0, t = ts[c];
    const double* Kt = K + ((size_t)q * (T - 1) + t) * m * n;
    const double* ft = f + ((size_t)q * (T - 1) + t) * m;
    const double* mt = mu + (size_t)q * T * n + (size_t)(t + 1) * n;
    for (int a = 0; a < m; a++) {
        double s = 0.0;
        for (int j = 0; j < n; j++) s += Kt[a * n + j] * (mt[j] - state[(size_t)c * n + j]);
        u[(size_t)c * m + a] = s + ft[a];
    }
}

template <int NT>
int launch_lqt(int nq, int n, int m, int T, int ab_per, const double* A, const double* B, const double* Qs, const double* mu, double r, double* Ps, double* ds, double* K,
               double* f, double* kk, double* u, double* xhat, int* status) {
    const size_t smem = (size_t)LqtSmem<NT>::DOUBLES * sizeof(double);
    k_lqt<NT><<<nq, 32, smem>>>(n, m, T, ab_per, A, B, Qs, mu, r, Ps, ds, K, f, kk, u, xhat, status);
    CU(cudaGetLastError());
    return 0;
}

}  // namespace
}  // namespace ilqrb

using namespace ilqrb;

extern "C" int ilqrb_lqt_solve(int device, const ilqrb_lqt_desc* d, double* Ps, double* ds, double* K, double* f, double* u, double* xhat, int32_t* status) {
    if (!d || !d->A || !d->B || !d->Qs || !d->mu) return fail(ILQRB_EINVAL, "ilqrb_lqt_solve: null descriptor or operand");
    const int nq = d->nq, n = d->n, m = d->m, T = d->T;
    if (nq < 1 || n < 1 || m < 1 || n > 16 || m > 16) return fail(ILQRB_EINVAL, "ilqrb_lqt_solve: 1 <= nb_state_var, nb_ctrl_var <= 16 and at least one problem expected");
    if (T < 2) return fail(ILQRB_EINVAL, "ilqrb_lqt_solve: at least two states expected");
    CU(cudaSetDevice(device));
    const size_t N = (size_t)nq, nA = (d->ab_per_problem ? N : 1) * n * n, nB = (d->ab_per_problem ? N : 1) * n * m, nQ = N * T * n * n, nMu = N * T * n;
    const size_t nK = N * (T - 1) * m * n, nf = N * (T - 1) * m, nd = N * T * n;
    const size_t tot = nA + nB + nQ + nMu + nQ /*Ps*/ + nd /*ds*/ + nK + 2 * nf + nf /*u*/ + nd /*xhat*/;
    DevBytes buf, sbuf;
    CU(cudaMalloc(&buf.p, tot * sizeof(double)));
    CU(cudaMalloc(&sbuf.p, N * sizeof(int)));
    double* dA = static_cast<double*>(buf.p);
    double *dB = dA + nA, *dQ = dB + nB, *dMu = dQ + nQ, *dPs = dMu + nMu, *dds = dPs + nQ, *dK = dds + nd, *df = dK + nK, *dk = df + nf, *du = dk + nf, *dx = du + nf;
    CU(cudaMemcpy(dA, d->A, nA * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(dB, d->B, nB * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(dQ, d->Qs, nQ * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(dMu, d->mu, nMu * sizeof(double), cudaMemcpyHostToDevice));
    int* dst = static_cast<int*>(sbuf.p);
    const int rc = (n <= 8 && m <= 8) ? launch_lqt<1>(nq, n, m, T, d->ab_per_problem, dA, dB, dQ, dMu, d->r, dPs, dds, dK, df, dk, du, dx, dst)
                                      : launch_lqt<2>(nq, n, m, T, d->ab_per_problem, dA, dB, dQ, dMu, d->r, dPs, dds, dK, df, dk, du, dx, dst);
    if (rc) return rc;
    if (Ps) CU(cudaMemcpy(Ps, dPs, nQ * sizeof(double), cudaMemcpyDeviceToHost));
    if (ds) CU(cudaMemcpy(ds, dds, nd * sizeof(double), cudaMemcpyDeviceToHost));
    if (K) CU(cudaMemcpy(K, dK, nK * sizeof(double), cudaMemcpyDeviceToHost));
    if (f) CU(cudaMemcpy(f, df, nf * sizeof(double), cudaMemcpyDeviceToHost));
    if (u) CU(cudaMemcpy(u, du, nf * sizeof(double), cudaMemcpyDeviceToHost));
    if (xhat) CU(cudaMemcpy(xhat, dx, nd * sizeof(double), cudaMemcpyDeviceToHost));
    if (status) CU(cudaMemcpy(status, dst, N * sizeof(int), cudaMemcpyDeviceToHost));
    CU(cudaDeviceSynchronize());
    return 0;
}

extern "C" int ilqrb_lqt_command(int device, int nq, int n, int m, int T, const double* K, const double* f, const double* mu, int nc, const int32_t* problem,
                                 const int32_t* timestep, const double* state, double* u) {
    if (!K || !f || !mu || !timestep || !state || !u) return fail(ILQRB_EINVAL, "ilqrb_lqt_command: null operand");
    if (nq < 1 || nc < 1 || n < 1 || m < 1 || T < 2) return fail(ILQRB_EINVAL, "ilqrb_lqt_command: empty problem or query set");
    for (int c = 0; c < nc; c++) {
        // the reference reads Ps.at(nb_states - timestep - 2) (lqt.cpp:106-114): anything outside [0, T-2] throws there
        if (timestep[c] < 0 || timestep[c] > T - 2) return fail(ILQRB_EINVAL, "ilqrb_lqt_command: timestep outside [0, nb_states - 2]");
        if (problem && (problem[c] < 0 || problem[c] >= nq)) return fail(ILQRB_EINVAL, "ilqrb_lqt_command: problem index out of range");
    }
    CU(cudaSetDevice(device));
    const size_t N = (size_t)nq, nK = N * (T - 1) * m * n, nf = N * (T - 1) * m, nMu = N * T * n, nS = (size_t)nc * n, nU = (size_t)nc * m;
    DevBytes buf, ibuf;
    CU(cudaMalloc(&buf.p, (nK + nf + nMu + nS + nU) * sizeof(double)));
    CU(cudaMalloc(&ibuf.p, 2 * (size_t)nc * sizeof(int)));
    double* dK = static_cast<double*>(buf.p);
    double *df = dK + nK, *dMu = df + nf, *dS = dMu + nMu, *dU = dS + nS;
    int* dts = static_cast<int*>(ibuf.p);
    int* dpr = dts + nc;
    CU(cudaMemcpy(dK, K, nK * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(df, f, nf * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(dMu, mu, nMu * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(dS, state, nS * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(dts, timestep, (size_t)nc * sizeof(int), cudaMemcpyHostToDevice));
    if (problem) CU(cudaMemcpy(dpr, problem, (size_t)nc * sizeof(int), cudaMemcpyHostToDevice));
    k_lqt_command<<<(nc + 63) / 64, 64>>>(n, m, T, nc, dK, df, dMu, problem ? dpr : nullptr, dts, dS, dU);
    CU(cudaGetLastError());
    CU(cudaMemcpy(u, dU, nU * sizeof(double), cudaMemcpyDeviceToHost));
    return 0;
}
