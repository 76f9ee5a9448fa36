// wmat.cuh — warp-level dense FP64 algebra on small shared-memory matrices (one problem per warp).
//
// Matrices live in shared memory, row-major, dimensions padded to multiples of 8, leading dimension
// LD = cols + 4 doubles (LD mod 16 == 4 or 12 makes the mma fragment loads bank-conflict free).
// Products run on the FP64 tensor pipe: mma.sync.aligned.m8n8k4.row.col.f64 (DMMA).  tcgen05 has no f64 kind, so
// this warp-level shape is the Blackwell FP64 tensor path.
//
// Fragment ownership for lane = 4*g + t (g = lane>>2, t = lane&3):
//   A (8x4, row):  a  = A[g][t]
//   B (4x8, col):  b  = B[t][g]
//   C (8x8):       c0 = C[g][2t], c1 = C[g][2t+1]
#pragma once
#include <cuda_runtime.h>

namespace ilqrb {

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// element of op(M): M[i][j] or M[j][i]
template <bool TR>
__device__ __forceinline__ double wm_ld(const double* M, int ld, int i, int j) {
    return TR ? M[j * ld + i] : M[i * ld + j];
}

// acc(8x8 tile at (rt,ct)) += op(A)[rt, :] * op(B)[:, ct] over KT k-tiles
template <int KT, bool TA, bool TB>
__device__ __forceinline__ void wm_tile_acc(double& c0, double& c1, const double* A, int lda, const double* B, int ldb, int rt, int ct, int g, int t) {
#pragma unroll
    for (int kt = 0; kt < KT; kt++) {
#pragma unroll
        for (int s = 0; s < 2; s++) {
            const int kk = 8 * kt + 4 * s + t;
            const double a = wm_ld<TA>(A, lda, 8 * rt + g, kk);
            const double b = wm_ld<TB>(B, ldb, kk, 8 * ct + g);
            dmma884(c0, c1, a, b);
        }
    }
}

// C = [negate](op(A) * op(B)) [+ C_old if ACC]   — the product is summed from zero and then added to the old
// value (keeps the reference's  X = l + (A*B)  association).  RT x CT output tiles, KT inner tiles.
// Ends with __syncwarp().
template <int RT, int CT, int KT, bool TA, bool TB, bool ACC, bool NEG>
__device__ __forceinline__ void wm_gemm(double* C, int ldc, const double* A, int lda, const double* B, int ldb, int lane) {
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int rt = 0; rt < RT; rt++)
#pragma unroll
        for (int ct = 0; ct < CT; ct++) {
            double c0 = 0.0, c1 = 0.0;
            wm_tile_acc<KT, TA, TB>(c0, c1, A, lda, B, ldb, rt, ct, g, t);
            double2* dst = reinterpret_cast<double2*>(C + (8 * rt + g) * ldc + 8 * ct + 2 * t);
            if (NEG) {
                c0 = -c0;
                c1 = -c1;
            }
            if (ACC) {
                double2 old = *dst;
                c0 = old.x + c0;
                c1 = old.y + c1;
            }
            *dst = make_double2(c0, c1);
        }
    __syncwarp();
}

// every lane visits its share of an R x C element range: f(i, j)
template <class F>
__device__ __forceinline__ void wm_foreach(int R, int C, int lane, F f) {
    for (int e = lane; e < R * C; e += 32) f(e / C, e % C);
}

// In-place inverse of the leading m x m block of an (8*NT) x (8*NT) shared-memory matrix by Gauss-Jordan
// elimination without pivoting (the matrix is Quu + 1e-6 I, symmetric positive definite in normal operation —
// replaces Eigen's PartialPivLU inverse at ILQRRecursive.cpp:89 / AL-ILQR.cpp:136).  Rows/cols >= m must hold the
// identity pattern; they are left untouched.  Returns false (warp-uniform) if a pivot was <= 0 or NaN.
template <int NT>
__device__ __forceinline__ bool wm_inverse_spd(double* M, int ld, int m, int lane) {
    const int g = lane >> 2, t = lane & 3;
    double e[NT][NT][2];
#pragma unroll
    for (int rt = 0; rt < NT; rt++)
#pragma unroll
        for (int ct = 0; ct < NT; ct++) {
            double2 v = *reinterpret_cast<const double2*>(M + (8 * rt + g) * ld + 8 * ct + 2 * t);
            e[rt][ct][0] = v.x;
            e[rt][ct][1] = v.y;
        }
    bool ok = true;
    for (int k = 0; k < m; k++) {
        const double pkk = M[k * ld + k];
        double rowk[NT][2], colk[NT];
#pragma unroll
        for (int ct = 0; ct < NT; ct++) {
            double2 v = *reinterpret_cast<const double2*>(M + k * ld + 8 * ct + 2 * t);
            rowk[ct][0] = v.x;
            rowk[ct][1] = v.y;
        }
#pragma unroll
        for (int rt = 0; rt < NT; rt++) colk[rt] = M[(8 * rt + g) * ld + k];
        __syncwarp();
        if (!(pkk > 0.0)) ok = false;
        const double piv = 1.0 / pkk;
#pragma unroll
        for (int rt = 0; rt < NT; rt++) {
            const int i = 8 * rt + g;
#pragma unroll
            for (int ct = 0; ct < NT; ct++) {
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int j = 8 * ct + 2 * t + h;
                    const double rkj = rowk[ct][h] * piv;
                    double v;
                    if (i == k) v = (j == k) ? piv : rkj;
                    else if (j == k) v = -colk[rt] * piv;
                    else v = e[rt][ct][h] - colk[rt] * rkj;
                    if (i >= m || j >= m) v = e[rt][ct][h];
                    e[rt][ct][h] = v;
                }
                *reinterpret_cast<double2*>(M + i * ld + 8 * ct + 2 * t) = make_double2(e[rt][ct][0], e[rt][ct][1]);
            }
        }
        __syncwarp();
    }
    return ok;
}

// Rare path: in-place inverse of the leading m x m block of a shared-memory matrix WITH partial pivoting — what the reference
// computes at ILQRRecursive.cpp:89 / AL-ILQR.cpp:136 (Eigen's PartialPivLU(m).inverse()): unblocked right-looking LU, the first
// maximum of |a_ik| as pivot, then unit-lower forward and upper backward substitution on P I.  Taken when the unpivoted
// elimination of the fast kernels met a pivot that was not positive (Quu + 1e-6 I indefinite or nearly singular), so that such a
// step returns the reference's answer instead of an unpivoted one.  Lane 0 works, the warp waits; X: m x ld scratch.
// Rows / columns >= m are left untouched.
static __device__ __noinline__ void wm_inverse_pivoted(double* M, int ld, int m, double* X, int lane) {
    __syncwarp();
    if (lane == 0) {
        int perm[16];
        for (int i = 0; i < m; i++) perm[i] = i;
        for (int k = 0; k < m; k++) {
            int piv = k;
            double best = fabs(M[k * ld + k]);
            for (int i = k + 1; i < m; i++) {
                const double v = fabs(M[i * ld + k]);
                if (v > best) {
                    best = v;
                    piv = i;
                }
            }
            if (piv != k) {
                for (int j = 0; j < m; j++) {
                    const double tmp = M[k * ld + j];
                    M[k * ld + j] = M[piv * ld + j];
                    M[piv * ld + j] = tmp;
                }
                const int tp = perm[k];
                perm[k] = perm[piv];
                perm[piv] = tp;
            }
            const double pk = M[k * ld + k];
            for (int i = k + 1; i < m; i++) M[i * ld + k] /= pk;
            for (int i = k + 1; i < m; i++) {
                const double lik = M[i * ld + k];
                for (int j = k + 1; j < m; j++) M[i * ld + j] -= lik * M[k * ld + j];
            }
        }
        for (int i = 0; i < m; i++)
            for (int j = 0; j < m; j++) X[i * ld + j] = (perm[i] == j) ? 1.0 : 0.0;  // P I
        for (int i = 0; i < m; i++)  // L y = P I (unit diagonal)
            for (int k = 0; k < i; k++) {
                const double lik = M[i * ld + k];
                if (lik == 0.0) continue;
                for (int j = 0; j < m; j++) X[i * ld + j] -= lik * X[k * ld + j];
            }
        for (int i = m - 1; i >= 0; i--) {  // U x = y
            for (int k = i + 1; k < m; k++) {
                const double uik = M[i * ld + k];
                if (uik == 0.0) continue;
                for (int j = 0; j < m; j++) X[i * ld + j] -= uik * X[k * ld + j];
            }
            const double uii = M[i * ld + i];
            for (int j = 0; j < m; j++) X[i * ld + j] /= uii;
        }
        for (int i = 0; i < m; i++)
            for (int j = 0; j < m; j++) M[i * ld + j] = X[i * ld + j];
    }
    __syncwarp();
}

}  // namespace ilqrb
