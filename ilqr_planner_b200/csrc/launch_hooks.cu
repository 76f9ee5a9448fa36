// launch_hooks.cu — single-point entry points behind the reference's debugging API: System::getFxJac(xk), System::diff,
// System::cost / cost_x / cost_xx (and the cost_F* forms), Keypoint::diff and the utils::Sd functions, each evaluated by the same
// device functions the solvers call (kin.cuh), one thread per query.  Nothing here is on the solve path; every call stages its
// operands in a buffer of its own and synchronises.
#include <cstring>
#include <string>
#include <vector>

#include "handle.hpp"
#include "kin.cuh"

namespace ilqrb {
namespace {

struct DevBuf {  // one cudaMalloc per call, freed on scope exit
    double* p = nullptr;
    ~DevBuf() {
        if (p) cudaFree(p);
    }
};

// one thread per query point: x [n], u [m] (nullptr = zeros), timestep k, problem b (whose keypoint targets / precisions apply)
__global__ void k_eval_point(const DevProblem* __restrict__ Pp, int nq, const int* __restrict__ prob, const int* __restrict__ ks, const double* __restrict__ xin,
                             const double* __restrict__ uin, const double* __restrict__ fx_in, double* __restrict__ fx_out, double* __restrict__ J_out,
                             double* __restrict__ diff_out, double* __restrict__ c_out, double* __restrict__ cx_out, double* __restrict__ cxx_out) {
    const DevProblem& P = *Pp;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    const int n = P.n, m = P.m, dof = P.dof, b = prob ? prob[i] : 0, k = ks ? ks[i] : -1;
    double xs[MAXN], us[MAXM], dqz[MAXM];
    for (int j = 0; j < n; j++) xs[j] = xin[(size_t)i * n + j];
    for (int j = 0; j < m; j++) us[j] = uin ? uin[(size_t)i * m + j] : 0.0;
    for (int j = 0; j < dof; j++) dqz[j] = 0.0;
    double* fxo = fx_out + (size_t)i * P.nt;
    double* Jo = J_out + (size_t)i * P.r * n;
    double* dfo = diff_out + (size_t)i * P.r;
    double* cx = cx_out + (size_t)i * n;
    double* cxx = cxx_out + (size_t)i * n * n;
    for (int j = 0; j < P.r * n; j++) Jo[j] = 0.0;
    for (int j = 0; j < P.r; j++) dfo[j] = 0.0;
    for (int j = 0; j < n; j++) cx[j] = 0.0;
    for (int j = 0; j < n * n; j++) cxx[j] = 0.0;
    double c = 0.0;
    for (int t = 0; t < P.n_tasks; t++) {
        const DevTask& tk = P.tasks[t];
        EEState e;
        task_ee(P, tk, xs, P.nb_deriv == 2 ? xs + dof : dqz, e);
        double fx[MAX_NTT], res[MAX_RT];
        task_fx(P, tk, e, P.time ? xs[n - 1] : 0.0, fx);
        for (int j = 0; j < tk.nt_t; j++) fxo[tk.fx_off + j] = fx[j];
        double* J = Jo + (size_t)tk.r_off * n;  // System::getFxJac(xk): rows of this sub-system (SequentialSystem.cpp:93-113)
        task_jacobian(P, tk, e, J, n);
        const int kp = task_kp_at(tk, k);
        const int r = tk.r_t;
        double ct = 0.0;
        if (kp >= 0) {
            const double* tg = tk.kp_target + ((size_t)b * tk.n_keypoints + kp) * ILQRB_KP_STRIDE;
            const double* Q = tk.kp_precision + ((size_t)(tk.precision_per_problem ? b : 0) * tk.n_keypoints + kp) * r * r;
            // System::diff(actual_state, k) on the caller's task vector when one is given (System.cpp:103-109)
            if (fx_in) {
                double dres[MAX_RT];
                task_residual(P, tk, fx_in + (size_t)i * P.nt + tk.fx_off, tg, dres, task_deadzone(tk, kp));
                for (int j = 0; j < r; j++) dfo[tk.r_off + j] = dres[j];
            }
            task_residual(P, tk, fx, tg, res, task_deadzone(tk, kp));
            if (!fx_in)
                for (int j = 0; j < r; j++) dfo[tk.r_off + j] = res[j];
            // System::cost (System.cpp:213-234)
            double dQd = 0.0;
            for (int j = 0; j < r; j++) {
                double v = 0.0;
                for (int a = 0; a < r; a++) v += res[a] * Q[a * r + j];
                dQd += v * res[j];
            }
            double uRu = 0.0;
            for (int j = 0; j < m; j++) uRu += (us[j] * tk.r_diag[j]) * us[j];
            ct += dQd + uRu;
            // cost_x = (-J^T Q) diff ; cost_xx = (J^T Q) J (System.cpp:248-272, 286-308)
            for (int a = 0; a < n; a++) {
                double jq[MAX_RT], njq[MAX_RT];
                for (int cc = 0; cc < r; cc++) {
                    double s = 0.0, ns_ = 0.0;
                    for (int a2 = 0; a2 < r; a2++) {
                        s += J[a2 * n + a] * Q[a2 * r + cc];
                        ns_ += (-1.0 * J[a2 * n + a]) * Q[a2 * r + cc];
                    }
                    jq[cc] = s;
                    njq[cc] = ns_;
                }
                double g = 0.0;
                for (int cc = 0; cc < r; cc++) g += njq[cc] * res[cc];
                cx[a] += g;
                for (int j = 0; j < n; j++) {
                    double hh = 0.0;
                    for (int cc = 0; cc < r; cc++) hh += jq[cc] * J[cc * n + j];
                    cxx[a * n + j] += hh;
                }
            }
        }  // no keypoint at k: System::diff returns zeros, the cost has no task term
        if (tk.limits_set) {  // System::inspectJointLimit terms (System.cpp:121-142, 226-232, 263-269, 299-305)
            double qLq = 0.0;
            for (int a = 0; a < n; a++) {
                bool act;
                const double q = limit_q(tk, a, xs[a], act);
                if (act) {
                    qLq += (q * 1.0) * q;
                    cx[a] += -1.0 * q;
                    cxx[a * n + a] += 1.0 * 1.0;
                }
            }
            ct += qLq;
        }
        c += ct;
    }
    c_out[i] = c;
}

// Keypoint::diff without a system: the caller's record, dead zone and task vectors; P holds only the fields task_residual reads
__global__ void k_keypoint_diff(const DevProblem* __restrict__ Pp, const double* __restrict__ tg, const double* __restrict__ dz, int nq, int nt, int r,
                                const double* __restrict__ fx, double* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    double f[MAX_NTT], res[MAX_RT];
    for (int j = 0; j < nt; j++) f[j] = fx[(size_t)i * nt + j];
    task_residual(*Pp, Pp->tasks[0], f, tg, res, dz);
    for (int j = 0; j < r; j++) out[(size_t)i * r + j] = res[j];
}

// sd.h:34-45 (expMap): the point reached from `base` along the tangent vector u
__device__ void sd_expmap(const double* base_in, const double* u, double* out) {
    const double nb = sqrt(dot4(base_in, base_in));
    double base[4];
    for (int i = 0; i < 4; i++) base[i] = base_in[i] / nb;
    const double nu = sqrt(dot4(u, u));
    if (nu == 0) {
        for (int i = 0; i < 4; i++) out[i] = base[i];
        return;
    }
    double sn, cs, v[4];
    sincos(nu, &sn, &cs);
    for (int i = 0; i < 4; i++) v[i] = base[i] * cs + u[i] / nu * sn;
    const double nv = sqrt(dot4(v, v));
    for (int i = 0; i < 4; i++) out[i] = v[i] / nv;
}

__global__ void k_sd(int op, int nq, const double* __restrict__ a, const double* __restrict__ b, const double* __restrict__ c, double* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq) return;
    const double *ai = a + 4 * (size_t)i, *bi = b ? b + 4 * (size_t)i : nullptr, *ci = c ? c + 4 * (size_t)i : nullptr;
    switch (op) {
        case ILQRB_SD_LOGMAP: sd_logmap(ai, bi, out + 4 * (size_t)i); break;
        case ILQRB_SD_EXPMAP: sd_expmap(ai, bi, out + 4 * (size_t)i); break;
        case ILQRB_SD_DISTANCE: out[i] = sd_distance(ai, bi); break;
        case ILQRB_SD_TRANSPORT: sd_transport(ai, bi, ci, out + 4 * (size_t)i); break;
        default: {  // ILQRB_SD_DQUAT_TO_W_JAC, sd.h:23-27: the 3 x 4 matrix H(q)
            double* o = out + 12 * (size_t)i;
            const double q0 = ai[0], q1 = ai[1], q2 = ai[2], q3 = ai[3];
            o[0] = -q1, o[1] = q0, o[2] = -q3, o[3] = q2;
            o[4] = -q2, o[5] = q3, o[6] = q0, o[7] = -q1;
            o[8] = -q3, o[9] = -q2, o[10] = q1, o[11] = q0;
        }
    }
}

}  // namespace
}  // namespace ilqrb

using namespace ilqrb;

extern "C" int ilqrb_eval_point(ilqrb_handle h, int nq, const int32_t* problem, const int32_t* k, const double* x, const double* u, const double* fx_in, double* fx,
                                double* J, double* diff, double* cost, double* cost_x, double* cost_xx) {
    if (!h) return fail(ILQRB_EINVAL, "null handle");
    if (nq < 1 || !x) return fail(ILQRB_EINVAL, "no query points");
    const DevProblem& P = h->hp;
    for (int i = 0; i < nq; i++) {
        if (problem && (problem[i] < 0 || problem[i] >= P.B)) return fail(ILQRB_EINVAL, "problem index outside the batch");
        if (k && (k[i] < 0 || k[i] >= P.H)) return fail(ILQRB_EINVAL, "timestep outside the horizon");
    }
    CU(cudaSetDevice(h->device));
    const size_t n = P.n, m = P.m, N = (size_t)nq;
    // doubles: x, u, fx_in | fx, J, diff, c, cx, cxx ; then two int arrays
    const size_t o_x = 0, o_u = o_x + N * n, o_fi = o_u + N * m, o_fx = o_fi + N * P.nt, o_J = o_fx + N * P.nt, o_d = o_J + N * P.r * n, o_c = o_d + N * P.r,
                 o_cx = o_c + N, o_cxx = o_cx + N * n, o_i = o_cxx + N * n * n, total = o_i + N + 1;  // ints packed two per double
    DevBuf buf;
    CU(cudaMalloc(&buf.p, total * sizeof(double)));
    double* d = buf.p;
    int* di = reinterpret_cast<int*>(d + o_i);
    cudaStream_t st = h->stream;
    CU(cudaMemcpyAsync(d + o_x, x, N * n * sizeof(double), cudaMemcpyHostToDevice, st));
    if (u) CU(cudaMemcpyAsync(d + o_u, u, N * m * sizeof(double), cudaMemcpyHostToDevice, st));
    if (fx_in) CU(cudaMemcpyAsync(d + o_fi, fx_in, N * P.nt * sizeof(double), cudaMemcpyHostToDevice, st));
    if (problem) CU(cudaMemcpyAsync(di, problem, N * sizeof(int), cudaMemcpyHostToDevice, st));
    if (k) CU(cudaMemcpyAsync(di + N, k, N * sizeof(int), cudaMemcpyHostToDevice, st));
    k_eval_point<<<(nq + 63) / 64, 64, 0, st>>>(h->dp, nq, problem ? di : nullptr, k ? di + N : nullptr, d + o_x, u ? d + o_u : nullptr, fx_in ? d + o_fi : nullptr,
                                                d + o_fx, d + o_J, d + o_d, d + o_c, d + o_cx, d + o_cxx);
    CU(cudaGetLastError());
    h->launches++;
    if (fx) CU(cudaMemcpyAsync(fx, d + o_fx, N * P.nt * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (J) CU(cudaMemcpyAsync(J, d + o_J, N * P.r * n * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (diff) CU(cudaMemcpyAsync(diff, d + o_d, N * P.r * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (cost) CU(cudaMemcpyAsync(cost, d + o_c, N * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (cost_x) CU(cudaMemcpyAsync(cost_x, d + o_cx, N * n * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (cost_xx) CU(cudaMemcpyAsync(cost_xx, d + o_cxx, N * n * n * sizeof(double), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return 0;
}

extern "C" int ilqrb_keypoint_diff(int device, int sys_kind, int task_kind, int nb_deriv, int dof, const double* target, const double* deadzone, int nq, const double* fx,
                                   double* diff) {
    if (!target || !fx || !diff || nq < 1) return fail(ILQRB_EINVAL, "ilqrb_keypoint_diff: null argument");
    if (nb_deriv != 1 && nb_deriv != 2) return fail(ILQRB_EINVAL, "nb_deriv must be 1 or 2");
    if (sys_kind != ILQRB_SYS_POS_ORN && sys_kind != ILQRB_SYS_POS_ORN_TIME && sys_kind != ILQRB_SYS_POSITION) return fail(ILQRB_EINVAL, "unknown sys_kind");
    if (task_kind != ILQRB_TASK_POSE && task_kind != ILQRB_TASK_JOINT) return fail(ILQRB_EINVAL, "unknown task_kind");
    const int time = sys_kind == ILQRB_SYS_POS_ORN_TIME ? 1 : 0;
    int nt, r;
    if (task_kind == ILQRB_TASK_JOINT) {
        if (dof < 1 || dof > 7) return fail(ILQRB_EINVAL, "joint-space keypoint records hold 1..7 joints");
        nt = r = nb_deriv * dof + time;
    } else if (sys_kind == ILQRB_SYS_POSITION) {
        if (dof != 2 && dof != 3) return fail(ILQRB_EINVAL, "POSITION keypoints: dof = number of cartesian axes (2 or 3)");
        nt = r = dof * nb_deriv;
    } else {
        nt = 7 * nb_deriv + time;
        r = 6 * nb_deriv + time;
    }
    if (deadzone && (task_kind != ILQRB_TASK_POSE || sys_kind != ILQRB_SYS_POS_ORN)) return fail(ILQRB_EINVAL, "dead zones apply to PosOrnKeypoint only");
    CU(cudaSetDevice(device));
    std::vector<DevProblem> hp(1);
    DevProblem& P = hp[0];
    memset(&P, 0, sizeof(P));
    P.sys_kind = sys_kind;
    P.nb_deriv = nb_deriv;
    P.time = time;
    P.n_tasks = 1;
    P.n_robots = 1;
    P.robots[0].dof = dof;
    P.tasks[0].kind = task_kind;
    P.tasks[0].nt_t = nt;
    P.tasks[0].r_t = r;
    const size_t pd = (sizeof(DevProblem) + 7) / 8, N = (size_t)nq;
    const size_t o_t = pd, o_z = o_t + ILQRB_KP_STRIDE, o_f = o_z + 4, o_o = o_f + N * nt, total = o_o + N * r;
    DevBuf buf;
    CU(cudaMalloc(&buf.p, total * sizeof(double)));
    double* d = buf.p;
    CU(cudaMemcpy(d, &P, sizeof(DevProblem), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d + o_t, target, ILQRB_KP_STRIDE * sizeof(double), cudaMemcpyHostToDevice));
    if (deadzone) CU(cudaMemcpy(d + o_z, deadzone, 4 * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d + o_f, fx, N * nt * sizeof(double), cudaMemcpyHostToDevice));
    k_keypoint_diff<<<(nq + 63) / 64, 64>>>(reinterpret_cast<const DevProblem*>(d), d + o_t, deadzone ? d + o_z : nullptr, nq, nt, r, d + o_f, d + o_o);
    CU(cudaGetLastError());
    CU(cudaMemcpy(diff, d + o_o, N * r * sizeof(double), cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int ilqrb_sd(int device, int op, int nq, const double* a, const double* b, const double* c, double* out) {
    if (op < ILQRB_SD_LOGMAP || op > ILQRB_SD_DQUAT_TO_W_JAC) return fail(ILQRB_EINVAL, "unknown Sd operation");
    const bool need_b = op != ILQRB_SD_DQUAT_TO_W_JAC, need_c = op == ILQRB_SD_TRANSPORT;
    if (nq < 1 || !a || !out || (need_b && !b) || (need_c && !c)) return fail(ILQRB_EINVAL, "ilqrb_sd: missing operand");
    CU(cudaSetDevice(device));
    const size_t N = (size_t)nq, on = op == ILQRB_SD_DISTANCE ? N : op == ILQRB_SD_DQUAT_TO_W_JAC ? 12 * N : 4 * N;
    DevBuf buf;
    CU(cudaMalloc(&buf.p, (12 * N + on) * sizeof(double)));
    double* d = buf.p;
    CU(cudaMemcpy(d, a, 4 * N * sizeof(double), cudaMemcpyHostToDevice));
    if (need_b) CU(cudaMemcpy(d + 4 * N, b, 4 * N * sizeof(double), cudaMemcpyHostToDevice));
    if (need_c) CU(cudaMemcpy(d + 8 * N, c, 4 * N * sizeof(double), cudaMemcpyHostToDevice));
    k_sd<<<(nq + 63) / 64, 64>>>(op, nq, d, need_b ? d + 4 * N : nullptr, need_c ? d + 8 * N : nullptr, d + 12 * N);
    CU(cudaGetLastError());
    CU(cudaMemcpy(out, d + 12 * N, on * sizeof(double), cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int ilqrb_host_alloc(size_t bytes, void** out) {
    if (!out || bytes == 0) return fail(ILQRB_EINVAL, "ilqrb_host_alloc: null output or zero size");
    *out = nullptr;
    CU(cudaHostAlloc(out, bytes, cudaHostAllocPortable));
    return 0;
}
extern "C" int ilqrb_host_free(void* p) {
    if (p) CU(cudaFreeHost(p));
    return 0;
}
