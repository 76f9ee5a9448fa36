// ilqrb_api.cu — C ABI of libilqrb.so (include/ilqrb.h): handle management, descriptor flattening, kernel
// launch sequences of the four solvers, result download.  Host logic only orchestrates; all arithmetic of the
// solve path runs in the CUDA kernels behind the launchers of handle.hpp.  There is no CPU fallback.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "handle.hpp"

using namespace ilqrb;

namespace {
thread_local std::string g_err;
const char* kProfNames[ILQRB_PROF_SLOTS] = {"rollout_init", "kp_derivs",  "riccati_backward", "linesearch_rollout", "accept",     "final_fx",
                                            "batch_fp",     "batch_assemble", "batch_syrk",    "batch_cholesky",     "batch_linesearch", "misc",
                                            "deviation_rollout"};
}  // namespace

namespace ilqrb {
int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}
}  // namespace ilqrb

int ilqrb_handle_s::ensure_groups(int G) {
    if (!ev_start && cudaEventCreateWithFlags(&ev_start, cudaEventDisableTiming) != cudaSuccess) return fail(ILQRB_ECUDA, "cudaEventCreate");
    int least = 0, greatest = 0;
    cudaDeviceGetStreamPriorityRange(&least, &greatest);  // numerically lower = higher priority
    for (; n_groups < G; n_groups++) {
        // earlier groups get higher priority: they finish first and their results travel while later groups compute
        const int prio = std::min(least, greatest + n_groups);
        if (cudaStreamCreateWithPriority(&gstream[n_groups], cudaStreamNonBlocking, prio) != cudaSuccess ||
            cudaEventCreateWithFlags(&ev_stagger[n_groups], cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&ev_done[n_groups], cudaEventDisableTiming) != cudaSuccess)
            return fail(ILQRB_ECUDA, "could not create the group streams");
    }
    return 0;
}

namespace {

int derive_dims(const ilqrb_problem_desc* d, ilqrb_dims* o, int* task_nt, int* task_r) {
    if (!d) return fail(ILQRB_EINVAL, "null descriptor");
    if (d->abi_version != ILQRB_ABI_VERSION) return fail(ILQRB_EINVAL, "descriptor abi_version mismatch");
    if (d->nb_deriv != 1 && d->nb_deriv != 2) return fail(ILQRB_EINVAL, "nb_deriv must be 1 or 2");
    if (d->sys_kind < 0 || d->sys_kind > 2) return fail(ILQRB_EUNSUPPORTED, "unsupported System kind");
    if (d->n_tasks < 1 || d->n_tasks > ILQRB_MAX_TASKS) return fail(ILQRB_EINVAL, "n_tasks out of range");
    if (d->n_robots < 1 || d->n_robots > ILQRB_MAX_TASKS) return fail(ILQRB_EINVAL, "n_robots out of range");
    if (d->horizon < 2) return fail(ILQRB_EINVAL, "horizon must be >= 2");
    if (d->batch < 1) return fail(ILQRB_EINVAL, "batch must be >= 1");
    const int time = d->sys_kind == ILQRB_SYS_POS_ORN_TIME ? 1 : 0;
    o->n = d->nb_deriv * d->dof + time;
    o->m = d->dof + time;
    if (o->n > MAXN - 1 || o->m > MAXM) return fail(ILQRB_EUNSUPPORTED, "state/control size exceeds the compiled bounds (n <= 31, m <= 16)");
    o->nt = 0;
    o->r = 0;
    o->n_kp_total = 0;
    for (int t = 0; t < d->n_tasks; t++) {
        const ilqrb_task& tk = d->tasks[t];
        if (tk.robot < 0 || tk.robot >= d->n_robots) return fail(ILQRB_EINVAL, "task.robot out of range");
        if (tk.n_keypoints < 1 || tk.n_keypoints > ILQRB_MAX_KEYPOINTS) return fail(ILQRB_EINVAL, "n_keypoints out of range (1..8)");
        const ilqrb_robot& rb = d->robots[tk.robot];
        if (rb.dof < 1 || rb.dof > ILQRB_MAX_ROBOT_DOF) return fail(ILQRB_EUNSUPPORTED, "robot dof must be 1..7 (KDLRobot.cpp:112 limit)");
        if (tk.joint_offset < 0 || tk.joint_offset + rb.dof > d->dof) return fail(ILQRB_EINVAL, "task joint range outside the state");
        int nt, r;
        if (tk.task_kind != ILQRB_TASK_POSE && tk.task_kind != ILQRB_TASK_JOINT) return fail(ILQRB_EINVAL, "unknown task_kind");
        if (tk.kp_deadzone && (tk.task_kind != ILQRB_TASK_POSE || d->sys_kind != ILQRB_SYS_POS_ORN))
            return fail(ILQRB_EINVAL, "kp_deadzone (PosOrnKeypointDistFunct) applies to the end-effector tasks of PosOrnPlannerSys only (the reference has no time variant)");
        if (tk.task_kind == ILQRB_TASK_JOINT) {
            if (d->sys_kind == ILQRB_SYS_POSITION) return fail(ILQRB_EUNSUPPORTED, "joint-space tasks need a PosOrn or PosOrnTime dynamics class (sys_kind)");
            if (d->nb_deriv != 1)
                return fail(ILQRB_EUNSUPPORTED, "JointSpacePlannerSys with nb_deriv = 2 is not built: the reference's Jacobian Identity(2 dof, dof) "
                                                "(JointSpacePlannerSys.cpp:79) has the wrong shape for its own cost_x (undefined behaviour there)");
            if (tk.has_transform) return fail(ILQRB_EINVAL, "joint-space tasks do not take a transform");
            if (tk.joint_offset != 0 || rb.dof != d->dof) return fail(ILQRB_EUNSUPPORTED, "a joint-space task must span the whole joint vector");
            if (rb.dof > 7) return fail(ILQRB_EUNSUPPORTED, "joint-space keypoint records hold 7 joints");
            nt = r = d->nb_deriv * rb.dof + time;
        } else if (d->sys_kind == ILQRB_SYS_POSITION) {
            if (tk.has_transform) return fail(ILQRB_EUNSUPPORTED, "POSITION tasks do not take a transform");
            const int car = rb.kind == ILQRB_ROBOT_2D ? 2 : 3;
            nt = car * d->nb_deriv;
            r = nt;
        } else {
            if (rb.kind == ILQRB_ROBOT_2D) return fail(ILQRB_EUNSUPPORTED, "Robot2D only supports the POSITION system (SURVEY 8g C1)");
            nt = 7 * d->nb_deriv + time;
            r = 6 * d->nb_deriv + time;
        }
        if (task_nt) task_nt[t] = nt;
        if (task_r) task_r[t] = r;
        o->nt += nt;
        o->r += r;
        o->n_kp_total += tk.n_keypoints;
        for (int i = 0; i < tk.n_keypoints; i++) {
            if (tk.kp_timestep[i] < 0 || tk.kp_timestep[i] >= d->horizon) return fail(ILQRB_EINVAL, "keypoint timestep outside the horizon");
            if (i > 0 && tk.kp_timestep[i] <= tk.kp_timestep[i - 1]) return fail(ILQRB_EINVAL, "keypoint timesteps must be strictly ascending");
        }
    }
    o->H = d->horizon;
    o->B = d->batch;
    return 0;
}

template <class T>
int upload(ilqrb_handle h, T** dst, const T* src, size_t count) {
    int rc = h->alloc(dst, count);
    if (rc) return rc;
    CU(cudaMemcpyAsync(*dst, src, count * sizeof(T), cudaMemcpyHostToDevice, h->stream));
    return 0;
}

unsigned env_uint(const char* name, unsigned dflt) {
    const char* v = getenv(name);
    return v ? (unsigned)strtoul(v, nullptr, 10) : dflt;
}

int ensure_iter_capacity(ilqrb_handle h, int nb_iter) {
    if (nb_iter <= h->cap_iter) return 0;
    int rc;
    if ((rc = h->alloc(&h->W.cost_hist, (size_t)h->hp.B * nb_iter))) return rc;
    if ((rc = h->alloc(&h->W.alpha_hist, (size_t)h->hp.B * nb_iter))) return rc;
    h->cap_iter = nb_iter;
    return 0;
}

__global__ void k_fill(double* p, size_t n, double v) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}
// strided pack: dst[row][0..w) = src[row*stride + 0..w) * (scale ? scale[row / rows_per_problem] : 1)
__global__ void k_pack(double* __restrict__ dst, const double* __restrict__ src, size_t rows, int w, int stride, const double* __restrict__ scale, int rows_per_problem) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * w) return;
    size_t row = i / w;
    int c = (int)(i % w);
    double v = src[row * stride + c];
    if (scale) v *= scale[row / rows_per_problem];
    dst[i] = v;
}
// dst[i] = src[(i mod per_problem) mod src_len]: one small array replicated over the batch (and over the steps of a problem)
__global__ void k_bcast_rows(double* __restrict__ dst, const double* __restrict__ src, size_t tot, size_t per_problem, size_t src_len) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < tot) dst[i] = src[(i % per_problem) % src_len];
}
__global__ void k_unpack(double* __restrict__ dst, const double* __restrict__ src, size_t rows, int w, int stride) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * stride) return;
    size_t row = i / stride;
    int c = (int)(i % stride);
    dst[i] = (c < w && src) ? src[row * w + c] : 0.0;
}

int ensure_pack(ilqrb_handle h, size_t elems) {
    if (elems <= h->pack_elems) return 0;
    int rc = h->alloc(&h->d_pack, elems);
    if (rc) return rc;
    h->pack_elems = elems;
    return 0;
}

int download_packed(ilqrb_handle h, double* dst_host, const double* src_dev, size_t rows, int w, int stride, const double* scale, int rpp) {
    if (!dst_host) return 0;
    int rc = ensure_pack(h, rows * w);
    if (rc) return rc;
    const size_t tot = rows * w;
    k_pack<<<(unsigned)((tot + 255) / 256), 256, 0, h->stream>>>(h->d_pack, src_dev, rows, w, stride, scale, rpp);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(dst_host, h->d_pack, tot * sizeof(double), cudaMemcpyDefault, h->stream));
    CU(cudaStreamSynchronize(h->stream));  // staging buffer is reused
    return 0;
}

// Results of problems [b0, b0+nB) packed and sent to the caller's buffers, all asynchronous on `st`.  Every output has
// its own slice of one staging buffer (problem-major), so ranges never wait for one another.
int enqueue_download_range(ilqrb_handle h, const ilqrb_result* out, int b0, int nB, int nit, cudaStream_t st, bool gains = true) {
    const DevProblem& P = h->hp;
    const size_t B = P.B;
    const int H = P.H, n = P.n, m = P.m;
    const size_t eX = (size_t)H * n, eU = (size_t)(H - 1) * m, eK = (size_t)(H - 1) * m * n;
    int rc = ensure_pack(h, B * (eX + 2 * eU + eK));
    if (rc) return rc;
    double* sX = h->d_pack;
    double* sU = sX + B * eX;
    double* sK = sU + B * eU;
    double* sD = sK + B * eK;
    auto pack = [&](double* host, double* stage, const double* src, size_t per_problem_rows, int w, int stride, const double* scale, int rpp) -> int {
        if (!host) return 0;
        const size_t rows = (size_t)nB * per_problem_rows, tot = rows * w;
        double* dst = stage + (size_t)b0 * per_problem_rows * w;
        k_pack<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(dst, src + (size_t)b0 * per_problem_rows * stride, rows, w, stride, scale ? scale + b0 : nullptr, rpp);
        CU(cudaGetLastError());
        CU(cudaMemcpyAsync(host + (size_t)b0 * per_problem_rows * w, dst, tot * sizeof(double), cudaMemcpyDefault, st));
        return 0;
    };
    if ((rc = pack(out->X, sX, h->W.X, H, n, h->W.ns, nullptr, 1))) return rc;
    if ((rc = pack(out->U, sU, h->W.U, H - 1, m, h->W.ms, nullptr, 1))) return rc;
    if (gains && (rc = pack(out->K, sK, h->W.K, (size_t)(H - 1) * m, n, h->W.ns, nullptr, 1))) return rc;
    // ds are returned multiplied by the accepted alpha (ILQRRecursive.cpp:128,144,162)
    if (gains && (rc = pack(out->d, sD, h->W.d, H - 1, m, h->W.ms, h->last_nb_iter > 0 ? h->W.alpha_last : nullptr, H - 1))) return rc;
    if (out->fX) {
        if ((rc = launch_final_fx(h, b0, nB, st))) return rc;
        const size_t off = (size_t)b0 * H * P.nt;
        CU(cudaMemcpyAsync(out->fX + off, h->d_fX + off, (size_t)nB * H * P.nt * sizeof(double), cudaMemcpyDefault, st));
    }
    if (out->cost) CU(cudaMemcpyAsync(out->cost + b0, h->W.cost0 + b0, nB * sizeof(double), cudaMemcpyDefault, st));
    if (out->cost_hist) CU(cudaMemcpyAsync(out->cost_hist + (size_t)b0 * nit, h->W.cost_hist + (size_t)b0 * nit, (size_t)nB * nit * sizeof(double), cudaMemcpyDefault, st));
    if (out->alpha_hist) CU(cudaMemcpyAsync(out->alpha_hist + (size_t)b0 * nit, h->W.alpha_hist + (size_t)b0 * nit, (size_t)nB * nit * sizeof(double), cudaMemcpyDefault, st));
    if (out->iters) CU(cudaMemcpyAsync(out->iters + b0, h->W.iters + b0, nB * sizeof(int), cudaMemcpyDefault, st));
    if (out->status) CU(cudaMemcpyAsync(out->status + b0, h->W.status + b0, nB * sizeof(int), cudaMemcpyDefault, st));
    return 0;
}

// out_async != nullptr: the caller wants the results in host buffers; the fast path then sends each problem group's
// results as soon as that group has finished (sets h->downloaded).
// marks "the solve's kernels are queued on st" for ilqrb_chain_after_compute (slot i of the handle's compute events)
static int record_compute_done(ilqrb_handle h, int i, cudaStream_t st) {
    if (!h->ev_compute[i]) CU(cudaEventCreateWithFlags(&h->ev_compute[i], cudaEventDisableTiming));
    CU(cudaEventRecord(h->ev_compute[i], st));
    if (h->n_ev_compute < i + 1) h->n_ev_compute = i + 1;
    return 0;
}

int run_recursive_like(ilqrb_handle h, const ilqrb_al_desc* al, const ilqrb_solve_opts* opts, const ilqrb_result* out_async = nullptr) {
    if (!opts) return fail(ILQRB_EINVAL, "null opts");
    if (opts->nb_iter < 0) return fail(ILQRB_EINVAL, "nb_iter < 0");
    CU(cudaSetDevice(h->device));
    int rc;
    const int B = h->hp.B, H = h->hp.H, n = h->hp.n, m = h->hp.m;
    if ((rc = ensure_iter_capacity(h, std::max(opts->nb_iter, 1)))) return rc;
    const size_t nh = (size_t)B * std::max(opts->nb_iter, 1);
    k_fill<<<(unsigned)((nh + 255) / 256), 256, 0, h->stream>>>(h->W.cost_hist, nh, NAN);
    k_fill<<<(unsigned)((nh + 255) / 256), 256, 0, h->stream>>>(h->W.alpha_hist, nh, NAN);
    CU(cudaMemsetAsync(h->W.K, 0, (size_t)B * (H - 1) * m * h->W.ns * sizeof(double), h->stream));
    CU(cudaMemsetAsync(h->W.d, 0, (size_t)B * (H - 1) * h->W.ms * sizeof(double), h->stream));
    double pen = 0, pen_frozen = 0;
    if (al) {
        if (al->nc < 1 || al->nc > MAX_NC) return fail(ILQRB_EUNSUPPORTED, "AL: constraint rows per step must be 1..32");
        if (al->lag_update_step < 1) return fail(ILQRB_EINVAL, "AL: lag_update_step must be >= 1");
        const int nc = al->nc, steps = al->per_step ? H - 1 : 1;
        const size_t ea = (size_t)steps * nc * (n + m), eb = (size_t)steps * nc;
        if (ea > h->alA_elems) {
            if ((rc = h->alloc(&h->d_alA, ea))) return rc;
            h->alA_elems = ea;
        }
        if (eb > h->alb_elems) {
            if ((rc = h->alloc(&h->d_alb, eb))) return rc;
            h->alb_elems = eb;
        }
        CU(cudaMemcpyAsync(h->d_alA, al->A, ea * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        CU(cudaMemcpyAsync(h->d_alb, al->b, eb * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        if (h->W.nc != nc || !h->W.lam) {
            if ((rc = h->alloc(&h->W.lam, (size_t)B * (H - 1) * nc))) return rc;
            if ((rc = h->alloc(&h->W.Cs, (size_t)B * (H - 1) * nc))) return rc;
            if (!h->W.mask && (rc = h->alloc(&h->W.mask, (size_t)B * (H - 1)))) return rc;
        }
        h->W.nc = nc;
        h->W.al_per_step = al->per_step;
        h->W.alA = h->d_alA;
        h->W.alb = h->d_alb;
        // initial multipliers: one copy to the device, replicated over the batch (and over steps when given once) by a kernel
        if (eb > h->lam0_elems) {
            if ((rc = h->alloc(&h->d_lam0, eb))) return rc;
            h->lam0_elems = eb;
        }
        CU(cudaMemcpyAsync(h->d_lam0, al->lambda0, eb * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        {
            const size_t tot = (size_t)B * (H - 1) * nc;
            k_bcast_rows<<<(unsigned)((tot + 255) / 256), 256, 0, h->stream>>>(h->W.lam, h->d_lam0, tot, (size_t)(H - 1) * nc, al->per_step ? (size_t)(H - 1) * nc : (size_t)nc);
            CU(cudaGetLastError());
        }
        pen = al->penalty;
        pen_frozen = pen;
    }
    if ((rc = launch_init(h))) return rc;
    if (al && (rc = launch_al_init(h))) return rc;
    const bool fast1 = fast_n1_applies(h, al != nullptr), fastt = !fast1 && fast_time_applies(h, al != nullptr);
    if (fast1 || fastt) {
        // one iteration of problems [b0, b0 + nB) on stream st.  Single-integrator path: backward, forward (phase F of the forward pass
        // leaves l_x / l_xx for the next backward pass; the first ones come from k_kp_derivs below).  Time path: k_kp_derivs, backward, forward.
        auto iteration = [&](int b0, int nB, cudaStream_t st, int it, cudaEvent_t after_backward) -> int {
            int r;
            if (fastt && (r = launch_derivs_time(h, b0, nB, st))) return r;
            if ((r = fast1 ? launch_backward_n1(h, b0, nB, st) : launch_backward_time(h, b0, nB, st))) return r;
            if (after_backward) CU(cudaEventRecord(after_backward, st));
            return fast1 ? launch_forward_n1(h, b0, nB, st, it, opts->nb_iter, opts->line_search, opts->early_stop)
                         : launch_forward_time(h, b0, nB, st, it, opts->nb_iter, opts->line_search, opts->early_stop);
        };
        if (fast1 && opts->nb_iter > 0 && (rc = launch_derivs(h))) return rc;
        // The batch is split into G groups of problems, each on its own stream, the first backward pass of group g+1
        // held back until group g has finished its own: from then on the FP64-bound backward pass of one group shares
        // the SMs with the latency-bound forward pass of another.  (Profiling mode: one group, per-kernel event pairs.)
        // measured on B200 (C2, B = 4096; ILQRB_GROUPS sweeps in DESIGN.md): 3 groups for device-resident results, 8 when one blocking call
        // must hide its own downloads, 5 when the caller streams batches through two handles (the other handle's solve covers the
        // downloads, and fewer, larger groups keep the kernels closer to a full wave)
        // device-resident results: fast_n1_default_groups (launch_fused1.cu); time path: about three groups per two waves of 16 warps per SM
        const int wave = 148 * 16;
        const int g_dev = fast1 ? fast_n1_default_groups(h, B) : std::min(8, std::max(3, (3 * B + 2 * wave - 1) / (2 * wave)));
        // host outputs on the column-per-lane backward path (C2; profiles/r2o_e2e_groups.txt, r2p_e2e_chain.txt): a blocking call 352 / 371 /
        // 360 / 349 k solves/s with 1 / 2 / 3 / 5 groups; two streaming handles half a cycle apart 593 / 611 / 602 / 576 k with gains,
        // 660 / 666 / 632 / 636 k without
        const bool col = fast1 && fast_n1_col_path(h);
        const int g_host = col ? 2 : (h->streaming ? 5 : 8);
        int G = h->prof ? 1 : (int)env_uint("ILQRB_GROUPS", out_async ? g_host : g_dev);
        if (G > ilqrb_handle_s::MAX_GROUPS) G = ilqrb_handle_s::MAX_GROUPS;
        if (G < 1 || B < 64 * G) G = 1;
        if (G == 1) {
            for (int it = 0; it < opts->nb_iter; it++)
                if ((rc = iteration(0, B, h->stream, it, nullptr))) return rc;
            h->n_ev_compute = 0;
            if ((rc = record_compute_done(h, 0, h->stream))) return rc;
            if (out_async && opts->nb_iter > 0) {  // the copies are queued behind the solve: the streaming entry point returns without waiting
                h->last_nb_iter = opts->nb_iter;
                if ((rc = enqueue_download_range(h, out_async, 0, B, std::max(opts->nb_iter, 1), h->stream))) return rc;
                h->downloaded = true;
            }
        } else if (opts->nb_iter > 0) {
            if ((rc = h->ensure_groups(G))) return rc;
            CU(cudaEventRecord(h->ev_start, h->stream));
            for (int gi = 0; gi < G; gi++) CU(cudaStreamWaitEvent(h->gstream[gi], h->ev_start, 0));
            // stagger: group g+1 starts once group g has finished `stag` iterations (0: its first backward pass only).
            // Measured on B200 (C2, B=4096): a longer stagger buys PCIe/compute overlap for host outputs but loses the
            // same time to SMs running below the ~4096 resident problems they need, so the default stays 0.
            int stag = (int)env_uint("ILQRB_STAGGER_ITERS", 0);
            if (stag >= opts->nb_iter) stag = opts->nb_iter - 1;
            for (int gi = 0; gi < G; gi++) {
                const int b0 = (int)((long long)B * gi / G), b1 = (int)((long long)B * (gi + 1) / G);
                cudaStream_t st = h->gstream[gi];
                if (gi > 0) CU(cudaStreamWaitEvent(st, h->ev_stagger[gi - 1], 0));
                for (int it = 0; it < opts->nb_iter; it++) {
                    if ((rc = iteration(b0, b1 - b0, st, it, (stag == 0 && it == 0 && gi + 1 < G) ? h->ev_stagger[gi] : nullptr))) return rc;
                    if (stag > 0 && it == stag - 1 && gi + 1 < G) CU(cudaEventRecord(h->ev_stagger[gi], st));
                }
            }
            h->last_nb_iter = opts->nb_iter;
            h->n_ev_compute = 0;
            for (int gi = 0; gi < G; gi++) {
                if ((rc = record_compute_done(h, gi, h->gstream[gi]))) return rc;
                if (out_async) {
                    const int b0 = (int)((long long)B * gi / G), b1 = (int)((long long)B * (gi + 1) / G);
                    if ((rc = enqueue_download_range(h, out_async, b0, b1 - b0, std::max(opts->nb_iter, 1), h->gstream[gi]))) return rc;
                    h->downloaded = true;
                }
                CU(cudaEventRecord(h->ev_done[gi], h->gstream[gi]));
                CU(cudaStreamWaitEvent(h->stream, h->ev_done[gi], 0));
            }
        }
        h->last_nb_iter = opts->nb_iter;
        h->last_solver = 1;
        return 0;
    }
    for (int it = 0; it < opts->nb_iter; it++) {
        if ((rc = launch_derivs(h))) return rc;
        if ((rc = fast_n2_applies(h, al != nullptr) ? launch_backward_n2(h, al != nullptr, pen_frozen) : launch_backward_generic(h, al != nullptr, pen_frozen))) return rc;
        if ((rc = launch_linesearch(h, opts->line_search, al != nullptr))) return rc;
        double pen_next = pen;
        bool upd = false;
        if (al) {
            upd = ((it + 1) % al->lag_update_step) == 0;
            if (upd) pen_next = pen * al->scaling_factor;
        }
        if ((rc = launch_accept(h, it, std::max(opts->nb_iter, 1), opts->line_search, opts->early_stop, al != nullptr, al ? al->lag_update_step : 1, pen_next))) return rc;
        pen_frozen = pen;  // Is[k] keeps the penalty of the rollout that produced it (AL-ILQR.cpp:190)
        pen = pen_next;
    }
    h->last_nb_iter = opts->nb_iter;
    h->last_solver = al ? 2 : 1;
    return 0;
}

}  // namespace

extern "C" {

int ilqrb_abi_version(void) { return ILQRB_ABI_VERSION; }
void ilqrb_set_error_(const char* msg) { g_err = msg ? msg : ""; }  // used by host_utils.cpp
const char* ilqrb_last_error(void) { return g_err.c_str(); }

int ilqrb_get_dims(const ilqrb_problem_desc* desc, ilqrb_dims* out) {
    if (!out) return fail(ILQRB_EINVAL, "null out");
    return derive_dims(desc, out, nullptr, nullptr);
}

int ilqrb_create(const ilqrb_problem_desc* d, int device, ilqrb_handle* out) {
    if (!out) return fail(ILQRB_EINVAL, "null out");
    ilqrb_dims dm;
    int task_nt[ILQRB_MAX_TASKS], task_r[ILQRB_MAX_TASKS];
    int rc = derive_dims(d, &dm, task_nt, task_r);
    if (rc) return rc;
    int ndev = 0;
    cudaError_t ce = cudaGetDeviceCount(&ndev);
    if (ce != cudaSuccess || ndev == 0)
        return fail(ILQRB_ECUDA, std::string("no CUDA device available (there is no CPU fallback): ") + (ce != cudaSuccess ? cudaGetErrorString(ce) : "device count 0"));
    if (device < 0 || device >= ndev) return fail(ILQRB_EINVAL, "device index out of range");
    CU(cudaSetDevice(device));
    ilqrb_handle h = new ilqrb_handle_s();
    h->device = device;
    h->dims = dm;
    std::memset(&h->W, 0, sizeof(Work));
    cudaError_t se = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    if (se != cudaSuccess) {
        delete h;
        return fail(ILQRB_ECUDA, std::string("cudaStreamCreate: ") + cudaGetErrorString(se));
    }
    h->own_stream = true;
    DevProblem& P = h->hp;
    std::memset(&P, 0, sizeof(P));
    P.sys_kind = d->sys_kind;
    P.nb_deriv = d->nb_deriv;
    P.H = d->horizon;
    P.B = d->batch;
    P.dof = d->dof;
    P.n_robots = d->n_robots;
    P.n_tasks = d->n_tasks;
    P.time = d->sys_kind == ILQRB_SYS_POS_ORN_TIME ? 1 : 0;
    P.n = dm.n;
    P.m = dm.m;
    P.nt = dm.nt;
    P.r = dm.r;
    P.dt = d->dt;
    for (int i = 0; i < dm.m; i++) P.r_diag[i] = d->r_diag[i];
    for (int i = 0; i < d->n_robots; i++) {
        const ilqrb_robot& rb = d->robots[i];
        DevRobot& o = P.robots[i];
        o.kind = rb.kind;
        o.dof = rb.dof;
        o.n_segments = rb.n_segments;
        if (rb.kind == ILQRB_ROBOT_KDL) {
            if (rb.n_segments < 1 || rb.n_segments > ILQRB_MAX_SEGMENTS) {
                ilqrb_destroy(h);
                return fail(ILQRB_EINVAL, "n_segments out of range (1..24)");
            }
            int nj = 0;
            for (int s = 0; s < rb.n_segments; s++) {
                o.segs[s] = rb.segments[s];
                if (rb.segments[s].joint_type != ILQRB_JOINT_FIXED) nj++;
            }
            if (nj != rb.dof) {
                ilqrb_destroy(h);
                return fail(ILQRB_EINVAL, "robot.dof does not match the chain's movable joints");
            }
            // compacted chain: a fixed segment's pose is its tip frame, so it folds into the tip of the segment before it
            o.n_lean = 0;
            for (int s = 0; s < rb.n_segments; s++) {
                const ilqrb_segment& sg = rb.segments[s];
                if (sg.joint_type == ILQRB_JOINT_FIXED && o.n_lean > 0) {
                    ilqrb_segment& pv = o.lean[o.n_lean - 1];
                    double R[9], p[3];
                    for (int i = 0; i < 3; i++) {
                        for (int j = 0; j < 3; j++) R[3 * i + j] = pv.tip_R[3 * i] * sg.tip_R[j] + pv.tip_R[3 * i + 1] * sg.tip_R[3 + j] + pv.tip_R[3 * i + 2] * sg.tip_R[6 + j];
                        p[i] = pv.tip_R[3 * i] * sg.tip_p[0] + pv.tip_R[3 * i + 1] * sg.tip_p[1] + pv.tip_R[3 * i + 2] * sg.tip_p[2] + pv.tip_p[i];
                    }
                    std::memcpy(pv.tip_R, R, sizeof(R));
                    std::memcpy(pv.tip_p, p, sizeof(p));
                } else {
                    o.lean[o.n_lean++] = sg;
                }
            }
        } else if (rb.kind == ILQRB_ROBOT_2D) {
            for (int c = 0; c < rb.dof; c++) o.lengths[c] = rb.lengths[c];
        } else {
            ilqrb_destroy(h);
            return fail(ILQRB_EUNSUPPORTED, "unsupported SimulationInterface kind");
        }
    }
    const size_t B = d->batch;
    int fx_off = 0, r_off = 0;
    std::vector<int> steps;
    for (int t = 0; t < d->n_tasks; t++) {
        const ilqrb_task& tk = d->tasks[t];
        DevTask& o = P.tasks[t];
        o.robot = tk.robot;
        o.joint_offset = tk.joint_offset;
        o.has_transform = tk.has_transform;
        o.n_keypoints = tk.n_keypoints;
        std::memcpy(o.T, tk.T, sizeof(o.T));
        for (int i = 0; i < tk.n_keypoints; i++) {
            o.kp_timestep[i] = tk.kp_timestep[i];
            bool seen = false;
            for (int s : steps) seen = seen || s == tk.kp_timestep[i];
            if (!seen) steps.push_back(tk.kp_timestep[i]);
        }
        o.precision_per_problem = tk.precision_per_problem;
        o.kind = tk.task_kind;
        o.kp_deadzone = nullptr;
        if (tk.kp_deadzone) {
            double* ddz = nullptr;
            if ((rc = upload(h, &ddz, tk.kp_deadzone, (size_t)tk.n_keypoints * 4))) {
                ilqrb_destroy(h);
                return rc;
            }
            o.kp_deadzone = ddz;
        }
        o.limits_set = tk.limits_set;
        o.nt_t = task_nt[t];
        o.r_t = task_r[t];
        o.fx_off = fx_off;
        o.r_off = r_off;
        fx_off += task_nt[t];
        r_off += task_r[t];
        for (int i = 0; i < dm.m; i++) o.r_diag[i] = tk.r_diag[i];
        if (tk.limits_set) {
            P.any_limits = 1;
            for (int i = 0; i < dm.n; i++) {
                o.state_max[i] = tk.state_max[i];
                o.state_min[i] = tk.state_min[i];
                o.limit_weight[i] = tk.limit_weight[i];
            }
        }
        double* dtg = nullptr;
        double* dpr = nullptr;
        if ((rc = upload(h, &dtg, tk.kp_target, B * tk.n_keypoints * ILQRB_KP_STRIDE)) ||
            (rc = upload(h, &dpr, tk.kp_precision, (size_t)(tk.precision_per_problem ? B : 1) * tk.n_keypoints * task_r[t] * task_r[t]))) {
            ilqrb_destroy(h);
            return rc;
        }
        o.kp_target = dtg;
        o.kp_precision = dpr;
    }
    std::sort(steps.begin(), steps.end());
    P.n_kp_steps = (int)steps.size();
    for (size_t i = 0; i < steps.size(); i++) P.kp_steps[i] = steps[i];
    double *dq0 = nullptr, *ddq0 = nullptr;
    if ((rc = upload(h, &dq0, d->q0, B * d->dof)) || (rc = upload(h, &ddq0, d->dq0, B * d->dof))) {
        ilqrb_destroy(h);
        return rc;
    }
    P.q0 = dq0;
    P.dq0 = ddq0;
    if ((rc = upload(h, &h->dp, &P, 1))) {
        ilqrb_destroy(h);
        return rc;
    }
    // work buffers
    Work& W = h->W;
    const int H = d->horizon, n = dm.n, m = dm.m;
    W.ns = (n + 1) & ~1;
    W.ms = (m + 1) & ~1;
    const size_t S = P.n_kp_steps;
    if ((rc = h->alloc(&W.X, B * H * W.ns)) || (rc = h->alloc(&W.U, B * (H - 1) * W.ms)) || (rc = h->alloc(&W.K, B * (H - 1) * m * W.ns)) ||
        (rc = h->alloc(&W.d, B * (H - 1) * W.ms)) || (rc = h->alloc(&W.Uc, B * (H - 1) * NCAND * W.ms)) || (rc = h->alloc(&W.cost_c, B * NCAND)) ||
        (rc = h->alloc(&W.dun_c, B * NCAND)) || (rc = h->alloc(&W.cost0, B)) || (rc = h->alloc(&W.LX, B * S * n)) || (rc = h->alloc(&W.LXX, B * S * n * n)) ||
        (rc = h->alloc(&W.alpha_last, B)) || (rc = h->alloc(&W.done, B)) || (rc = h->alloc(&W.iters, B)) || (rc = h->alloc(&W.status, B)) ||
        (rc = h->alloc(&h->d_fX, B * H * dm.nt)) || (rc = h->alloc(&W.DU, n == m ? B * (H - 1) * W.ms : 1)) || (rc = h->alloc(&W.DXk, n == m ? B * S * W.ns : 1)) ||
        (rc = h->alloc(&W.viol, B))) {
        ilqrb_destroy(h);
        return rc;
    }
    cudaMemsetAsync(W.U, 0, B * (H - 1) * W.ms * sizeof(double), h->stream);
    cudaMemsetAsync(W.X, 0, B * H * W.ns * sizeof(double), h->stream);
    cudaMemsetAsync(W.done, 0, B * sizeof(int), h->stream);
    cudaMemsetAsync(W.iters, 0, B * sizeof(int), h->stream);
    cudaMemsetAsync(W.status, 0, B * sizeof(int), h->stream);
    cudaError_t e2 = cudaStreamSynchronize(h->stream);
    if (e2 != cudaSuccess) {
        ilqrb_destroy(h);
        return fail(ILQRB_ECUDA, std::string("create: ") + cudaGetErrorString(e2));
    }
    *out = h;
    return 0;
}

int ilqrb_destroy(ilqrb_handle h) {
    if (!h) return 0;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    for (void* p : h->allocs) cudaFree(p);
    batch_free(h->BW);
    for (auto e : h->ev_pool) cudaEventDestroy(e);
    for (int gi = 0; gi < h->n_groups; gi++) {
        cudaStreamSynchronize(h->gstream[gi]);
        cudaStreamDestroy(h->gstream[gi]);
        cudaEventDestroy(h->ev_stagger[gi]);
        cudaEventDestroy(h->ev_done[gi]);
    }
    if (h->ev_start) cudaEventDestroy(h->ev_start);
    for (auto e : h->ev_compute)
        if (e) cudaEventDestroy(e);
    if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return 0;
}

int ilqrb_set_stream(ilqrb_handle h, void* s) {
    if (!h) return fail(ILQRB_EINVAL, "null handle");
    cudaStreamSynchronize(h->stream);
    if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
    h->stream = (cudaStream_t)s;
    h->own_stream = false;
    return 0;
}

int ilqrb_chain_after_compute(ilqrb_handle next, ilqrb_handle prev) {
    if (!next || !prev) return fail(ILQRB_EINVAL, "null handle");
    CU(cudaSetDevice(next->device));
    if (prev->n_ev_compute == 0) {  // no solve on the fast paths yet: order behind everything prev has queued
        if (!prev->ev_compute[0]) CU(cudaEventCreateWithFlags(&prev->ev_compute[0], cudaEventDisableTiming));
        CU(cudaEventRecord(prev->ev_compute[0], prev->stream));
        CU(cudaStreamWaitEvent(next->stream, prev->ev_compute[0], 0));
        return 0;
    }
    for (int i = 0; i < prev->n_ev_compute; i++) CU(cudaStreamWaitEvent(next->stream, prev->ev_compute[i], 0));
    return 0;
}

int ilqrb_synchronize(ilqrb_handle h) {
    if (!h) return fail(ILQRB_EINVAL, "null handle");
    CU(cudaStreamSynchronize(h->stream));
    return 0;
}

int ilqrb_update_problems(ilqrb_handle h, const double* q0, const double* dq0, int task, const double* kp_target) {
    if (!h) return fail(ILQRB_EINVAL, "null handle");
    CU(cudaSetDevice(h->device));
    const size_t B = h->hp.B;
    if (q0) CU(cudaMemcpyAsync((void*)h->hp.q0, q0, B * h->hp.dof * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    if (dq0) CU(cudaMemcpyAsync((void*)h->hp.dq0, dq0, B * h->hp.dof * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    if (kp_target) {
        if (task < 0 || task >= h->hp.n_tasks) return fail(ILQRB_EINVAL, "task out of range");
        const DevTask& tk = h->hp.tasks[task];
        CU(cudaMemcpyAsync((void*)tk.kp_target, kp_target, B * tk.n_keypoints * ILQRB_KP_STRIDE * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    }
    CU(cudaStreamSynchronize(h->stream));
    return 0;
}

int ilqrb_upload_u0(ilqrb_handle h, const double* U0) {
    if (!h) return fail(ILQRB_EINVAL, "null handle");
    CU(cudaSetDevice(h->device));
    const size_t rows = (size_t)h->hp.B * (h->hp.H - 1);
    const int m = h->hp.m, ms = h->W.ms;
    if (!U0) {
        CU(cudaMemsetAsync(h->W.U, 0, rows * ms * sizeof(double), h->stream));
        return 0;
    }
    int rc = ensure_pack(h, rows * m);
    if (rc) return rc;
    CU(cudaMemcpyAsync(h->d_pack, U0, rows * m * sizeof(double), cudaMemcpyDefault, h->stream));  // host or device memory
    const size_t tot = rows * ms;
    k_unpack<<<(unsigned)((tot + 255) / 256), 256, 0, h->stream>>>(h->W.U, h->d_pack, rows, m, ms);
    CU(cudaGetLastError());
    return 0;
}

int ilqrb_run_recursive(ilqrb_handle h, const ilqrb_solve_opts* opts) {
    if (!h) return fail(ILQRB_EINVAL, "null handle");
    return run_recursive_like(h, nullptr, opts);
}
int ilqrb_run_al(ilqrb_handle h, const ilqrb_al_desc* al, const ilqrb_solve_opts* opts) {
    if (!h) return fail(ILQRB_EINVAL, "null handle");
    if (!al || !al->A || !al->b || !al->lambda0) return fail(ILQRB_EINVAL, "AL descriptor incomplete");
    return run_recursive_like(h, al, opts);
}

int ilqrb_download(ilqrb_handle h, ilqrb_result* out) {
    if (!h || !out) return fail(ILQRB_EINVAL, "null argument");
    CU(cudaSetDevice(h->device));
    const DevProblem& P = h->hp;
    const size_t B = P.B;
    const int H = P.H, n = P.n, m = P.m, nit = std::max(h->last_nb_iter, 1);
    int rc;
    if (h->last_solver == 3) return batch_download_h(h, out, nit);
    // every output has its own slice of the staging buffer: all packs and copies are queued, one synchronisation at the end
    if ((rc = enqueue_download_range(h, out, 0, (int)B, nit, h->stream, h->last_solver == 1))) return rc;
    if (out->lambda && h->last_solver == 2) CU(cudaMemcpyAsync(out->lambda, h->W.lam, B * (H - 1) * h->W.nc * sizeof(double), cudaMemcpyDefault, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    (void)n;
    (void)m;
    return 0;
}

int ilqrb_solve_recursive(ilqrb_handle h, const double* U0, const ilqrb_solve_opts* opts, ilqrb_result* out) {
    int rc;
    if (!h || !out) return fail(ILQRB_EINVAL, "null argument");
    if ((rc = ilqrb_upload_u0(h, U0))) return rc;
    h->downloaded = false;
    if ((rc = run_recursive_like(h, nullptr, opts, out))) return rc;
    if (h->downloaded) {  // every group has queued its own device-to-host copies; the solve stream waits for them
        CU(cudaStreamSynchronize(h->stream));
        return 0;
    }
    return ilqrb_download(h, out);
}
int ilqrb_solve_recursive_async(ilqrb_handle h, const double* U0, const ilqrb_solve_opts* opts, ilqrb_result* out) {
    int rc;
    if (!h || !out) return fail(ILQRB_EINVAL, "null argument");
    if ((rc = ilqrb_upload_u0(h, U0))) return rc;
    h->downloaded = false;
    h->streaming = true;
    rc = run_recursive_like(h, nullptr, opts, out);
    h->streaming = false;
    if (rc) return rc;
    if (h->downloaded) return 0;  // every group has queued its own device-to-host copies; ilqrb_synchronize waits for them
    return ilqrb_download(h, out);
}
int ilqrb_solve_al(ilqrb_handle h, const double* U0, const ilqrb_al_desc* al, const ilqrb_solve_opts* opts, ilqrb_result* out) {
    int rc;
    if ((rc = ilqrb_upload_u0(h, U0))) return rc;
    if ((rc = ilqrb_run_al(h, al, opts))) return rc;
    return ilqrb_download(h, out);
}

int ilqrb_run_batch(ilqrb_handle h, const ilqrb_batch_desc* bd, const ilqrb_solve_opts* opts) {
    if (!h || !opts) return fail(ILQRB_EINVAL, "null argument");
    CU(cudaSetDevice(h->device));
    int rc = ensure_iter_capacity(h, std::max(opts->nb_iter, 1));
    if (rc) return rc;
    if ((rc = batch_run_h(h, bd, opts))) return rc;
    h->last_nb_iter = opts->nb_iter;
    h->last_solver = 3;
    return 0;
}
int ilqrb_solve_batch(ilqrb_handle h, const double* u0, const ilqrb_batch_desc* bd, const ilqrb_solve_opts* opts, ilqrb_result* out) {
    int rc;
    if ((rc = ilqrb_upload_u0(h, u0))) return rc;
    if ((rc = ilqrb_run_batch(h, bd, opts))) return rc;
    return ilqrb_download(h, out);
}

int ilqrb_rollout(ilqrb_handle h, const double* U, double* X, double* fX, double* J, double* cost) {
    if (!h) return fail(ILQRB_EINVAL, "null handle");
    CU(cudaSetDevice(h->device));
    int rc;
    if ((rc = ilqrb_upload_u0(h, U))) return rc;
    if ((rc = launch_init(h))) return rc;
    const DevProblem& P = h->hp;
    const size_t B = P.B;
    const int H = P.H;
    if ((rc = download_packed(h, X, h->W.X, B * H, P.n, h->W.ns, nullptr, 1))) return rc;
    if (fX) {
        if ((rc = launch_final_fx(h, 0, (int)B, h->stream))) return rc;
        CU(cudaMemcpyAsync(fX, h->d_fX, B * H * P.nt * sizeof(double), cudaMemcpyDefault, h->stream));
    }
    if (J) {
        const size_t el = B * H * P.r * P.n;
        if ((rc = ensure_pack(h, el))) return rc;
        if ((rc = launch_rollout_jac(h, h->d_pack))) return rc;
        CU(cudaMemcpyAsync(J, h->d_pack, el * sizeof(double), cudaMemcpyDefault, h->stream));
    }
    if (cost) CU(cudaMemcpyAsync(cost, h->W.cost0, B * sizeof(double), cudaMemcpyDefault, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return 0;
}

int ilqrb_kinematics(ilqrb_handle h, int robot, int nq, const double* q, const double* dq, double* x, double* quat, double* J, double* dx, double* w) {
    if (!h) return fail(ILQRB_EINVAL, "null handle");
    if (robot < 0 || robot >= h->hp.n_robots) return fail(ILQRB_EINVAL, "robot out of range");
    if (nq < 1 || !q) return fail(ILQRB_EINVAL, "no configurations");
    CU(cudaSetDevice(h->device));
    const int dof = h->hp.robots[robot].dof;
    const size_t in = (size_t)nq * dof, outn = (size_t)nq * (3 + 4 + 6 * dof + 3 + 3);
    double* buf = nullptr;
    CU(cudaMalloc(&buf, (2 * in + outn) * sizeof(double)));
    double *dq_d = buf + in, *ox = buf + 2 * in, *oq = ox + (size_t)nq * 3, *oJ = oq + (size_t)nq * 4, *odx = oJ + (size_t)nq * 6 * dof, *ow = odx + (size_t)nq * 3;
    cudaMemcpyAsync(buf, q, in * sizeof(double), cudaMemcpyHostToDevice, h->stream);
    if (dq) cudaMemcpyAsync(dq_d, dq, in * sizeof(double), cudaMemcpyHostToDevice, h->stream);
    else cudaMemsetAsync(dq_d, 0, in * sizeof(double), h->stream);
    launch_kinematics(h, robot, nq, buf, dq_d, ox, oq, oJ, odx, ow);
    if (x) cudaMemcpyAsync(x, ox, (size_t)nq * 3 * sizeof(double), cudaMemcpyDeviceToHost, h->stream);
    if (quat) cudaMemcpyAsync(quat, oq, (size_t)nq * 4 * sizeof(double), cudaMemcpyDeviceToHost, h->stream);
    if (J) cudaMemcpyAsync(J, oJ, (size_t)nq * 6 * dof * sizeof(double), cudaMemcpyDeviceToHost, h->stream);
    if (dx) cudaMemcpyAsync(dx, odx, (size_t)nq * 3 * sizeof(double), cudaMemcpyDeviceToHost, h->stream);
    if (w) cudaMemcpyAsync(w, ow, (size_t)nq * 3 * sizeof(double), cudaMemcpyDeviceToHost, h->stream);
    cudaError_t e = cudaStreamSynchronize(h->stream);
    cudaFree(buf);
    if (e != cudaSuccess) return fail(ILQRB_ECUDA, std::string("kinematics: ") + cudaGetErrorString(e));
    return 0;
}

int ilqrb_set_profiling(ilqrb_handle h, int enabled) {
    if (!h) return fail(ILQRB_EINVAL, "null handle");
    h->resolve();
    h->prof = enabled != 0;
    return 0;
}
int ilqrb_get_profile(ilqrb_handle h, ilqrb_profile* out) {
    if (!h || !out) return fail(ILQRB_EINVAL, "null argument");
    h->resolve();
    for (int i = 0; i < ILQRB_PROF_SLOTS; i++) {
        out->ms[i] = h->prof_ms[i];
        out->launches[i] = h->prof_n[i];
        out->name[i] = kProfNames[i];
    }
    return 0;
}
int ilqrb_reset_profile(ilqrb_handle h) {
    if (!h) return fail(ILQRB_EINVAL, "null handle");
    h->resolve();
    for (int i = 0; i < ILQRB_PROF_SLOTS; i++) {
        h->prof_ms[i] = 0;
        h->prof_n[i] = 0;
    }
    return 0;
}
int64_t ilqrb_launch_count(ilqrb_handle h) { return h ? h->launches : 0; }

}  // extern "C"
