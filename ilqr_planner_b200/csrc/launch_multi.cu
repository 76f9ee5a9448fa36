// launch_multi.cu — one batch sharded over several GPUs of one box (include/ilqrb.h: ilqrb_multi_*).
//
// Problems are independent, so the batch splits into contiguous slices [lo_i, hi_i), one ilqrb_handle per device; every device runs all
// iterations on its slice with no exchange, and the only "collective" is the final gather: each handle writes its slice of the caller's
// result buffers itself (host buffers: D2H per device, all devices concurrently; device buffers: peer copies over NVLink).  Host side:
// one worker thread per device for the blocking calls, so that descriptor uploads, launch sequences and result downloads of the devices
// overlap.  Built on the public single-device C ABI only.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "handle.hpp"

struct ilqrb_multi_s {
    int n = 0;
    std::vector<int> device, lo, hi;
    std::vector<ilqrb_handle> h;
    ilqrb_dims dims;  // of the whole batch
};

using namespace ilqrb;

extern "C" int ilqrb_multi_plan(int batch, int n_slices, int* lo, int* hi) {
    if (batch < 1 || n_slices < 1 || !lo || !hi) return fail(ILQRB_EINVAL, "multi plan: batch and n_slices must be >= 1");
    if (n_slices > batch) return fail(ILQRB_EINVAL, "multi plan: more slices than problems");
    for (int i = 0; i < n_slices; i++) {  // contiguous, sizes differ by at most one, larger slices first
        lo[i] = (int)((long long)batch * i / n_slices);
        hi[i] = (int)((long long)batch * (i + 1) / n_slices);
    }
    return 0;
}

namespace {

// per-problem strides of the descriptor arrays (host pointers): a slice is the same descriptor with offset pointers
void slice_desc(const ilqrb_problem_desc& d, const int* task_r, int lo, int hi, ilqrb_problem_desc& o, std::vector<ilqrb_task>& tasks) {
    o = d;
    o.batch = hi - lo;
    o.q0 = d.q0 + (size_t)lo * d.dof;
    o.dq0 = d.dq0 + (size_t)lo * d.dof;
    tasks.assign(d.tasks, d.tasks + d.n_tasks);
    for (int t = 0; t < d.n_tasks; t++) {
        tasks[t].kp_target = d.tasks[t].kp_target + (size_t)lo * d.tasks[t].n_keypoints * ILQRB_KP_STRIDE;
        if (d.tasks[t].precision_per_problem) tasks[t].kp_precision = d.tasks[t].kp_precision + (size_t)lo * d.tasks[t].n_keypoints * task_r[t] * task_r[t];
    }
    o.tasks = tasks.data();
}

// residual size of one task (the arithmetic of ilqrb_get_dims, per task)
int task_residual_size(const ilqrb_problem_desc& d, int t) {
    const int time = d.sys_kind == ILQRB_SYS_POS_ORN_TIME ? 1 : 0;
    if (d.tasks[t].task_kind == ILQRB_TASK_JOINT) return d.nb_deriv * d.robots[d.tasks[t].robot].dof + time;
    if (d.sys_kind == ILQRB_SYS_POSITION) return (d.robots[d.tasks[t].robot].kind == ILQRB_ROBOT_2D ? 2 : 3) * d.nb_deriv;
    return 6 * d.nb_deriv + time;
}

ilqrb_result slice_result(const ilqrb_multi_s& m, const ilqrb_result& r, int i, int nit, int nc) {
    const size_t lo = m.lo[i];
    const int H = m.dims.H, n = m.dims.n, mm = m.dims.m, nt = m.dims.nt;
    ilqrb_result o = r;
    if (r.X) o.X = r.X + lo * H * n;
    if (r.fX) o.fX = r.fX + lo * H * nt;
    if (r.U) o.U = r.U + lo * (H - 1) * mm;
    if (r.K) o.K = r.K + lo * (H - 1) * mm * n;
    if (r.d) o.d = r.d + lo * (H - 1) * mm;
    if (r.cost) o.cost = r.cost + lo;
    if (r.cost_hist) o.cost_hist = r.cost_hist + lo * nit;
    if (r.alpha_hist) o.alpha_hist = r.alpha_hist + lo * nit;
    if (r.iters) o.iters = r.iters + lo;
    if (r.status) o.status = r.status + lo;
    if (r.lambda) o.lambda = r.lambda + lo * (H - 1) * nc;
    return o;
}

// f(i) on one worker thread per slice; the first failure's code and message are handed to the calling thread
template <class F>
int for_each_device(ilqrb_multi m, F f) {
    std::vector<int> rc(m->n, 0);
    std::vector<std::string> msg(m->n);
    std::vector<std::thread> th;
    for (int i = 0; i < m->n; i++)
        th.emplace_back([&, i]() {
            rc[i] = f(i);
            if (rc[i]) msg[i] = ilqrb_last_error();  // thread-local in the worker
        });
    for (auto& t : th) t.join();
    for (int i = 0; i < m->n; i++)
        if (rc[i]) return fail(rc[i], "device " + std::to_string(m->device[i]) + " (problems " + std::to_string(m->lo[i]) + ".." + std::to_string(m->hi[i]) + "): " + msg[i]);
    return 0;
}

}  // namespace

extern "C" {

int ilqrb_multi_destroy(ilqrb_multi m) {
    if (!m) return 0;
    for (auto h : m->h) ilqrb_destroy(h);
    delete m;
    return 0;
}

int ilqrb_create_multi(const ilqrb_problem_desc* desc, int n_devices, const int* devices, ilqrb_multi* out) {
    if (!out || !desc) return fail(ILQRB_EINVAL, "null argument");
    ilqrb_dims dm;
    int rc = ilqrb_get_dims(desc, &dm);
    if (rc) return rc;
    if (n_devices < 1 || n_devices > MAX_DEVICES) return fail(ILQRB_EINVAL, "n_devices out of range");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(ILQRB_ECUDA, "no CUDA device available (there is no CPU fallback)");
    ilqrb_multi m = new ilqrb_multi_s();
    m->n = n_devices;
    m->dims = dm;
    m->device.resize(n_devices);
    m->lo.resize(n_devices);
    m->hi.resize(n_devices);
    m->h.assign(n_devices, nullptr);
    for (int i = 0; i < n_devices; i++) {
        m->device[i] = devices ? devices[i] : i;  // NULL: devices 0 .. n_devices-1
        if (m->device[i] < 0 || m->device[i] >= ndev) {
            delete m;
            return fail(ILQRB_EINVAL, "device index out of range");
        }
    }
    if ((rc = ilqrb_multi_plan(desc->batch, n_devices, m->lo.data(), m->hi.data()))) {
        delete m;
        return rc;
    }
    std::vector<int> task_r(desc->n_tasks);
    for (int t = 0; t < desc->n_tasks; t++) task_r[t] = task_residual_size(*desc, t);
    // peer access between the devices of the set, for device-resident gathers (ilqrb_multi_download to device buffers)
    for (int i = 0; i < n_devices; i++)
        for (int j = 0; j < n_devices; j++) {
            if (m->device[i] == m->device[j]) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, m->device[i], m->device[j]);
            if (can) {
                cudaSetDevice(m->device[i]);
                cudaError_t e = cudaDeviceEnablePeerAccess(m->device[j], 0);
                if (e != cudaSuccess) cudaGetLastError();  // already enabled: fine
            }
        }
    rc = for_each_device(m, [&](int i) {
        ilqrb_problem_desc sd;
        std::vector<ilqrb_task> tasks;
        slice_desc(*desc, task_r.data(), m->lo[i], m->hi[i], sd, tasks);
        return ilqrb_create(&sd, m->device[i], &m->h[i]);
    });
    if (rc) {
        const std::string keep = ilqrb_last_error();
        ilqrb_multi_destroy(m);
        return fail(rc, keep);
    }
    *out = m;
    return 0;
}

int ilqrb_multi_count(ilqrb_multi m) { return m ? m->n : 0; }
int ilqrb_multi_slice(ilqrb_multi m, int i, int* device, int* lo, int* hi, ilqrb_handle* h) {
    if (!m || i < 0 || i >= m->n) return fail(ILQRB_EINVAL, "slice index out of range");
    if (device) *device = m->device[i];
    if (lo) *lo = m->lo[i];
    if (hi) *hi = m->hi[i];
    if (h) *h = m->h[i];
    return 0;
}

int ilqrb_multi_solve_recursive(ilqrb_multi m, const double* U0, const ilqrb_solve_opts* opts, ilqrb_result* out) {
    if (!m || !opts || !out) return fail(ILQRB_EINVAL, "null argument");
    const int nit = std::max(opts->nb_iter, 1);
    const size_t eu = (size_t)(m->dims.H - 1) * m->dims.m;
    return for_each_device(m, [&](int i) {
        ilqrb_result r = slice_result(*m, *out, i, nit, 0);
        return ilqrb_solve_recursive(m->h[i], U0 ? U0 + (size_t)m->lo[i] * eu : nullptr, opts, &r);
    });
}
int ilqrb_multi_solve_al(ilqrb_multi m, const double* U0, const ilqrb_al_desc* al, const ilqrb_solve_opts* opts, ilqrb_result* out) {
    if (!m || !opts || !out || !al) return fail(ILQRB_EINVAL, "null argument");
    const int nit = std::max(opts->nb_iter, 1);
    const size_t eu = (size_t)(m->dims.H - 1) * m->dims.m;
    return for_each_device(m, [&](int i) {
        ilqrb_result r = slice_result(*m, *out, i, nit, al->nc);
        return ilqrb_solve_al(m->h[i], U0 ? U0 + (size_t)m->lo[i] * eu : nullptr, al, opts, &r);
    });
}
int ilqrb_multi_solve_batch(ilqrb_multi m, const double* u0, const ilqrb_batch_desc* bd, const ilqrb_solve_opts* opts, ilqrb_result* out) {
    if (!m || !opts || !out) return fail(ILQRB_EINVAL, "null argument");
    const int nit = std::max(opts->nb_iter, 1);
    const size_t eu = (size_t)(m->dims.H - 1) * m->dims.m;
    return for_each_device(m, [&](int i) {
        ilqrb_result r = slice_result(*m, *out, i, nit, 0);
        return ilqrb_solve_batch(m->h[i], u0 ? u0 + (size_t)m->lo[i] * eu : nullptr, bd, opts, &r);
    });
}

// ---- split entry points: upload once, run (asynchronous on every device), download / gather once --------------------------------
int ilqrb_multi_upload_u0(ilqrb_multi m, const double* U0) {
    if (!m) return fail(ILQRB_EINVAL, "null argument");
    const size_t eu = (size_t)(m->dims.H - 1) * m->dims.m;
    for (int i = 0; i < m->n; i++) {
        int rc = ilqrb_upload_u0(m->h[i], U0 ? U0 + (size_t)m->lo[i] * eu : nullptr);
        if (rc) return rc;
    }
    return 0;
}
int ilqrb_multi_run_recursive(ilqrb_multi m, const ilqrb_solve_opts* opts) {
    if (!m || !opts) return fail(ILQRB_EINVAL, "null argument");
    for (int i = 0; i < m->n; i++) {  // launches only: the devices run concurrently
        int rc = ilqrb_run_recursive(m->h[i], opts);
        if (rc) return rc;
    }
    return 0;
}
int ilqrb_multi_synchronize(ilqrb_multi m) {
    if (!m) return fail(ILQRB_EINVAL, "null argument");
    for (int i = 0; i < m->n; i++) {
        int rc = ilqrb_synchronize(m->h[i]);
        if (rc) return rc;
    }
    return 0;
}
// The final gather: every device packs its slice and copies it into the whole-batch buffers of `out`, which may be host memory or
// device memory of any GPU of the process (unified addressing: peer copies travel over NVLink between the devices of the set).
int ilqrb_multi_download(ilqrb_multi m, ilqrb_result* out, int nb_iter) {
    if (!m || !out) return fail(ILQRB_EINVAL, "null argument");
    const int nit = std::max(nb_iter, 1);
    return for_each_device(m, [&](int i) {
        ilqrb_result r = slice_result(*m, *out, i, nit, m->h[i]->W.nc);
        return ilqrb_download(m->h[i], &r);
    });
}

}  // extern "C"
