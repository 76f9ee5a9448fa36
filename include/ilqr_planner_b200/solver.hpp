// solver.hpp — host mirror of ilqr_planner::solver::{ILQRRecursive, AL_ILQR, BatchILQR, BatchILQRCP} and CallBackMessage
// (ILQRRecursive.h:18-41, AL-ILQR.h:19-70, BatchILQR.h:18-77, BatchILQRCP.h:18-77, CallbackMessage.h:12-16).
//
// `solve(...)` has the reference's signature, return type and per-iteration message; it runs a batch of one problem on the
// GPU through the C ABI (include/ilqrb.h).  `solveBatch(...)` is the batched overload: B systems of one structure (same
// chain, horizon, keypoint timesteps) that differ in initial state, keypoint targets and precisions.
#pragma once
#include <iostream>
#include <memory>
#include <sstream>
#include <string>

#include "system.hpp"

namespace ilqr_planner {

class CallBackMessage {
public:
    virtual ~CallBackMessage() {}
    virtual void notify(const std::string& message) = 0;
};

namespace detail {

// contiguous arrays + ilqrb_problem_desc for B flattened systems
class DescBuilder {
public:
    explicit DescBuilder(const std::vector<Flat>& fl) {
        if (fl.empty()) throw std::runtime_error("[solver] empty batch");
        const Flat& f0 = fl[0];
        const size_t B = fl.size(), T = f0.tasks.size();
        robots_.reserve(T);
        tasks_.resize(T);
        ts_.resize(T);
        tg_.resize(T);
        pr_.resize(T);
        std::vector<const void*> ids;
        for (size_t t = 0; t < T; t++) {
            const TaskFlat& k0 = f0.tasks[t];
            size_t ri = 0;
            for (; ri < ids.size(); ri++)
                if (ids[ri] == k0.view.base_id) break;
            if (ri == ids.size()) {
                ids.push_back(k0.view.base_id);
                views_.push_back(k0.view);
            }
            ts_[t] = k0.timestep;
            bool same_prec = true;
            for (size_t b = 0; b < B; b++) {
                const Flat& f = fl[b];
                if (f.tasks.size() != T || f.sys_kind != f0.sys_kind || f.nb_deriv != f0.nb_deriv || f.horizon != f0.horizon || f.dof != f0.dof || f.dt != f0.dt ||
                    f.tasks[t].timestep != k0.timestep)
                    throw std::runtime_error("[solver] the systems of a batch must share structure (kind, order, horizon, dt, keypoint timesteps)");
                tg_[t].insert(tg_[t].end(), f.tasks[t].target.begin(), f.tasks[t].target.end());
                if (f.tasks[t].precision != k0.precision) same_prec = false;
            }
            if (same_prec) pr_[t] = k0.precision;
            else
                for (size_t b = 0; b < B; b++) pr_[t].insert(pr_[t].end(), fl[b].tasks[t].precision.begin(), fl[b].tasks[t].precision.end());
            ilqrb_task& o = tasks_[t];
            o = ilqrb_task{};
            o.robot = int32_t(ri);
            o.has_transform = k0.view.has_transform ? 1 : 0;
            for (int i = 0; i < 16; i++) o.T[i] = k0.view.T[i];
            o.n_keypoints = int32_t(k0.timestep.size());
            o.precision_per_problem = same_prec ? 0 : 1;
            o.limits_set = k0.limits_set ? 1 : 0;
            for (size_t b = 1; b < B; b++)
                if (fl[b].tasks[t].task_kind != k0.task_kind || fl[b].tasks[t].deadzone != k0.deadzone)
                    throw std::runtime_error("[solver] the systems of a batch must share their task kinds and keypoint dead zones");
        }
        for (size_t b = 0; b < B; b++) {
            q0_.insert(q0_.end(), fl[b].q0.begin(), fl[b].q0.end());
            dq0_.insert(dq0_.end(), fl[b].dq0.begin(), fl[b].dq0.end());
        }
        flat0_ = f0;
        finish(int32_t(B), q0_.data(), dq0_.data());
    }
    // B problems that share the structure (and precisions) of one system and differ in the initial joint state and the keypoint
    // targets; the caller's arrays enter the descriptor as they are: q0 [B][dof], dq0 [B][dof] or nullptr (the system's own, for all),
    // kp_target[t] [B][Kp_t][ILQRB_KP_STRIDE] or nullptr (the system's own, for all)
    DescBuilder(const Flat& f0, int B, const double* q0, const double* dq0, const std::vector<const double*>& kp_target) {
        if (B < 1) throw std::runtime_error("[solver] empty batch");
        const size_t T = f0.tasks.size();
        if (kp_target.size() != T) throw std::runtime_error("[solver] one keypoint-target array per sub-system expected");
        tasks_.resize(T);
        ts_.resize(T);
        tg_.resize(T);
        pr_.resize(T);
        tg_ext_.assign(T, nullptr);
        std::vector<const void*> ids;
        for (size_t t = 0; t < T; t++) {
            const TaskFlat& k0 = f0.tasks[t];
            size_t ri = 0;
            for (; ri < ids.size(); ri++)
                if (ids[ri] == k0.view.base_id) break;
            if (ri == ids.size()) {
                ids.push_back(k0.view.base_id);
                views_.push_back(k0.view);
            }
            ts_[t] = k0.timestep;
            pr_[t] = k0.precision;
            if (kp_target[t]) tg_ext_[t] = kp_target[t];
            else
                for (int b = 0; b < B; b++) tg_[t].insert(tg_[t].end(), k0.target.begin(), k0.target.end());
            ilqrb_task& o = tasks_[t];
            o = ilqrb_task{};
            o.robot = int32_t(ri);
            o.has_transform = k0.view.has_transform ? 1 : 0;
            for (int i = 0; i < 16; i++) o.T[i] = k0.view.T[i];
            o.n_keypoints = int32_t(k0.timestep.size());
            o.precision_per_problem = 0;
            o.limits_set = k0.limits_set ? 1 : 0;
        }
        if (!q0)
            for (int b = 0; b < B; b++) q0_.insert(q0_.end(), f0.q0.begin(), f0.q0.end());
        if (!dq0)
            for (int b = 0; b < B; b++) dq0_.insert(dq0_.end(), f0.dq0.begin(), f0.dq0.end());
        flat0_ = f0;
        finish(B, q0 ? q0 : q0_.data(), dq0 ? dq0 : dq0_.data());
    }
    DescBuilder(const DescBuilder&) = delete;
    DescBuilder& operator=(const DescBuilder&) = delete;
    const ilqrb_problem_desc& desc() const { return desc_; }
    const ilqrb_dims& dims() const { return dims_; }
    // everything a handle keeps fixed: two batches with equal signatures differ only in q0, dq0 and the keypoint targets, which
    // ilqrb_update_problems replaces in place
    std::string signature() const {
        std::string s;
        auto put = [&s](const void* p, size_t bytes) { s.append(static_cast<const char*>(p), bytes); };
        auto puti = [&](int64_t v) { put(&v, sizeof v); };
        const int32_t head[] = {desc_.sys_kind, desc_.nb_deriv, desc_.horizon, desc_.batch, desc_.dof, desc_.n_robots, desc_.n_tasks};
        put(head, sizeof head);
        put(&desc_.dt, sizeof(double));
        put(flat0_.r_diag.data(), flat0_.r_diag.size() * sizeof(double));
        for (auto& v : views_) {
            puti(v.kind);
            puti(v.dof);
            put(v.segments.data(), v.segments.size() * sizeof(ilqrb_segment));
            put(v.lengths.data(), v.lengths.size() * sizeof(double));
        }
        for (size_t t = 0; t < tasks_.size(); t++) {
            const ilqrb_task& o = tasks_[t];
            const TaskFlat& k0 = flat0_.tasks[t];
            const int32_t th[] = {o.robot, o.joint_offset, o.has_transform, o.n_keypoints, o.precision_per_problem, o.limits_set, o.task_kind};
            put(th, sizeof th);
            put(o.T, sizeof o.T);
            put(ts_[t].data(), ts_[t].size() * sizeof(int32_t));
            put(pr_[t].data(), pr_[t].size() * sizeof(double));
            put(k0.r_diag.data(), k0.r_diag.size() * sizeof(double));
            put(k0.state_max.data(), k0.state_max.size() * sizeof(double));
            put(k0.state_min.data(), k0.state_min.size() * sizeof(double));
            put(k0.weight.data(), k0.weight.size() * sizeof(int32_t));
            put(k0.deadzone.data(), k0.deadzone.size() * sizeof(double));
        }
        return s;
    }
    size_t taskCount() const { return tasks_.size(); }

private:
    // robots, pointers and dims once the per-task arrays are in place
    void finish(int32_t B, const double* q0, const double* dq0) {
        const Flat& f0 = flat0_;
        const size_t T = tasks_.size();
        for (auto& v : views_) {
            ilqrb_robot rb{};
            rb.kind = v.kind;
            rb.dof = v.dof;
            rb.n_segments = int32_t(v.segments.size());
            rb.segments = v.segments.data();
            rb.lengths = v.lengths.empty() ? nullptr : v.lengths.data();
            robots_.push_back(rb);
        }
        for (size_t t = 0; t < T; t++) {
            ilqrb_task& o = tasks_[t];
            const TaskFlat& k0 = flat0_.tasks[t];
            o.kp_timestep = ts_[t].data();
            o.kp_target = (!tg_ext_.empty() && tg_ext_[t]) ? tg_ext_[t] : tg_[t].data();
            o.task_kind = k0.task_kind;
            o.kp_deadzone = k0.deadzone.empty() ? nullptr : k0.deadzone.data();
            o.kp_precision = pr_[t].data();
            o.r_diag = k0.r_diag.data();
            if (k0.limits_set) {
                o.state_max = k0.state_max.data();
                o.state_min = k0.state_min.data();
                o.limit_weight = k0.weight.data();
            }
        }
        desc_ = ilqrb_problem_desc{};
        desc_.abi_version = ILQRB_ABI_VERSION;
        desc_.sys_kind = f0.sys_kind;
        desc_.nb_deriv = f0.nb_deriv;
        desc_.horizon = f0.horizon;
        desc_.batch = B;
        desc_.dof = f0.dof;
        desc_.n_robots = int32_t(robots_.size());
        desc_.n_tasks = int32_t(T);
        desc_.dt = f0.dt;
        desc_.robots = robots_.data();
        desc_.tasks = tasks_.data();
        desc_.r_diag = flat0_.r_diag.data();
        desc_.q0 = q0;
        desc_.dq0 = dq0;
        check(ilqrb_get_dims(&desc_, &dims_));
    }
    Flat flat0_;
    std::vector<RobotView> views_;
    std::vector<ilqrb_robot> robots_;
    std::vector<ilqrb_task> tasks_;
    std::vector<std::vector<int32_t>> ts_;
    std::vector<std::vector<double>> tg_, pr_;
    std::vector<const double*> tg_ext_;  // caller-owned target arrays (arrays constructor)
    std::vector<double> q0_, dq0_;
    ilqrb_problem_desc desc_;
    ilqrb_dims dims_;
};

inline std::vector<Flat> flatten_all(const std::vector<std::shared_ptr<sys::System>>& systems) {
    std::vector<Flat> fl(systems.size());
    for (size_t i = 0; i < systems.size(); i++) systems[i]->flatten(fl[i]);
    return fl;
}

struct HandleGuard {
    ilqrb_handle h = nullptr;
    HandleGuard(const ilqrb_problem_desc& d, int device) { check(ilqrb_create(&d, device, &h)); }
    ~HandleGuard() {
        if (h) ilqrb_destroy(h);
    }
    HandleGuard(const HandleGuard&) = delete;
    HandleGuard& operator=(const HandleGuard&) = delete;
};

// The handle (or, for several devices, the multi-handle) a solver object keeps between solves.  A batch whose signature equals
// the previous one's only replaces q0, dq0 and the keypoint targets on the device (ilqrb_update_problems); anything else
// (another structure, batch size, precision, device list) builds a new handle.
class SolverCache {
public:
    SolverCache() {}
    ~SolverCache() { drop(); }
    SolverCache(const SolverCache&) = delete;
    SolverCache& operator=(const SolverCache&) = delete;
    // after the call: handle() or multi() is ready for `db`'s problems; `db` must outlive the solve
    void prepare(const DescBuilder& db, const std::vector<int>& devices) {
        const std::string sig = db.signature();
        const ilqrb_problem_desc& d = db.desc();
        if ((h_ || m_) && sig == sig_ && devices == devices_) {
            if (h_) {
                refresh(h_, d, 0);
            } else {
                const int ns = ilqrb_multi_count(m_);
                for (int i = 0; i < ns; i++) {
                    int lo = 0, hi = 0;
                    ilqrb_handle hs = nullptr;
                    check(ilqrb_multi_slice(m_, i, nullptr, &lo, &hi, &hs));
                    refresh(hs, d, lo);
                }
            }
            reused_++;
            return;
        }
        drop();
        if (devices.size() <= 1) check(ilqrb_create(&d, devices.empty() ? 0 : devices[0], &h_));
        else check(ilqrb_create_multi(&d, int(devices.size()), devices.data(), &m_));
        sig_ = sig;
        devices_ = devices;
        created_++;
    }
    ilqrb_handle handle() const { return h_; }
    ilqrb_multi multi() const { return m_; }
    int created() const { return created_; }
    int reused() const { return reused_; }

private:
    static void refresh(ilqrb_handle h, const ilqrb_problem_desc& d, int lo) {
        check(ilqrb_update_problems(h, d.q0 + size_t(lo) * d.dof, d.dq0 + size_t(lo) * d.dof, -1, nullptr));
        for (int t = 0; t < d.n_tasks; t++)
            check(ilqrb_update_problems(h, nullptr, nullptr, t, d.tasks[t].kp_target + size_t(lo) * d.tasks[t].n_keypoints * ILQRB_KP_STRIDE));
    }
    void drop() {
        if (h_) ilqrb_destroy(h_);
        if (m_) ilqrb_multi_destroy(m_);
        h_ = nullptr;
        m_ = nullptr;
    }
    ilqrb_handle h_ = nullptr;
    ilqrb_multi m_ = nullptr;
    std::string sig_;
    std::vector<int> devices_;
    int created_ = 0, reused_ = 0;
};

// page-locked host memory for the results of the array-based batch solves (ilqrb_host_alloc).  One block, reused by every solve
// it is large enough for; a larger request takes a new block and keeps the old one until the arena dies, so that result views
// handed out earlier never dangle (they go stale, not invalid).
class PinnedArena {
public:
    PinnedArena() {}
    ~PinnedArena() {
        for (void* q : blocks_) ilqrb_host_free(q);
    }
    PinnedArena(const PinnedArena&) = delete;
    PinnedArena& operator=(const PinnedArena&) = delete;
    void* get(size_t bytes) {
        if (bytes > cap_) {
            void* q = nullptr;
            check(ilqrb_host_alloc(bytes + bytes / 4, &q));
            blocks_.push_back(q);
            p_ = q;
            cap_ = bytes + bytes / 4;
        }
        return p_;
    }

private:
    void* p_ = nullptr;
    size_t cap_ = 0;
    std::vector<void*> blocks_;
};

// keypoint states in the layout of Keypoint::getState() -> ILQRB_KP_STRIDE records
//   POSE   nd=1: [p3 quat4 (t)]           nd=2: [p3 dp3 quat4 dquat4 (t)]   (PosOrnKeypoint.cpp:13-22, SpacetimeKeypoint.cpp:12-17)
//   JOINT  nd=1: [q(d) (t)]               nd=2: [q(d) dq(d) (t)]            (AngularKeypoint.cpp:13-20, AngularTimeKeypoint.cpp:12-17)
//   POSITION (Robot2D planner): [p(car) (dp(car))]
inline void states_to_records(const Flat& f, const TaskFlat& tk, size_t count, const double* st, double* rec) {
    const int nd = f.nb_deriv, tt = f.sys_kind == ILQRB_SYS_POS_ORN_TIME ? 1 : 0;
    const bool joint = tk.task_kind == ILQRB_TASK_JOINT, position = f.sys_kind == ILQRB_SYS_POSITION;
    const int d = joint ? tk.view.dof : position ? (tk.view.kind == ILQRB_ROBOT_2D ? 2 : 3) : 0;
    const int w = joint || position ? nd * d + tt : 7 * nd + tt;
    for (size_t i = 0; i < count; i++, st += w, rec += ILQRB_KP_STRIDE) {
        for (int j = 0; j < ILQRB_KP_STRIDE; j++) rec[j] = 0.0;
        if (joint || position) {
            for (int j = 0; j < d; j++) rec[j] = st[j];
            if (nd == 2)
                for (int j = 0; j < d; j++) rec[7 + j] = st[d + j];
        } else if (nd == 1) {
            for (int j = 0; j < 7; j++) rec[j] = st[j];
        } else {
            for (int j = 0; j < 3; j++) rec[j] = st[j], rec[7 + j] = st[3 + j];
            for (int j = 0; j < 4; j++) rec[3 + j] = st[6 + j], rec[10 + j] = st[10 + j];
        }
        if (tt) rec[14] = st[w - 1];
    }
}
inline int state_width(const Flat& f, const TaskFlat& tk) {
    const int nd = f.nb_deriv, tt = f.sys_kind == ILQRB_SYS_POS_ORN_TIME ? 1 : 0;
    if (tk.task_kind == ILQRB_TASK_JOINT) return nd * tk.view.dof + tt;
    if (f.sys_kind == ILQRB_SYS_POSITION) return nd * (tk.view.kind == ILQRB_ROBOT_2D ? 2 : 3);
    return 7 * nd + tt;
}

// "%g"-like shortest form the reference's operator<< produces for doubles (6 significant digits)
inline std::string fmt(double v) {
    std::ostringstream s;
    s << v;
    return s.str();
}

inline void emit(CallBackMessage* cb, const std::string& line) {
    if (cb) cb->notify(line);
    else std::cout << line << std::endl;
}

}  // namespace detail

namespace solver {

// flat, batch-major results of solveBatch (layouts of ilqrb_result)
struct BatchResult {
    int B = 0, H = 0, n = 0, m = 0, nt = 0, nc = 0, nb_iter = 0;
    std::vector<double> X, fX, U, K, d, cost, cost_hist, alpha_hist, lambda;
    std::vector<int32_t> iters, status;
};

// results of the array-based batch solves: pointers into the solver's page-locked arena, valid until its next batch solve
struct BatchView {
    int B = 0, H = 0, n = 0, m = 0, nt = 0, nc = 0, nb_iter = 0;
    double *X = nullptr, *fX = nullptr, *U = nullptr, *K = nullptr, *d = nullptr, *cost = nullptr, *cost_hist = nullptr, *alpha_hist = nullptr, *lambda = nullptr;
    int32_t *iters = nullptr, *status = nullptr;
    bool anyQuuNotSpd() const {
        for (int b = 0; b < B; b++)
            if (status[b] & ILQRB_ST_QUU_NOT_SPD) return true;
        return false;
    }
};

// per-problem inputs of an array-based batch solve; every array is the caller's and is only read
struct BatchArrays {
    int B = 0;
    const double* q0 = nullptr;              // [B][dof]; nullptr: the system's initial joint positions for all
    const double* dq0 = nullptr;             // [B][dof]; nullptr: the system's initial joint velocities for all
    std::vector<const double*> kp_states;    // per sub-system [B][Kp][w] in Keypoint::getState() layout; nullptr: the system's own keypoints
    const double* U0 = nullptr;              // [B][H-1][m] (Batch solvers: [B][(H-1) m]); nullptr: zeros
};

namespace detail_solver {

// the state shared by the four solver classes for their batch entry points
struct BatchState {
    detail::SolverCache cache;
    detail::PinnedArena arena;
    std::unique_ptr<detail::DescBuilder> db;
    std::vector<std::vector<double>> records;  // [task][B][Kp][ILQRB_KP_STRIDE]
};

// descriptor + handle for `in`, result pointers carved from the arena
inline BatchView prepare_arrays(BatchState& st, const sys::System& s, const BatchArrays& in, int nb_iter, bool gains, int nc, bool batch_solver, const std::vector<int>& devices,
                                ilqrb_result& o) {
    detail::Flat f;
    s.flatten(f);
    const size_t T = f.tasks.size();
    if (!in.kp_states.empty() && in.kp_states.size() != T) throw std::runtime_error("[solver] one keypoint-state array per sub-system expected");
    st.records.resize(T);
    std::vector<const double*> tg(T, nullptr);
    for (size_t t = 0; t < T; t++) {
        if (in.kp_states.empty() || !in.kp_states[t]) continue;
        const size_t count = size_t(in.B) * f.tasks[t].timestep.size();
        st.records[t].resize(count * ILQRB_KP_STRIDE);
        detail::states_to_records(f, f.tasks[t], count, in.kp_states[t], st.records[t].data());
        tg[t] = st.records[t].data();
    }
    st.db.reset(new detail::DescBuilder(f, in.B, in.q0, in.dq0, tg));
    st.cache.prepare(*st.db, devices);
    const ilqrb_dims& dm = st.db->dims();
    BatchView v;
    v.B = dm.B, v.H = dm.H, v.n = dm.n, v.m = dm.m, v.nt = dm.nt, v.nc = nc, v.nb_iter = nb_iter;
    const size_t B = size_t(dm.B), H = size_t(dm.H), it = size_t(nb_iter > 0 ? nb_iter : 1);
    const size_t nU = B * (H - 1) * dm.m, nX = batch_solver ? 0 : B * H * dm.n, nfX = batch_solver ? 0 : B * H * dm.nt, nK = gains ? B * (H - 1) * dm.m * dm.n : 0,
                 nd = gains ? nU : 0, nl = B * (H - 1) * size_t(nc);
    const size_t doubles = nU + nX + nfX + nK + nd + nl + B + 2 * B * it, ints = 2 * B;
    double* p = static_cast<double*>(st.arena.get(doubles * sizeof(double) + ints * sizeof(int32_t)));
    auto take = [&p](size_t count) {
        double* q = count ? p : nullptr;
        p += count;
        return q;
    };
    v.K = take(nK);
    v.X = take(nX);
    v.fX = take(nfX);
    v.U = take(nU);
    v.d = take(nd);
    v.lambda = take(nl);
    v.cost = take(B);
    v.cost_hist = take(B * it);
    v.alpha_hist = take(B * it);
    v.iters = reinterpret_cast<int32_t*>(p);
    v.status = v.iters + B;
    o = ilqrb_result{};
    o.X = v.X, o.fX = v.fX, o.U = v.U, o.K = v.K, o.d = v.d, o.lambda = v.lambda;
    o.cost = v.cost, o.cost_hist = v.cost_hist, o.alpha_hist = v.alpha_hist, o.iters = v.iters, o.status = v.status;
    return v;
}

inline void alloc(BatchResult& r, const ilqrb_dims& dm, int nb_iter, bool gains, int nc, bool batch_solver, ilqrb_result& o) {
    r.B = dm.B;
    r.H = dm.H;
    r.n = dm.n;
    r.m = dm.m;
    r.nt = dm.nt;
    r.nc = nc;
    r.nb_iter = nb_iter;
    const size_t B = size_t(dm.B), H = size_t(dm.H), it = size_t(nb_iter > 0 ? nb_iter : 1);
    o = ilqrb_result{};
    r.U.assign(B * (H - 1) * dm.m, 0.0);
    o.U = r.U.data();
    if (!batch_solver) {
        r.X.assign(B * H * dm.n, 0.0);
        r.fX.assign(B * H * dm.nt, 0.0);
        o.X = r.X.data();
        o.fX = r.fX.data();
    }
    if (gains) {
        r.K.assign(B * (H - 1) * dm.m * dm.n, 0.0);
        r.d.assign(B * (H - 1) * dm.m, 0.0);
        o.K = r.K.data();
        o.d = r.d.data();
    }
    if (nc > 0) {
        r.lambda.assign(B * (H - 1) * size_t(nc), 0.0);
        o.lambda = r.lambda.data();
    }
    r.cost.assign(B, 0.0);
    r.cost_hist.assign(B * it, 0.0);
    r.alpha_hist.assign(B * it, 0.0);
    r.iters.assign(B, 0);
    r.status.assign(B, 0);
    o.cost = r.cost.data();
    o.cost_hist = r.cost_hist.data();
    o.alpha_hist = r.alpha_hist.data();
    o.iters = r.iters.data();
    o.status = r.status.data();
}

inline std::vector<double> stack_u(const std::vector<std::vector<VectorXd>>& U0s, int H, int m) {
    std::vector<double> u;
    u.reserve(U0s.size() * size_t(H - 1) * m);
    for (auto& U0 : U0s) {
        if (int(U0.size()) != H - 1) throw std::runtime_error("[solver] U0 must hold horizon-1 control vectors");
        for (auto& v : U0) {
            if (v.size() != m) throw std::runtime_error("[solver] control vector of wrong size");
            for (long i = 0; i < m; i++) u.push_back(v(i));
        }
    }
    return u;
}

inline std::vector<VectorXd> rows(const std::vector<double>& a, size_t off, int count, int width) {
    std::vector<VectorXd> o;
    o.reserve(size_t(count));
    for (int k = 0; k < count; k++) o.push_back(detail::from_buffer(a.data() + off + size_t(k) * width, width));
    return o;
}

}  // namespace detail_solver
}  // namespace solver
}  // namespace ilqr_planner

#include "system_hooks.hpp"
#include "solver_classes.hpp"
