// solver.hpp — host mirror of ilqr_planner::solver::{ILQRRecursive, AL_ILQR, BatchILQR, BatchILQRCP} and CallBackMessage
// (ILQRRecursive.h:18-41, AL-ILQR.h:19-70, BatchILQR.h:18-77, BatchILQRCP.h:18-77, CallbackMessage.h:12-16).
//
// `solve(...)` has the reference's signature, return type and per-iteration message; it runs a batch of one problem on the
// GPU through the C ABI (include/ilqrb.h).  `solveBatch(...)` is the batched overload: B systems of one structure (same
// chain, horizon, keypoint timesteps) that differ in initial state, keypoint targets and precisions.
#pragma once
#include <iostream>
#include <sstream>

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
        }
        for (size_t b = 0; b < B; b++) {
            q0_.insert(q0_.end(), fl[b].q0.begin(), fl[b].q0.end());
            dq0_.insert(dq0_.end(), fl[b].dq0.begin(), fl[b].dq0.end());
        }
        flat0_ = f0;
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
            o.kp_target = tg_[t].data();
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
        desc_.batch = int32_t(B);
        desc_.dof = f0.dof;
        desc_.n_robots = int32_t(robots_.size());
        desc_.n_tasks = int32_t(T);
        desc_.dt = f0.dt;
        desc_.robots = robots_.data();
        desc_.tasks = tasks_.data();
        desc_.r_diag = flat0_.r_diag.data();
        desc_.q0 = q0_.data();
        desc_.dq0 = dq0_.data();
        check(ilqrb_get_dims(&desc_, &dims_));
    }
    DescBuilder(const DescBuilder&) = delete;
    DescBuilder& operator=(const DescBuilder&) = delete;
    const ilqrb_problem_desc& desc() const { return desc_; }
    const ilqrb_dims& dims() const { return dims_; }

private:
    Flat flat0_;
    std::vector<RobotView> views_;
    std::vector<ilqrb_robot> robots_;
    std::vector<ilqrb_task> tasks_;
    std::vector<std::vector<int32_t>> ts_;
    std::vector<std::vector<double>> tg_, pr_;
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

namespace detail_solver {

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

#include "solver_classes.hpp"
