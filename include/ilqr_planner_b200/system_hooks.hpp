// system_hooks.hpp — bodies of sys::System's single-point methods (System.h:62-150): each evaluates on the GPU through the one-problem
// handle the system caches on first use (ilqrb_eval_point, ilqrb_rollout; include/ilqrb.h).  Included by solver.hpp after the
// descriptor builder.
#pragma once

namespace ilqr_planner {
namespace detail {

struct HookHandle {
    DescBuilder db;
    HandleGuard g;
    explicit HookHandle(const std::vector<Flat>& fl) : db(fl), g(db.desc(), 0) {}
};

}  // namespace detail

namespace sys {

inline System::EvalOut System::evalPoint(const VectorXd& xk, const VectorXd* uk, int k, const VectorXd* fx_in) {
    if (!hook_) {
        std::vector<detail::Flat> fl(1);
        flatten(fl[0]);
        hook_ = std::make_shared<detail::HookHandle>(fl);
    }
    const ilqrb_dims& dm = hook_->db.dims();
    if (xk.size() != dm.n) throw std::runtime_error("[System] state of size " + std::to_string(dm.n) + " expected");
    if (uk && uk->size() != dm.m) throw std::runtime_error("[System] control of size " + std::to_string(dm.m) + " expected");
    if (fx_in && fx_in->size() != dm.nt) throw std::runtime_error("[System] task vector of size " + std::to_string(dm.nt) + " expected");
    EvalOut o;
    o.fx.assign(size_t(dm.nt), 0.0);
    o.J.assign(size_t(dm.r) * dm.n, 0.0);
    o.diff.assign(size_t(dm.r), 0.0);
    o.cost.assign(1, 0.0);
    o.cost_x.assign(size_t(dm.n), 0.0);
    o.cost_xx.assign(size_t(dm.n) * dm.n, 0.0);
    const int32_t kk = k;
    // a timestep outside the horizon has no keypoint (System::getKeypoint returns null): same as "no timestep"
    const int32_t* kp = (k >= 0 && k < dm.H) ? &kk : nullptr;
    detail::check(ilqrb_eval_point(hook_->g.h, 1, nullptr, kp, xk.data(), uk ? uk->data() : nullptr, fx_in ? fx_in->data() : nullptr, o.fx.data(), o.J.data(), o.diff.data(),
                                   o.cost.data(), o.cost_x.data(), o.cost_xx.data()));
    return o;
}

inline VectorXd System::diff(const VectorXd& actual_state, int k) {
    EvalOut o = evalPoint(x0_, nullptr, k, &actual_state);
    return detail::from_buffer(o.diff.data(), long(o.diff.size()));
}

inline VectorXd System::diffBatch(const VectorXd& x) {
    const long nk = long(keypoints.size());
    if (x.size() != nk * nb_target_var_) throw std::runtime_error("[System] diffBatch: one task vector per keypoint expected");
    VectorXd residual = VectorXd::Zero(nk * nb_Q_var_);
    for (long i = 0; i < nk; i++) {
        const VectorXd d = diff(x.segment(i * nb_target_var_, nb_target_var_), keypoints[size_t(i)]->getTimestep());
        for (long j = 0; j < nb_Q_var_; j++) residual(i * nb_Q_var_ + j) = d(j);
    }
    return residual;
}

inline std::tuple<VectorXd, MatrixXd> System::getFxJac(VectorXd xk) {
    EvalOut o = evalPoint(xk, nullptr, -1, nullptr);
    return std::make_tuple(detail::from_buffer(o.fx.data(), long(o.fx.size())), detail::from_row_major(o.J.data(), long(o.diff.size()), xk.size()));
}

inline VectorXd System::cost(const VectorXd& xk, const VectorXd& uk, int k) {
    EvalOut o = evalPoint(xk, &uk, k, nullptr);
    return detail::from_buffer(o.cost.data(), 1);
}
inline VectorXd System::cost_x(const VectorXd& xk, const VectorXd& uk, int k) {
    EvalOut o = evalPoint(xk, &uk, k, nullptr);
    return detail::from_buffer(o.cost_x.data(), long(o.cost_x.size()));
}
inline MatrixXd System::cost_xx(const VectorXd& xk, const VectorXd& uk, int k) {
    EvalOut o = evalPoint(xk, &uk, k, nullptr);
    return detail::from_row_major(o.cost_xx.data(), xk.size(), xk.size());
}

// the comparison against the limits is all there is to it (System.cpp:121-142); time systems: the time entry has weight 0
inline std::pair<MatrixXd, VectorXd> System::inspectJointLimit(const VectorXd& x_k) {
    MatrixXd L = MatrixXd::Zero(nb_state_var_, nb_state_var_);
    VectorXd q = VectorXd::Zero(nb_state_var_);
    if (limits_set_) {
        for (long i = 0; i < nb_state_var_ && i < state_max_.size(); i++) {
            if (joint_limits_weight_[size_t(i)] == 0) continue;
            if (x_k(i) > state_max_(i)) {
                q(i) = state_max_(i) - x_k(i);
                L(i, i) = penalty_;
            } else if (x_k(i) < state_min_(i)) {
                q(i) = state_min_(i) - x_k(i);
                L(i, i) = penalty_;
            }
        }
    }
    return std::make_pair(L, q);
}

inline std::tuple<VectorXd, VectorXd, VectorXd, VectorXd, MatrixXd, MatrixXd, MatrixXd, MatrixXd> System::forwardPassWithLimits(const VectorXd& xk, const VectorXd& uk, int k) {
    auto f = forwardPass(xk, uk, k);
    auto Lq = inspectJointLimit(xk);
    return std::make_tuple(std::get<0>(f), std::get<1>(f), Lq.second, VectorXd::Zero(nb_ctrl_var_), std::get<2>(f), std::get<3>(f), std::get<4>(f), Lq.first);
}

// System.cpp:181-211 with the rollout, f(X) and the Jacobians of all timesteps from one ilqrb_rollout; like the reference, the
// limit terms of step i+1 are those of the state BEFORE that step
inline std::tuple<VectorXd, VectorXd, std::vector<std::tuple<MatrixXd, MatrixXd, MatrixXd, MatrixXd>>> System::fpBatch(const VectorXd& u) {
    if (!hook_) {
        std::vector<detail::Flat> fl(1);
        flatten(fl[0]);
        hook_ = std::make_shared<detail::HookHandle>(fl);
    }
    const ilqrb_dims& dm = hook_->db.dims();
    const int H = dm.H, n = dm.n, m = dm.m, nt = dm.nt, rq = dm.r;
    if (u.size() != long(H - 1) * m) throw std::runtime_error("[System] fpBatch: u must have (horizon-1)*nb_ctrl_var entries");
    reset();
    std::vector<double> X(size_t(H) * n), fXb(size_t(H) * nt), Jb(size_t(H) * rq * n);
    detail::check(ilqrb_rollout(hook_->g.h, u.data(), X.data(), fXb.data(), Jb.data(), nullptr));
    VectorXd fX = detail::from_buffer(fXb.data(), long(fXb.size())), qL = VectorXd::Zero(long(H) * n);
    std::vector<std::tuple<MatrixXd, MatrixXd, MatrixXd, MatrixXd>> ABJLs;
    ABJLs.push_back(std::make_tuple(MatrixXd::Identity(n, n), MatrixXd::Zero(n, m), detail::from_row_major(Jb.data(), rq, n), MatrixXd::Zero(n, n)));
    for (int i = 0; i < H - 1; i++) {
        const VectorXd xb = detail::from_buffer(X.data() + size_t(i) * n, n), xa = detail::from_buffer(X.data() + size_t(i + 1) * n, n);
        MatrixXd A, B;
        linearize(xa, u.segment(long(i) * m, m), A, B);
        auto Lq = inspectJointLimit(xb);
        for (int j = 0; j < n; j++) qL(long(i + 1) * n + j) = Lq.second(j);
        ABJLs.push_back(std::make_tuple(A, B, detail::from_row_major(Jb.data() + size_t(i + 1) * rq * n, rq, n), Lq.first));
    }
    // like the reference's step-by-step rollout, the simulation is left at the final state
    const int dof = r->getDOF();
    const VectorXd xf = detail::from_buffer(X.data() + size_t(H - 1) * n, n);
    r->setConfiguration(xf.head(dof), nb_deriv_ == 2 ? xf.segment(dof, dof) : VectorXd::Zero(dof));
    if (n > nb_deriv_ * dof) r->setTime(xf(n - 1));
    return std::make_tuple(fX, qL, ABJLs);
}

}  // namespace sys
}  // namespace ilqr_planner
