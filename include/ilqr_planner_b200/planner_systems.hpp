// planner_systems.hpp — sys::PosOrnPlannerSys (PosOrnPlannerSys.h:18-77), sys::PosOrnTimePlannerSys (PosOrnTimePlannerSys.h:18-72),
// sys::JointSpacePlannerSys (JointSpacePlannerSys.h:17-66), sys::JointSpaceTimePlannerSys (JointSpaceTimePlannerSys.h),
// sys::SequentialSystem (SequentialSystem.h:15-62) and the position-only planner of README.md:15 (SURVEY §8g note C1).
// Included by system.hpp.
#pragma once

namespace ilqr_planner {
namespace sys {
namespace detail_sys {

// [p, quat (, dp, dquat) (, t)] and the block Jacobian of one robot view (PosOrnPlannerSys.cpp:80-102, PosOrnTimePlannerSys.cpp:85-112)
inline std::tuple<VectorXd, MatrixXd> fx_jac(sim::SimulationInterface& rb, int nb_deriv, bool time) {
    const VectorXd p = rb.getEEPosition(), e = rb.getEEOrnQuat();
    const MatrixXd Jk = rb.J();
    const long jr = Jk.rows(), jc = Jk.cols(), tt = time ? 1 : 0;
    VectorXd fx = VectorXd::Zero(7 * nb_deriv + tt);
    for (int i = 0; i < 3; i++) fx(i) = p(i);
    for (int i = 0; i < 4; i++) fx(3 + i) = e(i);
    MatrixXd J = MatrixXd::Zero(nb_deriv * jr + tt, nb_deriv * jc + tt);
    for (int b = 0; b < nb_deriv; b++)
        for (long i = 0; i < jr; i++)
            for (long j = 0; j < jc; j++) J(b * jr + i, b * jc + j) = Jk(i, j);
    if (nb_deriv == 2) {
        const VectorXd dp = rb.getEEVelocity(), de = rb.getEEAngVelQuat();
        for (int i = 0; i < 3; i++) fx(7 + i) = dp(i);
        for (int i = 0; i < 4; i++) fx(10 + i) = de(i);
    }
    if (time) {
        fx(fx.size() - 1) = rb.getTime();
        J(J.rows() - 1, J.cols() - 1) = 1;
    }
    return std::make_tuple(fx, J);
}

inline VectorXd joint_state(sim::SimulationInterface& rb, int nb_deriv, bool time) {
    const VectorXd q = rb.getJointsPos(), dq = rb.getJointsVel();
    const long d = q.size();
    VectorXd x = VectorXd::Zero(nb_deriv * d + (time ? 1 : 0));
    for (long i = 0; i < d; i++) {
        x(i) = q(i);
        if (nb_deriv == 2) x(d + i) = dq(i);
    }
    if (time) x(x.size() - 1) = rb.getTime();
    return x;
}

// A, B of the fixed-dt integrators (PosOrnPlannerSys.cpp:114-128, JointSpacePlannerSys.cpp:93-107)
inline void linearize_fixed_dt(int d, int nb_deriv, int n, int m, double dt, MatrixXd& A, MatrixXd& B) {
    A = MatrixXd::Identity(n, n);
    B = MatrixXd::Zero(n, m);
    for (int i = 0; i < d; i++) {
        if (nb_deriv == 1) {
            B(i, i) = dt;
        } else {
            A(i, d + i) = dt;
            B(i, i) = dt * dt / 2;
            B(d + i, i) = dt;
        }
    }
}
// A, B of the time systems, dt = s^2 with s the last control (PosOrnTimePlannerSys.cpp:149-176, JointSpaceTimePlannerSys.cpp:116-146);
// v = joint velocity AFTER the step (reference behaviour)
inline void linearize_time(int d, int nb_deriv, int n, int m, const VectorXd& x_after, const VectorXd& uk, MatrixXd& A, MatrixXd& B) {
    const int mc = m - 1;
    const double s = uk(mc), dt = s * s;
    A = MatrixXd::Identity(n, n);
    B = MatrixXd::Zero(n, m);
    for (int i = 0; i < d; i++) {
        if (nb_deriv == 1) {
            B(i, i) = dt;
            B(i, mc) = 2 * s * uk(i);
        } else {
            A(i, d + i) = dt;
            B(i, i) = dt * dt / 2;
            B(d + i, i) = dt;
            B(i, mc) = 2 * s * x_after(d + i) + 2 * s * s * s * uk(i);
            B(d + i, mc) = 2 * s * uk(i);
        }
    }
    B(n - 1, mc) = 2 * s;
}

}  // namespace detail_sys

class PosOrnPlannerSys : public System {
public:
    PosOrnPlannerSys(const std::shared_ptr<sim::SimulationInterface>& r_, const std::vector<std::shared_ptr<Keypoint>>& kps, const VectorXd& RtDiag, const VectorXd& qMax,
                     const VectorXd& qMin, const VectorXd& dqMax, const VectorXd& dqMin, int horizon, int nbDeriv, double dt)
        : System(r_, kps, RtDiag, qMax, qMin, dqMax, dqMin, horizon, nbDeriv, {"POS_ORN"}) {
        localInit(dt);
    }
    PosOrnPlannerSys(const std::shared_ptr<sim::SimulationInterface>& r_, const std::vector<std::shared_ptr<Keypoint>>& kps, const VectorXd& RtDiag, const VectorXd& qMax,
                     const VectorXd& qMin, int horizon, int nbDeriv, double dt)
        : System(r_, kps, RtDiag, qMax, qMin, horizon, nbDeriv, {"POS_ORN"}) {
        localInit(dt);
    }
    PosOrnPlannerSys(const std::shared_ptr<sim::SimulationInterface>& r_, const std::vector<std::shared_ptr<Keypoint>>& kps, const VectorXd& RtDiag, int horizon, int nbDeriv, double dt)
        : System(r_, kps, RtDiag, horizon, nbDeriv, {"POS_ORN"}) {
        localInit(dt);
    }
    // PosOrnPlannerSys.cpp:114-138; the integration itself is SimulationInterface::sendVel / sendAcc
    std::tuple<VectorXd, VectorXd, MatrixXd, MatrixXd, MatrixXd> forwardPass(const VectorXd&, const VectorXd& uk, int) override {
        MatrixXd A, B;
        if (nb_deriv_ == 1) r->sendVel(dt_, uk);
        else r->sendAcc(dt_, uk);
        linearize(getState(), uk, A, B);
        auto fj = getFxJac();
        return std::make_tuple(getState(), std::get<0>(fj), A, B, std::get<1>(fj));
    }
    void linearize(const VectorXd&, const VectorXd&, MatrixXd& A, MatrixXd& B) const override {
        detail_sys::linearize_fixed_dt(r->getDOF(), nb_deriv_, nb_state_var_, nb_ctrl_var_, dt_, A, B);
    }
    using System::getFxJac;
    std::tuple<VectorXd, MatrixXd> getFxJac() override { return detail_sys::fx_jac(*r, nb_deriv_, false); }
    VectorXd getState() override { return detail_sys::joint_state(*r, nb_deriv_, false); }
    void reset() override { r->setConfiguration(q0_, dq0_); }
    double getDt() const { return dt_; }

    void flatten(detail::Flat& f) const override {
        f.sys_kind = ILQRB_SYS_POS_ORN;
        f.nb_deriv = nb_deriv_;
        f.horizon = horizon_;
        f.dt = dt_;
        f.dof = r->getDOF();
        f.tasks.emplace_back();
        flattenTask(f.tasks.back(), false);
        f.r_diag = f.tasks.back().r_diag;
        f.q0 = detail::to_std(q0_);
        f.dq0 = detail::to_std(dq0_);
    }

private:
    // PosOrnPlannerSys.cpp:54-78
    void localInit(double dt) {
        dt_ = dt;
        q0_ = r->getJointsPos();
        dq0_ = r->getJointsVel();
        x0_ = detail_sys::joint_state(*r, nb_deriv_, false);
        nb_state_var_ = int(x0_.rows());
        nb_ctrl_var_ = r->getDOF();
        nb_target_var_ = 7 * nb_deriv_;
        nb_Q_var_ = nb_target_var_ - nb_deriv_;
    }
    double dt_ = 0;
    VectorXd q0_, dq0_;
};

class PosOrnTimePlannerSys : public System {
public:
    PosOrnTimePlannerSys(const std::shared_ptr<sim::SimulationInterface>& r_, const std::vector<std::shared_ptr<Keypoint>>& kps, const VectorXd& RtDiag, const VectorXd& qMax,
                         const VectorXd& qMin, const VectorXd& dqMax, const VectorXd& dqMin, int horizon, int nbDeriv)
        : System(r_, kps, RtDiag, qMax, qMin, dqMax, dqMin, horizon, nbDeriv, {"POS_ORN_TIME"}) {
        localInit();
    }
    PosOrnTimePlannerSys(const std::shared_ptr<sim::SimulationInterface>& r_, const std::vector<std::shared_ptr<Keypoint>>& kps, const VectorXd& RtDiag, const VectorXd& qMax,
                         const VectorXd& qMin, int horizon, int nbDeriv)
        : System(r_, kps, RtDiag, qMax, qMin, horizon, nbDeriv, {"POS_ORN_TIME"}) {
        localInit();
    }
    PosOrnTimePlannerSys(const std::shared_ptr<sim::SimulationInterface>& r_, const std::vector<std::shared_ptr<Keypoint>>& kps, const VectorXd& RtDiag, int horizon, int nbDeriv)
        : System(r_, kps, RtDiag, horizon, nbDeriv, {"POS_ORN_TIME"}) {
        localInit();
    }
    // PosOrnTimePlannerSys.cpp:149-184: dt = s^2, the last control column of B is d x+/d s
    std::tuple<VectorXd, VectorXd, MatrixXd, MatrixXd, MatrixXd> forwardPass(const VectorXd&, const VectorXd& uk, int) override {
        const int mc = nb_ctrl_var_ - 1;
        const double s = uk(mc), dt = s * s;
        if (nb_deriv_ == 1) r->sendVel(dt, uk.head(mc));
        else r->sendAcc(dt, uk.head(mc));
        MatrixXd A, B;
        linearize(getState(), uk, A, B);
        auto fj = getFxJac();
        return std::make_tuple(getState(), std::get<0>(fj), A, B, std::get<1>(fj));
    }
    void linearize(const VectorXd& x_after, const VectorXd& uk, MatrixXd& A, MatrixXd& B) const override {
        detail_sys::linearize_time(r->getDOF(), nb_deriv_, nb_state_var_, nb_ctrl_var_, x_after, uk, A, B);
    }
    std::tuple<VectorXd, MatrixXd> getFxJac() override { return detail_sys::fx_jac(*r, nb_deriv_, true); }
    using System::getFxJac;  // getFxJac(xk), PosOrnTimePlannerSys.cpp:114-137: on the device (system_hooks.hpp)
    VectorXd getState() override { return detail_sys::joint_state(*r, nb_deriv_, true); }
    void reset() override { r->setConfiguration(q0_, dq0_); }

    void flatten(detail::Flat& f) const override {
        f.sys_kind = ILQRB_SYS_POS_ORN_TIME;
        f.nb_deriv = nb_deriv_;
        f.horizon = horizon_;
        f.dt = 0;
        f.dof = r->getDOF();
        f.tasks.emplace_back();
        flattenTask(f.tasks.back(), true);
        f.r_diag = f.tasks.back().r_diag;
        f.q0 = detail::to_std(q0_);
        f.dq0 = detail::to_std(dq0_);
    }

private:
    // PosOrnTimePlannerSys.cpp:50-83
    void localInit() {
        q0_ = r->getJointsPos();
        dq0_ = r->getJointsVel();
        const int d = r->getDOF();
        x0_ = VectorXd::Zero(nb_deriv_ * d + 1);
        for (int i = 0; i < d; i++) {
            x0_(i) = q0_(i);
            if (nb_deriv_ == 2) x0_(d + i) = dq0_(i);
        }
        nb_state_var_ = int(x0_.rows());
        nb_ctrl_var_ = d + 1;
        nb_target_var_ = 7 * nb_deriv_ + 1;
        nb_Q_var_ = nb_target_var_ - nb_deriv_;
    }
    VectorXd q0_, dq0_;
};

// JointSpacePlannerSys.h:17-66, JointSpacePlannerSys.cpp:13-111: the task is the joint state itself (f(x) = x, J = Identity), keypoints
// are AngularKeypoints; the integrator is PosOrnPlannerSys's
class JointSpacePlannerSys : public System {
public:
    JointSpacePlannerSys(const std::shared_ptr<sim::SimulationInterface>& r_, const std::vector<std::shared_ptr<Keypoint>>& kps, const VectorXd& RtDiag, const VectorXd& qMax,
                         const VectorXd& qMin, const VectorXd& dqMax, const VectorXd& dqMin, int horizon, int nbDeriv, double dt)
        : System(r_, kps, RtDiag, qMax, qMin, dqMax, dqMin, horizon, nbDeriv, {"JNT"}) {
        localInit(dt);
    }
    JointSpacePlannerSys(const std::shared_ptr<sim::SimulationInterface>& r_, const std::vector<std::shared_ptr<Keypoint>>& kps, const VectorXd& RtDiag, const VectorXd& qMax,
                         const VectorXd& qMin, int horizon, int nbDeriv, double dt)
        : System(r_, kps, RtDiag, qMax, qMin, horizon, nbDeriv, {"JNT"}) {
        localInit(dt);
    }
    JointSpacePlannerSys(const std::shared_ptr<sim::SimulationInterface>& r_, const std::vector<std::shared_ptr<Keypoint>>& kps, const VectorXd& RtDiag, int horizon, int nbDeriv,
                         double dt)
        : System(r_, kps, RtDiag, horizon, nbDeriv, {"JNT"}) {
        localInit(dt);
    }
    // JointSpacePlannerSys.cpp:93-111
    std::tuple<VectorXd, VectorXd, MatrixXd, MatrixXd, MatrixXd> forwardPass(const VectorXd&, const VectorXd& uk, int) override {
        MatrixXd A, B;
        if (nb_deriv_ == 1) r->sendVel(dt_, uk);
        else r->sendAcc(dt_, uk);
        linearize(getState(), uk, A, B);
        auto fj = getFxJac();
        return std::make_tuple(getState(), std::get<0>(fj), A, B, std::get<1>(fj));
    }
    void linearize(const VectorXd&, const VectorXd&, MatrixXd& A, MatrixXd& B) const override {
        detail_sys::linearize_fixed_dt(r->getDOF(), nb_deriv_, nb_state_var_, nb_ctrl_var_, dt_, A, B);
    }
    using System::getFxJac;
    // JointSpacePlannerSys.cpp:77-81 (Identity(nb_Q_var, nb_ctrl_var): only nb_deriv = 1 has a shape the cost can use)
    std::tuple<VectorXd, MatrixXd> getFxJac() override { return std::make_tuple(getState(), MatrixXd::Identity(nb_Q_var_, nb_ctrl_var_)); }
    VectorXd getState() override { return detail_sys::joint_state(*r, nb_deriv_, false); }
    void reset() override { r->setConfiguration(q0_, dq0_); }
    double getDt() const { return dt_; }
    void flatten(detail::Flat& f) const override {
        f.sys_kind = ILQRB_SYS_POS_ORN;
        f.nb_deriv = nb_deriv_;
        f.horizon = horizon_;
        f.dt = dt_;
        f.dof = r->getDOF();
        f.tasks.emplace_back();
        flattenTask(f.tasks.back(), false, ILQRB_TASK_JOINT);
        f.r_diag = f.tasks.back().r_diag;
        f.q0 = detail::to_std(q0_);
        f.dq0 = detail::to_std(dq0_);
    }

private:
    // JointSpacePlannerSys.cpp:54-75
    void localInit(double dt) {
        dt_ = dt;
        q0_ = r->getJointsPos();
        dq0_ = r->getJointsVel();
        x0_ = detail_sys::joint_state(*r, nb_deriv_, false);
        nb_state_var_ = int(x0_.rows());
        nb_ctrl_var_ = r->getDOF();
        nb_target_var_ = nb_state_var_;
        nb_Q_var_ = nb_target_var_;
    }
    double dt_ = 0;
    VectorXd q0_, dq0_;
};

// JointSpaceTimePlannerSys.h, JointSpaceTimePlannerSys.cpp:13-150: joint-space task with the time state; keypoints are AngularTimeKeypoints
class JointSpaceTimePlannerSys : public System {
public:
    JointSpaceTimePlannerSys(const std::shared_ptr<sim::SimulationInterface>& r_, const std::vector<std::shared_ptr<Keypoint>>& kps, const VectorXd& RtDiag, const VectorXd& qMax,
                             const VectorXd& qMin, const VectorXd& dqMax, const VectorXd& dqMin, int horizon, int nbDeriv)
        : System(r_, kps, RtDiag, qMax, qMin, dqMax, dqMin, horizon, nbDeriv, {"JNT_TIME"}) {
        localInit();
    }
    JointSpaceTimePlannerSys(const std::shared_ptr<sim::SimulationInterface>& r_, const std::vector<std::shared_ptr<Keypoint>>& kps, const VectorXd& RtDiag, const VectorXd& qMax,
                             const VectorXd& qMin, int horizon, int nbDeriv)
        : System(r_, kps, RtDiag, qMax, qMin, horizon, nbDeriv, {"JNT_TIME"}) {
        localInit();
    }
    JointSpaceTimePlannerSys(const std::shared_ptr<sim::SimulationInterface>& r_, const std::vector<std::shared_ptr<Keypoint>>& kps, const VectorXd& RtDiag, int horizon, int nbDeriv)
        : System(r_, kps, RtDiag, horizon, nbDeriv, {"JNT_TIME"}) {
        localInit();
    }
    // JointSpaceTimePlannerSys.cpp:116-150
    std::tuple<VectorXd, VectorXd, MatrixXd, MatrixXd, MatrixXd> forwardPass(const VectorXd&, const VectorXd& uk, int) override {
        const int mc = nb_ctrl_var_ - 1;
        const double s = uk(mc), dt = s * s;
        if (nb_deriv_ == 1) r->sendVel(dt, uk.head(mc));
        else r->sendAcc(dt, uk.head(mc));
        MatrixXd A, B;
        linearize(getState(), uk, A, B);
        auto fj = getFxJac();
        return std::make_tuple(getState(), std::get<0>(fj), A, B, std::get<1>(fj));
    }
    void linearize(const VectorXd& x_after, const VectorXd& uk, MatrixXd& A, MatrixXd& B) const override {
        detail_sys::linearize_time(r->getDOF(), nb_deriv_, nb_state_var_, nb_ctrl_var_, x_after, uk, A, B);
    }
    using System::getFxJac;
    // JointSpaceTimePlannerSys.cpp:73-77
    std::tuple<VectorXd, MatrixXd> getFxJac() override { return std::make_tuple(getState(), MatrixXd::Identity(nb_target_var_, nb_ctrl_var_)); }
    VectorXd getState() override { return detail_sys::joint_state(*r, nb_deriv_, true); }
    void reset() override { r->setConfiguration(q0_, dq0_); }
    void flatten(detail::Flat& f) const override {
        f.sys_kind = ILQRB_SYS_POS_ORN_TIME;
        f.nb_deriv = nb_deriv_;
        f.horizon = horizon_;
        f.dt = 0;
        f.dof = r->getDOF();
        f.tasks.emplace_back();
        flattenTask(f.tasks.back(), true, ILQRB_TASK_JOINT);
        f.r_diag = f.tasks.back().r_diag;
        f.q0 = detail::to_std(q0_);
        f.dq0 = detail::to_std(dq0_);
    }

private:
    // JointSpaceTimePlannerSys.cpp:41-59
    void localInit() {
        q0_ = r->getJointsPos();
        dq0_ = r->getJointsVel();
        const int d = r->getDOF();
        x0_ = VectorXd::Zero(nb_deriv_ * d + 1);
        for (int i = 0; i < d; i++) {
            x0_(i) = q0_(i);
            if (nb_deriv_ == 2) x0_(d + i) = dq0_(i);
        }
        nb_state_var_ = int(x0_.rows());
        nb_ctrl_var_ = d + 1;
        nb_target_var_ = nb_state_var_;
        nb_Q_var_ = nb_target_var_;
    }
    VectorXd q0_, dq0_;
};

// SequentialSystem.cpp:13-76: several planners on one joint state (end-effector and joint-space sub-systems may mix); tasks stack, costs add
class SequentialSystem : public System {
public:
    SequentialSystem(const std::shared_ptr<sim::SimulationInterface>& r_, const std::vector<std::shared_ptr<System>>& systems, const VectorXd& RtDiag, int, int) : systems_(systems) {
        r = r_;
        penalty_ = 0;
        limits_set_ = false;
        R = detail::diag(RtDiag);
        auto& s0 = systems_.at(0);
        nb_state_var_ = s0->getNbStateVar();
        nb_ctrl_var_ = s0->getNbCtrlVar();
        horizon_ = s0->getHorizon();
        nb_deriv_ = s0->getNbDeriv();
        x0_ = s0->getInitState();
        nb_target_var_ = 0;
        nb_Q_var_ = 0;
        for (auto& s : systems_) {
            nb_target_var_ += s->getNbTargetVar();
            nb_Q_var_ += s->getNbQVar();
            if (nb_state_var_ != s->getNbStateVar()) throw std::runtime_error(" All the systems does not have the same number of state variable ");
            if (nb_ctrl_var_ != s->getNbCtrlVar()) throw std::runtime_error(" All the systems does not have the same number of control variable ");
            if (horizon_ != s->getHorizon()) throw std::runtime_error(" All the systems does not have the same horizon ");
            if (nb_deriv_ != s->getNbDeriv()) throw std::runtime_error(" All the systems does not have the same number of derivatives ");
            if (x0_ != s->getInitState()) throw std::runtime_error(" All the systems does not have the same initState ");
        }
        q0_ = r->getJointsPos();
        dq0_ = r->getJointsVel();
        for (auto& s : systems_) keypoints.insert(keypoints.end(), s->keypoints.begin(), s->keypoints.end());
        init();
    }
    // SequentialSystem.cpp:78-91: the first sub-system integrates, every sub-system contributes its task rows
    std::tuple<VectorXd, VectorXd, MatrixXd, MatrixXd, MatrixXd> forwardPass(const VectorXd& xk, const VectorXd& uk, int k) override {
        auto first = systems_.at(0)->forwardPass(xk, uk, k);
        auto fj = getFxJac();
        return std::make_tuple(std::get<0>(first), std::get<0>(fj), std::get<2>(first), std::get<3>(first), std::get<1>(fj));
    }
    void linearize(const VectorXd& x_after, const VectorXd& uk, MatrixXd& A, MatrixXd& B) const override { systems_.at(0)->linearize(x_after, uk, A, B); }
    using System::getFxJac;
    // SequentialSystem.cpp:93-113
    std::tuple<VectorXd, MatrixXd> getFxJac() override {
        VectorXd fx = VectorXd::Zero(nb_target_var_);
        MatrixXd J = MatrixXd::Zero(nb_Q_var_, nb_state_var_);
        long o = 0, ro = 0;
        for (auto& s : systems_) {
            auto fj = s->getFxJac();
            const VectorXd& f = std::get<0>(fj);
            const MatrixXd& Js = std::get<1>(fj);
            for (long i = 0; i < f.size(); i++) fx(o + i) = f(i);
            for (long i = 0; i < Js.rows(); i++)
                for (long j = 0; j < Js.cols(); j++) J(ro + i, j) = Js(i, j);
            o += f.size();
            ro += Js.rows();
        }
        return std::make_tuple(fx, J);
    }
    VectorXd getState() override { return systems_.at(0)->getState(); }
    void reset() override { r->setConfiguration(q0_, dq0_); }
    // SequentialSystem.cpp:186-272: one slot per timestep (dense) or per keypoint of the merged list (sparse); inside a slot the
    // sub-systems' blocks follow each other; a sub-system without a keypoint at that timestep leaves zeros
    VectorXd getMuVector(bool sparse = false) override {
        const long slots = sparse ? long(keypoints.size()) : long(horizon_);
        VectorXd mu = VectorXd::Zero(slots * nb_target_var_);
        for (long sl = 0; sl < slots; sl++) {
            const int k = sparse ? keypoints[size_t(sl)]->getTimestep() : int(sl);
            long o = sl * nb_target_var_;
            for (auto& s : systems_) {
                auto kp = s->getKeypoint(k);
                if (kp) {
                    const VectorXd st = kp->getState();
                    for (long i = 0; i < st.size(); i++) mu(o + i) = st(i);
                }
                o += s->getNbTargetVar();
            }
        }
        return mu;
    }
    MatrixXd getQMatrix(bool sparse = false) override {
        const long slots = sparse ? long(keypoints.size()) : long(horizon_);
        MatrixXd Q = MatrixXd::Zero(slots * nb_Q_var_, slots * nb_Q_var_);
        for (long sl = 0; sl < slots; sl++) {
            const int k = sparse ? keypoints[size_t(sl)]->getTimestep() : int(sl);
            long o = sl * nb_Q_var_;
            for (auto& s : systems_) {
                auto kp = s->getKeypoint(k);
                if (kp) {
                    const MatrixXd P = kp->getPrecision();
                    for (long i = 0; i < P.rows(); i++)
                        for (long j = 0; j < P.cols(); j++) Q(o + i, o + j) = P(i, j);
                }
                o += s->getNbQVar();
            }
        }
        return Q;
    }

    void flatten(detail::Flat& f) const override {
        const void* base = nullptr;
        for (size_t i = 0; i < systems_.size(); i++) {
            detail::Flat sub;
            systems_[i]->flatten(sub);
            if (sub.tasks.size() != 1) throw std::runtime_error("[SequentialSystem] nested sequential systems are not supported");
            if (i == 0) {
                f.sys_kind = sub.sys_kind;
                f.nb_deriv = sub.nb_deriv;
                f.horizon = sub.horizon;
                f.dt = sub.dt;
                f.dof = sub.dof;
                base = sub.tasks[0].view.base_id;
            } else if (sub.sys_kind != f.sys_kind || sub.dt != f.dt || sub.tasks[0].view.base_id != base) {
                throw std::runtime_error("[SequentialSystem] sub-systems must be of one kind, share dt and view the same robot");
            }
            f.tasks.push_back(sub.tasks[0]);
        }
        f.r_diag.clear();
        for (long i = 0; i < R.rows(); i++) f.r_diag.push_back(R(i, i));
        f.q0 = detail::to_std(q0_);
        f.dq0 = detail::to_std(dq0_);
    }

private:
    std::vector<std::shared_ptr<System>> systems_;
    VectorXd q0_, dq0_;
};

}  // namespace sys
}  // namespace ilqr_planner
