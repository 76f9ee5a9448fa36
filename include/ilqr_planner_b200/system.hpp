// system.hpp — host mirror of sys::System and its concrete planners (System.h:27-205, PosOrnPlannerSys.h, PosOrnTimePlannerSys.h,
// SequentialSystem.h).  Same constructors, size arithmetic (localInit), keypoint checks and error messages; the object is a
// description that `flatten`s into detail::Flat, from which the solvers build `ilqrb_problem_desc` (include/ilqrb.h).
// The cost model (System.cpp:213-312) and the rollout run on the device only: the single-point methods of the reference's
// debugging API (diff, diffBatch, cost*, cost_F*, getFxJac(xk), fpBatch) evaluate through a cached one-problem handle
// (ilqrb_eval_point / ilqrb_rollout); their bodies are in system_hooks.hpp, after the descriptor builder they need.
#pragma once
#include <algorithm>
#include <map>
#include <tuple>
#include <utility>

#include "keypoint.hpp"
#include "sim.hpp"

namespace ilqr_planner {
namespace detail {

struct TaskFlat {
    RobotView view;
    std::vector<int32_t> timestep;
    std::vector<double> target;     // [Kp][ILQRB_KP_STRIDE]
    std::vector<double> precision;  // [Kp][r][r]
    std::vector<double> r_diag;     // [m]
    bool limits_set = false;
    std::vector<double> state_max, state_min;  // [n]
    std::vector<int32_t> weight;               // [n]
    int task_kind = ILQRB_TASK_POSE;
    std::vector<double> deadzone;              // [Kp][4] (PosOrnKeypointDistFunct; radius -1 = plain keypoint) or empty
};
struct HookHandle;  // one-problem device handle of a System (system_hooks.hpp)

struct Flat {
    int sys_kind = ILQRB_SYS_POS_ORN, nb_deriv = 1, horizon = 0, dof = 0;
    double dt = 0;
    std::vector<TaskFlat> tasks;
    std::vector<double> r_diag, q0, dq0;
};

// Eigen's isApprox with the default precision 1e-12 (System.cpp:57)
inline bool is_approx(const VectorXd& a, const VectorXd& b) {
    double d = 0, na = 0, nb = 0;
    for (long i = 0; i < a.size(); i++) {
        d += (a(i) - b(i)) * (a(i) - b(i));
        na += a(i) * a(i);
        nb += b(i) * b(i);
    }
    return d <= 1e-24 * std::min(na, nb);
}

}  // namespace detail

namespace sys {

class System {
public:
    System() {}
    // System.cpp:29-61 (limits) ; the 5-argument form without velocity limits delegates with zeros (System.cpp:18-27)
    System(const std::shared_ptr<sim::SimulationInterface>& r_, const std::vector<std::shared_ptr<Keypoint>>& keypoints_, const VectorXd& RtDiag, const VectorXd& qMax,
           const VectorXd& qMin, const VectorXd& dqMax, const VectorXd& dqMin, const int& horizon, const int& nbDeriv, const std::vector<std::string>& allowed_kp_tags)
        : r(r_), keypoints(keypoints_), horizon_(horizon), nb_deriv_(nbDeriv), EXPECTED_KP_TAGS_(allowed_kp_tags) {
        limits_set_ = true;
        penalty_ = 1;
        R = detail::diag(RtDiag);
        init();
        const int dof = r->getDOF(), n = nb_deriv_ * dof;
        state_max_ = VectorXd::Zero(n);
        state_min_ = VectorXd::Zero(n);
        joint_limits_weight_.assign(size_t(n), 1);
        for (int i = 0; i < dof; i++) {
            state_max_(i) = qMax(i);
            state_min_(i) = qMin(i);
            if (nb_deriv_ == 2) {
                state_max_(dof + i) = dqMax(i);
                state_min_(dof + i) = dqMin(i);
            }
        }
        if (nb_deriv_ == 2 && detail::is_approx(dqMax, dqMin))
            for (int i = 0; i < dof; i++) joint_limits_weight_[size_t(dof + i)] = 0;
    }
    System(const std::shared_ptr<sim::SimulationInterface>& r_, const std::vector<std::shared_ptr<Keypoint>>& keypoints_, const VectorXd& RtDiag, const VectorXd& qMax,
           const VectorXd& qMin, const int& horizon, const int& nbDeriv, const std::vector<std::string>& allowed_kp_tags)
        : System(r_, keypoints_, RtDiag, qMax, qMin, VectorXd::Zero(r_->getDOF()), VectorXd::Zero(r_->getDOF()), horizon, nbDeriv, allowed_kp_tags) {}
    // System.cpp:63-75 (no limits)
    System(const std::shared_ptr<sim::SimulationInterface>& r_, const std::vector<std::shared_ptr<Keypoint>>& keypoints_, const VectorXd& RtDiag, const int& horizon,
           const int& nbDeriv, const std::vector<std::string>& allowed_kp_tags)
        : r(r_), keypoints(keypoints_), horizon_(horizon), nb_deriv_(nbDeriv), EXPECTED_KP_TAGS_(allowed_kp_tags) {
        limits_set_ = false;
        penalty_ = 0;
        R = detail::diag(RtDiag);
        init();
    }
    virtual ~System() {}

    virtual std::tuple<VectorXd, VectorXd, MatrixXd, MatrixXd, MatrixXd> forwardPass(const VectorXd& xk, const VectorXd& uk, int k) = 0;
    virtual VectorXd getState() = 0;
    virtual std::tuple<VectorXd, MatrixXd> getFxJac() = 0;
    virtual void reset() = 0;

    // System.cpp:321-361
    virtual VectorXd getMuVector(bool sparse = false) {
        const long nt = nb_target_var_;
        VectorXd mu = VectorXd::Zero((sparse ? long(keypoints.size()) : long(horizon_)) * nt);
        for (size_t i = 0; i < keypoints.size(); i++) {
            const VectorXd s = keypoints[i]->getState();
            const long at = (sparse ? long(i) : long(keypoints[i]->getTimestep())) * nt;
            for (long j = 0; j < nt && j < s.size(); j++) mu(at + j) = s(j);
        }
        return mu;
    }
    virtual MatrixXd getQMatrix(bool sparse = false) {
        const long nq = nb_Q_var_, N = (sparse ? long(keypoints.size()) : long(horizon_)) * nq;
        MatrixXd Q = MatrixXd::Zero(N, N);
        for (size_t i = 0; i < keypoints.size(); i++) {
            const MatrixXd P = keypoints[i]->getPrecision();
            const long at = (sparse ? long(i) : long(keypoints[i]->getTimestep())) * nq;
            for (long a = 0; a < nq && a < P.rows(); a++)
                for (long b = 0; b < nq && b < P.cols(); b++) Q(at + a, at + b) = P(a, b);
        }
        return Q;
    }
    MatrixXd getRt() { return R; }
    int getNbStateVar() { return nb_state_var_; }
    int getNbCtrlVar() { return nb_ctrl_var_; }
    int getNbTargetVar() { return nb_target_var_; }
    int getNbQVar() { return nb_Q_var_; }
    int getHorizon() { return horizon_; }
    int getNbDeriv() { return nb_deriv_; }
    std::shared_ptr<Keypoint> getKeypoint(int k) {
        auto it = keypoints_map_.find(k);
        return it == keypoints_map_.end() ? nullptr : it->second;
    }
    // System.cpp:363-374 (the reference prints the same prefix for every system class)
    void checkKeypoints() {
        for (auto& kp : keypoints) {
            if (std::find(EXPECTED_KP_TAGS_.begin(), EXPECTED_KP_TAGS_.end(), kp->getTAG()) == EXPECTED_KP_TAGS_.end())
                throw std::runtime_error("[PosOrnPlannerSys] Wrong keypoint type: got " + kp->getTAG());
            if (int(kp->getType()) != nb_deriv_)
                throw std::runtime_error("[PosOrnPlannerSys] Wrong keypoint order (nb_deriv_): Expecting " + std::to_string(nb_deriv_) + " got " + std::to_string(int(kp->getType())));
        }
    }
    virtual VectorXd getInitState() { return x0_; }
    virtual VectorXd getInitFoXState() {
        reset();
        return std::get<0>(getFxJac());
    }
    std::vector<int> getKpIndexes() {
        std::vector<int> idx;
        for (auto const& kp : keypoints) idx.push_back(kp->getTimestep());
        return idx;
    }

    // ---- single-point debugging API, evaluated on the device (system_hooks.hpp) ----
    // System.cpp:103-109 / SequentialSystem.cpp:167-184
    virtual VectorXd diff(const VectorXd& actual_state, int k);
    // System.cpp:111-119: x = the task vectors at the keypoints, stacked
    VectorXd diffBatch(const VectorXd& x);
    // System.cpp:163-179: f(x) and its Jacobian at a given state, the simulation left untouched
    virtual std::tuple<VectorXd, MatrixXd> getFxJac(VectorXd xk);
    // System.cpp:121-142
    std::pair<MatrixXd, VectorXd> inspectJointLimit(const VectorXd& x_k);
    // System.cpp:144-161
    std::tuple<VectorXd, VectorXd, VectorXd, VectorXd, MatrixXd, MatrixXd, MatrixXd, MatrixXd> forwardPassWithLimits(const VectorXd& xk, const VectorXd& uk, int k);
    // System.cpp:181-211: (fX, qL, [A, B, J, L] per timestep) of the rollout of u from the initial state
    std::tuple<VectorXd, VectorXd, std::vector<std::tuple<MatrixXd, MatrixXd, MatrixXd, MatrixXd>>> fpBatch(const VectorXd& u);
    // System.cpp:213-312
    VectorXd cost(const VectorXd& xk, const VectorXd& uk, int k);
    VectorXd cost_x(const VectorXd& xk, const VectorXd& uk, int k);
    MatrixXd cost_xx(const VectorXd& xk, const VectorXd& uk, int k);
    VectorXd cost_u(const VectorXd&, const VectorXd& uk, int) {  // R uk (System.cpp:274-276)
        VectorXd o = VectorXd::Zero(R.rows());
        for (long i = 0; i < R.rows(); i++)
            for (long j = 0; j < R.cols(); j++) o(i) += R(i, j) * uk(j);
        return o;
    }
    MatrixXd cost_uu(const VectorXd&, const VectorXd&, int) { return R; }
    MatrixXd cost_ux(const VectorXd&, const VectorXd& uk, int) { return MatrixXd::Zero(uk.rows(), nb_state_var_); }
    MatrixXd cost_xu(const VectorXd&, const VectorXd& uk, int) { return MatrixXd::Zero(nb_state_var_, uk.rows()); }
    VectorXd cost_F(const VectorXd& xk) { return cost(xk, VectorXd::Zero(nb_ctrl_var_), horizon_ - 1); }
    VectorXd cost_F_x(const VectorXd& xk) { return cost_x(xk, VectorXd::Zero(nb_ctrl_var_), horizon_ - 1); }
    MatrixXd cost_F_xx(const VectorXd& xk) { return cost_xx(xk, VectorXd::Zero(nb_ctrl_var_), horizon_ - 1); }
    // A, B of the step that took the state to x_after under uk (the time systems' B depends on both)
    virtual void linearize(const VectorXd& x_after, const VectorXd& uk, MatrixXd& A, MatrixXd& B) const = 0;

    // descriptor side
    virtual void flatten(detail::Flat& f) const = 0;

    std::shared_ptr<sim::SimulationInterface> r;
    std::vector<std::shared_ptr<Keypoint>> keypoints;

protected:
    // System.cpp:77-86
    virtual void init() {
        for (auto& kp : keypoints) keypoints_map_[kp->getTimestep()] = kp;
        std::stable_sort(keypoints.begin(), keypoints.end(), [](const std::shared_ptr<Keypoint>& a, const std::shared_ptr<Keypoint>& b) { return a->getTimestep() < b->getTimestep(); });
        if (!EXPECTED_KP_TAGS_.empty()) checkKeypoints();
    }
    // one task of the descriptor from this system's own robot view, keypoints, R and limits; `time` appends the zero-weight entry
    void flattenTask(detail::TaskFlat& t, bool time, int task_kind = ILQRB_TASK_POSE) const {
        r->describe(t.view);
        t.task_kind = task_kind;
        const int rq = nb_Q_var_;
        bool any_dz = false;
        for (auto const& kp : keypoints) {
            double dz[4] = {-1, 0, 0, 0};
            if (kp->deadzone(dz)) any_dz = true;
            t.deadzone.insert(t.deadzone.end(), dz, dz + 4);
            t.timestep.push_back(kp->getTimestep());
            t.target.resize(t.target.size() + ILQRB_KP_STRIDE);
            kp->writeRecord(&t.target[t.target.size() - ILQRB_KP_STRIDE]);
            const MatrixXd P = kp->getPrecision();
            if (P.rows() != rq || P.cols() != rq) throw std::runtime_error("[System] keypoint precision must be " + std::to_string(rq) + " x " + std::to_string(rq));
            for (int a = 0; a < rq; a++)
                for (int b = 0; b < rq; b++) t.precision.push_back(P(a, b));
        }
        if (!any_dz) t.deadzone.clear();
        for (long i = 0; i < R.rows(); i++) t.r_diag.push_back(R(i, i));
        t.limits_set = limits_set_;
        if (limits_set_) {
            t.state_max = detail::to_std(state_max_);
            t.state_min = detail::to_std(state_min_);
            t.weight = joint_limits_weight_;
            if (time && int(t.weight.size()) < nb_state_var_) {  // PosOrnTimePlannerSys.cpp:72-82
                t.state_max.push_back(0);
                t.state_min.push_back(0);
                t.weight.push_back(0);
            }
        }
    }

    struct EvalOut {
        std::vector<double> fx, J, diff, cost, cost_x, cost_xx;
    };
    // one query point through ilqrb_eval_point on the cached one-problem handle
    EvalOut evalPoint(const VectorXd& xk, const VectorXd* uk, int k, const VectorXd* fx_in);
    std::shared_ptr<detail::HookHandle> hook_;

    VectorXd x0_;
    int nb_state_var_ = 0, nb_target_var_ = 0, nb_ctrl_var_ = 0, nb_Q_var_ = 0, horizon_ = 0, nb_deriv_ = 1;
    bool limits_set_ = false;
    VectorXd state_max_, state_min_;
    std::vector<int32_t> joint_limits_weight_;
    double penalty_ = 0;
    MatrixXd R;
    std::map<int, std::shared_ptr<Keypoint>> keypoints_map_;
    std::vector<std::string> EXPECTED_KP_TAGS_;
};

}  // namespace sys
}  // namespace ilqr_planner

#include "planner_systems.hpp"
