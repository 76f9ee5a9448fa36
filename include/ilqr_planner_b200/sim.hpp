// sim.hpp — host mirror of ilqr_planner::sim (SimulationInterface.h, KDLRobot.h, 2DRobot.h, TransformedSimulationInterface.h).
//
// Same class names, constructor arguments, accessors and error messages as the reference, but the objects are
// DESCRIPTIONS: they hold the joint state and the chain, and know how to write themselves into an `ilqrb_robot`
// (include/ilqrb.h).  Kinematics is never evaluated on the host.  The single-configuration accessors the tutorials use
// (getEEPosition, getEEOrnQuat, J ...) run the device kinematics through a one-problem "probe" handle, lazily, the
// first time a value is read after the joint state changed.
#pragma once
#include <cmath>
#include <memory>
#include <string>
#include <vector>

#include "../ilqrb.h"
#include "dense.hpp"

namespace ilqr_planner {
namespace detail {

inline void check(int rc) {
    if (rc != 0) throw std::runtime_error(ilqrb_last_error());
}

// what a SimulationInterface contributes to a problem descriptor
struct RobotView {
    int kind = ILQRB_ROBOT_KDL;
    int dof = 0;
    std::vector<ilqrb_segment> segments;  // KDL
    std::vector<double> lengths;          // 2D
    bool has_transform = false;
    double T[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};  // row-major, base -> view frame
    const void* base_id = nullptr;  // identity of the underlying robot (views of one robot share it)
};

// one-problem device handle used by the single-configuration accessors
class Probe {
public:
    explicit Probe(const RobotView& v) : view_(v) {
        const int car = v.kind == ILQRB_ROBOT_2D ? 2 : 3;
        nt_ = v.kind == ILQRB_ROBOT_2D ? car : 7;
        r_ = v.kind == ILQRB_ROBOT_2D ? car : 6;
        ilqrb_robot rb{};
        rb.kind = v.kind;
        rb.dof = v.dof;
        rb.n_segments = int(view_.segments.size());
        rb.segments = view_.segments.data();
        rb.lengths = view_.lengths.empty() ? nullptr : view_.lengths.data();
        std::vector<double> tgt(ILQRB_KP_STRIDE, 0.0), prec(size_t(r_ * r_), 0.0), ones(size_t(v.dof), 1.0), q0(size_t(v.dof), 0.0);
        tgt[3] = 1.0;
        int32_t ts = 1;
        ilqrb_task tk{};
        tk.robot = 0;
        tk.has_transform = v.has_transform ? 1 : 0;
        for (int i = 0; i < 16; i++) tk.T[i] = v.T[i];
        tk.n_keypoints = 1;
        tk.kp_timestep = &ts;
        tk.kp_target = tgt.data();
        tk.kp_precision = prec.data();
        tk.r_diag = ones.data();
        ilqrb_problem_desc d{};
        d.abi_version = ILQRB_ABI_VERSION;
        d.sys_kind = v.kind == ILQRB_ROBOT_2D ? ILQRB_SYS_POSITION : ILQRB_SYS_POS_ORN;
        d.nb_deriv = 1;
        d.horizon = 2;
        d.batch = 1;
        d.dof = v.dof;
        d.n_robots = 1;
        d.n_tasks = 1;
        d.dt = 1.0;
        d.robots = &rb;
        d.tasks = &tk;
        d.r_diag = ones.data();
        d.q0 = q0.data();
        d.dq0 = q0.data();
        check(ilqrb_create(&d, 0, &h_));
    }
    ~Probe() {
        if (h_) ilqrb_destroy(h_);
    }
    Probe(const Probe&) = delete;
    Probe& operator=(const Probe&) = delete;

    // task vector [nt] and Jacobian [r][dof] (row-major) of configuration q
    void eval(const double* q, std::vector<double>& fx, std::vector<double>& J) {
        const int n = view_.dof;
        std::vector<double> zeros(size_t(n), 0.0), fX(size_t(2 * nt_)), Jb(size_t(2 * r_ * n));
        check(ilqrb_update_problems(h_, q, zeros.data(), -1, nullptr));
        check(ilqrb_rollout(h_, zeros.data(), nullptr, fX.data(), Jb.data(), nullptr));
        fx.assign(fX.begin(), fX.begin() + nt_);
        J.assign(Jb.begin(), Jb.begin() + r_ * n);
    }
    int nt() const { return nt_; }
    int r() const { return r_; }

private:
    RobotView view_;
    ilqrb_handle h_ = nullptr;
    int nt_, r_;
};

}  // namespace detail

namespace sim {

// SimulationInterface.h:15-169
class SimulationInterface {
public:
    virtual ~SimulationInterface() {}

    // KDLRobot.cpp:83-115 / 2DRobot.cpp:44-70 / TransformedSimulationInterface.cpp:31-46 — on the device, see file header
    virtual void updateKinematics() { dirty_ = true; }

    virtual MatrixXd J() {
        refresh();
        return Jac;
    }
    MatrixXd Jt() { return J().topRows(nbCarDim); }
    MatrixXd Jr() { return J().bottomRows(nbCarDim); }
    // SimulationInterface.cpp:13-17
    MatrixXd dQuatToDxJac(const VectorXd& quat) {
        MatrixXd H = MatrixXd::Zero(3, 4);
        const double w = quat(0), x = quat(1), y = quat(2), z = quat(3);
        const double v[12] = {-x, w, -z, y, -y, z, w, -x, -z, -y, x, w};
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 4; j++) H(i, j) = v[i * 4 + j];
        return H;
    }
    // SimulationInterface.cpp:19-31: q += dt dq + dt^2/2 ddq ; dq += dt ddq ; t += dt
    virtual void sendAcc(double dt, const VectorXd& ddq_, bool updateKin = true) {
        for (long i = 0; i < q.size(); i++) {
            q(i) += dt * dq(i) + dt * dt / 2 * ddq_(i);
            dq(i) += dt * ddq_(i);
        }
        t += dt;
        if (updateKin) updateKinematics();
        ddq = ddq_;
    }
    virtual void sendVel(double dt, const VectorXd& dq_, bool updateKin = true) {
        dq = dq_;
        sendAcc(dt, VectorXd::Zero(dq_.size()), updateKin);
    }
    virtual VectorXd getEEPosition() {
        refresh();
        return x;
    }
    virtual VectorXd getEEVelocity() {
        refresh();
        return dx;
    }
    virtual VectorXd getEEAngVel() {
        refresh();
        return w;
    }
    // SimulationInterface.cpp:69-73: 1/2 H(quat)^T w
    VectorXd getEEAngVelQuat() {
        VectorXd wv = getEEAngVel(), et = getEEOrnQuat();
        MatrixXd H = dQuatToDxJac(et);
        VectorXd o = VectorXd::Zero(4);
        for (int j = 0; j < 4; j++)
            for (int i = 0; i < 3 && i < wv.size(); i++) o(j) += 0.5 * H(i, j) * wv(i);
        return o;
    }
    virtual VectorXd getEEOrnQuat() {
        refresh();
        return ornQuat;
    }
    virtual VectorXd getJointsPos() { return q; }
    virtual VectorXd getJointsVel() { return dq; }
    int getDOF() { return dof; }
    int getNbCarDim() { return nbCarDim; }
    virtual double getTime() { return t; }
    virtual void setTime(double time) { t = time; }
    virtual void setConfiguration(const VectorXd& q_, const VectorXd& dq_, bool reset_time = true) {
        q = q_;
        dq = dq_;
        updateKinematics();
        if (reset_time) t = 0;
    }

    // descriptor side (not in the reference): what this object contributes to ilqrb_problem_desc
    virtual void describe(detail::RobotView& v) const = 0;

protected:
    void refresh() {
        if (!dirty_) return;
        detail::RobotView v;
        describe(v);
        if (!probe_) probe_ = std::make_shared<detail::Probe>(v);
        VectorXd qq = getJointsPos(), dqq = getJointsVel();
        std::vector<double> fx, Jb;
        probe_->eval(qq.data(), fx, Jb);
        const int car = nbCarDim, r = probe_->r();
        x = detail::from_buffer(fx.data(), car);
        ornQuat = VectorXd::Zero(4);
        ornQuat(0) = 1;
        if (v.kind == ILQRB_ROBOT_KDL)
            for (int i = 0; i < 4; i++) ornQuat(i) = fx[size_t(3 + i)];
        Jac = MatrixXd::Zero(2 * car, dof);  // 2DRobot.cpp:58-61 pads [Jt; 0]
        for (int i = 0; i < r; i++)
            for (int j = 0; j < dof; j++) Jac(i, j) = Jb[size_t(i * dof + j)];
        dx = VectorXd::Zero(car);  // dx = Jt dq, w = Jr dq (KDLRobot.cpp:104-105)
        w = VectorXd::Zero(car);
        for (int j = 0; j < dof && j < dqq.size(); j++)
            for (int i = 0; i < car; i++) {
                dx(i) += Jac(i, j) * dqq(j);
                w(i) += Jac(car + i, j) * dqq(j);
            }
        dirty_ = false;
    }

    VectorXd q, dq, ddq;
    VectorXd x, dx;
    VectorXd ornQuat, w;
    MatrixXd Jac;
    int dof = 0, nbCarDim = 3;
    double t = 0;
    bool dirty_ = true;
    std::shared_ptr<detail::Probe> probe_;
};

// KDLRobot.h:23-64, KDLRobot.cpp:17-81
class KDLRobot : public SimulationInterface {
public:
    KDLRobot(const std::string& urdf, const std::string& baseFrame, const std::string& tipFrame, const VectorXd& q_, const VectorXd& dq_, const VectorXd& transform_rpy,
             const VectorXd& transform_xyz, const bool& is_path) {
        segments_.resize(ILQRB_MAX_SEGMENTS);
        int ns = 0, nj = 0;
        double rpy[3] = {transform_rpy(0), transform_rpy(1), transform_rpy(2)}, xyz[3] = {transform_xyz(0), transform_xyz(1), transform_xyz(2)};
        detail::check(ilqrb_chain_from_urdf(urdf.c_str(), is_path ? 1 : 0, baseFrame.c_str(), tipFrame.c_str(), rpy, xyz, segments_.data(), ILQRB_MAX_SEGMENTS, &ns, &nj));
        segments_.resize(size_t(ns));
        dof = nj;
        nbCarDim = 3;
        q = q_;
        dq = dq_;
        ddq = VectorXd::Zero(dof);
    }
    KDLRobot(const std::string& urdf, const std::string& baseFrame, const std::string& tipFrame, const VectorXd& q_, const VectorXd& dq_)
        : KDLRobot(urdf, baseFrame, tipFrame, q_, dq_, VectorXd::Zero(3), VectorXd::Zero(3), true) {}
    KDLRobot(const std::string& urdf, const std::string& baseFrame, const std::string& tipFrame, const VectorXd& q_, const VectorXd& dq_, const VectorXd& transform_rpy,
             const VectorXd& transform_xyz)
        : KDLRobot(urdf, baseFrame, tipFrame, q_, dq_, transform_rpy, transform_xyz, true) {}

    void describe(detail::RobotView& v) const override {
        v.kind = ILQRB_ROBOT_KDL;
        v.dof = dof;
        v.segments = segments_;
        v.base_id = this;
    }

private:
    std::vector<ilqrb_segment> segments_;
};

// 2DRobot.h:17-38, 2DRobot.cpp:13-26
class Robot2D : public SimulationInterface {
public:
    Robot2D(const VectorXd& lengths_, const VectorXd& default_q) : lengths(lengths_) {
        dof = int(default_q.rows());
        nbCarDim = 2;
        q = default_q;
        dq = VectorXd::Zero(dof);
        ddq = VectorXd::Zero(dof);
    }
    VectorXd fkine(const VectorXd& q_) {
        VectorXd keep = q;
        q = q_;
        dirty_ = true;
        VectorXd out = getEEPosition();
        q = keep;
        dirty_ = true;
        return out;
    }
    VectorXd fkine() { return getEEPosition(); }
    void describe(detail::RobotView& v) const override {
        v.kind = ILQRB_ROBOT_2D;
        v.dof = dof;
        v.lengths = detail::to_std(lengths);
        v.base_id = this;
    }

private:
    VectorXd lengths;
};

// TransformedSimulationInterface.h:19-62, TransformedSimulationInterface.cpp:14-122: a view of r in frame T; the joint state
// lives in r
class TransformedSimulationInterface : public SimulationInterface {
public:
    TransformedSimulationInterface(const std::shared_ptr<SimulationInterface>& r_, const MatrixXd& T_) : r(r_), T(T_), initialized(true) {
        nbCarDim = r->getNbCarDim();
        dof = r->getDOF();
    }
    explicit TransformedSimulationInterface(const MatrixXd& T_) : T(T_), initialized(false) {}
    void subscribe(const std::shared_ptr<SimulationInterface>& r_) {
        r = r_;
        nbCarDim = r->getNbCarDim();
        dof = r->getDOF();
        initialized = true;
        dirty_ = true;
    }
    void updateKinematics() override {
        if (initialized) r->updateKinematics();
        dirty_ = true;
    }
    MatrixXd J() override {
        dirty_ = true;  // the wrapped robot may have moved
        return SimulationInterface::J();
    }
    VectorXd getEEPosition() override {
        dirty_ = true;
        return SimulationInterface::getEEPosition();
    }
    VectorXd getEEVelocity() override {
        dirty_ = true;
        return SimulationInterface::getEEVelocity();
    }
    VectorXd getEEAngVel() override {
        dirty_ = true;
        return SimulationInterface::getEEAngVel();
    }
    VectorXd getEEOrnQuat() override {
        dirty_ = true;
        return SimulationInterface::getEEOrnQuat();
    }
    VectorXd getJointsPos() override { return r->getJointsPos(); }
    VectorXd getJointsVel() override { return r->getJointsVel(); }
    void sendAcc(double dt, const VectorXd& ddq_, bool updateKin = true) override { r->sendAcc(dt, ddq_, updateKin); }
    void sendVel(double dt, const VectorXd& dq_, bool updateKin = true) override { r->sendVel(dt, dq_, updateKin); }
    void setConfiguration(const VectorXd& q_, const VectorXd& dq_, bool reset_time = true) override { r->setConfiguration(q_, dq_, reset_time); }
    void setTime(double time) override { r->setTime(time); }
    double getTime() override { return r->getTime(); }

    void describe(detail::RobotView& v) const override {
        if (!initialized) throw std::runtime_error("[TransformedSimulationInterface] no robot subscribed");
        r->describe(v);
        if (v.has_transform) throw std::runtime_error("[TransformedSimulationInterface] nested transforms are not supported by the device descriptor");
        v.has_transform = true;
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++) v.T[i * 4 + j] = T(i, j);
    }

private:
    std::shared_ptr<SimulationInterface> r;
    MatrixXd T;
    bool initialized;
};

}  // namespace sim
}  // namespace ilqr_planner
