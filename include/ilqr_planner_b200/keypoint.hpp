// keypoint.hpp — host mirror of sys::Keypoint, sys::PosOrnKeypoint, sys::SpacetimeKeypoint, sys::PosOrnKeypointDistFunct,
// sys::AngularKeypoint, sys::AngularTimeKeypoint (Keypoint.h:15-43, PosOrnKeypoint.h:15-58, SpacetimeKeypoint.h:14-44,
// PosOrnKeypointDistFunct.h:15-44, AngularKeypoint.h:15-41, AngularTimeKeypoint.h:14-36).  Value objects: target, precision,
// timestep.  The residual (Keypoint::diff) is evaluated on the device (csrc/kin.cuh task_residual): by the solvers for whole
// batches, and by `diff` below for one task vector through ilqrb_keypoint_diff.  `writeRecord` is how a keypoint enters the
// descriptor.
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "../ilqrb.h"
#include "dense.hpp"
#include "sim.hpp"  // detail::check

namespace ilqr_planner {
namespace sys {

class Keypoint {
public:
    enum KpType { FIRST_ORDER = 1, SECOND_ORDER = 2 };
    Keypoint(const int& timestep, const KpType& type, const std::string& TAG) : TAG_(TAG), type_(type), timestep_(timestep) {}
    virtual ~Keypoint() {}
    virtual VectorXd getState() const = 0;
    virtual MatrixXd getPrecision() const = 0;
    int getTimestep() const { return timestep_; }
    std::string getTAG() const { return TAG_; }
    KpType getType() { return type_; }
    virtual bool isPartOfEuclideanSpace() const = 0;
    // descriptor side: the ILQRB_KP_STRIDE record pos[3] quat[4] dpos[3] dquat[4] time (joint-space keypoints: q*[7] dq*[7] time)
    virtual void writeRecord(double* rec) const = 0;
    virtual int taskKind() const { return ILQRB_TASK_POSE; }
    virtual bool hasTime() const { return false; }
    virtual int jointCount() const { return 0; }  // joints of a joint-space keypoint
    // [pos_radius, orn_thresh[3]] of a PosOrnKeypointDistFunct; false = no dead zone
    virtual bool deadzone(double*) const { return false; }
    // Keypoint::diff (PosOrnKeypoint.cpp:24-45, SpacetimeKeypoint.cpp:19-24, PosOrnKeypointDistFunct.cpp:13-35, AngularKeypoint.cpp:22-25,
    // AngularTimeKeypoint.cpp:19-24), one task vector, on the device
    virtual VectorXd diff(const VectorXd& state) const {
        double rec[ILQRB_KP_STRIDE], dz[4];
        writeRecord(rec);
        const bool has_dz = deadzone(dz);
        const int nd = int(type_), tt = hasTime() ? 1 : 0, joint = taskKind() == ILQRB_TASK_JOINT;
        const int nt = joint ? nd * jointCount() + tt : 7 * nd + tt, r = joint ? nt : 6 * nd + tt;
        if (state.size() != nt) throw std::runtime_error("[Keypoint] diff: state of size " + std::to_string(nt) + " expected");
        std::vector<double> fx = detail::to_std(state), out(size_t(r), 0.0);
        detail::check(ilqrb_keypoint_diff(0, tt ? ILQRB_SYS_POS_ORN_TIME : ILQRB_SYS_POS_ORN, taskKind(), nd, jointCount(), rec, has_dz ? dz : nullptr, 1, fx.data(), out.data()));
        return detail::from_buffer(out.data(), r);
    }

protected:
    std::string TAG_;
    KpType type_;
    int timestep_;
};

class PosOrnKeypoint : public Keypoint {
public:
    PosOrnKeypoint(const VectorXd& position, const VectorXd& orientation, const MatrixXd& precision, const int& timestep)
        : Keypoint(timestep, KpType::FIRST_ORDER, "POS_ORN"), position_(position), orientation_(orientation), precision_(precision), state_size_(7) {}
    PosOrnKeypoint(const VectorXd& position, const VectorXd& dposition, const VectorXd& orientation, const VectorXd& dorientation, const MatrixXd& precision, const int& timestep)
        : Keypoint(timestep, KpType::SECOND_ORDER, "POS_ORN"),
          position_(position),
          orientation_(orientation),
          dposition_(dposition),
          dorientation_(dorientation),
          precision_(precision),
          state_size_(14) {}
    VectorXd getPosition() const { return position_; }
    VectorXd getOrientation() const { return orientation_; }
    MatrixXd getPrecision() const override { return precision_; }
    // PosOrnKeypoint.cpp:13-22: [p, quat] or [p, dp, quat, dquat]
    VectorXd getState() const override {
        VectorXd s = VectorXd::Zero(state_size_);
        long o = 0;
        auto put = [&](const VectorXd& v) {
            for (long i = 0; i < v.size(); i++) s(o++) = v(i);
        };
        put(position_);
        if (type_ == KpType::SECOND_ORDER) put(dposition_);
        put(orientation_);
        if (type_ == KpType::SECOND_ORDER) put(dorientation_);
        return s;
    }
    bool isPartOfEuclideanSpace() const override { return false; }
    void writeRecord(double* rec) const override {
        for (int i = 0; i < ILQRB_KP_STRIDE; i++) rec[i] = 0.0;
        for (int i = 0; i < 3 && i < position_.size(); i++) rec[i] = position_(i);
        for (int i = 0; i < 4 && i < orientation_.size(); i++) rec[3 + i] = orientation_(i);
        if (type_ == KpType::SECOND_ORDER) {
            for (int i = 0; i < 3 && i < dposition_.size(); i++) rec[7 + i] = dposition_(i);
            for (int i = 0; i < 4 && i < dorientation_.size(); i++) rec[10 + i] = dorientation_(i);
        }
    }

protected:
    VectorXd position_, orientation_, dposition_, dorientation_;
    MatrixXd precision_;
    int state_size_;
};

class SpacetimeKeypoint : public PosOrnKeypoint {
public:
    SpacetimeKeypoint(const VectorXd& position, const VectorXd& orientation, const MatrixXd& precision, const double& continuous_time, const int& timestep)
        : PosOrnKeypoint(position, orientation, precision, timestep), continuous_time_(continuous_time) {
        TAG_ = "POS_ORN_TIME";
    }
    SpacetimeKeypoint(const VectorXd& position, const VectorXd& dposition, const VectorXd& orientation, const VectorXd& dorientation, const MatrixXd& precision,
                      const double& continuous_time, const int& timestep)
        : PosOrnKeypoint(position, dposition, orientation, dorientation, precision, timestep), continuous_time_(continuous_time) {
        TAG_ = "POS_ORN_TIME";
    }
    double getContinuousTime() { return continuous_time_; }
    bool hasTime() const override { return true; }
    // SpacetimeKeypoint.cpp:12-17
    VectorXd getState() const override {
        VectorXd s = PosOrnKeypoint::getState();
        VectorXd a = VectorXd::Zero(s.size() + 1);
        for (long i = 0; i < s.size(); i++) a(i) = s(i);
        a(s.size()) = continuous_time_;
        return a;
    }
    void writeRecord(double* rec) const override {
        PosOrnKeypoint::writeRecord(rec);
        rec[14] = continuous_time_;
    }

protected:
    double continuous_time_;
};

// PosOrnKeypointDistFunct.h:15-44: a PosOrnKeypoint whose residual vanishes inside a ball (position) / box (orientation, tangent space)
class PosOrnKeypointDistFunct : public PosOrnKeypoint {
public:
    PosOrnKeypointDistFunct(const VectorXd& position, const VectorXd& orientation, const MatrixXd& precision, const double& pos_radius, const VectorXd& orn_thresh, const int& timestep)
        : PosOrnKeypoint(position, orientation, precision, timestep), pos_radius_(pos_radius), orn_thresh_(orn_thresh) {
        checkThresh();
    }
    PosOrnKeypointDistFunct(const VectorXd& position, const VectorXd& dposition, const VectorXd& orientation, const VectorXd& dorientation, const MatrixXd& precision,
                            const double& pos_radius, const VectorXd& orn_thresh, const int& timestep)
        : PosOrnKeypoint(position, dposition, orientation, dorientation, precision, timestep), pos_radius_(pos_radius), orn_thresh_(orn_thresh) {
        checkThresh();
    }
    bool deadzone(double* dz) const override {
        dz[0] = pos_radius_;
        for (int i = 0; i < 3; i++) dz[1 + i] = orn_thresh_(i);
        return true;
    }

protected:
    void checkThresh() const {
        if (orn_thresh_.size() != 3) throw std::runtime_error("[PosOrnKeypointDistFunct] orn_thresh must have 3 entries");
        if (pos_radius_ < 0) throw std::runtime_error("[PosOrnKeypointDistFunct] pos_radius must not be negative");
    }
    double pos_radius_;
    VectorXd orn_thresh_;
};

// AngularKeypoint.h:15-41: a target in joint space, residual = target - state (AngularKeypoint.cpp:22-25)
class AngularKeypoint : public Keypoint {
public:
    AngularKeypoint(const VectorXd& position, const MatrixXd& precision, const int& timestep)
        : Keypoint(timestep, KpType::FIRST_ORDER, "JNT"), position_(position), precision_(precision), state_size_(int(position.size())) {}
    AngularKeypoint(const VectorXd& position, const VectorXd& dposition, const MatrixXd& precision, const int& timestep)
        : Keypoint(timestep, KpType::SECOND_ORDER, "JNT"), position_(position), dposition_(dposition), precision_(precision), state_size_(int(position.size()) * 2) {}
    VectorXd getPosition() const { return position_; }
    MatrixXd getPrecision() const override { return precision_; }
    // AngularKeypoint.cpp:13-20
    VectorXd getState() const override {
        VectorXd s = VectorXd::Zero(state_size_);
        const long d = position_.size();
        for (long i = 0; i < d; i++) s(i) = position_(i);
        if (type_ == KpType::SECOND_ORDER)
            for (long i = 0; i < d && i < dposition_.size(); i++) s(d + i) = dposition_(i);
        return s;
    }
    bool isPartOfEuclideanSpace() const override { return true; }
    int taskKind() const override { return ILQRB_TASK_JOINT; }
    int jointCount() const override { return int(position_.size()); }
    void writeRecord(double* rec) const override {
        if (position_.size() > 7) throw std::runtime_error("[AngularKeypoint] joint-space keypoint records hold 7 joints");
        for (int i = 0; i < ILQRB_KP_STRIDE; i++) rec[i] = 0.0;
        for (long i = 0; i < position_.size(); i++) rec[i] = position_(i);
        if (type_ == KpType::SECOND_ORDER)
            for (long i = 0; i < dposition_.size() && i < 7; i++) rec[7 + i] = dposition_(i);
    }

protected:
    VectorXd position_, dposition_;
    MatrixXd precision_;
    int state_size_;
};

// AngularTimeKeypoint.h:14-36
class AngularTimeKeypoint : public AngularKeypoint {
public:
    AngularTimeKeypoint(const VectorXd& position, const MatrixXd& precision, const double& continuous_time, const int& timestep)
        : AngularKeypoint(position, precision, timestep), continuous_time_(continuous_time) {
        TAG_ = "JNT_TIME";
    }
    AngularTimeKeypoint(const VectorXd& position, const VectorXd& dposition, const MatrixXd& precision, const double& continuous_time, const int& timestep)
        : AngularKeypoint(position, dposition, precision, timestep), continuous_time_(continuous_time) {
        TAG_ = "JNT_TIME";
    }
    double getContinuousTime() { return continuous_time_; }
    bool hasTime() const override { return true; }
    // AngularTimeKeypoint.cpp:12-17
    VectorXd getState() const override {
        VectorXd s = AngularKeypoint::getState();
        VectorXd a = VectorXd::Zero(s.size() + 1);
        for (long i = 0; i < s.size(); i++) a(i) = s(i);
        a(s.size()) = continuous_time_;
        return a;
    }
    void writeRecord(double* rec) const override {
        AngularKeypoint::writeRecord(rec);
        rec[14] = continuous_time_;
    }

protected:
    double continuous_time_;
};

}  // namespace sys
}  // namespace ilqr_planner
