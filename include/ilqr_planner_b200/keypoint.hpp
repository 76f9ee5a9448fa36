// keypoint.hpp — host mirror of sys::Keypoint, sys::PosOrnKeypoint, sys::SpacetimeKeypoint
// (Keypoint.h:15-43, PosOrnKeypoint.h:15-58, SpacetimeKeypoint.h:14-44).  Value objects: target, precision, timestep.
// The residual (Keypoint::diff: PosOrnKeypoint.cpp:24-45, SpacetimeKeypoint.cpp:19-24) is evaluated on the device
// (csrc/kin.cuh task_residual); `writeRecord` is how a keypoint enters the descriptor.
#pragma once
#include <memory>
#include <string>

#include "../ilqrb.h"
#include "dense.hpp"

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
    // descriptor side: the ILQRB_KP_STRIDE record pos[3] quat[4] dpos[3] dquat[4] time
    virtual void writeRecord(double* rec) const = 0;

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

}  // namespace sys
}  // namespace ilqr_planner
