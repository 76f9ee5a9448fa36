// sd.hpp — ilqr_planner::Sd, the sphere-manifold helpers of the reference (utils/sd.h:19-99), for S^3 (unit quaternions, the only
// sphere the planners use).  Each call evaluates on the GPU through ilqrb_sd — the same device functions the keypoint residuals
// use — so what a user inspects with these is what the solver computes.
#pragma once
#include "sim.hpp"

namespace ilqr_planner {
namespace Sd {
namespace detail_sd {
inline const double* q4(const VectorXd& v, const char* what) {
    if (v.size() != 4) throw std::runtime_error(std::string("[Sd] ") + what + ": vectors of size 4 expected (S^3)");
    return v.data();
}
inline VectorXd call(int op, const VectorXd& a, const VectorXd* b, const VectorXd* c, int out_size, const char* what) {
    std::vector<double> o(size_t(out_size), 0.0);
    detail::check(ilqrb_sd(0, op, 1, q4(a, what), b ? q4(*b, what) : nullptr, c ? q4(*c, what) : nullptr, o.data()));
    return detail::from_buffer(o.data(), out_size);
}
}  // namespace detail_sd

// sd.h:67-82
inline VectorXd logMap(const VectorXd& base, const VectorXd& y) { return detail_sd::call(ILQRB_SD_LOGMAP, base, &y, nullptr, 4, "logMap"); }
// sd.h:34-45
inline VectorXd expMap(const VectorXd& base, const VectorXd& u) { return detail_sd::call(ILQRB_SD_EXPMAP, base, &u, nullptr, 4, "expMap"); }
// sd.h:48-62
inline double distance(const VectorXd& x, const VectorXd& y) { return detail_sd::call(ILQRB_SD_DISTANCE, x, &y, nullptr, 1, "distance")(0); }
// sd.h:87-99
inline VectorXd transport(const VectorXd& v, const VectorXd& base1, const VectorXd& base2) { return detail_sd::call(ILQRB_SD_TRANSPORT, v, &base1, &base2, 4, "transport"); }
// sd.h:23-27
inline MatrixXd dQuatToDxJac(const VectorXd& q) {
    const VectorXd o = detail_sd::call(ILQRB_SD_DQUAT_TO_W_JAC, q, nullptr, nullptr, 12, "dQuatToDxJac");
    return detail::from_row_major(o.data(), 3, 4);
}

}  // namespace Sd
}  // namespace ilqr_planner
