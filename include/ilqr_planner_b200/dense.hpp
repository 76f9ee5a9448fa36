// dense.hpp — the vector / matrix value types of the C++ host API (include/ilqr_planner_b200/*.hpp).
//
// The reference passes Eigen::VectorXd / Eigen::MatrixXd through every public signature (System.h, ILQRRecursive.h ...).
// With -DILQRB_WITH_EIGEN (and Eigen3 on the include path) the host API uses exactly those types.  Eigen is not part of
// this image, so the default is the two small value types below: same names, same column-major storage, and the handful
// of members the reference's callers use (size/rows/cols, operator(), data, Zero/Ones/Identity/Constant, asDiagonal-like
// helpers live in detail::).  They are containers only — no arithmetic happens on the host side of the solve path.
#pragma once
#include <cstddef>
#include <initializer_list>
#include <stdexcept>
#include <vector>

#ifdef ILQRB_WITH_EIGEN
#include <eigen3/Eigen/Dense>
namespace ilqr_planner {
using VectorXd = Eigen::VectorXd;
using VectorXi = Eigen::VectorXi;
using MatrixXd = Eigen::MatrixXd;
}  // namespace ilqr_planner
#else
namespace ilqr_planner {

class VectorXd {
public:
    VectorXd() {}
    explicit VectorXd(long n) : v_(size_t(n), 0.0) {}
    VectorXd(std::initializer_list<double> l) : v_(l) {}
    explicit VectorXd(const std::vector<double>& v) : v_(v) {}
    static VectorXd Zero(long n) { return VectorXd(n); }
    static VectorXd Ones(long n) { return Constant(n, 1.0); }
    static VectorXd Constant(long n, double c) {
        VectorXd r(n);
        for (auto& x : r.v_) x = c;
        return r;
    }
    long size() const { return long(v_.size()); }
    long rows() const { return long(v_.size()); }
    long cols() const { return 1; }
    double& operator()(long i) { return v_[size_t(i)]; }
    double operator()(long i) const { return v_[size_t(i)]; }
    double& operator[](long i) { return v_[size_t(i)]; }
    double operator[](long i) const { return v_[size_t(i)]; }
    double* data() { return v_.data(); }
    const double* data() const { return v_.data(); }
    bool operator==(const VectorXd& o) const { return v_ == o.v_; }
    bool operator!=(const VectorXd& o) const { return !(v_ == o.v_); }
    VectorXd segment(long start, long n) const {
        VectorXd r(n);
        for (long i = 0; i < n; i++) r(i) = v_[size_t(start + i)];
        return r;
    }
    VectorXd head(long n) const { return segment(0, n); }
    VectorXd tail(long n) const { return segment(size() - n, n); }

private:
    std::vector<double> v_;
};

// column-major like Eigen's default, so data() can be handed to code written against Eigen
class MatrixXd {
public:
    MatrixXd() : r_(0), c_(0) {}
    MatrixXd(long r, long c) : r_(r), c_(c), v_(size_t(r * c), 0.0) {}
    static MatrixXd Zero(long r, long c) { return MatrixXd(r, c); }
    static MatrixXd Identity(long r, long c) {
        MatrixXd m(r, c);
        for (long i = 0; i < (r < c ? r : c); i++) m(i, i) = 1.0;
        return m;
    }
    long rows() const { return r_; }
    long cols() const { return c_; }
    long size() const { return r_ * c_; }
    double& operator()(long i, long j) { return v_[size_t(j * r_ + i)]; }
    double operator()(long i, long j) const { return v_[size_t(j * r_ + i)]; }
    double* data() { return v_.data(); }
    const double* data() const { return v_.data(); }
    VectorXd col(long j) const {
        VectorXd r(r_);
        for (long i = 0; i < r_; i++) r(i) = (*this)(i, j);
        return r;
    }
    MatrixXd topRows(long n) const { return block(0, 0, n, c_); }
    MatrixXd bottomRows(long n) const { return block(r_ - n, 0, n, c_); }
    MatrixXd block(long i0, long j0, long nr, long nc) const {
        MatrixXd m(nr, nc);
        for (long j = 0; j < nc; j++)
            for (long i = 0; i < nr; i++) m(i, j) = (*this)(i0 + i, j0 + j);
        return m;
    }

private:
    long r_, c_;
    std::vector<double> v_;
};

}  // namespace ilqr_planner
#endif

namespace ilqr_planner {
namespace detail {

// row-major [r][c] buffer -> MatrixXd
inline MatrixXd from_row_major(const double* p, long r, long c) {
    MatrixXd m = MatrixXd::Zero(r, c);
    for (long i = 0; i < r; i++)
        for (long j = 0; j < c; j++) m(i, j) = p[i * c + j];
    return m;
}
inline std::vector<double> to_row_major(const MatrixXd& m) {
    std::vector<double> o(size_t(m.rows() * m.cols()));
    for (long i = 0; i < m.rows(); i++)
        for (long j = 0; j < m.cols(); j++) o[size_t(i * m.cols() + j)] = m(i, j);
    return o;
}
inline VectorXd from_buffer(const double* p, long n) {
    VectorXd v = VectorXd::Zero(n);
    for (long i = 0; i < n; i++) v(i) = p[i];
    return v;
}
inline std::vector<double> to_std(const VectorXd& v) {
    std::vector<double> o(size_t(v.size()));
    for (long i = 0; i < v.size(); i++) o[size_t(i)] = v(i);
    return o;
}
inline MatrixXd diag(const VectorXd& d) {
    MatrixXd m = MatrixXd::Zero(d.size(), d.size());
    for (long i = 0; i < d.size(); i++) m(i, i) = d(i);
    return m;
}

}  // namespace detail
}  // namespace ilqr_planner
