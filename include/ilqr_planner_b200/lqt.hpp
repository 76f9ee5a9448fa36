// lqt.hpp — ilqr_planner::solver::LQT with the reference's constructor and methods (lqt.h:24-88), evaluated on the GPU through
// ilqrb_lqt_solve / ilqrb_lqt_command (include/ilqrb.h; kernels in csrc/launch_lqt.cu).  One device call produces what solveDP and
// solveLinAl compute; each method still answers only after "its" solver was called, with the reference's messages (lqt.cpp:99-103,
// 111-113, 127-131).  LQT::solveBatch is the array-level form for many problems of one shape.
#pragma once
#include <cmath>

#include "sim.hpp"

namespace ilqr_planner {
namespace solver {

class LQT {
public:
    // lqt.cpp:16-26: R = Identity * pow(rfactor, nb_deriv) with rfactor held as a float; nb_states = states.size() / A.cols()
    LQT(const MatrixXd& A, const MatrixXd& B, const std::vector<MatrixXd>& Qs, const VectorXd& states, float rfactor = 0.1f, int nb_deriv = 2)
        : n_(int(A.cols())), m_(int(B.cols())), mu_(states.data(), states.data() + states.size()) {
        if (A.rows() != A.cols() || B.rows() != A.rows()) throw std::invalid_argument("[LQT] A must be square and B must have as many rows");
        if (n_ < 1 || states.size() % n_ != 0) throw std::invalid_argument("[LQT] states must stack vectors of size A.cols()");
        T_ = int(states.size() / n_);
        if (long(Qs.size()) != T_) throw std::invalid_argument("[LQT] one weight matrix per state expected");
        r_ = std::pow(double(rfactor), nb_deriv);
        A_ = detail::to_row_major(A);
        B_ = detail::to_row_major(B);
        Qs_.reserve(size_t(T_) * n_ * n_);
        for (const MatrixXd& Q : Qs) {
            if (Q.rows() != n_ || Q.cols() != n_) throw std::invalid_argument("[LQT] weight matrices of size nb_state_var x nb_state_var expected");
            const std::vector<double> q = detail::to_row_major(Q);
            Qs_.insert(Qs_.end(), q.begin(), q.end());
        }
    }
    LQT() {}

    void solveDP() {
        run();
        dp_ = true;
    }
    void solveLinAl() {
        run();
        linal_ = true;
    }
    int getNbStates() const { return T_; }

    // lqt.cpp:98-103
    VectorXd getCommand(int timestep) {
        if (!linal_) throw std::runtime_error("solveLinal() or solveQP() first");
        if (timestep < 0 || timestep > T_ - 2) throw std::out_of_range("[LQT] timestep outside the control horizon");
        return detail::from_buffer(u_.data() + size_t(timestep) * m_, m_);
    }
    // lqt.cpp:105-124; like the reference's Ps.at(), a timestep outside [0, nb_states - 2] throws std::out_of_range
    VectorXd getCommand(int timestep, const VectorXd& curr_state) {
        if (!dp_) throw std::runtime_error("solveDP() first");
        if (timestep < 0 || timestep > T_ - 2) throw std::out_of_range("vector::_M_range_check: timestep outside [0, nb_states - 2]");
        if (curr_state.size() != n_) throw std::invalid_argument("[LQT] state of size nb_state_var expected");
        std::vector<double> u(size_t(m_), 0.0);
        const int32_t ts = timestep;
        detail::check(ilqrb_lqt_command(0, 1, n_, m_, T_, K_.data(), f_.data(), mu_.data(), 1, nullptr, &ts, curr_state.data(), u.data()));
        return detail::from_buffer(u.data(), m_);
    }
    // lqt.cpp:126-132
    VectorXd getPredictedStates() {
        if (!linal_) throw std::runtime_error("solveLinal() or solveQP() first");
        return detail::from_buffer(xhat_.data(), long(xhat_.size()));
    }

    // solveDP's matrices in TIME order (the reference keeps them private, last timestep first)
    std::vector<MatrixXd> getPs() {
        if (!dp_) throw std::runtime_error("solveDP() first");
        std::vector<MatrixXd> o;
        for (int t = 0; t < T_; t++) o.push_back(detail::from_row_major(Ps_.data() + size_t(t) * n_ * n_, n_, n_));
        return o;
    }
    std::vector<VectorXd> getds() {
        if (!dp_) throw std::runtime_error("solveDP() first");
        std::vector<VectorXd> o;
        for (int t = 0; t < T_; t++) o.push_back(detail::from_buffer(ds_.data() + size_t(t) * n_, n_));
        return o;
    }

    // Many problems of one shape in one device call.  Qs [nq][T][n][n], mu [nq][T*n] row-major; returns u [nq][(T-1)m] and xhat [nq][T*n].
    struct BatchOut {
        std::vector<double> u, xhat, Ps, ds, K, f;
        std::vector<int32_t> status;
    };
    static BatchOut solveBatch(const MatrixXd& A, const MatrixXd& B, int nq, int T, const std::vector<double>& Qs, const std::vector<double>& mu, float rfactor = 0.1f,
                               int nb_deriv = 2, int device = 0) {
        const int n = int(A.cols()), m = int(B.cols());
        if (nq < 1 || T < 2 || Qs.size() != size_t(nq) * T * n * n || mu.size() != size_t(nq) * T * n) throw std::invalid_argument("[LQT] solveBatch: Qs [nq][T][n][n] and mu [nq][T*n] expected");
        const std::vector<double> a = detail::to_row_major(A), b = detail::to_row_major(B);
        BatchOut o;
        o.u.resize(size_t(nq) * (T - 1) * m);
        o.xhat.resize(size_t(nq) * T * n);
        o.Ps.resize(size_t(nq) * T * n * n);
        o.ds.resize(size_t(nq) * T * n);
        o.K.resize(size_t(nq) * (T - 1) * m * n);
        o.f.resize(size_t(nq) * (T - 1) * m);
        o.status.resize(size_t(nq));
        ilqrb_lqt_desc d{nq, n, m, T, 0, 0, a.data(), b.data(), Qs.data(), mu.data(), std::pow(double(rfactor), nb_deriv)};
        detail::check(ilqrb_lqt_solve(device, &d, o.Ps.data(), o.ds.data(), o.K.data(), o.f.data(), o.u.data(), o.xhat.data(), o.status.data()));
        return o;
    }

private:
    void run() {
        if (solved_) return;
        if (T_ < 2) throw std::runtime_error("[LQT] at least two states expected");
        Ps_.resize(size_t(T_) * n_ * n_);
        ds_.resize(size_t(T_) * n_);
        K_.resize(size_t(T_ - 1) * m_ * n_);
        f_.resize(size_t(T_ - 1) * m_);
        u_.resize(size_t(T_ - 1) * m_);
        xhat_.resize(size_t(T_) * n_);
        int32_t st = 0;
        ilqrb_lqt_desc d{1, n_, m_, T_, 0, 0, A_.data(), B_.data(), Qs_.data(), mu_.data(), r_};
        detail::check(ilqrb_lqt_solve(0, &d, Ps_.data(), ds_.data(), K_.data(), f_.data(), u_.data(), xhat_.data(), &st));
        solved_ = true;
    }

    int n_ = 0, m_ = 0, T_ = 0;
    double r_ = 0.0;
    bool solved_ = false, dp_ = false, linal_ = false;
    std::vector<double> A_, B_, Qs_, mu_, Ps_, ds_, K_, f_, u_, xhat_;
};

}  // namespace solver
}  // namespace ilqr_planner
