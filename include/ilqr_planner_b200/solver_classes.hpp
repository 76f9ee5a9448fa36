// solver_classes.hpp — the four solver classes; included by solver.hpp.
#pragma once
#include <chrono>

namespace ilqr_planner {
namespace solver {

namespace detail_solver {
inline void report(CallBackMessage* cb, const BatchResult& r, bool with_time, double seconds) {
    const int it = r.iters[0];
    for (int i = 0; i < it; i++) {
        std::string line = "Iteration " + std::to_string(i + 1) + ", Cost: " + detail::fmt(r.cost_hist[size_t(i)]) + ", alpha= " + detail::fmt(r.alpha_hist[size_t(i)]);
        if (with_time) line += ", time= " + detail::fmt(seconds / (it > 0 ? it : 1));
        detail::emit(cb, line);
    }
}
inline double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
}  // namespace detail_solver

// ILQRRecursive.h:18-41, ILQRRecursive.cpp:21-181
class ILQRRecursive {
public:
    explicit ILQRRecursive(const std::shared_ptr<sys::System>& s_) : s(s_) {}

    std::tuple<std::vector<VectorXd>, std::vector<VectorXd>, std::vector<VectorXd>, std::vector<MatrixXd>, std::vector<VectorXd>, double>
    solve(const std::vector<VectorXd>& U0, int nb_iter, bool line_search = true, bool early_stop = false, CallBackMessage* cb = nullptr) {
        const double t0 = detail_solver::now();
        BatchResult r = solveBatch({s}, {U0}, nb_iter, line_search, early_stop);
        detail_solver::report(cb, r, true, detail_solver::now() - t0);
        const int H = r.H, n = r.n, m = r.m;
        std::vector<MatrixXd> Ks;
        for (int k = 0; k < H - 1; k++) Ks.push_back(detail::from_row_major(r.K.data() + size_t(k) * m * n, m, n));
        return std::make_tuple(detail_solver::rows(r.X, 0, H, n), detail_solver::rows(r.fX, 0, H, r.nt), detail_solver::rows(r.U, 0, H - 1, m), Ks,
                               detail_solver::rows(r.d, 0, H - 1, m), r.cost[0]);
    }

    // batched overload: systems[b] with initial controls U0s[b] (U0s empty = zeros)
    static BatchResult solveBatch(const std::vector<std::shared_ptr<sys::System>>& systems, const std::vector<std::vector<VectorXd>>& U0s, int nb_iter, bool line_search = true,
                                  bool early_stop = false, int device = 0) {
        detail::DescBuilder db(detail::flatten_all(systems));
        detail::HandleGuard g(db.desc(), device);
        BatchResult r;
        ilqrb_result o;
        detail_solver::alloc(r, db.dims(), nb_iter, true, 0, false, o);
        std::vector<double> u0 = U0s.empty() ? std::vector<double>(size_t(r.B) * (r.H - 1) * r.m, 0.0) : detail_solver::stack_u(U0s, r.H, r.m);
        if (u0.size() != size_t(r.B) * (r.H - 1) * r.m) throw std::runtime_error("[ILQRRecursive] one U0 per system expected");
        ilqrb_solve_opts op{nb_iter, line_search ? 1 : 0, early_stop ? 1 : 0, 0};
        detail::check(ilqrb_solve_recursive(g.h, u0.data(), &op, &o));
        return r;
    }

private:
    std::shared_ptr<sys::System> s;
};

// AL-ILQR.h:19-22
struct Constraint {
    MatrixXd A;
    VectorXd b;
};

// AL-ILQR.h:24-70, AL-ILQR.cpp:50-232
class AL_ILQR {
public:
    AL_ILQR(const std::shared_ptr<sys::System>& s_, const std::vector<Constraint>& inequality_, const std::vector<VectorXd>& initLambda) : s(s_), inequality(inequality_), multipliers(initLambda) {}

    std::tuple<std::vector<VectorXd>, std::vector<VectorXd>, std::vector<VectorXd>> solve(const std::vector<VectorXd>& U0, int nb_iter, int lag_update_step, double penalty,
                                                                                         double scaling_factor, bool line_search = true, bool early_stop = false,
                                                                                         CallBackMessage* cb = nullptr) {
        const double t0 = detail_solver::now();
        BatchResult r = solveBatch({s}, {U0}, nb_iter, lag_update_step, penalty, scaling_factor, line_search, early_stop);
        detail_solver::report(cb, r, true, detail_solver::now() - t0);
        multipliers = detail_solver::rows(r.lambda, 0, r.H - 1, r.nc);  // the reference keeps the updated multipliers in the solver
        return std::make_tuple(detail_solver::rows(r.X, 0, r.H, r.n), detail_solver::rows(r.fX, 0, r.H, r.nt), detail_solver::rows(r.U, 0, r.H - 1, r.m));
    }

    BatchResult solveBatch(const std::vector<std::shared_ptr<sys::System>>& systems, const std::vector<std::vector<VectorXd>>& U0s, int nb_iter, int lag_update_step, double penalty,
                           double scaling_factor, bool line_search = true, bool early_stop = false, int device = 0) const {
        detail::DescBuilder db(detail::flatten_all(systems));
        const ilqrb_dims& dm = db.dims();
        const int steps = dm.H - 1, w = dm.n + dm.m;
        if (int(inequality.size()) != steps || int(multipliers.size()) != steps) throw std::runtime_error("[AL_ILQR] one Constraint and one multiplier vector per timestep expected");
        const int nc = int(inequality[0].b.size());
        std::vector<double> A(size_t(steps) * nc * w), b(size_t(steps) * nc), lam(size_t(steps) * nc);
        for (int k = 0; k < steps; k++) {
            const Constraint& c = inequality[size_t(k)];
            if (c.A.rows() != nc || c.A.cols() != w || c.b.size() != nc || multipliers[size_t(k)].size() != nc)
                throw std::runtime_error("[AL_ILQR] constraint rows must have the same shape at every timestep: A nc x (n+m), b nc, lambda nc");
            for (int i = 0; i < nc; i++) {
                for (int j = 0; j < w; j++) A[(size_t(k) * nc + i) * w + j] = c.A(i, j);
                b[size_t(k) * nc + i] = c.b(i);
                lam[size_t(k) * nc + i] = multipliers[size_t(k)](i);
            }
        }
        ilqrb_al_desc al{};
        al.nc = nc;
        al.per_step = 1;
        al.A = A.data();
        al.b = b.data();
        al.lambda0 = lam.data();
        al.lag_update_step = lag_update_step;
        al.penalty = penalty;
        al.scaling_factor = scaling_factor;
        detail::HandleGuard g(db.desc(), device);
        BatchResult r;
        ilqrb_result o;
        detail_solver::alloc(r, dm, nb_iter, false, nc, false, o);
        std::vector<double> u0 = U0s.empty() ? std::vector<double>(size_t(r.B) * steps * r.m, 0.0) : detail_solver::stack_u(U0s, r.H, r.m);
        ilqrb_solve_opts op{nb_iter, line_search ? 1 : 0, early_stop ? 1 : 0, 0};
        detail::check(ilqrb_solve_al(g.h, u0.data(), &al, &op, &o));
        return r;
    }

private:
    std::shared_ptr<sys::System> s;
    std::vector<Constraint> inequality;
    std::vector<VectorXd> multipliers;
};

namespace detail_solver {
// shared body of BatchILQR / BatchILQRCP
inline BatchResult run_batch(const std::vector<std::shared_ptr<sys::System>>& systems, const std::vector<VectorXd>& u0s, int nb_iter, bool early_stop, const MatrixXd* Q,
                             const MatrixXd* psi, int device) {
    detail::DescBuilder db(detail::flatten_all(systems));
    const ilqrb_dims& dm = db.dims();
    const size_t N = size_t(dm.H - 1) * dm.m;
    std::vector<double> Qr, Pr;
    ilqrb_batch_desc bd{};
    if (Q) {
        if (Q->rows() != Q->cols()) throw std::runtime_error("[BatchILQR] Q must be square");
        Qr = detail::to_row_major(*Q);
        bd.Q = Qr.data();
    }
    if (psi) {
        if (size_t(psi->rows()) != N) throw std::runtime_error("[BatchILQRCP] psi must have (horizon-1)*nb_ctrl_var rows");
        Pr = detail::to_row_major(*psi);
        bd.psi = Pr.data();
        bd.nw = int32_t(psi->cols());
    }
    std::vector<double> u0;
    for (auto& v : u0s) {
        if (size_t(v.size()) != N) throw std::runtime_error("[BatchILQR] u0 must have (horizon-1)*nb_ctrl_var entries");
        for (size_t i = 0; i < N; i++) u0.push_back(v(long(i)));
    }
    if (u0.empty()) u0.assign(size_t(dm.B) * N, 0.0);
    if (u0.size() != size_t(dm.B) * N) throw std::runtime_error("[BatchILQR] one u0 per system expected");
    detail::HandleGuard g(db.desc(), device);
    BatchResult r;
    ilqrb_result o;
    alloc(r, dm, nb_iter, false, 0, true, o);
    ilqrb_solve_opts op{nb_iter, 1, early_stop ? 1 : 0, 0};
    detail::check(ilqrb_solve_batch(g.h, u0.data(), &bd, &op, &o));
    return r;
}
}  // namespace detail_solver

// BatchILQR.h:18-77, BatchILQR.cpp:22-33,110-173
class BatchILQR {
public:
    explicit BatchILQR(const std::shared_ptr<sys::System>& s_) : s(s_), has_Q(false) {}
    BatchILQR(const std::shared_ptr<sys::System>& s_, const MatrixXd& Q_) : Q(Q_), s(s_), has_Q(true) {}
    VectorXd solve(int nb_iter, const VectorXd& u0, bool early_stop = false, CallBackMessage* cb = nullptr) {
        BatchResult r = solveBatch({s}, {u0}, nb_iter, early_stop);
        detail_solver::report(cb, r, false, 0);
        return detail::from_buffer(r.U.data(), long(r.U.size()));
    }
    BatchResult solveBatch(const std::vector<std::shared_ptr<sys::System>>& systems, const std::vector<VectorXd>& u0s, int nb_iter, bool early_stop = false, int device = 0) const {
        return detail_solver::run_batch(systems, u0s, nb_iter, early_stop, has_Q ? &Q : nullptr, nullptr, device);
    }

private:
    MatrixXd Q;
    std::shared_ptr<sys::System> s;
    bool has_Q;
};

// BatchILQRCP.h:18-77, BatchILQRCP.cpp:21-32,109-174
class BatchILQRCP {
public:
    BatchILQRCP(const std::shared_ptr<sys::System>& s_, const MatrixXd& Q_, const MatrixXd& psi) : PSI(psi), Q(Q_), s(s_), has_Q(true) {}
    BatchILQRCP(const std::shared_ptr<sys::System>& s_, const MatrixXd& psi) : PSI(psi), s(s_), has_Q(false) {}
    VectorXd solve(int nb_iter, const VectorXd& u0, bool early_stop = false, CallBackMessage* cb = nullptr) {
        BatchResult r = solveBatch({s}, {u0}, nb_iter, early_stop);
        detail_solver::report(cb, r, false, 0);
        return detail::from_buffer(r.U.data(), long(r.U.size()));
    }
    BatchResult solveBatch(const std::vector<std::shared_ptr<sys::System>>& systems, const std::vector<VectorXd>& u0s, int nb_iter, bool early_stop = false, int device = 0) const {
        return detail_solver::run_batch(systems, u0s, nb_iter, early_stop, has_Q ? &Q : nullptr, &PSI, device);
    }

private:
    MatrixXd PSI, Q;
    std::shared_ptr<sys::System> s;
    bool has_Q;
};

}  // namespace solver

// primitives.h / primitives.cpp:19-96
// the reference declares these directly in ilqr_planner:: (primitives.h:12-17)
inline MatrixXd build_psi(const char* kind, int dim, int K) {
    const int cols = std::string(kind) == "linear" ? 2 * K : K;
    std::vector<double> o(size_t(dim) * cols);
    detail::check(ilqrb_build_psi(kind, dim, K, o.data()));
    return detail::from_row_major(o.data(), dim, cols);
}
inline MatrixXd buildPsiRBF(int dim, int K) { return build_psi("rbf", dim, K); }
inline MatrixXd buildPsiBernstein(int dim, int K) { return build_psi("bernstein", dim, K); }
inline MatrixXd buildPsiUnitstep(int dim, int K) { return build_psi("unitstep", dim, K); }
inline MatrixXd buildPsiSawtooth(int dim, int K) { return build_psi("sawtooth", dim, K); }
inline MatrixXd buildPsiLinear(int dim, int K) { return build_psi("linear", dim, K); }

}  // namespace ilqr_planner
