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
// ILQRB_ST_QUU_NOT_SPD of any problem: the backward pass met a control Hessian that is not positive definite (the step taken then
// is not a descent step of a convex model); the reference inverts such a Quu silently (ILQRRecursive.cpp:89), here it is said once per solve
inline void warn_status(const int32_t* status, int B) {
    int bad = 0;
    for (int b = 0; b < B; b++) bad += (status[b] & ILQRB_ST_QUU_NOT_SPD) ? 1 : 0;
    // to stderr, never through the callback: the callback carries the reference's per-iteration lines and nothing else
    if (bad)
        std::cerr << "[ilqr_planner_b200] note: Quu + 1e-6 I was not positive definite in " << bad << " of " << B
                  << " problems (ILQRB_ST_QUU_NOT_SPD): those steps used the pivoted LU inverse, as the reference does" << std::endl;
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
        detail_solver::warn_status(r.status.data(), r.B);
        const int H = r.H, n = r.n, m = r.m;
        std::vector<MatrixXd> Ks;
        for (int k = 0; k < H - 1; k++) Ks.push_back(detail::from_row_major(r.K.data() + size_t(k) * m * n, m, n));
        return std::make_tuple(detail_solver::rows(r.X, 0, H, n), detail_solver::rows(r.fX, 0, H, r.nt), detail_solver::rows(r.U, 0, H - 1, m), Ks,
                               detail_solver::rows(r.d, 0, H - 1, m), r.cost[0]);
    }

    // batched overload: systems[b] with initial controls U0s[b] (U0s empty = zeros).  The handle stays with the solver object:
    // a later batch of the same structure only replaces its per-problem data on the device.
    BatchResult solveBatch(const std::vector<std::shared_ptr<sys::System>>& systems, const std::vector<std::vector<VectorXd>>& U0s, int nb_iter, bool line_search = true,
                           bool early_stop = false, const std::vector<int>& devices = {0}) {
        detail::DescBuilder db(detail::flatten_all(systems));
        st_->cache.prepare(db, devices);
        BatchResult r;
        ilqrb_result o;
        detail_solver::alloc(r, db.dims(), nb_iter, true, 0, false, o);
        std::vector<double> u0 = U0s.empty() ? std::vector<double>(size_t(r.B) * (r.H - 1) * r.m, 0.0) : detail_solver::stack_u(U0s, r.H, r.m);
        if (u0.size() != size_t(r.B) * (r.H - 1) * r.m) throw std::runtime_error("[ILQRRecursive] one U0 per system expected");
        ilqrb_solve_opts op{nb_iter, line_search ? 1 : 0, early_stop ? 1 : 0, 0};
        if (st_->cache.multi()) detail::check(ilqrb_multi_solve_recursive(st_->cache.multi(), u0.data(), &op, &o));
        else detail::check(ilqrb_solve_recursive(st_->cache.handle(), u0.data(), &op, &o));
        return r;
    }
    // array form: B problems of this solver's system that differ in q0 / dq0 / keypoint states; no per-problem objects, results
    // in the solver's page-locked arena (valid until the next batch solve of this object); gains = false leaves K and d on the device
    BatchView solveBatch(const BatchArrays& in, int nb_iter, bool line_search = true, bool early_stop = false, bool gains = true, const std::vector<int>& devices = {0}) {
        ilqrb_result o;
        BatchView v = detail_solver::prepare_arrays(*st_, *s, in, nb_iter, gains, 0, false, devices, o);
        ilqrb_solve_opts op{nb_iter, line_search ? 1 : 0, early_stop ? 1 : 0, 0};
        if (st_->cache.multi()) detail::check(ilqrb_multi_solve_recursive(st_->cache.multi(), in.U0, &op, &o));
        else detail::check(ilqrb_solve_recursive(st_->cache.handle(), in.U0, &op, &o));
        return v;
    }
    // (handles created, handles reused) so far — what the cache did
    std::pair<int, int> cacheStats() const { return {st_->cache.created(), st_->cache.reused()}; }
    // owner of the memory a BatchView points into (bindings tie the lifetime of their array views to it)
    std::shared_ptr<void> resultOwner() const { return st_; }
    const std::shared_ptr<sys::System>& system() const { return s; }

private:
    std::shared_ptr<sys::System> s;
    std::shared_ptr<detail_solver::BatchState> st_ = std::make_shared<detail_solver::BatchState>();
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
        detail_solver::warn_status(r.status.data(), r.B);
        multipliers = detail_solver::rows(r.lambda, 0, r.H - 1, r.nc);  // the reference keeps the updated multipliers in the solver
        return std::make_tuple(detail_solver::rows(r.X, 0, r.H, r.n), detail_solver::rows(r.fX, 0, r.H, r.nt), detail_solver::rows(r.U, 0, r.H - 1, r.m));
    }

    BatchResult solveBatch(const std::vector<std::shared_ptr<sys::System>>& systems, const std::vector<std::vector<VectorXd>>& U0s, int nb_iter, int lag_update_step, double penalty,
                           double scaling_factor, bool line_search = true, bool early_stop = false, const std::vector<int>& devices = {0}) const {
        detail::DescBuilder db(detail::flatten_all(systems));
        const ilqrb_dims& dm = db.dims();
        ConstraintArrays ca;
        ilqrb_al_desc al = alDesc(dm, lag_update_step, penalty, scaling_factor, ca);
        st_->cache.prepare(db, devices);
        BatchResult r;
        ilqrb_result o;
        detail_solver::alloc(r, dm, nb_iter, false, al.nc, false, o);
        std::vector<double> u0 = U0s.empty() ? std::vector<double>(size_t(r.B) * (dm.H - 1) * r.m, 0.0) : detail_solver::stack_u(U0s, r.H, r.m);
        ilqrb_solve_opts op{nb_iter, line_search ? 1 : 0, early_stop ? 1 : 0, 0};
        if (st_->cache.multi()) detail::check(ilqrb_multi_solve_al(st_->cache.multi(), u0.data(), &al, &op, &o));
        else detail::check(ilqrb_solve_al(st_->cache.handle(), u0.data(), &al, &op, &o));
        return r;
    }
    // array form (see ILQRRecursive::solveBatch); every problem starts from the solver's constraints and multipliers
    BatchView solveBatch(const BatchArrays& in, int nb_iter, int lag_update_step, double penalty, double scaling_factor, bool line_search = true, bool early_stop = false,
                         const std::vector<int>& devices = {0}) const {
        ilqrb_result o;
        const int nc = inequality.empty() ? 0 : int(inequality[0].b.size());
        BatchView v = detail_solver::prepare_arrays(*st_, *s, in, nb_iter, false, nc, false, devices, o);
        ConstraintArrays ca;
        ilqrb_al_desc al = alDesc(st_->db->dims(), lag_update_step, penalty, scaling_factor, ca);
        ilqrb_solve_opts op{nb_iter, line_search ? 1 : 0, early_stop ? 1 : 0, 0};
        if (st_->cache.multi()) detail::check(ilqrb_multi_solve_al(st_->cache.multi(), in.U0, &al, &op, &o));
        else detail::check(ilqrb_solve_al(st_->cache.handle(), in.U0, &al, &op, &o));
        return v;
    }
    std::pair<int, int> cacheStats() const { return {st_->cache.created(), st_->cache.reused()}; }
    // owner of the memory a BatchView points into (bindings tie the lifetime of their array views to it)
    std::shared_ptr<void> resultOwner() const { return st_; }
    const std::shared_ptr<sys::System>& system() const { return s; }

private:
    struct ConstraintArrays {
        std::vector<double> A, b, lam;
    };
    // the constraint list as the [H-1][nc][n+m] / [H-1][nc] arrays of ilqrb_al_desc
    ilqrb_al_desc alDesc(const ilqrb_dims& dm, int lag_update_step, double penalty, double scaling_factor, ConstraintArrays& ca) const {
        const int steps = dm.H - 1, w = dm.n + dm.m;
        if (int(inequality.size()) != steps || int(multipliers.size()) != steps) throw std::runtime_error("[AL_ILQR] one Constraint and one multiplier vector per timestep expected");
        const int nc = int(inequality[0].b.size());
        std::vector<double>&A = ca.A, &b = ca.b, &lam = ca.lam;
        A.assign(size_t(steps) * nc * w, 0.0);
        b.assign(size_t(steps) * nc, 0.0);
        lam.assign(size_t(steps) * nc, 0.0);
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
        return al;
    }
    std::shared_ptr<sys::System> s;
    std::vector<Constraint> inequality;
    std::vector<VectorXd> multipliers;
    std::shared_ptr<detail_solver::BatchState> st_ = std::make_shared<detail_solver::BatchState>();
};

namespace detail_solver {
// Q / psi of BatchILQR / BatchILQRCP as the row-major arrays of ilqrb_batch_desc
struct BatchDescArrays {
    std::vector<double> Q, psi;
};
inline ilqrb_batch_desc batch_desc(const ilqrb_dims& dm, const MatrixXd* Q, const MatrixXd* psi, BatchDescArrays& a) {
    const size_t N = size_t(dm.H - 1) * dm.m;
    ilqrb_batch_desc bd{};
    if (Q) {
        if (Q->rows() != Q->cols()) throw std::runtime_error("[BatchILQR] Q must be square");
        a.Q = detail::to_row_major(*Q);
        bd.Q = a.Q.data();
    }
    if (psi) {
        if (size_t(psi->rows()) != N) throw std::runtime_error("[BatchILQRCP] psi must have (horizon-1)*nb_ctrl_var rows");
        a.psi = detail::to_row_major(*psi);
        bd.psi = a.psi.data();
        bd.nw = int32_t(psi->cols());
    }
    return bd;
}
inline void call_batch(BatchState& st, const double* u0, const ilqrb_batch_desc& bd, const ilqrb_solve_opts& op, ilqrb_result& o) {
    if (st.cache.multi()) detail::check(ilqrb_multi_solve_batch(st.cache.multi(), u0, &bd, &op, &o));
    else detail::check(ilqrb_solve_batch(st.cache.handle(), u0, &bd, &op, &o));
}
// shared bodies of BatchILQR / BatchILQRCP
inline BatchResult run_batch(BatchState& st, const std::vector<std::shared_ptr<sys::System>>& systems, const std::vector<VectorXd>& u0s, int nb_iter, bool early_stop,
                             const MatrixXd* Q, const MatrixXd* psi, const std::vector<int>& devices) {
    detail::DescBuilder db(detail::flatten_all(systems));
    const ilqrb_dims& dm = db.dims();
    const size_t N = size_t(dm.H - 1) * dm.m;
    BatchDescArrays arr;
    ilqrb_batch_desc bd = batch_desc(dm, Q, psi, arr);
    std::vector<double> u0;
    for (auto& v : u0s) {
        if (size_t(v.size()) != N) throw std::runtime_error("[BatchILQR] u0 must have (horizon-1)*nb_ctrl_var entries");
        for (size_t i = 0; i < N; i++) u0.push_back(v(long(i)));
    }
    if (u0.empty()) u0.assign(size_t(dm.B) * N, 0.0);
    if (u0.size() != size_t(dm.B) * N) throw std::runtime_error("[BatchILQR] one u0 per system expected");
    st.cache.prepare(db, devices);
    BatchResult r;
    ilqrb_result o;
    alloc(r, dm, nb_iter, false, 0, true, o);
    ilqrb_solve_opts op{nb_iter, 1, early_stop ? 1 : 0, 0};
    call_batch(st, u0.data(), bd, op, o);
    return r;
}
inline BatchView run_batch(BatchState& st, const sys::System& s, const BatchArrays& in, int nb_iter, bool early_stop, const MatrixXd* Q, const MatrixXd* psi,
                           const std::vector<int>& devices) {
    ilqrb_result o;
    BatchView v = prepare_arrays(st, s, in, nb_iter, false, 0, true, devices, o);
    BatchDescArrays arr;
    ilqrb_batch_desc bd = batch_desc(st.db->dims(), Q, psi, arr);
    ilqrb_solve_opts op{nb_iter, 1, early_stop ? 1 : 0, 0};
    call_batch(st, in.U0, bd, op, o);
    return v;
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
    BatchResult solveBatch(const std::vector<std::shared_ptr<sys::System>>& systems, const std::vector<VectorXd>& u0s, int nb_iter, bool early_stop = false,
                           const std::vector<int>& devices = {0}) const {
        return detail_solver::run_batch(*st_, systems, u0s, nb_iter, early_stop, has_Q ? &Q : nullptr, nullptr, devices);
    }
    // array form (see ILQRRecursive::solveBatch); in.U0 is [B][(H-1) m]
    BatchView solveBatch(const BatchArrays& in, int nb_iter, bool early_stop = false, const std::vector<int>& devices = {0}) const {
        return detail_solver::run_batch(*st_, *s, in, nb_iter, early_stop, has_Q ? &Q : nullptr, nullptr, devices);
    }
    std::pair<int, int> cacheStats() const { return {st_->cache.created(), st_->cache.reused()}; }
    // owner of the memory a BatchView points into (bindings tie the lifetime of their array views to it)
    std::shared_ptr<void> resultOwner() const { return st_; }
    const std::shared_ptr<sys::System>& system() const { return s; }

private:
    MatrixXd Q;
    std::shared_ptr<sys::System> s;
    bool has_Q;
    std::shared_ptr<detail_solver::BatchState> st_ = std::make_shared<detail_solver::BatchState>();
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
    BatchResult solveBatch(const std::vector<std::shared_ptr<sys::System>>& systems, const std::vector<VectorXd>& u0s, int nb_iter, bool early_stop = false,
                           const std::vector<int>& devices = {0}) const {
        return detail_solver::run_batch(*st_, systems, u0s, nb_iter, early_stop, has_Q ? &Q : nullptr, &PSI, devices);
    }
    // array form (see ILQRRecursive::solveBatch); in.U0 is [B][(H-1) m]
    BatchView solveBatch(const BatchArrays& in, int nb_iter, bool early_stop = false, const std::vector<int>& devices = {0}) const {
        return detail_solver::run_batch(*st_, *s, in, nb_iter, early_stop, has_Q ? &Q : nullptr, &PSI, devices);
    }
    std::pair<int, int> cacheStats() const { return {st_->cache.created(), st_->cache.reused()}; }
    // owner of the memory a BatchView points into (bindings tie the lifetime of their array views to it)
    std::shared_ptr<void> resultOwner() const { return st_; }
    const std::shared_ptr<sys::System>& system() const { return s; }

private:
    MatrixXd PSI, Q;
    std::shared_ptr<sys::System> s;
    bool has_Q;
    std::shared_ptr<detail_solver::BatchState> st_ = std::make_shared<detail_solver::BatchState>();
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
