// pylqr_bindings.cpp — the `PyLQR` Python module of the reference (pylqr_planner/src/bindings.cpp:48-907) bound over the C++ host API
// of include/ilqr_planner_b200/*.hpp, so that the tutorial notebooks' code runs against the GPU solvers: same sub-modules
// (sim, system, solver, utils, utils.primitives), class names, constructor overloads, snake_case method names and
// keyword arguments.  Built by __graft_entry__.build() into PyLQR/_pylqr*.so; PyLQR/__init__.py registers the sub-modules.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <ilqr_planner_b200/ilqr_planner.hpp>

namespace py = pybind11;
using namespace ilqr_planner;

// numpy <-> VectorXd / MatrixXd (the reference relies on pybind11/eigen.h for the same conversions)
namespace pybind11 {
namespace detail {
template <>
struct type_caster<ilqr_planner::VectorXd> {
    PYBIND11_TYPE_CASTER(ilqr_planner::VectorXd, const_name("numpy.ndarray[numpy.float64[m, 1]]"));
    bool load(handle src, bool) {
        auto a = array_t<double, array::c_style | array::forcecast>::ensure(src);
        if (!a) {
            PyErr_Clear();
            return false;
        }
        if (a.ndim() > 2 || (a.ndim() == 2 && a.shape(0) != 1 && a.shape(1) != 1)) return false;
        value = ilqr_planner::VectorXd(long(a.size()));
        const double* p = a.data();
        for (ssize_t i = 0; i < a.size(); i++) value(long(i)) = p[i];
        return true;
    }
    static handle cast(const ilqr_planner::VectorXd& v, return_value_policy, handle) {
        array_t<double> a(static_cast<ssize_t>(v.size()));
        double* p = a.mutable_data();
        for (long i = 0; i < v.size(); i++) p[i] = v(i);
        return a.release();
    }
};
template <>
struct type_caster<ilqr_planner::MatrixXd> {
    PYBIND11_TYPE_CASTER(ilqr_planner::MatrixXd, const_name("numpy.ndarray[numpy.float64[m, n]]"));
    bool load(handle src, bool) {
        auto a = array_t<double, array::c_style | array::forcecast>::ensure(src);
        if (!a) {
            PyErr_Clear();
            return false;
        }
        if (a.ndim() != 2) return false;
        const long r = long(a.shape(0)), c = long(a.shape(1));
        value = ilqr_planner::MatrixXd(r, c);
        const double* p = a.data();
        for (long i = 0; i < r; i++)
            for (long j = 0; j < c; j++) value(i, j) = p[i * c + j];
        return true;
    }
    static handle cast(const ilqr_planner::MatrixXd& m, return_value_policy, handle) {
        array_t<double> a({static_cast<ssize_t>(m.rows()), static_cast<ssize_t>(m.cols())});
        double* p = a.mutable_data();
        for (long i = 0; i < m.rows(); i++)
            for (long j = 0; j < m.cols(); j++) p[i * m.cols() + j] = m(i, j);
        return a.release();
    }
};
}  // namespace detail
}  // namespace pybind11

// pylqr_planner/src/PythonCallbackMessage.h: prints through Python so notebooks show the trace; also keeps the lines
class PythonCallbackMessage : public CallBackMessage {
public:
    void notify(const std::string& msg) override {
        py::gil_scoped_acquire gil;  // the solvers run with the GIL released
        messages.push_back(msg);
        py::print(msg);
    }
    std::vector<std::string> messages;
};


// ---- array-based batch solves: numpy in, numpy views of the solver's page-locked result arena out -------------------------------
using DArr = py::array_t<double, py::array::c_style | py::array::forcecast>;

struct ArrayInputs {
    solver::BatchArrays in;
    std::vector<DArr> keep;  // the (possibly converted) arrays the pointers refer to
};

// q0 [B, dof]; kp_states: None, one array [B, Kp, w] or a list with one array (or None) per sub-system; U0 [B, ...] or None
inline ArrayInputs array_inputs(const py::object& q0, const py::object& kp_states, const py::object& U0, const py::object& dq0, const std::shared_ptr<sys::System>& s,
                                long u0_row) {
    ArrayInputs a;
    int B = -1;
    auto lead = [&B](const DArr& x, const char* what) {
        if (x.ndim() < 1) throw std::invalid_argument(std::string("solve_batch: ") + what + " needs a leading batch dimension");
        if (B < 0) B = int(x.shape(0));
        else if (B != int(x.shape(0))) throw std::invalid_argument(std::string("solve_batch: ") + what + " has another batch size than the arrays before it");
    };
    const long dof = s->r->getDOF();
    if (!q0.is_none()) {
        DArr x = DArr::ensure(q0);
        if (!x || x.ndim() != 2 || x.shape(1) != dof) throw std::invalid_argument("solve_batch: q0 must be [B, dof]");
        lead(x, "q0");
        a.in.q0 = x.data();
        a.keep.push_back(x);
    }
    if (!dq0.is_none()) {
        DArr x = DArr::ensure(dq0);
        if (!x || x.ndim() != 2 || x.shape(1) != dof) throw std::invalid_argument("solve_batch: dq0 must be [B, dof]");
        lead(x, "dq0");
        a.in.dq0 = x.data();
        a.keep.push_back(x);
    }
    if (!kp_states.is_none()) {
        detail::Flat f;
        s->flatten(f);
        std::vector<py::object> per;
        if (py::isinstance<py::list>(kp_states) || py::isinstance<py::tuple>(kp_states))
            for (auto h : kp_states) per.push_back(py::reinterpret_borrow<py::object>(h));
        else per.push_back(kp_states);
        if (per.size() != f.tasks.size()) throw std::invalid_argument("solve_batch: kp_states needs one array per sub-system (" + std::to_string(f.tasks.size()) + ")");
        for (size_t t = 0; t < per.size(); t++) {
            if (per[t].is_none()) {
                a.in.kp_states.push_back(nullptr);
                continue;
            }
            DArr x = DArr::ensure(per[t]);
            const long kp = long(f.tasks[t].timestep.size()), w = detail::state_width(f, f.tasks[t]);
            if (!x || x.ndim() != 3 || x.shape(1) != kp || x.shape(2) != w)
                throw std::invalid_argument("solve_batch: kp_states[" + std::to_string(t) + "] must be [B, " + std::to_string(kp) + ", " + std::to_string(w) +
                                            "] (the keypoints' get_state() vectors in timestep order)");
            lead(x, "kp_states");
            a.in.kp_states.push_back(x.data());
            a.keep.push_back(x);
        }
    }
    if (!U0.is_none()) {
        DArr x = DArr::ensure(U0);
        if (!x || x.ndim() < 2 || x.size() != x.shape(0) * u0_row) throw std::invalid_argument("solve_batch: U0 must be [B, horizon-1, nb_ctrl_var]");
        lead(x, "U0");
        a.in.U0 = x.data();
        a.keep.push_back(x);
    }
    if (B < 1) throw std::invalid_argument("solve_batch: give at least one of q0, dq0, kp_states, U0 (their leading dimension is the batch size)");
    a.in.B = B;
    return a;
}

inline py::dict view_dict(const solver::BatchView& v, const std::shared_ptr<void>& owner, bool batch_solver) {
    // the arrays alias the solver's arena: the capsule keeps the arena alive, the NEXT solve_batch of the same solver overwrites them
    auto* keep = new std::shared_ptr<void>(owner);
    py::capsule base(keep, [](void* p) { delete static_cast<std::shared_ptr<void>*>(p); });
    auto arr = [&base](double* p, std::vector<ssize_t> shape) -> py::object {
        if (!p) return py::none();
        return py::array_t<double>(shape, p, base);
    };
    const ssize_t B = v.B, H = v.H, n = v.n, m = v.m, it = v.nb_iter > 0 ? v.nb_iter : 1;
    py::dict d;
    if (batch_solver) {
        d["U"] = arr(v.U, {B, (H - 1) * m});
    } else {
        d["X"] = arr(v.X, {B, H, n});
        d["fX"] = arr(v.fX, {B, H, ssize_t(v.nt)});
        d["U"] = arr(v.U, {B, H - 1, m});
        d["K"] = arr(v.K, {B, H - 1, m, n});
        d["d"] = arr(v.d, {B, H - 1, m});
        if (v.nc > 0) d["lambda"] = arr(v.lambda, {B, H - 1, ssize_t(v.nc)});
    }
    d["cost"] = arr(v.cost, {B});
    d["cost_hist"] = arr(v.cost_hist, {B, it});
    d["alpha_hist"] = arr(v.alpha_hist, {B, it});
    d["iters"] = py::array_t<int32_t>(std::vector<ssize_t>{B}, v.iters, base);
    d["status"] = py::array_t<int32_t>(std::vector<ssize_t>{B}, v.status, base);
    if (v.anyQuuNotSpd())
        PyErr_WarnEx(PyExc_RuntimeWarning, "Quu + 1e-6 I was not positive definite in some problems (status & ILQRB_ST_QUU_NOT_SPD): those steps used the pivoted LU inverse, as the reference does", 1);
    return d;
}

inline py::dict result_dict(const solver::BatchResult& r, bool batch_solver) {
    auto arr = [](const std::vector<double>& v, std::vector<ssize_t> shape) -> py::object {
        if (v.empty()) return py::none();
        py::array_t<double> a(shape);
        std::copy(v.begin(), v.end(), a.mutable_data());
        return a;
    };
    const ssize_t B = r.B, H = r.H, n = r.n, m = r.m, it = r.nb_iter > 0 ? r.nb_iter : 1;
    py::dict d;
    if (batch_solver) {
        d["U"] = arr(r.U, {B, (H - 1) * m});
    } else {
        d["X"] = arr(r.X, {B, H, n});
        d["fX"] = arr(r.fX, {B, H, ssize_t(r.nt)});
        d["U"] = arr(r.U, {B, H - 1, m});
        d["K"] = arr(r.K, {B, H - 1, m, n});
        d["d"] = arr(r.d, {B, H - 1, m});
        if (r.nc > 0) d["lambda"] = arr(r.lambda, {B, H - 1, ssize_t(r.nc)});
    }
    d["cost"] = arr(r.cost, {B});
    d["cost_hist"] = arr(r.cost_hist, {B, it});
    d["alpha_hist"] = arr(r.alpha_hist, {B, it});
    py::array_t<int32_t> iters(std::vector<ssize_t>{B}), status(std::vector<ssize_t>{B});
    std::copy(r.iters.begin(), r.iters.end(), iters.mutable_data());
    std::copy(r.status.begin(), r.status.end(), status.mutable_data());
    d["iters"] = iters;
    d["status"] = status;
    for (int32_t st : r.status)
        if (st & ILQRB_ST_QUU_NOT_SPD) {
            PyErr_WarnEx(PyExc_RuntimeWarning, "Quu + 1e-6 I was not positive definite in some problems (status & ILQRB_ST_QUU_NOT_SPD): those steps used the pivoted LU inverse, as the reference does", 1);
            break;
        }
    return d;
}

using SysList = std::vector<std::shared_ptr<sys::System>>;
static const char* kBatchDoc =
    "Batched solve on the GPU: B problems that share this solver's system structure and differ in the initial joint state (q0, dq0: [B, dof]) and the keypoint "
    "states (kp_states: [B, Kp, w] in Keypoint.get_state() layout, or a list with one such array per sub-system). Anything left None takes the system's own value "
    "for all problems. Returns a dict of numpy arrays with a leading B; they alias page-locked memory owned by the solver and are overwritten by its next solve_batch. "
    "The GIL is released while the GPU works.";

using KpList = std::vector<std::shared_ptr<sys::Keypoint>>;
using SimPtr = std::shared_ptr<sim::SimulationInterface>;

PYBIND11_MODULE(_pylqr, m) {
    m.doc() = "PyLQR on the B200 batched iLQR solver (C ABI: include/ilqrb.h)";

    // ---------------------------------------------------------------- sim (bindings.cpp:65-216)
    py::module m_sim = m.def_submodule("sim");
    py::class_<sim::SimulationInterface, SimPtr>(m_sim, "SimulationInterface")
        .def("update_kinematics", &sim::SimulationInterface::updateKinematics)
        .def("Jt", &sim::SimulationInterface::Jt)
        .def("Jr", &sim::SimulationInterface::Jr)
        .def("J", &sim::SimulationInterface::J)
        .def("get_ee_pos", &sim::SimulationInterface::getEEPosition)
        .def("get_ee_orn", &sim::SimulationInterface::getEEOrnQuat)
        .def("get_ee_vel", &sim::SimulationInterface::getEEVelocity)
        .def("get_ee_ang_vel", &sim::SimulationInterface::getEEAngVel)
        .def("get_ee_ang_vel_quat", &sim::SimulationInterface::getEEAngVelQuat)
        .def("get_q", &sim::SimulationInterface::getJointsPos)
        .def("get_dq", &sim::SimulationInterface::getJointsVel)
        .def("get_time", &sim::SimulationInterface::getTime)
        .def("set_time", &sim::SimulationInterface::setTime, py::arg("time"))
        .def("dquat_to_w_jac", &sim::SimulationInterface::dQuatToDxJac, py::arg("quat"))
        .def("set_conf", &sim::SimulationInterface::setConfiguration, py::arg("q"), py::arg("dq"), py::arg("reset_time") = true)
        .def("send_acc", &sim::SimulationInterface::sendAcc, py::arg("dt"), py::arg("ddq"), py::arg("update_kin") = true)
        .def("send_vel", &sim::SimulationInterface::sendVel, py::arg("dt"), py::arg("dq"), py::arg("update_kin") = true)
        .def("get_dof", &sim::SimulationInterface::getDOF);
    py::class_<sim::KDLRobot, sim::SimulationInterface, std::shared_ptr<sim::KDLRobot>>(m_sim, "KDLRobot")
        .def(py::init<const std::string&, const std::string&, const std::string&, const VectorXd&, const VectorXd&, const VectorXd&, const VectorXd&, const bool&>(), py::arg("urdf"),
             py::arg("base_frame"), py::arg("tip_frame"), py::arg("q"), py::arg("dq"), py::arg("transform_rpy"), py::arg("transform_xyz"), py::arg("is_path"))
        .def(py::init<const std::string&, const std::string&, const std::string&, const VectorXd&, const VectorXd&, const VectorXd&, const VectorXd&>(), py::arg("urdf"),
             py::arg("base_frame"), py::arg("tip_frame"), py::arg("q"), py::arg("dq"), py::arg("transform_rpy"), py::arg("transform_xyz"))
        .def(py::init<const std::string&, const std::string&, const std::string&, const VectorXd&, const VectorXd&>(), py::arg("urdf"), py::arg("base_frame"), py::arg("tip_frame"),
             py::arg("q"), py::arg("dq"));
    py::class_<sim::Robot2D, sim::SimulationInterface, std::shared_ptr<sim::Robot2D>>(m_sim, "Robot2D")
        .def(py::init<const VectorXd&, const VectorXd&>(), py::arg("lengths"), py::arg("default_q"))
        .def("fkine", static_cast<VectorXd (sim::Robot2D::*)()>(&sim::Robot2D::fkine))
        .def("fkine", static_cast<VectorXd (sim::Robot2D::*)(const VectorXd&)>(&sim::Robot2D::fkine), py::arg("q"));
    py::class_<sim::TransformedSimulationInterface, sim::SimulationInterface, std::shared_ptr<sim::TransformedSimulationInterface>>(m_sim, "TransformedSimulationInterface")
        .def(py::init<const SimPtr&, const MatrixXd&>(), py::arg("r"), py::arg("T"))
        .def(py::init<const MatrixXd&>(), py::arg("T"))
        .def("subscribe", &sim::TransformedSimulationInterface::subscribe, py::arg("r"));

    // ---------------------------------------------------------------- system (bindings.cpp:219-692)
    py::module m_sys = m.def_submodule("system");
    py::class_<sys::Keypoint, std::shared_ptr<sys::Keypoint>>(m_sys, "Keypoint")
        .def("diff", &sys::Keypoint::diff, py::arg("state"))
        .def("get_state", &sys::Keypoint::getState)
        .def("get_precision", &sys::Keypoint::getPrecision)
        .def("get_timestep", &sys::Keypoint::getTimestep);
    py::class_<sys::PosOrnKeypoint, sys::Keypoint, std::shared_ptr<sys::PosOrnKeypoint>>(m_sys, "PosOrnKeypoint")
        .def(py::init<const VectorXd&, const VectorXd&, const MatrixXd&, const int&>(), py::arg("position"), py::arg("orientation"), py::arg("precision"), py::arg("timestep"))
        .def(py::init<const VectorXd&, const VectorXd&, const VectorXd&, const VectorXd&, const MatrixXd&, const int&>(), py::arg("position"), py::arg("dposition"),
             py::arg("orientation"), py::arg("dorientation"), py::arg("precision"), py::arg("timestep"))
        .def("get_position", &sys::PosOrnKeypoint::getPosition)
        .def("get_orientation", &sys::PosOrnKeypoint::getOrientation);
    py::class_<sys::SpacetimeKeypoint, sys::PosOrnKeypoint, std::shared_ptr<sys::SpacetimeKeypoint>>(m_sys, "SpacetimeKeypoint")
        .def(py::init<const VectorXd&, const VectorXd&, const MatrixXd&, const double&, const int&>(), py::arg("position"), py::arg("orientation"), py::arg("precision"),
             py::arg("continuous_time"), py::arg("timestep"))
        .def(py::init<const VectorXd&, const VectorXd&, const VectorXd&, const VectorXd&, const MatrixXd&, const double&, const int&>(), py::arg("position"), py::arg("dposition"),
             py::arg("orientation"), py::arg("dorientation"), py::arg("precision"), py::arg("continuous_time"), py::arg("timestep"))
        .def("get_continuous_time", &sys::SpacetimeKeypoint::getContinuousTime);

    py::class_<sys::PosOrnKeypointDistFunct, sys::PosOrnKeypoint, std::shared_ptr<sys::PosOrnKeypointDistFunct>>(m_sys, "PosOrnKeypointDistFunct")
        .def(py::init<const VectorXd&, const VectorXd&, const MatrixXd&, const double&, const VectorXd&, const int&>(), py::arg("position"), py::arg("orientation"),
             py::arg("precision"), py::arg("pos_thresh"), py::arg("orn_thresh"), py::arg("timestep"))
        .def(py::init<const VectorXd&, const VectorXd&, const VectorXd&, const VectorXd&, const MatrixXd&, const double&, const VectorXd&, const int&>(), py::arg("position"),
             py::arg("dposition"), py::arg("orientation"), py::arg("dorientation"), py::arg("precision"), py::arg("pos_thresh"), py::arg("orn_thresh"), py::arg("timestep"));
    py::class_<sys::AngularKeypoint, sys::Keypoint, std::shared_ptr<sys::AngularKeypoint>>(m_sys, "AngularKeypoint")
        .def(py::init<const VectorXd&, const MatrixXd&, const int&>(), py::arg("position"), py::arg("precision"), py::arg("timestep"))
        .def(py::init<const VectorXd&, const VectorXd&, const MatrixXd&, const int&>(), py::arg("position"), py::arg("dposition"), py::arg("precision"), py::arg("timestep"))
        .def("get_position", &sys::AngularKeypoint::getPosition);
    py::class_<sys::AngularTimeKeypoint, sys::AngularKeypoint, std::shared_ptr<sys::AngularTimeKeypoint>>(m_sys, "AngularTimeKeypoint")
        .def(py::init<const VectorXd&, const MatrixXd&, const double&, const int&>(), py::arg("position"), py::arg("precision"), py::arg("continuous_time"), py::arg("timestep"))
        .def(py::init<const VectorXd&, const VectorXd&, const MatrixXd&, const double&, const int&>(), py::arg("position"), py::arg("dposition"), py::arg("precision"),
             py::arg("continuous_time"), py::arg("timestep"))
        .def("get_continuous_time", &sys::AngularTimeKeypoint::getContinuousTime);

    py::class_<sys::System, std::shared_ptr<sys::System>>(m_sys, "System")
        .def("diff", &sys::System::diff, py::arg("actual_state"), py::arg("k"))
        .def("diff_batch", &sys::System::diffBatch, py::arg("x"))
        .def("forward_pass_with_limits", &sys::System::forwardPassWithLimits, py::arg("xk"), py::arg("uk"), py::arg("k"))
        .def("forward_pass_batch", &sys::System::fpBatch, py::arg("u"))
        .def("cost_F", &sys::System::cost_F, py::arg("xk"))
        .def("cost_F_x", &sys::System::cost_F_x, py::arg("xk"))
        .def("cost_F_xx", &sys::System::cost_F_xx, py::arg("xk"))
        .def("cost", &sys::System::cost, py::arg("xk"), py::arg("uk"), py::arg("k"))
        .def("cost_x", &sys::System::cost_x, py::arg("xk"), py::arg("uk"), py::arg("k"))
        .def("cost_xx", &sys::System::cost_xx, py::arg("xk"), py::arg("uk"), py::arg("k"))
        .def("cost_u", &sys::System::cost_u, py::arg("xk"), py::arg("uk"), py::arg("k"))
        .def("cost_uu", &sys::System::cost_uu, py::arg("xk"), py::arg("uk"), py::arg("k"))
        .def("cost_xu", &sys::System::cost_xu, py::arg("xk"), py::arg("uk"), py::arg("k"))
        .def("cost_ux", &sys::System::cost_ux, py::arg("xk"), py::arg("uk"), py::arg("k"))
        .def("get_fx_jac", static_cast<std::tuple<VectorXd, MatrixXd> (sys::System::*)(VectorXd)>(&sys::System::getFxJac), py::arg("xk"))
        .def("forward_pass", &sys::System::forwardPass, py::arg("xk"), py::arg("uk"), py::arg("k"))
        .def("get_mu_vector", &sys::System::getMuVector, py::arg("sparse") = false)
        .def("get_Q_matrix", &sys::System::getQMatrix, py::arg("sparse") = false)
        .def("get_Rt", &sys::System::getRt)
        .def("get_nb_state_var", &sys::System::getNbStateVar)
        .def("get_nb_ctrl_var", &sys::System::getNbCtrlVar)
        .def("get_nb_target_var", &sys::System::getNbTargetVar)
        .def("get_nb_Q_var", &sys::System::getNbQVar)
        .def("get_horizon", &sys::System::getHorizon)
        .def("get_nb_deriv", &sys::System::getNbDeriv)
        .def("get_kp_indexes", &sys::System::getKpIndexes)
        .def("get_state", &sys::System::getState)
        .def("get_fx_jac", static_cast<std::tuple<VectorXd, MatrixXd> (sys::System::*)()>(&sys::System::getFxJac))
        .def("get_init_state", &sys::System::getInitState)
        .def("get_init_fx_state", &sys::System::getInitFoXState)
        .def("reset", &sys::System::reset);
    py::class_<sys::SequentialSystem, sys::System, std::shared_ptr<sys::SequentialSystem>>(m_sys, "SequentialSystem")
        .def(py::init<const SimPtr&, const std::vector<std::shared_ptr<sys::System>>&, const VectorXd&, int, int>(), py::arg("r"), py::arg("systems"), py::arg("RtDiag"),
             py::arg("horizon"), py::arg("nb_deriv"));
    py::class_<sys::PosOrnPlannerSys, sys::System, std::shared_ptr<sys::PosOrnPlannerSys>>(m_sys, "PosOrnPlannerSys")
        .def(py::init<const SimPtr&, const KpList&, const VectorXd&, const VectorXd&, const VectorXd&, const VectorXd&, const VectorXd&, int, int, double>(), py::arg("r"),
             py::arg("keypoints"), py::arg("RtDiag"), py::arg("qMax"), py::arg("qMin"), py::arg("dqMax"), py::arg("dqMin"), py::arg("horizon"), py::arg("nb_deriv"), py::arg("dt"))
        .def(py::init<const SimPtr&, const KpList&, const VectorXd&, const VectorXd&, const VectorXd&, int, int, double>(), py::arg("r"), py::arg("keypoints"), py::arg("RtDiag"),
             py::arg("qMax"), py::arg("qMin"), py::arg("horizon"), py::arg("nb_deriv"), py::arg("dt"))
        .def(py::init<const SimPtr&, const KpList&, const VectorXd&, int, int, double>(), py::arg("r"), py::arg("keypoints"), py::arg("RtDiag"), py::arg("horizon"),
             py::arg("nb_deriv"), py::arg("dt"));
    py::class_<sys::PosOrnTimePlannerSys, sys::System, std::shared_ptr<sys::PosOrnTimePlannerSys>>(m_sys, "PosOrnTimePlannerSys")
        .def(py::init<const SimPtr&, const KpList&, const VectorXd&, const VectorXd&, const VectorXd&, const VectorXd&, const VectorXd&, int, int>(), py::arg("r"), py::arg("keypoints"),
             py::arg("RtDiag"), py::arg("qMax"), py::arg("qMin"), py::arg("dqMax"), py::arg("dqMin"), py::arg("horizon"), py::arg("nb_deriv"))
        .def(py::init<const SimPtr&, const KpList&, const VectorXd&, const VectorXd&, const VectorXd&, int, int>(), py::arg("r"), py::arg("keypoints"), py::arg("RtDiag"), py::arg("qMax"),
             py::arg("qMin"), py::arg("horizon"), py::arg("nb_deriv"))
        .def(py::init<const SimPtr&, const KpList&, const VectorXd&, int, int>(), py::arg("r"), py::arg("keypoints"), py::arg("RtDiag"), py::arg("horizon"), py::arg("nb_deriv"))
        ;
    py::class_<sys::JointSpacePlannerSys, sys::System, std::shared_ptr<sys::JointSpacePlannerSys>>(m_sys, "JointSpacePlannerSys")
        .def(py::init<const SimPtr&, const KpList&, const VectorXd&, const VectorXd&, const VectorXd&, const VectorXd&, const VectorXd&, int, int, double>(), py::arg("r"),
             py::arg("keypoints"), py::arg("RtDiag"), py::arg("qMax"), py::arg("qMin"), py::arg("dqMax"), py::arg("dqMin"), py::arg("horizon"), py::arg("nb_deriv"), py::arg("dt"))
        .def(py::init<const SimPtr&, const KpList&, const VectorXd&, const VectorXd&, const VectorXd&, int, int, double>(), py::arg("r"), py::arg("keypoints"), py::arg("RtDiag"),
             py::arg("qMax"), py::arg("qMin"), py::arg("horizon"), py::arg("nb_deriv"), py::arg("dt"))
        .def(py::init<const SimPtr&, const KpList&, const VectorXd&, int, int, double>(), py::arg("r"), py::arg("keypoints"), py::arg("RtDiag"), py::arg("horizon"),
             py::arg("nb_deriv"), py::arg("dt"));
    py::class_<sys::JointSpaceTimePlannerSys, sys::System, std::shared_ptr<sys::JointSpaceTimePlannerSys>>(m_sys, "JointSpaceTimePlannerSys")
        .def(py::init<const SimPtr&, const KpList&, const VectorXd&, const VectorXd&, const VectorXd&, const VectorXd&, const VectorXd&, int, int>(), py::arg("r"), py::arg("keypoints"),
             py::arg("RtDiag"), py::arg("qMax"), py::arg("qMin"), py::arg("dqMax"), py::arg("dqMin"), py::arg("horizon"), py::arg("nb_deriv"))
        .def(py::init<const SimPtr&, const KpList&, const VectorXd&, const VectorXd&, const VectorXd&, int, int>(), py::arg("r"), py::arg("keypoints"), py::arg("RtDiag"), py::arg("qMax"),
             py::arg("qMin"), py::arg("horizon"), py::arg("nb_deriv"))
        .def(py::init<const SimPtr&, const KpList&, const VectorXd&, int, int>(), py::arg("r"), py::arg("keypoints"), py::arg("RtDiag"), py::arg("horizon"), py::arg("nb_deriv"));

    // ---------------------------------------------------------------- utils (bindings.cpp:872-907); bound before the solvers, whose
    // default arguments mention CallBackMessage
    py::module m_ut = m.def_submodule("utils");
    py::class_<CallBackMessage>(m_ut, "CallBackMessage");
    py::class_<PythonCallbackMessage, CallBackMessage>(m_ut, "PythonCallbackMessage").def(py::init<>()).def_readonly("messages", &PythonCallbackMessage::messages);
    py::module m_sd = m_ut.def_submodule("Sd");  // bindings.cpp:891-898
    m_sd.def("logMap", &Sd::logMap, py::arg("base"), py::arg("y"));
    m_sd.def("expMap", &Sd::expMap, py::arg("base"), py::arg("u"));
    m_sd.def("distance", &Sd::distance, py::arg("x"), py::arg("y"));
    m_sd.def("transport", &Sd::transport, py::arg("v"), py::arg("base1"), py::arg("base2"));
    m_sd.def("dquat_to_w_jac", &Sd::dQuatToDxJac, py::arg("q"));
    py::module m_prim = m_ut.def_submodule("primitives");
    m_prim.def("build_psi_RBF", &buildPsiRBF, py::arg("dim"), py::arg("K"));
    m_prim.def("build_psi_bernstein", &buildPsiBernstein, py::arg("dim"), py::arg("K"));
    m_prim.def("build_psi_unitstep", &buildPsiUnitstep, py::arg("dim"), py::arg("K"));
    m_prim.def("build_psi_sawtooth", &buildPsiSawtooth, py::arg("dim"), py::arg("K"));
    m_prim.def("build_psi_linear", &buildPsiLinear, py::arg("dim"), py::arg("K"));

    // ---------------------------------------------------------------- solver (bindings.cpp:695-869)
    py::module m_sol = m.def_submodule("solver");
    py::class_<solver::Constraint>(m_sol, "Constraint").def(py::init<>()).def_readwrite("A", &solver::Constraint::A).def_readwrite("b", &solver::Constraint::b);
    const auto no_cb = static_cast<CallBackMessage*>(nullptr);
    using Devs = std::vector<int>;
    py::class_<solver::AL_ILQR>(m_sol, "AL_ILQR")
        .def(py::init<const std::shared_ptr<sys::System>&, const std::vector<solver::Constraint>&, const std::vector<VectorXd>&>(), py::arg("s"), py::arg("inequality"),
             py::arg("initLambda"))
        .def("solve", &solver::AL_ILQR::solve, py::arg("U0"), py::arg("nb_iter"), py::arg("lag_update_step"), py::arg("penalty"), py::arg("scaling_factor"),
             py::arg("line_search") = true, py::arg("early_stop") = false, py::arg("cb") = no_cb, py::call_guard<py::gil_scoped_release>())
        .def("solve_batch",
             [](solver::AL_ILQR& self, py::object q0, py::object kp_states, py::object U0, int nb_iter, int lag_update_step, double penalty,
                double scaling_factor, bool line_search, bool early_stop, py::object dq0, const Devs& devices) {
                 const std::shared_ptr<sys::System>& s = self.system();
                 ArrayInputs a = array_inputs(q0, kp_states, U0, dq0, s, long(s->getHorizon() - 1) * s->getNbCtrlVar());
                 solver::BatchView v;
                 {
                     py::gil_scoped_release rel;
                     v = self.solveBatch(a.in, nb_iter, lag_update_step, penalty, scaling_factor, line_search, early_stop, devices);
                 }
                 return view_dict(v, self.resultOwner(), false);
             },
             py::arg("q0") = py::none(), py::arg("kp_states") = py::none(), py::arg("U0") = py::none(), py::arg("nb_iter") = 10, py::arg("lag_update_step") = 3,
             py::arg("penalty") = 1.0, py::arg("scaling_factor") = 2.0, py::arg("line_search") = true, py::arg("early_stop") = false, py::arg("dq0") = py::none(),
             py::arg("devices") = Devs{0}, kBatchDoc)
        .def("cache_stats", &solver::AL_ILQR::cacheStats);
    py::class_<solver::BatchILQR>(m_sol, "BatchILQR")
        .def(py::init<const std::shared_ptr<sys::System>&>(), py::arg("s"))
        .def(py::init<const std::shared_ptr<sys::System>&, const MatrixXd&>(), py::arg("s"), py::arg("Q"))
        .def("solve", &solver::BatchILQR::solve, py::arg("nb_iter"), py::arg("u0"), py::arg("early_stop") = false, py::arg("cb") = no_cb, py::call_guard<py::gil_scoped_release>())
        .def("solve_batch",
             [](solver::BatchILQR& self, py::object q0, py::object kp_states, py::object u0, int nb_iter, bool early_stop, py::object dq0,
                const Devs& devices) {
                 const std::shared_ptr<sys::System>& s = self.system();
                 ArrayInputs a = array_inputs(q0, kp_states, u0, dq0, s, long(s->getHorizon() - 1) * s->getNbCtrlVar());
                 solver::BatchView v;
                 {
                     py::gil_scoped_release rel;
                     v = self.solveBatch(a.in, nb_iter, early_stop, devices);
                 }
                 return view_dict(v, self.resultOwner(), true);
             },
             py::arg("q0") = py::none(), py::arg("kp_states") = py::none(), py::arg("u0") = py::none(), py::arg("nb_iter") = 10, py::arg("early_stop") = false,
             py::arg("dq0") = py::none(), py::arg("devices") = Devs{0}, kBatchDoc)
        .def("cache_stats", &solver::BatchILQR::cacheStats);
    py::class_<solver::BatchILQRCP>(m_sol, "BatchILQRCP")
        .def(py::init<const std::shared_ptr<sys::System>&, const MatrixXd&, const MatrixXd&>(), py::arg("s"), py::arg("Q"), py::arg("psi"))
        .def(py::init<const std::shared_ptr<sys::System>&, const MatrixXd&>(), py::arg("s"), py::arg("psi"))
        .def("solve", &solver::BatchILQRCP::solve, py::arg("nb_iter"), py::arg("u0"), py::arg("early_stop") = false, py::arg("cb") = no_cb, py::call_guard<py::gil_scoped_release>())
        .def("solve_batch",
             [](solver::BatchILQRCP& self, py::object q0, py::object kp_states, py::object u0, int nb_iter, bool early_stop, py::object dq0,
                const Devs& devices) {
                 const std::shared_ptr<sys::System>& s = self.system();
                 ArrayInputs a = array_inputs(q0, kp_states, u0, dq0, s, long(s->getHorizon() - 1) * s->getNbCtrlVar());
                 solver::BatchView v;
                 {
                     py::gil_scoped_release rel;
                     v = self.solveBatch(a.in, nb_iter, early_stop, devices);
                 }
                 return view_dict(v, self.resultOwner(), true);
             },
             py::arg("q0") = py::none(), py::arg("kp_states") = py::none(), py::arg("u0") = py::none(), py::arg("nb_iter") = 10, py::arg("early_stop") = false,
             py::arg("dq0") = py::none(), py::arg("devices") = Devs{0}, kBatchDoc)
        .def("cache_stats", &solver::BatchILQRCP::cacheStats);
    py::class_<solver::ILQRRecursive>(m_sol, "ILQRRecursive")
        .def(py::init<const std::shared_ptr<sys::System>&>(), py::arg("s"))
        .def("solve", &solver::ILQRRecursive::solve, py::arg("U0"), py::arg("nb_iter"), py::arg("line_search") = true, py::arg("early_stop") = false, py::arg("cb") = no_cb,
             py::call_guard<py::gil_scoped_release>())
        .def("solve_batch",
             [](solver::ILQRRecursive& self, py::object q0, py::object kp_states, py::object U0, int nb_iter, bool line_search,
                bool early_stop, py::object dq0, bool gains, const Devs& devices) {
                 const std::shared_ptr<sys::System>& s = self.system();
                 ArrayInputs a = array_inputs(q0, kp_states, U0, dq0, s, long(s->getHorizon() - 1) * s->getNbCtrlVar());
                 solver::BatchView v;
                 {
                     py::gil_scoped_release rel;
                     v = self.solveBatch(a.in, nb_iter, line_search, early_stop, gains, devices);
                 }
                 return view_dict(v, self.resultOwner(), false);
             },
             py::arg("q0") = py::none(), py::arg("kp_states") = py::none(), py::arg("U0") = py::none(), py::arg("nb_iter") = 10, py::arg("line_search") = true,
             py::arg("early_stop") = false, py::arg("dq0") = py::none(), py::arg("gains") = true, py::arg("devices") = Devs{0}, kBatchDoc)
        .def("solve_systems",
             [](solver::ILQRRecursive& self, const SysList& systems, py::object U0, int nb_iter, bool line_search, bool early_stop, const Devs& devices) {
                 std::vector<std::vector<VectorXd>> U0s;
                 if (!U0.is_none()) {
                     DArr x = DArr::ensure(U0);
                     if (!x || x.ndim() != 3 || size_t(x.shape(0)) != systems.size()) throw std::invalid_argument("solve_systems: U0 must be [len(systems), horizon-1, nb_ctrl_var]");
                     for (ssize_t b = 0; b < x.shape(0); b++) {
                         U0s.emplace_back();
                         for (ssize_t k = 0; k < x.shape(1); k++) U0s.back().push_back(detail::from_buffer(x.data(b, k, 0), long(x.shape(2))));
                     }
                 }
                 solver::BatchResult r;
                 {
                     py::gil_scoped_release rel;
                     r = self.solveBatch(systems, U0s, nb_iter, line_search, early_stop, devices);
                 }
                 return result_dict(r, false);
             },
             py::arg("systems"), py::arg("U0") = py::none(), py::arg("nb_iter") = 10, py::arg("line_search") = true, py::arg("early_stop") = false, py::arg("devices") = Devs{0},
             "Batched solve of a list of system objects of one structure (they may differ in robot state, keypoint targets and precisions); results are copies.")
        .def("cache_stats", &solver::ILQRRecursive::cacheStats);
    // bindings.cpp:862-869
    py::class_<solver::LQT>(m_sol, "LQT")
        .def(py::init<const MatrixXd&, const MatrixXd&, const std::vector<MatrixXd>&, const VectorXd&, float, int>(), py::arg("A"), py::arg("B"), py::arg("Qs"), py::arg("states"),
             py::arg("rfactor") = 0.1f, py::arg("nb_deriv") = 2)
        .def("solve_DP", &solver::LQT::solveDP, py::call_guard<py::gil_scoped_release>())
        .def("solve_lin_al", &solver::LQT::solveLinAl, py::call_guard<py::gil_scoped_release>())
        .def("get_nb_states", &solver::LQT::getNbStates)
        .def("get_predicted_states", &solver::LQT::getPredictedStates)
        .def("get_command", static_cast<VectorXd (solver::LQT::*)(int)>(&solver::LQT::getCommand))
        .def("get_command", static_cast<VectorXd (solver::LQT::*)(int, const VectorXd&)>(&solver::LQT::getCommand))
        .def("get_Ps", &solver::LQT::getPs, "solve_DP's matrices in time order (private in the reference)")
        .def("get_ds", &solver::LQT::getds)
        .def_static(
            "solve_batch",
            [](const MatrixXd& A, const MatrixXd& B, DArr Qs, DArr mu, float rfactor, int nb_deriv, int device) {
                if (Qs.ndim() != 4 || mu.ndim() != 2 || Qs.shape(0) != mu.shape(0)) throw std::invalid_argument("LQT.solve_batch: Qs [nq, T, n, n] and mu [nq, T*n] expected");
                const int nq = int(Qs.shape(0)), T = int(Qs.shape(1));
                const std::vector<double> q(Qs.data(), Qs.data() + Qs.size()), muv(mu.data(), mu.data() + mu.size());
                solver::LQT::BatchOut o;
                {
                    py::gil_scoped_release rel;
                    o = solver::LQT::solveBatch(A, B, nq, T, q, muv, rfactor, nb_deriv, device);
                }
                const ssize_t n = A.cols(), m = B.cols();
                auto arr = [](const std::vector<double>& v, std::vector<ssize_t> shape) {
                    py::array_t<double> a(shape);
                    std::copy(v.begin(), v.end(), a.mutable_data());
                    return a;
                };
                py::dict d;
                d["u"] = arr(o.u, {nq, (T - 1) * m});
                d["xhat"] = arr(o.xhat, {nq, T * n});
                d["Ps"] = arr(o.Ps, {nq, T, n, n});
                d["ds"] = arr(o.ds, {nq, T, n});
                d["K"] = arr(o.K, {nq, T - 1, m, n});
                d["f"] = arr(o.f, {nq, T - 1, m});
                return d;
            },
            py::arg("A"), py::arg("B"), py::arg("Qs"), py::arg("mu"), py::arg("rfactor") = 0.1f, py::arg("nb_deriv") = 2, py::arg("device") = 0,
            "Many tracking problems of one shape in one device call: Qs [nq, T, n, n], mu [nq, T*n].");
}
