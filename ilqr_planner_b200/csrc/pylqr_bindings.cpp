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
        messages.push_back(msg);
        py::print(msg);
    }
    std::vector<std::string> messages;
};

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

    py::class_<sys::System, std::shared_ptr<sys::System>>(m_sys, "System")
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
        .def("get_fx_jac", static_cast<std::tuple<VectorXd, MatrixXd> (sys::PosOrnTimePlannerSys::*)(VectorXd)>(&sys::PosOrnTimePlannerSys::getFxJac), py::arg("xk"))
        .def("get_fx_jac", static_cast<std::tuple<VectorXd, MatrixXd> (sys::PosOrnTimePlannerSys::*)()>(&sys::PosOrnTimePlannerSys::getFxJac));

    // ---------------------------------------------------------------- utils (bindings.cpp:872-907); bound before the solvers, whose
    // default arguments mention CallBackMessage
    py::module m_ut = m.def_submodule("utils");
    py::class_<CallBackMessage>(m_ut, "CallBackMessage");
    py::class_<PythonCallbackMessage, CallBackMessage>(m_ut, "PythonCallbackMessage").def(py::init<>()).def_readonly("messages", &PythonCallbackMessage::messages);
    py::module m_prim = m_ut.def_submodule("primitives");
    m_prim.def("build_psi_RBF", &buildPsiRBF, py::arg("dim"), py::arg("K"));
    m_prim.def("build_psi_bernstein", &buildPsiBernstein, py::arg("dim"), py::arg("K"));
    m_prim.def("build_psi_unitstep", &buildPsiUnitstep, py::arg("dim"), py::arg("K"));
    m_prim.def("build_psi_sawtooth", &buildPsiSawtooth, py::arg("dim"), py::arg("K"));
    m_prim.def("build_psi_linear", &buildPsiLinear, py::arg("dim"), py::arg("K"));

    // ---------------------------------------------------------------- solver (bindings.cpp:695-869)
    py::module m_sol = m.def_submodule("solver");
    py::class_<solver::Constraint>(m_sol, "Constraint").def(py::init<>()).def_readwrite("A", &solver::Constraint::A).def_readwrite("b", &solver::Constraint::b);
    py::class_<solver::AL_ILQR>(m_sol, "AL_ILQR")
        .def(py::init<const std::shared_ptr<sys::System>&, const std::vector<solver::Constraint>&, const std::vector<VectorXd>&>(), py::arg("s"), py::arg("inequality"),
             py::arg("initLambda"))
        .def("solve", &solver::AL_ILQR::solve, py::arg("U0"), py::arg("nb_iter"), py::arg("lag_update_step"), py::arg("penalty"), py::arg("scaling_factor"),
             py::arg("line_search") = true, py::arg("early_stop") = false, py::arg("cb") = static_cast<CallBackMessage*>(nullptr));
    py::class_<solver::BatchILQR>(m_sol, "BatchILQR")
        .def(py::init<const std::shared_ptr<sys::System>&>(), py::arg("s"))
        .def(py::init<const std::shared_ptr<sys::System>&, const MatrixXd&>(), py::arg("s"), py::arg("Q"))
        .def("solve", &solver::BatchILQR::solve, py::arg("nb_iter"), py::arg("u0"), py::arg("early_stop") = false, py::arg("cb") = static_cast<CallBackMessage*>(nullptr));
    py::class_<solver::BatchILQRCP>(m_sol, "BatchILQRCP")
        .def(py::init<const std::shared_ptr<sys::System>&, const MatrixXd&, const MatrixXd&>(), py::arg("s"), py::arg("Q"), py::arg("psi"))
        .def(py::init<const std::shared_ptr<sys::System>&, const MatrixXd&>(), py::arg("s"), py::arg("psi"))
        .def("solve", &solver::BatchILQRCP::solve, py::arg("nb_iter"), py::arg("u0"), py::arg("early_stop") = false, py::arg("cb") = static_cast<CallBackMessage*>(nullptr));
    py::class_<solver::ILQRRecursive>(m_sol, "ILQRRecursive")
        .def(py::init<const std::shared_ptr<sys::System>&>(), py::arg("s"))
        .def("solve", &solver::ILQRRecursive::solve, py::arg("U0"), py::arg("nb_iter"), py::arg("line_search") = true, py::arg("early_stop") = false,
             py::arg("cb") = static_cast<CallBackMessage*>(nullptr));
}
