// kin_lean.cuh — register-only kinematics for the single-integrator fast path (fused1.cuh).
//
// Same mathematics as kin.cuh (KDLRobot.cpp:83-115 via orocos_kdl JntToCart / JntToJac / GetQuaternion, 2DRobot.cpp,
// TransformedSimulationInterface.cpp:53-103) arranged for a warp that evaluates many poses at once:
//   - the chain is the COMPACTED chain built at create time (runs of fixed segments folded into the tip frame of the
//     segment before them: the same frame product, associated differently);
//   - joint values are fetched through a functor (no locally indexed arrays -> no local memory);
//   - instead of KDL's incremental changeRefPoint sweep over all columns, ONE Jacobian column is formed in closed form
//     from the joint's world axis z and a point o on it:  Jv = z x (p_tip - o), Jw = z  (ROT) ;  Jv = z, Jw = 0 (TRANS).
//     Each lane asks for the column it owns.
// Differences to the KDL operation order are rounding-level (~1e-16) and covered by the 1e-9 parity bar.
#pragma once
#include "kin.cuh"

namespace ilqrb {

struct LeanPose {
    double p[3];
    double R[9];
    double z[3];  // world axis of the requested joint
    double o[3];  // world point on that axis
    int jt;       // joint type of the requested joint (FIXED = no such joint)
};

// q(j): value of joint j of this robot.  cj: joint whose column is wanted (-1: none)
// TR <- TR * M, one row at a time (no second 3x3 buffer)
__device__ __forceinline__ void mat3_rmul_inplace(double* TR, const double* M) {
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const double a = TR[3 * i], b = TR[3 * i + 1], c = TR[3 * i + 2];
        TR[3 * i] = a * M[0] + b * M[3] + c * M[6];
        TR[3 * i + 1] = a * M[1] + b * M[4] + c * M[7];
        TR[3 * i + 2] = a * M[2] + b * M[5] + c * M[8];
    }
}

template <class QF>
__device__ __forceinline__ void lean_kdl_fk(const ilqrb_segment* __restrict__ segs, int nseg, QF q, int cj, LeanPose& o) {
    double TR[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, Tp[3] = {0, 0, 0};
    o.jt = ILQRB_JOINT_FIXED;
    o.z[0] = o.z[1] = o.z[2] = 0.0;
    o.o[0] = o.o[1] = o.o[2] = 0.0;
    int j = 0;
    for (int s = 0; s < nseg; s++) {
        const ilqrb_segment& sg = segs[s];
        const int jt = sg.joint_type;
        // pose = (Rj * tipR, Rj * tip_p + pj) ;  T <- T * pose :  Tp += TR * (Rj tip_p + pj) ; TR <- (TR Rj) tipR
        double ps[3] = {sg.tip_p[0], sg.tip_p[1], sg.tip_p[2]};
        if (jt != ILQRB_JOINT_FIXED) {
            const double ax[3] = {sg.axis[0], sg.axis[1], sg.axis[2]};
            const double og[3] = {sg.origin[0], sg.origin[1], sg.origin[2]};
            if (j == cj) {
                mat3_vec(TR, ax, o.z);
                double t3[3];
                mat3_vec(TR, og, t3);
                o.o[0] = t3[0] + Tp[0];
                o.o[1] = t3[1] + Tp[1];
                o.o[2] = t3[2] + Tp[2];
                o.jt = jt;
            }
            const double qj = q(j);
            if (jt == ILQRB_JOINT_ROT) {
                double Rj[9], t3[3];
                kdl_rot2(ax, qj, Rj);
                mat3_vec(Rj, ps, t3);
                ps[0] = t3[0] + og[0];
                ps[1] = t3[1] + og[1];
                ps[2] = t3[2] + og[2];
                double w3[3];
                mat3_vec(TR, ps, w3);
                Tp[0] += w3[0];
                Tp[1] += w3[1];
                Tp[2] += w3[2];
                mat3_rmul_inplace(TR, Rj);
            } else {
                ps[0] += og[0] + ax[0] * qj;
                ps[1] += og[1] + ax[1] * qj;
                ps[2] += og[2] + ax[2] * qj;
                double w3[3];
                mat3_vec(TR, ps, w3);
                Tp[0] += w3[0];
                Tp[1] += w3[1];
                Tp[2] += w3[2];
            }
            j++;
        } else {
            double w3[3];
            mat3_vec(TR, ps, w3);
            Tp[0] += w3[0];
            Tp[1] += w3[1];
            Tp[2] += w3[2];
        }
        mat3_rmul_inplace(TR, sg.tip_R);
    }
#pragma unroll
    for (int i = 0; i < 9; i++) o.R[i] = TR[i];
    o.p[0] = Tp[0];
    o.p[1] = Tp[1];
    o.p[2] = Tp[2];
}

// Task vector fx = [p, quat] (POSITION: [p]) of task tk and, if cj >= 0, column cj (robot-local joint index) of its
// 6 x dof geometric Jacobian (rows v then w; POSITION uses the v rows).  nb_deriv == 1 only (no velocity part).
// Joint-space tasks (ILQRB_TASK_JOINT, first order): fx = q and column cj of the identity — the caller then provides 7 rows for Jc.
// RR: rows of Jc the caller provides; the joint-space branch exists in the RR = 7 instantiations only (systems with a joint-space task).
template <int RR = 6, class QF>
__device__ __forceinline__ void lean_task_eval(const DevProblem& P, const DevTask& tk, QF q, int cj, double* fx, double* Jc) {
    const DevRobot& rb = P.robots[tk.robot];
    if constexpr (RR >= 7) {
        if (tk.kind == ILQRB_TASK_JOINT) {  // JointSpacePlannerSys.cpp:77-81
#pragma unroll
            for (int a = 0; a < 7; a++) {
                fx[a] = (a < rb.dof) ? q(a) : 0.0;
                Jc[a] = (a == cj) ? 1.0 : 0.0;
            }
            return;
        }
    }
#pragma unroll
    for (int a = 0; a < 6; a++) Jc[a] = 0.0;
    if (rb.kind == ILQRB_ROBOT_2D) {
        // 2DRobot.cpp:28-38,50-61: absolute-angle FK, forward-difference Jacobian with delta = pi * 1e-3
        const double dl = M_PI * 1e-3;
        double x0 = 0, x1 = 0, b0 = 0, b1 = 0;
        for (int i = 0; i < rb.dof; i++) {
            const double qi = q(i);
            double s, c;
            sincos(qi, &s, &c);
            x0 += rb.lengths[i] * c;
            x1 += rb.lengths[i] * s;
            if (i == cj) sincos(qi + dl, &s, &c);
            b0 += rb.lengths[i] * c;
            b1 += rb.lengths[i] * s;
        }
        fx[0] = x0;
        fx[1] = x1;
        if (P.sys_kind != ILQRB_SYS_POSITION) {
            fx[2] = 0;
            fx[3] = 1;
            fx[4] = fx[5] = fx[6] = 0;
        }
        if (cj >= 0 && cj < rb.dof) {
            Jc[0] = (b0 - x0) / dl;
            Jc[1] = (b1 - x1) / dl;
        }
        return;
    }
    LeanPose e;
    lean_kdl_fk(rb.lean, rb.n_lean, q, cj, e);
    double Jv[3] = {0, 0, 0}, Jw[3] = {0, 0, 0};
    if (e.jt == ILQRB_JOINT_ROT) {
        const double dpv[3] = {e.p[0] - e.o[0], e.p[1] - e.o[1], e.p[2] - e.o[2]};
        cross3(e.z, dpv, Jv);
        Jw[0] = e.z[0];
        Jw[1] = e.z[1];
        Jw[2] = e.z[2];
    } else if (e.jt == ILQRB_JOINT_TRANS) {
        Jv[0] = e.z[0];
        Jv[1] = e.z[1];
        Jv[2] = e.z[2];
    }
    double quat[4];
    kdl_get_quaternion(e.R, quat);
    if (tk.has_transform) {  // TransformedSimulationInterface.cpp:53-103
        const double* T = tk.T;
        const double R[9] = {T[0], T[1], T[2], T[4], T[5], T[6], T[8], T[9], T[10]};
        const double dd[3] = {e.p[0] - T[3], e.p[1] - T[7], e.p[2] - T[11]};
        double o3[3];
        mat3T_vec(R, dd, o3);
        e.p[0] = o3[0];
        e.p[1] = o3[1];
        e.p[2] = o3[2];
        double Rb[9], Rd[9];
        eig_quat_to_rot(quat, Rb);
        const double Rt[9] = {R[0], R[3], R[6], R[1], R[4], R[7], R[2], R[5], R[8]};
        mat3_mul(Rt, Rb, Rd);
        eig_rot_to_quat(Rd, quat);
        mat3T_vec(R, Jv, o3);
        Jv[0] = o3[0];
        Jv[1] = o3[1];
        Jv[2] = o3[2];
        mat3T_vec(R, Jw, o3);
        Jw[0] = o3[0];
        Jw[1] = o3[1];
        Jw[2] = o3[2];
    }
    fx[0] = e.p[0];
    fx[1] = e.p[1];
    fx[2] = e.p[2];
    if (P.sys_kind != ILQRB_SYS_POSITION) {
        fx[3] = quat[0];
        fx[4] = quat[1];
        fx[5] = quat[2];
        fx[6] = quat[3];
    }
    Jc[0] = Jv[0];
    Jc[1] = Jv[1];
    Jc[2] = Jv[2];
    Jc[3] = Jw[0];
    Jc[4] = Jw[1];
    Jc[5] = Jw[2];
}

// FK sweep that also accumulates the tip twist of joint rates dq(j) without forming the Jacobian:
//   w = sum_j dq_j z_j (ROT) ,  v = w x p_tip - sum_j dq_j (z_j x o_j) + sum_j dq_j z_j (TRANS)
// which is J dq with the closed-form columns above (KDLRobot.cpp:104-105: dx = Jt dq, w = Jr dq).
template <class QF, class DQF>
__device__ __forceinline__ void lean_kdl_fk_twist(const ilqrb_segment* __restrict__ segs, int nseg, QF q, DQF dq, double* p, double* R, double* v, double* w) {
    double TR[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, Tp[3] = {0, 0, 0};
    double ws[3] = {0, 0, 0}, ms[3] = {0, 0, 0}, ls[3] = {0, 0, 0};
    int j = 0;
    for (int s = 0; s < nseg; s++) {
        const ilqrb_segment& sg = segs[s];
        const int jt = sg.joint_type;
        double ps[3] = {sg.tip_p[0], sg.tip_p[1], sg.tip_p[2]};
        if (jt != ILQRB_JOINT_FIXED) {
            const double ax[3] = {sg.axis[0], sg.axis[1], sg.axis[2]};
            const double og[3] = {sg.origin[0], sg.origin[1], sg.origin[2]};
            double z[3], o3[3];
            mat3_vec(TR, ax, z);
            mat3_vec(TR, og, o3);
            o3[0] += Tp[0];
            o3[1] += Tp[1];
            o3[2] += Tp[2];
            const double qj = q(j), dqj = dq(j);
            if (jt == ILQRB_JOINT_ROT) {
                double zo[3];
                cross3(z, o3, zo);
                for (int a = 0; a < 3; a++) {
                    ws[a] += dqj * z[a];
                    ms[a] += dqj * zo[a];
                }
                double Rj[9], t3[3];
                kdl_rot2(ax, qj, Rj);
                mat3_vec(Rj, ps, t3);
                ps[0] = t3[0] + og[0];
                ps[1] = t3[1] + og[1];
                ps[2] = t3[2] + og[2];
                double w3[3];
                mat3_vec(TR, ps, w3);
                Tp[0] += w3[0];
                Tp[1] += w3[1];
                Tp[2] += w3[2];
                mat3_rmul_inplace(TR, Rj);
            } else {
                for (int a = 0; a < 3; a++) ls[a] += dqj * z[a];
                ps[0] += og[0] + ax[0] * qj;
                ps[1] += og[1] + ax[1] * qj;
                ps[2] += og[2] + ax[2] * qj;
                double w3[3];
                mat3_vec(TR, ps, w3);
                Tp[0] += w3[0];
                Tp[1] += w3[1];
                Tp[2] += w3[2];
            }
            j++;
        } else {
            double w3[3];
            mat3_vec(TR, ps, w3);
            Tp[0] += w3[0];
            Tp[1] += w3[1];
            Tp[2] += w3[2];
        }
        mat3_rmul_inplace(TR, sg.tip_R);
    }
    for (int i = 0; i < 9; i++) R[i] = TR[i];
    double wp[3];
    cross3(ws, Tp, wp);
    for (int a = 0; a < 3; a++) {
        p[a] = Tp[a];
        w[a] = ws[a];
        v[a] = wp[a] - ms[a] + ls[a];
    }
}

// d^T Q d of task tk at keypoint kp for task vector fx   (System.cpp:216-220)
__device__ __forceinline__ double lean_task_cost(const DevProblem& P, const DevTask& tk, int b, int kp, const double* fx, double* res) {
    const double* tg = tk.kp_target + ((size_t)b * tk.n_keypoints + kp) * ILQRB_KP_STRIDE;
    task_residual(P, tk, fx, tg, res, task_deadzone(tk, kp));
    const int r = tk.r_t;
    const double* Q = tk.kp_precision + ((size_t)(tk.precision_per_problem ? b : 0) * tk.n_keypoints + kp) * r * r;
    double dQd = 0;
    for (int j = 0; j < r; j++) {
        double v = 0;
        for (int i = 0; i < r; i++) v += res[i] * Q[i * r + j];
        dQd += v * res[j];
    }
    return dQd;
}

// System::cost at timestep k (System.cpp:213-234) like kin.cuh stage_cost, with the lean sweeps for KDL robots under a
// PosOrn(-Time) system (1st and 2nd order); anything else goes to stage_cost.  x: state (n), u: control or nullptr.
template <int N>
__device__ __forceinline__ double lean_stage_cost(const DevProblem& P, int b, const double* x, const double* u, int k) {
    bool lean_ok = P.sys_kind != ILQRB_SYS_POSITION;
    for (int t = 0; t < P.n_tasks; t++) lean_ok = lean_ok && P.robots[P.tasks[t].robot].kind == ILQRB_ROBOT_KDL && P.tasks[t].kind == ILQRB_TASK_POSE;
    if (!lean_ok) return stage_cost<N>(P, b, x, u, k);
    double total = 0;
    const int dof = P.dof;
    for (int t = 0; t < P.n_tasks; t++) {
        const DevTask& tk = P.tasks[t];
        double c = 0;
        const int kp = task_kp_at(tk, k);
        if (kp >= 0) {
            const DevRobot& rb = P.robots[tk.robot];
            const double* qb = x + tk.joint_offset;
            const double* vb = x + dof + tk.joint_offset;
            const bool second = P.nb_deriv == 2;
            double p[3], R[9], v[3], w[3], quat[4];
            lean_kdl_fk_twist(rb.lean, rb.n_lean, [&](int j) { return qb[j]; }, [&](int j) { return second ? vb[j] : 0.0; }, p, R, v, w);
            kdl_get_quaternion(R, quat);
            if (tk.has_transform) {  // TransformedSimulationInterface.cpp:53-103
                const double* T = tk.T;
                const double Rt[9] = {T[0], T[1], T[2], T[4], T[5], T[6], T[8], T[9], T[10]};
                const double dd[3] = {p[0] - T[3], p[1] - T[7], p[2] - T[11]};
                double o3[3];
                mat3T_vec(Rt, dd, o3);
                p[0] = o3[0], p[1] = o3[1], p[2] = o3[2];
                double Rb[9], Rd[9];
                eig_quat_to_rot(quat, Rb);
                const double RtT[9] = {Rt[0], Rt[3], Rt[6], Rt[1], Rt[4], Rt[7], Rt[2], Rt[5], Rt[8]};
                mat3_mul(RtT, Rb, Rd);
                eig_rot_to_quat(Rd, quat);
                mat3T_vec(Rt, v, o3);
                v[0] = o3[0], v[1] = o3[1], v[2] = o3[2];
                mat3T_vec(Rt, w, o3);
                w[0] = o3[0], w[1] = o3[1], w[2] = o3[2];
            }
            double fx[MAX_NTT], res[MAX_RT];
            fx[0] = p[0], fx[1] = p[1], fx[2] = p[2];
            fx[3] = quat[0], fx[4] = quat[1], fx[5] = quat[2], fx[6] = quat[3];
            int o = 7;
            if (second) {  // SimulationInterface::getEEAngVelQuat = (.5 * H(quat)^T) * w   (SimulationInterface.cpp:69-73)
                fx[7] = v[0], fx[8] = v[1], fx[9] = v[2];
                fx[10] = (.5 * -quat[1]) * w[0] + (.5 * -quat[2]) * w[1] + (.5 * -quat[3]) * w[2];
                fx[11] = (.5 * quat[0]) * w[0] + (.5 * quat[3]) * w[1] + (.5 * -quat[2]) * w[2];
                fx[12] = (.5 * -quat[3]) * w[0] + (.5 * quat[0]) * w[1] + (.5 * quat[1]) * w[2];
                fx[13] = (.5 * quat[2]) * w[0] + (.5 * -quat[1]) * w[1] + (.5 * quat[0]) * w[2];
                o = 14;
            }
            if (P.time) fx[o] = x[P.n - 1];
            c += lean_task_cost(P, tk, b, kp, fx, res);
            double uRu = 0;
            if (u)
                for (int j = 0; j < P.m; j++) uRu += (u[j] * tk.r_diag[j]) * u[j];
            c = c + uRu;
        }
        if (tk.limits_set) {
            double qLq = 0;
            for (int i = 0; i < P.n; i++) {
                bool act;
                double qi = limit_q(tk, i, x[i], act);
                if (act) qLq += (qi * 1.0) * qi;
            }
            c += qLq;
        }
        total += c;
    }
    return total;
}

}  // namespace ilqrb
