// batch_types.hpp — workspace description of the Batch / control-primitive Batch solvers (batch.cuh), plain C++ so that the
// handle (handle.hpp) can own one without seeing the kernels.
#pragma once
#include <vector>

#include "work.cuh"

namespace ilqrb {

constexpr int BNC = NCAND + 1;  // candidate 0 = current controls (alpha = 0), 1..11 = alpha 1, 1/2, .., 2^-10

struct BatchDev {  // POD view passed to the kernels
    // sizes
    int N = 0, Np = 0, S = 0, Sn = 0, SnP = 0, Sr = 0, nw = 0, cp = 0, chunk = 0;  // Np = N padded to 64, SnP = Sn padded to 4
    // S = distinct keypoint timesteps (row blocks of J, L, Su: Sn = S n); nk = keypoints incl. duplicates (slots of e, Q: Sr = nk r)
    int nk = 0;
    int vp[MAX_KP_STEPS];   // slot -> timestep (sorted, duplicates kept)
    int vds[MAX_KP_STEPS];  // slot -> index of that timestep in P.kp_steps
    // whole batch
    double* u = nullptr;      // [B][N]
    double* du = nullptr;     // [B][N]
    double* cc = nullptr;     // [B][BNC] candidate costs (0: cost0)
    double* Qfull = nullptr;  // [Sr][Sr] shared precision override, or nullptr
    double* psi = nullptr;    // [N][nw]
    double* psiRpsi = nullptr;  // [nw][nw]  Psi^T R Psi
    // per chunk of problems
    double* X = nullptr;    // [C][H][n]  states of the current controls (double-integrator time systems: B needs the post-step velocity)
    double* Jt = nullptr;   // [C][S][r][n]
    double* e = nullptr;    // [C][Sr]
    double* Ld = nullptr;   // [C][S][n]  diagonal of L (0 / penalty)
    double* ql = nullptr;   // [C][S][n]
    double* Wm = nullptr;   // [C][Sn][Sn]
    double* g = nullptr;    // [C][Sn]
    double* Su = nullptr;   // [C][SnP][Np]  (rows >= Sn and columns >= N are zero)
    double* T = nullptr;    // [C][SnP][Np]
    double* A = nullptr;    // [C][Np][Np]
    double* rhs = nullptr;  // [C][Np]
};
struct BatchWork : BatchDev {
    bool allocated = false;
    std::vector<void*> allocs;
};

}  // namespace ilqrb
