// batch.cuh — BatchILQR / BatchILQRCP path (BatchILQR.cpp:110-173, BatchILQRCP.cpp:109-174).  (stub: filled in next)
#pragma once
#include <functional>
#include <string>

#include "kin.cuh"
#include "recursive.cuh"

namespace ilqrb {

struct BatchWork {
    double* u = nullptr;  // [B][N]
};

inline int batch_run(cudaStream_t, const DevProblem&, const DevProblem*, Work&, BatchWork&, const ilqrb_batch_desc*, const ilqrb_solve_opts*, std::string& err,
                     std::function<void(int)>, std::function<void()>) {
    err = "Batch solvers not built yet";
    return ILQRB_EUNSUPPORTED;
}
inline int batch_download(cudaStream_t, const DevProblem&, BatchWork&, ilqrb_result*, int) { return ILQRB_EUNSUPPORTED; }

}  // namespace ilqrb
