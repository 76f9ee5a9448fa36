// launch_batch.cu — Batch / control-primitive Batch solvers (batch.cuh) behind the handle.
#include "batch.cuh"
#include "handle.hpp"

namespace ilqrb {

int batch_run_h(ilqrb_handle h, const ilqrb_batch_desc* bd, const ilqrb_solve_opts* opts) {
    std::string err;
    const int rc = batch_run(h->stream, h->device, h->hp, h->dp, h->W, h->BW, bd, opts, err, [h](int slot) { h->begin(PS_BATCH_FP + slot); }, [h]() { h->end(); });
    return rc ? fail(rc, err) : 0;
}
int batch_download_h(ilqrb_handle h, ilqrb_result* out, int nit) {
    std::string err;
    const int rc = batch_download(h->stream, h->hp, h->BW, out, h->W, nit, err);
    return rc ? fail(rc, err) : 0;
}
void batch_free(BatchWork& BW) { batch_free_impl(BW); }

}  // namespace ilqrb
