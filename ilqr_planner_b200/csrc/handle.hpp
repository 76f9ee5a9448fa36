// handle.hpp — the object behind ilqrb_handle (include/ilqrb.h) and the launcher interface between the C ABI
// (ilqrb_api.cu: descriptor flattening, solver sequences, result download) and the kernel families, each compiled as its
// own translation unit:
//   launch_generic.cu  recursive.cuh  rollout, keypoint derivatives, generic backward pass, line searches, accept, f(X)
//   launch_fused1.cu   fused1.cuh     single-integrator fast path (backward + forward kernels)
//   launch_fused2.cu   fused2.cuh     double-integrator fast backward pass
//   launch_time.cu     fused_time.cuh time-system fast path
//   launch_batch.cu    batch.cuh      Batch / control-primitive Batch solvers
// Host logic only orchestrates; all arithmetic of the solve path runs in CUDA kernels.  There is no CPU fallback.
#pragma once
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <string>
#include <vector>

#include "batch_types.hpp"
#include "work.cuh"

namespace ilqrb {

int fail(int code, const std::string& msg);  // sets the thread-local message of ilqrb_last_error, returns code

#define CU(call)                                                                                                      \
    do {                                                                                                              \
        cudaError_t e_ = (call);                                                                                      \
        if (e_ != cudaSuccess) return ::ilqrb::fail(ILQRB_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

enum ProfSlot { PS_ROLLOUT_INIT = 0, PS_KP_DERIVS, PS_BACKWARD, PS_LINESEARCH, PS_ACCEPT, PS_FINAL, PS_BATCH_FP, PS_BATCH_ASM, PS_BATCH_SYRK, PS_BATCH_CHOL, PS_BATCH_LS, PS_MISC, PS_FWD_ROLLOUT };

constexpr int MAX_DEVICES = 64;

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device (per-context) setting: the "already raised to" value is kept
// per device (one DynSmem object per kernel instantiation, a function-local static of its launcher), so a second handle on
// another device of the same process raises it there too.  Racing host threads at worst both make the (idempotent) call.
struct DynSmem {
    std::atomic<size_t> raised[MAX_DEVICES];
    template <class Kern>
    int ensure(Kern kern, size_t smem, int device) {
        if (smem <= 48 * 1024) return 0;
        const bool tracked = device >= 0 && device < MAX_DEVICES;
        if (!tracked || smem > raised[device].load(std::memory_order_relaxed)) {
            CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            if (tracked) raised[device].store(smem, std::memory_order_relaxed);
        }
        return 0;
    }
};

}  // namespace ilqrb

struct ilqrb_handle_s {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    ilqrb::DevProblem hp;  // host copy holding device pointers
    ilqrb::DevProblem* dp = nullptr;
    ilqrb_dims dims;
    ilqrb::Work W;
    std::vector<void*> allocs;
    int cap_iter = 0;
    int last_nb_iter = 0;
    int last_solver = 0;  // 1 recursive, 2 al, 3 batch
    double* d_fX = nullptr;
    double* d_pack = nullptr;  // staging for packed downloads
    size_t pack_elems = 0;
    double* d_alA = nullptr;
    double* d_alb = nullptr;
    double* d_lam0 = nullptr;
    size_t alA_elems = 0, alb_elems = 0, lam0_elems = 0;
    ilqrb::BatchWork BW;
    // profiling
    bool prof = false;
    std::vector<cudaEvent_t> ev_pool;
    struct Rec {
        int slot;
        cudaEvent_t a, b;
    };
    std::vector<Rec> recs;
    size_t ev_used = 0;
    double prof_ms[ILQRB_PROF_SLOTS] = {0};
    int64_t prof_n[ILQRB_PROF_SLOTS] = {0};
    int64_t launches = 0;
    bool downloaded = false;  // the last run already sent its results to the host (per group)
    bool streaming = false;   // the current solve came through ilqrb_solve_recursive_async
    // problem groups of the fast paths (one stream each)
    static constexpr int MAX_GROUPS = 8;
    int n_groups = 0;
    cudaStream_t gstream[MAX_GROUPS] = {nullptr};
    cudaEvent_t ev_start = nullptr, ev_stagger[MAX_GROUPS] = {nullptr}, ev_done[MAX_GROUPS] = {nullptr};
    int ensure_groups(int G);
    // recorded when the last solve's kernels are queued, before its device-to-host copies (ilqrb_chain_after_compute)
    cudaEvent_t ev_compute[MAX_GROUPS + 1] = {nullptr};
    int n_ev_compute = 0;

    template <class T>
    int alloc(T** p, size_t count) {
        void* q = nullptr;
        cudaError_t e = cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T));
        if (e != cudaSuccess) return ilqrb::fail(ILQRB_ENOMEM, std::string("cudaMalloc: ") + cudaGetErrorString(e));
        allocs.push_back(q);
        *p = (T*)q;
        return 0;
    }
    cudaEvent_t ev() {
        if (ev_used == ev_pool.size()) {
            cudaEvent_t e;
            cudaEventCreate(&e);
            ev_pool.push_back(e);
        }
        return ev_pool[ev_used++];
    }
    void begin(int slot) {
        launches++;
        if (!prof) return;
        Rec r{slot, ev(), ev()};
        cudaEventRecord(r.a, stream);
        recs.push_back(r);
    }
    void end() {
        if (!prof) return;
        cudaEventRecord(recs.back().b, stream);
    }
    void resolve() {
        if (recs.empty()) return;
        cudaStreamSynchronize(stream);
        for (auto& r : recs) {
            float ms = 0;
            cudaEventElapsedTime(&ms, r.a, r.b);
            prof_ms[r.slot] += ms;
            prof_n[r.slot]++;
        }
        recs.clear();
        ev_used = 0;
    }
};

namespace ilqrb {

// ---- launch_generic.cu ----------------------------------------------------------------------------------------------
int launch_init(ilqrb_handle h, int b0 = 0, int nB = -1, cudaStream_t st = nullptr);
int launch_derivs(ilqrb_handle h, int b0 = 0, int nB = -1, cudaStream_t st = nullptr);
int launch_backward_generic(ilqrb_handle h, bool al, double penalty_frozen);
int launch_linesearch(ilqrb_handle h, int line_search, bool al);
int launch_accept(ilqrb_handle h, int it, int nb_iter, int line_search, int early_stop, int al, int lag, double pen_next);
int launch_al_init(ilqrb_handle h);
int launch_final_fx(ilqrb_handle h, int b0, int nB, cudaStream_t st);  // f(X) of problems [b0, b0 + nB) into h->d_fX
int launch_rollout_jac(ilqrb_handle h, double* Jout);
int launch_kinematics(ilqrb_handle h, int robot, int nq, const double* q, const double* dq, double* x, double* quat, double* J, double* dx, double* w);
bool generic_forced();  // ILQRB_GENERIC set: A/B switch for tests and profiling

// ---- launch_time.cu -------------------------------------------------------------------------------------------------
bool fast_time_applies(ilqrb_handle h, bool al);
int launch_derivs_time(ilqrb_handle h, int b0, int nB, cudaStream_t st);
int launch_backward_time(ilqrb_handle h, int b0, int nB, cudaStream_t st);
int launch_forward_time(ilqrb_handle h, int b0, int nB, cudaStream_t st, int it, int nb_iter, int line_search, int early_stop);
// ---- launch_fused1.cu -----------------------------------------------------------------------------------------------
bool fast_n1_applies(ilqrb_handle h, bool al);
int fast_n1_wave(ilqrb_handle h);  // problems one full wave of the forward kernel holds
bool fast_n1_col_path(ilqrb_handle h);            // the backward pass runs on the column-per-lane kernels (fused1_col.cuh)
int fast_n1_default_groups(ilqrb_handle h, int B);  // problem groups for device-resident results
int launch_backward_n1(ilqrb_handle h, int b0, int nB, cudaStream_t st);
int launch_forward_n1(ilqrb_handle h, int b0, int nB, cudaStream_t st, int it, int nb_iter, int line_search, int early_stop);

// ---- launch_fused2.cu -----------------------------------------------------------------------------------------------
bool fast_n2_applies(ilqrb_handle h, bool al);
int launch_backward_n2(ilqrb_handle h, bool al, double penalty_frozen);

// ---- launch_batch.cu ------------------------------------------------------------------------------------------------
int batch_run_h(ilqrb_handle h, const ilqrb_batch_desc* bd, const ilqrb_solve_opts* opts);
int batch_download_h(ilqrb_handle h, ilqrb_result* out, int nit);
void batch_free(BatchWork& BW);

}  // namespace ilqrb
