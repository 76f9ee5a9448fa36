// launch_fused1.cu — launchers of the single-integrator fast path (fused1.cuh): k_backward_n1 / k_backward_n1_blocks / k_forward_n1.
#include <cstdlib>
#include <cstring>

#include "fused1_col.cuh"
#include "fused1_fwd.cuh"
#include "handle.hpp"

namespace ilqrb {

bool fast_n1_applies(ilqrb_handle h, bool al) {
    const DevProblem& P = h->hp;
    for (int t = 0; t < P.n_tasks; t++)  // the forward kernel holds 7 residual rows for a joint-space task
        if (P.tasks[t].kind == ILQRB_TASK_JOINT && P.robots[P.tasks[t].robot].dof > 7) return false;
    return !generic_forced() && !al && P.nb_deriv == 1 && !P.time && P.n == P.m && P.n + 1 <= 16 &&
           (size_t)fwd_n1_smem_doubles(P.H, h->W.ns, P.n_kp_steps) * sizeof(double) <= 200 * 1024;
}
// resident problems of one full wave of the forward kernel: 28 warps per SM with one state tile, 12 with two
int fast_n1_wave(ilqrb_handle h) { return 148 * (h->hp.n + 1 <= 8 ? 28 : 12); }

namespace {

template <int NT, int DOF, int WARPS, int MINB>
int launch_backward_n1_t(ilqrb_handle h, int b0, int nB, cudaStream_t st) {
    const size_t smem = (size_t)WARPS * BwdN1Smem<NT>::PER_WARP * sizeof(double);
    auto kern = k_backward_n1<NT, DOF, WARPS, MINB>;
    static DynSmem attr;
    int rc = attr.ensure(kern, smem, h->device);
    if (rc) return rc;
    kern<<<(nB + WARPS - 1) / WARPS, 32 * WARPS, smem, st>>>(h->dp, h->W, b0, nB);
    return 0;
}
// four problems per warp, a column of the augmented matrices per lane (fused1_col.cuh); n = m = 7, at most one limit-carrying task
template <int N>
int launch_backward_n1_col_t(ilqrb_handle h, int b0, int nB, cudaStream_t st) {
    const size_t smem = (size_t)BwdColSmem::PER_WARP * sizeof(double);
    auto kern = k_backward_n1_col<N, 1>;  // 1 / 2 / 4 warps per block measured alike
    static DynSmem attr;
    int rc = attr.ensure(kern, smem, h->device);
    if (rc) return rc;
    kern<<<(nB + 3) / 4, 32, smem, st>>>(h->dp, h->W, b0, nB);
    return 0;
}
int backward_form() {  // A/B switch: ILQRB_BWD = warp (one problem per warp, fused1.cuh) | col (four per warp, fused1_col.cuh; default)
    static const int form = [] {
        const char* e = getenv("ILQRB_BWD");
        return (e && !strcmp(e, "warp")) ? 0 : 1;
    }();
    return form;
}
// blocks: the state splits into diagonal blocks of 7 joints each (find_state_blocks), one group of lanes per block
bool col_backward_applies(const DevProblem& P, bool blocks) {
    if (backward_form() == 0 || P.n != P.m || (!blocks && (P.n < 2 || P.n > 7))) return false;
    for (int i = 0; i < P.n; i++) {  // a lane checks its state against ONE sub-system's limits
        int nlim = 0;
        for (int t = 0; t < P.n_tasks; t++) nlim += (P.tasks[t].limits_set && P.tasks[t].limit_weight[i] != 0) ? 1 : 0;
        if (nlim > 1) return false;
    }
    return true;
}
template <int WARPS>
int launch_backward_n1_col_blocks_t(ilqrb_handle h, int b0, int nB, cudaStream_t st, const StateBlocks& sb) {
    const size_t smem = (size_t)WARPS * BwdColSmem::PER_WARP * sizeof(double);
    auto kern = k_backward_n1_col_blocks<7, WARPS>;
    static DynSmem attr;
    int rc = attr.ensure(kern, smem, h->device);
    if (rc) return rc;
    const int warps = (nB * sb.nblk + 3) / 4;
    kern<<<(warps + WARPS - 1) / WARPS, 32 * WARPS, smem, st>>>(h->dp, h->W, b0, nB, sb);
    return 0;
}
// Diagonal blocks of the state that no task couples: every task reads the joints [joint_offset, joint_offset + dof of its
// robot) only, so l_x / l_xx (and with R diagonal, A = I, B = dt I the whole Riccati recursion) are block diagonal over
// the connected groups of those ranges.  Returns nblk >= 2 only if the groups tile [0, n) and fit one 8x8 tile each.
StateBlocks find_state_blocks(const DevProblem& P) {
    StateBlocks sb{};
    sb.nblk = 0;
    static const bool off_switch = getenv("ILQRB_NO_BLOCKS") != nullptr;  // A/B switch for tests and profiling
    if (off_switch || P.nb_deriv != 1 || P.time || P.n != P.m || P.n_tasks < 2) return sb;
    int lo[ILQRB_MAX_TASKS], hi[ILQRB_MAX_TASKS], cnt = 0;
    for (int ti = 0; ti < P.n_tasks; ti++) {
        const int a = P.tasks[ti].joint_offset, e = a + P.robots[P.tasks[ti].robot].dof;
        int k = 0;
        for (; k < cnt; k++)
            if (a < hi[k] && lo[k] < e) break;  // overlaps group k
        if (k == cnt) {
            lo[cnt] = a;
            hi[cnt] = e;
            cnt++;
        } else {
            lo[k] = std::min(lo[k], a);
            hi[k] = std::max(hi[k], e);
        }
    }
    for (bool merged = true; merged;) {  // groups that grew into one another
        merged = false;
        for (int i = 0; i < cnt && !merged; i++)
            for (int j = i + 1; j < cnt && !merged; j++)
                if (lo[i] < hi[j] && lo[j] < hi[i]) {
                    lo[i] = std::min(lo[i], lo[j]);
                    hi[i] = std::max(hi[i], hi[j]);
                    lo[j] = lo[cnt - 1];
                    hi[j] = hi[cnt - 1];
                    cnt--;
                    merged = true;
                }
    }
    if (cnt < 2 || cnt > 4) return sb;
    for (int i = 0; i < cnt; i++)  // sort by offset
        for (int j = i + 1; j < cnt; j++)
            if (lo[j] < lo[i]) {
                std::swap(lo[i], lo[j]);
                std::swap(hi[i], hi[j]);
            }
    int at = 0;
    for (int i = 0; i < cnt; i++) {
        if (lo[i] != at || hi[i] - lo[i] > 7) return sb;
        at = hi[i];
    }
    if (at != P.n) return sb;
    sb.nblk = cnt;
    for (int i = 0; i < cnt; i++) {
        sb.off[i] = lo[i];
        sb.size[i] = hi[i] - lo[i];
    }
    return sb;
}
template <int DOF, int WARPS, int MINB>
int launch_backward_n1_blocks_t(ilqrb_handle h, int b0, int nB, cudaStream_t st, const StateBlocks& sb) {
    const size_t smem = (size_t)WARPS * BwdN1Smem<1>::PER_WARP * sizeof(double);
    auto kern = k_backward_n1_blocks<DOF, WARPS, MINB>;
    static DynSmem attr;
    int rc = attr.ensure(kern, smem, h->device);
    if (rc) return rc;
    const int warps = nB * sb.nblk;
    kern<<<(warps + WARPS - 1) / WARPS, 32 * WARPS, smem, st>>>(h->dp, h->W, b0, nB, sb);
    return 0;
}
template <int NT, int DOF, int MINB>
int launch_forward_n1_t(ilqrb_handle h, int b0, int nB, cudaStream_t st, int it, int nb_iter, int line_search, int early_stop) {
    const DevProblem& P = h->hp;
    const size_t smem = (size_t)fwd_n1_smem_doubles(P.H, h->W.ns, P.n_kp_steps) * sizeof(double);
    auto kern = k_forward_n1<NT, DOF, MINB>;
    static DynSmem attr;
    int rc = attr.ensure(kern, smem, h->device);
    if (rc) return rc;
    kern<<<nB, 32, smem, st>>>(h->dp, h->W, b0, it, nb_iter, line_search, early_stop);
    return 0;
}

// two-kernel forward pass (fused1_fwd.cuh)
template <int NT, int DOF, int RING, int SPS, int WARPS, int MINB>
int launch_fwd_rollout_t(ilqrb_handle h, int b0, int nB, cudaStream_t st) {
    const size_t smem = fwd_ring_smem_bytes<RING, SPS>(h->hp.m, h->W.ns, WARPS);
    auto kern = k_fwd_rollout_n1<NT, DOF, RING, SPS, WARPS, MINB>;
    static DynSmem attr;
    int rc = attr.ensure(kern, smem, h->device);
    if (rc) return rc;
    kern<<<(nB + WARPS - 1) / WARPS, 32 * WARPS, smem, st>>>(h->dp, h->W, b0, nB);
    return 0;
}
template <int NT, int DOF, int MINB>
int launch_fwd_select_t(ilqrb_handle h, int b0, int nB, cudaStream_t st, int it, int nb_iter, int line_search, int early_stop) {
    const size_t smem = (size_t)fwd_select_smem_doubles(h->hp.n_kp_steps) * sizeof(double);
    k_fwd_select_n1<NT, DOF, MINB><<<nB, 32, smem, st>>>(h->dp, h->W, b0, it, nb_iter, line_search, early_stop);
    return 0;
}
template <int NT, int DOF, int RING, int SPS, int WARPS, int MINB, int RR = 6>
int launch_forward_ring_t(ilqrb_handle h, int b0, int nB, cudaStream_t st, int it, int nb_iter, int line_search, int early_stop) {
    const size_t smem = (size_t)WARPS * fwd_fused_warp_bytes<RING, SPS>(h->hp.m, h->W.ns, h->hp.n_kp_steps);
    auto kern = k_forward_ring_n1<NT, DOF, RING, SPS, WARPS, MINB, RR>;
    static DynSmem attr;
    int rc = attr.ensure(kern, smem, h->device);
    if (rc) return rc;
    kern<<<(nB + WARPS - 1) / WARPS, 32 * WARPS, smem, st>>>(h->dp, h->W, b0, nB, it, nb_iter, line_search, early_stop);
    return 0;
}
bool has_joint_task(const DevProblem& P) {
    for (int t = 0; t < P.n_tasks; t++)
        if (P.tasks[t].kind == ILQRB_TASK_JOINT) return true;
    return false;
}
// A/B switches: ILQRB_FWD_V1 = the single-kernel forward pass of round 1 (gains through a register ring), ILQRB_FWD_SPLIT = the
// bulk-copy rollout and phases B-F as two launches; default = both in one kernel (k_forward_ring_n1)
int fwd_mode() {
    static const int mode = getenv("ILQRB_FWD_V1") ? 1 : (getenv("ILQRB_FWD_SPLIT") ? 2 : 0);
    return mode;
}

}  // namespace

// Problem groups (streams) for device-resident results.  The warp-per-problem backward kernel fills the SMs' warp slots and gains
// from sharing them with another group's forward pass: about three groups per two waves of the forward kernel, at least 3 (C2
// B = 4096 / 8192 / 16384 -> 3 / 3 / 6, C5 B = 8192 -> 7).  The column-per-lane kernel holds 4 problems per warp and is bound by latency
// below ~8 warps per SM: it wants large launches (B = 4096: 636 / 620 / 589 / 572 k solves/s with 1 / 2 / 3 / 4 groups in
// profiles/r2f_backward_col_groups.txt, 673 / 686 / 625 / 581 k with 1 / 2 / 3 / 5 in r2p_e2e_chain.txt on another box; 8192: flat; 16384: 768 /
// 778 / 750 / 744 k): two groups from 4096 problems, one per 8192 beyond.
bool fast_n1_col_path(ilqrb_handle h) {
    const StateBlocks sb = find_state_blocks(h->hp);
    if (sb.nblk >= 2) {
        for (int i = 0; i < sb.nblk; i++)
            if (sb.size[i] != 7) return false;
        return col_backward_applies(h->hp, true);
    }
    return col_backward_applies(h->hp, false);
}
int fast_n1_default_groups(ilqrb_handle h, int B) {
    if (col_backward_applies(h->hp, false) && find_state_blocks(h->hp).nblk < 2) return B < 4096 ? 1 : std::max(2, B / 8192);
    const int wave = fast_n1_wave(h);
    return std::min(8, std::max(3, (3 * B + 2 * wave - 1) / (2 * wave)));
}

int launch_backward_n1(ilqrb_handle h, int b0, int nB, cudaStream_t st) {
    h->begin(PS_BACKWARD);
    const int n = h->hp.n;
    int rc;
    const StateBlocks sb = find_state_blocks(h->hp);
    if (sb.nblk >= 2) {
        bool all7 = true;
        for (int i = 0; i < sb.nblk; i++) all7 = all7 && sb.size[i] == 7;
        if (all7 && col_backward_applies(h->hp, true)) rc = launch_backward_n1_col_blocks_t<1>(h, b0, nB, st, sb);
        else rc = all7 ? launch_backward_n1_blocks_t<7, 4, 7>(h, b0, nB, st, sb) : launch_backward_n1_blocks_t<0, 4, 7>(h, b0, nB, st, sb);
    } else if (col_backward_applies(h->hp, false)) {
        switch (n) {
            case 2: rc = launch_backward_n1_col_t<2>(h, b0, nB, st); break;
            case 3: rc = launch_backward_n1_col_t<3>(h, b0, nB, st); break;
            case 4: rc = launch_backward_n1_col_t<4>(h, b0, nB, st); break;
            case 5: rc = launch_backward_n1_col_t<5>(h, b0, nB, st); break;
            case 6: rc = launch_backward_n1_col_t<6>(h, b0, nB, st); break;
            default: rc = launch_backward_n1_col_t<7>(h, b0, nB, st); break;
        }
    }
    else if (n == 7) rc = launch_backward_n1_t<1, 7, 4, 7>(h, b0, nB, st);
    else if (n == 14) rc = launch_backward_n1_t<2, 14, 4, 2>(h, b0, nB, st);
    else if (n + 1 <= 8) rc = launch_backward_n1_t<1, 0, 4, 7>(h, b0, nB, st);
    else rc = launch_backward_n1_t<2, 0, 4, 2>(h, b0, nB, st);
    h->end();
    if (rc) return rc;
    CU(cudaGetLastError());
    return 0;
}
int launch_forward_n1(ilqrb_handle h, int b0, int nB, cudaStream_t st, int it, int nb_iter, int line_search, int early_stop) {
    const int n = h->hp.n;
    int rc;
    if (has_joint_task(h->hp)) {  // joint-space sub-systems (f(x) = x, J = I; joint tasks have at most 7 joints): 7 residual rows
        h->begin(PS_LINESEARCH);
        rc = launch_forward_ring_t<1, 0, 2, 4, 1, 24, 7>(h, b0, nB, st, it, nb_iter, line_search, early_stop);
        h->end();
        if (rc) return rc;
        CU(cudaGetLastError());
        return 0;
    }
    if (fwd_mode() == 0) {
        h->begin(PS_LINESEARCH);
        // (RING, SPS, WARPS per block, blocks per SM) measured on B200 at C2 (tools/sweep_fwd.sh, profiles/r2_sweep_forward.txt), us per
        // launch of 4096 problems: (2,4,1,28) 121 | (3,4,1,28) 127 | (2,6,1,28) 127 | (2,6,4,7) 138 | (3,4,4,7) 144 | round-1 kernel 133 |
        // rollout and phases B-F as two launches 55 + 96
        static const int variant = getenv("ILQRB_FWD_VARIANT") ? atoi(getenv("ILQRB_FWD_VARIANT")) : 0;
        if (n == 7 && variant == 1) rc = launch_forward_ring_t<1, 7, 3, 4, 4, 7>(h, b0, nB, st, it, nb_iter, line_search, early_stop);
        else if (n == 7 && variant == 2) rc = launch_forward_ring_t<1, 7, 3, 4, 1, 28>(h, b0, nB, st, it, nb_iter, line_search, early_stop);
        else if (n == 7) rc = launch_forward_ring_t<1, 7, 2, 4, 1, 28>(h, b0, nB, st, it, nb_iter, line_search, early_stop);
        else if (n == 14) rc = launch_forward_ring_t<2, 14, 3, 2, 4, 3>(h, b0, nB, st, it, nb_iter, line_search, early_stop);
        else if (n + 1 <= 8) rc = launch_forward_ring_t<1, 0, 2, 4, 1, 28>(h, b0, nB, st, it, nb_iter, line_search, early_stop);
        else rc = launch_forward_ring_t<2, 0, 3, 2, 4, 3>(h, b0, nB, st, it, nb_iter, line_search, early_stop);
        h->end();
        if (rc) return rc;
        CU(cudaGetLastError());
        return 0;
    }
    if (fwd_mode() == 2) {
        h->begin(PS_FWD_ROLLOUT);
        if (n == 7) rc = launch_fwd_rollout_t<1, 7, 3, 4, 4, 7>(h, b0, nB, st);
        else if (n == 14) rc = launch_fwd_rollout_t<2, 14, 3, 2, 4, 3>(h, b0, nB, st);
        else if (n + 1 <= 8) rc = launch_fwd_rollout_t<1, 0, 3, 4, 4, 7>(h, b0, nB, st);
        else rc = launch_fwd_rollout_t<2, 0, 3, 2, 4, 3>(h, b0, nB, st);
        h->end();
        if (rc) return rc;
        CU(cudaGetLastError());
        h->begin(PS_LINESEARCH);
        if (n == 7) rc = launch_fwd_select_t<1, 7, 16>(h, b0, nB, st, it, nb_iter, line_search, early_stop);
        else if (n == 14) rc = launch_fwd_select_t<2, 14, 12>(h, b0, nB, st, it, nb_iter, line_search, early_stop);
        else if (n + 1 <= 8) rc = launch_fwd_select_t<1, 0, 16>(h, b0, nB, st, it, nb_iter, line_search, early_stop);
        else rc = launch_fwd_select_t<2, 0, 12>(h, b0, nB, st, it, nb_iter, line_search, early_stop);
        h->end();
        if (rc) return rc;
        CU(cudaGetLastError());
        return 0;
    }
    h->begin(PS_LINESEARCH);
    if (n == 7) rc = launch_forward_n1_t<1, 7, 28>(h, b0, nB, st, it, nb_iter, line_search, early_stop);
    else if (n == 14) rc = launch_forward_n1_t<2, 14, 12>(h, b0, nB, st, it, nb_iter, line_search, early_stop);
    else if (n + 1 <= 8) rc = launch_forward_n1_t<1, 0, 28>(h, b0, nB, st, it, nb_iter, line_search, early_stop);
    else rc = launch_forward_n1_t<2, 0, 12>(h, b0, nB, st, it, nb_iter, line_search, early_stop);
    h->end();
    if (rc) return rc;
    CU(cudaGetLastError());
    return 0;
}

}  // namespace ilqrb
