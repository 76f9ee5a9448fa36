#!/usr/bin/env python
"""bench.py — iLQR problems solved/sec on B200 (BASELINE.json metric), one JSON line on stdout.

A "step" = one batched Recursive-iLQR solve (20 iterations, line search on, no early stop) of configs[1]:
4096 synthetic 7-DoF position+quaternion reaching problems, T=100, per GPU (weak scaling: every rank solves its own
4096 problems; no collective on the data path, one final gather of the costs).

  value     whole-job solves/s with U0 already resident in HBM (device-timed with CUDA events on the solve stream)
  e2e       same metric through the C ABI call ilqrb_solve_recursive with pinned HOST buffers (H2D of U0 and D2H of
            X, f(X), U, K, d, cost, traces inside the timed region)
  roofline  dominant kernel (riccati_backward): algorithmic bytes / measured launch duration vs measured HBM peak;
            roofline_fp64 adds algorithmic FP64 flops vs a DFMA peak measured here
  cpu_baseline / --impl reference: the CPU oracle (restatement of the reference's Eigen/KDL path; the reference itself
            cannot be built here) on all host cores, bounded sample of the same workload
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NB_ITER = 20
BATCH_PER_GPU = 4096
HORIZON = 100
METRIC = "iLQR problems solved/sec (7-DoF, T=100, Recursive, 20 iterations)"
UNIT = "solves/s"


def workload_config(n_gpus, batch):
    return {"workload": "configs[1]: synthetic 7-DoF KDL chain, PosOrn system, Recursive iLQR, T=100, 20 iterations, line search on, "
                        "no early stop, batch %d per GPU" % batch,
            "batch_per_gpu": batch, "global_batch": batch * n_gpus, "horizon": HORIZON, "nb_iter": NB_ITER, "parallelism": "problem-batch dp%d" % n_gpus,
            "l2": "no flush: every iteration streams the 159 MB of gains (written by the backward pass, read by the forward pass) plus "
                  "26 MB of trajectories through the 126 MB L2, so no launch re-times a cached working set"}


class ClockSampler:
    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.samples = []
        self.reasons = set()
        self._stop = threading.Event()
        self._t = None

    def _run(self):
        # NVML directly (a sample costs ~0.1 ms); nvidia-smi as the fallback
        try:
            import pynvml
            pynvml.nvmlInit()
            hdl = pynvml.nvmlDeviceGetHandleByIndex(self.idx)
            mx = float(pynvml.nvmlDeviceGetMaxClockInfo(hdl, pynvml.NVML_CLOCK_SM))
            bits = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}
            while not self._stop.is_set():
                self.samples.append((float(pynvml.nvmlDeviceGetClockInfo(hdl, pynvml.NVML_CLOCK_SM)), mx))
                try:
                    r = pynvml.nvmlDeviceGetCurrentClocksEventReasons(hdl)
                except Exception:
                    r = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(hdl)
                for name, bit in bits.items():
                    if r & bit:
                        self.reasons.add(name)
                self._stop.wait(0.005)
            return
        except Exception:
            pass
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self._stop.is_set():
            try:
                o = subprocess.run(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + q, "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in o.strip().split(",")]
                self.samples.append((float(f[0]), float(f[1])))
                for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[2:6]):
                    if v.lower().startswith("active"):
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.1)

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=6)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": sorted(self.reasons)}
        sm = sorted(s[0] for s in self.samples)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.samples[0][1], "reasons": sorted(self.reasons), "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_figures():
    """per-kernel DRAM traffic and executed FP64 flops per launch from the committed ncu capture (profiles/ncu_kernels.json,
    written by tools/ncu_summary.py --json from the `ncu --set full` report of tools/profile_run.py)"""
    p = os.path.join(ROOT, "profiles", "ncu_kernels.json")
    try:
        return json.load(open(p))
    except Exception:
        return {}


def ncu_kernel(figs, key, batch):
    """entry of kernel `key` if the capture was taken at this batch size (figures are per launch)"""
    e = figs.get("kernels", {}).get(key)
    if not e or e.get("batch") != batch:
        return {}
    return e


def cpu_baseline_run(batch, threads=None):
    """the oracle over `batch` problems of the workload on all host threads -> (solves/s, threads, seconds)"""
    from ilqr_planner_b200 import synth
    from oracle import oracle
    threads = threads or oracle.hardware_threads() or os.cpu_count() or 1
    pb = synth.config2(batch=batch, horizon=HORIZON)
    t0 = time.perf_counter()
    out = oracle.solve_recursive(pb, None, NB_ITER, True, False, threads=threads)
    dt = time.perf_counter() - t0
    return batch / dt, threads, dt, out


def use_test_side_chain():
    """the robot chain from the test-side URDF reader (oracle/urdf_chain.py) instead of the product's ilqrb_chain_from_urdf:
    the reference arm's process then never maps libilqrb.so"""
    from ilqr_planner_b200 import synth
    from oracle import urdf_chain
    synth._chain_cache["panda"] = urdf_chain.kdl_chain_from_urdf(open(synth.PANDA_URDF).read(), "panda_link0", "panda_tip")


def run_reference(args, rank, world):
    if rank != 0:
        return
    use_test_side_chain()
    sample = 16 * max(1, (os.cpu_count() or 8) // 4)
    for _ in range(args.warmup):
        cpu_baseline_run(max(sample // 4, 8))
    tot_t, tot_n, threads = 0.0, 0, 1
    for _ in range(args.steps):
        v, threads, dt, _ = cpu_baseline_run(sample)
        tot_t += dt
        tot_n += sample
    value = tot_n / tot_t
    desc = "each step = %d problems of the workload (of %d) solved by the CPU oracle on %d host threads" % (sample, BATCH_PER_GPU, threads)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * tot_t / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args.gpus, BATCH_PER_GPU),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": desc},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def w_bw(n, m):
    """SURVEY 8d dense-generic flop count of one backward Riccati step"""
    return 4 * m * n * n + 2 * m * n * n + 2 * m * m * n + 4 * n ** 3 + 2 * n ** 3 + 2 * n * n * m + 2 * m * n + 2 * n * n + 2 * m ** 3 + 2 * m * m * n + 2 * m * m \
        + (2 * n * m * m + 6 * n * n * m + 3 * n * n) + 6 * n * m


def other_configs(torch, device, fp64_peak, hbm_peak):
    """The other BASELINE.json configs on one GPU (parity-test shapes; the bench line proper is configs[1]): device-resident
    solves/s (CUDA events on the solve stream, U0 reset inside the timed region), the dominant kernel with its algorithmic
    roofline fraction, and the same solve through the blocking host-buffer C-ABI call (e2e: H2D of U0 and D2H of every result
    inside the timed region)."""
    import numpy as np
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.problem import ResultBuffers
    from ilqr_planner_b200.solver import Handle, build_psi, make_al_desc, make_batch_desc
    res = {}

    def run(name, pb, nit, kind="recursive", al=None, u0=None, psi=None, reps=3, flops_of=None, note=None):
        stream = torch.cuda.Stream()
        with Handle(pb, device=device, stream=stream.cuda_stream) as h:
            ald = make_al_desc(**al) if al else None
            bd = make_batch_desc(None, psi) if kind == "batch" else None

            # the initial controls are reset inside the timed region from a device-resident copy (a host array would put a 26 MB
            # pageable H2D copy of U0 into the device-resident number of the time-system configs)
            u0d = None if u0 is None else torch.from_numpy(np.ascontiguousarray(u0)).to("cuda")

            def go():
                if u0d is None:
                    h.upload_u0(None)
                else:
                    h.upload_u0_device(u0d.data_ptr())
                if kind == "al":
                    h.run_al(ald, nit, True, False)
                elif kind == "batch":
                    h.run_batch(bd, nit, False)
                else:
                    h.run_recursive(nit, True, False)
            go()
            h.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(reps):
                go()
            e1.record(stream)
            h.synchronize()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            h.set_profiling(True)
            h.reset_profile()
            go()
            prof = h.profile()
            h.set_profiling(False)
            # e2e: blocking host-buffer call
            out = ResultBuffers(pb, nit, kind, nc=(ald.nc if ald else 0))
            u0h = None if u0 is None else np.ascontiguousarray(u0)

            def call():
                if kind == "al":
                    h.solve_al(u0h, al["A"], al["b"], al["lambda0"], nit, al["lag_update_step"], al["penalty"], al["scaling_factor"], True, False, out=out)
                elif kind == "batch":
                    h.solve_batch(u0h, nit, False, psi=psi, out=out)
                else:
                    h.solve_recursive(u0h, nit, True, False, out=out)
            call()
            t0 = time.perf_counter()
            call()
            e2e_s = time.perf_counter() - t0
        dom = max(prof.items(), key=lambda kv: kv[1][0])
        entry = {"batch": pb.batch, "nb_iter": nit, "solves_per_s": pb.batch / (ms * 1e-3), "ms_per_solve_batch": ms, "ms_per_iteration": ms / nit,
                 "kernel_ms": {k: round(v[0], 3) for k, v in prof.items()}, "dominant_kernel": dom[0], "e2e_solves_per_s": pb.batch / e2e_s,
                 "e2e_mode": "one blocking host-buffer C-ABI call (pageable numpy buffers), H2D + solve + D2H"}
        if flops_of:  # (kernel family, algorithmic flops of the whole batch per ITERATION, peak, unit)
            kname, flops, peak, unit = flops_of
            t = prof.get(kname, (0.0, 0))
            if t[1]:
                ach = flops / (t[0] / nit * 1e-3) / (1e12 if unit == "TFLOP/s" else 1e9)
                entry["roofline"] = {"kernel": kname, "achieved": ach, "peak": peak, "unit": unit, "frac": ach / peak if peak else None, "ms_per_iteration": t[0] / nit}
        if note:
            entry["note"] = note
        res[name] = entry

    H = 100
    pb = synth.config1(batch=4096)
    run("C1 Robot2D position T=50 Recursive", pb, 20, flops_of=("riccati_backward", w_bw(3, 3) * 49.0 * 4096, fp64_peak, "TFLOP/s"))
    pb = synth.config3_active(batch=16384)
    al = synth.al_config3_active("qbox", lag_update_step=5)
    run("C3 AL-iLQR nd=2 T=100, 14 joint-box rows ACTIVE (config3_active)", pb, 40, kind="al", al=al, reps=2,
        flops_of=("riccati_backward", (w_bw(14, 7) + 2 * 14 * 21 * 21) * 99.0 * 16384, fp64_peak, "TFLOP/s"),
        note="dense-generic algorithmic count (SURVEY 8d); the kernel exploits the integrator structure, so a fraction near or above 1 means "
             "the counted work is not executed - see executed_pipe_utilisation of profiles/")
    pb = synth.config3(batch=16384)
    run("C3 AL-iLQR nd=2 T=100, rows = Panda joint limits (inert for these targets)", pb, 40, kind="al", al=synth.al_config3(), reps=2)
    pb = synth.config5(batch=8192)
    run("C5 dual arm n=m=14 T=100 Recursive", pb, 20, flops_of=("riccati_backward", w_bw(14, 14) * 99.0 * 8192, fp64_peak, "TFLOP/s"),
        note="block-diagonal backward pass: two 7x7 recursions per problem instead of the dense 14x14 the algorithmic figure counts")
    pb = synth.config4(batch=4096, horizon=H)
    run("PosOrnTime Recursive T=100", pb, 20, u0=synth.u0_time(pb), flops_of=("riccati_backward", w_bw(8, 8) * 99.0 * 4096, fp64_peak, "TFLOP/s"))
    pb = synth.config4(batch=1024, horizon=200)
    N = 199 * 8
    run("C4 BatchILQR PosOrnTime T=200 N=1592, one 1024-problem chunk", pb, 3, kind="batch", u0=synth.u0_time(pb), reps=1,
        flops_of=("batch_cholesky", (N ** 3 / 3.0 + 2.0 * N * N) * 1024, fp64_peak, "TFLOP/s"),
        note="batch_cholesky = blocked Cholesky + the two triangular solves, N^3/3 + 2 N^2 flop per problem and iteration")
    psi = np.kron(build_psi("unitstep", 199, 2), np.identity(8))
    run("C4 BatchILQRCP PosOrnTime T=200, Psi = 2 unit steps", pb, 20, kind="batch", u0=synth.u0_time(pb), psi=psi, reps=1)
    return res


def stream_e2e(torch, hs, outs, u0s, steps, nb_iter, streams, barrier):
    """the streaming host-buffer form: batches alternate between two handles, batch i's D2H overlapping batch i+1's solve -> ms"""
    for i in range(4):
        hs[i % 2].synchronize()
        hs[i % 2].solve_recursive_async(u0s[i % 2], nb_iter, True, False, out=outs[i % 2])
    hs[0].synchronize()
    hs[1].synchronize()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(streams[0])
    streams[1].wait_event(e0)
    chain = os.environ.get("ILQRB_BENCH_CHAIN", "once")  # once | always | never
    for i in range(steps):
        k = i % 2
        hs[k].synchronize()
        if (chain == "once" and i == 1) or chain == "always":
            hs[k].chain_after_compute(hs[1 - k])
        hs[k].solve_recursive_async(u0s[k], nb_iter, True, False, out=outs[k])
    streams[0].wait_stream(streams[1])
    e1.record(streams[0])
    barrier()
    return e0.elapsed_time(e1)


def d2h_ceiling_gbs(torch, dist, world, barrier, mbytes=256, reps=4):
    """aggregate device-to-host bandwidth of this box with all `world` GPUs copying into pinned host memory at the same time
    (plain cudaMemcpyAsync, one copy per call): the ceiling of every host-buffer e2e number at this N"""
    n = mbytes * (1 << 20) // 8
    src = torch.zeros(n, dtype=torch.float64, device="cuda")
    dst = torch.empty(n, dtype=torch.float64).pin_memory()
    dst.copy_(src, non_blocking=True)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        dst.copy_(src, non_blocking=True)
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return world * reps * n * 8 / (float(t.item()) * 1e-3) / 1e9


def dfma_peak_tflops(torch):
    """FP64 peak proxy measured here: torch fp64 matmul 4096^3 (cuBLAS DGEMM), best of 5 — MEASURED_PEAKS.json has no FP64 entry"""
    try:
        a = torch.randn(4096, 4096, dtype=torch.float64, device="cuda")
        b = torch.randn(4096, 4096, dtype=torch.float64, device="cuda")
        best = 1e9
        for _ in range(6):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch.matmul(a, b)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return 2 * 4096 ** 3 / (best * 1e-3) / 1e12
    except Exception:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=BATCH_PER_GPU, help="problems per GPU (default: the BASELINE config)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the other BASELINE configs (N=1 line only)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    from ilqr_planner_b200 import synth
    from ilqr_planner_b200.problem import ResultBuffers
    from ilqr_planner_b200.solver import Handle

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        # NCCL announces its version on stdout when the communicator comes up (NCCL_DEBUG=VERSION on some boxes); stdout carries
        # the one JSON line only, so file descriptor 1 points at stderr until the first collective has run
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)
    B = args.batch
    pb = synth.config2(batch=B, horizon=HORIZON, offset=rank * B)
    stream = torch.cuda.Stream()
    h = Handle(pb, device=local_rank, stream=stream.cuda_stream)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput ("value") -----------------------------------------------------------------
    # every solve starts from U0 = 0: the reset (a cudaMemsetAsync on the solve stream) is part of the timed step
    for _ in range(args.warmup):
        h.upload_u0(None)
        h.run_recursive(NB_ITER, True, False)
    h.synchronize()
    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0:
        sampler.start()
    l0 = h.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(args.steps):
            h.upload_u0(None)
            h.run_recursive(NB_ITER, True, False)
        e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = h.launch_count() - l0
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = world * B * args.steps / (ms_max * 1e-3)

    # ---- end to end through the host-buffer C ABI calls ("e2e") -------------------------------------------------
    # (a) one synchronous call per batch (ilqrb_solve_recursive): H2D -> solve -> D2H, nothing overlaps across batches;
    # (b) the streaming form a throughput user writes (ilqrb_solve_recursive_async on two handles used alternately):
    #     batch i's result transfer (PCIe) overlaps batch i+1's solve.  Every batch still uploads its U0 from pinned host
    #     memory and lands all of X, f(X), U, K, d, cost and the traces in pinned host memory inside the timed region, and
    #     the region ends when the last batch's results are on the host.  (b) is the headline e2e, (a) is kept beside it.
    out = ResultBuffers(pb, NB_ITER, "recursive", pinned=True)
    u0_pinned = torch.zeros((B, (HORIZON - 1) * pb.m), dtype=torch.float64).pin_memory()
    u0 = u0_pinned.numpy()
    h.solve_recursive(u0, NB_ITER, True, False, out=out)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        h.solve_recursive(u0, NB_ITER, True, False, out=out)
    e1.record(stream)
    barrier()
    ms_sync = e0.elapsed_time(e1)

    stream2 = torch.cuda.Stream()
    h2 = Handle(pb, device=local_rank, stream=stream2.cuda_stream)
    out2 = ResultBuffers(pb, NB_ITER, "recursive", pinned=True)
    u0b = torch.zeros((B, (HORIZON - 1) * pb.m), dtype=torch.float64).pin_memory().numpy()
    hs, outs, u0s = (h, h2), (out, out2), (u0, u0b)
    for i in range(4):  # warm-up of both handles
        hs[i % 2].synchronize()
        hs[i % 2].solve_recursive_async(u0s[i % 2], NB_ITER, True, False, out=outs[i % 2])
    h.synchronize()
    h2.synchronize()
    out2.cost[:] = np.nan
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record(stream)
    stream2.wait_event(e0)
    stagger_s = float(os.environ.get("ILQRB_BENCH_STAGGER_MS", "0")) * 1e-3
    chain = os.environ.get("ILQRB_BENCH_CHAIN", "once")  # once | always | never (A/B switch)
    for i in range(args.steps):
        k = i % 2
        hs[k].synchronize()  # batch i-2 is on the host (a consumer would read outs[k] here) before its buffers are reused
        if (chain == "once" and i == 1) or chain == "always":
            # The second handle's first solve starts when the first handle's kernels have finished (ilqrb_chain_after_compute): that puts
            # the handles half a cycle apart — one computes while the other copies — instead of in phase (compute together, then copy
            # together).  From then on each handle reissues as soon as its previous batch is on the host.
            hs[k].chain_after_compute(hs[1 - k])
        hs[k].solve_recursive_async(u0s[k], NB_ITER, True, False, out=outs[k])
        if i == 0 and stagger_s > 0:
            time.sleep(stagger_s)  # A/B switch: the second handle starts late instead of being chained
    stream.wait_stream(stream2)
    e1.record(stream)
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - t0)
    ms_e2e = e0.elapsed_time(e1)
    assert args.steps < 2 or np.array_equal(out.cost, out2.cost), "the two pipelined handles disagree"
    t = torch.tensor([ms_e2e, ms_sync], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_e2e, ms_sync = float(t[0].item()), float(t[1].item())
    e2e_value = world * B * args.steps / (ms_e2e * 1e-3)
    e2e_sync_value = world * B * args.steps / (ms_sync * 1e-3)
    h2d = u0.nbytes
    d2h = sum(a.nbytes for a in (out.X, out.fX, out.U, out.K, out.d, out.cost, out.cost_hist, out.alpha_hist, out.iters, out.status))
    # the same streaming form for a caller that does not take the gains back (out->K = out->d = NULL: 4x fewer bytes), and the
    # box's aggregate D2H ceiling with all ranks copying at once: the host-buffer numbers at N >= 4 are bounded by it
    outs_ng = (ResultBuffers(pb, NB_ITER, "recursive", pinned=True, gains=False), ResultBuffers(pb, NB_ITER, "recursive", pinned=True, gains=False))
    ms_ng = stream_e2e(torch, hs, outs_ng, u0s, args.steps, NB_ITER, (stream, stream2), barrier)
    t = torch.tensor([ms_ng], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ng_value = world * B * args.steps / (float(t.item()) * 1e-3)
    d2h_ng = sum(a.nbytes for a in (outs_ng[0].X, outs_ng[0].fX, outs_ng[0].U, outs_ng[0].cost, outs_ng[0].cost_hist, outs_ng[0].alpha_hist, outs_ng[0].iters, outs_ng[0].status))
    d2h_ceiling = d2h_ceiling_gbs(torch, dist, world, barrier)
    h2.close()
    del outs_ng

    # ---- north_star target workloads (device-resident, max over ranks): 8192 problems of C2 per GPU (65,536 on 8 GPUs) and the
    # 65,536-problem dual-arm batch of configs[4] split evenly over the ranks
    def target_run(pbt, reps=2):
        st = torch.cuda.Stream()
        with Handle(pbt, device=local_rank, stream=st.cuda_stream) as ht:
            ht.upload_u0(None)
            ht.run_recursive(NB_ITER, True, False)
            ht.synchronize()
            barrier()
            a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(st)
            for _ in range(reps):
                ht.upload_u0(None)
                ht.run_recursive(NB_ITER, True, False)
            b_.record(st)
            ht.synchronize()
            barrier()
            tt = torch.tensor([a.elapsed_time(b_) / reps], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())
    north_star = None
    if not args.no_configs and B == BATCH_PER_GPU:
        ms_c2 = target_run(synth.config2(batch=8192, horizon=HORIZON, offset=rank * 8192))
        per = 65536 // world
        ms_c5 = target_run(synth.config5(batch=per, horizon=HORIZON, offset=rank * per))
        north_star = {"c2_8192_per_gpu": {"global_batch": 8192 * world, "solves_per_s": 8192 * world / (ms_c2 * 1e-3), "ms_per_solve_batch": ms_c2},
                      "configs4_dual_arm_65536_total": {"global_batch": per * world, "batch_per_gpu": per, "solves_per_s": per * world / (ms_c5 * 1e-3), "ms_per_solve_batch": ms_c5,
                                                         "scaling": "strong (65,536 problems split over the ranks)"},
                      "note": "device-resident, CUDA events on the solve stream, U0 reset inside the timed region, max over ranks; the north_star batch "
                              "(65,536 C2 problems on 8 GPUs) is the c2 entry at n_gpus = 8"}

    # ---- final gather of the per-problem costs (the only collective; outside the data path) -----------------------
    cost_t = torch.from_numpy(out.cost.copy()).cuda()
    if world > 1:
        gathered = [torch.empty_like(cost_t) for _ in range(world)] if rank == 0 else None
        dist.gather(cost_t, gathered, dst=0)
        all_cost = torch.cat(gathered).cpu().numpy() if rank == 0 else None
    else:
        all_cost = out.cost
    # ---- per-kernel durations with event pairs (separate pass, not part of `value`) -------------------------------
    h.set_profiling(True)
    h.reset_profile()
    h.upload_u0(None)
    h.run_recursive(NB_ITER, True, False)
    prof = h.profile()
    h.set_profiling(False)

    if rank == 0:
        n, m, H = pb.n, pb.m, pb.horizon
        hbm_peak, peak_src = measured_peaks()
        bw_ms, bw_n = prof.get("riccati_backward", (0.0, 1))
        bw_avg_s = bw_ms / max(bw_n, 1) * 1e-3
        # algorithmic bytes of one backward launch: write K,d once; read X,U once; read keypoint l_x,l_xx (SURVEY 8d share of the 133 KB/iteration)
        alg_bytes = B * 8.0 * ((H - 1) * (m * n + m) + H * n + (H - 1) * m + 2 * (n * n + n))
        flops_step = 4 * m * n * n + 2 * m * n * n + 2 * m * m * n + 4 * n ** 3 + 2 * n ** 3 + 2 * n * n * m + 2 * m * n + 2 * n * n + 2 * m ** 3 + 2 * m * m * n + 2 * m * m \
            + (2 * n * m * m + 6 * n * n * m + 3 * n * n) + 6 * n * m
        alg_flops = B * (H - 1) * float(flops_step)
        achieved = alg_bytes / bw_avg_s / 1e9 if bw_avg_s > 0 else None
        fp64_peak = dfma_peak_tflops(torch)
        from ilqr_planner_b200.solver import fp64_peaks
        dfma_peak, dmma_peak = fp64_peaks(local_rank)  # micro-benchmarks of the library itself (csrc/ubench.cu)
        figs = ncu_figures()
        nb, nf = (ncu_kernel(figs, "k_backward_n1_col", B) or ncu_kernel(figs, "k_backward_n1", B)), (ncu_kernel(figs, "k_forward_ring_n1", B) or ncu_kernel(figs, "k_forward_n1", B))
        tot_ms = sum(v[0] for v in prof.values())
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps,
                "ms_per_iteration": ms_max / args.steps / NB_ITER, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": workload_config(world, B), "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "ms_per_step": ms_e2e / args.steps,
                        "host_wall_ms_per_step": wall_ms / args.steps,
                        "mode": "ilqrb_solve_recursive_async on two handles used alternately, the second handle's first solve chained behind the first one's kernels (ilqrb_chain_after_compute) so that one handle computes while the other copies; every batch's "
                                "U0 H2D and full result D2H (pinned host memory) inside the timed region, which ends with the last batch on the host",
                        "sync_call": {"value": e2e_sync_value, "unit": UNIT, "ms_per_step": ms_sync / args.steps,
                                      "mode": "one blocking ilqrb_solve_recursive per batch (H2D -> solve -> D2H, no overlap across batches)"},
                        "no_gains": {"value": e2e_ng_value, "unit": UNIT, "d2h_bytes_per_step": int(d2h_ng),
                                     "mode": "the streaming form with out->K = out->d = NULL (the caller keeps X, f(X), U, cost and the traces only)"},
                        # what bounds the host-buffer numbers: all ranks' result copies converge on one host
                        "d2h_ceiling_gbs": d2h_ceiling, "d2h_gbs": e2e_value * d2h / B / 1e9, "frac_of_d2h_ceiling": e2e_value * d2h / B / 1e9 / d2h_ceiling,
                        "d2h_ceiling_note": "aggregate D2H rate of this box with all %d GPUs copying 256 MB blocks into pinned host memory at once" % world},
                "north_star_target": north_star,
                "gpu_launches": int(launches),
                # dominant kernel = the backward Riccati pass; SURVEY 8d: a compute roofline on the FP64 pipe.  k_backward_n1_col issues DFMA
                # (four problems per warp); DFMA, DMMA m8n8k4 and cuBLAS DGEMM peak at the same rate on B200 (peaks_measured_here_tflops)
                "roofline": {"kernel": "riccati_backward (k_backward_n1_col: FP64 FMA, four problems per warp)", "bound": "tensor",
                             "achieved": alg_flops / bw_avg_s / 1e12 if bw_avg_s > 0 else None, "peak": fp64_peak, "unit": "TFLOP/s",
                             "frac": (alg_flops / bw_avg_s / 1e12 / fp64_peak) if (bw_avg_s > 0 and fp64_peak) else None,
                             "traffic": nb.get("dram_bytes"),
                             "traffic_note": "dram__bytes_read+write per launch, %s" % figs.get("source", "profiles/ncu_kernels.json absent"),
                             "peak_source": "FP64 tensor peak measured in this run: cuBLAS DGEMM 4096^3 via torch.matmul (MEASURED_PEAKS.json has no FP64 entry; "
                                            "the bf16 figure there does not apply to an FP64 path)",
                             "peaks_measured_here_tflops": {"cublas_dgemm_4096": fp64_peak, "dfma_ubench": dfma_peak, "dmma_m8n8k4_ubench": dmma_peak},
                             # what the kernel executes (ncu: DFMA x 2 + DADD + DMUL thread instructions + DMMA m8n8k4 x 512), against the DFMA peak:
                             # the structure of A, B removes most of the dense-generic work the algorithmic figure counts
                             "executed_flops_per_launch": nb.get("executed_flops"),
                             "executed_pipe_utilisation": (nb["executed_flops"] / bw_avg_s / 1e12 / dfma_peak) if (nb.get("executed_flops") and bw_avg_s > 0 and dfma_peak) else None,
                             "algorithmic_flops_per_launch": alg_flops, "algorithmic_flops_per_problem_step": flops_step,
                             "avg_launch_ms": bw_avg_s * 1e3, "share_of_step": bw_ms / tot_ms if tot_ms else None},
                "roofline_hbm": {"kernel": "riccati_backward", "bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                                 "frac": (achieved / hbm_peak) if achieved else None, "peak_source": peak_src,
                                 "algorithmic_bytes_per_launch": alg_bytes},
                # second kernel of the iteration: line search + accept (k_forward_n1).  It streams K, d (written by the backward pass), X, U
                # once and writes X, U once: the byte view against the measured copy peak; it is bound by the latency of its serial 99-step
                # deviation rollout and the FK chains of the candidates (ncu: 51 % of stall samples on the long scoreboard), not by HBM
                "roofline_forward": (lambda fms, fn: {"kernel": "linesearch_rollout (k_forward_ring_n1)", "bound": "hbm",
                                     "achieved": (B * 8.0 * ((H - 1) * (m * n + m) + 2 * (H * n + (H - 1) * m)) / (fms / max(fn, 1) * 1e-3) / 1e9) if fms > 0 else None,
                                     "peak": hbm_peak, "unit": "GB/s", "peak_source": peak_src,
                                     "algorithmic_bytes_per_launch": B * 8.0 * ((H - 1) * (m * n + m) + 2 * (H * n + (H - 1) * m)),
                                     "avg_launch_ms": fms / max(fn, 1), "traffic": nf.get("dram_bytes"),
                                     "executed_flops_per_launch": nf.get("executed_flops"),
                                     "traffic_note": "dram__bytes_read+write per launch, cold-cache ncu capture, %s" % figs.get("source", "profiles/ncu_kernels.json absent")})(*prof.get("linesearch_rollout", (0.0, 1))),
                "kernel_ms_per_solve": {k: v[0] for k, v in prof.items()},
                "final_cost_mean": float(np.nanmean(all_cost)), "problems_with_nan": int(np.isnan(all_cost).sum())}
        if world == 1 and not args.no_cpu_baseline:
            sample = min(B, 4096)
            v, threads, dt, ref = cpu_baseline_run(sample)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": "first %d problems of the workload, all %d iterations, CPU oracle on %d threads, %.1f s" % (sample, NB_ITER, threads, dt)}
            # same problems -> same answers (sanity, not a timing)
            rel = np.abs(ref.cost - out.cost[:sample]) / np.maximum(np.abs(ref.cost), 1e-300)
            line["cpu_gpu_cost_max_rel_diff"] = float(np.nanmax(rel))  # after all 20 iterations, worst of the sample
            line["cpu_gpu_cost_median_rel_diff"] = float(np.nanmedian(rel))
            # the parity tests' measure: per-iteration cost traces against the scale of the trace, and the accepted step sequences
            gh, rh = out.cost_hist[:sample], ref.cost_hist
            line["cpu_gpu_trace_max_diff_over_trace_scale"] = float(np.nanmax(np.abs(gh - rh) / np.nanmax(np.abs(rh), axis=1, keepdims=True)))
            # the oracle reports each iteration's line-search decision margin; below the 1e-9 cost tolerance a decision is a
            # tie made by rounding (tests/test_gpu_parity.py: decisive_prefix), after which two runs may take different branches
            ga, ra = out.alpha_hist[:sample], ref.alpha_hist
            tie = ~(ref.margin_hist >= 1e-9) | np.isnan(ref.cost_hist)
            first_tie = np.where(tie.any(axis=1), tie.argmax(axis=1), NB_ITER)
            upto = np.array([np.array_equal(ga[i, :first_tie[i]], ra[i, :first_tie[i]]) for i in range(sample)])
            full = np.all((ga == ra) | (np.isnan(ga) & np.isnan(ra)), axis=1)
            line["cpu_gpu_step_sequences"] = {"problems": sample, "identical_up_to_first_tie": int(upto.sum()), "identical_all_%d_iterations" % NB_ITER: int(full.sum()),
                                              "problems_with_a_tie": int(tie.any(axis=1).sum())}
            # where a sequence parts before a declared tie: how far the cost has fallen by then (a cost 1e-6 of its start is a sum of
            # squared residuals whose 1e-12 trajectory differences move it by more than the 1e-9 margin)
            early = np.nonzero(~upto)[0]
            if early.size:
                it_d = [int(np.argmax(~((ga[i] == ra[i]) | (np.isnan(ga[i]) & np.isnan(ra[i]))))) for i in early]
                fall = [float(ref.cost_hist[i, max(k - 1, 0)] / np.nanmax(ref.cost_hist[i])) for i, k in zip(early, it_d)]
                line["cpu_gpu_step_sequences"]["parting_before_a_tie"] = {"count": int(early.size), "earliest_iteration": int(min(it_d)) + 1,
                                                                           "largest_cost_over_trace_scale_when_parting": max(fall)}
        if world == 1 and not args.no_configs and B == BATCH_PER_GPU:
            try:
                line["configs"] = other_configs(torch, local_rank, fp64_peak, hbm_peak)
            except Exception as e:  # the headline line must survive a failure here
                line["configs"] = {"error": repr(e)}
        rf = line["roofline_forward"]
        if rf.get("achieved"):
            rf["frac"] = rf["achieved"] / hbm_peak
        print(json.dumps(line), flush=True)
    h.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
