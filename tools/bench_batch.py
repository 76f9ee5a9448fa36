"""C4 timing: BatchILQR (dense N x N) and BatchILQRCP on PosOrnTime, T=200 (N = 1592), per-kernel device times."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from ilqr_planner_b200 import synth
from ilqr_planner_b200.solver import Handle, build_psi, make_batch_desc
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
it_b = int(sys.argv[2]) if len(sys.argv) > 2 else 3
it_cp = int(sys.argv[3]) if len(sys.argv) > 3 else 20
H = int(sys.argv[4]) if len(sys.argv) > 4 else 200
pb = synth.config4(batch=B, horizon=H)
u0 = synth.u0_time(pb)
psi = np.kron(build_psi("unitstep", pb.horizon - 1, 2), np.identity(pb.m))
with Handle(pb) as h:
    for name, kw, nit in (("BatchILQRCP", dict(psi=psi), it_cp), ("BatchILQR", dict(), it_b)):
        h.solve_batch(u0, 1, False, **kw)  # warm-up (allocations)
        h.set_profiling(True); h.reset_profile()
        t0 = time.perf_counter()
        out = h.solve_batch(u0, nit, False, **kw)
        dt = time.perf_counter() - t0
        prof = h.profile(); h.set_profiling(False)
        print(name, "B=%d N=%d iterations=%d: %.3f s -> %.1f solves/s, %.2f ms/iteration/batch; final cost mean %.4g, NaN %d, notSPD %d" % (
            B, (H - 1) * pb.m, nit, dt, B / dt, 1e3 * dt / nit, np.nanmean(out.cost), int(np.isnan(out.cost).sum()), int((out.status & 2 != 0).sum())))
        print("   kernel ms:", {k: round(v[0], 2) for k, v in prof.items()})
