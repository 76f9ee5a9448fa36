"""N>1 host logic on CPU: world_size-2 gloo run of the batch sharding used by bench.py (contiguous slices, per-rank
offsets, no data-path collective, one final gather).  The solve itself is stood in by the CPU oracle here (test only)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
sys.path.insert(0, %r)
import numpy as np, torch, torch.distributed as dist
from ilqr_planner_b200 import synth
from oracle import oracle
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
B = 6
pb = synth.config2(batch=B, horizon=30, offset=rank * B)            # what bench.py gives each rank
out = oracle.solve_recursive(pb, None, 3, True, False)
cost = torch.from_numpy(out.cost.copy())
gathered = [torch.empty_like(cost) for _ in range(world)] if rank == 0 else None
dist.gather(cost, gathered, dst=0)                                   # the final gather
t = torch.tensor([float(rank + 1)], dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)                             # max-over-ranks timing reduction
if rank == 0:
    allc = torch.cat(gathered).numpy()
    full = synth.config2(batch=world * B, horizon=30)                # the unsharded batch
    ref = oracle.solve_recursive(full, None, 3, True, False)
    assert np.array_equal(allc, ref.cost), "sharded != unsharded"
    sl = full.slice(B, 2 * B)
    assert np.array_equal(sl.q0, synth.config2(batch=B, horizon=30, offset=B).q0)
    assert float(t.item()) == world
    print("GLOO_OK")
dist.destroy_process_group()
'''


def test_two_rank_sharding_matches_single_batch(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % ROOT)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1", "--master-port",
                        "29517", str(script)], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "GLOO_OK" in r.stdout
