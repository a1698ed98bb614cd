"""N>1 path of bench.py on CPU: two gloo ranks, the reduce helpers (max-over-ranks time, summed steps) and the
reference arm's "rank 0 alone prints, the others exit 0" rule.  No data-path collective exists (replicas only)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
sys.path.insert(0, %r)
import torch, torch.distributed as dist
import bench
d, world, rank, local = bench.dist_setup(2)
assert world == 2 and d is not None
dev = torch.device("cpu")
t = bench.reduce_max(d, 1.0 + rank, dev)        # slowest rank defines the time
n = bench.reduce_sum(d, 10.0, dev)              # every rank contributes its own steps (weak scaling)
assert t == 2.0 and n == 20.0, (t, n)
d.barrier()
if rank == 0:
    print("GLOO_OK", n / t)
d.destroy_process_group()
'''


def _run(args, extra_env=None):
    env = dict(os.environ)
    env.update(extra_env or {})
    return subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                           "--master-port", "29617"] + args, cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)


def test_reduce_helpers_two_ranks(tmp_path):
    w = tmp_path / "worker.py"
    w.write_text(WORKER % ROOT)
    r = _run([str(w)])
    assert r.returncode == 0, r.stderr[-2000:]
    assert "GLOO_OK 10.0" in r.stdout


def test_reference_arm_two_ranks_prints_once():
    r = _run(["bench.py", "--impl", "reference", "--gpus", "2", "--steps", "3", "--warmup", "3", "--frames", "3"])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["value"] > 0 and d["cpu_baseline"]["kind"] == "port"
