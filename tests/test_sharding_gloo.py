"""N > 1 path on CPU: two processes over gloo.  Each rank derives its image block and positions
its class-order stream on the host exactly as bench.py does under torchrun; the gathered order
tables must equal the single sequential stream (the E-step itself has no collective)."""
import os
import socket
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r"""
import os, sys
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import numpy as np, torch, torch.distributed as dist
import cases, em_adapt_tensorflow_b200 as wl
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
B = 30
sets = cases.voc_label_sets(B, np.random.default_rng(4321), forced={0: [0], 1: list(range(21))})
present = np.zeros((B, 21), np.uint8)
for b, s in enumerate(sets): present[b, s] = 1
lo, hi = wl.shard_range(B, world, rank)
mine = wl.draw_order(present[lo:hi], 5, wl.positioned_stream(present, 5, lo))
pad = np.full((B, 5, 21), -7, np.int32); pad[:hi - lo] = mine
got = [torch.zeros((B, 5, 21), dtype=torch.int32) for _ in range(world)]
dist.all_gather(got, torch.from_numpy(pad))
sizes = [wl.shard_range(B, world, r) for r in range(world)]
whole = np.concatenate([g.numpy()[: b - a] for g, (a, b) in zip(got, sizes)])
want = wl.draw_order(present, 5, wl.OrderStream(1))
# max-over-ranks timing reduction as in bench.py
t = torch.tensor([float(rank + 1)], dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX)
ok = np.array_equal(whole, want) and t.item() == world
dist.barrier(); dist.destroy_process_group()
print("RANK%%d %%s" %% (rank, "OK" if ok else "BAD"))
sys.exit(0 if ok else 1)
"""


def test_two_rank_sharding_over_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % dict(root=ROOT))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "RANK0 OK" in r.stdout and "RANK1 OK" in r.stdout


WORKER_LOSS = r"""
import os, sys
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import numpy as np, torch, torch.distributed as dist
import em_adapt_tensorflow_b200 as wl
from em_adapt_tensorflow_b200.loss import EMAdaptLoss
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
# the DDP-style wrapper: every rank owns its criterion (own plans, own order stream); there is no collective in it,
# and on a machine without a CUDA device it must refuse loudly instead of falling back to a CPU path
crit = EMAdaptLoss(seed=1 + rank)
ok = True
try:
    crit(torch.zeros(2, 8, 8, 21), torch.zeros(2, 8, 8, dtype=torch.uint8))
    ok = torch.cuda.is_available()
except (AssertionError, wl.WeakLabelError, RuntimeError):
    pass
# per-rank streams are independent and reproducible
a = [wl.OrderStream(1 + rank).next() for _ in range(3)]
got = [None] * world
dist.all_gather_object(got, a)
ok = ok and got[0] != got[1] and got[rank] == [wl.OrderStream(1 + rank).next() for _ in range(1)] + a[1:]
s = wl.OrderStream(1 + rank); b = [s.next() for _ in range(3)]
ok = ok and a[0] == b[0]
dist.barrier(); dist.destroy_process_group()
print("RANK%%d %%s" %% (rank, "OK" if ok else "BAD"))
sys.exit(0 if ok else 1)
"""


def test_loss_wrapper_two_ranks_over_gloo(tmp_path):
    """EMAdaptLoss under a 2-process group: no collective inside, no CPU fallback (it refuses CPU tensors)."""
    script = tmp_path / "worker_loss.py"
    script.write_text(WORKER_LOSS % dict(root=ROOT))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "RANK0 OK" in r.stdout and "RANK1 OK" in r.stdout
