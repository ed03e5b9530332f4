"""Row-partitioned solve vs the single-GPU solve of the same QP (run under torchrun, N >= 2):
   python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dist_check.py [n] [eps]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", rank))
os.environ["OSQP_CUDA_DEVICE"] = str(local)
import numpy as np
import torch, torch.distributed as dist
from osqp_b200 import Solver, problems
from osqp_b200.capi import Plugin
from osqp_b200 import dist as qd

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
eps = float(sys.argv[2]) if len(sys.argv) > 2 else 1e-5
polish = int(sys.argv[3]) if len(sys.argv) > 3 else 0
seed = int(sys.argv[4]) if len(sys.argv) > 4 else 7
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("gloo")
plug = Plugin()
p = problems.random_qp(n=n, seed=seed)
kw = dict(eps_abs=eps, eps_rel=eps, adaptive_rho_interval=50, polish=polish)

# single-GPU solve of the whole problem on every rank (the comparison value)
s1 = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], **kw)
assert s1.setup_flag == 0 and s1.solve() == 0
x1, y1, it1, obj1, st1 = s1.x, s1.y, int(s1.info.iter), float(s1.info.obj_val), s1.status
t1 = float(s1.info.solve_time)
pol1 = int(s1.info.status_polish)
s1.cleanup()

qd.init_partition(plug, rank, world)
sl = qd.slab(p, rank, world)
s = Solver(sl["P"], sl["q"], sl["A"], sl["l"], sl["u"], **kw)
assert s.setup_flag == 0, s.setup_flag
assert s.solve() == 0
x, y = s.x, qd.gather_rows(s.y, world)
part = plug.get_partition()
ok = (s.status == st1 and abs(float(s.info.obj_val) - obj1) <= 1e-4 * max(1.0, abs(obj1))
      and np.max(np.abs(x - x1)) <= 1e-3 * max(1.0, np.max(np.abs(x1))))
print(f"[rank {rank}/{world}] partitioned: {s.status} iter={int(s.info.iter)} obj={float(s.info.obj_val):.10g} "
      f"solve={float(s.info.solve_time):.4f}s allreduces={part['allreduces']} | single: {st1} iter={it1} obj={obj1:.10g} "
      f"solve={t1:.4f}s polish={int(s.info.status_polish)}/{pol1} | dx={np.max(np.abs(x - x1)):.2e} dy={np.max(np.abs(y - y1)):.2e} ok={ok}", flush=True)
s.cleanup()
qd.init_partition(plug, 0, 1)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
sys.exit(0 if ok else 1)
