"""BASELINE config 4: a batch of independent MPC QPs (same Ad, Bd; different x0 => only l, u differ)
through osqp_solve_batch.   python tools/config4_batch.py [count] [threads...]
Several GPUs: torchrun --nproc-per-node N tools/config4_batch.py [count] [threads...] -- the batch is
split across the ranks (one process per GPU, no communication on the data path; gloo is only used
for the barrier and for the max-over-ranks of the batch time)."""
import sys, os, time, json
RANK, WORLD = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
if WORLD > 1:
    os.environ["OSQP_CUDA_DEVICE"] = os.environ.get("LOCAL_RANK", "0")
    import torch.distributed as dist
    dist.init_process_group("gloo")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from osqp_b200 import Solver, problems
from osqp_b200.capi import solve_batch, Plugin
plug = Plugin()

count_all = int(sys.argv[1]) if len(sys.argv) > 1 else 512
count = count_all // WORLD          # this rank's share
threads = [int(a) for a in sys.argv[2:]] or [1, 8, 16, 32]
base = problems.mpc(horizon=50, nx=20, nu=10, seed=1)
nx = base["nx"]
rng = np.random.default_rng(123 + RANK)
kw = dict(eps_abs=1e-3, eps_rel=1e-3, adaptive_rho_interval=50, warm_start=0)
t = time.time()
solvers = []
for k in range(count):
    x0 = rng.standard_normal(nx)
    l, u = base["l"].copy(), base["u"].copy()
    l[:nx] = -x0
    u[:nx] = -x0
    s = Solver(base["P"], base["q"], base["A"], l, u, **kw)
    assert s.setup_flag == 0
    solvers.append(s)
setup_s = time.time() - t
n, m = base["A"].shape[1], base["A"].shape[0]
# one at a time (the comparison): first `ref_n` problems
ref_n = min(count, 64)
t = time.time()
for s in solvers[:ref_n]:
    assert s.solve() == 0
seq_s = (time.time() - t) / ref_n
ref = [(s.x.copy(), s.y.copy(), int(s.info.iter), s.status) for s in solvers[:ref_n]]
print(json.dumps(dict(workload=f"{count} x {base['name']} (n={n}, m={m}, nnzA={base['A'].nnz})", setup_s_total=round(setup_s, 2),
                      sequential_ms_per_problem=round(1e3 * seq_s, 3), iters_first=[r[2] for r in ref[:8]])), flush=True)
for T in threads:
    for s in solvers:
        s.update_rho(0.1)        # adaptive rho persists across solves (as in the reference): same start for every run
    for s in solvers:
        plug.sync(s.engine_handle())     # the refactorisations update_rho enqueued are not part of the batch's time
    if WORLD > 1: dist.barrier()
    t = time.time()
    failed = solve_batch(solvers, T)
    dt = time.time() - t
    if WORLD > 1:
        import torch
        tt = torch.tensor([dt], dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt_all = float(tt[0])
        if RANK == 0:
            print(json.dumps(dict(gpus=WORLD, problems=count * WORLD, batch_s_max_over_ranks=round(dt_all, 3),
                                  problems_per_s_all_gpus=round(count * WORLD / dt_all, 1))), flush=True)
    same = all(np.array_equal(s.x, r[0]) and np.array_equal(s.y, r[1]) and int(s.info.iter) == r[2]
               for s, r in zip(solvers[:ref_n], ref))
    worst = max(float(np.max(np.abs(s.x - r[0]))) for s, r in zip(solvers[:ref_n], ref))
    if os.environ.get("BATCH_DEBUG"):
        for k, (sv, r) in enumerate(zip(solvers[:ref_n], ref)):
            print("   ", k, "dx=%.2e" % float(np.max(np.abs(sv.x - r[0]))), "iter", int(sv.info.iter), r[2],
                  "rho_updates", int(sv.info.rho_updates), "rho", sv.current_setting("rho"), sv.stats()["pcg_iters"])
    diff_it = sum(1 for s, r in zip(solvers[:ref_n], ref) if int(s.info.iter) != r[2])
    iters = sum(int(s.info.iter) for s in solvers)
    solved = sum(1 for s in solvers if s.status == "solved")
    print(json.dumps(dict(threads=T, failed=failed, solved=solved, batch_s=round(dt, 3),
                          problems_per_s=round(count / dt, 1), admm_iters_per_s=round(iters / dt, 1),
                          identical_to_osqp_solve=bool(same), max_abs_dx_vs_osqp_solve=worst, iter_count_differs=diff_it)), flush=True)
for s in solvers:
    s.cleanup()
