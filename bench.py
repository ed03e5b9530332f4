#!/usr/bin/env python
"""bench.py -- ADMM iterations/s and time-to-solution of osqp_solve on the 1e7-nnz random QP.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

A "step" is one cold-started osqp_solve (device-resident ADMM + PCG) to eps_abs = eps_rel =
1e-3 on BASELINE.json's config 2 (random QP, n = 1e5, m = 1e6, nnz(P)+nnz(A) ~ 1e7, fp64).
  value : ADMM iterations/s with P, A, q, l, u already resident in HBM (osqp_setup done),
          timed with CUDA events on the engine's own stream, max over ranks.
  e2e   : the same metric through the public C API from HOST buffers: every step does
          osqp_setup (H2D of P, A, q, l, u from pinned memory + device format conversion +
          Ruiz scaling) + osqp_solve + solution readback + osqp_cleanup.
N > 1 : one process per GPU, each rank solves its own copy of the SAME QP (same seed, so the
        iteration counts are equal and the curve measures the hardware; north_star: "batches of
        independent QPs are split across GPUs with no communication");
        value = total iterations of all ranks / max time  -> weak scaling.
  partitioned : at every N the line also carries a `partitioned` record: BASELINE config 5 (ONE QP,
        n = 1e6, m = 1e7, 2e8 nnz) with the rows of A split over the N GPUs and an NCCL allreduce of
        the [P | A'] partial sums inside every CG iteration (strong scaling), with the share of the
        collective.  `--partitioned` makes that workload the headline line instead.
--impl reference : the reference's own CPU path (oracle/_ref: the unmodified reference
        compiled from source, QDLDL backend, 1 thread -- QDLDL has no threading) on a bounded
        sample of the same generator (QDLDL's fill-in makes n = 1e5 a multi-hour job).
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

REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libosqp_ref.so")
# dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernels on this workload:
# written by tools/ncu_traffic.py from an `ncu --set full` capture of this build (never typed in)
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "ncu_traffic.json")
PARTITIONED_N1_ITS = 158.6     # config 5 on one GPU, round 1 (profiles/r1_bench_partitioned_config5_sell_n1.json)
SETTINGS = dict(eps_abs=1e-3, eps_rel=1e-3, adaptive_rho_interval=50, warm_start=0, verbose=0)
CPU_SAMPLE_N = 3000            # random_qp(n=3000): ~8 s of QDLDL setup + solve per step


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([t.strip() for t in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                  "sw_power_cap"), r[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def pinned_problem(prob):
    """Copy the CSC arrays of the problem into pinned host buffers (torch) and wrap them."""
    import numpy as np
    import scipy.sparse as sp
    import torch
    from osqp_b200.capi import CscHolder

    def pin(a):
        t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
        return t, t.numpy()
    keep = []
    Pu = sp.triu(sp.csc_matrix(prob["P"]), format="csc"); Pu.sort_indices()
    A = sp.csc_matrix(prob["A"]); A.sort_indices()
    out, nbytes = {}, 0
    for name, M in (("P", Pu), ("A", A)):
        arrs = []
        for a, dt in ((M.indptr, np.int64), (M.indices, np.int64), (M.data, np.float64)):
            t, v = pin(a.astype(dt)); keep.append(t); arrs.append(v); nbytes += v.nbytes
        out[name] = CscHolder.from_arrays(M.shape, *arrs)
    for name in ("q", "l", "u"):
        t, v = pin(prob[name].astype(np.float64)); keep.append(t); out[name] = v; nbytes += v.nbytes
    return out, keep, nbytes


def run_reference(args, rank, world):
    """--impl reference: the unmodified reference (QDLDL, CPU) on a bounded sample."""
    if rank != 0:
        return
    from osqp_b200 import Solver, problems
    if not os.path.exists(REF_LIB):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libosqp_ref.so not built"}))
        return
    n = args.cpu_n
    prob = problems.random_qp(n=n, seed=1)
    kw = dict(SETTINGS, linsys_solver=0)

    def step():
        t0 = time.perf_counter()
        s = Solver(prob["P"], prob["q"], prob["A"], prob["l"], prob["u"], lib_path=REF_LIB, **kw)
        assert s.setup_flag == 0
        s.solve()
        x = s.x  # noqa: F841  (solution read back, like the e2e leg of our arm)
        it, st, sol = int(s.info.iter), int(s.info.status_val), float(s.info.solve_time)
        s.cleanup()
        return it, time.perf_counter() - t0, st, sol
    warm = min(args.warmup, 1)       # a CPU factorisation has nothing to warm beyond caches
    for _ in range(warm):
        step()
    iters, secs, solve_secs, status = 0, 0.0, 0.0, None
    for _ in range(args.steps):
        it, dt, status, sol = step()
        iters += it; secs += dt; solve_secs += sol
    # value: iterations / solve time (what our arm's device-resident `value` measures);
    # e2e: iterations / (setup + solve + read-back), like our arm's e2e leg
    value = iters / solve_secs
    e2e_value = iters / secs
    nnz = int(prob["A"].nnz + (prob["P"].nnz + n) // 2)
    sample = (f"random_qp(n={n}, m={10 * n}, nnz~{nnz}) -- same generator and settings as config 2 at "
              f"1/{100000 // n} of its size: QDLDL's fill-in (nnz(L) ~ n^2/3) makes n=1e5 a multi-hour, "
              f">25 GB factorisation; setup+solve per step, {iters // args.steps} iters/step, status {status}")
    line = {"impl": "reference", "metric": "admm_iters_per_s", "value": value, "unit": "ADMM iterations/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": warm, "ms_per_step": 1e3 * solve_secs / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": {"workload": "random_qp_n1e5_m1e6_nnz1e7 (config 2) -- bounded sample",
                                            "sample_n": n, "eps": 1e-3, "settings": "defaults + adaptive_rho_interval=50"},
            "cpu_baseline": {"value": value, "unit": "ADMM iterations/s", "cores": 1, "kind": "reference",
                             "sample": sample},
            "e2e": {"value": e2e_value, "unit": "ADMM iterations/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0, "ms_per_step": 1e3 * secs / args.steps},
            "like_for_like": "this arm runs 1/%d of the headline size; our arm's line carries the same-size pair "
                             "(cpu_baseline.same_size_gpu)" % (100000 // n),
            "gpu_launches": 0}
    print(json.dumps(line))


def load_traffic(workload="random_qp"):
    """{bench kernel name: {dram_bytes_per_launch, source}} of the given workload's capture; the
    top-level entries are config 2's (random_qp), other workloads sit under their own key."""
    try:
        t = json.load(open(TRAFFIC_FILE))
    except (OSError, ValueError):
        return {}
    return t if workload == "random_qp" else t.get(workload, {})


def run_partitioned(args, rank, world, local, settings, barrier, allmax, json_out):
    """ONE QP, rows of A / l / u / z / y split over the ranks, x replicated, P's nonzeros spread over
    the ranks; every [P | A'] product is joined by an ncclAllReduce inside the CUDA library.
    Returns the JSON record on rank 0 (None elsewhere); json_out (if given) prints it."""
    import numpy as np
    from osqp_b200 import Solver, problems
    from osqp_b200 import dist as qd
    from osqp_b200.capi import Plugin
    plug = Plugin()
    qd.init_partition(plug, rank, world)
    n, m = args.pn, 10 * args.pn
    b = qd.row_bounds(m, world)
    t0 = time.perf_counter()
    sl = problems.random_qp_rows(n=n, nnz_per_row=args.pnnz, seed=1, rows=(b[rank], b[rank + 1]))
    sl["P"] = qd.p_share(sl["P"], rank, world)      # every rank carries 1/world of P's nonzeros
    gen_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    s = Solver(sl["P"], sl["q"], sl["A"], sl["l"], sl["u"], **settings)
    if s.setup_flag != 0:
        raise SystemExit(f"osqp_setup failed on rank {rank}: {s.setup_flag}")
    setup_s = time.perf_counter() - t0
    eng = s.engine_handle()
    for _ in range(max(args.warmup, 3)):
        assert s.solve() == 0
    st0, p0 = s.stats(), plug.get_partition()
    clocks = ClockSampler(local)
    barrier()
    clocks.start()
    plug.timer_start(eng)
    iters = 0
    for _ in range(args.steps):
        assert s.solve() == 0
        iters += int(s.info.iter)
    ms = plug.timer_stop(eng)
    barrier()
    clk = clocks.stop()
    st1 = s.stats()
    t_max = allmax(ms)
    nnz_local = int(sl["A"].nnz)
    import torch
    tt = torch.tensor([float(nnz_local)], dtype=torch.float64, device="cuda")
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(tt)
    nnz_total = int(tt.item())
    pcg_per_admm = (st1["pcg_iters"] - st0["pcg_iters"]) / max(1, st1["admm_iters"] - st0["admm_iters"])
    # one more solve with every launch and collective bracketed by CUDA events (plain launches, not the
    # timed run): where the time of an iteration goes on this rank, the collective's share included
    plug.profile(eng, True)
    assert s.solve() == 0
    prof = plug.profile_get(eng)
    plug.profile(eng, False)
    tot = sum(k["total_ms"] for k in prof) or 1.0
    breakdown = [{"kernel": k["kernel"], "launches": k["launches"],
                  "avg_us": round(1e3 * k["total_ms"] / k["launches"], 2), "share": round(k["total_ms"] / tot, 4)}
                 for k in prof if k["launches"] and not k["kernel"].startswith("tile_")]
    coll = sum(k["total_ms"] for k in prof if k["kernel"].startswith("allreduce")) / tot
    line = None
    if rank == 0:
        # algorithmic bytes of one CG iteration over the WHOLE problem (B_K), for the roofline
        nnzP = 7 * n
        b_k = 12.0 * (2 * nnz_total + nnzP) + 8.0 * (2 * m + 12 * n)
        b_it = 12.0 * (2 * nnz_total + nnzP) + 8.0 * (12 * m + 8 * n) + pcg_per_admm * b_k
        peak = 6539.2
        try:
            peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        except (OSError, ValueError, KeyError):
            pass
        ach = b_it * iters / (t_max * 1e-3) / 1e9
        line = {"metric": "admm_iters_per_s", "value": iters / (t_max * 1e-3), "unit": "ADMM iterations/s",
                "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": t_max / args.steps, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"random_qp_rows n={n} m={m} {args.pnnz} nnz/row, row-partitioned over "
                                       f"{world} GPU(s) (BASELINE config 5 at the default size)",
                           "nnzA": nnz_total, "eps_abs": args.eps, "eps_rel": args.eps,
                           "multi_gpu": "rows of A, l, u, z, y partitioned; x, P, CG scalars replicated; "
                                        "ncclAllReduce(sum, n doubles) after every [P|A'] product",
                           "l2": "inputs larger than L2"},
                "solve": {"status": s.status, "iters_per_step": iters / args.steps, "obj": float(s.info.obj_val),
                          "pri_res": float(s.info.pri_res), "dua_res": float(s.info.dua_res),
                          "pcg_iters_per_admm_iter": round(pcg_per_admm, 3),
                          "allreduces_per_admm_iter": round(2.0 + pcg_per_admm, 2) if world > 1 else 0.0,
                          "allreduces_enqueued_outside_graphs": plug.get_partition()["allreduces"] - p0["allreduces"],
                          "setup_s": round(setup_s, 2), "generate_s": round(gen_s, 2)},
                "speedup_vs_1gpu": round(iters / (t_max * 1e-3) / PARTITIONED_N1_ITS, 3),
                "n1_its": PARTITIONED_N1_ITS,
                "n1_source": "profiles/r1_bench_partitioned_config5_sell_n1.json (same workload, one GPU); "
                             "the N = 1 line of this run's `partitioned` record is the live figure",
                "collective_share_rank0": round(coll, 4),
                "rank0_breakdown": breakdown,
                "roofline": {"bound": "hbm", "kernel": "whole ADMM iteration, all GPUs",
                             "achieved": round(ach, 1), "peak": peak * world, "unit": "GB/s",
                             "frac": round(ach / (peak * world), 4), "traffic": None,
                             "algorithmic_bytes_per_iteration": b_it},
                "cpu_baseline": None, "e2e": None,
                "gpu_launches": st1["kernel_launches"] - st0["kernel_launches"], "clocks": clk}
        if json_out:
            json_out(line)
    s.cleanup()
    qd.init_partition(plug, 0, 1)
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=100000, help="variables (config 2: 1e5; m = 10 n)")
    ap.add_argument("--cpu-n", type=int, default=CPU_SAMPLE_N)
    ap.add_argument("--eps", type=float, default=1e-3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-profile", action="store_true", help="skip per-kernel event timing")
    ap.add_argument("--partitioned", action="store_true",
                    help="ONE QP row-partitioned over the N GPUs (NCCL allreduce inside the CG), strong scaling; "
                         "default workload BASELINE config 5 (n=1e6, m=1e7, 20 nnz/row)")
    ap.add_argument("--pn", type=int, default=1000000, help="variables of the partitioned workload")
    ap.add_argument("--pnnz", type=int, default=20, help="nonzeros per row of A in the partitioned workload")
    ap.add_argument("--no-partitioned-record", action="store_true",
                    help="do not attach the row-partitioned config-5 record to the line")
    ap.add_argument("--workload", default="random_qp", choices=["random_qp", "lasso"],
                    help="random_qp = BASELINE config 2 (the headline); lasso = config 3 (n_f=1e5, 1e6 samples, 5e7 nnz)")
    args = ap.parse_args()
    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    # stdout carries exactly ONE JSON line: anything libraries print on fd 1 meanwhile (NCCL's
    # version banner, CUDA warnings) goes to stderr; json_out() restores fd 1 for the line
    sys.stdout.flush()
    saved_fd = os.dup(1)
    os.dup2(2, 1)

    def json_out(line):
        sys.stdout.flush()
        os.dup2(saved_fd, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)

    os.environ["OSQP_CUDA_DEVICE"] = str(local)
    if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"      # keep NCCL's version banner off stdout (one JSON line only)
    import numpy as np
    import torch
    import torch.distributed as dist
    from osqp_b200 import Solver, problems
    from osqp_b200.capi import Plugin

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the CUDA backend has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(v):
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(v):
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    settings = dict(SETTINGS, eps_abs=args.eps, eps_rel=args.eps)
    if args.partitioned:
        run_partitioned(args, rank, world, local, settings, barrier, allmax, json_out)
        if world > 1:
            dist.destroy_process_group()
        return
    # every rank solves its own copy of the SAME problem (same seed): equal iteration counts, so the
    # N-GPU value measures the hardware and not the spread of ADMM iterations across random seeds
    if args.workload == "lasso":
        prob = problems.lasso(n_feat=args.n, m_samp=10 * args.n, nnz_per_row=50, seed=1)
        wl_name = f"lasso_nf{args.n}_ms{10 * args.n} (BASELINE config 3: ~5e7 nnz)"
    else:
        prob = problems.random_qp(n=args.n, seed=1)
        wl_name = f"random_qp_n{args.n}_m{10 * args.n} (BASELINE config 2: ~1e7 nnz)"
    ab = problems.algorithmic_bytes(prob)
    host, keep, h2d_bytes = pinned_problem(prob)
    plug = Plugin()

    # ---------------------------------------------------------------- resident leg (value)
    s = Solver(host["P"], host["q"], host["A"], host["l"], host["u"], **settings)
    if s.setup_flag != 0:
        raise SystemExit(f"osqp_setup failed on the CUDA backend: {s.setup_flag}")
    eng = s.engine_handle()
    for _ in range(max(args.warmup, 3)):
        assert s.solve() == 0
    st0 = s.stats()
    clocks = ClockSampler(local)
    barrier()
    clocks.start()
    plug.timer_start(eng)
    iters = 0
    for _ in range(args.steps):
        assert s.solve() == 0
        iters += int(s.info.iter)
    ms = plug.timer_stop(eng)
    barrier()
    clk = clocks.stop()
    st1 = s.stats()
    status, obj = s.status, float(s.info.obj_val)
    pri, dua = float(s.info.pri_res), float(s.info.dua_res)
    t_max = allmax(ms)
    total_iters = allsum(iters)
    value = total_iters / (t_max * 1e-3)
    launches = st1["kernel_launches"] - st0["kernel_launches"]
    pcg_per_admm = (st1["pcg_iters"] - st0["pcg_iters"]) / max(1, st1["admm_iters"] - st0["admm_iters"])
    # Kernel breakdown: the SAME steps once more with every launch bracketed by CUDA events on
    # the engine's stream (plain launches instead of graph replays, so it is not the timed run).
    prof, full = [], []
    if not args.no_profile:
        plug.profile(eng, True)
        for _ in range(args.steps):
            assert s.solve() == 0
        prof = plug.profile_get(eng)
        plug.profile(eng, False)
    s.cleanup()
    # Roofline pass: the CG early exit leaves part of the enqueued launches as device-side
    # no-ops, which would dilute a per-launch average; here the CG is forced to exactly 4
    # iterations per ADMM iteration (no convergence test), so every launch is a full pass.
    if not args.no_profile:
        sf = Solver(host["P"], host["q"], host["A"], host["l"], host["u"],
                    **dict(settings, max_iter=60, check_termination=0))
        engf = sf.engine_handle()
        plug.set_pcg(engf, 1e-30, 0.0, 4)
        sf.solve()
        plug.profile(engf, True)
        sf.solve()
        full = plug.profile_get(engf)
        plug.profile(engf, False)
        sf.cleanup()

    # -------------------------------------------------------------------- e2e leg
    def e2e_step():
        s2 = Solver(host["P"], host["q"], host["A"], host["l"], host["u"], **settings)
        assert s2.setup_flag == 0
        assert s2.solve() == 0
        x, y = s2.x, s2.y          # D2H already done by osqp_solve (store_solution); copies here
        it = int(s2.info.iter)
        s2.cleanup()
        return it, x.nbytes + y.nbytes
    for _ in range(max(1, min(args.warmup, 2))):      # untimed: allocation cache, pinned pages, lazy module loads
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    e2e_iters, d2h = 0, 0
    for _ in range(args.steps):
        it, d2h = e2e_step()
        e2e_iters += it
    torch.cuda.synchronize()
    e2e_s = allmax(time.perf_counter() - t0)
    e2e_value = allsum(e2e_iters) / e2e_s
    # D2H per step: solution x, y + scaling vectors D, Dinv, E, Einv and scaled q, l, u mirrors
    d2h_bytes = d2h + 8 * (2 * ab["n"] + 2 * ab["m"]) + 8 * (ab["n"] + 2 * ab["m"])

    # ------------------------------------------------------------------- roofline
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    roof, kernels = None, []
    if prof:
        # "tile_*" records time the product kernel of a column-blocked pass on its own; the same launches
        # are inside spmv_M / rhs / spmv_A, so they count neither towards the total nor as the dominant kernel
        main = [k for k in prof if not k["kernel"].startswith("tile_")]
        tot = sum(k["total_ms"] for k in main) or 1.0
        for k in prof:
            if k["launches"]:
                avg_us = 1e3 * k["total_ms"] / k["launches"]
                kernels.append({"kernel": k["kernel"], "launches": k["launches"], "avg_us": round(avg_us, 2),
                                "share": round(k["total_ms"] / tot, 4),
                                "GBps": round(k["bytes_per_launch"] / (avg_us * 1e-6) / 1e9, 1)})
        dom = max(main, key=lambda k: k["total_ms"])
        domf = [k for k in full if k["kernel"] == dom["kernel"] and k["launches"]]
        src = domf[0] if domf else dom
        avg_s = 1e-3 * src["total_ms"] / max(1, src["launches"])
        ach = src["bytes_per_launch"] / avg_s / 1e9
        roof = {"bound": "hbm", "kernel": dom["kernel"], "achieved": round(ach, 1), "peak": peak,
                "unit": "GB/s", "frac": round(ach / peak, 4),
                "traffic": (load_traffic(args.workload).get(dom["kernel"]) or {}).get("dram_bytes_per_launch"),
                "traffic_source": (load_traffic(args.workload).get(dom["kernel"]) or {}).get("source"),
                "peak_source": peak_src,
                "algorithmic_bytes_per_launch": src["bytes_per_launch"],
                "avg_launch_us": round(avg_s * 1e6, 2),
                "how": "CUDA events around every launch on the engine stream; 'achieved' from a pass with "
                       "the CG forced to 4 full iterations per ADMM iteration (no no-op launches); "
                       "'kernels' = the timed steps re-run with events (avg includes no-op launches)",
                "full_pass_kernels": [
                    {"kernel": k["kernel"], "launches": k["launches"],
                     "avg_us": round(1e3 * k["total_ms"] / k["launches"], 2),
                     "GBps": round(k["bytes_per_launch"] / (1e-3 * k["total_ms"] / k["launches"]) / 1e9, 1)}
                    for k in full if k["launches"]],
                "kernels": kernels}

    # --------------------------------------------------------- CPU baseline (rank 0, N = 1)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and os.path.exists(REF_LIB):
        n = args.cpu_n
        cp = problems.random_qp(n=n, seed=1)
        t0 = time.perf_counter()
        r = Solver(cp["P"], cp["q"], cp["A"], cp["l"], cp["u"], lib_path=REF_LIB, linsys_solver=0, **SETTINGS)
        r.solve()
        dt = time.perf_counter() - t0
        r_it, r_solve, r_setup = int(r.info.iter), float(r.info.solve_time), float(r.info.setup_time)
        r.cleanup()
        # the SAME sample through this library, so that one like-for-like pair exists in the line
        t0 = time.perf_counter()
        g = Solver(cp["P"], cp["q"], cp["A"], cp["l"], cp["u"], **SETTINGS)
        assert g.setup_flag == 0 and g.solve() == 0
        g_e2e = time.perf_counter() - t0
        g.solve()                                 # second solve: resident, warmed
        g_it, g_solve = int(g.info.iter), float(g.info.solve_time)
        g.cleanup()
        cpu = {"value": r_it / r_solve, "unit": "ADMM iterations/s", "cores": 1, "kind": "reference",
               "e2e_value": r_it / dt, "setup_s": round(r_setup, 3), "solve_s": round(r_solve, 3),
               "sample": f"unmodified reference (QDLDL) on random_qp(n={n}, m={10 * n}): {r_it} iters, solve "
                         f"{r_solve:.2f} s after {r_setup:.2f} s of setup (value = iters / solve time, e2e_value = "
                         f"iters / (setup + solve)); n=1e5 does not finish (fill-in)",
               "host_cpus": os.cpu_count(),
               "same_size_gpu": {"workload": f"random_qp(n={n}, m={10 * n})", "iters": g_it,
                                 "value": g_it / g_solve, "e2e_value": g_it / g_e2e,
                                 "ratio_resident": round((g_it / g_solve) / (r_it / r_solve), 2),
                                 "ratio_e2e": round((g_it / g_e2e) / (r_it / dt), 2),
                                 "note": "like-for-like: same problem, same settings, both arms; the headline "
                                         "config (n=1e5) has no CPU counterpart"}}

    part = None
    if not args.no_partitioned_record and args.workload == "random_qp":
        try:
            part = run_partitioned(args, rank, world, local, settings, barrier, allmax, None)
        except Exception as exc:      # the headline line must survive a failure of the attachment
            part = {"error": repr(exc)} if rank == 0 else None
    per_rank = [iters]
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, (iters, ms))
        per_rank = gathered
    if rank == 0:
        line = {"metric": "admm_iters_per_s", "value": value, "unit": "ADMM iterations/s",
                "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": t_max / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": wl_name,
                           "n": ab["n"], "m": ab["m"], "nnzA": ab["nnzA"], "nnzP_full": ab["nnzPf"],
                           "eps_abs": args.eps, "eps_rel": args.eps,
                           "settings": "defaults + adaptive_rho_interval=50, cold start per step",
                           "l2": "inputs larger than L2 (A + [P|A'] = %.0f MB > 126 MB)" % ((ab["spmv_A"] + ab["spmv_M"]) / 1e6),
                           "multi_gpu": "every rank solves its own copy of the same QP, no data-path communication "
                                        "(the row-partitioned single-QP mode is the `partitioned` record)"},
                "solve": {"status": status, "iters_per_step": iters / args.steps, "obj": obj,
                          "pri_res": pri, "dua_res": dua, "time_to_solution_ms": t_max / args.steps,
                          "pcg_iters_per_admm_iter": round(pcg_per_admm, 3),
                          "algorithmic_GBps_whole_solve": round(
                              (iters * (ab["admm_fixed"] + pcg_per_admm * ab["pcg_iter"]) +
                               (iters / 25.0) * ab["check"]) / (ms * 1e-3) / 1e9, 1)},
                "roofline": roof, "cpu_baseline": cpu,
                "e2e": {"value": e2e_value, "unit": "ADMM iterations/s", "h2d_bytes_per_step": h2d_bytes,
                        "d2h_bytes_per_step": d2h_bytes, "ms_per_step": 1e3 * e2e_s / args.steps},
                "gpu_launches": launches, "clocks": clk,
                "per_rank_iters_ms": per_rank, "partitioned": part}
        json_out(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
