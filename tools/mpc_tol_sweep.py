"""CG tolerance vs ADMM / CG iteration counts on the config-4 MPC problem (fused single-block path)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from osqp_b200 import Solver, problems
from osqp_b200.capi import Plugin
plug = Plugin()
base = problems.mpc(horizon=50, nx=20, nu=10, seed=1)
for interval in (25, 50):
    for rel, lam in ((0.3, 0.05), (0.1, 0.05), (0.03, 0.05), (0.01, 0.05), (1e-3, 0.05), (1e-6, 0.0), (0.1, 0.0), (0.1, 0.01), (0.1, 0.2)):
        s = Solver(base["P"], base["q"], base["A"], base["l"], base["u"], eps_abs=1e-3, eps_rel=1e-3,
                   adaptive_rho_interval=interval, warm_start=0)
        eng = s.engine_handle()
        plug.set_pcg(eng, rel, 0.0, 0)
        plug.set_pcg_lambda(eng, lam)
        s.solve(); st0 = s.stats()
        t = time.time(); s.update_rho(0.1); s.solve(); dt = time.time() - t
        st = s.stats()
        print(f"interval={interval} rel={rel:g} lambda={lam:g}: {s.status} iters={int(s.info.iter)} cg_total={st['pcg_iters']-st0['pcg_iters']} "
              f"rho_updates={int(s.info.rho_updates)} solve={1e3*float(s.info.solve_time):.1f} ms", flush=True)
        s.cleanup()
