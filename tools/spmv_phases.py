"""Which phase of the stream SpMV costs what: run config 2 with the gathers and/or the row reduction disabled."""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from osqp_b200 import Solver, problems
from osqp_b200.capi import Plugin, PLUGIN_LIB
plug = Plugin()
n = 100000
flag_list = [int(a) for a in sys.argv[1:]] or [0, 1, 2, 3]   # the flags only act on the stream kernel
p = problems.random_qp(n=n, seed=1)
for flags in flag_list:
    s = Solver(p['P'], p['q'], p['A'], p['l'], p['u'], adaptive_rho_interval=50, warm_start=0, max_iter=100, check_termination=0)
    plug.lib.osqp_cuda_debug_set(flags)
    eng = s.engine_handle()
    plug.set_pcg(eng, 1e-30, 0.0, 4)       # exactly 4 CG iterations per ADMM iteration, no early exit
    s.solve()
    plug.profile(eng, True)
    s.solve()
    prof = plug.profile_get(eng)
    st = s.stats()
    print(f"flags={flags} (bit0 no gathers, bit1 no row reduction) pcg_iters={st['pcg_iters']}")
    for k in prof:
        if k['launches']:
            print(f"    {k['kernel']:28s} launches={k['launches']:5d} avg={1e3*k['total_ms']/k['launches']:8.2f} us")
    plug.lib.osqp_cuda_debug_set(0)
    s.cleanup()
