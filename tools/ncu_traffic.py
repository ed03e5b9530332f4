"""DRAM traffic per launch of the bench's dominant kernels from an `ncu --set full` report ->
profiles/ncu_traffic.json (read by bench.py for roofline.traffic; nothing is typed in by hand).
    python tools/ncu_traffic.py X.ncu-rep "spmv_A(rho.*A p)=spmv_tile_rows_kernel<EpiW>" \
        "spmv_M([P|A'][p;w])=spmv_tile_kernel<EpiKp>+k_tile_epi<EpiKp"
A bench kernel made of several launches (tile kernel + combine kernel) sums its parts; of several
captured launches of one function the largest (a full pass, not a device-side no-op) is taken."""
import csv, io, json, os, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h, units = rows[0], rows[1]


def mb(d, u, k):
    v = float(d[k].replace(",", ""))
    return v * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u[k]]


launches = []
for r in rows[2:]:
    d = dict(zip(h, r)); u = dict(zip(h, units))
    launches.append((d["Kernel Name"], mb(d, u, "dram__bytes_read.sum") + mb(d, u, "dram__bytes_write.sum")))
path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ncu_traffic.json")
allres = json.load(open(path)) if os.path.exists(path) else {}
# --workload NAME: the entries go under that key (config 2 = the top level, what bench.py reads by default)
args = sys.argv[2:]
wl = None
if args and args[0] == "--workload":
    wl, args = args[1], args[2:]
res = allres.setdefault(wl, {}) if wl else allres
for spec in args:
    name, pats = spec.split("=", 1)
    total, parts = 0.0, []
    for pat in pats.split("+"):
        cand = [b for k, b in launches if pat in k]
        if not cand:
            raise SystemExit(f"no launch matches {pat!r}")
        total += max(cand); parts.append(f"{pat}: {max(cand) / 1e6:.1f} MB")
    res[name] = {"dram_bytes_per_launch": total, "source": f"{os.path.basename(rep)} ({'; '.join(parts)})"}
    print(name, res[name])
json.dump(allres, open(path, "w"), indent=1)
