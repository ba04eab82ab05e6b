"""Development check: responsibilities after each of the first iterations, per layout, against the flat layout."""
import sys
import numpy as np
sys.path.insert(0, ".")
import melissa_b200 as mb
from melissa_b200 import synth

K = int(sys.argv[1]) if len(sys.argv) > 1 else 5
mb.set_option("debug_estep", 1)
d, tr = synth.generate(48, 37, K, n_basis=4, seed=15, sites_per_region=40, cov_lo=0.02, cov_hi=0.25)
delta_0 = np.full(K, 0.5) + 0.01 * np.arange(K)
out = {}
for name, opts in (("flat", dict(sliced=0)), ("b8", dict(sliced=1, cell_block=8)),
                   ("b8again", dict(sliced=1, cell_block=8)), ("b8w16", dict(sliced=1, cell_block=8, sweep_warps=16)),
                   ("b8w8", dict(sliced=1, cell_block=8, sweep_warps=8)), ("b8g3", dict(sliced=1, cell_block=8, sweep_group=3)),
                   ("b8g1", dict(sliced=1, cell_block=8, sweep_group=1)), ("b8g1again", dict(sliced=1, cell_block=8, sweep_group=1))):
    mb.set_option("sweep_warps", 0); mb.set_option("sweep_group", 2)
    for k, v in opts.items():
        mb.set_option(k, v)
    with mb.Engine(d, tr["basis"]) as eng:
        eng.begin(K, tr["w0"], delta_0)
        rs = []
        for it in range(12):
            eng.phase_a()
            rs.append(eng.fetch()["r_nk"].copy())
            eng.phase_b(False)
        out[name] = rs
for name in out:
    if name == "flat":
        continue
    print(name, ["%.1e" % np.abs(a - b).max() for a, b in zip(out[name], out["flat"])])
