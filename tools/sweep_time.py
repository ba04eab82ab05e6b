"""Per-kernel CUDA-event timing of the VB loop on a device-generated synthetic shape (development aid).

    python tools/sweep_time.py N M K [iters] [key=value ...]      # key: any melissa_set_option knob
"""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import melissa_b200 as mb
from melissa_b200 import synth

pos = [a for a in sys.argv[1:] if "=" not in a]
N, M, K = (int(a) for a in pos[:3])
iters = int(pos[3]) if len(pos) > 3 else 10
for kv in (a for a in sys.argv[1:] if "=" in a):
    k, v = kv.split("=")
    mb.set_option(k, int(v))
d, tr = synth.generate_device(N, M, K, n_basis=4, seed=1, chunk_cells=25)
torch.cuda.synchronize()
t0 = time.time()
eng = mb.Engine(d, tr["basis"], K_hint=K)
t_pack = time.time() - t0
P = int(d["pair_cell"].shape[0])
nnz = int(d["x"].shape[0])
del d
eng.begin(K, tr["w0"], np.full(K, 0.51))
for _ in range(3):
    eng.phase_a(); eng.phase_b(False)
eng.sync()
eng.profile(True)
for i in range(iters):
    eng.phase_a(); eng.phase_b(False)
prof = {k: round(v[0] / max(1, v[1]), 4) for k, v in eng.profile_read().items()}
eng.profile(False)
eng.profile(True)
eng.phase_a(); eng.phase_b(True)
prof_lb = {k: round(v[0] / max(1, v[1]), 4) for k, v in eng.profile_read().items()}
eng.profile(False)
lb = eng.lower_bound()
res = eng.fetch()
info = eng.layout_info()
alg = 9 * nnz + P * (16 * 5 + 8)
print(json.dumps(dict(N=N, M=M, K=K, P=P, nnz=nnz, pack_s=round(t_pack, 2), kernels_ms=prof, sweep_lb_ms=prof_lb["sweep"],
                      hbm_frac=alg / (prof["sweep"] * 1e-3) / 6551e9, layout=info, lb=lb,
                      ari=mb.cluster_ari(tr["C_true"], res["r_nk"]))))
