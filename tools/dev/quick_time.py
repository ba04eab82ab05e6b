"""Per-phase CUDA-event timing of the VB loop on a synthetic shape (development aid)."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import melissa_b200 as mb
from melissa_b200 import synth

N, M, K = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (1000, 1000, 4)
t0 = time.time()
d, tr = synth.generate(N, M, K, n_basis=4, seed=1)
t_gen = time.time() - t0
import os
if os.environ.get('MELISSA_CMMA'):
    mb.set_option('contract_mma', int(os.environ['MELISSA_CMMA']))
if os.environ.get('MELISSA_TN'):
    mb.set_option('cell_block', int(os.environ['MELISSA_TN']))
t0 = time.time()
eng = mb.Engine(d, tr["basis"], K_hint=K)
t_pack = time.time() - t0
eng.begin(K, tr["w0"], np.full(K, 0.51))
ev = lambda: torch.cuda.Event(enable_timing=True)
for _ in range(3):
    eng.phase_a(); eng.phase_b(False)
torch.cuda.synchronize()
ta, tb, tl = [], [], []
for i in range(10):
    e0, e1, e2 = ev(), ev(), ev()
    e0.record(); eng.phase_a(); e1.record(); eng.phase_b(i == 9); e2.record()
    torch.cuda.synchronize()
    ta.append(e0.elapsed_time(e1)); (tl if i == 9 else tb).append(e1.elapsed_time(e2))
eng.profile(True)
for i in range(10):
    eng.phase_a(); eng.phase_b(False)
prof = {k: round(v[0] / max(1, v[1]), 4) for k, v in eng.profile_read().items()}
eng.profile(False)
lb = eng.lower_bound()
res = eng.fetch()
print(json.dumps(dict(N=N, M=M, K=K, P=d.P, nnz=d.nnz, gen_s=round(t_gen, 2), pack_s=round(t_pack, 2),
                      phase_a_ms=float(np.median(ta)), phase_b_ms=float(np.median(tb)), phase_b_lb_ms=tl[0],
                      pairs_per_s=d.P / ((np.median(ta) + np.median(tb)) * 1e-3), lb=lb, kernels_ms=prof,
                      ari=mb.cluster_ari(tr["C_true"], res["r_nk"]))))
