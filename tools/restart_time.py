"""Default initialisation with vb_init_nstart restarts: sequential on one GPU against replicas over all GPUs of the box
(development aid; the numbers in DESIGN.md section 5 come from here).

    python tools/restart_time.py N M K [nstart] [iters]
"""
import json
import sys
import threading
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import melissa_b200 as mb
from melissa_b200 import synth
from melissa_b200.init import initial_w

N, M, K = (int(a) for a in sys.argv[1:4])
nstart = int(sys.argv[4]) if len(sys.argv) > 4 else 10
iters = int(sys.argv[5]) if len(sys.argv) > 5 else 20
G = torch.cuda.device_count()
d, tr = synth.generate(N, M, K, n_basis=4, seed=3)
args = dict(K=K, delta_0=np.full(K, 0.55), alpha_0=0.5, beta_0=None, epsilon_conv=1e-5, is_kmeans=True,
            vb_init_nstart=nstart, vb_init_max_iter=iters)
engines = [None] * G
t0 = time.perf_counter()
th = [threading.Thread(target=lambda i=i: engines.__setitem__(i, mb.Engine(d, tr["basis"], device=i, K_hint=K))) for i in range(G)]
[t.start() for t in th]
[t.join() for t in th]
t_pack = time.perf_counter() - t0
out = {}
for name, reps in (("sequential", []), ("replicas", engines[1:])):
    for e in engines:
        e.sync()
    t0 = time.perf_counter()
    w = initial_w(engines[0], d, tr["basis"], rng=np.random.default_rng(1), replicas=reps, **args)
    out[name] = dict(seconds=round(time.perf_counter() - t0, 3), checksum=float(np.abs(w).sum()))
print(json.dumps(dict(N=N, M=M, K=K, gpus=G, nstart=nstart, mini_iters=iters, pack_all_s=round(t_pack, 2), **out,
                      same_start=out["sequential"]["checksum"] == out["replicas"]["checksum"])))
