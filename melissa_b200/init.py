"""Default initialisation of the cluster means when `w` is not supplied
(reference R/melissa_vb.R:134-194): per-(cell, region) penalised probit MLE,
k-means on the stacked profiles, `vb_init_nstart` short VB runs to rank the
restarts.  The restarts run on the GPU through the same library call as the
final fit, and so does the MLE (melissa_profiles_mle_device); the profile
matrix and Lloyd's k-means live on the GPU as torch tensors (SURVEY.md 8f rank 2:
parity here is statistical, the reference path depends on R's RNG, `optim` and
`stats::kmeans`).
"""
from __future__ import annotations

import math

import numpy as np



def _kmeans(X, K, gen, n_iter=10):
    """Plain Lloyd k-means on the GPU, one random start (role of stats::kmeans(W_tmp, K, nstart = 1)).
    X: torch CUDA tensor [N, M D]; returns the labels as a CUDA tensor."""
    import torch
    N = X.shape[0]
    centres = X[torch.randperm(N, generator=gen, device=X.device)[:K]].clone()
    x2 = (X * X).sum(1)
    lab = torch.zeros(N, dtype=torch.int64, device=X.device)
    for it in range(n_iter):
        d2 = x2[:, None] - 2.0 * (X @ centres.T) + (centres * centres).sum(1)[None]
        new = d2.argmin(1)
        if it > 0 and torch.equal(new, lab):
            break
        lab = new
        sums = torch.zeros_like(centres).index_add_(0, lab, X)
        cnt = torch.bincount(lab, minlength=K).to(X.dtype)
        keep = cnt > 0
        centres[keep] = sums[keep] / cnt[keep, None]
    return lab


def initial_w(engine, data, basis, K, delta_0, alpha_0, beta_0, epsilon_conv, is_kmeans,
              vb_init_nstart, vb_init_max_iter, rng, replicas=()):
    """R/melissa_vb.R:134-194.  Keeps the reference's two quirks: the restart
    whose mini-VB ends with the best lower bound contributes its INITIAL w
    (:187-193), and vb_init_nstart == 1 skips the mini-VB (:174-177).
    The per-pair MLE, the N x (M D) profile matrix and the k-means all stay on
    the GPU; what comes back to the host are K profiles per restart.

    `replicas`: further Engines that hold the SAME data (normally one per extra GPU).  The restarts' mini-VB runs
    are independent -- same data, different w0 -- so they are handed out over `engine` and the replicas, one
    host thread each, no exchange between the GPUs.  The candidates are still drawn in order on `engine`'s GPU
    and the winner is chosen by the reference's rule in restart order, so the result is the one the sequential
    loop gives (every run is bit-reproducible)."""
    import torch
    N, M, D = data.N, data.M, int(basis["M"]) + 1
    dev = torch.device(f"cuda:{engine.device}")
    gen = torch.Generator(device=dev)
    gen.manual_seed(int(rng.integers(0, 2 ** 62)))
    w_mle = engine.profiles_mle_device(0.5, 15)          # [P, D] on the GPU; no CPU path (hand-placed centres: pass `w`)
    X3 = 0.01 * torch.randn((N, D, M), generator=gen, device=dev, dtype=torch.float64)          # :148, as c(ww[,,n])
    pc = torch.as_tensor(np.asarray(data.pair_cell), device=dev).long()
    pr = torch.as_tensor(np.asarray(data.pair_region), device=dev).long()
    X3[pc, :, pr] = w_mle                                                     # :151
    del w_mle, pc, pr
    W_tmp = X3.view(N, D * M)
    profile = lambda n: X3[int(n)].T.cpu().numpy()                            # ww[, , n]: M x D
    candidates = []
    for _t in range(int(vb_init_nstart)):
        w = np.zeros((M, D, K))
        if is_kmeans:
            lab = _kmeans(W_tmp, K, gen)
            for k in range(K):
                members = torch.nonzero(lab == k).flatten()
                if members.numel() == 0:                                      # the reference leaves this as a TODO (:165)
                    members = torch.arange(N, device=dev)
                pick = members[torch.randint(members.numel(), (1,), generator=gen, device=dev)]
                w[:, :, k] = profile(pick)                                    # :168
        else:
            for k, n in enumerate(torch.randperm(N, generator=gen, device=dev)[:K].tolist()):
                w[:, :, k] = profile(n)                                       # :171
        if vb_init_nstart == 1:
            return w
        candidates.append(w)
    del X3, W_tmp
    n_it = max(2, int(vb_init_max_iter))
    mini_lb = lambda eng, w: eng.run(K, w, delta_0, alpha_0, beta_0, n_it, epsilon_conv)["lb"][-1]
    engines = [engine] + list(replicas)
    if len(engines) == 1:
        lbs = [mini_lb(engine, w) for w in candidates]
    else:
        import queue
        import threading
        todo = queue.SimpleQueue()
        for t in range(len(candidates)):
            todo.put(t)
        lbs, errors = [None] * len(candidates), []

        def work(eng):                                   # (the library call releases the GIL)
            try:
                while True:
                    try:
                        t = todo.get_nowait()
                    except queue.Empty:
                        return
                    lbs[t] = mini_lb(eng, candidates[t])
            except Exception as exc:                     # reported below, never swallowed
                errors.append(exc)
        threads = [threading.Thread(target=work, args=(e,)) for e in engines]
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        if errors:
            raise errors[0]
    lb_prev = -1e120
    optimal_w = None
    for w, lb_cur in zip(candidates, lbs):               # :187-193, in restart order
        if lb_cur > lb_prev:
            optimal_w, lb_prev = w, lb_cur
    if optimal_w is None:
        raise FloatingPointError("no restart of the default initialisation ended with a finite lower bound")
    return optimal_w
