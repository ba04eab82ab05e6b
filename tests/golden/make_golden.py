"""Regenerate the golden fixtures under tests/golden/ (run in the BUILD container only).

Inputs: the reference's bundled datasets, read straight from
/root/reference/data/*.rda with tools/rdata.py (no R needed).  Outputs: small
.npz files holding the flattened data, a frozen initialisation (w0, delta_0), a
frozen train/test split for the ENCODE set, and the CPU oracle's results on
them (oracle/melissa_oracle.py).  Nothing here is read at test time except the
.npz files; /root/reference does not exist on the GPU box.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import melissa_oracle as mo  # noqa: E402
import rdata  # noqa: E402

REF = "/root/reference/data"
OUT = os.path.dirname(os.path.abspath(__file__))


def pack_positions(x):
    """x = (pos + 1000) / 2000 * 2 - 1 (min-max scaling of integer loci,
    reference inst/data-raw/create_synth_data.R:145-149); stored as int16."""
    pos = np.rint(x * 1000.0).astype(np.int16)
    assert np.array_equal((pos.astype(np.float64) + 1000.0) / 2000.0 * 2.0 - 1.0, x), "positions are not on the grid"
    return pos


def partition(d, rng, data_train_prcg, cpg_train_prcg):
    """reference R/utils.R:130-178 with is_synth = TRUE, region_train_prcg = 1:
    each pair is kept whole with probability data_train_prcg, otherwise
    round(cpg_train_prcg * I) of its CpGs stay in the training set.  The R RNG
    cannot be reproduced here, so the split is frozen in the fixture."""
    pp = d["pair_ptr"]
    is_train = np.ones(d["x"].shape[0], dtype=bool)
    for p in range(pp.shape[0] - 1):
        if rng.random_sample() < data_train_prcg:
            continue
        I = int(pp[p + 1] - pp[p])
        keep = np.sort(rng.choice(I, int(round(cpg_train_prcg * I)), replace=False))
        mask = np.zeros(I, dtype=bool)
        mask[keep] = True
        is_train[pp[p]:pp[p + 1]] = mask
    return is_train


def subset(d, mask):
    pp = d["pair_ptr"]
    lens = np.add.reduceat(mask.astype(np.int64), pp[:-1])
    keep = lens > 0
    out = dict(N=d["N"], M=d["M"], pair_cell=d["pair_cell"][keep], pair_region=d["pair_region"][keep],
               pair_ptr=np.concatenate([[0], np.cumsum(lens[keep])]).astype(np.int64),
               x=d["x"][mask], y=d["y"][mask])
    return out


def run(d, K, basis, w0, delta_0, iters):
    tr = {}
    o = mo.melissa_inner(d, K, basis, w0, delta_0, vb_max_iter=iters, trace=tr)
    o = mo.finalize(o, K)
    return o, tr


def main():
    # ---------------- C1: melissa_synth_dt --------------------------------
    d = rdata.melissa_data_to_csr(rdata.read_rda(f"{REF}/melissa_synth_dt.rda")["melissa_synth_dt"])
    basis = dict(M=d["basis_M"], mus=d["basis_mus"], gamma=d["basis_gamma"])
    rng = np.random.RandomState(20190601)
    w0_k2 = 0.5 * d["W_true"] + 0.05 * rng.standard_normal(d["W_true"].shape)
    # K = 3 is melissa()'s default (R/melissa_vb.R:86): two seeds near the truth, one diffuse
    w0_k3 = np.concatenate([w0_k2 + 0.05 * rng.standard_normal(w0_k2.shape),
                            0.3 * rng.standard_normal((d["M"], basis["M"] + 1, 1))], axis=2)
    out = dict(N=d["N"], M=d["M"], pair_cell=d["pair_cell"], pair_region=d["pair_region"],
               pair_ptr=d["pair_ptr"], pos=pack_positions(d["x"]), y=d["y"],
               basis_M=basis["M"], basis_mus=basis["mus"], basis_gamma=basis["gamma"],
               C_true=d["C_true"], pi_true=d["pi_true"], W_true=d["W_true"],
               w0_k2=w0_k2, w0_k3=w0_k3)
    for tag, K, w0, iters in (("k2", 2, w0_k2, 60), ("k3", 3, w0_k3, 45)):
        delta_0 = np.full(K, 0.51)
        o, tr = run(d, K, basis, w0, delta_0, iters)
        print("C1", tag, "lb", o["lb"], "flags", o["flags"], "ari", mo.cluster_ari(d["C_true"], o["r_nk"]))
        for f in ("W", "W_Sigma", "r_nk", "delta", "beta", "pi_k", "lb", "labels"):
            out[f"{tag}_{f}"] = o[f]
        out[f"{tag}_flags"] = o["flags"]
        out[f"{tag}_n_iter"] = o["n_iter"]
        out[f"{tag}_iters"] = iters
        out[f"{tag}_HtEz_init"] = tr["HtEz_init"]
        out[f"{tag}_delta_0"] = delta_0
    np.savez_compressed(f"{OUT}/synth_c1.npz", **out)

    # ---------------- C2: melissa_encode_dt, 50 % CpG hold-out ------------
    d = rdata.melissa_data_to_csr(rdata.read_rda(f"{REF}/melissa_encode_dt.rda")["melissa_encode_dt"])
    basis = dict(M=d["basis_M"], mus=d["basis_mus"], gamma=d["basis_gamma"])
    rng = np.random.RandomState(20190602)
    is_train = partition(d, rng, data_train_prcg=0.2, cpg_train_prcg=0.5)   # vignettes/run_melissa.Rmd:121-123
    train, test = subset(d, is_train), subset(d, ~is_train)
    K = 4
    w0 = 0.5 * d["W_true"] + 0.05 * rng.standard_normal(d["W_true"].shape)
    delta_0 = np.full(K, 0.51)
    iters = 60
    o, tr = run(train, K, basis, w0, delta_0, iters)
    act, pred = mo.impute_test_met(o, test)
    perf = mo.eval_imputation_performance(act, pred)
    print("C2 lb", o["lb"], "flags", o["flags"], "ari", mo.cluster_ari(d["C_true"], o["r_nk"]), perf,
          "pi", o["pi_k"], d["pi_true"])
    out = dict(N=d["N"], M=d["M"], pair_cell=d["pair_cell"], pair_region=d["pair_region"],
               pair_ptr=d["pair_ptr"], pos=pack_positions(d["x"]), y=d["y"], is_train=np.packbits(is_train),
               basis_M=basis["M"], basis_mus=basis["mus"], basis_gamma=basis["gamma"],
               C_true=d["C_true"], pi_true=d["pi_true"], W_true=d["W_true"], w0=w0, delta_0=delta_0,
               iters=iters, flags=o["flags"], n_iter=o["n_iter"], auc=perf["auc"], f_measure=perf["f_measure"],
               pred_head=pred[:2000])
    for f in ("W", "r_nk", "delta", "beta", "pi_k", "lb", "labels"):
        out[f] = o[f]
    out["W_Sigma_head"] = o["W_Sigma"][:, :5]
    out["HtEz_init_head"] = tr["HtEz_init"][:512]
    np.savez_compressed(f"{OUT}/encode_c2.npz", **out)
    for f in ("synth_c1.npz", "encode_c2.npz"):
        print(f, os.path.getsize(f"{OUT}/{f}"))


if __name__ == "__main__":
    main()
