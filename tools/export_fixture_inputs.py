"""Writes what tools/rebaseline_with_R.R needs from the golden fixtures as raw little-endian doubles (no R needed):
w0 (R's column-major array c(M, D, K)), delta_0, the frozen CpG hold-out mask of config 2 and a runs.csv."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
OUT = os.path.join(GOLD, "r_inputs")


def main():
    os.makedirs(OUT, exist_ok=True)
    rows = ["tag,dataset,K,iters,mask,nnz"]
    c1 = dict(np.load(os.path.join(GOLD, "synth_c1.npz")))
    for tag, K in (("k2", 2), ("k3", 3)):
        np.asfortranarray(c1[f"w0_{tag}"]).ravel(order="F").astype("<f8").tofile(os.path.join(OUT, f"c1_{tag}_w0.bin"))
        c1[f"{tag}_delta_0"].astype("<f8").tofile(os.path.join(OUT, f"c1_{tag}_delta_0.bin"))
        rows.append(f"c1_{tag},synth,{K},{int(c1[f'{tag}_iters'])},,0")
    c2 = dict(np.load(os.path.join(GOLD, "encode_c2.npz")))
    nnz = int(c2["pos"].shape[0])
    np.asfortranarray(c2["w0"]).ravel(order="F").astype("<f8").tofile(os.path.join(OUT, "c2_w0.bin"))
    c2["delta_0"].astype("<f8").tofile(os.path.join(OUT, "c2_delta_0.bin"))
    np.unpackbits(c2["is_train"])[:nnz].astype(np.int8).tofile(os.path.join(OUT, "c2_train_mask.bin"))
    rows.append(f"c2,encode,4,{int(c2['iters'])},c2_train_mask.bin,{nnz}")
    open(os.path.join(OUT, "runs.csv"), "w").write("\n".join(rows) + "\n")
    print("wrote", OUT)


if __name__ == "__main__":
    main()
