"""Generate the committed fixtures under tests/golden/.

Run in the build container only (needs /root/reference, which the GPU box does
not have):

    python tests/golden/make_golden.py

1. ``criteo_cfg1.npz`` -- BASELINE config 1 inputs, produced by IMPORTING the
   reference's own loader (DL/utils/dataset.py:36-75; it needs only
   pandas/sklearn, not TensorFlow) on the bundled ``train_1w.txt``.  The loader's
   split is unseeded (dataset.py:63); sklearn draws from numpy's global RNG when
   ``random_state`` is None, so ``np.random.seed(0)`` before each call makes it
   reproducible without touching the reference.  The 'fm' mode returns a dense
   one-hot matrix [N, 43604] (349 MB for the test split) that cannot be
   committed; the script PROVES that it equals the expansion of the 'DeepFM'-mode
   label-encoded ids (column = 13 + offset_f + id) and stores ids + dense only.
2. ``layers_tiny.npz`` -- seeded tiny weights/inputs and the float64 oracle
   outputs for every hot-path layer.  These come from ``oracle/ctr_oracle.py``
   (NOT from the reference -- TensorFlow is unavailable, parity is unpinned);
   they pin the oracle against regressions and travel to the GPU box.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF_DL = "/root/reference/algorithm/deep_learning"
CRITEO = "/root/reference/algorithm/data/criteo/train_1w.txt"


def make_criteo():
    sys.path.insert(0, REF_DL)
    from utils.dataset import create_criteo_dataset, features_dict  # the reference's loader

    np.random.seed(0)
    (Xtr_d, ytr_d), (Xte_d, yte_d) = create_criteo_dataset("DeepFM", CRITEO, test_size=0.2)
    np.random.seed(0)
    (Xtr_o, ytr_o), (Xte_o, yte_o) = create_criteo_dataset("fm", CRITEO, test_size=0.2)
    assert np.array_equal(yte_d, yte_o), "seeded splits differ"
    Xte_d = np.asarray(Xte_d, dtype=np.float64)
    Xte_o = np.asarray(Xte_o, dtype=np.float32)          # object dtype under pandas 3
    fc = features_dict(CRITEO)
    feat_onehot_dim = np.array([f["feat_onehot_dim"] for f in fc[1]], dtype=np.int64)

    dense = Xte_d[:, :13].astype(np.float32)
    ids = Xte_d[:, 13:].astype(np.int32)
    # get_dummies widths per field = number of distinct labels over ALL rows
    Xall = np.concatenate([np.asarray(Xtr_d, dtype=np.float64), Xte_d])[:, 13:].astype(np.int64)
    onehot_width = Xall.max(axis=0) + 1
    offsets = 13 + np.concatenate([[0], np.cumsum(onehot_width)[:-1]])
    assert 13 + onehot_width.sum() == Xte_o.shape[1], (onehot_width.sum(), Xte_o.shape)
    # prove: dense one-hot matrix == expansion of (dense, ids)
    rebuilt = np.zeros_like(Xte_o)
    rebuilt[:, :13] = dense
    rows = np.arange(ids.shape[0])
    for f in range(26):
        rebuilt[rows, offsets[f] + ids[:, f]] = 1.0
    assert np.array_equal(rebuilt, Xte_o), "one-hot expansion mismatch"
    np.savez_compressed(
        os.path.join(HERE, "criteo_cfg1.npz"),
        dense=dense, ids=ids, y=np.asarray(yte_d, dtype=np.int8),
        onehot_width=onehot_width.astype(np.int64), onehot_offsets=offsets.astype(np.int64),
        feat_onehot_dim=feat_onehot_dim, onehot_cols=np.int64(Xte_o.shape[1]))
    print("criteo_cfg1.npz: X_test", Xte_o.shape, "nnz/row", float((Xte_o != 0).sum(1).mean()))


def make_criteo_raw():
    """``criteo_raw.npz``: the RAW columns of the bundled train_1w.txt (10 000 rows) as integers -- labels, the 13 count
    features (with a missing mask) and the 26 categorical tokens as the 32-bit values of their 8 hex digits (with a missing
    mask) -- plus ``test_index``: which raw rows the reference loader's seeded split (see make_criteo) put in X_test, in order.
    Data, not code: this is what the GPU input step (csrc/encode.cu) is fed on the GPU box, where /root/reference does not
    exist; tests/test_gpu_encode.py re-serialises it to TSV text and must reproduce criteo_cfg1.npz (the reference loader's
    own output) from it."""
    import pandas as pd
    from sklearn.model_selection import train_test_split
    names = ["label"] + ["I%d" % i for i in range(1, 14)] + ["C%d" % i for i in range(1, 27)]
    df = pd.read_csv(CRITEO, sep="\t", header=None, names=names, dtype={n: str for n in names if n.startswith("C")})
    N = len(df)
    dense = df[names[1:14]]
    dense_missing = dense.isna().to_numpy()
    dense_raw = dense.fillna(0).to_numpy().astype(np.int64)
    assert np.array_equal(dense_raw.astype(np.float64), dense.fillna(0).to_numpy().astype(np.float64))   # counts are integers
    tok = df[names[14:]]
    tok_missing = tok.isna().to_numpy()
    toks = np.zeros((N, 26), dtype=np.uint32)
    for j, c in enumerate(names[14:]):
        col = tok[c]
        ok = ~col.isna()
        assert col[ok].str.fullmatch(r"[0-9a-f]{8}").all(), c        # fixed-width lowercase hex: string order == numeric order
        toks[ok.to_numpy(), j] = col[ok].map(lambda t: int(t, 16)).to_numpy().astype(np.uint32)
    np.random.seed(0)
    _, test_index = train_test_split(np.arange(N), test_size=0.2)    # same RNG draw as make_criteo's split of X
    # proof that this IS the reference's split: its X_test dense block equals the MinMax-scaled rows at test_index
    z = np.load(os.path.join(HERE, "criteo_cfg1.npz"))
    filled = dense.fillna(0).to_numpy().astype(np.float64)
    lo, hi = filled.min(0), filled.max(0)
    scaled = (filled - lo) / np.where(hi > lo, hi - lo, 1.0)
    assert np.allclose(scaled[test_index].astype(np.float32), z["dense"], atol=1e-7), "test_index does not reproduce the split"
    np.savez_compressed(os.path.join(HERE, "criteo_raw.npz"), label=df["label"].to_numpy().astype(np.int8),
                        dense_raw=dense_raw, dense_missing=dense_missing, tokens=toks, token_missing=tok_missing,
                        test_index=test_index.astype(np.int64))
    print("criteo_raw.npz:", N, "rows;", int(dense_missing.sum()), "missing counts,", int(tok_missing.sum()), "missing tokens")


def make_layers():
    from oracle import ctr_oracle as O
    from oracle.synth import tiny_case
    case = tiny_case(seed=7)
    out = {}
    for name, (fn, args, arrays) in case.items():
        out.update({"%s/%s" % (name, k): a for k, a in arrays.items()})
        out["%s/expected" % name] = fn(*args)
    np.savez_compressed(os.path.join(HERE, "layers_tiny.npz"), **out)
    print("layers_tiny.npz:", len(out), "arrays")


if __name__ == "__main__":
    if os.path.isdir(REF_DL):
        make_criteo()
        make_criteo_raw()
    else:
        print("reference not mounted; skipping criteo_cfg1.npz")
    make_layers()
