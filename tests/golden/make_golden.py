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
    else:
        print("reference not mounted; skipping criteo_cfg1.npz")
    make_layers()
