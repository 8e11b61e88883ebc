"""Synthetic Criteo-shaped / Amazon-history-shaped inputs for the BASELINE configs.

No network, no datasets: shapes and cardinalities follow the configs, values are seeded.  The
``kaggle`` preset uses the public Criteo-Kaggle cardinalities (builder's choice, not from the
reference) so the tables (2.16 GB at k=16) are HBM- not L2-resident; ``sample`` is the bundled
10k-row file's own ``feat_onehot_dim`` list (L2-resident, 43 605 rows).
"""
from __future__ import annotations

import numpy as np

N_DENSE, N_SPARSE = 13, 26

KAGGLE_VOCAB = [1460, 583, 10131227, 2202608, 305, 24, 12517, 633, 3, 93145, 5683, 8351593, 3194, 27, 14992,
                5461306, 10, 5652, 2173, 4, 7046547, 18, 15, 286181, 105, 142572]
SAMPLE_VOCAB = [176, 387, 5521, 4033, 57, 8, 3185, 94, 4, 2987, 2085, 5284, 1726, 25, 2036, 4724, 10, 1150, 547, 4,
                5037, 8, 13, 2525, 40, 1939]


def dense_feature(feat):
    return {"feat": feat}


def sparse_feature(feat, feat_onehot_dim, embed_dim):
    return {"feat": feat, "feat_onehot_dim": int(feat_onehot_dim), "embed_dim": int(embed_dim)}


def criteo_feature_columns(vocab=SAMPLE_VOCAB, embed_dim=8):
    """Same structure as the reference's ``features_dict`` (DL/utils/dataset.py:69-75)."""
    return [[dense_feature("I%d" % i) for i in range(1, N_DENSE + 1)],
            [sparse_feature("C%d" % (i + 1), v, embed_dim) for i, v in enumerate(vocab)]]


def criteo_batch(B, vocab=SAMPLE_VOCAB, seed=1234, dist="uniform", zipf_a=1.05):
    """-> (dense float32 [B,13] in [0,1), ids int32 [B,26])."""
    rng = np.random.default_rng(seed)
    dense = rng.random((B, N_DENSE), dtype=np.float32)
    ids = np.empty((B, len(vocab)), dtype=np.int32)
    for f, v in enumerate(vocab):
        if dist == "uniform":
            ids[:, f] = rng.integers(0, v, size=B, dtype=np.int64)
        elif dist == "zipf":
            z = rng.zipf(zipf_a, size=B)
            ids[:, f] = (z - 1) % v
        else:
            raise ValueError(dist)
    return dense, ids


def criteo_matrix(dense, ids):
    """The reference's single [B,39] matrix with ids stored as floats (DL/utils/dataset.py:49-52,61)."""
    return np.concatenate([dense.astype(np.float64), ids.astype(np.float64)], axis=1)


def din_feature_columns(v_item=63002, v_user=192404, embed_dim=64):
    """One dense, one other-sparse ('user_id'), the candidate key 'movie_id' is looked up in the
    behaviour table 'movies_seq' (DL/model/din.py:74-79)."""
    return [[dense_feature("age")],
            [sparse_feature("user_id", v_user, embed_dim), sparse_feature("movies_seq", v_item, embed_dim)]]


def din_batch(B, T=100, v_item=63002, v_user=192404, seed=1234):
    """History lengths ~ U{0..T} (about 1% all-padding rows), ids uniform in [1,V), zero-padded at the tail."""
    rng = np.random.default_rng(seed)
    lens = rng.integers(0, T + 1, size=B)
    hist = rng.integers(1, v_item, size=(B, T), dtype=np.int64).astype(np.int32)
    hist[np.arange(T)[None, :] >= lens[:, None]] = 0
    return {"age": rng.random((B, 1), dtype=np.float32),
            "user_id": rng.integers(0, v_user, size=(B, 1), dtype=np.int64).astype(np.int32),
            "movie_id": rng.integers(1, v_item, size=(B, 1), dtype=np.int64).astype(np.int32),
            "movies_seq": hist}
