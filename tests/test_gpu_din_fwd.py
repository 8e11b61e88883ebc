"""GPU parity for the one-launch DIN front end (ctrhot_din_fwd_tf32, csrc/din_fwd.cu) against the float64 oracle of
DIN.call's first half (DL/model/din.py:56-90): candidate / history lookups, Attention.call (DL/layer/interaction.py:369-406),
the other sparse lookups and the dense columns, concatenated as the tower's input ``[att | item | other | dense]``.

The lookups and the dense columns are copies: bit-exact.  The attention output goes through TF32 MMAs with the first layer
regrouped as ``u + k (Wb + diag(q) Wd)``: 2e-3 absolute (BASELINE north_star's tolerance for tensor-core paths), and it must
agree with the older two-launch unit (ctrhot_din_attention_gather_tf32) to the same bar.  Cases: the BASELINE config-4 shape,
ragged batch sizes (1 .. more samples per CTA than one prologue batch), all-padding and full histories, K = 32 and 64, the
largest hidden sizes, T = 128 and a short T, float ids, out-of-range ids.
"""
import numpy as np
import pytest
import torch

from oracle import ctr_oracle as O
from recommender_system_b200.layer import Attention
from test_gpu_din_tc import make_case, npy

pytestmark = pytest.mark.gpu
TOL = 2e-3


def front(table, hist, cand, ws, T, K, hidden, extras=(), dense=None, ids_float=False):
    L = Attention(hidden, "prelu", precision="tf32")
    L.build_for(T, K)
    L.set_weights(ws)
    assert L.front_fits(T, K)
    dev = L.device
    B = hist.shape[0]
    width = 2 * K + sum(t.shape[1] for t, _ in extras) + (0 if dense is None else dense.shape[1])
    ld = (width + 3) // 4 * 4
    emb = torch.full((B, ld), float("nan"), dtype=torch.float32, device=dev)
    conv = (lambda a: torch.from_numpy(a.astype(np.float32)).to(dev)) if ids_float else (lambda a: torch.from_numpy(a).to(dev))
    ex, col = [], 2 * K
    for t, ids in extras:
        ex.append((torch.from_numpy(t).to(dev), conv(ids), col))
        col += t.shape[1]
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    L.front(torch.from_numpy(table).to(dev), conv(hist), conv(cand), emb, ex,
            None if dense is None else torch.from_numpy(dense).to(dev), col, flag)
    torch.cuda.synchronize()
    return npy(emb)[:, :width], int(flag.item()), L


def check(B, T, K, hidden, V, seed, n_extra=1, n_dense=1, ids_float=False, scale=1.0, tol=TOL):
    table, hist, cand, ws, want, q = make_case(B, T, K, hidden, V, seed, scale)
    rng = np.random.default_rng(seed + 1)
    extras = []
    for j in range(n_extra):
        dim = [K, 8, 5, 12][j % 4]
        rows = 50 + 37 * j
        extras.append((np.asarray(rng.normal(0, 1, (rows, dim)), np.float32), rng.integers(0, rows, (B, 1)).astype(np.int32)))
    dense = np.asarray(rng.random((B, n_dense)), np.float32) if n_dense else None
    got, flag, L = front(table, hist, cand, ws, T, K, hidden, extras, dense, ids_float)
    assert flag == 0
    np.testing.assert_array_equal(got[:, K:2 * K], q)                              # candidate rows: bit-exact copies
    col = 2 * K
    for t, ids in extras:
        np.testing.assert_array_equal(got[:, col:col + t.shape[1]], t[ids[:, 0]].astype(np.float64))
        col += t.shape[1]
    if dense is not None:
        np.testing.assert_array_equal(got[:, col:], dense.astype(np.float64))
    err = np.max(np.abs(got[:, :K] - want))
    assert err <= tol, err
    return got, table, hist, cand, ws, L


@pytest.mark.parametrize("B", [1, 3, 200, 1000, 148 * 130 + 7])
def test_config4_shape(lib, B):
    T, K, hidden = 100, 64, (80, 40)
    got, table, hist, cand, ws, L = check(B, T, K, hidden, 5000, seed=B)
    if B > 2:                                                    # all-padding history -> uniform average of the T keys
        np.testing.assert_allclose(got[0, :K], table[hist[0]].astype(np.float64).mean(0), rtol=1e-5, atol=1e-6)
    # the two-launch unit of round 1 computes the same thing with the (q * k) product in the A operand
    dev = L.device
    att, _ = L.call_gather(torch.from_numpy(table).to(dev), torch.from_numpy(hist).to(dev), torch.from_numpy(cand).to(dev), None)
    assert np.max(np.abs(npy(att) - got[:, :K])) <= TOL


@pytest.mark.parametrize("T,K,hidden,n_extra,n_dense", [(128, 32, (96, 64), 3, 2), (37, 32, (24, 8), 0, 0), (8, 64, (16, 16), 8, 3),
                                                         (100, 64, (64, 32), 2, 0), (1, 32, (8, 1), 1, 1), (128, 64, (96, 64), 1, 1)])
def test_shapes(lib, T, K, hidden, n_extra, n_dense):
    check(300, T, K, hidden, 999, seed=T + K, n_extra=n_extra, n_dense=n_dense)


def test_float_ids(lib):
    """The reference feeds ids as float columns (DL/utils/dataset.py:61): truncation toward zero, same results."""
    a = check(257, 100, 64, (80, 40), 4000, seed=9, ids_float=True)[0]
    b = check(257, 100, 64, (80, 40), 4000, seed=9, ids_float=False)[0]
    np.testing.assert_array_equal(a, b)


def test_large_scores_keep_the_softmax_sharp(lib):
    check(500, 100, 64, (80, 40), 3000, seed=77, scale=4.0, tol=8 * TOL)          # keys are O(2): relative to their size


def test_bad_ids_raise_the_flag_and_read_nothing(lib):
    T, K, hidden = 20, 32, (16, 8)
    table, hist, cand, ws, want, q = make_case(64, T, K, hidden, 100, seed=3)
    hist[5, 2] = 100                                                               # one past the table
    got, flag, _ = front(table, hist, cand, ws, T, K, hidden)
    assert flag == 1 and np.isfinite(got[:, :2 * K]).all()
    hist[5, 2] = 1
    cand[7, 0] = -3
    got, flag, _ = front(table, hist, cand, ws, T, K, hidden)
    assert flag == 1 and np.all(got[7, K:2 * K] == 0)
    cand[7, 0] = 1
    ex = [(np.ones((10, 4), np.float32), np.full((64, 1), 10, np.int32))]
    got, flag, _ = front(table, hist, cand, ws, T, K, hidden, ex)
    assert flag == 1 and np.all(got[:, 2 * K:2 * K + 4] == 0)


def test_full_batch_properties(lib):
    """B = 16384 (BASELINE config 4): deterministic, sample-permutation equivariant bit for bit, a strided subset matches
    the oracle."""
    T, K, hidden, B, V = 100, 64, (80, 40), 16384, 63002
    table, _, _, ws, _, _ = make_case(8, T, K, hidden, V, seed=5)
    rng = np.random.default_rng(1)
    lens = rng.integers(0, T + 1, size=B)
    hist = rng.integers(1, V, size=(B, T)).astype(np.int32)
    hist[np.arange(T)[None, :] >= lens[:, None]] = 0
    cand = rng.integers(1, V, size=(B, 1)).astype(np.int32)
    got, flag, _ = front(table, hist, cand, ws, T, K, hidden)
    again, _, _ = front(table, hist, cand, ws, T, K, hidden)
    np.testing.assert_array_equal(got, again)
    perm = rng.permutation(B)
    got_p, _, _ = front(table, hist[perm], cand[perm], ws, T, K, hidden)
    np.testing.assert_array_equal(got[perm], got_p)
    sub = np.arange(0, B, 331)
    dense = [(ws[3 * i].astype(np.float64), ws[3 * i + 1].astype(np.float64), ws[3 * i + 2].astype(np.float64)) for i in range(2)]
    key = table[hist[sub]].astype(np.float64)
    want = O.din_attention(table[cand[sub, 0]].astype(np.float64), key, key, (hist[sub] != 0).astype(np.float64), dense,
                           (ws[6].astype(np.float64), ws[7].astype(np.float64)))
    assert np.max(np.abs(got[sub, :K] - want)) <= TOL


@pytest.mark.parametrize("prec", ["fp32", "tf32"])
def test_din_with_two_behaviour_features(lib, prec):
    """behavior_feature_list of two features (DL/model/din.py:76-79: key = concat of their rows, mask from the first): the rows
    go through the gather kernel and the attention unit, no eager indexing; against the oracle's DIN.call."""
    from recommender_system_b200.layer import set_seed
    from recommender_system_b200.model.din import DIN
    set_seed(3)
    B, T = 97, 9
    fc = [[{"feat": "d0"}],
          [{"feat": "o0", "feat_onehot_dim": 11, "embed_dim": 8}, {"feat": "seq_a", "feat_onehot_dim": 40, "embed_dim": 8},
           {"feat": "seq_b", "feat_onehot_dim": 23, "embed_dim": 4}]]
    m = DIN(fc, ["seq_a", "seq_b"], att_hidden_units=(16, 8), dnn_hidden_units=(16, 8), precision=prec)
    m.build(None)
    m.att_layer.build_for(T, 12)
    rng = np.random.default_rng(5)
    upd = {}
    for n, w in m.named_weights().items():                       # non-trivial alphas / BN statistics / biases
        if n.endswith("prelu_alpha"):
            upd[n] = rng.uniform(0, 0.3, tuple(w.shape))
        elif n.endswith("moving_variance") or n.endswith("gamma"):
            upd[n] = rng.uniform(0.5, 1.5, tuple(w.shape))
        elif n.endswith("bias") or n.endswith("beta") or n.endswith("moving_mean"):
            upd[n] = rng.normal(0, 0.1, tuple(w.shape))
    m.set_weights({n: np.asarray(v, np.float32) for n, v in upd.items()})
    lens = rng.integers(0, T + 1, size=B)
    lens[0], lens[1] = 0, T
    ha = rng.integers(1, 40, size=(B, T)).astype(np.int32)
    ha[np.arange(T)[None, :] >= lens[:, None]] = 0
    hb = rng.integers(0, 23, size=(B, T)).astype(np.int32)          # the second feature does not carry the mask
    cand = np.stack([rng.integers(1, 40, size=B), rng.integers(0, 23, size=B)], axis=1).astype(np.int32)
    inputs = {"d0": np.asarray(rng.random((B, 1)), np.float32), "o0": rng.integers(0, 11, size=(B, 1)).astype(np.int32),
              "seq_a": ha, "seq_b": hb, "movie_id": cand}
    got = npy(m(inputs))
    w = {n: npy(t) for n, t in m.named_weights().items()}
    att = [(w["Attention_0/dense_%d/kernel" % i], w["Attention_0/dense_%d/bias" % i], w["Attention_0/dense_%d/prelu_alpha" % i])
           for i in range(2)]
    tw = [(w["dense_%d/kernel" % i], w["dense_%d/bias" % i], w["dense_%d/prelu_alpha" % i]) for i in range(2)]
    want = O.din_model(inputs["d0"].astype(np.float64), inputs["o0"], np.stack([ha, hb], axis=2), cand,
                       [w["embed_sparse_0/embeddings"]], [w["embed_seq_0/embeddings"], w["embed_seq_1/embeddings"]], att,
                       (w["Attention_0/out/kernel"], w["Attention_0/out/bias"]),
                       (w["bn/gamma"], w["bn/beta"], w["bn/moving_mean"], w["bn/moving_variance"]), tw,
                       (w["out/kernel"], w["out/bias"]))
    tol = 5e-6 if prec == "fp32" else TOL / 4
    assert np.max(np.abs(got - want)) <= tol, np.max(np.abs(got - want))
