"""CPU: the oracle against the REFERENCE'S OWN outputs (tests/golden/ref_layers.npz: the reference's unmodified
layer / model files executed under tests/golden/tf_shim by make_ref_golden.py), against its frozen vectors, and against
algebraic identities: one-hot == gather, pair-sum identity, Cross with w=0, Attention with W=0 -> masked
mean, all-padding row -> uniform softmax."""
import numpy as np
import pytest

from oracle import ctr_oracle as O
from oracle import synth


@pytest.mark.parametrize("name", list(synth.CASES))
def test_oracle_matches_the_reference(golden_oracle, ref_golden, name):
    """THE PIN: float64 oracle == the reference's own call() output on the same inputs and weights, <= 1e-12."""
    arrays, _ = synth.load_case(golden_oracle, name)
    got = synth.CASES[name][1](arrays)
    want = ref_golden[name + "/expected"]
    assert got.shape == want.shape
    np.testing.assert_allclose(got, want, rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("name", list(synth.CASES))
def test_o32_matches_the_reference_in_float32(golden_oracle, ref_golden, name):
    """The oracle run in float32 against the reference run in float32 (its own precision): same op order, so the two agree
    to a few float32 roundings."""
    arrays, _ = synth.load_case(golden_oracle, name)
    a32 = {k: (v.astype(np.float32) if v.dtype == np.float64 else v) for k, v in arrays.items()}
    got = synth.CASES[name][1](a32)
    want = ref_golden[name + "/expected_f32"]
    np.testing.assert_allclose(got, want, rtol=2e-5, atol=2e-6)


@pytest.mark.parametrize("name,fn", [("fm", lambda a: O.fm_model(a["x"], a["w0"], a["w1"], a["v"])),
                                     ("ffm", lambda a: O.ffm_model(a["inputs"], [int(v) for v in a["vocab"]], a["w0"],
                                                                   a["w"], a["v"]))])
def test_sigmoid_model_wrappers_match_the_reference(golden_oracle, ref_golden, name, fn):
    """R11: FM.call (model/fm.py:19-23) and FFM.call (model/ffm.py:20-23)."""
    arrays, _ = synth.load_case(golden_oracle, name)
    np.testing.assert_allclose(fn(arrays), ref_golden[name + "_model/expected"], rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("name", list(synth.CASES))
def test_oracle_matches_frozen_golden(golden_oracle, name):
    arrays, expected = synth.load_case(golden_oracle, name)
    got = synth.CASES[name][1](arrays)
    np.testing.assert_allclose(got, expected, rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("name", list(synth.CASES))
def test_o32_close_to_o64(golden_oracle, name):
    """O32 (fp32, reference op order) stays within the north_star tolerances of O64."""
    arrays, expected = synth.load_case(golden_oracle, name)
    a32 = {k: (v.astype(np.float32) if v.dtype == np.float64 else v) for k, v in arrays.items()}
    got = synth.CASES[name][1](a32)
    np.testing.assert_allclose(got, expected, rtol=2e-5, atol=2e-6)


def test_fm_onehot_equals_gather_on_reference_data(criteo_cfg1):
    """Config 1: FM on the reference loader's dense one-hot matrix == 39-row gather form."""
    z = criteo_cfg1
    dense, ids = z["dense"][:256].astype(np.float64), z["ids"][:256]
    n = int(z["onehot_cols"])
    rng = np.random.default_rng(0)
    w0 = np.zeros(1)
    w1 = rng.normal(0, 0.05, (n, 1))
    v = rng.normal(0, 0.05, (n, 8))
    X = np.zeros((dense.shape[0], n))
    X[:, :13] = dense
    for f in range(26):
        X[np.arange(X.shape[0]), z["onehot_offsets"][f] + ids[:, f]] = 1.0
    assert abs((X != 0).sum(1).mean() - 35) < 1.0
    ref = O.fm_layer(X, w0, w1, v)
    got = O.fm_onehot_sparse(dense, np.arange(13), ids, z["onehot_offsets"], w0, w1, v)
    np.testing.assert_allclose(got, ref, rtol=1e-10, atol=1e-12)


def test_ffm_sparse_identity(golden):
    a, expected = synth.load_case(golden, "ffm")
    got = O.ffm_sparse(a["inputs"], [int(x) for x in a["vocab"]], a["w0"], a["w"], a["v"])
    np.testing.assert_allclose(got, expected, rtol=1e-10, atol=1e-12)


def test_pair_sum_identity():
    rng = np.random.default_rng(1)
    e = rng.normal(size=(5, 26, 8))
    inner = O.inner_product_layer(e).sum(axis=1)
    tot = e.sum(axis=1)
    np.testing.assert_allclose(inner, 0.5 * ((tot ** 2).sum(1) - (e ** 2).sum((1, 2))), rtol=1e-10)


def test_cross_zero_w():
    rng = np.random.default_rng(2)
    x = rng.normal(size=(4, 20))
    bs = [rng.normal(size=(20, 1)) for _ in range(3)]
    out = O.cross_layer(x, [np.zeros((20, 1))] * 3, bs)
    np.testing.assert_allclose(out, x + sum(b[:, 0] for b in bs), rtol=1e-12)


def test_attention_zero_weights_is_masked_mean():
    rng = np.random.default_rng(3)
    B, T, K = 4, 6, 8
    q, key = rng.normal(size=(B, K)), rng.normal(size=(B, T, K))
    mask = (np.arange(T)[None, :] < np.array([0, 6, 3, 1])[:, None]).astype(float)
    dense = [(np.zeros((4 * K, 5)), np.zeros(5), np.full((T, 5), 0.1))]
    out = O.din_attention(q, key, key, mask, dense, (np.zeros((5, 1)), np.zeros(1)))
    for b, n in enumerate([0, 6, 3, 1]):
        want = key[b].mean(0) if n == 0 else key[b, :n].mean(0)   # all-padding row -> uniform 1/T (interaction.py:399-404)
        np.testing.assert_allclose(out[b], want, rtol=1e-9, atol=1e-12)


def test_outer_product_closed_form():
    rng = np.random.default_rng(4)
    e = rng.normal(size=(3, 5, 4))
    W = rng.normal(size=(4, 10, 4))
    out = O.outer_product_layer(e, W)
    row, col = O.pair_index(5)
    want = np.einsum("bpa,apc,bpc->bp", e[:, col], W, e[:, row])
    np.testing.assert_allclose(out, want, rtol=1e-10)


def test_embed_out_of_range_raises():
    with pytest.raises(IndexError):
        O.embed_layer(np.array([[3.0]]), [np.zeros((3, 2))])


def test_embed_float_ids_truncate():
    t = np.arange(10.0).reshape(5, 2)
    out = O.embed_layer(np.array([[2.9]]), [t])
    np.testing.assert_array_equal(out, t[2:3])


def test_bi_interaction_is_the_sum_of_pairwise_products():
    """NFM's pooling (nfm.py:28) equals sum_{i<j} e_i * e_j elementwise -- the identity the kernel's sum / sum-of-squares
    form relies on."""
    rng = np.random.default_rng(0)
    e = rng.normal(size=(5, 7, 4))
    want = np.zeros((5, 4))
    for i in range(7):
        for j in range(i + 1, 7):
            want += e[:, i] * e[:, j]
    np.testing.assert_allclose(O.bi_interaction(e), want, rtol=1e-12, atol=1e-12)


def test_deepfm_train_step_gradients():
    """the oracle's analytic SGD step against central finite differences of its own loss (the forward
    half of that loss is the pinned deepfm_model).  Duplicate ids, dropout masks and both regularisers are in play."""
    rng = np.random.default_rng(0)
    B, D, F, k, kf = 9, 3, 4, 4, 5
    vocab = [7, 5, 11, 3]
    tables = [rng.normal(0, 0.3, (v, k)) for v in vocab]
    ids = np.stack([rng.integers(0, v, B) for v in vocab], 1)
    ids[1] = ids[0]
    dense, y = rng.random((B, D)), rng.integers(0, 2, B).astype(float)
    n = D + F * k
    w0, w1, v = rng.normal(0, 0.1, 1), rng.normal(0, 0.1, (n, 1)), rng.normal(0, 0.1, (n, kf))
    hidden = [(rng.normal(0, 0.3, (n, 8)), rng.normal(0, 0.1, 8)), (rng.normal(0, 0.3, (8, 6)), rng.normal(0, 0.1, 6))]
    out = (rng.normal(0, 0.3, (6, 1)), rng.normal(0, 0.1, 1))
    masks = [(rng.random((B, 8)) >= 0.2) / 0.8, (rng.random((B, 6)) >= 0.2) / 0.8]
    lr, rw, rv = 0.1, 1e-2, 2e-2
    new, bce, correct = O.deepfm_train_step(dense, ids, tables, w0, w1, v, hidden, out, y, lr, rw, rv, masks)
    base = dict(tables=tables, w0=w0, w1=w1, v=v, hidden=hidden, out=out)

    def loss(**kw):
        a = dict(base, **kw)
        return O.deepfm_loss(dense, ids, a["tables"], a["w0"], a["w1"], a["v"], a["hidden"], a["out"], y, rw, rv, masks)[0]

    def fd(arr, put, idx, eps=1e-6):
        a = arr.copy(); a[idx] += eps; lp = loss(**put(a)); a[idx] -= 2 * eps; lm = loss(**put(a))
        return (lp - lm) / (2 * eps)

    assert abs(loss() - (bce + rw * np.sum(w1 ** 2) + rv * np.sum(v ** 2))) < 1e-12 and 0 <= correct <= B
    checks = [(v, new[3], lambda a: {"v": a}, [(0, 0), (5, 2), (18, 4)]),
              (w1, new[2], lambda a: {"w1": a}, [(0, 0), (4, 0), (18, 0)]),
              (w0, new[1], lambda a: {"w0": a}, [(0,)]),
              (hidden[0][0], new[4][0][0], lambda a: {"hidden": [(a, hidden[0][1]), hidden[1]]}, [(0, 0), (7, 3), (18, 7)]),
              (hidden[1][1], new[4][1][1], lambda a: {"hidden": [hidden[0], (hidden[1][0], a)]}, [(j,) for j in range(6)]),
              (out[0], new[5][0], lambda a: {"out": (a, out[1])}, [(j, 0) for j in range(6)]),
              (out[1], new[5][1], lambda a: {"out": (out[0], a)}, [(0,)]),
              (tables[0], new[0][0], lambda a: {"tables": [a] + tables[1:]}, [(i, j) for i in range(7) for j in range(4)]),
              (tables[2], new[0][2], lambda a: {"tables": tables[:2] + [a] + tables[3:]}, [(i, 1) for i in range(11)])]
    for arr, upd, put, idxs in checks:
        g = (arr - upd) / lr
        for idx in idxs:
            assert abs(fd(arr, put, idx) - g[idx]) < 1e-8, (idx, fd(arr, put, idx), g[idx])


def test_fm_train_step_gradients():
    rng = np.random.default_rng(1)
    B, n, k = 11, 9, 4
    x, y = rng.normal(0, 1, (B, n)), rng.integers(0, 2, B).astype(float)
    w0, w1, v = rng.normal(0, 0.1, 1), rng.normal(0, 0.2, (n, 1)), rng.normal(0, 0.2, (n, k))
    lr, rw, rv = 0.1, 1e-2, 2e-2
    new, bce, _ = O.fm_train_step(x, w0, w1, v, y, lr, rw, rv)

    def loss(w0_, w1_, v_):
        p = np.clip(O.sigmoid(O.fm_layer(x, w0_, w1_, v_)[:, 0]), 1e-7, 1 - 1e-7)
        return np.mean(-(y * np.log(p) + (1 - y) * np.log(1 - p))) + rw * np.sum(w1_ ** 2) + rv * np.sum(v_ ** 2)
    eps = 1e-6
    for idx in [(0, 0), (4, 2), (8, 3)]:
        a, b = v.copy(), v.copy(); a[idx] += eps; b[idx] -= eps
        assert abs((loss(w0, w1, a) - loss(w0, w1, b)) / (2 * eps) - ((v - new[2]) / lr)[idx]) < 1e-8
    for i in (0, 5):
        a, b = w1.copy(), w1.copy(); a[i, 0] += eps; b[i, 0] -= eps
        assert abs((loss(w0, a, v) - loss(w0, b, v)) / (2 * eps) - ((w1 - new[1]) / lr)[i, 0]) < 1e-8
    assert abs((loss(w0 + eps, w1, v) - loss(w0 - eps, w1, v)) / (2 * eps) - ((w0 - new[0]) / lr)[0]) < 1e-8


def test_farm_fingerprint_pins():
    """TensorFlow's documented example of to_hash_bucket_fast and FarmHash's value for the empty string."""
    assert O.hash_layer([b"Hello", b"TensorFlow", b"2.x"], 3).tolist() == [0, 2, 2]
    assert O.farm_fingerprint64(b"") == 0x9ae16a3b2f90404f
    got = O.hash_layer(np.array([[0, 5], [123456789012, 7]]), 10, True)
    assert got.shape == (2, 2) and got[0, 0] == 0 and np.all(got[got != 0] >= 1) and np.all(got <= 9)


def test_dcn_train_step_gradients():
    """oracle.dcn_train_step's analytic update against central finite differences of oracle.dcn_loss (duplicate ids, dropout
    masks, both regularisers, 3 cross layers)."""
    rng = np.random.default_rng(7)
    B, D, F, k, L = 9, 3, 4, 4, 3
    vocab = [5, 7, 3, 6]
    tables = [rng.normal(0, 0.3, (v, k)) for v in vocab]
    ids = np.stack([rng.integers(0, v, B) for v in vocab], axis=1)
    ids[1] = ids[0]
    dense = rng.random((B, D))
    n = D + F * k
    ws = [rng.normal(0, 0.2, (n, 1)) for _ in range(L)]
    bs = [rng.normal(0, 0.2, (n, 1)) for _ in range(L)]
    hidden = [(rng.normal(0, 0.3, (n, 8)), rng.normal(0, 0.1, 8)), (rng.normal(0, 0.3, (8, 5)), rng.normal(0, 0.1, 5))]
    out = (rng.normal(0, 0.3, (5, 1)), rng.normal(0, 0.1, 1))
    final = (rng.normal(0, 0.3, (n + 1, 1)), rng.normal(0, 0.1, 1))
    y = rng.integers(0, 2, B).astype(float)
    masks = [(rng.random((B, 8)) >= 0.25) / 0.75, (rng.random((B, 5)) >= 0.25) / 0.75]
    lr, rw, rb = 0.05, 1e-2, 2e-2
    new, bce, _ = O.dcn_train_step(dense, ids, tables, ws, bs, hidden, out, final, y, lr, rw, rb, masks)
    base = O.dcn_loss(dense, ids, tables, ws, bs, hidden, out, final, y, rw, rb, masks)
    assert abs(base[1] - bce) < 1e-12
    eps = 1e-6

    def fd(obj, idx, rebuild):
        a, b = obj.copy(), obj.copy()
        a[idx] += eps; b[idx] -= eps
        return (O.dcn_loss(*rebuild(a), y, rw, rb, masks)[0] - O.dcn_loss(*rebuild(b), y, rw, rb, masks)[0]) / (2 * eps)

    def args(tables_=tables, ws_=ws, bs_=bs, hidden_=hidden, out_=out, final_=final):
        return dense, ids, tables_, ws_, bs_, hidden_, out_, final_

    for l in range(L):
        for idx in [(0, 0), (n - 1, 0), (5, 0)]:
            gw = fd(ws[l], idx, lambda a: args(ws_=ws[:l] + [a] + ws[l + 1:]))
            assert abs(gw - ((ws[l] - new[1][l]) / lr)[idx]) < 1e-7
            gb = fd(bs[l], idx, lambda a: args(bs_=bs[:l] + [a] + bs[l + 1:]))
            assert abs(gb - ((bs[l] - new[2][l]) / lr)[idx]) < 1e-7
    for idx in [(0, 0), (n, 0), (7, 0)]:
        gf = fd(final[0], idx, lambda a: args(final_=(a, final[1])))
        assert abs(gf - ((final[0] - new[5][0]) / lr)[idx]) < 1e-7
    gk = fd(hidden[0][0], (4, 3), lambda a: args(hidden_=[(a, hidden[0][1]), hidden[1]]))
    assert abs(gk - ((hidden[0][0] - new[3][0][0]) / lr)[4, 3]) < 1e-7
    go = fd(out[0], (2, 0), lambda a: args(out_=(a, out[1])))
    assert abs(go - ((out[0] - new[4][0]) / lr)[2, 0]) < 1e-7
    for f, r in [(0, int(ids[0, 0])), (2, int(ids[3, 2]))]:
        gt = fd(tables[f], (r, 1), lambda a: args(tables_=tables[:f] + [a] + tables[f + 1:]))
        assert abs(gt - ((tables[f] - new[0][f]) / lr)[r, 1]) < 1e-7


@pytest.mark.parametrize("mode", ["both", "inner", "outer"])
def test_pnn_train_step_gradients(mode):
    """oracle.pnn_train_step against central finite differences of oracle.pnn_loss: duplicate ids, dropout masks, raw outputs on
    both sides of keras' clip (inside (0, 1): gradient, outside: none)."""
    rng = np.random.default_rng(11)
    B, F, K = 10, 5, 4
    vocab = [6, 4, 7, 5, 3]
    tables = [rng.normal(0, 0.5, (v, K)) for v in vocab]
    ids = np.stack([rng.integers(0, v, B) for v in vocab], axis=1)
    ids[1] = ids[0]
    P = F * (F - 1) // 2
    W = rng.normal(0, 0.3, (K, P, K))
    n = F * K + P * (2 if mode == "both" else 1)
    hidden = [(rng.normal(0, 0.3, (n, 7)), rng.normal(0, 0.1, 7)), (rng.normal(0, 0.4, (7, 5)), rng.normal(0, 0.1, 5))]
    out = (rng.normal(0, 0.5, (5, 1)), np.array([0.45]))
    y = rng.integers(0, 2, B).astype(float)
    masks = [(rng.random((B, 7)) >= 0.25) / 0.75, (rng.random((B, 5)) >= 0.25) / 0.75]
    lr = 0.05
    new, loss, _ = O.pnn_train_step(ids, tables, W, hidden, out, y, lr, mode, masks)
    base, pre = O.pnn_loss(ids, tables, W, hidden, out, y, mode, masks)
    assert abs(base - loss) < 1e-12
    assert np.any((pre > 0.01) & (pre < 0.99)) and np.all(np.abs(pre - 1e-7) > 1e-3) and np.all(np.abs(pre - 1 + 1e-7) > 1e-3)
    eps = 1e-6

    def fd(obj, idx, rebuild):
        a, b = obj.copy(), obj.copy()
        a[idx] += eps; b[idx] -= eps
        return (O.pnn_loss(*rebuild(a), y, mode, masks)[0] - O.pnn_loss(*rebuild(b), y, mode, masks)[0]) / (2 * eps)

    for f, r, c in [(0, int(ids[0, 0]), 1), (3, int(ids[4, 3]), 2), (4, int(ids[2, 4]), 0)]:
        gt = fd(tables[f], (r, c), lambda a: (ids, tables[:f] + [a] + tables[f + 1:], W, hidden, out))
        assert abs(gt - ((tables[f] - new[0][f]) / lr)[r, c]) < 1e-6
    if mode != "inner":
        for idx in [(0, 0, 0), (3, P - 1, 2), (1, 4, 3)]:
            gw = fd(W, idx, lambda a: (ids, tables, a, hidden, out))
            assert abs(gw - ((W - new[1]) / lr)[idx]) < 1e-6
    gk = fd(hidden[0][0], (n - 1, 3), lambda a: (ids, tables, W, [(a, hidden[0][1]), hidden[1]], out))
    assert abs(gk - ((hidden[0][0] - new[2][0][0]) / lr)[n - 1, 3]) < 1e-6
    go = fd(out[0], (2, 0), lambda a: (ids, tables, W, hidden, (a, out[1])))
    assert abs(go - ((out[0] - new[3][0]) / lr)[2, 0]) < 1e-6


def test_widedeep_train_step_gradients():
    rng = np.random.default_rng(13)
    B, D, F, K = 9, 3, 4, 4
    vocab, widths = [6, 4, 7, 5], [6, 3, 7, 5]                    # field 1: ids 3 (not seen by get_dummies) have no wide column
    tables = [rng.normal(0, 0.5, (v, K)) for v in vocab]
    ids = np.stack([rng.integers(0, v, B) for v in vocab], axis=1)
    ids[1] = ids[0]
    ids[2, 1] = 3
    dense = rng.random((B, D))
    nw = 2 * D + sum(widths)
    w0, w = rng.normal(0, 0.1, 1), rng.normal(0, 0.3, (nw, 1))
    hidden = [(rng.normal(0, 0.3, (F * K, 7)), rng.normal(0, 0.1, 7)), (rng.normal(0, 0.4, (7, 5)), rng.normal(0, 0.1, 5))]
    out = (rng.normal(0, 0.5, (5, 1)), rng.normal(0, 0.1, 1))
    y = rng.integers(0, 2, B).astype(float)
    masks = [(rng.random((B, 7)) >= 0.25) / 0.75, (rng.random((B, 5)) >= 0.25) / 0.75]
    lr, reg = 0.05, 1e-2
    new, bce, _ = O.widedeep_train_step(dense, ids, tables, w0, w, widths, hidden, out, y, lr, reg, masks)
    eps = 1e-6

    def fd(obj, idx, rebuild):
        a, b = obj.copy(), obj.copy()
        a[idx] += eps; b[idx] -= eps
        return (O.widedeep_loss(*rebuild(a), y, reg, masks)[0] - O.widedeep_loss(*rebuild(b), y, reg, masks)[0]) / (2 * eps)

    for idx in [(0, 0), (D + 1, 0), (2 * D + int(ids[0, 0]), 0), (nw - 1, 0)]:
        gw = fd(w, idx, lambda a: (dense, ids, tables, w0, a, widths, hidden, out))
        assert abs(gw - ((w - new[2]) / lr)[idx]) < 1e-7
    g0 = fd(w0, (0,), lambda a: (dense, ids, tables, a, w, widths, hidden, out))
    assert abs(g0 - ((w0 - new[1]) / lr)[0]) < 1e-7
    for f, r, c in [(0, int(ids[0, 0]), 1), (3, int(ids[4, 3]), 2)]:
        gt = fd(tables[f], (r, c), lambda a: (dense, ids, tables[:f] + [a] + tables[f + 1:], w0, w, widths, hidden, out))
        assert abs(gt - ((tables[f] - new[0][f]) / lr)[r, c]) < 1e-7
    gk = fd(hidden[0][0], (5, 3), lambda a: (dense, ids, tables, w0, w, widths, [(a, hidden[0][1]), hidden[1]], out))
    assert abs(gk - ((hidden[0][0] - new[3][0][0]) / lr)[5, 3]) < 1e-7


def test_deepcrossing_train_step_gradients():
    rng = np.random.default_rng(17)
    B, D, F, K, R = 8, 3, 4, 4, 3
    vocab = [6, 4, 7, 5]
    tables = [rng.normal(0, 0.5, (v, K)) for v in vocab]
    ids = np.stack([rng.integers(0, v, B) for v in vocab], axis=1)
    ids[1] = ids[0]
    dense = rng.random((B, D))
    n = D + F * K
    res = [([(rng.normal(0, 0.3, (n, 6)), rng.normal(0, 0.1, 6)), (rng.normal(0, 0.3, (6, 5)), rng.normal(0, 0.1, 5))],
            (rng.normal(0, 0.3, (5, n)), rng.normal(0, 0.1, n))) for _ in range(R)]
    final = (rng.normal(0, 0.3, (n, 1)), rng.normal(0, 0.1, 1))
    y = rng.integers(0, 2, B).astype(float)
    lr = 0.05
    new, bce, _ = O.deepcrossing_train_step(dense, ids, tables, res, final, y, lr)

    def loss(tables_, res_, final_):
        X = np.concatenate([dense, ids.astype(float)], axis=1)
        p = np.clip(O.deepcrossing_model(X, tables_, res_, final_, n_dense=D)[:, 0], 1e-7, 1 - 1e-7)
        return np.mean(-(y * np.log(p) + (1 - y) * np.log(1 - p)))
    assert abs(loss(tables, res, final) - bce) < 1e-12
    eps = 1e-6

    def fd(obj, idx, rebuild):
        a, b = obj.copy(), obj.copy()
        a[idx] += eps; b[idx] -= eps
        return (loss(*rebuild(a)) - loss(*rebuild(b))) / (2 * eps)

    def with_res(r, hidden=None, out=None):
        h, o = res[r]
        return res[:r] + [(hidden if hidden is not None else h, out if out is not None else o)] + res[r + 1:]
    for r in range(R):
        h, o = res[r]
        gk = fd(h[0][0], (2, 3), lambda a: (tables, with_res(r, hidden=[(a, h[0][1]), h[1]]), final))
        assert abs(gk - ((h[0][0] - new[1][r][0][0][0]) / lr)[2, 3]) < 1e-7
        gb = fd(h[1][1], (4,), lambda a: (tables, with_res(r, hidden=[h[0], (h[1][0], a)]), final))
        assert abs(gb - ((h[1][1] - new[1][r][0][1][1]) / lr)[4]) < 1e-7
        go = fd(o[0], (1, n - 1), lambda a: (tables, with_res(r, out=(a, o[1])), final))
        assert abs(go - ((o[0] - new[1][r][1][0]) / lr)[1, n - 1]) < 1e-7
    gf = fd(final[0], (3, 0), lambda a: (tables, res, (a, final[1])))
    assert abs(gf - ((final[0] - new[2][0]) / lr)[3, 0]) < 1e-7
    for f, rr, c in [(0, int(ids[0, 0]), 1), (3, int(ids[4, 3]), 2)]:
        gt = fd(tables[f], (rr, c), lambda a: (tables[:f] + [a] + tables[f + 1:], res, final))
        assert abs(gt - ((tables[f] - new[0][f]) / lr)[rr, c]) < 1e-7


def test_nfm_train_step_gradients():
    """Finite differences go through the batch statistics of the training-mode BatchNormalization."""
    rng = np.random.default_rng(19)
    B, D, F, K = 9, 3, 4, 4
    vocab = [6, 4, 7, 5]
    tables = [rng.normal(0, 0.5, (v, K)) for v in vocab]
    ids = np.stack([rng.integers(0, v, B) for v in vocab], axis=1)
    ids[1] = ids[0]
    dense = rng.random((B, D))
    n = D + K
    bn = (rng.normal(0, 0.1, n), rng.uniform(0.5, 1.5, n), rng.uniform(0.5, 1.5, n), rng.normal(0, 0.1, n))
    hidden = [(rng.normal(0, 0.4, (n, 7)), rng.normal(0, 0.1, 7)), (rng.normal(0, 0.4, (7, 5)), rng.normal(0, 0.1, 5))]
    out = (rng.normal(0, 0.5, (5, 2)), rng.normal(0, 0.1, 2))
    final = (rng.normal(0, 0.5, (2, 1)), rng.normal(0, 0.1, 1))
    y = rng.integers(0, 2, B).astype(float)
    masks = [(rng.random((B, 7)) >= 0.25) / 0.75, (rng.random((B, 5)) >= 0.25) / 0.75]
    lr = 0.05
    new, bce, _ = O.nfm_train_step(dense, ids, tables, bn, hidden, out, final, y, lr, masks)

    def loss(tables_, bn_, hidden_, out_, final_):
        p = np.clip(O.nfm_train_forward(dense, ids, tables_, bn_, hidden_, out_, final_, masks)[0], 1e-7, 1 - 1e-7)
        return np.mean(-(y * np.log(p) + (1 - y) * np.log(1 - p)))
    assert abs(loss(tables, bn, hidden, out, final) - bce) < 1e-12
    eps = 1e-6

    def fd(obj, idx, rebuild):
        a, b = obj.copy(), obj.copy()
        a[idx] += eps; b[idx] -= eps
        return (loss(*rebuild(a)) - loss(*rebuild(b))) / (2 * eps)
    for f, r, c in [(0, int(ids[0, 0]), 1), (3, int(ids[4, 3]), 2), (2, int(ids[7, 2]), 0)]:
        gt = fd(tables[f], (r, c), lambda a: (tables[:f] + [a] + tables[f + 1:], bn, hidden, out, final))
        assert abs(gt - ((tables[f] - new[0][f]) / lr)[r, c]) < 1e-7
    for j in (0, D, n - 1):
        gg = fd(bn[2], (j,), lambda a: (tables, (bn[0], bn[1], a, bn[3]), hidden, out, final))
        assert abs(gg - ((bn[2] - new[1][2]) / lr)[j]) < 1e-7
        gb = fd(bn[3], (j,), lambda a: (tables, (bn[0], bn[1], bn[2], a), hidden, out, final))
        assert abs(gb - ((bn[3] - new[1][3]) / lr)[j]) < 1e-7
    gk = fd(hidden[0][0], (2, 3), lambda a: (tables, bn, [(a, hidden[0][1]), hidden[1]], out, final))
    assert abs(gk - ((hidden[0][0] - new[2][0][0]) / lr)[2, 3]) < 1e-7
    go = fd(out[0], (3, 1), lambda a: (tables, bn, hidden, (a, out[1]), final))
    assert abs(go - ((out[0] - new[3][0]) / lr)[3, 1]) < 1e-7
    gf = fd(final[0], (1, 0), lambda a: (tables, bn, hidden, out, (a, final[1])))
    assert abs(gf - ((final[0] - new[4][0]) / lr)[1, 0]) < 1e-7
    # moving statistics: keras' momentum rule on the batch mean / biased variance of x = [dense | bi-interaction]
    x = O.nfm_train_forward(dense, ids, tables, bn, hidden, out, final, masks)[1][2]
    np.testing.assert_allclose(new[1][0], 0.99 * bn[0] + 0.01 * x.mean(0), rtol=0, atol=1e-15)
    np.testing.assert_allclose(new[1][1], 0.99 * bn[1] + 0.01 * x.var(0), rtol=0, atol=1e-15)


@pytest.mark.parametrize("mode,ns", [("att", 2), ("avg", 2), ("max", 2), ("att", 1)])
def test_afm_train_step_gradients(mode, ns):
    rng = np.random.default_rng(23)
    B, F, K = 9, 5, 4
    vocab = [6, 4, 7, 5, 3]
    tables = [rng.normal(0, 0.7, (v, K)) for v in vocab]
    ids = np.stack([rng.integers(0, v, B) for v in vocab], axis=1)
    ids[1] = ids[0]
    out = (rng.normal(0, 0.7, (K, 1)), rng.normal(0, 0.1, 1))
    y = rng.integers(0, 2, B).astype(float)
    lr = 0.05
    new, bce, _ = O.afm_train_step(ids, tables, mode, out, y, lr, ns)
    X = np.concatenate([np.zeros((B, 13)), ids.astype(float)], axis=1)

    def loss(tables_, out_):
        p = O.afm_layer(X, tables_, mode, att_w=(np.zeros((K, 10)), np.zeros(10)), att_h=(np.zeros((10, 1)), np.zeros(1)), out=out_)[:, 0]
        p = np.clip(O.sigmoid(p) if ns == 2 else p, 1e-7, 1 - 1e-7)
        return np.mean(-(y * np.log(p) + (1 - y) * np.log(1 - p)))
    assert abs(loss(tables, out) - bce) < 1e-12
    eps = 1e-6

    def fd(obj, idx, rebuild):
        a, b = obj.copy(), obj.copy()
        a[idx] += eps; b[idx] -= eps
        return (loss(*rebuild(a)) - loss(*rebuild(b))) / (2 * eps)
    for f, r, c in [(0, int(ids[0, 0]), 1), (3, int(ids[4, 3]), 2), (4, int(ids[2, 4]), 0)]:
        gt = fd(tables[f], (r, c), lambda a: (tables[:f] + [a] + tables[f + 1:], out))
        assert abs(gt - ((tables[f] - new[0][f]) / lr)[r, c]) < 1e-7
    go = fd(out[0], (2, 0), lambda a: (tables, (a, out[1])))
    assert abs(go - ((out[0] - new[1][0]) / lr)[2, 0]) < 1e-7


def test_ffm_train_step_gradients():
    rng = np.random.default_rng(29)
    B, D, K = 7, 3, 4
    vocab = [5, 3, 6, 4]
    field = D + len(vocab)
    n_feat = D + sum(vocab)
    ids = np.stack([rng.integers(0, v, B) for v in vocab], axis=1)
    ids[1] = ids[0]
    X = np.concatenate([rng.random((B, D)), ids.astype(float)], axis=1)
    w0, w, v = rng.normal(0, 0.1, 1), rng.normal(0, 0.3, (n_feat, 1)), rng.normal(0, 0.3, (n_feat, field, K))
    y = rng.integers(0, 2, B).astype(float)
    lr, rw, rv = 0.05, 1e-2, 3e-2
    new, bce, _ = O.ffm_train_step(X, vocab, w0, w, v, y, lr, rw, rv, n_dense=D)

    def loss(w0_, w_, v_):
        p = np.clip(O.sigmoid(O.ffm_layer(X, vocab, w0_, w_, v_, D)[:, 0]), O._BCE_LO, O._BCE_HI)
        return np.mean(-(y * np.log(p) + (1 - y) * np.log(1 - p))) + rw * np.sum(w_ ** 2) + rv * np.sum(v_ ** 2)
    assert abs(loss(w0, w, v) - rw * np.sum(w ** 2) - rv * np.sum(v ** 2) - bce) < 1e-12
    eps = 1e-6

    def fd(obj, idx, rebuild):
        a, b = obj.copy(), obj.copy()
        a[idx] += eps; b[idx] -= eps
        return (loss(*rebuild(a)) - loss(*rebuild(b))) / (2 * eps)
    hit = D + int(ids[0, 0])                                                           # a row two samples share
    for idx in [(0, 2, 1), (hit, 0, 3), (hit, 5, 0), (D + vocab[0] + int(ids[3, 1]), 6, 2), (n_feat - 1, 1, 1)]:
        g = fd(v, idx, lambda a: (w0, w, a))
        assert abs(g - ((v - new[2]) / lr)[idx]) < 1e-7
    for idx in [(1, 0), (hit, 0), (n_feat - 1, 0)]:
        g = fd(w, idx, lambda a: (w0, a, v))
        assert abs(g - ((w - new[1]) / lr)[idx]) < 1e-7
    g = fd(w0, (0,), lambda a: (a, w, v))
    assert abs(g - ((w0 - new[0]) / lr)[0]) < 1e-7


def _din_case(seed=31, B=6, T=5, K=4, all_masked_row=True):
    rng = np.random.default_rng(seed)
    v_seq, other_v, other_k, nd = 9, [5, 4], [3, 2], 2
    seq = rng.normal(0, 0.6, (v_seq, K))
    other_tables = [rng.normal(0, 0.6, (v, k)) for v, k in zip(other_v, other_k)]
    hist = rng.integers(0, v_seq, (B, T))
    hist[0, 3:] = 0
    if all_masked_row:
        hist[2] = 0
    cand = rng.integers(1, v_seq, B)
    cand[1] = hist[1, 0] if hist[1, 0] else 1                     # candidate row also in the history
    other_ids = np.stack([rng.integers(0, v, B) for v in other_v], axis=1)
    dense = rng.random((B, nd))
    dims = [4 * K, 6, 5]
    att = [(rng.normal(0, 0.5, (dims[i], dims[i + 1])), rng.normal(0, 0.1, dims[i + 1]), rng.normal(0.2, 0.2, (T, dims[i + 1])))
           for i in range(2)]
    att_out = (rng.normal(0, 0.5, (dims[-1], 1)), rng.normal(0, 0.1, 1))
    n = 2 * K + sum(other_k) + nd
    bn = (rng.normal(1, 0.2, n), rng.normal(0, 0.2, n), rng.normal(0, 0.1, n), rng.random(n) + 0.5)
    td = [n, 7, 5]
    tower = [(rng.normal(0, 0.5, (td[i], td[i + 1])), rng.normal(0, 0.1, td[i + 1]), rng.normal(0.2, 0.2, td[i + 1])) for i in range(2)]
    out = (rng.normal(0, 0.5, (td[-1], 1)), rng.normal(0, 0.1, 1))
    y = rng.integers(0, 2, B).astype(float)
    return dict(dense_inputs=dense, other_ids=other_ids, hist_ids=hist, cand_ids=cand, other_tables=other_tables, seq_table=seq,
                att_dense=att, att_out=att_out, bn=bn, tower=tower, out=out), y


def test_din_train_step_gradients():
    P, y = _din_case()
    lr = 0.05
    new, bce, _ = O.din_train_step(**P, y=y, lr=lr)
    assert abs(O.din_loss(O.din_train_forward(**P)[0], y) - bce) < 1e-12
    # the training-mode forward with the batch statistics put in as moving statistics equals the inference model
    _, c = O.din_train_forward(**P)
    inf = O.din_model(P["dense_inputs"], P["other_ids"], P["hist_ids"][:, :, None], P["cand_ids"][:, None], P["other_tables"],
                      [P["seq_table"]], P["att_dense"], P["att_out"], (P["bn"][0], P["bn"][1], c["mu"], c["var"]), P["tower"], P["out"])
    np.testing.assert_allclose(inf[:, 0], O.din_train_forward(**P)[0], rtol=0, atol=1e-12)
    eps = 1e-6

    def fd(get, put, idx):
        vals = []
        for sgn in (1, -1):
            Q = dict(P)
            a = get(P).copy()
            a[idx] += sgn * eps
            put(Q, a)
            vals.append(O.din_loss(O.din_train_forward(**Q)[0], y))
        return (vals[0] - vals[1]) / (2 * eps)

    def check(get, put, got_new, idxs, what):
        for idx in idxs:
            g = fd(get, put, idx)
            assert abs(g - ((get(P) - got_new) / lr)[idx]) < 2e-7, (what, idx, g, ((get(P) - got_new) / lr)[idx])
    hist, cand = P["hist_ids"], P["cand_ids"]
    check(lambda D: D["seq_table"], lambda D, a: D.__setitem__("seq_table", a), new["seq_table"],
          [(int(hist[1, 0]), 1), (int(cand[3]), 2), (0, 0), (int(hist[4, 2]), 3)], "seq")
    for i in range(2):
        check(lambda D: D["other_tables"][i], lambda D, a: D.__setitem__("other_tables", [a if j == i else t for j, t in enumerate(P["other_tables"])]),
              new["other_tables"][i], [(int(P["other_ids"][0, i]), 1)], "other")
    for l in range(2):
        for part, idxs in ((0, [(1, 2), (3, 0)]), (1, [(2,)]), (2, [(1, 3), (4, 0)])):
            check(lambda D: D["att_dense"][l][part],
                  lambda D, a: D.__setitem__("att_dense", [tuple(a if (j == l and q_ == part) else t[q_] for q_ in range(3)) for j, t in enumerate(P["att_dense"])]),
                  new["att_dense"][l][part], idxs, "att %d %d" % (l, part))
        for part, idxs in ((0, [(1, 2), (3, 0)]), (1, [(2,)]), (2, [(1,), (4,)])):
            check(lambda D: D["tower"][l][part],
                  lambda D, a: D.__setitem__("tower", [tuple(a if (j == l and q_ == part) else t[q_] for q_ in range(3)) for j, t in enumerate(P["tower"])]),
                  new["tower"][l][part], idxs, "tower %d %d" % (l, part))
    for part, idxs in ((0, [(2, 0)]), (1, [(0,)])):
        check(lambda D: D["att_out"][part], lambda D, a: D.__setitem__("att_out", tuple(a if q_ == part else P["att_out"][q_] for q_ in range(2))),
              new["att_out"][part], idxs, "att_out")
        check(lambda D: D["out"][part], lambda D, a: D.__setitem__("out", tuple(a if q_ == part else P["out"][q_] for q_ in range(2))),
              new["out"][part], idxs, "out")
    for part in (0, 1):
        check(lambda D: D["bn"][part], lambda D, a: D.__setitem__("bn", tuple(a if q_ == part else P["bn"][q_] for q_ in range(4))),
              new["bn"][part], [(0,), (5,), (9,)], "bn")
    np.testing.assert_allclose(new["bn"][2], 0.99 * P["bn"][2] + 0.01 * c["mu"], rtol=0, atol=1e-15)
    np.testing.assert_allclose(new["bn"][3], 0.99 * P["bn"][3] + 0.01 * c["var"], rtol=0, atol=1e-15)
