"""GPU parity for SURVEY 8(f) rank 3: one SGD step of DeepFM (ctrhot_deepfm_train_step_f32, csrc/train.cu) and the
compile_fit drop-in (DL/utils/compile_fit.py:9-15) against the float64 oracle `deepfm_train_step`, whose gradients are
pinned by finite differences (tests/test_oracle.py).  PARITY UNPINNED against the reference itself: TensorFlow's autodiff
and optimizer cannot run here and the reference ships no training fixtures.

Checked: every updated weight after one step (fp32 arithmetic: 1e-5 relative to the size of the weights' update), rows that
were not looked up bit-unchanged, duplicate ids summed, the reported loss / accuracy, dropout with the kernel's own stateless
masks reproduced in NumPy, ragged last batches and a decreasing loss over epochs through compile_fit.
"""
import numpy as np
import pytest
import torch

from oracle import ctr_oracle as O
from recommender_system_b200.layer import set_seed
from recommender_system_b200.model.criteo import DeepFM
from recommender_system_b200.utils.compile_fit import _DeepFMTrainer, compile_fit

pytestmark = pytest.mark.gpu


def npy(t):
    return t.detach().cpu().numpy().astype(np.float64)


def make_model(vocab, k_embed, k_fm, hidden, dropout=0.0, seed=1):
    set_seed(seed)
    fc = [[{"feat": "I%d" % i} for i in range(13)],
          [{"feat": "C%d" % i, "feat_onehot_dim": int(v), "embed_dim": k_embed} for i, v in enumerate(vocab)]]
    m = DeepFM(fc, k_fm, 1e-2, 2e-2, list(hidden), 1, "relu", embed_dim=k_embed, precision="fp32")
    m.build(None)
    m.dnn.dropout = dropout
    rng = np.random.default_rng(seed)
    upd = {}
    for n, w in m.named_weights().items():                       # keras' zero biases / w0 would hide mistakes in their gradients
        if n.endswith("bias") or n.endswith("w0"):
            upd[n] = rng.normal(0, 0.1, tuple(w.shape))
        elif n.endswith("/v") or n.endswith("/w1"):
            upd[n] = rng.normal(0, 0.1, tuple(w.shape))
    m.set_weights({n: np.asarray(v, np.float32) for n, v in upd.items()})
    return m


def params(m):
    E = m.embed_layer
    tables = [npy(t) for t in E.emb_layers]
    hidden = [(npy(k), npy(b)) for k, b in m.dnn.hidden]
    return tables, npy(m.fm.w0), npy(m.fm.w1), npy(m.fm.v), hidden, (npy(m.dnn.output[0]), npy(m.dnn.output[1]))


def batch(vocab, B, seed, dup=True):
    rng = np.random.default_rng(seed)
    ids = np.stack([rng.integers(0, v, B) for v in vocab], axis=1).astype(np.int32)
    if dup and B > 3:
        ids[1], ids[2, :5] = ids[0], ids[0, :5]                  # whole-row and partial duplicates
    dense = rng.random((B, 13)).astype(np.float32)
    y = rng.integers(0, 2, B).astype(np.float32)
    return dense, ids, y


def kernel_masks(seed, hidden, B, rate):
    """The stateless dropout masks of csrc/train.cu (splitmix64 of (seed, layer, element)) in NumPy."""
    out = []
    with np.errstate(over="ignore"):
        for l, h in enumerate(hidden):
            idx = np.arange(B * h, dtype=np.uint64)
            z = np.uint64(seed) + np.uint64(0x9E3779B97F4A7C15) * (idx + np.uint64(1)) + (np.uint64(l) << np.uint64(56))
            z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
            z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
            z = z ^ (z >> np.uint64(31))
            u = (z >> np.uint64(40)).astype(np.float32) * np.float32(1.0 / 16777216.0)
            out.append(((u >= np.float32(rate)).astype(np.float64) / (1.0 - rate)).reshape(B, h))
    return out


def close(got, want, before, what, floor=2e-7):
    scale = max(np.max(np.abs(want - before)), 1e-12)            # size of the update itself
    err = np.max(np.abs(got - want))
    # floor: fp32 round-off of the weights themselves (an arena row hit by n samples takes n atomic adds, each rounded at the
    # row's own magnitude -- 1e-6 for the PNN cases, whose rows are O(0.5) and whose smallest table has 7 rows for 257 samples)
    assert err <= 2e-5 * scale + floor * max(np.max(np.abs(want)), 1.0), (what, err, scale)


@pytest.mark.parametrize("B,vocab,k_embed,k_fm,hidden,rate", [
    (32, [50, 7, 1000, 3] + [20] * 22, 8, 10, (256, 128, 64), 0.0),      # the reference's compile_fit batch and tower
    (1, [11] * 26, 4, 4, (16,), 0.0),
    (257, [50, 7, 1000, 3] + [20] * 22, 16, 8, (64, 32), 0.2),
    (4096, [5000] * 26, 8, 10, (256, 128, 64), 0.2),
    (96, [50, 7, 1000, 3] + [20] * 22, 8, 10, (256, 128, 64), 0.2),      # three 32-row blocks of the small-batch product
    (128, [5000] * 26, 8, 10, (256, 128, 64), 0.0),
])
def test_one_step_against_the_oracle(lib, B, vocab, k_embed, k_fm, hidden, rate):
    m = make_model(vocab, k_embed, k_fm, hidden, rate)
    dense, ids, y = batch(vocab, B, seed=B)
    before = params(m)
    lr, seed = 0.05, 12345
    tr = _DeepFMTrainer(m)
    dev = m.device
    out = tr.train_on_batch(torch.from_numpy(dense).to(dev), torch.from_numpy(ids).to(dev), torch.from_numpy(y).to(dev), lr, seed=seed)
    torch.cuda.synchronize()
    tr.done()
    masks = kernel_masks(seed, hidden, B, rate) if rate > 0 else None
    want, bce, correct = O.deepfm_train_step(dense.astype(np.float64), ids, before[0], before[1], before[2], before[3], before[4],
                                             before[5], y.astype(np.float64), lr, 1e-2, 2e-2, masks)
    got = params(m)
    assert abs(float(out[0]) - bce) <= 2e-6 * max(1.0, bce) and int(out[1]) == correct
    for f in range(len(vocab)):
        close(got[0][f], want[0][f], before[0][f], "table %d" % f)
        untouched = np.setdiff1d(np.arange(vocab[f]), ids[:, f])
        np.testing.assert_array_equal(got[0][f][untouched], before[0][f][untouched])      # rows not looked up do not move
    close(got[1], want[1], before[1], "w0"); close(got[2], want[2], before[2], "w1"); close(got[3], want[3], before[3], "v")
    for l in range(len(hidden)):
        close(got[4][l][0], want[4][l][0], before[4][l][0], "kernel %d" % l)
        close(got[4][l][1], want[4][l][1], before[4][l][1], "bias %d" % l)
    close(got[5][0], want[5][0], before[5][0], "out kernel"); close(got[5][1], want[5][1], before[5][1], "out bias")


def test_float_ids_and_bad_ids(lib):
    vocab = [9] * 26
    m1, m2 = make_model(vocab, 4, 4, (8,)), make_model(vocab, 4, 4, (8,))
    dense, ids, y = batch(vocab, 40, seed=3)
    dev = m1.device
    for m, conv in ((m1, lambda a: a), (m2, lambda a: a.astype(np.float32))):
        tr = _DeepFMTrainer(m)
        tr.train_on_batch(torch.from_numpy(dense).to(dev), torch.from_numpy(conv(ids)).to(dev), torch.from_numpy(y).to(dev), 0.1)
    torch.cuda.synchronize()
    for a, b in zip(params(m1)[0], params(m2)[0]):
        np.testing.assert_allclose(a, b, rtol=0, atol=1e-7)      # (atomic adds: the order of duplicate rows may differ)
    ids[3, 5] = 9                                                # one past the table: raises like the forward path
    with pytest.raises(IndexError):
        compile_fit(m1, np.concatenate([dense, ids.astype(np.float32)], axis=1), y, batch_size=16, epochs=1, verbose=False)


def test_compile_fit_learns_and_predicts(lib):
    """compile_fit on data with a planted rule (the label is decided by which half of its vocabulary one id falls in): keras'
    printed loss falls, accuracy rises, and the trained weights are what the forward path then uses."""
    vocab = [40] * 26
    m = make_model(vocab, 8, 10, (64, 32), dropout=0.2, seed=2)
    rng = np.random.default_rng(0)
    N = 4000 + 17                                                # ragged last batch
    ids = np.stack([rng.integers(0, v, N) for v in vocab], axis=1)
    dense = rng.random((N, 13))
    y = (ids[:, 0] < 20).astype(np.float32)
    X = np.concatenate([dense, ids], axis=1).astype(np.float32)
    p0 = npy(m(X))[:, 0]
    compile_fit(m, X, y, batch_size=32, epochs=8, sgd=0.1, verbose=False)
    losses = [l for l, _ in m.history]
    assert losses[-1] < 0.8 * losses[0] and m.history[-1][1] > m.history[0][1] + 0.05, m.history
    p1 = npy(m(X))[:, 0]                                          # the forward path sees the trained weights (versions bumped)
    assert np.mean((p1 > 0.5) == (y > 0.5)) > np.mean((p0 > 0.5) == (y > 0.5)) + 0.05
    want = O.deepfm_loss(dense, ids, *params(m)[:5], params(m)[5], y.astype(np.float64), 0.0, 0.0)[2]
    assert np.max(np.abs(p1 - want)) <= 2e-5


@pytest.mark.parametrize("rate", [0.0, 0.2])
def test_graph_replay_equals_call_by_call(lib, rate):
    """compile_fit replays one captured CUDA graph per full batch (ctrhot_deepfm_train_step_ctl_f32: the batch's first row and the
    step number live in a device control block): same weights and the same epoch loss / accuracy as launching every batch by
    hand, incl. the ragged last batch and the dropout masks."""
    vocab = [30] * 26
    rng = np.random.default_rng(1)
    N = 32 * 20 + 11
    ids = np.stack([rng.integers(0, v, N) for v in vocab], axis=1)
    X = np.concatenate([rng.random((N, 13)), ids], axis=1).astype(np.float32)
    y = (ids[:, 3] < 15).astype(np.float32)
    ms = []
    for use_graph in (True, False):
        m = make_model(vocab, 8, 10, (32, 16), dropout=rate, seed=4)
        compile_fit(m, X, y, batch_size=32, epochs=3, sgd=0.05, seed=7, verbose=False, use_graph=use_graph)
        ms.append(m)
    for (l1, a1), (l2, a2) in zip(ms[0].history, ms[1].history):
        assert abs(l1 - l2) <= 1e-5 * max(1.0, abs(l2)) and abs(a1 - a2) <= 1.5 / N, (ms[0].history, ms[1].history)
    pa, pb = params(ms[0]), params(ms[1])
    for ta, tb in zip(pa[0], pb[0]):
        np.testing.assert_allclose(ta, tb, rtol=0, atol=2e-6)            # atomic adds: the order of duplicate rows may differ
    np.testing.assert_allclose(pa[3], pb[3], rtol=0, atol=2e-6)
    for (ka, ba), (kb, bb) in zip(pa[4], pb[4]):
        np.testing.assert_allclose(ka, kb, rtol=0, atol=2e-6)
        np.testing.assert_allclose(ba, bb, rtol=0, atol=2e-6)


def test_fm_step_and_compile_fit(lib):
    """FM (DL/model/fm.py:14-23) on a dense one-hot matrix shaped like the reference's (13 dense columns + one-hot blocks):
    one step against the oracle's fm_train_step, then compile_fit learns a planted rule."""
    from recommender_system_b200.model.criteo import FM
    from recommender_system_b200.utils.compile_fit import _FMTrainer
    rng = np.random.default_rng(2)
    vocab = [30, 7, 120, 50]
    N, n = 2000 + 13, 13 + sum(vocab)
    ids = np.stack([rng.integers(0, v, N) for v in vocab], axis=1)
    X = np.zeros((N, n), np.float32)
    X[:, :13] = rng.random((N, 13))
    off = 13
    for f, v in enumerate(vocab):
        X[np.arange(N), off + ids[:, f]] = 1.0
        off += v
    y = (ids[:, 0] < 15).astype(np.float32)
    set_seed(9)
    m = FM(8, 1e-2, 2e-2)
    m.build((None, n))
    m.set_weights({"FMLayer_0/w0": np.asarray([0.05], np.float32)})
    w = {k: npy(t) for k, t in m.named_weights().items()}
    before = (w["FMLayer_0/w0"], w["FMLayer_0/w1"], w["FMLayer_0/v"])
    tr = _FMTrainer(m)
    dev = m.device
    out = tr.train_on_batch(torch.from_numpy(X[:257]).to(dev), torch.from_numpy(y[:257]).to(dev), 0.1)
    torch.cuda.synchronize()
    want, bce, correct = O.fm_train_step(X[:257].astype(np.float64), *before, y[:257].astype(np.float64), 0.1, 1e-2, 2e-2)
    assert abs(float(out[0]) - bce) <= 2e-6 and int(out[1]) == correct
    w = {k: npy(t) for k, t in m.named_weights().items()}
    close(w["FMLayer_0/w0"], want[0], before[0], "w0"); close(w["FMLayer_0/w1"], want[1], before[1], "w1")
    close(w["FMLayer_0/v"], want[2], before[2], "v")
    compile_fit(m, X, y, batch_size=32, epochs=6, sgd=0.1, verbose=False)
    assert m.history[-1][0] < 0.8 * m.history[0][0] and m.history[-1][1] > 0.85, m.history
    p = npy(m(X))[:, 0]
    assert np.mean((p > 0.5) == (y > 0.5)) > 0.85


# ------------------------------------------------------------------ DCN (DL/model/dcn.py under compile_fit)
def make_dcn(vocab, k_embed, L, hidden, dropout=0.0, seed=1):
    from recommender_system_b200.model.criteo import DCN
    set_seed(seed)
    fc = [[{"feat": "I%d" % i} for i in range(13)],
          [{"feat": "C%d" % i, "feat_onehot_dim": int(v), "embed_dim": k_embed} for i, v in enumerate(vocab)]]
    m = DCN(fc, list(hidden), 1, "relu", L, 1e-2, 2e-2, embed_dim=k_embed, precision="fp32")
    m.build(None)
    m.dense_layer.dropout = dropout
    rng = np.random.default_rng(seed)
    upd = {}
    for n, w in m.named_weights().items():
        if n.endswith("bias"):                                    # keras' zero biases would hide mistakes in their gradients
            upd[n] = rng.normal(0, 0.1, tuple(w.shape))
    m.set_weights({n: np.asarray(v, np.float32) for n, v in upd.items()})
    return m


def dcn_params(m):
    tables = [npy(t) for t in m.embed_layer.emb_layers]
    ws = [npy(t) for t in m.cross_layer.cross_weight]
    bs = [npy(t) for t in m.cross_layer.cross_bias]
    hidden = [(npy(k), npy(b)) for k, b in m.dense_layer.hidden]
    out = (npy(m.dense_layer.output[0]), npy(m.dense_layer.output[1]))
    return tables, ws, bs, hidden, out, (npy(m.out_kernel), npy(m.out_bias))


@pytest.mark.parametrize("B,vocab,k_embed,L,hidden,rate", [
    (32, [50, 7, 1000, 3] + [20] * 22, 8, 6, (256, 128, 64), 0.0),        # the reference's __main__: layer_num 6, batch 32
    (1, [11] * 26, 4, 1, (16,), 0.0),
    (257, [50, 7, 1000, 3] + [20] * 22, 16, 3, (64, 32), 0.2),
    (4096, [5000] * 26, 16, 3, (400, 400, 400), 0.2),                     # BASELINE config 3's cross depth and tower
])
def test_dcn_one_step_against_the_oracle(lib, B, vocab, k_embed, L, hidden, rate):
    from recommender_system_b200.utils.compile_fit import _DCNTrainer
    m = make_dcn(vocab, k_embed, L, hidden, rate)
    dense, ids, y = batch(vocab, B, seed=B + 1)
    before = dcn_params(m)
    lr, seed = 0.05, 4321
    tr = _DCNTrainer(m)
    dev = m.device
    out = tr.train_on_batch(torch.from_numpy(dense).to(dev), torch.from_numpy(ids).to(dev), torch.from_numpy(y).to(dev), lr, seed=seed)
    torch.cuda.synchronize()
    masks = kernel_masks(seed, hidden, B, rate) if rate > 0 else None
    want, bce, correct = O.dcn_train_step(dense.astype(np.float64), ids, *before, y.astype(np.float64), lr, 1e-2, 2e-2, masks)
    got = dcn_params(m)
    assert abs(float(out[0]) - bce) <= 2e-6 * max(1.0, bce) and int(out[1]) == correct
    for f in range(len(vocab)):
        close(got[0][f], want[0][f], before[0][f], "table %d" % f)
        untouched = np.setdiff1d(np.arange(vocab[f]), ids[:, f])
        np.testing.assert_array_equal(got[0][f][untouched], before[0][f][untouched])
    for l in range(L):
        close(got[1][l], want[1][l], before[1][l], "cross w %d" % l)
        close(got[2][l], want[2][l], before[2][l], "cross b %d" % l)
    for l in range(len(hidden)):
        close(got[3][l][0], want[3][l][0], before[3][l][0], "kernel %d" % l)
        close(got[3][l][1], want[3][l][1], before[3][l][1], "bias %d" % l)
    close(got[4][0], want[4][0], before[4][0], "dnn out kernel"); close(got[4][1], want[4][1], before[4][1], "dnn out bias")
    close(got[5][0], want[5][0], before[5][0], "final kernel"); close(got[5][1], want[5][1], before[5][1], "final bias")


def test_dcn_compile_fit_learns_and_serves(lib):
    """compile_fit(DCN, X, y) lowers the loss on a planted rule; the forward path (whole-model kernel, re-packed after the
    weights moved) then agrees with the float64 oracle on the trained weights."""
    vocab = [30, 7, 120, 50] + [9] * 22
    rng = np.random.default_rng(3)
    N = 1500 + 17
    ids = np.stack([rng.integers(0, v, N) for v in vocab], axis=1)
    X = np.concatenate([rng.random((N, 13)), ids], axis=1).astype(np.float32)
    y = (ids[:, 0] < 15).astype(np.float32)
    m = make_dcn(vocab, 8, 3, (64, 32), 0.0, seed=5)
    compile_fit(m, X, y, batch_size=32, epochs=6, sgd=0.1, verbose=False)
    assert m.history[-1][0] < 0.7 * m.history[0][0] and m.history[-1][1] > 0.9, m.history
    p = dcn_params(m)
    want = O.dcn_model(X[:300].astype(np.float64), p[0], p[1], p[2], p[3], p[4], p[5])
    got = npy(m(X[:300]))
    assert np.max(np.abs(got - want)) <= 2e-5, np.max(np.abs(got - want))


# ------------------------------------------------------------------ PNN (the reference's own loop, DL/model/pnn.py:74-85)
def make_pnn(vocab, mode, hidden, dropout=0.0, seed=1, out_bias=0.45):
    from recommender_system_b200.model.criteo import PNN
    set_seed(seed)
    fc = [[{"feat": "I%d" % i} for i in range(13)],
          [{"feat": "C%d" % i, "feat_onehot_dim": int(v), "embed_dim": 8} for i, v in enumerate(vocab)]]
    m = PNN(fc, mode, list(hidden), 1, "relu", dropout, embed_dim=8, precision="fp32")
    m.build(None)
    rng = np.random.default_rng(seed)
    upd = {}
    for n, w in m.named_weights().items():
        if n.endswith("bias"):
            upd[n] = rng.normal(0, 0.05, tuple(w.shape))
        elif n.endswith("embeddings"):                            # rows large enough for the products to matter
            upd[n] = rng.normal(0, 0.3, tuple(w.shape))
    m.set_weights({n: np.asarray(v, np.float32) for n, v in upd.items()})
    m.dnn_layer.output[1].fill_(out_bias)                         # raw outputs around 0.45: inside keras' clip for most samples
    return m


def pnn_params(m):
    tables = [npy(t) for t in m.embed_layer.emb_layers]
    hidden = [(npy(k), npy(b)) for k, b in m.dnn_layer.hidden]
    return tables, npy(m.outer_product_layer.w), hidden, (npy(m.dnn_layer.output[0]), npy(m.dnn_layer.output[1]))


@pytest.mark.parametrize("B,F,mode,hidden,rate", [
    (32, 26, "both", (256, 128, 64), 0.0),                               # the reference's batch, tower and mode
    (1, 3, "inner", (16,), 0.0),
    (257, 26, "outer", (64, 32), 0.3),
    (1000, 13, "both", (64, 32), 0.3),
])
def test_pnn_one_step_against_the_oracle(lib, B, F, mode, hidden, rate):
    from recommender_system_b200.utils.compile_fit import _PNNTrainer
    vocab = ([50, 7, 1000, 3] + [20] * 22)[:F]
    m = make_pnn(vocab, mode, hidden, rate)
    _, ids, y = batch(vocab, B, seed=B + 2)
    before = pnn_params(m)
    lr, seed = 0.05, 999
    tr = _PNNTrainer(m)
    dev = m.device
    out = tr.train_on_batch(torch.from_numpy(ids).to(dev), torch.from_numpy(y).to(dev), lr, seed=seed)
    torch.cuda.synchronize()
    masks = kernel_masks(seed, hidden, B, rate) if rate > 0 else None
    want, loss, correct = O.pnn_train_step(ids, *before, y.astype(np.float64), lr, mode, masks)
    _, pre = O.pnn_loss(ids, *before, y.astype(np.float64), mode, masks)
    assert np.mean((pre > 1e-3) & (pre < 1 - 1e-3)) > 0.5                 # the step is not vacuous: most samples carry gradient
    got = pnn_params(m)
    assert abs(float(out[0]) - loss) <= 1e-5 * max(1.0, loss) and abs(int(out[1]) - correct) <= 1
    for f in range(F):
        close(got[0][f], want[0][f], before[0][f], "table %d" % f, floor=1e-6)
        untouched = np.setdiff1d(np.arange(vocab[f]), ids[:, f])
        np.testing.assert_array_equal(got[0][f][untouched], before[0][f][untouched])
    if mode != "inner":
        close(got[1], want[1], before[1], "outer W")
    for l in range(len(hidden)):
        close(got[2][l][0], want[2][l][0], before[2][l][0], "kernel %d" % l)
        close(got[2][l][1], want[2][l][1], before[2][l][1], "bias %d" % l)
    close(got[3][0], want[3][0], before[3][0], "out kernel"); close(got[3][1], want[3][1], before[3][1], "out bias")


def test_pnn_compile_fit_lowers_the_loss(lib):
    """compile_fit on PNN runs the reference's loop; the loss of its (sigmoid-free) output goes down on a planted rule and the
    one-launch forward path then serves the trained weights."""
    vocab = [30, 7, 120, 50] + [9] * 22
    rng = np.random.default_rng(3)
    N = 1200 + 5
    ids = np.stack([rng.integers(0, v, N) for v in vocab], axis=1)
    X = np.concatenate([rng.random((N, 13)), ids], axis=1).astype(np.float32)
    y = (ids[:, 0] < 15).astype(np.float32)
    m = make_pnn(vocab, "both", (64, 32), 0.0, seed=5)
    compile_fit(m, X, y, batch_size=32, epochs=5, sgd=0.02, verbose=False)
    assert m.history[-1][0] < 0.9 * m.history[0][0], m.history
    p = pnn_params(m)
    want = O.pnn_model(X[:300].astype(np.float64), p[0], "both", p[2], p[3], outer_w=p[1])
    got = npy(m(X[:300]))
    assert np.max(np.abs(got - want)) <= 5e-5, np.max(np.abs(got - want))


# ------------------------------------------------------------------ WideDeep (DL/model/wideDeep.py under compile_fit)
def make_widedeep(vocab, widths, k_embed, hidden, dropout=0.0, seed=1):
    from recommender_system_b200.model.criteo import WideDeep
    set_seed(seed)
    fc = [[{"feat": "I%d" % i} for i in range(13)],
          [{"feat": "C%d" % i, "feat_onehot_dim": int(v), "embed_dim": k_embed} for i, v in enumerate(vocab)]]
    m = WideDeep(fc, list(hidden), 1, "relu", embed_dim=k_embed, precision="fp32", onehot_width=widths)
    m.build(None)
    m.deep.dropout = dropout
    rng = np.random.default_rng(seed)
    upd = {n: rng.normal(0, 0.1, tuple(w.shape)) for n, w in m.named_weights().items() if n.endswith("bias") or n.endswith("w0")}
    m.set_weights({n: np.asarray(v, np.float32) for n, v in upd.items()})
    return m


def widedeep_params(m):
    tables = [npy(t) for t in m.embed_layer.emb_layers]
    hidden = [(npy(k), npy(b)) for k, b in m.deep.hidden]
    return tables, npy(m.wide.w0), npy(m.wide.w), hidden, (npy(m.deep.output[0]), npy(m.deep.output[1]))


@pytest.mark.parametrize("B,k_embed,hidden,rate", [(32, 8, (256, 128, 64), 0.0), (1, 4, (16,), 0.0), (515, 16, (64, 32), 0.2)])
def test_widedeep_one_step_against_the_oracle(lib, B, k_embed, hidden, rate):
    from recommender_system_b200.utils.compile_fit import _WideDeepTrainer
    vocab = [50, 7, 1000, 3] + [20] * 22
    widths = [50, 5, 900, 3] + [20] * 22                          # get_dummies saw fewer labels in fields 1 and 2
    m = make_widedeep(vocab, widths, k_embed, hidden, rate)
    dense, ids, y = batch(vocab, B, seed=B + 3)
    t, w0, w, hid, out = widedeep_params(m)
    lr, seed = 0.05, 2024
    tr = _WideDeepTrainer(m)
    dev = m.device
    res = tr.train_on_batch(torch.from_numpy(dense).to(dev), torch.from_numpy(ids).to(dev), torch.from_numpy(y).to(dev), lr, seed=seed)
    torch.cuda.synchronize()
    masks = kernel_masks(seed, hidden, B, rate) if rate > 0 else None
    want, bce, correct = O.widedeep_train_step(dense.astype(np.float64), ids, t, w0, w, widths, hid, out, y.astype(np.float64),
                                               lr, 1e-4, masks)
    g = widedeep_params(m)
    assert abs(float(res[0]) - bce) <= 2e-6 * max(1.0, bce) and int(res[1]) == correct
    for f in range(len(vocab)):
        close(g[0][f], want[0][f], t[f], "table %d" % f)
    close(g[1], want[1], w0, "w0"); close(g[2], want[2], w, "wide w")
    for l in range(len(hidden)):
        close(g[3][l][0], want[3][l][0], hid[l][0], "kernel %d" % l)
        close(g[3][l][1], want[3][l][1], hid[l][1], "bias %d" % l)
    close(g[4][0], want[4][0], out[0], "out kernel"); close(g[4][1], want[4][1], out[1], "out bias")


def test_widedeep_compile_fit_learns_and_serves(lib):
    vocab = [30, 7, 120, 50] + [9] * 22
    rng = np.random.default_rng(3)
    N = 1500 + 9
    ids = np.stack([rng.integers(0, v, N) for v in vocab], axis=1)
    X = np.concatenate([rng.random((N, 13)), ids], axis=1).astype(np.float32)
    y = (ids[:, 0] < 15).astype(np.float32)
    m = make_widedeep(vocab, vocab, 8, (64, 32), 0.0, seed=5)
    compile_fit(m, X, y, batch_size=32, epochs=12, sgd=0.1, verbose=False)
    # sigmoid(0.5 (wide + deep)) halves every gradient: slower than DeepFM / DCN on the same rule, but monotone
    assert m.history[-1][0] < 0.85 * m.history[0][0] and m.history[-1][1] > 0.8, m.history
    assert all(b[0] < a[0] for a, b in zip(m.history, m.history[1:])), m.history
    t, w0, w, hid, out = widedeep_params(m)
    xw = O._widedeep_wide_input(X[:300, :13].astype(np.float64), ids[:300], vocab)
    want = O.widedeep_model(np.concatenate([X[:300].astype(np.float64), xw[:, 13:]], axis=1), t, (w0, w), hid, out)
    got = npy(m((X[:300, :13], ids[:300].astype(np.int32))))
    assert np.max(np.abs(got - want)) <= 2e-5, np.max(np.abs(got - want))


# ------------------------------------------------------------------ DeepCrossing (DL/model/deepCrossing.py under compile_fit)
def make_deepcrossing(vocab, k_embed, hidden, n_res, seed=1):
    from recommender_system_b200.model.deepCrossing import DeepCrossing
    set_seed(seed)
    fc = [[{"feat": "I%d" % i} for i in range(13)],
          [{"feat": "C%d" % i, "feat_onehot_dim": int(v), "embed_dim": k_embed} for i, v in enumerate(vocab)]]
    m = DeepCrossing(fc, 32, list(hidden), n_res, embed_dim=k_embed, precision="fp32")
    m.build(None)
    rng = np.random.default_rng(seed)
    upd = {n: rng.normal(0, 0.1, tuple(w.shape)) for n, w in m.named_weights().items() if n.endswith("bias")}
    m.set_weights({n: np.asarray(v, np.float32) for n, v in upd.items()})
    return m


def deepcrossing_params(m):
    tables = [npy(t) for t in m.embed_layer.emb_layers]
    res = [([(npy(k), npy(b)) for k, b in r.dense_layer], (npy(r.output_layer[0]), npy(r.output_layer[1]))) for r in m.res_layer]
    return tables, res, (npy(m.output_layer[0]), npy(m.output_layer[1]))


@pytest.mark.parametrize("B,k_embed,hidden,n_res", [(32, 8, (256, 256), 4), (1, 4, (16,), 1), (300, 16, (64, 32), 2)])
def test_deepcrossing_one_step_against_the_oracle(lib, B, k_embed, hidden, n_res):
    """(32, 8, (256, 256), 4) is the reference's __main__ (deepCrossing.py:37-40) at compile_fit's batch size."""
    from recommender_system_b200.utils.compile_fit import _DeepCrossingTrainer
    vocab = [50, 7, 1000, 3] + [20] * 22
    m = make_deepcrossing(vocab, k_embed, hidden, n_res)
    dense, ids, y = batch(vocab, B, seed=B + 4)
    t, res, fin = deepcrossing_params(m)
    lr = 0.05
    tr = _DeepCrossingTrainer(m)
    dev = m.device
    out = tr.train_on_batch(torch.from_numpy(dense).to(dev), torch.from_numpy(ids).to(dev), torch.from_numpy(y).to(dev), lr)
    torch.cuda.synchronize()
    want, bce, correct = O.deepcrossing_train_step(dense.astype(np.float64), ids, t, res, fin, y.astype(np.float64), lr)
    g = deepcrossing_params(m)
    assert abs(float(out[0]) - bce) <= 2e-6 * max(1.0, bce) and int(out[1]) == correct
    for f in range(len(vocab)):
        close(g[0][f], want[0][f], t[f], "table %d" % f)
    for r in range(n_res):
        for l in range(len(hidden)):
            close(g[1][r][0][l][0], want[1][r][0][l][0], res[r][0][l][0], "unit %d kernel %d" % (r, l))
            close(g[1][r][0][l][1], want[1][r][0][l][1], res[r][0][l][1], "unit %d bias %d" % (r, l))
        close(g[1][r][1][0], want[1][r][1][0], res[r][1][0], "unit %d out kernel" % r)
        close(g[1][r][1][1], want[1][r][1][1], res[r][1][1], "unit %d out bias" % r)
    close(g[2][0], want[2][0], fin[0], "final kernel"); close(g[2][1], want[2][1], fin[1], "final bias")


def test_deepcrossing_compile_fit_learns_and_serves(lib):
    vocab = [30, 7, 120, 50] + [9] * 22
    rng = np.random.default_rng(3)
    N = 1500 + 11
    ids = np.stack([rng.integers(0, v, N) for v in vocab], axis=1)
    X = np.concatenate([rng.random((N, 13)), ids], axis=1).astype(np.float32)
    y = (ids[:, 0] < 15).astype(np.float32)
    m = make_deepcrossing(vocab, 8, (64, 32), 2, seed=5)
    compile_fit(m, X, y, batch_size=32, epochs=10, sgd=0.2, verbose=False)
    assert m.history[-1][0] < 0.8 * m.history[0][0] and m.history[-1][1] > 0.8, m.history
    t, res, fin = deepcrossing_params(m)
    want = O.deepcrossing_model(X[:300].astype(np.float64), t, res, fin)
    got = npy(m(X[:300]))
    assert np.max(np.abs(got - want)) <= 2e-5, np.max(np.abs(got - want))


# ------------------------------------------------------------------ NFM (DL/model/nfm.py under compile_fit, BN in training mode)
def make_nfm(vocab, k_embed, hidden, od=1, dropout=0.0, seed=1):
    from recommender_system_b200.model.criteo import NFM
    set_seed(seed)
    fc = [[{"feat": "I%d" % i} for i in range(13)],
          [{"feat": "C%d" % i, "feat_onehot_dim": int(v), "embed_dim": k_embed} for i, v in enumerate(vocab)]]
    m = NFM(fc, list(hidden), od, "relu", dropout, embed_dim=k_embed, precision="fp32")
    m.build(None)
    rng = np.random.default_rng(seed)
    upd = {}
    for n, w in m.named_weights().items():
        if n.endswith("bias") or n.endswith("beta") or n.endswith("moving_mean"):
            upd[n] = rng.normal(0, 0.1, tuple(w.shape))
        elif n.endswith("gamma") or n.endswith("moving_variance"):
            upd[n] = rng.uniform(0.5, 1.5, tuple(w.shape))
        elif n.endswith("embeddings"):
            upd[n] = rng.normal(0, 0.3, tuple(w.shape))
    m.set_weights({n: np.asarray(v, np.float32) for n, v in upd.items()})
    return m


def nfm_params(m):
    tables = [npy(t) for t in m.emb_layers.emb_layers]
    bn = (npy(m.bn_mean), npy(m.bn_var), npy(m.bn_gamma), npy(m.bn_beta))
    hidden = [(npy(k), npy(b)) for k, b in m.dnn_layers.hidden]
    return tables, bn, hidden, (npy(m.dnn_layers.output[0]), npy(m.dnn_layers.output[1])), (npy(m.out_kernel), npy(m.out_bias))


@pytest.mark.parametrize("B,k_embed,hidden,od,rate", [(32, 8, (256, 128, 64), 1, 0.0), (5, 4, (16,), 3, 0.0), (700, 16, (64, 32), 1, 0.3)])
def test_nfm_one_step_against_the_oracle(lib, B, k_embed, hidden, od, rate):
    from recommender_system_b200.utils.compile_fit import _NFMTrainer
    vocab = [50, 7, 1000, 3] + [20] * 22
    m = make_nfm(vocab, k_embed, hidden, od, rate)
    dense, ids, y = batch(vocab, B, seed=B + 5)
    t, bn, hid, out, fin = nfm_params(m)
    lr, seed = 0.05, 77
    tr = _NFMTrainer(m)
    dev = m.device
    res = tr.train_on_batch(torch.from_numpy(dense).to(dev), torch.from_numpy(ids).to(dev), torch.from_numpy(y).to(dev), lr, seed=seed)
    torch.cuda.synchronize()
    masks = kernel_masks(seed, hidden, B, rate) if rate > 0 else None
    want, bce, correct = O.nfm_train_step(dense.astype(np.float64), ids, t, bn, hid, out, fin, y.astype(np.float64), lr, masks)
    g = nfm_params(m)
    assert abs(float(res[0]) - bce) <= 5e-6 * max(1.0, bce) and abs(int(res[1]) - correct) <= 1
    for f in range(len(vocab)):
        close(g[0][f], want[0][f], t[f], "table %d" % f, floor=1e-6)
    for i, name in enumerate(("moving mean", "moving var", "gamma", "beta")):
        close(g[1][i], want[1][i], bn[i], name, floor=1e-6)
    for l in range(len(hidden)):
        close(g[2][l][0], want[2][l][0], hid[l][0], "kernel %d" % l, floor=1e-6)
        close(g[2][l][1], want[2][l][1], hid[l][1], "bias %d" % l, floor=1e-6)
    close(g[3][0], want[3][0], out[0], "dnn out kernel", floor=1e-6); close(g[3][1], want[3][1], out[1], "dnn out bias", floor=1e-6)
    close(g[4][0], want[4][0], fin[0], "final kernel", floor=1e-6); close(g[4][1], want[4][1], fin[1], "final bias", floor=1e-6)


def test_nfm_compile_fit_learns_and_serves(lib):
    """After training, the forward path (inference BatchNormalization on the MOVING statistics the steps maintained) agrees with
    the float64 oracle on the trained weights."""
    vocab = [30, 7, 120, 50] + [9] * 22
    rng = np.random.default_rng(3)
    N = 1500 + 13
    ids = np.stack([rng.integers(0, v, N) for v in vocab], axis=1)
    X = np.concatenate([rng.random((N, 13)), ids], axis=1).astype(np.float32)
    y = (ids[:, 0] < 15).astype(np.float32)
    m = make_nfm(vocab, 8, (64, 32), 1, 0.0, seed=5)
    compile_fit(m, X, y, batch_size=32, epochs=8, sgd=0.1, verbose=False)
    assert m.history[-1][0] < m.history[0][0], m.history
    t, bn, hid, out, fin = nfm_params(m)
    want = O.nfm_model(X[:300].astype(np.float64), t, bn, hid, out, fin)
    got = npy(m(X[:300]))
    assert np.max(np.abs(got - want)) <= 5e-5, np.max(np.abs(got - want))


# ------------------------------------------------------------------ AFM (DL/model/afm.py under compile_fit)
def make_afm(vocab, mode, seed=1):
    from recommender_system_b200.model.criteo import AFM
    set_seed(seed)
    fc = [[{"feat": "I%d" % i} for i in range(13)],
          [{"feat": "C%d" % i, "feat_onehot_dim": int(v), "embed_dim": 8} for i, v in enumerate(vocab)]]
    m = AFM(fc, mode)
    m.build(None)
    rng = np.random.default_rng(seed)
    upd = {}
    for n, w in m.named_weights().items():
        if n.endswith("embeddings"):
            upd[n] = rng.normal(0, 0.4, tuple(w.shape))
        elif n.startswith("output_layer"):
            upd[n] = rng.normal(0, 0.5, tuple(w.shape))
    m.set_weights({n: np.asarray(v, np.float32) for n, v in upd.items()})
    return m


@pytest.mark.parametrize("B,F,mode", [(32, 26, "att"), (1, 3, "avg"), (257, 26, "max"), (600, 13, "avg")])
def test_afm_one_step_against_the_oracle(lib, B, F, mode):
    from recommender_system_b200.utils.compile_fit import _AFMTrainer
    vocab = ([50, 7, 1000, 3] + [20] * 22)[:F]
    m = make_afm(vocab, mode)
    _, ids, y = batch(vocab, B, seed=B + 6)
    tabs = [npy(t) for t in m.embed_layer.emb_layers]
    out = (npy(m.out_kernel), npy(m.out_bias))
    att = [npy(w).copy() for n, w in m.named_weights().items() if n.startswith("attention")]
    lr = 0.5                                                      # the double sigmoid makes gradients small
    tr = _AFMTrainer(m)
    dev = m.device
    res = tr.train_on_batch(None, torch.from_numpy(ids).to(dev), torch.from_numpy(y).to(dev), lr)
    torch.cuda.synchronize()
    want, bce, correct = O.afm_train_step(ids, tabs, mode, out, y.astype(np.float64), lr, 2)
    assert abs(float(res[0]) - bce) <= 2e-6 * max(1.0, bce) and int(res[1]) == correct
    got = [npy(t) for t in m.embed_layer.emb_layers]
    for f in range(F):
        close(got[f], want[0][f], tabs[f], "table %d" % f, floor=1e-6)
    close(npy(m.out_kernel), want[1][0], out[0], "out kernel"); close(npy(m.out_bias), want[1][1], out[1], "out bias")
    for a, (n, w) in zip(att, [(n, w) for n, w in m.named_weights().items() if n.startswith("attention")]):
        np.testing.assert_array_equal(npy(w), a)                  # zero gradient: AttentionLayer's weights do not move


def test_afm_compile_fit_lowers_the_loss_and_serves(lib):
    vocab = [30, 7, 120, 50] + [9] * 22
    rng = np.random.default_rng(3)
    N = 1000 + 3
    ids = np.stack([rng.integers(0, v, N) for v in vocab], axis=1)
    X = np.concatenate([rng.random((N, 13)), ids], axis=1).astype(np.float32)
    y = (ids[:, 0] < 15).astype(np.float32)
    m = make_afm(vocab, "att", seed=5)
    compile_fit(m, X, y, batch_size=32, epochs=5, sgd=0.5, verbose=False)
    assert m.history[-1][0] < m.history[0][0], m.history
    tabs = [npy(t) for t in m.embed_layer.emb_layers]
    want = O.afm_model(X[:300].astype(np.float64), tabs, "att", att_w=(np.zeros((8, 325)), np.zeros(325)),
                       att_h=(np.zeros((325, 1)), np.zeros(1)), out=(npy(m.out_kernel), npy(m.out_bias)))
    got = npy(m(X[:300]))
    assert np.max(np.abs(got - want)) <= 2e-5, np.max(np.abs(got - want))


# ------------------------------------------------------------------ FFM (DL/model/ffm.py under compile_fit)
@pytest.mark.parametrize("B,vocab,k", [(32, [50, 7, 1000, 3] + [20] * 22, 8), (1, [11] * 26, 4), (257, [30, 7, 120, 50], 8),
                                       (600, [5000] * 26, 8)])
def test_ffm_one_step_against_the_oracle(lib, B, vocab, k):
    """One SGD step of FFM against oracle.ffm_train_step: every entry of w0, w and v (rows the batch did not touch move by the
    l2 decay alone), the loss and the accuracy.  Ids arrive as floats inside the [B, 13 + F] matrix, like the reference's."""
    from recommender_system_b200.model.criteo import FFM
    from recommender_system_b200.utils.compile_fit import _FFMTrainer
    set_seed(4)
    fc = [[{"feat": "I%d" % i} for i in range(13)],
          [{"feat": "C%d" % i, "feat_onehot_dim": int(v)} for i, v in enumerate(vocab)]]
    m = FFM(fc, k, 1e-2, 2e-2)
    m.build((None, 13 + len(vocab)))
    L = m.ffm
    rng = np.random.default_rng(B)
    with torch.no_grad():
        L.w0.copy_(torch.tensor([0.05]))
        L.v.copy_(torch.from_numpy(rng.normal(0, 0.02, tuple(L.v.shape)).astype(np.float32)))
    dense, ids, y = batch(vocab, B, seed=B + 7)
    X = np.concatenate([dense, ids.astype(np.float32)], axis=1)
    before = (npy(L.w0), npy(L.w), npy(L.v))
    tr = _FFMTrainer(m)
    dev = m.device
    lr = 0.1
    out = tr.train_on_batch(torch.from_numpy(X).to(dev), torch.from_numpy(y).to(dev), lr)
    torch.cuda.synchronize()
    want, bce, correct = O.ffm_train_step(X.astype(np.float64), vocab, *before, y.astype(np.float64), lr, 1e-2, 2e-2)
    # (1e-5: z = 0.5 (|S|^2 - |ff|^2) over 39 x 8 terms cancels in fp32, as it does in the reference's own fp32 graph)
    assert abs(float(out[0]) - bce) <= 1e-5 * max(1.0, bce) and int(out[1]) == correct
    close(npy(L.w0), want[0], before[0], "w0"); close(npy(L.w), want[1], before[1], "w")
    close(npy(L.v), want[2], before[2], "v", floor=1e-6)
    # the serving kernel reads the same updated weights
    got = npy(m(X))
    ref = O.sigmoid(O.ffm_layer(X.astype(np.float64), vocab, npy(L.w0), npy(L.w), npy(L.v)))
    assert np.max(np.abs(got - ref)) <= 2e-5


def test_ffm_compile_fit_lowers_the_loss(lib):
    from recommender_system_b200.model.criteo import FFM
    vocab = [30, 7, 120, 50] + [9] * 22
    rng = np.random.default_rng(3)
    N = 1000 + 3
    ids = np.stack([rng.integers(0, v, N) for v in vocab], axis=1)
    X = np.concatenate([rng.random((N, 13)), ids], axis=1).astype(np.float32)
    y = (ids[:, 0] < 15).astype(np.float32)
    set_seed(6)
    fc = [[{"feat": "I%d" % i} for i in range(13)],
          [{"feat": "C%d" % i, "feat_onehot_dim": int(v)} for i, v in enumerate(vocab)]]
    m = FFM(fc, 8, 1e-4, 1e-4)
    m.build((None, X.shape[1]))
    with torch.no_grad():                    # keras' random_normal(0.05) start saturates the sigmoid (39 fields x 8): the clipped BCE
        m.ffm.v.mul_(0.2)                    # then has zero gradient on the wrong-side samples, in the reference as well
    compile_fit(m, X, y, batch_size=32, epochs=6, sgd=0.01, verbose=False)      # (sgd = 0.05 diverges, in the oracle as well)
    assert m.history[-1][0] < 0.5 * m.history[0][0] and m.history[-1][1] > 0.95, m.history


# ------------------------------------------------------------------ DIN (DL/model/din.py under compile_fit)
def make_din(v_item, v_user, K, T, att_hidden, dnn_hidden, seed=1):
    from recommender_system_b200.model.din import DIN
    from recommender_system_b200.utils import synth
    set_seed(seed)
    m = DIN(synth.din_feature_columns(v_item=v_item, v_user=v_user, embed_dim=K), ["movies_seq"], tuple(att_hidden), tuple(dnn_hidden),
            "prelu", "prelu", 0.0, precision="fp32")
    m.build(None)
    m.att_layer.build_for(T, K)
    rng = np.random.default_rng(seed)
    upd = {}
    for n, w in m.named_weights().items():                       # keras' zero biases / alphas would hide mistakes in their gradients
        if n.endswith("bias") or n.endswith("beta"):
            upd[n] = rng.normal(0, 0.1, tuple(w.shape))
        elif n.endswith("prelu_alpha"):
            upd[n] = rng.normal(0.2, 0.2, tuple(w.shape))
        elif n.endswith("gamma"):
            upd[n] = rng.normal(1.0, 0.2, tuple(w.shape))
        elif n.endswith("embeddings"):
            upd[n] = rng.normal(0, 0.5, tuple(w.shape))
    m.set_weights({n: np.asarray(v, np.float32) for n, v in upd.items()})
    return m


def din_params(m):
    A = m.att_layer
    return dict(other_tables=[npy(t) for t in m.embed_sparse_layers], seq_table=npy(m.embed_seq_layers[0]),
                att_dense=[(npy(k), npy(b), npy(a)) for k, b, a in A.layers], att_out=(npy(A.out_kernel), npy(A.out_bias)),
                bn=(npy(m.bn_gamma), npy(m.bn_beta), npy(m.bn_mean), npy(m.bn_var)),
                tower=[(npy(k), npy(b), npy(a)) for k, b, a in m.dense_layer], out=(npy(m.out_kernel), npy(m.out_bias)))


@pytest.mark.parametrize("B,T,K,att_hidden,dnn_hidden,v_item", [
    (32, 10, 8, (80, 40), (256, 128, 64), 50), (1, 3, 4, (6,), (8,), 50), (257, 7, 16, (20, 10), (64, 32), 50),
    (600, 40, 8, (80, 40), (64, 32), 3000)])      # (3000 items: at 50 every row would take ~480 atomic adds, 3e-6 of fp32 round-off)
def test_din_one_step_against_the_oracle(lib, B, T, K, att_hidden, dnn_hidden, v_item):
    """One SGD step of DIN against oracle.din_train_step: the behaviour table (history + candidate rows, duplicates, padding id
    0, one all-padding history), the other table, every Dense / PReLU alpha of the attention unit and the tower, both out
    layers, BatchNormalization's gamma / beta and moving statistics, the loss and the accuracy."""
    from recommender_system_b200.utils import synth
    from recommender_system_b200.utils.compile_fit import _DINTrainer
    v_user = 9
    m = make_din(v_item, v_user, K, T, att_hidden, dnn_hidden, seed=B)
    d = synth.din_batch(B, T, v_item=v_item, v_user=v_user, seed=B + 3)
    if B > 2:
        d["movies_seq"][2] = 0                                   # whole history masked: uniform softmax over the padding row
        d["movie_id"][1, 0] = max(1, int(d["movies_seq"][1, 0]))  # the candidate's row also sits in its history
    rng = np.random.default_rng(B)
    y = rng.integers(0, 2, B).astype(np.float32)
    P = din_params(m)
    before = din_params(m)
    lr = 0.1
    tr = _DINTrainer(m)
    dev = m.device
    res = tr.train_on_batch({k: torch.from_numpy(v).to(dev) for k, v in d.items()}, torch.from_numpy(y).to(dev), lr)
    torch.cuda.synchronize()
    m.check_ids()
    want, bce, correct = O.din_train_step(d["age"].astype(np.float64), d["user_id"], d["movies_seq"], d["movie_id"][:, 0], y=y.astype(np.float64),
                                          lr=lr, **P)
    assert abs(float(res[0]) - bce) <= 2e-6 * max(1.0, bce) and int(res[1]) == correct
    got = din_params(m)
    close(got["seq_table"], want["seq_table"], before["seq_table"], "seq table", floor=1e-6)
    close(got["other_tables"][0], want["other_tables"][0], before["other_tables"][0], "other table", floor=1e-6)
    for name in ("att_dense", "tower"):
        for l, (g3, w3, b3) in enumerate(zip(got[name], want[name], before[name])):
            for part, what in enumerate(("kernel", "bias", "alpha")):
                close(g3[part], w3[part], b3[part], "%s %d %s" % (name, l, what))
    for name in ("att_out", "out"):
        for part in (0, 1):
            close(got[name][part], want[name][part], before[name][part], "%s %d" % (name, part))
    for part, what in enumerate(("gamma", "beta", "moving mean", "moving variance")):
        close(got["bn"][part], want["bn"][part], before["bn"][part], "bn " + what)


def test_din_compile_fit_learns_and_serves(lib):
    """compile_fit on the reference's dict input (DL/model/din.py:98-108): the loss falls on a planted rule (the label says
    whether the candidate occurs in the history), and the forward path then serves the trained weights with the moving
    statistics the steps maintained."""
    from recommender_system_b200.utils import synth
    v_item, v_user, T, K = 40, 9, 8, 8
    N = 1500 + 7
    d = synth.din_batch(N, T, v_item=v_item, v_user=v_user, seed=5)
    rng = np.random.default_rng(8)
    pos = rng.random(N) < 0.5
    for i in np.nonzero(pos)[0]:
        d["movies_seq"][i, 0] = d["movie_id"][i, 0]
    y = np.array([float(d["movie_id"][i, 0] in d["movies_seq"][i]) for i in range(N)], np.float32)
    m = make_din(v_item, v_user, K, T, (16, 8), (32, 16), seed=2)
    X = {k: list(v) if k == "movies_seq" else v for k, v in d.items()}
    compile_fit(m, X, y, batch_size=32, epochs=8, sgd=0.05, verbose=False)
    assert m.history[-1][0] < 0.9 * m.history[0][0] and m.history[-1][1] > m.history[0][1] + 0.1, m.history
    P = din_params(m)
    sub = {k: v[:300] for k, v in d.items()}
    want = O.din_model(sub["age"].astype(np.float64), sub["user_id"], sub["movies_seq"][:, :, None], sub["movie_id"], P["other_tables"],
                       [P["seq_table"]], P["att_dense"], P["att_out"], P["bn"], P["tower"], P["out"])
    got = npy(m(sub))
    assert np.max(np.abs(got - want)) <= 2e-5, np.max(np.abs(got - want))


# ------------------------------------------------------------------ the staged graph step (trainers without dropout)
def test_staged_graph_equals_call_by_call(lib):
    """compile_fit(use_graph=True) replays one captured step over staging buffers for the trainers whose step has no dropout;
    it must leave the same weights and history as the call-by-call loop (the atomics' order aside: 1e-6)."""
    from recommender_system_b200.utils import synth
    vocab = [30, 7, 120, 50] + [9] * 22
    rng = np.random.default_rng(3)
    N = 32 * 9 + 5
    ids = np.stack([rng.integers(0, v, N) for v in vocab], axis=1)
    X = np.concatenate([rng.random((N, 13)), ids], axis=1).astype(np.float32)
    y = (ids[:, 0] < 15).astype(np.float32)
    pairs = []
    for g in (True, False):
        m = make_afm(vocab, "att", seed=3)
        compile_fit(m, X, y, batch_size=32, epochs=3, sgd=0.5, verbose=False, use_graph=g)
        pairs.append((m.history, {n: npy(w) for n, w in m.named_weights().items()}))
    for (ha, hb) in zip(pairs[0][0], pairs[1][0]):
        assert abs(ha[0] - hb[0]) <= 1e-6 and ha[1] == hb[1]
    for n in pairs[0][1]:
        assert np.max(np.abs(pairs[0][1][n] - pairs[1][1][n])) <= 1e-6, n
    # DIN on its dict input
    v_item, v_user, T, K = 40, 9, 8, 8
    d = synth.din_batch(N, T, v_item=v_item, v_user=v_user, seed=5)
    pairs = []
    for g in (True, False):
        m = make_din(v_item, v_user, K, T, (16, 8), (32, 16), seed=2)
        compile_fit(m, dict(d), y, batch_size=32, epochs=3, sgd=0.05, verbose=False, use_graph=g)
        pairs.append((m.history, {n: npy(w) for n, w in m.named_weights().items()}))
    for (ha, hb) in zip(pairs[0][0], pairs[1][0]):
        assert abs(ha[0] - hb[0]) <= 1e-6 and ha[1] == hb[1]
    for n in pairs[0][1]:
        assert np.max(np.abs(pairs[0][1][n] - pairs[1][1][n])) <= 1e-6, n
