"""GPU parity at the REAL sizes of BASELINE.json's configs (VERDICT r01 'configs with no parity evidence at their real size').

The float64 oracle cannot run 65 536 samples against multi-GB tables in seconds, so each test runs the CUDA path on the full
batch and full arena, then checks a strided subset of the samples against the oracle fed with exactly the table rows those
samples touch (fetched from the device arena, re-indexed).  Tolerances are BASELINE north_star's: rows bit-exact, fp32
interaction outputs <= 1e-5 relative, logits <= 2e-3 absolute behind TF32 towers.

  cfg 3  DCN 3 cross + 400-400-400 at B = 65 536 on the kaggle arena (2.16 GB)         DL/model/dcn.py:24-34
  cfg 4  the whole DIN model at B = 16 384, T = 100, K = 64, 80-40 / 256-128-64, 10 M items   DL/model/din.py:56-95
  FFM    the reference's own shape: 43 618 features, 39 fields, k = 8                  DL/model/ffm.py:27-40
  PNN    mode 'both' at B = 65 536, k = 16                                             DL/model/pnn.py:28-53
  R1     gather + whole-model kernel on an arena > 2^32 bytes, ids at the very end     DL/layer/core.py:273-280
"""
import numpy as np
import pytest
import torch

from oracle import ctr_oracle as O
from recommender_system_b200.layer import EmbedLayer, set_seed
from recommender_system_b200.model.criteo import DCN, FFM, PNN, DeepFM
from recommender_system_b200.model.din import DIN
from recommender_system_b200.utils import synth

pytestmark = pytest.mark.gpu

LOGIT = 2e-3
REL = 1e-5


def npy(t):
    return t.detach().cpu().numpy().astype(np.float64)


def logit(p):
    return np.log(p) - np.log1p(-p)


def logit_close(got, want, what=""):
    """Same rule as tests/test_gpu_fused.py: 2e-3 on the logit plus the fp32 sigmoid's own round-trip cost."""
    got, want = np.asarray(got, np.float64).ravel(), np.asarray(want, np.float64).ravel()
    fin = (got > 0) & (got < 1)
    assert np.all(np.abs(logit(want[~fin])) > 16.0)
    tol = LOGIT + 2.0 ** -22 / (want[fin] * (1.0 - want[fin]))
    err = np.abs(logit(got[fin]) - logit(want[fin]))
    assert np.all(err <= tol), "%s: max logit err %.3g (tolerance %.3g)" % (what, err.max(), LOGIT)
    return float(err.max())


def rows_touched(tables_dev, ids_sub):
    """Per field: the float64 rows the subset touches and the ids re-indexed into them."""
    tabs, new_ids = [], np.empty_like(ids_sub)
    for f, t in enumerate(tables_dev):
        u, inv = np.unique(ids_sub[:, f], return_inverse=True)
        tabs.append(npy(t[torch.from_numpy(u.astype(np.int64)).to(t.device)]))
        new_ids[:, f] = inv.reshape(-1)
    return tabs, new_ids


def w64(m):
    return {n: npy(t) for n, t in m.named_weights().items() if "embeddings" not in n}


def tower(w, prefix, n_hidden):
    hidden = [(w["%s/dense_%d/kernel" % (prefix, i)], w["%s/dense_%d/bias" % (prefix, i)]) for i in range(n_hidden)]
    return hidden, (w["%s/output/kernel" % prefix], w["%s/output/bias" % prefix])


def test_dcn_cfg3_full_batch_on_the_kaggle_arena(lib):
    set_seed(3)
    B, k, vocab = 65536, 16, synth.KAGGLE_VOCAB
    m = DCN(synth.criteo_feature_columns(vocab, k), [400, 400, 400], 1, "relu", 3, embed_dim=k, precision="tf32")
    m.build(None)
    assert m.embed_layer.arena.numel() * 4 > 2.1e9
    rng = np.random.default_rng(5)
    n = 13 + 26 * k
    # non-trivial cross and final weights (the defaults are N(0, 0.05^2): the cross term would be ~1e-3 of the logit)
    m.cross_layer.set_weights([rng.normal(0, 0.1, (n, 1)).astype(np.float32) for _ in range(6)])
    m.set_weights({"output_layer/kernel": rng.normal(0, 0.3, (n + 1, 1)).astype(np.float32),
                   "output_layer/bias": rng.normal(0, 0.1, (1,)).astype(np.float32)})
    dense, ids = synth.criteo_batch(B, vocab, seed=17)
    ids[-1, :] = np.asarray(vocab) - 1                               # the last row of every table
    got = npy(m((dense, ids)))[:, 0]
    assert np.isfinite(got).all()
    sub = np.unique(np.concatenate([np.arange(0, B, 127), [B - 1]]))
    tabs, sid = rows_touched(m.embed_layer.emb_layers, ids[sub])
    w = w64(m)
    hidden, out = tower(w, "DNNLayer_1", 3)
    want = O.dcn_model(synth.criteo_matrix(dense[sub], sid), tabs, [w["CrossLayer_2/w%d" % i] for i in range(3)],
                       [w["CrossLayer_2/b%d" % i] for i in range(3)], hidden, out,
                       (w["output_layer/kernel"], w["output_layer/bias"]))[:, 0]
    logit_close(got[sub], want, "DCN cfg 3")
    perm = np.random.default_rng(1).permutation(B)                   # size-independent property: tile placement is irrelevant
    got_p = npy(m((dense[perm], ids[perm])))[:, 0]
    np.testing.assert_array_equal(got[perm], got_p)


def test_dcn_cfg3_chained_launches_equal_stream_order(lib):
    """The three launches of DCN cfg 3 overlap at their edges under per-tile ready flags (DenseStack.chain_launches): twelve
    back-to-back forwards over alternating batches -- so that every launch runs next to its neighbours, also across calls --
    must give bit for bit what plain stream order gives, and a ragged batch (partial last tile) too."""
    set_seed(3)
    k, vocab = 16, [5000] * 26
    m = DCN(synth.criteo_feature_columns(vocab, k), [400, 400, 400], 1, "relu", 3, embed_dim=k, precision="tf32", check_ids=False)
    m.build(None)
    dev = m.device
    for B in (65536, 148 * 128 + 77):
        batches = []
        for i in range(3):
            d, ids = synth.criteo_batch(B, vocab, seed=40 + i)
            batches.append((torch.from_numpy(d).to(dev), torch.from_numpy(ids).to(dev)))
        stack = m.dense_layer.stack
        stack.chain_launches = False
        want = [m(b).clone() for b in batches]
        torch.cuda.synchronize()
        stack.chain_launches = True
        outs = [m(batches[i % 3]) for i in range(12)]
        torch.cuda.synchronize()
        for i, o in enumerate(outs):
            assert torch.equal(o, want[i % 3]), (B, i, float((o - want[i % 3]).abs().max()))
    # batch sizes changing from call to call (a ragged last batch, then a full one): the two workspaces must not move
    sizes = [40000, 65536, 148 * 128 + 77, 65536, 30000, 65536]
    seq = []
    for j, Bj in enumerate(sizes):
        d, ids = synth.criteo_batch(Bj, vocab, seed=60 + j)
        seq.append((torch.from_numpy(d).to(dev), torch.from_numpy(ids).to(dev)))
    stack.chain_launches = False
    want = [m(b).clone() for b in seq]
    torch.cuda.synchronize()
    stack.chain_launches = True
    outs = [m(b) for b in seq + seq]
    torch.cuda.synchronize()
    for i, o in enumerate(outs):
        assert torch.equal(o, want[i % len(seq)]), (i, sizes[i % len(seq)])


def test_din_cfg4_whole_model(lib):
    set_seed(4)
    B, T, K, v_item, v_user = 16384, 100, 64, 10_000_000, 192404
    m = DIN(synth.din_feature_columns(v_item=v_item, v_user=v_user, embed_dim=K), ["movies_seq"], (80, 40), (256, 128, 64),
            "prelu", "prelu", 0.0, precision="tf32")
    m.build(None)
    m.att_layer.build_for(T, K)
    rng = np.random.default_rng(6)
    dcat = m.dcat
    upd = {"bn/gamma": rng.uniform(0.5, 1.5, dcat), "bn/beta": rng.normal(0, 0.1, dcat), "bn/moving_mean": rng.normal(0, 0.1, dcat),
           "bn/moving_variance": rng.uniform(0.5, 1.5, dcat)}
    for i, u in enumerate((256, 128, 64)):
        upd["dense_%d/prelu_alpha" % i] = rng.uniform(0, 0.3, u)
    for i, h in enumerate((80, 40)):
        upd["Attention_0/dense_%d/prelu_alpha" % i] = rng.uniform(0, 0.3, (T, h))       # alpha is [T, h] (interaction.py:366)
        upd["Attention_0/dense_%d/bias" % i] = rng.normal(0, 0.1, h)
    m.set_weights({n: np.asarray(v, np.float32) for n, v in upd.items()})
    # keys of magnitude ~0.5 so that the scores spread and the softmax is not uniform
    m.embed_seq_layers[0].uniform_(-0.5, 0.5)
    m._bump()
    d = synth.din_batch(B, T, v_item=v_item, v_user=v_user, seed=21)
    d["movies_seq"][0, :] = 0                                                           # all padding
    d["movies_seq"][1, :] = np.arange(v_item - T, v_item)                               # full history, the table's last rows
    d["movie_id"][1, 0] = v_item - 1
    got = npy(m(d))[:, 0]
    assert np.isfinite(got).all()
    sub = np.unique(np.concatenate([np.arange(0, B, 61), [0, 1]]))
    hist = d["movies_seq"][sub].astype(np.int64)
    cand = d["movie_id"][sub].astype(np.int64)
    u, inv = np.unique(np.concatenate([hist.ravel(), cand.ravel(), [0]]), return_inverse=True)
    assert u[0] == 0                                                                    # padding id keeps index 0 (the mask tests != 0)
    seq_tab = npy(m.embed_seq_layers[0][torch.from_numpy(u).to(m.device)])
    inv = inv.reshape(-1)
    hist_c = inv[:hist.size].reshape(hist.shape)[:, :, None]
    cand_c = inv[hist.size:hist.size + cand.size].reshape(cand.shape)
    uu, uinv = np.unique(d["user_id"][sub].astype(np.int64), return_inverse=True)
    user_tab = npy(m.embed_sparse_layers[0][torch.from_numpy(uu).to(m.device)])
    w = w64(m)
    att = [(w["Attention_0/dense_%d/kernel" % i], w["Attention_0/dense_%d/bias" % i], w["Attention_0/dense_%d/prelu_alpha" % i])
           for i in range(2)]
    tw = [(w["dense_%d/kernel" % i], w["dense_%d/bias" % i], w["dense_%d/prelu_alpha" % i]) for i in range(3)]
    want = O.din_model(d["age"][sub].astype(np.float64), uinv.reshape(-1, 1), hist_c, cand_c, [user_tab], [seq_tab], att,
                       (w["Attention_0/out/kernel"], w["Attention_0/out/bias"]),
                       (w["bn/gamma"], w["bn/beta"], w["bn/moving_mean"], w["bn/moving_variance"]), tw,
                       (w["out/kernel"], w["out/bias"]))[:, 0]
    logit_close(got[sub], want, "DIN cfg 4")


def test_ffm_at_the_reference_shape(lib):
    """FFM exactly as DL/model/ffm.py:27-40 builds it on the bundled sample: feat_onehot_dims of train_1w.txt (sum 43 605) + 13
    dense = 43 618 features, 39 fields, k = 8 (v holds 13.6 M floats).  Gather-form float64 oracle on all rows, and the
    reference's literal dense one-hot form (interaction.py:141-163) on the first 48."""
    set_seed(5)
    B, k, vocab = 4096, 8, synth.SAMPLE_VOCAB
    fc = synth.criteo_feature_columns(vocab, k)
    m = FFM(fc, k)
    m.build((None, 39))
    rng = np.random.default_rng(8)
    n_feat = 13 + sum(vocab)
    assert n_feat == 43618
    w0 = rng.normal(0, 0.1, (1,)).astype(np.float32)
    w = rng.normal(0, 0.1, (n_feat, 1)).astype(np.float32)
    v = rng.normal(0, 0.05, (n_feat, 39, k)).astype(np.float32)
    m.ffm.set_weights([w0, w, v])
    dense, ids = synth.criteo_batch(B, vocab, seed=23)
    ids[0, :] = np.asarray(vocab) - 1
    ids[1, 4] = vocab[4] + 3                       # tf.one_hot gives an all-zero row for an out-of-range index: no error
    x = synth.criteo_matrix(dense, ids)
    got = npy(m(x.astype(np.float32)))[:, 0]
    lay = O.ffm_sparse(x, vocab, w0.astype(np.float64), w.astype(np.float64), v.astype(np.float64))[:, 0]
    lit = O.ffm_layer(x[:48], vocab, w0.astype(np.float64), w.astype(np.float64), v.astype(np.float64))[:, 0]
    np.testing.assert_allclose(lay[:48], lit, rtol=1e-10, atol=1e-12)
    # fp32 bar on the LAYER output (FFMLayer.call, no sigmoid): norm-wise <= 1e-5, and the element-wise figure beside it
    z = npy(m.ffm(x.astype(np.float32)))[:, 0]
    scale = np.maximum(np.abs(lay), np.sqrt(np.mean(lay ** 2)))
    err = np.abs(z - lay) / scale
    elementwise = np.abs(z - lay) / np.maximum(np.abs(lay), 1e-30)
    print("FFM reference shape: max norm-wise rel err %.3g, max element-wise rel err %.3g" % (err.max(), elementwise.max()))
    assert err.max() <= REL, err.max()
    np.testing.assert_allclose(got, O.sigmoid(lay), rtol=0, atol=2e-6)     # FFM.call = sigmoid of it (ffm.py:20-23)


@pytest.mark.parametrize("k", [16, 8])
def test_pnn_both_full_batch(lib, k):
    """k = 16: gather + product kernels + tower as separate launches; k = 8 (the reference's embed_dim): the one-launch front
    (csrc/pnn_fwd.cu) with the tower chained behind it under per-tile flags, and OuterProductLayer on the tensor pipe."""
    set_seed(6)
    B = 65536
    rng = np.random.default_rng(9)
    vocab = [int(x) for x in rng.integers(1000, 200000, 26)]
    m = PNN(synth.criteo_feature_columns(vocab, k), "both", [256, 128, 64], 1, embed_dim=k, precision="tf32")
    m.build(None)
    m.embed_layer.arena.uniform_(-0.5, 0.5)         # products of 0.05-scale rows would be ~1e-3: make them count
    m.embed_layer._bump()
    W = rng.normal(0, 0.2, (k, 325, k)).astype(np.float32)
    m.outer_product_layer.set_weights([W])
    dense, ids = synth.criteo_batch(B, vocab, seed=29)
    got = npy(m((dense, ids)))[:, 0]
    assert np.isfinite(got).all()
    sub = np.arange(0, B, 129)
    tabs, sid = rows_touched(m.embed_layer.emb_layers, ids[sub])
    w = w64(m)
    hidden, out = tower(w, "DNNLayer_0", 3)
    want = O.pnn_model(synth.criteo_matrix(dense[sub], sid), tabs, "both", hidden, out, W.astype(np.float64))[:, 0]
    err = np.max(np.abs(got[sub] - want))
    assert err <= LOGIT, err                        # PNN has no sigmoid (pnn.py:52-53): the output IS the logit
    # the product layers on their own, fp32 bar, full batch
    E = EmbedLayer(m.sparse_feature_columns, k=k, flatten=False)
    E.arena.copy_(m.embed_layer.arena)
    e3 = E(ids)
    inner = npy(m.inner_product_layer(e3))[sub]
    outer = npy(m.outer_product_layer(e3))[sub]
    e64 = O.embed_layer(sid, tabs, flatten=False)
    for name, g, wnt in (("inner", inner, O.inner_product_layer(e64)), ("outer", outer, O.outer_product_layer(e64, W.astype(np.float64)))):
        scale = np.maximum(np.abs(wnt), np.sqrt(np.mean(wnt ** 2)))
        r = np.max(np.abs(g - wnt) / scale)
        assert r <= REL, (name, r)


def test_arena_beyond_4_gib(lib):
    """One arena of 4.6 GB (> 2^32 bytes, > 2^30 floats): the last table alone is 70 M rows x 64 B.  Rows gathered from its end
    must be bit-exact and the whole-model kernel must read the same rows (64-bit address arithmetic everywhere)."""
    set_seed(7)
    k = 16
    vocab = [1000] * 25 + [70_000_000]
    fc = synth.criteo_feature_columns(vocab, k)
    m = DeepFM(fc, 10, 1e-4, 1e-4, [256, 128, 64], 1, "relu", embed_dim=k, precision="tf32")
    m.build(None)
    arena = m.embed_layer.arena
    assert arena.numel() * 4 > (1 << 32) and arena.numel() > (1 << 30)
    B = 8192
    dense, ids = synth.criteo_batch(B, vocab, seed=31)
    ids[:, 25] = vocab[25] - 1 - (np.arange(B) * 7919) % 1_000_000          # the last million rows of the arena
    ids[0, 25] = vocab[25] - 1
    rows = npy(m.embed_layer(ids))
    tabs, sid = rows_touched(m.embed_layer.emb_layers, ids)
    np.testing.assert_array_equal(rows, O.embed_layer(sid, tabs))           # R1: bit-exact
    # make the last field matter in the model output: a large FM weight on its columns
    n = 13 + 26 * k
    rng = np.random.default_rng(10)
    w1 = rng.normal(0, 0.05, (n, 1)).astype(np.float32)
    w1[13 + 25 * k:] = rng.normal(0, 5.0, (k, 1)).astype(np.float32)
    m.fm.set_weights([np.zeros(1, np.float32), w1, rng.normal(0, 0.05, (n, 10)).astype(np.float32)])
    got = npy(m((dense, ids)))[:, 0]
    sub = np.arange(0, B, 17)
    tabs, sid = rows_touched(m.embed_layer.emb_layers, ids[sub])
    w = w64(m)
    hidden, out = tower(w, "DNNLayer_2", 3)
    want = O.deepfm_model(synth.criteo_matrix(dense[sub], sid), tabs, (w["FMLayer_1/w0"], w["FMLayer_1/w1"], w["FMLayer_1/v"]),
                          hidden, out)[:, 0]
    logit_close(got[sub], want, "DeepFM on a 4.6 GB arena")
    # an id one past the end of the last table is reported, not wrapped
    bad = ids.copy()
    bad[5, 25] = vocab[25]
    with pytest.raises(IndexError):
        m((dense, bad))
