"""GPU parity for the whole-model kernel (ctrhot_ctr_fused_tf32, csrc/fused.cu): gather -> FM / cross -> tower ->
sigmoid in one persistent launch, against the float64 oracle.

* the FM / cross terms are formed on the tensor pipe by a split-TF32 product and must still meet the fp32 bar
  (<= 1e-5 relative, BASELINE north_star) -- tested with the tower zeroed so the term is observable on its own;
* the tower is TF32: logits <= 2e-3 absolute;
* ragged batches (B not a multiple of 128, B < 128, many tiles), every cluster size, k = 8 / 16 / 32.
"""
import numpy as np
import pytest
import torch

from oracle import ctr_oracle as O
from oracle import synth as S
from recommender_system_b200.model.criteo import AFM, DCN, NFM, AFMLayer, DeepFM, WideDeep
from recommender_system_b200.utils import synth

pytestmark = pytest.mark.gpu

LOGIT = 2e-3
REL = 1e-5


def npy(t):
    return t.detach().cpu().numpy().astype(np.float64)


def logit(p):
    return np.log(p) - np.log1p(-p)


def logit_close(got, want):
    """|logit(got) - logit(want)| <= 2e-3 plus what the model's own fp32 sigmoid costs on the way back: one ulp of p
    (2^-23 relative, 2^-24 absolute near 1) is 2^-23 / (p (1 - p)) in the logit -- 2.6e-3 at |logit| = 10."""
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    fin = (got > 0) & (got < 1)
    sat = ~fin                                            # fp32 sigmoid rounded to exactly 0 / 1: needs |logit| > 16.6
    assert np.all(np.abs(logit(want[sat])) > 16.0)
    tol = LOGIT + 2.0 ** -22 / (want[fin] * (1.0 - want[fin]))
    err = np.abs(logit(got[fin]) - logit(want[fin]))
    worst = int(np.argmax(err - tol)) if err.size else 0
    assert np.all(err <= tol), "logit err %.3g > tol %.3g at logit %.3g" % (err[worst], tol[worst], logit(want[fin][worst]))


def small_vocab(rng, F=26):
    return [int(v) for v in rng.integers(3, 400, F)]


def deepfm_case(B, k, kfm, hidden, seed, zero_tower=False):
    rng = np.random.default_rng(seed)
    vocab = small_vocab(rng)
    m = DeepFM(synth.criteo_feature_columns(vocab, k), kfm, 1e-4, 1e-4, hidden, 1, "relu", embed_dim=k, precision="tf32")
    m.build(None)
    n = 13 + 26 * k
    # bigger-than-default weights so that the FM term is O(1) and errors are visible
    m.fm.set_weights([S.f32(rng.normal(0, 0.3, (1,))), S.f32(rng.normal(0, 0.2, (n, 1))), S.f32(rng.normal(0, 0.2, (n, kfm)))])
    m.embed_layer.set_tables([S.f32(rng.uniform(-0.5, 0.5, (v, k))) for v in vocab])
    ws = []
    dims = [n] + list(hidden) + [1]
    for i in range(len(dims) - 1):
        lim = 0.0 if zero_tower else np.sqrt(6.0 / (dims[i] + dims[i + 1]))
        ws += [S.f32(rng.uniform(-lim, lim, (dims[i], dims[i + 1]))),
               S.f32(np.zeros(dims[i + 1]) if zero_tower else rng.normal(0, 0.1, (dims[i + 1],)))]
    m.dnn.set_weights(ws)
    dense, ids = synth.criteo_batch(B, vocab, seed=seed + 1)
    return m, vocab, dense, ids, ws


def deepfm_oracle(m, dense, ids, ws):
    tables = [npy(t) for t in m.embed_layer.emb_layers]
    fm = (npy(m.fm.w0), npy(m.fm.w1), npy(m.fm.v))
    pairs = [(ws[2 * i].astype(np.float64), ws[2 * i + 1].astype(np.float64)) for i in range(len(ws) // 2)]
    return O.deepfm_model(synth.criteo_matrix(dense, ids).astype(np.float64), tables, fm, pairs[:-1], pairs[-1])


@pytest.mark.parametrize("cluster", [1, 2, 4])
@pytest.mark.parametrize("B", [1, 100, 128, 1000, 5000])
def test_deepfm_whole_kernel(lib, B, cluster):
    m, vocab, dense, ids, ws = deepfm_case(B, 16, 10, [256, 128, 64], seed=B + cluster)
    m.cluster = cluster
    want = deepfm_oracle(m, dense, ids, ws)
    got = npy(m((dense, ids)))
    assert got.shape == (B, 1)
    logit_close(got, want)
    # the two-kernel path (front + tower) agrees
    m.whole = False
    got2 = npy(m((dense, ids)))
    logit_close(got2, want)


@pytest.mark.parametrize("k,kfm,hidden", [(8, 10, [32, 16, 8]), (16, 4, [256, 128]), (32, 15, [64]), (16, 10, [200, 80, 24])])
def test_deepfm_whole_kernel_shapes(lib, k, kfm, hidden):
    m, vocab, dense, ids, ws = deepfm_case(777, k, kfm, hidden, seed=k + kfm)
    want = deepfm_oracle(m, dense, ids, ws)
    got = npy(m((dense, ids)))
    logit_close(got, want)
    # float id columns (the reference's [B,39] matrix) give the same bits as the int32 fast path
    got2 = npy(m(synth.criteo_matrix(dense, ids).astype(np.float32)))
    np.testing.assert_array_equal(got, got2)


@pytest.mark.parametrize("k", [8, 16])
def test_fm_term_on_tensor_pipe_meets_fp32_bar(lib, k):
    """Tower zeroed: out = sigmoid(0.5 * FM(x)); the split-TF32 FM must be within 1e-5 relative of the fp64 oracle."""
    m, vocab, dense, ids, ws = deepfm_case(3000, k, 10, [64, 32], seed=11 + k, zero_tower=True)
    x = np.concatenate([dense.astype(np.float64), O.embed_layer(ids, [npy(t) for t in m.embed_layer.emb_layers])], 1)
    want = O.fm_layer(x, npy(m.fm.w0), npy(m.fm.w1), npy(m.fm.v))[:, 0]
    got = 2.0 * logit(npy(m((dense, ids))))[:, 0]
    scale = np.maximum(np.abs(want), np.sqrt(np.mean(want ** 2)))
    # sigmoid/logit round trip in fp32 costs ~1e-7 / (p (1-p)); keep samples away from saturation
    keep = np.abs(want) < 6.0
    err = np.max(np.abs(got - want)[keep] / scale[keep])
    assert keep.sum() > 1000 and err <= 2 * REL, err


def dcn_case(B, k, L, hidden, seed):
    rng = np.random.default_rng(seed)
    vocab = small_vocab(rng)
    m = DCN(synth.criteo_feature_columns(vocab, k), hidden, 1, "relu", L, embed_dim=k, precision="tf32")
    m.build(None)
    n = 13 + 26 * k
    m.embed_layer.set_tables([S.f32(rng.uniform(-0.5, 0.5, (v, k))) for v in vocab])
    cw = [S.f32(rng.normal(0, 0.1, (n, 1))) for _ in range(L)]
    cb = [S.f32(rng.normal(0, 0.1, (n, 1))) for _ in range(L)]
    m.cross_layer.set_weights(cw + cb)
    ws = []
    dims = [n] + list(hidden) + [1]
    for i in range(len(dims) - 1):
        lim = np.sqrt(6.0 / (dims[i] + dims[i + 1]))
        ws += [S.f32(rng.uniform(-lim, lim, (dims[i], dims[i + 1]))), S.f32(rng.normal(0, 0.1, (dims[i + 1],)))]
    m.dense_layer.set_weights(ws)
    fk, fb = S.f32(rng.normal(0, 0.1, (n + 1, 1))), S.f32(rng.normal(0, 0.1, (1,)))
    m.set_weights({"output_layer/kernel": fk, "output_layer/bias": fb})
    dense, ids = synth.criteo_batch(B, vocab, seed=seed + 1)
    tables = [npy(t) for t in m.embed_layer.emb_layers]
    pairs = [(ws[2 * i].astype(np.float64), ws[2 * i + 1].astype(np.float64)) for i in range(len(ws) // 2)]
    want = O.dcn_model(synth.criteo_matrix(dense, ids).astype(np.float64), tables, [w.astype(np.float64) for w in cw],
                       [b.astype(np.float64) for b in cb], pairs[:-1], pairs[-1], (fk.astype(np.float64), fb.astype(np.float64)))
    return m, dense, ids, want


@pytest.mark.parametrize("B,k,L,hidden", [(900, 16, 3, [256, 128, 64]), (130, 8, 6, [64, 32]), (2000, 16, 1, [128]),
                                          (1500, 16, 3, [400, 400, 400]), (129, 8, 2, [300, 72]), (20000, 16, 1, [416, 264, 40])])
def test_dcn_whole_kernel(lib, B, k, L, hidden):
    """The last three cases have towers too wide to chain through TMEM (cfg 3's 400-400-400 among them): the cross layers
    run as exact fp32 scalars on the CUDA cores of the first launch, the layers go through HBM one launch each."""
    m, dense, ids, want = dcn_case(B, k, L, hidden, seed=B + L)
    got = npy(m((dense, ids)))
    logit_close(got, want)
    m.whole = False
    got2 = npy(m((dense, ids)))
    logit_close(got2, want)


@pytest.mark.parametrize("B,k,hidden", [(900, 16, [256, 128, 64]), (129, 8, [64, 32]), (3000, 32, [128]), (65, 64, [40, 24])])
def test_widedeep_one_launch(lib, B, k, hidden):
    """Wide&Deep with the wide logit formed inside the persistent kernel (ctrhot_widedeep_fused_tf32: 4-byte weights fetched by
    the gather warps next to the embedding rows) against the float64 oracle in its sparse form, and against the two-launch path."""
    rng = np.random.default_rng(B + k)
    vocab = small_vocab(rng)
    D, F = 13, 26
    m = WideDeep(synth.criteo_feature_columns(vocab, k), hidden, 1, "relu", embed_dim=k, precision="tf32")
    m.build(None)
    tables = [S.f32(rng.uniform(-0.5, 0.5, (v, k))) for v in vocab]
    m.embed_layer.set_tables(tables)
    n_wide = 2 * D + sum(vocab)
    w0, w = S.f32(rng.normal(0, 0.1, (1,))), S.f32(rng.normal(0, 0.3, (n_wide, 1)))
    m.wide.set_weights([w0, w])
    ws = []
    dims = [F * k] + list(hidden) + [1]
    for i in range(len(dims) - 1):
        lim = np.sqrt(6.0 / (dims[i] + dims[i + 1]))
        ws += [S.f32(rng.uniform(-lim, lim, (dims[i], dims[i + 1]))), S.f32(rng.normal(0, 0.1, (dims[i + 1],)))]
    m.deep.set_weights(ws)
    dense, ids = synth.criteo_batch(B, vocab, seed=B)
    off = 2 * D + np.concatenate([[0], np.cumsum(vocab)])
    w64 = w[:, 0].astype(np.float64)
    wide = w0.astype(np.float64)[0] + dense.astype(np.float64) @ (w64[:D] + w64[D:2 * D])
    for f in range(F):
        wide = wide + w64[off[f] + ids[:, f]]
    pairs = [(ws[2 * i].astype(np.float64), ws[2 * i + 1].astype(np.float64)) for i in range(len(ws) // 2)]
    emb = O.embed_layer(ids, [t.astype(np.float64) for t in tables])
    deep = O.dnn_layer(emb, pairs[:-1], pairs[-1], "relu")[:, 0]
    want = O.sigmoid(0.5 * (wide + deep))
    got = npy(m((dense, ids)))[:, 0]
    logit_close(got, want)
    m.one_launch = False
    got2 = npy(m((dense, ids)))[:, 0]
    logit_close(got2, want)
    # an id past its field's vocabulary is reported, as on the two-launch path
    bad = ids.copy()
    bad[B // 2, 5] = vocab[5]
    m.one_launch = True
    with pytest.raises(IndexError):
        m((dense, bad))


@pytest.mark.parametrize("B,k,hidden,out_dim,prec", [(700, 8, [256, 128, 64], 1, "tf32"), (129, 16, [64, 32], 1, "tf32"),
                                                     (1, 8, [32], 1, "fp32"), (333, 32, [40, 24], 3, "fp32"),
                                                     (65536, 8, [256, 128, 64], 1, "tf32")])
def test_nfm(lib, B, k, hidden, out_dim, prec):
    """NFM: Bi-Interaction pooled in registers by the gather kernel, BatchNormalization folded into the
    tower's first layer, DNN output layer and final Dense(1) multiplied out -- against the float64 oracle of nfm.py:24-34."""
    rng = np.random.default_rng(B + k)
    vocab = small_vocab(rng)
    D, F = 13, 26
    m = NFM(synth.criteo_feature_columns(vocab, k), hidden, out_dim, "relu", embed_dim=k, precision=prec)
    m.build(None)
    tables = [S.f32(rng.uniform(-0.5, 0.5, (v, k))) for v in vocab]
    m.emb_layers.set_tables(tables)
    n = D + k
    bn = [S.f32(rng.normal(0, 0.2, n)), S.f32(rng.uniform(0.5, 1.5, n)), S.f32(rng.uniform(0.8, 1.2, n)), S.f32(rng.normal(0, 0.1, n))]
    ws = []
    dims = [n] + list(hidden) + [out_dim]
    for i in range(len(dims) - 1):
        lim = np.sqrt(6.0 / (dims[i] + dims[i + 1]))
        ws += [S.f32(rng.uniform(-lim, lim, (dims[i], dims[i + 1]))), S.f32(rng.normal(0, 0.1, (dims[i + 1],)))]
    m.dnn_layers.set_weights(ws)
    fk, fb = S.f32(rng.normal(0, 0.5, (out_dim, 1))), S.f32(rng.normal(0, 0.1, (1,)))
    m.set_weights({"bn/moving_mean": bn[0], "bn/moving_variance": bn[1], "bn/gamma": bn[2], "bn/beta": bn[3],
                   "output_layer/kernel": fk, "output_layer/bias": fb})
    dense, ids = synth.criteo_batch(B, vocab, seed=B)
    if B > 4096:                                            # full size: oracle on a strided subset
        sub = np.arange(0, B, 131)
    else:
        sub = np.arange(B)
    pairs = [(ws[2 * i].astype(np.float64), ws[2 * i + 1].astype(np.float64)) for i in range(len(ws) // 2)]
    want = O.nfm_model(synth.criteo_matrix(dense[sub], ids[sub]).astype(np.float64), [t.astype(np.float64) for t in tables],
                       [b.astype(np.float64) for b in bn], pairs[:-1], pairs[-1],
                       (fk.astype(np.float64), fb.astype(np.float64)))[:, 0]
    got = npy(m((dense, ids)))[:, 0][sub]
    if prec == "fp32":
        assert np.max(np.abs(got - want)) <= 5e-6, np.max(np.abs(got - want))
    else:
        logit_close(got, want)
    bad = ids.copy()
    bad[B // 2, 7] = vocab[7]
    with pytest.raises(IndexError):
        m((dense, bad))


@pytest.mark.parametrize("mode", ["avg", "max", "att"])
@pytest.mark.parametrize("B,F,k", [(1, 26, 8), (515, 26, 8), (300, 26, 16), (64, 2, 8), (100, 5, 32), (65536, 26, 8)])
def test_afm(lib, mode, B, F, k):
    """AFMLayer / AFM against the literal float64 restatement (all P pair products built, pooled,
    attention included -- whose softmax over a size-1 axis makes every weight 1), <= 1e-5 relative on the logit."""
    rng = np.random.default_rng(B + F + k)
    vocab = small_vocab(rng, F)
    fc = [[{"feat": "I%d" % i} for i in range(13)],
          [{"feat": "C%d" % i, "feat_onehot_dim": v, "embed_dim": k} for i, v in enumerate(vocab)]]
    tables = [S.f32(rng.uniform(-1.0, 1.0, (v, k))) for v in vocab]
    P = F * (F - 1) // 2
    att_w = (S.f32(rng.normal(0, 0.3, (k, P))), S.f32(rng.normal(0, 0.1, (P,))))
    att_h = (S.f32(rng.normal(0, 0.3, (P, 1))), S.f32(rng.normal(0, 0.1, (1,))))
    out = (S.f32(rng.normal(0, 0.5, (k, 1))), S.f32(rng.normal(0, 0.1, (1,))))
    dense = S.f32(rng.uniform(0, 1, (B, 13)))
    ids = np.stack([rng.integers(0, v, B) for v in vocab], axis=1).astype(np.int32)
    sub = np.arange(0, B, 257) if B > 4096 else np.arange(B)
    x64 = np.concatenate([dense[sub], ids[sub]], axis=1).astype(np.float64)
    f64 = lambda t: tuple(a.astype(np.float64) for a in t)                                    # noqa: E731
    for cls, ref in ((AFMLayer, O.afm_layer), (AFM, O.afm_model)):
        m = cls(fc, mode)
        m.build(None)
        m.embed_layer.set_tables(tables)
        w = {"output_layer/kernel": out[0], "output_layer/bias": out[1]}
        if mode == "att":
            w.update({"attention_w/kernel": att_w[0], "attention_w/bias": att_w[1],
                      "attention_h/kernel": att_h[0], "attention_h/bias": att_h[1]})
        m.set_weights(w)
        want = ref(x64, [t.astype(np.float64) for t in tables], mode, f64(att_w), f64(att_h), f64(out))[:, 0]
        got = npy(m((dense, ids)))[:, 0][sub]
        assert np.max(np.abs(got - want)) <= 1e-5 * np.maximum(np.abs(want), 0.25).max(), (cls.__name__, np.max(np.abs(got - want)))
    bad = ids.copy()
    bad[B // 2, F - 1] = vocab[F - 1]
    with pytest.raises(IndexError):
        m((dense, bad))


def test_whole_kernel_out_of_range_id(lib):
    m, vocab, dense, ids, ws = deepfm_case(300, 16, 10, [64, 32], seed=5)
    ids = ids.copy()
    ids[17, 3] = vocab[3]
    with pytest.raises(IndexError):
        m((dense, ids))


def test_whole_kernel_full_size_properties(lib):
    """B = 65536 (BASELINE config 2a size): size-independent checks -- permutation equivariance over samples and
    agreement with the fp32 CUDA-core path on a strided subset."""
    m, vocab, dense, ids, ws = deepfm_case(65536, 16, 10, [256, 128, 64], seed=3)
    got = npy(m((dense, ids)))[:, 0]
    perm = np.random.default_rng(0).permutation(65536)
    got_p = npy(m((dense[perm], ids[perm])))[:, 0]
    np.testing.assert_array_equal(got[perm], got_p)          # row results do not depend on tile / CTA placement
    sub = np.arange(0, 65536, 97)
    want = deepfm_oracle(m, dense[sub], ids[sub], ws)[:, 0]
    logit_close(got[sub], want)


def test_widedeep_onehot_narrower_than_vocab(lib):
    """WideDeep built with onehot_width[f] < feat_onehot_dim[f] (get_dummies keeps only the labels it saw,
    dataset.py:53-58).  An id in [width, vocab) has an embedding row but no one-hot column: it must add 0 to the wide logit on
    BOTH paths (the one-launch kernel used to add the next field's weight, and read past the end of w for the last field)."""
    rng = np.random.default_rng(77)
    B, k, D, F, hidden = 700, 16, 13, 26, [64, 32]
    vocab = small_vocab(rng)
    widths = [max(1, v - int(rng.integers(0, 3))) for v in vocab]
    widths[F - 1] = max(1, vocab[F - 1] - 2)                       # the last field: the old code read beyond the weight vector
    m = WideDeep(synth.criteo_feature_columns(vocab, k), hidden, 1, "relu", embed_dim=k, precision="tf32", onehot_width=widths)
    m.build(None)
    tables = [S.f32(rng.uniform(-0.5, 0.5, (v, k))) for v in vocab]
    m.embed_layer.set_tables(tables)
    n_wide = 2 * D + sum(widths)
    w0, w = S.f32(rng.normal(0, 0.1, (1,))), S.f32(rng.normal(0, 1.0, (n_wide, 1)))
    m.wide.set_weights([w0, w])
    ws = []
    dims = [F * k] + hidden + [1]
    for i in range(len(dims) - 1):
        lim = np.sqrt(6.0 / (dims[i] + dims[i + 1]))
        ws += [S.f32(rng.uniform(-lim, lim, (dims[i], dims[i + 1]))), S.f32(rng.normal(0, 0.1, (dims[i + 1],)))]
    m.deep.set_weights(ws)
    dense, ids = synth.criteo_batch(B, vocab, seed=78)
    for f in range(F):                                             # make sure every field has ids in [width, vocab)
        if widths[f] < vocab[f]:
            ids[f::F, f][: 8] = vocab[f] - 1
    assert any((ids[:, f] >= widths[f]).any() for f in range(F))
    off = 2 * D + np.concatenate([[0], np.cumsum(widths)])
    w64_ = w[:, 0].astype(np.float64)
    wide = w0.astype(np.float64)[0] + dense.astype(np.float64) @ (w64_[:D] + w64_[D:2 * D])
    for f in range(F):
        inside = ids[:, f] < widths[f]
        wide = wide + np.where(inside, w64_[off[f] + np.minimum(ids[:, f], widths[f] - 1)], 0.0)
    pairs = [(ws[2 * i].astype(np.float64), ws[2 * i + 1].astype(np.float64)) for i in range(len(ws) // 2)]
    deep = O.dnn_layer(O.embed_layer(ids, [t.astype(np.float64) for t in tables]), pairs[:-1], pairs[-1], "relu")[:, 0]
    want = O.sigmoid(0.5 * (wide + deep))
    for one in (True, False):
        m.one_launch = one
        logit_close(npy(m((dense, ids)))[:, 0], want)


@pytest.mark.parametrize("packed", [False, True])
def test_overlapped_launches_give_the_same_results(lib, packed):
    """CTRHOT_LAUNCH_OVERLAP (programmatic dependent launch): 12 back-to-back forward calls on independent batches, each
    allowed to start while the previous one drains its last tiles, must produce exactly what ordinary launches produce."""
    from recommender_system_b200.utils.transport import PackedRows, RowLayout
    vocab = synth.SAMPLE_VOCAB
    m = DeepFM(synth.criteo_feature_columns(vocab, 16), 10, 1e-4, 1e-4, [256, 128, 64], 1, "relu", embed_dim=16, check_ids=False)
    m.build(None)
    lay = RowLayout(vocab, 13)
    B = 148 * 128 * 2 + 128 * 40 + 17                      # an uneven last wave: some SMs finish a tile earlier than others
    batches = [synth.criteo_batch(B, vocab, seed=100 + i) for i in range(4)]
    if packed:
        ins = [PackedRows(lay, torch.from_numpy(lay.pack(d, i)).cuda()) for d, i in batches]
    else:
        ins = [(torch.from_numpy(d).cuda(), torch.from_numpy(i).cuda()) for d, i in batches]
    want = [m(x).clone() for x in ins]
    torch.cuda.synchronize()
    m.overlap_launches = True
    outs = [m(ins[j % 4]) for j in range(12)]
    torch.cuda.synchronize()
    for j, o in enumerate(outs):
        assert torch.equal(o, want[j % 4]), j
    # an ordinary launch after overlapped ones is ordered as usual
    m.overlap_launches = False
    assert torch.equal(m(ins[0]), want[0])
