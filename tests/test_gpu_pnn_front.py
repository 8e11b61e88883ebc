"""PNN front (csrc/pnn_fwd.cu: gather -> [flat | inner | outer] in one launch) through the C-ABI against the float64 oracle's
InnerProductLayer / OuterProductLayer (DL/layer/interaction.py:170-183, 201-215) on the same tables, ids and W.
Bars: flat columns bit-exact, products <= 1e-5 relative (norm-wise, as in test_gpu_parity.close_rel)."""
import numpy as np
import pytest
import torch

from oracle import ctr_oracle as O
from oracle import synth as S
from recommender_system_b200.model.criteo import PNN
from recommender_system_b200.utils import synth
from test_gpu_parity import close_rel, npy

pytestmark = pytest.mark.gpu


def make(F, vocab_hi, seed, mode="both", hidden=(64, 32), prec="tf32"):
    rng = np.random.default_rng(seed)
    vocab = [int(v) for v in rng.integers(3, vocab_hi, F)]
    m = PNN(synth.criteo_feature_columns(vocab, 8), mode, list(hidden), 1, embed_dim=8, precision=prec)
    m.build(None)
    tabs = [S.f32(rng.normal(0, 0.3, (v, 8))) for v in vocab]
    m.embed_layer.set_tables(tabs)
    P = F * (F - 1) // 2
    W = S.f32(rng.normal(0, 0.1, (8, P, 8)))
    m.outer_product_layer.set_weights([W])
    return m, vocab, tabs, W, rng


def front(m, ids, do_inner, ld=None):
    E = m.embed_layer
    P = E.F * (E.F - 1) // 2
    width = E.F * 8 + P * (2 if do_inner else 1)
    ld = ld or (width + 31) // 32 * 32
    buf = torch.full((ids.shape[0], ld), 7.0, dtype=torch.float32, device="cuda")
    m.outer_product_layer.front(E, ids, do_inner, buf, ld)
    torch.cuda.synchronize()
    return buf, width


@pytest.mark.parametrize("B,F", [(1, 26), (127, 26), (128, 26), (300, 26), (5000, 26), (148 * 128 * 2 + 77, 26),
                                 (333, 2), (200, 3), (513, 13), (260, 27), (129, 28)])
def test_front_against_oracle(lib, B, F):
    m, vocab, tabs, W, rng = make(F, 5000, B + F)
    ids = np.stack([rng.integers(0, v, B) for v in vocab], axis=1).astype(np.int32)
    buf, width = front(m, torch.from_numpy(ids).cuda(), True)
    got = buf.cpu().numpy()
    e = np.stack([tabs[f][ids[:, f]] for f in range(F)], axis=1)                     # [B, F, 8]
    P = F * (F - 1) // 2
    assert np.array_equal(got[:, :F * 8], e.reshape(B, F * 8))                        # gathered rows: bit-exact
    assert np.all(got[:, width:] == 0.0)
    close_rel(got[:, F * 8:F * 8 + P], O.inner_product_layer(e.astype(np.float64)))
    close_rel(got[:, F * 8 + P:width], O.outer_product_layer(e.astype(np.float64), W.astype(np.float64)))


def test_front_outer_only_float_ids_and_wide_rows(lib):
    B, F = 777, 26
    m, vocab, tabs, W, rng = make(F, 300, 5, mode="outer")
    ids = np.stack([rng.integers(0, v, B) for v in vocab], axis=1)
    buf, width = front(m, torch.from_numpy(ids.astype(np.float32)).cuda(), False, ld=640)
    got = buf.cpu().numpy()
    e = np.stack([tabs[f][ids[:, f]] for f in range(F)], axis=1)
    assert width == F * 8 + 325 and np.array_equal(got[:, :F * 8], e.reshape(B, -1)) and np.all(got[:, width:] == 0.0)
    close_rel(got[:, F * 8:width], O.outer_product_layer(e.astype(np.float64), W.astype(np.float64)))


def test_front_large_magnitudes_keep_fp32_accuracy(lib):
    """The three-term TF32 split must hold 1e-5 when rows and W are O(1)-O(10), not only at initialiser scale."""
    B, F = 1000, 26
    m, vocab, tabs, W, rng = make(F, 100, 11)
    tabs = [S.f32(rng.uniform(-8, 8, t.shape)) for t in tabs]
    W = S.f32(rng.uniform(-3, 3, W.shape))
    m.embed_layer.set_tables(tabs)
    m.outer_product_layer.set_weights([W])                                            # re-pack on the version bump
    ids = np.stack([rng.integers(0, v, B) for v in vocab], axis=1).astype(np.int32)
    buf, width = front(m, torch.from_numpy(ids).cuda(), True)
    got = buf.cpu().numpy()
    e = np.stack([tabs[f][ids[:, f]] for f in range(F)], axis=1)
    close_rel(got[:, F * 8 + 325:width], O.outer_product_layer(e.astype(np.float64), W.astype(np.float64)))
    close_rel(got[:, F * 8:F * 8 + 325], O.inner_product_layer(e.astype(np.float64)))


def test_front_bad_ids_raise(lib):
    m, vocab, tabs, W, rng = make(26, 50, 3)
    ids = np.stack([rng.integers(0, v, 40) for v in vocab], axis=1).astype(np.int32)
    ids[7, 3] = vocab[3]
    with pytest.raises(Exception):
        front(m, torch.from_numpy(ids).cuda(), True)
    m.embed_layer.check_ids() if False else None


@pytest.mark.parametrize("mode", ["both", "outer"])
def test_pnn_model_fused_front_equals_layer_path_and_oracle(lib, mode):
    """PNN.call (DL/model/pnn.py:28-53) at a Criteo-shaped batch: one-launch front + tower vs the separate
    gather / inner / outer launches (same tower) and vs the float64 oracle's pnn_model."""
    B, F = 4099, 26
    m, vocab, tabs, W, rng = make(F, 2000, 21, mode=mode, hidden=(256, 128, 64))
    ids = np.stack([rng.integers(0, v, B) for v in vocab], axis=1).astype(np.float32)
    x = np.concatenate([S.f32(rng.random((B, 13))), ids], axis=1)
    got = npy(m(x))
    m.fused_front = False
    ref = npy(m(x))
    assert np.max(np.abs(got - ref)) <= 2e-3
    tw = m.dnn_layer.get_weights()
    hidden = [(tw[2 * i].astype(np.float64), tw[2 * i + 1].astype(np.float64)) for i in range(3)]
    out = (tw[6].astype(np.float64), tw[7].astype(np.float64))
    want = O.pnn_model(x.astype(np.float64), [t.astype(np.float64) for t in tabs], mode, hidden, out,
                       outer_w=W.astype(np.float64))
    assert np.max(np.abs(got - want)) <= 2e-3, np.max(np.abs(got - want))


def test_pnn_chained_launches_equal_stream_order(lib):
    """Front -> tower under per-tile flags with overlapping launches (two alternating product buffers, a new epoch per call)
    gives bit-identical logits to plain stream order, call after call on different batches."""
    B, F = 148 * 128 * 2 + 77, 26
    m, vocab, tabs, W, rng = make(F, 3000, 31, mode="both", hidden=(256, 128, 64))
    xs = []
    for i in range(3):
        ids = np.stack([rng.integers(0, v, B) for v in vocab], axis=1).astype(np.float32)
        xs.append(torch.from_numpy(np.concatenate([S.f32(rng.random((B, 13))), ids], axis=1)).cuda())
    m.chain_launches = False
    want = [m(x).clone() for x in xs]
    m.chain_launches = True
    got = [m(xs[i % 3]).clone() for i in range(7)]
    torch.cuda.synchronize()
    assert m._chain is not None and m._chain[2] == 7
    for i in range(7):
        assert torch.equal(got[i], want[i % 3]), i
    # a larger batch after a smaller one re-sizes the two product buffers and the flags; a smaller one reuses them
    for Bn in (B + 5000, B - 3000):
        ids = np.stack([rng.integers(0, v, Bn) for v in vocab], axis=1).astype(np.float32)
        x = torch.from_numpy(np.concatenate([S.f32(rng.random((Bn, 13))), ids], axis=1)).cuda()
        a = m(x).clone()
        m.chain_launches = False
        b = m(x).clone()
        m.chain_launches = True
        assert a.shape == (Bn, 1) and torch.equal(a, b), Bn


@pytest.mark.parametrize("B,F", [(128, 26), (4099, 26), (1000, 5), (300, 28)])
def test_outer_product_layer_k8_on_the_tensor_pipe(lib, B, F):
    """OuterProductLayer.call (DL/layer/interaction.py:201-215) on a gathered [B, F, 8] tensor: the tcgen05 path against the float64
    oracle and against the CUDA-core kernel; written into a strided view without touching the columns around it."""
    from recommender_system_b200.layer import OuterProductLayer
    rng = np.random.default_rng(B + F)
    e = S.f32(rng.uniform(-0.5, 0.5, (B, F, 8)))
    P = F * (F - 1) // 2
    W = S.f32(rng.normal(0, 0.1, (8, P, 8)))
    op = OuterProductLayer()
    op.build(e.shape)
    op.set_weights([W])
    want = O.outer_product_layer(e.astype(np.float64), W.astype(np.float64))
    ed = torch.from_numpy(e).cuda()
    big = torch.full((B, P + 40), 7.0, dtype=torch.float32, device="cuda")
    l0 = __import__("recommender_system_b200")._lib.LAUNCHES[0]
    op(ed, out=big[:, 8:8 + P])
    close_rel(big[:, 8:8 + P].cpu().numpy(), want)
    assert torch.all(big[:, :8] == 7.0) and torch.all(big[:, 8 + P:] == 7.0)
    op.tensor_pipe = False
    close_rel(npy(op(ed)), want)
