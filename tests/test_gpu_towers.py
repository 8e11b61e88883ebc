"""GPU parity for SURVEY 8(f) rank 4, the remaining towers: ResLayer / DeepCrossing, mmoe_layer / tower_layer / MMOE, and the
DeepCTR-lineage DNN / LocalActivationUnit of layer/core.py -- against the REFERENCE'S OWN outputs on the tiny golden cases
(tests/golden/ref_layers.npz, its unmodified classes run under the numpy shim) and against the float64 oracle at larger, ragged
shapes.  fp32 CUDA-core path <= 2e-5; TF32 tensor-core path <= 2e-3 absolute (BASELINE north_star)."""
import numpy as np
import pytest
import torch

from oracle import ctr_oracle as O
from oracle import synth as S
from recommender_system_b200.layer import DNN, LocalActivationUnit, ResLayer, mmoe_layer
from recommender_system_b200.model.deepCrossing import DeepCrossing
from recommender_system_b200.model.mmoe import MMOE
from recommender_system_b200.utils import synth

pytestmark = pytest.mark.gpu
TOL = {"fp32": 2e-5, "tf32": 2e-3}


def npy(t):
    return t.detach().cpu().numpy().astype(np.float64)


def stack_w(a, prefix):
    out, i = [], 0
    while "%s_k%d" % (prefix, i) in a:
        out += [a["%s_k%d" % (prefix, i)], a["%s_b%d" % (prefix, i)]]
        i += 1
    return out


def err(got, want):
    return float(np.max(np.abs(np.asarray(got, np.float64) - np.asarray(want, np.float64))))


@pytest.mark.parametrize("prec", ["fp32", "tf32"])
def test_res_layer(golden, lib, prec):
    a, want = S.load_case(golden, "res")
    L = ResLayer([24, 16], precision=prec)
    L.build(a["x"].shape)
    L.set_weights(stack_w(a, "res"))
    assert err(npy(L(a["x"])), want) <= TOL[prec]


@pytest.mark.parametrize("prec", ["fp32", "tf32"])
def test_deepcrossing(golden, lib, prec):
    a, want = S.load_case(golden, "deepcrossing")
    fc = synth.criteo_feature_columns([int(v) for v in a["vocab"]], 8)
    m = DeepCrossing(fc, 32, [24, 16], 2, embed_dim=8, precision=prec)
    m.build(None)
    m.embed_layer.set_tables(S._tables_of(a))
    for r, lyr in enumerate(m.res_layer):
        lyr.set_weights(stack_w(a, "res%d" % r))
    m.set_weights({"output_layer/kernel": a["final_k"], "output_layer/bias": a["final_b"]})
    assert err(npy(m(a["inputs"])), want) <= TOL[prec] / 4            # probabilities: the logit tolerance shrinks by p(1-p)


@pytest.mark.parametrize("prec", ["fp32", "tf32"])
def test_mmoe(golden, lib, prec):
    def weights(a):
        T, gm, gb = S._mmoe_parts(a)
        w = {"expert_matrix": a["expert_matrix"], "expert_bias": a["expert_bias"]}
        for t in range(T):
            w["gate_matrix%d" % t], w["gate_bias%d" % t] = gm[t], gb[t]
        return T, w
    a, want_layer = S.load_case(golden, "mmoe_layer")          # the two cases are drawn with different seeds
    n_in, H, E = a["expert_matrix"].shape
    T, w = weights(a)
    L = mmoe_layer(H, E, T, precision=prec)
    L.build(a["x"].shape)
    L.set_weights(w)
    outs = L(a["x"])
    assert len(outs) == T and tuple(outs[0].shape) == (a["x"].shape[0], H)
    assert err(np.stack([npy(o) for o in outs]), want_layer) <= TOL[prec]
    a, want = S.load_case(golden, "mmoe")
    T, w = weights(a)
    m = MMOE(H, E, T, [20, 10], 1, precision=prec)
    m.build(a["x"].shape)
    m.mmoe_layer.set_weights(w)
    for t, tw in enumerate(m.tower_layer):
        tw.set_weights(stack_w(a, "tower%d" % t))
    assert err(np.stack([npy(o) for o in m(a["x"])]), want) <= TOL[prec]


@pytest.mark.parametrize("prec", ["fp32", "tf32"])
@pytest.mark.parametrize("case,act,use_bn,out_act", [("deepctr_dnn_relu", "relu", False, None),
                                                     ("deepctr_dnn_bn_sigmoid", "sigmoid", True, "tanh")])
def test_deepctr_dnn(golden, lib, prec, case, act, use_bn, out_act):
    a, want = S.load_case(golden, case)
    ks, bs, bn = S._dnn_parts(a)
    L = DNN([k.shape[1] for k in ks], act, use_bn=use_bn, output_activation=out_act, precision=prec)
    L.build(a["x"].shape)
    w = {}
    for i in range(len(ks)):
        w["kernel%d" % i], w["bias%d" % i] = ks[i], bs[i]
        if use_bn:
            for nm, v in zip(("moving_mean", "moving_variance", "gamma", "beta"), bn[i]):
                w["bn_%d/%s" % (i, nm)] = v
    L.set_weights(w)
    got = npy(L(a["x"]))
    assert got.shape == want.shape and err(got, want) <= TOL[prec]


@pytest.mark.parametrize("prec", ["fp32", "tf32"])
@pytest.mark.parametrize("act", ["sigmoid", "dice"])
def test_local_activation_unit(golden, lib, prec, act):
    a, want = S.load_case(golden, "lau_" + act)
    ks, bs, dice = S._lau_parts(a)
    L = LocalActivationUnit([k.shape[1] for k in ks], act, precision=prec)
    L.build([a["query"].shape, a["keys"].shape])
    w = {"kernel": a["out_k"], "bias": a["out_b"]}
    for i in range(len(ks)):
        w["DNN_0/kernel%d" % i], w["DNN_0/bias%d" % i] = ks[i], bs[i]
        if act == "dice":
            for nm, v in zip(("alpha", "moving_mean", "moving_variance"), dice[i]):
                w["DNN_0/dice_%d/%s" % (i, nm)] = v
    L.set_weights(w)
    got = npy(L([a["query"], a["keys"]]))
    assert got.shape == want.shape and err(got, want) <= TOL[prec]


@pytest.mark.parametrize("prec", ["fp32", "tf32"])
@pytest.mark.parametrize("case", ["seqpool_softmax", "seqpool_plain", "seqpool_score"])
def test_attention_sequence_pooling(golden, lib, prec, case):
    """AttentionSequencePoolingLayer (DL/layer/sequence.py:180-276) against the reference's own output: with the keras mask
    (the path that runs as written) and with keys_length (the path the reference means; it raises there as written)."""
    from recommender_system_b200.layer.sequence import AttentionSequencePoolingLayer
    a, want = S.load_case(golden, case)
    ks, bs, _ = S._lau_parts(a)
    wn, ret = case != "seqpool_plain", case == "seqpool_score"
    w = {"LocalActivationUnit_0/kernel": a["out_k"], "LocalActivationUnit_0/bias": a["out_b"]}
    for i in range(len(ks)):
        w["LocalActivationUnit_0/DNN_0/kernel%d" % i], w["LocalActivationUnit_0/DNN_0/bias%d" % i] = ks[i], bs[i]
    for masking in (True, False):
        L = AttentionSequencePoolingLayer([k.shape[1] for k in ks], "sigmoid", weight_normalization=wn, return_score=ret,
                                          supports_masking=masking, precision=prec)
        L.build([a["query"].shape, a["keys"].shape, a["keys_length"].shape])
        L.set_weights(w)
        if masking:
            got = npy(L([a["query"], a["keys"]], mask=[None, a["key_mask"]]))
        else:
            got = npy(L([a["query"], a["keys"], a["keys_length"]]))
        assert got.shape == want.shape and err(got, want) <= TOL[prec], (masking, err(got, want))


def test_attention_sequence_pooling_at_din_shape(lib):
    """B = 3 000, T = 100, E = 64, 80-40 against the float64 oracle, ragged lengths incl. empty and full sequences."""
    from recommender_system_b200.layer.sequence import AttentionSequencePoolingLayer
    rng = np.random.default_rng(11)
    B, T, E = 3000, 100, 64
    q = np.asarray(rng.normal(0, 0.5, (B, 1, E)), np.float32)
    keys = np.asarray(rng.normal(0, 0.5, (B, T, E)), np.float32)
    lens = rng.integers(0, T + 1, size=(B, 1)).astype(np.int32)
    lens[0], lens[1] = 0, T
    L = AttentionSequencePoolingLayer((80, 40), "sigmoid", weight_normalization=True, precision="fp32")
    L.build([q.shape, keys.shape, lens.shape])
    got = npy(L([q, keys, lens]))
    wt = {n: npy(t) for n, t in L.named_weights().items()}
    ks = [wt["LocalActivationUnit_0/DNN_0/kernel%d" % i] for i in range(2)]
    bs = [wt["LocalActivationUnit_0/DNN_0/bias%d" % i] for i in range(2)]
    want = O.attention_sequence_pooling(q.astype(np.float64), keys.astype(np.float64), np.arange(T)[None, :] < lens, ks, bs,
                                        wt["LocalActivationUnit_0/kernel"], wt["LocalActivationUnit_0/bias"], "sigmoid", True, False)
    assert np.max(np.abs(got - want)) <= 2e-5
    np.testing.assert_allclose(got[0, 0], keys[0].astype(np.float64).mean(0), rtol=1e-5, atol=1e-6)     # empty: uniform over T


def test_augru_against_the_reference_cell(golden, lib):
    """AUGRU (DL/layer/sequence.py:293-395) against the reference's own AUGRUCell.call stepped by its rnn_augru (both unmodified
    under the shim): the sequence of hidden states and the last one."""
    from recommender_system_b200.layer.sequence import AUGRU
    a, want = S.load_case(golden, "augru")
    for seq in (True, False):
        L = AUGRU(a["rkernel"].shape[0], return_sequence=seq)
        L.build(a["x"].shape)
        L.set_weights({"kernel": a["kernel"], "recurrent_kernel": a["rkernel"]})
        got = npy(L(a["x"], att_score=a["att"]))
        assert err(got, want if seq else want[:, -1]) <= 2e-6, err(got, want if seq else want[:, -1])


@pytest.mark.parametrize("B,T,E,H,hard", [(1, 1, 4, 4, 0), (77, 9, 5, 7, 1), (3000, 100, 64, 64, 0), (1025, 50, 18, 36, 1)])
def test_augru_shapes(lib, B, T, E, H, hard):
    """Ragged batches, odd widths, DIEN's shape (T = 100, 64 -> 64), both recurrent activations, a non-zero initial state."""
    from recommender_system_b200.layer.sequence import AUGRU
    rng = np.random.default_rng(B + T)
    x = np.asarray(rng.normal(0, 0.7, (B, T, E)), np.float32)
    att = np.asarray(rng.random((B, T, 1)), np.float32)
    h0 = np.asarray(rng.normal(0, 0.3, (B, H)), np.float32)
    L = AUGRU(H, return_sequence=True, recurrent_activation="hard_sigmoid" if hard else "sigmoid", as_written=False)
    L.build(x.shape)
    got = npy(L(x, att_score=att, initial_state=h0))
    k, rk = npy(L.kernel), npy(L.recurrent_kernel)
    want = O.augru(x.astype(np.float64), att.astype(np.float64), k, rk, h0.astype(np.float64),
                   "hard_sigmoid" if hard else "sigmoid")[1]
    assert np.max(np.abs(got - want)) <= 2e-5, np.max(np.abs(got - want))


def test_towers_at_size(lib):
    """Ragged, larger shapes against the float64 oracle: DeepCrossing at the reference script's own settings (hidden 256-256,
    4 residual units, deepCrossing.py:37-39) on 5 000 samples; MMoE with 8 experts / 3 tasks on 10 001 rows; the local
    activation unit at DIN's shape (T = 100, K = 64, 80-40)."""
    rng = np.random.default_rng(3)
    # DeepCrossing
    vocab = [int(v) for v in rng.integers(3, 400, 26)]
    m = DeepCrossing(synth.criteo_feature_columns(vocab, 8), 32, [256, 256], 4, embed_dim=8)
    m.build(None)
    dense, ids = synth.criteo_batch(5000, vocab, seed=4)
    w = {n: npy(t) for n, t in m.named_weights().items()}
    tabs = [npy(t) for t in m.embed_layer.emb_layers]
    res = []
    for r in range(4):
        p = "ResLayer_%d" % (r + 1)
        res.append(([(w["%s/dense_%d/kernel" % (p, i)], w["%s/dense_%d/bias" % (p, i)]) for i in range(2)],
                    (w[p + "/output/kernel"], w[p + "/output/bias"])))
    want = O.deepcrossing_model(synth.criteo_matrix(dense, ids), tabs, res, (w["output_layer/kernel"], w["output_layer/bias"]))
    assert err(npy(m((dense, ids))), want) <= 2e-3
    # MMoE
    x = rng.normal(0, 1, (10001, 37)).astype(np.float32)
    L = mmoe_layer(16, 8, 3)
    L.build(x.shape)
    w = {n: npy(t) for n, t in L.named_weights().items()}
    want = O.mmoe_layer(x.astype(np.float64), w["expert_matrix"], w["expert_bias"], [w["gate_matrix%d" % t] for t in range(3)],
                        [w["gate_bias%d" % t] for t in range(3)])
    got = L(x)
    assert err(np.stack([npy(o) for o in got]), np.stack(want)) <= 2e-3
    # LocalActivationUnit
    B, T, K = 257, 100, 64
    q = rng.normal(0, 0.5, (B, 1, K)).astype(np.float32)
    k = rng.normal(0, 0.5, (B, T, K)).astype(np.float32)
    U = LocalActivationUnit((80, 40), "sigmoid")
    U.build([q.shape, k.shape])
    w = {n: npy(t) for n, t in U.named_weights().items()}
    want = O.local_activation_unit(q.astype(np.float64), k.astype(np.float64), [w["DNN_0/kernel0"], w["DNN_0/kernel1"]],
                                   [w["DNN_0/bias0"], w["DNN_0/bias1"]], w["kernel"], w["bias"], "sigmoid")
    assert err(npy(U([q, k])), want) <= 2e-3
