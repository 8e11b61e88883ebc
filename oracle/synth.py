"""Seeded tiny cases for every hot-path layer/model -- TEST INFRASTRUCTURE ONLY.

Each case is ``(make(rng) -> dict[str, ndarray], run(arrays) -> ndarray)``.
``make`` draws float32-representable values (stored as float64 so the O64
oracle and the fp32 CUDA path see the *same* numbers); ``run`` calls
``oracle.ctr_oracle`` on them.  ``tests/golden/make_golden.py`` freezes
``make``'s arrays and ``run``'s float64 output into ``layers_tiny.npz``.

Shapes follow SURVEY.md 8(c): B=7, F=26, D=13, T=5 with one all-padding row
and one full row, non-zero PReLU/Dice alphas so every branch is exercised.
"""
from __future__ import annotations

import numpy as np

from . import ctr_oracle as O

B, F, D, T = 7, 26, 13, 5


def f32(a):
    return np.asarray(a, dtype=np.float32).astype(np.float64)


def _vocab(rng):
    return rng.integers(3, 50, size=F)


def _ids(rng, vocab, n=B):
    return np.stack([rng.integers(0, v, size=n) for v in vocab], axis=1).astype(np.float64)


def _tables(rng, vocab, k):
    return [f32(rng.uniform(-0.05, 0.05, size=(v, k))) for v in vocab]


def _dense_stack(rng, dims, prefix, out):
    for i in range(len(dims) - 1):
        lim = np.sqrt(6.0 / (dims[i] + dims[i + 1]))
        out["%s_k%d" % (prefix, i)] = f32(rng.uniform(-lim, lim, size=(dims[i], dims[i + 1])))
        out["%s_b%d" % (prefix, i)] = f32(rng.normal(0, 0.1, size=(dims[i + 1],)))


def _get_stack(a, prefix):
    pairs = []
    i = 0
    while "%s_k%d" % (prefix, i) in a:
        pairs.append((a["%s_k%d" % (prefix, i)], a["%s_b%d" % (prefix, i)]))
        i += 1
    return pairs[:-1], pairs[-1]


def _tables_of(a, prefix="table"):
    out = []
    i = 0
    while "%s%d" % (prefix, i) in a:
        out.append(a["%s%d" % (prefix, i)])
        i += 1
    return out


# ------------------------------------------------------------------ layers
def make_embed(rng, k=8):
    vocab = _vocab(rng)
    a = {"vocab": vocab.astype(np.int64), "ids": _ids(rng, vocab)}
    for i, t in enumerate(_tables(rng, vocab, k)):
        a["table%d" % i] = t
    return a


def run_embed(a):
    return O.embed_layer(a["ids"], _tables_of(a))


def make_wide(rng):
    x = f32(rng.uniform(0, 1, size=(B, 40))) * (rng.uniform(size=(B, 40)) < 0.4)
    return {"x": x, "w0": f32(rng.normal(0, 0.05, (1,))), "w": f32(rng.normal(0, 0.05, (40, 1)))}


def run_wide(a):
    return O.wide_layer(a["x"], a["w0"], a["w"])


def make_fm(rng, n=D + F * 8, k=10):
    return {"x": f32(rng.uniform(-0.5, 1.0, size=(B, n))), "w0": f32(rng.normal(0, 0.05, (1,))),
            "w1": f32(rng.normal(0, 0.05, (n, 1))), "v": f32(rng.normal(0, 0.05, (n, k)))}


def run_fm(a):
    return O.fm_layer(a["x"], a["w0"], a["w1"], a["v"])


def make_dnn(rng, n=D + F * 8):
    a = {"x": f32(rng.uniform(-0.5, 1.0, size=(B, n)))}
    _dense_stack(rng, [n, 32, 16, 1], "dnn", a)
    return a


def run_dnn(a):
    hidden, out = _get_stack(a, "dnn")
    return O.dnn_layer(a["x"], hidden, out, "relu")


def make_cross(rng, n=D + F * 8, L=3):
    a = {"x": f32(rng.uniform(-0.5, 1.0, size=(B, n)))}
    for i in range(L):
        a["w%d" % i] = f32(rng.normal(0, 0.05, (n, 1)))
        a["b%d" % i] = f32(rng.normal(0, 0.05, (n, 1)))
    return a


def run_cross(a):
    L = sum(1 for k in a if k.startswith("w"))
    return O.cross_layer(a["x"], [a["w%d" % i] for i in range(L)], [a["b%d" % i] for i in range(L)])


def make_ffm(rng, k=4):
    vocab = _vocab(rng)
    n_feat = int(vocab.sum()) + D
    ids = _ids(rng, vocab)
    ids[0, 3] = vocab[3] + 5          # out of range -> tf.one_hot gives a zero row
    inputs = np.concatenate([f32(rng.uniform(0, 1, (B, D))), ids], axis=1)
    return {"inputs": inputs, "vocab": vocab.astype(np.int64), "w0": f32(rng.normal(0, 0.05, (1,))),
            "w": f32(rng.normal(0, 0.05, (n_feat, 1))), "v": f32(rng.normal(0, 0.05, (n_feat, D + F, k)))}


def run_ffm(a):
    return O.ffm_layer(a["inputs"], [int(v) for v in a["vocab"]], a["w0"], a["w"], a["v"])


def make_pnn_layers(rng, k=8):
    P = F * (F - 1) // 2
    return {"e": f32(rng.uniform(-0.5, 0.5, (B, F, k))), "W": f32(rng.normal(0, 0.05, (k, P, k)))}


def run_inner(a):
    return O.inner_product_layer(a["e"])


def run_outer(a):
    return O.outer_product_layer(a["e"], a["W"])


def make_dice(rng, h=12):
    return {"x": f32(rng.normal(0, 1, (B, T, h))), "alphas": f32(rng.uniform(0, 0.3, (h,))),
            "mean": f32(rng.normal(0, 0.1, (h,))), "var": f32(rng.uniform(0.5, 1.5, (h,)))}


def run_dice(a):
    return O.dice(a["x"], a["alphas"], a["mean"], a["var"])


def _hist_mask(rng, n, t):
    lens = rng.integers(0, t + 1, size=n)
    lens[0] = 0          # all-padding row
    lens[1] = t          # full row
    return (np.arange(t)[None, :] < lens[:, None]).astype(np.float64)


def make_att(rng, K=8, hidden=(16, 8)):
    a = {"query": f32(rng.normal(0, 0.5, (B, K))), "key": f32(rng.normal(0, 0.5, (B, T, K))),
         "mask": _hist_mask(rng, B, T)}
    dims = [4 * K] + list(hidden)
    for i in range(len(hidden)):
        lim = np.sqrt(6.0 / (dims[i] + dims[i + 1]))
        a["att_k%d" % i] = f32(rng.uniform(-lim, lim, (dims[i], dims[i + 1])))
        a["att_b%d" % i] = f32(rng.normal(0, 0.1, (dims[i + 1],)))
        a["att_a%d" % i] = f32(rng.uniform(0, 0.3, (T, dims[i + 1])))
    a["att_ko"] = f32(rng.uniform(-0.5, 0.5, (dims[-1], 1)))
    a["att_bo"] = f32(rng.normal(0, 0.1, (1,)))
    return a


def _att_params(a):
    dense = []
    i = 0
    while "att_k%d" % i in a:
        dense.append((a["att_k%d" % i], a["att_b%d" % i], a["att_a%d" % i]))
        i += 1
    return dense, (a["att_ko"], a["att_bo"])


def run_att(a):
    dense, out = _att_params(a)
    return O.din_attention(a["query"], a["key"], a["key"], a["mask"], dense, out, "prelu")


def make_att_dice(rng, K=8):
    a = {"query": f32(rng.normal(0, 0.5, (B, K))), "key": f32(rng.normal(0, 0.5, (B, T, K))),
         "mask": _hist_mask(rng, B, T)}
    for i in range(2):
        a["dice_a%d" % i] = f32(rng.uniform(0, 0.3, (4 * K,)))
        a["dice_m%d" % i] = f32(rng.normal(0, 0.1, (4 * K,)))
        a["dice_v%d" % i] = f32(rng.uniform(0.5, 1.5, (4 * K,)))
    a["att_ko"] = f32(rng.uniform(-0.5, 0.5, (4 * K, 1)))
    a["att_bo"] = f32(rng.normal(0, 0.1, (1,)))
    return a


def run_att_dice(a):
    dp = [(a["dice_a%d" % i], a["dice_m%d" % i], a["dice_v%d" % i]) for i in range(2)]
    return O.din_attention(a["query"], a["key"], a["key"], a["mask"], None, (a["att_ko"], a["att_bo"]),
                           "dice", dp)


# ------------------------------------------------------------------ models
def _criteo_inputs(rng, vocab):
    return np.concatenate([f32(rng.uniform(0, 1, (B, D))), _ids(rng, vocab)], axis=1)


def make_deepfm(rng, k=8, k_fm=10):
    vocab = _vocab(rng)
    d = D + F * k
    a = {"inputs": _criteo_inputs(rng, vocab), "vocab": vocab.astype(np.int64),
         "w0": f32(rng.normal(0, 0.05, (1,))), "w1": f32(rng.normal(0, 0.05, (d, 1))),
         "v": f32(rng.normal(0, 0.05, (d, k_fm)))}
    for i, t in enumerate(_tables(rng, vocab, k)):
        a["table%d" % i] = t
    _dense_stack(rng, [d, 32, 16, 8, 1], "dnn", a)
    return a


def run_deepfm(a):
    hidden, out = _get_stack(a, "dnn")
    return O.deepfm_model(a["inputs"], _tables_of(a), (a["w0"], a["w1"], a["v"]), hidden, out)


def make_widedeep(rng, k=8):
    vocab = _vocab(rng)
    ids = _ids(rng, vocab)
    dense = f32(rng.uniform(0, 1, (B, D)))
    # the reference's one-hot block repeats the 13 dense columns (dataset.py:53-58)
    raw = np.concatenate([dense, ids], axis=1)
    onehot = O.onehot_expand(raw, [int(v) for v in vocab], D, np.float64)
    n_wide = D + onehot.shape[1]
    a = {"inputs": np.concatenate([dense, ids, onehot], axis=1), "vocab": vocab.astype(np.int64),
         "w0": f32(rng.normal(0, 0.05, (1,))), "w": f32(rng.normal(0, 0.05, (n_wide, 1)))}
    for i, t in enumerate(_tables(rng, vocab, k)):
        a["table%d" % i] = t
    _dense_stack(rng, [F * k, 32, 16, 8, 1], "dnn", a)
    return a


def run_widedeep(a):
    hidden, out = _get_stack(a, "dnn")
    return O.widedeep_model(a["inputs"], _tables_of(a), (a["w0"], a["w"]), hidden, out)


def make_dcn(rng, k=8, L=3):
    vocab = _vocab(rng)
    d = D + F * k
    a = {"inputs": _criteo_inputs(rng, vocab), "vocab": vocab.astype(np.int64)}
    for i, t in enumerate(_tables(rng, vocab, k)):
        a["table%d" % i] = t
    for i in range(L):
        a["cw%d" % i] = f32(rng.normal(0, 0.05, (d, 1)))
        a["cb%d" % i] = f32(rng.normal(0, 0.05, (d, 1)))
    _dense_stack(rng, [d, 32, 16, 8, 1], "dnn", a)
    a["final_k"] = f32(rng.uniform(-0.1, 0.1, (d + 1, 1)))
    a["final_b"] = f32(rng.normal(0, 0.1, (1,)))
    return a


def run_dcn(a):
    hidden, out = _get_stack(a, "dnn")
    L = sum(1 for k in a if k.startswith("cw"))
    return O.dcn_model(a["inputs"], _tables_of(a), [a["cw%d" % i] for i in range(L)],
                       [a["cb%d" % i] for i in range(L)], hidden, out, (a["final_k"], a["final_b"]))


def make_pnn(rng, k=8):
    vocab = _vocab(rng)
    P = F * (F - 1) // 2
    a = {"inputs": _criteo_inputs(rng, vocab), "vocab": vocab.astype(np.int64),
         "W": f32(rng.normal(0, 0.05, (k, P, k)))}
    for i, t in enumerate(_tables(rng, vocab, k)):
        a["table%d" % i] = t
    _dense_stack(rng, [F * k + 2 * P, 32, 16, 1], "dnn", a)
    return a


def run_pnn(a):
    hidden, out = _get_stack(a, "dnn")
    return O.pnn_model(a["inputs"], _tables_of(a), "both", hidden, out, a["W"])


def make_din(rng, K=8, n_other=2, n_dense=2, hidden=(16, 8), tower=(16, 8)):
    v_item, v_other = 40, [11, 7]
    mask = _hist_mask(rng, B, T)
    hist = (rng.integers(1, v_item, size=(B, T)) * mask).astype(np.float64)[:, :, None]
    a = {"dense": f32(rng.uniform(0, 1, (B, n_dense))),
         "other_ids": np.stack([rng.integers(0, v, B) for v in v_other], 1).astype(np.float64),
         "hist_ids": hist, "cand_ids": rng.integers(1, v_item, (B, 1)).astype(np.float64),
         "seq_table0": f32(rng.uniform(-0.5, 0.5, (v_item, K)))}
    for i, v in enumerate(v_other):
        a["other_table%d" % i] = f32(rng.uniform(-0.5, 0.5, (v, K)))
    att = make_att(rng, K, hidden)
    for key in att:
        if key.startswith("att_"):
            a[key] = att[key]
    dcat = 2 * K + n_other * K + n_dense
    a["bn_gamma"] = f32(rng.uniform(0.5, 1.5, (dcat,)))
    a["bn_beta"] = f32(rng.normal(0, 0.1, (dcat,)))
    a["bn_mean"] = f32(rng.normal(0, 0.1, (dcat,)))
    a["bn_var"] = f32(rng.uniform(0.5, 1.5, (dcat,)))
    dims = [dcat] + list(tower)
    for i in range(len(tower)):
        lim = np.sqrt(6.0 / (dims[i] + dims[i + 1]))
        a["tw_k%d" % i] = f32(rng.uniform(-lim, lim, (dims[i], dims[i + 1])))
        a["tw_b%d" % i] = f32(rng.normal(0, 0.1, (dims[i + 1],)))
        a["tw_a%d" % i] = f32(rng.uniform(0, 0.3, (dims[i + 1],)))
    a["out_k"] = f32(rng.uniform(-0.5, 0.5, (dims[-1], 1)))
    a["out_b"] = f32(rng.normal(0, 0.1, (1,)))
    return a


def run_din(a):
    dense, att_out = _att_params(a)
    tower = []
    i = 0
    while "tw_k%d" % i in a:
        tower.append((a["tw_k%d" % i], a["tw_b%d" % i], a["tw_a%d" % i]))
        i += 1
    return O.din_model(a["dense"], a["other_ids"], a["hist_ids"], a["cand_ids"],
                       _tables_of(a, "other_table"), _tables_of(a, "seq_table"), dense, att_out,
                       (a["bn_gamma"], a["bn_beta"], a["bn_mean"], a["bn_var"]), tower,
                       (a["out_k"], a["out_b"]))


def make_nfm(rng, k=8):
    vocab = _vocab(rng)
    n = D + k
    a = {"inputs": _criteo_inputs(rng, vocab), "vocab": vocab.astype(np.int64),
         "bn_mean": f32(rng.normal(0, 0.1, (n,))), "bn_var": f32(rng.uniform(0.5, 1.5, (n,))),
         "bn_gamma": f32(rng.uniform(0.5, 1.5, (n,))), "bn_beta": f32(rng.normal(0, 0.1, (n,)))}
    for i, t in enumerate(_tables(rng, vocab, k)):
        a["table%d" % i] = t
    _dense_stack(rng, [n, 32, 16, 3], "dnn", a)
    a["final_k"] = f32(rng.uniform(-0.5, 0.5, (3, 1)))
    a["final_b"] = f32(rng.normal(0, 0.1, (1,)))
    return a


def run_nfm(a):
    hidden, out = _get_stack(a, "dnn")
    return O.nfm_model(a["inputs"], _tables_of(a), (a["bn_mean"], a["bn_var"], a["bn_gamma"], a["bn_beta"]),
                       hidden, out, (a["final_k"], a["final_b"]))


def make_afm(rng, k=8):
    vocab = _vocab(rng)
    P = F * (F - 1) // 2
    a = {"inputs": _criteo_inputs(rng, vocab), "vocab": vocab.astype(np.int64),
         "att_wk": f32(rng.uniform(-0.3, 0.3, (k, P))), "att_wb": f32(rng.normal(0, 0.1, (P,))),
         "att_hk": f32(rng.uniform(-0.3, 0.3, (P, 1))), "att_hb": f32(rng.normal(0, 0.1, (1,))),
         "out_k": f32(rng.uniform(-0.5, 0.5, (k, 1))), "out_b": f32(rng.normal(0, 0.1, (1,)))}
    for i, t in enumerate(_tables(rng, vocab, k)):
        a["table%d" % i] = f32(t * 10.0)          # pair products of 0.05-scale rows would all be ~1e-3
    return a


def _run_afm(mode):
    def run(a):
        return O.afm_model(a["inputs"], _tables_of(a), mode, (a["att_wk"], a["att_wb"]), (a["att_hk"], a["att_hb"]),
                           (a["out_k"], a["out_b"]))
    return run



# ------------------------------------------------------------------ the remaining towers
def make_res(rng, n=D + F * 8):
    a = {"x": f32(rng.uniform(-0.5, 1.0, size=(B, n)))}
    _dense_stack(rng, [n, 24, 16, n], "res", a)
    return a


def run_res(a):
    hidden, out = _get_stack(a, "res")
    return O.res_layer(a["x"], hidden, out)


def make_deepcrossing(rng, k=8):
    vocab = _vocab(rng)
    d = D + F * k
    a = {"inputs": _criteo_inputs(rng, vocab), "vocab": vocab.astype(np.int64)}
    for i, t in enumerate(_tables(rng, vocab, k)):
        a["table%d" % i] = f32(t * 10.0)
    for r in range(2):
        _dense_stack(rng, [d, 24, 16, d], "res%d" % r, a)
    a["final_k"] = f32(rng.uniform(-0.2, 0.2, (d, 1)))
    a["final_b"] = f32(rng.normal(0, 0.1, (1,)))
    return a


def run_deepcrossing(a):
    res = [_get_stack(a, "res%d" % r) for r in range(2)]
    return O.deepcrossing_model(a["inputs"], _tables_of(a), res, (a["final_k"], a["final_b"]))


def make_mmoe(rng, n_in=12, hidden=10, E=3, T=3):
    a = {"x": f32(rng.uniform(-1.0, 2.0, (B, n_in))), "expert_matrix": f32(rng.normal(0, 0.3, (n_in, hidden, E))),
         "expert_bias": f32(rng.normal(0, 0.2, (hidden, E)))}
    for t in range(T):
        a["gate_matrix%d" % t] = f32(rng.normal(0, 0.5, (n_in, E)))
        a["gate_bias%d" % t] = f32(rng.normal(0, 0.3, (E,)))
        _dense_stack(rng, [hidden, 20, 10, 1], "tower%d" % t, a)
    return a


def _mmoe_parts(a):
    T = sum(1 for k in a if k.startswith("gate_matrix"))
    return T, [a["gate_matrix%d" % t] for t in range(T)], [a["gate_bias%d" % t] for t in range(T)]


def run_mmoe_layer(a):
    T, gm, gb = _mmoe_parts(a)
    return np.stack(O.mmoe_layer(a["x"], a["expert_matrix"], a["expert_bias"], gm, gb), axis=0)       # [T, B, hidden]


def run_mmoe(a):
    T, gm, gb = _mmoe_parts(a)
    towers = [_get_stack(a, "tower%d" % t) for t in range(T)]
    return np.stack(O.mmoe_model(a["x"], a["expert_matrix"], a["expert_bias"], gm, gb, towers), axis=0)  # [T, B, 1]


def make_deepctr_dnn(rng, n=20, hidden=(16, 8)):
    a = {"x": f32(rng.normal(0, 1.0, (B, T, n)))}
    dims = [n] + list(hidden)
    for i in range(len(hidden)):
        lim = np.sqrt(6.0 / (dims[i] + dims[i + 1]))
        a["kernel%d" % i] = f32(rng.uniform(-lim, lim, (dims[i], dims[i + 1])))
        a["bias%d" % i] = f32(rng.normal(0, 0.1, (dims[i + 1],)))
        a["bn_mean%d" % i] = f32(rng.normal(0, 0.1, (dims[i + 1],)))
        a["bn_var%d" % i] = f32(rng.uniform(0.5, 1.5, (dims[i + 1],)))
        a["bn_gamma%d" % i] = f32(rng.uniform(0.5, 1.5, (dims[i + 1],)))
        a["bn_beta%d" % i] = f32(rng.normal(0, 0.1, (dims[i + 1],)))
    return a


def _dnn_parts(a):
    n = sum(1 for k in a if k.startswith("kernel"))
    return ([a["kernel%d" % i] for i in range(n)], [a["bias%d" % i] for i in range(n)],
            [(a["bn_mean%d" % i], a["bn_var%d" % i], a["bn_gamma%d" % i], a["bn_beta%d" % i]) for i in range(n)])


def run_deepctr_dnn_relu(a):
    k, b, _ = _dnn_parts(a)
    return O.deepctr_dnn(a["x"], k, b, "relu")


def run_deepctr_dnn_bn_sigmoid(a):
    k, b, bn = _dnn_parts(a)
    return O.deepctr_dnn(a["x"], k, b, "sigmoid", bn=bn, output_activation="tanh")


def make_lau(rng, K=8, hidden=(16, 8)):
    a = {"query": f32(rng.normal(0, 0.5, (B, 1, K))), "keys": f32(rng.normal(0, 0.5, (B, T, K)))}
    dims = [4 * K] + list(hidden)
    for i in range(len(hidden)):
        lim = np.sqrt(6.0 / (dims[i] + dims[i + 1]))
        a["kernel%d" % i] = f32(rng.uniform(-lim, lim, (dims[i], dims[i + 1])))
        a["bias%d" % i] = f32(rng.normal(0, 0.1, (dims[i + 1],)))
        a["dice_a%d" % i] = f32(rng.uniform(0, 0.3, (dims[i + 1],)))
        a["dice_m%d" % i] = f32(rng.normal(0, 0.1, (dims[i + 1],)))
        a["dice_v%d" % i] = f32(rng.uniform(0.5, 1.5, (dims[i + 1],)))
    a["out_k"] = f32(rng.uniform(-0.5, 0.5, (dims[-1], 1)))
    a["out_b"] = f32(rng.normal(0, 0.1, (1,)))
    return a


def _lau_parts(a):
    n = sum(1 for k in a if k.startswith("kernel"))
    return ([a["kernel%d" % i] for i in range(n)], [a["bias%d" % i] for i in range(n)],
            [(a["dice_a%d" % i], a["dice_m%d" % i], a["dice_v%d" % i]) for i in range(n)])


def run_lau_sigmoid(a):
    k, b, _ = _lau_parts(a)
    return O.local_activation_unit(a["query"], a["keys"], k, b, a["out_k"], a["out_b"], "sigmoid")


def run_lau_dice(a):
    k, b, dice = _lau_parts(a)
    return O.local_activation_unit(a["query"], a["keys"], k, b, a["out_k"], a["out_b"], "dice", dice=dice)


def make_seqpool(rng, K=8, hidden=(16, 8)):
    a = make_lau(rng, K, hidden)
    lens = rng.integers(0, T + 1, size=B)
    lens[0], lens[1 % B] = 0, T                                # an empty and a full sequence
    a["key_mask"] = (np.arange(T)[None, :] < lens[:, None])
    a["keys_length"] = lens.reshape(B, 1).astype(np.int32)
    return a


def _run_seqpool(wn, ret, act="sigmoid"):
    def run(a):
        k, b, dice = _lau_parts(a)
        return O.attention_sequence_pooling(a["query"], a["keys"], a["key_mask"], k, b, a["out_k"], a["out_b"], act,
                                            weight_normalization=wn, return_score=ret, dice=dice if act == "dice" else None)
    return run


def make_augru(rng, E=6, H=8):
    lim = np.sqrt(6.0 / (E + 3 * H))
    return {"x": f32(rng.normal(0, 0.7, (B, T, E))), "att": f32(rng.random((B, T, 1))),
            "kernel": f32(rng.uniform(-lim, lim, (E, 3 * H))), "rkernel": f32(rng.uniform(-0.6, 0.6, (H, 3 * H)))}


def run_augru(a):
    return O.augru(a["x"], a["att"], a["kernel"], a["rkernel"])[1]


CASES = {
    "embed": (make_embed, run_embed),
    "wide": (make_wide, run_wide),
    "fm": (make_fm, run_fm),
    "dnn": (make_dnn, run_dnn),
    "cross": (make_cross, run_cross),
    "ffm": (make_ffm, run_ffm),
    "inner": (make_pnn_layers, run_inner),
    "outer": (make_pnn_layers, run_outer),
    "dice": (make_dice, run_dice),
    "att": (make_att, run_att),
    "att_dice": (make_att_dice, run_att_dice),
    "deepfm": (make_deepfm, run_deepfm),
    "widedeep": (make_widedeep, run_widedeep),
    "dcn": (make_dcn, run_dcn),
    "pnn": (make_pnn, run_pnn),
    "din": (make_din, run_din),
    "nfm": (make_nfm, run_nfm),
    "afm_avg": (make_afm, _run_afm("avg")),
    "afm_max": (make_afm, _run_afm("max")),
    "afm_att": (make_afm, _run_afm("att")),
    "res": (make_res, run_res),
    "deepcrossing": (make_deepcrossing, run_deepcrossing),
    "mmoe_layer": (make_mmoe, run_mmoe_layer),
    "mmoe": (make_mmoe, run_mmoe),
    "deepctr_dnn_relu": (make_deepctr_dnn, run_deepctr_dnn_relu),
    "deepctr_dnn_bn_sigmoid": (make_deepctr_dnn, run_deepctr_dnn_bn_sigmoid),
    "lau_sigmoid": (make_lau, run_lau_sigmoid),
    "lau_dice": (make_lau, run_lau_dice),
    "seqpool_softmax": (make_seqpool, _run_seqpool(True, False)),
    "seqpool_plain": (make_seqpool, _run_seqpool(False, False)),
    "seqpool_score": (make_seqpool, _run_seqpool(True, True)),
    "augru": (make_augru, run_augru),
}


def tiny_case(seed=7):
    """name -> (run, (arrays,), arrays) for make_golden.py."""
    out = {}
    for i, (name, (make, run)) in enumerate(CASES.items()):
        arrays = make(np.random.default_rng(seed * 1000 + i))
        out[name] = (run, (arrays,), arrays)
    return out


def load_case(npz, name):
    """Rebuild one case's arrays dict from the frozen npz."""
    pre = name + "/"
    a = {k[len(pre):]: npz[k] for k in npz.files if k.startswith(pre) and k != pre + "expected"}
    return a, npz[pre + "expected"]


def hash_table_value(seed, f, grow, col):
    """NumPy twin of csrc/api.cu::shard_hash_value (synthetic config-5 tables): U(-0.05, 0.05) float32 from a
    counter-based hash of (field, global row, column).  Arguments broadcast."""
    M = np.uint64(0xFFFFFFFFFFFFFFFF)
    with np.errstate(over="ignore"):
        z = (np.uint64(seed) + np.uint64(0x9E3779B97F4A7C15) * (np.asarray(f, dtype=np.uint64) + np.uint64(1))
             + np.uint64(0xBF58476D1CE4E5B9) * np.asarray(grow, dtype=np.uint64)
             + np.uint64(0x94D049BB133111EB) * (np.asarray(col, dtype=np.uint64) + np.uint64(1))) & M
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & M
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & M
        z = z ^ (z >> np.uint64(31))
    u = (z >> np.uint64(40)).astype(np.float32) * np.float32(1.0 / 16777216.0)
    return ((u - np.float32(0.5)) * np.float32(0.1)).astype(np.float32)
