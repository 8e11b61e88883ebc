"""Golden outputs of the REFERENCE'S OWN call() bodies -> tests/golden/ref_layers.npz.  TEST INFRASTRUCTURE ONLY.

    python tests/golden/make_ref_golden.py          (build container only: needs /root/reference)

The reference's files are imported UNMODIFIED from /root/reference/algorithm/deep_learning (they are pure Python);
``tests/golden/tf_shim`` supplies numpy stand-ins for the TensorFlow/Keras symbols they import.  For every case of
``oracle/synth.py`` (same seeded inputs and weights that ``layers_tiny.npz`` freezes) the script builds the
reference's Layer / Model object, lets it create its weights through its own ``build()``, overwrites them with the
case's arrays (checking that every shape the reference created matches), runs its ``call()`` and stores the result:

    <case>/expected       float64 weights and inputs  -> what tests/test_oracle.py compares the float64 oracle with
    <case>/expected_f32   float32 weights and inputs  -> the reference's own precision
    <case>_model/expected sigmoid model wrappers FM / FFM (R11)

What this pins: the reference's op order, axes, pair order, one-hot layout, mask fill, weight shapes (PReLU alpha
[T,h] in the attention unit, Dice's BatchNormalization without gamma/beta, cross weights [d,1]) -- everything its own
Python decides.  What it cannot pin: TensorFlow's kernels themselves; the shim restates their documented semantics.

Two reference models raise as written: ``PNN.call`` indexes ``embed.shape[2]`` and ``NFM.call`` pools over
axis 1, but ``EmbedLayer.call`` returns the flattened [B, F*k].  The script first asserts that they do raise, then
runs them with ``EmbedLayer`` subclassed to skip its final reshape (the [B, F, k] view both files' comments name);
nothing else of the two files is touched.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
REF_DL = "/root/reference/algorithm/deep_learning"

import tf_shim  # noqa: E402

tf_shim.install()
sys.path.insert(0, REF_DL)

from layer import interaction as RI  # noqa: E402  (the reference's layer/interaction.py)
from layer import core as RC  # noqa: E402         (the reference's layer/core.py)
from model.fm import FM  # noqa: E402
from model.ffm import FFM  # noqa: E402
from model.deepFM import DeepFM  # noqa: E402
from model.wideDeep import WideDeep  # noqa: E402
from model.dcn import DCN  # noqa: E402
from model.pnn import PNN  # noqa: E402
from model.din import DIN  # noqa: E402
from model.nfm import NFM  # noqa: E402
from model.afm import AFM  # noqa: E402
from model.deepCrossing import DeepCrossing  # noqa: E402
from model.mmoe import MMOE  # noqa: E402
from tf_shim import ops as SO  # noqa: E402
import model.pnn as ref_pnn  # noqa: E402
import model.nfm as ref_nfm  # noqa: E402

from oracle import synth  # noqa: E402
from tf_shim import layers as SL  # noqa: E402


def put(obj, attr, value, index=None):
    """Overwrite a weight the reference's build() created, insisting on the shape it chose."""
    cur = getattr(obj, attr)
    if index is not None:
        assert tuple(cur[index].shape) == tuple(value.shape), (attr, index, cur[index].shape, value.shape)
        cur[index] = value
    else:
        assert tuple(cur.shape) == tuple(value.shape), (type(obj).__name__, attr, cur.shape, value.shape)
        setattr(obj, attr, value)


def put_dense(layer, kernel, bias):
    put(layer, "kernel", kernel)
    put(layer, "bias", bias)


def put_dnn(dnn, a, prefix="dnn"):
    hidden, out = synth._get_stack(a, prefix)
    assert len(dnn.hidden_layer) == len(hidden)
    for lyr, (k, b) in zip(dnn.hidden_layer, hidden):
        put_dense(lyr, k, b)
    put_dense(dnn.output_layer, *out)


def put_tables(layers, tables):
    assert len(layers) == len(tables)
    for lyr, t in zip(layers, tables):
        put(lyr, "embeddings", t)


def fcols(vocab, k=8, n_dense=13):
    """The reference's feature_columns structure (utils/dataset.py:19-26, 69-75)."""
    return [[{"feat": "I%d" % (i + 1)} for i in range(n_dense)],
            [{"feat": "C%d" % (i + 1), "feat_onehot_dim": int(v), "embed_dim": k} for i, v in enumerate(vocab)]]


def two_pass(obj, x, inject):
    obj(x)                 # the reference's own build() creates the weights
    inject(obj)
    return np.asarray(obj(x))


# ------------------------------------------------------------------ one runner per case
def r_embed(a):
    k = a["table0"].shape[1]
    lyr = RC.EmbedLayer(fcols(a["vocab"], k)[1], k)
    return two_pass(lyr, a["ids"], lambda o: put_tables(o.emb_layers, synth._tables_of(a)))


def r_wide(a):
    def inj(o):
        put(o, "w0", a["w0"]); put(o, "w", a["w"])
    return two_pass(RI.WideLayer(), a["x"], inj)


def _inj_fm(a):
    def inj(o):
        put(o, "w0", a["w0"]); put(o, "w1", a["w1"]); put(o, "v", a["v"])
    return inj


def r_fm(a):
    return two_pass(RI.FMLayer(a["v"].shape[1]), a["x"], _inj_fm(a))


def r_fm_model(a):
    m = FM(a["v"].shape[1])
    return two_pass(m, a["x"], lambda o: _inj_fm(a)(o.fm))


def r_dnn(a):
    hidden, _ = synth._get_stack(a, "dnn")
    lyr = RI.DNNLayer([k.shape[1] for k, _ in hidden], 1, "relu")
    return two_pass(lyr, a["x"], lambda o: put_dnn(o, a))


def r_cross(a):
    L = sum(1 for k in a if k.startswith("w"))

    def inj(o):
        for i in range(L):
            put(o, "cross_weight", a["w%d" % i], i)
            put(o, "cross_bias", a["b%d" % i], i)
    return two_pass(RI.CrossLayer(L), a["x"], inj)


def _inj_ffm(a):
    def inj(o):
        put(o, "w0", a["w0"]); put(o, "w", a["w"]); put(o, "v", a["v"])
    return inj


def r_ffm(a):
    return two_pass(RI.FFMLayer(fcols(a["vocab"]), a["v"].shape[2]), a["inputs"], _inj_ffm(a))


def r_ffm_model(a):
    return two_pass(FFM(fcols(a["vocab"]), a["v"].shape[2]), a["inputs"], lambda o: _inj_ffm(a)(o.ffm))


def r_inner(a):
    return np.asarray(RI.InnerProductLayer()(a["e"]))


def r_outer(a):
    return two_pass(RI.OuterProductLayer(), a["e"], lambda o: put(o, "w", a["W"]))


def _inj_dice(o, alphas, mean, var):
    put(o, "alphas", alphas)
    assert o.bn.gamma is None and o.bn.beta is None          # center=False, scale=False (interaction.py:417)
    put(o.bn, "moving_mean", mean)
    put(o.bn, "moving_variance", var)


def r_dice(a):
    return two_pass(RI.Dice(), a["x"], lambda o: _inj_dice(o, a["alphas"], a["mean"], a["var"]))


def _inj_att(o, a):
    dense, out = synth._att_params(a)
    assert len(o.activation_layer) == len(dense)
    for lyr, (k, b, alpha) in zip(o.activation_layer, dense):
        put_dense(lyr, k, b)
        put(lyr.activation, "alpha", alpha)                   # PReLU alpha: [T, h] (interaction.py:366)
    put_dense(o.out_layer, *out)


def r_att(a):
    dense, _ = synth._att_params(a)
    lyr = RI.Attention([k.shape[1] for k, _, _ in dense], "prelu")
    x = [a["query"], a["key"], a["key"], a["mask"]]
    return two_pass(lyr, x, lambda o: _inj_att(o, a))


def r_att_dice(a):
    lyr = RI.Attention((16, 8), "dice")                       # the hidden sizes only set HOW MANY Dice layers (N3)
    x = [a["query"], a["key"], a["key"], a["mask"]]

    def inj(o):
        for i, d in enumerate(o.activation_layer):
            _inj_dice(d, a["dice_a%d" % i], a["dice_m%d" % i], a["dice_v%d" % i])
        put_dense(o.out_layer, a["att_ko"], a["att_bo"])
    return two_pass(lyr, x, inj)


def r_deepfm(a):
    hidden, _ = synth._get_stack(a, "dnn")
    m = DeepFM(fcols(a["vocab"]), a["v"].shape[1], 1e-4, 1e-4, [k.shape[1] for k, _ in hidden], 1, "relu")

    def inj(o):
        put_tables(o.embed_layer.emb_layers, synth._tables_of(a))
        _inj_fm(a)(o.fm)
        put_dnn(o.dnn, a)
    return two_pass(m, a["inputs"], inj)


def r_widedeep(a):
    hidden, _ = synth._get_stack(a, "dnn")
    m = WideDeep(fcols(a["vocab"]), [k.shape[1] for k, _ in hidden], 1, "relu")

    def inj(o):
        put_tables(o.embed_layer.emb_layers, synth._tables_of(a))
        put(o.wide, "w0", a["w0"]); put(o.wide, "w", a["w"])
        put_dnn(o.deep, a)
    return two_pass(m, a["inputs"], inj)


def r_dcn(a):
    hidden, _ = synth._get_stack(a, "dnn")
    L = sum(1 for k in a if k.startswith("cw"))
    m = DCN(fcols(a["vocab"]), [k.shape[1] for k, _ in hidden], 1, "relu", L)

    def inj(o):
        put_tables(o.embed_layer.emb_layers, synth._tables_of(a))
        for i in range(L):
            put(o.cross_layer, "cross_weight", a["cw%d" % i], i)
            put(o.cross_layer, "cross_bias", a["cb%d" % i], i)
        put_dnn(o.dense_layer, a)
        put_dense(o.output_layer, a["final_k"], a["final_b"])
    return two_pass(m, a["inputs"], inj)


class _Embed3D(RC.EmbedLayer):
    """The reference's EmbedLayer minus its last line (core.py:279): the [B, F, k] view PNN / NFM are written for."""

    def call(self, inputs, *args, **kwargs):
        flat = super().call(inputs, *args, **kwargs)
        k = self.emb_layers[0].output_dim
        return flat.reshape(flat.shape[0], flat.shape[1] // k, k)


def _raises_as_written(make, x):
    try:
        make()(x)
    except Exception:                                           # IndexError (PNN) / ValueError (NFM)
        return True
    return False


def r_pnn(a):
    hidden, _ = synth._get_stack(a, "dnn")
    mk = lambda: PNN(fcols(a["vocab"]), "both", [k.shape[1] for k, _ in hidden], 1, "relu")  # noqa: E731
    assert _raises_as_written(mk, a["inputs"]), "PNN.call ran as written: SURVEY N2 no longer holds"
    saved = ref_pnn.EmbedLayer
    ref_pnn.EmbedLayer = _Embed3D
    try:
        m = mk()
    finally:
        ref_pnn.EmbedLayer = saved

    def inj(o):
        put_tables(o.embed_layer.emb_layers, synth._tables_of(a))
        put(o.outer_product_layer, "w", a["W"])
        put_dnn(o.dnn_layer, a)
    return two_pass(m, a["inputs"], inj)


def r_nfm(a):
    hidden, _ = synth._get_stack(a, "dnn")
    mk = lambda: NFM(fcols(a["vocab"]), [k.shape[1] for k, _ in hidden], a["dnn_k%d" % len(hidden)].shape[1])  # noqa: E731
    assert _raises_as_written(mk, a["inputs"]), "NFM.call ran as written: SURVEY N2 no longer holds"
    saved = ref_nfm.EmbedLayer
    ref_nfm.EmbedLayer = _Embed3D
    try:
        m = mk()
    finally:
        ref_nfm.EmbedLayer = saved

    def inj(o):
        put_tables(o.emb_layers.emb_layers, synth._tables_of(a))
        put(o.bn_layer, "moving_mean", a["bn_mean"]); put(o.bn_layer, "moving_variance", a["bn_var"])
        put(o.bn_layer, "gamma", a["bn_gamma"]); put(o.bn_layer, "beta", a["bn_beta"])
        put_dnn(o.dnn_layers, a)
        put_dense(o.output_layer, a["final_k"], a["final_b"])
    return two_pass(m, a["inputs"], inj)


def _r_afm(mode):
    def run(a):
        m = AFM(fcols(a["vocab"], a["table0"].shape[1]), mode)

        def inj(o):
            lay = o.afm_layer
            put_tables([lay.embed_layer["emb_%d" % i] for i in range(len(a["vocab"]))], synth._tables_of(a))
            if mode == "att":
                put_dense(lay.attention_layer.attention_w, a["att_wk"], a["att_wb"])
                put_dense(lay.attention_layer.attention_h, a["att_hk"], a["att_hb"])
            put_dense(lay.output_layer, a["out_k"], a["out_b"])
        return two_pass(m, a["inputs"], inj)
    return run


def r_din(a):
    n_dense, n_other = a["dense"].shape[1], a["other_ids"].shape[1]
    K = a["seq_table0"].shape[1]
    dense_cols = [{"feat": "d%d" % i} for i in range(n_dense)]
    sparse_cols = [{"feat": "o%d" % i, "feat_onehot_dim": a["other_table%d" % i].shape[0], "embed_dim": K}
                   for i in range(n_other)]
    sparse_cols.append({"feat": "hist", "feat_onehot_dim": a["seq_table0"].shape[0], "embed_dim": K})
    dense_p, _ = synth._att_params(a)
    tower = []
    while "tw_k%d" % len(tower) in a:
        i = len(tower)
        tower.append((a["tw_k%d" % i], a["tw_b%d" % i], a["tw_a%d" % i]))
    m = DIN([dense_cols, sparse_cols], ["hist"], att_hidden_units=[k.shape[1] for k, _, _ in dense_p],
            dnn_hidden_units=[k.shape[1] for k, _, _ in tower])
    x = {"d%d" % i: a["dense"][:, i:i + 1] for i in range(n_dense)}
    x.update({"o%d" % i: a["other_ids"][:, i:i + 1] for i in range(n_other)})
    x["hist"] = a["hist_ids"][:, :, 0]                         # [B, T] (din.py:108-114)
    x["movie_id"] = a["cand_ids"]                              # [B, 1], the literal key of din.py:74

    def inj(o):
        put_tables(o.embed_sparse_layers, synth._tables_of(a, "other_table"))
        put_tables(o.embed_seq_layers, synth._tables_of(a, "seq_table"))
        _inj_att(o.att_layer, a)
        put(o.bn_layer, "gamma", a["bn_gamma"]); put(o.bn_layer, "beta", a["bn_beta"])
        put(o.bn_layer, "moving_mean", a["bn_mean"]); put(o.bn_layer, "moving_variance", a["bn_var"])
        assert o.bn_layer.epsilon == 1e-3
        for lyr, (k, b, alpha) in zip(o.dense_layer, tower):
            put_dense(lyr, k, b)
            put(lyr.activation, "alpha", alpha)               # PReLU alpha: [u] (din.py:51)
        put_dense(o.out_layer, a["out_k"], a["out_b"])
    return two_pass(m, x, inj)



# ------------------------------------------------------------------ the remaining towers
def _inj_res(o, a, prefix):
    hidden, out = synth._get_stack(a, prefix)
    assert len(o.dense_layer) == len(hidden)
    for lyr, (k, b) in zip(o.dense_layer, hidden):
        put_dense(lyr, k, b)
    put_dense(o.output_layer, *out)                            # Dense(input width), built in ResLayer.build (:268-269)


def r_res(a):
    hidden, _ = synth._get_stack(a, "res")
    return two_pass(RI.ResLayer([k.shape[1] for k, _ in hidden]), a["x"], lambda o: _inj_res(o, a, "res"))


def r_deepcrossing(a):
    hidden, _ = synth._get_stack(a, "res0")
    m = DeepCrossing(fcols(a["vocab"]), 32, [k.shape[1] for k, _ in hidden], 2)

    def inj(o):
        put_tables(o.embed_layer.emb_layers, synth._tables_of(a))
        for r, lyr in enumerate(o.res_layer):
            _inj_res(lyr, a, "res%d" % r)
        put_dense(o.output_layer, a["final_k"], a["final_b"])
    return two_pass(m, a["inputs"], inj)


def _inj_mmoe(o, a):
    T, gm, gb = synth._mmoe_parts(a)
    put(o, "expert_matrix", a["expert_matrix"])
    put(o, "expert_bias", a["expert_bias"])
    for t in range(T):
        put(o, "gate_matrix", gm[t], t)
        put(o, "gate_bias", gb[t], t)


def r_mmoe_layer(a):
    n_in, hidden, E = a["expert_matrix"].shape
    T, _, _ = synth._mmoe_parts(a)
    lyr = RI.mmoe_layer(hidden, E, T)
    lyr(a["x"])
    _inj_mmoe(lyr, a)
    return np.stack([np.asarray(t) for t in lyr(a["x"])], axis=0)


def r_mmoe(a):
    n_in, hidden, E = a["expert_matrix"].shape
    T, _, _ = synth._mmoe_parts(a)
    h0, _ = synth._get_stack(a, "tower0")
    m = MMOE(hidden, E, T, [k.shape[1] for k, _ in h0], 1)
    m(a["x"])
    _inj_mmoe(m.mmoe_layer, a)
    for t, tw in enumerate(m.tower_layer):
        put_dnn(tw, a, "tower%d" % t)
    return np.stack([np.asarray(t) for t in m(a["x"])], axis=0)


def _r_deepctr_dnn(act, use_bn, out_act):
    def run(a):
        ks, bs, bn = synth._dnn_parts(a)
        lyr = RC.DNN([k.shape[1] for k in ks], act, use_bn=use_bn, output_activation=out_act)

        def inj(o):
            for i in range(len(ks)):
                put(o, "kernels", ks[i], i)
                put(o, "bias", bs[i], i)
                if use_bn:
                    put(o.bn_layers[i], "moving_mean", bn[i][0]); put(o.bn_layers[i], "moving_variance", bn[i][1])
                    put(o.bn_layers[i], "gamma", bn[i][2]); put(o.bn_layers[i], "beta", bn[i][3])
        return two_pass(lyr, SO._t(a["x"]), inj)
    return run


def _r_lau(act):
    def run(a):
        ks, bs, dice = synth._lau_parts(a)
        lyr = RC.LocalActivationUnit([k.shape[1] for k in ks], act)
        x = [SO._t(a["query"]), SO._t(a["keys"])]

        def inj(o):
            put(o, "kernel", a["out_k"]); put(o, "bias", a["out_b"])
            for i in range(len(ks)):
                put(o.dnn, "kernels", ks[i], i)
                put(o.dnn, "bias", bs[i], i)
                if act == "dice":
                    _inj_dice(o.dnn.activation_layers[i], *dice[i])
        return two_pass(lyr, x, inj)
    return run


def _r_seqpool(wn, ret):
    """AttentionSequencePoolingLayer, DL/layer/sequence.py:180-276.  With ``supports_masking=False`` the reference's call()
    raises as written (``keys.get_shape().as_list[1]``: a method is subscripted, :252): asserted here.  The same body runs
    unmodified with ``supports_masking=True`` and the key mask handed in as keras would (``mask[-1]``)."""
    import layer.sequence as RS

    def run(a):
        ks, bs, _ = synth._lau_parts(a)
        x = [SO._t(a["query"]), SO._t(a["keys"])]
        broken = RS.AttentionSequencePoolingLayer([k.shape[1] for k in ks], "sigmoid", weight_normalization=wn, return_score=ret)
        try:
            broken(x + [SO._t(a["keys_length"])])
            raise AssertionError("the reference's keys_length path was expected to raise")
        except (TypeError, AttributeError):                    # TensorShape.as_list is a method (TypeError in TF); the shim's shape is a tuple
            pass
        lyr = RS.AttentionSequencePoolingLayer([k.shape[1] for k in ks], "sigmoid", weight_normalization=wn, return_score=ret,
                                               supports_masking=True)
        mask = [None, SO._t(a["key_mask"])]

        def inj(o):
            put(o.local_att, "kernel", a["out_k"]); put(o.local_att, "bias", a["out_b"])
            for i in range(len(ks)):
                put(o.local_att.dnn, "kernels", ks[i], i)
                put(o.local_att.dnn, "bias", bs[i], i)
        lyr(x, mask=mask)                                      # first pass builds the weights
        inj(lyr)
        return np.asarray(lyr(x, mask=mask))
    return run


def r_augru(a):
    """AUGRUCell.call (DL/layer/activation.py:113-142) stepped by the reference's own rnn_augru (:145-263), both unmodified, with
    the step closure of AUGRU.call (DL/layer/sequence.py:354-360).  The AUGRU layer object itself needs keras' RNN base class
    (state handling, ragged inputs) and is not constructed; its call() does nothing else on this path."""
    import layer.activation as RA
    cell = RA.AUGRUCell(a["rkernel"].shape[0], recurrent_activation="hard_sigmoid")      # (the base class resets it: see tf_shim)
    cell.build((None, a["kernel"].shape[0]))
    put(cell, "kernel", a["kernel"]); put(cell, "recurrent_kernel", a["rkernel"])

    def step(inputs, states, att_score):
        output, new_states = cell.call(inputs, states, att_score=att_score)
        return output, (new_states if isinstance(new_states, (list, tuple)) else [new_states])
    x, att = SO._t(a["x"]), SO._t(a["att"])
    h0 = [SO._t(np.zeros((x.shape[0], a["rkernel"].shape[0]), x.dtype))]
    last, outs, _ = RA.rnn_augru(step, x, h0, att_score=att)
    assert np.array_equal(np.asarray(last), np.asarray(outs)[:, -1])
    return np.asarray(outs)


RUNNERS = {
    "embed": r_embed, "wide": r_wide, "fm": r_fm, "dnn": r_dnn, "cross": r_cross, "ffm": r_ffm, "inner": r_inner,
    "outer": r_outer, "dice": r_dice, "att": r_att, "att_dice": r_att_dice, "deepfm": r_deepfm,
    "widedeep": r_widedeep, "dcn": r_dcn, "pnn": r_pnn, "din": r_din, "nfm": r_nfm,
    "afm_avg": _r_afm("avg"), "afm_max": _r_afm("max"), "afm_att": _r_afm("att"),
    "res": r_res, "deepcrossing": r_deepcrossing, "mmoe_layer": r_mmoe_layer, "mmoe": r_mmoe,
    "deepctr_dnn_relu": _r_deepctr_dnn("relu", False, None), "deepctr_dnn_bn_sigmoid": _r_deepctr_dnn("sigmoid", True, "tanh"),
    "lau_sigmoid": _r_lau("sigmoid"), "lau_dice": _r_lau("dice"),
    "seqpool_softmax": _r_seqpool(True, False), "seqpool_plain": _r_seqpool(False, False), "seqpool_score": _r_seqpool(True, True),
    "augru": r_augru,
}
MODEL_WRAPPERS = {"fm": r_fm_model, "ffm": r_ffm_model}


def as_dtype(a, dt):
    return {k: (v.astype(dt) if v.dtype == np.float64 else v) for k, v in a.items()}


def main():
    frozen = np.load(os.path.join(HERE, "layers_tiny.npz"))
    cases = synth.tiny_case(seed=7)
    assert set(cases) == set(RUNNERS), set(cases) ^ set(RUNNERS)
    out = {}
    for name, (_, _, arrays) in cases.items():
        for k, v in arrays.items():                            # same inputs / weights as the frozen oracle fixture
            assert np.array_equal(frozen["%s/%s" % (name, k)], v), (name, k)
        for dt, key in ((np.float64, "expected"), (np.float32, "expected_f32")):
            SL.FLOATX[0] = dt
            a = as_dtype(arrays, dt)
            out["%s/%s" % (name, key)] = RUNNERS[name](a)
            if name in MODEL_WRAPPERS:
                out["%s_model/%s" % (name, key)] = MODEL_WRAPPERS[name](a)
        print("%-10s ref %s  max|ref - oracle| = %.3g   f32-vs-f64 %.3g" % (
            name, out[name + "/expected"].shape,
            np.max(np.abs(out[name + "/expected"] - frozen[name + "/expected"])),
            np.max(np.abs(out[name + "/expected_f32"].astype(np.float64) - out[name + "/expected"]))))
    path = os.path.join(HERE, "ref_layers.npz")
    if "--check" in sys.argv:                                  # tests/test_ref_golden.py: the committed file IS this output
        have = np.load(path)
        assert set(have.files) == set(out), set(have.files) ^ set(out)
        for k, v in out.items():
            assert have[k].dtype == v.dtype and np.array_equal(have[k], v), k
        print("ref_layers.npz: committed fixture reproduced bit for bit (%d arrays)" % len(out))
        return
    np.savez_compressed(path, **out)
    print("ref_layers.npz:", len(out), "arrays")


if __name__ == "__main__":
    main()
