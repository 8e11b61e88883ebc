"""CPU oracle for the CTR forward hot path -- TEST INFRASTRUCTURE ONLY.

This module is the checker, never the product: only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it.  Nothing under
``recommender_system_b200/`` imports anything from ``oracle/``.

PARITY: PINNED AGAINST THE REFERENCE'S OWN PYTHON, not against TensorFlow's kernels.  The reference
(Hcyand/recommender_system) ships no tests, fixtures or golden vectors and delegates its arithmetic to TensorFlow/Keras
(third party, not vendored, version unpinned; import style implies TF 2.6-2.11), which cannot be installed here.  But
the reference itself is pure Python: ``tests/golden/make_ref_golden.py`` imports its UNMODIFIED
``layer/interaction.py``, ``layer/core.py`` and ``model/{fm,ffm,deepFM,wideDeep,dcn,pnn,din,nfm,afm}.py`` under
``tests/golden/tf_shim`` (numpy stand-ins for the ~40 TF/Keras symbols those files use), injects the seeded weights of
``oracle/synth.py`` and stores what their ``call()`` bodies return in ``tests/golden/ref_layers.npz``.
``tests/test_oracle.py::test_oracle_matches_the_reference`` holds every function below to that file at <= 1e-12
(measured: <= 2.3e-16) -- op order, axes, pair order, one-hot layout, mask fill, the [T,h] PReLU alpha, Dice's
BatchNormalization without gamma/beta are the reference's own.  Not pinned: the TF kernels under the shim (their
documented semantics are restated in tf_shim/ops.py and layers.py) and two models that raise as written (PNN, NFM:
SURVEY N2), which run with EmbedLayer's final reshape skipped and nothing else changed.

All citations are relative to /root/reference/algorithm/deep_learning/ ("DL/").

Every function is dtype-parametric: call with float64 arrays for the O64
tolerance judge, float32 arrays for O32 (the reference's own precision and op
order).  Keras defaults restated here:

* ``Embedding``: ids cast to int32 (truncation), then a row gather.
* ``Dense``: ``x @ kernel + bias`` on the last axis, ``kernel`` is [in, out].
* ``PReLU``: ``max(x,0) - alpha*max(-x,0)``, alpha shaped like the activation
  input minus the batch axis.
* ``BatchNormalization`` (inference): ``gamma*(x-mean)/sqrt(var+eps)+beta``,
  eps 1e-3 by default; inside Dice eps=1e-9 and no gamma/beta.
* ``Dropout``: identity at inference.  ``softmax``: last axis.
"""
from __future__ import annotations

import numpy as np

MASK_FILL = float(-2 ** 32 + 1)  # DL/layer/interaction.py:399


# --------------------------------------------------------------------------- helpers
def sigmoid(x):
    x = np.asarray(x)
    out = np.empty_like(x)
    pos = x >= 0
    out[pos] = 1.0 / (1.0 + np.exp(-x[pos]))
    ex = np.exp(x[~pos])
    out[~pos] = ex / (1.0 + ex)
    return out


def softmax_last(x):
    m = np.max(x, axis=-1, keepdims=True)
    e = np.exp(x - m)
    return e / np.sum(e, axis=-1, keepdims=True)


def ids_to_int(ids):
    """keras Embedding casts its input to int32 (truncation toward zero)."""
    return np.asarray(ids).astype(np.int64)


def activation(name, x):
    if name is None or name == "linear":
        return x
    if name == "relu":
        return np.maximum(x, 0)
    if name == "sigmoid":
        return sigmoid(x)
    if name == "tanh":
        return np.tanh(x)
    raise ValueError("unknown activation %r" % (name,))


def prelu(x, alpha):
    """keras.layers.PReLU.call: relu(x) - alpha*relu(-x)."""
    return np.maximum(x, 0) - alpha * np.maximum(-x, 0)


def batchnorm_infer(x, mean, var, gamma=None, beta=None, eps=1e-3):
    y = (x - mean) / np.sqrt(var + eps)
    if gamma is not None:
        y = y * gamma
    if beta is not None:
        y = y + beta
    return y


def pair_index(field_num):
    """Row-major (i<j) pair order, DL/layer/interaction.py:174-177, 203-206."""
    row, col = [], []
    for i in range(field_num - 1):
        for j in range(i + 1, field_num):
            row.append(i)
            col.append(j)
    return np.asarray(row), np.asarray(col)


# --------------------------------------------------------------------------- R1
def embed_layer(sparse_inputs, tables, flatten=True):
    """EmbedLayer.call, DL/layer/core.py:273-280.

    26 separate gathers -> stack [F,B,k] -> transpose [B,F,k] -> reshape
    [B,F*k].  Out-of-range ids raise IndexError (CPU TF raises
    InvalidArgumentError).
    """
    ids = ids_to_int(sparse_inputs)
    embs = []
    for f, table in enumerate(tables):
        col = ids[:, f]
        if col.size and (col.min() < 0 or col.max() >= table.shape[0]):
            raise IndexError("id out of range in field %d" % f)
        embs.append(table[col])
    emb = np.transpose(np.stack(embs, axis=0), (1, 0, 2))
    if flatten:
        emb = emb.reshape(emb.shape[0], emb.shape[1] * emb.shape[2])
    return emb


# --------------------------------------------------------------------------- R2
def wide_layer(x, w0, w):
    """WideLayer.call, DL/layer/interaction.py:24-26."""
    return w0 + x @ w


# --------------------------------------------------------------------------- R3
def fm_layer(x, w0, w1, v):
    """FMLayer.call, DL/layer/interaction.py:106-114."""
    linear_part = x @ w1 + w0
    inter_part1 = np.power(x @ v, 2)
    inter_part2 = np.power(x, 2) @ np.power(v, 2)
    inter_part = 0.5 * np.sum(inter_part1 - inter_part2, axis=-1, keepdims=True)
    return linear_part + inter_part


# --------------------------------------------------------------------------- R4
def dnn_layer(x, hidden, out, act="relu"):
    """DNNLayer.call, DL/layer/interaction.py:40-46.

    ``hidden`` is a list of (kernel[in,out], bias[out]); ``out`` one such pair.
    Dropout (:44) is the identity at inference.
    """
    for kernel, bias in hidden:
        x = activation(act, x @ kernel + bias)
    kernel, bias = out
    return x @ kernel + bias


# --------------------------------------------------------------------------- R8
def cross_layer(x, ws, bs):
    """CrossLayer.call, DL/layer/interaction.py:75-83.  ws/bs: lists of [d,1]."""
    x0 = x[:, :, None]
    x1 = x0
    for w, b in zip(ws, bs):
        x1_w = np.matmul(np.transpose(x1, (0, 2, 1)), w)        # [B,1,1]
        x1 = np.matmul(x0, x1_w) + b + x1
    return x1[:, :, 0]


# --------------------------------------------------------------------------- R5
def onehot_expand(inputs, vocab_sizes, n_dense=13, dtype=None):
    """FFMLayer.call's one-hot construction, DL/layer/interaction.py:142-150.

    tf.one_hot yields an all-zero row for an out-of-range index.
    """
    inputs = np.asarray(inputs)
    dtype = dtype or inputs.dtype
    dense = inputs[:, :n_dense].astype(dtype)
    ids = ids_to_int(inputs[:, n_dense:])
    cols = [dense]
    for f, vsz in enumerate(vocab_sizes):
        oh = np.zeros((inputs.shape[0], vsz), dtype=dtype)
        c = ids[:, f]
        ok = (c >= 0) & (c < vsz)
        oh[np.nonzero(ok)[0], c[ok]] = 1
        cols.append(oh)
    return np.concatenate(cols, axis=1)


def ffm_layer(inputs, vocab_sizes, w0, w, v, n_dense=13):
    """FFMLayer.call, DL/layer/interaction.py:141-163 (dense one-hot form).

    w:[n_feat,1], v:[n_feat, field_num, k].  NOT textbook FFM.
    """
    x = onehot_expand(inputs, vocab_sizes, n_dense, dtype=v.dtype)
    linear_part = w0 + x @ w
    field_f = np.tensordot(x, v, axes=1)                         # [B, field_num, k]
    field_num = v.shape[1]
    inter_part = np.zeros((x.shape[0], 1), dtype=v.dtype)
    for i in range(field_num):
        for j in range(i + 1, field_num):
            inter_part = inter_part + np.sum(field_f[:, i] * field_f[:, j], axis=1, keepdims=True)
    return linear_part + inter_part


# --------------------------------------------------------------------------- R6 / R7
def inner_product_layer(e):
    """InnerProductLayer.call, DL/layer/interaction.py:170-183.  e:[B,F,k]."""
    row, col = pair_index(e.shape[1])
    et = np.transpose(e, (1, 0, 2))
    p = np.transpose(et[row], (1, 0, 2))
    q = np.transpose(et[col], (1, 0, 2))
    return np.sum(p * q, axis=-1)


def outer_product_layer(e, w):
    """OuterProductLayer.call, DL/layer/interaction.py:201-215.  w:[k,P,k]."""
    row, col = pair_index(e.shape[1])
    et = np.transpose(e, (1, 0, 2))
    p = np.transpose(et[row], (1, 0, 2))                         # [B,P,k]
    q = np.transpose(et[col], (1, 0, 2))
    p = p[:, None, :, :]                                         # [B,1,P,k]
    tmp = p * w                                                  # [B,k,P,k]
    tmp = np.sum(tmp, axis=-1)                                   # [B,k,P]
    tmp = np.transpose(tmp, (0, 2, 1)) * q                       # [B,P,k]
    return np.sum(tmp, axis=-1)


# --------------------------------------------------------------------------- R10
def dice(x, alphas, mean, var, eps=1e-9):
    """Dice.call, DL/layer/interaction.py:422-425 (inference BN, no gamma/beta)."""
    xn = batchnorm_infer(x, mean, var, eps=eps)
    x_p = sigmoid(xn)
    return alphas * (1.0 - x_p) * x + x_p * x


# --------------------------------------------------------------------------- R9
def din_attention(query, key, value, mask, dense, out, act="prelu", dice_params=None):
    """DIN Attention.call, DL/layer/interaction.py:369-406.

    ``act='prelu'``: ``dense`` is a list of (kernel[in,h], bias[h], alpha[T,h]).
    ``act='dice'`` : no Dense projection at all; ``dice_params`` is a
    list of (alphas[4K], mean[4K], var[4K]), applied on the 4K-wide concat.
    ``out`` = (kernel[h_last,1], bias[1]).
    """
    T = key.shape[1]
    q = np.tile(query[:, None, :], (1, T, 1))
    emb = np.concatenate([q, key, q - key, q * key], axis=-1)    # [B,T,4K]
    if act == "prelu":
        for kernel, bias, alpha in dense:
            emb = prelu(emb @ kernel + bias, alpha)
    elif act == "dice":
        for alphas, mean, var in dice_params:
            emb = dice(emb, alphas, mean, var)
    else:
        raise ValueError(act)
    kernel, bias = out
    score = (emb @ kernel + bias)[..., 0]                        # [B,T]
    padding = np.ones_like(score) * score.dtype.type(MASK_FILL)
    score = np.where(mask == 0, padding, score)
    score = softmax_last(score)
    return np.matmul(score[:, None, :], value)[:, 0, :]


# --------------------------------------------------------------------------- R11-R16 models
def fm_model(x, w0, w1, v):
    """FM.call, DL/model/fm.py:19-23."""
    return sigmoid(fm_layer(x, w0, w1, v))


def ffm_model(inputs, vocab_sizes, w0, w, v):
    """FFM.call, DL/model/ffm.py:20-23."""
    return sigmoid(ffm_layer(inputs, vocab_sizes, w0, w, v))


def deepfm_model(inputs, tables, fm, dnn_hidden, dnn_out, act="relu", n_dense=13):
    """DeepFM.call, DL/model/deepFM.py:23-31.  fm=(w0,w1,v) over d=n_dense+F*k."""
    dense_inputs, sparse_inputs = inputs[:, :n_dense], inputs[:, n_dense:]
    sparse_embed = embed_layer(sparse_inputs, tables)
    x = np.concatenate([dense_inputs.astype(sparse_embed.dtype), sparse_embed], axis=-1)
    fm_output = fm_layer(x, *fm)
    dnn_output = dnn_layer(x, dnn_hidden, dnn_out, act)
    return sigmoid(0.5 * (fm_output + dnn_output))


def widedeep_model(inputs, tables, wide, dnn_hidden, dnn_out, act="relu", n_dense=13):
    """WideDeep.call, DL/model/wideDeep.py:22-34.

    inputs = [dense(13), ids(26), onehot(...)]; the deep side sees embeddings
    only.  wide=(w0,w) with w over 13+len(onehot).
    """
    F = len(tables)
    dense_inputs = inputs[:, :n_dense]
    sparse_inputs = inputs[:, n_dense:n_dense + F]
    onehot_inputs = inputs[:, n_dense + F:]
    wide_input = np.concatenate([dense_inputs, onehot_inputs], axis=1)
    wide_output = wide_layer(wide_input.astype(wide[1].dtype), *wide)
    sparse_embed = embed_layer(sparse_inputs, tables)
    deep_output = dnn_layer(sparse_embed, dnn_hidden, dnn_out, act)
    return sigmoid(0.5 * (wide_output + deep_output))


def dcn_model(inputs, tables, cross_ws, cross_bs, dnn_hidden, dnn_out, final, act="relu", n_dense=13):
    """DCN.call, DL/model/dcn.py:24-34.  final=(kernel[d+od,1], bias[1])."""
    dense_inputs, sparse_inputs = inputs[:, :n_dense], inputs[:, n_dense:]
    sparse_embed = embed_layer(sparse_inputs, tables)
    x = np.concatenate([dense_inputs.astype(sparse_embed.dtype), sparse_embed], axis=1)
    cross_output = cross_layer(x, cross_ws, cross_bs)
    dnn_output = dnn_layer(x, dnn_hidden, dnn_out, act)
    x = np.concatenate([cross_output, dnn_output], axis=1)
    kernel, bias = final
    return sigmoid(x @ kernel + bias)


def bi_interaction(emb):
    """NFM's Bi-Interaction pooling, DL/model/nfm.py:28: 0.5 * ((sum_f e_f)^2 - sum_f e_f^2) over the field axis of [B,F,k]."""
    return 0.5 * (np.sum(emb, axis=1) ** 2 - np.sum(emb ** 2, axis=1))


def nfm_model(inputs, tables, bn, dnn_hidden, dnn_out, final, act="relu", n_dense=13):
    """NFM.call, DL/model/nfm.py:24-34.  The reference pools ``emb`` over ``axis=1`` of what its own
    comment calls ``[None, 26, embed_dim]``; its EmbedLayer returns the flattened [B, F*k] (core.py:279), so the model as
    written does not run -- this follows the comment: the [B, F, k] view.
    bn = (mean, var, gamma, beta) of keras BatchNormalization (eps 1e-3); final = (kernel [out_dim,1], bias [1])."""
    dense_inputs, sparse_inputs = inputs[:, :n_dense], inputs[:, n_dense:n_dense + len(tables)]
    emb = embed_layer(sparse_inputs, tables, flatten=False)
    x = np.concatenate([dense_inputs, bi_interaction(emb)], axis=-1)
    x = batchnorm_infer(x, bn[0], bn[1], bn[2], bn[3], eps=1e-3)
    y = dnn_layer(x, dnn_hidden, dnn_out, act)
    return sigmoid(y @ final[0] + final[1])


def afm_layer(inputs, tables, mode, att_w=None, att_h=None, out=None, n_dense=13):
    """AFMLayer.call, DL/layer/interaction.py:337-351, literally: InteractionLayer (:291-297) builds every e_i * e_j, i < j;
    'avg' / 'max' pool over the pairs; 'att' runs AttentionLayer (:313-319): Dense(P, relu) and Dense(1) on the pairs, then
    ``tf.nn.softmax`` over the LAST axis of the [B, P, 1] scores -- a size-1 axis, so every attention weight is exactly 1 and the
    result is the sum over pairs.  att_w = (kernel [k,P], bias [P]), att_h = (kernel [P,1], bias [1]), out = (kernel [k,1], bias [1])."""
    sparse_inputs = inputs[:, n_dense:n_dense + len(tables)]
    e = embed_layer(sparse_inputs, tables, flatten=False)
    F = e.shape[1]
    pairs = np.stack([e[:, i] * e[:, j] for i in range(F) for j in range(i + 1, F)], axis=1)      # [B, P, k]
    if mode == "avg":
        x = pairs.mean(axis=1)
    elif mode == "max":
        x = pairs.max(axis=1)
    else:
        h = np.maximum(pairs @ att_w[0] + att_w[1], 0.0)                                         # [B, P, P]
        sc = h @ att_h[0] + att_h[1]                                                              # [B, P, 1]
        a = softmax_last(sc)                                                                      # over the size-1 axis
        x = (np.transpose(a, (0, 2, 1)) @ pairs)[:, 0, :]
    return sigmoid(x @ out[0] + out[1])


def afm_model(inputs, tables, mode, att_w=None, att_h=None, out=None, n_dense=13):
    """AFM.call, DL/model/afm.py: sigmoid of the layer's (already sigmoid) output."""
    return sigmoid(afm_layer(inputs, tables, mode, att_w, att_h, out, n_dense))


def pnn_model(inputs, tables, mode, dnn_hidden, dnn_out, outer_w=None, act="relu", n_dense=13):
    """PNN.call, DL/model/pnn.py:28-53 with the [B,F,k] view the product layers
    are specified on.  No sigmoid (:52-53)."""
    sparse_inputs = inputs[:, n_dense:]
    z = embed_layer(sparse_inputs, tables, flatten=False)
    embed = z.reshape(z.shape[0], z.shape[1] * z.shape[2])
    if mode == "inner":
        feats = [embed, inner_product_layer(z)]
    elif mode == "outer":
        feats = [embed, outer_product_layer(z, outer_w)]
    elif mode == "both":
        feats = [embed, inner_product_layer(z), outer_product_layer(z, outer_w)]
    else:
        raise ValueError("Please choice mode's value in 'inner', 'outer', 'both'.")
    return dnn_layer(np.concatenate(feats, axis=1), dnn_hidden, dnn_out, act)


def din_model(dense_inputs, other_ids, hist_ids, cand_ids,
              other_tables, seq_tables, att_dense, att_out,
              bn, tower, out, att_act="prelu", att_dice=None, dnn_act="prelu"):
    """DIN.call, DL/model/din.py:56-95.

    dense_inputs [B,nd]; other_ids [B,no]; hist_ids [B,T,nb]; cand_ids [B,nb];
    seq_tables are shared by history and candidate (:76-79).
    bn=(gamma,beta,mean,var) eps 1e-3 (:50,:89).
    tower: list of (kernel,bias,alpha[u]) for 'prelu', or
    (kernel,bias,(alphas,mean,var)) for Dice (:51).
    """
    dt = seq_tables[0].dtype
    other_ids = ids_to_int(other_ids)
    hist_ids_i = ids_to_int(hist_ids)
    cand_ids = ids_to_int(cand_ids)
    parts = [other_tables[i][other_ids[:, i]] for i in range(len(other_tables))]
    parts.append(np.asarray(dense_inputs, dtype=dt))
    other_feat = np.concatenate(parts, axis=-1)
    seq_embed = np.concatenate([seq_tables[i][hist_ids_i[:, :, i]] for i in range(len(seq_tables))], axis=-1)
    item_embed = np.concatenate([seq_tables[i][cand_ids[:, i]] for i in range(len(seq_tables))], axis=-1)
    mask = (np.asarray(hist_ids)[:, :, 0] != 0).astype(dt)
    att_emb = din_attention(item_embed, seq_embed, seq_embed, mask, att_dense, att_out, att_act, att_dice)
    emb = np.concatenate([att_emb, item_embed, other_feat], axis=-1)
    gamma, beta, mean, var = bn
    emb = batchnorm_infer(emb, mean, var, gamma, beta, eps=1e-3)
    for kernel, bias, a in tower:
        h = emb @ kernel + bias
        emb = prelu(h, a) if dnn_act == "prelu" else dice(h, *a)
    kernel, bias = out
    return sigmoid(emb @ kernel + bias)


# --------------------------------------------------------------------------- sparse restatements
# The B200 kernels never materialise the one-hot matrix.  These are the
# gather-form equivalents, kept here so the tests can prove dense == sparse
# on the oracle itself before any GPU is involved.
def fm_onehot_sparse(dense_x, dense_cols, ids, offsets, w0, w1, v):
    """FM over [dense, one-hot] as a gather.  Column of (field f, id i) is
    offsets[f]+i; dense feature j sits at column dense_cols[j]."""
    ids = ids_to_int(ids)
    B = ids.shape[0]
    cols = ids + np.asarray(offsets)[None, :]
    lin = w0 + dense_x @ w1[dense_cols] + np.sum(w1[cols, 0], axis=1, keepdims=True)
    s = dense_x @ v[dense_cols] + np.sum(v[cols], axis=1)                       # [B,k]
    s2 = (dense_x ** 2) @ (v[dense_cols] ** 2) + np.sum(v[cols] ** 2, axis=1)
    assert s.shape[0] == B
    return lin + 0.5 * np.sum(s * s - s2, axis=-1, keepdims=True)


def ffm_sparse(inputs, vocab_sizes, w0, w, v, n_dense=13):
    """FFMLayer as gathers + the pair-sum identity
    sum_{i<j}<a_i,a_j> = 0.5*(|sum a|^2 - sum |a|^2)."""
    dense = np.asarray(inputs[:, :n_dense], dtype=v.dtype)
    ids = ids_to_int(inputs[:, n_dense:])
    offs = n_dense + np.concatenate([[0], np.cumsum(vocab_sizes)[:-1]])
    lin = w0 + dense @ w[:n_dense]
    field_f = np.tensordot(dense, v[:n_dense], axes=1)
    for f, vsz in enumerate(vocab_sizes):
        c = ids[:, f]
        ok = ((c >= 0) & (c < vsz)).astype(v.dtype)
        cc = np.clip(c, 0, vsz - 1) + offs[f]
        lin = lin + ok[:, None] * w[cc]
        field_f = field_f + ok[:, None, None] * v[cc]
    tot = np.sum(field_f, axis=1)
    inter = 0.5 * (np.sum(tot * tot, axis=1, keepdims=True) - np.sum(field_f * field_f, axis=(1, 2))[:, None])
    return lin + inter


# --------------------------------------------------------------------------- the remaining towers
def res_layer(x, hidden, out):
    """ResLayer.call, DL/layer/interaction.py:271-278: ``relu(x + Dense_out(relu-Dense chain(x)))``; ``out`` maps back to
    the input width (:269)."""
    h = x
    for kernel, bias in hidden:
        h = np.maximum(h @ kernel + bias, 0)
    h = h @ out[0] + out[1]
    return np.maximum(x + h, 0)


def deepcrossing_model(inputs, tables, res_layers, final, n_dense=13):
    """DeepCrossing.call, DL/model/deepCrossing.py:25-32: x = [dense, Embed(ids)]; the residual units; Dense(1, sigmoid).
    res_layers: list of (hidden pairs, out pair); final = (kernel [d,1], bias [1])."""
    dense_inputs, sparse_inputs = inputs[:, :n_dense], inputs[:, n_dense:]
    emb = embed_layer(sparse_inputs, tables)
    x = np.concatenate([dense_inputs.astype(emb.dtype), emb], axis=-1)
    for hidden, out in res_layers:
        x = res_layer(x, hidden, out)
    return sigmoid(x @ final[0] + final[1])


def mmoe_layer(x, expert_matrix, expert_bias, gate_matrix, gate_bias):
    """mmoe_layer.call, DL/layer/interaction.py:481-509.  expert_matrix [in, hidden, E], expert_bias [hidden, E] or None,
    gate_matrix: list over tasks of [in, E], gate_bias: list of [E] or None.  -> list over tasks of [B, hidden]."""
    E = expert_matrix.shape[2]
    expert_output = np.transpose(np.stack([x @ expert_matrix[:, :, i] for i in range(E)], axis=0), (1, 2, 0))  # [B, hidden, E]
    if expert_bias is not None:
        expert_output = expert_output + expert_bias
    expert_output = np.maximum(expert_output, 0)
    outs = []
    for t, gate in enumerate(gate_matrix):
        g = x @ gate
        if gate_bias is not None:
            g = g + gate_bias[t]
        g = softmax_last(g)                                                   # [B, E]
        g = np.tile(g[:, None, :], (1, expert_output.shape[1], 1))            # [B, hidden, E]
        outs.append(np.sum(g * expert_output, axis=-1))
    return outs


def tower_layer(x, hidden, out, act="relu"):
    """tower_layer.call, DL/layer/interaction.py:518-523: DNNLayer without the dropout."""
    return dnn_layer(x, hidden, out, act)


def mmoe_model(x, expert_matrix, expert_bias, gate_matrix, gate_bias, towers, act="relu"):
    """MMOE.call, DL/model/mmoe.py:24-32.  towers: list over tasks of (hidden pairs, out pair)."""
    outs = mmoe_layer(x, expert_matrix, expert_bias, gate_matrix, gate_bias)
    return [tower_layer(o, h, w, act) for o, (h, w) in zip(outs, towers)]


def deepctr_dnn(x, kernels, biases, act="relu", bn=None, output_activation=None, dice=None):
    """DNN.call, DL/layer/core.py:181-199 (the DeepCTR multi-layer perceptron): per layer ``tensordot(x, kernel) + bias``,
    optional BatchNormalization (inference, eps 1e-3) BEFORE the activation, then the activation (the last layer takes
    ``output_activation`` when given), Dropout = identity.  Works on the last axis of an nD input.
    bn: list of (mean, var, gamma, beta) or None; dice: list of (alphas, mean, var) when act == 'dice'."""
    n = len(kernels)
    for i in range(n):
        x = np.tensordot(x, kernels[i], axes=(-1, 0)) + biases[i]
        if bn is not None:
            x = batchnorm_infer(x, bn[i][0], bn[i][1], bn[i][2], bn[i][3], eps=1e-3)
        a = output_activation if (output_activation and i == n - 1) else act
        x = globals()["dice"](x, *dice[i]) if a in ("dice", "Dice") else activation(a, x)
    return x


def local_activation_unit(query, keys, kernels, biases, out_kernel, out_bias, act="sigmoid", dice=None):
    """LocalActivationUnit.call, DL/layer/core.py:92-107: query [B,1,K] repeated over T, ``[q, k, q-k, q*k]`` -> DNN ->
    ``tensordot(., kernel [h,1]) + bias`` -> [B,T,1] (no mask, no softmax: those belong to the pooling layer)."""
    T = keys.shape[1]
    q = np.repeat(query, T, axis=1)
    att_input = np.concatenate([q, keys, q - keys, q * keys], axis=-1)
    att_out = deepctr_dnn(att_input, kernels, biases, act, dice=dice)
    return np.tensordot(att_out, out_kernel, axes=(-1, 0)) + out_bias


def attention_sequence_pooling(query, keys, key_mask, kernels, biases, out_kernel, out_bias, act="sigmoid",
                               weight_normalization=False, return_score=False, dice=None):
    """AttentionSequencePoolingLayer.call, DL/layer/sequence.py:242-276: scores = LocalActivationUnit([query, keys]) [B,T,1]
    -> transpose [B,1,T] -> where(key_mask, scores, paddings) with paddings -2^32+1 (weight_normalization) or 0 -> softmax over T
    (weight_normalization only) -> @ keys [B,1,E] (or the scores themselves when return_score).  key_mask [B,T] bool: the
    layer's ``mask[-1]`` (supports_masking) or ``sequence_mask(keys_length, T)``."""
    score = np.transpose(local_activation_unit(query, keys, kernels, biases, out_kernel, out_bias, act, dice=dice), (0, 2, 1))
    pad = (-2.0 ** 32 + 1) if weight_normalization else 0.0
    out = np.where(np.asarray(key_mask, bool)[:, None, :], score, pad)
    if weight_normalization:
        out = softmax_last(out)
    return out if return_score else out @ keys


def hard_sigmoid(x):
    """keras.backend.hard_sigmoid: clip(0.2 x + 0.5, 0, 1)."""
    return np.clip(0.2 * x + 0.5, 0.0, 1.0)


def augru_cell(x, h, att, kernel, recurrent_kernel, recurrent_activation="sigmoid", activation="tanh"):
    """AUGRUCell.call, DL/layer/activation.py:113-142 (no biases): z = ra(x Wz + h Uz) * att; r = ra(x Wr + h Ur);
    hh = act(x Wh + (r * h) Uh); h' = z h + (1 - z) hh.  kernel [E, 3H] and recurrent_kernel [H, 3H] hold the z | r | h blocks.
    Under TF 2 the cell's activations are tanh / SIGMOID whatever its constructor is given: keras' GRUCell.__init__ overwrites
    the attributes AUGRUCell sets before calling it (activation.py:97-100)."""
    H = recurrent_kernel.shape[0]
    ra = sigmoid if recurrent_activation == "sigmoid" else hard_sigmoid
    act = np.tanh if activation == "tanh" else (lambda v: v)
    z = ra(x @ kernel[:, :H] + h @ recurrent_kernel[:, :H]) * att
    r = ra(x @ kernel[:, H:2 * H] + h @ recurrent_kernel[:, H:2 * H])
    hh = act(x @ kernel[:, 2 * H:] + (r * h) @ recurrent_kernel[:, 2 * H:])
    return z * h + (1 - z) * hh


def augru(x, att, kernel, recurrent_kernel, h0=None, recurrent_activation="sigmoid", activation="tanh"):
    """rnn_augru without a mask, DL/layer/activation.py:237-244 (what AUGRU.call, DL/layer/sequence.py:293-395, unrolls):
    h_t = cell(x[:, t], h_{t-1}, att[:, t, :]); -> (last output [B,H], outputs [B,T,H])."""
    B, T, _ = x.shape
    h = np.zeros((B, recurrent_kernel.shape[0]), x.dtype) if h0 is None else h0
    outs = []
    for t in range(T):
        h = augru_cell(x[:, t], h, att[:, t, :], kernel, recurrent_kernel, recurrent_activation, activation)
        outs.append(h)
    return h, np.stack(outs, axis=1)


# --------------------------------------------------------------------------- one SGD step of DeepFM
# PARITY UNPINNED for this function: the reference trains through TensorFlow's autodiff and optimizer
# (DL/utils/compile_fit.py:9-15), which cannot run in this container, and it ships no training fixtures.  The forward
# half is the pinned `deepfm_model` above; the gradients are pinned by central finite differences of this file's own
# loss (tests/test_oracle.py::test_deepfm_train_step_gradients).
def deepfm_loss(dense, ids, tables, fm_w0, fm_w1, fm_v, hidden, out, y, reg_w, reg_v, drop_masks=None):
    """keras' training loss of DeepFM for one batch: mean BCE(y, clip(p, 1e-7, 1-1e-7)) + reg_w*sum(w1^2) + reg_v*sum(v^2)
    (binary_crossentropy; l2 regularisers of FMLayer.build, DL/layer/interaction.py:97-104).  hidden: [(kernel, bias)] with
    ReLU; out: (kernel [h,1], bias [1]); drop_masks: per hidden layer a [B, h_l] array of 0 or 1/(1-rate) (Dropout in
    training mode, interaction.py:40-46) or None.  -> (loss, mean_bce, p)"""
    ids = ids_to_int(ids)
    x = np.concatenate([np.asarray(dense, np.float64)] + [tables[f][ids[:, f]] for f in range(len(tables))], axis=1)
    fm = fm_layer(x, fm_w0, fm_w1, fm_v)[:, 0]
    h = x
    for l, (k, b) in enumerate(hidden):
        h = np.maximum(h @ k + b, 0.0)
        if drop_masks is not None:
            h = h * drop_masks[l]
    dnn = (h @ out[0] + out[1])[:, 0]
    p = sigmoid(0.5 * (fm + dnn))
    pc = np.clip(p, 1e-7, 1 - 1e-7)
    bce = float(np.mean(-(y * np.log(pc) + (1 - y) * np.log(1 - pc))))
    return bce + reg_w * float(np.sum(fm_w1 ** 2)) + reg_v * float(np.sum(fm_v ** 2)), bce, p


def deepfm_train_step(dense, ids, tables, fm_w0, fm_w1, fm_v, hidden, out, y, lr, reg_w, reg_v, drop_masks=None):
    """One step of keras SGD(lr) on `deepfm_loss` (what compile_fit's model.fit does per batch).  Analytic gradients;
    returns the UPDATED (tables, fm_w0, fm_w1, fm_v, hidden, out), the mean BCE before the update and the number of correct
    predictions.  Embedding rows looked up by several samples receive the sum of their gradients (np.add.at), rows that
    were not looked up do not move."""
    ids = ids_to_int(ids)
    B, D, F = ids.shape[0], np.asarray(dense).shape[1], len(tables)
    x = np.concatenate([np.asarray(dense, np.float64)] + [tables[f][ids[:, f]] for f in range(F)], axis=1)
    xv = x @ fm_v
    fm = (fm_w0 + x @ fm_w1)[:, 0] + 0.5 * (np.sum(xv ** 2, axis=1) - (x ** 2) @ np.sum(fm_v ** 2, axis=1))
    acts = [x]
    for l, (k, b) in enumerate(hidden):
        h = np.maximum(acts[-1] @ k + b, 0.0)
        if drop_masks is not None:
            h = h * drop_masks[l]
        acts.append(h)
    dnn = (acts[-1] @ out[0] + out[1])[:, 0]
    p = sigmoid(0.5 * (fm + dnn))
    pc = np.clip(p, 1e-7, 1 - 1e-7)
    bce = float(np.mean(-(y * np.log(pc) + (1 - y) * np.log(1 - pc))))
    correct = int(np.sum((p > 0.5) == (y > 0.5)))
    dz = np.where((p > 1e-7) & (p < 1 - 1e-7), pc - y, 0.0) / B          # d loss / d logit, logit = (fm + dnn) / 2
    dhalf = 0.5 * dz
    # tower
    g = dhalf[:, None] @ out[0].T                                        # [B, h_L]
    new_out = (out[0] - lr * (acts[-1].T @ dhalf[:, None]), out[1] - lr * np.sum(dhalf))
    new_hidden = [None] * len(hidden)
    for l in range(len(hidden) - 1, -1, -1):
        k, b = hidden[l]
        m = (acts[l + 1] > 0).astype(np.float64)
        if drop_masks is not None:
            m = m * np.max(drop_masks[l])                                # kept elements carry the 1/(1-rate) scale
        g = g * m
        gin = g @ k.T
        new_hidden[l] = (k - lr * (acts[l].T @ g), b - lr * np.sum(g, axis=0))
        g = gin
    dx = g
    # FM
    vsq = np.sum(fm_v ** 2, axis=1)
    dx = dx + dhalf[:, None] * (fm_w1[:, 0][None, :] + xv @ fm_v.T - x * vsq[None, :])
    g_w1 = x.T @ dhalf[:, None] + 2 * reg_w * fm_w1
    g_v = x.T @ (dhalf[:, None] * xv) - fm_v * ((x ** 2).T @ dhalf)[:, None] + 2 * reg_v * fm_v
    new_w0, new_w1, new_v = fm_w0 - lr * np.sum(dhalf), fm_w1 - lr * g_w1, fm_v - lr * g_v
    # embedding rows
    new_tables = [t.copy() for t in tables]
    off = D
    for f in range(F):
        kdim = tables[f].shape[1]
        np.add.at(new_tables[f], ids[:, f], -lr * dx[:, off:off + kdim])
        off += kdim
    return (new_tables, new_w0, new_w1, new_v, new_hidden, new_out), bce, correct


def fm_train_step(x, fm_w0, fm_w1, fm_v, y, lr, reg_w, reg_v):
    """One step of keras SGD(lr) on FM (DL/model/fm.py:14-23: p = sigmoid(FMLayer(x))): mean clipped BCE + the l2 regularisers
    of FMLayer.build.  PARITY UNPINNED against the reference like deepfm_train_step; gradients pinned by finite differences.
    -> ((w0, w1, v) updated, mean BCE before the update, number of correct predictions)."""
    B = x.shape[0]
    xv = x @ fm_v
    p = sigmoid(fm_layer(x, fm_w0, fm_w1, fm_v)[:, 0])
    pc = np.clip(p, 1e-7, 1 - 1e-7)
    bce = float(np.mean(-(y * np.log(pc) + (1 - y) * np.log(1 - pc))))
    dz = np.where((p > 1e-7) & (p < 1 - 1e-7), pc - y, 0.0) / B
    g_w1 = x.T @ dz[:, None] + 2 * reg_w * fm_w1
    g_v = x.T @ (dz[:, None] * xv) - fm_v * ((x ** 2).T @ dz)[:, None] + 2 * reg_v * fm_v
    return (fm_w0 - lr * np.sum(dz), fm_w1 - lr * g_w1, fm_v - lr * g_v), bce, int(np.sum((p > 0.5) == (y > 0.5)))


# --------------------------------------------------------------------------- 8(f) rank 2: Hash (DL/layer/utils.py:21-56)
# tf.strings.to_hash_bucket_fast is FarmHash Fingerprint64 (farmhashna::Hash64, google/farmhash, not vendored by the reference:
# TensorFlow links it) modulo num_buckets.  Restated from the published algorithm.  PIN: the example in TensorFlow's own
# documentation of the op, to_hash_bucket_fast(["Hello", "TensorFlow", "2.x"], 3) == [0, 2, 2], which exercises the 4-7, 8-16 and
# 1-3 byte branches (every decimal int32 / most int64 ids), and Fingerprint64("") == k2.  The 17-32 and 33-64 byte branches follow
# the same source but have no vector here: "parity unpinned" for ids of 17+ digits and long strings.
_M64 = (1 << 64) - 1
_K0, _K1, _K2 = 0xc3a5c85c97cb3127, 0xb492b66fbe98f273, 0x9ae16a3b2f90404f


def _rot(v, s):
    return ((v >> s) | (v << (64 - s))) & _M64 if s else v


def _hl16(u, v, mul):
    a = ((u ^ v) * mul) & _M64
    a ^= a >> 47
    b = ((v ^ a) * mul) & _M64
    b ^= b >> 47
    return (b * mul) & _M64


def farm_fingerprint64(s):
    """farmhashna::Hash64 for len(s) <= 64 bytes (HashLen0to16 / HashLen17to32 / HashLen33to64)."""
    n = len(s)
    f64 = lambda i: int.from_bytes(s[i:i + 8], "little")
    f32 = lambda i: int.from_bytes(s[i:i + 4], "little")
    mul = (_K2 + 2 * n) & _M64
    if n == 0:
        return _K2
    if n < 4:
        y = (s[0] + (s[n >> 1] << 8)) & 0xffffffff
        z = (n + (s[n - 1] << 2)) & 0xffffffff
        v = ((y * _K2) & _M64) ^ ((z * _K0) & _M64)
        return ((v ^ (v >> 47)) * _K2) & _M64
    if n < 8:
        return _hl16((n + (f32(0) << 3)) & _M64, f32(n - 4), mul)
    if n <= 16:
        a, b = (f64(0) + _K2) & _M64, f64(n - 8)
        return _hl16((_rot(b, 37) * mul + a) & _M64, ((_rot(a, 25) + b) * mul) & _M64, mul)
    if n <= 32:
        a, b, c, d = (f64(0) * _K1) & _M64, f64(8), (f64(n - 8) * mul) & _M64, (f64(n - 16) * _K2) & _M64
        return _hl16((_rot((a + b) & _M64, 43) + _rot(c, 30) + d) & _M64, (a + _rot((b + _K2) & _M64, 18) + c) & _M64, mul)
    if n <= 64:
        a, b, c, d = (f64(0) * _K2) & _M64, f64(8), (f64(n - 8) * mul) & _M64, (f64(n - 16) * _K2) & _M64
        y = (_rot((a + b) & _M64, 43) + _rot(c, 30) + d) & _M64
        z = _hl16(y, (a + _rot((b + _K2) & _M64, 18) + c) & _M64, mul)
        e, f = (f64(16) * mul) & _M64, f64(24)
        g, h = ((y + f64(n - 32)) * mul) & _M64, ((z + f64(n - 24)) * mul) & _M64
        return _hl16((_rot((e + f) & _M64, 43) + _rot(g, 30) + h) & _M64, (e + _rot((f + a) & _M64, 18) + g) & _M64, mul)
    raise NotImplementedError("strings over 64 bytes")


def hash_layer(x, num_buckets, mask_zero=False):
    """Hash.call without a vocabulary file, DL/layer/utils.py:35-56: integers go through tf.as_string (decimal), strings as they
    are; bucket = Fingerprint64 % (num_buckets - mask_zero); with mask_zero the buckets shift up by one and the value "0" maps
    to 0.  -> int64 array of x's shape."""
    a = np.asarray(x)
    flat = a.reshape(-1)
    nb = num_buckets - 1 if mask_zero else num_buckets
    out = np.empty(flat.shape, np.int64)
    for i, v in enumerate(flat):
        s = v if isinstance(v, bytes) else (v.encode() if isinstance(v, str) else str(int(v)).encode())
        h = farm_fingerprint64(s) % nb
        out[i] = (h + 1) * (s != b"0") if mask_zero else h
    return out.reshape(a.shape)


def dcn_loss(dense, ids, tables, cross_ws, cross_bs, hidden, out, final, y, reg_w, reg_b, drop_masks=None):
    """What keras minimises for DCN (DL/model/dcn.py:15-34 under compile_fit): mean clipped BCE + CrossLayer's l2 regularisers
    (interaction.py:59-70: reg_w sum w_l^2 + reg_b sum b_l^2 per layer).  -> (total, bce, p)."""
    ids = ids_to_int(ids)
    x = np.concatenate([np.asarray(dense, np.float64)] + [tables[f][ids[:, f]] for f in range(len(tables))], axis=1)
    xl = x
    for w, b in zip(cross_ws, cross_bs):
        xl = x * (xl @ w) + b[:, 0][None, :] + xl
    h = x
    for l, (k, bb) in enumerate(hidden):
        h = np.maximum(h @ k + bb, 0.0)
        if drop_masks is not None:
            h = h * drop_masks[l]
    d = h @ out[0] + out[1]
    p = sigmoid(np.concatenate([xl, d], axis=1) @ final[0] + final[1])[:, 0]
    pc = np.clip(p, 1e-7, 1 - 1e-7)
    bce = float(np.mean(-(y * np.log(pc) + (1 - y) * np.log(1 - pc))))
    reg = sum(reg_w * float(np.sum(w ** 2)) for w in cross_ws) + sum(reg_b * float(np.sum(b ** 2)) for b in cross_bs)
    return bce + reg, bce, p


def dcn_train_step(dense, ids, tables, cross_ws, cross_bs, hidden, out, final, y, lr, reg_w, reg_b, drop_masks=None):
    """One step of keras SGD(lr) on `dcn_loss`; analytic gradients (PARITY UNPINNED against the reference like deepfm_train_step,
    pinned by finite differences).  Cross layer l: x_{l+1} = x_0 s_l + b_l + x_l with s_l = x_l . w_l, so with g = dL/dx_{l+1} and
    c = g . x_0:  dw_l = sum_b c x_l,  db_l = sum_b g,  dL/dx_l = g + c w_l,  dL/dx_0 += g s_l.
    -> ((tables, cross_ws, cross_bs, hidden, out, final) updated, mean BCE before the update, number of correct predictions)."""
    ids = ids_to_int(ids)
    B, D, F = ids.shape[0], np.asarray(dense).shape[1], len(tables)
    x = np.concatenate([np.asarray(dense, np.float64)] + [tables[f][ids[:, f]] for f in range(F)], axis=1)
    n = x.shape[1]
    xs, ss = [x], []
    for w, b in zip(cross_ws, cross_bs):
        s = xs[-1] @ w                                                   # [B, 1]
        ss.append(s)
        xs.append(x * s + b[:, 0][None, :] + xs[-1])
    acts = [x]
    for l, (k, bb) in enumerate(hidden):
        h = np.maximum(acts[-1] @ k + bb, 0.0)
        if drop_masks is not None:
            h = h * drop_masks[l]
        acts.append(h)
    d = acts[-1] @ out[0] + out[1]                                       # [B, 1]
    fk, fb = final
    p = sigmoid(xs[-1] @ fk[:n] + d * fk[n] + fb)[:, 0]
    pc = np.clip(p, 1e-7, 1 - 1e-7)
    bce = float(np.mean(-(y * np.log(pc) + (1 - y) * np.log(1 - pc))))
    correct = int(np.sum((p > 0.5) == (y > 0.5)))
    dz = (np.where((p > 1e-7) & (p < 1 - 1e-7), pc - y, 0.0) / B)[:, None]   # [B, 1]
    new_final = (fk - lr * np.concatenate([xs[-1].T @ dz, d.T @ dz], axis=0), fb - lr * np.sum(dz))
    # tower, from d dnn = dz * fk[n]
    gd = dz * fk[n, 0]
    g = gd @ out[0].T
    new_out = (out[0] - lr * (acts[-1].T @ gd), out[1] - lr * np.sum(gd))
    new_hidden = [None] * len(hidden)
    for l in range(len(hidden) - 1, -1, -1):
        k, bb = hidden[l]
        m = (acts[l + 1] > 0).astype(np.float64)
        if drop_masks is not None:
            m = m * np.max(drop_masks[l])
        g = g * m
        gin = g @ k.T
        new_hidden[l] = (k - lr * (acts[l].T @ g), bb - lr * np.sum(g, axis=0))
        g = gin
    dx = g
    # cross layers, last to first
    g = dz @ fk[:n].T                                                    # dL/dx_L
    dx0 = np.zeros_like(x)
    new_ws, new_bs = [None] * len(cross_ws), [None] * len(cross_bs)
    for l in range(len(cross_ws) - 1, -1, -1):
        c = np.sum(g * x, axis=1, keepdims=True)
        new_ws[l] = cross_ws[l] - lr * (xs[l].T @ c + 2 * reg_w * cross_ws[l])
        new_bs[l] = cross_bs[l] - lr * (np.sum(g, axis=0)[:, None] + 2 * reg_b * cross_bs[l])
        dx0 = dx0 + g * ss[l]
        g = g + c @ cross_ws[l].T
    dx = dx + g + dx0
    new_tables = [t.copy() for t in tables]
    off = D
    for f in range(F):
        kdim = tables[f].shape[1]
        np.add.at(new_tables[f], ids[:, f], -lr * dx[:, off:off + kdim])
        off += kdim
    return (new_tables, new_ws, new_bs, new_hidden, new_out, new_final), bce, correct


# keras' binary_crossentropy clips in the output's dtype: epsilon_ = float32(1e-7) and the upper bound is the float32 value of
# 1 - epsilon_ = 0.99999988 (not 1 - 1e-7).  It matters for PNN, whose raw output reaches the clip on a fair share of samples.
_BCE_LO = float(np.float32(1e-7))
_BCE_HI = float(np.float32(1.0) - np.float32(1e-7))


def pnn_loss(ids, tables, outer_w, hidden, out, y, mode, drop_masks=None):
    """What the reference's own PNN training loop minimises (DL/model/pnn.py:74-85): tf.reduce_mean(binary_crossentropy(y, pre))
    with pre = model(X) -- PNN.call ends WITHOUT a sigmoid (:52-53), so keras' BCE clips the raw output to [1e-7, 1 - 1e-7] and
    takes logs of that; tape.gradient(loss, ...) leaves the layers' regularisers out.  -> (bce, pre)."""
    ids = ids_to_int(ids)
    z = np.stack([tables[f][ids[:, f]] for f in range(len(tables))], axis=1)
    feats = [z.reshape(z.shape[0], -1)]
    if mode in ("inner", "both"):
        feats.append(inner_product_layer(z))
    if mode in ("outer", "both"):
        feats.append(outer_product_layer(z, outer_w))
    h = np.concatenate(feats, axis=1)
    for l, (k, bb) in enumerate(hidden):
        h = np.maximum(h @ k + bb, 0.0)
        if drop_masks is not None:
            h = h * drop_masks[l]
    pre = (h @ out[0] + out[1])[:, 0]
    pc = np.clip(pre, _BCE_LO, _BCE_HI)
    return float(np.mean(-(y * np.log(pc) + (1 - y) * np.log(1 - pc)))), pre


def pnn_train_step(ids, tables, outer_w, hidden, out, y, lr, mode, drop_masks=None):
    """One step of SGD(lr) on `pnn_loss` (PARITY UNPINNED against the reference, pinned by finite differences).  With pair p = (i, j),
    i < j:  inner_p = <e_i, e_j>,  outer_p = sum_a e_j[a] sum_c W[a, p, c] e_i[c]; for upstream gradients gi_p, go_p:
    de_i += gi e_j + go W_p^T e_j,  de_j += gi e_i + go W_p e_i,  dW[a, p, c] = sum_b go e_j[a] e_i[c].  The clipped BCE on a raw
    output has gradient (pre - y) / (pre (1 - pre)) / B inside (1e-7, 1 - 1e-7) and zero outside.
    -> ((tables, outer_w, hidden, out) updated, loss before the update, number of correct predictions (pre > 0.5))."""
    ids = ids_to_int(ids)
    B, F = ids.shape[0], len(tables)
    K = tables[0].shape[1]
    z = np.stack([tables[f][ids[:, f]] for f in range(F)], axis=1)       # [B, F, K]
    row, col = pair_index(F)
    P = len(row)
    feats = [z.reshape(B, -1)]
    if mode in ("inner", "both"):
        feats.append(inner_product_layer(z))
    if mode in ("outer", "both"):
        feats.append(outer_product_layer(z, outer_w))
    acts = [np.concatenate(feats, axis=1)]
    for l, (k, bb) in enumerate(hidden):
        h = np.maximum(acts[-1] @ k + bb, 0.0)
        if drop_masks is not None:
            h = h * drop_masks[l]
        acts.append(h)
    pre = (acts[-1] @ out[0] + out[1])[:, 0]
    pc = np.clip(pre, _BCE_LO, _BCE_HI)
    loss = float(np.mean(-(y * np.log(pc) + (1 - y) * np.log(1 - pc))))
    correct = int(np.sum((pre > 0.5) == (y > 0.5)))
    dz = (np.where((pre > _BCE_LO) & (pre < _BCE_HI), (pc - y) / (pc * (1 - pc)), 0.0) / B)[:, None]
    g = dz @ out[0].T
    new_out = (out[0] - lr * (acts[-1].T @ dz), out[1] - lr * np.sum(dz))
    new_hidden = [None] * len(hidden)
    for l in range(len(hidden) - 1, -1, -1):
        k, bb = hidden[l]
        m = (acts[l + 1] > 0).astype(np.float64)
        if drop_masks is not None:
            m = m * np.max(drop_masks[l])
        g = g * m
        gin = g @ k.T
        new_hidden[l] = (k - lr * (acts[l].T @ g), bb - lr * np.sum(g, axis=0))
        g = gin
    de = g[:, :F * K].reshape(B, F, K).copy()
    colp = F * K
    new_w = outer_w
    ei, ej = z[:, row, :], z[:, col, :]                                  # [B, P, K]
    if mode in ("inner", "both"):
        gi = g[:, colp:colp + P]
        colp += P
        np.add.at(de, (slice(None), row), gi[:, :, None] * ej)
        np.add.at(de, (slice(None), col), gi[:, :, None] * ei)
    if mode in ("outer", "both"):
        go = g[:, colp:colp + P]
        # W[a, p, c]: t_i[b, p, c] = sum_a W[a, p, c] e_j[b, p, a];  t_j[b, p, a] = sum_c W[a, p, c] e_i[b, p, c]
        t_i = np.einsum("apc,bpa->bpc", outer_w, ej)
        t_j = np.einsum("apc,bpc->bpa", outer_w, ei)
        np.add.at(de, (slice(None), row), go[:, :, None] * t_i)
        np.add.at(de, (slice(None), col), go[:, :, None] * t_j)
        new_w = outer_w - lr * np.einsum("bp,bpa,bpc->apc", go, ej, ei)
    new_tables = [t.copy() for t in tables]
    for f in range(F):
        np.add.at(new_tables[f], ids[:, f], -lr * de[:, f, :])
    return (new_tables, new_w, new_hidden, new_out), loss, correct


def _widedeep_wide_input(dense, ids, widths):
    """The wide input as the reference's loader lays it out for Wide&Deep (DL/utils/dataset.py:53-58): the dense block, the
    get_dummies frame -- which starts with the dense block AGAIN -- and one 0/1 block per sparse field (width = labels seen)."""
    B = ids.shape[0]
    blocks = [np.asarray(dense, np.float64), np.asarray(dense, np.float64)]
    for f, wd in enumerate(widths):
        oh = np.zeros((B, wd))
        ok = ids[:, f] < wd
        oh[np.arange(B)[ok], ids[ok, f]] = 1.0
        blocks.append(oh)
    return np.concatenate(blocks, axis=1)


def widedeep_loss(dense, ids, tables, wide_w0, wide_w, widths, hidden, out, y, reg_w, drop_masks=None):
    """keras' objective for WideDeep (DL/model/wideDeep.py:14-34 under compile_fit): mean clipped BCE of
    sigmoid(0.5 (wide + deep)) + WideLayer's l2(1e-4) on w (interaction.py:19-22).  -> (total, bce, p)."""
    ids = ids_to_int(ids)
    xw = _widedeep_wide_input(dense, ids, widths)
    wide = (wide_w0 + xw @ wide_w)[:, 0]
    h = np.concatenate([tables[f][ids[:, f]] for f in range(len(tables))], axis=1)
    for l, (k, bb) in enumerate(hidden):
        h = np.maximum(h @ k + bb, 0.0)
        if drop_masks is not None:
            h = h * drop_masks[l]
    deep = (h @ out[0] + out[1])[:, 0]
    p = sigmoid(0.5 * (wide + deep))
    pc = np.clip(p, 1e-7, 1 - 1e-7)
    bce = float(np.mean(-(y * np.log(pc) + (1 - y) * np.log(1 - pc))))
    return bce + reg_w * float(np.sum(wide_w ** 2)), bce, p


def widedeep_train_step(dense, ids, tables, wide_w0, wide_w, widths, hidden, out, y, lr, reg_w, drop_masks=None):
    """One step of keras SGD(lr) on `widedeep_loss` (PARITY UNPINNED against the reference, pinned by finite differences).
    -> ((tables, wide_w0, wide_w, hidden, out) updated, mean BCE before the update, number of correct predictions)."""
    ids = ids_to_int(ids)
    B, F = ids.shape[0], len(tables)
    xw = _widedeep_wide_input(dense, ids, widths)
    wide = (wide_w0 + xw @ wide_w)[:, 0]
    acts = [np.concatenate([tables[f][ids[:, f]] for f in range(F)], axis=1)]
    for l, (k, bb) in enumerate(hidden):
        h = np.maximum(acts[-1] @ k + bb, 0.0)
        if drop_masks is not None:
            h = h * drop_masks[l]
        acts.append(h)
    deep = (acts[-1] @ out[0] + out[1])[:, 0]
    p = sigmoid(0.5 * (wide + deep))
    pc = np.clip(p, 1e-7, 1 - 1e-7)
    bce = float(np.mean(-(y * np.log(pc) + (1 - y) * np.log(1 - pc))))
    correct = int(np.sum((p > 0.5) == (y > 0.5)))
    dh = (0.5 * np.where((p > 1e-7) & (p < 1 - 1e-7), pc - y, 0.0) / B)[:, None]
    new_w0 = wide_w0 - lr * np.sum(dh)
    new_w = wide_w - lr * (xw.T @ dh + 2 * reg_w * wide_w)
    g = dh @ out[0].T
    new_out = (out[0] - lr * (acts[-1].T @ dh), out[1] - lr * np.sum(dh))
    new_hidden = [None] * len(hidden)
    for l in range(len(hidden) - 1, -1, -1):
        k, bb = hidden[l]
        m = (acts[l + 1] > 0).astype(np.float64)
        if drop_masks is not None:
            m = m * np.max(drop_masks[l])
        g = g * m
        gin = g @ k.T
        new_hidden[l] = (k - lr * (acts[l].T @ g), bb - lr * np.sum(g, axis=0))
        g = gin
    new_tables = [t.copy() for t in tables]
    off = 0
    for f in range(F):
        kdim = tables[f].shape[1]
        np.add.at(new_tables[f], ids[:, f], -lr * g[:, off:off + kdim])
        off += kdim
    return (new_tables, new_w0, new_w, new_hidden, new_out), bce, correct


def deepcrossing_train_step(dense, ids, tables, res_layers, final, y, lr):
    """One step of keras SGD(lr) on the mean clipped BCE of DeepCrossing (DL/model/deepCrossing.py:14-32 under compile_fit; no
    regularisers, no dropout).  res_layers: list of (hidden pairs, out pair).  PARITY UNPINNED against the reference, pinned by
    finite differences of `deepcrossing_model`'s loss.
    -> ((tables, res_layers, final) updated, mean BCE before the update, number of correct predictions)."""
    ids = ids_to_int(ids)
    B, D, F = ids.shape[0], np.asarray(dense).shape[1], len(tables)
    x = np.concatenate([np.asarray(dense, np.float64)] + [tables[f][ids[:, f]] for f in range(F)], axis=1)
    xs, hs = [x], []
    for hidden, out in res_layers:
        acts = [xs[-1]]
        for k, b in hidden:
            acts.append(np.maximum(acts[-1] @ k + b, 0.0))
        hs.append(acts)
        xs.append(np.maximum(xs[-1] + acts[-1] @ out[0] + out[1], 0.0))
    p = sigmoid(xs[-1] @ final[0] + final[1])[:, 0]
    pc = np.clip(p, 1e-7, 1 - 1e-7)
    bce = float(np.mean(-(y * np.log(pc) + (1 - y) * np.log(1 - pc))))
    correct = int(np.sum((p > 0.5) == (y > 0.5)))
    dz = (np.where((p > 1e-7) & (p < 1 - 1e-7), pc - y, 0.0) / B)[:, None]
    new_final = (final[0] - lr * (xs[-1].T @ dz), final[1] - lr * np.sum(dz))
    g = dz @ final[0].T
    new_res = [None] * len(res_layers)
    for r in range(len(res_layers) - 1, -1, -1):
        hidden, out = res_layers[r]
        acts = hs[r]
        gm = g * (xs[r + 1] > 0)
        new_out = (out[0] - lr * (acts[-1].T @ gm), out[1] - lr * np.sum(gm, axis=0))
        gh = gm @ out[0].T
        new_hidden = [None] * len(hidden)
        for l in range(len(hidden) - 1, -1, -1):
            k, b = hidden[l]
            gh = gh * (acts[l + 1] > 0)
            gin = gh @ k.T
            new_hidden[l] = (k - lr * (acts[l].T @ gh), b - lr * np.sum(gh, axis=0))
            gh = gin
        new_res[r] = (new_hidden, new_out)
        g = gm + gh
    new_tables = [t.copy() for t in tables]
    off = D
    for f in range(F):
        kdim = tables[f].shape[1]
        np.add.at(new_tables[f], ids[:, f], -lr * g[:, off:off + kdim])
        off += kdim
    return (new_tables, new_res, new_final), bce, correct


def nfm_train_forward(dense, ids, tables, bn, hidden, out, final, drop_masks=None, eps=1e-3):
    """NFM.call in TRAINING mode (DL/model/nfm.py:24-34 as model.fit runs it): keras BatchNormalization normalises with the
    batch's own mean and (biased) variance.  -> (p, cache)."""
    ids = ids_to_int(ids)
    emb = np.stack([tables[f][ids[:, f]] for f in range(len(tables))], axis=1)          # [B, F, k]
    S = np.sum(emb, axis=1)
    x = np.concatenate([np.asarray(dense, np.float64), 0.5 * (S ** 2 - np.sum(emb ** 2, axis=1))], axis=1)
    mu, var = np.mean(x, axis=0), np.var(x, axis=0)
    xhat = (x - mu) / np.sqrt(var + eps)
    acts = [xhat * bn[2] + bn[3]]
    for l, (k, b) in enumerate(hidden):
        h = np.maximum(acts[-1] @ k + b, 0.0)
        if drop_masks is not None:
            h = h * drop_masks[l]
        acts.append(h)
    d = acts[-1] @ out[0] + out[1]
    p = sigmoid(d @ final[0] + final[1])[:, 0]
    return p, (emb, S, x, mu, var, xhat, acts, d)


def nfm_train_step(dense, ids, tables, bn, hidden, out, final, y, lr, drop_masks=None, momentum=0.99, eps=1e-3):
    """One step of keras SGD(lr) on the mean clipped BCE of NFM in training mode (PARITY UNPINNED against the reference, pinned by
    finite differences THROUGH the batch statistics).  bn = (moving_mean, moving_var, gamma, beta); the moving statistics move by
    keras' rule  m <- momentum m + (1 - momentum) batch_stat  (biased batch variance: the non-fused 2-D path).
    -> ((tables, bn, hidden, out, final) updated, mean BCE before the update, number of correct predictions)."""
    ids = ids_to_int(ids)
    B, D, F = ids.shape[0], np.asarray(dense).shape[1], len(tables)
    p, (emb, S, x, mu, var, xhat, acts, d) = nfm_train_forward(dense, ids, tables, bn, hidden, out, final, drop_masks, eps)
    pc = np.clip(p, 1e-7, 1 - 1e-7)
    bce = float(np.mean(-(y * np.log(pc) + (1 - y) * np.log(1 - pc))))
    correct = int(np.sum((p > 0.5) == (y > 0.5)))
    dz = (np.where((p > 1e-7) & (p < 1 - 1e-7), pc - y, 0.0) / B)[:, None]
    new_final = (final[0] - lr * (d.T @ dz), final[1] - lr * np.sum(dz))
    gd = dz @ final[0].T                                                               # [B, od]
    g = gd @ out[0].T
    new_out = (out[0] - lr * (acts[-1].T @ gd), out[1] - lr * np.sum(gd, axis=0))
    new_hidden = [None] * len(hidden)
    for l in range(len(hidden) - 1, -1, -1):
        k, b = hidden[l]
        m = (acts[l + 1] > 0).astype(np.float64)
        if drop_masks is not None:
            m = m * np.max(drop_masks[l])
        g = g * m
        gin = g @ k.T
        new_hidden[l] = (k - lr * (acts[l].T @ g), b - lr * np.sum(g, axis=0))
        g = gin
    # BatchNormalization backward (batch statistics are functions of x)
    dgamma, dbeta = np.sum(g * xhat, axis=0), np.sum(g, axis=0)
    rstd = 1.0 / np.sqrt(var + eps)
    gx = bn[2] * rstd * (g - dbeta / B - xhat * dgamma / B)
    new_bn = (momentum * bn[0] + (1 - momentum) * mu, momentum * bn[1] + (1 - momentum) * var, bn[2] - lr * dgamma, bn[3] - lr * dbeta)
    # Bi-Interaction backward: d bi_c / d e_fc = S_c - e_fc
    de = gx[:, None, D:] * (S[:, None, :] - emb)
    new_tables = [t.copy() for t in tables]
    for f in range(F):
        np.add.at(new_tables[f], ids[:, f], -lr * de[:, f, :])
    return (new_tables, new_bn, new_hidden, new_out, new_final), bce, correct


def afm_train_step(ids, tables, mode, out, y, lr, n_sigmoid=2):
    """One step of keras SGD(lr) on the mean clipped BCE of AFM (DL/model/afm.py: sigmoid of AFMLayer's already-sigmoid output,
    n_sigmoid = 2; AFMLayer alone: 1).  In 'att' mode the attention weights are identically 1 (softmax over a size-1 axis,
    interaction.py:316), so AttentionLayer's two Dense layers receive zero gradient and do not move; the pooled vector is the sum
    over pairs.  PARITY UNPINNED against the reference, pinned by finite differences of `afm_layer`'s loss.
    -> ((tables, out) updated, mean BCE before the update, number of correct predictions)."""
    ids = ids_to_int(ids)
    B, F = ids.shape[0], len(tables)
    e = np.stack([tables[f][ids[:, f]] for f in range(F)], axis=1)                     # [B, F, k]
    row, col = pair_index(F)
    pairs = e[:, row, :] * e[:, col, :]                                                # [B, P, k]
    P = len(row)
    if mode == "avg":
        x = pairs.mean(axis=1)
    elif mode == "max":
        x = pairs.max(axis=1)
    else:
        x = pairs.sum(axis=1)
    q = sigmoid(x @ out[0] + out[1])[:, 0]
    p = sigmoid(q) if n_sigmoid == 2 else q
    pc = np.clip(p, 1e-7, 1 - 1e-7)
    bce = float(np.mean(-(y * np.log(pc) + (1 - y) * np.log(1 - pc))))
    correct = int(np.sum((p > 0.5) == (y > 0.5)))
    if n_sigmoid == 2:
        dq = np.where((p > 1e-7) & (p < 1 - 1e-7), pc - y, 0.0) / B                    # d/d(logit of the outer sigmoid) = d/dq
        dz = (dq * q * (1 - q))[:, None]
    else:
        dz = (np.where((p > 1e-7) & (p < 1 - 1e-7), pc - y, 0.0) / B)[:, None]
    new_out = (out[0] - lr * (x.T @ dz), out[1] - lr * np.sum(dz))
    gx = dz @ out[0].T                                                                 # [B, k]
    if mode == "max":
        arg = pairs.argmax(axis=1)                                                     # [B, k]
        gp = np.zeros_like(pairs)
        bi, ci = np.meshgrid(np.arange(B), np.arange(pairs.shape[2]), indexing="ij")
        gp[bi, arg, ci] = gx
    else:
        gp = np.broadcast_to(gx[:, None, :] / (P if mode == "avg" else 1.0), pairs.shape)
    de = np.zeros_like(e)
    np.add.at(de, (slice(None), row), gp * e[:, col, :])
    np.add.at(de, (slice(None), col), gp * e[:, row, :])
    new_tables = [t.copy() for t in tables]
    for f in range(F):
        np.add.at(new_tables[f], ids[:, f], -lr * de[:, f, :])
    return (new_tables, new_out), bce, correct


def ffm_train_step(inputs, vocab_sizes, w0, w, v, y, lr, reg_w, reg_v, n_dense=13):
    """One step of keras SGD(lr) on FFM (DL/model/ffm.py:14-23: p = sigmoid(FFMLayer(x))): mean clipped BCE + l2(w_reg)(w) +
    l2(v_reg)(v) (FFMLayer.build, DL/layer/interaction.py:131-138).  With ff = tensordot(x, v) [B, field, k] and S = sum over
    fields, inter = 0.5 (|S|^2 - |ff|^2), so d inter / d v[i, f, :] = x_i (S - ff[f]).  PARITY UNPINNED against the reference like
    deepfm_train_step; gradients pinned by finite differences of `ffm_layer`'s loss.
    -> ((w0, w, v) updated, mean BCE before the update, number of correct predictions)."""
    x = onehot_expand(inputs, vocab_sizes, n_dense, dtype=v.dtype)
    B = x.shape[0]
    p = sigmoid(ffm_layer(inputs, vocab_sizes, w0, w, v, n_dense)[:, 0])
    pc = np.clip(p, _BCE_LO, _BCE_HI)                                                  # keras clips in fp32
    bce = float(np.mean(-(y * np.log(pc) + (1 - y) * np.log(1 - pc))))
    dz = np.where((p > _BCE_LO) & (p < _BCE_HI), pc - y, 0.0) / B
    ff = np.tensordot(x, v, axes=1)                                                    # [B, field, k]
    G = ff.sum(axis=1, keepdims=True) - ff                                             # d inter / d ff
    g_v = np.tensordot(x * dz[:, None], G, axes=([0], [0])) + 2 * reg_v * v            # [n_feat, field, k]
    g_w = x.T @ dz[:, None] + 2 * reg_w * w
    return (w0 - lr * np.sum(dz), w - lr * g_w, v - lr * g_v), bce, int(np.sum((p > 0.5) == (y > 0.5)))


# --------------------------------------------------------------------------- 8(f) rank 3: DIN under compile_fit
def din_train_forward(dense_inputs, other_ids, hist_ids, cand_ids, other_tables, seq_table, att_dense, att_out, bn, tower, out):
    """DIN.call (DL/model/din.py:56-95) in TRAINING mode for one behaviour feature: Attention (DL/layer/interaction.py:369-406,
    Dense + PReLU with alpha [T, h]), BatchNormalization on the batch's own statistics (biased variance, eps 1e-3), the PReLU
    tower (alpha [h]); dnn_dropout = 0 (the reference's default).  -> (p [B], cache)."""
    hist = ids_to_int(hist_ids)
    cand = ids_to_int(cand_ids).reshape(-1)
    other_ids = ids_to_int(other_ids) if len(other_tables) else None
    k = seq_table[hist]                                                                # [B, T, K]
    q = seq_table[cand]                                                                # [B, K]
    qt = np.broadcast_to(q[:, None, :], k.shape)
    xs = [np.concatenate([qt, k, qt - k, qt * k], axis=-1)]                            # [B, T, 4K]
    pre = []
    for kernel, bias, alpha in att_dense:
        pre.append(xs[-1] @ kernel + bias)
        xs.append(prelu(pre[-1], alpha))
    s = (xs[-1] @ att_out[0] + att_out[1])[..., 0]                                     # [B, T]
    mask = hist != 0
    pw = softmax_last(np.where(mask, s, s.dtype.type(MASK_FILL)))
    att = np.einsum("bt,btk->bk", pw, k)
    parts = [att, q] + [other_tables[i][other_ids[:, i]] for i in range(len(other_tables))]
    if dense_inputs is not None and np.asarray(dense_inputs).shape[1] > 0:
        parts.append(np.asarray(dense_inputs, dtype=k.dtype))
    emb = np.concatenate(parts, axis=-1)
    gamma, beta = bn[0], bn[1]
    mu, var = emb.mean(axis=0), emb.var(axis=0)
    rstd = 1.0 / np.sqrt(var + 1e-3)
    xhat = (emb - mu) * rstd
    es = [gamma * xhat + beta]
    tpre = []
    for kernel, bias, alpha in tower:
        tpre.append(es[-1] @ kernel + bias)
        es.append(prelu(tpre[-1], alpha))
    p = sigmoid(es[-1] @ out[0] + out[1])[:, 0]
    return p, dict(hist=hist, cand=cand, other_ids=other_ids, k=k, q=q, xs=xs, pre=pre, mask=mask, pw=pw, emb=emb, mu=mu, var=var,
                   rstd=rstd, xhat=xhat, es=es, tpre=tpre)


def din_loss(p, y):
    pc = np.clip(p, _BCE_LO, _BCE_HI)
    return float(np.mean(-(y * np.log(pc) + (1 - y) * np.log(1 - pc))))


def din_train_step(dense_inputs, other_ids, hist_ids, cand_ids, other_tables, seq_table, att_dense, att_out, bn, tower, out, y,
                   lr, momentum=0.99):
    """One step of keras SGD(lr) on DIN's mean clipped BCE (compile_fit, DL/utils/compile_fit.py:9-15, on DL/model/din.py).
    Every variable moves: the behaviour table (history rows AND the candidate's row: one table, din.py:76-79), the other tables,
    the attention unit's Dense / PReLU alpha [T, h] / out layer, BatchNormalization's gamma / beta (and its moving statistics by
    momentum 0.99), the tower's Dense / PReLU alpha [h], the out layer.  Masked positions (history id 0) get no score gradient
    (tf.where) but still feed the value sum when the whole history is masked (uniform softmax).  PARITY UNPINNED against the
    reference; gradients pinned by finite differences of `din_train_forward`'s loss.
    -> (dict of updated parameters, mean BCE before the update, number of correct predictions)."""
    p, c = din_train_forward(dense_inputs, other_ids, hist_ids, cand_ids, other_tables, seq_table, att_dense, att_out, bn, tower, out)
    B = p.shape[0]
    K = seq_table.shape[1]
    bce = din_loss(p, y)
    pc = np.clip(p, _BCE_LO, _BCE_HI)
    dz = (np.where((p > _BCE_LO) & (p < _BCE_HI), pc - y, 0.0) / B)[:, None]
    new_out = (out[0] - lr * (c["es"][-1].T @ dz), out[1] - lr * np.sum(dz))
    g = dz @ out[0].T
    new_tower = [None] * len(tower)
    for l in range(len(tower) - 1, -1, -1):
        kernel, bias, alpha = tower[l]
        P = c["tpre"][l]
        g_alpha = np.sum(g * np.where(P < 0, P, 0.0), axis=0)
        gp = g * np.where(P > 0, 1.0, alpha)
        new_tower[l] = (kernel - lr * (c["es"][l].T @ gp), bias - lr * gp.sum(axis=0), alpha - lr * g_alpha)
        g = gp @ kernel.T
    gamma, beta, mean, var = bn
    xhat = c["xhat"]
    g_gamma, g_beta = np.sum(g * xhat, axis=0), np.sum(g, axis=0)
    gx = gamma * c["rstd"] * (g - g_beta / B - xhat * g_gamma / B)
    new_bn = (gamma - lr * g_gamma, beta - lr * g_beta, momentum * mean + (1 - momentum) * c["mu"],
              momentum * var + (1 - momentum) * c["var"])
    g_att, g_q = gx[:, :K], gx[:, K:2 * K].copy()
    k, q, pw = c["k"], c["q"], c["pw"]
    g_k = pw[:, :, None] * g_att[:, None, :]
    g_p = np.einsum("bk,btk->bt", g_att, k)
    g_s = np.where(c["mask"], pw * (g_p - np.sum(pw * g_p, axis=1, keepdims=True)), 0.0)
    new_att_out = (att_out[0] - lr * np.einsum("bth,bt->h", c["xs"][-1], g_s)[:, None], att_out[1] - lr * np.sum(g_s))
    g = g_s[:, :, None] * att_out[0][:, 0]
    new_att = [None] * len(att_dense)
    for l in range(len(att_dense) - 1, -1, -1):
        kernel, bias, alpha = att_dense[l]
        P = c["pre"][l]
        g_alpha = np.sum(g * np.where(P < 0, P, 0.0), axis=0)                          # [T, h]
        gp = g * np.where(P > 0, 1.0, alpha)
        new_att[l] = (kernel - lr * np.einsum("bti,bth->ih", c["xs"][l], gp), bias - lr * gp.sum(axis=(0, 1)), alpha - lr * g_alpha)
        g = gp @ kernel.T
    gA = g                                                                             # [B, T, 4K]
    g_q += np.sum(gA[..., :K] + gA[..., 2 * K:3 * K] + gA[..., 3 * K:] * k, axis=1)
    g_k = g_k + gA[..., K:2 * K] - gA[..., 2 * K:3 * K] + gA[..., 3 * K:] * q[:, None, :]
    new_seq = seq_table.copy()
    np.add.at(new_seq, c["hist"], -lr * g_k)
    np.add.at(new_seq, c["cand"], -lr * g_q)
    new_other = [t.copy() for t in other_tables]
    col = 2 * K
    for i, t in enumerate(other_tables):
        np.add.at(new_other[i], c["other_ids"][:, i], -lr * gx[:, col:col + t.shape[1]])
        col += t.shape[1]
    new = dict(seq_table=new_seq, other_tables=new_other, att_dense=new_att, att_out=new_att_out, bn=new_bn, tower=new_tower,
               out=new_out)
    return new, bce, int(np.sum((p > 0.5) == (y > 0.5)))
