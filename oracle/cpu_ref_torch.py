"""Timed CPU baseline ("port") -- TEST/BENCH INFRASTRUCTURE ONLY, never the product path.

torch-CPU fp32 restatements that keep the reference's op STRUCTURE (26 separate embedding lookups
-> stack -> transpose -> reshape, DL/layer/core.py:273-280; the 3-GEMM FM with two pow passes,
DL/layer/interaction.py:106-114; Dense chain, :40-46; tiled query + 4K concat for DIN, :379-404),
because the reference's own TensorFlow code cannot run here (TF not installed, no network).
Only bench.py's cpu_baseline / --impl reference legs and tests/ import this.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def embed_layer(sparse_inputs, tables):
    embs = [F.embedding(sparse_inputs[:, i].to(torch.int32).long(), t) for i, t in enumerate(tables)]
    emb = torch.stack(embs, 0).permute(1, 0, 2)
    return emb.reshape(emb.shape[0], emb.shape[1] * emb.shape[2])


def fm_layer(x, w0, w1, v):
    linear = x @ w1 + w0
    p1 = torch.pow(x @ v, 2)
    p2 = torch.pow(x, 2) @ torch.pow(v, 2)
    return linear + 0.5 * torch.sum(p1 - p2, dim=-1, keepdim=True)


def dnn_layer(x, hidden, out):
    for k, b in hidden:
        x = torch.relu(x @ k + b)
    return x @ out[0] + out[1]


def cross_layer(x, ws, bs):
    x0 = x.unsqueeze(2)
    x1 = x0
    for w, b in zip(ws, bs):
        x1_w = torch.matmul(x1.transpose(1, 2), w)
        x1 = torch.matmul(x0, x1_w) + b + x1
    return x1.squeeze(2)


def deepfm(inputs, tables, fm, hidden, out):
    """DeepFM.call, DL/model/deepFM.py:23-31, on the [B,39] float matrix."""
    dense, sparse = inputs[:, :13], inputs[:, 13:]
    x = torch.cat([dense, embed_layer(sparse, tables)], dim=-1)
    return torch.sigmoid(0.5 * (fm_layer(x, *fm) + dnn_layer(x, hidden, out)))


def dcn(inputs, tables, ws, bs, hidden, out, final):
    dense, sparse = inputs[:, :13], inputs[:, 13:]
    x = torch.cat([dense, embed_layer(sparse, tables)], dim=1)
    cat = torch.cat([cross_layer(x, ws, bs), dnn_layer(x, hidden, out)], dim=1)
    return torch.sigmoid(cat @ final[0] + final[1])


def din_attention(q, key, mask, dense, out):
    T = key.shape[1]
    qt = q.unsqueeze(1).repeat(1, T, 1)
    emb = torch.cat([qt, key, qt - key, qt * key], dim=-1)
    for k, b, a in dense:
        h = emb @ k + b
        emb = torch.relu(h) - a * torch.relu(-h)
    score = (emb @ out[0] + out[1]).squeeze(-1)
    score = torch.where(mask == 0, torch.full_like(score, float(-2 ** 32 + 1)), score)
    score = torch.softmax(score, dim=-1)
    return torch.matmul(score.unsqueeze(1), key).squeeze(1)
