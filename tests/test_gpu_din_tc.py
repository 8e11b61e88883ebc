"""GPU parity for the tcgen05 DIN attention unit (ctrhot_din_attention_gather_tf32, csrc/din_tc.cu) against the float64
oracle of Attention.call (DL/layer/interaction.py:369-406) fed by the history / candidate lookups of DIN.call
(DL/model/din.py:76-81).

TF32 multiply + the hoisted first layer change rounding only: the attention output (a convex combination of the gathered
keys) must agree to 2e-3 absolute (BASELINE north_star's tolerance for tensor-core paths), and it must agree with the exact
fp32 CUDA-core kernel of the same unit to the same bar.  Cases: the BASELINE config-4 shape (T=100, K=64, (80,40)), ragged
batch sizes, an all-padding history (uniform average, interaction.py:399-404), a full history, the smallest and the largest
shapes the kernel takes, relu-free PReLU with non-zero alpha of shape [T, h].
"""
import numpy as np
import pytest
import torch

from oracle import ctr_oracle as O
from recommender_system_b200.layer import Attention

pytestmark = pytest.mark.gpu
TOL = 2e-3


def npy(t):
    return t.detach().cpu().numpy().astype(np.float64)


def make_case(B, T, K, hidden, V, seed, scale=1.0):
    rng = np.random.default_rng(seed)
    f32 = lambda a: np.asarray(a, dtype=np.float32)
    table = f32(rng.uniform(-0.5, 0.5, (V, K)) * scale)
    lens = rng.integers(0, T + 1, size=B)
    if B > 2:
        lens[0], lens[1] = 0, T                                  # one all-padding row, one full row
    hist = rng.integers(1, V, size=(B, T)).astype(np.int32)
    hist[np.arange(T)[None, :] >= lens[:, None]] = 0
    cand = rng.integers(1, V, size=(B, 1)).astype(np.int32)
    dims = [4 * K] + list(hidden)
    ws, dense = [], []
    for i in range(len(hidden)):
        lim = np.sqrt(6.0 / (dims[i] + dims[i + 1]))
        k, b, a = f32(rng.uniform(-lim, lim, (dims[i], dims[i + 1]))), f32(rng.normal(0, 0.1, dims[i + 1])), \
            f32(rng.uniform(0, 0.3, (T, dims[i + 1])))
        ws += [k, b, a]
        dense.append((k.astype(np.float64), b.astype(np.float64), a.astype(np.float64)))
    ko, bo = f32(rng.uniform(-0.5, 0.5, (dims[-1], 1))), f32(rng.normal(0, 0.1, 1))
    ws += [ko, bo]
    key = table[hist].astype(np.float64)
    q = table[cand[:, 0]].astype(np.float64)
    mask = (hist != 0).astype(np.float64)
    want = O.din_attention(q, key, key, mask, dense, (ko.astype(np.float64), bo.astype(np.float64)))
    return table, hist, cand, ws, want, q


def run(table, hist, cand, ws, T, K, hidden, prec):
    L = Attention(hidden, "prelu", precision=prec)
    L.build_for(T, K)
    L.set_weights(ws)
    tab = torch.from_numpy(table).to(L.device)
    flag = torch.zeros(1, dtype=torch.int32, device=L.device)
    att, item = L.call_gather(tab, torch.from_numpy(hist).to(L.device), torch.from_numpy(cand).to(L.device), flag)
    torch.cuda.synchronize()
    assert int(flag.item()) == 0
    return npy(att), npy(item)


@pytest.mark.parametrize("B", [1, 3, 200, 1000])
def test_config4_shape(lib, B):
    T, K, hidden = 100, 64, (80, 40)
    table, hist, cand, ws, want, q = make_case(B, T, K, hidden, 5000, seed=B)
    got, item = run(table, hist, cand, ws, T, K, hidden, "tf32")
    np.testing.assert_array_equal(item, q)                       # candidate rows: bit-exact copies
    assert np.max(np.abs(got - want)) <= TOL, np.max(np.abs(got - want))
    exact, _ = run(table, hist, cand, ws, T, K, hidden, "fp32")
    assert np.max(np.abs(exact - want)) <= 2e-5
    if B > 2:                                                    # all-padding history -> uniform average of the T keys
        np.testing.assert_allclose(got[0], table[hist[0]].astype(np.float64).mean(0), rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("T,K,hidden", [(5, 8, (16, 8)), (128, 32, (96, 64)), (37, 16, (24, 8)), (100, 64, (64, 32))])
def test_shapes(lib, T, K, hidden):
    table, hist, cand, ws, want, q = make_case(300, T, K, hidden, 999, seed=T + K)
    got, item = run(table, hist, cand, ws, T, K, hidden, "tf32")
    np.testing.assert_array_equal(item, q)
    assert np.max(np.abs(got - want)) <= TOL, np.max(np.abs(got - want))


def test_large_scores_keep_the_softmax_sharp(lib):
    """Scores of O(10): the softmax is nearly one-hot, so TF32 score errors move the weights the most here."""
    T, K, hidden = 100, 64, (80, 40)
    table, hist, cand, ws, want, q = make_case(500, T, K, hidden, 3000, seed=77, scale=4.0)
    got, _ = run(table, hist, cand, ws, T, K, hidden, "tf32")
    # keys are O(2) here: compare relative to their size
    assert np.max(np.abs(got - want)) <= 8 * TOL, np.max(np.abs(got - want))


def test_full_batch_properties(lib):
    """B = 16384 (BASELINE config 4): sample-permutation equivariance is bit-exact, a strided subset matches the oracle."""
    T, K, hidden, B = 100, 64, (80, 40), 16384
    table, hist, cand, ws, _, _ = make_case(8, T, K, hidden, 63002, seed=5)
    rng = np.random.default_rng(1)
    lens = rng.integers(0, T + 1, size=B)
    hist = rng.integers(1, 63002, size=(B, T)).astype(np.int32)
    hist[np.arange(T)[None, :] >= lens[:, None]] = 0
    cand = rng.integers(1, 63002, size=(B, 1)).astype(np.int32)
    got, _ = run(table, hist, cand, ws, T, K, hidden, "tf32")
    perm = rng.permutation(B)
    got_p, _ = run(table, hist[perm], cand[perm], ws, T, K, hidden, "tf32")
    np.testing.assert_array_equal(got[perm], got_p)
    sub = np.arange(0, B, 331)
    dense = [(ws[3 * i].astype(np.float64), ws[3 * i + 1].astype(np.float64), ws[3 * i + 2].astype(np.float64)) for i in range(2)]
    key = table[hist[sub]].astype(np.float64)
    want = O.din_attention(table[cand[sub, 0]].astype(np.float64), key, key, (hist[sub] != 0).astype(np.float64), dense,
                           (ws[6].astype(np.float64), ws[7].astype(np.float64)))
    assert np.max(np.abs(got[sub] - want)) <= TOL
