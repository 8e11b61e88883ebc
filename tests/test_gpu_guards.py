"""GPU: home-made memcheck / racecheck for the warp-specialised kernels (compute-sanitizer is closed on this pool,
profiles/r02_sanitizer.txt).

* guard bands: ids, dense features, transport records, the embedding arena and the output are carved out of allocations with
  poisoned bands on both sides.  A write outside a buffer breaks a band; a read outside the arena / inputs picks up NaN (or an
  id far outside every table) and shows in the output or the id flag.  Results must equal the unguarded run bit for bit.
* determinism: 20 back-to-back runs on the same input must give identical bits -- the kernels hand tiles between roles
  through ~40 mbarriers and aliased TMEM columns, and a missing wait shows up as run-to-run noise long before it shows as a
  wrong mean.
"""
import numpy as np
import pytest
import torch

from recommender_system_b200 import _lib
from recommender_system_b200.layer.base import ids_arg
from recommender_system_b200.model.criteo import DCN, DeepFM, WideDeep
from recommender_system_b200.model.din import DIN
from recommender_system_b200.utils import synth
from recommender_system_b200.utils.transport import PackedRows, RowLayout

pytestmark = pytest.mark.gpu
G = 4096            # guard elements on each side


def guarded(t, poison):
    """-> (view equal to t, whole buffer, check()) with `poison` on both sides of the view (first dimension)."""
    t = t.contiguous()
    row = t[0].numel() if t.dim() > 1 else 1
    g = ((G + row - 1) // row + 15) // 16 * 16   # (a multiple of 16 rows: the view keeps the base alignment the C-ABI asks for)
    big = torch.empty((t.shape[0] + 2 * g,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    big.fill_(poison)
    big[g:g + t.shape[0]].copy_(t)
    view = big[g:g + t.shape[0]]

    def check():
        for band in (big[:g], big[g + t.shape[0]:]):
            if t.dtype.is_floating_point and poison != poison:
                assert bool(torch.isnan(band).all()), "guard band overwritten"
            else:
                assert bool((band == poison).all()), "guard band overwritten"
    return view, big, check


def put_arena_behind_guards(E):
    view, big, check = guarded(E.arena, float("nan"))
    E.arena = view
    E._weights["embeddings"] = view
    off = E.row_off_host
    E.emb_layers = [view[off[i]:off[i + 1]] for i in range(E.F)]
    E._bump()
    return big, check


@pytest.mark.parametrize("B,cluster", [(1, 1), (1000, 1), (1000, 2), (5001, 4)])
@pytest.mark.parametrize("packed", [False, True])
def test_whole_model_kernel_stays_inside_its_buffers(lib, B, cluster, packed):
    vocab = synth.SAMPLE_VOCAB
    m = DeepFM(synth.criteo_feature_columns(vocab, 16), 10, 1e-4, 1e-4, [256, 128, 64], 1, "relu", embed_dim=16)
    m.build(None)
    m.cluster = cluster
    dense, ids = synth.criteo_batch(B, vocab, seed=B + cluster)
    ids[B - 1] = np.asarray(vocab) - 1
    want = m((dense, ids))
    E = m.embed_layer
    keep, check_arena = put_arena_behind_guards(E)
    D = 13
    stack = m.dnn.stack
    stack.set_layers(m._tower_layers(m.dnn, D), (m.dnn._version, D))
    side, sq_w, cb = m._fm_side(D)
    out, _, check_out = guarded(torch.zeros((B, 1), dtype=torch.float32, device=m.device), float("nan"))
    if packed:
        lay = RowLayout(vocab, D)
        rec, _, check_in = guarded(torch.from_numpy(lay.pack(dense, ids)).to(m.device), 0xFF)     # ids of 0xFFFFFF.. if read
        stack.run_ctr_fused(E, None, _lib.IDS_PACKED, 0, rec, B, m.dnn.hidden[0][0], side, sq_w, mode=1, n_side=m.fm.k,
                            combine=cb, sigmoid=True, err_flag=m.err_flag(), out=out, cluster=cluster,
                            side_key=("fm", m.fm._version), records=PackedRows(lay, rec), n_dense=D)
        checks = [check_in]
    else:
        ids_d, _, check_ids = guarded(torch.from_numpy(ids).to(m.device), 0x7FFFFFF0)             # far outside every table
        dense_d, _, check_dense = guarded(torch.from_numpy(dense).to(m.device), float("nan"))
        i2, kind, ld = ids_arg(ids_d)
        stack.run_ctr_fused(E, i2, kind, ld, dense_d, B, m.dnn.hidden[0][0], side, sq_w, mode=1, n_side=m.fm.k, combine=cb,
                            sigmoid=True, err_flag=m.err_flag(), out=out, cluster=cluster, side_key=("fm", m.fm._version))
        checks = [check_ids, check_dense]
    torch.cuda.synchronize()
    m.check_ids()                                             # a guard id that had been read would raise IndexError here
    for c in checks + [check_out, check_arena]:
        c()
    assert bool(torch.isfinite(out).all())
    assert torch.equal(out, want)
    del keep


def _repeat_identical(fn, n=20):
    first = fn().clone()
    for _ in range(n - 1):
        assert torch.equal(fn(), first)
    return first


@pytest.mark.parametrize("cluster", [1, 2, 4])
def test_deepfm_whole_kernel_is_deterministic(lib, cluster):
    vocab = synth.SAMPLE_VOCAB
    m = DeepFM(synth.criteo_feature_columns(vocab, 16), 10, 1e-4, 1e-4, [256, 128, 64], 1, "relu", embed_dim=16, check_ids=False)
    m.build(None)
    m.cluster = cluster
    dense, ids = synth.criteo_batch(30011, vocab, seed=9)
    d, i = torch.from_numpy(dense).to(m.device), torch.from_numpy(ids).to(m.device)
    a = _repeat_identical(lambda: m((d, i)))
    lay = RowLayout(vocab, 13)
    rec = PackedRows(lay, torch.from_numpy(lay.pack(dense, ids)).to(m.device))
    b = _repeat_identical(lambda: m(rec))
    assert torch.equal(a, b)


def test_dcn_wide_tower_and_widedeep_are_deterministic(lib):
    vocab = synth.SAMPLE_VOCAB
    dense, ids = synth.criteo_batch(20000, vocab, seed=10)
    m = DCN(synth.criteo_feature_columns(vocab, 16), [400, 400, 400], 1, "relu", 3, embed_dim=16, check_ids=False)
    m.build(None)
    d, i = torch.from_numpy(dense).to(m.device), torch.from_numpy(ids).to(m.device)
    _repeat_identical(lambda: m((d, i)))
    w = WideDeep(synth.criteo_feature_columns(vocab, 16), [256, 128, 64], 1, "relu", embed_dim=16, check_ids=False)
    w.build(None)
    _repeat_identical(lambda: w((d, i)))


def test_din_is_deterministic_and_reads_only_its_table(lib):
    B, T, K, v_item, v_user = 3000, 100, 64, 50000, 1000
    m = DIN(synth.din_feature_columns(v_item=v_item, v_user=v_user, embed_dim=K), ["movies_seq"], (80, 40), (256, 128, 64),
            "prelu", "prelu", 0.0, check_ids=False)
    m.build(None)
    m.att_layer.build_for(T, K)
    d = synth.din_batch(B, T, v_item=v_item, v_user=v_user, seed=12)
    d["movies_seq"][1, :] = np.arange(v_item - T, v_item)       # the table's last rows
    dev = {k: torch.from_numpy(v).to(m.device) for k, v in d.items()}
    want = _repeat_identical(lambda: m(dev))
    # the behaviour table behind NaN guards: same bits, no NaN picked up
    view, big, check = guarded(m.embed_seq_layers[0], float("nan"))
    m.embed_seq_layers[0] = view
    m._weights["embed_seq_0/embeddings"] = view
    m._bump()
    got = m(dev)
    torch.cuda.synchronize()
    check()
    assert bool(torch.isfinite(got).all()) and torch.equal(got, want)
