"""GPU + host: compact transport records (csrc/transport.cu, utils/transport.py) -- pack on the host, decode either inside the
whole-model kernel (CTRHOT_IDS_PACKED) or by ctrhot_rows_unpack -- against the plain (dense, ids) path: the records carry the
same numbers, so every result must be IDENTICAL, bit for bit (ids are integers, dense floats are copied)."""
import numpy as np
import pytest
import torch

from recommender_system_b200 import _lib
from recommender_system_b200.model.criteo import DCN, DeepFM
from recommender_system_b200.utils import synth
from recommender_system_b200.utils.pipeline import HostPipeline
from recommender_system_b200.utils.transport import PackedRows, RowLayout


def test_layout_widths_follow_the_vocabulary(lib):
    lay = RowLayout([256, 257, 65536, 65537, 1 << 24, (1 << 24) + 1, 3], n_dense=13, bits=False)
    assert lay.widths == [1, 2, 2, 3, 3, 4, 1]
    assert lay.rec_bytes == (52 + 16 + 3) // 4 * 4
    k = RowLayout(synth.KAGGLE_VOCAB, 13, bits=False)
    assert k.rec_bytes == 104 and sum(k.widths) == 52           # 156 B per sample as int32
    # bit-granular (the default): ceil(log2(vocab)) bits per id, at least one
    lb = RowLayout([256, 257, 65536, 65537, 1 << 24, (1 << 24) + 1, 3, 1, 2], n_dense=13)
    assert lb.widths == [8, 9, 16, 17, 24, 25, 2, 1, 1]
    assert lb.rec_bytes == (52 * 8 + sum(lb.widths) + 31) // 32 * 4
    kb = RowLayout(synth.KAGGLE_VOCAB, 13)
    assert kb.rec_bytes == 96 and sum(kb.widths) == 321


@pytest.mark.parametrize("bits", [True, False])
def test_host_pack_round_trips_in_both_layouts(lib, bits):
    """Decoded on the host independently of the C code, bit by bit."""
    rng = np.random.default_rng(1)
    vocab = [256, 300, 70000, 20_000_000, 5, 65536, 1, 2, (1 << 31) - 1, 3]
    lay = RowLayout(vocab, n_dense=3, bits=bits)
    B = 4097
    dense = rng.random((B, 3), dtype=np.float32)
    ids = np.stack([rng.integers(0, v, B) for v in vocab], 1).astype(np.int32)
    ids[0] = np.asarray(vocab) - 1
    rec = lay.pack(dense, ids)
    np.testing.assert_array_equal(rec[:, :12].copy().view(np.float32).reshape(B, 3), dense)
    allbits = np.unpackbits(rec, axis=1, bitorder="little").astype(np.int64)          # [B, 8 rec_bytes], LSB first
    for f in range(len(vocab)):
        d = int(lay.desc_host[f]) & 0xFFFFFFFF
        off, w = (d & 0xFFFFFF, (d >> 24) & 0x7F) if d >> 31 else ((d & 0xFFFF) * 8, (d >> 16) * 8)
        v = sum(allbits[:, off + b] << b for b in range(w))
        np.testing.assert_array_equal(v, ids[:, f])
    np.testing.assert_array_equal(rec, lay.pack(dense, ids, threads=1))
    bad = ids.copy()
    bad[5, 1] = 512                                             # 300 needs 9 bits / 2 bytes: 512 fits neither... in bits
    if bits:
        with pytest.raises(IndexError):
            lay.pack(dense, bad)


def test_host_pack_round_trips_and_reports_ids_that_do_not_fit(lib):
    rng = np.random.default_rng(0)
    vocab = [256, 300, 70000, 20_000_000, 5, 65536]
    lay = RowLayout(vocab, n_dense=3, bits=False)
    B = 10_001                                                  # several threads, ragged last chunk
    dense = rng.random((B, 3), dtype=np.float32)
    ids = np.stack([rng.integers(0, v, B) for v in vocab], 1).astype(np.int32)
    ids[0] = np.asarray(vocab) - 1
    rec = lay.pack(dense, ids)
    assert rec.shape == (B, lay.rec_bytes)
    # decode on the host, independently of the C code
    got_dense = rec[:, :12].copy().view(np.float32).reshape(B, 3)
    np.testing.assert_array_equal(got_dense, dense)
    for f in range(len(vocab)):
        off, w = int(lay.desc_host[f]) & 0xFFFF, int(lay.desc_host[f]) >> 16
        v = sum(rec[:, off + b].astype(np.int64) << (8 * b) for b in range(w))
        np.testing.assert_array_equal(v, ids[:, f])
    np.testing.assert_array_equal(rec, lay.pack(dense, ids, threads=1))
    bad = ids.copy()
    bad[77, 0] = 256                                            # one past what a byte holds
    with pytest.raises(IndexError):
        lay.pack(dense, bad)


@pytest.mark.gpu
@pytest.mark.parametrize("B", [1, 127, 1000, 65536])
def test_unpack_kernel_is_exact(lib, B):
    rng = np.random.default_rng(B)
    vocab = synth.KAGGLE_VOCAB
    lay = RowLayout(vocab, 13)
    dense, ids = synth.criteo_batch(B, vocab, seed=B)
    ids[-1] = np.asarray(vocab) - 1
    rec = torch.from_numpy(lay.pack(dense, ids)).cuda()
    d, i = lay.unpack(rec)
    np.testing.assert_array_equal(d.cpu().numpy(), dense)
    np.testing.assert_array_equal(i.cpu().numpy(), ids)


@pytest.mark.gpu
@pytest.mark.parametrize("B,vocab_name", [(1, "sample"), (300, "sample"), (5000, "sample"), (65536, "kaggle")])
def test_deepfm_on_records_equals_deepfm_on_dense_and_ids(lib, B, vocab_name):
    vocab = synth.KAGGLE_VOCAB if vocab_name == "kaggle" else synth.SAMPLE_VOCAB
    m = DeepFM(synth.criteo_feature_columns(vocab, 16), 10, 1e-4, 1e-4, [256, 128, 64], 1, "relu", embed_dim=16)
    m.build(None)
    lay = RowLayout.from_feature_columns(m.feature_columns if hasattr(m, "feature_columns")
                                         else [m.dense_feature_columns, m.sparse_feature_columns])
    dense, ids = synth.criteo_batch(B, vocab, seed=7 + B)
    ids[B // 2] = np.asarray(vocab) - 1                          # the last row of every table, every width saturated
    want = m((dense, ids)).cpu().numpy()
    launches = _lib.LAUNCHES[0]
    got = m(PackedRows(lay, torch.from_numpy(lay.pack(dense, ids)).cuda())).cpu().numpy()
    assert _lib.LAUNCHES[0] - launches == 1                     # decoded inside the whole-model kernel: no unpack pass
    np.testing.assert_array_equal(got, want)
    # an id beyond its table (but inside its byte width) is reported like on the int32 path
    f = 1 if vocab_name == "kaggle" else 0                      # a field whose vocabulary does not fill its width
    bad = ids.copy()
    bad[0, f] = vocab[f]
    with pytest.raises(IndexError):
        m(PackedRows(lay, torch.from_numpy(lay.pack(dense, bad)).cuda()))


@pytest.mark.gpu
def test_other_models_take_records_through_the_unpack_kernel(lib):
    vocab = synth.SAMPLE_VOCAB
    m = DCN(synth.criteo_feature_columns(vocab, 16), [64, 32], 1, "relu", 3, embed_dim=16)
    m.build(None)
    lay = RowLayout(vocab, 13)
    dense, ids = synth.criteo_batch(777, vocab, seed=3)
    want = m((dense, ids)).cpu().numpy()
    got = m(PackedRows(lay, torch.from_numpy(lay.pack(dense, ids)).cuda())).cpu().numpy()
    np.testing.assert_array_equal(got, want)


@pytest.mark.gpu
def test_host_pipeline_with_records(lib):
    vocab = synth.SAMPLE_VOCAB
    m = DeepFM(synth.criteo_feature_columns(vocab, 16), 10, 1e-4, 1e-4, [256, 128, 64], 1, "relu", embed_dim=16, check_ids=False)
    m.build(None)
    lay = RowLayout(vocab, 13)
    B = 4096
    pipe = HostPipeline(m, B, depth=3, layout=lay)
    batches = [synth.criteo_batch(B, vocab, seed=40 + i) for i in range(5)]
    recs = [torch.from_numpy(lay.pack(d, i)).pin_memory() for d, i in batches]
    outs = []
    tickets = []
    for j, r in enumerate(recs):
        tickets.append(pipe.submit_packed(r))
        if j >= 2:
            outs.append(pipe.result(tickets[j - 2]).copy())
    outs += [r.copy() for r in pipe.drain()]
    assert len(outs) == 5
    for (d, i), o in zip(batches, outs):
        np.testing.assert_array_equal(o, m((d, i)).cpu().numpy())
    assert pipe.h2d_bytes_packed == B * lay.rec_bytes
