"""GPU: the input step (csrc/encode.cu, utils/dataset.py) against the REFERENCE LOADER'S OWN OUTPUT.

tests/golden/criteo_cfg1.npz is what the reference's create_criteo_dataset returned on its bundled train_1w.txt (made by
tests/golden/make_golden.py by importing the reference's utils/dataset.py); tests/golden/criteo_raw.npz holds that file's raw
columns as integers plus the row numbers of the reference's seeded test split.  The test re-serialises the raw columns to TSV
text, runs parse -> fit -> transform on the GPU and must reproduce the reference: ids bit-exact, scaled counts <= 1e-7,
labels, and the feat_onehot_dim list of features_dict."""
import os
import time

import numpy as np
import pytest
import torch

from recommender_system_b200.utils.dataset import CriteoEncoder, create_criteo_dataset, features_dict

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def tsv_from_raw(z, rows=None, with_label=True):
    label, dm, dr, tm, tk = z["label"], z["dense_missing"], z["dense_raw"], z["token_missing"], z["tokens"]
    rows = np.arange(label.shape[0]) if rows is None else rows
    out = []
    for r in rows:
        f = [str(int(label[r]))] if with_label else []
        f += ["" if dm[r, c] else str(int(dr[r, c])) for c in range(13)]
        f += ["" if tm[r, c] else "%08x" % int(tk[r, c]) for c in range(26)]
        out.append("\t".join(f))
    return ("\n".join(out) + "\n").encode()


@pytest.fixture(scope="module")
def raw():
    z = np.load(os.path.join(HERE, "golden", "criteo_raw.npz"))
    return {k: z[k] for k in z.files}           # plain arrays: an NpzFile decompresses the member again on every access


def test_reproduces_the_reference_loader(lib, raw, criteo_cfg1, tmp_path):
    text = tsv_from_raw(raw)
    enc = CriteoEncoder()
    label, dense, ids = enc.fit_transform_text(text)
    ti = raw["test_index"]
    np.testing.assert_array_equal(ids.cpu().numpy()[ti], criteo_cfg1["ids"])                       # bit-exact
    assert np.max(np.abs(dense.cpu().numpy()[ti].astype(np.float64) - criteo_cfg1["dense"].astype(np.float64))) <= 1e-7
    np.testing.assert_array_equal(label.cpu().numpy()[ti], criteo_cfg1["y"])
    np.testing.assert_array_equal(np.asarray(enc.feat_onehot_dim), criteo_cfg1["feat_onehot_dim"])    # features_dict
    np.testing.assert_array_equal(np.asarray(enc.n_classes), criteo_cfg1["onehot_width"])             # get_dummies widths
    # the drop-in functions on a file
    p = tmp_path / "train.txt"
    p.write_bytes(text)
    fc = features_dict(str(p))
    assert [f["feat_onehot_dim"] for f in fc[1]] == list(criteo_cfg1["feat_onehot_dim"]) and len(fc[0]) == 13
    np.random.seed(0)                                           # the reference's split draws from numpy's global RNG
    (Xtr, ytr), (Xte, yte) = create_criteo_dataset("DeepFM", str(p), test_size=0.2)
    assert Xtr.shape == (8000, 39) and Xte.shape == (2000, 39) and Xte.dtype == np.float64
    np.testing.assert_array_equal(Xte[:, 13:].astype(np.int32), criteo_cfg1["ids"])
    assert np.max(np.abs(Xte[:, :13] - criteo_cfg1["dense"].astype(np.float64))) <= 1e-7
    np.testing.assert_array_equal(yte, criteo_cfg1["y"])
    np.random.seed(0)
    (_, _), (Xo, _) = create_criteo_dataset("fm", str(p), test_size=0.2)
    assert Xo.shape == (2000, int(criteo_cfg1["onehot_cols"]))
    rows = np.arange(2000)
    for f in (0, 7, 25):
        assert np.all(Xo[rows, criteo_cfg1["onehot_offsets"][f] + criteo_cfg1["ids"][:, f]] == 1.0)
    assert np.all(Xo[:, 13:].sum(1) == 26)


def test_edge_cases(lib):
    enc = CriteoEncoder()
    line = lambda lab, d, t: "\t".join([lab] + d + t)
    d0 = [""] * 13
    t0 = [""] * 26
    text = "\n".join([line("1", ["-3"] + [""] * 12, ["0000000a"] + [""] * 25),
                      line("0", ["7"] + ["5"] * 12, ["ffffffff"] + ["00000001"] * 25),
                      line("0", d0, t0)]).encode()                 # no trailing newline, an all-missing row
    label, dr, tok = enc.parse(text)
    assert label.tolist() == [1, 0, 0]
    assert dr[0, 0].item() == -3.0 and torch.isnan(dr[0, 1]) and torch.isnan(dr[2]).all()
    assert tok[0, 0].item() == 10 and tok[1, 0].item() == 0xFFFFFFFF and (tok[2] == -1).all()
    enc.fit(dr, tok)
    dense, ids = enc.transform(dr, tok)
    # column 0: values (-3, 7, fill 0) -> min -3, max 7;  columns 1..: (fill 0, 5, fill 0)
    np.testing.assert_allclose(dense[:, 0].cpu().numpy(), [0.0, 1.0, 0.3], atol=1e-7)
    np.testing.assert_allclose(dense[:, 1].cpu().numpy(), [0.0, 1.0, 0.0], atol=0)
    assert ids[:, 0].tolist() == [1, 2, 0]                         # '-1' (missing) sorts first, then 0000000a, then ffffffff
    assert ids[:, 1].tolist() == [0, 1, 0]
    assert enc.feat_onehot_dim[0] == 3 and enc.feat_onehot_dim[1] == 2
    # unseen token at transform time, like LabelEncoder.transform
    _, dr2, tok2 = enc.parse(line("0", d0, ["0000000b"] + [""] * 25).encode())
    with pytest.raises(ValueError):
        enc.transform(dr2, tok2)
    with pytest.raises(ValueError):
        enc.parse(line("0", ["12x"] + [""] * 12, t0).encode())


def test_throughput_report(lib, raw):
    """Rows per second of the whole step (bytes on the host -> dense / ids on the device) next to the reference loader's
    17 s ('fm') / 1.0 s ('DeepFM') for these 10 000 rows: printed, and written for profiles/."""
    text = tsv_from_raw(raw) * 20                                   # 200 000 rows, 40 MB of text
    enc = CriteoEncoder()
    enc.fit_transform_text(text)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    enc.fit_transform_text(text)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    label, dr, tok = enc.parse(text)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        enc.transform(dr, tok)
    e1.record()
    torch.cuda.synchronize()
    tr = e0.elapsed_time(e1) / 10 * 1e-3
    msg = ("input step: %d rows, %.1f MB of TSV: parse + fit + transform %.1f ms (%.2f M rows/s, copy of the text included); "
           "transform alone %.3f ms (%.1f M rows/s); reference loader: 1.0 s for 10 000 rows in 'DeepFM' mode = 0.01 M rows/s"
           % (dr.shape[0], len(text) / 1e6, dt * 1e3, dr.shape[0] / dt / 1e6, tr * 1e3, dr.shape[0] / tr / 1e6))
    # the Hash layer (DL/layer/utils.py:21-56) on a [65 536, 26] int32 id matrix: ids in, int64 buckets out
    from recommender_system_b200.layer import Hash
    hx = torch.randint(0, 10 ** 9, (65536, 26), dtype=torch.int32, device="cuda")
    hl = Hash(1000003, mask_zero=True)
    hl(hx)
    e0.record()
    for _ in range(20):
        hl(hx)
    e1.record()
    torch.cuda.synchronize()
    th = e0.elapsed_time(e1) / 20 * 1e-3
    msg += ("\nHash layer: %d ids (decimal string -> FarmHash Fingerprint64 -> bucket) in %.3f ms = %.1f G ids/s, %.0f GB/s of "
            "ids + buckets" % (hx.numel(), th * 1e3, hx.numel() / th / 1e9, hx.numel() * 12 / th / 1e9))
    print(msg)
    out = os.path.join(os.path.dirname(HERE), "gpurun_out")
    if os.path.isdir(out):
        open(os.path.join(out, "r02_input_step.txt"), "w").write(msg + "\n")
    assert dr.shape[0] == 200000


# ------------------------------------------------------------------ Hash (DL/layer/utils.py:21-56)
def test_hash_layer_matches_tensorflows_documented_example_and_the_oracle(lib):
    """tf.strings.to_hash_bucket_fast(["Hello", "TensorFlow", "2.x"], 3) == [0, 2, 2] is the example in TensorFlow's documentation
    of the op (the pin of the FarmHash restatement); then integers as their decimal strings, mask_zero, negatives, 17-19 digit
    values and strings up to 64 bytes against oracle.hash_layer."""
    import torch
    from oracle import ctr_oracle as O
    from recommender_system_b200.layer import Hash
    got = Hash(3)(np.array([b"Hello", b"TensorFlow", b"2.x"])).cpu().numpy()
    assert got.tolist() == [0, 2, 2]
    rng = np.random.default_rng(0)
    x32 = np.concatenate([[0, 1, 9, 10, 99, 100, 2147483647, -1, -2147483648], rng.integers(0, 10 ** 9, 5000)]).astype(np.int32)
    for nb, mz in ((1000, False), (1000, True), (7, True), (2 ** 40, False)):
        np.testing.assert_array_equal(Hash(nb, mask_zero=mz)(x32).cpu().numpy(), O.hash_layer(x32, nb, mz))
    x64 = np.concatenate([[0, 10 ** 15, 10 ** 16, 10 ** 17 + 3, 9223372036854775807, -9223372036854775808, -(10 ** 18)],
                          rng.integers(0, 2 ** 62, 3000)]).astype(np.int64)
    x64 = x64[:x64.size // 7 * 7].reshape(-1, 7)                                      # a [B, 7] id matrix, already on the device
    np.testing.assert_array_equal(Hash(100003, mask_zero=True)(torch.from_numpy(x64).cuda()).cpu().numpy(),
                                  O.hash_layer(x64, 100003, True))
    words = [bytes(rng.integers(33, 127, n).astype(np.uint8)) for n in list(range(1, 65)) * 3] + [b"0", b"00", b""]
    np.testing.assert_array_equal(Hash(65536, mask_zero=True)(np.array(words, dtype=object)).cpu().numpy(),
                                  O.hash_layer(words, 65536, True))
    with pytest.raises(NotImplementedError):
        Hash(10, vocabulary_path="vocab.csv")
    with pytest.raises(NotImplementedError):
        Hash(10)(np.array([1.5, 2.0], np.float32))
