"""CPU: host-side mirror of the reference interface -- constructor signatures, weight names/shapes,
set_weights/load_npz, feature-column handling, and the loud failure without a CUDA device."""
import inspect

import numpy as np
import pytest
import torch

from recommender_system_b200 import _lib
from recommender_system_b200.layer import (Attention, CrossLayer, Dice, DNNLayer, EmbedLayer, FFMLayer, FMLayer,
                                           InnerProductLayer, OuterProductLayer, WideLayer)
from recommender_system_b200.model.dcn import DCN
from recommender_system_b200.model.deepFM import DeepFM
from recommender_system_b200.model.afm import AFM, AFMLayer
from recommender_system_b200.model.din import DIN
from recommender_system_b200.model.nfm import NFM
from recommender_system_b200.model.ffm import FFM
from recommender_system_b200.model.fm import FM
from recommender_system_b200.model.pnn import PNN
from recommender_system_b200.model.wideDeep import WideDeep
from recommender_system_b200.utils import synth

FC = synth.criteo_feature_columns(synth.SAMPLE_VOCAB, 8)


def _params(fn):
    return [p for p in inspect.signature(fn).parameters if p != "self"]


def test_constructor_signatures_start_like_the_reference():
    assert _params(EmbedLayer.__init__)[:2] == ["sparse_feature_columns", "k"]                      # core.py:268
    assert _params(DNNLayer.__init__)[:4] == ["hidden_units", "output_dim", "activation", "dropout"]  # interaction.py:31
    assert _params(CrossLayer.__init__)[:3] == ["layer_num", "reg_w", "reg_b"]                      # :50
    assert _params(FMLayer.__init__)[:3] == ["k", "reg_w", "reg_b"]                                 # :87
    assert _params(FFMLayer.__init__)[:4] == ["feature_columns", "k", "w_reg", "v_reg"]             # :118
    assert _params(Attention.__init__)[:2] == ["hidden_units", "activation"]                        # :356
    assert _params(Dice.__init__)[:2] == ["axis", "epsilon"]                                        # :411
    assert _params(FM.__init__)[:3] == ["k", "w_reg", "v_reg"]
    assert _params(FFM.__init__)[:4] == ["feature_columns", "k", "w_reg", "v_reg"]
    assert _params(DeepFM.__init__)[:7] == ["feature_columns", "k", "w_reg", "v_reg", "hidden_units", "output_dim",
                                            "activation"]
    assert _params(WideDeep.__init__)[:4] == ["feature_columns", "hidden_units", "output_dim", "activation"]
    assert _params(DCN.__init__)[:7] == ["feature_columns", "hidden_units", "output_dim", "activation", "layer_num",
                                         "reg_w", "reg_b"]
    assert _params(PNN.__init__)[:7] == ["feature_columns", "mode", "hidden_units", "output_dim", "activation",
                                         "dropout", "use_fgcnn"]
    assert _params(DIN.__init__)[:7] == ["feature_columns", "behavior_feature_list", "att_hidden_units",
                                         "dnn_hidden_units", "att_attention", "dnn_activation", "dnn_dropout"]
    assert _params(NFM.__init__)[:5] == ["feature_columns", "hidden_units", "output_dim", "activation", "dropout"]  # nfm.py:16
    assert _params(AFMLayer.__init__)[:2] == ["feature_columns", "mode"]                            # interaction.py:323
    assert _params(AFM.__init__)[:2] == ["feature_columns", "mode"]                                 # afm.py


def test_weight_names_and_shapes_follow_add_weight():
    fm = FMLayer(8)
    fm.build((None, 221))
    assert [(k, tuple(v.shape)) for k, v in fm.named_weights().items()] == [("w0", (1,)), ("w1", (221, 1)), ("v", (221, 8))]
    cr = CrossLayer(2)
    cr.build((None, 221))
    assert list(cr.named_weights()) == ["w0", "w1", "b0", "b1"]
    ffm = FFMLayer(FC, 4)
    ffm.build(None)
    assert tuple(ffm.v.shape) == (sum(synth.SAMPLE_VOCAB) + 13, 39, 4)
    assert ffm.feature_num == 43605 + 13 and ffm.field_num == 39
    op = OuterProductLayer()
    op.build((None, 26, 8))
    assert tuple(op.w.shape) == (8, 325, 8)
    att = Attention((80, 40))
    att.build([(None, 64), (None, 100, 64), (None, 100, 64), (None, 100)])
    w = att.named_weights()
    assert tuple(w["dense_0/kernel"].shape) == (256, 80) and tuple(w["dense_0/prelu_alpha"].shape) == (100, 80)
    assert tuple(w["dense_1/prelu_alpha"].shape) == (100, 40) and tuple(w["out/kernel"].shape) == (40, 1)


def test_embed_layer_ignores_feature_embed_dim_like_the_reference():
    fc = synth.criteo_feature_columns(synth.SAMPLE_VOCAB, 32)
    e = EmbedLayer(fc[1])            # k defaults to 8 regardless of feat['embed_dim']
    assert e.k == 8 and tuple(e.arena.shape) == (43605, 8) and len(e.emb_layers) == 26
    assert tuple(e.emb_layers[2].shape) == (5521, 8)
    assert e.emb_layers[1].data_ptr() == e.arena[176:].data_ptr()


def test_set_weights_roundtrip_and_shape_check(tmp_path):
    fm = FMLayer(4)
    fm.build((None, 10))
    ws = [np.full(s, i + 1, np.float32) for i, s in enumerate([(1,), (10, 1), (10, 4)])]
    v0 = fm._version
    fm.set_weights(ws)
    assert fm._version > v0
    for a, b in zip(fm.get_weights(), ws):
        np.testing.assert_array_equal(a, b)
    with pytest.raises(ValueError):
        fm.set_weights([ws[0], ws[1], np.zeros((3, 3), np.float32)])
    np.savez(tmp_path / "w.npz", w0=ws[0] * 2, w1=ws[1] * 2, v=ws[2] * 2)
    fm.load_npz(tmp_path / "w.npz")
    np.testing.assert_array_equal(fm.get_weights()[2], ws[2] * 2)


def test_models_construct_and_build_on_cpu():
    m = DeepFM(FC, 10, 1e-4, 1e-4, [256, 128, 64], 1, "relu")
    m.build(None)
    assert tuple(m.fm.v.shape) == (221, 10)
    d = DCN(FC, [32, 16], 1, "relu", 3)
    d.build(None)
    assert tuple(d.out_kernel.shape) == (222, 1)
    w = WideDeep(FC, [32], 1, "relu")
    w.build(None)
    assert w.wide.w.shape[0] == 13 + 13 + 43605
    p = PNN(FC, "both", [32], 1)
    p.build(None)
    assert p.width == 208 + 650
    with pytest.raises(ValueError):
        PNN(FC, "nope", [32], 1)
    din = DIN(synth.din_feature_columns(100, 50, 8), ["movies_seq"])
    din.build(None)
    assert din.dcat == 8 * 2 + 8 + 1


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_call_without_cuda_fails_loudly():
    fm = FMLayer(4)
    with pytest.raises(_lib.CtrhotError, match="no CPU fallback"):
        fm(np.zeros((2, 10), np.float32))
    m = DeepFM(FC, 10, 1e-4, 1e-4, [32], 1, "relu")
    with pytest.raises(_lib.CtrhotError):
        m(np.zeros((2, 39)))


def test_product_code_never_imports_the_oracle():
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for dp, _, fs in os.walk(os.path.join(root, "recommender_system_b200")):
        for f in fs:
            if f.endswith(".py"):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, os.path.join(dp, f)


def test_synth_batches():
    dense, ids = synth.criteo_batch(64, synth.KAGGLE_VOCAB, seed=1)
    assert dense.shape == (64, 13) and ids.shape == (64, 26) and ids.dtype == np.int32
    assert (ids >= 0).all() and (ids < np.array(synth.KAGGLE_VOCAB)).all()
    _, z = synth.criteo_batch(64, synth.KAGGLE_VOCAB, seed=1, dist="zipf")
    assert (z >= 0).all() and (z < np.array(synth.KAGGLE_VOCAB)).all()
    b = synth.din_batch(512, 100, 1000, 50, seed=2)
    assert b["movies_seq"].shape == (512, 100) and (b["movies_seq"].sum(1) == 0).any()


def test_compile_fit_signature_matches_the_reference():
    """DL/utils/compile_fit.py:9: compile_fit(model, X, y, batch_size=32, epochs=10, sgd=0.01)."""
    import inspect
    from recommender_system_b200.utils.compile_fit import compile_fit
    sig = inspect.signature(compile_fit)
    names = list(sig.parameters)
    assert names[:6] == ["model", "X", "y", "batch_size", "epochs", "sgd"]
    assert [sig.parameters[n].default for n in names[3:6]] == [32, 10, 0.01]
