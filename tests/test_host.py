"""CPU tests of the host side: the C-ABI library loads and exports every symbol include/b2srgan.h declares (no
compute calls without a GPU), weight-packing index maps, and the Keras-compatible Model surface."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    with open(os.path.join(ROOT, "include", "b2srgan.h")) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b2_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from upscaler import _lib
    lib = _lib.load()
    declared = _declared_symbols()
    assert "b2_conv2d_fprop" in declared and len(declared) >= 7
    raw = ctypes.CDLL(_lib.lib_path())
    for name in declared:
        assert hasattr(raw, name), "libb2srgan.so does not export %s" % name
    assert sorted(_lib.exported_symbols()) == declared, "ctypes signatures and header disagree"
    assert lib.b2_version() >= 1


def test_null_pointer_calls_fail_loudly_without_touching_the_gpu():
    from upscaler import _lib
    lib = _lib.load()
    ep = _lib.Epilogue()
    rc = lib.b2_conv2d_fprop(None, 1, 8, 8, 64, None, 64, 3, 3, 1, ctypes.byref(ep), None)
    assert rc != 0 and b"null" in lib.b2_last_error()
    with pytest.raises(_lib.B2Error):
        _lib.check(rc, "b2_conv2d_fprop")


def test_pack_index_maps():
    from upscaler import ops
    rng = np.random.default_rng(0)
    k = rng.standard_normal((3, 3, 64, 128)).astype(np.float32)
    idx = ops.conv_pack_index(3, 3, 64, 128)
    packed = k.reshape(-1)[idx]
    assert packed.shape == (128, 9, 64)
    np.testing.assert_array_equal(packed, k.reshape(9, 64, 128).transpose(2, 0, 1))
    kt = rng.standard_normal((5, 5, 256, 64)).astype(np.float32)  # Conv2DTranspose layout (kh,kw,Cout,Cin)
    np.testing.assert_array_equal(kt.reshape(-1)[ops.convT_pack_index(5, 5, 64, 256)], kt.reshape(25, 256, 64).transpose(1, 0, 2))
    kh = rng.standard_normal((9, 9, 3, 64)).astype(np.float32)
    idx = ops.conv_rowfold_pack_index(9, 9, 3, 64)
    flat = np.concatenate([kh.reshape(-1), [0.0]]).astype(np.float32)
    p = flat[np.where(idx < 0, flat.size - 1, idx)]
    assert p.shape == (64, 9, 64) and (p[:, :, 27:] == 0).all()
    np.testing.assert_array_equal(p[:, :, :27], kh.transpose(3, 0, 1, 2).reshape(64, 9, 27))
    kf = rng.standard_normal((9, 9, 256, 3)).astype(np.float32)
    idx = ops.conv_rows_pack_index(9, 9, 256, 3)
    flat = np.concatenate([kf.reshape(-1), [0.0]]).astype(np.float32)
    p = flat[np.where(idx < 0, flat.size - 1, idx)]
    assert p.shape == (32, 9, 256) and (p[27:] == 0).all()
    np.testing.assert_array_equal(p[:27], kf.transpose(0, 3, 1, 2).reshape(27, 9, 256))
    v = rng.standard_normal((3, 3, 3, 64)).astype(np.float32)
    idx = ops.conv_fullfold_pack_index(3, 3, 3, 64)
    flat = np.concatenate([v.reshape(-1), [0.0]]).astype(np.float32)
    p = flat[np.where(idx < 0, flat.size - 1, idx)]
    np.testing.assert_array_equal(p[:, 0, :27], v.reshape(27, 64).T)


def test_model_surface_matches_reference_usage():
    """Shapes / params / layers / re-binding as used in upscaler_mini_testing.ipynb cells 5-7 and model.py:1132."""
    from upscaler.model import Input, Model, make_upscaler_orig
    up = make_upscaler_orig((512, 512, 3), kernel_size=3)
    assert up.input_shape == (None, 128, 128, 3) and up.output_shape == (None, 512, 512, 3)
    assert up.count_params() == 2043779
    assert make_upscaler_orig((384, 384, 3)).count_params() == 5517187  # kernel_size default 5 (model.py:267)
    names = [l.name for l in up.layers]
    assert names[0] == "initial/input" and "res_block/15/conv_post" in names and "final/conv" in names
    up.layers.pop(0)
    inp = Input(shape=(270, 480, 3))
    full = Model(inp, up(inp))
    assert full.input_shape == (None, 270, 480, 3) and full.output_shape == (None, 1080, 1920, 3)
    assert full.leaf_nets()[0] is up.net
    up.name = "upscaler"
    assert up.name == "upscaler"
    with pytest.raises(ValueError):
        make_upscaler_orig((512, 512, 1))


def test_data_contract():
    from upscaler import data
    u = np.arange(256, dtype=np.uint8).reshape(16, 16, 1).repeat(3, 2)
    a = data.convert_image_series_to_array([u, u])
    assert a.shape == (2, 16, 16, 3) and a.min() == -1 and a.max() == 1
    assert (data.convert_array_to_uint8(a[0]) == u).all()
