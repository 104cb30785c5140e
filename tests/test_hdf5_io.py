"""Keras-HDF5 persistence without an HDF5 library (upscaler/hdf5_io.py).

The reader is pinned against a file written by the real libhdf5: scipy ships a MATLAB v7.3 test file (HDF5 with a
512-byte user block, symbol-table group, version-1 object headers) whose content is known (9 doubles k*pi/4 and the
attribute MATLAB_class = 'double').  The writer follows the same on-disk conventions as that file (checked field by
field below) and is exercised through writer -> reader round trips and Model.save() -> load_model()."""
import os
import struct
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
from upscaler import hdf5_io as H  # noqa: E402


def _real_file():
    try:
        import scipy.io.matlab as m
    except ImportError:
        return None
    p = os.path.join(os.path.dirname(m.__file__), "tests", "data", "testhdf5_7.4_GLNX86.mat")
    return p if os.path.exists(p) else None


@pytest.mark.skipif(_real_file() is None, reason="scipy's HDF5 test file is not installed")
def test_reader_on_a_file_written_by_libhdf5():
    g = H.read_h5(_real_file())
    assert list(g.children) == ["testdouble"]
    d = g.children["testdouble"]
    np.testing.assert_allclose(d.data.ravel(), np.arange(9) * np.pi / 4, rtol=1e-15)
    assert d.data.shape == (9, 1) and d.data.dtype == np.float64
    assert H.attrs_dict(d)["MATLAB_class"] == b"double"


@pytest.mark.skipif(_real_file() is None, reason="scipy's HDF5 test file is not installed")
def test_writer_uses_the_on_disk_conventions_of_libhdf5(tmp_path):
    """Same structures, same field layout: superblock v0, local heap (free-list terminator 1), B-tree key order,
    SNOD entries, object header prefix, datatype / dataspace / attribute encodings."""
    real = open(_real_file(), "rb").read()
    base = 512
    root = H.H5Group()
    ds = root.create_dataset("testdouble", (np.arange(9) * np.pi / 4).reshape(9, 1))
    ds.attrs.append(("MATLAB_class", b"double"))
    p = str(tmp_path / "mine.h5")
    H.write_h5(p, root)
    mine = open(p, "rb").read()
    # superblock: signature, versions, sizes, then (after K values / flags) base, free-space, eof, driver addresses
    assert mine[:16] == real[base:base + 16]
    assert struct.unpack_from("<Q", mine, 40)[0] == len(mine)
    assert struct.unpack_from("<II", mine, 72) == struct.unpack_from("<II", real, base + 72) == (1, 0)   # root entry cache type
    r_hdr, r_bt, r_hp = [struct.unpack_from("<Q", real, base + o)[0] + base for o in (64, 80, 88)]
    m_hdr, m_bt, m_hp = [struct.unpack_from("<Q", mine, o)[0] for o in (64, 80, 88)]
    # local heap header and data segment: empty name at 0, name at 8, free block {next = 1, size}
    assert real[r_hp:r_hp + 8] == mine[m_hp:m_hp + 8] == b"HEAP\x00\x00\x00\x00"
    r_seg, m_seg = struct.unpack_from("<Q", real, r_hp + 24)[0] + base, struct.unpack_from("<Q", mine, m_hp + 24)[0]
    assert real[r_seg:r_seg + 24] == mine[m_seg:m_seg + 24]
    r_free, m_free = struct.unpack_from("<Q", real, r_hp + 16)[0], struct.unpack_from("<Q", mine, m_hp + 16)[0]
    assert r_free == m_free == 24
    assert struct.unpack_from("<Q", real, r_seg + r_free)[0] == struct.unpack_from("<Q", mine, m_seg + m_free)[0] == 1
    # B-tree leaf: header, key 0 = 0, child, key 1 = heap offset of the largest name
    assert real[r_bt:r_bt + 32] == mine[m_bt:m_bt + 32]
    assert struct.unpack_from("<Q", real, r_bt + 24)[0] == struct.unpack_from("<Q", mine, m_bt + 24)[0] == 0
    assert struct.unpack_from("<Q", real, r_bt + 40)[0] == struct.unpack_from("<Q", mine, m_bt + 40)[0] == 8
    r_sn, m_sn = struct.unpack_from("<Q", real, r_bt + 32)[0] + base, struct.unpack_from("<Q", mine, m_bt + 32)[0]
    assert real[r_sn:r_sn + 16] == mine[m_sn:m_sn + 16]                  # "SNOD", version, count, name offset of entry 0
    # object header prefix of the root group + its symbol-table message header
    assert real[r_hdr:r_hdr + 2] == mine[m_hdr:m_hdr + 2] == b"\x01\x00"
    assert real[r_hdr + 16:r_hdr + 20] == mine[m_hdr + 16:m_hdr + 20] == b"\x11\x00\x10\x00"
    # datatype (IEEE double), dataspace (9 x 1) and attribute message bodies are byte-identical
    r_ds = struct.unpack_from("<Q", real, r_sn + 16)[0] + base
    m_ds = struct.unpack_from("<Q", mine, m_sn + 16)[0]

    def msgs(buf, hdr):
        n, q, out = struct.unpack_from("<H", buf, hdr + 2)[0], hdr + 16, {}
        for _ in range(n):
            t, s = struct.unpack_from("<HH", buf, q)
            out[t] = buf[q + 8:q + 8 + s]
            q += 8 + s
        return out
    rm, mm = msgs(real, r_ds), msgs(mine, m_ds)
    assert rm[3][:20] == mm[3][:20]
    assert rm[1] == mm[1]
    # attribute: same layout; libhdf5 wrote a null-terminated string type (bit field 0), h5py/this writer null-padded (1)
    assert rm[12][:25] == mm[12][:25] and rm[12][26:] == mm[12][26:] and (rm[12][25], mm[12][25]) == (0, 1)


def test_writer_reader_round_trip_edge_cases(tmp_path):
    root = H.H5Group()
    root.attrs += [("s", b"abc"), ("u", "hé"), ("i", np.int64(-7)), ("f", np.float32(1.5)),
                   ("names", [b"a", b"bcd/e:0"]), ("empty", np.zeros((0,), "S1"))]
    root.require_group("empty_group")
    rng = np.random.default_rng(0)
    arrays = {"g/a:0": rng.standard_normal((3, 3, 64, 64)).astype(np.float32), "g/h/i/j": np.arange(5, dtype=np.int32),
              "g/d64": rng.standard_normal((7,)), "g/scalar": np.float32(3.0), "g/zero": np.zeros((0, 4), np.float32)}
    for k, v in arrays.items():
        root.create_dataset(k, v)
    for i in range(300):                                   # > 128 members: several symbol-table nodes under one B-tree
        root.create_dataset("many/l%03d" % i, np.full((2,), i, np.float32))
    p = str(tmp_path / "t.h5")
    H.write_h5(p, root)
    g = H.read_h5(p)
    a = H.attrs_dict(g)
    assert a["s"] == b"abc" and a["u"] == "hé".encode("utf8") and a["i"] == -7 and a["f"] == 1.5
    assert list(a["names"]) == [b"a", b"bcd/e:0"] and a["empty"].shape == (0,)
    assert g.children["empty_group"].children == {}
    for k, v in arrays.items():
        got = H.get_path(g, k).data
        assert np.array_equal(got, v) and got.dtype == np.asarray(v).dtype and got.shape == np.asarray(v).shape
    many = H.get_path(g, "many").children
    assert list(many) == ["l%03d" % i for i in range(300)]
    assert all(many["l%03d" % i].data[0] == i for i in range(300))
    with pytest.raises(ValueError):
        H.read_h5(__file__)


def test_long_attribute_lists_are_split_like_keras(tmp_path):
    g = H.H5Group()
    names = [("layer_%05d/with/a/long/name/kernel:0" % i).encode() for i in range(4000)]
    H._split_attr(g, "weight_names", names)
    assert len(g.attrs) > 1 and all(n.startswith("weight_names") for n, _ in g.attrs)
    p = str(tmp_path / "t.h5")
    H.write_h5(p, g)
    assert H._load_attr_list(H.read_h5(p), "weight_names") == [n.decode() for n in names]


def test_model_save_and_load_model_round_trip(tmp_path):
    from upscaler.model import load_model, make_discriminator_simple_512, make_upscaler_orig
    up = make_upscaler_orig((64, 96, 3), kernel_size=3, res_block_num=3)
    p = str(tmp_path / "model_upscaler.h5")
    up.save(p)
    attrs, weights = H.read_keras_weights(p)
    assert attrs["keras_version"] == b"2.2.4" and attrs["backend"] == b"tensorflow"
    assert weights["res_block/0/conv_pre"][0].shape == (3, 3, 64, 64) and weights["initial/prelu"][0].shape == (1, 1, 64)
    assert len(weights["res_block/2/batch_norm_post"]) == 4 and weights["final/tanh"] == []
    root = H.read_h5(p)
    # the nested-group layout h5py produces for Keras' slash-carrying names
    assert H.get_path(root, "model_weights/res_block/0/conv_pre/res_block/0/conv_pre/kernel:0").data.shape == (3, 3, 64, 64)
    m2 = load_model(p, custom_objects={"loss": None})
    assert m2.input_shape == (None, 16, 24, 3) and m2.output_shape == (None, 64, 96, 3)
    assert m2.net.kernel_size == 3 and m2.net.res_block_num == 3
    a, b = up.net.host_weights(), m2.net.host_weights()
    assert all(np.array_equal(x, y) for k in a for x, y in zip(a[k], b[k]))
    d = make_discriminator_simple_512((32, 32, 3), activation="sigmoid")
    pd = str(tmp_path / "disc.h5")
    d.save(pd)
    d2 = load_model(pd)
    assert d2.input_shape == (None, 32, 32, 3) and d2.net.activation == "sigmoid" and d2.count_params() == d.count_params()
    a, b = d.net.host_weights(), d2.net.host_weights()
    assert all(np.array_equal(x, y) for k in a for x, y in zip(a[k], b[k]))


def test_save_and_load_model_pick_the_variant_from_the_weight_shapes(tmp_path):
    """skip-con generator (final conv over 227 channels) and thin discriminator (128-wide blocks) survive save -> load_model."""
    from upscaler.model import load_model, make_discriminator_thin_512, make_upscaler_skip_con
    from upscaler import networks
    up = make_upscaler_skip_con((64, 96, 3), kernel_size=3)
    p = str(tmp_path / "skip_con.h5")
    up.save(p)
    m2 = load_model(p)
    assert isinstance(m2.net, networks.UpscalerSkipConNet) and m2.output_shape == (None, 64, 96, 3)
    a, b = up.net.host_weights(), m2.net.host_weights()
    assert a["final/conv"][0].shape == (9, 9, 227, 3) and a["upscaling/1/block/conv_transp"][0].shape == (3, 3, 224, 224)
    assert all(np.array_equal(x, y) for k in a for x, y in zip(a[k], b[k]))
    d = make_discriminator_thin_512((32, 32, 3))
    pd = str(tmp_path / "thin.h5")
    d.save(pd)
    d2 = load_model(pd)
    assert isinstance(d2.net, networks.DiscriminatorThin512Net) and d2.count_params() == d.count_params()
    from upscaler.model import make_discriminator_sparse_512
    ds = make_discriminator_sparse_512((512, 512, 3), activation="tanh")
    assert ds.count_params() == 5987777          # Keras summary of model.py:964-1012 at (512, 512, 3)
    ps = str(tmp_path / "sparse.h5")
    ds.save(ps)
    ds2 = load_model(ps)
    assert isinstance(ds2.net, networks.DiscriminatorSparse512Net) and ds2.net.activation == "tanh"
    a, b = ds.net.host_weights(), ds2.net.host_weights()
    assert a["discriminator/block_2/Conv2d"][0].shape == (5, 5, 64, 128) and a["discriminator/final/Dense_1"][0].shape == (256, 128)
    assert all(np.array_equal(x, y) for k in a for x, y in zip(a[k], b[k]))
    with pytest.raises(ValueError):
        make_discriminator_sparse_512((256, 256, 3))


def test_load_model_resolves_keras_auto_named_generator_files(tmp_path):
    """example_movie_frames/model_upscaler_orig_vgg-mse_best.h5 (example_frame_test.ipynb cell 9) was written by
    old/train_orig_vgg-mse.py, which names no layer: the file carries Keras' per-class counters (conv2d_N,
    batch_normalization_N, p_re_lu_N, conv2d_transpose_N, plus weightless add_N / leaky_re_lu_N / activation_N) and a
    width-first (480, 270, 3) input.  A synthetic file in exactly that layout -- counters NOT starting at 1, as after an
    earlier model in the same process, layer_names in graph order -- must load into make_upscaler_orig with every weight in
    the right layer."""
    import json
    from upscaler.model import load_model, make_upscaler_orig
    rng = np.random.default_rng(5)
    up = make_upscaler_orig((64, 96, 3), kernel_size=3, res_block_num=2)
    host = {k: [rng.standard_normal(a.shape).astype(np.float32) for a in v] for k, v in up.net.host_weights().items()}
    up.net.set_weights(host)
    counters = {"conv2d": 6, "batch_normalization": 3, "p_re_lu": 11, "conv2d_transpose": 0, "add": 1, "leaky_re_lu": 4, "activation": 0}
    suffixes = {"conv": ("conv2d", ["kernel:0", "bias:0"]), "convT": ("conv2d_transpose", ["kernel:0", "bias:0"]),
                "bn": ("batch_normalization", ["gamma:0", "beta:0", "moving_mean:0", "moving_variance:0"]), "prelu": ("p_re_lu", ["alpha:0"])}

    def auto(cls):
        counters[cls] += 1
        return "%s_%d" % (cls, counters[cls])

    root = H.H5Group()
    cfg = {"class_name": "Model", "config": {"name": "model_1", "layers": [
        {"class_name": "InputLayer", "name": "input_1", "config": {"batch_input_shape": [None, 480, 270, 3], "dtype": "float32", "name": "input_1"}}]}}
    root.attrs += [("keras_version", b"2.2.4"), ("backend", b"tensorflow"), ("model_config", json.dumps(cfg).encode("utf8"))]
    mw = root.require_group("model_weights")
    layer_names, mapping = ["input_1"], {}
    mw.require_group("input_1").attrs.append(("weight_names", np.zeros((0,), dtype="S1")))

    def weightless(cls):
        n = auto(cls)
        layer_names.append(n)
        mw.require_group(n).attrs.append(("weight_names", np.zeros((0,), dtype="S1")))

    for name, kind, _ in up.net.specs:
        cls, sfx = suffixes[kind]
        n = auto(cls)
        mapping[name] = n
        layer_names.append(n)
        g = mw.require_group(n)
        wn = [("%s/%s" % (n, s_)).encode("utf8") for s_ in sfx]
        g.attrs.append(("weight_names", np.array(wn)))
        for w, a in zip(wn, host[name]):
            g.create_dataset(w.decode("utf8"), a)
        if name.endswith("batch_norm_post") or name == "prefinal/batch_norm":
            weightless("add")
        if kind == "convT":
            weightless("leaky_re_lu")
    weightless("activation")
    mw.attrs.append(("layer_names", np.array([n.encode("utf8") for n in layer_names])))
    p = str(tmp_path / "model_upscaler_orig_vgg-mse_best.h5")
    H.write_h5(p, root)
    assert mapping["initial/conv"] == "conv2d_7" and mapping["final/conv"] == "conv2d_13" and mapping["res_block/1/prelu"] == "p_re_lu_14"
    m = load_model(p, custom_objects={"loss": None})
    assert m.net.kernel_size == 3 and m.net.res_block_num == 2
    assert m.input_shape == (None, 480, 270, 3) and m.output_shape == (None, 1920, 1080, 3)      # width-first, as the file says
    got = m.net.host_weights()
    for k in host:
        assert all(np.array_equal(x, y) for x, y in zip(host[k], got[k])), k
    # a file whose layer counts do not form the generator's graph is refused, not mis-assigned
    del mw.children[mapping["res_block/1/batch_norm_post"]]
    mw.attrs[-1] = ("layer_names", np.array([n.encode("utf8") for n in layer_names if n != mapping["res_block/1/batch_norm_post"]]))
    H.write_h5(p, root)
    with pytest.raises(ValueError, match="auto-named"):
        load_model(p)
