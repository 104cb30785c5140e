"""CPU: the ctypes mirror of b2_epilogue has the size and field offsets of the C struct (a binder following a
stale struct definition hands the library a short struct), and the overlay shim of INTEGRATION.md resolves the package."""
import ctypes
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_epilogue_struct_layout_matches_the_library():
    from upscaler import _lib
    lib = _lib.load()
    assert lib.b2_sizeof_epilogue() == ctypes.sizeof(_lib.Epilogue)
    n = lib.b2_epilogue_field_count()
    assert n == len(_lib.Epilogue._fields_)
    for i, (name, _) in enumerate(_lib.Epilogue._fields_):
        assert lib.b2_epilogue_field_offset(i) == getattr(_lib.Epilogue, name).offset, name


def test_integration_md_shows_the_current_struct():
    text = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    from upscaler import _lib
    for name, _ in _lib.Epilogue._fields_:
        assert '("%s"' % name in text, "INTEGRATION.md's b2_epilogue ctypes stub lacks field %r" % name


def test_overlay_shim_imports_every_name_the_reference_scripts_import(tmp_path):
    """integration/upscaler_model_shim.py as upscaler/model.py of a namespace package, in a fresh interpreter without the
    test harness's sys.path: the import lines of train.py:5-7, train_gan.py:4-7 and train_gan3.py:4-7."""
    pkg = tmp_path / "upscaler"
    pkg.mkdir()
    (pkg / "model.py").write_text(open(os.path.join(ROOT, "integration", "upscaler_model_shim.py")).read())
    (tmp_path / "s.py").write_text(
        "from upscaler.model import make_upscaler_skip_con, make_upscaler_orig, make_upscaler_unetish, make_upscaler_unetish_add, make_upscaler_attention, make_upscaler_incep_resnet\n"
        "from upscaler.model import make_discriminator_simple_512, make_discriminator_sparse_512, make_discriminator_thin_512\n"
        "from upscaler.model import VGG_LOSS, VGG_MSE_LOSS, VGG_MAE_LOSS, WassersteinLosses, RelativisticLosses, wasserstein_loss\n"
        "from upscaler.model import make_and_compile_gan, make_and_compile_gan2, compile_training_model, load_model\n"
        "print('IMPORTS_OK')\n")
    env = dict(os.environ, B2_UPSCALER_ROOT=os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
    env.pop("PYTHONPATH", None)
    res = subprocess.run([sys.executable, "s.py"], cwd=str(tmp_path), env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0 and "IMPORTS_OK" in res.stdout, res.stderr[-2000:]
    env.pop("B2_UPSCALER_ROOT")
    res = subprocess.run([sys.executable, "s.py"], cwd=str(tmp_path), env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode != 0 and "B2_UPSCALER_ROOT" in res.stderr
