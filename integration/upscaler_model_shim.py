"""Drop-in overlay for the reference checkout: save this file as  upscaling/upscaler/model.py  (the ONLY file replaced).

The reference's scripts do `from upscaler.model import make_upscaler_orig, VGG_MSE_LOSS, compile_training_model, ...`
(train.py:5-7, train_mini.py:4-6, train_gan.py:4-7, train_gan3.py:4-7, train_cgc*.py) next to
`from upscaler.data import ...` / `from upscaler.json import PandasEncoder`.  Dataset loading, cropping, image saving and
the JSON encoder are not part of the accelerated path and stay the reference's own files; only the model builders, losses
and compile helpers -- and through them Model.predict / train_on_batch / save -- are replaced.

The B200 package is loaded under the private name `b2_upscaler` (its own sub-modules use relative imports), so it does
not collide with the reference's `upscaler` namespace package.  Point B2_UPSCALER_ROOT at the directory that holds it:
    export B2_UPSCALER_ROOT=/path/to/repo/video-cycle_gan-upscaling_b200
The scripts' direct Keras imports are the remaining two-line edit per script:
    from keras.utils.vis_utils import plot_model   ->   plot_model = lambda *a, **k: None
    from keras.models import load_model            ->   from upscaler.model import load_model
"""
import importlib.util
import os
import sys

_root = os.environ.get("B2_UPSCALER_ROOT")
if not _root or not os.path.isfile(os.path.join(_root, "upscaler", "model.py")):
    raise ImportError("upscaler.model (B200 overlay): set B2_UPSCALER_ROOT to the directory containing the B200 `upscaler` "
                      "package (video-cycle_gan-upscaling_b200)")
if "b2_upscaler" not in sys.modules:
    _pkg = os.path.join(_root, "upscaler")
    _spec = importlib.util.spec_from_file_location("b2_upscaler", os.path.join(_pkg, "__init__.py"),
                                                   submodule_search_locations=[_pkg])
    _mod = importlib.util.module_from_spec(_spec)
    sys.modules["b2_upscaler"] = _mod
    _spec.loader.exec_module(_mod)

from b2_upscaler.model import *          # noqa: E402,F401,F403  every public name of the reference's model.py
from b2_upscaler.model import load_model, Adam, Input, Model  # noqa: E402,F401
