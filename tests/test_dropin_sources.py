"""CPU: drift check of the restated script snippets in tests/test_dropin_scripts_gpu.py against the reference's own
scripts, WHEN the reference checkout is present (it is in the build container, not on the GPU box -- there the GPU test runs
the restated text, and this test skips).  Every `from upscaler...` import line and every statement of the loop bodies that
touches the model API (train_on_batch / predict / the batch assembly around them) must appear, whitespace-normalised, in
the reference file it is cited from; a reference script that changes its imports or its loop makes this fail here instead
of silently passing on stale text there.  Nothing is copied: the reference text is only read and compared."""
import importlib.util
import os
import re

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("B2_REFERENCE_DIR", "/root/reference")
SCRIPTS = {"TRAIN_MINI": "upscaling/train_mini.py", "TRAIN_GAN3": "upscaling/train_gan3.py"}
# statements of the restated loop bodies that must be the reference's own (anything that calls into upscaler.*)
API_WORDS = ("train_on_batch", ".predict(", "select_random_rows", "convert_image_series_to_array", "pd.concat", "np.ones(")


def _norm(line):
    line = line.split("#")[0].replace("values.", "")       # argparse namespace of the scripts; plain variables in the snippets
    return re.sub(r"\s+", "", line)


def _load_snippets():
    spec = importlib.util.spec_from_file_location("_dropin_snippets", os.path.join(HERE, "test_dropin_scripts_gpu.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.parametrize("which", sorted(SCRIPTS))
def test_restated_snippets_follow_the_reference_scripts(which):
    path = os.path.join(REF, SCRIPTS[which])
    if not os.path.exists(path):
        pytest.skip("reference checkout not present (%s): the GPU test runs the restated snippets" % path)
    with open(path, encoding="utf-8", errors="replace") as f:
        ref_lines = {_norm(l) for l in f if l.strip()}
    text = getattr(_load_snippets(), which)
    checked = 0
    in_loop = False
    for raw in text.splitlines():
        if raw.startswith("for b in range("):
            in_loop = True
            continue
        if in_loop and raw and not raw.startswith(" "):
            in_loop = False
        is_import = raw.startswith("from upscaler.")
        is_api = in_loop and any(w in raw for w in API_WORDS)
        if not (is_import or is_api):
            continue
        assert _norm(raw) in ref_lines, "%s: restated line not found in %s:\n    %s" % (which, SCRIPTS[which], raw.strip())
        checked += 1
    assert checked >= 10, checked
