import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "video-cycle_gan-upscaling_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")
    config.addinivalue_line("markers", "slow: minutes of fp32 CPU-oracle autograd beside the GPU run; skipped unless B2_SLOW=1 "
                                       "(their logs are committed under profiles/)")


def pytest_collection_modifyitems(config, items):
    import torch
    if os.environ.get("B2_SLOW", "0") in ("", "0"):
        skip_slow = pytest.mark.skip(reason="slow oracle comparison: set B2_SLOW=1 (log of the last run under profiles/)")
        for item in items:
            if "slow" in item.keywords:
                item.add_marker(skip_slow)
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
