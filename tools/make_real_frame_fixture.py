"""Writes tests/golden/BtoA_00286_crop480x270.png: the reference's config-1 input frame, decoded ONCE here.

example_movie_frames/example_frame_test.ipynb cells 2-4 open BtoA_00286.jpg, convert to RGB, crop (0, 0, 480, 270) and
feed the crop (after swapaxes(1, 2) and x / 127.5 - 1) to upscaler.predict; ukiyo_lowres.jpg is a bicubic enlargement
of that same crop.  /root/reference does not travel to the GPU box, and JPEG decoding may differ between libjpeg
builds, so the decoded crop is committed as a lossless PNG (uint8, 270 x 480 x 3, ~0.3 MB).  Run in the build container:
    python tools/make_real_frame_fixture.py
"""
import os
import sys

import numpy as np
from PIL import Image

SRC = "/root/reference/example_movie_frames/BtoA_00286.jpg"
DST = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "BtoA_00286_crop480x270.png")

if __name__ == "__main__":
    img = Image.open(SRC).convert("RGB")                    # cell 2
    crop = img.crop((0, 0, 480, 270))
    a = np.array(crop)
    assert a.shape == (270, 480, 3) and a.dtype == np.uint8, a.shape
    Image.fromarray(a).save(DST, optimize=True)
    back = np.array(Image.open(DST))
    assert np.array_equal(a, back)
    print("wrote %s (%d bytes): mean %.2f, saturated pixels (0 or 255): %.2f %%" % (
        DST, os.path.getsize(DST), a.mean(), 100.0 * np.mean((a == 0) | (a == 255))))
    sys.exit(0)
