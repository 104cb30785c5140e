"""Generates tests/golden/*.npz: small seeded input/output vectors of the hot path, produced by the CPU oracle
(oracle/srgan_oracle.py, PyTorch fp32).  The reference itself (Keras 2.2 / TensorFlow 1.14) cannot run in this container
(SURVEY.md 8c), so these vectors pin the ORACLE (any later change to it shows up in `-m "not gpu"` tests) and give the
GPU tests a fixed target that does not depend on re-running the oracle.  Re-run: python tools/make_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import srgan_oracle as O  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def generator_case(ks, shape, seed):
    w = O.init_upscaler_orig_weights(seed=seed, kernel_size=ks)
    x = np.random.default_rng(seed + 1).uniform(-1, 1, shape).astype(np.float32)
    with torch.no_grad():
        y = O.upscaler_orig_forward(O.to_torch(w), x, training=False).numpy()
    return {"x": x, "y": y.astype(np.float32), "kernel_size": np.int32(ks), "weight_seed": np.int32(seed)}


def vgg_case(shape, seed):
    vw = O.to_torch(O.init_vgg19_weights(seed=seed))
    x = np.random.default_rng(seed + 1).uniform(-1, 1, shape).astype(np.float32)
    with torch.no_grad():
        f = O.vgg19_block5_conv4(vw, torch.from_numpy(x)).numpy()
    return {"x": x, "features": f.astype(np.float32), "weight_seed": np.int32(seed)}


def discriminator_case(shape, seed):
    dw = O.init_discriminator_simple_512_weights(seed=seed, input_shape=shape[1:])
    x = np.random.default_rng(seed + 1).uniform(-1, 1, shape).astype(np.float32)
    with torch.no_grad():
        s = O.discriminator_simple_512_forward(O.to_torch(dw), x, "none", False).numpy()
    return {"x": x, "scores": s.astype(np.float32), "weight_seed": np.int32(seed)}


def frame_io_case():
    u8 = np.arange(256, dtype=np.uint8).reshape(1, 16, 16, 1).repeat(3, axis=3)
    f = O.convert_image_to_array(u8).astype(np.float32)
    back = O.convert_array_to_u8(f)
    probe = np.linspace(-1.0, 1.0, 2049, dtype=np.float32)   # tanh range; outside it NumPy wraps, the kernel saturates
    return {"u8": u8, "f32": f, "back": back, "probe": probe, "probe_u8": O.convert_array_to_u8(probe)}


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    cases = {
        "generator_k3_1x12x20": generator_case(3, (1, 12, 20, 3), 11),
        "generator_k5_2x9x7": generator_case(5, (2, 9, 7, 3), 12),
        "generator_k3_1x33x17": generator_case(3, (1, 33, 17, 3), 13),     # large enough for the fused-phase ConvT kernels
        "vgg19_2x32x48": vgg_case((2, 32, 48, 3), 21),
        "discriminator_4x64x64": discriminator_case((4, 64, 64, 3), 31),
        "frame_io": frame_io_case(),
    }
    for name, arrs in cases.items():
        path = os.path.join(OUT, name + ".npz")
        np.savez_compressed(path, **arrs)
        print("%-28s %7.1f KB" % (name, os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()
