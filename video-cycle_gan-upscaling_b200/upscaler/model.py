"""Drop-in for the reference's upscaler.model (upscaling/upscaler/model.py) on the SRGAN hot path.

Same builder / loss / compile-helper names and argument meaning; the returned objects honour the Keras Model surface
the reference uses (predict, train_on_batch, save, input_shape, output_shape, layers, name, trainable).  Underneath,
every layer runs as hand-written sm_100a CUDA through libb2srgan.so — there is no Keras, TensorFlow or CPU path.
"""
import math
from abc import ABCMeta, abstractmethod

from .keras_compat import Adam, Input, Model  # noqa: F401  (re-exported: scripts import these from keras)
from . import networks


def make_upscaler_orig(output_image_shape, kernel_size=5, filters=64, upscale_factor=4, res_block_num=16):
    """model.py:267-295 — SRGAN generator; returns a Model whose input is output_image_shape // upscale_factor."""
    input_image_shape = (output_image_shape[0] // upscale_factor, output_image_shape[1] // upscale_factor,
                         output_image_shape[2])
    if output_image_shape[2] != 3:
        raise ValueError("make_upscaler_orig: the final conv emits 3 channels (model.py:290)")
    int(math.log(upscale_factor, 2))
    net = networks.UpscalerOrigNet(kernel_size=kernel_size, filters=filters, upscale_factor=upscale_factor,
                                   res_block_num=res_block_num)
    return Model(net=net, input_shape=input_image_shape, name="model_1")
