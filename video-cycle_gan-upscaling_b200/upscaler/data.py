"""Tensor I/O contract of the reference's upscaler.data (upscaling/upscaler/data.py:253-277): uint8 HWC images
<-> float NHWC arrays in [-1, 1].  Dataset loading / cropping / JPEG writers are host-side I/O outside the hot path.
"""
import numpy as np


def convert_array_to_image(array):
    """data.py:253-256 — float [-1,1] HWC -> PIL image (uint8, round-half-even like np.around)."""
    from PIL import Image
    return Image.fromarray(convert_array_to_uint8(array))


def convert_array_to_uint8(array):
    return np.uint8(np.around((np.asarray(array) + 1) * 127.5))


def convert_image_to_array(img):
    """data.py:259-263."""
    array = np.array(img)
    return (array / 127.5) - 1


def convert_image_series_to_array(image_series):
    """data.py:266-270."""
    array = np.array([np.array(img) for img in image_series])
    return (array / 127.5) - 1


def convert_imagesdf_to_arrays(images_df):
    """data.py:273-277."""
    return convert_image_series_to_array(images_df.image_hr), convert_image_series_to_array(images_df.image_lr)


def upscale_images(images_series, upscaler, batch_size=8):
    """The loop of save_img_predict / save_images_predicted (data.py:314, 338, 358-363) without the file writes:
    PIL images (or uint8 HWC arrays) of one size -> list of upscaled PIL images.  The uint8 <-> float conversions run on
    the device (Model.predict_uint8), the arithmetic and rounding are those of convert_image_series_to_array and
    convert_array_to_image."""
    from PIL import Image
    frames = np.stack([np.asarray(img, dtype=np.uint8) for img in images_series])
    out = upscaler.predict_uint8(frames, batch_size=batch_size)
    return [Image.fromarray(o) for o in out]
