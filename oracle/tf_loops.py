"""Pure-Python/NumPy loop restatement of TensorFlow's documented conv rules — TEST INFRASTRUCTURE ONLY.

Independent of the vectorised oracle (srgan_oracle.py): used by tests/test_oracle.py to pin the oracle's padding and
tap alignment on small cases.
  * tf.nn.conv2d, padding='SAME': out = ceil(in/stride); pad_total = max((out-1)*stride + k - in, 0);
    pad_before = pad_total // 2 (the extra pixel goes after).
  * keras Conv2DTranspose(strides=2, padding='same') = tf.nn.conv2d_transpose with output 2*in, i.e. the gradient
    w.r.t. the input of the SAME stride-2 conv above: every input pixel i scatters x[i] * W[k] into
    y = 2*i + k - pad_before, where pad_before is the forward conv's (input size 2*in).
"""
import numpy as np


def same_pad(size, k, s):
    out = -(-size // s)
    total = max((out - 1) * s + k - size, 0)
    return total // 2, out


def conv2d_same_loops(x, kernel, stride=1):
    """x (H,W,Cin), kernel (kh,kw,Cin,Cout) -> (OH,OW,Cout)."""
    H, W, _ = x.shape
    kh, kw, _, cout = kernel.shape
    pt, OH = same_pad(H, kh, stride)
    pl, OW = same_pad(W, kw, stride)
    out = np.zeros((OH, OW, cout), np.float64)
    for oy in range(OH):
        for ox in range(OW):
            for ky in range(kh):
                for kx in range(kw):
                    iy, ix = oy * stride + ky - pt, ox * stride + kx - pl
                    if 0 <= iy < H and 0 <= ix < W:
                        out[oy, ox] += x[iy, ix].astype(np.float64) @ kernel[ky, kx].astype(np.float64)
    return out


def conv2d_transpose_same_s2_loops(x, kernel):
    """x (H,W,Cin), Keras Conv2DTranspose kernel (kh,kw,Cout,Cin) -> (2H,2W,Cout)."""
    H, W, _ = x.shape
    kh, kw, cout, _ = kernel.shape
    pt, _ = same_pad(2 * H, kh, 2)
    pl, _ = same_pad(2 * W, kw, 2)
    out = np.zeros((2 * H, 2 * W, cout), np.float64)
    for iy in range(H):
        for ix in range(W):
            for ky in range(kh):
                for kx in range(kw):
                    y, xx = 2 * iy + ky - pt, 2 * ix + kx - pl
                    if 0 <= y < 2 * H and 0 <= xx < 2 * W:
                        out[y, xx] += kernel[ky, kx].astype(np.float64) @ x[iy, ix].astype(np.float64)
    return out


def conv2d_transpose_same_loops(x, kernel, stride):
    """Keras Conv2DTranspose(strides=stride, padding='same') for any stride <= kernel size (make_upscaler_attention uses
    kernel = stride + 1, reference model.py:95): the gradient w.r.t. the input of the SAME conv that maps stride*in -> in;
    input pixel i scatters x[i] * W[k] into y = stride*i + k - pad_before.  x (H,W,Cin), kernel (kh,kw,Cout,Cin)."""
    H, W, _ = x.shape
    kh, kw, cout, _ = kernel.shape
    pt, oh = same_pad(stride * H, kh, stride)
    pl, ow = same_pad(stride * W, kw, stride)
    assert (oh, ow) == (H, W)
    out = np.zeros((stride * H, stride * W, cout), np.float64)
    for iy in range(H):
        for ix in range(W):
            for ky in range(kh):
                for kx in range(kw):
                    y, xx = stride * iy + ky - pt, stride * ix + kx - pl
                    if 0 <= y < stride * H and 0 <= xx < stride * W:
                        out[y, xx] += kernel[ky, kx].astype(np.float64) @ x[iy, ix].astype(np.float64)
    return out


def resize_nearest_tf1_loops(img, factor):
    """tf.image.resize_nearest_neighbor (TF 1.x, align_corners=False): source index = floor(out * in / out_size)."""
    H, W = len(img), len(img[0])
    return [[img[int(oy * (H / float(H * factor)))][int(ox * (W / float(W * factor)))] for ox in range(W * factor)]
            for oy in range(H * factor)]


def resize_bilinear_tf1_loops(img, factor):
    """tf.image.resize_bilinear (TF 1.x, align_corners=False, no half-pixel centres) as plain loops over an (H, W)
    list-of-lists image: the independent restatement oracle.resize_bilinear_tf1 is checked against."""
    H, W = len(img), len(img[0])
    out = [[0.0] * (W * factor) for _ in range(H * factor)]
    for oy in range(H * factor):
        sy = oy * (H / float(H * factor))
        y0 = int(sy)
        y1 = min(y0 + 1, H - 1)
        fy = sy - y0
        for ox in range(W * factor):
            sx = ox * (W / float(W * factor))
            x0 = int(sx)
            x1 = min(x0 + 1, W - 1)
            fx = sx - x0
            top = img[y0][x0] + (img[y0][x1] - img[y0][x0]) * fx
            bot = img[y1][x0] + (img[y1][x1] - img[y1][x0]) * fx
            out[oy][ox] = top + (bot - top) * fy
    return out
