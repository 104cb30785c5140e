"""Execution of Model.predict / compile / train_on_batch over the device networks.

predict: host NumPy -> pinned -> device, the network's inference kernels, device -> host.  Training entry points are
resolved at compile() time from the structure the reference's compile helpers build (model.py:1017-1137).
"""
import numpy as np
import torch

from .keras_compat import as_batch


def _device():
    if not torch.cuda.is_available():
        raise RuntimeError("upscaler (B200 build) needs a CUDA device: the hot path has no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _check_input(model, a, idx=0):
    want = model._inputs[idx].shape
    got = a.shape[1:]
    for w, g in zip(want, got):
        if w is not None and w != g:
            raise ValueError("Error when checking input: expected %s to have shape %s but got array with shape %s"
                             % (model._inputs[idx].name, (None,) + tuple(want), a.shape))


def _run_chain_infer(nets, x):
    for net in nets:
        x = net.forward_infer(x)
    return x


def predict_device(model, x_dev):
    """Inference on a device tensor (float32 NHWC); returns device tensor(s)."""
    if model._is_leaf:
        return model.net.forward_infer(x_dev)
    outs = []
    cache = {}
    for o in model._outputs:
        src, chain = o.chain()
        y, key = x_dev, ()
        for m in chain:
            key = key + (id(m),)
            if key not in cache:
                cache[key] = _run_chain_infer(m.leaf_nets(), y)
            y = cache[key]
        outs.append(y)
    return outs[0] if len(outs) == 1 else outs


# Frames per pipeline stage of predict(): the device->host copy of stage i (24.9 MB per 1080p frame, PCIe) runs on a copy
# stream while stage i+1 computes, and the host-side staging copy of stage i+1 overlaps the kernels of stage i.
# (Measured alternative: stages shrinking towards the end of the batch, 8 frames -> 4, 2, 1, 1, to expose less of the last
# copy: 584 frames/s against 610 with uniform stages of 2 -- one-frame stages run the kernels at 560 frames/s.)
PIPELINE_FRAMES = 2

_pinned_in = {}
_dev_out = {}
_copy_stream = {}


def _stage_input(a, slot):
    """Host ndarray -> pinned staging buffer (cached per shape, type and pipeline slot) so the H2D copy is one async DMA."""
    key = (a.shape, slot, a.dtype.str)
    buf = _pinned_in.get(key)
    if buf is None:
        if len(_pinned_in) > 16:
            _pinned_in.clear()
        buf = torch.empty(a.shape, dtype=torch.uint8 if a.dtype == np.uint8 else torch.float32, pin_memory=True)
        _pinned_in[key] = buf
    buf.numpy()[...] = a
    return buf


def _device_out(model, shape, dev):
    """Device-side result buffer, cached per model and shape: the copy stream reads it after the compute stream has
    moved on, so it must not go back to the caching allocator between calls."""
    key = (id(model), tuple(shape), dev.index)
    t = _dev_out.get(key)
    if t is None:
        if len(_dev_out) > 4:
            _dev_out.clear()
        t = torch.empty(shape, dtype=torch.float32, device=dev)
        _dev_out[key] = t
    return t


def predict_uint8(model, frames, batch_size=None):
    """uint8 frames (N,H,W,3) -> uint8 upscaled frames: convert_image_series_to_array -> Model.predict ->
    convert_array_to_image's arithmetic (data.py:253-270, save_img_predict data.py:358-363) with both conversions on
    the device, so the PCIe traffic is the uint8 frames (a quarter of the float32 arrays) and the result is bit-identical
    to doing the two conversions in NumPy around predict()."""
    a = np.asarray(frames)
    if a.dtype != np.uint8 or a.ndim != 4:
        raise ValueError("predict_uint8 expects a uint8 array (frames, height, width, channels)")
    return predict(model, np.ascontiguousarray(a), batch_size, u8=True)


def predict(model, x, batch_size=None, u8=False):
    """Model.predict: host ndarray in, fresh host ndarray(s) out.  Results are copied device->host straight into a
    pinned allocation (torch's caching host allocator recycles it once the caller drops the array).  Frames are
    independent in inference (BatchNorm on moving statistics), so the batch is walked in stages of PIPELINE_FRAMES:
    H2D + kernels on the current stream, D2H on a copy stream one stage behind."""
    if isinstance(x, (list, tuple)):
        if len(x) != 1:
            raise ValueError("predict on multi-input models is not part of the reference path")
        x = x[0]
    a = x if u8 else as_batch(x)
    _check_input(model, a)
    dev = _device()
    from . import ops
    n = a.shape[0]
    if n == 0:                                   # Keras returns an empty array of the output shape
        shapes = model.output_shape if isinstance(model.output_shape, list) else [model.output_shape]
        res = []
        for shp in shapes:
            dims = tuple(shp[1:])
            if any(d is None for d in dims):     # fully-convolutional model: derive from the input's spatial size
                dims = tuple(model.net.output_shape_for(tuple(a.shape[1:]))) if model._is_leaf else dims
            res.append(np.zeros((0,) + tuple(int(d) for d in dims), np.uint8 if u8 else np.float32))
        return res[0] if len(res) == 1 else res
    bs = int(batch_size) if batch_size else 32  # Keras' predict default
    step = max(1, min(bs, PIPELINE_FRAMES))
    main = torch.cuda.current_stream()
    cs = _copy_stream.get(dev.index)
    if cs is None:
        cs = _copy_stream[dev.index] = torch.cuda.Stream(device=dev)
    outs = devs = None
    staged = []          # (pinned buffer, event after its H2D) of the two slots in flight
    for ci, i in enumerate(range(0, n, step)):
        slot = ci & 1
        if len(staged) > slot and staged[slot] is not None:
            staged[slot].synchronize()      # the H2D that last read this staging slot has finished
        h = _stage_input(a[i:i + step], slot)
        xd = h.to(dev, non_blocking=True)
        if u8:
            xd = ops.u8_to_f32_norm(xd)
        ev_in = torch.cuda.Event()
        ev_in.record(main)
        if len(staged) <= slot:
            staged.append(ev_in)
        else:
            staged[slot] = ev_in
        if model._is_leaf and devs is not None:
            y = model.net.forward_infer(xd, out=devs[0][i:i + xd.shape[0]])
        else:
            y = predict_device(model, xd)
        ys = y if isinstance(y, list) else [y]
        if u8:
            ys = [ops.f32_to_u8_denorm(t.float().contiguous()) for t in ys]
        if outs is None:
            outs = [torch.empty((n,) + tuple(t.shape[1:]), dtype=torch.uint8 if u8 else torch.float32, pin_memory=True)
                    for t in ys]
            if model._is_leaf and n > step and getattr(model.net, 'writes_out', False) and not u8:
                # later stages write straight into slices of one cached device buffer
                devs = [_device_out(model, (n,) + tuple(ys[0].shape[1:]), dev)]
        done = torch.cuda.Event()
        done.record(main)
        cs.wait_event(done)
        with torch.cuda.stream(cs):
            for o, t in zip(outs, ys):
                o[i:i + t.shape[0]].copy_(t, non_blocking=True)
                t.record_stream(cs)
    cs.synchronize()
    res = [o.numpy() for o in outs]
    return res[0] if len(res) == 1 else res


def compile_model(model, loss, optimizer, loss_weights):
    from . import training
    return training.compile_model(model, loss, optimizer, loss_weights)


def train_on_batch(model, x, y):
    from . import training
    return training.train_on_batch(model, x, y)
