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


_pinned_in = {}


def _stage_input(a):
    """Host ndarray -> pinned staging buffer (cached per shape) so the H2D copy is a single async DMA."""
    key = a.shape
    buf = _pinned_in.get(key)
    if buf is None:
        if len(_pinned_in) > 8:
            _pinned_in.clear()
        buf = torch.empty(a.shape, dtype=torch.float32, pin_memory=True)
        _pinned_in[key] = buf
    buf.numpy()[...] = a
    return buf


def predict(model, x, batch_size=None):
    """Model.predict: host ndarray in, fresh host ndarray(s) out.  Results are copied device->host straight into a
    pinned allocation (torch's caching host allocator recycles it once the caller drops the array)."""
    if isinstance(x, (list, tuple)):
        if len(x) != 1:
            raise ValueError("predict on multi-input models is not part of the reference path")
        x = x[0]
    a = as_batch(x)
    _check_input(model, a)
    dev = _device()
    n = a.shape[0]
    bs = int(batch_size) if batch_size else 32  # Keras' predict default
    outs = None
    for i in range(0, n, bs):
        h = _stage_input(a[i:i + bs])
        y = predict_device(model, h.to(dev, non_blocking=True))
        ys = y if isinstance(y, list) else [y]
        if outs is None:
            outs = [torch.empty((n,) + tuple(t.shape[1:]), dtype=torch.float32, pin_memory=True) for t in ys]
        for o, t in zip(outs, ys):
            o[i:i + t.shape[0]].copy_(t, non_blocking=True)
        torch.cuda.current_stream().synchronize()  # the staging buffer is reused by the next chunk
    res = [o.numpy() for o in outs]
    return res[0] if len(res) == 1 else res


def compile_model(model, loss, optimizer, loss_weights):
    from . import training
    return training.compile_model(model, loss, optimizer, loss_weights)


def train_on_batch(model, x, y):
    from . import training
    return training.train_on_batch(model, x, y)
