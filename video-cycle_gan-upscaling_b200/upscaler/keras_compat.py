"""The slice of the Keras functional API the reference's hot path touches, backed by the B200 networks.

The reference (upscaling/upscaler/model.py, train*.py, the *_testing notebooks) only ever uses: Input(shape=...),
calling a Model on an Input / on another Model's output, Model(inputs=..., outputs=...), .compile(loss, optimizer,
loss_weights), .predict, .train_on_batch, .save, .layers (pop(0)), .input_shape / .output_shape, .name, .trainable,
.summary(), count_params().  This module provides exactly that surface; the arithmetic is in networks.py / ops.py.
"""
import numpy as np
import torch


class Input:
    """keras.layers.Input(shape=...) placeholder (model.py:273, 1028, 1132)."""

    def __init__(self, shape=None, name=None, **_):
        if shape is None:
            raise ValueError("Input needs a shape")
        self.shape = tuple(shape)
        self.name = name or "input"


class Applied:
    """Symbolic result of calling a Model on an Input or on another Applied (e.g. discriminator(generator(x)))."""

    def __init__(self, model, source):
        self.model = model
        self.source = source

    def chain(self):
        """[(model, ...)] from the Input outwards."""
        node, out = self, []
        while isinstance(node, Applied):
            out.append(node.model)
            node = node.source
        return node, out[::-1]


class Adam:
    """keras.optimizers.Adam defaults (lr 1e-3, beta_1 0.9, beta_2 0.999, epsilon 1e-7 = K.epsilon()).  One instance
    may be shared by several compiled models: `iterations` is then shared too while the moment slots stay per model
    (the default-argument quirk of model.py:1026, 1066, 1130 — SURVEY.md 5.9)."""

    def __init__(self, lr=1e-3, beta_1=0.9, beta_2=0.999, epsilon=None, decay=0.0, **_):
        self.lr, self.beta_1, self.beta_2 = float(lr), float(beta_1), float(beta_2)
        self.epsilon = 1e-7 if epsilon is None else float(epsilon)
        self.decay = float(decay)
        self.iterations = 0


class Layer:
    """Minimal stand-in for the entries of Model.layers (name + weights), enough for .pop(0), len() and iteration."""

    def __init__(self, name, kind, weights=None):
        self.name, self.kind = name, kind
        self._weights = weights or []
        self.trainable = True

    def get_weights(self):
        return [w.detach().cpu().numpy() for w in self._weights]

    def count_params(self):
        return int(sum(w.numel() for w in self._weights))


class Model:
    """A network bound to an input shape.  `net` owns weights and kernels; several Models may share one net (the
    reference re-wraps the generator with a new Input to change the spatial size: upscaler_mini_testing.ipynb
    cells 6-7, model.py:1132-1134)."""

    def __init__(self, inputs=None, outputs=None, name=None, net=None, input_shape=None):
        self.trainable = True
        self.name = name or "model"
        self._compiled = None
        if net is not None:
            self.net = net
            self._inputs = [Input(shape=input_shape, name=getattr(net, "input_name", "input"))]
            self._outputs = None
            self._chain = [self]
            self._is_leaf = True
        else:
            if inputs is None or outputs is None:
                raise ValueError("Model needs inputs and outputs")
            self._inputs = list(inputs) if isinstance(inputs, (list, tuple)) else [inputs]
            self._outputs = list(outputs) if isinstance(outputs, (list, tuple)) else [outputs]
            self._is_leaf = False
            for o in self._outputs:
                if not isinstance(o, Applied):
                    raise ValueError("outputs must come from calling a Model on an Input")
            src, chain = self._outputs[0].chain()
            self.net = chain[0].net if len(chain) == 1 and len(self._outputs) == 1 else None
            self._chain = chain
        self.layers = self._make_layers()

    # ------------------------------------------------------------------ structure
    def _make_layers(self):
        if self._is_leaf:
            ls = [Layer(self._inputs[0].name, "input")]
            for name, kind, _ in self.net.specs:
                ls.append(Layer(name, kind, self.net.weights.get(name)))
            ls += [Layer(n, k) for n, k in getattr(self.net, "extra_layers", [])]
            return ls
        ls = [Layer(i.name, "input") for i in self._inputs]
        seen = []
        for o in self._outputs:
            for m in o.chain()[1]:
                if m not in seen:
                    seen.append(m)
                    ls.append(m)
        return ls

    def __call__(self, x):
        if isinstance(x, (Input, Applied)):
            return Applied(self, x)
        raise TypeError("a Model can be called on an Input or on another model's output")

    def leaf_nets(self):
        """Underlying networks in application order (wrappers of wrappers are flattened)."""
        if self._is_leaf:
            return [self.net]
        nets = []
        for m in self._chain:
            nets += m.leaf_nets()
        return nets

    def leaf_models(self):
        """The leaf Models (the ones that own a network) in application order."""
        if self._is_leaf:
            return [self]
        out = []
        for m in self._chain:
            out += m.leaf_models()
        return out

    @property
    def input_shape(self):
        shapes = [(None,) + tuple(i.shape) for i in self._inputs]
        return shapes[0] if len(shapes) == 1 else shapes

    @property
    def output_shape(self):
        if self._is_leaf:
            return (None,) + tuple(self.net.output_shape_for(self._inputs[0].shape))
        outs = []
        for o in self._outputs:
            src, chain = o.chain()
            shape = tuple(src.shape)
            for m in chain:
                for net in m.leaf_nets():
                    shape = tuple(net.output_shape_for(shape))
            outs.append((None,) + shape)
        return outs[0] if len(outs) == 1 else outs

    def count_params(self):
        return int(sum(net.count_params() for net in self.leaf_nets()))

    def get_weights(self):
        out = []
        for net in self.leaf_nets():
            for ts in net.weights.values():
                out += [t.detach().cpu().numpy() for t in ts]
        return out

    def summary(self, line_length=None, print_fn=print):
        print_fn("Model: %s" % self.name)
        print_fn("input_shape %s -> output_shape %s" % (self.input_shape, self.output_shape))
        for net in self.leaf_nets():
            for name, kind, cfg in net.specs:
                n = sum(t.numel() for t in net.weights.get(name, []))
                print_fn("%-48s %-8s %10d" % (name, kind, n))
        print_fn("Total params: %d" % self.count_params())

    # ------------------------------------------------------------------ inference
    def predict(self, x, batch_size=None, verbose=0):
        """Model.predict: NumPy NHWC float in, float32 NumPy out; BatchNorm in inference mode (moving statistics)."""
        from . import engine
        return engine.predict(self, x, batch_size)

    def predict_uint8(self, frames, batch_size=None):
        """Extension used by the frame-upscaling helpers (upscaler.data.upscale_images): uint8 frames in, uint8 frames
        out, bit-identical to convert_image_series_to_array -> predict -> convert_array_to_image's arithmetic."""
        from . import engine
        return engine.predict_uint8(self, frames, batch_size)

    # ------------------------------------------------------------------ training
    def compile(self, loss=None, optimizer=None, loss_weights=None, **_):
        from . import engine
        self._compiled = engine.compile_model(self, loss, optimizer if optimizer is not None else Adam(), loss_weights)

    def train_on_batch(self, x, y=None, **_):
        from . import engine
        if self._compiled is None:
            raise RuntimeError("You must compile a model before training/testing. Use `model.compile(optimizer, loss)`.")
        return engine.train_on_batch(self, x, y)

    # ------------------------------------------------------------------ persistence
    def save(self, filepath, overwrite=True, include_optimizer=True):
        from . import hdf5_io
        hdf5_io.save_model(self, filepath)

    def save_weights(self, filepath):
        from . import hdf5_io
        hdf5_io.save_model(self, filepath)


def as_batch(x):
    a = np.asarray(x)
    if a.ndim != 4:
        raise ValueError("Error when checking input: expected 4 dimensions (batch, height, width, channels), got array "
                         "with shape %s" % (a.shape,))
    return np.ascontiguousarray(a, dtype=np.float32)


def to_device(a, device):
    """Host array -> device float32 through pinned memory."""
    t = torch.from_numpy(a)
    if torch.cuda.is_available():
        t = t.pin_memory()
    return t.to(device, non_blocking=True)
