"""Model.compile / Model.train_on_batch for the models the reference's compile helpers build
(upscaling/upscaler/model.py:1017-1137), executed as hand-written sm_100a kernels through the C ABI.

Recognised structures (everything else raises at compile time):
  'gen'    outputs G(x),               loss = content loss                        compile_training_model, gen_train
  'disc'   outputs D(x),               loss = wasserstein_loss                    make_and_compile_gan  disc_train
  'gan'    outputs [G(x), D(G(x))],    loss = [content, wasserstein_loss]         make_and_compile_gan  gan_train
  'disc2'  inputs [real, fake], out D(fake), loss = GanLosses.discriminator_loss  make_and_compile_gan2 disc_train
  'gan2'   inputs [x, real], outputs [G(x), D(G(x))], loss = [content, GanLosses.generator_loss]   ...  gan_train

Keras semantics kept: losses are returned as computed in the forward pass (before the update); a model's `trainable`
flag is snapshotted at compile time for the WEIGHTS (model.py:1034, 1040: `_collected_trainable_weights`); BatchNormalization
of a frozen discriminator still normalises with batch statistics inside a training step (Keras 2.2.x keeps
learning_phase = 1 for every layer of a train function); one Adam instance shared between compiled models shares
`iterations` while each compiled model owns its moment slots.

Moving statistics of a frozen model — UNPINNED third-party behaviour, exposed as a switch.  In Keras >= 2.1.3
`Layer.updates` / `Network.updates` return [] while the layer's or nested model's `trainable` is False
(keras/engine/base_layer.py `updates`, keras/engine/network.py `updates`: "if not self.trainable and not self.stateful:
return []"), and a compiled model collects its updates when it builds the train function, which happens lazily at the
FIRST train_on_batch (keras/engine/training.py `_make_train_function`).  The reference's helpers leave
`discriminator.trainable = False` behind (model.py:1040, 1101) and its scripts never set it back, so under that reading
the discriminator's moving mean / variance never move in either compiled model — the saved model_discriminator_*.h5
and D.predict differ, losses do not.  `FROZEN_MODEL_UPDATES_MOVING_STATS = False` (default) follows that reading: the
flag is sampled from the leaf model's `trainable` at the first train_on_batch of each compiled model.  Set it to True
for the pre-2.1.3 behaviour (moving statistics follow every training-phase pass).  No Keras is installed here to pin
either.

Data parallelism: when torch.distributed is initialised, every rank runs the same step on its shard of the batch;
BatchNorm statistics (forward sums and backward sums) are all-reduced so that N replicas normalise like one big batch,
loss gradients are scaled by 1/world at the source and the flat gradient buffer is sum-all-reduced over NCCL in
buckets, launched on a side stream as soon as backward has finished the slice (the flat layout follows the graph
order, so the tail of the network — whose gradients are ready first — is the last slice).
"""
import math

import numpy as np
import torch

from . import ops
from .keras_compat import as_batch


# ------------------------------------------------------------------------------------------------ loss descriptors
class ContentLoss:
    """What VGG_LOSS / VGG_MSE_LOSS / VGG_MAE_LOSS(...).loss hands to compile (model.py:101-157)."""

    def __init__(self, kind, rate, vgg_net):
        self.kind, self.rate, self.vgg = kind, float(rate), vgg_net

    @property
    def mae(self):
        return self.kind == "vgg-mae"


class GanLossFn:
    """The closures GanLosses.discriminator_loss / generator_loss return (model.py:215-261): they ignore
    (y_true, y_pred) and read the symbolic discriminator outputs stored on the GanLosses object."""

    def __init__(self, owner, role):
        self.owner, self.role = owner, role

    def __call__(self, y_true, y_pred):
        raise TypeError("GAN loss closures are evaluated inside train_on_batch")


def loss_activation(name, x):
    """(value, derivative) of GanLosses' loss activations at the scalar x (model.py:170-181).  Saturating forms: the
    un-clipped critic reaches |D| ~ 1e11 (gan_test.ipynb cell 17), where K.sigmoid / K.log saturate instead of
    overflowing: sigmoid -> 0 / 1, log(sigmoid) -> x (Keras' two-step log(sigmoid(x)) gives -inf there; the value is
    only reported, the derivative 1 - sigmoid(x) is what drives the update and is identical)."""
    if name in ("sigmoid", "log-sigm"):
        e = math.exp(-abs(x))                       # in (0, 1]: never overflows
        s = 1.0 / (1.0 + e) if x >= 0 else e / (1.0 + e)
        if name == "sigmoid":
            return s, s * (1 - s)
        return min(x, 0.0) - math.log1p(e), 1 - s
    if name == "tanh":
        t = math.tanh(x)
        return t, 1 - t * t
    if name == "bi-log":
        a = abs(x)
        f = x / (1 + a)
        return f * math.log(a + 2), math.log(a + 2) / (1 + a) ** 2 + f * (1.0 if x >= 0 else -1.0) / (a + 2)
    return x, 1.0


FROZEN_MODEL_UPDATES_MOVING_STATS = False      # see the module docstring

# CUDA-graph replay of the generator-only training step.  A step is ~360 launches of which two thirds last a few microseconds;
# issued from Python the stream runs dry whenever the host falls behind (measured on 8 GPUs sharing 32 host cores: the ranks
# wait for each other's hosts at every one of the 66 cross-replica BatchNorm exchanges, 17.5 -> 22.3 ms per step,
# profiles/r02_s9_peer_probe_8gpu.log).  After GRAPH_WARMUP eager calls with the same shapes the step is captured as two
# graphs -- [weight repack + forward] and [VGG19 loss + backward] -- split where the target batch is needed, so its host
# staging and DMA still overlap the forward pass; the gradient all-reduce (one NCCL call), Adam and the loss read stay
# eager.  Captured launches bake their arguments: the cross-replica exchanges take their epoch from a device word
# (peer.PeerReducer.before_replay).  Any failure to capture falls back to the eager path for good.
GRAPH_STEPS = True
GRAPH_WARMUP = 3
GRAPH_MAX_SHAPES = 4          # captured steps kept per compiled model (each owns its activations: GBs at full-frame sizes)

# Host-side accounting of the steps (bench.py, tools/): seconds the host spent blocked in the step's one device -> host
# read (the loss).  A step whose wait is ~0 is bound by the host's launch rate, not by the GPU.
HOST_STATS = {"wait_s": 0.0, "reads": 0}

# Gradient scaling of the training step.  Every loss gradient is multiplied by S at its source, the whole backward pass
# (linear in that gradient) carries S x the true values, and Adam divides S out again (b2_adam's gscale).  With bf16
# storage S = 1.  With fp16 storage S lifts the gradients out of the subnormal range: "numel" = the number of elements of
# the generated batch (cancels the 1/n of the mean-type content losses: per-pixel gradients become O(1)); for a
# discriminator step S = D_SCALE x batch size (the score gradients +-1/n become +-D_SCALE).
GRAD_SCALE = {"G": "numel", "D": 256.0}


# ------------------------------------------------------------------------------------------------ distributed helpers
def _dist():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        return dist
    return None


def agree(flag, device=None, mode="any"):
    """One decision for all data-parallel replicas: True when `flag` holds on ANY rank (mode "any") or on ALL ranks ("all").
    Decisions that change which collectives a step issues -- leaving fp16 storage after a range hit (a new graph key, eager
    warm-up steps with bucketed all-reduces), giving up on CUDA-graph capture -- must fall the same way everywhere, or one
    rank waits in a collective its peers never post.  Every rank has to call this at the same point of the step sequence;
    a single process returns `flag` unchanged."""
    dist = _dist()
    if dist is None:
        return bool(flag)
    on_dev = device is not None and torch.device(device).type == "cuda" and dist.get_backend() == "nccl"
    t = torch.tensor([1.0 if flag else 0.0], dtype=torch.float32, device=device if on_dev else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX if mode == "any" else dist.ReduceOp.MIN)
    return bool(t.item() != 0.0)


class BnSync:
    """Cross-replica exchange of the BatchNorm sums.  Where the replicas can map each other's memory the exchange is
    folded into the BatchNorm finalize kernels (csrc/b2_bn_train.cu over csrc/b2_peer.cuh): `peer_for(device)` returns
    the PeerReducer whose regions / epoch those launches take.  Otherwise `all_reduce(buf)` runs on the process group
    (NCCL, or gloo in the CPU tests).  Calling the object keeps the older protocol: sync(buf, rows) all-reduces buf in
    place and returns the global row count (or the world size when rows is None)."""

    def __init__(self, dist):
        self.dist, self.world = dist, dist.get_world_size()
        self._peer, self._asked = None, False

    def peer_for(self, device):
        if not self._asked:
            self._asked = True
            if torch.device(device).type == "cuda":
                from . import peer
                self._peer = peer.reducer(self.dist, device)
        return self._peer

    def all_reduce(self, buf):
        r = self.peer_for(buf.device) if buf.is_cuda else None
        if r is not None and buf.dtype == torch.float32 and buf.is_contiguous() and buf.numel() <= r.max_elems:
            r.all_reduce(buf)
        else:
            self.dist.all_reduce(buf)
        return buf

    def __call__(self, buf, rows):
        self.all_reduce(buf)
        return rows * self.world if rows is not None else self.world


_sync_state = {}

# Data-parallel BatchNorm semantics.  True (default): the replicas exchange their BatchNorm sums, so N replicas of B/N samples
# normalise exactly like the reference's single process does over the batch of B (parity; 66 exchanges per generator step).
# False: every replica normalises over its own shard (what torch's DistributedDataParallel does without SyncBatchNorm) --
# no exchange on the critical path, results differ from the single-process step.
SYNC_BATCHNORM = True


def _make_sync():
    """The process-wide BnSync when torch.distributed runs with more than one rank, else None."""
    dist = _dist()
    if dist is None or not SYNC_BATCHNORM:
        _sync_state.clear()
        return None
    if _sync_state.get("dist") is not dist or _sync_state.get("world") != dist.get_world_size():
        _sync_state.update(dist=dist, world=dist.get_world_size(), sync=BnSync(dist))
    return _sync_state["sync"]


class GradReducer:
    """Bucketed sum-all-reduce of a flat gradient buffer, overlapped with the rest of backward: the networks' backward
    passes call `hook(offset)` when every gradient at flat offsets >= offset is final; the slice [offset, previous
    offset) is then all-reduced on a side stream (ordered after the kernels already enqueued on the compute stream)
    while backward keeps running.  Adam waits for the buckets' events.  Slices smaller than `min_elems` are merged
    into the next bucket (launch latency, not link count, sizes a bucket on NVSwitch)."""

    def __init__(self, min_elems=1 << 17):
        self.min_elems = min_elems
        self.stream = None
        self.pending = []
        self.log = []          # (lo, hi) of the buckets of the last step (tests)

    def hook(self, flat, n):
        """Returns on_ready(offset) for one backward pass over `flat[:n]`."""
        state = {"hi": n}
        self.log = []

        def on_ready(offset):
            lo, hi = int(offset), state["hi"]
            if hi - lo < self.min_elems and lo > 0:
                return
            if hi > lo:
                self.reduce_slice(flat, lo, hi)
                state["hi"] = lo
        return on_ready

    def reduce_slice(self, flat, lo, hi):
        self.log.append((lo, hi))
        dist = _dist()
        if dist is None or hi <= lo:
            return
        if flat.is_cuda:
            if self.stream is None:
                self.stream = torch.cuda.Stream()
            ev = torch.cuda.Event()
            ev.record()
            self.stream.wait_event(ev)
            with torch.cuda.stream(self.stream):
                dist.all_reduce(flat[lo:hi])
                done = torch.cuda.Event()
                done.record()
            self.pending.append(done)
        else:       # gloo (CPU tests of the host logic): synchronous
            dist.all_reduce(flat[lo:hi])

    def wait(self):
        for ev in self.pending:
            torch.cuda.current_stream().wait_event(ev)
        self.pending = []


# ------------------------------------------------------------------------------------------------ compiled models
class _Uploader:
    """Host -> device staging of train_on_batch's NumPy arguments.  Persistent pinned buffers (keyed by slot and shape) are
    filled by a small thread pool in row chunks -- NumPy's memcpy releases the GIL -- because torch's own pageable -> pinned
    copy runs on the intra-op thread pool, which `torchrun` pins to ONE thread (OMP_NUM_THREADS=1): measured at 2 GPUs, the
    30 MB batch then took 4.8 ms per step instead of 1.1 ms, most of the data-parallel overhead.  `begin()` starts
    staging in the background so the caller can launch kernels meanwhile; `finish()` issues the DMA on a copy stream and
    makes the compute stream wait for it."""

    THREADS = 4

    def __init__(self):
        self.pinned, self.events, self.stream, self.pool = {}, {}, None, None

    def begin(self, a, slot=0):
        a = np.asarray(a)       # any numeric dtype: the reference's convert_image_series_to_array hands over float64
        if not a.flags.c_contiguous:
            a = np.ascontiguousarray(a)
        key = (slot, tuple(a.shape))
        buf = self.pinned.get(key)
        if buf is None:
            buf = self.pinned[key] = torch.empty(tuple(a.shape), dtype=torch.float32).pin_memory()
        ev = self.events.get(key)
        if ev is not None:
            ev.synchronize()                 # the previous DMA out of this buffer has finished
        dst = buf.numpy().reshape(-1)
        src = a.reshape(-1)
        n = src.size
        if n < (1 << 20):
            np.copyto(dst, src, casting="unsafe")
            return key, buf, []
        if self.pool is None:
            from concurrent.futures import ThreadPoolExecutor
            self.pool = ThreadPoolExecutor(self.THREADS)
        # Twice as many chunks as threads: finish_into() sends a chunk's DMA off as soon as it is staged, so the first half of
        # the batch crosses PCIe while the second half is still being copied.  The cast to float32 (float64 batches are
        # twice the bytes) happens inside the chunked, GIL-free copies.
        step = -(-n // (2 * self.THREADS))
        step += (-step) % 1024
        self.ranges = getattr(self, "ranges", {})
        futs = [self.pool.submit(np.copyto, dst[i:i + step], src[i:i + step], "unsafe") for i in range(0, n, step)]
        self.ranges[key] = [(i, min(i + step, n)) for i in range(0, n, step)]
        return key, buf, futs

    def finish(self, handle, dev, side_stream=False):
        key, buf, futs = handle
        for f in futs:
            f.result()
        if side_stream:
            if self.stream is None:
                self.stream = torch.cuda.Stream(device=dev)
            with torch.cuda.stream(self.stream):
                t = buf.to(dev, non_blocking=True)
                ev = self.events[key] = torch.cuda.Event()
                ev.record()
            t.record_stream(torch.cuda.current_stream())
            torch.cuda.current_stream().wait_event(ev)
            return t
        t = buf.to(dev, non_blocking=True)
        ev = self.events[key] = torch.cuda.Event()
        ev.record()
        return t


def _finish_into(self, handle, dst, side_stream=False, after=None):
    """Like finish(), but the DMA lands in the caller's persistent device tensor `dst` (CUDA-graph replay reads fixed
    addresses).  side_stream: the copy runs on the copy stream once event `after` (recorded on the compute stream) has
    passed, and the compute stream waits for it."""
    key, buf, futs = handle
    import time
    HOST_STATS.setdefault("stage_wait_s", 0.0)
    if not side_stream:
        t0 = time.perf_counter()
        for f in futs:
            f.result()
        HOST_STATS["stage_wait_s"] += time.perf_counter() - t0
    if side_stream:
        if self.stream is None:
            self.stream = torch.cuda.Stream(device=dst.device)
        if after is not None:
            self.stream.wait_event(after)
        with torch.cuda.stream(self.stream):
            if futs:                 # chunk by chunk, in staging order
                flat_d, flat_s = dst.view(-1), buf.view(-1)
                for f, (lo, hi) in zip(futs, self.ranges[key]):
                    t1 = time.perf_counter()
                    f.result()
                    HOST_STATS["stage_wait_s"] += time.perf_counter() - t1
                    flat_d[lo:hi].copy_(flat_s[lo:hi], non_blocking=True)
            else:
                dst.copy_(buf, non_blocking=True)
            ev = self.events[key] = torch.cuda.Event()
            ev.record()
        torch.cuda.current_stream().wait_event(ev)
        return dst
    dst.copy_(buf, non_blocking=True)
    ev = self.events[key] = torch.cuda.Event()
    ev.record()
    return dst


_Uploader.finish_into = _finish_into
_uploader = _Uploader()


def _to_dev(a, dev, slot="a"):
    """Synchronous form: stage + DMA on the current stream.  Every argument of a step names its own slot: arguments of the
    same shape must not share a staging buffer within a step, and a fixed name per call site keeps the set of pinned
    buffers fixed (a rotating slot would keep allocating 40 MB pinned blocks for the first steps, ~5 ms each)."""
    return _uploader.finish(_uploader.begin(a, slot), dev)


class Compiled:
    def __init__(self, kind, model, optimizer):
        self.kind, self.model, self.opt = kind, model, optimizer
        self.G = self.D = None
        self.G_model = self.D_model = None     # the leaf Model objects (their `trainable` flags; see _moves)
        self.content = None
        self.g_frozen = self.d_frozen = False
        self.weights = (1.0, 1.0)
        self.slots = {}
        self.gan_losses = None
        self.reducer = GradReducer()
        self._ticks = {}
        self._warm, self._graphs, self._graph_off = {}, {}, False
        self._guarded, self._undone = set(), False

    def _moves(self, which):
        """Whether this compiled model's training-phase passes update the moving statistics of G / D: sampled once, at
        the first train_on_batch, from the leaf model's `trainable` (module docstring)."""
        key = "_moves_" + which
        if key not in self.__dict__:
            m = self.G_model if which == "G" else self.D_model
            self.__dict__[key] = bool(FROZEN_MODEL_UPDATES_MOVING_STATS or m is None or m.trainable)
        return self.__dict__[key]

    # -------------------------------------------------------------- fp16 storage range guard (networks.RANGE_LIMIT)
    def _range_tick(self, net, tape=None):
        """Every step, after the update was launched: read the network's guard words (the BatchNorm range flag and the count
        of saturated activation gradients; the stream is about to be drained for the loss anyway).  If they were raised the
        guarded Adam launch of this step did nothing: the optimizer's step count is taken back and the network continues
        in bf16 storage (or the step raises, networks.ON_RANGE).  First step and every RANGE_CHECK_EVERY-th additionally
        sample the stored forward tensors no BatchNorm watches (head conv, Conv2DTranspose outputs)."""
        import time
        from . import networks
        if net.train_dtype != torch.float16:
            return
        n = self._ticks[id(net)] = self._ticks.get(id(net), 0) + 1
        extra = []
        if (n == 1 or n % networks.RANGE_CHECK_EVERY == 0) and tape is not None and "ups" in tape:
            extra = [tape["head"][1]] + list(tape["ups"][1:])
        t0 = time.perf_counter()
        try:
            net.check_range(extra, agree=lambda hit: agree(hit, net.device, "any"))
        finally:
            HOST_STATS["wait_s"] += time.perf_counter() - t0
            if net.guard_hit and id(net) in self._guarded and not self._undone:
                self.opt.iterations -= 1        # the update was dropped on the device: Adam's t does not advance
                self._undone = True

    # -------------------------------------------------------------- Adam
    @staticmethod
    def _scale_for(net, numel, n):
        """S for a step whose backward starts in `net` ('numel' elements generated / `n` scores)."""
        if net.train_dtype != torch.float16:
            return 1.0
        from .networks import UpscalerOrigNet
        if isinstance(net, UpscalerOrigNet):
            g = GRAD_SCALE["G"]
            return float(numel) if g == "numel" else float(g)
        return float(GRAD_SCALE["D"]) * float(n)

    @staticmethod
    def _inner_scale(D, max_abs_ds):
        """Extra factor for the score gradient of a (frozen) fp16 discriminator inside a generator step, so that it enters
        D's backward at magnitude GRAD_SCALE['D'] whatever the adversarial loss weight is; D.backward divides it out."""
        if D.train_dtype != torch.float16 or not (max_abs_ds > 0.0) or not math.isfinite(max_abs_ds):
            return 1.0
        return float(2.0 ** round(math.log2(float(GRAD_SCALE["D"]) / max_abs_ds)))

    def _adam(self, net, S=1.0, also=None):
        """Adam update of `net` from its flat gradient.  The launch is guarded by the guard words of `net` and of `also` (the
        frozen discriminator a generator step differentiates through): if the fp16 passes that produced the gradient raised
        one, the update is dropped on the device (see _range_tick).  Replicas max-reduce the words first, so all drop or none."""
        opt = self.opt
        opt.iterations += 1
        t = opt.iterations
        lr = opt.lr * (1.0 / (1.0 + opt.decay * (t - 1))) if opt.decay else opt.lr
        lr_t = lr * math.sqrt(1.0 - opt.beta_2 ** t) / (1.0 - opt.beta_1 ** t)
        key = id(net)
        if key not in self.slots:
            self.slots[key] = (torch.zeros_like(net.grad_flat), torch.zeros_like(net.grad_flat))
        m, v = self.slots[key]
        self.reducer.wait()
        guards = [g for g in (n_.guard_words() for n_ in (net, also) if n_ is not None and n_.on_gpu) if g is not None]
        self._guarded = {id(n_) for n_ in (net, also) if n_ is not None}
        self._undone = False
        dist = _dist()
        if dist is not None:
            for g in guards:
                dist.all_reduce(g, op=dist.ReduceOp.MAX)
            for n_ in (net, also):
                if n_ is not None:
                    n_._guard_reduced = True
        ops.adam(net.flat[:net.n_trainable], net.grad_flat, m, v, lr_t, opt.beta_1, opt.beta_2, opt.epsilon, gscale=1.0 / S, guards=guards)
        net._train_dirty = True
        net._infer_dirty = True

    @staticmethod
    def _fresh(net, sets=("train",)):
        """Packed 16-bit weights are re-derived lazily after an update of the fp32 masters."""
        for s in sets:
            flag = "_%s_dirty" % s
            if getattr(net, flag, True) or s not in net._packed:
                net.refresh((s,))
                setattr(net, flag, False)

    # -------------------------------------------------------------- content loss (+ gradient w.r.t. the generated image)
    def _content(self, y, y_true, loss_buf, weight, S=1.0, pre=None):
        """loss_buf += content loss (unweighted); returns the gradient tensors (fp32 images) scaled by S * `weight` / world.
        `pre` = (features, activations) of vgg.forward(y, save=True) when the caller already ran it (the graph-replayed steps
        run the VGG19 pass over the GENERATED batch before the target batch has arrived on the device)."""
        cl = self.content
        dist = _dist()
        gs = weight / (dist.get_world_size() if dist else 1)
        vgg = cl.vgg
        self._fresh(vgg)
        ft = vgg.forward(y_true)
        fp, acts = pre if pre is not None else vgg.forward(y, save=True)
        dfeat = torch.empty_like(fp)
        Sv = float(fp.numel())      # carry the feature gradient without the 1/n of the mean (fp16 range), see Vgg19Net
        ops.loss_pair(fp, ft, loss_buf, mae=cl.mae, weight=1.0, gscale=gs * Sv, grad=dfeat)
        grads = [vgg.backward(acts, dfeat, Sv / S)]   # the last data-gradient conv multiplies by S / Sv in fp32
        if cl.kind in ("vgg-mse", "vgg-mae"):
            gpix = torch.empty_like(y)
            ops.loss_pair(y, y_true, loss_buf, mae=cl.mae, weight=cl.rate, gscale=gs * cl.rate * S, grad=gpix)
            grads.append(gpix)
        return grads

    # -------------------------------------------------------------- steps
    # -------------------------------------------------------------- CUDA-graph replay of step_gen (see GRAPH_STEPS)
    def _graph_lookup(self, key, device, capture):
        """The captured step stored under `key`, capturing it (capture(sync, peer) -> state) when the key has been seen
        GRAPH_WARMUP times; None = run the eager path."""
        if not GRAPH_STEPS or self._graph_off or device.type != "cuda":
            return None
        st = self._graphs.get(key)
        if st is not None:
            return st
        n = self._warm[key] = self._warm.get(key, 0) + 1
        if n <= GRAPH_WARMUP or len(self._graphs) >= GRAPH_MAX_SHAPES:
            return None
        sync = _make_sync()
        peer = sync.peer_for(device) if sync is not None else None
        if sync is not None and peer is None:
            self._graph_off = True          # BatchNorm sums over NCCL: collectives stay out of captured graphs
            return None
        st = err = None
        # Python's cyclic collector must not run inside a capture: if it finalises an older CUDAGraph (a compiled model of
        # an earlier leg that was dropped), the cudaGraphExecDestroy lands in the capturing stream's window and invalidates
        # the capture (seen on one rank of two, profiles/r03_s4_bench_n2.json.log).  Collect first, then hold it off.
        import gc
        gc.collect()
        gc_was_on = gc.isenabled()
        gc.disable()
        try:
            st = capture(sync, peer)
        except Exception as e:  # noqa: BLE001
            err = e
            if peer is not None:
                peer.end_capture()
        finally:
            if gc_was_on:
                gc.enable()
        # replayed and eager steps issue different collectives: all replicas replay, or none does (every rank reaches this
        # point on the same step: the key and its warm-up count are functions of shapes and storage types only)
        if not agree(st is not None, device, "all"):
            self._graph_off = True
            import warnings
            why = "%s: %s" % (type(err).__name__, err) if err is not None else "another replica could not capture it"
            warnings.warn("CUDA-graph capture of the training step failed (%s); staying on the eager path" % why, RuntimeWarning)
            return None
        dts = tuple(n_.train_dtype for n_ in (self.G, self.D) if n_ is not None)
        for k in [k for k in self._graphs if k[-1] != dts]:
            del self._graphs[k]             # graphs of a storage type a network has left (range guard)
        self._graphs[key] = st
        return st

    def _dtypes(self):
        return tuple(n_.train_dtype for n_ in (self.G, self.D) if n_ is not None)

    @staticmethod
    def _world():
        dist = _dist()
        return dist.get_world_size() if dist else 1

    def _graph_state(self, x, y_true):
        G = self.G
        if not G.on_gpu:
            return None
        key = ("gen", tuple(np.shape(x)), tuple(np.shape(y_true)), self.content.vgg.train_dtype, self._world(), self._dtypes())
        return self._graph_lookup(key, G.device, lambda sync, peer: self._capture_gen(key, x, y_true, sync, peer))

    @staticmethod
    def _replay(st, graph, n_exchanges):
        if st.peer is not None:
            st.peer.before_replay()
        graph.replay()
        if st.peer is not None:
            st.peer.after_replay(n_exchanges)

    def _capture_gen(self, key, x, y_true, sync, peer):
        from types import SimpleNamespace
        G = self.G
        dev = G.device
        st = SimpleNamespace(peer=peer, n1=0, n2=0)
        st.x = torch.empty(tuple(np.shape(x)), dtype=torch.float32, device=dev)
        st.y = torch.empty(tuple(np.shape(y_true)), dtype=torch.float32, device=dev)
        G.ensure_grads()
        G._train_dirty = True               # the captured forward always starts with the repack of the fp32 masters
        pool = torch.cuda.graph_pool_handle()
        torch.cuda.synchronize()
        n0 = ops.launch_count()
        st.g1, st.g2 = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        if peer is not None:
            peer.begin_capture()
        with torch.cuda.graph(st.g1, pool=pool, capture_error_mode="thread_local"):
            self._fresh(G)
            y, tape = G.forward_train(st.x, sync, update_moving=self._moves('G'))
            self._fresh(self.content.vgg)
            pre = self.content.vgg.forward(y, save=True)     # needs no target: more cover for the target batch's staging + DMA
        if peer is not None:
            st.n1 = peer.end_capture()
            peer.begin_capture()
        with torch.cuda.graph(st.g2, pool=pool, capture_error_mode="thread_local"):
            loss = torch.zeros(1, device=dev)
            S = self._scale_for(G, y.numel(), y.shape[0])
            grads = self._content(y, st.y, loss, 1.0, S, pre=pre)
            dz = ops.tanh_bwd(y, grads[0], torch.empty_like(y), g2=grads[1] if len(grads) > 1 else None)
            G.grad_flat.zero_()
            G.backward(tape, dz, sync, on_ready=None)
        if peer is not None:
            st.n2 = peer.end_capture()
        st.launches = ops.launch_count() - n0
        ops._count(-st.launches)            # nothing ran yet: replays are what counts
        st.loss, st.S = loss, S
        st.tape = {"head": tape["head"], "ups": tape["ups"]}       # tensors the range guard samples (graph-owned memory)
        return st

    def _step_gen_graph(self, st, x, y_true):
        G = self.G
        peer = st.peer
        if peer is not None and not peer.can_replay(st.n1 + st.n2):
            return None
        pending = _uploader.begin(y_true, "y")          # the long pole first: staged by the pool threads while graph 1 runs
        hx = _uploader.begin(x, "x")
        _uploader.finish_into(hx, st.x)
        started = torch.cuda.Event()
        started.record()
        if peer is not None:
            peer.before_replay()
        st.g1.replay()
        if peer is not None:
            peer.after_replay(st.n1)
        G._train_dirty = False
        _uploader.finish_into(pending, st.y, side_stream=True, after=started)
        if peer is not None:
            peer.before_replay()
        st.g2.replay()
        if peer is not None:
            peer.after_replay(st.n2)
        ops._count(st.launches)
        dist = _dist()
        if dist is not None:
            dist.all_reduce(G.grad_flat)                # loss gradients carry 1/world: the sum is the global-batch gradient
        self._adam(G, st.S)
        self._range_tick(G, st.tape)
        return self._mean_over_ranks(st.loss)[0]

    def step_gen(self, x, y_true):
        G = self.G
        dev = G.device
        st = self._graph_state(x, y_true)
        if st is not None:
            r = self._step_gen_graph(st, x, y_true)
            if r is not None:
                return r
        self._fresh(G)
        xd = _to_dev(x, dev, "x")
        pending = _uploader.begin(y_true, "y")        # the 16x larger target batch is staged while the forward pass is launched
        sync = _make_sync()
        y, tape = G.forward_train(xd, sync, update_moving=self._moves('G'))
        yt = _uploader.finish(pending, dev, side_stream=True)
        loss = torch.zeros(1, device=dev)
        S = self._scale_for(G, y.numel(), y.shape[0])
        grads = self._content(y, yt, loss, 1.0, S)
        dz = ops.tanh_bwd(y, grads[0], torch.empty_like(y), g2=grads[1] if len(grads) > 1 else None)
        G.ensure_grads().zero_()
        G.backward(tape, dz, sync, on_ready=self.reducer.hook(G.grad_flat, G.n_trainable))
        self._adam(G, S)
        self._range_tick(G, tape)
        return self._mean_over_ranks(loss)[0]

    def _capture_disc(self, x, nlab, sync, peer):
        from types import SimpleNamespace
        D = self.D
        dev = D.device
        st = SimpleNamespace(peer=peer, n1=0)
        st.x = torch.empty(tuple(np.shape(x)), dtype=torch.float32, device=dev)
        st.lab = torch.empty((nlab,), dtype=torch.float32, device=dev)
        D.ensure_grads()
        D._train_dirty = True
        torch.cuda.synchronize()
        n0 = ops.launch_count()
        st.g1 = torch.cuda.CUDAGraph()
        if peer is not None:
            peer.begin_capture()
        with torch.cuda.graph(st.g1, capture_error_mode="thread_local"):
            self._fresh(D)
            scores, tape = D.forward(st.x, True, sync, update_moving=self._moves('D'))
            loss = torch.zeros(1, device=dev)
            ds = torch.empty_like(scores)
            S = self._scale_for(D, 0, scores.shape[0])
            ops.mean_prod(scores, st.lab, loss, 1.0, S / self._world(), ds)
            D.grad_flat.zero_()
            D.backward(tape, ds, True, False, sync, on_ready=None)
        if peer is not None:
            st.n1 = peer.end_capture()
        st.launches = ops.launch_count() - n0
        ops._count(-st.launches)
        st.loss, st.S = loss, S
        return st

    def _step_disc_graph(self, st, x, y):
        D = self.D
        if st.peer is not None and not st.peer.can_replay(st.n1):
            return None
        _uploader.finish_into(_uploader.begin(x, "x"), st.x)
        _uploader.finish_into(_uploader.begin(np.asarray(y, np.float32).reshape(-1), "labels"), st.lab)
        self._replay(st, st.g1, st.n1)
        D._train_dirty = False
        ops._count(st.launches)
        dist = _dist()
        if dist is not None:
            dist.all_reduce(D.grad_flat)
        self._adam(D, st.S)
        self._range_tick(D)
        return self._mean_over_ranks(st.loss)[0]

    def step_disc(self, x, y):
        D = self.D
        dev = D.device
        if D.on_gpu:
            key = ("disc", tuple(np.shape(x)), int(np.size(y)), self._world(), self._dtypes())
            st = self._graph_lookup(key, dev, lambda sync, peer: self._capture_disc(x, int(np.size(y)), sync, peer))
            if st is not None:
                r = self._step_disc_graph(st, x, y)
                if r is not None:
                    return r
        self._fresh(D)
        xd = _to_dev(x, dev, "x")
        yd = _to_dev(np.asarray(y, np.float32).reshape(-1), dev, "labels")
        sync = _make_sync()
        scores, tape = D.forward(xd, True, sync, update_moving=self._moves('D'))
        loss = torch.zeros(1, device=dev)
        ds = torch.empty_like(scores)
        dist = _dist()
        S = self._scale_for(D, 0, scores.shape[0])
        ops.mean_prod(scores, yd, loss, 1.0, S / (dist.get_world_size() if dist else 1), ds)
        D.ensure_grads().zero_()
        D.backward(tape, ds, True, False, sync, on_ready=self.reducer.hook(D.grad_flat, D.n_trainable))
        self._adam(D, S)
        self._range_tick(D)
        return self._mean_over_ranks(loss)[0]

    def _capture_gan(self, x, y_true, nadv, amax, sync, peer):
        from types import SimpleNamespace
        G, D = self.G, self.D
        dev = G.device
        cw, dw = self.weights
        world = self._world()
        st = SimpleNamespace(peer=peer, n1=0, n2=0)
        st.x = torch.empty(tuple(np.shape(x)), dtype=torch.float32, device=dev)
        st.y = torch.empty(tuple(np.shape(y_true)), dtype=torch.float32, device=dev)
        st.adv = torch.empty((nadv,), dtype=torch.float32, device=dev)
        G.ensure_grads()
        if not self.d_frozen:
            D.ensure_grads()
        G._train_dirty = D._train_dirty = True      # both repacks are part of the captured forward
        pool = torch.cuda.graph_pool_handle()
        torch.cuda.synchronize()
        n0 = ops.launch_count()
        st.g1, st.g2 = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        if peer is not None:
            peer.begin_capture()
        with torch.cuda.graph(st.g1, pool=pool, capture_error_mode="thread_local"):
            self._fresh(G)
            self._fresh(D)
            y, gtape = G.forward_train(st.x, sync, update_moving=self._moves('G'))
            scores, dtape = D.forward(y, True, sync, update_moving=self._moves('D'))
            self._fresh(self.content.vgg)
            pre = self.content.vgg.forward(y, save=True)
        if peer is not None:
            st.n1 = peer.end_capture()
            peer.begin_capture()
        with torch.cuda.graph(st.g2, pool=pool, capture_error_mode="thread_local"):
            losses = torch.zeros(2, device=dev)
            S = self._scale_for(G, y.numel(), y.shape[0])
            grads = self._content(y, st.y, losses[0:1], cw, S, pre=pre)
            ds = torch.empty_like(scores)
            kD = self._inner_scale(D, abs(dw * S / world) * amax / max(scores.shape[0], 1)) if self.d_frozen else 1.0
            ops.mean_prod(scores, st.adv, losses[1:2], 1.0, dw * S * kD / world, ds)
            grads.append(D.backward(dtape, ds, not self.d_frozen, True, sync, grad_scale=kD))
            dz = ops.tanh_bwd(y, grads[0], torch.empty_like(y), g2=grads[1], g3=grads[2] if len(grads) > 2 else None)
            G.grad_flat.zero_()
            G.backward(gtape, dz, sync, on_ready=None)
        if peer is not None:
            st.n2 = peer.end_capture()
        st.launches = ops.launch_count() - n0
        ops._count(-st.launches)
        st.losses, st.S = losses, S
        st.tape = {"head": gtape["head"], "ups": gtape["ups"]}
        return st

    def _step_gan_graph(self, st, x, y_true, adv_y):
        G, D = self.G, self.D
        cw, dw = self.weights
        if st.peer is not None and not st.peer.can_replay(st.n1 + st.n2):
            return None
        pending = _uploader.begin(y_true, "y")
        hx = _uploader.begin(x, "x")
        _uploader.finish_into(hx, st.x)
        _uploader.finish_into(_uploader.begin(np.asarray(adv_y, np.float32).reshape(-1), "adv"), st.adv)
        started = torch.cuda.Event()
        started.record()
        self._replay(st, st.g1, st.n1)
        G._train_dirty = D._train_dirty = False
        _uploader.finish_into(pending, st.y, side_stream=True, after=started)
        self._replay(st, st.g2, st.n2)
        ops._count(st.launches)
        dist = _dist()
        if dist is not None:
            dist.all_reduce(G.grad_flat)
        self._adam(G, st.S, also=D)
        self._range_tick(G, st.tape)
        self._range_tick(D)
        lc, la = self._mean_over_ranks(st.losses)
        return [cw * lc + dw * la, lc, la]

    def step_gan(self, x, y_true, adv_y):
        G, D = self.G, self.D
        dev = G.device
        if G.on_gpu:
            amax_k = float(np.abs(np.asarray(adv_y, np.float32)).max()) if np.size(adv_y) else 0.0
            key = ("gan", tuple(np.shape(x)), tuple(np.shape(y_true)), int(np.size(adv_y)), amax_k, self.content.vgg.train_dtype,
                   self._world(), self._dtypes())
            st = self._graph_lookup(key, dev, lambda sync, peer: self._capture_gan(x, y_true, int(np.size(adv_y)), amax_k, sync, peer))
            if st is not None:
                r = self._step_gan_graph(st, x, y_true, adv_y)
                if r is not None:
                    return r
        self._fresh(G)
        self._fresh(D)
        cw, dw = self.weights
        xd = _to_dev(x, dev, "x")
        ad = _to_dev(np.asarray(adv_y, np.float32).reshape(-1), dev, "adv")
        pending = _uploader.begin(y_true, "y")
        sync = _make_sync()
        dist = _dist()
        world = dist.get_world_size() if dist else 1
        y, gtape = G.forward_train(xd, sync, update_moving=self._moves('G'))
        yt = _uploader.finish(pending, dev, side_stream=True)
        scores, dtape = D.forward(y, True, sync, update_moving=self._moves('D'))
        losses = torch.zeros(2, device=dev)
        S = self._scale_for(G, y.numel(), y.shape[0])
        grads = self._content(y, yt, losses[0:1], cw, S)
        ds = torch.empty_like(scores)
        amax = float(np.abs(np.asarray(adv_y, np.float32)).max()) if np.size(adv_y) else 0.0
        kD = self._inner_scale(D, abs(dw * S / world) * amax / max(scores.shape[0], 1)) if self.d_frozen else 1.0
        ops.mean_prod(scores, ad, losses[1:2], 1.0, dw * S * kD / world, ds)
        grads.append(D.backward(dtape, ds, not self.d_frozen, True, sync, grad_scale=kD))
        dz = ops.tanh_bwd(y, grads[0], torch.empty_like(y), g2=grads[1], g3=grads[2] if len(grads) > 2 else None)
        G.ensure_grads().zero_()
        G.backward(gtape, dz, sync, on_ready=self.reducer.hook(G.grad_flat, G.n_trainable))
        self._adam(G, S, also=D)
        self._range_tick(G, gtape)
        self._range_tick(D)
        lc, la = self._mean_over_ranks(losses)
        return [cw * lc + dw * la, lc, la]

    def _relativistic(self, mean_real, mean_fake, role):
        gl = self.gan_losses
        if gl.family == "wasserstein":
            if role == "disc":
                return mean_real - mean_fake, 1.0, -1.0
            return mean_fake, 0.0, 1.0
        delta = (mean_real - mean_fake) if role == "disc" else (mean_fake - mean_real)
        val, der = loss_activation(gl.loss_activation_name, delta)
        return (val, der, -der) if role == "disc" else (val, -der, der)

    def _capture_disc2(self, real, fake, sync, peer):
        """Two graphs around the step's host read of the two critic means (the loss activation is evaluated on the host):
        [repack, both forward passes, means] -> host -> the two score gradients are FILLED into fixed tensors -> [both
        backward passes]."""
        from types import SimpleNamespace
        D = self.D
        dev = D.device
        st = SimpleNamespace(peer=peer, n1=0, n2=0)
        st.real = torch.empty(tuple(np.shape(real)), dtype=torch.float32, device=dev)
        st.fake = torch.empty(tuple(np.shape(fake)), dtype=torch.float32, device=dev)
        D.ensure_grads()
        D._train_dirty = True
        pool = torch.cuda.graph_pool_handle()
        torch.cuda.synchronize()
        n0 = ops.launch_count()
        st.g1, st.g2 = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        if peer is not None:
            peer.begin_capture()
        with torch.cuda.graph(st.g1, pool=pool, capture_error_mode="thread_local"):
            self._fresh(D)
            sr, tr = D.forward(st.real, True, sync, update_moving=self._moves('D'))
            sf, tf = D.forward(st.fake, True, sync, update_moving=self._moves('D'))
            st.means = torch.zeros(2, device=dev)
            ops.mean_prod(sr, None, st.means[0:1], 1.0, 0.0, None, 1.0)
            ops.mean_prod(sf, None, st.means[1:2], 1.0, 0.0, None, 1.0)
            st.ds_r, st.ds_f = torch.empty_like(sr), torch.empty_like(sf)
        if peer is not None:
            st.n1 = peer.end_capture()
            peer.begin_capture()
        with torch.cuda.graph(st.g2, pool=pool, capture_error_mode="thread_local"):
            D.grad_flat.zero_()
            D.backward(tr, st.ds_r, True, False, sync, on_ready=None)
            D.backward(tf, st.ds_f, True, False, sync, on_ready=None)
        if peer is not None:
            st.n2 = peer.end_capture()
        st.launches = ops.launch_count() - n0
        ops._count(-st.launches)
        st.S, st.n = self._scale_for(D, 0, sr.shape[0]), int(sr.shape[0])
        return st

    def _step_disc2_graph(self, st, real, fake):
        D = self.D
        if st.peer is not None and not st.peer.can_replay(st.n1 + st.n2):
            return None
        world = self._world()
        _uploader.finish_into(_uploader.begin(real, "real"), st.real)
        _uploader.finish_into(_uploader.begin(fake, "fake"), st.fake)
        self._replay(st, st.g1, st.n1)
        D._train_dirty = False
        mr, mf = self._mean_over_ranks(st.means)
        val, d_real, d_fake = self._relativistic(mr, mf, "disc")
        st.ds_r.fill_(d_real * st.S / world / st.n)          # = mean_prod's gradient of coef * mean(scores)
        st.ds_f.fill_(d_fake * st.S / world / st.n)
        self._replay(st, st.g2, st.n2)
        ops._count(st.launches)
        dist = _dist()
        if dist is not None:
            dist.all_reduce(D.grad_flat)
        self._adam(D, st.S)
        self._range_tick(D)
        return val

    def step_disc2(self, real, fake):
        """train_gan3.py:353 — loss = act(mean(D(real)) - mean(D(fake))); real and fake are separate BatchNorm batches."""
        D = self.D
        dev = D.device
        if D.on_gpu:
            key = ("disc2", tuple(np.shape(real)), tuple(np.shape(fake)), self._world(), self._dtypes())
            st = self._graph_lookup(key, dev, lambda sync, peer: self._capture_disc2(real, fake, sync, peer))
            if st is not None:
                r = self._step_disc2_graph(st, real, fake)
                if r is not None:
                    return r
        self._fresh(D)
        rd, fd = _to_dev(real, dev, "real"), _to_dev(fake, dev, "fake")
        sync = _make_sync()
        dist = _dist()
        world = dist.get_world_size() if dist else 1
        sr, tr = D.forward(rd, True, sync, update_moving=self._moves('D'))
        sf, tf = D.forward(fd, True, sync, update_moving=self._moves('D'))
        means = torch.zeros(2, device=dev)
        ops.mean_prod(sr, None, means[0:1], 1.0, 0.0, None, 1.0)
        ops.mean_prod(sf, None, means[1:2], 1.0, 0.0, None, 1.0)
        mr, mf = self._mean_over_ranks(means)
        val, d_real, d_fake = self._relativistic(mr, mf, "disc")
        D.ensure_grads().zero_()
        S = self._scale_for(D, 0, sr.shape[0])
        for scores, tape, coef in ((sr, tr, d_real), (sf, tf, d_fake)):
            ds = torch.empty_like(scores)
            ops.mean_prod(scores, None, None, 0.0, coef * S / world, ds, 1.0)
            # the two passes accumulate into the same buffer: only the second one's slices are final
            D.backward(tape, ds, True, False, sync,
                       on_ready=self.reducer.hook(D.grad_flat, D.n_trainable) if scores is sf else None)
        self._adam(D, S)
        self._range_tick(D)
        return val

    def _capture_gan2(self, x, real, y_true, sync, peer):
        """Three graphs: [repacks, D(real), G(x), D(G(x)), the two means, VGG19(G(x))] -> (target batch arrives) -> [content
        loss and its gradient] -> host read of the means and the content loss -> the score gradient is FILLED into a fixed
        tensor -> [critic backward, generator backward].  The internal critic scale kD is the one of the capturing step."""
        from types import SimpleNamespace
        G, D = self.G, self.D
        dev = G.device
        cw, dw = self.weights
        world = self._world()
        st = SimpleNamespace(peer=peer, n1=0, n2=0, n3=0)
        st.x = torch.empty(tuple(np.shape(x)), dtype=torch.float32, device=dev)
        st.real = torch.empty(tuple(np.shape(real)), dtype=torch.float32, device=dev)
        st.y = torch.empty(tuple(np.shape(y_true)), dtype=torch.float32, device=dev)
        G.ensure_grads()
        if not self.d_frozen:
            D.ensure_grads()
        G._train_dirty = D._train_dirty = True
        pool = torch.cuda.graph_pool_handle()
        torch.cuda.synchronize()
        n0 = ops.launch_count()
        st.g1, st.g2, st.g3 = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        if peer is not None:
            peer.begin_capture()
        with torch.cuda.graph(st.g1, pool=pool, capture_error_mode="thread_local"):
            self._fresh(G)
            self._fresh(D)
            sr, _ = D.forward(st.real, True, sync, update_moving=self._moves('D'))
            y, gtape = G.forward_train(st.x, sync, update_moving=self._moves('G'))
            sf, dtape = D.forward(y, True, sync, update_moving=self._moves('D'))
            st.buf = torch.zeros(3, device=dev)
            ops.mean_prod(sr, None, st.buf[0:1], 1.0, 0.0, None, 1.0)
            ops.mean_prod(sf, None, st.buf[1:2], 1.0, 0.0, None, 1.0)
            self._fresh(self.content.vgg)
            pre = self.content.vgg.forward(y, save=True)
            st.ds = torch.empty_like(sf)
        if peer is not None:
            st.n1 = peer.end_capture()
            peer.begin_capture()
        S = self._scale_for(G, y.numel(), y.shape[0])
        with torch.cuda.graph(st.g2, pool=pool, capture_error_mode="thread_local"):
            grads = self._content(y, st.y, st.buf[2:3], cw, S, pre=pre)
        if peer is not None:
            st.n2 = peer.end_capture()
            peer.begin_capture()
        st.n = int(sf.shape[0])
        # the inner scale of the frozen critic depends on |d_fake|, a host value of each step (it shrinks as D wins): the
        # captured backward reads 1 / kD from a device vector that every replay fills, like the eager step computes it
        st.inv_kD = torch.ones(64, dtype=torch.float32, device=dev) if self.d_frozen else None
        with torch.cuda.graph(st.g3, pool=pool, capture_error_mode="thread_local"):
            grads.append(D.backward(dtape, st.ds, not self.d_frozen, True, sync, grad_scale=st.inv_kD if self.d_frozen else 1.0))
            dz = ops.tanh_bwd(y, grads[0], torch.empty_like(y), g2=grads[1], g3=grads[2] if len(grads) > 2 else None)
            G.grad_flat.zero_()
            G.backward(gtape, dz, sync, on_ready=None)
        if peer is not None:
            st.n3 = peer.end_capture()
        st.launches = ops.launch_count() - n0
        ops._count(-st.launches)
        st.S = S
        st.tape = {"head": gtape["head"], "ups": gtape["ups"]}
        return st

    def _step_gan2_graph(self, st, x, real, y_true):
        G, D = self.G, self.D
        cw, dw = self.weights
        world = self._world()
        if st.peer is not None and not st.peer.can_replay(st.n1 + st.n2 + st.n3):
            return None
        pending = _uploader.begin(y_true, "y")
        _uploader.finish_into(_uploader.begin(x, "x"), st.x)
        _uploader.finish_into(_uploader.begin(real, "real"), st.real)
        started = torch.cuda.Event()
        started.record()
        self._replay(st, st.g1, st.n1)
        G._train_dirty = D._train_dirty = False
        _uploader.finish_into(pending, st.y, side_stream=True, after=started)
        self._replay(st, st.g2, st.n2)
        mr, mf, lc = self._mean_over_ranks(st.buf)
        la, _, d_fake = self._relativistic(mr, mf, "gen")
        kD = self._inner_scale(D, abs(dw * d_fake * st.S / world) / max(st.n, 1)) if self.d_frozen else 1.0
        if st.inv_kD is not None:
            st.inv_kD.fill_(1.0 / kD)
        st.ds.fill_(dw * d_fake * st.S * kD / world / st.n)
        self._replay(st, st.g3, st.n3)
        ops._count(st.launches)
        dist = _dist()
        if dist is not None:
            dist.all_reduce(G.grad_flat)
        self._adam(G, st.S, also=D)
        self._range_tick(G, st.tape)
        self._range_tick(D)
        return [cw * lc + dw * la, lc, la]

    def step_gan2(self, x, real, y_true):
        """train_gan3.py:354 — outputs [G(x), D(G(x))]; generator loss from D(fake) and D(real), D frozen."""
        G, D = self.G, self.D
        dev = G.device
        if G.on_gpu:
            key = ("gan2", tuple(np.shape(x)), tuple(np.shape(real)), tuple(np.shape(y_true)), self.content.vgg.train_dtype,
                   self._world(), self._dtypes())
            st = self._graph_lookup(key, dev, lambda sync, peer: self._capture_gan2(x, real, y_true, sync, peer))
            if st is not None:
                r = self._step_gan2_graph(st, x, real, y_true)
                if r is not None:
                    return r
        self._fresh(G)
        self._fresh(D)
        cw, dw = self.weights
        xd, rd, yt = _to_dev(x, dev, "x"), _to_dev(real, dev, "real"), _to_dev(y_true, dev, "y")
        sync = _make_sync()
        dist = _dist()
        world = dist.get_world_size() if dist else 1
        sr, _ = D.forward(rd, True, sync, update_moving=self._moves('D'))
        y, gtape = G.forward_train(xd, sync, update_moving=self._moves('G'))
        sf, dtape = D.forward(y, True, sync, update_moving=self._moves('D'))
        buf = torch.zeros(3, device=dev)
        ops.mean_prod(sr, None, buf[0:1], 1.0, 0.0, None, 1.0)
        ops.mean_prod(sf, None, buf[1:2], 1.0, 0.0, None, 1.0)
        S = self._scale_for(G, y.numel(), y.shape[0])
        grads = self._content(y, yt, buf[2:3], cw, S)
        mr, mf, lc = self._mean_over_ranks(buf)
        la, _, d_fake = self._relativistic(mr, mf, "gen")
        ds = torch.empty_like(sf)
        kD = self._inner_scale(D, abs(dw * d_fake * S / world) / max(sf.shape[0], 1)) if self.d_frozen else 1.0
        ops.mean_prod(sf, None, None, 0.0, dw * d_fake * S * kD / world, ds, 1.0)
        grads.append(D.backward(dtape, ds, not self.d_frozen, True, sync, grad_scale=kD))
        dz = ops.tanh_bwd(y, grads[0], torch.empty_like(y), g2=grads[1], g3=grads[2] if len(grads) > 2 else None)
        G.ensure_grads().zero_()
        G.backward(gtape, dz, sync, on_ready=self.reducer.hook(G.grad_flat, G.n_trainable))
        self._adam(G, S, also=D)
        self._range_tick(G, gtape)
        self._range_tick(D)
        return [cw * lc + dw * la, lc, la]

    def _mean_over_ranks(self, buf):
        """Device scalars -> host floats (the step's single D2H read); averaged over replicas so every rank reports
        the global-batch loss."""
        import time
        dist = _dist()
        if dist is not None:
            dist.all_reduce(buf)
            buf = buf / dist.get_world_size()
        t0 = time.perf_counter()
        host = buf.cpu().numpy()
        HOST_STATS["wait_s"] += time.perf_counter() - t0
        HOST_STATS["reads"] += 1
        return [float(v) for v in host]


def evaluate_content_loss(loss_obj, y_true, y_pred):
    """Content loss of two host batches (float): the eager form of VGG_*_LOSS.loss."""
    c = Compiled("eval", None, None)
    c.content = ContentLoss(loss_obj.kind, loss_obj.rate, loss_obj.net)
    dev = loss_obj.net.device
    buf = torch.zeros(1, device=dev)
    vgg = loss_obj.net
    yt, yp = _to_dev(as_batch(y_true), dev, "eval_true"), _to_dev(as_batch(y_pred), dev, "eval_pred")
    c._fresh(vgg)
    ft, fp = vgg.forward(yt), vgg.forward(yp)
    ops.loss_pair(fp, ft, buf, mae=c.content.mae, weight=1.0, gscale=0.0, grad=None)
    if loss_obj.kind in ("vgg-mse", "vgg-mae"):
        ops.loss_pair(yp, yt, buf, mae=c.content.mae, weight=c.content.rate, gscale=0.0, grad=None)
    return float(buf.item())


# ------------------------------------------------------------------------------------------------ compile
def _leafs(applied):
    src, chain = applied.chain()
    nets, models = [], []
    for m in chain:
        for leaf in m.leaf_models():
            nets.append(leaf.net)
            models.append(leaf)
    return src, nets, models


def compile_model(model, loss, optimizer, loss_weights):
    from .networks import DiscriminatorSimple512Net, UpscalerOrigNet
    from . import model as M
    if model._is_leaf:
        outs = None
        nets_per_out = [[model.net]]
        models_per_out = [[model]]
        srcs = [model._inputs[0]]
    else:
        outs = model._outputs
        info = [_leafs(o) for o in outs]
        srcs = [i[0] for i in info]
        nets_per_out = [i[1] for i in info]
        models_per_out = [i[2] for i in info]
    losses = list(loss) if isinstance(loss, (list, tuple)) else [loss]
    if len(losses) != len(nets_per_out):
        raise ValueError("When passing a list as loss, it should have one entry per model output")

    def content_of(l):
        owner = getattr(l, "__self__", None)
        if isinstance(owner, M._VggLossBase):
            return ContentLoss(owner.kind, owner.rate, owner.net)
        raise ValueError("unsupported content loss %r: the B200 path implements VGG_LOSS / VGG_MSE_LOSS / VGG_MAE_LOSS" % (l,))

    def isG(n):
        if isinstance(n, UpscalerOrigNet) and not n.supports_training:
            raise NotImplementedError("%s: the inference path of this generator is built on the B200 kernels, its training "
                                      "step is not (train make_upscaler_orig / make_upscaler_skip_con)" % type(n).__name__)
        return isinstance(n, UpscalerOrigNet)
    isD = lambda n: isinstance(n, DiscriminatorSimple512Net)
    n_in = len(model._inputs)
    c = None
    if len(nets_per_out) == 1 and len(nets_per_out[0]) == 1 and isG(nets_per_out[0][0]) and n_in == 1:
        c = Compiled("gen", model, optimizer)
        c.G, c.content = nets_per_out[0][0], content_of(losses[0])
        c.G_model = models_per_out[0][0]
        c.g_frozen = not models_per_out[0][0].trainable
    elif len(nets_per_out) == 1 and len(nets_per_out[0]) == 1 and isD(nets_per_out[0][0]):
        D = nets_per_out[0][0]
        if n_in == 1 and losses[0] is M.wasserstein_loss:
            c = Compiled("disc", model, optimizer)
        elif n_in == 2 and isinstance(losses[0], GanLossFn) and losses[0].role == "disc":
            c = Compiled("disc2", model, optimizer)
            c.gan_losses = losses[0].owner
            c.fake_input_index = model._inputs.index(srcs[0])
        else:
            raise ValueError("unsupported discriminator loss %r" % (losses[0],))
        c.D = D
        c.D_model = models_per_out[0][0]
        c.d_frozen = not models_per_out[0][0].trainable
    elif len(nets_per_out) == 2 and len(nets_per_out[0]) == 1 and isG(nets_per_out[0][0]) and \
            len(nets_per_out[1]) == 2 and nets_per_out[1][0] is nets_per_out[0][0] and isD(nets_per_out[1][1]):
        if n_in == 1 and losses[1] is M.wasserstein_loss:
            c = Compiled("gan", model, optimizer)
        elif n_in == 2 and isinstance(losses[1], GanLossFn) and losses[1].role == "gen":
            c = Compiled("gan2", model, optimizer)
            c.gan_losses = losses[1].owner
            c.lr_input_index = model._inputs.index(srcs[0])
        else:
            raise ValueError("unsupported adversarial loss %r" % (losses[1],))
        c.G, c.D = nets_per_out[0][0], nets_per_out[1][1]
        c.G_model, c.D_model = models_per_out[0][0], models_per_out[1][1]
        c.content = content_of(losses[0])
        c.d_frozen = not models_per_out[1][1].trainable
        lw = list(loss_weights) if loss_weights is not None else [1.0, 1.0]
        c.weights = (float(lw[0]), float(lw[1]))
    if c is None:
        raise ValueError("compile: this model structure is not one the reference's compile helpers build "
                         "(model.py:1017-1137); see upscaler/training.py")
    return c


def _as4d(x):
    """Shape check of a batch argument WITHOUT a dtype conversion: float64 batches (what the reference's
    convert_image_series_to_array produces) are cast to float32 inside the multi-threaded staging copy (_Uploader.begin)."""
    a = np.asarray(x)
    if a.ndim != 4:
        raise ValueError("Error when checking input: expected 4 dimensions (batch, height, width, channels), got array "
                         "with shape %s" % (a.shape,))
    return a


def train_on_batch(model, x, y):
    c = model._compiled
    xs = list(x) if isinstance(x, (list, tuple)) else [x]
    ys = list(y) if isinstance(y, (list, tuple)) else [y]
    xs = [_as4d(a) for a in xs]
    if c.kind == "gen":
        return c.step_gen(xs[0], _as4d(ys[0]))
    if c.kind == "disc":
        return c.step_disc(xs[0], ys[0])
    if c.kind == "gan":
        return c.step_gan(xs[0], _as4d(ys[0]), ys[1])
    if c.kind == "disc2":
        fi = c.fake_input_index
        return c.step_disc2(xs[1 - fi], xs[fi])
    if c.kind == "gan2":
        li = c.lr_input_index
        return c.step_gan2(xs[li], xs[1 - li], _as4d(ys[0]))
    raise RuntimeError(c.kind)
