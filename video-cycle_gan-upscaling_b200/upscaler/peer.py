"""Cross-replica sums of small fp32 vectors over NVLink peer memory (csrc/b2_peer.cu) for the data-parallel training
step: the BatchNormalization statistics of N replicas, ~70 latency-bound reductions per step that sit on the
critical path.  torch.distributed carries the one-time exchange of the CUDA IPC handles; every reduction afterwards is one
single-CTA kernel on the compute stream.  B2_DP_NO_PEER=1 keeps these reductions on NCCL (A/B timing)."""
import ctypes
import os

from . import _lib

_state = {"reducer": None, "tried": False}


class PeerReducer:
    def __init__(self, dist, device):
        import torch
        lib = _lib.load()
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.max_elems = lib.b2_peer_max_elems()
        with torch.cuda.device(device):
            own = ctypes.c_void_p()
            handle = ctypes.create_string_buffer(64)
            _lib.check(lib.b2_peer_alloc(ctypes.byref(own), handle), "b2_peer_alloc")
            handles = [None] * self.world
            dist.all_gather_object(handles, bytes(handle.raw))
            self.regions = (ctypes.c_void_p * self.world)()
            for p in range(self.world):
                if p == self.rank:
                    self.regions[p] = own.value
                else:
                    q = ctypes.c_void_p()
                    _lib.check(lib.b2_peer_open(handles[p], ctypes.byref(q)), "b2_peer_open")
                    self.regions[p] = q.value
        dist.barrier()
        self.epoch = 0
        self.calls = 0
        # CUDA-graph replay of a training step (training.py): a captured launch bakes its arguments, so while a step is being
        # captured next_epoch() hands out the exchange's POSITION in the step (1, 2, ...) and the kernels add the device word
        # `base_dev`, which the host sets to `epoch` (exchanges completed so far) before every replay.
        self.base_dev = torch.zeros(1, dtype=torch.int32, device=device)
        self._cap = None

    def next_epoch(self):
        """Epoch of the next exchange (identical call sequence on every rank); 1 .. 0x7FFFFFFE, never 0."""
        if self._cap is not None:
            self._cap += 1
            return self._cap
        self.epoch = (self.epoch % 0x7FFFFFFE) + 1
        self.calls += 1
        return self.epoch

    def epoch_base_ptr(self):
        return _lib.ptr(self.base_dev) if self._cap is not None else None

    def begin_capture(self):
        self._cap = 0

    def end_capture(self):
        n, self._cap = self._cap or 0, None
        return n

    def can_replay(self, n):
        return self.epoch + n < 0x7FFFFFF0

    def before_replay(self):
        """Enqueue (current stream) the epoch base of the captured exchanges that follow."""
        self.base_dev.fill_(self.epoch)

    def after_replay(self, n):
        self.epoch += n
        self.calls += n

    def all_reduce(self, buf):
        """In-place sum over the ranks of a contiguous float32 CUDA tensor of at most max_elems elements."""
        import torch
        if buf.dtype != torch.float32 or not buf.is_cuda or not buf.is_contiguous() or buf.numel() > self.max_elems:
            raise ValueError("peer all-reduce takes contiguous float32 CUDA tensors of at most %d elements" % self.max_elems)
        epoch = self.next_epoch()
        rc = _lib.load().b2_peer_allreduce_f32(_lib.ptr(buf), buf.numel(), self.regions, self.rank, self.world, epoch,
                                               self.epoch_base_ptr(), _lib.stream_ptr())
        _lib.check(rc, "b2_peer_allreduce_f32")
        from . import ops
        ops._count(1)
        return buf


    def exchange_probe(self, n, iters):
        """Measurement aid: `iters` dependent exchanges of n floats inside one launch; returns SM clock cycles per exchange."""
        import torch
        buf = torch.ones(n, dtype=torch.float32, device="cuda")
        cyc = torch.zeros(2, dtype=torch.int64, device="cuda")
        e0 = self.next_epoch()
        for _ in range(iters - 1):
            self.next_epoch()
        rc = _lib.load().b2_peer_exchange_probe(_lib.ptr(buf), n, self.regions, self.rank, self.world, e0, iters, _lib.ptr(cyc),
                                                _lib.stream_ptr())
        _lib.check(rc, "b2_peer_exchange_probe")
        torch.cuda.synchronize()
        return float(cyc[0].item()) / iters


def reducer(dist, device):
    """The process-wide PeerReducer, or None when the ranks cannot reach each other's memory (then NCCL is used).  The
    decision is collective: every rank reports whether it could open all regions, and all follow the minimum."""
    if _state["tried"]:
        return _state["reducer"]
    _state["tried"] = True
    import torch
    ok, red = 1, None
    if os.environ.get("B2_DP_NO_PEER") or dist.get_backend() != "nccl" or dist.get_world_size() > 8:
        ok = 0
    if ok:
        dev = torch.device(device)
        n = torch.cuda.device_count()
        if dist.get_world_size() > n or not all(torch.cuda.can_device_access_peer(dev.index, j) for j in range(n) if j != dev.index):
            ok = 0
    flag = torch.tensor([ok], device=device, dtype=torch.int32)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if int(flag.item()) == 1:
        try:
            red = PeerReducer(dist, device)
            good = 1
        except _lib.B2Error as e:            # e.g. IPC disabled in the container: say so once, stay on NCCL
            print("[b2] peer all-reduce unavailable (%s); BatchNorm sums stay on NCCL" % e, flush=True)
            good = 0
        flag.fill_(good)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) != 1:
            red = None
    _state["reducer"] = red
    return red
