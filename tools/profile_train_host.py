"""Development aid: host-side timeline of one generator training step (where the CPU spends its time between launches)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
import numpy as np, torch
from upscaler.model import VGG_MSE_LOSS, compile_training_model, make_upscaler_orig
from upscaler import ops, training
if int(os.environ.get("WORLD_SIZE", "1")) > 1:
    import torch.distributed as dist
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
RANK = int(os.environ.get("RANK", "0"))
B, lr = 16, 96
up = make_upscaler_orig((4 * lr, 4 * lr, 3), kernel_size=3)
tm = compile_training_model(up, VGG_MSE_LOSS((4 * lr, 4 * lr, 3), 1e-3).loss)
rng = np.random.default_rng(0)
hr = rng.uniform(-1, 1, (B, 4 * lr, 4 * lr, 3)).astype(np.float32)
x = hr.reshape(B, lr, 4, lr, 4, 3).mean(axis=(2, 4)).astype(np.float32)
for _ in range(3):
    tm.train_on_batch(x, hr)
C = tm._compiled if hasattr(tm, "_compiled") else None
marks = []
def wrap(obj, name):
    f = getattr(obj, name)
    def g(*a, **k):
        t0 = time.perf_counter(); r = f(*a, **k); marks.append((name, (time.perf_counter() - t0) * 1e3)); return r
    setattr(obj, name, g)
comp = [v for v in vars(tm).values() if isinstance(v, training.Compiled)]
comp = comp[0] if comp else None
print("compiled:", comp)
if comp is not None:
    wrap(comp.G, "forward_train"); wrap(comp.G, "backward"); wrap(comp, "_content"); wrap(comp, "_adam")
    wrap(comp, "_mean_over_ranks"); wrap(comp.reducer, "wait")
for it in range(4):
    marks.clear()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    l = tm.train_on_batch(x, hr)
    tot = (time.perf_counter() - t0) * 1e3
    if RANK == 0:
        print("step %.2f ms: " % tot + ", ".join("%s %.2f" % m for m in marks))
# GPU-only time of the same step: launch everything, then sync
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); l = tm.train_on_batch(x, hr); e1.record(); torch.cuda.synchronize()
ops.reset_launch_count(); tm.train_on_batch(x, hr)      # every rank: the step holds collectives
if RANK == 0:
    print("event-timed step: %.2f ms" % e0.elapsed_time(e1))
    print("launches per step:", ops.launch_count())
