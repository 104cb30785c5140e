"""Development aid: end-to-end Model.predict(host ndarray) -> host ndarray throughput for several pipeline stage sizes,
plus the raw pinned D2H bandwidth of the box (the floor of the result copy)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
import numpy as np, torch
from upscaler import engine
from upscaler.model import make_upscaler_orig
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
model = make_upscaler_orig((1080, 1920, 3), kernel_size=3)
x = np.random.default_rng(0).uniform(-1, 1, (B, 270, 480, 3)).astype(np.float32)
d = torch.empty((B, 1080, 1920, 3), dtype=torch.float32, device="cuda")
h = torch.empty(d.shape, dtype=torch.float32, pin_memory=True)
for _ in range(2):
    h.copy_(d, non_blocking=True); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    h.copy_(d, non_blocking=True)
torch.cuda.synchronize()
print("pinned D2H: %.1f GB/s" % (5 * d.numel() * 4 / (time.perf_counter() - t0) / 1e9))
ref = None
for pf in (8, 4, 2, 1):
    engine.PIPELINE_FRAMES = pf
    for _ in range(3):
        y = model.predict(x, batch_size=B)
    t0 = time.perf_counter()
    for _ in range(8):
        y = model.predict(x, batch_size=B)
    dt = (time.perf_counter() - t0) / 8
    if ref is None:
        ref = y.copy()
    print("PIPELINE_FRAMES=%d: %.2f ms per %d frames, %.1f frames/s e2e, max diff vs unpipelined %.2g" % (
        pf, dt * 1e3, B, B / dt, float(np.abs(y - ref).max())))
