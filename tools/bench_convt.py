"""Development aid: times the second Conv2DTranspose (256 -> 256 at 540x960, B frames) alone through ops.conv2d_transpose_s2."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
import numpy as np, torch
from upscaler import ops
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
cin = int(sys.argv[2]) if len(sys.argv) > 2 else 256
H, W = (540, 960) if cin == 256 else (270, 480)
k = 3
x = (torch.randn(B, H, W, cin, device="cuda") * 0.5).half()
kern = torch.randn(k, k, 256, cin, device="cuda") / (k * k * cin / 4) ** 0.5
wp = ops.gather_cast(kern, torch.from_numpy(ops.convT_pack_index(k, k, cin, 256)).cuda(), dtype=torch.float16)
bias = torch.randn(256, device="cuda") * 0.1
out = torch.empty(B, 2 * H, 2 * W, 256, device="cuda", dtype=torch.float16)
for pair in (True, False):
    ops.set_pair_kernels(pair)
    for _ in range(3):
        ops.conv2d_transpose_s2(x, wp, 256, k, k, out=out, shift=bias, act=ops.ACT_LEAKY, leaky=0.2)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 20
    e0.record()
    for _ in range(n):
        ops.conv2d_transpose_s2(x, wp, 256, k, k, out=out, shift=bias, act=ops.ACT_LEAKY, leaky=0.2)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    fl = 2.0 * B * H * W * 9 * cin * 256
    print("ConvT %d->256 B=%d pair=%s: %.3f ms, %.1f TFLOP/s" % (cin, B, pair, ms, fl / ms / 1e9))
