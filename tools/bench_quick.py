"""Quick device-side timing of generator inference (not the bench contract; a development aid)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
import numpy as np, torch
from upscaler.networks import UpscalerOrigNet
from upscaler import ops
if os.environ.get('B2_NO_PDL'):
    ops.set_pdl(False)      # A/B: plain stream order instead of programmatic dependent launch
ks = int(sys.argv[1]) if len(sys.argv) > 1 else 3
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1
net = UpscalerOrigNet(kernel_size=ks, seed=1)
if os.environ.get('B2_TRUNK_CHUNK'):
    net.TRUNK_CHUNK = int(os.environ['B2_TRUNK_CHUNK'])      # A/B: frames per trunk chunk of the inference pass
x = torch.rand(B, 270, 480, 3, device="cuda") * 2 - 1
out = torch.empty(B, 1080, 1920, 3, device="cuda")
for _ in range(3): net.forward_infer(x, out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 10
e0.record()
for _ in range(n): net.forward_infer(x, out)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
fl = {3: 1227.09e9, 5: 2942.77e9}[ks] * B
print("k=%d B=%d: %.3f ms/step, %.1f frames/s, %.1f TFLOP/s" % (ks, B, ms, B * 1000 / ms, fl / ms / 1e9))
