"""Development aid: layer-by-layer comparison of the VGG19 backward pass with the bf16-emulating CPU oracle."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "video-cycle_gan-upscaling_b200"))
import numpy as np, torch
import torch.nn.functional as F
from oracle import srgan_oracle as O
from upscaler.networks import Vgg19Net, VGG19_CFG, VGG19_NAMES
from upscaler import ops

def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30)

vw = O.init_vgg19_weights()
tw = O.to_torch(vw)
rng = np.random.default_rng(1)
x = rng.uniform(-1, 1, (2, 64, 48, 3)).astype(np.float32)
target = rng.uniform(0, 1, (2, 4, 3, 512)).astype(np.float32)
q = O.f16_ste
xt = torch.from_numpy(x).requires_grad_(True)
h = q(xt.permute(0, 3, 1, 2))
acts_ref, i = [], 0
for c in VGG19_CFG:
    if c == "M":
        h = F.max_pool2d(h, 2, 2)
    else:
        k, b = tw[VGG19_NAMES[i]]
        h = q(F.relu(O.conv2d_same(h, q(k), b)))
        i += 1
    h.retain_grad()
    acts_ref.append(h)
loss = torch.mean((h.permute(0, 2, 3, 1) - torch.from_numpy(target).to(torch.float16).float()) ** 2)
loss.backward()

net = Vgg19Net(weights=vw)
fp, acts = net.forward(torch.from_numpy(x).cuda(), save=True)
for j, (a, r) in enumerate(zip(acts, acts_ref)):
    print("fwd %2d %-5s rel err %.5f" % (j, a[0], rel(a[2].float().cpu().numpy(), r.detach().permute(0, 2, 3, 1).numpy())))
ft = torch.from_numpy(target).cuda().to(torch.float16)
buf = torch.zeros(1, device="cuda")
dfeat = torch.empty_like(fp)
S = float(fp.numel()); ops.loss_pair(fp, ft, buf, gscale=S, grad=dfeat)
print("dfeat rel err %.5f" % rel(dfeat.float().cpu().numpy(), acts_ref[-1].grad.permute(0, 2, 3, 1).numpy()))
# replay backward layer by layer (same code as Vgg19Net.backward) printing the gradient w.r.t. each activation's input
pk = net._packed["train"]
top = acts[-1][2]
d = ops.bn_bwd_apply(top, dfeat, None, None, None, ops.ACT_RELU, 0.0, None, None, None, None, 0, torch.empty_like(top))
for j in range(len(acts) - 1, 0, -1):
    kind, name, _ = acts[j]
    below = acts[j - 1]
    if kind == "pool":
        continue
    cin = pk[name + "#dgrad"].shape[0]
    if below[0] == "conv":
        d = ops.conv2d(d, pk[name + "#dgrad"], cin, 3, 3, gate=below[2], gate_slope=0.0)
        # reference: grad wrt pre-activation of layer j-1 = grad(act_{j-1}) * (act>0)
        r = acts_ref[j - 1]
        ref = (r.grad * (r.detach() > 0)).permute(0, 2, 3, 1).numpy()
        print("bwd into %2d (conv) rel err %.5f" % (j - 1, rel(d.float().cpu().numpy(), ref)))
    else:
        _, a_in, p_out = below
        dp = ops.conv2d(d, pk[name + "#dgrad"], cin, 3, 3)
        print("bwd into %2d (pool out) rel err %.5f" % (j - 1, rel(dp.float().cpu().numpy(), acts_ref[j - 1].grad.permute(0, 2, 3, 1).numpy())))
        d = ops.maxpool2x2_bwd(a_in, p_out, dp, relu_gate=True)
        r = acts_ref[j - 2]
        ref = (r.grad * (r.detach() > 0)).permute(0, 2, 3, 1).numpy()
        print("bwd into %2d (conv below pool) rel err %.5f" % (j - 2, rel(d.float().cpu().numpy(), ref)))
