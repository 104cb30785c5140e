import torch
n = 8*540*960*256
a = torch.empty(n, dtype=torch.float16, device="cuda")
b = torch.empty(n, dtype=torch.float16, device="cuda")
def t(f, reps=20):
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
ms = t(lambda: a.zero_()); print("zero_ %.1f GB: %.3f ms, %.2f TB/s write" % (n*2/1e9, ms, n*2/ms/1e9))
ms = t(lambda: a.fill_(1.5)); print("fill_: %.3f ms, %.2f TB/s write" % (ms, n*2/ms/1e9))
ms = t(lambda: b.copy_(a)); print("copy: %.3f ms, %.2f TB/s r+w" % (ms, 2*n*2/ms/1e9))
ms = t(lambda: a.sum()); print("sum(read): %.3f ms, %.2f TB/s read" % (ms, n*2/ms/1e9))
