import time, numpy as np, torch
from concurrent.futures import ThreadPoolExecutor
a = np.random.default_rng(0).uniform(-1, 1, (12, 512, 512, 3)).astype(np.float32)
buf = torch.empty(a.shape, dtype=torch.float32).pin_memory()
dst = buf.numpy().reshape(-1); src = a.reshape(-1); n = src.size
def t(f, reps=10):
    f(); t0 = time.perf_counter()
    for _ in range(reps): f()
    return (time.perf_counter() - t0) / reps * 1e3
print("MB", a.nbytes / 1e6)
print("np.copyto 1 thread: %.2f ms" % t(lambda: np.copyto(dst, src)))
for T in (2, 4, 8):
    pool = ThreadPoolExecutor(T); step = -(-n // T)
    def par():
        fs = [pool.submit(np.copyto, dst[i:i + step], src[i:i + step]) for i in range(0, n, step)]
        for f in fs: f.result()
    print("np.copyto %d threads: %.2f ms" % (T, t(par)))
print("torch copy_ (threads=%d): %.2f ms" % (torch.get_num_threads(), t(lambda: buf.copy_(torch.from_numpy(a)))))
torch.set_num_threads(1)
print("torch copy_ (threads=1): %.2f ms" % t(lambda: buf.copy_(torch.from_numpy(a))))
print("pin_memory() fresh: %.2f ms" % t(lambda: torch.from_numpy(a).pin_memory()))
d = torch.empty(a.shape, device="cuda")
def h2d():
    d.copy_(buf, non_blocking=True); torch.cuda.synchronize()
print("H2D DMA: %.2f ms" % t(h2d))
def pageable():
    d.copy_(torch.from_numpy(a)); torch.cuda.synchronize()
print("H2D from pageable: %.2f ms" % t(pageable))
c = np.concatenate((a[:6], a[6:]), axis=0)
print("np.concatenate: %.2f ms" % t(lambda: np.concatenate((a[:6], a[6:]), axis=0)))
