"""In-tree build of the CUDA extension: nvcc -> csrc/libb2srgan.so (sm_100a only).

nvcc cross-compiles without a GPU, so this runs in the build container; the resulting .so travels to the GPU box
with the repo snapshot.  Nothing here touches torch's JIT cache.
"""
import hashlib
import os
import subprocess

CSRC = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "csrc"))
LIB_PATH = os.path.join(CSRC, "libb2srgan.so")
SOURCES = ["b2_api.cu", "b2_host.cu", "b2_conv_igemm.cu", "b2_conv_direct.cu", "b2_conv_tail.cu", "b2_conv_halo.cu",
           "b2_elementwise.cu", "b2_conv_wgrad.cu", "b2_conv_scatter.cu", "b2_train_elementwise.cu", "b2_conv_stack.cu", "b2_conv_phases.cu", "b2_conv_phases2.cu", "b2_peer.cu", "b2_bn_train.cu", "b2_attention.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC"]
if os.environ.get("B2_BUILD_PROFILING"):       # tools/ only: compiles the environment-driven A/B switches in (b2_host.h)
    NVCC_FLAGS.append("-DB2_PROFILING")


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def _fingerprint():
    h = hashlib.sha256()
    for name in sorted(os.listdir(CSRC)):
        if name.endswith((".cu", ".cuh", ".h")):
            with open(os.path.join(CSRC, name), "rb") as f:
                h.update(name.encode())
                h.update(f.read())
    inc = os.path.normpath(os.path.join(CSRC, "..", "..", "include", "b2srgan.h"))
    with open(inc, "rb") as f:
        h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force=False, verbose=False):
    """Compile every .cu under csrc/ for sm_100a and link libb2srgan.so.  Skips work when sources are unchanged."""
    stamp = os.path.join(CSRC, ".build_stamp")
    fp = _fingerprint()
    if not force and os.path.exists(LIB_PATH) and os.path.exists(stamp):
        with open(stamp) as f:
            if f.read().strip() == fp:
                return LIB_PATH
    objdir = os.path.join(CSRC, "build")
    os.makedirs(objdir, exist_ok=True)
    nvcc = _nvcc()
    procs = []
    objs = []
    for src in SOURCES:
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        objs.append(obj)
        cmd = [nvcc] + NVCC_FLAGS + ["-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            print(" ".join(cmd))
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for src, pr in procs:
        out, _ = pr.communicate()
        if pr.returncode != 0:
            raise RuntimeError("nvcc failed on %s:\n%s" % (src, out.decode(errors="replace")))
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB_PATH] + objs + ["-lcudart"]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if res.returncode != 0:
        raise RuntimeError("link failed:\n%s" % res.stdout.decode(errors="replace"))
    with open(stamp, "w") as f:
        f.write(fp)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force=True, verbose=True))
