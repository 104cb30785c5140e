// b2_conv_direct.cu — CUDA-core convolution that follows the SAME tap list / packed-weight / output-mapping
// parameter block as the tcgen05 kernel, one thread per (pixel, output channel), fp32 accumulate.
// It is the on-device cross-check for the tensor-core kernel (selftest and -m gpu tests); the model path does not
// route through it.
#include "b2_conv_igemm.cuh"
#include "b2_host.h"

namespace b2 {

__device__ __forceinline__ float direct_act(float v, int act, float leaky, float alpha) {
    switch (act) {
        case ACT_RELU: return fmaxf(v, 0.f);
        case ACT_LEAKY: return v > 0.f ? v : v * leaky;
        case ACT_PRELU: return v > 0.f ? v : v * alpha;
        case ACT_TANH: return tanhf(v);
        default: return v;
    }
}

__global__ void conv_direct_kernel(const __nv_bfloat16* __restrict__ x, int xB, int xH, int xW, int xC,
                                   const __nv_bfloat16* __restrict__ w, int w_rows, int w_k, IgemmParams p) {
    b2::pdl_entry();
    const long total = long(p.B) * p.H * p.W * p.n_valid;
    for (long idx = blockIdx.x * long(blockDim.x) + threadIdx.x; idx < total; idx += long(gridDim.x) * blockDim.x) {
        const int n = int(idx % p.n_valid);
        long t = idx / p.n_valid;
        const int gx = int(t % p.W); t /= p.W;
        const int gy = int(t % p.H); t /= p.H;
        const int gb = int(t);
        float acc = 0.f;
        for (int tap = 0; tap < p.num_taps; ++tap) {
            const int ix = gx * p.in_mul + p.dx[tap], iy = gy * p.in_mul + p.dy[tap];
            if (ix < 0 || iy < 0 || ix >= xW || iy >= xH) continue;
            const __nv_bfloat16* xp = x + ((size_t(gb) * xH + iy) * xW + ix) * xC;
            const __nv_bfloat16* wp = w + size_t(n) * w_k + size_t(p.wtap[tap]) * p.cin;
            if (p.f16) {
                const __half* xh = reinterpret_cast<const __half*>(xp);
                const __half* wh = reinterpret_cast<const __half*>(wp);
                for (int c = 0; c < p.cin; ++c) acc = fmaf(__half2float(xh[c]), __half2float(wh[c]), acc);
            } else {
                for (int c = 0; c < p.cin; ++c) acc = fmaf(__bfloat162float(xp[c]), __bfloat162float(wp[c]), acc);
            }
        }
        const float sc = p.scale ? p.scale[n] : 1.f;
        const float sh = p.shift ? p.shift[n] : 0.f;
        const float al = p.act == ACT_PRELU ? p.alpha[n] : 0.f;
        float v = direct_act(fmaf(acc, sc, sh), p.act, p.leaky, al);
        const size_t pix = (size_t(gb) * p.OH + size_t(gy * p.out_sy + p.out_oy)) * p.OW + size_t(gx * p.out_sx + p.out_ox);
        const size_t o = pix * size_t(p.out_ld) + n;
        if (p.gate) {
            const float gv = p.f16 ? __half2float(reinterpret_cast<const __half*>(p.gate)[o])
                                   : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.gate)[o]);
            if (!(gv > 0.f)) v *= p.gate_slope;
        }
        if (p.residual)
            v += p.f16 ? __half2float(reinterpret_cast<const __half*>(p.residual)[o]) : __bfloat162float(p.residual[o]);
        if (p.out_fp32) reinterpret_cast<float*>(p.out)[o] = v;
        else if (p.f16) reinterpret_cast<__half*>(p.out)[o] = __float2half_rn(v);
        else reinterpret_cast<__nv_bfloat16*>(p.out)[o] = __float2bfloat16_rn(v);
    }
}

int launch_conv_direct(const ConvIgemmArgs& a, cudaStream_t stream) {
    const long total = long(a.p.B) * a.p.H * a.p.W * a.p.n_valid;
    long blocks = (total + 255) / 256;
    if (blocks > 148L * 64) blocks = 148L * 64;
    if (blocks < 1) blocks = 1;
    b2::launch_k(conv_direct_kernel, int(blocks), 256, 0, stream, a.x, a.xB, a.xH, a.xW, a.xC, a.w, a.w_rows, a.w_k, a.p);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("conv_direct launch: %s", cudaGetErrorString(e));
    return 0;
}

}  // namespace b2
