// b2_elementwise.cu — memory-bound helper kernels of the path: tap folding for 3-channel inputs, weight repacking.
// All are vectorised (16-byte stores) grid-stride kernels sized to a multiple of the SM count.
#include "../../include/b2srgan.h"
#include <cuda_fp16.h>

#include "b2_host.h"

using namespace b2;

namespace {

constexpr int kFoldMaxTaps = 21;  // 21 taps x 3 channels = 63 <= 64

struct FoldParams {
    int B, H, W, Cin, OH, OW, stride, ntaps, f16, out_ld, out_c0;
    int dy[kFoldMaxTaps], dx[kFoldMaxTaps];
};

// One CTA walks whole output rows (no per-element 64-bit divisions); the channel -> (tap, colour) map is decoded once
// per CTA into shared memory.  Thread = (pixel, group of 8 channels): one 16-byte store, 8 L1-served scalar loads.
template <typename TIn>
__global__ void __launch_bounds__(256) fold_taps_kernel(const TIn* __restrict__ in, void* __restrict__ out, FoldParams p) {
    b2::pdl_entry();
    // per output channel: x = element offset (dy * W + dx) * Cin + c from the stride-scaled output pixel,
    //                     y = dy, z = dx (dy = -2^20 marks an unused channel: always out of range -> zero)
    __shared__ int4 s_tab[64];     // index = (ch % 8) * 8 + ch / 8
    if (threadIdx.x < 64) {
        const int ch = threadIdx.x;
        if (ch < p.ntaps * p.Cin) {
            const int tap = ch / p.Cin, c = ch - tap * p.Cin;
            s_tab[(ch & 7) * 8 + (ch >> 3)] = make_int4((p.dy[tap] * p.W + p.dx[tap]) * p.Cin + c, p.dy[tap], p.dx[tap], 0);
        } else {
            s_tab[(ch & 7) * 8 + (ch >> 3)] = make_int4(0, -(1 << 20), 0, 0);
        }
    }
    __syncthreads();
    const int rows = p.B * p.OH;
    const int per_row = p.OW * 8;
    for (int row = blockIdx.x; row < rows; row += gridDim.x) {
        const int b = row / p.OH, oy = row - b * p.OH;
        const int iy0 = oy * p.stride;
        const TIn* src = in + (size_t(b) * p.H + iy0) * p.W * p.Cin;
        uint16_t* dst = reinterpret_cast<uint16_t*>(out) + size_t(row) * p.OW * p.out_ld + p.out_c0;
        for (int i = threadIdx.x; i < per_row; i += blockDim.x) {
            const int g = i & 7, ox = i >> 3;
            const int ix0 = ox * p.stride;
            const TIn* px = src + ix0 * p.Cin;
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int ch = g * 8 + e;
                const int4 t = s_tab[e * 8 + g];      // transposed: the 8 groups of a warp hit 8 different bank quads
                const int iy = iy0 + t.y, ix = ix0 + t.z;
                float f = 0.f;
                if (iy >= 0 && iy < p.H && ix >= 0 && ix < p.W) f = float(px[t.x]);
                v[e] = f;
            }
            uint4 o;
            if (p.f16) {
                __half2* h = reinterpret_cast<__half2*>(&o);
#pragma unroll
                for (int e = 0; e < 4; ++e) h[e] = __floats2half2_rn(v[2 * e], v[2 * e + 1]);
            } else {
                __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
                for (int e = 0; e < 4; ++e) h[e] = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
            }
            *reinterpret_cast<uint4*>(dst + size_t(ox) * p.out_ld + g * 8) = o;
        }
    }
}

__global__ void gather_cast_kernel(const float* __restrict__ src, const int32_t* __restrict__ idx, long n,
                                   void* __restrict__ out, int f16) {
    b2::pdl_entry();
    for (long i = blockIdx.x * long(blockDim.x) + threadIdx.x; i < n; i += long(gridDim.x) * blockDim.x) {
        const int32_t j = idx[i];
        const float v = j >= 0 ? src[j] : 0.f;
        if (f16) reinterpret_cast<__half*>(out)[i] = __float2half_rn(v);
        else reinterpret_cast<__nv_bfloat16*>(out)[i] = __float2bfloat16_rn(v);
    }
}

// K.resize_images(x, f, f, "channels_last", "bilinear") of make_upscaler_skip_con (reference model.py:357) = TensorFlow 1.x
// tf.image.resize_bilinear with align_corners = False and no half-pixel centres: source coordinate = out * (in / out),
// y0 = floor, y1 = min(y0 + 1, in - 1), top/bottom rows lerped in x, then in y -- all in fp32.  The result is written as
// 16-bit elements into channels [c0, c0 + C) of an NHWC tensor with `ld` channels per pixel (the Concatenate of model.py:358
// becomes a channel slice of the tensor the final conv reads).  One thread per output pixel.
__global__ void __launch_bounds__(256)
resize_bilinear_into_kernel(const float* __restrict__ in, int B, int H, int W, int C, int factor, void* __restrict__ out,
                            int ld, int c0, int f16) {
    b2::pdl_entry();
    const int OH = H * factor, OW = W * factor;
    const float inv = 1.0f / float(factor);
    for (int row = blockIdx.x; row < B * OH; row += gridDim.x) {
        const int b = row / OH, oy = row - b * OH;
        const float sy = float(oy) * inv;
        const int y0 = int(floorf(sy)), y1 = min(y0 + 1, H - 1);
        const float fy = sy - float(y0);
        const float* r0 = in + (size_t(b) * H + y0) * W * C;
        const float* r1 = in + (size_t(b) * H + y1) * W * C;
        uint16_t* dst = reinterpret_cast<uint16_t*>(out) + size_t(row) * OW * ld + c0;
        for (int ox = threadIdx.x; ox < OW; ox += blockDim.x) {
            const float sx = float(ox) * inv;
            const int x0 = int(floorf(sx)), x1 = min(x0 + 1, W - 1);
            const float fx = sx - float(x0);
            for (int c = 0; c < C; ++c) {
                const float tl = r0[x0 * C + c], tr = r0[x1 * C + c], bl = r1[x0 * C + c], br = r1[x1 * C + c];
                const float top = tl + (tr - tl) * fx, bot = bl + (br - bl) * fx;
                const float v = top + (bot - top) * fy;
                uint16_t h;
                if (f16) { const __half t = __float2half_rn(v); h = *reinterpret_cast<const uint16_t*>(&t); }
                else { const __nv_bfloat16 t = __float2bfloat16_rn(v); h = *reinterpret_cast<const uint16_t*>(&t); }
                dst[size_t(ox) * ld + c] = h;
            }
        }
    }
}

inline int grid_for(long work, int threads) {
    long blocks = (work + threads - 1) / threads;
    const long cap = long(device_sm_count()) * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return int(blocks);
}

}  // namespace

// Frame I/O edges (reference: upscaling/upscaler/data.py:253-270).  Vectorised, coalesced: 16 pixels-channels per thread.
//   u8 -> f32:  x / 127.5 - 1, evaluated in double like NumPy does on a uint8 array, then rounded once to float (what
//               Keras feeds the network); bit-exact with convert_image_series_to_array(...).astype(float32).
//   f32 -> u8:  uint8(around((a + 1) * 127.5)) in float32 with separate add and multiply (no FMA contraction) and
//               round-half-to-even, bit-exact with convert_array_to_image on a float32 prediction.
__global__ void u8_to_f32_norm_kernel(const uint8_t* __restrict__ in, float* __restrict__ out, long n) {
    b2::pdl_entry();
    const long nv = n >> 4;
    for (long i = blockIdx.x * long(blockDim.x) + threadIdx.x; i < nv; i += long(gridDim.x) * blockDim.x) {
        const uint4 u = __ldg(reinterpret_cast<const uint4*>(in) + i);
        const uint8_t* b = reinterpret_cast<const uint8_t*>(&u);
        float4 o[4];
        float* f = reinterpret_cast<float*>(o);
#pragma unroll
        for (int e = 0; e < 16; ++e) f[e] = float(__dsub_rn(__ddiv_rn(double(b[e]), 127.5), 1.0));
#pragma unroll
        for (int e = 0; e < 4; ++e) reinterpret_cast<float4*>(out)[i * 4 + e] = o[e];
    }
    for (long i = (nv << 4) + blockIdx.x * long(blockDim.x) + threadIdx.x; i < n; i += long(gridDim.x) * blockDim.x)
        out[i] = float(__dsub_rn(__ddiv_rn(double(in[i]), 127.5), 1.0));
}

__device__ __forceinline__ uint8_t denorm_u8(float a) {
    const float v = rintf(__fmul_rn(__fadd_rn(a, 1.0f), 127.5f));
    return uint8_t(int(fminf(fmaxf(v, 0.f), 255.f)));
}
__global__ void f32_to_u8_denorm_kernel(const float* __restrict__ in, uint8_t* __restrict__ out, long n) {
    b2::pdl_entry();
    const long nv = n >> 4;
    for (long i = blockIdx.x * long(blockDim.x) + threadIdx.x; i < nv; i += long(gridDim.x) * blockDim.x) {
        float4 v[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) v[e] = __ldg(reinterpret_cast<const float4*>(in) + i * 4 + e);
        const float* f = reinterpret_cast<const float*>(v);
        uint4 o;
        uint8_t* b = reinterpret_cast<uint8_t*>(&o);
#pragma unroll
        for (int e = 0; e < 16; ++e) b[e] = denorm_u8(f[e]);
        reinterpret_cast<uint4*>(out)[i] = o;
    }
    for (long i = (nv << 4) + blockIdx.x * long(blockDim.x) + threadIdx.x; i < n; i += long(gridDim.x) * blockDim.x)
        out[i] = denorm_u8(in[i]);
}

// Counts the elements of a 16-bit tensor that sit at the edge of their type: fp16 |x| >= 65504 (what the conv epilogues'
// cvt.rn.satfinite stores for anything larger), inf or NaN; bf16 inf or NaN.  One atomic per CTA, and only when non-zero.
__global__ void count_saturated_kernel(const uint4* __restrict__ x, long n8, const uint16_t* __restrict__ tail, int ntail,
                                       uint32_t limit, unsigned int* __restrict__ count) {
    b2::pdl_entry();
    unsigned int c = 0;
    for (long i = blockIdx.x * long(blockDim.x) + threadIdx.x; i < n8; i += long(gridDim.x) * blockDim.x) {
        const uint4 v = __ldg(x + i);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            c += ((w[e] & 0x7FFFu) >= limit) ? 1u : 0u;
            c += (((w[e] >> 16) & 0x7FFFu) >= limit) ? 1u : 0u;
        }
    }
    if (blockIdx.x == 0 && int(threadIdx.x) < ntail) c += ((tail[threadIdx.x] & 0x7FFFu) >= limit) ? 1u : 0u;
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    __shared__ unsigned int sh[8];
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int t = 0;
        for (int w = 0; w < int(blockDim.x >> 5); ++w) t += sh[w];
        if (t) atomicAdd(count, t);
    }
}

extern "C" {

int b2_count_saturated(const void* x, int64_t n, int f16, unsigned int* count, void* stream) {
    if (!x || !count) return set_error("b2_count_saturated: null pointer");
    if (reinterpret_cast<uintptr_t>(x) & 15) return set_error("b2_count_saturated: tensor must be 16-byte aligned");
    if (n <= 0) return 0;
    const long n8 = n >> 3;
    b2::launch_k(count_saturated_kernel, grid_for(n8 > 0 ? n8 : 1, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<const uint4*>(x), n8, reinterpret_cast<const uint16_t*>(x) + (n8 << 3), int(n & 7),
        f16 ? 0x7BFFu : 0x7F80u, count);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("b2_count_saturated launch: %s", cudaGetErrorString(e));
    return 0;
}

int b2_fold_taps(const void* in, int in_fp32, int B, int H, int W, int Cin, int OH, int OW, int stride, int ntaps,
                 const int* dy, const int* dx, void* out, int f16, int out_ld, int out_c0, void* stream) {
    if (!in || !out || !dy || !dx) return set_error("b2_fold_taps: null pointer");
    if (ntaps < 1 || ntaps > kFoldMaxTaps || ntaps * Cin > 64) return set_error("b2_fold_taps: %d taps x %d channels do not fit 64", ntaps, Cin);
    FoldParams p;
    if (out_ld < 64 || (out_ld & 7) || (out_c0 & 7) || out_c0 + 64 > out_ld) return set_error("b2_fold_taps: bad output pitch %d / offset %d", out_ld, out_c0);
    p.B = B; p.H = H; p.W = W; p.Cin = Cin; p.OH = OH; p.OW = OW; p.stride = stride; p.ntaps = ntaps; p.f16 = f16;
    p.out_ld = out_ld; p.out_c0 = out_c0;
    for (int t = 0; t < ntaps; ++t) { p.dy[t] = dy[t]; p.dx[t] = dx[t]; }
    const int grid = grid_for(long(B) * OH * 256, 256);     // one CTA per output row, capped at 16 CTAs per SM
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    if (in_fp32)
        b2::launch_k(fold_taps_kernel<float>, grid, 256, 0, s, reinterpret_cast<const float*>(in), out, p);
    else if (f16)
        b2::launch_k(fold_taps_kernel<__half>, grid, 256, 0, s, reinterpret_cast<const __half*>(in), out, p);
    else
        b2::launch_k(fold_taps_kernel<__nv_bfloat16>, grid, 256, 0, s, reinterpret_cast<const __nv_bfloat16*>(in), out, p);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("b2_fold_taps launch: %s", cudaGetErrorString(e));
    return 0;
}

int b2_gather_cast(const float* src, const int32_t* idx, int64_t n, void* out, int f16, void* stream) {
    if (!src || !idx || !out) return set_error("b2_gather_cast: null pointer");
    b2::launch_k(gather_cast_kernel, grid_for(n, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream), src, idx, long(n), out, f16);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("b2_gather_cast launch: %s", cudaGetErrorString(e));
    return 0;
}

int b2_u8_to_f32_norm(const uint8_t* in, float* out, int64_t n, void* stream) {
    if (!in || !out) return set_error("b2_u8_to_f32_norm: null pointer");
    if ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) return set_error("b2_u8_to_f32_norm: buffers must be 16-byte aligned");
    if (n <= 0) return 0;
    b2::launch_k(u8_to_f32_norm_kernel, grid_for((n + 15) / 16, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream), in, out, long(n));
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("b2_u8_to_f32_norm launch: %s", cudaGetErrorString(e));
    return 0;
}

int b2_resize_bilinear_into(const float* in, int B, int H, int W, int C, int factor, void* out, int ld, int c0, int f16,
                            void* stream) {
    if (!in || !out) return set_error("b2_resize_bilinear_into: null pointer");
    if (B < 1 || H < 1 || W < 1 || C < 1 || factor < 1 || c0 < 0 || c0 + C > ld)
        return set_error("b2_resize_bilinear_into: bad geometry (%d,%d,%d,%d) x%d into channels [%d,%d) of %d", B, H, W, C, factor, c0, c0 + C, ld);
    b2::launch_k(resize_bilinear_into_kernel, grid_for(long(B) * H * factor * 256, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream), in, B, H, W, C, factor, out, ld, c0, f16);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("b2_resize_bilinear_into launch: %s", cudaGetErrorString(e));
    return 0;
}

int b2_f32_to_u8_denorm(const float* in, uint8_t* out, int64_t n, void* stream) {
    if (!in || !out) return set_error("b2_f32_to_u8_denorm: null pointer");
    if ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) return set_error("b2_f32_to_u8_denorm: buffers must be 16-byte aligned");
    if (n <= 0) return 0;
    b2::launch_k(f32_to_u8_denorm_kernel, grid_for((n + 15) / 16, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream), in, out, long(n));
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("b2_f32_to_u8_denorm launch: %s", cudaGetErrorString(e));
    return 0;
}

}  // extern "C"
