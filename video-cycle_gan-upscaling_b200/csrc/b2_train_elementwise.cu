// b2_train_elementwise.cu — the memory-bound kernels of the training step: train-mode BatchNormalization (batch
// statistics, normalise + activation + residual, backward reductions and backward apply), PReLU / LeakyReLU backward,
// 2x2 max-pooling forward / backward (VGG19), the MSE / MAE content-loss reductions with their gradients, the tanh
// backward of the generator output, fp32 gathers for folded-layer weight gradients, and the fused flat Adam update.
//
// Reference semantics (Keras 2.2.x / TF 1.14 behind upscaling/upscaler/model.py): BatchNormalization(axis=-1, eps 1e-3,
// momentum 0.99) at model.py:20, 23, 284, 840-884 normalises with the biased batch variance in the training phase;
// PReLU(shared_axes=[1,2]) model.py:21, 276; LeakyReLU model.py:73, 841; VGG losses model.py:117, 137, 157;
// Adam() defaults model.py:1026, 1066, 1130.
//
// Layout: activations are NHWC 16-bit (bf16 for training) viewed as [rows = B*H*W][C], C a multiple of 8.  Every
// thread owns one group of 8 consecutive channels (one 16-byte load) and strides over rows, so per-channel parameters
// sit in registers and reductions over rows are private until a final shared-memory / atomic combine.  Grids are
// sized to a multiple of the SM count.
#include <cuda_fp16.h>

#include "../../include/b2srgan.h"
#include "b2_conv_epilogue.cuh"
#include "b2_host.h"

using namespace b2;

namespace {

constexpr int kEwThreads = 256;

struct RowGrid {
    int groups;        // C / 8
    int rows_per_blk;  // kEwThreads / groups
    int blocks;
};

// C/8 must divide kEwThreads (C in {8,...,2048} powers of two times ...): checked by the launchers.
inline bool row_grid(long rows, int C, RowGrid& g, int blocks_per_sm = 8) {
    if (C % 8 != 0) return false;
    g.groups = C / 8;
    if (g.groups > kEwThreads || kEwThreads % g.groups != 0) return false;
    g.rows_per_blk = kEwThreads / g.groups;
    long need = (rows + g.rows_per_blk - 1) / g.rows_per_blk;
    long cap = long(device_sm_count()) * blocks_per_sm;
    g.blocks = int(need < cap ? (need < 1 ? 1 : need) : cap);
    return true;
}

__device__ __forceinline__ void load8(const void* base, size_t elem, int f16, float (&v)[8]) {
    const uint4 u = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(base) + elem);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(&u);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        const float2 f = unpack2(w[e], f16);
        v[2 * e] = f.x;
        v[2 * e + 1] = f.y;
    }
}
__device__ __forceinline__ void store8(void* base, size_t elem, int f16, const float (&v)[8]) {
    uint4 u;
    uint32_t* w = reinterpret_cast<uint32_t*>(&u);
#pragma unroll
    for (int e = 0; e < 4; ++e) w[e] = pack2(v[2 * e], v[2 * e + 1], f16);
    *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(base) + elem) = u;
}
__device__ __forceinline__ void loadp8(const float* p, int c0, float (&v)[8], float dflt) {
    if (p) {
        const float4 a = *reinterpret_cast<const float4*>(p + c0), b = *reinterpret_cast<const float4*>(p + c0 + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = dflt;
    }
}

// Combine per-thread partial sums of NQ quantities x 8 channels.  Threads with the same channel group sit `groups`
// apart (groups is a power of two): lanes of one warp that share a group are folded with xor shuffles first, the
// per-warp partials go through shared memory, and each block issues two 16-byte vector reductions per quantity and
// channel group.  All blocks add into the same few cache lines and those reductions serialise in one L2 slice, so
// their count (blocks x C/4 per quantity), not the bytes, is what this step costs: grids are kept at 2 blocks per SM.
__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
template <int NQ>
__device__ __forceinline__ void block_combine(float (&acc)[NQ][8], int groups, int c0, float* const (&dst)[NQ]) {
    __shared__ float sh[NQ][kEwThreads][8];
    for (int off = 16; off >= groups; off >>= 1)
#pragma unroll
        for (int q = 0; q < NQ; ++q)
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[q][e] += __shfl_xor_sync(0xffffffffu, acc[q][e], off);
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        *reinterpret_cast<float4*>(&sh[q][threadIdx.x][0]) = make_float4(acc[q][0], acc[q][1], acc[q][2], acc[q][3]);
        *reinterpret_cast<float4*>(&sh[q][threadIdx.x][4]) = make_float4(acc[q][4], acc[q][5], acc[q][6], acc[q][7]);
    }
    __syncthreads();
    if (int(threadIdx.x) < groups) {
        const int step = groups < 32 ? 32 : groups;      // after the shuffles one partial per (warp, group) is left
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            if (!dst[q]) continue;
            float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            for (int t = threadIdx.x; t < kEwThreads; t += step) {
                const float4 lo = *reinterpret_cast<const float4*>(&sh[q][t][0]);
                const float4 hi = *reinterpret_cast<const float4*>(&sh[q][t][4]);
                s[0] += lo.x; s[1] += lo.y; s[2] += lo.z; s[3] += lo.w;
                s[4] += hi.x; s[5] += hi.y; s[6] += hi.z; s[7] += hi.w;
            }
            float* d = dst[q] + c0;
            if ((reinterpret_cast<uintptr_t>(d) & 15) == 0) {
                red_add_v4(d, s[0], s[1], s[2], s[3]);
                red_add_v4(d + 4, s[4], s[5], s[6], s[7]);
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) atomicAdd(d + e, s[e]);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ batch statistics
__global__ void __launch_bounds__(kEwThreads)
colstats_kernel(const void* __restrict__ x, long rows, int C, int f16, float* __restrict__ sum, float* __restrict__ sumsq) {
    b2::pdl_entry();
    const int groups = C >> 3, g = threadIdx.x % groups, c0 = g * 8;
    const int rpb = kEwThreads / groups;
    float acc[2][8] = {};
    const long stride = long(gridDim.x) * rpb;
    long r = long(blockIdx.x) * rpb + threadIdx.x / groups;
    for (; r + 3 * stride < rows; r += 4 * stride) {       // four independent 16-byte loads in flight per thread
        float v[4][8];
#pragma unroll
        for (int u = 0; u < 4; ++u) load8(x, size_t(r + u * stride) * C + c0, f16, v[u]);
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int e = 0; e < 8; ++e) { acc[0][e] += v[u][e]; acc[1][e] = fmaf(v[u][e], v[u][e], acc[1][e]); }
    }
    for (; r < rows; r += stride) {
        float v[8];
        load8(x, size_t(r) * C + c0, f16, v);
#pragma unroll
        for (int e = 0; e < 8; ++e) { acc[0][e] += v[e]; acc[1][e] = fmaf(v[e], v[e], acc[1][e]); }
    }
    float* const dst[2] = {sum, sumsq};
    block_combine<2>(acc, groups, c0, dst);
}

// mean / biased variance -> the affine (scale, shift) the apply kernels use, plus the Keras moving-statistics update.
__global__ void bn_finalize_kernel(const float* sum, const float* sumsq, float count, const float* gamma, const float* beta,
                                   float eps, float momentum, float* moving_mean, float* moving_var, float* scale,
                                   float* shift, float* mean_out, float* invstd_out, int C) {
    b2::pdl_entry();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float m = sum[c] / count;
    float var = sumsq[c] / count - m * m;
    var = fmaxf(var, 0.f);
    const float inv = rsqrtf(var + eps);
    const float sc = gamma[c] * inv;
    scale[c] = sc;
    shift[c] = beta[c] - m * sc;
    if (mean_out) mean_out[c] = m;
    if (invstd_out) invstd_out[c] = inv;
    if (moving_mean) {
        // keras/layers/normalization.py (2.2.x): sample variance = var * n / (n - (1 + eps))
        const float adj = var * (count / (count - (1.f + eps)));
        moving_mean[c] = moving_mean[c] * momentum + m * (1.f - momentum);
        moving_var[c] = moving_var[c] * momentum + adj * (1.f - momentum);
    }
}

// inference-mode fold: scale = gamma / sqrt(var + eps); shift = beta + (bias - mean) * scale
__global__ void bn_fold_kernel(const float* gamma, const float* beta, const float* mean, const float* var, const float* bias,
                               float eps, float* scale, float* shift, int C) {
    b2::pdl_entry();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float sc = gamma[c] * rsqrtf(var[c] + eps);
    scale[c] = sc;
    shift[c] = beta[c] + ((bias ? bias[c] : 0.f) - mean[c]) * sc;
}

__device__ __forceinline__ float act_fwd(float u, int act, float leaky, float alpha) {
    switch (act) {
        case ACT_RELU: return fmaxf(u, 0.f);
        case ACT_LEAKY: return u > 0.f ? u : u * leaky;
        case ACT_PRELU: return u > 0.f ? u : u * alpha;
        case ACT_TANH: return tanhf(u);
        default: return u;
    }
}
__device__ __forceinline__ float act_slope(float u, int act, float leaky, float alpha) {
    switch (act) {
        case ACT_RELU: return u > 0.f ? 1.f : 0.f;
        case ACT_LEAKY: return u > 0.f ? 1.f : leaky;
        case ACT_PRELU: return u > 0.f ? 1.f : alpha;
        default: return 1.f;
    }
}

// y = act(z * scale + shift) + residual
__global__ void __launch_bounds__(kEwThreads)
affine_act_kernel(const void* __restrict__ z, long rows, int C, int f16, const float* __restrict__ scale,
                  const float* __restrict__ shift, const float* __restrict__ alpha, int act, float leaky,
                  const void* __restrict__ residual, void* __restrict__ out) {
    b2::pdl_entry();
    const int groups = C >> 3, g = threadIdx.x % groups, c0 = g * 8;
    const int rpb = kEwThreads / groups;
    float sc[8], sh[8], al[8];
    loadp8(scale, c0, sc, 1.f);
    loadp8(shift, c0, sh, 0.f);
    loadp8(alpha, c0, al, 0.f);
    for (long r = long(blockIdx.x) * rpb + threadIdx.x / groups; r < rows; r += long(gridDim.x) * rpb) {
        float v[8];
        load8(z, size_t(r) * C + c0, f16, v);
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = act_fwd(fmaf(v[e], sc[e], sh[e]), act, leaky, al[e]);
        if (residual) {
            float rv[8];
            load8(residual, size_t(r) * C + c0, f16, rv);
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] += rv[e];
        }
        store8(out, size_t(r) * C + c0, f16, v);
    }
}

// Backward reductions of  y = act(u),  u = z * scale + shift  (BatchNorm output; scale = gamma * invstd):
//   dy = g * act'(u);  s1 = sum dy;  s2 = sum dy * xhat  (xhat = (z - mean) * invstd);  dalpha = sum g * min(u, 0)
__global__ void __launch_bounds__(kEwThreads)
bn_bwd_reduce_kernel(const void* __restrict__ z, const void* __restrict__ gin, long rows, int C, int f16,
                     const float* __restrict__ scale, const float* __restrict__ shift, const float* __restrict__ alpha,
                     int act, float leaky, const float* __restrict__ mean, const float* __restrict__ invstd,
                     float* __restrict__ s1, float* __restrict__ s2, float* __restrict__ dalpha) {
    b2::pdl_entry();
    const int groups = C >> 3, g = threadIdx.x % groups, c0 = g * 8;
    const int rpb = kEwThreads / groups;
    float sc[8], sh[8], al[8], mu[8], is[8];
    loadp8(scale, c0, sc, 1.f);
    loadp8(shift, c0, sh, 0.f);
    loadp8(alpha, c0, al, 0.f);
    loadp8(mean, c0, mu, 0.f);
    loadp8(invstd, c0, is, 1.f);
    float acc[3][8] = {};
    const long stride = long(gridDim.x) * rpb;
    long r = long(blockIdx.x) * rpb + threadIdx.x / groups;
    auto body = [&](const float (&zv)[8], const float (&gv)[8]) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const float u = fmaf(zv[e], sc[e], sh[e]);
            const float dy = gv[e] * act_slope(u, act, leaky, al[e]);
            acc[0][e] += dy;
            acc[1][e] = fmaf(dy, (zv[e] - mu[e]) * is[e], acc[1][e]);
            acc[2][e] = fmaf(gv[e], fminf(u, 0.f), acc[2][e]);
        }
    };
    for (; r + stride < rows; r += 2 * stride) {          // four independent 16-byte loads in flight per thread
        float z0[8], g0[8], z1[8], g1[8];
        load8(z, size_t(r) * C + c0, f16, z0);
        load8(gin, size_t(r) * C + c0, f16, g0);
        load8(z, size_t(r + stride) * C + c0, f16, z1);
        load8(gin, size_t(r + stride) * C + c0, f16, g1);
        body(z0, g0);
        body(z1, g1);
    }
    for (; r < rows; r += stride) {
        float zv[8], gv[8];
        load8(z, size_t(r) * C + c0, f16, zv);
        load8(gin, size_t(r) * C + c0, f16, gv);
        body(zv, gv);
    }
    float* const dst[3] = {s1, s2, dalpha};
    block_combine<3>(acc, groups, c0, dst);
}

// dz = scale * (dy - s1/n - xhat * s2/n)   (train-mode BatchNorm backward; with stats == null: dz = scale * dy)
__global__ void __launch_bounds__(kEwThreads)
bn_bwd_apply_kernel(const void* __restrict__ z, const void* __restrict__ gin, long rows, int C, int f16,
                    const float* __restrict__ scale, const float* __restrict__ shift, const float* __restrict__ alpha,
                    int act, float leaky, const float* __restrict__ mean, const float* __restrict__ invstd,
                    const float* __restrict__ s1, const float* __restrict__ s2, float inv_count, void* __restrict__ dz) {
    b2::pdl_entry();
    const int groups = C >> 3, g = threadIdx.x % groups, c0 = g * 8;
    const int rpb = kEwThreads / groups;
    float sc[8], sh[8], al[8], mu[8], is[8], m1[8], m2[8];
    loadp8(scale, c0, sc, 1.f);
    loadp8(shift, c0, sh, 0.f);
    loadp8(alpha, c0, al, 0.f);
    loadp8(mean, c0, mu, 0.f);
    loadp8(invstd, c0, is, 1.f);
    loadp8(s1, c0, m1, 0.f);
    loadp8(s2, c0, m2, 0.f);
#pragma unroll
    for (int e = 0; e < 8; ++e) { m1[e] *= inv_count; m2[e] *= inv_count; }
    const long stride = long(gridDim.x) * rpb;
    long r = long(blockIdx.x) * rpb + threadIdx.x / groups;
    auto body = [&](const float (&zv)[8], float (&gv)[8]) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const float u = fmaf(zv[e], sc[e], sh[e]);
            const float dy = gv[e] * act_slope(u, act, leaky, al[e]);
            const float xh = (zv[e] - mu[e]) * is[e];
            gv[e] = sc[e] * (dy - m1[e] - xh * m2[e]);
        }
    };
    for (; r + stride < rows; r += 2 * stride) {
        float z0[8], g0[8], z1[8], g1[8];
        load8(z, size_t(r) * C + c0, f16, z0);
        load8(gin, size_t(r) * C + c0, f16, g0);
        load8(z, size_t(r + stride) * C + c0, f16, z1);
        load8(gin, size_t(r + stride) * C + c0, f16, g1);
        body(z0, g0);
        body(z1, g1);
        store8(dz, size_t(r) * C + c0, f16, g0);
        store8(dz, size_t(r + stride) * C + c0, f16, g1);
    }
    for (; r < rows; r += stride) {
        float zv[8], gv[8];
        load8(z, size_t(r) * C + c0, f16, zv);
        load8(gin, size_t(r) * C + c0, f16, gv);
        body(zv, gv);
        store8(dz, size_t(r) * C + c0, f16, gv);
    }
}

// ------------------------------------------------------------------------------------------------ max pooling (VGG19)
// One CTA walks whole output rows (b, oy): no per-element 64-bit divisions, consecutive threads = consecutive 16-byte
// channel groups of consecutive pixels, four independent loads per thread.
__global__ void __launch_bounds__(kEwThreads)
maxpool_fwd_kernel(const void* __restrict__ x, int B, int H, int W, int C, int f16, void* __restrict__ y) {
    b2::pdl_entry();
    const int OH = H >> 1, OW = W >> 1, groups = C >> 3;
    const int per_row = OW * groups;
    for (int row = blockIdx.x; row < B * OH; row += gridDim.x) {
        const int b = row / OH, oy = row - b * OH;
        const size_t in_row = (size_t(b) * H + 2 * oy) * W * C;
        const size_t out_row = size_t(row) * OW * C;
        for (int i = threadIdx.x; i < per_row; i += blockDim.x) {
            const int ox = i / groups, g = i - ox * groups;
            const size_t base = in_row + size_t(2 * ox) * C + g * 8;
            float a[8], b1[8], c[8], d[8];
            load8(x, base, f16, a);
            load8(x, base + C, f16, b1);
            load8(x, base + size_t(W) * C, f16, c);
            load8(x, base + size_t(W) * C + C, f16, d);
#pragma unroll
            for (int e = 0; e < 8; ++e) a[e] = fmaxf(fmaxf(a[e], b1[e]), fmaxf(c[e], d[e]));
            store8(y, out_row + size_t(ox) * C + g * 8, f16, a);
        }
    }
}

// dx = dy routed to the first element of each 2x2 window that equals the window maximum (TensorFlow's MaxPoolGrad);
// rows / columns beyond the floor(H/2)*2 x floor(W/2)*2 region get zero.  Optional ReLU gate: dx *= (x > 0).
// Thread = (window, 8 channels): reads the four inputs and dy once, writes the four gradients; the window maximum is
// recomputed from the inputs (it equals the stored y bit for bit), so y is not read.
__global__ void __launch_bounds__(kEwThreads)
maxpool_bwd_kernel(const void* __restrict__ x, const void* __restrict__ y, const void* __restrict__ dy, int B, int H, int W,
                   int C, int f16, int relu_gate, void* __restrict__ dx) {
    b2::pdl_entry();
    const int OH = H >> 1, OW = W >> 1, groups = C >> 3;
    const int WH = (H + 1) >> 1, WW = (W + 1) >> 1;       // windows including the zero-filled odd edge
    const int per_row = WW * groups;
    for (int row = blockIdx.x; row < B * WH; row += gridDim.x) {
        const int b = row / WH, oy = row - b * WH;
        const size_t in_row = (size_t(b) * H + 2 * oy) * W * C;
        for (int i = threadIdx.x; i < per_row; i += blockDim.x) {
            const int ox = i / groups, g = i - ox * groups;
            const size_t base = in_row + size_t(2 * ox) * C + g * 8;
            const bool has_x1 = 2 * ox + 1 < W, has_y1 = 2 * oy + 1 < H;
            float o[4][8] = {};
            if (oy < OH && ox < OW) {
                float w[4][8], d[8];
                load8(x, base, f16, w[0]);
                load8(x, base + C, f16, w[1]);
                load8(x, base + size_t(W) * C, f16, w[2]);
                load8(x, base + size_t(W) * C + C, f16, w[3]);
                load8(dy, ((size_t(b) * OH + oy) * OW + ox) * C + g * 8, f16, d);
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const float m = fmaxf(fmaxf(w[0][e], w[1][e]), fmaxf(w[2][e], w[3][e]));
                    bool found = false;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const bool hit = !found && w[q][e] == m;
                        found = found || hit;
                        o[q][e] = (hit && (!relu_gate || w[q][e] > 0.f)) ? d[e] : 0.f;
                    }
                }
            }
            store8(dx, base, f16, o[0]);
            if (has_x1) store8(dx, base + C, f16, o[1]);
            if (has_y1) {
                store8(dx, base + size_t(W) * C, f16, o[2]);
                if (has_x1) store8(dx, base + size_t(W) * C + C, f16, o[3]);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ losses
// loss += weight * sum(f(a - b)) / n ;  grad = gscale * f'(a - b) / n   (f = square or abs), block-reduced, one atomic per block.
template <typename T>
__device__ __forceinline__ float ld_as_float(const T* p, size_t i);
template <> __device__ __forceinline__ float ld_as_float<float>(const float* p, size_t i) { return p[i]; }
template <> __device__ __forceinline__ float ld_as_float<__nv_bfloat16>(const __nv_bfloat16* p, size_t i) { return __bfloat162float(p[i]); }
template <> __device__ __forceinline__ float ld_as_float<__half>(const __half* p, size_t i) { return __half2float(p[i]); }
template <typename T>
__device__ __forceinline__ void st_from_float(T* p, size_t i, float v);
template <> __device__ __forceinline__ void st_from_float<float>(float* p, size_t i, float v) { p[i] = v; }
template <> __device__ __forceinline__ void st_from_float<__nv_bfloat16>(__nv_bfloat16* p, size_t i, float v) { p[i] = __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ void st_from_float<__half>(__half* p, size_t i, float v) { p[i] = __float2half_rn(v); }

template <typename T>
__global__ void __launch_bounds__(kEwThreads)
loss_pair_kernel(const T* __restrict__ a, const T* __restrict__ b, long n, int mae, float weight, float gscale,
                 float* __restrict__ loss, T* __restrict__ grad) {
    b2::pdl_entry();
    const float inv_n = 1.f / float(n);
    float acc = 0.f;
    for (long i = blockIdx.x * long(blockDim.x) + threadIdx.x; i < n; i += long(gridDim.x) * blockDim.x) {
        const float d = ld_as_float(a, i) - ld_as_float(b, i);
        acc += mae ? fabsf(d) : d * d;
        if (grad) {
            const float gr = mae ? (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f)) : 2.f * d;
            st_from_float(grad, i, gr * gscale * inv_n);
        }
    }
    __shared__ float sh[kEwThreads / 32];
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        float v = threadIdx.x < kEwThreads / 32 ? sh[threadIdx.x] : 0.f;
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (threadIdx.x == 0 && loss) atomicAdd(loss, v * weight * inv_n);
    }
}

// loss += weight * sum(x * y) / n  with y broadcast per row (Wasserstein: mean(y_true * y_pred), model.py:159-160);
// grad[i] = gscale * y[i] / n.  x, y, grad fp32; tiny.
__global__ void mean_prod_kernel(const float* x, const float* y, float yconst, long n, float weight, float gscale, float* loss,
                                 float* grad) {
    b2::pdl_entry();
    float acc = 0.f;
    for (long i = threadIdx.x; i < n; i += blockDim.x) {
        const float yv = y ? y[i] : yconst;
        acc += x[i] * yv;
        if (grad) grad[i] = gscale * yv / float(n);
    }
    __shared__ float sh[32];
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float v = 0.f;
        for (int w = 0; w < int(blockDim.x) / 32; ++w) v += sh[w];
        if (loss) atomicAdd(loss, v * weight / float(n));
    }
}

// Discriminator output activation (model.py:885-892, 950-957, 1003-1010) on the (n,1) scores and its backward.
// kind: 1 sigmoid, 2 log(sigmoid), 3 tanh, 4 bi-log = x/(1+|x|) * log(|x|+2).  dy == nullptr: out = f(z); else
// out = dy * f'(z).  fp32; n = batch size.
__global__ void score_act_kernel(const float* z, const float* dy, float* out, long n, int kind) {
    b2::pdl_entry();
    for (long i = blockIdx.x * long(blockDim.x) + threadIdx.x; i < n; i += long(gridDim.x) * blockDim.x) {
        const float x = z[i];
        float f, d;
        if (kind == 1) {
            f = 1.f / (1.f + expf(-x));
            d = f * (1.f - f);
        } else if (kind == 2) {
            f = fminf(x, 0.f) - log1pf(expf(-fabsf(x)));
            d = 1.f / (1.f + expf(x));
        } else if (kind == 3) {
            f = tanhf(x);
            d = 1.f - f * f;
        } else {
            const float a = fabsf(x), r = x / (1.f + a), l = logf(a + 2.f);
            f = r * l;
            d = l / ((1.f + a) * (1.f + a)) + a / ((1.f + a) * (a + 2.f));
        }
        out[i] = dy ? dy[i] * d : f;
    }
}

// dz = (g1 + g2 + g3) * (1 - y^2): backward of the generator's final tanh (model.py:291) with up to three incoming
// gradients (pixel loss, VGG image gradient, discriminator image gradient); all fp32.
__global__ void tanh_bwd_kernel(const float* __restrict__ y, const float* __restrict__ g1, const float* __restrict__ g2,
                                const float* __restrict__ g3, long n, float* __restrict__ dz) {
    b2::pdl_entry();
    for (long i = blockIdx.x * long(blockDim.x) + threadIdx.x; i < n; i += long(gridDim.x) * blockDim.x) {
        float g = g1[i];
        if (g2) g += g2[i];
        if (g3) g += g3[i];
        const float yv = y[i];
        dz[i] = g * (1.f - yv * yv);
    }
}

__global__ void __launch_bounds__(kEwThreads)
colsum_f32_kernel(const float* __restrict__ x, long rows, int C, float* __restrict__ sum) {
    b2::pdl_entry();
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (long r = blockIdx.x * long(blockDim.x) + threadIdx.x; r < rows; r += long(gridDim.x) * blockDim.x)
        for (int c = 0; c < C; ++c) acc[c] += x[r * C + c];
    __shared__ float sh[8][kEwThreads / 32];
    for (int c = 0; c < C; ++c) {
        float v = acc[c];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) sh[c][threadIdx.x >> 5] = v;
    }
    __syncthreads();
    if (int(threadIdx.x) < C) {
        float v = 0.f;
        for (int w = 0; w < kEwThreads / 32; ++w) v += sh[threadIdx.x][w];
        atomicAdd(sum + threadIdx.x, v);
    }
}

__device__ __forceinline__ float ld_any(const void* p, size_t i, int dtype) {
    if (dtype == 2) return reinterpret_cast<const float*>(p)[i];
    if (dtype == 1) return __half2float(reinterpret_cast<const __half*>(p)[i]);
    return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]);
}
__device__ __forceinline__ void st_any(void* p, size_t i, int dtype, float v) {
    if (dtype == 2) reinterpret_cast<float*>(p)[i] = v;
    else if (dtype == 1) reinterpret_cast<__half*>(p)[i] = __float2half_rn(v);
    else reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16_rn(v);
}

// Dense head of the discriminator: batch-sized M, so plain CUDA-core loops over the fp32 master weights.
// Forward: a block owns 32 output columns; its 8 warps stride over K (each warp-load of W is one 128-byte row piece), every
// thread keeps up to 8 batch rows in registers so W is read once per 8 rows, partial sums meet in shared memory.  (The
// thread-per-output form needed K = 2048 dependent L2 round trips: 380 us for 12 x 2048 x 1024.)
__global__ void __launch_bounds__(256)
dense_fwd_kernel(const void* x, int xd, int n, int K, int N, const float* __restrict__ W, const float* b, void* out, int od) {
    b2::pdl_entry();
    __shared__ float red[8][8][33];
    const int col = threadIdx.x & 31, ks = threadIdx.x >> 5;
    const int o = blockIdx.x * 32 + col;
    for (int r0 = 0; r0 < n; r0 += 8) {
        const int nr = min(8, n - r0);
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (o < N) {
#pragma unroll 4
            for (int k = ks; k < K; k += 8) {
                const float w = W[size_t(k) * N + o];
#pragma unroll
                for (int r = 0; r < 8; ++r)
                    if (r < nr) acc[r] = fmaf(ld_any(x, size_t(r0 + r) * K + k, xd), w, acc[r]);
            }
        }
#pragma unroll
        for (int r = 0; r < 8; ++r) red[ks][r][col] = acc[r];
        __syncthreads();
        const int r = ks;                     // thread (r, col) finishes row r0 + r
        if (o < N && r < nr) {
            float sum = b ? b[o] : 0.f;
#pragma unroll
            for (int q = 0; q < 8; ++q) sum += red[q][r][col];
            st_any(out, size_t(r0 + r) * N + o, od, sum);
        }
        __syncthreads();
    }
}
__global__ void dense_bwd_data_kernel(const void* dy, int yd, int n, int K, int N, const float* __restrict__ W, void* dx, int xd) {
    b2::pdl_entry();
    const long total = long(n) * K;
    for (long i = blockIdx.x * long(blockDim.x) + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
        const int k = int(i % K), r = int(i / K);
        float acc = 0.f;
        for (int o = 0; o < N; ++o) acc = fmaf(ld_any(dy, size_t(r) * N + o, yd), W[size_t(k) * N + o], acc);
        st_any(dx, i, xd, acc);
    }
}
__global__ void dense_bwd_weight_kernel(const void* x, int xd, const void* dy, int yd, int n, int K, int N, float* dW, float* db) {
    b2::pdl_entry();
    const long total = long(K) * N;
    for (long i = blockIdx.x * long(blockDim.x) + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
        const int o = int(i % N), k = int(i / N);
        float acc = 0.f;
        for (int r = 0; r < n; ++r) acc = fmaf(ld_any(x, size_t(r) * K + k, xd), ld_any(dy, size_t(r) * N + o, yd), acc);
        dW[i] += acc;
        if (k == 0 && db) {
            float s = 0.f;
            for (int r = 0; r < n; ++r) s += ld_any(dy, size_t(r) * N + o, yd);
            db[o] += s;
        }
    }
}

// out[i] (+)= idx[i] >= 0 ? src[idx[i]] : 0   (fp32; folded-layer weight gradients back to the Keras layout)
__global__ void gather_f32_kernel(const float* __restrict__ src, const int32_t* __restrict__ idx, long n, float* __restrict__ out,
                                  int accumulate) {
    b2::pdl_entry();
    for (long i = blockIdx.x * long(blockDim.x) + threadIdx.x; i < n; i += long(gridDim.x) * blockDim.x) {
        const int32_t j = idx[i];
        const float v = j >= 0 ? src[j] : 0.f;
        out[i] = accumulate ? out[i] + v : v;
    }
}

// 16-bit <-> fp32 casts with a scale (gradient hand-off between fp32 image tensors and 16-bit feature tensors)
template <typename TI, typename TO>
__global__ void cast_scale_kernel(const TI* __restrict__ in, long n, float scale, TO* __restrict__ out) {
    b2::pdl_entry();
    for (long i = blockIdx.x * long(blockDim.x) + threadIdx.x; i < n; i += long(gridDim.x) * blockDim.x)
        st_from_float(out, i, ld_as_float(in, i) * scale);
}

// Keras Adam: m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2; p -= lr_t * m / (sqrt(v) + eps), lr_t precomputed on the host.
// grads are multiplied by gscale first (1 / world_size after a sum all-reduce).
__global__ void __launch_bounds__(kEwThreads)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, long n, float lr_t,
            float b1, float b2, float eps, float gscale, const unsigned* __restrict__ guard0, const unsigned* __restrict__ guard1,
            int guard_words) {
    b2::pdl_entry();
    // guarded form: any non-zero guard word (range flags / saturation counts raised by the passes that produced g) turns the
    // whole launch into a no-op -- parameters and both moments keep their bits, the host sees the words with the step's loss
    for (int i = 0; i < guard_words; ++i)
        if ((guard0 && guard0[i] != 0u) || (guard1 && guard1[i] != 0u)) return;
    const long n4 = n >> 2;
    for (long i = blockIdx.x * long(blockDim.x) + threadIdx.x; i < n4; i += long(gridDim.x) * blockDim.x) {
        float4 pp = reinterpret_cast<float4*>(p)[i], gg = reinterpret_cast<const float4*>(g)[i];
        float4 mm = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
        float* pa = &pp.x; float* ga = &gg.x; float* ma = &mm.x; float* va = &vv.x;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float gr = ga[e] * gscale;
            ma[e] = b1 * ma[e] + (1.f - b1) * gr;
            va[e] = b2 * va[e] + (1.f - b2) * gr * gr;
            pa[e] -= lr_t * ma[e] / (sqrtf(va[e]) + eps);
        }
        reinterpret_cast<float4*>(p)[i] = pp;
        reinterpret_cast<float4*>(m)[i] = mm;
        reinterpret_cast<float4*>(v)[i] = vv;
    }
    for (long i = (n4 << 2) + blockIdx.x * long(blockDim.x) + threadIdx.x; i < n; i += long(gridDim.x) * blockDim.x) {
        const float gr = g[i] * gscale;
        m[i] = b1 * m[i] + (1.f - b1) * gr;
        v[i] = b2 * v[i] + (1.f - b2) * gr * gr;
        p[i] -= lr_t * m[i] / (sqrtf(v[i]) + eps);
    }
}

inline int flat_grid(long work, int per_sm = 8) {
    long blocks = (work + kEwThreads - 1) / kEwThreads;
    const long cap = long(device_sm_count()) * per_sm;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return int(blocks);
}

inline int launch_ok(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("%s launch: %s", what, cudaGetErrorString(e));
    return 0;
}

}  // namespace

#define B2_STREAM reinterpret_cast<cudaStream_t>(stream)

extern "C" {

int b2_colstats(const void* x, int64_t rows, int C, int f16, float* sum, float* sumsq, void* stream) {
    if (!x || !sum) return set_error("b2_colstats: null pointer");
    RowGrid g;
    if (!row_grid(rows, C, g, 2)) return set_error("b2_colstats: C=%d unsupported (C/8 must divide %d)", C, kEwThreads);
    b2::launch_k(colstats_kernel, g.blocks, kEwThreads, 0, B2_STREAM, x, rows, C, f16, sum, sumsq);
    return launch_ok("b2_colstats");
}

int b2_bn_finalize(const float* sum, const float* sumsq, float count, const float* gamma, const float* beta, float eps,
                   float momentum, float* moving_mean, float* moving_var, float* scale, float* shift, float* mean_out,
                   float* invstd_out, int C, void* stream) {
    if (!sum || !sumsq || !gamma || !beta || !scale || !shift) return set_error("b2_bn_finalize: null pointer");
    b2::launch_k(bn_finalize_kernel, (C + 127) / 128, 128, 0, B2_STREAM, sum, sumsq, count, gamma, beta, eps, momentum, moving_mean,
                                                                moving_var, scale, shift, mean_out, invstd_out, C);
    return launch_ok("b2_bn_finalize");
}

int b2_bn_fold(const float* gamma, const float* beta, const float* mean, const float* var, const float* bias, float eps,
               float* scale, float* shift, int C, void* stream) {
    if (!gamma || !beta || !mean || !var || !scale || !shift) return set_error("b2_bn_fold: null pointer");
    b2::launch_k(bn_fold_kernel, (C + 127) / 128, 128, 0, B2_STREAM, gamma, beta, mean, var, bias, eps, scale, shift, C);
    return launch_ok("b2_bn_fold");
}

int b2_affine_act(const void* z, int64_t rows, int C, int f16, const float* scale, const float* shift, const float* alpha,
                  int act, float leaky, const void* residual, void* out, void* stream) {
    if (!z || !out) return set_error("b2_affine_act: null pointer");
    RowGrid g;
    if (!row_grid(rows, C, g)) return set_error("b2_affine_act: C=%d unsupported", C);
    b2::launch_k(affine_act_kernel, g.blocks, kEwThreads, 0, B2_STREAM, z, rows, C, f16, scale, shift, alpha, act, leaky, residual, out);
    return launch_ok("b2_affine_act");
}

int b2_bn_bwd_reduce(const void* z, const void* g, int64_t rows, int C, int f16, const float* scale, const float* shift,
                     const float* alpha, int act, float leaky, const float* mean, const float* invstd, float* s1, float* s2,
                     float* dalpha, void* stream) {
    if (!z || !g || !s1) return set_error("b2_bn_bwd_reduce: null pointer");
    RowGrid rg;
    if (!row_grid(rows, C, rg, 2)) return set_error("b2_bn_bwd_reduce: C=%d unsupported", C);
    b2::launch_k(bn_bwd_reduce_kernel, rg.blocks, kEwThreads, 0, B2_STREAM, z, g, rows, C, f16, scale, shift, alpha, act, leaky, mean,
                                                                   invstd, s1, s2, dalpha);
    return launch_ok("b2_bn_bwd_reduce");
}

int b2_bn_bwd_apply(const void* z, const void* g, int64_t rows, int C, int f16, const float* scale, const float* shift,
                    const float* alpha, int act, float leaky, const float* mean, const float* invstd, const float* s1,
                    const float* s2, float count, void* dz, void* stream) {
    if (!z || !g || !dz) return set_error("b2_bn_bwd_apply: null pointer");
    RowGrid rg;
    if (!row_grid(rows, C, rg)) return set_error("b2_bn_bwd_apply: C=%d unsupported", C);
    b2::launch_k(bn_bwd_apply_kernel, rg.blocks, kEwThreads, 0, B2_STREAM, z, g, rows, C, f16, scale, shift, alpha, act, leaky, mean,
                                                                  invstd, s1, s2, count > 0.f ? 1.f / count : 0.f, dz);
    return launch_ok("b2_bn_bwd_apply");
}

int b2_maxpool2x2(const void* x, int B, int H, int W, int C, int f16, void* y, void* stream) {
    if (!x || !y) return set_error("b2_maxpool2x2: null pointer");
    if (C % 8 != 0 || H < 2 || W < 2) return set_error("b2_maxpool2x2: bad geometry %dx%dx%d", H, W, C);
    b2::launch_k(maxpool_fwd_kernel, flat_grid(long(B) * (H / 2) * kEwThreads), kEwThreads, 0, B2_STREAM, x, B, H, W, C, f16, y);
    return launch_ok("b2_maxpool2x2");
}

int b2_maxpool2x2_bwd(const void* x, const void* y, const void* dy, int B, int H, int W, int C, int f16, int relu_gate,
                      void* dx, void* stream) {
    if (!x || !y || !dy || !dx) return set_error("b2_maxpool2x2_bwd: null pointer");
    if (C % 8 != 0 || H < 2 || W < 2) return set_error("b2_maxpool2x2_bwd: bad geometry %dx%dx%d", H, W, C);
    b2::launch_k(maxpool_bwd_kernel, flat_grid(long(B) * ((H + 1) / 2) * kEwThreads), kEwThreads, 0, B2_STREAM, x, y, dy, B, H, W, C, f16, relu_gate, dx);
    return launch_ok("b2_maxpool2x2_bwd");
}

int b2_loss_pair(const void* a, const void* b, int64_t n, int dtype, int mae, float weight, float gscale, float* loss,
                 void* grad, void* stream) {
    if (!a || !b) return set_error("b2_loss_pair: null pointer");
    const int grid = flat_grid(n, 4);
    if (dtype == 2)
        b2::launch_k(loss_pair_kernel<float>, grid, kEwThreads, 0, B2_STREAM, reinterpret_cast<const float*>(a), reinterpret_cast<const float*>(b), n, mae, weight, gscale, loss, reinterpret_cast<float*>(grad));
    else if (dtype == 1)
        b2::launch_k(loss_pair_kernel<__half>, grid, kEwThreads, 0, B2_STREAM, reinterpret_cast<const __half*>(a), reinterpret_cast<const __half*>(b), n, mae, weight, gscale, loss, reinterpret_cast<__half*>(grad));
    else
        b2::launch_k(loss_pair_kernel<__nv_bfloat16>, grid, kEwThreads, 0, B2_STREAM, reinterpret_cast<const __nv_bfloat16*>(a), reinterpret_cast<const __nv_bfloat16*>(b), n, mae, weight, gscale, loss, reinterpret_cast<__nv_bfloat16*>(grad));
    return launch_ok("b2_loss_pair");
}

int b2_score_act(const float* z, const float* dy, float* out, int64_t n, int kind, void* stream) {
    if (!z || !out) return set_error("b2_score_act: null pointer");
    if (kind < 1 || kind > 4) return set_error("b2_score_act: kind must be 1 (sigmoid), 2 (log-sigm), 3 (tanh) or 4 (bi-log)");
    if (n == 0) return 0;
    b2::launch_k(score_act_kernel, int((n + 255) / 256 < 148 ? (n + 255) / 256 : 148), 256, 0, B2_STREAM, z, dy, out, n, kind);
    return launch_ok("b2_score_act");
}

int b2_mean_prod(const float* x, const float* y, float yconst, int64_t n, float weight, float gscale, float* loss, float* grad,
                 void* stream) {
    if (!x) return set_error("b2_mean_prod: null pointer");
    b2::launch_k(mean_prod_kernel, 1, 256, 0, B2_STREAM, x, y, yconst, n, weight, gscale, loss, grad);
    return launch_ok("b2_mean_prod");
}

int b2_tanh_bwd(const float* y, const float* g1, const float* g2, const float* g3, int64_t n, float* dz, void* stream) {
    if (!y || !g1 || !dz) return set_error("b2_tanh_bwd: null pointer");
    b2::launch_k(tanh_bwd_kernel, flat_grid(n), kEwThreads, 0, B2_STREAM, y, g1, g2, g3, n, dz);
    return launch_ok("b2_tanh_bwd");
}

int b2_colsum_f32(const float* x, int64_t rows, int C, float* sum, void* stream) {
    if (!x || !sum) return set_error("b2_colsum_f32: null pointer");
    if (C < 1 || C > 8) return set_error("b2_colsum_f32: C=%d unsupported", C);
    b2::launch_k(colsum_f32_kernel, flat_grid(rows, 2), kEwThreads, 0, B2_STREAM, x, rows, C, sum);
    return launch_ok("b2_colsum_f32");
}

int b2_dense_fwd(const void* x, int x_dtype, int n, int K, int N, const float* W, const float* b, void* out, int out_dtype,
                 void* stream) {
    if (!x || !W || !out) return set_error("b2_dense_fwd: null pointer");
    b2::launch_k(dense_fwd_kernel, (N + 31) / 32, 256, 0, B2_STREAM, x, x_dtype, n, K, N, W, b, out, out_dtype);
    return launch_ok("b2_dense_fwd");
}
int b2_dense_bwd_data(const void* dy, int dy_dtype, int n, int K, int N, const float* W, void* dx, int dx_dtype, void* stream) {
    if (!dy || !W || !dx) return set_error("b2_dense_bwd_data: null pointer");
    b2::launch_k(dense_bwd_data_kernel, flat_grid(long(n) * K), kEwThreads, 0, B2_STREAM, dy, dy_dtype, n, K, N, W, dx, dx_dtype);
    return launch_ok("b2_dense_bwd_data");
}
int b2_dense_bwd_weight(const void* x, int x_dtype, const void* dy, int dy_dtype, int n, int K, int N, float* dW, float* db,
                        void* stream) {
    if (!x || !dy || !dW) return set_error("b2_dense_bwd_weight: null pointer");
    b2::launch_k(dense_bwd_weight_kernel, flat_grid(long(K) * N), kEwThreads, 0, B2_STREAM, x, x_dtype, dy, dy_dtype, n, K, N, dW, db);
    return launch_ok("b2_dense_bwd_weight");
}

int b2_gather_f32(const float* src, const int32_t* idx, int64_t n, float* out, int accumulate, void* stream) {
    if (!src || !idx || !out) return set_error("b2_gather_f32: null pointer");
    b2::launch_k(gather_f32_kernel, flat_grid(n), kEwThreads, 0, B2_STREAM, src, idx, n, out, accumulate);
    return launch_ok("b2_gather_f32");
}

int b2_cast_scale(const void* in, int in_dtype, int64_t n, float scale, void* out, int out_dtype, void* stream) {
    if (!in || !out) return set_error("b2_cast_scale: null pointer");
    const int grid = flat_grid(n);
    cudaStream_t s = B2_STREAM;
#define B2_CAST(TI, TO) b2::launch_k(cast_scale_kernel<TI, TO>, grid, kEwThreads, 0, s, reinterpret_cast<const TI*>(in), n, scale, reinterpret_cast<TO*>(out))
    if (in_dtype == 2 && out_dtype == 0) B2_CAST(float, __nv_bfloat16);
    else if (in_dtype == 2 && out_dtype == 1) B2_CAST(float, __half);
    else if (in_dtype == 0 && out_dtype == 2) B2_CAST(__nv_bfloat16, float);
    else if (in_dtype == 1 && out_dtype == 2) B2_CAST(__half, float);
    else if (in_dtype == 2 && out_dtype == 2) B2_CAST(float, float);
    else return set_error("b2_cast_scale: unsupported dtype pair %d -> %d", in_dtype, out_dtype);
#undef B2_CAST
    return launch_ok("b2_cast_scale");
}

int b2_adam(float* params, const float* grads, float* m, float* v, int64_t n, float lr_t, float beta1, float beta2, float eps,
            float gscale, void* stream) {
    if (!params || !grads || !m || !v) return set_error("b2_adam: null pointer");
    b2::launch_k(adam_kernel, flat_grid((n + 3) / 4), kEwThreads, 0, B2_STREAM, params, grads, m, v, n, lr_t, beta1, beta2, eps, gscale,
                 static_cast<const unsigned*>(nullptr), static_cast<const unsigned*>(nullptr), 0);
    return launch_ok("b2_adam");
}

int b2_adam_guarded(float* params, const float* grads, float* m, float* v, int64_t n, float lr_t, float beta1, float beta2, float eps,
                    float gscale, const void* guard0, const void* guard1, int guard_words, void* stream) {
    if (!params || !grads || !m || !v) return set_error("b2_adam_guarded: null pointer");
    if (guard_words < 0 || guard_words > 16) return set_error("b2_adam_guarded: guard_words=%d (0..16)", guard_words);
    b2::launch_k(adam_kernel, flat_grid((n + 3) / 4), kEwThreads, 0, B2_STREAM, params, grads, m, v, n, lr_t, beta1, beta2, eps, gscale,
                 static_cast<const unsigned*>(guard0), static_cast<const unsigned*>(guard1), guard_words);
    return launch_ok("b2_adam_guarded");
}

}  // extern "C"
