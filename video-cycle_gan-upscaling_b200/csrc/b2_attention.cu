// b2_attention.cu — the memory-bound layers make_upscaler_attention adds to the SRGAN generator (reference
// upscaling/upscaler/model.py:30-48 residual_block_attention, :78-98 upsampling_block_attention):
//   * Multiply([sigmoid(attention conv), stream])                     -> gate_mul_kernel (one 16-byte load of each operand,
//                                                                         one 16-byte store per thread; HBM-bound)
//   * Concatenate([nearest, bilinear] resize of the network input)   -> resize_pair_kernel (fp32, 2*C channels per pixel)
//   * Conv2DTranspose(F, scale + 1, strides = scale)(atanh(0.99999 x)) on the 3-channel network input
//                                                                     -> atanh_convT_kernel: every output pixel has at most
//       2 x 2 contributing input pixels (kernel = stride + 1), 3 channels each: 12 MACs per output value, so the layer is
//       bound by its F x 16-bit output write, not by arithmetic; weights (<= 38 KB fp32) live in shared memory.
// All three are grid-stride kernels with 16-byte coalesced stores, grids sized in multiples of the SM count.
#include "../../include/b2srgan.h"
#include <cuda_fp16.h>

#include "b2_host.h"

using namespace b2;

namespace {

__device__ __forceinline__ float2 unpack16(uint32_t u, int f16) {
    if (f16) return __half22float2(*reinterpret_cast<const __half2*>(&u));
    return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u));
}
__device__ __forceinline__ uint32_t pack16(float a, float b, int f16) {
    if (f16) {
        uint32_t r;
        asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
        return r;
    }
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
}

inline int grid_for(long work, int threads) {
    long blocks = (work + threads - 1) / threads;
    const long cap = long(device_sm_count()) * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return int(blocks);
}

// out[r, c] = sigmoid(att[r * att_ld + att_c0 + c]) * m[r * C + c]; 8 channels per thread.
__global__ void __launch_bounds__(256)
gate_mul_kernel(const uint4* __restrict__ att, int att_ld8, int att_c08, const uint4* __restrict__ m, uint4* __restrict__ out,
                long rows, int C8, int f16) {
    b2::pdl_entry();
    const long total = rows * C8;
    for (long i = blockIdx.x * long(blockDim.x) + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
        const long r = i / C8;
        const int g = int(i - r * C8);
        const uint4 a = att[r * att_ld8 + att_c08 + g], v = m[i];
        const uint32_t au[4] = {a.x, a.y, a.z, a.w}, vu[4] = {v.x, v.y, v.z, v.w};
        uint32_t o[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float2 fa = unpack16(au[e], f16), fv = unpack16(vu[e], f16);
            o[e] = pack16(fv.x / (1.f + __expf(-fa.x)), fv.y / (1.f + __expf(-fa.y)), f16);
        }
        out[i] = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

// out[b, oy, ox, 0..C) = in[b, oy / f, ox / f, :] (tf.image.resize_nearest_neighbor, align_corners = False);
// out[b, oy, ox, C..2C) = TF 1.x resize_bilinear (source = out / f, no half-pixel centres), fp32 throughout.
__global__ void __launch_bounds__(256)
resize_pair_kernel(const float* __restrict__ in, int B, int H, int W, int C, int factor, float* __restrict__ out) {
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
        const float* rn = in + (size_t(b) * H + oy / factor) * W * C;
        float* dst = out + size_t(row) * OW * 2 * C;
        for (int ox = threadIdx.x; ox < OW; ox += blockDim.x) {
            const float sx = float(ox) * inv;
            const int x0 = int(floorf(sx)), x1 = min(x0 + 1, W - 1);
            const float fx = sx - float(x0);
            for (int c = 0; c < C; ++c) {
                const float tl = r0[x0 * C + c], tr = r0[x1 * C + c], bl = r1[x0 * C + c], br = r1[x1 * C + c];
                const float top = tl + (tr - tl) * fx, bot = bl + (br - bl) * fx;
                dst[size_t(ox) * 2 * C + c] = rn[(ox / factor) * C + c];
                dst[size_t(ox) * 2 * C + C + c] = top + (bot - top) * fy;
            }
        }
    }
}

// y[b, oy, ox, co] = bias[co] + sum over the (<= 2 x 2) input pixels (iy, ix) with taps ty = oy - iy*s, tx = ox - ix*s in
// [0, s] and the Cin channels of  atanh(0.99999 * in[b, iy, ix, c]) * w[ty, tx, co, c]   (Keras Conv2DTranspose kernel layout
// (kh, kw, Cout, Cin), padding 'same': pad_total = k - s = 1, pad_before = 0).  Thread = (output pixel, 8 channels).
// Shared memory: weights re-laid as [ty][tx][c][co] fp32 so the 8 channels of a thread are two float4 loads and the 16
// threads of a pixel read 512 contiguous bytes.
template <int CIN>
__global__ void __launch_bounds__(256)
atanh_convT_kernel(const float* __restrict__ in, int B, int H, int W, const float* __restrict__ w, const float* __restrict__ bias,
                   int s, int Cout, void* __restrict__ out, int f16) {
    b2::pdl_entry();
    extern __shared__ float s_w[];
    const int k = s + 1;
    float* s_b = s_w + k * k * CIN * Cout;
    for (int i = threadIdx.x; i < k * k * CIN * Cout; i += blockDim.x) {
        const int co = i % Cout, c = (i / Cout) % CIN, t = i / (Cout * CIN);      // destination [t][c][co]
        s_w[i] = w[(size_t(t) * Cout + co) * CIN + c];                                // source [t][co][c]
    }
    for (int i = threadIdx.x; i < Cout; i += blockDim.x) s_b[i] = bias ? bias[i] : 0.f;
    __syncthreads();
    const int OH = H * s, OW = W * s, G = Cout / 8;
    const long total = long(B) * OH * OW * G;
    for (long i = blockIdx.x * long(blockDim.x) + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
        const int g = int(i % G);
        long pix = i / G;
        const int ox = int(pix % OW); pix /= OW;
        const int oy = int(pix % OH);
        const int b = int(pix / OH);
        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = s_b[g * 8 + e];
        const int iy0 = oy / s, ty0 = oy - iy0 * s, ix0 = ox / s, tx0 = ox - ix0 * s;
#pragma unroll
        for (int dy = 0; dy < 2; ++dy) {
            const int iy = iy0 - dy, ty = ty0 + dy * s;
            if (iy < 0 || ty > s) continue;               // second row only when oy is a multiple of s (tap index s)
#pragma unroll
            for (int dx = 0; dx < 2; ++dx) {
                const int ix = ix0 - dx, tx = tx0 + dx * s;
                if (ix < 0 || tx > s) continue;
                const float* px = in + ((size_t(b) * H + iy) * W + ix) * CIN;
                const float* wt = s_w + size_t(ty * k + tx) * CIN * Cout + g * 8;
#pragma unroll
                for (int c = 0; c < CIN; ++c) {
                    const float a = atanhf(0.99999f * px[c]);
                    const float4 w0 = *reinterpret_cast<const float4*>(wt + c * Cout);
                    const float4 w1 = *reinterpret_cast<const float4*>(wt + c * Cout + 4);
                    acc[0] = fmaf(a, w0.x, acc[0]); acc[1] = fmaf(a, w0.y, acc[1]);
                    acc[2] = fmaf(a, w0.z, acc[2]); acc[3] = fmaf(a, w0.w, acc[3]);
                    acc[4] = fmaf(a, w1.x, acc[4]); acc[5] = fmaf(a, w1.y, acc[5]);
                    acc[6] = fmaf(a, w1.z, acc[6]); acc[7] = fmaf(a, w1.w, acc[7]);
                }
            }
        }
        reinterpret_cast<uint4*>(out)[i] = make_uint4(pack16(acc[0], acc[1], f16), pack16(acc[2], acc[3], f16),
                                                      pack16(acc[4], acc[5], f16), pack16(acc[6], acc[7], f16));
    }
}

}  // namespace

extern "C" {

int b2_gate_mul(const void* att, int att_ld, int att_c0, const void* m, void* out, int64_t rows, int C, int f16, void* stream) {
    if (!att || !m || !out) return set_error("b2_gate_mul: null pointer");
    if (rows < 1 || C < 8 || (C & 7) || (att_ld & 7) || (att_c0 & 7) || att_c0 < 0 || att_c0 + C > att_ld)
        return set_error("b2_gate_mul: channels [%d, %d) of %d must be multiples of 8", att_c0, att_c0 + C, att_ld);
    if ((reinterpret_cast<uintptr_t>(att) | reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(out)) & 15)
        return set_error("b2_gate_mul: buffers must be 16-byte aligned");
    b2::launch_k(gate_mul_kernel, grid_for(rows * (C / 8), 256), 256, 0, reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<const uint4*>(att), att_ld / 8, att_c0 / 8, reinterpret_cast<const uint4*>(m), reinterpret_cast<uint4*>(out),
        long(rows), C / 8, f16);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("b2_gate_mul launch: %s", cudaGetErrorString(e));
    return 0;
}

int b2_resize_nearest_bilinear_pair(const float* in, int B, int H, int W, int C, int factor, float* out, void* stream) {
    if (!in || !out) return set_error("b2_resize_nearest_bilinear_pair: null pointer");
    if (B < 1 || H < 1 || W < 1 || C < 1 || factor < 1) return set_error("b2_resize_nearest_bilinear_pair: bad geometry");
    b2::launch_k(resize_pair_kernel, grid_for(long(B) * H * factor * 256, 256), 256, 0, reinterpret_cast<cudaStream_t>(stream), in, B, H, W, C, factor, out);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("b2_resize_nearest_bilinear_pair launch: %s", cudaGetErrorString(e));
    return 0;
}

int b2_atanh_conv2d_transpose_small(const float* in, int B, int H, int W, int Cin, const float* w, const float* bias, int stride,
                                    int Cout, void* out, int f16, void* stream) {
    if (!in || !w || !out) return set_error("b2_atanh_conv2d_transpose_small: null pointer");
    if (Cin != 3) return set_error("b2_atanh_conv2d_transpose_small: built for the 3-channel network input, got Cin=%d", Cin);
    if (stride < 1 || stride > 8 || Cout < 8 || (Cout & 7)) return set_error("b2_atanh_conv2d_transpose_small: stride %d / Cout %d unsupported", stride, Cout);
    const int k = stride + 1;
    const size_t smem = (size_t(k) * k * Cin * Cout + Cout) * sizeof(float);
    if (smem > 200 * 1024) return set_error("b2_atanh_conv2d_transpose_small: weights (%zu bytes) do not fit in shared memory", smem);
    static size_t attr = 0;
    if (smem > attr) {
        cudaError_t e = cudaFuncSetAttribute(atanh_convT_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(atanh_convT): %s", cudaGetErrorString(e));
        attr = smem;
    }
    const long work = long(B) * H * stride * W * stride * (Cout / 8);
    long blocks = (work + 255) / 256;
    const long cap = long(device_sm_count()) * 4;         // each block re-lays the weights once: keep the grid persistent
    if (blocks > cap) blocks = cap;
    b2::launch_k(atanh_convT_kernel<3>, int(blocks), 256, smem, reinterpret_cast<cudaStream_t>(stream), in, B, H, W, w, bias, stride, Cout, out, f16);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("b2_atanh_conv2d_transpose_small launch: %s", cudaGetErrorString(e));
    return 0;
}

}  // extern "C"
