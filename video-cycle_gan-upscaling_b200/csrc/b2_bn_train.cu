// b2_bn_train.cu — train-mode BatchNormalization statistics in a deterministic two-stage form (Keras
// BatchNormalization(axis=-1, eps 1e-3, momentum 0.99) in the training phase: upscaling/upscaler/model.py:20, 23, 284,
// 840-884 as lowered by Keras 2.2.x / TF 1.14).
//
// Stage 1 writes PARTIAL per-channel sums, one row per CTA, with plain stores: part[row][q*C + c].  It is either one of
// the two kernels here (forward: sum z, sum z^2; backward: sum dy, sum dy*xhat, sum g*min(u,0)) or the epilogue of the
// convolution that produces the tensor (b2_epilogue.stats_part: sums of the fp32 accumulator values, no extra pass).
// Stage 2 is ONE small launch per BatchNorm (a CTA per 64 channels): it adds the partial rows in a fixed order, optionally exchanges the sums
// with the other data-parallel replicas over NVLink peer memory (b2_peer.cuh; summed in rank order), and finishes the
// layer's small vectors: forward -> scale / shift / mean / invstd (+ the Keras moving-statistics update); backward ->
// the two sums the apply kernel needs, and dgamma / dbeta / dalpha ACCUMULATED into the gradient buffer.
//
// No atomics and no zero-fill anywhere: the result does not depend on CTA scheduling, so a step is bit-reproducible
// and N replicas hold identical bits; the tape of a forward pass owns its statistics (a second forward through the
// same layers — real and fake batches of the two-input discriminator model, train_gan3.py:353 — cannot disturb them).
#include <cuda_fp16.h>

#include <mutex>

#include "../../include/b2srgan.h"
#include "b2_conv_epilogue.cuh"
#include "b2_host.h"
#include "b2_peer.cuh"

using namespace b2;

namespace {

constexpr int kEwThreads = 256;
constexpr int kFinThreads = 1024;
constexpr int kFinMaxElems = 3 * 4096 / 2;  // widest layer the peer staging area takes: 2 sums x C <= kPeerMaxElems

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
__device__ __forceinline__ void loadp8(const float* p, int c0, float (&v)[8], float dflt) {
    if (p) {
        const float4 a = *reinterpret_cast<const float4*>(p + c0), b = *reinterpret_cast<const float4*>(p + c0 + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = dflt;
    }
}

// Per-thread partial sums of NQ quantities x 8 channels -> this block's row of `part`.  Threads that share a channel
// group sit `groups` apart; lanes of one warp are folded with xor shuffles, warps meet in shared memory, summed in
// thread order.
template <int NQ>
__device__ __forceinline__ void block_combine_store(float (&acc)[NQ][8], int groups, int c0, int C, float* row) {
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
            float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            for (int t = threadIdx.x; t < kEwThreads; t += step) {
                const float4 lo = *reinterpret_cast<const float4*>(&sh[q][t][0]);
                const float4 hi = *reinterpret_cast<const float4*>(&sh[q][t][4]);
                s[0] += lo.x; s[1] += lo.y; s[2] += lo.z; s[3] += lo.w;
                s[4] += hi.x; s[5] += hi.y; s[6] += hi.z; s[7] += hi.w;
            }
            float* d = row + q * C + c0;
            *reinterpret_cast<float4*>(d) = make_float4(s[0], s[1], s[2], s[3]);
            *reinterpret_cast<float4*>(d + 4) = make_float4(s[4], s[5], s[6], s[7]);
        }
    }
}

// ------------------------------------------------------------------------------------------------ stage 1, forward
__global__ void __launch_bounds__(kEwThreads)
bn_stats_partial_kernel(const void* __restrict__ x, long rows, int C, int f16, float* __restrict__ part, int zero_rows) {
    b2::pdl_entry();
    // rows of `part` beyond this launch's grid, up to zero_rows, are cleared (callers that read a fixed number of rows)
    for (int r = int(gridDim.x) + int(blockIdx.x); r < zero_rows; r += int(gridDim.x))
        for (int i = threadIdx.x; i < 2 * C; i += kEwThreads) part[size_t(r) * 2 * C + i] = 0.f;
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
    block_combine_store<2>(acc, groups, c0, C, part + size_t(blockIdx.x) * 2 * C);
}

__device__ __forceinline__ float act_slope(float u, int act, float leaky, float alpha) {
    switch (act) {
        case ACT_RELU: return u > 0.f ? 1.f : 0.f;
        case ACT_LEAKY: return u > 0.f ? 1.f : leaky;
        case ACT_PRELU: return u > 0.f ? 1.f : alpha;
        default: return 1.f;
    }
}

// ------------------------------------------------------------------------------------------------ stage 1, backward
// y = act(u), u = z*scale + shift:  dy = g*act'(u);  part = [sum dy | sum dy*xhat | sum g*min(u,0)],  xhat = (z-mean)*invstd
__global__ void __launch_bounds__(kEwThreads)
bn_bwd_partial_kernel(const void* __restrict__ z, const void* __restrict__ gin, long rows, int C, int f16,
                      const float* __restrict__ aff, const float* __restrict__ alpha, int act, float leaky,
                      float* __restrict__ part) {
    b2::pdl_entry();
    const int groups = C >> 3, g = threadIdx.x % groups, c0 = g * 8;
    const int rpb = kEwThreads / groups;
    float sc[8], sh[8], al[8], mu[8], is[8];
    loadp8(aff, c0, sc, 1.f);
    loadp8(aff + C, c0, sh, 0.f);
    loadp8(aff + 2 * C, c0, mu, 0.f);
    loadp8(aff + 3 * C, c0, is, 1.f);
    loadp8(alpha, c0, al, 0.f);
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
    block_combine_store<3>(acc, groups, c0, C, part + size_t(blockIdx.x) * 3 * C);
}

// ------------------------------------------------------------------------------------------------ stage 2
// One CTA per block of kFinChan channels (a single CTA for the 64-channel layers of the generator): the partial rows of
// its channels are added in a fixed order -- S interleaved row slices run concurrently, each in ascending row order, then
// the slices in ascending order -- so the sums do not depend on scheduling.  tot[q * cb + cl] = quantity q of channel
// c0 + cl, cb = channels of this CTA.
constexpr int kFinChan = 64;

template <int NQ>
__device__ __forceinline__ void sum_partial_rows(const float* __restrict__ part, int nparts, int C, int c0, int cb, float* sh,
                                                 float* tot) {
    // 16-byte loads: thread = (group of 4 consecutive channels of one quantity, row slice).  With 64 channels and three
    // quantities that is 48 groups x 21 slices, ~14 rows per thread, four loads in flight: the 296 partial rows of a layer
    // are in after ~4 dependent L2 round trips (the scalar form needed ~15 and cost 12-15 us per BatchNorm launch).
    const int n = NQ * cb, n4 = n >> 2;                    // cb is a multiple of 4 (C % 8 == 0)
    const int S = kFinThreads / n4;                        // row slices summed concurrently
    const int v = threadIdx.x % n4, sl = threadIdx.x / n4;
    const int e4 = v * 4, q = e4 / cb, cl = e4 - q * cb;
    const size_t col = size_t(q) * C + c0 + cl, pitch = size_t(NQ) * C;
    if (sl < S) {
        float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0, a2 = a0, a3 = a0;
        auto add = [](float4& a, const float4 b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; };
        int r = sl;
        for (; r + 3 * S < nparts; r += 4 * S) {           // four loads in flight; the order of the adds is fixed
            const float4 v0 = *reinterpret_cast<const float4*>(part + size_t(r) * pitch + col);
            const float4 v1 = *reinterpret_cast<const float4*>(part + size_t(r + S) * pitch + col);
            const float4 v2 = *reinterpret_cast<const float4*>(part + size_t(r + 2 * S) * pitch + col);
            const float4 v3 = *reinterpret_cast<const float4*>(part + size_t(r + 3 * S) * pitch + col);
            add(a0, v0); add(a1, v1); add(a2, v2); add(a3, v3);
        }
        for (; r < nparts; r += S) add(a0, *reinterpret_cast<const float4*>(part + size_t(r) * pitch + col));
        float* d = sh + sl * n + e4;
        d[0] = (a0.x + a1.x) + (a2.x + a3.x);
        d[1] = (a0.y + a1.y) + (a2.y + a3.y);
        d[2] = (a0.z + a1.z) + (a2.z + a3.z);
        d[3] = (a0.w + a1.w) + (a2.w + a3.w);
    }
    __syncthreads();
    if (int(threadIdx.x) < n) {
        float s = 0.f;
        for (int k = 0; k < S; ++k) s += sh[k * n + threadIdx.x];
        tot[threadIdx.x] = s;
    }
    __syncthreads();
}

// In-place cross-replica sum of quantities [0, nq) of this CTA's channels (no-op for world <= 1).  Slot of an element in
// the peer staging area = its index in the layer's [nq][C] sum vector, so CTAs never share a slot.
__device__ __forceinline__ void exchange(const PeerBases& pb, float* tot, int nq, int C, int c0, int cb) {
    if (pb.world <= 1) return;
    const int n = nq * cb;
    if (int(threadIdx.x) < n) {
        const int q = threadIdx.x / cb, cl = threadIdx.x - q * cb;
        const int slot = q * C + c0 + cl;
        peer_put(pb, slot, tot[threadIdx.x]);
        tot[threadIdx.x] = peer_get_sum(pb, slot);
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kFinThreads)
bn_fwd_finalize_kernel(const float* __restrict__ part, int nparts, int C, float count, PeerBases pb, float* sums_out,
                       const float* __restrict__ gamma, const float* __restrict__ beta, float eps, float momentum,
                       float* moving_mean, float* moving_var, float* aff, float* range_watch, float range_limit_sq) {
    b2::pdl_entry();
    __shared__ float sh[4 * kFinThreads];
    __shared__ float tot[3 * kFinChan];
    peer_resolve_epoch(pb);
    const int c0 = blockIdx.x * kFinChan, cb = min(kFinChan, C - c0);
    sum_partial_rows<2>(part, nparts, C, c0, cb, sh, tot);
    exchange(pb, tot, 2, C, c0, cb);
    const int cl = threadIdx.x;
    if (cl >= cb) return;
    const int c = c0 + cl;
    if (sums_out) { sums_out[c] = tot[cl]; sums_out[C + c] = tot[cb + cl]; }
    if (!aff) return;
    const float m = tot[cl] / count;
    float var = tot[cb + cl] / count - m * m;
    var = fmaxf(var, 0.f);
    // range watch of 16-bit storage: a channel whose mean square comes within the caller's margin of the type's edge (or is
    // not finite) raises the flag; the host reads it with the step's loss (upscaler/training.py)
    if (range_watch && !(m * m + var < range_limit_sq)) *range_watch = 1.f;
    const float inv = rsqrtf(var + eps);
    const float sc = gamma[c] * inv;
    aff[c] = sc;
    aff[C + c] = beta[c] - m * sc;
    aff[2 * C + c] = m;
    aff[3 * C + c] = inv;
    if (moving_mean) {
        // keras/layers/normalization.py (2.2.x): sample variance = var * n / (n - (1 + eps))
        const float adj = var * (count / (count - (1.f + eps)));
        moving_mean[c] = moving_mean[c] * momentum + m * (1.f - momentum);
        moving_var[c] = moving_var[c] * momentum + adj * (1.f - momentum);
    }
}

// part rows hold [s1 | s2 | dalpha_local]; only s1, s2 cross the replicas.  sums = [s1 | s2] (global) for bn_bwd_apply;
// dbeta += s1*gscale, dgamma += s2*gscale (gscale = 1/world: the gradient all-reduce sums the replicas again),
// dalpha += dalpha_local (a per-replica share, like every other weight gradient).
__global__ void __launch_bounds__(kFinThreads)
bn_bwd_finalize_kernel(const float* __restrict__ part, int nparts, int C, PeerBases pb, float* raw_out, float* sums,
                       float* dgamma, float* dbeta, float* dalpha, float gscale) {
    b2::pdl_entry();
    __shared__ float sh[4 * kFinThreads];
    __shared__ float tot[3 * kFinChan];
    peer_resolve_epoch(pb);
    const int c0 = blockIdx.x * kFinChan, cb = min(kFinChan, C - c0);
    sum_partial_rows<3>(part, nparts, C, c0, cb, sh, tot);
    const int cl = threadIdx.x;
    if (raw_out) {          // NCCL fallback: hand the local sums to the host-side all-reduce, finish in a second call
        if (cl < cb) {
            raw_out[c0 + cl] = tot[cl];
            raw_out[C + c0 + cl] = tot[cb + cl];
            raw_out[2 * C + c0 + cl] = tot[2 * cb + cl];
        }
        return;
    }
    exchange(pb, tot, 2, C, c0, cb);
    if (cl >= cb) return;
    const int c = c0 + cl;
    const float s1 = tot[cl], s2 = tot[cb + cl];
    if (sums) { sums[c] = s1; sums[C + c] = s2; }
    if (dbeta) dbeta[c] += s1 * gscale;
    if (dgamma) dgamma[c] += s2 * gscale;
    if (dalpha) dalpha[c] += tot[2 * cb + cl];
}

inline int launch_ok(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("%s launch: %s", what, cudaGetErrorString(e));
    return 0;
}

// Stage-1 grid: one block per kEwThreads / (C/8) rows, at most 2 blocks per SM (= b2_bn_part_rows()).
inline int part_grid(long rows, int C, int* blocks) {
    if (C % 8 != 0 || C / 8 > kEwThreads || kEwThreads % (C / 8) != 0) return -1;
    const int rpb = kEwThreads / (C / 8);
    long need = (rows + rpb - 1) / rpb;
    const long cap = 2L * device_sm_count();
    *blocks = int(need < cap ? (need < 1 ? 1 : need) : cap);
    return 0;
}

}  // namespace

namespace b2 {
// Statistics of a finished 16-bit tensor into ALL bn_part_rows() rows of `part` (the convolution launchers' fallback when
// the kernel that ran could not fuse b2_epilogue.stats_part into its epilogue).
int launch_bn_stats_all_rows(const void* x, long rows, int C, int f16, float* part, cudaStream_t stream) {
    int blocks;
    if (part_grid(rows, C, &blocks)) return set_error("BatchNorm statistics: C=%d unsupported (C/8 must divide %d)", C, kEwThreads);
    b2::launch_k(bn_stats_partial_kernel, blocks, kEwThreads, 0, stream, x, rows, C, f16, part, 2 * device_sm_count());
    return launch_ok("bn_stats_partial");
}
}  // namespace b2

#define B2_STREAM reinterpret_cast<cudaStream_t>(stream)

extern "C" {

int b2_bn_part_rows(void) { return 2 * device_sm_count(); }
int b2_sm_count(void) { return device_sm_count(); }

int b2_bn_stats_partial(const void* x, int64_t rows, int C, int f16, float* part, int* nparts, void* stream) {
    if (!x || !part || !nparts) return set_error("b2_bn_stats_partial: null pointer");
    int blocks;
    if (part_grid(rows, C, &blocks)) return set_error("b2_bn_stats_partial: C=%d unsupported (C/8 must divide %d)", C, kEwThreads);
    b2::launch_k(bn_stats_partial_kernel, blocks, kEwThreads, 0, B2_STREAM, x, rows, C, f16, part, 0);
    *nparts = blocks;
    return launch_ok("b2_bn_stats_partial");
}

int b2_bn_fwd_finalize(const float* part, int nparts, int C, float count, void* const* peer_regions, int rank, int world,
                       unsigned epoch, const unsigned* epoch_base, float* sums_out, const float* gamma, const float* beta, float eps, float momentum,
                       float* moving_mean, float* moving_var, float* aff, float* range_watch, float range_limit, void* stream) {
    if (!part || nparts < 1) return set_error("b2_bn_fwd_finalize: no partial sums");
    if (aff && (!gamma || !beta)) return set_error("b2_bn_fwd_finalize: gamma / beta missing");
    if (C < 1 || 3 * C > kFinMaxElems) return set_error("b2_bn_fwd_finalize: C=%d exceeds %d", C, kFinMaxElems / 3);
    if ((C & 3) || (reinterpret_cast<uintptr_t>(part) & 15)) return set_error("b2_bn_fwd_finalize: C must be a multiple of 4 and part 16-byte aligned");
    PeerBases pb;
    if (int rc = make_peer_bases(peer_regions, rank, world, epoch, &pb, epoch_base)) return rc;
    if (pb.world > 1 && 2 * C > kPeerMaxElems) return set_error("b2_bn_fwd_finalize: %d sums exceed the peer staging area", 2 * C);
    b2::launch_k(bn_fwd_finalize_kernel, (C + kFinChan - 1) / kFinChan, kFinThreads, 0, B2_STREAM, part, nparts, C, count, pb, sums_out, gamma, beta, eps, momentum,
                                                             moving_mean, moving_var, aff, range_watch, range_limit * range_limit);
    return launch_ok("b2_bn_fwd_finalize");
}

int b2_bn_bwd_partial(const void* z, const void* g, int64_t rows, int C, int f16, const float* aff, const float* alpha, int act,
                      float leaky, float* part, int* nparts, void* stream) {
    if (!z || !g || !aff || !part || !nparts) return set_error("b2_bn_bwd_partial: null pointer");
    int blocks;
    if (part_grid(rows, C, &blocks)) return set_error("b2_bn_bwd_partial: C=%d unsupported", C);
    b2::launch_k(bn_bwd_partial_kernel, blocks, kEwThreads, 0, B2_STREAM, z, g, rows, C, f16, aff, alpha, act, leaky, part);
    *nparts = blocks;
    return launch_ok("b2_bn_bwd_partial");
}

int b2_bn_bwd_finalize(const float* part, int nparts, int C, void* const* peer_regions, int rank, int world, unsigned epoch,
                       const unsigned* epoch_base, float* raw_out, float* sums, float* dgamma, float* dbeta, float* dalpha, float gscale, void* stream) {
    if (!part || nparts < 1) return set_error("b2_bn_bwd_finalize: no partial sums");
    if (C < 1 || 3 * C > kFinMaxElems) return set_error("b2_bn_bwd_finalize: C=%d exceeds %d", C, kFinMaxElems / 3);
    if ((C & 3) || (reinterpret_cast<uintptr_t>(part) & 15)) return set_error("b2_bn_bwd_finalize: C must be a multiple of 4 and part 16-byte aligned");
    PeerBases pb;
    if (int rc = make_peer_bases(peer_regions, rank, world, epoch, &pb, epoch_base)) return rc;
    if (pb.world > 1 && 2 * C > kPeerMaxElems) return set_error("b2_bn_bwd_finalize: %d sums exceed the peer staging area", 2 * C);
    b2::launch_k(bn_bwd_finalize_kernel, (C + kFinChan - 1) / kFinChan, kFinThreads, 0, B2_STREAM, part, nparts, C, pb, raw_out, sums, dgamma, dbeta, dalpha, gscale);
    return launch_ok("b2_bn_bwd_finalize");
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------ allocation-free variant
// b2_bn_fwd_stats_static: both stages of the forward statistics in one call, with the partial rows kept in a buffer that
// is part of the library image instead of one the caller provides.  It exists for hosts that may not allocate device
// memory when they reach this code: a plugin running inside another framework's memory arena (every cudaMalloc there
// fragments or bypasses the arena's accounting), a real-time capture / playback loop where an allocation is an unbounded
// stall, or a process whose device-memory quota was handed out at start-up and is fully spoken for.  Such callers own
// their input and output tensors already; the only thing they lack is the [b2_bn_part_rows()][2*C] scratch tensor that
// b2_bn_stats_partial / b2_bn_fwd_finalize need between the two launches.
//
// The buffer costs kStaticPartRows * 2 * kStaticMaxC floats (1 212 416 bytes) on every GPU that loads the library, for the
// life of the process.  It is shared by all calls: nothing in it needs clearing (stage 1 rewrites every row stage 2 reads,
// each call), but two calls that overlap on one GPU -- different streams, or different host threads -- write the same rows
// and corrupt each other.  Keep calls on a device ordered.
constexpr int kStaticPartRows = 296;            // rows stage 1 may write: 2 CTAs per SM of a 148-SM B200
constexpr int kStaticMaxC = 512;                // widest BatchNormalization on the path (discriminator blocks, model.py:840-884)
static __device__ __align__(256) float g_static_part[size_t(kStaticPartRows) * 2 * kStaticMaxC];

namespace {
// Device address of g_static_part on the current device.  In host code the symbol's own name is only a handle the runtime
// looks the array up by -- it is not a pointer a kernel may dereference -- so the address is asked for once per device,
// with that device current, and remembered.  Returns null (error text set) when the lookup fails.
float* static_part_ptr() {
    static std::mutex mu;
    static float* ptrs[64] = {nullptr};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) {
        set_error("b2_bn_fwd_stats_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_part);
        if (e != cudaSuccess || !p) {
            set_error("b2_bn_fwd_stats_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
            return nullptr;
        }
        ptrs[dev] = static_cast<float*>(p);
    }
    return ptrs[dev];
}
}  // namespace

extern "C" int b2_bn_fwd_stats_static(const void* x, int64_t rows, int C, int f16, float count, const float* gamma, const float* beta,
                                      float eps, float* sums_out, float* aff, void* stream) {
    if (!x) return set_error("b2_bn_fwd_stats_static: x is null");
    if (!sums_out && !aff) return set_error("b2_bn_fwd_stats_static: neither sums_out nor aff given");
    if (aff && (!gamma || !beta)) return set_error("b2_bn_fwd_stats_static: gamma / beta missing");
    if (rows < 1) return set_error("b2_bn_fwd_stats_static: rows=%lld", static_cast<long long>(rows));
    if (C < 8 || C > kStaticMaxC)
        return set_error("b2_bn_fwd_stats_static: C=%d is outside the built-in scratch buffer's range (8..%d channels); "
                         "use b2_bn_stats_partial + b2_bn_fwd_finalize with a buffer of your own", C, kStaticMaxC);
    int blocks;
    if (part_grid(rows, C, &blocks)) return set_error("b2_bn_fwd_stats_static: C=%d unsupported (C/8 must divide %d)", C, kEwThreads);
    if (blocks > kStaticPartRows) blocks = kStaticPartRows;      // a GPU with more SMs than a B200 still fits the buffer
    float* part = static_part_ptr();                             // first call on a device: symbol lookup (not capturable)
    if (!part) return -1;
    b2::launch_k(bn_stats_partial_kernel, blocks, kEwThreads, 0, B2_STREAM, x, long(rows), C, f16, part, 0);
    if (int rc = launch_ok("b2_bn_fwd_stats_static (stage 1)")) return rc;
    PeerBases pb;
    make_peer_bases(nullptr, 0, 1, 0u, &pb, nullptr);            // single replica: no exchange, nothing to fail
    b2::launch_k(bn_fwd_finalize_kernel, (C + kFinChan - 1) / kFinChan, kFinThreads, 0, B2_STREAM, static_cast<const float*>(part), blocks, C,
                 count, pb, sums_out, gamma, beta, eps, 0.f, static_cast<float*>(nullptr), static_cast<float*>(nullptr), aff,
                 static_cast<float*>(nullptr), 0.f);
    return launch_ok("b2_bn_fwd_stats_static (stage 2)");
}
