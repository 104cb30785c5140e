// b2_conv_stack.cu — stride-1 NHWC convolution with Cout = 64 (the residual-trunk convs, model.py:19, 22, 283; the
// folded 9x1 head conv; their data gradients; VGG19 block1_conv2) as a tcgen05 implicit GEMM whose N dimension stacks
// R vertically adjacent OUTPUT rows.
//
// Why: a tcgen05.mma with M = 128 costs ~80-92 cycles for every N <= 128 and 128 cycles at N = 256 when both operands
// come from shared memory (measured: tools/b2_probe.cu "mma2"/"mma3", independent of the number of accumulators), so
// the N = 64 main loop of b2_conv_halo.cu cannot exceed ~35-40 % of the tensor peak.  Here accumulator row p holds R
// output pixels: D[p, j*64 + c] = out(y_p + j, x_p, c), j < R.  The A operand of a "shift" (s, dx) is the halo tile
// viewed at rows y_p + s: that input pixel feeds out(y_p + j) through filter row ky = s - j, so the B operand of the
// shift is the slab [W(ky = s - j_lo, dx) | W(ky = s - j_lo - 1, dx) | ...] — consecutive 8 KB blocks of a per-dx weight
// strip stored in DESCENDING ky — and the MMA runs at N = 64 * (number of valid j), into accumulator columns
// [64*j_lo, ...).  A 3x3 filter with R = 3 needs 5 shifts per dx with N = 64/128/192/128/64 instead of 9 N=64 MMAs for
// the same 3 output rows: 15 instead of 27 MMA groups, none of them wasted on zero weights.
//   M tile   = 16 row-groups x 8 pixels; group g covers output rows y0 + R*g .. y0 + R*g + R-1 -> 16R x 8 output pixels
//   A tile   = (16R + kh - 1) x (8 + kw - 1) pixels x 64 channels (128-byte swizzled rows, one TMA box per 64 channels),
//              view of shift (s, dx): start + (s*HW + dx)*128 B, 8-row-group stride R*HW*128 B
//   weights  = all kh*kw*(Cin/64) blocks of [64 x 64] resident in shared memory (<= 96 KB), or -- 5x5 filters: 200 KB --
//              STREAMED: a ring of strip slots (one per-dx strip of kh blocks + R-1 zero blocks each), filled by a
//              second producer warp in the order the MMA warp walks (cb, dx); 200 KB per 32x8-pixel tile from L2 is
//              ~20 B/clk/SM, far below what the MMAs of the tile take
// The very first MMA group of a tile must overwrite all R*64 accumulator columns (tcgen05.mma has one accumulate flag
// per instruction): the (dx = 0, cb = 0) strip is placed last in shared memory and followed by R-1 blocks of zeros, so
// that group runs at N = 64R with B = [W(ky=0) | 0 | ... | 0]; every other MMA accumulates.
//
// Warp roles: 0 = TMA producer (input tiles; resident weights), 1 = MMA issuer + TMEM owner, 2..9 = epilogue (TMEM lane
// quarter = warp % 4, channel half = (warp - 2) / 4), 10 = TMA producer of streamed weight strips.
#include <string.h>

#include "b2_conv_epilogue.cuh"
#include "b2_conv_igemm.cuh"
#include "b2_host.h"
#include "b2_ptx.cuh"

namespace b2 {

constexpr int kStackThreads = 320;          // TMA warp, MMA warp, 8 epilogue warps (+ a weight-strip warp when streaming: 352)
constexpr int kStackTW = 8;                    // pixels per row group (one 8-row swizzle atom)
constexpr int kStackGroups = 16;               // row groups per M = 128 tile
constexpr int kStackMaxK = 9;                  // filter rows / columns supported
constexpr int kStackMaxShifts = kStackMaxK + 3;
constexpr int kStackSmemMax = 227 * 1024;

struct StackGeom {
    int kh, kw;          // filter rectangle (all kh*kw taps present)
    int dy_min, dx_min;  // most negative tap offsets
    int HH, HW;          // halo tile (pixels)
    int a_block;         // bytes per halo block, multiple of 1024
    int a_stages;
    int b_bytes;         // resident weights + zero blocks
    int nshift;          // kh + R - 1
    int b_slots;         // 0: weights resident; >= 2: weight strips streamed through a ring of this many slots
    int slot_bytes;      // (kh + R - 1) blocks of 8 KB
    int16_t wt[kStackMaxK][kStackMaxK];   // [ky][dxi] -> index of the tap's [Cout x Cin] slab in the packed weights
    uint8_t cnt[kStackMaxShifts];         // per shift: number of stacked output rows it feeds
    uint8_t jlo[kStackMaxShifts];         //            first of them
    uint8_t i0[kStackMaxShifts];          //            first weight block of the descending-ky strip
};

// Sum of v[e] over the 32 lanes of a warp for all 32 elements at once ("reduce-scatter" butterfly: 16 + 8 + 4 + 2 + 1
// shuffles): on return lane l holds the total of element l in v[0].  A lane with bit b set keeps the upper half.
__device__ __forceinline__ void warp_reduce_scatter32(float (&v)[32], int lane) {
#pragma unroll
    for (int half = 16; half >= 1; half >>= 1) {
        const bool hi = (lane & half) != 0;
#pragma unroll
        for (int i = 0; i < half; ++i) {
            const float send = hi ? v[i] : v[i + half];
            const float keep = hi ? v[i + half] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, half);
        }
    }
}

// Epilogue warp main loop.  Eight epilogue warps: warp (quarter, half) owns TMEM lanes [32*quarter, +32) and channels
// [32*half, +32) of each of the R stacked output pixels of an accumulator row, so two warps per scheduler overlap one
// another's TMEM-load and global-load latency.  The residual row of block j+1 is prefetched while block j is processed.
// F16 / ACT are compile-time so the per-element path carries no uniform branches; the j loop stays rolled (a fully
// unrolled epilogue measured 24 % of its stall samples as instruction-cache misses).
// STATS (train-mode BatchNorm statistics fused into the conv that feeds it, b2_epilogue.stats_part): every thread keeps
// running sums of v and v*v for its 32 channels over all the pixels it stores (fp32 values, before the 16-bit rounding);
// at the end the lanes of a warp are combined with the reduce-scatter above and lane l leaves channel 32*half + l's
// (sum, sum of squares) in s_stats[warp][l] for the CTA-level combine at the end of the kernel.  Residual / gate do not
// occur in front of a BatchNorm, so they are compiled out of this variant.
template <int R, int ACT, int F16, bool STATS>
__device__ __forceinline__ void stack_epilogue_loop(const IgemmParams& p, uint32_t tmem_base, uint64_t* tfull, uint64_t* tempty,
                                                    int quarter, int half, int lane, const float* s_scale,
                                                    const float* s_shift, const float* s_alpha, float2* s_stats) {
    constexpr int kAccCols = 64 * R;
    const int row = quarter * 32 + lane;
    const int r_tx = row % kStackTW, r_g = row / kStackTW;
    const uint32_t sc_a = smem_u32(s_scale + 32 * half), sh_a = smem_u32(s_shift + 32 * half), al_a = smem_u32(s_alpha + 32 * half);
    const bool has_res = !STATS && p.residual != nullptr, has_gate = !STATS && p.gate != nullptr;
    float st_sum[STATS ? 32 : 1], st_sq[STATS ? 32 : 1];
    if constexpr (STATS) {
#pragma unroll
        for (int e = 0; e < 32; ++e) { st_sum[e] = 0.f; st_sq[e] = 0.f; }
    }
    const size_t row_pitch = size_t(p.OW) * size_t(p.out_ld);
    int as = 0; uint32_t aph = 0;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
        int t = tile;
        const int xt = t % p.tiles_x; t /= p.tiles_x;
        const int yt = t % p.tiles_y; t /= p.tiles_y;
        const int gx = xt * kStackTW + r_tx;
        const int gy0 = yt * (kStackGroups * R) + r_g * R;
        size_t obase = ((size_t(t) * p.OH + size_t(gy0)) * p.OW + size_t(gx)) * size_t(p.out_ld) + size_t(32 * half);
        uint8v nxt[2];
#pragma unroll
        for (int g = 0; g < 2; ++g)
#pragma unroll
            for (int e = 0; e < 8; ++e) nxt[g].v[e] = 0;
        if (has_res && gx < p.W && gy0 < p.H) {
            nxt[0] = ldg256(p.residual + obase);
            nxt[1] = ldg256(p.residual + obase + 16);
        }
        if (has_res && half == 0 && tile + int(gridDim.x) < p.num_tiles) {
            // pull the NEXT tile's residual rows into L2 a whole tile ahead (they were written two layers ago and have
            // usually left L2); the register prefetch one block ahead then only has to cover an L2 hit
            int t2 = tile + int(gridDim.x);
            const int xt2 = t2 % p.tiles_x; t2 /= p.tiles_x;
            const int yt2 = t2 % p.tiles_y; t2 /= p.tiles_y;
            const int gx2 = xt2 * kStackTW + r_tx, gy2 = yt2 * (kStackGroups * R) + r_g * R;
            if (gx2 < p.W) {
                const __nv_bfloat16* rp = p.residual + ((size_t(t2) * p.OH + size_t(gy2)) * p.OW + size_t(gx2)) * size_t(p.out_ld);
#pragma unroll
                for (int j = 0; j < R; ++j)
                    if (gy2 + j < p.H) asm volatile("prefetch.global.L2 [%0];" ::"l"(rp + size_t(j) * row_pitch));
            }
        }
        mbar_wait(&tfull[as], aph);
        tc_fence_after();
        uint32_t ta = tmem_base + (uint32_t(quarter * 32) << 16) + uint32_t(as * kAccCols + 32 * half);
#pragma unroll 1
        for (int j = 0; j < R; ++j, obase += row_pitch, ta += 64) {
            uint32_t r[32];
            __syncwarp();       // rows outside the image skip the tail of the previous iteration
            tmem_ld32(ta, r);
            uint8v cur[2];
            cur[0] = nxt[0];
            cur[1] = nxt[1];
            const bool valid = gx < p.W && gy0 + j < p.H;
            if (has_res && j + 1 < R && gx < p.W && gy0 + j + 1 < p.H) {
                nxt[0] = ldg256(p.residual + obase + row_pitch);
                nxt[1] = ldg256(p.residual + obase + row_pitch + 16);
            }
            uint8v gv[2];
            if (has_gate && valid) {
                gv[0] = ldg256(reinterpret_cast<const __nv_bfloat16*>(p.gate) + obase);
                gv[1] = ldg256(reinterpret_cast<const __nv_bfloat16*>(p.gate) + obase + 16);
            }
            tmem_ld_wait();
            if (!valid) continue;
            float v[32];
#pragma unroll
            for (int g = 0; g < 8; ++g) {
                const float4 sc = lds_f4(sc_a + 16 * g), sh = lds_f4(sh_a + 16 * g);
                float4 al = make_float4(0.f, 0.f, 0.f, 0.f);
                if constexpr (ACT == ACT_PRELU) al = lds_f4(al_a + 16 * g);
                v[4 * g + 0] = apply_act_t<ACT>(fmaf(__uint_as_float(r[4 * g + 0]), sc.x, sh.x), p.leaky, al.x);
                v[4 * g + 1] = apply_act_t<ACT>(fmaf(__uint_as_float(r[4 * g + 1]), sc.y, sh.y), p.leaky, al.y);
                v[4 * g + 2] = apply_act_t<ACT>(fmaf(__uint_as_float(r[4 * g + 2]), sc.z, sh.z), p.leaky, al.z);
                v[4 * g + 3] = apply_act_t<ACT>(fmaf(__uint_as_float(r[4 * g + 3]), sc.w, sh.w), p.leaky, al.w);
            }
            if (has_gate) {
#pragma unroll
                for (int g = 0; g < 4; ++g)
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float2 f = unpack2(gv[g >> 1].v[(g & 1) * 4 + e], F16);
                        if (!(f.x > 0.f)) v[g * 8 + 2 * e] *= p.gate_slope;
                        if (!(f.y > 0.f)) v[g * 8 + 2 * e + 1] *= p.gate_slope;
                    }
            }
            if (has_res) {
#pragma unroll
                for (int g = 0; g < 4; ++g)
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float2 f = unpack2(cur[g >> 1].v[(g & 1) * 4 + e], F16);
                        v[g * 8 + 2 * e] += f.x;
                        v[g * 8 + 2 * e + 1] += f.y;
                    }
            }
            if constexpr (STATS) {
#pragma unroll
                for (int e = 0; e < 32; ++e) { st_sum[e] += v[e]; st_sq[e] = fmaf(v[e], v[e], st_sq[e]); }
            }
            if (!p.out_fp32) {
                __nv_bfloat16* op = reinterpret_cast<__nv_bfloat16*>(p.out) + obase;
#pragma unroll
                for (int g = 0; g < 2; ++g) {
                    uint8v o;
#pragma unroll
                    for (int e = 0; e < 8; ++e) o.v[e] = pack2(v[g * 16 + 2 * e], v[g * 16 + 2 * e + 1], F16);
                    stg256(op + 16 * g, o);
                }
            } else {
                float* op = reinterpret_cast<float*>(p.out) + obase;
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    uint8v o;
#pragma unroll
                    for (int e = 0; e < 8; ++e) o.v[e] = __float_as_uint(v[8 * g + e]);
                    stg256(op + 8 * g, o);
                }
            }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[as]);
        if (++as == 2) { as = 0; aph ^= 1; }
    }
    if constexpr (STATS) {
        __syncwarp();
        warp_reduce_scatter32(st_sum, lane);
        warp_reduce_scatter32(st_sq, lane);
        s_stats[(half * 4 + quarter) * 32 + lane] = make_float2(st_sum[0], st_sq[0]);
    }
}

template <int R, int F16>
__device__ __forceinline__ void stack_epilogue_dispatch(const IgemmParams& p, uint32_t tmem_base, uint64_t* tfull, uint64_t* tempty,
                                                        int quarter, int half, int lane, const float* s_scale,
                                                        const float* s_shift, const float* s_alpha, float2* s_stats) {
    if (p.stats_part) {      // only offered for act == none without residual / gate (checked by the launcher)
        stack_epilogue_loop<R, ACT_NONE, F16, true>(p, tmem_base, tfull, tempty, quarter, half, lane, s_scale, s_shift, s_alpha, s_stats);
        return;
    }
    switch (p.act) {
        case ACT_RELU: stack_epilogue_loop<R, ACT_RELU, F16, false>(p, tmem_base, tfull, tempty, quarter, half, lane, s_scale, s_shift, s_alpha, s_stats); break;
        case ACT_LEAKY: stack_epilogue_loop<R, ACT_LEAKY, F16, false>(p, tmem_base, tfull, tempty, quarter, half, lane, s_scale, s_shift, s_alpha, s_stats); break;
        case ACT_PRELU: stack_epilogue_loop<R, ACT_PRELU, F16, false>(p, tmem_base, tfull, tempty, quarter, half, lane, s_scale, s_shift, s_alpha, s_stats); break;
        case ACT_TANH: stack_epilogue_loop<R, ACT_TANH, F16, false>(p, tmem_base, tfull, tempty, quarter, half, lane, s_scale, s_shift, s_alpha, s_stats); break;
        default: stack_epilogue_loop<R, ACT_NONE, F16, false>(p, tmem_base, tfull, tempty, quarter, half, lane, s_scale, s_shift, s_alpha, s_stats); break;
    }
}

template <int R, bool STREAM>
__global__ void __launch_bounds__(STREAM ? kStackThreads + 32 : kStackThreads, 1)
conv_stack_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                  const __grid_constant__ IgemmParams p, const __grid_constant__ StackGeom g) {
    b2::pdl_entry();
    constexpr int kBlk = 64 * 128;            // one [64 cout x 64 k] weight block
    constexpr int kThreads = STREAM ? kStackThreads + 32 : kStackThreads;
    constexpr int kAccCols = 64 * R;
    constexpr uint32_t kTmemCols = (2 * kAccCols) <= 128 ? 128 : ((2 * kAccCols) <= 256 ? 256 : 512);
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smemB = smem;
    uint8_t* smemA = smem + g.b_bytes;
    uint8_t* tail = smemA + g.a_stages * g.a_block;
    uint64_t* bars = reinterpret_cast<uint64_t*>(tail);
    uint64_t* afull = bars;            // [4]
    uint64_t* aempty = bars + 4;       // [4]
    uint64_t* bfull = bars + 8;        // [1]
    uint64_t* tfull = bars + 9;        // [2]
    uint64_t* tempty = bars + 11;      // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 13);
    uint64_t* sfull = bars + 16;       // [4] streamed weight strips
    uint64_t* sempty = bars + 20;      // [4]
    float* s_scale = reinterpret_cast<float*>(tail + 256);
    float* s_shift = s_scale + 64;
    float* s_alpha = s_shift + 64;
    float2* s_stats = reinterpret_cast<float2*>(s_alpha + 64);     // [8 epilogue warps][32 lanes] (sum, sum of squares)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cbs = p.cin >> 6;
    const int nstrips = g.kw * cbs;
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&mapA);
        tma_prefetch_desc(&mapB);
        for (int s = 0; s < 4; ++s) { mbar_init(&afull[s], 1); mbar_init(&aempty[s], 1); }
        mbar_init(bfull, 1);
        for (int s = 0; s < 4; ++s) { mbar_init(&sfull[s], 1); mbar_init(&sempty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&tfull[s], 1); mbar_init(&tempty[s], 8); }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<kTmemCols>(tmem_slot);
    if constexpr (!STREAM) {   // zero blocks behind the last strip (B operand of the tile-opening MMA group)
        uint4* z = reinterpret_cast<uint4*>(smemB + nstrips * g.kh * kBlk);
        for (int i = threadIdx.x; i < (R - 1) * kBlk / 16; i += kThreads) z[i] = make_uint4(0, 0, 0, 0);
        fence_proxy_async();
    } else {                // streamed: every ring slot ends in R-1 zero blocks the strip loads never touch
        for (int sl = 0; sl < g.b_slots; ++sl) {
            uint4* z = reinterpret_cast<uint4*>(smemB + sl * g.slot_bytes + g.kh * kBlk);
            for (int i = threadIdx.x; i < (R - 1) * kBlk / 16; i += kThreads) z[i] = make_uint4(0, 0, 0, 0);
        }
        fence_proxy_async();
    }
    for (int c = threadIdx.x; c < 64; c += kThreads) {
        s_scale[c] = p.scale ? p.scale[c] : 1.f;
        s_shift[c] = p.shift ? p.shift[c] : 0.f;
        s_alpha[c] = p.alpha ? p.alpha[c] : 0.f;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    const uint32_t a_bytes = uint32_t(g.HH * g.HW) * 128u;

    if (warp == 0) {
        // ------------------------------------------------------------ producer: resident weights, then halo tiles
        if (lane == 0 && !STREAM) {
            mbar_expect_tx(bfull, uint32_t(nstrips * g.kh) * kBlk);
            for (int dxi = 0; dxi < g.kw; ++dxi)
                for (int cb = 0; cb < cbs; ++cb) {
                    const int q = dxi * cbs + cb;
                    const int pos = q == 0 ? nstrips - 1 : q - 1;      // strip (dx 0, cb 0) sits last, before the zeros
                    for (int i = 0; i < g.kh; ++i)                      // descending ky
                        tma_load_2d(smemB + (pos * g.kh + i) * kBlk, &mapB, bfull,
                                    int(g.wt[g.kh - 1 - i][dxi]) * p.cin + cb * 64, 0);
                }
        }
        __syncwarp();
        int sa = 0; uint32_t pa = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
            int t = tile;
            const int xt = t % p.tiles_x; t /= p.tiles_x;
            const int yt = t % p.tiles_y; t /= p.tiles_y;
            const int x0 = xt * kStackTW + g.dx_min, y0 = yt * (kStackGroups * R) + g.dy_min;
            for (int cb = 0; cb < cbs; ++cb) {
                mbar_wait(&aempty[sa], pa ^ 1);
                if (lane == 0) {
                    mbar_expect_tx(&afull[sa], a_bytes);
                    tma_load_4d(smemA + sa * g.a_block, &mapA, &afull[sa], cb * 64, x0, y0, t);
                }
                __syncwarp();
                if (++sa == g.a_stages) { sa = 0; pa ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        const uint32_t idesc0 = p.f16 ? make_idesc_f16(128, 0) : make_idesc_bf16(128, 0);
        const uint32_t a_hi = sdesc_hi_k_sw128(uint32_t(R * g.HW) * 128u);
        constexpr uint32_t b_hi = sdesc_hi_k_sw128(1024);
        const uint32_t a_lo0 = sdesc_lo(smem_u32(smemA)), b_lo0 = sdesc_lo(smem_u32(smemB));
        int sa = 0; uint32_t pa = 0;
        int as = 0; uint32_t aph = 0;
        int sb = 0; uint32_t pb = 0;
        constexpr bool streamed = STREAM;
        if (!streamed) mbar_wait(bfull, 0);
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
            mbar_wait(&tempty[as], aph ^ 1);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + uint32_t(as * kAccCols);
            for (int cb = 0; cb < cbs; ++cb) {
                mbar_wait(&afull[sa], pa);
                tc_fence_after();
                const uint32_t a_lo_blk = a_lo0 + uint32_t(sa) * uint32_t(g.a_block >> 4);
                for (int dxi = 0; dxi < g.kw; ++dxi) {
                    const int q = dxi * cbs + cb;
                    const int pos = q == 0 ? nstrips - 1 : q - 1;
                    uint32_t b_lo_strip = b_lo0 + uint32_t(pos * g.kh) * (kBlk >> 4);
                    if (streamed) {
                        mbar_wait(&sfull[sb], pb);
                        tc_fence_after();
                        b_lo_strip = b_lo0 + uint32_t(sb) * uint32_t(g.slot_bytes >> 4);
                    }
                    for (int s = 0; s < g.nshift; ++s) {
                        const bool opening = (q | s) == 0;          // first MMA group of the tile: overwrite all R blocks
                        const uint32_t cnt = opening ? uint32_t(R) : uint32_t(g.cnt[s]);
                        const uint32_t idesc = idesc0 | ((cnt * 8u) << 17);            // N = 64*cnt -> N>>3 = 8*cnt
                        const uint32_t d = tmem_d + uint32_t(g.jlo[s]) * 64u;
                        const uint32_t a_lo = a_lo_blk + uint32_t(s * g.HW + dxi) * 8u;  // 128 B per pixel = 8 x 16 B
                        const uint32_t b_lo = b_lo_strip + uint32_t(g.i0[s]) * (kBlk >> 4);
                        const bool last = (dxi == g.kw - 1) && (s == g.nshift - 1);
                        if (elect_one()) {
                            umma_f16_lohi(d, a_lo, a_hi, b_lo, b_hi, idesc, opening ? 0u : 1u);
                            umma_f16_lohi(d, a_lo + 2, a_hi, b_lo + 2, b_hi, idesc, 1);
                            umma_f16_lohi(d, a_lo + 4, a_hi, b_lo + 4, b_hi, idesc, 1);
                            umma_f16_lohi(d, a_lo + 6, a_hi, b_lo + 6, b_hi, idesc, 1);
                            if (streamed && s == g.nshift - 1) umma_commit(&sempty[sb]);
                            if (last) {
                                umma_commit(&aempty[sa]);
                                if (cb == cbs - 1) umma_commit(&tfull[as]);
                            }
                        }
                        __syncwarp();
                    }
                    if (streamed && ++sb == g.b_slots) { sb = 0; pb ^= 1; }
                }
                if (++sa == g.a_stages) { sa = 0; pa ^= 1; }
            }
            if (++as == 2) { as = 0; aph ^= 1; }
        }
    } else if (STREAM && warp == 10) {
        // ------------------------------------------------------------ producer: streamed weight strips, (cb, dx) order
        {
            int sb = 0; uint32_t pb = 0;
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x)
                for (int cb = 0; cb < cbs; ++cb)
                    for (int dxi = 0; dxi < g.kw; ++dxi) {
                        mbar_wait(&sempty[sb], pb ^ 1);
                        if (lane == 0) {
                            mbar_expect_tx(&sfull[sb], uint32_t(g.kh) * kBlk);
                            for (int i = 0; i < g.kh; ++i)                      // descending ky
                                tma_load_2d(smemB + sb * g.slot_bytes + i * kBlk, &mapB, &sfull[sb],
                                            int(g.wt[g.kh - 1 - i][dxi]) * p.cin + cb * 64, 0);
                        }
                        __syncwarp();
                        if (++sb == g.b_slots) { sb = 0; pb ^= 1; }
                    }
        }
    } else {
        // ------------------------------------------------------------ epilogue: R output pixels per accumulator row
        const int quarter = warp & 3, half = (warp - 2) >> 2;
        if (p.f16) stack_epilogue_dispatch<R, 1>(p, tmem_base, tfull, tempty, quarter, half, lane, s_scale, s_shift, s_alpha, s_stats);
        else stack_epilogue_dispatch<R, 0>(p, tmem_base, tfull, tempty, quarter, half, lane, s_scale, s_shift, s_alpha, s_stats);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<kTmemCols>(tmem_base);
    }
    if (p.stats_part) {
        // this CTA's row of partial sums: the four lane-quarter warps of each channel half added in a fixed order; rows
        // of CTAs that do not exist in this launch (grid < stats_rows) are written as zeros
        if (threadIdx.x < 64) {
            const int c = threadIdx.x, hf = c >> 5, l = c & 31;
            float sm = 0.f, sq = 0.f;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float2 t = s_stats[(hf * 4 + q) * 32 + l];
                sm += t.x; sq += t.y;
            }
            float* rowp = p.stats_part + size_t(blockIdx.x) * 128;
            rowp[c] = sm;
            rowp[64 + c] = sq;
        }
        for (int r = int(gridDim.x) + int(blockIdx.x); r < p.stats_rows; r += int(gridDim.x))
            for (int i = threadIdx.x; i < 128; i += kThreads) p.stats_part[size_t(r) * 128 + i] = 0.f;
    }
}

// Filter geometry of a stride-1 conv whose taps form a full kh x kw rectangle; false otherwise.
static bool stack_rect(const IgemmParams& p, int& kh, int& kw, int& dy0, int& dx0) {
    int dy1 = -127, dx1 = -127;
    dy0 = 127; dx0 = 127;
    for (int t = 0; t < p.num_taps; ++t) {
        dy0 = p.dy[t] < dy0 ? p.dy[t] : dy0; dy1 = p.dy[t] > dy1 ? p.dy[t] : dy1;
        dx0 = p.dx[t] < dx0 ? p.dx[t] : dx0; dx1 = p.dx[t] > dx1 ? p.dx[t] : dx1;
    }
    kh = dy1 - dy0 + 1; kw = dx1 - dx0 + 1;
    if (kh > kStackMaxK || kw > kStackMaxK || kh * kw != p.num_taps) return false;
    bool seen[kStackMaxK][kStackMaxK] = {};
    for (int t = 0; t < p.num_taps; ++t) {
        bool& s = seen[p.dy[t] - dy0][p.dx[t] - dx0];
        if (s) return false;
        s = true;
    }
    return true;
}

// Rows stacked per accumulator row for an image of H rows: the largest R in {3, 2} whose 16R-row tiles waste at most
// ~15 % of the MMA rows and whose operands fit in shared memory; 0 when the kernel does not apply.
static int stack_pick_r(const ConvIgemmArgs& a, int kh, int kw, bool* streamed = nullptr) {
    const IgemmParams& p = a.p;
    const int cbs = p.cin / 64;
    const int wbytes = kh * kw * cbs * 64 * 128;
    const int tw = (p.W + kStackTW - 1) / kStackTW * kStackTW;
    for (int pass = 0; pass < 2; ++pass) {           // resident weights first, then streamed strips
        if (pass == 0 && wbytes > 96 * 1024) continue;
        for (int R = 3; R >= 2; --R) {
            const int th = kStackGroups * R;
            const long covered = long((p.H + th - 1) / th) * th * tw;
            if (long(p.H) * p.W * 100 < covered * 85) continue;
            const int a_block = (((th + kh - 1) * (kStackTW + kw - 1) * 128) + 1023) & ~1023;
            const int b_bytes = pass == 0 ? wbytes + (R - 1) * 8192 : 2 * (kh + R - 1) * 8192;
            if (b_bytes + 2 * a_block + 4096 > kStackSmemMax) continue;
            if (streamed) *streamed = pass == 1;
            return R;
        }
    }
    return 0;
}

bool conv_stack_applicable(const ConvIgemmArgs& a) {
    const IgemmParams& p = a.p;
    if (p.stats_part && (p.act != ACT_NONE || p.residual || p.gate || p.out_fp32)) return false;
    if (p.in_mul != 1 || p.n_valid != 64 || p.out_ld != 64 || a.w_rows != 64) return false;
    if (p.out_sy != 1 || p.out_sx != 1 || p.out_oy != 0 || p.out_ox != 0) return false;
    if (a.xC % 64 != 0 || p.cin != a.xC || p.num_taps < 2) return false;
    int kh, kw, dy0, dx0;
    if (!stack_rect(p, kh, kw, dy0, dx0)) return false;
    return stack_pick_r(a, kh, kw) != 0;
}

template <int R, bool STREAM>
static int launch_stack_r(const CUtensorMap& mapA, const CUtensorMap& mapB, const IgemmParams& p, const StackGeom& g,
                          int smem_bytes, cudaStream_t stream) {
    static int attr_smem = 0;
    if (attr_smem < smem_bytes) {
        cudaError_t e = cudaFuncSetAttribute(conv_stack_kernel<R, STREAM>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
        if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(conv_stack): %s", cudaGetErrorString(e));
        attr_smem = smem_bytes;
    }
    const int sms = device_sm_count();
    const int grid = p.num_tiles < sms ? p.num_tiles : sms;
    b2::launch_k(conv_stack_kernel<R, STREAM>, grid, STREAM ? kStackThreads + 32 : kStackThreads, smem_bytes, stream, mapA, mapB, p, g);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("conv_stack launch: %s", cudaGetErrorString(e));
    return 0;
}

int launch_conv_stack(const ConvIgemmArgs& a, cudaStream_t stream) {
    IgemmParams p = a.p;
    StackGeom g;
    memset(&g, 0, sizeof(g));
    if (!stack_rect(p, g.kh, g.kw, g.dy_min, g.dx_min)) return set_error("conv_stack: taps are not a full rectangle");
    bool streamed = false;
    const int R = stack_pick_r(a, g.kh, g.kw, &streamed);
    if (R == 0) return set_error("conv_stack: not applicable to %dx%d, Cin=%d, %dx%d taps", p.H, p.W, p.cin, g.kh, g.kw);
    for (int t = 0; t < p.num_taps; ++t) g.wt[p.dy[t] - g.dy_min][p.dx[t] - g.dx_min] = p.wtap[t];
    g.nshift = g.kh + R - 1;
    for (int s = 0; s < g.nshift; ++s) {
        const int jlo = s - (g.kh - 1) > 0 ? s - (g.kh - 1) : 0;
        const int jhi = s < R - 1 ? s : R - 1;
        g.jlo[s] = uint8_t(jlo);
        g.cnt[s] = uint8_t(jhi - jlo + 1);
        g.i0[s] = uint8_t(g.kh - 1 - (s - jlo));
    }
    const int cbs = p.cin / 64;
    const int th = kStackGroups * R;
    g.HH = th + g.kh - 1;
    g.HW = kStackTW + g.kw - 1;
    g.a_block = ((g.HH * g.HW * 128) + 1023) & ~1023;
    g.b_bytes = (g.kh * g.kw * cbs + (R - 1)) * 64 * 128;
    if (streamed) {
        g.slot_bytes = (g.kh + R - 1) * 64 * 128;
        g.b_slots = (kStackSmemMax - 4096 - 2 * g.a_block) / g.slot_bytes;
        if (g.b_slots > 4) g.b_slots = 4;
        if (g.b_slots > g.kw * cbs) g.b_slots = g.kw * cbs < 2 ? 2 : g.kw * cbs;
        g.b_bytes = g.b_slots * g.slot_bytes;
    }
    g.a_stages = (kStackSmemMax - 4096 - g.b_bytes) / g.a_block;
    if (g.a_stages > 4) g.a_stages = 4;
    if (g.a_stages < 2) return set_error("conv_stack: no room for double buffering");
    const int smem_bytes = g.b_bytes + g.a_stages * g.a_block + 1024 + 3072;   // alignment slack + barriers, epilogue vectors, statistics
    p.TB = 1; p.TH = th; p.TW = kStackTW;
    p.tiles_x = (p.W + kStackTW - 1) / kStackTW;
    p.tiles_y = (p.H + th - 1) / th;
    p.tiles_b = p.B;
    p.tiles_n = 1;
    p.num_tiles = p.tiles_x * p.tiles_y * p.tiles_b;

    CUtensorMap mapA, mapB;
    {
        cuuint64_t dims[4] = {cuuint64_t(a.xC), cuuint64_t(a.xW), cuuint64_t(a.xH), cuuint64_t(a.xB)};
        cuuint64_t strides[3] = {cuuint64_t(a.xC) * 2, cuuint64_t(a.xW) * a.xC * 2, cuuint64_t(a.xH) * a.xW * a.xC * 2};
        cuuint32_t box[4] = {64, cuuint32_t(g.HW), cuuint32_t(g.HH), 1};
        cuuint32_t estr[4] = {1, 1, 1, 1};
        int rc = encode_tensor_map(&mapA, a.x, 4, dims, strides, box, estr);
        if (rc) return rc;
    }
    {
        cuuint64_t dims[2] = {cuuint64_t(a.w_k), cuuint64_t(a.w_rows)};
        cuuint64_t strides[1] = {cuuint64_t(a.w_k) * 2};
        cuuint32_t box[2] = {64, 64};
        cuuint32_t estr[2] = {1, 1};
        int rc = encode_tensor_map(&mapB, a.w, 2, dims, strides, box, estr);
        if (rc) return rc;
    }
    if (streamed)
        return R == 3 ? launch_stack_r<3, true>(mapA, mapB, p, g, smem_bytes, stream) : launch_stack_r<2, true>(mapA, mapB, p, g, smem_bytes, stream);
    return R == 3 ? launch_stack_r<3, false>(mapA, mapB, p, g, smem_bytes, stream) : launch_stack_r<2, false>(mapA, mapB, p, g, smem_bytes, stream);
}

}  // namespace b2
