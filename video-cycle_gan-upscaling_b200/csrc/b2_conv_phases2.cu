// b2_conv_phases2.cu — the CTA-pair (tcgen05 cta_group::2) form of the fused-phase Conv2DTranspose kernel
// (b2_conv_phases.cu; upsampling_block, model.py:70-75).
//
// Why: ncu on the 256 -> 256 block showed the single-CTA kernel bound by L2 -> shared-memory traffic, not by the tensor
// core: every 128-pixel tile streams the whole weight set (one [256 x 64] block = 32 KB per four MMAs = 64 B/clk/SM
// at full MMA rate, against ~44 B/clk/SM achieved), tensor pipe 64 % active.  Here two CTAs of a cluster (one TPC) run
// ONE M = 256 MMA per K step: each CTA keeps its own 128-pixel halo tile (A rows 0..127 / 128..255) and its own TMEM
// accumulator, and loads only HALF of every weight block (N rows 128 * rank ...), which the pair's MMA reads from both
// shared memories.  Weight traffic per SM halves (32 B/clk at full rate) and the ring holds 6 blocks instead of 3.
//   * rank 0 issues all MMAs and commits; commits are multicast to the same barrier offset in both CTAs;
//   * both CTAs' TMA loads credit the rank-0 "full" barriers (cp.async.bulk.tensor ... cta_group::2);
//   * the epilogue warps of both CTAs arrive on rank 0's "accumulator empty" barriers (count 16).
// Everything else (halo-tile taps as shifted descriptor views, phase order, TMA-store epilogue) is as in the single-CTA
// kernel, which stays as the fallback when the geometry does not fit.
#include <stdlib.h>
#include <string.h>

#include "b2_conv_epilogue.cuh"
#include "b2_conv_igemm.cuh"
#include "b2_host.h"
#include "b2_ptx.cuh"

namespace b2 {
namespace pair {

constexpr int kPhThreads = 352;
constexpr int kPh2BHalf = 128;                    // weight rows per CTA
constexpr int kPhTH = 16, kPhTW = 8;
constexpr int kPhBN = 256;
constexpr int kPhBBlock = kPh2BHalf * 128;  // this CTA's half of one [256 x 64] weight block
constexpr int kPhBStages = 4;     // 4 x 16 KB: leaves room for two staging blocks per epilogue warp at Cin = 256
constexpr int kPhMaxTaps = 32;             // k = 5: 9 + 6 + 6 + 4 = 25
constexpr int kPhSmemMax = 227 * 1024;

struct PhaseGeom {
    int HH, HW;              // halo tile (pixels)
    int dy_min, dx_min;      // most negative input offset over all phases
    int a_block;             // bytes per halo block, multiple of 1024
    int a_stages;            // >= Cin/64 + 1
    int nph;                 // phases (4), in execution order
    int ph_first[5];         // tap range [ph_first[i], ph_first[i+1]) of phase i in the arrays below
    int8_t oy[4], ox[4];     // output parity of the phase
    int8_t dy[kPhMaxTaps], dx[kPhMaxTaps];
    int16_t wtap[kPhMaxTaps];
    int stage_bufs;          // 0: direct global stores; 1 / 2: staging blocks per epilogue warp for TMA stores
    int dbg;                 // profiling only (env B2_PH2_DEBUG): bit 0 = epilogue releases the accumulator without reading it
};

constexpr int kPhStageBytes = 32 * 64;     // 32 pixels x 32 channels x 2 bytes
struct alignas(64) PhaseOutMaps { CUtensorMap m[4]; };   // per phase (execution order): [B][H][W][256] view of one output parity

// TMA-store epilogue of one (tile, phase) for one epilogue warp: 128 accumulator columns of 32 pixel rows, in chunks of
// 32 channels.  Staging block = TMA box {32 channels, 8 x, 4 y} with the 64-byte swizzle: pixel row `lane` is 64 bytes,
// its 16-byte piece j sits at piece j ^ ((lane >> 1) & 3), which also makes the warp's shared stores conflict-free.
template <int ACT, bool UNIT>
__device__ __forceinline__ void phase_chunk_tma(const IgemmParams& p, const uint32_t (&r)[32], int n, int lane,
                                                const float* s_scale, const float* s_shift, const float* s_alpha,
                                                uint8_t* my_stage, int nbuf, int& sbuf, const CUtensorMap* map, int bx,
                                                int by, int b) {
    if (lane == 0) {                       // the store that last used this staging block must have read it
        if (nbuf >= 2) tma_store_wait_read<1>(); else tma_store_wait_read<0>();
    }
    __syncwarp();
    const uint32_t swz = uint32_t((lane >> 1) & 3);
    const uint32_t sc_a = smem_u32(s_scale + n), sh_a = smem_u32(s_shift + n), al_a = smem_u32(s_alpha + n);
    const uint32_t row_a = smem_u32(my_stage + sbuf * kPhStageBytes) + uint32_t(lane) * 64u;
    // all per-channel vector loads first (independent, back to back), then the arithmetic, then the four staging stores:
    // the asm statements keep their program order, so interleaving them would expose one LDS latency per group
    float4 sh[8], sc[UNIT ? 1 : 8], al[ACT == ACT_PRELU ? 8 : 1];
#pragma unroll
    for (int g4 = 0; g4 < 8; ++g4) {
        sh[g4] = lds_f4(sh_a + 16 * g4);
        if constexpr (!UNIT) sc[g4] = lds_f4(sc_a + 16 * g4);
        if constexpr (ACT == ACT_PRELU) al[g4] = lds_f4(al_a + 16 * g4);
    }
    uint32_t o[16];
#pragma unroll
    for (int g4 = 0; g4 < 8; ++g4) {
        float v[4] = {__uint_as_float(r[4 * g4 + 0]), __uint_as_float(r[4 * g4 + 1]), __uint_as_float(r[4 * g4 + 2]),
                      __uint_as_float(r[4 * g4 + 3])};
        const float s4[4] = {sh[g4].x, sh[g4].y, sh[g4].z, sh[g4].w};
        if constexpr (UNIT) {
#pragma unroll
            for (int e = 0; e < 4; ++e) v[e] += s4[e];
        } else {
            const float c4[4] = {sc[g4].x, sc[g4].y, sc[g4].z, sc[g4].w};
#pragma unroll
            for (int e = 0; e < 4; ++e) v[e] = fmaf(v[e], c4[e], s4[e]);
        }
        if constexpr (ACT == ACT_LEAKY) {
            // max(v, leaky * v) == (v > 0 ? v : leaky * v) for 0 <= leaky <= 1 (the launcher checks), two instructions
#pragma unroll
            for (int e = 0; e < 4; ++e) v[e] = fmaxf(v[e], v[e] * p.leaky);
        } else if constexpr (ACT == ACT_PRELU) {
            const float a4[4] = {al[g4].x, al[g4].y, al[g4].z, al[g4].w};
#pragma unroll
            for (int e = 0; e < 4; ++e) v[e] = apply_act_t<ACT>(v[e], p.leaky, a4[e]);
        } else {
#pragma unroll
            for (int e = 0; e < 4; ++e) v[e] = apply_act_t<ACT>(v[e], p.leaky, 0.f);
        }
        o[2 * g4] = pack2(v[0], v[1], p.f16);
        o[2 * g4 + 1] = pack2(v[2], v[3], p.f16);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_a + ((uint32_t(j) ^ swz) << 4)), "r"(o[4 * j]),
                     "r"(o[4 * j + 1]), "r"(o[4 * j + 2]), "r"(o[4 * j + 3]) : "memory");
    fence_proxy_async();
    __syncwarp();
    if (lane == 0) {
        tma_store_4d(map, my_stage + sbuf * kPhStageBytes, n, bx, by, b);
        tma_store_commit();
    }
    if (++sbuf >= nbuf) sbuf = 0;
}

// The four 32-column chunks of this warp's half of the accumulator.  The accumulator is handed back to the MMA warp
// (rank 0's `tempty`) as soon as the last TMEM load has landed -- before the last chunk's arithmetic and stores.
template <int ACT, bool UNIT>
__device__ __forceinline__ void phase_epilogue_tma_u(const IgemmParams& p, uint32_t taddr, int half, int lane,
                                                     const float* s_scale, const float* s_shift, const float* s_alpha,
                                                     uint8_t* my_stage, int nbuf, int& sbuf, const CUtensorMap* map,
                                                     int bx, int by, int b, uint64_t* tempty_bar) {
    const int n0 = 128 * half;
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
        uint32_t r[32];
        tmem_ld32(taddr + 32 * c, r);
        tmem_ld_wait();
        if (c == 3) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_rank0(tempty_bar);
        }
        phase_chunk_tma<ACT, UNIT>(p, r, n0 + 32 * c, lane, s_scale, s_shift, s_alpha, my_stage, nbuf, sbuf, map, bx, by, b);
    }
}
template <int ACT>
__device__ __forceinline__ void phase_epilogue_tma(const IgemmParams& p, uint32_t taddr, int half, int lane,
                                                   const float* s_scale, const float* s_shift, const float* s_alpha,
                                                   uint8_t* my_stage, int nbuf, int& sbuf, const CUtensorMap* map,
                                                   int bx, int by, int b, uint64_t* tempty_bar) {
    if (p.scale == nullptr)
        phase_epilogue_tma_u<ACT, true>(p, taddr, half, lane, s_scale, s_shift, s_alpha, my_stage, nbuf, sbuf, map, bx, by, b, tempty_bar);
    else
        phase_epilogue_tma_u<ACT, false>(p, taddr, half, lane, s_scale, s_shift, s_alpha, my_stage, nbuf, sbuf, map, bx, by, b, tempty_bar);
}

__global__ void __launch_bounds__(kPhThreads, 1)
conv_phases2_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                   const __grid_constant__ PhaseOutMaps mapO, const __grid_constant__ IgemmParams p,
                   const __grid_constant__ PhaseGeom g) {
    b2::pdl_entry();
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smemB = smem;
    uint8_t* smemA = smem + kPhBStages * kPhBBlock;
    uint8_t* stage = smemA + g.a_stages * g.a_block;            // 8 warps x stage_bufs x 2 KB, 1024-aligned
    uint8_t* tail = stage + 8 * g.stage_bufs * kPhStageBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(tail);
    uint64_t* afull = bars;            // [8]
    uint64_t* aempty = bars + 8;       // [8]
    uint64_t* bfull = bars + 16;       // [8]
    uint64_t* bempty = bars + 24;      // [8]
    uint64_t* tfull = bars + 32;       // [2]
    uint64_t* tempty = bars + 34;      // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 36);
    float* s_scale = reinterpret_cast<float*>(tail + 512);
    float* s_shift = s_scale + kPhBN;
    float* s_alpha = s_shift + kPhBN;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&mapA);
        tma_prefetch_desc(&mapB);
        for (int s = 0; s < 8; ++s) { mbar_init(&afull[s], 1); mbar_init(&aempty[s], 1); }
        for (int s = 0; s < 8; ++s) { mbar_init(&bfull[s], 1); mbar_init(&bempty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&tfull[s], 1); mbar_init(&tempty[s], 16); }   // 8 epilogue warps x 2 CTAs
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc_2cta<512>(tmem_slot);
    for (int c = threadIdx.x; c < kPhBN; c += kPhThreads) {
        s_scale[c] = p.scale ? p.scale[c] : 1.f;
        s_shift[c] = p.shift ? p.shift[c] : 0.f;
        s_alpha[c] = p.alpha ? p.alpha[c] : 0.f;
    }
    tc_fence_before();
    cluster_sync_all();          // both CTAs' barriers are initialised before any remote arrive / multicast commit
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    const int rank = int(cluster_ctarank());
    const int pair0 = blockIdx.x >> 1, npairs = gridDim.x >> 1;
    const int num_pairs = (p.num_tiles + 1) >> 1;
    const int cbs = p.cin >> 6;
    const uint32_t a_bytes = uint32_t(g.HH * g.HW) * 128u;

    if (warp == 0) {
        // ------------------------------------------------------------ producer: input halo blocks
        int sa = 0; uint32_t pa = 0;
        for (int pt = pair0; pt < num_pairs; pt += npairs) {
            int t = min(2 * pt + rank, p.num_tiles - 1);      // an odd tile count: the last pair's second CTA redoes the last tile
            const int xt = t % p.tiles_x; t /= p.tiles_x;
            const int yt = t % p.tiles_y; t /= p.tiles_y;
            const int x0 = xt * kPhTW + g.dx_min, y0 = yt * kPhTH + g.dy_min;
            for (int cb = 0; cb < cbs; ++cb) {
                mbar_wait(&aempty[sa], pa ^ 1);
                if (lane == 0) {
                    if (rank == 0) mbar_expect_tx(&afull[sa], 2 * a_bytes);     // both CTAs' halo blocks
                    tma2_load_4d(smemA + sa * g.a_block, &mapA, &afull[sa], cb * 64, x0, y0, t);
                }
                __syncwarp();
                if (++sa == g.a_stages) { sa = 0; pa ^= 1; }
            }
        }
    } else if (warp == 10) {
        // ------------------------------------------------------------ producer: weight blocks in (phase, cb, tap) order
        int sb = 0; uint32_t pb = 0;
        for (int pt = pair0; pt < num_pairs; pt += npairs)
            for (int ph = 0; ph < g.nph; ++ph)
                for (int cb = 0; cb < cbs; ++cb)
                    for (int t = g.ph_first[ph]; t < g.ph_first[ph + 1]; ++t) {
                        mbar_wait(&bempty[sb], pb ^ 1);
                        if (lane == 0) {
                            if (rank == 0) mbar_expect_tx(&bfull[sb], 2 * kPhBBlock);
                            tma2_load_2d(smemB + sb * kPhBBlock, &mapB, &bfull[sb], int(g.wtap[t]) * p.cin + cb * 64, kPh2BHalf * rank);
                        }
                        __syncwarp();
                        if (++sb == kPhBStages) { sb = 0; pb ^= 1; }
                    }
    } else if (warp == 1 && rank == 0) {
        // ------------------------------------------------------------ MMA issuer (rank 0 only, M = 256 over the pair)
        const uint32_t idesc = p.f16 ? make_idesc_f16(256, kPhBN) : make_idesc_bf16(256, kPhBN);
        const uint32_t a_hi = sdesc_hi_k_sw128(uint32_t(g.HW) * 128u);
        constexpr uint32_t b_hi = sdesc_hi_k_sw128(1024);
        const uint32_t a_lo0 = sdesc_lo(smem_u32(smemA)), b_lo0 = sdesc_lo(smem_u32(smemB));
        int sa0 = 0; uint32_t pa0 = 0;       // ring slot / parity of this tile's first block
        int sb = 0; uint32_t pb = 0;
        int as = 0; uint32_t aph = 0;
        for (int pt = pair0; pt < num_pairs; pt += npairs) {
            for (int ph = 0; ph < g.nph; ++ph) {
                mbar_wait(&tempty[as], aph ^ 1);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + uint32_t(as * kPhBN);
                int sa = sa0; uint32_t pa = pa0;
                for (int cb = 0; cb < cbs; ++cb) {
                    if (ph == 0) {                               // later phases find the block already resident
                        mbar_wait(&afull[sa], pa);
                        tc_fence_after();
                    }
                    const uint32_t a_lo_blk = a_lo0 + uint32_t(sa) * uint32_t(g.a_block >> 4);
                    const int t0 = g.ph_first[ph], t1 = g.ph_first[ph + 1];
                    for (int t = t0; t < t1; ++t) {
                        mbar_wait(&bfull[sb], pb);
                        tc_fence_after();
                        const uint32_t b_lo = b_lo0 + uint32_t(sb) * (kPhBBlock >> 4);
                        const uint32_t a_lo = a_lo_blk + uint32_t((int(g.dy[t]) - g.dy_min) * g.HW + (int(g.dx[t]) - g.dx_min)) * 8u;
                        if (elect_one()) {
                            umma2_f16_lohi(tmem_d, a_lo, a_hi, b_lo, b_hi, idesc, (cb | (t - t0)) != 0);
                            umma2_f16_lohi(tmem_d, a_lo + 2, a_hi, b_lo + 2, b_hi, idesc, 1);
                            umma2_f16_lohi(tmem_d, a_lo + 4, a_hi, b_lo + 4, b_hi, idesc, 1);
                            umma2_f16_lohi(tmem_d, a_lo + 6, a_hi, b_lo + 6, b_hi, idesc, 1);
                            umma2_commit_mc(&bempty[sb]);
                            if (t == t1 - 1) {
                                if (ph == g.nph - 1) umma2_commit_mc(&aempty[sa]);
                                if (cb == cbs - 1) umma2_commit_mc(&tfull[as]);
                            }
                        }
                        __syncwarp();
                        if (++sb == kPhBStages) { sb = 0; pb ^= 1; }
                    }
                    if (++sa == g.a_stages) { sa = 0; pa ^= 1; }
                }
                if (++as == 2) { as = 0; aph ^= 1; }
                if (ph == g.nph - 1) { sa0 = sa; pa0 = pa; }
            }
        }
    } else if (warp >= 2 && warp <= 9) {
        // ------------------------------------------------------------ epilogue: (tile, phase) -> interleaved output pixels
        const int q = warp & 3, half = (warp - 2) >> 2;
        const int row = q * 32 + lane;
        const int r_tx = row % kPhTW, r_ty = row / kPhTW;
        uint8_t* my_stage = stage + (warp - 2) * g.stage_bufs * kPhStageBytes;
        int sbuf = 0;
        int as = 0; uint32_t aph = 0;
        for (int pt = pair0; pt < num_pairs; pt += npairs) {
            int t = 2 * pt + rank;
            const bool tile_ok = t < p.num_tiles;
            t = min(t, p.num_tiles - 1);
            const int xt = t % p.tiles_x; t /= p.tiles_x;
            const int yt = t % p.tiles_y; t /= p.tiles_y;
            const int gx = xt * kPhTW + r_tx, gy = yt * kPhTH + r_ty;
            const bool valid = tile_ok && gx < p.W && gy < p.H;
            for (int ph = 0; ph < g.nph; ++ph) {
                mbar_wait(&tfull[as], aph);
                tc_fence_after();
                const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(as * kPhBN + 128 * half);
                if (!tile_ok || (g.dbg & 1)) {
                    // duplicate of the last tile: nothing to store
                } else if (g.stage_bufs) {
                    const int bx = xt * kPhTW, by = yt * kPhTH + 4 * q;
                    switch (p.act) {
                        case ACT_RELU: phase_epilogue_tma<ACT_RELU>(p, taddr, half, lane, s_scale, s_shift, s_alpha, my_stage, g.stage_bufs, sbuf, &mapO.m[ph], bx, by, t, &tempty[as]); break;
                        case ACT_LEAKY: phase_epilogue_tma<ACT_LEAKY>(p, taddr, half, lane, s_scale, s_shift, s_alpha, my_stage, g.stage_bufs, sbuf, &mapO.m[ph], bx, by, t, &tempty[as]); break;
                        case ACT_PRELU: phase_epilogue_tma<ACT_PRELU>(p, taddr, half, lane, s_scale, s_shift, s_alpha, my_stage, g.stage_bufs, sbuf, &mapO.m[ph], bx, by, t, &tempty[as]); break;
                        case ACT_TANH: phase_epilogue_tma<ACT_TANH>(p, taddr, half, lane, s_scale, s_shift, s_alpha, my_stage, g.stage_bufs, sbuf, &mapO.m[ph], bx, by, t, &tempty[as]); break;
                        default: phase_epilogue_tma<ACT_NONE>(p, taddr, half, lane, s_scale, s_shift, s_alpha, my_stage, g.stage_bufs, sbuf, &mapO.m[ph], bx, by, t, &tempty[as]); break;
                    }
                } else {
                    const size_t pix = (size_t(t) * p.OH + size_t(gy * 2 + g.oy[ph])) * p.OW + size_t(gx * 2 + g.ox[ph]);
                    const size_t obase = pix * size_t(p.out_ld) + size_t(128 * half);
                    switch (p.act) {
                        case ACT_RELU: epilogue_tile<128, ACT_RELU>(p, taddr, valid, 128 * half, obase, s_scale, s_shift, s_alpha); break;
                        case ACT_LEAKY: epilogue_tile<128, ACT_LEAKY>(p, taddr, valid, 128 * half, obase, s_scale, s_shift, s_alpha); break;
                        case ACT_PRELU: epilogue_tile<128, ACT_PRELU>(p, taddr, valid, 128 * half, obase, s_scale, s_shift, s_alpha); break;
                        case ACT_TANH: epilogue_tile<128, ACT_TANH>(p, taddr, valid, 128 * half, obase, s_scale, s_shift, s_alpha); break;
                        default: epilogue_tile<128, ACT_NONE>(p, taddr, valid, 128 * half, obase, s_scale, s_shift, s_alpha); break;
                    }
                }
                if (!tile_ok || (g.dbg & 1) || !g.stage_bufs) {        // the TMA-store epilogue has released it already
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_rank0(&tempty[as]);
                }
                if (++as == 2) { as = 0; aph ^= 1; }
            }
        }
        if (g.stage_bufs && lane == 0) tma_store_wait_all<0>();   // the staging blocks must outlive the stores' reads
    }

    tc_fence_before();
    cluster_sync_all();          // the pair's MMAs read both shared memories and write both TMEMs: leave together
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_2cta<512>(tmem_base);
    }
}

// `ph[0..nph)`: the per-phase arguments b2_api.cu builds for a stride-2 transposed conv (same input, weights and
// epilogue; taps, output parity and pixel grid per phase).
}  // namespace pair
using namespace pair;

bool conv_phases2_applicable(const ConvIgemmArgs* ph, int nph) {
    if (nph != 4) return false;
    const IgemmParams& p0 = ph[0].p;
    if (p0.n_valid != kPhBN || p0.out_ld != kPhBN || ph[0].w_rows != kPhBN) return false;
    if (p0.residual || p0.gate || p0.out_fp32) return false;       // their TMA-store epilogue carries neither
    if (ph[0].xC % 64 != 0 || p0.cin != ph[0].xC || p0.cin / 64 > 7) return false;
    if (p0.H < kPhTH || p0.W < kPhTW) return false;
    int total = 0;
    for (int i = 0; i < nph; ++i) {
        const IgemmParams& p = ph[i].p;
        if (p.H != p0.H || p.W != p0.W || p.B != p0.B || p.in_mul != 1 || p.out_sy != 2 || p.out_sx != 2) return false;
        if (p.num_taps < 1) return false;
        total += p.num_taps;
    }
    if (total > kPhMaxTaps) return false;
    const long covered = long((p0.H + kPhTH - 1) / kPhTH) * kPhTH * ((p0.W + kPhTW - 1) / kPhTW) * kPhTW;
    return long(p0.H) * p0.W * 100 >= covered * 88;
}

int launch_conv_phases2(const ConvIgemmArgs* ph, int nph, cudaStream_t stream) {
    if (!conv_phases2_applicable(ph, nph)) return set_error("conv_phases2: not applicable");
    IgemmParams p = ph[0].p;
    PhaseGeom g;
    memset(&g, 0, sizeof(g));
    // execution order: a short phase first, the longest last (its MMAs cover the next tile's input loads)
    int order[4] = {0, 1, 2, 3};
    for (int i = 0; i < 4; ++i)
        for (int j = i + 1; j < 4; ++j)
            if (ph[order[j]].p.num_taps < ph[order[i]].p.num_taps) { int t = order[i]; order[i] = order[j]; order[j] = t; }
    { int t = order[0]; order[0] = order[1]; order[1] = t; }
    int dy0 = 127, dy1 = -127, dx0 = 127, dx1 = -127, nt = 0;
    g.nph = 4;
    for (int i = 0; i < 4; ++i) {
        const IgemmParams& q = ph[order[i]].p;
        g.ph_first[i] = nt;
        g.oy[i] = int8_t(q.out_oy); g.ox[i] = int8_t(q.out_ox);
        for (int t = 0; t < q.num_taps; ++t, ++nt) {
            g.dy[nt] = q.dy[t]; g.dx[nt] = q.dx[t]; g.wtap[nt] = q.wtap[t];
            dy0 = q.dy[t] < dy0 ? q.dy[t] : dy0; dy1 = q.dy[t] > dy1 ? q.dy[t] : dy1;
            dx0 = q.dx[t] < dx0 ? q.dx[t] : dx0; dx1 = q.dx[t] > dx1 ? q.dx[t] : dx1;
        }
    }
    g.ph_first[4] = nt;
    g.dy_min = dy0; g.dx_min = dx0;
    g.HH = kPhTH + (dy1 - dy0);
    g.HW = kPhTW + (dx1 - dx0);
    g.a_block = ((g.HH * g.HW * 128) + 1023) & ~1023;
    const int cbs = p.cin / 64;
    const int tail_bytes = 512 + 3 * kPhBN * 4;
    const int avail = kPhSmemMax - 1024 - kPhBStages * kPhBBlock - tail_bytes;
    // staging blocks for the TMA-store epilogue: two per epilogue warp if they fit beside the minimum input ring, else one
    const bool can_stage = !p.out_fp32 && !p.gate && !p.residual && p.OH == 2 * p.H && p.OW == 2 * p.W &&
                           (p.act != ACT_LEAKY || (p.leaky >= 0.f && p.leaky <= 1.f));
    static const int force_bufs = B2_PROF_ENV("B2_PHASES2_STAGE_BUFS") ? atoi(B2_PROF_ENV("B2_PHASES2_STAGE_BUFS")) : -1;
    static const int dbg = B2_PROF_ENV("B2_PH2_DEBUG") ? atoi(B2_PROF_ENV("B2_PH2_DEBUG")) : 0;
    g.dbg = dbg;
    g.stage_bufs = 0;
    if (can_stage) {
        for (int nb = 2; nb >= 1; --nb)
            if (avail - 8 * nb * kPhStageBytes >= (cbs + 1) * g.a_block) { g.stage_bufs = nb; break; }
        if (force_bufs >= 0 && force_bufs < g.stage_bufs) g.stage_bufs = force_bufs;
    }
    g.a_stages = (avail - 8 * g.stage_bufs * kPhStageBytes) / g.a_block;
    if (g.a_stages > 8) g.a_stages = 8;
    if (g.a_stages < cbs + 1) { set_error("conv_phases2: %d input blocks of %d bytes do not fit", cbs, g.a_block); return -2; }
    const int smem_bytes = 1024 + kPhBStages * kPhBBlock + g.a_stages * g.a_block + 8 * g.stage_bufs * kPhStageBytes + tail_bytes;
    p.TB = 1; p.TH = kPhTH; p.TW = kPhTW;
    p.tiles_x = (p.W + kPhTW - 1) / kPhTW;
    p.tiles_y = (p.H + kPhTH - 1) / kPhTH;
    p.tiles_b = p.B;
    p.tiles_n = 1;
    p.num_tiles = p.tiles_x * p.tiles_y * p.tiles_b;

    const ConvIgemmArgs& a = ph[0];
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
        cuuint32_t box[2] = {64, cuuint32_t(kPh2BHalf)};      // this CTA's half of the N rows
        cuuint32_t estr[2] = {1, 1};
        int rc = encode_tensor_map(&mapB, a.w, 2, dims, strides, box, estr);
        if (rc) return rc;
    }
    PhaseOutMaps mapO;
    memset(&mapO, 0, sizeof(mapO));
    if (g.stage_bufs) {
        for (int i = 0; i < 4; ++i) {
            const char* base = reinterpret_cast<const char*>(p.out) + (size_t(g.oy[i]) * p.OW + size_t(g.ox[i])) * size_t(p.out_ld) * 2;
            cuuint64_t dims[4] = {cuuint64_t(kPhBN), cuuint64_t(p.W), cuuint64_t(p.H), cuuint64_t(p.B)};
            cuuint64_t strides[3] = {cuuint64_t(p.out_ld) * 4, cuuint64_t(p.OW) * p.out_ld * 4, cuuint64_t(p.OH) * p.OW * p.out_ld * 2};
            cuuint32_t box[4] = {32, cuuint32_t(kPhTW), 4, 1};
            cuuint32_t estr[4] = {1, 1, 1, 1};
            int rc = encode_tensor_map(&mapO.m[i], base, 4, dims, strides, box, estr, 64);
            if (rc) return rc;
        }
    }
    static int attr_smem = 0;
    if (attr_smem < smem_bytes) {
        cudaError_t e = cudaFuncSetAttribute(conv_phases2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
        if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(conv_phases2): %s", cudaGetErrorString(e));
        attr_smem = smem_bytes;
    }
    const int sms = device_sm_count();
    int grid = 2 * ((p.num_tiles + 1) / 2);
    if (grid > sms) grid = sms & ~1;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(grid, 1, 1);
    cfg.blockDim = dim3(kPhThreads, 1, 1);
    cfg.dynamicSmemBytes = smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;       // see pdl_entry() in b2_host.h
    attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    cudaError_t e = cudaLaunchKernelEx(&cfg, conv_phases2_kernel, mapA, mapB, mapO, p, g);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("conv_phases2 launch: %s", cudaGetErrorString(e));
    return 0;
}

}  // namespace b2
