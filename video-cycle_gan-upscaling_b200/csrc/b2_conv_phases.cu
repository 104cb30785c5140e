// b2_conv_phases.cu — Conv2DTranspose(stride 2, 'same') with Cout = 256 (upsampling_block, model.py:70-75) as ONE
// persistent tcgen05 kernel over all four sub-pixel phases.
//
// A stride-2 transposed conv is four ordinary convs, one per output parity (py, px), each using the filter taps of that
// parity (k=3: 4/2/2/1 taps, k=5: 9/6/6/4).  Launching them separately (b2_api.cu run_transposed) reads the input four
// times (8.5 GB instead of 2.1 GB per 8 frames for the second block) and leaves the 1- and 2-tap phases bound by their
// epilogue (LSU store issue / HBM write), not by the tensor core.  Here one CTA keeps the (16+h) x (8+h) halo tile of
// an input tile in shared memory (all Cin/64 channel blocks) and runs the phases back to back into the two TMEM
// accumulator stages: the epilogue of phase i overlaps the MMAs of phase i+1, the input crosses HBM once, and the
// store-heavy short phases average out against the long one.
//   * A ring: halo blocks of one 64-channel slice each; a tile owns Cin/64 consecutive ring slots which are released
//     block by block during its LAST phase, so the next tile's blocks stream in while that phase still runs.  The phase
//     with the most taps is scheduled last and a short one first, which gives the loads the longest head start.
//   * B ring: [256 x 64] weight blocks streamed from L2 in (phase, channel block, tap) order.
//   * taps are shifted UMMA-descriptor views of the halo block, as in b2_conv_halo.cu.
//   * Output: each epilogue warp owns a 4 x 8-pixel block of the tile and 128 of the 256 channels.  Writing a pixel row
//     per thread straight to global memory costs one LSU wavefront per (thread, 32-byte piece) -- ncu showed the LSU data
//     pipe 93 % busy on the 64 -> 256 block, half of it these stores.  Instead the warp packs 32 channels x 32 pixels into
//     a 64-byte-swizzled 2 KB staging block (conflict-free 16-byte shared stores, 4 wavefronts per instruction) and one
//     lane hands it to the TMA (cp.async.bulk.tensor store through a per-phase tensor map whose x / y strides skip the
//     other parities).  Used when the staging blocks fit next to the rings and the layer has no gate / residual / fp32
//     output; otherwise the shared register epilogue (b2_conv_epilogue.cuh) stores directly.
// Warp roles: 0 = TMA producer (input), 1 = MMA issuer + TMEM owner, 2..9 = epilogue (lane quarter = warp % 4, column
// half = (warp - 2) / 4), 10 = TMA producer (weights).
#include <stdlib.h>
#include <string.h>

#include "b2_conv_epilogue.cuh"
#include "b2_conv_igemm.cuh"
#include "b2_host.h"
#include "b2_ptx.cuh"

namespace b2 {

constexpr int kPhThreads = 352;
constexpr int kPhTH = 16, kPhTW = 8;
constexpr int kPhBN = 256;
constexpr int kPhBBlock = kPhBN * 128;     // one [256 x 64] weight block
constexpr int kPhBStages = 3;
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
};

constexpr int kPhStageBytes = 32 * 64;     // 32 pixels x 32 channels x 2 bytes
struct alignas(64) PhaseOutMaps { CUtensorMap m[4]; };   // per phase (execution order): [B][H][W][256] view of one output parity

// TMA-store epilogue of one (tile, phase) for one epilogue warp: 128 accumulator columns of 32 pixel rows, in chunks of
// 32 channels.  Staging block = TMA box {32 channels, 8 x, 4 y} with the 64-byte swizzle: pixel row `lane` is 64 bytes,
// its 16-byte piece j sits at piece j ^ ((lane >> 1) & 3), which also makes the warp's shared stores conflict-free.
template <int ACT>
__device__ __forceinline__ void phase_epilogue_tma(const IgemmParams& p, uint32_t taddr, int half, int lane,
                                                   const float* s_scale, const float* s_shift, const float* s_alpha,
                                                   uint8_t* my_stage, int nbuf, int& sbuf, const CUtensorMap* map,
                                                   int bx, int by, int b) {
    const bool unit_scale = p.scale == nullptr;
    const uint32_t swz = uint32_t((lane >> 1) & 3);
#pragma unroll 1
    for (int c0 = 0; c0 < 128; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(taddr + c0, r);
        if (lane == 0) {                       // the store that last used this staging block must have read it
            if (nbuf >= 2) tma_store_wait_read<1>(); else tma_store_wait_read<0>();
        }
        __syncwarp();
        tmem_ld_wait();
        const int n = 128 * half + c0;
        const uint32_t sc_a = smem_u32(s_scale + n), sh_a = smem_u32(s_shift + n), al_a = smem_u32(s_alpha + n);
        const uint32_t row_a = smem_u32(my_stage + sbuf * kPhStageBytes) + uint32_t(lane) * 64u;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            uint32_t o[4];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int g4 = 2 * j + h;
                float4 sc = make_float4(1.f, 1.f, 1.f, 1.f);
                if (!unit_scale) sc = lds_f4(sc_a + 16 * g4);
                const float4 sh = lds_f4(sh_a + 16 * g4);
                float4 al = make_float4(0.f, 0.f, 0.f, 0.f);
                if constexpr (ACT == ACT_PRELU) al = lds_f4(al_a + 16 * g4);
                const float v0 = apply_act_t<ACT>(fmaf(__uint_as_float(r[4 * g4 + 0]), sc.x, sh.x), p.leaky, al.x);
                const float v1 = apply_act_t<ACT>(fmaf(__uint_as_float(r[4 * g4 + 1]), sc.y, sh.y), p.leaky, al.y);
                const float v2 = apply_act_t<ACT>(fmaf(__uint_as_float(r[4 * g4 + 2]), sc.z, sh.z), p.leaky, al.z);
                const float v3 = apply_act_t<ACT>(fmaf(__uint_as_float(r[4 * g4 + 3]), sc.w, sh.w), p.leaky, al.w);
                o[2 * h] = pack2(v0, v1, p.f16);
                o[2 * h + 1] = pack2(v2, v3, p.f16);
            }
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_a + ((uint32_t(j) ^ swz) << 4)), "r"(o[0]),
                         "r"(o[1]), "r"(o[2]), "r"(o[3]) : "memory");
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
            tma_store_4d(map, my_stage + sbuf * kPhStageBytes, n, bx, by, b);
            tma_store_commit();
        }
        if (++sbuf >= nbuf) sbuf = 0;
    }
}

__global__ void __launch_bounds__(kPhThreads, 1)
conv_phases_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
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
    uint64_t* bfull = bars + 16;       // [4]
    uint64_t* bempty = bars + 20;      // [4]
    uint64_t* tfull = bars + 24;       // [2]
    uint64_t* tempty = bars + 26;      // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 28);
    float* s_scale = reinterpret_cast<float*>(tail + 256);
    float* s_shift = s_scale + kPhBN;
    float* s_alpha = s_shift + kPhBN;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&mapA);
        tma_prefetch_desc(&mapB);
        for (int s = 0; s < 8; ++s) { mbar_init(&afull[s], 1); mbar_init(&aempty[s], 1); }
        for (int s = 0; s < 4; ++s) { mbar_init(&bfull[s], 1); mbar_init(&bempty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&tfull[s], 1); mbar_init(&tempty[s], 8); }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<512>(tmem_slot);
    for (int c = threadIdx.x; c < kPhBN; c += kPhThreads) {
        s_scale[c] = p.scale ? p.scale[c] : 1.f;
        s_shift[c] = p.shift ? p.shift[c] : 0.f;
        s_alpha[c] = p.alpha ? p.alpha[c] : 0.f;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    const int cbs = p.cin >> 6;
    const uint32_t a_bytes = uint32_t(g.HH * g.HW) * 128u;

    if (warp == 0) {
        // ------------------------------------------------------------ producer: input halo blocks
        int sa = 0; uint32_t pa = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
            int t = tile;
            const int xt = t % p.tiles_x; t /= p.tiles_x;
            const int yt = t % p.tiles_y; t /= p.tiles_y;
            const int x0 = xt * kPhTW + g.dx_min, y0 = yt * kPhTH + g.dy_min;
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
    } else if (warp == 10) {
        // ------------------------------------------------------------ producer: weight blocks in (phase, cb, tap) order
        int sb = 0; uint32_t pb = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x)
            for (int ph = 0; ph < g.nph; ++ph)
                for (int cb = 0; cb < cbs; ++cb)
                    for (int t = g.ph_first[ph]; t < g.ph_first[ph + 1]; ++t) {
                        mbar_wait(&bempty[sb], pb ^ 1);
                        if (lane == 0) {
                            mbar_expect_tx(&bfull[sb], kPhBBlock);
                            tma_load_2d(smemB + sb * kPhBBlock, &mapB, &bfull[sb], int(g.wtap[t]) * p.cin + cb * 64, 0);
                        }
                        __syncwarp();
                        if (++sb == kPhBStages) { sb = 0; pb ^= 1; }
                    }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        const uint32_t idesc = p.f16 ? make_idesc_f16(128, kPhBN) : make_idesc_bf16(128, kPhBN);
        const uint32_t a_hi = sdesc_hi_k_sw128(uint32_t(g.HW) * 128u);
        constexpr uint32_t b_hi = sdesc_hi_k_sw128(1024);
        const uint32_t a_lo0 = sdesc_lo(smem_u32(smemA)), b_lo0 = sdesc_lo(smem_u32(smemB));
        int sa0 = 0; uint32_t pa0 = 0;       // ring slot / parity of this tile's first block
        int sb = 0; uint32_t pb = 0;
        int as = 0; uint32_t aph = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
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
                            umma_f16_lohi(tmem_d, a_lo, a_hi, b_lo, b_hi, idesc, (cb | (t - t0)) != 0);
                            umma_f16_lohi(tmem_d, a_lo + 2, a_hi, b_lo + 2, b_hi, idesc, 1);
                            umma_f16_lohi(tmem_d, a_lo + 4, a_hi, b_lo + 4, b_hi, idesc, 1);
                            umma_f16_lohi(tmem_d, a_lo + 6, a_hi, b_lo + 6, b_hi, idesc, 1);
                            umma_commit(&bempty[sb]);
                            if (t == t1 - 1) {
                                if (ph == g.nph - 1) umma_commit(&aempty[sa]);
                                if (cb == cbs - 1) umma_commit(&tfull[as]);
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
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
            int t = tile;
            const int xt = t % p.tiles_x; t /= p.tiles_x;
            const int yt = t % p.tiles_y; t /= p.tiles_y;
            const int gx = xt * kPhTW + r_tx, gy = yt * kPhTH + r_ty;
            const bool valid = gx < p.W && gy < p.H;
            for (int ph = 0; ph < g.nph; ++ph) {
                mbar_wait(&tfull[as], aph);
                tc_fence_after();
                const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(as * kPhBN + 128 * half);
                if (g.stage_bufs) {
                    const int bx = xt * kPhTW, by = yt * kPhTH + 4 * q;
                    switch (p.act) {
                        case ACT_RELU: phase_epilogue_tma<ACT_RELU>(p, taddr, half, lane, s_scale, s_shift, s_alpha, my_stage, g.stage_bufs, sbuf, &mapO.m[ph], bx, by, t); break;
                        case ACT_LEAKY: phase_epilogue_tma<ACT_LEAKY>(p, taddr, half, lane, s_scale, s_shift, s_alpha, my_stage, g.stage_bufs, sbuf, &mapO.m[ph], bx, by, t); break;
                        case ACT_PRELU: phase_epilogue_tma<ACT_PRELU>(p, taddr, half, lane, s_scale, s_shift, s_alpha, my_stage, g.stage_bufs, sbuf, &mapO.m[ph], bx, by, t); break;
                        case ACT_TANH: phase_epilogue_tma<ACT_TANH>(p, taddr, half, lane, s_scale, s_shift, s_alpha, my_stage, g.stage_bufs, sbuf, &mapO.m[ph], bx, by, t); break;
                        default: phase_epilogue_tma<ACT_NONE>(p, taddr, half, lane, s_scale, s_shift, s_alpha, my_stage, g.stage_bufs, sbuf, &mapO.m[ph], bx, by, t); break;
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
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tempty[as]);
                if (++as == 2) { as = 0; aph ^= 1; }
            }
        }
        if (g.stage_bufs && lane == 0) tma_store_wait_all<0>();   // the staging blocks must outlive the stores' reads
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

// `ph[0..nph)`: the per-phase arguments b2_api.cu builds for a stride-2 transposed conv (same input, weights and
// epilogue; taps, output parity and pixel grid per phase).
bool conv_phases_applicable(const ConvIgemmArgs* ph, int nph) {
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

int launch_conv_phases(const ConvIgemmArgs* ph, int nph, cudaStream_t stream) {
    if (!conv_phases_applicable(ph, nph)) return set_error("conv_phases: not applicable");
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
    const int tail_bytes = 256 + 3 * kPhBN * 4;
    const int avail = kPhSmemMax - 1024 - kPhBStages * kPhBBlock - tail_bytes;
    // staging blocks for the TMA-store epilogue: two per epilogue warp if they fit beside the minimum input ring, else one
    const bool can_stage = !p.out_fp32 && !p.gate && !p.residual && p.OH == 2 * p.H && p.OW == 2 * p.W;
    static const int force_bufs = B2_PROF_ENV("B2_PHASES_STAGE_BUFS") ? atoi(B2_PROF_ENV("B2_PHASES_STAGE_BUFS")) : -1;
    g.stage_bufs = 0;
    if (can_stage) {
        for (int nb = 2; nb >= 1; --nb)
            if (avail - 8 * nb * kPhStageBytes >= (cbs + 1) * g.a_block) { g.stage_bufs = nb; break; }
        if (force_bufs >= 0 && force_bufs < g.stage_bufs) g.stage_bufs = force_bufs;
    }
    g.a_stages = (avail - 8 * g.stage_bufs * kPhStageBytes) / g.a_block;
    if (g.a_stages > 8) g.a_stages = 8;
    if (g.a_stages < cbs + 1) return set_error("conv_phases: %d input blocks of %d bytes do not fit", cbs, g.a_block);
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
        cuuint32_t box[2] = {64, cuuint32_t(kPhBN)};
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
        cudaError_t e = cudaFuncSetAttribute(conv_phases_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
        if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(conv_phases): %s", cudaGetErrorString(e));
        attr_smem = smem_bytes;
    }
    const int sms = device_sm_count();
    const int grid = p.num_tiles < sms ? p.num_tiles : sms;
    b2::launch_k(conv_phases_kernel, grid, kPhThreads, smem_bytes, stream, mapA, mapB, mapO, p, g);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("conv_phases launch: %s", cudaGetErrorString(e));
    return 0;
}

}  // namespace b2
