// b2_conv_halo.cu — stride-1 NHWC implicit-GEMM convolution on tcgen05 with the input tile (plus filter halo) staged
// ONCE in shared memory and every filter tap expressed as a shifted view of it.
//
// The plain kernel (b2_conv_igemm.cu) issues one TMA box per tap, so a 3x3 conv pulls every input pixel through
// L2 -> shared memory nine times; for the 64-channel residual-block convs (model.py:19, 22, 283) that traffic, not the
// tensor core, sets the speed.  Here an output tile is 16 rows x 8 pixels (M = 128).  For each block of 64 input
// channels one TMA box brings the (16 + kh - 1) x (8 + kw - 1) pixel halo tile into shared memory as rows of 128 bytes
// (128-byte swizzle).  The A operand of tap (dy, dx) is that block read through a UMMA descriptor whose start address
// is advanced by ((dy - dy_min) * HW + (dx - dx_min)) * 128 bytes and whose 8-row-group stride (SBO) is the halo row
// pitch HW * 128: the hardware applies the swizzle to absolute shared-memory address bits, so any 128-byte-aligned
// shifted view of a TMA-written block is a valid operand (measured: tools/b2_probe.cu "shift").
// Weights: when all taps of the layer fit (<= 96 KB, e.g. 3x3x64x64 = 72 KB) they are loaded once per CTA and stay
// resident; otherwise they stream through a ring of [BN x 64] blocks, each CTA starting at a different tap so the
// 148 SMs do not hit the same L2 lines together.
//
// Warp roles: 0 = TMA producer (input halo blocks, resident weights), 1 = MMA issuer + TMEM owner, 2..5 = epilogue,
// 6 = TMA producer for streamed weight blocks.
#include "b2_conv_epilogue.cuh"
#include "b2_conv_igemm.cuh"
#include "b2_host.h"
#include "b2_ptx.cuh"

namespace b2 {

constexpr int kHaloThreads = 224;
constexpr int kHaloTH = 16, kHaloTW = 8;
constexpr int kHaloSmemBudget = 205 * 1024;   // operands + rings; barriers and per-channel vectors come on top
constexpr int kHaloResidentMax = 96 * 1024;

struct HaloGeom {
    int HH, HW;          // halo tile (pixels)
    int dy_min, dx_min;  // most negative tap offsets
    int a_block;         // bytes per halo block, multiple of 1024
    int a_stages;
    int b_stages;        // streamed mode
    int resident;        // 1: all weight blocks resident in shared memory
    int b_bytes;         // resident region or ring size
};

template <int BN>
__global__ void __launch_bounds__(kHaloThreads, 1)
conv_halo_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                 const __grid_constant__ IgemmParams p, const __grid_constant__ HaloGeom g) {
    b2::pdl_entry();
    constexpr int kBBlock = BN * 128;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smemB = smem;
    uint8_t* smemA = smem + g.b_bytes;
    uint8_t* tail = smemA + g.a_stages * g.a_block;
    uint64_t* bars = reinterpret_cast<uint64_t*>(tail);
    uint64_t* afull = bars;            // [8]
    uint64_t* aempty = bars + 8;       // [8]
    uint64_t* bfull = bars + 16;       // [16] (bfull[0] doubles as the "resident weights landed" barrier)
    uint64_t* bempty = bars + 32;      // [16]
    uint64_t* tfull = bars + 48;       // [2]
    uint64_t* tempty = bars + 50;      // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 52);
    uint32_t* s_tapoff = reinterpret_cast<uint32_t*>(tail + 512);   // [kMaxTaps] descriptor-lo offset of each tap's view
    float* s_scale = reinterpret_cast<float*>(tail + 1024);
    float* s_shift = s_scale + kMaxChan;
    float* s_alpha = s_shift + kMaxChan;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    constexpr uint32_t kTmemCols = (2 * BN) < 32 ? 32 : 2 * BN;
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&mapA);
        tma_prefetch_desc(&mapB);
        for (int s = 0; s < 8; ++s) { mbar_init(&afull[s], 1); mbar_init(&aempty[s], 1); }
        for (int s = 0; s < 16; ++s) { mbar_init(&bfull[s], 1); mbar_init(&bempty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&tfull[s], 1); mbar_init(&tempty[s], 4); }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<kTmemCols>(tmem_slot);
    for (int t = threadIdx.x; t < p.num_taps; t += kHaloThreads)
        s_tapoff[t] = uint32_t((int(p.dy[t]) - g.dy_min) * g.HW + (int(p.dx[t]) - g.dx_min)) * 8u;   // 128 B per pixel = 8 x 16 B
    for (int c = threadIdx.x; c < p.tiles_n * BN; c += kHaloThreads) {
        const bool cv = c < p.n_valid;
        s_scale[c] = (p.scale && cv) ? p.scale[c] : 1.f;
        s_shift[c] = (p.shift && cv) ? p.shift[c] : 0.f;
        s_alpha[c] = (p.alpha && cv) ? p.alpha[c] : 0.f;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);   // warp-uniform for the MMA issue path

    const int cbs = p.cin >> 6;
    const int taps = p.num_taps;
    const int tap_rot = int(blockIdx.x % uint32_t(taps));
    const uint32_t a_bytes = uint32_t(g.HH * g.HW) * 128u;

    if (warp == 0) {
        // ------------------------------------------------------------ producer: resident weights + input halo blocks
        if (g.resident && lane == 0) {
            mbar_expect_tx(&bfull[0], uint32_t(taps * cbs) * kBBlock);
            for (int t = 0; t < taps; ++t)
                for (int cb = 0; cb < cbs; ++cb)
                    tma_load_2d(smemB + (t * cbs + cb) * kBBlock, &mapB, &bfull[0], int(p.wtap[t]) * p.cin + cb * 64, 0);
        }
        __syncwarp();
        int sa = 0; uint32_t pa = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
            int t = tile / p.tiles_n;
            const int xt = t % p.tiles_x; t /= p.tiles_x;
            const int yt = t % p.tiles_y; t /= p.tiles_y;
            const int x0 = xt * kHaloTW + g.dx_min, y0 = yt * kHaloTH + g.dy_min;
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
    } else if (warp == 6) {
        // ------------------------------------------------------------ producer: streamed weight blocks
        if (!g.resident) {
            int sb = 0; uint32_t pb = 0;
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
                const int n0 = (tile % p.tiles_n) * BN;
                for (int cb = 0; cb < cbs; ++cb)
                    for (int i = 0; i < taps; ++i) {
                        int t = i + tap_rot;
                        if (t >= taps) t -= taps;
                        mbar_wait(&bempty[sb], pb ^ 1);
                        if (lane == 0) {
                            mbar_expect_tx(&bfull[sb], kBBlock);
                            tma_load_2d(smemB + sb * kBBlock, &mapB, &bfull[sb], int(p.wtap[t]) * p.cin + cb * 64, n0);
                        }
                        __syncwarp();
                        if (++sb == g.b_stages) { sb = 0; pb ^= 1; }
                    }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        const uint32_t idesc = p.f16 ? make_idesc_f16(128, BN) : make_idesc_bf16(128, BN);
        const uint32_t a_hi = sdesc_hi_k_sw128(uint32_t(g.HW) * 128u);
        constexpr uint32_t b_hi = sdesc_hi_k_sw128(1024);
        const uint32_t a_lo0 = sdesc_lo(smem_u32(smemA)), b_lo0 = sdesc_lo(smem_u32(smemB));
        int sa = 0; uint32_t pa = 0;
        int sb = 0; uint32_t pb = 0;
        int as = 0; uint32_t aph = 0;
        if (g.resident) mbar_wait(&bfull[0], 0);
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
            mbar_wait(&tempty[as], aph ^ 1);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + uint32_t(as * BN);
            for (int cb = 0; cb < cbs; ++cb) {
                mbar_wait(&afull[sa], pa);
                tc_fence_after();
                const uint32_t a_lo_blk = a_lo0 + uint32_t(sa) * uint32_t(g.a_block >> 4);
                for (int i = 0; i < taps; ++i) {
                    int t = i + tap_rot;
                    if (t >= taps) t -= taps;
                    uint32_t b_lo;
                    if (g.resident) {
                        b_lo = b_lo0 + uint32_t(t * cbs + cb) * (kBBlock >> 4);
                    } else {
                        mbar_wait(&bfull[sb], pb);
                        tc_fence_after();
                        b_lo = b_lo0 + uint32_t(sb) * (kBBlock >> 4);
                    }
                    // tap offset straight from the kernel parameters (constant bank) so it stays warp-uniform
                    const uint32_t a_lo = a_lo_blk + uint32_t((int(p.dy[t]) - g.dy_min) * g.HW + (int(p.dx[t]) - g.dx_min)) * 8u;
                    if (elect_one()) {
                        umma_f16_lohi(tmem_d, a_lo, a_hi, b_lo, b_hi, idesc, (cb | i) != 0);
                        umma_f16_lohi(tmem_d, a_lo + 2, a_hi, b_lo + 2, b_hi, idesc, 1);
                        umma_f16_lohi(tmem_d, a_lo + 4, a_hi, b_lo + 4, b_hi, idesc, 1);
                        umma_f16_lohi(tmem_d, a_lo + 6, a_hi, b_lo + 6, b_hi, idesc, 1);
                        if (!g.resident) umma_commit(&bempty[sb]);
                        if (i == taps - 1) {
                            umma_commit(&aempty[sa]);
                            if (cb == cbs - 1) umma_commit(&tfull[as]);
                        }
                    }
                    __syncwarp();
                    if (!g.resident) { if (++sb == g.b_stages) { sb = 0; pb ^= 1; } }
                }
                if (++sa == g.a_stages) { sa = 0; pa ^= 1; }
            }
            if (++as == 2) { as = 0; aph ^= 1; }
        }
    } else if (warp >= 2 && warp <= 5) {
        // ------------------------------------------------------------ epilogue
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int r_tx = row % kHaloTW, r_ty = row / kHaloTW;
        int as = 0; uint32_t aph = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
            const int nt = tile % p.tiles_n;
            int t = tile / p.tiles_n;
            const int xt = t % p.tiles_x; t /= p.tiles_x;
            const int yt = t % p.tiles_y; t /= p.tiles_y;
            const int gx = xt * kHaloTW + r_tx, gy = yt * kHaloTH + r_ty, gb = t;
            const bool valid = gx < p.W && gy < p.H;
            const int n0 = nt * BN;
            const size_t pix = (size_t(gb) * p.OH + size_t(gy * p.out_sy + p.out_oy)) * p.OW + size_t(gx * p.out_sx + p.out_ox);
            const size_t obase = pix * size_t(p.out_ld) + n0;
            mbar_wait(&tfull[as], aph);
            tc_fence_after();
            const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(as * BN);
            switch (p.act) {
                case ACT_RELU: epilogue_tile<BN, ACT_RELU>(p, taddr, valid, n0, obase, s_scale, s_shift, s_alpha); break;
                case ACT_LEAKY: epilogue_tile<BN, ACT_LEAKY>(p, taddr, valid, n0, obase, s_scale, s_shift, s_alpha); break;
                case ACT_PRELU: epilogue_tile<BN, ACT_PRELU>(p, taddr, valid, n0, obase, s_scale, s_shift, s_alpha); break;
                case ACT_TANH: epilogue_tile<BN, ACT_TANH>(p, taddr, valid, n0, obase, s_scale, s_shift, s_alpha); break;
                default: epilogue_tile<BN, ACT_NONE>(p, taddr, valid, n0, obase, s_scale, s_shift, s_alpha); break;
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[as]);
            if (++as == 2) { as = 0; aph ^= 1; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<kTmemCols>(tmem_base);
    }
}

template <int BN>
static int launch_halo_bn(const CUtensorMap& mapA, const CUtensorMap& mapB, const IgemmParams& p, const HaloGeom& g,
                          int smem_bytes, cudaStream_t stream) {
    static int attr_smem = 0;
    if (attr_smem < smem_bytes) {
        cudaError_t e = cudaFuncSetAttribute(conv_halo_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
        if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(conv_halo): %s", cudaGetErrorString(e));
        attr_smem = smem_bytes;
    }
    const int sms = device_sm_count();
    const int grid = p.num_tiles < sms ? p.num_tiles : sms;
    b2::launch_k(conv_halo_kernel<BN>, grid, kHaloThreads, smem_bytes, stream, mapA, mapB, p, g);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("conv_halo launch: %s", cudaGetErrorString(e));
    return 0;
}

// True when the halo kernel applies: stride-1 taps, an image at least one tile large, Cout in {64, 128, 256, 512}.
bool conv_halo_applicable(const ConvIgemmArgs& a) {
    const IgemmParams& p = a.p;
    if (p.in_mul != 1 || p.num_taps < 2) return false;
    if (p.H < kHaloTH || p.W < kHaloTW) return false;
    if (p.n_valid != 64 && p.n_valid != 128 && p.n_valid != 256 && p.n_valid != 512) return false;
    int dy0 = 127, dy1 = -127, dx0 = 127, dx1 = -127;
    for (int t = 0; t < p.num_taps; ++t) {
        dy0 = p.dy[t] < dy0 ? p.dy[t] : dy0; dy1 = p.dy[t] > dy1 ? p.dy[t] : dy1;
        dx0 = p.dx[t] < dx0 ? p.dx[t] : dx0; dx1 = p.dx[t] > dx1 ? p.dx[t] : dx1;
    }
    if ((dy1 - dy0) > 8 || (dx1 - dx0) > 8) return false;
    // tile quantisation: 16x8 tiles must cover the image without wasting more than ~12 % of the MMA rows
    const long covered = long((p.H + kHaloTH - 1) / kHaloTH) * kHaloTH * ((p.W + kHaloTW - 1) / kHaloTW) * kHaloTW;
    return long(p.H) * p.W * 100 >= covered * 88;
}

int launch_conv_halo(const ConvIgemmArgs& a, cudaStream_t stream) {
    IgemmParams p = a.p;
    if (a.xC % 64 != 0 || p.cin != a.xC) return set_error("conv_halo: Cin=%d must be a multiple of 64", a.xC);
    int dy0 = 127, dy1 = -127, dx0 = 127, dx1 = -127;
    for (int t = 0; t < p.num_taps; ++t) {
        dy0 = p.dy[t] < dy0 ? p.dy[t] : dy0; dy1 = p.dy[t] > dy1 ? p.dy[t] : dy1;
        dx0 = p.dx[t] < dx0 ? p.dx[t] : dx0; dx1 = p.dx[t] > dx1 ? p.dx[t] : dx1;
    }
    const int BN = p.n_valid >= 256 ? 256 : p.n_valid;
    HaloGeom g;
    g.HH = kHaloTH + (dy1 - dy0);
    g.HW = kHaloTW + (dx1 - dx0);
    g.dy_min = dy0; g.dx_min = dx0;
    g.a_block = ((g.HH * g.HW * 128) + 1023) & ~1023;
    const int cbs = p.cin / 64;
    const int wbytes = p.num_taps * cbs * BN * 128;
    p.TB = 1; p.TH = kHaloTH; p.TW = kHaloTW;
    p.tiles_x = (p.W + kHaloTW - 1) / kHaloTW;
    p.tiles_y = (p.H + kHaloTH - 1) / kHaloTH;
    p.tiles_b = p.B;
    p.tiles_n = (p.n_valid + BN - 1) / BN;
    p.num_tiles = p.tiles_x * p.tiles_y * p.tiles_b * p.tiles_n;
    g.resident = (wbytes <= kHaloResidentMax && p.tiles_n == 1) ? 1 : 0;
    if (g.resident) {
        g.b_bytes = wbytes;
        g.b_stages = 0;
    } else {
        g.b_stages = (BN == 256) ? 3 : (BN == 128 ? 6 : 8);
        g.b_bytes = g.b_stages * BN * 128;
    }
    g.a_stages = (kHaloSmemBudget - g.b_bytes) / g.a_block;
    if (g.a_stages > 8) g.a_stages = 8;
    if (g.a_stages < 2) return set_error("conv_halo: halo block of %d bytes leaves no room for double buffering", g.a_block);
    const int smem_bytes = g.b_bytes + g.a_stages * g.a_block + 1024 + 3 * kMaxChan * 4 + 1024;

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
        cuuint32_t box[2] = {64, cuuint32_t(BN)};
        cuuint32_t estr[2] = {1, 1};
        int rc = encode_tensor_map(&mapB, a.w, 2, dims, strides, box, estr);
        if (rc) return rc;
    }
    switch (BN) {
        case 64: return launch_halo_bn<64>(mapA, mapB, p, g, smem_bytes, stream);
        case 128: return launch_halo_bn<128>(mapA, mapB, p, g, smem_bytes, stream);
        default: return launch_halo_bn<256>(mapA, mapB, p, g, smem_bytes, stream);
    }
}

}  // namespace b2
