// b2_conv_igemm.cu — NHWC bf16 implicit-GEMM convolution on tcgen05 tensor cores (sm_100a).
//
// GEMM view:  D[pixel, cout] = sum_{tap, cin}  X[pixel + offset(tap), cin] * Wp[cout, tap*Cin + cin]
//   M = 128 output pixels (a TB x TH x TW box of the pixel grid), N = BN output channels, K walked tap by tap in
//   blocks of 64 input channels.
// Data movement: one TMA 4-D box load per (tap, channel block) lands the shifted pixel box in shared memory as
//   128 rows x 128 bytes with the 128-byte swizzle; out-of-image coordinates are zero-filled by TMA, which IS the
//   "same" padding of the reference's Conv2D (upscaling/upscaler/model.py:19, 22, 275, 283, 290). Stride-2 convs
//   use the tensor map's element strides; the four sub-pixel phases of Conv2DTranspose(stride 2) (model.py:72) are
//   four launches with their own tap lists and an interleaved output mapping.
// Compute: one thread issues tcgen05.mma (M=128, N=BN, K=16) with both operands from shared memory and the fp32
//   accumulator in TMEM (double-buffered so the epilogue of tile i overlaps the main loop of tile i+1).
// Epilogue: 4 warps read TMEM (tcgen05.ld 32x32b), apply scale/shift (bias or folded BatchNorm), activation
//   (ReLU / LeakyReLU / PReLU / tanh), residual add, and store bf16 or fp32 NHWC.
// Warp roles: warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer, warps 2..5 = epilogue.
#include <stdlib.h>

#include "b2_conv_epilogue.cuh"
#include "b2_conv_igemm.cuh"
#include "b2_host.h"
#include "b2_ptx.cuh"

namespace b2 {

constexpr int kThreads = 192;
constexpr int kATileBytes = 128 * 128;  // 128 pixel rows x 64 bf16

template <int BN>
struct IgemmCfg {
    static constexpr int kBTileBytes = BN * 128;
    static constexpr int kStageBytes = kATileBytes + kBTileBytes;
    static constexpr int kStagesRaw = (200 * 1024) / kStageBytes;
    static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
    static constexpr int kTmemCols = (2 * BN) < 32 ? 32 : 2 * BN;
    static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/ + 3 * kMaxChan * 4;
};

template <int BN>
__global__ void __launch_bounds__(kThreads, 1)
conv_igemm_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                  const __grid_constant__ IgemmParams p) {
    b2::pdl_entry();
    using Cfg = IgemmCfg<BN>;
    constexpr int STAGES = Cfg::kStages;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smemA = smem;
    uint8_t* smemB = smem + STAGES * kATileBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::kStageBytes);
    uint64_t* full = bars;                    // [STAGES] TMA -> MMA
    uint64_t* empty = bars + STAGES;          // [STAGES] MMA -> TMA
    uint64_t* tfull = bars + 2 * STAGES;      // [2] MMA -> epilogue
    uint64_t* tempty = bars + 2 * STAGES + 2; // [2] epilogue -> MMA
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
    float* s_scale = reinterpret_cast<float*>(smem + STAGES * Cfg::kStageBytes + 256);
    float* s_shift = s_scale + kMaxChan;
    float* s_alpha = s_shift + kMaxChan;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&mapA);
        tma_prefetch_desc(&mapB);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tfull[s], 1);
            mbar_init(&tempty[s], 4);
        }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
    for (int c = threadIdx.x; c < p.tiles_n * BN; c += kThreads) {
        const bool cv = c < p.n_valid;
        s_scale[c] = (p.scale && cv) ? p.scale[c] : 1.f;
        s_shift[c] = (p.shift && cv) ? p.shift[c] : 0.f;
        s_alpha[c] = (p.alpha && cv) ? p.alpha[c] : 0.f;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

    const int cin_blocks = p.cin >> 6;
    const int num_kb = p.num_taps * cin_blocks;
    const int rows = p.TB * p.TH * p.TW;
    const int kb_rot = int(blockIdx.x % uint32_t(num_kb));
    const uint32_t tx_bytes = uint32_t(rows) * 128u + uint32_t(Cfg::kBTileBytes);

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        int stage = 0;
        uint32_t phase = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
            int t = tile;
            const int nt = t % p.tiles_n; t /= p.tiles_n;
            const int xt = t % p.tiles_x; t /= p.tiles_x;
            const int yt = t % p.tiles_y; t /= p.tiles_y;
            const int bt = t;
            const int x0 = xt * p.TW * p.in_mul, y0 = yt * p.TH * p.in_mul, b0 = bt * p.TB, n0 = nt * BN;
            // Every CTA walks the same K blocks; starting each CTA at a different block (the accumulation order is
            // free) keeps the 148 SMs from hammering the same few L2 lines of the weight matrix at the same time.
            for (int i = 0; i < num_kb; ++i) {
                int kb = i + kb_rot;
                if (kb >= num_kb) kb -= num_kb;
                const int tap = kb / cin_blocks, cb = kb - tap * cin_blocks;
                const int cx = x0 + p.dx[tap], cy = y0 + p.dy[tap];
                const int kbase = int(p.wtap[tap]) * p.cin;
                mbar_wait(&empty[stage], phase ^ 1);
                if (lane == 0) {
                    mbar_expect_tx(&full[stage], tx_bytes);
                    tma_load_4d(smemA + stage * kATileBytes, &mapA, &full[stage], cb * 64, cx, cy, b0);
                    tma_load_2d(smemB + stage * Cfg::kBTileBytes, &mapB, &full[stage], kbase + cb * 64, n0);
                }
                __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        const uint32_t idesc = p.f16 ? make_idesc_f16(128, BN) : make_idesc_bf16(128, BN);
        constexpr uint32_t desc_hi = sdesc_hi_k_sw128(1024);
        const uint32_t a_lo0 = sdesc_lo(smem_u32(smemA)), b_lo0 = sdesc_lo(smem_u32(smemB));
        int stage = 0;
        uint32_t phase = 0;
        int as = 0;
        uint32_t aphase = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
            mbar_wait(&tempty[as], aphase ^ 1);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + uint32_t(as * BN);
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(&full[stage], phase);
                tc_fence_after();
                // elect.sync (not `lane == 0`): ptxas then knows a single thread runs the block and keeps the descriptors
                // in uniform registers, so the four UTCHMMA issue back to back
                if (elect_one()) {
                    const uint32_t a_lo = a_lo0 + uint32_t(stage) * (kATileBytes >> 4);
                    const uint32_t b_lo = b_lo0 + uint32_t(stage) * (Cfg::kBTileBytes >> 4);
                    umma_f16_lohi(tmem_d, a_lo, desc_hi, b_lo, desc_hi, idesc, kb != 0);
                    umma_f16_lohi(tmem_d, a_lo + 2, desc_hi, b_lo + 2, desc_hi, idesc, 1);
                    umma_f16_lohi(tmem_d, a_lo + 4, desc_hi, b_lo + 4, desc_hi, idesc, 1);
                    umma_f16_lohi(tmem_d, a_lo + 6, desc_hi, b_lo + 6, desc_hi, idesc, 1);
                    umma_commit(&empty[stage]);
                    if (kb == num_kb - 1) umma_commit(&tfull[as]);
                }
                __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
            if (++as == 2) { as = 0; aphase ^= 1; }
        }
    } else {
        // ------------------------------------------------------------ epilogue (warps 2..5)
        const int q = warp & 3;            // TMEM lane quarter this warp may read
        const int row = q * 32 + lane;     // row of the 128-row tile == TMEM lane
        const int tw = p.TW, th = p.TH;
        const int r_tx = row % tw;
        const int r_ty = (row / tw) % th;
        const int r_tb = row / (tw * th);
        int as = 0;
        uint32_t aphase = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
            int t = tile;
            const int nt = t % p.tiles_n; t /= p.tiles_n;
            const int xt = t % p.tiles_x; t /= p.tiles_x;
            const int yt = t % p.tiles_y; t /= p.tiles_y;
            const int bt = t;
            const int gx = xt * tw + r_tx, gy = yt * th + r_ty, gb = bt * p.TB + r_tb;
            const bool valid = (row < rows) && gx < p.W && gy < p.H && gb < p.B;
            const int n0 = nt * BN;
            const size_t pix = (size_t(gb) * p.OH + size_t(gy * p.out_sy + p.out_oy)) * p.OW +
                               size_t(gx * p.out_sx + p.out_ox);
            const size_t obase = pix * size_t(p.out_ld) + n0;

            mbar_wait(&tfull[as], aphase);
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
            if (++as == 2) { as = 0; aphase ^= 1; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<Cfg::kTmemCols>(tmem_base);
    }
}

// ---------------------------------------------------------------------------------------------- host side

static void choose_tile(int B, int H, int W, int& TB, int& TH, int& TW, int ext_max = 128) {
    double best = -1.0;
    TB = TH = TW = 1;
    const int wmax = W < ext_max ? W : ext_max;    // ext_max: a strided TMA box spans tile * stride <= 256 elements
    for (int tw = 1; tw <= wmax; ++tw) {
        int th = 128 / tw;
        if (th > H) th = H;
        if (th > ext_max) th = ext_max;
        if (th < 1) continue;
        int tb = 128 / (tw * th);
        if (tb > B) tb = B;
        if (tb < 1) tb = 1;
        const long tiles = long((W + tw - 1) / tw) * ((H + th - 1) / th) * ((B + tb - 1) / tb);
        double eff = double(B) * H * W / (double(tiles) * 128.0);
        if ((tw & 7) == 0) eff += 1e-6;  // tie-break: rows of 8 pixels keep 1024-byte atoms inside one image row
        eff += 1e-9 * tw;
        if (eff > best) { best = eff; TB = tb; TH = th; TW = tw; }
    }
}

template <int BN>
static int launch_bn(const CUtensorMap& mapA, const CUtensorMap& mapB, const IgemmParams& p, int num_sms,
                     cudaStream_t stream) {
    using Cfg = IgemmCfg<BN>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(conv_igemm_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Cfg::kSmemBytes);
        if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(igemm): %s", cudaGetErrorString(e));
        attr_set = true;
    }
    const int grid = p.num_tiles < num_sms ? p.num_tiles : num_sms;
    b2::launch_k(conv_igemm_kernel<BN>, grid, kThreads, Cfg::kSmemBytes, stream, mapA, mapB, p);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("conv_igemm launch: %s", cudaGetErrorString(e));
    return 0;
}

static int launch_conv_igemm_nostats(const ConvIgemmArgs& a, cudaStream_t stream);

int launch_conv_igemm(const ConvIgemmArgs& a, cudaStream_t stream) {
    if (!a.p.stats_part) return launch_conv_igemm_nostats(a, stream);
    // Train-mode BatchNorm statistics of the output: fused into the row-stacked kernel's epilogue where that kernel runs
    // (the residual trunk), otherwise one statistics pass over the finished tensor -- same contract for the caller.
    if (a.p.out_fp32 || a.p.out_sy != 1 || a.p.out_sx != 1 || a.p.out_ld != a.p.n_valid)
        return set_error("conv: BatchNorm statistics need a dense 16-bit output");
    if (conv_stack_applicable(a)) return launch_conv_stack(a, stream);
    ConvIgemmArgs b = a;
    b.p.stats_part = nullptr;
    if (int rc = launch_conv_igemm_nostats(b, stream)) return rc;
    return launch_bn_stats_all_rows(a.p.out, long(a.p.B) * a.p.OH * a.p.OW, a.p.n_valid, a.p.f16, a.p.stats_part, stream);
}

static int launch_conv_igemm_nostats(const ConvIgemmArgs& a, cudaStream_t stream) {
    static const bool no_halo = B2_PROF_ENV("B2_NO_HALO") != nullptr;   // A/B switch for profiling the two main loops
    static const bool no_stack = B2_PROF_ENV("B2_NO_STACK") != nullptr;
    if (!no_halo && !no_stack && conv_stack_applicable(a)) return launch_conv_stack(a, stream);
    if (!no_halo && conv_halo_applicable(a)) return launch_conv_halo(a, stream);
    IgemmParams p = a.p;
    if (a.xC % 64 != 0 || p.cin != a.xC) return set_error("igemm: Cin=%d must be a multiple of 64", a.xC);
    if (p.num_taps < 1 || p.num_taps > kMaxTaps) return set_error("igemm: bad tap count %d", p.num_taps);
    if (p.in_mul < 1 || p.in_mul > 3) return set_error("igemm: in_mul must be 1, 2 or 3");
    choose_tile(p.B, p.H, p.W, p.TB, p.TH, p.TW, p.in_mul <= 2 ? 128 : 256 / p.in_mul);
    int BN = 256;
    if (p.n_valid <= 16) BN = 16;
    else if (p.n_valid <= 32) BN = 32;
    else if (p.n_valid <= 64) BN = 64;
    else if (p.n_valid <= 128) BN = 128;
    p.tiles_x = (p.W + p.TW - 1) / p.TW;
    p.tiles_y = (p.H + p.TH - 1) / p.TH;
    p.tiles_b = (p.B + p.TB - 1) / p.TB;
    p.tiles_n = (p.n_valid + BN - 1) / BN;
    p.num_tiles = p.tiles_x * p.tiles_y * p.tiles_b * p.tiles_n;
    if (p.tiles_n * BN > kMaxChan) return set_error("igemm: Cout=%d exceeds %d", p.n_valid, kMaxChan);

    CUtensorMap mapA, mapB;
    {
        cuuint64_t dims[4] = {cuuint64_t(a.xC), cuuint64_t(a.xW), cuuint64_t(a.xH), cuuint64_t(a.xB)};
        cuuint64_t strides[3] = {cuuint64_t(a.xC) * 2, cuuint64_t(a.xW) * a.xC * 2,
                                 cuuint64_t(a.xH) * a.xW * a.xC * 2};
        cuuint32_t box[4] = {64, cuuint32_t(p.TW * p.in_mul), cuuint32_t(p.TH * p.in_mul), cuuint32_t(p.TB)};
        cuuint32_t estr[4] = {1, cuuint32_t(p.in_mul), cuuint32_t(p.in_mul), 1};
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
    const int sms = device_sm_count();
    switch (BN) {
        case 16: return launch_bn<16>(mapA, mapB, p, sms, stream);
        case 32: return launch_bn<32>(mapA, mapB, p, sms, stream);
        case 64: return launch_bn<64>(mapA, mapB, p, sms, stream);
        case 128: return launch_bn<128>(mapA, mapB, p, sms, stream);
        default: return launch_bn<256>(mapA, mapB, p, sms, stream);
    }
}

}  // namespace b2
