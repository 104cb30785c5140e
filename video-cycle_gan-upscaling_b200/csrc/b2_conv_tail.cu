// b2_conv_tail.cu — the generator's final Conv2D(3, 9x9, 'same') + tanh (reference: upscaling/upscaler/model.py:290-291)
// as a row-streaming tcgen05 kernel.
//
// A plain implicit GEMM wastes the tensor core here (N = 3 output channels) and re-reads each 256-channel input
// pixel 81 times.  Instead the filter ROWS are folded into the GEMM N dimension and the filter COLUMNS into K:
//     Q[y, x, (dy, co)] = sum_{dx, c} X[y, x + dx - pw, c] * W[dy, dx, c, co]          (N = KH*Cout = 27 -> 32)
//     out[y', x, co]    = tanh(bias[co] + sum_dy Q[y' + dy - ph, x, (dy, co)])
// One M tile is 128 consecutive pixels of ONE input row.  The row (plus KW-1 halo pixels) is loaded once by TMA as
// Cin/64 boxes of [128+KW-1 pixels x 64 channels] (128-byte swizzle); the dx shifts are expressed purely through the
// UMMA shared-memory descriptor start address (+128 B per pixel), so every input element is fetched exactly once.
// The sum over dy is a 9-row ring of fp32 partial output rows kept in shared memory by the epilogue warps while the
// CTA walks down a 128-pixel-wide strip of the image; a row is finished (bias, tanh, store) as soon as its last
// contributing input row has been consumed.  The flattened (image, strip, row) space is split evenly over the CTAs.
//
// Warp roles: 0 = TMA producer for input rows, 1 = MMA issuer (+TMEM alloc), 2..5 = epilogue, 6 = TMA producer for
// weight k-blocks.
#include "b2_conv_igemm.cuh"
#include "b2_host.h"
#include "b2_ptx.cuh"

namespace b2 {

constexpr int kTailThreads = 224;
constexpr int kTailMaxKW = 9;
constexpr int kTailRowsA = 128 + kTailMaxKW - 1;   // pixels per loaded row box
constexpr int kTailABlock = kTailRowsA * 128;      // 17408 B, a multiple of 1024
constexpr int kTailMaxCb = 4;                      // Cin <= 256
constexpr int kTailAStage = kTailMaxCb * kTailABlock;
constexpr int kTailAStages = 2;
constexpr int kTailBBlock = 32 * 128;              // N = 32 rows x 64 channels
constexpr int kTailBStages = 8;
constexpr int kTailMaxKH = 9;
constexpr int kTailRing = kTailMaxKH * 128 * 3 * 4;
constexpr int kTailSmem = kTailAStages * kTailAStage + kTailBStages * kTailBBlock + kTailRing + 512 + 1024;

struct TailParams {
    int B, H, W, cin;
    int KH, KW, ph, pw;   // filter and 'same' pad-before
    int cout;             // <= 3, KH*cout <= 32
    int strips;           // ceil(W / 128)
    long total_tiles;     // B * strips * H
    int act;              // ACT_TANH or ACT_NONE
    int f16;
    const float* bias;
    float* out;           // [B, H, W, cout] fp32
};

// A segment is a run of output rows [y0, y1) inside one (image, strip); its input rows are y0-ph .. y1-1+(KH-1-ph).
struct Seg { int y0, y1, img, x0; };
__device__ __forceinline__ bool seg_next(long& t, long t_end, const TailParams& p, Seg& s) {
    if (t >= t_end) return false;
    const long col = t / p.H;
    long seg_end = (col + 1) * p.H;
    if (seg_end > t_end) seg_end = t_end;
    s.y0 = int(t - col * p.H);
    s.y1 = s.y0 + int(seg_end - t);
    s.img = int(col / p.strips);
    s.x0 = int(col % p.strips) * 128;
    t = seg_end;
    return true;
}

template <int KH, int COUT>
__global__ void __launch_bounds__(kTailThreads, 1)
conv_tail_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                 const __grid_constant__ TailParams p) {
    b2::pdl_entry();
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smemA = smem;
    uint8_t* smemB = smem + kTailAStages * kTailAStage;
    float* ring = reinterpret_cast<float*>(smemB + kTailBStages * kTailBBlock);
    uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(ring) + kTailRing);
    uint64_t* afull = bars;                        // [2]
    uint64_t* aempty = bars + 2;                   // [2]
    uint64_t* bfull = bars + 4;                    // [8]
    uint64_t* bempty = bars + 12;                  // [8]
    uint64_t* tfull = bars + 20;                   // [2]
    uint64_t* tempty = bars + 22;                  // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 24);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&mapA);
        tma_prefetch_desc(&mapB);
        for (int s = 0; s < kTailAStages; ++s) { mbar_init(&afull[s], 1); mbar_init(&aempty[s], 1); }
        for (int s = 0; s < kTailBStages; ++s) { mbar_init(&bfull[s], 1); mbar_init(&bempty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&tfull[s], 1); mbar_init(&tempty[s], 4); }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<64>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

    // This CTA's contiguous range of output-row tiles in the flattened (image, strip, row) order.
    const long t_begin = p.total_tiles * blockIdx.x / gridDim.x;
    const long t_end = p.total_tiles * (blockIdx.x + 1) / gridDim.x;
    const int cbs = p.cin >> 6;
    const int kblocks = p.KW * cbs;
    // each CTA starts its K walk at a different weight block so the SMs do not all hit the same L2 lines at once
    const int kb_rot = int(blockIdx.x % uint32_t(kblocks));
    const int halo_lo = p.ph, halo_hi = KH - 1 - p.ph;   // input rows needed above / below an output row

    Seg sg;

    if (warp == 0) {
        // ---------------------------------------------------------------- input-row producer
        int st = 0; uint32_t ph = 0;
        for (long t = t_begin; seg_next(t, t_end, p, sg);) {
            const int y0 = sg.y0, y1 = sg.y1, img = sg.img, x0 = sg.x0;
            for (int r = y0 - halo_lo; r < y1 + halo_hi; ++r) {
                mbar_wait(&aempty[st], ph ^ 1);
                if (lane == 0) {
                    mbar_expect_tx(&afull[st], uint32_t(cbs) * kTailABlock);
                    for (int cb = 0; cb < cbs; ++cb)
                        tma_load_4d(smemA + st * kTailAStage + cb * kTailABlock, &mapA, &afull[st], cb * 64, x0 - p.pw, r, img);
                }
                __syncwarp();
                if (++st == kTailAStages) { st = 0; ph ^= 1; }
            }
            (void)img; (void)x0;
        }
    } else if (warp == 6) {
        // ---------------------------------------------------------------- weight k-block producer
        int st = 0; uint32_t ph = 0;
        for (long t = t_begin; seg_next(t, t_end, p, sg);) {
            const int y0 = sg.y0, y1 = sg.y1, img = sg.img, x0 = sg.x0;
            const int nrows = (y1 + halo_hi) - (y0 - halo_lo);
            for (int r = 0; r < nrows; ++r)
                for (int i = 0; i < kblocks; ++i) {
                    int kb = i + kb_rot;
                    if (kb >= kblocks) kb -= kblocks;
                    mbar_wait(&bempty[st], ph ^ 1);
                    if (lane == 0) {
                        mbar_expect_tx(&bfull[st], kTailBBlock);
                        tma_load_2d(smemB + st * kTailBBlock, &mapB, &bfull[st], kb * 64, 0);
                    }
                    __syncwarp();
                    if (++st == kTailBStages) { st = 0; ph ^= 1; }
                }
            (void)img; (void)x0;
        }
    } else if (warp == 1) {
        // ---------------------------------------------------------------- MMA issuer
        const uint32_t idesc = p.f16 ? make_idesc_f16(128, 32) : make_idesc_bf16(128, 32);
        constexpr uint32_t desc_hi = sdesc_hi_k_sw128(1024);
        const uint32_t a_lo0 = sdesc_lo(smem_u32(smemA)), b_lo0 = sdesc_lo(smem_u32(smemB));
        int sa = 0; uint32_t pa = 0;
        int sb = 0; uint32_t pb = 0;
        int as = 0; uint32_t aph = 0;
        for (long t = t_begin; seg_next(t, t_end, p, sg);) {
            const int y0 = sg.y0, y1 = sg.y1, img = sg.img, x0 = sg.x0;
            const int nrows = (y1 + halo_hi) - (y0 - halo_lo);
            for (int r = 0; r < nrows; ++r) {
                mbar_wait(&tempty[as], aph ^ 1);
                mbar_wait(&afull[sa], pa);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + uint32_t(as * 32);
                const uint32_t a_lo_stage = a_lo0 + uint32_t(sa) * (kTailAStage >> 4);
                for (int i = 0; i < kblocks; ++i) {
                        int kb = i + kb_rot;
                        if (kb >= kblocks) kb -= kblocks;
                        const int dx = kb / cbs, cb = kb - dx * cbs;
                        mbar_wait(&bfull[sb], pb);
                        tc_fence_after();
                        if (elect_one()) {
                            const uint32_t a_lo = a_lo_stage + uint32_t(cb) * (kTailABlock >> 4) + uint32_t(dx) * 8u;
                            const uint32_t b_lo = b_lo0 + uint32_t(sb) * (kTailBBlock >> 4);
                            umma_f16_lohi(tmem_d, a_lo, desc_hi, b_lo, desc_hi, idesc, i != 0);
                            umma_f16_lohi(tmem_d, a_lo + 2, desc_hi, b_lo + 2, desc_hi, idesc, 1);
                            umma_f16_lohi(tmem_d, a_lo + 4, desc_hi, b_lo + 4, desc_hi, idesc, 1);
                            umma_f16_lohi(tmem_d, a_lo + 6, desc_hi, b_lo + 6, desc_hi, idesc, 1);
                            umma_commit(&bempty[sb]);
                            if (i == kblocks - 1) {
                                umma_commit(&aempty[sa]);
                                umma_commit(&tfull[as]);
                            }
                        }
                        __syncwarp();
                        if (++sb == kTailBStages) { sb = 0; pb ^= 1; }
                    }
                if (++sa == kTailAStages) { sa = 0; pa ^= 1; }
                if (++as == 2) { as = 0; aph ^= 1; }
            }
            (void)img; (void)x0;
        }
    } else if (warp >= 2 && warp <= 5) {
        // ---------------------------------------------------------------- epilogue: ring of partial output rows
        const int q = warp & 3;
        const int px = q * 32 + lane;                 // pixel within the strip == TMEM lane
        float* myring = ring + px * 3;                // ring[slot][px][c], slot stride 128*3
        float bias[COUT];
#pragma unroll
        for (int c = 0; c < COUT; ++c) bias[c] = p.bias ? p.bias[c] : 0.f;
        int as = 0; uint32_t aph = 0;
        for (long t = t_begin; seg_next(t, t_end, p, sg);) {
            const int y0 = sg.y0, y1 = sg.y1, img = sg.img, x0 = sg.x0;
            for (int s = 0; s < KH; ++s)
#pragma unroll
                for (int c = 0; c < COUT; ++c) myring[s * 384 + c] = 0.f;
            const int gx = x0 + px;
            for (int r = y0 - halo_lo; r < y1 + halo_hi; ++r) {
                mbar_wait(&tfull[as], aph);
                tc_fence_after();
                uint32_t v[32];
                tmem_ld32(tmem_base + (uint32_t(q * 32) << 16) + uint32_t(as * 32), v);
                tmem_ld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tempty[as]);
                if (++as == 2) { as = 0; aph ^= 1; }
                // input row r feeds output rows r - dy + ph
#pragma unroll
                for (int dy = 0; dy < KH; ++dy) {
                    const int yo = r - dy + p.ph;
                    if (yo >= y0 && yo < y1) {
                        float* slot = myring + (yo % KH) * 384;
#pragma unroll
                        for (int c = 0; c < COUT; ++c) slot[c] += __uint_as_float(v[dy * COUT + c]);
                    }
                }
                // output row r - halo_hi has now received its last contribution
                const int yd = r - halo_hi;
                if (yd >= y0 && yd < y1) {
                    float* slot = myring + (yd % KH) * 384;
                    if (gx < p.W) {
                        float* o = p.out + ((size_t(img) * p.H + yd) * p.W + gx) * COUT;
#pragma unroll
                        for (int c = 0; c < COUT; ++c) {
                            const float a = slot[c] + bias[c];
                            o[c] = p.act == ACT_TANH ? tanhf(a) : a;
                        }
                    }
#pragma unroll
                    for (int c = 0; c < COUT; ++c) slot[c] = 0.f;
                }
            }
            (void)img; (void)x0;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<64>(tmem_base);
    }
}

int launch_conv_tail(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int KH, int KW,
                     const float* bias, int act, int f16, float* out, cudaStream_t stream) {
    if (Cin % 64 != 0 || Cin > 64 * kTailMaxCb) return set_error("conv_tail: Cin=%d unsupported", Cin);
    if (KH != 9 || Cout != 3 || KW > kTailMaxKW || (KW & 1) == 0)
        return set_error("conv_tail: filter %dx%d -> %d unsupported", KH, KW, Cout);
    TailParams p;
    p.B = B; p.H = H; p.W = W; p.cin = Cin;
    p.KH = KH; p.KW = KW; p.ph = (KH - 1) / 2; p.pw = (KW - 1) / 2;   // stride-1 'same' padding is symmetric
    p.cout = Cout;
    p.strips = (W + 127) / 128;
    p.total_tiles = long(B) * p.strips * H;
    p.act = act; p.f16 = f16; p.bias = bias; p.out = out;
    CUtensorMap mapA, mapB;
    {
        cuuint64_t dims[4] = {cuuint64_t(Cin), cuuint64_t(W), cuuint64_t(H), cuuint64_t(B)};
        cuuint64_t strides[3] = {cuuint64_t(Cin) * 2, cuuint64_t(W) * Cin * 2, cuuint64_t(H) * W * Cin * 2};
        cuuint32_t box[4] = {64, cuuint32_t(kTailRowsA), 1, 1};
        cuuint32_t estr[4] = {1, 1, 1, 1};
        int rc = encode_tensor_map(&mapA, x, 4, dims, strides, box, estr);
        if (rc) return rc;
    }
    {
        cuuint64_t dims[2] = {cuuint64_t(KW) * Cin, 32};
        cuuint64_t strides[1] = {cuuint64_t(KW) * Cin * 2};
        cuuint32_t box[2] = {64, 32};
        cuuint32_t estr[2] = {1, 1};
        int rc = encode_tensor_map(&mapB, w, 2, dims, strides, box, estr);
        if (rc) return rc;
    }
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(conv_tail_kernel<9, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTailSmem);
        if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(conv_tail): %s", cudaGetErrorString(e));
        attr_set = true;
    }
    long grid = device_sm_count();
    if (grid > p.total_tiles) grid = p.total_tiles;
    b2::launch_k(conv_tail_kernel<9, 3>, int(grid), kTailThreads, kTailSmem, stream, mapA, mapB, p);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("conv_tail launch: %s", cudaGetErrorString(e));
    return 0;
}

}  // namespace b2
