// b2_conv_wgrad.cu — weight gradient of an NHWC convolution on tcgen05 tensor cores (sm_100a).
//
//   dW[tap][cx][cz] += sum over pixels p of the dZ grid:  X[p * s + d(tap), cx] * dZ[p, cz]
//
// which is what TensorFlow's conv2d_backprop_filter computes behind keras.layers.Conv2D in training (reference:
// upscaling/upscaler/model.py:19, 22, 275, 283, 290, 839-871) and, with the roles of the two tensors swapped
// (X := gradient of the big output, dZ := the small input), behind Conv2DTranspose (model.py:72).  With cz the fastest
// index the result is laid out exactly like the Keras kernels: (kh, kw, Cin, Cout) for Conv2D and (kh, kw, Cout, Cin)
// for Conv2DTranspose, so it accumulates straight into the fp32 gradient buffer next to the fp32 master weights.
//
// GEMM view: the contraction (K) dimension is PIXELS, so both operands are "MN-major" in shared memory: a TMA box of
// 128 pixels x 64 channels lands as 128 rows of 128 bytes (128-byte swizzle), one row per k.  M = 128 = two "slots",
// a slot being the 64 channels of one (tap, 64-channel block) of X, loaded as its own shifted TMA box (zero fill
// outside the image = 'same' padding); N = BN channels of dZ.  Each CTA keeps the accumulators of up to 512 / BN slot
// pairs resident in TMEM while it streams over its share of the pixel tiles (split-K over pixels), then flushes them
// with vectorised fp32 reductions (red.global.add.v4.f32) into dW.  Jobs = slot groups x dZ channel blocks x pixel
// splits, sized to a multiple of the SM count.
//
// Warp roles: 0 = TMA producer, 1 = TMEM owner + MMA issuer, 2..5 = flush.
#include <string.h>

#include "b2_conv_igemm.cuh"
#include "b2_host.h"
#include "b2_ptx.cuh"

namespace b2 {

constexpr int kWgThreads = 192;
constexpr int kWgBox = 128 * 128;        // one 128-pixel x 64-channel box
constexpr int kWgXStage = 2 * kWgBox;    // a slot pair
constexpr int kWgXStages = 4;
constexpr int kWgZStages = 2;

struct WgradParams {
    int B, H, W;                 // pixel grid of dZ
    int TB, TH, TW;              // tile box, TB*TH*TW <= 128 pixels
    int tiles_x, tiles_y, tiles_b, num_ptiles;
    int cx, cz;                  // channels of X and dZ, multiples of 64
    int num_taps, in_mul;
    int8_t dy[kMaxTaps], dx[kMaxTaps];
    int16_t wtap[kMaxTaps];      // slab of dW each tap accumulates into
    int slots_per_job;           // even, <= 2 * (512 / BN)
    int num_slots;               // num_taps * cx / 64
    int slot_groups, n_blocks, splits, num_jobs;
    int ksteps;                  // ceil(rows / 16) MMAs per box
    int f16;
    float* dw;
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// lo word of an MN-major SWIZZLE_128B descriptor: start address >> 4, LBO (distance between 64-element MN blocks) >> 4
__device__ __forceinline__ uint32_t sdesc_lo_mn(uint32_t saddr, uint32_t lbo_bytes) {
    return ((saddr & 0x3FFFF) >> 4) | (((lbo_bytes >> 4) & 0x3FFF) << 16);
}

template <int BN>
__global__ void __launch_bounds__(kWgThreads, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapZ,
                  const __grid_constant__ WgradParams p) {
    b2::pdl_entry();
    constexpr int kZStage = (BN / 64) * kWgBox;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smemX = smem;
    uint8_t* smemZ = smem + kWgXStages * kWgXStage;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smemZ + kWgZStages * kZStage);
    uint64_t* xfull = bars;            // [4]
    uint64_t* xempty = bars + 4;       // [4]
    uint64_t* zfull = bars + 8;        // [2]
    uint64_t* zempty = bars + 10;      // [2]
    uint64_t* tfull = bars + 12;       // accumulators complete -> flush warps
    uint64_t* tempty = bars + 13;      // flushed -> MMA warp may overwrite TMEM
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 14);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // Boxes smaller than 128 pixels leave the tail rows of every stage untouched: zero them once so they contribute 0.
    {
        uint4* z = reinterpret_cast<uint4*>(smem);
        const int n16 = (kWgXStages * kWgXStage + kWgZStages * kZStage) / 16;
        for (int i = threadIdx.x; i < n16; i += kWgThreads) z[i] = make_uint4(0, 0, 0, 0);
    }
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&mapX);
        tma_prefetch_desc(&mapZ);
        for (int s = 0; s < kWgXStages; ++s) { mbar_init(&xfull[s], 1); mbar_init(&xempty[s], 1); }
        for (int s = 0; s < kWgZStages; ++s) { mbar_init(&zfull[s], 1); mbar_init(&zempty[s], 1); }
        mbar_init(tfull, 1);
        mbar_init(tempty, 4);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<512>(tmem_slot);
    fence_proxy_async();   // generic-proxy zero fill must be visible to the TMA / tensor-core (async) proxy
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

    const int cxb = p.cx >> 6;
    const int rows = p.TB * p.TH * p.TW;
    const uint32_t box_bytes = uint32_t(rows) * 128u;

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        int sx = 0; uint32_t px = 0;
        int sz = 0; uint32_t pz = 0;
        for (int job = blockIdx.x; job < p.num_jobs; job += gridDim.x) {
            int j = job;
            const int g = j % p.slot_groups; j /= p.slot_groups;
            const int nb = j % p.n_blocks; j /= p.n_blocks;
            const int sp = j;
            const int slot0 = g * p.slots_per_job;
            int nslots = p.num_slots - slot0;
            if (nslots > p.slots_per_job) nslots = p.slots_per_job;
            const int t0 = int(long(p.num_ptiles) * sp / p.splits), t1 = int(long(p.num_ptiles) * (sp + 1) / p.splits);
            for (int tile = t0; tile < t1; ++tile) {
                int t = tile;
                const int xt = t % p.tiles_x; t /= p.tiles_x;
                const int yt = t % p.tiles_y; t /= p.tiles_y;
                const int x0 = xt * p.TW, y0 = yt * p.TH, b0 = t * p.TB;
                mbar_wait(&zempty[sz], pz ^ 1);
                if (lane == 0) {
                    mbar_expect_tx(&zfull[sz], box_bytes * uint32_t(BN / 64));
                    for (int h = 0; h < BN / 64; ++h)
                        tma_load_4d(smemZ + sz * kZStage + h * kWgBox, &mapZ, &zfull[sz], nb * BN + h * 64, x0, y0, b0);
                }
                __syncwarp();
                if (++sz == kWgZStages) { sz = 0; pz ^= 1; }
                for (int s = 0; s < nslots; s += 2) {
                    const int nbox = (nslots - s) >= 2 ? 2 : 1;
                    mbar_wait(&xempty[sx], px ^ 1);
                    if (lane == 0) {
                        mbar_expect_tx(&xfull[sx], box_bytes * uint32_t(nbox));
                        for (int h = 0; h < nbox; ++h) {
                            const int slot = slot0 + s + h;
                            const int tap = slot / cxb, cb = slot - tap * cxb;
                            tma_load_4d(smemX + sx * kWgXStage + h * kWgBox, &mapX, &xfull[sx], cb * 64,
                                        x0 * p.in_mul + p.dx[tap], y0 * p.in_mul + p.dy[tap], b0);
                        }
                    }
                    __syncwarp();
                    if (++sx == kWgXStages) { sx = 0; px ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        const uint32_t idesc = p.f16 ? make_idesc_f16(128, BN, 1, 1) : make_idesc_bf16(128, BN, 1, 1);
        constexpr uint32_t desc_hi = sdesc_hi_k_sw128(1024);   // SBO = 1024 B between 8-pixel groups, SWIZZLE_128B
        const uint32_t a_lo0 = sdesc_lo_mn(smem_u32(smemX), kWgBox), b_lo0 = sdesc_lo_mn(smem_u32(smemZ), kWgBox);
        int sx = 0; uint32_t px = 0;
        int sz = 0; uint32_t pz = 0;
        uint32_t tph = 0;
        for (int job = blockIdx.x; job < p.num_jobs; job += gridDim.x) {
            int j = job;
            const int g = j % p.slot_groups; j /= p.slot_groups;
            j /= p.n_blocks;
            const int sp = j;
            const int slot0 = g * p.slots_per_job;
            int nslots = p.num_slots - slot0;
            if (nslots > p.slots_per_job) nslots = p.slots_per_job;
            const int t0 = int(long(p.num_ptiles) * sp / p.splits), t1 = int(long(p.num_ptiles) * (sp + 1) / p.splits);
            mbar_wait(tempty, tph ^ 1);
            tc_fence_after();
            for (int tile = t0; tile < t1; ++tile) {
                mbar_wait(&zfull[sz], pz);
                tc_fence_after();
                const uint32_t b_lo = b_lo0 + uint32_t(sz) * (kZStage >> 4);
                for (int s = 0; s < nslots; s += 2) {
                    mbar_wait(&xfull[sx], px);
                    tc_fence_after();
                    if (elect_one()) {
                        const uint32_t a_lo = a_lo0 + uint32_t(sx) * (kWgXStage >> 4);
                        const uint32_t tmem_d = tmem_base + uint32_t((s >> 1) * BN);
                        for (int k = 0; k < p.ksteps; ++k)   // 16 pixels = two 8-row groups = 2048 B per step
                            umma_f16_lohi(tmem_d, a_lo + uint32_t(k) * 128u, desc_hi, b_lo + uint32_t(k) * 128u, desc_hi, idesc,
                                          uint32_t(tile != t0) | uint32_t(k != 0));
                        umma_commit(&xempty[sx]);
                        if (s + 2 >= nslots) {
                            umma_commit(&zempty[sz]);
                            if (tile == t1 - 1) umma_commit(tfull);
                        }
                    }
                    __syncwarp();
                    if (++sx == kWgXStages) { sx = 0; px ^= 1; }
                }
                if (++sz == kWgZStages) { sz = 0; pz ^= 1; }
            }
            if (t1 > t0) tph ^= 1;
        }
    } else {
        // ------------------------------------------------------------ flush (warps 2..5)
        const int q = warp & 3;              // TMEM lane quarter
        const int half = q >> 1;             // which slot of the pair these lanes hold
        const int ci_in = (q & 1) * 32 + lane;
        uint32_t tph = 0;
        for (int job = blockIdx.x; job < p.num_jobs; job += gridDim.x) {
            int j = job;
            const int g = j % p.slot_groups; j /= p.slot_groups;
            const int nb = j % p.n_blocks; j /= p.n_blocks;
            const int sp = j;
            const int slot0 = g * p.slots_per_job;
            int nslots = p.num_slots - slot0;
            if (nslots > p.slots_per_job) nslots = p.slots_per_job;
            const int t0 = int(long(p.num_ptiles) * sp / p.splits), t1 = int(long(p.num_ptiles) * (sp + 1) / p.splits);
            if (t1 <= t0) continue;
            mbar_wait(tfull, tph);
            tc_fence_after();
            for (int s = 0; s < nslots; s += 2) {
                const int sl = s + half;
                const bool valid = sl < nslots;
                const int slot = slot0 + (valid ? sl : 0);
                const int tap = slot / cxb, cb = slot - tap * cxb;
                float* base = p.dw + (size_t(p.wtap[tap]) * p.cx + size_t(cb * 64 + ci_in)) * p.cz + size_t(nb) * BN;
#pragma unroll 1
                for (int c0 = 0; c0 < BN; c0 += 32) {
                    uint32_t r[32];
                    tmem_ld32(tmem_base + (uint32_t(q * 32) << 16) + uint32_t((s >> 1) * BN + c0), r);
                    tmem_ld_wait();
                    if (valid) {
#pragma unroll
                        for (int v = 0; v < 8; ++v)
                            red_add_v4(base + c0 + 4 * v, __uint_as_float(r[4 * v]), __uint_as_float(r[4 * v + 1]),
                                       __uint_as_float(r[4 * v + 2]), __uint_as_float(r[4 * v + 3]));
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty);
            tph ^= 1;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

template <int BN>
static int launch_wgrad_bn(const CUtensorMap& mapX, const CUtensorMap& mapZ, const WgradParams& p, cudaStream_t stream) {
    constexpr int smem_bytes = kWgXStages * kWgXStage + kWgZStages * (BN / 64) * kWgBox + 1024 + 256;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(conv_wgrad_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
        if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(conv_wgrad): %s", cudaGetErrorString(e));
        attr_set = true;
    }
    const int sms = device_sm_count();
    const int grid = p.num_jobs < sms ? p.num_jobs : sms;
    b2::launch_k(conv_wgrad_kernel<BN>, grid, kWgThreads, smem_bytes, stream, mapX, mapZ, p);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("conv_wgrad launch: %s", cudaGetErrorString(e));
    return 0;
}

// Same tile chooser as the forward kernel: the box of <= 128 pixels that wastes the fewest MMA rows.
static void wgrad_choose_tile(int B, int H, int W, int& TB, int& TH, int& TW, int ext_max = 128) {
    double best = -1.0;
    TB = TH = TW = 1;
    const int wmax = W < ext_max ? W : ext_max;
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
        if ((tw & 7) == 0) eff += 1e-6;
        eff += 1e-9 * tw;
        if (eff > best) { best = eff; TB = tb; TH = th; TW = tw; }
    }
}

int launch_conv_wgrad(const WgradArgs& a, cudaStream_t stream) {
    WgradParams p;
    memset(&p, 0, sizeof(p));
    if (a.cx % 64 != 0 || a.cz % 64 != 0) return set_error("wgrad: channel counts %d / %d must be multiples of 64", a.cx, a.cz);
    if (a.num_taps < 1 || a.num_taps > kMaxTaps) return set_error("wgrad: bad tap count %d", a.num_taps);
    if (a.in_mul < 1 || a.in_mul > 3) return set_error("wgrad: stride must be 1, 2 or 3");
    p.B = a.zB; p.H = a.zH; p.W = a.zW;
    p.cx = a.cx; p.cz = a.cz;
    p.num_taps = a.num_taps; p.in_mul = a.in_mul;
    for (int t = 0; t < a.num_taps; ++t) { p.dy[t] = a.dy[t]; p.dx[t] = a.dx[t]; p.wtap[t] = a.wtap[t]; }
    p.dw = a.dw;
    p.f16 = a.f16;
    wgrad_choose_tile(p.B, p.H, p.W, p.TB, p.TH, p.TW, a.in_mul <= 2 ? 128 : 256 / a.in_mul);
    p.tiles_x = (p.W + p.TW - 1) / p.TW;
    p.tiles_y = (p.H + p.TH - 1) / p.TH;
    p.tiles_b = (p.B + p.TB - 1) / p.TB;
    p.num_ptiles = p.tiles_x * p.tiles_y * p.tiles_b;
    const int rows = p.TB * p.TH * p.TW;
    p.ksteps = (rows + 15) / 16;
    const int BN = (a.cz % 128 == 0) ? 128 : 64;
    p.num_slots = a.num_taps * (a.cx / 64);
    const int max_slots = 2 * (512 / BN);
    // spread the slots evenly over the groups (an even count per group keeps the M = 128 pairs full)
    p.slot_groups = (p.num_slots + max_slots - 1) / max_slots;
    int spj = (p.num_slots + p.slot_groups - 1) / p.slot_groups;
    spj = (spj + 1) & ~1;
    p.slots_per_job = spj;
    p.slot_groups = (p.num_slots + spj - 1) / spj;
    p.n_blocks = a.cz / BN;
    const int base = p.slot_groups * p.n_blocks;
    const int sms = device_sm_count();
    int splits = (sms * ((base + sms - 1) / sms)) / base;
    if (splits > p.num_ptiles) splits = p.num_ptiles;
    if (splits < 1) splits = 1;
    p.splits = splits;
    p.num_jobs = base * splits;

    CUtensorMap mapX, mapZ;
    {
        cuuint64_t dims[4] = {cuuint64_t(a.cx), cuuint64_t(a.xW), cuuint64_t(a.xH), cuuint64_t(a.xB)};
        cuuint64_t strides[3] = {cuuint64_t(a.cx) * 2, cuuint64_t(a.xW) * a.cx * 2, cuuint64_t(a.xH) * a.xW * a.cx * 2};
        cuuint32_t box[4] = {64, cuuint32_t(p.TW * p.in_mul), cuuint32_t(p.TH * p.in_mul), cuuint32_t(p.TB)};
        cuuint32_t estr[4] = {1, cuuint32_t(p.in_mul), cuuint32_t(p.in_mul), 1};
        int rc = encode_tensor_map(&mapX, a.x, 4, dims, strides, box, estr);
        if (rc) return rc;
    }
    {
        cuuint64_t dims[4] = {cuuint64_t(a.cz), cuuint64_t(a.zW), cuuint64_t(a.zH), cuuint64_t(a.zB)};
        cuuint64_t strides[3] = {cuuint64_t(a.cz) * 2, cuuint64_t(a.zW) * a.cz * 2, cuuint64_t(a.zH) * a.zW * a.cz * 2};
        cuuint32_t box[4] = {64, cuuint32_t(p.TW), cuuint32_t(p.TH), cuuint32_t(p.TB)};
        cuuint32_t estr[4] = {1, 1, 1, 1};
        int rc = encode_tensor_map(&mapZ, a.dz, 4, dims, strides, box, estr);
        if (rc) return rc;
    }
    if (BN == 128) return launch_wgrad_bn<128>(mapX, mapZ, p, stream);
    return launch_wgrad_bn<64>(mapX, mapZ, p, stream);
}

}  // namespace b2
