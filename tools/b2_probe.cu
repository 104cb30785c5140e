// b2_probe.cu — standalone GPU probe: hardware experiments that decide kernel design (descriptor row shifts,
// MN-major descriptors, TMA L2->smem bandwidth, tcgen05.mma issue rate) and a correctness + timing sweep of the
// implicit-GEMM convolution against the CUDA-core cross-check.  Usage: b2_probe <shift|mn|tma|mma|conv|time>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../video-cycle_gan-upscaling_b200/csrc/b2_host.h"
#include "../video-cycle_gan-upscaling_b200/csrc/b2_ptx.cuh"

using namespace b2;

#define CK(x)                                                                                  \
    do {                                                                                       \
        cudaError_t e_ = (x);                                                                  \
        if (e_ != cudaSuccess) {                                                               \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);    \
            exit(2);                                                                           \
        }                                                                                      \
    } while (0)

static uint32_t g_seed = 12345;
static float frand() {
    g_seed = g_seed * 1664525u + 1013904223u;
    return ((g_seed >> 8) & 0xFFFF) / 65536.0f * 2.f - 1.f;
}
static __nv_bfloat16 f2bf(float f) { return __float2bfloat16_rn(f); }
static float bf2f(__nv_bfloat16 b) { return __bfloat162float(b); }

// ============================================================== experiment 1: descriptor row shift
constexpr int SH_ROWS = 160;
constexpr int SH_NVAR = 12;
__global__ void __launch_bounds__(128, 1)
probe_shift_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, float* out,
                   const int* shifts, const int* modes) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;                   // 160 x 128 B
    uint8_t* sB = smem + 24 * 1024;       // 64 x 128 B
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 36 * 1024);
    uint64_t* mbar = bar + 1;
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        mbar_init(mbar, 1);
        fence_mbar_init();
    }
    if (warp == 0) tmem_alloc<64>(slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *slot;
    if (threadIdx.x == 0) {
        mbar_expect_tx(bar, SH_ROWS * 128 + 64 * 128);
        tma_load_2d(sA, &mapA, bar, 0, 0);
        tma_load_2d(sB, &mapB, bar, 0, 0);
    }
    mbar_wait(bar, 0);
    tc_fence_after();
    uint32_t ph = 0;
    for (int v = 0; v < SH_NVAR; ++v) {
        if (threadIdx.x == 0) {
            const uint32_t a0 = smem_u32(sA) + shifts[v] * 128, b0 = smem_u32(sB);
            const uint32_t bo = modes[v] ? ((a0 >> 7) & 7) : 0;
            for (int k = 0; k < 4; ++k)
                umma_bf16(tmem, make_sdesc_k_sw128(a0 + k * 32, 1024, bo), make_sdesc_k_sw128(b0 + k * 32, 1024),
                          make_idesc_bf16(128, 64), k != 0);
            umma_commit(mbar);
        }
        mbar_wait(mbar, ph);
        ph ^= 1;
        tc_fence_after();
        for (int c0 = 0; c0 < 64; c0 += 32) {
            uint32_t r[32];
            tmem_ld32(tmem + (uint32_t(warp * 32) << 16) + c0, r);
            tmem_ld_wait();
            for (int j = 0; j < 32; ++j) out[(size_t(v) * 128 + warp * 32 + lane) * 64 + c0 + j] = __uint_as_float(r[j]);
        }
        tc_fence_before();
        __syncthreads();
        tc_fence_after();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<64>(tmem);
}

static int run_shift() {
    std::vector<__nv_bfloat16> hA(SH_ROWS * 64), hB(64 * 64);
    for (auto& v : hA) v = f2bf(frand());
    for (auto& v : hB) v = f2bf(frand());
    __nv_bfloat16 *dA, *dB;
    float* dO;
    int *dS, *dM;
    CK(cudaMalloc(&dA, hA.size() * 2));
    CK(cudaMalloc(&dB, hB.size() * 2));
    CK(cudaMalloc(&dO, SH_NVAR * 128 * 64 * 4));
    CK(cudaMalloc(&dS, SH_NVAR * 4));
    CK(cudaMalloc(&dM, SH_NVAR * 4));
    CK(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice));
    int shifts[SH_NVAR] = {0, 1, 1, 2, 2, 3, 3, 7, 7, 8, 9, 9};
    int modes[SH_NVAR] = {0, 0, 1, 0, 1, 0, 1, 0, 1, 0, 0, 1};
    CK(cudaMemcpy(dS, shifts, sizeof(shifts), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dM, modes, sizeof(modes), cudaMemcpyHostToDevice));
    CUtensorMap mA, mB;
    {
        cuuint64_t dims[2] = {64, SH_ROWS};
        cuuint64_t str[1] = {128};
        cuuint32_t box[2] = {64, SH_ROWS};
        cuuint32_t es[2] = {1, 1};
        if (encode_tensor_map(&mA, dA, 2, dims, str, box, es)) { printf("%s\n", last_error()); return 1; }
        cuuint64_t dimsb[2] = {64, 64};
        cuuint32_t boxb[2] = {64, 64};
        if (encode_tensor_map(&mB, dB, 2, dimsb, str, boxb, es)) { printf("%s\n", last_error()); return 1; }
    }
    CK(cudaFuncSetAttribute(probe_shift_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 40 * 1024));
    probe_shift_kernel<<<1, 128, 40 * 1024>>>(mA, mB, dO, dS, dM);
    CK(cudaDeviceSynchronize());
    std::vector<float> hO(SH_NVAR * 128 * 64);
    CK(cudaMemcpy(hO.data(), dO, hO.size() * 4, cudaMemcpyDeviceToHost));
    for (int v = 0; v < SH_NVAR; ++v) {
        double maxerr = 0;
        for (int i = 0; i < 128; ++i)
            for (int n = 0; n < 64; ++n) {
                double ref = 0;
                for (int k = 0; k < 64; ++k) ref += double(bf2f(hA[(shifts[v] + i) * 64 + k])) * bf2f(hB[n * 64 + k]);
                maxerr = fmax(maxerr, fabs(ref - hO[(size_t(v) * 128 + i) * 64 + n]));
            }
        printf("SHIFT rows=%d base_offset_mode=%d  max_err=%.5f  %s\n", shifts[v], modes[v], maxerr,
               maxerr < 1e-2 ? "OK" : "MISMATCH");
    }
    return 0;
}

// ============================================================== experiment 2: MN-major operands (wgrad shape)
__global__ void __launch_bounds__(128, 1)
probe_mn_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapY, float* out, int variant) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sX = smem;               // 2 x (64 pixels x 128 B): channels 0..63, 64..127
    uint8_t* sY = smem + 16 * 1024;   // 64 pixels x 128 B
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 24 * 1024);
    uint64_t* mbar = bar + 1;
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        mbar_init(mbar, 1);
        fence_mbar_init();
    }
    if (warp == 0) tmem_alloc<64>(slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *slot;
    if (threadIdx.x == 0) {
        mbar_expect_tx(bar, 3 * 64 * 128);
        tma_load_2d(sX, &mapX, bar, 0, 0);
        tma_load_2d(sX + 8192, &mapX, bar, 64, 0);
        tma_load_2d(sY, &mapY, bar, 0, 0);
    }
    mbar_wait(bar, 0);
    tc_fence_after();
    uint32_t ph = 0;
    for (int v = variant; v < variant + 1; ++v) {
        if (threadIdx.x == 0) {
            const uint32_t lbo = v == 0 ? 8192 : 1024, sbo = v == 0 ? 1024 : 8192;
            for (int k = 0; k < 4; ++k)
                umma_bf16(tmem, make_sdesc_mn_sw128(smem_u32(sX) + k * 2048, lbo, sbo),
                          make_sdesc_mn_sw128(smem_u32(sY) + k * 2048, lbo, sbo), make_idesc_bf16(128, 64, 1, 1), k != 0);
            umma_commit(mbar);
        }
        mbar_wait(mbar, ph);
        ph ^= 1;
        tc_fence_after();
        for (int c0 = 0; c0 < 64; c0 += 32) {
            uint32_t r[32];
            tmem_ld32(tmem + (uint32_t(warp * 32) << 16) + c0, r);
            tmem_ld_wait();
            for (int j = 0; j < 32; ++j) out[(size_t(v) * 128 + warp * 32 + lane) * 64 + c0 + j] = __uint_as_float(r[j]);
        }
        tc_fence_before();
        __syncthreads();
        tc_fence_after();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<64>(tmem);
}

static int run_mn() {
    const int P = 64;
    std::vector<__nv_bfloat16> hX(P * 128), hY(P * 64);
    for (auto& v : hX) v = f2bf(frand());
    for (auto& v : hY) v = f2bf(frand());
    __nv_bfloat16 *dX, *dY;
    float* dO;
    CK(cudaMalloc(&dX, hX.size() * 2));
    CK(cudaMalloc(&dY, hY.size() * 2));
    CK(cudaMalloc(&dO, 2 * 128 * 64 * 4));
    CK(cudaMemcpy(dX, hX.data(), hX.size() * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dY, hY.data(), hY.size() * 2, cudaMemcpyHostToDevice));
    CUtensorMap mX, mY;
    cuuint32_t es[2] = {1, 1};
    {
        cuuint64_t dims[2] = {128, cuuint64_t(P)};
        cuuint64_t str[1] = {256};
        cuuint32_t box[2] = {64, cuuint32_t(P)};
        if (encode_tensor_map(&mX, dX, 2, dims, str, box, es)) { printf("%s\n", last_error()); return 1; }
        cuuint64_t dimy[2] = {64, cuuint64_t(P)};
        cuuint64_t stry[1] = {128};
        if (encode_tensor_map(&mY, dY, 2, dimy, stry, box, es)) { printf("%s\n", last_error()); return 1; }
    }
    CK(cudaFuncSetAttribute(probe_mn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    const int variant = getenv("MN_VARIANT") ? atoi(getenv("MN_VARIANT")) : 0;
    CK(cudaMemset(dO, 0, 2 * 128 * 64 * 4));
    probe_mn_kernel<<<1, 128, 200 * 1024>>>(mX, mY, dO, variant);
    CK(cudaDeviceSynchronize());
    std::vector<float> hO(2 * 128 * 64);
    CK(cudaMemcpy(hO.data(), dO, hO.size() * 4, cudaMemcpyDeviceToHost));
    for (int v = variant; v < variant + 1; ++v) {
        double maxerr = 0;
        for (int m = 0; m < 128; ++m)
            for (int n = 0; n < 64; ++n) {
                double ref = 0;
                for (int p = 0; p < P; ++p) ref += double(bf2f(hX[p * 128 + m])) * bf2f(hY[p * 64 + n]);
                maxerr = fmax(maxerr, fabs(ref - hO[(size_t(v) * 128 + m) * 64 + n]));
            }
        printf("MNMAJOR variant=%d (lbo=%d sbo=%d) max_err=%.5f %s\n", v, v == 0 ? 8192 : 1024, v == 0 ? 1024 : 8192,
               maxerr, maxerr < 1e-2 ? "OK" : "MISMATCH");
    }
    return 0;
}

// ============================================================== experiment 3: TMA L2 -> smem feed rate
__global__ void __launch_bounds__(64, 1)
probe_tma_kernel(const __grid_constant__ CUtensorMap mapA, int iters, int H, int W, int spread) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    constexpr int ST = 8;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + ST * 16384);
    uint64_t* empty = full + ST;
    if (threadIdx.x == 0) {
        for (int s = 0; s < ST; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        fence_mbar_init();
    }
    __syncthreads();
    const int tiles_x = W / 16, tiles_y = H / 8;
    if (threadIdx.x == 0) {
        int stage = 0; uint32_t ph = 0;
        for (int i = 0; i < iters; ++i) {
            const int t = (blockIdx.x * spread + i) % (tiles_x * tiles_y);
            mbar_wait(&empty[stage], ph ^ 1);
            mbar_expect_tx(&full[stage], 16384);
            tma_load_4d(smem + stage * 16384, &mapA, &full[stage], 0, (t % tiles_x) * 16, (t / tiles_x) * 8, 0);
            if (++stage == ST) { stage = 0; ph ^= 1; }
        }
    } else if (threadIdx.x == 32) {
        int stage = 0; uint32_t ph = 0;
        for (int i = 0; i < iters; ++i) {
            mbar_wait(&full[stage], ph);
            mbar_arrive(&empty[stage]);
            if (++stage == ST) { stage = 0; ph ^= 1; }
        }
    }
}

static int run_tma() {
    const int H = 272, W = 480, C = 64;  // 16.7 MB, L2 resident
    __nv_bfloat16* dX;
    CK(cudaMalloc(&dX, size_t(H) * W * C * 2));
    CK(cudaMemset(dX, 0, size_t(H) * W * C * 2));
    CUtensorMap mA;
    cuuint64_t dims[4] = {cuuint64_t(C), cuuint64_t(W), cuuint64_t(H), 1};
    cuuint64_t str[3] = {cuuint64_t(C) * 2, cuuint64_t(W) * C * 2, cuuint64_t(H) * W * C * 2};
    cuuint32_t box[4] = {64, 16, 8, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    if (encode_tensor_map(&mA, dX, 4, dims, str, box, es)) { printf("%s\n", last_error()); return 1; }
    CK(cudaFuncSetAttribute(probe_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 16384 + 2048));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int spread : {1, 7, 9}) {
        for (int rep = 0; rep < 2; ++rep) {
            const int iters = 4096;
            CK(cudaEventRecord(e0));
            probe_tma_kernel<<<148, 64, 8 * 16384 + 2048>>>(mA, iters, H, W, spread);
            CK(cudaEventRecord(e1));
            CK(cudaDeviceSynchronize());
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            const double bytes = 148.0 * iters * 16384;
            if (rep) printf("TMA L2->smem spread=%d: %.1f GB/s (%.3f ms, %.1f B/clk/SM @1.9GHz)\n", spread, bytes / ms / 1e6, ms,
                            bytes / 148 / (ms * 1e-3 * 1.9e9));
        }
    }
    return 0;
}

// ============================================================== experiment 4: tcgen05.mma issue rate from smem
template <int N>
__global__ void __launch_bounds__(128, 1) probe_mma_kernel(int iters, long long* cycles) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;               // 128 x 128 B
    uint8_t* sB = smem + 16384;       // N x 128 B
    uint64_t* mbar = reinterpret_cast<uint64_t*>(smem + 16384 + 256 * 128);
    uint32_t* slot = reinterpret_cast<uint32_t*>(mbar + 1);
    for (int i = threadIdx.x; i < (16384 + N * 128) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    fence_proxy_async();
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { mbar_init(mbar, 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc<256>(slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *slot;
    if (threadIdx.x == 0) {
        const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
        const long long t0 = clock64();
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                umma_bf16(tmem, make_sdesc_k_sw128(a0 + k * 32, 1024), make_sdesc_k_sw128(b0 + k * 32, 1024),
                          make_idesc_bf16(128, N), 1);
        }
        umma_commit(mbar);
        mbar_wait(mbar, 0);
        const long long t1 = clock64();
        if (blockIdx.x == 0) *cycles = t1 - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<256>(tmem);
}

template <int N>
static void run_mma_n() {
    long long* dC;
    CK(cudaMalloc(&dC, 8));
    const int smem = 16384 + 256 * 128 + 2048;
    CK(cudaFuncSetAttribute(probe_mma_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int iters = 2048;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int rep = 0; rep < 2; ++rep) {
        CK(cudaEventRecord(e0));
        probe_mma_kernel<N><<<148, 128, smem>>>(iters, dC);
        CK(cudaEventRecord(e1));
        CK(cudaDeviceSynchronize());
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        long long cyc;
        CK(cudaMemcpy(&cyc, dC, 8, cudaMemcpyDeviceToHost));
        const double flops = 148.0 * iters * 4 * 2.0 * 128 * N * 16;
        if (rep) printf("MMA M=128 N=%d K=16 SS: %.1f cycles/MMA, %.1f TFLOP/s chip (%.3f ms)\n", N,
                        double(cyc) / (iters * 4), flops / ms / 1e9, ms);
    }
}
static int run_mma() {
    run_mma_n<16>();
    run_mma_n<32>();
    run_mma_n<64>();
    run_mma_n<128>();
    run_mma_n<256>();
    return 0;
}


// ============================================================== experiment 4b: MMA rate vs operand view alignment
// A is a [512 rows x 128 B] region; each MMA group reads a view at `shift` rows with group stride `sbo`; `rotate`
// cycles the view start over 9 different row offsets (like the taps of a 3x3 conv).
template <int N>
__global__ void __launch_bounds__(128, 1) probe_mma2_kernel(int iters, int shift, int sbo, int rotate, int bshift, long long* cycles) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;               // 64 KB
    uint8_t* sB = smem + 65536;       // up to 9 x N x 128 B
    uint64_t* mbar = reinterpret_cast<uint64_t*>(smem + 65536 + 98304);
    uint32_t* slot = reinterpret_cast<uint32_t*>(mbar + 1);
    constexpr int NB = 98304 / (N * 128) > 9 ? 9 : 98304 / (N * 128);
    for (int i = threadIdx.x; i < (65536 + 98304) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + i;
    fence_proxy_async();
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { mbar_init(mbar, 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc<256>(slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *slot;
    if (threadIdx.x == 0) {
        const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
        const long long t0 = clock64();
        for (int i = 0; i < iters; ++i) {
            const int tap = rotate ? (i % 9) : 0;
            const uint32_t a = a0 + uint32_t(shift + (rotate ? (tap / 3) * (sbo / 128) + (tap % 3) : 0)) * 128u;
            const uint32_t b = b0 + (bshift ? uint32_t(tap % NB) * N * 128 : 0);
#pragma unroll
            for (int k = 0; k < 4; ++k)
                umma_bf16(tmem, make_sdesc_k_sw128(a + k * 32, sbo), make_sdesc_k_sw128(b + k * 32, 1024),
                          make_idesc_bf16(128, N), 1);
        }
        umma_commit(mbar);
        mbar_wait(mbar, 0);
        const long long t1 = clock64();
        if (blockIdx.x == 0) *cycles = t1 - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<256>(tmem);
}

template <int N>
static void run_mma2_n() {
    long long* dC;
    CK(cudaMalloc(&dC, 8));
    const int smem = 65536 + 98304 + 2048;
    CK(cudaFuncSetAttribute(probe_mma2_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int iters = 2048;
    struct V { int shift, sbo, rotate, bshift; } vs[] = {{0, 1024, 0, 0}, {1, 1024, 0, 0}, {8, 1024, 0, 0}, {0, 1280, 0, 0},
                                                          {1, 1280, 0, 0}, {0, 2048, 0, 0}, {0, 1280, 1, 0}, {0, 1280, 1, 1},
                                                          {0, 1024, 1, 1}, {0, 2048, 1, 1}};
    for (auto v : vs) {
        for (int rep = 0; rep < 2; ++rep) {
            probe_mma2_kernel<N><<<148, 128, smem>>>(iters, v.shift, v.sbo, v.rotate, v.bshift, dC);
            CK(cudaDeviceSynchronize());
        }
        long long cyc;
        CK(cudaMemcpy(&cyc, dC, 8, cudaMemcpyDeviceToHost));
        printf("MMA2 N=%d shift=%d sbo=%d rotate=%d bshift=%d: %.1f cycles/MMA\n", N, v.shift, v.sbo, v.rotate, v.bshift,
               double(cyc) / (iters * 4));
    }
}
// ============================================================== experiment 4c: MMA rate vs number of accumulators
// Same as 4b, but consecutive MMA groups rotate over `nacc` different TMEM accumulators: tells whether the ~80-cycle
// floor for N <= 128 is a dependency on the accumulator (then interleaving two tiles hides it) or an operand-fetch limit.
// mode 0: rotate the accumulator per group of 4 K-steps; mode 1: rotate per single MMA.
template <int N>
__global__ void __launch_bounds__(128, 1) probe_mma3_kernel(int iters, int nacc, int mode, long long* cycles) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sA = smem;
    uint8_t* sB = smem + 65536;
    uint64_t* mbar = reinterpret_cast<uint64_t*>(smem + 65536 + 98304);
    uint32_t* slot = reinterpret_cast<uint32_t*>(mbar + 1);
    for (int i = threadIdx.x; i < (65536 + 98304) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + i;
    fence_proxy_async();
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { mbar_init(mbar, 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc<512>(slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *slot;
    if (threadIdx.x == 0) {
        const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
        const long long t0 = clock64();
        int acc = 0;
        for (int i = 0; i < iters; ++i) {
            const int tap = i % 9;
            const uint32_t a = a0 + uint32_t((tap / 3) * 10 + (tap % 3)) * 128u;
            const uint32_t b = b0 + uint32_t(tap % 6) * N * 128;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                umma_bf16(tmem + uint32_t(acc) * N, make_sdesc_k_sw128(a + k * 32, 1280), make_sdesc_k_sw128(b + k * 32, 1024),
                          make_idesc_bf16(128, N), 1);
                if (mode == 1) { if (++acc == nacc) acc = 0; }
            }
            if (mode == 0) { if (++acc == nacc) acc = 0; }
        }
        umma_commit(mbar);
        mbar_wait(mbar, 0);
        const long long t1 = clock64();
        if (blockIdx.x == 0) *cycles = t1 - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tmem);
}

template <int N>
static void run_mma3_n() {
    long long* dC;
    CK(cudaMalloc(&dC, 8));
    const int smem = 65536 + 98304 + 2048;
    CK(cudaFuncSetAttribute(probe_mma3_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int iters = 2048;
    for (int mode = 0; mode < 2; ++mode)
        for (int nacc = 1; nacc <= 4 && nacc * N <= 512; nacc *= 2) {
            for (int rep = 0; rep < 2; ++rep) {
                probe_mma3_kernel<N><<<148, 128, smem>>>(iters, nacc, mode, dC);
                CK(cudaDeviceSynchronize());
            }
            long long cyc;
            CK(cudaMemcpy(&cyc, dC, 8, cudaMemcpyDeviceToHost));
            printf("MMA3 N=%d nacc=%d mode=%d: %.1f cycles/MMA\n", N, nacc, mode, double(cyc) / (iters * 4));
        }
}
static int run_mma3() {
    run_mma3_n<64>();
    run_mma3_n<128>();
    run_mma3_n<192>();
    run_mma3_n<256>();
    return 0;
}

static int run_mma2() {
    run_mma2_n<32>();
    run_mma2_n<64>();
    run_mma2_n<128>();
    run_mma2_n<256>();
    return 0;
}

// ============================================================== igemm vs direct
struct ConvCase {
    const char* name;
    int B, H, W, Cin, Cout, k, stride;   // conv: output grid = ceil(H/stride)
    int convT_phase;                     // -1: plain conv; 0..3: sub-pixel phase (py*2+px) of ConvT stride 2
    int act, residual, out_fp32;
};

static void fill_taps(IgemmParams& p, const ConvCase& c, int& OHg, int& OWg) {
    memset(&p, 0, sizeof(p));
    p.cin = c.Cin;
    p.n_valid = c.Cout;
    p.act = c.act;
    p.leaky = 0.2f;
    p.out_fp32 = c.out_fp32;
    p.out_ld = c.Cout;
    if (c.convT_phase < 0) {
        // TF "same": out = ceil(in/s); pad_total = max((out-1)*s + k - in, 0); pad_before = pad_total/2
        const int OH = (c.H + c.stride - 1) / c.stride, OW = (c.W + c.stride - 1) / c.stride;
        const int pth = (OH - 1) * c.stride + c.k - c.H, ptw = (OW - 1) * c.stride + c.k - c.W;
        const int pbh = (pth > 0 ? pth : 0) / 2, pbw = (ptw > 0 ? ptw : 0) / 2;
        p.B = c.B; p.H = OH; p.W = OW;
        p.in_mul = c.stride;
        p.num_taps = c.k * c.k;
        for (int ky = 0; ky < c.k; ++ky)
            for (int kx = 0; kx < c.k; ++kx) {
                const int t = ky * c.k + kx;
                p.dy[t] = int8_t(ky - pbh);
                p.dx[t] = int8_t(kx - pbw);
                p.wtap[t] = int16_t(t);
            }
        p.OH = OH; p.OW = OW; p.out_sy = 1; p.out_sx = 1; p.out_oy = 0; p.out_ox = 0;
        OHg = OH; OWg = OW;
    } else {
        // ConvT stride 2 "same": out[y] = sum_i x[i] * W[y - 2i + pb], pb = (k-2)/2  
        const int py = c.convT_phase >> 1, px = c.convT_phase & 1, pb = (c.k - 2) / 2;
        p.B = c.B; p.H = c.H; p.W = c.W;
        p.in_mul = 1;
        int nt = 0;
        for (int ky = 0; ky < c.k; ++ky) {
            if (((py + pb - ky) & 1) != 0) continue;   // y = 2j+py, i = j + d  =>  ky = py + pb - 2d
            const int dy = (py + pb - ky) / 2;
            for (int kx = 0; kx < c.k; ++kx) {
                if (((px + pb - kx) & 1) != 0) continue;
                const int dx = (px + pb - kx) / 2;
                p.dy[nt] = int8_t(dy);
                p.dx[nt] = int8_t(dx);
                p.wtap[nt] = int16_t(ky * c.k + kx);
                ++nt;
            }
        }
        p.num_taps = nt;
        p.OH = 2 * c.H; p.OW = 2 * c.W; p.out_sy = 2; p.out_sx = 2; p.out_oy = py; p.out_ox = px;
        OHg = 2 * c.H; OWg = 2 * c.W;
    }
}

static int run_conv_case(const ConvCase& c, bool check, int time_iters) {
    ConvIgemmArgs a;
    int OH, OW;
    fill_taps(a.p, c, OH, OW);
    const size_t nx = size_t(c.B) * c.H * c.W * c.Cin, nw = size_t(c.Cout) * c.k * c.k * c.Cin;
    const size_t no = size_t(c.B) * OH * OW * c.Cout;
    std::vector<__nv_bfloat16> hx(nx), hw(nw), hres(no);
    const float wscale = 1.0f / sqrtf(float(c.k * c.k * c.Cin));
    for (auto& v : hx) v = f2bf(frand());
    for (auto& v : hw) v = f2bf(frand() * wscale * 1.7f);
    for (auto& v : hres) v = f2bf(frand());
    std::vector<float> hs(c.Cout), hb(c.Cout), ha(c.Cout);
    for (int i = 0; i < c.Cout; ++i) { hs[i] = 0.5f + 0.5f * fabsf(frand()); hb[i] = 0.1f * frand(); ha[i] = 0.3f * fabsf(frand()); }
    __nv_bfloat16 *dx, *dw, *dres;
    float *ds, *db, *da;
    void *o1, *o2;
    const size_t obytes = no * (c.out_fp32 ? 4 : 2);
    CK(cudaMalloc(&dx, nx * 2)); CK(cudaMalloc(&dw, nw * 2)); CK(cudaMalloc(&dres, no * 2));
    CK(cudaMalloc(&ds, c.Cout * 4)); CK(cudaMalloc(&db, c.Cout * 4)); CK(cudaMalloc(&da, c.Cout * 4));
    CK(cudaMalloc(&o1, obytes)); CK(cudaMalloc(&o2, obytes));
    CK(cudaMemset(o1, 0, obytes)); CK(cudaMemset(o2, 0, obytes));
    CK(cudaMemcpy(dx, hx.data(), nx * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dw, hw.data(), nw * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dres, hres.data(), no * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ds, hs.data(), c.Cout * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(db, hb.data(), c.Cout * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(da, ha.data(), c.Cout * 4, cudaMemcpyHostToDevice));
    a.x = dx; a.xB = c.B; a.xH = c.H; a.xW = c.W; a.xC = c.Cin;
    a.w = dw; a.w_rows = c.Cout; a.w_k = c.k * c.k * c.Cin;
    a.p.scale = ds; a.p.shift = db; a.p.alpha = da;
    a.p.residual = c.residual ? dres : nullptr;
    int rc = 0;
    if (check) {
        a.p.out = o1;
        if (launch_conv_igemm(a, 0)) { printf("CONV %-28s igemm launch failed: %s\n", c.name, last_error()); return 1; }
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("CONV %-28s igemm run failed: %s\n", c.name, cudaGetErrorString(e)); exit(3); }
        a.p.out = o2;
        if (launch_conv_direct(a, 0)) { printf("direct launch failed: %s\n", last_error()); return 1; }
        CK(cudaDeviceSynchronize());
        std::vector<uint8_t> h1(obytes), h2(obytes);
        CK(cudaMemcpy(h1.data(), o1, obytes, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(h2.data(), o2, obytes, cudaMemcpyDeviceToHost));
        double maxerr = 0, maxref = 0;
        size_t bad = 0;
        for (size_t i = 0; i < no; ++i) {
            const float v1 = c.out_fp32 ? reinterpret_cast<float*>(h1.data())[i] : bf2f(reinterpret_cast<__nv_bfloat16*>(h1.data())[i]);
            const float v2 = c.out_fp32 ? reinterpret_cast<float*>(h2.data())[i] : bf2f(reinterpret_cast<__nv_bfloat16*>(h2.data())[i]);
            const double d = fabs(double(v1) - v2);
            if (!(d <= 0.02 + 0.01 * fabs(v2))) ++bad;
            if (d > maxerr) maxerr = d;
            if (fabs(v2) > maxref) maxref = fabs(v2);
        }
        printf("CONV %-28s B%d %dx%d %d->%d k%d s%d ph%d taps=%d: max_err=%.4f (max_ref %.3f) bad=%zu/%zu %s\n", c.name, c.B,
               c.H, c.W, c.Cin, c.Cout, c.k, c.stride, c.convT_phase, a.p.num_taps, maxerr, maxref, bad, no,
               bad == 0 ? "OK" : "MISMATCH");
        rc = bad != 0;
    }
    if (time_iters > 0) {
        a.p.out = o1;
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        for (int i = 0; i < 3; ++i) launch_conv_igemm(a, 0);
        CK(cudaEventRecord(e0));
        for (int i = 0; i < time_iters; ++i) launch_conv_igemm(a, 0);
        CK(cudaEventRecord(e1));
        CK(cudaDeviceSynchronize());
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        ms /= time_iters;
        const double flops = 2.0 * a.p.B * a.p.H * a.p.W * double(c.Cout) * a.p.num_taps * c.Cin;
        printf("TIME %-28s B%d %dx%d %d->%d taps=%d: %.3f ms  %.1f TFLOP/s\n", c.name, c.B, c.H, c.W, c.Cin, c.Cout,
               a.p.num_taps, ms, flops / ms / 1e9);
    }
    cudaFree(dx); cudaFree(dw); cudaFree(dres); cudaFree(ds); cudaFree(db); cudaFree(da); cudaFree(o1); cudaFree(o2);
    return rc;
}

static int run_conv() {
    const ConvCase cases[] = {
        {"res3x3_prelu", 1, 20, 33, 64, 64, 3, 1, -1, ACT_PRELU, 0, 0},
        {"res3x3_add", 2, 17, 40, 64, 64, 3, 1, -1, ACT_NONE, 1, 0},
        {"res5x5", 1, 24, 24, 64, 64, 5, 1, -1, ACT_LEAKY, 0, 0},
        {"vgg_64_128_relu", 2, 24, 24, 64, 128, 3, 1, -1, ACT_RELU, 0, 0},
        {"vgg_256_512", 3, 12, 12, 256, 512, 3, 1, -1, ACT_RELU, 0, 0},
        {"disc_s2_even", 2, 32, 32, 64, 128, 3, 2, -1, ACT_LEAKY, 0, 0},
        {"disc_s2_odd", 1, 27, 15, 128, 256, 3, 2, -1, ACT_LEAKY, 0, 0},
        {"convT_k3_ph0", 1, 12, 20, 64, 256, 3, 1, 0, ACT_LEAKY, 0, 0},
        {"convT_k3_ph3", 1, 12, 20, 64, 256, 3, 1, 3, ACT_LEAKY, 0, 0},
        {"convT_k5_ph1", 1, 12, 20, 256, 256, 5, 1, 1, ACT_LEAKY, 0, 0},
        {"tail9x9_tanh", 1, 24, 40, 256, 3, 9, 1, -1, ACT_TANH, 0, 1},
        {"small_n32", 1, 16, 16, 64, 32, 3, 1, -1, ACT_NONE, 0, 0},
    };
    int fails = 0;
    for (const auto& c : cases) fails += run_conv_case(c, true, 0);
    printf("CONV summary: %d failing cases\n", fails);
    return fails;
}

static int run_time(int only = -1) {
    const ConvCase cases[] = {
        {"res3x3 270x480", 1, 270, 480, 64, 64, 3, 1, -1, ACT_PRELU, 0, 0},
        {"res3x3 270x480 B8", 8, 270, 480, 64, 64, 3, 1, -1, ACT_PRELU, 0, 0},
        {"res5x5 270x480", 1, 270, 480, 64, 64, 5, 1, -1, ACT_PRELU, 0, 0},
        {"convT1 k3 ph3 64->256", 1, 270, 480, 64, 256, 3, 1, 3, ACT_LEAKY, 0, 0},
        {"convT2 k3 ph0 256->256", 1, 540, 960, 256, 256, 3, 1, 0, ACT_LEAKY, 0, 0},
        {"convT2 k3 ph3 256->256", 1, 540, 960, 256, 256, 3, 1, 3, ACT_LEAKY, 0, 0},
        {"vgg 384 64->64", 4, 384, 384, 64, 64, 3, 1, -1, ACT_RELU, 0, 0},
        {"vgg 96 256->256", 16, 96, 96, 256, 256, 3, 1, -1, ACT_RELU, 0, 0},
        {"vgg 48 512->512", 16, 48, 48, 512, 512, 3, 1, -1, ACT_RELU, 0, 0},
        {"tail9x9 540x960", 1, 540, 960, 256, 3, 9, 1, -1, ACT_TANH, 0, 1},
    };
    int i = 0;
    for (const auto& c : cases) {
        if (only < 0 || only == i) run_conv_case(c, false, only < 0 ? 5 : 2);
        ++i;
    }
    return 0;
}

int main(int argc, char** argv) {
    const char* what = argc > 1 ? argv[1] : "conv";
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    printf("device: %s, %d SMs, cc %d.%d\n", prop.name, prop.multiProcessorCount, prop.major, prop.minor);
    if (!strcmp(what, "shift")) return run_shift();
    if (!strcmp(what, "mn")) return run_mn();
    if (!strcmp(what, "tma")) return run_tma();
    if (!strcmp(what, "mma")) return run_mma();
    if (!strcmp(what, "mma2")) return run_mma2();
    if (!strcmp(what, "mma3")) return run_mma3();
    if (!strcmp(what, "conv")) return run_conv();
    if (!strcmp(what, "time")) return run_time(argc > 2 ? atoi(argv[2]) : -1);
    printf("unknown experiment %s\n", what);
    return 1;
}
