// b2_conv_scatter.cu — the generator's final Conv2D(3, 9x9, 'same') + tanh (reference: upscaling/upscaler/model.py:290-291)
// as a full-rate tensor-core GEMM followed by a register-resident scatter ("col2im") epilogue.
//
// A direct implicit GEMM has N = 3 output channels; folding the filter rows into N (b2_conv_tail.cu) reaches N = 27 and
// is bound by the shared-memory A-operand read of the 144 MMAs a pixel row needs (~80 cycles each whatever N is).
// Here ALL 81 taps go into N:
//     P[x, (dy, dx, co)] = sum_c X[r, x, c] * W[dy, dx, c, co]          M = 128 pixels, N = 243 -> 256, K = Cin
//     out[r + 4 - dy, x + 4 - dx, co] += P[x, (dy, dx, co)]
// so one 128-pixel row tile costs Cin/16 = 16 MMAs at N = 256 (the shape that runs at the tensor core's full rate)
// instead of 144, every input element is fetched once, and the kernel becomes bound by reading the input from HBM.
// The scatter runs in the epilogue warps without shared memory or atomics: a warp's 32 TMEM lanes are 32 consecutive
// input pixels; output pixel (lane l) needs P from lanes l + dx - 4, which is a warp shuffle; the sum over dy is a ring
// of 9 x 3 fp32 partial output rows held in REGISTERS while the CTA walks down a strip of the image (input row r feeds
// output rows r-4 .. r+4; output row r-4 is finished, bias + tanh applied and stored, after row r).  To keep the
// shuffles inside one warp each warp's 32 pixels overlap the next warp's by 8: a tile is 4 TMA boxes of 32 pixels at
// x0 + 24 q - 4, producing 96 finished output pixels (lanes 4..27 of each warp).  Out-of-image pixels are zero-filled
// by TMA, which is the 'same' padding.
//
// Weights ([256][Cin] K-major, 128 KB for Cin = 256) stay resident in shared memory; input row tiles stream through a
// ring of 64-channel blocks.  Warp roles: 0 = TMA producer, 1 = TMEM owner + MMA issuer, 2..5 = scatter epilogue.
#include "b2_conv_igemm.cuh"
#include "b2_host.h"
#include "b2_ptx.cuh"

namespace b2 {

constexpr int kScThreads = 192;
constexpr int kScOutPerWarp = 24;                 // finished output pixels per warp (32 lanes - 2 x 4 halo)
constexpr int kScOutPerTile = 4 * kScOutPerWarp;  // 96
constexpr int kScABlock = 128 * 128;              // 128 pixels x 64 channels
constexpr int kScAStages = 5;
constexpr int kScBBlock = 256 * 128;              // 256 rows (dy,dx,co) x 64 channels
constexpr int kScMaxCb = 4;                       // Cin <= 256

struct ScatterParams {
    int B, H, W, cin;
    int strips;           // ceil(W / 96)
    long total_tiles;     // B * strips * H output-row tiles
    int act, f16;
    const float* bias;
    float* out;           // [B, H, W, 3] fp32
};

struct ScSeg { int y0, y1, img, x0; };
// The flattened (image, strip, output row) space is cut into one contiguous range per CTA; a segment is the part of a
// range that lies in one (image, strip).
__device__ __forceinline__ bool sc_seg_next(long& t, long t_end, const ScatterParams& p, ScSeg& s) {
    if (t >= t_end) return false;
    const long col = t / p.H;
    long seg_end = (col + 1) * p.H;
    if (seg_end > t_end) seg_end = t_end;
    s.y0 = int(t - col * p.H);
    s.y1 = s.y0 + int(seg_end - t);
    s.img = int(col / p.strips);
    s.x0 = int(col % p.strips) * kScOutPerTile;
    t = seg_end;
    return true;
}

// Adds the 32 accumulator columns [C0, C0+32) of this thread's pixel to the ring: column n = (dy*9 + dx)*3 + co goes to
// ring row 8 - dy of the output pixel 4 - dx lanes away.
template <int C0>
__device__ __forceinline__ void scatter_chunk(const uint32_t (&v)[32], float (&R)[9][3], int lane) {
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        const int n = C0 + j;
        if (n < 243) {
            const int dy = n / 27, dx = (n % 27) / 3, co = n % 3;
            float t = __uint_as_float(v[j]);
            if (dx != 4) t = __shfl_sync(0xffffffffu, t, (lane + dx - 4) & 31);
            R[8 - dy][co] += t;
        }
    }
}

template <int COUT>
__global__ void __launch_bounds__(kScThreads, 1)
conv_scatter_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                    const __grid_constant__ ScatterParams p) {
    b2::pdl_entry();
    static_assert(COUT == 3, "column mapping is written for 3 output channels");
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int cbs = p.cin >> 6;
    uint8_t* smemB = smem;                                  // cbs x 32 KB, resident
    uint8_t* smemA = smem + cbs * kScBBlock;                // ring of 16 KB blocks
    uint64_t* bars = reinterpret_cast<uint64_t*>(smemA + kScAStages * kScABlock);
    uint64_t* afull = bars;                   // [kScAStages]
    uint64_t* aempty = bars + 8;              // [kScAStages]
    uint64_t* bfull = bars + 16;              // weights landed
    uint64_t* tfull = bars + 17;              // [2]
    uint64_t* tempty = bars + 19;             // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 21);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        tma_prefetch_desc(&mapA);
        tma_prefetch_desc(&mapB);
        for (int s = 0; s < kScAStages; ++s) { mbar_init(&afull[s], 1); mbar_init(&aempty[s], 1); }
        mbar_init(bfull, 1);
        for (int s = 0; s < 2; ++s) { mbar_init(&tfull[s], 1); mbar_init(&tempty[s], 4); }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

    const long t_begin = p.total_tiles * blockIdx.x / gridDim.x;
    const long t_end = p.total_tiles * (blockIdx.x + 1) / gridDim.x;
    ScSeg sg;

    if (warp == 0) {
        // ---------------------------------------------------------------- producer: weights once, then input row tiles
        if (lane == 0) {
            mbar_expect_tx(bfull, uint32_t(cbs) * kScBBlock);
            for (int cb = 0; cb < cbs; ++cb) tma_load_2d(smemB + cb * kScBBlock, &mapB, bfull, cb * 64, 0);
        }
        __syncwarp();
        int st = 0; uint32_t ph = 0;
        for (long t = t_begin; sc_seg_next(t, t_end, p, sg);) {
            for (int r = sg.y0 - 4; r < sg.y1 + 4; ++r) {
                for (int cb = 0; cb < cbs; ++cb) {
                    mbar_wait(&aempty[st], ph ^ 1);
                    if (lane == 0) {
                        mbar_expect_tx(&afull[st], kScABlock);
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            tma_load_4d(smemA + st * kScABlock + q * 4096, &mapA, &afull[st], cb * 64,
                                        sg.x0 + kScOutPerWarp * q - 4, r, sg.img);
                    }
                    __syncwarp();
                    if (++st == kScAStages) { st = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ---------------------------------------------------------------- MMA issuer
        const uint32_t idesc = p.f16 ? make_idesc_f16(128, 256) : make_idesc_bf16(128, 256);
        constexpr uint32_t desc_hi = sdesc_hi_k_sw128(1024);
        const uint32_t a_lo0 = sdesc_lo(smem_u32(smemA)), b_lo0 = sdesc_lo(smem_u32(smemB));
        int st = 0; uint32_t ph = 0;
        int as = 0; uint32_t aph = 0;
        mbar_wait(bfull, 0);
        for (long t = t_begin; sc_seg_next(t, t_end, p, sg);) {
            for (int r = sg.y0 - 4; r < sg.y1 + 4; ++r) {
                mbar_wait(&tempty[as], aph ^ 1);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + uint32_t(as * 256);
                for (int cb = 0; cb < cbs; ++cb) {
                    mbar_wait(&afull[st], ph);
                    tc_fence_after();
                    if (elect_one()) {
                        const uint32_t a_lo = a_lo0 + uint32_t(st) * (kScABlock >> 4);
                        const uint32_t b_lo = b_lo0 + uint32_t(cb) * (kScBBlock >> 4);
                        umma_f16_lohi(tmem_d, a_lo, desc_hi, b_lo, desc_hi, idesc, cb != 0);
                        umma_f16_lohi(tmem_d, a_lo + 2, desc_hi, b_lo + 2, desc_hi, idesc, 1);
                        umma_f16_lohi(tmem_d, a_lo + 4, desc_hi, b_lo + 4, desc_hi, idesc, 1);
                        umma_f16_lohi(tmem_d, a_lo + 6, desc_hi, b_lo + 6, desc_hi, idesc, 1);
                        umma_commit(&aempty[st]);
                        if (cb == cbs - 1) umma_commit(&tfull[as]);
                    }
                    __syncwarp();
                    if (++st == kScAStages) { st = 0; ph ^= 1; }
                }
                if (++as == 2) { as = 0; aph ^= 1; }
            }
        }
    } else {
        // ---------------------------------------------------------------- scatter epilogue (warps 2..5)
        const int q = warp & 3;                       // TMEM lane quarter == which 32-pixel box of the tile
        const bool lane_ok = lane >= 4 && lane < 4 + kScOutPerWarp;
        float bias[COUT];
#pragma unroll
        for (int c = 0; c < COUT; ++c) bias[c] = p.bias ? p.bias[c] : 0.f;
        int as = 0; uint32_t aph = 0;
        for (long t = t_begin; sc_seg_next(t, t_end, p, sg);) {
            float R[9][3];
#pragma unroll
            for (int i = 0; i < 9; ++i)
#pragma unroll
                for (int c = 0; c < COUT; ++c) R[i][c] = 0.f;
            const int gx = sg.x0 + kScOutPerWarp * q + lane - 4;      // output pixel this lane finishes
            const bool col_ok = lane_ok && gx < p.W;
            for (int r = sg.y0 - 4; r < sg.y1 + 4; ++r) {
                mbar_wait(&tfull[as], aph);
                tc_fence_after();
                const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(as * 256);
                uint32_t v[32];
                tmem_ld32(taddr + 0, v);   tmem_ld_wait(); scatter_chunk<0>(v, R, lane);
                tmem_ld32(taddr + 32, v);  tmem_ld_wait(); scatter_chunk<32>(v, R, lane);
                tmem_ld32(taddr + 64, v);  tmem_ld_wait(); scatter_chunk<64>(v, R, lane);
                tmem_ld32(taddr + 96, v);  tmem_ld_wait(); scatter_chunk<96>(v, R, lane);
                tmem_ld32(taddr + 128, v); tmem_ld_wait(); scatter_chunk<128>(v, R, lane);
                tmem_ld32(taddr + 160, v); tmem_ld_wait(); scatter_chunk<160>(v, R, lane);
                tmem_ld32(taddr + 192, v); tmem_ld_wait(); scatter_chunk<192>(v, R, lane);
                tmem_ld32(taddr + 224, v); tmem_ld_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tempty[as]);      // accumulator free for the MMA warp
                if (++as == 2) { as = 0; aph ^= 1; }
                scatter_chunk<224>(v, R, lane);
                // output row r - 4 has now received its last contribution (from input row r through dy = 8)
                const int yo = r - 4;
                if (yo >= sg.y0 && yo < sg.y1 && col_ok) {
                    float* o = p.out + ((size_t(sg.img) * p.H + yo) * p.W + gx) * COUT;
#pragma unroll
                    for (int c = 0; c < COUT; ++c) {
                        const float a = R[0][c] + bias[c];
                        o[c] = p.act == ACT_TANH ? tanhf(a) : a;
                    }
                }
#pragma unroll
                for (int i = 0; i < 8; ++i)
#pragma unroll
                    for (int c = 0; c < COUT; ++c) R[i][c] = R[i + 1][c];
#pragma unroll
                for (int c = 0; c < COUT; ++c) R[8][c] = 0.f;
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<512>(tmem_base);
    }
}

int launch_conv_scatter(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int KH, int KW,
                        const float* bias, int act, int f16, float* out, cudaStream_t stream) {
    if (Cin % 64 != 0 || Cin > 64 * kScMaxCb) return set_error("conv_scatter: Cin=%d unsupported", Cin);
    if (KH != 9 || KW != 9 || Cout != 3) return set_error("conv_scatter: filter %dx%d -> %d unsupported", KH, KW, Cout);
    ScatterParams p;
    p.B = B; p.H = H; p.W = W; p.cin = Cin;
    p.strips = (W + kScOutPerTile - 1) / kScOutPerTile;
    p.total_tiles = long(B) * p.strips * H;
    p.act = act; p.f16 = f16; p.bias = bias; p.out = out;
    CUtensorMap mapA, mapB;
    {
        cuuint64_t dims[4] = {cuuint64_t(Cin), cuuint64_t(W), cuuint64_t(H), cuuint64_t(B)};
        cuuint64_t strides[3] = {cuuint64_t(Cin) * 2, cuuint64_t(W) * Cin * 2, cuuint64_t(H) * W * Cin * 2};
        cuuint32_t box[4] = {64, 32, 1, 1};
        cuuint32_t estr[4] = {1, 1, 1, 1};
        int rc = encode_tensor_map(&mapA, x, 4, dims, strides, box, estr);
        if (rc) return rc;
    }
    {
        cuuint64_t dims[2] = {cuuint64_t(Cin), 256};
        cuuint64_t strides[1] = {cuuint64_t(Cin) * 2};
        cuuint32_t box[2] = {64, 256};
        cuuint32_t estr[2] = {1, 1};
        int rc = encode_tensor_map(&mapB, w, 2, dims, strides, box, estr);
        if (rc) return rc;
    }
    const int smem_bytes = (Cin / 64) * kScBBlock + kScAStages * kScABlock + 1024 + 512;
    static int attr_smem = 0;
    if (attr_smem < smem_bytes) {
        cudaError_t e = cudaFuncSetAttribute(conv_scatter_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
        if (e != cudaSuccess) return set_error("cudaFuncSetAttribute(conv_scatter): %s", cudaGetErrorString(e));
        attr_smem = smem_bytes;
    }
    long grid = device_sm_count();
    if (grid > p.total_tiles) grid = p.total_tiles;
    b2::launch_k(conv_scatter_kernel<3>, int(grid), kScThreads, smem_bytes, stream, mapA, mapB, p);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("conv_scatter launch: %s", cudaGetErrorString(e));
    return 0;
}

}  // namespace b2
