// b2_host.h — host-side helpers shared by the launchers: error reporting for the C ABI, device queries,
// and TMA tensor-map encoding through the driver entry point (no link-time dependency on libcuda).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "b2_conv_igemm.cuh"

// A/B switches read from the environment exist only in profiling builds (B2_BUILD_PROFILING=1 python -m upscaler._build ->
// -DB2_PROFILING, used by tools/): the shipped library reads no environment variables and has no hidden state.
#ifdef B2_PROFILING
#include <stdlib.h>
#define B2_PROF_ENV(name) getenv(name)
#else
#define B2_PROF_ENV(name) (static_cast<const char*>(nullptr))
#endif

namespace b2 {

// Records a printf-style message retrievable through b2_last_error() and returns -1.
int set_error(const char* fmt, ...);
const char* last_error();

int device_sm_count();

// rank-`rank` bf16 tensor, 128-byte swizzle, zero fill for out-of-bounds elements.
int encode_tensor_map(CUtensorMap* map, const void* base, int rank, const cuuint64_t* dims,
                      const cuuint64_t* strides_bytes /* rank-1 entries */, const cuuint32_t* box,
                      const cuuint32_t* elem_strides, int swizzle_bytes = 128 /* 128 or 64 */);

struct ConvIgemmArgs {
    const __nv_bfloat16* x;  // input NHWC
    int xB, xH, xW, xC;
    const __nv_bfloat16* w;  // packed weights [w_rows][w_k], k = tap*Cin + cin contiguous
    int w_rows, w_k;
    IgemmParams p;  // grid, taps, output mapping and epilogue filled by the caller; tiling filled by the launcher
};

int launch_conv_igemm(const ConvIgemmArgs& a, cudaStream_t stream);
// b2_bn_train.cu: [sum | sum of squares] of a finished [rows][C] 16-bit tensor into all 2 * SM-count rows of `part`.
int launch_bn_stats_all_rows(const void* x, long rows, int C, int f16, float* part, cudaStream_t stream);
// CUDA-core cross-check with identical semantics (tests only).
int launch_conv_direct(const ConvIgemmArgs& a, cudaStream_t stream);

// Halo-tile variant (b2_conv_halo.cu): input tile staged once in shared memory, taps as shifted descriptor views.
bool conv_halo_applicable(const ConvIgemmArgs& a);
int launch_conv_halo(const ConvIgemmArgs& a, cudaStream_t stream);

// Row-stacked variant for Cout = 64 (b2_conv_stack.cu): R vertically adjacent output rows share one accumulator row, so
// the MMAs run at N = 64..192 instead of 64.
bool conv_stack_applicable(const ConvIgemmArgs& a);
int launch_conv_stack(const ConvIgemmArgs& a, cudaStream_t stream);

// All sub-pixel phases of a stride-2 transposed conv with Cout = 256 in one persistent kernel (b2_conv_phases.cu).
bool conv_phases_applicable(const ConvIgemmArgs* phases, int nph);
int launch_conv_phases(const ConvIgemmArgs* phases, int nph, cudaStream_t stream);
// The same over CTA pairs (tcgen05 cta_group::2, b2_conv_phases2.cu); returns -2 when its rings do not fit.
bool conv_phases2_applicable(const ConvIgemmArgs* phases, int nph);
int launch_conv_phases2(const ConvIgemmArgs* phases, int nph, cudaStream_t stream);

// Weight gradient (b2_conv_wgrad.cu): dw[wtap[t]][cx][cz] += sum_p x[p*in_mul + d(t), cx] * dz[p, cz], fp32 accumulate.
struct WgradArgs {
    const void* x;   // [xB, xH, xW, cx] 16-bit
    int xB, xH, xW, cx;
    const void* dz;  // [zB, zH, zW, cz] 16-bit
    int zB, zH, zW, cz;
    int num_taps, in_mul;
    int8_t dy[kMaxTaps], dx[kMaxTaps];
    int16_t wtap[kMaxTaps];
    float* dw;
    int f16;
};
int launch_conv_wgrad(const WgradArgs& a, cudaStream_t stream);

// Final 9xKW conv with Cout=3 as a row-streaming kernel (b2_conv_tail.cu).
int launch_conv_tail(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int KH, int KW,
                     const float* bias, int act, int f16, float* out, cudaStream_t stream);

// Final 9x9 conv with Cout=3 as one N=256 GEMM per pixel row + register scatter epilogue (b2_conv_scatter.cu).
int launch_conv_scatter(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int KH, int KW,
                        const float* bias, int act, int f16, float* out, cudaStream_t stream);

}  // namespace b2
