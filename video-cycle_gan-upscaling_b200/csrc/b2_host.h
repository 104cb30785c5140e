// b2_host.h — host-side helpers shared by the launchers: error reporting for the C ABI, device queries,
// and TMA tensor-map encoding through the driver entry point (no link-time dependency on libcuda).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

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

// ---------------------------------------------------------------------------------------------------------------------
// Programmatic dependent launch.  A training step is ~360 launches (an adversarial iteration ~800), many of them a few
// microseconds long, so the grid-to-grid hand-off is a visible share of the step.  Every kernel of this library starts
// with pdl_entry() -- wait until the preceding grid of the stream has completed and flushed, THEN allow the next grid of the
// stream to be scheduled -- and every launch goes through launch_k(), which sets programmatic stream serialization: the
// next kernel's CTAs become resident (where registers / shared memory allow) while this one still runs and block in
// griddepcontrol.wait, so the launch latency and CTA scheduling of kernel N+1 overlap kernel N.  Because the trigger follows
// the wait, a kernel can never run ahead of anything it may read, whatever came before it in the stream; nothing precedes
// pdl_entry() in any kernel.  Measured on B200 (profiles/r02_s8_pdl_ab.log) the wait-first form is neutral -- inference
// 11.32 vs 11.19 ms, training step 17.73 vs 18.06 ms, adversarial iteration 32.67 vs 32.77 ms with / without -- so the
// attribute is OFF by default and b2_set_pdl(1) turns it on; griddepcontrol.wait is a no-op in a plain launch.
int pdl_enabled();
int set_pdl_enabled(int enable);

#ifdef __CUDACC__
__device__ __forceinline__ void pdl_entry() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

template <typename K, typename... Args>
inline cudaError_t launch_k(K kernel, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args&&... args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<Args&&>(args)...);
}
#endif

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
