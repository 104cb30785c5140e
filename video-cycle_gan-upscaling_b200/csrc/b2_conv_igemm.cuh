// b2_conv_igemm.cuh — parameter block shared by the tcgen05 implicit-GEMM convolution kernels and their host
// launchers. One "tap" is one (dy, dx) offset of the filter; the GEMM K dimension is taps x input channels and
// is walked tap by tap in blocks of 64 channels (one 128-byte swizzle row per pixel).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace b2 {

constexpr int kMaxTaps = 96;  // 9x9 = 81 is the largest filter on the path (model.py:275, 290)

enum Act : int { ACT_NONE = 0, ACT_RELU = 1, ACT_LEAKY = 2, ACT_PRELU = 3, ACT_TANH = 4 };

struct IgemmParams {
    // pixel grid the M dimension iterates over (the output grid of a conv; the input grid of one
    // sub-pixel phase of a stride-2 transposed conv)
    int B, H, W;
    int TB, TH, TW;  // tile box, TB*TH*TW <= 128 rows of the UMMA M=128 tile
    int tiles_x, tiles_y, tiles_b, tiles_n, num_tiles;
    int cin;         // input channels (multiple of 64)
    int num_taps;
    int in_mul;      // input coordinate = grid coordinate * in_mul + d (2 for a stride-2 conv)
    int8_t dy[kMaxTaps];
    int8_t dx[kMaxTaps];
    int16_t wtap[kMaxTaps];  // index of the tap's [Cout x Cin] slab in the packed weight matrix
    // output tensor (NHWC) and where a grid pixel lands in it
    int OH, OW, out_sy, out_oy, out_sx, out_ox;
    int out_ld;   // channels per output pixel (row pitch in elements)
    int n_valid;  // real output channels (columns >= n_valid are computed but never stored)
    int act;
    float leaky;
    const float* scale;  // per output channel, may be null (=1)
    const float* shift;  // per output channel, may be null (=0)
    const float* alpha;  // PReLU slope per channel (act == ACT_PRELU)
    const __nv_bfloat16* residual;  // added after the activation; same geometry as out; may be null
    const void* gate;     // backward of ReLU / LeakyReLU fused into a dgrad: value *= (gate > 0 ? 1 : gate_slope); same
    float gate_slope;     //   geometry and element type as out's 16-bit form; may be null
    float* stats_part;    // train-mode BatchNorm statistics fused into the epilogue: per-CTA rows [2 * n_valid]; may be null
    int stats_rows;       //   rows the producer must leave valid (b2_bn_part_rows(); rows beyond the grid are zeroed)
    void* out;
    int out_fp32;  // 1: out is float, 0: out is the 16-bit element type
    int f16;       // 16-bit element type of x / w / residual / out: 0 = bf16, 1 = fp16 (IEEE half)
};

}  // namespace b2
