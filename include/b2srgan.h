/* b2srgan.h — C ABI of the B200-native SRGAN upscaler hot path (libb2srgan.so).
 *
 * The reference (kjedrzejewski/video-cycle_gan-upscaling) has no FFI: every FLOP of its hot path runs inside Keras /
 * TensorFlow library kernels reached from upscaling/upscaler/model.py.  Each entry point below replaces the
 * TensorFlow op(s) that one Keras layer / loss of that file lowers to; the Python host code in
 * video-cycle_gan-upscaling_b200/upscaler/ mirrors model.py's builders and calls these through ctypes.
 *
 * Conventions: every pointer is a DEVICE pointer owned by the caller (PyTorch tensors on the host side); activations
 * are NHWC bf16 unless a parameter says otherwise; per-channel vectors are fp32; all launches go to the `stream`
 * argument (a cudaStream_t passed as void*); functions return 0 on success and a negative value on error, the
 * message being available from b2_last_error() (thread-local).  No hidden allocations, no global state apart from
 * per-kernel attribute caching.
 */
#ifndef B2SRGAN_H_
#define B2SRGAN_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Activations and packed weights are 16-bit: bf16 (training; gradient range) or IEEE fp16 (inference: 3 more
 * mantissa bits are what it takes to meet the 2e-2 max-abs bar through 37 conv layers).  Both run the
 * same tcgen05 kind::f16 instruction at the same rate; the `f16` flags below select the element type per call. */

/* activation codes of the convolution epilogue */
enum { B2_ACT_NONE = 0, B2_ACT_RELU = 1, B2_ACT_LEAKY = 2, B2_ACT_PRELU = 3, B2_ACT_TANH = 4 };

const char* b2_last_error(void);
int b2_version(void);

/* Epilogue shared by every convolution entry point:
 *   out = act(acc * scale[c] + shift[c]) + residual
 * scale/shift carry the bias and (at inference) the folded BatchNormalization (model.py:20, 23, 284); alpha is the
 * PReLU slope per channel (model.py:21, 276); leaky the LeakyReLU slope (model.py:73, 841). Null pointers mean
 * scale=1, shift=0, no residual. */
typedef struct {
    const float* scale;
    const float* shift;
    const float* alpha;
    int act;
    float leaky;
    const void* residual; /* bf16, same geometry as out */
    void* out;            /* bf16, or float when out_fp32 != 0 */
    int out_fp32;
    int f16; /* 16-bit element type of x / w / residual / out: 0 = bf16, 1 = IEEE fp16 */
} b2_epilogue;

/* Conv2D(padding='same') forward on tcgen05 tensor cores — replaces the TF conv2d behind keras.layers.Conv2D at
 * model.py:19, 22, 275, 283, 290, 839-871 and the VGG19 convs behind model.py:108-113.
 *   x: [B,H,W,Cin] bf16, Cin a multiple of 64.   w: packed [Cout][kh*kw*Cin] bf16 (see b2_pack_conv_weights).
 *   out: [B,ceil(H/stride),ceil(W/stride),Cout].  TensorFlow SAME padding: pad_before = pad_total / 2. */
int b2_conv2d_fprop(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int kh, int kw, int stride,
                    const b2_epilogue* ep, void* stream);

/* Conv2DTranspose(strides=2, padding='same') forward — replaces TF conv2d_backprop_input behind
 * keras.layers.Conv2DTranspose at model.py:72.  Runs as four sub-pixel phases of dense convolutions.
 *   x: [B,H,W,Cin] bf16;  w: packed [Cout][kh*kw*Cin] bf16 (b2_pack_convT_weights);  out: [B,2H,2W,Cout]. */
int b2_conv2d_transpose_s2_fprop(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int kh,
                                 int kw, const b2_epilogue* ep, void* stream);

/* Conv2D(Cout=3, 9 x KW, 'same', stride 1) + bias + optional tanh, fp32 NHWC output — the generator's final/conv +
 * final/tanh (model.py:290-291).  Row-streaming kernel: filter rows folded into the GEMM N dimension, filter columns
 * into K through descriptor shifts, every input element fetched once.
 *   x: [B,H,W,Cin] 16-bit, Cin in {64,128,192,256};  w: packed [32][KW*Cin] with row n = dy*3 + co (rows 27..31 zero),
 *   column k = dx*Cin + c;  bias: [3] fp32 or null;  out: [B,H,W,3] fp32. */
int b2_conv2d_rows_fprop(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int kh, int kw,
                         const float* bias, int act, int f16, float* out, void* stream);

/* Same semantics as b2_conv2d_fprop / b2_conv2d_transpose_s2_fprop on CUDA cores, one thread per output element.
 * Cross-check for the tests; not used by the model path. transposed != 0 selects the Conv2DTranspose form. */
int b2_conv2d_direct(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int kh, int kw, int stride,
                     int transposed, const b2_epilogue* ep, void* stream);

/* Gathers `ntaps` shifted copies of a narrow NHWC tensor into the channels of a 64-channel bf16 tensor so that a
 * small-Cin convolution (Cin=3 head conv model.py:275, VGG19 block1_conv1, discriminator block_1) becomes a
 * 64-channel one:  out[b,y,x, t*Cin + c] = in[b, y*stride + dy[t], x*stride + dx[t], c]  (0 outside the image),
 * channels >= ntaps*Cin are zero.  `in` is float when in_fp32 != 0, else bf16.  out: [B,OH,OW,64]. */
int b2_fold_taps(const void* in, int in_fp32, int B, int H, int W, int Cin, int OH, int OW, int stride, int ntaps,
                 const int* dy, const int* dx, void* out, int f16, void* stream);

/* Weight (re)packing: out[i] = idx[i] >= 0 ? half_or_bf16(src[idx[i]]) : 0.  The index map is built once per layer on the
 * host (Keras kernel layout (kh,kw,Cin,Cout) / (kh,kw,Cout,Cin) -> packed [Cout][tap*Cin_pad + c], its dgrad
 * transpose, or the tap-folded layouts); this call refreshes the bf16 shadow copy from the fp32 master weights. */
int b2_gather_cast(const float* src, const int32_t* idx, int64_t n, void* out, int f16, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B2SRGAN_H_ */
