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
/* Switches the CTA-pair (tcgen05 cta_group::2) kernels on or off for subsequent calls; returns the previous setting.
 * Both settings compute the same results; the switch exists for A/B measurements and tests. */
int b2_set_pair_kernels(int enable);
/* Programmatic dependent launch.  Every kernel of the library begins with griddepcontrol.wait (a no-op in a plain launch);
 * b2_set_pdl(1) makes every launch carry programmatic stream serialization, so a kernel's CTAs may become resident while the
 * preceding grid of its stream drains and the launch latency of the ~360 launches of a training step overlaps the previous
 * kernel.  Stream order as the host sees it is unchanged.  Off by default (measured neutral on B200); returns the previous
 * setting. */
int b2_set_pdl(int enable);

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
    /* Backward of ReLU / LeakyReLU fused into a data-gradient convolution: after the activation step the value is
     * multiplied by (gate > 0 ? 1 : gate_slope), gate being the forward OUTPUT of the layer whose input gradient this
     * is (same geometry as out, 16-bit).  Null = no gating. */
    const void* gate;
    float gate_slope;
    /* Train-mode BatchNormalization statistics fused into the producing convolution (model.py:19-20, 22-23, 283-284,
     * 839-872): when set, every CTA adds its share of the per-channel sums of the fp32 values it stores (after scale / shift /
     * activation / residual) and writes them to ITS row stats_part[cta][0..Cout) = sum, [Cout..2*Cout) = sum of squares;
     * rows beyond the launch's grid up to b2_sm_count() are written as zeros, so b2_bn_fwd_finalize(part, b2_sm_count(), ...)
     * follows without a separate statistics pass.  Supported by the stride-1 Cout = 64 kernels (the residual trunk) and the
     * generic implicit-GEMM kernels; entry points that cannot honour it return an error.  Null = off. */
    float* stats_part;
} b2_epilogue;

/* Layout self-description for foreign-function binders (checked against the ctypes mirror in tests/test_abi_layout.py). */
int b2_sizeof_epilogue(void);
int b2_epilogue_field_count(void);
int b2_epilogue_field_offset(int field); /* fields in declaration order; -1 when out of range */

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

/* The same layer (Conv2D(3, 9x9, 'same') + bias + optional tanh, fp32 NHWC output; model.py:290-291) in scatter form:
 * all 81 taps folded into the GEMM N dimension (P[x,(dy,dx,co)] = sum_c X[x,c] W[dy,dx,c,co], N = 243 -> 256, K = Cin),
 * then out[y+4-dy, x+4-dx, co] += P summed in the epilogue through warp shuffles and a register ring of 9 partial
 * rows.  16 full-rate MMAs per 128-pixel row instead of 144 narrow ones; bound by reading x from HBM.
 *   x: [B,H,W,Cin] 16-bit, Cin in {64,128,192,256};  w: packed [256][Cin], row n = (dy*9+dx)*3+co (rows 243..255 zero);
 *   bias: [3] fp32 or null;  out: [B,H,W,3] fp32. */
int b2_conv2d_scatter_fprop(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int kh, int kw,
                            const float* bias, int act, int f16, float* out, void* stream);

/* Same semantics as b2_conv2d_fprop / b2_conv2d_transpose_s2_fprop on CUDA cores, one thread per output element.
 * Cross-check for the tests; not used by the model path. transposed != 0 selects the Conv2DTranspose form. */
int b2_conv2d_direct(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int kh, int kw, int stride,
                     int transposed, const b2_epilogue* ep, void* stream);

/* Gathers `ntaps` shifted copies of a narrow NHWC tensor into the channels of a 64-channel bf16 tensor so that a
 * small-Cin convolution (Cin=3 head conv model.py:275, VGG19 block1_conv1, discriminator block_1) becomes a
 * 64-channel one:  out[b,y,x, t*Cin + c] = in[b, y*stride + dy[t], x*stride + dx[t], c]  (0 outside the image),
 * channels >= ntaps*Cin are zero.  `in` is float when in_fp32 != 0, else bf16.  out: [B,OH,OW,out_ld] with the 64 folded
 * channels written at channel offset out_c0 (out_ld = 64, out_c0 = 0 for a dense result; the backward of the final
 * 9x9 conv, model.py:290, gathers its 81 taps into four 64-channel groups of one 256-channel tensor). */
int b2_fold_taps(const void* in, int in_fp32, int B, int H, int W, int Cin, int OH, int OW, int stride, int ntaps,
                 const int* dy, const int* dx, void* out, int f16, int out_ld, int out_c0, void* stream);

/* Weight (re)packing: out[i] = idx[i] >= 0 ? half_or_bf16(src[idx[i]]) : 0.  The index map is built once per layer on the
 * host (Keras kernel layout (kh,kw,Cin,Cout) / (kh,kw,Cout,Cin) -> packed [Cout][tap*Cin_pad + c], its dgrad
 * transpose, or the tap-folded layouts); this call refreshes the bf16 shadow copy from the fp32 master weights. */
int b2_gather_cast(const float* src, const int32_t* idx, int64_t n, void* out, int f16, void* stream);

/* Frame I/O edges (reference: upscaling/upscaler/data.py:253-270, used around Model.predict in data.py:314-363 and the
 * frame-upscaling notebooks).  b2_u8_to_f32_norm: out[i] = float(in[i] / 127.5 - 1) (convert_image_series_to_array);
 * b2_f32_to_u8_denorm: out[i] = uint8(around((in[i] + 1) * 127.5)), round-half-even (convert_array_to_image).  Both are
 * bit-exact with the NumPy expressions; device buffers 16-byte aligned, n elements. */
int b2_u8_to_f32_norm(const uint8_t* in, float* out, int64_t n, void* stream);
/* Bilinear x`factor` enlargement of an fp32 NHWC image (TensorFlow 1.x resize_bilinear, align_corners = False: what
 * K.resize_images(..., "bilinear") lowers to in make_upscaler_skip_con, reference upscaling/upscaler/model.py:357),
 * written as 16-bit elements into channels [c0, c0 + C) of an NHWC tensor with `ld` channels per pixel -- the
 * Concatenate of model.py:358 becomes a channel slice of the final conv's input. */
int b2_resize_bilinear_into(const float* in, int B, int H, int W, int C, int factor, void* out, int ld, int c0, int f16,
                            void* stream);
int b2_f32_to_u8_denorm(const float* in, uint8_t* out, int64_t n, void* stream);

/* Layers make_upscaler_attention adds to the generator (reference upscaling/upscaler/model.py:30-48, 78-98, 296-328).
 * b2_gate_mul: Multiply([Activation('sigmoid')(attention conv), stream]) (model.py:34-36, 88-91):
 *   out[r, c] = sigmoid(att[r * att_ld + att_c0 + c]) * m[r * C + c], 16-bit in / out, rows pixels, C channels (the
 *   attention convs of all residual blocks are evaluated as one wide convolution; att_c0 selects a block's slice).
 * b2_resize_nearest_bilinear_pair: Concatenate([K.resize_images(x, f, f, 'nearest'), K.resize_images(x, f, f, 'bilinear')])
 *   (model.py:80-82; TensorFlow 1.x kernels, align_corners = False, no half-pixel centres): fp32 [B,H,W,C] -> fp32
 *   [B, f*H, f*W, 2*C].
 * b2_atanh_conv2d_transpose_small: Conv2DTranspose(Cout, kernel_size = stride + 1, strides = stride, padding = 'same') of
 *   atanh(0.99999 * x) on the Cin = 3 network input (model.py:94-95); w fp32 in the Keras layout (k, k, Cout, Cin), bias fp32,
 *   out 16-bit [B, stride*H, stride*W, Cout] (it then enters the block's Conv2DTranspose as b2_epilogue.residual:
 *   Add, model.py:97). */
int b2_gate_mul(const void* att, int att_ld, int att_c0, const void* m, void* out, int64_t rows, int C, int f16, void* stream);
int b2_resize_nearest_bilinear_pair(const float* in, int B, int H, int W, int C, int factor, float* out, void* stream);
int b2_atanh_conv2d_transpose_small(const float* in, int B, int H, int W, int Cin, const float* w, const float* bias, int stride,
                                    int Cout, void* out, int f16, void* stream);

/* Saturation guard of the 16-bit storage: *count += number of elements of x[n] at the edge of their type -- fp16 (f16 != 0):
 * |x| >= 65504 (the value the convolution epilogues' saturating conversion stores for anything larger), inf or NaN; bf16:
 * inf or NaN.  The host runs it over every layer output of the first inference pass after a weight change (and on
 * request), so a network whose activations leave the fp16 range is reported instead of silently clipped. */
int b2_count_saturated(const void* x, int64_t n, int f16, unsigned int* count, void* stream);

/* ---------------------------------------------------------------------------------------------------------------
 * Training step (Model.train_on_batch of the models compile_training_model / make_and_compile_gan build,
 * model.py:1017-1137; call sites train.py:220, train_mini.py:232, train_gan.py:316-317, train_gan3.py:353-354).
 * Data gradients of stride-1 convolutions are b2_conv2d_fprop calls on weights packed flipped/transposed
 * (upscaler/ops.py conv_dgrad_pack_index); the data gradient of Conv2DTranspose is b2_conv2d_fprop with stride 2.
 * ------------------------------------------------------------------------------------------------------------- */

/* Data gradient of Conv2D(strides=2, padding='same') — TF conv2d_backprop_input behind the discriminator convs
 * (model.py:843-871).  dz: [B,OH,OW,Cout] 16-bit; w: packed [Cin][kh*kw*Cout] (un-flipped taps, transposed channels);
 * ep->out: [B,H,W,Cin] where (OH,OW) = ceil((H,W)/2).  Runs as up to four sub-pixel phases. */
int b2_conv2d_s2_dgrad(const void* dz, int B, int OH, int OW, int Cout, const void* w, int Cin, int kh, int kw, int H, int W,
                       const b2_epilogue* ep, void* stream);

/* Conv2D with explicit zero padding and output grid — keras Conv2D(padding='valid', strides 1 or 3) of
 * make_discriminator_sparse_512 (model.py:967-989: pad 0, OH = (H - kh) / stride + 1), and, with pad = k - 1 and
 * flipped/transposed weights, the data gradient of its stride-1 first block.
 *   out[b,oy,ox,:] = sum_{ky,kx} x[b, oy*stride + ky - pad_h, ox*stride + kx - pad_w, :] . w[:, ky, kx, :]   (0 outside x)
 *   x: [B,H,W,Cin] 16-bit, Cin a multiple of 64; w packed [Cout][kh*kw*Cin]; ep->out: [B,OH,OW,Cout]; stride 1..3. */
int b2_conv2d_fprop_pad(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int kh, int kw, int stride,
                        int pad_h, int pad_w, int OH, int OW, const b2_epilogue* ep, void* stream);

/* Data gradient of that convolution for stride 2 or 3 (TF conv2d_backprop_input behind model.py:971-989), run as
 * stride*stride sub-pixel phases.  dz: [B,OH,OW,Cout]; w packed [Cin][kh*kw*Cout] (taps un-flipped, channels
 * transposed: conv_s2_dgrad_pack_index); ep->out: [B,H,W,Cin], every element written (rows the forward never read get 0). */
int b2_conv2d_strided_dgrad(const void* dz, int B, int OH, int OW, int Cout, const void* w, int Cin, int kh, int kw, int stride,
                            int pad_h, int pad_w, int H, int W, const b2_epilogue* ep, void* stream);

/* Weight gradient of that convolution (see b2_conv2d_wgrad): dw[ky][kx][cx][cz] += sum_p x[p*stride + (ky,kx) - pad, cx] *
 * dz[p, cz] over the (OH, OW) grid of dz. */
int b2_conv2d_wgrad_pad(const void* x, int B, int H, int W, int Cx, const void* dz, int OH, int OW, int Cz, int kh, int kw,
                        int stride, int pad_h, int pad_w, float* dw, int f16, void* stream);

/* Weight gradient — TF conv2d_backprop_filter:  dw[ky][kx][cx][cz] += sum_p x[p*stride + (ky,kx) - pad, cx] * dz[p, cz],
 * fp32, ACCUMULATED into dw (zero it first), tcgen05 with the pixel dimension as GEMM K and split-K over pixels.
 *   Conv2D:           x = layer input [B,H,W,Cin],  dz = gradient of the layer output [B,ceil(H/s),ceil(W/s),Cout]
 *                     -> dw in Keras layout (kh,kw,Cin,Cout).
 *   Conv2DTranspose:  x = gradient of the layer output [B,2H,2W,Cout], dz = layer input [B,H,W,Cin], stride 2
 *                     -> dw in Keras layout (kh,kw,Cout,Cin).
 * Cx and Cz must be multiples of 64. */
int b2_conv2d_wgrad(const void* x, int B, int H, int W, int Cx, const void* dz, int Cz, int kh, int kw, int stride, float* dw,
                    int f16, void* stream);

/* Per-channel sums over the rows of a [rows][C] 16-bit tensor: sum[c] += x, sumsq[c] += x^2 (sumsq may be null).
 * Batch statistics of BatchNormalization in the training phase (model.py:20, 23, 284, 840-884) and bias gradients. */
int b2_colstats(const void* x, int64_t rows, int C, int f16, float* sum, float* sumsq, void* stream);

/* mean = sum/count, var = sumsq/count - mean^2 (biased) -> scale = gamma/sqrt(var+eps), shift = beta - mean*scale;
 * optional Keras moving-statistics update (momentum 0.99; variance corrected by n/(n-(1+eps))); optional mean/invstd
 * outputs for the backward pass. */
int b2_bn_finalize(const float* sum, const float* sumsq, float count, const float* gamma, const float* beta, float eps,
                   float momentum, float* moving_mean, float* moving_var, float* scale, float* shift, float* mean_out,
                   float* invstd_out, int C, void* stream);

/* Inference fold of BatchNormalization after a biased conv: scale = gamma/sqrt(var+eps), shift = beta+(bias-mean)*scale. */
int b2_bn_fold(const float* gamma, const float* beta, const float* mean, const float* var, const float* bias, float eps,
               float* scale, float* shift, int C, void* stream);

/* out = act(z*scale + shift) + residual over [rows][C] 16-bit: BatchNorm apply fused with PReLU / LeakyReLU and the
 * residual Add (model.py:20-25, 284-285, 840-841). */
int b2_affine_act(const void* z, int64_t rows, int C, int f16, const float* scale, const float* shift, const float* alpha,
                  int act, float leaky, const void* residual, void* out, void* stream);

/* Backward reductions through act(BN(z)): with u = z*scale+shift, dy = g*act'(u), xhat = (z-mean)*invstd:
 *   s1[c] += dy (= dbeta), s2[c] += dy*xhat (= dgamma), dalpha[c] += g*min(u,0) (PReLU slope gradient; may be null). */
int b2_bn_bwd_reduce(const void* z, const void* g, int64_t rows, int C, int f16, const float* scale, const float* shift,
                     const float* alpha, int act, float leaky, const float* mean, const float* invstd, float* s1, float* s2,
                     float* dalpha, void* stream);

/* dz = scale*(dy - s1/count - xhat*s2/count): gradient w.r.t. the conv output under train-mode BatchNormalization.
 * With s1 == s2 == null (count 0) it is the plain activation backward dz = scale*dy. */
int b2_bn_bwd_apply(const void* z, const void* g, int64_t rows, int C, int f16, const float* scale, const float* shift,
                    const float* alpha, int act, float leaky, const float* mean, const float* invstd, const float* s1,
                    const float* s2, float count, void* dz, void* stream);

/* ---------------------------------------------------------------------------------------------------------------
 * Train-mode BatchNormalization statistics, deterministic two-stage form (what the training step uses; b2_colstats /
 * b2_bn_finalize / b2_bn_bwd_reduce above remain for one-off reductions such as bias gradients).
 * Stage 1 writes partial per-channel sums with plain stores, one row per CTA: part[row][q*C + c].  Producers: the two
 * kernels below, or a convolution's epilogue (b2_epilogue.stats_part, rows = b2_sm_count()).  `*nparts` receives the
 * number of rows written (<= b2_bn_part_rows()).  Stage 2 is one single-CTA launch that adds the rows in a fixed order,
 * optionally sum-exchanges the result with the data-parallel replicas over NVLink peer memory (peer_regions / rank /
 * world / epoch as in b2_peer_allreduce_f32; peer_regions == NULL or world <= 1: no exchange) and finishes the layer.
 * No atomics: results are bit-reproducible and identical on every replica.
 * ------------------------------------------------------------------------------------------------------------- */
int b2_sm_count(void);
int b2_bn_part_rows(void);
/* forward stage 1: part[row] = [sum x | sum x^2] over the rows of a [rows][C] 16-bit tensor. */
int b2_bn_stats_partial(const void* x, int64_t rows, int C, int f16, float* part, int* nparts, void* stream);
/* forward stage 2: sums (x world) -> mean, biased variance -> aff = [scale | shift | mean | invstd] (4*C floats, owned by
 * the caller's tape), Keras moving-statistics update when moving_mean / moving_var are given (momentum 0.99, variance
 * n/(n-(1+eps))).  `count` = rows over ALL replicas.  sums_out (2*C, nullable) receives the summed statistics; with
 * aff == NULL nothing else is computed (NCCL fallback: all-reduce sums_out, call again with nparts = 1).
 * range_watch (nullable, one float): set to 1 when a channel's root mean square reaches range_limit or is not finite --
 * the guard of fp16 training storage (the tensor was stored with a saturating conversion; the host switches to bf16). */
int b2_bn_fwd_finalize(const float* part, int nparts, int C, float count, void* const* peer_regions, int rank, int world,
                       unsigned epoch, const unsigned* epoch_base, float* sums_out, const float* gamma, const float* beta, float eps,
                       float momentum, float* moving_mean, float* moving_var, float* aff, float* range_watch, float range_limit,
                       void* stream);
/* backward stage 1 through y = act(u), u = z*scale + shift (aff from the forward tape): with dy = g*act'(u),
 * part[row] = [sum dy | sum dy*xhat | sum g*min(u,0)]. */
int b2_bn_bwd_partial(const void* z, const void* g, int64_t rows, int C, int f16, const float* aff, const float* alpha, int act,
                      float leaky, float* part, int* nparts, void* stream);
/* backward stage 2: sums = [s1 | s2] over all replicas (input of b2_bn_bwd_apply); dbeta += s1*gscale, dgamma += s2*gscale,
 * dalpha += this replica's third sum (each nullable).  raw_out != NULL: only write the 3*C local sums there and return
 * (NCCL fallback; call again on the all-reduced buffer with nparts = 1). */
int b2_bn_bwd_finalize(const float* part, int nparts, int C, void* const* peer_regions, int rank, int world, unsigned epoch,
                       const unsigned* epoch_base, float* raw_out, float* sums, float* dgamma, float* dbeta, float* dalpha, float gscale, void* stream);

/* Forward BatchNormalization statistics for callers that cannot allocate device memory.  (No counterpart in the
 * reference: Keras / TensorFlow allocate freely; this serves hosts that embed the library.)
 * Same arithmetic and same bits as b2_bn_stats_partial followed by a single-replica b2_bn_fwd_finalize (no moving-statistics
 * update, no range watch), but the per-CTA partial rows live in a 1 212 416-byte array compiled into the library
 * (296 rows x 2 x 512 floats, resident on every GPU that loads it until the process exits) instead of a tensor the caller
 * has to provide.  Meant for a plugin inside another framework's memory arena, a real-time frame loop that must not stall
 * in the allocator, or a process whose device-memory quota is already committed: those callers own x and the outputs but
 * have nowhere to get the scratch rows from.  The call performs no cudaMalloc / cudaFree and uses no allocating library.
 *   x: [rows][C] 16-bit (f16 = 1 fp16, 0 bf16), C a multiple of 8 with 256 % (C/8) == 0 and 8 <= C <= 512; wider layers are
 *      refused with an error (b2_last_error) -- use the two-call form with your own buffer for those.
 *   sums_out (2*C floats, nullable): [sum x | sum x^2].  aff (4*C floats, nullable; needs gamma, beta): [scale | shift |
 *      mean | invstd] for `count` elements per channel.  At least one of the two must be given.
 * The first call on each device looks up the array's device address (cudaGetSymbolAddress, with that device current) and
 * keeps it; make that first call outside any stream capture.  All calls share the one array: nothing has to be cleared
 * between calls (every row stage 2 reads is rewritten by stage 1 of the same call), but calls that run concurrently on one
 * GPU -- two streams, two threads -- overwrite each other's rows.  Issue them in order on a device. */
int b2_bn_fwd_stats_static(const void* x, int64_t rows, int C, int f16, float count, const float* gamma, const float* beta,
                           float eps, float* sums_out, float* aff, void* stream);

/* MaxPooling2D(2,2) 'valid' of VGG19 (keras.applications behind model.py:108) and its gradient (first maximum of each
 * window takes the gradient; relu_gate != 0 additionally zeroes it where x <= 0, i.e. the ReLU of the conv below). */
int b2_maxpool2x2(const void* x, int B, int H, int W, int C, int f16, void* y, void* stream);
int b2_maxpool2x2_bwd(const void* x, const void* y, const void* dy, int B, int H, int W, int C, int f16, int relu_gate,
                      void* dx, void* stream);

/* Content-loss reductions (model.py:117, 137, 157): *loss += weight * mean(f(a-b)), f = square (mae=0) or abs (mae=1);
 * grad (nullable, same type as a) = gscale * f'(a-b) / n.  dtype: 0 bf16, 1 fp16, 2 fp32. */
int b2_loss_pair(const void* a, const void* b, int64_t n, int dtype, int mae, float weight, float gscale, float* loss,
                 void* grad, void* stream);

/* Wasserstein-type reductions (model.py:159-160, 215-261): *loss += weight * mean(x*y) with y an array or the constant
 * yconst; grad[i] = gscale * y[i] / n.  fp32, n = batch size. */
int b2_mean_prod(const float* x, const float* y, float yconst, int64_t n, float weight, float gscale, float* loss, float* grad,
                 void* stream);

/* Discriminator output activation (model.py:885-892, 950-957, 1003-1010; train_gan3.py -da) on the (n,1) scores.
 * kind: 1 sigmoid, 2 log-sigm, 3 tanh, 4 bi-log.  dy == NULL: out = f(z); otherwise out = dy * f'(z).  fp32. */
int b2_score_act(const float* z, const float* dy, float* out, int64_t n, int kind, void* stream);

/* Backward of the generator's final tanh (model.py:291): dz = (g1 + g2 + g3) * (1 - y*y); g2, g3 may be null. fp32. */
int b2_tanh_bwd(const float* y, const float* g1, const float* g2, const float* g3, int64_t n, float* dz, void* stream);

/* sum[c] += sum over rows of x[rows][C], fp32, C <= 8 (bias gradient of the 3-channel final conv, model.py:290). */
int b2_colsum_f32(const float* x, int64_t rows, int C, float* sum, void* stream);

/* ---------------------------------------------------------------------------------------------------------------
 * Data-parallel training: sum-all-reduce of small fp32 vectors (the cross-replica BatchNormalization sums,
 * model.py:20, 23, 284, 840-872 seen by N replicas) over NVLink peer memory, one single-CTA launch on the compute
 * stream.  Every rank allocates one region (b2_peer_alloc), sends the 64-byte CUDA IPC handle to its peers (the host
 * side uses torch.distributed for that), opens theirs (b2_peer_open) and passes the `world` region pointers, indexed by
 * rank, to every call.  `epoch` starts at 1 and increases by one per call, identically on all ranks.  n <= b2_peer_max_elems().
 * ------------------------------------------------------------------------------------------------------------- */
int b2_peer_max_elems(void);
int b2_peer_alloc(void** region, unsigned char* handle64);
int b2_peer_open(const unsigned char* handle64, void** region);
int b2_peer_close(void* region, int opened);
/* Measurement aid (tools/peer_probe.py): `iters` back-to-back exchanges of n floats inside ONE launch, epochs epoch0 ..
 * epoch0 + iters - 1 (the caller advances its epoch counter by iters); cycles[0] receives the SM clock cycles the loop took,
 * cycles[1] the SM clock in kHz.  Isolates the exchange latency from launch latency and host effects. */
int b2_peer_exchange_probe(float* data, int n, void* const* peer_regions, int rank, int world, unsigned epoch0, int iters,
                           long long* cycles, void* stream);
/* epoch_base (device pointer, nullable; also on b2_bn_fwd_finalize / b2_bn_bwd_finalize): when the launch is captured into a
 * CUDA graph its arguments are baked, so the exchange uses epoch + *epoch_base -- `epoch` = the launch's position inside the
 * captured step (1, 2, ...), *epoch_base = the number of exchanges completed before the replay, set by the host on the same
 * stream before every replay.  NULL in a plain launch. */
int b2_peer_allreduce_f32(float* data, int n, void* const* regions, int rank, int world, unsigned epoch, const unsigned* epoch_base,
                          void* stream);

/* Dense layers of the discriminator head (model.py:876-884; batch-sized M, negligible FLOPs) on CUDA cores, reading
 * the fp32 master weights in Keras layout W[K][N] directly.  x / dy / dx / out are 16-bit unless the dtype code
 * (0 bf16, 1 fp16, 2 fp32) says otherwise.
 *   fwd:        out[n][o] = sum_i x[n][i] W[i][o] + b[o]
 *   bwd_data:   dx[n][i]  = sum_o dy[n][o] W[i][o]
 *   bwd_weight: dW[i][o] += sum_n x[n][i] dy[n][o];  db[o] += sum_n dy[n][o] */
int b2_dense_fwd(const void* x, int x_dtype, int n, int K, int N, const float* W, const float* b, void* out, int out_dtype,
                 void* stream);
int b2_dense_bwd_data(const void* dy, int dy_dtype, int n, int K, int N, const float* W, void* dx, int dx_dtype, void* stream);
int b2_dense_bwd_weight(const void* x, int x_dtype, const void* dy, int dy_dtype, int n, int K, int N, float* dW, float* db,
                        void* stream);

/* out[i] (+)= idx[i] >= 0 ? src[idx[i]] : 0 in fp32 (weight gradients of tap-folded layers back to the Keras layout). */
int b2_gather_f32(const float* src, const int32_t* idx, int64_t n, float* out, int accumulate, void* stream);

/* out = cast(in * scale); dtype codes as in b2_loss_pair. */
int b2_cast_scale(const void* in, int in_dtype, int64_t n, float scale, void* out, int out_dtype, void* stream);

/* keras.optimizers.Adam update over a flat fp32 parameter buffer (model.py:1026, 1066, 1130):
 *   g *= gscale; m = b1*m + (1-b1)*g; v = b2*v + (1-b2)*g*g; p -= lr_t * m / (sqrt(v) + eps),
 * lr_t = lr*sqrt(1-b2^t)/(1-b1^t) computed by the caller. */
int b2_adam(float* params, const float* grads, float* m, float* v, int64_t n, float lr_t, float beta1, float beta2, float eps,
            float gscale, void* stream);
/* The same update behind a device-side guard: guard0 / guard1 (each nullable) point at `guard_words` 32-bit words; if any
 * word is non-zero when the kernel runs, nothing is written (parameters, m, v keep their bits).  The training step points
 * them at the words its fp16 passes raise -- the BatchNormalization range flag (b2_bn_fwd_finalize range_watch) and the
 * saturation count of sampled activation gradients (b2_count_saturated) -- so an update computed from clipped 16-bit
 * values is dropped without a host round trip; the host reads the words with the step's loss and switches storage type.
 * (Keras has no counterpart: the reference trains in fp32, model.py:1026-1137.) */
int b2_adam_guarded(float* params, const float* grads, float* m, float* v, int64_t n, float lr_t, float beta1, float beta2, float eps,
                    float gscale, const void* guard0, const void* guard1, int guard_words, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B2SRGAN_H_ */
