// b2_api.cu — the extern "C" boundary (include/b2srgan.h): turns layer-level geometry (kernel size, stride,
// TensorFlow "same" padding, Conv2DTranspose phases) into tap lists for the implicit-GEMM kernels and launches them.
#include <stddef.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/b2srgan.h"
#include "b2_host.h"

using namespace b2;

// CTA-pair (cta_group::2) kernels on/off: A/B switch for tests and profiling (env B2_NO_PAIR=1 starts with them off)
static int g_pair_kernels = B2_PROF_ENV("B2_NO_PAIR") ? 0 : 1;

static void fill_epilogue(IgemmParams& p, const b2_epilogue* ep, int Cout) {
    p.scale = ep->scale;
    p.shift = ep->shift;
    p.alpha = ep->alpha;
    p.act = ep->act;
    p.leaky = ep->leaky;
    p.residual = reinterpret_cast<const __nv_bfloat16*>(ep->residual);
    p.gate = ep->gate;
    p.gate_slope = ep->gate_slope;
    p.stats_part = ep->stats_part;
    p.stats_rows = 2 * device_sm_count();
    p.out = ep->out;
    p.out_fp32 = ep->out_fp32;
    p.f16 = ep->f16;
    p.out_ld = Cout;
    p.n_valid = Cout;
}

// TensorFlow SAME: out = ceil(in/s); pad_total = max((out-1)*s + k - in, 0); pad_before = pad_total / 2.
static void tf_same(int in, int k, int s, int& out, int& pad_before) {
    out = (in + s - 1) / s;
    int total = (out - 1) * s + k - in;
    if (total < 0) total = 0;
    pad_before = total / 2;
}

static int conv_taps(IgemmParams& p, int B, int H, int W, int Cin, int kh, int kw, int stride) {
    if (kh * kw > kMaxTaps) return set_error("conv: %dx%d filter exceeds %d taps", kh, kw, kMaxTaps);
    int OH, OW, pbh, pbw;
    tf_same(H, kh, stride, OH, pbh);
    tf_same(W, kw, stride, OW, pbw);
    p.B = B; p.H = OH; p.W = OW;
    p.cin = Cin;
    p.in_mul = stride;
    p.num_taps = kh * kw;
    for (int ky = 0; ky < kh; ++ky)
        for (int kx = 0; kx < kw; ++kx) {
            const int t = ky * kw + kx;
            p.dy[t] = int8_t(ky - pbh);
            p.dx[t] = int8_t(kx - pbw);
            p.wtap[t] = int16_t(t);
        }
    p.OH = OH; p.OW = OW;
    p.out_sy = 1; p.out_sx = 1; p.out_oy = 0; p.out_ox = 0;
    return 0;
}

// Explicit-padding form (Conv2D(padding='valid'), and the 'full' correlation that is its stride-1 data gradient):
// out[oy] = sum_ky x[oy*stride + ky - pad]; the output grid (OH, OW) is the caller's.
static int conv_taps_pad(IgemmParams& p, int B, int Cin, int kh, int kw, int stride, int pad_h, int pad_w, int OH, int OW) {
    if (kh * kw > kMaxTaps) return set_error("conv: %dx%d filter exceeds %d taps", kh, kw, kMaxTaps);
    if (pad_h < 0 || pad_w < 0 || pad_h > 100 || pad_w > 100) return set_error("conv: bad padding %d,%d", pad_h, pad_w);
    p.B = B; p.H = OH; p.W = OW;
    p.cin = Cin;
    p.in_mul = stride;
    p.num_taps = kh * kw;
    for (int ky = 0; ky < kh; ++ky)
        for (int kx = 0; kx < kw; ++kx) {
            const int t = ky * kw + kx;
            p.dy[t] = int8_t(ky - pad_h);
            p.dx[t] = int8_t(kx - pad_w);
            p.wtap[t] = int16_t(t);
        }
    p.OH = OH; p.OW = OW;
    p.out_sy = 1; p.out_sx = 1; p.out_oy = 0; p.out_ox = 0;
    return 0;
}

// Data gradient of a stride-s convolution as s*s sub-pixel phases: x position y = s*j + py receives
// sum over ky = py + pad - s*d of dz[j + d] * W[ky]  (the stride-2 SAME case above is convT_phase_taps).
static void strided_phase_taps(IgemmParams& p, int B, int H, int W, int Cz, int kh, int kw, int s, int pad_h, int pad_w,
                               int py, int px) {
    p.B = B; p.H = (H - py + s - 1) / s; p.W = (W - px + s - 1) / s;
    p.cin = Cz;
    p.in_mul = 1;
    int nt = 0;
    for (int ky = 0; ky < kh; ++ky) {
        const int ny = py + pad_h - ky;
        if (((ny % s) + s) % s != 0) continue;
        for (int kx = 0; kx < kw; ++kx) {
            const int nx = px + pad_w - kx;
            if (((nx % s) + s) % s != 0) continue;
            p.dy[nt] = int8_t(ny >= 0 ? ny / s : -((-ny) / s));
            p.dx[nt] = int8_t(nx >= 0 ? nx / s : -((-nx) / s));
            p.wtap[nt] = int16_t(ky * kw + kx);
            ++nt;
        }
    }
    p.num_taps = nt;
    p.OH = H; p.OW = W;
    p.out_sy = s; p.out_sx = s; p.out_oy = py; p.out_ox = px;
}

// Conv2DTranspose(stride 2, same) is the input-gradient of a SAME stride-2 conv: out[y] = sum_i x[i] * W[y - 2i + pb],
// pb = the forward conv's pad_before ((k-2)/2 when out = 2*in).  For output parity py: y = 2j + py, i = j + d
// =>  ky = py + pb - 2d.  The same tap lists give the data gradient of a stride-2 Conv2D (discriminator, model.py:843-871)
// with (Hout, Wout) = the forward conv's input size, which need not be even.
static void convT_phase_taps(IgemmParams& p, int B, int Hout, int Wout, int Cin, int kh, int kw, int pbh, int pbw, int py,
                             int px) {
    p.B = B; p.H = (Hout - py + 1) / 2; p.W = (Wout - px + 1) / 2;
    p.cin = Cin;
    p.in_mul = 1;
    int nt = 0;
    for (int ky = 0; ky < kh; ++ky) {
        if (((py + pbh - ky) & 1) != 0) continue;
        for (int kx = 0; kx < kw; ++kx) {
            if (((px + pbw - kx) & 1) != 0) continue;
            p.dy[nt] = int8_t((py + pbh - ky) / 2);
            p.dx[nt] = int8_t((px + pbw - kx) / 2);
            p.wtap[nt] = int16_t(ky * kw + kx);
            ++nt;
        }
    }
    p.num_taps = nt;
    p.OH = Hout; p.OW = Wout;
    p.out_sy = 2; p.out_sx = 2; p.out_oy = py; p.out_ox = px;
}

// Launches the (up to) four sub-pixel phases of a stride-2 transposed convolution; phases without taps or pixels write
// nothing (their outputs are zero only through the bias-free epilogue, so callers needing them pre-zero `out`).
static int run_transposed(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int kh, int kw, int Hout,
                          int Wout, int pbh, int pbw, const b2_epilogue* ep, void* stream, bool direct) {
    if (ep->stats_part) return set_error("b2_epilogue.stats_part is only honoured by b2_conv2d_fprop");
    ConvIgemmArgs args[4];
    int n = 0;
    bool all = true;
    for (int ph = 0; ph < 4; ++ph) {
        ConvIgemmArgs& a = args[n];
        memset(&a, 0, sizeof(a));
        convT_phase_taps(a.p, B, Hout, Wout, Cin, kh, kw, pbh, pbw, ph >> 1, ph & 1);
        if (a.p.H <= 0 || a.p.W <= 0) { all = false; continue; }
        fill_epilogue(a.p, ep, Cout);
        a.x = reinterpret_cast<const __nv_bfloat16*>(x);
        a.xB = B; a.xH = H; a.xW = W; a.xC = Cin;
        a.w = reinterpret_cast<const __nv_bfloat16*>(w);
        a.w_rows = Cout; a.w_k = kh * kw * Cin;
        if (a.p.num_taps == 0) return set_error("transposed conv: phase %d has no taps (filter %dx%d)", ph, kh, kw);
        ++n;
    }
    static const bool no_fuse = B2_PROF_ENV("B2_NO_PHASE_FUSION") != nullptr;      // A/B switch for profiling
    if (!direct && !no_fuse && all) {
        // CTA-pair kernel first (half the weight traffic per SM); -2 = its rings do not fit this geometry
        if (g_pair_kernels && conv_phases2_applicable(args, n)) {
            const int rc = launch_conv_phases2(args, n, reinterpret_cast<cudaStream_t>(stream));
            if (rc != -2) return rc;
        }
        if (conv_phases_applicable(args, n)) return launch_conv_phases(args, n, reinterpret_cast<cudaStream_t>(stream));
    }
    for (int i = 0; i < n; ++i) {
        int rc = direct ? launch_conv_direct(args[i], reinterpret_cast<cudaStream_t>(stream))
                        : launch_conv_igemm(args[i], reinterpret_cast<cudaStream_t>(stream));
        if (rc) return rc;
    }
    return 0;
}

extern "C" {

const char* b2_last_error(void) { return last_error(); }
int b2_version(void) { return 2; }

int b2_sizeof_epilogue(void) { return int(sizeof(b2_epilogue)); }
int b2_epilogue_field_count(void) { return 12; }
int b2_epilogue_field_offset(int field) {
    switch (field) {
        case 0: return int(offsetof(b2_epilogue, scale));
        case 1: return int(offsetof(b2_epilogue, shift));
        case 2: return int(offsetof(b2_epilogue, alpha));
        case 3: return int(offsetof(b2_epilogue, act));
        case 4: return int(offsetof(b2_epilogue, leaky));
        case 5: return int(offsetof(b2_epilogue, residual));
        case 6: return int(offsetof(b2_epilogue, out));
        case 7: return int(offsetof(b2_epilogue, out_fp32));
        case 8: return int(offsetof(b2_epilogue, f16));
        case 9: return int(offsetof(b2_epilogue, gate));
        case 10: return int(offsetof(b2_epilogue, gate_slope));
        case 11: return int(offsetof(b2_epilogue, stats_part));
        default: return -1;
    }
}

int b2_set_pdl(int enable) { return set_pdl_enabled(enable); }

int b2_set_pair_kernels(int enable) {
    const int old = g_pair_kernels;
    g_pair_kernels = enable ? 1 : 0;
    return old;
}

int b2_conv2d_fprop(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int kh, int kw, int stride,
                    const b2_epilogue* ep, void* stream) {
    if (!x || !w || !ep || !ep->out) return set_error("b2_conv2d_fprop: null pointer");
    if (stride != 1 && stride != 2) return set_error("b2_conv2d_fprop: stride %d unsupported", stride);
    ConvIgemmArgs a;
    memset(&a, 0, sizeof(a));
    if (conv_taps(a.p, B, H, W, Cin, kh, kw, stride)) return -1;
    fill_epilogue(a.p, ep, Cout);
    a.x = reinterpret_cast<const __nv_bfloat16*>(x);
    a.xB = B; a.xH = H; a.xW = W; a.xC = Cin;
    a.w = reinterpret_cast<const __nv_bfloat16*>(w);
    a.w_rows = Cout; a.w_k = kh * kw * Cin;
    return launch_conv_igemm(a, reinterpret_cast<cudaStream_t>(stream));
}

int b2_conv2d_transpose_s2_fprop(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int kh,
                                 int kw, const b2_epilogue* ep, void* stream) {
    if (!x || !w || !ep || !ep->out) return set_error("b2_conv2d_transpose_s2_fprop: null pointer");
    if (kh < 2 || kw < 2 || kh * kw > kMaxTaps) return set_error("b2_conv2d_transpose_s2_fprop: bad filter %dx%d", kh, kw);
    return run_transposed(x, B, H, W, Cin, w, Cout, kh, kw, 2 * H, 2 * W, (kh - 2) / 2, (kw - 2) / 2, ep, stream, false);
}

int b2_conv2d_s2_dgrad(const void* dz, int B, int OH, int OW, int Cout, const void* w, int Cin, int kh, int kw, int H, int W,
                       const b2_epilogue* ep, void* stream) {
    if (!dz || !w || !ep || !ep->out) return set_error("b2_conv2d_s2_dgrad: null pointer");
    int oh, ow, pbh, pbw;
    tf_same(H, kh, 2, oh, pbh);
    tf_same(W, kw, 2, ow, pbw);
    if (oh != OH || ow != OW) return set_error("b2_conv2d_s2_dgrad: %dx%d is not the stride-2 SAME output of %dx%d", OH, OW, H, W);
    if (kh < 2 || kw < 2 || kh * kw > kMaxTaps) return set_error("b2_conv2d_s2_dgrad: bad filter %dx%d", kh, kw);
    return run_transposed(dz, B, OH, OW, Cout, w, Cin, kh, kw, H, W, pbh, pbw, ep, stream, false);
}

int b2_conv2d_wgrad(const void* x, int B, int H, int W, int Cx, const void* dz, int Cz, int kh, int kw, int stride, float* dw,
                    int f16, void* stream) {
    if (!x || !dz || !dw) return set_error("b2_conv2d_wgrad: null pointer");
    if (stride != 1 && stride != 2) return set_error("b2_conv2d_wgrad: stride %d unsupported", stride);
    if (kh * kw > kMaxTaps) return set_error("b2_conv2d_wgrad: %dx%d filter exceeds %d taps", kh, kw, kMaxTaps);
    WgradArgs a;
    memset(&a, 0, sizeof(a));
    int OH, OW, pbh, pbw;
    tf_same(H, kh, stride, OH, pbh);
    tf_same(W, kw, stride, OW, pbw);
    a.x = x; a.xB = B; a.xH = H; a.xW = W; a.cx = Cx;
    a.dz = dz; a.zB = B; a.zH = OH; a.zW = OW; a.cz = Cz;
    a.num_taps = kh * kw; a.in_mul = stride;
    for (int ky = 0; ky < kh; ++ky)
        for (int kx = 0; kx < kw; ++kx) {
            const int t = ky * kw + kx;
            a.dy[t] = int8_t(ky - pbh);
            a.dx[t] = int8_t(kx - pbw);
            a.wtap[t] = int16_t(t);
        }
    a.dw = dw;
    a.f16 = f16;
    return launch_conv_wgrad(a, reinterpret_cast<cudaStream_t>(stream));
}

int b2_conv2d_fprop_pad(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int kh, int kw, int stride,
                        int pad_h, int pad_w, int OH, int OW, const b2_epilogue* ep, void* stream) {
    if (!x || !w || !ep || !ep->out) return set_error("b2_conv2d_fprop_pad: null pointer");
    if (ep->stats_part) return set_error("b2_epilogue.stats_part is only honoured by b2_conv2d_fprop");
    if (stride < 1 || stride > 3) return set_error("b2_conv2d_fprop_pad: stride %d unsupported", stride);
    if (OH < 1 || OW < 1) return set_error("b2_conv2d_fprop_pad: empty output %dx%d", OH, OW);
    ConvIgemmArgs a;
    memset(&a, 0, sizeof(a));
    if (conv_taps_pad(a.p, B, Cin, kh, kw, stride, pad_h, pad_w, OH, OW)) return -1;
    fill_epilogue(a.p, ep, Cout);
    a.x = reinterpret_cast<const __nv_bfloat16*>(x);
    a.xB = B; a.xH = H; a.xW = W; a.xC = Cin;
    a.w = reinterpret_cast<const __nv_bfloat16*>(w);
    a.w_rows = Cout; a.w_k = kh * kw * Cin;
    return launch_conv_igemm(a, reinterpret_cast<cudaStream_t>(stream));
}

int b2_conv2d_strided_dgrad(const void* dz, int B, int OH, int OW, int Cout, const void* w, int Cin, int kh, int kw, int stride,
                            int pad_h, int pad_w, int H, int W, const b2_epilogue* ep, void* stream) {
    if (!dz || !w || !ep || !ep->out) return set_error("b2_conv2d_strided_dgrad: null pointer");
    if (ep->stats_part) return set_error("b2_epilogue.stats_part is only honoured by b2_conv2d_fprop");
    if (stride < 2 || stride > 3) return set_error("b2_conv2d_strided_dgrad: stride %d unsupported", stride);
    if (kh < stride || kw < stride || kh * kw > kMaxTaps)
        return set_error("b2_conv2d_strided_dgrad: filter %dx%d must cover the stride and have at most %d taps", kh, kw, kMaxTaps);
    for (int ph = 0; ph < stride * stride; ++ph) {
        ConvIgemmArgs a;
        memset(&a, 0, sizeof(a));
        strided_phase_taps(a.p, B, H, W, Cout, kh, kw, stride, pad_h, pad_w, ph / stride, ph % stride);
        if (a.p.H <= 0 || a.p.W <= 0) continue;
        fill_epilogue(a.p, ep, Cin);
        a.x = reinterpret_cast<const __nv_bfloat16*>(dz);
        a.xB = B; a.xH = OH; a.xW = OW; a.xC = Cout;
        a.w = reinterpret_cast<const __nv_bfloat16*>(w);
        a.w_rows = Cin; a.w_k = kh * kw * Cout;
        const int rc = launch_conv_igemm(a, reinterpret_cast<cudaStream_t>(stream));
        if (rc) return rc;
    }
    return 0;
}

int b2_conv2d_wgrad_pad(const void* x, int B, int H, int W, int Cx, const void* dz, int OH, int OW, int Cz, int kh, int kw,
                        int stride, int pad_h, int pad_w, float* dw, int f16, void* stream) {
    if (!x || !dz || !dw) return set_error("b2_conv2d_wgrad_pad: null pointer");
    if (stride < 1 || stride > 3) return set_error("b2_conv2d_wgrad_pad: stride %d unsupported", stride);
    if (kh * kw > kMaxTaps) return set_error("b2_conv2d_wgrad_pad: %dx%d filter exceeds %d taps", kh, kw, kMaxTaps);
    if (pad_h < 0 || pad_w < 0 || pad_h > 100 || pad_w > 100) return set_error("b2_conv2d_wgrad_pad: bad padding");
    WgradArgs a;
    memset(&a, 0, sizeof(a));
    a.x = x; a.xB = B; a.xH = H; a.xW = W; a.cx = Cx;
    a.dz = dz; a.zB = B; a.zH = OH; a.zW = OW; a.cz = Cz;
    a.num_taps = kh * kw; a.in_mul = stride;
    for (int ky = 0; ky < kh; ++ky)
        for (int kx = 0; kx < kw; ++kx) {
            const int t = ky * kw + kx;
            a.dy[t] = int8_t(ky - pad_h);
            a.dx[t] = int8_t(kx - pad_w);
            a.wtap[t] = int16_t(t);
        }
    a.dw = dw;
    a.f16 = f16;
    return launch_conv_wgrad(a, reinterpret_cast<cudaStream_t>(stream));
}

int b2_conv2d_rows_fprop(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int kh, int kw,
                         const float* bias, int act, int f16, float* out, void* stream) {
    if (!x || !w || !out) return set_error("b2_conv2d_rows_fprop: null pointer");
    return launch_conv_tail(x, B, H, W, Cin, w, Cout, kh, kw, bias, act, f16, out, reinterpret_cast<cudaStream_t>(stream));
}

int b2_conv2d_scatter_fprop(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int kh, int kw,
                            const float* bias, int act, int f16, float* out, void* stream) {
    if (!x || !w || !out) return set_error("b2_conv2d_scatter_fprop: null pointer");
    return launch_conv_scatter(x, B, H, W, Cin, w, Cout, kh, kw, bias, act, f16, out, reinterpret_cast<cudaStream_t>(stream));
}

int b2_conv2d_direct(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int kh, int kw, int stride,
                     int transposed, const b2_epilogue* ep, void* stream) {
    if (!x || !w || !ep || !ep->out) return set_error("b2_conv2d_direct: null pointer");
    if (ep->stats_part) return set_error("b2_epilogue.stats_part is only honoured by b2_conv2d_fprop");
    if (transposed)
        return run_transposed(x, B, H, W, Cin, w, Cout, kh, kw, 2 * H, 2 * W, (kh - 2) / 2, (kw - 2) / 2, ep, stream, true);
    ConvIgemmArgs a;
    memset(&a, 0, sizeof(a));
    if (conv_taps(a.p, B, H, W, Cin, kh, kw, stride)) return -1;
    fill_epilogue(a.p, ep, Cout);
    a.x = reinterpret_cast<const __nv_bfloat16*>(x);
    a.xB = B; a.xH = H; a.xW = W; a.xC = Cin;
    a.w = reinterpret_cast<const __nv_bfloat16*>(w);
    a.w_rows = Cout; a.w_k = kh * kw * Cin;
    return launch_conv_direct(a, reinterpret_cast<cudaStream_t>(stream));
}

}  // extern "C"
