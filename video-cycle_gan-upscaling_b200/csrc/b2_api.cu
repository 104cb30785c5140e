// b2_api.cu — the extern "C" boundary (include/b2srgan.h): turns layer-level geometry (kernel size, stride,
// TensorFlow "same" padding, Conv2DTranspose phases) into tap lists for the implicit-GEMM kernels and launches them.
#include <string.h>

#include "../../include/b2srgan.h"
#include "b2_host.h"

using namespace b2;

static void fill_epilogue(IgemmParams& p, const b2_epilogue* ep, int Cout) {
    p.scale = ep->scale;
    p.shift = ep->shift;
    p.alpha = ep->alpha;
    p.act = ep->act;
    p.leaky = ep->leaky;
    p.residual = reinterpret_cast<const __nv_bfloat16*>(ep->residual);
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

// Conv2DTranspose(stride 2, same) is the input-gradient of a SAME stride-2 conv: out[y] = sum_i x[i] * W[y - 2i + pb],
// pb = (k-2)/2.  For output parity py: y = 2j + py, i = j + d  =>  ky = py + pb - 2d.
static void convT_phase_taps(IgemmParams& p, int B, int H, int W, int Cin, int kh, int kw, int py, int px) {
    const int pbh = (kh - 2) / 2, pbw = (kw - 2) / 2;
    p.B = B; p.H = H; p.W = W;
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
    p.OH = 2 * H; p.OW = 2 * W;
    p.out_sy = 2; p.out_sx = 2; p.out_oy = py; p.out_ox = px;
}

extern "C" {

const char* b2_last_error(void) { return last_error(); }
int b2_version(void) { return 1; }

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
    for (int ph = 0; ph < 4; ++ph) {
        ConvIgemmArgs a;
        memset(&a, 0, sizeof(a));
        convT_phase_taps(a.p, B, H, W, Cin, kh, kw, ph >> 1, ph & 1);
        fill_epilogue(a.p, ep, Cout);
        a.x = reinterpret_cast<const __nv_bfloat16*>(x);
        a.xB = B; a.xH = H; a.xW = W; a.xC = Cin;
        a.w = reinterpret_cast<const __nv_bfloat16*>(w);
        a.w_rows = Cout; a.w_k = kh * kw * Cin;
        if (a.p.num_taps == 0) return set_error("b2_conv2d_transpose_s2_fprop: phase %d has no taps", ph);
        int rc = launch_conv_igemm(a, reinterpret_cast<cudaStream_t>(stream));
        if (rc) return rc;
    }
    return 0;
}

int b2_conv2d_rows_fprop(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int kh, int kw,
                         const float* bias, int act, int f16, float* out, void* stream) {
    if (!x || !w || !out) return set_error("b2_conv2d_rows_fprop: null pointer");
    return launch_conv_tail(x, B, H, W, Cin, w, Cout, kh, kw, bias, act, f16, out, reinterpret_cast<cudaStream_t>(stream));
}

int b2_conv2d_direct(const void* x, int B, int H, int W, int Cin, const void* w, int Cout, int kh, int kw, int stride,
                     int transposed, const b2_epilogue* ep, void* stream) {
    if (!x || !w || !ep || !ep->out) return set_error("b2_conv2d_direct: null pointer");
    for (int ph = 0; ph < (transposed ? 4 : 1); ++ph) {
        ConvIgemmArgs a;
        memset(&a, 0, sizeof(a));
        if (transposed) convT_phase_taps(a.p, B, H, W, Cin, kh, kw, ph >> 1, ph & 1);
        else if (conv_taps(a.p, B, H, W, Cin, kh, kw, stride)) return -1;
        fill_epilogue(a.p, ep, Cout);
        a.x = reinterpret_cast<const __nv_bfloat16*>(x);
        a.xB = B; a.xH = H; a.xW = W; a.xC = Cin;
        a.w = reinterpret_cast<const __nv_bfloat16*>(w);
        a.w_rows = Cout; a.w_k = kh * kw * Cin;
        int rc = launch_conv_direct(a, reinterpret_cast<cudaStream_t>(stream));
        if (rc) return rc;
    }
    return 0;
}

}  // extern "C"
