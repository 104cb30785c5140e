// b2_conv_epilogue.cuh — the epilogue shared by the tcgen05 convolution kernels: TMEM -> registers, per-channel
// scale/shift (bias or folded BatchNorm), activation, residual add, 16-bit or fp32 NHWC stores.
#pragma once
#include "b2_conv_igemm.cuh"
#include "b2_ptx.cuh"

namespace b2 {

constexpr int kMaxChan = 1024;  // widest layer on the path (discriminator Dense_1, model.py:876)

// 16-bit element helpers: the storage type is bf16 or IEEE fp16, chosen per launch (IgemmParams::f16).
__device__ __forceinline__ uint32_t pack2(float a, float b, int f16) {
    if (f16) {
        // saturate to the finite half range (one F2FP.SATFINITE) so a stray large activation never becomes inf
        uint32_t r;
        asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
        return r;
    }
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
}
__device__ __forceinline__ float2 unpack2(uint32_t u, int f16) {
    if (f16) return __half22float2(*reinterpret_cast<const __half2*>(&u));
    return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u));
}
__device__ __forceinline__ float load1(const void* base, size_t i, int f16) {
    if (f16) return __half2float(reinterpret_cast<const __half*>(base)[i]);
    return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[i]);
}
__device__ __forceinline__ void store1(void* base, size_t i, float v, int f16) {
    if (f16) reinterpret_cast<__half*>(base)[i] = __float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f));
    else reinterpret_cast<__nv_bfloat16*>(base)[i] = __float2bfloat16_rn(v);
}

template <int ACT>
__device__ __forceinline__ float apply_act_t(float v, float leaky, float alpha) {
    if constexpr (ACT == ACT_RELU) return fmaxf(v, 0.f);
    else if constexpr (ACT == ACT_LEAKY) return v > 0.f ? v : v * leaky;
    else if constexpr (ACT == ACT_PRELU) return v > 0.f ? v : v * alpha;
    else if constexpr (ACT == ACT_TANH) return tanhf(v);
    else return v;
}

// Epilogue for one 128 x BN accumulator tile: this thread owns one pixel row (TMEM lane) and walks the columns in
// chunks.  Per-channel scale / shift / alpha come from shared memory (staged once per CTA, padded to the N tiling).
template <int BN, int ACT>
__device__ __forceinline__ void epilogue_tile(const IgemmParams& p, uint32_t taddr, bool valid, int n0, size_t obase,
                                              const float* s_scale, const float* s_shift, const float* s_alpha) {
    constexpr int CH = BN < 32 ? 16 : 32;
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += CH) {
        uint32_t r[32];
        if constexpr (CH == 32) {
            tmem_ld32(taddr + c0, r);
        } else {
            uint32_t r16[16];
            tmem_ld16(taddr + c0, r16);
#pragma unroll
            for (int j = 0; j < 16; ++j) r[j] = r16[j];
        }
        tmem_ld_wait();
        if (!valid || n0 + c0 >= p.n_valid) continue;
        float v[CH];
        const uint32_t sc_a = smem_u32(s_scale + n0 + c0), sh_a = smem_u32(s_shift + n0 + c0), al_a = smem_u32(s_alpha + n0 + c0);
        const bool unit_scale = p.scale == nullptr;      // bias-only layers (Conv2DTranspose): skip the scale loads
#pragma unroll
        for (int g = 0; g < CH / 4; ++g) {
            float4 sc = make_float4(1.f, 1.f, 1.f, 1.f);
            if (!unit_scale) sc = lds_f4(sc_a + 16 * g);
            const float4 sh = lds_f4(sh_a + 16 * g);
            float4 al = make_float4(0.f, 0.f, 0.f, 0.f);
            if constexpr (ACT == ACT_PRELU) al = lds_f4(al_a + 16 * g);
            v[4 * g + 0] = apply_act_t<ACT>(fmaf(__uint_as_float(r[4 * g + 0]), sc.x, sh.x), p.leaky, al.x);
            v[4 * g + 1] = apply_act_t<ACT>(fmaf(__uint_as_float(r[4 * g + 1]), sc.y, sh.y), p.leaky, al.y);
            v[4 * g + 2] = apply_act_t<ACT>(fmaf(__uint_as_float(r[4 * g + 2]), sc.z, sh.z), p.leaky, al.z);
            v[4 * g + 3] = apply_act_t<ACT>(fmaf(__uint_as_float(r[4 * g + 3]), sc.w, sh.w), p.leaky, al.w);
        }
        // 32-byte accesses: this thread's pixel row shares no 128-byte line with its neighbours', so the LSU cost is per
        // (instruction, line); 256-bit LDG/STG halve the instruction count (measured on the residual-trunk convs).
        const bool vec = (!p.out_fp32) && ((p.out_ld & 15) == 0) && (n0 + c0 + CH <= p.n_valid);
        if (vec) {
            if (p.gate) {
                const __nv_bfloat16* gp = reinterpret_cast<const __nv_bfloat16*>(p.gate) + obase + c0;
                uint8v gv[CH / 16];
#pragma unroll
                for (int g = 0; g < CH / 16; ++g) gv[g] = ldg256(gp + 16 * g);
#pragma unroll
                for (int g = 0; g < CH / 16; ++g) {
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const float2 f = unpack2(gv[g].v[e], p.f16);
                        if (!(f.x > 0.f)) v[g * 16 + 2 * e] *= p.gate_slope;
                        if (!(f.y > 0.f)) v[g * 16 + 2 * e + 1] *= p.gate_slope;
                    }
                }
            }
            if (p.residual) {
                const __nv_bfloat16* rp = p.residual + obase + c0;
                uint8v rv[CH / 16];
#pragma unroll
                for (int g = 0; g < CH / 16; ++g) rv[g] = ldg256(rp + 16 * g);
#pragma unroll
                for (int g = 0; g < CH / 16; ++g) {
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const float2 f = unpack2(rv[g].v[e], p.f16);
                        v[g * 16 + 2 * e] += f.x;
                        v[g * 16 + 2 * e + 1] += f.y;
                    }
                }
            }
            __nv_bfloat16* op = reinterpret_cast<__nv_bfloat16*>(p.out) + obase + c0;
#pragma unroll
            for (int g = 0; g < CH / 16; ++g) {
                uint8v o;
#pragma unroll
                for (int e = 0; e < 8; ++e) o.v[e] = pack2(v[g * 16 + 2 * e], v[g * 16 + 2 * e + 1], p.f16);
                stg256(op + 16 * g, o);
            }
        } else {
#pragma unroll
            for (int j = 0; j < CH; ++j) {
                const int c = n0 + c0 + j;
                if (c < p.n_valid) {
                    float o = v[j];
                    if (p.gate && !(load1(p.gate, obase + c0 + j, p.f16) > 0.f)) o *= p.gate_slope;
                    if (p.residual) o += load1(p.residual, obase + c0 + j, p.f16);
                    if (p.out_fp32) reinterpret_cast<float*>(p.out)[obase + c0 + j] = o;
                    else store1(p.out, obase + c0 + j, o, p.f16);
                }
            }
        }
    }
}

}  // namespace b2
