// Per-row epilogue math of the fused FFN kernels (thread ↔ row; values come from TMEM in registers).
//   ffn_e1_store: Hs = bf16(live ? ReLU(D1 + b1) : 0)                      PositionWiseConvFF model.py:228-229 (+ H1 live padding)
//   ffn_e2_store: out = mask(FiLM(LN(D2 + b2 + x)))                        model.py:230-235, FFTBlock pad mask model.py:258
#pragma once
#include "dx_common.cuh"

namespace dx {

// 64 hidden channels of one row → 8 chunks of the K-major A operand of G2 (chunk stride `chunk_stride` bytes)
__device__ __forceinline__ void ffn_e1_store(const float (&v)[64], const float* bias, bool hid_live, uint8_t* dst, uint32_t chunk_stride) {
#pragma unroll
    for (int kc = 0; kc < 8; ++kc) {
        const float4 ba = reinterpret_cast<const float4*>(bias)[kc * 2];
        const float4 bb = reinterpret_cast<const float4*>(bias)[kc * 2 + 1];
        const float bz[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
        __align__(16) __nv_bfloat162 o2[4];
#pragma unroll
        for (int e = 0; e < 8; e += 2) {
            float t0 = fmaxf(v[kc * 8 + e] + bz[e], 0.f);
            float t1 = fmaxf(v[kc * 8 + e + 1] + bz[e + 1], 0.f);
            if (!hid_live) { t0 = 0.f; t1 = 0.f; }
            o2[e / 2] = __floats2bfloat162_rn(t0, t1);
        }
        *reinterpret_cast<uint4*>(dst + kc * chunk_stride) = *reinterpret_cast<const uint4*>(o2);
    }
}

// v: the row's 128 conv2 outputs (modified in place); xres: the row's residual in the X tile (chunk stride 4096 B);
// fp: the utterance's FiLM vector [gamma 128 | beta 128] or null; out_row: &out[chunk 0][row][0], chunk stride out_chunk elements
__device__ __forceinline__ void ffn_e2_store(float (&v)[128], const uint8_t* xres, const float* s_b2, const float* s_g, const float* s_be,
                                             const float* fp, bool out_live, __nv_bfloat16* out_row, int64_t out_chunk) {
    float sum = 0.f;
#pragma unroll
    for (int kc = 0; kc < 16; ++kc) {
        const float4 ba = reinterpret_cast<const float4*>(s_b2)[kc * 2], bb = reinterpret_cast<const float4*>(s_b2)[kc * 2 + 1];
        const float bz[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
        const uint4 pk = *reinterpret_cast<const uint4*>(xres + kc * 4096);
        const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&pk);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float t0 = v[kc * 8 + 2 * e] + bz[2 * e] + __low2float(h2[e]);
            const float t1 = v[kc * 8 + 2 * e + 1] + bz[2 * e + 1] + __high2float(h2[e]);
            v[kc * 8 + 2 * e] = t0;
            v[kc * 8 + 2 * e + 1] = t1;
            sum += t0 + t1;
        }
    }
    const float mean = sum * (1.f / 128.f);
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < 128; ++i) {
        const float d = v[i] - mean;
        sq = fmaf(d, d, sq);
    }
    const float rstd = rsqrtf(sq * (1.f / 128.f) + 1e-5f);
#pragma unroll
    for (int kc = 0; kc < 16; ++kc) {
        const float4 g0 = reinterpret_cast<const float4*>(s_g)[kc * 2], g1 = reinterpret_cast<const float4*>(s_g)[kc * 2 + 1];
        const float4 e0 = reinterpret_cast<const float4*>(s_be)[kc * 2], e1 = reinterpret_cast<const float4*>(s_be)[kc * 2 + 1];
        const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
        const float ee[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
        float fg[8] = {1.f, 1.f, 1.f, 1.f, 1.f, 1.f, 1.f, 1.f}, fb[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (fp) {
            const float4 a0 = reinterpret_cast<const float4*>(fp)[kc * 2], a1 = reinterpret_cast<const float4*>(fp)[kc * 2 + 1];
            const float4 c0 = reinterpret_cast<const float4*>(fp + 128)[kc * 2], c1 = reinterpret_cast<const float4*>(fp + 128)[kc * 2 + 1];
            fg[0] = a0.x; fg[1] = a0.y; fg[2] = a0.z; fg[3] = a0.w; fg[4] = a1.x; fg[5] = a1.y; fg[6] = a1.z; fg[7] = a1.w;
            fb[0] = c0.x; fb[1] = c0.y; fb[2] = c0.z; fb[3] = c0.w; fb[4] = c1.x; fb[5] = c1.y; fb[6] = c1.z; fb[7] = c1.w;
        }
        __align__(16) __nv_bfloat162 o2[4];
#pragma unroll
        for (int e = 0; e < 8; e += 2) {
            float y0 = fg[e] * ((v[kc * 8 + e] - mean) * rstd * gg[e] + ee[e]) + fb[e];
            float y1 = fg[e + 1] * ((v[kc * 8 + e + 1] - mean) * rstd * gg[e + 1] + ee[e + 1]) + fb[e + 1];
            if (!out_live) { y0 = 0.f; y1 = 0.f; }
            o2[e / 2] = __floats2bfloat162_rn(y0, y1);
        }
        *reinterpret_cast<uint4*>(out_row + kc * out_chunk) = *reinterpret_cast<const uint4*>(o2);
    }
}

}  // namespace dx
