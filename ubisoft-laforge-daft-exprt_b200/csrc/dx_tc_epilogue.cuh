// Epilogues shared by the tcgen05 GEMM kernels (dx_tc_gemm.cu, dx_tc_linear.cu): one thread owns one output row
// (TMEM lane) of a 128 x NT fp32 accumulator tile.
//   TC_EPI_F32   out fp32 [rows][N] = acc + bias
//   TC_EPI_BF16  out bf16 interleaved = act(acc + bias), liveness-masked (+ per-tile LayerNorm partial sums)
//   TC_EPI_LN    out bf16 interleaved = mask(film(LN(act(acc + bias) + res) * g + b))   (NT == N == 128)
// `tempty` is arrived on (one arrival per warp) as soon as the accumulator has been read out of TMEM.
#pragma once
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_ptx.cuh"

namespace dx {

template <int NT, int EPI>
__device__ __forceinline__ void tc_epilogue_tile(const TcGemmArgs& g, const Rows& R, int64_t total, int64_t r, int n_tile,
                                                 uint32_t taddr, uint64_t* tempty, int lane, bool live, int b_idx,
                                                 const float* s_bias, const float* s_lng, const float* s_lnb) {
    const int n0 = n_tile * NT;
    if constexpr (EPI == TC_EPI_LN) {
        static_assert(EPI != TC_EPI_LN || NT == 128, "LN epilogue needs the whole 128-wide row in one tile");
        float v[128];
#pragma unroll
        for (int q = 0; q < 4; ++q) ptx::tmem_ld32(taddr + q * 32, v + q * 32);
        ptx::tmem_ld_wait();
        // accumulator is in registers: release TMEM early
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0 && tempty != nullptr) ptx::mbar_arrive(tempty);
        if (r < total) {
            // statistics are summed per 64-channel half and combined as (low + high): the order the fused block
            // kernel uses (two threads per row), so both paths give identical bits
            float sum_lo = 0.f, sum_hi = 0.f;
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                float4 b0 = reinterpret_cast<const float4*>(s_bias)[c * 2];
                float4 b1 = reinterpret_cast<const float4*>(s_bias)[c * 2 + 1];
                const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                float rr[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                if (g.res != nullptr && live) {
                    uint4 pk = *reinterpret_cast<const uint4*>(g.res + ((int64_t)c * g.res_rows_phys + DX_HALO + r) * 8);
                    const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&pk);
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        rr[2 * e] = __low2float(h2[e]);
                        rr[2 * e + 1] = __high2float(h2[e]);
                    }
                }
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    float t = v[c * 8 + e] + bb[e];
                    if (g.relu) t = fmaxf(t, 0.f);
                    t += rr[e];
                    v[c * 8 + e] = t;
                    if (c < 8) sum_lo += t;
                    else sum_hi += t;
                }
            }
            const float mean = (sum_lo + sum_hi) * (1.f / 128.f);
            float sq_lo = 0.f, sq_hi = 0.f;
#pragma unroll
            for (int i = 0; i < 64; ++i) {
                const float d0 = __fsub_rn(v[i], mean), d1 = __fsub_rn(v[64 + i], mean);
                sq_lo = __fmaf_rn(d0, d0, sq_lo);
                sq_hi = __fmaf_rn(d1, d1, sq_hi);
            }
            const float rstd = rsqrtf((sq_lo + sq_hi) * (1.f / 128.f) + 1e-5f);
            const float* fp = g.film ? g.film + (int64_t)b_idx * g.film_stride : nullptr;
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                __align__(16) __nv_bfloat162 o2[4];
                const float4 g0 = reinterpret_cast<const float4*>(s_lng)[c * 2], g1 = reinterpret_cast<const float4*>(s_lng)[c * 2 + 1];
                const float4 e0 = reinterpret_cast<const float4*>(s_lnb)[c * 2], e1 = reinterpret_cast<const float4*>(s_lnb)[c * 2 + 1];
                const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
                const float ee[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
                float fg[8] = {1.f, 1.f, 1.f, 1.f, 1.f, 1.f, 1.f, 1.f}, fb[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                if (fp) {
                    const float4 a0 = reinterpret_cast<const float4*>(fp)[c * 2], a1 = reinterpret_cast<const float4*>(fp)[c * 2 + 1];
                    const float4 c0 = reinterpret_cast<const float4*>(fp + 128)[c * 2], c1 = reinterpret_cast<const float4*>(fp + 128)[c * 2 + 1];
                    fg[0] = a0.x; fg[1] = a0.y; fg[2] = a0.z; fg[3] = a0.w; fg[4] = a1.x; fg[5] = a1.y; fg[6] = a1.z; fg[7] = a1.w;
                    fb[0] = c0.x; fb[1] = c0.y; fb[2] = c0.z; fb[3] = c0.w; fb[4] = c1.x; fb[5] = c1.y; fb[6] = c1.z; fb[7] = c1.w;
                }
#pragma unroll
                for (int e = 0; e < 8; e += 2) {
                    const int i = c * 8 + e;
                    float y0 = ln_film(v[i], mean, rstd, gg[e], ee[e], fg[e], fb[e]);
                    float y1 = ln_film(v[i + 1], mean, rstd, gg[e + 1], ee[e + 1], fg[e + 1], fb[e + 1]);
                    if (!live) { y0 = 0.f; y1 = 0.f; }
                    o2[e / 2] = __floats2bfloat162_rn(y0, y1);
                }
                *reinterpret_cast<uint4*>(g.out_bf + ((int64_t)c * g.out_rows_phys + DX_HALO + r) * 8) =
                    *reinterpret_cast<const uint4*>(o2);
            }
        }
    } else {
        float ssum = 0.f, ssq = 0.f;
#pragma unroll 1
        for (int q = 0; q < NT / 32; ++q) {
            float v[32];
            ptx::tmem_ld32(taddr + q * 32, v);
            ptx::tmem_ld_wait();
            const int nb = n0 + q * 32;
            if (nb < g.N && r < total) {
#pragma unroll
                for (int e = 0; e < 32; ++e) {
                    float t = v[e] + s_bias[nb + e];
                    if (g.relu) t = fmaxf(t, 0.f);
                    if (!live) t = 0.f;
                    v[e] = t;
                    ssum += t;
                    ssq = fmaf(t, t, ssq);
                }
                if constexpr (EPI == TC_EPI_F32) {
                    float* orow = g.out_f32 + r * g.N + nb;
                    if (g.relu_mask != nullptr && nb + 32 <= g.N) {
                        const float4* m4 = reinterpret_cast<const float4*>(g.relu_mask + r * g.N + nb);
#pragma unroll
                        for (int e = 0; e < 32; e += 4) {
                            const float4 m = m4[e / 4];
                            v[e] = m.x > 0.f ? v[e] : 0.f; v[e + 1] = m.y > 0.f ? v[e + 1] : 0.f;
                            v[e + 2] = m.z > 0.f ? v[e + 2] : 0.f; v[e + 3] = m.w > 0.f ? v[e + 3] : 0.f;
                        }
                    }
                    if (nb + 32 <= g.N) {
#pragma unroll
                        for (int e = 0; e < 32; e += 4)
                            *reinterpret_cast<float4*>(orow + e) = make_float4(v[e], v[e + 1], v[e + 2], v[e + 3]);
                    } else {
                        for (int e = 0; e < 32; ++e)
                            if (nb + e < g.N) orow[e] = v[e];
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        __align__(16) __nv_bfloat162 o2[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) o2[e] = __floats2bfloat162_rn(v[c * 8 + 2 * e], v[c * 8 + 2 * e + 1]);
                        *reinterpret_cast<uint4*>(g.out_bf + ((int64_t)(nb / 8 + c) * g.out_rows_phys + DX_HALO + r) * 8) =
                            *reinterpret_cast<const uint4*>(o2);
                    }
                }
            }
        }
        // per-(row, N tile) partial sums, written without atomics: the consumer adds the tiles in a fixed
        // order, so the LayerNorm statistics (and everything downstream) are run-to-run deterministic
        if (g.row_stats != nullptr && r < total)
            *reinterpret_cast<float2*>(g.row_stats + (r * TC_STAT_TILES + n_tile) * 2) = make_float2(ssum, ssq);
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0 && tempty != nullptr) ptx::mbar_arrive(tempty);
    }
}

}  // namespace dx
