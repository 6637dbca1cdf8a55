// Per-row epilogue math of the fused FFT-block kernel (dx_ffn_tc.cu).  A row of 128 channels is owned by TWO threads
// (same TMEM lane, two warps): "part" 0 holds channels 0-63, part 1 channels 64-127, each as 64 fp32 values read
// from TMEM.  LayerNorm statistics are combined through shared memory as (part 0) + (part 1), each part summed in
// channel order - the same order tc_epilogue_tile<128, TC_EPI_LN> uses, so fused and unfused blocks agree bit for bit.
//   ffn_e1_store   Hs = bf16(live ? ReLU(D1 + b1) : 0)                 PositionWiseConvFF model.py:228-229 (+ H1 live padding)
//   ffn_row_add    t = acc + bias + residual, partial sum              model.py:190 / :233 (residual)
//   ffn_row_sq     partial sum of (t - mean)^2                         nn.LayerNorm model.py:191 / :234
//   ffn_row_store  out = mask(FiLM(LN(t)))  → global and/or shared     model.py:235 (FiLM), :256 / :258 (pad mask)
//   ffn_qkv_store  out = bf16(acc + bias)                              in_proj model.py:165,182
#pragma once
#include "dx_common.cuh"
#include "dx_ptx.cuh"

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

// v[64] += bias[64] + residual (8 chunks of 8 bf16 at res + kc * res_stride bytes; shared or global memory);
// returns the sum of the 64 results in channel order
__device__ __forceinline__ float ffn_row_add(float (&v)[64], const float* bias, const uint8_t* res, int64_t res_stride) {
    float sum = 0.f;
#pragma unroll
    for (int kc = 0; kc < 8; ++kc) {
        const float4 ba = reinterpret_cast<const float4*>(bias)[kc * 2], bb = reinterpret_cast<const float4*>(bias)[kc * 2 + 1];
        const float bz[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
        const uint4 pk = *reinterpret_cast<const uint4*>(res + kc * res_stride);
        const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&pk);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            float t0 = v[kc * 8 + 2 * e] + bz[2 * e];
            float t1 = v[kc * 8 + 2 * e + 1] + bz[2 * e + 1];
            t0 += __low2float(h2[e]);
            t1 += __high2float(h2[e]);
            v[kc * 8 + 2 * e] = t0;
            v[kc * 8 + 2 * e + 1] = t1;
            sum += t0;
            sum += t1;
        }
    }
    return sum;
}

__device__ __forceinline__ float ffn_row_sq(const float (&v)[64], float mean) {
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < 64; ++i) {
        const float d = __fsub_rn(v[i], mean);
        sq = __fmaf_rn(d, d, sq);
    }
    return sq;
}

// g, be: LayerNorm gamma / beta of these 64 channels (shared memory); fp: FiLM [gamma 128 | beta 128] of the row's
// utterance pre-offset to these 64 channels, or null; out_g: global &out[first chunk][row][0] or null (chunk stride
// out_chunk elements); out_s: shared-memory copy or null (chunk stride smem_chunk bytes)
__device__ __forceinline__ void ffn_row_store(const float (&v)[64], float mean, float rstd, const float* g, const float* be, const float* fp,
                                              bool live, __nv_bfloat16* out_g, int64_t out_chunk, uint8_t* out_s, uint32_t smem_chunk) {
#pragma unroll
    for (int kc = 0; kc < 8; ++kc) {
        const float4 g0 = reinterpret_cast<const float4*>(g)[kc * 2], g1 = reinterpret_cast<const float4*>(g)[kc * 2 + 1];
        const float4 e0 = reinterpret_cast<const float4*>(be)[kc * 2], e1 = reinterpret_cast<const float4*>(be)[kc * 2 + 1];
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
            float y0 = ln_film(v[kc * 8 + e], mean, rstd, gg[e], ee[e], fg[e], fb[e]);
            float y1 = ln_film(v[kc * 8 + e + 1], mean, rstd, gg[e + 1], ee[e + 1], fg[e + 1], fb[e + 1]);
            if (!live) { y0 = 0.f; y1 = 0.f; }
            o2[e / 2] = __floats2bfloat162_rn(y0, y1);
        }
        if (out_g != nullptr) *reinterpret_cast<uint4*>(out_g + kc * out_chunk) = *reinterpret_cast<const uint4*>(o2);
        if (out_s != nullptr) *reinterpret_cast<uint4*>(out_s + kc * smem_chunk) = *reinterpret_cast<const uint4*>(o2);
    }
}

// 64 columns of one 128-column tile of the next block's QKV projection
__device__ __forceinline__ void ffn_qkv_store(const float (&v)[64], const float* bias, bool in_utt, __nv_bfloat16* out_row, int64_t out_chunk) {
#pragma unroll
    for (int kc = 0; kc < 8; ++kc) {
        const float4 b0 = reinterpret_cast<const float4*>(bias)[kc * 2], b1 = reinterpret_cast<const float4*>(bias)[kc * 2 + 1];
        const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
        __align__(16) __nv_bfloat162 o2[4];
#pragma unroll
        for (int e = 0; e < 8; e += 2) {
            float t0 = v[kc * 8 + e] + bb[e], t1 = v[kc * 8 + e + 1] + bb[e + 1];
            if (!in_utt) { t0 = 0.f; t1 = 0.f; }
            o2[e / 2] = __floats2bfloat162_rn(t0, t1);
        }
        *reinterpret_cast<uint4*>(out_row + kc * out_chunk) = *reinterpret_cast<const uint4*>(o2);
    }
}


// ---------------------------------------------------------------------------------------------------------------------
// Rolled store epilogues.  Code that runs once per tile is fetched from L2 every time (ncu: stall_no_inst was the top stall
// reason of the row epilogues; clock64 traces showed 9-19 k cycles for ~800 instructions of work per warp), so the P3 store
// and the partial-sum dump visit the thread's 64 accumulator columns in TMEM in four groups of 16 with a short loop.  The
// LayerNorm epilogues stay unrolled (their global loads must be in flight together: a rolled three-pass version that
// kept the row in TMEM measured 10.8 k cycles against 8.4 k) but exist ONCE in the kernel, shared by P0 and E2.
// ---------------------------------------------------------------------------------------------------------------------
// ---- rolled LayerNorm epilogue: the thread's 64 accumulator columns stay in TMEM and are visited in four groups of 16:
// pass A adds bias + residual (+ the partial sums of a split tile), accumulates the row sum and writes the values back
// (tcgen05.st); pass B the squared deviations; pass C normalises, applies FiLM and the mask and stores.  Per-element
// operations and their order are those of ffn_row_add / ffn_row_sq / ffn_row_store, so the results are the same bits.
struct FfnRowEpi {
    uint32_t taddr;              // TMEM address of this thread's 64 accumulator columns
    const float* bias;           // shared memory, this part's 64 values
    const uint8_t* res;          // residual chunk 0 of this part at this row (8 bf16, shared memory), chunk kc at + kc * res_stride
    uint32_t res_stride;
    const float *g, *be;         // LayerNorm gamma / beta of these 64 channels (shared memory)
    const float* film;           // FiLM [gamma 128 | beta 128] of the row's utterance pre-offset to these channels, or null
    bool live;
    __nv_bfloat16* out_g;        // global &out[first chunk][row][0] or null
    int64_t out_chunk;
    uint8_t* out_s;              // shared-memory copy or null
    uint32_t smem_chunk;
    const float* part_base;      // owner of a split tile: partial-sum slot 0 + this thread's offset
    int64_t part_slot;           // floats per slot
    int k_hi, n_contrib;         // slots k_hi, k_hi - 1, ... (n_contrib of them) are added, in this order
};

__device__ __forceinline__ void ffn_ln_rolled(const FfnRowEpi& E, float* my_red, const float* peer_red, int pair_bar, int part) {
    float sum = 0.f;
#pragma unroll 1
    for (int q4 = 0; q4 < 4; ++q4) {
        float t[16];
        ptx::tmem_ld16(E.taddr + 16 * q4, t);
        ptx::tmem_ld_wait();
        for (int q = 0; q < E.n_contrib; ++q) {
            const float* src = E.part_base + (int64_t)(E.k_hi - q) * E.part_slot + (16 * q4) * 128;
            float u[16];
#pragma unroll
            for (int e = 0; e < 16; ++e) u[e] = __ldcg(src + e * 128);
#pragma unroll
            for (int e = 0; e < 16; ++e) t[e] += u[e];
        }
#pragma unroll
        for (int c2 = 0; c2 < 2; ++c2) {
            const int kc = 2 * q4 + c2;
            const float4 ba = reinterpret_cast<const float4*>(E.bias)[kc * 2], bb = reinterpret_cast<const float4*>(E.bias)[kc * 2 + 1];
            const float bz[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
            const uint4 pk = *reinterpret_cast<const uint4*>(E.res + kc * E.res_stride);
            const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&pk);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                float t0 = t[c2 * 8 + 2 * e] + bz[2 * e];
                float t1 = t[c2 * 8 + 2 * e + 1] + bz[2 * e + 1];
                t0 += __low2float(h2[e]);
                t1 += __high2float(h2[e]);
                t[c2 * 8 + 2 * e] = t0;
                t[c2 * 8 + 2 * e + 1] = t1;
                sum += t0;
                sum += t1;
            }
        }
        ptx::tmem_st16(E.taddr + 16 * q4, t);
    }
    ptx::tmem_st_wait();
    my_red[0] = sum;
    ptx::named_bar_sync(pair_bar, 64);
    const float o_ = peer_red[0];
    const float mean = (part == 0 ? sum + o_ : o_ + sum) * (1.f / 128.f);
    float sq = 0.f;
#pragma unroll 1
    for (int q4 = 0; q4 < 4; ++q4) {
        float t[16];
        ptx::tmem_ld16(E.taddr + 16 * q4, t);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 16; ++e) {
            const float d = __fsub_rn(t[e], mean);
            sq = __fmaf_rn(d, d, sq);
        }
    }
    my_red[512] = sq;
    ptx::named_bar_sync(pair_bar, 64);
    const float p_ = peer_red[512];
    const float rstd = rsqrtf((part == 0 ? sq + p_ : p_ + sq) * (1.f / 128.f) + 1e-5f);
#pragma unroll 1
    for (int q4 = 0; q4 < 4; ++q4) {
        float t[16];
        ptx::tmem_ld16(E.taddr + 16 * q4, t);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int c2 = 0; c2 < 2; ++c2) {
            const int kc = 2 * q4 + c2;
            const float4 g0 = reinterpret_cast<const float4*>(E.g)[kc * 2], g1 = reinterpret_cast<const float4*>(E.g)[kc * 2 + 1];
            const float4 e0 = reinterpret_cast<const float4*>(E.be)[kc * 2], e1 = reinterpret_cast<const float4*>(E.be)[kc * 2 + 1];
            const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
            const float ee[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
            float fg[8] = {1.f, 1.f, 1.f, 1.f, 1.f, 1.f, 1.f, 1.f}, fb[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (E.film) {
                const float4 a0 = reinterpret_cast<const float4*>(E.film)[kc * 2], a1 = reinterpret_cast<const float4*>(E.film)[kc * 2 + 1];
                const float4 c0 = reinterpret_cast<const float4*>(E.film + 128)[kc * 2], c1 = reinterpret_cast<const float4*>(E.film + 128)[kc * 2 + 1];
                fg[0] = a0.x; fg[1] = a0.y; fg[2] = a0.z; fg[3] = a0.w; fg[4] = a1.x; fg[5] = a1.y; fg[6] = a1.z; fg[7] = a1.w;
                fb[0] = c0.x; fb[1] = c0.y; fb[2] = c0.z; fb[3] = c0.w; fb[4] = c1.x; fb[5] = c1.y; fb[6] = c1.z; fb[7] = c1.w;
            }
            __align__(16) __nv_bfloat162 o2[4];
#pragma unroll
            for (int e = 0; e < 8; e += 2) {
                float y0 = ln_film(t[c2 * 8 + e], mean, rstd, gg[e], ee[e], fg[e], fb[e]);
                float y1 = ln_film(t[c2 * 8 + e + 1], mean, rstd, gg[e + 1], ee[e + 1], fg[e + 1], fb[e + 1]);
                if (!E.live) { y0 = 0.f; y1 = 0.f; }
                o2[e / 2] = __floats2bfloat162_rn(y0, y1);
            }
            if (E.out_g != nullptr) *reinterpret_cast<uint4*>(E.out_g + kc * E.out_chunk) = *reinterpret_cast<const uint4*>(o2);
            if (E.out_s != nullptr) *reinterpret_cast<uint4*>(E.out_s + kc * E.smem_chunk) = *reinterpret_cast<const uint4*>(o2);
        }
    }
}

// 64 columns of one 128-column tile of the next block's QKV projection, straight from TMEM (in_proj model.py:165,182)
__device__ __forceinline__ void ffn_qkv_rolled(uint32_t taddr, const float* bias, bool in_utt, __nv_bfloat16* out_row, int64_t out_chunk,
                                               bool do_store) {
#pragma unroll 1
    for (int q4 = 0; q4 < 4; ++q4) {
        float t[16];
        ptx::tmem_ld16(taddr + 16 * q4, t);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int c2 = 0; c2 < 2; ++c2) {
            const int kc = 2 * q4 + c2;
            const float4 b0 = reinterpret_cast<const float4*>(bias)[kc * 2], b1 = reinterpret_cast<const float4*>(bias)[kc * 2 + 1];
            const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
            __align__(16) __nv_bfloat162 o2[4];
#pragma unroll
            for (int e = 0; e < 8; e += 2) {
                float t0 = t[c2 * 8 + e] + bb[e], t1 = t[c2 * 8 + e + 1] + bb[e + 1];
                if (!in_utt) { t0 = 0.f; t1 = 0.f; }
                o2[e / 2] = __floats2bfloat162_rn(t0, t1);
            }
            if (do_store) *reinterpret_cast<uint4*>(out_row + kc * out_chunk) = *reinterpret_cast<const uint4*>(o2);
        }
    }
}

// this thread's 64 columns of a segment's partial D2 sum → its slot in global memory ([col][128 rows]: coalesced)
__device__ __forceinline__ void ffn_dump_rolled(uint32_t taddr, float* dst) {
#pragma unroll 1
    for (int q4 = 0; q4 < 4; ++q4) {
        float t[16];
        ptx::tmem_ld16(taddr + 16 * q4, t);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 16; ++e) dst[(16 * q4 + e) * 128] = t[e];
    }
}

}  // namespace dx
