// Masked multi-head self-attention block on tcgen05 (bf16 operands, fp32 scores/softmax/accumulation), fused with
// the output projection, the residual, the post-attention LayerNorm and the pad mask.
//
// Reference: MultiHeadAttention.forward (model.py:171-193) = nn.MultiheadAttention slow path (:182-186):
//   S_h = (Q_h · d^-0.5) K_h^T, -inf at padded keys, softmax over keys, O_h = A_h V_h, O = [O_0|O_1] W_out^T + b_out,
//   then LayerNorm(O + x) (:191) and FFTBlock's masked_fill (:256).  The head-averaged weights torch also returns
//   are discarded by FFTBlock (model.py:255) and never materialised here.
//
// One CTA = (utterance, 128-query tile), both heads one after the other.  K/V are streamed in 64-key tiles.
//   S  = Q K^T      : tcgen05.mma 128x64x16 x4, A = Q_h K-major, B = K tile K-major → TMEM cols 0..63
//   P  = softmax    : 4 warps, thread ↔ query row; the exponent reference m_ref is only raised when the row maximum
//                     exceeds it by more than 8 (log2 units), so P <= 256 and the rescale below is rare; P written to
//                     smem as the bf16 K-major A operand
//   O_h += P V      : tcgen05.mma 128x64x16 x4, B = V tile MN-major (the channel-interleaved global layout IS that
//                     canonical layout) → accumulates in TMEM cols 64(1+h) .. +63 across all KV tiles
//   rescale         : when a warp raises m_ref it multiplies its 32 O rows in TMEM (tcgen05.ld / st) before P is
//                     handed to the MMA; s_full(j) already implies that P V of tile j-1 has completed.
//   Y  = LN(O W + b + x) : O_h / l_h → bf16 → smem (over the Q tiles) as the A operand, W_out (32 KB) bulk-copied over
//                     the K/V ring, 8 MMAs 128x128x16 → TMEM cols 0..127, LayerNorm epilogue (dx_tc_epilogue.cuh).
// FUSED = false (unit-test surface dx_test_attention): O is written to global instead of the projection.
// 2 CTAs per SM (80 KB smem, 256 TMEM columns): the S → softmax → P·V chain of one CTA is latency-bound
// (profiles/r01c_attention_tc_kernel_full.txt) and is hidden by the other.
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_ptx.cuh"
#include "dx_tc_epilogue.cuh"
#include <string.h>

namespace dx {

constexpr int ATC_BQ = 128, ATC_BK = 64;
constexpr int ATC_THREADS = 192;
constexpr uint32_t ATC_QH_BYTES = 8 * 128 * 16;        // 16 KB per head: [8 dh chunk][128 row][8]; FUSED: both heads, later the O tile
constexpr uint32_t ATC_K_BYTES = 8 * ATC_BK * 16;      // 8 KB:  [8 dh chunk][64 key][8]
constexpr uint32_t ATC_P_BYTES = 8 * 128 * 16;         // 16 KB: [8 key chunk][128 row][8]
constexpr uint32_t ATC_PAR_BYTES = 3 * 128 * 4;        // out-proj bias, LN gamma, LN beta
template <bool FUSED> struct AtcCfg {
    static constexpr uint32_t Q_BYTES = (FUSED ? 2 : 1) * ATC_QH_BYTES;
    static constexpr uint32_t TMEM_COLS = FUSED ? 256 : 128;       // S | O_h0 [| O_h1 | -]
    static constexpr uint32_t SMEM = Q_BYTES + 4 * ATC_K_BYTES + ATC_P_BYTES + ATC_PAR_BYTES + 12 * 8 + 16;
    static constexpr int CTAS_PER_SM = FUSED ? 2 : 3;
};
constexpr float ATC_LAZY = 8.f;                        // raise the exponent reference only beyond this margin

struct AtcParams {
    const __nv_bfloat16* qkv;   // [48][rows_phys][8]
    __nv_bfloat16* out;         // FUSED = false: attention output [16][rows_phys][8]
    int64_t rows_phys;
    Rows R;
    uint32_t v_lbo, v_sbo;      // MN-major descriptor strides of the V operand (debug knob 1 swaps them)
    TcGemmArgs g;               // FUSED: out-proj weights (packed nt 128, ks 64), bias, LN, residual, output
    long long* trace;           // bring-up: clock64 timeline of CTA (0,0,0) (dx_set_trace_buffer) or null
};
// timeline slots: [128] kernel start, [129] Q landed, [130] end, [131] out-proj issued; per KV step g < 8
// (base 132 + 10 g): issuer +0 kv_full ok, +1 S issued, +2 p_full ok, +3 PV issued;  softmax +4 s_full ok,
// +5 max done, +6 P stored
#define AT_TRACE(slot) do { if (tr) tr[(slot)] = clock64(); } while (0)

// FUSED: grid (q tiles, 1, B), both heads per CTA.  Otherwise grid (q tiles, 2, B), one head per CTA, 3 CTAs per SM.
template <bool FUSED>
__global__ void __launch_bounds__(ATC_THREADS, AtcCfg<FUSED>::CTAS_PER_SM) attention_tc_kernel(const AtcParams p) {
    constexpr uint32_t ATC_Q_BYTES = AtcCfg<FUSED>::Q_BYTES;
    constexpr uint32_t ATC_TMEM_COLS = AtcCfg<FUSED>::TMEM_COLS;
    extern __shared__ __align__(1024) uint8_t at_smem[];
    ptx::griddep_wait();      // PDL: the lengths and QKV come from earlier kernels
    ptx::griddep_launch();
    const int b = blockIdx.z;
    const int len = (int)min((long long)p.R.len[b], (long long)p.R.n_max);
    const int q0 = blockIdx.x * ATC_BQ;
    if (q0 >= len) return;
    const int64_t rs = p.R.row_start[b];
    const int n_kv = (len + ATC_BK - 1) / ATC_BK;
    const int h0 = FUSED ? 0 : (int)blockIdx.y;
    const int n_steps = FUSED ? 2 * n_kv : n_kv;        // FUSED: head 0 tiles, then head 1 tiles

    uint8_t* sQ = at_smem;
    uint8_t* sKV = sQ + ATC_Q_BYTES;                    // 2 stages x (K | V); later W_out (32 KB)
    uint8_t* sP = sKV + 4 * ATC_K_BYTES;
    float* s_bias = reinterpret_cast<float*>(sP + ATC_P_BYTES);
    float* s_lng = s_bias + 128;
    float* s_lnb = s_lng + 128;
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_lnb + 128);
    uint64_t* q_full = bars;            // 1
    uint64_t* kv_full = bars + 1;       // 2
    uint64_t* kv_empty = bars + 3;      // 2
    uint64_t* s_full = bars + 5;
    uint64_t* p_full = bars + 6;
    uint64_t* o_done = bars + 7;
    uint64_t* w_full = bars + 8;
    uint64_t* oo_full = bars + 9;       // O tile (bf16) in smem
    uint64_t* d_full = bars + 10;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 12);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    long long* tr = (p.trace != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0 && (warp == 1 || warp == 2)) ? p.trace : nullptr;
    if (warp == 1) AT_TRACE(128);
    if (FUSED && threadIdx.x < 128) {
        s_bias[threadIdx.x] = p.g.bias[threadIdx.x];
        s_lng[threadIdx.x] = p.g.ln_g[threadIdx.x];
        s_lnb[threadIdx.x] = p.g.ln_b[threadIdx.x];
    }
    if (threadIdx.x == 0) {
        ptx::mbar_init(q_full, 1);
        for (int i = 0; i < 2; ++i) { ptx::mbar_init(&kv_full[i], 1); ptx::mbar_init(&kv_empty[i], 1); }
        ptx::mbar_init(s_full, 1);
        ptx::mbar_init(p_full, 4);
        ptx::mbar_init(o_done, 1);
        ptx::mbar_init(w_full, 1);
        ptx::mbar_init(oo_full, 4);
        ptx::mbar_init(d_full, 1);
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<ATC_TMEM_COLS>(tmem_ptr);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    const uint32_t tm_s = tmem_base, tm_o = tmem_base + 64;     // O_h at tm_o + 64 h; out-proj D at cols 0..127

    if (warp == 0) {
        if (ptx::elect_one()) {
            ptx::mbar_expect_tx(q_full, ATC_Q_BYTES);
            for (int c = 0; c < (FUSED ? 16 : 8); ++c) {      // Q occupies qkv chunks 0..15: 0-7 = head 0, 8-15 = head 1
                const __nv_bfloat16* src = p.qkv + ((int64_t)(h0 * 8 + c) * p.rows_phys + DX_HALO + rs + q0) * 8;
                ptx::bulk_g2s(sQ + c * 2048, src, 2048, q_full);
            }
            for (int g = 0; g < n_steps; ++g) {
                const int hs = (FUSED && g >= n_kv) ? 1 : 0, j = g - hs * n_kv, h = h0 + hs;
                const int st = g & 1;
                ptx::mbar_wait(&kv_empty[st], ((g >> 1) & 1) ^ 1);
                uint8_t* sk = sKV + (size_t)st * 2 * ATC_K_BYTES;
                uint8_t* sv = sk + ATC_K_BYTES;
                ptx::mbar_expect_tx(&kv_full[st], 2 * ATC_K_BYTES);
                const int64_t krow = DX_HALO + rs + (int64_t)j * ATC_BK;
                for (int c = 0; c < 8; ++c) {
                    ptx::bulk_g2s(sk + c * 1024, p.qkv + ((int64_t)(16 + h * 8 + c) * p.rows_phys + krow) * 8, 1024, &kv_full[st]);
                    ptx::bulk_g2s(sv + c * 1024, p.qkv + ((int64_t)(32 + h * 8 + c) * p.rows_phys + krow) * 8, 1024, &kv_full[st]);
                }
            }
            if (FUSED) {
                // W_out over the whole K/V ring once its last two tiles have been consumed
                ptx::mbar_wait(&kv_empty[n_steps & 1], (((n_steps) >> 1) & 1) ^ 1);
                ptx::mbar_wait(&kv_empty[(n_steps + 1) & 1], (((n_steps + 1) >> 1) & 1) ^ 1);
                ptx::mbar_expect_tx(w_full, 32768);
                ptx::bulk_g2s(sKV, p.g.w, 32768, w_full);
            }
        }
    } else if (warp == 1) {
        // MMA issuer.  idesc_s: M=128, N=64, A/B K-major.  idesc_o: M=128, N=64, B MN-major (bit 16).
        constexpr uint32_t idesc_s = ptx::make_idesc_bf16(128, 64);
        constexpr uint32_t idesc_o = ptx::make_idesc_bf16(128, 64) | (1u << 16);
        const uint32_t aQ = ptx::smem_u32(sQ), aP = ptx::smem_u32(sP), aKV = ptx::smem_u32(sKV);
        auto issue_s = [&](int g) {
            const int h = (FUSED && g >= n_kv) ? 1 : 0;     // head slot inside this CTA
            const int st = g & 1;
            ptx::mbar_wait(&kv_full[st], (g >> 1) & 1);
            ptx::tc_fence_after();
            if (g < 8) AT_TRACE(132 + 10 * g + 0);
            if (ptx::elect_one()) {
                const uint32_t aK = aKV + (uint32_t)st * 2 * ATC_K_BYTES;
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {   // dh = 64 = 4 x 16
                    const uint64_t da = ptx::make_desc(aQ + h * 16384 + kk * 2 * 2048, 2048, 128);
                    const uint64_t db = ptx::make_desc(aK + kk * 2 * 1024, 1024, 128);
                    ptx::mma_bf16(tm_s, da, db, idesc_s, kk != 0);
                }
                ptx::mma_commit(s_full);
            }
            __syncwarp();
            if (g < 8) AT_TRACE(132 + 10 * g + 1);
        };
        ptx::mbar_wait(q_full, 0);
        AT_TRACE(129);
        issue_s(0);
        for (int g = 0; g < n_steps; ++g) {
            const int h = (FUSED && g >= n_kv) ? 1 : 0, j = g - h * n_kv;
            const int st = g & 1;
            ptx::mbar_wait(p_full, g & 1);         // P_g in smem, S_g consumed, O rescaled if needed
            ptx::tc_fence_after();
            if (g < 8) AT_TRACE(132 + 10 * g + 2);
            if (ptx::elect_one()) {
                const uint32_t aV = aKV + (uint32_t)st * 2 * ATC_K_BYTES + ATC_K_BYTES;
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {   // 64 keys = 4 x 16
                    const uint64_t da = ptx::make_desc(aP + kk * 2 * 2048, 2048, 128);
                    const uint64_t db = ptx::make_desc(aV + kk * 256, p.v_lbo, p.v_sbo);
                    ptx::mma_bf16(tm_o + 64 * h, da, db, idesc_o, (j | kk) != 0);
                }
                ptx::mma_commit(&kv_empty[st]);
                if (g == n_steps - 1) ptx::mma_commit(o_done);
            }
            __syncwarp();
            if (g < 8) AT_TRACE(132 + 10 * g + 3);
            if (g + 1 < n_steps) issue_s(g + 1);   // queued behind P V_g: s_full(g+1) implies P V_g has completed
        }
        if (FUSED) {
            // out-proj: D[128 x 128] = O (bf16, over the Q tiles) · W_out (over the K/V ring)
            constexpr uint32_t idesc_d = ptx::make_idesc_bf16(128, 128);
            ptx::mbar_wait(w_full, 0);
            ptx::mbar_wait(oo_full, 0);
            ptx::tc_fence_after();
            if (ptx::elect_one()) {
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    const uint64_t da = ptx::make_desc(aQ + kk * 2 * 2048, 2048, 128);
                    const uint64_t db = ptx::make_desc(aKV + (kk >> 2) * 16384 + (kk & 3) * 2 * 2048, 2048, 128);
                    ptx::mma_bf16(tmem_base, da, db, idesc_d, kk != 0);
                }
                ptx::mma_commit(d_full);
            }
            __syncwarp();
            AT_TRACE(131);
        }
    } else {
        // softmax: thread ↔ query row
        const int lane_grp = warp & 3;
        const int row = lane_grp * 32 + lane;
        const uint32_t lane_off = (uint32_t)(lane_grp * 32) << 16;
        const uint32_t t_s = tm_s + lane_off;
        const float sl2 = 0.125f * 1.4426950408889634f;   // d^-0.5 * log2(e)
        float m_ref = -INFINITY, l = 0.f, l_head0 = 1.f;
        for (int g = 0; g < n_steps; ++g) {
            const int h = (FUSED && g >= n_kv) ? 1 : 0, j = g - h * n_kv;
            if (FUSED && g == n_kv) { l_head0 = l; l = 0.f; m_ref = -INFINITY; }     // head 1 starts
            const uint32_t t_o = tm_o + 64 * h + lane_off;
            const int nvalid = len - j * ATC_BK;               // keys of this tile that exist (>= 1)
            const bool full = nvalid >= ATC_BK;
            ptx::mbar_wait(s_full, g & 1);                     // also: P V_{g-1} complete → sP free, O stable
            ptx::tc_fence_after();
            if (g < 8) AT_TRACE(132 + 10 * g + 4);
            // pass A: row maximum (S is read from TMEM twice, 32 columns at a time: holding all 64 scores across both
            // passes costs registers that the second resident CTA needs)
            float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
            {
                float va[32], vb[32];
                ptx::tmem_ld32(t_s, va);
                ptx::tmem_ld32(t_s + 32, vb);
                ptx::tmem_ld_wait_dep(va);
                if (full) {
#pragma unroll
                    for (int e = 0; e < 32; e += 4) {
                        m0 = fmaxf(m0, va[e]); m1 = fmaxf(m1, va[e + 1]); m2 = fmaxf(m2, va[e + 2]); m3 = fmaxf(m3, va[e + 3]);
                    }
#pragma unroll
                    for (int e = 0; e < 32; e += 4) {
                        m0 = fmaxf(m0, vb[e]); m1 = fmaxf(m1, vb[e + 1]); m2 = fmaxf(m2, vb[e + 2]); m3 = fmaxf(m3, vb[e + 3]);
                    }
                } else {
#pragma unroll
                    for (int e = 0; e < 32; ++e) {
                        if (e < nvalid) m0 = fmaxf(m0, va[e]);
                        if (32 + e < nvalid) m1 = fmaxf(m1, vb[e]);
                    }
                }
            }
            const float mxs = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)) * sl2;
            if (g < 8) AT_TRACE(132 + 10 * g + 5);
            // lazy reference update: keep m_ref while the new maximum stays within 2^ATC_LAZY of it
            const bool raise = mxs > m_ref + ATC_LAZY;         // always true on a head's first tile (m_ref = -inf)
            float corr = 1.f;
            if (raise) {
                corr = ptx::ex2_approx(m_ref - mxs);           // 0 on the first tile
                m_ref = mxs;
                l *= corr;
            }
            if (j > 0 && __any_sync(0xffffffffu, raise)) {
                // rescale this warp's 32 rows of O_h in TMEM (lanes that keep their reference use corr = 1)
                float o[32];
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    ptx::tmem_ld32(t_o + q * 32, o);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int e = 0; e < 32; ++e) o[e] *= corr;
                    ptx::tmem_st32(t_o + q * 32, o);
                }
                ptx::tmem_st_wait();
            }
            float ps0 = 0.f, ps1 = 0.f, ps2 = 0.f, ps3 = 0.f;
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                float v[32];
                ptx::tmem_ld32(t_s + q * 32, v);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    __align__(16) __nv_bfloat162 o2[4];
                    float pe[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        pe[e] = ptx::ex2_approx(fmaf(v[c * 8 + e], sl2, -m_ref));
                        if (!full && q * 32 + c * 8 + e >= nvalid) pe[e] = 0.f;
                    }
                    ps0 += pe[0] + pe[4]; ps1 += pe[1] + pe[5]; ps2 += pe[2] + pe[6]; ps3 += pe[3] + pe[7];
#pragma unroll
                    for (int e = 0; e < 4; ++e) o2[e] = __floats2bfloat162_rn(pe[2 * e], pe[2 * e + 1]);
                    *reinterpret_cast<uint4*>(sP + (size_t)(q * 4 + c) * 2048 + row * 16) = *reinterpret_cast<const uint4*>(o2);
                }
            }
            l += (ps0 + ps1) + (ps2 + ps3);
            // make the generic-proxy smem writes visible to the tensor core (async proxy), then signal
            ptx::fence_proxy_async();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(p_full);
            if (g < 8) AT_TRACE(132 + 10 * g + 6);
        }
        // O_h = (sum_j P_j V_j) / l_h
        ptx::mbar_wait(o_done, 0);
        ptx::tc_fence_after();
        const int64_t r = rs + q0 + row;
        const bool row_ok = q0 + row < len;
#pragma unroll
        for (int q = 0; q < (FUSED ? 4 : 2); ++q) {     // 32 columns each: FUSED head 0 dims 0-63, head 1 dims 0-63
            const float inv = 1.f / ((FUSED && q < 2) ? l_head0 : l);
            float o[32];
            ptx::tmem_ld32(tm_o + lane_off + q * 32, o);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                __align__(16) __nv_bfloat162 o2[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) o2[e] = __floats2bfloat162_rn(o[c * 8 + 2 * e] * inv, o[c * 8 + 2 * e + 1] * inv);
                if (FUSED) {
                    // A operand of the out-proj: [16 chunk][128 row][8] over the Q tiles (all S MMAs have completed)
                    *reinterpret_cast<uint4*>(sQ + (size_t)(q * 4 + c) * 2048 + row * 16) = *reinterpret_cast<const uint4*>(o2);
                } else if (row_ok) {
                    *reinterpret_cast<uint4*>(p.out + ((int64_t)(h0 * 8 + q * 4 + c) * p.rows_phys + DX_HALO + r) * 8) =
                        *reinterpret_cast<const uint4*>(o2);
                }
            }
        }
        if (FUSED) {
            ptx::fence_proxy_async();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(oo_full);
            ptx::mbar_wait(d_full, 0);
            ptx::tc_fence_after();
            // rows of this utterance only: rows >= len of the tile belong to the gap / the next utterance
            tc_epilogue_tile<128, TC_EPI_LN>(p.g, p.R, rs + len, r, 0, tmem_base + lane_off, nullptr, lane, true, b, s_bias, s_lng, s_lnb);
            // the gap rows after the utterance are conv halos of the FFN: keep them zero
            if (q0 + ATC_BQ >= len && row < DX_GAP * 16) {
                const int gr = row >> 4, c = row & 15;
                *reinterpret_cast<uint4*>(p.g.out_bf + ((int64_t)c * p.g.out_rows_phys + DX_HALO + rs + len + gr) * 8) = make_uint4(0u, 0u, 0u, 0u);
            }
        }
        ptx::tc_fence_before();
    }
    if (warp == 2) AT_TRACE(130);
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<ATC_TMEM_COLS>(tmem_base);
    }
}

// zero `n_rows` rows starting at the stream's total (tiles of the last utterance read a little past the end of
// the stream; those rows must hold finite values because 0 * NaN would poison the P·V product)
__global__ void zero_tail_rows_kernel(__nv_bfloat16* buf, int nchunks, int64_t rows_phys, Rows R, int n_rows) {
    ptx::griddep_wait();      // PDL: predecessor complete + visible
    ptx::griddep_launch();    // let the next kernel start its prologue
    const int64_t total = R.row_start[R.B];
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)nchunks * n_rows) return;
    const int c = (int)(i / n_rows);
    const int64_t row = DX_HALO + total + (i % n_rows);
    if (row >= rows_phys) return;
    *reinterpret_cast<uint4*>(buf + ((int64_t)c * rows_phys + row) * 8) = make_uint4(0u, 0u, 0u, 0u);
}

int zero_tail_rows(__nv_bfloat16* buf, int nchunks, int64_t rows_phys, const Rows& R, int n_rows, cudaStream_t s) {
    const int64_t n = (int64_t)nchunks * n_rows;
    DX_LAUNCH_PDL((zero_tail_rows_kernel), (int)((n + 255) / 256), 256, 0, s, buf, nchunks, rows_phys, R, n_rows);
    DX_CHECK_LAUNCH();
    return 0;
}

static int attention_launch(const AtcParams& p, bool fused, cudaStream_t s) {
    static bool configured = false;
    if (!configured) {
        DX_CHECK_CUDA(cudaFuncSetAttribute(attention_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)AtcCfg<true>::SMEM));
        DX_CHECK_CUDA(cudaFuncSetAttribute(attention_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)AtcCfg<false>::SMEM));
        configured = true;
    }
    dim3 grid((unsigned)((p.R.n_max + ATC_BQ - 1) / ATC_BQ), fused ? 1 : DX_HEADS, (unsigned)p.R.B);
    ProfScope prof(fused ? "attn_block_tc" : "attn_tc", s);
    if (fused) DX_LAUNCH_PDL((attention_tc_kernel<true>), grid, ATC_THREADS, AtcCfg<true>::SMEM, s, p);
    else DX_LAUNCH_PDL((attention_tc_kernel<false>), grid, ATC_THREADS, AtcCfg<false>::SMEM, s, p);
    DX_CHECK_LAUNCH();
    return 0;
}

static void attention_common(AtcParams& p, const __nv_bfloat16* qkv, int64_t rows_phys, const Rows& R) {
    memset(&p, 0, sizeof(p));
    p.qkv = qkv; p.rows_phys = rows_phys; p.R = R;
    p.trace = reinterpret_cast<long long*>(trace_buffer(256 * sizeof(long long)));
    p.v_lbo = 128; p.v_sbo = ATC_BK * 16;
    if (debug_knob(1)) { p.v_lbo = ATC_BK * 16; p.v_sbo = 128; }
}

// attention alone: out [16][rows_phys][8] = concat_h softmax(Q_h K_h^T / 8) V_h
int attention_bf16(const __nv_bfloat16* qkv, __nv_bfloat16* out, int64_t rows_phys, const Rows& R, cudaStream_t s) {
    AtcParams p;
    attention_common(p, qkv, rows_phys, R);
    p.out = out;
    return attention_launch(p, false, s);
}

// attention + out-proj + residual + LayerNorm + pad mask: g carries the out-proj (w packed nt 128 / ks 64, bias,
// ln_g, ln_b, res, out_bf and their row strides)
int attention_block_bf16(const __nv_bfloat16* qkv, int64_t rows_phys, const TcGemmArgs& g, const Rows& R, cudaStream_t s) {
    DX_REQUIRE(g.K == 128 && g.N == 128 && g.nt == 128 && g.ks == 64 && g.taps == 1, "attention_block_bf16: out-proj must be 128 x 128 packed (nt 128, ks 64)");
    DX_REQUIRE(g.w && g.bias && g.ln_g && g.ln_b && g.out_bf, "attention_block_bf16: out-proj weights / LN parameters / output required");
    AtcParams p;
    attention_common(p, qkv, rows_phys, R);
    p.g = g;
    return attention_launch(p, true, s);
}

}  // namespace dx
