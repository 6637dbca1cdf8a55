// Masked multi-head self-attention on tcgen05 (bf16 operands, fp32 scores / softmax / accumulation).
//
// Reference: MultiHeadAttention.forward (model.py:171-193) = nn.MultiheadAttention slow path (:182-186):
//   S_h = (Q_h · d^-0.5) K_h^T, -inf at padded keys, softmax over keys, O_h = A_h V_h.  The output projection, the
//   residual and the LayerNorm (:187-191) live in the fused block kernel (dx_ffn_tc.cu); the head-averaged weights
//   torch also returns are discarded by FFTBlock (model.py:255) and never materialised.
//
// One CTA = (utterance, head, 128-query tile).  K/V are streamed in 64-key tiles through a 3-stage ring.
//   S_g = Q K_g^T   : tcgen05.mma 128x64x16 x4, A = Q_h K-major, B = K tile K-major → TMEM S[g & 1] (64 columns each)
//   P_g = softmax   : 4 warps, thread ↔ query row, the row's 64 scores in registers; the exponent reference m_ref is
//                     only raised when the row maximum exceeds it by more than 8 (log2 units), so P <= 256 and the O
//                     rescale is rare; P written to smem buffer g & 1 as the bf16 K-major A operand
//   O  += P_g V_g   : tcgen05.mma 128x64x16 x4, B = V tile MN-major (the channel-interleaved global layout IS that
//                     canonical layout) → accumulates in TMEM across all K/V tiles
// The chain S → softmax → P·V of one tile is serial, the longest utterance's CTAs are the kernel's critical path, and
// the SM as a whole is bound by the softmax (MUFU ex2: 512 cycles per 128 x 64 tile; measured ~1000 with 2-3 warps per
// scheduler).  So S is double-buffered and issued TWO tiles ahead, and TWO softmax warp groups work on alternating
// tiles (group A even, group B odd), each with its own exponent reference, row sum, P buffer and O accumulator in
// TMEM; the two partial results are merged at the end (O = O_A 2^(m_A - m) + O_B 2^(m_B - m)).  A CTA thus has two
// tiles in flight and the SM four softmax warps per scheduler at 2 CTAs per SM (97 KB smem, 256 TMEM columns each).
// (The first version waited S → softmax → P·V → S: 2.6 k cycles per tile, profiles/r01c_ffn_attn_trace_after.txt.)
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_ptx.cuh"
#include "dx_dropout.cuh"
#include <string.h>

namespace dx {

constexpr int ATC_BQ = 128, ATC_BK = 64;
constexpr int ATC_THREADS = 352;                       // producer, 2 issuers (one per softmax group), 2 x 4 softmax warps
constexpr int ATC_SM_WARP0 = 3;                        // first softmax warp
constexpr int ATC_K_STAGES = 3;                        // K tile: live from its load to the end of its S MMAs (short)
constexpr int ATC_V_STAGES = 4;                        // V tile: live until its P·V has completed; tiles g .. g+3 are in flight
constexpr uint32_t ATC_Q_BYTES = 8 * 128 * 16;         // 16 KB: [8 dh chunk][128 row][8]
constexpr uint32_t ATC_K_BYTES = 8 * ATC_BK * 16;      // 8 KB:  [8 dh chunk][64 key][8]
constexpr uint32_t ATC_P_BYTES = 8 * 128 * 16;         // 16 KB: [8 key chunk][128 row][8]
constexpr uint32_t ATC_TMEM_COLS = 256;                // S0 | S1 | O_A | O_B
constexpr uint32_t ATC_NBARS = 1 + 2 * ATC_K_STAGES + 2 * ATC_V_STAGES + 2 + 2 + 2 + 2 + 1;
constexpr uint32_t ATC_SMEM = ATC_Q_BYTES + (ATC_K_STAGES + ATC_V_STAGES) * ATC_K_BYTES + 2 * ATC_P_BYTES + 2 * 128 * 8 + ATC_NBARS * 8 + 16;
constexpr float ATC_LAZY = 8.f;                        // raise the exponent reference only beyond this margin

struct AtcParams {
    const __nv_bfloat16* qkv;   // [48][rows_phys][8]
    __nv_bfloat16* out;         // attention output [16][rows_phys][8]
    int64_t rows_phys;
    Rows R;
    float* lse;                 // training: natural-log log-sum-exp of the scaled scores per (row, head), [rows][2], or null
    DropSpec drop;              // training (DROP instantiation): dropout of the attention weights inside nn.MultiheadAttention
    uint32_t drop_site;
    uint32_t v_lbo, v_sbo;      // MN-major descriptor strides of the V operand (debug knob 1 swaps them)
    long long* trace;           // bring-up: clock64 timeline of CTA (0,0,0) (dx_set_trace_buffer) or null
    int trace_ctas;             // > 0: the buffer also holds [256 + 4 i ..] = (start ns, end ns, SM id, K/V tiles) of CTA i
};
// timeline slots: [128] kernel start, [129] Q landed, [130] end; per K/V tile g < 8 (base 132 + 10 g):
// issuer +0 kv_full ok, +1 S issued, +2 p_full ok, +3 PV issued;  softmax +4 s_full ok, +5 max done, +6 P stored
#define AT_TRACE(slot) do { if (tr) tr[(slot)] = clock64(); } while (0)

template <bool DROP>
__global__ void __launch_bounds__(ATC_THREADS, 2) attention_tc_kernel(const AtcParams p) {
    extern __shared__ __align__(1024) uint8_t at_smem[];
    ptx::griddep_wait();      // PDL: the lengths and QKV come from earlier kernels
    ptx::griddep_launch();
    const int b = blockIdx.z;
    const int len = (int)min((long long)p.R.len[b], (long long)p.R.n_max);
    const int q0 = blockIdx.x * ATC_BQ;
    if (q0 >= len) return;
    if (p.trace != nullptr && threadIdx.x == 0) {       // [250] = -(earliest CTA start), [251] = latest CTA end (globaltimer ns), [252] = CTAs
        long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        atomicMax(reinterpret_cast<long long*>(p.trace) + 250, -t);
        atomicAdd(reinterpret_cast<unsigned long long*>(p.trace) + 252, 1ull);
        const int cta = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        if (cta < p.trace_ctas) {
            uint32_t smid;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            p.trace[256 + 4 * cta] = t;
            p.trace[256 + 4 * cta + 2] = smid;
            p.trace[256 + 4 * cta + 3] = (len + ATC_BK - 1) / ATC_BK;
        }
    }
    const int64_t rs = p.R.row_start[b];
    const int n_kv = (len + ATC_BK - 1) / ATC_BK;
    const int h = (int)blockIdx.y;

    uint8_t* sQ = at_smem;
    uint8_t* sK = sQ + ATC_Q_BYTES;                     // K ring
    uint8_t* sV = sK + ATC_K_STAGES * ATC_K_BYTES;      // V ring
    uint8_t* sP = sV + ATC_V_STAGES * ATC_K_BYTES;      // 2 buffers
    float2* s_ml = reinterpret_cast<float2*>(sP + 2 * ATC_P_BYTES);      // [2 group][128 row] (m_ref, l) for the merge
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_ml + 2 * 128);
    uint64_t* q_full = bars;                            // 1
    uint64_t* k_full = bars + 1;                        // [ATC_K_STAGES]
    uint64_t* k_empty = k_full + ATC_K_STAGES;          // [ATC_K_STAGES]  S of the tile has completed
    uint64_t* v_full = k_empty + ATC_K_STAGES;          // [ATC_V_STAGES]
    uint64_t* v_empty = v_full + ATC_V_STAGES;          // [ATC_V_STAGES]  P·V of the tile has completed
    uint64_t* s_full = v_empty + ATC_V_STAGES;          // [2]  S buffer written
    uint64_t* p_full = s_full + 2;                      // [2]  P buffer written (and S buffer consumed, O rescaled)
    uint64_t* pv_done = p_full + 2;                     // [2]  P·V reading P buffer has completed
    uint64_t* s_free = pv_done + 2;                     // [2]  the group holds S in registers: the next S may overwrite the buffer
    uint64_t* o_done = s_free + 2;                      // both groups' last P·V have completed
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + ATC_NBARS);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    long long* tr = (p.trace != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0 && (warp == 1 || warp == ATC_SM_WARP0)) ? p.trace : nullptr;
    if (warp == 1) AT_TRACE(128);
    if (threadIdx.x == 0) {
        ptx::mbar_init(q_full, 1);
        for (int i = 0; i < ATC_K_STAGES; ++i) { ptx::mbar_init(&k_full[i], 1); ptx::mbar_init(&k_empty[i], 1); }
        for (int i = 0; i < ATC_V_STAGES; ++i) { ptx::mbar_init(&v_full[i], 1); ptx::mbar_init(&v_empty[i], 1); }
        for (int i = 0; i < 2; ++i) {
            ptx::mbar_init(&s_full[i], 1); ptx::mbar_init(&p_full[i], 4); ptx::mbar_init(&pv_done[i], 1); ptx::mbar_init(&s_free[i], 4);
        }
        ptx::mbar_init(o_done, 2);
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<ATC_TMEM_COLS>(tmem_ptr);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    const uint32_t tm_s = tmem_base, tm_o = tmem_base + 128;

    if (warp == 0) {
        // ===================== producer: Q tile, then the K/V tiles =====================
        if (ptx::elect_one()) {
            ptx::mbar_expect_tx(q_full, ATC_Q_BYTES);
            for (int c = 0; c < 8; ++c)       // Q occupies qkv chunks 0..15: 0-7 = head 0, 8-15 = head 1
                ptx::bulk_g2s(sQ + c * 2048, p.qkv + ((int64_t)(h * 8 + c) * p.rows_phys + DX_HALO + rs + q0) * 8, 2048, q_full);
            for (int g = 0; g < n_kv; ++g) {
                const int ks = g % ATC_K_STAGES, vs = g % ATC_V_STAGES;
                const int64_t krow = DX_HALO + rs + (int64_t)g * ATC_BK;
                ptx::mbar_wait(&k_empty[ks], ((uint32_t)(g / ATC_K_STAGES) & 1) ^ 1);
                ptx::mbar_expect_tx(&k_full[ks], ATC_K_BYTES);
                for (int c = 0; c < 8; ++c)
                    ptx::bulk_g2s(sK + (size_t)ks * ATC_K_BYTES + c * 1024, p.qkv + ((int64_t)(16 + h * 8 + c) * p.rows_phys + krow) * 8, 1024, &k_full[ks]);
                ptx::mbar_wait(&v_empty[vs], ((uint32_t)(g / ATC_V_STAGES) & 1) ^ 1);
                ptx::mbar_expect_tx(&v_full[vs], ATC_K_BYTES);
                for (int c = 0; c < 8; ++c)
                    ptx::bulk_g2s(sV + (size_t)vs * ATC_K_BYTES + c * 1024, p.qkv + ((int64_t)(32 + h * 8 + c) * p.rows_phys + krow) * 8, 1024, &v_full[vs]);
            }
        }
    } else if (warp <= 2) {
        // ===================== MMA issuers, one per softmax group (warp 1: even tiles, warp 2: odd tiles) =============
        // A single issuer thread needs ~1100 cycles per tile (mbarrier waits ~170 each, 8 MMAs + 4 commits) and was the
        // CTA's bottleneck (profiles/r01f_attn_cta_times.txt); each group's chain S -> softmax -> P·V is private.
        // idesc_s: M=128, N=64, A/B K-major.  idesc_o: B MN-major (bit 16)
        const int grp = warp - 1;
        constexpr uint32_t idesc_s = ptx::make_idesc_bf16(128, 64);
        constexpr uint32_t idesc_o = ptx::make_idesc_bf16(128, 64) | (1u << 16);
        const uint32_t aQ = ptx::smem_u32(sQ), aP = ptx::smem_u32(sP) + grp * ATC_P_BYTES, aK0 = ptx::smem_u32(sK), aV0 = ptx::smem_u32(sV);
        const uint32_t t_s = tm_s + 64 * grp, t_o = tm_o + 64 * grp;
        auto issue_s = [&](int g) {
            const int st = g % ATC_K_STAGES;
            ptx::mbar_wait(&k_full[st], (uint32_t)(g / ATC_K_STAGES) & 1);
            ptx::tc_fence_after();
            if (g < 8) AT_TRACE(132 + 10 * g + 0);
            if (ptx::elect_one()) {
                const uint32_t aK = aK0 + (uint32_t)st * ATC_K_BYTES;
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {   // dh = 64 = 4 x 16
                    const uint64_t da = ptx::make_desc(aQ + kk * 2 * 2048, 2048, 128);
                    const uint64_t db = ptx::make_desc(aK + kk * 2 * 1024, 1024, 128);
                    ptx::mma_bf16(t_s, da, db, idesc_s, kk != 0);
                }
                ptx::mma_commit(&k_empty[st]);
                ptx::mma_commit(&s_full[grp]);
            }
            __syncwarp();
            if (g < 8) AT_TRACE(132 + 10 * g + 1);
        };
        if (grp < n_kv) {
            ptx::mbar_wait(q_full, 0);
            AT_TRACE(129);
            issue_s(grp);
            uint32_t it = 0;
            for (int g = grp; g < n_kv; g += 2, ++it) {
                if (g + 2 < n_kv) {
                    ptx::mbar_wait(&s_free[grp], it & 1);       // S_g is in the group's registers
                    ptx::tc_fence_after();
                    issue_s(g + 2);
                }
                const int st = g % ATC_V_STAGES;
                ptx::mbar_wait(&v_full[st], (uint32_t)(g / ATC_V_STAGES) & 1);
                ptx::mbar_wait(&p_full[grp], it & 1);           // P_g in smem, O rescaled if needed
                ptx::tc_fence_after();
                if (g < 8) AT_TRACE(132 + 10 * g + 2);
                if (ptx::elect_one()) {
                    const uint32_t aV = aV0 + (uint32_t)st * ATC_K_BYTES;
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {   // 64 keys = 4 x 16
                        const uint64_t da = ptx::make_desc(aP + kk * 2 * 2048, 2048, 128);
                        const uint64_t db = ptx::make_desc(aV + kk * 256, p.v_lbo, p.v_sbo);
                        ptx::mma_bf16(t_o, da, db, idesc_o, it != 0 || kk != 0);
                    }
                    ptx::mma_commit(&v_empty[st]);
                    ptx::mma_commit(&pv_done[grp]);
                    if (g + 2 >= n_kv) ptx::mma_commit(o_done);
                }
                __syncwarp();
                if (g < 8) AT_TRACE(132 + 10 * g + 3);
            }
        } else if (lane == 0) {
            ptx::mbar_arrive(o_done);                           // a group without tiles (n_kv == 1)
        }
    } else {
        // ===================== softmax: two groups of 4 warps on alternating tiles; thread ↔ query row ================
        const int grp = (warp - ATC_SM_WARP0) >> 2;         // 0: tiles 0, 2, 4, ...   1: tiles 1, 3, 5, ...
        const int lane_grp = warp & 3;
        const int row = lane_grp * 32 + lane;
        const uint32_t lane_off = (uint32_t)(lane_grp * 32) << 16;
        const uint32_t t_s = tm_s + 64 * grp + lane_off;
        const uint32_t t_o = tm_o + 64 * grp + lane_off;
        uint8_t* sPg = sP + grp * ATC_P_BYTES;
        const float sl2 = 0.125f * 1.4426950408889634f;   // d^-0.5 * log2(e)
        float m_ref = -INFINITY, l = 0.f;
        const uint64_t drop_row = DROP ? drop_pair_attn_row(b, h, q0 + row, p.R.n_max) : 0;
        for (int g = grp; g < n_kv; g += 2) {
            const uint32_t par = (g >> 1) & 1;
            const int nvalid = len - g * ATC_BK;               // keys of this tile that exist (>= 1)
            const bool full = nvalid >= ATC_BK;
            ptx::mbar_wait(&s_full[grp], par);
            ptx::tc_fence_after();
            if (g < 8) AT_TRACE(132 + 10 * g + 4);
            // pass A: row maximum.  S is read from TMEM twice (both 32-column halves in flight at once): holding all
            // 64 scores across both passes costs registers (spills at 2 CTAs x 320 threads per SM)
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
            const bool raise = mxs > m_ref + ATC_LAZY;         // always true on the group's first tile (m_ref = -inf)
            float corr = 1.f;
            if (raise) {
                corr = ptx::ex2_approx(m_ref - mxs);           // 0 on the first tile
                m_ref = mxs;
                l *= corr;
            }
            // the group's previous P·V (tile g-2) read this P buffer and wrote this O accumulator
            if (g >= 2) {
                ptx::mbar_wait(&pv_done[grp], par ^ 1);
                ptx::tc_fence_after();
                if (__any_sync(0xffffffffu, raise)) {
                    // rescale this warp's 32 rows of O in TMEM (lanes that keep their reference use corr = 1); P·V of
                    // tile g cannot start before this warp arrives on p_full
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
            }
            float ps0 = 0.f, ps1 = 0.f, ps2 = 0.f, ps3 = 0.f;
            {
                // pass B: the second 32-column load is in flight while the first half is processed; once it has landed
                // the S buffer is handed back to the issuer (S of the group's next tile overlaps the rest of this pass)
                float va[32], vb[32];
                ptx::tmem_ld32(t_s, va);
                ptx::tmem_ld_wait_dep(va);
                ptx::tmem_ld32(t_s + 32, vb);
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    if (q == 1) {
                        ptx::tmem_ld_wait_dep(vb);
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) ptx::mbar_arrive(&s_free[grp]);
                    }
                    const float* v = q == 0 ? va : vb;
                    if (full) {
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            __align__(16) __nv_bfloat162 o2[4];
                            float pe[8];
#pragma unroll
                            for (int e = 0; e < 8; ++e) pe[e] = ptx::ex2_approx(fmaf(v[c * 8 + e], sl2, -m_ref));
                            ps0 += pe[0] + pe[4]; ps1 += pe[1] + pe[5]; ps2 += pe[2] + pe[6]; ps3 += pe[3] + pe[7];
                            if constexpr (DROP) {       // the row sum above is of the UNdropped weights (softmax, then dropout)
#pragma unroll
                                for (int e = 0; e < 8; e += 2) {
                                    const uint32_t z = drop_hash(p.drop.seed, p.drop_site, drop_row + ((g * ATC_BK + q * 32 + c * 8 + e) >> 1));
                                    pe[e] = drop_keep(z, 0, p.drop.thr) ? pe[e] * p.drop.scale : 0.f;
                                    pe[e + 1] = drop_keep(z, 1, p.drop.thr) ? pe[e + 1] * p.drop.scale : 0.f;
                                }
                            }
#pragma unroll
                            for (int e = 0; e < 4; ++e) o2[e] = __floats2bfloat162_rn(pe[2 * e], pe[2 * e + 1]);
                            *reinterpret_cast<uint4*>(sPg + (size_t)(q * 4 + c) * 2048 + row * 16) = *reinterpret_cast<const uint4*>(o2);
                        }
                    } else {
#pragma unroll
                        for (int c = 0; c < 4; ++c) {
                            __align__(16) __nv_bfloat162 o2[4];
                            float pe[8];
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                pe[e] = ptx::ex2_approx(fmaf(v[c * 8 + e], sl2, -m_ref));
                                if (q * 32 + c * 8 + e >= nvalid) pe[e] = 0.f;
                            }
                            ps0 += pe[0] + pe[4]; ps1 += pe[1] + pe[5]; ps2 += pe[2] + pe[6]; ps3 += pe[3] + pe[7];
                            if constexpr (DROP) {
#pragma unroll
                                for (int e = 0; e < 8; e += 2) {
                                    const uint32_t z = drop_hash(p.drop.seed, p.drop_site, drop_row + ((g * ATC_BK + q * 32 + c * 8 + e) >> 1));
                                    pe[e] = drop_keep(z, 0, p.drop.thr) ? pe[e] * p.drop.scale : 0.f;
                                    pe[e + 1] = drop_keep(z, 1, p.drop.thr) ? pe[e + 1] * p.drop.scale : 0.f;
                                }
                            }
#pragma unroll
                            for (int e = 0; e < 4; ++e) o2[e] = __floats2bfloat162_rn(pe[2 * e], pe[2 * e + 1]);
                            *reinterpret_cast<uint4*>(sPg + (size_t)(q * 4 + c) * 2048 + row * 16) = *reinterpret_cast<const uint4*>(o2);
                        }
                    }
                }
            }
            l += (ps0 + ps1) + (ps2 + ps3);
            // make the generic-proxy smem writes visible to the tensor core (async proxy), then signal
            ptx::fence_proxy_async();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&p_full[grp]);
            if (g < 8) AT_TRACE(132 + 10 * g + 6);
        }
        // merge the two groups: m = max(m_A, m_B), O = (O_A 2^(m_A - m) + O_B 2^(m_B - m)) / (l_A 2^(m_A - m) + l_B 2^(m_B - m));
        // group A writes head dims 0-31, group B dims 32-63 (a group that saw no tile has m = -inf, l = 0: weight 0)
        s_ml[grp * 128 + row] = make_float2(m_ref, l);
        ptx::named_bar_sync(1, 256);
        const float2 mlA = s_ml[row], mlB = s_ml[128 + row];
        const float m = fmaxf(mlA.x, mlB.x);
        const float wa = ptx::ex2_approx(mlA.x - m), wb = n_kv > 1 ? ptx::ex2_approx(mlB.x - m) : 0.f;
        const float l_tot = mlA.y * wa + mlB.y * wb;
        const float inv = 1.f / l_tot;
        ptx::mbar_wait(o_done, 0);
        ptx::tc_fence_after();
        const int64_t r = rs + q0 + row;
        const bool row_ok = q0 + row < len;
        // P = 2^(s sl2 - m) whatever reference m the lazy update kept, so ln sum_k exp(s_k / 8) = m ln 2 + ln l exactly
        if (p.lse != nullptr && grp == 0 && row_ok) p.lse[r * DX_HEADS + h] = m * 0.6931471805599453f + __logf(l_tot);
        {
            float oa[32], ob[32];
            ptx::tmem_ld32(tm_o + lane_off + 32 * grp, oa);
            if (n_kv > 1) ptx::tmem_ld32(tm_o + 64 + lane_off + 32 * grp, ob);
            ptx::tmem_ld_wait();
            if (row_ok) {
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    __align__(16) __nv_bfloat162 o2[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        float y0 = oa[c * 8 + 2 * e] * wa, y1 = oa[c * 8 + 2 * e + 1] * wa;
                        if (n_kv > 1) { y0 = fmaf(ob[c * 8 + 2 * e], wb, y0); y1 = fmaf(ob[c * 8 + 2 * e + 1], wb, y1); }
                        o2[e] = __floats2bfloat162_rn(y0 * inv, y1 * inv);
                    }
                    *reinterpret_cast<uint4*>(p.out + ((int64_t)(h * 8 + grp * 4 + c) * p.rows_phys + DX_HALO + r) * 8) =
                        *reinterpret_cast<const uint4*>(o2);
                }
            }
        }
        ptx::tc_fence_before();
    }
    if (warp == ATC_SM_WARP0) AT_TRACE(130);
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<ATC_TMEM_COLS>(tmem_base);
    }
    if (p.trace != nullptr && threadIdx.x == 0) {
        long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        atomicMax(reinterpret_cast<long long*>(p.trace) + 251, t);
        const int cta = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        if (cta < p.trace_ctas) p.trace[256 + 4 * cta + 1] = t;
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

// out [16][rows_phys][8] = concat_h softmax(Q_h K_h^T / 8) V_h
int attention_bf16(const __nv_bfloat16* qkv, __nv_bfloat16* out, int64_t rows_phys, const Rows& R, cudaStream_t s, float* lse,
                   const DropSpec* drop, uint32_t drop_site_id) {
    static SmemConfigured configured, configured_drop;
    const bool use_drop = drop != nullptr && drop->thr != 0;
    if (use_drop) DX_TRY(ensure_dyn_smem(attention_tc_kernel<true>, ATC_SMEM, configured_drop));
    else DX_TRY(ensure_dyn_smem(attention_tc_kernel<false>, ATC_SMEM, configured));
    AtcParams p;
    memset(&p, 0, sizeof(p));
    p.qkv = qkv; p.out = out; p.rows_phys = rows_phys; p.R = R; p.lse = lse;
    if (use_drop) { p.drop = *drop; p.drop_site = drop_site_id; }
    p.trace = reinterpret_cast<long long*>(trace_buffer(256 * sizeof(long long)));
    if (p.trace != nullptr && debug_knob(10) % 100 > 0) {       // knob 10 = n: trace only the n-th attention launch of a step (1..12)
        static int n_traced = 0;
        if ((n_traced++ % 12) + 1 != debug_knob(10) % 100) p.trace = nullptr;
    }
    dim3 grid((unsigned)((R.n_max + ATC_BQ - 1) / ATC_BQ), DX_HEADS, (unsigned)R.B);
    const size_t n_cta = (size_t)grid.x * grid.y * grid.z;
    p.trace_ctas = (p.trace != nullptr && trace_buffer((256 + 4 * n_cta) * sizeof(long long)) != nullptr) ? (int)n_cta : 0;
    p.v_lbo = 128; p.v_sbo = ATC_BK * 16;
    if (debug_knob(1)) { p.v_lbo = ATC_BK * 16; p.v_sbo = 128; }
    ProfScope prof("attn_tc", s);
    if (use_drop) DX_LAUNCH_PDL((attention_tc_kernel<true>), grid, ATC_THREADS, ATC_SMEM, s, p);
    else DX_LAUNCH_PDL((attention_tc_kernel<false>), grid, ATC_THREADS, ATC_SMEM, s, p);
    DX_CHECK_LAUNCH();
    return 0;
}

}  // namespace dx
