// Attention backward on tcgen05 for the training step's tensor-core mode (DX_TRAIN_TC_GEMM).
//
// Reference: autograd's backward of nn.MultiheadAttention inside MultiHeadAttention.forward (model.py:182-186) under
// loss.backward() (train.py:434).  With S = Q K^T / 8, P = softmax(S) (keys masked by length), O = P V:
//     dV = P^T dO      dP = dO V^T      dS = P o (dP - D),  D = rowsum(dO o O)      dQ = dS K / 8      dK = dS^T Q / 8
// No N x N tensor is kept by the forward: P is recomputed from Q, K and the saved log-sum-exp (as the fp32 kernels of
// dx_train_kernels.cu do, which this replaces: 6.4 of the step's 18 ms).  Two kernels from one template:
//   DQ   one CTA per (utterance, head, 128-query tile), loops over 128-key tiles:   S, dP -> dS -> dQ += dS K
//   DKV  one CTA per (utterance, head, 128-key tile),   loops over 128-query tiles: S, dP -> P, dS -> dV += P^T dO, dK += dS^T Q
// Operands are bf16 copies in the channel-interleaved layout [C/8][rows][8] (rows_to_il of qkv and dO), fp32 accumulation
// in TMEM; lse and D stay fp32.  Every operand form comes straight from that layout:
//   Q K^T, dO V^T   A and B K-major   (tile [8 d-chunk][128 rows][16 B]: LBO = chunk stride, SBO = 128)
//   dS K            A = dS K-major    (the epilogue writes P / dS as [16 key-chunk][128 query rows][16 B]), B = K MN-major
//   P^T dO, dS^T Q  A MN-major        (the same P / dS tile read with rows as K: LBO = 128, SBO = chunk stride), B MN-major
// Per tile pair: 8 MMAs (S, dP), a thread-per-query-row epilogue (exp, mask, dS; bf16 tiles to shared memory), 8 or 16 MMAs.
// The streamed tiles are double buffered; everything else is one step at a time (the kernels are ~8x faster than the fp32
// ones as they are).  Deterministic: each output element has one owner CTA and a fixed accumulation order.
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_ptx.cuh"
#include "dx_train.cuh"

namespace dx {

constexpr int ABT_THREADS = 192;                    // producer, issuer, 4 epilogue warps
constexpr uint32_t ABT_TILE = 8 * 128 * 16;         // 16 KB: [8 d-chunk][128 rows][8 bf16]
constexpr uint32_t ABT_PT = 16 * 128 * 16;          // 32 KB: [16 key-chunk][128 query rows][8 bf16]
// smem: fixed pair (2 tiles) | streamed pair x 2 stages (4 tiles) | P | dS | barriers
constexpr uint32_t ABT_OFF_ST = 2 * ABT_TILE;
constexpr uint32_t ABT_OFF_P = ABT_OFF_ST + 4 * ABT_TILE;
constexpr uint32_t ABT_OFF_DS = ABT_OFF_P + ABT_PT;
constexpr uint32_t ABT_OFF_BAR = ABT_OFF_DS + ABT_PT;
constexpr uint32_t ABT_NBARS = 1 + 2 + 2 + 4 + 1;   // fx_full, st_full[2], st_empty[2], s_full, s_empty, p_full, p_empty, acc_full
constexpr uint32_t ABT_SMEM = ABT_OFF_BAR + ABT_NBARS * 8 + 16;

struct AbtParams {
    const __nv_bfloat16* qkv;   // [48][rows_phys][8]
    const __nv_bfloat16* dO;    // [16][rows_phys][8]
    int64_t rows_phys;
    const float *lse, *Dr;      // [rows][2]
    DropSpec drop;              // DROP instantiations: the forward's attention-weight dropout, regenerated here
    uint32_t drop_site;
    Rows R;
    float* dqkv;                // [rows][384] fp32
};

template <bool DQ, bool DROP>
__global__ void __launch_bounds__(ABT_THREADS, 1) attn_bwd_tc_kernel(const AbtParams p) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    extern __shared__ __align__(1024) uint8_t ab_smem[];
    const int b = blockIdx.z, h = blockIdx.y;
    const int len = (int)min((long long)p.R.len[b], (long long)p.R.n_max);
    const int f0 = blockIdx.x * 128;                 // first row of the CTA's fixed tile (queries for DQ, keys for DKV)
    if (f0 >= len) return;
    const int64_t rs = p.R.row_start[b];
    const int n_st = (len + 127) / 128;              // streamed tiles (keys for DQ, queries for DKV)

    uint8_t* sFx = ab_smem;                          // DQ: Q | dO     DKV: K | V
    uint8_t* sSt = ab_smem + ABT_OFF_ST;             // DQ: K | V      DKV: Q | dO   (x 2 stages)
    uint8_t* sP = ab_smem + ABT_OFF_P;
    uint8_t* sDS = ab_smem + ABT_OFF_DS;
    uint64_t* bars = reinterpret_cast<uint64_t*>(ab_smem + ABT_OFF_BAR);
    uint64_t* fx_full = bars;
    uint64_t* st_full = bars + 1;                    // [2]
    uint64_t* st_empty = bars + 3;                   // [2]
    uint64_t* s_full = bars + 5;                     // S, dP accumulators written
    uint64_t* s_empty = bars + 6;                    // ... read by the epilogue (4 warps)
    uint64_t* p_full = bars + 7;                     // P / dS tiles written (4 warps)
    uint64_t* p_empty = bars + 8;                    // ... consumed by the MMAs
    uint64_t* acc_full = bars + 9;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + ABT_NBARS);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        ptx::mbar_init(fx_full, 1);
        for (int i = 0; i < 2; ++i) { ptx::mbar_init(&st_full[i], 1); ptx::mbar_init(&st_empty[i], 1); }
        ptx::mbar_init(s_full, 1); ptx::mbar_init(s_empty, 4);
        ptx::mbar_init(p_full, 4); ptx::mbar_init(p_empty, 1);
        ptx::mbar_init(acc_full, 1);
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<512>(tmem_ptr);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    // TMEM columns: S 0..127, dP 128..255, accumulators 256.. (DQ: dQ 64; DKV: dV 64 | dK 64)
    const uint32_t t_s = tmem_base, t_dp = tmem_base + 128, t_acc = tmem_base + 256;
    // chunk offsets inside qkv_il: Q 0, K 16, V 32 (+ 8 h); dO: 8 h
    const int cq = h * 8, ck = 16 + h * 8, cv = 32 + h * 8;

    if (warp == 0) {
        if (ptx::elect_one()) {
            const int64_t fr = DX_HALO + rs + f0;
            ptx::mbar_expect_tx(fx_full, 2 * ABT_TILE);
            for (int c = 0; c < 8; ++c) {
                if (DQ) {
                    ptx::bulk_g2s(sFx + c * 2048, p.qkv + ((int64_t)(cq + c) * p.rows_phys + fr) * 8, 2048, fx_full);
                    ptx::bulk_g2s(sFx + ABT_TILE + c * 2048, p.dO + ((int64_t)(cq + c) * p.rows_phys + fr) * 8, 2048, fx_full);
                } else {
                    ptx::bulk_g2s(sFx + c * 2048, p.qkv + ((int64_t)(ck + c) * p.rows_phys + fr) * 8, 2048, fx_full);
                    ptx::bulk_g2s(sFx + ABT_TILE + c * 2048, p.qkv + ((int64_t)(cv + c) * p.rows_phys + fr) * 8, 2048, fx_full);
                }
            }
            for (int it = 0; it < n_st; ++it) {
                const int s = it & 1;
                ptx::mbar_wait(&st_empty[s], ((it >> 1) & 1) ^ 1);
                uint8_t* d0 = sSt + (size_t)s * 2 * ABT_TILE;
                const int64_t sr = DX_HALO + rs + (int64_t)it * 128;
                ptx::mbar_expect_tx(&st_full[s], 2 * ABT_TILE);
                for (int c = 0; c < 8; ++c) {
                    if (DQ) {
                        ptx::bulk_g2s(d0 + c * 2048, p.qkv + ((int64_t)(ck + c) * p.rows_phys + sr) * 8, 2048, &st_full[s]);
                        ptx::bulk_g2s(d0 + ABT_TILE + c * 2048, p.qkv + ((int64_t)(cv + c) * p.rows_phys + sr) * 8, 2048, &st_full[s]);
                    } else {
                        ptx::bulk_g2s(d0 + c * 2048, p.qkv + ((int64_t)(cq + c) * p.rows_phys + sr) * 8, 2048, &st_full[s]);
                        ptx::bulk_g2s(d0 + ABT_TILE + c * 2048, p.dO + ((int64_t)(cq + c) * p.rows_phys + sr) * 8, 2048, &st_full[s]);
                    }
                }
            }
        }
    } else if (warp == 1) {
        constexpr uint32_t id_s = ptx::make_idesc_bf16(128, 128);                               // A, B K-major
        constexpr uint32_t id_q = ptx::make_idesc_bf16(128, 64) | (1u << 16);                   // A K-major, B MN-major
        constexpr uint32_t id_kv = ptx::make_idesc_bf16(128, 64) | (1u << 15) | (1u << 16);     // A, B MN-major
        const uint32_t aFx = ptx::smem_u32(sFx), aP = ptx::smem_u32(sP), aDS = ptx::smem_u32(sDS);
        ptx::mbar_wait(fx_full, 0);
        for (int it = 0; it < n_st; ++it) {
            const int s = it & 1;
            const uint32_t aSt = ptx::smem_u32(sSt) + (uint32_t)s * 2 * ABT_TILE;
            // the tiles of this step: Q, dO (query rows) and K, V (key rows)
            const uint32_t aQ = DQ ? aFx : aSt, adO = DQ ? aFx + ABT_TILE : aSt + ABT_TILE;
            const uint32_t aK = DQ ? aSt : aFx, aV = DQ ? aSt + ABT_TILE : aFx + ABT_TILE;
            ptx::mbar_wait(&st_full[s], (it >> 1) & 1);
            ptx::mbar_wait(s_empty, (it & 1) ^ 1);                   // the epilogue has read the previous S / dP
            ptx::tc_fence_after();
            if (ptx::elect_one()) {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)       // S = Q K^T, d = 64 = 4 x 16
                    ptx::mma_bf16(t_s, ptx::make_desc(aQ + kk * 4096, 2048, 128), ptx::make_desc(aK + kk * 4096, 2048, 128), id_s, kk != 0);
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)       // dP = dO V^T
                    ptx::mma_bf16(t_dp, ptx::make_desc(adO + kk * 4096, 2048, 128), ptx::make_desc(aV + kk * 4096, 2048, 128), id_s, kk != 0);
                ptx::mma_commit(s_full);
            }
            __syncwarp();
            ptx::mbar_wait(p_full, it & 1);                          // P / dS tiles of this step are in shared memory
            ptx::tc_fence_after();
            if (ptx::elect_one()) {
                if (DQ) {
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk)   // dQ += dS K: reduction over the 128 keys
                        ptx::mma_bf16(t_acc, ptx::make_desc(aDS + kk * 4096, 2048, 128), ptx::make_desc(aK + kk * 256, 128, 2048), id_q,
                                      it != 0 || kk != 0);
                } else {
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk)   // dV += P^T dO: reduction over the 128 queries
                        ptx::mma_bf16(t_acc, ptx::make_desc(aP + kk * 256, 128, 2048), ptx::make_desc(adO + kk * 256, 128, 2048), id_kv,
                                      it != 0 || kk != 0);
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk)   // dK += dS^T Q
                        ptx::mma_bf16(t_acc + 64, ptx::make_desc(aDS + kk * 256, 128, 2048), ptx::make_desc(aQ + kk * 256, 128, 2048), id_kv,
                                      it != 0 || kk != 0);
                }
                ptx::mma_commit(&st_empty[s]);
                ptx::mma_commit(p_empty);
                if (it == n_st - 1) ptx::mma_commit(acc_full);
            }
            __syncwarp();
        }
    } else {
        const int lane_grp = warp & 3;
        const int row = lane_grp * 32 + lane;                        // TMEM lane: query row of the S / dP tiles
        const uint32_t lane_off = (uint32_t)(lane_grp * 32) << 16;
        for (int it = 0; it < n_st; ++it) {
            const int q0 = DQ ? f0 : it * 128, k0 = DQ ? it * 128 : f0;
            const int q = q0 + row;
            const bool q_ok = q < len;
            float lse = 0.f, dr = 0.f;
            const uint64_t drop_row = DROP ? drop_pair_attn_row(b, h, q, p.R.n_max) + (uint64_t)(k0 >> 1) : 0;
            if (q_ok) {
                lse = p.lse[(rs + q) * DX_HEADS + h];
                dr = p.Dr[(rs + q) * DX_HEADS + h];
            }
            ptx::mbar_wait(s_full, it & 1);
            ptx::tc_fence_after();
            ptx::mbar_wait(p_empty, (it & 1) ^ 1);                   // the previous step's MMAs have consumed the P / dS tiles
#pragma unroll 1
            for (int g4 = 0; g4 < 4; ++g4) {
                float sv[32], dp[32];
                ptx::tmem_ld32(t_s + lane_off + g4 * 32, sv);
                ptx::tmem_ld32(t_dp + lane_off + g4 * 32, dp);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    __align__(16) __nv_bfloat162 p2[4], d2[4];
#pragma unroll
                    for (int e = 0; e < 8; e += 2) {
                        float pv[2], dv[2];
                        uint32_t z = 0;
                        if constexpr (DROP) z = drop_hash(p.drop.seed, p.drop_site, drop_row + ((g4 * 32 + c * 8 + e) >> 1));
#pragma unroll
                        for (int u = 0; u < 2; ++u) {
                            const int col = g4 * 32 + c * 8 + e + u;
                            const bool ok = q_ok && (k0 + col < len);
                            const float pe = ok ? __expf(fmaf(sv[c * 8 + e + u], 0.125f, -lse)) : 0.f;
                            if constexpr (DROP) {
                                // O = (mask * scale * P) V: dV takes the dropped weights, dP = mask * scale * (dO Vᵀ); the row
                                // term D = sum_k P dP = dO · O is unchanged
                                const bool keep = drop_keep(z, u, p.drop.thr);
                                pv[u] = keep ? pe * p.drop.scale : 0.f;
                                const float dpm = keep ? dp[c * 8 + e + u] * p.drop.scale : 0.f;
                                dv[u] = pe * (dpm - dr) * 0.125f;
                            } else {
                                pv[u] = pe;
                                dv[u] = pe * (dp[c * 8 + e + u] - dr) * 0.125f;
                            }
                        }
                        p2[e / 2] = __floats2bfloat162_rn(pv[0], pv[1]);
                        d2[e / 2] = __floats2bfloat162_rn(dv[0], dv[1]);
                    }
                    const uint32_t off = (uint32_t)(g4 * 4 + c) * 2048 + row * 16;
                    if (!DQ) *reinterpret_cast<uint4*>(sP + off) = *reinterpret_cast<const uint4*>(p2);
                    *reinterpret_cast<uint4*>(sDS + off) = *reinterpret_cast<const uint4*>(d2);
                }
            }
            ptx::tc_fence_before();
            ptx::fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
                ptx::mbar_arrive(s_empty);
                ptx::mbar_arrive(p_full);
            }
        }
        // accumulators -> global fp32: thread <-> fixed-tile row (query row for dQ, key row for dK / dV)
        ptx::mbar_wait(acc_full, 0);
        ptx::tc_fence_after();
        const int fr = f0 + row;
        constexpr int NACC = DQ ? 1 : 2;
#pragma unroll
        for (int a = 0; a < NACC; ++a) {
#pragma unroll
            for (int g2 = 0; g2 < 2; ++g2) {
                float v[32];
                ptx::tmem_ld32(t_acc + lane_off + a * 64 + g2 * 32, v);
                ptx::tmem_ld_wait();
                if (fr < len) {
                    // dQ -> columns 0..127, dK -> 128..255 (accumulator 1 of DKV), dV -> 256..383 (accumulator 0)
                    const int colbase = DQ ? 0 : (a == 0 ? 256 : 128);
                    float* o = p.dqkv + (rs + fr) * 384 + colbase + h * 64 + g2 * 32;
#pragma unroll
                    for (int e = 0; e < 32; e += 4) *reinterpret_cast<float4*>(o + e) = make_float4(v[e], v[e + 1], v[e + 2], v[e + 3]);
                }
            }
        }
        ptx::tc_fence_before();
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<512>(tmem_base);
    }
}

// qkv_il / dO_il: bf16 channel-interleaved copies of the block's saved QKV rows and of dO (rows_to_il); lse, Dr fp32 [rows][2];
// dqkv fp32 [rows][384] is fully overwritten at valid positions and zeroed elsewhere.
int attention_bwd_tc(const __nv_bfloat16* qkv_il, const __nv_bfloat16* dO_il, int64_t rows_phys, const float* lse, const float* Dr,
                     const Rows& R, float* dqkv, cudaStream_t s, const DropSpec* drop, uint32_t drop_site_id) {
    static SmemConfigured c1, c2, c3, c4;
    const bool use_drop = drop != nullptr && drop->thr != 0;
    if (use_drop) {
        DX_TRY(ensure_dyn_smem(attn_bwd_tc_kernel<true, true>, (size_t)ABT_SMEM, c3));
        DX_TRY(ensure_dyn_smem(attn_bwd_tc_kernel<false, true>, (size_t)ABT_SMEM, c4));
    } else {
        DX_TRY(ensure_dyn_smem(attn_bwd_tc_kernel<true, false>, (size_t)ABT_SMEM, c1));
        DX_TRY(ensure_dyn_smem(attn_bwd_tc_kernel<false, false>, (size_t)ABT_SMEM, c2));
    }
    AbtParams p;
    if (use_drop) { p.drop = *drop; p.drop_site = drop_site_id; }
    p.qkv = qkv_il; p.dO = dO_il; p.rows_phys = rows_phys; p.lse = lse; p.Dr = Dr; p.R = R; p.dqkv = dqkv;
    ProfScope prof("attn_bwd_tc", s);
    DX_CHECK_CUDA(cudaMemsetAsync(dqkv, 0, (size_t)R.cap * 384 * sizeof(float), s));
    dim3 grid((unsigned)((R.n_max + 127) / 128), DX_HEADS, (unsigned)R.B);
    if (use_drop) {
        DX_LAUNCH_PDL((attn_bwd_tc_kernel<true, true>), grid, ABT_THREADS, ABT_SMEM, s, p);
        DX_CHECK_LAUNCH();
        DX_LAUNCH_PDL((attn_bwd_tc_kernel<false, true>), grid, ABT_THREADS, ABT_SMEM, s, p);
    } else {
        DX_LAUNCH_PDL((attn_bwd_tc_kernel<true, false>), grid, ABT_THREADS, ABT_SMEM, s, p);
        DX_CHECK_LAUNCH();
        DX_LAUNCH_PDL((attn_bwd_tc_kernel<false, false>), grid, ABT_THREADS, ABT_SMEM, s, p);
    }
    DX_CHECK_LAUNCH();
    return 0;
}

}  // namespace dx
