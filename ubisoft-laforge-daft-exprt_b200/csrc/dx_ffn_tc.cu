// Fused position-wise conv feed-forward for sm_100a (tcgen05 + TMEM + bulk-copy pipeline):
//
//   y = mask( FiLM( LN( conv1d_{C→D,k3}( ReLU( conv1d_{D→C,k3}(x) + b1 ) ) + b2 + x ) ) )
//
// = PositionWiseConvFF.forward (model.py:220-235) + the FFTBlock pad mask (model.py:258), evaluated on the packed row
// stream with the reference's "live padding" semantics (H1): the hidden activation of the first row after an
// utterance is live, everything beyond is true zero padding.  The 1024-wide hidden activation never leaves the SM.
//
// One CTA = 252 output rows = two "halves" of 128 hidden rows (126 outputs each, 1 halo row per side).
//   X   [16 chunk][256 row][8]  bf16  rows R0-2 .. R0+253, loaded once per tile by 16 bulk copies; also the residual
//   for hidden chunk c (128 channels, 8 chunks):
//     G1  D1[h] (128x128 fp32, TMEM) = sum_tap X_h[. + tap] · W1[c][tap]        24 MMAs (128x128x16) per half
//     E1  Hs[h][sub] = bf16(live ? ReLU(D1[h][:, 64 sub ..] + b1) : 0)          sub = 0,1; thread ↔ row
//     G2  D2[h] (128x128 fp32, TMEM) += sum_tap Hs[h][sub][. + tap] · W2[tap][c, sub]   12 MMAs per half and sub
//   E2  out = mask(FiLM(LN(D2[h] + b2 + X_res)))                                 thread ↔ row, LN in registers
// Issue order G1(c), G2(c-1, 0), G2(c-1, 1): E1(c) runs under G2(c-1).
//
// Why these shapes (profiles/r01b_mma_bench.txt, r01b_ffn_trace_before.txt): a 128xNx16 SS MMA reads (128+N)*32 B
// of shared memory in N/2 cycles, so N = 64 needs 192 B/clk against the SM's 128 B/clk; N = 128 runs at the
// tensor-pipe rate.  Every 16 KB weight stage is used by both halves.  Conv taps are +16 B shifts of the
// no-swizzle K-major descriptors (see dx_tc_gemm.cu; measured free, profiles/r01d_mma_ring_bench.txt).
//
// Why TWO MMA issuer warps (one per half): tcgen05.mma issue blocks while the tensor pipe's short queue is full, and
// an mbarrier wait costs the issuing thread ~170 cycles even when the barrier has already completed, so a single
// issuer leaves the pipe idle for ~180 cycles per weight stage (88-100 cycles per MMA instead of 64, whatever the
// ring depth; tools/ring_bench.cu).  With one issuer per half the two halves are independent pipelines that share
// only the X tile and the weight ring: while one warp sits in a barrier wait the other's MMAs keep the pipe busy
// (72 cycles per MMA in the microbenchmark).  All D1 / Hs / D2 barriers are per half.
// Warps: 0 = bulk-copy producer, 1 = TMEM owner + issuer of half 0, 2 = issuer of half 1, 3-6 = epilogue of half 0,
// 7-10 = epilogue of half 1 (warp w may only touch TMEM lanes 32 (w % 4) .. +31).
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_ptx.cuh"
#include "dx_ffn_epilogue.cuh"

namespace dx {

constexpr int FF_THREADS = 352;
constexpr int FF_OUT = 126;                 // outputs per half
constexpr int FF_TILE = 2 * FF_OUT;         // outputs per CTA tile
constexpr int FF_NCH = 8;                   // hidden chunks of 128 channels
constexpr int FF_WSTAGES = 5;
constexpr uint32_t FF_X_BYTES = 16 * 256 * 16;          // 65536
constexpr uint32_t FF_HS_ROWS = 136;
constexpr uint32_t FF_HS_BYTES = 8 * FF_HS_ROWS * 16;   // 17408 per (half, sub)
constexpr uint32_t FF_W_BYTES = 16384;
constexpr uint32_t FF_OFF_HS = FF_X_BYTES;
constexpr uint32_t FF_OFF_W = FF_OFF_HS + 4 * FF_HS_BYTES;
constexpr uint32_t FF_OFF_PAR = FF_OFF_W + FF_WSTAGES * FF_W_BYTES;
constexpr uint32_t FF_PAR_FLOATS = 1024 + 128 + 128 + 128;          // b1, b2, gamma, beta
constexpr uint32_t FF_OFF_BAR = FF_OFF_PAR + FF_PAR_FLOATS * 4;
constexpr uint32_t FF_NBARS = 2 * FF_WSTAGES + 2 + 2 * 8;
constexpr uint32_t FF_SMEM = FF_OFF_BAR + FF_NBARS * 8 + 16;

struct FfnParams {
    const __nv_bfloat16* x;     // [16][rows_phys][8] block input (also the residual)
    __nv_bfloat16* out;         // [16][rows_phys][8]
    int64_t rows_phys;
    const __nv_bfloat16 *w1, *w2;   // packed: W1 (taps 3, K 128, N 1024, nt 128, ks 64), W2 (taps 3, K 1024, N 128, nt 128, ks 64)
    const float *b1, *b2, *ln_g, *ln_b;
    const float* film;          // pre-offset to the block, or null
    int film_stride;
    Rows R;
    int dbg;                    // timing experiments (debug knob 8): bit 0 = weight stages are not copied (arrive only),
                                // bit 1 = issue no MMAs, bit 2 = no Hs stores, bit 3 = no TMEM loads in E1 (results are wrong)
    long long* trace;           // bring-up: clock64 timeline of CTA 0 (dx_set_trace_buffer) or null
};

// timeline slots (CTA 0, half 0 only): [0] kernel start, [1] X landed, then per chunk c (base 8 + 12 c):
//   issuer  +0 d1_empty ok, +1 G1 issued, +2 hs_full[0] ok, +3 G2 issued (both subs), +4 first w_full wait of G1 done
//   epilog  +6 d1_full ok, +7 TMEM loaded (sub 1), +8 hs_empty[1] ok, +9 Hs stored + fenced (sub 1)
// [2] d2_full ok (E2 start), [3] E2 done, [4] CTA done
#define FF_TRACE(slot) do { if (tr) tr[(slot)] = clock64(); } while (0)

__global__ void __launch_bounds__(FF_THREADS, 1) ffn_tc_kernel(const FfnParams p) {
    extern __shared__ __align__(1024) uint8_t ff_smem[];
    uint8_t* sX = ff_smem;
    uint8_t* sHs = ff_smem + FF_OFF_HS;         // [half][sub]
    uint8_t* sW = ff_smem + FF_OFF_W;
    float* s_b1 = reinterpret_cast<float*>(ff_smem + FF_OFF_PAR);
    float* s_b2 = s_b1 + 1024;
    float* s_g = s_b2 + 128;
    float* s_be = s_g + 128;
    uint64_t* bars = reinterpret_cast<uint64_t*>(ff_smem + FF_OFF_BAR);
    uint64_t* w_full = bars;                       // [S]   producer → both issuers
    uint64_t* w_empty = bars + FF_WSTAGES;         // [S]   one commit per active issuer
    uint64_t* x_full = bars + 2 * FF_WSTAGES;
    uint64_t* x_empty = x_full + 1;
    uint64_t* hb = x_full + 2;                     // per half h (8 barriers each):
    // hb[8 h + 0] d1_full, +1 d1_empty, +2 hs_full[sub 0], +3 hs_full[sub 1], +4 hs_empty[0], +5 hs_empty[1], +6 d2_full, +7 d2_empty
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + FF_NBARS);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    long long* tr = (p.trace != nullptr && blockIdx.x == 0 && lane == 0 && (warp == 1 || warp == 3)) ? p.trace : nullptr;
    if (warp == 1) FF_TRACE(0);

    // ---- prologue that does not depend on the previous kernel (PDL: overlaps its tail) ----
    for (int i = threadIdx.x; i < 1024; i += FF_THREADS) s_b1[i] = p.b1[i];
    if (threadIdx.x < 128) {
        s_b2[threadIdx.x] = p.b2[threadIdx.x];
        s_g[threadIdx.x] = p.ln_g[threadIdx.x];
        s_be[threadIdx.x] = p.ln_b[threadIdx.x];
    }
    // rows 128..135 of the Hs buffers are read (taps of the two discarded output rows) but never written: zero once
    for (int i = threadIdx.x; i < 4 * 8 * 8; i += FF_THREADS) {
        const int q = i >> 6, kc = (i >> 3) & 7, rr = 128 + (i & 7);
        *reinterpret_cast<uint4*>(sHs + q * FF_HS_BYTES + kc * FF_HS_ROWS * 16 + rr * 16) = make_uint4(0u, 0u, 0u, 0u);
    }
    if (threadIdx.x == 0) {
        for (int i = 0; i < FF_WSTAGES; ++i) ptx::mbar_init(&w_full[i], 1);
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<512>(tmem_ptr);
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    // the first FF_WSTAGES weight stages are the same for every tile (W1 chunk 0): fetch them now
    if (warp == 0 && ptx::elect_one()) {
        for (int s0 = 0; s0 < FF_WSTAGES; ++s0) {
            if (p.dbg & 1) { ptx::mbar_arrive(&w_full[s0]); continue; }
            ptx::mbar_expect_tx(&w_full[s0], FF_W_BYTES);
            ptx::bulk_g2s(sW + s0 * FF_W_BYTES, reinterpret_cast<const uint8_t*>(p.w1) + (size_t)s0 * FF_W_BYTES, FF_W_BYTES, &w_full[s0]);
        }
    }
    __syncwarp();
    ptx::griddep_launch();
    ptx::griddep_wait();
    // ---- from here on the previous kernel's output (row map, activations) may be read ----
    const int64_t total = p.R.row_start[p.R.B];
    // Short streams (symbol side): when the 252-row tiles would leave more than half of the CTAs idle, every CTA
    // takes ONE half (126 outputs) instead, which halves the kernel's critical path.  Decided on the device: the
    // valid row count is only known there.
    const int halves = ((total + FF_TILE - 1) / FF_TILE) * 2 <= (int64_t)gridDim.x ? 1 : 2;
    const int tile_rows = FF_OUT * halves;
    const int n_tiles = (int)((total + tile_rows - 1) / tile_rows);
    const uint32_t x_chunk_bytes = halves == 2 ? 4096u : 2176u;      // 256 / 136 rows of one 8-channel chunk
    if (threadIdx.x == 0) {
        for (int i = 0; i < FF_WSTAGES; ++i) ptx::mbar_init(&w_empty[i], halves);
        ptx::mbar_init(x_full, 1);
        ptx::mbar_init(x_empty, 4 * halves);
        for (int h = 0; h < 2; ++h) {
            ptx::mbar_init(&hb[8 * h + 0], 1);          // d1_full
            ptx::mbar_init(&hb[8 * h + 1], 4);          // d1_empty
            ptx::mbar_init(&hb[8 * h + 2], 4);          // hs_full[0]
            ptx::mbar_init(&hb[8 * h + 3], 4);          // hs_full[1]
            ptx::mbar_init(&hb[8 * h + 4], 1);          // hs_empty[0]
            ptx::mbar_init(&hb[8 * h + 5], 1);          // hs_empty[1]
            ptx::mbar_init(&hb[8 * h + 6], 1);          // d2_full
            ptx::mbar_init(&hb[8 * h + 7], 4);          // d2_empty
        }
        ptx::fence_barrier_init();
    }
    __syncthreads();
    // TMEM columns: D1[h] at 128 h, D2[h] at 256 + 128 h
    const uint32_t tm_d1 = tmem_base, tm_d2 = tmem_base + 256;

    if (warp == 0) {
        // ===================== producer: X tile + weight stages =====================
        if (ptx::elect_one()) {
            uint32_t wit = 0, tile_it = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++tile_it) {
                const int64_t R0 = (int64_t)tile * tile_rows;
                ptx::mbar_wait(x_empty, (tile_it & 1) ^ 1);
                ptx::mbar_expect_tx(x_full, 16 * x_chunk_bytes);
                for (int kc = 0; kc < 16; ++kc)
                    ptx::bulk_g2s(sX + kc * 4096, p.x + ((int64_t)kc * p.rows_phys + DX_HALO + R0 - 2) * 8, x_chunk_bytes, x_full);
                // weight order must match the MMA issue order: W1(c) [2 K halves x 3 taps], then W2(c-1) [2 subs x 3 taps]
                for (int c = 0; c <= FF_NCH; ++c) {
                    for (int which = 0; which < 2; ++which) {
                        const int cc = which == 0 ? c : c - 1;
                        if (cc < 0 || cc >= FF_NCH) continue;
                        const uint8_t* base = reinterpret_cast<const uint8_t*>(which == 0 ? p.w1 : p.w2) + (size_t)cc * 6 * FF_W_BYTES;
                        for (int st = 0; st < 6; ++st, ++wit) {
                            if (wit < FF_WSTAGES) continue;          // fetched in the prologue
                            const int s = wit % FF_WSTAGES;
                            ptx::mbar_wait(&w_empty[s], ((wit / FF_WSTAGES) & 1) ^ 1);
                            if (p.dbg & 1) { ptx::mbar_arrive(&w_full[s]); continue; }
                            ptx::mbar_expect_tx(&w_full[s], FF_W_BYTES);
                            ptx::bulk_g2s(sW + s * FF_W_BYTES, base + (size_t)st * FF_W_BYTES, FF_W_BYTES, &w_full[s]);
                        }
                    }
                }
            }
        }
    } else if (warp <= 2) {
        // ===================== MMA issuers: warp 1 ↔ half 0, warp 2 ↔ half 1 =====================
        const int h = warp - 1;
        if ((int)blockIdx.x >= n_tiles && h == 0)      // no tile: the prefetched weight stages must land before the CTA exits
            for (int s0 = 0; s0 < FF_WSTAGES; ++s0) ptx::mbar_wait(&w_full[s0], 0);
        if (h < halves) {
            constexpr uint32_t idesc = ptx::make_idesc_bf16(128, 128);
            const uint32_t aX = ptx::smem_u32(sX) + FF_OUT * h * 16, aHs = ptx::smem_u32(sHs) + h * 2 * FF_HS_BYTES, aW = ptx::smem_u32(sW);
            const uint32_t td1 = tm_d1 + 128 * h, td2 = tm_d2 + 128 * h;
            uint64_t* d1_full = &hb[8 * h + 0];
            uint64_t* d1_empty = &hb[8 * h + 1];
            uint64_t* hs_full = &hb[8 * h + 2];
            uint64_t* hs_empty = &hb[8 * h + 4];
            uint64_t* d2_full = &hb[8 * h + 6];
            uint64_t* d2_empty = &hb[8 * h + 7];
            uint32_t wit = 0, tile_it = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++tile_it) {
                ptx::mbar_wait(x_full, tile_it & 1);
                if (tile_it == 0) FF_TRACE(1);
                ptx::mbar_wait(d2_empty, (tile_it & 1) ^ 1);
                ptx::tc_fence_after();
                for (int c = 0; c <= FF_NCH; ++c) {
                    if (c < FF_NCH) {
                        // ---- G1(c): D1[h] = X_h (*) W1[c]   (stage = (K half, tap): 64 input channels x 128 hidden)
                        const uint32_t n = tile_it * FF_NCH + c;
                        ptx::mbar_wait(d1_empty, (n & 1) ^ 1);
                        ptx::tc_fence_after();
                        if (tile_it == 0) FF_TRACE(8 + 12 * c + 0);
                        for (int st = 0; st < 6; ++st, ++wit) {
                            const int kh = st / 3, tap = st % 3;
                            const int s = wit % FF_WSTAGES;
                            ptx::mbar_wait(&w_full[s], (wit / FF_WSTAGES) & 1);
                            ptx::tc_fence_after();
                            if (tile_it == 0 && st == 0) FF_TRACE(8 + 12 * c + 4);
                            if (ptx::elect_one()) {
#pragma unroll
                                for (int kk = 0; kk < 4; ++kk) {
                                    const uint64_t da = ptx::make_desc(aX + (kh * 8 + kk * 2) * 4096 + tap * 16, 4096, 128);
                                    const uint64_t db = ptx::make_desc(aW + s * FF_W_BYTES + kk * 2 * 2048, 2048, 128);
                                    if (!(p.dbg & 2)) ptx::mma_bf16(td1, da, db, idesc, (st | kk) != 0);
                                }
                                ptx::mma_commit(&w_empty[s]);
                                if (st == 5) ptx::mma_commit(d1_full);
                            }
                            __syncwarp();
                        }
                        if (tile_it == 0) FF_TRACE(8 + 12 * c + 1);
                    }
                    if (c >= 1) {
                        // ---- G2(c-1, sub): D2[h] += Hs[h][sub] (*) W2[c-1, sub]   (stage = (sub, tap): 64 hidden x 128 out)
                        const int cc = c - 1;
                        const uint32_t k = tile_it * FF_NCH + cc;
                        for (int st = 0; st < 6; ++st, ++wit) {
                            const int sub = st / 3, tap = st % 3;
                            if (tap == 0) {
                                ptx::mbar_wait(&hs_full[sub], k & 1);
                                ptx::tc_fence_after();
                                if (tile_it == 0 && sub == 0) FF_TRACE(8 + 12 * cc + 2);
                            }
                            const int s = wit % FF_WSTAGES;
                            ptx::mbar_wait(&w_full[s], (wit / FF_WSTAGES) & 1);
                            ptx::tc_fence_after();
                            if (ptx::elect_one()) {
#pragma unroll
                                for (int kk = 0; kk < 4; ++kk) {
                                    const uint64_t da = ptx::make_desc(aHs + sub * FF_HS_BYTES + kk * 2 * FF_HS_ROWS * 16 + tap * 16, FF_HS_ROWS * 16, 128);
                                    const uint64_t db = ptx::make_desc(aW + s * FF_W_BYTES + kk * 2 * 2048, 2048, 128);
                                    if (!(p.dbg & 2)) ptx::mma_bf16(td2, da, db, idesc, (cc | st | kk) != 0);
                                }
                                ptx::mma_commit(&w_empty[s]);
                                if (tap == 2) {
                                    ptx::mma_commit(&hs_empty[sub]);
                                    if (cc == FF_NCH - 1 && sub == 1) ptx::mma_commit(d2_full);
                                }
                            }
                            __syncwarp();
                        }
                        if (tile_it == 0) FF_TRACE(8 + 12 * cc + 3);
                    }
                }
            }
        }
    } else if (((warp - 3) >> 2) < halves) {
        // ===================== epilogue: warps 3-6 own half 0, warps 7-10 own half 1; thread ↔ row =====================
        const int h = (warp - 3) >> 2;
        const int lane_grp = warp & 3;
        const int j = lane_grp * 32 + lane;                         // row inside the half
        const uint32_t lane_off = (uint32_t)(lane_grp * 32) << 16;
        const uint32_t td1 = tm_d1 + 128 * h + lane_off, td2 = tm_d2 + 128 * h + lane_off;
        uint64_t* d1_full = &hb[8 * h + 0];
        uint64_t* d1_empty = &hb[8 * h + 1];
        uint64_t* hs_full = &hb[8 * h + 2];
        uint64_t* hs_empty = &hb[8 * h + 4];
        uint64_t* d2_full = &hb[8 * h + 6];
        uint64_t* d2_empty = &hb[8 * h + 7];
        if (h != 0) tr = nullptr;
        uint32_t tile_it = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++tile_it) {
            const int64_t R0 = (int64_t)tile * tile_rows;
            const int64_t sh = R0 + FF_OUT * h - 1 + j;             // stream row of hidden row j
            const bool hid_live = sh >= 0 && sh < total && row_live(p.R, sh, 1);
            const int64_t so = R0 + FF_OUT * h + j;                 // stream row of output row j
            const bool out_write = j < FF_OUT && so < total;
            const bool out_live = out_write && row_live(p.R, so, 0);
            const int b_idx = out_write ? max(p.R.row_b[so], 0) : 0;
            ptx::mbar_wait(x_full, tile_it & 1);                     // residual rows are read from sX in E2
            for (int c = 0; c < FF_NCH; ++c) {
                const uint32_t n = tile_it * FF_NCH + c;
                ptx::mbar_wait(d1_full, n & 1);
                ptx::tc_fence_after();
                if (tile_it == 0) FF_TRACE(8 + 12 * c + 6);
#pragma unroll
                for (int sub = 0; sub < 2; ++sub) {
                    float v[64];
                    if (!(p.dbg & 8)) {
                        ptx::tmem_ld32(td1 + 64 * sub, v);
                        ptx::tmem_ld32(td1 + 64 * sub + 32, v + 32);
                        ptx::tmem_ld_wait();
                    } else {
#pragma unroll
                        for (int e = 0; e < 64; ++e) v[e] = (float)(e + j);
                    }
                    if (sub == 1) {
                        // D1 is in registers: G1(c+1) may overwrite it
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) ptx::mbar_arrive(d1_empty);
                        if (tile_it == 0) FF_TRACE(8 + 12 * c + 7);
                    }
                    ptx::mbar_wait(&hs_empty[sub], (n & 1) ^ 1);     // G2(c-1, sub) has finished reading this buffer
                    if (tile_it == 0 && sub == 1) FF_TRACE(8 + 12 * c + 8);
                    if (!(p.dbg & 4))
                        ffn_e1_store(v, s_b1 + c * 128 + sub * 64, hid_live, sHs + (h * 2 + sub) * FF_HS_BYTES + j * 16, FF_HS_ROWS * 16);
                    ptx::fence_proxy_async();        // Hs writes → visible to the tensor core's async proxy
                    ptx::tc_fence_before();
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(&hs_full[sub]);
                    if (tile_it == 0 && sub == 1) FF_TRACE(8 + 12 * c + 9);
                }
            }
            // ---- E2: LN epilogue of this half
            ptx::mbar_wait(d2_full, tile_it & 1);
            ptx::tc_fence_after();
            if (tile_it == 0) FF_TRACE(2);
            {
                float v[128];
#pragma unroll
                for (int q = 0; q < 4; ++q) ptx::tmem_ld32(td2 + q * 32, v + q * 32);
                ptx::tmem_ld_wait();
                if (out_write)
                    ffn_e2_store(v, sX + (FF_OUT * h + j + 2) * 16, s_b2, s_g, s_be, p.film ? p.film + (int64_t)b_idx * p.film_stride : nullptr,
                                 out_live, p.out + (DX_HALO + so) * 8, p.rows_phys * 8);
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (tile_it == 0) FF_TRACE(3);
            if (lane == 0) {
                ptx::mbar_arrive(d2_empty);
                ptx::mbar_arrive(x_empty);
            }
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        FF_TRACE(4);
        ptx::tc_fence_after();
        ptx::tmem_dealloc<512>(tmem_base);
    }
}

int ffn_bf16(const __nv_bfloat16* x, __nv_bfloat16* out, int64_t rows_phys, const DxBlockWeights& bw, const float* film,
             int film_stride, const Rows& R, cudaStream_t s) {
    DX_REQUIRE(bw.w_ffn1_bf && bw.w_ffn2_bf, "ffn_bf16: fused-FFN packed weights missing");
    DX_REQUIRE(x != out, "ffn_bf16: in-place operation is not supported (rows of neighbouring tiles are read as halos)");
    static int num_sms = 0;
    static bool configured = false;
    if (!configured) {
        int dev = 0;
        DX_CHECK_CUDA(cudaGetDevice(&dev));
        DX_CHECK_CUDA(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
        DX_CHECK_CUDA(cudaFuncSetAttribute(ffn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FF_SMEM));
        configured = true;
    }
    FfnParams p;
    p.x = x; p.out = out; p.rows_phys = rows_phys;
    p.w1 = reinterpret_cast<const __nv_bfloat16*>(bw.w_ffn1_bf);
    p.w2 = reinterpret_cast<const __nv_bfloat16*>(bw.w_ffn2_bf);
    p.b1 = bw.b_c1; p.b2 = bw.b_c2; p.ln_g = bw.ln2_g; p.ln_b = bw.ln2_b;
    p.film = film; p.film_stride = film_stride; p.R = R;
    p.dbg = debug_knob(8);
    p.trace = reinterpret_cast<long long*>(trace_buffer(256 * sizeof(long long)));
    int64_t tiles = (R.cap + FF_OUT - 1) / FF_OUT;     // enough CTAs for the one-half-per-CTA mode of short streams
    int grid = (int)(tiles < num_sms ? tiles : num_sms);
    ProfScope prof("tc_ffn_fused", s);
    DX_LAUNCH_PDL((ffn_tc_kernel), grid, FF_THREADS, FF_SMEM, s, p);
    DX_CHECK_LAUNCH();
    return 0;
}

}  // namespace dx
