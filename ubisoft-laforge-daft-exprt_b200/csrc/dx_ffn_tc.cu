// Fused tail of an FFT block for sm_100a (tcgen05 + TMEM + bulk-copy pipeline), one kernel for
//
//   [P0]  x1 = mask( LN1( attn · W_out + b_out + x_in ) )                       MultiHeadAttention out_proj + residual
//                                                                               + LayerNorm (model.py:187-191), pad mask (:256)
//   [FFN] y  = mask( FiLM( LN2( conv1d_{C→D,k3}( ReLU( conv1d_{D→C,k3}(x1) + b1 ) ) + b2 + x1 ) ) )
//                                                                               PositionWiseConvFF.forward (model.py:220-235), mask (:258)
//   [P3]  qkv' = y · W_qkv' + b_qkv'                                            in_proj of the NEXT block (model.py:165,182)
//
// evaluated on the packed row stream with the reference's "live padding" semantics (H1): the hidden activation of
// the first row after an utterance is live, everything beyond is true zero padding.  P0 and P3 are optional: without
// P0 the kernel reads x1 from global memory; without P3 it stops after the FFN (last block of a stack).  Fusing them
// removes two launches per block whose fixed cost (launch, prologue, tail: ~10 us each) exceeded their math (~1 us).
// x1 and the 1024-wide hidden activation never leave the SM.
//
// One CTA = 252 output rows = two "halves" of 128 hidden rows (126 outputs each, 1 halo row per side).
//   X   [16 chunk][256 row][8]  bf16  rows R0-2 .. R0+253: written by the P0 epilogue (or 16 bulk copies); also the residual
//   P0  A = attn rows R0-2 .. R0+253 (bulk-copied into the Hs region), D2[h] = A[128 h ..] · W_out, 8 MMAs per half
//   for hidden chunk c (128 channels, 8 chunks):
//     G1  D1[h] (128x128 fp32, TMEM) = sum_tap X_h[. + tap] · W1[c][tap]        24 MMAs (128x128x16) per half
//     E1  Hs[h][sub] = bf16(live ? ReLU(D1[h][:, 64 sub ..] + b1) : 0)          sub = 0,1; thread ↔ row
//     G2  D2[h] (128x128 fp32, TMEM) += sum_tap Hs[h][sub][. + tap] · W2[tap][c, sub]   12 MMAs per half and sub
//   E2  y = mask(FiLM(LN2(D2[h] + b2 + X_res)))  → global, and → Y[h] (the Hs region of the half) as the A operand of P3
//   P3  D1[h], D2[h], D1[h] = Y[h] · W_qkv'[:, 128 nt ..]   8 MMAs per 128-column tile; epilogue + bias → bf16 → global
// Issue order G1(c), G2(c-1, 0), G2(c-1, 1): E1(c) runs under G2(c-1).  Short streams use ONE half per CTA.
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
// Warps: 0 = bulk-copy producer, 1 = TMEM owner + issuer of half 0, 2 = issuer of half 1, 3 idle, 4-11 = epilogue of
// half 0, 12-19 = epilogue of half 1.  A warp may only touch TMEM lanes 32 (w % 4) .. +31, so every 32-row group
// is served by TWO warps that split the 128 channels (E1: one 64-channel K sub-block of G2 each; the LayerNorm
// epilogues exchange their partial sums through shared memory): the row epilogues were latency-bound with one
// thread per row (P0 8 k cycles, E2 14 k cycles of a 108 k-cycle tile).
//
// TMEM accumulator protocol (per half): D1 is "used" once per hidden chunk (+2 with P3), D2 once per segment (+1 with
// P0, +1 with P3); use number u (a running count kept identically by issuer and epilogue): issuer waits empty (parity
// (u & 1) ^ 1), issues, commits full; epilogue waits full (parity u & 1), loads the accumulator into registers, arrives
// on empty.
//
// Work split (stream-K over the hidden dimension) for SHORT streams.  One tile is ~50 us of serial work; a cfg2 symbol
// stream has 46 tiles for 148 SMs and a single utterance 3, so whole tiles per CTA leave most of the GPU idle and set the
// latency floor of a launch.  When whole tiles would occupy about a third of the grid or less (decided on the device), the
// unit of work becomes ONE HIDDEN CHUNK OF ONE 124-ROW TILE (8 units per tile): CTA i takes the contiguous unit range
// [i U / G', (i + 1) U / G'), G' = min(grid, U / min_units), and walks it tile by tile ("segments", chunks a..b of a tile).
// Every segment needs the tile's x1 (P0 is recomputed: 16 of a tile's 784 MMAs).  The segment that holds chunk 7 OWNS the
// tile: it adds the partial D2 sums of the tile's other segments (fp32, written to global memory = L2 by their CTAs, one
// slot per CTA, published with a release store of the launch's epoch into flags[cta]) to its own accumulator, in
// descending CTA order - a fixed order, so results are deterministic for a given (row count, grid) - and runs E2 / P3.  A
// CTA walks its range from the TOP: its only possible non-owner segment (the head of a tile whose tail belongs to CTA
// i + 1) comes first and waits for nothing, so an owner only ever waits for CTAs with a LOWER index that are already past
// their publishing point (no deadlock as long as CTAs are dispatched in index order, also next to another stream's
// kernels; every wait is bounded and traps).  Longer streams keep whole tiles: there each extra segment costs more (P0 +
// partial exchange, ~20 k cycles) than the chunks it takes over (~7 k each); measured, cfg2 frame stream: 0.489 -> 0.548 ms.
// Without a partial buffer (debug knob 6) units are always whole tiles.  The split is a function of (row count, grid)
// only, so the fused and the four-launch block (knob 9) split alike and stay bit-identical.
//
// Code size matters here: the once-per-tile epilogues are straight-line code that is fetched from L2 each time (ncu:
// stall_no_inst was their top stall reason).  The mbarrier time-out report is one out-of-line function (it was 40 KB of
// inlined printf set-up), and the E2 / P3 / dump epilogues keep the row in TMEM and loop over 16-column groups.
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_ptx.cuh"
#include "dx_ffn_epilogue.cuh"

namespace dx {

constexpr int FF_THREADS = 640;             // 20 warps: producer, 2 issuers, 1 idle, 16 epilogue
constexpr int FF_EPI_WARP0 = 4;
constexpr int FF_OUT = 126;                 // outputs per half (two-half tiles)
constexpr int FF_OUT1 = 124;                // outputs of a one-half tile: its 128 X rows are exactly one P0 M tile
constexpr int FF_TILE = 2 * FF_OUT;         // outputs per two-half CTA tile
constexpr int FF_NCH = 8;                   // hidden chunks of 128 channels
constexpr int FF_WSTAGES = 5;
constexpr int FF_PARTIAL_FLOATS = 2 * 2 * 64 * 128;     // one CTA's partial D2 of a tile: [half][part][col 64][row 128] fp32 = 128 KB
constexpr uint32_t FF_X_BYTES = 16 * 256 * 16;          // 65536
constexpr uint32_t FF_HS_ROWS = 136;
constexpr uint32_t FF_HS_BYTES = 8 * FF_HS_ROWS * 16;   // 17408 per (half, sub)
constexpr uint32_t FF_W_BYTES = 16384;
constexpr uint32_t FF_OFF_HS = FF_X_BYTES;
constexpr uint32_t FF_OFF_W = FF_OFF_HS + 4 * FF_HS_BYTES;
constexpr uint32_t FF_OFF_PAR = FF_OFF_W + FF_WSTAGES * FF_W_BYTES;
constexpr uint32_t FF_PAR_FLOATS = 1024 + 3 * 128 + 3 * 128 + 384 + 1024;   // b1 | b2, ln2 g, b | b_out, ln1 g, b | b_qkv' | LN exchange
constexpr uint32_t FF_OFF_BAR = FF_OFF_PAR + FF_PAR_FLOATS * 4;
constexpr uint32_t FF_NBARS = 2 * FF_WSTAGES + 3 + 2 * 9;
constexpr uint32_t FF_SMEM = FF_OFF_BAR + FF_NBARS * 8 + 16;
static_assert(FF_SMEM <= 232448, "fused FFN kernel: shared memory");
static_assert(2 * FF_HS_BYTES >= 32768 && 4 * FF_HS_BYTES >= FF_X_BYTES, "Hs region must hold the P0 A tile and the P3 A tiles");

struct FfnParams {
    const __nv_bfloat16* x;     // [16][rows_phys][8] x1 (FFN input and residual); unused with P0
    __nv_bfloat16* out;         // [16][rows_phys][8]
    int64_t rows_phys;
    const __nv_bfloat16 *w1, *w2;   // packed: W1 (taps 3, K 128, N 1024, nt 128, ks 64), W2 (taps 3, K 1024, N 128, nt 128, ks 64)
    const float *b1, *b2, *ln_g, *ln_b;
    const float* film;          // pre-offset to the block, or null
    int film_stride;
    // P0 (attn != null): attention output, block input (residual), packed W_out (K 128, N 128, nt 128, ks 64)
    const __nv_bfloat16 *attn, *x_in, *w_out;
    const float *b_out, *ln1_g, *ln1_b;
    // P3 (qkv_out != null): packed W_qkv of the next block (K 128, N 384, nt 128, ks 64), output [48][rows_phys][8]
    const __nv_bfloat16* w_qkv;
    const float* b_qkv;
    __nv_bfloat16* qkv_out;
    Rows R;
    // split over the hidden dimension (null: whole tiles per CTA)
    float* partial;             // [FF_MAX_CTAS][FF_PARTIAL_FLOATS]
    uint32_t* flags;            // [FF_MAX_CTAS]: flags[i] == epoch once CTA i's partial of THIS launch is complete
    uint32_t epoch;
    int min_units;              // fewest hidden chunks per CTA; 0 = 4 when there are >= 2 chunks per CTA, else 2
    int split_pct;              // split when whole tiles would occupy at most this percentage of the grid
    long long* trace;           // bring-up: clock64 timeline of one CTA (dx_set_trace_buffer) or null
    int trace_last;             // trace the last working CTA (always a tile owner) instead of CTA 0
    int exp;                    // timing experiments (debug knob 15; results are wrong): bit 0 no E2 global stores, bit 1 no P3
                                // global stores, bit 2 no P0 residual loads, bit 3 no FiLM loads, bit 4 no FiLM L1 prefetch
};

__device__ __forceinline__ uint32_t ff_ld_acquire(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void ff_st_release(uint32_t* p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// bounded: a protocol bug must trap, not hang the box
__device__ __forceinline__ void ff_wait_flag(const uint32_t* p, uint32_t epoch) {
    uint32_t spins = 0;
    while (ff_ld_acquire(p) != epoch) {
        __nanosleep(100);
        if (++spins > (1u << 24)) {
            printf("dx: ffn partial-sum flag timeout (block %d waits for epoch %u)\n", (int)blockIdx.x, epoch);
            __trap();
        }
    }
}

// Segment walk shared by the three roles: the CTA's unit range [w_lo, w_hi) from the top; per segment: tile, chunks
// a..b, t0 = first unit of the tile, owner = holds chunk 7.
#define FF_SEG_BEGIN                                                                                        \
    for (int64_t u_hi = w_hi; u_hi > w_lo;) {                                                               \
        const int tile = (int)((u_hi - 1) / FF_NCH);                                                        \
        const int64_t t0 = (int64_t)tile * FF_NCH;                                                          \
        const int b = (int)(u_hi - 1 - t0);                                                                 \
        const int a = (int)((w_lo > t0 ? w_lo : t0) - t0);                                                  \
        const bool owner = b == FF_NCH - 1;                                                                 \
        const bool seg_p3 = has_p3 && owner;
#define FF_SEG_END                                                                                          \
        u_hi = t0 + a;                                                                                      \
    }

// timeline slots (CTA 0, half 0 only): [0] kernel start, [1] X ready, then per chunk c (base 8 + 12 c):
//   issuer  +0 d1_empty ok, +1 G1 issued, +2 hs_full[0] ok, +3 G2 issued (both subs), +4 first w_full wait of G1 done
//   epilog  +6 d1_full ok, +7 TMEM loaded (sub 1), +8 hs_empty[1] ok, +9 Hs stored + fenced (sub 1)
// [2] d2_full ok (E2 start), [3] E2 done, [4] CTA done, [5] P0 A tile landed, [6] P0 accumulator ready, [7] P3 done
#define FF_TRACE(slot) do { if (tr) tr[(slot)] = clock64(); } while (0)

// Timing experiments (results are wrong; only the kernel time is meaningful): compile with -DFF_EXPERIMENT=<bits>,
// bit 0 = weight stages are not copied (arrive only), bit 1 = issue no MMAs, bit 2 = no Hs stores.  Compile time on
// purpose: the MMA issuers are bound by the latency of their own instruction stream (tools/ring_bench.cu), so even a
// uniform branch per MMA costs throughput.  profiles/r01d_ffn_timing_knobs.txt has the numbers.
#ifndef FF_EXPERIMENT
#define FF_EXPERIMENT 0
#endif

__global__ void __launch_bounds__(FF_THREADS, 1) ffn_tc_kernel(const FfnParams p) {
    extern __shared__ __align__(1024) uint8_t ff_smem[];
    uint8_t* sX = ff_smem;
    uint8_t* sHs = ff_smem + FF_OFF_HS;         // [half][sub]; also the P0 A tile and the P3 A tiles
    uint8_t* sW = ff_smem + FF_OFF_W;
    float* s_b1 = reinterpret_cast<float*>(ff_smem + FF_OFF_PAR);
    float* s_b2 = s_b1 + 1024;
    float* s_g = s_b2 + 128;
    float* s_be = s_g + 128;
    float* s_bo = s_be + 128;                   // P0: b_out, ln1 gamma, ln1 beta
    float* s_g1 = s_bo + 128;
    float* s_be1 = s_g1 + 128;
    float* s_bq = s_be1 + 128;                  // P3: b_qkv' [384]
    float* s_red = s_bq + 384;                  // LayerNorm partial sums [2 buf][2 half][2 part][128 row]
    uint64_t* bars = reinterpret_cast<uint64_t*>(ff_smem + FF_OFF_BAR);
    uint64_t* w_full = bars;                       // [S]   producer → both issuers
    uint64_t* w_empty = bars + FF_WSTAGES;         // [S]   one commit per active issuer
    uint64_t* x_full = bars + 2 * FF_WSTAGES;      // bulk copies of the X tile (or of the P0 A tile) have landed
    uint64_t* x_empty = x_full + 1;                // epilogue warps are done with the segment's shared memory
    uint64_t* x_ready = x_full + 2;                // P0: the epilogue warps have written the X tile
    uint64_t* hb = x_full + 3;                     // per half h (9 barriers each):
    // hb[9 h + 0] d1_full, +1 d1_empty, +2 hs_full[sub 0], +3 hs_full[sub 1], +4 hs_empty[0], +5 hs_empty[1], +6 d2_full, +7 d2_empty,
    // +8 y_full (P3: E2 has written the half's A tile)
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + FF_NBARS);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool has_p0 = p.attn != nullptr, has_p3 = p.qkv_out != nullptr;
    const int n_pre = has_p0 ? 2 : 0;              // weight stages every segment starts with (W_out): fetched before the PDL wait
    long long* tr = nullptr;
    const long long t_start = p.trace != nullptr ? clock64() : 0;

    // ---- prologue that does not depend on the previous kernel (PDL: overlaps its tail) ----
    for (int i = threadIdx.x; i < 1024; i += FF_THREADS) s_b1[i] = p.b1[i];
    if (threadIdx.x < 128) {
        s_b2[threadIdx.x] = p.b2[threadIdx.x];
        s_g[threadIdx.x] = p.ln_g[threadIdx.x];
        s_be[threadIdx.x] = p.ln_b[threadIdx.x];
        if (has_p0) {
            s_bo[threadIdx.x] = p.b_out[threadIdx.x];
            s_g1[threadIdx.x] = p.ln1_g[threadIdx.x];
            s_be1[threadIdx.x] = p.ln1_b[threadIdx.x];
        }
    }
    if (has_p3)
        for (int i = threadIdx.x; i < 384; i += FF_THREADS) s_bq[i] = p.b_qkv[i];
    // rows 128..135 of the Hs buffers are read (taps of the two discarded output rows) but never written by E1: they
    // only feed discarded rows, so any finite value will do (zero here; later whatever P0 / P3 left there)
    for (int i = threadIdx.x; i < 4 * 8 * 8; i += FF_THREADS) {
        const int q = i >> 6, kc = (i >> 3) & 7, rr = 128 + (i & 7);
        *reinterpret_cast<uint4*>(sHs + q * FF_HS_BYTES + kc * FF_HS_ROWS * 16 + rr * 16) = make_uint4(0u, 0u, 0u, 0u);
    }
    for (int i = threadIdx.x; i < 16 * 8; i += FF_THREADS)       // X rows 128..135: only written by two-half tiles
        *reinterpret_cast<uint4*>(sX + (i >> 3) * 4096 + (128 + (i & 7)) * 16) = make_uint4(0u, 0u, 0u, 0u);
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
    // the W_out stages are the same for every segment: fetch them now
    if (warp == 0 && ptx::elect_one()) {
        for (int s0 = 0; s0 < n_pre; ++s0) {
            if (FF_EXPERIMENT & 1) { ptx::mbar_arrive(&w_full[s0]); continue; }
            ptx::mbar_expect_tx(&w_full[s0], FF_W_BYTES);
            ptx::bulk_g2s(sW + s0 * FF_W_BYTES, reinterpret_cast<const uint8_t*>(p.w_out) + (size_t)s0 * FF_W_BYTES, FF_W_BYTES, &w_full[s0]);
        }
    }
    __syncwarp();
    ptx::griddep_launch();
    ptx::griddep_wait();
    // ---- from here on the previous kernel's output (row map, activations) may be read ----
    const int64_t total = p.R.row_start[p.R.B];
    const int64_t n252 = (total + FF_TILE - 1) / FF_TILE;
    // Short streams: 124-row tiles (ONE half per CTA) shorten the critical path when 252-row tiles would leave more than
    // half of the CTAs idle.  Decided on the device: the valid row count is only known there.
    const int halves_whole = n252 * 2 <= (int64_t)gridDim.x ? 1 : 2;
    const int64_t tiles_whole = halves_whole == 2 ? n252 : (total + FF_OUT1 - 1) / FF_OUT1;
    // Split the hidden dimension only when whole tiles would use about a third of the grid or less (measured on B200:
    // cfg2 symbol stream, 46 tiles: 0.180 -> 0.161 ms per stack; one utterance, 3 tiles: 0.343 -> 0.283 ms; cfg2 frame
    // stream, 109 tiles: 0.489 -> 0.548 ms - every extra segment repeats P0 and the partial-sum exchange, ~20 k cycles
    // against ~7 k per chunk)
    const bool split = p.partial != nullptr && tiles_whole * 100 <= (int64_t)p.split_pct * gridDim.x;
    const int halves = split ? 1 : halves_whole;
    const int out_rows = halves == 2 ? FF_OUT : FF_OUT1;               // outputs per half
    const int tile_rows = out_rows * halves;
    const int n_tiles = (int)((total + tile_rows - 1) / tile_rows);
    // this CTA's unit range (unit = one hidden chunk of one tile; whole tiles without the split)
    const int gran = split ? 1 : FF_NCH;
    const int64_t n_groups = (int64_t)n_tiles * (FF_NCH / gran);
    const int min_units = p.min_units > 0 ? p.min_units : (n_groups >= 2 * (int64_t)gridDim.x ? 4 : 2);
    int64_t ctas = split ? n_groups / min_units : n_groups;
    if (ctas > (int64_t)gridDim.x) ctas = gridDim.x;
    if (ctas < 1) ctas = 1;
    const bool has_work = (int64_t)blockIdx.x < ctas;
    const int64_t w_lo = has_work ? (int64_t)blockIdx.x * n_groups / ctas * gran : 0;
    const int64_t w_hi = has_work ? ((int64_t)blockIdx.x + 1) * n_groups / ctas * gran : 0;
    const uint32_t x_chunk_bytes = halves == 2 ? 4096u : (has_p0 ? 2048u : 2176u);      // 256 / 128 / 136 rows of one 8-channel chunk
    if (p.trace != nullptr && lane == 0 && (warp == 1 || warp == FF_EPI_WARP0) &&
        (int64_t)blockIdx.x == (p.trace_last ? (n_groups > 0 ? ctas - 1 : 0) : 0)) {
        tr = p.trace;
        if (warp == 1) tr[0] = t_start;
    }
    if (threadIdx.x == 0) {
        for (int i = 0; i < FF_WSTAGES; ++i) ptx::mbar_init(&w_empty[i], halves);
        ptx::mbar_init(x_full, 1);
        ptx::mbar_init(x_empty, 8 * halves);
        ptx::mbar_init(x_ready, 8 * halves);
        for (int h = 0; h < 2; ++h) {
            ptx::mbar_init(&hb[9 * h + 0], 1);          // d1_full
            ptx::mbar_init(&hb[9 * h + 1], 8);          // d1_empty
            ptx::mbar_init(&hb[9 * h + 2], 4);          // hs_full[0]
            ptx::mbar_init(&hb[9 * h + 3], 4);          // hs_full[1]
            ptx::mbar_init(&hb[9 * h + 4], 1);          // hs_empty[0]
            ptx::mbar_init(&hb[9 * h + 5], 1);          // hs_empty[1]
            ptx::mbar_init(&hb[9 * h + 6], 1);          // d2_full
            ptx::mbar_init(&hb[9 * h + 7], 8);          // d2_empty
            ptx::mbar_init(&hb[9 * h + 8], 8);          // y_full
        }
        ptx::fence_barrier_init();
    }
    __syncthreads();
    // TMEM columns: D1[h] at 128 h, D2[h] at 256 + 128 h
    const uint32_t tm_d1 = tmem_base, tm_d2 = tmem_base + 256;

    if (warp < FF_EPI_WARP0) {
    if (warp == 0) {
        // ===================== producer: X tile (or P0 A tile) + weight stages =====================
        if (ptx::elect_one()) {
            uint32_t wit = 0, seg_it = 0;
            const __nv_bfloat16* src = has_p0 ? p.attn : p.x;
            uint8_t* dst = has_p0 ? sHs : sX;
            const uint32_t dst_chunk = (has_p0 && halves == 1) ? 2048u : 4096u;
            // weight stages in MMA issue order: [W_out k 0,1]  G1(a)  { G1(c) G2(c-1) }  G2(b)  [W_qkv' (nt, k)]
#define FF_EMIT(srcptr)                                                                                     \
            do {                                                                                            \
                if ((int)wit >= n_pre) {                 /* the first n_pre stages were fetched in the prologue */ \
                    const int s_ = wit % FF_WSTAGES;                                                        \
                    ptx::mbar_wait(&w_empty[s_], ((wit / FF_WSTAGES) & 1) ^ 1);                             \
                    if (FF_EXPERIMENT & 1) ptx::mbar_arrive(&w_full[s_]);                                   \
                    else {                                                                                  \
                        ptx::mbar_expect_tx(&w_full[s_], FF_W_BYTES);                                       \
                        ptx::bulk_g2s(sW + s_ * FF_W_BYTES, (srcptr), FF_W_BYTES, &w_full[s_]);             \
                    }                                                                                       \
                }                                                                                           \
                ++wit;                                                                                      \
            } while (0)
            FF_SEG_BEGIN
                const int64_t R0 = (int64_t)tile * tile_rows;
                ptx::mbar_wait(x_empty, (seg_it & 1) ^ 1);
                ptx::mbar_expect_tx(x_full, 16 * x_chunk_bytes);
                for (int kc = 0; kc < 16; ++kc)
                    ptx::bulk_g2s(dst + kc * dst_chunk, src + ((int64_t)kc * p.rows_phys + DX_HALO + R0 - 2) * 8, x_chunk_bytes, x_full);
                for (int q = 0; q < n_pre; ++q) FF_EMIT(reinterpret_cast<const uint8_t*>(p.w_out) + (size_t)q * FF_W_BYTES);
                for (int c = a; c <= b + 1; ++c) {
                    if (c <= b)
                        for (int q = 0; q < 6; ++q) FF_EMIT(reinterpret_cast<const uint8_t*>(p.w1) + ((size_t)c * 6 + q) * FF_W_BYTES);
                    if (c > a)
                        for (int q = 0; q < 6; ++q) FF_EMIT(reinterpret_cast<const uint8_t*>(p.w2) + ((size_t)(c - 1) * 6 + q) * FF_W_BYTES);
                }
                if (seg_p3)
                    for (int q = 0; q < 6; ++q) FF_EMIT(reinterpret_cast<const uint8_t*>(p.w_qkv) + (size_t)q * FF_W_BYTES);
                ++seg_it;
            FF_SEG_END
#undef FF_EMIT
        }
    } else if (warp <= 2) {
        // ===================== MMA issuers: warp 1 ↔ half 0, warp 2 ↔ half 1 =====================
        const int h = warp - 1;
        if (w_hi <= w_lo && h == 0)      // no work: the prefetched weight stages must land before the CTA exits
            for (int s0 = 0; s0 < n_pre; ++s0) ptx::mbar_wait(&w_full[s0], 0);
        if (h < halves) {
            constexpr uint32_t idesc = ptx::make_idesc_bf16(128, 128);
            const uint32_t aX = ptx::smem_u32(sX) + FF_OUT * h * 16, aHs = ptx::smem_u32(sHs) + h * 2 * FF_HS_BYTES, aW = ptx::smem_u32(sW);
            const uint32_t aP0 = ptx::smem_u32(sHs) + 128 * h * 16;          // P0 A tile: rows 128 h .. of [16][256 | 128 rows][16 B]
            const uint32_t p0_lbo = halves == 2 ? 4096u : 2048u;
            const uint32_t td1 = tm_d1 + 128 * h, td2 = tm_d2 + 128 * h;
            uint64_t* d1_full = &hb[9 * h + 0];
            uint64_t* d1_empty = &hb[9 * h + 1];
            uint64_t* hs_full = &hb[9 * h + 2];
            uint64_t* hs_empty = &hb[9 * h + 4];
            uint64_t* d2_full = &hb[9 * h + 6];
            uint64_t* d2_empty = &hb[9 * h + 7];
            uint64_t* y_full = &hb[9 * h + 8];
            uint32_t seg_it = 0, w_phase = 0;
            uint32_t u1 = 0, u2 = 0, n_hs = 0, n_y = 0;  // running use counts: D1, D2, Hs buffers, P3 A tiles
            int s = 0;                                   // ring slot and phase of the next weight stage
            // one weight stage = 4 MMAs of this half: A at a_base + kk * a_step (K-major, LBO a_lbo), B = the stage
#define FF_STAGE(a_base, a_step, a_lbo, d_tmem, acc_first, ...)                                                         \
            do {                                                                                                        \
                ptx::mbar_wait(&w_full[s], w_phase);                                                                    \
                ptx::tc_fence_after();                                                                                  \
                if (ptx::elect_one()) {                                                                                 \
                    _Pragma("unroll") for (int kk = 0; kk < 4; ++kk) {                                                  \
                        const uint64_t da = ptx::make_desc((a_base) + kk * (a_step), (a_lbo), 128);                     \
                        const uint64_t db = ptx::make_desc(aW + s * FF_W_BYTES + kk * 2 * 2048, 2048, 128);             \
                        if (!(FF_EXPERIMENT & 2)) ptx::mma_bf16((d_tmem), da, db, idesc, (acc_first) || kk != 0);     \
                    }                                                                                                   \
                    ptx::mma_commit(&w_empty[s]);                                                                       \
                    __VA_ARGS__;                                                                                        \
                }                                                                                                       \
                __syncwarp();                                                                                           \
                if (++s == FF_WSTAGES) { s = 0; w_phase ^= 1; }                                                         \
            } while (0)
            FF_SEG_BEGIN
                ptx::mbar_wait(x_full, seg_it & 1);
                if (seg_it == 0 && has_p0) FF_TRACE(5);
                if (has_p0) {
                    // ---- P0: D2[h] = attn[128 h .. +127] · W_out   (stage = K half)
                    ptx::mbar_wait(d2_empty, (u2 & 1) ^ 1);
                    ptx::tc_fence_after();
                    for (int ks = 0; ks < 2; ++ks)
                        FF_STAGE(aP0 + ks * 8 * p0_lbo, 2 * p0_lbo, p0_lbo, td2, ks != 0, if (ks == 1) ptx::mma_commit(d2_full));
                    ++u2;
                    ptx::mbar_wait(x_ready, seg_it & 1);       // X tile written by the epilogue warps of both halves
                    ptx::tc_fence_after();
                }
                if (seg_it == 0) FF_TRACE(1);
                ptx::mbar_wait(d2_empty, (u2 & 1) ^ 1);
                ptx::tc_fence_after();
                for (int c = a; c <= b + 1; ++c) {
                    if (c <= b) {
                        // ---- G1(c): D1[h] = X_h (*) W1[c]   (stage = (K half, tap): 64 input channels x 128 hidden)
                        ptx::mbar_wait(d1_empty, (u1 & 1) ^ 1);
                        ptx::tc_fence_after();
                        if (seg_it == 0) FF_TRACE(8 + 12 * c + 0);
#pragma unroll
                        for (int st = 0; st < 6; ++st) {
                            const int kh = st / 3, tap = st % 3;
                            FF_STAGE(aX + kh * 8 * 4096 + tap * 16, 2 * 4096, 4096, td1, st != 0, if (st == 5) ptx::mma_commit(d1_full));
                        }
                        ++u1;
                        if (seg_it == 0) FF_TRACE(8 + 12 * c + 1);
                    }
                    if (c > a) {
                        // ---- G2(c-1, sub): D2[h] += Hs[h][sub] (*) W2[c-1, sub]   (stage = (sub, tap): 64 hidden x 128 out)
                        const int cc = c - 1;
                        const bool first = cc == a, last = cc == b;
#pragma unroll
                        for (int st = 0; st < 6; ++st) {
                            const int sub = st / 3, tap = st % 3;
                            if (tap == 0) {
                                ptx::mbar_wait(&hs_full[sub], n_hs & 1);
                                ptx::tc_fence_after();
                                if (seg_it == 0 && sub == 0) FF_TRACE(8 + 12 * cc + 2);
                            }
                            FF_STAGE(aHs + sub * FF_HS_BYTES + tap * 16, 2 * FF_HS_ROWS * 16, FF_HS_ROWS * 16, td2, !first || st != 0,
                                     if (tap == 2) {
                                         ptx::mma_commit(&hs_empty[sub]);
                                         if (last && sub == 1) ptx::mma_commit(d2_full);
                                     });
                        }
                        ++n_hs;
                        if (seg_it == 0) FF_TRACE(8 + 12 * cc + 3);
                    }
                }
                ++u2;
                if (seg_p3) {
                    // ---- P3: next block's QKV projection of this half's outputs: tiles 0, 2 → D1[h], tile 1 → D2[h]
                    ptx::mbar_wait(y_full, n_y & 1);
                    ++n_y;
                    ptx::tc_fence_after();
                    for (int nt = 0; nt < 3; ++nt) {
                        const bool on_d1 = nt != 1;
                        ptx::mbar_wait(on_d1 ? d1_empty : d2_empty, ((on_d1 ? u1 : u2) & 1) ^ 1);
                        ptx::tc_fence_after();
                        for (int ks = 0; ks < 2; ++ks)
                            FF_STAGE(aHs + ks * 8 * 2048, 2 * 2048, 2048, on_d1 ? td1 : td2, ks != 0,
                                     if (ks == 1) ptx::mma_commit(on_d1 ? d1_full : d2_full));
                        if (on_d1) ++u1; else ++u2;
                    }
                }
                ++seg_it;
            FF_SEG_END
#undef FF_STAGE
        }
    }
    } else {
    if (((warp - FF_EPI_WARP0) >> 3) < halves) {
        // ===================== epilogue: 16 warps; (half, channel part, lane quadrant); thread ↔ (row, 64 channels) ==========
        const int e = warp - FF_EPI_WARP0;
        const int h = e >> 3;
        const int part = (e >> 2) & 1;                               // channels 64 part .. +63 of the row
        const int lane_grp = warp & 3;
        const int j = lane_grp * 32 + lane;                         // row inside the half
        const uint32_t lane_off = (uint32_t)(lane_grp * 32) << 16;
        const uint32_t td1 = tm_d1 + 128 * h + lane_off + 64 * part, td2 = tm_d2 + 128 * h + lane_off + 64 * part;
        uint64_t* d1_full = &hb[9 * h + 0];
        uint64_t* d1_empty = &hb[9 * h + 1];
        uint64_t* hs_full = &hb[9 * h + 2];
        uint64_t* hs_empty = &hb[9 * h + 4];
        uint64_t* d2_full = &hb[9 * h + 6];
        uint64_t* d2_empty = &hb[9 * h + 7];
        uint64_t* y_full = &hb[9 * h + 8];
        uint8_t* sY = sHs + h * 2 * FF_HS_BYTES;                     // P3 A tile of this half: [16 chunk][128 row][16 B]
        // LayerNorm statistics exchange between the two threads of a row: red[buf][h][part][row]
        float* my_red = s_red + (h * 2 + part) * 128 + j;
        const float* peer_red = s_red + (h * 2 + (part ^ 1)) * 128 + j;
        const int pair_bar = 1 + h * 4 + lane_grp;                   // named barrier of the two warps that share these 32 rows
        const int64_t part_off = (int64_t)((h * 2 + part) * 64) * 128 + j;     // this thread's column 0 inside a partial slot
        if (h != 0 || part != 0) tr = nullptr;
        uint32_t seg_it = 0, u1 = 0, u2 = 0, n_hs = 0;
        // this part's 64 accumulator columns → registers, then hand the accumulator back to the issuer
#define FF_LOAD64(v, taddr, full_bar, empty_bar, use)                                                       \
        do {                                                                                                \
            ptx::mbar_wait((full_bar), (use) & 1);                                                          \
            ptx::tc_fence_after();                                                                          \
            ptx::tmem_ld32((taddr), (v));                                                                   \
            ptx::tmem_ld32((taddr) + 32, (v) + 32);                                                         \
            ptx::tmem_ld_wait();                                                                            \
            ptx::tc_fence_before();                                                                         \
            __syncwarp();                                                                                   \
            if (lane == 0) ptx::mbar_arrive(empty_bar);                                                     \
        } while (0)
        FF_SEG_BEGIN
            const int64_t R0 = (int64_t)tile * tile_rows;
            const int64_t sh = R0 + FF_OUT * h - 1 + j;             // stream row of hidden row j
            const bool hid_live = sh >= 0 && sh < total && row_live(p.R, sh, 1);
            const int64_t so = R0 + FF_OUT * h + j;                 // stream row of output row j
            const bool out_write = owner && j < out_rows && so < total;
            const bool out_live = out_write && row_live(p.R, so, 0);
            const int row_b = out_write ? p.R.row_b[so] : -1;
            const int b_idx = max(row_b, 0);
            if (has_p0) {
                // ---- P0 epilogue: X row 128 h + j = mask(LN1(D2 + b_out + x_in)), stream row R0 - 2 + 128 h + j
                const int64_t sx = R0 - 2 + 128 * h + j;
                const bool x_live = sx >= 0 && sx < total && row_live(p.R, sx, 0);
                float v[64];
                FF_LOAD64(v, td2, d2_full, d2_empty, u2);
                ++u2;
                if (seg_it == 0) FF_TRACE(6);
                float psum = 0.f;
                if (x_live && !(p.exp & 4))
                    psum = ffn_row_add(v, s_bo + 64 * part, reinterpret_cast<const uint8_t*>(p.x_in + ((int64_t)8 * part * p.rows_phys + DX_HALO + sx) * 8),
                                       p.rows_phys * 16);
                my_red[0] = psum;
                ptx::named_bar_sync(pair_bar, 64);
                const float o_ = peer_red[0];
                const float mean = (part == 0 ? psum + o_ : o_ + psum) * (1.f / 128.f);
                const float sq_ = ffn_row_sq(v, mean);
                my_red[512] = sq_;
                ptx::named_bar_sync(pair_bar, 64);
                const float p_ = peer_red[512];
                const float rstd = rsqrtf((part == 0 ? sq_ + p_ : p_ + sq_) * (1.f / 128.f) + 1e-5f);
                ffn_row_store(v, mean, rstd, s_g1 + 64 * part, s_be1 + 64 * part, nullptr, x_live, nullptr, 0,
                              sX + 8 * part * 4096 + (128 * h + j) * 16, 4096);
                ptx::fence_proxy_async();        // X writes → visible to the tensor core's async proxy
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(x_ready);
            } else {
                ptx::mbar_wait(x_full, seg_it & 1);                  // residual rows are read from sX in E2
            }
            if (owner && p.film != nullptr && !(p.exp & 16)) {
                // E2 reads this row's FiLM gamma / beta (2 x 256 B) in pass C, group by group: bring the lines into L1 now
                const float* fp = p.film + (int64_t)b_idx * p.film_stride + 64 * part;
                asm volatile("prefetch.global.L1 [%0];" ::"l"(fp));
                asm volatile("prefetch.global.L1 [%0];" ::"l"(fp + 32));
                asm volatile("prefetch.global.L1 [%0];" ::"l"(fp + 128));
                asm volatile("prefetch.global.L1 [%0];" ::"l"(fp + 160));
            }
            for (int c = a; c <= b; ++c, ++u1, ++n_hs) {
                // ---- E1(c): this part's 64 hidden channels (= G2's K sub-block `part`)
                float v[64];
                if (seg_it == 0) FF_TRACE(8 + 12 * c + 6);
                FF_LOAD64(v, td1, d1_full, d1_empty, u1);            // D1 is in registers: G1(c+1) may overwrite it
                if (seg_it == 0) FF_TRACE(8 + 12 * c + 7);
                ptx::mbar_wait(&hs_empty[part], (n_hs & 1) ^ 1);     // G2 of the previous chunk has finished reading this buffer
                if (seg_it == 0) FF_TRACE(8 + 12 * c + 8);
                if (!(FF_EXPERIMENT & 4))
                    ffn_e1_store(v, s_b1 + c * 128 + part * 64, hid_live, sHs + (h * 2 + part) * FF_HS_BYTES + j * 16, FF_HS_ROWS * 16);
                ptx::fence_proxy_async();        // Hs writes → visible to the tensor core's async proxy
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&hs_full[part]);
                if (seg_it == 0) FF_TRACE(8 + 12 * c + 9);
            }
            {   // ---- E2: the segment's D2 (all its G2 MMAs of the half have completed: the Hs buffers are free)
                ptx::mbar_wait(d2_full, u2 & 1);
                ++u2;
                ptx::tc_fence_after();
                if (seg_it == 0) FF_TRACE(2);
                if (!owner) {
                    // partial sum of chunks a..b → this CTA's slot; published once every epilogue thread has stored
                    ffn_dump_rolled(td2, p.partial + (int64_t)blockIdx.x * FF_PARTIAL_FLOATS + part_off);
                    __threadfence();
                    ptx::named_bar_sync(9, 256 * halves);
                    if (e == 0 && lane == 0) {
                        __threadfence();
                        ff_st_release(p.flags + blockIdx.x, p.epoch);
                    }
                } else {
                    // y = mask(FiLM(LN2(D2 (+ the tile's other segments) + b2 + x1))): the row stays in TMEM (rolled, see
                    // dx_ffn_epilogue.cuh: the unrolled form took 11-18 k cycles here, mostly instruction fetch)
                    FfnRowEpi E;
                    E.taddr = td2;
                    E.part_base = p.partial + part_off; E.part_slot = FF_PARTIAL_FLOATS; E.k_hi = (int)blockIdx.x - 1; E.n_contrib = 0;
                    if (a > 0) {
                        // CTAs i-1, i-2, .. down to the one that holds chunk 0 (fixed order)
                        int n = 0;
                        for (int64_t k = (int64_t)blockIdx.x - 1; k >= 0 && (k + 1) * n_groups / ctas > t0; --k, ++n)
                            if (lane == 0) ff_wait_flag(p.flags + k, p.epoch);
                        __syncwarp();
                        E.n_contrib = n;
                    }
                    E.bias = s_b2 + 64 * part;
                    E.res = sX + 8 * part * 4096 + (FF_OUT * h + j + 2) * 16;
                    E.res_stride = 4096;
                    E.g = s_g + 64 * part; E.be = s_be + 64 * part;
                    E.film = (p.film && !(p.exp & 8)) ? p.film + (int64_t)b_idx * p.film_stride + 64 * part : nullptr;
                    E.live = out_live;
                    E.out_g = (out_write && !(p.exp & 1)) ? p.out + ((int64_t)8 * part * p.rows_phys + DX_HALO + so) * 8 : nullptr;
                    E.out_chunk = p.rows_phys * 8;
                    E.out_s = seg_p3 ? sY + 8 * part * 2048 + j * 16 : nullptr; E.smem_chunk = 2048;
                    ffn_ln_rolled(E, my_red, peer_red, pair_bar, part);
                }
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(d2_empty);           // the accumulator goes back to the issuer
                if (seg_it == 0) FF_TRACE(3);
            }
            if (seg_p3) {
                ptx::fence_proxy_async();            // Y writes → async proxy
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(y_full);
                const bool in_utt = row_b >= 0;
                for (int nt = 0; nt < 3; ++nt) {
                    const bool on_d1 = nt != 1;
                    ptx::mbar_wait(on_d1 ? d1_full : d2_full, (on_d1 ? u1 : u2) & 1);
                    if (on_d1) ++u1; else ++u2;
                    ptx::tc_fence_after();
                    ffn_qkv_rolled(on_d1 ? td1 : td2, s_bq + nt * 128 + 64 * part, in_utt,
                                   p.qkv_out + ((int64_t)(nt * 16 + 8 * part) * p.rows_phys + DX_HALO + so) * 8, p.rows_phys * 8,
                                   out_write && !(p.exp & 2));
                    ptx::tc_fence_before();
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(on_d1 ? d1_empty : d2_empty);
                }
                if (seg_it == 0) FF_TRACE(7);
            }
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(x_empty);
            ++seg_it;
        FF_SEG_END
#undef FF_LOAD64
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
             int film_stride, const Rows& R, cudaStream_t s, const FfnFuse* fuse, const FfnSplit* split) {
    DX_REQUIRE(bw.w_ffn1_bf && bw.w_ffn2_bf, "ffn_bf16: fused-FFN packed weights missing");
    int num_sms = 0;
    DX_TRY(device_sm_count(&num_sms));
    static SmemConfigured configured;
    DX_TRY(ensure_dyn_smem(ffn_tc_kernel, (size_t)FF_SMEM, configured));
    FfnParams p;
    memset(&p, 0, sizeof(p));
    p.x = x; p.out = out; p.rows_phys = rows_phys;
    p.w1 = reinterpret_cast<const __nv_bfloat16*>(bw.w_ffn1_bf);
    p.w2 = reinterpret_cast<const __nv_bfloat16*>(bw.w_ffn2_bf);
    p.b1 = bw.b_c1; p.b2 = bw.b_c2; p.ln_g = bw.ln2_g; p.ln_b = bw.ln2_b;
    p.film = film; p.film_stride = film_stride; p.R = R;
    if (fuse != nullptr && fuse->attn != nullptr) {
        DX_REQUIRE(fuse->x_in && bw.w_out_bf && bw.b_out && bw.ln1_g && bw.ln1_b, "ffn_bf16: fused out-proj needs x_in, W_out, b_out, LN1");
        DX_REQUIRE(fuse->x_in != out && fuse->attn != out, "ffn_bf16: the output must not alias the block input or the attention output");
        p.attn = fuse->attn; p.x_in = fuse->x_in;
        p.w_out = reinterpret_cast<const __nv_bfloat16*>(bw.w_out_bf);
        p.b_out = bw.b_out; p.ln1_g = bw.ln1_g; p.ln1_b = bw.ln1_b;
    } else {
        DX_REQUIRE(x != nullptr && x != out, "ffn_bf16: in-place operation is not supported (rows of neighbouring tiles are read as halos)");
    }
    if (fuse != nullptr && fuse->qkv_out != nullptr) {
        DX_REQUIRE(fuse->next && fuse->next->w_qkv_bf && fuse->next->b_qkv, "ffn_bf16: fused QKV projection needs the next block's weights");
        p.w_qkv = reinterpret_cast<const __nv_bfloat16*>(fuse->next->w_qkv_bf);
        p.b_qkv = fuse->next->b_qkv;
        p.qkv_out = fuse->qkv_out;
    }
    p.trace = reinterpret_cast<long long*>(trace_buffer(256 * sizeof(long long)));
    if (p.trace != nullptr && debug_knob(10) % 100 > 0) {       // knob 10 = n (+ 100: trace the last working CTA): trace only the n-th launch of a step (1..12)
        static int n_traced = 0;
        if ((n_traced++ % 12) + 1 != debug_knob(10) % 100) p.trace = nullptr;
    }
    p.trace_last = debug_knob(10) >= 100;
    p.exp = debug_knob(15);
    int64_t tiles = (R.cap + FF_OUT1 - 1) / FF_OUT1;     // enough CTAs for the one-half-per-CTA mode of short streams
    if (split != nullptr && split->partial != nullptr && split->flags != nullptr && !debug_knob(6)) {
        // split over the hidden dimension: every launch gets its own epoch (flags never have to be cleared between
        // launches; the caller zeroes them once per pass against stale memory)
        static std::atomic<uint32_t> g_epoch{0};
        uint32_t ep = g_epoch.fetch_add(1, std::memory_order_relaxed) + 1;
        if (ep == 0) ep = g_epoch.fetch_add(1, std::memory_order_relaxed) + 1;
        p.partial = split->partial; p.flags = split->flags; p.epoch = ep;
        p.min_units = debug_knob(7) > 0 ? debug_knob(7) : 0;
        p.split_pct = debug_knob(8) > 0 ? debug_knob(8) : 35;
        tiles *= FF_NCH;
    }
    int grid = (int)(tiles < num_sms ? tiles : num_sms);
    if (grid > FF_MAX_CTAS) grid = FF_MAX_CTAS;
    ProfScope prof(fuse != nullptr && fuse->tag != nullptr ? fuse->tag : "tc_ffn_fused", s);
    DX_LAUNCH_PDL((ffn_tc_kernel), grid, FF_THREADS, FF_SMEM, s, p);
    DX_CHECK_LAUNCH();
    return 0;
}

int64_t ffn_partial_floats() { return (int64_t)FF_MAX_CTAS * FF_PARTIAL_FLOATS; }

}  // namespace dx
