// Tensor-core conv-as-GEMM for sm_100a: tcgen05.mma (bf16 x bf16 → fp32 in TMEM), operands staged into shared
// memory by the bulk-copy (TMA) engine, warp-specialised persistent kernel.
//
//   out[r][n] = epi( bias[n] + sum_{tap} sum_k A[r + tap - taps/2][k] * W[tap][k][n] )
//
// Reference ops replaced: nn.Conv1d(k=3,"same") inside PositionWiseConvFF (model.py:206-217) and the accent
// pre-net (model.py:649-671); nn.Linear for in_proj / out_proj of nn.MultiheadAttention (model.py:165) and the
// mel projection (model.py:546,561); with their LayerNorm / residual / FiLM / pad-mask epilogues fused
// (model.py:191,226-233,256,258).
//
// Layouts (DESIGN.md §"data layout"):
//   A   bf16 "channel-interleaved" [K/8][rows_phys][8]: for a fixed 8-channel chunk, rows are contiguous 16-byte
//       units.  A 128-row tile (+1 halo row each side) of one chunk is ONE contiguous 2080-byte run, so a stage
//       is filled by KS/8 bulk copies, and in shared memory it is exactly the no-swizzle K-major UMMA canonical
//       layout with SBO = 128 B (rows 16 B apart): a conv tap is a +16 B shift of the descriptor start address,
//       so the 3 taps reuse one staged tile.
//   W   bf16 packed [n_tile][k_stage][tap][KS/8][NT][8]: one contiguous block per stage → one bulk copy.
//   D   fp32 accumulators in TMEM, 2 x NT columns (double buffered: epilogue of tile i overlaps MMAs of i+1).
// Warp roles (192 threads): warp 0 = copy producer, warp 1 = TMEM owner + MMA issuer, warps 2-5 = epilogue
// (warp w reads TMEM lanes 32*(w%4)..+31: thread ↔ output row).
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_ptx.cuh"
#include "dx_tc_epilogue.cuh"

namespace dx {

constexpr int TC_BM = 128;          // rows per tile (UMMA M)
constexpr int TC_RA = 136;          // staged A rows per chunk (130 used; stride kept 128-B aligned)
constexpr int TC_THREADS = 192;
constexpr int TC_MAX_N = 1024;      // widest output (conv_channels)
constexpr size_t TC_PAR_BYTES = (TC_MAX_N + 256) * sizeof(float);

struct TcParams {
    TcGemmArgs g;
    Rows R;                 // R.row_b == nullptr → plain mode (all rows < total_val live)
    int64_t total_val;      // plain mode
    int n_tiles, nks, kc;   // N tiles, K stages, chunks per stage
    int nstages;
    uint32_t a_stage_bytes, w_stage_bytes;
    uint32_t a_lbo, a_sbo, b_lbo, b_sbo;   // descriptor strides (host-computed; debug knob can swap them)
    uint32_t par_off;                      // byte offset of the staged epilogue parameters in dynamic smem
};

__device__ __forceinline__ int64_t tc_total_rows(const TcParams& p) {
    return p.R.row_b ? (int64_t)p.R.row_start[p.R.B] : p.total_val;
}

template <int NT, int EPI>
__global__ void __launch_bounds__(TC_THREADS, 1) tc_gemm_kernel(const TcParams p) {
    extern __shared__ __align__(1024) uint8_t tc_smem[];
    // carve: [stages x (A | W)] [barriers] [tmem ptr]
    const uint32_t stage_bytes = p.a_stage_bytes + p.w_stage_bytes;
    uint8_t* stage_base = tc_smem;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(tc_smem + (size_t)p.nstages * stage_bytes);
    uint64_t* empty_bar = full_bar + p.nstages;
    uint64_t* tfull_bar = empty_bar + p.nstages;   // [2]
    uint64_t* tempty_bar = tfull_bar + 2;          // [2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty_bar + 2);
    // epilogue parameters staged once per CTA (scalar global loads per element made the epilogue the bottleneck):
    // bias zero-padded to n_tiles * NT (<= TC_MAX_N) floats, then LayerNorm gamma / beta (128 each)
    float* s_bias = reinterpret_cast<float*>(tc_smem + p.par_off);
    float* s_lng = s_bias + TC_MAX_N;
    float* s_lnb = s_lng + 128;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < TC_MAX_N; i += TC_THREADS) s_bias[i] = (p.g.bias != nullptr && i < p.g.N) ? p.g.bias[i] : 0.f;
    if (EPI == TC_EPI_LN && threadIdx.x < 128) {
        s_lng[threadIdx.x] = p.g.ln_g[threadIdx.x];
        s_lnb[threadIdx.x] = p.g.ln_b[threadIdx.x];
    }
    if (threadIdx.x == 0) {
        for (int i = 0; i < p.nstages; ++i) {
            ptx::mbar_init(&full_bar[i], 1);
            ptx::mbar_init(&empty_bar[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            ptx::mbar_init(&tfull_bar[i], 1);
            ptx::mbar_init(&tempty_bar[i], 4);
        }
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<2 * NT>(tmem_ptr);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    // PDL: everything above (parameter staging, barriers, TMEM) overlapped the previous kernel's tail
    ptx::griddep_launch();
    ptx::griddep_wait();
    const int64_t total = tc_total_rows(p);
    const int m_tiles = (int)((total + TC_BM - 1) / TC_BM);
    const int64_t n_items = (int64_t)m_tiles * p.n_tiles;

    if (warp == 0) {
        // ===================== producer =====================
        if (ptx::elect_one()) {
            uint32_t it = 0;
            for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x) {
                const int m_tile = (int)(item / p.n_tiles), n_tile = (int)(item % p.n_tiles);
                const int64_t r0 = (int64_t)m_tile * TC_BM;
                for (int ksg = 0; ksg < p.nks; ++ksg, ++it) {
                    const int s = it % p.nstages;
                    const uint32_t ph = (it / p.nstages) & 1;
                    ptx::mbar_wait(&empty_bar[s], ph ^ 1);
                    uint8_t* sa = stage_base + (size_t)s * stage_bytes;
                    uint8_t* sw = sa + p.a_stage_bytes;
                    ptx::mbar_expect_tx(&full_bar[s], (uint32_t)p.kc * (TC_BM + 2) * 16 + p.w_stage_bytes);
                    for (int c = 0; c < p.kc; ++c) {
                        const __nv_bfloat16* src =
                            p.g.a + ((int64_t)(ksg * p.kc + c) * p.g.a_rows_phys + DX_HALO + r0 - 1) * 8;
                        ptx::bulk_g2s(sa + (size_t)c * TC_RA * 16, src, (TC_BM + 2) * 16, &full_bar[s]);
                    }
                    const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(p.g.w) +
                                          ((int64_t)n_tile * p.nks + ksg) * (int64_t)p.w_stage_bytes;
                    ptx::bulk_g2s(sw, wsrc, p.w_stage_bytes, &full_bar[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        constexpr uint32_t idesc = ptx::make_idesc_bf16(TC_BM, NT);
        const int half = p.g.taps / 2;
        uint32_t it = 0, tile_it = 0;
        for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x, ++tile_it) {
            const int acc = tile_it & 1;
            const uint32_t acc_ph = (tile_it >> 1) & 1;
            ptx::mbar_wait(&tempty_bar[acc], acc_ph ^ 1);
            ptx::tc_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)acc * NT;
            for (int ksg = 0; ksg < p.nks; ++ksg, ++it) {
                const int s = it % p.nstages;
                const uint32_t ph = (it / p.nstages) & 1;
                ptx::mbar_wait(&full_bar[s], ph);
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
                    const uint32_t sa = ptx::smem_u32(stage_base + (size_t)s * stage_bytes);
                    const uint32_t sw = sa + p.a_stage_bytes;
                    for (int tap = 0; tap < p.g.taps; ++tap) {
                        for (int kk = 0; kk < p.kc / 2; ++kk) {
                            const uint64_t da = ptx::make_desc(sa + (uint32_t)(tap - half + 1) * 16 + (uint32_t)kk * 2 * TC_RA * 16,
                                                               p.a_lbo, p.a_sbo);
                            const uint64_t db = ptx::make_desc(sw + (uint32_t)(tap * p.kc + kk * 2) * NT * 16, p.b_lbo, p.b_sbo);
                            ptx::mma_bf16(d_tmem, da, db, idesc, (ksg | tap | kk) != 0);
                        }
                    }
                    ptx::mma_commit(&empty_bar[s]);
                    if (ksg == p.nks - 1) ptx::mma_commit(&tfull_bar[acc]);
                }
                __syncwarp();
            }
        }
    } else {
        // ===================== epilogue (warps 2..5) =====================
        const int lane_grp = warp & 3;                 // TMEM lanes 32*lane_grp .. +31
        uint32_t tile_it = 0;
        for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x, ++tile_it) {
            const int m_tile = (int)(item / p.n_tiles), n_tile = (int)(item % p.n_tiles);
            const int acc = tile_it & 1;
            const uint32_t acc_ph = (tile_it >> 1) & 1;
            const int64_t r = (int64_t)m_tile * TC_BM + lane_grp * 32 + lane;
            bool live = r < total;
            int b_idx = 0;
            if (live && p.R.row_b) {
                b_idx = p.R.row_b[r];
                if (p.g.live_ext >= 0) live = row_live(p.R, r, p.g.live_ext);
                else live = b_idx >= 0;
                if (b_idx < 0) b_idx = 0;
            }
            ptx::mbar_wait(&tfull_bar[acc], acc_ph);
            ptx::tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(lane_grp * 32) << 16) + (uint32_t)acc * NT;

            tc_epilogue_tile<NT, EPI>(p.g, p.R, total, r, n_tile, taddr, &tempty_bar[acc], lane, live, b_idx, s_bias, s_lng, s_lnb);
        }
    }
    // teardown
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<2 * NT>(tmem_base);
    }
}

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------

template <int NT, int EPI>
static int tc_launch(const TcParams& p, size_t smem, int grid, cudaStream_t s) {
    static SmemConfigured configured;
    DX_TRY(ensure_dyn_smem(tc_gemm_kernel<NT, EPI>, smem, configured));
    ProfScope prof(p.g.tag ? p.g.tag : "tc_gemm", s);
    DX_LAUNCH_PDL((tc_gemm_kernel<NT, EPI>), grid, TC_THREADS, smem, s, p);
    DX_CHECK_LAUNCH();
    return 0;
}

static int tc_gemm_impl(const TcGemmArgs& g, const Rows* R, int64_t M, cudaStream_t s) {
    DX_REQUIRE(g.nt == 128 || g.nt == 256, "tc_gemm: nt=%d (128 or 256)", g.nt);
    DX_REQUIRE(g.ks % 16 == 0 && g.K % g.ks == 0, "tc_gemm: K=%d must be a multiple of ks=%d (multiple of 16)", g.K, g.ks);
    DX_REQUIRE(g.taps == 1 || g.taps == 3, "tc_gemm: taps=%d", g.taps);
    int g_num_sms = 0;
    DX_TRY(device_sm_count(&g_num_sms));
    TcParams p;
    p.g = g;
    if (R) {
        p.R = *R;
        p.total_val = 0;
    } else {
        p.R = Rows{nullptr, nullptr, nullptr, nullptr, 0, 0, round_up(M, TC_BM)};
        p.total_val = M;
    }
    p.n_tiles = (g.N + g.nt - 1) / g.nt;
    p.nks = g.K / g.ks;
    p.kc = g.ks / 8;
    p.a_stage_bytes = (uint32_t)p.kc * TC_RA * 16;
    p.w_stage_bytes = (uint32_t)g.taps * p.kc * g.nt * 16;
    p.a_lbo = TC_RA * 16; p.a_sbo = 128;
    p.b_lbo = (uint32_t)g.nt * 16; p.b_sbo = 128;
    if (debug_knob(0)) {   // swap the meaning of LBO / SBO (bring-up aid for the descriptor convention)
        uint32_t t = p.a_lbo; p.a_lbo = p.a_sbo; p.a_sbo = t;
        t = p.b_lbo; p.b_lbo = p.b_sbo; p.b_sbo = t;
    }
    const size_t stage = (size_t)p.a_stage_bytes + p.w_stage_bytes;
    DX_REQUIRE(p.n_tiles * g.nt <= TC_MAX_N, "tc_gemm: N=%d wider than %d", g.N, TC_MAX_N);
    DX_REQUIRE(g.epilogue != TC_EPI_LN || (g.bias && g.ln_g && g.ln_b), "tc_gemm: LN epilogue needs bias, gamma, beta");
    DX_REQUIRE(g.row_stats == nullptr || p.n_tiles == TC_STAT_TILES, "tc_gemm: row_stats needs exactly %d N tiles (got %d)", TC_STAT_TILES, p.n_tiles);
    const size_t budget = 220 * 1024;
    int nst = (int)(budget / stage);
    if (nst > 6) nst = 6;
    DX_REQUIRE(nst >= 2, "tc_gemm: stage of %zu bytes does not allow 2 stages (taps=%d ks=%d nt=%d)", stage, g.taps, g.ks, g.nt);
    p.nstages = nst;
    p.par_off = (uint32_t)round_up((int64_t)((size_t)nst * stage + (2 * nst + 4) * sizeof(uint64_t) + 16), 16);
    const size_t smem = p.par_off + TC_PAR_BYTES;
    const int64_t max_rows = R ? R->cap : M;
    int64_t items = ((max_rows + TC_BM - 1) / TC_BM) * p.n_tiles;
    int grid = (int)(items < g_num_sms ? items : g_num_sms);
    if (grid < 1) grid = 1;
    if (g.epilogue == TC_EPI_LN) {
        DX_REQUIRE(g.nt == 128 && g.N == 128, "tc_gemm: LN epilogue needs N == nt == 128");
        return tc_launch<128, TC_EPI_LN>(p, smem, grid, s);
    }
    if (g.epilogue == TC_EPI_F32)
        return g.nt == 128 ? tc_launch<128, TC_EPI_F32>(p, smem, grid, s) : tc_launch<256, TC_EPI_F32>(p, smem, grid, s);
    if (g.epilogue == TC_EPI_BF16)
        return g.nt == 128 ? tc_launch<128, TC_EPI_BF16>(p, smem, grid, s) : tc_launch<256, TC_EPI_BF16>(p, smem, grid, s);
    DX_REQUIRE(false, "tc_gemm: unknown epilogue %d", g.epilogue);
    return 0;
}

int tc_gemm(const TcGemmArgs& g, const Rows& R, cudaStream_t s) { return tc_gemm_impl(g, &R, 0, s); }
int tc_gemm_plain(const TcGemmArgs& g, int64_t M, cudaStream_t s) { return tc_gemm_impl(g, nullptr, M, s); }

// ---------------------------------------------------------------------------------------------------------
// weight packing: fp32 [taps][K][N] → bf16 [n_tile][k_stage][tap][ks/8][nt][8], N zero-padded to n_tiles*nt
// ---------------------------------------------------------------------------------------------------------
int64_t packed_weight_elems(int taps, int K, int N, int nt, int ks) {
    if (nt <= 0 || ks <= 0 || K % ks != 0) return -1;
    const int64_t n_tiles = (N + nt - 1) / nt;
    return n_tiles * (K / ks) * taps * (ks / 8) * (int64_t)nt * 8;
}

__global__ void pack_weight_kernel(const float* __restrict__ w, int taps, int K, int N, int nt, int ks,
                                   __nv_bfloat16* __restrict__ out, int64_t n_elems) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_elems) return;
    const int kc = ks / 8, nks = K / ks;
    int64_t t = i;
    const int e = (int)(t % 8); t /= 8;
    const int n_in = (int)(t % nt); t /= nt;
    const int c = (int)(t % kc); t /= kc;
    const int tap = (int)(t % taps); t /= taps;
    const int ksg = (int)(t % nks); t /= nks;
    const int n_tile = (int)t;
    const int k = ksg * ks + c * 8 + e;
    const int n = n_tile * nt + n_in;
    float v = (n < N) ? w[((int64_t)tap * K + k) * N + n] : 0.f;
    out[i] = __float2bfloat16(v);
}

// Many weights in one launch (training step: every dense layer's weight, and its tap-flipped transpose for the data gradient,
// is packed once per step).  blockIdx.y = entry; an entry with `transposed` packs Wt[tap][n][k] = W[taps - 1 - tap][k][n],
// i.e. the GEMM's (K, N) are the weight's (N, K).
__global__ void __launch_bounds__(256) pack_weight_batch_kernel(const PackEntry* __restrict__ tab) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const PackEntry en = tab[blockIdx.y];
    const int kc = en.ks / 8, nks = en.K / en.ks;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < en.n_elems; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t t = i;
        const int e = (int)(t % 8); t /= 8;
        const int n_in = (int)(t % en.nt); t /= en.nt;
        const int c = (int)(t % kc); t /= kc;
        const int tap = (int)(t % en.taps); t /= en.taps;
        const int ksg = (int)(t % nks); t /= nks;
        const int n_tile = (int)t;
        const int k = ksg * en.ks + c * 8 + e;
        const int n = n_tile * en.nt + n_in;
        float v = 0.f;
        if (n < en.N)
            v = en.transposed ? en.src[((int64_t)(en.taps - 1 - tap) * en.N + n) * en.K + k]     // W is [taps][N][K] in GEMM terms
                              : en.src[((int64_t)tap * en.K + k) * en.N + n];
        en.dst[i] = __float2bfloat16(v);
    }
}

int pack_weight_batch(const PackEntry* tab_dev, int n_entries, cudaStream_t s) {
    if (n_entries <= 0) return 0;
    DX_LAUNCH_PDL((pack_weight_batch_kernel), dim3(192, (unsigned)n_entries), 256, 0, s, tab_dev);
    DX_CHECK_LAUNCH();
    return 0;
}

int pack_weight_bf16(const float* w, int taps, int K, int N, int nt, int ks, __nv_bfloat16* out, cudaStream_t s) {
    const int64_t n = packed_weight_elems(taps, K, N, nt, ks);
    DX_REQUIRE(n > 0, "pack_weight_bf16: bad shape taps=%d K=%d N=%d nt=%d ks=%d", taps, K, N, nt, ks);
    DX_LAUNCH_PDL((pack_weight_kernel), (int)((n + 255) / 256), 256, 0, s, w, taps, K, N, nt, ks, out, n);
    DX_CHECK_LAUNCH();
    return 0;
}

}  // namespace dx
