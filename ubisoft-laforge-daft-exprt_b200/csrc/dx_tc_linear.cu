// Weight-resident tcgen05 linear layer for K = 128 (sm_100a): out = epi(A[rows x 128] · W[128 x N] + b), N <= 384.
//
// Reference ops replaced: in_proj / out_proj of nn.MultiheadAttention (model.py:165, called at :182-186) with the
// post-attention residual + LayerNorm + pad mask (model.py:191, 256) fused into the out-proj epilogue, and the mel
// projection (model.py:546, 561).
//
// Why a second GEMM kernel: these layers are 2-3 % of the FLOPs but were 20 % of the step in the generic
// conv-GEMM kernel (dx_tc_gemm.cu), which streams a weight stage per (M tile, N tile) item and re-reads the A
// tile for every N tile.  Here the whole packed weight (32 KB per 128 output columns, at most 96 KB) is loaded
// into shared memory once per CTA, an A tile (128 rows x 128 channels, 32 KB, 16 bulk copies) is loaded once and
// used for all N tiles, and two groups of four epilogue warps drain the two TMEM accumulators alternately.
//   warp 0 producer | warp 1 TMEM owner + MMA issuer | warps 2-5 epilogue of accumulator 0 | warps 6-9 of accumulator 1
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_ptx.cuh"
#include "dx_tc_epilogue.cuh"

namespace dx {

constexpr int TL_THREADS = 320;
constexpr int TL_BM = 128;
constexpr int TL_ASTAGES = 3;
constexpr uint32_t TL_A_BYTES = 16 * 128 * 16;      // 32 KB: [16 chunk][128 row][8]
constexpr uint32_t TL_W_TILE = 32768;               // one 128-column N tile: [2 k_stage][8 chunk][128][8]
constexpr int TL_MAX_NT = 3;

struct TlParams {
    TcGemmArgs g;
    Rows R;
    int n_tiles;
    uint32_t off_a, off_par, off_bar;               // dynamic smem offsets (W at 0)
};

template <int EPI>
__global__ void __launch_bounds__(TL_THREADS, 1) tc_linear_kernel(const TlParams p) {
    extern __shared__ __align__(1024) uint8_t tl_smem[];
    uint8_t* sW = tl_smem;
    uint8_t* sA = tl_smem + p.off_a;
    float* s_bias = reinterpret_cast<float*>(tl_smem + p.off_par);     // 384
    float* s_lng = s_bias + TL_MAX_NT * 128;
    float* s_lnb = s_lng + 128;
    uint64_t* bars = reinterpret_cast<uint64_t*>(tl_smem + p.off_bar);
    uint64_t* w_full = bars;                 // 1
    uint64_t* a_full = bars + 1;             // [TL_ASTAGES]
    uint64_t* a_empty = a_full + TL_ASTAGES; // [TL_ASTAGES]
    uint64_t* tfull = a_empty + TL_ASTAGES;  // [2]
    uint64_t* tempty = tfull + 2;            // [2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < TL_MAX_NT * 128; i += TL_THREADS)
        s_bias[i] = (p.g.bias != nullptr && i < p.g.N) ? p.g.bias[i] : 0.f;
    if (EPI == TC_EPI_LN && threadIdx.x < 128) {
        s_lng[threadIdx.x] = p.g.ln_g[threadIdx.x];
        s_lnb[threadIdx.x] = p.g.ln_b[threadIdx.x];
    }
    if (threadIdx.x == 0) {
        ptx::mbar_init(w_full, 1);
        for (int i = 0; i < TL_ASTAGES; ++i) { ptx::mbar_init(&a_full[i], 1); ptx::mbar_init(&a_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { ptx::mbar_init(&tfull[i], 1); ptx::mbar_init(&tempty[i], 4); }
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<256>(tmem_ptr);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    // PDL: the weights do not depend on the previous kernel: fetch them while it drains
    if (warp == 0 && ptx::elect_one()) {
        const uint32_t wbytes = (uint32_t)p.n_tiles * TL_W_TILE;
        ptx::mbar_expect_tx(w_full, wbytes);
        for (int t = 0; t < p.n_tiles; ++t)
            ptx::bulk_g2s(sW + (size_t)t * TL_W_TILE, reinterpret_cast<const uint8_t*>(p.g.w) + (size_t)t * TL_W_TILE, TL_W_TILE, w_full);
    }
    __syncwarp();
    ptx::griddep_launch();
    ptx::griddep_wait();
    const int64_t total = p.R.row_start[p.R.B];
    const int m_tiles = (int)((total + TL_BM - 1) / TL_BM);

    if (warp == 0) {
        // ===================== producer =====================
        if (ptx::elect_one()) {
            uint32_t it = 0;
            for (int m_tile = blockIdx.x; m_tile < m_tiles; m_tile += gridDim.x, ++it) {
                const int s = it % TL_ASTAGES;
                ptx::mbar_wait(&a_empty[s], ((it / TL_ASTAGES) & 1) ^ 1);
                ptx::mbar_expect_tx(&a_full[s], TL_A_BYTES);
                const int64_t r0 = (int64_t)m_tile * TL_BM;
                for (int c = 0; c < 16; ++c)
                    ptx::bulk_g2s(sA + (size_t)s * TL_A_BYTES + c * 2048, p.g.a + ((int64_t)c * p.g.a_rows_phys + DX_HALO + r0) * 8, 2048,
                                  &a_full[s]);
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        constexpr uint32_t idesc = ptx::make_idesc_bf16(TL_BM, 128);
        const uint32_t aW = ptx::smem_u32(sW), aA = ptx::smem_u32(sA);
        ptx::mbar_wait(w_full, 0);        // unconditional: the CTA must not exit with the weight copy in flight
        ptx::tc_fence_after();
        uint32_t it = 0, item = 0;
        for (int m_tile = blockIdx.x; m_tile < m_tiles; m_tile += gridDim.x, ++it) {
            const int s = it % TL_ASTAGES;
            ptx::mbar_wait(&a_full[s], (it / TL_ASTAGES) & 1);
            ptx::tc_fence_after();
            for (int t = 0; t < p.n_tiles; ++t, ++item) {
                const int acc = item & 1;
                ptx::mbar_wait(&tempty[acc], ((item >> 1) & 1) ^ 1);
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk) {
                        const uint64_t da = ptx::make_desc(aA + s * TL_A_BYTES + kk * 2 * 2048, 2048, 128);
                        const uint64_t db = ptx::make_desc(aW + t * TL_W_TILE + (kk >> 2) * 16384 + (kk & 3) * 2 * 2048, 2048, 128);
                        ptx::mma_bf16(tmem_base + (uint32_t)acc * 128, da, db, idesc, kk != 0);
                    }
                    ptx::mma_commit(&tfull[acc]);
                    if (t == p.n_tiles - 1) ptx::mma_commit(&a_empty[s]);
                }
                __syncwarp();
            }
        }
    } else {
        // ===================== epilogue: group g = (warp - 2) / 4 drains accumulator g =====================
        const int grp = (warp - 2) >> 2;
        const int lane_grp = warp & 3;                 // TMEM lanes 32*lane_grp .. +31
        uint32_t item = 0;
        for (int m_tile = blockIdx.x; m_tile < m_tiles; m_tile += gridDim.x) {
            const int64_t r = (int64_t)m_tile * TL_BM + lane_grp * 32 + lane;
            bool live = r < total;
            int b_idx = 0;
            if (live) {
                b_idx = p.R.row_b[r];
                if (p.g.live_ext >= 0) live = row_live(p.R, r, p.g.live_ext);
                else live = b_idx >= 0;
                if (b_idx < 0) b_idx = 0;
            }
            for (int t = 0; t < p.n_tiles; ++t, ++item) {
                if ((int)(item & 1) != grp) continue;
                ptx::mbar_wait(&tfull[grp], (item >> 1) & 1);
                ptx::tc_fence_after();
                const uint32_t taddr = tmem_base + ((uint32_t)(lane_grp * 32) << 16) + (uint32_t)grp * 128;
                tc_epilogue_tile<128, EPI>(p.g, p.R, total, r, t, taddr, &tempty[grp], lane, live, b_idx, s_bias, s_lng, s_lnb);
            }
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<256>(tmem_base);
    }
}

template <int EPI>
static int tl_launch(const TlParams& p, size_t smem, int grid, cudaStream_t s) {
    static SmemConfigured configured;
    DX_TRY(ensure_dyn_smem(tc_linear_kernel<EPI>, smem, configured));
    ProfScope prof(p.g.tag ? p.g.tag : "tc_linear", s);
    DX_LAUNCH_PDL((tc_linear_kernel<EPI>), grid, TL_THREADS, smem, s, p);
    DX_CHECK_LAUNCH();
    return 0;
}

int tc_linear(const TcGemmArgs& g, const Rows& R, cudaStream_t s) {
    DX_REQUIRE(g.K == 128 && g.taps == 1 && g.nt == 128 && g.ks == 64, "tc_linear: needs K=128, taps=1, nt=128, ks=64 (K=%d taps=%d nt=%d ks=%d)",
               g.K, g.taps, g.nt, g.ks);
    DX_REQUIRE(g.row_stats == nullptr, "tc_linear: row_stats not supported");
    int num_sms = 0;
    DX_TRY(device_sm_count(&num_sms));
    TlParams p;
    p.g = g;
    p.R = R;
    p.n_tiles = (g.N + 127) / 128;
    DX_REQUIRE(p.n_tiles >= 1 && p.n_tiles <= TL_MAX_NT, "tc_linear: N=%d wider than %d", g.N, TL_MAX_NT * 128);
    DX_REQUIRE(g.epilogue != TC_EPI_LN || (g.N == 128 && g.bias && g.ln_g && g.ln_b), "tc_linear: LN epilogue needs N == 128, bias, gamma, beta");
    p.off_a = (uint32_t)p.n_tiles * TL_W_TILE;
    p.off_par = p.off_a + TL_ASTAGES * TL_A_BYTES;
    p.off_bar = p.off_par + (TL_MAX_NT * 128 + 256) * sizeof(float);
    const size_t smem = p.off_bar + (2 * TL_ASTAGES + 5) * sizeof(uint64_t) + 16;
    const int64_t m_tiles = (R.cap + TL_BM - 1) / TL_BM;
    int grid = (int)(m_tiles < num_sms ? m_tiles : num_sms);
    if (grid < 1) grid = 1;
    if (g.epilogue == TC_EPI_LN) return tl_launch<TC_EPI_LN>(p, smem, grid, s);
    if (g.epilogue == TC_EPI_F32) return tl_launch<TC_EPI_F32>(p, smem, grid, s);
    if (g.epilogue == TC_EPI_BF16) return tl_launch<TC_EPI_BF16>(p, smem, grid, s);
    DX_REQUIRE(false, "tc_linear: unknown epilogue %d", g.epilogue);
    return 0;
}

}  // namespace dx
