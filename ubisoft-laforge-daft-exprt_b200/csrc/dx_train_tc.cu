// Tensor-core GEMMs for the training step (BASELINE config 5, DX_TRAIN_TC_GEMM): the step keeps its fp32 row-stream
// activations and gradients ([rows][C], dx_train.cu), and every large dense contraction runs on tcgen05 with bf16 copies
// of its operands and fp32 accumulation:
//   forward / dgrad   out = conv(A, W)                    rows_to_il(A) + pack_weight_bf16(W) + tc_gemm (dx_tc_gemm.cu)
//   wgrad             dW[tap][k][n] = sum_r A[r + tap - half][k] G[r][n]     rows_to_il(A), rows_to_il(G) + wgrad_tc_kernel (here)
// They replace gemm_f32 / wgrad_f32 (fp32 SIMT, 39 of the step's 53 ms), i.e. autograd's backward of nn.Conv1d / nn.Linear
// in model.py:165,187,206-217,546,649-671 under train.py:434.
//
// rows_to_il: fp32 [rows][C] -> the bf16 channel-interleaved layout [C/8][rows_phys][8] of the inference kernels (DESIGN.md
// section 3), zero halo in front, zeros for DX_TAIL rows behind the stream's total row count (no reader goes further): the buffer
// may hold anything before.
//
// wgrad_tc_kernel: the reduction runs over ROWS, so both operands are "MN-major" for the tensor core, and the interleaved
// layout is exactly the no-swizzle MN-major canonical layout with rows as K (the way attention feeds V): for one 8-channel
// chunk, 64 rows are one contiguous 1 KB run = one bulk copy; core-matrix groups of 8 rows are 128 B apart (LBO), chunks
// one tile apart (SBO).  One CTA owns a (128 k-channels) x (128 n-channels) tile of dW for all taps (3 accumulators in
// TMEM, the conv tap is a +16 B shift of the A descriptor, one staged A tile serves all taps) over one contiguous share of
// the row blocks; the S shares of a tile are summed afterwards in a fixed order (reduce_partials): deterministic, no atomics.
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_ptx.cuh"
#include "dx_train.cuh"

namespace dx {

// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
rows_to_il_kernel(const float* __restrict__ src, int C, int C_src, const int32_t* __restrict__ total_ptr,
                  __nv_bfloat16* __restrict__ dst, int64_t rows_phys) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    __shared__ __align__(16) __nv_bfloat16 tile[32][128 + 8];
    const int64_t total = *total_ptr;
    const int64_t pr0 = (int64_t)blockIdx.x * 32;        // physical row of the tile: stream row = physical - DX_HALO
    // readers (GEMM row tiles of 128 + taps, 64-row weight-gradient stages, attention key tiles) stop within DX_TAIL rows of the
    // stream's end; the buffer is sized for the worst case (every utterance at full length), typically 1.5-8x the stream
    if (pr0 >= DX_HALO + total + DX_TAIL) return;
    const int c0 = blockIdx.y * 128;
    const int t = threadIdx.x;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int idx = t + i * 256, row = idx >> 5, c4 = idx & 31;
        const int64_t r = pr0 + row - DX_HALO;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r >= 0 && r < total && c0 + c4 * 4 < C_src) v = *reinterpret_cast<const float4*>(src + r * C_src + c0 + c4 * 4);
        __nv_bfloat162* d2 = reinterpret_cast<__nv_bfloat162*>(&tile[row][c4 * 4]);
        d2[0] = __floats2bfloat162_rn(v.x, v.y);
        d2[1] = __floats2bfloat162_rn(v.z, v.w);
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int idx = t + i * 256, chunk = idx >> 5, row = idx & 31;
        if (pr0 + row < rows_phys)
            *reinterpret_cast<uint4*>(dst + ((int64_t)(c0 / 8 + chunk) * rows_phys + pr0 + row) * 8) =
                *reinterpret_cast<const uint4*>(&tile[row][chunk * 8]);
    }
}

// C_src < C: the source rows have C_src channels (a multiple of 4); channels C_src .. C - 1 of the copy are zero
int rows_to_il(const float* src, int C, const Rows& R, __nv_bfloat16* dst, int64_t rows_phys, cudaStream_t s, int C_src) {
    if (C_src <= 0) C_src = C;
    DX_REQUIRE(C % 128 == 0 && C_src % 4 == 0 && C_src <= C, "rows_to_il: C=%d (multiple of 128), C_src=%d (multiple of 4, <= C)", C, C_src);
    DX_REQUIRE(rows_phys >= R.cap + DX_HALO + DX_TAIL, "rows_to_il: rows_phys too small");
    dim3 grid((unsigned)((rows_phys + 31) / 32), (unsigned)(C / 128));
    ProfScope prof("rows_to_il", s);
    DX_LAUNCH_PDL((rows_to_il_kernel), grid, 256, 0, s, src, C, C_src, R.row_start + R.B, dst, rows_phys);
    DX_CHECK_LAUNCH();
    return 0;
}

// dst[t][k][n] = src[t][k][n] inside the source's extent, 0 outside: zero-pads (Kd >= Ks, Nd >= Ns) or crops a weight tensor
__global__ void copy_pad3_kernel(const float* __restrict__ src, int Ks, int Ns, float* __restrict__ dst, int Kd, int Nd, int64_t n) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int nn = (int)(i % Nd);
    const int k = (int)((i / Nd) % Kd);
    const int t = (int)(i / ((int64_t)Nd * Kd));
    dst[i] = (k < Ks && nn < Ns) ? src[((int64_t)t * Ks + k) * Ns + nn] : 0.f;
}
int copy_pad3(const float* src, int T, int Ks, int Ns, float* dst, int Kd, int Nd, cudaStream_t s) {
    const int64_t n = (int64_t)T * Kd * Nd;
    DX_LAUNCH_PDL((copy_pad3_kernel), (unsigned)((n + 255) / 256), 256, 0, s, src, Ks, Ns, dst, Kd, Nd, n);
    DX_CHECK_LAUNCH();
    return 0;
}

// il_to_rows: bf16 [C/8][rows_phys][8] -> fp32 [rows][C]; rows that are not valid positions (gap rows, rows beyond the
// stream) get zeros: the producer (attention) only writes valid rows and the buffer may hold anything elsewhere
__global__ void __launch_bounds__(256)
il_to_rows_kernel(const __nv_bfloat16* __restrict__ src, int C, Rows R, int64_t rows_phys, float* __restrict__ dst) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    __shared__ __align__(16) __nv_bfloat16 tile[32][128 + 8];
    const int64_t total = R.row_start[R.B];
    const int64_t r0 = (int64_t)blockIdx.x * 32;
    const int c0 = blockIdx.y * 128;
    const int t = threadIdx.x;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int idx = t + i * 256, chunk = idx >> 5, row = idx & 31;
        *reinterpret_cast<uint4*>(&tile[row][chunk * 8]) =
            *reinterpret_cast<const uint4*>(src + ((int64_t)(c0 / 8 + chunk) * rows_phys + DX_HALO + r0 + row) * 8);
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int idx = t + i * 256, row = idx >> 5, c4 = idx & 31;
        const int64_t r = r0 + row;
        if (r >= R.cap) continue;
        bool ok = r < total;
        if (ok) {
            const int b = R.row_b[r];
            ok = b >= 0 && R.row_p[r] < R.len[b];
        }
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ok) {
            const __nv_bfloat162* s2 = reinterpret_cast<const __nv_bfloat162*>(&tile[row][c4 * 4]);
            v = make_float4(__low2float(s2[0]), __high2float(s2[0]), __low2float(s2[1]), __high2float(s2[1]));
        }
        *reinterpret_cast<float4*>(dst + r * C + c0 + c4 * 4) = v;
    }
}

int il_to_rows(const __nv_bfloat16* src, int C, const Rows& R, int64_t rows_phys, float* dst, cudaStream_t s) {
    DX_REQUIRE(C % 128 == 0, "il_to_rows: C=%d must be a multiple of 128", C);
    dim3 grid((unsigned)((R.cap + 31) / 32), (unsigned)(C / 128));
    DX_LAUNCH_PDL((il_to_rows_kernel), grid, 256, 0, s, src, C, R, rows_phys, dst);
    DX_CHECK_LAUNCH();
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
constexpr int WT_RB = 64;                              // rows per pipeline stage
constexpr int WT_STAGES = 4;
constexpr uint32_t WT_A_STRIDE = 72 * 16;              // staged A rows per chunk: 66 used (64 + one halo row per side)
constexpr uint32_t WT_A_BYTES = 16 * WT_A_STRIDE;      // 18,432
constexpr uint32_t WT_G_BYTES = 16 * WT_RB * 16;       // 16,384
constexpr uint32_t WT_STAGE_BYTES = WT_A_BYTES + WT_G_BYTES;
constexpr uint32_t WT_TR_FLOATS = 4 * 32 * 33;         // per-warp transpose tiles of the epilogue
constexpr uint32_t WT_SMEM = WT_STAGES * WT_STAGE_BYTES + WT_TR_FLOATS * 4 + (2 * WT_STAGES + 1) * 8 + 16;
constexpr int WT_THREADS = 192;

struct WgradTcParams {
    const __nv_bfloat16 *a, *g;     // [K/8][rows_phys][8], [N/8][rows_phys][8]
    int64_t rows_phys;
    int K, N, taps, S, n_nt;
    const int32_t* total_ptr;
    float* part;                    // [S][taps][K][N]
};

__global__ void __launch_bounds__(WT_THREADS, 1) wgrad_tc_kernel(const WgradTcParams p) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    extern __shared__ __align__(1024) uint8_t wt_smem[];
    uint8_t* stage_base = wt_smem;
    float* s_tr = reinterpret_cast<float*>(wt_smem + WT_STAGES * WT_STAGE_BYTES);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(s_tr + WT_TR_FLOATS);
    uint64_t* empty_bar = full_bar + WT_STAGES;
    uint64_t* tfull_bar = empty_bar + WT_STAGES;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tfull_bar + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < WT_STAGES; ++i) { ptx::mbar_init(&full_bar[i], 1); ptx::mbar_init(&empty_bar[i], 1); }
        ptx::mbar_init(tfull_bar, 1);
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<512>(tmem_ptr);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    const int sp = (int)(blockIdx.x % p.S), tile_id = (int)(blockIdx.x / p.S);
    const int mt = tile_id / p.n_nt, nt = tile_id % p.n_nt;
    const int64_t total = *p.total_ptr;
    const int64_t n_blocks = (total + WT_RB - 1) / WT_RB;
    const int64_t b0 = sp * n_blocks / p.S, b1 = (sp + 1) * n_blocks / p.S;
    const int half = p.taps / 2;

    if (warp == 0) {
        if (ptx::elect_one()) {
            uint32_t it = 0;
            for (int64_t blk = b0; blk < b1; ++blk, ++it) {
                const int s = it % WT_STAGES;
                ptx::mbar_wait(&empty_bar[s], ((it / WT_STAGES) & 1) ^ 1);
                uint8_t* sa = stage_base + (size_t)s * WT_STAGE_BYTES;
                uint8_t* sg = sa + WT_A_BYTES;
                const int64_t r0 = blk * WT_RB;
                ptx::mbar_expect_tx(&full_bar[s], 16 * (WT_RB + 2) * 16 + WT_G_BYTES);
                for (int c = 0; c < 16; ++c)
                    ptx::bulk_g2s(sa + c * WT_A_STRIDE, p.a + ((int64_t)(mt * 16 + c) * p.rows_phys + DX_HALO + r0 - 1) * 8,
                                  (WT_RB + 2) * 16, &full_bar[s]);
                for (int c = 0; c < 16; ++c)
                    ptx::bulk_g2s(sg + c * WT_RB * 16, p.g + ((int64_t)(nt * 16 + c) * p.rows_phys + DX_HALO + r0) * 8, WT_RB * 16,
                                  &full_bar[s]);
            }
        }
    } else if (warp == 1) {
        // A and B both MN-major (bits 15, 16): the contraction runs over rows
        constexpr uint32_t idesc = ptx::make_idesc_bf16(128, 128) | (1u << 15) | (1u << 16);
        uint32_t it = 0;
        for (int64_t blk = b0; blk < b1; ++blk, ++it) {
            const int s = it % WT_STAGES;
            ptx::mbar_wait(&full_bar[s], (it / WT_STAGES) & 1);
            ptx::tc_fence_after();
            if (ptx::elect_one()) {
                const uint32_t sa = ptx::smem_u32(stage_base + (size_t)s * WT_STAGE_BYTES);
                const uint32_t sg = sa + WT_A_BYTES;
                for (int tap = 0; tap < p.taps; ++tap) {
#pragma unroll
                    for (int kk = 0; kk < WT_RB / 16; ++kk) {
                        // G row r0 + i pairs with A row r0 + i + tap - half = staged A row i + tap - half + 1
                        const uint64_t da = ptx::make_desc(sa + (uint32_t)(tap - half + 1) * 16 + kk * 256, 128, WT_A_STRIDE);
                        const uint64_t db = ptx::make_desc(sg + kk * 256, 128, WT_RB * 16);
                        ptx::mma_bf16(tmem_base + (uint32_t)tap * 128, da, db, idesc, blk != b0 || kk != 0);
                    }
                }
                ptx::mma_commit(&empty_bar[s]);
                if (blk == b1 - 1) ptx::mma_commit(tfull_bar);
            }
            __syncwarp();
        }
    } else {
        // epilogue: thread <-> TMEM lane = k row of the tile; columns = n.  Through a per-warp 32 x 32 transpose so that a
        // warp's stores are 128 contiguous bytes of one dW row.
        const int lane_grp = warp & 3;
        float* tr = s_tr + (warp - 2) * 32 * 33;
        const bool any = b1 > b0;
        if (any) {
            ptx::mbar_wait(tfull_bar, 0);
            ptx::tc_fence_after();
        }
        for (int tap = 0; tap < p.taps; ++tap) {
            float* out = p.part + (((int64_t)sp * p.taps + tap) * p.K + mt * 128 + lane_grp * 32) * p.N + nt * 128;
            for (int q = 0; q < 4; ++q) {
                float v[32];
                if (any) {
                    ptx::tmem_ld32(tmem_base + ((uint32_t)(lane_grp * 32) << 16) + (uint32_t)tap * 128 + q * 32, v);
                    ptx::tmem_ld_wait();
                } else {
#pragma unroll
                    for (int e = 0; e < 32; ++e) v[e] = 0.f;
                }
#pragma unroll
                for (int e = 0; e < 32; ++e) tr[lane * 33 + e] = v[e];
                __syncwarp();
#pragma unroll 8
                for (int j = 0; j < 32; ++j) out[(int64_t)j * p.N + q * 32 + lane] = tr[j * 33 + lane];
                __syncwarp();
            }
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<512>(tmem_base);
    }
}

static int wgrad_tc_splits(int K, int N) {
    const int tiles = (K / 128) * (N / 128);
    // CTAs per launch.  One per SM: with two (296) every CTA's share of the rows halves while its fixed cost - TMEM allocation, the
    // transposed fp32 dump of its 3 x 128 x 128 partial tile and the partial's trip through reduce_partials - does not (measured,
    // B = 32 step: 296 -> 8.14 ms, 222 -> 8.15, 148 -> 7.77).  Debug knob 18 overrides.
    const int budget = debug_knob(18) > 0 ? debug_knob(18) : 148;
    int S = budget / tiles;
    if (S < 1) S = 1;
    if (S > 64) S = 64;
    return S;
}

int64_t wgrad_tc_scratch_floats() { return (int64_t)296 * 3 * 128 * 128; }

int wgrad_tc(const __nv_bfloat16* a_il, int K, const __nv_bfloat16* g_il, int N, int taps, const Rows& R, int64_t rows_phys, float* dW,
             float* part, cudaStream_t s) {
    DX_REQUIRE(K % 128 == 0 && N % 128 == 0 && (taps == 1 || taps == 3), "wgrad_tc: K=%d N=%d taps=%d", K, N, taps);
    static SmemConfigured configured;
    DX_TRY(ensure_dyn_smem(wgrad_tc_kernel, (size_t)WT_SMEM, configured));
    WgradTcParams p;
    p.a = a_il; p.g = g_il; p.rows_phys = rows_phys; p.K = K; p.N = N; p.taps = taps;
    p.S = wgrad_tc_splits(K, N);
    p.n_nt = N / 128;
    p.total_ptr = R.row_start + R.B;
    p.part = part;
    DX_REQUIRE((int64_t)p.S * taps * K * N <= wgrad_tc_scratch_floats(), "wgrad_tc: scratch too small");
    const int grid = (K / 128) * (N / 128) * p.S;
    {
        ProfScope prof("wgrad_tc", s);
        DX_LAUNCH_PDL((wgrad_tc_kernel), grid, WT_THREADS, WT_SMEM, s, p);
        DX_CHECK_LAUNCH();
    }
    return reduce_partials(part, p.S, (int64_t)taps * K * N, dW, nullptr, 0, 1.f, s);
}

}  // namespace dx
