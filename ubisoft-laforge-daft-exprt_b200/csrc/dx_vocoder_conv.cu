// The convolution kernels of the HiFi-GAN generator (dx_vocoder.cu has the layout, the plan and the citations).
//
//   out[r][n] = epi( sum_{tap} sum_k A[r + (tap - taps/2) * dil][k] * W[tap][k][n] )
//
// vc_conv_tc_kernel<NT>  bf16 operands on tcgen05 (fp32 accumulators in TMEM, double buffered), operands staged by the
//                        bulk-copy engine; persistent, warp-specialised: warp 0 producer, warp 1 TMEM owner + MMA
//                        issuer, warps 2-5 epilogue (thread ↔ output row).  One pipeline stage = `ks` input channels x
//                        all taps: the A tile (128 + 2 * halo rows) is staged once and every tap reads it through a
//                        descriptor start address shifted by tap * dil rows.
// vc_conv_f32_kernel     fp32 SIMT, 128 x 32 output tile per block, 4 x 4 per thread: the parity mode.
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_ptx.cuh"
#include "dx_vocoder.cuh"

namespace dx {

struct VcTcParams {
    VcConvArgs c;
    int n_tiles, nks, kc, nstages, halo, ra;
    uint32_t a_stage_bytes, w_stage_bytes, par_off;
    int dbg;   // timing experiments (debug knob 11): 1 no epilogue, 2 no A copies, 4 no W copies, 8 no MMAs
};

// ---- tensor-core epilogue: 32 consecutive columns (one phase) of one input-rate row per thread ---------------------
// Split in two so that nothing with DRAM latency sits behind the accumulator wait: vc_tc_prefetch (row validity and
// the residual, both independent of the accumulator) is issued BEFORE the thread waits for the MMAs; vc_tc_finish
// does loads-then-math-then-stores in that program order (res / xs / x_out may alias, so the compiler cannot hoist a
// load above a store by itself — interleaved, the 8 residual loads became 8 serialised DRAM round trips: 18 k cycles
// per tile, 3x the rest of the kernel; knob-11 timing experiments).
struct VcPre {
    float4 res[8];
};

__device__ __forceinline__ uint32_t vc_pack2(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

__device__ __forceinline__ void vc_tc_prefetch(const VcEpi& e, int64_t r, int n, bool valid, VcPre& pre) {
    if (e.res == nullptr || !valid) return;
    const int ph = n / e.c_out, co = n - ph * e.c_out;
    const float* src = e.res + ((int64_t)(co >> 2) * e.x_rows + r * e.up + ph) * 4;
#pragma unroll
    for (int g = 0; g < 8; ++g) pre.res[g] = *reinterpret_cast<const float4*>(src + (int64_t)g * e.x_rows * 4);
}

__device__ __forceinline__ void vc_tc_finish(const VcEpi& e, int64_t r, int n, bool valid, float* v, const VcPre& pre,
                                             const float* s_bias_n) {
    const int ph = n / e.c_out, co = n - ph * e.c_out;
    const int64_t ro = r * e.up + ph;
    if (valid) {
        const int64_t off = ((int64_t)(co >> 2) * e.x_rows + ro) * 4;
        const int64_t gs = e.x_rows * 4;
        float4 xs[8];
        if (e.xs_mode >= 2) {
#pragma unroll
            for (int g = 0; g < 8; ++g) xs[g] = *reinterpret_cast<const float4*>(e.xs + off + g * gs);
        }
#pragma unroll
        for (int g = 0; g < 8; ++g) {
            const float4 bb = *reinterpret_cast<const float4*>(s_bias_n + 4 * g);
            float4 y = make_float4(v[4 * g] + bb.x, v[4 * g + 1] + bb.y, v[4 * g + 2] + bb.z, v[4 * g + 3] + bb.w);
            if (e.res) { y.x += pre.res[g].x; y.y += pre.res[g].y; y.z += pre.res[g].z; y.w += pre.res[g].w; }
            if (e.xs_mode >= 2) {                                              // xs += resblock(x), hifigan.py:170
                y.x += xs[g].x; y.y += xs[g].y; y.z += xs[g].z; y.w += xs[g].w;
                if (e.xs_mode == 3) { y.x /= e.n_kernels; y.y /= e.n_kernels; y.z /= e.n_kernels; y.w /= e.n_kernels; }   // :171
            }
            v[4 * g] = y.x; v[4 * g + 1] = y.y; v[4 * g + 2] = y.z; v[4 * g + 3] = y.w;
        }
        if (e.xs_mode == 1 || e.xs_mode == 2) {
#pragma unroll
            for (int g = 0; g < 8; ++g)
                *reinterpret_cast<float4*>(e.xs + off + g * gs) = make_float4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
        }
        if (e.x_out) {
#pragma unroll
            for (int g = 0; g < 8; ++g)
                *reinterpret_cast<float4*>(e.x_out + off + g * gs) = make_float4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
        }
    }
    if (e.act_out == nullptr) return;
    __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(e.act_out) + ((int64_t)(co >> 3) * e.act_rows_phys + DX_VC_HALO + ro) * 8;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        uint4 pk;
        uint32_t* w = reinterpret_cast<uint32_t*>(&pk);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float a = v[c * 8 + 2 * q], b = v[c * 8 + 2 * q + 1];
            if (e.act) { a = vc_lrelu(a); b = vc_lrelu(b); }
            w[q] = valid ? vc_pack2(a, b) : 0u;
        }
        *reinterpret_cast<uint4*>(out + (int64_t)c * e.act_rows_phys * 8) = pk;
    }
}

template <int NT>
__global__ void __launch_bounds__(VC_THREADS, NT == 128 ? 2 : 3) vc_conv_tc_kernel(const VcTcParams p) {
    extern __shared__ __align__(1024) uint8_t vc_smem[];
    const uint32_t stage_bytes = p.a_stage_bytes + p.w_stage_bytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(vc_smem + (size_t)p.nstages * stage_bytes);
    uint64_t* empty_bar = full_bar + p.nstages;
    uint64_t* tfull_bar = empty_bar + p.nstages;   // [2]
    uint64_t* tempty_bar = tfull_bar + 2;          // [2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty_bar + 2);
    float* s_bias = reinterpret_cast<float*>(vc_smem + p.par_off);   // [N]: bias of column n = bias[n % c_out]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < p.c.N; i += VC_THREADS) s_bias[i] = p.c.e.bias[i % p.c.e.c_out];
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
    ptx::griddep_launch();
    ptx::griddep_wait();
    const int64_t total = (int64_t)(*p.c.total_frames) << p.c.in_shift;
    const int m_tiles = (int)((total + VC_BM - 1) / VC_BM);
    const int64_t n_items = (int64_t)m_tiles * p.n_tiles;
    const uint32_t a_rows_bytes = (uint32_t)(VC_BM + 2 * p.halo) * 16;

    if (warp == 0) {
        // ===================== producer =====================
        if (ptx::elect_one()) {
            const __nv_bfloat16* a = reinterpret_cast<const __nv_bfloat16*>(p.c.a);
            uint32_t it = 0;
            for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x) {
                const int m_tile = (int)(item / p.n_tiles), n_tile = (int)(item % p.n_tiles);
                const int64_t r0 = (int64_t)m_tile * VC_BM;
                for (int ksg = 0; ksg < p.nks; ++ksg, ++it) {
                    const int s = it % p.nstages;
                    const uint32_t ph = (it / p.nstages) & 1;
                    ptx::mbar_wait_sleep(&empty_bar[s], ph ^ 1, 64);
                    uint8_t* sa = vc_smem + (size_t)s * stage_bytes;
                    uint8_t* sw = sa + p.a_stage_bytes;
                    ptx::mbar_expect_tx(&full_bar[s], ((p.dbg & 2) ? 0u : (uint32_t)p.kc * a_rows_bytes) + ((p.dbg & 4) ? 0u : p.w_stage_bytes));
                    for (int c = 0; c < ((p.dbg & 2) ? 0 : p.kc); ++c) {
                        const __nv_bfloat16* src =
                            a + ((int64_t)(ksg * p.kc + c) * p.c.a_rows_phys + DX_VC_HALO + r0 - p.halo) * 8;
                        ptx::bulk_g2s(sa + (size_t)c * p.ra * 16, src, a_rows_bytes, &full_bar[s]);
                    }
                    const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(p.c.w_bf) +
                                          ((int64_t)n_tile * p.nks + ksg) * (int64_t)p.w_stage_bytes;
                    if (!(p.dbg & 4)) ptx::bulk_g2s(sw, wsrc, p.w_stage_bytes, &full_bar[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        constexpr uint32_t idesc = ptx::make_idesc_bf16(VC_BM, NT);
        const uint32_t a_lbo = (uint32_t)p.ra * 16, b_lbo = NT * 16;
        uint32_t it = 0, tile_it = 0;
        for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x, ++tile_it) {
            const int acc = tile_it & 1;
            const uint32_t acc_ph = (tile_it >> 1) & 1;
            ptx::mbar_wait_sleep(&tempty_bar[acc], acc_ph ^ 1, 64);
            ptx::tc_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)acc * NT;
            for (int ksg = 0; ksg < p.nks; ++ksg, ++it) {
                const int s = it % p.nstages;
                const uint32_t ph = (it / p.nstages) & 1;
                ptx::mbar_wait_sleep(&full_bar[s], ph, 32);
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
                    const uint32_t sa = ptx::smem_u32(vc_smem + (size_t)s * stage_bytes);
                    const uint32_t sw = sa + p.a_stage_bytes;
                    for (int tap = 0; tap < ((p.dbg & 8) ? 0 : p.c.taps); ++tap) {
                        for (int kk = 0; kk < p.kc / 2; ++kk) {
                            const uint64_t da = ptx::make_desc(sa + (uint32_t)(tap * p.c.dil) * 16 + (uint32_t)kk * 2 * a_lbo, a_lbo, 128);
                            const uint64_t db = ptx::make_desc(sw + (uint32_t)(tap * p.kc + kk * 2) * b_lbo, b_lbo, 128);
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
            const int64_t r = (int64_t)m_tile * VC_BM + lane_grp * 32 + lane;
            const bool inr = r < total;
            const bool valid = inr && p.c.e.frame_utt[r >> p.c.in_shift] >= 0;   // all phases of a row share its frame
            VcPre pre;
            vc_tc_prefetch(p.c.e, r, n_tile * NT, valid, pre);
            ptx::mbar_wait_sleep(&tfull_bar[acc], acc_ph, 100);
            ptx::tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(lane_grp * 32) << 16) + (uint32_t)acc * NT;
#pragma unroll 1
            for (int q = 0; q < NT / 32; ++q) {
                float v[32];
                ptx::tmem_ld32(taddr + q * 32, v);
                const int n = n_tile * NT + q * 32;
                if (q > 0) vc_tc_prefetch(p.c.e, r, n, valid, pre);
                ptx::tmem_ld_wait();
                if (inr && n < p.c.N && !(p.dbg & 1)) vc_tc_finish(p.c.e, r, n, valid, v, pre, s_bias + n);
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&tempty_bar[acc]);
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<2 * NT>(tmem_base);
    }
}

// nt: tile width; ks: input channels per pipeline stage — the largest of 64 / 32 / 16 that divides K and keeps one
// stage (A tile + all taps of the weights) within 36 KB, so that 2-3 CTAs (each with 2-3 stages) share an SM and
// their copy / MMA / epilogue phases overlap
void vc_conv_tiling(int taps, int dil, int K, int N, int* nt, int* ks) {
    const int t = N >= 128 ? 128 : N;
    const int halo = (taps / 2) * dil;
    const int ra = (int)round_up(VC_BM + 2 * halo, 8);
    int k = 16;
    for (int cand = 64; cand >= 16; cand /= 2) {
        if (K % cand != 0) continue;
        const int64_t stage = (int64_t)(cand / 8) * ra * 16 + (int64_t)taps * (cand / 8) * t * 16;
        if (stage <= 36 * 1024) { k = cand; break; }
    }
    *nt = t;
    *ks = k;
}

static int g_vc_sms = 0;

template <int NT>
static int vc_tc_launch(const VcTcParams& p, size_t smem, int grid, cudaStream_t s) {
    static size_t configured = 0;
    if (smem > configured) {
        DX_CHECK_CUDA(cudaFuncSetAttribute(vc_conv_tc_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    ProfScope prof(p.c.tag ? p.c.tag : "vc_conv", s);
    DX_LAUNCH_PDL((vc_conv_tc_kernel<NT>), grid, VC_THREADS, smem, s, p);
    DX_CHECK_LAUNCH();
    return 0;
}

int vc_conv_bf16(const VcConvArgs& c, cudaStream_t s) {
    DX_REQUIRE(c.K % 16 == 0 && c.N % 32 == 0 && c.N <= VC_MAX_N, "vc_conv: K=%d N=%d (K %% 16, N %% 32, N <= %d)", c.K, c.N, VC_MAX_N);
    DX_REQUIRE(c.taps % 2 == 1 && (c.taps / 2) * c.dil <= DX_VC_HALO, "vc_conv: taps=%d dil=%d exceed the %d-row halo", c.taps, c.dil, DX_VC_HALO);
    DX_REQUIRE(c.e.c_out % 32 == 0 && c.N == c.e.c_out * c.e.up, "vc_conv: N=%d != up=%d * c_out=%d", c.N, c.e.up, c.e.c_out);
    DX_REQUIRE(c.w_bf != nullptr, "vc_conv: packed bf16 weights missing (tag %s)", c.tag ? c.tag : "?");
    if (g_vc_sms == 0) {
        int dev = 0;
        DX_CHECK_CUDA(cudaGetDevice(&dev));
        DX_CHECK_CUDA(cudaDeviceGetAttribute(&g_vc_sms, cudaDevAttrMultiProcessorCount, dev));
    }
    VcTcParams p;
    p.c = c;
    p.dbg = debug_knob(11);
    int nt, ks;
    vc_conv_tiling(c.taps, c.dil, c.K, c.N, &nt, &ks);
    p.n_tiles = (c.N + nt - 1) / nt;
    p.nks = c.K / ks;
    p.kc = ks / 8;
    p.halo = (c.taps / 2) * c.dil;
    p.ra = (int)round_up(VC_BM + 2 * p.halo, 8);
    p.a_stage_bytes = (uint32_t)p.kc * p.ra * 16;
    p.w_stage_bytes = (uint32_t)c.taps * p.kc * nt * 16;
    const size_t stage = (size_t)p.a_stage_bytes + p.w_stage_bytes;
    int nst = (int)((104 * 1024) / stage);
    if (nst > 3) nst = 3;
    if (nst < 2) nst = 2;
    DX_REQUIRE(2 * stage <= 200 * 1024, "vc_conv: stage of %zu bytes does not allow 2 stages", stage);
    p.nstages = nst;
    p.par_off = (uint32_t)round_up((int64_t)((size_t)nst * stage + (2 * nst + 4) * sizeof(uint64_t) + 16), 16);
    const size_t smem = p.par_off + (size_t)c.N * sizeof(float);
    int64_t items = ((c.rows_bound + VC_BM - 1) / VC_BM) * p.n_tiles;
    // CTAs per SM: bounded by shared memory (227 KB) and by TMEM (2 * nt of 512 columns each)
    int per_sm = (int)((220 * 1024) / (smem + 1024));
    if (per_sm > 512 / (2 * nt)) per_sm = 512 / (2 * nt);
    if (per_sm > 3) per_sm = 3;
    if (per_sm < 1) per_sm = 1;
    int grid = (int)(items < (int64_t)g_vc_sms * per_sm ? items : (int64_t)g_vc_sms * per_sm);
    if (grid < 1) grid = 1;
    if (nt == 128) return vc_tc_launch<128>(p, smem, grid, s);
    if (nt == 64) return vc_tc_launch<64>(p, smem, grid, s);
    if (nt == 32) return vc_tc_launch<32>(p, smem, grid, s);
    DX_REQUIRE(false, "vc_conv: tile width %d", nt);
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// fp32 SIMT path
// ---------------------------------------------------------------------------------------------------------
constexpr int VF_BM = 128, VF_BN = 32, VF_BK = 16;

__global__ void __launch_bounds__(256) vc_conv_f32_kernel(const VcConvArgs c) {
    __shared__ float As[VF_BK][VF_BM + 4];
    __shared__ float Ws[VF_BK][VF_BN];
    ptx::griddep_launch();
    ptx::griddep_wait();
    const int64_t total = (int64_t)(*c.total_frames) << c.in_shift;
    const int64_t r0 = (int64_t)blockIdx.x * VF_BM;
    if (r0 >= total) return;
    const int n0 = blockIdx.y * VF_BN;
    const int tid = threadIdx.x, ty = tid >> 3, tx = tid & 7;
    const float* A = reinterpret_cast<const float*>(c.a);
    const int half = c.taps / 2;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    for (int tap = 0; tap < c.taps; ++tap) {
        const int64_t shift = (int64_t)(tap - half) * c.dil;
        for (int k0 = 0; k0 < c.K; k0 += VF_BK) {
            // A tile: 128 rows x 16 k (2 float4 per thread), stored transposed
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int idx = tid + q * 256;
                const int row = idx >> 2, kq = (idx & 3) * 4;
                const int64_t r = r0 + row + shift;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r >= 0 && r < total) v = *reinterpret_cast<const float4*>(A + r * c.K + k0 + kq);
                As[kq][row] = v.x; As[kq + 1][row] = v.y; As[kq + 2][row] = v.z; As[kq + 3][row] = v.w;
            }
            if (tid < VF_BK * VF_BN / 4) {
                const int k = tid >> 3, nq = (tid & 7) * 4;
                *reinterpret_cast<float4*>(&Ws[k][nq]) =
                    *reinterpret_cast<const float4*>(c.w_f32 + ((int64_t)tap * c.K + k0 + k) * c.N + n0 + nq);
            }
            __syncthreads();
#pragma unroll
            for (int k = 0; k < VF_BK; ++k) {
                const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
                const float4 b = *reinterpret_cast<const float4*>(&Ws[k][tx * 4]);
                const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
            }
            __syncthreads();
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int64_t r = r0 + ty * 4 + i;
        if (r < total) vc_epilogue<4, false>(c.e, r, n0 + tx * 4, acc[i], nullptr);
    }
}

int vc_conv_f32(const VcConvArgs& c, cudaStream_t s) {
    DX_REQUIRE(c.K % 16 == 0 && c.N % 32 == 0, "vc_conv_f32: K=%d N=%d", c.K, c.N);
    DX_REQUIRE(c.e.c_out % 32 == 0 && c.N == c.e.c_out * c.e.up, "vc_conv_f32: N=%d != up=%d * c_out=%d", c.N, c.e.up, c.e.c_out);
    const int64_t m_tiles = (c.rows_bound + VF_BM - 1) / VF_BM;
    DX_REQUIRE(m_tiles > 0 && m_tiles < (1ll << 31), "vc_conv_f32: %lld row tiles", (long long)m_tiles);
    ProfScope prof(c.tag ? c.tag : "vc_conv_f32", s);
    DX_LAUNCH_PDL(vc_conv_f32_kernel, dim3((unsigned)m_tiles, (unsigned)(c.N / VF_BN)), 256, 0, s, c);
    DX_CHECK_LAUNCH();
    return 0;
}

}  // namespace dx
