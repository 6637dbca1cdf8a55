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
template <bool XBF>
struct VcPre {
    float4 res[XBF ? 4 : 8];         // one row's 32 residual channels: bf16 (4 x 16 B) or fp32 (8 x 16 B)
};

__device__ __forceinline__ uint32_t vc_pack2(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

template <bool XBF>
__device__ __forceinline__ void vc_tc_prefetch(const VcEpi& e, int64_t r, int n, bool valid, VcPre<XBF>& pre) {
    if (e.res == nullptr || !valid) return;
    const int ph = n / e.c_out, co = n - ph * e.c_out;
    if constexpr (XBF) {
        const __nv_bfloat16* src = reinterpret_cast<const __nv_bfloat16*>(e.res) + ((int64_t)(co >> 3) * e.x_rows + r * e.up + ph) * 8;
#pragma unroll
        for (int g = 0; g < 4; ++g) pre.res[g] = *reinterpret_cast<const float4*>(src + (int64_t)g * e.x_rows * 8);
    } else {
        const float* src = e.res + ((int64_t)(co >> 2) * e.x_rows + r * e.up + ph) * 4;
#pragma unroll
        for (int g = 0; g < 8; ++g) pre.res[g] = *reinterpret_cast<const float4*>(src + (int64_t)g * e.x_rows * 4);
    }
}

template <bool XBF>
__device__ __forceinline__ void vc_tc_finish(const VcEpi& e, int64_t r, int n, bool valid, float* v, const VcPre<XBF>& pre,
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
            if (e.res) {
                if constexpr (XBF) {                                           // 8 bf16 per float4: group g = half (g & 1) of chunk g / 2
                    const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&pre.res[g >> 1]) + (g & 1) * 2;
                    y.x += __low2float(h2[0]); y.y += __high2float(h2[0]); y.z += __low2float(h2[1]); y.w += __high2float(h2[1]);
                } else {
                    y.x += pre.res[g].x; y.y += pre.res[g].y; y.z += pre.res[g].z; y.w += pre.res[g].w;
                }
            }
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
        if (e.x_out && XBF) {
            __nv_bfloat16* xo = reinterpret_cast<__nv_bfloat16*>(e.x_out) + ((int64_t)(co >> 3) * e.x_rows + ro) * 8;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint4 pk;
                pk.x = vc_pack2(v[8 * c], v[8 * c + 1]); pk.y = vc_pack2(v[8 * c + 2], v[8 * c + 3]);
                pk.z = vc_pack2(v[8 * c + 4], v[8 * c + 5]); pk.w = vc_pack2(v[8 * c + 6], v[8 * c + 7]);
                *reinterpret_cast<uint4*>(xo + (int64_t)c * e.x_rows * 8) = pk;
            }
        } else if (e.x_out) {
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

// (8 epilogue warps for NT >= 64 - two per TMEM lane group, alternating 32-column groups - were measured slower:
// 320 threads x 2 CTAs cap the kernel at 96 registers and the NT = 128 epilogue spills; stage 1 conv2 141 -> 157 us.)
template <int NT, bool XBF>
__global__ void __launch_bounds__(VC_THREADS, NT == 128 ? 2 : 3) vc_conv_tc_kernel(const VcTcParams p) {
    constexpr int EW = 4;                            // epilogue warps: one per TMEM lane group
    constexpr int QS = EW / 4;
    constexpr int NTHREADS = 64 + EW * 32;
    extern __shared__ __align__(1024) uint8_t vc_smem[];
    const uint32_t stage_bytes = p.a_stage_bytes + p.w_stage_bytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(vc_smem + (size_t)p.nstages * stage_bytes);
    uint64_t* empty_bar = full_bar + p.nstages;
    uint64_t* tfull_bar = empty_bar + p.nstages;   // [2]
    uint64_t* tempty_bar = tfull_bar + 2;          // [2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty_bar + 2);
    float* s_bias = reinterpret_cast<float*>(vc_smem + p.par_off);   // [N]: bias of column n = bias[n % c_out]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < p.c.N; i += NTHREADS) s_bias[i] = p.c.e.bias[i % p.c.e.c_out];
    if (threadIdx.x == 0) {
        for (int i = 0; i < p.nstages; ++i) {
            ptx::mbar_init(&full_bar[i], 1);
            ptx::mbar_init(&empty_bar[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            ptx::mbar_init(&tfull_bar[i], 1);
            ptx::mbar_init(&tempty_bar[i], EW);
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
        // ===================== epilogue (warps 2 .. 2+EW-1) =====================
        const int lane_grp = warp & 3;                 // TMEM lanes 32*lane_grp .. +31 (a warp may only touch lanes 32 * (warp % 4) ..)
        const int qoff = (warp - 2) >> 2;              // this warp's 32-column groups: q = qoff, qoff + QS, ... (0 with EW = 4)
        uint32_t tile_it = 0;
        for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x, ++tile_it) {
            const int m_tile = (int)(item / p.n_tiles), n_tile = (int)(item % p.n_tiles);
            const int acc = tile_it & 1;
            const uint32_t acc_ph = (tile_it >> 1) & 1;
            const int64_t r = (int64_t)m_tile * VC_BM + lane_grp * 32 + lane;
            const bool inr = r < total;
            const bool valid = inr && p.c.e.frame_utt[r >> p.c.in_shift] >= 0;   // all phases of a row share its frame
            VcPre<XBF> pre;
            vc_tc_prefetch(p.c.e, r, n_tile * NT + qoff * 32, valid, pre);
            ptx::mbar_wait_sleep(&tfull_bar[acc], acc_ph, 100);
            ptx::tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(lane_grp * 32) << 16) + (uint32_t)acc * NT;
#pragma unroll 1
            for (int q = qoff; q < NT / 32; q += QS) {
                float v[32];
                ptx::tmem_ld32(taddr + q * 32, v);
                const int n = n_tile * NT + q * 32;
                if (q > qoff) vc_tc_prefetch(p.c.e, r, n, valid, pre);
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

template <int NT, bool XBF>
static int vc_tc_launch2(const VcTcParams& p, size_t smem, int grid, cudaStream_t s) {
    static SmemConfigured configured;
    DX_TRY(ensure_dyn_smem(vc_conv_tc_kernel<NT, XBF>, smem, configured));
    ProfScope prof(p.c.tag ? p.c.tag : "vc_conv", s);
    DX_LAUNCH_PDL((vc_conv_tc_kernel<NT, XBF>), grid, VC_THREADS, smem, s, p);
    DX_CHECK_LAUNCH();
    return 0;
}
template <int NT>
static int vc_tc_launch(const VcTcParams& p, size_t smem, int grid, cudaStream_t s) {
    return p.c.e.x_bf16 ? vc_tc_launch2<NT, true>(p, smem, grid, s) : vc_tc_launch2<NT, false>(p, smem, grid, s);
}

int vc_conv_bf16(const VcConvArgs& c, cudaStream_t s) {
    DX_REQUIRE(c.K % 16 == 0 && c.N % 32 == 0 && c.N <= VC_MAX_N, "vc_conv: K=%d N=%d (K %% 16, N %% 32, N <= %d)", c.K, c.N, VC_MAX_N);
    DX_REQUIRE(c.taps % 2 == 1 && (c.taps / 2) * c.dil <= DX_VC_HALO, "vc_conv: taps=%d dil=%d exceed the %d-row halo", c.taps, c.dil, DX_VC_HALO);
    DX_REQUIRE(c.e.c_out % 32 == 0 && c.N == c.e.c_out * c.e.up, "vc_conv: N=%d != up=%d * c_out=%d", c.N, c.e.up, c.e.c_out);
    DX_REQUIRE(c.w_bf != nullptr, "vc_conv: packed bf16 weights missing (tag %s)", c.tag ? c.tag : "?");
    int g_vc_sms = 0;
    DX_TRY(device_sm_count(&g_vc_sms));
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
// fused ResBlock layer: convs1 (dilated) → leaky_relu → convs2 → + x, hidden tile never leaves the SM
// ---------------------------------------------------------------------------------------------------------
// One tile = 128 hidden rows H0 .. H0+127 with H0 = t * VM - hk (hk = taps / 2, VM = 128 - 2 hk valid output rows).
//   M1: D1[b] = convs1 over the staged A tile (rows H0 - hk*dil .. H0 + 127 + hk*dil)
//   P1: D1[b] → + bias1, leaky_relu, zero outside the utterances (= convs2's zero padding) → bf16 → Hs[b] in shared
//       memory, already in the K-major UMMA layout (chunk stride RH rows; rows 128.. stay zero)
//   M2: D2[b] = convs2 with A = Hs[b] (tap = +16 B) and B = the streamed W2 stages
//   P2: D2[b] → the usual epilogue (bias2 + residual, xs bookkeeping, x_out, act_out) for rows m < VM
// Issue order M1(0 .. NB-2), { M1(i+NB-1), M2(i) }: P1(i) runs under the following M1s, P2(i) under the next ones; D1, D2
// and Hs have NB buffers (NB tiles in flight per CTA: the per-tile chain copy → M1 → P1 → M2 → P2 is ~7 µs long, its
// bare barrier skeleton alone 1.5 µs).  Warps: 0 producer, 1 issuer, then EW epilogue warps (4 per 32 columns pair-up: warp w handles TMEM lane
// group w % 4 and the 32-column groups q ≡ w / 4 mod EW / 4).
#define VP_TRACE(j, slot)                                                                            \
    do {                                                                                             \
        if constexpr (TR) {                                                                          \
            if (blockIdx.x == 0 && (j) < 64) p.trace[8 + 16 * (j) + (slot)] = clock64();                 \
        }                                                                                            \
    } while (0)
constexpr int VP_RH = 144;           // rows per chunk of the hidden tile in shared memory (128 + taps - 1 <= 144)

struct VcPairParams {
    VcPairArgs c;
    int nks1, kc1, nks2, kc2, nstages, halo1, ra, hk, vm;
    uint32_t a_stage_bytes, w1_stage_bytes, w2_stage_bytes, stage_bytes, hs_bytes, hs_off, bar_off, par_off;
    int dbg;   // timing experiments (debug knob 11): 1 no P2 body, 2 no A copies, 8 no MMAs, 16 no P1 body
    long long* trace;    // bring-up: clock64 timeline of CTA 0 (dx_set_trace_buffer): 16 slots per tile, see tools/voc_trace.py
    int wres;            // both weight sets stay in shared memory for the whole kernel (<= 50 KB): the ring carries A only
    uint32_t w_off;
};

// TR: with the clock64 timeline of CTA 0 (tools/voc_trace.py); a separate instantiation because the 96-register C = 64
// variant pays ~8 % for the extra live state
template <int C, bool TR, bool XBF>
__global__ void __launch_bounds__(64 + (C >= 128 ? 256 : 128), C == 128 ? 1 : 2) vc_pair_tc_kernel(const VcPairParams p) {
    constexpr int EW = C >= 128 ? 8 : 4;             // epilogue warps
    constexpr int QS = EW / 4;                       // 32-column groups handled in parallel
    constexpr int NTHREADS = 64 + EW * 32;
    constexpr int NB = 2;              // tiles in flight per CTA: D1, D2 and Hs buffers (2 NB C TMEM columns)
    extern __shared__ __align__(1024) uint8_t vp_smem[];
    uint8_t* hs = vp_smem + p.hs_off;                // [NB][C/8][VP_RH][16 B]
    uint64_t* bars = reinterpret_cast<uint64_t*>(vp_smem + p.bar_off);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = full_bar + p.nstages;
    uint64_t* d1_full = empty_bar + p.nstages;       // [NB] each below
    uint64_t* d1_empty = d1_full + NB;
    uint64_t* hs_full = d1_empty + NB;
    uint64_t* hs_empty = hs_full + NB;
    uint64_t* d2_full = hs_empty + NB;
    uint64_t* d2_empty = d2_full + NB;
    uint64_t* w_bar = d2_empty + NB;                 // resident weights have landed (p.wres)
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(w_bar + 1);
    float* s_b1 = reinterpret_cast<float*>(vp_smem + p.par_off);
    float* s_b2 = s_b1 + C;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < C; i += NTHREADS) {
        s_b1[i] = p.c.bias1[i];
        s_b2[i] = p.c.bias2[i];
    }
    for (uint32_t i = threadIdx.x; i < NB * p.hs_bytes / 16; i += NTHREADS) reinterpret_cast<uint4*>(hs)[i] = make_uint4(0, 0, 0, 0);
    if (threadIdx.x == 0) {
        for (int i = 0; i < p.nstages; ++i) {
            ptx::mbar_init(&full_bar[i], 1);
            ptx::mbar_init(&empty_bar[i], 1);
        }
        ptx::mbar_init(w_bar, 1);
        for (int i = 0; i < NB; ++i) {
            ptx::mbar_init(&d1_full[i], 1);
            ptx::mbar_init(&d1_empty[i], EW);
            ptx::mbar_init(&hs_full[i], EW);
            ptx::mbar_init(&hs_empty[i], 1);
            ptx::mbar_init(&d2_full[i], 1);
            ptx::mbar_init(&d2_empty[i], EW);
        }
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc<2 * NB * C>(tmem_ptr);
    ptx::fence_proxy_async();                        // the zeroed pad rows of Hs are read by the MMA (async proxy)
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;            // columns: D1[0..NB) D2[0..NB)
    if (p.wres && warp == 0 && ptx::elect_one()) {   // weights do not depend on the previous kernel: fetch them under its tail
        const uint32_t b1 = (uint32_t)p.nks1 * p.w1_stage_bytes, b2 = (uint32_t)p.nks2 * p.w2_stage_bytes;
        ptx::mbar_expect_tx(w_bar, b1 + b2);
        ptx::bulk_g2s(vp_smem + p.w_off, p.c.w1_bf, b1, w_bar);
        ptx::bulk_g2s(vp_smem + p.w_off + b1, p.c.w2_bf, b2, w_bar);
    }
    ptx::griddep_launch();
    ptx::griddep_wait();
    if (TR && blockIdx.x == 0 && threadIdx.x == 0) p.trace[0] = clock64();
    const int64_t total = (int64_t)(*p.c.total_frames) << p.c.shift;
    const int64_t m_tiles = (total + p.vm - 1) / p.vm;
    // tiles of this CTA: t = blockIdx.x + j * gridDim.x, j = 0 .. n_mine-1
    const int64_t n_mine = m_tiles > blockIdx.x ? (m_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const uint32_t a_rows_bytes = (uint32_t)(VC_BM + 2 * p.halo1) * 16;

    if (warp == 0) {
        // ===================== producer: stages in the issuer's order M1(0), { M1(j+1), M2(j) } =====================
        if (ptx::elect_one()) {
            uint32_t it = 0;
            auto load1 = [&](int64_t j) {
                const int64_t h0 = (blockIdx.x + j * gridDim.x) * (int64_t)p.vm - p.hk;
                for (int ksg = 0; ksg < p.nks1; ++ksg, ++it) {
                    const int s = it % p.nstages;
                    ptx::mbar_wait_sleep(&empty_bar[s], ((it / p.nstages) & 1) ^ 1, 64);
                    if (ksg == 0) VP_TRACE(j, 0);
                    uint8_t* sa = vp_smem + (size_t)s * p.stage_bytes;
                    ptx::mbar_expect_tx(&full_bar[s], ((p.dbg & 2) ? 0u : (uint32_t)p.kc1 * a_rows_bytes) + (p.wres ? 0u : p.w1_stage_bytes));
                    for (int c = 0; c < ((p.dbg & 2) ? 0 : p.kc1); ++c)
                        ptx::bulk_g2s(sa + (size_t)c * p.ra * 16,
                                      p.c.a + ((int64_t)(ksg * p.kc1 + c) * p.c.a_rows_phys + DX_VC_HALO + h0 - p.halo1) * 8,
                                      a_rows_bytes, &full_bar[s]);
                    if (!p.wres)
                        ptx::bulk_g2s(sa + p.a_stage_bytes, reinterpret_cast<const uint8_t*>(p.c.w1_bf) + (int64_t)ksg * p.w1_stage_bytes,
                                      p.w1_stage_bytes, &full_bar[s]);
                }
            };
            auto load2 = [&]() {
                for (int ksg = 0; ksg < (p.wres ? 0 : p.nks2); ++ksg, ++it) {
                    const int s = it % p.nstages;
                    ptx::mbar_wait_sleep(&empty_bar[s], ((it / p.nstages) & 1) ^ 1, 64);
                    uint8_t* sa = vp_smem + (size_t)s * p.stage_bytes;
                    ptx::mbar_expect_tx(&full_bar[s], p.w2_stage_bytes);
                    ptx::bulk_g2s(sa + p.a_stage_bytes, reinterpret_cast<const uint8_t*>(p.c.w2_bf) + (int64_t)ksg * p.w2_stage_bytes,
                                  p.w2_stage_bytes, &full_bar[s]);
                }
            };
            for (int64_t j = 0; j < NB - 1 && j < n_mine; ++j) load1(j);
            for (int64_t j = 0; j < n_mine; ++j) {
                if (j + NB - 1 < n_mine) load1(j + NB - 1);
                load2();
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        constexpr uint32_t idesc = ptx::make_idesc_bf16(VC_BM, C);
        const uint32_t a_lbo = (uint32_t)p.ra * 16, b_lbo = C * 16, h_lbo = VP_RH * 16;
        uint32_t it = 0;
        const uint32_t wres1 = ptx::smem_u32(vp_smem + p.w_off), wres2 = wres1 + (uint32_t)p.nks1 * p.w1_stage_bytes;
        if (p.wres) ptx::mbar_wait_sleep(w_bar, 0, 32);
        auto mma1 = [&](int64_t j) {
            const int b = (int)(j % NB);
            ptx::mbar_wait_sleep(&d1_empty[b], ((uint32_t)(j / NB) & 1) ^ 1, 64);
            ptx::tc_fence_after();
            if (lane == 0) VP_TRACE(j, 2);
            const uint32_t d_tmem = tmem_base + (uint32_t)b * C;
            for (int ksg = 0; ksg < p.nks1; ++ksg, ++it) {
                const int s = it % p.nstages;
                ptx::mbar_wait_sleep(&full_bar[s], (it / p.nstages) & 1, 32);
                ptx::tc_fence_after();
                if (ksg == 0 && lane == 0) VP_TRACE(j, 3);
                if (ptx::elect_one()) {
                    const uint32_t sa = ptx::smem_u32(vp_smem + (size_t)s * p.stage_bytes);
                    const uint32_t sw = p.wres ? wres1 + (uint32_t)ksg * p.w1_stage_bytes : sa + p.a_stage_bytes;
                    for (int tap = 0; tap < ((p.dbg & 8) ? 0 : p.c.taps); ++tap)
                        for (int kk = 0; kk < p.kc1 / 2; ++kk)
                            ptx::mma_bf16(d_tmem, ptx::make_desc(sa + (uint32_t)(tap * p.c.dil) * 16 + (uint32_t)kk * 2 * a_lbo, a_lbo, 128),
                                          ptx::make_desc(sw + (uint32_t)(tap * p.kc1 + kk * 2) * b_lbo, b_lbo, 128), idesc, (ksg | tap | kk) != 0);
                    ptx::mma_commit(&empty_bar[s]);
                    if (ksg == p.nks1 - 1) ptx::mma_commit(&d1_full[b]);
                }
                __syncwarp();
            }
            if (lane == 0) VP_TRACE(j, 4);
        };
        auto mma2 = [&](int64_t j) {
            const int b = (int)(j % NB);
            const uint32_t ph = (uint32_t)(j / NB) & 1;
            ptx::mbar_wait_sleep(&hs_full[b], ph, 32);
            if (lane == 0) VP_TRACE(j, 5);
            ptx::mbar_wait_sleep(&d2_empty[b], ph ^ 1, 64);
            ptx::tc_fence_after();
            if (lane == 0) VP_TRACE(j, 6);
            const uint32_t d_tmem = tmem_base + (uint32_t)(NB + b) * C;
            const uint32_t sh = ptx::smem_u32(hs + (size_t)b * p.hs_bytes);
            for (int ksg = 0; ksg < p.nks2; ++ksg) {
                int s = 0;
                if (!p.wres) {
                    s = it % p.nstages;
                    ptx::mbar_wait_sleep(&full_bar[s], (it / p.nstages) & 1, 32);
                    ptx::tc_fence_after();
                    ++it;
                }
                if (ptx::elect_one()) {
                    const uint32_t sw = p.wres ? wres2 + (uint32_t)ksg * p.w2_stage_bytes
                                               : ptx::smem_u32(vp_smem + (size_t)s * p.stage_bytes) + p.a_stage_bytes;
                    for (int tap = 0; tap < ((p.dbg & 8) ? 0 : p.c.taps); ++tap)
                        for (int kk = 0; kk < p.kc2 / 2; ++kk)
                            ptx::mma_bf16(d_tmem, ptx::make_desc(sh + (uint32_t)tap * 16 + (uint32_t)(ksg * p.kc2 + kk * 2) * h_lbo, h_lbo, 128),
                                          ptx::make_desc(sw + (uint32_t)(tap * p.kc2 + kk * 2) * b_lbo, b_lbo, 128), idesc, (ksg | tap | kk) != 0);
                    if (!p.wres) ptx::mma_commit(&empty_bar[s]);
                    if (ksg == p.nks2 - 1) {
                        ptx::mma_commit(&d2_full[b]);
                        ptx::mma_commit(&hs_empty[b]);
                    }
                }
                __syncwarp();
            }
            if (lane == 0) VP_TRACE(j, 7);
        };
        for (int64_t j = 0; j < NB - 1 && j < n_mine; ++j) mma1(j);
        for (int64_t j = 0; j < n_mine; ++j) {
            if (j + NB - 1 < n_mine) mma1(j + NB - 1);
            mma2(j);
        }
    } else {
        // ===================== epilogue warps =====================
        const int ew = warp - 2;
        const int lane_grp = warp & 3, qoff = ew >> 2;        // a warp may only touch TMEM lanes 32 * (warp % 4) .. +31
        const int m = lane_grp * 32 + lane;                                   // tile row (TMEM lane)
        const uint32_t lane_addr = tmem_base + ((uint32_t)(lane_grp * 32) << 16);
        // row validity comes from frame_utt (a dependent global load, ~1 k cycles even on an L2 hit): every lookup is
        // issued one loop iteration before it is needed
        // (the looked-up utterance id stays an int until it is used: comparing it at once would wait for the load)
        auto row_utt = [&](int64_t row) -> int { return (row >= 0 && row < total) ? p.c.e.frame_utt[row >> p.c.shift] : -1; };
        auto hrow = [&](int64_t j) -> int64_t { return (blockIdx.x + j * gridDim.x) * (int64_t)p.vm - p.hk + m; };
        auto orow = [&](int64_t j) -> int64_t { return (blockIdx.x + j * gridDim.x) * (int64_t)p.vm + m; };
        auto p1 = [&](int64_t j, bool hv) {
            const int b = (int)(j % NB);
            const uint32_t ph = (uint32_t)(j / NB) & 1;
            if (threadIdx.x == 64) VP_TRACE(j, 8);
            ptx::mbar_wait_sleep(&d1_full[b], ph, 64);
            ptx::tc_fence_after();
            if (threadIdx.x == 64) VP_TRACE(j, 9);
            ptx::mbar_wait_sleep(&hs_empty[b], ph ^ 1, 64);                    // M2(j - 2) has finished reading Hs[b]
            uint8_t* dst = hs + (size_t)b * p.hs_bytes + (size_t)m * 16;
#pragma unroll 1
            for (int q = qoff; q < ((p.dbg & 16) ? 0 : C / 32); q += QS) {
                float v[32];
                ptx::tmem_ld32(lane_addr + (uint32_t)b * C + q * 32, v);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint4 pk;
                    uint32_t* w = reinterpret_cast<uint32_t*>(&pk);
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float x0 = vc_lrelu(v[c * 8 + 2 * e] + s_b1[q * 32 + c * 8 + 2 * e]);
                        const float x1 = vc_lrelu(v[c * 8 + 2 * e + 1] + s_b1[q * 32 + c * 8 + 2 * e + 1]);
                        w[e] = hv ? vc_pack2(x0, x1) : 0u;
                    }
                    *reinterpret_cast<uint4*>(dst + (size_t)(q * 4 + c) * VP_RH * 16) = pk;
                }
            }
            ptx::tc_fence_before();
            ptx::fence_proxy_async();                                          // Hs stores → visible to the MMA
            __syncwarp();
            if (lane == 0) {
                ptx::mbar_arrive(&d1_empty[b]);
                ptx::mbar_arrive(&hs_full[b]);
            }
            if (threadIdx.x == 64) VP_TRACE(j, 10);
        };
        // P2 in two parts: the residual loads of tile j are requested BEFORE P1(j + NB - 1) runs, so their DRAM latency
        // (the timeline showed ~2.5 k of P2's 4.2 k cycles waiting for them) hides under that P1
        VcPre<XBF> pre;
        auto p2 = [&](int64_t j, bool valid) {
            const int b = (int)(j % NB);
            const uint32_t ph = (uint32_t)(j / NB) & 1;
            const int64_t r = orow(j);
            const bool inr = m < p.vm && r < total;
            if (threadIdx.x == 64) VP_TRACE(j, 11);
            ptx::mbar_wait_sleep(&d2_full[b], ph, 64);
            ptx::tc_fence_after();
            if (threadIdx.x == 64) VP_TRACE(j, 12);
#pragma unroll 1
            for (int q = qoff; q < C / 32; q += QS) {
                float v[32];
                ptx::tmem_ld32(lane_addr + (uint32_t)(NB + b) * C + q * 32, v);
                if (q > qoff) vc_tc_prefetch(p.c.e, r, q * 32, valid, pre);
                ptx::tmem_ld_wait();
                if (inr && !(p.dbg & 1)) vc_tc_finish(p.c.e, r, q * 32, valid, v, pre, s_b2 + q * 32);
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&d2_empty[b]);
            if (threadIdx.x == 64) VP_TRACE(j, 13);
        };
        static_assert(NB == 2, "the epilogue loop below is written for two tiles in flight");
        if constexpr (C <= 64) {
            // software-pipelined epilogue (163 registers available at 192 threads x 2 CTAs): per iteration the residual of
            // tile j and the validity lookups of the NEXT iteration are requested first, then P1(j + 1), then P2(j)
            int hu_next = n_mine > 1 ? row_utt(hrow(1)) : -1;                      // for P1(1)
            int vu_cur = n_mine > 0 ? row_utt(orow(0)) : -1;                       // for P2(0)
            if (n_mine > 0) p1(0, row_utt(hrow(0)) >= 0);
            for (int64_t j = 0; j < n_mine; ++j) {
                const bool valid = m < p.vm && vu_cur >= 0;
                vc_tc_prefetch(p.c.e, orow(j), qoff * 32, valid, pre);
                const int hu_after = j + 2 < n_mine ? row_utt(hrow(j + 2)) : -1;   // consumed in the next iteration
                const int vu_next = j + 1 < n_mine ? row_utt(orow(j + 1)) : -1;
                if (j + 1 < n_mine) p1(j + 1, hu_next >= 0);
                p2(j, valid);
                hu_next = hu_after;
                vu_cur = vu_next;
            }
        } else {
            // (8 epilogue warps x 2 CTAs cap a kernel at 96 registers: the pipelined loop then spills and is slower;
            // measured on the 64-channel variant, stage 2: 2.10 -> 2.32 ms)
            if (n_mine > 0) p1(0, row_utt(hrow(0)) >= 0);
            for (int64_t j = 0; j < n_mine; ++j) {
                if (j + 1 < n_mine) p1(j + 1, row_utt(hrow(j + 1)) >= 0);
                const bool valid = m < p.vm && row_utt(orow(j)) >= 0;
                vc_tc_prefetch(p.c.e, orow(j), qoff * 32, valid, pre);
                p2(j, valid);
            }
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (TR && blockIdx.x == 0 && threadIdx.x == 0) {
        p.trace[1] = clock64();
        p.trace[2] = n_mine;
    }
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<2 * NB * C>(tmem_base);
    }
}

template <int C, bool TR, bool XBF>
static int vc_pair_launch3(const VcPairParams& p, size_t smem, int grid, cudaStream_t s) {
    static SmemConfigured configured;
    DX_TRY(ensure_dyn_smem(vc_pair_tc_kernel<C, TR, XBF>, smem, configured));
    ProfScope prof(p.c.tag ? p.c.tag : "vc_pair", s);
    DX_LAUNCH_PDL((vc_pair_tc_kernel<C, TR, XBF>), grid, 64 + (C >= 128 ? 256 : 128), smem, s, p);
    DX_CHECK_LAUNCH();
    return 0;
}
template <int C>
static int vc_pair_launch(const VcPairParams& p, size_t smem, int grid, cudaStream_t s) {
    if (p.trace != nullptr)
        return p.c.e.x_bf16 ? vc_pair_launch3<C, true, true>(p, smem, grid, s) : vc_pair_launch3<C, true, false>(p, smem, grid, s);
    return p.c.e.x_bf16 ? vc_pair_launch3<C, false, true>(p, smem, grid, s) : vc_pair_launch3<C, false, false>(p, smem, grid, s);
}

int vc_pair_bf16(const VcPairArgs& c, cudaStream_t s) {
    DX_REQUIRE(c.C == 32 || c.C == 64 || c.C == 128, "vc_pair: C=%d (32, 64 or 128)", c.C);
    DX_REQUIRE(c.taps % 2 == 1 && c.taps <= 17 && (c.taps / 2) * c.dil <= DX_VC_HALO - c.taps / 2, "vc_pair: taps=%d dil=%d", c.taps, c.dil);
    DX_REQUIRE(c.w1_bf && c.w2_bf && c.bias1 && c.bias2, "vc_pair: weights missing (tag %s)", c.tag ? c.tag : "?");
    DX_REQUIRE(c.e.up == 1 && c.e.c_out == c.C, "vc_pair: epilogue must be up = 1, c_out = C");
    int g_vc_sms = 0;
    DX_TRY(device_sm_count(&g_vc_sms));
    VcPairParams p;
    p.c = c;
    p.dbg = debug_knob(11);
    p.trace = reinterpret_cast<long long*>(trace_buffer((8 + 16 * 64) * sizeof(long long)));
    if (debug_knob(10) != 0 && (debug_knob(10) != c.C || (debug_knob(15) != 0 && debug_knob(15) != c.taps))) p.trace = nullptr;   // trace filter: C, taps
    int nt1, ks1, nt2, ks2;
    vc_conv_tiling(c.taps, c.dil, c.C, c.C, &nt1, &ks1);      // how the weights were packed (vocoder.py)
    vc_conv_tiling(c.taps, 1, c.C, c.C, &nt2, &ks2);
    DX_REQUIRE(nt1 == c.C && nt2 == c.C, "vc_pair: tile width %d / %d != C", nt1, nt2);
    p.nks1 = c.C / ks1; p.kc1 = ks1 / 8;
    p.nks2 = c.C / ks2; p.kc2 = ks2 / 8;
    p.hk = c.taps / 2;
    p.vm = VC_BM - 2 * p.hk;
    p.halo1 = p.hk * c.dil;
    p.ra = (int)round_up(VC_BM + 2 * p.halo1, 8);
    p.a_stage_bytes = (uint32_t)p.kc1 * p.ra * 16;
    p.w1_stage_bytes = (uint32_t)c.taps * p.kc1 * c.C * 16;
    p.w2_stage_bytes = (uint32_t)c.taps * p.kc2 * c.C * 16;
    const uint32_t w_total = (uint32_t)p.nks1 * p.w1_stage_bytes + (uint32_t)p.nks2 * p.w2_stage_bytes;
    p.wres = (w_total <= 50 * 1024 && !debug_knob(14)) ? 1 : 0;
    p.stage_bytes = p.a_stage_bytes + (p.wres ? 0u : (p.w1_stage_bytes > p.w2_stage_bytes ? p.w1_stage_bytes : p.w2_stage_bytes));
    p.hs_bytes = (uint32_t)(c.C / 8) * VP_RH * 16;
    const int nb = 2;                                       // = NB of the kernel
    const size_t budget = (c.C == 128 ? 216 : 112) * 1024 - nb * (size_t)p.hs_bytes - 2048 - (p.wres ? w_total : 0);
    int nst = (int)(budget / p.stage_bytes);
    if (nst > (p.wres ? 6 : 4)) nst = p.wres ? 6 : 4;
    DX_REQUIRE(nst >= 2, "vc_pair: stage of %u bytes does not allow 2 stages", p.stage_bytes);
    p.nstages = nst;
    p.hs_off = (uint32_t)nst * p.stage_bytes;
    p.w_off = p.hs_off + nb * p.hs_bytes;
    p.bar_off = p.w_off + (p.wres ? w_total : 0);
    p.par_off = (uint32_t)round_up((int64_t)(p.bar_off + (2 * nst + 6 * nb + 1) * sizeof(uint64_t) + 16), 16);
    const size_t smem = p.par_off + 2 * (size_t)c.C * sizeof(float);
    const int per_sm = c.C == 128 ? 1 : 2;
    const int64_t items = (c.rows_bound + p.vm - 1) / p.vm;
    int grid = (int)(items < (int64_t)g_vc_sms * per_sm ? items : (int64_t)g_vc_sms * per_sm);
    if (grid < 1) grid = 1;
    if (c.C == 128) return vc_pair_launch<128>(p, smem, grid, s);
    if (c.C == 64) return vc_pair_launch<64>(p, smem, grid, s);
    return vc_pair_launch<32>(p, smem, grid, s);
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
