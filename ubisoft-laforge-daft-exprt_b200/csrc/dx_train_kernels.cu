// Backward (training) kernels of the fp32 path — BASELINE config 5: the reference's training step is
// `outputs = model(inputs); loss = criterion(...); loss.backward(); optimizer.step()` (train.py:421-445) with torch
// autograd doing the backward of every module of model.py.  These are the hand-written gradients of the same
// modules on the packed row stream ([rows][C] fp32, see dx_f32.cu): every kernel is deterministic (fixed summation
// order, two-stage reductions, no atomics), so a step is reproducible bit for bit and the data-parallel ranks stay in
// lock step after the gradient all-reduce.
//
//   ln_bwd_rows      nn.LayerNorm backward (+ FiLM scale, + pad / liveness mask)       model.py:191,234-235,256-258,655-669
//   film_bwd         d(gamma_f), d(beta_f) of FiLM per utterance                        model.py:228-233
//   colsum_rows      bias gradients (sum over rows)                                    nn.Conv1d / nn.Linear bias
//   wgrad_f32        dW[tap][k][n] = sum_r A[r + tap - half][k] * G[r][n]              nn.Conv1d / nn.Linear weight
//   (dgrad is gemm_f32 of dx_f32.cu with the tap-flipped, transposed weight: transpose_weight)
//   relu_bwd         ReLU mask                                                          model.py:229,652-667
//   attn_bwd_*       nn.MultiheadAttention backward, recomputing P from Q, K and the saved log-sum-exp (no N x N tensor)
//   lenreg_bwd_*     GaussianUpsamplingModule backward (weights, range predictor, projections)   model.py:417-510
//   embed / scalar-conv / masked-mean / style / loss kernels                            model.py:597-604,640-646,714,764-830; loss.py:57-159
#include <math.h>
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_ptx.cuh"
#include "dx_train.cuh"

namespace dx {

// =========================================================================================================
// ln_bwd_rows<C>.  Forward (ln_rows_f32): y = live ? film_g * (g * xh + b) + film_b : 0,  xh = (v - mean) * rstd,
// v = in + res.  Given dy = dy1 + dy2:  dz = dy * film_g;  dg += dz * xh;  db += dz;  dxh = dz * g;
// dv = rstd * (dxh - mean(dxh) - xh * mean(dxh * xh));  non-live rows: dv = 0.
// One CTA = 8 warps = LNB_ROWS rows; per-lane accumulators of dg / db, combined warp by warp in a fixed order into
// part[cta][2][C].
// =========================================================================================================
constexpr int LNB_ROWS = 64;

template <int C, bool DROP>
__global__ void __launch_bounds__(256)
ln_bwd_rows_kernel(const float* __restrict__ in, const float* __restrict__ res, const float* __restrict__ g,
                   const float* __restrict__ film, int film_stride, const float* __restrict__ dy1,
                   const float* __restrict__ dy2, Rows R, int live_ext, float* __restrict__ dv,
                   float* __restrict__ part, LnDrop drop) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    constexpr int V = C / 128;
    __shared__ __align__(16) float s_acc[2 * C];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t total = R.row_start[R.B];
    float4 ag[V], ab[V];
#pragma unroll
    for (int i = 0; i < V; ++i) ag[i] = ab[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int k = 0; k < LNB_ROWS / 8; ++k) {
        const int64_t r = (int64_t)blockIdx.x * LNB_ROWS + k * 8 + warp;
        if (r >= R.cap) break;
        float4* orow = reinterpret_cast<float4*>(dv + r * C);
        const bool live = (r < total) && row_live(R, r, live_ext);
        if (!live) {
#pragma unroll
            for (int i = 0; i < V; ++i) orow[lane + 32 * i] = make_float4(0.f, 0.f, 0.f, 0.f);
            if constexpr (DROP) {
                if (drop.where == LN_DROP_IN) {
#pragma unroll
                    for (int i = 0; i < V; ++i) reinterpret_cast<float4*>(drop.dvm + r * C)[lane + 32 * i] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
            continue;
        }
        const int b_idx = R.row_b[r];
        const int pos = DROP ? R.row_p[r] : 0;
        float4 v[V];
        float sum = 0.f;
#pragma unroll
        for (int i = 0; i < V; ++i) {
            v[i] = reinterpret_cast<const float4*>(in + r * C)[lane + 32 * i];
            if (res != nullptr) {
                const float4 q = reinterpret_cast<const float4*>(res + r * C)[lane + 32 * i];
                v[i].x += q.x; v[i].y += q.y; v[i].z += q.z; v[i].w += q.w;
            }
            sum += (v[i].x + v[i].y) + (v[i].z + v[i].w);
        }
        const float mean = warp_sum(sum) * (1.f / C);
        float sq = 0.f;
#pragma unroll
        for (int i = 0; i < V; ++i) {
            const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
            sq += (a * a + b * b) + (c * c + d * d);
        }
        const float rstd = rsqrtf(warp_sum(sq) * (1.f / C) + 1e-5f);
        const float* fp = film ? film + (int64_t)b_idx * film_stride : nullptr;
        float4 dxh[V];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int i = 0; i < V; ++i) {
            const int c4 = lane + 32 * i;
            float4 d = reinterpret_cast<const float4*>(dy1 + r * C)[c4];
            if (dy2 != nullptr) {
                const float4 e = reinterpret_cast<const float4*>(dy2 + r * C)[c4];
                d.x += e.x; d.y += e.y; d.z += e.z; d.w += e.w;
            }
            if constexpr (DROP) {           // the forward dropped this LayerNorm's OUTPUT: mask the incoming gradient
                if (drop.where == LN_DROP_OUT) d = drop_apply4(d, drop.d, drop.site, b_idx, pos, R.n_max, C, c4 * 4);
            }
            if (fp != nullptr) {
                const float4 fg = reinterpret_cast<const float4*>(fp)[c4];
                d.x *= fg.x; d.y *= fg.y; d.z *= fg.z; d.w *= fg.w;
            }
            const float4 xh = make_float4((v[i].x - mean) * rstd, (v[i].y - mean) * rstd, (v[i].z - mean) * rstd, (v[i].w - mean) * rstd);
            ag[i].x += d.x * xh.x; ag[i].y += d.y * xh.y; ag[i].z += d.z * xh.z; ag[i].w += d.w * xh.w;
            ab[i].x += d.x; ab[i].y += d.y; ab[i].z += d.z; ab[i].w += d.w;
            const float4 gg = reinterpret_cast<const float4*>(g)[c4];
            dxh[i] = make_float4(d.x * gg.x, d.y * gg.y, d.z * gg.z, d.w * gg.w);
            s1 += (dxh[i].x + dxh[i].y) + (dxh[i].z + dxh[i].w);
            s2 += (dxh[i].x * xh.x + dxh[i].y * xh.y) + (dxh[i].z * xh.z + dxh[i].w * xh.w);
            v[i] = xh;
        }
        const float m1 = warp_sum(s1) * (1.f / C), m2 = warp_sum(s2) * (1.f / C);
#pragma unroll
        for (int i = 0; i < V; ++i) {
            const float4 o = make_float4(rstd * (dxh[i].x - m1 - v[i].x * m2), rstd * (dxh[i].y - m1 - v[i].y * m2),
                                         rstd * (dxh[i].z - m1 - v[i].z * m2), rstd * (dxh[i].w - m1 - v[i].w * m2));
            orow[lane + 32 * i] = o;
            if constexpr (DROP) {           // the forward dropped the branch `in`: its gradient is the masked copy
                if (drop.where == LN_DROP_IN)
                    reinterpret_cast<float4*>(drop.dvm + r * C)[lane + 32 * i] =
                        drop_apply4(o, drop.d, drop.site, b_idx, pos, R.n_max, C, (lane + 32 * i) * 4);
            }
        }
    }
    // ordered combination of the 8 warps' partial sums
    for (int w = 0; w < 8; ++w) {
        if (warp == w) {
#pragma unroll
            for (int i = 0; i < V; ++i) {
                float4* sg = reinterpret_cast<float4*>(s_acc) + lane + 32 * i;
                float4* sb = reinterpret_cast<float4*>(s_acc + C) + lane + 32 * i;
                if (w == 0) { *sg = ag[i]; *sb = ab[i]; }
                else {
                    float4 a = *sg, b = *sb;
                    a.x += ag[i].x; a.y += ag[i].y; a.z += ag[i].z; a.w += ag[i].w;
                    b.x += ab[i].x; b.y += ab[i].y; b.z += ab[i].z; b.w += ab[i].w;
                    *sg = a; *sb = b;
                }
            }
        }
        __syncthreads();
    }
    for (int i = threadIdx.x; i < 2 * C; i += 256) part[(int64_t)blockIdx.x * 2 * C + i] = s_acc[i];
}

// part [S][n] -> out[n] = scale * sum_s part[s][n] (s ascending); S_valid chunks carry data when `total_rows` is given.
// With out2: columns >= split go to out2[i - split] (two gradients reduced by one launch).
__global__ void reduce_partials_kernel(const float* __restrict__ part, int S, int64_t n, float* __restrict__ out,
                                       const int32_t* __restrict__ total_rows, int chunk_rows, float scale,
                                       float* __restrict__ out2, int64_t split) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int s_lim = S;
    if (total_rows != nullptr) {
        const int v = (int)((*total_rows + chunk_rows - 1) / chunk_rows);
        s_lim = v < S ? v : S;
    }
    float acc = 0.f;
    for (int s = 0; s < s_lim; ++s) acc += part[(int64_t)s * n + i];
    if (out2 != nullptr && i >= split) out2[i - split] = acc * scale;
    else out[i] = acc * scale;
}

// Same result layout, for FEW columns and MANY partials (bias sums, LayerNorm gains: n <= a few thousand, S in the hundreds): a
// thread per column would walk all S partials alone (one CTA busy, S dependent steps).  Here 16 lanes share a column (lane l adds
// partials l, l + 16, ... in ascending order) and lane 0 adds the 16 lane sums in lane order: a different, still fixed, summation
// order.
constexpr int RP_LANES = 16;
__global__ void __launch_bounds__(32 * RP_LANES)
reduce_partials_lanes_kernel(const float* __restrict__ part, int S, int64_t n, float* __restrict__ out,
                             const int32_t* __restrict__ total_rows, int chunk_rows, float scale, float* __restrict__ out2,
                             int64_t split) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    __shared__ float s_acc[RP_LANES][33];
    const int c = threadIdx.x & 31, l = threadIdx.x >> 5;
    const int64_t i = (int64_t)blockIdx.x * 32 + c;
    int s_lim = S;
    if (total_rows != nullptr) {
        const int v = (int)((*total_rows + chunk_rows - 1) / chunk_rows);
        s_lim = v < S ? v : S;
    }
    float acc = 0.f;
    if (i < n) {
#pragma unroll 4
        for (int s = l; s < s_lim; s += RP_LANES) acc += part[(int64_t)s * n + i];
    }
    s_acc[l][c] = acc;
    __syncthreads();
    if (l == 0 && i < n) {
        float t = s_acc[0][c];
#pragma unroll
        for (int k = 1; k < RP_LANES; ++k) t += s_acc[k][c];
        if (out2 != nullptr && i >= split) out2[i - split] = t * scale;
        else out[i] = t * scale;
    }
}

int reduce_partials(const float* part, int S, int64_t n, float* out, const int32_t* total_rows, int chunk_rows, float scale,
                    cudaStream_t s, float* out2, int64_t split) {
    if (n <= 8192 && S >= 2 * RP_LANES) {
        DX_LAUNCH_PDL((reduce_partials_lanes_kernel), (unsigned)((n + 31) / 32), 32 * RP_LANES, 0, s, part, S, n, out, total_rows,
                      chunk_rows, scale, out2, split);
        DX_CHECK_LAUNCH();
        return 0;
    }
    DX_LAUNCH_PDL((reduce_partials_kernel), (unsigned)((n + 255) / 256), 256, 0, s, part, S, n, out, total_rows, chunk_rows, scale,
                  out2, split);
    DX_CHECK_LAUNCH();
    return 0;
}

int ln_bwd_rows(const float* in, const float* res, const float* g, int C, const float* film, int film_stride, const float* dy1,
                const float* dy2, const Rows& R, int live_ext, float* dv, float* dg, float* db, float* scratch, cudaStream_t s,
                const LnDrop* drop) {
    const int S = (int)((R.cap + LNB_ROWS - 1) / LNB_ROWS);
    ProfScope prof("ln_bwd", s);
    LnDrop nd;
    memset(&nd, 0, sizeof(nd));
    if (drop != nullptr && drop->d.thr != 0) {
        DX_REQUIRE(drop->where == LN_DROP_OUT || (drop->dvm != nullptr && drop->dvm != dv && drop->dvm != dy1 && drop->dvm != dy2),
                   "ln_bwd_rows: the masked gradient needs its own buffer");
        if (C == 128)
            DX_LAUNCH_PDL((ln_bwd_rows_kernel<128, true>), S, 256, 0, s, in, res, g, film, film_stride, dy1, dy2, R, live_ext, dv, scratch, *drop);
        else if (C == 1024)
            DX_LAUNCH_PDL((ln_bwd_rows_kernel<1024, true>), S, 256, 0, s, in, res, g, film, film_stride, dy1, dy2, R, live_ext, dv, scratch, *drop);
        else
            DX_REQUIRE(false, "ln_bwd_rows: unsupported width %d", C);
    } else if (C == 128)
        DX_LAUNCH_PDL((ln_bwd_rows_kernel<128, false>), S, 256, 0, s, in, res, g, film, film_stride, dy1, dy2, R, live_ext, dv, scratch, nd);
    else if (C == 1024)
        DX_LAUNCH_PDL((ln_bwd_rows_kernel<1024, false>), S, 256, 0, s, in, res, g, film, film_stride, dy1, dy2, R, live_ext, dv, scratch, nd);
    else
        DX_REQUIRE(false, "ln_bwd_rows: unsupported width %d", C);
    DX_CHECK_LAUNCH();
    // part[s] = [dg(C) | db(C)]: one reduction over 2C columns, split between the two gradients
    return reduce_partials(scratch, S, 2 * C, dg, R.row_start + R.B, LNB_ROWS, 1.f, s, db, C);
}
int64_t ln_bwd_scratch_floats(int64_t cap, int C) { return ((cap + LNB_ROWS - 1) / LNB_ROWS + 1) * 2 * C; }

// =========================================================================================================
// film_bwd (C = 128): dfilm[b] = [ sum_rows dy * z | sum_rows dy ],  z = g * xh + b (the LayerNorm output FiLM scales).
// CTA (b, y) takes the utterance's rows y * 8 + w, + 8 FB_SLICES, ... (w = warp) and writes one partial [256]; the warps are
// combined in warp order and the FB_SLICES partials in slice order (film_bwd_finish): a fixed summation order.  (One CTA per
// utterance - 32 CTAs on 148 SMs, each walking up to 750 rows - took 67 us per launch.)
// =========================================================================================================
constexpr int FB_SLICES = 8;

__global__ void __launch_bounds__(256)
film_bwd_kernel(const float* __restrict__ in, const float* __restrict__ res, const float* __restrict__ g,
                const float* __restrict__ bta, const float* __restrict__ dy1, const float* __restrict__ dy2, Rows R,
                float* __restrict__ part) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    __shared__ __align__(16) float s_acc[256];
    const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int len = (int)min((long long)R.len[b], (long long)R.n_max);
    const int64_t rs = R.row_start[b];
    const float4 gg = reinterpret_cast<const float4*>(g)[lane], bb = reinterpret_cast<const float4*>(bta)[lane];
    float4 ag = make_float4(0.f, 0.f, 0.f, 0.f), ab = ag;
    for (int p = blockIdx.y * 8 + warp; p < len; p += 8 * FB_SLICES) {
        const int64_t r = rs + p;
        float4 v = reinterpret_cast<const float4*>(in + r * DX_D)[lane];
        if (res != nullptr) {
            const float4 q = reinterpret_cast<const float4*>(res + r * DX_D)[lane];
            v.x += q.x; v.y += q.y; v.z += q.z; v.w += q.w;
        }
        const float mean = warp_sum((v.x + v.y) + (v.z + v.w)) * (1.f / DX_D);
        const float a0 = v.x - mean, a1 = v.y - mean, a2 = v.z - mean, a3 = v.w - mean;
        const float rstd = rsqrtf(warp_sum((a0 * a0 + a1 * a1) + (a2 * a2 + a3 * a3)) * (1.f / DX_D) + 1e-5f);
        float4 d = reinterpret_cast<const float4*>(dy1 + r * DX_D)[lane];
        if (dy2 != nullptr) {
            const float4 e = reinterpret_cast<const float4*>(dy2 + r * DX_D)[lane];
            d.x += e.x; d.y += e.y; d.z += e.z; d.w += e.w;
        }
        ag.x += d.x * (a0 * rstd * gg.x + bb.x); ag.y += d.y * (a1 * rstd * gg.y + bb.y);
        ag.z += d.z * (a2 * rstd * gg.z + bb.z); ag.w += d.w * (a3 * rstd * gg.w + bb.w);
        ab.x += d.x; ab.y += d.y; ab.z += d.z; ab.w += d.w;
    }
    for (int w = 0; w < 8; ++w) {
        if (warp == w) {
            float4* sg = reinterpret_cast<float4*>(s_acc) + lane;
            float4* sb = reinterpret_cast<float4*>(s_acc + 128) + lane;
            if (w == 0) { *sg = ag; *sb = ab; }
            else {
                float4 a = *sg, c = *sb;
                a.x += ag.x; a.y += ag.y; a.z += ag.z; a.w += ag.w;
                c.x += ab.x; c.y += ab.y; c.z += ab.z; c.w += ab.w;
                *sg = a; *sb = c;
            }
        }
        __syncthreads();
    }
    part[((int64_t)blockIdx.y * gridDim.x + b) * 256 + threadIdx.x] = s_acc[threadIdx.x];
}

__global__ void __launch_bounds__(256)
film_bwd_finish_kernel(const float* __restrict__ part, int B, float* __restrict__ dfilm, int film_stride) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const int b = blockIdx.x;
    float acc = 0.f;
#pragma unroll
    for (int y = 0; y < FB_SLICES; ++y) acc += part[((int64_t)y * B + b) * 256 + threadIdx.x];
    dfilm[(int64_t)b * film_stride + threadIdx.x] = acc;
}

int64_t film_bwd_scratch_floats(int B) { return (int64_t)FB_SLICES * B * 256; }

int film_bwd(const float* in, const float* res, const float* g, const float* bta, const float* dy1, const float* dy2,
             const Rows& R, float* dfilm, int film_stride, float* scratch, cudaStream_t s) {
    DX_LAUNCH_PDL((film_bwd_kernel), dim3((unsigned)R.B, FB_SLICES), 256, 0, s, in, res, g, bta, dy1, dy2, R, scratch);
    DX_CHECK_LAUNCH();
    DX_LAUNCH_PDL((film_bwd_finish_kernel), R.B, 256, 0, s, scratch, R.B, dfilm, film_stride);
    DX_CHECK_LAUNCH();
    return 0;
}

// =========================================================================================================
// colsum_rows: part[chunk][n] = sum over the chunk's rows (< total) of in[r][n]; then reduce_partials.
// =========================================================================================================
constexpr int CS_ROWS = 128;

// 256 threads = 32 column quads (float4, 128 columns) x 8 row lanes; a row lane adds rows r0 + lane, r0 + lane + 8, ...
// (16 independent 16-byte loads in flight per thread), the 8 lanes are combined in lane order: a fixed summation order.
__global__ void __launch_bounds__(256)
colsum_rows_kernel(const float* __restrict__ in, int N, Rows R, float* __restrict__ part) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    __shared__ float4 s_acc[8][32];
    const int64_t total = R.row_start[R.B];
    const int q = threadIdx.x & 31, rl = threadIdx.x >> 5;
    const int c = blockIdx.y * 128 + q * 4;
    const int64_t r0 = (int64_t)blockIdx.x * CS_ROWS;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < N) {
#pragma unroll
        for (int k = 0; k < CS_ROWS / 8; ++k) {
            const int64_t r = r0 + rl + 8 * k;
            if (r < total) {
                const float4 v = *reinterpret_cast<const float4*>(in + r * N + c);
                acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
            }
        }
    }
    s_acc[rl][q] = acc;
    __syncthreads();
    if (rl == 0 && c < N) {
        float4 t = s_acc[0][q];
#pragma unroll
        for (int k = 1; k < 8; ++k) {
            const float4 v = s_acc[k][q];
            t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
        }
        *reinterpret_cast<float4*>(part + (int64_t)blockIdx.x * N + c) = t;
    }
}

int colsum_rows(const float* in, int N, const Rows& R, float* out, float* scratch, cudaStream_t s) {
    DX_REQUIRE(N % 4 == 0, "colsum_rows: N=%d must be a multiple of 4", N);
    const int S = (int)((R.cap + CS_ROWS - 1) / CS_ROWS);
    dim3 grid(S, (N + 127) / 128);
    DX_LAUNCH_PDL((colsum_rows_kernel), grid, 256, 0, s, in, N, R, scratch);
    DX_CHECK_LAUNCH();
    return reduce_partials(scratch, S, N, out, R.row_start + R.B, CS_ROWS, 1.f, s);
}
int64_t colsum_scratch_floats(int64_t cap, int N) { return ((cap + CS_ROWS - 1) / CS_ROWS) * (int64_t)N; }

// =========================================================================================================
// wgrad_f32: dW[j][k][n] = sum_{r < total} A[r + j - half][k] * G[r][n]   (rows outside [0, total) of A read as 0).
// 128 x 128 tile of (k, n) per CTA for one tap j and one chunk of WG_ROWS rows; 8 x 8 outputs per thread;
// part[chunk][taps][K][N], then reduce_partials.
// =========================================================================================================
constexpr int WG_ROWS = 1024, WG_BR = 16;

__global__ void __launch_bounds__(256)
wgrad_f32_kernel(const float* __restrict__ A, int K, const float* __restrict__ G, int N, int taps, Rows R,
                 float* __restrict__ part) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    __shared__ __align__(16) float As[WG_BR][128];
    __shared__ __align__(16) float Gs[WG_BR][128];
    const int64_t total = R.row_start[R.B];
    const int n0 = blockIdx.x * 128, k0 = blockIdx.y * 128;
    const int chunk = blockIdx.z / taps, j = blockIdx.z % taps;
    const int half = taps / 2;
    const int64_t r_begin = (int64_t)chunk * WG_ROWS;
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int q = 0; q < 8; ++q) acc[i][q] = 0.f;
    if (r_begin < total) {
        const int64_t r_end = min((int64_t)(r_begin + WG_ROWS), total);
        // loader: 16 rows x 128 columns = 512 float4; thread loads 2 float4 of A and 2 of G
        for (int64_t rb = r_begin; rb < r_end; rb += WG_BR) {
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int idx = tid + 256 * u;            // 0..511
                const int rr = idx >> 5, c4 = idx & 31;
                const int64_t rg = rb + rr;
                float4 a = make_float4(0.f, 0.f, 0.f, 0.f), gq = a;
                if (rg < r_end) {
                    const int64_t ra = rg + j - half;
                    const int k = k0 + c4 * 4, n = n0 + c4 * 4;
                    if (ra >= 0 && ra < total) {
                        if (k + 3 < K) a = *reinterpret_cast<const float4*>(A + ra * K + k);
                        else { a.x = k < K ? A[ra * K + k] : 0.f; a.y = k + 1 < K ? A[ra * K + k + 1] : 0.f; a.z = k + 2 < K ? A[ra * K + k + 2] : 0.f; }
                    }
                    if (n + 3 < N) gq = *reinterpret_cast<const float4*>(G + rg * N + n);
                    else { gq.x = n < N ? G[rg * N + n] : 0.f; gq.y = n + 1 < N ? G[rg * N + n + 1] : 0.f; gq.z = n + 2 < N ? G[rg * N + n + 2] : 0.f; }
                }
                *reinterpret_cast<float4*>(&As[rr][c4 * 4]) = a;
                *reinterpret_cast<float4*>(&Gs[rr][c4 * 4]) = gq;
            }
            __syncthreads();
#pragma unroll
            for (int rr = 0; rr < WG_BR; ++rr) {
                const float4 a0 = *reinterpret_cast<const float4*>(&As[rr][ty * 4]);
                const float4 a1 = *reinterpret_cast<const float4*>(&As[rr][64 + ty * 4]);
                const float4 g0 = *reinterpret_cast<const float4*>(&Gs[rr][tx * 4]);
                const float4 g1 = *reinterpret_cast<const float4*>(&Gs[rr][64 + tx * 4]);
                const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
                const float gv[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
#pragma unroll
                for (int i = 0; i < 8; ++i)
#pragma unroll
                    for (int q = 0; q < 8; ++q) acc[i][q] = fmaf(av[i], gv[q], acc[i][q]);
            }
            __syncthreads();
        }
    }
    float* out = part + ((int64_t)chunk * taps + j) * K * N;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int k = k0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (k >= K) continue;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int n = n0 + (q < 4 ? tx * 4 + q : 64 + tx * 4 + (q - 4));
            if (n < N) out[(int64_t)k * N + n] = acc[i][q];
        }
    }
}

int wgrad_f32(const float* A, int K, const float* G, int N, int taps, const Rows& R, float* dW, float* scratch, cudaStream_t s) {
    const int S = (int)((R.cap + WG_ROWS - 1) / WG_ROWS);
    dim3 grid((N + 127) / 128, (K + 127) / 128, S * taps);
    ProfScope prof("wgrad_f32", s);
    DX_LAUNCH_PDL((wgrad_f32_kernel), grid, 256, 0, s, A, K, G, N, taps, R, scratch);
    DX_CHECK_LAUNCH();
    return reduce_partials(scratch, S, (int64_t)taps * K * N, dW, R.row_start + R.B, WG_ROWS, 1.f, s);
}
int64_t wgrad_scratch_floats(int64_t cap, int K, int N, int taps) { return ((cap + WG_ROWS - 1) / WG_ROWS) * (int64_t)taps * K * N; }

// dgrad weight: Wt[taps - 1 - j][n][k] = W[j][k][n]
__global__ void transpose_weight_kernel(const float* __restrict__ W, int taps, int K, int N, float* __restrict__ Wt) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    __shared__ float tile[32][33];
    const int j = blockIdx.z;
    const int k0 = blockIdx.y * 32, n0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;      // 256 threads: ty 0..7
    for (int i = ty; i < 32; i += 8) {
        const int k = k0 + i, n = n0 + tx;
        tile[i][tx] = (k < K && n < N) ? W[((int64_t)j * K + k) * N + n] : 0.f;
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        const int n = n0 + i, k = k0 + tx;
        if (n < N && k < K) Wt[((int64_t)(taps - 1 - j) * N + n) * K + k] = tile[tx][i];
    }
}

int transpose_weight(const float* W, int taps, int K, int N, float* Wt, cudaStream_t s) {
    dim3 grid((N + 31) / 32, (K + 31) / 32, taps);
    DX_LAUNCH_PDL((transpose_weight_kernel), grid, 256, 0, s, W, taps, K, N, Wt);
    DX_CHECK_LAUNCH();
    return 0;
}

__global__ void relu_bwd_kernel(float* __restrict__ d, const float* __restrict__ act, int64_t n4) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    float4 g = reinterpret_cast<float4*>(d)[i];
    const float4 a = reinterpret_cast<const float4*>(act)[i];
    g.x = a.x > 0.f ? g.x : 0.f; g.y = a.y > 0.f ? g.y : 0.f; g.z = a.z > 0.f ? g.z : 0.f; g.w = a.w > 0.f ? g.w : 0.f;
    reinterpret_cast<float4*>(d)[i] = g;
}

int relu_bwd(float* d, const float* act, int64_t n, cudaStream_t s) {
    DX_LAUNCH_PDL((relu_bwd_kernel), (unsigned)((n / 4 + 255) / 256), 256, 0, s, d, act, n / 4);
    DX_CHECK_LAUNCH();
    return 0;
}

// =========================================================================================================
// Attention backward.  Forward (attention_f32): S = (0.125 Q) K^T, P = softmax over valid keys, O = P V.
//   Dr[i] = dO[i] . O[i];  P[i][j] = exp(S[i][j] - lse[i]);  dP = dO V^T;  dS = P (dP - Dr)
//   dQ = 0.125 dS K;  dK = 0.125 dS^T Q;  dV = P^T dO
// attn_bwd_dq: one CTA per (utterance, head, 64-query tile), loops over key tiles.
// attn_bwd_dkv: one CTA per (utterance, head, 64-key tile), loops over query tiles.  Both recompute S and P.
// =========================================================================================================
__global__ void attn_rowdot_kernel(const float* __restrict__ dO, const float* __restrict__ O, Rows R, float* __restrict__ Dr) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= R.cap) return;
    const float4 a = reinterpret_cast<const float4*>(dO + r * DX_D)[lane], b = reinterpret_cast<const float4*>(O + r * DX_D)[lane];
    float v = (a.x * b.x + a.y * b.y) + (a.z * b.z + a.w * b.w);
    // lanes 0-15 hold head 0 (channels 0-63), lanes 16-31 head 1
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((lane & 15) == 0) Dr[r * DX_HEADS + (lane >> 4)] = v;
}

constexpr int AB_T = 64;
constexpr int AB_DQ_SMEM = (4 * 64 * 64 + 2 * 64 * 65 + 64 * 68) * (int)sizeof(float);

__global__ void __launch_bounds__(256)
attn_bwd_dq_kernel(const float* __restrict__ qkv, const float* __restrict__ dO, const float* __restrict__ lse,
                   const float* __restrict__ Dr, Rows R, float* __restrict__ dqkv) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    extern __shared__ __align__(16) float ab_smem[];
    float(*Qt)[64] = reinterpret_cast<float(*)[64]>(ab_smem);                 // [d][q], pre-scaled
    float(*dOt)[64] = reinterpret_cast<float(*)[64]>(ab_smem + 64 * 64);      // [d][q]
    float(*Ks)[64] = reinterpret_cast<float(*)[64]>(ab_smem + 2 * 64 * 64);   // [k][d]
    float(*Kt)[65] = reinterpret_cast<float(*)[65]>(ab_smem + 3 * 64 * 64);   // [d][k]
    float(*Vt)[65] = reinterpret_cast<float(*)[65]>(ab_smem + 3 * 64 * 64 + 64 * 65);
    float(*dSt)[68] = reinterpret_cast<float(*)[68]>(ab_smem + 3 * 64 * 64 + 2 * 64 * 65);   // [k][q]
    const int b = blockIdx.z, h = blockIdx.y;
    const int len = (int)min((long long)R.len[b], (long long)R.n_max);
    const int q0 = blockIdx.x * AB_T;
    if (q0 >= len) return;
    const int64_t rs = R.row_start[b];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    for (int i = tid; i < AB_T * 16; i += 256) {
        const int q = i >> 4, d4 = i & 15;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f), g = v;
        if (q0 + q < len) {
            v = *reinterpret_cast<const float4*>(qkv + (rs + q0 + q) * 384 + h * DX_DH + d4 * 4);
            g = *reinterpret_cast<const float4*>(dO + (rs + q0 + q) * DX_D + h * DX_DH + d4 * 4);
        }
        Qt[d4 * 4 + 0][q] = v.x * 0.125f; Qt[d4 * 4 + 1][q] = v.y * 0.125f; Qt[d4 * 4 + 2][q] = v.z * 0.125f; Qt[d4 * 4 + 3][q] = v.w * 0.125f;
        dOt[d4 * 4 + 0][q] = g.x; dOt[d4 * 4 + 1][q] = g.y; dOt[d4 * 4 + 2][q] = g.z; dOt[d4 * 4 + 3][q] = g.w;
    }
    float lse_i[4], dr_i[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int q = q0 + ty * 4 + i;
        lse_i[i] = q < len ? lse[(rs + q) * DX_HEADS + h] : 0.f;
        dr_i[i] = q < len ? Dr[(rs + q) * DX_HEADS + h] : 0.f;
    }
    float dq[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) dq[i][j] = 0.f;
    for (int k0 = 0; k0 < len; k0 += AB_T) {
        __syncthreads();
        for (int i = tid; i < AB_T * 16; i += 256) {
            const int kk = i >> 4, d4 = i & 15;
            float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
            if (k0 + kk < len) {
                const float* base = qkv + (rs + k0 + kk) * 384 + h * DX_DH + d4 * 4;
                kv = *reinterpret_cast<const float4*>(base + 128);
                vv = *reinterpret_cast<const float4*>(base + 256);
            }
            *reinterpret_cast<float4*>(&Ks[kk][d4 * 4]) = kv;
            Kt[d4 * 4 + 0][kk] = kv.x; Kt[d4 * 4 + 1][kk] = kv.y; Kt[d4 * 4 + 2][kk] = kv.z; Kt[d4 * 4 + 3][kk] = kv.w;
            Vt[d4 * 4 + 0][kk] = vv.x; Vt[d4 * 4 + 1][kk] = vv.y; Vt[d4 * 4 + 2][kk] = vv.z; Vt[d4 * 4 + 3][kk] = vv.w;
        }
        __syncthreads();
        float sc[4][4], dp[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) { sc[i][j] = 0.f; dp[i][j] = 0.f; }
#pragma unroll 8
        for (int d = 0; d < DX_DH; ++d) {
            const float4 qa = *reinterpret_cast<const float4*>(&Qt[d][ty * 4]);
            const float4 ga = *reinterpret_cast<const float4*>(&dOt[d][ty * 4]);
            const float qv[4] = {qa.x, qa.y, qa.z, qa.w}, gv[4] = {ga.x, ga.y, ga.z, ga.w};
            const float kv[4] = {Kt[d][tx * 4], Kt[d][tx * 4 + 1], Kt[d][tx * 4 + 2], Kt[d][tx * 4 + 3]};
            const float vv[4] = {Vt[d][tx * 4], Vt[d][tx * 4 + 1], Vt[d][tx * 4 + 2], Vt[d][tx * 4 + 3]};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) { sc[i][j] = fmaf(qv[i], kv[j], sc[i][j]); dp[i][j] = fmaf(gv[i], vv[j], dp[i][j]); }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const bool ok = (k0 + tx * 4 + j < len) && (q0 + ty * 4 + i < len);
                const float p = ok ? expf(sc[i][j] - lse_i[i]) : 0.f;
                dSt[tx * 4 + j][ty * 4 + i] = p * (dp[i][j] - dr_i[i]);
            }
        __syncthreads();
#pragma unroll 8
        for (int kk = 0; kk < AB_T; ++kk) {
            const float4 sa = *reinterpret_cast<const float4*>(&dSt[kk][ty * 4]);
            const float4 ka = *reinterpret_cast<const float4*>(&Ks[kk][tx * 4]);
            const float sv[4] = {sa.x, sa.y, sa.z, sa.w}, kv[4] = {ka.x, ka.y, ka.z, ka.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dq[i][j] = fmaf(sv[i], kv[j], dq[i][j]);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int q = q0 + ty * 4 + i;
        if (q < len)
            *reinterpret_cast<float4*>(dqkv + (rs + q) * 384 + h * DX_DH + tx * 4) =
                make_float4(dq[i][0] * 0.125f, dq[i][1] * 0.125f, dq[i][2] * 0.125f, dq[i][3] * 0.125f);
    }
}

constexpr int AB_DKV_SMEM = (2 * 64 * 64 + 2 * 64 * 65 + 2 * 64 * 64 + 2 * 64 * 68) * (int)sizeof(float);

__global__ void __launch_bounds__(256)
attn_bwd_dkv_kernel(const float* __restrict__ qkv, const float* __restrict__ dO, const float* __restrict__ lse,
                    const float* __restrict__ Dr, Rows R, float* __restrict__ dqkv) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    extern __shared__ __align__(16) float ab_smem[];
    float(*Qt)[64] = reinterpret_cast<float(*)[64]>(ab_smem);                       // [d][q], pre-scaled
    float(*dOt)[64] = reinterpret_cast<float(*)[64]>(ab_smem + 64 * 64);            // [d][q]
    float(*Kt)[65] = reinterpret_cast<float(*)[65]>(ab_smem + 2 * 64 * 64);         // [d][k]
    float(*Vt)[65] = reinterpret_cast<float(*)[65]>(ab_smem + 2 * 64 * 64 + 64 * 65);
    float(*Qs)[64] = reinterpret_cast<float(*)[64]>(ab_smem + 2 * 64 * 64 + 2 * 64 * 65);            // [q][d], pre-scaled
    float(*dOs)[64] = reinterpret_cast<float(*)[64]>(ab_smem + 3 * 64 * 64 + 2 * 64 * 65);           // [q][d]
    float(*Ps)[68] = reinterpret_cast<float(*)[68]>(ab_smem + 4 * 64 * 64 + 2 * 64 * 65);            // [q][k]
    float(*dSs)[68] = reinterpret_cast<float(*)[68]>(ab_smem + 4 * 64 * 64 + 2 * 64 * 65 + 64 * 68); // [q][k]
    const int b = blockIdx.z, h = blockIdx.y;
    const int len = (int)min((long long)R.len[b], (long long)R.n_max);
    const int k0 = blockIdx.x * AB_T;
    if (k0 >= len) return;
    const int64_t rs = R.row_start[b];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    for (int i = tid; i < AB_T * 16; i += 256) {
        const int kk = i >> 4, d4 = i & 15;
        float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
        if (k0 + kk < len) {
            const float* base = qkv + (rs + k0 + kk) * 384 + h * DX_DH + d4 * 4;
            kv = *reinterpret_cast<const float4*>(base + 128);
            vv = *reinterpret_cast<const float4*>(base + 256);
        }
        Kt[d4 * 4 + 0][kk] = kv.x; Kt[d4 * 4 + 1][kk] = kv.y; Kt[d4 * 4 + 2][kk] = kv.z; Kt[d4 * 4 + 3][kk] = kv.w;
        Vt[d4 * 4 + 0][kk] = vv.x; Vt[d4 * 4 + 1][kk] = vv.y; Vt[d4 * 4 + 2][kk] = vv.z; Vt[d4 * 4 + 3][kk] = vv.w;
    }
    float dk[4][4], dvv[4][4];          // rows: key ty*4+i, columns: dim tx*4+j
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) { dk[i][j] = 0.f; dvv[i][j] = 0.f; }
    for (int q0 = 0; q0 < len; q0 += AB_T) {
        __syncthreads();
        for (int i = tid; i < AB_T * 16; i += 256) {
            const int q = i >> 4, d4 = i & 15;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f), g = v;
            if (q0 + q < len) {
                v = *reinterpret_cast<const float4*>(qkv + (rs + q0 + q) * 384 + h * DX_DH + d4 * 4);
                g = *reinterpret_cast<const float4*>(dO + (rs + q0 + q) * DX_D + h * DX_DH + d4 * 4);
            }
            v.x *= 0.125f; v.y *= 0.125f; v.z *= 0.125f; v.w *= 0.125f;
            *reinterpret_cast<float4*>(&Qs[q][d4 * 4]) = v;
            *reinterpret_cast<float4*>(&dOs[q][d4 * 4]) = g;
            Qt[d4 * 4 + 0][q] = v.x; Qt[d4 * 4 + 1][q] = v.y; Qt[d4 * 4 + 2][q] = v.z; Qt[d4 * 4 + 3][q] = v.w;
            dOt[d4 * 4 + 0][q] = g.x; dOt[d4 * 4 + 1][q] = g.y; dOt[d4 * 4 + 2][q] = g.z; dOt[d4 * 4 + 3][q] = g.w;
        }
        __syncthreads();
        // S / dP block: queries ty*4+i, keys tx*4+j
        float sc[4][4], dp[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) { sc[i][j] = 0.f; dp[i][j] = 0.f; }
#pragma unroll 8
        for (int d = 0; d < DX_DH; ++d) {
            const float4 qa = *reinterpret_cast<const float4*>(&Qt[d][ty * 4]);
            const float4 ga = *reinterpret_cast<const float4*>(&dOt[d][ty * 4]);
            const float qv[4] = {qa.x, qa.y, qa.z, qa.w}, gv[4] = {ga.x, ga.y, ga.z, ga.w};
            const float kv[4] = {Kt[d][tx * 4], Kt[d][tx * 4 + 1], Kt[d][tx * 4 + 2], Kt[d][tx * 4 + 3]};
            const float vv[4] = {Vt[d][tx * 4], Vt[d][tx * 4 + 1], Vt[d][tx * 4 + 2], Vt[d][tx * 4 + 3]};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) { sc[i][j] = fmaf(qv[i], kv[j], sc[i][j]); dp[i][j] = fmaf(gv[i], vv[j], dp[i][j]); }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int q = q0 + ty * 4 + i;
            const float l_i = q < len ? lse[(rs + q) * DX_HEADS + h] : 0.f;
            const float d_i = q < len ? Dr[(rs + q) * DX_HEADS + h] : 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const bool ok = (k0 + tx * 4 + j < len) && (q < len);
                const float p = ok ? expf(sc[i][j] - l_i) : 0.f;
                Ps[ty * 4 + i][tx * 4 + j] = p;
                dSs[ty * 4 + i][tx * 4 + j] = p * (dp[i][j] - d_i);
            }
        }
        __syncthreads();
        // dV[k][d] += sum_q P[q][k] dO[q][d];  dK[k][d] += sum_q dS[q][k] Q[q][d]   (Q pre-scaled by 0.125)
#pragma unroll 8
        for (int q = 0; q < AB_T; ++q) {
            const float4 pa = *reinterpret_cast<const float4*>(&Ps[q][ty * 4]);
            const float4 sa = *reinterpret_cast<const float4*>(&dSs[q][ty * 4]);
            const float4 ga = *reinterpret_cast<const float4*>(&dOs[q][tx * 4]);
            const float4 qa = *reinterpret_cast<const float4*>(&Qs[q][tx * 4]);
            const float pv[4] = {pa.x, pa.y, pa.z, pa.w}, sv[4] = {sa.x, sa.y, sa.z, sa.w};
            const float gv[4] = {ga.x, ga.y, ga.z, ga.w}, qv[4] = {qa.x, qa.y, qa.z, qa.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) { dvv[i][j] = fmaf(pv[i], gv[j], dvv[i][j]); dk[i][j] = fmaf(sv[i], qv[j], dk[i][j]); }
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int kk = k0 + ty * 4 + i;
        if (kk < len) {
            float* base = dqkv + (rs + kk) * 384 + h * DX_DH + tx * 4;
            *reinterpret_cast<float4*>(base + 128) = make_float4(dk[i][0], dk[i][1], dk[i][2], dk[i][3]);
            *reinterpret_cast<float4*>(base + 256) = make_float4(dvv[i][0], dvv[i][1], dvv[i][2], dvv[i][3]);
        }
    }
}

int attn_rowdot(const float* dO, const float* O, const Rows& R, float* Dr, cudaStream_t s) {
    DX_LAUNCH_PDL((attn_rowdot_kernel), (unsigned)((R.cap + 7) / 8), 256, 0, s, dO, O, R, Dr);
    DX_CHECK_LAUNCH();
    return 0;
}

// dqkv must be zero at rows that are not valid positions (gap rows); the caller memsets it once per call.
int attention_bwd_f32(const float* qkv, const float* O, const float* lse, const float* dO, const Rows& R, float* Dr, float* dqkv,
                      cudaStream_t s) {
    static SmemConfigured c1, c2;
    DX_TRY(ensure_dyn_smem(attn_bwd_dq_kernel, (size_t)AB_DQ_SMEM, c1));
    DX_TRY(ensure_dyn_smem(attn_bwd_dkv_kernel, (size_t)AB_DKV_SMEM, c2));
    ProfScope prof("attn_bwd", s);
    DX_CHECK_CUDA(cudaMemsetAsync(dqkv, 0, (size_t)R.cap * 384 * sizeof(float), s));
    DX_LAUNCH_PDL((attn_rowdot_kernel), (unsigned)((R.cap + 7) / 8), 256, 0, s, dO, O, R, Dr);
    DX_CHECK_LAUNCH();
    dim3 grid((unsigned)((R.n_max + AB_T - 1) / AB_T), DX_HEADS, (unsigned)R.B);
    DX_LAUNCH_PDL((attn_bwd_dq_kernel), grid, 256, AB_DQ_SMEM, s, qkv, dO, lse, Dr, R, dqkv);
    DX_CHECK_LAUNCH();
    DX_LAUNCH_PDL((attn_bwd_dkv_kernel), grid, 256, AB_DKV_SMEM, s, qkv, dO, lse, Dr, R, dqkv);
    DX_CHECK_LAUNCH();
    return 0;
}

// =========================================================================================================
// Small pieces around the stacks
// =========================================================================================================
// masked mean pool backward (model.py:714): dx[r] = daccent[b] / len[b] at valid rows, 0 elsewhere
__global__ void masked_mean_bwd_kernel(const float* __restrict__ dacc, Rows R, float* __restrict__ dx) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= R.cap) return;
    const int b = R.row_b[r];
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (b >= 0 && R.row_p[r] < R.len[b]) {
        const float inv = 1.f / (float)R.len[b];
        const float4 a = reinterpret_cast<const float4*>(dacc + (int64_t)b * DX_D)[lane];
        v = make_float4(a.x * inv, a.y * inv, a.z * inv, a.w * inv);
    }
    reinterpret_cast<float4*>(dx + r * DX_D)[lane] = v;
}
int masked_mean_bwd(const float* dacc, const Rows& R, float* dx, cudaStream_t s) {
    DX_LAUNCH_PDL((masked_mean_bwd_kernel), (unsigned)((R.cap + 7) / 8), 256, 0, s, dacc, R, dx);
    DX_CHECK_LAUNCH();
    return 0;
}

// out[r] = (a[r] + b[r]) at valid rows (p < len), 0 elsewhere: the pad mask of a stack input (model.py:604,706,557)
__global__ void add_mask_rows_kernel(const float* __restrict__ a, const float* __restrict__ b2, Rows R, float* __restrict__ out) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= R.cap) return;
    const int b = R.row_b[r];
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (b >= 0 && R.row_p[r] < R.len[b]) {
        v = reinterpret_cast<const float4*>(a + r * DX_D)[lane];
        if (b2 != nullptr) {
            const float4 q = reinterpret_cast<const float4*>(b2 + r * DX_D)[lane];
            v.x += q.x; v.y += q.y; v.z += q.z; v.w += q.w;
        }
    }
    reinterpret_cast<float4*>(out + r * DX_D)[lane] = v;
}
int add_mask_rows(const float* a, const float* b, const Rows& R, float* out, cudaStream_t s) {
    DX_LAUNCH_PDL((add_mask_rows_kernel), (unsigned)((R.cap + 7) / 8), 256, 0, s, a, b, R, out);
    DX_CHECK_LAUNCH();
    return 0;
}

// Scalar-input conv k=3 (ConvNorm1D(1 -> 128), model.py:402-410,640-646): y[r][c] = b[c] + sum_j w[c][j] x[b, p + j - 1].
// dW[c][j] = sum_{valid rows} g[r][c] x[b, p + j - 1];  db[c] = sum g[r][c].  part[chunk][128][4] (3 taps + bias).
constexpr int SC_ROWS = 32;
__global__ void __launch_bounds__(128)
scalar_conv_wgrad_kernel(const float* __restrict__ g, const float* __restrict__ x, Rows R, float* __restrict__ part) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const int c = threadIdx.x;
    const int64_t r0 = (int64_t)blockIdx.x * SC_ROWS;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, ab = 0.f;
    for (int k = 0; k < SC_ROWS; ++k) {
        const int64_t r = r0 + k;
        if (r >= R.cap) break;
        const int b = R.row_b[r];
        if (b < 0) continue;
        const int p = R.row_p[r];
        if (p >= R.len[b] || p >= R.n_max) continue;
        const float* xr = x + (int64_t)b * R.n_max;
        const float gv = g[r * DX_D + c];
        a0 += gv * (p - 1 >= 0 ? xr[p - 1] : 0.f);
        a1 += gv * xr[p];
        a2 += gv * (p + 1 < R.n_max ? xr[p + 1] : 0.f);
        ab += gv;
    }
    float* o = part + ((int64_t)blockIdx.x * DX_D + c) * 4;
    o[0] = a0; o[1] = a1; o[2] = a2; o[3] = ab;
}
__global__ void scalar_conv_split_kernel(const float* __restrict__ red, float* __restrict__ dw, float* __restrict__ db) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const int c = threadIdx.x;
    dw[c * 3 + 0] = red[c * 4 + 0]; dw[c * 3 + 1] = red[c * 4 + 1]; dw[c * 3 + 2] = red[c * 4 + 2];
    db[c] = red[c * 4 + 3];
}
int scalar_conv_wgrad(const float* g, const float* x, const Rows& R, float* dw, float* db, float* scratch, cudaStream_t s) {
    const int S = (int)((R.cap + SC_ROWS - 1) / SC_ROWS);
    DX_LAUNCH_PDL((scalar_conv_wgrad_kernel), S, 128, 0, s, g, x, R, scratch);
    DX_CHECK_LAUNCH();
    float* red = scratch + (int64_t)S * DX_D * 4;
    DX_TRY(reduce_partials(scratch, S, DX_D * 4, red, nullptr, 0, 1.f, s));
    DX_LAUNCH_PDL((scalar_conv_split_kernel), 1, 128, 0, s, red, dw, db);
    DX_CHECK_LAUNCH();
    return 0;
}
int64_t scalar_conv_scratch_floats(int64_t cap) { return ((cap + SC_ROWS - 1) / SC_ROWS + 1) * DX_D * 4; }

// Embedding backward (nn.Embedding, model.py:597): dEmb[s] = sum over valid rows whose symbol is s of g[r].  One CTA
// per symbol scans the row stream in order (deterministic); symbols are clamped like embed_symbols clamps them.
__global__ void __launch_bounds__(128)
embed_bwd_kernel(const float* __restrict__ g, const int64_t* __restrict__ symbols, int n_symbols, Rows R, float* __restrict__ demb) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    __shared__ int s_match[128];
    const int sym = blockIdx.x, c = threadIdx.x;
    const int64_t total = R.row_start[R.B];
    float acc = 0.f;
    // chunks of 128 rows: every thread classifies one row (parallel loads), then all threads add the matching rows of
    // the chunk in row order (the summation order of the serial scan)
    for (int64_t r0 = 0; r0 < total; r0 += 128) {
        const int64_t r = r0 + c;
        int m = 0;
        if (r < total) {
            const int b = R.row_b[r];
            const int p = R.row_p[r];
            if (b >= 0 && p < R.len[b] && p < R.n_max) {
                long long sv = symbols[(int64_t)b * R.n_max + p];
                sv = sv < 0 ? 0 : (sv >= n_symbols ? n_symbols - 1 : sv);
                m = (int)sv == sym;
            }
        }
        s_match[c] = m;
        __syncthreads();
        const int lim = (int)min((int64_t)128, total - r0);
        for (int k = 0; k < lim; ++k)
            if (s_match[k]) acc += g[(r0 + k) * DX_D + c];
        __syncthreads();
    }
    demb[(int64_t)sym * DX_D + c] = acc;
}
int embed_bwd(const float* g, const int64_t* symbols, int n_symbols, const Rows& R, float* demb, cudaStream_t s) {
    DX_LAUNCH_PDL((embed_bwd_kernel), n_symbols, 128, 0, s, g, symbols, n_symbols, R, demb);
    DX_CHECK_LAUNCH();
    return 0;
}

// dW[n][k] = sum_b dOut[b][n] In[b][k],  db[n] = sum_b dOut[b][n]   (nn.Linear in torch layout [out][in]; B utterances)
__global__ void outer_wgrad_kernel(const float* __restrict__ dOut, int N, const float* __restrict__ In, int K, int B,
                                   float* __restrict__ dW, float* __restrict__ db) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (int64_t)N * K) {
        const int n = (int)(i / K), k = (int)(i % K);
        float acc = 0.f;
        for (int b = 0; b < B; ++b) acc += dOut[(int64_t)b * N + n] * In[(int64_t)b * K + k];
        dW[i] = acc;
    }
    if (db != nullptr && i < N) {
        float acc = 0.f;
        for (int b = 0; b < B; ++b) acc += dOut[(int64_t)b * N + i];
        db[i] = acc;
    }
}
int outer_wgrad(const float* dOut, int N, const float* In, int K, int B, float* dW, float* db, cudaStream_t s) {
    DX_LAUNCH_PDL((outer_wgrad_kernel), (unsigned)(((int64_t)N * K + 255) / 256), 256, 0, s, dOut, N, In, K, B, dW, db);
    DX_CHECK_LAUNCH();
    return 0;
}

// =========================================================================================================
// Length regulator backward (GaussianUpsamplingModule.forward, model.py:417-510; forward kernels in dx_lenreg.cu).
// With w[l,t] the saved alignment, g[t] the gradient of the decoder input at valid frames and x_up[t] = sum_l w x'[l]:
//   dw[l,t] = g[t] . x'[l];  c[t] = sum_l w dw = g[t] . x_up[t];  dlogp[l,t] = w (dw - c[t])      (softmax-like normalisation;
//   the 1e-20 in the denominator only matters where every p underflows, where w = 0 as well)
//   dx'[l] = sum_t w g[t];  dsigma[l] = sum_t dlogp ((t + .5 - mu)^2 / sigma^3 - 1 / sigma)
//   sigma = max(softplus(pre), 1e-3), pre = w_r . (x' + dproj) + b_r
// =========================================================================================================
__global__ void lenreg_bwd_ct_kernel(const float* __restrict__ g, const float* __restrict__ xdec, const float* __restrict__ pos,
                                     Rows RT, float* __restrict__ ct) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= RT.cap) return;
    const int b = RT.row_b[r];
    float v = 0.f;
    if (b >= 0 && RT.row_p[r] < RT.len[b]) {
        const int t = RT.row_p[r];
        const float4 a = reinterpret_cast<const float4*>(g + r * DX_D)[lane];
        const float4 x = reinterpret_cast<const float4*>(xdec + r * DX_D)[lane];
        const float4 q = reinterpret_cast<const float4*>(pos + (int64_t)t * DX_D)[lane];
        v = warp_sum((a.x * (x.x - q.x) + a.y * (x.y - q.y)) + (a.z * (x.z - q.z) + a.w * (x.w - q.w)));
    }
    if (lane == 0) ct[r] = v;
}

// one warp per symbol row; lane <-> 4 channels.  Writes gx = dx' (total, incl. the range-predictor path), gd = d(dproj),
// gr = dpre * (x' + dproj) (for d w_r), dpre[row].
__global__ void __launch_bounds__(256)
lenreg_bwd_sym_kernel(const float* __restrict__ g, const float* __restrict__ ct, const float* __restrict__ weights,
                      const float* __restrict__ xs, const float* __restrict__ mu, const float* __restrict__ sigma,
                      const float* __restrict__ dur, DxModelWeights w, Rows RL, Rows RT, float* __restrict__ gx,
                      float* __restrict__ gd, float* __restrict__ gr, float* __restrict__ dpre_out) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= RL.cap) return;
    const int b = RL.row_b[r];
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    const bool valid = b >= 0 && RL.row_p[r] < RL.len[b] && RL.row_p[r] < RL.n_max;
    if (!valid) {
        reinterpret_cast<float4*>(gx + r * DX_D)[lane] = z4;
        reinterpret_cast<float4*>(gd + r * DX_D)[lane] = z4;
        reinterpret_cast<float4*>(gr + r * DX_D)[lane] = z4;
        if (lane == 0) dpre_out[r] = 0.f;
        return;
    }
    const int l = RL.row_p[r], L = RL.n_max, T = RT.n_max;
    const int len_out = (int)min((long long)RT.len[b], (long long)T);
    const int64_t rsT = RT.row_start[b];
    const float4 x = reinterpret_cast<const float4*>(xs + r * DX_D)[lane];
    const float m = mu[(int64_t)b * L + l], sg = sigma[(int64_t)b * L + l];
    const float inv_sg = 1.f / sg, inv_sg3 = inv_sg * inv_sg * inv_sg;
    const float* wrow = weights + ((int64_t)b * L + l) * T;
    float4 acc = z4;
    float dsig = 0.f;
    for (int t0 = 0; t0 < len_out; t0 += 32) {
        const int t = t0 + lane;
        const float wv = (t < len_out) ? wrow[t] : 0.f;
        unsigned mask = __ballot_sync(0xffffffffu, wv != 0.f);
        while (mask) {
            const int j = __ffs(mask) - 1;
            mask &= mask - 1;
            const float wj = __shfl_sync(0xffffffffu, wv, j);
            const int tj = t0 + j;
            const float4 gv = reinterpret_cast<const float4*>(g + (rsT + tj) * DX_D)[lane];
            const float dw = warp_sum((gv.x * x.x + gv.y * x.y) + (gv.z * x.z + gv.w * x.w));
            acc.x = fmaf(wj, gv.x, acc.x); acc.y = fmaf(wj, gv.y, acc.y); acc.z = fmaf(wj, gv.z, acc.z); acc.w = fmaf(wj, gv.w, acc.w);
            const float d = (float)tj + 0.5f - m;
            dsig += wj * (dw - ct[rsT + tj]) * (d * d * inv_sg3 - inv_sg);
        }
    }
    // range predictor: recompute pre = w_r . (x' + dproj) + b_r
    const float* dr = dur + (int64_t)b * L;
    const float dm = (l - 1 >= 0) ? dr[l - 1] : 0.f, d0 = dr[l], dp = (l + 1 < L) ? dr[l + 1] : 0.f;
    float dj[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int c = lane * 4 + i;
        dj[i] = w.gu_dur_b[c] + w.gu_dur_w[c * 3 + 0] * dm + w.gu_dur_w[c * 3 + 1] * d0 + w.gu_dur_w[c * 3 + 2] * dp;
    }
    const float4 wr = reinterpret_cast<const float4*>(w.gu_range_w)[lane];
    const float4 xd = make_float4(x.x + dj[0], x.y + dj[1], x.z + dj[2], x.w + dj[3]);
    const float pre = warp_sum((xd.x * wr.x + xd.y * wr.y) + (xd.z * wr.z + xd.w * wr.w)) + w.gu_range_b[0];
    const float sp = (pre > 20.f) ? pre : log1pf(expf(pre));
    float dpre = 0.f;
    if (sp > 1e-3f) dpre = dsig * ((pre > 20.f) ? 1.f : 1.f / (1.f + expf(-pre)));      // clamp(min) passes the gradient above the floor
    reinterpret_cast<float4*>(gx + r * DX_D)[lane] = make_float4(acc.x + dpre * wr.x, acc.y + dpre * wr.y, acc.z + dpre * wr.z, acc.w + dpre * wr.w);
    reinterpret_cast<float4*>(gd + r * DX_D)[lane] = make_float4(dpre * wr.x, dpre * wr.y, dpre * wr.z, dpre * wr.w);
    reinterpret_cast<float4*>(gr + r * DX_D)[lane] = make_float4(dpre * xd.x, dpre * xd.y, dpre * xd.z, dpre * xd.w);
    if (lane == 0) dpre_out[r] = dpre;
}

int lenreg_bwd(const float* g, const float* xdec, const float* pos, const float* weights, const float* xs, const float* mu,
               const float* sigma, const float* dur, const DxModelWeights* w, const Rows& RL, const Rows& RT, float* ct, float* gx,
               float* gd, float* gr, float* dpre, cudaStream_t s) {
    ProfScope prof("lenreg_bwd", s);
    DX_LAUNCH_PDL((lenreg_bwd_ct_kernel), (unsigned)((RT.cap + 7) / 8), 256, 0, s, g, xdec, pos, RT, ct);
    DX_CHECK_LAUNCH();
    DX_LAUNCH_PDL((lenreg_bwd_sym_kernel), (unsigned)((RL.cap + 7) / 8), 256, 0, s, g, ct, weights, xs, mu, sigma, dur, *w, RL, RT, gx, gd, gr, dpre);
    DX_CHECK_LAUNCH();
    return 0;
}

// rows [cap] -> single sum (d b_r): part by chunks of 1024 rows, ordered
__global__ void sum_rows_scalar_kernel(const float* __restrict__ v, Rows R, float* __restrict__ out) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    __shared__ float s_w[32];
    const int64_t total = R.row_start[R.B];
    float acc = 0.f;
    for (int64_t r = threadIdx.x; r < total; r += blockDim.x) acc += v[r];
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += s_w[i];
        out[0] = t;
    }
}
int sum_rows_scalar(const float* v, const Rows& R, float* out, cudaStream_t s) {
    DX_LAUNCH_PDL((sum_rows_scalar_kernel), 1, 1024, 0, s, v, R, out);
    DX_CHECK_LAUNCH();
    return 0;
}

// =========================================================================================================
// Conditioning heads backward (forward: style_kernel in dx_style.cu; model.py:904-920, 764-806, 829-830 and the
// gradient-reversal layer :27-40).  One CTA per utterance recomputes the small forward and writes the per-utterance
// vectors the weight-gradient reductions (outer_wgrad) need.
//   v layout per utterance (floats): e[192] | c[128] | dg[1024] | dq[1024] | dc[128] | h1[128] | dh1[128] | h2[128] | dh2[128] |
//   dpost[16] | daccent[128]
// =========================================================================================================
__global__ void __launch_bounds__(256)
style_bwd_kernel(const float* __restrict__ spk_embs, const float* __restrict__ accent, DxModelWeights w,
                 const float* __restrict__ dfilm, const float* __restrict__ dlogits, float lambda_rev,
                 float* __restrict__ v) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    __shared__ __align__(16) float s_e[DX_SPK], s_c[DX_D], s_a[DX_D], s_h1[DX_D], s_h2[DX_D], s_t[DX_D], s_dh2[DX_D], s_dh1[DX_D];
    __shared__ float s_g[1024], s_q[1024], s_dg[1024], s_dq[1024];
    __shared__ float s_red[8];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_film = n_blocks_of(w.n_pe_blocks) + n_blocks_of(w.n_fd_blocks), nf = n_film * DX_D;     // FiLM rows (8 by default)
    float* o = v + (int64_t)b * SB_STRIDE;
    if (tid < DX_D) s_a[tid] = accent[b * DX_D + tid];
    {
        const float x = (tid < DX_SPK) ? spk_embs[b * DX_SPK + tid] : 0.f;
        const float ss = warp_sum(x * x);
        if (lane == 0) s_red[warp] = ss;
        __syncthreads();
        float tot = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) tot += s_red[i];
        const float nrm = fmaxf(sqrtf(tot), 1e-12f);
        if (tid < DX_SPK) { s_e[tid] = x / nrm; o[SB_E + tid] = x / nrm; }
    }
    __syncthreads();
    for (int n = warp; n < DX_D; n += 8) {
        const float* wr = w.spk_w + n * DX_SPK;
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < DX_SPK / 32; ++k) acc += wr[lane + 32 * k] * s_e[lane + 32 * k];
        acc = warp_sum(acc);
        if (lane == 0) s_c[n] = s_a[n] + (acc + w.spk_b[n]);
    }
    __syncthreads();
    if (tid < DX_D) o[SB_C + tid] = s_c[tid];
    // raw FiLM predictor outputs and their gradients
    {
        const float4 x = reinterpret_cast<const float4*>(s_c)[lane];
        for (int n = warp; n < nf; n += 8) {
            const float4 wg = reinterpret_cast<const float4*>(w.film_g_w + (int64_t)n * DX_D)[lane];
            const float4 wb = reinterpret_cast<const float4*>(w.film_b_w + (int64_t)n * DX_D)[lane];
            const float gq = warp_sum((wg.x * x.x + wg.y * x.y) + (wg.z * x.z + wg.w * x.w));
            const float qq = warp_sum((wb.x * x.x + wb.y * x.y) + (wb.z * x.z + wb.w * x.w));
            if (lane == 0) { s_g[n] = gq + w.film_g_b[n]; s_q[n] = qq + w.film_b_b[n]; }
        }
    }
    __syncthreads();
    for (int n = tid; n < 1024; n += 256) {
        if (n >= nf) {           // unused rows of the record: the weight-gradient gathers stop at nf, keep them defined anyway
            o[SB_DG + n] = 0.f; o[SB_DQ + n] = 0.f;
            continue;
        }
        const int k = n >> 7, i = n & 127;
        const float pg = w.post_mult ? w.post_mult[k] : 1.f, pb = w.post_mult ? w.post_mult[n_film + k] : 1.f;
        const float df_g = dfilm[((int64_t)b * 8 + k) * 256 + i], df_b = dfilm[((int64_t)b * 8 + k) * 256 + 128 + i];
        s_dg[n] = df_g * pg; s_dq[n] = df_b * pb;
        o[SB_DG + n] = s_dg[n]; o[SB_DQ + n] = s_dq[n];
        s_g[n] = df_g * s_g[n];          // contributions to d post_mult[0][k]
        s_q[n] = df_b * s_q[n];
    }
    __syncthreads();
    if (tid < 16) {              // record slot tid = element tid of the flat (2, n_film) post-multiplier gradient
        float acc = 0.f;
        if (tid < 2 * n_film) {
            const int k = tid % n_film;
            const float* src = (tid < n_film ? s_g : s_q) + k * 128;
            for (int i = 0; i < 128; ++i) acc += src[i];
        }
        o[SB_DPOST + tid] = acc;
    }
    // dc[k] = sum_n Wg[n][k] dg[n] + Wb[n][k] dq[n]: thread k (< 128) walks the columns (coalesced over k)
    if (tid < DX_D) {
        float acc = 0.f;
        for (int n = 0; n < nf; ++n) acc += w.film_g_w[(int64_t)n * DX_D + tid] * s_dg[n] + w.film_b_w[(int64_t)n * DX_D + tid] * s_dq[n];
        s_t[tid] = acc;
        o[SB_DC + tid] = acc;
    }
    __syncthreads();
    // classifier forward (GRL = identity) and backward
    for (int n = warp; n < DX_D; n += 8) {
        const float4 wv = reinterpret_cast<const float4*>(w.cls_w1 + n * DX_D)[lane], x = reinterpret_cast<const float4*>(s_a)[lane];
        const float a = warp_sum((wv.x * x.x + wv.y * x.y) + (wv.z * x.z + wv.w * x.w));
        if (lane == 0) s_h1[n] = fmaxf(a + w.cls_b1[n], 0.f);
    }
    __syncthreads();
    for (int n = warp; n < DX_D; n += 8) {
        const float4 wv = reinterpret_cast<const float4*>(w.cls_w2 + n * DX_D)[lane], x = reinterpret_cast<const float4*>(s_h1)[lane];
        const float a = warp_sum((wv.x * x.x + wv.y * x.y) + (wv.z * x.z + wv.w * x.w));
        if (lane == 0) s_h2[n] = fmaxf(a + w.cls_b2[n], 0.f);
    }
    __syncthreads();
    if (tid < DX_D) {
        float acc = 0.f;
        for (int n = 0; n < w.n_speakers; ++n) acc += w.cls_w3[n * DX_D + tid] * dlogits[(int64_t)b * w.n_speakers + n];
        s_dh2[tid] = s_h2[tid] > 0.f ? acc : 0.f;
        o[SB_H2 + tid] = s_h2[tid]; o[SB_DH2 + tid] = s_dh2[tid];
    }
    __syncthreads();
    if (tid < DX_D) {
        float acc = 0.f;
        for (int n = 0; n < DX_D; ++n) acc += w.cls_w2[n * DX_D + tid] * s_dh2[n];
        s_dh1[tid] = s_h1[tid] > 0.f ? acc : 0.f;
        o[SB_H1 + tid] = s_h1[tid]; o[SB_DH1 + tid] = s_dh1[tid];
    }
    __syncthreads();
    if (tid < DX_D) {
        float acc = 0.f;
        for (int n = 0; n < DX_D; ++n) acc += w.cls_w1[n * DX_D + tid] * s_dh1[n];
        o[SB_DACC + tid] = s_t[tid] - lambda_rev * acc;       // d accent = d c (FiLM path) + GRL(-lambda) of the classifier path
    }
}
int style_bwd(const float* spk_embs, const float* accent, const DxModelWeights* w, const float* dfilm, const float* dlogits,
              float lambda_rev, int B, float* v, cudaStream_t s) {
    DX_LAUNCH_PDL((style_bwd_kernel), B, 256, 0, s, spk_embs, accent, *w, dfilm, dlogits, lambda_rev, v);
    DX_CHECK_LAUNCH();
    return 0;
}

// gather a strided per-utterance vector into a dense [B][n] matrix
__global__ void gather_strided_kernel(const float* __restrict__ v, int64_t stride, int off, int n, int B, float* __restrict__ out) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)B * n) return;
    out[i] = v[(i / n) * stride + off + (i % n)];
}
int gather_strided(const float* v, int64_t stride, int off, int n, int B, float* out, cudaStream_t s) {
    DX_LAUNCH_PDL((gather_strided_kernel), (unsigned)(((int64_t)B * n + 255) / 256), 256, 0, s, v, stride, off, n, B, out);
    DX_CHECK_LAUNCH();
    return 0;
}

// =========================================================================================================
// Loss (DaftExprtLoss.forward, loss.py:57-159; pitch-consistency off) and its gradient w.r.t. the mel rows.
// One CTA per utterance: mel L1 + L2 over the PADDED (80, T_max) plane divided by 80 len_b (loss.py:99-104), the
// energy-consistency MSE of 5-frame average-pooled ||exp(mel)||_2 over valid frames (:116-131), the speaker
// cross-entropy row (:85-88); partial sums per utterance, combined in order by loss_finish_kernel.
// part[b] = { l1_b, l2_b, e_b, ce_b }.   dmel rows [cap_t][80]: gradient at valid rows, 0 elsewhere.
// =========================================================================================================
__global__ void __launch_bounds__(256)
loss_utt_kernel(const float* __restrict__ melo, const float* __restrict__ tgt, Rows RT, const float* __restrict__ logits,
                const int64_t* __restrict__ speaker_ids, int n_spk, DxLossParams lp, float* __restrict__ part,
                float* __restrict__ dmel, float* __restrict__ dlogits) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    extern __shared__ __align__(16) float ls[];
    const int T = RT.n_max, B = RT.B;
    float* s_ep = ls;             // [T + 4] pooled-energy inputs of the prediction (2 zero frames each side)
    float* s_et = s_ep + T + 4;   // [T + 4] target
    float* s_de = s_et + T + 4;   // [T + 4] d loss / d e_p
    __shared__ float s_w[3][8];
    __shared__ float s_tot;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int len = (int)min((long long)RT.len[b], (long long)T);
    const int64_t rs = RT.row_start[b];
    if (tid == 0) {
        float tl = 0.f;
        for (int i = 0; i < B; ++i) tl += (float)min((long long)RT.len[i], (long long)T);
        s_tot = tl;
    }
    for (int i = tid; i < 2; i += 256) { s_ep[i] = 0.f; s_et[i] = 0.f; s_ep[T + 2 + i] = 0.f; s_et[T + 2 + i] = 0.f; }
    // pass 1: per frame (thread <-> frame): L1 / L2 partial sums and the energies
    float l1 = 0.f, l2 = 0.f;
    const float* tb = tgt + (int64_t)b * DX_MEL * T;
    for (int t = tid; t < T; t += 256) {
        float ep = 0.f, et = 0.f;
        for (int c = 0; c < DX_MEL; ++c) {
            const float p = (t < len) ? melo[(rs + t) * DX_MEL + c] : 0.f;
            const float y = tb[(int64_t)c * T + t];
            const float d = p - y;
            l1 += fabsf(d);
            l2 += d * d;
            const float xp = expf(p), xt = expf(y);
            ep += xp * xp;
            et += xt * xt;
        }
        s_ep[t + 2] = sqrtf(ep);
        s_et[t + 2] = sqrtf(et);
    }
    __syncthreads();
    // pass 2: pooled difference at valid frames -> e loss and d/ds
    float el = 0.f;
    for (int t = tid; t < T; t += 256) {
        float ds = 0.f;
        if (t < len) {
            const float sp = (((s_ep[t] + s_ep[t + 1]) + s_ep[t + 2]) + s_ep[t + 3] + s_ep[t + 4]) * 0.2f;
            const float st = (((s_et[t] + s_et[t + 1]) + s_et[t + 2]) + s_et[t + 3] + s_et[t + 4]) * 0.2f;
            const float d = sp - st;
            el += d * d;
            ds = 2.f * d;
        }
        s_de[t + 2] = ds;
    }
    for (int i = tid; i < 2; i += 256) { s_de[i] = 0.f; s_de[T + 2 + i] = 0.f; }
    l1 = warp_sum(l1); l2 = warp_sum(l2); el = warp_sum(el);
    if (lane == 0) { s_w[0][warp] = l1; s_w[1][warp] = l2; s_w[2][warp] = el; }
    __syncthreads();
    const float inv_norm = 1.f / ((float)DX_MEL * (float)RT.len[b]);
    if (tid == 0) {
        float a = 0.f, c = 0.f, e = 0.f;
        for (int i = 0; i < 8; ++i) { a += s_w[0][i]; c += s_w[1][i]; e += s_w[2][i]; }
        part[b * 4 + 0] = a * inv_norm;
        part[b * 4 + 1] = c * inv_norm;
        part[b * 4 + 2] = e;
        // cross-entropy row
        float ce = 0.f;
        if (logits != nullptr) {
            const float* lg = logits + (int64_t)b * n_spk;
            float mx = -INFINITY;
            for (int n = 0; n < n_spk; ++n) mx = fmaxf(mx, lg[n]);
            float se = 0.f;
            for (int n = 0; n < n_spk; ++n) se += expf(lg[n] - mx);
            const float lz = mx + logf(se);
            const int y = (int)speaker_ids[b];
            ce = lz - lg[y];
            for (int n = 0; n < n_spk; ++n)
                dlogits[(int64_t)b * n_spk + n] = lp.scale * lp.speaker_weight * (expf(lg[n] - lz) - (n == y ? 1.f : 0.f)) / (float)B;
        }
        part[b * 4 + 3] = ce;
    }
    // pass 3: gradient rows.  d/dpred = w_mel / (B 80 len) (sign(d) + 2 d) + w_e / sum_len * de[t] * exp(2 p) / e_p[t]
    const float cm = lp.scale * lp.mel_spec_weight * inv_norm / (float)B;
    const float ce_w = lp.scale * lp.energy_weight / s_tot;
    for (int i = tid; i < len * DX_MEL; i += 256) {
        const int t = i / DX_MEL, c = i % DX_MEL;
        const float p = melo[(rs + t) * DX_MEL + c];
        const float d = p - tb[(int64_t)c * T + t];
        float gsum = cm * ((d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f)) + 2.f * d);
        if (lp.energy_weight > 0.f) {
            const float de = (((s_de[t] + s_de[t + 1]) + s_de[t + 2]) + s_de[t + 3] + s_de[t + 4]) * 0.2f;
            const float xp = expf(p);
            gsum += ce_w * de * xp * xp / s_ep[t + 2];
        }
        dmel[(rs + t) * DX_MEL + c] = gsum;
    }
}

// losses[8]: total, speaker_loss, speaker_ce_raw, post_mult_loss, mel_l1, mel_l2, energy_consistency, pitch_consistency (0)
__global__ void loss_finish_kernel(const float* __restrict__ part, int B, Rows RT, const float* __restrict__ post_mult, int n_post,
                                   DxLossParams lp, float* __restrict__ losses, float* __restrict__ dpost_reg) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    float l1 = 0.f, l2 = 0.f, e = 0.f, ce = 0.f, tl = 0.f;
    for (int b = 0; b < B; ++b) {
        l1 += part[b * 4 + 0]; l2 += part[b * 4 + 1]; e += part[b * 4 + 2]; ce += part[b * 4 + 3];
        tl += (float)min((long long)RT.len[b], (long long)RT.n_max);
    }
    l1 = lp.mel_spec_weight * (l1 / (float)B);
    l2 = lp.mel_spec_weight * (l2 / (float)B);
    ce = ce / (float)B;
    e = e / tl;
    float pm = 0.f;
    if (post_mult != nullptr && lp.post_mult_weight != 0.f) {
        float ss = 0.f;
        for (int i = 0; i < n_post; ++i) ss += post_mult[i] * post_mult[i];
        const float nrm = sqrtf(ss);
        pm = lp.post_mult_weight * nrm;
        for (int i = 0; i < 16; ++i) dpost_reg[i] = (i < n_post && nrm > 0.f) ? lp.scale * lp.post_mult_weight * post_mult[i] / nrm : 0.f;
    } else if (dpost_reg != nullptr) {
        for (int i = 0; i < 16; ++i) dpost_reg[i] = 0.f;
    }
    const float spk = lp.speaker_weight * ce;
    float total = spk + pm + l1 + l2;
    if (lp.energy_weight > 0.f) total += lp.energy_weight * e;
    losses[0] = total * lp.scale; losses[1] = spk; losses[2] = ce; losses[3] = pm; losses[4] = l1; losses[5] = l2;
    losses[6] = lp.energy_weight > 0.f ? e : 0.f; losses[7] = 0.f;
}

int loss_forward_backward(const float* melo, const float* tgt, const Rows& RT, const float* logits, const int64_t* speaker_ids,
                          int n_spk, const float* post_mult, int n_post, const DxLossParams& lp, float* part, float* dmel,
                          float* dlogits, float* losses, float* dpost_reg, cudaStream_t s) {
    const size_t smem = (size_t)3 * (RT.n_max + 4) * sizeof(float);
    DX_REQUIRE(smem <= 160 * 1024, "loss: T_max=%d too large", RT.n_max);
    static SmemConfigured cfg;
    if (smem > 40 * 1024) DX_TRY(ensure_dyn_smem(loss_utt_kernel, smem, cfg));
    DX_CHECK_CUDA(cudaMemsetAsync(dmel, 0, (size_t)RT.cap * DX_MEL * sizeof(float), s));
    ProfScope prof("loss", s);
    DX_LAUNCH_PDL((loss_utt_kernel), RT.B, 256, smem, s, melo, tgt, RT, logits, speaker_ids, n_spk, lp, part, dmel, dlogits);
    DX_CHECK_LAUNCH();
    DX_LAUNCH_PDL((loss_finish_kernel), 1, 32, 0, s, part, RT.B, RT, post_mult, n_post, lp, losses, dpost_reg);
    DX_CHECK_LAUNCH();
    return 0;
}

// d post_mult = sum_b dpost_b (FiLM path) + the regulariser's gradient
__global__ void post_mult_grad_kernel(const float* __restrict__ v, int B, const float* __restrict__ dreg, int n_post,
                                      float* __restrict__ out) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const int i = threadIdx.x;
    if (i >= n_post) return;
    float acc = 0.f;
    for (int b = 0; b < B; ++b) acc += v[(int64_t)b * SB_STRIDE + SB_DPOST + i];
    out[i] = acc + dreg[i];
}
int post_mult_grad(const float* v, int B, const float* dreg, int n_post, float* out, cudaStream_t s) {
    DX_LAUNCH_PDL((post_mult_grad_kernel), 1, 32, 0, s, v, B, dreg, n_post, out);
    DX_CHECK_LAUNCH();
    return 0;
}

// =========================================================================================================
// Adam (torch.optim.Adam, amsgrad=False, L2 weight decay folded into the gradient; train.py:278-280) on flat buffers.
// grad_scale: 1 / world_size after a SUM all-reduce (DDP averages gradients).
// =========================================================================================================
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                            int64_t n, float lr, float b1, float b2, float eps, float wd, float bc1, float bc2_sqrt, float gscale) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float gi = g[i] * gscale;
    const float pi = p[i];
    gi = fmaf(wd, pi, gi);
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi; v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = pi - (lr / bc1) * (mi / denom);
}
int adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float b1, float b2, float eps, float wd, int step,
              float gscale, cudaStream_t s) {
    DX_REQUIRE(step >= 1, "adam_step: step counts from 1");
    const float bc1 = 1.f - powf(b1, (float)step), bc2 = 1.f - powf(b2, (float)step);
    ProfScope prof("adam", s);
    DX_LAUNCH_PDL((adam_kernel), (unsigned)((n + 255) / 256), 256, 0, s, p, g, m, v, n, lr, b1, b2, eps, wd, bc1, sqrtf(bc2), gscale);
    DX_CHECK_LAUNCH();
    return 0;
}

}  // namespace dx
