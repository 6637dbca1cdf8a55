// fp32 (parity-mode) kernels: SIMT conv-as-GEMM on the packed row stream, row LayerNorm/FiLM/mask, and a
// flash-style masked multi-head attention.  These back DX_MODE_F32, whose gate is mel max-abs <= 1e-3 against
// the reference CPU forward (TF32 tensor-core inputs miss that gate: SURVEY.md §7.3 H2).
#include "dx_common.cuh"
#include "dx_dropout.cuh"
#include "dx_ptx.cuh"
#include "dx_kernels.cuh"

namespace dx {

// =========================================================================================================
// gemm_f32: out[r][n] = act(bias[n] + sum_j sum_k A[r + j - taps/2][k] * W[j][k][n])
// Reference ops: nn.Conv1d k=3 "same" (ConvNorm1D, model.py:75-94) and nn.Linear (LinearNorm, model.py:57-72).
// 128x64 tile, BK=16, 256 threads, 8x4 outputs per thread, register-prefetch double buffering.
// =========================================================================================================
constexpr int GF_BM = 128, GF_BN = 64, GF_BK = 16;

__global__ void __launch_bounds__(256)
gemm_f32_kernel(const float* __restrict__ A, int K, const float* __restrict__ W, const float* __restrict__ bias,
                int N, int taps, int relu, Rows R, int live_ext, float* __restrict__ out, const float* __restrict__ addend) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    __shared__ __align__(16) float As[2][GF_BK][GF_BM + 4];
    __shared__ __align__(16) float Bs[2][GF_BK][GF_BN];
    const int64_t total = R.row_start[R.B];
    const int64_t r0 = (int64_t)blockIdx.x * GF_BM;
    if (r0 >= total) return;
    const int n0 = blockIdx.y * GF_BN;
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int half = taps / 2;
    const int ksteps = K / GF_BK;
    const int nsteps = taps * ksteps;

    // loader coordinates
    const int a_row[2] = {(tid) >> 2, (tid + 256) >> 2};  // 0..127
    const int a_kq = tid & 3;                             // float4 index along K (same for both)
    const int b_kr = tid >> 4, b_nq = tid & 15;           // 16 x 16 float4

    float4 a_reg[2], b_reg;
    auto load_step = [&](int step) {
        const int j = step / ksteps, k0 = (step % ksteps) * GF_BK;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            int64_t r = r0 + a_row[i] + j - half;
            a_reg[i] = (r >= 0 && r < total)
                           ? *reinterpret_cast<const float4*>(A + r * K + k0 + a_kq * 4)
                           : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        const int n = n0 + b_nq * 4;
        const float* wp = W + ((int64_t)j * K + k0 + b_kr) * N + n;
        if (n + 3 < N) b_reg = *reinterpret_cast<const float4*>(wp);
        else {
            b_reg.x = (n < N) ? wp[0] : 0.f;
            b_reg.y = (n + 1 < N) ? wp[1] : 0.f;
            b_reg.z = (n + 2 < N) ? wp[2] : 0.f;
            b_reg.w = 0.f;
        }
    };
    auto store_step = [&](int buf) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            As[buf][a_kq * 4 + 0][a_row[i]] = a_reg[i].x;
            As[buf][a_kq * 4 + 1][a_row[i]] = a_reg[i].y;
            As[buf][a_kq * 4 + 2][a_row[i]] = a_reg[i].z;
            As[buf][a_kq * 4 + 3][a_row[i]] = a_reg[i].w;
        }
        *reinterpret_cast<float4*>(&Bs[buf][b_kr][b_nq * 4]) = b_reg;
    };

    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    load_step(0);
    store_step(0);
    __syncthreads();
    for (int step = 0; step < nsteps; ++step) {
        const int buf = step & 1;
        if (step + 1 < nsteps) load_step(step + 1);
#pragma unroll
        for (int k = 0; k < GF_BK; ++k) {
            float4 a0 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 8]);
            float4 a1 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 8 + 4]);
            float4 b = *reinterpret_cast<const float4*>(&Bs[buf][k][tx * 4]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        if (step + 1 < nsteps) {
            store_step(buf ^ 1);
            __syncthreads();
        }
    }
    // epilogue
    const int n = n0 + tx * 4;
    float bz[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) bz[j] = (bias != nullptr && n + j < N) ? bias[n + j] : 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int64_t r = r0 + ty * 8 + i;
        if (r >= R.cap) continue;
        const bool live = (live_ext < 0) ? (r < total) : (r < total && row_live(R, r, live_ext));
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float t = acc[i][j] + bz[j];
            if (relu) t = fmaxf(t, 0.f);
            if (addend != nullptr && n + j < N) t += addend[r * N + n + j];      // training: residual-gradient accumulation
            v[j] = live ? t : 0.f;
        }
        if (n + 3 < N) *reinterpret_cast<float4*>(out + r * N + n) = make_float4(v[0], v[1], v[2], v[3]);
        else
            for (int j = 0; j < 4; ++j)
                if (n + j < N) out[r * N + n + j] = v[j];
    }
}

int gemm_f32(const float* A, int K, const float* W, const float* bias, int N, int taps, bool relu,
             const Rows& R, int live_ext, float* out, cudaStream_t s, const float* addend) {
    DX_REQUIRE(K % GF_BK == 0 && N % 4 == 0, "gemm_f32: K=%d must be a multiple of 16 and N=%d of 4", K, N);
    dim3 grid((unsigned)(R.cap / GF_BM), (unsigned)((N + GF_BN - 1) / GF_BN));
    ProfScope prof("gemm_f32", s);
    DX_LAUNCH_PDL((gemm_f32_kernel), grid, 256, 0, s, A, K, W, bias, N, taps, relu ? 1 : 0, R, live_ext, out, addend);
    DX_CHECK_LAUNCH();
    return 0;
}

// =========================================================================================================
// ln_rows_f32: out = live ? film(LN(in + res) * g + b) : 0.   One warp per row, values held in registers,
// two-pass mean / variance (biased, eps 1e-5) like torch's native_layer_norm.
// Reference: MultiHeadAttention.forward model.py:191 + FFTBlock mask :256; PositionWiseConvFF.forward
// model.py:226-233 + mask :258; accent pre-net LayerNorms model.py:655,662,669.
// =========================================================================================================
template <int C, bool DROP>
__global__ void ln_rows_f32_kernel(const float* __restrict__ in, const float* __restrict__ res,
                                   const float* __restrict__ g, const float* __restrict__ bta,
                                   const float* __restrict__ film, int film_stride, Rows R, int live_ext,
                                   float* __restrict__ out, LnDrop drop) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    constexpr int V = C / 128;  // float4 per lane
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= R.cap) return;
    const int64_t total = R.row_start[R.B];
    float4* orow = reinterpret_cast<float4*>(out + r * C);
    const bool live = (r < total) && row_live(R, r, live_ext);
    if (!live) {
#pragma unroll
        for (int i = 0; i < V; ++i) orow[lane + 32 * i] = make_float4(0.f, 0.f, 0.f, 0.f);
        return;
    }
    float4 v[V];
    const float4* irow = reinterpret_cast<const float4*>(in + r * C);
    float sum = 0.f;
    const int b_idx = R.row_b[r];
    const int pos = DROP ? R.row_p[r] : 0;
#pragma unroll
    for (int i = 0; i < V; ++i) {
        v[i] = irow[lane + 32 * i];
        if constexpr (DROP) {
            if (drop.where == LN_DROP_IN) {         // training: dropout of the branch before the residual add, kept for the backward
                v[i] = drop_apply4(v[i], drop.d, drop.site, b_idx, pos, R.n_max, C, (lane + 32 * i) * 4);
                reinterpret_cast<float4*>(const_cast<float*>(in) + r * C)[lane + 32 * i] = v[i];
            }
        }
        if (res != nullptr) {
            float4 q = reinterpret_cast<const float4*>(res + r * C)[lane + 32 * i];
            v[i].x += q.x; v[i].y += q.y; v[i].z += q.z; v[i].w += q.w;
        }
        sum += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    }
    const float mean = warp_sum(sum) * (1.f / C);
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < V; ++i) {
        float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
        sq += (a * a + b * b) + (c * c + d * d);
    }
    const float rstd = rsqrtf(warp_sum(sq) * (1.f / C) + 1e-5f);
#pragma unroll
    for (int i = 0; i < V; ++i) {
        const int c4 = lane + 32 * i;
        float4 gg = reinterpret_cast<const float4*>(g)[c4];
        float4 bb = reinterpret_cast<const float4*>(bta)[c4];
        float4 y;
        y.x = (v[i].x - mean) * rstd * gg.x + bb.x;
        y.y = (v[i].y - mean) * rstd * gg.y + bb.y;
        y.z = (v[i].z - mean) * rstd * gg.z + bb.z;
        y.w = (v[i].w - mean) * rstd * gg.w + bb.w;
        if (film != nullptr) {
            const float* fp = film + (int64_t)b_idx * film_stride;
            float4 ga = reinterpret_cast<const float4*>(fp)[c4];
            float4 be = reinterpret_cast<const float4*>(fp + C)[c4];
            y.x = ga.x * y.x + be.x; y.y = ga.y * y.y + be.y; y.z = ga.z * y.z + be.z; y.w = ga.w * y.w + be.w;
        }
        if constexpr (DROP) {
            if (drop.where == LN_DROP_OUT) y = drop_apply4(y, drop.d, drop.site, b_idx, pos, R.n_max, C, c4 * 4);
        }
        orow[c4] = y;
    }
}

int ln_rows_f32(const float* in, const float* res, const float* g, const float* b, int C, const float* film,
                int film_stride, const Rows& R, int live_ext, float* out, cudaStream_t s, const LnDrop* drop) {
    int grid = (int)((R.cap + 7) / 8);
    ProfScope prof("ln_rows_f32", s);
    LnDrop nd;
    memset(&nd, 0, sizeof(nd));
    if (drop != nullptr && drop->d.thr != 0) {
        DX_REQUIRE(in != out && in != res, "ln_rows_f32: the dropped input is rewritten in place and cannot alias res / out");
        if (C == 128)
            DX_LAUNCH_PDL((ln_rows_f32_kernel<128, true>), grid, 256, 0, s, in, res, g, b, film, film_stride, R, live_ext, out, *drop);
        else if (C == 1024)
            DX_LAUNCH_PDL((ln_rows_f32_kernel<1024, true>), grid, 256, 0, s, in, res, g, b, film, film_stride, R, live_ext, out, *drop);
        else
            DX_REQUIRE(false, "ln_rows_f32: unsupported width %d", C);
    } else if (C == 128)
        DX_LAUNCH_PDL((ln_rows_f32_kernel<128, false>), grid, 256, 0, s, in, res, g, b, film, film_stride, R, live_ext, out, nd);
    else if (C == 1024)
        DX_LAUNCH_PDL((ln_rows_f32_kernel<1024, false>), grid, 256, 0, s, in, res, g, b, film, film_stride, R, live_ext, out, nd);
    else
        DX_REQUIRE(false, "ln_rows_f32: unsupported width %d", C);
    DX_CHECK_LAUNCH();
    return 0;
}

// =========================================================================================================
// attention_f32: per (utterance, head, 64-query tile), keys streamed in tiles of 64 with online softmax.
// Reference: nn.MultiheadAttention slow path as called at model.py:182-186 — scores = (q * d^-0.5) k^T with
// -inf at padded keys, softmax, · v.  (q * 0.125 is exact, so scaling the score instead is bit-equivalent.)
// The head-averaged weights torch returns are discarded by the reference (model.py:255) and not computed.
// qkv rows: [q(128) | k(128) | v(128)], head h uses columns [64h, 64h+64) of each.
// =========================================================================================================
constexpr int AT_Q = 64, AT_K = 64;

__global__ void __launch_bounds__(256)
attention_f32_kernel(const float* __restrict__ qkv, Rows R, float* __restrict__ out, float* __restrict__ lse) {
    ptx::griddep_wait();      // PDL: the previous kernel of the stream is complete and visible
    ptx::griddep_launch();    // let the next kernel start launching
    extern __shared__ __align__(16) float smem[];
    float(*Qt)[AT_Q] = reinterpret_cast<float(*)[AT_Q]>(smem);                          // [64 d][64 q]
    float(*Kt)[AT_K + 1] = reinterpret_cast<float(*)[AT_K + 1]>(smem + 64 * AT_Q);      // [64 d][65]
    float(*Vs)[DX_DH] = reinterpret_cast<float(*)[DX_DH]>(smem + 64 * AT_Q + 64 * (AT_K + 1));   // [64 key][64 d]
    float(*Pt)[AT_Q + 4] = reinterpret_cast<float(*)[AT_Q + 4]>(smem + 64 * AT_Q + 64 * (AT_K + 1) + AT_K * DX_DH);

    const int b = blockIdx.z, h = blockIdx.y;
    const int len = (int)min((long long)R.len[b], (long long)R.n_max);
    const int q0 = blockIdx.x * AT_Q;
    if (q0 >= len) return;
    const int64_t rs = R.row_start[b];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;

    // load Q tile transposed, pre-scaled
    for (int i = tid; i < AT_Q * 16; i += 256) {
        int q = i >> 4, d4 = i & 15;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (q0 + q < len) v = *reinterpret_cast<const float4*>(qkv + (rs + q0 + q) * 384 + h * DX_DH + d4 * 4);
        Qt[d4 * 4 + 0][q] = v.x * 0.125f;
        Qt[d4 * 4 + 1][q] = v.y * 0.125f;
        Qt[d4 * 4 + 2][q] = v.z * 0.125f;
        Qt[d4 * 4 + 3][q] = v.w * 0.125f;
    }
    float o[4][4], m[4], l[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        m[i] = -INFINITY;
        l[i] = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) o[i][j] = 0.f;
    }
    for (int k0 = 0; k0 < len; k0 += AT_K) {
        __syncthreads();  // previous tile's Pt / Vs reads done (and Qt visible on first pass)
        for (int i = tid; i < AT_K * 16; i += 256) {
            int kk = i >> 4, d4 = i & 15;
            float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
            if (k0 + kk < len) {
                const float* base = qkv + (rs + k0 + kk) * 384 + h * DX_DH + d4 * 4;
                kv = *reinterpret_cast<const float4*>(base + 128);
                vv = *reinterpret_cast<const float4*>(base + 256);
            }
            Kt[d4 * 4 + 0][kk] = kv.x; Kt[d4 * 4 + 1][kk] = kv.y; Kt[d4 * 4 + 2][kk] = kv.z; Kt[d4 * 4 + 3][kk] = kv.w;
            *reinterpret_cast<float4*>(&Vs[kk][d4 * 4]) = vv;
        }
        __syncthreads();
        // S block: rows ty*4+i, keys tx*4+j
        float sc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) sc[i][j] = 0.f;
#pragma unroll 8
        for (int d = 0; d < DX_DH; ++d) {
            float4 qa = *reinterpret_cast<const float4*>(&Qt[d][ty * 4]);
            const float qv[4] = {qa.x, qa.y, qa.z, qa.w};
            const float kv[4] = {Kt[d][tx * 4], Kt[d][tx * 4 + 1], Kt[d][tx * 4 + 2], Kt[d][tx * 4 + 3]};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) sc[i][j] = fmaf(qv[i], kv[j], sc[i][j]);
        }
        // online softmax per row (row owned by the 16 threads sharing ty → xor-shuffles over 1,2,4,8)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float mx = -INFINITY;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (k0 + tx * 4 + j >= len) sc[i][j] = -INFINITY;
                mx = fmaxf(mx, sc[i][j]);
            }
#pragma unroll
            for (int ofs = 8; ofs > 0; ofs >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, ofs));
            const float m_new = fmaxf(m[i], mx);   // finite: every key tile holds >= 1 valid key
            const float corr = expf(m[i] - m_new);
            float ps = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float p = expf(sc[i][j] - m_new);
                ps += p;
                Pt[tx * 4 + j][ty * 4 + i] = p;
            }
#pragma unroll
            for (int ofs = 8; ofs > 0; ofs >>= 1) ps += __shfl_xor_sync(0xffffffffu, ps, ofs);
            l[i] = l[i] * corr + ps;
            m[i] = m_new;
#pragma unroll
            for (int j = 0; j < 4; ++j) o[i][j] *= corr;
        }
        __syncthreads();
        // O block: rows ty*4+i, dims tx*4+j
#pragma unroll 8
        for (int kk = 0; kk < AT_K; ++kk) {
            float4 pa = *reinterpret_cast<const float4*>(&Pt[kk][ty * 4]);
            float4 va = *reinterpret_cast<const float4*>(&Vs[kk][tx * 4]);
            const float pv[4] = {pa.x, pa.y, pa.z, pa.w};
            const float vv[4] = {va.x, va.y, va.z, va.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) o[i][j] = fmaf(pv[i], vv[j], o[i][j]);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int q = q0 + ty * 4 + i;
        if (q < len) {
            const float inv = 1.f / l[i];
            *reinterpret_cast<float4*>(out + (rs + q) * DX_D + h * DX_DH + tx * 4) =
                make_float4(o[i][0] * inv, o[i][1] * inv, o[i][2] * inv, o[i][3] * inv);
            // training: log-sum-exp of the (scaled) scores of this query row, for the backward's recomputation of P
            if (lse != nullptr && tx == 0) lse[(rs + q) * DX_HEADS + h] = m[i] + logf(l[i]);
        }
    }
    // the utterance's gap rows get zeros: the training step's weight gradients run over every row of the stream
    // (d out_proj.weight = O^T dT1 with dT1 = 0 there), and 0 x <whatever the workspace held> must stay 0
    if (blockIdx.x == 0 && tid < DX_GAP * 16) {
        const int64_t r = rs + len + tid / 16;
        *reinterpret_cast<float4*>(out + r * DX_D + h * DX_DH + (tid & 15) * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
        if (lse != nullptr && (tid & 15) == 0) lse[r * DX_HEADS + h] = 0.f;
    }
}

static const int AT_SMEM = (64 * AT_Q + 64 * (AT_K + 1) + AT_K * DX_DH + AT_K * (AT_Q + 4)) * (int)sizeof(float);

int attention_f32(const float* qkv, const Rows& R, float* out, cudaStream_t s, float* lse) {
    static SmemConfigured configured;
    DX_TRY(ensure_dyn_smem(attention_f32_kernel, (size_t)AT_SMEM, configured));
    dim3 grid((unsigned)((R.n_max + AT_Q - 1) / AT_Q), DX_HEADS, (unsigned)R.B);
    ProfScope prof("attn_f32", s);
    DX_LAUNCH_PDL((attention_f32_kernel), grid, 256, AT_SMEM, s, qkv, R, out, lse);
    DX_CHECK_LAUNCH();
    return 0;
}

}  // namespace dx
