// Length regulator = GaussianUpsamplingModule.forward (model.py:417-510) and the integer plumbing around it
// (get_int_durations model.py:950-973 / duration_to_integer extract_features.py:69-125, output_lengths
// model.py:1096-1100).  Integer quantities are int64 and bit-exact; σ, the Gaussian weights, their
// normalisation and the contraction stay fp32 in both arithmetic modes.
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_ptx.cuh"

namespace dx {

__device__ __forceinline__ float4 conv3_scalar_in(const float* __restrict__ x, int b, int p, int n_max,
                                                  const float* __restrict__ w, const float* __restrict__ bias,
                                                  int lane) {
    const float* xr = x + (int64_t)b * n_max;
    float xm = (p - 1 >= 0) ? xr[p - 1] : 0.f;
    float x0 = xr[p];
    float xp = (p + 1 < n_max) ? xr[p + 1] : 0.f;
    float o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int c = lane * 4 + i;
        o[i] = bias[c] + w[c * 3 + 0] * xm + w[c * 3 + 1] * x0 + w[c * 3 + 2] * xp;
    }
    return make_float4(o[0], o[1], o[2], o[3]);
}

// ---------------------------------------------------------------------------------------------------------
// Stage 1 (per symbol row): x' = x + conv(energy) + conv(pitch); σ = max(softplus(w·(x' + conv(dur)) + b), 1e-3)
// (model.py:430-482; pad symbols get σ = 1 but never contribute because their probs are masked, :503).
// One warp per row.  xs: fp32 [rowsL][128];  sigma: fp32 (B, L_max), 1 at pads.
// ---------------------------------------------------------------------------------------------------------
template <bool BF16>
__global__ void lenreg_symbols_kernel(const float* __restrict__ x_f32, const __nv_bfloat16* __restrict__ x_bf,
                                      int64_t rows_phys, Rows R, const float* __restrict__ dur,
                                      const float* __restrict__ energy, const float* __restrict__ pitch,
                                      DxModelWeights w, float* __restrict__ xs, float* __restrict__ sigma) {
    ptx::griddep_wait();      // PDL: predecessor complete + visible
    ptx::griddep_launch();    // let the next kernel start its prologue
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= R.cap) return;
    const int b = R.row_b[r];
    if (b < 0) return;
    const int p = R.row_p[r];
    if (p >= R.n_max) return;
    if (p >= R.len[b]) {
        if (lane == 0) sigma[(int64_t)b * R.n_max + p] = 1.f;
        return;
    }
    float4 x;
    if (!BF16) {
        x = reinterpret_cast<const float4*>(x_f32 + r * DX_D)[lane];
    } else {
        int64_t off = ((int64_t)(lane >> 1) * rows_phys + DX_HALO + r) * 8 + (lane & 1) * 4;
        uint2 pk = *reinterpret_cast<const uint2*>(x_bf + off);
        __nv_bfloat162 lo = *reinterpret_cast<__nv_bfloat162*>(&pk.x);
        __nv_bfloat162 hi = *reinterpret_cast<__nv_bfloat162*>(&pk.y);
        x = make_float4(__low2float(lo), __high2float(lo), __low2float(hi), __high2float(hi));
    }
    float4 e = conv3_scalar_in(energy, b, p, R.n_max, w.gu_energy_w, w.gu_energy_b, lane);
    float4 q = conv3_scalar_in(pitch, b, p, R.n_max, w.gu_pitch_w, w.gu_pitch_b, lane);
    float4 d = conv3_scalar_in(dur, b, p, R.n_max, w.gu_dur_w, w.gu_dur_b, lane);
    x = make_float4((x.x + e.x) + q.x, (x.y + e.y) + q.y, (x.z + e.z) + q.z, (x.w + e.w) + q.w);
    reinterpret_cast<float4*>(xs + r * DX_D)[lane] = x;
    float4 wr = reinterpret_cast<const float4*>(w.gu_range_w)[lane];
    float dot = ((x.x + d.x) * wr.x + (x.y + d.y) * wr.y) + ((x.z + d.z) * wr.z + (x.w + d.w) * wr.w);
    dot = warp_sum(dot) + w.gu_range_b[0];
    if (lane == 0) {
        float sp = (dot > 20.f) ? dot : log1pf(expf(dot));   // nn.Softplus(beta=1, threshold=20)
        sigma[(int64_t)b * R.n_max + p] = fmaxf(sp, 1e-3f);
    }
}

int lenreg_symbols(const float* x_f32, const __nv_bfloat16* x_bf, int64_t rows_phys, const Rows& RL,
                   const float* dur, const float* energy, const float* pitch, const DxModelWeights* w, float* xs,
                   float* sigma, cudaStream_t s) {
    int grid = (int)((RL.cap + 7) / 8);
    if (x_bf)
        DX_LAUNCH_PDL((lenreg_symbols_kernel<true>), grid, 256, 0, s, nullptr, x_bf, rows_phys, RL, dur, energy, pitch, *w, xs, sigma);
    else
        DX_LAUNCH_PDL((lenreg_symbols_kernel<false>), grid, 256, 0, s, x_f32, nullptr, 0, RL, dur, energy, pitch, *w, xs, sigma);
    DX_CHECK_LAUNCH();
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Stage 2 (per utterance): int64 inclusive scan of durations_int over the whole padded row (as torch.cumsum,
// model.py:486), μ_l = float(d_l)/2 + float(cs_{l-1}) (:485-487), output_lengths = Σ d (:1096; clamp >= 1 when
// `clamp_one`, :1098).  One warp per utterance, warp-shuffle scan in int64.
// ---------------------------------------------------------------------------------------------------------
__global__ void lenreg_scan_kernel(const int64_t* __restrict__ dint, int B, int L, int clamp_one,
                                   int64_t* __restrict__ cumsum, float* __restrict__ mu,
                                   int64_t* __restrict__ out_len, int64_t* __restrict__ two_mu) {
    ptx::griddep_wait();      // PDL: predecessor complete + visible
    ptx::griddep_launch();    // let the next kernel start its prologue
    const int lane = threadIdx.x & 31;
    const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (b >= B) return;
    const int64_t* d = dint + (int64_t)b * L;
    long long carry = 0;
    for (int l0 = 0; l0 < L; l0 += 32) {
        const int l = l0 + lane;
        long long v = (l < L) ? d[l] : 0;
        long long incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            long long t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        incl += carry;
        if (l < L) {
            const long long prev = incl - v;
            if (cumsum) cumsum[(int64_t)b * L + l] = incl;
            if (two_mu) two_mu[(int64_t)b * L + l] = v + 2 * prev;
            if (mu) mu[(int64_t)b * L + l] = (float)v / 2.f + (float)prev;
        }
        carry = __shfl_sync(0xffffffffu, incl, 31);
    }
    if (lane == 0 && out_len) {
        long long t = carry;
        if (clamp_one && t == 0) t = 1;
        out_len[b] = t;
    }
}

int lenreg_scan(const int64_t* dint, int B, int L, bool clamp_one, int64_t* cumsum, float* mu, int64_t* out_len,
                int64_t* two_mu, cudaStream_t s) {
    DX_LAUNCH_PDL((lenreg_scan_kernel), (B + 7) / 8, 256, 0, s, dint, B, L, clamp_one ? 1 : 0, cumsum, mu, out_len, two_mu);
    DX_CHECK_LAUNCH();
    return 0;
}

// frame → symbol map: searchsorted(cumsum[b, :len_b], t, right=True) for t < out_len[b], else -1.
__global__ void frame2symbol_kernel(const int64_t* __restrict__ cumsum, const int64_t* __restrict__ in_len,
                                    const int64_t* __restrict__ out_len, int B, int L, int T,
                                    int64_t* __restrict__ f2s) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)B * T) return;
    const int b = (int)(i / T), t = (int)(i % T);
    long long res = -1;
    if (t < out_len[b]) {
        int n = (int)min((long long)in_len[b], (long long)L);
        const int64_t* cs = cumsum + (int64_t)b * L;
        int lo = 0, hi = n;  // first index with cs > t
        while (lo < hi) {
            int mid = (lo + hi) >> 1;
            if (cs[mid] > t) hi = mid; else lo = mid + 1;
        }
        res = lo;
    }
    f2s[i] = res;
}

int frame2symbol(const int64_t* cumsum, const int64_t* in_len, const int64_t* out_len, int B, int L, int T,
                 int64_t* f2s, cudaStream_t s) {
    int64_t n = (int64_t)B * T;
    if (n == 0) return 0;
    frame2symbol_kernel<<<(int)((n + 255) / 256), 256, 0, s>>>(cumsum, in_len, out_len, B, L, T, f2s);
    DX_CHECK_LAUNCH();
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Stage 3 (per utterance, 32-frame tile): p[l,t] = exp(N(μ_l, σ_l).log_prob(t + 0.5)), masked at pad symbols;
// w = p / (Σ_l p + 1e-20) → `weights` (B, L_max, T_max); x_up[t] = Σ_l w[l,t] x'[l] (model.py:495-507), then the
// FrameDecoder input x_up + P[t], masked (model.py:554-557), written into the frame row stream.
// Pass 1 sums the column denominators; pass 2 recomputes p (cheaper than holding an L x 32 tile), writes the
// weights coalesced along t and contracts in chunks of 32 symbols, skipping symbols whose 32 weights are all
// exactly 0 (fp32 exp underflows beyond ~14.4 σ, so the contraction is banded without changing the result).
// Symbols whose Gaussian cannot reach the tile at all (exponent below LR_ZERO_ARG at the nearest frame centre, where
// expf returns exactly 0) are flagged once per tile: they cost no expf in either pass, and a 32-symbol chunk without
// any flagged-in symbol only writes its zero weights (no barrier, no contraction).
// ---------------------------------------------------------------------------------------------------------
constexpr int LR_FT = 32;
constexpr float LR_LOG_SQRT_2PI = 0.91893853320467274178f;
constexpr float LR_ZERO_ARG = -110.f;      // expf(x) == 0 exactly for x < -103.98; margin for rounding of the argument

__device__ __forceinline__ float gauss_prob(float t_center, float mu, float sg, float log_sg) {
    const float d = t_center - mu;
    return expf(-(d * d) / (2.f * sg * sg) - log_sg - LR_LOG_SQRT_2PI);
}

template <bool BF16>
__global__ void __launch_bounds__(256)
lenreg_frames_kernel(const float* __restrict__ xs, Rows RL, Rows RT, const float* __restrict__ mu,
                     const float* __restrict__ sigma, const float* __restrict__ pos,
                     float* __restrict__ weights, float* __restrict__ xo_f32, __nv_bfloat16* __restrict__ xo_bf,
                     int64_t rows_phys_t) {
    ptx::griddep_wait();      // PDL: predecessor complete + visible
    ptx::griddep_launch();    // let the next kernel start its prologue
    extern __shared__ __align__(16) float lr_smem[];
    const int L = RL.n_max, T = RT.n_max;
    const int Lp = (L + 3) & ~3;      // keep the float4-read tile below 16-byte aligned
    float* s_mu = lr_smem;            // [Lp]
    float* s_sg = s_mu + Lp;          // [Lp]
    float* s_lsg = s_sg + Lp;         // [Lp]
    float* s_part = s_lsg + Lp;       // [8][32]
    float* s_den = s_part + 8 * LR_FT;  // [32]
    float* s_w = s_den + LR_FT;       // [32 l][32 t]
    int* s_any = reinterpret_cast<int*>(s_w + 32 * LR_FT);  // [32]
    int* s_band = s_any + 32;         // [Lp + 32]: 1 = the symbol's Gaussian can be non-zero somewhere in this tile

    const int b = blockIdx.y;
    const int t0 = blockIdx.x * LR_FT;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int len_in = (int)min((long long)RL.len[b], (long long)L);
    const long long len_out = RT.len[b];
    for (int l = tid; l < len_in; l += 256) {
        float sg = sigma[(int64_t)b * L + l];
        s_mu[l] = mu[(int64_t)b * L + l];
        s_sg[l] = sg;
        s_lsg[l] = logf(sg);
        // nearest frame centre of the tile to μ_l; the largest exponent any frame of the tile can see
        const float m = s_mu[l];
        const float dmin = fmaxf(fmaxf((float)t0 + 0.5f - m, m - ((float)t0 + (float)LR_FT - 0.5f)), 0.f);
        s_band[l] = (-(dmin * dmin) / (2.f * sg * sg) - s_lsg[l] - LR_LOG_SQRT_2PI) >= LR_ZERO_ARG;
    }
    for (int l = len_in + tid; l < Lp + 32; l += 256) s_band[l] = 0;
    __syncthreads();
    const float tc = (float)(t0 + lane) + 0.5f;
    // pass 1: denominators
    float part = 0.f;
    for (int l = warp; l < len_in; l += 8)
        if (s_band[l]) part += gauss_prob(tc, s_mu[l], s_sg[l], s_lsg[l]);      // else exactly + 0.f
    s_part[warp * LR_FT + lane] = part;
    __syncthreads();
    if (tid < LR_FT) {
        float d = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) d += s_part[i * LR_FT + tid];
        s_den[tid] = d + 1e-20f;
    }
    __syncthreads();
    const float den = s_den[lane];
    // pass 2
    const int c = tid & 127, th = tid >> 7;   // contraction: channel c, frames th*16 .. th*16+15
    float acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = 0.f;
    const int64_t rsL = RL.row_start[b];
    for (int l0 = 0; l0 < L; l0 += 32) {
        if (__ballot_sync(0xffffffffu, s_band[l0 + lane]) == 0u) {
            // no symbol of this chunk reaches the tile: its weights are exactly 0
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int l = l0 + warp * 4 + i;
                if (l < L && t0 + lane < T) weights[((int64_t)b * L + l) * T + t0 + lane] = 0.f;
            }
            continue;
        }
        // 8 warps x 4 symbols each: weights for (l, t0 + lane)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int ll = warp * 4 + i, l = l0 + ll;
            float wv = 0.f;
            if (l < len_in && s_band[l]) wv = gauss_prob(tc, s_mu[l], s_sg[l], s_lsg[l]) / den;
            s_w[ll * LR_FT + lane] = wv;
            const int any = __any_sync(0xffffffffu, wv != 0.f);
            if (lane == 0) s_any[ll] = any;
            if (l < L && t0 + lane < T) weights[((int64_t)b * L + l) * T + t0 + lane] = wv;
        }
        __syncthreads();
        if (l0 < len_in) {
            const int lim = min(32, len_in - l0);
            for (int ll = 0; ll < lim; ++ll) {
                if (!s_any[ll]) continue;
                const float xv = xs[(rsL + l0 + ll) * DX_D + c];
                const float4* wp = reinterpret_cast<const float4*>(s_w + ll * LR_FT + th * 16);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float4 w4 = wp[q];
                    acc[q * 4 + 0] = fmaf(w4.x, xv, acc[q * 4 + 0]);
                    acc[q * 4 + 1] = fmaf(w4.y, xv, acc[q * 4 + 1]);
                    acc[q * 4 + 2] = fmaf(w4.z, xv, acc[q * 4 + 2]);
                    acc[q * 4 + 3] = fmaf(w4.w, xv, acc[q * 4 + 3]);
                }
            }
        }
        __syncthreads();
    }
    // decoder input rows: valid frames get x_up + P[t]; the DX_GAP gap rows get zeros
    const int64_t rsT = RT.row_start[b];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const int t = t0 + th * 16 + i;
        if (t >= len_out + DX_GAP) continue;
        float v = 0.f;
        if (t < len_out && t < T) v = acc[i] + pos[(int64_t)t * DX_D + c];
        const int64_t r = rsT + t;
        if (!BF16) xo_f32[r * DX_D + c] = v;
        else xo_bf[((int64_t)(c >> 3) * rows_phys_t + DX_HALO + r) * 8 + (c & 7)] = __float2bfloat16(v);
    }
}

int lenreg_frames(const float* xs, const Rows& RL, const Rows& RT, const float* mu, const float* sigma,
                  const float* pos, float* weights, float* xo_f32, __nv_bfloat16* xo_bf, int64_t rows_phys_t,
                  cudaStream_t s) {
    const int L = RL.n_max;
    const size_t smem = (size_t)(3 * ((L + 3) & ~3) + 8 * LR_FT + LR_FT + 32 * LR_FT) * sizeof(float) + (32 + ((L + 3) & ~3) + 32) * sizeof(int);
    DX_REQUIRE(smem <= 200 * 1024, "lenreg_frames: L_max=%d too large", L);
    static SmemConfigured configured_bf, configured_f32;
    if (smem > 48 * 1024) {
        DX_TRY(ensure_dyn_smem(lenreg_frames_kernel<true>, smem, configured_bf));
        DX_TRY(ensure_dyn_smem(lenreg_frames_kernel<false>, smem, configured_f32));
    }
    dim3 grid((unsigned)((RT.n_max + DX_GAP + LR_FT - 1) / LR_FT), (unsigned)RT.B);
    ProfScope prof("lenreg_frames", s);
    if (xo_bf)
        DX_LAUNCH_PDL((lenreg_frames_kernel<true>), grid, 256, smem, s, xs, RL, RT, mu, sigma, pos, weights, nullptr, xo_bf, rows_phys_t);
    else
        DX_LAUNCH_PDL((lenreg_frames_kernel<false>), grid, 256, smem, s, xs, RL, RT, mu, sigma, pos, weights, xo_f32, nullptr, 0);
    DX_CHECK_LAUNCH();
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// int_durations: get_int_durations (model.py:950-973) + duration_to_integer (extract_features.py:69-125) on the
// device, one thread per utterance, IEEE double arithmetic in the reference's (Python) order, no FMA
// contraction (explicit __dadd_rn / __dmul_rn).  `duration` is thresholded in place (model.py:954-957).
// status[b] = 1 where the reference would raise (IndexError / shape mismatch / ValueError).
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ long long frames_upto(long long x, long long half, long long hop, long long nb) {
    // number of k in [0, nb) with half + hop * k <= x
    long long num = x - half;
    if (num < 0) return 0;
    long long k = num / hop + 1;
    return k > nb ? nb : k;
}

__global__ void int_durations_kernel(float* __restrict__ duration, int B, int L, int sr, int flt, int hop,
                                     int centered, int64_t* __restrict__ dint, int32_t* __restrict__ status) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    float* dur = duration + (int64_t)b * L;
    int64_t* out = dint + (int64_t)b * L;
    // dur_min = (filter_length / sampling_rate) / 2 in double; the comparison happens in fp32 (torch casts the
    // Python scalar to the tensor's dtype)
    const float dur_min = (float)(((double)flt / (double)sr) / 2.0);
    double end_prev = 0.0, total = 0.0;
    int nnz = 0;
    for (int l = 0; l < L; ++l) {
        float d = dur[l];
        if (d < dur_min) { d = 0.f; dur[l] = 0.f; }
        out[l] = 0;
        if (d != 0.f) {
            const double bg = end_prev;
            const double en = __dadd_rn(end_prev, (double)d);
            end_prev = en;                                   // end_prev += symb_dur (same sum)
            total = __dadd_rn(total, __dsub_rn(en, bg));      // sum([(x[1] - x[0]) ...]) left to right from 0
            ++nnz;
        }
    }
    const long long nb_samples = (long long)__dmul_rn(total, (double)sr);           // int(): trunc toward zero
    const long long nb_frames = 1 + (long long)((double)(nb_samples - flt) / (double)hop);
    const long long half = flt / 2;
    int st = 0;
    long long curr = 1;
    int consumed = 0, first_l = -1, last_l = -1, l = 0;
    end_prev = 0.0;
    while (curr <= nb_frames) {
        // pop next non-zero symbol
        while (l < L && dur[l] == 0.f) ++l;
        if (l >= L) { st = 1; break; }                       // IndexError: pop from empty list
        const double bg = end_prev;
        const double en = __dadd_rn(end_prev, (double)dur[l]);
        end_prev = en;
        if (bg == en) { st = 1; break; }                     // ValueError
        const long long bs = (long long)__dmul_rn(bg, (double)sr);
        const long long es = (long long)__dmul_rn(en, (double)sr);
        long long n = frames_upto(es, half, hop, nb_frames) - frames_upto(bs, half, hop, nb_frames);
        if (n < 0) n = 0;
        out[l] = n;
        if (first_l < 0) first_l = l;
        last_l = l;
        curr += n;
        ++consumed;
        ++l;
    }
    if (st == 0) {
        if (consumed == 0) st = 1;                           // int_durations[0] on an empty list
        else {
            long long left, right;
            if (centered) { left = (flt / 2) / hop; right = left; }
            else { const long long extra = (flt - hop) / hop; left = extra / 2; right = extra - left; }
            out[first_l] += left;
            const int remaining = nnz - consumed;
            if (remaining != 0) {
                while (l < L && dur[l] == 0.f) ++l;          // the next non-zero symbol takes the edge frames
                out[l] = right;
                ++consumed;
            } else {
                out[last_l] += right;
            }
            if (consumed != nnz) st = 1;                     // index-assign shape mismatch (model.py:969)
        }
    }
    status[b] = st;
}

int int_durations(float* duration, int B, int L, int sr, int flt, int hop, int centered, int64_t* dint,
                  int32_t* status, cudaStream_t s) {
    int_durations_kernel<<<(B + 63) / 64, 64, 0, s>>>(duration, B, L, sr, flt, hop, centered, dint, status);
    DX_CHECK_LAUNCH();
    return 0;
}

}  // namespace dx
