// Conditioning heads: speaker projection, FiLM style adapter, speaker classifier (all tiny: one CTA per
// utterance, warp-per-output dot products with coalesced weight reads), and the inference-side prosody
// transforms (pitch shift / multiply, zero-duration masking).
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_ptx.cuh"

namespace dx {

__device__ __forceinline__ float dot_row_128(const float* __restrict__ wrow, const float* __restrict__ x_smem,
                                             int lane) {
    float4 w = reinterpret_cast<const float4*>(wrow)[lane];
    float4 x = reinterpret_cast<const float4*>(x_smem)[lane];
    return warp_sum((w.x * x.x + w.y * x.y) + (w.z * x.z + w.w * x.w));
}

// Reference: model.py:904-905 (F.normalize p=2 eps 1e-12 + spk_projection), :919 (accent + spk),
// StyleAdapter.forward :764-806, SpeakerClassifier.forward :829-830 (GRL is the identity in forward).
// Grid (B, n_film + 1): blockIdx.y = 0..n_film-1 → FiLM block (128 gammas + 128 betas; n_film = phoneme-encoder + decoder
// blocks, 8 by default), n_film → speaker classifier.  Every CTA
// recomputes the 128-wide conditioning vector (24 K MACs) rather than synchronising across CTAs.
__global__ void __launch_bounds__(256)
style_kernel(const float* __restrict__ spk_embs, const float* __restrict__ spk_emb_in,
             const float* __restrict__ accent, DxModelWeights w, float* __restrict__ film,
             float* __restrict__ speaker_preds) {
    ptx::griddep_wait();      // PDL: predecessor complete + visible
    ptx::griddep_launch();    // let the next kernel start its prologue
    __shared__ __align__(16) float s_e[DX_SPK];
    __shared__ __align__(16) float s_c[DX_D];     // accent + spk
    __shared__ __align__(16) float s_a[DX_D];     // accent
    __shared__ __align__(16) float s_h[DX_D];
    __shared__ __align__(16) float s_h2[DX_D];
    __shared__ float s_red[8];
    const int b = blockIdx.x, job = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_film = (int)gridDim.y - 1;

    if (tid < DX_D) s_a[tid] = accent[b * DX_D + tid];
    if (job == n_film) {
        if (speaker_preds == nullptr) return;
        __syncthreads();
        for (int n = warp; n < DX_D; n += 8) {
            float v = dot_row_128(w.cls_w1 + n * DX_D, s_a, lane);
            if (lane == 0) s_h[n] = fmaxf(v + w.cls_b1[n], 0.f);
        }
        __syncthreads();
        for (int n = warp; n < DX_D; n += 8) {
            float v = dot_row_128(w.cls_w2 + n * DX_D, s_h, lane);
            if (lane == 0) s_h2[n] = fmaxf(v + w.cls_b2[n], 0.f);
        }
        __syncthreads();
        for (int n = warp; n < w.n_speakers; n += 8) {
            float v = dot_row_128(w.cls_w3 + n * DX_D, s_h2, lane);
            if (lane == 0) speaker_preds[b * w.n_speakers + n] = v + w.cls_b3[n];
        }
        return;
    }
    if (spk_emb_in != nullptr) {
        if (tid < DX_D) s_c[tid] = spk_emb_in[b * DX_D + tid];
        __syncthreads();
    } else {
        float v = (tid < DX_SPK) ? spk_embs[b * DX_SPK + tid] : 0.f;
        float ss = warp_sum(v * v);
        if (lane == 0) s_red[warp] = ss;
        __syncthreads();
        float tot = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) tot += s_red[i];
        const float nrm = fmaxf(sqrtf(tot), 1e-12f);
        if (tid < DX_SPK) s_e[tid] = v / nrm;
        __syncthreads();
        // spk projection: 128 outputs, warp per output, 192-long rows (6 coalesced loads per lane)
        for (int n = warp; n < DX_D; n += 8) {
            const float* wr = w.spk_w + n * DX_SPK;
            float acc = 0.f;
#pragma unroll
            for (int k = 0; k < DX_SPK / 32; ++k) acc += wr[lane + 32 * k] * s_e[lane + 32 * k];
            acc = warp_sum(acc);
            if (lane == 0) s_c[n] = acc + w.spk_b[n];
        }
        __syncthreads();
    }
    if (tid < DX_D) s_c[tid] = s_a[tid] + s_c[tid];
    __syncthreads();
    // FiLM block `job`: outputs n = 128*job + i.  Each warp owns 16 outputs, 4 at a time for load-level parallelism.
    const float pg = w.post_mult ? w.post_mult[job] : 1.f;
    const float pb = w.post_mult ? w.post_mult[n_film + job] : 1.f;
    const float4 x = reinterpret_cast<const float4*>(s_c)[lane];
    for (int i0 = warp * 16; i0 < warp * 16 + 16; i0 += 4) {
        float g[4], q[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int n = job * DX_D + i0 + u;
            const float4 wg = reinterpret_cast<const float4*>(w.film_g_w + (int64_t)n * DX_D)[lane];
            const float4 wb = reinterpret_cast<const float4*>(w.film_b_w + (int64_t)n * DX_D)[lane];
            g[u] = (wg.x * x.x + wg.y * x.y) + (wg.z * x.z + wg.w * x.w);
            q[u] = (wb.x * x.x + wb.y * x.y) + (wb.z * x.z + wb.w * x.w);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            g[u] = warp_sum(g[u]);
            q[u] = warp_sum(q[u]);
        }
        if (lane < 4) {
            const int i = i0 + lane, n = job * DX_D + i;
            const float gv = (lane == 0 ? g[0] : lane == 1 ? g[1] : lane == 2 ? g[2] : g[3]) + w.film_g_b[n];
            const float qv = (lane == 0 ? q[0] : lane == 1 ? q[1] : lane == 2 ? q[2] : q[3]) + w.film_b_b[n];
            film[((int64_t)b * 8 + job) * 256 + i] = pg * gv + 1.f;
            film[((int64_t)b * 8 + job) * 256 + 128 + i] = pb * qv;
        }
    }
}

int style(const DxModelWeights* w, int B, const float* spk_embs, const float* spk_emb_in, const float* accent,
          float* film, float* speaker_preds, cudaStream_t s) {
    DX_REQUIRE(spk_embs != nullptr || spk_emb_in != nullptr, "style: speaker embeddings required");
    ProfScope prof("style", s);
    const int n_film = n_blocks_of(w->n_pe_blocks) + n_blocks_of(w->n_fd_blocks);
    DX_LAUNCH_PDL((style_kernel), dim3(B, n_film + 1), 256, 0, s, spk_embs, spk_emb_in, accent, *w, film, speaker_preds);
    DX_CHECK_LAUNCH();
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// prosody_transform: DaftExprt.inference model.py:1077-1087 with pitch_shift :975-994 / pitch_multiply
// :996-1024.  One CTA per utterance; all fp32 like the reference's tensor ops.
// ---------------------------------------------------------------------------------------------------------
__global__ void prosody_transform_kernel(float* __restrict__ energy, float* __restrict__ pitch,
                                         const float* __restrict__ energy_factors,
                                         const float* __restrict__ pitch_factors,
                                         const int64_t* __restrict__ dint, const float* __restrict__ pmean,
                                         const float* __restrict__ pstd, int L, int transform) {
    __shared__ float s_sum[32];
    __shared__ float s_cnt[32];
    const int b = blockIdx.x;
    float* e = energy + (int64_t)b * L;
    float* p = pitch + (int64_t)b * L;
    const int64_t* d = dint + (int64_t)b * L;
    const float* ef = energy_factors + (int64_t)b * L;
    const float* pf = pitch_factors + (int64_t)b * L;
    // masking first (model.py:1077-1080)
    for (int l = threadIdx.x; l < L; l += blockDim.x) {
        float ev = e[l] * ef[l];
        float pv = p[l];
        if (d[l] == 0) { ev = 0.f; pv = 0.f; }
        e[l] = ev;
        p[l] = pv;
    }
    __syncthreads();
    if (transform == 0) {
        const float mean = pmean[b], std = pstd[b];
        for (int l = threadIdx.x; l < L; l += blockDim.x) {
            float pv = p[l];
            if (pv != 0.f) {
                float hz = expf(std * pv + mean);
                hz += pf[l];
                pv = (logf(hz) - mean) / std;
            }
            p[l] = pv;
        }
    } else {
        float sm = 0.f, cnt = 0.f;
        for (int l = threadIdx.x; l < L; l += blockDim.x) {
            float pv = p[l];
            if (pv != 0.f) { sm += pv; cnt += 1.f; }
        }
        sm = warp_sum(sm);
        cnt = warp_sum(cnt);
        if ((threadIdx.x & 31) == 0) { s_sum[threadIdx.x >> 5] = sm; s_cnt[threadIdx.x >> 5] = cnt; }
        __syncthreads();
        float tot = 0.f, n = 0.f;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) { tot += s_sum[i]; n += s_cnt[i]; }
        const float mean = tot / n;   // NaN when no voiced symbol, as torch.mean of an empty tensor
        for (int l = threadIdx.x; l < L; l += blockDim.x) {
            float pv = p[l];
            if (pv != 0.f) {
                float dev = (pv - mean) * pf[l];
                pv = pv + dev;
            }
            p[l] = pv;
        }
    }
}

int prosody_transform(float* energy, float* pitch, const float* energy_factors, const float* pitch_factors,
                      const int64_t* dint, const float* pmean, const float* pstd, int B, int L, int transform,
                      cudaStream_t s) {
    DX_REQUIRE(transform == 0 || transform == 1, "prosody_transform: unknown transform %d", transform);
    prosody_transform_kernel<<<B, 256, 0, s>>>(energy, pitch, energy_factors, pitch_factors, dint, pmean, pstd, L,
                                               transform);
    DX_CHECK_LAUNCH();
    return 0;
}

}  // namespace dx
