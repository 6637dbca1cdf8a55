// HiFi-GAN generator for a packed batch of utterances.
//
// Reference ops replaced: HiFiGANGenerator.forward (vocoder/hifigan.py:159-176), ResBlock1.forward (hifigan.py:90-97),
// weight-norm-free Conv1d / ConvTranspose1d (after remove_weight_norm, hifigan.py:178-184), and the one-utterance-at-
// a-time loop of HiFiGanVocoder.infer (hifigan.py:211-226): every utterance is vocoded as if alone on its cropped mel.
//
// Data layout.  Utterance b owns frames [frame_start[b], frame_start[b] + len[b]) of one packed frame stream;
// DX_VC_GAP zero frames sit in front of, between and behind utterances.  At rate u (1, 8, 64, 128, 256) activation
// row r is sample (r mod u) of frame r / u, so frame_utt[r >> log2(u)] >= 0 is the row's validity.  Every epilogue
// writes zeros at invalid rows of the tensor the next convolution reads, which IS the reference's zero padding at the
// utterance ends (DX_VC_GAP * 8 >= 25 = the widest half receptive field, k = 11 with dilation 5).
//   bf16 mode: conv inputs are bf16 channel-interleaved [C/8][DX_VC_HALO + rows + DX_VC_HALO][8] (one 128-row tile +
//              halo of an 8-channel chunk = one bulk copy that lands as the no-swizzle K-major UMMA layout; a dilated
//              tap is a +dil*16 B shift of the descriptor start address); residual stream x fp32 [rows][C].
//   fp32 mode: everything fp32 [rows][C].
// Every convolution is ONE launch with its pointwise neighbours fused into the epilogue:
//   y = acc + bias (+ residual); xs bookkeeping of the three parallel ResBlocks (init / add / mean);
//   x_out = y (fp32 residual stream); act_out = mask(leaky_relu(y)) = the operand of the next convolution.
// A transposed convolution with kernel = 2 * stride is its 3-tap polyphase form at the INPUT rate with
// N = stride * C_out: column n = phase * C_out + channel lands in output row r * stride + phase.
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_ptx.cuh"
#include "dx_vocoder.cuh"

namespace dx {

// ---- frame map ----------------------------------------------------------------------------------------------
// frame_start[b] = GAP + sum_{b' < b} (len[b'] + GAP); frame_start[B] = end of the stream incl. the trailing gap
// (clamped to frame_cap: utterances that do not fit become invalid); frame_utt[f] = b or -1.
constexpr int VC_MAX_B = 4096;
__global__ void vc_build_frames_kernel(const int64_t* __restrict__ lengths, int B, int T_max, int64_t frame_cap,
                                       int32_t* __restrict__ frame_start, int32_t* __restrict__ frame_utt) {
    __shared__ int32_t s_start[VC_MAX_B + 1];
    __shared__ int32_t s_len[VC_MAX_B];
    if (threadIdx.x == 0) {
        int64_t f = DX_VC_GAP;
        for (int b = 0; b < B; ++b) {
            int64_t n = lengths[b];
            if (n < 0) n = 0;
            if (n > T_max) n = T_max;
            if (f + n + DX_VC_GAP > frame_cap) n = 0;
            s_start[b] = (int32_t)f;
            s_len[b] = (int32_t)n;
            if (f + n + DX_VC_GAP <= frame_cap) f += n + DX_VC_GAP;
        }
        s_start[B] = (int32_t)f;
    }
    __syncthreads();
    if (blockIdx.x == 0)
        for (int b = threadIdx.x; b <= B; b += blockDim.x) frame_start[b] = s_start[b];
    for (int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; f < frame_cap; f += (int64_t)gridDim.x * blockDim.x) {
        int lo = 0, hi = B - 1;                        // largest b with s_start[b] <= f
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (s_start[mid] <= f) lo = mid; else hi = mid - 1;
        }
        const bool in = f >= s_start[lo] && f < s_start[lo] + s_len[lo];
        frame_utt[f] = in ? lo : -1;
    }
}

// mel (B, C, T_max) fp32 → packed frame rows: bf16 interleaved [C/8][HALO + cap + HALO][8] and / or fp32 [cap][C]
__global__ void vc_pack_mel_kernel(const float* __restrict__ mel, int C, int T_max, int64_t frame_cap,
                                   const int32_t* __restrict__ frame_start, const int32_t* __restrict__ frame_utt,
                                   __nv_bfloat16* __restrict__ out_bf, int64_t rows_phys, float* __restrict__ out_f32) {
    const int chunks = C / 8;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= frame_cap * chunks) return;
    const int64_t f = i % frame_cap;                   // consecutive threads ↔ consecutive frames (coalesced in t)
    const int c = (int)(i / frame_cap);
    const int b = frame_utt[f];
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = 0.f;
    if (b >= 0) {
        const int t = (int)(f - frame_start[b]);
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = mel[((int64_t)b * C + c * 8 + e) * T_max + t];
    }
    if (out_bf) {
        __align__(16) __nv_bfloat162 o2[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) o2[q] = __floats2bfloat162_rn(v[2 * q], v[2 * q + 1]);
        *reinterpret_cast<uint4*>(out_bf + ((int64_t)c * rows_phys + DX_VC_HALO + f) * 8) = *reinterpret_cast<const uint4*>(o2);
    }
    if (out_f32) {
#pragma unroll
        for (int e = 0; e < 8; ++e) out_f32[f * C + c * 8 + e] = v[e];
    }
}

// conv_post (C → 1, hifigan.py:152) + tanh (hifigan.py:175), written straight into the padded (B, T_max * hop) output.
// The input already carries the leaky_relu in front of it (hifigan.py:173).  HBM-bound: 2 C bytes read per sample.
template <bool BF>
__global__ void vc_post_kernel(const void* __restrict__ act, int64_t rows_phys, int C, int taps, const float* __restrict__ w,
                               const float* __restrict__ bias, const int32_t* __restrict__ frame_start,
                               int hop_shift, int T_max, int B, float* __restrict__ audio) {
    extern __shared__ float s_w[];                      // [taps][C]
    for (int i = threadIdx.x; i < taps * C; i += blockDim.x) s_w[i] = w[i];
    ptx::griddep_launch();
    ptx::griddep_wait();
    __syncthreads();
    const int64_t per = (int64_t)T_max << hop_shift;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= per * B) return;
    const int b = (int)(i / per);
    const int64_t j = i - (int64_t)b * per;
    const int64_t n = ((int64_t)(frame_start[b + 1] - frame_start[b]) - DX_VC_GAP) << hop_shift;   // valid samples
    float y = 0.f;
    if (j < n) {
        const int64_t r = ((int64_t)frame_start[b] << hop_shift) + j;
        y = bias[0];
        for (int tap = 0; tap < taps; ++tap) {
            const int64_t rr = r + tap - taps / 2;
            if constexpr (BF) {
                const __nv_bfloat16* a = reinterpret_cast<const __nv_bfloat16*>(act);
                for (int c = 0; c < C / 8; ++c) {
                    const uint4 pk = *reinterpret_cast<const uint4*>(a + ((int64_t)c * rows_phys + DX_VC_HALO + rr) * 8);
                    const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&pk);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        y = fmaf(__low2float(h2[q]), s_w[tap * C + c * 8 + 2 * q], y);
                        y = fmaf(__high2float(h2[q]), s_w[tap * C + c * 8 + 2 * q + 1], y);
                    }
                }
            } else {
                const float* a = reinterpret_cast<const float*>(act) + rr * C;
                for (int c = 0; c < C; ++c) y = fmaf(a[c], s_w[tap * C + c], y);
            }
        }
        y = tanhf(y);
    }
    audio[i] = y;
}


// ---------------------------------------------------------------------------------------------------------
// plan
// ---------------------------------------------------------------------------------------------------------
static int ilog2(int v) {
    int s = 0;
    while ((1 << s) < v) ++s;
    return s;
}

struct VcLayout {
    int32_t *frame_utt, *frame_start;
    void *melp, *next, *a0, *ar, *h;       // conv operands (bf16 interleaved / fp32 rows)
    float *x0, *xr, *xs;                   // fp32 residual stream
    int64_t bytes;
};

// elements (not bytes) of the widest activation per frame over the stages, and the widest channel count
static void vc_extents(const DxVocoderWeights* w, int64_t* per_frame, int* c_max) {
    int64_t pf = w->pre.N;
    int cm = w->pre.N > w->mel_channels ? w->pre.N : w->mel_channels;
    int rate = 1;
    for (int s = 0; s < w->n_stages; ++s) {
        rate *= w->up[s];
        const int64_t e = (int64_t)rate * w->channels[s];
        if (e > pf) pf = e;
        if (w->channels[s] > cm) cm = w->channels[s];
    }
    *per_frame = pf;
    *c_max = cm;
}

static int vc_carve(const DxVocoderWeights* w, int B, int64_t frame_cap, int mode, void* base, VcLayout* L) {
    int64_t pf;
    int cm;
    vc_extents(w, &pf, &cm);
    int64_t off = 0;
    auto take = [&](int64_t bytes) {
        void* p = base ? (void*)((uint8_t*)base + off) : nullptr;
        off += round_up(bytes, 256);
        return p;
    };
    L->frame_utt = (int32_t*)take(frame_cap * 4);
    L->frame_start = (int32_t*)take((int64_t)(B + 1) * 4);
    // an operand buffer holds C/8 chunks of (rows + 2 * HALO) rows of 16 bytes; C * rows <= pf * frame_cap
    const int64_t opnd = mode == DX_MODE_BF16 ? (pf * frame_cap + 2ll * DX_VC_HALO * cm) * 2 : pf * frame_cap * 4;
    const int64_t melb = mode == DX_MODE_BF16 ? ((int64_t)w->mel_channels * (frame_cap + 2 * DX_VC_HALO)) * 2
                                              : (int64_t)w->mel_channels * frame_cap * 4;
    L->melp = take(melb);
    L->next = take(opnd);
    L->a0 = take(opnd);
    L->ar = take(opnd);
    L->h = take(opnd);
    L->x0 = (float*)take(pf * frame_cap * 4);
    L->xr = (float*)take(pf * frame_cap * 4);
    L->xs = (float*)take(pf * frame_cap * 4);
    L->bytes = off;
    return 0;
}

static int vc_check_weights(const DxVocoderWeights* w) {
    DX_REQUIRE(w != nullptr, "vocoder: weights == NULL");
    DX_REQUIRE(w->n_stages >= 1 && w->n_stages <= DX_VC_MAX_STAGES && w->n_kernels >= 1 && w->n_kernels <= DX_VC_MAX_KERNELS &&
                   w->n_layers >= 1 && w->n_layers <= DX_VC_MAX_LAYERS,
               "vocoder: %d stages x %d kernels x %d layers not supported", w->n_stages, w->n_kernels, w->n_layers);
    DX_REQUIRE(w->mel_channels % 16 == 0, "vocoder: mel_channels=%d must be a multiple of 16", w->mel_channels);
    for (int s = 0; s < w->n_stages; ++s) {
        DX_REQUIRE((w->up[s] & (w->up[s] - 1)) == 0 && w->up[s] >= 2, "vocoder: upsample rate %d is not a power of two", w->up[s]);
        DX_REQUIRE(w->channels[s] % 32 == 0, "vocoder: %d channels at stage %d (multiple of 32)", w->channels[s], s);
    }
    return 0;
}

}  // namespace dx

using namespace dx;

extern "C" {

int64_t dx_vocoder_frame_cap(int B, int64_t total_frames) {
    if (B < 1 || total_frames < 0) return -1;
    return round_up(total_frames + (int64_t)(B + 1) * DX_VC_GAP, 128);
}

int dx_vocoder_conv_tiling(int taps, int dil, int K, int N, int* nt, int* ks) {
    DX_REQUIRE(nt && ks && taps >= 1 && dil >= 1 && K >= 16 && N >= 32, "dx_vocoder_conv_tiling: bad arguments");
    vc_conv_tiling(taps, dil, K, N, nt, ks);
    return 0;
}

int64_t dx_vocoder_workspace_bytes(const DxVocoderWeights* w, int B, int64_t frame_cap, int mode) {
    if (vc_check_weights(w) != 0) return -1;
    if (B < 1 || B > VC_MAX_B || frame_cap < 128 || frame_cap % 128 != 0 || (mode != DX_MODE_F32 && mode != DX_MODE_BF16)) {
        set_error("dx_vocoder_workspace_bytes: B=%d frame_cap=%lld mode=%d", B, (long long)frame_cap, mode);
        return -1;
    }
    VcLayout L;
    vc_carve(w, B, frame_cap, mode, nullptr, &L);
    return L.bytes;
}

int dx_vocoder_forward(const DxVocoderWeights* w, const DxVocoderArgs* a, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    DX_TRY(vc_check_weights(w));
    DX_REQUIRE(a != nullptr && a->mel && a->lengths && a->audio && a->workspace, "dx_vocoder_forward: NULL argument");
    DX_REQUIRE(a->B >= 1 && a->B <= VC_MAX_B && a->T_max >= 1, "dx_vocoder_forward: B=%d T_max=%d", a->B, a->T_max);
    DX_REQUIRE(a->mode == DX_MODE_F32 || a->mode == DX_MODE_BF16, "dx_vocoder_forward: mode=%d", a->mode);
    DX_REQUIRE(a->frame_cap >= 128 && a->frame_cap % 128 == 0, "dx_vocoder_forward: frame_cap=%lld (multiple of 128)", (long long)a->frame_cap);
    const bool bf = a->mode == DX_MODE_BF16;
    const int x_bf16 = bf && !(a->flags & DX_VC_RESIDUAL_F32) ? 1 : 0;     // residual stream x: bf16 [C/8][rows][8] or fp32 [C/4][rows][4]
    const int64_t cap = a->frame_cap;
    int hop = 1;
    for (int i = 0; i < w->n_stages; ++i) hop *= w->up[i];
    DX_REQUIRE(cap < (1ll << 31) / hop * 128, "dx_vocoder_forward: frame_cap=%lld too large (row tiles are counted in int32)", (long long)cap);
    VcLayout L;
    vc_carve(w, a->B, cap, a->mode, a->workspace, &L);
    DX_REQUIRE(L.bytes <= a->workspace_bytes, "dx_vocoder_forward: workspace %lld < %lld bytes", (long long)a->workspace_bytes, (long long)L.bytes);

    {
        ProfScope prof("vc_frames", s);
        int grid = (int)((cap + 1023) / 1024);
        if (grid > 148) grid = 148;
        vc_build_frames_kernel<<<grid, 1024, 0, s>>>(a->lengths, a->B, a->T_max, cap, L.frame_start, L.frame_utt);
        DX_CHECK_LAUNCH();
        const int64_t n = cap * (w->mel_channels / 8);
        vc_pack_mel_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(a->mel, w->mel_channels, a->T_max, cap, L.frame_start, L.frame_utt,
                                                                      bf ? (__nv_bfloat16*)L.melp : nullptr, cap + 2 * DX_VC_HALO,
                                                                      bf ? nullptr : (float*)L.melp);
        DX_CHECK_LAUNCH();
    }
    // one convolution launch
    auto conv = [&](const DxVcConv& cw, const void* in, int in_rate, int c_out, int up, const float* res, float* x_out,
                    int xs_mode, void* act_out, const char* tag) -> int {
        VcConvArgs c;
        c.a = in;
        c.a_rows_phys = cap * in_rate + 2 * DX_VC_HALO;
        c.K = cw.K; c.N = cw.N; c.taps = cw.taps; c.dil = cw.dil;
        c.w_f32 = cw.w;
        c.w_bf = (const __nv_bfloat16*)cw.w_bf;
        c.total_frames = L.frame_start + a->B;
        c.in_shift = ilog2(in_rate);
        c.rows_bound = cap * in_rate;
        c.e.bias = cw.bias; c.e.c_out = c_out; c.e.up = up;
        c.e.frame_utt = L.frame_utt;
        c.e.out_shift = ilog2(in_rate * up);
        c.e.res = res; c.e.x_out = x_out; c.e.xs = L.xs; c.e.xs_mode = xs_mode;
        c.e.x_rows = cap * in_rate * up;
        c.e.x_bf16 = x_bf16;
        c.e.n_kernels = (float)w->n_kernels;
        c.e.act = 1;
        c.e.act_out = act_out;
        c.e.act_rows_phys = cap * in_rate * up + 2 * DX_VC_HALO;
        c.tag = tag;
        DX_REQUIRE(cw.w != nullptr && cw.bias != nullptr, "vocoder: weights of %s missing", tag);
        return bf ? vc_conv_bf16(c, s) : vc_conv_f32(c, s);
    };
    static const char* const kStageTags[4][3] = {{"vc_up0", "vc_s0_conv1", "vc_s0_conv2"}, {"vc_up1", "vc_s1_conv1", "vc_s1_conv2"},
                                                {"vc_up2", "vc_s2_conv1", "vc_s2_conv2"}, {"vc_up3", "vc_s3_conv1", "vc_s3_conv2"}};
    static const char* const kPairTags[4] = {"vc_s0_pair", "vc_s1_pair", "vc_s2_pair", "vc_s3_pair"};
    const bool fuse_pairs = bf && !debug_knob(12);      // knob 12: every ResBlock layer as two launches (A/B, bit-identical)
    // conv_pre (hifigan.py:160); its only consumer is leaky_relu → ups[0] (hifigan.py:162-163)
    DX_TRY(conv(w->pre, L.melp, 1, w->pre.N, 1, nullptr, nullptr, 0, L.next, "vc_pre"));
    int rate = 1;
    for (int st = 0; st < w->n_stages; ++st) {
        const int C = w->channels[st];
        // x = ups[i](leaky_relu(x)) (hifigan.py:162-163): x0 fp32 + a0 = leaky_relu(x0) for the three ResBlocks
        DX_TRY(conv(w->ups[st], L.next, rate, C, w->up[st], nullptr, L.x0, 0, L.a0, kStageTags[st][0]));
        rate *= w->up[st];
        for (int j = 0; j < w->n_kernels; ++j) {
            for (int l = 0; l < w->n_layers; ++l) {
                const bool last = l == w->n_layers - 1;
                const float* res = l == 0 ? L.x0 : L.xr;
                const bool fin = last && j == w->n_kernels - 1;
                const int xs_mode = !last ? 0 : (fin ? (w->n_kernels == 1 ? 0 : 3) : (j == 0 ? 1 : 2));
                float* x_out = last ? nullptr : L.xr;
                void* act_out = last ? (fin ? L.next : nullptr) : L.ar;
                const void* in = l == 0 ? L.a0 : L.ar;
                if (fuse_pairs && C <= (debug_knob(13) ? 128 : 64)) {
                    // one launch: x = convs2[l](leaky_relu(convs1[l](leaky_relu(x)))) + x, hidden tile on chip.  The
                    // operand it reads (with halo rows of neighbouring tiles) and the one it writes must differ:
                    // ar and h (free in this mode) alternate
                    void* pp[2] = {L.ar, L.h};
                    const void* pin = l == 0 ? L.a0 : pp[(l - 1) & 1];
                    void* pout = last ? act_out : pp[l & 1];
                    const DxVcConv& c1 = w->rb[st][j][l][0];
                    const DxVcConv& c2 = w->rb[st][j][l][1];
                    DX_REQUIRE(c1.w_bf && c2.w_bf && c1.bias && c2.bias && c1.taps == c2.taps && c2.dil == 1 && c1.K == C && c1.N == C &&
                                   c2.K == C && c2.N == C, "vocoder: ResBlock layer %d/%d/%d is not a fusable pair", st, j, l);
                    VcPairArgs pa;
                    pa.a = (const __nv_bfloat16*)pin;
                    pa.a_rows_phys = cap * rate + 2 * DX_VC_HALO;
                    pa.C = C; pa.taps = c1.taps; pa.dil = c1.dil;
                    pa.w1_bf = (const __nv_bfloat16*)c1.w_bf; pa.w2_bf = (const __nv_bfloat16*)c2.w_bf;
                    pa.bias1 = c1.bias; pa.bias2 = c2.bias;
                    pa.total_frames = L.frame_start + a->B;
                    pa.shift = ilog2(rate);
                    pa.rows_bound = cap * rate;
                    pa.e.bias = c2.bias; pa.e.c_out = C; pa.e.up = 1;
                    pa.e.frame_utt = L.frame_utt; pa.e.out_shift = pa.shift;
                    pa.e.res = res; pa.e.x_out = x_out; pa.e.xs = L.xs; pa.e.xs_mode = xs_mode;
                    pa.e.x_rows = cap * rate;
                    pa.e.x_bf16 = x_bf16;
                    pa.e.n_kernels = (float)w->n_kernels;
                    pa.e.act = 1; pa.e.act_out = pout; pa.e.act_rows_phys = cap * rate + 2 * DX_VC_HALO;
                    pa.tag = kPairTags[st];
                    DX_TRY(vc_pair_bf16(pa, s));
                    continue;
                }
                // xt = leaky_relu(convs1[l](leaky_relu(x)))  (hifigan.py:92-94)
                DX_TRY(conv(w->rb[st][j][l][0], in, rate, C, 1, nullptr, nullptr, 0, L.h, kStageTags[st][1]));
                // x = convs2[l](xt) + x (hifigan.py:95-96); the last layer feeds xs (hifigan.py:167-171)
                DX_TRY(conv(w->rb[st][j][l][1], L.h, rate, C, 1, res, x_out, xs_mode, act_out, kStageTags[st][2]));
            }
        }
    }
    // leaky_relu → conv_post → tanh (hifigan.py:173-175)
    {
        DX_REQUIRE(w->post.w && w->post.bias && w->post.K == w->channels[w->n_stages - 1], "vocoder: conv_post weights");
        const int64_t n = (int64_t)a->B * a->T_max * hop;
        ProfScope prof("vc_post", s);
        const size_t smem = (size_t)w->post.taps * w->post.K * sizeof(float);
        if (bf) {
            DX_LAUNCH_PDL(vc_post_kernel<true>, dim3((unsigned)((n + 255) / 256)), dim3(256), smem, s, (const void*)L.next,
                          cap * hop + 2 * DX_VC_HALO, w->post.K, w->post.taps, w->post.w, w->post.bias, (const int32_t*)L.frame_start,
                          ilog2(hop), a->T_max, a->B, a->audio);
        } else {
            DX_LAUNCH_PDL(vc_post_kernel<false>, dim3((unsigned)((n + 255) / 256)), dim3(256), smem, s, (const void*)L.next,
                          (int64_t)0, w->post.K, w->post.taps, w->post.w, w->post.bias, (const int32_t*)L.frame_start,
                          ilog2(hop), a->T_max, a->B, a->audio);
        }
        DX_CHECK_LAUNCH();
    }
    return 0;
}

}  // extern "C"
