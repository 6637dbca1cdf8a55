// Whole-pass planner: workspace carving and the launch sequences of DaftExprt.forward / .inference
// (model.py:889-948, 1026-1114) in both arithmetic modes.  Everything is enqueued on the caller's stream; there is
// no host synchronisation (shapes come from the caller's tensor shapes, lengths stay on the device).
#include <string.h>
#include "dx_plan.cuh"

namespace dx {

namespace {
struct Carver {
    uint8_t* base;
    int64_t off;
    template <typename T>
    T* take(int64_t n) {
        off = round_up(off, 256);
        T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
        off += n * (int64_t)sizeof(T);
        return p;
    }
};
}  // namespace

int plan_layout(Plan* p, int B, int L_max, int T_max, int mode, void* workspace, int64_t workspace_bytes) {
    DX_REQUIRE(B >= 1 && L_max >= 1 && T_max >= 1, "plan: bad shape B=%d L_max=%d T_max=%d", B, L_max, T_max);
    DX_REQUIRE(mode == DX_MODE_F32 || mode == DX_MODE_BF16, "plan: unknown mode %d", mode);
    DX_REQUIRE(T_max <= 5000 && L_max <= 5000, "plan: sequence longer than the 5000-row positional table (model.py:123)");
    memset(p, 0, sizeof(*p));
    p->B = B; p->L_max = L_max; p->T_max = T_max; p->mode = mode;
    p->cap_t = row_cap(B, T_max);
    p->cap_l = row_cap(B, L_max);
    p->cap = p->cap_t > p->cap_l ? p->cap_t : p->cap_l;
    p->rows_phys = p->cap + DX_HALO + DX_TAIL;
    Carver c{reinterpret_cast<uint8_t*>(workspace), 0};
    p->t_row_start = c.take<int32_t>(B + 1);
    p->t_row_b = c.take<int32_t>(p->cap_t);
    p->t_row_p = c.take<int32_t>(p->cap_t);
    p->l_row_start = c.take<int32_t>(B + 1);
    p->l_row_b = c.take<int32_t>(p->cap_l);
    p->l_row_p = c.take<int32_t>(p->cap_l);
    p->accent = c.take<float>((int64_t)B * DX_D);
    p->sigma = c.take<float>((int64_t)B * L_max);
    p->mu = c.take<float>((int64_t)B * L_max);
    p->out_len = c.take<int64_t>(B);
    p->xs = c.take<float>(p->cap_l * DX_D);
    p->melo = c.take<float>(p->cap_t * DX_MEL);
    if (mode == DX_MODE_F32) {
        p->x = c.take<float>(p->cap * DX_D);
        p->y = c.take<float>(p->cap * DX_D);
        p->o = c.take<float>(p->cap * DX_D);
        p->tmp = c.take<float>(p->cap * DX_D);
        p->qkv = c.take<float>(p->cap * 3 * DX_D);
        p->h1 = c.take<float>(p->cap * DX_C);
        p->h2 = c.take<float>(p->cap * DX_C);
        p->melp = c.take<float>(p->cap_t * DX_MEL);
    } else {
        const int64_t rp = p->rows_phys;
        p->row_stats = c.take<float>(p->cap * TC_STAT_TILES * 2);
        p->xb = c.take<__nv_bfloat16>(rp * DX_D);
        p->yb = c.take<__nv_bfloat16>(rp * DX_D);
        p->ob = c.take<__nv_bfloat16>(rp * DX_D);
        p->qkvb = c.take<__nv_bfloat16>(rp * 3 * DX_D);
        p->h1b = c.take<__nv_bfloat16>(rp * DX_C);
        p->h2b = c.take<__nv_bfloat16>(rp * DX_C);
        p->melb = c.take<__nv_bfloat16>(rp * DX_MEL_PAD);
        p->ffn_partial = c.take<float>(ffn_partial_floats());
        p->ffn_flags = c.take<uint32_t>(FF_MAX_CTAS);
    }
    p->bytes = round_up(c.off, 256);
    if (workspace != nullptr) {
        DX_REQUIRE(workspace_bytes >= p->bytes, "plan: workspace of %lld bytes < required %lld", (long long)workspace_bytes,
                   (long long)p->bytes);
        DX_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "plan: workspace must be 256-byte aligned");
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// one FFT block (FFTBlock.forward, model.py:249-259) on the row stream; x is both input and output.
// film: (B, 8, 256) table or null; film_block: global block index (0-3 encoder, 4-7 decoder).
// ---------------------------------------------------------------------------------------------------------
static int fft_block_f32(const DxBlockWeights& bw, const Plan& p, const Rows& R, const float* film, int film_block,
                         cudaStream_t s) {
    const float* fp = film ? film + film_block * 256 : nullptr;
    DX_TRY(gemm_f32(p.x, DX_D, bw.w_qkv, bw.b_qkv, 3 * DX_D, 1, false, R, -1, p.qkv, s));
    DX_TRY(attention_f32(p.qkv, R, p.o, s));
    DX_TRY(gemm_f32(p.o, DX_D, bw.w_out, bw.b_out, DX_D, 1, false, R, -1, p.tmp, s));
    DX_TRY(ln_rows_f32(p.tmp, p.x, bw.ln1_g, bw.ln1_b, DX_D, nullptr, 0, R, 0, p.y, s));
    DX_TRY(gemm_f32(p.y, DX_D, bw.w_c1, bw.b_c1, DX_C, DX_TAPS, true, R, 1, p.h1, s));
    DX_TRY(gemm_f32(p.h1, DX_C, bw.w_c2, bw.b_c2, DX_D, DX_TAPS, false, R, -1, p.tmp, s));
    DX_TRY(ln_rows_f32(p.tmp, p.y, bw.ln2_g, bw.ln2_b, DX_D, fp, 8 * 256, R, 0, p.x, s));
    return 0;
}

static TcGemmArgs tc_args(const __nv_bfloat16* a, int64_t rp, int K, int N, int taps, const void* w, int nt, int ks,
                          const float* bias, int relu, int epi, int live_ext) {
    TcGemmArgs g;
    memset(&g, 0, sizeof(g));
    g.a = a; g.a_rows_phys = rp; g.K = K; g.N = N; g.taps = taps;
    g.w = reinterpret_cast<const __nv_bfloat16*>(w); g.nt = nt; g.ks = ks;
    g.bias = bias; g.relu = relu; g.epilogue = epi; g.live_ext = live_ext;
    g.out_rows_phys = rp; g.res_rows_phys = rp;
    return g;
}

// K = 128 linear layers go to the weight-resident kernel (debug knob 4: use the generic conv-GEMM kernel instead)
static int linear_bf16(const TcGemmArgs& g, const Rows& R, cudaStream_t s) {
    return debug_knob(4) ? tc_gemm(g, R, s) : tc_linear(g, R, s);
}

// One FFT block in four launches (debug knob 9; also the A/B reference of the fused stack below)
static int fft_block_bf16(const DxBlockWeights& bw, const Plan& p, const Rows& R, const float* film, int film_block,
                          cudaStream_t s) {
    const int64_t rp = p.rows_phys;
    DX_REQUIRE(bw.w_qkv_bf && bw.w_out_bf && bw.w_c1_bf && bw.w_c2_bf, "bf16 mode: packed weights missing (model not prepared for DX_MODE_BF16)");
    {   // QKV projection
        TcGemmArgs g = tc_args(p.xb, rp, DX_D, 3 * DX_D, 1, bw.w_qkv_bf, 128, 64, bw.b_qkv, 0, TC_EPI_BF16, -1);
        g.out_bf = p.qkvb; g.tag = "tc_qkv";
        DX_TRY(linear_bf16(g, R, s));
    }
    {   // attention, then out-proj + residual + LN + mask
        TcGemmArgs g = tc_args(p.ob, rp, DX_D, DX_D, 1, bw.w_out_bf, 128, 64, bw.b_out, 0, TC_EPI_LN, 0);
        g.out_bf = p.yb; g.res = p.xb; g.ln_g = bw.ln1_g; g.ln_b = bw.ln1_b; g.tag = "tc_outproj_ln";
        DX_TRY(attention_bf16(p.qkvb, p.ob, rp, R, s));
        DX_TRY(linear_bf16(g, R, s));
    }
    if (!debug_knob(3)) {
        const float* fp = film ? film + film_block * 256 : nullptr;
        const FfnSplit sp = p.ffn_split();
        return ffn_bf16(p.yb, p.xb, rp, bw, fp, 8 * 256, R, s, nullptr, &sp);
    }
    {   // conv1 + ReLU (hidden live one row past the utterance: H1)
        TcGemmArgs g = tc_args(p.yb, rp, DX_D, DX_C, DX_TAPS, bw.w_c1_bf, 256, 32, bw.b_c1, 1, TC_EPI_BF16, 1);
        g.out_bf = p.h1b; g.tag = "tc_ffn_conv1";
        DX_TRY(tc_gemm(g, R, s));
    }
    {   // conv2 + residual + LN + FiLM + mask
        TcGemmArgs g = tc_args(p.h1b, rp, DX_C, DX_D, DX_TAPS, bw.w_c2_bf, 128, 64, bw.b_c2, 0, TC_EPI_LN, 0);
        g.out_bf = p.xb; g.res = p.yb; g.ln_g = bw.ln2_g; g.ln_b = bw.ln2_b;
        g.film = film ? film + film_block * 256 : nullptr;
        g.film_stride = 8 * 256; g.tag = "tc_ffn_conv2_ln";
        DX_TRY(tc_gemm(g, R, s));
    }
    return 0;
}

// A stack of four FFT blocks (FFTBlock.forward x 4, model.py:249-259, called at :299-300 / :531-532 / :704-705) on the
// row stream in p.xb, result back in p.xb.  Two launches per block: attention, then ONE kernel for out-proj + LN1 +
// FFN + LN2/FiLM + the next block's QKV projection; the stack's first QKV projection is its own launch.  The block
// input and output alternate between p.xb and p.yb (neighbouring tiles read the input rows as residual halos).
static int fft_stack_bf16(const DxBlockWeights* blocks, int n_blocks, const Plan& p, const Rows& R, const float* film,
                          int film_block0, const char* tag, cudaStream_t s) {
    if (debug_knob(9) || debug_knob(3)) {
        for (int i = 0; i < n_blocks; ++i) DX_TRY(fft_block_bf16(blocks[i], p, R, film, film_block0 + i, s));
        return 0;
    }
    const int64_t rp = p.rows_phys;
    __nv_bfloat16 *cur = p.xb, *nxt = p.yb;
    for (int i = 0; i < n_blocks; ++i) {
        const DxBlockWeights& bw = blocks[i];
        DX_REQUIRE(bw.w_qkv_bf && bw.w_out_bf && bw.w_ffn1_bf && bw.w_ffn2_bf, "bf16 mode: packed weights missing (model not prepared for DX_MODE_BF16)");
        if (i == 0) {
            TcGemmArgs g = tc_args(cur, rp, DX_D, 3 * DX_D, 1, bw.w_qkv_bf, 128, 64, bw.b_qkv, 0, TC_EPI_BF16, -1);
            g.out_bf = p.qkvb; g.tag = "tc_qkv";
            DX_TRY(linear_bf16(g, R, s));
        }
        DX_TRY(attention_bf16(p.qkvb, p.ob, rp, R, s));
        FfnFuse f;
        f.attn = p.ob; f.x_in = cur;
        f.next = i < n_blocks - 1 ? &blocks[i + 1] : nullptr;
        f.qkv_out = i < n_blocks - 1 ? p.qkvb : nullptr;
        f.tag = tag;
        const float* fp = film ? film + (film_block0 + i) * 256 : nullptr;
        const FfnSplit sp = p.ffn_split();
        DX_TRY(ffn_bf16(nullptr, nxt, rp, bw, fp, 8 * 256, R, s, &f, &sp));
        __nv_bfloat16* t = cur; cur = nxt; nxt = t;
    }
    // an odd number of blocks leaves the result in p.yb: the consumers read p.xb
    if (cur != p.xb) DX_CHECK_CUDA(cudaMemcpyAsync(p.xb, cur, (size_t)rp * DX_D * sizeof(__nv_bfloat16), cudaMemcpyDeviceToDevice, s));
    return 0;
}

int zero_plan_halos(const Plan& p, cudaStream_t s) {
    // flags of the fused FFN kernel's hidden-dimension split: every launch publishes under its own epoch, this only guards
    // against stale memory that happens to hold a future epoch
    DX_CHECK_CUDA(cudaMemsetAsync(p.ffn_flags, 0, FF_MAX_CTAS * sizeof(uint32_t), s));
    HaloList hl;
    memset(&hl, 0, sizeof(hl));
    __nv_bfloat16* bufs[7] = {p.xb, p.yb, p.ob, p.qkvb, p.h1b, p.h2b, p.melb};
    const int chunks[7] = {16, 16, 16, 48, 128, 128, DX_MEL_PAD / 8};
    for (int i = 0; i < 7; ++i) { hl.buf[i] = bufs[i]; hl.chunks[i] = chunks[i]; }
    hl.n = 7;
    return zero_halos(hl, p.rows_phys, p.cap, s);
}

// ---------------------------------------------------------------------------------------------------------
// AccentEncoder.forward (model.py:679-716)
// ---------------------------------------------------------------------------------------------------------
int run_accent_encoder(const DxModelWeights& w, const Plan& p, const Rows& RT, const float* frames_energy,
                       const float* frames_pitch, const float* mel_specs, float* accent_emb, cudaStream_t s,
                       bool packed_frames) {
    DX_REQUIRE(frames_energy && frames_pitch && mel_specs, "accent encoder: frames_energy / frames_pitch / mel_specs required");
    if (p.mode == DX_MODE_F32) {
        DX_TRY(pack_mel(mel_specs, RT, p.melp, nullptr, 0, s, packed_frames));
        // pre-net: 3 x (conv k3 → ReLU → LN), evaluated on the padded layout: ghost rows live (H1)
        DX_TRY(gemm_f32(p.melp, DX_MEL, w.ae_c0_w, w.ae_c0_b, DX_C, DX_TAPS, true, RT, -1, p.h1, s));
        DX_TRY(ln_rows_f32(p.h1, nullptr, w.ae_ln0_g, w.ae_ln0_b, DX_C, nullptr, 0, RT, 2, p.h2, s));
        DX_TRY(gemm_f32(p.h2, DX_C, w.ae_c1_w, w.ae_c1_b, DX_C, DX_TAPS, true, RT, -1, p.h1, s));
        DX_TRY(ln_rows_f32(p.h1, nullptr, w.ae_ln1_g, w.ae_ln1_b, DX_C, nullptr, 0, RT, 1, p.h2, s));
        DX_TRY(gemm_f32(p.h2, DX_C, w.ae_c2_w, w.ae_c2_b, DX_D, DX_TAPS, true, RT, -1, p.tmp, s));
        DX_TRY(ln_rows_f32(p.tmp, nullptr, w.ae_ln2_g, w.ae_ln2_b, DX_D, nullptr, 0, RT, 0, p.o, s));
        DX_TRY(accent_input(p.o, nullptr, frames_energy, frames_pitch, w.ae_energy_w, w.ae_energy_b, w.ae_pitch_w,
                            w.ae_pitch_b, w.pos_table, RT, p.x, nullptr, 0, s, packed_frames));
        for (int i = 0; i < n_blocks_of(w.n_ae_blocks); ++i) DX_TRY(fft_block_f32(w.ae_blocks[i], p, RT, nullptr, 0, s));
        DX_TRY(masked_mean(p.x, nullptr, 0, RT, accent_emb, s));
    } else {
        const int64_t rp = p.rows_phys;
        DX_REQUIRE(w.ae_c0_w_bf && w.ae_c1_w_bf && w.ae_c2_w_bf, "bf16 mode: packed pre-net weights missing");
        DX_TRY(pack_mel(mel_specs, RT, nullptr, p.melb, rp, s, packed_frames));
        {
            TcGemmArgs g = tc_args(p.melb, rp, DX_MEL_PAD, DX_C, DX_TAPS, w.ae_c0_w_bf, 256, 32, w.ae_c0_b, 1, TC_EPI_BF16, 2);
            g.out_bf = p.h1b; g.row_stats = p.row_stats; g.tag = "tc_prenet0";
            DX_TRY(tc_gemm(g, RT, s));
        }
        DX_TRY(ln_apply_bf16(p.h1b, rp, DX_C, p.row_stats, w.ae_ln0_g, w.ae_ln0_b, RT, 2, s));
        {
            TcGemmArgs g = tc_args(p.h1b, rp, DX_C, DX_C, DX_TAPS, w.ae_c1_w_bf, 256, 32, w.ae_c1_b, 1, TC_EPI_BF16, 1);
            g.out_bf = p.h2b; g.row_stats = p.row_stats; g.tag = "tc_prenet1";
            DX_TRY(tc_gemm(g, RT, s));
        }
        DX_TRY(ln_apply_bf16(p.h2b, rp, DX_C, p.row_stats, w.ae_ln1_g, w.ae_ln1_b, RT, 1, s));
        {
            TcGemmArgs g = tc_args(p.h2b, rp, DX_C, DX_D, DX_TAPS, w.ae_c2_w_bf, 128, 64, w.ae_c2_b, 1, TC_EPI_LN, 0);
            g.out_bf = p.ob; g.ln_g = w.ae_ln2_g; g.ln_b = w.ae_ln2_b; g.tag = "tc_prenet2_ln";
            DX_TRY(tc_gemm(g, RT, s));
        }
        DX_TRY(accent_input(nullptr, p.ob, frames_energy, frames_pitch, w.ae_energy_w, w.ae_energy_b, w.ae_pitch_w,
                            w.ae_pitch_b, w.pos_table, RT, nullptr, p.xb, rp, s, packed_frames));
        DX_TRY(fft_stack_bf16(w.ae_blocks, n_blocks_of(w.n_ae_blocks), p, RT, nullptr, 0, "tc_block_tail_frames", s));
        DX_TRY(masked_mean(nullptr, p.xb, rp, RT, accent_emb, s));
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// DaftExprt.forward (model.py:889-948) / the device part of DaftExprt.inference (model.py:1039-1107)
// ---------------------------------------------------------------------------------------------------------
int run_forward(const DxModelWeights& w, const DxForwardArgs& a, cudaStream_t s) {
    Plan p;
    DX_TRY(plan_layout(&p, a.B, a.L_max, a.T_max, a.mode, a.workspace, a.workspace_bytes));
    DX_REQUIRE(a.workspace != nullptr, "dx_forward: workspace required");
    DX_REQUIRE(a.symbols && a.durations_float && a.durations_int && a.symbols_energy && a.symbols_pitch && a.input_lengths,
               "dx_forward: symbol-side inputs required");
    DX_REQUIRE(a.film && a.mel_out && a.weights_out && a.output_lengths_out, "dx_forward: film / mel_out / weights_out / output_lengths_out required");
    DX_REQUIRE((a.io_flags & ~(DX_IO_PACKED_FRAMES | DX_IO_PACKED_MEL_OUT)) == 0, "dx_forward: unknown io_flags 0x%x", a.io_flags);
    DX_REQUIRE(w.n_ae_blocks >= 0 && w.n_ae_blocks <= 4 && w.n_pe_blocks >= 0 && w.n_pe_blocks <= 4 && w.n_fd_blocks >= 0 &&
                   w.n_fd_blocks <= 4, "dx_forward: block counts %d / %d / %d out of range (1..4, 0 = 4)", w.n_ae_blocks,
               w.n_pe_blocks, w.n_fd_blocks);
    const bool bf = a.mode == DX_MODE_BF16;
    const int64_t rp = p.rows_phys;
    if (bf) DX_TRY(zero_plan_halos(p, s));

    // frame-side lengths: given (forward) or Σ durations_int clamped to >= 1 (inference, model.py:1096-1098)
    DX_TRY(lenreg_scan(a.durations_int, a.B, a.L_max, a.output_lengths == nullptr, nullptr, p.mu, p.out_len, nullptr, s));
    const int64_t* out_len = a.output_lengths ? a.output_lengths : p.out_len;
    DX_CHECK_CUDA(cudaMemcpyAsync(a.output_lengths_out, out_len, a.B * sizeof(int64_t), cudaMemcpyDeviceToDevice, s));
    Rows RT = p.rows_t(out_len);
    Rows RL = p.rows_l(a.input_lengths);
    DX_TRY(build_rows(out_len, a.B, a.T_max, p.t_row_start, p.t_row_b, p.t_row_p, p.cap_t, s));
    DX_TRY(build_rows(a.input_lengths, a.B, a.L_max, p.l_row_start, p.l_row_b, p.l_row_p, p.cap_l, s));
    if (bf) {   // attention K/V tiles of the last utterance read up to 127 rows past the stream: keep them finite
        DX_TRY(zero_tail_rows(p.qkvb, 48, rp, RT, 128, s));
        DX_TRY(zero_tail_rows(p.qkvb, 48, rp, RL, 128, s));
    }

    // 1-2. accent embedding
    float* accent = a.accent_emb ? a.accent_emb : p.accent;
    if (a.accent_emb_in != nullptr) {
        DX_CHECK_CUDA(cudaMemcpyAsync(accent, a.accent_emb_in, (size_t)a.B * DX_D * sizeof(float), cudaMemcpyDeviceToDevice, s));
    } else {
        DX_TRY(run_accent_encoder(w, p, RT, a.frames_energy, a.frames_pitch, a.mel_specs, accent, s,
                                  (a.io_flags & DX_IO_PACKED_FRAMES) != 0));
    }
    // 3-4. speaker projection, classifier, FiLM table
    DX_TRY(style(&w, a.B, a.spk_embs, a.spk_emb_in, accent, a.film, a.speaker_preds, s));
    // 5. phoneme encoder
    DX_TRY(embed_symbols(a.symbols, w.symbols_embedding, w.n_symbols, w.pos_table, RL, bf ? nullptr : p.x,
                         bf ? p.xb : nullptr, rp, s));
    const int n_pe = n_blocks_of(w.n_pe_blocks), n_fd = n_blocks_of(w.n_fd_blocks);
    if (bf) DX_TRY(fft_stack_bf16(w.pe_blocks, n_pe, p, RL, a.film, 0, "tc_block_tail_symbols", s));
    else for (int i = 0; i < n_pe; ++i) DX_TRY(fft_block_f32(w.pe_blocks[i], p, RL, a.film, i, s));
    if (a.enc_out) DX_TRY(unpack_rows(bf ? nullptr : p.x, bf ? p.xb : nullptr, rp, RL, a.enc_out, s));
    // 6. length regulator → decoder input rows (x + P, masked)
    DX_TRY(lenreg_symbols(bf ? nullptr : p.x, bf ? p.xb : nullptr, rp, RL, a.durations_float, a.symbols_energy,
                          a.symbols_pitch, &w, p.xs, p.sigma, s));
    DX_TRY(lenreg_frames(p.xs, RL, RT, p.mu, p.sigma, w.pos_table, a.weights_out, bf ? nullptr : p.x,
                         bf ? p.xb : nullptr, rp, s));
    // 7. frame decoder
    if (bf) DX_TRY(fft_stack_bf16(w.fd_blocks, n_fd, p, RT, a.film, n_pe, "tc_block_tail_frames", s));
    else for (int i = 0; i < n_fd; ++i) DX_TRY(fft_block_f32(w.fd_blocks[i], p, RT, a.film, n_pe + i, s));
    if (bf) {
        DX_REQUIRE(w.mel_w_bf, "bf16 mode: packed mel projection missing");
        TcGemmArgs g = tc_args(p.xb, rp, DX_D, DX_MEL, 1, w.mel_w_bf, 128, 64, w.mel_b, 0, TC_EPI_F32, -1);
        g.out_f32 = p.melo; g.tag = "tc_melproj";
        DX_TRY(linear_bf16(g, RT, s));
    } else {
        DX_TRY(gemm_f32(p.x, DX_D, w.mel_w, w.mel_b, DX_MEL, 1, false, RT, -1, p.melo, s));
    }
    DX_TRY(unpack_mel(p.melo, RT, a.mel_out, s, (a.io_flags & DX_IO_PACKED_MEL_OUT) != 0));
    return 0;
}

}  // namespace dx
