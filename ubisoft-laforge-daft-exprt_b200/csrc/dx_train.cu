// Whole training step of BASELINE config 5 on one GPU: forward with saved activations, the reference's loss and
// the hand-written backward of every module, enqueued on the caller's stream without host synchronisation.
//
// Replaces, for one rank, `outputs = model(inputs); loss, _ = criterion(outputs, targets, iteration); loss.backward()`
// (train.py:421-434: DaftExprt.forward model.py:889-948 under autograd + DaftExprtLoss.forward loss.py:57-159).  Dropout is
// the identity here (the parity configuration of SURVEY.md section 8d cfg5; the reference's dropout sites are model.py:167-168,
// 216, 656, 663, 670).  The gradient of every one of the 184 parameters is written in the ENGINE layout of its weight
// (DxModelWeights: conv weights [tap][Cin][Cout], in/out projections transposed, everything else torch layout); the
// data-parallel exchange is one all-reduce of that flat buffer, done by the caller between this call and
// dx_adam_step.  Arithmetic: fp32 SIMT kernels (the 1e-3 parity mode); the forward half is the same kernels DX_MODE_F32
// inference runs, so the loss is the inference forward's to the last bit.
#include <string.h>
#include "dx_train.cuh"
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

constexpr int TRAIN_MAX_PACKED = 128;
// bf16 elements of the packed forms of the step's dense weights: per block QKV, out-proj, conv1, conv2; the pre-net's second and
// third convs; each in forward and transposed form (both have taps * K * N elements: every width is a multiple of 128)
constexpr int64_t TRAIN_PACKED_ELEMS =
    2 * (12 * ((int64_t)DX_D * 3 * DX_D + DX_D * DX_D + 2 * (int64_t)DX_TAPS * DX_D * DX_C) + (int64_t)DX_TAPS * DX_C * DX_C +
         (int64_t)DX_TAPS * DX_C * DX_D);

struct BlockSave {
    float *xin, *qkv, *lse, *o, *t1, *x1, *h, *t2;
    // tensor-core mode: the bf16 channel-interleaved copies the forward's contractions made of their inputs, kept for the
    // backward (A operand of the weight gradients, Q / K / V of the attention backward)
    __nv_bfloat16 *xin_il, *qkv_il, *o_il, *x1_il, *h_il;
};
struct StackSave {
    BlockSave blk[4];
    float* out;
    int64_t cap;
};

struct TrainPlan {
    int B, L_max, T_max;
    int64_t cap_t, cap_l, cap, bytes;
    int32_t *t_row_start, *t_row_b, *t_row_p, *l_row_start, *l_row_b, *l_row_p;
    StackSave acc, enc, dec;
    float *melp, *a0, *z0, *a1, *z1, *a2, *z2;
    __nv_bfloat16 *z0_il, *z1_il;               // as BlockSave's copies, for the accent pre-net
    float *accent, *film, *logits, *xs, *sigma, *mu, *weights, *melo;
    int64_t* out_len;
    // backward
    float *dA, *dB, *dC, *dQKV, *dH1, *dH2, *dmel, *Dr, *wt, *scratch;
    float *fb_part;
    float *dfilm, *dlogits, *sv, *gat1, *gat2, *lpart, *dpost_reg, *ct, *gd, *gr, *dpre;
    int64_t scratch_floats;
    int n_blk[3];                               // FFT blocks of the accent encoder, phoneme encoder, frame decoder (1..4)
    DropSpec drop_attn[3], drop_conv[3];       // per stack (accent encoder, phoneme encoder, frame decoder); thr = 0: off
    // DX_TRAIN_TC_GEMM: every dense weight packed to bf16 once per step, forward form and transposed (data gradient) form
    PackEntry pk[TRAIN_MAX_PACKED];
    int n_pk;
    PackEntry* pk_dev;
    __nv_bfloat16* pk_store;
    int64_t pk_store_elems;
    // DX_TRAIN_TC_GEMM: bf16 channel-interleaved copies of a GEMM's operands, packed bf16 weight, row stride of the copies
    bool tc;
    __nv_bfloat16 *il_a, *il_g, *w_bf;
    float* wpad;                // zero-padded weight / padded weight gradient (widths that are not multiples of 128)
    int64_t rows_phys;
};

// the step's pre-packed bf16 form of weight W (forward, or transposed for the data gradient), or null
const __nv_bfloat16* find_packed(const TrainPlan& p, const float* W, bool transposed) {
    for (int i = 0; i < p.n_pk; ++i)
        if (p.pk[i].src == W && (p.pk[i].transposed != 0) == transposed) return p.pk[i].dst;
    return nullptr;
}

// out = conv(A, W) (+ bias, ReLU, liveness mask): gemm_f32, or - tensor-core mode - tcgen05 on bf16 copies (dx_train_tc.cu).  An input
// width that is not a multiple of 128 (the 80 mel channels) is zero-padded in the copies.
// il_keep: where the bf16 copy of A goes when the backward will want it again (default: the plan's scratch copy)
int gemm(const TrainPlan& p, const float* A, int K, const float* W, const float* bias, int N, int taps, bool relu, const Rows& R,
         int live_ext, float* out, cudaStream_t s, __nv_bfloat16* il_keep = nullptr) {
    if (!(p.tc && K % 4 == 0 && K <= 1024 && N <= 1024 && (taps == 1 || taps == 3)))
        return gemm_f32(A, K, W, bias, N, taps, relu, R, live_ext, out, s);
    const int Kp = (K + 127) / 128 * 128;
    const int nt = (N % 256 == 0) ? 256 : 128, ks = nt == 256 ? 32 : 64;
    __nv_bfloat16* a_il = il_keep ? il_keep : p.il_a;
    DX_TRY(rows_to_il(A, Kp, R, a_il, p.rows_phys, s, K));
    const __nv_bfloat16* w_bf = find_packed(p, W, false);
    if (w_bf == nullptr) {
        if (Kp != K) {
            DX_TRY(copy_pad3(W, taps, K, N, p.wpad, Kp, N, s));
            W = p.wpad;
        }
        DX_TRY(pack_weight_bf16(W, taps, Kp, N, nt, ks, p.w_bf, s));
        w_bf = p.w_bf;
    }
    TcGemmArgs g;
    memset(&g, 0, sizeof(g));
    g.a = a_il; g.a_rows_phys = p.rows_phys; g.K = Kp; g.N = N; g.taps = taps;
    g.w = w_bf; g.nt = nt; g.ks = ks;
    g.bias = bias; g.relu = relu ? 1 : 0; g.epilogue = TC_EPI_F32; g.live_ext = live_ext;
    g.out_f32 = out; g.tag = "tc_gemm_train";
    return tc_gemm(g, R, s);
}

// dW[tap][k][n] = sum_r A[r + tap - half][k] G[r][n]: wgrad_f32, or - tensor-core mode - wgrad_tc on bf16 copies (widths padded to
// multiples of 128, the padded gradient cropped afterwards)
int wgrad(const TrainPlan& p, const float* A, int K, const float* G, int N, int taps, const Rows& R, float* dW, cudaStream_t s) {
    if (!(p.tc && K % 4 == 0 && N % 4 == 0 && K <= 1024 && N <= 1024)) return wgrad_f32(A, K, G, N, taps, R, dW, p.scratch, s);
    const int Kp = (K + 127) / 128 * 128, Np = (N + 127) / 128 * 128;
    DX_TRY(rows_to_il(A, Kp, R, p.il_a, p.rows_phys, s, K));
    DX_TRY(rows_to_il(G, Np, R, p.il_g, p.rows_phys, s, N));
    if (Kp == K && Np == N) return wgrad_tc(p.il_a, K, p.il_g, N, taps, R, p.rows_phys, dW, p.scratch, s);
    DX_TRY(wgrad_tc(p.il_a, Kp, p.il_g, Np, taps, R, p.rows_phys, p.wpad, p.scratch, s));
    return copy_pad3(p.wpad, taps, Kp, Np, dW, K, N, s);
}

void carve_stack(Carver& c, StackSave& st, int64_t cap, int64_t rows_phys) {
    st.cap = cap;
    for (int i = 0; i < 4; ++i) {
        BlockSave& b = st.blk[i];
        b.xin = c.take<float>(cap * DX_D);
        b.qkv = c.take<float>(cap * 3 * DX_D);
        b.lse = c.take<float>(cap * DX_HEADS);
        b.o = c.take<float>(cap * DX_D);
        b.t1 = c.take<float>(cap * DX_D);
        b.x1 = c.take<float>(cap * DX_D);
        b.h = c.take<float>(cap * DX_C);
        b.t2 = c.take<float>(cap * DX_D);
        b.xin_il = c.take<__nv_bfloat16>(rows_phys * DX_D);
        b.qkv_il = c.take<__nv_bfloat16>(rows_phys * 3 * DX_D);
        b.o_il = c.take<__nv_bfloat16>(rows_phys * DX_D);
        b.x1_il = c.take<__nv_bfloat16>(rows_phys * DX_D);
        b.h_il = c.take<__nv_bfloat16>(rows_phys * DX_C);
    }
    st.out = c.take<float>(cap * DX_D);
}

int64_t max64(int64_t a, int64_t b) { return a > b ? a : b; }

int train_layout(TrainPlan* p, int B, int L_max, int T_max, int n_speakers, void* ws, int64_t ws_bytes) {
    DX_REQUIRE(B >= 1 && L_max >= 1 && T_max >= 1, "train plan: bad shape B=%d L_max=%d T_max=%d", B, L_max, T_max);
    DX_REQUIRE(T_max <= 5000 && L_max <= 5000, "train plan: sequence longer than the 5000-row positional table (model.py:123)");
    memset(p, 0, sizeof(*p));
    p->B = B; p->L_max = L_max; p->T_max = T_max;
    p->cap_t = row_cap(B, T_max);
    p->cap_l = row_cap(B, L_max);
    p->cap = max64(p->cap_t, p->cap_l);
    Carver c{reinterpret_cast<uint8_t*>(ws), 0};
    p->t_row_start = c.take<int32_t>(B + 1);
    p->t_row_b = c.take<int32_t>(p->cap_t);
    p->t_row_p = c.take<int32_t>(p->cap_t);
    p->l_row_start = c.take<int32_t>(B + 1);
    p->l_row_b = c.take<int32_t>(p->cap_l);
    p->l_row_p = c.take<int32_t>(p->cap_l);
    p->rows_phys = p->cap + DX_HALO + DX_TAIL;
    carve_stack(c, p->acc, p->cap_t, p->rows_phys);
    carve_stack(c, p->enc, p->cap_l, p->rows_phys);
    carve_stack(c, p->dec, p->cap_t, p->rows_phys);
    p->z0_il = c.take<__nv_bfloat16>(p->rows_phys * DX_C);
    p->z1_il = c.take<__nv_bfloat16>(p->rows_phys * DX_C);
    p->melp = c.take<float>(p->cap_t * DX_MEL);
    p->a0 = c.take<float>(p->cap_t * DX_C);
    p->z0 = c.take<float>(p->cap_t * DX_C);
    p->a1 = c.take<float>(p->cap_t * DX_C);
    p->z1 = c.take<float>(p->cap_t * DX_C);
    p->a2 = c.take<float>(p->cap_t * DX_D);
    p->z2 = c.take<float>(p->cap_t * DX_D);
    p->accent = c.take<float>((int64_t)B * DX_D);
    p->film = c.take<float>((int64_t)B * 8 * 256);
    p->logits = c.take<float>((int64_t)B * (n_speakers > 0 ? n_speakers : 1));
    p->xs = c.take<float>(p->cap_l * DX_D);
    p->sigma = c.take<float>((int64_t)B * L_max);
    p->mu = c.take<float>((int64_t)B * L_max);
    p->weights = c.take<float>((int64_t)B * L_max * T_max);
    p->melo = c.take<float>(p->cap_t * DX_MEL);
    p->out_len = c.take<int64_t>(B);
    p->dA = c.take<float>(p->cap * DX_D);
    p->dB = c.take<float>(p->cap * DX_D);
    p->dC = c.take<float>(p->cap * DX_D);
    p->dQKV = c.take<float>(p->cap * 3 * DX_D);
    p->dH1 = c.take<float>(p->cap * DX_C);
    p->dH2 = c.take<float>(p->cap * DX_C);
    p->dmel = c.take<float>(p->cap_t * DX_MEL);
    p->Dr = c.take<float>(p->cap * DX_HEADS);
    p->wt = c.take<float>((int64_t)DX_TAPS * DX_C * DX_C);
    int64_t sf = wgrad_scratch_floats(p->cap_t, DX_C, DX_C, DX_TAPS);
    sf = max64(sf, ln_bwd_scratch_floats(p->cap, DX_C));
    sf = max64(sf, colsum_scratch_floats(p->cap, DX_C));
    sf = max64(sf, scalar_conv_scratch_floats(p->cap));
    sf = max64(sf, wgrad_tc_scratch_floats());
    p->scratch_floats = sf;
    p->scratch = c.take<float>(sf);
    p->dfilm = c.take<float>((int64_t)B * 8 * 256);
    p->fb_part = c.take<float>(film_bwd_scratch_floats(B));
    p->dlogits = c.take<float>((int64_t)B * (n_speakers > 0 ? n_speakers : 1));
    p->sv = c.take<float>((int64_t)B * SB_STRIDE);
    p->gat1 = c.take<float>((int64_t)B * 1024);
    p->gat2 = c.take<float>((int64_t)B * 1024);
    p->lpart = c.take<float>((int64_t)B * 4);
    p->dpost_reg = c.take<float>(16);
    p->ct = c.take<float>(p->cap_t);
    p->gd = c.take<float>(p->cap_l * DX_D);
    p->gr = c.take<float>(p->cap_l * DX_D);
    p->dpre = c.take<float>(p->cap_l);
    p->il_a = c.take<__nv_bfloat16>(p->rows_phys * DX_C);
    p->il_g = c.take<__nv_bfloat16>(p->rows_phys * DX_C);
    p->w_bf = c.take<__nv_bfloat16>((int64_t)DX_TAPS * DX_C * DX_C);
    p->pk_dev = c.take<PackEntry>(TRAIN_MAX_PACKED);
    p->pk_store = c.take<__nv_bfloat16>(TRAIN_PACKED_ELEMS);
    p->wpad = c.take<float>((int64_t)DX_TAPS * DX_C * DX_C);
    p->bytes = round_up(c.off, 256);
    if (ws != nullptr) {
        DX_REQUIRE(ws_bytes >= p->bytes, "train plan: workspace of %lld bytes < required %lld", (long long)ws_bytes, (long long)p->bytes);
        DX_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 255) == 0, "train plan: workspace must be 256-byte aligned");
    }
    return 0;
}

inline float* G(const float* p) { return const_cast<float*>(p); }

// ---- forward of one stack with saved activations (FFTBlock.forward x 4, model.py:249-259) -----------------------------
// stack: 0 accent encoder, 1 phoneme encoder, 2 frame decoder (the dropout site numbering of dx_dropout.cuh)
int stack_forward(const TrainPlan& p, const DxBlockWeights* blocks, StackSave& st, const Rows& R, const float* film, int film_block0,
                  int stack, cudaStream_t s) {
    const DropSpec &da = p.drop_attn[stack], &dc = p.drop_conv[stack];
    const int nb = p.n_blk[stack];
    for (int i = 0; i < nb; ++i) {
        const DxBlockWeights& bw = blocks[i];
        BlockSave& b = st.blk[i];
        float* y = i < nb - 1 ? st.blk[i + 1].xin : st.out;
        const float* fp = film ? film + (film_block0 + i) * 256 : nullptr;
        DX_TRY(gemm(p, b.xin, DX_D, bw.w_qkv, bw.b_qkv, 3 * DX_D, 1, false, R, -1, b.qkv, s, b.xin_il));
        if (p.tc && !debug_knob(17)) {
            // the inference path's tcgen05 attention on a bf16 copy of the QKV rows; it also writes the log-sum-exp
            DX_TRY(rows_to_il(b.qkv, 3 * DX_D, R, b.qkv_il, p.rows_phys, s));
            DX_TRY(attention_bf16(b.qkv_il, p.il_g, p.rows_phys, R, s, b.lse, &da, drop_site(DROP_ATTN_W, stack, i)));
            DX_TRY(il_to_rows(p.il_g, DX_D, R, p.rows_phys, b.o, s));
        } else {
            DX_TRY(attention_f32(b.qkv, R, b.o, s, b.lse));
        }
        DX_TRY(gemm(p, b.o, DX_D, bw.w_out, bw.b_out, DX_D, 1, false, R, -1, b.t1, s, b.o_il));
        // train mode: the dropouts behind the attention (model.py:189) and the feed-forward (:216) are applied by the LayerNorm
        // kernel that reads the branch, which also writes the dropped branch back for the backward
        const LnDrop d1{da, drop_site(DROP_ATTN_OUT, stack, i), LN_DROP_IN, nullptr};
        const LnDrop d2{dc, drop_site(DROP_FFN_OUT, stack, i), LN_DROP_IN, nullptr};
        DX_TRY(ln_rows_f32(b.t1, b.xin, bw.ln1_g, bw.ln1_b, DX_D, nullptr, 0, R, 0, b.x1, s, &d1));
        DX_TRY(gemm(p, b.x1, DX_D, bw.w_c1, bw.b_c1, DX_C, DX_TAPS, true, R, 1, b.h, s, b.x1_il));
        DX_TRY(gemm(p, b.h, DX_C, bw.w_c2, bw.b_c2, DX_D, DX_TAPS, false, R, -1, b.t2, s, b.h_il));
        DX_TRY(ln_rows_f32(b.t2, b.x1, bw.ln2_g, bw.ln2_b, DX_D, fp, 8 * 256, R, 0, y, s, &d2));
    }
    return 0;
}

// dgrad of a conv / linear layer: dX = sum_taps dY[r + j - half] Wt[j]   with Wt the tap-flipped transpose of W
int dgrad(const TrainPlan& p, const float* dY, int N_out, const float* W, int K_in, int taps, const Rows& R, float* dX,
          const float* addend, cudaStream_t s) {
    DX_TRY(transpose_weight(W, taps, K_in, N_out, p.wt, s));
    if (addend == nullptr) return gemm(p, dY, N_out, p.wt, nullptr, K_in, taps, false, R, -1, dX, s);
    return gemm_f32(dY, N_out, p.wt, nullptr, K_in, taps, false, R, -1, dX, s, addend);
}

// Backward of one conv / linear layer y = conv(act, W): dW = wgrad(act, G) and dX = dgrad(G).  In tensor-core mode the bf16 copy
// of G serves both contractions (B operand of the weight gradient, A operand of the data gradient).
// act_il: the forward's bf16 copy of `act` when it was kept (gemm's il_keep)
// relu_src: `act` is the output of a ReLU whose input gradient is wanted: dX = (act > 0) * dgrad (fused into the dgrad's epilogue)
int conv_bwd(const TrainPlan& p, const float* act, int K, const float* Gr, int N, const float* W, int taps, const Rows& R, float* dW,
             float* dX, cudaStream_t s, const __nv_bfloat16* act_il = nullptr, bool relu_src = false) {
    if (!(p.tc && K % 128 == 0 && N % 128 == 0 && K <= 1024 && N <= 1024)) {
        DX_TRY(wgrad(p, act, K, Gr, N, taps, R, dW, s));
        DX_TRY(dgrad(p, Gr, N, W, K, taps, R, dX, nullptr, s));
        return relu_src ? relu_bwd(dX, act, R.cap * K, s) : 0;
    }
    if (act_il == nullptr) {
        DX_TRY(rows_to_il(act, K, R, p.il_a, p.rows_phys, s));
        act_il = p.il_a;
    }
    DX_TRY(rows_to_il(Gr, N, R, p.il_g, p.rows_phys, s));
    DX_TRY(wgrad_tc(act_il, K, p.il_g, N, taps, R, p.rows_phys, dW, p.scratch, s));
    const int nt = (K % 256 == 0) ? 256 : 128, ks = nt == 256 ? 32 : 64;
    const __nv_bfloat16* w_bf = find_packed(p, W, true);
    if (w_bf == nullptr) {
        DX_TRY(transpose_weight(W, taps, K, N, p.wt, s));
        DX_TRY(pack_weight_bf16(p.wt, taps, N, K, nt, ks, p.w_bf, s));
        w_bf = p.w_bf;
    }
    TcGemmArgs g;
    memset(&g, 0, sizeof(g));
    g.a = p.il_g; g.a_rows_phys = p.rows_phys; g.K = N; g.N = K; g.taps = taps;
    g.w = w_bf; g.nt = nt; g.ks = ks;
    g.epilogue = TC_EPI_F32; g.live_ext = -1;
    g.out_f32 = dX; g.tag = "tc_gemm_train";
    g.relu_mask = relu_src ? act : nullptr;
    return tc_gemm(g, R, s);
}

// ---- backward of one stack.  In: gradient of the stack output = dy1 (+ dy2).  Out: gradient of the stack input = *g1 + *g2
// (two buffers of the plan; the caller masks and adds them).  Parameter gradients go to gb[i].
int stack_backward(const TrainPlan& p, const DxBlockWeights* blocks, const DxBlockWeights* gb, const StackSave& st, const Rows& R,
                   const float* film, float* dfilm, int film_block0, int stack, const float* dy1, const float* dy2, const float** g1,
                   const float** g2, cudaStream_t s) {
    const DropSpec &da = p.drop_attn[stack], &dc = p.drop_conv[stack];
    const bool tcs = p.tc;                           // the forward kept bf16 copies of the contractions' inputs
    float *A = p.dA, *Bf = p.dB, *C = p.dC;
    // the incoming pair may live in two of A / B / C: the first LayerNorm backward must write into the third
    auto is = [](const float* x, const float* y) { return x == y; };
    float* bufs[3] = {p.dA, p.dB, p.dC};
    int freei = 0;
    for (int k = 0; k < 3; ++k)
        if (!is(bufs[k], dy1) && !is(bufs[k], dy2)) { freei = k; break; }
    A = bufs[freei];
    // the two other plan buffers, in an order that keeps dy1 / dy2 alive until the first two kernels have read them
    Bf = bufs[(freei + 1) % 3];
    C = bufs[(freei + 2) % 3];
    for (int i = p.n_blk[stack] - 1; i >= 0; --i) {
        const DxBlockWeights& bw = blocks[i];
        const DxBlockWeights& g = gb[i];
        const BlockSave& b = st.blk[i];
        const float* fp = film ? film + (film_block0 + i) * 256 : nullptr;
        // LN2 + FiLM + mask.  A = gradient of t2 + x1 (the residual branch's share); in train mode the kernel also writes
        // dT2 = mask * A / (1 - p), the gradient of conv2's output, into dH2 (unused by the stacks)
        const LnDrop d2{dc, drop_site(DROP_FFN_OUT, stack, i), LN_DROP_IN, p.dH2};
        const LnDrop d1{da, drop_site(DROP_ATTN_OUT, stack, i), LN_DROP_IN, p.dH2};
        DX_TRY(ln_bwd_rows(b.t2, b.x1, bw.ln2_g, DX_D, fp, 8 * 256, dy1, dy2, R, 0, A, G(g.ln2_g), G(g.ln2_b), p.scratch, s, &d2));
        if (film) DX_TRY(film_bwd(b.t2, b.x1, bw.ln2_g, bw.ln2_b, dy1, dy2, R, dfilm + (film_block0 + i) * 256, 8 * 256, p.fb_part, s));
        const float* dT2 = dc.thr ? p.dH2 : A;
        DX_TRY(colsum_rows(dT2, DX_D, R, G(g.b_c2), p.scratch, s));
        DX_TRY(conv_bwd(p, b.h, DX_C, dT2, DX_D, bw.w_c2, DX_TAPS, R, G(g.w_c2), p.dH1, s, tcs ? b.h_il : nullptr, true));
        // conv1
        DX_TRY(colsum_rows(p.dH1, DX_C, R, G(g.b_c1), p.scratch, s));
        DX_TRY(conv_bwd(p, b.x1, DX_D, p.dH1, DX_C, bw.w_c1, DX_TAPS, R, G(g.w_c1), Bf, s, tcs ? b.x1_il : nullptr));   // Bf = d x1 through the FFN
        // LN1 + mask: d x1 = Bf + A (residual of the FFN)
        DX_TRY(ln_bwd_rows(b.t1, b.xin, bw.ln1_g, DX_D, nullptr, 0, Bf, A, R, 0, C, G(g.ln1_g), G(g.ln1_b), p.scratch, s, &d1));
        // out-proj (C = gradient of t1 + x; dT1 = its masked copy in train mode)
        const float* dT1 = da.thr ? p.dH2 : C;
        DX_TRY(colsum_rows(dT1, DX_D, R, G(g.b_out), p.scratch, s));
        DX_TRY(conv_bwd(p, b.o, DX_D, dT1, DX_D, bw.w_out, 1, R, G(g.w_out), A, s, tcs ? b.o_il : nullptr));    // A = dO
        if (p.tc && !debug_knob(16)) {
            // attention backward on tcgen05 (dx_attn_bwd_tc.cu) from bf16 copies of the saved QKV rows and of dO
            DX_TRY(attn_rowdot(A, b.o, R, p.Dr, s));
            const __nv_bfloat16* qkv_il = b.qkv_il;
            if (debug_knob(17)) {                    // the fp32 attention forward made no copy
                DX_TRY(rows_to_il(b.qkv, 3 * DX_D, R, p.il_a, p.rows_phys, s));
                qkv_il = p.il_a;
            }
            DX_TRY(rows_to_il(A, DX_D, R, p.il_g, p.rows_phys, s));
            DX_TRY(attention_bwd_tc(qkv_il, p.il_g, p.rows_phys, b.lse, p.Dr, R, p.dQKV, s, &da, drop_site(DROP_ATTN_W, stack, i)));
        } else {
            DX_TRY(attention_bwd_f32(b.qkv, b.o, b.lse, A, R, p.Dr, p.dQKV, s));
        }
        DX_TRY(colsum_rows(p.dQKV, 3 * DX_D, R, G(g.b_qkv), p.scratch, s));
        DX_TRY(conv_bwd(p, b.xin, DX_D, p.dQKV, 3 * DX_D, bw.w_qkv, 1, R, G(g.w_qkv), Bf, s, tcs ? b.xin_il : nullptr));   // Bf = d x through QKV
        // gradient of the block input = Bf + C (residual of the attention); becomes (dy1, dy2) of block i - 1
        dy1 = Bf; dy2 = C;          // the next block's first LayerNorm backward writes into A, which holds neither
    }
    *g1 = dy1; *g2 = dy2;
    return 0;
}

}  // namespace

int64_t train_workspace_bytes(int B, int L_max, int T_max) {
    TrainPlan p;
    if (train_layout(&p, B, L_max, T_max, 64, nullptr, 0) != 0) return -1;
    return p.bytes;
}

int run_train_step(const DxModelWeights& w, const DxModelWeights& g, const DxTrainArgs& a, cudaStream_t s) {
    TrainPlan p;
    DX_REQUIRE(w.n_speakers >= 1 && w.n_speakers <= 64, "train: n_speakers=%d out of range (1..64)", w.n_speakers);
    DX_REQUIRE(w.n_ae_blocks >= 0 && w.n_ae_blocks <= 4 && w.n_pe_blocks >= 0 && w.n_pe_blocks <= 4 && w.n_fd_blocks >= 0 &&
                   w.n_fd_blocks <= 4, "dx_train_step: block counts %d / %d / %d out of range (1..4, 0 = 4)", w.n_ae_blocks,
               w.n_pe_blocks, w.n_fd_blocks);
    DX_TRY(train_layout(&p, a.B, a.L_max, a.T_max, w.n_speakers, a.workspace, a.workspace_bytes));
    DX_REQUIRE(a.workspace != nullptr, "dx_train_step: workspace required");
    DX_REQUIRE((a.flags & ~(DX_TRAIN_FORWARD_ONLY | DX_TRAIN_TC_GEMM)) == 0, "dx_train_step: unknown flags 0x%x", a.flags);
    p.tc = (a.flags & DX_TRAIN_TC_GEMM) != 0;
    p.n_blk[0] = n_blocks_of(w.n_ae_blocks); p.n_blk[1] = n_blocks_of(w.n_pe_blocks); p.n_blk[2] = n_blocks_of(w.n_fd_blocks);
    const int n_pe = p.n_blk[1], n_film = p.n_blk[1] + p.n_blk[2];       // FiLM rows: phoneme encoder first, then the decoder
    bool any_drop = false;
    for (int k = 0; k < 3; ++k) {
        DX_REQUIRE(a.attn_dropout[k] >= 0.f && a.attn_dropout[k] < 1.f && a.conv_dropout[k] >= 0.f && a.conv_dropout[k] < 1.f,
                   "dx_train_step: dropout probabilities must be in [0, 1)");
        p.drop_attn[k] = make_drop(a.dropout_seed, a.attn_dropout[k]);
        p.drop_conv[k] = make_drop(a.dropout_seed, a.conv_dropout[k]);
        any_drop = any_drop || p.drop_attn[k].thr || p.drop_conv[k].thr;
    }
    DX_REQUIRE(!any_drop || (p.tc && !debug_knob(16) && !debug_knob(17)),
               "dx_train_step: the train-mode forward (dropout) is implemented on the DX_TRAIN_TC_GEMM path only");
    DX_REQUIRE(a.symbols && a.durations_float && a.durations_int && a.symbols_energy && a.symbols_pitch && a.input_lengths &&
                   a.frames_energy && a.frames_pitch && a.mel_specs && a.output_lengths && a.speaker_ids && a.spk_embs && a.losses,
               "dx_train_step: every input of DaftExprt.forward, speaker_ids and the losses buffer are required");
    const int B = a.B;
    Rows RT{p.t_row_start, p.t_row_b, p.t_row_p, a.output_lengths, B, a.T_max, p.cap_t};
    Rows RL{p.l_row_start, p.l_row_b, p.l_row_p, a.input_lengths, B, a.L_max, p.cap_l};
    DX_TRY(build_rows(a.output_lengths, B, a.T_max, p.t_row_start, p.t_row_b, p.t_row_p, p.cap_t, s));
    DX_TRY(build_rows(a.input_lengths, B, a.L_max, p.l_row_start, p.l_row_b, p.l_row_p, p.cap_l, s));

    if (p.tc && !debug_knob(19)) {
        // pack every dense weight (forward form, and transposed for the data gradient) in ONE launch; knob 19: per use, as before
        auto add = [&](const float* W, int taps, int K, int N) {
            for (int tr = 0; tr < 2; ++tr) {
                PackEntry& e = p.pk[p.n_pk++];
                e.src = W; e.transposed = tr; e.taps = taps;
                e.K = tr ? N : K; e.N = tr ? K : N;
                e.nt = (e.N % 256 == 0) ? 256 : 128; e.ks = e.nt == 256 ? 32 : 64;
                e.n_elems = packed_weight_elems(taps, e.K, e.N, e.nt, e.ks);
                e.dst = p.pk_store + p.pk_store_elems;
                p.pk_store_elems += e.n_elems;
            }
        };
        const DxBlockWeights* stacks[3] = {w.ae_blocks, w.pe_blocks, w.fd_blocks};
        for (int k = 0; k < 3; ++k)
            for (int i = 0; i < p.n_blk[k]; ++i) {
                const DxBlockWeights& bw = stacks[k][i];
                add(bw.w_qkv, 1, DX_D, 3 * DX_D);
                add(bw.w_out, 1, DX_D, DX_D);
                add(bw.w_c1, DX_TAPS, DX_D, DX_C);
                add(bw.w_c2, DX_TAPS, DX_C, DX_D);
            }
        add(w.ae_c1_w, DX_TAPS, DX_C, DX_C);
        add(w.ae_c2_w, DX_TAPS, DX_C, DX_D);
        DX_REQUIRE(p.n_pk <= TRAIN_MAX_PACKED && p.pk_store_elems <= TRAIN_PACKED_ELEMS, "train: packed-weight table overflow");
        DX_CHECK_CUDA(cudaMemcpyAsync(p.pk_dev, p.pk, sizeof(PackEntry) * p.n_pk, cudaMemcpyHostToDevice, s));
        DX_TRY(pack_weight_batch(p.pk_dev, p.n_pk, s));
    }

    // ======================================= forward (model.py:889-948) =======================================
    DX_TRY(lenreg_scan(a.durations_int, B, a.L_max, false, nullptr, p.mu, p.out_len, nullptr, s));
    // accent encoder (model.py:679-716)
    DX_TRY(pack_mel(a.mel_specs, RT, p.melp, nullptr, 0, s));
    // train mode: the pre-net's three dropouts (model.py:656,663,670) act on the LayerNorm outputs, inside the LayerNorm kernels
    LnDrop dpn[3];
    for (int k = 0; k < 3; ++k) dpn[k] = LnDrop{p.drop_conv[0], drop_site(DROP_PRENET, 0, k), LN_DROP_OUT, nullptr};
    DX_TRY(gemm(p, p.melp, DX_MEL, w.ae_c0_w, w.ae_c0_b, DX_C, DX_TAPS, true, RT, -1, p.a0, s));
    DX_TRY(ln_rows_f32(p.a0, nullptr, w.ae_ln0_g, w.ae_ln0_b, DX_C, nullptr, 0, RT, 2, p.z0, s, &dpn[0]));
    DX_TRY(gemm(p, p.z0, DX_C, w.ae_c1_w, w.ae_c1_b, DX_C, DX_TAPS, true, RT, -1, p.a1, s, p.z0_il));
    DX_TRY(ln_rows_f32(p.a1, nullptr, w.ae_ln1_g, w.ae_ln1_b, DX_C, nullptr, 0, RT, 1, p.z1, s, &dpn[1]));
    DX_TRY(gemm(p, p.z1, DX_C, w.ae_c2_w, w.ae_c2_b, DX_D, DX_TAPS, true, RT, -1, p.a2, s, p.z1_il));
    DX_TRY(ln_rows_f32(p.a2, nullptr, w.ae_ln2_g, w.ae_ln2_b, DX_D, nullptr, 0, RT, 0, p.z2, s, &dpn[2]));
    DX_TRY(accent_input(p.z2, nullptr, a.frames_energy, a.frames_pitch, w.ae_energy_w, w.ae_energy_b, w.ae_pitch_w, w.ae_pitch_b,
                        w.pos_table, RT, p.acc.blk[0].xin, nullptr, 0, s));
    DX_TRY(stack_forward(p, w.ae_blocks, p.acc, RT, nullptr, 0, 0, s));
    DX_TRY(masked_mean(p.acc.out, nullptr, 0, RT, p.accent, s));
    // conditioning (model.py:904-920)
    DX_TRY(style(&w, B, a.spk_embs, nullptr, p.accent, p.film, p.logits, s));
    // phoneme encoder (model.py:590-610)
    DX_TRY(embed_symbols(a.symbols, w.symbols_embedding, w.n_symbols, w.pos_table, RL, p.enc.blk[0].xin, nullptr, 0, s));
    DX_TRY(stack_forward(p, w.pe_blocks, p.enc, RL, p.film, 0, 1, s));
    // length regulator (model.py:417-510)
    DX_TRY(lenreg_symbols(p.enc.out, nullptr, 0, RL, a.durations_float, a.symbols_energy, a.symbols_pitch, &w, p.xs, p.sigma, s));
    DX_TRY(lenreg_frames(p.xs, RL, RT, p.mu, p.sigma, w.pos_table, p.weights, p.dec.blk[0].xin, nullptr, 0, s));
    // frame decoder (model.py:548-564)
    DX_TRY(stack_forward(p, w.fd_blocks, p.dec, RT, p.film, n_pe, 2, s));
    DX_TRY(gemm(p, p.dec.out, DX_D, w.mel_w, w.mel_b, DX_MEL, 1, false, RT, -1, p.melo, s));
    if (a.mel_out) DX_TRY(unpack_mel(p.melo, RT, a.mel_out, s));

    // ======================================= loss (loss.py:57-159) ============================================
    DxLossParams lp{a.mel_spec_weight, a.post_mult_weight, a.energy_consistency_weight, a.speaker_weight, a.loss_scale};
    DX_TRY(loss_forward_backward(p.melo, a.mel_specs, RT, p.logits, a.speaker_ids, w.n_speakers, w.post_mult, 2 * n_film, lp, p.lpart, p.dmel,
                                 p.dlogits, a.losses, p.dpost_reg, s));
    if (a.flags & DX_TRAIN_FORWARD_ONLY) return 0;

    // ======================================= backward =========================================================
    const float *g1 = nullptr, *g2 = nullptr;
    // mel projection (model.py:561)
    DX_TRY(colsum_rows(p.dmel, DX_MEL, RT, G(g.mel_b), p.scratch, s));
    DX_TRY(conv_bwd(p, p.dec.out, DX_D, p.dmel, DX_MEL, w.mel_w, 1, RT, G(g.mel_w), p.dB, s));
    DX_CHECK_CUDA(cudaMemsetAsync(p.dfilm, 0, (size_t)B * 8 * 256 * sizeof(float), s));
    // frame decoder
    DX_TRY(stack_backward(p, w.fd_blocks, g.fd_blocks, p.dec, RT, p.film, p.dfilm, n_pe, 2, p.dB, nullptr, &g1, &g2, s));
    // decoder input = (x_up + P) * valid: mask, then the length regulator
    float* gdec = (g1 != p.dA && g2 != p.dA) ? p.dA : ((g1 != p.dB && g2 != p.dB) ? p.dB : p.dC);
    DX_TRY(add_mask_rows(g1, g2, RT, gdec, s));
    float* gx = (gdec == p.dA) ? p.dB : p.dA;          // d x' on the symbol stream (cap_l <= cap rows)
    DX_TRY(lenreg_bwd(gdec, p.dec.blk[0].xin, w.pos_table, p.weights, p.xs, p.mu, p.sigma, a.durations_float, &w, RL, RT, p.ct, gx,
                      p.gd, p.gr, p.dpre, s));
    DX_TRY(colsum_rows(p.gr, DX_D, RL, G(g.gu_range_w), p.scratch, s));
    DX_TRY(sum_rows_scalar(p.dpre, RL, G(g.gu_range_b), s));
    DX_TRY(scalar_conv_wgrad(gx, a.symbols_energy, RL, G(g.gu_energy_w), G(g.gu_energy_b), p.scratch, s));
    DX_TRY(scalar_conv_wgrad(gx, a.symbols_pitch, RL, G(g.gu_pitch_w), G(g.gu_pitch_b), p.scratch, s));
    DX_TRY(scalar_conv_wgrad(p.gd, a.durations_float, RL, G(g.gu_dur_w), G(g.gu_dur_b), p.scratch, s));
    // phoneme encoder
    DX_TRY(stack_backward(p, w.pe_blocks, g.pe_blocks, p.enc, RL, p.film, p.dfilm, 0, 1, gx, nullptr, &g1, &g2, s));
    float* genc = (g1 != p.dA && g2 != p.dA) ? p.dA : ((g1 != p.dB && g2 != p.dB) ? p.dB : p.dC);
    DX_TRY(add_mask_rows(g1, g2, RL, genc, s));
    DX_TRY(embed_bwd(genc, a.symbols, w.n_symbols, RL, G(g.symbols_embedding), s));
    // conditioning heads (model.py:904-920, 764-806, 829-830)
    DX_TRY(style_bwd(a.spk_embs, p.accent, &w, p.dfilm, p.dlogits, a.lambda_reversal, B, p.sv, s));
    DX_TRY(gather_strided(p.sv, SB_STRIDE, SB_DG, n_film * DX_D, B, p.gat1, s));
    DX_TRY(gather_strided(p.sv, SB_STRIDE, SB_C, DX_D, B, p.gat2, s));
    DX_TRY(outer_wgrad(p.gat1, n_film * DX_D, p.gat2, DX_D, B, G(g.film_g_w), G(g.film_g_b), s));
    DX_TRY(gather_strided(p.sv, SB_STRIDE, SB_DQ, n_film * DX_D, B, p.gat1, s));
    DX_TRY(outer_wgrad(p.gat1, n_film * DX_D, p.gat2, DX_D, B, G(g.film_b_w), G(g.film_b_b), s));
    DX_TRY(gather_strided(p.sv, SB_STRIDE, SB_DC, DX_D, B, p.gat1, s));
    DX_TRY(gather_strided(p.sv, SB_STRIDE, SB_E, DX_SPK, B, p.gat2, s));
    DX_TRY(outer_wgrad(p.gat1, DX_D, p.gat2, DX_SPK, B, G(g.spk_w), G(g.spk_b), s));
    DX_TRY(gather_strided(p.sv, SB_STRIDE, SB_DH1, DX_D, B, p.gat1, s));
    DX_TRY(outer_wgrad(p.gat1, DX_D, p.accent, DX_D, B, G(g.cls_w1), G(g.cls_b1), s));
    DX_TRY(gather_strided(p.sv, SB_STRIDE, SB_DH2, DX_D, B, p.gat1, s));
    DX_TRY(gather_strided(p.sv, SB_STRIDE, SB_H1, DX_D, B, p.gat2, s));
    DX_TRY(outer_wgrad(p.gat1, DX_D, p.gat2, DX_D, B, G(g.cls_w2), G(g.cls_b2), s));
    DX_TRY(gather_strided(p.sv, SB_STRIDE, SB_H2, DX_D, B, p.gat2, s));
    DX_TRY(outer_wgrad(p.dlogits, w.n_speakers, p.gat2, DX_D, B, G(g.cls_w3), G(g.cls_b3), s));
    if (g.post_mult) DX_TRY(post_mult_grad(p.sv, B, p.dpost_reg, 2 * n_film, G(g.post_mult), s));
    // accent encoder: mean pool, stack, pre-net
    DX_TRY(gather_strided(p.sv, SB_STRIDE, SB_DACC, DX_D, B, p.gat1, s));
    DX_TRY(masked_mean_bwd(p.gat1, RT, p.dB, s));
    DX_TRY(stack_backward(p, w.ae_blocks, g.ae_blocks, p.acc, RT, nullptr, nullptr, 0, 0, p.dB, nullptr, &g1, &g2, s));
    float* gacc = (g1 != p.dA && g2 != p.dA) ? p.dA : ((g1 != p.dB && g2 != p.dB) ? p.dB : p.dC);
    DX_TRY(add_mask_rows(g1, g2, RT, gacc, s));
    DX_TRY(scalar_conv_wgrad(gacc, a.frames_energy, RT, G(g.ae_energy_w), G(g.ae_energy_b), p.scratch, s));
    DX_TRY(scalar_conv_wgrad(gacc, a.frames_pitch, RT, G(g.ae_pitch_w), G(g.ae_pitch_b), p.scratch, s));
    float* d2 = (gacc == p.dA) ? p.dB : p.dA;
    DX_TRY(ln_bwd_rows(p.a2, nullptr, w.ae_ln2_g, DX_D, nullptr, 0, gacc, nullptr, RT, 0, d2, G(g.ae_ln2_g), G(g.ae_ln2_b), p.scratch, s, &dpn[2]));
    DX_TRY(relu_bwd(d2, p.a2, p.cap_t * DX_D, s));
    DX_TRY(colsum_rows(d2, DX_D, RT, G(g.ae_c2_b), p.scratch, s));
    DX_TRY(conv_bwd(p, p.z1, DX_C, d2, DX_D, w.ae_c2_w, DX_TAPS, RT, G(g.ae_c2_w), p.dH1, s, p.tc ? p.z1_il : nullptr));
    DX_TRY(ln_bwd_rows(p.a1, nullptr, w.ae_ln1_g, DX_C, nullptr, 0, p.dH1, nullptr, RT, 1, p.dH2, G(g.ae_ln1_g), G(g.ae_ln1_b), p.scratch, s, &dpn[1]));
    DX_TRY(relu_bwd(p.dH2, p.a1, p.cap_t * DX_C, s));
    DX_TRY(colsum_rows(p.dH2, DX_C, RT, G(g.ae_c1_b), p.scratch, s));
    DX_TRY(conv_bwd(p, p.z0, DX_C, p.dH2, DX_C, w.ae_c1_w, DX_TAPS, RT, G(g.ae_c1_w), p.dH1, s, p.tc ? p.z0_il : nullptr));
    DX_TRY(ln_bwd_rows(p.a0, nullptr, w.ae_ln0_g, DX_C, nullptr, 0, p.dH1, nullptr, RT, 2, p.dH2, G(g.ae_ln0_g), G(g.ae_ln0_b), p.scratch, s, &dpn[0]));
    DX_TRY(relu_bwd(p.dH2, p.a0, p.cap_t * DX_C, s));
    DX_TRY(colsum_rows(p.dH2, DX_C, RT, G(g.ae_c0_b), p.scratch, s));
    DX_TRY(wgrad(p, p.melp, DX_MEL, p.dH2, DX_C, DX_TAPS, RT, G(g.ae_c0_w), s));
    return 0;
}

}  // namespace dx
