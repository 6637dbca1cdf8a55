// Training step (BASELINE config 5): launcher declarations of the backward kernels (dx_train_kernels.cu) and the
// whole-step planner (dx_train.cu).  fp32 row-stream layout throughout ([rows][C], see dx_f32.cu).
#pragma once
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_dropout.cuh"

namespace dx {

struct DxLossParams {
    float mel_spec_weight, post_mult_weight, energy_weight, speaker_weight, scale;
};

// per-utterance record of style_bwd (floats)
constexpr int SB_E = 0, SB_C = 192, SB_DG = 320, SB_DQ = 1344, SB_DC = 2368, SB_H1 = 2496, SB_DH1 = 2624, SB_H2 = 2752,
              SB_DH2 = 2880, SB_DPOST = 3008, SB_DACC = 3024, SB_STRIDE = 3152;

int reduce_partials(const float* part, int S, int64_t n, float* out, const int32_t* total_rows, int chunk_rows, float scale,
                    cudaStream_t s, float* out2 = nullptr, int64_t split = 0);
int ln_bwd_rows(const float* in, const float* res, const float* g, int C, const float* film, int film_stride, const float* dy1,
                const float* dy2, const Rows& R, int live_ext, float* dv, float* dg, float* db, float* scratch, cudaStream_t s,
                const LnDrop* drop = nullptr);
int64_t ln_bwd_scratch_floats(int64_t cap, int C);
int film_bwd(const float* in, const float* res, const float* g, const float* bta, const float* dy1, const float* dy2,
             const Rows& R, float* dfilm, int film_stride, float* scratch, cudaStream_t s);
int64_t film_bwd_scratch_floats(int B);
int colsum_rows(const float* in, int N, const Rows& R, float* out, float* scratch, cudaStream_t s);
int64_t colsum_scratch_floats(int64_t cap, int N);
int wgrad_f32(const float* A, int K, const float* G, int N, int taps, const Rows& R, float* dW, float* scratch, cudaStream_t s);
int64_t wgrad_scratch_floats(int64_t cap, int K, int N, int taps);
int transpose_weight(const float* W, int taps, int K, int N, float* Wt, cudaStream_t s);
int relu_bwd(float* d, const float* act, int64_t n, cudaStream_t s);
int attention_bwd_f32(const float* qkv, const float* O, const float* lse, const float* dO, const Rows& R, float* Dr, float* dqkv,
                      cudaStream_t s);
int masked_mean_bwd(const float* dacc, const Rows& R, float* dx, cudaStream_t s);
int add_mask_rows(const float* a, const float* b, const Rows& R, float* out, cudaStream_t s);
int scalar_conv_wgrad(const float* g, const float* x, const Rows& R, float* dw, float* db, float* scratch, cudaStream_t s);
int64_t scalar_conv_scratch_floats(int64_t cap);
int embed_bwd(const float* g, const int64_t* symbols, int n_symbols, const Rows& R, float* demb, cudaStream_t s);
int outer_wgrad(const float* dOut, int N, const float* In, int K, int B, float* dW, float* db, cudaStream_t s);
int lenreg_bwd(const float* g, const float* xdec, const float* pos, const float* weights, const float* xs, const float* mu,
               const float* sigma, const float* dur, const DxModelWeights* w, const Rows& RL, const Rows& RT, float* ct, float* gx,
               float* gd, float* gr, float* dpre, cudaStream_t s);
int sum_rows_scalar(const float* v, const Rows& R, float* out, cudaStream_t s);
int style_bwd(const float* spk_embs, const float* accent, const DxModelWeights* w, const float* dfilm, const float* dlogits,
              float lambda_rev, int B, float* v, cudaStream_t s);
int gather_strided(const float* v, int64_t stride, int off, int n, int B, float* out, cudaStream_t s);
int loss_forward_backward(const float* melo, const float* tgt, const Rows& RT, const float* logits, const int64_t* speaker_ids,
                          int n_spk, const float* post_mult, int n_post, const DxLossParams& lp, float* part, float* dmel,
                          float* dlogits, float* losses, float* dpost_reg, cudaStream_t s);
int post_mult_grad(const float* v, int B, const float* dreg, int n_post, float* out, cudaStream_t s);
int adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float b1, float b2, float eps, float wd, int step,
              float gscale, cudaStream_t s);

// dx_train_tc.cu: tensor-core GEMMs of the training step (DX_TRAIN_TC_GEMM)
int rows_to_il(const float* src, int C, const Rows& R, __nv_bfloat16* dst, int64_t rows_phys, cudaStream_t s, int C_src = 0);
int copy_pad3(const float* src, int T, int Ks, int Ns, float* dst, int Kd, int Nd, cudaStream_t s);
int il_to_rows(const __nv_bfloat16* src, int C, const Rows& R, int64_t rows_phys, float* dst, cudaStream_t s);
int wgrad_tc(const __nv_bfloat16* a_il, int K, const __nv_bfloat16* g_il, int N, int taps, const Rows& R, int64_t rows_phys, float* dW,
             float* part, cudaStream_t s);
int64_t wgrad_tc_scratch_floats();
// dx_attn_bwd_tc.cu: attention backward on tcgen05 from bf16 interleaved copies of QKV and dO
int attention_bwd_tc(const __nv_bfloat16* qkv_il, const __nv_bfloat16* dO_il, int64_t rows_phys, const float* lse, const float* Dr,
                     const Rows& R, float* dqkv, cudaStream_t s, const DropSpec* drop = nullptr, uint32_t drop_site_id = 0);
int attn_rowdot(const float* dO, const float* O, const Rows& R, float* Dr, cudaStream_t s);      // Dr[r][h] = dO[r] . O[r] per head

int64_t train_workspace_bytes(int B, int L_max, int T_max);
int run_train_step(const DxModelWeights& w, const DxModelWeights& g, const DxTrainArgs& a, cudaStream_t s);

}  // namespace dx
