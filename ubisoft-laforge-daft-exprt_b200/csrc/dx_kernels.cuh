// Internal launcher declarations (one per kernel family).  See the .cu files for the reference citations.
#pragma once
#include "dx_common.cuh"

namespace dx {

// dx_rows.cu
// packed: the frame-side tensor is in the DX_IO_PACKED_* layout (per-utterance (80, T_b) blocks / float[sum T_b])
int pack_mel(const float* mel, const Rows& R, float* out_f32, __nv_bfloat16* out_bf, int64_t rows_phys, cudaStream_t s, bool packed = false);
int unpack_mel(const float* rows, const Rows& R, float* mel, cudaStream_t s, bool packed = false);
int embed_symbols(const int64_t* symbols, const float* emb, int n_symbols, const float* pos, const Rows& R,
                  float* out_f32, __nv_bfloat16* out_bf, int64_t rows_phys, cudaStream_t s);
int accent_input(const float* z2_f32, const __nv_bfloat16* z2_bf, const float* energy, const float* pitch,
                 const float* ew, const float* eb, const float* pw, const float* pb, const float* pos,
                 const Rows& R, float* out_f32, __nv_bfloat16* out_bf, int64_t rows_phys, cudaStream_t s, bool packed = false);
int masked_mean(const float* x_f32, const __nv_bfloat16* x_bf, int64_t rows_phys, const Rows& R, float* out,
                cudaStream_t s);
int unpack_rows(const float* x_f32, const __nv_bfloat16* x_bf, int64_t rows_phys, const Rows& R, float* out,
                cudaStream_t s);

struct HaloList {
    __nv_bfloat16* buf[8];
    int chunks[8];
    int n;
};
int zero_halos(const HaloList& hl, int64_t rows_phys, int64_t cap, cudaStream_t s);
int ln_apply_bf16(__nv_bfloat16* x, int64_t rows_phys, int C, const float* stats, const float* g, const float* b,
                  const Rows& R, int live_ext, cudaStream_t s);

// dx_attn_tc.cu: masked multi-head self-attention, bf16 operands on tcgen05, fp32 softmax
int zero_tail_rows(__nv_bfloat16* buf, int nchunks, int64_t rows_phys, const Rows& R, int n_rows, cudaStream_t s);
// lse (training): natural-log log-sum-exp of the scaled scores, fp32 [rows][2], written at valid rows
// drop (training): dropout of the attention weights (after the softmax normalisation, as nn.MultiheadAttention applies it)
struct DropSpec;
int attention_bf16(const __nv_bfloat16* qkv, __nv_bfloat16* out, int64_t rows_phys, const Rows& R, cudaStream_t s, float* lse = nullptr,
                   const DropSpec* drop = nullptr, uint32_t drop_site_id = 0);

// dx_ffn_tc.cu: fused conv1 + ReLU + conv2 + residual + LN + FiLM + mask; optionally with the attention output
// projection + residual + LN1 in front (attn != null: x is ignored, x_in is the block input) and the next block's QKV
// projection behind (qkv_out != null, next = that block's weights)
struct FfnFuse {
    const __nv_bfloat16* attn;
    const __nv_bfloat16* x_in;
    const DxBlockWeights* next;
    __nv_bfloat16* qkv_out;
    const char* tag;            // profiler tag (null: "tc_ffn_fused")
};
// split of a tile's hidden dimension across CTAs (stream-K, see dx_ffn_tc.cu): scratch for the partial sums
// (ffn_partial_floats() floats) and FF_MAX_CTAS flags the caller zeroes once per pass; null = whole tiles per CTA
constexpr int FF_MAX_CTAS = 160;
struct FfnSplit {
    float* partial;
    uint32_t* flags;
};
int64_t ffn_partial_floats();
int ffn_bf16(const __nv_bfloat16* x, __nv_bfloat16* out, int64_t rows_phys, const DxBlockWeights& bw, const float* film,
             int film_stride, const Rows& R, cudaStream_t s, const FfnFuse* fuse = nullptr, const FfnSplit* split = nullptr);

// dx_style.cu
int style(const DxModelWeights* w, int B, const float* spk_embs, const float* spk_emb_in, const float* accent,
          float* film, float* speaker_preds, cudaStream_t s);
int prosody_transform(float* energy, float* pitch, const float* energy_factors, const float* pitch_factors,
                      const int64_t* dint, const float* pmean, const float* pstd, int B, int L, int transform,
                      cudaStream_t s);

// dx_lenreg.cu
int lenreg_symbols(const float* x_f32, const __nv_bfloat16* x_bf, int64_t rows_phys, const Rows& RL,
                   const float* dur, const float* energy, const float* pitch, const DxModelWeights* w, float* xs,
                   float* sigma, cudaStream_t s);
int lenreg_scan(const int64_t* dint, int B, int L, bool clamp_one, int64_t* cumsum, float* mu, int64_t* out_len,
                int64_t* two_mu, cudaStream_t s);
int frame2symbol(const int64_t* cumsum, const int64_t* in_len, const int64_t* out_len, int B, int L, int T,
                 int64_t* f2s, cudaStream_t s);
int lenreg_frames(const float* xs, const Rows& RL, const Rows& RT, const float* mu, const float* sigma,
                  const float* pos, float* weights, float* xo_f32, __nv_bfloat16* xo_bf, int64_t rows_phys_t,
                  cudaStream_t s);
int int_durations(float* duration, int B, int L, int sr, int flt, int hop, int centered, int64_t* dint,
                  int32_t* status, cudaStream_t s);

// dx_tc_gemm.cu (tcgen05 tensor-core path)
constexpr int TC_STAT_TILES = 4;   // N = 1024 as four 256-wide tiles (accent pre-net)
enum TcEpilogue : int {
    TC_EPI_F32 = 0,        // out fp32 [rows][N] = acc (+bias)                        (tests, mel projection)
    TC_EPI_BF16 = 1,       // out bf16 interleaved = act(acc + bias), liveness-masked (QKV, conv1 hidden, pre-net)
    TC_EPI_LN = 2,         // out bf16 interleaved = mask(film(LN(act(acc + bias) + res) * g + b))   (N == 128)
};
struct TcGemmArgs {
    const __nv_bfloat16* a;     // channel-interleaved activations [K/8][a_rows_phys][8]
    int64_t a_rows_phys;
    int K, N, taps;
    const __nv_bfloat16* w;     // packed weights, see dx_pack_weight_bf16
    int nt, ks;                 // packing parameters
    const float* bias;          // [N] or null
    int relu;
    int epilogue;
    int live_ext;               // -1: no liveness masking
    // TC_EPI_F32
    float* out_f32;             // [rows][N]
    const float* relu_mask;     // [rows][N] or null: out = relu_mask > 0 ? out : 0 (training: ReLU backward fused into the dgrad)
    // TC_EPI_BF16 / TC_EPI_LN
    __nv_bfloat16* out_bf;      // [N/8][out_rows_phys][8]
    int64_t out_rows_phys;
    // TC_EPI_LN
    const __nv_bfloat16* res;   // residual, interleaved [16][res_rows_phys][8] or null
    int64_t res_rows_phys;
    const float *ln_g, *ln_b;
    const float* film;          // (B, 8, 256) pre-offset to the block, or null
    int film_stride;
    // optional per-row statistics for a following LayerNorm over N > 128 (pre-net): sum / sum of squares
    float* row_stats;           // [rows][TC_STAT_TILES][2] or null: per-N-tile partial sums (deterministic, no atomics)
    const char* tag;            // profiler tag (dx_profile_report)
};
int tc_gemm(const TcGemmArgs& g, const Rows& R, cudaStream_t s);
// dx_tc_linear.cu: weight-resident variant for K = 128, taps = 1, N <= 384 (QKV / out-proj + LN / mel projection)
int tc_linear(const TcGemmArgs& g, const Rows& R, cudaStream_t s);
int tc_gemm_plain(const TcGemmArgs& g, int64_t M, cudaStream_t s);   // no row map: rows [0, M) all live
int pack_weight_bf16(const float* w, int taps, int K, int N, int nt, int ks, __nv_bfloat16* out, cudaStream_t s);
struct PackEntry {              // one weight of pack_weight_batch; (taps, K, N) are the GEMM's
    const float* src;
    __nv_bfloat16* dst;
    int64_t n_elems;
    int taps, K, N, nt, ks, transposed;
};
int pack_weight_batch(const PackEntry* tab_dev, int n_entries, cudaStream_t s);
int64_t packed_weight_elems(int taps, int K, int N, int nt, int ks);

}  // namespace dx
