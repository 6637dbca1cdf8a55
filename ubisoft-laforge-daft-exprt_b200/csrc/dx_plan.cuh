// Workspace layout and the whole-pass launch sequences.
#pragma once
#include "dx_common.cuh"
#include "dx_kernels.cuh"

namespace dx {

int64_t row_cap(int B, int n_max);

// All scratch of one forward pass, carved from the caller's workspace (256-byte aligned slices).
struct Plan {
    int B, L_max, T_max, mode;
    int64_t cap_t, cap_l, cap;          // row capacities: frame stream, symbol stream, max of both
    int64_t rows_phys;                  // DX_HALO + cap + DX_TAIL (bf16 channel-interleaved buffers)
    int64_t bytes;
    // row maps
    int32_t *t_row_start, *t_row_b, *t_row_p, *l_row_start, *l_row_b, *l_row_p;
    // small per-utterance buffers
    float *accent, *sigma, *mu;
    int64_t* out_len;
    float* row_stats;                   // [cap][TC_STAT_TILES][2]
    // fp32 mode activations (row-major [cap][C])
    float *x, *y, *o, *tmp, *qkv, *h1, *h2, *melp;
    // length regulator
    float* xs;                          // [cap_l][128] fp32 (both modes)
    // bf16 mode activations (channel-interleaved [C/8][rows_phys][8])
    __nv_bfloat16 *xb, *yb, *ob, *qkvb, *h1b, *h2b, *melb;
    float* melo;                        // [cap_t][80] fp32 mel projection rows (both modes)
    // fused FFN kernel, bf16 mode: partial sums + flags of the hidden-dimension split (dx_ffn_tc.cu)
    float* ffn_partial;
    uint32_t* ffn_flags;
    FfnSplit ffn_split() const { return FfnSplit{ffn_partial, ffn_flags}; }

    Rows rows_t(const int64_t* lengths) const {
        return Rows{t_row_start, t_row_b, t_row_p, lengths, B, T_max, cap_t};
    }
    Rows rows_l(const int64_t* lengths) const {
        return Rows{l_row_start, l_row_b, l_row_p, lengths, B, L_max, cap_l};
    }
};

// workspace == nullptr: only compute sizes.
int plan_layout(Plan* p, int B, int L_max, int T_max, int mode, void* workspace, int64_t workspace_bytes);

// bf16 mode, once per pass before the first kernel: zero rows around the row range of every channel-interleaved buffer,
// cleared split flags of the fused FFN kernel
int zero_plan_halos(const Plan& p, cudaStream_t s);
int run_accent_encoder(const DxModelWeights& w, const Plan& p, const Rows& RT, const float* frames_energy,
                       const float* frames_pitch, const float* mel_specs, float* accent_emb, cudaStream_t s,
                       bool packed_frames = false);
int run_forward(const DxModelWeights& w, const DxForwardArgs& a, cudaStream_t s);

}  // namespace dx
