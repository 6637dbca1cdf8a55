/*
 * daft_b200.h — C ABI of libdaft_b200.so: Daft-Exprt's batched synthesis forward pass on B200 (sm_100a).
 *
 * The reference (claussss/ubisoft-laforge-daft-exprt) has no FFI/plugin interface: its boundary is the
 * Python class daft_exprt.model.DaftExprt (src/daft_exprt/model.py:833-1114).  This header is the
 * boundary a binding for that class calls into; the Python mirror of the class lives in
 * ubisoft-laforge-daft-exprt_b200/model.py and INTEGRATION.md shows the ctypes stub a maintainer of the
 * reference would add.  Every entry point cites the reference code it replaces.
 *
 * Conventions
 *   - plain C: device pointers, ints, a cudaStream_t passed as void*; no torch types, no C++ exceptions.
 *   - every function returns 0 on success, non-zero on error; dx_last_error_string() explains.
 *   - no allocation inside: callers pass outputs and scratch.
 *   - all pointers are DEVICE pointers unless the name ends in _host.
 *   - "rows": activations live in a packed row stream (see DESIGN.md "data layout"): utterance b owns rows
 *     row_start[b] .. row_start[b]+len[b]+DX_GAP-1; the last DX_GAP rows of each utterance are gap rows.
 *
 * Two arithmetic modes (BASELINE.json north_star): DX_MODE_F32 = fp32 SIMT kernels (parity gate: mel max-abs
 * <= 1e-3 vs the reference CPU forward) and DX_MODE_BF16 = bf16 operands on tcgen05 tensor cores with fp32
 * accumulation, LayerNorm, softmax and Gaussian weights (parity gate: relative L1 <= 2e-2).
 */
#ifndef DAFT_B200_H
#define DAFT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DX_GAP 3            /* gap rows after every utterance in the packed row stream (H1 ghost rows) */
#define DX_HALO 8           /* zero rows in front of every bf16 channel-interleaved buffer (conv tap -1)  */
#define DX_TAIL 264         /* zero rows behind the row range (tile overruns of the last utterance)           */
#define DX_D 128            /* hidden_embed_dim  (hparams.py:108)                                         */
#define DX_HEADS 2          /* attn_nb_heads     (hparams.py:109)                                         */
#define DX_DH 64            /* head dim                                                                   */
#define DX_C 1024           /* conv_channels     (hparams.py:112)                                         */
#define DX_TAPS 3           /* conv_kernel       (hparams.py:111)                                         */
#define DX_MEL_PAD 96       /* mel channels padded to a multiple of the K stage (bf16 tensor-core path)      */
#define DX_MEL 80           /* n_mel_channels    (hparams.py:44)                                          */
#define DX_SPK 192          /* ECAPA embedding size (model.py:855)                                        */
#define DX_MODE_F32 0
#define DX_MODE_BF16 1

/* ---- library ------------------------------------------------------------------------------------------- */
int         dx_version(void);
const char* dx_last_error_string(void);
/* number of kernels launched by this library since load (bench.py's "gpu_launches") */
int64_t     dx_launch_count(void);
/* bring-up switches (all 0 in production).  key 0: swap LBO/SBO in the UMMA shared-memory descriptors;
 * key 1: same for the MN-major V operand of attention; key 2: synchronise after every launch (fault location);
 * key 3: run the FFN as two GEMM launches instead of the fused kernel (A/B testing);
 * key 4: run the K = 128 linear layers in the generic conv-GEMM kernel instead of the weight-resident one;
 * key 5: launch without programmatic dependent launch (plain stream order between kernels);
 * key 6: fused block kernel: never split a tile's hidden dimension across CTAs (whole tiles per CTA);
 * key 7: its fewest hidden chunks per CTA when it splits (default 0 = 2 or 4 by the stream's size);
 * key 8: it splits when whole tiles would occupy at most this percentage of the grid (default 0 = 35; 1000 = always);
 * key 15: its timing experiments (wrong results; bit 0 no E2 stores, 1 no P3 stores, 2 no P0 residual, 3 no FiLM);
 * key 16 / 17: training step in tensor-core mode: keep the fp32 attention backward / forward (A/B);
 * key 18: its weight-gradient kernel's CTA budget per launch (default 0 = 148); key 19: pack each weight where it is used
 *         instead of all of them in one launch at the start of the step (A/B);
 * key 9: run every FFT block as four launches (QKV, attention, out-proj + LN, FFN) instead of two;
 * key 10: with a trace buffer set, trace only the n-th block-tail launch of a step (1..12; 0 = every launch);
 *         + 100: trace the last working CTA (always a tile owner) instead of CTA 0. */
int         dx_set_debug(int key, int value);

/* Bring-up: while a device buffer (>= 2 KB) is set, the fused FFN kernel writes a clock64 timeline of its CTA 0
 * into it (tools/ffn_trace.py decodes it).  NULL switches it off (the default). */
int         dx_set_trace_buffer(void* device_ptr, int64_t bytes);

/* Per-launch timing for bench.py's roofline: while enabled, the library records a CUDA event pair around every
 * launch of its main kernel families on the launching stream.  dx_profile_report synchronises the device and
 * writes one line per tag: "<tag> <launches> <total_ms>\n" (NUL-terminated), then clears the records. */
int dx_profile_enable(int on);
int dx_profile_report(char* buf, int64_t buf_bytes);

/* ---- row stream (replaces get_mask_from_lengths, model.py:14-24, and the padded layouts) --------------- */
/* lengths: int64[B].  n_max: width of the caller's padded tensor.  Writes row_start int32[B+1] (row_start[B] =
 * total rows), row_b int32[row_cap], row_p int32[row_cap] (utterance / position of every row; -1 / 0 beyond
 * the total).  row_cap must be >= dx_row_cap(B, n_max). */
int64_t dx_row_cap(int B, int n_max);
int dx_build_rows(const int64_t* lengths, int B, int n_max, int32_t* row_start, int32_t* row_b,
                  int32_t* row_p, int64_t row_cap, void* stream);

/* ---- weights -------------------------------------------------------------------------------------------- */
/* One FFT block (model.py:238-259).  fp32 tensors are in GEMM layout (prepared once by the host mirror):
 *   w_qkv  [128][384]  = in_proj_weight^T           b_qkv [384]
 *   w_out  [128][128]  = out_proj.weight^T          b_out [128]
 *   w_c1   [3][128][1024] = convs.0.conv.weight permuted (tap, cin, cout)     b_c1 [1024]
 *   w_c2   [3][1024][128] = convs.2.conv.weight permuted                      b_c2 [128]
 * bf16 tensors (tensor-core mode) are in the tile-packed layout dx_pack_weight_bf16 produces. */
typedef struct {
    const float *w_qkv, *b_qkv, *w_out, *b_out, *ln1_g, *ln1_b;
    const float *w_c1, *b_c1, *w_c2, *b_c2, *ln2_g, *ln2_b;
    const void  *w_qkv_bf, *w_out_bf, *w_c1_bf, *w_c2_bf;
    /* fused FFN kernel: conv1 packed with (nt 128, ks 64), conv2 packed with (nt 128, ks 64) */
    const void  *w_ffn1_bf, *w_ffn2_bf;
} DxBlockWeights;

typedef struct {
    /* accent encoder (model.py:614-716) */
    const float *ae_energy_w, *ae_energy_b, *ae_pitch_w, *ae_pitch_b;     /* [128][3], [128] */
    const float *ae_c0_w, *ae_c0_b, *ae_ln0_g, *ae_ln0_b;                 /* [3][80][1024]   */
    const float *ae_c1_w, *ae_c1_b, *ae_ln1_g, *ae_ln1_b;                 /* [3][1024][1024] */
    const float *ae_c2_w, *ae_c2_b, *ae_ln2_g, *ae_ln2_b;                 /* [3][1024][128]  */
    const void  *ae_c0_w_bf, *ae_c1_w_bf, *ae_c2_w_bf;
    DxBlockWeights ae_blocks[4];
    /* speaker projection + classifier + style adapter (model.py:719-830, 855-856) */
    const float *spk_w, *spk_b;                                           /* [128][192] (torch layout) */
    const float *cls_w1, *cls_b1, *cls_w2, *cls_b2, *cls_w3, *cls_b3;     /* torch layouts */
    const float *film_g_w, *film_g_b, *film_b_w, *film_b_b;               /* [1024][128], [1024] (default block counts) */
    const float *post_mult;                                               /* [2][8] (default) or NULL (== 1.0) */
    /* phoneme encoder (model.py:567-610) */
    const float *symbols_embedding;                                       /* [n_symbols][128] */
    DxBlockWeights pe_blocks[4];
    /* gaussian upsampling (model.py:385-510) */
    const float *gu_dur_w, *gu_dur_b, *gu_energy_w, *gu_energy_b, *gu_pitch_w, *gu_pitch_b;  /* [128][3],[128] */
    const float *gu_range_w, *gu_range_b;                                 /* [128], [1] */
    /* frame decoder (model.py:513-564) */
    DxBlockWeights fd_blocks[4];
    const float *mel_w, *mel_b;                                           /* [128][80] = projection^T, [80] */
    const void  *mel_w_bf;
    /* shared */
    const float *pos_table;                                               /* [5000][128], model.py:126-130 */
    int32_t n_symbols, n_speakers, pos_rows, pad_;
    /* FFT blocks per stack, 1..4 each (hparams.{accent_encoder, phoneme_encoder, frame_decoder}['nb_blocks']); 0 = 4, the
     * reference's default.  With P = n_pe_blocks and F = n_fd_blocks the style adapter predicts (P + F) * 128 gammas and
     * betas (film_g_w / film_b_w are [(P + F) * 128][128], model.py:731-747), post_mult is [2][P + F], and block i of the
     * decoder uses FiLM row P + i of the (B, 8, 256) film output.  dx_forward and dx_train_step both follow the counts. */
    int32_t n_ae_blocks, n_pe_blocks, n_fd_blocks, pad2_;
} DxModelWeights;

/* Pack an fp32 GEMM-layout weight [taps][K][N] into the bf16 tile layout the tensor-core kernels stream with
 * one bulk copy per stage:  [n_tile][k_stage][tap][ks/8][nt][8]  (N zero-padded to a multiple of nt). */
int64_t dx_packed_weight_elems(int taps, int K, int N, int nt, int ks);
int dx_pack_weight_bf16(const float* w, int taps, int K, int N, int nt, int ks, void* out, void* stream);

/* ---- whole passes (replace DaftExprt.forward model.py:889-948 / .inference model.py:1026-1114) ---------- */
typedef struct {
    int32_t B, L_max, T_max, mode;
    /* inputs, reference layouts (model.py:893-896) */
    const int64_t *symbols;            /* (B, L_max) */
    const float   *durations_float;    /* (B, L_max) seconds */
    const int64_t *durations_int;      /* (B, L_max) frames  */
    const float   *symbols_energy, *symbols_pitch;   /* (B, L_max) */
    const int64_t *input_lengths;      /* (B,) */
    const float   *frames_energy, *frames_pitch;     /* (B, T_max) or NULL when accent_emb_in is given */
    const float   *mel_specs;          /* (B, 80, T_max) or NULL */
    const int64_t *output_lengths;     /* (B,)  forward: given; inference: NULL → computed = max(sum dur,1) */
    const float   *spk_embs;           /* (B, 192) raw ECAPA, or NULL when spk_emb_in is given */
    const float   *accent_emb_in;      /* (B, 128) external accent embedding or NULL (run the accent encoder) */
    const float   *spk_emb_in;         /* (B, 128) external projected speaker embedding or NULL */
    /* outputs */
    float   *accent_emb;               /* (B, 128)      */
    float   *speaker_preds;            /* (B, n_speakers) or NULL */
    float   *film;                     /* (B, 8, 256): rows 0..P-1 phoneme encoder, P..P+F-1 frame decoder (P = F = 4 by default) */
    float   *enc_out;                  /* (B, L_max, 128) phoneme-encoder output or NULL (debug/tests) */
    float   *mel_out;                  /* (B, 80, T_max) */
    float   *weights_out;              /* (B, L_max, T_max) alignments */
    int64_t *output_lengths_out;       /* (B,) */
    /* scratch: dx_forward_workspace_bytes(B, L_max, T_max, mode) bytes, 256-byte aligned */
    void    *workspace;
    int64_t  workspace_bytes;
    /* DX_IO_* bits; 0 = the reference's padded layouts everywhere */
    int32_t  io_flags, pad_;
} DxForwardArgs;

/* Packed frame-side I/O (bulk generation, fine_tune.py:62-129: only valid frames cross PCIe).  With
 * fs[b] = sum of output_lengths[0..b-1] and T_b = output_lengths[b]:
 *   DX_IO_PACKED_FRAMES  frames_energy / frames_pitch are float[sum T_b] (utterance b at fs[b]) and mel_specs is
 *                        float[80 * sum T_b]: utterance b is the (80, T_b) row-major block at 80 * fs[b] - the layout of
 *                        the per-utterance .npy files the reference's loader reads (data_loader.py:150-160).  Frames
 *                        beyond T_b read as zeros, which is what the reference's collate pads with (data_loader.py:246-259).
 *   DX_IO_PACKED_MEL_OUT mel_out is float[80 * sum T_b] in the same per-utterance (80, T_b) blocks - exactly the
 *                        crop fine_tune.py:79 takes from the padded output; nothing is written for padded frames. */
#define DX_IO_PACKED_FRAMES 1
#define DX_IO_PACKED_MEL_OUT 2

int64_t dx_forward_workspace_bytes(int B, int L_max, int T_max, int mode);
int dx_forward(const DxModelWeights* w, const DxForwardArgs* a, void* stream);

/* ---- training step (BASELINE.json config 5) ------------------------------------------------------------------
 * One rank's share of the reference's training iteration (train.py:421-445): `outputs = model(inputs)` (model.py:889-948;
 * dropout = identity unless the attn_dropout / conv_dropout fields are set), `loss, individual = criterion(outputs, targets,
 * iteration)` (loss.py:57-159, pitch-consistency term
 * off), `loss.backward()`.  `grads` is a DxModelWeights whose fp32 pointers say where the gradient of each weight is
 * written, in the same ENGINE layout as the weight (the *_bf fields are ignored).  Every gradient is overwritten (no
 * accumulation); the kernels are deterministic.  The data-parallel exchange (DDP's gradient averaging, train.py:269-272) is
 * one all-reduce of the caller's flat gradient buffer between dx_train_step and dx_adam_step.  fp32 arithmetic. */
#define DX_TRAIN_FORWARD_ONLY 1    /* flags: stop after the loss (validation: train.py:155-209) */
#define DX_TRAIN_TC_GEMM 2         /* flags: run the large dense contractions (forward convs / linears, their dgrad and wgrad)
                                      on tcgen05 with bf16 copies of the fp32 operands and fp32 accumulation; everything else
                                      (LayerNorm, attention, length regulator, loss, reductions, Adam) stays fp32 */
typedef struct {
    int32_t B, L_max, T_max, flags;
    /* DaftExprt.forward inputs (model.py:893-896), reference layouts */
    const int64_t *symbols;  const float *durations_float;  const int64_t *durations_int;
    const float   *symbols_energy, *symbols_pitch;  const int64_t *input_lengths;
    const float   *frames_energy, *frames_pitch, *mel_specs;  const int64_t *output_lengths;
    const int64_t *speaker_ids;  const float *spk_embs;
    /* DaftExprtLoss (loss.py:22-55): speaker_weight = update_adversarial_weight(iteration) (:51-55), computed by the host;
     * loss_scale = 1 / accumulation_steps (train.py:423) */
    float mel_spec_weight, post_mult_weight, energy_consistency_weight, speaker_weight, lambda_reversal, loss_scale;
    /* outputs */
    float *losses;        /* [8]: loss * loss_scale, speaker_loss, speaker_ce_raw, post_mult_loss, mel_spec_l1_loss,
                                  mel_spec_l2_loss, energy_consistency_loss, pitch_consistency_loss (= 0)  (loss.py:148-157) */
    float *mel_out;       /* (B, 80, T_max) mel prediction or NULL */
    void  *workspace;     /* dx_train_workspace_bytes(B, L_max, T_max) bytes, 256-byte aligned */
    int64_t workspace_bytes;
    /* train-mode forward (model.train(), train.py:411): dropout probabilities per stack - [0] accent encoder (its pre-net
     * uses conv_dropout[0], model.py:656-670), [1] phoneme encoder, [2] frame decoder (hparams.py:105-127; attn_dropout is
     * both nn.MultiheadAttention's weight dropout, model.py:165-167, and the nn.Dropout behind it, :168,189; conv_dropout
     * is the one that ends PositionWiseConvFF.convs, :216).  All zero = eval-mode forward (the parity configuration).  Any
     * non-zero value needs DX_TRAIN_TC_GEMM.  Masks are a pure function of (dropout_seed, site, position): pass a fresh
     * seed per step; dx_dropout_mask returns the mask of a site for checking against the reference. */
    float attn_dropout[3], conv_dropout[3];
    uint64_t dropout_seed;
} DxTrainArgs;

int64_t dx_train_workspace_bytes(int B, int L_max, int T_max);
/* Keep-mask (uint8, 1 = kept) of one dropout site of dx_train_step, in the reference's padded coordinates.  kind 0: the
 * nn.Dropout behind the attention (model.py:189), 1: the one ending the conv feed-forward (:216), 2: layer `block` (0..2)
 * of the accent encoder's pre-net (:656,663,670; stack must be 0) - out is (B, n_max, C); kind 3: the attention-weight
 * dropout inside nn.MultiheadAttention - out is (B, 2 heads, n_max, n_max), C ignored.  stack: 0 accent encoder,
 * 1 phoneme encoder, 2 frame decoder; block 0..3; p: the site's probability; n_max: T_max (stacks 0, 2) or L_max (1). */
int dx_dropout_mask(uint64_t seed, int kind, int stack, int block, float p, int B, int n_max, int C, uint8_t* out,
                    void* stream);
int dx_train_step(const DxModelWeights* w, const DxModelWeights* grads, const DxTrainArgs* a, void* stream);
/* torch.optim.Adam(betas, eps, weight_decay, amsgrad=False) (train.py:278-280) on flat fp32 buffers of n elements:
 * g = grad * grad_scale + weight_decay * p; m, v moments; step counts from 1.  grad_scale = 1 / world_size after a SUM
 * all-reduce. */
int dx_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1,
                 float beta2, float eps, float weight_decay, int step, float grad_scale, void* stream);

/* ---- single stages (unit-test surface; same kernels dx_forward launches) -------------------------------- */
/* AccentEncoder.forward, model.py:679-716 */
int dx_accent_encoder(const DxModelWeights* w, int mode, int B, int T_max, const float* frames_energy,
                      const float* frames_pitch, const float* mel_specs, const int64_t* output_lengths,
                      float* accent_emb, void* workspace, int64_t workspace_bytes, void* stream);
/* speaker projection + classifier + StyleAdapter.forward, model.py:904-920, 764-806, 829-830 */
int dx_style(const DxModelWeights* w, int B, const float* spk_embs, const float* spk_emb_in,
             const float* accent_emb, float* film, float* speaker_preds, void* stream);
/* integer side of the length regulator, model.py:485-497, 1096-1100 (bit-exact contract, int64) */
int dx_length_regulator_integers(const int64_t* durations_int, const int64_t* input_lengths, int B, int L_max,
                                 int T_max, int64_t* cumsum, int64_t* two_mu, int64_t* output_lengths,
                                 int64_t* frame2symbol, void* stream);
/* get_int_durations + duration_to_integer, model.py:950-973, extract_features.py:69-125.  One thread per
 * utterance, IEEE double arithmetic in the reference's order.  duration (B,L) fp32 is thresholded in place.
 * status int32[B]: 0 ok, 1 = reference would raise (count mismatch / empty pop). */
int dx_int_durations(float* duration, int B, int L_max, int sampling_rate, int filter_length, int hop_length,
                     int centered, int64_t* durations_int, int32_t* status, void* stream);
/* pitch_shift / pitch_multiply + the zero-duration masking, model.py:1077-1087, 975-1024 (in place) */
int dx_prosody_transform(float* energy, float* pitch, const float* energy_factors, const float* pitch_factors,
                         const int64_t* durations_int, const float* pitch_mean, const float* pitch_std,
                         int B, int L_max, int transform /*0=add,1=multiply*/, void* stream);

/* ---- tensor-core self tests (GPU unit tests of the tcgen05 building blocks) ----------------------------- */
/* out[M][N] fp32 = sum_taps A[row+tap-(taps/2)] * W[tap]  with A bf16 channel-interleaved, W packed bf16. */
int dx_tc_gemm_test(const void* a_il, int64_t a_rows_phys, int M, int K, int N, int taps, const void* w_packed,
                    int nt, int ks, float* out, void* stream);

/* Masked multi-head self-attention alone (MultiHeadAttention.forward's nn.MultiheadAttention call, model.py:182-186).
 * mode F32: qkv fp32 [row_cap][384] → out fp32 [row_cap][128].  mode BF16: qkv bf16 channel-interleaved
 * [48][rows_phys][8] → out [16][rows_phys][8].  rowmap_scratch: int32[(B + 1) + 2 * row_cap]. */
int dx_test_attention(int mode, const void* qkv, void* out, int64_t rows_phys, const int64_t* lengths, int B, int n_max,
                      int32_t* rowmap_scratch, int64_t row_cap, void* stream);

/* Allocation-free variant of dx_test_attention.  It has NO counterpart in the reference: it was added
 * for callers that cannot allocate device memory (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  Same kernels, same arguments and results, but the int32 row map lives in a
 * file-scope __device__ array inside the library (257 KB, held by every GPU that loads the library until exit)
 * instead of caller scratch; the call performs no cudaMalloc / cudaFree / allocating library routine.
 * Limits: 1 <= B <= 255 and dx_row_cap(B, n_max) <= DX_ATTENTION_STATIC_MAX_ROWS, else status 2 and a message.
 * All calls on one device share the array (it keeps the last call's contents; every entry a call reads is rewritten
 * by that call first, nothing needs to start at zero): calls that overlap on the GPU race — serialise them. */
#define DX_ATTENTION_STATIC_MAX_ROWS 32768
int64_t dx_attention_static_max_rows(void);
int dx_attention_static(int mode, const void* qkv, void* out, int64_t rows_phys, const int64_t* lengths, int B, int n_max,
                        void* stream);

/* ---- HiFi-GAN generator, batched -------------------------------------------------
 * Replaces HiFiGANGenerator.forward (vocoder/hifigan.py:159-176) with ResBlock1 (hifigan.py:77-97) for a whole batch
 * of utterances at once, each vocoded as if alone on its cropped mel (what HiFiGanVocoder.infer does one utterance
 * at a time, hifigan.py:211-226).  Utterances are packed into one frame stream separated by DX_VC_GAP zero frames;
 * at stage s every activation row is (frame, phase) at rate prod(up[0..s]).  Every convolution is one implicit-GEMM
 * launch: out[r] = sum_tap A[r + (tap - taps/2) * dil] · W[tap]; transposed convolutions (kernel = 2 · stride) are
 * their 3-tap polyphase form at the input rate with N = stride · C_out.  DX_MODE_F32: fp32 SIMT kernels;
 * DX_MODE_BF16: bf16 operands on tcgen05 with fp32 accumulation; the residual stream x is bf16 (default) or fp32
 * (DX_VC_RESIDUAL_F32), the ResBlock sums xs are always fp32. */
#define DX_VC_GAP 4          /* zero frames in front of, between and behind utterances in the packed frame stream  */
#define DX_VC_HALO 32        /* rows in front of / behind every bf16 activation buffer (>= (taps/2) * dil)          */
#define DX_VC_RESIDUAL_F32 1  /* DX_MODE_BF16 only: keep the residual stream x in fp32 instead of bf16 (rel. L1 1.06e-2
                                instead of 1.22e-2 against the reference on the golden cases, 8 % slower)           */
#define DX_VC_MAX_STAGES 4
#define DX_VC_MAX_KERNELS 3
#define DX_VC_MAX_LAYERS 3

typedef struct {
    const float* w;          /* fp32 [taps][K][N]                                                                  */
    const void*  w_bf;       /* bf16 packed by dx_pack_weight_bf16 with the (nt, ks) of dx_vocoder_conv_tiling, or NULL */
    const float* bias;       /* [C_out]                                                                            */
    int32_t taps, dil, K, N; /* N = stride * C_out for the upsamplers                                              */
} DxVcConv;

typedef struct {
    int32_t n_stages, n_kernels, n_layers, mel_channels;      /* 4, 3, 3, 80 in the reference's DEFAULT_CONFIG     */
    int32_t up[DX_VC_MAX_STAGES];                             /* upsample_rates   (hifigan.py:30)                  */
    int32_t channels[DX_VC_MAX_STAGES];                       /* C_out of stage s (hifigan.py:141)                 */
    DxVcConv pre;                                             /* conv_pre   (hifigan.py:131)                       */
    DxVcConv ups[DX_VC_MAX_STAGES];                           /* ups[s]     (hifigan.py:133-143), polyphase form   */
    DxVcConv rb[DX_VC_MAX_STAGES][DX_VC_MAX_KERNELS][DX_VC_MAX_LAYERS][2];   /* convs1[l], convs2[l] (hifigan.py:80-87) */
    DxVcConv post;                                            /* conv_post  (hifigan.py:152): w fp32 [taps][K], N = 1 */
} DxVocoderWeights;

typedef struct {
    int32_t B, T_max, mode, flags;     /* flags: DX_VC_RESIDUAL_F32 or 0                                              */
    int64_t frame_cap;       /* >= dx_vocoder_frame_cap(B, sum of lengths)                                         */
    const float* mel;        /* (B, mel_channels, T_max) fp32                                                      */
    const int64_t* lengths;  /* (B,) frames per utterance, each <= T_max                                           */
    float* audio;            /* (B, T_max * hop) fp32, zeros behind every utterance                                */
    void* workspace;
    int64_t workspace_bytes;
} DxVocoderArgs;

/* rows of the packed frame stream needed for `total_frames` valid frames in B utterances (multiple of 128) */
int64_t dx_vocoder_frame_cap(int B, int64_t total_frames);
/* tile width nt and channels per pipeline stage ks the tensor-core conv kernel uses for a (taps, dil, K, N) conv */
int dx_vocoder_conv_tiling(int taps, int dil, int K, int N, int* nt, int* ks);
int64_t dx_vocoder_workspace_bytes(const DxVocoderWeights* w, int B, int64_t frame_cap, int mode);
int dx_vocoder_forward(const DxVocoderWeights* w, const DxVocoderArgs* a, void* stream);

/* ---- host-side batch collation for bulk generation ---------------------------------
 * The padded-batch assembly of DaftExprtDataCollate.__call__ (data_loader.py:214-263) as memcpy / memset into ONE
 * (pinned) staging block, outside the Python GIL and on `threads` host threads.  All pointers are HOST pointers.
 * ptrs_host[b * 10 + f]: field f of utterance b — 0 symbols i64 (L_b), 1 durations_float f32 (L_b), 2 durations_int
 * i64 (L_b), 3 symbols_energy f32 (L_b), 4 symbols_pitch f32 (L_b), 5 frames_energy f32 (T_b), 6 frames_pitch f32
 * (T_b), 7 mel f32 (M, T_b) row-major, 8 speaker_id i64 (1), 9 spk_emb f32 (E).  lens_host[2 b] = L_b, [2 b + 1] = T_b.
 * offs_host[12]: byte offsets inside the block of the 12 tensors of DaftExprt.forward (model.py:893-896 order), each
 * padded to (B, L) / (B, T) / (B, M, T) / (B,) / (B, E); pad regions are zeroed. */
int dx_host_collate(int B, int L, int T, int M, int E, const void* const* ptrs_host, const int32_t* lens_host,
                    void* block_host, const int64_t* offs_host, int threads);
/* Same, with the frame side in the DX_IO_PACKED_FRAMES layout: offs_host[6..8] are the offsets of float[sum T_b],
 * float[sum T_b] and float[M * sum T_b]; every utterance's mel is one memcpy of its (M, T_b) array, nothing is padded.
 * T is only the bound the lengths are checked against. */
int dx_host_collate_packed(int B, int L, int T, int M, int E, const void* const* ptrs_host, const int32_t* lens_host,
                           void* block_host, const int64_t* offs_host, int threads);

#ifdef __cplusplus
}
#endif
#endif /* DAFT_B200_H */
