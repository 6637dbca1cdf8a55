// Shared by dx_vocoder.cu (plan, frame map, mel packing, conv_post) and dx_vocoder_conv.cu (the convolution kernels):
// the fused epilogue of every HiFi-GAN convolution.  See dx_vocoder.cu for the layout and the reference citations.
#pragma once
#include "dx_common.cuh"
#include "dx_ptx.cuh"

namespace dx {

constexpr float VC_SLOPE = 0.1f;     // LRELU_SLOPE, hifigan.py:21
constexpr int VC_BM = 128;           // rows per tile (UMMA M)
constexpr int VC_THREADS = 192;      // producer warp, issuer warp, 4 epilogue warps
constexpr int VC_MAX_N = 2048;       // widest GEMM (first upsampler: 8 * 256)

struct VcEpi {
    const float* bias;               // [c_out]
    int c_out, up;                   // column n = ph * c_out + co → output row r * up + ph
    const int32_t* frame_utt;
    int out_shift;                   // log2(rate of the output rows)
    // fp32 residual stream, "quad-interleaved" [c_out/4][x_rows][4]: lane ↔ row accesses are 16 B at consecutive
    // addresses (one coalesced 512 B request per warp and 4-channel group)
    const float* res;                // or null
    float* x_out;                    // or null
    float* xs;                       // (xs_mode != 0)
    int64_t x_rows;                  // rows of the output-rate tensors
    int x_bf16;                      // res / x_out are bf16 [c_out/8][x_rows][8] (bf16 mode default) instead of fp32
    int xs_mode;                     // 0 none, 1 xs = y, 2 xs += y, 3 y = (xs + y) / n_kernels
    float n_kernels;
    int act;                         // 1: leaky_relu on act_out, 0: identity
    void* act_out;                   // bf16 interleaved / fp32 [rows_out][c_out], or null
    int64_t act_rows_phys;           // bf16 interleaved only
};

__device__ __forceinline__ float vc_lrelu(float v) { return v > 0.f ? v : v * VC_SLOPE; }

// NV consecutive columns n .. n+NV-1 (one phase) of input-rate row r.  BF: act_out is bf16 interleaved (NV % 8 == 0).
template <int NV, bool BF>
__device__ __forceinline__ void vc_epilogue(const VcEpi& e, int64_t r, int n, float* v, const float* s_bias) {
    const int ph = n / e.c_out, co = n - ph * e.c_out;
    const int64_t ro = r * e.up + ph;
    const bool valid = e.frame_utt[ro >> e.out_shift] >= 0;
    if (valid) {
        const int64_t off = ((int64_t)(co >> 2) * e.x_rows + ro) * 4;   // 4-channel group co/4 of row ro
        const int64_t gstride = e.x_rows * 4 - 4;                        // off + i + (i/4) * gstride = group (co + i)/4
#pragma unroll
        for (int i = 0; i < NV; i += 4) {
            float4 y = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
            const float* bb = s_bias ? s_bias + n + i : e.bias + co + i;
            y.x += bb[0]; y.y += bb[1]; y.z += bb[2]; y.w += bb[3];
            if (e.res) {
                const float4 rr = *reinterpret_cast<const float4*>(e.res + off + i + (i >> 2) * gstride);
                y.x += rr.x; y.y += rr.y; y.z += rr.z; y.w += rr.w;
            }
            if (e.xs_mode == 1) {
                *reinterpret_cast<float4*>(e.xs + off + i + (i >> 2) * gstride) = y;
            } else if (e.xs_mode >= 2) {
                const float4 a = *reinterpret_cast<const float4*>(e.xs + off + i + (i >> 2) * gstride);
                y.x += a.x; y.y += a.y; y.z += a.z; y.w += a.w;     // xs += resblock(x), hifigan.py:170
                if (e.xs_mode == 2) *reinterpret_cast<float4*>(e.xs + off + i + (i >> 2) * gstride) = y;
                else { y.x /= e.n_kernels; y.y /= e.n_kernels; y.z /= e.n_kernels; y.w /= e.n_kernels; }   // :171
            }
            if (e.x_out) *reinterpret_cast<float4*>(e.x_out + off + i + (i >> 2) * gstride) = y;
            v[i] = y.x; v[i + 1] = y.y; v[i + 2] = y.z; v[i + 3] = y.w;
        }
    }
    if (e.act_out == nullptr) return;
#pragma unroll
    for (int i = 0; i < NV; ++i) v[i] = valid ? (e.act ? vc_lrelu(v[i]) : v[i]) : 0.f;
    if constexpr (BF) {
        __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(e.act_out);
#pragma unroll
        for (int c = 0; c < NV / 8; ++c) {
            __align__(16) __nv_bfloat162 o2[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) o2[q] = __floats2bfloat162_rn(v[c * 8 + 2 * q], v[c * 8 + 2 * q + 1]);
            *reinterpret_cast<uint4*>(out + ((int64_t)(co / 8 + c) * e.act_rows_phys + DX_VC_HALO + ro) * 8) =
                *reinterpret_cast<const uint4*>(o2);
        }
    } else {
        float* out = reinterpret_cast<float*>(e.act_out) + ro * e.c_out + co;
#pragma unroll
        for (int i = 0; i < NV; i += 4) *reinterpret_cast<float4*>(out + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
    }
}


// One convolution launch (either mode).  a: conv input at the input rate; total rows = *total_frames << in_shift.
struct VcConvArgs {
    const void* a;                   // bf16 interleaved [K/8][a_rows_phys][8] or fp32 [rows][K]
    int64_t a_rows_phys;
    int K, N, taps, dil;
    const float* w_f32;              // [taps][K][N]
    const __nv_bfloat16* w_bf;       // packed [n_tile][k_stage][tap][ks/8][nt][8]
    const int32_t* total_frames;     // device: frame_start[B]
    int in_shift;                    // log2(rate of the input rows)
    int64_t rows_bound;              // host bound on the input rows (frame_cap << in_shift), sizes the grid
    VcEpi e;
    const char* tag;
};
int vc_conv_bf16(const VcConvArgs& c, cudaStream_t s);
int vc_conv_f32(const VcConvArgs& c, cudaStream_t s);
void vc_conv_tiling(int taps, int dil, int K, int N, int* nt, int* ks);

}  // namespace dx

namespace dx {
// One ResBlock1 layer (hifigan.py:92-96) in ONE launch: x_new = convs2(leaky_relu(convs1_dilated(a))) + x, with the
// hidden tile in shared memory (dx_vocoder_conv.cu, vc_pair_tc_kernel).  bf16 mode, C = K = N <= 128.
struct VcPairArgs {
    const __nv_bfloat16* a;          // leaky_relu(x), bf16 interleaved [C/8][a_rows_phys][8]
    int64_t a_rows_phys;
    int C, taps, dil;                // convs1: taps x dil; convs2: taps x 1
    const __nv_bfloat16 *w1_bf, *w2_bf;
    const float *bias1, *bias2;
    const int32_t* total_frames;
    int shift;                       // log2(rate)
    int64_t rows_bound;
    VcEpi e;                         // epilogue of convs2 (bias = bias2, up = 1)
    const char* tag;
};
int vc_pair_bf16(const VcPairArgs& c, cudaStream_t s);
}  // namespace dx
