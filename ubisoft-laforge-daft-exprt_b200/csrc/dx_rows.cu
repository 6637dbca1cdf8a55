// Row-stream plumbing: build the packed row map from lengths and move data between the reference's padded
// layouts ((B, C, T) / (B, L)) and the packed row stream.  All HBM-bound, coalesced kernels.
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_ptx.cuh"

namespace dx {

// ---------------------------------------------------------------------------------------------------------
// build_rows.  Replaces get_mask_from_lengths (model.py:14-24): every later kernel derives its pad
// mask from (row_b, row_p, len).  No host sync: the total row count stays on the device (row_start[B]).
// ---------------------------------------------------------------------------------------------------------
__global__ void build_rows_kernel(const int64_t* __restrict__ lengths, int B, int n_max,
                                  int32_t* __restrict__ row_start, int32_t* __restrict__ row_b,
                                  int32_t* __restrict__ row_p, int64_t cap) {
    ptx::griddep_wait();      // PDL: predecessor complete + visible
    ptx::griddep_launch();    // let the next kernel start its prologue
    extern __shared__ int32_t s_start[];  // B + 1
    // every CTA rebuilds the (tiny) prefix sum: parallel clamped loads, then one serial pass over shared memory
    for (int b = threadIdx.x; b < B; b += blockDim.x) {
        long long l = lengths[b];
        if (l < 0) l = 0;
        if (l > n_max) l = n_max;
        s_start[b + 1] = (int32_t)l + DX_GAP;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int32_t acc = 0;
        s_start[0] = 0;
        for (int b = 1; b <= B; ++b) {
            acc += s_start[b];
            s_start[b] = acc;
        }
    }
    __syncthreads();
    if (blockIdx.x == 0)
        for (int b = threadIdx.x; b <= B; b += blockDim.x) row_start[b] = s_start[b];
    const int32_t total = s_start[B];
    // fill (b, p) for every row: binary search of the owning utterance
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < cap; r += (int64_t)gridDim.x * blockDim.x) {
        if (r >= total) {
            row_b[r] = -1;
            row_p[r] = 0;
            continue;
        }
        int lo = 0, hi = B - 1;
        while (lo < hi) {
            int mid = (lo + hi + 1) >> 1;
            if (s_start[mid] <= r) lo = mid; else hi = mid - 1;
        }
        row_b[r] = lo;
        row_p[r] = (int32_t)(r - s_start[lo]);
    }
}

int build_rows(const int64_t* lengths, int B, int n_max, int32_t* row_start, int32_t* row_b, int32_t* row_p,
               int64_t cap, cudaStream_t s) {
    DX_REQUIRE(B >= 1 && B <= 8192, "build_rows: B=%d out of range", B);
    int grid = (int)((cap + 1023) / 1024);
    if (grid > 148) grid = 148;
    if (grid < 1) grid = 1;
    DX_LAUNCH_PDL((build_rows_kernel), grid, 256, (B + 1) * sizeof(int32_t), s, lengths, B, n_max, row_start, row_b, row_p, cap);
    DX_CHECK_LAUNCH();
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// pack_mel: (B, 80, T) fp32 → rows.  Row (b, p) gets mel[b, :, p] for p < T (gap rows copy the caller's
// padding content, like the reference's convs see it), 0 for p >= T.
// Tile transpose through smem: reads coalesced along T, writes coalesced along channels.
// OUT_BF16: write the bf16 channel-interleaved layout [C/8][rows_phys][8] instead of fp32 [rows][C].
// ---------------------------------------------------------------------------------------------------------
// Frame-side source of utterance b: element (channel c, frame p) at base[c * stride + p], valid for p < n.
// Padded (B, C, n_max) tensors: n = n_max (the caller's padding content is read like the reference's convs see it).
// Packed (DX_IO_PACKED_FRAMES): per-utterance (C, T_b) blocks at C * fs[b], fs[b] = row_start[b] - b * DX_GAP; frames
// beyond T_b read as zeros (= the reference collate's zero padding).
struct FrameView {
    int64_t base;
    int stride, n;
};
__device__ __forceinline__ FrameView frame_view(const Rows& R, int b, int C, bool packed) {
    if (!packed) return FrameView{(int64_t)b * C * R.n_max, R.n_max, R.n_max};
    const int tb = (int)min((long long)R.len[b], (long long)R.n_max);
    return FrameView{(int64_t)C * (R.row_start[b] - b * DX_GAP), tb, tb};
}

template <bool OUT_BF16>
__global__ void pack_mel_kernel(const float* __restrict__ mel, Rows R, int C, float* __restrict__ out_f32,
                                __nv_bfloat16* __restrict__ out_bf, int64_t rows_phys, bool packed) {
    ptx::griddep_wait();      // PDL: predecessor complete + visible
    ptx::griddep_launch();    // let the next kernel start its prologue
    __shared__ float tile[32][DX_MEL + 1];
    const int64_t r0 = (int64_t)blockIdx.x * 32;
    const int32_t total = R.row_start[R.B];
    if (r0 >= total) return;
    // load: thread (ty, tx): row r0+tx, channels ty, ty+8, ...
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 256 threads → ty 0..7
    {
        int64_t r = r0 + tx;
        int b = (r < total) ? R.row_b[r] : -1;
        int p = (b >= 0) ? R.row_p[r] : 0;
        FrameView fv = (b >= 0) ? frame_view(R, b, C, packed) : FrameView{0, 0, 0};
        bool ok = (b >= 0) && (p < fv.n);
        for (int c = ty; c < C; c += 8)
            tile[tx][c] = ok ? mel[fv.base + (int64_t)c * fv.stride + p] : 0.f;
    }
    __syncthreads();
    if (!OUT_BF16) {
        for (int i = threadIdx.x; i < 32 * C; i += blockDim.x) {
            int rr = i / C, c = i % C;
            if (r0 + rr < R.cap) out_f32[(r0 + rr) * C + c] = tile[rr][c];
        }
    } else {
        // one thread per (chunk, row): 16-byte store; consecutive threads → consecutive rows.  The bf16 buffer has
        // DX_MEL_PAD channels (K of the first pre-net GEMM must be a multiple of its K stage): extra ones are 0.
        const int nchunk = DX_MEL_PAD / 8;
        for (int i = threadIdx.x; i < 32 * nchunk; i += blockDim.x) {
            int rr = i & 31, ch = i >> 5;
            __align__(16) __nv_bfloat16 v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = __float2bfloat16(ch * 8 + e < C ? tile[rr][ch * 8 + e] : 0.f);
            int64_t off = ((int64_t)ch * rows_phys + DX_HALO + r0 + rr) * 8;
            *reinterpret_cast<uint4*>(out_bf + off) = *reinterpret_cast<const uint4*>(v);
        }
    }
}

int pack_mel(const float* mel, const Rows& R, float* out_f32, __nv_bfloat16* out_bf, int64_t rows_phys,
             cudaStream_t s, bool packed) {
    int grid = (int)(R.cap / 32);
    ProfScope prof("pack_mel", s);
    if (out_bf)
        DX_LAUNCH_PDL((pack_mel_kernel<true>), grid, 256, 0, s, mel, R, DX_MEL, nullptr, out_bf, rows_phys, packed);
    else
        DX_LAUNCH_PDL((pack_mel_kernel<false>), grid, 256, 0, s, mel, R, DX_MEL, out_f32, nullptr, 0, packed);
    DX_CHECK_LAUNCH();
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// unpack_mel: rows [*][80] fp32 → (B, 80, T) with zeros at t >= len (FrameDecoder mask + transpose,
// model.py:562-563).  Every element of the output is written.
// ---------------------------------------------------------------------------------------------------------
__global__ void unpack_mel_kernel(const float* __restrict__ rows, Rows R, int C, float* __restrict__ mel, bool packed) {
    ptx::griddep_wait();      // PDL: predecessor complete + visible
    ptx::griddep_launch();    // let the next kernel start its prologue
    __shared__ float tile[32][DX_MEL + 1];
    const int b = blockIdx.y;
    const int t0 = blockIdx.x * 32;
    const int len = (int)min((long long)R.len[b], (long long)R.n_max);
    const int64_t rs = R.row_start[b];
    for (int i = threadIdx.x; i < 32 * C; i += blockDim.x) {
        int tt = i / C, c = i % C;
        int t = t0 + tt;
        tile[tt][c] = (t < len) ? rows[(rs + t) * C + c] : 0.f;
    }
    __syncthreads();
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    int t = t0 + tx;
    const FrameView fv = frame_view(R, b, C, packed);      // packed: only the len valid frames exist in the output
    if (t < fv.n)
        for (int c = ty; c < C; c += 8) mel[fv.base + (int64_t)c * fv.stride + t] = tile[tx][c];
}

int unpack_mel(const float* rows, const Rows& R, float* mel, cudaStream_t s, bool packed) {
    dim3 grid((R.n_max + 31) / 32, R.B);
    ProfScope prof("unpack_mel", s);
    DX_LAUNCH_PDL((unpack_mel_kernel), grid, 256, 0, s, rows, R, DX_MEL, mel, packed);
    DX_CHECK_LAUNCH();
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// embed_symbols: x[r] = valid ? Emb[symbols[b,p]] + P[p] : 0   (PhonemeEncoder.forward, model.py:597-604 and
// PositionalEncoding.forward, model.py:132-150: the gather index for (b, p) is p).  One warp per row, float4.
// ---------------------------------------------------------------------------------------------------------
template <bool OUT_BF16>
__global__ void embed_symbols_kernel(const int64_t* __restrict__ symbols, const float* __restrict__ emb,
                                     int n_symbols, const float* __restrict__ pos, Rows R,
                                     float* __restrict__ out_f32, __nv_bfloat16* __restrict__ out_bf,
                                     int64_t rows_phys) {
    ptx::griddep_wait();      // PDL: predecessor complete + visible
    ptx::griddep_launch();    // let the next kernel start its prologue
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= R.cap) return;
    const int b = R.row_b[r];
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (b >= 0) {
        const int p = R.row_p[r];
        if (p < R.len[b]) {
            long long sym = symbols[(int64_t)b * R.n_max + p];
            sym = sym < 0 ? 0 : (sym >= n_symbols ? n_symbols - 1 : sym);
            float4 e = reinterpret_cast<const float4*>(emb + sym * DX_D)[lane];
            float4 q = reinterpret_cast<const float4*>(pos + (int64_t)p * DX_D)[lane];
            v = make_float4(e.x + q.x, e.y + q.y, e.z + q.z, e.w + q.w);
        }
    }
    if (!OUT_BF16) {
        reinterpret_cast<float4*>(out_f32 + r * DX_D)[lane] = v;
    } else {
        // lane holds channels 4*lane..4*lane+3 → half of chunk lane/2
        __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
        uint2 pk = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
        int ch = lane >> 1;
        int64_t off = ((int64_t)ch * rows_phys + DX_HALO + r) * 8 + (lane & 1) * 4;
        *reinterpret_cast<uint2*>(out_bf + off) = pk;
    }
}

int embed_symbols(const int64_t* symbols, const float* emb, int n_symbols, const float* pos, const Rows& R,
                  float* out_f32, __nv_bfloat16* out_bf, int64_t rows_phys, cudaStream_t s) {
    int grid = (int)((R.cap + 7) / 8);
    if (out_bf)
        DX_LAUNCH_PDL((embed_symbols_kernel<true>), grid, 256, 0, s, symbols, emb, n_symbols, pos, R, nullptr, out_bf, rows_phys);
    else
        DX_LAUNCH_PDL((embed_symbols_kernel<false>), grid, 256, 0, s, symbols, emb, n_symbols, pos, R, out_f32, nullptr, 0);
    DX_CHECK_LAUNCH();
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// scalar_conv3: y[c] = bias[c] + sum_j w[c][j] * x[b, p + j - 1]   (ConvNorm1D(1 → 128, k=3), model.py:640-646,
// 402-410) evaluated straight from the padded (B, n_max) input.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 scalar_conv3(const float* __restrict__ xr, int p, int n_max,
                                               const float* __restrict__ w, const float* __restrict__ bias,
                                               int lane) {
    float xm = (p - 1 >= 0) ? xr[p - 1] : 0.f;
    float x0 = (p < n_max) ? xr[p] : 0.f;
    float xp = (p + 1 < n_max) ? xr[p + 1] : 0.f;
    float o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int c = lane * 4 + i;
        // same tap order as a direct convolution: ((b + w0*x-1) + w1*x0) + w2*x+1 is not guaranteed by
        // the reference's MKL-DNN either; fp32 tolerance applies
        o[i] = bias[c] + w[c * 3 + 0] * xm + w[c * 3 + 1] * x0 + w[c * 3 + 2] * xp;
    }
    return make_float4(o[0], o[1], o[2], o[3]);
}

// accent_input: x = valid ? z2 + conv(energy) + conv(pitch) + P[p] : 0   (AccentEncoder.forward, model.py:687-706)
template <bool BF16>
__global__ void accent_input_kernel(const float* __restrict__ z2_f32, const __nv_bfloat16* __restrict__ z2_bf,
                                    const float* __restrict__ energy, const float* __restrict__ pitch,
                                    const float* __restrict__ ew, const float* __restrict__ eb,
                                    const float* __restrict__ pw, const float* __restrict__ pb,
                                    const float* __restrict__ pos, Rows R, float* __restrict__ out_f32,
                                    __nv_bfloat16* __restrict__ out_bf, int64_t rows_phys, bool packed) {
    ptx::griddep_wait();      // PDL: predecessor complete + visible
    ptx::griddep_launch();    // let the next kernel start its prologue
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (r >= R.cap) return;
    const int b = R.row_b[r];
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    const int ch = lane >> 1;
    const int64_t il_off = ((int64_t)ch * rows_phys + DX_HALO + r) * 8 + (lane & 1) * 4;
    if (b >= 0) {
        const int p = R.row_p[r];
        if (p < R.len[b]) {
            float4 z;
            if (!BF16) {
                z = reinterpret_cast<const float4*>(z2_f32 + r * DX_D)[lane];
            } else {
                uint2 pk = *reinterpret_cast<const uint2*>(z2_bf + il_off);
                __nv_bfloat162 lo = *reinterpret_cast<__nv_bfloat162*>(&pk.x);
                __nv_bfloat162 hi = *reinterpret_cast<__nv_bfloat162*>(&pk.y);
                z = make_float4(__low2float(lo), __high2float(lo), __low2float(hi), __high2float(hi));
            }
            const FrameView fv = frame_view(R, b, 1, packed);
            float4 e = scalar_conv3(energy + fv.base, p, fv.n, ew, eb, lane);
            float4 q = scalar_conv3(pitch + fv.base, p, fv.n, pw, pb, lane);
            float4 ps = reinterpret_cast<const float4*>(pos + (int64_t)p * DX_D)[lane];
            v = make_float4(((z.x + e.x) + q.x) + ps.x, ((z.y + e.y) + q.y) + ps.y, ((z.z + e.z) + q.z) + ps.z,
                            ((z.w + e.w) + q.w) + ps.w);
        }
    }
    if (!BF16) {
        reinterpret_cast<float4*>(out_f32 + r * DX_D)[lane] = v;
    } else {
        __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
        *reinterpret_cast<uint2*>(out_bf + il_off) =
            make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
    }
}

int accent_input(const float* z2_f32, const __nv_bfloat16* z2_bf, const float* energy, const float* pitch,
                 const float* ew, const float* eb, const float* pw, const float* pb, const float* pos,
                 const Rows& R, float* out_f32, __nv_bfloat16* out_bf, int64_t rows_phys, cudaStream_t s, bool packed) {
    int grid = (int)((R.cap + 7) / 8);
    ProfScope prof("accent_input", s);
    if (out_bf)
        DX_LAUNCH_PDL((accent_input_kernel<true>), grid, 256, 0, s, nullptr, z2_bf, energy, pitch, ew, eb, pw, pb, pos, R,
                                                       nullptr, out_bf, rows_phys, packed);
    else
        DX_LAUNCH_PDL((accent_input_kernel<false>), grid, 256, 0, s, z2_f32, nullptr, energy, pitch, ew, eb, pw, pb, pos, R,
                                                        out_f32, nullptr, 0, packed);
    DX_CHECK_LAUNCH();
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// masked_mean: accent[b] = sum_{p < len} x[b, p] / len   (model.py:714).  Grid (B, 16): one CTA per utterance and
// 8-channel chunk; 256 threads stride over the rows (16-byte / 32-byte coalesced loads), block reduction.
// ---------------------------------------------------------------------------------------------------------
template <bool BF16>
__global__ void __launch_bounds__(256)
masked_mean_kernel(const float* __restrict__ x_f32, const __nv_bfloat16* __restrict__ x_bf,
                   int64_t rows_phys, Rows R, float* __restrict__ out) {
    ptx::griddep_wait();      // PDL: predecessor complete + visible
    ptx::griddep_launch();    // let the next kernel start its prologue
    __shared__ float s_part[8][8];
    const int b = blockIdx.x, c = blockIdx.y;   // chunk of 8 channels
    const int len = (int)min((long long)R.len[b], (long long)R.n_max);
    const int64_t rs = R.row_start[b];
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int p = threadIdx.x; p < len; p += 256) {
        if (!BF16) {
            const float4* src = reinterpret_cast<const float4*>(x_f32 + (rs + p) * DX_D + c * 8);
            const float4 a = src[0], d = src[1];
            acc[0] += a.x; acc[1] += a.y; acc[2] += a.z; acc[3] += a.w;
            acc[4] += d.x; acc[5] += d.y; acc[6] += d.z; acc[7] += d.w;
        } else {
            const uint4 pk = *reinterpret_cast<const uint4*>(x_bf + ((int64_t)c * rows_phys + DX_HALO + rs + p) * 8);
            const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&pk);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                acc[2 * e] += __low2float(h2[e]);
                acc[2 * e + 1] += __high2float(h2[e]);
            }
        }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = warp_sum(acc[e]);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0)
        for (int e = 0; e < 8; ++e) s_part[warp][e] = acc[e];
    __syncthreads();
    if (threadIdx.x < 8) {
        float t = 0.f;
#pragma unroll
        for (int wv = 0; wv < 8; ++wv) t += s_part[wv][threadIdx.x];
        out[b * DX_D + c * 8 + threadIdx.x] = t / (float)R.len[b];
    }
}

int masked_mean(const float* x_f32, const __nv_bfloat16* x_bf, int64_t rows_phys, const Rows& R, float* out,
                cudaStream_t s) {
    ProfScope prof("masked_mean", s);
    if (x_bf)
        DX_LAUNCH_PDL((masked_mean_kernel<true>), dim3(R.B, 16), 256, 0, s, nullptr, x_bf, rows_phys, R, out);
    else
        DX_LAUNCH_PDL((masked_mean_kernel<false>), dim3(R.B, 16), 256, 0, s, x_f32, nullptr, 0, R, out);
    DX_CHECK_LAUNCH();
    return 0;
}

// rows [*][128] fp32 → padded (B, n_max, 128) with zeros at pads (debug / enc_out)
__global__ void unpack_rows_kernel(const float* __restrict__ x_f32, const __nv_bfloat16* __restrict__ x_bf,
                                   int64_t rows_phys, Rows R, float* __restrict__ out) {
    ptx::griddep_wait();      // PDL: predecessor complete + visible
    ptx::griddep_launch();    // let the next kernel start its prologue
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= (int64_t)R.B * R.n_max) return;
    const int b = (int)(i / R.n_max), p = (int)(i % R.n_max);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p < R.len[b]) {
        int64_t r = R.row_start[b] + p;
        if (x_bf) {
            int64_t off = ((int64_t)(lane >> 1) * rows_phys + DX_HALO + r) * 8 + (lane & 1) * 4;
            uint2 pk = *reinterpret_cast<const uint2*>(x_bf + off);
            __nv_bfloat162 lo = *reinterpret_cast<__nv_bfloat162*>(&pk.x);
            __nv_bfloat162 hi = *reinterpret_cast<__nv_bfloat162*>(&pk.y);
            v = make_float4(__low2float(lo), __high2float(lo), __low2float(hi), __high2float(hi));
        } else {
            v = reinterpret_cast<const float4*>(x_f32 + r * DX_D)[lane];
        }
    }
    reinterpret_cast<float4*>(out + i * DX_D)[lane] = v;
}

int unpack_rows(const float* x_f32, const __nv_bfloat16* x_bf, int64_t rows_phys, const Rows& R, float* out,
                cudaStream_t s) {
    int64_t n = (int64_t)R.B * R.n_max;
    DX_LAUNCH_PDL((unpack_rows_kernel), (int)((n + 7) / 8), 256, 0, s, x_f32, x_bf, rows_phys, R, out);
    DX_CHECK_LAUNCH();
    return 0;
}


// ---------------------------------------------------------------------------------------------------------
// zero_halos: the DX_HALO rows in front of and behind the row range of every channel-interleaved bf16 buffer must
// read as zeros (conv taps and tile overruns land there).  One launch for all buffers of a plan.
// ---------------------------------------------------------------------------------------------------------
__global__ void zero_halos_kernel(HaloList hl, int64_t rows_phys, int64_t cap) {
    ptx::griddep_wait();      // PDL: predecessor complete + visible
    ptx::griddep_launch();    // let the next kernel start its prologue
    const int bi = blockIdx.y;
    if (bi >= hl.n) return;
    __nv_bfloat16* buf = hl.buf[bi];
    const int nchunks = hl.chunks[bi];
    const int64_t tail = rows_phys - (DX_HALO + cap);
    const int64_t per_chunk = DX_HALO + tail;
    const int64_t n = (int64_t)nchunks * per_chunk;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i / per_chunk);
        const int64_t j = i % per_chunk;
        const int64_t row = (j < DX_HALO) ? j : (DX_HALO + cap + (j - DX_HALO));
        *reinterpret_cast<uint4*>(buf + ((int64_t)c * rows_phys + row) * 8) = make_uint4(0u, 0u, 0u, 0u);
    }
}

int zero_halos(const HaloList& hl, int64_t rows_phys, int64_t cap, cudaStream_t s) {
    if (hl.n == 0) return 0;
    dim3 grid(8, hl.n);
    DX_LAUNCH_PDL((zero_halos_kernel), grid, 256, 0, s, hl, rows_phys, cap);
    DX_CHECK_LAUNCH();
    return 0;
}

// ---------------------------------------------------------------------------------------------------------
// ln_apply_bf16: LayerNorm over C = 1024 channels of a channel-interleaved bf16 buffer, in place, using the per-row
// sum / sum-of-squares the producing GEMM's epilogue accumulated (accent pre-net, model.py:655,662).
// Thread ↔ (row, LNA_CHUNKS consecutive 8-channel chunks): the row's liveness and statistics are loaded once and
// reused for LNA_CHUNKS 16-byte loads/stores (all issued before the first use); consecutive threads on consecutive rows.
// ---------------------------------------------------------------------------------------------------------
constexpr int LNA_CHUNKS = 16;

__global__ void __launch_bounds__(256) ln_apply_bf16_kernel(__nv_bfloat16* __restrict__ x, int64_t rows_phys, int C,
                                                            const float* __restrict__ stats, const float* __restrict__ g,
                                                            const float* __restrict__ b, Rows R, int live_ext) {
    ptx::griddep_wait();      // PDL: predecessor complete + visible
    ptx::griddep_launch();    // let the next kernel start its prologue
    const int64_t total = R.row_start[R.B];
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int c0 = blockIdx.y * LNA_CHUNKS;
    if (r >= total) return;
    __nv_bfloat16* ptr = x + ((int64_t)c0 * rows_phys + DX_HALO + r) * 8;
    const int64_t cs = rows_phys * 8;
    if (!row_live(R, r, live_ext)) {
#pragma unroll
        for (int i = 0; i < LNA_CHUNKS; ++i) *reinterpret_cast<uint4*>(ptr + i * cs) = make_uint4(0u, 0u, 0u, 0u);
        return;
    }
    uint4 pk[LNA_CHUNKS];
#pragma unroll
    for (int i = 0; i < LNA_CHUNKS; ++i) pk[i] = *reinterpret_cast<const uint4*>(ptr + i * cs);
    const float4 s01 = reinterpret_cast<const float4*>(stats + r * (TC_STAT_TILES * 2))[0];
    const float4 s23 = reinterpret_cast<const float4*>(stats + r * (TC_STAT_TILES * 2))[1];
    const float mean = (((s01.x + s01.z) + s23.x) + s23.z) / (float)C;      // fixed summation order
    float var = (((s01.y + s01.w) + s23.y) + s23.w) / (float)C - mean * mean;
    var = fmaxf(var, 0.f);
    const float rstd = rsqrtf(var + 1e-5f);
#pragma unroll
    for (int i = 0; i < LNA_CHUNKS; ++i) {
        const int c = c0 + i;
        __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&pk[i]);
        const float4 g0 = reinterpret_cast<const float4*>(g)[c * 2], g1 = reinterpret_cast<const float4*>(g)[c * 2 + 1];
        const float4 b0 = reinterpret_cast<const float4*>(b)[c * 2], b1 = reinterpret_cast<const float4*>(b)[c * 2 + 1];
        const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
        const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            float v0 = (__low2float(h2[e]) - mean) * rstd * gg[2 * e] + bb[2 * e];
            float v1 = (__high2float(h2[e]) - mean) * rstd * gg[2 * e + 1] + bb[2 * e + 1];
            h2[e] = __floats2bfloat162_rn(v0, v1);
        }
        *reinterpret_cast<uint4*>(ptr + i * cs) = pk[i];
    }
}

int ln_apply_bf16(__nv_bfloat16* x, int64_t rows_phys, int C, const float* stats, const float* g, const float* b,
                  const Rows& R, int live_ext, cudaStream_t s) {
    DX_REQUIRE(C % (8 * LNA_CHUNKS) == 0, "ln_apply_bf16: C=%d must be a multiple of %d", C, 8 * LNA_CHUNKS);
    dim3 grid((unsigned)((R.cap + 255) / 256), (unsigned)(C / (8 * LNA_CHUNKS)));
    ProfScope prof("ln_apply_bf16", s);
    DX_LAUNCH_PDL((ln_apply_bf16_kernel), grid, 256, 0, s, x, rows_phys, C, stats, g, b, R, live_ext);
    DX_CHECK_LAUNCH();
    return 0;
}

}  // namespace dx
