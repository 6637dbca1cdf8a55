// Shared declarations of libdaft_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>
#include "../../include/daft_b200.h"

namespace dx {

// ---- error handling: no exceptions cross the C ABI -------------------------------------------------------
void set_error(const char* fmt, ...);
void count_launch(int n = 1);
void* trace_buffer(size_t min_bytes);   // dx_set_trace_buffer: device scratch for kernel timelines, or null
int debug_knob(int key);   // dx_set_debug(key, value): bring-up switches, all 0 in production

// Optional per-launch timing (dx_profile_enable): events recorded around a launch on its stream, keyed by a tag.
struct ProfScope {
    cudaStream_t s;
    int slot;
    ProfScope(const char* tag, cudaStream_t stream);
    ~ProfScope();
};

#define DX_CHECK_CUDA(expr)                                                                   \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess) {                                                              \
            dx::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return 1;                                                                         \
        }                                                                                     \
    } while (0)

#define DX_CHECK_LAUNCH()                                                                     \
    do {                                                                                      \
        dx::count_launch();                                                                   \
        cudaError_t _e = cudaGetLastError();                                                  \
        if (_e == cudaSuccess && dx::debug_knob(2)) _e = cudaDeviceSynchronize();             \
        if (_e != cudaSuccess) {                                                              \
            dx::set_error("%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(_e)); \
            return 1;                                                                         \
        }                                                                                     \
    } while (0)

#define DX_REQUIRE(cond, ...)                                                                 \
    do {                                                                                      \
        if (!(cond)) {                                                                        \
            dx::set_error(__VA_ARGS__);                                                       \
            return 2;                                                                         \
        }                                                                                     \
    } while (0)

#define DX_TRY(expr)                                                                          \
    do {                                                                                      \
        int _s = (expr);                                                                      \
        if (_s != 0) return _s;                                                               \
    } while (0)

// ---- packed row stream -----------------------------------------------------------------------------------
// Utterance b owns rows [row_start[b], row_start[b] + len[b] + DX_GAP).  Row r of the stream is
// (row_b[r], row_p[r]); p >= len marks a gap row.  "live(ext)" is the H1 rule: a conv
// intermediate at position p carries the reference's value iff p < min(len + ext, n_max); elsewhere the
// reference sees true zero padding (p >= n_max, or p = -1 which is the previous utterance's last gap row).
struct Rows {
    const int32_t* row_start;  // [B+1]
    const int32_t* row_b;      // [cap]
    const int32_t* row_p;      // [cap]
    const int64_t* len;        // [B]
    int32_t B, n_max;
    int64_t cap;               // allocated rows (multiple of 256)
};

__device__ __forceinline__ bool row_live(const Rows& R, int64_t r, int ext) {
    int b = R.row_b[r];
    if (b < 0) return false;
    int p = R.row_p[r];
    long long lim = (long long)R.len[b] + ext;
    if (lim > R.n_max) lim = R.n_max;
    return p < lim;
}

// Launch with programmatic stream serialization (PDL): the kernel may start while the previous kernel of the
// stream drains; it must call ptx::griddep_wait() before touching anything that kernel wrote.  debug knob 5
// switches the attribute off (plain stream order) for A/B measurements.
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = debug_knob(5) ? 0 : 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#define DX_LAUNCH_PDL(kernel, grid, block, smem, stream, ...)                                  \
    do {                                                                                      \
        cudaError_t _le = dx::launch_pdl(kernel, grid, block, smem, stream, __VA_ARGS__);     \
        if (_le != cudaSuccess) {                                                             \
            dx::set_error("%s:%d: launch of %s -> %s", __FILE__, __LINE__, #kernel, cudaGetErrorString(_le)); \
            return 1;                                                                         \
        }                                                                                     \
    } while (0)

static inline int64_t round_up(int64_t a, int64_t m) { return (a + m - 1) / m * m; }

// ---- per-device one-time kernel configuration --------------------------------------------------------------
// cudaFuncAttributeMaxDynamicSharedMemorySize and the SM count are PER DEVICE: a process that drives several GPUs
// (or several host threads: bulk.BulkRunner, the writer) must not share one "configured" flag.  The caches are keyed by
// cudaGetDevice(); entries are atomics, and two threads racing on the same entry only set the same attribute twice.
constexpr int DX_MAX_DEVICES = 64;
struct SmemConfigured {
    std::atomic<size_t> bytes[DX_MAX_DEVICES];
};
int device_sm_count(int* sms);      // SM count of the current device (dx_api.cu)

template <typename Kern>
static inline int ensure_dyn_smem(Kern kernel, size_t smem, SmemConfigured& cache) {
    int dev = 0;
    DX_CHECK_CUDA(cudaGetDevice(&dev));
    const bool cached = dev >= 0 && dev < DX_MAX_DEVICES;
    if (cached && smem <= cache.bytes[dev].load(std::memory_order_acquire)) return 0;
    DX_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (cached) cache.bytes[dev].store(smem, std::memory_order_release);
    return 0;
}

// ---- warp helpers ----------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// LayerNorm affine + FiLM of one value with the rounding points pinned (no compiler-chosen contraction), so that every
// kernel that finishes a row this way (fused block kernel, tc_gemm / tc_linear LN epilogue) produces the same bits
__device__ __forceinline__ float ln_film(float v, float mean, float rstd, float g, float be, float fg, float fb) {
    const float y = __fmaf_rn(__fmul_rn(__fsub_rn(v, mean), rstd), g, be);
    return __fmaf_rn(fg, y, fb);
}

// ---- fp32 path (dx_f32.cu) -------------------------------------------------------------------------------
// out[r][n] = act(bias[n] + sum_j sum_k A[r + j - taps/2][k] * W[j][k][n]); rows outside [0, total) read as 0;
// rows that are not live(ext) are written as 0.  ext < 0: no liveness masking (plain linear layers).
int gemm_f32(const float* A, int K, const float* W, const float* bias, int N, int taps, bool relu,
             const Rows& R, int live_ext, float* out, cudaStream_t s, const float* addend = nullptr);
// out = live(ext) ? film(LN(in + res) * g + b) : 0   (one warp per row; C = 128 or 1024)
struct LnDrop;      // dx_dropout.cuh (training step)
// block counts of the three stacks (DxModelWeights.n_*_blocks: 0 = the reference's default of 4)
__host__ __device__ static inline int n_blocks_of(int32_t n) { return n == 0 ? 4 : (int)n; }
int ln_rows_f32(const float* in, const float* res, const float* g, const float* b, int C, const float* film,
                int film_stride, const Rows& R, int live_ext, float* out, cudaStream_t s, const LnDrop* drop = nullptr);
// masked multi-head self attention on packed rows; qkv [rows][384] → out [rows][128]
int attention_f32(const float* qkv, const Rows& R, float* out, cudaStream_t s, float* lse = nullptr);

// ---- row-stream plumbing (dx_rows.cu) --------------------------------------------------------------------
int build_rows(const int64_t* lengths, int B, int n_max, int32_t* row_start, int32_t* row_b, int32_t* row_p,
               int64_t cap, cudaStream_t s);

}  // namespace dx
