// dx_attention_static: masked multi-head self-attention (same kernels as dx_test_attention) for callers that cannot
// allocate device memory (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The row map the
// attention kernels need (row_start / row_b / row_p, normally int32 scratch the caller passes) lives in a file-scope
// __device__ array of a fixed, modest size instead.  No counterpart in the
// reference.
//
// Rules this file keeps:
//   * the array's NAME is only a host-side stand-in in host code: the device address is fetched once per device, with
//     that device current, by cudaGetSymbolAddress (return code checked) and only that pointer reaches a kernel;
//   * no cudaMalloc / cudaFree / allocating library routine anywhere on the path;
//   * every call on a device shares the one array, which keeps the previous call's contents.  Nothing in it has to
//     start at zero: build_rows rewrites row_start[0..B] and every row_b / row_p entry up to the call's row_cap before
//     the attention kernel (same stream) reads them.  Two calls that overlap on the GPU (different streams, or a
//     graph replay next to a live call) RACE on the array: callers serialise them.
#include <mutex>
#include "dx_common.cuh"
#include "dx_kernels.cuh"

namespace dx {

int64_t row_cap(int B, int n_max);   // dx_api.cu

constexpr int     NOALLOC_MAX_B   = 255;                          // row_start needs B + 1 <= 256 ints
constexpr int64_t NOALLOC_MAX_CAP = DX_ATTENTION_STATIC_MAX_ROWS;  // rows of the packed stream (multiple of 256)
constexpr int64_t NOALLOC_INTS    = (NOALLOC_MAX_B + 1) + 2 * NOALLOC_MAX_CAP;   // 65,792 ints = 257 KB per GPU

static __device__ __align__(256) int32_t g_rowmap_static[NOALLOC_INTS];

// device address of g_rowmap_static on the current device (looked up once per device, under a mutex)
static int rowmap_static_ptr(int32_t** out) {
    static std::mutex mu;
    static int32_t* cache[DX_MAX_DEVICES] = {nullptr};
    int dev = 0;
    DX_CHECK_CUDA(cudaGetDevice(&dev));
    DX_REQUIRE(dev >= 0 && dev < DX_MAX_DEVICES, "dx_attention_static: device %d beyond the %d supported", dev, DX_MAX_DEVICES);
    std::lock_guard<std::mutex> lock(mu);
    if (cache[dev] == nullptr) {
        void* p = nullptr;
        DX_CHECK_CUDA(cudaGetSymbolAddress(&p, g_rowmap_static));
        DX_REQUIRE(p != nullptr, "dx_attention_static: cudaGetSymbolAddress returned a null address");
        cache[dev] = (int32_t*)p;
    }
    *out = cache[dev];
    return 0;
}

}  // namespace dx

using namespace dx;

extern "C" {

int64_t dx_attention_static_max_rows(void) { return NOALLOC_MAX_CAP; }

int dx_attention_static(int mode, const void* qkv, void* out, int64_t rows_phys, const int64_t* lengths, int B, int n_max,
                        void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    DX_REQUIRE(mode == DX_MODE_F32 || mode == DX_MODE_BF16, "dx_attention_static: unknown mode %d", mode);
    DX_REQUIRE(qkv != nullptr && out != nullptr && lengths != nullptr, "dx_attention_static: null argument");
    DX_REQUIRE(B >= 1 && B <= NOALLOC_MAX_B, "dx_attention_static: B = %d outside 1..%d (the static row map's limit)", B,
               NOALLOC_MAX_B);
    DX_REQUIRE(n_max >= 1, "dx_attention_static: n_max = %d", n_max);
    const int64_t cap = row_cap(B, n_max);
    DX_REQUIRE(cap <= NOALLOC_MAX_CAP,
               "dx_attention_static: B * (n_max + %d) rounded up to 256 = %lld rows exceeds the static row map's %lld rows; "
               "use dx_test_attention with caller scratch",
               DX_GAP, (long long)cap, (long long)NOALLOC_MAX_CAP);
    int32_t* rs = nullptr;
    DX_TRY(rowmap_static_ptr(&rs));
    int32_t* rb = rs + (NOALLOC_MAX_B + 1);
    int32_t* rpos = rb + NOALLOC_MAX_CAP;
    DX_TRY(build_rows(lengths, B, n_max, rs, rb, rpos, cap, s));
    Rows R{rs, rb, rpos, lengths, B, n_max, cap};
    if (mode == DX_MODE_F32) return attention_f32((const float*)qkv, R, (float*)out, s);
    DX_REQUIRE(rows_phys >= cap + DX_HALO + DX_TAIL, "dx_attention_static: rows_phys too small");
    return attention_bf16((const __nv_bfloat16*)qkv, (__nv_bfloat16*)out, rows_phys, R, s);
}

}  // extern "C"
