// C ABI of libdaft_b200.so (include/daft_b200.h) and the whole-pass planner: DaftExprt.forward
// (model.py:889-948) / DaftExprt.inference (model.py:1026-1114) as one sequence of kernel launches on the
// caller's stream, with no host synchronisation and no allocation (all scratch comes from the workspace).
#include <stdarg.h>
#include <string.h>
#include <atomic>
#include <map>
#include <string>
#include <vector>
#include "dx_common.cuh"
#include "dx_kernels.cuh"
#include "dx_plan.cuh"
#include "dx_train.cuh"

namespace dx {

static thread_local char g_error[1024] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
static int g_knobs[32] = {0};
int debug_knob(int key) { return (key >= 0 && key < 32) ? g_knobs[key] : 0; }
static void* g_trace = nullptr;
static size_t g_trace_bytes = 0;
void* trace_buffer(size_t min_bytes) { return (g_trace != nullptr && g_trace_bytes >= min_bytes) ? g_trace : nullptr; }

// ---- launch profiler ----------------------------------------------------------------------------------------
static bool g_prof_on = false;
struct ProfRec { const char* tag; cudaEvent_t a, b; };
static std::vector<ProfRec> g_prof;
static size_t g_prof_used = 0;

ProfScope::ProfScope(const char* tag, cudaStream_t stream) : s(stream), slot(-1) {
    if (!g_prof_on) return;
    if (g_prof_used == g_prof.size()) {
        ProfRec r{tag, nullptr, nullptr};
        if (cudaEventCreate(&r.a) != cudaSuccess || cudaEventCreate(&r.b) != cudaSuccess) return;
        g_prof.push_back(r);
    }
    slot = (int)g_prof_used++;
    g_prof[slot].tag = tag;
    cudaEventRecord(g_prof[slot].a, s);
}
ProfScope::~ProfScope() {
    if (slot >= 0) cudaEventRecord(g_prof[slot].b, s);
}

int device_sm_count(int* sms) {
    static std::atomic<int> cache[DX_MAX_DEVICES];
    int dev = 0;
    DX_CHECK_CUDA(cudaGetDevice(&dev));
    const bool cached = dev >= 0 && dev < DX_MAX_DEVICES;
    int n = cached ? cache[dev].load(std::memory_order_acquire) : 0;
    if (n == 0) {
        DX_CHECK_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
        if (cached) cache[dev].store(n, std::memory_order_release);
    }
    *sms = n;
    return 0;
}

int64_t row_cap(int B, int n_max) { return round_up((int64_t)B * (n_max + DX_GAP), 256); }

}  // namespace dx

using namespace dx;

extern "C" {

int dx_version(void) { return 100; }
const char* dx_last_error_string(void) { return g_error; }
int64_t dx_launch_count(void) { return g_launches.load(); }
int dx_set_debug(int key, int value) {
    if (key < 0 || key >= 32) return 2;
    g_knobs[key] = value;
    return 0;
}

int dx_set_trace_buffer(void* device_ptr, int64_t bytes) {
    g_trace = device_ptr;
    g_trace_bytes = device_ptr ? (size_t)bytes : 0;
    return 0;
}

int dx_profile_enable(int on) {
    g_prof_on = on != 0;
    g_prof_used = 0;
    return 0;
}

int dx_profile_report(char* buf, int64_t cap) {
    DX_REQUIRE(buf != nullptr && cap > 0, "dx_profile_report: buffer required");
    DX_CHECK_CUDA(cudaDeviceSynchronize());
    std::map<std::string, std::pair<int, double>> acc;
    for (size_t i = 0; i < g_prof_used; ++i) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, g_prof[i].a, g_prof[i].b) != cudaSuccess) continue;
        auto& e = acc[g_prof[i].tag];
        e.first += 1;
        e.second += ms;
    }
    g_prof_used = 0;
    std::string out;
    char line[160];
    for (auto& kv : acc) {
        snprintf(line, sizeof(line), "%s %d %.6f\n", kv.first.c_str(), kv.second.first, kv.second.second);
        out += line;
    }
    DX_REQUIRE((int64_t)out.size() + 1 <= cap, "dx_profile_report: buffer of %lld bytes too small", (long long)cap);
    memcpy(buf, out.c_str(), out.size() + 1);
    return 0;
}

int64_t dx_row_cap(int B, int n_max) { return row_cap(B, n_max); }

int dx_build_rows(const int64_t* lengths, int B, int n_max, int32_t* row_start, int32_t* row_b, int32_t* row_p,
                  int64_t cap, void* stream) {
    DX_REQUIRE(cap >= row_cap(B, n_max), "dx_build_rows: row_cap %lld < required %lld", (long long)cap,
               (long long)row_cap(B, n_max));
    return build_rows(lengths, B, n_max, row_start, row_b, row_p, cap, (cudaStream_t)stream);
}

int64_t dx_packed_weight_elems(int taps, int K, int N, int nt, int ks) { return packed_weight_elems(taps, K, N, nt, ks); }

int dx_pack_weight_bf16(const float* w, int taps, int K, int N, int nt, int ks, void* out, void* stream) {
    return pack_weight_bf16(w, taps, K, N, nt, ks, (__nv_bfloat16*)out, (cudaStream_t)stream);
}

int64_t dx_forward_workspace_bytes(int B, int L_max, int T_max, int mode) {
    Plan p;
    if (plan_layout(&p, B, L_max, T_max, mode, nullptr, 0) != 0) return -1;
    return p.bytes;
}

int dx_forward(const DxModelWeights* w, const DxForwardArgs* a, void* stream) {
    DX_REQUIRE(w != nullptr && a != nullptr, "dx_forward: null argument");
    return run_forward(*w, *a, (cudaStream_t)stream);
}

int64_t dx_train_workspace_bytes(int B, int L_max, int T_max) { return train_workspace_bytes(B, L_max, T_max); }

int dx_train_step(const DxModelWeights* w, const DxModelWeights* grads, const DxTrainArgs* a, void* stream) {
    DX_REQUIRE(w != nullptr && grads != nullptr && a != nullptr, "dx_train_step: null argument");
    return run_train_step(*w, *grads, *a, (cudaStream_t)stream);
}

int dx_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1,
                 float beta2, float eps, float weight_decay, int step, float grad_scale, void* stream) {
    DX_REQUIRE(params && grads && exp_avg && exp_avg_sq && n >= 0, "dx_adam_step: null argument");
    return adam_step(params, grads, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay, step, grad_scale,
                     (cudaStream_t)stream);
}

int dx_accent_encoder(const DxModelWeights* w, int mode, int B, int T_max, const float* frames_energy,
                      const float* frames_pitch, const float* mel_specs, const int64_t* output_lengths,
                      float* accent_emb, void* workspace, int64_t workspace_bytes, void* stream) {
    DX_REQUIRE(w != nullptr, "dx_accent_encoder: null weights");
    Plan p;
    DX_TRY(plan_layout(&p, B, 1, T_max, mode, workspace, workspace_bytes));
    cudaStream_t s = (cudaStream_t)stream;
    Rows RT = p.rows_t(output_lengths);
    if (mode == DX_MODE_BF16) DX_TRY(zero_plan_halos(p, s));
    DX_TRY(build_rows(output_lengths, B, T_max, p.t_row_start, p.t_row_b, p.t_row_p, p.cap_t, s));
    return run_accent_encoder(*w, p, RT, frames_energy, frames_pitch, mel_specs, accent_emb, s);
}

int dx_style(const DxModelWeights* w, int B, const float* spk_embs, const float* spk_emb_in,
             const float* accent_emb, float* film, float* speaker_preds, void* stream) {
    DX_REQUIRE(w != nullptr, "dx_style: null weights");
    return style(w, B, spk_embs, spk_emb_in, accent_emb, film, speaker_preds, (cudaStream_t)stream);
}

int dx_length_regulator_integers(const int64_t* durations_int, const int64_t* input_lengths, int B, int L_max,
                                 int T_max, int64_t* cumsum, int64_t* two_mu, int64_t* output_lengths,
                                 int64_t* frame2symbol_out, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    DX_REQUIRE(cumsum != nullptr && output_lengths != nullptr, "dx_length_regulator_integers: cumsum/output_lengths required");
    DX_TRY(lenreg_scan(durations_int, B, L_max, false, cumsum, nullptr, output_lengths, two_mu, s));
    if (frame2symbol_out)
        DX_TRY(frame2symbol(cumsum, input_lengths, output_lengths, B, L_max, T_max, frame2symbol_out, s));
    return 0;
}

int dx_int_durations(float* duration, int B, int L_max, int sampling_rate, int filter_length, int hop_length,
                     int centered, int64_t* durations_int, int32_t* status, void* stream) {
    return int_durations(duration, B, L_max, sampling_rate, filter_length, hop_length, centered, durations_int,
                         status, (cudaStream_t)stream);
}

int dx_prosody_transform(float* energy, float* pitch, const float* energy_factors, const float* pitch_factors,
                         const int64_t* durations_int, const float* pitch_mean, const float* pitch_std, int B,
                         int L_max, int transform, void* stream) {
    return prosody_transform(energy, pitch, energy_factors, pitch_factors, durations_int, pitch_mean, pitch_std, B,
                             L_max, transform, (cudaStream_t)stream);
}

int dx_tc_gemm_test(const void* a_il, int64_t a_rows_phys, int M, int K, int N, int taps, const void* w_packed,
                    int nt, int ks, float* out, void* stream) {
    TcGemmArgs g;
    memset(&g, 0, sizeof(g));
    g.a = (const __nv_bfloat16*)a_il;
    g.a_rows_phys = a_rows_phys;
    g.K = K; g.N = N; g.taps = taps;
    g.w = (const __nv_bfloat16*)w_packed;
    g.nt = nt; g.ks = ks;
    g.epilogue = TC_EPI_F32;
    g.live_ext = -1;
    g.out_f32 = out;
    return tc_gemm_plain(g, M, (cudaStream_t)stream);
}

int dx_test_attention(int mode, const void* qkv, void* out, int64_t rows_phys, const int64_t* lengths, int B, int n_max,
                      int32_t* rowmap_scratch, int64_t cap, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    DX_REQUIRE(cap >= row_cap(B, n_max), "dx_test_attention: row_cap too small");
    int32_t* rs = rowmap_scratch;
    int32_t* rb = rs + (B + 1);
    int32_t* rpos = rb + cap;
    DX_TRY(build_rows(lengths, B, n_max, rs, rb, rpos, cap, s));
    Rows R{rs, rb, rpos, lengths, B, n_max, cap};
    if (mode == DX_MODE_F32) return attention_f32((const float*)qkv, R, (float*)out, s);
    DX_REQUIRE(rows_phys >= cap + DX_HALO + DX_TAIL, "dx_test_attention: rows_phys too small");
    return attention_bf16((const __nv_bfloat16*)qkv, (__nv_bfloat16*)out, rows_phys, R, s);
}

}  // extern "C"
