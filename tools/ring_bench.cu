// Weight-ring microbenchmark for sm_100a: how fast can a warp-specialised CTA feed tcgen05.mma from a ring of
// bulk-copied weight stages?  Mirrors the main loop of the fused FFN kernel (dx_ffn_tc.cu) without its epilogue:
//   warp 0 (producer): wait w_empty[s] -> bulk copy of one stage (or a plain arrive) -> w_full[s]
//   warp 1 (issuer)  : wait w_full[s]  -> MPS MMAs (128x128x16) -> tcgen05.commit -> w_empty[s]
// Variants: ring depth, stage size, copies on/off, and "self refill" (the issuer warp refills the ring itself,
// LAG stages behind its own issue point, so no second warp sits in the handshake).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/ring_bench tools/ring_bench.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include "../ubisoft-laforge-daft-exprt_b200/csrc/dx_ptx.cuh"

using namespace dx;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s -> %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

constexpr int N_STAGES_TOTAL = 96 * 8;       // stages streamed per CTA (8 "tiles" of 96 stages)
constexpr int W_REGION_STAGES = 96;          // the weight set that is cycled through (stays in L2)

// NS: ring depth, SB: stage bytes (16384: 8 MMAs per stage, 8192: 4 MMAs), MODE: 0 = producer warp + copies,
// 1 = producer warp, arrive only, 2 = issuer refills the ring itself (copies), LAG stages behind
template <int NS, int SB, int MODE, int LAG, int FENCE = 1, int PIPE = 0, int LANE0 = 0, int TWO = 0, int RELAY = 0, int FINE = 0>
__global__ void __launch_bounds__(128, 1) ring_bench_kernel(const uint8_t* w, long long* out, long long* stamps) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t w_full[NS], w_empty[NS];
    __shared__ uint32_t tmem_ptr;
    __shared__ volatile int landed;       // RELAY: number of stages a waiter warp has seen land
    constexpr int MPS = SB / 2048;                    // MMAs per stage and half (K = 16 each: 2 chunks of 2048 B... x 2 halves)
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        for (int i = 0; i < NS; ++i) { ptx::mbar_init(&w_full[i], 1); ptx::mbar_init(&w_empty[i], TWO ? 2 : 1); }
        ptx::fence_barrier_init();
    }
    if (threadIdx.x == 0) landed = 0;
    if (warp == 0) ptx::tmem_alloc<512>(&tmem_ptr);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = tmem_ptr;
    uint8_t* sA = smem;                               // 64 KB "X tile" (garbage)
    uint8_t* sW = smem + 65536;
    constexpr uint32_t idesc = ptx::make_idesc_bf16(128, 128);
    if (warp == 0 && MODE != 2) {
        if (ptx::elect_one()) {
            for (int it = 0; it < N_STAGES_TOTAL; ++it) {
                const int s = it % NS;
                ptx::mbar_wait(&w_empty[s], ((it / NS) & 1) ^ 1);
                if (MODE == 0) {
                    ptx::mbar_expect_tx(&w_full[s], SB);
                    ptx::bulk_g2s(sW + s * SB, w + (size_t)(it % W_REGION_STAGES) * SB, SB, &w_full[s]);
                } else {
                    ptx::mbar_arrive(&w_full[s]);
                }
            }
        }
    } else if (RELAY && warp == 3) {
        // waiter: turns mbarrier completions into a plain shared-memory counter the issuers can poll with a 30-cycle load
        for (int it = 0; it < N_STAGES_TOTAL; ++it) {
            ptx::mbar_wait(&w_full[it % NS], (it / NS) & 1);
            if ((threadIdx.x & 31) == 0) asm volatile("st.release.cta.shared::cta.s32 [%0], %1;" ::"r"(ptx::smem_u32((const void*)&landed)), "r"(it + 1) : "memory");
        }
    } else if (warp == 1 || (TWO && warp == 2)) {
        const uint32_t aA = ptx::smem_u32(sA), aW = ptx::smem_u32(sW);
        long long t0 = clock64();
        if (MODE == 2 && ptx::elect_one()) {
            for (int it = 0; it < NS - LAG && it < N_STAGES_TOTAL; ++it) {       // initial fill
                ptx::mbar_expect_tx(&w_full[it % NS], SB);
                ptx::bulk_g2s(sW + (it % NS) * SB, w + (size_t)(it % W_REGION_STAGES) * SB, SB, &w_full[it % NS]);
            }
        }
        __syncwarp();
        for (int it = 0; it < N_STAGES_TOTAL; ++it) {
            const int s = it % NS;
            if (MODE == 2) {
                // refill: stage it + NS - LAG goes into the slot of stage it - LAG (whose MMAs were issued LAG stages ago)
                const int nx = it + NS - LAG;
                if (nx < N_STAGES_TOTAL && ptx::elect_one()) {
                    const int sn = nx % NS;
                    if (nx >= NS) ptx::mbar_wait(&w_empty[sn], ((nx / NS) & 1) ^ 1);     // MMAs of stage it - LAG have completed
                    ptx::mbar_expect_tx(&w_full[sn], SB);
                    ptx::bulk_g2s(sW + sn * SB, w + (size_t)(nx % W_REGION_STAGES) * SB, SB, &w_full[sn]);
                }
                __syncwarp();
            }
            const bool rec = stamps != nullptr && blockIdx.x == 0 && warp == 1 && it >= 200 && it < 216 && (threadIdx.x & 31) == 0;
            if (rec) stamps[(it - 200) * 3] = clock64();
            if (RELAY) {
                int seen;
                do {
                    asm volatile("ld.acquire.cta.shared::cta.s32 %0, [%1];" : "=r"(seen) : "r"(ptx::smem_u32((const void*)&landed)) : "memory");
                } while (seen <= it);
            } else if (PIPE == 0 || it == 0) {
                ptx::mbar_wait(&w_full[s], (it / NS) & 1);
                if (FENCE) ptx::tc_fence_after();
            }
            if (rec) stamps[(it - 200) * 3 + 1] = clock64();
            const bool issue = LANE0 ? (threadIdx.x & 31) == 0 : ptx::elect_one();
            if (issue) {
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    if (TWO && h != warp - 1) continue;       // one issuer warp per half
#pragma unroll
                    for (int kk = 0; kk < MPS / 2; ++kk) {
                        if (PIPE && h == 1 && kk == MPS / 2 - 2 && it + 1 < N_STAGES_TOTAL) {
                            // software pipelining: the wait for the next stage happens while MMAs of this one are queued
                            ptx::mbar_wait(&w_full[(it + 1) % NS], ((it + 1) / NS) & 1);
                            if (FENCE) ptx::tc_fence_after();
                        }
                        const uint64_t da = ptx::make_desc(aA + kk * 2 * 4096 + h * 126 * 16, 4096, 128);
                        const uint64_t db = ptx::make_desc(aW + s * SB + kk * 2 * 2048, 2048, 128);
                        ptx::mma_bf16(tmem_base + 128 * h, da, db, idesc, (it | kk) != 0);
                        if (FINE && stamps != nullptr && blockIdx.x == 0 && warp == 1 && it >= 200 && it < 204)
                            stamps[64 + (it - 200) * 12 + h * (MPS / 2) + kk] = clock64();
                    }
                }
                ptx::mma_commit(&w_empty[s]);
                if (FINE && stamps != nullptr && blockIdx.x == 0 && warp == 1 && it >= 200 && it < 204) stamps[64 + (it - 200) * 12 + 8] = clock64();
            }
            if (!PIPE) __syncwarp();
            if (rec) stamps[(it - 200) * 3 + 2] = clock64();
        }
        // drain: the last NS commits
        for (int it = N_STAGES_TOTAL - NS; it < N_STAGES_TOTAL; ++it)
            if (it >= 0) ptx::mbar_wait(&w_empty[it % NS], (it / NS) & 1);
        long long t1 = clock64();
        if ((threadIdx.x & 31) == 0 && warp == 1) out[blockIdx.x] = t1 - t0;
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<512>(tmem_base);
    }
}

template <int NS, int SB, int MODE, int LAG, int FENCE = 1, int PIPE = 0, int LANE0 = 0, int TWO = 0, int RELAY = 0, int FINE = 0>
static void run(const uint8_t* d_w, int grid, bool show = false) {
    long long* d_out;
    CK(cudaMalloc(&d_out, grid * sizeof(long long)));
    const size_t smem = 65536 + (size_t)NS * SB + 1024;
    auto kern = ring_bench_kernel<NS, SB, MODE, LAG, FENCE, PIPE, LANE0, TWO, RELAY, FINE>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long* d_st;
    CK(cudaMalloc(&d_st, 128 * sizeof(long long)));
    CK(cudaMemset(d_st, 0, 128 * sizeof(long long)));
    for (int rep = 0; rep < 2; ++rep) {
        kern<<<grid, 128, smem>>>(d_w, d_out, show ? d_st : nullptr);
        CK(cudaDeviceSynchronize());
    }
    std::vector<long long> h(grid);
    CK(cudaMemcpy(h.data(), d_out, grid * sizeof(long long), cudaMemcpyDeviceToHost));
    double tot = 0;
    for (int i = 0; i < grid; ++i) tot += h[i];
    tot /= grid;
    const int mmas = N_STAGES_TOTAL * (SB / 2048);
    static const char* mode_name[] = {"producer warp, copies", "producer warp, arrive only", "issuer refills itself"};
    printf("ring %2d x %5d B  %-28s lag %d fence %d pipe %d lane0 %d two %d relay %d  grid %3d  %7.1f cyc/MMA  %7.0f cyc/stage\n", NS, SB, mode_name[MODE], LAG, FENCE, PIPE, LANE0, TWO, RELAY, grid,
           tot / mmas, tot / N_STAGES_TOTAL);
    if (show && FINE) {
        long long st[128];
        CK(cudaMemcpy(st, d_st, sizeof st, cudaMemcpyDeviceToHost));
        for (int i = 0; i < 4; ++i) {
            printf("   stage %d: wait start +%lld | wait done +%lld | MMA issued", 200 + i, st[i * 3] - st[0], st[i * 3 + 1] - st[0]);
            for (int k = 0; k < 9; ++k) printf(" +%lld", st[64 + i * 12 + k] ? st[64 + i * 12 + k] - st[0] : 0);
            printf(" (last = commit)\n");
        }
    } else if (show) {
        long long st[48];
        CK(cudaMemcpy(st, d_st, sizeof st, cudaMemcpyDeviceToHost));
        for (int i = 0; i < 16; ++i)
            printf("   stage %d: loop start +%lld, wait done +%lld, issued +%lld\n", 200 + i, st[i * 3] - st[0], st[i * 3 + 1] - st[i * 3],
                   st[i * 3 + 2] - st[i * 3 + 1]);
    }
    fflush(stdout);
    CK(cudaFree(d_st));
    CK(cudaFree(d_out));
}

// ---- lean variant: what does the issuing thread spend per stage, and which trims matter?  VAR bits: 1 = leader lane
// chosen once (no elect.sync per stage), 2 = no __syncwarp per stage, 4 = no tcgen05 fence after the weight wait,
// 8 = slot / phase kept incrementally (no modulo), 16 = wait with an inlined first try, 32 = only lane 0 runs the loop
template <int NS, int TWO, int VAR>
__global__ void __launch_bounds__(128, 1) ring_opt_kernel(const uint8_t* w, long long* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t w_full[NS], w_empty[NS];
    __shared__ uint32_t tmem_ptr;
    constexpr int SB = 16384;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < NS; ++i) { ptx::mbar_init(&w_full[i], 1); ptx::mbar_init(&w_empty[i], TWO ? 2 : 1); }
        ptx::fence_barrier_init();
    }
    if (warp == 0) ptx::tmem_alloc<512>(&tmem_ptr);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = tmem_ptr;
    uint8_t* sA = smem;
    uint8_t* sW = smem + 65536;
    constexpr uint32_t idesc = ptx::make_idesc_bf16(128, 128);
    if (warp == 0) {
        if (ptx::elect_one()) {
            for (int it = 0; it < N_STAGES_TOTAL; ++it) {
                const int s = it % NS;
                ptx::mbar_wait(&w_empty[s], ((it / NS) & 1) ^ 1);
                ptx::mbar_expect_tx(&w_full[s], SB);
                ptx::bulk_g2s(sW + s * SB, w + (size_t)(it % W_REGION_STAGES) * SB, SB, &w_full[s]);
            }
        }
    } else if ((warp == 1 || (TWO && warp == 2)) && (!(VAR & 32) || lane == 0)) {
        const int h0 = TWO ? warp - 1 : 0, h1 = TWO ? warp : 2;
        const uint32_t aA = ptx::smem_u32(sA), aW = ptx::smem_u32(sW);
        const long long t0 = clock64();
        const bool leader = (VAR & 32) ? true : ((VAR & 1) ? lane == 0 : false);
        int s = 0;
        uint32_t phase = 0;
        for (int it = 0; it < N_STAGES_TOTAL; ++it) {
            if (!(VAR & 8)) { s = it % NS; phase = (it / NS) & 1; }
            if (VAR & 16) {
                if (!ptx::mbar_try_wait(&w_full[s], phase)) ptx::mbar_wait(&w_full[s], phase);
            } else {
                ptx::mbar_wait(&w_full[s], phase);
            }
            if (!(VAR & 4)) ptx::tc_fence_after();
            if ((VAR & (1 | 32)) ? leader : ptx::elect_one()) {
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    if (h < h0 || h >= h1) continue;
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {
                        const uint64_t da = ptx::make_desc(aA + kk * 2 * 4096 + h * 126 * 16, 4096, 128);
                        const uint64_t db = ptx::make_desc(aW + s * SB + kk * 2 * 2048, 2048, 128);
                        ptx::mma_bf16(tmem_base + 128 * h, da, db, idesc, (it | kk) != 0);
                    }
                }
                ptx::mma_commit(&w_empty[s]);
            }
            if (!(VAR & (2 | 32))) __syncwarp();
            if (VAR & 8) { if (++s == NS) { s = 0; phase ^= 1; } }
        }
        for (int it = N_STAGES_TOTAL - NS; it < N_STAGES_TOTAL; ++it) ptx::mbar_wait(&w_empty[it % NS], (it / NS) & 1);
        const long long t1 = clock64();
        if (lane == 0 && warp == 1) out[blockIdx.x] = t1 - t0;
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<512>(tmem_base);
    }
}

template <int NS, int TWO, int VAR>
static void run_opt(const uint8_t* d_w, int grid) {
    long long* d_out;
    CK(cudaMalloc(&d_out, grid * sizeof(long long)));
    const size_t smem = 65536 + (size_t)NS * 16384 + 1024;
    auto kern = ring_opt_kernel<NS, TWO, VAR>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (int rep = 0; rep < 2; ++rep) {
        kern<<<grid, 128, smem>>>(d_w, d_out);
        CK(cudaDeviceSynchronize());
    }
    std::vector<long long> h(grid);
    CK(cudaMemcpy(h.data(), d_out, grid * sizeof(long long), cudaMemcpyDeviceToHost));
    double tot = 0;
    for (int i = 0; i < grid; ++i) tot += h[i];
    tot /= grid;
    printf("lean ring %d x 16 KB, %d issuer warp(s), trims 0x%02x  %7.1f cyc/MMA  %7.0f cyc/stage\n", NS, TWO ? 2 : 1, VAR,
           tot / (N_STAGES_TOTAL * 8), tot / N_STAGES_TOTAL);
    fflush(stdout);
    CK(cudaFree(d_out));
}

int main() {
    uint8_t* d_w;
    CK(cudaMalloc(&d_w, (size_t)W_REGION_STAGES * 16384));
    CK(cudaMemset(d_w, 0, (size_t)W_REGION_STAGES * 16384));
    run_opt<5, 0, 0>(d_w, 109);
    run_opt<5, 0, 1>(d_w, 109);
    run_opt<5, 0, 2>(d_w, 109);
    run_opt<5, 0, 4>(d_w, 109);
    run_opt<5, 0, 8>(d_w, 109);
    run_opt<5, 0, 16>(d_w, 109);
    run_opt<5, 0, 31>(d_w, 109);
    run_opt<5, 0, 32 + 4 + 8 + 16>(d_w, 109);
    run_opt<5, 1, 0>(d_w, 109);
    run_opt<5, 1, 31>(d_w, 109);
    run_opt<5, 1, 32 + 4 + 8 + 16>(d_w, 109);
    return 0;
}
