// tcgen05.mma throughput microbenchmark for sm_100a: cycles per MMA for the operand sources / layouts / CTA-group
// modes the GEMM kernels can choose between.  Operand data is whatever is in shared memory (timing only).
// The issuing code mirrors the production kernels: warp 1, elect_one, fully unrolled groups of 16 MMAs whose
// descriptors are compile-time offsets from a uniform base (so ptxas emits straight UTCHMMA sequences).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/mma_bench tools/mma_bench.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include "../ubisoft-laforge-daft-exprt_b200/csrc/dx_ptx.cuh"

using namespace dx;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s -> %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

enum { A_NOSW = 0, A_SW128 = 1, A_TMEM = 2 };

__device__ __forceinline__ uint64_t desc_sw128(uint32_t addr) {
    uint64_t d = 0;
    d |= (uint64_t)((addr >> 4) & 0x3FFF);
    d |= (uint64_t)1 << 16;                    // LBO (ignored for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;          // SBO: 8 rows x 128 B
    d |= (uint64_t)1 << 46;                    // version
    d |= (uint64_t)2 << 61;                    // SWIZZLE_128B
    return d;
}
__device__ __forceinline__ void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, bool acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc),
        "r"((uint32_t)acc)
        : "memory");
}
__device__ __forceinline__ void mma_ss2(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, bool acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc),
        "r"((uint32_t)acc)
        : "memory");
}
__device__ __forceinline__ void commit2(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                     ptx::smem_u32(bar)), "h"((uint16_t)3)
                 : "memory");
}
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}

// CG = 1: one CTA, M = 128.   CG = 2: CTA pair, M = 256 (128 rows per CTA), B split N/2 per CTA.
template <int CG, int N, int AMODE, int BSW, int CE = 0, int DSW = 0>
__global__ void __launch_bounds__(128, 1) mma_bench_kernel(int ngroups, long long* out, int a_off, int a_lbo) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint64_t dummy_bar[4];
    __shared__ uint32_t tmem_ptr;
    const int warp = threadIdx.x >> 5;
    const uint32_t rank = CG == 2 ? cluster_rank() : 0;
    if (threadIdx.x == 0) {
        ptx::mbar_init(&bar, 1);
        for (int i = 0; i < 4; ++i) ptx::mbar_init(&dummy_bar[i], 1);
        ptx::fence_barrier_init();
    }
    if (warp == 0) {
        if (CG == 1) {
            ptx::tmem_alloc<512>(&tmem_ptr);
        } else {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(ptx::smem_u32(&tmem_ptr)), "n"(512) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync();
    ptx::tc_fence_after();
    const uint32_t tmem_base = tmem_ptr;
    const uint32_t sbase = ptx::smem_u32(smem);
    const uint32_t sA = sbase + (uint32_t)a_off, sB = sbase + 65536 + 4096;     // A: 4 stages x 16 KB, B: 4 stages x 32 KB
    constexpr int NB = CG == 2 ? N / 2 : N;            // B rows held by this CTA
    constexpr uint32_t idesc = ptx::make_idesc_bf16(CG == 2 ? 256 : 128, N);
    constexpr uint32_t acc_stride = N <= 128 ? 128u : (AMODE == A_TMEM ? 0u : 256u);
    if (warp == 1 && rank == 0) {
        long long t0 = 0, t1 = 0, t2 = 0;
        t0 = clock64();
        for (int g = 0; g < ngroups; ++g) {
            const uint32_t d = tmem_base + (uint32_t)(g & 1) * acc_stride;
            if (ptx::elect_one()) {
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int k = i & 3, st = i >> 2;
                    uint64_t da, db;
                    if (AMODE == A_SW128) da = desc_sw128(sA + st * 16384 + k * 32) | ((uint64_t)((a_off >> 7) & 7) << 49);
                    else da = ptx::make_desc(sA + st * 16384 + k * 2 * a_lbo, a_lbo, 128);
                    if (BSW) db = desc_sw128(sB + st * 32768 + k * 32);
                    else db = ptx::make_desc(sB + st * 32768 + k * 2 * NB * 16, NB * 16, 128);
                    const uint32_t dd = DSW ? d + (uint32_t)((i >> 2) & 1) * 256u : d;     // production: 4 MMAs per accumulator, then the other half's
                    if (CG == 2) mma_ss2(dd, da, db, idesc, DSW ? (i & 3) != 0 || i > 7 : i != 0);
                    else if (AMODE == A_TMEM) mma_ts(d, tmem_base + 384 + k * 8, db, idesc, i != 0);
                    else ptx::mma_bf16(dd, da, db, idesc, DSW ? (i & 3) != 0 || i > 7 : i != 0);
                    if (CE > 0 && (i + 1) % CE == 0) {
                        if (CG == 2) commit2(&dummy_bar[(i / CE) & 3]);
                        else ptx::mma_commit(&dummy_bar[(i / CE) & 3]);
                    }
                }
            }
            __syncwarp();
        }
        if (ptx::elect_one()) {
            if (CG == 2) commit2(&bar);
            else ptx::mma_commit(&bar);
        }
        __syncwarp();
        t1 = clock64();
        ptx::mbar_wait(&bar, 0);
        t2 = clock64();
        if ((threadIdx.x & 31) == 0) {
            out[blockIdx.x * 2] = t2 - t0;
            out[blockIdx.x * 2 + 1] = t1 - t0;
        }
    } else if (CG == 2 && warp == 1) {
        ptx::mbar_wait(&bar, 0);
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync();
    if (warp == 0) {
        ptx::tc_fence_after();
        if (CG == 1) ptx::tmem_dealloc<512>(tmem_base);
        else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
    }
}

template <int CG, int N, int AMODE, int BSW, int CE = 0, int DSW = 0>
static void run(const char* name, int grid, int a_off = 0, int a_lbo = 2048) {
    const int ngroups = 256, nmma = ngroups * 16;
    if (grid < CG) grid = CG;
    long long* d_out;
    CK(cudaMalloc(&d_out, grid * 2 * sizeof(long long)));
    CK(cudaMemset(d_out, 0, grid * 2 * sizeof(long long)));
    const size_t smem = 200 * 1024;
    auto kern = mma_bench_kernel<CG, N, AMODE, BSW, CE, DSW>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (int rep = 0; rep < 2; ++rep) {
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(grid);
        cfg.blockDim = dim3(128);
        cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = CG;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        CK(cudaLaunchKernelEx(&cfg, kern, ngroups, d_out, a_off, a_lbo));
        CK(cudaDeviceSynchronize());
    }
    std::vector<long long> h(grid * 2);
    CK(cudaMemcpy(h.data(), d_out, grid * 2 * sizeof(long long), cudaMemcpyDeviceToHost));
    double tot = 0, iss = 0;
    int n = 0;
    for (int i = 0; i < grid; i += CG) { tot += h[i * 2]; iss += h[i * 2 + 1]; ++n; }
    tot /= n; iss /= n;
    const double cyc = tot / nmma;
    const double macs = (double)(CG == 2 ? 256 : 128) * N * 16 / cyc / CG;    // MAC / clk / SM (peak 4096)
    printf("%-46s N=%3d grid=%3d  %7.1f cyc/MMA (issue %6.1f)  %6.0f MAC/clk/SM = %5.1f%% of 4096\n", name, N, grid, cyc,
           iss / nmma, macs, 100.0 * macs / 4096.0);
    fflush(stdout);
    CK(cudaFree(d_out));
}

template <int N>
static void run_n(int grid) {
    run<1, N, A_NOSW, 0>("cg1 SS  A no-swizzle  B no-swizzle", grid);
    run<1, N, A_SW128, 1>("cg1 SS  A SW128       B SW128", grid);
    run<1, N, A_TMEM, 0>("cg1 TS  A TMEM        B no-swizzle", grid);
    run<1, N, A_TMEM, 1>("cg1 TS  A TMEM        B SW128", grid);
    run<2, N, A_NOSW, 0>("cg2 SS  A no-swizzle  B no-swizzle (M=256)", grid);
    run<2, N, A_SW128, 1>("cg2 SS  A SW128       B SW128      (M=256)", grid);
}

// conv taps are row shifts of the A operand: +16 B per row in the no-swizzle K-major layout (core matrices then straddle
// 128-byte lines), +128 B per row in the SWIZZLE_128B layout (descriptor base_offset = row & 7)
static void run_shift() {
    char name[96];
    for (int lbo : {2048, 4096})
        for (int off : {0, 16, 32, 64, 112}) {
            snprintf(name, sizeof name, "cg1 SS A no-swizzle +%3d B LBO %d", off, lbo);
            run<1, 128, A_NOSW, 0>(name, 148, off, lbo);
            snprintf(name, sizeof name, "cg2 SS A no-swizzle +%3d B LBO %d", off, lbo);
            run<2, 128, A_NOSW, 0>(name, 148, off, lbo);
        }
    for (int off : {0, 128, 256, 896}) {
        snprintf(name, sizeof name, "cg1 SS A SW128 +%3d B (row shift)", off);
        run<1, 128, A_SW128, 1>(name, 148, off);
        snprintf(name, sizeof name, "cg2 SS A SW128 +%3d B (row shift)", off);
        run<2, 128, A_SW128, 1>(name, 148, off);
    }
}

// production issue pattern: a tcgen05.commit after every stage (8 MMAs), accumulator switch every 4 MMAs
template <int CE, int DSW>
static void run_commit1() {
    char name[96];
    snprintf(name, sizeof name, "cg1 SS commit every %2d MMAs, D switch %d", CE, DSW);
    run<1, 128, A_NOSW, 0, CE, DSW>(name, 148);
    snprintf(name, sizeof name, "cg2 SS commit every %2d MMAs, D switch %d", CE, DSW);
    run<2, 128, A_NOSW, 0, CE, DSW>(name, 148);
}
static void run_commit() {
    run_commit1<0, 0>(); run_commit1<16, 0>(); run_commit1<8, 0>(); run_commit1<4, 0>(); run_commit1<2, 0>(); run_commit1<1, 0>();
    run_commit1<0, 1>(); run_commit1<8, 1>(); run_commit1<4, 1>();
}

int main(int argc, char** argv) {
    if (argc > 1 && argv[1][0] == 'c') { run_commit(); return 0; }
    if (argc > 1) { run_shift(); return 0; }
    for (int grid : {2, 148}) {
        run_n<64>(grid);
        run_n<128>(grid);
        run_n<256>(grid);
    }
    return 0;
}
