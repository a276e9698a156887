// mma_pace.cu - microbenchmark: how fast does ONE thread per CTA (pair) get tcgen05.mma.kind::f16
// instructions through the tensor pipe, for the shapes the scan kernels use?  No loads, no epilogue:
// operands sit in shared memory (or TMEM for the A operand in the "TS" form) and are re-used.
// Prints SM cycles per MMA next to the arithmetic floor (M*N*16 MACs at 4096 MACs/clk/SM).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr
//        -I../django-assistant-bot_b200/csrc -I../include tools/mma_pace.cu -o tools/_build/mma_pace -lcuda
// Not part of the product; numbers quoted in DESIGN.md section 4.3b come from here.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "umma.cuh"

namespace dab {
// error plumbing the headers expect (the library's own lives in store.cu)
int fail(int code, const std::string &msg) {
    fprintf(stderr, "error %d: %s\n", code, msg.c_str());
    return code;
}
}  // namespace dab

using namespace dab;
using namespace dab::umma;

__device__ __forceinline__ void mma_ts_1cta(uint32_t d, uint32_t a, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void mma_ts_2cta(uint32_t d, uint32_t a, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(db), "r"(idesc), "r"(acc) : "memory");
}

// one lane of a converged warp; ptxas knows that exactly one thread runs the guarded region, so a
// tcgen05 instruction inside it is emitted once (after `lane == 0` it is wrapped in a loop over the active lanes)
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}

struct Result {
    long long cycles;
    long long mmas;
};

// one "k-block" = 4 MMAs (K = 64 bf16 = one 128-byte swizzle atom row); a_stages A buffers are cycled
template <bool CTA2, bool TS, bool UB>
__global__ void __launch_bounds__(128, 1)
pace_kernel(int n, int kblocks, int a_stages, int commit_every, int d_stages, Result *out) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    unsigned char *a_smem = smem;                                   // [a_stages][16 KiB]
    unsigned char *b_smem = a_smem + (size_t)a_stages * kAStageBytes;  // [n rows (n/2 for CTA2) x 128 B]
    const int b_rows = CTA2 ? n / 2 : n;
    uint64_t *bar = reinterpret_cast<uint64_t *>(b_smem + (size_t)(TS ? 4 * b_rows : b_rows) * 128);
    uint64_t *bar2 = bar + 1;
    uint32_t *tmem_base_s = reinterpret_cast<uint32_t *>(bar + 2);
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (int)((unsigned char *)bar - smem) / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0x3c003c00u;
    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_init(bar2, 1);
        fence_barrier_init();
    }
    if (warp == 0) {
        if (CTA2) tmem_alloc2(tmem_base_s, 512u); else tmem_alloc(tmem_base_s, 512u);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    if (CTA2) cluster_sync_all(); else __syncthreads();
    tc_fence_after();
    // the tcgen05 instructions take their operands from UNIFORM registers: a value the compiler cannot
    // prove warp-uniform (a shared-memory load) costs a ELECT/R2UR.BROADCAST/branch loop per instruction.
    // A warp reduction's result is uniform by construction.
    const uint32_t tmem_base = UB ? __reduce_or_sync(0xffffffffu, *tmem_base_s) : *tmem_base_s;
    const bool leader = !CTA2 || (blockIdx.x & 1) == 0;
    if (warp == 1 && leader && (UB ? elect_one() : (tid & 31) == 0)) {
        const int m = CTA2 ? 256 : 128;
        const uint32_t idesc = make_idesc(n, m);
        const uint32_t tmem_a = tmem_base + 480;   // 32 columns of A (TS form): 4 K-steps x 8 columns
        const long long t0 = clock64();
        int s = 0, d = 0;
        for (int kb = 0; kb < kblocks; kb++) {
            const uint32_t tmem_d = tmem_base + (uint32_t)(d * n);
            const uint64_t da = make_smem_desc(smem_u32(a_smem + (size_t)s * kAStageBytes));
            const uint64_t db = make_smem_desc(smem_u32(b_smem));
#pragma unroll
            for (int k = 0; k < 4; k++) {
                if constexpr (TS) {
                    // A in TMEM, B = [n rows] in smem; use the A ring as extra B k-blocks so smem addresses vary
                    const uint64_t dbb = make_smem_desc(smem_u32(b_smem + (size_t)(kb & 3) * b_rows * 128));
                    if constexpr (CTA2) mma_ts_2cta(tmem_d, tmem_a + 8 * k, dbb + (uint64_t)(2 * k), idesc, (uint32_t)(kb | k));
                    else mma_ts_1cta(tmem_d, tmem_a + 8 * k, dbb + (uint64_t)(2 * k), idesc, (uint32_t)(kb | k));
                } else {
                    if constexpr (CTA2) umma_f16_2cta(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)(kb | k));
                    else umma_f16(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)(kb | k));
                }
            }
            if (commit_every > 0 && kb % commit_every == commit_every - 1) {
                if constexpr (CTA2) umma_commit_mc2(bar2, (uint16_t)0x3u); else umma_commit(bar2);
            }
            if (++s == a_stages) s = 0;
            if ((kb & 15) == 15 && ++d == d_stages) d = 0;
        }
        if constexpr (CTA2) umma_commit_mc2(bar, (uint16_t)0x1u); else umma_commit(bar);
        mbar_wait_bounded(bar, 0);
        const long long t1 = clock64();
        if (blockIdx.x == 0) {
            out->cycles = t1 - t0;
            out->mmas = 4ll * kblocks;
        }
    }
    tc_fence_before();
    if (CTA2) cluster_sync_all(); else __syncthreads();
    if (warp == 0) {
        if (CTA2) tmem_dealloc2(tmem_base, 512u); else tmem_dealloc(tmem_base, 512u);
    }
}

template <bool CTA2, bool TS, bool UB = true>
void run(int n, int kblocks, int a_stages, int commit_every, int grid) {
    const int uniform_base = UB;
    Result *out;
    cudaMalloc(&out, sizeof(Result));
    cudaMemset(out, 0, sizeof(Result));
    const int d_stages = 448 / n >= 2 ? 2 : 1;
    const size_t smem = 1024 + (size_t)a_stages * kAStageBytes + (size_t)(TS ? 4 : 1) * (CTA2 ? n / 2 : n) * 128 + 64;
    auto kern = pace_kernel<CTA2, TS, UB>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(128);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CTA2 ? 2 : 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float ms = 0;
    for (int rep = 0; rep < 2; rep++) {
        cudaEventRecord(e0);
        cudaError_t err = cudaLaunchKernelEx(&cfg, kern, n, kblocks, a_stages, commit_every, d_stages, out);
        cudaEventRecord(e1);
        cudaError_t err2 = cudaDeviceSynchronize();
        if (err != cudaSuccess || err2 != cudaSuccess) {
            printf("cta%d %s N=%3d: FAILED %s / %s\n", CTA2 ? 2 : 1, TS ? "TS" : "SS", n, cudaGetErrorString(err), cudaGetErrorString(err2));
            exit(1);
        }
        cudaEventElapsedTime(&ms, e0, e1);
    }
    Result r;
    cudaMemcpy(&r, out, sizeof(r), cudaMemcpyDeviceToHost);
    const int m = CTA2 ? 256 : 128;
    const double floor_clk = (double)m * n * 16 / (4096.0 * (CTA2 ? 2 : 1));
    const double per = (double)r.cycles / (double)r.mmas;
    const double tflops = 2.0 * m * n * 16 * (double)r.mmas * (grid / (CTA2 ? 2 : 1)) / (ms * 1e-3) / 1e12;
    printf("%s cta%d %s M=%d N=%3d a_stages=%d commit_every=%2d grid=%3d: %.1f clk/MMA (floor %.0f, %.0f%%)  %.3f ms  ~%.2f GHz  %.0f TF/s\n",
           uniform_base ? "uniform" : "lds-base", CTA2 ? 2 : 1, TS ? "TS" : "SS", m, n, a_stages, commit_every, grid, per, floor_clk, 100.0 * floor_clk / per, ms,
           (double)r.cycles / (ms * 1e-3) / 1e9, tflops);
    cudaFree(out);
}

int main(int argc, char **argv) {
    const int kblocks = argc > 1 ? atoi(argv[1]) : 20000;
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int grid = sms / 2 * 2;
    for (int ce : {1}) {
        for (int n : {64, 128, 256}) run<false, false>(n, kblocks, 4, ce, grid);
        for (int n : {64, 128, 256}) run<true, false>(n, kblocks, 4, ce, grid);
        for (int n : {64, 128, 256}) run<false, true>(n, kblocks, 4, ce, grid);
        for (int n : {64, 128, 256}) run<true, true>(n, kblocks, 4, ce, grid);
    }
    // one CTA (pair) alone: no power / clock effects from the rest of the chip
    run<true, false, false>(128, kblocks, 4, 1, grid);
    run<false, false, false>(64, kblocks, 4, 1, grid);
    return 0;
}
