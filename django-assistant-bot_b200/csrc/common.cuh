// common.cuh - shared device helpers for libdab_b200 (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <string>

#include "../../include/dab_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libdab_b200 is written for sm_100a (B200) only"
#endif

namespace dab {

// ---------------------------------------------------------------- host-side error plumbing
void set_error(const std::string &msg);
int fail(int code, const std::string &msg);
#define DAB_CUDA_OK(expr)                                                                      \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess)                                                                 \
            return ::dab::fail(DAB_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
    } while (0)

// every kernel launch of the library is counted (bench.py reports it as gpu_launches)
void count_launch();
#define DAB_LAUNCH_OK()                      \
    do {                                     \
        ::dab::count_launch();               \
        DAB_CUDA_OK(cudaGetLastError());     \
    } while (0)

// Tuning switches (environment variables) exist only in builds made with -DDAB_TUNING (tools/); the product
// library reads no environment variable: the compiler sees a constant nullptr and drops the branches.
#ifdef DAB_TUNING
#include <stdlib.h>
static inline const char *tuning_env(const char *name) { return getenv(name); }
#else
static inline const char *tuning_env(const char *) { return nullptr; }
#endif

// cudaFuncAttributeMaxDynamicSharedMemorySize, raised once per (kernel, device) to the largest size asked so far.
struct SmemAttrOnce {
    int set[64] = {};
    std::mutex mu;
    template <typename K>
    cudaError_t ensure(K kernel, int bytes) {
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return e;
        if (dev < 0 || dev >= 64) return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        std::lock_guard<std::mutex> g(mu);
        if (set[dev] >= bytes) return cudaSuccess;
        e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        if (e == cudaSuccess) set[dev] = bytes;
        return e;
    }
};

// Where a producer kernel writes the per-row store columns when its output rows land directly in a dab_store.
struct StoreSink {
    float *inv_norm = nullptr;   // [B] 1/|row| from the stored values (nullptr: the output is not a store)
    int64_t *ids = nullptr;      // [B]
    int64_t id_first = 0;
};

static inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---------------------------------------------------------------- candidate keys
// One 64-bit key orders candidates by (score desc, row index asc): larger key = better.
//   high 32 bits: order-preserving image of the fp32 score; low 32 bits: ~row
__host__ __device__ __forceinline__ uint32_t f32_orderable(float s) {
#ifdef __CUDA_ARCH__
    uint32_t u = __float_as_uint(s);
#else
    union { float f; uint32_t u; } c; c.f = s; uint32_t u = c.u;
#endif
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float orderable_f32(uint32_t o) {
    uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}
__host__ __device__ __forceinline__ uint64_t make_key(float score, uint32_t row) {
    return (static_cast<uint64_t>(f32_orderable(score)) << 32) | static_cast<uint64_t>(~row);
}
__host__ __device__ __forceinline__ float key_score(uint64_t key) { return orderable_f32(static_cast<uint32_t>(key >> 32)); }
__host__ __device__ __forceinline__ uint32_t key_row(uint64_t key) { return ~static_cast<uint32_t>(key); }

// ---------------------------------------------------------------- per-search device state
constexpr int kCandCap = 8192;      // candidate slots per query
constexpr int kMaxLevels = 8;

// Sampling levels of the threshold scan (see DESIGN.md "levels"): level l visits the tiles
// t with t % stride[l] == 0 that no earlier level visited.
struct LevelPlan {
    int num_levels;
    int growth;                      // stride[l] = growth^(num_levels-1-l)
    int stride[kMaxLevels];
    long long tiles_in_level[kMaxLevels];
    long long num_tiles;             // ceil(n_rows / tile_rows)
    int tile_rows;
};

struct SearchCtrl {
    int cur_level;                   // next level to scan; == num_levels when the scan is complete
    int retry;                       // 1 = cur_level is being re-scanned for the overflowed queries only
    unsigned int ticket;             // compact kernel: CTAs done
    int any_overflow;                // set by compact CTAs of this round
    int rounds_run;
    int overflow_events;             // statistics
    int inexact;                     // tensor path: more rows within the rescoring margin than can be kept
    int sticky_incomplete;           // some chunk of this search call ended before its last level
    int sticky_inexact;              // some chunk of this search call raised `inexact`
    int pad[3];
};

// ---------------------------------------------------------------- PTX: mbarrier + bulk async copy (TMA)
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
// 1-D bulk copy global -> shared through the TMA unit (SASS: UBLKCP); bytes % 16 == 0.
__device__ __forceinline__ void tma_load_1d(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// Single-thread roles (TMA producer, MMA issuer) must be entered through elect_one(), not `lane == 0`:
// tcgen05.mma / tcgen05.commit / cp.async.bulk.tensor execute once per WARP and take their operands from
// uniform registers.  After `lane == 0` ptxas cannot know how many threads are active and wraps every such
// instruction in an ELECT / R2UR.BROADCAST / branch loop - measured at ~113 cycles per tcgen05.mma instead
// of the 64-cycle floor of a 256x128x16 MMA (tools/mma_pace.cu).  elect.sync tells it exactly one thread runs.
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
// ---------------------------------------------------------------- programmatic dependent launch
// A search is a chain of short dependent kernels (begin, then scan + compact per level, then finish).  Launched
// with cudaLaunchAttributeProgrammaticStreamSerialization, kernel N+1 is placed on the SMs while kernel N still
// runs and parks at pdl_wait() - which returns when kernel N has completed and its writes are visible - so the
// hand-over costs no launch latency.  Every kernel of the chain calls pdl_enter() before it touches global memory.
__device__ __forceinline__ void pdl_enter() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
#ifdef __CUDACC__
template <typename... KArgs, typename... Args>
inline cudaError_t launch_chained(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#endif

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float bf16_lo(uint32_t packed) { return __uint_as_float(packed << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t packed) { return __uint_as_float(packed & 0xffff0000u); }

}  // namespace dab
