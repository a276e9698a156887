// scan_umma_qa.cu - tensor-core scoring for WIDE query blocks with the QUERIES as the A operand in TMEM.
//
// The other tensor kernels keep the bf16 query block in shared memory as the B operand, which is what
// limits them: ~96 queries per SM (scan_umma), or 256 queries over four SMs at the price of delivering
// every corpus tile to two SMs (scan_umma_quad) - and SM ingest, not HBM, then becomes the ceiling.
// Here the roles are swapped:
//
//   A = queries   M = 256 (two CTAs x 128 lanes), resident in TENSOR MEMORY: lane = query, 32-bit column c
//                 holds dims 2c, 2c+1 (bf16 pair) - 384 of the 512 TMEM columns at dim 768.  Written once
//                 per launch with tcgen05.st by the epilogue warps.
//   B = corpus    N = 64 rows per step, K-major 128-byte-swizzled tiles streamed by TMA; with
//                 cta_group::2 each CTA loads only HALF of the rows (32) into its own shared memory and
//                 the pair's tensor cores share them - so every corpus byte is read from HBM once AND
//                 delivered to exactly one SM, while both SMs score it against their 128 queries.
//   D             [128 queries x 64 rows] fp32 per CTA, two accumulator stages in the remaining 128 columns.
//
// Shared memory holds nothing but the corpus ring (16 stages x 8 KiB per CTA), so the pipeline is as deep
// as HBM latency needs.  In the epilogue a thread IS a query: its admission threshold lives in a
// register, the rows' inverse norms are broadcast from shared memory, and survivors go through the
// shared-memory staging list to the flusher warp (as in scan_umma_quad).
// tcgen05.mma with A in TMEM ("TS" form): tcgen05.mma.cta_group::2.kind::f16 [d], [a], b_desc, idesc, p.
#include <stdio.h>
#include <stdlib.h>

#include "umma.cuh"

namespace dab {

namespace {

using namespace umma;
constexpr int kThreads = 256;
constexpr int kStepRows = 64;                 // corpus rows per accumulator (UMMA N)
constexpr int kHalfRows = kStepRows / 2;      // rows each CTA of the pair loads
constexpr int kKbPerStage = 2;                // k-blocks (64 dims each) per shared-memory stage
constexpr int kSubBytes = kHalfRows * kBlockK * 2;          // 4 KiB: 32 rows x 128 B
constexpr int kStageBytes = kKbPerStage * kSubBytes;        // 8 KiB per CTA per stage
constexpr int kMaxStages = 20;
constexpr int kEpiWarps = 4;
constexpr int kWarpCap = 128;                 // staged candidates per epilogue warp per step
constexpr int kPendCap = kEpiWarps * kWarpCap;
constexpr int kAccCols = kStepRows;           // fp32 accumulator columns per stage
constexpr int kSmemBudget = 227 * 1024 - 2048;

__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
        ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
          "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// D[tmem] (+)= A[tmem] * B[smem], both CTAs of the pair
__device__ __forceinline__ void umma_f16_ts_2cta(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__global__ void __launch_bounds__(kThreads, 1)
scan_umma_qa_kernel(const __grid_constant__ CUtensorMap tmap_rows, StoreView sv, const uint32_t *__restrict__ allow,
                    SearchScratch sc, int q0_all, int nq_all, int stages, const __grid_constant__ LevelPlan plan) {
    extern __shared__ unsigned char smem_raw[];
    const int lvl = (int)warp_uniform((uint32_t)sc.ctrl->cur_level);
    if (lvl >= plan.num_levels) return;
    const long long ntiles = plan.tiles_in_level[lvl];  // steps of 64 rows
    const int pair_id = blockIdx.x / 2;
    const int num_pairs = gridDim.x / 2;
    if ((long long)pair_id >= ntiles) return;           // same decision in both CTAs of the pair

    const uint32_t rank = warp_uniform(cluster_ctarank());   // 0 = leader
    const bool leader = rank == 0;
    const int num_kb = sv.dim / kBlockK;
    const int stages_per_step = num_kb / kKbPerStage;
    const int a_cols = sv.dim / 2;                      // TMEM columns of the query operand

    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    unsigned char *b_smem = smem;                                                   // [stages][2][32 rows x 128 B]
    float *s_inv = reinterpret_cast<float *>(b_smem + (size_t)stages * kStageBytes);   // [2][64] inverse norms per accumulator stage
    uint64_t *full = reinterpret_cast<uint64_t *>(s_inv + 2 * kStepRows);           // leader: both halves of the stage landed
    uint64_t *empty = full + kMaxStages;               // each CTA: the pair's MMAs are done with my half
    uint64_t *a_ready = empty + kMaxStages;            // leader: all 8 epilogue warps of the pair wrote their queries
    uint64_t *tmem_full = a_ready + 1;                 // [2]
    uint64_t *tmem_empty = tmem_full + 2;              // [2] leader: both CTAs drained the accumulator stage
    uint64_t *pend_full = tmem_empty + 2;              // [2]
    uint64_t *pend_free = pend_full + 2;               // [2]
    uint64_t *pend_key = pend_free + 2;                // [2][kPendCap]
    uint32_t *tmem_base_s = reinterpret_cast<uint32_t *>(pend_key + 2 * kPendCap);
    uint32_t *pend_n = tmem_base_s + 1;                // [2][kEpiWarps]
    uint32_t *pend_q = pend_n + 2 * kEpiWarps;         // [2][kPendCap]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < stages; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        mbar_init(a_ready, 8);
        for (int a = 0; a < 2; a++) {
            mbar_init(&tmem_full[a], 1);
            mbar_init(&tmem_empty[a], 8);
            mbar_init(&pend_full[a], 4);
            mbar_init(&pend_free[a], 1);
        }
        fence_barrier_init();
        prefetch_tmap(&tmap_rows);
    }
    if (warp == 2) tmem_alloc2(tmem_base_s, 512u);
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = warp_uniform(*tmem_base_s);
    const uint32_t tmem_acc0 = tmem_base + (uint32_t)a_cols;   // accumulators live after the query columns
    const long long my_steps = (ntiles - pair_id + num_pairs - 1) / num_pairs;

    if (warp == 0) {
        if (elect_one()) {
            // ------------------------------------------------------------ TMA producer: my 32 rows of every step
            int s = 0;
            uint32_t ph = 0;
            for (long long it = 0; it < my_steps; it++) {
                const long long t = level_tile(plan, lvl, pair_id + it * num_pairs);
                const int row0 = (int)(t * kStepRows) + (int)rank * kHalfRows;
                for (int st = 0; st < stages_per_step; st++) {
                    mbar_wait_bounded(&empty[s], ph ^ 1u);
                    unsigned char *dst = b_smem + (size_t)s * kStageBytes;
#pragma unroll
                    for (int kk = 0; kk < kKbPerStage; kk++)
                        tma_load_2d_2cta(dst + kk * kSubBytes, &tmap_rows, (st * kKbPerStage + kk) * kBlockK, row0, &full[s]);
                    if (++s == stages) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        if (leader && elect_one()) {
            // ------------------------------------------------------------ MMA issuer of the pair
            const uint32_t idesc = make_idesc(kStepRows, 2 * kTileRows);   // N = 64 corpus rows, M = 256 queries
            mbar_wait_bounded(a_ready, 0);
            tc_fence_after();
            int s = 0;
            uint32_t ph = 0;
            for (long long it = 0; it < my_steps; it++) {
                const int a = (int)(it & 1);
                const uint32_t aph = (uint32_t)((it >> 1) & 1);
                mbar_wait_bounded(&tmem_empty[a], aph ^ 1u);
                tc_fence_after();
                const uint32_t tmem_d = tmem_acc0 + (uint32_t)(a * kAccCols);
                for (int st = 0; st < stages_per_step; st++) {
                    mbar_arrive_expect_tx(&full[s], 2 * kStageBytes);      // my half + my peer's half
                    mbar_wait_bounded(&full[s], ph);
                    tc_fence_after();
                    const uint32_t sbase = smem_u32(b_smem + (size_t)s * kStageBytes);
#pragma unroll
                    for (int kk = 0; kk < kKbPerStage; kk++) {
                        const uint64_t db = make_smem_desc(sbase + kk * kSubBytes);
                        const int kb = st * kKbPerStage + kk;
#pragma unroll
                        for (int k = 0; k < kBlockK / kUmmaK; k++) {
                            // A: 16 dims = 8 TMEM columns per K step
                            const uint32_t ta = tmem_base + (uint32_t)((kb * (kBlockK / kUmmaK) + k) * (kUmmaK / 2));
                            umma_f16_ts_2cta(tmem_d, ta, db + (uint64_t)(2 * k), idesc, (kb | k) != 0 ? 1u : 0u);
                        }
                    }
                    umma_commit_mc2(&empty[s], (uint16_t)0x3u);
                    if (++s == stages) { s = 0; ph ^= 1u; }
                }
                umma_commit_mc2(&tmem_full[a], (uint16_t)0x3u);
            }
        }
    } else if (warp == 3) {
        // ------------------------------------------------------------ flusher: staged candidates -> global buffers
        for (long long it = 0; it < my_steps; it++) {
            const int pl = (int)(it & 1);
            const uint32_t plph = (uint32_t)((it >> 1) & 1);
            mbar_wait_bounded(&pend_full[pl], plph);
            flush_warp_lists(pend_key + pl * kPendCap, pend_q + pl * kPendCap, pend_n + pl * kEpiWarps, kEpiWarps, kWarpCap, allow, sc);
            __syncwarp();
            if (lane == 0) mbar_arrive(&pend_free[pl]);
        }
    } else if (warp >= 4) {
        const int wq = warp & 3;                               // TMEM lane quarter this warp may touch
        const int et = tid - 128;                              // 0..127 within the epilogue warps
        const int qloc = wq * 32 + lane;                       // my query within this CTA (= my TMEM lane)
        const int q = q0_all + (int)rank * kTileRows + qloc;   // global query index
        const bool live = (int)rank * kTileRows + qloc < nq_all && sc.active[q];
        const float tau = live ? key_score(sc.tau_key[q]) : INFINITY;
        // ---- my query's bf16 row -> TMEM lane (once per launch): column c = dims 2c, 2c+1
        {
            const uint4 *src = reinterpret_cast<const uint4 *>(sc.qsplit + (size_t)q * sv.dim);
            const uint32_t tl = tmem_base + ((uint32_t)(wq * 32) << 16);
            for (int c0 = 0; c0 < a_cols; c0 += 16) {
                uint32_t v[16];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const uint4 x = src[c0 / 4 + j];
                    v[4 * j + 0] = x.x; v[4 * j + 1] = x.y; v[4 * j + 2] = x.z; v[4 * j + 3] = x.w;
                }
                tmem_st16(tl + (uint32_t)c0, v);
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (leader) mbar_arrive(a_ready); else mbar_arrive_remote(a_ready, 0);
            }
        }
        // ---- epilogue: a thread is a query; columns are corpus rows
        auto step_row0 = [&](long long it_) { return level_tile(plan, lvl, pair_id + it_ * num_pairs) * kStepRows; };
        if (et < kStepRows) {   // inverse norms of the first step
            const long long r = step_row0(0) + et;
            const float x = r < sv.n ? sv.inv_norm[r] : 0.f;
            s_inv[et] = x > 0.f ? x : __int_as_float(0x7fc00000);   // zero rows / rows past the end: NaN, never >= tau
        }
        for (long long it = 0; it < my_steps; it++) {
            const int a = (int)(it & 1);
            const uint32_t aph = (uint32_t)((it >> 1) & 1);
            const long long row0 = step_row0(it);
            float inv_next = 0.f;
            if (et < kStepRows && it + 1 < my_steps) {          // fetched now, stored after this step is drained
                const long long r = step_row0(it + 1) + et;
                const float x = r < sv.n ? sv.inv_norm[r] : 0.f;
                inv_next = x > 0.f ? x : __int_as_float(0x7fc00000);
            }
            named_bar_sync(1, 128);                             // s_inv[a] written by all epilogue threads
            mbar_wait_bounded(&pend_free[a], aph ^ 1u);
            mbar_wait_bounded(&tmem_full[a], aph);
            tc_fence_after();
            const uint32_t taddr = tmem_acc0 + ((uint32_t)(wq * 32) << 16) + (uint32_t)(a * kAccCols);
            const float *inv = s_inv + a * kStepRows;
            uint32_t v0[16], v1[16];
            uint32_t hits = 0;
            const uint32_t lt = (1u << lane) - 1u;
            uint64_t *wkey = pend_key + a * kPendCap + wq * kWarpCap;
            uint32_t *wq_ = pend_q + a * kPendCap + wq * kWarpCap;
            tmem_ld16(taddr, v0);
#pragma unroll
            for (int c0 = 0; c0 < kStepRows; c0 += 32) {
                tmem_ld_wait();
                tmem_ld16(taddr + (uint32_t)(c0 + 16), v1);
#pragma unroll
                for (int half = 0; half < 2; half++) {
                    const uint32_t(&v)[16] = half ? v1 : v0;
                    const int cb = c0 + 16 * half;
                    if (half) {
                        tmem_ld_wait();
                        if (c0 + 32 < kStepRows) tmem_ld16(taddr + (uint32_t)(c0 + 32), v0);
                    }
                    uint32_t pass = 0;
#pragma unroll
                    for (int j = 0; j < 16; j += 4) {
                        const float4 iv = *reinterpret_cast<const float4 *>(inv + cb + j);
                        pass |= (__uint_as_float(v[j + 0]) * iv.x >= tau ? 1u : 0u) << (j + 0);
                        pass |= (__uint_as_float(v[j + 1]) * iv.y >= tau ? 1u : 0u) << (j + 1);
                        pass |= (__uint_as_float(v[j + 2]) * iv.z >= tau ? 1u : 0u) << (j + 2);
                        pass |= (__uint_as_float(v[j + 3]) * iv.w >= tau ? 1u : 0u) << (j + 3);
                    }
                    // hits: slot = the warp's running count + hitting lanes below me (no atomics, warp converged)
                    while (__any_sync(0xffffffffu, pass != 0)) {
                        const bool has = pass != 0;
                        const uint32_t b = __ballot_sync(0xffffffffu, has);
                        if (has) {
                            const int j = __ffs(pass) - 1;
                            pass &= pass - 1;
                            const uint64_t key = make_key(__uint_as_float(pick16(v, j)) * inv[cb + j], (uint32_t)(row0 + cb + j));
                            const uint32_t slot = hits + __popc(b & lt);
                            if (slot < (uint32_t)kWarpCap) {
                                wkey[slot] = key;
                                wq_[slot] = (uint32_t)q;
                            } else {
                                append_candidate(sc, allow, q, key);
                            }
                        }
                        hits += __popc(b);
                    }
                }
            }
            tc_fence_before();
            if (et < kStepRows) s_inv[(a ^ 1) * kStepRows + et] = inv_next;   // for the next step (other stage)
            __syncwarp();
            if (lane == 0) {
                if (leader) mbar_arrive(&tmem_empty[a]); else mbar_arrive_remote(&tmem_empty[a], 0);
                pend_n[a * kEpiWarps + wq] = hits;
                mbar_arrive(&pend_full[a]);
            }
        }
    }
    // ---------------------------------------------------------------- teardown
    tc_fence_before();
    cluster_sync_all();
    if (warp == 2) tmem_dealloc2(tmem_base, 512u);
}

struct QaPlan {
    int stages;
    size_t smem;
};

bool plan_qa(int dim, QaPlan *out) {
    if (dim % (kBlockK * kKbPerStage) || dim / 2 + 2 * kAccCols > 512) return false;   // queries + 2 accumulators must fit TMEM
    const size_t fixed = 2 * kStepRows * sizeof(float) + (2 * kMaxStages + 16) * sizeof(uint64_t) + 16 +
                         2 * kPendCap * (sizeof(uint64_t) + sizeof(uint32_t)) + 2 * kEpiWarps * sizeof(uint32_t) + 1024;
    int stages = (int)(((size_t)kSmemBudget - fixed) / kStageBytes);
    if (stages > kMaxStages) stages = kMaxStages;
    if (stages < 4) return false;
    out->stages = stages;
    out->smem = (size_t)stages * kStageBytes + fixed;
    return true;
}

}  // namespace

int qa_tile_rows() { return kStepRows; }

bool qa_supported(int dim, int dtype, int nq) {
    QaPlan p;
    return dtype == DAB_BF16 && nq >= 1 && nq <= 2 * kTileRows && plan_qa(dim, &p);
}

int launch_scan_umma_qa(const StoreView &sv, const uint32_t *allow, const SearchScratch &sc, int q0, int nq, const LevelPlan &plan,
                        int num_sms, cudaStream_t st) {
    QaPlan qp;
    if (sv.dtype != DAB_BF16 || !plan_qa(sv.dim, &qp) || nq < 1 || nq > 2 * kTileRows)
        return fail(DAB_E_INVALID, "scan_umma_qa: needs a bf16 store, dim % 128 == 0, dim <= 768 and <= 256 queries");
    if (plan.tile_rows != kStepRows) return fail(DAB_E_INVALID, "scan_umma_qa: level plan built for another tile size");
    alignas(64) CUtensorMap tmap_rows;
    if (int rc = cached_map(&tmap_rows, sv.rows, (uint64_t)sv.n, (uint64_t)sv.dim, kHalfRows)) return rc;
    static SmemAttrOnce attr_once;
    static ClusterOccupancyCache occ_cache;
    DAB_CUDA_OK(attr_once.ensure(scan_umma_qa_kernel, (int)qp.smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(num_sms / 2 * 2), 1, 1);
    cfg.blockDim = dim3(kThreads, 1, 1);
    cfg.dynamicSmemBytes = qp.smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int max_clusters = 0;
    DAB_CUDA_OK(occ_cache.get(&max_clusters, scan_umma_qa_kernel, &cfg, 2));
    if (max_clusters < 1) return fail(DAB_E_CUDA, "scan_umma_qa: no CTA pair fits on this device");
    long long max_tiles = 0;
    for (int l = 0; l < plan.num_levels; l++) max_tiles = plan.tiles_in_level[l] > max_tiles ? plan.tiles_in_level[l] : max_tiles;
    const long long pairs = max_tiles < max_clusters ? max_tiles : max_clusters;
    cfg.gridDim = dim3((unsigned)(pairs * 2), 1, 1);
    static const bool debug = tuning_env("DAB_DEBUG") != nullptr;
    if (debug) fprintf(stderr, "[dab] qa: nq=%d stages=%d smem=%zu pairs=%lld levels=%d\n", nq, qp.stages, qp.smem, pairs, plan.num_levels);
    DAB_CUDA_OK(cudaLaunchKernelEx(&cfg, scan_umma_qa_kernel, tmap_rows, sv, allow, sc, q0, nq, qp.stages, plan));
    count_launch();
    return DAB_OK;
}

}  // namespace dab
