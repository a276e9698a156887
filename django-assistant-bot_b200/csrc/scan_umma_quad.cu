// scan_umma_quad.cu - tensor-core scoring for WIDE query blocks (up to 256 queries in one corpus pass).
//
// Why a second kernel.  scan_umma.cu keeps the bf16 query block resident in one SM's shared memory,
// which caps it at ~96 queries for dim 768; wider blocks had to be cut into slices that each re-read
// the corpus tiles through L2, and L2 (~6.9 TB/s measured) is no faster than HBM on this chip.  Here a
// thread-block CLUSTER of 4 CTAs shares ONE stream of corpus tiles:
//
//   * the query block is cut in two halves, one per CTA PAIR; inside a pair the two SMs run
//     tcgen05.mma.cta_group::2 (M = 256 corpus rows, N = half the queries): each SM holds only a
//     QUARTER of the query block in shared memory (98 KiB at dim 768, Q = 256) and 128 of the 256
//     corpus rows of a super-tile; the accumulator halves live in the two SMs' TMEM;
//   * both pairs need the same corpus rows, so the A tiles are loaded once per cluster and
//     TMA-MULTICAST to the matching CTA of the other pair (cp.async.bulk.tensor ... .multicast::cluster,
//     .cta_group::2 so that the byte counts land on the pair leader's mbarrier);
//   * tcgen05.commit ... .multicast::cluster releases the shared-memory stage in the loading CTAs and
//     publishes the accumulator to both CTAs of the pair; epilogue warps of all four CTAs drain
//     their own TMEM half (threshold test, rare appends) while the next super-tile's MMAs run.
//
// HBM and L2 traffic = the corpus bytes, once per launch.  Roles per CTA (12 warps):
//   warp 0, one elected lane   TMA producer (A tiles: ranks 0 and 1 only; B quarter: every rank)
//   warp 1, one elected lane   MMA issuer (pair leaders = even ranks only)
//   warp 2                     TMEM allocator (cta_group::2, executed in both CTAs of a pair)
//   warp 3                     tells the pair leader when this CTA's query quarter has landed, then
//                              flushes the epilogue warps' staged candidates to the global buffers
//   warps 4-11                 epilogue: two warps per TMEM lane quarter, half of the pair's columns each
// The single-lane roles are entered through elect_one() and every tcgen05/TMA operand is warp-uniform
// (umma.cuh explains what it costs otherwise).
#include <stdio.h>
#include <stdlib.h>

#include "umma.cuh"

namespace dab {

namespace {

using namespace umma;
constexpr int kThreads = 384;         // warps 0-3: roles; warps 4-11: epilogue (two per TMEM lane quarter)
// 128-byte k-blocks per shared-memory stage.  Measured on a 12.5M x 768 shard, Q = 256, 300 searches back to back:
// one k-block per stage (seven 16 KiB stages) 5.22 ms; two per stage (three 32 KiB stages, half the barrier rounds
// of the issuer) 5.37 ms; arming a stage's barrier one use ahead of its wait 5.37 ms.  The issuer's barrier
// round is not what bounds this kernel; the finer ring is.
constexpr int kKbPerStage = 1;
constexpr int kStageBytes = kKbPerStage * kAStageBytes;
constexpr int kMaxStages = 12;
constexpr int kMinStages = kKbPerStage == 1 ? 3 : 2;
constexpr int kMaxAcc = 4;         // TMEM accumulator stages (512 columns / n_pair, at most 4)
constexpr int kEpiWarps = 8;
constexpr int kWarpCap = 64;        // staged candidates per epilogue warp per tile (more go straight to the candidate buffer)
constexpr int kPendCap = kEpiWarps * kWarpCap;
constexpr int kSuperRows = 2 * kTileRows;  // corpus rows per MMA pair step (UMMA M = 256)
constexpr int kSmemBudget = 227 * 1024 - 2048;
template <bool F32>   // F32: fp32 store and fp32 queries through kind::tf32 (a k-block is 128 bytes = 32 elements)
__global__ void __launch_bounds__(kThreads, 1)
scan_umma_quad_kernel(const __grid_constant__ CUtensorMap tmap_rows, const __grid_constant__ CUtensorMap tmap_q, StoreView sv,
                      const uint32_t *__restrict__ allow, SearchScratch sc, int q0_all, int nq_all, int n_pair, int stages,
                      int tmem_cols, int num_pairs, int acc_stages, const __grid_constant__ LevelPlan plan) {
    extern __shared__ unsigned char smem_raw[];
    pdl_enter();
    const int lvl = (int)warp_uniform((uint32_t)sc.ctrl->cur_level);
    if (lvl >= plan.num_levels) return;
    const long long ntiles = plan.tiles_in_level[lvl];  // super-tiles of 256 rows
    const int cluster_size = 2 * num_pairs;             // 4 = two pairs sharing one tile stream; 2 = one pair
    const int cluster_id = blockIdx.x / cluster_size;
    const int num_clusters = gridDim.x / cluster_size;
    if ((long long)cluster_id >= ntiles) return;       // same decision in every CTA of the cluster

    const uint32_t rank = warp_uniform(cluster_ctarank());
    const int pair = (int)(rank >> 1), half = (int)(rank & 1);
    const bool leader = half == 0;
    const int n_half = n_pair / 2;                      // query rows of the B operand this CTA holds
    const int q0 = q0_all + pair * n_pair;              // first query of this pair
    const int nq = max(0, min(n_pair, nq_all - pair * n_pair));

    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    constexpr int kb_elems = F32 ? kBlockK / 2 : kBlockK;
    const int num_kb = (sv.dim + kb_elems - 1) / kb_elems;   // partial last k-block: zero-filled by TMA
    const int b_kb_bytes = n_half * 128;
    unsigned char *b_smem = smem;                                              // [num_kb][n_half rows x 128 B]
    unsigned char *a_smem = b_smem + (size_t)num_kb * b_kb_bytes;              // [stages][kKbPerStage][128 rows x 128 B]
    float *tau_s = reinterpret_cast<float *>(a_smem + (size_t)stages * kStageBytes);   // [256]
    float *ntau_s = tau_s + DAB_MAX_QUERY_BATCH;                                        // [256] negated thresholds
    uint64_t *full = reinterpret_cast<uint64_t *>(ntau_s + DAB_MAX_QUERY_BATCH);       // leaders: both A halves landed
    uint64_t *empty = full + kMaxStages;               // ranks 0/1: both pairs are done with the stage
    uint64_t *b_full = empty + kMaxStages;             // this CTA's query quarter landed
    uint64_t *b_pair = b_full + 1;                     // leaders: both quarters of the pair landed
    uint64_t *tmem_full = b_pair + 1;                  // [kMaxAcc] accumulator stage complete (both CTAs of the pair)
    uint64_t *tmem_empty = tmem_full + kMaxAcc;        // [kMaxAcc] leaders: both CTAs drained the accumulator stage
    uint64_t *pend_full = tmem_empty + kMaxAcc;              // [2] all 4 epilogue warps finished appending to list a
    uint64_t *pend_free = pend_full + 2;               // [2] the flusher drained list a
    uint64_t *pend_key = pend_free + 2;                // [2][kPendCap]
    uint32_t *tmem_base_s = reinterpret_cast<uint32_t *>(pend_key + 2 * kPendCap);
    uint32_t *pend_n = tmem_base_s + 1;                // [2][kEpiWarps] hits of each epilogue warp in the tile
    uint32_t *pend_q = pend_n + 2 * kEpiWarps;         // [2][kPendCap]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < stages; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], (uint32_t)num_pairs);
        }
        mbar_init(b_full, 1);
        mbar_init(b_pair, 2);
        for (int a = 0; a < kMaxAcc; a++) {
            mbar_init(&tmem_full[a], 1);
            mbar_init(&tmem_empty[a], 16);   // 8 epilogue warps x 2 CTAs
        }
        for (int a = 0; a < 2; a++) {
            mbar_init(&pend_full[a], 8);
            mbar_init(&pend_free[a], 1);
        }
        fence_barrier_init();
        prefetch_tmap(&tmap_rows);
        prefetch_tmap(&tmap_q);
    }
    if (warp == 2) tmem_alloc2(tmem_base_s, (uint32_t)tmem_cols);
    for (int q = tid; q < DAB_MAX_QUERY_BATCH; q += kThreads) {
        const bool live = q < nq && sc.active[q0 + q];
        tau_s[q] = live ? key_score(sc.tau_key[q0 + q]) : INFINITY;
        ntau_s[q] = -tau_s[q];
    }
    tc_fence_before();
    cluster_sync_all();   // barriers of every CTA are initialised before any remote arrive / multicast
    tc_fence_after();
    const uint32_t tmem_base = warp_uniform(*tmem_base_s);
    const long long my_tiles = (ntiles - cluster_id + num_clusters - 1) / num_clusters;
    const int spt = (num_kb + kKbPerStage - 1) / kKbPerStage;   // stage uses per super-tile (the last may hold fewer k-blocks)

    if (warp == 0) {
        if (elect_one()) {
            // this CTA's quarter of the query block: rows [pair*n_pair + half*n_half, +n_half) of the bf16 copy
            mbar_arrive_expect_tx(b_full, (uint32_t)(num_kb * b_kb_bytes));
            for (int kb = 0; kb < num_kb; kb++)
                tma_load_2d(b_smem + (size_t)kb * b_kb_bytes, &tmap_q, kb * kb_elems, pair * n_pair + half * n_half, b_full);
            if (rank < 2) {
                // A tiles for the whole cluster: my 128-row half of every super-tile, multicast to my
                // twin in the other pair
                const uint16_t mask = (uint16_t)((1u << rank) | (num_pairs == 2 ? (1u << (rank + 2)) : 0u));
                int s = 0;
                uint32_t ph = 0;
                for (long long it = 0; it < my_tiles; it++) {
                    const long long t = level_tile(plan, lvl, cluster_id + it * num_clusters);
                    const int row0 = (int)(t * kSuperRows) + half * kTileRows;
                    for (int j = 0; j < spt; j++) {
                        mbar_wait_bounded(&empty[s], ph ^ 1u);
                        unsigned char *dst = a_smem + (size_t)s * kStageBytes;
#pragma unroll
                        for (int kk = 0; kk < kKbPerStage; kk++) {
                            const int kb = j * kKbPerStage + kk;
                            if (kb < num_kb) {
                                if (num_pairs == 2)
                                    tma_load_2d_mc2(dst + kk * kAStageBytes, &tmap_rows, kb * kb_elems, row0, &full[s], mask);
                                else
                                    tma_load_2d_2cta(dst + kk * kAStageBytes, &tmap_rows, kb * kb_elems, row0, &full[s]);
                            }
                        }
                        if (++s == stages) { s = 0; ph ^= 1u; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (leader && elect_one()) {
            // ------------------------------------------------------------ MMA issuer of the pair
            const uint32_t idesc = F32 ? make_idesc_tf32(n_pair, kSuperRows) : make_idesc(n_pair, kSuperRows);
            const uint16_t pair_mask = (uint16_t)(0x3u << (2 * pair));
            mbar_wait_bounded(b_pair, 0);
            tc_fence_after();
            int s = 0;
            uint32_t ph = 0;
            // both halves of a stage (mine and my peer's), every k-block of it, report to the leader's barrier
            auto stage_bytes = [&](int j) { return (uint32_t)(2 * kAStageBytes * min(kKbPerStage, num_kb - j * kKbPerStage)); };
            for (long long it = 0; it < my_tiles; it++) {
                const int a = (int)(it % acc_stages);
                const uint32_t aph = (uint32_t)((it / acc_stages) & 1);
                mbar_wait_bounded(&tmem_empty[a], aph ^ 1u);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)(a * n_pair);
                for (int j = 0; j < spt; j++) {
                    mbar_arrive_expect_tx(&full[s], stage_bytes(j));
                    mbar_wait_bounded(&full[s], ph);
                    tc_fence_after();
                    const uint32_t a_base = smem_u32(a_smem + (size_t)s * kStageBytes);
#pragma unroll
                    for (int kk = 0; kk < kKbPerStage; kk++) {
                        const int kb = j * kKbPerStage + kk;
                        if (kb < num_kb) {
                            const uint64_t da = make_smem_desc(a_base + kk * kAStageBytes);
                            const uint64_t db = make_smem_desc(smem_u32(b_smem + (size_t)kb * b_kb_bytes));
#pragma unroll
                            for (int k = 0; k < kBlockK / kUmmaK; k++)
                                if constexpr (F32)
                                    umma_tf32_2cta(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0 ? 1u : 0u);
                                else
                                    umma_f16_2cta(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0 ? 1u : 0u);
                        }
                    }
                    umma_commit_mc2(&empty[s], (uint16_t)0x3u);        // the loaders are ranks 0 and 1
                    if (++s == stages) { s = 0; ph ^= 1u; }
                }
                umma_commit_mc2(&tmem_full[a], pair_mask);
            }
        }
    } else if (warp == 3) {
        if (lane == 0) {
            mbar_wait_bounded(b_full, 0);
            if (leader) mbar_arrive(b_pair); else mbar_arrive_remote(b_pair, rank & ~1u);
        }
        // ------------------------------------------------------------ flusher: staged candidates -> global buffers
        for (long long it = 0; it < my_tiles; it++) {
            const int a = (int)(it & 1);
            const uint32_t aph = (uint32_t)((it >> 1) & 1);
            mbar_wait_bounded(&pend_full[a], aph);
            flush_warp_lists(pend_key + a * kPendCap, pend_q + a * kPendCap, pend_n + a * kEpiWarps, kEpiWarps, kWarpCap, allow, sc);
            __syncwarp();
            if (lane == 0) mbar_arrive(&pend_free[a]);
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------------ epilogue (own TMEM half)
        const int wq = warp & 3;                               // TMEM lane quarter this warp may read
        // two warps share a lane quarter and split the pair's query columns (whole 16-column slabs each):
        // one warp per quarter was the critical path once the MMAs ran at their floor
        const bool dense = lvl == 0 && warp_uniform((uint32_t)sc.ctrl->retry) == 0u;   // first visit of level 0: thresholds are all -inf
        const int ncols = (nq + 15) & ~15;
        const int c_split = (ncols / 16 + 1) / 2 * 16;
        const int c_begin = warp >= 8 ? c_split : 0;
        const int c_count = warp >= 8 ? ncols - c_split : c_split;
        // the row's 1/|row| for the NEXT tile is fetched while the current one is drained
        auto row_of = [&](long long it_) {
            const long long t = level_tile(plan, lvl, cluster_id + it_ * num_clusters);
            return t * kSuperRows + half * kTileRows + wq * 32 + lane;
        };
        long long row = row_of(0);
        float rinv = row < sv.n ? sv.inv_norm[row] : 0.f;
        for (long long it = 0; it < my_tiles; it++) {
            const int a = (int)(it % acc_stages);                  // accumulator stage
            const uint32_t aph = (uint32_t)((it / acc_stages) & 1);
            const int pl = (int)(it & 1);                           // staging list (always two)
            const uint32_t plph = (uint32_t)((it >> 1) & 1);
            const long long row_next = it + 1 < my_tiles ? row_of(it + 1) : sv.n;
            const float rinv_next = row_next < sv.n ? sv.inv_norm[row_next] : 0.f;
            mbar_wait_bounded(&pend_free[pl], plph ^ 1u);     // the flusher is done with this list (tile it-2)
            mbar_wait_bounded(&tmem_full[a], aph);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16) + (uint32_t)(a * n_pair + c_begin);
            const int ew = warp - 4;                                // 0..7
            // level 0 admits every row (no threshold exists yet): the lists would overflow at once, so the hits go
            // straight to the candidate buffers, one reservation per column (epilogue_slab_direct)
            const WarpList wl = {pend_key + pl * kPendCap + ew * kWarpCap, pend_q + pl * kPendCap + ew * kWarpCap, dense ? 0 : kWarpCap};
            uint32_t hits = 0;
            hits = epilogue_tile_wl(taddr, c_count, rinv, row, tau_s + c_begin, ntau_s + c_begin, q0 + c_begin, allow, sc, wl);
            row = row_next;
            rinv = rinv_next;
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (leader) mbar_arrive(&tmem_empty[a]); else mbar_arrive_remote(&tmem_empty[a], rank & ~1u);
                pend_n[pl * kEpiWarps + ew] = hits;
                mbar_arrive(&pend_full[pl]);                        // release: the flusher sees the list and its count
            }
        }
    }
    // ---------------------------------------------------------------- teardown
    tc_fence_before();
    cluster_sync_all();   // nobody may free TMEM or leave while a peer still multicasts into / reads from it
    if (warp == 2) tmem_dealloc2(tmem_base, (uint32_t)tmem_cols);
}

struct QuadPlan {
    int n_pair;
    int stages;
    int tmem_cols;
    size_t smem;
};

// n_pair = queries per CTA pair (multiple of 32: each CTA's share of B must be a multiple of 16 rows)
bool plan_quad(int dim, int nq, int num_pairs, int elem_bytes, QuadPlan *out) {
    int n_pair = ((nq + num_pairs - 1) / num_pairs + 31) / 32 * 32;
    if (n_pair < 32) n_pair = 32;
    if (n_pair > 128) return false;
    const size_t b_bytes = (size_t)(n_pair / 2) * (((size_t)dim * elem_bytes + 127) / 128 * 128);
    const size_t fixed = 2 * DAB_MAX_QUERY_BATCH * sizeof(float) + (2 * kMaxStages + 16) * sizeof(uint64_t) + 16 + 2 * kEpiWarps * sizeof(uint32_t) +
                         2 * kPendCap * (sizeof(uint64_t) + sizeof(uint32_t)) + 1024;
    if (b_bytes + fixed + (size_t)kMinStages * kStageBytes > (size_t)kSmemBudget) return false;
    int stages = (int)(((size_t)kSmemBudget - b_bytes - fixed) / kStageBytes);
    if (stages > kMaxStages) stages = kMaxStages;
    // all 512 TMEM columns: as many accumulator stages as fit (2..4) decouple the MMAs from epilogue jitter
    int cols = 512;
    out->n_pair = n_pair;
    out->stages = stages;
    out->tmem_cols = cols;
    out->smem = b_bytes + (size_t)stages * kStageBytes + fixed;
    return true;
}

}  // namespace

int quad_tile_rows() { return kSuperRows; }

static int quad_elem_bytes(int dtype) { return dtype == DAB_BF16 ? 2 : 4; }
static bool quad_layout_ok(int dim, int dtype) {
    return (dtype == DAB_BF16 || dtype == DAB_F32) && (dim * quad_elem_bytes(dtype)) % 16 == 0;
}

// bf16 stores (kind::f16) and fp32 stores (kind::tf32); a partial last k-block is zero-filled by TMA
bool quad_supported(int dim, int dtype, int nq) {
    QuadPlan p;
    return quad_layout_ok(dim, dtype) && nq >= 1 && nq <= DAB_MAX_QUERY_BATCH && plan_quad(dim, nq, 2, quad_elem_bytes(dtype), &p);
}

// the most queries one 4-CTA cluster (two pairs) can score per corpus pass; 0 if the layout is unsupported
int quad_queries_per_pass(int dim, int dtype) {
    int best = 0;
    QuadPlan p;
    if (!quad_layout_ok(dim, dtype)) return 0;
    for (int n = 64; n <= DAB_MAX_QUERY_BATCH; n += 64)
        if (plan_quad(dim, n, 2, quad_elem_bytes(dtype), &p)) best = n;
    return best;
}

// the most queries ONE pair (cluster of 2, no multicast) can score per corpus pass
int pair_queries_per_pass(int dim, int dtype) {
    int best = 0;
    QuadPlan p;
    if (!quad_layout_ok(dim, dtype)) return 0;
    for (int n = 32; n <= 256; n += 32)
        if (plan_quad(dim, n, 1, quad_elem_bytes(dtype), &p)) best = n;
    return best;
}

int launch_scan_umma_quad(const StoreView &sv, const uint32_t *allow, const SearchScratch &sc, int q0, int nq,
                          const LevelPlan &plan, int num_sms, cudaStream_t st, int num_pairs) {
    QuadPlan qp;
    if (num_pairs != 1 && num_pairs != 2) return fail(DAB_E_INVALID, "scan_umma_quad: 1 or 2 pairs per cluster");
    const int kCluster = 2 * num_pairs;
    const bool f32 = sv.dtype == DAB_F32;
    if (!quad_layout_ok(sv.dim, sv.dtype) || !plan_quad(sv.dim, nq, num_pairs, quad_elem_bytes(sv.dtype), &qp))
        return fail(DAB_E_INVALID, "scan_umma_quad: needs a bf16 or fp32 store and a query block that fits shared memory");
    if (plan.tile_rows != kSuperRows) return fail(DAB_E_INVALID, "scan_umma_quad: level plan built for another tile size");
    alignas(64) CUtensorMap tmap_rows, tmap_q;
    if (int rc = cached_map(&tmap_rows, sv.rows, (uint64_t)sv.n, (uint64_t)sv.dim, kTileRows, f32)) return rc;
    // B = the unit-norm fp32 queries (fp32 stores) or their bf16 copy (bf16 stores); rows past the buffer read as zero
    const void *qbase = f32 ? static_cast<const void *>(sc.qhat + (size_t)q0 * sv.dim)
                            : static_cast<const void *>(sc.qbf16 + (size_t)q0 * sv.dim);
    const uint64_t q_rows = (uint64_t)(DAB_MAX_QUERY_BATCH - q0);
    if (int rc = cached_map(&tmap_q, qbase, q_rows, (uint64_t)sv.dim, (uint32_t)(qp.n_pair / 2), f32)) return rc;
    auto kernel = f32 ? scan_umma_quad_kernel<true> : scan_umma_quad_kernel<false>;
    static SmemAttrOnce attr_once[2];
    static ClusterOccupancyCache occ_cache[2];
    DAB_CUDA_OK(attr_once[f32 ? 1 : 0].ensure(kernel, (int)qp.smem));
    // how many 4-CTA clusters can be resident at once (GPC shapes strand a few SMs)
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(num_sms / kCluster * kCluster), 1, 1);
    cfg.blockDim = dim3(kThreads, 1, 1);
    cfg.dynamicSmemBytes = qp.smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kCluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;   // see pdl_enter (common.cuh)
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    int max_clusters = 0;
    DAB_CUDA_OK(occ_cache[f32 ? 1 : 0].get(&max_clusters, kernel, &cfg, kCluster));
    if (max_clusters < 1) return fail(DAB_E_CUDA, "scan_umma_quad: no 4-CTA cluster fits on this device");
    long long max_tiles = 0;
    for (int l = 0; l < plan.num_levels; l++) max_tiles = plan.tiles_in_level[l] > max_tiles ? plan.tiles_in_level[l] : max_tiles;
    const long long clusters = max_tiles < max_clusters ? max_tiles : max_clusters;
    int acc = qp.tmem_cols / qp.n_pair;
    if (acc > kMaxAcc) acc = kMaxAcc;
    if (const char *e = tuning_env("DAB_QUAD_ACC")) acc = atoi(e) >= 2 && atoi(e) <= acc ? atoi(e) : acc;  // tuning aid
    if (const char *e = tuning_env("DAB_QUAD_STAGES")) qp.stages = atoi(e) >= 2 && atoi(e) <= qp.stages ? atoi(e) : qp.stages;
    static const bool debug = tuning_env("DAB_DEBUG") != nullptr;
    if (debug)
        fprintf(stderr, "[dab] quad: pairs=%d nq=%d n_pair=%d stages=%d tmem_cols=%d smem=%zu max_clusters=%d clusters=%lld levels=%d\n", num_pairs, nq,
                qp.n_pair, qp.stages, qp.tmem_cols, qp.smem, max_clusters, clusters, plan.num_levels);
    cfg.gridDim = dim3((unsigned)(clusters * kCluster), 1, 1);
    DAB_CUDA_OK(cudaLaunchKernelEx(&cfg, kernel, tmap_rows, tmap_q, sv, allow, sc, q0, nq, qp.n_pair, qp.stages,
                                   qp.tmem_cols, num_pairs, acc, plan));
    count_launch();
    return DAB_OK;
}

}  // namespace dab
