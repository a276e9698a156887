// scan_umma.cu - tensor-core scoring for larger query batches (tcgen05 + TMEM + TMA): bf16 stores with
// kind::f16 MMAs, fp32 stores with kind::tf32 MMAs (the kernel is a template on the store type).
//
// Scoring Q queries against N rows is a dense [N x D] x [D x Q] contraction; beyond a handful of
// queries the CUDA cores cannot keep up with HBM, the 5th-gen tensor cores can.  Per CTA
// (persistent, one per SM, cta_group::1):
//   warp 0      TMA producer (one elected lane): the query block B [n_pad x D] bf16 is loaded ONCE into shared memory
//               (it stays resident for the whole launch); corpus tiles A [128 rows x 64 cols] stream
//               through a multi-stage ring (cp.async.bulk.tensor.2d, 128-byte swizzle, mbarriers)
//   warp 1      MMA issuer: one elected lane issues tcgen05.mma.cta_group::1.kind::f16 (M=128, N=n_pad, K=16)
//               D/64 x 4 times per tile into one of two TMEM accumulator stages, then
//               tcgen05.commit -> mbarrier frees the smem stage / publishes the accumulator
//   warp 2      TMEM allocator (tcgen05.alloc / dealloc)
//   warps 4-7   epilogue: tcgen05.ld the 128 x n_pad fp32 accumulator (lane = corpus row, column =
//               query), scale by the row's 1/|row|, compare with the query's admission threshold,
//               append the rare survivors to the candidate buffer.  Runs concurrently with the MMAs
//               of the next tile (double-buffered TMEM).
// The query is rounded to bf16 for this pass, so these scores are approximate: the selection keeps
// everything within a rigorous margin of the k-th best and finish_rescore_kernel re-scores the kept
// rows exactly in fp32 (select.cu).
// HBM traffic = the corpus tile bytes, once; B is read from L2 once per CTA per launch.
#include "umma.cuh"

namespace dab {

namespace {

using namespace umma;
constexpr int kThreads = 256;
constexpr int kMaxStages = 8;
constexpr int kMinStages = 4;
constexpr int kSmemBudget = 227 * 1024 - 2048;  // keep room for alignment slack

template <bool F32>
__global__ void __launch_bounds__(kThreads, 1)
scan_umma_kernel(const __grid_constant__ CUtensorMap tmap_rows, const __grid_constant__ CUtensorMap tmap_q, StoreView sv,
                 const uint32_t *__restrict__ allow, SearchScratch sc, int q0_all, int nq_all, int n_pad, int stages,
                 int tmem_cols, int groups, const __grid_constant__ LevelPlan plan) {
    // F32: fp32 store and fp32 queries, scored with kind::tf32 (K = 8 per MMA); a k-block is
    // 128 bytes of a row either way: 64 bf16 or 32 fp32
    extern __shared__ unsigned char smem_raw[];
    pdl_enter();
    const int lvl = (int)warp_uniform((uint32_t)sc.ctrl->cur_level);
    if (lvl >= plan.num_levels) return;
    const long long ntiles = plan.tiles_in_level[lvl];
    // Query-slice groups: when the query block is too wide to stay resident in one SM's shared
    // memory it is cut into `groups` slices of n_pad columns.  CTA b scores slice b % groups over
    // the tile sequence b / groups: the `groups` CTAs of a sequence stream the same corpus tiles at
    // the same time, so HBM still sees every byte once and the other reads are L2 hits.
    const int slice = blockIdx.x % groups;
    const int seq = blockIdx.x / groups;
    const int num_seq = gridDim.x / groups;
    if ((long long)seq >= ntiles) return;
    const int q0 = q0_all + slice * n_pad;
    const int nq = max(0, min(n_pad, nq_all - slice * n_pad));

    unsigned char *smem = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    constexpr int kb_elems = F32 ? kBlockK / 2 : kBlockK;    // elements per 128-byte k-block
    const int num_kb = (sv.dim + kb_elems - 1) / kb_elems;   // a partial last k-block is zero-filled by TMA (both operands)
    const int b_kb_bytes = n_pad * 128;                   // one k-block of the query operand
    unsigned char *b_smem = smem;                          // [num_kb][n_pad rows x 128 B]
    unsigned char *a_smem = b_smem + (size_t)num_kb * b_kb_bytes;  // [stages][128 rows x 128 B]
    float *tau_s = reinterpret_cast<float *>(a_smem + (size_t)stages * kAStageBytes);  // [256]
    float *ntau_s = tau_s + DAB_MAX_QUERY_BATCH;                                        // [256] negated thresholds
    uint64_t *tau_key_s = reinterpret_cast<uint64_t *>(ntau_s + DAB_MAX_QUERY_BATCH);  // [256] exact threshold keys
    uint64_t *full = tau_key_s + DAB_MAX_QUERY_BATCH;
    uint64_t *empty = full + kMaxStages;
    uint64_t *b_full = empty + kMaxStages;
    uint64_t *tmem_full = b_full + 1;                      // [2]
    uint64_t *tmem_empty = tmem_full + 2;                  // [2]
    uint32_t *tmem_base_s = reinterpret_cast<uint32_t *>(tmem_empty + 2);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < stages; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 1);
        }
        mbar_init(b_full, 1);
        for (int a = 0; a < 2; a++) {
            mbar_init(&tmem_full[a], 1);
            mbar_init(&tmem_empty[a], 4);
        }
        fence_barrier_init();
        prefetch_tmap(&tmap_rows);
        prefetch_tmap(&tmap_q);
    }
    if (warp == 2) tmem_alloc(tmem_base_s, (uint32_t)tmem_cols);
    for (int q = tid; q < DAB_MAX_QUERY_BATCH; q += kThreads) {
        const bool live = q < nq && sc.active[q0 + q];
        tau_s[q] = live ? key_score(sc.tau_key[q0 + q]) : INFINITY;
        ntau_s[q] = -tau_s[q];
        tau_key_s[q] = live ? sc.tau_key[q0 + q] : ~0ull;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = warp_uniform(*tmem_base_s);
    const long long my_tiles = (ntiles - seq + num_seq - 1) / num_seq;

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        if (elect_one()) {
            mbar_arrive_expect_tx(b_full, (uint32_t)(num_kb * b_kb_bytes));
            for (int kb = 0; kb < num_kb; kb++)
                tma_load_2d(b_smem + (size_t)kb * b_kb_bytes, &tmap_q, kb * kb_elems, slice * n_pad, b_full);
            int s = 0;
            uint32_t ph = 0;
            for (long long it = 0; it < my_tiles; it++) {
                const long long t = level_tile(plan, lvl, seq + it * num_seq);
                const int row0 = (int)(t * kTileRows);
                for (int kb = 0; kb < num_kb; kb++) {
                    mbar_wait_bounded(&empty[s], ph ^ 1u);
                    mbar_arrive_expect_tx(&full[s], kAStageBytes);
                    tma_load_2d(a_smem + (size_t)s * kAStageBytes, &tmap_rows, kb * kb_elems, row0, &full[s]);
                    if (++s == stages) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        if (elect_one()) {
            const uint32_t idesc = F32 ? make_idesc_tf32(n_pad) : make_idesc(n_pad);
            mbar_wait_bounded(b_full, 0);
            tc_fence_after();
            int s = 0;
            uint32_t ph = 0;
            for (long long it = 0; it < my_tiles; it++) {
                const int a = (int)(it & 1);
                const uint32_t aph = (uint32_t)((it >> 1) & 1);
                mbar_wait_bounded(&tmem_empty[a], aph ^ 1u);   // epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)(a * n_pad);
                for (int kb = 0; kb < num_kb; kb++) {
                    mbar_wait_bounded(&full[s], ph);
                    tc_fence_after();
                    const uint64_t da = make_smem_desc(smem_u32(a_smem + (size_t)s * kAStageBytes));
                    const uint64_t db = make_smem_desc(smem_u32(b_smem + (size_t)kb * b_kb_bytes));
#pragma unroll
                    for (int k = 0; k < kBlockK / kUmmaK; k++) {
                        // advancing K by one MMA (16 bf16 or 8 fp32) = 32 bytes inside the swizzle atom: +2 in the address field
                        if constexpr (F32)
                            umma_tf32(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0 ? 1u : 0u);
                        else
                            umma_f16(tmem_d, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    umma_commit(&empty[s]);                    // smem stage reusable once these MMAs finish
                    if (++s == stages) { s = 0; ph ^= 1u; }
                }
                umma_commit(&tmem_full[a]);                    // accumulator complete
            }
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------------ epilogue
        const int wq = warp & 3;                               // TMEM lane quarter this warp may read
        // the row's 1/|row| for the NEXT tile is fetched while the current one is drained
        auto row_of = [&](long long it_) {
            const long long t = level_tile(plan, lvl, seq + it_ * num_seq);
            return t * kTileRows + wq * 32 + lane;
        };
        long long row = row_of(0);
        float rinv = row < sv.n ? sv.inv_norm[row] : 0.f;
        for (long long it = 0; it < my_tiles; it++) {
            const int a = (int)(it & 1);
            const uint32_t aph = (uint32_t)((it >> 1) & 1);
            const long long row_next = it + 1 < my_tiles ? row_of(it + 1) : sv.n;
            const float rinv_next = row_next < sv.n ? sv.inv_norm[row_next] : 0.f;
            mbar_wait_bounded(&tmem_full[a], aph);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16) + (uint32_t)(a * n_pad);
            epilogue_tile_wl(taddr, (nq + 15) & ~15, rinv, row, tau_s, ntau_s, q0, allow, sc, WarpList{nullptr, nullptr, 0}, tau_key_s);
            row = row_next;
            rinv = rinv_next;
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[a]);
        }
    }
    // ---------------------------------------------------------------- teardown
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, (uint32_t)tmem_cols);
}

struct UmmaPlan {
    int n_pad;
    int stages;
    int tmem_cols;
    size_t smem;
};

bool plan_umma(int dim, int nq, int elem_bytes, UmmaPlan *out) {
    const int n_pad = nq <= 16 ? 16 : (nq + 15) / 16 * 16;
    const size_t b_bytes = (size_t)n_pad * (((size_t)dim * elem_bytes + 127) / 128 * 128);   // whole 128-byte k-blocks
    const size_t fixed = 2 * DAB_MAX_QUERY_BATCH * sizeof(float) + DAB_MAX_QUERY_BATCH * sizeof(uint64_t) + (2 * kMaxStages + 8) * sizeof(uint64_t) + 1024;
    if (b_bytes + fixed + (size_t)kMinStages * kAStageBytes > (size_t)kSmemBudget) return false;
    int stages = (int)(((size_t)kSmemBudget - b_bytes - fixed) / kAStageBytes);
    if (stages > kMaxStages) stages = kMaxStages;
    int cols = 32;
    while (cols < 2 * n_pad) cols <<= 1;
    if (cols > 512) return false;
    out->n_pad = n_pad;
    out->stages = stages;
    out->tmem_cols = cols;
    out->smem = b_bytes + (size_t)stages * kAStageBytes + fixed;
    return true;
}

int elem_bytes_of(int dtype) { return dtype == DAB_BF16 ? 2 : 4; }

}  // namespace

int umma_tile_rows() { return kTileRows; }

// bf16 stores (kind::f16) and fp32 stores (kind::tf32).  Rows need not be whole 128-byte k-blocks: the tensor maps
// describe the true row length and TMA zero-fills the tail of the last k-block of both operands (row pitch % 16 == 0
// is what the tensor map needs, and dab_store_create guarantees it)
bool umma_supported(int dim, int dtype) {
    UmmaPlan p;
    if (dtype != DAB_BF16 && dtype != DAB_F32) return false;
    return (dim * elem_bytes_of(dtype)) % 16 == 0 && plan_umma(dim, 16, elem_bytes_of(dtype), &p);
}

// the widest query block whose copy stays resident in shared memory next to >= 4 corpus stages
int umma_queries_per_pass(int dim, int dtype) {
    int best = 0;
    UmmaPlan p;
    for (int n = 16; n <= DAB_MAX_QUERY_BATCH; n += 16)
        if (plan_umma(dim, n, elem_bytes_of(dtype), &p)) best = n;
    return best;
}

int launch_scan_umma(const StoreView &sv, const uint32_t *allow, const SearchScratch &sc, int q0, int nq, const LevelPlan &plan,
                     int num_sms, cudaStream_t st) {
    if (!umma_supported(sv.dim, sv.dtype))
        return fail(DAB_E_INVALID, "scan_umma: needs a bf16 or fp32 store whose query block fits shared memory");
    if (plan.tile_rows != kTileRows) return fail(DAB_E_INVALID, "scan_umma: level plan built for another tile size");
    const bool f32 = sv.dtype == DAB_F32;
    const int eb = elem_bytes_of(sv.dtype);
    // one slice if the whole query block stays resident in shared memory, else `groups` equal slices
    const int qpass = umma_queries_per_pass(sv.dim, sv.dtype);
    if (qpass <= 0 || nq < 1) return fail(DAB_E_INVALID, "scan_umma: query block does not fit shared memory");
    const int groups = (nq + qpass - 1) / qpass;
    const int per_slice = (nq + groups - 1) / groups;
    // B comes from the unit-norm fp32 queries (fp32 stores) or from their bf16 copy (bf16 stores); rows past
    // the live queries hold stale data whose columns the epilogue ignores, rows past the buffer read as zero
    const int q_rows = DAB_MAX_QUERY_BATCH - q0;
    UmmaPlan up;
    if (!plan_umma(sv.dim, per_slice, eb, &up) || groups > num_sms)
        return fail(DAB_E_INVALID, "scan_umma: query block does not fit shared memory");
    alignas(64) CUtensorMap tmap_rows, tmap_q;
    if (int rc = cached_map(&tmap_rows, sv.rows, (uint64_t)sv.n, (uint64_t)sv.dim, kTileRows, f32)) return rc;
    const void *qbase = f32 ? static_cast<const void *>(sc.qhat + (size_t)q0 * sv.dim)
                            : static_cast<const void *>(sc.qbf16 + (size_t)q0 * sv.dim);
    if (int rc = cached_map(&tmap_q, qbase, (uint64_t)q_rows, (uint64_t)sv.dim, (uint32_t)up.n_pad, f32)) return rc;
    auto kernel = f32 ? scan_umma_kernel<true> : scan_umma_kernel<false>;
    static SmemAttrOnce attr_once[2];
    DAB_CUDA_OK(attr_once[f32 ? 1 : 0].ensure(kernel, (int)up.smem));
    long long max_tiles = 0;
    for (int l = 0; l < plan.num_levels; l++) max_tiles = plan.tiles_in_level[l] > max_tiles ? plan.tiles_in_level[l] : max_tiles;
    const long long seqs = max_tiles < num_sms / groups ? max_tiles : num_sms / groups;
    const int grid = (int)seqs * groups;
    DAB_CUDA_OK(launch_chained(kernel, dim3(grid), dim3(kThreads), up.smem, st, tmap_rows, tmap_q, sv, allow, sc, q0, nq, up.n_pad, up.stages,
                               up.tmem_cols, groups, plan));
    count_launch();
    return DAB_OK;
}

}  // namespace dab
