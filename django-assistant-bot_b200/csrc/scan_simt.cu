// scan_simt.cu - bandwidth-bound exact scoring for small query batches (CUDA cores, fp32 FMA).
//
// The corpus side of `ORDER BY embedding <=> q` (assistant/rag/services/search_service.py:192-194):
// every stored row is dotted with up to 4 unit-norm fp32 queries and scaled by its stored 1/|row|.
//
// Structure (per CTA, persistent over the tiles of the current level):
//   warp 0      producer: one lane issues 1-D TMA bulk copies (cp.async.bulk, SASS UBLKCP) of
//               [tile_rows x dim] contiguous row bytes + their inverse norms into a 4-stage
//               shared-memory ring, signalled through mbarriers (full/empty per stage)
//   warps 1..8  consumers: each warp owns rows w, w+8, ... of a tile; lanes read 16-byte
//               column vectors (bank-conflict free), FMA against query registers, butterfly-
//               reduce, scale, and compare with the query's admission threshold.  Rows that
//               pass (rare once a threshold exists) are appended to the query's candidate
//               buffer with one atomicAdd.
// HBM traffic = the tile bytes, once.  No corpus byte is re-read within a search.
#include "search.cuh"

namespace dab {

namespace {

constexpr int kStages = 4;
constexpr int kConsumerWarps = 8;
constexpr int kScanThreads = 32 * (1 + kConsumerWarps);

template <bool BF16>
struct RowVec;
template <>
struct RowVec<false> {
    static constexpr int kElems = 4;
    __device__ static __forceinline__ void unpack(const uint4 &v, float *x) {
        x[0] = __uint_as_float(v.x);
        x[1] = __uint_as_float(v.y);
        x[2] = __uint_as_float(v.z);
        x[3] = __uint_as_float(v.w);
    }
};
template <>
struct RowVec<true> {
    static constexpr int kElems = 8;
    __device__ static __forceinline__ void unpack(const uint4 &v, float *x) {
        x[0] = bf16_lo(v.x);
        x[1] = bf16_hi(v.x);
        x[2] = bf16_lo(v.y);
        x[3] = bf16_hi(v.y);
        x[4] = bf16_lo(v.z);
        x[5] = bf16_hi(v.z);
        x[6] = bf16_lo(v.w);
        x[7] = bf16_hi(v.w);
    }
};

// tile index of the j-th tile of level `lvl`
__device__ __forceinline__ long long level_tile(const LevelPlan &plan, int lvl, long long j) {
    const long long i = lvl == 0 ? j : j + j / (plan.growth - 1) + 1;
    return i * plan.stride[lvl];
}

// QB queries per pass, NCH 16-byte column chunks per lane (dim <= NCH*32*kElems)
template <bool BF16, int QB, int NCH>
__global__ void __launch_bounds__(kScanThreads, 1)
scan_simt_kernel(StoreView sv, const uint32_t *__restrict__ allow, SearchScratch sc, int q0, int nq,
                 const __grid_constant__ LevelPlan plan) {
    constexpr int VE = RowVec<BF16>::kElems;
    extern __shared__ __align__(128) unsigned char smem[];
    pdl_enter();
    const int lvl = sc.ctrl->cur_level;
    if (lvl >= plan.num_levels) return;
    const long long ntiles = plan.tiles_in_level[lvl];
    if ((long long)blockIdx.x >= ntiles) return;

    const int R = plan.tile_rows;
    const int row_bytes = sv.dim * (BF16 ? 2 : 4);
    const int stage_bytes = R * row_bytes;
    const int inv_bytes = R * 4;
    unsigned char *tiles = smem;                                              // [kStages][stage_bytes]
    float *invs = reinterpret_cast<float *>(smem + (size_t)kStages * stage_bytes);  // [kStages][R]
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + (size_t)kStages * (stage_bytes + inv_bytes));
    uint64_t *empty = full + kStages;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < kStages; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], kConsumerWarps);
        }
        fence_barrier_init();
    }
    __syncthreads();

    const long long my_tiles = (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x;

    if (warp == 0) {
        // ------------------------------------------------ producer (one elected lane: see elect_one in common.cuh)
        if (elect_one()) {
            for (long long it = 0; it < my_tiles; it++) {
                const int s = (int)(it % kStages);
                const uint32_t ph = (uint32_t)((it / kStages) & 1);
                mbar_wait(&empty[s], ph ^ 1u);
                const long long t = level_tile(plan, lvl, blockIdx.x + it * gridDim.x);
                const long long row0 = t * R;
                const int nrows = (int)min((long long)R, sv.n - row0);
                const uint32_t bytes = (uint32_t)nrows * row_bytes;
                const uint32_t ibytes = (uint32_t)((nrows + 3) & ~3) * 4u;
                mbar_arrive_expect_tx(&full[s], bytes + ibytes);
                tma_load_1d(tiles + (size_t)s * stage_bytes, static_cast<const unsigned char *>(sv.rows) + (size_t)row0 * row_bytes,
                            bytes, &full[s]);
                tma_load_1d(invs + (size_t)s * R, sv.inv_norm + row0, ibytes, &full[s]);
            }
        }
        return;
    }

    // ---------------------------------------------------- consumers
    const int cw = warp - 1;
    float qreg[QB][NCH * VE];
    float tau[QB];
#pragma unroll
    for (int q = 0; q < QB; q++) {
        const bool live = q < nq && sc.active[q0 + q];
        tau[q] = live ? key_score(sc.tau_key[q0 + q]) : INFINITY;
        const float *qv = sc.qhat + (size_t)(q0 + min(q, nq - 1)) * sv.dim;
#pragma unroll
        for (int c = 0; c < NCH; c++) {
            const int col = (c * 32 + lane) * VE;
#pragma unroll
            for (int e = 0; e < VE; e++) qreg[q][c * VE + e] = (col + e < sv.dim) ? qv[col + e] : 0.f;
        }
    }

    for (long long it = 0; it < my_tiles; it++) {
        const int s = (int)(it % kStages);
        const uint32_t ph = (uint32_t)((it / kStages) & 1);
        const long long t = level_tile(plan, lvl, blockIdx.x + it * gridDim.x);
        const long long row0 = t * R;
        const int nrows = (int)min((long long)R, sv.n - row0);
        mbar_wait(&full[s], ph);
        const unsigned char *tile = tiles + (size_t)s * stage_bytes;
        const float *inv = invs + (size_t)s * R;
#pragma unroll 2
        for (int r = cw; r < nrows; r += kConsumerWarps) {
            const uint4 *rowp = reinterpret_cast<const uint4 *>(tile + (size_t)r * row_bytes);
            float acc[QB];
#pragma unroll
            for (int q = 0; q < QB; q++) acc[q] = 0.f;
#pragma unroll
            for (int c = 0; c < NCH; c++) {
                const int vi = c * 32 + lane;
                if (vi * VE < sv.dim) {
                    float x[VE];
                    RowVec<BF16>::unpack(rowp[vi], x);
#pragma unroll
                    for (int q = 0; q < QB; q++)
#pragma unroll
                        for (int e = 0; e < VE; e++) acc[q] = fmaf(x[e], qreg[q][c * VE + e], acc[q]);
                }
            }
            const float rinv = inv[r];
#pragma unroll
            for (int q = 0; q < QB; q++) {
                const float sco = warp_sum(acc[q]) * rinv;
                if (sco >= tau[q]) {
                    // slow path (rare once a threshold exists): exact key test, filter, append
                    if (lane == 0) {
                        const long long row = row0 + r;
                        const bool ok = rinv > 0.f && (allow == nullptr || ((allow[row >> 5] >> (row & 31)) & 1u));
                        const uint64_t key = make_key(sco, (uint32_t)row);
                        if (ok && key > sc.tau_key[q0 + q]) {
                            const uint32_t pos = atomicAdd(&sc.cand_count[q0 + q], 1u);
                            if (pos < (uint32_t)kCandCap) sc.cand[(size_t)(q0 + q) * kCandCap + pos] = key;
                        }
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }
}

size_t simt_smem_bytes(int tile_rows, int row_bytes) {
    return (size_t)kStages * ((size_t)tile_rows * row_bytes + (size_t)tile_rows * 4) + 2 * kStages * sizeof(uint64_t);
}

template <bool BF16, int QB, int NCH>
int launch_one(const StoreView &sv, const uint32_t *allow, const SearchScratch &sc, int q0, int nq, const LevelPlan &plan,
               int num_sms, cudaStream_t st) {
    auto kern = scan_simt_kernel<BF16, QB, NCH>;
    const int row_bytes = sv.dim * (BF16 ? 2 : 4);
    const size_t smem = simt_smem_bytes(plan.tile_rows, row_bytes);
    static SmemAttrOnce attr_once;   // one per template instance
    DAB_CUDA_OK(attr_once.ensure(kern, (int)smem));
    long long max_tiles = 0;
    for (int l = 0; l < plan.num_levels; l++) max_tiles = plan.tiles_in_level[l] > max_tiles ? plan.tiles_in_level[l] : max_tiles;
    const int grid = (int)(max_tiles < num_sms ? max_tiles : num_sms);
    DAB_CUDA_OK(launch_chained(kern, dim3(grid), dim3(kScanThreads), smem, st, sv, allow, sc, q0, nq, plan));
    count_launch();
    return DAB_OK;
}

// template NCH the dispatcher rounds a chunk count to
int round_nch(int nch) { return nch <= 4 ? nch : nch <= 6 ? 6 : nch <= 8 ? 8 : 16; }

template <bool BF16, int QB, int NCH>
int launch_checked(const StoreView &sv, const uint32_t *allow, const SearchScratch &sc, int q0, int nq, const LevelPlan &plan,
                   int num_sms, cudaStream_t st) {
    constexpr int VE = BF16 ? 8 : 4;
    if constexpr (QB * NCH * VE <= 128) {  // query registers per lane
        return launch_one<BF16, QB, NCH>(sv, allow, sc, q0, nq, plan, num_sms, st);
    } else {
        return fail(DAB_E_INVALID, "scan_simt: too many queries per pass for this row width");
    }
}

template <bool BF16, int QB>
int launch_nch(const StoreView &sv, const uint32_t *allow, const SearchScratch &sc, int q0, int nq, const LevelPlan &plan,
               int num_sms, cudaStream_t st) {
    constexpr int VE = BF16 ? 8 : 4;
    const int nch = (sv.dim + 32 * VE - 1) / (32 * VE);
    if (nch > 16) return fail(DAB_E_INVALID, "scan: dim too large for the CUDA-core path");
    switch (round_nch(nch)) {
        case 1: return launch_checked<BF16, QB, 1>(sv, allow, sc, q0, nq, plan, num_sms, st);
        case 2: return launch_checked<BF16, QB, 2>(sv, allow, sc, q0, nq, plan, num_sms, st);
        case 3: return launch_checked<BF16, QB, 3>(sv, allow, sc, q0, nq, plan, num_sms, st);
        case 4: return launch_checked<BF16, QB, 4>(sv, allow, sc, q0, nq, plan, num_sms, st);
        case 6: return launch_checked<BF16, QB, 6>(sv, allow, sc, q0, nq, plan, num_sms, st);
        case 8: return launch_checked<BF16, QB, 8>(sv, allow, sc, q0, nq, plan, num_sms, st);
        default: return launch_checked<BF16, QB, 16>(sv, allow, sc, q0, nq, plan, num_sms, st);
    }
}

}  // namespace

// rows per tile: ~24-32 KiB stages, a multiple of the 8 consumer warps
int simt_tile_rows(int dim, int dtype) {
    const int row_bytes = dim * (dtype == DAB_BF16 ? 2 : 4);
    int r = (32 * 1024 / row_bytes) / kConsumerWarps * kConsumerWarps;
    if (r < kConsumerWarps) r = kConsumerWarps;
    if (r > 64) r = 64;
    // 4 stages must fit in ~200 KiB
    while (r > 4 && simt_smem_bytes(r, row_bytes) > 200 * 1024) r -= 4;
    return r;
}

// queries one pass can keep in registers (<= 128 fp32 query registers per lane)
int simt_queries_per_pass(int dim, int dtype) {
    const int ve = dtype == DAB_BF16 ? 8 : 4;
    const int nch = round_nch((dim + 32 * ve - 1) / (32 * ve));
    int qb = 4;
    while (qb > 1 && qb * nch * ve > 128) qb >>= 1;
    return qb;
}

int launch_scan_simt(const StoreView &sv, const uint32_t *allow, const SearchScratch &sc, int q0, int nq,
                     const LevelPlan &plan, int num_sms, cudaStream_t st) {
    const bool bf = sv.dtype == DAB_BF16;
    if (nq < 1 || nq > 4) return fail(DAB_E_INVALID, "scan_simt: 1..4 queries per pass");
    if (nq > simt_queries_per_pass(sv.dim, sv.dtype)) return fail(DAB_E_INVALID, "scan_simt: too many queries per pass for this row width");
#define DAB_SCAN(BF, QB) return launch_nch<BF, QB>(sv, allow, sc, q0, nq, plan, num_sms, st)
    if (nq == 1) { if (bf) DAB_SCAN(true, 1); else DAB_SCAN(false, 1); }
    if (nq == 2) { if (bf) DAB_SCAN(true, 2); else DAB_SCAN(false, 2); }
    // 3 queries run in the 4-wide kernel with one idle slot
    if (bf) DAB_SCAN(true, 4); else DAB_SCAN(false, 4);
#undef DAB_SCAN
}

}  // namespace dab
