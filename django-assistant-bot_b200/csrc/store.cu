// store.cu - the device-resident chunk-embedding store and the search orchestration.
//
// One dab_store is one shard of the table behind assistant/storage/models.py:19-29
// (BaseEmbeddingModel.embedding) held in HBM: rows [capacity, dim] (fp32 like pgvector, or
// bf16), 1/|row| per row, the row ids (Django pks), and the scratch a search needs.
// dab_store_search is what `_objects_embedding_search` (assistant/rag/services/
// search_service.py:185-196) becomes: exact cosine top-k, (score desc, row asc).
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <new>
#include <utility>
#include <vector>

#include <cub/device/device_scan.cuh>

#include "store.cuh"

namespace dab {

static thread_local std::string g_last_error;
void set_error(const std::string &msg) { g_last_error = msg; }
int fail(int code, const std::string &msg) {
    g_last_error = msg;
    return code;
}

static std::atomic<unsigned long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

int simt_queries_per_pass(int dim, int dtype);

}  // namespace dab

namespace dab {

namespace {

constexpr int kVerdictSlots = 256;

// What this thread searched last: dab_store_search_status / dab_store_status_dev answer for THAT call, not for
// whatever another thread enqueued on the same store in between.
struct LastSearch {
    const dab_store *store = nullptr;
    uint64_t seq = 0;
    bool empty = true;          // nothing was scanned (empty store): the verdict is DAB_OK
    bool small_tensor = false;  // a small (<= CUDA-core queries per pass) auto search that took the tensor path
};
thread_local LastSearch t_last;

// ---------------------------------------------------------------- append: copy/convert rows, 1/|row|
// One warp per row; 128-bit loads/stores; the norm is taken from the values as STORED.
template <bool IN_BF16, bool OUT_BF16>
__global__ void __launch_bounds__(256)
append_rows_kernel(const void *__restrict__ src, void *__restrict__ dst, float *__restrict__ inv_norm,
                   int64_t *__restrict__ ids_dst, const int64_t *__restrict__ ids_src, int64_t id_first,
                   int64_t n, int dim) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= n) return;
    float ss = 0.f;
    if constexpr (IN_BF16 && OUT_BF16) {
        const uint4 *s = reinterpret_cast<const uint4 *>(static_cast<const __nv_bfloat16 *>(src) + row * dim);
        uint4 *d = reinterpret_cast<uint4 *>(static_cast<__nv_bfloat16 *>(dst) + row * dim);
        for (int c = lane; c < dim / 8; c += 32) {
            const uint4 v = s[c];
            d[c] = v;
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const float a = bf16_lo(w[i]), b = bf16_hi(w[i]);
                ss = fmaf(a, a, ss);
                ss = fmaf(b, b, ss);
            }
        }
    } else if constexpr (!IN_BF16 && !OUT_BF16) {
        const float4 *s = reinterpret_cast<const float4 *>(static_cast<const float *>(src) + row * dim);
        float4 *d = reinterpret_cast<float4 *>(static_cast<float *>(dst) + row * dim);
        for (int c = lane; c < dim / 4; c += 32) {
            const float4 v = s[c];
            d[c] = v;
            ss = fmaf(v.x, v.x, ss);
            ss = fmaf(v.y, v.y, ss);
            ss = fmaf(v.z, v.z, ss);
            ss = fmaf(v.w, v.w, ss);
        }
    } else if constexpr (!IN_BF16 && OUT_BF16) {
        const float4 *s = reinterpret_cast<const float4 *>(static_cast<const float *>(src) + row * dim);
        uint2 *d = reinterpret_cast<uint2 *>(static_cast<__nv_bfloat16 *>(dst) + row * dim);
        for (int c = lane; c < dim / 4; c += 32) {
            const float4 v = s[c];
            const __nv_bfloat16 b0 = __float2bfloat16_rn(v.x), b1 = __float2bfloat16_rn(v.y);
            const __nv_bfloat16 b2 = __float2bfloat16_rn(v.z), b3 = __float2bfloat16_rn(v.w);
            uint2 o;
            o.x = (uint32_t)__bfloat16_as_ushort(b0) | ((uint32_t)__bfloat16_as_ushort(b1) << 16);
            o.y = (uint32_t)__bfloat16_as_ushort(b2) | ((uint32_t)__bfloat16_as_ushort(b3) << 16);
            d[c] = o;
            const float f0 = __bfloat162float(b0), f1 = __bfloat162float(b1), f2 = __bfloat162float(b2), f3 = __bfloat162float(b3);
            ss = fmaf(f0, f0, ss);
            ss = fmaf(f1, f1, ss);
            ss = fmaf(f2, f2, ss);
            ss = fmaf(f3, f3, ss);
        }
    }
    ss = warp_sum(ss);
    if (lane == 0) {
        // NaN/inf rows get 0 as well: they can never be ranked (PostgreSQL sorts NaN distances last)
        const float inv = (ss > 0.f && ss < INFINITY) ? 1.0f / sqrtf(ss) : 0.f;
        inv_norm[row] = inv;
        ids_dst[row] = ids_src ? ids_src[row] : id_first + row;
    }
}

int check_store(const dab_store *s, const char *who) {
    if (!s) return fail(DAB_E_INVALID, std::string(who) + ": store is NULL");
    return DAB_OK;
}

}  // namespace
}  // namespace dab

using namespace dab;

extern "C" DAB_API const char *dab_version(void) { return "dab_b200 0.1 sm_100a"; }
extern "C" DAB_API const char *dab_last_error(void) { return g_last_error.c_str(); }

extern "C" DAB_API int dab_store_create(dab_store **out, int device, int64_t capacity_rows, int dim, int dtype) {
    if (!out) return fail(DAB_E_INVALID, "dab_store_create: out is NULL");
    *out = nullptr;
    if (dtype != DAB_F32 && dtype != DAB_BF16) return fail(DAB_E_INVALID, "dab_store_create: dtype must be DAB_F32 or DAB_BF16");
    const int ve = dtype == DAB_BF16 ? 8 : 4;
    if (dim <= 0 || dim % ve) return fail(DAB_E_INVALID, "dab_store_create: dim must be a positive multiple of 4 (fp32) / 8 (bf16)");
    const size_t row_bytes = (size_t)dim * (dtype == DAB_BF16 ? 2 : 4);
    if (row_bytes > DAB_MAX_ROW_BYTES) return fail(DAB_E_INVALID, "dab_store_create: row wider than DAB_MAX_ROW_BYTES");
    // row indices travel as int32 TMA coordinates and as the low 32 bits of a candidate key
    if (capacity_rows <= 0 || capacity_rows > (int64_t)0x7fffffffll) return fail(DAB_E_INVALID, "dab_store_create: capacity must be in 1..2^31-1");
    DAB_CUDA_OK(cudaSetDevice(device));
    dab_store *s = new (std::nothrow) dab_store();
    if (!s) return fail(DAB_E_INVALID, "dab_store_create: out of host memory");
    s->device = device;
    s->dim = dim;
    s->dtype = dtype;
    s->capacity = capacity_rows;
    cudaError_t e;
    auto bail = [&](const char *what, cudaError_t err) {
        std::string m = std::string("dab_store_create: ") + what + ": " + cudaGetErrorString(err);
        dab_store_destroy(s);
        return fail(DAB_E_CUDA, m);
    };
    // tile reads may run past the last row by < one tile of inverse norms: pad
    const size_t pad_rows = 256;
    if ((e = cudaMalloc(&s->rows, (size_t)capacity_rows * row_bytes)) != cudaSuccess) return bail("rows", e);
    if ((e = cudaMalloc(&s->inv_norm, ((size_t)capacity_rows + pad_rows) * sizeof(float))) != cudaSuccess) return bail("inv_norm", e);
    if ((e = cudaMemset(s->inv_norm, 0, ((size_t)capacity_rows + pad_rows) * sizeof(float))) != cudaSuccess) return bail("memset", e);
    if ((e = cudaMalloc(&s->ids, (size_t)capacity_rows * sizeof(int64_t))) != cudaSuccess) return bail("ids", e);
    if ((e = cudaMalloc(&s->scratch_base, search_scratch_bytes(dim))) != cudaSuccess) return bail("scratch", e);
    if ((e = cudaMemset(s->scratch_base, 0, search_scratch_bytes(dim))) != cudaSuccess) return bail("memset", e);
    s->scratch = carve_scratch(s->scratch_base, dim);
    int sms = 0;
    if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device)) != cudaSuccess) return bail("attr", e);
    s->num_sms = sms;
    if ((e = cudaMalloc(&s->verdict_ring_dev, (kVerdictSlots + 1) * 2 * sizeof(int))) != cudaSuccess) return bail("verdict ring", e);
    if ((e = cudaMemset(s->verdict_ring_dev, 0, (kVerdictSlots + 1) * 2 * sizeof(int))) != cudaSuccess) return bail("memset", e);
    if ((e = cudaHostAlloc(reinterpret_cast<void **>(&s->verdict_ring_host), (kVerdictSlots + 1) * 2 * sizeof(int), cudaHostAllocMapped)) != cudaSuccess)
        return bail("verdict ring (pinned)", e);
    memset(s->verdict_ring_host, 0, (kVerdictSlots + 1) * 2 * sizeof(int));
    *out = s;
    return DAB_OK;
}

extern "C" DAB_API void dab_store_destroy(dab_store *s) {
    if (!s) return;
    cudaSetDevice(s->device);
    cudaFree(s->rows);
    cudaFree(s->inv_norm);
    cudaFree(s->ids);
    cudaFree(s->scratch_base);
    cudaFree(s->verdict_ring_dev);
    if (s->verdict_ring_host) cudaFreeHost(s->verdict_ring_host);
    if (s->scratch_free) cudaEventDestroy(s->scratch_free);
    if (s->io_host) cudaFreeHost(s->io_host);
    cudaFree(s->io_dev);
    for (int c = 0; c < DAB_MAX_COLUMNS; c++) cudaFree(s->columns[c]);
    delete s;
}

extern "C" DAB_API int64_t dab_store_size(const dab_store *s) { return s ? s->size : -1; }
extern "C" DAB_API int64_t dab_store_capacity(const dab_store *s) { return s ? s->capacity : -1; }
extern "C" DAB_API int dab_store_dim(const dab_store *s) { return s ? s->dim : -1; }
extern "C" DAB_API int dab_store_dtype(const dab_store *s) { return s ? s->dtype : -1; }

extern "C" DAB_API int dab_store_reset(dab_store *s) {
    if (int rc = check_store(s, "dab_store_reset")) return rc;
    std::lock_guard<std::mutex> g(s->mu);
    s->size = 0;
    s->explicit_ids = false;
    return DAB_OK;
}

extern "C" DAB_API int dab_store_set_id_base(dab_store *s, int64_t id_base) {
    if (int rc = check_store(s, "dab_store_set_id_base")) return rc;
    std::lock_guard<std::mutex> g(s->mu);
    if (s->size) return fail(DAB_E_INVALID, "dab_store_set_id_base: store already holds rows");
    s->id_base = id_base;
    return DAB_OK;
}

extern "C" DAB_API int dab_store_set_spare_rounds(dab_store *s, int rounds) {
    if (int rc = check_store(s, "dab_store_set_spare_rounds")) return rc;
    if (rounds < 0 || rounds > 64) return fail(DAB_E_INVALID, "dab_store_set_spare_rounds: 0..64");
    std::lock_guard<std::mutex> g(s->mu);
    s->spare_rounds = rounds;
    return DAB_OK;
}

extern "C" DAB_API int dab_store_set_path(dab_store *s, int path) {
    if (int rc = check_store(s, "dab_store_set_path")) return rc;
    if (path < 0 || path > 2) return fail(DAB_E_INVALID, "dab_store_set_path: 0 auto, 1 CUDA cores, 2 tcgen05");
    if (path == 2 && !umma_supported(s->dim, s->dtype)) return fail(DAB_E_INVALID, "dab_store_set_path: the query block of this store does not fit shared memory on the tcgen05 path (rows too wide)");
    std::lock_guard<std::mutex> g(s->mu);
    s->path = path;
    return DAB_OK;
}

extern "C" DAB_API int dab_store_append(dab_store *s, const void *rows_dev, int rows_dtype, const int64_t *ids_dev, int64_t n,
                                void *stream) {
    if (int rc = check_store(s, "dab_store_append")) return rc;
    if (n < 0) return fail(DAB_E_INVALID, "dab_store_append: n < 0");
    if (n == 0) return DAB_OK;
    if (!rows_dev || !aligned16(rows_dev)) return fail(DAB_E_INVALID, "dab_store_append: rows must be a 16-byte aligned device pointer");
    if (rows_dtype != DAB_F32 && rows_dtype != DAB_BF16) return fail(DAB_E_INVALID, "dab_store_append: rows_dtype must be DAB_F32 or DAB_BF16");
    if (rows_dtype == DAB_BF16 && s->dtype == DAB_F32) return fail(DAB_E_INVALID, "dab_store_append: bf16 rows into an fp32 store are not widened; append fp32");
    std::lock_guard<std::mutex> g(s->mu);
    if (s->size + n > s->capacity) return fail(DAB_E_CAPACITY, "dab_store_append: store is full");
    if (s->size > 0 && (ids_dev != nullptr) != s->explicit_ids)
        return fail(DAB_E_INVALID, "dab_store_append: explicit and implicit ids cannot be mixed in one store");
    DAB_CUDA_OK(cudaSetDevice(s->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t row_bytes = (size_t)s->dim * (s->dtype == DAB_BF16 ? 2 : 4);
    void *dst = static_cast<char *>(s->rows) + (size_t)s->size * row_bytes;
    float *inv = s->inv_norm + s->size;
    int64_t *idd = s->ids + s->size;
    const int64_t id_first = s->id_base + s->size;
    const unsigned blocks = (unsigned)((n + 7) / 8);
    if (rows_dtype == DAB_BF16)
        append_rows_kernel<true, true><<<blocks, 256, 0, st>>>(rows_dev, dst, inv, idd, ids_dev, id_first, n, s->dim);
    else if (s->dtype == DAB_F32)
        append_rows_kernel<false, false><<<blocks, 256, 0, st>>>(rows_dev, dst, inv, idd, ids_dev, id_first, n, s->dim);
    else
        append_rows_kernel<false, true><<<blocks, 256, 0, st>>>(rows_dev, dst, inv, idd, ids_dev, id_first, n, s->dim);
    DAB_LAUNCH_OK();
    s->explicit_ids = ids_dev != nullptr;
    s->size += n;
    return DAB_OK;
}

namespace dab {

// Brackets one kernel launch with cudaEvents on its own stream when profiling is on.
struct ScanTimer {
    dab_profile &p;
    cudaStream_t st;
    cudaEvent_t a = nullptr, b = nullptr;
    ScanTimer(dab_profile &p_, cudaStream_t st_, bool whole_search = false) : p(p_), st(st_), whole(whole_search) {
        if (!p.on) return;
        if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) { a = b = nullptr; return; }
        cudaEventRecord(a, st);
    }
    void stop() {
        if (!a) return;
        cudaEventRecord(b, st);
        (whole ? p.search_events : p.scan_events).emplace_back(a, b);
        a = b = nullptr;
    }
    bool whole;
};

// The rounds of one query chunk (<= 256 queries), all enqueued on `st` without a host sync.
// Everything a search of nq queries / top-k over n rows decides before launching anything.
struct SearchPlan {
    int path;        // 1 = CUDA cores, 2 = tcgen05
    bool quad;       // tcgen05: cluster kernel (cta_group::2 pairs) for query blocks wider than one SM holds
    int pairs;       // cluster kernel: 2 = 4-CTA cluster, tiles multicast to both pairs; 1 = one pair per cluster
    int quad_chunk;  // cluster kernel: queries per launch (the whole block unless it exceeds one cluster's shared memory)
    int k_cap;       // candidates kept per query between levels
    float margin;    // tensor path: approximate-score margin that is kept for rescoring
    LevelPlan levels;
};

SearchPlan choose_plan(const dab_store *s, int path, int64_t n, int dim, int dtype, int nq, int k) {
    SearchPlan sp = {};
    sp.path = path;
    // auto: the tensor path whenever the row layout allows it - measured faster than the CUDA-core kernel at every
    // query count, 1 included (12.5M x 768 bf16: 2.7 ms vs 3.3 / 6.5 / 9.3 ms at Q = 1 / 2 / 4; fp32 via tf32:
    // 2.5 ms vs 2.8 / 4.1 / 5.7 ms), and no slower on stores of a few thousand rows.  The CUDA-core kernel
    // is the exact fallback for DAB_E_INEXACT and serves rows too wide for a resident query block.
    if (sp.path == 0) {
        sp.path = umma_supported(dim, dtype) ? 2 : 1;
        if (sp.path == 2 && s->small_simt_credit > 0 && nq <= simt_queries_per_pass(dim, dtype)) sp.path = 1;
    }
    if (sp.path == 2) {
        // Tensor path.  The pass scores with the query rounded to bf16: per element |dq| <= 2^-8 |q|, so by
        // Cauchy-Schwarz |approx - exact| <= 2^-8 for unit vectors, plus fp32 accumulation noise
        // (<= dim * 2^-24).  Everything within twice that bound of the k-th best approximate score
        // can still belong to the exact top-k: it is kept (up to k_cap) and rescored in fp32.
        // (fp32 stores go through kind::tf32: both operands lose their low 13 mantissa bits, |d| < 2^-10 each,
        // so the product is off by < 2^-9 relative.)
        const float eps = (dtype == DAB_F32 ? 0.001953125f : 0.00390625f) + (float)dim * 6e-8f;
        sp.margin = 2.0f * eps;
        sp.k_cap = kCandCap / 4;  // 2048 kept candidates per query at most
        // wide blocks that do not fit one SM's shared memory go to the 4-CTA cluster kernel
        static const bool no_quad = tuning_env("DAB_NO_QUAD") != nullptr;  // tuning builds only: query-slice groups instead
        const bool wide = nq > umma_queries_per_pass(dim, dtype);
        sp.quad = wide && !no_quad && quad_supported(dim, dtype, nq);
        sp.quad_chunk = nq;
        // fp32 rows are twice as wide: one cluster holds 192 queries at dim 768, so a 256-query block takes two
        // corpus passes of 128 (still far better than slicing it six ways through the single-CTA kernel)
        const int quad_cap = quad_queries_per_pass(dim, dtype);
        if (wide && !no_quad && !sp.quad && quad_cap > 0 && nq > quad_cap) {
            const int passes = (nq + quad_cap - 1) / quad_cap;
            sp.quad = true;
            sp.quad_chunk = ((nq + passes - 1) / passes + 63) / 64 * 64;
        }
        // up to pair_queries_per_pass (128 at dim 768) ONE pair holds the whole block: 74 two-CTA clusters use every
        // SM, every corpus byte goes to one SM, N = 128 MMAs run at their floor - 2.9 ms per 12.5M x 768 shard at
        // Q = 100..128 against 3.6-3.7 ms for the 4-CTA form (whose N = 64 MMAs are issue-bound); two pair passes
        // for Q = 256 cost 6.0 ms against 4.7 ms.  It also beats the single-CTA kernel once that one's epilogue gets
        // long (Q = 80: 2.89 vs 3.00 ms, Q = 96: 2.91 vs 3.25 ms; equal at Q <= 64).  DAB_PAIR_MIN_Q moves the switch.
        static const int pair_min_q = tuning_env("DAB_PAIR_MIN_Q") ? atoi(tuning_env("DAB_PAIR_MIN_Q")) : 65;
        if (!no_quad && nq <= pair_queries_per_pass(dim, dtype) && quad_supported(dim, dtype, nq) &&
            (wide || (dtype == DAB_BF16 && nq >= pair_min_q))) {
            sp.quad = true;
            sp.pairs = 1;
        } else {
            sp.pairs = 2;
        }
        // levels are sized for ~4k admitted rows per k (random unit vectors put ~3k inside the margin)
        sp.levels = make_level_plan(n, sp.quad ? quad_tile_rows() : umma_tile_rows(),
                                    4 * k < sp.k_cap ? 4 * k : sp.k_cap);
    } else {
        sp.margin = 0.f;
        sp.k_cap = k;
        sp.levels = make_level_plan(n, simt_tile_rows(dim, dtype), k);
    }
    return sp;
}

// The rounds of one query chunk (<= 256 queries), all enqueued on `st` without a host sync.
int run_search_chunk(dab_store *s, int path, int spare_rounds, const StoreView &sv, const SearchScratch &sc, const float *q_dev,
                     int nq, int k, const uint32_t *allow, float *out_scores, int64_t *out_ids, cudaStream_t st,
                     bool first_chunk) {
    const SearchPlan sp = choose_plan(s, path, sv.n, sv.dim, sv.dtype, nq, k);
    const LevelPlan &plan = sp.levels;
    if (sp.path == 2) {
        if (int rc = launch_search_begin(sc, q_dev, nq, sv.dim, sv.dtype == DAB_BF16, st, first_chunk)) return rc;
        for (int r = 0; r < plan.num_levels + spare_rounds; r++) {
            {
                // one launch scores all nq queries
                ScanTimer tm(s->prof, st);
                static const bool pair_mode = tuning_env("DAB_WIDE_PAIRS") != nullptr;  // tuning builds only
                if (sp.quad && sp.pairs == 1) {
                    if (int rc = launch_scan_umma_quad(sv, allow, sc, 0, nq, plan, s->num_sms, st, 1)) return rc;
                } else if (sp.quad && sp.quad_chunk < nq) {
                    for (int q0 = 0; q0 < nq; q0 += sp.quad_chunk)
                        if (int rc = launch_scan_umma_quad(sv, allow, sc, q0, nq - q0 < sp.quad_chunk ? nq - q0 : sp.quad_chunk, plan,
                                                           s->num_sms, st))
                            return rc;
                } else if (sp.quad && pair_mode) {
                    const int per = pair_queries_per_pass(sv.dim, sv.dtype);
                    for (int q0 = 0; q0 < nq; q0 += per)
                        if (int rc = launch_scan_umma_quad(sv, allow, sc, q0, nq - q0 < per ? nq - q0 : per, plan, s->num_sms, st, 1))
                            return rc;
                } else if (int rc = sp.quad ? launch_scan_umma_quad(sv, allow, sc, 0, nq, plan, s->num_sms, st)
                                            : launch_scan_umma(sv, allow, sc, 0, nq, plan, s->num_sms, st))
                    return rc;
                tm.stop();
            }
            if (int rc = launch_compact_margin(sc, nq, k, sp.k_cap, sp.margin, plan.num_levels, st)) return rc;
        }
        return launch_finish_rescore(sc, sv, nq, k, sp.k_cap, plan.num_levels, out_scores, out_ids, st);
    }
    const int qb = simt_queries_per_pass(sv.dim, sv.dtype);
    if (int rc = launch_search_begin(sc, q_dev, nq, sv.dim, false, st, first_chunk)) return rc;
    for (int r = 0; r < plan.num_levels + spare_rounds; r++) {
        for (int q0 = 0; q0 < nq; q0 += qb) {
            const int n_sub = nq - q0 < qb ? nq - q0 : qb;
            ScanTimer tm(s->prof, st);
            if (int rc = launch_scan_simt(sv, allow, sc, q0, n_sub, plan, s->num_sms, st)) return rc;
            tm.stop();
        }
        if (int rc = launch_compact(sc, nq, k, plan.num_levels, st)) return rc;
    }
    return launch_finish_exact(sc, sv, nq, k, plan.num_levels, out_scores, out_ids, st);
}

int validate_search(const dab_store *s, const float *q_dev, int Q, int k, const float *out_scores, const int64_t *out_ids) {
    if (Q < 0) return fail(DAB_E_INVALID, "dab_store_search: Q < 0");
    if (k < 1 || k > DAB_MAX_K) return fail(DAB_E_INVALID, "dab_store_search: k must be in 1..DAB_MAX_K");
    if (Q == 0) return DAB_OK;
    if (!q_dev || !out_scores || !out_ids) return fail(DAB_E_INVALID, "dab_store_search: NULL pointer");
    if (!aligned16(q_dev)) return fail(DAB_E_INVALID, "dab_store_search: queries must be 16-byte aligned");
    (void)s;
    return DAB_OK;
}

}  // namespace dab

namespace dab {
// One search call: all query chunks enqueued on `st` under the store's mutex.  `path` / `spare_rounds` < 0 = the
// store's settings.  The call's verdict lands in slot seq % kVerdictSlots of the store's rings.
static int search_impl(dab_store *s, const float *q_dev, int Q, int k, const uint32_t *allow, float *out_scores, int64_t *out_ids,
                       int path, int spare_rounds, cudaStream_t st) {
    std::lock_guard<std::mutex> g(s->mu);
    DAB_CUDA_OK(cudaSetDevice(s->device));
    if (path < 0) path = s->path;
    if (spare_rounds < 0) spare_rounds = s->spare_rounds;
    t_last.store = s;
    t_last.seq = ++s->search_seq;
    t_last.empty = s->size == 0;
    t_last.small_tensor = false;
    if (s->size == 0) return launch_fill_empty(out_scores, out_ids, (int64_t)Q * k, st);
    StoreView sv;
    sv.rows = s->rows;
    sv.inv_norm = s->inv_norm;
    sv.ids = s->explicit_ids ? s->ids : nullptr;
    sv.id_base = s->id_base;
    sv.n = s->size;
    sv.dim = s->dim;
    sv.dtype = s->dtype;
    SearchScratch sc = s->scratch;
    const int slot = (int)(t_last.seq % kVerdictSlots);
    sc.verdict_dev = s->verdict_ring_dev + 2 * slot;
    sc.verdict_host = s->verdict_ring_host + 2 * slot;
    if (!s->scratch_free) DAB_CUDA_OK(cudaEventCreateWithFlags(&s->scratch_free, cudaEventDisableTiming));
    DAB_CUDA_OK(cudaStreamWaitEvent(st, s->scratch_free, 0));   // no-op for the stream that recorded it last
    const bool small = Q <= simt_queries_per_pass(s->dim, s->dtype);
    t_last.small_tensor = path == 0 && small && s->small_simt_credit == 0 && umma_supported(s->dim, s->dtype);
    ScanTimer whole(s->prof, st, true);
    struct Stop {
        ScanTimer &t;
        ~Stop() { t.stop(); }
    } stop_on_exit{whole};
    for (int q0 = 0; q0 < Q; q0 += DAB_MAX_QUERY_BATCH) {
        const int nq = Q - q0 < DAB_MAX_QUERY_BATCH ? Q - q0 : DAB_MAX_QUERY_BATCH;
        if (int rc = run_search_chunk(s, path, spare_rounds, sv, sc, q_dev + (size_t)q0 * s->dim, nq, k, allow,
                                      out_scores + (size_t)q0 * k, out_ids + (size_t)q0 * k, st, q0 == 0))
            return rc;
    }
    DAB_CUDA_OK(cudaEventRecord(s->scratch_free, st));
    if (path == 0 && small && s->small_simt_credit > 0) s->small_simt_credit--;   // consumed after the plan was chosen
    return DAB_OK;
}

// Verdict of this thread's last search on `s` (after `st` has been synchronised).
static int read_verdict(dab_store *s) {
    if (t_last.store != s || t_last.empty) return DAB_OK;   // nothing searched by this thread, or the store was empty
    const volatile int *v = s->verdict_ring_host + 2 * (t_last.seq % kVerdictSlots);
    const int incomplete = v[0], inexact = v[1];
    if (incomplete)
        return fail(DAB_E_INCOMPLETE, "search incomplete: a scan round overflowed its candidate buffer more often than the spare "
                                      "rounds allow; raise dab_store_set_spare_rounds and search again");
    if (inexact) {
        if (t_last.small_tensor) {
            std::lock_guard<std::mutex> g(s->mu);
            s->small_simt_credit = 64;
        }
        return fail(DAB_E_INEXACT, "more than " + std::to_string(kCandCap / 4) +
                                       " rows lie within the bf16 rescoring margin of the k-th best; search again on the CUDA-core path");
    }
    return DAB_OK;
}
}  // namespace dab

extern "C" DAB_API int dab_store_search(dab_store *s, const float *q_dev, int Q, int k, const uint32_t *allow, float *out_scores,
                                int64_t *out_ids, void *stream) {
    return dab_store_search_ex(s, q_dev, Q, k, allow, out_scores, out_ids, -1, -1, stream);
}

extern "C" DAB_API int dab_store_search_ex(dab_store *s, const float *q_dev, int Q, int k, const uint32_t *allow, float *out_scores,
                                           int64_t *out_ids, int path, int spare_rounds, void *stream) {
    if (int rc = check_store(s, "dab_store_search")) return rc;
    if (int rc = validate_search(s, q_dev, Q, k, out_scores, out_ids)) return rc;
    if (path > 2 || spare_rounds > 64) return fail(DAB_E_INVALID, "dab_store_search_ex: path must be -1..2, spare_rounds -1..64");
    if (path == 2 && !umma_supported(s->dim, s->dtype)) return fail(DAB_E_INVALID, "dab_store_search_ex: rows too wide for the tcgen05 path");
    if (Q == 0) return DAB_OK;
    return search_impl(s, q_dev, Q, k, allow, out_scores, out_ids, path, spare_rounds, static_cast<cudaStream_t>(stream));
}

extern "C" DAB_API int dab_store_search_status(dab_store *s, void *stream) {
    if (int rc = check_store(s, "dab_store_search_status")) return rc;
    DAB_CUDA_OK(cudaSetDevice(s->device));
    // the verdict was written to pinned memory by the search's last kernel: waiting for the stream is all it takes
    DAB_CUDA_OK(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
    return read_verdict(s);
}

extern "C" DAB_API uint64_t dab_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" DAB_API int dab_store_set_profile(dab_store *s, int on) {
    if (int rc = check_store(s, "dab_store_set_profile")) return rc;
    std::lock_guard<std::mutex> g(s->mu);
    s->prof.on = on != 0;
    return DAB_OK;
}

extern "C" DAB_API int dab_store_get_profile(dab_store *s, float *scan_ms, int *scan_launches, float *search_ms,
                                             int *searches) {
    if (int rc = check_store(s, "dab_store_get_profile")) return rc;
    std::lock_guard<std::mutex> g(s->mu);
    DAB_CUDA_OK(cudaSetDevice(s->device));
    DAB_CUDA_OK(cudaDeviceSynchronize());
    auto drain = [](std::vector<std::pair<cudaEvent_t, cudaEvent_t>> &v, float *ms, int *n) {
        float tot = 0.f;
        for (auto &e : v) {
            float t = 0.f;
            if (cudaEventElapsedTime(&t, e.first, e.second) == cudaSuccess) tot += t;
            cudaEventDestroy(e.first);
            cudaEventDestroy(e.second);
        }
        if (ms) *ms = tot;
        if (n) *n = (int)v.size();
        v.clear();
    };
    drain(s->prof.scan_events, scan_ms, scan_launches);
    drain(s->prof.search_events, search_ms, searches);
    return DAB_OK;
}

extern "C" DAB_API int dab_store_describe_plan(dab_store *s, int Q, int k, int *path, int *tile_rows, int *num_levels, int *growth,
                                               int *strides /* [8] */) {
    if (int rc = check_store(s, "dab_store_describe_plan")) return rc;
    if (Q < 1 || k < 1 || k > DAB_MAX_K) return fail(DAB_E_INVALID, "dab_store_describe_plan: bad Q or k");
    std::lock_guard<std::mutex> g(s->mu);
    const int nq = Q < DAB_MAX_QUERY_BATCH ? Q : DAB_MAX_QUERY_BATCH;
    const SearchPlan sp = choose_plan(s, s->path, s->size > 0 ? s->size : 1, s->dim, s->dtype, nq, k);
    if (path) *path = sp.path == 2 && sp.quad ? 3 : sp.path;
    if (tile_rows) *tile_rows = sp.levels.tile_rows;
    if (num_levels) *num_levels = sp.levels.num_levels;
    if (growth) *growth = sp.levels.growth;
    if (strides)
        for (int l = 0; l < kMaxLevels; l++) strides[l] = l < sp.levels.num_levels ? sp.levels.stride[l] : 0;
    return DAB_OK;
}

// device address of the {incomplete, inexact} verdict of the calling thread's last search on `s` (for
// dab_exchange_merge_topk, enqueued behind that search); before any search: a slot that reads {0, 0}
extern "C" DAB_API const int *dab_store_status_dev(dab_store *s) {
    if (!s) return nullptr;
    if (t_last.store != s || t_last.empty) return s->verdict_ring_dev + 2 * kVerdictSlots;   // the extra slot no search writes
    return s->verdict_ring_dev + 2 * (t_last.seq % kVerdictSlots);
}

// ---------------------------------------------------------------- host-buffer search (the e2e entry point)
namespace dab {
struct IoLayout {
    size_t q, scores, ids, total;
};
static IoLayout io_layout(int dim) {
    IoLayout l;
    auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
    l.q = 0;
    l.scores = up((size_t)DAB_MAX_QUERY_BATCH * dim * sizeof(float));
    l.ids = l.scores + up((size_t)DAB_MAX_QUERY_BATCH * DAB_MAX_K * sizeof(float));
    l.total = l.ids + up((size_t)DAB_MAX_QUERY_BATCH * DAB_MAX_K * sizeof(int64_t));
    return l;
}
}  // namespace dab

extern "C" DAB_API int dab_store_search_host(dab_store *s, const float *q_host, int Q, int k, const uint32_t *allow_bitmap_dev,
                                             float *out_scores_host, int64_t *out_ids_host, void *stream) {
    if (int rc = check_store(s, "dab_store_search_host")) return rc;
    if (Q < 0) return fail(DAB_E_INVALID, "dab_store_search_host: Q < 0");
    if (k < 1 || k > DAB_MAX_K) return fail(DAB_E_INVALID, "dab_store_search_host: k must be in 1..DAB_MAX_K");
    if (Q == 0) return DAB_OK;
    if (!q_host || !out_scores_host || !out_ids_host) return fail(DAB_E_INVALID, "dab_store_search_host: NULL pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const IoLayout L = io_layout(s->dim);
    std::lock_guard<std::mutex> io_guard(s->io_mu);
    {
        std::lock_guard<std::mutex> g(s->mu);
        DAB_CUDA_OK(cudaSetDevice(s->device));
        if (!s->io_host) {
            DAB_CUDA_OK(cudaHostAlloc(&s->io_host, L.total, cudaHostAllocDefault));
            DAB_CUDA_OK(cudaMalloc(&s->io_dev, L.total));
        }
    }
    char *hbase = static_cast<char *>(s->io_host), *dbase = static_cast<char *>(s->io_dev);
    int rc = DAB_OK;
    for (int q0 = 0; q0 < Q && rc == DAB_OK; q0 += DAB_MAX_QUERY_BATCH) {
        const int nq = Q - q0 < DAB_MAX_QUERY_BATCH ? Q - q0 : DAB_MAX_QUERY_BATCH;
        memcpy(hbase + L.q, q_host + (size_t)q0 * s->dim, (size_t)nq * s->dim * sizeof(float));
        int spare = -1;             // -1 = the store's settings (dab_store_set_spare_rounds / dab_store_set_path)
        int path = -1;
        bool forced = false;
        int nrun = nq;              // queries in the staging block: the chunk, or its poisoned queries on the exact redo
        std::vector<int> redo;
        for (;;) {
            if ((rc = cudaMemcpyAsync(dbase + L.q, hbase + L.q, (size_t)nrun * s->dim * sizeof(float), cudaMemcpyHostToDevice, st)) != cudaSuccess) {
                rc = fail(DAB_E_CUDA, std::string("dab_store_search_host: H2D: ") + cudaGetErrorString((cudaError_t)rc));
                break;
            }
            rc = dab_store_search_ex(s, reinterpret_cast<const float *>(dbase + L.q), nrun, k, allow_bitmap_dev,
                                     reinterpret_cast<float *>(dbase + L.scores), reinterpret_cast<int64_t *>(dbase + L.ids), path, spare, st);
            if (rc != DAB_OK) break;
            cudaMemcpyAsync(hbase + L.scores, dbase + L.scores, (size_t)nrun * k * sizeof(float), cudaMemcpyDeviceToHost, st);
            cudaMemcpyAsync(hbase + L.ids, dbase + L.ids, (size_t)nrun * k * sizeof(int64_t), cudaMemcpyDeviceToHost, st);
            rc = dab_store_search_status(s, st);   // synchronises the stream: results and verdict are on the host
            if (rc == DAB_E_INEXACT && !forced) {
                // Only the queries whose margin overflowed came back poisoned (NaN score in front); the others are
                // final.  Keep those, pack the poisoned ones to the front of the staging block and answer them on
                // the exact CUDA-core path.
                forced = true;
                path = 1;
                const float *hs = reinterpret_cast<const float *>(hbase + L.scores);
                const int64_t *hi = reinterpret_cast<const int64_t *>(hbase + L.ids);
                redo.clear();
                for (int q = 0; q < nq; q++) {
                    if (hs[(size_t)q * k] != hs[(size_t)q * k]) {
                        redo.push_back(q);
                    } else {
                        memcpy(out_scores_host + (size_t)(q0 + q) * k, hs + (size_t)q * k, (size_t)k * sizeof(float));
                        memcpy(out_ids_host + (size_t)(q0 + q) * k, hi + (size_t)q * k, (size_t)k * sizeof(int64_t));
                    }
                }
                if (redo.empty() || (int)redo.size() == nq) {   // nothing to tell apart: redo the whole chunk
                    redo.clear();
                    continue;
                }
                for (size_t j = 0; j < redo.size(); j++)
                    memcpy(hbase + L.q + j * s->dim * sizeof(float), q_host + (size_t)(q0 + redo[j]) * s->dim, (size_t)s->dim * sizeof(float));
                nrun = (int)redo.size();
                continue;
            }
            if (rc == DAB_E_INCOMPLETE && spare < 64) {   // adversarial row order: retry with spare rounds
                spare = spare <= 0 ? 4 : spare * 2;
                continue;
            }
            break;
        }
        if (rc == DAB_OK && redo.empty()) {
            memcpy(out_scores_host + (size_t)q0 * k, hbase + L.scores, (size_t)nq * k * sizeof(float));
            memcpy(out_ids_host + (size_t)q0 * k, hbase + L.ids, (size_t)nq * k * sizeof(int64_t));
        } else if (rc == DAB_OK) {
            for (size_t j = 0; j < redo.size(); j++) {
                memcpy(out_scores_host + (size_t)(q0 + redo[j]) * k, hbase + L.scores + j * k * sizeof(float), (size_t)k * sizeof(float));
                memcpy(out_ids_host + (size_t)(q0 + redo[j]) * k, hbase + L.ids + j * k * sizeof(int64_t), (size_t)k * sizeof(int64_t));
            }
        }
    }
    return rc;
}

// ---------------------------------------------------------------- maintenance: tombstones and export
namespace dab {
namespace {
__global__ void delete_rows_kernel(float *__restrict__ inv_norm, const int64_t *__restrict__ rows, int64_t n, int64_t size) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const int64_t r = rows[i];
        // an inverse norm that is not > 0 is never returned by any scan; the sign bit tells a tombstone (-0.0)
        // from a genuine zero vector (+0.0), which dab_store_compact must keep
        if (r >= 0 && r < size) inv_norm[r] = __int_as_float((int)0x80000000);
    }
}
}  // namespace
}  // namespace dab

extern "C" DAB_API int dab_store_delete_rows(dab_store *s, const int64_t *rows_dev, int64_t n, void *stream) {
    if (int rc = check_store(s, "dab_store_delete_rows")) return rc;
    if (n < 0 || (n > 0 && !rows_dev)) return fail(DAB_E_INVALID, "dab_store_delete_rows: bad argument");
    if (n == 0) return DAB_OK;
    std::lock_guard<std::mutex> g(s->mu);
    DAB_CUDA_OK(cudaSetDevice(s->device));
    delete_rows_kernel<<<(unsigned)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(s->inv_norm, rows_dev, n, s->size);
    DAB_LAUNCH_OK();
    return DAB_OK;
}

extern "C" DAB_API int dab_store_export(dab_store *s, int64_t row0, int64_t n, void *rows_out_dev, float *inv_norm_out_dev,
                                        int64_t *ids_out_dev, void *stream) {
    if (int rc = check_store(s, "dab_store_export")) return rc;
    std::lock_guard<std::mutex> g(s->mu);
    if (row0 < 0 || n < 0 || row0 + n > s->size) return fail(DAB_E_INVALID, "dab_store_export: row range outside the store");
    if (n == 0) return DAB_OK;
    DAB_CUDA_OK(cudaSetDevice(s->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t row_bytes = (size_t)s->dim * (s->dtype == DAB_BF16 ? 2 : 4);
    if (rows_out_dev)
        DAB_CUDA_OK(cudaMemcpyAsync(rows_out_dev, static_cast<char *>(s->rows) + (size_t)row0 * row_bytes, (size_t)n * row_bytes,
                                    cudaMemcpyDeviceToDevice, st));
    if (inv_norm_out_dev)
        DAB_CUDA_OK(cudaMemcpyAsync(inv_norm_out_dev, s->inv_norm + row0, (size_t)n * sizeof(float), cudaMemcpyDeviceToDevice, st));
    if (ids_out_dev)
        DAB_CUDA_OK(cudaMemcpyAsync(ids_out_dev, s->ids + row0, (size_t)n * sizeof(int64_t), cudaMemcpyDeviceToDevice, st));
    return DAB_OK;
}


// ---------------------------------------------------------------------------------------------
// Physical compaction: drop the tombstones, keep the order of the surviving rows.
// Positions come from one exclusive scan of the "alive" flags; rows then move chunk by chunk
// through a staging buffer (gather the chunk's survivors, then copy them down), because the
// destination range [pos[c0], pos[c1]) may overlap the chunk's own source range but never a later one.
namespace dab {
namespace {

constexpr uint32_t kTombstoneBits = 0x80000000u;

__global__ void alive_flags_kernel(const float *__restrict__ inv_norm, int64_t n, int *__restrict__ flags) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flags[i] = __float_as_uint(inv_norm[i]) != kTombstoneBits ? 1 : 0;
    else if (i == n) flags[i] = 0;   // pos[n] = survivors
}

// one warp per row of the chunk [c0, c1): survivors go to stage[pos[i] - pos[c0]]
__global__ void compact_gather_kernel(const unsigned char *__restrict__ rows, const float *__restrict__ inv_norm,
                                      const int64_t *__restrict__ ids, const int *__restrict__ flags, const int *__restrict__ pos,
                                      int64_t c0, int64_t c1, size_t row_bytes, unsigned char *__restrict__ stage_rows,
                                      float *__restrict__ stage_inv, int64_t *__restrict__ stage_ids) {
    const int lane = threadIdx.x & 31;
    const int64_t i = c0 + (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= c1 || !flags[i]) return;
    const int64_t j = pos[i] - pos[c0];
    const uint4 *src = reinterpret_cast<const uint4 *>(rows + (size_t)i * row_bytes);
    uint4 *dst = reinterpret_cast<uint4 *>(stage_rows + (size_t)j * row_bytes);
    for (size_t u = lane; u < row_bytes / sizeof(uint4); u += 32) dst[u] = src[u];
    if (lane == 0) {
        stage_inv[j] = inv_norm[i];
        stage_ids[j] = ids[i];
    }
}

__global__ void compact_scatter_kernel(unsigned char *__restrict__ rows, float *__restrict__ inv_norm, int64_t *__restrict__ ids,
                                       const int *__restrict__ pos, int64_t c0, int64_t c1, size_t row_bytes,
                                       const unsigned char *__restrict__ stage_rows, const float *__restrict__ stage_inv,
                                       const int64_t *__restrict__ stage_ids) {
    const int64_t d0 = pos[c0];
    const int64_t cnt = pos[c1] - d0;
    const size_t units = (size_t)cnt * row_bytes / sizeof(uint4);
    const uint4 *src = reinterpret_cast<const uint4 *>(stage_rows);
    uint4 *dst = reinterpret_cast<uint4 *>(rows + (size_t)d0 * row_bytes);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (size_t u = t; u < units; u += stride) dst[u] = src[u];
    for (size_t j = t; j < (size_t)cnt; j += stride) {
        inv_norm[d0 + j] = stage_inv[j];
        ids[d0 + j] = stage_ids[j];
    }
}

// filter columns follow the same survivors (one int64 per row)
__global__ void compact_gather_col_kernel(const int64_t *__restrict__ col, const int *__restrict__ flags, const int *__restrict__ pos,
                                          int64_t c0, int64_t c1, int64_t *__restrict__ stage) {
    const int64_t i = c0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < c1 && flags[i]) stage[pos[i] - pos[c0]] = col[i];
}
__global__ void compact_scatter_col_kernel(int64_t *__restrict__ col, const int *__restrict__ pos, int64_t c0, int64_t c1,
                                           const int64_t *__restrict__ stage) {
    const int64_t d0 = pos[c0];
    const int64_t cnt = pos[c1] - d0;
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < cnt) col[d0 + j] = stage[j];
}

__global__ void compact_tail_kernel(float *__restrict__ inv_norm, int64_t *__restrict__ ids, const int *__restrict__ pos, int64_t n,
                                    int64_t id_base, int renumber) {
    const int64_t kept = pos[n];
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (i >= kept) inv_norm[i] = 0.f;              // rows past the new end read as zero vectors
    else if (renumber) ids[i] = id_base + i;       // implicit ids stay "id_base + row"
}

struct CompactTemps {
    int *flags = nullptr, *pos = nullptr;
    void *cub_tmp = nullptr;
    unsigned char *stage_rows = nullptr;
    float *stage_inv = nullptr;
    int64_t *stage_ids = nullptr;
    ~CompactTemps() {
        cudaFree(flags);
        cudaFree(pos);
        cudaFree(cub_tmp);
        cudaFree(stage_rows);
        cudaFree(stage_inv);
        cudaFree(stage_ids);
    }
};

}  // namespace
}  // namespace dab

extern "C" DAB_API int dab_store_compact(dab_store *s, int64_t *removed_out, void *stream) {
    using namespace dab;
    if (int rc = check_store(s, "dab_store_compact")) return rc;
    std::lock_guard<std::mutex> g(s->mu);
    if (removed_out) *removed_out = 0;
    const int64_t n = s->size;
    if (n == 0) return DAB_OK;
    if (n >= (int64_t)1 << 31) return fail(DAB_E_CAPACITY, "dab_store_compact: more than 2^31 - 1 rows");
    DAB_CUDA_OK(cudaSetDevice(s->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (s->scratch_free) DAB_CUDA_OK(cudaStreamWaitEvent(st, s->scratch_free, 0));   // after every search already enqueued
    const size_t row_bytes = (size_t)s->dim * (s->dtype == DAB_BF16 ? 2 : 4);
    int64_t chunk = (int64_t)((size_t)256 << 20) / (int64_t)row_bytes;   // <= 256 MiB of staging
    if (chunk > 131072) chunk = 131072;
    if (chunk < 1) chunk = 1;
    if (chunk > n) chunk = n;
    CompactTemps t;
    DAB_CUDA_OK(cudaMalloc(&t.flags, (size_t)(n + 1) * sizeof(int)));
    DAB_CUDA_OK(cudaMalloc(&t.pos, (size_t)(n + 1) * sizeof(int)));
    DAB_CUDA_OK(cudaMalloc(&t.stage_rows, (size_t)chunk * row_bytes));
    DAB_CUDA_OK(cudaMalloc(&t.stage_inv, (size_t)chunk * sizeof(float)));
    DAB_CUDA_OK(cudaMalloc(&t.stage_ids, (size_t)chunk * sizeof(int64_t)));
    size_t cub_bytes = 0;
    DAB_CUDA_OK(cub::DeviceScan::ExclusiveSum(nullptr, cub_bytes, t.flags, t.pos, (int)(n + 1), st));
    DAB_CUDA_OK(cudaMalloc(&t.cub_tmp, cub_bytes ? cub_bytes : 1));
    alive_flags_kernel<<<(unsigned)((n + 1 + 255) / 256), 256, 0, st>>>(s->inv_norm, n, t.flags);
    DAB_LAUNCH_OK();
    DAB_CUDA_OK(cub::DeviceScan::ExclusiveSum(t.cub_tmp, cub_bytes, t.flags, t.pos, (int)(n + 1), st));
    unsigned char *rows = static_cast<unsigned char *>(s->rows);
    // dab_store_create only admits rows that are a multiple of 16 bytes, and the base is 256-byte aligned
    for (int64_t c0 = 0; c0 < n; c0 += chunk) {
        const int64_t c1 = c0 + chunk < n ? c0 + chunk : n;
        compact_gather_kernel<<<(unsigned)((c1 - c0 + 7) / 8), 256, 0, st>>>(rows, s->inv_norm, s->ids, t.flags, t.pos, c0, c1, row_bytes,
                                                                            t.stage_rows, t.stage_inv, t.stage_ids);
        compact_scatter_kernel<<<(unsigned)(4 * s->num_sms), 256, 0, st>>>(rows, s->inv_norm, s->ids, t.pos, c0, c1, row_bytes,
                                                                          t.stage_rows, t.stage_inv, t.stage_ids);
        DAB_LAUNCH_OK();
        for (int c = 0; c < DAB_MAX_COLUMNS; c++) {
            if (!s->columns[c]) continue;
            const unsigned blocks = (unsigned)((c1 - c0 + 255) / 256);
            compact_gather_col_kernel<<<blocks, 256, 0, st>>>(s->columns[c], t.flags, t.pos, c0, c1, t.stage_ids);
            compact_scatter_col_kernel<<<blocks, 256, 0, st>>>(s->columns[c], t.pos, c0, c1, t.stage_ids);
            DAB_LAUNCH_OK();
        }
    }
    compact_tail_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(s->inv_norm, s->ids, t.pos, n, s->id_base, s->explicit_ids ? 0 : 1);
    DAB_LAUNCH_OK();
    int kept = 0;
    DAB_CUDA_OK(cudaMemcpyAsync(&kept, t.pos + n, sizeof(int), cudaMemcpyDeviceToHost, st));
    DAB_CUDA_OK(cudaStreamSynchronize(st));   // the new size is host state: this call is synchronous
    s->size = kept;
    if (removed_out) *removed_out = n - kept;
    return DAB_OK;
}
