// filter.cu - queryset filters evaluated on the device.
//
// Every search the reference issues runs over a filtered queryset
// (assistant/bot/services/context_service/steps/embeddings.py:26-29: bot and processing status;
// assistant/processing/documents/steps/questions.py:124: document__id__lt).  PostgreSQL applies such a
// WHERE clause row by row inside the scan.  Here the tested columns sit in HBM beside the rows and one
// streaming kernel evaluates the conjunction for every row, writing the allow bitmap the scan kernels read
// (one warp ballot = one 32-bit word).  HBM-bound: 8 bytes per row and referenced column + 4 bytes of
// inverse norm in, 1 bit out; 100M rows and two columns = 2 GB = ~0.3 ms against a 21 ms single-query scan.
#include "store.cuh"

namespace dab {
namespace {

struct PredArgs {
    const int64_t *col[DAB_MAX_PREDICATES];   // column values, or nullptr = the row id
    const int64_t *set[DAB_MAX_PREDICATES];
    int64_t set_n[DAB_MAX_PREDICATES];
    int64_t value[DAB_MAX_PREDICATES];
    int op[DAB_MAX_PREDICATES];
    int n;
};

__device__ __forceinline__ bool in_sorted(const int64_t *__restrict__ set, int64_t n, int64_t v) {
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        const int64_t x = set[mid];
        if (x < v) lo = mid + 1; else hi = mid;
    }
    return lo < n && set[lo] == v;
}

__global__ void __launch_bounds__(256)
build_allow_kernel(const __grid_constant__ PredArgs pa, const float *__restrict__ inv_norm, const int64_t *__restrict__ ids,
                   int64_t id_base, int64_t n_rows, uint32_t *__restrict__ bitmap) {
    const int64_t words = (n_rows + 31) >> 5;
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t w = warp0; w < words; w += nwarps) {
        const int64_t r = (w << 5) + lane;
        bool ok = false;
        if (r < n_rows) {
            ok = __float_as_uint(inv_norm[r]) != 0x80000000u;   // -0.0 marks a tombstone (dab_store_delete_rows)
            for (int p = 0; p < pa.n && ok; p++) {
                const int64_t v = pa.col[p] ? pa.col[p][r] : (ids ? ids[r] : id_base + r);
                const int64_t x = pa.value[p];
                switch (pa.op[p]) {
                    case DAB_OP_EQ: ok = v == x; break;
                    case DAB_OP_NE: ok = v != x; break;
                    case DAB_OP_LT: ok = v < x; break;
                    case DAB_OP_LE: ok = v <= x; break;
                    case DAB_OP_GT: ok = v > x; break;
                    case DAB_OP_GE: ok = v >= x; break;
                    case DAB_OP_IN: ok = in_sorted(pa.set[p], pa.set_n[p], v); break;
                    default: ok = !in_sorted(pa.set[p], pa.set_n[p], v); break;   // DAB_OP_NOT_IN
                }
            }
        }
        const uint32_t word = __ballot_sync(0xffffffffu, ok);
        if (lane == 0) bitmap[w] = word;
    }
}

// queryset given as a list of store rows: one bit per listed row (tombstones and out-of-range rows are skipped)
__global__ void scatter_allow_kernel(const int64_t *__restrict__ rows, int64_t n, const float *__restrict__ inv_norm, int64_t n_rows,
                                     uint32_t *__restrict__ bitmap) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t r = rows[i];
    if (r < 0 || r >= n_rows || __float_as_uint(inv_norm[r]) == 0x80000000u) return;
    atomicOr(&bitmap[r >> 5], 1u << (r & 31));
}

}  // namespace
}  // namespace dab

using namespace dab;

extern "C" DAB_API int dab_store_set_column(dab_store *s, int column, const int64_t *values_dev, int64_t row0, int64_t n, void *stream) {
    if (!s) return fail(DAB_E_INVALID, "dab_store_set_column: store is NULL");
    if (column < 0 || column >= DAB_MAX_COLUMNS) return fail(DAB_E_INVALID, "dab_store_set_column: column must be in 0..DAB_MAX_COLUMNS-1");
    if (n < 0 || row0 < 0 || (n > 0 && !values_dev)) return fail(DAB_E_INVALID, "dab_store_set_column: bad argument");
    std::lock_guard<std::mutex> g(s->mu);
    if (row0 + n > s->capacity) return fail(DAB_E_CAPACITY, "dab_store_set_column: rows beyond the store's capacity");
    if (n == 0) return DAB_OK;
    DAB_CUDA_OK(cudaSetDevice(s->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (!s->columns[column]) {
        DAB_CUDA_OK(cudaMalloc(&s->columns[column], (size_t)s->capacity * sizeof(int64_t)));
        DAB_CUDA_OK(cudaMemsetAsync(s->columns[column], 0, (size_t)s->capacity * sizeof(int64_t), st));
    }
    DAB_CUDA_OK(cudaMemcpyAsync(s->columns[column] + row0, values_dev, (size_t)n * sizeof(int64_t), cudaMemcpyDeviceToDevice, st));
    return DAB_OK;
}

extern "C" DAB_API int dab_store_build_allow(dab_store *s, const dab_predicate *preds, int n_pred, uint32_t *bitmap_out_dev,
                                             void *stream) {
    if (!s) return fail(DAB_E_INVALID, "dab_store_build_allow: store is NULL");
    if (n_pred < 0 || n_pred > DAB_MAX_PREDICATES) return fail(DAB_E_INVALID, "dab_store_build_allow: at most DAB_MAX_PREDICATES predicates");
    if ((n_pred > 0 && !preds) || !bitmap_out_dev) return fail(DAB_E_INVALID, "dab_store_build_allow: NULL pointer");
    std::lock_guard<std::mutex> g(s->mu);
    if (s->size == 0) return DAB_OK;
    PredArgs pa = {};
    pa.n = n_pred;
    for (int p = 0; p < n_pred; p++) {
        const dab_predicate &d = preds[p];
        if (d.op < DAB_OP_EQ || d.op > DAB_OP_NOT_IN) return fail(DAB_E_INVALID, "dab_store_build_allow: unknown operator");
        if (d.column == DAB_COL_ID) {
            pa.col[p] = nullptr;
        } else if (d.column >= 0 && d.column < DAB_MAX_COLUMNS) {
            if (!s->columns[d.column]) return fail(DAB_E_INVALID, "dab_store_build_allow: column " + std::to_string(d.column) + " was never set");
            pa.col[p] = s->columns[d.column];
        } else {
            return fail(DAB_E_INVALID, "dab_store_build_allow: column must be DAB_COL_ID or 0..DAB_MAX_COLUMNS-1");
        }
        if (d.op >= DAB_OP_IN && (d.set_n < 0 || (d.set_n > 0 && !d.set_dev)))
            return fail(DAB_E_INVALID, "dab_store_build_allow: IN / NOT_IN need an ascending device array");
        pa.op[p] = d.op;
        pa.value[p] = d.value;
        pa.set[p] = d.set_dev;
        pa.set_n[p] = d.set_n;
    }
    DAB_CUDA_OK(cudaSetDevice(s->device));
    const int64_t words = (s->size + 31) / 32;
    int64_t blocks = (words + 7) / 8;                 // 8 warps per block, one word per warp per pass
    const int64_t cap = (int64_t)s->num_sms * 8;      // resident: 8 blocks of 256 threads per SM
    if (blocks > cap) blocks = cap;
    build_allow_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(pa, s->inv_norm, s->explicit_ids ? s->ids : nullptr,
                                                                                       s->id_base, s->size, bitmap_out_dev);
    DAB_LAUNCH_OK();
    return DAB_OK;
}

extern "C" DAB_API int dab_store_allow_from_rows(dab_store *s, const int64_t *rows_dev, int64_t n, uint32_t *bitmap_out_dev, void *stream) {
    if (!s) return fail(DAB_E_INVALID, "dab_store_allow_from_rows: store is NULL");
    if (n < 0 || (n > 0 && !rows_dev) || !bitmap_out_dev) return fail(DAB_E_INVALID, "dab_store_allow_from_rows: bad argument");
    std::lock_guard<std::mutex> g(s->mu);
    if (s->size == 0) return DAB_OK;
    DAB_CUDA_OK(cudaSetDevice(s->device));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    DAB_CUDA_OK(cudaMemsetAsync(bitmap_out_dev, 0, (size_t)((s->size + 31) / 32) * sizeof(uint32_t), st));
    if (n == 0) return DAB_OK;
    scatter_allow_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(rows_dev, n, s->inv_norm, s->size, bitmap_out_dev);
    DAB_LAUNCH_OK();
    return DAB_OK;
}
