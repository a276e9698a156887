// noalloc.cu - dab_search_noalloc: exact cosine top-k over caller-owned rows with NO allocation.
//
// Who this is for.  dab_store_search needs a dab_store, and dab_store_create calls cudaMalloc for
// the rows and for ~17 MB of search scratch.  Some hosts of this library must not do that:
//   * a search issued from inside the embedding service, where the encoder's framework
//     allocator (or a serving runtime started with "use all HBM") has already reserved the
//     device: cudaMalloc there fails, or succeeds only by fragmenting somebody else's pool;
//   * a request handler on a latency budget: cudaMalloc / cudaFree synchronise the whole
//     device, stalling every other stream (the encoder forward of the next request included);
//   * a caller that keeps its embeddings in memory it manages itself (a framework tensor, a
//     mapped file) and only wants them ranked in place.
// Such a caller passes its own row matrix, its own 1/|row| vector and its own output buffers;
// everything the search needs in between lives in ONE file-scope __device__ array below, which
// the CUDA module loader places on every GPU that loads this library and keeps until exit.
//
// The price of a shared, static scratch:
//   * sizes are capped (DAB_NOALLOC_MAX_QUERIES queries, DAB_NOALLOC_MAX_DIM columns); larger
//     requests are refused with DAB_E_CAPACITY rather than served slowly;
//   * every call on a device uses the same bytes.  The array still holds the previous call's
//     candidates; the first kernel of each call resets the counters, thresholds, flags and the
//     control block it is going to read, and nothing else is read before it is written.  Two
//     calls that overlap ON THE GPU (different streams, or different threads/processes' streams
//     on one device) corrupt each other's candidates: serialise them (one stream, or events).
//     A call is a SEQUENCE of launches (begin, then scan + compact per round, then finish); a per-device
//     host mutex is held while that sequence is enqueued, so two host threads that share a stream cannot
//     interleave their launches.  The mutex orders the enqueue only - it does not wait for the GPU - so
//     threads using DIFFERENT streams still have to order those streams themselves.
//   * the first call on each device looks up the array's device address
//     (cudaGetSymbolAddress) and the SM count.  Make that first call outside any stream capture.
//
// It reuses the library's kernels unchanged (search_begin / scan_simt / compact / finish_exact);
// only the place the scratch pointers point to differs.  No cudaMalloc, cudaFree, or any other
// allocating call is made here or in the launchers it uses.
#include <mutex>

#include "search.cuh"

namespace dab {
int simt_queries_per_pass(int dim, int dtype);

namespace {

constexpr int kMaxQ = DAB_NOALLOC_MAX_QUERIES;
constexpr int kMaxDim = DAB_NOALLOC_MAX_DIM;
constexpr size_t up256(size_t x) { return (x + 255) & ~size_t(255); }

// layout of the static scratch (each piece 256-byte aligned)
constexpr size_t kOffQhat = 0;
constexpr size_t kOffCand = kOffQhat + up256((size_t)kMaxQ * kMaxDim * sizeof(float));
constexpr size_t kOffCount = kOffCand + up256((size_t)kMaxQ * kCandCap * sizeof(uint64_t));
constexpr size_t kOffTau = kOffCount + up256(kMaxQ * sizeof(uint32_t));
constexpr size_t kOffActive = kOffTau + up256(kMaxQ * sizeof(uint64_t));
constexpr size_t kOffOverflow = kOffActive + up256(kMaxQ * sizeof(int));
constexpr size_t kOffCtrl = kOffOverflow + up256(kMaxQ * sizeof(int));
constexpr size_t kScratchBytes = kOffCtrl + up256(sizeof(SearchCtrl));
static_assert(kScratchBytes <= DAB_NOALLOC_SCRATCH_BYTES, "header states a smaller scratch than the layout needs");

// The scratch itself.  In host code this name is only a handle for cudaGetSymbolAddress - it is
// NOT a device address and must never be handed to a kernel as a pointer.
static __device__ __align__(256) unsigned char g_noalloc_scratch[DAB_NOALLOC_SCRATCH_BYTES];

struct DeviceSetup {
    std::mutex enqueue_mu;   // one call at a time enqueues its launches on this device (host side; see below)
    bool ready = false;
    unsigned char *scratch = nullptr;  // device address of g_noalloc_scratch on this device
    int num_sms = 0;
};
constexpr int kMaxDevices = 64;
DeviceSetup g_setup[kMaxDevices];
std::mutex g_setup_mu;

// Once per device, with that device current: resolve the device address of the static array.
int setup_current_device(DeviceSetup **out) {
    int dev = -1;
    DAB_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(DAB_E_INVALID, "dab_search_noalloc: device index out of range");
    std::lock_guard<std::mutex> g(g_setup_mu);
    DeviceSetup &d = g_setup[dev];
    if (!d.ready) {
        void *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || ptr == nullptr)
            return fail(DAB_E_CUDA, std::string("dab_search_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
        int sms = 0;
        DAB_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        d.scratch = static_cast<unsigned char *>(ptr);
        d.num_sms = sms;
        d.ready = true;
    }
    *out = &d;
    return DAB_OK;
}

}  // namespace
}  // namespace dab

using namespace dab;

extern "C" DAB_API int dab_search_noalloc(const void *rows_dev, int rows_dtype, const float *inv_norm_dev, int64_t n_rows,
                                          int dim, int64_t id_base, const float *q_dev, int Q, int k,
                                          const uint32_t *allow_bitmap_dev_or_null, float *out_scores_dev,
                                          int64_t *out_ids_dev, void *stream) {
    if (rows_dtype != DAB_F32 && rows_dtype != DAB_BF16) return fail(DAB_E_INVALID, "dab_search_noalloc: rows_dtype must be DAB_F32 or DAB_BF16");
    const int ve = rows_dtype == DAB_BF16 ? 8 : 4;
    if (dim <= 0 || dim % ve) return fail(DAB_E_INVALID, "dab_search_noalloc: dim must be a positive multiple of 4 (fp32) / 8 (bf16)");
    if (dim > kMaxDim)
        return fail(DAB_E_CAPACITY, "dab_search_noalloc: dim " + std::to_string(dim) + " exceeds DAB_NOALLOC_MAX_DIM (" +
                                        std::to_string(kMaxDim) + "); use a dab_store for wider rows");
    if (Q < 0 || k < 1 || k > DAB_MAX_K) return fail(DAB_E_INVALID, "dab_search_noalloc: need Q >= 0 and 1 <= k <= DAB_MAX_K");
    if (Q > kMaxQ)
        return fail(DAB_E_CAPACITY, "dab_search_noalloc: " + std::to_string(Q) + " queries exceed DAB_NOALLOC_MAX_QUERIES (" +
                                        std::to_string(kMaxQ) + "); split the batch or use a dab_store");
    if (n_rows < 0 || n_rows > (int64_t)0x7fffffffll) return fail(DAB_E_INVALID, "dab_search_noalloc: n_rows must be in 0..2^31-1");
    if (Q == 0) return DAB_OK;
    if (!q_dev || !out_scores_dev || !out_ids_dev) return fail(DAB_E_INVALID, "dab_search_noalloc: NULL query/output pointer");
    if (n_rows > 0 && (!rows_dev || !inv_norm_dev)) return fail(DAB_E_INVALID, "dab_search_noalloc: NULL rows/inv_norm pointer");
    if (!aligned16(rows_dev) || !aligned16(inv_norm_dev) || !aligned16(q_dev))
        return fail(DAB_E_INVALID, "dab_search_noalloc: rows, inv_norm and queries must be 16-byte aligned");

    DeviceSetup *d = nullptr;
    if (int rc = setup_current_device(&d)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    std::lock_guard<std::mutex> enqueue_guard(d->enqueue_mu);   // host threads enqueue whole calls, never interleaved launches
    if (n_rows == 0) return launch_fill_empty(out_scores_dev, out_ids_dev, (int64_t)Q * k, st);

    SearchScratch sc{};
    sc.qhat = reinterpret_cast<float *>(d->scratch + kOffQhat);
    sc.qbf16 = nullptr;  // only the tensor path uses it; this entry point scores on CUDA cores
    sc.cand = reinterpret_cast<uint64_t *>(d->scratch + kOffCand);
    sc.cand_count = reinterpret_cast<uint32_t *>(d->scratch + kOffCount);
    sc.tau_key = reinterpret_cast<uint64_t *>(d->scratch + kOffTau);
    sc.active = reinterpret_cast<int *>(d->scratch + kOffActive);
    sc.overflowed = reinterpret_cast<int *>(d->scratch + kOffOverflow);
    sc.ctrl = reinterpret_cast<SearchCtrl *>(d->scratch + kOffCtrl);

    StoreView sv;
    sv.rows = rows_dev;
    sv.inv_norm = inv_norm_dev;
    sv.ids = nullptr;
    sv.id_base = id_base;
    sv.n = n_rows;
    sv.dim = dim;
    sv.dtype = rows_dtype;

    const int qb = simt_queries_per_pass(dim, rows_dtype);
    const LevelPlan plan = make_level_plan(n_rows, simt_tile_rows(dim, rows_dtype), k);
    // search_begin_kernel rewrites, for each of the Q queries, its counter, threshold and flags, and
    // the control block: the only state a round reads before writing.  Stale candidates from the
    // previous call sit beyond the reset counters and are never read.
    if (int rc = launch_search_begin(sc, q_dev, Q, dim, false, st)) return rc;
    for (int r = 0; r < plan.num_levels + DAB_NOALLOC_SPARE_ROUNDS; r++) {
        for (int q0 = 0; q0 < Q; q0 += qb) {
            const int n_sub = Q - q0 < qb ? Q - q0 : qb;
            if (int rc = launch_scan_simt(sv, allow_bitmap_dev_or_null, sc, q0, n_sub, plan, d->num_sms, st)) return rc;
        }
        if (int rc = launch_compact(sc, Q, k, plan.num_levels, st)) return rc;
    }
    return launch_finish_exact(sc, sv, Q, k, plan.num_levels, out_scores_dev, out_ids_dev, st);
}
