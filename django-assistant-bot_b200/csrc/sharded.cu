// sharded.cu - the row-sharded store driven by ONE process.
//
// The reference searches from inside a Django / Celery / FastAPI process: `_objects_embedding_search` runs in an
// asgiref worker thread (assistant/rag/services/search_service.py:191), the embedding service is two gunicorn
// workers (gpu_service/gunicorn_conf.py:9).  Such a caller has no launcher and no torch.distributed: it needs the
// G GPUs of the box behind one call.  dab_sharded owns one dab_store per device (rows [g*N/G, (g+1)*N/G), ids =
// global row index), one stream and one peer-mapped staging block per device (cudaMalloc + cudaDeviceEnablePeerAccess),
// and answers a search with the same kernels the one-process-per-GPU mode uses: every shard's local top-k is written
// straight into its staging block and dab_exchange_merge_topk (publish / wait / pull over NVLink / merge) runs on every
// device; shard 0's merged lists and the OR-ed verdict come back in one device-to-host copy.  No NCCL, no collective
// library: the exchange is this library's own kernel.  Reads of peer memory are plain loads of the staging blocks.
#include <stdlib.h>
#include <string.h>

#include <condition_variable>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

#include "search.cuh"

namespace dab {
namespace {
struct Shard {
    int device = 0;
    dab_store *store = nullptr;
    cudaStream_t stream = nullptr;
    unsigned char *staging = nullptr;   // dab_exchange_bytes(R, max_q, max_k), zero-initialised, mapped into every peer
    void **bases_dev = nullptr;         // [R] staging base addresses, on this device
    float *q_dev = nullptr;             // [max_q, dim]
    unsigned char *result_dev = nullptr;  // [scores | ids | verdict] of the merged search (read back from shard 0 only)
    int64_t lo = 0, hi = 0;
};
}  // namespace
}  // namespace dab

struct dab_sharded {
    std::vector<dab::Shard> shards;
    int64_t total_rows = 0;
    int dim = 0, dtype = DAB_BF16, max_q = 0, max_k = 0;
    uint32_t epoch = 0;
    float *q_host = nullptr;              // pinned [max_q, dim]
    unsigned char *result_host = nullptr; // pinned, same layout as result_dev
    size_t off_ids = 0, off_status = 0, result_bytes = 0;
    std::mutex mu;                        // one search at a time (the staging blocks hold two epochs)
    // One host thread per shard enqueues that shard's launches: serially, G x ~13 launches would put the last
    // device ~0.5 ms behind the first at G = 8 (measured: 3.54 ms against 3.25 ms per 64-query search).
    std::vector<std::thread> workers;
    std::mutex job_mu;
    std::condition_variable job_cv, done_cv;
    uint64_t job_gen = 0;
    int job_done = 0;
    bool quit = false;
    struct Job { int Q, k, path, spare; uint32_t epoch; } job = {};
    std::vector<int> job_rc;
    std::vector<std::string> job_err;
};

using namespace dab;

static void shard_worker(dab_sharded *sh, int g);

extern "C" DAB_API void dab_sharded_destroy(dab_sharded *sh) {
    if (!sh) return;
    {
        std::lock_guard<std::mutex> g(sh->job_mu);
        sh->quit = true;
    }
    sh->job_cv.notify_all();
    for (std::thread &t : sh->workers)
        if (t.joinable()) t.join();
    for (Shard &s : sh->shards) {
        cudaSetDevice(s.device);
        if (s.stream) cudaStreamSynchronize(s.stream);
        if (s.store) dab_store_destroy(s.store);
        cudaFree(s.staging);
        cudaFree(s.bases_dev);
        cudaFree(s.q_dev);
        cudaFree(s.result_dev);
        if (s.stream) cudaStreamDestroy(s.stream);
    }
    if (sh->q_host) cudaFreeHost(sh->q_host);
    if (sh->result_host) cudaFreeHost(sh->result_host);
    delete sh;
}

extern "C" DAB_API int dab_sharded_create(dab_sharded **out, const int *devices, int n_devices, int64_t total_rows, int dim, int dtype,
                                          int max_q, int max_k) {
    if (!out) return fail(DAB_E_INVALID, "dab_sharded_create: out is NULL");
    *out = nullptr;
    if (!devices || n_devices < 1 || n_devices > 64) return fail(DAB_E_INVALID, "dab_sharded_create: 1..64 devices");
    if (total_rows < n_devices) return fail(DAB_E_INVALID, "dab_sharded_create: fewer rows than shards");
    if (max_q < 1 || max_q > DAB_MAX_QUERY_BATCH || max_k < 1 || max_k > DAB_MAX_K || (long long)n_devices * max_k > 8192)
        return fail(DAB_E_INVALID, "dab_sharded_create: need 1 <= max_q <= 256, 1 <= max_k <= 2048 and devices * max_k <= 8192");
    const int R = n_devices;
    const int64_t ex_bytes = dab_exchange_bytes(R, max_q, max_k);
    dab_sharded *sh = new (std::nothrow) dab_sharded();
    if (!sh) return fail(DAB_E_INVALID, "dab_sharded_create: out of host memory");
    sh->total_rows = total_rows;
    sh->dim = dim;
    sh->dtype = dtype;
    sh->max_q = max_q;
    sh->max_k = max_k;
    sh->shards.resize(R);
    auto up16 = [](size_t x) { return (x + 15) & ~size_t(15); };
    sh->off_ids = up16((size_t)max_q * max_k * sizeof(float));
    sh->off_status = sh->off_ids + (size_t)max_q * max_k * sizeof(int64_t);
    sh->result_bytes = sh->off_status + 16;
    auto bail = [&](int rc) {
        const std::string msg = dab_last_error();
        dab_sharded_destroy(sh);
        return fail(rc, msg);
    };
#define SH_CUDA(expr)                                                                                         \
    do {                                                                                                      \
        cudaError_t e_ = (expr);                                                                              \
        if (e_ != cudaSuccess) {                                                                              \
            fail(DAB_E_CUDA, std::string("dab_sharded_create: " #expr ": ") + cudaGetErrorString(e_));       \
            return bail(DAB_E_CUDA);                                                                          \
        }                                                                                                     \
    } while (0)
    for (int g = 0; g < R; g++) {
        Shard &s = sh->shards[g];
        s.device = devices[g];
        s.lo = (int64_t)g * total_rows / R;
        s.hi = (int64_t)(g + 1) * total_rows / R;
        SH_CUDA(cudaSetDevice(s.device));
        if (int rc = dab_store_create(&s.store, s.device, s.hi - s.lo, dim, dtype)) return bail(rc);
        if (int rc = dab_store_set_id_base(s.store, s.lo)) return bail(rc);
        SH_CUDA(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        SH_CUDA(cudaMalloc(&s.staging, (size_t)ex_bytes));
        SH_CUDA(cudaMemset(s.staging, 0, (size_t)ex_bytes));
        SH_CUDA(cudaMalloc(&s.bases_dev, (size_t)R * sizeof(void *)));
        SH_CUDA(cudaMalloc(&s.q_dev, (size_t)max_q * dim * sizeof(float)));
        SH_CUDA(cudaMalloc(&s.result_dev, sh->result_bytes));
        SH_CUDA(cudaMemset(s.result_dev, 0, sh->result_bytes));
    }
    // every device maps every other device's memory (the staging blocks are what crosses NVLink)
    for (int g = 0; g < R; g++) {
        SH_CUDA(cudaSetDevice(sh->shards[g].device));
        for (int p = 0; p < R; p++) {
            const int a = sh->shards[g].device, b = sh->shards[p].device;
            if (a == b) continue;
            int can = 0;
            SH_CUDA(cudaDeviceCanAccessPeer(&can, a, b));
            if (!can) {
                fail(DAB_E_CUDA, "dab_sharded_create: device " + std::to_string(a) + " cannot map device " + std::to_string(b) + " (no peer access)");
                return bail(DAB_E_CUDA);
            }
            const cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) {
                cudaGetLastError();
            } else if (e != cudaSuccess) {
                fail(DAB_E_CUDA, std::string("dab_sharded_create: cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e));
                return bail(DAB_E_CUDA);
            }
        }
    }
    std::vector<void *> bases(R);
    for (int g = 0; g < R; g++) bases[g] = sh->shards[g].staging;
    for (int g = 0; g < R; g++) {
        SH_CUDA(cudaSetDevice(sh->shards[g].device));
        SH_CUDA(cudaMemcpy(sh->shards[g].bases_dev, bases.data(), (size_t)R * sizeof(void *), cudaMemcpyHostToDevice));
        SH_CUDA(cudaDeviceSynchronize());
    }
    SH_CUDA(cudaHostAlloc(reinterpret_cast<void **>(&sh->q_host), (size_t)max_q * dim * sizeof(float), cudaHostAllocPortable));
    SH_CUDA(cudaHostAlloc(reinterpret_cast<void **>(&sh->result_host), sh->result_bytes, cudaHostAllocPortable));
#undef SH_CUDA
    sh->job_rc.assign(R, DAB_OK);
    sh->job_err.assign(R, std::string());
    for (int g = 0; g < R; g++) sh->workers.emplace_back(shard_worker, sh, g);
    *out = sh;
    return DAB_OK;
}

extern "C" DAB_API int dab_sharded_num_shards(const dab_sharded *sh) { return sh ? (int)sh->shards.size() : -1; }

extern "C" DAB_API dab_store *dab_sharded_shard(dab_sharded *sh, int g, int *device_out, int64_t *row_lo_out, int64_t *row_hi_out,
                                                void **stream_out) {
    if (!sh || g < 0 || g >= (int)sh->shards.size()) return nullptr;
    const Shard &s = sh->shards[g];
    if (device_out) *device_out = s.device;
    if (row_lo_out) *row_lo_out = s.lo;
    if (row_hi_out) *row_hi_out = s.hi;
    if (stream_out) *stream_out = s.stream;
    return s.store;
}

// shard g's part of one attempt: queries in, local search into the staging slot, exchange kernel, (shard 0) results out
static int shard_attempt(dab_sharded *sh, int g, const dab_sharded::Job &j) {
    const int R = (int)sh->shards.size();
    Shard &s = sh->shards[g];
    DAB_CUDA_OK(cudaSetDevice(s.device));
    DAB_CUDA_OK(cudaMemcpyAsync(s.q_dev, sh->q_host, (size_t)j.Q * sh->dim * sizeof(float), cudaMemcpyHostToDevice, s.stream));
    float *slot_s = nullptr;
    int64_t *slot_i = nullptr;
    if (int rc = dab_exchange_slot(s.staging, R, sh->max_q, sh->max_k, j.epoch, &slot_s, &slot_i)) return rc;
    if (int rc = dab_store_search_ex(s.store, s.q_dev, j.Q, j.k, nullptr, slot_s, slot_i, j.path, j.spare, s.stream)) return rc;
    const int *verdict = dab_store_status_dev(s.store);   // the slot of the search this thread just enqueued
    float *out_s = reinterpret_cast<float *>(s.result_dev);
    int64_t *out_i = reinterpret_cast<int64_t *>(s.result_dev + sh->off_ids);
    int *out_st = reinterpret_cast<int *>(s.result_dev + sh->off_status);
    if (int rc = dab_exchange_merge_topk(s.bases_dev, g, R, sh->max_q, sh->max_k, j.Q, j.k, j.epoch, verdict, out_s, out_i, out_st, s.stream))
        return rc;
    if (g == 0) DAB_CUDA_OK(cudaMemcpyAsync(sh->result_host, s.result_dev, sh->result_bytes, cudaMemcpyDeviceToHost, s.stream));
    // every device must be past its exchange kernel before the next search reuses the slots two epochs on
    DAB_CUDA_OK(cudaStreamSynchronize(s.stream));
    return DAB_OK;
}

static void shard_worker(dab_sharded *sh, int g) {
    uint64_t seen = 0;
    for (;;) {
        dab_sharded::Job j;
        {
            std::unique_lock<std::mutex> lk(sh->job_mu);
            sh->job_cv.wait(lk, [&] { return sh->quit || sh->job_gen != seen; });
            if (sh->quit) return;
            seen = sh->job_gen;
            j = sh->job;
        }
        const int rc = shard_attempt(sh, g, j);
        std::string err = rc ? dab_last_error() : "";
        {
            std::lock_guard<std::mutex> lk(sh->job_mu);
            sh->job_rc[g] = rc;
            sh->job_err[g] = err;
            sh->job_done++;
        }
        sh->done_cv.notify_one();
    }
}

// one attempt: every shard's thread enqueues its own device's work; shard 0 brings the answer back
static int sharded_search_once(dab_sharded *sh, int Q, int k, int path, int spare, bool *incomplete, bool *inexact) {
    const int R = (int)sh->shards.size();
    {
        std::lock_guard<std::mutex> lk(sh->job_mu);
        sh->job = dab_sharded::Job{Q, k, path, spare, ++sh->epoch};
        sh->job_done = 0;
        sh->job_gen++;
    }
    sh->job_cv.notify_all();
    {
        std::unique_lock<std::mutex> lk(sh->job_mu);
        sh->done_cv.wait(lk, [&] { return sh->job_done == R; });
    }
    for (int g = 0; g < R; g++)
        if (sh->job_rc[g] != DAB_OK) return fail(sh->job_rc[g], sh->job_err[g]);
    const int *st = reinterpret_cast<const int *>(sh->result_host + sh->off_status);
    *incomplete = st[0] != 0;
    *inexact = st[1] != 0;
    return DAB_OK;
}

extern "C" DAB_API int dab_sharded_search_host(dab_sharded *sh, const float *q_host, int Q, int k, float *out_scores_host,
                                               int64_t *out_ids_host) {
    if (!sh) return fail(DAB_E_INVALID, "dab_sharded_search_host: handle is NULL");
    if (Q < 0 || k < 1) return fail(DAB_E_INVALID, "dab_sharded_search_host: bad Q or k");
    if (Q > sh->max_q || k > sh->max_k) return fail(DAB_E_CAPACITY, "dab_sharded_search_host: Q or k beyond what dab_sharded_create was given");
    if (Q == 0) return DAB_OK;
    if (!q_host || !out_scores_host || !out_ids_host) return fail(DAB_E_INVALID, "dab_sharded_search_host: NULL pointer");
    std::lock_guard<std::mutex> g(sh->mu);
    int prev_dev = 0;
    cudaGetDevice(&prev_dev);
    memcpy(sh->q_host, q_host, (size_t)Q * sh->dim * sizeof(float));
    int path = -1, spare = -1, rc = DAB_OK;
    for (;;) {
        bool incomplete = false, inexact = false;
        rc = sharded_search_once(sh, Q, k, path, spare, &incomplete, &inexact);
        if (rc != DAB_OK) break;
        if (inexact && path != 1) {   // some shard's tensor pass lost exactness for some query: the exact CUDA-core path answers
            path = 1;
            continue;
        }
        if (incomplete) {
            if (spare >= 64) {
                rc = fail(DAB_E_INCOMPLETE, "dab_sharded_search_host: search did not complete within 64 spare rounds");
                break;
            }
            spare = spare <= 0 ? 4 : spare * 2;
            continue;
        }
        memcpy(out_scores_host, sh->result_host, (size_t)Q * k * sizeof(float));
        memcpy(out_ids_host, sh->result_host + sh->off_ids, (size_t)Q * k * sizeof(int64_t));
        break;
    }
    cudaSetDevice(prev_dev);
    return rc;
}
