// exchange.cu - the one exchange step of the sharded search as ONE kernel over NVLink peer memory.
//
// After the local scans every GPU holds its exact top-k per query; the global answer is the top-k of
// the union (what `ORDER BY embedding <=> q LIMIT n` gives over the single PostgreSQL table,
// assistant/rag/services/search_service.py:192-194).  Instead of two NCCL all-gathers followed by
// a merge kernel, each rank's search writes its lists straight into a peer-mapped staging block
// (e.g. torch symmetric memory), and this kernel
//   1. publishes "rank r finished search number `epoch`" with one release-store per peer,
//   2. waits (acquire loads on its OWN flag words) until every rank has published,
//   3. pulls the R lists of each query straight out of the peers' HBM with NVLink loads and merges
//      them under (score desc, id asc) - one CTA per query,
//   4. ORs the ranks' search verdicts, so every rank takes the same retry decision.
// All ranks must launch it once per search (SPMD).  Spins are bounded: a lost peer traps instead of
// hanging the GPU.
//
// Staging block layout (identical on every rank; sizes from dab_exchange_bytes):
//   flags   [2][R]            uint32   flags[p][r] = last epoch of parity p that rank r published HERE
//   status  [2][2]            int32    (incomplete, inexact) of this rank's search of parity p
//   scores  [2][max_Q*max_k]  float
//   ids     [2][max_Q*max_k]  int64
// Two parities: a rank can run at most one search ahead of the slowest peer, so the lists of search
// e are never overwritten (by search e+2) before every peer has merged them.
#include "search.cuh"

namespace dab {

namespace {

constexpr size_t up256(size_t x) { return (x + 255) & ~size_t(255); }

struct ExLayout {
    size_t flags, status, scores, ids, total;
    size_t slot_elems;
};
__host__ __device__ inline ExLayout ex_layout(int R, int max_q, int max_k) {
    ExLayout l;
    l.slot_elems = (size_t)max_q * max_k;
    l.flags = 0;
    l.status = up256(2 * (size_t)R * sizeof(uint32_t));
    l.scores = l.status + 256;
    l.ids = l.scores + up256(2 * l.slot_elems * sizeof(float));
    l.total = l.ids + up256(2 * l.slot_elems * sizeof(int64_t));
    return l;
}

__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

struct MergeItem {
    uint32_t s;
    int64_t id;
};
__device__ __forceinline__ bool better(const MergeItem &a, const MergeItem &b) {
    const bool ae = a.id < 0, be = b.id < 0;
    if (ae != be) return be;
    if (a.s != b.s) return a.s > b.s;
    return a.id < b.id;
}

constexpr int kThreads = 256;
constexpr int kMaxItems = 8192;

__global__ void __launch_bounds__(kThreads)
exchange_merge_kernel(unsigned char *const *__restrict__ peer_bases, int rank, int R, int max_q, int max_k, int Q, int k,
                      uint32_t epoch, const int *__restrict__ local_status, float *__restrict__ out_scores,
                      int64_t *__restrict__ out_ids, int *__restrict__ out_status) {
    extern __shared__ unsigned char smem_raw[];
    const ExLayout L = ex_layout(R, max_q, max_k);
    const int parity = (int)(epoch & 1u);
    const int qi = blockIdx.x, tid = threadIdx.x;
    unsigned char *mine = peer_bases[rank];

    if (qi == 0) {
        // 1. publish: my verdict first, then one flag per peer (release = my lists and verdict are visible)
        if (tid == 0) {
            int *st = reinterpret_cast<int *>(mine + L.status) + 2 * parity;
            st[0] = local_status ? local_status[0] : 0;
            st[1] = local_status ? local_status[1] : 0;
            __threadfence_system();
        }
        __syncthreads();
        if (tid < R) {
            uint32_t *flag = reinterpret_cast<uint32_t *>(peer_bases[tid] + L.flags) + (size_t)parity * R + rank;
            st_release_sys(flag, epoch);
        }
    }
    // 2. wait until every rank has published this epoch in MY block
    if (tid < R) {
        const uint32_t *flag = reinterpret_cast<const uint32_t *>(mine + L.flags) + (size_t)parity * R + tid;
        for (uint32_t spins = 0; ld_acquire_sys(flag) != epoch; spins++) {
            if (spins > (1u << 26)) __trap();
            __nanosleep(64);
        }
    }
    __syncthreads();

    // 3. gather this query's R lists over NVLink and merge
    const int n = R * k;
    int P = 32;
    while (P < n) P <<= 1;
    uint32_t *ss = reinterpret_cast<uint32_t *>(smem_raw);
    int64_t *si = reinterpret_cast<int64_t *>(smem_raw + (size_t)P * sizeof(uint32_t));
    for (int i = tid; i < P; i += kThreads) {
        if (i < n) {
            const int r = i / k, j = i % k;
            const unsigned char *base = peer_bases[r];
            const size_t off = (size_t)parity * L.slot_elems + (size_t)qi * k + j;
            const float sc = reinterpret_cast<const float *>(base + L.scores)[off];
            const int64_t id = reinterpret_cast<const int64_t *>(base + L.ids)[off];
            si[i] = (id >= 0 && sc == sc) ? id : -1;
            ss[i] = f32_orderable(sc);
        } else {
            si[i] = -1;
            ss[i] = 0;
        }
    }
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (int i = tid; i < (P >> 1); i += kThreads) {
                const int lo = 2 * i - (i & (stride - 1));
                const int hi = lo + stride;
                const bool desc = (lo & size) == 0;
                const MergeItem a{ss[lo], si[lo]}, b{ss[hi], si[hi]};
                if (better(b, a) == desc && (a.s != b.s || a.id != b.id)) {
                    ss[lo] = b.s; si[lo] = b.id;
                    ss[hi] = a.s; si[hi] = a.id;
                }
            }
        }
    }
    __syncthreads();
    for (int j = tid; j < k; j += kThreads) {
        const bool ok = j < n && si[j] >= 0;
        out_scores[(size_t)qi * k + j] = ok ? orderable_f32(ss[j]) : -INFINITY;
        out_ids[(size_t)qi * k + j] = ok ? si[j] : -1;
    }
    // 4. the same verdict on every rank
    if (qi == 0 && tid == 0 && out_status) {
        int inc = 0, inex = 0;
        for (int r = 0; r < R; r++) {
            const int *st = reinterpret_cast<const int *>(peer_bases[r] + L.status) + 2 * parity;
            inc |= st[0];
            inex |= st[1];
        }
        out_status[0] = inc;
        out_status[1] = inex;
    }
}

}  // namespace
}  // namespace dab

using namespace dab;

extern "C" DAB_API int64_t dab_exchange_bytes(int R, int max_q, int max_k) {
    if (R < 1 || max_q < 1 || max_k < 1) return -1;
    return (int64_t)ex_layout(R, max_q, max_k).total;
}

extern "C" DAB_API int dab_exchange_slot(void *base, int R, int max_q, int max_k, uint32_t epoch, float **scores, int64_t **ids) {
    if (!base || R < 1 || max_q < 1 || max_k < 1 || !scores || !ids) return fail(DAB_E_INVALID, "dab_exchange_slot: bad argument");
    const ExLayout L = ex_layout(R, max_q, max_k);
    const size_t p = epoch & 1u;
    *scores = reinterpret_cast<float *>(static_cast<unsigned char *>(base) + L.scores) + p * L.slot_elems;
    *ids = reinterpret_cast<int64_t *>(static_cast<unsigned char *>(base) + L.ids) + p * L.slot_elems;
    return DAB_OK;
}

extern "C" DAB_API int dab_exchange_merge_topk(void *const *peer_bases_dev, int rank, int R, int max_q, int max_k, int Q, int k,
                                               uint32_t epoch, const int *local_status_dev_or_null, float *out_scores_dev,
                                               int64_t *out_ids_dev, int *out_status_dev_or_null, void *stream) {
    if (!peer_bases_dev || !out_scores_dev || !out_ids_dev) return fail(DAB_E_INVALID, "dab_exchange_merge_topk: NULL pointer");
    if (R < 1 || R > 64 || rank < 0 || rank >= R) return fail(DAB_E_INVALID, "dab_exchange_merge_topk: bad rank / world size");
    if (Q < 0 || Q > max_q || k < 1 || k > max_k) return fail(DAB_E_CAPACITY, "dab_exchange_merge_topk: Q or k beyond the staging block's capacity");
    if ((long long)R * k > kMaxItems) return fail(DAB_E_CAPACITY, "dab_exchange_merge_topk: R*k exceeds 8192");
    if (epoch == 0) return fail(DAB_E_INVALID, "dab_exchange_merge_topk: epochs start at 1 (the block is zero-initialised)");
    if (Q == 0) return DAB_OK;
    int P = 32;
    while (P < R * k) P <<= 1;
    const int smem = P * (int)(sizeof(uint32_t) + sizeof(int64_t));
    static SmemAttrOnce attr_once;
    DAB_CUDA_OK(attr_once.ensure(exchange_merge_kernel, kMaxItems * (int)(sizeof(uint32_t) + sizeof(int64_t))));
    exchange_merge_kernel<<<Q, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<unsigned char *const *>(peer_bases_dev), rank, R, max_q, max_k, Q, k, epoch, local_status_dev_or_null,
        out_scores_dev, out_ids_dev, out_status_dev_or_null);
    DAB_LAUNCH_OK();
    return DAB_OK;
}
