// docagg.cu - document-level aggregation of search hits on the device.
//
// Stands behind assistant/rag/services/search_service.py:133-150 of the reference: the top-n chunk
// hits (n = max_scores_n * top_n * 10, :129-131) are grouped by document in rank order, documents
// with fewer than max_scores_n hits are dropped, a document scores
// 1 - sum(first max_scores_n distances) / max_scores_n, and the best top_n documents are returned.
// Doing this next to the search output means the host fetches <= top_n documents instead of
// materialising n = 250 chunk rows first.
//
// One CTA per query.  Hits arrive sorted by (score desc, row asc); a stable grouping is obtained by
// sorting keys (document id, rank).  Sums are taken in double, like the reference's Python floats.
// Ties between document scores are broken by the rank of the document's best hit (first appearance).
#include "search.cuh"

namespace dab {

namespace {

constexpr int kThreads = 256;
constexpr int kMaxHits = 4096;

struct DocItem {
    long long doc;
    int rank;  // position in the hit list; < 0 = empty
};
__device__ __forceinline__ bool doc_less(const DocItem &a, const DocItem &b) {  // empty items sort last
    const bool ae = a.rank < 0, be = b.rank < 0;
    if (ae != be) return be;
    if (a.doc != b.doc) return a.doc < b.doc;
    return a.rank < b.rank;
}

__global__ void __launch_bounds__(kThreads)
doc_aggregate_kernel(const float *__restrict__ scores, const int64_t *__restrict__ rows, const int64_t *__restrict__ doc_of_row,
                     int64_t n_rows, int n, int max_scores_n, int top_n, int64_t *__restrict__ out_docs,
                     float *__restrict__ out_scores) {
    extern __shared__ unsigned char smem_raw[];
    int P = 32;
    while (P < n) P <<= 1;
    long long *s_doc = reinterpret_cast<long long *>(smem_raw);                 // [P]
    double *s_val = reinterpret_cast<double *>(s_doc + P);                      // [P] distance by rank, later doc score
    int *s_rank = reinterpret_cast<int *>(s_val + P);                           // [P]
    const int qi = blockIdx.x, tid = threadIdx.x;
    const float *sc = scores + (size_t)qi * n;
    const int64_t *rw = rows + (size_t)qi * n;

    for (int i = tid; i < P; i += kThreads) {
        long long d = 0;
        int r = -1;
        double dist = 0.0;
        if (i < n) {
            const int64_t row = rw[i];
            if (row >= 0 && row < n_rows) {
                d = doc_of_row[row];
                r = i;
                dist = 1.0 - (double)sc[i];
            }
        }
        s_doc[i] = d;
        s_rank[i] = r;
        s_val[i] = dist;   // indexed by rank i
    }
    // sort (doc, rank) ascending; distances stay addressable by rank
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (int i = tid; i < (P >> 1); i += kThreads) {
                const int lo = 2 * i - (i & (stride - 1));
                const int hi = lo + stride;
                const bool asc = (lo & size) == 0;
                const DocItem a{s_doc[lo], s_rank[lo]}, b{s_doc[hi], s_rank[hi]};
                if (doc_less(b, a) == asc && (a.doc != b.doc || a.rank != b.rank)) {
                    s_doc[lo] = b.doc; s_rank[lo] = b.rank;
                    s_doc[hi] = a.doc; s_rank[hi] = a.rank;
                }
            }
        }
    }
    __syncthreads();
    // segment heads: a document with >= max_scores_n hits gets its score, everything else is dropped.
    // Result key per head: (score desc, best rank asc) packed for a second sort.
    double *s_score = reinterpret_cast<double *>(s_rank + P);                   // [P]
    int *s_first = reinterpret_cast<int *>(s_score + P);                        // [P]
    for (int i = tid; i < P; i += kThreads) {
        double out = -1e300;
        int first = -1;
        if (s_rank[i] >= 0 && (i == 0 || s_rank[i - 1] < 0 || s_doc[i - 1] != s_doc[i])) {
            int cnt = 0;
            double sum = 0.0;
            for (int j = i; j < P && s_rank[j] >= 0 && s_doc[j] == s_doc[i]; j++) {
                if (cnt < max_scores_n) sum += s_val[s_rank[j]];
                cnt++;
            }
            if (cnt >= max_scores_n) {
                out = 1.0 - sum / (double)max_scores_n;
                first = s_rank[i];
            }
        }
        s_score[i] = out;
        s_first[i] = first;
    }
    // second sort: (score desc, first rank asc), dropped entries last
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (int i = tid; i < (P >> 1); i += kThreads) {
                const int lo = 2 * i - (i & (stride - 1));
                const int hi = lo + stride;
                const bool desc = (lo & size) == 0;
                const double sa = s_score[lo], sb = s_score[hi];
                const int fa = s_first[lo], fb = s_first[hi];
                const bool ae = fa < 0, be = fb < 0;
                // is b better than a?
                const bool b_better = ae != be ? ae : (sa != sb ? sb > sa : fb < fa);
                if (b_better == desc && (fa != fb || sa != sb)) {
                    const long long da = s_doc[lo];
                    s_doc[lo] = s_doc[hi]; s_doc[hi] = da;
                    s_score[lo] = sb; s_score[hi] = sa;
                    s_first[lo] = fb; s_first[hi] = fa;
                }
            }
        }
    }
    __syncthreads();
    for (int j = tid; j < top_n; j += kThreads) {
        const bool ok = j < P && s_first[j] >= 0;
        out_docs[(size_t)qi * top_n + j] = ok ? s_doc[j] : -1;
        out_scores[(size_t)qi * top_n + j] = ok ? (float)s_score[j] : -INFINITY;
    }
}

}  // namespace
}  // namespace dab

using namespace dab;

extern "C" DAB_API int dab_doc_aggregate(const float *scores_dev, const int64_t *rows_dev, const int64_t *doc_of_row_dev,
                                         int64_t n_rows, int Q, int n, int max_scores_n, int top_n, int64_t *out_doc_ids_dev,
                                         float *out_doc_scores_dev, void *stream) {
    if (!scores_dev || !rows_dev || !doc_of_row_dev || !out_doc_ids_dev || !out_doc_scores_dev)
        return fail(DAB_E_INVALID, "dab_doc_aggregate: NULL pointer");
    if (Q < 0 || n < 1 || max_scores_n < 1 || top_n < 1 || n_rows < 0) return fail(DAB_E_INVALID, "dab_doc_aggregate: bad size");
    if (n > kMaxHits) return fail(DAB_E_CAPACITY, "dab_doc_aggregate: more than 4096 hits per query");
    if (Q == 0) return DAB_OK;
    int P = 32;
    while (P < n) P <<= 1;
    const size_t smem = (size_t)P * (sizeof(long long) + sizeof(double) + sizeof(int) + sizeof(double) + sizeof(int));
    DAB_CUDA_OK(cudaFuncSetAttribute(doc_aggregate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)((size_t)kMaxHits * 32)));
    doc_aggregate_kernel<<<Q, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(scores_dev, rows_dev, doc_of_row_dev, n_rows, n,
                                                                                  max_scores_n, top_n, out_doc_ids_dev,
                                                                                  out_doc_scores_dev);
    DAB_LAUNCH_OK();
    return DAB_OK;
}
