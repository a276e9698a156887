/*
 * oracle/scan_ref.c - CPU restatement of the exact cosine scan + top-k and of the
 * masked mean-pool.  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): the checker
 * in tests/ and the timed CPU baseline in bench.py.  Never linked into the product.
 *
 * What it follows (paths relative to /root/reference):
 *   - assistant/rag/services/search_service.py:185-196  `_objects_embedding_search`:
 *     ORDER BY (field <=> q) ASC LIMIT n.  The arithmetic of `<=>` is pgvector's
 *     cosine distance (third-party C, NOT in the reference tree; server version
 *     unpinned by assistant/storage/migrations/0001_initial.py:14-16, client pinned
 *     pgvector==0.2.5 in requirements.txt:4).  PARITY UNPINNED for this half: no
 *     PostgreSQL here, no golden vector in the reference.  Published algorithm,
 *     restated: three fp32 running sums over i (a.b, a.a, b.b); similarity =
 *     (double)dot / sqrt((double)na * (double)nb); clamp to [-1,1]; distance =
 *     1 - similarity; zero vector => NaN.  PostgreSQL orders NaN after every number.
 *   - assistant/storage/models.py:13,23: rows are fp32[dim] (dim 768), nullable; the
 *     build also stores bf16 rows, upcast exactly here.
 *   - assistant/ai/embedders/transformers.py:25: mean over the token axis (fp32).
 *
 * Compiled twice by oracle/Makefile into one library:
 *   SUFFIX=_strict : -O2, no reassociation -> the sums run in index order (i = 0..dim-1)
 *   SUFFIX=_vec    : -O3 -mavx2 -mfma -fassociative-math ... -> what a pgvector build
 *                    does ("auto-vectorized"), the timed CPU baseline.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifndef SUFFIX
#define SUFFIX _strict
#endif
#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)
#define FN(name) CAT(name, SUFFIX)

static inline float bf16_up(uint16_t b) {
    uint32_t u = ((uint32_t)b) << 16;
    float f;
    memcpy(&f, &u, 4);
    return f;
}

/* one pgvector-style cosine distance, fp32 rows */
static inline double dist_f32(const float *a, const float *b, int dim) {
    float dot = 0.0f, na = 0.0f, nb = 0.0f;
    for (int i = 0; i < dim; i++) {
        dot += a[i] * b[i];
        na += a[i] * a[i];
        nb += b[i] * b[i];
    }
    double sim = (double)dot / sqrt((double)na * (double)nb);
    if (isnan(sim)) return NAN;
    if (sim > 1.0) sim = 1.0;
    else if (sim < -1.0) sim = -1.0;
    return 1.0 - sim;
}

static inline double dist_bf16(const uint16_t *a, const float *b, int dim) {
    float dot = 0.0f, na = 0.0f, nb = 0.0f;
    for (int i = 0; i < dim; i++) {
        float x = bf16_up(a[i]);
        dot += x * b[i];
        na += x * x;
        nb += b[i] * b[i];
    }
    double sim = (double)dot / sqrt((double)na * (double)nb);
    if (isnan(sim)) return NAN;
    if (sim > 1.0) sim = 1.0;
    else if (sim < -1.0) sim = -1.0;
    return 1.0 - sim;
}

/* distance of every row to one query; dtype 0 = fp32 rows, 1 = bf16 rows */
void FN(dab_oracle_distances)(const void *rows, int dtype, int64_t n, int dim, const float *q,
                             double *dist_out, int nthreads) {
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < n; r++) {
        dist_out[r] = dtype == 0 ? dist_f32((const float *)rows + r * dim, q, dim)
                                 : dist_bf16((const uint16_t *)rows + r * dim, q, dim);
    }
}

/* ---- bounded selection: keep the k best under (distance asc, NaN last, index asc) ---- */
typedef struct {
    double d;
    int64_t i;
} hit_t;

/* a "worse" than b ? (worse = later in the ORDER BY) */
static inline int worse(hit_t a, hit_t b) {
    int an = isnan(a.d), bn = isnan(b.d);
    if (an != bn) return an;
    if (!an && a.d != b.d) return a.d > b.d;
    return a.i > b.i;
}

static void heap_sift_down(hit_t *h, int n, int p) { /* max-heap on "worse" */
    for (;;) {
        int l = 2 * p + 1, r = l + 1, m = p;
        if (l < n && worse(h[l], h[m])) m = l;
        if (r < n && worse(h[r], h[m])) m = r;
        if (m == p) return;
        hit_t t = h[p];
        h[p] = h[m];
        h[m] = t;
        p = m;
    }
}

static void heap_offer(hit_t *h, int *n, int k, hit_t x) {
    if (*n < k) {
        int p = (*n)++;
        h[p] = x;
        while (p > 0) {
            int up = (p - 1) / 2;
            if (!worse(h[p], h[up])) break;
            hit_t t = h[p];
            h[p] = h[up];
            h[up] = t;
            p = up;
        }
    } else if (worse(h[0], x)) {
        h[0] = x;
        heap_sift_down(h, k, 0);
    }
}

static int cmp_hit(const void *pa, const void *pb) {
    hit_t a = *(const hit_t *)pa, b = *(const hit_t *)pb;
    if (worse(a, b)) return 1;
    if (worse(b, a)) return -1;
    return 0;
}

/*
 * Exact scan + top-k for nq queries (search_service.py:185-196 semantics per query).
 * allow: optional byte mask [n] (the queryset filter), NULL = all rows.
 * NaN-distance rows (zero vectors) are kept out of the result, as the device store does.
 * Writes ids_out/dist_out [nq, k] (padded with -1 / NaN) and returns rows returned for
 * the LAST query (callers use counts_out for per-query counts when not NULL).
 */
int64_t FN(dab_oracle_search)(const void *rows, int dtype, int64_t n, int dim, const float *queries,
                             int nq, int k, const uint8_t *allow, int nthreads, int64_t *ids_out,
                             double *dist_out, int64_t *counts_out) {
    int nt = 1;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
    nt = nthreads > 0 ? nthreads : omp_get_max_threads();
#endif
    if (k <= 0 || nq <= 0) return 0;
    hit_t *heaps = (hit_t *)malloc(sizeof(hit_t) * (size_t)nt * (size_t)k);
    int *sizes = (int *)malloc(sizeof(int) * (size_t)nt);
    int64_t last = 0;
    for (int qi = 0; qi < nq; qi++) {
        const float *q = queries + (int64_t)qi * dim;
        for (int t = 0; t < nt; t++) sizes[t] = 0;
#pragma omp parallel num_threads(nt)
        {
            int t = 0;
#ifdef _OPENMP
            t = omp_get_thread_num();
#endif
            hit_t *h = heaps + (size_t)t * k;
            int sz = 0;
#pragma omp for schedule(static)
            for (int64_t r = 0; r < n; r++) {
                if (allow && !allow[r]) continue;
                hit_t x;
                x.d = dtype == 0 ? dist_f32((const float *)rows + r * dim, q, dim)
                                 : dist_bf16((const uint16_t *)rows + r * dim, q, dim);
                x.i = r;
                if (isnan(x.d)) continue;
                heap_offer(h, &sz, k, x);
            }
            sizes[t] = sz;
        }
        /* merge the per-thread heaps */
        int total = 0;
        for (int t = 0; t < nt; t++) total += sizes[t];
        hit_t *all = (hit_t *)malloc(sizeof(hit_t) * (size_t)(total > 0 ? total : 1));
        int p = 0;
        for (int t = 0; t < nt; t++)
            for (int j = 0; j < sizes[t]; j++) all[p++] = heaps[(size_t)t * k + j];
        qsort(all, (size_t)total, sizeof(hit_t), cmp_hit);
        int got = total < k ? total : k;
        for (int j = 0; j < k; j++) {
            ids_out[(int64_t)qi * k + j] = j < got ? all[j].i : -1;
            dist_out[(int64_t)qi * k + j] = j < got ? all[j].d : NAN;
        }
        if (counts_out) counts_out[qi] = got;
        last = got;
        free(all);
    }
    free(heaps);
    free(sizes);
    return last;
}

/*
 * Masked mean-pool + L2 normalise: e[b] = sum_t m[b,t] h[b,t,:] / max(sum m, 1); e /= max(|e|, 1e-12).
 * fp32 accumulation like the reference's torch mean (transformers.py:25).  hidden fp32 [B,T,D],
 * mask int64 [B,T] (the HF tokenizer's attention_mask dtype).
 */
void FN(dab_oracle_pool)(const float *hidden, const int64_t *mask, int64_t B, int64_t T, int64_t D,
                        float *out, int nthreads) {
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < B; b++) {
        float *o = out + b * D;
        for (int64_t d = 0; d < D; d++) o[d] = 0.0f;
        int64_t cnt = 0;
        for (int64_t t = 0; t < T; t++) {
            if (!mask[b * T + t]) continue;
            const float *h = hidden + (b * T + t) * D;
            for (int64_t d = 0; d < D; d++) o[d] += h[d];
            cnt++;
        }
        float inv = 1.0f / (float)(cnt > 0 ? cnt : 1);
        float ss = 0.0f;
        for (int64_t d = 0; d < D; d++) {
            o[d] *= inv;
            ss += o[d] * o[d];
        }
        float nrm = sqrtf(ss);
        float s = 1.0f / (nrm > 1e-12f ? nrm : 1e-12f);
        for (int64_t d = 0; d < D; d++) o[d] *= s;
    }
}
