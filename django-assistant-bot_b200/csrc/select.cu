// select.cu - the selection side of the search: query preparation, per-round candidate
// compaction (sort + dedupe + threshold update + on-device round control), the final
// write-out (with exact fp32 rescoring for the tensor path) and the cross-shard k-way merge.
//
// Together with a scan kernel these implement what PostgreSQL does for
// `ORDER BY embedding <=> q LIMIT n` (assistant/rag/services/search_service.py:192-194).
#include "search.cuh"

namespace dab {

size_t search_scratch_bytes(int dim) {
    auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
    size_t b = 0;
    b += up((size_t)DAB_MAX_QUERY_BATCH * dim * sizeof(float));               // qhat
    b += up((size_t)DAB_MAX_QUERY_BATCH * dim * sizeof(__nv_bfloat16));       // qbf16
    b += up((size_t)DAB_MAX_QUERY_BATCH * kCandCap * sizeof(uint64_t));       // cand
    b += up(DAB_MAX_QUERY_BATCH * sizeof(uint32_t));                          // cand_count
    b += up(DAB_MAX_QUERY_BATCH * sizeof(uint64_t));                          // tau_key
    b += up(DAB_MAX_QUERY_BATCH * sizeof(int));                               // active
    b += up(DAB_MAX_QUERY_BATCH * sizeof(int));                               // overflowed
    b += up(DAB_MAX_QUERY_BATCH * sizeof(int));                               // inexact_q
    b += up(DAB_MAX_QUERY_BATCH * sizeof(unsigned int));                      // finish_ticket
    b += up(sizeof(SearchCtrl));
    return b;
}

SearchScratch carve_scratch(void *base, int dim) {
    auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
    char *p = static_cast<char *>(base);
    SearchScratch s;
    s.qhat = reinterpret_cast<float *>(p);
    p += up((size_t)DAB_MAX_QUERY_BATCH * dim * sizeof(float));
    s.qbf16 = reinterpret_cast<__nv_bfloat16 *>(p);
    p += up((size_t)DAB_MAX_QUERY_BATCH * dim * sizeof(__nv_bfloat16));
    s.cand = reinterpret_cast<uint64_t *>(p);
    p += up((size_t)DAB_MAX_QUERY_BATCH * kCandCap * sizeof(uint64_t));
    s.cand_count = reinterpret_cast<uint32_t *>(p);
    p += up(DAB_MAX_QUERY_BATCH * sizeof(uint32_t));
    s.tau_key = reinterpret_cast<uint64_t *>(p);
    p += up(DAB_MAX_QUERY_BATCH * sizeof(uint64_t));
    s.active = reinterpret_cast<int *>(p);
    p += up(DAB_MAX_QUERY_BATCH * sizeof(int));
    s.overflowed = reinterpret_cast<int *>(p);
    p += up(DAB_MAX_QUERY_BATCH * sizeof(int));
    s.inexact_q = reinterpret_cast<int *>(p);
    p += up(DAB_MAX_QUERY_BATCH * sizeof(int));
    s.finish_ticket = reinterpret_cast<unsigned int *>(p);
    p += up(DAB_MAX_QUERY_BATCH * sizeof(unsigned int));
    s.ctrl = reinterpret_cast<SearchCtrl *>(p);
    s.verdict_dev = nullptr;
    s.verdict_host = nullptr;
    return s;
}

// Levels: level l scans the tiles t with t % stride[l] == 0 not scanned before, stride[l] =
// growth^(L-1-l).  Level 0 is a strided sample small enough that every row may pass
// (<= kCandCap/2 rows); each later level is (growth-1) times everything before it, so with a
// threshold equal to the k-th best seen so far about k*(growth-1) <= kCandCap/2 rows pass.
// Growth is capped at 32: measured on B200, fewer but fatter levels (growth 163-245 for k = 10,
// 13 instead of 8 for the tensor path) were SLOWER - every admitted row costs an atomic append in
// the scan and the compaction sort grows with the admitted count - than one more kernel boundary.
LevelPlan make_level_plan(int64_t n_rows, int tile_rows, int k_keep) {
    LevelPlan p = {};
    p.tile_rows = tile_rows;
    p.num_tiles = (n_rows + tile_rows - 1) / tile_rows;
    int g = 32;
    while (g > 2 && (long long)g * k_keep > kCandCap / 2) g >>= 1;
    p.growth = g;
    int L = 1;
    long long s0 = 1;
    while (L < kMaxLevels && ((p.num_tiles + s0 - 1) / s0) * tile_rows > kCandCap / 2) {
        s0 *= g;
        L++;
    }
    p.num_levels = L;
    long long s = s0;
    for (int l = 0; l < L; l++) {
        p.stride[l] = (int)s;
        const long long m = (p.num_tiles + s - 1) / s;
        p.tiles_in_level[l] = l == 0 ? m : m - (m + g - 1) / g;
        s /= g;
    }
    return p;
}

namespace {

// ---------------------------------------------------------------- begin: normalise queries, reset state
__global__ void __launch_bounds__(128)
search_begin_kernel(SearchScratch sc, const float *__restrict__ q, int dim, int want_split, int first_chunk) {
    pdl_enter();
    const int qi = blockIdx.x;
    const int tid = threadIdx.x;
    __shared__ float red[4];
    const float *src = q + (size_t)qi * dim;
    float ss = 0.f;
    for (int d = tid; d < dim; d += 128) {
        const float v = src[d];
        ss += v * v;
    }
    ss = warp_sum(ss);
    if ((tid & 31) == 0) red[tid >> 5] = ss;
    __syncthreads();
    const float tot = red[0] + red[1] + red[2] + red[3];
    const float inv = tot > 0.f ? 1.0f / sqrtf(tot) : 0.f;
    for (int d = tid; d < dim; d += 128) {
        const float v = src[d] * inv;
        sc.qhat[(size_t)qi * dim + d] = v;
        // the tensor pass scores with the query rounded to bf16 (the rescoring uses qhat itself)
        if (want_split) sc.qbf16[(size_t)qi * dim + d] = __float2bfloat16_rn(v);
    }
    if (tid == 0) {
        sc.cand_count[qi] = 0;
        sc.tau_key[qi] = kTauMin;
        sc.active[qi] = 1;
        sc.overflowed[qi] = 0;
        if (sc.inexact_q) sc.inexact_q[qi] = 0;
        if (sc.finish_ticket) sc.finish_ticket[qi] = 0;
        if (qi == 0) {
            SearchCtrl c = {};
            if (!first_chunk) {  // a later chunk of the same search call keeps the sticky verdicts
                c.sticky_incomplete = sc.ctrl->sticky_incomplete;
                c.sticky_inexact = sc.ctrl->sticky_inexact;
                c.overflow_events = sc.ctrl->overflow_events;
            }
            *sc.ctrl = c;
        }
    }
}

// ---------------------------------------------------------------- bitonic sort of 64-bit keys, descending
template <int THREADS>
__device__ void bitonic_sort_desc(uint64_t *keys, int P) {
    for (int size = 2; size <= P; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (int i = threadIdx.x; i < (P >> 1); i += THREADS) {
                const int lo = 2 * i - (i & (stride - 1));
                const int hi = lo + stride;
                const bool desc = (lo & size) == 0;
                const uint64_t a = keys[lo], b = keys[hi];
                if ((a < b) == desc) {
                    keys[lo] = b;
                    keys[hi] = a;
                }
            }
        }
    }
    __syncthreads();
}

constexpr int kCompactThreads = 1024;
constexpr int kSelBins = 2 * kCompactThreads;  // two histogram bins per thread

// Selection instead of sorting (tensor path, ordinary rounds).  `keys` is the shared-memory copy
// buffer, `mine` the query's candidate list in global memory (n entries, n >= k).  On success the
// kept candidates (every key whose score is >= [lower edge of the histogram bin that holds the k-th
// best score, minus one bin, minus margin]) are compacted into mine[0..kept) in index order, the
// count and the threshold are published and true is returned.  Returns false (block-uniform, nothing
// modified) when the scores are degenerate or the kept set would exceed k_cap: the caller then takes
// the exact sort path.
__device__ bool compact_select(uint64_t *keys, uint64_t *mine, int n, int k, int k_cap, float margin, uint32_t *cand_count,
                               uint64_t *tau_key) {
    __shared__ int s_hist[kSelBins];
    __shared__ float s_min[kCompactThreads / 32], s_max[kCompactThreads / 32];
    __shared__ int s_part[kCompactThreads / 32];
    __shared__ int s_bstar;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float lo = INFINITY, hi = -INFINITY;
    for (int i = tid; i < n; i += kCompactThreads) {
        const uint64_t key = mine[i];
        keys[i] = key;
        const float s = key_score(key);
        lo = fminf(lo, s);
        hi = fmaxf(hi, s);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if (lane == 0) {
        s_min[warp] = lo;
        s_max[warp] = hi;
    }
    for (int b = tid; b < kSelBins; b += kCompactThreads) s_hist[b] = 0;
    __syncthreads();
    lo = s_min[0];
    hi = s_max[0];
    for (int w = 1; w < kCompactThreads / 32; w++) {
        lo = fminf(lo, s_min[w]);
        hi = fmaxf(hi, s_max[w]);
    }
    if (!(hi > lo) || !(hi - lo < INFINITY)) return false;
    const float scale = (float)kSelBins / (hi - lo);
    for (int i = tid; i < n; i += kCompactThreads) {
        const int b = min(kSelBins - 1, max(0, (int)((key_score(keys[i]) - lo) * scale)));
        atomicAdd(&s_hist[b], 1);
    }
    __syncthreads();
    // suffix counts from the top bin down: thread t owns bins 2p and 2p+1 with p = last - t
    const int p = kCompactThreads - 1 - tid;
    const int h0 = s_hist[2 * p], h1 = s_hist[2 * p + 1];
    int incl = h0 + h1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) s_part[warp] = incl;
    __syncthreads();
    int before = 0;
    for (int w = 0; w < warp; w++) before += s_part[w];
    incl += before;                     // keys in bins >= 2p
    const int excl = incl - (h0 + h1);  // keys in bins > 2p+1
    if (excl < k && k <= incl) s_bstar = (excl + h1 >= k) ? 2 * p + 1 : 2 * p;
    __syncthreads();
    const float thr = lo + (float)(s_bstar - 1) / scale - margin;
    const uint64_t thr_key = make_key(thr, 0xffffffffu);
    // first count, then (only if it fits) compact: a refused call must leave mine[] untouched
    int kept = 0;
    for (int i = tid; i < n; i += kCompactThreads) kept += keys[i] > thr_key ? 1 : 0;
    kept = __reduce_add_sync(0xffffffffu, kept);
    __syncthreads();                    // s_part is reused
    if (lane == 0) s_part[warp] = kept;
    __syncthreads();
    kept = 0;
    for (int w = 0; w < kCompactThreads / 32; w++) kept += s_part[w];
    if (kept > k_cap) return false;
    int done = 0;
    for (int base = 0; base < n; base += kCompactThreads) {
        const int i = base + tid;
        const uint64_t me = i < n ? keys[i] : 0ull;
        const int flag = (i < n && me > thr_key) ? 1 : 0;
        const unsigned ballot = __ballot_sync(0xffffffffu, flag);
        __syncthreads();
        if (lane == 0) s_part[warp] = __popc(ballot);
        __syncthreads();
        int offset = done, block_total = 0;
        for (int w = 0; w < kCompactThreads / 32; w++) {
            const int c = s_part[w];
            if (w < warp) offset += c;
            block_total += c;
        }
        if (flag) mine[offset + __popc(ballot & ((1u << lane) - 1u))] = me;
        done += block_total;
    }
    if (tid == 0) {
        *cand_count = (uint32_t)kept;
        *tau_key = thr_key;
    }
    return true;
}

// One CTA per query: sort this round's candidates, drop duplicates (a retried level re-adds
// rows), keep the best k_keep, publish the new threshold; the last CTA to finish advances
// (or repeats) the level - the host never has to look.
__global__ void __launch_bounds__(kCompactThreads)
compact_kernel(SearchScratch sc, int nq, int k, int k_cap, float margin, int num_levels) {
    extern __shared__ uint64_t keys[];  // [kCandCap]
    __shared__ int s_scan[kCompactThreads / 32];
    __shared__ int s_is_last;
    pdl_enter();
    const int qi = blockIdx.x;
    const int tid = threadIdx.x;
    SearchCtrl *ctrl = sc.ctrl;
    const int level = ctrl->cur_level;
    if (level < num_levels && sc.active[qi]) {
        const uint32_t raw = sc.cand_count[qi];
        const int n = (int)min(raw, (uint32_t)kCandCap);
        const bool ovf = raw > (uint32_t)kCandCap;
        uint64_t *mine = sc.cand + (size_t)qi * kCandCap;
        // Tensor path, ordinary round: nothing here needs ORDER, only "the k-th best score, minus the
        // margin, as a threshold" - and any threshold at or below that is still correct (it keeps a
        // superset).  So select instead of sort: histogram of the scores, the bin where the k-th best
        // falls, its lower edge (one bin of slack) minus the margin.  Falls back to the exact sort
        // path below when the round overflowed or is a retry (duplicates), when fewer than k
        // candidates exist, or when the kept set would not fit.
        bool fast_done = false;
        bool gave_up = false;   // thread 0: exactness is lost for this query, retrying the round cannot help
        if (margin > 0.f && !ovf && !ctrl->retry && n >= k) {
            fast_done = compact_select(keys, mine, n, k, k_cap, margin, &sc.cand_count[qi], &sc.tau_key[qi]);
        }
        if (!fast_done) {
        int P = 32;
        while (P < n) P <<= 1;
        for (int i = tid; i < P; i += kCompactThreads) keys[i] = i < n ? mine[i] : 0ull;
        bitonic_sort_desc<kCompactThreads>(keys, P);
        // dedupe + keep the first k_cap unique keys (block-wide exclusive scan of the "is new" flags)
        int kept_total = 0;
        int base = 0;
        for (; base < n && kept_total < k_cap; base += kCompactThreads) {
            const int i = base + tid;
            const uint64_t me = i < n ? keys[i] : 0ull;
            const int flag = (i < n && (i == 0 || keys[i - 1] != me)) ? 1 : 0;
            const unsigned ballot = __ballot_sync(0xffffffffu, flag);
            const int lane = tid & 31, warp = tid >> 5;
            if (lane == 0) s_scan[warp] = __popc(ballot);
            __syncthreads();
            int offset = kept_total;
            for (int w = 0; w < warp; w++) offset += s_scan[w];
            int block_total = 0;
            for (int w = 0; w < kCompactThreads / 32; w++) block_total += s_scan[w];
            const int pos = offset + __popc(ballot & ((1u << lane) - 1u));
            if (flag && pos < k_cap) mine[pos] = me;
            kept_total += block_total;
            __syncthreads();
        }
        __syncthreads();  // mine[0..u) written by this block
        const int u = min(kept_total, k_cap);
        if (margin <= 0.f) {
            // exact scores: keep the k best; the k-th key itself is the admission threshold
            if (tid == 0) {
                const int keep = min(u, k);
                sc.cand_count[qi] = keep;
                sc.tau_key[qi] = keep == k ? mine[k - 1] : kTauMin;  // valid only once k candidates exist
            }
        } else {
            // approximate scores (tensor pass): everything within `margin` of the k-th best may still
            // be in the exact top-k, so it is kept (up to k_cap) and admitted from now on
            __shared__ int s_keep;
            if (tid == 0) s_keep = u;
            __syncthreads();
            if (u >= k) {
                const uint64_t thr_key = make_key(key_score(mine[k - 1]) - margin, 0xffffffffu);
                for (int i = tid; i < u; i += kCompactThreads) {
                    if (mine[i] > thr_key && (i + 1 == u || mine[i + 1] <= thr_key)) s_keep = i + 1;
                }
                __syncthreads();
                if (tid == 0) {
                    sc.tau_key[qi] = thr_key;
                    // more rows inside the margin than we can keep: the exactness guarantee is gone
                    // (sorted keys were left unread, or more unique keys than slots, or the buffer itself overflowed)
                    if (s_keep == u && u == k_cap && (base < n || kept_total > k_cap || ovf)) {
                        atomicExch(&ctrl->inexact, 1);
                        if (sc.inexact_q) sc.inexact_q[qi] = 1;   // this query's outputs get poisoned; the others stay valid
                        // Everything kept sits inside the margin, so the threshold cannot tighten: a re-scan would
                        // overflow again, forever.  Let the search finish; its verdict is DAB_E_INEXACT and the
                        // caller repeats it on the CUDA-core path.
                        gave_up = true;
                    }
                }
            } else if (tid == 0) {
                sc.tau_key[qi] = kTauMin;
            }
            __syncthreads();
            if (tid == 0) sc.cand_count[qi] = s_keep;
        }
        }  // !fast_done
        if (tid == 0) {
            const bool again = ovf && !gave_up;
            sc.overflowed[qi] = again ? 1 : 0;
            if (again) atomicExch(&ctrl->any_overflow, 1);
        }
    }
    // ---- round control by the last CTA
    __threadfence();
    __syncthreads();
    if (tid == 0) s_is_last = (atomicAdd(&ctrl->ticket, 1u) == gridDim.x - 1) ? 1 : 0;
    __syncthreads();
    if (!s_is_last) return;
    __threadfence();
    const int any = *(volatile int *)&ctrl->any_overflow;
    if (level < num_levels) {
        for (int q = tid; q < nq; q += kCompactThreads) sc.active[q] = any ? sc.overflowed[q] : 1;
    }
    if (tid == 0) {
        if (level < num_levels) {
            if (any) {
                ctrl->retry = 1;
                ctrl->overflow_events += 1;
            } else {
                ctrl->retry = 0;
                ctrl->cur_level = level + 1;
            }
            ctrl->rounds_run += 1;
        }
        ctrl->any_overflow = 0;
        ctrl->ticket = 0;
    }
}

// the finish kernels' single bookkeeping thread: publish this call's verdict so far (see SearchScratch)
__device__ __forceinline__ void publish_verdict(const SearchScratch &sc) {
    const int inc = sc.ctrl->sticky_incomplete, inx = sc.ctrl->sticky_inexact;
    if (sc.verdict_dev) {
        sc.verdict_dev[0] = inc;
        sc.verdict_dev[1] = inx;
    }
    if (sc.verdict_host) {
        sc.verdict_host[0] = inc;
        sc.verdict_host[1] = inx;
        __threadfence_system();
    }
}

// ---------------------------------------------------------------- finish (exact paths): keys -> outputs
__global__ void __launch_bounds__(128)
finish_exact_kernel(SearchScratch sc, StoreView sv, int k, int num_levels, float *__restrict__ out_scores,
                    int64_t *__restrict__ out_ids) {
    pdl_enter();
    const int qi = blockIdx.x;
    const bool done = sc.ctrl->cur_level >= num_levels;
    const int cnt = done ? min((int)sc.cand_count[qi], k) : 0;
    const uint64_t *mine = sc.cand + (size_t)qi * kCandCap;
    if (qi == 0 && threadIdx.x == 0) {
        if (!done) sc.ctrl->sticky_incomplete = 1;
        publish_verdict(sc);
    }
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        float s;
        int64_t id;
        if (j < cnt) {
            const uint64_t key = mine[j];
            const uint32_t row = key_row(key);
            s = key_score(key);
            id = sv.ids ? sv.ids[row] : sv.id_base + (int64_t)row;
        } else {
            s = done ? -INFINITY : __int_as_float(0x7fc00000);
            id = -1;
        }
        out_scores[(size_t)qi * k + j] = s;
        out_ids[(size_t)qi * k + j] = id;
    }
}

// ---------------------------------------------------------------- finish (tensor path): exact fp32 rescoring
// The tensor pass ranks with approximate scores; the kept k_keep >= k candidates are rescored
// here with the fp32 query on CUDA cores (fp32 FMA over the stored bf16 row, upcast exactly),
// re-sorted under (score desc, row asc) and cut to k.
constexpr int kRescoreThreads = 256;
constexpr int kRescoreSplit = 8;    // CTAs that share the rescoring of one query's kept candidates

// exact fp32 score of a stored row with the unit-norm fp32 query (bf16 rows upcast exactly); all lanes return it
__device__ __forceinline__ float rescore_row(const StoreView &sv, const float *__restrict__ q, uint32_t row, int lane) {
    float acc = 0.f;
    if (sv.dtype == DAB_BF16) {
        const uint4 *r = reinterpret_cast<const uint4 *>(static_cast<const __nv_bfloat16 *>(sv.rows) + (size_t)row * sv.dim);
        for (int c = lane; c < sv.dim / 8; c += 32) {
            const uint4 v = r[c];
            const float4 q0 = reinterpret_cast<const float4 *>(q)[2 * c];
            const float4 q1 = reinterpret_cast<const float4 *>(q)[2 * c + 1];
            acc = fmaf(bf16_lo(v.x), q0.x, acc);
            acc = fmaf(bf16_hi(v.x), q0.y, acc);
            acc = fmaf(bf16_lo(v.y), q0.z, acc);
            acc = fmaf(bf16_hi(v.y), q0.w, acc);
            acc = fmaf(bf16_lo(v.z), q1.x, acc);
            acc = fmaf(bf16_hi(v.z), q1.y, acc);
            acc = fmaf(bf16_lo(v.w), q1.z, acc);
            acc = fmaf(bf16_hi(v.w), q1.w, acc);
        }
    } else {
        const float4 *r = reinterpret_cast<const float4 *>(static_cast<const float *>(sv.rows) + (size_t)row * sv.dim);
        for (int c = lane; c < sv.dim / 4; c += 32) {
            const float4 v = r[c];
            const float4 qq = reinterpret_cast<const float4 *>(q)[c];
            acc = fmaf(v.x, qq.x, acc);
            acc = fmaf(v.y, qq.y, acc);
            acc = fmaf(v.z, qq.z, acc);
            acc = fmaf(v.w, qq.w, acc);
        }
    }
    return warp_sum(acc);
}

// Grid (kRescoreSplit, Q).  The kept candidates of a query (up to 2048 random rows, ~1.5 KB each) are re-scored by
// kRescoreSplit CTAs - one CTA per query left Q SMs chasing row latencies one warp-row at a time, 80-130 us at
// Q = 64 - which overwrite the approximate keys in place with exact ones; the CTA that finishes last (ticket)
// sorts them under (score desc, row asc) and writes the k outputs.
__global__ void __launch_bounds__(kRescoreThreads)
finish_rescore_kernel(SearchScratch sc, StoreView sv, int k, int k_keep, int num_levels,
                      float *__restrict__ out_scores, int64_t *__restrict__ out_ids) {
    extern __shared__ uint64_t keys[];  // [pow2 >= k_keep]
    __shared__ int s_last;
    pdl_enter();
    const int qi = blockIdx.y;
    const int part = blockIdx.x, parts = gridDim.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // a query with more rows inside the rescoring margin than could be kept is answered with poison (NaN, -1):
    // the caller sees DAB_E_INEXACT and repeats exactly those queries on the CUDA-core path; the rest are final
    const bool lost = sc.inexact_q != nullptr && sc.inexact_q[qi] != 0;
    const bool done = sc.ctrl->cur_level >= num_levels && !lost;
    const int cnt = done ? min((int)sc.cand_count[qi], k_keep) : 0;
    uint64_t *mine = sc.cand + (size_t)qi * kCandCap;
    const float *q = sc.qhat + (size_t)qi * sv.dim;
    // ---- my share of the candidates: warp w of part p takes j = p * warps + w, + parts * warps, ... (two rows in flight)
    constexpr int kWarps = kRescoreThreads / 32;
    const int stride = parts * kWarps;
    for (int j = part * kWarps + warp; j < cnt; j += 2 * stride) {
        const int j2 = j + stride;
        const uint32_t row_a = key_row(mine[j]);
        const uint32_t row_b = j2 < cnt ? key_row(mine[j2]) : row_a;
        const float sa = rescore_row(sv, q, row_a, lane);
        const float sb = rescore_row(sv, q, row_b, lane);
        if (lane == 0) {
            mine[j] = make_key(sa * sv.inv_norm[row_a], row_a);
            if (j2 < cnt) mine[j2] = make_key(sb * sv.inv_norm[row_b], row_b);
        }
    }
    // ---- the last CTA of this query's group finishes
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = (parts == 1 || atomicAdd(&sc.finish_ticket[qi], 1u) == (unsigned)(parts - 1)) ? 1 : 0;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (qi == 0 && tid == 0) {
        if (sc.ctrl->cur_level < num_levels) sc.ctrl->sticky_incomplete = 1;
        if (sc.ctrl->inexact) sc.ctrl->sticky_inexact = 1;
        publish_verdict(sc);
    }
    int P = 32;
    while (P < cnt) P <<= 1;
    for (int i = tid; i < P; i += kRescoreThreads) keys[i] = i < cnt ? __ldcg(&mine[i]) : 0ull;
    bitonic_sort_desc<kRescoreThreads>(keys, P);
    const int outn = min(cnt, k);
    for (int j = tid; j < k; j += kRescoreThreads) {
        float s;
        int64_t id;
        if (j < outn) {
            const uint64_t key = keys[j];
            const uint32_t row = key_row(key);
            s = key_score(key);
            id = sv.ids ? sv.ids[row] : sv.id_base + (int64_t)row;
        } else {
            s = done ? -INFINITY : __int_as_float(0x7fc00000);
            id = -1;
        }
        out_scores[(size_t)qi * k + j] = s;
        out_ids[(size_t)qi * k + j] = id;
    }
}

__global__ void fill_empty_kernel(float *out_scores, int64_t *out_ids, int64_t count) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) {
        out_scores[i] = -INFINITY;
        out_ids[i] = -1;
    }
}

// ---------------------------------------------------------------- k-way merge of per-shard lists
struct MergeItem {
    uint32_t s;   // orderable score
    int64_t id;   // < 0 = empty
};
__device__ __forceinline__ bool merge_better(const MergeItem &a, const MergeItem &b) {
    const bool ae = a.id < 0, be = b.id < 0;
    if (ae != be) return be;
    if (a.s != b.s) return a.s > b.s;
    return a.id < b.id;
}
constexpr int kMergeThreads = 256;
constexpr int kMergeMax = 8192;
__global__ void __launch_bounds__(kMergeThreads)
merge_topk_kernel(const float *__restrict__ gs, const int64_t *__restrict__ gi, int R, int Q, int k,
                  float *__restrict__ out_scores, int64_t *__restrict__ out_ids) {
    extern __shared__ unsigned char smem_raw[];
    uint32_t *ss = reinterpret_cast<uint32_t *>(smem_raw);
    int n = R * k;
    int P = 32;
    while (P < n) P <<= 1;
    int64_t *si = reinterpret_cast<int64_t *>(smem_raw + (size_t)P * sizeof(uint32_t));
    const int qi = blockIdx.x;
    const int tid = threadIdx.x;
    for (int i = tid; i < P; i += kMergeThreads) {
        if (i < n) {
            const int r = i / k, j = i % k;
            const size_t src = ((size_t)r * Q + qi) * k + j;
            const int64_t id = gi[src];
            const float sc = gs[src];
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
            for (int i = tid; i < (P >> 1); i += kMergeThreads) {
                const int lo = 2 * i - (i & (stride - 1));
                const int hi = lo + stride;
                const bool desc = (lo & size) == 0;
                MergeItem a{ss[lo], si[lo]}, b{ss[hi], si[hi]};
                if (merge_better(b, a) == desc && (a.s != b.s || a.id != b.id)) {
                    ss[lo] = b.s; si[lo] = b.id;
                    ss[hi] = a.s; si[hi] = a.id;
                }
            }
        }
    }
    __syncthreads();
    for (int j = tid; j < k; j += kMergeThreads) {
        const bool ok = j < n && si[j] >= 0;
        out_scores[(size_t)qi * k + j] = ok ? orderable_f32(ss[j]) : -INFINITY;
        out_ids[(size_t)qi * k + j] = ok ? si[j] : -1;
    }
}

}  // namespace

int launch_search_begin(const SearchScratch &sc, const float *q_dev, int nq, int dim, bool want_split, cudaStream_t st,
                        bool first_chunk) {
    DAB_CUDA_OK(launch_chained(search_begin_kernel, dim3(nq), dim3(128), 0, st, sc, q_dev, dim, want_split ? 1 : 0, first_chunk ? 1 : 0));
    count_launch();
    return DAB_OK;
}

int launch_compact(const SearchScratch &sc, int nq, int k, int num_levels, cudaStream_t st) {
    return launch_compact_margin(sc, nq, k, k, 0.f, num_levels, st);
}

int launch_compact_margin(const SearchScratch &sc, int nq, int k, int k_cap, float margin, int num_levels, cudaStream_t st) {
    const int smem = kCandCap * sizeof(uint64_t);
    static SmemAttrOnce attr_once;
    DAB_CUDA_OK(attr_once.ensure(compact_kernel, smem));
    DAB_CUDA_OK(launch_chained(compact_kernel, dim3(nq), dim3(kCompactThreads), smem, st, sc, nq, k, k_cap, margin, num_levels));
    count_launch();
    return DAB_OK;
}

int launch_finish_exact(const SearchScratch &sc, const StoreView &sv, int nq, int k, int num_levels, float *out_scores,
                        int64_t *out_ids, cudaStream_t st) {
    DAB_CUDA_OK(launch_chained(finish_exact_kernel, dim3(nq), dim3(128), 0, st, sc, sv, k, num_levels, out_scores, out_ids));
    count_launch();
    return DAB_OK;
}

int launch_finish_rescore(const SearchScratch &sc, const StoreView &sv, int nq, int k, int k_keep, int num_levels,
                          float *out_scores, int64_t *out_ids, cudaStream_t st) {
    int P = 32;
    while (P < k_keep) P <<= 1;
    const int parts = sc.finish_ticket ? kRescoreSplit : 1;
    DAB_CUDA_OK(launch_chained(finish_rescore_kernel, dim3(parts, nq), dim3(kRescoreThreads), P * sizeof(uint64_t), st, sc, sv, k, k_keep,
                               num_levels, out_scores, out_ids));
    count_launch();
    return DAB_OK;
}

int launch_fill_empty(float *out_scores, int64_t *out_ids, int64_t count, cudaStream_t st) {
    if (count <= 0) return DAB_OK;
    fill_empty_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(out_scores, out_ids, count);
    DAB_LAUNCH_OK();
    return DAB_OK;
}

}  // namespace dab

extern "C" DAB_API int dab_merge_topk(const float *gathered_scores, const int64_t *gathered_ids, int R, int Q, int k,
                              float *out_scores, int64_t *out_ids, void *stream) {
    using namespace dab;
    if (!gathered_scores || !gathered_ids || !out_scores || !out_ids) return fail(DAB_E_INVALID, "dab_merge_topk: NULL pointer");
    if (R <= 0 || Q < 0 || k <= 0) return fail(DAB_E_INVALID, "dab_merge_topk: R, k must be positive");
    if ((long long)R * k > kMergeMax) return fail(DAB_E_CAPACITY, "dab_merge_topk: R*k exceeds 8192");
    if (Q == 0) return DAB_OK;
    int P = 32;
    while (P < R * k) P <<= 1;
    const int smem = P * (int)(sizeof(uint32_t) + sizeof(int64_t));
    static SmemAttrOnce attr_once;
    DAB_CUDA_OK(attr_once.ensure(merge_topk_kernel, kMergeMax * (int)(sizeof(uint32_t) + sizeof(int64_t))));
    merge_topk_kernel<<<Q, kMergeThreads, smem, static_cast<cudaStream_t>(stream)>>>(gathered_scores, gathered_ids, R, Q, k,
                                                                                      out_scores, out_ids);
    DAB_LAUNCH_OK();
    return DAB_OK;
}
