// search.cuh - state shared by the scan kernels, the selection kernels and the store.
#pragma once
#include "common.cuh"

namespace dab {

// smallest key any real (finite-score) candidate beats: score -inf, row 0xffffffff
constexpr uint64_t kTauMin = 0x007FFFFF00000000ull;

// Per-store device scratch used by a search (allocated once in dab_store_create, or handed
// in by the caller of the no-allocation entry point).  All arrays are sized for
// DAB_MAX_QUERY_BATCH queries.
struct SearchScratch {
    float *qhat;            // [256, dim] unit-norm fp32 queries
    __nv_bfloat16 *qbf16;   // [256, dim] qhat rounded to bf16: the B operand of the tensor pass on bf16 stores
    uint64_t *cand;         // [256, kCandCap] candidate keys
    uint32_t *cand_count;   // [256] appended this far (may exceed kCandCap: overflow)
    uint64_t *tau_key;      // [256] current admission threshold (a kept candidate's key)
    int *active;            // [256] query takes part in the current round
    int *overflowed;        // [256] query overflowed in the current round
    int *inexact_q;         // [256] tensor path: more rows inside the margin than kept -> outputs poisoned (may be null)
    unsigned int *finish_ticket;  // [256] tensor path: CTAs of a query's rescoring group that are done (may be null)
    SearchCtrl *ctrl;
    // Per-CALL verdict slots {incomplete, inexact}: the finish kernel of every query chunk copies the sticky flags
    // there, so the verdict of a search survives the next search's reset of `ctrl` (two threads on one store).
    // verdict_dev lives in HBM (read by dab_exchange_merge_topk), verdict_host in pinned host memory (read by
    // dab_store_search_status after a stream synchronise - no copy).  Either may be null.
    int *verdict_dev;
    int *verdict_host;
};

size_t search_scratch_bytes(int dim);
// carve `base` (256-byte aligned, search_scratch_bytes(dim) long) into a SearchScratch
SearchScratch carve_scratch(void *base, int dim);

struct StoreView {
    const void *rows;        // [n, dim] fp32 or bf16
    const float *inv_norm;   // [n] (padded to a multiple of 4 with zeros)
    const int64_t *ids;      // [n] or nullptr (implicit id_base + row)
    int64_t id_base;
    int64_t n;
    int dim;
    int dtype;
};

LevelPlan make_level_plan(int64_t n_rows, int tile_rows, int k_keep);

// ---- kernels' host launchers (each returns DAB_OK or an error code) ----
int launch_search_begin(const SearchScratch &sc, const float *q_dev, int nq, int dim, bool want_split, cudaStream_t st,
                        bool first_chunk = true);
int launch_compact(const SearchScratch &sc, int nq, int k, int num_levels, cudaStream_t st);
// tensor path: keep every candidate within `margin` of the k-th best approximate score (<= k_cap)
int launch_compact_margin(const SearchScratch &sc, int nq, int k, int k_cap, float margin, int num_levels, cudaStream_t st);
int launch_finish_exact(const SearchScratch &sc, const StoreView &sv, int nq, int k, int num_levels, float *out_scores,
                        int64_t *out_ids, cudaStream_t st);
int launch_finish_rescore(const SearchScratch &sc, const StoreView &sv, int nq, int k, int k_keep, int num_levels,
                          float *out_scores, int64_t *out_ids, cudaStream_t st);
int launch_fill_empty(float *out_scores, int64_t *out_ids, int64_t count, cudaStream_t st);

int simt_tile_rows(int dim, int dtype);
int launch_scan_simt(const StoreView &sv, const uint32_t *allow, const SearchScratch &sc, int q0, int nq,
                     const LevelPlan &plan, int num_sms, cudaStream_t st);

int umma_tile_rows();
bool umma_supported(int dim, int dtype);
int umma_queries_per_pass(int dim, int dtype = DAB_BF16);
int launch_scan_umma(const StoreView &sv, const uint32_t *allow, const SearchScratch &sc, int q0, int nq,
                     const LevelPlan &plan, int num_sms, cudaStream_t st);


// wide query blocks (up to 256): 4-CTA clusters, cta_group::2 MMA pairs, multicast corpus tiles
int quad_tile_rows();
bool quad_supported(int dim, int dtype, int nq);
int quad_queries_per_pass(int dim, int dtype);
int pair_queries_per_pass(int dim, int dtype = DAB_BF16);
// num_pairs = 2: one 4-CTA cluster scores up to 256 queries per pass (tiles multicast to both pairs);
// num_pairs = 1: a 2-CTA cluster scores up to pair_queries_per_pass(dim) queries, every tile byte delivered once
int launch_scan_umma_quad(const StoreView &sv, const uint32_t *allow, const SearchScratch &sc, int q0, int nq,
                          const LevelPlan &plan, int num_sms, cudaStream_t st, int num_pairs = 2);


}  // namespace dab
