// store.cuh - the dab_store object shared by store.cu (lifecycle, search) and filter.cu (queryset filters).
#pragma once
#include <mutex>
#include <utility>
#include <vector>

#include "search.cuh"

// cudaEvent pairs around the scan launches of a search (only when profiling is switched on)
struct dab_profile {
    bool on = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> scan_events;  // per scan launch
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> search_events;  // per search call
};

struct dab_store {
    int device = 0;
    int dim = 0;
    int dtype = DAB_F32;
    int64_t capacity = 0;
    int64_t size = 0;
    int64_t id_base = 0;
    bool explicit_ids = false;
    void *rows = nullptr;
    float *inv_norm = nullptr;
    int64_t *ids = nullptr;
    void *scratch_base = nullptr;
    dab::SearchScratch scratch{};
    int num_sms = 148;
    int spare_rounds = 0;
    int path = 0;
    dab_profile prof;
    // Per-call verdicts: search number `seq` owns slot seq % kVerdictSlots of both rings; its finish kernels write
    // {incomplete, inexact} there (HBM ring for the exchange kernel, pinned ring for dab_store_search_status).
    int *verdict_ring_dev = nullptr;
    int *verdict_ring_host = nullptr;
    uint64_t search_seq = 0;
    // staging of dab_store_search_host (allocated on first use): pinned host + device buffers for one
    // chunk of DAB_MAX_QUERY_BATCH queries and their DAB_MAX_K results
    void *io_host = nullptr;
    void *io_dev = nullptr;
    std::mutex io_mu;  // one host-buffer search at a time per store (they share the staging buffers)
    // Searches on one store share its scratch: the mutex orders their ENQUEUE, this event orders their
    // EXECUTION when callers use different streams (each search waits for the previous one's last kernel).
    cudaEvent_t scratch_free = nullptr;
    // Auto path, few queries: the tensor pass is the faster choice unless the corpus is so dense around the
    // top-k that the rescoring margin overflows (DAB_E_INEXACT) and the call has to be repeated on the CUDA
    // cores.  After such a verdict the next `small_simt_credit` small searches go to the CUDA cores directly.
    int small_simt_credit = 0;
    // Filter columns (dab_store_set_column): int64 per row, allocated on first use, aligned with the store rows
    int64_t *columns[DAB_MAX_COLUMNS] = {};
    std::mutex mu;
};

