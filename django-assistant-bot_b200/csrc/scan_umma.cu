// scan_umma.cu - tcgen05 bf16 tensor-core scoring path (placeholder until the kernel lands).
#include "search.cuh"

namespace dab {
int umma_tile_rows() { return 128; }
bool umma_supported(int, int) { return false; }
void umma_release_maps(void *) {}
int launch_scan_umma(const StoreView &, const uint32_t *, const SearchScratch &, int, bool, const LevelPlan &, int, void *, void *,
                     cudaStream_t) {
    return fail(DAB_E_INVALID, "tcgen05 path not built");
}
}  // namespace dab
