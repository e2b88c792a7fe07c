// Instances of the scan kernel (scan_kernel.cuh): <NQ=1, KPL=8>
#include "scan_kernel.cuh"

namespace b200r {
cudaError_t scan_launch_1_8(const ScanParams& p, bool l2, bool bf16, int grid, size_t smem, cudaStream_t st, bool flagged) {
    return launch_mk<1, 8>(p, l2, bf16, grid, smem, st, flagged);
}
}  // namespace b200r
