// Instances of the scan kernel (scan_kernel.cuh): <NQ=4, KPL=1>, <NQ=4, KPL=4>
#include "scan_kernel.cuh"

namespace b200r {
cudaError_t scan_launch_4_1(const ScanParams& p, bool l2, bool bf16, int grid, size_t smem, cudaStream_t st, bool flagged) {
    return launch_mk<4, 1>(p, l2, bf16, grid, smem, st, flagged);
}
cudaError_t scan_launch_4_4(const ScanParams& p, bool l2, bool bf16, int grid, size_t smem, cudaStream_t st, bool flagged) {
    return launch_mk<4, 4>(p, l2, bf16, grid, smem, st, flagged);
}
}  // namespace b200r
