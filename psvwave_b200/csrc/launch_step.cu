// Variant 2 (fused march from global memory) and the optional FP32 mode: kernel instantiations + launchers.
#include "fused.cuh"
#include "fused32.cuh"
#include "launchers.hpp"
#include "model.hpp"

namespace psv {

namespace {
template <class M>
void launch_fused_m(int mode, const FusedCommon &fc, const FusedModel *models, dim3 grid, int block, cudaStream_t st) {
  if (mode == MODE_SNAPSHOT) fused_step_kernel<M, MODE_SNAPSHOT><<<grid, block, 0, st>>>(fc, models);
  else if (mode == MODE_ACCUMULATE) fused_step_kernel<M, MODE_ACCUMULATE><<<grid, block, 0, st>>>(fc, models);
  else fused_step_kernel<M, MODE_PLAIN><<<grid, block, 0, st>>>(fc, models);
}
}  // namespace

void launch_fused_step(bool exact, int mode, const FusedCommon &fc, const FusedModel *models, dim3 grid, int block,
                       cudaStream_t st) {
  if (exact) launch_fused_m<ExactMath>(mode, fc, models, grid, block, st);
  else launch_fused_m<FastMath>(mode, fc, models, grid, block, st);
}

int fused_step_occupancy(int block) {
  int per_sm = 0;
  PSV_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fused_step_kernel<FastMath, MODE_PLAIN>, block, 0));
  return per_sm;
}

void launch_fused32(int mode, const FusedCommon &fc, const FusedModel32 *m32, dim3 grid, cudaStream_t st) {
  if (mode == MODE_SNAPSHOT) fused32_step_kernel<MODE_SNAPSHOT><<<grid, 128, 0, st>>>(fc, m32);
  else if (mode == MODE_ACCUMULATE) fused32_step_kernel<MODE_ACCUMULATE><<<grid, 128, 0, st>>>(fc, m32);
  else fused32_step_kernel<MODE_PLAIN><<<grid, 128, 0, st>>>(fc, m32);
}

int fused32_occupancy() {
  int per_sm = 0;
  PSV_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fused32_step_kernel<MODE_PLAIN>, 128, 0));
  return per_sm;
}

}  // namespace psv
