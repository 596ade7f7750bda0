// Variant 4 (two time steps per launch, pair.cuh): kernel instantiations + launchers.
#include "launchers.hpp"
#include "model.hpp"
#include "pair.cuh"

namespace psv {

namespace {
template <class M, int MODE, int NW>
void launch_k(const FusedCommon &fc, const FusedModel *models, dim3 grid, int stages, size_t smem, cudaStream_t st) {
  static SmemOptIn opt_in;  // per instantiation
  opt_in.ensure(fused_pair_kernel<M, MODE, NW>, smem);
  fused_pair_kernel<M, MODE, NW><<<grid, 2 * NW * 32, smem, st>>>(fc, models, stages);
}
template <class M, int NW>
void launch_m(int mode, const FusedCommon &fc, const FusedModel *models, dim3 grid, int stages, size_t smem,
              cudaStream_t st) {
  if (mode == MODE_SNAPSHOT) launch_k<M, MODE_SNAPSHOT, NW>(fc, models, grid, stages, smem, st);
  else if (mode == MODE_ACCUMULATE) launch_k<M, MODE_ACCUMULATE, NW>(fc, models, grid, stages, smem, st);
  else launch_k<M, MODE_PLAIN, NW>(fc, models, grid, stages, smem, st);
}
}  // namespace

size_t pair_smem(int nw, int stages) { return nw == 4 ? pair_smem_bytes<4>(stages) : pair_smem_bytes<6>(stages); }

void launch_pair(bool exact, int nw, int mode, const FusedCommon &fc, const FusedModel *models, dim3 grid, int stages,
                 size_t smem, cudaStream_t st) {
  if (nw == 4) {
    if (exact) launch_m<ExactMath, 4>(mode, fc, models, grid, stages, smem, st);
    else launch_m<FastMath, 4>(mode, fc, models, grid, stages, smem, st);
  } else {
    if (exact) launch_m<ExactMath, 6>(mode, fc, models, grid, stages, smem, st);
    else launch_m<FastMath, 6>(mode, fc, models, grid, stages, smem, st);
  }
}

}  // namespace psv
