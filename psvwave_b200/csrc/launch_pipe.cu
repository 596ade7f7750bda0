// Variant 3 (the march fed by a TMA bulk-copy ring, fused.cuh): kernel instantiations + launchers.
#include <type_traits>

#include "fused.cuh"
#include "fused32.cuh"
#include "launchers.hpp"
#include "model.hpp"

namespace psv {

namespace {
template <class F>
auto with_pipe_warps(int nw, F &&f) {
  switch (nw) {
    case 8: return f(std::integral_constant<int, 8>{});
    case 4: return f(std::integral_constant<int, 4>{});
    default: return f(std::integral_constant<int, 3>{});
  }
}

template <int NW>
size_t smem_bytes(int stages, int arrays) {
  return 128 * ((2 * (size_t)stages * 8 + 127) / 128) + (size_t)stages * PipeGeom<NW>::kSeg * arrays;
}

template <class M, int MODE, int NW>
void launch_k(const FusedCommon &fc, const FusedModel *models, dim3 grid, int stages, size_t smem, cudaStream_t st) {
  static SmemOptIn opt_in;  // per instantiation
  opt_in.ensure(fused_pipe_kernel<M, MODE, NW>, smem);
  if (fc.pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3((NW + 1) * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    PSV_CUDA(cudaLaunchKernelEx(&cfg, fused_pipe_kernel<M, MODE, NW>, fc, models, stages));
    return;
  }
  fused_pipe_kernel<M, MODE, NW><<<grid, (NW + 1) * 32, smem, st>>>(fc, models, stages);
}

template <class M, int NW>
void launch_m(int mode, const FusedCommon &fc, const FusedModel *models, dim3 grid, int stages, size_t smem,
              cudaStream_t st) {
  if (mode == MODE_SNAPSHOT) launch_k<M, MODE_SNAPSHOT, NW>(fc, models, grid, stages, smem, st);
  else if (mode == MODE_ACCUMULATE) launch_k<M, MODE_ACCUMULATE, NW>(fc, models, grid, stages, smem, st);
  else launch_k<M, MODE_PLAIN, NW>(fc, models, grid, stages, smem, st);
}
}  // namespace

size_t pipe_smem_bytes(int nw, int stages, bool corr) {
  const int arrays = corr ? kPipeArraysCorr : kPipeArrays;
  return with_pipe_warps(nw, [&](auto w) { return smem_bytes<decltype(w)::value>(stages, arrays); });
}

int pipe_corr_stages(int nw, int stages) {
  while (stages > 2 && pipe_smem_bytes(nw, stages, true) > kMaxDynamicSmem) --stages;
  return stages;
}

int pipe_occupancy(int nw, size_t smem) {
  return with_pipe_warps(nw, [&](auto w) {
    constexpr int NW = decltype(w)::value;
    int per_sm = 0;
    PSV_CUDA(cudaFuncSetAttribute(fused_pipe_kernel<FastMath, MODE_PLAIN, NW>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    PSV_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fused_pipe_kernel<FastMath, MODE_PLAIN, NW>,
                                                           (NW + 1) * 32, smem));
    return per_sm;
  });
}

// ---- the FP32 march on the same ring (fused32.cuh) ----
namespace {
template <int MODE, int NW>
void launch32_k(const FusedCommon &fc, const FusedModel32 *models, dim3 grid, int stages, size_t smem, cudaStream_t st) {
  static SmemOptIn opt_in;  // per instantiation
  opt_in.ensure(fused32_pipe_kernel<MODE, NW>, smem);
  fused32_pipe_kernel<MODE, NW><<<grid, (NW + 1) * 32, smem, st>>>(fc, models, stages);
}
}  // namespace

size_t pipe32_smem_bytes(int nw, int stages) {
  return with_pipe_warps(nw, [&](auto w) {
    return 128 * ((2 * (size_t)stages * 8 + 127) / 128) + (size_t)stages * PipeGeom32<decltype(w)::value>::kStage;
  });
}

void launch_pipe32(int nw, int mode, const FusedCommon &fc, const FusedModel32 *models, dim3 grid, int stages,
                   size_t smem, cudaStream_t st) {
  with_pipe_warps(nw, [&](auto w) {
    constexpr int NW = decltype(w)::value;
    if (mode == MODE_SNAPSHOT) launch32_k<MODE_SNAPSHOT, NW>(fc, models, grid, stages, smem, st);
    else if (mode == MODE_ACCUMULATE) launch32_k<MODE_ACCUMULATE, NW>(fc, models, grid, stages, smem, st);
    else launch32_k<MODE_PLAIN, NW>(fc, models, grid, stages, smem, st);
    return 0;
  });
}

void launch_pipe(bool exact, int nw, int mode, const FusedCommon &fc, const FusedModel *models, dim3 grid, int stages,
                 size_t smem, cudaStream_t st) {
  with_pipe_warps(nw, [&](auto w) {
    constexpr int NW = decltype(w)::value;
    if (exact) launch_m<ExactMath, NW>(mode, fc, models, grid, stages, smem, st);
    else launch_m<FastMath, NW>(mode, fc, models, grid, stages, smem, st);
    return 0;
  });
}

}  // namespace psv
