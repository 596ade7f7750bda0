// Host-side launch entry points of the step kernels.  Every kernel family is compiled in its own
// translation unit (launch_step.cu, launch_pipe.cu, launch_pair.cu, launch_resident.cu) so that the
// library builds in parallel; model.cu sees only these declarations.
#pragma once

#include <cuda_runtime.h>

#include "fused_types.cuh"

namespace psv {

constexpr size_t kMaxDynamicSmem = 232448;  // 227 KB per CTA on sm_100
constexpr int kMaxDevices = 64;

// Opt-in to more than 48 KB of dynamic shared memory is an attribute of (kernel, device): cache it per
// device, not per process (a handle may live on any device of the box).
struct SmemOptIn {
  size_t configured[kMaxDevices] = {};
  template <class K>
  void ensure(K kernel, size_t smem) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= kMaxDevices || configured[dev] < smem) {
      const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) {
        cudaGetLastError();
        return;  // the launch that follows reports the failure
      }
      if (dev >= 0 && dev < kMaxDevices) configured[dev] = smem;
    }
  }
};

// ---- variant 2: fused march, loads straight from global memory (fused.cuh) ----
void launch_fused_step(bool exact, int mode, const FusedCommon &fc, const FusedModel *models, dim3 grid, int block,
                       cudaStream_t st);
int fused_step_occupancy(int block);

// ---- optional FP32 wavefield mode (fused32.cuh) ----
void launch_fused32(int mode, const FusedCommon &fc, const FusedModel32 *models, dim3 grid, cudaStream_t st);
int fused32_occupancy();

// ---- variant 3: the march fed by a TMA bulk-copy ring (fused.cuh) ----
constexpr int kPipeWarpChoices[] = {8, 4, 3};  // compiled consumer-warp counts
size_t pipe_smem_bytes(int nw, int stages, bool corr);
int pipe_corr_stages(int nw, int stages);  // correlation launches stage 16 segments per row: as many slots as fit
int pipe_occupancy(int nw, size_t smem);
void launch_pipe(bool exact, int nw, int mode, const FusedCommon &fc, const FusedModel *models, dim3 grid, int stages,
                 size_t smem, cudaStream_t st);

// ---- the optional FP32 wavefield mode on the same ring (fused32.cuh), its default ----
size_t pipe32_smem_bytes(int nw, int stages);
void launch_pipe32(int nw, int mode, const FusedCommon &fc, const FusedModel32 *models, dim3 grid, int stages,
                   size_t smem, cudaStream_t st);

// ---- variant 4: two time steps per launch (pair.cuh) ----
size_t pair_smem(int nw, int stages);
void launch_pair(bool exact, int nw, int mode, const FusedCommon &fc, const FusedModel *models, dim3 grid, int stages,
                 size_t smem, cudaStream_t st);

// ---- resident whole-loop kernel for grids that fit on chip (resident.cuh) ----
size_t resident_smem_bytes(int ld, int rows);
int resident_max_ctas(size_t smem);  // co-resident CTAs on the current device (0: no cooperative launch / too large)
void launch_resident(bool exact, const ResidentArgs &a, dim3 grid, int threads, size_t smem, cudaStream_t st);

}  // namespace psv
