// Resident whole-loop kernel (resident.cuh): instantiations + cooperative launch.
#include "launchers.hpp"
#include "model.hpp"
#include "resident.cuh"

namespace psv {

size_t resident_smem_bytes(int ld, int rows) { return ResidentLayout(ld, rows).bytes(); }

// CTAs of the loop kernel that can be co-resident on the current device with `smem` bytes of tiles each.
int resident_max_ctas(size_t smem) {
  int dev = 0, sms = 0, per_sm = 0, coop = 0;
  PSV_CUDA(cudaGetDevice(&dev));
  PSV_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
  if (!coop || smem > kMaxDynamicSmem) return 0;
  PSV_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  PSV_CUDA(cudaFuncSetAttribute(resident_loop_kernel<FastMath>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  PSV_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, resident_loop_kernel<FastMath>, kResidentThreads,
                                                         smem));
  return per_sm * sms;
}

namespace {
template <class M>
void launch_m(const ResidentArgs &a, dim3 grid, int threads, size_t smem, cudaStream_t st) {
  static SmemOptIn opt_in;  // per instantiation
  opt_in.ensure(resident_loop_kernel<M>, smem);
  void *args[] = {const_cast<ResidentArgs *>(&a)};
  // cooperative launch: all CTAs are guaranteed co-resident (they wait for each other inside the loop)
  PSV_CUDA(cudaLaunchCooperativeKernel((const void *)resident_loop_kernel<M>, grid, dim3(threads), args, smem,
                                       st));
}
}  // namespace

void launch_resident(bool exact, const ResidentArgs &a, dim3 grid, int threads, size_t smem, cudaStream_t st) {
  if (exact) launch_m<ExactMath>(a, grid, threads, smem, st);
  else launch_m<FastMath>(a, grid, threads, smem, st);
}

}  // namespace psv
