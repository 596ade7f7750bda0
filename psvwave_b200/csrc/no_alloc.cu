// psv_l2_misfit_static_scratch: the L2 misfit without a single device allocation.
//
// Who this is for.  Model::misfit_of_range (model.cu) keeps its block partial sums in a buffer that it grows with
// cudaMalloc.  Some users of this library cannot have that happen underneath them:
//   * a host application that hands every byte of GPU memory to its own arena at start-up and forbids
//     plug-ins to go to the driver behind its back (the arena's budget would be wrong, or the driver has nothing
//     left to give);
//   * a sampler thread that evaluates one misfit per proposal inside a loop that must never stall:
//     cudaMalloc / cudaFree synchronise the whole device, so one call freezes every other stream of the process.
// Such a caller already owns the four seismogram arrays on the device (its own allocator made them) and owns the
// double the result goes to; the only thing left that needed memory was the scratch for the partial sums.  Here that
// scratch is a fixed array in the library image itself.
//
// What that costs.  The array is part of the module: EVERY GPU on which this library's kernels get loaded carries
// kMaxPartials * 8 B = 64 KiB of it until the process exits, used or not.  It is one array per device, shared
// by all callers: two calls whose kernels overlap on the same GPU (different streams, different host threads) write
// the same slots and both results are garbage.  Order the calls on one stream, or serialise them yourself.
//
// Stale contents.  The array keeps whatever the previous call left in it.  Nothing in it has to start at zero:
// a call with nb blocks writes slots [0, nb) in its first kernel and reads exactly those slots in its second, so
// older values in [nb, kMaxPartials) are never looked at.  (tests/test_zz_no_alloc.py runs a long call and then a
// short one to prove it.)
//
// Same numbers as the allocating path: the first kernel is the arithmetic of misfit_partial_kernel
// (pointwise.cuh: 256 threads x 16 consecutive samples, serial per thread, fixed tree per block), the second adds
// the block partials serially in block order, which is what the host loop of Model::misfit_of_range does.  Same
// terms, same association, so the result is bit-identical to psv_calculate_l2_misfit on the same arrays.
#include <cstdio>
#include <mutex>

#include <cuda_runtime.h>

#include "psvwave_b200.h"

namespace psv {
void set_last_error(const char *message);  // capi.cu: the text psv_last_error() returns on this thread
}

namespace {

constexpr int kThreads = 256;    // = kMisfitThreads   (pointwise.cuh) -- must stay equal for bit-identity
constexpr int kPerThread = 16;   // = kMisfitPerThread (pointwise.cuh)
constexpr size_t kMaxPartials = PSV_MISFIT_STATIC_MAX_SAMPLES / ((size_t)kThreads * kPerThread);
static_assert(kMaxPartials * kThreads * kPerThread == PSV_MISFIT_STATIC_MAX_SAMPLES, "limit must be whole blocks");
static_assert(kMaxPartials * sizeof(double) == 64 * 1024, "the documented footprint is 64 KiB per GPU");

// The scratch itself.  In HOST code this name is only a stand-in the runtime uses to look the symbol up; it is not a
// device address and must never be handed to a kernel as one.  Host code below asks cudaGetSymbolAddress for the
// real address, once per device.
static __device__ __align__(256) double g_partials[kMaxPartials];

__global__ void misfit_static_partial_kernel(const double *ux, const double *uz, const double *ox, const double *oz,
                                             size_t n, double half_dt, double *partial) {
  __shared__ double sh[kThreads];
  const size_t base = ((size_t)blockIdx.x * kThreads + threadIdx.x) * kPerThread;
  double acc = 0.0;
  for (int j = 0; j < kPerThread; ++j) {
    const size_t i = base + j;
    if (i < n) {
      const double dx = __dsub_rn(ox[i], ux[i]), dz = __dsub_rn(oz[i], uz[i]);
      acc = __dadd_rn(acc, __dmul_rn(half_dt, __dmul_rn(dx, dx)));
      acc = __dadd_rn(acc, __dmul_rn(half_dt, __dmul_rn(dz, dz)));
    }
  }
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = kThreads / 2; s > 0; s >>= 1) {
    if (threadIdx.x < s) sh[threadIdx.x] = __dadd_rn(sh[threadIdx.x], sh[threadIdx.x + s]);
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = sh[0];
}

// Block partials added one after the other in block order (the order of the host loop in Model::misfit_of_range).
// One warp stages 32 partials at a time with coalesced loads; lane 0 does the ordered additions.
__global__ void misfit_static_finish_kernel(const double *partial, unsigned nb, double *out) {
  __shared__ double stage[32];
  double acc = 0.0;
  for (unsigned b0 = 0; b0 < nb; b0 += 32) {
    const unsigned b = b0 + threadIdx.x;
    stage[threadIdx.x] = b < nb ? partial[b] : 0.0;
    __syncwarp();
    if (threadIdx.x == 0) {
      const unsigned m = nb - b0 < 32u ? nb - b0 : 32u;
      for (unsigned k = 0; k < m; ++k) acc = __dadd_rn(acc, stage[k]);
    }
    __syncwarp();
  }
  if (threadIdx.x == 0) *out = acc;
}

// Device address of g_partials, looked up once per device with that device current.
constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;
double *g_partials_address[kMaxDevices] = {};

int fail(int code, const char *text) {
  psv::set_last_error(text);
  return code;
}

int fail_cuda(const char *what, cudaError_t e) {
  char text[256];
  std::snprintf(text, sizeof text, "psv_l2_misfit_static_scratch: %s: %s", what, cudaGetErrorString(e));
  cudaGetLastError();
  return fail(PSV_ERR_CUDA, text);
}

}  // namespace

extern "C" int psv_l2_misfit_static_scratch(const double *device_ux, const double *device_uz,
                                            const double *device_obs_ux, const double *device_obs_uz, size_t n,
                                            double dt, double *device_misfit_out, void *stream) {
  if (!device_misfit_out) return fail(PSV_ERR_INVALID_ARGUMENT, "psv_l2_misfit_static_scratch: null result pointer.");
  if (n > 0 && (!device_ux || !device_uz || !device_obs_ux || !device_obs_uz))
    return fail(PSV_ERR_INVALID_ARGUMENT, "psv_l2_misfit_static_scratch: null seismogram pointer.");
  if (n > (size_t)PSV_MISFIT_STATIC_MAX_SAMPLES) {
    char text[256];
    std::snprintf(text, sizeof text,
                  "psv_l2_misfit_static_scratch: %zu samples per array exceed the fixed scratch "
                  "(PSV_MISFIT_STATIC_MAX_SAMPLES = %zu); split the range or use psv_calculate_l2_misfit.",
                  n, (size_t)PSV_MISFIT_STATIC_MAX_SAMPLES);
    return fail(PSV_ERR_LENGTH, text);
  }

  int device = -1;
  cudaError_t e = cudaGetDevice(&device);
  if (e != cudaSuccess) return fail_cuda("cudaGetDevice", e);
  if (device < 0 || device >= kMaxDevices)
    return fail(PSV_ERR_INVALID_ARGUMENT, "psv_l2_misfit_static_scratch: device index beyond the 64 this entry point tracks.");

  // Set-up on the first call per device: resolve the symbol to this device's copy of the array.  The look-up may make
  // the runtime load the module onto the device, which is why that first call belongs before any stream capture.
  double *scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_partials_address[device]) {
      double *p = nullptr;
      e = cudaGetSymbolAddress((void **)&p, g_partials);
      if (e != cudaSuccess) return fail_cuda("cudaGetSymbolAddress(scratch)", e);
      if (!p) return fail(PSV_ERR_CUDA, "psv_l2_misfit_static_scratch: the scratch symbol resolved to a null address.");
      g_partials_address[device] = p;
    }
    scratch = g_partials_address[device];
  }

  const size_t per_block = (size_t)kThreads * kPerThread;
  const unsigned nb = (unsigned)(n == 0 ? 1 : (n + per_block - 1) / per_block);  // n = 0 -> one idle block -> 0.0
  cudaStream_t s = (cudaStream_t)stream;
  misfit_static_partial_kernel<<<nb, kThreads, 0, s>>>(device_ux, device_uz, device_obs_ux, device_obs_uz, n, 0.5 * dt,
                                                       scratch);
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail_cuda("launch of the partial-sum kernel", e);
  misfit_static_finish_kernel<<<1, 32, 0, s>>>(scratch, nb, device_misfit_out);
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail_cuda("launch of the finishing kernel", e);
  return PSV_OK;
}
