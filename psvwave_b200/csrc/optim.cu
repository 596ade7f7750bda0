// Device-resident smoothing and L-BFGS (see optim.hpp).
#include "optim.hpp"

#include <algorithm>
#include <cmath>
#include <mutex>
#include <stdexcept>

#include "model.hpp"

namespace psv {

// ---------------------------------------------------------------------------------------------
// Gaussian smoothing (scipy.ndimage.gaussian_filter, mode "reflect")
// ---------------------------------------------------------------------------------------------
std::vector<double> gaussian_weights(double sigma, int *radius) {
  // scipy/ndimage/_filters.py: lw = int(truncate * sd + 0.5); x = arange(-lw, lw + 1);
  // phi = exp(-0.5 / sd^2 * x^2); phi / phi.sum()
  const int lw = (int)(4.0 * sigma + 0.5);
  std::vector<double> w(2 * lw + 1);
  const double sigma2 = sigma * sigma;
  for (int i = -lw; i <= lw; ++i) w[i + lw] = std::exp(-0.5 / sigma2 * (double)(i * i));
  // numpy's pairwise sum of a short array: 8 running sums, combined as a tree, then the tail
  const size_t n = w.size();
  double total;
  if (n < 8) {
    total = 0.0;
    for (double v : w) total += v;
  } else {
    double r[8];
    for (int k = 0; k < 8; ++k) r[k] = w[k];
    size_t i = 8;
    for (; i + 8 <= n; i += 8)
      for (int k = 0; k < 8; ++k) r[k] += w[i + k];
    total = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
    for (; i < n; ++i) total += w[i];
  }
  for (double &v : w) v /= total;
  *radius = lw;
  return w;
}

namespace {
constexpr int kMaxRadius = 256;
__constant__ double c_gauss[2 * kMaxRadius + 1];

__device__ __forceinline__ int reflect(int i, int n) {  // d c b a | a b c d | d c b a
  const int period = 2 * n;
  i %= period;
  if (i < 0) i += period;
  return i < n ? i : period - 1 - i;
}

// one thread per output element; `stride` between consecutive samples of a line, `len` samples per line
__global__ void gauss_axis_kernel(const double *__restrict__ in, double *__restrict__ out, int panels, int n0, int n1,
                                  int axis, int radius) {
  const size_t total = (size_t)panels * n0 * n1;
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const size_t per = (size_t)n0 * n1;
  const size_t p = idx / per, r = idx - p * per;
  const int i0 = (int)(r / n1), i1 = (int)(r - (size_t)i0 * n1);
  const int len = axis == 0 ? n0 : n1, l = axis == 0 ? i0 : i1;
  const size_t stride = axis == 0 ? n1 : 1;
  const double *line = in + p * per + (axis == 0 ? (size_t)i1 : (size_t)i0 * n1);
  double acc = __dmul_rn(line[(size_t)l * stride], c_gauss[radius]);
  for (int ii = -radius; ii < 0; ++ii) {
    const double a = line[(size_t)reflect(l + ii, len) * stride], b = line[(size_t)reflect(l - ii, len) * stride];
    acc = __dadd_rn(acc, __dmul_rn(__dadd_rn(a, b), c_gauss[ii + radius]));
  }
  out[idx] = acc;
}

constexpr int kDotThreads = 256, kDotPerThread = 16;
__global__ void dot_partial_kernel(const double *__restrict__ a, const double *__restrict__ b, size_t n,
                                   double *__restrict__ partial) {
  __shared__ double sh[kDotThreads];
  const size_t base = ((size_t)blockIdx.x * kDotThreads + threadIdx.x) * kDotPerThread;
  double acc = 0.0;
  for (int j = 0; j < kDotPerThread; ++j)
    if (base + j < n) acc = fma(a[base + j], b[base + j], acc);
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = kDotThreads / 2; s > 0; s >>= 1) {
    if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = sh[0];
}
__global__ void maxabs_partial_kernel(const double *__restrict__ a, size_t n, double *__restrict__ partial) {
  __shared__ double sh[kDotThreads];
  const size_t base = ((size_t)blockIdx.x * kDotThreads + threadIdx.x) * kDotPerThread;
  double acc = 0.0;
  for (int j = 0; j < kDotPerThread; ++j)
    if (base + j < n) acc = fmax(acc, fabs(a[base + j]));
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = kDotThreads / 2; s > 0; s >>= 1) {
    if (threadIdx.x < s) sh[threadIdx.x] = fmax(sh[threadIdx.x], sh[threadIdx.x + s]);
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = sh[0];
}
__global__ void axpby_kernel(double alpha, const double *__restrict__ x, double beta, double *__restrict__ y, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = beta == 0.0 ? alpha * x[i] : fma(alpha, x[i], beta * y[i]);
}
__global__ void step_kernel(const double *__restrict__ x, double alpha, const double *__restrict__ d,
                            double *__restrict__ z, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) z[i] = fma(alpha, d[i], x[i]);
}
unsigned blocks(size_t n, int per) { return (unsigned)std::max<size_t>(1, (n + per - 1) / per); }
}  // namespace

constexpr int kGaussMutexDevices = 64;

void gaussian_smooth_panels(double *data, double *tmp, int panels, int n0, int n1, double sigma, cudaStream_t stream) {
  if (!(sigma > 0.0)) return;
  int radius = 0;
  const std::vector<double> w = gaussian_weights(sigma, &radius);
  if (radius > kMaxRadius) throw invalid_argument("Gaussian smoothing: sigma too large (radius > 256 samples).");
  if (radius == 0) return;  // a single weight of 1
  // c_gauss is ONE constant array per device: two models on the same GPU smoothing at once (different streams or host
  // threads, different sigma) would read each other's weights.  The call is synchronous anyway, so it holds the
  // device's lock from the upload to the end of the second pass.
  static std::mutex gauss_mutex[kGaussMutexDevices];
  int device = 0;
  PSV_CUDA(cudaGetDevice(&device));
  std::lock_guard<std::mutex> lock(gauss_mutex[device >= 0 && device < kGaussMutexDevices ? device : 0]);
  PSV_CUDA(cudaMemcpyToSymbolAsync(c_gauss, w.data(), w.size() * sizeof(double), 0, cudaMemcpyHostToDevice, stream));
  const size_t total = (size_t)panels * n0 * n1;
  gauss_axis_kernel<<<blocks(total, 256), 256, 0, stream>>>(data, tmp, panels, n0, n1, 0, radius);
  gauss_axis_kernel<<<blocks(total, 256), 256, 0, stream>>>(tmp, data, panels, n0, n1, 1, radius);
  PSV_CUDA(cudaGetLastError());
  PSV_CUDA(cudaStreamSynchronize(stream));  // `w` is pageable host memory
}

size_t dot_scratch_doubles(size_t n) { return blocks(n, kDotThreads * kDotPerThread); }

double device_dot(const double *a, const double *b, size_t n, double *scratch, cudaStream_t stream) {
  const unsigned nb = blocks(n, kDotThreads * kDotPerThread);
  dot_partial_kernel<<<nb, kDotThreads, 0, stream>>>(a, b, n, scratch);
  PSV_CUDA(cudaGetLastError());
  std::vector<double> part(nb);
  PSV_CUDA(cudaMemcpyAsync(part.data(), scratch, nb * sizeof(double), cudaMemcpyDeviceToHost, stream));
  PSV_CUDA(cudaStreamSynchronize(stream));
  double acc = 0.0;
  for (double v : part) acc += v;
  return acc;
}

double device_max_abs(const double *a, size_t n, double *scratch, cudaStream_t stream) {
  const unsigned nb = blocks(n, kDotThreads * kDotPerThread);
  maxabs_partial_kernel<<<nb, kDotThreads, 0, stream>>>(a, n, scratch);
  PSV_CUDA(cudaGetLastError());
  std::vector<double> part(nb);
  PSV_CUDA(cudaMemcpyAsync(part.data(), scratch, nb * sizeof(double), cudaMemcpyDeviceToHost, stream));
  PSV_CUDA(cudaStreamSynchronize(stream));
  double acc = 0.0;
  for (double v : part) acc = std::max(acc, v);
  return acc;
}

void device_axpby(double alpha, const double *x, double beta, double *y, size_t n, cudaStream_t stream) {
  axpby_kernel<<<blocks(n, 256), 256, 0, stream>>>(alpha, x, beta, y, n);
  PSV_CUDA(cudaGetLastError());
}
void device_step(const double *x, double alpha, const double *d, double *z, size_t n, cudaStream_t stream) {
  step_kernel<<<blocks(n, 256), 256, 0, stream>>>(x, alpha, d, z, n);
  PSV_CUDA(cudaGetLastError());
}

// ---------------------------------------------------------------------------------------------
// L-BFGS: two-loop recursion over a ring of (s, y) pairs in device memory, initial scaling s.y / y.y,
// backtracking line search on the sufficient-decrease condition starting from the unit step (what L-BFGS-B accepts
// on almost every iteration of the reference's notebook: 53 evaluations for 50 iterations), first step 1 / |g|
// like L-BFGS-B; pairs without positive curvature are dropped.
// ---------------------------------------------------------------------------------------------
LbfgsResult lbfgs_minimize(double *x, size_t n, const std::function<double(const double *, double *)> &evaluate,
                           const LbfgsOptions &opt, double *misfits, cudaStream_t stream) {
  if (opt.history < 1 || opt.max_iterations < 0) throw invalid_argument("L-BFGS: history >= 1, iterations >= 0.");
  const int m = opt.history;
  const size_t ns = dot_scratch_doubles(n);
  DeviceBuffer work((5 * n + 2 * (size_t)m * n + ns) * sizeof(double));
  double *g = work.as<double>(), *d = g + n, *xn = d + n, *gn = xn + n, *q = gn + n;
  double *S = q + n, *Y = S + (size_t)m * n, *scratch = Y + (size_t)m * n;
  std::vector<double> rho(m, 0.0), alpha(m, 0.0);
  int stored = 0, head = 0;  // pairs held; slot of the next pair
  LbfgsResult res;
  double f = evaluate(x, g);
  ++res.evaluations;
  if (misfits) misfits[0] = f;
  for (int k = 0; k < opt.max_iterations; ++k) {
    res.gradient_max = device_max_abs(g, n, scratch, stream);
    if (res.gradient_max <= opt.gtol) break;
    // d = -H g
    PSV_CUDA(cudaMemcpyAsync(q, g, n * sizeof(double), cudaMemcpyDeviceToDevice, stream));
    for (int i = 0; i < stored; ++i) {  // newest pair first
      const int j = (head - 1 - i + 2 * m) % m;
      alpha[j] = rho[j] * device_dot(S + (size_t)j * n, q, n, scratch, stream);
      device_axpby(-alpha[j], Y + (size_t)j * n, 1.0, q, n, stream);
    }
    double gamma = 1.0;
    if (stored > 0) {
      const int j = (head - 1 + m) % m;
      gamma = 1.0 / (rho[j] * device_dot(Y + (size_t)j * n, Y + (size_t)j * n, n, scratch, stream));
    }
    device_axpby(gamma, q, 0.0, d, n, stream);
    for (int i = stored - 1; i >= 0; --i) {  // oldest pair first
      const int j = (head - 1 - i + 2 * m) % m;
      const double beta = rho[j] * device_dot(Y + (size_t)j * n, d, n, scratch, stream);
      device_axpby(alpha[j] - beta, S + (size_t)j * n, 1.0, d, n, stream);
    }
    device_axpby(0.0, d, -1.0, d, n, stream);  // d <- -d
    double dg = device_dot(d, g, n, scratch, stream);
    if (!(dg < 0.0)) {  // not a descent direction: restart from steepest descent
      stored = 0;
      device_axpby(-1.0, g, 0.0, d, n, stream);
      dg = -device_dot(g, g, n, scratch, stream);
    }
    double step = 1.0;
    if (stored == 0) step = std::min(1.0, 1.0 / std::sqrt(-dg));  // L-BFGS-B's first step: 1 / |g|
    double fn = f;
    bool ok = false;
    for (int ls = 0; ls < opt.max_line_search; ++ls) {
      device_step(x, step, d, xn, n, stream);
      fn = evaluate(xn, gn);
      ++res.evaluations;
      if (std::isfinite(fn) && fn <= f + opt.c1 * step * dg) {
        ok = true;
        break;
      }
      // minimiser of the parabola through f, dg and fn, kept inside [0.1, 0.5] of the step
      double t = std::isfinite(fn) ? -dg * step * step / (2.0 * (fn - f - dg * step)) : 0.1 * step;
      step = std::min(0.5 * step, std::max(0.1 * step, t));
    }
    if (!ok) {  // no decrease along a descent direction: the gradient is as good as it gets
      // the last evaluation was a REJECTED trial point: evaluate x once more so that everything the evaluation leaves
      // behind in the caller's object (synthetics, sensitivity kernels, misfit) belongs to the point that is returned
      evaluate(x, gn);
      ++res.evaluations;
      break;
    }
    // s = xn - x, y = gn - g
    double *s = S + (size_t)head * n, *y = Y + (size_t)head * n;
    PSV_CUDA(cudaMemcpyAsync(s, xn, n * sizeof(double), cudaMemcpyDeviceToDevice, stream));
    device_axpby(-1.0, x, 1.0, s, n, stream);
    PSV_CUDA(cudaMemcpyAsync(y, gn, n * sizeof(double), cudaMemcpyDeviceToDevice, stream));
    device_axpby(-1.0, g, 1.0, y, n, stream);
    const double sy = device_dot(s, y, n, scratch, stream), yy = device_dot(y, y, n, scratch, stream);
    if (sy > 2.2e-16 * yy) {  // L-BFGS-B's curvature test (epsmch * |y|^2)
      rho[head] = 1.0 / sy;
      head = (head + 1) % m;
      stored = std::min(stored + 1, m);
    }
    PSV_CUDA(cudaMemcpyAsync(x, xn, n * sizeof(double), cudaMemcpyDeviceToDevice, stream));
    PSV_CUDA(cudaMemcpyAsync(g, gn, n * sizeof(double), cudaMemcpyDeviceToDevice, stream));
    f = fn;
    ++res.iterations;
    if (misfits) misfits[res.iterations] = f;
  }
  PSV_CUDA(cudaStreamSynchronize(stream));
  res.misfit = f;
  res.gradient_max = device_max_abs(g, n, scratch, stream);
  return res;
}

}  // namespace psv
