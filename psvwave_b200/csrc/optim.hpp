// Device-resident outer loop of the FWI: Gaussian smoothing of the gradient vector and L-BFGS on the
// model vector, both in device memory.  The reference hands get_model_vector / get_gradient_vector to SciPy
// (docs-source/source/typical-workflow.rst:16-72; Notebook 2 cells 15-17: scipy.optimize.minimize(method="L-BFGS-B",
// maxcor=10) with scipy.ndimage.gaussian_filter on the three (nz_free, nx_free) gradient panels); here the vectors
// never leave the GPU and only scalars (misfit, dot products) reach the host.
#pragma once

#include <cuda_runtime.h>

#include <functional>
#include <vector>

namespace psv {

// scipy.ndimage.gaussian_filter(panel, sigma) for `panels` arrays of shape (n0, n1) stored back to back (C order):
// separable, axis 0 then axis 1, mode "reflect", truncate 4.0 -> radius int(4 sigma + 0.5); the correlation adds the
// centre term first, then the symmetric pairs from the far end inwards (scipy's correlate1d).  `tmp` has the size of
// `data`.  Stream-ordered.
void gaussian_smooth_panels(double *data, double *tmp, int panels, int n0, int n1, double sigma, cudaStream_t stream);
std::vector<double> gaussian_weights(double sigma, int *radius);  // host: scipy's _gaussian_kernel1d, order 0

// Deterministic dot product of two device vectors (fixed-shape partial sums on the device, added on the host).
double device_dot(const double *a, const double *b, size_t n, double *scratch /* >= dot_scratch_doubles(n) */,
                  cudaStream_t stream);
size_t dot_scratch_doubles(size_t n);
double device_max_abs(const double *a, size_t n, double *scratch, cudaStream_t stream);
// y <- alpha * x + beta * y   /   z <- x + alpha * d
void device_axpby(double alpha, const double *x, double beta, double *y, size_t n, cudaStream_t stream);
void device_step(const double *x, double alpha, const double *d, double *z, size_t n, cudaStream_t stream);

struct LbfgsOptions {
  int max_iterations = 50;
  int history = 10;      // maxcor of the reference's notebook
  double gtol = 1e-6;    // stop when max |g| <= gtol (L-BFGS-B's projected-gradient test without bounds)
  double c1 = 1e-4;      // sufficient-decrease constant
  int max_line_search = 20;
};
struct LbfgsResult {
  int iterations = 0, evaluations = 0;
  double misfit = 0.0, gradient_max = 0.0;
};
// evaluate(x, g): sets g to the gradient at x (both device vectors of length n) and returns the misfit.
// x holds the starting model on entry and the final one on return; misfits[0..iterations] (may be null).
LbfgsResult lbfgs_minimize(double *x, size_t n, const std::function<double(const double *, double *)> &evaluate,
                           const LbfgsOptions &opt, double *misfits, cudaStream_t stream);

}  // namespace psv
