/* The reference's C++ tests re-expressed over the C ABI, in plain C (no pybind11, no PyTorch, no CUDA headers):
 *   tests/test_file_constructor.cpp:13-73   construct from the .ini file, run forward_simulate(shot, false, ...) twice,
 *                                           the sum of rtf_ux must be EXACTLY equal; write_receivers / write_sources;
 *   tests/test_copy_constructor.cpp:14-69   a copy of the object gives the identical sum.
 * This is how a C or C++ user of the reference's fdModel class switches to libpsvwave_b200.so.
 *
 *   gcc -std=c99 -I include examples/c_api_determinism.c -L psvwave_b200 -lpsvwave_b200 -Wl,-rpath,$PWD/psvwave_b200 -o demo
 *   ./demo configs/default_testing_configuration.ini [write]
 * prints "sum <value>" and exits 0 when every comparison holds. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "psvwave_b200.h"

#define CHECK(call)                                                          \
  do {                                                                       \
    int rc_ = (call);                                                        \
    if (rc_ != PSV_OK) {                                                     \
      fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, psv_last_error()); \
      return 2;                                                              \
    }                                                                        \
  } while (0)

static int sum_of_rtf_ux(psv_handle h, double *sum) {
  size_t n = 0;
  CHECK(psv_array_size(h, PSV_RTF_UX, &n));
  double *u = (double *)malloc(n * sizeof(double));
  if (!u) return 2;
  CHECK(psv_get_array(h, PSV_RTF_UX, u, n));
  double s = 0.0; /* the reference sums shot-major, receiver, time (tests/test_file_constructor.cpp:40-49) */
  for (size_t i = 0; i < n; ++i) s += u[i];
  free(u);
  *sum = s;
  return 0;
}

static int simulate_all_shots(psv_handle h) {
  long long n_shots = 0;
  CHECK(psv_get_int(h, PSV_INT_N_SHOTS, &n_shots));
  for (int s = 0; s < (int)n_shots; ++s) CHECK(psv_forward_simulate(h, s, /*store_fields=*/0, /*verbose=*/0, 0));
  return 0;
}

int main(int argc, char **argv) {
  if (argc < 2) {
    fprintf(stderr, "usage: %s configuration.ini [write]\n", argv[0]);
    return 2;
  }
  psv_handle model = NULL, copy = NULL;
  CHECK(psv_create_from_ini(argv[1], &model));
  double first = 0, second = 0, copied = 0;
  if (simulate_all_shots(model) || sum_of_rtf_ux(model, &first)) return 2;
  if (simulate_all_shots(model) || sum_of_rtf_ux(model, &second)) return 2;
  CHECK(psv_clone(model, &copy));
  if (simulate_all_shots(copy) || sum_of_rtf_ux(copy, &copied)) return 2;
  if (argc > 2 && strcmp(argv[2], "write") == 0) {
    CHECK(psv_write_receivers(model));
    CHECK(psv_write_sources(model));
  }
  CHECK(psv_destroy(copy));
  CHECK(psv_destroy(model));
  printf("sum %.17g\n", first);
  if (first != second) {
    fprintf(stderr, "two runs differ: %.17g vs %.17g\n", first, second);
    return 1;
  }
  if (first != copied) {
    fprintf(stderr, "the copy differs: %.17g vs %.17g\n", first, copied);
    return 1;
  }
  return 0;
}
