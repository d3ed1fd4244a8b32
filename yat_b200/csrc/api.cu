// Library-level pieces of the C-ABI: version, error text, host-side bucket assignment.
#include <cmath>
#include <cstdarg>
#include <cstdio>

#include "common.cuh"

namespace yat {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_launch(const char* what) {
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: CUDA error %d (%s)", what, (int)e, cudaGetErrorString(e));
    return YAT_ERR_CUDA;
  }
  return YAT_OK;
}

}  // namespace yat

extern "C" int yat_version(void) { return 100; }  // 0.1.0

extern "C" const char* yat_last_error(void) { return yat::g_err; }

// common/trainer.py:109-120, common/bucket_sampler.py:314-322: first key (table order) with the
// strictly smallest |key - ratio| in float64; -1 when the table is empty.
extern "C" int yat_bucket_assign(const double* ratios, int n, const double* keys, int nkeys, int32_t* out) {
  YAT_REQUIRE(n >= 0 && nkeys >= 0 && (n == 0 || (ratios && out)) && (nkeys == 0 || keys), YAT_ERR_BAD_ARG,
              "yat_bucket_assign: bad arguments");
  for (int i = 0; i < n; ++i) {
    double best = INFINITY;
    int32_t arg = -1;
    for (int j = 0; j < nkeys; ++j) {
      const double d = std::fabs(keys[j] - ratios[i]);
      if (d < best) {
        best = d;
        arg = j;
      }
    }
    out[i] = arg;
  }
  return YAT_OK;
}
