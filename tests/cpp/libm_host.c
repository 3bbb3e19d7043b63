/* TEST INFRASTRUCTURE: the host libm (the one the reference links) over arrays, so the GPU
 * test can compare the device routines of csrc/sgx_libm.cuh against it bit for bit.
 * Built by tests/test_libm_parity.py with -fno-builtin so every call reaches libm.so. */
#include <gnu/libc-version.h>
#include <math.h>
const char *host_libm_version(void) { return gnu_get_libc_version(); }
void host_libm_eval(int which, const double *x, const double *y, double *out, long n) {
  for (long i = 0; i < n; ++i) {
    switch (which) {
    case 0: out[i] = exp(x[i]); break;
    case 1: out[i] = log(x[i]); break;
    case 2: out[i] = log10(x[i]); break;
    case 3: out[i] = cbrt(x[i]); break;
    default: out[i] = pow(x[i], y[i]); break;
    }
  }
}
