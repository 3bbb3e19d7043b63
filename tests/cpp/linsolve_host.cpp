// TEST INFRASTRUCTURE: csrc/sgx_linsolve.cuh compiled for the host, so the CPU test can
// compare it bit for bit with the reference's solve_Ax_b_wscr (Eigen) without a GPU.
#include "../../singularity-eos_b200/csrc/sgx_linsolve.cuh"
extern "C" void host_solve_Ax_b_batch(int n, double *a, double *b, int *ok, long batch) {
  for (long i = 0; i < batch; ++i)
    ok[i] = sgx::mix::solve_Ax_b<6>(n, a + i * n * n, b + i * n) ? 1 : 0;
}
