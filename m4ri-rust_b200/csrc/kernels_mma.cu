// kernels_mma.cu — path (b): bits -> int8, tcgen05 MMA, parity. (under construction)
#include "common.cuh"
namespace m4b {
bool mma_available() { return false; }
bool mma_shape_ok(const DMat *, const DMat *, const DMat *) { return false; }
void k_mul_abt_mma(DMat *, const DMat *, const DMat *, int, cudaStream_t) {
  die("m4ri_b200: the tcgen05 MMA path is not built into this library yet\n");
}
}  // namespace m4b
