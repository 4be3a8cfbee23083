#include <cstdio>
#include <cstdint>
__global__ void k(uint32_t a, uint32_t b, int *o) { bool hi, lo; uint32_t m = __vibmax_s16x2(a, b, &hi, &lo); o[0] = m; o[1] = hi; o[2] = lo; }
int main() { int *d; cudaMalloc(&d, 12); int h[3];
  uint32_t A[3] = {0x00050001u, 0x00010005u, 0x00030003u}, B[3] = {0x00020002u, 0x00020002u, 0x00030003u};
  for (int i = 0; i < 3; ++i) { k<<<1,1>>>(A[i], B[i], d); cudaMemcpy(h, d, 12, cudaMemcpyDeviceToHost); printf("a=%08x b=%08x max=%08x pred_hi=%d pred_lo=%d\n", A[i], B[i], h[0], h[1], h[2]); }
  return 0; }
