// DEBUG HARNESS (not product): one block of cooperative T1 compositions with progress marks in mapped host memory.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DCG_DEBUG -Icofhe_b200/csrc -Iinclude tools/coop_one.cu -o tools/_coop_one
//   tools/_coop_one tests/golden/coop_case_t1.bin      (case.bin: u32 dbits, nc, n, then f1, f2, expected as SoA [2N+nc][n] limbs)
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "coop_launch_impl.cuh"
using namespace cofhe;
int main(int argc, char **argv) {
  FILE *f = fopen(argv[1], "rb");
  unsigned hdr[3];
  if (fread(hdr, 4, 3, f) != 3) return 2;
  int dbits = hdr[0], nc = hdr[1], n = hdr[2];
  constexpr int N = 33;
  size_t w = (size_t)(2 * N + nc) * n;
  std::vector<unsigned> h1(w), h2(w), hw(w), ho(w);
  if (fread(h1.data(), 4, w, f) != w || fread(h2.data(), 4, w, f) != w || fread(hw.data(), 4, w, f) != w) return 2;
  int *hd; cudaHostAlloc(&hd, 256, cudaHostAllocMapped); for (int i = 0; i < 64; i++) hd[i] = -1;
  int *dd; cudaHostGetDevicePointer(&dd, hd, 0);
  cudaMemcpyToSymbol(cg_dbg, &dd, sizeof(dd));
  unsigned *d1, *d2, *d3, *bl; int *nb;
  cudaMalloc(&d1, w * 4); cudaMalloc(&d2, w * 4); cudaMalloc(&d3, w * 4); cudaMalloc(&bl, 4096); cudaMalloc(&nb, 8); cudaMemset(nb, 0, 8);
  cudaMemcpy(d1, h1.data(), w * 4, cudaMemcpyHostToDevice); cudaMemcpy(d2, h2.data(), w * 4, cudaMemcpyHostToDevice);
  k_coop_compose<4, 9, N, 136, false, 3><<<(n + 31) / 32, kCoopBlock>>>(dbits, nc, n, d1, d2, d3, 0, bl, nb);
  cudaError_t e = cudaDeviceSynchronize();
  int hnb[2] = {-1, -1};
  int bad = -1;
  if (e == cudaSuccess) {
    cudaMemcpy(hnb, nb, 8, cudaMemcpyDeviceToHost); cudaMemcpy(ho.data(), d3, w * 4, cudaMemcpyDeviceToHost);
    bad = 0; for (size_t i = 0; i < w; i++) bad += ho[i] != hw[i];
  }
  printf("sync: %s ; bailed %d ; wrong words %d ; marks per warp", cudaGetErrorString(e), hnb[0], bad);
  for (int i = 0; i < 52; i++) printf(" %d", hd[i]);
  printf("\n");
  return 0;
}
