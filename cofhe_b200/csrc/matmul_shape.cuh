// cofhe_b200 -- shape of a batched multi-exponentiation (matmul / fixed base): shared by host code and kernels.
#pragma once
#include <stddef.h>

#include "fw.cuh"

namespace cofhe {

struct MMShape {
  unsigned n, m, p;     // ct is n x m, scalars m x p
  int nwin, W;          // windows per scalar, window width
  int el;               // limbs per scalar
  unsigned chunks, jper;  // j range split: chunks * jper >= m
  int block;              // threads per block of the accumulate kernel (0: chosen at launch)
};

QF_HD size_t mm_tab_index(const MMShape &s, unsigned y, size_t base, int win, int d) {
  return ((((size_t)y * s.n * s.m + base) * s.nwin + win) * ((1u << s.W) - 1) + (d - 1));
}

}  // namespace cofhe
