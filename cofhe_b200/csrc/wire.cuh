// cofhe_b200 -- the reference's binary tensor wire format <-> SoA limb batches, on the device.
//
// Format (CPUCryptoSystem::serialize_ciphertext_tensor / deserialize_ciphertext_tensor, reference
// include/x86_64/cpu_cryptosystem.inl:322-520): u32 ndim | u32 dims[ndim] | u64 table[6 n] | magnitudes.
// table[6 i + j] = byte offset of integer j of ciphertext i (order c1.a c1.b c1.c c2.a c2.b c2.c) inside the
// magnitude area, bit 63 set when the integer is <= 0 (the reference tests sgn() != 1, so zero carries the flag too).
// An integer x owns mpz_sizeinbase(x, 2) / 8 + 1 bytes (bits(0) counts as 1), little endian, zero padded; on reading,
// its length is the distance to the next offset (the last one runs to the end of the buffer).
//
// One thread per integer slot.  Slot t of the grid is (j = t / n, i = t % n): the threads of a warp handle the same
// coefficient of 32 consecutive ciphertexts, so their limb rows in the SoA batch are coalesced 128-byte lines; the
// wire side is byte-granular and unaligned by construction (offsets are running sums of odd sizes) and goes through
// L1/L2 as byte accesses -- this path is bound by the byte traffic of the wire buffer, a few MB per tensor.
#pragma once
#include "fw.cuh"

namespace cofhe {

// The three tensor formats of the reference share this layout and differ in the integers per element:
//   kind 0  ciphertext tensor   6 per element: c1.a c1.b c1.c c2.a c2.b c2.c   (cpu_cryptosystem.inl:322-520)
//   kind 1  part-decryption-result tensor (forms)   3 per element: a b c      (cpu_cryptosystem.inl:510-635)
//   kind 2  plaintext tensor    1 per element, any sign                        (cpu_cryptosystem.inl:229-318)
// SoA side: ciphertext batch [2][2na+nc][n]; form batch [2na+nc][n]; plaintexts as magnitudes [ml][n] plus one sign byte
// per element (plaintexts are not reduced modulo 2^k by the reference: sums and products of plaintexts travel too).
struct WireShape {
  int na, nc;       // limbs of a / b and of c in the SoA batch
  size_t n;         // elements
  int kind;         // 0 ciphertexts, 1 forms, 2 plaintexts
  int ml;           // limbs per plaintext magnitude (kind 2)
};
QF_HD int wire_slots(const WireShape &s) { return s.kind == 0 ? 6 : (s.kind == 1 ? 3 : 1); }

// (row offset inside the batch, limb count, two's complement?) of integer j of an element
QF_HD void wire_slot(const WireShape &s, int j, size_t *row, int *limbs, bool *is_b) {
  if (s.kind == 2) { *row = 0; *limbs = s.ml; *is_b = false; return; }
  int f = j / 3, c = j % 3, fl = 2 * s.na + s.nc;
  *row = (size_t)f * fl + (c == 0 ? 0 : c == 1 ? s.na : 2 * s.na);
  *limbs = c == 2 ? s.nc : s.na;
  *is_b = c == 1;
}

#if defined(__CUDACC__)
constexpr int kWireBlock = 256;

// 1. per slot (wire order k = 6 i + j): reserved size and sign flag
__global__ void __launch_bounds__(kWireBlock) k_wire_sizes(WireShape s, const u32 *ct, const unsigned char *sgn, u32 *sizes, unsigned char *flags) {
  const int S = wire_slots(s);
  size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= S * s.n) return;
  int j = (int)(t / s.n);
  size_t i = t % s.n, row;
  int nl;
  bool is_b;
  wire_slot(s, j, &row, &nl, &is_b);
  const u32 *p = ct + row * s.n + i;
  bool neg = is_b && (p[(size_t)(nl - 1) * s.n] >> 31);
  // bit length of the magnitude; for negative b: |b| = ~b + 1
  int bits = 0;
  u32 carry = 1;
  bool nonzero = false;
  for (int l = 0; l < nl; l++) {
    u32 w = p[(size_t)l * s.n];
    if (neg) { u32 v = ~w + carry; carry = (carry && v == 0) ? 1u : 0u; w = v; }
    if (w) { bits = 32 * l + 32 - __clz(w); nonzero = true; }
  }
  size_t k = (size_t)S * i + j;
  sizes[k] = (u32)((bits > 0 ? bits : 1) / 8 + 1);
  flags[k] = (neg || !nonzero || (s.kind == 2 && sgn && sgn[i])) ? 1 : 0;
}

// 2. exclusive prefix sum of u32 sizes into u64 offsets: per-block scan, scan of the block totals, add.
constexpr int kScanBlock = 1024;
__global__ void __launch_bounds__(kScanBlock) k_scan_blocks(size_t m, const u32 *in, unsigned long long *out, unsigned long long *totals) {
  __shared__ unsigned long long sh[kScanBlock];
  size_t t = (size_t)blockIdx.x * kScanBlock + threadIdx.x;
  unsigned long long v = t < m ? in[t] : 0;
  sh[threadIdx.x] = v;
  __syncthreads();
  for (int d = 1; d < kScanBlock; d <<= 1) {
    unsigned long long add = threadIdx.x >= d ? sh[threadIdx.x - d] : 0;
    __syncthreads();
    sh[threadIdx.x] += add;
    __syncthreads();
  }
  if (t < m) out[t] = sh[threadIdx.x] - v;  // exclusive inside the block
  if (threadIdx.x == kScanBlock - 1) totals[blockIdx.x] = sh[kScanBlock - 1];
}
// one block: exclusive scan of `nb` block totals in place (sequential over chunks of 1024), grand total into *sum
__global__ void __launch_bounds__(kScanBlock) k_scan_totals(size_t nb, unsigned long long *totals, unsigned long long *sum) {
  __shared__ unsigned long long sh[kScanBlock];
  __shared__ unsigned long long base;
  if (threadIdx.x == 0) base = 0;
  __syncthreads();
  for (size_t c0 = 0; c0 < nb; c0 += kScanBlock) {
    size_t t = c0 + threadIdx.x;
    unsigned long long v = t < nb ? totals[t] : 0;
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int d = 1; d < kScanBlock; d <<= 1) {
      unsigned long long add = threadIdx.x >= d ? sh[threadIdx.x - d] : 0;
      __syncthreads();
      sh[threadIdx.x] += add;
      __syncthreads();
    }
    if (t < nb) totals[t] = base + sh[threadIdx.x] - v;
    __syncthreads();
    if (threadIdx.x == 0) base += sh[kScanBlock - 1];
    __syncthreads();
  }
  if (threadIdx.x == 0) *sum = base;
}

// 3. table entry + magnitude bytes of every slot.  body = wire + header + 48 n.
__global__ void __launch_bounds__(kWireBlock) k_wire_write(WireShape s, const u32 *ct, const u32 *sizes, const unsigned char *flags,
                                                           const unsigned long long *offs, const unsigned long long *totals,
                                                           unsigned char *table, unsigned char *body) {
  const int S = wire_slots(s);
  size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= S * s.n) return;
  int j = (int)(t / s.n);
  size_t i = t % s.n, row;
  int nl;
  bool is_b;
  wire_slot(s, j, &row, &nl, &is_b);
  size_t k = (size_t)S * i + j;
  unsigned long long off = offs[k] + totals[k / kScanBlock];
  unsigned long long entry = off | ((unsigned long long)flags[k] << 63);
  for (int b = 0; b < 8; b++) table[8 * k + b] = (unsigned char)(entry >> (8 * b));
  const u32 *p = ct + row * s.n + i;
  bool neg = is_b && (p[(size_t)(nl - 1) * s.n] >> 31);
  unsigned char *dst = body + off;
  u32 sz = sizes[k], carry = 1;
  for (int l = 0; l < nl; l++) {
    u32 w = p[(size_t)l * s.n];
    if (neg) { u32 v = ~w + carry; carry = (carry && v == 0) ? 1u : 0u; w = v; }
    for (int b = 0; b < 4; b++) {
      u32 pos = 4 * l + b;
      if (pos < sz) dst[pos] = (unsigned char)(w >> (8 * b));
    }
  }
  for (u32 pos = 4 * nl; pos < sz; pos++) dst[pos] = 0;  // the padding byte of a full-width integer
}

// wire -> SoA.  table/body as above, body_len = bytes of the magnitude area.  err: 1 = an integer does not fit its limbs
__global__ void __launch_bounds__(kWireBlock) k_wire_read(WireShape s, const unsigned char *table, const unsigned char *body,
                                                          unsigned long long body_len, u32 *ct, unsigned char *sgn, int *err) {
  const int S = wire_slots(s);
  size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= S * s.n) return;
  int j = (int)(t / s.n);
  size_t i = t % s.n, row;
  int nl;
  bool is_b;
  wire_slot(s, j, &row, &nl, &is_b);
  size_t k = (size_t)S * i + j;
  const size_t last = (size_t)S * s.n;
  unsigned long long e0 = 0, e1 = 0;
  for (int b = 0; b < 8; b++) e0 |= (unsigned long long)table[8 * k + b] << (8 * b);
  if (k + 1 < last) { for (int b = 0; b < 8; b++) e1 |= (unsigned long long)table[8 * (k + 1) + b] << (8 * b); }
  const unsigned long long mask = ~(1ull << 63);
  unsigned long long lo = e0 & mask, hi = (k + 1 < last) ? (e1 & mask) : body_len;
  bool flag = (e0 >> 63) != 0, bad = hi < lo || hi > body_len;
  unsigned long long len = bad ? 0 : hi - lo;
  const unsigned char *src = body + lo;
  u32 *p = ct + row * s.n + i;
  u32 carry = 1, top = 0;
  bool neg = flag && is_b;
  for (int l = 0; l < nl; l++) {
    u32 w = 0;
    for (int b = 0; b < 4; b++) {
      unsigned long long pos = 4ull * l + b;
      if (pos < len) w |= (u32)src[pos] << (8 * b);
    }
    if (neg) { u32 v = ~w + carry; carry = (carry && v == 0) ? 1u : 0u; w = v; }
    p[(size_t)l * s.n] = w;
    top = w;
  }
  // anything beyond the limbs must be zero padding; a signed b must keep its sign bit consistent
  for (unsigned long long pos = 4ull * nl; pos < len; pos++) bad = bad || src[pos] != 0;
  if (is_b && !bad) {
    bool msb = top >> 31;
    // positive b: sign bit clear; negative b (flag and non-zero magnitude): sign bit set
    bool zero_mag = carry != 0 && neg;  // ~0 + 1 rippled through every limb: the magnitude was 0
    if (!neg && msb) bad = true;
    if (neg && !zero_mag && !msb) bad = true;
  } else if (s.kind == 2) {
    // plaintexts are sign-magnitude on the SoA side: the flag with a non-zero magnitude is a negative value
    bool any = false;
    for (int l = 0; l < nl; l++) any = any || p[(size_t)l * s.n] != 0;
    if (sgn) sgn[i] = (flag && any) ? 1 : 0;
    else if (flag && any) bad = true;  // the caller has no room for signs
  } else if (!is_b && flag && !bad) {
    // a and c are positive in every valid form; the flag can only mean zero
    bool any = false;
    for (int l = 0; l < nl; l++) any = any || p[(size_t)l * s.n] != 0;
    if (any) bad = true;
  }
  if (bad) atomicMax(err, 1);
}
#endif

}  // namespace cofhe
