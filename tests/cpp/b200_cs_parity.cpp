// TEST DRIVER for the drop-in class include/cuda/cuda_cryptosystem.hpp: runs the same calls through
// CoFHE::B200CryptoSystem (device) and the reference's CoFHE::CPUCryptoSystem (host), both default-seeded, and counts
// mismatching results (ciphertext forms compared bit for bit, plaintexts as integers).  Built by tests/cpp/build.py
// where /root/reference exists (the reference headers are needed to compile it); the binary travels to the GPU box.
// usage: b200_cs_parity [N_bits k]
//        b200_cs_parity --bench matadd ROWS COLS ITERS CPU_ITERS     (benchmarks/local.cpp:65-138 at ROWS x COLS)
//        b200_cs_parity --bench matmul N M P ITERS CPU_ITERS         (benchmarks/local.cpp:140-215)
// The --bench modes time the chained calls of the reference's own benchmark THROUGH THE CLASS API (Tensor<CipherText*> in,
// Tensor<CipherText*> out, all marshalling inside the timed region) on the device backend and on the reference class
// (all OpenMP threads), compare the tensors after CPU_ITERS steps bit for bit, and print one JSON line (bench.py's
// `e2e_class`).
#include "cofhe.hpp"
#include "cuda/cuda_cryptosystem.hpp"
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <iostream>
using namespace CoFHE;
namespace {
typedef B200CryptoSystem::CipherText CT;
typedef B200CryptoSystem::PlainText PT;
double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
Tensor<PT*> pattern(size_t r, size_t c) {
  Tensor<PT*> t(r, c, nullptr); t.flatten();
  for (size_t i = 0; i < r * c; i++) t.at(i) = new PT((unsigned long)(i + 1));
  t.reshape({r, c});
  return t;
}
void free_cts(Tensor<CT*>& t) { t.flatten(); for (size_t i = 0; i < t.num_elements(); i++) delete t.at(i); }
size_t diff(Tensor<CT*> a, Tensor<CT*> b) {
  a.flatten(); b.flatten();
  if (a.num_elements() != b.num_elements()) return (size_t)-1;
  size_t bad = 0;
  for (size_t i = 0; i < a.num_elements(); i++) bad += !(a.at(i)->c1() == b.at(i)->c1() && a.at(i)->c2() == b.at(i)->c2());
  return bad;
}
// chain: res = f(ct0); res = f(res) ... `iters` calls in all; returns seconds; *at_k = deep copy after step k
template <class F>
double chain(F f, const Tensor<CT*>& ct0, size_t iters, size_t k, Tensor<CT*>* at_k) {
  double t0 = now();
  Tensor<CT*> res = f(ct0);
  double pause = 0;
  for (size_t i = 1;; i++) {
    if (i == k && at_k) {
      double p0 = now();
      auto shape = res.shape(); res.flatten();
      Tensor<CT*> cp({res.num_elements()}, nullptr);
      for (size_t j = 0; j < res.num_elements(); j++) cp.at(j) = new CT(*res.at(j));
      cp.reshape(shape); res.reshape(shape);
      *at_k = cp;
      pause += now() - p0;
    }
    if (i >= iters) break;
    Tensor<CT*> nx = f(res);
    free_cts(res);
    res = nx;
  }
  free_cts(res);
  return now() - t0 - pause;
}
int bench(int argc, char** argv) {
  std::string op = argv[2];
  bool mm = op == "matmul";
  if (argc < (mm ? 8 : 7)) { std::cerr << "bad --bench arguments" << std::endl; return 2; }
  size_t a1 = atol(argv[3]), a2 = atol(argv[4]), a3 = mm ? atol(argv[5]) : 0;
  size_t iters = atol(argv[mm ? 6 : 5]), cpu_iters = atol(argv[mm ? 7 : 6]);
  if (cpu_iters > iters) cpu_iters = iters;
  B200CryptoSystem cs(128, 128);
  CPUCryptoSystem ref(128, 128);
  auto sk = cs.keygen(); auto pk = cs.keygen(sk);
  auto sk2 = ref.keygen(); auto pk2 = ref.keygen(sk2);
  size_t bad = 0;
  double tb = 0, tc = 0;
  size_t units = 0;
  Tensor<CT*> gk(1, nullptr), ck(1, nullptr);
  if (!mm) {
    auto pt1 = pattern(a1, a2), pt2 = pattern(a1, a2);
    auto ct1 = cs.encrypt_tensor(pk, pt1), ct2 = cs.encrypt_tensor(pk, pt2);
    auto rt1 = ref.encrypt_tensor(pk2, pt1), rt2 = ref.encrypt_tensor(pk2, pt2);
    bad += diff(ct1, rt1) + diff(ct2, rt2);
    { chain([&](const Tensor<CT*>& x) { return cs.add_ciphertext_tensors(pk, x, ct2); }, ct1, 2, 0, nullptr); }  // warm-up
    tb = chain([&](const Tensor<CT*>& x) { return cs.add_ciphertext_tensors(pk, x, ct2); }, ct1, iters, cpu_iters, &gk);
    if (cpu_iters) tc = chain([&](const Tensor<CT*>& x) { return ref.add_ciphertext_tensors(pk2, x, rt2); }, rt1, cpu_iters, cpu_iters, &ck);
    units = a1 * a2;
  } else {
    auto pt1 = pattern(a1, a2), pt2 = pattern(a2, a3);
    auto ct1 = cs.encrypt_tensor(pk, pt1);
    auto rt1 = ref.encrypt_tensor(pk2, pt1);
    bad += diff(ct1, rt1);
    tb = chain([&](const Tensor<CT*>& x) { return cs.scal_ciphertext_tensors(pk, pt2, x); }, ct1, iters, cpu_iters, &gk);
    if (cpu_iters) tc = chain([&](const Tensor<CT*>& x) { return ref.scal_ciphertext_tensors(pk2, pt2, x); }, rt1, cpu_iters, cpu_iters, &ck);
    units = 2 * a1 * a2 * a3;
  }
  if (cpu_iters) bad += diff(gk, ck);
  std::cout << "{\"op\": \"" << op << "\", \"shape\": [" << a1 << ", " << a2 << ", " << a3 << "], \"iters\": " << iters
            << ", \"b200_s\": " << tb << ", \"b200_units_per_s\": " << units * iters / tb << ", \"cpu_iters\": " << cpu_iters
            << ", \"cpu_s\": " << tc << ", \"cpu_units_per_s\": " << (cpu_iters ? units * cpu_iters / tc : 0.0)
            << ", \"cpu_threads\": " << omp_get_max_threads() << ", \"mismatches\": " << bad << ", \"launches\": " << cs.gpu_launches() << "}" << std::endl;
  return bad != 0;
}
}  // namespace
int main(int argc, char** argv) {
  if (argc > 2 && !strcmp(argv[1], "--bench")) return bench(argc, argv);
  unsigned nbits = argc > 2 ? atoi(argv[1]) : 128, kk = argc > 2 ? atoi(argv[2]) : 128;
  B200CryptoSystem cs(nbits, kk);
  CPUCryptoSystem ref(nbits, kk);
  auto sk = cs.keygen(); auto pk = cs.keygen(sk);
  auto sk2 = ref.keygen(); auto pk2 = ref.keygen(sk2);
  size_t n = 4, m = 3, p = 2;
  Tensor<B200CryptoSystem::PlainText*> pt(n, m, nullptr); pt.flatten();
  for (size_t i = 0; i < n*m; i++) pt.at(i) = new B200CryptoSystem::PlainText((unsigned long)(i+1));
  pt.reshape({n, m});
  auto ct = cs.encrypt_tensor(pk, pt);
  auto ct2 = ref.encrypt_tensor(pk2, pt);
  ct.flatten(); ct2.flatten();
  int bad = 0;
  for (size_t i = 0; i < n*m; i++) bad += !(ct.at(i)->c1() == ct2.at(i)->c1() && ct.at(i)->c2() == ct2.at(i)->c2());
  ct.reshape({n, m}); ct2.reshape({n, m});
  auto sum = cs.add_ciphertext_tensors(pk, ct, ct); auto sum2 = ref.add_ciphertext_tensors(pk2, ct2, ct2);
  sum.flatten(); sum2.flatten();
  for (size_t i = 0; i < n*m; i++) bad += !(sum.at(i)->c1() == sum2.at(i)->c1() && sum.at(i)->c2() == sum2.at(i)->c2());
  Tensor<B200CryptoSystem::PlainText*> s(m, p, nullptr); s.flatten();
  for (size_t i = 0; i < m*p; i++) s.at(i) = new B200CryptoSystem::PlainText((unsigned long)(7*i+3));
  s.reshape({m, p});
  auto mm = cs.scal_ciphertext_tensors(pk, s, ct); auto mm2 = ref.scal_ciphertext_tensors(pk2, s, ct2);
  mm.flatten(); mm2.flatten();
  for (size_t i = 0; i < n*p; i++) bad += !(mm.at(i)->c1() == mm2.at(i)->c1() && mm.at(i)->c2() == mm2.at(i)->c2());
  mm.reshape({n, p});
  auto dec = cs.decrypt_tensor(sk, mm); dec.flatten();
  auto dec2 = ref.decrypt_tensor(sk2, mm2); dec2.flatten();
  for (size_t i = 0; i < n*p; i++) bad += !(*dec.at(i) == *dec2.at(i));
  auto shares = cs.keygen(sk, 3, 5);
  Vector<Tensor<B200CryptoSystem::PartDecryptionResult*>> pd;
  for (int q = 0; q < 3; q++) pd.push_back(cs.part_decrypt_tensor(shares[q][0], mm));
  auto comb = cs.combine_part_decryption_results_tensor(mm, pd); comb.flatten();
  for (size_t i = 0; i < n*p; i++) bad += !(*comb.at(i) == *dec2.at(i));
  auto neg = cs.negate_ciphertext_tensor(pk, ct); auto neg2 = ref.negate_ciphertext_tensor(pk2, ct2);
  neg.flatten(); neg2.flatten();
  for (size_t i = 0; i < n*m; i++) bad += !(neg.at(i)->c1() == neg2.at(i)->c1() && neg.at(i)->c2() == neg2.at(i)->c2());
  std::cout << "mismatches " << bad << " launches " << cs.gpu_launches() << std::endl;
  return bad != 0;
}
