// TEST DRIVER for the drop-in class include/cuda/cuda_cryptosystem.hpp: runs the same calls through
// CoFHE::B200CryptoSystem (device) and the reference's CoFHE::CPUCryptoSystem (host), both default-seeded, and counts
// mismatching results (ciphertext forms compared bit for bit, plaintexts as integers).  Built by tests/cpp/build.py
// where /root/reference exists (the reference headers are needed to compile it); the binary travels to the GPU box.
// usage: b200_cs_parity [N_bits k [compact]]
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
#include <thread>
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
  double tb = 0, tc = 0, tdev = 0;
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
    // the same chain with the tensors resident on the device: one upload of the operands, `iters` adds, one download
    {
      double t0 = now();
      DeviceBatch d1 = cs.upload_ciphertexts(ct1), d2 = cs.upload_ciphertexts(ct2);
      DeviceBatch acc = cs.add_ciphertext_tensors(pk, d1, d2);
      for (size_t i = 1; i < iters; i++) acc = cs.add_ciphertext_tensors(pk, acc, d2);
      auto fin = cs.download_ciphertexts(acc);
      tdev = now() - t0;
      // compare with the class-API chain at the last step
      Tensor<CT*> last(1, nullptr);
      chain([&](const Tensor<CT*>& x) { return cs.add_ciphertext_tensors(pk, x, ct2); }, ct1, iters, iters, &last);
      bad += diff(fin, last);
      free_cts(fin); free_cts(last);
    }
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
            << ", \"b200_resident_s\": " << tdev << ", \"b200_resident_units_per_s\": " << (tdev > 0 ? units * iters / tdev : 0.0)
            << ", \"cpu_threads\": " << omp_get_max_threads() << ", \"mismatches\": " << bad << ", \"launches\": " << cs.gpu_launches() << "}" << std::endl;
  return bad != 0;
}
}  // namespace
// --threads T ROWS COLS ITERS: T host threads share ONE B200CryptoSystem, as the reference's server does with its 8 worker
// threads (include/node/server.hpp:16,187-196); each chains ITERS add_ciphertext_tensors on its own ROWS x COLS tensor and
// part-decrypts the result.  Results are compared with the reference class; the same work is also timed from one thread.
int threads_mode(int argc, char** argv) {
  if (argc < 6) return 2;
  size_t T = atoi(argv[2]), rows = atoi(argv[3]), cols = atoi(argv[4]), iters = atoi(argv[5]);
  B200CryptoSystem cs(128, 128);
  CPUCryptoSystem ref(128, 128);
  auto sk = cs.keygen(); auto pk = cs.keygen(sk);
  auto sk2 = ref.keygen(); auto pk2 = ref.keygen(sk2);
  auto shares = cs.keygen(sk, 3, 5);
  const Tensor<CT*> e1(1, nullptr);
  const Tensor<B200CryptoSystem::PartDecryptionResult*> e2(1, nullptr);
  std::vector<Tensor<CT*>> in(T, e1), want(T, e1), got(T, e1);
  std::vector<Tensor<B200CryptoSystem::PartDecryptionResult*>> wantp(T, e2), gotp(T, e2);
  for (size_t t = 0; t < T; t++) {
    auto pt = pattern(rows, cols);
    in[t] = ref.encrypt_tensor(pk2, pt);
    Tensor<CT*> r = ref.add_ciphertext_tensors(pk2, in[t], in[t]);
    for (size_t i = 1; i < iters; i++) { auto nx = ref.add_ciphertext_tensors(pk2, r, in[t]); free_cts(r); r = nx; }
    want[t] = r;
    wantp[t] = ref.part_decrypt_tensor(shares[0][0], r);
  }
  auto work = [&](size_t t) {
    Tensor<CT*> r = cs.add_ciphertext_tensors(pk, in[t], in[t]);
    for (size_t i = 1; i < iters; i++) { auto nx = cs.add_ciphertext_tensors(pk, r, in[t]); free_cts(r); r = nx; }
    got[t] = r;
    gotp[t] = cs.part_decrypt_tensor(shares[0][0], r);
  };
  work(0);  // warm-up: tables, arenas
  free_cts(got[0]);
  double t0 = now();
  for (size_t t = 0; t < T; t++) { work(t); if (t + 1 < T) free_cts(got[t]); }
  double seq = now() - t0;
  free_cts(got[T - 1]);
  std::vector<std::exception_ptr> errs(T);
  t0 = now();
  {
    std::vector<std::thread> th;
    for (size_t t = 0; t < T; t++) th.emplace_back([&, t]() { try { work(t); } catch (...) { errs[t] = std::current_exception(); } });
    for (auto& x : th) x.join();
  }
  double par = now() - t0;
  size_t bad = 0;
  for (size_t t = 0; t < T; t++) {
    if (errs[t]) { bad += 1000; continue; }
    bad += diff(got[t], want[t]);
    auto a = gotp[t], b = wantp[t]; a.flatten(); b.flatten();
    for (size_t i = 0; i < a.num_elements(); i++) bad += !(*a.at(i) == *b.at(i));
  }
  std::cout << "{\"threads\": " << T << ", \"shape\": [" << rows << ", " << cols << "], \"iters\": " << iters << ", \"one_thread_s\": " << seq
            << ", \"threads_s\": " << par << ", \"lanes\": " << cs.num_lanes() << ", \"mismatches\": " << bad << "}" << std::endl;
  return bad != 0;
}
int main(int argc, char** argv) {
  if (argc > 2 && !strcmp(argv[1], "--bench")) return bench(argc, argv);
  if (argc > 2 && !strcmp(argv[1], "--threads")) return threads_mode(argc, argv);
  unsigned nbits = argc > 2 ? atoi(argv[1]) : 128, kk = argc > 2 ? atoi(argv[2]) : 128;
  // third argument 1: the compact variant of CL_HSM2k (c1, h, pk in Cl(DeltaK)) on both sides
  const bool compact = argc > 3 && atoi(argv[3]) != 0;
  B200CryptoSystem cs(nbits, kk, compact);
  CPUCryptoSystem ref(nbits, kk, compact);
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
  // single-ciphertext calls: same draws from the two default-seeded RandGens, results compared bit for bit
  { auto ks2 = ref.keygen(sk2, 3, 5); (void)ks2; }  // keep the reference's RandGen in step with cs.keygen(sk, 3, 5) above
  try {
    auto same = [](const B200CryptoSystem::CipherText& a, const B200CryptoSystem::CipherText& b) { return a.c1() == b.c1() && a.c2() == b.c2(); };
    B200CryptoSystem::PlainText m1((unsigned long)123456789), m2((unsigned long)42), sc((unsigned long)1000003), neg_s(-77L);
    auto e1 = cs.encrypt(pk, m1); auto r1 = ref.encrypt(pk2, m1);
    auto e2 = cs.encrypt(pk, m2); auto r2 = ref.encrypt(pk2, m2);
    bad += !same(e1, r1) + !same(e2, r2);
    auto a = cs.add_ciphertexts(pk, e1, e2); auto ar = ref.add_ciphertexts(pk2, r1, r2);
    bad += !same(a, ar);
    auto sm = cs.scal_ciphertext(pk, sc, e1); auto smr = ref.scal_ciphertext(pk2, sc, r1);
    bad += !same(sm, smr);
    auto sn = cs.scal_ciphertext(pk, neg_s, e2); auto snr = ref.scal_ciphertext(pk2, neg_s, r2);
    bad += !same(sn, snr);
    auto ng = cs.negate_ciphertext(pk, e1); auto ngr = ref.negate_ciphertext(pk2, r1);
    bad += !same(ng, ngr);
    bad += !(cs.decrypt(sk, sm) == ref.decrypt(sk2, smr)) + !(cs.decrypt(sk, ng) == ref.decrypt(sk2, ngr));
    Vector<B200CryptoSystem::PartDecryptionResult> ps, psr;
    Vector<B200CryptoSystem::PartDecryptionResult*> pp;
    for (int q = 0; q < 3; q++) { ps.push_back(cs.part_decrypt(shares[q][0], a)); psr.push_back(ref.part_decrypt(shares[q][0], ar)); bad += !(ps[q] == psr[q]); }
    bad += !(cs.combine_part_decryption_results(a, ps) == ref.combine_part_decryption_results(ar, psr));
    bad += !(cs.combine_part_decryption_results(a, ps) == ref.decrypt(sk2, ar));
    for (int q = 0; q < 3; q++) pp.push_back(&ps[q]);
    auto cv = cs.combine_part_decryption_results_vector(a, pp); auto cvr = ref.combine_part_decryption_results_vector(ar, pp);
    bad += cv.size() != cvr.size();
    for (size_t i = 0; i < cv.size() && i < cvr.size(); i++) { bad += !(*cv[i] == *cvr[i]); delete cv[i]; delete cvr[i]; }
    // 0-D tensors take the single-ciphertext paths (one draw each)
    Tensor<B200CryptoSystem::CipherText*> z1(new B200CryptoSystem::CipherText(e1)), z2(new B200CryptoSystem::CipherText(e2));
    Tensor<B200CryptoSystem::CipherText*> y1(new B200CryptoSystem::CipherText(r1)), y2(new B200CryptoSystem::CipherText(r2));
    auto zs = cs.add_ciphertext_tensors(pk, z1, z2); auto ys = ref.add_ciphertext_tensors(pk2, y1, y2);
    bad += !same(*zs.get_value(), *ys.get_value());
  } catch (const std::exception& e) { std::cout << "single-ciphertext block threw: " << e.what() << " after " << bad << " mismatches" << std::endl; bad += 1000; }
  // the reference's binary tensor formats through the device wire kernels: byte-identical to the reference's own
  // serialisers, and back; then the same tensors kept on the device across chained calls
  try {
    ct.reshape({n, m}); ct2.reshape({n, m});
    std::string wb = cs.serialize_ciphertext_tensor(ct), wr = ref.serialize_ciphertext_tensor(ct2);
    { size_t _b = (size_t)(wb != wr); if (_b) std::cerr << "wire-block check 1 failed (" << _b << "): wb != wr" << std::endl; bad += _b; }
    auto back = cs.deserialize_ciphertext_tensor(wr); back.flatten(); ct.flatten();
    for (size_t i = 0; i < n*m; i++) bad += !(back.at(i)->c1() == ct.at(i)->c1() && back.at(i)->c2() == ct.at(i)->c2());
    ct.reshape({n, m});
    mm.reshape({n, p});
    std::string pb = cs.serialize_part_decryption_result_tensor(pd[0]), pr = ref.serialize_part_decryption_result_tensor(pd[0]);
    { size_t _b = (size_t)(pb != pr); if (_b) std::cerr << "wire-block check 2 failed (" << _b << "): pb != pr" << std::endl; bad += _b; }
    auto pback = cs.deserialize_part_decryption_result_tensor(pr); pback.flatten();
    auto pd0 = pd[0]; pd0.flatten();
    for (size_t i = 0; i < n*p; i++) bad += !(*pback.at(i) == *pd0.at(i));
    Tensor<B200CryptoSystem::PlainText*> mixed(2, 3, nullptr); mixed.flatten();
    long vals[6] = {0, -7, 123456789, -1, 65536, 255};
    for (int i = 0; i < 6; i++) mixed.at(i) = new B200CryptoSystem::PlainText(vals[i]);
    mixed.reshape({2, 3});
    std::string tb = cs.serialize_plaintext_tensor(mixed), tr = ref.serialize_plaintext_tensor(mixed);
    { size_t _b = (size_t)(tb != tr); if (_b) std::cerr << "wire-block check 3 failed (" << _b << "): tb != tr" << std::endl; bad += _b; }
    auto tback = cs.deserialize_plaintext_tensor(tr); auto tref = ref.deserialize_plaintext_tensor(tr);
    { size_t _b = (size_t)(tback.shape() != tref.shape()); if (_b) std::cerr << "wire-block check 4 failed (" << _b << "): tback.shape() != tref.shape()" << std::endl; bad += _b; }
    tback.flatten(); tref.flatten();
    for (size_t i = 0; i < 6; i++) bad += !(*tback.at(i) == *tref.at(i));
    // device-resident chain: wire -> device -> 5 chained adds + matmul + threshold decryption -> wire, no BICYCL object in between
    DeviceBatch dct = cs.deserialize_tensor_device(wr, 0, 0);
    DeviceBatch acc = cs.add_ciphertext_tensors(pk, dct, dct);
    auto hacc = ref.add_ciphertext_tensors(pk2, ct2, ct2);
    for (int it = 0; it < 4; it++) {
      acc = cs.add_ciphertext_tensors(pk, acc, dct);
      auto nx = ref.add_ciphertext_tensors(pk2, hacc, ct2); free_cts(hacc); hacc = nx;
    }
    { size_t _b = (size_t)(cs.serialize_tensor(acc) != ref.serialize_ciphertext_tensor(hacc)); if (_b) std::cerr << "wire-block check 5 failed (" << _b << "): cs.serialize_tensor(acc) != ref.serialize_ciphertext_tensor(" << std::endl; bad += _b; }
    DeviceBatch ds = cs.upload_plaintexts(s, 0);
    DeviceBatch dmm = cs.scal_ciphertext_tensors(pk, ds, acc);
    auto hmm = ref.scal_ciphertext_tensors(pk2, s, hacc);
    { size_t _b = (size_t)(cs.serialize_tensor(dmm) != ref.serialize_ciphertext_tensor(hmm)); if (_b) std::cerr << "wire-block check 6 failed (" << _b << "): cs.serialize_tensor(dmm) != ref.serialize_ciphertext_tensor(" << std::endl; bad += _b; }
    DeviceBatch dneg = cs.negate_ciphertext_tensor(pk, dmm);
    auto hneg = ref.negate_ciphertext_tensor(pk2, hmm);
    { size_t _b = (size_t)(cs.serialize_tensor(dneg) != ref.serialize_ciphertext_tensor(hneg)); if (_b) std::cerr << "wire-block check 7 failed (" << _b << "): cs.serialize_tensor(dneg) != ref.serialize_ciphertext_tensor" << std::endl; bad += _b; }
    DeviceBatch dp0 = cs.part_decrypt_tensor(shares[0][0], dneg), dp1 = cs.part_decrypt_tensor(shares[1][0], dneg), dp2 = cs.part_decrypt_tensor(shares[2][0], dneg);
    { size_t _b = (size_t)(cs.serialize_tensor(dp0) != ref.serialize_part_decryption_result_tensor(ref.part_decrypt_tensor(shares[0][0], hneg))); if (_b) std::cerr << "wire-block check 8 failed (" << _b << "): cs.serialize_tensor(dp0) != ref.serialize_part_decryption_re" << std::endl; bad += _b; }
    Vector<const DeviceBatch*> dps; dps.push_back(&dp0); dps.push_back(&dp1); dps.push_back(&dp2);
    DeviceBatch dplain = cs.combine_part_decryption_results_tensor(dneg, dps);
    auto hplain = ref.decrypt_tensor(sk2, hneg);
    { size_t _b = (size_t)(cs.serialize_tensor(dplain) != ref.serialize_plaintext_tensor(hplain)); if (_b) std::cerr << "wire-block check 9 failed (" << _b << "): cs.serialize_tensor(dplain) != ref.serialize_plaintext_tenso" << std::endl; bad += _b; }
    DeviceBatch ddec = cs.decrypt_tensor(sk, dneg);
    { size_t _b = (size_t)(cs.serialize_tensor(ddec) != ref.serialize_plaintext_tensor(hplain)); if (_b) std::cerr << "wire-block check 10 failed (" << _b << "): cs.serialize_tensor(ddec) != ref.serialize_plaintext_tensor(" << std::endl; bad += _b; }
    // encrypt_tensor from a device plaintext batch: same draw on both sides.  Non-negative values only: the reference hands a
    // negative plaintext to power_of_f unreduced, and for some of them (-1) the result is a ciphertext that the reference
    // itself cannot decrypt ("the form is not in the kernel"); the device reduces modulo 2^k first, like Mpz::mod would
    for (int i = 0; i < 6; i++) { delete mixed.at(i / 3, i % 3); mixed.at(i / 3, i % 3) = new B200CryptoSystem::PlainText((unsigned long)(vals[i] < 0 ? -vals[i] * 1000003 : vals[i])); }
    DeviceBatch dpt = cs.upload_plaintexts(mixed, 0);
    DeviceBatch denc = cs.encrypt_tensor(pk, dpt);
    auto henc = ref.encrypt_tensor(pk2, mixed);
    { size_t _b = (size_t)(cs.serialize_tensor(denc) != ref.serialize_ciphertext_tensor(henc)); if (_b) std::cerr << "wire-block check 11 failed (" << _b << "): cs.serialize_tensor(denc) != ref.serialize_ciphertext_tensor" << std::endl; bad += _b; }
    auto dl = cs.download_ciphertexts(denc);
    { auto x = dl, y = henc; x.flatten(); y.flatten();
      for (size_t i = 0; i < x.num_elements(); i++) if (!(x.at(i)->c1() == y.at(i)->c1() && x.at(i)->c2() == y.at(i)->c2())) std::cerr << "encrypt_tensor(device plaintexts): element " << i << " differs (c1 equal: " << (x.at(i)->c1() == y.at(i)->c1()) << ")" << std::endl; }
    bad += diff(dl, henc);
    // float codec on the device against the host object's scalar codec (= the reference's)
    {
      float fv[8] = {0.0f, 1.0f, -1.0f, 2.5f, -2.5f, 123456.0f, -65536.75f, 1e20f};
      Tensor<float> ft(2, 4, 0.0f); ft.flatten();
      for (int i = 0; i < 8; i++) ft.at(i) = fv[i];
      ft.reshape({2, 4});
      DeviceBatch fp = cs.make_plaintext_tensor(ft);
      auto fpt = cs.download_plaintexts(fp); fpt.flatten();
      for (int i = 0; i < 8; i++) { size_t _b = !(*fpt.at(i) == ref.make_plaintext(fv[i])); if (_b) std::cerr << "make_plaintext_tensor differs at " << i << std::endl; bad += _b; }
      auto fb = cs.get_float_tensor(fp); fb.flatten();
      for (int i = 0; i < 8; i++) { size_t _b = !(fb.at(i) == ref.get_float_from_plaintext(*fpt.at(i))); if (_b) std::cerr << "get_float_tensor differs at " << i << std::endl; bad += _b; }
    }
  } catch (const std::exception& e) { std::cout << "wire / device-resident block threw: " << e.what() << " after " << bad << " mismatches" << std::endl; bad += 1000; }
  // a ciphertext of a larger discriminant does not fit the limb shape: the marshalling loops run under OpenMP, so the
  // error must come back as std::invalid_argument from the call, not as an exception escaping a parallel region
  if (nbits < 256) {
    try {
      CPUCryptoSystem other(256, kk, compact);
      auto osk = other.keygen(); auto opk = other.keygen(osk);
      Tensor<B200CryptoSystem::PlainText*> opt(2, nullptr);
      for (size_t i = 0; i < 2; i++) opt.at(i) = new B200CryptoSystem::PlainText((unsigned long)(i + 5));
      auto oct = other.encrypt_tensor(opk, opt);
      bool thrown = false;
      try { auto r = cs.add_ciphertext_tensors(pk, oct, oct); } catch (const std::invalid_argument&) { thrown = true; }
      if (!thrown) { std::cerr << "alien ciphertexts were not rejected" << std::endl; bad++; }
      // and the object is usable afterwards
      auto again = cs.add_ciphertext_tensors(pk, ct, ct);
      bad += diff(again, sum2);
    } catch (const std::exception& e) { std::cout << "alien-ciphertext block threw: " << e.what() << std::endl; bad += 1000; }
  }
  std::cout << "mismatches " << bad << " launches " << cs.gpu_launches() << std::endl;
  return bad != 0;
}
