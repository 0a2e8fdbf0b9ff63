// TEST DRIVER for the drop-in class include/cuda/cuda_cryptosystem.hpp: runs the same calls through
// CoFHE::B200CryptoSystem (device) and the reference's CoFHE::CPUCryptoSystem (host), both default-seeded, and counts
// mismatching results (ciphertext forms compared bit for bit, plaintexts as integers).  Built by tests/cpp/build.py
// where /root/reference exists (the reference headers are needed to compile it); the binary travels to the GPU box.
// usage: b200_cs_parity [N_bits k]
#include "cofhe.hpp"
#include "cuda/cuda_cryptosystem.hpp"
#include <cstdlib>
#include <iostream>
using namespace CoFHE;
int main(int argc, char** argv) {
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
