// ORACLE / TEST INFRASTRUCTURE ONLY -- never linked into, imported by or called from the product path.
//
// A small C ABI around the UNMODIFIED reference (CoFHE CPUCryptoSystem + BICYCL + system libgmp.so.10),
// compiled from the sources where they lie under /root/reference by oracle/Makefile into
// oracle/_ref/libcofhe_ref.so.  Three users:
//   * tools/gen_golden.py   -- generates the committed fixtures under tests/golden/
//   * tests/                -- validates oracle/qf_oracle.py (the restatement) against the real thing
//   * bench.py              -- cpu_baseline leg / --impl reference (times the reference on the host cores)
//
// All integers cross this ABI as decimal strings separated by single spaces (ref_call), or as SoA
// uint32 limb buffers identical to the product layout (ref_bulk_*; include/cofhe_b200.h "Layout").
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <sstream>
#include <string>
#include <vector>
#include <chrono>
#include <stdexcept>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "cofhe.hpp"

using namespace CoFHE;
using BICYCL::Mpz;
using BICYCL::QFI;
typedef CPUCryptoSystem CS;

namespace {

struct Ref {
  CS cs;
  CS::SecretKey *sk = nullptr;
  CS::PublicKey *pk = nullptr;
  Ref(unsigned nbits, unsigned k, bool compact = false) : cs(nbits, k, compact) {}
  ~Ref() { delete sk; delete pk; }
};

std::vector<Mpz> parse(const char *s) {
  std::vector<Mpz> v;
  std::istringstream in(s ? s : "");
  std::string tok;
  while (in >> tok) v.emplace_back(tok);
  return v;
}
void put(std::ostringstream &o, const Mpz &z) { o << z << ' '; }
void put(std::ostringstream &o, const QFI &f) { put(o, f.a()); put(o, f.b()); put(o, f.c()); }
void put(std::ostringstream &o, const CS::CipherText &c) { put(o, c.c1()); put(o, c.c2()); }
QFI form(const std::vector<Mpz> &v, size_t i) { return QFI(v[i], v[i + 1], v[i + 2], true); }
CS::CipherText ctext(const std::vector<Mpz> &v, size_t i) { return CS::CipherText(form(v, i), form(v, i + 3)); }
size_t to_sz(const Mpz &z) { return (size_t)mpz_get_ui((mpz_srcptr)z); }

QFI pow2k(const BICYCL::ClassGroup &G, QFI f, size_t n) { for (size_t i = 0; i < n; i++) G.nudupl(f, f); return f; }

char *dup(const std::string &s) { char *r = (char *)malloc(s.size() + 1); memcpy(r, s.c_str(), s.size() + 1); return r; }

template <class T> Tensor<T *> make_tensor(size_t r, size_t c) { return Tensor<T *>(r, c, nullptr); }

std::string call(Ref &R, const std::string &op, const std::vector<Mpz> &a) {
  std::ostringstream o;
  const BICYCL::CL_HSM2k &H = R.cs.get_hsm2k();
  const BICYCL::ClassGroup &G = H.Cl_G();
  if (op == "params") {
    // N DeltaK Delta k exponent_bound d e h he hd hde L large_message secretkey_bound cleartext_bound
    size_t d = (H.encrypt_randomness_bound().nbits() + 1) / 2, e = d / 2 + 1;
    put(o, H.N()); put(o, H.DeltaK()); put(o, H.Delta()); o << H.k() << ' ';
    put(o, H.encrypt_randomness_bound()); o << d << ' ' << e << ' ';
    put(o, H.h()); put(o, pow2k(G, H.h(), e)); put(o, pow2k(G, H.h(), d)); put(o, pow2k(G, H.h(), d + e));
    put(o, G.default_nucomp_bound()); o << (H.large_message_variant() ? 1 : 0) << ' ';
    put(o, H.secretkey_bound()); put(o, H.cleartext_bound());
  } else if (op == "set_seed") {
    R.cs.get_rand_gen().set_seed(a[0]);
  } else if (op == "rand_mpz") {
    put(o, R.cs.get_rand_gen().random_mpz(a[0]));
  } else if (op == "keygen") {               // draws sk from rand_gen, installs (sk, pk); returns sk pk
    delete R.sk; delete R.pk;
    R.sk = new CS::SecretKey(R.cs.keygen());
    R.pk = new CS::PublicKey(R.cs.keygen(*R.sk));
    put(o, *R.sk); put(o, R.pk->elt());
  } else if (op == "set_sk") {
    delete R.sk; delete R.pk;
    R.sk = new CS::SecretKey(H, a[0]);
    R.pk = new CS::PublicKey(R.cs.keygen(*R.sk));
    put(o, *R.sk); put(o, R.pk->elt());
  } else if (op == "pk_precomp") {           // pk^(2^e) pk^(2^d) pk^(2^(d+e))
    size_t d = (H.encrypt_randomness_bound().nbits() + 1) / 2, e = d / 2 + 1;
    put(o, pow2k(G, R.pk->elt(), e)); put(o, pow2k(G, R.pk->elt(), d)); put(o, pow2k(G, R.pk->elt(), d + e));
  } else if (op == "keygen_shares") {        // t n -> n_parties shares_per_party then shares row-major
    auto sh = R.cs.keygen(*R.sk, to_sz(a[0]), to_sz(a[1]));
    o << sh.size() << ' ' << sh[0].size() << ' ';
    for (auto &row : sh) for (auto &s : row) put(o, s);
  } else if (op == "one") {
    put(o, G.one());
  } else if (op == "reduce") {               // QFI ctor without bypass checks and reduces?  use compose with one
    QFI f(a[0], a[1], a[2], true), r;
    G.nucomp(r, G.one(), f);                 // second operand may be unreduced
    put(o, r);
  } else if (op == "nucomp") {
    QFI r; G.nucomp(r, form(a, 0), form(a, 3)); put(o, r);
  } else if (op == "nucompinv") {
    QFI r; G.nucompinv(r, form(a, 0), form(a, 3)); put(o, r);
  } else if (op == "nudupl") {
    QFI r; G.nudupl(r, form(a, 0)); put(o, r);
  } else if (op == "nupow") {
    QFI r; G.nupow(r, form(a, 0), a[3]); put(o, r);
  } else if (op == "lift") {                 // form of Cl(DeltaK) -> Cl(Delta) (compact variant: CL_HSM2k.inl:458-474)
    QFI f = form(a, 0); H.from_Cl_DeltaK_to_Cl_Delta(f); put(o, f);
  } else if (op == "nucomp_delta") {         // composition in Cl(Delta) whatever the variant
    QFI r; H.Cl_Delta().nucomp(r, form(a, 0), form(a, 3)); put(o, r);
  } else if (op == "power_of_h") {
    QFI r; H.power_of_h(r, a[0]); put(o, r);
  } else if (op == "pk_exp") {
    QFI r; R.pk->exponentiation(H, r, a[0]); put(o, r);
  } else if (op == "power_of_f") {
    put(o, H.power_of_f(a[0]));
  } else if (op == "dlog_in_F") {
    put(o, H.dlog_in_F(form(a, 0)));
  } else if (op == "encrypt_r") {            // m r -> ct (explicit randomness)
    CS::CipherText c(H, *R.pk, BICYCL::CL_HSM2k::ClearText(H, a[0]), a[1]); put(o, c);
  } else if (op == "encrypt") {              // m -> ct (1 draw from rand_gen)
    put(o, R.cs.encrypt(*R.pk, a[0]));
  } else if (op == "encrypt_tensor") {       // rows cols m... -> cts (1 draw per call)
    size_t r = to_sz(a[0]), c = to_sz(a[1]);
    auto pt = make_tensor<CS::PlainText>(r, c); pt.flatten();
    for (size_t i = 0; i < r * c; i++) pt.at(i) = new CS::PlainText(a[2 + i]);
    pt.reshape({r, c});
    auto ct = R.cs.encrypt_tensor(*R.pk, pt); ct.flatten(); pt.flatten();
    for (size_t i = 0; i < r * c; i++) { put(o, *ct.at(i)); delete ct.at(i); delete pt.at(i); }
  } else if (op == "decrypt") {
    put(o, R.cs.decrypt(*R.sk, ctext(a, 0)));
  } else if (op == "add_ciphertexts") {      // single, re-randomising (1 draw)
    put(o, R.cs.add_ciphertexts(*R.pk, ctext(a, 0), ctext(a, 6)));
  } else if (op == "scal_ciphertext") {      // s ct, single, re-randomising (1 draw)
    put(o, R.cs.scal_ciphertext(*R.pk, a[0], ctext(a, 1)));
  } else if (op == "add_tensors") {          // rows cols ct1... ct2...
    size_t r = to_sz(a[0]), c = to_sz(a[1]), n = r * c;
    auto t1 = make_tensor<CS::CipherText>(r, c), t2 = make_tensor<CS::CipherText>(r, c);
    t1.flatten(); t2.flatten();
    for (size_t i = 0; i < n; i++) { t1.at(i) = new CS::CipherText(ctext(a, 2 + 6 * i)); t2.at(i) = new CS::CipherText(ctext(a, 2 + 6 * (n + i))); }
    t1.reshape({r, c}); t2.reshape({r, c});
    auto res = R.cs.add_ciphertext_tensors(*R.pk, t1, t2); res.flatten(); t1.flatten(); t2.flatten();
    for (size_t i = 0; i < n; i++) { put(o, *res.at(i)); delete res.at(i); delete t1.at(i); delete t2.at(i); }
  } else if (op == "scal_matmul") {          // n m p ct[n*m]... s[m*p]... -> res[n*p] (1 draw: the zero ciphertext)
    size_t n = to_sz(a[0]), m = to_sz(a[1]), p = to_sz(a[2]);
    auto ct = make_tensor<CS::CipherText>(n, m); auto s = make_tensor<CS::PlainText>(m, p);
    ct.flatten(); s.flatten();
    for (size_t i = 0; i < n * m; i++) ct.at(i) = new CS::CipherText(ctext(a, 3 + 6 * i));
    for (size_t i = 0; i < m * p; i++) s.at(i) = new CS::PlainText(a[3 + 6 * n * m + i]);
    ct.reshape({n, m}); s.reshape({m, p});
    auto res = R.cs.scal_ciphertext_tensors(*R.pk, s, ct); res.flatten(); ct.flatten(); s.flatten();
    for (size_t i = 0; i < n * p; i++) { put(o, *res.at(i)); delete res.at(i); }
    for (size_t i = 0; i < n * m; i++) delete ct.at(i);
    for (size_t i = 0; i < m * p; i++) delete s.at(i);
  } else if (op == "part_decrypt") {         // share ct -> d
    put(o, R.cs.part_decrypt(a[0], ctext(a, 1)));
  } else if (op == "combine") {              // ct d0 d1 ... -> m
    Vector<CS::PartDecryptionResult> ds;
    for (size_t i = 6; i + 2 < a.size(); i += 3) ds.push_back(form(a, i));
    put(o, R.cs.combine_part_decryption_results(ctext(a, 0), ds));
  } else if (op == "make_plaintext") {       // integer-valued float -> plaintext through the float codec
    put(o, R.cs.make_plaintext((float)mpz_get_si((mpz_srcptr)a[0])));
  } else if (op == "negate_tensor") {        // count cts... -> cts  (ct^(2^k-1), no randomness)
    size_t n = to_sz(a[0]);
    auto t = make_tensor<CS::CipherText>(n, 1); t.flatten();
    for (size_t i = 0; i < n; i++) t.at(i) = new CS::CipherText(ctext(a, 1 + 6 * i));
    auto res = R.cs.negate_ciphertext_tensor(*R.pk, t); res.flatten();
    for (size_t i = 0; i < n; i++) { put(o, *res.at(i)); delete res.at(i); delete t.at(i); }
  } else if (op == "serialize_ct_tensor") {  // rows cols cts... -> the bytes of serialize_ciphertext_tensor, one number per byte
    size_t r = to_sz(a[0]), c = to_sz(a[1]), n = r * c;
    auto t = make_tensor<CS::CipherText>(r, c); t.flatten();
    for (size_t i = 0; i < n; i++) t.at(i) = new CS::CipherText(ctext(a, 2 + 6 * i));
    t.reshape({r, c});
    std::string bytes = R.cs.serialize_ciphertext_tensor(t);
    t.flatten();
    for (size_t i = 0; i < n; i++) delete t.at(i);
    for (unsigned char ch : bytes) o << (unsigned)ch << ' ';
  } else if (op == "deserialize_ct_tensor") {  // bytes... -> ndim dims... cts...
    std::string bytes(a.size(), 0);
    for (size_t i = 0; i < a.size(); i++) bytes[i] = (char)to_sz(a[i]);
    auto t = R.cs.deserialize_ciphertext_tensor(bytes);
    o << t.ndim() << ' ';
    for (size_t i = 0; i < t.ndim(); i++) o << t.shape()[i] << ' ';
    t.flatten();
    for (size_t i = 0; i < t.num_elements(); i++) { put(o, *t.at(i)); delete t.at(i); }
  } else if (op == "make_plaintext_bits") {  // IEEE-754 bit patterns of floats... -> make_plaintext of each
    for (auto &z : a) { uint32_t b = (uint32_t)to_sz(z); float x; memcpy(&x, &b, 4); put(o, R.cs.make_plaintext(x)); }
  } else if (op == "get_float_bits") {       // plaintexts... -> bit patterns of get_float_from_plaintext
    for (auto &z : a) { float x = R.cs.get_float_from_plaintext(z); uint32_t b; memcpy(&b, &x, 4); o << b << ' '; }
  } else if (op == "serialize_pdr_tensor") {  // rows cols forms... -> bytes of serialize_part_decryption_result_tensor
    size_t r = to_sz(a[0]), c = to_sz(a[1]), n = r * c;
    auto t = make_tensor<CS::PartDecryptionResult>(r, c); t.flatten();
    for (size_t i = 0; i < n; i++) t.at(i) = new CS::PartDecryptionResult(a[2 + 3 * i], a[3 + 3 * i], a[4 + 3 * i], true);
    t.reshape({r, c});
    std::string bytes = R.cs.serialize_part_decryption_result_tensor(t);
    t.flatten();
    for (size_t i = 0; i < n; i++) delete t.at(i);
    for (unsigned char ch : bytes) o << (unsigned)ch << ' ';
  } else if (op == "deserialize_pdr_tensor") {  // bytes... -> ndim dims... forms...
    std::string bytes(a.size(), 0);
    for (size_t i = 0; i < a.size(); i++) bytes[i] = (char)to_sz(a[i]);
    auto t = R.cs.deserialize_part_decryption_result_tensor(bytes);
    o << t.ndim() << ' ';
    for (size_t i = 0; i < t.ndim(); i++) o << t.shape()[i] << ' ';
    t.flatten();
    for (size_t i = 0; i < t.num_elements(); i++) { put(o, *t.at(i)); delete t.at(i); }
  } else if (op == "serialize_pt_tensor") {  // rows cols ints... -> bytes of serialize_plaintext_tensor
    size_t r = to_sz(a[0]), c = to_sz(a[1]), n = r * c;
    auto t = make_tensor<CS::PlainText>(r, c); t.flatten();
    for (size_t i = 0; i < n; i++) t.at(i) = new CS::PlainText(a[2 + i]);
    t.reshape({r, c});
    std::string bytes = R.cs.serialize_plaintext_tensor(t);
    t.flatten();
    for (size_t i = 0; i < n; i++) delete t.at(i);
    for (unsigned char ch : bytes) o << (unsigned)ch << ' ';
  } else if (op == "deserialize_pt_tensor") {  // bytes... -> ndim dims... ints...
    std::string bytes(a.size(), 0);
    for (size_t i = 0; i < a.size(); i++) bytes[i] = (char)to_sz(a[i]);
    auto t = R.cs.deserialize_plaintext_tensor(bytes);
    o << t.ndim() << ' ';
    for (size_t i = 0; i < t.ndim(); i++) o << t.shape()[i] << ' ';
    t.flatten();
    for (size_t i = 0; i < t.num_elements(); i++) { put(o, *t.at(i)); delete t.at(i); }
  } else {
    throw std::invalid_argument("unknown op " + op);
  }
  return o.str();
}

// ---- SoA limb export (same layout as the product C ABI: limb-major, element-minor, little-endian u32 limbs,
//      b in two's complement over its limb width).
void export_int(const Mpz &z, uint32_t *dst, size_t nl, size_t stride, bool twos) {
  std::vector<uint32_t> tmp(nl + 2, 0);
  size_t cnt = 0;
  if (z.nbits() > 32 * nl) throw std::range_error("export: integer does not fit");
  mpz_export(tmp.data(), &cnt, -1, 4, 0, 0, (mpz_srcptr)z);
  if (twos && z.sgn() < 0) {
    uint64_t c = 1;
    for (size_t i = 0; i < nl; i++) { uint64_t v = (uint64_t)(uint32_t)~tmp[i] + c; tmp[i] = (uint32_t)v; c = v >> 32; }
  }
  for (size_t i = 0; i < nl; i++) dst[i * stride] = tmp[i];
}
void export_form(const QFI &f, uint32_t *dst, size_t idx, size_t n, size_t na, size_t nc) {
  export_int(f.a(), dst + idx, na, n, false);
  export_int(f.b(), dst + na * n + idx, na, n, true);
  export_int(f.c(), dst + 2 * na * n + idx, nc, n, false);
}

} // namespace

extern "C" {

void *ref_new(unsigned nbits, unsigned k) {
  try { return new Ref(nbits, k); } catch (...) { return nullptr; }
}
// the compact variant of CL_HSM2k (CPUCryptoSystem(security_level, k, compact), cpu_cryptosystem.hpp:33)
void *ref_new_compact(unsigned nbits, unsigned k, int compact) {
  try { return new Ref(nbits, k, compact != 0); } catch (...) { return nullptr; }
}
void ref_free(void *h) { delete (Ref *)h; }
void ref_free_str(char *s) { free(s); }

// Returns a malloc'd string: "OK <numbers...>" or "ERR <what>".
char *ref_call(void *h, const char *op, const char *args) {
  try { return dup("OK " + call(*(Ref *)h, op, parse(args))); }
  catch (const std::exception &e) { return dup(std::string("ERR ") + e.what()); }
}

int ref_num_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
void ref_set_threads(int n) {
#ifdef _OPENMP
  omp_set_num_threads(n);
#else
  (void)n;
#endif
}

// ---------------------------------------------------------------------------------------------
// Bulk workloads for bench.py (cpu_baseline / --impl reference) and for full-size parity checks.
// Each builds its inputs exactly as benchmarks/local.cpp does (pt[i] = i+1, encrypt_tensor),
// runs the reference's own tensor entry point, optionally exports inputs/outputs as SoA limbs,
// and returns the wall time of the timed region in seconds (<0 on error).
// Buffers may be NULL (no export).  Layout per form: a[na] b[na] c[nc], per ciphertext c1 then c2:
// buffer is [2][na+na+nc][count] u32.
// ---------------------------------------------------------------------------------------------
static void export_cts(Tensor<CS::CipherText *> &t, size_t n, uint32_t *buf, size_t na, size_t nc) {
  if (!buf) return;
  size_t fl = 2 * na + nc;
  for (size_t i = 0; i < n; i++) {
    export_form(t.at(i)->c1(), buf, i, n, na, nc);
    export_form(t.at(i)->c2(), buf + fl * n, i, n, na, nc);
  }
}
static Tensor<CS::CipherText *> enc_pattern(Ref &R, size_t rows, size_t cols, size_t offset) {
  auto pt = make_tensor<CS::PlainText>(rows, cols); pt.flatten();
  for (size_t i = 0; i < rows * cols; i++) pt.at(i) = new CS::PlainText((unsigned long)(i + 1 + offset));
  pt.reshape({rows, cols});
  auto ct = R.cs.encrypt_tensor(*R.pk, pt);
  pt.flatten();
  for (size_t i = 0; i < rows * cols; i++) delete pt.at(i);
  return ct;
}
static void free_cts(Tensor<CS::CipherText *> &t) { t.flatten(); for (size_t i = 0; i < t.num_elements(); i++) delete t.at(i); }

// matadd: ct1, ct2 = encrypt_tensor(i+1) (two calls, as local.cpp), then `iters` chained adds res = res + ct2
// (res starts as ct1 + ct2 like benchmarks/local.cpp:65-138).  Exports ct1, ct2 and the final result.
double ref_bulk_matadd(void *h, unsigned rows, unsigned cols, unsigned iters,
                       uint32_t *in1, uint32_t *in2, uint32_t *out, unsigned na, unsigned nc) {
  try {
    Ref &R = *(Ref *)h; size_t n = (size_t)rows * cols;
    auto ct1 = enc_pattern(R, rows, cols, 0), ct2 = enc_pattern(R, rows, cols, 0);
    ct1.flatten(); ct2.flatten(); export_cts(ct1, n, in1, na, nc); export_cts(ct2, n, in2, na, nc);
    ct1.reshape({rows, cols}); ct2.reshape({rows, cols});
    auto t0 = std::chrono::steady_clock::now();
    auto res = R.cs.add_ciphertext_tensors(*R.pk, ct1, ct2);
    for (unsigned it = 1; it < iters; it++) {
      auto nxt = R.cs.add_ciphertext_tensors(*R.pk, res, ct2);
      free_cts(res); res = nxt;
    }
    double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    res.flatten(); export_cts(res, n, out, na, nc);
    free_cts(res); free_cts(ct1); free_cts(ct2);
    return dt;
  } catch (...) { return -1.0; }
}

// wire format: ct = encrypt_tensor(i+1) [rows x cols]; `iters` x serialize_ciphertext_tensor then `iters` x
// deserialize_ciphertext_tensor of the bytes (reference cpu_cryptosystem.inl:322-520).  Exports the ciphertexts (SoA limbs) and
// the serialised bytes (wire, capacity cap; *wire_len = length) so that the device kernels can be checked against them.
// Returns the serialise time, *deser_s the deserialise time (seconds, all iterations).
double ref_bulk_wire(void *h, unsigned rows, unsigned cols, unsigned iters, uint32_t *cts, unsigned na, unsigned nc,
                     unsigned char *wire, size_t cap, size_t *wire_len, double *deser_s) {
  try {
    Ref &R = *(Ref *)h; size_t n = (size_t)rows * cols;
    auto ct = enc_pattern(R, rows, cols, 0);
    ct.flatten(); if (cts) export_cts(ct, n, cts, na, nc); ct.reshape({rows, cols});
    std::string bytes;
    auto t0 = std::chrono::steady_clock::now();
    for (unsigned it = 0; it < iters; it++) bytes = R.cs.serialize_ciphertext_tensor(ct);
    double ds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    t0 = std::chrono::steady_clock::now();
    for (unsigned it = 0; it < iters; it++) { auto back = R.cs.deserialize_ciphertext_tensor(bytes); free_cts(back); }
    if (deser_s) *deser_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (wire_len) *wire_len = bytes.size();
    if (wire && bytes.size() <= cap) memcpy(wire, bytes.data(), bytes.size());
    free_cts(ct);
    return ds;
  } catch (...) { return -1.0; }
}

// scal_matmul: ct = encrypt_tensor(i+1) [n x m], scalars s[j] (decimal strings, m*p, row-major) or the
// benchmark pattern i+1 when scal == NULL; `iters` chained calls ct <- scal_ciphertext_tensors(pk, s, ct)
// needs p == m when iters > 1.  Exports the input ct, the `zero` ciphertext used by the LAST call is not
// observable, so parity checks use ref_call("scal_matmul") or decrypt; here we export in/out.
double ref_bulk_scal_matmul(void *h, unsigned n, unsigned m, unsigned p, unsigned iters, const char *scal,
                            uint32_t *in, uint32_t *out, unsigned na, unsigned nc) {
  try {
    Ref &R = *(Ref *)h;
    auto ct = enc_pattern(R, n, m, 0);
    ct.flatten(); export_cts(ct, (size_t)n * m, in, na, nc); ct.reshape({n, m});
    auto s = make_tensor<CS::PlainText>(m, p); s.flatten();
    if (scal) { auto v = parse(scal); for (size_t i = 0; i < (size_t)m * p; i++) s.at(i) = new CS::PlainText(v[i]); }
    else for (size_t i = 0; i < (size_t)m * p; i++) s.at(i) = new CS::PlainText((unsigned long)(i + 1));
    s.reshape({m, p});
    auto t0 = std::chrono::steady_clock::now();
    auto res = R.cs.scal_ciphertext_tensors(*R.pk, s, ct);
    for (unsigned it = 1; it < iters; it++) {
      auto nxt = R.cs.scal_ciphertext_tensors(*R.pk, s, res);
      free_cts(res); res = nxt;
    }
    double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    res.flatten(); export_cts(res, (size_t)n * p, out, na, nc);
    free_cts(res); free_cts(ct);
    s.flatten(); for (size_t i = 0; i < (size_t)m * p; i++) delete s.at(i);
    return dt;
  } catch (...) { return -1.0; }
}

// N sequential single encryptions (own r each, drawn in index order) of m_i = random_mpz(2^k) -- config 3 mode B.
// Exports plaintexts (k-bit, nm limbs each, [limb][elt]), randomness r_i (nr limbs) and the ciphertexts.
double ref_bulk_encrypt(void *h, unsigned count, uint32_t *m_out, unsigned nm, uint32_t *r_out, unsigned nr,
                        uint32_t *ct_out, unsigned na, unsigned nc) {
  try {
    Ref &R = *(Ref *)h; const auto &H = R.cs.get_hsm2k();
    std::vector<Mpz> ms(count), rs(count);
    for (unsigned i = 0; i < count; i++) ms[i] = R.cs.get_rand_gen().random_mpz(H.cleartext_bound());
    for (unsigned i = 0; i < count; i++) rs[i] = R.cs.get_rand_gen().random_mpz(H.encrypt_randomness_bound());
    std::vector<CS::CipherText> cts(count);
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(static)
    for (unsigned i = 0; i < count; i++)
      cts[i] = CS::CipherText(H, *R.pk, BICYCL::CL_HSM2k::ClearText(H, ms[i]), rs[i]);
    double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    size_t fl = 2 * na + nc;
    for (unsigned i = 0; i < count; i++) {
      if (m_out) export_int(ms[i], m_out + i, nm, count, false);
      if (r_out) export_int(rs[i], r_out + i, nr, count, false);
      if (ct_out) { export_form(cts[i].c1(), ct_out, i, count, na, nc); export_form(cts[i].c2(), ct_out + fl * count, i, count, na, nc); }
    }
    return dt;
  } catch (...) { return -1.0; }
}

// Threshold: part_decrypt_tensor with one share over `count` ciphertexts given as decimal text is too slow
// for 2^16; instead encrypt count plaintexts with encrypt_tensor(i+1), run part_decrypt_tensor(share) and export.
double ref_bulk_part_decrypt(void *h, unsigned count, const char *share, uint32_t *ct_out, uint32_t *d_out,
                             unsigned na, unsigned nc) {
  try {
    Ref &R = *(Ref *)h; Mpz sh{std::string(share)};
    auto ct = enc_pattern(R, count, 1, 0);
    ct.flatten(); export_cts(ct, count, ct_out, na, nc); ct.reshape({count, 1});
    auto t0 = std::chrono::steady_clock::now();
    auto ds = R.cs.part_decrypt_tensor(sh, ct);
    double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    ds.flatten();
    for (unsigned i = 0; i < count; i++) { if (d_out) export_form(*ds.at(i), d_out, i, count, na, nc); delete ds.at(i); }
    free_cts(ct);
    return dt;
  } catch (...) { return -1.0; }
}

// ---------------------------------------------------------------------------------------------
// "Given-input" bulk calls: the reference's own tensor entry points on SoA limb buffers supplied by the
// caller (full-size parity at the configurations bench.py times: matmul at p = 256, encryption /
// threshold decryption at 2^16).  Layouts as above.
// ---------------------------------------------------------------------------------------------
static Mpz import_int(const uint32_t *src, size_t nl, size_t stride, bool twos) {
  std::vector<uint32_t> tmp(nl);
  for (size_t i = 0; i < nl; i++) tmp[i] = src[i * stride];
  bool neg = twos && (tmp[nl - 1] >> 31);
  if (neg) {
    uint64_t c = 1;
    for (size_t i = 0; i < nl; i++) { uint64_t v = (uint64_t)(uint32_t)~tmp[i] + c; tmp[i] = (uint32_t)v; c = v >> 32; }
  }
  mpz_t raw;
  mpz_init(raw);
  mpz_import(raw, nl, -1, 4, 0, 0, tmp.data());
  Mpz z(std::move(raw));
  if (neg) z.neg();
  return z;
}
static QFI import_form(const uint32_t *buf, size_t idx, size_t n, size_t na, size_t nc) {
  return QFI(import_int(buf + idx, na, n, false), import_int(buf + na * n + idx, na, n, true),
             import_int(buf + 2 * na * n + idx, nc, n, false), true);
}
static Tensor<CS::CipherText *> import_cts(const uint32_t *buf, size_t rows, size_t cols, size_t na, size_t nc) {
  size_t n = rows * cols, fl = 2 * na + nc;
  auto t = make_tensor<CS::CipherText>(rows, cols); t.flatten();
  for (size_t i = 0; i < n; i++) t.at(i) = new CS::CipherText(import_form(buf, i, n, na, nc), import_form(buf + fl * n, i, n, na, nc));
  t.reshape({rows, cols});
  return t;
}

// scal_ciphertext_tensors (matmul branch) on given ciphertexts ct[n*m] and scalars (decimal text, m*p).  The call
// draws its `zero` from the RandGen: seed it first (ref_call set_seed) to know which.
double ref_bulk_scal_matmul_given(void *h, unsigned n, unsigned m, unsigned p, const char *scal, const uint32_t *ct_in,
                                  uint32_t *out, unsigned na, unsigned nc) {
  try {
    Ref &R = *(Ref *)h;
    auto ct = import_cts(ct_in, n, m, na, nc);
    auto s = make_tensor<CS::PlainText>(m, p); s.flatten();
    auto v = parse(scal);
    for (size_t i = 0; i < (size_t)m * p; i++) s.at(i) = new CS::PlainText(v[i]);
    s.reshape({m, p});
    auto t0 = std::chrono::steady_clock::now();
    auto res = R.cs.scal_ciphertext_tensors(*R.pk, s, ct);
    double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    res.flatten(); export_cts(res, (size_t)n * p, out, na, nc);
    free_cts(res); free_cts(ct);
    s.flatten(); for (size_t i = 0; i < (size_t)m * p; i++) delete s.at(i);
    return dt;
  } catch (...) { return -1.0; }
}

// The same matmul from the reference's exact primitives: out[i,k] = zero o prod_j nupow(ct[i,j], s[j,k]) with
// ClassGroup::nupow / nucomp (qfi.inl:1441-1477) -- the value scal_ciphertext_tensors is meant to return; its
// qfi_nupow is wrong for some scalars >= 2^65.
double ref_bulk_matmul_exact(void *h, unsigned n, unsigned m, unsigned p, const char *scal, const uint32_t *ct_in,
                             const uint32_t *zero_in, uint32_t *out, unsigned na, unsigned nc) {
  try {
    Ref &R = *(Ref *)h; const auto &H = R.cs.get_hsm2k(); const BICYCL::ClassGroup &G = H.Cl_G();
    size_t fl = 2 * na + nc, nb = (size_t)n * m, no = (size_t)n * p;
    auto v = parse(scal);
    std::vector<QFI> base(2 * nb);
    for (size_t i = 0; i < nb; i++) { base[i] = import_form(ct_in, i, nb, na, nc); base[nb + i] = import_form(ct_in + fl * nb, i, nb, na, nc); }
    QFI z[2] = {import_form(zero_in, 0, 1, na, nc), import_form(zero_in + fl, 0, 1, na, nc)};
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 4)
    for (size_t q = 0; q < 2 * no; q++) {
      size_t y = q / no, idx = q % no, i = idx / p, k = idx % p;
      QFI acc = z[y], t;
      for (size_t j = 0; j < m; j++) { G.nupow(t, base[y * nb + i * m + j], v[j * p + k]); G.nucomp(acc, acc, t); }
      export_form(acc, out + y * fl * no, idx, no, na, nc);
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  } catch (...) { return -1.0; }
}

// N single encryptions with the caller's plaintexts and randomness (config 3 mode B on given inputs)
double ref_bulk_encrypt_given(void *h, unsigned count, const uint32_t *m_in, unsigned nm, const uint32_t *r_in, unsigned nr,
                              uint32_t *ct_out, unsigned na, unsigned nc) {
  try {
    Ref &R = *(Ref *)h; const auto &H = R.cs.get_hsm2k();
    std::vector<CS::CipherText> cts(count);
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(static)
    for (unsigned i = 0; i < count; i++)
      cts[i] = CS::CipherText(H, *R.pk, BICYCL::CL_HSM2k::ClearText(H, import_int(m_in + i, nm, count, false)), import_int(r_in + i, nr, count, false));
    double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    size_t fl = 2 * na + nc;
    for (unsigned i = 0; i < count; i++) { export_form(cts[i].c1(), ct_out, i, count, na, nc); export_form(cts[i].c2(), ct_out + fl * count, i, count, na, nc); }
    return dt;
  } catch (...) { return -1.0; }
}

// part_decrypt_tensor with one share on given ciphertexts
double ref_bulk_part_decrypt_given(void *h, unsigned count, const char *share, const uint32_t *ct_in, uint32_t *d_out,
                                   unsigned na, unsigned nc) {
  try {
    Ref &R = *(Ref *)h; Mpz sh{std::string(share)};
    auto ct = import_cts(ct_in, count, 1, na, nc);
    auto t0 = std::chrono::steady_clock::now();
    auto ds = R.cs.part_decrypt_tensor(sh, ct);
    double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    ds.flatten();
    for (unsigned i = 0; i < count; i++) { export_form(*ds.at(i), d_out, i, count, na, nc); delete ds.at(i); }
    free_cts(ct);
    return dt;
  } catch (...) { return -1.0; }
}

// combine_part_decryption_results_tensor: ct[count], ds = [t][form batch] -> plaintexts [nm][count]
double ref_bulk_combine_given(void *h, unsigned count, unsigned t, const uint32_t *ct_in, const uint32_t *ds_in, uint32_t *m_out,
                              unsigned nm, unsigned na, unsigned nc) {
  try {
    Ref &R = *(Ref *)h; size_t fl = 2 * na + nc;
    auto ct = import_cts(ct_in, count, 1, na, nc);
    Vector<Tensor<CS::PartDecryptionResult *>> pd;
    for (unsigned j = 0; j < t; j++) {
      auto d = make_tensor<CS::PartDecryptionResult>(count, 1); d.flatten();
      for (unsigned i = 0; i < count; i++) d.at(i) = new CS::PartDecryptionResult(import_form(ds_in + (size_t)j * fl * count, i, count, na, nc));
      d.reshape({count, 1});
      pd.push_back(d);
    }
    auto t0 = std::chrono::steady_clock::now();
    auto pt = R.cs.combine_part_decryption_results_tensor(ct, pd);
    double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    pt.flatten();
    for (unsigned i = 0; i < count; i++) { export_int(*pt.at(i), m_out + i, nm, count, false); delete pt.at(i); }
    for (auto &d : pd) { d.flatten(); for (unsigned i = 0; i < count; i++) delete d.at(i); }
    free_cts(ct);
    return dt;
  } catch (...) { return -1.0; }
}

} // extern "C"
