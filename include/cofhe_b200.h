/* cofhe_b200 -- C ABI of the B200-native backend for CoFHE's data-parallel ciphertext hot path
 * (batched arithmetic on binary quadratic forms of a class group: CL_HSM2k over an imaginary quadratic order).
 *
 * The reference has no FFI: its seam is the C++ class CPUCryptoSystem used as a template argument
 * (reference include/x86_64/cpu_cryptosystem.hpp:23-172).  The drop-in class for that seam is
 * include/cuda/cuda_cryptosystem.hpp (it replaces the reference's 3-line "#error" stub of the same path); it
 * marshals Tensor<CipherText*> into the buffers below and calls these entry points.  Each entry point cites the
 * reference code whose results it reproduces bit for bit.
 *
 * Layout.  All big integers cross this ABI as structure-of-arrays 32-bit limbs, limb-major / element-minor:
 * limb i of element e of an n-element batch is word [i*n + e].  Limbs are little endian; `b` coefficients are
 * two's complement over their limbs.  A batch of forms is  a[na] | b[na] | c[nc]  = (2*na+nc) limb rows of n words;
 * a batch of ciphertexts is c1-forms followed by c2-forms: [2][2*na+nc][n].  (na, nc) come from
 * cofhe_b200_shape(): na = ceil((bits(|D|)/2 + 1)/32) holds every reduced a and b; nc = ceil(bits(|D|)/32)
 * holds every reduced c.  Exponents/scalars are magnitudes [el][n] plus an optional sign byte per element.
 *
 * Pointers.  Entry points without suffix take DEVICE pointers and a CUDA stream (cudaStream_t passed as
 * void*; NULL = legacy default stream) and are asynchronous.  Entry points ending in _host take HOST
 * pointers, do the transfers on an internal stream and return when the result is in the output buffer.
 *
 * Threads and streams.  A context may be shared: every entry point takes the context's lock for the time it issues
 * work, and because all calls of a context share its scratch buffers, a call on a stream other than the previous
 * call's stream is ordered after it on the device (event wait).  Calls on ONE stream from one thread at a time pay
 * nothing for this; callers that want concurrency between streams use one context per stream (contexts are cheap:
 * the fixed-base tables, ~0.6 MB at the reference's default size, are the only sizeable per-context state).
 *
 * Errors.  Every call returns a COFHE_B200_* code; nothing throws across this boundary.  Arithmetic failures
 * of individual elements (operand not a form of this discriminant, value outside the limb shape) are
 * sticky per context: cofhe_b200_status() synchronises the stream and reports/clears them; _host calls
 * report them directly.  There is NO CPU fallback: without a CUDA device every call fails with
 * COFHE_B200_ECUDA.
 */
#ifndef COFHE_B200_H
#define COFHE_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define COFHE_B200_OK 0
#define COFHE_B200_EINVAL 1 /* bad argument / shape mismatch (reference: std::invalid_argument) */
#define COFHE_B200_ECUDA 2  /* CUDA runtime failure, or no device */
#define COFHE_B200_EARITH 3 /* an element failed: not a form of this discriminant / does not fit the limb shape */
#define COFHE_B200_ENOMEM 4
#define COFHE_B200_ESTATE 5 /* scheme parameters / public key not set for an operation that needs them */
#define COFHE_B200_ETOOBIG 6 /* batch larger than the fixed scratch of a _static entry point (COFHE_B200_STATIC_MAX_N) */

typedef struct cofhe_b200_ctx cofhe_b200_ctx;

/* Library self-description: "cofhe_b200 <version> sm_100a". */
const char *cofhe_b200_version(void);

/* Class group of discriminant D = -abs_delta (nd limbs) on CUDA device `device`.
 * Replaces BICYCL::ClassGroup(D) (reference thirdparty/bicycl/src/bicycl/qfi.inl:1293-1331). */
int cofhe_b200_create(cofhe_b200_ctx **ctx, int device, const uint32_t *abs_delta, uint32_t nd);
void cofhe_b200_destroy(cofhe_b200_ctx *ctx);
/* Limb shape of form buffers for this context. */
int cofhe_b200_shape(const cofhe_b200_ctx *ctx, uint32_t *na, uint32_t *nc, uint32_t *nd);
/* Synchronise `stream`, return COFHE_B200_EARITH if any element failed since the last call, then clear. */
int cofhe_b200_status(cofhe_b200_ctx *ctx, void *stream);
/* Number of kernel launches issued through this context so far (bench.py reports it as gpu_launches). */
uint64_t cofhe_b200_launch_count(const cofhe_b200_ctx *ctx);
/* Coefficient limbs of the hand-written kernel instantiation serving this discriminant -- register-resident (one thread
 * per form) or cooperative (one lane group per form) -- or 0 when only the generic integer path exists. */
int cofhe_b200_fast_limbs(const cofhe_b200_ctx *ctx);
/* Coefficient limbs of the cooperative instantiation for this discriminant (0: none).  It serves the widths without a
 * register instantiation (2053-bit discriminants) and, with the option "coop" = 1, every form operation. */
int cofhe_b200_coop_limbs(const cofhe_b200_ctx *ctx);
/* Coefficient limbs of the register-resident instantiation (one thread per form) for this discriminant (0: none).  The
 * zero-copy variant of cofhe_b200_add_ciphertexts_host exists on these widths. */
int cofhe_b200_reg_limbs(const cofhe_b200_ctx *ctx);
/* Tuning knobs, -1 = automatic: "mm_w" (window bits of the matmul multi-exponentiation, 2..8), "mm_chunks" (split of
 * the inner dimension), "mm_block" (128/256/512 threads), "fb_w" (fixed-base window bits, before set_bases),
 * "host_staged", "host_loadc", "host_chunks", "host_trace" (variants of the host-pointer add), "coop" (1: cooperative
 * kernels for every form operation, 0: never), "coop_mb" (3/4/5: resident blocks per SM the cooperative composition is compiled for; 14: 4 with block barriers
 * between the phases, 24: the same in 256-thread blocks; default 14 / 24 by batch size), "coop_pb" (128 / 256: threads
 * per block of the cooperative exponentiation and table kernels), "fast_block" (128 / 256 / 512: threads per block of the
 * register composition kernel; default 128, 256 from 200 000 compositions per launch), "fast_flat" (1 / 0: branch-free
 * scalar Euclid rounds in that kernel; default on up to 49 152 compositions per launch).  They change how the work
 * is laid out, never the results.  The COFHE_B200_<NAME> environment variables give the defaults at creation. */
int cofhe_b200_set_option(cofhe_b200_ctx *ctx, const char *name, long value);

/* out[e] = f1[e] o f2[e]   (negf2 != 0: f1[e] o f2[e]^-1), reduced.
 * ClassGroup::nucomp / nucompinv (qfi.inl:1441-1451 -> QFI::nucomp qfi.inl:683-818). */
int cofhe_b200_nucomp(cofhe_b200_ctx *ctx, size_t n, const uint32_t *f1, const uint32_t *f2, uint32_t *out, int negf2,
                      void *stream);
/* out[e] = f[e]^2, reduced.  ClassGroup::nudupl (qfi.inl:1453-1471 -> QFI::nudupl qfi.inl:861-955). */
int cofhe_b200_nudupl(cofhe_b200_ctx *ctx, size_t n, const uint32_t *f, uint32_t *out, void *stream);
/* out[e] = the reduced form equivalent to f[e] (QFI::reduction, qfi.inl:523-548).  Input rows may be wider
 * than the context shape (unreduced forms): a/b have na_in limbs, c has nc_in limbs; output uses the context shape. */
int cofhe_b200_reduce(cofhe_b200_ctx *ctx, size_t n, const uint32_t *f, uint32_t na_in, uint32_t nc_in, uint32_t *out,
                      void *stream);
/* out[e] = f[e]^x.  per_elem != 0: x = exp[.][e] (el limb rows) with sign expneg[e] (NULL = all positive);
 * per_elem == 0: one shared exponent exp[0..el) (device memory) with sign shared_neg.
 * ClassGroup::nupow (qfi.inl:1473-1477 -> QFI::nupow qfi.inl:977-1064): x == 0 gives the neutral form, x < 0 gives
 * QFI::neg of f^|x|. */
int cofhe_b200_nupow(cofhe_b200_ctx *ctx, size_t n, const uint32_t *f, const uint32_t *exp, uint32_t el, int per_elem,
                     const uint8_t *expneg, int shared_neg, uint32_t *out, void *stream);

/* out[e] = ct1[e] + ct2[e] (homomorphic add, no re-randomisation): c1 o c1', c2 o c2'.
 * CPUCryptoSystem::add_ciphertext_tensors / add_ciphertext_vectors
 * (reference include/x86_64/cpu_cryptosystem_tensor_ops.inl:197-267, cpu_cryptosystem_vector_ops.inl:58-120). */
int cofhe_b200_add_ciphertexts(cofhe_b200_ctx *ctx, size_t n, const uint32_t *ct1, const uint32_t *ct2, uint32_t *out,
                               void *stream);
/* cofhe_b200_add_ciphertexts for callers that cannot tolerate an allocation inside the call (CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments, same kernels, bit-identical results; the call issues one memset and the kernels on
 * `stream` and calls no cudaMalloc / cudaFree or any other allocating routine.  The scratch that cofhe_b200_add_ciphertexts
 * grows on demand (the list of elements handed to the robust path, and its counters) lives in ONE fixed array of device
 * memory inside the library instead: 2 MiB + 256 bytes on every GPU that loads the library, held until the process exits.
 *   - n > COFHE_B200_STATIC_MAX_N is refused with COFHE_B200_ETOOBIG (nothing is launched); split the batch.
 *   - the array is shared by every call, every context and every stream of the process on that GPU: calls that overlap
 *     on the GPU (different streams or contexts without ordering between them) race on it.  Calls on one stream, or
 *     ordered by events, are fine -- each call clears the counters it needs at its start, on its stream.
 * The context itself (cofhe_b200_create) allocates as before; create it ahead of the region that must not allocate. */
#define COFHE_B200_STATIC_MAX_N 262144u /* ciphertexts per call */
int cofhe_b200_add_ciphertexts_static(cofhe_b200_ctx *ctx, size_t n, const uint32_t *ct1, const uint32_t *ct2, uint32_t *out,
                                      void *stream);
/* Same, on HOST buffers, blocking (the call a CryptoSystem backend makes per add_ciphertext_tensors).  Page-locked
 * buffers (cudaHostAlloc / cudaHostRegister / cofhe_b200_host_alloc) are read and written by the kernels directly over
 * PCIe, without staging (discriminants with a register-resident instantiation, cofhe_b200_reg_limbs() != 0; the others
 * are staged); on that path only the a and b rows of the operands are read (c = (b^2 - Delta) / 4a is
 * recomputed on the device, also for the elements that take the robust path), all rows of `out` are written.
 * Pageable buffers are staged through device memory by pipelined copies (all rows are read). */
int cofhe_b200_add_ciphertexts_host(cofhe_b200_ctx *ctx, size_t n, const uint32_t *ct1, const uint32_t *ct2,
                                    uint32_t *out);

/* out[e] = ct[e]^x (both forms), no re-randomisation: the 1-D branch of CPUCryptoSystem::scal_ciphertext_tensors and
 * scal_ciphertext_vector (reference include/x86_64/cpu_cryptosystem_tensor_ops.inl:280-340, cpu_cryptosystem_vector_ops.inl
 * :122-242) with per_elem != 0, negate_ciphertext_tensor (:135-195) with the shared exponent 2^k - 1.  Exponent
 * arguments as cofhe_b200_nupow. */
int cofhe_b200_scal_ciphertexts(cofhe_b200_ctx *ctx, size_t n, const uint32_t *ct, const uint32_t *exp, uint32_t el, int per_elem,
                                const uint8_t *expneg, int shared_neg, uint32_t *out, void *stream);

/* out[e] = (ct.c1[e] o h^r, ct.c2[e] o pk^r), r = r[.][e] per element (shared_r == 0) or one r[0..el) for the batch: the
 * re-randomisation that the reference's single-ciphertext add / scal / negate apply after the homomorphic operation
 * (CL_HSM_CipherText constructors, reference thirdparty/bicycl/src/bicycl/CL_HSM_utils.inl:188-215 and :218-246;
 * CPUCryptoSystem::add_ciphertexts / scal_ciphertext / negate_ciphertext, include/x86_64/cpu_cryptosystem.inl:26-34, 95-98).
 * add = add_ciphertexts + rerandomize, scal = scal_ciphertexts + rerandomize.  Needs set_bases; r in DEVICE memory. */
int cofhe_b200_rerandomize(cofhe_b200_ctx *ctx, size_t n, const uint32_t *ct, const uint32_t *r, uint32_t el, int shared_r, uint32_t *out,
                           void *stream);

/* Beaver-triple ciphertext x ciphertext multiplication, the device side of
 * SMPCCipherTextMultiplier::handle_vector_ciphertext_mul (reference include/smpc/ciphertext_multiplications.hpp:115-160),
 * as two fused pipelines around the network round that opens the masked operands (needs set_scheme / set_bases):
 *   open : out1 = ct1 + negate(a), out2 = ct2 + negate(b)             (2 negate + 2 add, :127-130)
 *   close: out  = pt1 * b + pt2 * a + c + encrypt_tensor(pt1 * pt2)   (2 scal + encrypt + 3 add, :133-139)
 * pt1, pt2: the opened plaintexts [ml][n]; r: the ONE randomness of the encrypt_tensor call (device memory, el limbs, as cofhe_b200_encrypt with shared_r);
 * (a, b, c): the triple, ciphertext batches.  Everything stays in device memory between the steps. */
int cofhe_b200_beaver_open(cofhe_b200_ctx *ctx, size_t n, const uint32_t *ct1, const uint32_t *ct2, const uint32_t *a,
                           const uint32_t *b, uint32_t *out1, uint32_t *out2, void *stream);
int cofhe_b200_beaver_close(cofhe_b200_ctx *ctx, size_t n, const uint32_t *pt1, const uint32_t *pt2, uint32_t ml, const uint32_t *a,
                            const uint32_t *b, const uint32_t *c, const uint32_t *r, uint32_t el, uint32_t *out, void *stream);

/* Beaver triplets: out = encrypt_tensor of the [n, 3] plaintext tensor (a_i, b_i, a_i * b_i), 3n ciphertexts in row-major
 * order, ONE randomness r (device memory, el limbs) for the whole tensor: BeaversTripletGenerator::generate (reference
 * include/smpc/beavers_triplet_generation.hpp:17-37), which fills the client's triplet cache (include/smpc/smpc_client.hpp
 * :17,218).  a, b: plaintexts [ml][n] drawn by the caller's RandGen (cpu_cryptosystem.inl:36-47). */
int cofhe_b200_beaver_triplets(cofhe_b200_ctx *ctx, size_t n, const uint32_t *a, const uint32_t *b, uint32_t ml, const uint32_t *r, uint32_t el,
                               uint32_t *out, void *stream);
/* Ciphertext x ciphertext MATRIX product around the Beaver multiplication (reference
 * include/smpc/ciphertext_multiplications.hpp:51-110), ct1: n x m, ct2: m x p:
 *   expand: out1[(i*m + j)*p + k] = ct1[i*m + j], out2[(i*m + j)*p + k] = ct2[j*p + k]   (:58-70; n*m*p elements each) --
 *           the two operands of cofhe_b200_beaver_open / _close on n*m*p elements;
 *   reduce: out[i*p + k] = init o prod_j prod[(i*m + j)*p + k]   (:79-100: zero = encrypt(pk, 0), then the direct nucomp
 *           accumulation over j), as a balanced tree: log2(m) batched composition launches. */
int cofhe_b200_ct_matmul_expand(cofhe_b200_ctx *ctx, size_t n, size_t m, size_t p, const uint32_t *ct1, const uint32_t *ct2, uint32_t *out1,
                                uint32_t *out2, void *stream);
int cofhe_b200_ct_matmul_reduce(cofhe_b200_ctx *ctx, size_t n, size_t m, size_t p, const uint32_t *prod, const uint32_t *init, uint32_t *out,
                                void *stream);

/* ---- Binary tensor wire format <-> SoA batches on the device.
 * The format is the one of CPUCryptoSystem::serialize_ciphertext_tensor / deserialize_ciphertext_tensor (reference
 * include/x86_64/cpu_cryptosystem.inl:322-391 and :393-520): u32 ndim | u32 dims[ndim] | u64 table[6n] | magnitudes;
 * table entries are byte offsets into the magnitude area with bit 63 set for integers <= 0; an integer x owns
 * mpz_sizeinbase(x,2)/8 + 1 little-endian bytes.  These calls replace the per-element mpz_import / mpz_export loops
 * either side of every network operation (reference include/node/compute_request_handler.hpp:414-431, 562-578).
 *
 * cofhe_b200_wire_header: parse ndim / dims / element count from the first bytes of a HOST buffer (no device work).
 * cofhe_b200_wire_bound : upper bound of the serialised size of n ciphertexts of this context's shape.
 * cofhe_b200_wire_to_cts: wire bytes in DEVICE memory (or page-locked host memory) -> ciphertext batch [2][2na+nc][n].
 *                         Malformed input (integer wider than its limbs, offsets out of range) raises the sticky
 *                         element status (COFHE_B200_EARITH at the next cofhe_b200_status).
 * cofhe_b200_cts_to_wire: ciphertext batch -> wire bytes in device memory; blocking for the size (one 8-byte read-back),
 *                         *written = bytes produced, or needed when `cap` is too small (COFHE_B200_EINVAL). */
int cofhe_b200_wire_header(const uint8_t *wire, size_t bytes, uint32_t *ndim, uint32_t *dims, uint32_t max_dims, size_t *n);
/* The same layout carries the reference's other two tensor formats; `kind` selects the integers per element:
 *   COFHE_B200_WIRE_CIPHERTEXTS (6: c1.a c1.b c1.c c2.a c2.b c2.c; batch [2][2na+nc][n])
 *   COFHE_B200_WIRE_FORMS       (3: a b c; serialize / deserialize_part_decryption_result_tensor, reference
 *                                include/x86_64/cpu_cryptosystem.inl:510-635; batch [2na+nc][n])
 *   COFHE_B200_WIRE_PLAINTEXTS  (1; serialize / deserialize_plaintext_tensor, cpu_cryptosystem.inl:229-318; magnitudes [ml][n] plus
 *                                one sign byte per element -- the reference ships unreduced, possibly negative plaintexts)
 * These replace the mpz_import / mpz_export loops of include/node/compute_request_handler.hpp:414-431, 562-578 and
 * include/node/partial_decryption_request_handler.hpp:143-158.  ml / sign buffers are used by the plaintext kind only
 * (sign_out == NULL: negative values raise the element status; sign_in == NULL: all non-negative). */
#define COFHE_B200_WIRE_CIPHERTEXTS 0
#define COFHE_B200_WIRE_FORMS 1
#define COFHE_B200_WIRE_PLAINTEXTS 2
size_t cofhe_b200_wire_bound_kind(const cofhe_b200_ctx *ctx, int kind, uint32_t ndim, size_t n, uint32_t ml);
int cofhe_b200_wire_to_batch(cofhe_b200_ctx *ctx, int kind, const uint8_t *wire, size_t bytes, uint32_t ndim, size_t n, uint32_t *out,
                             uint32_t ml, uint8_t *sign_out, void *stream);
int cofhe_b200_batch_to_wire(cofhe_b200_ctx *ctx, int kind, uint32_t ndim, const uint32_t *dims, size_t n, const uint32_t *in, uint32_t ml,
                             const uint8_t *sign_in, uint8_t *wire, size_t cap, size_t *written, void *stream);
size_t cofhe_b200_wire_bound(const cofhe_b200_ctx *ctx, uint32_t ndim, size_t n);
int cofhe_b200_wire_to_cts(cofhe_b200_ctx *ctx, const uint8_t *wire, size_t bytes, uint32_t ndim, size_t n, uint32_t *out,
                           void *stream);
int cofhe_b200_cts_to_wire(cofhe_b200_ctx *ctx, uint32_t ndim, const uint32_t *dims, size_t n, const uint32_t *ct, uint8_t *wire,
                           size_t cap, size_t *written, void *stream);

/* Scalar x ciphertext matmul: out[i,k] = init o prod_j ct[i,j]^{s[j,k]}  (i < n, j < m, k < p; row-major).
 * ct: ciphertext batch of n*m elements, scal: non-negative scalars [el][m*p] (limb-major), init: ONE ciphertext
 * (the reference's `zero` = encrypt(pk, 0)), out: ciphertext batch of n*p elements.  ebits: an upper bound on the
 * bit length of the scalars (0 = 32*el); it only sizes the work, not the result.
 * CPUCryptoSystem::scal_ciphertext_tensors, matmul branch (reference
 * include/x86_64/cpu_cryptosystem_tensor_ops.inl:346-461; qfi_nupow include/x86_64/qfi.inl:1-135).  Results are those
 * of ClassGroup::nupow / nucomp, i.e. the mathematically exact ones; the reference's qfi_nupow is wrong for some
 * scalars >= 2^65 and that defect is not reproduced. */
int cofhe_b200_scal_matmul(cofhe_b200_ctx *ctx, size_t n, size_t m, size_t p, const uint32_t *ct, const uint32_t *scal,
                           uint32_t el, uint32_t ebits, const uint32_t *init, uint32_t *out, void *stream);

/* ---- CL_HSM2k scheme level --------------------------------------------------------------------------------------
 * The context becomes a CL_HSM2k instance once the scheme constants are known: message space Z/2^k and the modulus N
 * with DeltaK = -8N, Delta = -2^(2k+5) N (reference thirdparty/bicycl/src/bicycl/CL_HSM2k.inl:31-103, 620-637; setup
 * itself -- random N, class-number bound, h -- stays on the host with the reference's RandGen).
 * Fails with COFHE_B200_EINVAL when 2^(2k+5) N is not the context's |Delta|.  Host pointers. */
int cofhe_b200_set_scheme(cofhe_b200_ctx *ctx, uint32_t k, const uint32_t *N, uint32_t nN);
/* The compact variant of CL_HSM2k (reference CL_HSM2k.inl:31-39, 76-80, 255: constructor argument `compact_variant`,
 * CPUCryptoSystem(security_level, k, compact), include/x86_64/cpu_cryptosystem.hpp:33).  With on != 0 the group G of
 * the generator is Cl(DeltaK) instead of Cl(Delta): h, pk and every c1 half are forms of discriminant DeltaK = -8N
 * (they occupy the same limb rows of the layout, with leading zero limbs); c2 halves, f^m and part-decryption results
 * stay in Cl(Delta).  Every ciphertext-level entry point then composes / exponentiates form block 0 in Cl(DeltaK) and
 * form block 1 in Cl(Delta) inside the same launch; encrypt / rerandomize / part_decrypt / decrypt apply
 * cofhe_b200_lift where the reference calls from_Cl_DeltaK_to_Cl_Delta.  Call after set_scheme and before set_bases
 * (the tables are dropped).  cofhe_b200_add_ciphertexts_host is not available in this variant (COFHE_B200_ESTATE). */
int cofhe_b200_set_compact(cofhe_b200_ctx *ctx, int on);
/* out[e] = the image in Cl(Delta) of the form f[e] of Cl(DeltaK), compact variant only:
 * CL_HSM2k::from_Cl_DeltaK_to_Cl_Delta (CL_HSM2k.inl:458-474) = QFI::lift_2exp(k+1) (qfi.inl:263-271, prime_to_2exp
 * :590-610), k squarings in Cl(Delta) (raise_to_power_M, :641-648) and, where a or c of the lifted form is 3 mod 4,
 * one composition with (4, 4, 1 - DeltaK 2^(2(k-1))).  Form batches of the context shape; in place allowed. */
int cofhe_b200_lift(cofhe_b200_ctx *ctx, size_t n, const uint32_t *f, uint32_t *out, void *stream);
/* Fixed bases for encryption: the generator h and the public key pk (host pointers to one form each, context shape;
 * pk may be NULL).  Builds their window tables on the device for exponents below 2^ebits.  Replaces the
 * precomputation of CL_HSM2k (CL_HSM2k.inl:92-102: h^(2^e), h^(2^d), h^(2^(d+e))) and of CL_HSM_PublicKey
 * (CL_HSM_utils.inl:43-80).  Synchronous. */
int cofhe_b200_set_bases(cofhe_b200_ctx *ctx, const uint32_t *h, const uint32_t *pk, uint32_t ebits);
/* out[e] = (h^x[e], pk^x[e]) as a ciphertext-shaped batch, x = exps[.][e] (el limb rows, x < 2^ebits).
 * CL_HSM2k::power_of_h (CL_HSM2k.inl:318-322 -> QFI::nupow with 3 precomputed powers, qfi.inl:1162-1251) and
 * CL_HSM_PublicKey::exponentiation (CL_HSM_utils.inl:89-95). */
int cofhe_b200_fixed_base_pow(cofhe_b200_ctx *ctx, size_t n, const uint32_t *exps, uint32_t el, uint32_t *out, void *stream);
/* out[e] = the REDUCED form of f^m[e], m = m[.][e] (ml limb rows, taken modulo 2^k).  CL_HSM2k::power_of_f
 * (CL_HSM2k.inl:326-405) returns the unreduced triple; every consumer composes it, which reduces it. */
int cofhe_b200_power_of_f(cofhe_b200_ctx *ctx, size_t n, const uint32_t *m, uint32_t ml, uint32_t *out, void *stream);
/* m[.][e] = dlog_F(f[e]) in [0, 2^k), ml limb rows (CL_HSM2k::dlog_in_F, CL_HSM2k.inl:408-455; both the
 * large-message branch through QFI::kernel_representative_2exp, qfi.inl:396-435, and the small-message branch). */
int cofhe_b200_dlog(cofhe_b200_ctx *ctx, size_t n, const uint32_t *f, uint32_t *m, uint32_t ml, void *stream);
/* The float <-> plaintext codec, batched: mag[.][e] / sign[e] = CPUCryptoSystem::make_plaintext(x[e]) and
 * out[e] = get_float_from_plaintext(plaintext e) (reference include/x86_64/cpu_cryptosystem.inl:49-87, 114-122), bit for bit
 * including what GMP's default-precision mpf arithmetic does to negative and fractional values (csrc/codec.cuh).  x and out
 * are DEVICE float arrays; plaintexts are sign-magnitude ([ml][n] limbs + one sign byte per element, as the wire kernels
 * produce them; sign == NULL: non-negative).  make: a value that does not fit ml limbs (ml >= 5 always fits), NaN or
 * infinity raises the element status.  Needs set_scheme (k). */
int cofhe_b200_make_plaintexts(cofhe_b200_ctx *ctx, size_t n, const float *x, uint32_t *mag, uint32_t ml, uint8_t *sign, void *stream);
int cofhe_b200_plaintexts_to_floats(cofhe_b200_ctx *ctx, size_t n, const uint32_t *mag, uint32_t ml, const uint8_t *sign, float *out,
                                    void *stream);
/* out[e] = (h^r, pk^r o f^m[e]).  shared_r != 0: ONE exponent r[0..el) for the whole batch, as
 * CPUCryptoSystem::encrypt_tensor / encrypt_vector do (reference include/x86_64/cpu_cryptosystem_tensor_ops.inl:1-19,
 * CL_HSM_utils.inl:248-255); shared_r == 0: r[.][e] per element, as N calls of CPUCryptoSystem::encrypt
 * (cpu_cryptosystem.inl:11-14 -> CL_HSM_utils.inl:161-185).  r is drawn by the caller (host RandGen). */
int cofhe_b200_encrypt(cofhe_b200_ctx *ctx, size_t n, const uint32_t *m, uint32_t ml, const uint32_t *r, uint32_t el, int shared_r,
                       uint32_t *out, void *stream);
/* out[e] = c1[e]^share  (threshold partial decryption, one share for the whole batch; share: HOST pointer, el limbs).
 * CPUCryptoSystem::part_decrypt_tensor (cpu_cryptosystem_tensor_ops.inl:35-47 -> partDecrypt,
 * cpu_cryptosystem_distributed.inl:244-254). */
int cofhe_b200_part_decrypt(cofhe_b200_ctx *ctx, size_t n, const uint32_t *ct, const uint32_t *share, uint32_t el, int share_neg,
                            uint32_t *out, void *stream);
/* m[.][e] = dlog_F( c2[e] o (d_0[e] o d_1[e]^-1 o ... o d_{t-1}[e]^-1)^-1 ), ds = [t][form batch].
 * CPUCryptoSystem::combine_part_decryption_results_tensor (cpu_cryptosystem_tensor_ops.inl:49-73 -> finalDecrypt,
 * compute_d, compute_lambda: cpu_cryptosystem_distributed.inl:214-270). */
int cofhe_b200_combine(cofhe_b200_ctx *ctx, size_t n, uint32_t t, const uint32_t *ct, const uint32_t *ds, uint32_t *m, uint32_t ml,
                       void *stream);
/* m[.][e] = dlog_F( c2[e] o (c1[e]^sk)^-1 )  (sk: HOST pointer).  CPUCryptoSystem::decrypt_tensor
 * (cpu_cryptosystem_tensor_ops.inl:21-33 -> CL_HSM_ClearText, CL_HSM_utils.inl:124-138). */
int cofhe_b200_decrypt(cofhe_b200_ctx *ctx, size_t n, const uint32_t *ct, const uint32_t *sk, uint32_t el, uint32_t *m, uint32_t ml,
                       void *stream);

/* ---- memory helpers, so that a caller (the header class include/cuda/cuda_cryptosystem.hpp) needs nothing but this
 * header: device buffers on the context's device, pinned host buffers, copies ordered on a stream, and the
 * context's own non-blocking stream (cofhe_b200_status(ctx, stream) synchronises it and reports element errors). */
int cofhe_b200_device_count(void);
int cofhe_b200_dev_alloc(cofhe_b200_ctx *ctx, size_t bytes, void **ptr);
void cofhe_b200_dev_free(cofhe_b200_ctx *ctx, void *ptr);
int cofhe_b200_host_alloc(size_t bytes, void **ptr);
void cofhe_b200_host_free(void *ptr);
void *cofhe_b200_stream(cofhe_b200_ctx *ctx);
int cofhe_b200_upload(cofhe_b200_ctx *ctx, void *dst, const void *src, size_t bytes, void *stream);
int cofhe_b200_download(cofhe_b200_ctx *ctx, void *dst, const void *src, size_t bytes, void *stream);
int cofhe_b200_copy(cofhe_b200_ctx *ctx, void *dst, const void *src, size_t bytes, void *stream); /* device to device */
/* out[.][e] = (sign[e] ? -mag[e] : mag[e]) mod 2^k, ml_out limb rows: the reduction CPUCryptoSystem applies to a plaintext
 * before it encrypts it (to_plaintext -> CL_HSM_ClearText, reference thirdparty/bicycl/src/bicycl/CL_HSM_utils.inl:107-121),
 * for plaintext batches in the sign-magnitude layout of the wire kernels.  sign may be NULL.  Needs set_scheme. */
int cofhe_b200_plaintexts_mod2k(cofhe_b200_ctx *ctx, size_t n, const uint32_t *mag, uint32_t ml, const uint8_t *sign, uint32_t *out,
                                uint32_t ml_out, void *stream);

/* Measurement helper (no reference counterpart): sustained rate of independent 32x32->64-bit multiply-accumulates
 * (wide != 0: IMAD.WIDE.U32) or 32-bit multiply-adds (wide == 0: IMAD) over all SMs of `device`, in MACs/s.
 * bench.py uses it as the integer-multiply roofline denominator. */
int cofhe_b200_imad_probe(int device, int wide, double *macs_per_s);

#ifdef __cplusplus
}
#endif
#endif /* COFHE_B200_H */
