/* Declarations-only gmp.h shim matching the GMP 6.3.0 ABI of the system libgmp.so.10 (no GMP headers are installed in this image). Test/oracle infrastructure only. */
#ifndef __GMP_H__
#define __GMP_H__
#include <stddef.h>
#include <limits.h>
#ifdef __cplusplus
#include <iosfwd>
#include <string>
#include <istream>
#include <ostream>
#endif
#define __GNU_MP_VERSION 6
#define __GNU_MP_VERSION_MINOR 3
#define __GNU_MP_VERSION_PATCHLEVEL 0
#define GMP_LIMB_BITS 64
#define GMP_NAIL_BITS 0
#define GMP_NUMB_BITS 64
#define GMP_NUMB_MASK (~(mp_limb_t)0)
#define __MPN(x) __gmpn_##x
typedef unsigned long mp_limb_t;
typedef long mp_limb_signed_t;
typedef unsigned long mp_bitcnt_t;
typedef struct { int _mp_alloc; int _mp_size; mp_limb_t *_mp_d; } __mpz_struct;
typedef __mpz_struct mpz_t[1];
typedef mp_limb_t *mp_ptr;
typedef const mp_limb_t *mp_srcptr;
typedef long mp_size_t;
typedef long mp_exp_t;
typedef struct { int _mp_prec; int _mp_size; mp_exp_t _mp_exp; mp_limb_t *_mp_d; } __mpf_struct;
typedef __mpf_struct mpf_t[1];
typedef enum { GMP_RAND_ALG_DEFAULT = 0, GMP_RAND_ALG_LC = GMP_RAND_ALG_DEFAULT } gmp_randalg_t;
typedef struct { mpz_t _mp_seed; gmp_randalg_t _mp_alg; union { void *_mp_lc; } _mp_algdata; } __gmp_randstate_struct;
typedef __gmp_randstate_struct gmp_randstate_t[1];
typedef const __mpz_struct *mpz_srcptr;
typedef __mpz_struct *mpz_ptr;
typedef const __mpf_struct *mpf_srcptr;
typedef __mpf_struct *mpf_ptr;
#ifdef __cplusplus
extern "C" {
#endif
/* mpz */
#define mpz_init __gmpz_init
void __gmpz_init (mpz_ptr);
#define mpz_clear __gmpz_clear
void __gmpz_clear (mpz_ptr);
#define mpz_init_set __gmpz_init_set
void __gmpz_init_set (mpz_ptr, mpz_srcptr);
#define mpz_init_set_ui __gmpz_init_set_ui
void __gmpz_init_set_ui (mpz_ptr, unsigned long);
#define mpz_init_set_si __gmpz_init_set_si
void __gmpz_init_set_si (mpz_ptr, long);
#define mpz_init_set_str __gmpz_init_set_str
int __gmpz_init_set_str (mpz_ptr, const char *, int);
#define mpz_set __gmpz_set
void __gmpz_set (mpz_ptr, mpz_srcptr);
#define mpz_set_ui __gmpz_set_ui
void __gmpz_set_ui (mpz_ptr, unsigned long);
#define mpz_set_si __gmpz_set_si
void __gmpz_set_si (mpz_ptr, long);
#define mpz_set_str __gmpz_set_str
int __gmpz_set_str (mpz_ptr, const char *, int);
#define mpz_set_f __gmpz_set_f
void __gmpz_set_f (mpz_ptr, mpf_srcptr);
#define mpz_get_str __gmpz_get_str
char *__gmpz_get_str (char *, int, mpz_srcptr);
#define mpz_get_ui __gmpz_get_ui
unsigned long __gmpz_get_ui (mpz_srcptr);
#define mpz_get_si __gmpz_get_si
long __gmpz_get_si (mpz_srcptr);
#define mpz_swap __gmpz_swap
void __gmpz_swap (mpz_ptr, mpz_ptr);
#define mpz_import __gmpz_import
void __gmpz_import (mpz_ptr, size_t, int, size_t, int, size_t, const void *);
#define mpz_export __gmpz_export
void *__gmpz_export (void *, size_t *, int, size_t, int, size_t, mpz_srcptr);
#define mpz_sizeinbase __gmpz_sizeinbase
size_t __gmpz_sizeinbase (mpz_srcptr, int);
#define mpz_init2 __gmpz_init2
void __gmpz_init2 (mpz_ptr, mp_bitcnt_t);
#define mpz_size __gmpz_size
size_t __gmpz_size (mpz_srcptr);
#define mpz_getlimbn __gmpz_getlimbn
mp_limb_t __gmpz_getlimbn (mpz_srcptr, mp_size_t);
#define mpz_limbs_read __gmpz_limbs_read
mp_srcptr __gmpz_limbs_read (mpz_srcptr);
#define mpz_limbs_write __gmpz_limbs_write
mp_ptr __gmpz_limbs_write (mpz_ptr, mp_size_t);
#define mpz_limbs_modify __gmpz_limbs_modify
mp_ptr __gmpz_limbs_modify (mpz_ptr, mp_size_t);
#define mpz_limbs_finish __gmpz_limbs_finish
void __gmpz_limbs_finish (mpz_ptr, mp_size_t);
#define _mpz_realloc __gmpz_realloc
void *__gmpz_realloc (mpz_ptr, mp_size_t);
#define mpz_fits_uint_p __gmpz_fits_uint_p
int __gmpz_fits_uint_p (mpz_srcptr);
#define mpz_fits_sint_p __gmpz_fits_sint_p
int __gmpz_fits_sint_p (mpz_srcptr);
#define mpz_tstbit __gmpz_tstbit
int __gmpz_tstbit (mpz_srcptr, mp_bitcnt_t);
#define mpz_setbit __gmpz_setbit
void __gmpz_setbit (mpz_ptr, mp_bitcnt_t);
#define mpz_cmp __gmpz_cmp
int __gmpz_cmp (mpz_srcptr, mpz_srcptr);
#define mpz_cmp_ui __gmpz_cmp_ui
int __gmpz_cmp_ui (mpz_srcptr, unsigned long);
#define mpz_cmp_si __gmpz_cmp_si
int __gmpz_cmp_si (mpz_srcptr, long);
#define mpz_cmpabs __gmpz_cmpabs
int __gmpz_cmpabs (mpz_srcptr, mpz_srcptr);
#define mpz_abs __gmpz_abs
void __gmpz_abs (mpz_ptr, mpz_srcptr);
#define mpz_neg __gmpz_neg
void __gmpz_neg (mpz_ptr, mpz_srcptr);
#define mpz_add __gmpz_add
void __gmpz_add (mpz_ptr, mpz_srcptr, mpz_srcptr);
#define mpz_add_ui __gmpz_add_ui
void __gmpz_add_ui (mpz_ptr, mpz_srcptr, unsigned long);
#define mpz_sub __gmpz_sub
void __gmpz_sub (mpz_ptr, mpz_srcptr, mpz_srcptr);
#define mpz_sub_ui __gmpz_sub_ui
void __gmpz_sub_ui (mpz_ptr, mpz_srcptr, unsigned long);
#define mpz_mul __gmpz_mul
void __gmpz_mul (mpz_ptr, mpz_srcptr, mpz_srcptr);
#define mpz_mul_ui __gmpz_mul_ui
void __gmpz_mul_ui (mpz_ptr, mpz_srcptr, unsigned long);
#define mpz_mul_2exp __gmpz_mul_2exp
void __gmpz_mul_2exp (mpz_ptr, mpz_srcptr, mp_bitcnt_t);
#define mpz_addmul __gmpz_addmul
void __gmpz_addmul (mpz_ptr, mpz_srcptr, mpz_srcptr);
#define mpz_submul __gmpz_submul
void __gmpz_submul (mpz_ptr, mpz_srcptr, mpz_srcptr);
#define mpz_fdiv_q_2exp __gmpz_fdiv_q_2exp
void __gmpz_fdiv_q_2exp (mpz_ptr, mpz_srcptr, mp_bitcnt_t);
#define mpz_fdiv_r_2exp __gmpz_fdiv_r_2exp
void __gmpz_fdiv_r_2exp (mpz_ptr, mpz_srcptr, mp_bitcnt_t);
#define mpz_divexact __gmpz_divexact
void __gmpz_divexact (mpz_ptr, mpz_srcptr, mpz_srcptr);
#define mpz_divexact_ui __gmpz_divexact_ui
void __gmpz_divexact_ui (mpz_ptr, mpz_srcptr, unsigned long);
#define mpz_cdiv_qr __gmpz_cdiv_qr
void __gmpz_cdiv_qr (mpz_ptr, mpz_ptr, mpz_srcptr, mpz_srcptr);
#define mpz_fdiv_qr __gmpz_fdiv_qr
void __gmpz_fdiv_qr (mpz_ptr, mpz_ptr, mpz_srcptr, mpz_srcptr);
#define mpz_mod __gmpz_mod
void __gmpz_mod (mpz_ptr, mpz_srcptr, mpz_srcptr);
#define mpz_invert __gmpz_invert
int __gmpz_invert (mpz_ptr, mpz_srcptr, mpz_srcptr);
#define mpz_powm __gmpz_powm
void __gmpz_powm (mpz_ptr, mpz_srcptr, mpz_srcptr, mpz_srcptr);
#define mpz_gcd __gmpz_gcd
void __gmpz_gcd (mpz_ptr, mpz_srcptr, mpz_srcptr);
#define mpz_gcdext __gmpz_gcdext
void __gmpz_gcdext (mpz_ptr, mpz_ptr, mpz_ptr, mpz_srcptr, mpz_srcptr);
#define mpz_lcm __gmpz_lcm
void __gmpz_lcm (mpz_ptr, mpz_srcptr, mpz_srcptr);
#define mpz_sqrt __gmpz_sqrt
void __gmpz_sqrt (mpz_ptr, mpz_srcptr);
#define mpz_root __gmpz_root
int __gmpz_root (mpz_ptr, mpz_srcptr, unsigned long);
#define mpz_remove __gmpz_remove
mp_bitcnt_t __gmpz_remove (mpz_ptr, mpz_srcptr, mpz_srcptr);
#define mpz_probab_prime_p __gmpz_probab_prime_p
int __gmpz_probab_prime_p (mpz_srcptr, int);
#define mpz_nextprime __gmpz_nextprime
void __gmpz_nextprime (mpz_ptr, mpz_srcptr);
#define mpz_jacobi __gmpz_jacobi
int __gmpz_jacobi (mpz_srcptr, mpz_srcptr);
#define mpz_legendre __gmpz_jacobi
#define mpz_kronecker __gmpz_jacobi
#define mpz_divisible_p __gmpz_divisible_p
int __gmpz_divisible_p (mpz_srcptr, mpz_srcptr);
#define mpz_urandomm __gmpz_urandomm
void __gmpz_urandomm (mpz_ptr, gmp_randstate_t, mpz_srcptr);
#define mpz_urandomb __gmpz_urandomb
void __gmpz_urandomb (mpz_ptr, gmp_randstate_t, mp_bitcnt_t);
#define mpz_sgn(Z) ((Z)->_mp_size < 0 ? -1 : (Z)->_mp_size > 0)
#define mpz_odd_p(z) (((z)->_mp_size != 0) & (int) (z)->_mp_d[0])
#define mpz_even_p(z) (! mpz_odd_p (z))
/* rand */
#define gmp_randinit_default __gmp_randinit_default
void __gmp_randinit_default (gmp_randstate_t);
#define gmp_randinit_set __gmp_randinit_set
void __gmp_randinit_set (gmp_randstate_t, const __gmp_randstate_struct *);
#define gmp_randclear __gmp_randclear
void __gmp_randclear (gmp_randstate_t);
#define gmp_randseed __gmp_randseed
void __gmp_randseed (gmp_randstate_t, mpz_srcptr);
#define gmp_urandomb_ui __gmp_urandomb_ui
unsigned long __gmp_urandomb_ui (gmp_randstate_t, unsigned long);
#define gmp_urandomm_ui __gmp_urandomm_ui
unsigned long __gmp_urandomm_ui (gmp_randstate_t, unsigned long);
/* mpn */
#define mpn_zero __gmpn_zero
void __gmpn_zero (mp_ptr, mp_size_t);
#define mpn_copyi __gmpn_copyi
void __gmpn_copyi (mp_ptr, mp_srcptr, mp_size_t);
#define mpn_com __gmpn_com
void __gmpn_com (mp_ptr, mp_srcptr, mp_size_t);
#define mpn_scan1 __gmpn_scan1
mp_bitcnt_t __gmpn_scan1 (mp_srcptr, mp_bitcnt_t);
#define mpn_add_n __gmpn_add_n
mp_limb_t __gmpn_add_n (mp_ptr, mp_srcptr, mp_srcptr, mp_size_t);
#define mpn_sub_n __gmpn_sub_n
mp_limb_t __gmpn_sub_n (mp_ptr, mp_srcptr, mp_srcptr, mp_size_t);
#define mpn_add __gmpn_add
mp_limb_t __gmpn_add (mp_ptr, mp_srcptr, mp_size_t, mp_srcptr, mp_size_t);
#define mpn_addmul_1 __gmpn_addmul_1
mp_limb_t __gmpn_addmul_1 (mp_ptr, mp_srcptr, mp_size_t, mp_limb_t);
#define mpn_mul __gmpn_mul
mp_limb_t __gmpn_mul (mp_ptr, mp_srcptr, mp_size_t, mp_srcptr, mp_size_t);
#define mpn_mul_n __gmpn_mul_n
void __gmpn_mul_n (mp_ptr, mp_srcptr, mp_srcptr, mp_size_t);
#define mpn_sqr __gmpn_sqr
void __gmpn_sqr (mp_ptr, mp_srcptr, mp_size_t);
#define mpn_tdiv_qr __gmpn_tdiv_qr
void __gmpn_tdiv_qr (mp_ptr, mp_ptr, mp_size_t, mp_srcptr, mp_size_t, mp_srcptr, mp_size_t);
#define mpn_cmp __gmpn_cmp
int __gmpn_cmp (mp_srcptr, mp_srcptr, mp_size_t);
#define mpn_sizeinbase __gmpn_sizeinbase
size_t __gmpn_sizeinbase (mp_srcptr, mp_size_t, int);
/* mpf */
#define mpf_init __gmpf_init
void __gmpf_init (mpf_ptr);
#define mpf_init2 __gmpf_init2
void __gmpf_init2 (mpf_ptr, mp_bitcnt_t);
#define mpf_inits __gmpf_inits
void __gmpf_inits (mpf_ptr, ...);
#define mpf_clear __gmpf_clear
void __gmpf_clear (mpf_ptr);
#define mpf_clears __gmpf_clears
void __gmpf_clears (mpf_ptr, ...);
#define mpf_set_prec __gmpf_set_prec
void __gmpf_set_prec (mpf_ptr, mp_bitcnt_t);
#define mpf_set_d __gmpf_set_d
void __gmpf_set_d (mpf_ptr, double);
#define mpf_set_ui __gmpf_set_ui
void __gmpf_set_ui (mpf_ptr, unsigned long);
#define mpf_set_z __gmpf_set_z
void __gmpf_set_z (mpf_ptr, mpz_srcptr);
#define mpf_get_d __gmpf_get_d
double __gmpf_get_d (mpf_srcptr);
#define mpf_add __gmpf_add
void __gmpf_add (mpf_ptr, mpf_srcptr, mpf_srcptr);
#define mpf_sub __gmpf_sub
void __gmpf_sub (mpf_ptr, mpf_srcptr, mpf_srcptr);
#define mpf_mul __gmpf_mul
void __gmpf_mul (mpf_ptr, mpf_srcptr, mpf_srcptr);
#define mpf_div __gmpf_div
void __gmpf_div (mpf_ptr, mpf_srcptr, mpf_srcptr);
#define mpf_div_ui __gmpf_div_ui
void __gmpf_div_ui (mpf_ptr, mpf_srcptr, unsigned long);
#define mpf_ui_div __gmpf_ui_div
void __gmpf_ui_div (mpf_ptr, unsigned long, mpf_srcptr);
#define mpf_ui_sub __gmpf_ui_sub
void __gmpf_ui_sub (mpf_ptr, unsigned long, mpf_srcptr);
#define mpf_mul_2exp __gmpf_mul_2exp
void __gmpf_mul_2exp (mpf_ptr, mpf_srcptr, mp_bitcnt_t);
#define mpf_pow_ui __gmpf_pow_ui
void __gmpf_pow_ui (mpf_ptr, mpf_srcptr, unsigned long);
#define mpf_sqrt __gmpf_sqrt
void __gmpf_sqrt (mpf_ptr, mpf_srcptr);
#define mpf_abs __gmpf_abs
void __gmpf_abs (mpf_ptr, mpf_srcptr);
#define mpf_ceil __gmpf_ceil
void __gmpf_ceil (mpf_ptr, mpf_srcptr);
#define mpf_cmp __gmpf_cmp
int __gmpf_cmp (mpf_srcptr, mpf_srcptr);
#define mpf_cmp_ui __gmpf_cmp_ui
int __gmpf_cmp_ui (mpf_srcptr, unsigned long);
#define mpf_cmp_d __gmpf_cmp_d
int __gmpf_cmp_d (mpf_srcptr, double);
void __gmp_get_memory_functions (void *(**)(size_t), void *(**)(void *, size_t, size_t), void (**)(void *, size_t));
#ifdef __cplusplus
}
/* libgmpxx is absent: provide the two stream operators BICYCL uses */
inline std::ostream &operator<< (std::ostream &o, mpz_srcptr z)
{ char *s = mpz_get_str (NULL, 10, z); o << s; void (*freefn)(void *, size_t); __gmp_get_memory_functions (NULL, NULL, &freefn); freefn (s, std::char_traits<char>::length(s) + 1); return o; }
inline std::istream &operator>> (std::istream &i, mpz_ptr z)
{ std::string s; i >> s; if (mpz_set_str (z, s.c_str(), 10)) i.setstate (std::ios::failbit); return i; }
#endif
#endif
