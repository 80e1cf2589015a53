/* Prototype-only stand-in for <gmp.h>: this container ships the GMP 6.3 runtime (libgmp.so.10) but not its headers.
 * Declares exactly the entry points libff's alt_bn128 / Fp_model / bigint code uses, with GMP's real symbol names
 * (__gmpn_*, __gmpz_*), so the reference sources compile unmodified and link with -l:libgmp.so.10.
 * TEST INFRASTRUCTURE ONLY (used by oracle/ref/Makefile to build oracle/_ref). */
#ifndef B200_GMP_SHIM_H
#define B200_GMP_SHIM_H
#include <stddef.h>
#include <stdio.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef unsigned long mp_limb_t;
typedef long mp_limb_signed_t;
typedef long mp_size_t;
typedef unsigned long mp_bitcnt_t;
typedef mp_limb_t* mp_ptr;
typedef const mp_limb_t* mp_srcptr;
typedef struct { int _mp_alloc; int _mp_size; mp_limb_t* _mp_d; } __mpz_struct;
typedef __mpz_struct mpz_t[1];
typedef __mpz_struct* mpz_ptr;
typedef const __mpz_struct* mpz_srcptr;
#define GMP_LIMB_BITS 64
#define GMP_NUMB_BITS 64
#define GMP_NAIL_BITS 0
#define __GNU_MP_VERSION 6
#define __GNU_MP_VERSION_MINOR 3

#define mpn_copyi __gmpn_copyi
#define mpn_copyd __gmpn_copyd
#define mpn_zero __gmpn_zero
#define mpn_cmp __gmpn_cmp
#define mpn_add __gmpn_add
#define mpn_add_n __gmpn_add_n
#define mpn_add_1 __gmpn_add_1
#define mpn_sub __gmpn_sub
#define mpn_sub_n __gmpn_sub_n
#define mpn_sub_1 __gmpn_sub_1
#define mpn_mul __gmpn_mul
#define mpn_mul_n __gmpn_mul_n
#define mpn_mul_1 __gmpn_mul_1
#define mpn_addmul_1 __gmpn_addmul_1
#define mpn_submul_1 __gmpn_submul_1
#define mpn_rshift __gmpn_rshift
#define mpn_lshift __gmpn_lshift
#define mpn_set_str __gmpn_set_str
#define mpn_get_str __gmpn_get_str
#define mpn_gcdext __gmpn_gcdext
#define mpn_tdiv_qr __gmpn_tdiv_qr
#define mpz_init __gmpz_init
#define mpz_init_set __gmpz_init_set
#define mpz_init_set_ui __gmpz_init_set_ui
#define mpz_init_set_str __gmpz_init_set_str
#define mpz_set __gmpz_set
#define mpz_set_ui __gmpz_set_ui
#define mpz_set_str __gmpz_set_str
#define mpz_get_ui __gmpz_get_ui
#define mpz_get_str __gmpz_get_str
#define mpz_add_ui __gmpz_add_ui
#define mpz_mul_2exp __gmpz_mul_2exp
#define mpz_fdiv_q_2exp __gmpz_fdiv_q_2exp
#define mpz_clear __gmpz_clear
#define mpz_cmp __gmpz_cmp
#define mpz_cmp_ui __gmpz_cmp_ui
#define mpz_sizeinbase __gmpz_sizeinbase
#define mpz_import __gmpz_import
#define mpz_jacobi __gmpz_jacobi
#define mpz_export __gmpz_export
#define gmp_printf __gmp_printf

void __gmpn_copyi(mp_ptr, mp_srcptr, mp_size_t);
void __gmpn_copyd(mp_ptr, mp_srcptr, mp_size_t);
void __gmpn_zero(mp_ptr, mp_size_t);
int __gmpn_cmp(mp_srcptr, mp_srcptr, mp_size_t);
mp_limb_t __gmpn_add(mp_ptr, mp_srcptr, mp_size_t, mp_srcptr, mp_size_t);
mp_limb_t __gmpn_add_n(mp_ptr, mp_srcptr, mp_srcptr, mp_size_t);
mp_limb_t __gmpn_add_1(mp_ptr, mp_srcptr, mp_size_t, mp_limb_t);
mp_limb_t __gmpn_sub(mp_ptr, mp_srcptr, mp_size_t, mp_srcptr, mp_size_t);
mp_limb_t __gmpn_sub_n(mp_ptr, mp_srcptr, mp_srcptr, mp_size_t);
mp_limb_t __gmpn_sub_1(mp_ptr, mp_srcptr, mp_size_t, mp_limb_t);
mp_limb_t __gmpn_mul(mp_ptr, mp_srcptr, mp_size_t, mp_srcptr, mp_size_t);
void __gmpn_mul_n(mp_ptr, mp_srcptr, mp_srcptr, mp_size_t);
mp_limb_t __gmpn_mul_1(mp_ptr, mp_srcptr, mp_size_t, mp_limb_t);
mp_limb_t __gmpn_addmul_1(mp_ptr, mp_srcptr, mp_size_t, mp_limb_t);
mp_limb_t __gmpn_submul_1(mp_ptr, mp_srcptr, mp_size_t, mp_limb_t);
mp_limb_t __gmpn_rshift(mp_ptr, mp_srcptr, mp_size_t, unsigned int);
mp_limb_t __gmpn_lshift(mp_ptr, mp_srcptr, mp_size_t, unsigned int);
mp_size_t __gmpn_set_str(mp_ptr, const unsigned char*, size_t, int);
size_t __gmpn_get_str(unsigned char*, int, mp_ptr, mp_size_t);
mp_size_t __gmpn_gcdext(mp_ptr, mp_ptr, mp_size_t*, mp_ptr, mp_size_t, mp_ptr, mp_size_t);
void __gmpn_tdiv_qr(mp_ptr, mp_ptr, mp_size_t, mp_srcptr, mp_size_t, mp_srcptr, mp_size_t);
void __gmpz_init(mpz_ptr);
int __gmpz_jacobi(mpz_srcptr, mpz_srcptr);
void __gmpz_import(mpz_ptr, size_t, int, size_t, int, size_t, const void*);
void* __gmpz_export(void*, size_t*, int, size_t, int, size_t, mpz_srcptr);
void __gmpz_init_set(mpz_ptr, mpz_srcptr);
void __gmpz_init_set_ui(mpz_ptr, unsigned long);
int __gmpz_init_set_str(mpz_ptr, const char*, int);
void __gmpz_set(mpz_ptr, mpz_srcptr);
void __gmpz_set_ui(mpz_ptr, unsigned long);
int __gmpz_set_str(mpz_ptr, const char*, int);
unsigned long __gmpz_get_ui(mpz_srcptr);
char* __gmpz_get_str(char*, int, mpz_srcptr);
void __gmpz_add_ui(mpz_ptr, mpz_srcptr, unsigned long);
void __gmpz_mul_2exp(mpz_ptr, mpz_srcptr, mp_bitcnt_t);
void __gmpz_fdiv_q_2exp(mpz_ptr, mpz_srcptr, mp_bitcnt_t);
void __gmpz_clear(mpz_ptr);
int __gmpz_cmp(mpz_srcptr, mpz_srcptr);
int __gmpz_cmp_ui(mpz_srcptr, unsigned long);
size_t __gmpz_sizeinbase(mpz_srcptr, int);
int __gmp_printf(const char*, ...);
static inline int mpz_sgn(mpz_srcptr z) { return z->_mp_size < 0 ? -1 : (z->_mp_size > 0); }
#ifdef __cplusplus
}
#endif
#endif
