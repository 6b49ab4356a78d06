// qe_ops.h -- per-row semantics of the 17 expression functions and the
// accumulators, as branch-free scalar device functions.
//
// Compiles under nvcc, under NVRTC (no standard headers are included here) and
// as plain host C++ (tests compile it for the host to check it against the
// oracle without a GPU).
//
// Reference semantics (M/ = src/main/java/net/jhorstmann/queryengine/):
//   M/evaluator/Interpreter.kt:29-109 is canonical; where the three evaluator
//   modes disagree the BYTECODE_COMPILER behaviour is
//   used: ordering comparisons follow java.lang.Double.compare
//   (M/evaluator/BytecodeCompiler.kt:300), NOT null is null (:346-351).
//
// A value is a pair (payload, valid).  `valid == false` is the Kotlin null.
// Payloads of null values are unspecified and must never influence a result.
#ifndef QE_OPS_H
#define QE_OPS_H

#if defined(__CUDACC__) || defined(__CUDACC_RTC__)
#define QE_HD __host__ __device__ __forceinline__
#else
#define QE_HD inline
#endif

typedef unsigned long long qe_u64;
typedef long long qe_i64;
typedef unsigned int qe_u32;
typedef int qe_i32;
typedef unsigned char qe_u8;

// ---- bit casts -------------------------------------------------------------
QE_HD qe_i64 qe_f64_bits(double d)
{
#if defined(__CUDA_ARCH__)
    return __double_as_longlong(d);
#else
    union { double d; qe_i64 i; } u;
    u.d = d;
    return u.i;
#endif
}

QE_HD double qe_bits_f64(qe_i64 i)
{
#if defined(__CUDA_ARCH__)
    return __longlong_as_double(i);
#else
    union { double d; qe_i64 i; } u;
    u.i = i;
    return u.d;
#endif
}

// ---- exact integer -> double widening (INT64 / DATE32 columns) -----------------
// The hardware conversion (I2F.F64.S64) runs on the quarter-rate XU pipe and was
// the limiter of the int64 scan (ncu: XU 78 % busy, HBM 19 %).  For |x| < 2^51
// the classic magic-number form is exact and uses one integer add and one DADD:
//   bits(2^52 + 2^51) + x  is the double  2^52 + 2^51 + x ; subtract the bias.
QE_HD double qe_i64_to_f64(qe_i64 x)
{
    const qe_i64 lim = (qe_i64)1 << 51;
    if ((qe_u64)(x + lim) < ((qe_u64)1 << 52))
        return qe_bits_f64(x + (qe_i64)0x4338000000000000LL) - 6755399441055744.0;
    return (double)x;
}

QE_HD double qe_i32_to_f64(qe_i32 x)
{
    return qe_bits_f64((qe_i64)x + (qe_i64)0x4338000000000000LL) - 6755399441055744.0;
}

// ---- java.lang.Double.compare as a monotone integer key ---------------------
// Double.compare orders  -Inf < ... < -0.0 < 0.0 < ... < +Inf < NaN  and treats
// every NaN as one value.  Canonicalise NaN to 0x7ff8000000000000 (what
// doubleToLongBits does), then flip the magnitude bits of negative numbers:
// the signed 64-bit keys compare exactly like Double.compare does.
QE_HD qe_i64 qe_total_order_key(double d)
{
    qe_i64 b = qe_f64_bits(d);
    b = (d != d) ? (qe_i64)0x7ff8000000000000LL : b;
    return b ^ ((b >> 63) & (qe_i64)0x7fffffffffffffffLL);
}

// M/evaluator/Interpreter.kt:104-107 via Double.compareTo
QE_HD bool qe_cmp_lt(double a, double b) { return qe_total_order_key(a) <  qe_total_order_key(b); }
QE_HD bool qe_cmp_le(double a, double b) { return qe_total_order_key(a) <= qe_total_order_key(b); }
QE_HD bool qe_cmp_ge(double a, double b) { return qe_total_order_key(a) >= qe_total_order_key(b); }
QE_HD bool qe_cmp_gt(double a, double b) { return qe_total_order_key(a) >  qe_total_order_key(b); }
// M/evaluator/Interpreter.kt:102-103: boxed equals == doubleToLongBits equality
QE_HD bool qe_cmp_eq(double a, double b) { return qe_total_order_key(a) == qe_total_order_key(b); }
QE_HD bool qe_cmp_ne(double a, double b) { return qe_total_order_key(a) != qe_total_order_key(b); }

// Comparison against a literal that is neither NaN nor a zero: Double.compare
// then agrees with the IEEE comparison except when the column value is NaN
// (NaN sorts above everything).  One DSETP instead of the 64-bit key dance.
// The code generator picks these only when it has checked the literal.
QE_HD bool qe_cmp_lt_k(double a, double k) { return a <  k; }                 // NaN < k  is false
QE_HD bool qe_cmp_le_k(double a, double k) { return a <= k; }                 // NaN <= k is false
QE_HD bool qe_cmp_ge_k(double a, double k) { return !(a <  k); }              // NaN >= k is true
QE_HD bool qe_cmp_gt_k(double a, double k) { return !(a <= k); }              // NaN > k  is true
QE_HD bool qe_cmp_eq_k(double a, double k) { return a == k; }
QE_HD bool qe_cmp_ne_k(double a, double k) { return !(a == k); }

// ---- arithmetic (M/evaluator/Interpreter.kt:94-100) --------------------------
// Plain IEEE binary64.  The translation unit is compiled with -fmad=false so
// a*b+c is never contracted, matching the JVM's separate DMUL / DADD.
QE_HD double qe_add(double a, double b) { return a + b; }
QE_HD double qe_sub(double a, double b) { return a - b; }
QE_HD double qe_mul(double a, double b) { return a * b; }
QE_HD double qe_div(double a, double b) { return a / b; }
QE_HD double qe_neg(double a) { return qe_bits_f64(qe_f64_bits(a) ^ (qe_i64)0x8000000000000000ULL); }

#if defined(__CUDA_ARCH__)
// JVM DREM == C fmod.  The library fmod is a shift-and-subtract loop (measured on B200: 8 iterations
// of 10 instructions per call on the benchmark tree, half of all instructions of the kernel).  When
// the quotient fits 53 bits the result can be had exactly in three steps:
//   q = trunc(a / b) with the division rounded TOWARD ZERO: rz(a/b) and a/b lie on the same side of
//       every integer, so q is the true integer quotient (it is exactly representable below 2^53);
//   r = fma(-q, b, a): the true remainder a - q*b is always representable, so the single rounding of
//       the fused multiply-add returns it exactly;
//   a zero result takes the sign of the dividend.
// Everything else (NaN, infinities, b == 0, quotients of 2^53 and more) goes to the library routine.
__device__ __forceinline__ double qe_mod(double a, double b)
{
    const double q = __ddiv_rz(a, b);
    // false for NaN / Inf quotients as well; an infinite divisor (quotient 0) must not reach the fma (0 * Inf)
    if (fabs(q) < 9007199254740992.0 && b != 0.0 && fabs(b) < qe_bits_f64((qe_i64)0x7ff0000000000000LL)) {
        const double r = __fma_rn(-trunc(q), b, a);
        return r == 0.0 ? copysign(0.0, a) : r;
    }
    return fmod(a, b);
}
// MOD by a literal (`MOD(x, 7.0)`): kk = |k| and inv = 1 / kk are compile-time constants (the code generator only
// takes this path for normal kk and inv), so the quotient costs a multiply instead of an IEEE division
// (~25 instructions).  q = |a| * inv is within 2^-52 relative of |a| / kk, so below 2^51 its truncation n is the
// true integer quotient or one off; r = fma(-n, kk, |a|) tells which (rounding is monotonic: an exact r < 0 or
// >= kk rounds to one), n is corrected and r recomputed -- with the true quotient the fused multiply-add is exact,
// as in qe_mod.  fmod(a, k) = sign(a) * fmod(|a|, |k|), and a zero takes the sign of the dividend.
__device__ __forceinline__ double qe_mod_k(double a, double k, double kk, double inv)
{
    const double aa = fabs(a);
    const double q = aa * inv;
    if (q < 2251799813685248.0) {                              // 2^51; false for NaN and Inf dividends
        double n = trunc(q);
        const double r0 = __fma_rn(-n, kk, aa);
        n += r0 < 0.0 ? -1.0 : (r0 >= kk ? 1.0 : 0.0);
        return copysign(__fma_rn(-n, kk, aa), a);
    }
    return qe_mod(a, k);
}
#elif defined(__CUDACC__) || defined(__CUDACC_RTC__)
QE_HD double qe_mod(double a, double b) { return fmod(a, b); }                // host pass of nvcc
QE_HD double qe_mod_k(double a, double k, double, double) { return fmod(a, k); }
#else
extern "C" double fmod(double, double);
QE_HD double qe_mod(double a, double b) { return fmod(a, b); }
QE_HD double qe_mod_k(double a, double k, double, double) { return fmod(a, k); }
#endif

// ---- Kleene logic (M/evaluator/Interpreter.kt:54-92) -------------------------
// Payload bits of null operands are masked first so they cannot leak.
QE_HD void qe_and(bool a, bool va, bool b, bool vb, bool &r, bool &vr)
{
    a = a && va;
    b = b && vb;
    r = a && b;
    vr = (va && vb) || (va && !a) || (vb && !b);
}

QE_HD void qe_or(bool a, bool va, bool b, bool vb, bool &r, bool &vr)
{
    a = a && va;
    b = b && vb;
    r = a || b;
    vr = (va && vb) || (va && a) || (vb && b);
}

// ---- java.lang.Math.min / max (M/evaluator/Accumulators.kt:62,80) ------------
// NaN if either is NaN; min(-0.0, 0.0) = -0.0, max(-0.0, 0.0) = 0.0.  Both are
// commutative and associative, so any reduction tree gives the sequential
// result bit for bit (NaN results are canonicalised by the caller).
QE_HD double qe_math_min(double a, double b)
{
    if (a != a) return a;
    if (b != b) return b;
    if (a == 0.0 && b == 0.0) return qe_bits_f64(qe_f64_bits(a) | qe_f64_bits(b));
    return a <= b ? a : b;
}

QE_HD double qe_math_max(double a, double b)
{
    if (a != a) return a;
    if (b != b) return b;
    if (a == 0.0 && b == 0.0) return qe_bits_f64(qe_f64_bits(a) & qe_f64_bits(b));
    return a >= b ? a : b;
}

#endif  // QE_OPS_H
