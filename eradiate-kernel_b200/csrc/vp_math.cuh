/*
 * vp_math.cuh -- device math for the B200 volpath kernels.
 *
 * Every transcendental the path needs (log, exp, pow, sincos, cbrt) is a fixed
 * sequence of IEEE-754 operations: +,-,*,/ (round-to-nearest, the kernels are
 * compiled with -fmad=false -prec-div=true -prec-sqrt=true -ftz=false), explicit
 * fma(), sqrt and bit casts.  The algorithms are the CEPHES range reductions and
 * minimax polynomials that enoki's array_math.h cites (reference:
 * ext/enoki/include/enoki/array_math.h:260-960), which is what the reference
 * evaluates for Color<Float,1> values; scalar Float values go through the host
 * libm in the reference, which no GPU can reproduce bit
 * for bit, so the kernels use this one family everywhere.  Because nothing here
 * depends on compiler contraction or hardware approximations, a result is a pure
 * function of its inputs and can be checked bit for bit from the CPU.
 *
 * Templated on the real type: float (scalar_mono) and double (scalar_mono_double).
 */
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

/* The kernel headers are compiled in several translation units (csrc/kern_*.cu).  The exact ones share namespace
   vp; the tolerance-mode unit (kern_q32_fast.cu: FMA contraction, hardware approximations, VP_FASTMATH = 1)
   compiles the same text in its own namespace so that the two sets of inline functions never meet at link time. */
#ifndef VP_NS
#define VP_NS vp
#endif
#ifndef VP_FASTMATH
#define VP_FASTMATH 0
#endif

namespace VP_NS {

#define VP_HD __host__ __device__ __forceinline__
#define VP_D __device__ __forceinline__

template <typename T> struct RealBits;
template <> struct RealBits<float> {
    using U = uint32_t;
    static constexpr int mant = 23, bias = 127;
};
template <> struct RealBits<double> {
    using U = uint64_t;
    static constexpr int mant = 52, bias = 1023;
};

VP_D uint32_t bits_of(float x) { return __float_as_uint(x); }
VP_D uint64_t bits_of(double x) { return (uint64_t) __double_as_longlong(x); }
template <typename T> VP_D T real_of(typename RealBits<T>::U u);
template <> VP_D float real_of<float>(uint32_t u) { return __uint_as_float(u); }
template <> VP_D double real_of<double>(uint64_t u) { return __longlong_as_double((long long) u); }

VP_D float fma_(float a, float b, float c) { return __fmaf_rn(a, b, c); }
VP_D double fma_(double a, double b, double c) { return __fma_rn(a, b, c); }
#if VP_FASTMATH
VP_D float sqrt_(float x) { return sqrtf(x); }   /* -prec-sqrt=false: MUFU.RSQ based */
#else
VP_D float sqrt_(float x) { return __fsqrt_rn(x); }
#endif
VP_D double sqrt_(double x) { return __dsqrt_rn(x); }
#if VP_FASTMATH
VP_D float div_(float a, float b) { return __fdividef(a, b); }
#else
VP_D float div_(float a, float b) { return __fdiv_rn(a, b); }
#endif
VP_D double div_(double a, double b) { return __ddiv_rn(a, b); }
VP_D float floor_(float x) { return floorf(x); }
VP_D double floor_(double x) { return floor(x); }
VP_D float ceil_(float x) { return ceilf(x); }
VP_D double ceil_(double x) { return ceil(x); }

template <typename T> struct Lim;
template <> struct Lim<float> {
    static VP_D float inf() { return __uint_as_float(0x7f800000u); }
    static VP_D float nan() { return __uint_as_float(0x7fc00000u); }
    static VP_D float min_normal() { return 1.17549435e-38f; }
};
template <> struct Lim<double> {
    static VP_D double inf() { return __longlong_as_double(0x7ff0000000000000ll); }
    static VP_D double nan() { return __longlong_as_double(0x7ff8000000000000ll); }
    static VP_D double min_normal() { return 2.2250738585072014e-308; }
};

template <typename T> VP_D bool finite_(T x) { return x - x == (T) 0; }

/* 2^k, k inside the normal exponent range */
template <typename T> VP_D T pow2i(int k) {
    using U = typename RealBits<T>::U;
    return real_of<T>((U) (k + RealBits<T>::bias) << RealBits<T>::mant);
}
/* z * 2^n in two exact-or-single-rounding steps */
template <typename T> VP_D T scale2(T z, int n) {
    int h = n >> 1;
    return z * pow2i<T>(h) * pow2i<T>(n - h);
}
/* x = m * 2^e, m in [0.5,1); x positive, finite, normal */
template <typename T> VP_D T split2(T x, int &e) {
    using U = typename RealBits<T>::U;
    U u = bits_of(x);
    e = (int) (u >> RealBits<T>::mant) - (RealBits<T>::bias - 1);
    U mant_mask = ((U) 1 << RealBits<T>::mant) - 1;
    return real_of<T>((u & mant_mask) | ((U) (RealBits<T>::bias - 1) << RealBits<T>::mant));
}

/* ---------------------------------------------------------------- log --- */
#if VP_FASTMATH
/* Tolerance mode: MUFU.LG2 / MUFU.EX2.  __logf has an absolute error of about 2^-21.4 near 1,
   which on the free-flight distance -log(1 - u) / sigma is 4e-7 mean free paths: the size of fp32 position rounding. */
VP_D float log_(float x) { return __logf(x); }
#else
VP_D float log_(float x) {
    if (!(x >= 0.f)) return Lim<float>::nan();
    if (x == 0.f) return -Lim<float>::inf();
    if (x == Lim<float>::inf()) return x;
    if (x < Lim<float>::min_normal()) x = Lim<float>::min_normal();
    int ei;
    float m = split2(x, ei);
    if (m < 0.70710678118654752440f) { ei -= 1; m = m + m; }
    m = m - 1.f;
    float e = (float) ei;
    float p = 7.0376836292e-2f;
    p = fma_(p, m, -1.1514610310e-1f);
    p = fma_(p, m, 1.1676998740e-1f);
    p = fma_(p, m, -1.2420140846e-1f);
    p = fma_(p, m, 1.4249322787e-1f);
    p = fma_(p, m, -1.6668057665e-1f);
    p = fma_(p, m, 2.0000714765e-1f);
    p = fma_(p, m, -2.4999993993e-1f);
    p = fma_(p, m, 3.3333331174e-1f);
    float z = m * m;
    float y = p * (m * z);
    y = fma_(e, -2.12194440e-4f, y);
    z = fma_(z, -0.5f, m + y);
    return fma_(e, 0.693359375f, z);
}
#endif

VP_D double log_(double x) {
    if (!(x >= 0.0)) return Lim<double>::nan();
    if (x == 0.0) return -Lim<double>::inf();
    if (x == Lim<double>::inf()) return x;
    if (x < Lim<double>::min_normal()) x = Lim<double>::min_normal();
    int ei;
    double m = split2(x, ei);
    bool ge = m >= 0.70710678118654752440;
    if (!ge) ei -= 1;
    double e = (double) ei, r;
    if (ei > 2 || ei < -2) {
        double z, y;
        if (ge) { z = (m - 0.5) - 0.5; y = fma_(0.5, m, 0.5); }
        else    { z = m - 0.5;         y = fma_(0.5, z, 0.5); }
        double x2 = z / y;
        z = x2 * x2;
        double P = -7.89580278884799154124e-1;
        P = fma_(P, z, 1.63866645699558079767e1);
        P = fma_(P, z, -6.41409952958715622951e1);
        double Q = 1.0;
        Q = fma_(Q, z, -3.56722798256324312549e1);
        Q = fma_(Q, z, 3.12093766372244180303e2);
        Q = fma_(Q, z, -7.69691943550460008604e2);
        z = x2 * (z * P / Q);
        r = fma_(-e, 2.121944400546905827679e-4, z) + x2;
    } else {
        double x2 = (ge ? m : m + m) - 1.0;
        double z = x2 * x2;
        double P = 1.01875663804580931796e-4;
        P = fma_(P, x2, 4.97494994976747001425e-1);
        P = fma_(P, x2, 4.70579119878881725854e0);
        P = fma_(P, x2, 1.44989225341610930846e1);
        P = fma_(P, x2, 1.79368678507819816313e1);
        P = fma_(P, x2, 7.70838733755885391666e0);
        double Q = 1.0;
        Q = fma_(Q, x2, 1.12873587189167450590e1);
        Q = fma_(Q, x2, 4.52279145837532221105e1);
        Q = fma_(Q, x2, 8.29875266912776603211e1);
        Q = fma_(Q, x2, 7.11544750618563894466e1);
        Q = fma_(Q, x2, 2.31251620126765340583e1);
        double y = x2 * (z * P / Q);
        y = fma_(-e, 2.121944400546905827679e-4, y);
        r = x2 + fma_(-0.5, z, y);
    }
    return fma_(e, 0.693359375, r);
}

/* ---------------------------------------------------------------- exp --- */
#if VP_FASTMATH
VP_D float exp_(float x) { return __expf(x); }
#else
VP_D float exp_(float x) {
    if (x > 88.3762588501f) return Lim<float>::inf();
    if (x < -88.3762588501f) return 0.f;
    if (x != x) return x;
    float n = floor_(fma_(1.4426950408889634073599f, x, 0.5f));
    float xr = fma_(-n, 0.693359375f, x);
    xr = fma_(-n, -2.12194440e-4f, xr);
    float z = 1.9875691500e-4f;
    z = fma_(z, xr, 1.3981999507e-3f);
    z = fma_(z, xr, 8.3334519073e-3f);
    z = fma_(z, xr, 4.1665795894e-2f);
    z = fma_(z, xr, 1.6666665459e-1f);
    z = fma_(z, xr, 5.0000001201e-1f);
    z = fma_(z, xr * xr, xr + 1.f);
    return scale2(z, (int) n);
}
#endif

VP_D double exp_(double x) {
    if (x > 7.0943613930310391424428e2) return Lim<double>::inf();
    if (x < -7.0943613930310391424428e2) return 0.0;
    if (x != x) return x;
    double n = floor_(fma_(1.4426950408889634073599, x, 0.5));
    double xr = fma_(-n, 6.93145751953125e-1, x);
    xr = fma_(-n, 1.42860682030941723212e-6, xr);
    double z = xr * xr;
    double P = 1.26177193074810590878e-4;
    P = fma_(P, z, 3.02994407707441961300e-2);
    P = fma_(P, z, 9.99999999999999999910e-1);
    double Q = 3.00198505138664455042e-6;
    Q = fma_(Q, z, 2.52448340349684104192e-3);
    Q = fma_(Q, z, 2.27265548208155028766e-1);
    Q = fma_(Q, z, 2.00000000000000000009e0);
    double p = P * xr;
    double pq = p / (Q - p);
    z = pq + pq + 1.0;
    return scale2(z, (int) n);
}

template <typename T> VP_D T pow_(T x, T y) { return exp_(log_(x) * y); }
#if VP_FASTMATH
/* pow (RPV), sincos (azimuths, sensor directions) and cbrt (Rayleigh) run once per scattering or surface event, not per
   tentative collision: the accurate library versions cost nothing measurable and keep the per-path error down.  Only
   what the STEP phase evaluates per collision -- log, exp, division, reciprocal, square root -- is approximate. */
VP_D float pow_(float x, float y) { return powf(x, y); }
#endif

/* ------------------------------------------------------------- sincos --- */
template <typename T> struct TrigK;
template <> struct TrigK<float> {
    static VP_D float dp1() { return 0.78515625f; }
    static VP_D float dp2() { return 2.4187564849853515625e-4f; }
    static VP_D float dp3() { return 3.77489497744594108e-8f; }
    static VP_D float sp(float z) {
        float r = -1.9515295891e-4f;
        r = fma_(r, z, 8.3321608736e-3f);
        return fma_(r, z, -1.6666654611e-1f);
    }
    static VP_D float cp(float z) {
        float r = 2.443315711809948e-5f;
        r = fma_(r, z, -1.388731625493765e-3f);
        return fma_(r, z, 4.166664568298827e-2f);
    }
};
template <> struct TrigK<double> {
    static VP_D double dp1() { return 7.85398125648498535156e-1; }
    static VP_D double dp2() { return 3.77489470793079817668e-8; }
    static VP_D double dp3() { return 2.69515142907905952645e-15; }
    static VP_D double sp(double z) {
        double r = 1.58962301576546568060e-10;
        r = fma_(r, z, -2.50507477628578072866e-8);
        r = fma_(r, z, 2.75573136213857245213e-6);
        r = fma_(r, z, -1.98412698295895385996e-4);
        r = fma_(r, z, 8.33333333332211858878e-3);
        return fma_(r, z, -1.66666666666666307295e-1);
    }
    static VP_D double cp(double z) {
        double r = -1.13585365213876817300e-11;
        r = fma_(r, z, 2.08757008419747316778e-9);
        r = fma_(r, z, -2.75573141792967388112e-7);
        r = fma_(r, z, 2.48015872888517045348e-5);
        r = fma_(r, z, -1.38888888888730564116e-3);
        return fma_(r, z, 4.16666666666665929218e-2);
    }
};

/* joint sine / cosine, |x| < ~8192 (arguments here are 2*pi*u, pi*u, disk angles) */
#if VP_FASTMATH
VP_D void sincos_(float x, float &s_out, float &c_out) { sincosf(x, &s_out, &c_out); }
#endif
template <typename T> VP_D void sincos_(T x, T &s_out, T &c_out) {
    using K = TrigK<T>;
    T xa = x < 0 ? -x : x;
    int j = (int) (xa * (T) 1.2732395447351626862);
    j = (j + 1) & ~1;
    T y = (T) j;
    y = ((xa - y * K::dp1()) - y * K::dp2()) - y * K::dp3();
    T z = y * y;
    T s = K::sp(z) * z;
    T c = K::cp(z) * z;
    s = fma_(s, y, y);
    c = fma_(c, z, fma_(z, (T) -0.5, (T) 1));
    bool swap = (j & 2) != 0;
    T rs = swap ? c : s, rc = swap ? s : c;
    if (((j & 4) != 0) != (x < 0)) rs = -rs;
    if (((j - 2) & 4) == 0) rc = -rc;
    s_out = rs;
    c_out = rc;
}

/* --------------------------------------------------------------- cbrt --- */
#if VP_FASTMATH
VP_D float cbrt_(float x) { return cbrtf(x); }
#endif
template <typename T> VP_D T cbrt_(T x) {
    if (x == 0 || x != x || x == Lim<T>::inf() || x == -Lim<T>::inf())
        return x;
    const T CBRT2 = (T) 1.25992104989487316477, CBRT4 = (T) 1.58740105196819947475,
            THIRD = (T) (1.0 / 3.0);
    T xa = x < 0 ? -x : x;
    if (xa < Lim<T>::min_normal()) xa = Lim<T>::min_normal();
    int e;
    T m = split2(xa, e);
    T p = (T) -0.13466110473359520655053;
    p = fma_(p, m, (T) 0.54664601366395524503440);
    p = fma_(p, m, (T) -0.95438224771509446525043);
    p = fma_(p, m, (T) 1.1399983354717293273738);
    p = fma_(p, m, (T) 0.40238979564544752126924);
    m = p;
    int ea = e < 0 ? -e : e;
    int q = ea / 3, rem = ea - 3 * q;
    if (rem == 1) m = m * (e >= 0 ? CBRT2 : (T) 1 / CBRT2);
    else if (rem == 2) m = m * (e >= 0 ? CBRT4 : (T) 1 / CBRT4);
    T r = scale2(m, e >= 0 ? q : -q);
    if (x < 0) r = -r;
    r = r - (r - (x / (r * r))) * THIRD;
    if (sizeof(T) == 8)
        r = r - (r - (x / (r * r))) * THIRD;
    return r;
}

template <typename T> VP_D T safe_sqrt_(T x) { return sqrt_(x > 0 ? x : (T) 0); }

} // namespace VP_NS
