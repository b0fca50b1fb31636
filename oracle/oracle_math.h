/*
 * oracle_math.h -- TEST INFRASTRUCTURE (CPU oracle), not product code.
 *
 * Deterministic IEEE-754 restatement of the transcendental functions the
 * volpath path uses.  The reference mixes glibc libm (scalar Float) and enoki's
 * CEPHES-derived polynomials (Color<Float,1>), hardware rcp/rsqrt estimates and
 * compiler-chosen FMA contraction (ext/enoki/include/enoki/
 * array_fallbacks.h:23-90, array_math.h:260-960), so its fp32 results are
 * platform dependent at the ulp level.  The oracle therefore fixes ONE
 * sequence of IEEE operations (+,-,*,/,fma,sqrt, bit casts) per function,
 * following the published CEPHES algorithms enoki's array_math.h cites, so that
 * the CUDA kernels (which implement the same sequences independently in
 * eradiate-kernel_b200/csrc/vp_math.cuh) can be compared bit for bit.
 *
 * Must be compiled with -ffp-contract=off: every fused operation is explicit.
 */
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <utility>

namespace omath {

template <typename T> struct Bits;
template <> struct Bits<float> {
    using U = uint32_t; using I = int32_t;
    static constexpr int mant = 23, bias = 127;
};
template <> struct Bits<double> {
    using U = uint64_t; using I = int64_t;
    static constexpr int mant = 52, bias = 1023;
};

template <typename T> inline typename Bits<T>::U to_bits(T x) {
    typename Bits<T>::U u; std::memcpy(&u, &x, sizeof(T)); return u;
}
template <typename T> inline T from_bits(typename Bits<T>::U u) {
    T x; std::memcpy(&x, &u, sizeof(T)); return x;
}

inline float  fma_(float a, float b, float c)    { return std::fmaf(a, b, c); }
inline double fma_(double a, double b, double c) { return std::fma(a, b, c); }

/* Horner evaluation c[0] + x*(c[1] + x*(... c[n])) with one fma per term. */
template <typename T, int N> inline T horner(T x, const T (&c)[N]) {
    T r = c[N - 1];
    for (int i = N - 2; i >= 0; --i)
        r = fma_(r, x, c[i]);
    return r;
}

/* 2^k for k in the normal exponent range. */
template <typename T> inline T pow2i(int k) {
    using U = typename Bits<T>::U;
    return from_bits<T>((U)(k + Bits<T>::bias) << Bits<T>::mant);
}

/* z * 2^n, split in two steps so that results in the subnormal range round once. */
template <typename T> inline T ldexp_(T z, int n) {
    int h = n >> 1;
    return z * pow2i<T>(h) * pow2i<T>(n - h);
}

/* x = m * 2^e with m in [0.5, 1); x must be positive, finite and normal. */
template <typename T> inline T frexp_(T x, int &e) {
    using U = typename Bits<T>::U;
    U u = to_bits(x);
    e = (int) (u >> Bits<T>::mant) - (Bits<T>::bias - 1);
    U mant_mask = ((U) 1 << Bits<T>::mant) - 1;
    u = (u & mant_mask) | ((U)(Bits<T>::bias - 1) << Bits<T>::mant);
    return from_bits<T>(u);
}

/* ------------------------------------------------------------------ log -- */
inline float log_(float x) {
    const float inf = std::numeric_limits<float>::infinity();
    if (!(x >= 0.f)) return std::numeric_limits<float>::quiet_NaN();
    if (x == 0.f) return -inf;
    if (x == inf) return inf;
    if (x < 1.17549435e-38f) x = 1.17549435e-38f;
    int ei; float m = frexp_(x, ei);
    if (m < 0.70710678118654752440f) { ei -= 1; m = m + m; }
    m = m - 1.f;
    float e = (float) ei;
    static const float c[9] = { 3.3333331174e-1f, -2.4999993993e-1f, 2.0000714765e-1f,
                                -1.6668057665e-1f, 1.4249322787e-1f, -1.2420140846e-1f,
                                1.1676998740e-1f, -1.1514610310e-1f, 7.0376836292e-2f };
    float z = m * m;
    float y = horner(m, c) * (m * z);
    y = fma_(e, -2.12194440e-4f, y);
    z = fma_(z, -0.5f, m + y);
    return fma_(e, 0.693359375f, z);
}

inline double log_(double x) {
    const double inf = std::numeric_limits<double>::infinity();
    if (!(x >= 0.0)) return std::numeric_limits<double>::quiet_NaN();
    if (x == 0.0) return -inf;
    if (x == inf) return inf;
    if (x < 2.2250738585072014e-308) x = 2.2250738585072014e-308;
    int ei; double m = frexp_(x, ei);
    bool ge = m >= 0.70710678118654752440;
    if (!ge) ei -= 1;
    double e = (double) ei, r;
    if (ei > 2 || ei < -2) {
        /* log(x) = z + z^3 P(z^2)/Q(z^2), z = 2(x-1)/(x+1) */
        double z, y;
        if (ge) { z = (m - 0.5) - 0.5; y = fma_(0.5, m, 0.5); }
        else    { z = m - 0.5;         y = fma_(0.5, z, 0.5); }
        double x2 = z / y;
        z = x2 * x2;
        static const double P[3] = { -6.41409952958715622951e1, 1.63866645699558079767e1,
                                     -7.89580278884799154124e-1 };
        static const double Q[4] = { -7.69691943550460008604e2, 3.12093766372244180303e2,
                                     -3.56722798256324312549e1, 1.0 };
        z = x2 * (z * horner(z, P) / horner(z, Q));
        r = fma_(-e, 2.121944400546905827679e-4, z) + x2;
    } else {
        /* log(1+x) = x - x^2/2 + x^3 P(x)/Q(x) */
        double x2 = (ge ? m : m + m) - 1.0;
        double z = x2 * x2;
        static const double P[6] = { 7.70838733755885391666e0, 1.79368678507819816313e1,
                                     1.44989225341610930846e1, 4.70579119878881725854e0,
                                     4.97494994976747001425e-1, 1.01875663804580931796e-4 };
        static const double Q[6] = { 2.31251620126765340583e1, 7.11544750618563894466e1,
                                     8.29875266912776603211e1, 4.52279145837532221105e1,
                                     1.12873587189167450590e1, 1.0 };
        double y = x2 * (z * horner(x2, P) / horner(x2, Q));
        y = fma_(-e, 2.121944400546905827679e-4, y);
        r = x2 + fma_(-0.5, z, y);
    }
    return fma_(e, 0.693359375, r);
}

/* ------------------------------------------------------------------ exp -- */
inline float exp_(float x) {
    if (x > 88.3762588501f) return std::numeric_limits<float>::infinity();
    if (x < -88.3762588501f) return 0.f;
    if (x != x) return x;
    float n = std::floor(fma_(1.4426950408889634073599f, x, 0.5f));
    float xr = fma_(-n, 0.693359375f, x);
    xr = fma_(-n, -2.12194440e-4f, xr);
    static const float c[6] = { 5.0000001201e-1f, 1.6666665459e-1f, 4.1665795894e-2f,
                                8.3334519073e-3f, 1.3981999507e-3f, 1.9875691500e-4f };
    float z = horner(xr, c);
    z = fma_(z, xr * xr, xr + 1.f);
    return ldexp_(z, (int) n);
}

inline double exp_(double x) {
    if (x > 7.0943613930310391424428e2) return std::numeric_limits<double>::infinity();
    if (x < -7.0943613930310391424428e2) return 0.0;
    if (x != x) return x;
    double n = std::floor(fma_(1.4426950408889634073599, x, 0.5));
    double xr = fma_(-n, 6.93145751953125e-1, x);
    xr = fma_(-n, 1.42860682030941723212e-6, xr);
    double z = xr * xr;
    static const double P[3] = { 9.99999999999999999910e-1, 3.02994407707441961300e-2,
                                 1.26177193074810590878e-4 };
    static const double Q[4] = { 2.00000000000000000009e0, 2.27265548208155028766e-1,
                                 2.52448340349684104192e-3, 3.00198505138664455042e-6 };
    double p = horner(z, P) * xr;
    double q = horner(z, Q);
    double pq = p / (q - p);
    z = pq + pq + 1.0;
    return ldexp_(z, (int) n);
}

template <typename T> inline T pow_(T x, T y) { return exp_(log_(x) * y); }

/* --------------------------------------------------------------- sincos -- */
/* CEPHES joint sine/cosine: octant reduction with a three-part pi/4, then the
   degree-3 (float) / degree-6 (double) minimax polynomials in y^2.           */
template <typename T> struct SinCosConst;
template <> struct SinCosConst<float> {
    static constexpr float dp1 = 0.78515625f, dp2 = 2.4187564849853515625e-4f,
                           dp3 = 3.77489497744594108e-8f;
    static float sp(float z) {
        static const float c[3] = { -1.6666654611e-1f, 8.3321608736e-3f, -1.9515295891e-4f };
        return horner(z, c);
    }
    static float cp(float z) {
        static const float c[3] = { 4.166664568298827e-2f, -1.388731625493765e-3f,
                                    2.443315711809948e-5f };
        return horner(z, c);
    }
};
template <> struct SinCosConst<double> {
    static constexpr double dp1 = 7.85398125648498535156e-1, dp2 = 3.77489470793079817668e-8,
                            dp3 = 2.69515142907905952645e-15;
    static double sp(double z) {
        static const double c[6] = { -1.66666666666666307295e-1, 8.33333333332211858878e-3,
                                     -1.98412698295895385996e-4, 2.75573136213857245213e-6,
                                     -2.50507477628578072866e-8, 1.58962301576546568060e-10 };
        return horner(z, c);
    }
    static double cp(double z) {
        static const double c[6] = { 4.16666666666665929218e-2, -1.38888888888730564116e-3,
                                     2.48015872888517045348e-5, -2.75573141792967388112e-7,
                                     2.08757008419747316778e-9, -1.13585365213876817300e-11 };
        return horner(z, c);
    }
};

/* Returns (sin x, cos x); |x| must stay below ~8192 (it does on this path:
   arguments are 2*pi*u, pi*u or the concentric-disk angle).                  */
template <typename T> inline std::pair<T, T> sincos_(T x) {
    using K = SinCosConst<T>;
    T xa = x < 0 ? -x : x;
    int j = (int) (xa * (T) 1.2732395447351626862);
    j = (j + 1) & ~1;
    T y = (T) j;
    y = ((xa - y * K::dp1) - y * K::dp2) - y * K::dp3;
    T z = y * y;
    T s = K::sp(z) * z;
    T c = K::cp(z) * z;
    s = fma_(s, y, y);
    c = fma_(c, z, fma_(z, (T) -0.5, (T) 1));
    bool swap = (j & 2) != 0;
    T rs = swap ? c : s, rc = swap ? s : c;
    /* sign of sine: octant bit 2 xor sign of x; cosine: octants 2..5 negative */
    if (((j & 4) != 0) != (x < 0)) rs = -rs;
    if (((j - 2) & 4) == 0) rc = -rc;
    return { rs, rc };
}

/* ----------------------------------------------------------------- cbrt -- */
template <typename T> inline T cbrt_(T x) {
    if (x == 0 || x != x || x == std::numeric_limits<T>::infinity() ||
        x == -std::numeric_limits<T>::infinity())
        return x;
    const T CBRT2 = (T) 1.25992104989487316477, CBRT4 = (T) 1.58740105196819947475,
            THIRD = (T) (1.0 / 3.0);
    T xa = x < 0 ? -x : x;
    if (xa < std::numeric_limits<T>::min()) xa = std::numeric_limits<T>::min();
    int e; T m = frexp_(xa, e);
    static const T c[5] = { (T) 0.40238979564544752126924, (T) 1.1399983354717293273738,
                            (T) -0.95438224771509446525043, (T) 0.54664601366395524503440,
                            (T) -0.13466110473359520655053 };
    m = horner(m, c);
    int ea = e < 0 ? -e : e;
    int q = ea / 3, rem = ea - 3 * q;
    if (rem == 1) m = m * (e >= 0 ? CBRT2 : (T) 1 / CBRT2);
    else if (rem == 2) m = m * (e >= 0 ? CBRT4 : (T) 1 / CBRT4);
    T r = ldexp_(m, e >= 0 ? q : -q);
    if (x < 0) r = -r;
    r = r - (r - (x / (r * r))) * THIRD;
    if (sizeof(T) == 8)
        r = r - (r - (x / (r * r))) * THIRD;
    return r;
}

template <typename T> inline T sqrt_(T x) { return std::sqrt(x); }
template <typename T> inline T safe_sqrt_(T x) { return std::sqrt(x > 0 ? x : (T) 0); }

} // namespace omath
