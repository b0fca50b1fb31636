// TEST INFRASTRUCTURE -- libm interposer for the reference build (oracle/_ref/libb2vp_refmath.so, LD_PRELOADed
// by oracle/ref_xml.py when exact_math=True).  glibc's logf / expf / sincosf / cbrtf are the one external
// arithmetic dependency of the reference's volpath path ("hard parts"): results are
// libm-version dependent at the ulp level, and no GPU has glibc.  With this library in front of libm the
// UNMODIFIED reference binaries evaluate those functions with the fixed IEEE operation sequences of
// oracle_math.h (the ones csrc/vp_math.cuh implements on the device), so that reference, oracle and CUDA
// kernel can be compared bit for bit.  Without it the reference runs on stock glibc and agrees to ~1e-7
// relative per path (tests/test_reference_parity.py checks both).
#include "oracle_math.h"

extern "C" {
float logf(float x) { return omath::log_(x); }
double log(double x) { return omath::log_(x); }
float expf(float x) { return omath::exp_(x); }
double exp(double x) { return omath::exp_(x); }
void sincosf(float x, float *s, float *c) { auto r = omath::sincos_(x); *s = r.first; *c = r.second; }
void sincos(double x, double *s, double *c) { auto r = omath::sincos_(x); *s = r.first; *c = r.second; }
float cbrtf(float x) { return omath::cbrt_(x); }
double cbrt(double x) { return omath::cbrt_(x); }
}
