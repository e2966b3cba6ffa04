// Host/device helpers shared by the CUDA kernels and the CPU oracle (the oracle includes the
// model headers of this directory so that both sides evaluate the same operation sequence).
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define PANTR_HD __host__ __device__ __forceinline__
#define PANTR_UNROLL _Pragma("unroll")
#define PANTR_NOUNROLL _Pragma("unroll 1")
#else
#define PANTR_HD inline
#define PANTR_UNROLL _Pragma("GCC unroll 16")
#define PANTR_NOUNROLL _Pragma("GCC unroll 1")
#endif

namespace pantr {

/// sin and cos of x in double precision, written out so that the GPU and the CPU oracle produce
/// the same bits (libdevice and glibc differ in the last ulp; CasADi's generated code calls libm's
/// sin/cos, from which this differs by <= 2 ulp, i.e. ~1e-16 relative, eight orders below the
/// parity tolerance).  Three-term Cody-Waite reduction by pi/2 (exact products for |x| < 1.6e6,
/// the constants carry 33 bits each) and the classic degree-13/14 minimax kernels on [-pi/4, pi/4].
/// |x| >= 1.6e6 or non-finite x gives NaN (the solver then reports NotFinite).
PANTR_HD void pantr_sincos(double x, double *s, double *c) {
    if (!(fabs(x) < 1.6e6)) { *s = x - x; *c = x - x; if (*s == 0) { *s = NAN; *c = NAN; } return; }
    const double fn = rint(x * 6.36619772367581382433e-01);
    const int q = (int)fn;
    double r = x - fn * 1.57079632673412561417e+00;
    r = r - fn * 6.07710050630396597660e-11;
    r = r - fn * 2.02226624871116645580e-21;
    const double z = r * r;
    const double ps = -1.66666666666666324348e-01 +
                      z * (8.33333333332248946124e-03 +
                           z * (-1.98412698298579493134e-04 +
                                z * (2.75573137070700676789e-06 +
                                     z * (-2.50507602534068634195e-08 + z * 1.58969099521155010221e-10))));
    const double pc = 4.16666666666666019037e-02 +
                      z * (-1.38888888888741095749e-03 +
                           z * (2.48015872894767294178e-05 +
                                z * (-2.75573143513906633035e-07 +
                                     z * (2.08757232129817482790e-09 + z * -1.13596475577881948265e-11))));
    const double sr = r + (r * z) * ps;
    const double cr = (1.0 - 0.5 * z) + (z * z) * pc;
    switch (q & 3) {
        case 0: *s = sr; *c = cr; break;
        case 1: *s = cr; *c = -sr; break;
        case 2: *s = -sr; *c = -cr; break;
        default: *s = -cr; *c = sr; break;
    }
}

/// The same evaluation with every multiply-add written as an explicit fma (reduction, both Horner chains, the final
/// combinations): half the dependent FP64 operations per call, and |error| <= 1 ulp of the exactly reduced argument's
/// sin / cos.  Used by the condensed quadcopter formulation (device and oracle mirror); bit-reproducible like
/// pantr_sincos.  On the device the 16 coefficients are read from constant memory (an FP64 instruction takes a constant
/// bank operand for free, a literal costs two UMOV issue slots); the host reads the same literals.
#define PANTR_SINCOS_COEFFS(X)                                                                                 \
    X(0, 6.36619772367581382433e-01)   /* 2 / pi */                                                            \
    X(1, 1.57079632673412561417e+00) X(2, 6.07710050630396597660e-11) X(3, 2.02226624871116645580e-21)         \
    X(4, 1.58969099521155010221e-10) X(5, -2.50507602534068634195e-08) X(6, 2.75573137070700676789e-06)        \
    X(7, -1.98412698298579493134e-04) X(8, 8.33333333332248946124e-03) X(9, -1.66666666666666324348e-01)       \
    X(10, -1.13596475577881948265e-11) X(11, 2.08757232129817482790e-09) X(12, -2.75573143513906633035e-07)    \
    X(13, 2.48015872894767294178e-05) X(14, -1.38888888888741095749e-03) X(15, 4.16666666666666019037e-02)
#if defined(__CUDACC__)
#define PANTR_SC_ENTRY(i, v) v,
static __constant__ double pantr_sincos_k[16] = {PANTR_SINCOS_COEFFS(PANTR_SC_ENTRY)};
#undef PANTR_SC_ENTRY
#endif
PANTR_HD double pantr_sincos_coeff(int i) {
#if defined(__CUDA_ARCH__)
    return pantr_sincos_k[i];
#else
#define PANTR_SC_CASE(k, v) case k: return v;
    switch (i) { PANTR_SINCOS_COEFFS(PANTR_SC_CASE) default: return 0; }
#undef PANTR_SC_CASE
#endif
}
PANTR_HD void pantr_sincos_fma(double x, double *s, double *c) {
#define K(i) pantr_sincos_coeff(i)
    const bool ok = fabs(x) < 1.6e6;
    const double fn = rint(x * K(0));
    const int q = ok ? (int)fn : 0;
    double r = fma(-fn, K(1), x);
    r = fma(-fn, K(2), r);
    r = fma(-fn, K(3), r);
    if (!ok) r = NAN;
    const double z = r * r;
    double ps = fma(z, K(4), K(5));
    ps = fma(z, ps, K(6));
    ps = fma(z, ps, K(7));
    ps = fma(z, ps, K(8));
    ps = fma(z, ps, K(9));
    double pc = fma(z, K(10), K(11));
    pc = fma(z, pc, K(12));
    pc = fma(z, pc, K(13));
    pc = fma(z, pc, K(14));
    pc = fma(z, pc, K(15));
#undef K
    const double sr = fma(r * z, ps, r);
    const double cr = fma(z * z, pc, fma(-0.5, z, 1.0));
    const bool swap = q & 1;
    const double sv = swap ? cr : sr, cv = swap ? sr : cr;
    *s = (q & 2) ? -sv : sv;
    *c = ((q + 1) & 2) ? -cv : cv;
}

} // namespace pantr
