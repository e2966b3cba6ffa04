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
/// combinations): 20 fewer FP64 instructions per call, and |error| <= 1 ulp of the exactly reduced argument's sin / cos.
/// Used by the condensed quadcopter formulation (device and oracle mirror); bit-reproducible like pantr_sincos.
PANTR_HD void pantr_sincos_fma(double x, double *s, double *c) {
    const bool ok = fabs(x) < 1.6e6;
    const double fn = rint(x * 6.36619772367581382433e-01);
    const int q = ok ? (int)fn : 0;
    double r = fma(-fn, 1.57079632673412561417e+00, x);
    r = fma(-fn, 6.07710050630396597660e-11, r);
    r = fma(-fn, 2.02226624871116645580e-21, r);
    if (!ok) r = NAN;
    const double z = r * r;
    double ps = fma(z, 1.58969099521155010221e-10, -2.50507602534068634195e-08);
    ps = fma(z, ps, 2.75573137070700676789e-06);
    ps = fma(z, ps, -1.98412698298579493134e-04);
    ps = fma(z, ps, 8.33333333332248946124e-03);
    ps = fma(z, ps, -1.66666666666666324348e-01);
    double pc = fma(z, -1.13596475577881948265e-11, 2.08757232129817482790e-09);
    pc = fma(z, pc, -2.75573143513906633035e-07);
    pc = fma(z, pc, 2.48015872894767294178e-05);
    pc = fma(z, pc, -1.38888888888741095749e-03);
    pc = fma(z, pc, 4.16666666666666019037e-02);
    const double sr = fma(r * z, ps, r);
    const double cr = fma(z * z, pc, fma(-0.5, z, 1.0));
    const bool swap = q & 1;
    const double sv = swap ? cr : sr, cv = swap ? sr : cr;
    *s = (q & 2) ? -sv : sv;
    *c = ((q + 1) & 2) ? -cv : cv;
}

} // namespace pantr
