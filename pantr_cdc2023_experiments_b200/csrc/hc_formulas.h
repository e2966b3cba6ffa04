// Scalar formulas of the hanging-chain sweeps with their fused multiply-adds written out, shared by
// the CUDA kernels (hanging_chain.cuh) and the CPU oracle (oracle/hanging_chain.hpp) so that both
// contract exactly the same operations (the compilers' own contraction is off on both sides:
// --fmad=false / -ffp-contract=off).  The sweeps are bound by the latency of dependent FP64
// operations, so every mul+add pair folded into one DFMA shortens the critical path.
// Model: python/alpaqa_mpc_benchmarks/problems/hanging_chain.py:9-18,47-80 (reference).
#pragma once
#include "pantr_hd.h"

namespace pantr {
namespace hcf {

constexpr double alpha = 25, beta = 1, gamma_u = 0.01, mass = 0.03, Dspring = 0.1, Lspring = 0.033,
                 g_grav = 9.81, Ts = 0.1;
constexpr double inv_m = 1.0 / mass, DL = Dspring * Lspring;
constexpr double h = Ts, hh = Ts / 2, h6 = Ts / 6, h3 = 2 * (Ts / 6);
constexpr double two_alpha = 2 * alpha, two_beta = 2 * beta, two_gamma = 2 * gamma_u;

/// c = D (1 - L / |dist|)
PANTR_HD double spring_c(double iw) { return Dspring * fma(-Lspring, iw, 1.0); }
/// e = D L / |dist|^3
PANTR_HD double spring_e(double iw, double iw2) { return (DL * iw) * iw2; }
/// (F_next - F) / m + g
PANTR_HD double accel(double Fu, double F, double grav) { return fma(Fu - F, inv_m, grav); }
/// x + a k
PANTR_HD double axpy(double a, double k, double x) { return fma(a, k, x); }
/// a b + c d
PANTR_HD double lin2(double a, double b, double c, double d) { return fma(c, d, a * b); }
/// RK4 combination x + h/6 (k1 + 2 k2 + 2 k3 + k4)
PANTR_HD double rk4c(double x, double k1, double k2, double k3, double k4) {
    return fma(h6, fma(2.0, k3, fma(2.0, k2, k1)) + k4, x);
}
/// c a + (e t) d : first-order force term (vjp: a = Fbar, t = dist.Fbar ; jvp: a = ddist, t = dist.ddist)
PANTR_HD double force_lin(double c, double a, double e, double t, double d) { return fma(e * t, d, c * a); }
/// second-order part of the stage adjoint
PANTR_HD double force_sec(double dc, double Fbar, double de, double dotv, double e, double df, double dist,
                          double ddist) {
    const double t1 = fma(de, dotv, e * df);
    const double t2 = fma(t1, dist, dc * Fbar);
    return fma(e * dotv, ddist, t2);
}

} // namespace hcf
} // namespace pantr
