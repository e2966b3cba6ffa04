// Warp-level primitives for the batched PANTR solver: one warp owns one MPC instance.
// Vectors of length n live in the instance's workspace; element i belongs to lane i % 32.
// Reductions use 32 strided partial sums + xor butterfly, so every lane ends up with the same
// bits (and the same bits as the CPU oracle, which emulates this order).
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace pantr {

constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ double shfl(double v, int src) {
    int lo = __double2loint(v), hi = __double2hiint(v);
    lo = __shfl_sync(FULL, lo, src);
    hi = __shfl_sync(FULL, hi, src);
    return __hiloint2double(hi, lo);
}

__device__ __forceinline__ double warp_sum(double p) {
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) p = p + __shfl_xor_sync(FULL, p, m);
    return p;
}
/// max with NaN propagation (matches the oracle's norm_inf)
__device__ __forceinline__ double nanmax(double a, double b) {
    return (isnan(a) || isnan(b)) ? nan("") : (a > b ? a : b);
}
__device__ __forceinline__ double warp_nanmax(double p) {
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) p = nanmax(p, __shfl_xor_sync(FULL, p, m));
    return p;
}

__device__ __forceinline__ double vdot(const double *a, const double *b, int n) {
    double p = 0;
    for (int i = lane_id(); i < n; i += 32) p = p + a[i] * b[i];
    return warp_sum(p);
}
__device__ __forceinline__ double vsqnorm(const double *a, int n) { return vdot(a, a, n); }
__device__ __forceinline__ double vnorm2(const double *a, int n) { return sqrt(vsqnorm(a, n)); }
__device__ __forceinline__ double vnorm_inf(const double *a, int n) {
    double m = 0;
    for (int i = lane_id(); i < n; i += 32) m = nanmax(m, fabs(a[i]));
    return warp_nanmax(m);
}
__device__ __forceinline__ bool vall_finite(const double *a, int n) {
    bool ok = true;
    for (int i = lane_id(); i < n; i += 32) ok = ok && isfinite(a[i]);
    return __all_sync(FULL, ok);
}
__device__ __forceinline__ void vcopy(double *dst, const double *src, int n) {
    for (int i = lane_id(); i < n; i += 32) dst[i] = src[i];
}

} // namespace pantr
