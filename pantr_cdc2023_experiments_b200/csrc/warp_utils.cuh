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

// ---- bulk asynchronous copies (TMA, cp.async.bulk) completing on a shared-memory mbarrier ----
__device__ __forceinline__ unsigned smem_addr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
/// orders this thread's (and, after a __syncwarp, the warp's) generic-proxy accesses -- the rows / AB written with plain
/// stores to global memory, the reads of a ring slot about to be refilled -- before its following async-proxy copies
__device__ __forceinline__ void proxy_async_fence() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
/// global -> shared bulk copy (TMA), bytes a multiple of 16, both addresses 16-byte aligned; completes on `bar`
__device__ __forceinline__ void bulk_g2s(double *dst, const double *src, unsigned bytes, unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(dst)),
                 "l"(src), "r"(bytes), "r"(smem_addr(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_a(unsigned bar_addr, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_addr), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s_a(unsigned dst_addr, const double *src, unsigned bytes, unsigned bar_addr) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_addr), "l"(src),
                 "r"(bytes), "r"(bar_addr)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait_a(unsigned bar_addr, unsigned parity) {
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "WAIT_%=:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                 "@p bra DONE_%=;\n"
                 "bra WAIT_%=;\n"
                 "DONE_%=:\n"
                 "}" ::"r"(bar_addr), "r"(parity)
                 : "memory");
}
__device__ __forceinline__ void cp_async_16_a(unsigned dst_addr, const double *g) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_addr), "l"(g) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "WAIT_%=:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                 "@p bra DONE_%=;\n"
                 "bra WAIT_%=;\n"
                 "DONE_%=:\n"
                 "}" ::"r"(smem_addr(bar)), "r"(parity)
                 : "memory");
}

// ---- shared-memory accesses by 32-bit shared-window address (the recurrences keep running addresses in registers;
// through generic pointers the compiler rebuilt the window base -- S2R SR_CgaCtaId + LEA -- in every step) ----
__device__ __forceinline__ double lds_f64(unsigned a) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ double2 lds_v2f64(unsigned a) {
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts_f64(unsigned a, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory"); }
/// COUNT consecutive doubles from a 16-byte aligned shared address
template <int COUNT>
__device__ __forceinline__ void lds_vec(unsigned a, double *out) {
#pragma unroll
    for (int i = 0; i + 1 < COUNT; i += 2) { const double2 v = lds_v2f64(a + 8 * i); out[i] = v.x; out[i + 1] = v.y; }
    if (COUNT & 1) out[COUNT - 1] = lds_f64(a + 8 * (COUNT - 1));
}

// ---- per-thread asynchronous copies (cp.async / LDGSTS: generic proxy, L2 -> shared) ----
__device__ __forceinline__ void cp_async_16(double *smem, const double *g) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr(smem)), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_commit_group() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait_group() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

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
