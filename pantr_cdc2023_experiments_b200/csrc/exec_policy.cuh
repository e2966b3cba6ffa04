// Execution policies of the batched solver: how one MPC instance maps onto the hardware.
//
//   WarpExec    one warp = one instance.  Vectors are plain arrays, element i belongs to lane i % 32,
//               reductions = 32 lane-strided partial sums + xor butterfly.  Used where the model
//               itself has ~32-way parallelism per stage (hanging chain: 30 position components).
//   ThreadExec  one thread = one instance, 32 instances per warp.  Vectors are interleaved across the
//               warp (element i of lane t at p[32 i + t]) so that every access is one coalesced 256-byte
//               transaction.  Used for the small-state models (quadcopters: nx = 9), whose sweeps are
//               sequential chains with no intra-instance parallelism worth a shuffle.
//
// Both policies produce the same bits: ThreadExec's reductions keep the 32 strided partial sums of
// the warp layout in registers and combine them in the butterfly's order (which is also the order
// the CPU oracle emulates).
#pragma once
#include "warp_utils.cuh"
#include <type_traits>

namespace pantr {

// (An element proxy that loaded / stored with ld/st.global.cg, to keep L1 for the threads' local memory,
// measured 6x SLOWER -- 0.90 M instead of 5.2 M Hessian-vector products / s: the plain loads' 35 % L1 hit
// rate on the interleaved vectors is worth far more than the spill traffic it displaces.)
/// Strided view: element i at p[32 i].
template <class T>
struct SVecT {
    T *p;
    __device__ __forceinline__ T &operator[](int i) const { return p[i * 32]; }
    __device__ __forceinline__ SVecT operator+(int k) const { return SVecT{p + k * 32}; }
    __device__ __forceinline__ SVecT &operator+=(int k) { p += k * 32; return *this; }
    __device__ __forceinline__ explicit operator bool() const { return p != nullptr; }
    template <class U, class = std::enable_if_t<std::is_same<U, const T>::value && !std::is_const<T>::value>>
    __device__ __forceinline__ operator SVecT<U>() const { return SVecT<U>{p}; }
};

struct WarpExec {
    static constexpr bool THREAD = false;
    static constexpr int STEP = 32;
    using Vec = double *;
    using CVec = const double *;
    using IVec = int *;
    __device__ static __forceinline__ int first() { return lane_id(); }
    __device__ static __forceinline__ void sync() { __syncwarp(); }
    __device__ static __forceinline__ bool leader() { return lane_id() == 0; }
    __device__ static __forceinline__ bool all(bool ok) { return __all_sync(FULL, ok); }
    template <class F>
    __device__ static __forceinline__ double sum(int n, F f) {
        double p = 0;
        for (int i = lane_id(); i < n; i += 32) p = p + f(i);
        return warp_sum(p);
    }
    template <class F>
    __device__ static __forceinline__ double nanmax_of(int n, F f) {
        double m = 0;
        for (int i = lane_id(); i < n; i += 32) m = nanmax(m, f(i));
        return warp_nanmax(m);
    }
    /// store(i, load(i)) for this lane's elements.  BATCHED issues four elements' loads before the first store (written
    /// as a plain loop, every store through a double* may alias the next element's operands, so the compiler keeps
    /// load -> store -> load order).  Measured on one box (profiles/r02_m_variants.log): hanging chain +1.2 %, quadcopters
    /// -3 % (larger code, more registers live), hence a per-model switch (Model::BATCHED_MAPS).
    /// load must not read what store writes at OTHER indices.
    template <bool BATCHED = false, class L, class S>
    __device__ static __forceinline__ void map(int n, L load, S store) {
        if constexpr (!BATCHED) {
            for (int i = lane_id(); i < n; i += 32) store(i, load(i));
        } else {
            for (int i = lane_id(); i < n; i += 128) {
                const bool b1 = i + 32 < n, b2 = i + 64 < n, b3 = i + 96 < n;
                const auto v0 = load(i);
                auto v1 = v0, v2 = v0, v3 = v0;
                if (b1) v1 = load(i + 32);
                if (b2) v2 = load(i + 64);
                if (b3) v3 = load(i + 96);
                store(i, v0);
                if (b1) store(i + 32, v1);
                if (b2) store(i + 64, v2);
                if (b3) store(i + 96, v3);
            }
        }
    }
    /// ascending compaction of {i < n : pred(i)} into out[off..]; returns the count
    template <class F>
    __device__ static __forceinline__ int compact(int n, int *out, int off, F pred) {
        const int lane = lane_id();
        const unsigned lt = (1u << lane) - 1u;
        int cnt = 0;
        for (int i0 = 0; i0 < n; i0 += 32) {
            int i = i0 + lane;
            bool in = i < n && pred(i);
            unsigned b = __ballot_sync(FULL, in);
            if (in) out[off + cnt + __popc(b & lt)] = i;
            cnt += __popc(b);
        }
        return cnt;
    }
};

struct ThreadExec {
    static constexpr bool THREAD = true;
    static constexpr int STEP = 1;
    using Vec = SVecT<double>;
    using CVec = SVecT<const double>;
    using IVec = SVecT<int>;
    __device__ static __forceinline__ int first() { return 0; }
    __device__ static __forceinline__ void sync() {}
    __device__ static __forceinline__ bool leader() { return true; }
    __device__ static __forceinline__ bool all(bool ok) { return ok; }
    template <class F>
    __device__ static __forceinline__ double sum(int n, F f) {
        double acc[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[j] = 0;
        int base = 0;
        for (; base + 32 <= n; base += 32) {
#pragma unroll
            for (int j = 0; j < 32; ++j) acc[j] = acc[j] + f(base + j);
        }
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (base + j < n) acc[j] = acc[j] + f(base + j);
        // xor butterfly: lane j ends with ((((p_j + p_{j^16}) + ..^8) + ..^4) + ..^2) + ..^1
#pragma unroll
        for (int m = 16; m >= 1; m >>= 1) {
#pragma unroll
            for (int j = 0; j < m; ++j) acc[j] = acc[j] + acc[j + m];
        }
        return acc[0];
    }
    template <class F>
    __device__ static __forceinline__ double nanmax_of(int n, F f) {
        double m = 0;
        for (int i = 0; i < n; ++i) m = nanmax(m, f(i));
        return m;
    }
    template <bool BATCHED = false, class L, class S>
    __device__ static __forceinline__ void map(int n, L load, S store) {
        for (int i = 0; i < n; ++i) store(i, load(i));
    }
    template <class F>
    __device__ static __forceinline__ int compact(int n, IVec out, int off, F pred) {
        int cnt = 0;
        for (int i = 0; i < n; ++i)
            if (pred(i)) { out[off + cnt] = i; ++cnt; }
        return cnt;
    }
};

// ---- vector helpers shared by both policies ----
template <class E, class A, class B>
__device__ __forceinline__ double xdot(A a, B b, int n) {
    return E::sum(n, [&](int i) { return a[i] * b[i]; });
}
template <class E, class A>
__device__ __forceinline__ double xsqnorm(A a, int n) { return xdot<E>(a, a, n); }
template <class E, class A>
__device__ __forceinline__ double xnorm2(A a, int n) { return sqrt(xsqnorm<E>(a, n)); }
template <class E, class A>
__device__ __forceinline__ double xnorm_inf(A a, int n) {
    return E::nanmax_of(n, [&](int i) { return fabs(a[i]); });
}
template <class E, class A>
__device__ __forceinline__ double xnorm1(A a, int n) {
    return E::sum(n, [&](int i) { return fabs(a[i]); });
}
template <class E, class A>
__device__ __forceinline__ bool xall_finite(A a, int n) {
    bool ok = true;
    for (int i = E::first(); i < n; i += E::STEP) ok = ok && isfinite(a[i]);
    return E::all(ok);
}
template <class E, bool BATCHED = false, class D, class S>
__device__ __forceinline__ void xcopy(D dst, S src, int n) {
    E::template map<BATCHED>(n, [&](int i) { return src[i]; }, [&](int i, double v) { dst[i] = v; });
}

} // namespace pantr
