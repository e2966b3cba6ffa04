// pantr_b200: batched PANTR+ALM on B200 (sm_100a, FP64) — kernels and C ABI (include/pantr_b200.h).
// One warp-sized thread block owns one MPC instance for the whole solve (ALM outer loop, PANTR inner
// loop, Steihaug-CG, problem evaluations); blocks are persistent and pull instances from a queue.
#include "pantr_device.cuh"
#include "hanging_chain.cuh"
#include "quadcopter.cuh"
#include "quadcopter_condensed.cuh"
#include "problem_face.cuh"
#include <cstdio>
#include <cstring>
#include <new>
#include <mutex>
#include <string>
#include <vector>

using namespace pantr;

namespace {

thread_local std::string g_last_error;
int fail(int code, const std::string &msg) { g_last_error = msg; return code; }
#define CUDA_TRY(expr)                                                                            \
    do {                                                                                          \
        cudaError_t e_ = (expr);                                                                  \
        if (e_ != cudaSuccess)                                                                    \
            return fail(PANTR_B200_CUDA_ERROR, std::string(#expr) + ": " + cudaGetErrorString(e_)); \
    } while (0)

struct KernelArgs {
    int N, n, m, nx;
    long long B;
    const double *x0;
    double *u, *y;
    pantr_b200_stats *stats;
    double *ws;
    size_t ws_stride; // doubles per block
    unsigned long long *counter;
    Trace trace;
    int smem_set; // which solver vectors live in dynamic shared memory (Pantr::smem_bytes)
    const int *order; // optional: queue ticket -> instance index (longest-expected-first scheduling)
};

// ---- the solve kernel: ALMSolver::operator() for every instance of the batch ----
template <class Model>
__global__ void __launch_bounds__(32, Model::MIN_BLOCKS) solve_kernel(const __grid_constant__ SolverParams P,
                                                   const __grid_constant__ KernelArgs A) {
    __shared__ IterScalars it[3];
    // The model and solver objects live in SHARED memory where the model asks for it (OBJECTS_IN_SHARED).  As locals they are address-taken (their routines are
    // __noinline__ members), i.e. in local memory, one private copy per lane: ~0.5 KB x 32 lanes = 16 KB of L1 footprint
    // per warp, re-read with LDL after every call and regularly evicted by the streamed workspace (the solver routines
    // held ~20 % of the stall samples with ~5 % of the instructions, profiles/r02_e_quadc_ncu_*).  All scalars are
    // warp-uniform, every lane writes the same values; read-modify-writes happen while the warp is converged.
    // (Measured, profiles/r02_j_*: quadcopters +1.5 %, hanging chain -1.9 % -- its 12 warps per SM keep their copies
    // in L1 and a generic load from shared memory is no faster than an LDL hit -- hence the per-model switch.)
    constexpr size_t MDL_BYTES = (sizeof(Model) + 15) / 16 * 16, OBJ_BYTES = MDL_BYTES + sizeof(Pantr<Model>);
    constexpr bool SH = Model::OBJECTS_IN_SHARED;
    __shared__ __align__(16) unsigned char obj_shared[SH ? OBJ_BYTES : 16];
    alignas(16) unsigned char obj_local[SH ? 16 : OBJ_BYTES];
    unsigned char *const obj_mem = SH ? obj_shared : obj_local;
    extern __shared__ __align__(16) double dyn_smem[];
    double *ws = A.ws + size_t(blockIdx.x) * A.ws_stride;
    for (;;) {
        unsigned long long b = 0;
        if (lane_id() == 0) b = atomicAdd(A.counter, 1ULL);
        b = __shfl_sync(FULL, b, 0);
        if (b >= (unsigned long long)A.B) break;
        if (A.order) { // a hint entry outside 0..B-1 is skipped, never dereferenced as an instance
            const int o = A.order[b];
            if (o < 0 || (long long)o >= A.B) continue;
            b = (unsigned long long)o;
        }
        long long t0 = clock64();
        const unsigned long long g0 = globaltimer_ns();
        __syncwarp();
        Model &mdl = *new (obj_mem) Model;
        mdl.init(A.N, A.x0 + b * A.nx, ws);
        Pantr<Model> &S = *new (obj_mem + MDL_BYTES) Pantr<Model>(mdl, P, it, A.n, A.m, ws + Model::traj_doubles(A.N), dyn_smem, A.smem_set);
        memset(&S.st, 0, sizeof(S.st));
        S.x = A.u + b * A.n;
        S.tracing = A.trace.rows != nullptr && (long long)b == A.trace.instance;
        S.tr = A.trace;
        __syncwarp();
        for (int i = lane_id(); i < A.m; i += 32) S.y[i] = A.y[b * A.m + i];
        __syncwarp();
        S.alm_solve();
        __syncwarp();
        for (int i = lane_id(); i < A.m; i += 32) A.y[b * A.m + i] = S.y[i];
        __syncwarp();
        S.st.cycles = (unsigned long long)(clock64() - t0);
        S.st.elapsed_us = (unsigned)((globaltimer_ns() - g0) / 1000);
        __syncwarp();
        if (lane_id() == 0) A.stats[b] = S.st;
        __syncwarp();
    }
}

// ---- thread-per-instance variant (ThreadExec, exec_policy.cuh): 32 instances per warp, vectors
// interleaved across the warp.  For the small-state models whose sweeps are sequential chains. ----
// (Round 1 also had an iteration-sliced variant of solve_kernel -- PANTR iterations handed out through a device-side
// MPMC ring, per-instance iterate slots in HBM: bit-identical, measured equal or slower, removed.)
#ifndef PANTR_THREAD_MIN_BLOCKS
#define PANTR_THREAD_MIN_BLOCKS 8 // resident warps per SM the thread-per-instance kernel is compiled for
#endif
template <class Model>
__global__ void __launch_bounds__(32, PANTR_THREAD_MIN_BLOCKS) solve_kernel_thread(const __grid_constant__ SolverParams P,
                                                                             const __grid_constant__ KernelArgs A) {
    IterScalars it[3];
    const int lane = lane_id();
    SVecT<double> ws{A.ws + size_t(blockIdx.x) * A.ws_stride + lane};
    for (;;) {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(A.counter, 32ULL);
        base = __shfl_sync(FULL, base, 0);
        if (base >= (unsigned long long)A.B) break;
        const long long b = (long long)base + lane;
        if (b < A.B) {
            long long t0 = clock64();
            const unsigned long long g0 = globaltimer_ns();
            Model mdl;
            mdl.init(A.N, A.x0 + b * A.nx, ws);
            Pantr<Model, ThreadExec> S(mdl, P, it, A.n, A.m, ws + Model::traj_doubles(A.N));
            memset(&S.st, 0, sizeof(S.st));
            for (int i = 0; i < A.n; ++i) S.x[i] = A.u[b * A.n + i];
            for (int i = 0; i < A.m; ++i) S.y[i] = A.y[b * A.m + i];
            S.tracing = A.trace.rows != nullptr && b == A.trace.instance;
            S.tr = A.trace;
            S.alm_solve();
            for (int i = 0; i < A.n; ++i) A.u[b * A.n + i] = S.x[i];
            for (int i = 0; i < A.m; ++i) A.y[b * A.m + i] = S.y[i];
            S.st.cycles = (unsigned long long)(clock64() - t0);
            S.st.elapsed_us = (unsigned)((globaltimer_ns() - g0) / 1000);
            A.stats[b] = S.st;
        }
        __syncwarp();
    }
}

// ---- batched problem-function evaluation (eval_psi_grad_psi / eval_psi / eval_hess_psi_prod) ----
struct EvalArgs {
    int N, n, m, nx;
    long long B;
    const double *x0, *u, *y, *Sigma, *v;
    double *psi, *grad, *yhat, *Hv;
    double *ws;
    size_t ws_stride;
};
template <class Model>
__global__ void __launch_bounds__(32) eval_kernel(const __grid_constant__ EvalArgs A) {
    double *ws = A.ws + size_t(blockIdx.x) * A.ws_stride;
    for (long long b = blockIdx.x; b < A.B; b += gridDim.x) {
        Model mdl;
        mdl.init(A.N, A.x0 + b * A.nx, ws);
        const double *u = A.u + b * A.n, *y = A.y ? A.y + b * A.m : nullptr,
                     *S = A.Sigma ? A.Sigma + b * A.m : nullptr;
        double psi = 0;
        if (A.grad) psi = mdl.eval_psi_grad_psi(u, y, S, A.grad + b * A.n);
        if (A.yhat || !A.grad) psi = mdl.eval_psi(u, y, S, A.yhat ? A.yhat + b * A.m : nullptr);
        if (A.psi && lane_id() == 0) A.psi[b] = psi;
        if (A.v && A.Hv) mdl.hess_psi_prod(u, y, S, A.v + b * A.n, A.Hv + b * A.n);
        __syncwarp();
    }
}

// ---- plant step: "simulate_dynamics" (problems/problems/formulation.hpp:258-261) ----
template <class Model>
__global__ void __launch_bounds__(32) simulate_kernel(long long B, const double *x, const double *u,
                                                      long long u_stride, double *xn) {
    for (long long b = blockIdx.x; b < B; b += gridDim.x) {
        Model mdl;
        mdl.init(1, x + b * Model::NX, nullptr);
        mdl.plant_step(x + b * Model::NX, u + b * u_stride, xn + b * Model::NX);
        __syncwarp();
    }
}

// ---- warm-start shift (drivers/mpc-driver.cpp:168-175): forward overlapping copy ----
__global__ void __launch_bounds__(32) shift_kernel(long long B, double *u, int n, int nu, double *y, int m,
                                                   int nc) {
    for (long long b = blockIdx.x; b < B; b += gridDim.x) {
        double *ub = u + b * n;
        for (int i0 = 0; i0 < n - nu; i0 += 32) {
            int i = i0 + lane_id();
            double v = i < n - nu ? ub[i + nu] : 0.0;
            __syncwarp();
            if (i < n - nu) ub[i] = v;
            __syncwarp();
        }
        if (y && m > nc) {
            double *yb = y + b * m;
            for (int i0 = 0; i0 < m - nc; i0 += 32) {
                int i = i0 + lane_id();
                double v = i < m - nc ? yb[i + nc] : 0.0;
                __syncwarp();
                if (i < m - nc) yb[i] = v;
                __syncwarp();
            }
        }
    }
}

// ---- FP64 peak probe: independent DFMA chains (the roofline denominator, measured live) ----
__global__ void __launch_bounds__(256) fp64_peak_kernel(double *out, int iters, double a, double b) {
    double c0 = threadIdx.x, c1 = c0 + 1, c2 = c0 + 2, c3 = c0 + 3, c4 = c0 + 4, c5 = c0 + 5, c6 = c0 + 6,
           c7 = c0 + 7;
    for (int i = 0; i < iters; ++i) {
        c0 = fma(c0, a, b); c1 = fma(c1, a, b); c2 = fma(c2, a, b); c3 = fma(c3, a, b);
        c4 = fma(c4, a, b); c5 = fma(c5, a, b); c6 = fma(c6, a, b); c7 = fma(c7, a, b);
    }
    double s = ((c0 + c1) + (c2 + c3)) + ((c4 + c5) + (c6 + c7));
    if (s == 123.456) out[0] = s; // never true; keeps the chains alive
}

} // namespace

#include "pantr_b200_host.inc"
#include "pantr_b200_problem.inc"
#include "pantr_b200_noalloc.inc"
