// N1-fd -- the whole primal L2 projection of bilinear quads WITHOUT a mass matrix: fast diagonalisation.
//
//   u_e = M_e^-1 b_e,   b_e[i] = sum_q f_e[q] W[q] det J_e(q) phi_i(q)       (projection_l2_primal,
//                                                                             /root/reference/python/interplib/integration.py:171-223)
//
// On a bilinear quad det J_e(xi0, xi1) = alpha + beta xi0 + gamma xi1 exactly (the xi0 xi1 terms of the 2x2 determinant
// cancel), hence M_e = alpha A0(x)A1 + beta X0(x)A1 + gamma A0(x)X1 with element-independent 1-D matrices, and with the
// generalised eigenvectors V_k of the pencils (X_k, A_k) (host: fastdiag.c)
//
//   u_e = (V0 (x) V1) diag(1 / (alpha + beta lambda0_i0 + gamma lambda1_i1)) (V0 (x) V1)^T b_e.
//
// The reference assembles M_e (O(N^2 Q) per entry) and calls LAPACK's LU per element; the factorisation kernels of
// kernels_generic.cu (spd_solve_*) do the same on the device at 0.06 .. 0.29 of the HBM bound.  Here the projection
// is four sum-factorised contractions and N reciprocals per element -- 552 bytes and ~1.3 kflop at BASELINE config 5
// instead of 5.5 KB of scratch traffic and ~17 kflop -- and the result agrees with solve(M, b) to cond(M) * eps.
//
// Mapping (the transpose of K1 followed by K1, see kernels_eval2d.cu): N lanes own one element, lane i1 owns column i1
// of the transformed right-hand side.
//   stage 1  S[q0] = sum_q1 f[q0][q1] (a_q0 + gamma xi1[q1]) Vw1[i1][q1],  a_q0 = alpha + beta xi0[q0]
//            (two lane-private tables Vw1 and xi1*Vw1 in registers, f as shared-memory broadcasts from the TMA tile)
//   stage 2  bt[i0] = sum_q0 Vw0[i0][q0] S[q0]                      (Vw0 in the kernel-parameter constant bank)
//   scale    ut[i0] = bt[i0] / (alpha + beta lambda0[i0] + gamma lambda1[i1])
//   stage 3  t[j0]  = sum_i0 V0[j0][i0] ut[i0]                      (constant bank), written to per-warp shared memory
//   stage 4  u[j0][j1] = sum_i1 V1[j1][i1] t[j0][i1]                (lane j1, V1 row in registers, t as broadcasts)
// f (288 B) and the corner coordinates (64 B) of a tile arrive by cp.async.bulk with mbarrier completion, double
// buffered; tiles are claimed dynamically (ctx.cuh).
#include <algorithm>
#include <memory>
#include <string>
#include <vector>

#include "ctx.cuh"

namespace
{

__device__ __forceinline__ unsigned fd_smem_addr(const void *p)
{
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void fd_mbar_init(unsigned long long *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(fd_smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void fd_mbar_expect_tx(unsigned long long *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fd_smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void fd_mbar_wait(unsigned long long *bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "FD_WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra FD_DONE;\n"
        "bra FD_WAIT_LOOP;\n"
        "FD_DONE:\n"
        "}\n" ::"r"(fd_smem_addr(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void fd_bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     fd_smem_addr(dst)),
                 "l"(src), "r"(bytes), "r"(fd_smem_addr(bar))
                 : "memory");
}

constexpr int kFdWarps = 8;

template <int N, int M>
struct FdTabs
{
    double Vw0[N][M]; // transformed, weighted test functions of dimension 0
    double V0[N][N];  // V0[j0][i0]
    double lam0[N];
    double xi0[M];
    double G[3][4];   // (x_xi0, x_xi1, x_xi0xi1) = G (c00, c01, c10, c11): coefficients of the bilinear coordinate map
    double Gs[3];     // row sums of G (zero for Lagrange / Bernstein geometry DoFs)
};

// 1 / d to an ulp or two: hardware seed (about 20 bits) and two Newton steps.  __drcp_rn costs about twice the FP64
// instructions for its exact rounding and special cases; d is a Jacobian determinant here, far from the subnormal range.
__device__ __forceinline__ double fd_rcp(double d)
{
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
    double e = fma(-d, r, 1.0);
    r = fma(r, e, r);
    e = fma(-d, r, 1.0);
    return fma(r, e, r);
}

// difference of products a*b - c*d with the rounding error of the second product recovered (Kahan)
__device__ __forceinline__ double fd_diff_of_products(double a, double b, double c, double d)
{
    const double p = c * d;
    const double err = fma(-c, d, p);
    return fma(a, b, -p) + err;
}

template <int N, int M>
__global__ void __launch_bounds__(kFdWarps * 32)
    l2proj_fd2d_kernel(const __grid_constant__ FdTabs<N, M> tab, const double *__restrict__ dim1, size_t E,
                       const double *__restrict__ fvals, const double *__restrict__ geom, double *__restrict__ out,
                       unsigned long long *__restrict__ sched, int *__restrict__ status)
{
    constexpr int EPW = 32 / N;
    constexpr int EPB = EPW * kFdWarps;
    constexpr int Q = M * M;
    constexpr int NP = N + (N & 1); // even row pitch of the stage-3 tile: 16-byte aligned rows
    // element pitch of the f tile: with an even Q the rows are read as 16-byte words, and a pitch whose half is odd puts the
    // words of the warp's elements into different bank groups (Q = 36: 288 B apart, elements 0/4 and 1/5 collided; ncu:
    // L1/shared pipe 82 % busy) -- the tile then arrives as one bulk copy per element instead of one per tile
    // (measured dead end: one 288-byte bulk copy per element instead of one per tile made C5 27 % slower)
    constexpr int QP = Q;
    constexpr bool PER_ELEMENT = false;

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int el = lane / N;
    const int i1 = lane % N;
    const bool active = el < EPW;
    const int slot = warp * EPW + el;

    // dynamic shared memory: f tiles | corner tiles | stage-3 tiles (per warp) | mbarriers
    extern __shared__ __align__(128) double dyn_smem[];
    auto ftile = [&](int b) -> double * { return dyn_smem + b * (EPB * QP); };
    auto gtile = [&](int b) -> double * { return dyn_smem + 2 * EPB * QP + b * (EPB * 8); };
    double *tsm = dyn_smem + 2 * EPB * QP + 2 * EPB * 8 + (size_t)(warp * EPW + (active ? el : 0)) * (N * NP);
    unsigned long long *bar = reinterpret_cast<unsigned long long *>(dyn_smem + 2 * EPB * QP + 2 * EPB * 8 + (size_t)EPB * N * NP);

    // lane-private tables of dimension 1: Vw1[i1][:], xi1[:] * Vw1[i1][:], V1[i1][:], lambda1[i1]
    // dim1 layout: Vw1[N][M] | V1[N][N] | lam1[N] | xi1[M]
    double vw1[M], xvw1[M], v1[N];
    {
        const double *Vw1 = dim1, *V1 = dim1 + N * M, *xi1 = dim1 + N * M + N * N + N;
#pragma unroll
        for (int q = 0; q < M; ++q)
        {
            vw1[q] = Vw1[i1 * M + q];
            xvw1[q] = __dmul_rn(xi1[q], vw1[q]);
        }
#pragma unroll
        for (int i = 0; i < N; ++i)
            v1[i] = V1[i1 * N + i];
    }
    const double lam1 = dim1[N * M + N * N + i1];

    const size_t n_tiles = (E + EPB - 1) / EPB;
    const bool base_aligned = ((reinterpret_cast<size_t>(fvals) | reinterpret_cast<size_t>(geom)) & 15) == 0;
    if (threadIdx.x == 0)
    {
        fd_mbar_init(&bar[0], 1);
        fd_mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    auto tile_elems = [&](size_t t) -> int { return (int)((t + 1) * EPB <= E ? EPB : E - t * EPB); };
    auto tile_is_bulk = [&](size_t t) -> bool {
        const int n = tile_elems(t);
        return base_aligned && ((n * Q * sizeof(double)) % 16 == 0);
    };
    auto issue = [&](size_t t, int b) {
        const int n = tile_elems(t);
        const unsigned fb = n * Q * sizeof(double), gb = n * 8 * sizeof(double);
        fd_mbar_expect_tx(&bar[b], fb + gb);
        if (PER_ELEMENT)
        {
            for (int k = 0; k < n; ++k)
                fd_bulk_g2s(ftile(b) + k * QP, fvals + (t * (size_t)EPB + k) * Q, Q * sizeof(double), &bar[b]);
        }
        else
            fd_bulk_g2s(ftile(b), fvals + t * (size_t)EPB * Q, fb, &bar[b]);
        fd_bulk_g2s(gtile(b), geom + t * (size_t)EPB * 8, gb, &bar[b]);
    };

    __shared__ size_t next_tile[2];
    unsigned phase = 0; // bit b: parity the next wait on bar[b] expects
    size_t t = blockIdx.x;
    size_t t_next = n_tiles, t_next2 = n_tiles; // meaningful in thread 0 only
    if (threadIdx.x == 0)
    {
        if (t < n_tiles && tile_is_bulk(t))
            issue(t, 0);
        t_next = gridDim.x + atomicAdd(sched, 1ull);
        t_next2 = gridDim.x + atomicAdd(sched, 1ull);
    }
    bool bad = false;
    for (int it = 0; t < n_tiles; ++it)
    {
        const int cur = it & 1;
        if (threadIdx.x == 0)
        {
            if (t_next < n_tiles && tile_is_bulk(t_next))
                issue(t_next, cur ^ 1);
            next_tile[cur ^ 1] = t_next;
            t_next = t_next2;
            if (t_next2 < n_tiles)
                t_next2 = gridDim.x + atomicAdd(sched, 1ull);
        }
        const int n_here = tile_elems(t);
        if (tile_is_bulk(t))
        {
            fd_mbar_wait(&bar[cur], (phase >> cur) & 1u);
            phase ^= 1u << cur;
        }
        else
        {
            const double *src = fvals + t * (size_t)EPB * Q;
            for (int idx = threadIdx.x; idx < n_here * Q; idx += blockDim.x)
                ftile(cur)[(idx / Q) * QP + idx % Q] = src[idx];
            const double *gsrc = geom + t * (size_t)EPB * 8;
            for (int idx = threadIdx.x; idx < n_here * 8; idx += blockDim.x)
                gtile(cur)[idx] = gsrc[idx];
            __syncthreads();
        }

        const bool work = active && slot < n_here;
        double tj[N];
        if (work)
        {
            // det J = alpha + beta xi0 + gamma xi1 from the corner coordinates [c][i0][i1]
            double gc[8]; // four 16-byte loads (the corner block of an element is 64-byte aligned)
            {
                const double2 *g2 = reinterpret_cast<const double2 *>(gtile(cur) + slot * 8);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                {
                    const double2 v = g2[k];
                    gc[2 * k] = v.x;
                    gc[2 * k + 1] = v.y;
                }
            }
            double dx[3], dy[3];
#pragma unroll
            for (int k = 0; k < 3; ++k)
            {
                double ax = __dmul_rn(tab.Gs[k], gc[0]), ay = __dmul_rn(tab.Gs[k], gc[4]);
#pragma unroll
                for (int c = 1; c < 4; ++c) // G c = G (c - c0) + rowsum * c0: exact differences, eps * h instead of eps * |x|
                {
                    ax = fma(tab.G[k][c], gc[c] - gc[0], ax);
                    ay = fma(tab.G[k][c], gc[4 + c] - gc[4], ay);
                }
                dx[k] = ax;
                dy[k] = ay;
            }
            const double alpha = fd_diff_of_products(dx[0], dy[1], dx[1], dy[0]);
            const double beta = fd_diff_of_products(dx[0], dy[2], dx[2], dy[0]);
            const double gamma = fd_diff_of_products(dx[2], dy[1], dx[1], dy[2]);

            // stages 1 and 2
            const double *f = ftile(cur) + slot * QP;
            double sv[M];
#pragma unroll
            for (int q0 = 0; q0 < M; ++q0)
            {
                double s1 = 0.0, s2 = 0.0;
#pragma unroll
                for (int q1 = 0; q1 < M; ++q1)
                {
                    const double fv = f[q0 * M + q1];
                    s1 = fma(fv, vw1[q1], s1);
                    s2 = fma(fv, xvw1[q1], s2);
                }
                sv[q0] = fma(fma(beta, tab.xi0[q0], alpha), s1, __dmul_rn(gamma, s2));
            }
            double ut[N];
            const double d1 = fma(gamma, lam1, alpha);
#pragma unroll
            for (int i0 = 0; i0 < N; ++i0)
            {
                double b = 0.0;
#pragma unroll
                for (int q0 = 0; q0 < M; ++q0)
                    b = fma(tab.Vw0[i0][q0], sv[q0], b);
                const double d = fma(beta, tab.lam0[i0], d1);
                bad = bad || !(fabs(d) > 0.0);
                ut[i0] = __dmul_rn(b, fd_rcp(d));
            }
            // stage 3
#pragma unroll
            for (int j0 = 0; j0 < N; ++j0)
            {
                double a = 0.0;
#pragma unroll
                for (int i0 = 0; i0 < N; ++i0)
                    a = fma(tab.V0[j0][i0], ut[i0], a);
                tj[j0] = a;
            }
        }
        __syncwarp(); // stage-4 readers of the previous tile are done with the warp's stage-3 tile
        if (work)
        {
#pragma unroll
            for (int j0 = 0; j0 < N; ++j0)
                tsm[j0 * NP + i1] = tj[j0];
        }
        __syncwarp();
        if (work)
        {
            // stage 4: lane j1 = i1
            double *o = out + (t * (size_t)EPB + slot) * (size_t)(N * N) + i1;
#pragma unroll
            for (int j0 = 0; j0 < N; ++j0)
            {
                double a = 0.0;
#pragma unroll
                for (int k = 0; k < N; ++k)
                    a = fma(v1[k], tsm[j0 * NP + k], a);
                __stcs(o + j0 * N, a);
            }
        }
        __syncthreads();
        t = next_tile[cur ^ 1];
    }
    if (bad)
        atomicOr(status, IB200_STATUS_BAD_SYSTEM);
    sched_finish(sched);
}

struct FdPlan
{
    bool usable = false;
    int n = 0, m = 0;
    std::vector<double> Vw0, V0, lam0, xi0;
    double G[3][4] = {};
    double *d_dim1 = nullptr; // Vw1[N][M] | V1[N][N] | lam1[N] | xi1[M]
    ~FdPlan()
    {
        if (d_dim1)
            cudaFree(d_dim1);
    }
};

template <int N, int M>
int launch_fd_one(const LaunchCfg &cfg, const FdPlan &plan, size_t E, const double *fvals, const double *geom, double *out)
{
    static_assert(sizeof(FdTabs<N, M>) <= 16 * 1024, "parameter block too large");
    std::unique_ptr<FdTabs<N, M>> holder(new FdTabs<N, M>);
    FdTabs<N, M> &tab = *holder;
    std::memcpy(tab.Vw0, plan.Vw0.data(), sizeof tab.Vw0);
    std::memcpy(tab.V0, plan.V0.data(), sizeof tab.V0);
    std::memcpy(tab.lam0, plan.lam0.data(), sizeof tab.lam0);
    std::memcpy(tab.xi0, plan.xi0.data(), sizeof tab.xi0);
    std::memcpy(tab.G, plan.G, sizeof tab.G);
    for (int k = 0; k < 3; ++k)
        tab.Gs[k] = plan.G[k][0] + plan.G[k][1] + plan.G[k][2] + plan.G[k][3];
    constexpr int EPW = 32 / N, EPB = EPW * kFdWarps, NP = N + (N & 1);
    const size_t n_tiles = (E + EPB - 1) / EPB;
    constexpr int Q = M * M, QP = Q;
    const size_t smem_bytes = sizeof(double) * (2 * EPB * QP + 2 * EPB * 8 + EPB * N * NP) + 16;
    if (smem_bytes > 200 * 1024)
        return IB200_ERROR_UNSUPPORTED;
    auto kernel = l2proj_fd2d_kernel<N, M>;
    IB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    int per_sm = 0;
    IB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kFdWarps * 32, smem_bytes));
    if (per_sm < 1)
        return IB200_ERROR_UNSUPPORTED;
    const size_t cap = (size_t)cfg.sm_count * per_sm;
    const int grid = (int)std::max<size_t>(1, n_tiles < cap ? n_tiles : cap);
    kernel<<<grid, kFdWarps * 32, smem_bytes, cfg.stream>>>(tab, plan.d_dim1, E, fvals, geom, out, cfg.sched, cfg.status);
    ++*cfg.launch_counter;
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess)
        return ib200_fail(IB200_ERROR_CUDA, "launch of l2proj_fd2d_kernel<%d,%d> failed: %s", N, M, cudaGetErrorString(err));
    return INTERP_SUCCESS;
}

} // namespace

// Primal L2 projection of E bilinear quads by fast diagonalisation: geom[E][2][2][2] corner coordinates (DoFs of the
// order-1 geometry basis), fvals[E][Q] samples of f at the mapped rule nodes -> out[E][N] primal DoFs.
// IB200_ERROR_UNSUPPORTED (no error text) outside the envelope: square 2-D spaces in the instantiation list with a rule
// that makes the 1-D mass matrices positive definite.  E == 0 with out == NULL only answers that question.
int launch_l2proj_fd2d(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, const basis_spec_t *basis,
                       const integration_spec_t *rules, const basis_spec_t *geom_basis, size_t E, const double *fvals,
                       const double *geom, double *out)
{
    if (basis[0].order != basis[1].order || rules[0].order != rules[1].order)
        return IB200_ERROR_UNSUPPORTED;
    const int n = (int)basis[0].order + 1, m = (int)rules[0].order + 1;
    if (n > 32 || m < n)
        return IB200_ERROR_UNSUPPORTED;
    auto it = ctx->plans.find(key);
    if (it == ctx->plans.end())
    {
        auto plan = std::make_shared<FdPlan>();
        plan->n = n;
        plan->m = m;
        plan->Vw0.resize((size_t)n * m);
        plan->V0.resize((size_t)n * n);
        plan->lam0.resize(n);
        plan->xi0.resize(m);
        std::vector<double> dim1((size_t)n * m + (size_t)n * n + n + m), wts(m);
        int res = ib200_tables_fastdiag(basis[0], rules[0], plan->V0.data(), plan->lam0.data(), plan->Vw0.data());
        if (res == INTERP_SUCCESS)
            res = ib200_tables_fastdiag(basis[1], rules[1], dim1.data() + (size_t)n * m, dim1.data() + (size_t)n * m + (size_t)n * n,
                                        dim1.data());
        if (res == INTERP_SUCCESS)
            res = ib200_tables_rule(rules[0], plan->xi0.data(), wts.data(), nullptr);
        if (res == INTERP_SUCCESS)
            res = ib200_tables_rule(rules[1], dim1.data() + (size_t)n * m + (size_t)n * n + n, wts.data(), nullptr);
        double a0[2], s0[2], a1[2], s1[2];
        if (res == INTERP_SUCCESS)
            res = ib200_tables_linear_basis(geom_basis[0], a0, s0);
        if (res == INTERP_SUCCESS)
            res = ib200_tables_linear_basis(geom_basis[1], a1, s1);
        if (res == INTERP_SUCCESS)
        {
            for (int i0 = 0; i0 < 2; ++i0)
                for (int k = 0; k < 2; ++k)
                {
                    plan->G[0][i0 * 2 + k] = s0[i0] * a1[k]; // d/dxi0 at the origin
                    plan->G[1][i0 * 2 + k] = a0[i0] * s1[k]; // d/dxi1 at the origin
                    plan->G[2][i0 * 2 + k] = s0[i0] * s1[k]; // mixed term
                }
            if (cudaMalloc((void **)&plan->d_dim1, dim1.size() * sizeof(double)) != cudaSuccess ||
                cudaMemcpy(plan->d_dim1, dim1.data(), dim1.size() * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess)
                return ib200_fail(IB200_ERROR_CUDA, "upload of the fast-diagonalisation tables failed");
            plan->usable = true;
        }
        else if (res != INTERP_ERROR_BAD_SYSTEM && res != IB200_ERROR_UNSUPPORTED)
            return res;
        it = ctx->plans.emplace(key, std::static_pointer_cast<void>(plan)).first;
    }
    const FdPlan &plan = *static_cast<const FdPlan *>(it->second.get());
    if (!plan.usable)
        return IB200_ERROR_UNSUPPORTED;
    const bool probe = E == 0 && out == nullptr; // "is this operation inside the envelope?" -- no launch
#define IB_CASE(NV, MV) \
    if (n == NV && m == MV) \
        return probe ? (int)INTERP_SUCCESS : launch_fd_one<NV, MV>(cfg, plan, E, fvals, geom, out);
    IB_CASE(2, 2) IB_CASE(2, 3) IB_CASE(2, 4)
    IB_CASE(3, 3) IB_CASE(3, 4) IB_CASE(3, 5)
    IB_CASE(4, 4) IB_CASE(4, 5) IB_CASE(4, 6)
    IB_CASE(5, 5) IB_CASE(5, 6) IB_CASE(5, 7)
    IB_CASE(6, 6) IB_CASE(6, 7) IB_CASE(6, 8)
    IB_CASE(7, 7) IB_CASE(7, 8) IB_CASE(7, 9)
    IB_CASE(8, 8) IB_CASE(8, 9) IB_CASE(8, 10)
    IB_CASE(9, 9) IB_CASE(9, 10) IB_CASE(9, 11)
    IB_CASE(13, 14) IB_CASE(17, 18)
#undef IB_CASE
    return IB200_ERROR_UNSUPPORTED;
}

// ------------------------------------------------------------------------------------------------------------------------
// 1-D: straight segments (BASELINE config 1: mesh1d, p = 4, 8-point Gauss rule, mass matrix + L2 projection of sin(x)).
// The Jacobian determinant of an order-1 coordinate map is one number per element, so everything the reference computes per
// element (python/interplib/integration.py:106-223 over src/python/mass_matrices.c:264-335 and mappings.c:304-556) is a
// table times that number:  M_e = det_e A,  b_e = det_e Bw f_e,  u_e = M_e^-1 b_e = P f_e  (P = A^-1 Bw, host, long double).
// ONE launch instead of the three (system) or four (projection) generic ones: thread (e, j) writes row j of M_e and entry j
// of b_e / u_e, consecutive threads write consecutive rows.  The configuration is launch-latency bound (1 024 elements).
namespace
{

// tabs: A[n][n] | Bw[n][m] | P[n][m]
__global__ void __launch_bounds__(256)
    l2_system_1d_kernel(int n, int m, double s0, double s1, const double *__restrict__ tabs, size_t E, const double *__restrict__ geom,
                        const double *__restrict__ f, double *__restrict__ M, double *__restrict__ b, double *__restrict__ u,
                        int *__restrict__ status)
{
    const double *A = tabs, *Bw = tabs + n * n, *P = Bw + n * m;
    const size_t total = E * (size_t)n;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x)
    {
        const size_t e = t / n;
        const int j = (int)(t - e * n);
        // dx/dxi = (0 + D0 c0) + D1 c1 with separate roundings, as compute_integration_point_values_derivatives does
        const double det = __dadd_rn(__dmul_rn(s0, geom[2 * e]), __dmul_rn(s1, geom[2 * e + 1]));
        if (M)
        {
            double *row = M + t * n;
            for (int i = 0; i < n; ++i)
                __stcs(row + i, __dmul_rn(det, A[j * n + i]));
        }
        if (b || u)
        {
            const double *fe = f + e * m;
            double sb = 0.0, su = 0.0;
            for (int q = 0; q < m; ++q)
            {
                const double fv = fe[q];
                sb = fma(Bw[j * m + q], fv, sb);
                su = fma(P[j * m + q], fv, su);
            }
            if (b)
                b[t] = __dmul_rn(det, sb);
            if (u)
            {
                u[t] = su;
                if (!(fabs(det) > 0.0))
                    atomicOr(status, IB200_STATUS_BAD_SYSTEM);
            }
        }
    }
}

struct Seg1dPlan
{
    bool usable = false;
    int n = 0, m = 0;
    double s[2] = {0.0, 0.0};
    double *d_tabs = nullptr;
    ~Seg1dPlan()
    {
        if (d_tabs)
            cudaFree(d_tabs);
    }
};

} // namespace

// geom[E][1][2] (DoFs of the order-1 geometry basis), fvals[E][m] -> any of mass[E][n][n], rhs[E][n], primal[E][n] (NULL: not
// wanted).  IB200_ERROR_UNSUPPORTED (no error text) when the rule under-integrates the space; E == 0 with all outputs NULL
// only answers whether the operation is inside the envelope.
int launch_l2_system_1d(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, basis_spec_t basis, integration_spec_t rule,
                        basis_spec_t geom_basis, size_t E, const double *geom, const double *fvals, double *mass, double *rhs,
                        double *primal)
{
    if (geom_basis.order != 1)
        return IB200_ERROR_UNSUPPORTED;
    auto it = ctx->plans.find(key);
    if (it == ctx->plans.end())
    {
        auto plan = std::make_shared<Seg1dPlan>();
        const int n = (int)basis.order + 1, m = (int)rule.order + 1;
        plan->n = n;
        plan->m = m;
        std::vector<double> tabs((size_t)n * n + 2 * (size_t)n * m), X((size_t)n * n), B((size_t)n * m), D((size_t)n * m), x(m), w(m);
        double a[2];
        int res = ib200_tables_pair_moments(basis, basis, rule, tabs.data(), X.data());
        if (res == INTERP_SUCCESS)
            res = ib200_tables_rule(rule, x.data(), w.data(), nullptr);
        if (res == INTERP_SUCCESS)
            res = ib200_tables_basis(basis, rule, B.data(), D.data());
        if (res == INTERP_SUCCESS)
            res = ib200_tables_projector_1d(basis, rule, tabs.data() + (size_t)n * n + (size_t)n * m);
        if (res == INTERP_SUCCESS)
            res = ib200_tables_linear_basis(geom_basis, a, plan->s);
        if (res == INTERP_SUCCESS)
        {
            for (int i = 0; i < n; ++i)
                for (int q = 0; q < m; ++q)
                    tabs[(size_t)n * n + (size_t)i * m + q] = B[(size_t)i * m + q] * w[q];
            if (cudaMalloc((void **)&plan->d_tabs, tabs.size() * sizeof(double)) != cudaSuccess ||
                cudaMemcpy(plan->d_tabs, tabs.data(), tabs.size() * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess)
                return ib200_fail(IB200_ERROR_CUDA, "upload of the 1-D projection tables failed");
            plan->usable = true;
        }
        else if (res != INTERP_ERROR_BAD_SYSTEM && res != IB200_ERROR_UNSUPPORTED)
            return res;
        it = ctx->plans.emplace(key, std::static_pointer_cast<void>(plan)).first;
    }
    const Seg1dPlan &plan = *static_cast<const Seg1dPlan *>(it->second.get());
    if (!plan.usable)
        return IB200_ERROR_UNSUPPORTED;
    if (E == 0)
        return INTERP_SUCCESS;
    const size_t total = E * (size_t)plan.n;
    const int grid = (int)std::max<size_t>(1, std::min<size_t>((total + 255) / 256, (size_t)cfg.sm_count * 8));
    l2_system_1d_kernel<<<grid, 256, 0, cfg.stream>>>(plan.n, plan.m, plan.s[0], plan.s[1], plan.d_tabs, E, geom, fvals, mass, rhs, primal,
                                                      cfg.status);
    ++*cfg.launch_counter;
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess)
        return ib200_fail(IB200_ERROR_CUDA, "launch of l2_system_1d_kernel failed: %s", cudaGetErrorString(err));
    return INTERP_SUCCESS;
}
