// K1 -- sum-factorised batched reconstruction on 2-D tensor-product elements.
//
//   out[e][f][q0][q1] = sum_{i0,i1} T0_f[i0][q0] * T1_f[i1][q1] * dofs[e][i0][i1]
//
// replaces the per-element loops of compute_integration_point_values[_derivatives]
// (/root/reference/src/python/degrees_of_freedom.c:232-303).  One launch produces the value and the two
// reference-space derivatives (fields 0, d/dxi0, d/dxi1) -- BASELINE.json config 2, "basis + gradient eval".
//
// Mapping (HBM-bound kernel, FP64 pipe co-critical: 2*(2 n^2 m + 3 n m^2) flops per 8*(n^2 + 3 m^2) bytes):
//   * a CTA walks tiles of EPB consecutive elements; the tile's DoFs ([EPB][N][N] doubles, contiguous in
//     HBM) are brought into shared memory by ONE bulk async copy (cp.async.bulk = TMA, mbarrier completion),
//     double buffered so the next tile streams in while the current one is contracted;
//   * M lanes own one element, lane q1 owns quadrature column q1: stage 1 contracts dimension 1 with the
//     lane's private columns B1[:,q1], D1[:,q1] (registers, loaded once per kernel), DoFs arrive as
//     shared-memory broadcasts; stage 2 contracts dimension 0 entirely in registers against B0/D0 that sit in
//     the kernel-parameter constant bank (compile-time indices -> constant operands of DFMA, no loads);
//   * each lane streams its results straight from registers (st.global.cs): for fixed (field, q0) the M
//     lanes of an element write one full 8*M-byte run.
// No thread is idle in either stage and the FMA count is exactly the sum-factorised minimum.
#include <algorithm>
#include <memory>
#include <string>

#include "ctx.cuh"

namespace
{

template <int N, int M>
struct Dim0Tabs
{
    double B[N][M];
    double D[N][M];
};

__device__ __forceinline__ unsigned smem_addr(const void *p)
{
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_addr(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_addr(dst)),
                 "l"(src), "r"(bytes), "r"(smem_addr(bar))
                 : "memory");
}

constexpr int kWarps = 8;

template <int N, int M, bool GRAD>
__global__ void __launch_bounds__(kWarps * 32)
    eval2d_kernel(const __grid_constant__ Dim0Tabs<N, M> tab0, const double *__restrict__ B1g,
                  const double *__restrict__ D1g, size_t E, const double *__restrict__ dofs, double *__restrict__ out,
                  unsigned long long *__restrict__ sched)
{
    constexpr int EPW = 32 / M;         // elements per warp
    constexpr int EPB = EPW * kWarps;   // elements per tile
    constexpr int NN = N * N;
    constexpr int Q = M * M;
    constexpr int NF = GRAD ? 3 : 1;

    __shared__ __align__(128) double tile[2][EPB * NN];
    __shared__ __align__(8) unsigned long long bar[2];

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int el = lane / M;  // element within the warp
    const int q1 = lane % M;
    const bool active = el < EPW;
    const int slot = warp * EPW + el; // element within the tile

    // lane-private columns of the dimension-1 tables
    double b1[N], d1[N];
#pragma unroll
    for (int i = 0; i < N; ++i)
    {
        b1[i] = B1g[i * M + q1];
        d1[i] = GRAD ? D1g[i * M + q1] : 0.0;
    }

    const size_t n_tiles = (E + EPB - 1) / EPB;
    const bool base_aligned = (reinterpret_cast<size_t>(dofs) & 15) == 0;
    if (threadIdx.x == 0)
    {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    auto tile_elems = [&](size_t t) -> int { return (int)((t + 1) * EPB <= E ? EPB : E - t * EPB); };
    auto tile_is_bulk = [&](size_t t) -> bool { return base_aligned && ((tile_elems(t) * NN * sizeof(double)) % 16 == 0); };

    // Dynamic tile scheduling: the first tile of a CTA is its block index, every further one is claimed from the
    // global counter sched[0] (SMs do not drain an HBM-bound stream at the same rate; with a static round-robin the
    // fast ones idle at the end).  Thread 0 keeps the claim two tiles ahead so the atomic's latency hides behind a
    // whole tile of work, and hands tile indices to the CTA through next_tile[].
    __shared__ size_t next_tile[2];
    unsigned phase[2] = {0, 0};
    size_t t = blockIdx.x;
    size_t t_next = n_tiles, t_next2 = n_tiles; // meaningful in thread 0 only
    if (threadIdx.x == 0)
    {
        if (t < n_tiles && tile_is_bulk(t))
        {
            const unsigned bytes = tile_elems(t) * NN * sizeof(double);
            mbar_expect_tx(&bar[0], bytes);
            bulk_g2s(tile[0], dofs + t * (size_t)EPB * NN, bytes, &bar[0]);
        }
        t_next = gridDim.x + atomicAdd(sched, 1ull);
        t_next2 = gridDim.x + atomicAdd(sched, 1ull);
    }
    for (int it = 0; t < n_tiles; ++it)
    {
        const int cur = it & 1;
        // buffer cur^1 was released by the __syncthreads() closing the previous iteration
        if (threadIdx.x == 0)
        {
            if (t_next < n_tiles && tile_is_bulk(t_next))
            {
                const unsigned bytes = tile_elems(t_next) * NN * sizeof(double);
                mbar_expect_tx(&bar[cur ^ 1], bytes);
                bulk_g2s(tile[cur ^ 1], dofs + t_next * (size_t)EPB * NN, bytes, &bar[cur ^ 1]);
            }
            next_tile[cur ^ 1] = t_next;
            t_next = t_next2;
            if (t_next2 < n_tiles)
                t_next2 = gridDim.x + atomicAdd(sched, 1ull);
        }
        const int n_here = tile_elems(t);
        if (tile_is_bulk(t))
        {
            mbar_wait(&bar[cur], phase[cur]);
            phase[cur] ^= 1u;
        }
        else
        {
            // ragged or unaligned tail: plain coalesced loads
            const double *src = dofs + t * (size_t)EPB * NN;
            for (int idx = threadIdx.x; idx < n_here * NN; idx += blockDim.x)
                tile[cur][idx] = src[idx];
            __syncthreads();
        }

        if (active && slot < n_here)
        {
            const double *c = tile[cur] + slot * NN;
            double tv[N], td[N];
#pragma unroll
            for (int i0 = 0; i0 < N; ++i0)
            {
                double a = 0.0, b = 0.0;
#pragma unroll
                for (int i1 = 0; i1 < N; ++i1)
                {
                    const double cv = c[i0 * N + i1];
                    a = fma(cv, b1[i1], a);
                    if (GRAD)
                        b = fma(cv, d1[i1], b);
                }
                tv[i0] = a;
                td[i0] = b;
            }
            double *o = out + (t * (size_t)EPB + slot) * (size_t)(NF * Q) + q1;
#pragma unroll
            for (int q0 = 0; q0 < M; ++q0)
            {
                double u = 0.0, g0 = 0.0, g1 = 0.0;
#pragma unroll
                for (int i0 = 0; i0 < N; ++i0)
                {
                    u = fma(tab0.B[i0][q0], tv[i0], u);
                    if (GRAD)
                    {
                        g0 = fma(tab0.D[i0][q0], tv[i0], g0);
                        g1 = fma(tab0.B[i0][q0], td[i0], g1);
                    }
                }
                __stcs(o + q0 * M, u);
                if (GRAD)
                {
                    __stcs(o + Q + q0 * M, g0);
                    __stcs(o + 2 * Q + q0 * M, g1);
                }
            }
        }
        __syncthreads();
        t = next_tile[cur ^ 1];
    }
    sched_finish(sched);
}

// ------------------------------------------------------------------------------------------------------------
// K3 -- batched L2-projection right-hand sides, the transpose of K1:
//   out[e][i0][i1] = sum_{q0} Bw0[i0][q0] * sum_{q1} Bw1[i1][q1] * (f * det)[e][q0][q1],   Bw = B * rule weight
// replaces projection_l2_dual's [Q][N] broadcast + column sums (/root/reference/python/interplib/integration.py:143-166).
// Lane (el, i1) owns output column i1: stage 1 contracts q1 against its private row Bw1[i1][:] (registers) with
// f*det read as shared-memory broadcasts, stage 2 contracts q0 against Bw0 from the constant bank.  f (and the
// geometry / det slab) of a tile arrive by bulk async copy, double buffered; det is either given, computed in
// place from the bilinear corner coordinates (reference operation order, see kernels_mass2d.cu) or absent.
template <int N, int M>
struct Dim0W
{
    double Bw[N][M];
};

__device__ __forceinline__ double bilinear_term(const double t0[2], const double t1[2], const double c[4])
{
    double val = 0.0;
#pragma unroll
    for (int i0 = 0; i0 < 2; ++i0)
#pragma unroll
        for (int i1 = 0; i1 < 2; ++i1)
            val = __dadd_rn(val, __dmul_rn(__dmul_rn(t0[i0], t1[i1]), c[i0 * 2 + i1]));
    return val;
}

template <int N, int M>
__global__ void __launch_bounds__(kWarps * 32)
    l2dual2d_kernel(const __grid_constant__ Dim0W<N, M> tab0, const double *__restrict__ Bw1g, const double *__restrict__ gtab,
                    int wsrc, size_t E, const double *__restrict__ fvals, const double *__restrict__ aux, double *__restrict__ out,
                    unsigned long long *__restrict__ sched)
{
    constexpr int EPW = 32 / N;
    constexpr int EPB = EPW * kWarps;
    constexpr int Q = M * M;

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int el = lane / N;
    const int i1 = lane % N;
    const bool active = el < EPW;
    const int slot = warp * EPW + el;
    const int aux_per = wsrc == 1 ? Q : (wsrc == 2 ? 8 : 0); // det[Q] or the 8 corner coordinates per element

    // dynamic shared memory: f tiles | aux tiles | geometry 1-D tables | mbarriers
    extern __shared__ __align__(128) double dyn_smem[];
    double *ftile[2] = {dyn_smem, dyn_smem + EPB * Q};
    double *atile[2] = {dyn_smem + 2 * EPB * Q, dyn_smem + 2 * EPB * Q + EPB * aux_per};
    double *gt = dyn_smem + 2 * EPB * Q + 2 * EPB * aux_per;
    unsigned long long *bar = reinterpret_cast<unsigned long long *>(gt + 8 * M);

    double bw1[M];
#pragma unroll
    for (int q = 0; q < M; ++q)
        bw1[q] = Bw1g[i1 * M + q];
    if (wsrc == 2)
        for (int i = threadIdx.x; i < 8 * M; i += blockDim.x)
            gt[i] = gtab[i];

    const size_t n_tiles = (E + EPB - 1) / EPB;
    const bool base_aligned = ((reinterpret_cast<size_t>(fvals) | reinterpret_cast<size_t>(aux)) & 15) == 0;
    if (threadIdx.x == 0)
    {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    auto tile_elems = [&](size_t t) -> int { return (int)((t + 1) * EPB <= E ? EPB : E - t * EPB); };
    auto tile_is_bulk = [&](size_t t) -> bool {
        const int n = tile_elems(t);
        return base_aligned && ((n * Q * sizeof(double)) % 16 == 0) && ((n * aux_per * sizeof(double)) % 16 == 0);
    };
    auto issue = [&](size_t t, int b) {
        const int n = tile_elems(t);
        const unsigned fb = n * Q * sizeof(double), ab = n * aux_per * sizeof(double);
        mbar_expect_tx(&bar[b], fb + ab);
        bulk_g2s(ftile[b], fvals + t * (size_t)EPB * Q, fb, &bar[b]);
        if (ab)
            bulk_g2s(atile[b], aux + t * (size_t)EPB * aux_per, ab, &bar[b]);
    };

    // dynamic tile scheduling, as in eval2d_kernel
    __shared__ size_t next_tile[2];
    unsigned phase[2] = {0, 0};
    size_t t = blockIdx.x;
    size_t t_next = n_tiles, t_next2 = n_tiles; // meaningful in thread 0 only
    if (threadIdx.x == 0)
    {
        if (t < n_tiles && tile_is_bulk(t))
            issue(t, 0);
        t_next = gridDim.x + atomicAdd(sched, 1ull);
        t_next2 = gridDim.x + atomicAdd(sched, 1ull);
    }
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
            mbar_wait(&bar[cur], phase[cur]);
            phase[cur] ^= 1u;
        }
        else
        {
            const double *src = fvals + t * (size_t)EPB * Q;
            for (int idx = threadIdx.x; idx < n_here * Q; idx += blockDim.x)
                ftile[cur][idx] = src[idx];
            const double *asrc = aux + t * (size_t)EPB * aux_per;
            for (int idx = threadIdx.x; idx < n_here * aux_per; idx += blockDim.x)
                atile[cur][idx] = asrc[idx];
            __syncthreads();
        }

        const bool work = active && slot < n_here;
        double *g = ftile[cur] + (work ? slot : 0) * Q;
        if (work && wsrc != 0)
        {
            // g = f * det, each lane of the element takes every N-th point
            if (wsrc == 1)
            {
                const double *dt = atile[cur] + slot * Q;
                for (int q = i1; q < Q; q += N)
                    g[q] *= dt[q];
            }
            else
            {
                const double *gc = atile[cur] + slot * 8;
                double cx[4], cy[4];
#pragma unroll
                for (int i = 0; i < 4; ++i)
                {
                    cx[i] = gc[i];
                    cy[i] = gc[4 + i];
                }
                const double *gB0 = gt, *gD0 = gt + 2 * M, *gB1 = gt + 4 * M, *gD1 = gt + 6 * M;
                for (int q = i1; q < Q; q += N)
                {
                    const int q0 = q / M, q1 = q % M;
                    const double b0[2] = {gB0[q0], gB0[M + q0]}, d0[2] = {gD0[q0], gD0[M + q0]};
                    const double b1[2] = {gB1[q1], gB1[M + q1]}, d1[2] = {gD1[q1], gD1[M + q1]};
                    const double J00 = bilinear_term(d0, b1, cx), J01 = bilinear_term(b0, d1, cx);
                    const double J10 = bilinear_term(d0, b1, cy), J11 = bilinear_term(b0, d1, cy);
                    const double p = J01 * J10;
                    const double err = fma(-J01, J10, p);
                    g[q] *= fma(J00, J11, -p) + err;
                }
            }
        }
        __syncwarp();
        if (work)
        {
            double sv[M];
#pragma unroll
            for (int q0 = 0; q0 < M; ++q0)
            {
                double a = 0.0;
#pragma unroll
                for (int q1 = 0; q1 < M; ++q1)
                    a = fma(g[q0 * M + q1], bw1[q1], a);
                sv[q0] = a;
            }
            double *o = out + (t * (size_t)EPB + slot) * (size_t)(N * N) + i1;
#pragma unroll
            for (int i0 = 0; i0 < N; ++i0)
            {
                double b = 0.0;
#pragma unroll
                for (int q0 = 0; q0 < M; ++q0)
                    b = fma(tab0.Bw[i0][q0], sv[q0], b);
                __stcs(o + i0 * N, b);
            }
        }
        __syncthreads();
        t = next_tile[cur ^ 1];
    }
    sched_finish(sched);
}

template <int N, int M>
int launch_l2_one(const LaunchCfg &cfg, const double *d_tabs, const OpHost &host, int wsrc, size_t E, const double *fvals,
                  const double *aux, double *out)
{
    Dim0W<N, M> tab0;
    for (int i = 0; i < N; ++i)
        for (int q = 0; q < M; ++q)
            tab0.Bw[i][q] = host.Bin[0][i * M + q] * host.w[0][q];
    constexpr int EPB = (32 / N) * kWarps;
    const size_t n_tiles = (E + EPB - 1) / EPB;
    const int aux_per = wsrc == 1 ? M * M : (wsrc == 2 ? 8 : 0);
    const size_t smem_bytes = sizeof(double) * (2 * EPB * M * M + 2 * EPB * aux_per + 8 * M) + 16;
    if (smem_bytes > 160 * 1024)
        return IB200_ERROR_UNSUPPORTED;
    IB_CUDA(cudaFuncSetAttribute(l2dual2d_kernel<N, M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    int per_sm = 0;
    IB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, l2dual2d_kernel<N, M>, kWarps * 32, smem_bytes));
    if (per_sm < 1)
        return IB200_ERROR_UNSUPPORTED;
    const size_t cap = (size_t)cfg.sm_count * per_sm;
    const int grid = (int)std::max<size_t>(1, n_tiles < cap ? n_tiles : cap);
    l2dual2d_kernel<N, M><<<grid, kWarps * 32, smem_bytes, cfg.stream>>>(tab0, d_tabs, d_tabs + N * M, wsrc, E, fvals, aux, out, cfg.sched);
    ++*cfg.launch_counter;
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess)
        return ib200_fail(IB200_ERROR_CUDA, "launch of l2dual2d_kernel<%d,%d> failed: %s", N, M, cudaGetErrorString(err));
    return INTERP_SUCCESS;
}

struct L2Plan
{
    double *d_tabs = nullptr; // Bw1[N][M] then gtab[8*M]
    ~L2Plan()
    {
        if (d_tabs)
            cudaFree(d_tabs);
    }
};

template <int N, int M, bool GRAD>
int launch_one(const LaunchCfg &cfg, const OpDesc &op, const OpHost &host, size_t E, const double *dofs, double *out)
{
    static_assert(sizeof(Dim0Tabs<N, M>) <= 8192, "dimension-0 tables must fit the kernel parameter space");
    Dim0Tabs<N, M> tab0;
    std::memcpy(tab0.B, host.Bin[0], sizeof(double) * N * M);
    std::memcpy(tab0.D, host.Din[0], sizeof(double) * N * M);
    constexpr int EPB = (32 / M) * kWarps;
    const size_t n_tiles = (E + EPB - 1) / EPB;
    int per_sm = 0;
    IB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, eval2d_kernel<N, M, GRAD>, kWarps * 32, 0));
    if (per_sm < 1)
        per_sm = 1;
    const size_t cap = (size_t)cfg.sm_count * per_sm;
    const int grid = (int)(n_tiles < cap ? n_tiles : cap);
    eval2d_kernel<N, M, GRAD><<<grid, kWarps * 32, 0, cfg.stream>>>(tab0, op.dim[1].Bin, op.dim[1].Din, E, dofs, out, cfg.sched);
    ++*cfg.launch_counter;
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess)
        return ib200_fail(IB200_ERROR_CUDA, "launch of eval2d_kernel<%d,%d> failed: %s", N, M, cudaGetErrorString(err));
    return INTERP_SUCCESS;
}

template <bool GRAD>
int dispatch(const LaunchCfg &cfg, const OpDesc &op, const OpHost &host, size_t E, const double *dofs, double *out)
{
    const int n = op.dim[0].n_in, m = op.dim[0].m;
#define IB_CASE(NV, MV) \
    if (n == NV && m == MV) \
        return launch_one<NV, MV, GRAD>(cfg, op, host, E, dofs, out);
    // (basis count, quadrature points) per dimension: p = 1..8 with p+1 .. p+3 points, plus the
    // over-integrated low orders the reference's Quad geometry uses
    IB_CASE(2, 2) IB_CASE(2, 3) IB_CASE(2, 4) IB_CASE(2, 6) IB_CASE(2, 8) IB_CASE(2, 10)
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

} // namespace

int launch_eval2d_fast(const LaunchCfg &cfg, const OpDesc &op, const OpHost &host, size_t E, const double *dofs,
                       unsigned n_fields, const unsigned *masks, double *out)
{
    if (op.ndim != 2 || op.dim[0].n_in != op.dim[1].n_in || op.dim[0].m != op.dim[1].m)
        return IB200_ERROR_UNSUPPORTED;
    if (n_fields == 1 && masks[0] == 0u)
        return dispatch<false>(cfg, op, host, E, dofs, out);
    if (n_fields == 3 && masks[0] == 0u && masks[1] == 1u && masks[2] == 2u)
        return dispatch<true>(cfg, op, host, E, dofs, out);
    return IB200_ERROR_UNSUPPORTED;
}

// wsrc: 0 = reference-domain integration, 1 = det[E][Q] given, 2 = bilinear geometry [E][2][2][2] fused.
int launch_l2dual2d_fast(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, const OpDesc &op, const OpHost &host,
                         int wsrc, const OpHost *geom_host, size_t E, const double *fvals, const double *aux, double *out)
{
    if (op.ndim != 2 || op.dim[0].n_in != op.dim[1].n_in || op.dim[0].m != op.dim[1].m)
        return IB200_ERROR_UNSUPPORTED;
    const int n = op.dim[0].n_in, m = op.dim[0].m;
    if (n > 32)
        return IB200_ERROR_UNSUPPORTED;
    auto it = ctx->plans.find(key);
    if (it == ctx->plans.end())
    {
        auto plan = std::make_shared<L2Plan>();
        std::vector<double> tabs((size_t)n * m + 8 * m, 0.0);
        for (int i = 0; i < n; ++i)
            for (int q = 0; q < m; ++q)
                tabs[(size_t)i * m + q] = host.Bin[1][i * m + q] * host.w[1][q];
        if (wsrc == 2)
        {
            if (!geom_host)
                return IB200_ERROR_UNSUPPORTED;
            double *gt = tabs.data() + (size_t)n * m;
            std::memcpy(gt + 0 * m, geom_host->Bin[0], sizeof(double) * 2 * m);
            std::memcpy(gt + 2 * m, geom_host->Din[0], sizeof(double) * 2 * m);
            std::memcpy(gt + 4 * m, geom_host->Bin[1], sizeof(double) * 2 * m);
            std::memcpy(gt + 6 * m, geom_host->Din[1], sizeof(double) * 2 * m);
        }
        IB_CUDA(cudaMalloc((void **)&plan->d_tabs, tabs.size() * sizeof(double)));
        IB_CUDA(cudaMemcpy(plan->d_tabs, tabs.data(), tabs.size() * sizeof(double), cudaMemcpyHostToDevice));
        it = ctx->plans.emplace(key, std::static_pointer_cast<void>(plan)).first;
    }
    const L2Plan &plan = *static_cast<const L2Plan *>(it->second.get());
#define IB_CASE(NV, MV) \
    if (n == NV && m == MV) \
        return launch_l2_one<NV, MV>(cfg, plan.d_tabs, host, wsrc, E, fvals, aux, out);
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
