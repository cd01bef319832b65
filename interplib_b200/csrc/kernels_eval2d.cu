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
                  const double *__restrict__ D1g, size_t E, const double *__restrict__ dofs, double *__restrict__ out)
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

    unsigned phase[2] = {0, 0};
    size_t t = blockIdx.x;
    if (t < n_tiles && tile_is_bulk(t) && threadIdx.x == 0)
    {
        const unsigned bytes = tile_elems(t) * NN * sizeof(double);
        mbar_expect_tx(&bar[0], bytes);
        bulk_g2s(tile[0], dofs + t * (size_t)EPB * NN, bytes, &bar[0]);
    }
    for (int it = 0; t < n_tiles; ++it, t += gridDim.x)
    {
        const int cur = it & 1;
        const size_t t_next = t + gridDim.x;
        // buffer cur^1 was released by the __syncthreads() closing the previous iteration
        if (t_next < n_tiles && tile_is_bulk(t_next) && threadIdx.x == 0)
        {
            const unsigned bytes = tile_elems(t_next) * NN * sizeof(double);
            mbar_expect_tx(&bar[cur ^ 1], bytes);
            bulk_g2s(tile[cur ^ 1], dofs + t_next * (size_t)EPB * NN, bytes, &bar[cur ^ 1]);
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
    }
}

template <int N, int M, bool GRAD>
int launch_one(const LaunchCfg &cfg, const OpDesc &op, const OpHost &host, size_t E, const double *dofs, double *out)
{
    static_assert(sizeof(Dim0Tabs<N, M>) <= 3072, "dimension-0 tables must fit the kernel parameter space");
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
    eval2d_kernel<N, M, GRAD><<<grid, kWarps * 32, 0, cfg.stream>>>(tab0, op.dim[1].Bin, op.dim[1].Din, E, dofs, out);
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
