// K2-lo -- element mass matrices of LOW-ORDER 2-D tensor-product elements, fully specialised at compile time
// (SURVEY.md 8a rows A7 + A9; BASELINE config 5: p = 4, 6x6 Gauss points, 16M quads).
//
//   out[e][(j0,j1)][(i0,i1)] = sum_{q0} P0[(j0,i0)][q0] * T_e[q0][(j1,i1)]
//   T_e[q0][(j1,i1)]         = sum_{q1} Wt_e[q0][q1] * P1[(j1,i1)][q1]
//
// Same mathematics and tables as kernels_mass2d.cu (reference loops /root/reference/src/python/mass_matrices.c:264-335,
// geometry /root/reference/src/python/mappings.c:448-542) for the case both spaces match (symmetric matrix, no
// derivative).  Why a second kernel: at p = 4 the roofline leaves 228 SM cycles per element (5 064 B at 6.45 TB/s over
// 148 SMs); the run-time-shaped kernel spends 765 warp instructions per element (issue slots 68 % busy, ncu
// profiles/r01_c5_mass2d_full.txt) on index arithmetic, constant loads through computed offsets and predicated
// stores.  Here basis count N, point count M and items per thread R are template parameters:
//   * stage 2 is fully unrolled over (j0,i0): the coefficients P0[(j0,i0)][q0] are compile-time offsets into the
//     kernel-parameter constant bank (direct DFMA operands, no load instructions), every staging store is one STS with
//     an immediate offset;
//   * a WARP is the unit of work (EPW elements per warp, TPE threads per element, thread (el,bg) owns R column pairs
//     (j1 >= i1) and writes the mirror entry too -> exactly symmetric result); there is no CTA-wide barrier after
//     set-up, warps never wait for each other;
//   * bilinear-quad geometry in the reference's operation order with the element-independent products
//     dphi0*phi1 / phi0*dphi1 tabulated once per CTA in shared memory;
//   * row panels (PJ values of j0 = one contiguous run of global memory per element) are staged in per-warp shared
//     memory and leave the SM as cp.async.bulk shared->global (TMA) copies, NBUF-deep.
#include <algorithm>
#include <memory>
#include <string>

#include "ctx.cuh"

namespace
{

__device__ __forceinline__ unsigned lo_smem_u32(const void *p)
{
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void lo_bulk_s2g(double *gdst, const double *ssrc, unsigned bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(lo_smem_u32(ssrc)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void lo_bulk_commit()
{
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int K>
__device__ __forceinline__ void lo_bulk_wait_read()
{
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(K) : "memory");
}
__device__ __forceinline__ void lo_bulk_wait_all()
{
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ void lo_fence_async_smem()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

template <int N, int M, int R>
struct LoShape
{
    static constexpr int NB = N * (N + 1) / 2;       // column pairs (j1 >= i1)
    static constexpr int TPE = (NB + R - 1) / R;     // threads per element
    static constexpr int EPW = 32 / TPE;             // elements per warp
    static constexpr int Q = M * M;
    static constexpr int Qp = (Q + 1) & ~1;
    static constexpr int ROWB = N * N * N;           // doubles per j0 (rows (j0, 0..N-1))
    static constexpr int NN = ROWB * N;              // doubles per element
    static constexpr int PJ_RAW = 8192 / (8 * ROWB * EPW);
    static constexpr int PJ = PJ_RAW < 1 ? 1 : (PJ_RAW > N ? N : PJ_RAW);
    static constexpr int NPANELS = (N + PJ - 1) / PJ;
    static constexpr int CAP = ((PJ * ROWB + 2 + 1) & ~1) + 2; // staged doubles per element and panel (+ parity slack, bank skew)
    static constexpr int NBUF = 2;
    static constexpr int WARPS = 4;
    static constexpr int WARP_DOUBLES = EPW * Qp + NBUF * EPW * CAP;
    static constexpr size_t SMEM_BYTES = sizeof(double) * (8 * Q + (size_t)WARPS * WARP_DOUBLES);
};

template <int N, int M>
struct LoParams
{
    double P0[N * (N + 1) / 2][M]; // pairs hi >= lo at hi*(hi+1)/2 + lo
    double gB0[2][M], gD0[2][M], gB1[2][M], gD1[2][M];
    int wsrc;                      // 0: unit weights, 1: weight[E][Q], 2: bilinear geometry [E][2][2][2] -> det
};

template <int N, int M, int R>
__global__ void __launch_bounds__(LoShape<N, M, R>::WARPS * 32)
    mass2d_lo_kernel(const __grid_constant__ LoParams<N, M> prm, const double *__restrict__ P1g, size_t E,
                     const double *__restrict__ wsrc, double *__restrict__ out)
{
    using S = LoShape<N, M, R>;
    extern __shared__ __align__(16) double smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int el = lane / S::TPE, bg = lane % S::TPE;
    const bool lane_active = el < S::EPW;
    const bool leader = lane_active && bg == 0;

    // CTA-shared: element-independent geometry products per point, pt[q][0..3] = dphi0[i0]*phi1[i1] (d/dxi0),
    // pt[q][4..7] = phi0[i0]*dphi1[i1] (d/dxi1), rounded like the reference's basis product
    double *pt = smem;
    double *Wsm = smem + 8 * S::Q + (size_t)warp * S::WARP_DOUBLES;
    double *stage = Wsm + S::EPW * S::Qp;
    if (prm.wsrc == 2)
    {
        for (int i = threadIdx.x; i < 8 * S::Q; i += blockDim.x)
        {
            const int q = i >> 3, k = i & 7, q0 = q / M, q1 = q % M, i0 = (k >> 1) & 1, i1 = k & 1;
            pt[i] = (k < 4) ? __dmul_rn(prm.gD0[i0][q0], prm.gB1[i1][q1]) : __dmul_rn(prm.gB0[i0][q0], prm.gD1[i1][q1]);
        }
    }
    __syncthreads();

    // the thread's column pairs b = bg + r*TPE -> (j1, i1), staging offsets of the entry and of its mirror
    int off_d[R], off_m[R];
    double p1[R][M];
#pragma unroll
    for (int r = 0; r < R; ++r)
    {
        const int b = bg + r * S::TPE;
        const bool ok = lane_active && b < S::NB;
        int j1 = 0;
        while ((j1 + 1) * (j1 + 2) / 2 <= (ok ? b : 0))
            ++j1;
        const int i1 = (ok ? b : 0) - j1 * (j1 + 1) / 2;
        off_d[r] = ok ? j1 * N * N + i1 : -1;
        off_m[r] = (ok && i1 != j1) ? i1 * N * N + j1 : -1;
#pragma unroll
        for (int q = 0; q < M; ++q)
            p1[r][q] = ok ? P1g[(size_t)b * M + q] : 0.0;
    }

    const size_t n_tiles = (E + S::EPW - 1) / S::EPW;
    const size_t n_warps = (size_t)gridDim.x * S::WARPS;
    unsigned pc = 0; // running panel counter of this warp

    for (size_t tile = (size_t)blockIdx.x * S::WARPS + warp; tile < n_tiles; tile += n_warps)
    {
        const size_t e = tile * S::EPW + el;
        const bool el_valid = lane_active && e < E;

        // ---- per-point weights of the tile's elements -> shared memory (per warp)
        __syncwarp(); // stage-1 readers of the previous tile are done
        if (el_valid)
        {
            double *W = Wsm + el * S::Qp;
            if (prm.wsrc == 2)
            {
                const double *g = wsrc + e * 8; // [c][i0][i1]
                double cx[4], cy[4];
#pragma unroll
                for (int i = 0; i < 4; ++i)
                {
                    cx[i] = g[i];
                    cy[i] = g[4 + i];
                }
                for (int q = bg; q < S::Q; q += S::TPE)
                {
                    const double2 *pq = reinterpret_cast<const double2 *>(pt + 8 * q);
                    const double2 a01 = pq[0], a23 = pq[1], b01 = pq[2], b23 = pq[3];
                    // val = ((0 + p0*c0) + p1*c1) + ... with separate roundings (degrees_of_freedom.c:264-303)
                    const double J00 = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(a01.x, cx[0]), __dmul_rn(a01.y, cx[1])), __dmul_rn(a23.x, cx[2])), __dmul_rn(a23.y, cx[3]));
                    const double J01 = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(b01.x, cx[0]), __dmul_rn(b01.y, cx[1])), __dmul_rn(b23.x, cx[2])), __dmul_rn(b23.y, cx[3]));
                    const double J10 = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(a01.x, cy[0]), __dmul_rn(a01.y, cy[1])), __dmul_rn(a23.x, cy[2])), __dmul_rn(a23.y, cy[3]));
                    const double J11 = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(b01.x, cy[0]), __dmul_rn(b01.y, cy[1])), __dmul_rn(b23.x, cy[2])), __dmul_rn(b23.y, cy[3]));
                    // difference of products with the rounding error of the second product recovered (Kahan)
                    const double p = J01 * J10;
                    const double err = fma(-J01, J10, p);
                    W[q] = fma(J00, J11, -p) + err;
                }
            }
            else if (prm.wsrc == 1)
            {
                const double *src = wsrc + e * S::Q;
                for (int q = bg; q < S::Q; q += S::TPE)
                    W[q] = src[q];
            }
            else
            {
                for (int q = bg; q < S::Q; q += S::TPE)
                    W[q] = 1.0;
            }
        }
        __syncwarp();

        // ---- stage 1: T[r][q0] = sum_q1 W[q0][q1] * P1[b_r][q1]
        double T[R][M];
        {
            const double *W = Wsm + (lane_active ? el : 0) * S::Qp;
#pragma unroll
            for (int q0 = 0; q0 < M; ++q0)
            {
                double w[M];
#pragma unroll
                for (int q1 = 0; q1 < M; ++q1)
                    w[q1] = W[q0 * M + q1];
#pragma unroll
                for (int r = 0; r < R; ++r)
                {
                    double acc = 0.0;
#pragma unroll
                    for (int q1 = 0; q1 < M; ++q1)
                        acc = fma(w[q1], p1[r][q1], acc);
                    T[r][q0] = acc;
                }
            }
        }

        // ---- stage 2, one row panel at a time (fully unrolled: all offsets and coefficients are immediates)
#pragma unroll
        for (int pj = 0; pj < S::NPANELS; ++pj)
        {
            constexpr int NBUF = S::NBUF;
            const unsigned pcc = pc + (unsigned)pj;
            const int buf = (int)(pcc % NBUF);
            if (pcc >= (unsigned)NBUF)
            {
                if (leader)
                    lo_bulk_wait_read<NBUF - 1>();
                __syncwarp();
            }
            const int j0s = pj * S::PJ;
            const int rows = (S::PJ < N - j0s) ? S::PJ : N - j0s;
            const size_t gstart = e * (size_t)S::NN + (size_t)j0s * S::ROWB;
            const int par = (int)(gstart & 1);
            double *dst = stage + ((size_t)buf * S::EPW + (lane_active ? el : 0)) * S::CAP + par;
#pragma unroll
            for (int jj = 0; jj < S::PJ; ++jj)
            {
                if (jj < rows)
                {
                    const int j0 = j0s + jj;
#pragma unroll
                    for (int i0 = 0; i0 < N; ++i0)
                    {
                        const int hi = j0 > i0 ? j0 : i0, lo = j0 > i0 ? i0 : j0;
                        const int a = hi * (hi + 1) / 2 + lo;
                        double v[R];
#pragma unroll
                        for (int r = 0; r < R; ++r)
                            v[r] = 0.0;
#pragma unroll
                        for (int q0 = 0; q0 < M; ++q0)
#pragma unroll
                            for (int r = 0; r < R; ++r)
                                v[r] = fma(prm.P0[a][q0], T[r][q0], v[r]);
                        const int base = jj * S::ROWB + i0 * N;
#pragma unroll
                        for (int r = 0; r < R; ++r)
                        {
                            if (off_d[r] >= 0)
                                dst[base + off_d[r]] = v[r];
                            if (off_m[r] >= 0)
                                dst[base + off_m[r]] = v[r];
                        }
                    }
                }
            }
            lo_fence_async_smem();
            __syncwarp();
            if (leader)
            {
                if (el_valid)
                {
                    const int cnt = rows * S::ROWB;
                    double *g = out + gstart;
                    const double *sp = dst; // sp[k] <-> g[k]
                    int k0 = 0;
                    if (par)
                    {
                        g[0] = sp[0];
                        k0 = 1;
                    }
                    const int body = (cnt - k0) & ~1;
                    if (body > 0)
                        lo_bulk_s2g(g + k0, sp + k0, (unsigned)body * 8u);
                    if (k0 + body < cnt)
                        g[cnt - 1] = sp[cnt - 1];
                }
                lo_bulk_commit();
            }
        }
        pc += S::NPANELS;
    }
    if (leader)
        lo_bulk_wait_all();
}

struct LoPlan
{
    double *d_P1 = nullptr; // [NB][M], pairs j1 >= i1 at j1*(j1+1)/2 + i1
    ~LoPlan()
    {
        if (d_P1)
            cudaFree(d_P1);
    }
};

template <int N, int M, int R>
int launch_lo(const LaunchCfg &cfg, const LoPlan &plan, const OpHost &host, int wsrc_kind, const OpHost *geom_host, size_t E,
              const double *wsrc, double *out)
{
    using S = LoShape<N, M, R>;
    static_assert(sizeof(LoParams<N, M>) <= 16 * 1024, "parameter block too large");
    LoParams<N, M> prm;
    std::memset(&prm, 0, sizeof prm);
    // the product of the two basis values first -> exactly symmetric in (j0,i0), same table as kernels_mass2d.cu
    for (int hi = 0; hi < N; ++hi)
        for (int lo = 0; lo <= hi; ++lo)
            for (int q = 0; q < M; ++q)
                prm.P0[hi * (hi + 1) / 2 + lo][q] = host.w[0][q] * (host.Bout[0][hi * M + q] * host.Bin[0][lo * M + q]);
    prm.wsrc = wsrc_kind;
    if (wsrc_kind == 2)
    {
        std::memcpy(prm.gB0, geom_host->Bin[0], sizeof(double) * 2 * M);
        std::memcpy(prm.gD0, geom_host->Din[0], sizeof(double) * 2 * M);
        std::memcpy(prm.gB1, geom_host->Bin[1], sizeof(double) * 2 * M);
        std::memcpy(prm.gD1, geom_host->Din[1], sizeof(double) * 2 * M);
    }
    auto kernel = mass2d_lo_kernel<N, M, R>;
    IB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::SMEM_BYTES));
    int per_sm = 0;
    IB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, S::WARPS * 32, S::SMEM_BYTES));
    if (per_sm < 1)
        return IB200_ERROR_UNSUPPORTED;
    const size_t n_tiles = (E + S::EPW - 1) / S::EPW;
    const size_t want = (n_tiles + S::WARPS - 1) / S::WARPS;
    const size_t cap = (size_t)cfg.sm_count * per_sm;
    const int grid = (int)std::max<size_t>(1, std::min(want, cap));
    kernel<<<grid, S::WARPS * 32, S::SMEM_BYTES, cfg.stream>>>(prm, plan.d_P1, E, wsrc, out);
    ++*cfg.launch_counter;
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess)
        return ib200_fail(IB200_ERROR_CUDA, "launch of mass2d_lo_kernel<%d,%d,%d> failed: %s", N, M, R, cudaGetErrorString(err));
    return INTERP_SUCCESS;
}

} // namespace

// Low-order symmetric path: both dimensions with matching spaces of the same basis count, no derivative.  Same contract as
// launch_mass2d_fast; IB200_ERROR_UNSUPPORTED (no error text) for every other shape.
int launch_mass2d_lo(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, const OpDesc &op, const OpHost &host,
                     bool match0, bool match1, int deriv_dim_in, int wsrc_kind, const OpHost *geom_host, int widx_in, size_t E,
                     const double *wsrc, double *out)
{
    if (op.ndim != 2 || !match0 || !match1 || deriv_dim_in >= 0 || widx_in >= 0)
        return IB200_ERROR_UNSUPPORTED;
    if (op.dim[0].m != op.dim[1].m || op.dim[0].n_in != op.dim[1].n_in)
        return IB200_ERROR_UNSUPPORTED;
    if ((reinterpret_cast<size_t>(out) & 15) != 0 || (wsrc_kind == 2 && !geom_host))
        return IB200_ERROR_UNSUPPORTED;
    const int n = op.dim[0].n_in, m = op.dim[0].m;
    if (n < 2 || n > 6 || m > 9)
        return IB200_ERROR_UNSUPPORTED;
    const std::string k = "lo:" + key;
    auto it = ctx->plans.find(k);
    if (it == ctx->plans.end())
    {
        auto plan = std::make_shared<LoPlan>();
        std::vector<double> P1;
        for (int j1 = 0; j1 < n; ++j1)
            for (int i1 = 0; i1 <= j1; ++i1)
                for (int q = 0; q < m; ++q)
                    P1.push_back(host.w[1][q] * (host.Bout[1][j1 * m + q] * host.Bin[1][i1 * m + q]));
        IB_CUDA(cudaMalloc((void **)&plan->d_P1, P1.size() * sizeof(double)));
        IB_CUDA(cudaMemcpy(plan->d_P1, P1.data(), P1.size() * sizeof(double), cudaMemcpyHostToDevice));
        it = ctx->plans.emplace(k, std::static_pointer_cast<void>(plan)).first;
    }
    const LoPlan &plan = *static_cast<const LoPlan *>(it->second.get());
#define IB_LO(NV, MV, RV) \
    if (n == NV && m == MV) \
        return launch_lo<NV, MV, RV>(cfg, plan, host, wsrc_kind, geom_host, E, wsrc, out);
    // (basis count, points, column pairs per thread)
    IB_LO(2, 2, 3) IB_LO(2, 3, 3) IB_LO(2, 4, 3)
    IB_LO(3, 3, 3) IB_LO(3, 4, 3) IB_LO(3, 5, 3)
    IB_LO(4, 4, 3) IB_LO(4, 5, 3) IB_LO(4, 6, 3)
    IB_LO(5, 5, 2) IB_LO(5, 6, 2) IB_LO(5, 7, 2)
    IB_LO(6, 6, 3) IB_LO(6, 7, 3) IB_LO(6, 8, 3)
#undef IB_LO
    return IB200_ERROR_UNSUPPORTED;
}
