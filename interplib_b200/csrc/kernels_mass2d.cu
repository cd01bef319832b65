// K2 -- sum-factorised batched element mass / gradient-mass matrices on 2-D tensor-product elements
// (geometry row A9 fused for bilinear quads).
//
//   out[e][(j0,j1)][(i0,i1)] = sum_{q0} P0[(j0,i0)][q0] * T_e[q0][(j1,i1)]
//   T_e[q0][(j1,i1)]         = sum_{q1} Wt_e[q0][q1] * P1[(j1,i1)][q1]
//   P_k[(j,i)][q]            = w_k[q] * Bout_k[j][q] * Tin_k[i][q]          (Tin = D in the derivative dimension)
//   Wt_e[q]                  = 1 | weight[e][q] | det J_e(q) | cofactor of J_e(q)
//
// replaces the O(N^2 Q) per-entry quadrature loops of compute_mass_matrix / compute_gradient_mass_matrix
// (/root/reference/src/python/mass_matrices.c:264-335, :486-560) and, when fused, the CoordinateMap/SpaceMap
// construction (/root/reference/src/python/mappings.c:8-90, 304-556) for bilinear quads.
//
// The kernel is write-bound (N_out*N_in*8 bytes out per element against 64 bytes in), FP64 co-critical at high
// order.  Structure:
//   * a CTA hosts several independent "engines" (1..3 warps, private named barrier); an engine walks tiles of EPE
//     consecutive elements.  Thread (el, bg) owns R column items b = (j1,i1) of element el -- when the dim-1
//     spaces match only the pairs i1 <= j1 (T is symmetric in (j1,i1); the mirror entry is the same number);
//   * stage 1: T[r][q0] in registers (Wt from shared memory as broadcast reads, P1 rows from L1);
//   * stage 2: for every (j0,i0) the M coefficients P0[(j0,i0)][:] are UNIFORM and live in the kernel-parameter
//     constant bank (LDCU -> uniform registers -> DFMA operand): R*M DFMAs per M uniform loads, no LSU traffic;
//   * results are staged in shared memory one row panel (PJ values of j0 = a contiguous run of global memory) at a
//     time and leave the SM as ONE bulk async copy per panel (cp.async.bulk shared->global = TMA store), double
//     buffered so the next panel is computed while the previous one drains.  Odd-sized elements make panel starts
//     only 8-byte aligned: the panel is staged with the same 16-byte phase as its global address and the (at most
//     one) head and tail doubles are stored with plain st.global.
// Exact symmetry of the result for matching spaces needs no mirroring pass: M[J][I] and M[I][J] are produced by
// the same DFMA sequence on bitwise-identical operands.
#include <algorithm>
#include <cstdlib>
#include <memory>
#include <string>

#include "ctx.cuh"

namespace
{

struct MassShape
{
    int n_out0, n_in0, n_out1, n_in1;
    int N_in;      // row length of the output matrix
    int NN;        // N_out * N_in, doubles per element
    int NB;        // number of column items b (pairs when symb)
    int BGp;       // threads per element
    int EPE;       // elements per engine tile
    int G;         // warps per engine
    int PJ;        // j0 values per panel
    int n_panels;
    int panel_cap; // doubles per staged panel incl. alignment slack (even)
    int nbuf;      // 1 or 2 staging buffers
    int syma;      // P0 stored for pairs i0 <= j0 only
    int wsrc;      // 0: unit weights, 1: weight[E][Q] from global memory, 2: fused bilinear geometry
    int widx_in, widx_out; // fused geometry: -1 -> det, else det * Jinv[widx_in][widx_out] (= a cofactor of J)
    int engines_per_cta;
    int Qp;        // padded points per element in shared memory (even)
};

template <int CAP>
struct MassParams
{
    MassShape s;
    double P0[CAP];
};

__device__ __forceinline__ unsigned smem_u32(const void *p)
{
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void engine_barrier(int id, int threads)
{
    if (threads == 32)
        __syncwarp();
    else
        asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void bulk_s2g(double *gdst, const double *ssrc, unsigned bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit()
{
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_read()
{
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void bulk_wait_all()
{
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ void fence_async_smem()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// Jacobian of a bilinear quad at one point in the reference's operation order (i0 outer, i1 inner,
// phi = T0*T1 rounded, val = val + phi*c with separate roundings; degrees_of_freedom.c:264-303): the entries
// are bit-identical to CoordinateMap.gradient().
__device__ __forceinline__ double bilinear_ref_order(const double t0[2], const double t1[2], const double c[4])
{
    double val = 0.0;
#pragma unroll
    for (int i0 = 0; i0 < 2; ++i0)
#pragma unroll
        for (int i1 = 0; i1 < 2; ++i1)
            val = __dadd_rn(val, __dmul_rn(__dmul_rn(t0[i0], t1[i1]), c[i0 * 2 + i1]));
    return val;
}

// Resident CTAs the register allocation is sized for.  The kernel is latency-bound at 8 warps per SM (ncu: issue slots
// 41 % busy, stall reason "wait"), so up to 10 points per dimension the registers are capped at 128 for two 256-thread
// CTAs per SM (a few hundred bytes of spills at R = 3, M = 10; net +25..40 % on the C3 blocks).
constexpr int min_blocks_for(int M, int R, bool /*symb*/)
{
    return (M <= 6 && R == 2) ? 3 : (M <= 10 ? 2 : 1);
}

template <int M, int R, bool SYMB, int CAP>
__global__ void __launch_bounds__(256, min_blocks_for(M, R, SYMB))
    mass2d_kernel(const __grid_constant__ MassParams<CAP> prm, const double *__restrict__ P1g, const int2 *__restrict__ bmap,
                  const double *__restrict__ gtab, size_t E, const double *__restrict__ wsrc, double *__restrict__ out,
                  unsigned long long *__restrict__ sched)
{
    extern __shared__ __align__(16) double smem[];
    const MassShape &s = prm.s;
    constexpr int Q = M * M;
    const int engine_threads = 32 * s.G;
    const int engine = threadIdx.x / engine_threads;
    const int t = threadIdx.x % engine_threads;
    const int el = t / s.BGp;
    const int bg = t % s.BGp;
    const bool lane_active = el < s.EPE;
    const bool issuer = lane_active && bg == 0;

    // shared memory: [geometry 1-D tables: 8*M doubles][per engine: Wsm EPE*Qp | stage nbuf*EPE*panel_cap]
    double *gt = smem;
    const int engine_doubles = s.EPE * s.Qp + s.nbuf * s.EPE * s.panel_cap;
    double *Wsm = smem + 8 * M + (size_t)engine * engine_doubles;
    double *stage = Wsm + s.EPE * s.Qp;
    if (s.wsrc == 2)
    {
        for (int i = threadIdx.x; i < 8 * M; i += blockDim.x)
            gt[i] = gtab[i];
    }
    __syncthreads();
    const double *gB0 = gt, *gD0 = gt + 2 * M, *gB1 = gt + 4 * M, *gD1 = gt + 6 * M;

    // the thread's column items
    int off_d[R], off_m[R];
    bool valid[R];
#pragma unroll
    for (int r = 0; r < R; ++r)
    {
        const int b = bg * R + r;
        valid[r] = lane_active && b < s.NB;
        const int2 ji = valid[r] ? bmap[b] : make_int2(0, 0);
        off_d[r] = ji.x * s.N_in + ji.y;
        off_m[r] = (SYMB && ji.x != ji.y) ? ji.y * s.N_in + ji.x : -1;
    }

    const size_t n_tiles = (E + s.EPE - 1) / s.EPE;
    const size_t n_engines = (size_t)gridDim.x * s.engines_per_cta;
    const int bar_id = 1 + engine;
    unsigned panel_counter = 0;

    // Dynamic scheduling per engine: the first tile is the engine's global index, every further one is claimed from
    // sched[0] by the engine's first thread one tile ahead and handed over through eng_tile[engine][parity] across the
    // barrier that opens a tile (written at the end of tile `it` into slot (it+1)&1, read after the next barrier).
    __shared__ size_t eng_tile[8][2];
    unsigned long long pend = 0;
    if (t == 0)
    {
        eng_tile[engine][0] = (size_t)blockIdx.x * s.engines_per_cta + engine;
        pend = atomicAdd(sched, 1ull);
    }
    for (int it = 0;; ++it)
    {
        engine_barrier(bar_id, engine_threads); // stage-1 readers of the previous tile are done, eng_tile is visible
        const size_t tile = eng_tile[engine][it & 1];
        if (tile >= n_tiles)
            break;
        if (t == 0)
        {
            const size_t nt = n_engines + pend;
            eng_tile[engine][(it + 1) & 1] = nt;
            if (nt < n_tiles)
                pend = atomicAdd(sched, 1ull);
        }
        const size_t e = tile * s.EPE + el;
        const bool el_valid = lane_active && e < E;

        // ---- per-point weights of the tile's elements -> shared memory
        if (el_valid)
        {
            double *W = Wsm + el * s.Qp;
            if (s.wsrc == 1)
            {
                const double *src = wsrc + e * Q;
                for (int q = bg; q < Q; q += s.BGp)
                    W[q] = src[q];
            }
            else if (s.wsrc == 2)
            {
                const double *g = wsrc + e * 8; // [c][i0][i1]
                double cx[4], cy[4];
#pragma unroll
                for (int i = 0; i < 4; ++i)
                {
                    cx[i] = g[i];
                    cy[i] = g[4 + i];
                }
                for (int q = bg; q < Q; q += s.BGp)
                {
                    const int q0 = q / M, q1 = q % M;
                    const double b0[2] = {gB0[q0], gB0[M + q0]}, d0[2] = {gD0[q0], gD0[M + q0]};
                    const double b1[2] = {gB1[q1], gB1[M + q1]}, d1[2] = {gD1[q1], gD1[M + q1]};
                    const double J00 = bilinear_ref_order(d0, b1, cx); // dx/dxi0
                    const double J01 = bilinear_ref_order(b0, d1, cx); // dx/dxi1
                    const double J10 = bilinear_ref_order(d0, b1, cy);
                    const double J11 = bilinear_ref_order(b0, d1, cy);
                    double wv;
                    if (s.widx_in < 0)
                    {
                        // difference of products with the rounding error of the second product recovered (Kahan)
                        const double p = J01 * J10;
                        const double err = fma(-J01, J10, p);
                        wv = fma(J00, J11, -p) + err;
                    }
                    else
                    {
                        // det * Jinv[a][b] = cofactor: Jinv = [[J11, -J01], [-J10, J00]] / det
                        wv = s.widx_in == 0 ? (s.widx_out == 0 ? J11 : -J01) : (s.widx_out == 0 ? -J10 : J00);
                    }
                    W[q] = wv;
                }
            }
            else
            {
                for (int q = bg; q < Q; q += s.BGp)
                    W[q] = 1.0;
            }
        }
        engine_barrier(bar_id, engine_threads);

        // ---- stage 1: T[r][q0] = sum_q1 W[q0][q1] * P1[b_r][q1]
        double T[R][M];
        {
            const double *W = Wsm + (lane_active ? el : 0) * s.Qp;
#pragma unroll
            for (int r = 0; r < R; ++r)
            {
                double p1[M];
                const double *src = P1g + (size_t)(valid[r] ? bg * R + r : 0) * M;
#pragma unroll
                for (int q1 = 0; q1 < M; ++q1)
                    p1[q1] = src[q1];
#pragma unroll
                for (int q0 = 0; q0 < M; ++q0)
                {
                    double acc = 0.0;
#pragma unroll
                    for (int q1 = 0; q1 < M; ++q1)
                        acc = fma(W[q0 * M + q1], p1[q1], acc);
                    T[r][q0] = acc;
                }
            }
        }

        // ---- stage 2, one row panel at a time
        for (int pj = 0; pj < s.n_panels; ++pj, ++panel_counter)
        {
            const int buf = (s.nbuf == 2) ? (panel_counter & 1) : 0;
            if (panel_counter >= (unsigned)s.nbuf)
            {
                if (issuer)
                {
                    if (s.nbuf == 2)
                        bulk_wait_read<1>();
                    else
                        bulk_wait_read<0>();
                }
                engine_barrier(bar_id, engine_threads);
            }
            const int j0s = pj * s.PJ;
            const int rows = min(s.PJ, s.n_out0 - j0s);
            const int row_block = s.n_out1 * s.N_in; // doubles per j0
            const size_t gstart = e * (size_t)s.NN + (size_t)j0s * row_block;
            const int par = (int)(gstart & 1);
            double *dst = stage + ((size_t)buf * s.EPE + (lane_active ? el : 0)) * s.panel_cap + par;

            for (int jj = 0; jj < rows; ++jj)
            {
                const int j0 = j0s + jj;
#pragma unroll 2
                for (int i0 = 0; i0 < s.n_in0; ++i0)
                {
                    int a;
                    if (s.syma)
                    {
                        const int hi = max(j0, i0), lo = min(j0, i0);
                        a = hi * (hi + 1) / 2 + lo;
                    }
                    else
                    {
                        a = j0 * s.n_in0 + i0;
                    }
                    const double *c = prm.P0 + a * M;
                    double v[R];
#pragma unroll
                    for (int r = 0; r < R; ++r)
                        v[r] = 0.0;
#pragma unroll
                    for (int q0 = 0; q0 < M; ++q0)
                    {
                        const double cq = c[q0];
#pragma unroll
                        for (int r = 0; r < R; ++r)
                            v[r] = fma(cq, T[r][q0], v[r]);
                    }
                    const int base = jj * row_block + i0 * s.n_in1;
#pragma unroll
                    for (int r = 0; r < R; ++r)
                    {
                        if (valid[r] && el_valid)
                        {
                            dst[base + off_d[r]] = v[r];
                            if (SYMB && off_m[r] >= 0)
                                dst[base + off_m[r]] = v[r];
                        }
                    }
                }
            }
            fence_async_smem();
            engine_barrier(bar_id, engine_threads);
            if (issuer)
            {
                if (el_valid)
                {
                    const int cnt = rows * row_block;
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
                        bulk_s2g(g + k0, sp + k0, (unsigned)body * 8u);
                    if (k0 + body < cnt)
                        g[cnt - 1] = sp[cnt - 1];
                }
                bulk_commit();
            }
        }
    }
    if (issuer)
        bulk_wait_all();
    if (t == 0)
        sched_finish_one(sched, n_engines);
}

constexpr int cap_for(int M)
{
    return M <= 6 ? 512 : (M <= 10 ? 1536 : 3840);
}

// Per-operation tables and thread mapping, built once per (spaces, rule, derivative, weight source) and cached
// in the context next to the 1-D table cache.
struct MassPlan
{
    MassShape s{};
    int M = 0, R = 0;
    bool symb = false;
    std::vector<double> P0;
    char *d_tab = nullptr; // P1 | bmap | gtab
    const double *d_P1 = nullptr;
    const int2 *d_bmap = nullptr;
    const double *d_gtab = nullptr;
    ~MassPlan()
    {
        if (d_tab)
            cudaFree(d_tab);
    }
};

template <int M, int R, bool SYMB>
int launch_mass(const LaunchCfg &cfg, const MassPlan &plan, size_t E, const double *wsrc, double *out)
{
    constexpr int CAP = cap_for(M);
    const MassShape &shape = plan.s;
    if ((int)plan.P0.size() > CAP)
        return IB200_ERROR_UNSUPPORTED;
    // kernel parameters are copied at launch; up to 30 KB, so keep them off the stack
    std::unique_ptr<MassParams<CAP>> prm_holder(new MassParams<CAP>);
    MassParams<CAP> &prm = *prm_holder;
    prm.s = shape;
    std::memset(prm.P0, 0, sizeof prm.P0);
    std::memcpy(prm.P0, plan.P0.data(), plan.P0.size() * sizeof(double));
    const int threads = shape.engines_per_cta * 32 * shape.G;
    const size_t smem_bytes =
        sizeof(double) * (8 * M + (size_t)shape.engines_per_cta * (shape.EPE * shape.Qp + shape.nbuf * shape.EPE * shape.panel_cap));
    auto kernel = mass2d_kernel<M, R, SYMB, CAP>;
    IB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    int per_sm = 0;
    IB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem_bytes));
    if (per_sm < 1)
        return IB200_ERROR_UNSUPPORTED;
    const size_t n_tiles = (E + shape.EPE - 1) / shape.EPE;
    const size_t want = (n_tiles + shape.engines_per_cta - 1) / shape.engines_per_cta;
    const size_t cap = (size_t)cfg.sm_count * per_sm;
    const int grid = (int)std::max<size_t>(1, std::min(want, cap));
    kernel<<<grid, threads, smem_bytes, cfg.stream>>>(prm, plan.d_P1, plan.d_bmap, plan.d_gtab, E, wsrc, out, cfg.sched);
    ++*cfg.launch_counter;
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess)
        return ib200_fail(IB200_ERROR_CUDA, "launch of mass2d_kernel<%d,%d,%d> failed: %s", M, R, (int)SYMB, cudaGetErrorString(err));
    return INTERP_SUCCESS;
}

template <int M>
int dispatch_m(const LaunchCfg &cfg, const MassPlan &plan, size_t E, const double *wsrc, double *out)
{
    if (plan.R == 2)
        return plan.symb ? launch_mass<M, 2, true>(cfg, plan, E, wsrc, out) : launch_mass<M, 2, false>(cfg, plan, E, wsrc, out);
    return plan.symb ? launch_mass<M, 3, true>(cfg, plan, E, wsrc, out) : launch_mass<M, 3, false>(cfg, plan, E, wsrc, out);
}

} // namespace

// Host side: build the per-operation tables and choose the thread mapping (once per cache key).
static int build_plan(const OpDesc &op, const OpHost &host, bool match0, bool match1, int deriv_dim_in, int wsrc_kind,
                      const OpHost *geom_host, int widx_in, int widx_out, MassPlan &plan)
{
    const int M = op.dim[0].m;
    const int n_out0 = op.dim[0].n_out, n_in0 = op.dim[0].n_in, n_out1 = op.dim[1].n_out, n_in1 = op.dim[1].n_in;
    const bool syma = match0 && deriv_dim_in != 0;
    const bool symb = match1 && deriv_dim_in != 1;
    plan.M = M;
    plan.symb = symb;

    // P0[(j0,i0)][q0] = w0 * (Bout0[j0] * Tin0[i0]); the product of the two basis values is formed first so that
    // the table is exactly symmetric in (j0,i0) for matching spaces
    const double *Tin0 = deriv_dim_in == 0 ? host.Din[0] : host.Bin[0];
    const double *Tin1 = deriv_dim_in == 1 ? host.Din[1] : host.Bin[1];
    auto p0_entry = [&](int j0, int i0, int q) { return host.w[0][q] * (host.Bout[0][j0 * M + q] * Tin0[i0 * M + q]); };
    if (syma)
    {
        for (int hi = 0; hi < n_out0; ++hi)
            for (int lo = 0; lo <= hi; ++lo)
                for (int q = 0; q < M; ++q)
                    plan.P0.push_back(p0_entry(hi, lo, q));
    }
    else
    {
        for (int j0 = 0; j0 < n_out0; ++j0)
            for (int i0 = 0; i0 < n_in0; ++i0)
                for (int q = 0; q < M; ++q)
                    plan.P0.push_back(p0_entry(j0, i0, q));
    }
    if ((int)plan.P0.size() > cap_for(M))
        return IB200_ERROR_UNSUPPORTED;
    std::vector<double> P1, gtab(8 * M, 0.0);
    std::vector<int2> bmap;
    for (int j1 = 0; j1 < n_out1; ++j1)
        for (int i1 = 0; i1 < (symb ? j1 + 1 : n_in1); ++i1)
        {
            bmap.push_back(make_int2(j1, i1));
            for (int q = 0; q < M; ++q)
                P1.push_back(host.w[1][q] * (host.Bout[1][j1 * M + q] * Tin1[i1 * M + q]));
        }
    if (wsrc_kind == 2)
    {
        if (!geom_host)
            return IB200_ERROR_UNSUPPORTED;
        std::memcpy(gtab.data() + 0 * M, geom_host->Bin[0], sizeof(double) * 2 * M);
        std::memcpy(gtab.data() + 2 * M, geom_host->Din[0], sizeof(double) * 2 * M);
        std::memcpy(gtab.data() + 4 * M, geom_host->Bin[1], sizeof(double) * 2 * M);
        std::memcpy(gtab.data() + 6 * M, geom_host->Din[1], sizeof(double) * 2 * M);
    }

    MassShape &s = plan.s;
    s.n_out0 = n_out0;
    s.n_in0 = n_in0;
    s.n_out1 = n_out1;
    s.n_in1 = n_in1;
    s.N_in = n_in0 * n_in1;
    s.NN = op.N_out * op.N_in;
    s.NB = (int)bmap.size();
    s.syma = syma ? 1 : 0;
    s.wsrc = wsrc_kind;
    s.widx_in = widx_in;
    s.widx_out = widx_out;
    s.Qp = (M * M + 1) & ~1;
    // thread mapping: R items per thread, an engine of G warps hosts EPE = floor(32 G / BGp) elements; choose (R, G)
    // for lane utilisation NB*EPE / (R*32*G) (e.g. NB = 72, R = 3: 24 threads per element -> 3 warps host 4 elements
    // with every lane busy, one warp alone would idle 8 of 32).  Ties: fewer warps per engine, then the larger R
    // from 9 points per dimension (it amortises the coefficient loads; below that it costs registers).
    const int row_block = n_out1 * s.N_in;
    int best_R = 0, best_G = 0;
    double best_util = -1.0;
    for (int R = 2; R <= 3; ++R)
    {
        const int bgp = (s.NB + R - 1) / R;
        for (int G = (bgp + 31) / 32; G <= 4; ++G)
        {
            const int epe = (32 * G) / bgp;
            if (epe < 1 || (size_t)epe * row_block * sizeof(double) > 40 * 1024)
                continue; // one j0 row block of every hosted element must fit a staging buffer
            const double util = (double)s.NB * epe / (double)(R * 32 * G);
            const bool tie = util >= best_util - 1e-9 && util <= best_util + 1e-9;
            const bool better = util > best_util + 1e-9 || (tie && (G < best_G || (G == best_G && M >= 9)));
            if (better)
            {
                best_util = util;
                best_R = R;
                best_G = G;
            }
        }
    }
    if (best_R == 0)
        return IB200_ERROR_UNSUPPORTED;
    const int R = plan.R = best_R;
    s.BGp = (s.NB + R - 1) / R;
    s.G = best_G;
    s.EPE = (32 * s.G) / s.BGp;
    // panels: about 6 KB of staged output per engine and buffer (all EPE elements together), at least one j0
    s.PJ = std::max(1, std::min(n_out0, (int)(6144 / (sizeof(double) * row_block * s.EPE))));
    s.n_panels = (n_out0 + s.PJ - 1) / s.PJ;
    s.panel_cap = ((s.PJ * row_block + 2) + 1) & ~1;
    const size_t panel_bytes = sizeof(double) * s.EPE * s.panel_cap;
    // The kernel is latency-bound below ~16 resident warps (measured on the C3 blocks: 8 -> 16 warps per SM is
    // +25..40 %), so the second staging buffer is only kept when shared memory still allows 16 warps with it.
    const size_t fixed_bytes = sizeof(double) * s.EPE * s.Qp;
    const double warps_with_2 = 220.0 * 1024 / (double)(fixed_bytes + 2 * panel_bytes) * s.G;
    s.nbuf = (warps_with_2 >= 16.0) ? 2 : 1;
    if (const char *env = std::getenv("IB200_MASS_NBUF"))
        s.nbuf = std::atoi(env) == 1 ? 1 : (std::atoi(env) == 2 ? 2 : s.nbuf);
    const size_t engine_bytes = fixed_bytes + s.nbuf * panel_bytes;
    // engines per CTA: up to 256 threads and ~100 KB of shared memory (two CTAs per SM)
    int epc = std::max(1, 256 / (32 * s.G));
    while (epc > 1 && (epc * engine_bytes + sizeof(double) * 8 * M > 100 * 1024 || epc > 15))
        --epc;
    if (epc * engine_bytes + sizeof(double) * 8 * M > 200 * 1024)
        return IB200_ERROR_UNSUPPORTED;
    s.engines_per_cta = epc;

    // upload the per-operation tables (small, once)
    const size_t p1_bytes = P1.size() * sizeof(double), bm_bytes = bmap.size() * sizeof(int2), gt_bytes = gtab.size() * sizeof(double);
    const size_t o_bm = (p1_bytes + 255) & ~(size_t)255, o_gt = o_bm + ((bm_bytes + 255) & ~(size_t)255);
    IB_CUDA(cudaMalloc((void **)&plan.d_tab, o_gt + gt_bytes));
    IB_CUDA(cudaMemcpy(plan.d_tab, P1.data(), p1_bytes, cudaMemcpyHostToDevice));
    IB_CUDA(cudaMemcpy(plan.d_tab + o_bm, bmap.data(), bm_bytes, cudaMemcpyHostToDevice));
    IB_CUDA(cudaMemcpy(plan.d_tab + o_gt, gtab.data(), gt_bytes, cudaMemcpyHostToDevice));
    plan.d_P1 = (const double *)plan.d_tab;
    plan.d_bmap = (const int2 *)(plan.d_tab + o_bm);
    plan.d_gtab = (const double *)(plan.d_tab + o_gt);
    return INTERP_SUCCESS;
}

// geom_host: host copies of the geometry basis tables (order 1 per dimension) when the geometry is fused, else null.
// `key` identifies the operation (all spec integers) for the plan cache.
int launch_mass2d_fast(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, const OpDesc &op, const OpHost &host,
                       bool match0, bool match1, int deriv_dim_in, int wsrc_kind, const OpHost *geom_host, int widx_in,
                       int widx_out, size_t E, const double *wsrc, double *out)
{
    if (op.ndim != 2 || op.dim[0].m != op.dim[1].m)
        return IB200_ERROR_UNSUPPORTED;
    if ((reinterpret_cast<size_t>(out) & 15) != 0)
        return IB200_ERROR_UNSUPPORTED;
    // the reference mirrors the lower triangle even for gradient matrices when the spaces match
    // (mass_matrices.c:551-560); that quirk is reproduced by the generic kernel only
    if (deriv_dim_in >= 0 && match0 && match1)
        return IB200_ERROR_UNSUPPORTED;
    const int M = op.dim[0].m;
    switch (M)
    {
    case 2: case 3: case 4: case 5: case 6: case 7: case 8: case 9: case 10: case 12: case 18:
        break;
    default:
        return IB200_ERROR_UNSUPPORTED;
    }
    auto it = ctx->plans.find(key);
    if (it == ctx->plans.end())
    {
        auto plan = std::make_shared<MassPlan>();
        const int res = build_plan(op, host, match0, match1, deriv_dim_in, wsrc_kind, geom_host, widx_in, widx_out, *plan);
        if (res != INTERP_SUCCESS && res != IB200_ERROR_UNSUPPORTED)
            return res;
        if (res == IB200_ERROR_UNSUPPORTED)
            plan.reset(); // remember that this operation has no fast path
        it = ctx->plans.emplace(key, std::static_pointer_cast<void>(plan)).first;
    }
    if (!it->second)
        return IB200_ERROR_UNSUPPORTED;
    const MassPlan &plan = *static_cast<const MassPlan *>(it->second.get());
#define IB_M(MV) \
    case MV:     \
        return dispatch_m<MV>(cfg, plan, E, wsrc, out);
    switch (M)
    {
        IB_M(2) IB_M(3) IB_M(4) IB_M(5) IB_M(6) IB_M(7) IB_M(8) IB_M(9) IB_M(10) IB_M(12) IB_M(18)
    default:
        return IB200_ERROR_UNSUPPORTED;
    }
#undef IB_M
}
