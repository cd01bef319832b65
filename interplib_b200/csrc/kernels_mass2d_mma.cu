// K2-hi -- element mass / gradient-mass matrices of HIGH-ORDER 2-D tensor-product elements with the stage-2
// contraction on the FP64 tensor instruction (mma.sync.aligned.m8n8k4.f64, "DMMA").
//
//   out[e][(j0,j1)][(i0,i1)] = sum_{q0} P0[(j0,i0)][q0] * T_e[q0][(j1,i1)]            (stage 2: one GEMM per element)
//   T_e[q0][(j1,i1)]         = sum_{q1} Wt_e[q0][q1] * P1[(j1,i1)][q1]                 (stage 1: DFMA)
//
// Same mathematics and tables as kernels_mass2d.cu (reference loops: /root/reference/src/python/mass_matrices.c:264-335,
// :486-560; geometry /root/reference/src/python/mappings.c:448-542).  Why a second kernel: the measured FP64 peaks of
// DFMA and DMMA on B200 are equal (36.7 vs 37.1 TFLOP/s, tools/fp64_peak.cu), but at p = 16 the DFMA kernel is
// issue/latency-bound (8 warps/SM at 237 registers, FP64 pipe 30 % busy, profiles/r01_c4_mass2d_full.txt): one DMMA
// retires 256 multiply-adds per issue slot against 32 for a DFMA, and its operands come from registers instead of one
// uniform constant load per 3 DFMAs.  Stage 2 is 95 % of the flops at p = 16.
//
// GEMM mapping (per element): C[a][b] = sum_k A[a][k] B[k][b] with a = (j0,i0) (rows, tiles of 8), b = (j1,i1) column
// items (tiles of 8; only the pairs i1 <= j1 when the dim-1 spaces match, the mirror entry is the same number), k = q0
// padded to a multiple of 4.  A = P0 is the same for every element and sits in shared memory for the whole kernel
// ([rows][KS*4] doubles: the fragment loads are bank-conflict free); B = T_e is written to shared memory by stage 1
// with a row pitch = 4 (mod 16) doubles, again conflict free.
//
// A CTA = 8 compute warps + 1 store warp, one element at a time, persistent over elements.  Compute warp (cg, rg) owns
// the column-tile chunks c = cg (mod CG) and the row tiles rt = rg (mod RG).  The C fragments go to a row-panel staging
// buffer in shared memory (a panel = PJ values of j0 = one contiguous run of global memory; mirror entries are written
// there too, so the result is exactly symmetric) and every finished panel leaves the SM as ONE bulk async copy
// (cp.async.bulk shared -> global, TMA) issued by the store warp.  Panels are double buffered and handed over with
// mbarriers (full: 8 warp arrivals -> store warp; empty: store warp -> compute warps after the bulk read completed), so
// compute warps never wait for a store unless they are two panels ahead.
#include <algorithm>
#include <cstdlib>
#include <memory>
#include <string>
#include <vector>

#include "ctx.cuh"

namespace
{

constexpr int kComputeWarps = 8;
constexpr int kThreadsMma = (kComputeWarps + 1) * 32;

struct MmaShape
{
    int M;          // quadrature points per dimension
    int Q;          // M*M
    int n_in0, n_out1, n_in1;
    int N_in;       // row length of the output matrix
    int NN;         // doubles per element
    int NA;         // rows a = n_out0*n_in0
    int RT;         // row tiles
    int NB;         // column items
    int CT;         // column tiles
    int ldT;        // pitch of T in shared memory (doubles)
    int n_chunks;   // column chunks of CTW tiles
    int CG, RG;     // warp grid: CG*RG = 8
    int PJ, n_panels, panel_cap, row_block;
    int nbuf;       // panel staging buffers (2 or 3)
    int s1_split;   // stage 1: q0 range split so that NB*s1_split items fill the compute threads
    int wsrc, widx_in, widx_out;
};

__device__ __forceinline__ unsigned smem_u32(const void *p)
{
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned long long *bar, unsigned parity)
{
    unsigned ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// Spin on the barrier phase; a hand-over that does not complete within ~2 s is a protocol bug: trap (the launch then
// fails with an error) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity)
{
    if (mbar_try_wait(bar, parity))
        return;
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity))
        if (clock64() - t0 > 4000000000ll)
            __trap();
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
__device__ __forceinline__ void compute_barrier()
{
    asm volatile("bar.sync 1, %0;" ::"n"(kComputeWarps * 32) : "memory");
}
__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// Jacobian entry of a bilinear quad in the reference's operation order (see kernels_mass2d.cu)
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

// tables in global memory (built once per operation): P0pad[RT*8][KS*4] | P1[NB][M] | boff[CT*8] (int2) | gtab[8*M]
template <int KS, int CTW>
__global__ void __launch_bounds__(kThreadsMma, 1)
    mass2d_mma_kernel(const __grid_constant__ MmaShape s, const double *__restrict__ P0g, const double *__restrict__ P1g,
                      const int2 *__restrict__ boffg, const int2 *__restrict__ rowinfog, const double *__restrict__ gtab, size_t E,
                      const double *__restrict__ wsrc, double *__restrict__ out)
{
    constexpr int LDA = KS * 4;
    extern __shared__ __align__(16) double smem[];
    // layout: P0s[RT*8*LDA] | T[LDA*ldT] | W[M*LDA] | gt[8*M] | panels[nbuf*panel_cap] | boff[CT*8] int2 |
    //         rowinfo[RT*8] int2 {panel, offset in panel} | tileinfo[RT] int2 {last panel written, panels finished after
    //         this tile for this row group} | barriers[2*nbuf]
    double *P0s = smem;
    double *Ts = P0s + s.RT * 8 * LDA;
    double *Ws = Ts + LDA * s.ldT; // per-point weights, rows q0 padded to LDA (padding stays zero)
    double *gt = Ws + s.M * LDA;
    double *panels = gt + 8 * s.M;
    int2 *boff = reinterpret_cast<int2 *>(panels + (size_t)s.nbuf * s.panel_cap);
    int2 *rowinfo = boff + s.CT * 8;
    int2 *tileinfo = rowinfo + s.RT * 8;
    unsigned long long *bars = reinterpret_cast<unsigned long long *>(tileinfo + s.RT);
    unsigned long long *full = bars, *empty = bars + s.nbuf;

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const bool is_store_warp = warp == kComputeWarps;

    for (int i = tid; i < s.RT * 8 * LDA; i += blockDim.x)
        P0s[i] = P0g[i];
    for (int i = tid; i < LDA * s.ldT; i += blockDim.x)
        Ts[i] = 0.0; // padding rows (q0 >= M) and columns (b >= NB) stay zero for the whole kernel
    for (int i = tid; i < s.M * LDA; i += blockDim.x)
        Ws[i] = 0.0;
    for (int i = tid; i < s.CT * 8; i += blockDim.x)
        boff[i] = boffg[i];
    for (int i = tid; i < s.RT * 9; i += blockDim.x)
        rowinfo[i] = rowinfog[i]; // rowinfo and tileinfo are contiguous
    if (s.wsrc == 2)
        for (int i = tid; i < 8 * s.M; i += blockDim.x)
            gt[i] = gtab[i];
    if (tid == 0)
    {
        for (int b = 0; b < s.nbuf; ++b)
        {
            mbar_init(&full[b], kComputeWarps);
            mbar_init(&empty[b], 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const size_t first = blockIdx.x, stride = gridDim.x;

    if (is_store_warp)
    {
        // ---- store warp: one bulk copy per finished panel -------------------------------------------------
        if (lane == 0)
        {
            unsigned buf = 0, phase = 0; // ring position of the running panel counter (same sequence as the compute warps')
            for (size_t e = first; e < E; e += stride)
            {
                for (int p = 0; p < s.n_panels; ++p)
                {
                    mbar_wait(&full[buf], phase);
                    const int j0s = p * s.PJ;
                    const int rows = min(s.PJ, s.NA / s.n_in0 - j0s);
                    const int cnt = rows * s.row_block;
                    const size_t gstart = e * (size_t)s.NN + (size_t)j0s * s.row_block;
                    const int par = (int)(gstart & 1);
                    const double *sp = panels + (size_t)buf * s.panel_cap + par; // sp[k] <-> gdst[k]
                    double *gdst = out + gstart;
                    int k0 = 0;
                    if (par)
                    {
                        gdst[0] = sp[0];
                        k0 = 1;
                    }
                    const int body = (cnt - k0) & ~1;
                    if (body > 0)
                        bulk_s2g(gdst + k0, sp + k0, (unsigned)body * 8u);
                    if (k0 + body < cnt)
                        gdst[cnt - 1] = sp[cnt - 1];
                    bulk_commit();
                    // hand the buffer back as soon as the copy engine has READ it (the global write continues
                    // asynchronously); waiting for the next panel first would deadlock: filling panel p+1 can require
                    // the buffer of panel p-1 ... p+2
                    bulk_wait_read<0>();
                    mbar_arrive(&empty[buf]);
                    if (++buf == (unsigned)s.nbuf)
                    {
                        buf = 0;
                        phase ^= 1u;
                    }
                }
            }
            bulk_wait_all();
        }
        return;
    }

    // ---- compute warps -----------------------------------------------------------------------------------------
    const int cg = warp % s.CG, rg = warp / s.CG;
    const int quad_row = lane >> 2, quad_col = lane & 3;
    const double *gB0 = gt, *gD0 = gt + 2 * s.M, *gB1 = gt + 4 * s.M, *gD1 = gt + 6 * s.M;
    // ring state of panel 0 of the current element: buffer index and phase of the `empty`/`full` barriers
    unsigned ring_buf0 = 0, ring_phase0 = 0;
    int acquired = -1;     // highest panel (of the current element) this warp may write to
    int released = -1;     // highest panel (of the current element) this warp has arrived on

    for (size_t e = first; e < E; e += stride)
    {
        // ---- per-point weights -> Ws
        if (s.wsrc == 1)
        {
            const double *src = wsrc + e * s.Q;
            for (int q = tid; q < s.Q; q += kComputeWarps * 32)
                Ws[(q / s.M) * LDA + q % s.M] = src[q];
        }
        else if (s.wsrc == 2)
        {
            const double *gq = wsrc + e * 8; // [c][i0][i1]
            double cx[4], cy[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
            {
                cx[i] = gq[i];
                cy[i] = gq[4 + i];
            }
            for (int q = tid; q < s.Q; q += kComputeWarps * 32)
            {
                const int q0 = q / s.M, q1 = q % s.M;
                const double b0[2] = {gB0[q0], gB0[s.M + q0]}, d0[2] = {gD0[q0], gD0[s.M + q0]};
                const double b1[2] = {gB1[q1], gB1[s.M + q1]}, d1[2] = {gD1[q1], gD1[s.M + q1]};
                const double J00 = bilinear_ref_order(d0, b1, cx), J01 = bilinear_ref_order(b0, d1, cx);
                const double J10 = bilinear_ref_order(d0, b1, cy), J11 = bilinear_ref_order(b0, d1, cy);
                double wv;
                if (s.widx_in < 0)
                {
                    const double p = J01 * J10;
                    const double err = fma(-J01, J10, p);
                    wv = fma(J00, J11, -p) + err;
                }
                else
                {
                    wv = s.widx_in == 0 ? (s.widx_out == 0 ? J11 : -J01) : (s.widx_out == 0 ? -J10 : J00);
                }
                Ws[q0 * LDA + q1] = wv;
            }
        }
        else
        {
            for (int q = tid; q < s.Q; q += kComputeWarps * 32)
                Ws[(q / s.M) * LDA + q % s.M] = 1.0;
        }
        compute_barrier(); // Ws complete; also: every warp finished reading Ts of the previous element

        // ---- stage 1 (DFMA): T[q0][b] = sum_q1 W[q0][q1] * P1[b][q1]
        {
            const int per = (s.M + s.s1_split - 1) / s.s1_split;
            for (int item = tid; item < s.NB * s.s1_split; item += kComputeWarps * 32)
            {
                const int b = item % s.NB, part = item / s.NB;
                const int q0_lo = part * per, q0_hi = min(s.M, q0_lo + per);
                const double2 *p1 = reinterpret_cast<const double2 *>(P1g + (size_t)b * LDA); // rows padded to LDA with zeros
                double2 p1r[LDA / 2];
#pragma unroll
                for (int k = 0; k < LDA / 2; ++k)
                    p1r[k] = __ldg(p1 + k);
                for (int q0 = q0_lo; q0 < q0_hi; ++q0)
                {
                    const double2 *w = reinterpret_cast<const double2 *>(Ws + q0 * LDA);
                    double acc = 0.0;
#pragma unroll
                    for (int k = 0; k < LDA / 2; ++k)
                    {
                        const double2 wv = w[k];
                        acc = fma(wv.x, p1r[k].x, acc);
                        acc = fma(wv.y, p1r[k].y, acc);
                    }
                    Ts[q0 * s.ldT + b] = acc;
                }
            }
        }
        compute_barrier();

        // ---- stage 2 (DMMA)
        acquired = -1;
        released = -1;
        // ring position of the next panel to acquire / to release (advance with the counters above)
        unsigned acq_buf = ring_buf0, acq_phase = ring_phase0, rel_buf = ring_buf0;
        auto ring_next = [&](unsigned &b, unsigned &ph) {
            if (++b == (unsigned)s.nbuf)
            {
                b = 0;
                ph ^= 1u;
            }
        };
        auto acquire_next = [&]() {
            ++acquired;
            mbar_wait(&empty[acq_buf], acq_phase ^ 1u);
            ring_next(acq_buf, acq_phase);
        };
        unsigned dummy_phase = 0;
        auto release_upto = [&](int last) {
            if (released >= last)
                return;
            fence_async_smem();
            __syncwarp();
            while (released < last)
            {
                ++released;
                if (acquired < released) // a panel this warp never wrote to is still acquired in order
                    acquire_next();
                if (lane == 0)
                    mbar_arrive(&full[rel_buf]);
                ring_next(rel_buf, dummy_phase);
            }
        };
        const int e_par = (int)((e * (size_t)s.NN) & 1);
        const int panel_par = (s.PJ * s.row_block) & 1;

        double Bf[CTW][KS];
        int2 off[CTW][2];
        int loaded_chunk = -1;
        for (int rt = rg; rt < s.RT; rt += s.RG)
        {
            // A fragments of this row tile: A[row = quad_row][k = ks*4 + quad_col]
            double Af[KS];
            const int a_row = rt * 8 + quad_row;
#pragma unroll
            for (int ks = 0; ks < KS; ++ks)
                Af[ks] = P0s[a_row * LDA + ks * 4 + quad_col];
            const int2 ti = tileinfo[rt]; // x: last panel this tile writes to, y: panels finished once this row group is past it
            const int2 ri = rowinfo[a_row]; // x: panel of this lane's row (-1: padding row), y: offset of the row in its panel
            while (acquired < ti.x)
                acquire_next();
            const bool row_ok = ri.x >= 0;
            const int p = row_ok ? ri.x : 0;
            unsigned pbuf = ring_buf0 + (unsigned)p;
            pbuf %= (unsigned)s.nbuf;
            double *dst = panels + (size_t)pbuf * s.panel_cap + ((e_par + (p & 1) * panel_par) & 1) + ri.y;

            for (int chunk = cg; chunk < s.n_chunks; chunk += s.CG)
            {
                if (chunk != loaded_chunk)
                {
                    loaded_chunk = chunk;
#pragma unroll
                    for (int ct = 0; ct < CTW; ++ct)
                    {
                        const int col = (chunk * CTW + ct) * 8;
                        const bool ct_ok = chunk * CTW + ct < s.CT;
#pragma unroll
                        for (int ks = 0; ks < KS; ++ks)
                            Bf[ct][ks] = ct_ok ? Ts[(ks * 4 + quad_col) * s.ldT + col + quad_row] : 0.0;
                        off[ct][0] = ct_ok ? boff[col + 2 * quad_col] : make_int2(-1, -1);
                        off[ct][1] = ct_ok ? boff[col + 2 * quad_col + 1] : make_int2(-1, -1);
                    }
                }
                double C[CTW][2];
#pragma unroll
                for (int ct = 0; ct < CTW; ++ct)
                    C[ct][0] = C[ct][1] = 0.0;
#pragma unroll
                for (int ks = 0; ks < KS; ++ks)
#pragma unroll
                    for (int ct = 0; ct < CTW; ++ct)
                        dmma884(C[ct][0], C[ct][1], Af[ks], Bf[ct][ks]);
#pragma unroll
                for (int ct = 0; ct < CTW; ++ct)
#pragma unroll
                    for (int v = 0; v < 2; ++v)
                    {
                        const int2 o = off[ct][v];
                        const bool px = row_ok && o.x >= 0, py = row_ok && o.y >= 0;
                        if (px)
                            dst[o.x] = C[ct][v];
                        if (py)
                            dst[o.y] = C[ct][v];
                    }
            }
            release_upto(ti.y - 1);
        }
        release_upto(s.n_panels - 1); // warps without any row tile (RT < RG) still owe their arrivals
        // advance the ring to panel 0 of the next element
        for (int k = 0; k < s.n_panels % s.nbuf; ++k)
            ring_next(ring_buf0, ring_phase0);
        if ((s.n_panels / s.nbuf) & 1)
            ring_phase0 ^= 1u;
    }
}

struct MmaPlan
{
    MmaShape s{};
    int KS = 0, CTW = 0;
    size_t smem_bytes = 0;
    char *d_tab = nullptr;
    const double *d_P0 = nullptr, *d_P1 = nullptr, *d_gtab = nullptr;
    const int2 *d_boff = nullptr, *d_rowinfo = nullptr;
    ~MmaPlan()
    {
        if (d_tab)
            cudaFree(d_tab);
    }
};

template <int KS, int CTW>
int launch_mma(const LaunchCfg &cfg, const MmaPlan &plan, size_t E, const double *wsrc, double *out)
{
    auto kernel = mass2d_mma_kernel<KS, CTW>;
    IB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.smem_bytes));
    int per_sm = 0;
    IB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kThreadsMma, plan.smem_bytes));
    if (per_sm < 1)
        return IB200_ERROR_UNSUPPORTED;
    const size_t cap = (size_t)cfg.sm_count * per_sm;
    const int grid = (int)std::max<size_t>(1, std::min(E, cap));
    kernel<<<grid, kThreadsMma, plan.smem_bytes, cfg.stream>>>(plan.s, plan.d_P0, plan.d_P1, plan.d_boff, plan.d_rowinfo, plan.d_gtab, E, wsrc,
                                                                out);
    ++*cfg.launch_counter;
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess)
        return ib200_fail(IB200_ERROR_CUDA, "launch of mass2d_mma_kernel<%d,%d> failed: %s", KS, CTW, cudaGetErrorString(err));
    return INTERP_SUCCESS;
}

template <int KS>
int dispatch_ctw(const LaunchCfg &cfg, const MmaPlan &plan, size_t E, const double *wsrc, double *out)
{
    switch (plan.CTW)
    {
    case 2: return launch_mma<KS, 2>(cfg, plan, E, wsrc, out);
    case 3: return launch_mma<KS, 3>(cfg, plan, E, wsrc, out);
    case 4: return launch_mma<KS, 4>(cfg, plan, E, wsrc, out);
    case 5: return launch_mma<KS, 5>(cfg, plan, E, wsrc, out);
    default: return IB200_ERROR_UNSUPPORTED;
    }
}

int build_mma_plan(const OpDesc &op, const OpHost &host, bool match0, bool match1, int deriv_dim_in, int wsrc_kind,
                   const OpHost *geom_host, int widx_in, int widx_out, MmaPlan &plan)
{
    const int M = op.dim[0].m;
    const int n_out0 = op.dim[0].n_out, n_in0 = op.dim[0].n_in, n_out1 = op.dim[1].n_out, n_in1 = op.dim[1].n_in;
    const bool symb = match1 && deriv_dim_in != 1;
    (void)match0;
    const int KS = (M + 3) / 4, LDA = KS * 4;
    if (KS < 3 || KS > 5)
        return IB200_ERROR_UNSUPPORTED;
    MmaShape &s = plan.s;
    s.M = M;
    s.Q = M * M;
    s.n_in0 = n_in0;
    s.n_out1 = n_out1;
    s.n_in1 = n_in1;
    s.N_in = n_in0 * n_in1;
    s.NN = op.N_out * op.N_in;
    s.NA = n_out0 * n_in0;
    s.RT = (s.NA + 7) / 8;
    s.row_block = n_out1 * s.N_in;
    s.wsrc = wsrc_kind;
    s.widx_in = widx_in;
    s.widx_out = widx_out;

    // tables: P0 as in kernels_mass2d.cu (the product of the two basis values first -> exactly symmetric in (j0,i0))
    const double *Tin0 = deriv_dim_in == 0 ? host.Din[0] : host.Bin[0];
    const double *Tin1 = deriv_dim_in == 1 ? host.Din[1] : host.Bin[1];
    std::vector<double> P0((size_t)s.RT * 8 * LDA, 0.0), P1, gtab(8 * M, 0.0);
    for (int j0 = 0; j0 < n_out0; ++j0)
        for (int i0 = 0; i0 < n_in0; ++i0)
            for (int q = 0; q < M; ++q)
                P0[(size_t)(j0 * n_in0 + i0) * LDA + q] = host.w[0][q] * (host.Bout[0][j0 * M + q] * Tin0[i0 * M + q]);
    // Column items b = (j1,i1) (pairs i1 <= j1 when the dimension-1 spaces match).  Their ORDER is free: eight consecutive
    // items form one DMMA column tile, lane (quad_row, quad_col) stores columns 2*quad_col + v of its row, and a 64-bit
    // shared-memory store is served per half warp (quad_row 0..3 x quad_col 0..3).  Rows of a tile sit 17 (= 1 mod 16)
    // doubles apart, so the staging stores of a half warp are conflict-free iff the four items at columns {v, 2+v, 4+v,
    // 6+v} have offsets that are pairwise >= 4 apart modulo 16 -- for the entry and for its mirror.  Greedy tile
    // construction on that cost (row-major order had ~2-way conflicts on both stores, 41 % of all wavefronts, ncu).
    struct Item
    {
        int j1, i1, od, om;
    };
    std::vector<Item> pool, items;
    for (int j1 = 0; j1 < n_out1; ++j1)
        for (int i1 = 0; i1 < (symb ? j1 + 1 : n_in1); ++i1)
            pool.push_back({j1, i1, j1 * s.N_in + i1, (symb && i1 != j1) ? i1 * s.N_in + j1 : -1});
    {
        const int row_step = n_in1 % 16; // bank-slot step between consecutive rows a = (j0,i0) of a tile
        auto store_cost = [&](const std::vector<Item> &tile, const Item &cand, bool mirror) {
            // extra wavefronts when `cand` takes the next column of `tile`: lanes of the same parity v collide when
            // (offset + quad_row*row_step) mod 16 coincide
            const int c = (int)tile.size(), v = c & 1;
            const int oc = mirror ? cand.om : cand.od;
            if (oc < 0)
                return 0;
            int cost = 0;
            for (int k = v; k < c; k += 2)
            {
                const int ok = mirror ? tile[k].om : tile[k].od;
                if (ok < 0)
                    continue;
                for (int ra = 0; ra < 4; ++ra)
                    for (int rb = 0; rb < 4; ++rb)
                        if (((oc + ra * row_step) - (ok + rb * row_step)) % 16 == 0)
                            ++cost;
            }
            return cost;
        };
        std::vector<Item> tile;
        while (!pool.empty())
        {
            size_t best_i = 0;
            int best_c = 1 << 30;
            for (size_t i = 0; i < pool.size(); ++i)
            {
                const int c = store_cost(tile, pool[i], false) + store_cost(tile, pool[i], true);
                if (c < best_c)
                {
                    best_c = c;
                    best_i = i;
                }
            }
            tile.push_back(pool[best_i]);
            items.push_back(pool[best_i]);
            pool.erase(pool.begin() + (long)best_i);
            if (tile.size() == 8)
                tile.clear();
        }
    }
    std::vector<int2> boff;
    for (const Item &it : items)
    {
        boff.push_back(make_int2(it.od, it.om));
        for (int q = 0; q < LDA; ++q)
            P1.push_back(q < M ? host.w[1][q] * (host.Bout[1][it.j1 * M + q] * Tin1[it.i1 * M + q]) : 0.0);
    }
    s.NB = (int)boff.size();
    s.CT = (s.NB + 7) / 8;
    boff.resize((size_t)s.CT * 8, make_int2(-1, -1));
    s.ldT = s.CT * 8 + ((s.CT & 1) ? 12 : 4);
    if (wsrc_kind == 2)
    {
        if (!geom_host)
            return IB200_ERROR_UNSUPPORTED;
        std::memcpy(gtab.data() + 0 * M, geom_host->Bin[0], sizeof(double) * 2 * M);
        std::memcpy(gtab.data() + 2 * M, geom_host->Din[0], sizeof(double) * 2 * M);
        std::memcpy(gtab.data() + 4 * M, geom_host->Bin[1], sizeof(double) * 2 * M);
        std::memcpy(gtab.data() + 6 * M, geom_host->Din[1], sizeof(double) * 2 * M);
    }

    // warp grid: minimise the DMMA count of the busiest warp
    long best = -1;
    for (int ctw = 2; ctw <= 5; ++ctw)
        for (int cgv = 1; cgv <= kComputeWarps; cgv *= 2)
        {
            const int chunks = (s.CT + ctw - 1) / ctw, rgv = kComputeWarps / cgv;
            const long cost = (long)((chunks + cgv - 1) / cgv) * ctw * ((s.RT + rgv - 1) / rgv);
            if (best < 0 || cost < best || (cost == best && ctw > plan.CTW))
            {
                best = cost;
                plan.CTW = ctw;
                s.CG = cgv;
                s.RG = rgv;
                s.n_chunks = chunks;
            }
        }
    plan.KS = KS;
    // panels: whole j0 blocks, about 40 KB per buffer
    s.PJ = std::max(1, std::min(n_out0, (int)(40960 / (sizeof(double) * s.row_block))));
    s.n_panels = (n_out0 + s.PJ - 1) / s.PJ;
    s.panel_cap = ((s.PJ * s.row_block + 2) + 1) & ~1;
    // stage 1 split: about two passes of the 256 compute threads
    s.s1_split = std::max(1, std::min(M, (2 * kComputeWarps * 32) / std::max(1, s.NB)));
    // a row tile (8 rows of a) must not span more than two panels (hand-over protocol of the kernel)
    while (s.n_panels > 1 && s.PJ * n_in0 < 8)
    {
        ++s.PJ;
        s.n_panels = (n_out0 + s.PJ - 1) / s.PJ;
        s.panel_cap = ((s.PJ * s.row_block + 2) + 1) & ~1;
    }
    if (const char *env = std::getenv("IB200_MMA_CG")) // experiment knob: force the number of column groups
    {
        const int cgv = std::atoi(env);
        if (cgv == 1 || cgv == 2 || cgv == 4 || cgv == 8)
        {
            s.CG = cgv;
            s.RG = kComputeWarps / cgv;
            long bestc = -1;
            for (int ctw = 2; ctw <= 5; ++ctw)
            {
                const int chunks = (s.CT + ctw - 1) / ctw;
                const long cost = (long)((chunks + cgv - 1) / cgv) * ctw;
                if (bestc < 0 || cost < bestc || (cost == bestc && ctw > plan.CTW))
                {
                    bestc = cost;
                    plan.CTW = ctw;
                    s.n_chunks = chunks;
                }
            }
        }
    }
    auto smem_for = [&](int nbuf) {
        return sizeof(double) * ((size_t)s.RT * 8 * LDA + (size_t)LDA * s.ldT + (size_t)M * LDA + 8 * M + (size_t)nbuf * s.panel_cap) +
               sizeof(int2) * ((size_t)s.CT * 8 + (size_t)s.RT * 9) + 2 * nbuf * sizeof(unsigned long long);
    };
    s.nbuf = (s.n_panels >= 3 && smem_for(3) <= 200 * 1024) ? 3 : 2;
    if (const char *env = std::getenv("IB200_MMA_NBUF"))
        if (std::atoi(env) == 2 || (std::atoi(env) == 3 && smem_for(3) <= 220 * 1024))
            s.nbuf = std::atoi(env);
    plan.smem_bytes = smem_for(s.nbuf);
    if (plan.smem_bytes > 220 * 1024)
        return IB200_ERROR_UNSUPPORTED;
    // per-row and per-tile bookkeeping (replaces integer divisions in the kernel)
    std::vector<int2> rowinfo((size_t)s.RT * 9, make_int2(-1, 0));
    for (int a = 0; a < s.NA; ++a)
    {
        const int j0 = a / n_in0, i0 = a % n_in0, p = j0 / s.PJ;
        rowinfo[a] = make_int2(p, (j0 - p * s.PJ) * s.row_block + i0 * n_in1);
    }
    for (int rt = 0; rt < s.RT; ++rt)
    {
        const int a_last = std::min(rt * 8 + 7, s.NA - 1), next_first = (rt + s.RG) * 8;
        const int p_hi = (a_last / n_in0) / s.PJ;
        const int done_below = next_first >= s.NA ? s.n_panels : (next_first / n_in0) / s.PJ;
        rowinfo[(size_t)s.RT * 8 + rt] = make_int2(p_hi, done_below);
    }

    const size_t b0 = P0.size() * sizeof(double), b1 = P1.size() * sizeof(double), b2 = boff.size() * sizeof(int2),
                 b3 = gtab.size() * sizeof(double), b4 = rowinfo.size() * sizeof(int2);
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const size_t o1 = up(b0), o2 = o1 + up(b1), o3 = o2 + up(b2), o4 = o3 + up(b3);
    IB_CUDA(cudaMalloc((void **)&plan.d_tab, o4 + b4));
    IB_CUDA(cudaMemcpy(plan.d_tab + o4, rowinfo.data(), b4, cudaMemcpyHostToDevice));
    plan.d_rowinfo = (const int2 *)(plan.d_tab + o4);
    IB_CUDA(cudaMemcpy(plan.d_tab, P0.data(), b0, cudaMemcpyHostToDevice));
    IB_CUDA(cudaMemcpy(plan.d_tab + o1, P1.data(), b1, cudaMemcpyHostToDevice));
    IB_CUDA(cudaMemcpy(plan.d_tab + o2, boff.data(), b2, cudaMemcpyHostToDevice));
    IB_CUDA(cudaMemcpy(plan.d_tab + o3, gtab.data(), b3, cudaMemcpyHostToDevice));
    plan.d_P0 = (const double *)plan.d_tab;
    plan.d_P1 = (const double *)(plan.d_tab + o1);
    plan.d_boff = (const int2 *)(plan.d_tab + o2);
    plan.d_gtab = (const double *)(plan.d_tab + o3);
    return INTERP_SUCCESS;
}

} // namespace

// High-order path: chosen by the caller for M >= kMmaMinPoints (see capi.cu); same contract as launch_mass2d_fast.
int launch_mass2d_mma(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, const OpDesc &op, const OpHost &host,
                      bool match0, bool match1, int deriv_dim_in, int wsrc_kind, const OpHost *geom_host, int widx_in,
                      int widx_out, size_t E, const double *wsrc, double *out)
{
    if (op.ndim != 2 || op.dim[0].m != op.dim[1].m)
        return IB200_ERROR_UNSUPPORTED;
    if ((reinterpret_cast<size_t>(out) & 15) != 0)
        return IB200_ERROR_UNSUPPORTED;
    if (deriv_dim_in >= 0 && match0 && match1)
        return IB200_ERROR_UNSUPPORTED; // mirrored gradient quirk: generic kernel only (see kernels_mass2d.cu)
    const std::string k = "mma:" + key;
    auto it = ctx->plans.find(k);
    if (it == ctx->plans.end())
    {
        auto plan = std::make_shared<MmaPlan>();
        const int res = build_mma_plan(op, host, match0, match1, deriv_dim_in, wsrc_kind, geom_host, widx_in, widx_out, *plan);
        if (res != INTERP_SUCCESS && res != IB200_ERROR_UNSUPPORTED)
            return res;
        if (res == IB200_ERROR_UNSUPPORTED)
            plan.reset();
        it = ctx->plans.emplace(k, std::static_pointer_cast<void>(plan)).first;
    }
    if (!it->second)
        return IB200_ERROR_UNSUPPORTED;
    const MmaPlan &plan = *static_cast<const MmaPlan *>(it->second.get());
    switch (plan.KS)
    {
    case 3: return dispatch_ctw<3>(cfg, plan, E, wsrc, out);
    case 4: return dispatch_ctw<4>(cfg, plan, E, wsrc, out);
    case 5: return dispatch_ctw<5>(cfg, plan, E, wsrc, out);
    default: return IB200_ERROR_UNSUPPORTED;
    }
}
