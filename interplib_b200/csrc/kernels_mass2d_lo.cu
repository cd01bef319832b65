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
#include <cstdlib>
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

// TILE = true: the whole matrices of the warp's EPW elements are staged back to back and leave as ONE bulk copy (a
// warp tile is one contiguous run of global memory).  Used for odd N, where the row panels of an element start and end
// in the middle of 32-byte sectors: bulk stores of 2 000 B with sector-unaligned ends reach 5.58 TB/s, 20 000 B ones
// 6.37 TB/s, aligned 2 KB ones 6.29 TB/s (tools/write_peak.cu, profiles/r01_write_peak.json).
#ifndef IB200_LO_SCHED_CHUNK
#define IB200_LO_SCHED_CHUNK 2
#endif

template <int N, int M, int R, bool TILE>
struct LoShape
{
    static constexpr int PB = 8192;                  // panel byte budget per warp and buffer (panel mode)
    static constexpr int NB = N * (N + 1) / 2;       // column pairs (j1 >= i1)
    static constexpr int TPE = (NB + R - 1) / R;     // threads per element
    static constexpr int EPW = 32 / TPE;             // elements per warp
    static constexpr int Q = M * M;
    static constexpr int Qp = (Q + 1) & ~1;
    static constexpr int ROWB = N * N * N;           // doubles per j0 (rows (j0, 0..N-1))
    static constexpr int NN = ROWB * N;              // doubles per element
    static constexpr int PJ_RAW = PB / (8 * ROWB * EPW);
    static constexpr int PJ = TILE ? N : (PJ_RAW < 1 ? 1 : (PJ_RAW > N ? N : PJ_RAW));
    static constexpr int NPANELS = (N + PJ - 1) / PJ;
    // staged doubles per element and panel: panel mode adds parity slack and (run time, LoParams::cap) an even bank
    // skew of 0..14; tile mode is linear like global memory
    static constexpr int CAP0 = TILE ? NN : ((PJ * ROWB + 2 + 1) & ~1);
    static constexpr int NBUF = (TILE && EPW * NN * 16 > 24 * 1024) ? 1 : 2;
    static constexpr int WARPS = TILE ? 3 : 4;
    static_assert(!TILE || (EPW % 2 == 0), "tile mode needs 16-byte aligned tile starts");
    static constexpr int SLOTS = R * TPE;                 // column-pair slots per element (>= NB)
    // per warp: W[EPW][Qp] | parked inputs (run-time size, LoParams::in_doubles) | panels NBUF*EPW*cap
};

template <int N, int M>
struct LoParams
{
    double P0[N * (N + 1) / 2][M]; // pairs hi >= lo at hi*(hi+1)/2 + lo
    double gB0[2][M], gD0[2][M], gB1[2][M], gD1[2][M];
    double Bw0[N][M], Bw1[N][M];   // basis value * rule weight per dimension (fused right-hand side)
    int wsrc;                      // 0: unit weights, 1: weight[E][Q], 2: bilinear geometry [E][2][2][2] -> det
    int aux_per;                   // prefetched doubles per element of the weight source: 0, Q or 8
    int in_doubles;                // doubles of the per-warp input buffer: EPW * (aux_per + (rhs ? Q : 0)), even
    int cap;                       // staged doubles per element and panel (CAP0 + bank skew, even)
};

template <int N, int M, int R, bool TILE>
__global__ void __launch_bounds__(LoShape<N, M, R, TILE>::WARPS * 32)
    mass2d_lo_kernel(const __grid_constant__ LoParams<N, M> prm, const double *__restrict__ P1g, const int2 *__restrict__ offs, size_t E,
                     const double *__restrict__ wsrc, double *__restrict__ out, const double *__restrict__ fvals,
                     double *__restrict__ rhs, unsigned long long *__restrict__ sched)
{
    using S = LoShape<N, M, R, TILE>;
    extern __shared__ __align__(16) double smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int el = lane / S::TPE, bg = lane % S::TPE;
    const bool lane_active = el < S::EPW;
    const bool leader = lane_active && bg == 0;
    const bool issuer = TILE ? lane == 0 : leader; // lanes that issue (and track) bulk copies

    // CTA-shared: element-independent geometry products per point, pt[q][0..3] = dphi0[i0]*phi1[i1] (d/dxi0),
    // pt[q][4..7] = phi0[i0]*dphi1[i1] (d/dxi1), rounded like the reference's basis product
    double *pt = smem;
    double *Wsm = smem + 8 * S::Q + (size_t)warp * (S::EPW * S::Qp + prm.in_doubles + S::NBUF * S::EPW * prm.cap);
    double *Insm = Wsm + S::EPW * S::Qp;
    double *stage = Insm + prm.in_doubles;
    const bool want_rhs = rhs != nullptr;
    if (prm.wsrc == 2)
    {
        for (int i = threadIdx.x; i < 8 * S::Q; i += blockDim.x)
        {
            // layout [k/2][q][k%2]: four arrays of Q double2 -> lanes with consecutive q read consecutive 16-byte words
            const int q = i >> 3, k = i & 7, q0 = q / M, q1 = q % M, i0 = (k >> 1) & 1, i1 = k & 1;
            pt[((k >> 1) * S::Q + q) * 2 + (k & 1)] =
                (k < 4) ? __dmul_rn(prm.gD0[i0][q0], prm.gB1[i1][q1]) : __dmul_rn(prm.gB0[i0][q0], prm.gD1[i1][q1]);
        }
    }
    __syncthreads();

    // the thread's column pairs: slot bg + r*TPE of the host-built assignment (conflict-minimising, see assign_slots):
    // staging offsets of the entry (j1,i1) and of its mirror (-1: none), and the slot's row of P1
    int off_d[R], off_m[R];
    double p1[R][M];
#pragma unroll
    for (int r = 0; r < R; ++r)
    {
        const int slot = bg + r * S::TPE;
        const int2 o = lane_active ? offs[slot] : make_int2(-1, -1);
        off_d[r] = o.x;
        off_m[r] = o.y;
#pragma unroll
        for (int q = 0; q < M; ++q)
            p1[r][q] = lane_active ? P1g[(size_t)slot * M + q] : 0.0;
    }

    const size_t n_tiles = (E + S::EPW - 1) / S::EPW;
    const size_t n_warps = (size_t)gridDim.x * S::WARPS;
    unsigned pc = 0; // running panel counter of this warp

    // The tile's inputs (8 corner coordinates or Q weights per element, and the Q sampled values) are loaded into
    // registers one tile ahead (lane-strided, coalesced) and parked in per-warp shared memory at the top of their tile,
    // so their global-memory latency hides behind stages 1 and 2.  (Not cp.async: its wait shares the scoreboard of
    // cp.async.bulk.wait_group.read in SASS -- DEPBAR.LE SB0 -- and would serialise the tile behind its predecessor's
    // bulk store; measured with ncu, 36 % of the stall samples.)
    constexpr int NPF = (S::EPW * S::Q + 31) / 32;
    constexpr int NPA = NPF > (S::EPW * 8 + 31) / 32 ? NPF : (S::EPW * 8 + 31) / 32;
    double pre_aux[NPA], pre_f[NPF];
    auto prefetch = [&](size_t t) {
        const size_t e0 = t * S::EPW;
        const int nv = t < n_tiles ? (int)((E - e0) < (size_t)S::EPW ? (E - e0) : (size_t)S::EPW) : 0;
        const int per = prm.aux_per;
        const double *src = wsrc + e0 * per;
#pragma unroll
        for (int k = 0; k < NPA; ++k)
            pre_aux[k] = (lane + 32 * k < nv * per) ? src[lane + 32 * k] : 0.0;
        if (want_rhs)
        {
            const double *fsrc = fvals + e0 * S::Q;
#pragma unroll
            for (int k = 0; k < NPF; ++k)
                pre_f[k] = (lane + 32 * k < nv * S::Q) ? fsrc[lane + 32 * k] : 0.0;
        }
    };
    // Dynamic scheduling per warp, in chunks of CH consecutive tiles: the first chunk is the warp's global index, every
    // further one is claimed from sched[0] by lane 0 one chunk ahead (the atomic's latency hides behind a whole chunk).
    // Warps are autonomous and the FP64 pipe is per SM sub-partition: with 9 warps per SM a static split leaves the
    // sub-partitions hosting two warps idle while the one hosting three finishes (16 % of the SMSP cycles, ncu).
    constexpr size_t CH = IB200_LO_SCHED_CHUNK;
    size_t tile = ((size_t)blockIdx.x * S::WARPS + warp) * CH, tile_end = tile + CH;
    unsigned long long pend = 0;
    if (lane == 0)
        pend = atomicAdd(sched, 1ull);
    prefetch(tile);

    while (tile < n_tiles)
    {
        size_t tile_next = tile + 1, tile_next_end = tile_end;
        if (tile_next >= tile_end)
        {
            const unsigned long long c = __shfl_sync(0xffffffffu, pend, 0);
            tile_next = (n_warps + c) * CH;
            tile_next_end = tile_next + CH;
            if (lane == 0 && tile_next < n_tiles)
                pend = atomicAdd(sched, 1ull);
        }
        const size_t e = tile * S::EPW + el;
        const bool el_valid = lane_active && e < E;

        // ---- per-point weights of the tile's elements -> shared memory (per warp)
        __syncwarp(); // readers of the previous tile's inputs and weights are done
#pragma unroll
        for (int k = 0; k < NPA; ++k)
            if (lane + 32 * k < S::EPW * prm.aux_per)
                Insm[lane + 32 * k] = pre_aux[k];
        if (want_rhs)
        {
#pragma unroll
            for (int k = 0; k < NPF; ++k)
                if (lane + 32 * k < S::EPW * S::Q)
                    Insm[S::EPW * prm.aux_per + lane + 32 * k] = pre_f[k];
        }
        __syncwarp();
        prefetch(tile_next);
        if (el_valid)
        {
            double *W = Wsm + el * S::Qp;
            const double *in_aux = Insm;
            double *G = Insm + S::EPW * prm.aux_per + el * S::Q; // f in, f*W out (in place)
            const double *in_f = G;
            if (prm.wsrc == 2)
            {
                double cx[4], cy[4];
#pragma unroll
                for (int i = 0; i < 4; ++i)
                {
                    cx[i] = in_aux[el * 8 + i]; // [c][i0][i1]
                    cy[i] = in_aux[el * 8 + 4 + i];
                }
#pragma unroll 1
                for (int q = bg; q < S::Q; q += S::TPE)
                {
                    const double2 *pq = reinterpret_cast<const double2 *>(pt) + q;
                    const double2 a01 = pq[0], a23 = pq[S::Q], b01 = pq[2 * S::Q], b23 = pq[3 * S::Q];
                    // val = ((0 + p0*c0) + p1*c1) + ... with separate roundings (degrees_of_freedom.c:264-303)
                    const double J00 = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(a01.x, cx[0]), __dmul_rn(a01.y, cx[1])), __dmul_rn(a23.x, cx[2])), __dmul_rn(a23.y, cx[3]));
                    const double J01 = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(b01.x, cx[0]), __dmul_rn(b01.y, cx[1])), __dmul_rn(b23.x, cx[2])), __dmul_rn(b23.y, cx[3]));
                    const double J10 = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(a01.x, cy[0]), __dmul_rn(a01.y, cy[1])), __dmul_rn(a23.x, cy[2])), __dmul_rn(a23.y, cy[3]));
                    const double J11 = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(b01.x, cy[0]), __dmul_rn(b01.y, cy[1])), __dmul_rn(b23.x, cy[2])), __dmul_rn(b23.y, cy[3]));
                    // difference of products with the rounding error of the second product recovered (Kahan)
                    const double p = J01 * J10;
                    const double err = fma(-J01, J10, p);
                    const double det = fma(J00, J11, -p) + err;
                    W[q] = det;
                    if (want_rhs)
                        G[q] = __dmul_rn(in_f[q], det);
                }
            }
            else if (prm.wsrc == 1)
            {
#pragma unroll 1
                for (int q = bg; q < S::Q; q += S::TPE)
                {
                    const double wv = in_aux[el * S::Q + q];
                    W[q] = wv;
                    if (want_rhs)
                        G[q] = __dmul_rn(in_f[q], wv);
                }
            }
            else
            {
#pragma unroll 1
                for (int q = bg; q < S::Q; q += S::TPE)
                {
                    W[q] = 1.0;
                    if (want_rhs)
                        G[q] = in_f[q];
                }
            }
        }
        __syncwarp();

        // ---- fused L2 right-hand side (K3, integration.py:143-166): b[i0][i1] = sum_q0 Bw0[i0][q0] sum_q1 Bw1[i1][q1] (f*W)[q0][q1]
        // same operation order as l2dual2d_kernel, so the fused and the separate results are bit-identical
        if (want_rhs && el_valid)
        {
            const double *G = Insm + S::EPW * prm.aux_per + el * S::Q;
            for (int i1 = bg; i1 < N; i1 += S::TPE)
            {
                double bw1[M], sv[M];
#pragma unroll
                for (int q1 = 0; q1 < M; ++q1)
                    bw1[q1] = prm.Bw1[i1][q1];
#pragma unroll
                for (int q0 = 0; q0 < M; ++q0)
                {
                    double a = 0.0;
#pragma unroll
                    for (int q1 = 0; q1 < M; ++q1)
                        a = fma(G[q0 * M + q1], bw1[q1], a);
                    sv[q0] = a;
                }
                double *o = rhs + e * (size_t)(N * N) + i1;
#pragma unroll
                for (int i0 = 0; i0 < N; ++i0)
                {
                    double b = 0.0;
#pragma unroll
                    for (int q0 = 0; q0 < M; ++q0)
                        b = fma(prm.Bw0[i0][q0], sv[q0], b);
                    __stcs(o + i0 * N, b);
                }
            }
        }

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
                if (issuer)
                    lo_bulk_wait_read<NBUF - 1>();
                __syncwarp();
            }
            const int j0s = pj * S::PJ;
            const int rows = (S::PJ < N - j0s) ? S::PJ : N - j0s;
            const size_t gstart = e * (size_t)S::NN + (size_t)j0s * S::ROWB;
            const int par = TILE ? 0 : (int)(gstart & 1);
            double *dst = stage + ((size_t)buf * S::EPW + (lane_active ? el : 0)) * prm.cap + par;
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
            if (TILE)
            {
                if (lane == 0)
                {
                    // the tile's valid elements are one contiguous, 16-byte aligned run of global memory
                    const size_t e0 = tile * S::EPW;
                    const int nv = (int)((E - e0) < (size_t)S::EPW ? (E - e0) : (size_t)S::EPW);
                    const int cnt = nv * S::NN;
                    double *g = out + e0 * (size_t)S::NN;
                    const double *sp = stage + (size_t)buf * S::EPW * prm.cap;
                    const int body = cnt & ~1;
                    if (body > 0)
                        lo_bulk_s2g(g, sp, (unsigned)body * 8u);
                    if (body < cnt)
                        g[cnt - 1] = sp[cnt - 1];
                    lo_bulk_commit();
                }
            }
            else if (leader)
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
        tile = tile_next;
        tile_end = tile_next_end;
    }
    if (issuer)
        lo_bulk_wait_all();
    if (lane == 0)
        sched_finish_one(sched, n_warps);
}

struct LoPlan
{
    bool built = false;
    double *d_P1 = nullptr; // [SLOTS][M]: row of the pair assigned to the slot (zeros for empty slots)
    int2 *d_offs = nullptr; // [SLOTS]: staging offsets (entry, mirror) of the slot's pair, -1 = none
    int cap = 0;            // staged doubles per element and panel
    double cost = 0.0;      // shared-memory wavefronts per staging store of the chosen assignment (diagnostic)
    ~LoPlan()
    {
        if (d_P1)
            cudaFree(d_P1);
        if (d_offs)
            cudaFree(d_offs);
    }
};

// Which thread stores which column pair decides the shared-memory bank conflicts of the staging stores: a warp stores,
// for one (j0,i0) and one r, the entries of pairs slot[bg + r*TPE] of EPW elements that sit `cap` doubles apart (plus
// the 16-byte phase of the element's global address).  64-bit shared stores are served per half-warp, one wavefront
// per distinct set of 8-byte banks (16 of them).  ncu on the natural ordering showed 2x the ideal wavefront count
// (l1tex 74 % busy, the kernel's limiter), so the assignment of pairs to slots and the even skew added to the element
// stride are chosen here by a small deterministic local search that minimises the wavefront count of the four store
// kinds (entry / mirror).  Returns wavefronts per store instruction (ideal 2.0 with all lanes busy).
double assign_slots(int N, int R, int TPE, int EPW, int cap0, bool tile, std::vector<int> &slot_pair, int &cap_out)
{
    const int NB = N * (N + 1) / 2, SLOTS = R * TPE, NN = N * N * N * N;
    std::vector<int> pj(NB), pi(NB);
    for (int j1 = 0, b = 0; j1 < N; ++j1)
        for (int i1 = 0; i1 <= j1; ++i1, ++b)
            pj[b] = j1, pi[b] = i1;
    auto cost_of = [&](const std::vector<int> &sp, int cap) {
        long total = 0, insts = 0;
        for (int tp = 0; tp < (!tile && (NN & 1) && (EPW & 1) ? 2 : 1); ++tp) // tile parity matters for odd NN and odd EPW
            for (int r = 0; r < R; ++r)
                for (int kind = 0; kind < 2; ++kind)
                {
                    bool any = false;
                    for (int half = 0; half < 2; ++half)
                    {
                        int mult[16] = {0}, worst = 0;
                        for (int lane = half * 16; lane < half * 16 + 16; ++lane)
                        {
                            const int el = lane / TPE, bg = lane % TPE;
                            if (el >= EPW)
                                continue;
                            const int b = sp[bg + r * TPE];
                            if (b < 0 || (kind == 1 && pj[b] == pi[b]))
                                continue;
                            const int off = kind == 0 ? pj[b] * N * N + pi[b] : pi[b] * N * N + pj[b];
                            const int par = tile ? 0 : ((tp * EPW + el) * NN) & 1;
                            const int bank = (el * cap + par + off) & 15;
                            worst = std::max(worst, ++mult[bank]);
                        }
                        total += worst;
                        any = any || worst > 0;
                    }
                    insts += any ? 1 : 0;
                }
        return insts ? (double)total / (double)insts : 0.0;
    };
    std::vector<int> best_sp;
    double best = 1e30;
    int best_cap = cap0;
    unsigned long long rng = 0x9E3779B97F4A7C15ull;
    auto next = [&]() {
        rng = rng * 6364136223846793005ull + 1442695040888963407ull;
        return (unsigned)(rng >> 33);
    };
    for (int skew = 0; skew < (tile ? 1 : 16); skew += 2)
        for (int restart = 0; restart < 6; ++restart)
        {
            std::vector<int> sp(SLOTS, -1);
            for (int b = 0; b < NB; ++b)
                sp[b] = b;
            for (int k = SLOTS - 1; k > 0 && restart > 0; --k)
                std::swap(sp[k], sp[next() % (k + 1)]);
            double cur = cost_of(sp, cap0 + skew);
            for (bool improved = true; improved;)
            {
                improved = false;
                for (int a = 0; a < SLOTS; ++a)
                    for (int b = a + 1; b < SLOTS; ++b)
                    {
                        if (sp[a] == sp[b])
                            continue;
                        std::swap(sp[a], sp[b]);
                        const double c = cost_of(sp, cap0 + skew);
                        if (c < cur - 1e-12)
                            cur = c, improved = true;
                        else
                            std::swap(sp[a], sp[b]);
                    }
            }
            if (cur < best - 1e-12)
                best = cur, best_sp = sp, best_cap = cap0 + skew;
        }
    slot_pair = best_sp;
    cap_out = best_cap;
    return best;
}

template <int N, int M, int R, bool TILE>
int launch_lo(const LaunchCfg &cfg, LoPlan &plan, const OpHost &host, int wsrc_kind, const OpHost *geom_host, size_t E,
              const double *wsrc, double *out, const double *fvals, double *rhs)
{
    using S = LoShape<N, M, R, TILE>;
    if (!plan.built)
    {
        std::vector<int> slot_pair;
        plan.cost = assign_slots(N, R, S::TPE, S::EPW, S::CAP0, TILE, slot_pair, plan.cap);
        std::vector<double> P1((size_t)S::SLOTS * M, 0.0);
        std::vector<int2> offs(S::SLOTS, make_int2(-1, -1));
        for (int sl = 0; sl < S::SLOTS; ++sl)
        {
            const int b = slot_pair[sl];
            if (b < 0)
                continue;
            int j1 = 0;
            while ((j1 + 1) * (j1 + 2) / 2 <= b)
                ++j1;
            const int i1 = b - j1 * (j1 + 1) / 2;
            offs[sl] = make_int2(j1 * N * N + i1, i1 != j1 ? i1 * N * N + j1 : -1);
            // the product of the two basis values first -> exactly symmetric in (j1,i1), same table as kernels_mass2d.cu
            for (int q = 0; q < M; ++q)
                P1[(size_t)sl * M + q] = host.w[1][q] * (host.Bout[1][j1 * M + q] * host.Bin[1][i1 * M + q]);
        }
        IB_CUDA(cudaMalloc((void **)&plan.d_P1, P1.size() * sizeof(double)));
        IB_CUDA(cudaMemcpy(plan.d_P1, P1.data(), P1.size() * sizeof(double), cudaMemcpyHostToDevice));
        IB_CUDA(cudaMalloc((void **)&plan.d_offs, offs.size() * sizeof(int2)));
        IB_CUDA(cudaMemcpy(plan.d_offs, offs.data(), offs.size() * sizeof(int2), cudaMemcpyHostToDevice));
        if (std::getenv("IB200_LO_VERBOSE"))
            fprintf(stderr, "mass2d_lo<%d,%d,%d,%s>: cap %d (+%d), %.2f wavefronts per staging store\n", N, M, R,
                    TILE ? "tile" : "panel", plan.cap, plan.cap - S::CAP0, plan.cost);
        plan.built = true;
    }
    static_assert(sizeof(LoParams<N, M>) <= 16 * 1024, "parameter block too large");
    LoParams<N, M> prm;
    std::memset(&prm, 0, sizeof prm);
    // the product of the two basis values first -> exactly symmetric in (j0,i0), same table as kernels_mass2d.cu
    for (int hi = 0; hi < N; ++hi)
        for (int lo = 0; lo <= hi; ++lo)
            for (int q = 0; q < M; ++q)
                prm.P0[hi * (hi + 1) / 2 + lo][q] = host.w[0][q] * (host.Bout[0][hi * M + q] * host.Bin[0][lo * M + q]);
    for (int i = 0; i < N; ++i)
        for (int q = 0; q < M; ++q)
        {
            prm.Bw0[i][q] = host.Bin[0][i * M + q] * host.w[0][q];
            prm.Bw1[i][q] = host.Bin[1][i * M + q] * host.w[1][q];
        }
    prm.wsrc = wsrc_kind;
    prm.aux_per = wsrc_kind == 2 ? 8 : (wsrc_kind == 1 ? S::Q : 0);
    prm.in_doubles = (S::EPW * (prm.aux_per + (rhs ? S::Q : 0)) + 1) & ~1;
    prm.cap = plan.cap;
    const size_t smem_bytes =
        sizeof(double) * (8 * S::Q + (size_t)S::WARPS * (S::EPW * S::Qp + prm.in_doubles + S::NBUF * S::EPW * prm.cap));
    if (wsrc_kind == 2)
    {
        std::memcpy(prm.gB0, geom_host->Bin[0], sizeof(double) * 2 * M);
        std::memcpy(prm.gD0, geom_host->Din[0], sizeof(double) * 2 * M);
        std::memcpy(prm.gB1, geom_host->Bin[1], sizeof(double) * 2 * M);
        std::memcpy(prm.gD1, geom_host->Din[1], sizeof(double) * 2 * M);
    }
    auto kernel = mass2d_lo_kernel<N, M, R, TILE>;
    IB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    int per_sm = 0;
    IB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, S::WARPS * 32, smem_bytes));
    if (per_sm < 1)
        return IB200_ERROR_UNSUPPORTED;
    const size_t n_tiles = (E + S::EPW - 1) / S::EPW;
    const size_t want = (n_tiles + S::WARPS * IB200_LO_SCHED_CHUNK - 1) / (S::WARPS * IB200_LO_SCHED_CHUNK);
    const size_t cap = (size_t)cfg.sm_count * per_sm;
    const int grid = (int)std::max<size_t>(1, std::min(want, cap));
    kernel<<<grid, S::WARPS * 32, smem_bytes, cfg.stream>>>(prm, plan.d_P1, plan.d_offs, E, wsrc, out, fvals, rhs, cfg.sched);
    ++*cfg.launch_counter;
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess)
        return ib200_fail(IB200_ERROR_CUDA, "launch of mass2d_lo_kernel<%d,%d,%d> failed: %s", N, M, R, cudaGetErrorString(err));
    return INTERP_SUCCESS;
}

} // namespace

// Low-order symmetric path: both dimensions with matching spaces of the same basis count, no derivative.  Same contract as
// launch_mass2d_fast; IB200_ERROR_UNSUPPORTED (no error text) for every other shape.  With fvals/rhs non-null the L2
// right-hand sides rhs[E][N*N] of the sampled values fvals[E][Q] are produced by the same launch (fused K3).
int launch_mass2d_lo(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, const OpDesc &op, const OpHost &host,
                     bool match0, bool match1, int deriv_dim_in, int wsrc_kind, const OpHost *geom_host, int widx_in, size_t E,
                     const double *wsrc, double *out, const double *fvals, double *rhs)
{
    if (op.ndim != 2 || !match0 || !match1 || deriv_dim_in >= 0 || widx_in >= 0)
        return IB200_ERROR_UNSUPPORTED;
    if (op.dim[0].m != op.dim[1].m || op.dim[0].n_in != op.dim[1].n_in)
        return IB200_ERROR_UNSUPPORTED;
    if ((reinterpret_cast<size_t>(out) & 15) != 0 || (wsrc_kind == 2 && !geom_host) || ((fvals == nullptr) != (rhs == nullptr)))
        return IB200_ERROR_UNSUPPORTED;
    const int n = op.dim[0].n_in, m = op.dim[0].m;
    if (n < 2 || n > 6 || m > 9)
        return IB200_ERROR_UNSUPPORTED;
    const std::string k = "lo:" + key;
    auto it = ctx->plans.find(k);
    if (it == ctx->plans.end())
    {
        it = ctx->plans.emplace(k, std::static_pointer_cast<void>(std::make_shared<LoPlan>())).first; // filled at first launch
    }
    LoPlan &plan = *static_cast<LoPlan *>(it->second.get());
#define IB_LO(NV, MV, RV) \
    if (n == NV && m == MV) \
        return launch_lo<NV, MV, RV, (NV % 2 == 1)>(cfg, plan, host, wsrc_kind, geom_host, E, wsrc, out, fvals, rhs);
    // (basis count, points, column pairs per thread)
    IB_LO(2, 2, 3) IB_LO(2, 3, 3) IB_LO(2, 4, 3)
    IB_LO(3, 3, 3) IB_LO(3, 4, 3) IB_LO(3, 5, 3)
    IB_LO(4, 4, 3) IB_LO(4, 5, 3) IB_LO(4, 6, 3)
    IB_LO(5, 5, 2) IB_LO(5, 6, 2) IB_LO(5, 7, 2)
    IB_LO(6, 6, 3) IB_LO(6, 7, 3) IB_LO(6, 8, 3)
#undef IB_LO
    return IB200_ERROR_UNSUPPORTED;
}
