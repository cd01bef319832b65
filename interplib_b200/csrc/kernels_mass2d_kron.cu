// K2-kron -- element mass matrices of bilinear quads as a sum of three Kronecker products (SURVEY.md 8a rows A7 + A9).
//
// On a bilinear quad det J_e(xi0, xi1) = alpha_e + beta_e xi0 + gamma_e xi1 exactly (the xi0 xi1 terms of the 2x2
// determinant cancel), so the quadrature sum the reference evaluates per entry (/root/reference/src/python/
// mass_matrices.c:264-335, weights from /root/reference/src/python/mappings.c:448-542) separates:
//
//   M_e[(j0,j1)][(i0,i1)] = sum_q0 sum_q1 w0 w1 (alpha + beta x_q0 + gamma x_q1) Bo0_j0 Bi0_i0 Bo1_j1 Bi1_i1
//                         = (alpha A0[j0][i0] + beta X0[j0][i0]) A1[j1][i1] + (gamma A0[j0][i0]) X1[j1][i1]
//
// with the element-independent 1-D moment tables A_k = sum_q w Bo Bi, X_k = sum_q w x Bo Bi (host, long double:
// fastdiag.c).  Two FP64 operations per entry instead of the m (quadrature points per dimension) FMAs of the
// sum-factorised stage 2 -- 18 at BASELINE config 4, 10 at config 3 -- which turns the high-order mass kernels from
// FP64/latency co-limited (0.76 and 0.68 of the HBM bound with DMMA / DFMA stage 2, profiles/r01e_c4_*, r01e_c3_*) into
// plain streaming writers.  Row groups x column groups cover block matrices of several tensor-product spaces (the 2-D
// 1-form basis of config 3: one launch writes the whole [144][144] element matrix, block (a,b) being
// compute_mass_matrix(space_b, space_a)); a single group each way is compute_mass_matrix itself, mixed spaces included.
// Matching spaces give exactly symmetric matrices: M[J][I] and M[I][J] are the same two operations on the same operands.
//
// Mapping: one CTA per element, thread I owns COLUMN I of the element matrix for the whole launch -- its 1-D factors
// A1[.][i1], X1[.][i1] stay in registers, (A0, X0)[j0][i0] come as one 16-byte L1-resident load per j0 -- and walks down
// the rows: every store instruction of a warp writes 256 consecutive bytes of one row (st.global.cs).  No shared-memory
// staging, no barrier inside an element; elements are claimed dynamically one ahead and the next element's corner
// coordinates are requested before the current one is written.
#include <algorithm>
#include <cstdlib>
#include <memory>
#include <string>
#include <vector>

#include "ctx.cuh"

namespace
{

template <int NG>
struct KronParams
{
    int n_out0[NG], n_out1[NG]; // basis counts of the row groups
    int N_in;                   // columns = sum over column groups of n_in0 * n_in1
    unsigned long long NN;      // doubles per element matrix
    int wmode;                  // weight: 0 = det J; 1..4 = det J * Jinv[idx_in][idx_out] for (0,0), (0,1), (1,0), (1,1)
    double G[3][4];             // (x_xi0, x_xi1, x_xi0xi1) = G (c00, c01, c10, c11)
    double Gs[3];               // row sums of G: zero for the geometry families whose equal DoFs mean a constant
};

__device__ __forceinline__ double kron_diff_of_products(double a, double b, double c, double d)
{
    const double p = c * d;
    const double err = fma(-c, d, p);
    return fma(a, b, -p) + err;
}

__device__ __forceinline__ unsigned kron_smem_u32(const void *p)
{
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void kron_bulk_s2g(double *gdst, const double *ssrc, unsigned bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(kron_smem_u32(ssrc)), "r"(bytes)
                 : "memory");
}

constexpr int kron_max_threads(int NG, int NO)
{
    const int t = (NG * NO * NO + 31) / 32 * 32;
    return t > 1024 ? 1024 : t;
}

// tab0: [NG][NO][N_in] of {A0, X0}[j0][i0(I)] (zero beyond the group's n_out0); tab1: [N_in][NG][NO] of {A1, X1}[j1][i1(I)]
//
// STAGED = false: rows are whole 32-byte sectors (N_in % 4 == 0), every warp store is 8 full sectors: 7.1-7.3 TB/s.
// STAGED = true:  N_in % 4 != 0 -- a direct warp store would straddle 9 sectors with partial ones at both ends, and the
//                 write stream falls to 0.63-0.85 of the copy bandwidth at one CTA per SM (tools/kron_sweep.py,
//                 profiles/r02h_kron_sweep.jsonl).  The n_out1 rows of one (group, j0) -- a contiguous run of global memory --
//                 are staged in shared memory (lane = column: conflict-free) with the 16-byte phase of their global address
//                 and leave as ONE cp.async.bulk shared->global copy; three buffers, one CTA barrier per run.
//
// Gradient matrices (compute_gradient_mass_matrix, mass_matrices.c:341-565): the weight det J * Jinv[idx_in][idx_out] is a
// cofactor of J -- for a bilinear map linear in ONE reference coordinate, a + b xi0 (idx_in = 0) or a + b xi1 (idx_in = 1) --
// so the same two operations per entry apply with (alpha, beta, gamma) = (a, b, 0) or (a, 0, b) and the moment tables of the
// differentiated dimension built from the derivative of the input basis.  MIRROR reproduces the reference's treatment of
// matching spaces: it evaluates the entries on and below the diagonal only and copies them into the upper triangle
// (mass_matrices.c:497-560) although a gradient matrix is not symmetric; thread I then also holds the transposed factors
// and writes f(row I, column J) into the rows J < I of its column.
template <int NG, int NO, bool STAGED, bool MIRROR>
__global__ void __launch_bounds__(kron_max_threads(NG, NO))
    mass2d_kron_kernel(const __grid_constant__ KronParams<NG> prm, const double2 *__restrict__ tab0, const double2 *__restrict__ tab1,
                       const double2 *__restrict__ tab0T, const double2 *__restrict__ tab1T, size_t E, const double *__restrict__ geom,
                       double *__restrict__ out, unsigned long long *__restrict__ sched)
{
    static_assert(!MIRROR || NG == 1, "the mirrored variant serves one space");
    const int I = threadIdx.x;
    const bool active = I < prm.N_in;
    const int col = active ? I : 0;

    double a1[NG][NO], x1[NG][NO], a1T[MIRROR ? NO : 1], x1T[MIRROR ? NO : 1];
#pragma unroll
    for (int g = 0; g < NG; ++g)
#pragma unroll
        for (int j1 = 0; j1 < NO; ++j1)
        {
            const double2 v = tab1[((size_t)col * NG + g) * NO + j1];
            a1[g][j1] = v.x;
            x1[g][j1] = v.y;
        }
    if (MIRROR)
    {
#pragma unroll
        for (int j1 = 0; j1 < NO; ++j1)
        {
            const double2 v = tab1T[(size_t)col * NO + j1];
            a1T[MIRROR ? j1 : 0] = v.x;
            x1T[MIRROR ? j1 : 0] = v.y;
        }
    }

    __shared__ size_t claim[2];
    size_t e = blockIdx.x, e_next = E;
    unsigned long long pend = 0;
    if (threadIdx.x == 0)
    {
        claim[0] = gridDim.x + atomicAdd(sched, 1ull);
        pend = atomicAdd(sched, 1ull);
    }
    __syncthreads();
    e_next = claim[0];
    double cc[8], cn[8];
#pragma unroll
    for (int k = 0; k < 8; ++k)
        cc[k] = e < E ? __ldg(geom + e * 8 + k) : 0.0;

    unsigned run = 0; // staged runs issued by this CTA
    for (int it = 0; e < E; ++it)
    {
#pragma unroll
        for (int k = 0; k < 8; ++k)
            cn[k] = e_next < E ? __ldg(geom + e_next * 8 + k) : 0.0;
        if (threadIdx.x == 0)
        {
            const size_t nx = gridDim.x + pend;
            claim[(it + 1) & 1] = nx;
            if (nx < E)
                pend = atomicAdd(sched, 1ull);
        }
        // det J = alpha + beta xi0 + gamma xi1 from the corner coordinates [c][i0][i1]
        // G c = G (c - c0) + (row sum of G) c0: for Lagrange / Bernstein geometry DoFs the row sums vanish (equal DoFs are a
        // constant, and every row is a derivative), the differences from corner 0 are exact for neighbouring coordinates,
        // and the coefficients carry eps * h instead of eps * |x|
        double dx[3], dy[3];
#pragma unroll
        for (int k = 0; k < 3; ++k)
        {
            double ax = __dmul_rn(prm.Gs[k], cc[0]), ay = __dmul_rn(prm.Gs[k], cc[4]);
#pragma unroll
            for (int c = 1; c < 4; ++c)
            {
                ax = fma(prm.G[k][c], cc[c] - cc[0], ax);
                ay = fma(prm.G[k][c], cc[4 + c] - cc[4], ay);
            }
            dx[k] = ax;
            dy[k] = ay;
        }
        double alpha, beta, gamma;
        if (prm.wmode == 0)
        {
            alpha = kron_diff_of_products(dx[0], dy[1], dx[1], dy[0]);
            beta = kron_diff_of_products(dx[0], dy[2], dx[2], dy[0]);
            gamma = kron_diff_of_products(dx[2], dy[1], dx[1], dy[2]);
        }
        else
        {
            // det J * Jinv = [[y_xi1, -x_xi1], [-y_xi0, x_xi0]]; x_xi0 = dx0 + dx2 xi1, x_xi1 = dx1 + dx2 xi0 (same for y)
            const int w = prm.wmode;
            alpha = w == 1 ? dy[1] : (w == 2 ? -dx[1] : (w == 3 ? -dy[0] : dx[0]));
            const double b = w == 1 ? dy[2] : (w == 2 ? -dx[2] : (w == 3 ? -dy[2] : dx[2]));
            beta = w <= 2 ? b : 0.0;
            gamma = w <= 2 ? 0.0 : b;
        }

        if (!STAGED)
        {
            if (active)
            {
                double *p = out + e * prm.NN + I;
#pragma unroll
                for (int g = 0; g < NG; ++g)
                {
                    const int n0 = prm.n_out0[g], n1 = prm.n_out1[g];
                    const double2 *t0 = tab0 + (size_t)g * NO * prm.N_in + I;
                    const double2 *t0T = MIRROR ? tab0T + I : t0;
                    double2 ax_next = t0[0], axT_next = t0T[0];
                    int J = 0; // row (MIRROR: rows above the diagonal of column I take the transposed factors)
#pragma unroll
                    for (int j0 = 0; j0 < NO; ++j0)
                    {
                        if (j0 < n0)
                        {
                            // the pair of the next j0 is requested before this j0's rows are written (the table has NO rows)
                            const double2 ax = ax_next, axT = axT_next;
                            if (j0 + 1 < NO)
                            {
                                ax_next = t0[(size_t)(j0 + 1) * prm.N_in];
                                if (MIRROR)
                                    axT_next = t0T[(size_t)(j0 + 1) * prm.N_in];
                            }
                            const double c1 = fma(beta, ax.y, __dmul_rn(alpha, ax.x));
                            const double c2 = __dmul_rn(gamma, ax.x);
                            const double c1T = fma(beta, axT.y, __dmul_rn(alpha, axT.x));
                            const double c2T = __dmul_rn(gamma, axT.x);
#pragma unroll
                            for (int j1 = 0; j1 < NO; ++j1)
                            {
                                if (j1 < n1)
                                {
                                    double v = fma(c1, a1[g][j1], __dmul_rn(c2, x1[g][j1]));
                                    if (MIRROR)
                                    {
                                        const double vT = fma(c1T, a1T[MIRROR ? j1 : 0], __dmul_rn(c2T, x1T[MIRROR ? j1 : 0]));
                                        v = J >= I ? v : vT;
                                        ++J;
                                    }
                                    __stcs(p, v);
                                    p += prm.N_in;
                                }
                            }
                        }
                    }
                }
            }
        }
        else
        {
            extern __shared__ __align__(16) double kron_stage[];
            const int cap = (NO * prm.N_in + 3) & ~1; // doubles per buffer: even, room for the phase double
            size_t gs = e * prm.NN;            // global index of the run's first double
#pragma unroll
            for (int g = 0; g < NG; ++g)
            {
                const int n0 = prm.n_out0[g], n1 = prm.n_out1[g];
                const double2 *t0 = tab0 + (size_t)g * NO * prm.N_in + col;
                const double2 *t0T = MIRROR ? tab0T + col : t0;
                double2 ax_next = t0[0], axT_next = t0T[0];
                int J = 0;
#pragma unroll
                for (int j0 = 0; j0 < NO; ++j0)
                {
                    if (j0 < n0)
                    {
                        const double2 ax = ax_next, axT = axT_next;
                        if (j0 + 1 < NO)
                        {
                            ax_next = t0[(size_t)(j0 + 1) * prm.N_in];
                            if (MIRROR)
                                axT_next = t0T[(size_t)(j0 + 1) * prm.N_in];
                        }
                        const double c1 = fma(beta, ax.y, __dmul_rn(alpha, ax.x));
                        const double c2 = __dmul_rn(gamma, ax.x);
                        const double c1T = fma(beta, axT.y, __dmul_rn(alpha, axT.x));
                        const double c2T = __dmul_rn(gamma, axT.x);
                        const int par = (int)(gs & 1);
                        double *sm = kron_stage + (run % 3u) * cap + par;
                        if (active)
                        {
                            double *d = sm + I;
#pragma unroll
                            for (int j1 = 0; j1 < NO; ++j1)
                            {
                                if (j1 < n1)
                                {
                                    double v = fma(c1, a1[g][j1], __dmul_rn(c2, x1[g][j1]));
                                    if (MIRROR)
                                    {
                                        const double vT = fma(c1T, a1T[MIRROR ? j1 : 0], __dmul_rn(c2T, x1T[MIRROR ? j1 : 0]));
                                        v = (J + j1) >= I ? v : vT;
                                    }
                                    *d = v;
                                    d += prm.N_in;
                                }
                            }
                        }
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        __syncthreads();
                        const int cnt = n1 * prm.N_in;
                        if (threadIdx.x == 0)
                        {
                            // sm[k] <-> out[gs + k]; the bulk copy needs 16-byte aligned ends: at most one head and one tail double
                            double *gp = out + gs;
                            int k0 = 0;
                            if (par)
                            {
                                gp[0] = sm[0];
                                k0 = 1;
                            }
                            const int body = (cnt - k0) & ~1;
                            if (body > 0)
                                kron_bulk_s2g(gp + k0, sm + k0, (unsigned)body * 8u);
                            if (k0 + body < cnt)
                                gp[cnt - 1] = sm[cnt - 1];
                            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                            // the buffer the NEXT run writes was read by the copy issued two runs ago: done once at most one
                            // group (this one) is still reading
                            asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                        }
                        gs += cnt;
                        ++run;
                        J += n1;
                    }
                }
            }
        }
        __syncthreads(); // claim[(it + 1) & 1] is visible; nobody still reads claim[it & 1] (read before the previous barrier)
        e = e_next;
        e_next = claim[(it + 1) & 1];
#pragma unroll
        for (int k = 0; k < 8; ++k)
            cc[k] = cn[k];
    }
    if (STAGED && threadIdx.x == 0)
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    sched_finish(sched);
}

struct KronPlan
{
    bool usable = false;
    int NG = 0, NO = 0, N_in = 0, N_out = 0;
    int n_out0[2] = {0, 0}, n_out1[2] = {0, 0};
    double G[3][4] = {};
    int wmode = 0;
    bool mirror = false;
    char *d_tab = nullptr;
    const double2 *d_tab0 = nullptr, *d_tab1 = nullptr, *d_tab0T = nullptr, *d_tab1T = nullptr;
    ~KronPlan()
    {
        if (d_tab)
            cudaFree(d_tab);
    }
};

template <int NG, int NO, bool STAGED, bool MIRROR>
int launch_kron(const LaunchCfg &cfg, const KronPlan &plan, size_t E, const double *geom, double *out)
{
    KronParams<NG> prm;
    std::memset(&prm, 0, sizeof prm);
    for (int g = 0; g < NG; ++g)
    {
        prm.n_out0[g] = plan.n_out0[g];
        prm.n_out1[g] = plan.n_out1[g];
    }
    prm.N_in = plan.N_in;
    prm.NN = (unsigned long long)plan.N_out * plan.N_in;
    prm.wmode = plan.wmode;
    std::memcpy(prm.G, plan.G, sizeof prm.G);
    for (int k = 0; k < 3; ++k)
        prm.Gs[k] = plan.G[k][0] + plan.G[k][1] + plan.G[k][2] + plan.G[k][3];
    const int threads = (plan.N_in + 31) / 32 * 32;
    if (threads > kron_max_threads(NG, NO))
        return IB200_ERROR_UNSUPPORTED;
    constexpr bool staged = STAGED;
    auto kernel = mass2d_kron_kernel<NG, NO, STAGED, MIRROR>;
    const size_t smem_bytes = staged ? sizeof(double) * 3 * (((size_t)NO * plan.N_in + 3) & ~(size_t)1) : 0;
    if (smem_bytes > 224 * 1024)
        return IB200_ERROR_UNSUPPORTED;
    if (smem_bytes)
        IB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    int per_sm = 0;
    IB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem_bytes));
    if (per_sm < 1)
        return IB200_ERROR_UNSUPPORTED;
    const size_t cap = (size_t)cfg.sm_count * per_sm;
    const int grid = (int)std::max<size_t>(1, std::min(E, cap));
    kernel<<<grid, threads, smem_bytes, cfg.stream>>>(prm, plan.d_tab0, plan.d_tab1, plan.d_tab0T, plan.d_tab1T, E, geom, out, cfg.sched);
    ++*cfg.launch_counter;
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess)
        return ib200_fail(IB200_ERROR_CUDA, "launch of mass2d_kron_kernel<%d,%d> failed: %s", NG, NO, cudaGetErrorString(err));
    return INTERP_SUCCESS;
}

// smallest instantiated table height that holds `need` basis functions per dimension (0: none)
int kron_pick(int NG, int need)
{
    static const int one[] = {3, 5, 7, 9, 11, 13, 17, 21}, two[] = {3, 5, 7, 9, 11};
    if (NG == 1)
    {
        for (int v : one)
            if (v >= need)
                return v;
    }
    else
        for (int v : two)
            if (v >= need)
                return v;
    return 0;
}

} // namespace

// Row groups `rows[g][2]` (output spaces) x column groups `cols[b][2]` (input spaces) of 2-D tensor-product spaces under
// `rules[2]`, bilinear geometry `geom_basis[2]` (order 1): out[E][N_out][N_in] with N_out = sum_g n0 n1 of the row groups,
// N_in likewise; block (g, b) = compute_mass_matrix(cols[b], rows[g], smap) with the determinant weight.
// deriv_in >= 0: compute_gradient_mass_matrix(cols[0], rows[0], smap, deriv_in, idx_out) instead (one group each way; `key`
// must tell the variants apart); `mirror`: the spaces match and the reference's lower-triangle mirroring is reproduced.
// IB200_ERROR_UNSUPPORTED (no error text) outside the envelope.  E == 0 with out == NULL only answers that question.
int launch_mass2d_kron(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, int n_row_groups, const basis_spec_t *rows,
                       int n_col_groups, const basis_spec_t *cols, const integration_spec_t *rules, const basis_spec_t *geom_basis,
                       size_t E, const double *geom, double *out, int deriv_in, int idx_out, bool mirror)
{
    if (n_row_groups < 1 || n_row_groups > 2 || n_col_groups < 1 || n_col_groups > 2)
        return IB200_ERROR_UNSUPPORTED;
    if (deriv_in >= 0 && (n_row_groups != 1 || n_col_groups != 1 || deriv_in > 1 || idx_out < 0 || idx_out > 1))
        return IB200_ERROR_UNSUPPORTED;
    if (mirror && deriv_in < 0)
        mirror = false; // plain mass matrices of matching spaces are symmetric as they are
    if (geom_basis[0].order != 1 || geom_basis[1].order != 1)
        return IB200_ERROR_UNSUPPORTED;
    auto it = ctx->plans.find(key);
    if (it == ctx->plans.end())
    {
        auto plan = std::make_shared<KronPlan>();
        int need = 0, N_in = 0, N_out = 0;
        for (int g = 0; g < n_row_groups; ++g)
        {
            need = std::max(need, (int)std::max(rows[2 * g].order, rows[2 * g + 1].order) + 1);
            N_out += (int)(rows[2 * g].order + 1) * (int)(rows[2 * g + 1].order + 1);
        }
        for (int b = 0; b < n_col_groups; ++b)
            N_in += (int)(cols[2 * b].order + 1) * (int)(cols[2 * b + 1].order + 1);
        const int NG = n_row_groups, NO = kron_pick(NG, need);
        double a0[2], s0[2], a1[2], s1[2];
        int res = NO ? (int)INTERP_SUCCESS : (int)IB200_ERROR_UNSUPPORTED;
        if (res == INTERP_SUCCESS && (N_in + 31) / 32 * 32 > kron_max_threads(NG, NO))
            res = IB200_ERROR_UNSUPPORTED;
        if (res == INTERP_SUCCESS)
            res = ib200_tables_linear_basis(geom_basis[0], a0, s0);
        if (res == INTERP_SUCCESS)
            res = ib200_tables_linear_basis(geom_basis[1], a1, s1);
        if (mirror && 2 * (4 * NO + 20) * ((N_in + 31) / 32 * 32) > 65536)
            res = IB200_ERROR_UNSUPPORTED; // the transposed factors double the register tables: does not fit one CTA
        std::vector<double2> tab0((size_t)NG * std::max(NO, 1) * N_in, make_double2(0.0, 0.0)),
            tab1((size_t)N_in * NG * std::max(NO, 1), make_double2(0.0, 0.0)), tab0T(mirror ? tab0.size() : 0, make_double2(0.0, 0.0)),
            tab1T(mirror ? tab1.size() : 0, make_double2(0.0, 0.0));
        for (int g = 0; g < NG && res == INTERP_SUCCESS; ++g)
        {
            const int no0 = (int)rows[2 * g].order + 1, no1 = (int)rows[2 * g + 1].order + 1;
            plan->n_out0[g] = no0;
            plan->n_out1[g] = no1;
            int col0 = 0;
            for (int b = 0; b < n_col_groups && res == INTERP_SUCCESS; ++b)
            {
                const int ni0 = (int)cols[2 * b].order + 1, ni1 = (int)cols[2 * b + 1].order + 1;
                std::vector<double> A0((size_t)no0 * ni0), X0(A0.size()), A1((size_t)no1 * ni1), X1(A1.size());
                res = (deriv_in == 0 ? ib200_tables_pair_moments_deriv : ib200_tables_pair_moments)(rows[2 * g], cols[2 * b], rules[0],
                                                                                                      A0.data(), X0.data());
                if (res == INTERP_SUCCESS)
                    res = (deriv_in == 1 ? ib200_tables_pair_moments_deriv : ib200_tables_pair_moments)(rows[2 * g + 1], cols[2 * b + 1],
                                                                                                          rules[1], A1.data(), X1.data());
                if (res != INTERP_SUCCESS)
                    break;
                for (int i0 = 0; i0 < ni0; ++i0)
                    for (int i1 = 0; i1 < ni1; ++i1)
                    {
                        const int I = col0 + i0 * ni1 + i1;
                        for (int j0 = 0; j0 < no0; ++j0)
                            tab0[((size_t)g * NO + j0) * N_in + I] = make_double2(A0[(size_t)j0 * ni0 + i0], X0[(size_t)j0 * ni0 + i0]);
                        for (int j1 = 0; j1 < no1; ++j1)
                            tab1[((size_t)I * NG + g) * NO + j1] = make_double2(A1[(size_t)j1 * ni1 + i1], X1[(size_t)j1 * ni1 + i1]);
                        if (mirror) // matching spaces (no == ni): factors of f(out = I, in = J), the thread's index on the OUT side
                        {
                            for (int j0 = 0; j0 < no0; ++j0)
                                tab0T[(size_t)j0 * N_in + I] = make_double2(A0[(size_t)i0 * ni0 + j0], X0[(size_t)i0 * ni0 + j0]);
                            for (int j1 = 0; j1 < no1; ++j1)
                                tab1T[(size_t)I * NO + j1] = make_double2(A1[(size_t)i1 * ni1 + j1], X1[(size_t)i1 * ni1 + j1]);
                        }
                    }
                col0 += ni0 * ni1;
            }
        }
        if (res == INTERP_SUCCESS)
        {
            for (int i0 = 0; i0 < 2; ++i0)
                for (int k = 0; k < 2; ++k)
                {
                    plan->G[0][i0 * 2 + k] = s0[i0] * a1[k];
                    plan->G[1][i0 * 2 + k] = a0[i0] * s1[k];
                    plan->G[2][i0 * 2 + k] = s0[i0] * s1[k];
                }
            const size_t b0 = tab0.size() * sizeof(double2), b1 = tab1.size() * sizeof(double2);
            const size_t o1 = (b0 + 255) & ~(size_t)255, o2 = o1 + ((b1 + 255) & ~(size_t)255), o3 = o2 + o1;
            if (cudaMalloc((void **)&plan->d_tab, mirror ? o3 + b1 : o1 + b1) != cudaSuccess ||
                cudaMemcpy(plan->d_tab, tab0.data(), b0, cudaMemcpyHostToDevice) != cudaSuccess ||
                cudaMemcpy(plan->d_tab + o1, tab1.data(), b1, cudaMemcpyHostToDevice) != cudaSuccess ||
                (mirror && (cudaMemcpy(plan->d_tab + o2, tab0T.data(), b0, cudaMemcpyHostToDevice) != cudaSuccess ||
                            cudaMemcpy(plan->d_tab + o3, tab1T.data(), b1, cudaMemcpyHostToDevice) != cudaSuccess)))
                return ib200_fail(IB200_ERROR_CUDA, "upload of the Kronecker mass tables failed");
            plan->d_tab0 = (const double2 *)plan->d_tab;
            plan->d_tab1 = (const double2 *)(plan->d_tab + o1);
            plan->d_tab0T = mirror ? (const double2 *)(plan->d_tab + o2) : plan->d_tab0;
            plan->d_tab1T = mirror ? (const double2 *)(plan->d_tab + o3) : plan->d_tab1;
            plan->wmode = deriv_in < 0 ? 0 : 1 + 2 * deriv_in + idx_out;
            plan->mirror = mirror;
            plan->NG = NG;
            plan->NO = NO;
            plan->N_in = N_in;
            plan->N_out = N_out;
            plan->usable = true;
        }
        else if (res != IB200_ERROR_UNSUPPORTED)
            return res;
        it = ctx->plans.emplace(key, std::static_pointer_cast<void>(plan)).first;
    }
    const KronPlan &plan = *static_cast<const KronPlan *>(it->second.get());
    if (!plan.usable)
        return IB200_ERROR_UNSUPPORTED;
    if (E == 0 && out == nullptr)
        return INTERP_SUCCESS;
    // rows that are not whole 32-byte sectors (or an output that does not start on one) go through the staged variant
    bool staged = plan.N_in % 4 != 0 || (reinterpret_cast<size_t>(out) & 31) != 0;
    if (const char *env = std::getenv("IB200_KRON_STAGED")) // measuring aid: 0 = always direct stores, 1 = always staged
        staged = std::atoi(env) != 0;
#define IB_KRON(G_, O_) \
    if (plan.NG == G_ && plan.NO == O_ && !plan.mirror) \
        return staged ? launch_kron<G_, O_, true, false>(cfg, plan, E, geom, out) : launch_kron<G_, O_, false, false>(cfg, plan, E, geom, out);
#define IB_KRON_M(O_) \
    if (plan.NG == 1 && plan.NO == O_ && plan.mirror) \
        return staged ? launch_kron<1, O_, true, true>(cfg, plan, E, geom, out) : launch_kron<1, O_, false, true>(cfg, plan, E, geom, out);
    IB_KRON(1, 3) IB_KRON(1, 5) IB_KRON(1, 7) IB_KRON(1, 9) IB_KRON(1, 11) IB_KRON(1, 13) IB_KRON(1, 17) IB_KRON(1, 21)
    IB_KRON(2, 3) IB_KRON(2, 5) IB_KRON(2, 7) IB_KRON(2, 9) IB_KRON(2, 11)
    IB_KRON_M(3) IB_KRON_M(5) IB_KRON_M(7) IB_KRON_M(9) IB_KRON_M(11) IB_KRON_M(13) IB_KRON_M(17)
#undef IB_KRON
#undef IB_KRON_M
    return IB200_ERROR_UNSUPPORTED;
}
