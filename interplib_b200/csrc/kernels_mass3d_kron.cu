// K2-kron3 -- element mass matrices of TRILINEAR HEXAHEDRA as a 27-term Kronecker sum (SURVEY.md 8a rows A7 + A9 in 3-D).
//
// A trilinear map x(xi) = sum_ijk c_ijk phi_i(xi0) phi_j(xi1) phi_k(xi2) (order-1 geometry basis per dimension) has partial
// derivatives that are bilinear in the other two coordinates, so its Jacobian determinant is a polynomial of degree 2 in
// each reference coordinate,  det J_e = sum_{a,b,c in 0..2} d_e[a][b][c] xi0^a xi1^b xi2^c,  and the quadrature sum the
// reference evaluates per entry (/root/reference/src/python/mass_matrices.c:264-335 with the weights of
// /root/reference/src/python/mappings.c:448-542) separates into element-independent 1-D moment tables:
//
//   M_e[(j0,j1,j2)][(i0,i1,i2)] = sum_abc d_e[a][b][c] X0[a][j0][i0] X1[b][j1][i1] X2[c][j2][i2],
//   Xk[a][j][i] = sum_q w_q x_q^a Bout_j(x_q) Bin_i(x_q)                       (host, long double: fastdiag.c)
//
// evaluated innermost-last: 27 FMAs per j0, 9 per (j0,j1), 3 per entry -- against the m^3 (3 d + 2) operations per entry of
// the quadrature loop that the generic kernel (one thread per output) restates.  Until round 2 every 3-D shape ran on that
// generic kernel.  Mapping as in kernels_mass2d_kron.cu: one CTA per element, thread I owns column I and walks down the rows,
// a warp's store writes 256 consecutive bytes of one row; rows that are not whole 32-byte sectors are staged in shared
// memory run by run and leave as cp.async.bulk copies.  The 27 coefficients of an element are formed cooperatively: 24
// threads expand the three coordinates into monomial coefficients, 27 threads sum the signed triple products.
#include <algorithm>
#include <cstdlib>
#include <memory>
#include <string>
#include <vector>

#include "ctx.cuh"

namespace
{

__device__ __forceinline__ unsigned k3_smem_u32(const void *p)
{
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void k3_bulk_s2g(double *gdst, const double *ssrc, unsigned bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(k3_smem_u32(ssrc)), "r"(bytes)
                 : "memory");
}

struct Kron3Params
{
    int n_out[3];
    int N_in;
    unsigned long long NN; // doubles per element matrix
    int run_level;         // staged variant: rows per run = 0: whole element, 1: one j0 slab, 2: one (j0, j1) slab
    int run_cap;           // doubles per staging buffer (even, room for the phase double)
    double T[8][8];        // monomial coefficients (1, xi2, xi1, xi1 xi2, xi0, ...) of a coordinate from its 8 geometry DoFs
    double Ts[8];          // row sums of T (zero for m >= 1 with Lagrange / Bernstein geometry DoFs)
};

constexpr int k3_max_threads(int NO)
{
    const int t = (NO * NO * NO + 31) / 32 * 32;
    return t < 32 ? 32 : t;
}

// tab: [N_in][3 dims][3 powers][NO] = Xk[a][j][i_k(I)] (zero beyond the dimension's n_out)
template <int NO, bool STAGED>
__global__ void __launch_bounds__(k3_max_threads(NO))
    mass3d_kron_kernel(const __grid_constant__ Kron3Params prm, const double *__restrict__ tab, size_t E,
                       const double *__restrict__ geom, double *__restrict__ out, unsigned long long *__restrict__ sched)
{
    extern __shared__ __align__(16) double k3_stage[];
    __shared__ double xm[3][8];  // monomial coefficients of x, y, z: index 4 alpha + 2 beta + gamma
    __shared__ double dsm[27];   // d[a][b][c]
    __shared__ size_t claim[2];
    const int I = threadIdx.x;
    const bool active = I < prm.N_in;
    const int col = active ? I : 0;
    const int n0 = prm.n_out[0], n1 = prm.n_out[1], n2 = prm.n_out[2];

    double X0[3][NO], X1[3][NO], X2[3][NO];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int j = 0; j < NO; ++j)
        {
            X0[a][j] = tab[(((size_t)col * 3 + 0) * 3 + a) * NO + j];
            X1[a][j] = tab[(((size_t)col * 3 + 1) * 3 + a) * NO + j];
            X2[a][j] = tab[(((size_t)col * 3 + 2) * 3 + a) * NO + j];
        }

    size_t e = blockIdx.x;
    unsigned long long pend = 0;
    if (threadIdx.x == 0)
    {
        claim[0] = gridDim.x + atomicAdd(sched, 1ull);
        pend = atomicAdd(sched, 1ull);
    }
    unsigned run = 0;
    for (int it = 0; e < E; ++it)
    {
        // ---- the 27 coefficients of det J_e
        if (threadIdx.x < 24)
        {
            const int p = threadIdx.x >> 3, m = threadIdx.x & 7;
            const double *c = geom + (e * 3 + p) * 8;
            // differences from corner 0 are exact for neighbouring coordinates: the coefficients carry eps * h, not eps * |x|
            const double c0 = __ldg(c);
            double acc = __dmul_rn(prm.Ts[m], c0); // T c = T (c - c0) + rowsum * c0
#pragma unroll
            for (int n = 1; n < 8; ++n)
                acc = fma(prm.T[m][n], __ldg(c + n) - c0, acc);
            xm[p][m] = acc;
        }
        if (threadIdx.x == 0)
        {
            const size_t nx = gridDim.x + pend;
            claim[(it + 1) & 1] = nx;
            if (nx < E)
                pend = atomicAdd(sched, 1ull);
        }
        __syncthreads();
        if (threadIdx.x < 27)
        {
            const int a = threadIdx.x / 9, b = (threadIdx.x / 3) % 3, c = threadIdx.x % 3;
            // d[a][b][c] = sum over the ways to split the exponents between the factors of
            //   det = sum_perm sgn * (dx_p0/dxi0)(xi1, xi2) * (dx_p1/dxi1)(xi0, xi2) * (dx_p2/dxi2)(xi0, xi1)
            double acc = 0.0;
            for (int a1 = 0; a1 < 2; ++a1)
            {
                const int a2 = a - a1;
                if (a2 < 0 || a2 > 1)
                    continue;
                for (int b0 = 0; b0 < 2; ++b0)
                {
                    const int b2 = b - b0;
                    if (b2 < 0 || b2 > 1)
                        continue;
                    for (int c0 = 0; c0 < 2; ++c0)
                    {
                        const int c1 = c - c0;
                        if (c1 < 0 || c1 > 1)
                            continue;
                        const int m0 = 4 + 2 * b0 + c0, m1 = 4 * a1 + 2 + c1, m2 = 4 * a2 + 2 * b2 + 1;
                        // cofactor expansion along the first factor, differences of products kept together
                        const double f0x = xm[0][m0], f0y = xm[1][m0], f0z = xm[2][m0];
                        const double f1x = xm[0][m1], f1y = xm[1][m1], f1z = xm[2][m1];
                        const double f2x = xm[0][m2], f2y = xm[1][m2], f2z = xm[2][m2];
                        acc += f0x * (f1y * f2z - f1z * f2y) + f0y * (f1z * f2x - f1x * f2z) + f0z * (f1x * f2y - f1y * f2x);
                    }
                }
            }
            dsm[threadIdx.x] = acc;
        }
        __syncthreads();
        double d[3][3][3];
#pragma unroll
        for (int k = 0; k < 27; ++k)
            d[k / 9][(k / 3) % 3][k % 3] = dsm[k];

        // ---- rows (j0, j1, j2) of column I
        size_t gs = e * prm.NN; // global index of the current run's first double (staged variant)
        double *p = out + e * prm.NN + I;
        // staged variant: where this thread's next entry goes (advanced by one row per entry, reset by flush)
        double *sp = k3_stage + (run % 3u) * prm.run_cap + (int)(gs & 1) + I;
        auto flush = [&](int rows) { // staged variant: ship the `rows` rows staged since the last flush
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            if (threadIdx.x == 0)
            {
                const int par = (int)(gs & 1);
                const double *sm = k3_stage + (run % 3u) * prm.run_cap + par;
                const int cnt = rows * prm.N_in;
                double *gp = out + gs;
                int k0 = 0;
                if (par)
                {
                    gp[0] = sm[0];
                    k0 = 1;
                }
                const int body = (cnt - k0) & ~1;
                if (body > 0)
                    k3_bulk_s2g(gp + k0, sm + k0, (unsigned)body * 8u);
                if (k0 + body < cnt)
                    gp[cnt - 1] = sm[cnt - 1];
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            }
            gs += (size_t)rows * prm.N_in;
            ++run;
            sp = k3_stage + (run % 3u) * prm.run_cap + (int)(gs & 1) + I;
        };
#pragma unroll
        for (int j0 = 0; j0 < NO; ++j0)
        {
            if (j0 < n0)
            {
                double u[3][3];
#pragma unroll
                for (int b = 0; b < 3; ++b)
#pragma unroll
                    for (int c = 0; c < 3; ++c)
                        u[b][c] = fma(d[2][b][c], X0[2][j0], fma(d[1][b][c], X0[1][j0], __dmul_rn(d[0][b][c], X0[0][j0])));
#pragma unroll
                for (int j1 = 0; j1 < NO; ++j1)
                {
                    if (j1 < n1)
                    {
                        double v[3];
#pragma unroll
                        for (int c = 0; c < 3; ++c)
                            v[c] = fma(u[2][c], X1[2][j1], fma(u[1][c], X1[1][j1], __dmul_rn(u[0][c], X1[0][j1])));
#pragma unroll
                        for (int j2 = 0; j2 < NO; ++j2)
                        {
                            if (j2 < n2)
                            {
                                const double val = fma(v[2], X2[2][j2], fma(v[1], X2[1][j2], __dmul_rn(v[0], X2[0][j2])));
                                if (STAGED)
                                {
                                    if (active)
                                        *sp = val;
                                    sp += prm.N_in;
                                }
                                else if (active)
                                {
                                    __stcs(p, val);
                                    p += prm.N_in;
                                }
                            }
                        }
                        if (STAGED && prm.run_level == 2)
                            flush(n2);
                    }
                }
                if (STAGED && prm.run_level == 1)
                    flush(n1 * n2);
            }
        }
        if (STAGED && prm.run_level == 0)
            flush(n0 * n1 * n2);
        // the next element's coefficient phases write xm / dsm: every thread has read dsm (before its row loop) and xm
        // (before the second barrier above); claim[(it + 1) & 1] was written before the first barrier
        e = claim[(it + 1) & 1];
    }
    if (STAGED && threadIdx.x == 0)
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    sched_finish(sched);
}

struct Kron3Plan
{
    bool usable = false;
    int NO = 0, N_in = 0, N_out = 0, n_out[3] = {0, 0, 0};
    double T[8][8] = {};
    double *d_tab = nullptr;
    ~Kron3Plan()
    {
        if (d_tab)
            cudaFree(d_tab);
    }
};

template <int NO, bool STAGED>
int launch_kron3(const LaunchCfg &cfg, const Kron3Plan &plan, size_t E, const double *geom, double *out)
{
    Kron3Params prm;
    std::memset(&prm, 0, sizeof prm);
    for (int k = 0; k < 3; ++k)
        prm.n_out[k] = plan.n_out[k];
    prm.N_in = plan.N_in;
    prm.NN = (unsigned long long)plan.N_out * plan.N_in;
    std::memcpy(prm.T, plan.T, sizeof prm.T);
    for (int m = 0; m < 8; ++m)
    {
        prm.Ts[m] = 0.0;
        for (int n = 0; n < 8; ++n)
            prm.Ts[m] += plan.T[m][n];
    }
    const int threads = std::max(32, (plan.N_in + 31) / 32 * 32);
    if (threads > k3_max_threads(NO))
        return IB200_ERROR_UNSUPPORTED;
    size_t smem_bytes = 0;
    if (STAGED)
    {
        // the largest run (whole element, one j0 slab, one (j0, j1) slab) of which three fit in 150 KB
        const size_t rows[3] = {(size_t)plan.N_out, (size_t)plan.n_out[1] * plan.n_out[2], (size_t)plan.n_out[2]};
        int level = 0;
        while (level < 2 && 3 * sizeof(double) * (rows[level] * plan.N_in + 3) > 150 * 1024)
            ++level;
        prm.run_level = level;
        prm.run_cap = (int)((rows[level] * plan.N_in + 3) & ~(size_t)1);
        smem_bytes = 3 * sizeof(double) * (size_t)prm.run_cap;
        if (smem_bytes > 200 * 1024)
            return IB200_ERROR_UNSUPPORTED;
    }
    auto kernel = mass3d_kron_kernel<NO, STAGED>;
    if (smem_bytes)
        IB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    int per_sm = 0;
    IB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem_bytes));
    if (per_sm < 1)
        return IB200_ERROR_UNSUPPORTED;
    const int grid = (int)std::max<size_t>(1, std::min(E, (size_t)cfg.sm_count * per_sm));
    kernel<<<grid, threads, smem_bytes, cfg.stream>>>(prm, plan.d_tab, E, geom, out, cfg.sched);
    ++*cfg.launch_counter;
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess)
        return ib200_fail(IB200_ERROR_CUDA, "launch of mass3d_kron_kernel<%d> failed: %s", NO, cudaGetErrorString(err));
    return INTERP_SUCCESS;
}

} // namespace

// Determinant-weighted mass matrices of trilinear hexahedra: geom[E][3][2][2][2] (DoFs of the order-1 geometry bases) ->
// out[E][N_out][N_in] for output spaces `rows[3]`, input spaces `cols[3]`.  IB200_ERROR_UNSUPPORTED (no error text) outside
// the envelope (more than 6 basis functions per dimension); E == 0 with out == NULL only answers that question.
int launch_mass3d_kron(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, const basis_spec_t *rows, const basis_spec_t *cols,
                       const integration_spec_t *rules, const basis_spec_t *geom_basis, size_t E, const double *geom, double *out)
{
    for (int k = 0; k < 3; ++k)
        if (geom_basis[k].order != 1)
            return IB200_ERROR_UNSUPPORTED;
    auto it = ctx->plans.find(key);
    if (it == ctx->plans.end())
    {
        auto plan = std::make_shared<Kron3Plan>();
        int need = 0, no[3], ni[3];
        for (int k = 0; k < 3; ++k)
        {
            no[k] = (int)rows[k].order + 1;
            ni[k] = (int)cols[k].order + 1;
            need = std::max(need, std::max(no[k], ni[k]));
            plan->n_out[k] = no[k];
        }
        static const int sizes[] = {2, 3, 4, 5, 6};
        int NO = 0;
        for (int v : sizes)
            if (!NO && v >= need)
                NO = v;
        const int N_in = ni[0] * ni[1] * ni[2], N_out = no[0] * no[1] * no[2];
        int res = (NO && std::max(32, (N_in + 31) / 32 * 32) <= k3_max_threads(NO)) ? (int)INTERP_SUCCESS : (int)IB200_ERROR_UNSUPPORTED;
        double a[3][2], s[3][2];
        for (int k = 0; k < 3 && res == INTERP_SUCCESS; ++k)
            res = ib200_tables_linear_basis(geom_basis[k], a[k], s[k]);
        std::vector<double> tab((size_t)N_in * 3 * 3 * std::max(NO, 1), 0.0);
        for (int k = 0; k < 3 && res == INTERP_SUCCESS; ++k)
        {
            std::vector<double> mom((size_t)3 * no[k] * ni[k]);
            res = ib200_tables_pair_moments_k(rows[k], cols[k], rules[k], 2, mom.data());
            if (res != INTERP_SUCCESS)
                break;
            for (int I = 0; I < N_in; ++I)
            {
                const int idx[3] = {I / (ni[1] * ni[2]), (I / ni[2]) % ni[1], I % ni[2]};
                for (int p = 0; p < 3; ++p)
                    for (int j = 0; j < no[k]; ++j)
                        tab[(((size_t)I * 3 + k) * 3 + p) * NO + j] = mom[((size_t)p * no[k] + j) * ni[k] + idx[k]];
            }
        }
        if (res == INTERP_SUCCESS)
        {
            // monomial m = 4 alpha + 2 beta + gamma (power of xi0, xi1, xi2) from geometry DoF n = 4 i0 + 2 i1 + i2
            for (int m = 0; m < 8; ++m)
                for (int n = 0; n < 8; ++n)
                {
                    const int al = m >> 2, be = (m >> 1) & 1, ga = m & 1, i0 = n >> 2, i1 = (n >> 1) & 1, i2 = n & 1;
                    plan->T[m][n] = (al ? s[0][i0] : a[0][i0]) * (be ? s[1][i1] : a[1][i1]) * (ga ? s[2][i2] : a[2][i2]);
                }
            if (cudaMalloc((void **)&plan->d_tab, tab.size() * sizeof(double)) != cudaSuccess ||
                cudaMemcpy(plan->d_tab, tab.data(), tab.size() * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess)
                return ib200_fail(IB200_ERROR_CUDA, "upload of the 3-D Kronecker mass tables failed");
            plan->NO = NO;
            plan->N_in = N_in;
            plan->N_out = N_out;
            plan->usable = true;
        }
        else if (res != IB200_ERROR_UNSUPPORTED)
            return res;
        it = ctx->plans.emplace(key, std::static_pointer_cast<void>(plan)).first;
    }
    const Kron3Plan &plan = *static_cast<const Kron3Plan *>(it->second.get());
    if (!plan.usable)
        return IB200_ERROR_UNSUPPORTED;
    if (E == 0 && out == nullptr)
        return INTERP_SUCCESS;
    bool staged = plan.N_in % 4 != 0 || (reinterpret_cast<size_t>(out) & 31) != 0;
    if (const char *env = std::getenv("IB200_KRON_STAGED")) // measuring aid: 0 = always direct stores, 1 = always staged
        staged = std::atoi(env) != 0;
#define IB_K3(O_) \
    if (plan.NO == O_) \
        return staged ? launch_kron3<O_, true>(cfg, plan, E, geom, out) : launch_kron3<O_, false>(cfg, plan, E, geom, out);
    IB_K3(2) IB_K3(3) IB_K3(4) IB_K3(5) IB_K3(6)
#undef IB_K3
    return IB200_ERROR_UNSUPPORTED;
}
