// Generic CUDA kernels: any reference dimension 1..3, any basis type / order per dimension, any rule.
//
// These are the "always correct" device path behind the C ABI: one thread per output value, the
// tensor-product structure is used only to nest the loops (so an output costs prod(n) or prod(m)
// fused multiply-adds, not d times that).  The specialised sum-factorised kernels in
// kernels_eval2d.cu / kernels_mass2d.cu take over for the shapes on BASELINE.json's configs; these
// kernels remain the device implementation for everything else and the cross-check for the fast
// paths (ib200_ctx_set_force_generic).  There is no host implementation of any of them.
//
// Reference loops restated here (per element; results agree to ~1e-15 relative, summation order
// differs): src/python/degrees_of_freedom.c:232-303 (reconstruct), src/python/mappings.c:448-542 +
// src/operations/matrices.c:15-141 (Jacobian, Givens QR, left inverse), src/python/mass_matrices.c:264-335
// and :486-560 (mass / gradient mass), python/interplib/integration.py:143-166 (L2 dual).
#include "ctx.cuh"

namespace
{

constexpr int kThreads = 256;

constexpr int MAXD = IB200_MAX_DIM;

__device__ __forceinline__ void unflatten(int flat, const OpDesc &op, bool use_m, bool use_out, int idx[MAXD])
{
    for (int k = 0; k < MAXD; ++k)
        idx[k] = 0;
    for (int k = op.ndim - 1; k >= 0; --k)
    {
        const int ext = use_m ? op.dim[k].m : (use_out ? op.dim[k].n_out : op.dim[k].n_in);
        idx[k] = flat % ext;
        flat /= ext;
    }
}

// sum_i prod_k T_k[i_k][q_k] * dof[i], nested so the work is N fused multiply-adds.
// More than three reference dimensions (the reference accepts any count, e.g. tests/test_dofs.py:79 uses five):
// plain loop over all basis functions with an incrementing multi-index, prod(n)*d multiplies per output.
__device__ double tensor_contract_nd(const OpDesc &op, unsigned mask, const int q[MAXD], const double *__restrict__ dof)
{
    int idx[MAXD] = {0};
    double acc = 0.0;
    for (int i = 0; i < op.N_in; ++i)
    {
        double phi = 1.0;
        for (int k = 0; k < op.ndim; ++k)
        {
            const DimTab &d = op.dim[k];
            phi *= (((mask >> k) & 1u) ? d.Din : d.Bin)[idx[k] * d.m + q[k]];
        }
        acc = fma(phi, dof[i], acc);
        for (int k = op.ndim - 1; k >= 0; --k)
        {
            if (++idx[k] < op.dim[k].n_in)
                break;
            idx[k] = 0;
        }
    }
    return acc;
}

__device__ double tensor_contract(const OpDesc &op, unsigned mask, const int q[MAXD], const double *__restrict__ dof)
{
    if (op.ndim > 3)
        return tensor_contract_nd(op, mask, q, dof);
    const DimTab &d0 = op.dim[0];
    const double *T0 = ((mask & 1u) ? d0.Din : d0.Bin) + q[0];
    if (op.ndim == 1)
    {
        double acc = 0.0;
        for (int i0 = 0; i0 < d0.n_in; ++i0)
            acc = fma(T0[i0 * d0.m], dof[i0], acc);
        return acc;
    }
    const DimTab &d1 = op.dim[1];
    const double *T1 = ((mask & 2u) ? d1.Din : d1.Bin) + q[1];
    if (op.ndim == 2)
    {
        double acc = 0.0;
        for (int i0 = 0; i0 < d0.n_in; ++i0)
        {
            double a1 = 0.0;
            for (int i1 = 0; i1 < d1.n_in; ++i1)
                a1 = fma(T1[i1 * d1.m], dof[i0 * d1.n_in + i1], a1);
            acc = fma(T0[i0 * d0.m], a1, acc);
        }
        return acc;
    }
    const DimTab &d2 = op.dim[2];
    const double *T2 = ((mask & 4u) ? d2.Din : d2.Bin) + q[2];
    double acc = 0.0;
    for (int i0 = 0; i0 < d0.n_in; ++i0)
    {
        double a1 = 0.0;
        for (int i1 = 0; i1 < d1.n_in; ++i1)
        {
            double a2 = 0.0;
            const double *row = dof + (size_t)(i0 * d1.n_in + i1) * d2.n_in;
            for (int i2 = 0; i2 < d2.n_in; ++i2)
                a2 = fma(T2[i2 * d2.m], row[i2], a2);
            a1 = fma(T1[i1 * d1.m], a2, a1);
        }
        acc = fma(T0[i0 * d0.m], a1, acc);
    }
    return acc;
}

// ---- K1 generic: out[e][f][q] -----------------------------------------------------------------------
struct MaskList
{
    unsigned n;
    unsigned mask[8];
};

__global__ void __launch_bounds__(kThreads) reconstruct_generic_kernel(const __grid_constant__ OpDesc op, size_t E,
                                                                       const double *__restrict__ dofs,
                                                                       const __grid_constant__ MaskList fields,
                                                                       double *__restrict__ out)
{
    const size_t per_elem = (size_t)fields.n * op.Q;
    const size_t total = E * per_elem;
    for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x)
    {
        const size_t e = t / per_elem;
        const int r = (int)(t % per_elem);
        const int f = r / op.Q;
        int q[MAXD];
        unflatten(r % op.Q, op, true, false, q);
        out[t] = tensor_contract(op, fields.mask[f], q, dofs + e * op.N_in);
    }
}

// ---- K0 generic: geometry per (element, point) --------------------------------------------------------
// Givens QR exactly as src/operations/matrices.c:15-86 orders its rotations (rows outer, columns inner),
// R in place, Q^T accumulated; then det = prod diag(R) and the left inverse by back-substitution
// (src/python/mappings.c:483-527).  Products and sums are rounded separately (no FMA contraction) like the
// reference build, so apart from hypot() (<= 1 ulp in both libm and CUDA) the results are bit-identical.
__device__ void qr_det_inverse(int rows, int cols, double *r, double *det_out, double *inv /*[cols][rows] or null*/)
{
    double qt[IB200_MAX_PHYS * IB200_MAX_PHYS];
    for (int i = 0; i < rows; ++i)
        for (int j = 0; j < rows; ++j)
            qt[i * rows + j] = (i == j) ? 1.0 : 0.0;
    if (rows > 1)
    {
        for (int row = 0; row < rows; ++row)
            for (int col = 0; col < cols && col < row; ++col)
            {
                double c = r[col * cols + col];
                double s = -r[row * cols + col];
                const double mag = hypot(c, s);
                if (mag < 1e-12)
                    continue;
                c /= mag;
                s /= mag;
                for (int k = col; k < cols; ++k)
                {
                    const double a = r[col * cols + k];
                    const double b = r[row * cols + k];
                    r[col * cols + k] = __dsub_rn(__dmul_rn(c, a), __dmul_rn(s, b));
                    r[row * cols + k] = __dadd_rn(__dmul_rn(s, a), __dmul_rn(c, b));
                }
                r[row * cols + col] = 0.0;
                for (int k = 0; k < rows; ++k)
                {
                    const double a = qt[col * rows + k];
                    const double b = qt[row * rows + k];
                    qt[col * rows + k] = __dsub_rn(__dmul_rn(c, a), __dmul_rn(s, b));
                    qt[row * rows + k] = __dadd_rn(__dmul_rn(s, a), __dmul_rn(c, b));
                }
            }
    }
    double det = 1.0;
    for (int i = 0; i < cols; ++i)
        det *= r[i * cols + i];
    *det_out = det;
    if (inv)
    {
        for (int bc = 0; bc < rows; ++bc)
            for (int row = cols; row > 0; --row)
            {
                double v = qt[(row - 1) * rows + bc];
                for (int col = row; col < cols; ++col)
                    v = __dsub_rn(v, __dmul_rn(r[(row - 1) * cols + col], inv[col * rows + bc]));
                inv[(row - 1) * rows + bc] = v / r[(row - 1) * cols + (row - 1)];
            }
    }
}

// Geometry is the ill-conditioned input of the path (coordinates are O(1), their differences O(h)), so the
// coordinate maps are evaluated in the reference's own operation order -- i row-major, phi = T0*T1*T2,
// val = val + phi*dof with separately rounded multiply and add (the reference is built without FMA
// contraction) -- which makes the Jacobian entries bit-identical to compute_integration_point_values[_derivatives]
// (src/python/degrees_of_freedom.c:232-303) instead of merely close.
__device__ double tensor_contract_reference_order(const OpDesc &op, unsigned mask, const int q[MAXD], const double *__restrict__ dof)
{
    int idx[MAXD] = {0};
    double val = 0.0;
    for (int i = 0; i < op.N_in; ++i)
    {
        double phi = 1.0;
        for (int k = 0; k < op.ndim; ++k)
        {
            const DimTab &d = op.dim[k];
            phi = __dmul_rn(phi, (((mask >> k) & 1u) ? d.Din : d.Bin)[idx[k] * d.m + q[k]]);
        }
        val = __dadd_rn(val, __dmul_rn(phi, dof[i]));
        for (int k = op.ndim - 1; k >= 0; --k)
        {
            if (++idx[k] < op.dim[k].n_in)
                break;
            idx[k] = 0;
        }
    }
    return val;
}

__global__ void __launch_bounds__(kThreads)
    geometry_generic_kernel(const __grid_constant__ OpDesc op, int n_phys, size_t E, const double *__restrict__ gdofs,
                            double *__restrict__ coords, double *__restrict__ det, double *__restrict__ jinv,
                            double *__restrict__ weight_out, int w_idx_in, int w_idx_out)
{
    const int ndim = op.ndim;
    const size_t total = E * (size_t)op.Q;
    for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x)
    {
        const size_t e = t / op.Q;
        const int qf = (int)(t % op.Q);
        int q[MAXD];
        unflatten(qf, op, true, false, q);
        double jac[IB200_MAX_PHYS * MAXD];
        for (int a = 0; a < n_phys; ++a)
        {
            const double *dof = gdofs + (e * n_phys + a) * op.N_in;
            if (coords)
                coords[((e * n_phys + a) * (1 + ndim)) * op.Q + qf] = tensor_contract_reference_order(op, 0u, q, dof);
            for (int k = 0; k < ndim; ++k)
            {
                const double g = tensor_contract_reference_order(op, 1u << k, q, dof);
                jac[a * ndim + k] = g;
                if (coords)
                    coords[((e * n_phys + a) * (1 + ndim) + 1 + k) * op.Q + qf] = g;
            }
        }
        if (det || jinv || weight_out)
        {
            double dv, inv[IB200_MAX_PHYS * MAXD];
            qr_det_inverse(n_phys, ndim, jac, &dv, inv);
            if (det)
                det[t] = dv;
            if (jinv)
                for (int i = 0; i < ndim * n_phys; ++i)
                    jinv[t * (size_t)(ndim * n_phys) + i] = inv[i];
            if (weight_out)
                weight_out[t] = (w_idx_in < 0) ? dv : dv * inv[w_idx_in * n_phys + w_idx_out];
        }
    }
}

// ---- K2 generic: out[e][j][i] --------------------------------------------------------------------------
// One thread per entry of the lower triangle (symmetric) or of the full matrix; the quadrature sum is
// nested dimension by dimension: sum_q0 P0 sum_q1 P1 sum_q2 P2 * wt[q], Pk = Bout_k[j_k][q_k]*Tin_k[i_k][q_k].
__global__ void __launch_bounds__(kThreads)
    mass_generic_kernel(const __grid_constant__ OpDesc op, int symmetric, size_t E, const double *__restrict__ weight,
                        int deriv_in, double *__restrict__ out)
{
    const size_t per_elem = (size_t)op.N_out * op.N_in;
    const size_t total = E * per_elem;
    const DimTab &d0 = op.dim[0];
    const DimTab &d1 = op.dim[1];
    const DimTab &d2 = op.dim[2];
    const int m1 = op.ndim > 1 ? d1.m : 1;
    const int m2 = op.ndim > 2 ? d2.m : 1;
    for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x)
    {
        const size_t e = t / per_elem;
        const int r = (int)(t % per_elem);
        const int jo = r / op.N_in, ii = r % op.N_in;
        if (symmetric && ii > jo)
            continue;
        int bo[MAXD], bi[MAXD];
        unflatten(jo, op, false, true, bo);
        unflatten(ii, op, false, false, bi);
        const double *wt = weight ? weight + e * op.Q : nullptr;
        if (op.ndim > 3)
        {
            // any dimension count: loop over all quadrature points with an incrementing multi-index
            int qi[MAXD] = {0};
            double acc = 0.0;
            for (int qf = 0; qf < op.Q; ++qf)
            {
                double term = wt ? wt[qf] : 1.0;
                for (int k = 0; k < op.ndim; ++k)
                {
                    const DimTab &d = op.dim[k];
                    const double *Ti = (deriv_in == k ? d.Din : d.Bin) + bi[k] * d.m;
                    term *= Ti[qi[k]] * d.Bout[bo[k] * d.m + qi[k]] * d.w[qi[k]];
                }
                acc += term;
                for (int k = op.ndim - 1; k >= 0; --k)
                {
                    if (++qi[k] < op.dim[k].m)
                        break;
                    qi[k] = 0;
                }
            }
            out[t] = acc;
            if (symmetric && ii != jo)
                out[e * per_elem + (size_t)ii * op.N_in + jo] = acc;
            continue;
        }
        const double *Ti0 = (deriv_in == 0 ? d0.Din : d0.Bin) + bi[0] * d0.m;
        const double *To0 = d0.Bout + bo[0] * d0.m;
        double acc = 0.0;
        for (int q0 = 0; q0 < d0.m; ++q0)
        {
            double a1 = 0.0;
            if (op.ndim == 1)
            {
                a1 = wt ? wt[q0] : 1.0;
            }
            else
            {
                const double *Ti1 = (deriv_in == 1 ? d1.Din : d1.Bin) + bi[1] * d1.m;
                const double *To1 = d1.Bout + bo[1] * d1.m;
                for (int q1 = 0; q1 < m1; ++q1)
                {
                    double a2 = 0.0;
                    if (op.ndim == 2)
                    {
                        a2 = wt ? wt[q0 * m1 + q1] : 1.0;
                    }
                    else
                    {
                        const double *Ti2 = (deriv_in == 2 ? d2.Din : d2.Bin) + bi[2] * d2.m;
                        const double *To2 = d2.Bout + bo[2] * d2.m;
                        for (int q2 = 0; q2 < m2; ++q2)
                        {
                            const double w = wt ? wt[(q0 * m1 + q1) * m2 + q2] : 1.0;
                            a2 = fma(Ti2[q2] * To2[q2] * d2.w[q2], w, a2);
                        }
                    }
                    a1 = fma(Ti1[q1] * To1[q1] * d1.w[q1], a2, a1);
                }
            }
            acc = fma(Ti0[q0] * To0[q0] * d0.w[q0], a1, acc);
        }
        out[t] = acc;
        if (symmetric && ii != jo)
            out[e * per_elem + (size_t)ii * op.N_in + jo] = acc;
    }
}

// ---- K3 generic: out[e][i] = sum_q f[q] W[q] det[q] Phi_i(q) ------------------------------------------------
__global__ void __launch_bounds__(kThreads)
    l2_generic_kernel(const __grid_constant__ OpDesc op, size_t E, const double *__restrict__ fvals,
                      const double *__restrict__ det, double *__restrict__ out)
{
    const size_t total = E * (size_t)op.N_in;
    const DimTab &d0 = op.dim[0];
    const DimTab &d1 = op.dim[1];
    const DimTab &d2 = op.dim[2];
    const int m1 = op.ndim > 1 ? d1.m : 1;
    const int m2 = op.ndim > 2 ? d2.m : 1;
    for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x)
    {
        const size_t e = t / op.N_in;
        int bi[MAXD];
        unflatten((int)(t % op.N_in), op, false, false, bi);
        const double *f = fvals + e * op.Q;
        const double *dt = det ? det + e * op.Q : nullptr;
        double acc = 0.0;
        if (op.ndim > 3)
        {
            int qi[MAXD] = {0};
            for (int qf = 0; qf < op.Q; ++qf)
            {
                double term = dt ? f[qf] * dt[qf] : f[qf];
                for (int k = 0; k < op.ndim; ++k)
                    term *= op.dim[k].w[qi[k]] * op.dim[k].Bin[bi[k] * op.dim[k].m + qi[k]];
                acc += term;
                for (int k = op.ndim - 1; k >= 0; --k)
                {
                    if (++qi[k] < op.dim[k].m)
                        break;
                    qi[k] = 0;
                }
            }
            out[t] = acc;
            continue;
        }
        for (int q0 = 0; q0 < d0.m; ++q0)
        {
            double a1 = 0.0;
            for (int q1 = 0; q1 < m1; ++q1)
            {
                double a2 = 0.0;
                for (int q2 = 0; q2 < m2; ++q2)
                {
                    const int qf = (q0 * m1 + q1) * m2 + q2;
                    double v = f[qf];
                    if (dt)
                        v *= dt[qf];
                    if (op.ndim > 2)
                        v *= d2.w[q2] * d2.Bin[bi[2] * d2.m + q2];
                    a2 += v;
                }
                if (op.ndim > 1)
                    a2 *= d1.w[q1] * d1.Bin[bi[1] * d1.m + q1];
                a1 += a2;
            }
            acc = fma(d0.w[q0] * d0.Bin[bi[0] * d0.m + q0], a1, acc);
        }
        out[t] = acc;
    }
}

// ---- N1: batched solve of M u = b --------------------------------------------------------------------------------
// The reference solves with LAPACK LU (np.linalg.solve, python/interplib/integration.py:221), which does not care about
// the sign of M: a negatively oriented element (clockwise quad corners, a line with decreasing knots) has det J < 0
// and a NEGATIVE definite mass matrix with a perfectly good solution.  Every kernel below therefore factorises
// sign(M[0][0]) * M and substitutes sign(M[0][0]) * b (same u), and raises IB200_STATUS_BAD_SYSTEM when the solution is
// not finite (a non-positive pivot of the signed matrix gives rsqrt -> NaN/inf): singular and indefinite systems
// surface as INTERP_ERROR_BAD_SYSTEM instead of silent NaNs.
__device__ __forceinline__ void spd_flag_if_bad(double u, int *status)
{
    if (!(fabs(u) <= 1.79769313486231570e308)) // NaN or infinite
        atomicOr(status, IB200_STATUS_BAD_SYSTEM);
}

// one warp per element, Cholesky in place in global/L2 (fallback for systems no other kernel takes)
// projection_l2_primal (python/interplib/integration.py:212-221) solves M u = b per element with LAPACK LU;
// M is SPD (cond <= ~1e3 on these spaces) so Cholesky agrees to ~1e-14.
__global__ void __launch_bounds__(128) spd_solve_kernel(int n, size_t E, double *__restrict__ mass, double *__restrict__ rhs,
                                                        int *__restrict__ status)
{
    const int lane = threadIdx.x & 31;
    const size_t warp = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    const size_t n_warps = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t e = warp; e < E; e += n_warps)
    {
        double *A = mass + e * (size_t)n * n;
        double *b = rhs + e * (size_t)n;
        if (A[0] < 0.0) // warp uniform: negative definite -> factorise -M, solve with -b
        {
            __syncwarp();
            for (int idx = lane; idx < n * n; idx += 32)
                A[idx] = -A[idx];
            for (int j = lane; j < n; j += 32)
                b[j] = -b[j];
        }
        // right-looking Cholesky, lower triangle
        for (int k = 0; k < n; ++k)
        {
            __syncwarp();
            const double akk = sqrt(A[k * n + k]);
            __syncwarp();
            if (lane == 0)
                A[k * n + k] = akk;
            for (int i = k + 1 + lane; i < n; i += 32)
                A[i * n + k] /= akk;
            __syncwarp();
            for (int idx = lane; idx < (n - k - 1) * (n - k - 1); idx += 32)
            {
                const int i = k + 1 + idx / (n - k - 1);
                const int j = k + 1 + idx % (n - k - 1);
                if (j <= i)
                    A[i * n + j] -= A[i * n + k] * A[j * n + k];
            }
        }
        __syncwarp();
        // forward then backward substitution (serial in k, parallel in rows)
        for (int k = 0; k < n; ++k)
        {
            __syncwarp();
            const double yk = b[k] / A[k * n + k];
            __syncwarp();
            if (lane == 0)
                b[k] = yk;
            for (int i = k + 1 + lane; i < n; i += 32)
                b[i] -= A[i * n + k] * yk;
        }
        for (int k = n - 1; k >= 0; --k)
        {
            __syncwarp();
            const double xk = b[k] / A[k * n + k];
            __syncwarp();
            if (lane == 0)
                b[k] = xk;
            for (int i = lane; i < k; i += 32)
                b[i] -= A[k * n + i] * xk;
        }
        __syncwarp();
        for (int j = lane; j < n; j += 32)
            spd_flag_if_bad(b[j], status);
    }
}

// Shared-memory variant for systems that fit (n up to ~100): one warp per element, the matrix is read from global memory
// ONCE (coalesced), factorised in shared memory and only the solution is written -- 8(n^2 + 2n) bytes per element instead
// of the read-modify-write traffic of the in-place kernel above.  Left-looking (Crout) Cholesky with the right-hand side
// appended as row n, so the forward substitution is the same recurrence: at column k the lane owning row i >= k forms
// s_i = A[i][k] - sum_{m<k} L[i][m] L[k][m] (its own row: stride ld, odd -> conflict-free; the pivot row: broadcast),
// then L[i][k] = s_i * rsqrt(s_k).  Back substitution is column oriented (no reductions).
__global__ void __launch_bounds__(128) spd_solve_smem_kernel(int n, int ld, size_t E, const double *__restrict__ mass,
                                                             double *__restrict__ rhs, int *__restrict__ status)
{
    extern __shared__ __align__(16) double chol_smem[];
    const int lane = threadIdx.x & 31, warp_in_cta = threadIdx.x >> 5;
    double *L = chol_smem + (size_t)warp_in_cta * (n + 1) * ld;
    const size_t warp = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    const size_t n_warps = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t e = warp; e < E; e += n_warps)
    {
        const double *A = mass + e * (size_t)n * n;
        double *b = rhs + e * (size_t)n;
        __syncwarp();
        const double sg = A[0] < 0.0 ? -1.0 : 1.0;
        for (int idx = lane, i = 0, j = lane; idx < n * n; idx += 32, j += 32)
        {
            while (j >= n)
                j -= n, ++i;
            L[i * ld + j] = sg * __ldcs(A + idx);
        }
        for (int j = lane; j < n; j += 32)
            L[n * ld + j] = sg * b[j];
        __syncwarp();
        for (int k = 0; k < n; ++k)
        {
            const double *pk = L + k * ld;
            double s_first = 0.0; // s of the first row this lane owns (row k for lane 0)
            double sv[4];         // up to 4 rows per lane: n + 1 <= 128
            int cnt = 0;
            for (int i = k + lane; i <= n; i += 32, ++cnt)
            {
                const double *pi = L + i * ld;
                double a0 = pi[k], a1 = 0.0, a2 = 0.0, a3 = 0.0;
                int m = 0;
                for (; m + 4 <= k; m += 4)
                {
                    a0 = fma(-pi[m], pk[m], a0);
                    a1 = fma(-pi[m + 1], pk[m + 1], a1);
                    a2 = fma(-pi[m + 2], pk[m + 2], a2);
                    a3 = fma(-pi[m + 3], pk[m + 3], a3);
                }
                for (; m < k; ++m)
                    a0 = fma(-pi[m], pk[m], a0);
                sv[cnt] = (a0 + a1) + (a2 + a3);
                if (cnt == 0)
                    s_first = sv[0];
            }
            // one reciprocal square root per column instead of a square root and a division per row (both are long
            // software sequences in FP64); the diagonal keeps 1 / L[k][k], which is all the substitutions need
            const double rkk = rsqrt(__shfl_sync(0xffffffffu, s_first, 0));
            cnt = 0;
            for (int i = k + lane; i <= n; i += 32, ++cnt)
                L[i * ld + k] = (i == k) ? rkk : sv[cnt] * rkk;
            __syncwarp();
        }
        // L[n][k] = y_k;  L^T u = y, column oriented: u_k = y_k / L[k][k], then y_i -= L[k][i] u_k for i < k
        double *y = L + n * ld;
        for (int k = n - 1; k >= 0; --k)
        {
            const double uk = y[k] * L[k * ld + k];
            __syncwarp();
            for (int i = lane; i < k; i += 32)
                y[i] = fma(-L[k * ld + i], uk, y[i]);
            if (lane == 0)
                y[k] = uk;
            __syncwarp();
        }
        for (int j = lane; j < n; j += 32)
        {
            b[j] = y[j];
            spd_flag_if_bad(y[j], status);
        }
    }
}

// Register-resident variant for n <= 32 (every 2-D space up to p = 4, 3-D up to p = 2 -- BASELINE config 5 is 25 x 25):
// one warp per element, lane i holds row i of the lower triangle in registers (all loops unrolled at compile time, so the
// indices are static).  Right-looking factorisation: at column k the pivot is broadcast by one shuffle, ONE reciprocal
// square root scales the column, the scaled column goes to shared memory (row k of L^T, later read by the back
// substitution with unit stride) and comes back as warp-uniform broadcasts for the rank-1 update of the trailing rows
// a[j] -= l_ik * l_jk.  The forward substitution rides along (b is one more register).  About 1/5 of the instructions of
// the shared-memory kernel above, which spends most of its issue slots on loop control and the FP64 division routine.
template <int N>
__global__ void __launch_bounds__(128) spd_solve_reg_kernel(size_t E, const double *__restrict__ mass, double *__restrict__ rhs,
                                                            int *__restrict__ status)
{
    constexpr int LD = N | 1, PANEL = (N * LD + 1) & ~1; // doubles; panels stay 16-byte aligned
    extern __shared__ __align__(16) double chol_reg_smem[]; // [4 warps][2 panels][PANEL]
    const int lane = threadIdx.x & 31;
    double *panels = chol_reg_smem + (size_t)(threadIdx.x >> 5) * 2 * PANEL;
    const size_t warp = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    const size_t n_warps = ((size_t)gridDim.x * blockDim.x) >> 5;
    const int row = lane < N ? lane : N - 1; // idle lanes shadow the last row (their results are never stored)
    // the matrix of the NEXT element streams into the other panel (cp.async, 8 bytes per lane and request: an element
    // starts only 8-byte aligned when N is odd) while this one is factorised
    auto stage = [&](size_t e, double *P) {
        const double *A = mass + e * (size_t)(N * N);
        for (int idx = lane, i = 0, j = lane; idx < N * N; idx += 32, j += 32)
        {
            while (j >= N)
                j -= N, ++i;
            asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(P + i * LD + j)),
                         "l"(A + idx)
                         : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    if (warp < E)
        stage(warp, panels);
    int cur = 0;
    for (size_t e = warp; e < E; e += n_warps, cur ^= 1)
    {
        double *S = panels + cur * PANEL;
        double *b = rhs + e * (size_t)N;
        double y = b[row];
        __syncwarp(); // every lane is done with the other panel (back substitution of the previous element)
        if (e + n_warps < E)
        {
            stage(e + n_warps, panels + (cur ^ 1) * PANEL);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        }
        else
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
        double a[N];
#pragma unroll
        for (int j = 0; j < N; ++j)
            a[j] = S[row * LD + j];
        if (S[0] < 0.0) // warp uniform: negative definite -> factorise -M, solve with -b
        {
#pragma unroll
            for (int j = 0; j < N; ++j)
                a[j] = -a[j];
            y = -y;
        }
        __syncwarp();
        double rinv = 0.0; // 1 / L[lane][lane]
#pragma unroll
        for (int k = 0; k < N; ++k)
        {
            const double r = rsqrt(__shfl_sync(0xffffffffu, a[k], k));
            a[k] *= r; // l_ik for the lanes i >= k (lane k: sqrt of the pivot)
            const double yk = __shfl_sync(0xffffffffu, y, k) * r;
            if (lane == k)
                y = yk, rinv = r;
            else if (lane > k)
                y = fma(-a[k], yk, y);
            if (k + 1 < N)
            {
                if (lane < N)
                    S[k * LD + lane] = a[k];
                __syncwarp();
                // rank-1 update of the trailing rows, a[j] -= l_ik l_jk (touches only entries that are still live for
                // the lanes i >= j); l_jk comes back as warp-uniform broadcasts, two per load where the address is
                // 16-byte aligned: k LD + k + 2 is even because LD is odd
                const double *c = S + k * LD;
                a[k + 1] = fma(-a[k], c[k + 1], a[k + 1]);
#pragma unroll
                for (int j = k + 2; j + 1 < N; j += 2)
                {
                    const double2 l = *reinterpret_cast<const double2 *>(c + j);
                    a[j] = fma(-a[k], l.x, a[j]);
                    a[j + 1] = fma(-a[k], l.y, a[j + 1]);
                }
                if (k + 2 < N && ((N - k) & 1))
                    a[N - 1] = fma(-a[k], c[N - 1], a[N - 1]);
            }
        }
        // L^T u = y, column oriented: u_k = y_k / l_kk, then y_i -= l_ki u_k for i < k (l_ki = S[i][k]: odd stride LD,
        // conflict free); the lanes i > k already hold u_i
        __syncwarp();
#pragma unroll
        for (int k = N - 1; k >= 0; --k)
        {
            const double uk = __shfl_sync(0xffffffffu, y * rinv, k);
            if (lane == k)
                y = uk;
            else if (lane < k)
                y = fma(-S[lane * LD + k], uk, y);
        }
        if (lane < N)
        {
            b[lane] = y;
            spd_flag_if_bad(y, status);
        }
    }
}

template <int N>
static void launch_spd_solve_reg(const LaunchCfg &cfg, size_t E, const double *mass, double *rhs)
{
    constexpr int LD = N | 1, PANEL = (N * LD + 1) & ~1;
    const int smem = (int)sizeof(double) * 4 * 2 * PANEL;
    if (smem > 48 * 1024) // per launch: the attribute is per device
        cudaFuncSetAttribute(spd_solve_reg_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, spd_solve_reg_kernel<N>, 128, smem) != cudaSuccess || per_sm < 1)
        per_sm = 1;
    const size_t blocks = std::min<size_t>((E + 3) / 4, (size_t)cfg.sm_count * per_sm);
    spd_solve_reg_kernel<N><<<(int)blocks, 128, smem, cfg.stream>>>(E, mass, rhs, cfg.status);
}

// Two-block register variant for 33 <= n <= 64 (the C2 space is 49 x 49, a C3 block pair 72 is beyond it): still ONE WARP
// PER ELEMENT and no CTA barrier.  Rows 0..31 are factorised as in spd_solve_reg_kernel (lane = row, columns of L11 to
// shared memory); then lane r takes row 32 + r: its 32 entries of A21 are solved against L11 in registers (x L11^T = a,
// L11 as warp-uniform broadcasts) and stored as panel row X[r]; its row of the Schur complement A22 - X X^T is formed
// in registers (X rows as broadcasts, two per LDS.128) and factorised by the same routine, padded with the identity to
// 32 x 32.  The right-hand side rides along in two registers (y1, y2); the back substitution runs block 2, then
// y1 -= X^T u2, then block 1.  Rows come straight from global memory (a lane reads its own row: every sector is used by
// four consecutive loads and stays in L1).
constexpr int kReg2LD = 33, kReg2XLD = 34;

// a[32]: row `lane` of a 32 x 32 SPD block (lower triangle valid) -> L in place, columns to S[k][i] = l_ik; forward
// substitution of y on the way; returns 1 / l_lane,lane
__device__ __forceinline__ double chol_factor32(double (&a)[32], double &y, double *S, int lane, int n_act)
{
    double rinv = 0.0;
#pragma unroll
    for (int k = 0; k < 32; ++k)
    {
        if (k >= n_act) // warp uniform: the identity padding needs no work (y = 0 there)
            break;
        const double r = rsqrt(__shfl_sync(0xffffffffu, a[k], k));
        a[k] *= r;
        const double yk = __shfl_sync(0xffffffffu, y, k) * r;
        if (lane == k)
            y = yk, rinv = r;
        else if (lane > k)
            y = fma(-a[k], yk, y);
        S[k * kReg2LD + lane] = a[k];
        if (k + 1 < 32)
        {
            __syncwarp();
#pragma unroll
            for (int j = k + 1; j < 32; ++j)
                a[j] = fma(-a[k], S[k * kReg2LD + j], a[j]);
        }
    }
    __syncwarp();
    return rinv;
}

// L^T u = y for a factor stored by chol_factor32 (l_ki = S[i][k]); y -> u in place
__device__ __forceinline__ void chol_back32(double &y, double rinv, const double *S, int lane, int n_act)
{
#pragma unroll
    for (int k = 31; k >= 0; --k)
    {
        if (k >= n_act)
            continue;
        const double uk = __shfl_sync(0xffffffffu, y * rinv, k);
        if (lane == k)
            y = uk;
        else if (lane < k)
            y = fma(-S[lane * kReg2LD + k], uk, y);
    }
}

__global__ void __launch_bounds__(128) spd_solve_reg2_kernel(int n, size_t E, const double *__restrict__ mass, double *__restrict__ rhs,
                                                             int *__restrict__ status)
{
    extern __shared__ __align__(16) double chol_reg2_smem[];
    constexpr int PER_WARP = 2 * 32 * kReg2LD + 32 * kReg2XLD + 32; // S1 | S2 | X | y1
    const int lane = threadIdx.x & 31;
    double *S1 = chol_reg2_smem + (size_t)(threadIdx.x >> 5) * PER_WARP;
    double *S2 = S1 + 32 * kReg2LD;
    double *X = S2 + 32 * kReg2LD; // 16-byte aligned: 2 * 32 * 33 doubles and PER_WARP are even
    double *Y1 = X + 32 * kReg2XLD;
    const size_t warp = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    const size_t n_warps = ((size_t)gridDim.x * blockDim.x) >> 5;
    const int n2 = n - 32;
    const bool valid = lane < n2;
    const int rr = valid ? 32 + lane : 32; // idle lanes shadow row 32 for the panel (results never used)
    for (size_t e = warp; e < E; e += n_warps)
    {
        const double *A = mass + e * (size_t)n * n;
        double *b = rhs + e * (size_t)n;
        __syncwarp();
        double a[32], x[32];
        const bool neg = A[0] < 0.0; // negative definite -> factorise -M, solve with -b (a predicate, not a register pair)
        auto sgn = [neg](double v) { return neg ? -v : v; };
#pragma unroll
        for (int j = 0; j < 32; ++j)
            a[j] = sgn(A[(size_t)lane * n + j]);
#pragma unroll
        for (int j = 0; j < 32; ++j)
            x[j] = sgn(A[(size_t)rr * n + j]);
        double y1 = sgn(b[lane]), y2 = valid ? sgn(b[32 + lane]) : 0.0;
        const double rinv1 = chol_factor32(a, y1, S1, lane, 32);
        Y1[lane] = y1;
        __syncwarp();
        // panel: x L11^T = a21 (l_jm = S1[m][j]), and the forward substitution of y2 through it
#pragma unroll
        for (int j = 0; j < 32; ++j)
        {
            double sacc = x[j];
#pragma unroll
            for (int m = 0; m < j; ++m)
                sacc = fma(-x[m], S1[m * kReg2LD + j], sacc);
            x[j] = sacc * __shfl_sync(0xffffffffu, rinv1, j);
            y2 = fma(-x[j], Y1[j], y2);
        }
#pragma unroll
        for (int j = 0; j < 32; ++j)
            X[lane * kReg2XLD + j] = x[j];
        __syncwarp();
        // Schur complement row: a22[j] - sum_m x[m] X[j][m] for the columns j < n2, identity beyond (and on idle lanes)
#pragma unroll
        for (int j = 0; j < 32; ++j)
        {
            double sacc = (j == lane) ? 1.0 : 0.0;
            if (j < n2) // warp uniform
            {
                sacc = valid ? sgn(A[(size_t)rr * n + 32 + j]) : sacc;
                const double *xj = X + j * kReg2XLD;
                double s0 = 0.0, s1 = 0.0;
#pragma unroll
                for (int m = 0; m < 32; m += 2)
                {
                    const double2 v = *reinterpret_cast<const double2 *>(xj + m);
                    s0 = fma(x[m], v.x, s0);
                    s1 = fma(x[m + 1], v.y, s1);
                }
                if (valid)
                    sacc -= s0 + s1;
            }
            a[j] = sacc;
        }
        if (!valid)
            y2 = 0.0;
        const double rinv2 = chol_factor32(a, y2, S2, lane, n2);
        // back substitution: block 2, y1 -= X^T u2, block 1
        chol_back32(y2, rinv2, S2, lane, n2);
#pragma unroll
        for (int r = 0; r < 32; ++r)
            if (r < n2) // warp uniform
                y1 = fma(-X[r * kReg2XLD + lane], __shfl_sync(0xffffffffu, y2, r), y1);
        chol_back32(y1, rinv1, S1, lane, 32);
        b[lane] = y1;
        if (valid)
            b[32 + lane] = y2;
        spd_flag_if_bad(y1 + (valid ? y2 : 0.0), status);
    }
}

// Blocked variant for n > 32: ONE CTA PER ELEMENT, right-looking Cholesky in 32-column block steps on the matrix extended
// by the right-hand side as row n (so the forward substitution is the same recurrence as the panel solve):
//   1. the 32 x 32 diagonal block goes to shared memory and is factorised by warp 0 in registers (lane = row, as in
//      spd_solve_reg_kernel; a ragged last block is padded with the identity);
//   2. panel: one thread per row below the block solves x L_kk^T = a (32 entries in registers, L_kk rows as broadcasts),
//      writes x back into the matrix (the factor) and, transposed, into the shared-memory panel PT[m][row];
//   3. trailing update A[i][j] -= sum_m P[i][m] P[j][m] on 4 x 4 tiles of the lower triangle, two LDS.128 per operand and m.
// Then L^T u = y in 32-row blocks from the bottom: warp 0 solves the diagonal block with shuffles, every thread i above
// it subtracts its 32 products (factor rows read with unit stride across threads).
// The matrix is staged in shared memory when it fits (n <= 127: read once, left untouched); larger systems are factorised
// in place in device memory (they stay L2 resident: 0.67 MB per CTA at n = 289), as the C-ABI contract says.
constexpr int kCholNB = 32, kCholLD = 33;

__device__ __forceinline__ void chol_diag32(double *D, double *rinv_out, int lane)
{
    // D[32][33] (lower triangle valid, padded with the identity) -> L in place (row major), 1 / L[k][k] to rinv_out[k]
    double a[kCholNB];
#pragma unroll
    for (int j = 0; j < kCholNB; ++j)
        a[j] = D[lane * kCholLD + j];
    __syncwarp();
    double rinv = 0.0;
#pragma unroll
    for (int k = 0; k < kCholNB; ++k)
    {
        const double r = rsqrt(__shfl_sync(0xffffffffu, a[k], k));
        a[k] *= r;
        if (lane == k)
            rinv = r;
        if (k + 1 < kCholNB)
        {
            D[lane * kCholLD + k] = a[k]; // column k of L (rows < k keep scaled garbage of the upper triangle: never read)
            __syncwarp();
#pragma unroll
            for (int j = k + 1; j < kCholNB; ++j)
                a[j] = fma(-a[k], D[j * kCholLD + k], a[j]);
        }
        else
            D[lane * kCholLD + k] = a[k];
    }
    rinv_out[lane] = rinv;
    __syncwarp();
}

__global__ void __launch_bounds__(256, 2)
    spd_solve_blocked_kernel(int n, int lda, bool staged, size_t E, double *__restrict__ mass, double *__restrict__ rhs,
                             int *__restrict__ status)
{
    extern __shared__ __align__(16) double chol_blk_smem[];
    const int tid = threadIdx.x, lane = tid & 31, nthr = blockDim.x;
    const int RP = (n + 1 + 3) & ~3; // panel row capacity (rows below the first block + the right-hand side), 4-aligned
    double *D = chol_blk_smem;                    // [32][33]
    double *PT = chol_blk_smem + ((kCholNB * kCholLD + 1) & ~1); // [32][RP] transposed panel, 16-byte aligned
    double *rinv = PT + kCholNB * RP;             // [n + 32]
    double *y = rinv + ((n + kCholNB + 1) & ~1);  // [n + 32]
    double *As = y + ((n + kCholNB + 1) & ~1);    // [(n + 1)][lda] when staged
    for (size_t e = blockIdx.x; e < E; e += gridDim.x)
    {
        double *Ag = mass + e * (size_t)n * n;
        double *bg = rhs + e * (size_t)n;
        double *A = staged ? As : Ag;
        double *brow = staged ? As + (size_t)n * lda : bg; // row n of the extended matrix
        __syncthreads();
        const double sg = Ag[0] < 0.0 ? -1.0 : 1.0; // negative definite -> factorise -M, solve with -b
        if (staged)
        {
            for (int idx = tid, i = 0, j = tid; idx < n * n; idx += nthr, j += nthr)
            {
                while (j >= n)
                    j -= n, ++i;
                As[i * lda + j] = sg * __ldcs(Ag + idx);
            }
            for (int j = tid; j < n; j += nthr)
                brow[j] = sg * bg[j];
        }
        else if (sg < 0.0) // in place (the matrix is overwritten by its factor anyway)
        {
            __syncthreads(); // every thread has read Ag[0]
            for (int idx = tid; idx < n * n; idx += nthr)
                Ag[idx] = -Ag[idx];
            for (int j = tid; j < n; j += nthr)
                bg[j] = -bg[j];
        }
        __syncthreads();
        for (int kb = 0; kb < n; kb += kCholNB)
        {
            const int nb = n - kb < kCholNB ? n - kb : kCholNB;
            const int k1 = kb + nb;      // first trailing row / column
            const int R = n + 1 - k1;    // rows below the block, the right-hand side included
            // 1. diagonal block (identity padding for a ragged block)
            for (int idx = tid; idx < kCholNB * kCholNB; idx += nthr)
            {
                const int i = idx >> 5, j = idx & 31;
                D[i * kCholLD + j] = (i < nb && j < nb) ? A[(size_t)(kb + i) * lda + kb + j] : (i == j ? 1.0 : 0.0);
            }
            __syncthreads();
            if (tid < 32)
                chol_diag32(D, rinv + kb, lane);
            __syncthreads();
            // the factor's diagonal block back into the matrix (the back substitution reads it there)
            for (int idx = tid; idx < nb * nb; idx += nthr)
            {
                const int i = idx / nb, j = idx - i * nb;
                if (j <= i)
                    A[(size_t)(kb + i) * lda + kb + j] = D[i * kCholLD + j];
            }
            // 2. panel rows k1 .. n (row n = right-hand side)
            for (int r = tid; r < R; r += nthr)
            {
                const int i = k1 + r;
                double *row = (i < n ? A + (size_t)i * lda : brow) + kb;
                double xv[kCholNB];
#pragma unroll
                for (int j = 0; j < kCholNB; ++j)
                    xv[j] = j < nb ? row[j] : 0.0;
#pragma unroll
                for (int j = 0; j < kCholNB; ++j)
                {
                    double sacc = xv[j];
#pragma unroll
                    for (int m = 0; m < j; ++m)
                        sacc = fma(-xv[m], D[j * kCholLD + m], sacc);
                    xv[j] = sacc * rinv[kb + j];
                }
#pragma unroll
                for (int j = 0; j < kCholNB; ++j)
                {
                    if (j < nb)
                        row[j] = xv[j];
                    PT[j * RP + r] = xv[j];
                }
            }
            __syncthreads();
            // 3. trailing update on 4 x 4 tiles (ti >= tj) of rows/columns k1 ..; row n is the right-hand side
            const int nt = (R + 3) >> 2;
            const int n_tiles = nt * (nt + 1) / 2;
            for (int t = tid; t < n_tiles; t += nthr)
            {
                int ti = (int)((sqrtf(8.0f * (float)t + 1.0f) - 1.0f) * 0.5f);
                while ((ti + 1) * (ti + 2) / 2 <= t)
                    ++ti;
                while (ti * (ti + 1) / 2 > t)
                    --ti;
                const int tj = t - ti * (ti + 1) / 2;
                // the tile of A is requested first: its latency (L2 or HBM for the in-place mode) hides behind the 32 steps
                double acc[4][4];
#pragma unroll
                for (int a = 0; a < 4; ++a)
                {
                    const int i = k1 + 4 * ti + a;
                    const double *row = i < n ? A + (size_t)i * lda : brow;
#pragma unroll
                    for (int b = 0; b < 4; ++b)
                    {
                        const int j = k1 + 4 * tj + b;
                        acc[a][b] = (i <= n && j < n && j <= i) ? row[j] : 0.0;
                    }
                }
                const double *pi = PT + 4 * ti, *pj = PT + 4 * tj;
#pragma unroll 8
                for (int m = 0; m < kCholNB; ++m)
                {
                    const double2 a01 = *reinterpret_cast<const double2 *>(pi + m * RP);
                    const double2 a23 = *reinterpret_cast<const double2 *>(pi + m * RP + 2);
                    const double2 b01 = *reinterpret_cast<const double2 *>(pj + m * RP);
                    const double2 b23 = *reinterpret_cast<const double2 *>(pj + m * RP + 2);
                    const double av[4] = {a01.x, a01.y, a23.x, a23.y}, bv[4] = {b01.x, b01.y, b23.x, b23.y};
#pragma unroll
                    for (int a = 0; a < 4; ++a)
#pragma unroll
                        for (int b = 0; b < 4; ++b)
                            acc[a][b] = fma(-av[a], bv[b], acc[a][b]);
                }
#pragma unroll
                for (int a = 0; a < 4; ++a)
                {
                    const int i = k1 + 4 * ti + a;
                    if (i > n)
                        continue;
                    double *row = i < n ? A + (size_t)i * lda : brow;
#pragma unroll
                    for (int b = 0; b < 4; ++b)
                    {
                        const int j = k1 + 4 * tj + b;
                        if (j < n && j <= i)
                            row[j] = acc[a][b];
                    }
                }
            }
            __syncthreads();
        }
        // y = L^-1 b sits in row n;  L^T u = y from the bottom block up
        for (int j = tid; j < n; j += nthr)
            y[j] = brow[j];
        __syncthreads();
        for (int kb = ((n - 1) / kCholNB) * kCholNB; kb >= 0; kb -= kCholNB)
        {
            const int nb = n - kb < kCholNB ? n - kb : kCholNB;
            // the factor's diagonal block into shared memory first: warp 0's 32 dependent steps then wait on LDS, not on L2
            for (int idx = tid; idx < nb * nb; idx += nthr)
            {
                const int i = idx / nb, j = idx - i * nb;
                if (j < i)
                    D[i * kCholLD + j] = A[(size_t)(kb + i) * lda + kb + j];
            }
            __syncthreads();
            if (tid < 32)
            {
                double yv = lane < nb ? y[kb + lane] : 0.0;
                const double ri = lane < nb ? rinv[kb + lane] : 0.0;
                for (int k = nb - 1; k >= 0; --k)
                {
                    const double uk = __shfl_sync(0xffffffffu, yv * ri, k);
                    if (lane == k)
                        yv = uk;
                    else if (lane < k)
                        yv = fma(-D[k * kCholLD + lane], uk, yv);
                }
                if (lane < nb)
                    y[kb + lane] = yv;
            }
            __syncthreads();
            for (int i = tid; i < kb; i += nthr)
            {
                double yv = y[i];
                for (int k = 0; k < nb; ++k)
                    yv = fma(-A[(size_t)(kb + k) * lda + i], y[kb + k], yv);
                y[i] = yv;
            }
            __syncthreads();
        }
        for (int j = tid; j < n; j += nthr)
        {
            bg[j] = y[j];
            spd_flag_if_bad(y[j], status);
        }
    }
}

// ---- projection norms: deterministic two-pass dot product sum_i a[i]*b[i] (fixed block -> partial mapping) ----
constexpr int kDotBlocks = 592; // 4 CTAs per SM on 148 SMs

__global__ void __launch_bounds__(kThreads) dot_partial_kernel(size_t n, const double *__restrict__ a, const double *__restrict__ b,
                                                               double *__restrict__ partial)
{
    __shared__ double red[kThreads / 32];
    double acc = 0.0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        acc = fma(a[i], b[i], acc);
    for (int off = 16; off > 0; off >>= 1)
        acc += __shfl_down_sync(0xffffffffu, acc, off);
    if ((threadIdx.x & 31) == 0)
        red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        double t = 0.0;
        for (int w = 0; w < kThreads / 32; ++w)
            t += red[w];
        partial[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(32) dot_final_kernel(int n_partial, const double *__restrict__ partial, double *__restrict__ out)
{
    double acc = 0.0;
    for (int i = threadIdx.x; i < n_partial; i += 32)
        acc += partial[i];
    for (int off = 16; off > 0; off >>= 1)
        acc += __shfl_down_sync(0xffffffffu, acc, off);
    if (threadIdx.x == 0)
        out[0] = acc;
}

// ---- lower-triangle mirroring of square per-element matrices, in place: out[e][r][c] = out[e][c][r] for r < c ----
// The reference fills a matching-space gradient mass matrix from its lower triangle (mass_matrices.c:494-495, :551-560)
// although that matrix is not symmetric.  The sum-factorised kernels compute the true full matrix; this pass then
// overwrites the strict upper triangle with the transposed lower one.  One 32 x 32 tile pair per CTA iteration: the
// lower tile is read row-wise (coalesced), transposed through shared memory and written row-wise into its mirror tile.
__global__ void __launch_bounds__(256) mirror_lower_kernel(size_t E, int N, double *__restrict__ out)
{
    __shared__ double tile[32][33];
    const int nt = (N + 31) / 32;
    const size_t pairs = (size_t)nt * (nt + 1) / 2, total = E * pairs;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (size_t w = blockIdx.x; w < total; w += gridDim.x)
    {
        const size_t e = w / pairs;
        const int p = (int)(w % pairs);
        int tr = 0;
        while ((tr + 1) * (tr + 2) / 2 <= p)
            ++tr;
        const int tc = p - tr * (tr + 1) / 2; // tile (tr, tc) of the lower triangle, tc <= tr
        double *m = out + e * (size_t)N * N;
        for (int r = ty; r < 32; r += 8)
        {
            const int gr = tr * 32 + r, gc = tc * 32 + tx;
            if (gr < N && gc < N)
                tile[r][tx] = m[(size_t)gr * N + gc];
        }
        __syncthreads();
        for (int r = ty; r < 32; r += 8)
        {
            const int gr = tc * 32 + r, gc = tr * 32 + tx; // entry of the mirror tile (tc, tr)
            if (gr < N && gc < N && gr < gc)
                m[(size_t)gr * N + gc] = tile[tx][r];
        }
        __syncthreads();
    }
}

// Small matrices (N < 64: at most 32 KB each): one warp per element, every lane copies entries (c, r) -> (r, c) of the
// strict upper triangle straight in global memory.  The strided reads stay inside the element's few KB, which the L1 serves
// after the first touch; reads (lower triangle) and writes (upper triangle) never overlap, so no barrier is needed.
__global__ void __launch_bounds__(256) mirror_lower_small_kernel(size_t E, int N, double *__restrict__ out)
{
    const int lane = threadIdx.x & 31;
    const size_t warp = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5, n_warps = ((size_t)gridDim.x * blockDim.x) >> 5;
    const int NN = N * N;
    for (size_t e = warp; e < E; e += n_warps)
    {
        double *m = out + e * (size_t)NN;
        for (int t = lane; t < NN; t += 32)
        {
            const int r = t / N, c = t - r * N;
            if (r < c)
                m[t] = m[c * N + r];
        }
    }
}

int grid_for(size_t total_threads, int sm_count)
{
    const size_t blocks = (total_threads + kThreads - 1) / kThreads;
    const size_t cap = (size_t)sm_count * 16; // grid-stride beyond 16 resident CTAs per SM
    return (int)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
}

int check_launch(const char *what)
{
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess)
        return ib200_fail(IB200_ERROR_CUDA, "launch of %s failed: %s", what, cudaGetErrorString(err));
    return INTERP_SUCCESS;
}

} // namespace

int launch_reconstruct_generic(const LaunchCfg &cfg, const OpDesc &op, size_t E, const double *dofs, unsigned n_fields,
                               const unsigned *masks, double *out)
{
    if (n_fields > 8)
        return ib200_fail(IB200_ERROR_UNSUPPORTED, "at most 8 fields per reconstruct call (got %u)", n_fields);
    MaskList ml{};
    ml.n = n_fields;
    for (unsigned f = 0; f < n_fields; ++f)
        ml.mask[f] = masks[f];
    if (E == 0 || n_fields == 0)
        return INTERP_SUCCESS;
    reconstruct_generic_kernel<<<grid_for(E * n_fields * op.Q, cfg.sm_count), kThreads, 0, cfg.stream>>>(op, E, dofs, ml, out);
    ++*cfg.launch_counter;
    return check_launch("reconstruct_generic_kernel");
}

int launch_geometry_generic(const LaunchCfg &cfg, const OpDesc &op, unsigned n_phys, size_t E, const double *geom_dofs,
                            double *coords, double *det, double *jinv, double *weight_out, int w_idx_in, int w_idx_out)
{
    if (E == 0)
        return INTERP_SUCCESS;
    geometry_generic_kernel<<<grid_for(E * op.Q, cfg.sm_count), kThreads, 0, cfg.stream>>>(
        op, (int)n_phys, E, geom_dofs, coords, det, jinv, weight_out, w_idx_in, w_idx_out);
    ++*cfg.launch_counter;
    return check_launch("geometry_generic_kernel");
}

int launch_mass_generic(const LaunchCfg &cfg, const OpDesc &op, bool symmetric, size_t E, const double *weight,
                        int deriv_dim_in, double *out)
{
    if (E == 0)
        return INTERP_SUCCESS;
    mass_generic_kernel<<<grid_for(E * op.N_out * op.N_in, cfg.sm_count), kThreads, 0, cfg.stream>>>(
        op, symmetric ? 1 : 0, E, weight, deriv_dim_in, out);
    ++*cfg.launch_counter;
    return check_launch("mass_generic_kernel");
}

int launch_l2_generic(const LaunchCfg &cfg, const OpDesc &op, size_t E, const double *fvals, const double *det, double *out)
{
    if (E == 0)
        return INTERP_SUCCESS;
    l2_generic_kernel<<<grid_for(E * op.N_in, cfg.sm_count), kThreads, 0, cfg.stream>>>(op, E, fvals, det, out);
    ++*cfg.launch_counter;
    return check_launch("l2_generic_kernel");
}

int launch_mirror_lower(const LaunchCfg &cfg, size_t E, int N, double *out)
{
    if (E == 0 || N < 2)
        return INTERP_SUCCESS;
    if (N < 64)
    {
        const size_t blocks = (E + 7) / 8, cap = (size_t)cfg.sm_count * 8;
        mirror_lower_small_kernel<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, cfg.stream>>>(E, N, out);
        ++*cfg.launch_counter;
        return check_launch("mirror_lower_small_kernel");
    }
    const int nt = (N + 31) / 32;
    const size_t total = E * ((size_t)nt * (nt + 1) / 2), cap = (size_t)cfg.sm_count * 8;
    mirror_lower_kernel<<<(unsigned)(total < cap ? total : cap), 256, 0, cfg.stream>>>(E, N, out);
    ++*cfg.launch_counter;
    return check_launch("mirror_lower_kernel");
}

int dot_scratch_doubles()
{
    return kDotBlocks;
}

int launch_dot(const LaunchCfg &cfg, size_t n, const double *a, const double *b, double *partial, double *out)
{
    dot_partial_kernel<<<kDotBlocks, kThreads, 0, cfg.stream>>>(n, a, b, partial);
    ++*cfg.launch_counter;
    IB_TRY(check_launch("dot_partial_kernel"));
    dot_final_kernel<<<1, 32, 0, cfg.stream>>>(kDotBlocks, partial, out);
    ++*cfg.launch_counter;
    return check_launch("dot_final_kernel");
}

// ---- ib200_dot_static: the same two-pass dot product for callers that must not allocate ---------------------------
// ib200_dot keeps its 592 block partials (and, for host vectors, a staged copy of a and b) in memory the context
// obtains from cudaMalloc.  Some users of this library cannot afford that call: a host application whose own pool
// allocator has already claimed the GPU (a raw cudaMalloc then fails, or punches a hole into the pool's accounting),
// and a time-stepping loop on a GPU shared with other streams (cudaMalloc / cudaFree synchronise the whole device).
// For them the scratch below is part of the loaded module: fixed size, present on every GPU that loads the library
// from module load until the process exits, never allocated or freed by any call.
//
// Layout (doubles): [0, 592) block partials | [592] result slot of a host-vector call | [1024, 1024 + MAX) staged a |
// [1024 + MAX, 1024 + 2 MAX) staged b, MAX = IB200_DOT_STATIC_MAX_HOST_N.  Nothing has to start at zero: every launch
// overwrites all 592 partials and the result slot before reading them, and the staged vectors are copied in before the
// kernel reads them, so what the previous call left behind is never observed.
constexpr size_t kDotStaticMaxN = IB200_DOT_STATIC_MAX_HOST_N;
constexpr size_t kDotStaticHead = 1024; // partials + result slot, padded so that the staged vectors stay 256-byte aligned
static_assert(kDotBlocks + 1 <= kDotStaticHead, "partials and result slot must fit the head of the static scratch");
static __device__ __align__(256) double g_dot_static_scratch[kDotStaticHead + 2 * kDotStaticMaxN];

namespace
{
// In host code `g_dot_static_scratch` is only a handle for the runtime, not an address a kernel could use: the device
// address is looked up once per GPU (with that GPU current) and remembered here.
constexpr int kDotStaticMaxDevices = 64;
std::mutex g_dot_static_mutex;
double *g_dot_static_ptr[kDotStaticMaxDevices] = {};

int dot_static_scratch(double **out)
{
    int dev = -1;
    IB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kDotStaticMaxDevices)
        return ib200_fail(IB200_ERROR_UNSUPPORTED, "ib200_dot_static: device ordinal %d is outside the %d this table holds", dev,
                          kDotStaticMaxDevices);
    if (!g_dot_static_ptr[dev])
    {
        void *p = nullptr;
        IB_CUDA(cudaGetSymbolAddress(&p, g_dot_static_scratch));
        if (!p)
            return ib200_fail(IB200_ERROR_CUDA, "ib200_dot_static: the runtime returned no address for the static scratch");
        g_dot_static_ptr[dev] = (double *)p;
    }
    *out = g_dot_static_ptr[dev];
    return INTERP_SUCCESS;
}
} // namespace

extern "C" int ib200_dot_static(size_t n, const double *a, const double *b, double *out, int where, void *cuda_stream)
{
    if (!out || (n && (!a || !b)))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "ib200_dot_static: a/b/out must not be NULL");
    if (where != IB200_DEVICE && where != IB200_HOST)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "ib200_dot_static: `where` must be IB200_HOST or IB200_DEVICE");
    if (where == IB200_HOST && n > kDotStaticMaxN)
        return ib200_fail(IB200_ERROR_UNSUPPORTED,
                          "ib200_dot_static: %zu host values per vector exceed the %zu the built-in scratch can stage "
                          "(IB200_DOT_STATIC_MAX_HOST_N); pass device pointers or use ib200_dot",
                          n, kDotStaticMaxN);
    // one host thread at a time: the scratch is shared by every caller in the process
    std::lock_guard<std::mutex> lock(g_dot_static_mutex);
    double *scratch = nullptr;
    IB_TRY(dot_static_scratch(&scratch));
    const cudaStream_t st = (cudaStream_t)cuda_stream;
    double *partial = scratch, *result = scratch + kDotBlocks;
    const double *da = a, *db = b;
    if (where == IB200_HOST)
    {
        double *sa = scratch + kDotStaticHead, *sb = sa + kDotStaticMaxN;
        if (n)
        {
            IB_CUDA(cudaMemcpyAsync(sa, a, n * sizeof(double), cudaMemcpyHostToDevice, st));
            IB_CUDA(cudaMemcpyAsync(sb, b, n * sizeof(double), cudaMemcpyHostToDevice, st));
        }
        da = sa;
        db = sb;
    }
    // same kernels, grid and reduction tree as launch_dot: the result is bit-identical to ib200_dot's
    dot_partial_kernel<<<kDotBlocks, kThreads, 0, st>>>(n, da, db, partial);
    IB_TRY(check_launch("dot_partial_kernel (static scratch)"));
    dot_final_kernel<<<1, 32, 0, st>>>(kDotBlocks, partial, where == IB200_HOST ? result : out);
    IB_TRY(check_launch("dot_final_kernel (static scratch)"));
    if (where == IB200_HOST)
    {
        IB_CUDA(cudaMemcpyAsync(out, result, sizeof(double), cudaMemcpyDeviceToHost, st));
        IB_CUDA(cudaStreamSynchronize(st));
    }
    return INTERP_SUCCESS;
}

int launch_spd_solve(const LaunchCfg &cfg, unsigned n, size_t E, double *mass, double *rhs)
{
    if (E == 0 || n == 0)
        return INTERP_SUCCESS;
    if (n <= 32 && !getenv("IB200_SPD_NO_REG"))
    {
#define IB_SPD_CASE(NN)                                                                                                         \
    case NN:                                                                                                                    \
        launch_spd_solve_reg<NN>(cfg, E, mass, rhs);                                                                             \
        break;
        switch (n)
        {
            IB_SPD_CASE(1) IB_SPD_CASE(2) IB_SPD_CASE(3) IB_SPD_CASE(4) IB_SPD_CASE(5) IB_SPD_CASE(6) IB_SPD_CASE(7) IB_SPD_CASE(8)
            IB_SPD_CASE(9) IB_SPD_CASE(10) IB_SPD_CASE(11) IB_SPD_CASE(12) IB_SPD_CASE(13) IB_SPD_CASE(14) IB_SPD_CASE(15)
            IB_SPD_CASE(16) IB_SPD_CASE(17) IB_SPD_CASE(18) IB_SPD_CASE(19) IB_SPD_CASE(20) IB_SPD_CASE(21) IB_SPD_CASE(22)
            IB_SPD_CASE(23) IB_SPD_CASE(24) IB_SPD_CASE(25) IB_SPD_CASE(26) IB_SPD_CASE(27) IB_SPD_CASE(28) IB_SPD_CASE(29)
            IB_SPD_CASE(30) IB_SPD_CASE(31) IB_SPD_CASE(32)
        }
#undef IB_SPD_CASE
        ++*cfg.launch_counter;
        return check_launch("spd_solve_reg_kernel");
    }
    if (n > 32 && n <= 64 && !getenv("IB200_SPD_NO_REG2"))
    {
        const int smem = (int)sizeof(double) * 4 * (2 * 32 * kReg2LD + 32 * kReg2XLD + 32);
        IB_CUDA(cudaFuncSetAttribute(spd_solve_reg2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        int per_sm = 0;
        IB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, spd_solve_reg2_kernel, 128, smem));
        const size_t blocks = std::min<size_t>((E + 3) / 4, (size_t)cfg.sm_count * std::max(per_sm, 1));
        spd_solve_reg2_kernel<<<(int)blocks, 128, smem, cfg.stream>>>((int)n, E, mass, rhs, cfg.status);
        ++*cfg.launch_counter;
        return check_launch("spd_solve_reg2_kernel");
    }
    const bool staged = n <= 127;
    const int lda = staged ? ((int)n | 1) : (int)n;
    const int RP = ((int)n + 1 + 3) & ~3, vec = ((int)n + kCholNB + 1) & ~1;
    const size_t smem_blocked = sizeof(double) * (((kCholNB * kCholLD + 1) & ~1) + (size_t)kCholNB * RP + 2 * vec +
                                                  (staged ? (size_t)(n + 1) * lda : 0));
    // the transposed panel of the blocked kernel holds 32 x (n + 1) doubles: systems beyond n ~ 740 do not fit 200 KB of
    // shared memory and take the warp-per-element in-place kernel below
    if (smem_blocked <= 200 * 1024 && !getenv("IB200_SPD_NO_BLOCKED"))
    {
        const size_t smem = smem_blocked;
        const int threads = n <= 64 ? 128 : 256;
        IB_CUDA(cudaFuncSetAttribute(spd_solve_blocked_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        int per_sm = 0;
        IB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, spd_solve_blocked_kernel, threads, smem));
        const size_t grid = std::min<size_t>(E, (size_t)cfg.sm_count * std::max(per_sm, 1));
        spd_solve_blocked_kernel<<<(int)grid, threads, smem, cfg.stream>>>((int)n, lda, staged, E, mass, rhs, cfg.status);
        ++*cfg.launch_counter;
        return check_launch("spd_solve_blocked_kernel");
    }
    const size_t warps = E;
    size_t blocks = (warps * 32 + 127) / 128;
    const int ld = (int)n | 1;
    int wpc = 4; // warps per CTA: as many as 200 KB of shared memory hold
    while (wpc > 1 && sizeof(double) * wpc * (n + 1) * ld > 200 * 1024)
        wpc /= 2;
    const size_t smem = sizeof(double) * wpc * (n + 1) * ld;
    if (n + 1 <= 128 && smem <= 200 * 1024)
    {
        blocks = (warps + wpc - 1) / wpc;
        // shared-memory resident factorisation (the matrix is left untouched)
        // per launch, not once per process: the attribute is per device and a process may hold contexts on several
        IB_CUDA(cudaFuncSetAttribute(spd_solve_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        int per_sm = 0;
        IB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, spd_solve_smem_kernel, 32 * wpc, smem));
        const size_t cap = (size_t)cfg.sm_count * std::max(per_sm, 1);
        if (blocks > cap)
            blocks = cap;
        spd_solve_smem_kernel<<<(int)blocks, 32 * wpc, smem, cfg.stream>>>((int)n, ld, E, mass, rhs, cfg.status);
        ++*cfg.launch_counter;
        return check_launch("spd_solve_smem_kernel");
    }
    const size_t cap = (size_t)cfg.sm_count * 16;
    if (blocks > cap)
        blocks = cap;
    spd_solve_kernel<<<(int)blocks, 128, 0, cfg.stream>>>((int)n, E, mass, rhs, cfg.status);
    ++*cfg.launch_counter;
    return check_launch("spd_solve_kernel");
}
