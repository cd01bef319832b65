// N2 -- tensor-product basis evaluation and reconstruction at ARBITRARY reference points, element batched.
//
//   evaluate:     out[e][p][i]   = ((1 * b0_{i0}(x0)) * b1_{i1}(x1)) * ...          i = (i0,i1,..) row-major
//   reconstruct:  out[e][f][p]   = sum_i dofs[e][i] * prod_k T^k_f,{i_k}(x_k[e][p]),    T^k_f = d/dxi_k basis if bit k of
//                                                                                      masks[f] is set, else the basis
//
// replaces FunctionSpace.evaluate (/root/reference/src/python/function_space_objects.c:142-276 ->
// basis_compute_outer_product_basis, src/basis/basis_set.c:577-633, 1-D evaluators :462-556) and
// reconstruct() (/root/reference/python/interplib/degrees_of_freedom.py:10-37) which the reference runs once per
// element on the host.  Points are given per element and dimension: x[e][k][p].
//
// One thread per (element, point): the 1-D values (and derivatives) of every dimension are formed in registers / local
// memory in the operation order of the reference's evaluators (Bonnet recursion, degree-raising Bernstein recursion,
// Lagrange numerator product in ascending node order divided by the host-built denominator), the contraction with the
// element's DoFs is sum-factorised per point, DoFs are read through the read-only path (the threads of a block share
// an element).  Not a roofline-tuned kernel: it is the caller-side companion of K1 (sampling geometry, plotting).
#include <algorithm>
#include <cstdint>
#include <cstdlib>

#include "ctx.cuh"

namespace
{

constexpr int PT_MAXN = 24; // 1-D basis functions per dimension (p <= 23)

struct PointMasks
{
    unsigned m[8];
};

__device__ void pt_basis_1d(const PointDim &pd, double x, double *v, double *dv, bool want_d)
{
    const int n = pd.n;
    if (pd.type == BASIS_LEGENDRE)
    {
        // legendre_eval_bonnet_all, src/polynomials/legendre.c:60-85; derivative recurrence of basis_set.c:494-513
        v[0] = 1.0;
        if (n > 1)
            v[1] = x;
        double lo = 1.0, hi = x;
        for (int k = 2; k < n; ++k)
        {
            const double a = __dmul_rn((double)(2 * k - 1), x);
            const double next = __ddiv_rn(__dsub_rn(__dmul_rn(a, hi), __dmul_rn((double)(k - 1), lo)), (double)k);
            v[k] = next;
            lo = hi;
            hi = next;
        }
        if (want_d)
        {
            double d = 0.0;
            dv[0] = 0.0;
            for (int i = 1; i < n; ++i)
            {
                d = __dadd_rn(__dmul_rn((double)i, v[i - 1]), __dmul_rn(x, d));
                dv[i] = d;
            }
        }
        return;
    }
    if (pd.type == BASIS_BERNSTEIN)
    {
        // bernstein_interpolation_vector, src/polynomials/bernstein.c:31-51 at t = (x + 1) / 2
        const double t = __dmul_rn(__dadd_rn(x, 1.0), 0.5);
        const double a = __dsub_rn(1.0, t), b = t;
        const int deg = n - 1;
        if (want_d)
        {
            // degree deg-1 polynomials first (they give the derivative), then one more raising step for the values
            if (deg == 0)
            {
                v[0] = 1.0;
                dv[0] = 0.0;
                return;
            }
            double w[PT_MAXN];
            w[0] = 1.0;
            for (int i = 0; i < deg - 1; ++i)
            {
                w[i + 1] = __dmul_rn(w[i], b);
                for (int j = i; j > 0; --j)
                    w[j] = __dadd_rn(__dmul_rn(b, w[j - 1]), __dmul_rn(a, w[j]));
                w[0] = __dmul_rn(w[0], a);
            }
            v[deg] = __dmul_rn(w[deg - 1], b);
            for (int j = deg - 1; j > 0; --j)
                v[j] = __dadd_rn(__dmul_rn(b, w[j - 1]), __dmul_rn(a, w[j]));
            v[0] = __dmul_rn(w[0], a);
            dv[deg] = __dmul_rn((double)deg, w[deg - 1]);
            for (int j = deg - 1; j > 0; --j)
                dv[j] = __dmul_rn((double)deg, __dsub_rn(w[j - 1], w[j]));
            dv[0] = __dmul_rn(w[0], -(double)deg);
            for (int j = 0; j < n; ++j)
                dv[j] = __dmul_rn(dv[j], 0.5);
            return;
        }
        v[0] = 1.0;
        for (int i = 0; i < deg; ++i)
        {
            v[i + 1] = __dmul_rn(v[i], b);
            for (int j = i; j > 0; --j)
                v[j] = __dadd_rn(__dmul_rn(b, v[j - 1]), __dmul_rn(a, v[j]));
            v[0] = __dmul_rn(v[0], a);
        }
        return;
    }
    // Lagrange family: lagrange_polynomial_values_2 / _first_derivative_2, src/polynomials/lagrange.c:108-183, 383-525
    double dif[PT_MAXN];
    for (int j = 0; j < n; ++j)
        dif[j] = __dsub_rn(x, pd.roots[j]);
    for (int j = 0; j < n; ++j)
    {
        // value: product of the differences in ascending node order, divided by the denominator last;
        // derivative: the same running product carried together with its x-derivative
        double acc = 1.0, der = 0.0;
        for (int k = 0; k < n; ++k)
        {
            if (k == j)
                continue;
            der = fma(der, dif[k], acc);
            acc = __dmul_rn(acc, dif[k]);
        }
        v[j] = n == 1 ? 1.0 : __ddiv_rn(acc, pd.den[j]);
        if (want_d)
            dv[j] = n == 1 ? 0.0 : __ddiv_rn(der, pd.den[j]);
    }
}

__global__ void __launch_bounds__(128)
    points_evaluate_kernel(const PointOp op, size_t E, size_t P, const double *__restrict__ x, double *__restrict__ out)
{
    const size_t total = E * P;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x)
    {
        const size_t e = idx / P, p = idx - e * P;
        double b[IB200_POINT_MAX_DIM][PT_MAXN];
        for (int k = 0; k < op.ndim; ++k)
            pt_basis_1d(op.dim[k], x[(e * op.ndim + k) * P + p], b[k], nullptr, false);
        double *o = out + idx * (size_t)op.N;
        const int n1 = op.ndim > 1 ? op.dim[1].n : 1, n2 = op.ndim > 2 ? op.dim[2].n : 1;
        for (int i0 = 0; i0 < op.dim[0].n; ++i0)
        {
            const double v0 = __dmul_rn(1.0, b[0][i0]);
            for (int i1 = 0; i1 < n1; ++i1)
            {
                const double v1 = op.ndim > 1 ? __dmul_rn(v0, b[1][i1]) : v0;
                for (int i2 = 0; i2 < n2; ++i2)
                    o[((size_t)i0 * n1 + i1) * n2 + i2] = op.ndim > 2 ? __dmul_rn(v1, b[2][i2]) : v1;
            }
        }
    }
}

__global__ void __launch_bounds__(128)
    points_reconstruct_kernel(const PointOp op, size_t E, size_t P, const double *__restrict__ dofs, const double *__restrict__ x,
                              int n_fields, const PointMasks masks, bool any_deriv, double *__restrict__ out)
{
    // a block walks the points of ONE element at a time (blockIdx.y strides over elements) so that the DoF reads of
    // its threads are broadcasts
    for (size_t e = blockIdx.y; e < E; e += gridDim.y)
    {
        const double *c = dofs + e * (size_t)op.N;
        for (size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x; p < P; p += (size_t)gridDim.x * blockDim.x)
        {
            double b[IB200_POINT_MAX_DIM][PT_MAXN], d[IB200_POINT_MAX_DIM][PT_MAXN];
            for (int k = 0; k < op.ndim; ++k)
                pt_basis_1d(op.dim[k], x[(e * op.ndim + k) * P + p], b[k], d[k], any_deriv);
            const int n0 = op.dim[0].n, n1 = op.ndim > 1 ? op.dim[1].n : 1, n2 = op.ndim > 2 ? op.dim[2].n : 1;
            for (int f = 0; f < n_fields; ++f)
            {
                const unsigned m = masks.m[f];
                const double *t0 = (m & 1u) ? d[0] : b[0];
                const double *t1 = (m & 2u) ? d[1] : b[1];
                const double *t2 = (m & 4u) ? d[2] : b[2];
                double acc0 = 0.0;
                for (int i0 = 0; i0 < n0; ++i0)
                {
                    double acc1 = 0.0;
                    for (int i1 = 0; i1 < n1; ++i1)
                    {
                        const double *row = c + ((size_t)i0 * n1 + i1) * n2;
                        double acc2 = 0.0;
                        if (op.ndim > 2)
                        {
                            for (int i2 = 0; i2 < n2; ++i2)
                                acc2 = fma(__ldg(row + i2), t2[i2], acc2);
                        }
                        else
                            acc2 = __ldg(row);
                        acc1 = op.ndim > 1 ? fma(acc2, t1[i1], acc1) : acc2;
                    }
                    acc0 = fma(acc1, t0[i0], acc0);
                }
                out[(e * n_fields + f) * P + p] = acc0;
            }
        }
    }
}

// ---- cooperative variants (the ones normally launched) ---------------------------------------------------------------
// A CTA takes one element and a chunk of PB of its points at a time.  Phase 1 tabulates the 1-D values (and derivatives)
// of every dimension at the chunk's points in shared memory -- one work item per (dimension, point, basis function) for
// the Lagrange family (the numerator product needs no scratch), per (dimension, point) for the recursions, which write
// their rows straight to shared memory; same operation order as pt_basis_1d.  Phase 2 contracts: one work item per
// (point, field) against the element's DoFs (staged in shared memory once per element) / per output entry of the tensor
// basis.  No per-thread local arrays (the thread-per-point kernels above keep 1.7 KB of them on the stack).
struct PointTabs
{
    int ld[IB200_POINT_MAX_DIM];  // padded row length of a point's table (odd)
    int off[IB200_POINT_MAX_DIM]; // offset of dimension k's value table in doubles; the derivative table follows it
    int total;                    // doubles of all tables for PB points
};

__device__ __forceinline__ PointTabs point_tabs_layout(const PointOp &op, int PB, bool with_der)
{
    PointTabs t{};
    int o = 0;
    for (int k = 0; k < op.ndim; ++k)
    {
        t.ld[k] = op.dim[k].n | 1;
        t.off[k] = o;
        o += (with_der ? 2 : 1) * PB * t.ld[k];
    }
    t.total = o;
    return t;
}

// phase 1 for the points [p0, p0 + np) of element e
__device__ void point_tabs_fill(const PointOp &op, const PointTabs &lay, int PB, bool with_der, double *tabs, const double *xe,
                                size_t P, size_t p0, int np)
{
    for (int k = 0; k < op.ndim; ++k)
    {
        const PointDim &pd = op.dim[k];
        const int n = pd.n, ld = lay.ld[k];
        double *val = tabs + lay.off[k], *der = val + PB * ld;
        const double *xk = xe + (size_t)k * P + p0;
        if (pd.type == BASIS_LEGENDRE || pd.type == BASIS_BERNSTEIN)
        {
            for (int p = threadIdx.x; p < np; p += blockDim.x)
            {
                const double x = xk[p];
                double *v = val + p * ld, *dv = der + p * ld;
                if (pd.type == BASIS_LEGENDRE)
                {
                    v[0] = 1.0;
                    if (n > 1)
                        v[1] = x;
                    double lo = 1.0, hi = x;
                    for (int i = 2; i < n; ++i)
                    {
                        const double a = __dmul_rn((double)(2 * i - 1), x);
                        const double next = __ddiv_rn(__dsub_rn(__dmul_rn(a, hi), __dmul_rn((double)(i - 1), lo)), (double)i);
                        v[i] = next;
                        lo = hi;
                        hi = next;
                    }
                    if (with_der)
                    {
                        double d = 0.0;
                        dv[0] = 0.0;
                        for (int i = 1; i < n; ++i)
                        {
                            d = __dadd_rn(__dmul_rn((double)i, v[i - 1]), __dmul_rn(x, d));
                            dv[i] = d;
                        }
                    }
                }
                else
                {
                    const double t = __dmul_rn(__dadd_rn(x, 1.0), 0.5);
                    const double a = __dsub_rn(1.0, t), b = t;
                    const int deg = n - 1;
                    // degree deg-1 polynomials in dv (they give the derivative), one more raising step for the values
                    double *w = with_der ? dv : v;
                    const int wdeg = with_der ? deg - 1 : deg;
                    w[0] = 1.0;
                    for (int i = 0; i < wdeg; ++i)
                    {
                        w[i + 1] = __dmul_rn(w[i], b);
                        for (int j = i; j > 0; --j)
                            w[j] = __dadd_rn(__dmul_rn(b, w[j - 1]), __dmul_rn(a, w[j]));
                        w[0] = __dmul_rn(w[0], a);
                    }
                    if (with_der)
                    {
                        if (deg == 0)
                        {
                            v[0] = 1.0;
                            dv[0] = 0.0;
                        }
                        else
                        {
                            v[deg] = __dmul_rn(w[deg - 1], b);
                            for (int j = deg - 1; j > 0; --j)
                                v[j] = __dadd_rn(__dmul_rn(b, w[j - 1]), __dmul_rn(a, w[j]));
                            v[0] = __dmul_rn(w[0], a);
                            // in place, descending: dv[j] needs the old w[j-1] and w[j] only
                            dv[deg] = __dmul_rn((double)deg, w[deg - 1]);
                            for (int j = deg - 1; j > 0; --j)
                                dv[j] = __dmul_rn((double)deg, __dsub_rn(w[j - 1], w[j]));
                            dv[0] = __dmul_rn(w[0], -(double)deg);
                            for (int j = 0; j < n; ++j)
                                dv[j] = __dmul_rn(dv[j], 0.5);
                        }
                    }
                }
            }
        }
        else
        {
            for (int item = threadIdx.x; item < np * n; item += blockDim.x)
            {
                const int p = item / n, j = item - p * n;
                const double x = xk[p];
                double acc = 1.0, dd = 0.0;
                for (int m = 0; m < n; ++m)
                {
                    if (m == j)
                        continue;
                    const double dif = __dsub_rn(x, pd.roots[m]);
                    dd = fma(dd, dif, acc);
                    acc = __dmul_rn(acc, dif);
                }
                // times the host-rounded reciprocal of the denominator (stored after roots and denominators): one
                // rounding more than the reference's division, a tenth of its instructions
                const double inv = pd.den[n + j];
                val[p * ld + j] = n == 1 ? 1.0 : __dmul_rn(acc, inv);
                if (with_der)
                    der[p * ld + j] = n == 1 ? 0.0 : __dmul_rn(dd, inv);
            }
        }
    }
}

__global__ void __launch_bounds__(128)
    points_reconstruct_coop_kernel(const PointOp op, int PB, size_t E, size_t P, const double *__restrict__ dofs,
                                   const double *__restrict__ x, int n_fields, const PointMasks masks, bool any_deriv,
                                   double *__restrict__ out)
{
    extern __shared__ __align__(16) double pt_smem[];
    const PointTabs lay = point_tabs_layout(op, PB, any_deriv);
    double *tabs = pt_smem, *c = pt_smem + lay.total;
    const int n0 = op.dim[0].n, n1 = op.ndim > 1 ? op.dim[1].n : 1, n2 = op.ndim > 2 ? op.dim[2].n : 1;
    for (size_t e = blockIdx.x; e < E; e += gridDim.x)
    {
        __syncthreads();
        for (int i = threadIdx.x; i < op.N; i += blockDim.x)
            c[i] = dofs[e * (size_t)op.N + i];
        const double *xe = x + e * (size_t)op.ndim * P;
        for (size_t p0 = (size_t)blockIdx.y * PB; p0 < P; p0 += (size_t)PB * gridDim.y)
        {
            const int np = (int)((P - p0) < (size_t)PB ? (P - p0) : (size_t)PB);
            __syncthreads(); // tables of the previous chunk consumed
            point_tabs_fill(op, lay, PB, any_deriv, tabs, xe, P, p0, np);
            __syncthreads();
            for (int item = threadIdx.x; item < np * n_fields; item += blockDim.x)
            {
                const int f = item / np, p = item - f * np; // consecutive threads -> consecutive points
                const unsigned m = masks.m[f];
                const double *t0 = tabs + lay.off[0] + ((m & 1u) ? PB * lay.ld[0] : 0) + p * lay.ld[0];
                const double *t1 = op.ndim > 1 ? tabs + lay.off[1] + ((m & 2u) ? PB * lay.ld[1] : 0) + p * lay.ld[1] : nullptr;
                const double *t2 = op.ndim > 2 ? tabs + lay.off[2] + ((m & 4u) ? PB * lay.ld[2] : 0) + p * lay.ld[2] : nullptr;
                double acc0 = 0.0;
                for (int i0 = 0; i0 < n0; ++i0)
                {
                    double acc1 = 0.0;
                    for (int i1 = 0; i1 < n1; ++i1)
                    {
                        const double *row = c + (i0 * n1 + i1) * n2;
                        double acc2 = 0.0;
                        if (op.ndim > 2)
                        {
                            for (int i2 = 0; i2 < n2; ++i2)
                                acc2 = fma(row[i2], t2[i2], acc2);
                        }
                        else
                            acc2 = row[0];
                        acc1 = op.ndim > 1 ? fma(acc2, t1[i1], acc1) : acc2;
                    }
                    acc0 = fma(acc1, t0[i0], acc0);
                }
                out[(e * n_fields + f) * P + p0 + p] = acc0;
            }
        }
    }
}

// 2-D Lagrange-family spaces with at most 8 basis functions per dimension (every BASELINE configuration): ONE
// THREAD PER POINT, everything in registers with compile-time shapes and unrolled loops -- the 1-D values and derivatives
// of both dimensions (same operations in the same order as point_tabs_fill), then the contraction in the order of the
// kernels above (inner sum over i1 for every i0, then over i0), so the result is bit-identical to theirs.  The inner sums
// with the values and with the derivatives of dimension 1 are formed once and shared by every requested field.  The
// DoFs of the TE elements a CTA pass covers sit in shared memory (a warp-uniform broadcast when P is a multiple of 32).
// About 1/12 of the instructions per point of the cooperative kernel, whose run-time loops and shared-memory tables
// made it issue-bound.
constexpr int kPtRegThreads = 128;

template <int N>
__device__ __forceinline__ void pt_lagrange_reg(const double *__restrict__ roots, const double *__restrict__ inv, double x,
                                                bool with_der, double (&val)[N], double (&der)[N])
{
    double dif[N];
#pragma unroll
    for (int m = 0; m < N; ++m)
        dif[m] = __dsub_rn(x, roots[m]);
#pragma unroll
    for (int j = 0; j < N; ++j)
    {
        double acc = 1.0, dd = 0.0;
#pragma unroll
        for (int m = 0; m < N; ++m)
            if (m != j)
            {
                dd = fma(dd, dif[m], acc);
                acc = __dmul_rn(acc, dif[m]);
            }
        val[j] = N == 1 ? 1.0 : __dmul_rn(acc, inv[j]);
        der[j] = (N == 1 || !with_der) ? 0.0 : __dmul_rn(dd, inv[j]);
    }
}

template <int N0, int N1>
__global__ void __launch_bounds__(kPtRegThreads)
    points_reconstruct_reg2d_kernel(const PointOp op, int TE, size_t E, size_t P, const double *__restrict__ dofs,
                                    const double *__restrict__ x, int n_fields, const PointMasks masks, double *__restrict__ out)
{
    extern __shared__ __align__(16) double pt_smem[];
    constexpr int N = N0 * N1;
    double *r0 = pt_smem, *v0 = r0 + N0, *r1 = v0 + N0, *v1 = r1 + N1; // roots | 1/denominators, per dimension
    double *c = v1 + N1;                                               // [TE][N]
    __shared__ unsigned fmask[8];
    if (threadIdx.x < N0)
    {
        r0[threadIdx.x] = op.dim[0].roots[threadIdx.x];
        v0[threadIdx.x] = op.dim[0].den[N0 + threadIdx.x];
    }
    if (threadIdx.x < N1)
    {
        r1[threadIdx.x] = op.dim[1].roots[threadIdx.x];
        v1[threadIdx.x] = op.dim[1].den[N1 + threadIdx.x];
    }
    if (threadIdx.x < 8)
        fmask[threadIdx.x] = masks.m[threadIdx.x];
    bool need_a = false, need_b = false, der0 = false;
#pragma unroll
    for (int f = 0; f < 8; ++f)
        if (f < n_fields)
        {
            need_a |= !(masks.m[f] & 2u);
            need_b |= (masks.m[f] & 2u) != 0;
            der0 |= (masks.m[f] & 1u) != 0;
        }
    const size_t n_tiles = (E + TE - 1) / TE;
    for (size_t t = blockIdx.x; t < n_tiles; t += gridDim.x)
    {
        const size_t e0 = t * TE;
        const int ne = (int)((E - e0) < (size_t)TE ? (E - e0) : (size_t)TE);
        const size_t items = (size_t)ne * P;
        // the coordinates of this thread's first point are requested together with the DoF tile: one memory latency per
        // tile instead of two
        double px0 = 0.0, px1 = 0.0;
        if (threadIdx.x < items)
        {
            const size_t el = threadIdx.x / P, p = threadIdx.x - el * P;
            px0 = __ldcs(x + (e0 + el) * 2 * P + p);
            px1 = __ldcs(x + (e0 + el) * 2 * P + P + p);
        }
        __syncthreads(); // DoFs of the previous tile consumed (and the root tables written)
        for (int i = threadIdx.x; i < ne * N; i += blockDim.x)
            c[i] = __ldcs(dofs + e0 * N + i);
        __syncthreads();
        for (size_t it = threadIdx.x; it < items; it += blockDim.x)
        {
            const int el = (int)(it / P);
            const size_t p = it - (size_t)el * P;
            const size_t e = e0 + el;
            const double *xe = x + e * 2 * P;
            const bool first = it == threadIdx.x;
            const double x0 = first ? px0 : __ldcs(xe + p), x1 = first ? px1 : __ldcs(xe + P + p);
            double t1v[N1], t1d[N1];
            pt_lagrange_reg<N1>(r1, v1, x1, need_b, t1v, t1d);
            const double *ce = c + el * N;
            double sa[N0], sb[N0]; // sum_i1 c[i0][i1] * (value | derivative) of dimension 1
#pragma unroll
            for (int i0 = 0; i0 < N0; ++i0)
            {
                double a = 0.0, b = 0.0;
#pragma unroll
                for (int i1 = 0; i1 < N1; ++i1)
                {
                    const double cv = ce[i0 * N1 + i1];
                    if (need_a)
                        a = fma(cv, t1v[i1], a);
                    if (need_b)
                        b = fma(cv, t1d[i1], b);
                }
                sa[i0] = a, sb[i0] = b;
            }
            double t0v[N0], t0d[N0];
            pt_lagrange_reg<N0>(r0, v0, x0, der0, t0v, t0d);
            for (int f = 0; f < n_fields; ++f)
            {
                const unsigned m = fmask[f];
                double acc = 0.0;
#pragma unroll
                for (int i0 = 0; i0 < N0; ++i0)
                    acc = fma((m & 2u) ? sb[i0] : sa[i0], (m & 1u) ? t0d[i0] : t0v[i0], acc);
                __stcs(out + (e * n_fields + f) * P + p, acc);
            }
        }
    }
}

template <int N0, int N1>
static void launch_points_reg2d(const LaunchCfg &cfg, const PointOp &op, size_t E, size_t P, const double *dofs, const double *x,
                                int n_fields, const PointMasks &pm, double *out)
{
    // elements per CTA pass: enough points to occupy every thread, at most 32 elements / 16 KB of DoFs
    const int TE = (int)std::min<size_t>(32, std::max<size_t>(1, (kPtRegThreads + P - 1) / P));
    const size_t smem = sizeof(double) * (2 * N0 + 2 * N1 + (size_t)TE * N0 * N1);
    const size_t n_tiles = (E + TE - 1) / TE;
    const int grid = (int)std::min<size_t>(n_tiles, (size_t)cfg.sm_count * 8);
    points_reconstruct_reg2d_kernel<N0, N1><<<grid, kPtRegThreads, smem, cfg.stream>>>(op, TE, E, P, dofs, x, n_fields, pm, out);
}

// FunctionSpace.evaluate for the same spaces: a warp takes 32 consecutive points of one element, every lane builds the
// 1-D values of its point in registers and writes the N products into the warp's shared-memory tile (odd row stride:
// conflict free); the tile -- 32 N contiguous doubles of the output -- then leaves with coalesced streaming stores.
// Same products as the cooperative kernel (1 * t0 * t1), bit-identical.
template <int N0, int N1>
__global__ void __launch_bounds__(kPtRegThreads)
    points_evaluate_reg2d_kernel(const PointOp op, size_t E, size_t P, const double *__restrict__ x, double *__restrict__ out)
{
    extern __shared__ __align__(16) double pt_smem[];
    constexpr int N = N0 * N1, NP = N | 1;
    double *r0 = pt_smem, *v0 = r0 + N0, *r1 = v0 + N0, *v1 = r1 + N1;
    const int lane = threadIdx.x & 31;
    double *tile = v1 + N1 + (size_t)(threadIdx.x >> 5) * 32 * NP; // [32][NP] per warp
    if (threadIdx.x < N0)
    {
        r0[threadIdx.x] = op.dim[0].roots[threadIdx.x];
        v0[threadIdx.x] = op.dim[0].den[N0 + threadIdx.x];
    }
    if (threadIdx.x < N1)
    {
        r1[threadIdx.x] = op.dim[1].roots[threadIdx.x];
        v1[threadIdx.x] = op.dim[1].den[N1 + threadIdx.x];
    }
    __syncthreads();
    const size_t groups = (P + 31) / 32, total = E * groups;
    const size_t warp = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    const size_t n_warps = ((size_t)gridDim.x * blockDim.x) >> 5;
    // the coordinates of the NEXT group are requested before this one's tile is written and copied out
    const auto fetch = [&](size_t it, double &c0, double &c1) {
        c0 = c1 = 0.0;
        if (it < total)
        {
            const size_t e = it / groups, p0 = (it - e * groups) * 32;
            if (p0 + lane < P)
            {
                c0 = __ldcs(x + e * 2 * P + p0 + lane);
                c1 = __ldcs(x + e * 2 * P + P + p0 + lane);
            }
        }
    };
    double nx0, nx1;
    fetch(warp, nx0, nx1);
    for (size_t it = warp; it < total; it += n_warps)
    {
        const size_t e = it / groups, p0 = (it - e * groups) * 32;
        const int np = (int)((P - p0) < 32 ? (P - p0) : 32);
        const double x0 = nx0, x1 = nx1;
        fetch(it + n_warps, nx0, nx1);
        double t0v[N0], t0d[N0], t1v[N1], t1d[N1];
        pt_lagrange_reg<N0>(r0, v0, x0, false, t0v, t0d);
        pt_lagrange_reg<N1>(r1, v1, x1, false, t1v, t1d);
        __syncwarp(); // the previous tile has been copied out
#pragma unroll
        for (int i0 = 0; i0 < N0; ++i0)
#pragma unroll
            for (int i1 = 0; i1 < N1; ++i1)
                tile[lane * NP + i0 * N1 + i1] = __dmul_rn(__dmul_rn(1.0, t0v[i0]), t1v[i1]);
        __syncwarp();
        double *o = out + (e * P + p0) * N;
        for (int idx = lane; idx < np * N; idx += 32)
        {
            const int row = idx / N, col = idx - row * N;
            __stcs(o + idx, tile[row * NP + col]);
        }
    }
}

template <int N0, int N1>
static void launch_points_eval_reg2d(const LaunchCfg &cfg, const PointOp &op, size_t E, size_t P, const double *x, double *out)
{
    constexpr int NP = (N0 * N1) | 1;
    const size_t smem = sizeof(double) * (2 * N0 + 2 * N1 + 4 * 32 * NP);
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(points_evaluate_reg2d_kernel<N0, N1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const size_t total_warps = E * ((P + 31) / 32);
    const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, 200 * 1024 / (smem + 1024)));
    const int grid = (int)std::min<size_t>((total_warps + 3) / 4, (size_t)cfg.sm_count * per_sm);
    points_evaluate_reg2d_kernel<N0, N1><<<grid, kPtRegThreads, smem, cfg.stream>>>(op, E, P, x, out);
}

// the instantiated shapes: n0, n1 <= 8 and |n0 - n1| <= 1 (isotropic spaces and the k-form component spaces p x (p+1))
static bool dispatch_points_reg2d(const LaunchCfg &cfg, const PointOp &op, size_t E, size_t P, const double *dofs, const double *x,
                                  int n_fields, const PointMasks &pm, double *out)
{
#define IB_PT_CASE(A, B)                                                                                                        \
    case (A) * 16 + (B):                                                                                                        \
        launch_points_reg2d<A, B>(cfg, op, E, P, dofs, x, n_fields, pm, out);                                                    \
        return true;
#define IB_PT_ROW(A) IB_PT_CASE(A, A) IB_PT_CASE(A, A + 1) IB_PT_CASE(A + 1, A)
    switch (op.dim[0].n * 16 + op.dim[1].n)
    {
        IB_PT_ROW(1) IB_PT_ROW(2) IB_PT_ROW(3) IB_PT_ROW(4) IB_PT_ROW(5) IB_PT_ROW(6) IB_PT_ROW(7) IB_PT_CASE(8, 8)
    }
#undef IB_PT_ROW
#undef IB_PT_CASE
    return false;
}

static bool dispatch_points_eval_reg2d(const LaunchCfg &cfg, const PointOp &op, size_t E, size_t P, const double *x, double *out)
{
#define IB_PT_CASE(A, B)                                                                                                        \
    case (A) * 16 + (B):                                                                                                        \
        launch_points_eval_reg2d<A, B>(cfg, op, E, P, x, out);                                                                   \
        return true;
#define IB_PT_ROW(A) IB_PT_CASE(A, A) IB_PT_CASE(A, A + 1) IB_PT_CASE(A + 1, A)
    switch (op.dim[0].n * 16 + op.dim[1].n)
    {
        IB_PT_ROW(1) IB_PT_ROW(2) IB_PT_ROW(3) IB_PT_ROW(4) IB_PT_ROW(5) IB_PT_ROW(6) IB_PT_ROW(7) IB_PT_CASE(8, 8)
    }
#undef IB_PT_ROW
#undef IB_PT_CASE
    return false;
}

__global__ void __launch_bounds__(128)
    points_evaluate_coop_kernel(const PointOp op, int PB, size_t E, size_t P, const double *__restrict__ x, double *__restrict__ out)
{
    extern __shared__ __align__(16) double pt_smem[];
    const PointTabs lay = point_tabs_layout(op, PB, false);
    double *tabs = pt_smem;
    const int n1 = op.ndim > 1 ? op.dim[1].n : 1, n2 = op.ndim > 2 ? op.dim[2].n : 1, n12 = n1 * n2;
    for (size_t e = blockIdx.x; e < E; e += gridDim.x)
    {
        const double *xe = x + e * (size_t)op.ndim * P;
        for (size_t p0 = (size_t)blockIdx.y * PB; p0 < P; p0 += (size_t)PB * gridDim.y)
        {
            const int np = (int)((P - p0) < (size_t)PB ? (P - p0) : (size_t)PB);
            __syncthreads();
            point_tabs_fill(op, lay, PB, false, tabs, xe, P, p0, np);
            __syncthreads();
            double *o = out + (e * P + p0) * (size_t)op.N;
            for (int p = 0; p < np; ++p)
            {
                const double *t0 = tabs + lay.off[0] + p * lay.ld[0];
                const double *t1 = tabs + lay.off[1] + p * lay.ld[1], *t2 = tabs + lay.off[2] + p * lay.ld[2];
                for (int i = threadIdx.x; i < op.N; i += blockDim.x)
                {
                    const int i0 = i / n12, r = i - i0 * n12, i1 = r / n2, i2 = r - i1 * n2;
                    double v = __dmul_rn(1.0, t0[i0]);
                    if (op.ndim > 1)
                        v = __dmul_rn(v, t1[i1]);
                    if (op.ndim > 2)
                        v = __dmul_rn(v, t2[i2]);
                    __stcs(o + (size_t)p * op.N + i, v);
                }
            }
        }
    }
}

// N3 -- DegreesOfFreedom.derivative(idim), /root/reference/src/python/degrees_of_freedom.c:738-945: the 1-D incidence
// operator [n][n+1] (src/python/incidence.c:38-110) applied along one dimension of every element's DoF tensor,
//   out[e][pre][row][post] = sum_col Inc[row][col] * in[e][pre][col][post].
// One thread per output value; the additions run in the reference's order with separate roundings (Bernstein: (0 - c*in[r])
// + c*in[r+1]; Legendre: columns r+1, r+3, .. descending, weight 2r+1; Lagrange: columns ascending from 0), so the
// result is bit-identical to the reference's.
__global__ void __launch_bounds__(256)
    dof_derivative_kernel(int type, int n, size_t pre, size_t post, const double *__restrict__ inc, size_t E,
                          const double *__restrict__ in, double *__restrict__ out)
{
    const size_t per_out = pre * (size_t)n * post, per_in = pre * (size_t)(n + 1) * post;
    const size_t total = E * per_out;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x)
    {
        const size_t e = idx / per_out, rem = idx - e * per_out;
        const size_t i_pre = rem / ((size_t)n * post), r2 = rem - i_pre * (size_t)n * post;
        const int row = (int)(r2 / post);
        const size_t i_post = r2 - (size_t)row * post;
        const double *src = in + e * per_in + i_pre * (size_t)(n + 1) * post + i_post;
        double v = 0.0;
        if (type == BASIS_BERNSTEIN)
        {
            const double c = (double)n / 2.0;
            v = __dadd_rn(__dsub_rn(0.0, __dmul_rn(c, src[(size_t)row * post])), __dmul_rn(c, src[(size_t)(row + 1) * post]));
        }
        else if (type == BASIS_LEGENDRE)
        {
            const double w = (double)(2 * row + 1);
            int col = row + 1 + 2 * ((n - row - 1) / 2); // largest column <= n with the parity of row + 1
            for (; col > row; col -= 2)
                v = __dadd_rn(v, __dmul_rn(w, src[(size_t)col * post]));
        }
        else
        {
            const double *m = inc + (size_t)row * (n + 1);
            for (int col = 0; col <= n; ++col)
                v = __dadd_rn(v, __dmul_rn(m[col], src[(size_t)col * post]));
        }
        out[idx] = v;
    }
}

// Tiled variant (the one normally launched): a CTA stages the DoFs of EPB consecutive elements (one contiguous run of
// global memory) in shared memory with coalesced loads, every thread then forms output values from shared memory with
// 32-bit index arithmetic and stores them coalesced.  Same additions in the same order as dof_derivative_kernel.
__global__ void __launch_bounds__(256)
    dof_derivative_tiled_kernel(int type, int n, int pre, int post, int epb, const double *__restrict__ inc, size_t E,
                                const double *__restrict__ in, double *__restrict__ out)
{
    extern __shared__ __align__(16) double dd_smem[];
    const int per_in = pre * (n + 1) * post, per_out = pre * n * post;
    double *tile = dd_smem;                  // [epb][per_in]
    double *incs = dd_smem + epb * per_in;   // [n][n+1] (Lagrange only)
    int2 *desc = reinterpret_cast<int2 *>(incs + (inc ? n * (n + 1) : 0)); // [per_out] {offset of column 0 in the element, row}
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    if (inc)
        for (int i = threadIdx.x; i < n * (n + 1); i += blockDim.x)
            incs[i] = inc[i];
    for (int r = threadIdx.x; r < per_out; r += blockDim.x)
    {
        const int i_pre = r / (n * post), r2 = r - i_pre * n * post;
        const int row = r2 / post, i_post = r2 - row * post;
        desc[r] = make_int2(i_pre * (n + 1) * post + i_post, row);
    }
    const size_t n_tiles = (E + epb - 1) / epb;
    for (size_t t = blockIdx.x; t < n_tiles; t += gridDim.x)
    {
        const size_t e0 = t * epb;
        const int ne = (int)((E - e0) < (size_t)epb ? (E - e0) : (size_t)epb);
        const double *src = in + e0 * per_in;
        __syncthreads(); // previous tile consumed (and incs written)
        for (int i = threadIdx.x; i < ne * per_in; i += blockDim.x)
            tile[i] = __ldcs(src + i);
        __syncthreads();
        double *dst = out + e0 * per_out;
        // a warp takes whole elements, lane r walks the element's outputs: (source offset, row) of output r come
        // from the table built once per CTA -> no integer division in the loop
        for (int el = warp; el < ne; el += n_warps)
        {
            const double *se = tile + el * per_in;
            for (int r = lane; r < per_out; r += 32)
            {
                const int2 d = desc[r];
                const double *sp = se + d.x;
                const int row = d.y;
                double v = 0.0;
                if (type == BASIS_BERNSTEIN)
                {
                    const double c = (double)n / 2.0;
                    v = __dadd_rn(__dsub_rn(0.0, __dmul_rn(c, sp[row * post])), __dmul_rn(c, sp[(row + 1) * post]));
                }
                else if (type == BASIS_LEGENDRE)
                {
                    const double w = (double)(2 * row + 1);
                    for (int col = row + 1 + 2 * ((n - row - 1) / 2); col > row; col -= 2)
                        v = __dadd_rn(v, __dmul_rn(w, sp[col * post]));
                }
                else
                {
                    const double *m = incs + row * (n + 1);
                    for (int col = 0; col <= n; ++col)
                        v = __dadd_rn(v, __dmul_rn(m[col], sp[col * post]));
                }
                __stcs(dst + el * per_out + r, v);
            }
        }
    }
}

// Column variant (orders up to kDerivColsMax, the one normally launched): a work item is one (element, pre, post) LINE of
// the DoF tensor -- a thread reads the n+1 values of the line from the staged input tile into registers ONCE, forms all n
// derivative DoFs from them (incidence row from shared memory, a warp-uniform broadcast; loops unrolled at compile time)
// and writes them to a staged output tile.  Tiles are contiguous runs of global memory, so they move as ONE bulk async
// copy each way (cp.async.bulk: mbarrier completion inbound, bulk group outbound), double buffered: the load of tile
// i+1 and the store of tile i-1 are in flight while tile i is computed.  The additions and their order are those of the
// kernels above, so the result is bit-identical.
constexpr int kDerivColsMax = 12;
constexpr int kDerivColsThreads = 128;

__device__ __forceinline__ unsigned dd_smem_u32(const void *p)
{
    return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void dd_mbar_wait(unsigned long long *bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "DD_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DD_DONE;\n"
        "bra DD_WAIT;\n"
        "DD_DONE:\n"
        "}\n" ::"r"(dd_smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void dd_load_tile(double *dst, const double *src, unsigned bytes, unsigned long long *bar)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(dd_smem_u32(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dd_smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(dd_smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void dd_store_tile(double *gdst, const double *ssrc, unsigned bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(dd_smem_u32(ssrc)), "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

// all lines of `ne` staged elements: tin [ne][per_in] -> tout [ne][per_out].  (Two lines per thread, to share the incidence
// loads, was measured slower: 74 registers and more shared-memory conflicts, 0.59 instead of 0.89 of the HBM roofline.)
template <int NN>
__device__ __forceinline__ void deriv_cols_tile(int type, int post, int n_lines, const int2 *desc, const double *incs,
                                                const double *tin, double *tout)
{
    for (int it = threadIdx.x; it < n_lines; it += blockDim.x)
    {
        const int2 d = desc[it];
        const double *sp = tin + d.x;
        double *op = tout + d.y;
        double x[NN + 1];
#pragma unroll
        for (int c = 0; c <= NN; ++c)
            x[c] = sp[c * post];
        if (type == BASIS_BERNSTEIN)
        {
            const double c = (double)NN / 2.0;
#pragma unroll
            for (int row = 0; row < NN; ++row)
                op[row * post] = __dadd_rn(__dsub_rn(0.0, __dmul_rn(c, x[row])), __dmul_rn(c, x[row + 1]));
        }
        else if (type == BASIS_LEGENDRE)
        {
#pragma unroll
            for (int row = 0; row < NN; ++row)
            {
                const double w = (double)(2 * row + 1);
                double v = 0.0;
#pragma unroll
                for (int col = row + 1 + 2 * ((NN - row - 1) / 2); col > row; col -= 2)
                    v = __dadd_rn(v, __dmul_rn(w, x[col]));
                op[row * post] = v;
            }
        }
        else
        {
#pragma unroll
            for (int row = 0; row < NN; ++row)
            {
                const double *m = incs + row * (NN + 1);
                double v = 0.0;
#pragma unroll
                for (int col = 0; col <= NN; ++col)
                    v = __dadd_rn(v, __dmul_rn(m[col], x[col]));
                op[row * post] = v;
            }
        }
    }
}

// epb is chosen by the launcher so that a full tile starts 16-byte aligned and is a multiple of 16 bytes both ways
template <int NN>
__global__ void __launch_bounds__(kDerivColsThreads)
    dof_derivative_cols_kernel(int type, int pre, int post, int epb, const double *__restrict__ inc, size_t E,
                               const double *__restrict__ in, double *__restrict__ out)
{
    extern __shared__ __align__(16) double dd_smem[];
    const int lines = pre * post, per_in = lines * (NN + 1), per_out = lines * NN;
    const int in_tile = (epb * per_in + 1) & ~1, out_tile = (epb * per_out + 1) & ~1; // doubles, 16-byte multiples
    double *tin = dd_smem;                  // [2][in_tile]
    double *tout = tin + 2 * in_tile;       // [2][out_tile]
    double *incs = tout + 2 * out_tile;     // [NN][NN+1] (Lagrange only)
    unsigned long long *bars = reinterpret_cast<unsigned long long *>(incs + (inc ? NN * (NN + 1) : 0)); // [2]
    int2 *desc = reinterpret_cast<int2 *>(bars + 2); // [epb * lines] {first input, first output} of a line
    if (threadIdx.x == 0)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(dd_smem_u32(bars)));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(dd_smem_u32(bars + 1)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (inc)
        for (int i = threadIdx.x; i < NN * (NN + 1); i += blockDim.x)
            incs[i] = inc[i];
    for (int it = threadIdx.x; it < epb * lines; it += blockDim.x)
    {
        const int el = it / lines, rem = it - el * lines;
        const int i_pre = rem / post, i_post = rem - i_pre * post;
        desc[it] = make_int2(el * per_in + i_pre * (NN + 1) * post + i_post, el * per_out + i_pre * NN * post + i_post);
    }
    __syncthreads();
    const size_t n_full = E / epb; // full tiles go through the bulk-copy pipeline
    const unsigned in_bytes = (unsigned)(epb * per_in * sizeof(double)), out_bytes = (unsigned)(epb * per_out * sizeof(double));
    if (threadIdx.x == 0 && blockIdx.x < n_full)
        dd_load_tile(tin, in + (size_t)blockIdx.x * epb * per_in, in_bytes, bars);
    int i = 0;
    for (size_t t = blockIdx.x; t < n_full; t += gridDim.x, ++i)
    {
        const int cur = i & 1;
        if (threadIdx.x == 0)
            asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); // the store that last used tout[cur] has read it
        __syncthreads(); // ... and every thread is done with tin[cur ^ 1] (tile i - 1)
        if (threadIdx.x == 0 && t + gridDim.x < n_full)
            dd_load_tile(tin + (cur ^ 1) * in_tile, in + (t + gridDim.x) * epb * per_in, in_bytes, bars + (cur ^ 1));
        dd_mbar_wait(bars + cur, (unsigned)(i >> 1) & 1u);
        deriv_cols_tile<NN>(type, post, epb * lines, desc, incs, tin + cur * in_tile, tout + cur * out_tile);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (threadIdx.x == 0)
            dd_store_tile(out + t * epb * per_out, tout + cur * out_tile, out_bytes);
    }
    if (threadIdx.x == 0)
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    // the ragged tail (E % epb elements) with plain loads and stores, by the last CTA
    const int ne = (int)(E - n_full * epb);
    if (ne > 0 && blockIdx.x == gridDim.x - 1)
    {
        __syncthreads();
        const double *src = in + n_full * epb * per_in;
        double *dst = out + n_full * epb * per_out;
        for (int k = threadIdx.x; k < ne * per_in; k += blockDim.x)
            tin[k] = src[k];
        __syncthreads();
        deriv_cols_tile<NN>(type, post, ne * lines, desc, incs, tin, tout);
        __syncthreads();
        for (int k = threadIdx.x; k < ne * per_out; k += blockDim.x)
            dst[k] = tout[k];
    }
}

template <int NN>
static void launch_deriv_cols(const LaunchCfg &cfg, int grid, size_t smem, int type, int pre, int post, int epb, const double *d_inc,
                              size_t E, const double *in, double *out)
{
    if (smem > 48 * 1024) // per launch: the attribute is per device
        cudaFuncSetAttribute(dof_derivative_cols_kernel<NN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    dof_derivative_cols_kernel<NN><<<grid, kDerivColsThreads, smem, cfg.stream>>>(type, pre, post, epb, d_inc, E, in, out);
}

} // namespace

int launch_dof_derivative(const LaunchCfg &cfg, int type, int n, size_t pre, size_t post, const double *d_inc, size_t E,
                          const double *in, double *out)
{
    const size_t total = E * pre * (size_t)n * post;
    if (total == 0)
        return INTERP_SUCCESS;
    const size_t per_in = pre * (size_t)(n + 1) * post;
    const size_t inc_bytes = d_inc ? sizeof(double) * n * (n + 1) : 0;
    const size_t lines = pre * post, per_out = lines * n;
    const size_t el_bytes = sizeof(double) * 2 * (per_in + per_out) + sizeof(int2) * lines; // two buffers each way + line table
    const bool aligned = (((uintptr_t)in | (uintptr_t)out) & 15u) == 0;
    if (n <= kDerivColsMax && aligned && 2 * el_bytes + inc_bytes <= 56 * 1024 && !getenv("IB200_DERIV_NO_COLS"))
    {
        // elements per tile: an EVEN count (tiles then start 16-byte aligned and are 16-byte multiples whatever per_in is),
        // 20 .. 52 KB of shared memory per CTA, chosen so that the lines of a tile fill whole passes of the CTA
        const int lo = (int)std::max<size_t>(2, (20 * 1024 / el_bytes) & ~(size_t)1);
        const int hi = (int)std::max<size_t>(lo, (52 * 1024 / el_bytes) & ~(size_t)1);
        int epb = lo;
        double best = 0.0;
        for (int c = lo; c <= hi; c += 2)
        {
            const size_t items = (size_t)c * lines, passes = (items + kDerivColsThreads - 1) / kDerivColsThreads;
            const double eff = (double)items / (double)(passes * kDerivColsThreads);
            if (eff > best + 1e-9)
                best = eff, epb = c;
        }
        const size_t smem = (size_t)epb * el_bytes + inc_bytes + 64;
        const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(16, 216 * 1024 / (smem + 1024)));
        const size_t n_tiles = (E + epb - 1) / epb;
        const int grid = (int)std::min<size_t>(n_tiles, (size_t)cfg.sm_count * per_sm);
#define IB_DERIV_CASE(NN)                                                                                                       \
    case NN:                                                                                                                    \
        launch_deriv_cols<NN>(cfg, grid, smem, type, (int)pre, (int)post, epb, d_inc, E, in, out);                               \
        break;
        switch (n)
        {
            IB_DERIV_CASE(1) IB_DERIV_CASE(2) IB_DERIV_CASE(3) IB_DERIV_CASE(4) IB_DERIV_CASE(5) IB_DERIV_CASE(6)
            IB_DERIV_CASE(7) IB_DERIV_CASE(8) IB_DERIV_CASE(9) IB_DERIV_CASE(10) IB_DERIV_CASE(11) IB_DERIV_CASE(12)
        }
#undef IB_DERIV_CASE
    }
    else if (2 * per_in * sizeof(double) + inc_bytes <= 44 * 1024)
    {
        // about 24 KB of staged DoFs per CTA (at least one element), 4+ CTAs per SM
        const int epb = (int)std::max<size_t>(1, std::min<size_t>(24 * 1024 / (per_in * sizeof(double)), 256));
        const size_t smem = sizeof(double) * epb * per_in + inc_bytes + sizeof(int2) * pre * n * post;
        const size_t n_tiles = (E + epb - 1) / epb;
        const int grid = (int)std::min<size_t>(n_tiles, (size_t)cfg.sm_count * 6);
        dof_derivative_tiled_kernel<<<grid, 256, smem, cfg.stream>>>(type, n, (int)pre, (int)post, epb, d_inc, E, in, out);
    }
    else
    {
        const int grid = (int)std::min<size_t>((total + 255) / 256, (size_t)cfg.sm_count * 8);
        dof_derivative_kernel<<<grid, 256, 0, cfg.stream>>>(type, n, pre, post, d_inc, E, in, out);
    }
    ++*cfg.launch_counter;
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess)
        return ib200_fail(IB200_ERROR_CUDA, "launch of dof_derivative_kernel failed: %s", cudaGetErrorString(err));
    return INTERP_SUCCESS;
}

// Shape of a cooperative launch: points per chunk (<= 64, shrunk until tables + DoFs fit 96 KB), grid (elements x point
// chunks, about 8 CTAs per SM in total) and the dynamic shared memory in bytes; 0 -> use the thread-per-point kernel.
static size_t coop_shape(const LaunchCfg &cfg, const PointOp &op, size_t E, size_t P, bool with_der, bool with_dofs, int *PB,
                         dim3 *grid)
{
    if (std::getenv("IB200_POINTS_NO_COOP"))
        return 0;
    int pb = (int)std::min<size_t>(64, P);
    auto bytes = [&](int b) {
        size_t d = with_dofs ? (size_t)op.N : 0;
        for (int k = 0; k < op.ndim; ++k)
            d += (size_t)(with_der ? 2 : 1) * b * (op.dim[k].n | 1);
        return d * sizeof(double);
    };
    while (pb > 8 && bytes(pb) > 96 * 1024)
        pb /= 2;
    if (bytes(pb) > 96 * 1024)
        return 0;
    // per launch, not once per process: the attribute is per device and a process may hold contexts on several
    if (cudaFuncSetAttribute(points_reconstruct_coop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024) != cudaSuccess ||
        cudaFuncSetAttribute(points_evaluate_coop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024) != cudaSuccess)
    {
        cudaGetLastError();
        return 0;
    }
    const size_t chunks = (P + pb - 1) / pb, want = (size_t)cfg.sm_count * 8;
    const size_t gx = std::min(E, want);
    const size_t gy = std::min<size_t>(chunks, std::max<size_t>(1, want / gx));
    *PB = pb;
    *grid = dim3((unsigned)gx, (unsigned)std::min<size_t>(gy, 65535));
    return bytes(pb);
}

int launch_points_evaluate(const LaunchCfg &cfg, const PointOp &op, size_t E, size_t P, const double *x, double *out)
{
    if (E == 0 || P == 0)
        return INTERP_SUCCESS;
    for (int k = 0; k < op.ndim; ++k)
        if (op.dim[k].n > PT_MAXN)
            return ib200_fail(IB200_ERROR_UNSUPPORTED, "arbitrary-point evaluation supports orders up to %d", PT_MAXN - 1);
    const auto has_roots = [](int type) { return type != BASIS_LEGENDRE && type != BASIS_BERNSTEIN; };
    if (op.ndim == 2 && has_roots(op.dim[0].type) && has_roots(op.dim[1].type) && !getenv("IB200_POINTS_NO_REG") &&
        dispatch_points_eval_reg2d(cfg, op, E, P, x, out))
    {
        ++*cfg.launch_counter;
        const cudaError_t err = cudaGetLastError();
        if (err != cudaSuccess)
            return ib200_fail(IB200_ERROR_CUDA, "launch of points_evaluate_reg2d_kernel failed: %s", cudaGetErrorString(err));
        return INTERP_SUCCESS;
    }
    int PB = 0;
    dim3 grid;
    const size_t smem = coop_shape(cfg, op, E, P, false, false, &PB, &grid);
    if (smem)
        points_evaluate_coop_kernel<<<grid, 128, smem, cfg.stream>>>(op, PB, E, P, x, out);
    else
    {
        const size_t total = E * P;
        const int g1 = (int)std::min<size_t>((total + 127) / 128, (size_t)cfg.sm_count * 16);
        points_evaluate_kernel<<<g1, 128, 0, cfg.stream>>>(op, E, P, x, out);
    }
    ++*cfg.launch_counter;
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess)
        return ib200_fail(IB200_ERROR_CUDA, "launch of points_evaluate_kernel failed: %s", cudaGetErrorString(err));
    return INTERP_SUCCESS;
}

int launch_points_reconstruct(const LaunchCfg &cfg, const PointOp &op, size_t E, size_t P, const double *dofs, const double *x,
                              unsigned n_fields, const unsigned *masks, double *out)
{
    PointMasks pm{};
    bool any_deriv = false;
    for (unsigned f = 0; f < n_fields && f < 8; ++f)
    {
        pm.m[f] = masks[f];
        any_deriv = any_deriv || masks[f] != 0;
    }
    if (E == 0 || P == 0 || n_fields == 0)
        return INTERP_SUCCESS;
    for (int k = 0; k < op.ndim; ++k)
        if (op.dim[k].n > PT_MAXN)
            return ib200_fail(IB200_ERROR_UNSUPPORTED, "arbitrary-point evaluation supports orders up to %d", PT_MAXN - 1);
    const auto has_roots = [](int type) { return type != BASIS_LEGENDRE && type != BASIS_BERNSTEIN; };
    if (op.ndim == 2 && has_roots(op.dim[0].type) && has_roots(op.dim[1].type) && !getenv("IB200_POINTS_NO_REG") &&
        dispatch_points_reg2d(cfg, op, E, P, dofs, x, (int)n_fields, pm, out))
    {
        ++*cfg.launch_counter;
        const cudaError_t err = cudaGetLastError();
        if (err != cudaSuccess)
            return ib200_fail(IB200_ERROR_CUDA, "launch of points_reconstruct_reg2d_kernel failed: %s", cudaGetErrorString(err));
        return INTERP_SUCCESS;
    }
    int PB = 0;
    dim3 grid;
    const size_t smem = coop_shape(cfg, op, E, P, any_deriv, true, &PB, &grid);
    if (smem)
        points_reconstruct_coop_kernel<<<grid, 128, smem, cfg.stream>>>(op, PB, E, P, dofs, x, (int)n_fields, pm, any_deriv, out);
    else
    {
        const unsigned gx = (unsigned)std::min<size_t>((P + 127) / 128, 1024);
        const unsigned gy = (unsigned)std::min<size_t>(E, std::max<size_t>(1, (size_t)cfg.sm_count * 16 / gx));
        points_reconstruct_kernel<<<dim3(gx, gy), 128, 0, cfg.stream>>>(op, E, P, dofs, x, (int)n_fields, pm, any_deriv, out);
    }
    ++*cfg.launch_counter;
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess)
        return ib200_fail(IB200_ERROR_CUDA, "launch of points_reconstruct_kernel failed: %s", cudaGetErrorString(err));
    return INTERP_SUCCESS;
}
