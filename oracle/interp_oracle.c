/* interp_oracle.c -- CPU restatement of the interplib hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This file is the parity checker of the repository: tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py are the only callers.  Nothing on the product
 * path (interplib_b200/, libinterp_b200.so) includes, links or loads it.
 *
 * It restates, per element and in the reference's own loop and floating-point order (so that it
 * compares bit-for-bit with the compiled reference in oracle/_ref, see tests/test_oracle_vs_ref.py),
 * the functions of SURVEY.md section 8a; paths below are into /root/reference:
 *
 *   orc_rule            integration_rule_for_order            src/integration/integration_rules.c:40-73
 *                       gauss_legendre_nodes_weights          src/integration/gauss_legendre.c:7-50
 *                       gauss_lobatto_nodes_weights           src/integration/gauss_lobatto.c:14-61
 *   orc_roots           generate_lagrange_roots               src/basis/basis_lagrange.c:11-42
 *   orc_basis_table     lagrange_/legendre_/bernstein_basis_create
 *                                                              src/basis/basis_lagrange.c:45-71,
 *                                                              src/basis/basis_legendre.c:9-40,
 *                                                              src/basis/basis_bernstein.c:9-23
 *   orc_reconstruct     compute_integration_point_values[_derivatives]
 *                                                              src/python/degrees_of_freedom.c:232-303
 *   orc_geometry        coordinate_map_new, space_map_new     src/python/mappings.c:8-90, 304-556
 *                       matrix_qr_decompose/_back_substitute  src/operations/matrices.c:15-86, 117-141
 *   orc_mass            compute_mass_matrix                   src/python/mass_matrices.c:161-339
 *                       compute_gradient_mass_matrix          src/python/mass_matrices.c:341-565
 *   orc_l2_dual         projection_l2_dual                    python/interplib/integration.py:106-168
 *   orc_space_weights   integration_space_weights             src/python/integration_objects.c:633-716
 *   orc_evaluate_points FunctionSpace.evaluate                src/python/function_space_objects.c:142-276,
 *                       basis_compute_outer_product_basis     src/basis/basis_set.c:462-633,
 *                       lagrange_polynomial_values_2 / _first_derivative_2
 *                                                              src/polynomials/lagrange.c:108-183, 383-525
 *   orc_reconstruct_points  reconstruct                       python/interplib/degrees_of_freedom.py:10-37
 *   orc_incidence_matrix    incidence_matrix                  src/python/incidence.c:38-110
 *   orc_dof_derivative      DegreesOfFreedom.derivative       src/python/degrees_of_freedom.c:738-945
 *
 * Parity status: PINNED -- validated `==` against the reference itself compiled in this
 * container (oracle/build_ref.py) and against the golden vectors in tests/golden/ that the same
 * compiled reference generated (tests/golden/make_golden.py).
 *
 * Build: gcc -std=gnu17 -O2 -fPIC -ffp-contract=off [-fopenmp] -shared interp_oracle.c -lm
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

enum
{
    ORC_MAX_DIM = 6
};
enum
{
    B_LEGENDRE = 1,
    B_LAG_GLL = 2,
    B_LAG_GAUSS = 3,
    B_LAG_UNIFORM = 4,
    B_LAG_CHEB = 5,
    B_BERNSTEIN = 6
};
enum
{
    R_GAUSS = 1,
    R_LOBATTO = 2
};

/* ------------------------------------------------------------------ Legendre recursion ---- */

static void bonnet_pair(unsigned n, double x, double out[2])
{
    double older = 1.0, newer = x;
    for (unsigned i = 2; i < n + 1; ++i)
    {
        const double t1 = (2 * i - 1) * x;
        const double t2 = (i - 1);
        const double val = (t1 * newer - t2 * older) / (double)(i);
        older = newer;
        newer = val;
    }
    out[0] = older;
    out[1] = newer;
}

/* ------------------------------------------------------------------ quadrature rules ------ */

static void rule_gauss(unsigned n, double tol, unsigned max_iter, double *x, double *w)
{
    if (n == 1)
    {
        x[0] = 0.0;
        if (w)
            w[0] = 2.0;
        return;
    }
    if (n == 2)
    {
        x[0] = -1.0, x[1] = 1.0;
        if (w)
            w[0] = 1.0, w[1] = 1.0;
        return;
    }
    for (unsigned i = 0; i < n; ++i)
    {
        double guess = cos(M_PI * (double)(4 * i + 3) / (double)(2 * (2 * (n) + 1)));
        double err = 1.0;
        double P[2];
        for (unsigned it = 0; it < max_iter && err > tol; ++it)
        {
            bonnet_pair(n, guess, P);
            const double q = 1 - guess * guess;
            const double slope = (n - 1) * (P[0] - guess * P[1]);
            const double delta = P[1] / slope * q;
            guess -= delta;
            err = fabs(delta);
        }
        x[n - 1 - i] = guess;
        if (w)
        {
            bonnet_pair(n - 1, guess, P);
            w[n - 1 - i] = 2.0 / (n * n * P[1] * P[1]) * (1 - guess * guess);
        }
    }
}

static void rule_lobatto(unsigned n, double tol, unsigned max_iter, double *x, double *w)
{
    if (n == 1)
    {
        x[0] = 0.0;
        if (w)
            w[0] = 2.0;
        return;
    }
    if (n == 2)
    {
        x[0] = -1.0, x[1] = +1.0;
        if (w)
            w[0] = 1.0, w[1] = 1.0;
        return;
    }
    x[0] = -1.0;
    x[n - 1] = +1.0;
    if (w)
        w[n - 1] = w[0] = 2.0 / (double)(n * (n - 1));
    const double k1 = (1.0 - 3.0 * (n - 2) / (double)(8.0 * (n - 1) * (n - 1) * (n - 1)));
    const double k2 = M_PI / (4.0 * (n - 1) + 1);
    for (unsigned i = 2; i < n; ++i)
    {
        double guess = k1 * cos(k2 * (4 * i - 3));
        double err = 1.0;
        double P[2];
        for (unsigned it = 0; it < max_iter && err > tol; ++it)
        {
            bonnet_pair(n - 1, guess, P);
            const double q = 1 - guess * guess;
            const double y1 = (n - 1) * (P[0] - guess * P[1]) / q;
            const double y2 = (2 * guess * y1 - (n - 1) * n * P[1]) / q;
            const double y3 = (4 * guess * y2 - ((n - 1) * n - 2) * y1) / q;
            const double delta = 2 * y1 * y2 / (2 * y2 * y2 - y1 * y3);
            guess -= delta;
            err = fabs(delta);
        }
        x[n - i] = guess;
        if (w)
        {
            bonnet_pair(n - 1, guess, P);
            w[n - i] = 2.0 / (n * (n - 1) * P[1] * P[1]);
        }
    }
}

int orc_rule(int type, unsigned order, double *x, double *w)
{
    if (type == R_GAUSS)
        rule_gauss(order + 1, 1e-14, 1000, x, w);
    else if (type == R_LOBATTO)
        rule_lobatto(order + 1, 1e-14, 1000, x, w);
    else
        return 5;
    return 0;
}

int orc_roots(int type, unsigned order, double *r)
{
    const unsigned n = order + 1;
    switch (type)
    {
    case B_LAG_UNIFORM:
        for (unsigned i = 0; i < n; ++i)
            r[i] = (2.0 * i) / (double)order - 1.0;
        return 0;
    case B_LAG_CHEB:
        for (unsigned i = 0; i < n; ++i)
            r[i] = -cos(M_PI * (double)(2 * i + 1) / (double)(2 * (order + 1)));
        return 0;
    case B_LAG_GAUSS:
        rule_gauss(n, 1e-12, 100, r, NULL);
        return 0;
    case B_LAG_GLL:
        rule_lobatto(n, 1e-12, 100, r, NULL);
        return 0;
    default:
        return 5;
    }
}

/* ------------------------------------------------------------------ 1-D basis tables ------ */

/* incremental pair loop of the reference: every pair (i<j) multiplies slot i by +d, slot j by -d */
static void lag_denominators(unsigned n, const double *r, double *den)
{
    den[0] = 1.0;
    for (unsigned j = 1; j < n; ++j)
    {
        const double d = r[0] - r[j];
        den[0] *= d;
        den[j] = -d;
    }
    for (unsigned i = 1; i < n; ++i)
        for (unsigned j = i + 1; j < n; ++j)
        {
            const double d = r[i] - r[j];
            den[i] *= +d;
            den[j] *= -d;
        }
}

static void lag_divide_pairs(unsigned n, const double *r, double *v)
{
    for (unsigned i = 0; i < n; ++i)
        for (unsigned j = i + 1; j < n; ++j)
        {
            const double d = r[i] - r[j];
            v[i] /= +d;
            v[j] /= -d;
        }
}

static void lag_tables(unsigned n, const double *r, unsigned m, const double *x, double *B, double *D)
{
    double *den = malloc(sizeof(double) * 3 * n);
    double *col = den + n, *dif = den + 2 * n;
    lag_denominators(n, r, den);
    /* values: the quotient by the denominator comes first, then the remaining differences */
    for (unsigned k = 0; k < m; ++k)
    {
        const double d0 = x[k] - r[0];
        col[0] = 1.0 / den[0];
        for (unsigned j = 1; j < n; ++j)
            col[j] = +d0 / den[j];
        for (unsigned i = 1; i < n; ++i)
        {
            const double di = x[k] - r[i];
            for (unsigned j = 0; j < n; ++j)
                if (j != i)
                    col[j] *= +di;
        }
        for (unsigned j = 0; j < n; ++j)
            B[j * m + k] = col[j];
    }
    /* derivatives */
    for (unsigned k = 0; k < m; ++k)
    {
        for (unsigned j = 0; j < n; ++j)
        {
            dif[j] = x[k] - r[j];
            col[j] = 0.0;
        }
        for (unsigned i = 0; i < n; ++i)
            for (unsigned j = 0; j < i; ++j)
            {
                double term = 1.0;
                for (unsigned l = 0; l < j; ++l)
                    term *= dif[l];
                for (unsigned l = j + 1; l < i; ++l)
                    term *= dif[l];
                for (unsigned l = i + 1; l < n; ++l)
                    term *= dif[l];
                col[j] += term;
                col[i] += term;
            }
        if (k + 2 < m)
        {
            for (unsigned j = 0; j < n; ++j)
                col[j] /= den[j];
        }
        else
        {
            lag_divide_pairs(n, r, col);
        }
        for (unsigned j = 0; j < n; ++j)
            D[j * m + k] = col[j];
    }
    free(den);
}

static void legendre_tables(unsigned order, unsigned m, const double *x, double *B, double *D)
{
    for (unsigned k = 0; k < m; ++k)
    {
        const double xv = x[k];
        B[0 * m + k] = 1.0;
        if (order >= 1)
            B[1 * m + k] = xv;
        double a = 1.0, b = xv;
        for (unsigned i = 2; i < order + 1; ++i)
        {
            const double t1 = (2 * i - 1) * xv;
            const double t2 = (i - 1);
            const double val = (t1 * b - t2 * a) / (double)(i);
            B[i * m + k] = val;
            a = b;
            b = val;
        }
        D[k] = 0;
        double dv = 0;
        for (unsigned i = 1; i <= order; ++i)
        {
            dv = i * B[k + (i - 1) * m] + xv * dv;
            D[k + i * m] = dv;
        }
    }
}

static void bernstein_tables(unsigned n, unsigned m, const double *x, double *B, double *D)
{
    if (n == 0)
    {
        for (unsigned k = 0; k < m; ++k)
            B[k] = 1.0, D[k] = 0.0;
        return;
    }
    for (unsigned k = 0; k < m; ++k)
    {
        const double t = (x[k] + 1) / 2;
        const double a = 1.0 - t;
        const double b = t;
        /* degree n-1 polynomials are parked in the derivative column */
        D[0 * m + k] = 1.0;
        for (unsigned i = 0; i < n - 1; ++i)
        {
            D[(i + 1) * m + k] = D[i * m + k] * b;
            for (unsigned j = i; j > 0; --j)
                D[j * m + k] = b * D[(j - 1) * m + k] + a * D[j * m + k];
            D[0 * m + k] *= a;
        }
        B[n * m + k] = D[(n - 1) * m + k] * b;
        for (unsigned j = n - 1; j > 0; --j)
            B[j * m + k] = b * D[(j - 1) * m + k] + a * D[j * m + k];
        B[k] = D[k] * a;
        D[n * m + k] = n * D[(n - 1) * m + k];
        for (unsigned i = n - 1; i > 0; --i)
            D[i * m + k] = n * (D[(i - 1) * m + k] - D[i * m + k]);
        D[0 * m + k] *= -(double)n;
        for (unsigned i = 0; i < n + 1; ++i)
            D[i * m + k] /= 2.0;
    }
}

int orc_basis_table(int btype, unsigned border, int rtype, unsigned rorder, double *B, double *D)
{
    const unsigned m = rorder + 1, n = border + 1;
    double *x = malloc(sizeof(double) * (2 * m + n));
    double *r = x + 2 * m;
    int res = orc_rule(rtype, rorder, x, x + m);
    if (res == 0)
    {
        if (btype == B_LEGENDRE)
            legendre_tables(border, m, x, B, D);
        else if (btype == B_BERNSTEIN)
            bernstein_tables(border, m, x, B, D);
        else if ((res = orc_roots(btype, border, r)) == 0)
            lag_tables(n, r, m, x, B, D);
    }
    free(x);
    return res;
}

/* ------------------------------------------------------------------ shared element setup -- */

typedef struct
{
    unsigned ndim;
    unsigned n[ORC_MAX_DIM], m[ORC_MAX_DIM];
    size_t N, Q;
    double *B[ORC_MAX_DIM], *D[ORC_MAX_DIM], *w[ORC_MAX_DIM];
} space_tab_t;

static int space_tab_init(space_tab_t *t, unsigned ndim, const int *btype, const unsigned *border, const int *rtype,
                          const unsigned *rorder)
{
    memset(t, 0, sizeof *t);
    if (ndim == 0 || ndim > ORC_MAX_DIM)
        return 102;
    t->ndim = ndim;
    t->N = t->Q = 1;
    for (unsigned k = 0; k < ndim; ++k)
    {
        t->n[k] = border[k] + 1;
        t->m[k] = rorder[k] + 1;
        t->N *= t->n[k];
        t->Q *= t->m[k];
        t->B[k] = malloc(sizeof(double) * (2 * t->n[k] * t->m[k] + 2 * t->m[k]));
        t->D[k] = t->B[k] + t->n[k] * t->m[k];
        t->w[k] = t->D[k] + t->n[k] * t->m[k] + t->m[k];
        int res = orc_basis_table(btype[k], border[k], rtype[k], rorder[k], t->B[k], t->D[k]);
        if (res == 0)
            res = orc_rule(rtype[k], rorder[k], t->w[k] - t->m[k], t->w[k]);
        if (res)
            return res;
    }
    return 0;
}

static void space_tab_free(space_tab_t *t)
{
    for (unsigned k = 0; k < t->ndim; ++k)
        free(t->B[k]);
}

/* row-major unflatten, last index fastest (multidim_iterator_t contract) */
static inline void unflatten(size_t flat, unsigned ndim, const unsigned *ext, unsigned *idx)
{
    for (unsigned k = ndim; k-- > 0;)
    {
        idx[k] = (unsigned)(flat % ext[k]);
        flat /= ext[k];
    }
}

/* ------------------------------------------------------------------ arbitrary points (N2) -- */

/* basis_compute_at_point_values / _derivatives for ONE call of cnt points, /root/reference/src/basis/basis_set.c:462-556:
 * out[p*n + j].  Lagrange: lagrange_polynomial_values_2 (src/polynomials/lagrange.c:108-183: the numerator product is
 * divided by the denominator last, except at the final point of the call where the quotient comes first) and
 * lagrange_polynomial_first_derivative_2 (:383-525: summed pair numerators over the denominator, the last two points
 * divided by the root differences pair by pair). */
static int point_basis(int btype, unsigned order, size_t cnt, const double *x, double *val, double *der)
{
    const unsigned n = order + 1;
    if (btype == B_LEGENDRE || btype == B_BERNSTEIN)
    {
        /* the table builders evaluate point by point with exactly the evaluators of basis_set.c:462-556; they write
         * basis-major [n][cnt], transpose into point-major */
        double *B = malloc(sizeof(double) * 2 * n * (cnt ? cnt : 1));
        double *D = B + n * (cnt ? cnt : 1);
        if (btype == B_LEGENDRE)
            legendre_tables(order, (unsigned)cnt, x, B, D);
        else
            bernstein_tables(order, (unsigned)cnt, x, B, D);
        for (size_t p = 0; p < cnt; ++p)
            for (unsigned j = 0; j < n; ++j)
            {
                if (val)
                    val[p * n + j] = B[j * cnt + p];
                if (der)
                    der[p * n + j] = D[j * cnt + p];
            }
        free(B);
        return 0;
    }
    double *r = malloc(sizeof(double) * 4 * n);
    double *den = r + n, *dif = r + 2 * n, *col = r + 3 * n;
    const int res = orc_roots(btype, order, r);
    if (res)
    {
        free(r);
        return res;
    }
    lag_denominators(n, r, den);
    for (size_t p = 0; p < cnt; ++p)
    {
        if (val)
        {
            double *row = val + p * n;
            if (n == 1)
                row[0] = 1.0;
            else
            {
                const int last = p + 1 == cnt;
                const double d0 = x[p] - r[0];
                row[0] = last ? 1.0 / den[0] : 1.0;
                for (unsigned j = 1; j < n; ++j)
                    row[j] = last ? +d0 / den[j] : +d0;
                for (unsigned i = 1; i < n; ++i)
                {
                    const double di = x[p] - r[i];
                    for (unsigned j = 0; j < n; ++j)
                        if (j != i)
                            row[j] *= +di;
                }
                if (!last)
                    for (unsigned j = 0; j < n; ++j)
                        row[j] /= den[j];
            }
        }
        if (der)
        {
            for (unsigned j = 0; j < n; ++j)
            {
                dif[j] = x[p] - r[j];
                col[j] = 0.0;
            }
            for (unsigned i = 0; i < n; ++i)
                for (unsigned j = 0; j < i; ++j)
                {
                    double term = 1.0;
                    for (unsigned l = 0; l < n; ++l)
                        if (l != i && l != j)
                            term *= dif[l];
                    col[j] += term;
                    col[i] += term;
                }
            if (p + 2 < cnt)
                for (unsigned j = 0; j < n; ++j)
                    col[j] /= den[j];
            else
                lag_divide_pairs(n, r, col);
            for (unsigned j = 0; j < n; ++j)
                der[p * n + j] = col[j];
        }
    }
    free(r);
    return 0;
}

/* FunctionSpace.evaluate per element (basis_compute_outer_product_basis, src/basis/basis_set.c:577-633):
 * x[E][ndim][P] -> out[E][P][N], running product over the dimensions in ascending order starting from 1 */
int orc_evaluate_points(unsigned ndim, const int *btype, const unsigned *border, size_t E, size_t P, const double *x,
                        double *out)
{
    if (ndim == 0 || ndim > ORC_MAX_DIM)
        return 102;
    unsigned n[ORC_MAX_DIM];
    size_t N = 1;
    for (unsigned k = 0; k < ndim; ++k)
        N *= (n[k] = border[k] + 1);
    int fail = 0;
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
    for (size_t e = 0; e < E; ++e)
    {
        double *tab[ORC_MAX_DIM] = {0};
        for (unsigned k = 0; k < ndim; ++k)
        {
            tab[k] = malloc(sizeof(double) * (P ? P : 1) * n[k]);
            if (point_basis(btype[k], border[k], P, x + (e * ndim + k) * P, tab[k], NULL))
                fail = 1;
        }
        for (size_t p = 0; p < P; ++p)
            for (size_t i = 0; i < N; ++i)
            {
                unsigned idx[ORC_MAX_DIM];
                unflatten(i, ndim, n, idx);
                double v = 1.0;
                for (unsigned k = 0; k < ndim; ++k)
                    v = v * tab[k][p * n[k] + idx[k]];
                out[(e * P + p) * N + i] = v;
            }
        for (unsigned k = 0; k < ndim; ++k)
            free(tab[k]);
    }
    return fail ? 104 : 0;
}

/* reconstruct(dof, *x), python/interplib/degrees_of_freedom.py:10-37: sum_i evaluate(x)[p][i] * dofs[i]; with a
 * derivative mask the flagged dimensions use BasisSpecs.derivatives (basis_compute_at_point_derivatives).  The sum over
 * i runs in flat order (NumPy sums the contiguous last axis pairwise in blocks; the difference is O(eps)).
 * dofs[E][N], x[E][ndim][P] -> out[E][n_fields][P] */
int orc_reconstruct_points(unsigned ndim, const int *btype, const unsigned *border, size_t E, const double *dofs, size_t P,
                           const double *x, unsigned n_fields, const unsigned *masks, double *out)
{
    if (ndim == 0 || ndim > ORC_MAX_DIM)
        return 102;
    unsigned n[ORC_MAX_DIM];
    size_t N = 1;
    for (unsigned k = 0; k < ndim; ++k)
        N *= (n[k] = border[k] + 1);
    int fail = 0;
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
    for (size_t e = 0; e < E; ++e)
    {
        double *val[ORC_MAX_DIM] = {0}, *der[ORC_MAX_DIM] = {0};
        for (unsigned k = 0; k < ndim; ++k)
        {
            val[k] = malloc(sizeof(double) * 2 * (P ? P : 1) * n[k]);
            der[k] = val[k] + (P ? P : 1) * n[k];
            if (point_basis(btype[k], border[k], P, x + (e * ndim + k) * P, val[k], der[k]))
                fail = 1;
        }
        for (unsigned f = 0; f < n_fields; ++f)
            for (size_t p = 0; p < P; ++p)
            {
                double acc = 0.0;
                for (size_t i = 0; i < N; ++i)
                {
                    unsigned idx[ORC_MAX_DIM];
                    unflatten(i, ndim, n, idx);
                    double v = 1.0;
                    for (unsigned k = 0; k < ndim; ++k)
                        v = v * (((masks[f] >> k) & 1u) ? der[k] : val[k])[p * n[k] + idx[k]];
                    acc += v * dofs[e * N + i];
                }
                out[(e * n_fields + f) * P + p] = acc;
            }
        for (unsigned k = 0; k < ndim; ++k)
            free(val[k]);
    }
    return fail ? 104 : 0;
}

/* ------------------------------------------------------------------ incidence (N3) -------- */

/* incidence_matrix, /root/reference/src/python/incidence.c:38-110: out[n][n+1], n = order >= 1 */
int orc_incidence_matrix(int btype, unsigned order, double *out)
{
    const unsigned n = order;
    if (n == 0)
        return 102;
    memset(out, 0, sizeof(double) * n * (n + 1));
    if (btype == B_BERNSTEIN)
    {
        const double coeff = (double)n / 2.0;
        out[0 * (n + 1) + 0] = -coeff;
        for (unsigned col = 1; col < n; ++col)
        {
            out[col * (n + 1) + col] = -coeff;
            out[(col - 1) * (n + 1) + col] = +coeff;
        }
        out[(n - 1) * (n + 1) + n] = +coeff;
        return 0;
    }
    if (btype == B_LEGENDRE)
    {
        for (unsigned col = n; col > 0; --col)
        {
            unsigned coeff = 2 * col - 1;
            for (unsigned c_row = 0; 2 * c_row < col; ++c_row)
            {
                const unsigned r = (col - 1 - 2 * c_row);
                out[r * (n + 1) + col] = coeff;
                coeff -= 4;
            }
        }
        return 0;
    }
    /* Lagrange: first derivative of the order-n basis at the nodes of the order n-1 basis */
    double *nodes = malloc(sizeof(double) * n);
    int res = orc_roots(btype, n - 1, nodes);
    if (res == 0)
        res = point_basis(btype, n, n, nodes, NULL, out);
    free(nodes);
    return res;
}

/* DegreesOfFreedom.derivative(idim), src/python/degrees_of_freedom.c:738-945 (the three apply routines :738-851),
 * per element: in[E][pre][n+1][post] -> out[E][pre][n][post] */
int orc_dof_derivative(unsigned ndim, const int *btype, const unsigned *border, unsigned idim, size_t E, const double *in,
                       double *out)
{
    if (ndim == 0 || ndim > ORC_MAX_DIM || idim >= ndim || border[idim] == 0)
        return 102;
    size_t pre = 1, post = 1;
    for (unsigned k = 0; k < idim; ++k)
        pre *= border[k] + 1;
    for (unsigned k = idim + 1; k < ndim; ++k)
        post *= border[k] + 1;
    const unsigned n = border[idim];
    const int type = btype[idim];
    double *mat = NULL;
    if (type != B_BERNSTEIN && type != B_LEGENDRE)
    {
        mat = malloc(sizeof(double) * n * (n + 1));
        const int res = orc_incidence_matrix(type, n, mat);
        if (res)
        {
            free(mat);
            return res;
        }
    }
    memset(out, 0, sizeof(double) * E * pre * n * post); /* a fresh DegreesOfFreedom object holds zeros */
    for (size_t e = 0; e < E; ++e)
        for (size_t i_pre = 0; i_pre < pre; ++i_pre)
            for (size_t i_post = 0; i_post < post; ++i_post)
            {
                double *po = out + e * pre * n * post + i_pre * n * post + i_post;
                const double *pi = in + e * pre * (n + 1) * post + i_pre * (n + 1) * post + i_post;
                if (type == B_BERNSTEIN)
                {
                    const double coeff = (double)n / 2.0;
                    po[0] -= coeff * pi[0];
                    for (unsigned col = 1; col < n; ++col)
                    {
                        const double x = coeff * pi[col * post];
                        po[col * post] -= x;
                        po[(col - 1) * post] += x;
                    }
                    po[(n - 1) * post] += coeff * pi[n * post];
                }
                else if (type == B_LEGENDRE)
                {
                    for (unsigned col = n; col > 0; --col)
                    {
                        unsigned coeff = 2 * col - 1;
                        for (unsigned c_row = 0; 2 * c_row < col; ++c_row)
                        {
                            const unsigned r = (col - 1 - 2 * c_row);
                            po[r * post] += coeff * pi[col * post];
                            coeff -= 4;
                        }
                    }
                }
                else
                {
                    for (unsigned row = 0; row < n; ++row)
                    {
                        double v = 0;
                        for (unsigned col = 0; col < n + 1; ++col)
                            v += mat[row * (n + 1) + col] * pi[col * post];
                        po[row * post] = v;
                    }
                }
            }
    free(mat);
    return 0;
}

int orc_space_weights(unsigned ndim, const int *rtype, const unsigned *rorder, double *W)
{
    int bt[ORC_MAX_DIM];
    unsigned bo[ORC_MAX_DIM];
    for (unsigned k = 0; k < ndim; ++k)
        bt[k] = B_LEGENDRE, bo[k] = 0;
    space_tab_t t;
    int res = space_tab_init(&t, ndim, bt, bo, rtype, rorder);
    if (res == 0)
    {
        for (size_t q = 0; q < t.Q; ++q)
        {
            unsigned qi[ORC_MAX_DIM];
            unflatten(q, ndim, t.m, qi);
            double v = 1.0;
            for (unsigned k = 0; k < ndim; ++k)
                v = v * t.w[k][qi[k]];
            W[q] = v;
        }
    }
    space_tab_free(&t);
    return res;
}

/* ------------------------------------------------------------------ A5 / A6 ---------------- */

static void reconstruct_one(const space_tab_t *t, unsigned deriv_mask, const double *dofs, double *out)
{
    for (size_t q = 0; q < t->Q; ++q)
    {
        unsigned qi[ORC_MAX_DIM], bi[ORC_MAX_DIM];
        unflatten(q, t->ndim, t->m, qi);
        double val = 0;
        for (size_t i = 0; i < t->N; ++i)
        {
            unflatten(i, t->ndim, t->n, bi);
            double phi = 1;
            for (unsigned k = 0; k < t->ndim; ++k)
            {
                const double *tab = ((deriv_mask >> k) & 1u) ? t->D[k] : t->B[k];
                phi *= tab[bi[k] * t->m[k] + qi[k]];
            }
            val += phi * dofs[i];
        }
        out[q] = val;
    }
}

int orc_reconstruct(unsigned ndim, const int *btype, const unsigned *border, const int *rtype,
                    const unsigned *rorder, size_t E, const double *dofs, unsigned n_fields,
                    const unsigned *deriv_masks, double *out)
{
    space_tab_t t;
    int res = space_tab_init(&t, ndim, btype, border, rtype, rorder);
    if (res == 0)
    {
#pragma omp parallel for schedule(static)
        for (long long e = 0; e < (long long)E; ++e)
            for (unsigned f = 0; f < n_fields; ++f)
                reconstruct_one(&t, deriv_masks[f], dofs + (size_t)e * t.N, out + ((size_t)e * n_fields + f) * t.Q);
    }
    space_tab_free(&t);
    return res;
}

/* ------------------------------------------------------------------ A9 geometry ------------ */

/* Givens QR of a rows x cols matrix (row major, in place -> R), qt accumulates Q^T */
static void givens_qr(unsigned rows, unsigned cols, double *r, double *qt)
{
    for (unsigned i = 0; i < rows; ++i)
        for (unsigned j = 0; j < rows; ++j)
            qt[i * rows + j] = (double)(i == j);
    if (rows == 1)
        return;
    for (unsigned row = 0; row < rows; ++row)
        for (unsigned col = 0; col < cols && col < row; ++col)
        {
            double c = r[col * cols + col];
            double s = -r[row * cols + col];
            const double mag = hypot(c, s);
            if (mag < 1e-12)
                continue;
            c /= mag;
            s /= mag;
            for (unsigned k = col; k < cols; ++k)
            {
                const double a = r[col * cols + k];
                const double b = r[row * cols + k];
                r[col * cols + k] = +c * a - s * b;
                r[row * cols + k] = +s * a + c * b;
            }
            r[row * cols + col] = 0;
            for (unsigned k = 0; k < rows; ++k)
            {
                const double a = qt[col * rows + k];
                qt[col * rows + k] = +c * a - s * qt[row * rows + k];
                qt[row * rows + k] = +s * a + c * qt[row * rows + k];
            }
        }
}

static void back_substitute(unsigned u_cols, const double *upper, unsigned b_cols, double *b)
{
    for (unsigned bc = 0; bc < b_cols; ++bc)
    {
        double *p = b + bc;
        for (unsigned row = u_cols; row > 0; --row)
        {
            double v = p[(row - 1) * b_cols];
            for (unsigned col = row; col < u_cols; ++col)
                v -= upper[(row - 1) * u_cols + col] * p[col * b_cols];
            p[(row - 1) * b_cols] = v / upper[(row - 1) * u_cols + (row - 1)];
        }
    }
}

/* coords[n_phys][1+ndim][Q] in, det[Q], jinv[Q][ndim][n_phys] out */
static void space_map_one(unsigned ndim, unsigned n_phys, size_t Q, const double *coords, double *det, double *jinv)
{
    double jac[ORC_MAX_DIM * ORC_MAX_DIM], qt[ORC_MAX_DIM * ORC_MAX_DIM];
    const unsigned rows = n_phys, cols = ndim;
    for (size_t q = 0; q < Q; ++q)
    {
        for (unsigned a = 0; a < rows; ++a)
            for (unsigned b = 0; b < cols; ++b)
                jac[a * cols + b] = coords[((size_t)a * (1 + ndim) + 1 + b) * Q + q];
        givens_qr(rows, cols, jac, qt);
        double d = 1;
        for (unsigned i = 0; i < cols; ++i)
            d *= jac[i * cols + i];
        if (det)
            det[q] = d;
        if (jinv)
        {
            double *inv = jinv + q * (size_t)(rows * cols);
            for (unsigned a = 0; a < cols; ++a)
                for (unsigned b = 0; b < rows; ++b)
                    inv[a * rows + b] = qt[a * rows + b];
            back_substitute(cols, jac, rows, inv);
        }
    }
}

int orc_geometry(unsigned ndim, unsigned n_phys, const int *gtype, const unsigned *gorder, const int *rtype,
                 const unsigned *rorder, size_t E, const double *geom_dofs, double *coords, double *det,
                 double *jinv)
{
    if (n_phys < ndim || n_phys > ORC_MAX_DIM)
        return 102;
    space_tab_t t;
    int res = space_tab_init(&t, ndim, gtype, gorder, rtype, rorder);
    if (res == 0)
    {
        const size_t cstride = (size_t)n_phys * (1 + ndim) * t.Q;
#pragma omp parallel for schedule(static)
        for (long long e = 0; e < (long long)E; ++e)
        {
            double *c = coords ? coords + (size_t)e * cstride : malloc(sizeof(double) * cstride);
            for (unsigned a = 0; a < n_phys; ++a)
            {
                const double *dof = geom_dofs + ((size_t)e * n_phys + a) * t.N;
                reconstruct_one(&t, 0u, dof, c + ((size_t)a * (1 + ndim)) * t.Q);
                for (unsigned k = 0; k < ndim; ++k)
                    reconstruct_one(&t, 1u << k, dof, c + ((size_t)a * (1 + ndim) + 1 + k) * t.Q);
            }
            if (det || jinv)
                space_map_one(ndim, n_phys, t.Q, c, det ? det + (size_t)e * t.Q : NULL,
                              jinv ? jinv + (size_t)e * t.Q * ndim * n_phys : NULL);
            if (!coords)
                free(c);
        }
    }
    space_tab_free(&t);
    return res;
}

/* ------------------------------------------------------------------ A7 / A8 mass ----------- */

/* One element.  det may be NULL (IntegrationSpace branch).  deriv_in < 0: mass matrix; else the
 * gradient mass matrix with the input basis differentiated along deriv_in and, when a map is
 * present, weight det*jinv[q][deriv_in][idx_out]. */
static void mass_one(const space_tab_t *tin, const space_tab_t *tout, int symmetric, const double *det,
                     const double *jinv, unsigned n_phys, int deriv_in, int idx_out, double *out)
{
    const unsigned ndim = tin->ndim;
    const size_t Nin = tin->N, Nout = tout->N, Q = tin->Q;
    for (size_t jo = 0; jo < Nout; ++jo)
    {
        unsigned bo[ORC_MAX_DIM];
        unflatten(jo, ndim, tout->n, bo);
        const size_t in_end = symmetric ? jo + 1 : Nin;
        for (size_t ii = 0; ii < in_end; ++ii)
        {
            unsigned bi[ORC_MAX_DIM];
            unflatten(ii, ndim, tin->n, bi);
            double acc = 0;
            for (size_t q = 0; q < Q; ++q)
            {
                unsigned qi[ORC_MAX_DIM];
                unflatten(q, ndim, tin->m, qi);
                double wgt;
                if (deriv_in < 0)
                    wgt = det ? det[q] : 1;
                else
                    wgt = det ? det[q] * jinv[q * (size_t)(ndim * n_phys) + (size_t)deriv_in * n_phys + idx_out] : 1;
                for (unsigned k = 0; k < ndim; ++k)
                    wgt *= tin->w[k][qi[k]];
                double pin = 1;
                for (unsigned k = 0; k < ndim; ++k)
                    pin *= ((int)k == deriv_in ? tin->D[k] : tin->B[k])[bi[k] * tin->m[k] + qi[k]];
                double pout = 1;
                for (unsigned k = 0; k < ndim; ++k)
                    pout *= tout->B[k][bo[k] * tout->m[k] + qi[k]];
                acc += wgt * pin * pout;
            }
            out[jo * Nin + ii] = acc;
        }
    }
    if (symmetric)
        for (size_t i = 0; i < Nout; ++i)
            for (size_t j = i + 1; j < Nin; ++j)
                out[i * Nin + j] = out[j * Nin + i];
}

/* det[E][Q] / jinv[E][Q][ndim][n_phys] may be NULL; deriv_in = -1 for the plain mass matrix.
 * Returns the reference's early all-zero answer when there is no map and deriv_in != idx_out. */
int orc_mass(unsigned ndim, const int *btype_in, const unsigned *border_in, const int *btype_out,
             const unsigned *border_out, const int *rtype, const unsigned *rorder, size_t E, const double *det,
             const double *jinv, unsigned n_phys, int deriv_in, int idx_out, double *out)
{
    space_tab_t tin, tout;
    int res = space_tab_init(&tin, ndim, btype_in, border_in, rtype, rorder);
    if (res == 0)
        res = space_tab_init(&tout, ndim, btype_out, border_out, rtype, rorder);
    else
        memset(&tout, 0, sizeof tout);
    if (res == 0)
    {
        int symmetric = 1;
        for (unsigned k = 0; k < ndim; ++k)
            if (btype_in[k] != btype_out[k] || border_in[k] != border_out[k])
                symmetric = 0;
        const size_t per = tin.N * tout.N;
        if (deriv_in >= 0 && det == NULL && deriv_in != idx_out)
        {
            memset(out, 0, sizeof(double) * per * E);
        }
        else
        {
#pragma omp parallel for schedule(static)
            for (long long e = 0; e < (long long)E; ++e)
                mass_one(&tin, &tout, symmetric, det ? det + (size_t)e * tin.Q : NULL,
                         jinv ? jinv + (size_t)e * tin.Q * ndim * n_phys : NULL, n_phys, deriv_in, idx_out,
                         out + (size_t)e * per);
        }
    }
    space_tab_free(&tin);
    space_tab_free(&tout);
    return res;
}

/* ------------------------------------------------------------------ A12 L2 dual ------------ */

/* numpy semantics of integration.py:143-166: fw = f * (W * det) elementwise, then the column sums
 * of fw[:,None] * Phi[Q][N] taken row after row (numpy reduces axis 0 of a C-contiguous array by
 * adding whole rows in order). */
int orc_l2_dual(unsigned ndim, const int *btype, const unsigned *border, const int *rtype, const unsigned *rorder,
                size_t E, const double *fvals, const double *det, double *out)
{
    space_tab_t t;
    int res = space_tab_init(&t, ndim, btype, border, rtype, rorder);
    if (res == 0)
    {
        double *W = malloc(sizeof(double) * t.Q);
        double *Phi = malloc(sizeof(double) * t.Q * t.N);
        for (size_t q = 0; q < t.Q; ++q)
        {
            unsigned qi[ORC_MAX_DIM], bi[ORC_MAX_DIM];
            unflatten(q, ndim, t.m, qi);
            double v = 1.0;
            for (unsigned k = 0; k < ndim; ++k)
                v = v * t.w[k][qi[k]];
            W[q] = v;
            for (size_t i = 0; i < t.N; ++i)
            {
                unflatten(i, ndim, t.n, bi);
                double phi = 1;
                for (unsigned k = 0; k < ndim; ++k)
                    phi *= t.B[k][bi[k] * t.m[k] + qi[k]];
                Phi[q * t.N + i] = phi;
            }
        }
#pragma omp parallel for schedule(static)
        for (long long e = 0; e < (long long)E; ++e)
        {
            double *b = out + (size_t)e * t.N;
            for (size_t q = 0; q < t.Q; ++q)
            {
                const double wq = det ? W[q] * det[(size_t)e * t.Q + q] : W[q];
                const double fw = fvals[(size_t)e * t.Q + q] * wq;
                for (size_t i = 0; i < t.N; ++i)
                {
                    const double term = fw * Phi[q * t.N + i];
                    b[i] = (q == 0) ? term : b[i] + term;
                }
            }
        }
        free(Phi);
        free(W);
    }
    space_tab_free(&t);
    return res;
}

/* number of OpenMP threads this build uses (1 when built without -fopenmp) */
#ifdef _OPENMP
#include <omp.h>
int orc_threads(void)
{
    return omp_get_max_threads();
}
#else
int orc_threads(void)
{
    return 1;
}
#endif
