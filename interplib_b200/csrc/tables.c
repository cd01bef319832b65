/* Host-side 1-D tables of the interplib hot path.
 *
 * These are the inputs the CUDA kernels consume: quadrature nodes/weights and the basis value /
 * derivative tables B[n][m], D[n][m].  They are computed once per (basis type, order, rule type,
 * rule order) on the host and uploaded; to make the kernels' inputs identical to the reference's,
 * every formula below keeps the reference's floating-point operation order (so the tables compare
 * `==` against the compiled reference), but the code is organised around closed forms of what the
 * reference's in-place loops compute, e.g. the Lagrange denominator of root j is the ascending
 * product prod_{k != j} (r_j - r_k), which is what src/polynomials/lagrange.c:233-258 builds
 * incrementally.
 *
 * Compiled with -ffp-contract=off: a fused multiply-add would change the last bit.
 */
#include "tables.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ---- Legendre three-term recursion (src/polynomials/legendre.c:7-24, 89-116) -------------- */

/* P_{deg-1}(x), P_deg(x) for deg >= 1 */
static void legendre_last_two(unsigned deg, double x, double *p_prev, double *p_last)
{
    double lo = 1.0, hi = x;
    for (unsigned k = 2; k <= deg; ++k)
    {
        const double a = (double)(2 * k - 1) * x;
        const double b = (double)(k - 1);
        const double next = (a * hi - b * lo) / (double)k;
        lo = hi;
        hi = next;
    }
    *p_prev = lo;
    *p_last = hi;
}

/* P_0..P_deg at x, written with a stride */
static void legendre_all(unsigned deg, double x, double *out, size_t stride)
{
    out[0] = 1.0;
    if (deg == 0)
        return;
    out[stride] = x;
    double lo = 1.0, hi = x;
    for (unsigned k = 2; k <= deg; ++k)
    {
        const double a = (double)(2 * k - 1) * x;
        const double b = (double)(k - 1);
        const double next = (a * hi - b * lo) / (double)k;
        out[(size_t)k * stride] = next;
        lo = hi;
        hi = next;
    }
}

/* ---- A1: Gauss-Legendre / Gauss-Lobatto (src/integration/gauss_legendre.c:7-50,
 *      src/integration/gauss_lobatto.c:14-61); w may be NULL (node-only variants :53-90 / :64-106) */

static int trivial_rule(unsigned n, double *x, double *w)
{
    if (n == 1)
    {
        x[0] = 0.0;
        if (w)
            w[0] = 2.0;
        return 1;
    }
    if (n == 2)
    {
        /* reference quirk kept for parity: the 2-point rule is (-1, +1) with unit weights */
        x[0] = -1.0;
        x[1] = +1.0;
        if (w)
            w[0] = w[1] = 1.0;
        return 1;
    }
    return 0;
}

static void gauss_legendre(unsigned n, double tol, unsigned max_iter, double *x, double *w)
{
    if (trivial_rule(n, x, w))
        return;
    for (unsigned i = 0; i < n; ++i)
    {
        double xi = cos(M_PI * (double)(4 * i + 3) / (double)(2 * (2 * n + 1)));
        double err = 1.0;
        for (unsigned it = 0; it < max_iter && err > tol; ++it)
        {
            double pm, pn;
            legendre_last_two(n, xi, &pm, &pn);
            const double one_minus_x2 = 1 - xi * xi;
            /* the reference scales the slope by (n-1), not n; kept */
            const double slope = (double)(n - 1) * (pm - xi * pn);
            const double step = pn / slope * one_minus_x2;
            xi -= step;
            err = fabs(step);
        }
        x[n - 1 - i] = xi;
        if (w)
        {
            double pm, pn;
            legendre_last_two(n - 1, xi, &pm, &pn);
            w[n - 1 - i] = 2.0 / ((double)(n * n) * pn * pn) * (1 - xi * xi);
        }
    }
}

static int gauss_lobatto(unsigned n, double tol, unsigned max_iter, double *x, double *w)
{
    int non_converged = 0; /* nodes whose last Newton step was still above tol (the reference's return value) */
    if (trivial_rule(n, x, w))
        return 0;
    x[0] = -1.0;
    x[n - 1] = +1.0;
    if (w)
        w[n - 1] = w[0] = 2.0 / (double)(n * (n - 1));
    const double c1 = (1.0 - 3.0 * (double)(n - 2) / (double)(8.0 * (double)(n - 1) * (double)(n - 1) * (double)(n - 1)));
    const double c2 = M_PI / (4.0 * (double)(n - 1) + 1);
    for (unsigned i = 2; i < n; ++i)
    {
        double xi = c1 * cos(c2 * (double)(4 * i - 3));
        double err = 1.0;
        for (unsigned it = 0; it < max_iter && err > tol; ++it)
        {
            double pm, pn;
            legendre_last_two(n - 1, xi, &pm, &pn);
            const double den = 1 - xi * xi;
            const double d1 = (double)(n - 1) * (pm - xi * pn) / den;
            const double d2 = (2 * xi * d1 - (double)((n - 1) * n) * pn) / den;
            const double d3 = (4 * xi * d2 - (double)((n - 1) * n - 2) * d1) / den;
            const double step = 2 * d1 * d2 / (2 * d2 * d2 - d1 * d3);
            xi -= step;
            err = fabs(step);
        }
        non_converged += (err > tol);
        x[n - i] = xi;
        if (w)
        {
            double pm, pn;
            legendre_last_two(n - 1, xi, &pm, &pn);
            w[n - i] = 2.0 / ((double)(n * (n - 1)) * pn * pn);
        }
    }
    return non_converged;
}

int ib200_tables_rule(integration_spec_t spec, double *nodes, double *weights, unsigned *accuracy)
{
    /* integration_rule_for_order, src/integration/integration_rules.c:40-73 */
    const unsigned order = spec.order;
    switch (spec.type)
    {
    case INTEGRATION_RULE_TYPE_GAUSS_LEGENDRE:
        if (accuracy)
            *accuracy = order > 0 ? 2 * order - 1 : 0;
        gauss_legendre(order + 1, 1e-14, 1000, nodes, weights);
        return INTERP_SUCCESS;
    case INTEGRATION_RULE_TYPE_GAUSS_LOBATTO:
        if (accuracy)
            *accuracy = order > 1 ? 2 * order - 3 : order;
        gauss_lobatto(order + 1, 1e-14, 1000, nodes, weights);
        return INTERP_SUCCESS;
    default:
        return INTERP_ERROR_INVALID_ENUM;
    }
}

int ib200_tables_order_for_accuracy(integration_rule_type_t type, unsigned accuracy, unsigned *order)
{
    /* integration_rule_for_accuracy, src/integration/integration_rules.c:15-39 */
    switch (type)
    {
    case INTEGRATION_RULE_TYPE_GAUSS_LEGENDRE:
        *order = accuracy < 2 ? 1 : accuracy / 2 + 1;
        return INTERP_SUCCESS;
    case INTEGRATION_RULE_TYPE_GAUSS_LOBATTO:
        *order = accuracy / 2 + 2;
        return INTERP_SUCCESS;
    default:
        return INTERP_ERROR_INVALID_ENUM;
    }
}

/* ---- A2: Lagrange nodes (src/basis/basis_lagrange.c:11-42, same as
 *      basis_compute_at_point_prepare, src/basis/basis_set.c:427-460) ------------------------- */

int ib200_tables_is_lagrange(basis_set_type_t type)
{
    return type == BASIS_LAGRANGE_UNIFORM || type == BASIS_LAGRANGE_CHEBYSHEV_GAUSS || type == BASIS_LAGRANGE_GAUSS ||
           type == BASIS_LAGRANGE_GAUSS_LOBATTO;
}

int ib200_tables_lagrange_roots(basis_spec_t spec, double *roots)
{
    const unsigned n = spec.order + 1;
    switch (spec.type)
    {
    case BASIS_LAGRANGE_UNIFORM:
        for (unsigned i = 0; i < n; ++i)
            roots[i] = (2.0 * i) / (double)spec.order - 1.0;
        return INTERP_SUCCESS;
    case BASIS_LAGRANGE_CHEBYSHEV_GAUSS:
        for (unsigned i = 0; i < n; ++i)
            roots[i] = -cos(M_PI * (double)(2 * i + 1) / (double)(2 * n));
        return INTERP_SUCCESS;
    case BASIS_LAGRANGE_GAUSS:
        gauss_legendre(n, 1e-12, 100, roots, NULL);
        return INTERP_SUCCESS;
    case BASIS_LAGRANGE_GAUSS_LOBATTO:
        gauss_lobatto(n, 1e-12, 100, roots, NULL);
        return INTERP_SUCCESS;
    default:
        return INTERP_ERROR_INVALID_ENUM;
    }
}

/* ---- A3: Lagrange value / derivative tables ------------------------------------------------ */

/* prod_{k != j, ascending} (r_j - r_k): what lagrange_polynomial_denominators[_stride]
 * (src/polynomials/lagrange.c:8-30, 233-258) accumulates; -(r_i - r_j) == r_j - r_i exactly. */
static double lagrange_denominator(unsigned n, const double *r, unsigned j)
{
    double d = 1.0;
    for (unsigned k = 0; k < n; ++k)
        if (k != j)
            d *= r[j] - r[k];
    return d;
}

/* roots[n] then denominators[n] of a Lagrange basis (device arbitrary-point evaluation, kernels_points.cu) */
int ib200_tables_lagrange_roots_denominators(basis_spec_t spec, double *roots, double *den)
{
    const unsigned n = spec.order + 1;
    const int res = ib200_tables_lagrange_roots(spec, roots);
    if (res != INTERP_SUCCESS)
        return res;
    for (unsigned j = 0; j < n; ++j)
        den[j] = lagrange_denominator(n, roots, j);
    return INTERP_SUCCESS;
}

/* ((v / (r_j - r_0)) / (r_j - r_1)) ... : lagrange_polynomial_denominators_apply_stride,
 * src/polynomials/lagrange.c:261-274 */
static double lagrange_divide_chain(unsigned n, const double *r, unsigned j, double v)
{
    for (unsigned k = 0; k < n; ++k)
        if (k != j)
            v /= r[j] - r[k];
    return v;
}

/* numerator of l_j(x) with the division by `den` placed either first ("early", what the _2
 * routines do while the denominators are still parked in the output) or last */
static double lagrange_value(unsigned n, const double *r, unsigned j, double x, double den, int early)
{
    double acc = (j == 0) ? 1.0 : x - r[0];
    if (early)
        acc = acc / den;
    for (unsigned i = 1; i < n; ++i)
        if (i != j)
            acc *= x - r[i];
    return early ? acc : acc / den;
}

/* sum over the symmetric pair numerators prod_{k != i,j} (x - r_k), accumulated in the order of
 * src/polynomials/lagrange.c:626-648: pairs (i, j<i) ascending in i then j, each numerator added
 * to slot j and then to slot i.  num[] has n entries. */
static void lagrange_derivative_numerators(unsigned n, const double *r, double x, double *dif, double *num)
{
    for (unsigned j = 0; j < n; ++j)
    {
        dif[j] = x - r[j];
        num[j] = 0.0;
    }
    for (unsigned i = 0; i < n; ++i)
    {
        for (unsigned j = 0; j < i; ++j)
        {
            double prod = 1.0;
            for (unsigned k = 0; k < n; ++k)
                if (k != i && k != j)
                    prod *= dif[k];
            num[j] += prod;
            num[i] += prod;
        }
    }
}

/* lagrange_polynomial_values_transposed_2 (src/polynomials/lagrange.c:277-315): B[j][k], every
 * position divides first.  lagrange_polynomial_first_derivative_transposed_2 (:594-737):
 * positions 0..m-3 divide the summed numerator by the product denominator, the last two
 * positions divide by the root differences one after another. */
static int lagrange_tables(basis_spec_t spec, unsigned m, const double *x, double *values, double *derivs)
{
    const unsigned n = spec.order + 1;
    double *buf = malloc(sizeof(double) * 4 * n);
    if (!buf)
        return INTERP_ERROR_FAILED_ALLOCATION;
    double *roots = buf, *den = buf + n, *dif = buf + 2 * n, *num = buf + 3 * n;
    const int res = ib200_tables_lagrange_roots(spec, roots);
    if (res != INTERP_SUCCESS)
    {
        free(buf);
        return res;
    }
    for (unsigned j = 0; j < n; ++j)
        den[j] = lagrange_denominator(n, roots, j);
    for (unsigned k = 0; k < m; ++k)
    {
        if (values)
            for (unsigned j = 0; j < n; ++j)
                values[(size_t)j * m + k] = lagrange_value(n, roots, j, x[k], den[j], 1);
        if (derivs)
        {
            lagrange_derivative_numerators(n, roots, x[k], dif, num);
            const int chained = (k + 2 >= m);
            for (unsigned j = 0; j < n; ++j)
                derivs[(size_t)j * m + k] = chained ? lagrange_divide_chain(n, roots, j, num[j]) : num[j] / den[j];
        }
    }
    free(buf);
    return INTERP_SUCCESS;
}

/* ---- A4: Legendre and Bernstein tables ------------------------------------------------------ */

/* legendre_basis_create, src/basis/basis_legendre.c:9-40 */
static void legendre_tables(unsigned order, unsigned m, const double *x, double *values, double *derivs)
{
    for (unsigned k = 0; k < m; ++k)
    {
        legendre_all(order, x[k], values + k, m);
        double d = 0;
        derivs[k] = 0;
        for (unsigned i = 1; i <= order; ++i)
        {
            d = (double)i * values[k + (size_t)(i - 1) * m] + x[k] * d;
            derivs[k + (size_t)i * m] = d;
        }
    }
}

/* Bernstein polynomials of degree `deg` at t in [0,1] by the degree-raising recursion
 * (src/polynomials/bernstein.c:31-51) */
static void bernstein_vector(double t, unsigned deg, double *out)
{
    const double a = 1.0 - t, b = t;
    out[0] = 1.0;
    for (unsigned i = 0; i < deg; ++i)
    {
        out[i + 1] = out[i] * b;
        for (unsigned j = i; j > 0; --j)
            out[j] = b * out[j - 1] + a * out[j];
        out[0] *= a;
    }
}

/* bernstein_interpolation_value_derivative_matrix, src/polynomials/bernstein.c:53-119 */
static int bernstein_tables(unsigned order, unsigned m, const double *x, double *values, double *derivs)
{
    if (order == 0)
    {
        for (unsigned k = 0; k < m; ++k)
        {
            values[k] = 1.0;
            derivs[k] = 0.0;
        }
        return INTERP_SUCCESS;
    }
    double *low = malloc(sizeof(double) * (order + 1));
    if (!low)
        return INTERP_ERROR_FAILED_ALLOCATION;
    const unsigned n = order;
    for (unsigned k = 0; k < m; ++k)
    {
        const double t = (x[k] + 1) / 2;
        const double a = 1.0 - t, b = t;
        bernstein_vector(t, n - 1, low); /* degree n-1 */
        /* one more degree-raising step gives the values */
        values[(size_t)n * m + k] = low[n - 1] * b;
        for (unsigned j = n - 1; j > 0; --j)
            values[(size_t)j * m + k] = b * low[j - 1] + a * low[j];
        values[k] = low[0] * a;
        /* differences of the degree n-1 polynomials give the derivative; /2 is the chain rule of
         * the map [-1,1] -> [0,1] */
        derivs[(size_t)n * m + k] = ((double)n * low[n - 1]) / 2.0;
        for (unsigned i = n - 1; i > 0; --i)
            derivs[(size_t)i * m + k] = ((double)n * (low[i - 1] - low[i])) / 2.0;
        derivs[k] = (low[0] * -(double)n) / 2.0;
    }
    free(low);
    return INTERP_SUCCESS;
}

int ib200_tables_basis(basis_spec_t basis, integration_spec_t rule, double *values, double *derivs)
{
    /* basis_set_create, src/basis/basis_set.c:105-131 */
    const unsigned m = rule.order + 1;
    double *x = malloc(sizeof(double) * 2 * m);
    if (!x)
        return INTERP_ERROR_FAILED_ALLOCATION;
    int res = ib200_tables_rule(rule, x, x + m, NULL);
    if (res == INTERP_SUCCESS)
    {
        switch (basis.type)
        {
        case BASIS_LEGENDRE:
            legendre_tables(basis.order, m, x, values, derivs);
            break;
        case BASIS_BERNSTEIN:
            res = bernstein_tables(basis.order, m, x, values, derivs);
            break;
        case BASIS_LAGRANGE_GAUSS:
        case BASIS_LAGRANGE_UNIFORM:
        case BASIS_LAGRANGE_GAUSS_LOBATTO:
        case BASIS_LAGRANGE_CHEBYSHEV_GAUSS:
            res = lagrange_tables(basis, m, x, values, derivs);
            break;
        default:
            res = INTERP_ERROR_INVALID_ENUM;
        }
    }
    free(x);
    return res;
}

/* ---- A10: tensor weights / nodes (src/python/integration_objects.c:633-716, 731-814) -------- */

static size_t rule_points(unsigned ndim, const integration_spec_t *rules, unsigned *m)
{
    size_t q = 1;
    for (unsigned k = 0; k < ndim; ++k)
    {
        m[k] = rules[k].order + 1;
        q *= m[k];
    }
    return q;
}

int ib200_tables_space_weights(unsigned ndim, const integration_spec_t *rules, double *weights)
{
    if (ndim == 0 || ndim > 8)
        return IB200_ERROR_BAD_ARGUMENT;
    unsigned m[8];
    const size_t total = rule_points(ndim, rules, m);
    unsigned m_max = 0;
    for (unsigned k = 0; k < ndim; ++k)
        m_max = m[k] > m_max ? m[k] : m_max;
    double *buf = malloc(sizeof(double) * 2 * (size_t)m_max * ndim);
    if (!buf)
        return INTERP_ERROR_FAILED_ALLOCATION;
    for (unsigned k = 0; k < ndim; ++k)
    {
        const int res = ib200_tables_rule(rules[k], buf + 2 * (size_t)m_max * k, buf + 2 * (size_t)m_max * k + m_max, NULL);
        if (res != INTERP_SUCCESS)
        {
            free(buf);
            return res;
        }
    }
    /* W[q] = ((1 * w0[q0]) * w1[q1]) * ... : the reference multiplies dimension by dimension */
    for (size_t q = 0; q < total; ++q)
    {
        size_t rem = q;
        unsigned idx[8];
        for (unsigned k = ndim; k-- > 0;)
        {
            idx[k] = (unsigned)(rem % m[k]);
            rem /= m[k];
        }
        double w = 1.0;
        for (unsigned k = 0; k < ndim; ++k)
            w = w * buf[2 * (size_t)m_max * k + m_max + idx[k]];
        weights[q] = w;
    }
    free(buf);
    return INTERP_SUCCESS;
}

int ib200_tables_space_nodes(unsigned ndim, const integration_spec_t *rules, double *nodes)
{
    if (ndim == 0 || ndim > 8)
        return IB200_ERROR_BAD_ARGUMENT;
    unsigned m[8];
    const size_t total = rule_points(ndim, rules, m);
    for (unsigned k = 0; k < ndim; ++k)
    {
        double *x = malloc(sizeof(double) * 2 * m[k]);
        if (!x)
            return INTERP_ERROR_FAILED_ALLOCATION;
        const int res = ib200_tables_rule(rules[k], x, x + m[k], NULL);
        if (res != INTERP_SUCCESS)
        {
            free(x);
            return res;
        }
        size_t inner = 1;
        for (unsigned l = k + 1; l < ndim; ++l)
            inner *= m[l];
        for (size_t q = 0; q < total; ++q)
            nodes[(size_t)k * total + q] = x[(q / inner) % m[k]];
        free(x);
    }
    return INTERP_SUCCESS;
}

/* ---- A11: function_space_values_at_integration_nodes
 *      (src/python/function_space_objects.c:301-474): out[q][i] = prod_dim B_dim[i_dim][q_dim] -- */

int ib200_tables_values_at_nodes(unsigned ndim, const basis_spec_t *basis, const integration_spec_t *rules,
                                 double *out)
{
    if (ndim == 0 || ndim > 8)
        return IB200_ERROR_BAD_ARGUMENT;
    unsigned m[8], n[8];
    double *tab[8] = {0};
    size_t Q = 1, N = 1;
    int res = INTERP_SUCCESS;
    for (unsigned k = 0; k < ndim && res == INTERP_SUCCESS; ++k)
    {
        m[k] = rules[k].order + 1;
        n[k] = basis[k].order + 1;
        Q *= m[k];
        N *= n[k];
        tab[k] = malloc(sizeof(double) * 2 * (size_t)m[k] * n[k]);
        res = tab[k] ? ib200_tables_basis(basis[k], rules[k], tab[k], tab[k] + (size_t)m[k] * n[k])
                     : INTERP_ERROR_FAILED_ALLOCATION;
    }
    if (res == INTERP_SUCCESS)
    {
        for (size_t q = 0; q < Q; ++q)
        {
            unsigned qi[8];
            size_t rem = q;
            for (unsigned k = ndim; k-- > 0;)
            {
                qi[k] = (unsigned)(rem % m[k]);
                rem /= m[k];
            }
            for (size_t i = 0; i < N; ++i)
            {
                unsigned bi[8];
                size_t r2 = i;
                for (unsigned k = ndim; k-- > 0;)
                {
                    bi[k] = (unsigned)(r2 % n[k]);
                    r2 /= n[k];
                }
                double v = 1.0;
                for (unsigned k = 0; k < ndim; ++k)
                    v *= tab[k][(size_t)bi[k] * m[k] + qi[k]];
                out[q * N + i] = v;
            }
        }
    }
    for (unsigned k = 0; k < ndim; ++k)
        free(tab[k]);
    return res;
}

/* ---- A13: basis at arbitrary points (src/basis/basis_set.c:462-556) -------------------------- */

int ib200_tables_basis_at_points(basis_spec_t basis, size_t cnt, const double *x, double *values, double *derivs)
{
    const unsigned order = basis.order;
    const unsigned n = order + 1;
    switch (basis.type)
    {
    case BASIS_LEGENDRE:
        for (size_t p = 0; p < cnt; ++p)
        {
            if (values)
                legendre_all(order, x[p], values + p * n, 1);
            if (derivs)
            {
                /* same recurrence as the table: d_i = i P_{i-1} + x d_{i-1} (:494-513) */
                double *row = derivs + p * n;
                legendre_all(order, x[p], row, 1);
                double prev_val = row[0];
                row[0] = 0;
                for (unsigned i = 0; i < order; ++i)
                {
                    const double keep = row[i + 1];
                    row[i + 1] = (double)(i + 1) * prev_val + x[p] * row[i];
                    prev_val = keep;
                }
            }
        }
        return INTERP_SUCCESS;
    case BASIS_BERNSTEIN:
        for (size_t p = 0; p < cnt; ++p)
        {
            const double t = (x[p] + 1) / 2;
            if (values)
                bernstein_vector(t, order, values + p * n);
            if (derivs)
            {
                double *row = derivs + p * n;
                if (order == 0)
                {
                    row[0] = 0.0; /* the reference underflows order-1 here; a constant has no slope */
                    continue;
                }
                bernstein_vector(t, order - 1, row);
                row[order] = (double)order * row[order - 1];
                for (unsigned j = order - 1; j > 0; --j)
                    row[j] = (double)order * (row[j - 1] - row[j]);
                row[0] *= -(double)order;
                for (unsigned j = 0; j < n; ++j)
                    row[j] /= 2.0;
            }
        }
        return INTERP_SUCCESS;
    case BASIS_LAGRANGE_GAUSS:
    case BASIS_LAGRANGE_UNIFORM:
    case BASIS_LAGRANGE_GAUSS_LOBATTO:
    case BASIS_LAGRANGE_CHEBYSHEV_GAUSS: {
        double *buf = malloc(sizeof(double) * 4 * n);
        if (!buf)
            return INTERP_ERROR_FAILED_ALLOCATION;
        double *roots = buf, *den = buf + n, *dif = buf + 2 * n, *num = buf + 3 * n;
        const int res = ib200_tables_lagrange_roots(basis, roots);
        if (res != INTERP_SUCCESS)
        {
            free(buf);
            return res;
        }
        for (unsigned j = 0; j < n; ++j)
            den[j] = lagrange_denominator(n, roots, j);
        for (size_t p = 0; p < cnt; ++p)
        {
            if (values)
            {
                /* lagrange_polynomial_values_2, src/polynomials/lagrange.c:108-183: every point
                 * divides last except the final one, which divides first */
                const int early = (p + 1 == cnt);
                for (unsigned j = 0; j < n; ++j)
                    values[p * n + j] = (n == 1) ? 1.0 : lagrange_value(n, roots, j, x[p], den[j], early);
            }
            if (derivs)
            {
                /* lagrange_polynomial_first_derivative_2, :383-525 */
                lagrange_derivative_numerators(n, roots, x[p], dif, num);
                const int chained = (p + 2 >= cnt);
                for (unsigned j = 0; j < n; ++j)
                    derivs[p * n + j] = chained ? lagrange_divide_chain(n, roots, j, num[j]) : num[j] / den[j];
            }
        }
        free(buf);
        return INTERP_SUCCESS;
    }
    default:
        return INTERP_ERROR_INVALID_ENUM;
    }
}

/* basis_compute_outer_product_basis, src/basis/basis_set.c:577-633: running product over the
 * dimensions in ascending order, out[p][i] = ((1 * b0[i0]) * b1[i1]) * ... */
int ib200_tables_function_space_evaluate(unsigned ndim, const basis_spec_t *basis, size_t cnt,
                                         const double *const *x, double *out)
{
    if (ndim == 0 || ndim > 8)
        return IB200_ERROR_BAD_ARGUMENT;
    unsigned n[8];
    size_t N = 1;
    double *tab[8] = {0};
    int res = INTERP_SUCCESS;
    for (unsigned k = 0; k < ndim && res == INTERP_SUCCESS; ++k)
    {
        n[k] = basis[k].order + 1;
        N *= n[k];
        tab[k] = malloc(sizeof(double) * (cnt ? cnt : 1) * n[k]);
        res = tab[k] ? ib200_tables_basis_at_points(basis[k], cnt, x[k], tab[k], NULL) : INTERP_ERROR_FAILED_ALLOCATION;
    }
    if (res == INTERP_SUCCESS)
    {
        for (size_t p = 0; p < cnt; ++p)
            for (size_t i = 0; i < N; ++i)
            {
                unsigned bi[8];
                size_t rem = i;
                for (unsigned k = ndim; k-- > 0;)
                {
                    bi[k] = (unsigned)(rem % n[k]);
                    rem /= n[k];
                }
                double v = 1.0;
                for (unsigned k = 0; k < ndim; ++k)
                    v = v * tab[k][p * n[k] + bi[k]];
                out[p * N + i] = v;
            }
    }
    for (unsigned k = 0; k < ndim; ++k)
        free(tab[k]);
    return res;
}

/* N3 -- incidence_matrix(basis_specs), /root/reference/src/python/incidence.c:38-110: the [n][n+1] matrix (n = order)
 * that maps the DoFs of a 1-D function to the DoFs of its derivative in the space one order lower.  Bernstein and
 * Legendre: closed-form integer patterns; Lagrange: l_col'(node_row of the order n-1 basis) by
 * lagrange_polynomial_first_derivative_2 (the last two rows divide pair by pair, as in basis_at_points above). */
int ib200_tables_incidence_matrix(basis_spec_t basis, double *out)
{
    const unsigned n = basis.order;
    if (n == 0)
        return IB200_ERROR_BAD_ARGUMENT;
    memset(out, 0, sizeof(double) * n * (n + 1));
    switch (basis.type)
    {
    case BASIS_BERNSTEIN: {
        const double coeff = (double)n / 2.0;
        out[0] = -coeff;
        for (unsigned col = 1; col < n; ++col)
        {
            out[col * (n + 1) + col] = -coeff;
            out[(col - 1) * (n + 1) + col] = +coeff;
        }
        out[(n - 1) * (n + 1) + n] = +coeff;
        return INTERP_SUCCESS;
    }
    case BASIS_LEGENDRE:
        /* d/dx P_col = sum over r = col-1, col-3, .. of (2r+1) P_r */
        for (unsigned col = 1; col <= n; ++col)
            for (unsigned r = col - 1;; r -= 2)
            {
                out[r * (n + 1) + col] = (double)(2 * r + 1);
                if (r < 2)
                    break;
            }
        return INTERP_SUCCESS;
    case BASIS_LAGRANGE_GAUSS:
    case BASIS_LAGRANGE_UNIFORM:
    case BASIS_LAGRANGE_GAUSS_LOBATTO:
    case BASIS_LAGRANGE_CHEBYSHEV_GAUSS: {
        double *nodes = malloc(sizeof(double) * n);
        if (!nodes)
            return INTERP_ERROR_FAILED_ALLOCATION;
        basis_spec_t lower = basis;
        lower.order = n - 1;
        int res = ib200_tables_lagrange_roots(lower, nodes);
        if (res == INTERP_SUCCESS)
            res = ib200_tables_basis_at_points(basis, n, nodes, NULL, out);
        free(nodes);
        return res;
    }
    default:
        return INTERP_ERROR_INVALID_ENUM;
    }
}


/* ---- src/polynomials on arbitrary roots (what lagrange1d / dlagrange1d / d2lagrange1d of the extension expose,
 *      src/python/module.c:102-560) ------------------------------------------------------------------------------
 * out[p][j] = d^k/dx^k l_j(x_p) for k = deriv in {0, 1, 2}.  The reference multiplies the numerator sums by the
 * RECIPROCAL of the product denominator here (lagrange_polynomial_values / _first_derivative / _second_derivative,
 * src/polynomials/lagrange.c:59-106, 317-381, 739-808), unlike the table routines above, which divide. */
int ib200_tables_lagrange_samples(unsigned n, const double *roots, size_t cnt, const double *x, unsigned deriv, double *out)
{
    if (deriv > 2)
        return INTERP_ERROR_BAD_SYSTEM;
    if (n == 0 || cnt == 0)
        return INTERP_SUCCESS;
    double *buf = malloc(sizeof(double) * 2 * n);
    if (!buf)
        return INTERP_ERROR_FAILED_ALLOCATION;
    double *rden = buf, *dif = buf + n;
    for (unsigned j = 0; j < n; ++j)
        rden[j] = 1.0 / lagrange_denominator(n, roots, j);
    for (size_t p = 0; p < cnt; ++p)
    {
        double *row = out + p * n;
        if (deriv == 0)
        {
            for (unsigned j = 0; j < n; ++j)
                row[j] = lagrange_value(n, roots, j, x[p], 1.0, 0) * rden[j];
            continue;
        }
        if (deriv == 1)
            lagrange_derivative_numerators(n, roots, x[p], dif, row);
        else
        {
            /* every unordered triple {i > j > k} contributes twice the product of the other differences to each of
             * its three members, in the order k, j, i */
            for (unsigned j = 0; j < n; ++j)
            {
                dif[j] = x[p] - roots[j];
                row[j] = 0.0;
            }
            for (unsigned i = 0; i < n; ++i)
                for (unsigned j = 0; j < i; ++j)
                    for (unsigned k = 0; k < j; ++k)
                    {
                        double prod = 1.0;
                        for (unsigned l = 0; l < n; ++l)
                            if (l != i && l != j && l != k)
                                prod *= dif[l];
                        row[k] += 2 * prod;
                        row[j] += 2 * prod;
                        row[i] += 2 * prod;
                    }
        }
        for (unsigned j = 0; j < n; ++j)
            row[j] *= rden[j];
    }
    free(buf);
    return INTERP_SUCCESS;
}

/* compute_gll (src/python/module.c:559-621 -> gauss_lobatto_nodes_weights, src/integration/gauss_lobatto.c:14-61) with
 * the caller's iteration limit and tolerance; *non_converged = nodes whose last Newton step was still above tol. */
int ib200_tables_gauss_lobatto(unsigned order, unsigned max_iter, double tol, double *nodes, double *weights,
                               int *non_converged)
{
    const int bad = gauss_lobatto(order + 1, tol, max_iter, nodes, weights);
    if (non_converged)
        *non_converged = bad;
    return INTERP_SUCCESS;
}

/* bernstein1d (src/python/module.c:646-686): out[p][0..order] = Bernstein polynomials of degree `order` at t = x[p]
 * in [0, 1] (NOT mapped from [-1, 1]: the extension function takes the unit-interval coordinate) */
int ib200_tables_bernstein_samples(unsigned order, size_t cnt, const double *t, double *out)
{
    for (size_t p = 0; p < cnt; ++p)
        bernstein_vector(t[p], order, out + p * (order + 1));
    return INTERP_SUCCESS;
}

/* bernstein_coefficients (src/python/module.c:706-718 -> bernstein_from_power_series, src/polynomials/bernstein.c:8-29):
 * power-series coefficients a_k of a degree m = n-1 polynomial -> Bernstein coefficients, in place, by forward
 * elimination: B^m_k(t) = C(m,k) sum_i (-1)^i C(m-k,i) t^(k+i), so once c_k = r_k / C(m,k) is known (r = running
 * residual) its higher powers are removed from the residual: r_(k+i) += r_k * (-1)^(i+1) C(m-k,i). */
int ib200_tables_bernstein_from_power(unsigned n, double *c)
{
    const int m = (int)n - 1;
    unsigned binom = 1; /* C(m, k) */
    for (int k = 0; k <= m; ++k)
    {
        const double resid = c[k];
        const int rest = m - k;
        int alt = rest; /* (-1)^(i+1) C(rest, i), i = 1 */
        for (int i = 1; i <= rest; ++i)
        {
            c[k + i] += resid * (double)alt;
            alt = (alt * (i - rest)) / (i + 1);
        }
        c[k] = resid / (double)binom;
        binom = (binom * (unsigned)rest) / (unsigned)(k + 1);
    }
    return INTERP_SUCCESS;
}
