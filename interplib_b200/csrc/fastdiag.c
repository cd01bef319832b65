/* Host tables of the fast-diagonalisation L2 projection on bilinear quads (and straight 1-D segments).
 *
 * What the reference does per element (python/interplib/integration.py:171-223): assemble M (mass_matrices.c:161-339),
 * assemble b (integration.py:106-168) and call np.linalg.solve(M, b).  On an element whose coordinate map is multilinear
 * of order 1 the Jacobian determinant is a LINEAR function of the reference coordinates,
 *
 *      det J(xi0, xi1) = alpha + beta xi0 + gamma xi1        (the xi0*xi1 terms of the 2x2 determinant cancel),
 *
 * so the element mass matrix of a tensor-product space is a sum of three Kronecker products of 1-D matrices that do not
 * depend on the element:
 *
 *      M = alpha A0 (x) A1 + beta X0 (x) A1 + gamma A0 (x) X1,
 *      A_k[j][i] = sum_q w_q B_k[j][q] B_k[i][q],        X_k[j][i] = sum_q w_q x_q B_k[j][q] B_k[i][q].
 *
 * With the generalised eigen-decomposition of the 1-D pencil, V_k^T A_k V_k = I and V_k^T X_k V_k = diag(lambda_k),
 *
 *      M^-1 = (V0 (x) V1) diag(1 / (alpha + beta lambda0_i + gamma lambda1_j)) (V0 (x) V1)^T,
 *
 * i.e. the primal DoFs are two sum-factorised transforms and N divisions per element instead of an O(N^3) factorisation
 * (and M never has to exist).  This file builds V, lambda and the transformed test functions once per (basis, rule); all
 * arithmetic is in long double so that the tables carry no error of their own at double precision.
 */
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "tables.h"

typedef long double real;

/* Cyclic Jacobi for a symmetric n x n matrix C (row-major, destroyed): C = Q diag(ev) Q^T, Q row-major [n][n]. */
static void jacobi_eigen(unsigned n, real *C, real *Q, real *ev)
{
    for (unsigned i = 0; i < n; ++i)
        for (unsigned j = 0; j < n; ++j)
            Q[i * n + j] = i == j ? 1.0L : 0.0L;
    for (unsigned sweep = 0; sweep < 64; ++sweep)
    {
        real off = 0.0L, diag = 0.0L;
        for (unsigned i = 0; i < n; ++i)
        {
            diag += C[i * n + i] * C[i * n + i];
            for (unsigned j = i + 1; j < n; ++j)
                off += C[i * n + j] * C[i * n + j];
        }
        if (off <= 1e-42L * (diag + 1e-300L))
            break;
        for (unsigned p = 0; p + 1 < n; ++p)
            for (unsigned q = p + 1; q < n; ++q)
            {
                const real apq = C[p * n + q];
                if (apq == 0.0L)
                    continue;
                const real theta = (C[q * n + q] - C[p * n + p]) / (2.0L * apq);
                const real t = (theta >= 0.0L ? 1.0L : -1.0L) / (fabsl(theta) + sqrtl(theta * theta + 1.0L));
                const real c = 1.0L / sqrtl(t * t + 1.0L), s = t * c;
                for (unsigned k = 0; k < n; ++k)
                {
                    const real akp = C[k * n + p], akq = C[k * n + q];
                    C[k * n + p] = c * akp - s * akq;
                    C[k * n + q] = s * akp + c * akq;
                }
                for (unsigned k = 0; k < n; ++k)
                {
                    const real apk = C[p * n + k], aqk = C[q * n + k];
                    C[p * n + k] = c * apk - s * aqk;
                    C[q * n + k] = s * apk + c * aqk;
                }
                for (unsigned k = 0; k < n; ++k)
                {
                    const real qkp = Q[k * n + p], qkq = Q[k * n + q];
                    Q[k * n + p] = c * qkp - s * qkq;
                    Q[k * n + q] = s * qkp + c * qkq;
                }
            }
    }
    for (unsigned i = 0; i < n; ++i)
        ev[i] = C[i * n + i];
}

/* V[j][i] (basis function j, mode i), lambda[i], Vw[i][q] = w_q sum_j V[j][i] B[j][q].
 * INTERP_ERROR_BAD_SYSTEM when the discrete 1-D mass matrix is not positive definite (fewer rule points than basis
 * functions): the caller then keeps the factorisation path, whose systems are singular in the reference too. */
int ib200_tables_fastdiag(basis_spec_t basis, integration_spec_t rule, double *V, double *lambda, double *Vw)
{
    const unsigned n = basis.order + 1, m = rule.order + 1;
    double *B = (double *)malloc(sizeof(double) * (2 * (size_t)n * m + 2 * m));
    real *work = (real *)malloc(sizeof(real) * (5 * (size_t)n * n + n));
    if (!B || !work)
    {
        free(B);
        free(work);
        return INTERP_ERROR_FAILED_ALLOCATION;
    }
    double *D = B + (size_t)n * m, *x = D + (size_t)n * m, *w = x + m;
    real *A = work, *X = A + n * n, *L = X + n * n, *C = L + n * n, *Q = C + n * n, *ev = Q + n * n;
    int res = ib200_tables_rule(rule, x, w, NULL);
    if (res == INTERP_SUCCESS)
        res = ib200_tables_basis(basis, rule, B, D);
    /* fewer nodes than basis functions: the quadrature mass matrix B W B^T has rank <= m < n whatever the roundoff says */
    if (res == INTERP_SUCCESS && m < n)
        res = INTERP_ERROR_BAD_SYSTEM;
    if (res != INTERP_SUCCESS)
        goto done;
    for (unsigned j = 0; j < n; ++j)
        for (unsigned i = 0; i < n; ++i)
        {
            real a = 0.0L, xx = 0.0L;
            for (unsigned q = 0; q < m; ++q)
            {
                const real p = (real)w[q] * ((real)B[j * m + q] * (real)B[i * m + q]);
                a += p;
                xx += p * (real)x[q];
            }
            A[j * n + i] = a;
            X[j * n + i] = xx;
        }
    /* A = L L^T */
    memset(L, 0, sizeof(real) * n * n);
    for (unsigned j = 0; j < n; ++j)
    {
        real d = A[j * n + j];
        for (unsigned k = 0; k < j; ++k)
            d -= L[j * n + k] * L[j * n + k];
        /* a pivot at roundoff level (~ LDBL_EPSILON * A_jj, either sign) means a numerically singular matrix */
        if (!(A[j * n + j] > 0.0L) || !(d > (real)n * 1e3L * LDBL_EPSILON * A[j * n + j]))
        {
            res = INTERP_ERROR_BAD_SYSTEM;
            goto done;
        }
        L[j * n + j] = sqrtl(d);
        for (unsigned i = j + 1; i < n; ++i)
        {
            real s = A[i * n + j];
            for (unsigned k = 0; k < j; ++k)
                s -= L[i * n + k] * L[j * n + k];
            L[i * n + j] = s / L[j * n + j];
        }
    }
    /* C = L^-1 X L^-T: forward substitution on the columns, then on the rows */
    for (unsigned c = 0; c < n; ++c) /* Y = L^-1 X, stored in C */
        for (unsigned i = 0; i < n; ++i)
        {
            real s = X[i * n + c];
            for (unsigned k = 0; k < i; ++k)
                s -= L[i * n + k] * C[k * n + c];
            C[i * n + c] = s / L[i * n + i];
        }
    for (unsigned r = 0; r < n; ++r) /* C = Y L^-T: row r solves z L^T = y */
        for (unsigned i = 0; i < n; ++i)
        {
            real s = C[r * n + i];
            for (unsigned k = 0; k < i; ++k)
                s -= C[r * n + k] * L[i * n + k];
            C[r * n + i] = s / L[i * n + i];
        }
    for (unsigned i = 0; i < n; ++i) /* symmetrise the rounding noise */
        for (unsigned j = i + 1; j < n; ++j)
            C[i * n + j] = C[j * n + i] = 0.5L * (C[i * n + j] + C[j * n + i]);
    jacobi_eigen(n, C, Q, ev);
    /* V = L^-T Q: back substitution per column (X reused as V in long double) */
    for (unsigned c = 0; c < n; ++c)
        for (unsigned ii = n; ii-- > 0;)
        {
            real s = Q[ii * n + c];
            for (unsigned k = ii + 1; k < n; ++k)
                s -= L[k * n + ii] * X[k * n + c];
            X[ii * n + c] = s / L[ii * n + ii];
        }
    for (unsigned i = 0; i < n; ++i)
        lambda[i] = (double)ev[i];
    for (unsigned j = 0; j < n; ++j)
        for (unsigned i = 0; i < n; ++i)
            V[j * n + i] = (double)X[j * n + i];
    for (unsigned i = 0; i < n; ++i)
        for (unsigned q = 0; q < m; ++q)
        {
            real s = 0.0L;
            for (unsigned j = 0; j < n; ++j)
                s += X[j * n + i] * (real)B[j * m + q];
            Vw[i * m + q] = (double)(s * (real)w[q]);
        }
done:
    free(B);
    free(work);
    return res;
}

/* The two functions of an order-1 basis as phi_i(xi) = a[i] + s[i] xi (every order-1 family of the reference is linear:
 * Lagrange on two nodes, Bernstein (1 -+ xi)/2, Legendre 1 and xi). */
int ib200_tables_linear_basis(basis_spec_t basis, double a[2], double s[2])
{
    if (basis.order != 1)
        return IB200_ERROR_UNSUPPORTED;
    const double x[2] = {0.0, 1.0};
    double v[4], d[4]; /* basis_at_points layout: [cnt][n] */
    const int res = ib200_tables_basis_at_points(basis, 2, x, v, d);
    if (res != INTERP_SUCCESS)
        return res;
    for (int i = 0; i < 2; ++i)
    {
        a[i] = v[0 * 2 + i];
        s[i] = v[1 * 2 + i] - v[0 * 2 + i];
    }
    return INTERP_SUCCESS;
}

/* Zeroth and first moments of the products of two 1-D bases under a rule:
 *   A[j][i] = sum_q w_q Bout_j(x_q) Bin_i(x_q),   X[j][i] = sum_q w_q x_q Bout_j(x_q) Bin_i(x_q)
 * accumulated in long double.  The product of the two basis values is formed first, so for equal specs both tables are
 * exactly symmetric.  They are all an element with a (multi)linear Jacobian determinant needs of the rule: its mass matrix
 * is alpha A0(x)A1 + beta X0(x)A1 + gamma A0(x)X1 (kernels_mass2d_kron.cu). */
int ib200_tables_pair_moments(basis_spec_t bout, basis_spec_t bin, integration_spec_t rule, double *A, double *X)
{
    const unsigned no = bout.order + 1, ni = bin.order + 1, m = rule.order + 1;
    double *buf = (double *)malloc(sizeof(double) * (2 * (size_t)(no + ni) * m + 2 * m));
    if (!buf)
        return INTERP_ERROR_FAILED_ALLOCATION;
    double *Bo = buf, *Do = Bo + (size_t)no * m, *Bi = Do + (size_t)no * m, *Di = Bi + (size_t)ni * m, *x = Di + (size_t)ni * m,
           *w = x + m;
    int res = ib200_tables_rule(rule, x, w, NULL);
    if (res == INTERP_SUCCESS)
        res = ib200_tables_basis(bout, rule, Bo, Do);
    if (res == INTERP_SUCCESS)
        res = ib200_tables_basis(bin, rule, Bi, Di);
    if (res == INTERP_SUCCESS)
        for (unsigned j = 0; j < no; ++j)
            for (unsigned i = 0; i < ni; ++i)
            {
                real a = 0.0L, xx = 0.0L;
                for (unsigned q = 0; q < m; ++q)
                {
                    const real p = (real)w[q] * ((real)Bo[j * m + q] * (real)Bi[i * m + q]);
                    a += p;
                    xx += p * (real)x[q];
                }
                A[j * ni + i] = (double)a;
                X[j * ni + i] = (double)xx;
            }
    free(buf);
    return res;
}

/* P[i][q] = sum_j (A^-1)[i][j] w_q B_j(x_q): the L2 projector of the 1-D space on an element with CONSTANT Jacobian
 * determinant (straight segments: M_e = det_e A, b_e = det_e Bw f_e, so u_e = P f_e whatever det_e is).  Long double
 * Cholesky; INTERP_ERROR_BAD_SYSTEM when A is not positive definite. */
int ib200_tables_projector_1d(basis_spec_t basis, integration_spec_t rule, double *P)
{
    const unsigned n = basis.order + 1, m = rule.order + 1;
    double *B = (double *)malloc(sizeof(double) * (2 * (size_t)n * m + 2 * m));
    real *work = (real *)malloc(sizeof(real) * (2 * (size_t)n * n + n));
    if (!B || !work)
    {
        free(B);
        free(work);
        return INTERP_ERROR_FAILED_ALLOCATION;
    }
    double *D = B + (size_t)n * m, *x = D + (size_t)n * m, *w = x + m;
    real *A = work, *L = A + n * n, *y = L + n * n;
    int res = ib200_tables_rule(rule, x, w, NULL);
    if (res == INTERP_SUCCESS)
        res = ib200_tables_basis(basis, rule, B, D);
    /* fewer nodes than basis functions: the quadrature mass matrix B W B^T has rank <= m < n whatever the roundoff says */
    if (res == INTERP_SUCCESS && m < n)
        res = INTERP_ERROR_BAD_SYSTEM;
    if (res != INTERP_SUCCESS)
        goto done;
    for (unsigned j = 0; j < n; ++j)
        for (unsigned i = 0; i < n; ++i)
        {
            real a = 0.0L;
            for (unsigned q = 0; q < m; ++q)
                a += (real)w[q] * ((real)B[j * m + q] * (real)B[i * m + q]);
            A[j * n + i] = a;
        }
    memset(L, 0, sizeof(real) * n * n);
    for (unsigned j = 0; j < n; ++j)
    {
        real d = A[j * n + j];
        for (unsigned k = 0; k < j; ++k)
            d -= L[j * n + k] * L[j * n + k];
        /* a pivot at roundoff level (~ LDBL_EPSILON * A_jj, either sign) means a numerically singular matrix */
        if (!(A[j * n + j] > 0.0L) || !(d > (real)n * 1e3L * LDBL_EPSILON * A[j * n + j]))
        {
            res = INTERP_ERROR_BAD_SYSTEM;
            goto done;
        }
        L[j * n + j] = sqrtl(d);
        for (unsigned i = j + 1; i < n; ++i)
        {
            real s = A[i * n + j];
            for (unsigned k = 0; k < j; ++k)
                s -= L[i * n + k] * L[j * n + k];
            L[i * n + j] = s / L[j * n + j];
        }
    }
    for (unsigned q = 0; q < m; ++q) /* column q of P solves A p = w_q B[:, q] */
    {
        for (unsigned i = 0; i < n; ++i)
        {
            real s = (real)w[q] * (real)B[i * m + q];
            for (unsigned k = 0; k < i; ++k)
                s -= L[i * n + k] * y[k];
            y[i] = s / L[i * n + i];
        }
        for (unsigned ii = n; ii-- > 0;)
        {
            real s = y[ii];
            for (unsigned k = ii + 1; k < n; ++k)
                s -= L[k * n + ii] * y[k];
            y[ii] = s / L[ii * n + ii];
        }
        for (unsigned i = 0; i < n; ++i)
            P[i * m + q] = (double)y[i];
    }
done:
    free(B);
    free(work);
    return res;
}

/* Moments 0..kmax of the products of two 1-D bases: out[k][j][i] = sum_q w_q x_q^k Bout_j(x_q) Bin_i(x_q), long double.
 * The Jacobian determinant of a trilinear hexahedron is a polynomial of degree 2 in each reference coordinate, so its mass
 * matrix is a 27-term Kronecker sum of these tables (kernels_mass3d_kron.cu). */
int ib200_tables_pair_moments_k(basis_spec_t bout, basis_spec_t bin, integration_spec_t rule, unsigned kmax, double *out)
{
    const unsigned no = bout.order + 1, ni = bin.order + 1, m = rule.order + 1;
    double *buf = (double *)malloc(sizeof(double) * (2 * (size_t)(no + ni) * m + 2 * m));
    if (!buf)
        return INTERP_ERROR_FAILED_ALLOCATION;
    double *Bo = buf, *Do = Bo + (size_t)no * m, *Bi = Do + (size_t)no * m, *Di = Bi + (size_t)ni * m, *x = Di + (size_t)ni * m,
           *w = x + m;
    int res = ib200_tables_rule(rule, x, w, NULL);
    if (res == INTERP_SUCCESS)
        res = ib200_tables_basis(bout, rule, Bo, Do);
    if (res == INTERP_SUCCESS)
        res = ib200_tables_basis(bin, rule, Bi, Di);
    if (res == INTERP_SUCCESS)
        for (unsigned j = 0; j < no; ++j)
            for (unsigned i = 0; i < ni; ++i)
                for (unsigned k = 0; k <= kmax; ++k)
                {
                    real acc = 0.0L;
                    for (unsigned q = 0; q < m; ++q)
                    {
                        real p = (real)w[q] * ((real)Bo[j * m + q] * (real)Bi[i * m + q]);
                        for (unsigned r = 0; r < k; ++r)
                            p *= (real)x[q];
                        acc += p;
                    }
                    out[((size_t)k * no + j) * ni + i] = (double)acc;
                }
    free(buf);
    return res;
}

/* As ib200_tables_pair_moments with the DERIVATIVE of the input basis: A[j][i] = sum_q w_q Bout_j(x_q) Bin_i'(x_q),
 * X[j][i] = sum_q w_q x_q Bout_j(x_q) Bin_i'(x_q) -- the 1-D factors of compute_gradient_mass_matrix along its
 * differentiated dimension (/root/reference/src/python/mass_matrices.c:341-565). */
int ib200_tables_pair_moments_deriv(basis_spec_t bout, basis_spec_t bin, integration_spec_t rule, double *A, double *X)
{
    const unsigned no = bout.order + 1, ni = bin.order + 1, m = rule.order + 1;
    double *buf = (double *)malloc(sizeof(double) * (2 * (size_t)(no + ni) * m + 2 * m));
    if (!buf)
        return INTERP_ERROR_FAILED_ALLOCATION;
    double *Bo = buf, *Do = Bo + (size_t)no * m, *Bi = Do + (size_t)no * m, *Di = Bi + (size_t)ni * m, *x = Di + (size_t)ni * m,
           *w = x + m;
    int res = ib200_tables_rule(rule, x, w, NULL);
    if (res == INTERP_SUCCESS)
        res = ib200_tables_basis(bout, rule, Bo, Do);
    if (res == INTERP_SUCCESS)
        res = ib200_tables_basis(bin, rule, Bi, Di);
    if (res == INTERP_SUCCESS)
        for (unsigned j = 0; j < no; ++j)
            for (unsigned i = 0; i < ni; ++i)
            {
                real a = 0.0L, xx = 0.0L;
                for (unsigned q = 0; q < m; ++q)
                {
                    const real pq = (real)w[q] * ((real)Bo[j * m + q] * (real)Di[i * m + q]);
                    a += pq;
                    xx += pq * (real)x[q];
                }
                A[j * ni + i] = (double)a;
                X[j * ni + i] = (double)xx;
            }
    free(buf);
    return res;
}
