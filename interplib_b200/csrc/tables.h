/* Internal interface of the host table layer (tables.c); the public wrappers are in capi.cu. */
#ifndef IB200_TABLES_H
#define IB200_TABLES_H
#include "../../include/interp_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

int ib200_tables_rule(integration_spec_t spec, double *nodes, double *weights, unsigned *accuracy);
int ib200_tables_order_for_accuracy(integration_rule_type_t type, unsigned accuracy, unsigned *order);
int ib200_tables_is_lagrange(basis_set_type_t type);
int ib200_tables_lagrange_roots(basis_spec_t spec, double *roots);
int ib200_tables_lagrange_roots_denominators(basis_spec_t spec, double *roots, double *den);
int ib200_tables_basis(basis_spec_t basis, integration_spec_t rule, double *values, double *derivs);
int ib200_tables_space_weights(unsigned ndim, const integration_spec_t *rules, double *weights);
int ib200_tables_space_nodes(unsigned ndim, const integration_spec_t *rules, double *nodes);
int ib200_tables_values_at_nodes(unsigned ndim, const basis_spec_t *basis, const integration_spec_t *rules,
                                 double *out);
int ib200_tables_basis_at_points(basis_spec_t basis, size_t cnt, const double *x, double *values, double *derivs);
int ib200_tables_function_space_evaluate(unsigned ndim, const basis_spec_t *basis, size_t cnt,
                                         const double *const *x, double *out);

int ib200_tables_incidence_matrix(basis_spec_t basis, double *out);
int ib200_tables_lagrange_samples(unsigned n, const double *roots, size_t cnt, const double *x, unsigned deriv, double *out);
int ib200_tables_gauss_lobatto(unsigned order, unsigned max_iter, double tol, double *nodes, double *weights,
                               int *non_converged);
int ib200_tables_bernstein_samples(unsigned order, size_t cnt, const double *t, double *out);
int ib200_tables_bernstein_from_power(unsigned n, double *c);
/* fastdiag.c */
int ib200_tables_fastdiag(basis_spec_t basis, integration_spec_t rule, double *V, double *lambda, double *Vw);
int ib200_tables_linear_basis(basis_spec_t basis, double a[2], double s[2]);
int ib200_tables_projector_1d(basis_spec_t basis, integration_spec_t rule, double *P);
int ib200_tables_pair_moments_deriv(basis_spec_t bout, basis_spec_t bin, integration_spec_t rule, double *A, double *X);
int ib200_tables_pair_moments_k(basis_spec_t bout, basis_spec_t bin, integration_spec_t rule, unsigned kmax, double *out);
int ib200_tables_pair_moments(basis_spec_t bout, basis_spec_t bin, integration_spec_t rule, double *A, double *X);

#ifdef __cplusplus
}
#endif
#endif
