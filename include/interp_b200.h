/* interp_b200.h -- flat C ABI of the B200-native interplib hot path.
 *
 * One shared library (libinterp_b200.so, hand-written CUDA for sm_100a) sits exactly where the
 * reference's CPython extension (the .c files of `src/python/`, L4 in SURVEY.md section 1) calls down into
 * its table layer and runs its per-element loops.  Every entry point cites the reference
 * interface it replaces (paths are into /root/reference).  Conventions:
 *
 *   - plain pointers and sizes only; all floating point is IEEE double; indices are unsigned /
 *     size_t; the element index is always the slowest dimension, and inside one element the
 *     reference's row-major layouts are kept verbatim (last index fastest,
 *     `multidim_iterator_t` contract, e.g. src/python/mass_matrices.c:264-323);
 *   - every function returns an `interp_result_t` value (same numbering as
 *     src/common/error.h:9-25, extended with IB200_* codes); nothing aborts, no C++ exception
 *     crosses the boundary; `ib200_last_error()` returns the text of the last failure;
 *   - the caller owns every host buffer; the library owns device scratch and the device-side
 *     table cache inside the context;
 *   - batched calls take `where`: IB200_HOST  = pointers are host memory, the call stages
 *     chunks through pinned buffers (H2D -> kernels -> D2H) and returns when `out` is
 *     complete; IB200_DEVICE = pointers are device memory on the context's device, work is
 *     queued on the context stream and the call returns immediately.
 *   - there is NO CPU fallback for the batched hot path: without a CUDA device the context
 *     cannot be created and the batched calls fail with IB200_ERROR_NO_DEVICE.
 */
#ifndef INTERP_B200_H
#define INTERP_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define IB200_API __attribute__((visibility("default")))
#else
#define IB200_API
#endif

/* ---- A14: POD descriptors, numbering identical to the reference ------------------------ */

/* src/common/error.h:9-25 */
typedef enum
{
    INTERP_SUCCESS = 0,
    INTERP_ERROR_NOT_IN_DOMAIN,
    INTERP_ERROR_NOT_INCREASING,
    INTERP_ERROR_FAILED_ALLOCATION,
    INTERP_ERROR_BAD_SYSTEM,
    INTERP_ERROR_INVALID_ENUM,
    INTERP_ERROR_NOT_IN_REGISTRY,
    INTERP_ERROR_GEOID_OUT_OF_RANGE,
    INTERP_ERROR_GEOID_NOT_VALID,
    INTERP_ERROR_SURFACE_NOT_CLOSED,
    INTERP_ERROR_OBJECT_CONNECTED_TWICE,
    INTERP_ERROR_MATRIX_DIMS_MISMATCH,
    INTERP_ERROR_COUNT,
    /* extensions of this library */
    IB200_ERROR_CUDA = 100,        /* a CUDA runtime call or kernel failed               */
    IB200_ERROR_NO_DEVICE = 101,   /* no usable sm_100 device / context not created      */
    IB200_ERROR_BAD_ARGUMENT = 102,/* NULL pointer, dimension or index out of range      */
    IB200_ERROR_UNSUPPORTED = 103, /* valid request outside the implemented envelope     */
    IB200_ERROR_NCCL = 104,        /* an NCCL call failed or libnccl could not be loaded */
} interp_result_t;

/* src/basis/basis_set.h:11-20 */
typedef enum
{
    BASIS_INVALID = 0,
    BASIS_LEGENDRE,
    BASIS_LAGRANGE_GAUSS_LOBATTO,
    BASIS_LAGRANGE_GAUSS,
    BASIS_LAGRANGE_UNIFORM,
    BASIS_LAGRANGE_CHEBYSHEV_GAUSS,
    BASIS_BERNSTEIN,
} basis_set_type_t;

/* src/integration/integration_rules.h:7-12 */
typedef enum
{
    INTEGRATION_RULE_TYPE_NONE = 0,
    INTEGRATION_RULE_TYPE_GAUSS_LEGENDRE,
    INTEGRATION_RULE_TYPE_GAUSS_LOBATTO,
} integration_rule_type_t;

/* src/basis/basis_set.h:22-26 */
typedef struct
{
    basis_set_type_t type;
    unsigned order;
} basis_spec_t;

/* src/integration/integration_rules.h:17-21 */
typedef struct
{
    integration_rule_type_t type;
    unsigned order;
} integration_spec_t;

enum
{
    IB200_MAX_DIM = 6,   /* reference dimensions supported by the batched kernels          */
    IB200_MAX_PHYS = 6,  /* physical coordinates per point                                 */
    IB200_HOST = 0,
    IB200_DEVICE = 1,
};

/* src/common/error.c:9-38 (interp_error_str / interp_error_msg) + the IB200_* codes */
IB200_API const char *ib200_error_str(int code);
IB200_API const char *ib200_error_msg(int code);
/* text of the last failure on the calling thread ("" if none) */
IB200_API const char *ib200_last_error(void);
IB200_API const char *ib200_version(void);

/* ---- A1-A4, A10, A11: host-side 1-D tables (bit-exact restatements, no GPU needed) ------ */

/* integration_rule_for_order, src/integration/integration_rules.c:40-73: nodes[order+1]
 * ascending, weights[order+1]; accuracy may be NULL.  Gauss-Legendre:
 * src/integration/gauss_legendre.c:7-50, Gauss-Lobatto: src/integration/gauss_lobatto.c:14-61
 * (tol 1e-14, 1000 iterations). */
IB200_API int ib200_rule_1d(integration_spec_t spec, double *nodes, double *weights, unsigned *accuracy);
/* integration_rule_for_accuracy, src/integration/integration_rules.c:15-39 */
IB200_API int ib200_rule_order_for_accuracy(integration_rule_type_t type, unsigned accuracy, unsigned *order);
/* generate_lagrange_roots, src/basis/basis_lagrange.c:11-42 (tol 1e-12, 100 iterations) */
IB200_API int ib200_lagrange_roots(basis_spec_t spec, double *roots);
/* basis_set_t payload, src/basis/basis_set.h:28-55 as filled by lagrange_basis_create
 * (src/basis/basis_lagrange.c:45-71), legendre_basis_create (src/basis/basis_legendre.c:9-40),
 * bernstein_basis_create (src/basis/basis_bernstein.c:9-23): values[n][m] and
 * derivatives[n][m], basis-major, n = basis order+1, m = rule order+1. */
IB200_API int ib200_basis_table_1d(basis_spec_t basis, integration_spec_t rule, double *values, double *derivatives);
/* integration_space_weights / _nodes, src/python/integration_objects.c:633-716, 731-814:
 * weights[Q] (product built in ascending dimension order), nodes[d][Q]. */
IB200_API int ib200_integration_space_weights(unsigned ndim, const integration_spec_t *rules, double *weights);
IB200_API int ib200_integration_space_nodes(unsigned ndim, const integration_spec_t *rules, double *nodes);
/* function_space_values_at_integration_nodes, src/python/function_space_objects.c:301-474:
 * out[Q][N] = prod_dim B_dim[i_dim][q_dim]. */
IB200_API int ib200_values_at_integration_nodes(unsigned ndim, const basis_spec_t *basis,
                                                const integration_spec_t *rules, double *out);
/* basis_compute_at_point_values / _derivatives, src/basis/basis_set.c:462-556 (BasisSpecs.values /
 * .derivatives, src/python/basis_objects.c:351-491): out[cnt][n] at arbitrary points. */
IB200_API int ib200_basis_at_points(basis_spec_t basis, size_t cnt, const double *x, double *values,
                                    double *derivatives);
/* basis_compute_outer_product_basis, src/basis/basis_set.c:577-633 (FunctionSpace.evaluate,
 * src/python/function_space_objects.c:142-276): out[cnt][N]; x[k] points to cnt coordinates. */
IB200_API int ib200_function_space_evaluate(unsigned ndim, const basis_spec_t *basis, size_t cnt,
                                            const double *const *x, double *out);
/* src/polynomials on caller-supplied nodes -- the module-level functions of the extension (src/python/module.c:102-560):
 * lagrange1d / dlagrange1d / d2lagrange1d = out[cnt][n_roots], derivative = 0 / 1 / 2
 * (lagrange_polynomial_values, _first_derivative, _second_derivative: src/polynomials/lagrange.c:59-106, 317-381, 739-808) */
IB200_API int ib200_lagrange_samples(unsigned n_roots, const double *roots, size_t cnt, const double *x, unsigned derivative,
                                     double *out);
/* compute_gll(order, max_iter, tol) (src/python/module.c:559-621; src/integration/gauss_lobatto.c:14-61) */
IB200_API int ib200_gauss_lobatto(unsigned order, unsigned max_iter, double tol, double *nodes, double *weights,
                                  int *non_converged);
/* bernstein1d(order, t) = out[cnt][order+1] on the unit interval (src/python/module.c:646-686; bernstein.c:31-51) and
 * bernstein_coefficients (in place power series -> Bernstein, src/python/module.c:706-718; bernstein.c:8-29) */
IB200_API int ib200_bernstein_samples(unsigned order, size_t cnt, const double *t, double *out);
IB200_API int ib200_bernstein_from_power_series(unsigned n, double *coefficients);

/* ---- context ---------------------------------------------------------------------------- */

typedef struct ib200_ctx ib200_ctx_t;

IB200_API int ib200_device_count(int *count);
/* One context per GPU per process: owns a stream, pinned staging buffers, the device table
 * cache keyed like the reference registries (basis type, order, rule type, rule order;
 * src/basis/basis_set.c:212-287) and scratch.  Thread-safe (internal mutex). */
IB200_API int ib200_ctx_create(int device, ib200_ctx_t **out);
IB200_API int ib200_ctx_destroy(ib200_ctx_t *ctx);
/* Waits for the context stream.  Also reports what the kernels queued by IB200_DEVICE calls found since the last
 * synchronize: INTERP_ERROR_BAD_SYSTEM when a batched solve met a system it could not factorise (singular, indefinite
 * or non-finite -- np.linalg.solve raises LinAlgError there, python/interplib/integration.py:221).  IB200_HOST calls
 * report the same condition from the call itself. */
IB200_API int ib200_ctx_synchronize(ib200_ctx_t *ctx);
/* adopt an external cudaStream_t (e.g. torch's current stream); NULL restores the own stream */
IB200_API int ib200_ctx_set_stream(ib200_ctx_t *ctx, void *cuda_stream);
/* name[len] = device name; sm_count, cc = major*10+minor, total device bytes (each may be NULL) */
IB200_API int ib200_ctx_info(ib200_ctx_t *ctx, char *name, size_t len, int *sm_count, int *cc, size_t *mem_bytes);
/* number of kernel launches / entries in the device table cache since creation (IntegrationRegistry.usage()
 * analogue, src/python/integration_objects.c:90-149) */
IB200_API int ib200_ctx_stats(ib200_ctx_t *ctx, unsigned long long *kernel_launches, unsigned *cached_tables);
IB200_API int ib200_ctx_clear_cache(ib200_ctx_t *ctx);
/* force the generic (any-dimension, any-order) kernels instead of the specialised ones: testing aid */
IB200_API int ib200_ctx_set_force_generic(ib200_ctx_t *ctx, int on);
/* quadrature points per dimension from which the 2-D mass kernels run stage 2 on FP64 DMMA tiles (mma.sync m8n8k4)
 * instead of DFMA; 0 disables the DMMA kernel.  Default 13, overridable with $IB200_MMA_MIN_POINTS.  Tuning/testing aid. */
IB200_API int ib200_ctx_set_mma_min_points(ib200_ctx_t *ctx, int points);
/* compile-time specialised low-order mass kernel (kernels_mass2d_lo.cu) on/off; default on, $IB200_NO_LO_KERNEL=1
 * turns it off.  Tuning/testing aid: with it off the run-time-shaped kernel serves the same shapes. */
IB200_API int ib200_ctx_set_lo_kernel(ib200_ctx_t *ctx, int on);
/* determinant-weighted mass matrices of bilinear quads as a sum of three Kronecker products of 1-D moment tables
 * (kernels_mass2d_kron.cu: 2 FP64 operations per entry) on/off; default on, $IB200_NO_KRON=1 turns it off: the
 * sum-factorised kernels (DFMA / DMMA stage 2) then serve the same shapes, as they do for every other weight. */
IB200_API int ib200_ctx_set_kron(ib200_ctx_t *ctx, int on);
/* ... and for the gradient matrices (compute_gradient_mass_matrix on bilinear quads: the cofactor weight is linear in one
 * reference coordinate; matching spaces included, with the reference's lower-triangle mirroring).  Default OFF,
 * $IB200_KRON_GRADIENT=1 turns it on: 50-150x the generic kernel that serves matching spaces otherwise, but a cofactor that
 * is small against the others (x_xi1 of a nearly axis-aligned quad) is then rounded differently from the reference, which
 * forms it from O(|x|) products -- agreement 1e-12 of the largest of the four gradient matrices of an element instead of
 * 1e-12 of each. */
IB200_API int ib200_ctx_set_kron_gradient(ib200_ctx_t *ctx, int on);
/* primal L2 projections of bilinear quads by fast diagonalisation (kernels_fastdiag.cu) on/off; default on,
 * $IB200_NO_FASTDIAG=1 turns it off: ib200_l2_project_mapped_batch then assembles and factorises every mass matrix. */
IB200_API int ib200_ctx_set_fastdiag(ib200_ctx_t *ctx, int on);
/* host-pointer calls: elements per staged chunk (0 = automatic) */
IB200_API int ib200_ctx_set_chunk_elements(ib200_ctx_t *ctx, size_t elements);

/* device / pinned memory and stream-ordered copies for callers that keep results resident */
IB200_API int ib200_device_alloc(ib200_ctx_t *ctx, size_t bytes, void **dptr);
IB200_API int ib200_device_free(ib200_ctx_t *ctx, void *dptr);
IB200_API int ib200_pinned_alloc(ib200_ctx_t *ctx, size_t bytes, void **hptr);
IB200_API int ib200_pinned_free(ib200_ctx_t *ctx, void *hptr);
IB200_API int ib200_copy_h2d(ib200_ctx_t *ctx, void *dst_dev, const void *src_host, size_t bytes);
IB200_API int ib200_copy_d2h(ib200_ctx_t *ctx, void *dst_host, const void *src_dev, size_t bytes);
IB200_API int ib200_device_memset(ib200_ctx_t *ctx, void *dptr, int byte, size_t bytes);
/* CUDA-event stopwatch on the context stream (slot 0..7): start, stop, elapsed (syncs on stop event) */
IB200_API int ib200_timer_start(ib200_ctx_t *ctx, unsigned slot);
IB200_API int ib200_timer_stop(ib200_ctx_t *ctx, unsigned slot);
IB200_API int ib200_timer_elapsed_ms(ib200_ctx_t *ctx, unsigned slot, float *ms);

/* ---- A5/A6 (K1): batched reconstruction at integration points ---------------------------- */
/* compute_integration_point_values[_derivatives], src/python/degrees_of_freedom.c:232-303
 * (entries :426-510, :622-722).  For element e and field f:
 *   out[e][f][q] = sum_i ( prod_dim T_dim[i_dim][q_dim] ) * dofs[e][i],
 *   T_dim = D (derivative table) if bit `dim` of deriv_masks[f] is set, else B.
 * deriv_masks = {0} is reconstruct_at_integration_points, {1<<k} the derivative along k;
 * {0, 1, 2} is the 2-D "basis + gradient" evaluation (u, du/dxi0, du/dxi1) in one pass.
 * dofs[E][N], out[E][n_fields][Q]. */
IB200_API int ib200_reconstruct_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis,
                                      const integration_spec_t *rules, size_t n_elements, const double *dofs,
                                      unsigned n_fields, const unsigned *deriv_masks, double *out, int where);

/* ---- N2: arbitrary reference points, element batched --------------------------- */
/* FunctionSpace.evaluate (src/python/function_space_objects.c:142-276 ->
 * basis_compute_outer_product_basis, src/basis/basis_set.c:577-633) for n_points points per
 * element: x[E][ndim][P] -> out[E][P][N], out = ((1 * b0[i0]) * b1[i1]) * ...  (1..3 dimensions). */
IB200_API int ib200_evaluate_points_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis, size_t n_elements,
                                          size_t n_points, const double *x, double *out, int where);
/* reconstruct(dof, *x), python/interplib/degrees_of_freedom.py:10-37, with optional reference
 * derivatives (BasisSpecs.derivatives, src/basis/basis_set.c:491-556, per flagged dimension):
 *   out[e][f][p] = sum_i dofs[e][i] * prod_k T_k[i_k](x[e][k][p]),  T_k = basis derivative if bit k
 * of deriv_masks[f] is set.  dofs[E][N], x[E][ndim][P], out[E][n_fields][P]. */
IB200_API int ib200_reconstruct_points_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis, size_t n_elements,
                                             const double *dofs, size_t n_points, const double *x, unsigned n_fields,
                                             const unsigned *deriv_masks, double *out, int where);

/* ---- N3: incidence matrix and derivative DoFs ---------------------------------- */
/* incidence_matrix(basis_specs), src/python/incidence.c:38-110: out[n][n+1], n = basis.order >= 1
 * (host; Bernstein / Legendre integer patterns, Lagrange by lagrange_polynomial_first_derivative_2). */
IB200_API int ib200_incidence_matrix(basis_spec_t basis, double *out);
/* DegreesOfFreedom.derivative(idim), src/python/degrees_of_freedom.c:738-945, for a batch of elements:
 * dofs[E][n_0+1]..[n_idim+1]..  ->  out[E][n_0+1]..[n_idim]..  (the space one order lower along idim);
 * additions in the reference's order, bit-identical results. */
IB200_API int ib200_dof_derivative_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis, unsigned idim,
                                         size_t n_elements, const double *dofs, double *out, int where);

/* ---- N4: k-form bookkeeping (host, integers only) ------------------------------- */
/* covector_basis_t, src/basis/covector_basis.h:21-26 (same bit layout): a wedge product of dx_i
 * as a bit set plus a sign; dimension 0 is the zero bundle. */
enum
{
    IB200_COVECTOR_MAX_DIM = sizeof(unsigned) * 8
};
typedef struct
{
    unsigned dimension : 31;
    unsigned sign : 1;
    unsigned basis_bits;
} ib200_covector_t;
/* covector_basis_make / _rank / _is_zero / _wedge / _hodge / _apply_contra_basis / _determine_order,
 * src/basis/covector_basis.c:27-158.  ib200_covector_order: 0 before, 1 equal, 2 after, 3 not comparable. */
IB200_API int ib200_covector_make(unsigned dimension, int sign, unsigned rank, const unsigned *indices, ib200_covector_t *out);
IB200_API unsigned ib200_covector_rank(ib200_covector_t b);
IB200_API int ib200_covector_is_zero(ib200_covector_t b);
IB200_API int ib200_covector_wedge(ib200_covector_t b1, ib200_covector_t b2, ib200_covector_t *out);
IB200_API int ib200_covector_hodge(ib200_covector_t b, ib200_covector_t *out);
IB200_API int ib200_covector_apply_contra(ib200_covector_t b, unsigned dim, ib200_covector_t *out);
IB200_API int ib200_covector_order(ib200_covector_t b1, ib200_covector_t b2);
/* Components of a k-form in ascending order, the function space of one component (order - 1 along
 * every dimension it contains), the DoF count of a k-form (differential_form_degrees_of_freedom_count,
 * src/basis/de_rham_sequence.h:66-110) and per-element DoF offsets (prefix sums, offsets[E+1];
 * element_collection_form_offsets, src/basis/de_rham_sequence.c:62-88). */
IB200_API unsigned ib200_kform_component_count(unsigned ndim, unsigned k);
IB200_API int ib200_kform_components(unsigned ndim, unsigned k, ib200_covector_t *out);
IB200_API int ib200_kform_component_space(unsigned ndim, const basis_spec_t *basis, ib200_covector_t component,
                                          basis_spec_t *out);
IB200_API int ib200_kform_dof_count(unsigned ndim, const basis_spec_t *basis, unsigned k, size_t *count);
IB200_API int ib200_dof_offsets(size_t n_elements, const size_t *counts, size_t *offsets);

/* ---- A9 (K0): batched element geometry ---------------------------------------------------- */
/* coordinate_map_new + space_map_new, src/python/mappings.c:8-90, 304-556 with
 * matrix_qr_decompose / matrix_back_substitute (src/operations/matrices.c:15-86, 117-141).
 * geom_dofs[E][n_phys][N_g] are the coordinate DoFs in the space `geom_basis`.
 * Outputs (each may be NULL): coords[E][n_phys][1+ndim][Q]  (CoordinateMap storage: values then
 * the ndim reference gradients, src/python/mappings.h:6-12), det[E][Q] (= prod diag(R), the Gram
 * determinant when n_phys > ndim), jinv[E][Q][ndim][n_phys] (left inverse). */
IB200_API int ib200_geometry_batch(ib200_ctx_t *ctx, unsigned ndim, unsigned n_phys, const basis_spec_t *geom_basis,
                                   const integration_spec_t *rules, size_t n_elements, const double *geom_dofs,
                                   double *coords, double *det, double *jinv, int where);

/* ---- A7/A8 (K2): batched element mass matrices --------------------------------------------- */
/* compute_mass_matrix, src/python/mass_matrices.c:161-339 and compute_gradient_mass_matrix,
 * :341-565.  out[E][N_out][N_in]:
 *   out[e][j][i] = sum_q weight[e][q] * W[q] * phi_in_i(q) * phi_out_j(q)
 * W = tensor rule weights; `weight` is NULL (reference-domain integration, IntegrationSpace
 * branch) or the per-point factor [E][Q] (det for the SpaceMap branch of A7, det*Jinv[idx_in][idx_out]
 * for A8).  deriv_dim_in >= 0 replaces the input basis by its derivative along that dimension (A8),
 * -1 = plain mass matrix.  When the two spaces match (function_spaces_match, :142-159) only the
 * lower triangle is integrated and mirrored exactly, like the reference (:326-335, :551-560 --
 * the reference mirrors the gradient matrix too; reproduced). */
IB200_API int ib200_mass_matrix_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis_in,
                                      const basis_spec_t *basis_out, const integration_spec_t *rules,
                                      size_t n_elements, const double *weight, int deriv_dim_in, double *out,
                                      int where);
/* Same with the geometry prologue fused (K0 + K2): the SpaceMap is never materialised.
 * idx_in < 0: mass matrix weighted by det (A7 with a SpaceMap); idx_in >= 0: gradient mass matrix
 * with weight det*Jinv[idx_in][idx_out] (A8 with a SpaceMap). */
IB200_API int ib200_mass_matrix_mapped_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis_in,
                                             const basis_spec_t *basis_out, const integration_spec_t *rules,
                                             unsigned n_phys, const basis_spec_t *geom_basis, size_t n_elements,
                                             const double *geom_dofs, int idx_in, int idx_out, double *out,
                                             int where);

/* The whole k-form element mass matrix in one call: out[E][N][N], N = ib200_kform_dof_count, block (a, b) over the
 * component spaces in component order = compute_mass_matrix(space_b, space_a, smap) with the determinant weight
 * (src/python/mass_matrices.c:264-335 non-matching-space branch; DoF layout src/basis/de_rham_sequence.h:66-110).  The
 * reference has no such function -- a caller loops over the blocks -- so this is the batched form of that
 * loop.  Bilinear quads in the plane (every BASELINE configuration) take ONE launch of the Kronecker kernel; other
 * elements one mapped mass launch per block plus a strided device copy. */
IB200_API int ib200_kform_mass_mapped_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis, unsigned k,
                                            const integration_spec_t *rules, unsigned n_phys,
                                            const basis_spec_t *geom_basis, size_t n_elements, const double *geom_dofs,
                                            double *out, int where);

/* ---- A12 (K3): batched L2-projection right-hand sides -------------------------------------- */
/* projection_l2_dual, python/interplib/integration.py:106-168 (+ _prepare_integration :44-67):
 *   out[e][i] = sum_q fvals[e][q] * W[q] * det[e][q] * Phi[q][i]
 * fvals[E][Q] = the callable sampled at the (mapped) integration nodes; det may be NULL. */
IB200_API int ib200_l2_dual_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis,
                                  const integration_spec_t *rules, size_t n_elements, const double *fvals,
                                  const double *det, double *out, int where);
/* geometry fused: det computed from geom_dofs in the same kernel */
IB200_API int ib200_l2_dual_mapped_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis,
                                         const integration_spec_t *rules, unsigned n_phys,
                                         const basis_spec_t *geom_basis, size_t n_elements,
                                         const double *geom_dofs, const double *fvals, double *out, int where);

/* K2 + K3 in one pass: the pair (M, b) that projection_l2_primal assembles per element
 * (python/interplib/integration.py:171-223: compute_mass_matrix(space, space, smap) and
 * projection_l2_dual(...)): mass[E][N][N] and rhs[E][N] from geom_dofs[E][n_phys][N_g] and
 * fvals[E][Q].  Low-order bilinear quads take ONE fused kernel (geometry, mass matrix and
 * right-hand side share the determinant); every other shape runs the two operations back to
 * back.  Results are bit-identical to the separate calls. */
IB200_API int ib200_l2_system_mapped_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis,
                                           const integration_spec_t *rules, unsigned n_phys,
                                           const basis_spec_t *geom_basis, size_t n_elements,
                                           const double *geom_dofs, const double *fvals, double *mass,
                                           double *rhs, int where);

/* ---- N1 ("next"): batched primal solve M u = b --------------------------------- */
/* projection_l2_primal, python/interplib/integration.py:171-223 (np.linalg.solve per element).
 * mass[E][N][N] symmetric positive definite, lower triangle read; rhs[E][N] overwritten by the solution.
 * Systems up to N = 127 are factorised in shared memory (matrix read once, left untouched); larger ones
 * in place in device memory (a DEVICE-resident mass is then overwritten by its Cholesky factor; a HOST
 * array is never written back). */
IB200_API int ib200_spd_solve_batch(ib200_ctx_t *ctx, unsigned n, size_t n_elements, double *mass, double *rhs,
                                    int where);

/* Host tables of the fast-diagonalisation projection (csrc/fastdiag.c): with A[j][i] = sum_q w_q B_j(x_q) B_i(x_q) and
 * X[j][i] = sum_q w_q x_q B_j(x_q) B_i(x_q) of the 1-D basis under the 1-D rule, V[n][n] and lambda[n] satisfy V^T A V = I,
 * V^T X V = diag(lambda); Vw[n][m] = w_q sum_j V[j][i] B_j(x_q).  On an element whose Jacobian determinant is
 * alpha + beta xi0 + gamma xi1 (every bilinear quad) the inverse mass matrix of the tensor-product space is
 * (V0 (x) V1) diag(1 / (alpha + beta lambda0_i + gamma lambda1_j)) (V0 (x) V1)^T.  INTERP_ERROR_BAD_SYSTEM when A is not
 * positive definite (rule with fewer points than basis functions). */
IB200_API int ib200_fastdiag_tables(basis_spec_t basis, integration_spec_t rule, double *V, double *lambda, double *Vw);

/* Host tables of the Kronecker-sum mass kernels (csrc/fastdiag.c): out[k][j][i] = sum_q w_q x_q^k Bout_j(x_q) T_i(x_q) for
 * k = 0..kmax, T = Bin or, with deriv_in != 0, its derivative (kmax <= 1 then).  On an element whose Jacobian determinant is
 * a polynomial in the reference coordinates -- linear on a bilinear quad, degree 2 per coordinate on a trilinear hexahedron,
 * constant on a straight segment -- the mass matrix is the Kronecker sum of these tables weighted by the polynomial's
 * coefficients (kernels_mass2d_kron.cu, kernels_mass3d_kron.cu). */
IB200_API int ib200_pair_moments(basis_spec_t basis_out, basis_spec_t basis_in, integration_spec_t rule, unsigned kmax,
                                 int deriv_in, double *out);

/* The whole projection_l2_primal of a batch in one call (python/interplib/integration.py:171-223: dual projection, mass
 * matrix, np.linalg.solve): geometry + f samples in, primal[E][N] out.  Bilinear quads (2-D, order-1 geometry basis, square
 * space) are projected by fast diagonalisation -- no mass matrix exists anywhere (kernels_fastdiag.cu,
 * ib200_ctx_set_fastdiag); for every other element the mass matrices are assembled into the context's scratch, solved
 * there (signed Cholesky) and never leave the device.  A HOST call moves 8(n_phys*N_geom + Q) bytes in and 8 N bytes out
 * per element instead of the 8 N^2 of ib200_l2_system_mapped_batch. */
IB200_API int ib200_l2_project_mapped_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis,
                                            const integration_spec_t *rules, unsigned n_phys,
                                            const basis_spec_t *geom_basis, size_t n_elements, const double *geom_dofs,
                                            const double *fvals, double *primal, int where);

/* ---- 8(e): projection norms -------------------------------------------------------------------- */
/* out[0] = sum_i a[i]*b[i] in FP64 with a fixed (deterministic) reduction tree.  The per-rank partial of the
 * norms a partitioned projection reports -- sum_e b_e.u_e (= ||Pi f||^2), sum_e sum_q f^2 W det (= ||f||^2, with
 * a = f*f and b = W*det) -- which the ranks then all-reduce (NCCL).  The reference computes such sums with
 * np.sum on the host (python/interplib/integration.py:96-103, domains.py:117). */
IB200_API int ib200_dot(ib200_ctx_t *ctx, size_t n, const double *a, const double *b, double *out, int where);

/* ib200_dot for callers that must not allocate GPU memory.  (No counterpart in the reference: it has no device code
 * and no allocator to avoid; this entry point exists for the two kinds of caller named below.)
 *
 * Who needs it: (1) a plug-in living inside an application whose own pool allocator has already taken the GPU's
 * memory -- a raw cudaMalloc then fails, or succeeds behind the pool's back; (2) a solver's time loop on a GPU it
 * shares with other streams -- cudaMalloc and cudaFree stall the whole device.  ib200_dot needs a context (whose
 * creation allocates) and takes its block partials and its staging of host vectors from cudaMalloc.  This variant
 * needs no context and calls no cudaMalloc, cudaFree or other allocating routine, ever: its scratch (592 block
 * partials, a result slot and room to stage two host vectors) is a fixed array of 16 MiB + 8 KiB compiled into the
 * library's device code.  Every GPU that loads the library carries that array until the process exits.
 *
 * out[0] = sum_i a[i]*b[i], same kernels and reduction tree as ib200_dot, hence the same bits.
 *   where == IB200_DEVICE: a, b, out are device pointers on the CURRENT device; the work is queued on `cuda_stream`
 *       (a cudaStream_t; NULL = the default stream) and the call returns without waiting.  Any n.
 *   where == IB200_HOST:   a, b, out are host pointers; the vectors are staged in the built-in scratch, so
 *       n <= IB200_DOT_STATIC_MAX_HOST_N -- larger n is refused with IB200_ERROR_UNSUPPORTED and a message naming
 *       the limit.  Returns after the result has reached out[0] (synchronises `cuda_stream`).
 * The first call on each GPU looks up the device address of the scratch (once per device, with that device current);
 * make that first call before any stream capture begins.  The scratch is one array per GPU shared by all callers and
 * it keeps whatever the last call wrote; each call overwrites everything it later reads, so nothing needs clearing,
 * but two calls whose kernels overlap on the GPU (different streams, no ordering between them) race on it -- order them
 * with events or use one stream.  Host threads are serialised by an internal mutex for the duration of the call only. */
#define IB200_DOT_STATIC_MAX_HOST_N ((size_t)1 << 20)
IB200_API int ib200_dot_static(size_t n, const double *a, const double *b, double *out, int where, void *cuda_stream);

/* ---- 8(e): collectives over the element partition (NCCL over NVLink / NVSwitch) ------------------------------------ */
/* The reference is single process and has no communication layer; the element-partitioned path
 * needs two collectives and nothing else: the all-gather of per-rank [E_g][N] DoF / right-hand-side slabs and the
 * all-reduce(sum) of a few FP64 norms.  One communicator per context (= per GPU, one process per GPU or several
 * contexts in one process); libnccl.so.2 is loaded with dlopen at the first call (the library has no link-time
 * dependency on NCCL), $INTERP_B200_NCCL overrides the path.  All calls are queued on the context stream unless stated
 * otherwise; nothing synchronises with the host.  NCCL failures, including asynchronous ones
 * (ncclCommGetAsyncError), come back as IB200_ERROR_NCCL with the text in ib200_last_error(). */
typedef struct ib200_comm ib200_comm_t;
enum
{
    IB200_COMM_ID_BYTES = 128 /* ncclUniqueId */
};
/* rank 0 creates the id and hands the 128 bytes to the other ranks by any means (file, socket, torch.distributed ...) */
IB200_API int ib200_comm_unique_id(void *id);
/* collective over all ranks of the new communicator; ctx's device is the rank's GPU */
IB200_API int ib200_comm_create(ib200_ctx_t *ctx, int n_ranks, int rank, const void *id, ib200_comm_t **out);
IB200_API int ib200_comm_destroy(ib200_comm_t *comm);
IB200_API int ib200_comm_info(ib200_comm_t *comm, int *n_ranks, int *rank, int *nccl_version);
/* ncclCommGetAsyncError: INTERP_SUCCESS while the communicator is healthy */
IB200_API int ib200_comm_check(ib200_comm_t *comm);
/* Block partition of the leading element index: rank g owns [g*ceil(E/G), min(E, (g+1)*ceil(E/G))). */
IB200_API int ib200_partition_range(size_t n_elements, int n_ranks, int rank, size_t *first, size_t *count);
/* All-gather of per-rank slabs: local[count_of_this_rank][elem_doubles] -> out[n_elements][elem_doubles] on every rank
 * (device pointers; `local` may alias its own place inside `out`).  Uneven tails need no padding. */
IB200_API int ib200_allgather_slabs(ib200_comm_t *comm, const double *local, size_t n_elements, size_t elem_doubles,
                                    double *out);
/* General form: rank r contributes counts[r] doubles which land at out + offsets[r] on every rank (counts / offsets are
 * host arrays of n_ranks entries, identical on all ranks).  `on_side_stream` != 0 runs the collective on the
 * communicator's own stream, ordered after the work already queued on the context stream, so that kernels queued
 * afterwards overlap the transfer; ib200_comm_join makes the context stream wait for everything sent that way. */
IB200_API int ib200_allgatherv(ib200_comm_t *comm, const double *send, double *out, const size_t *counts,
                               const size_t *offsets, int on_side_stream);
IB200_API int ib200_comm_join(ib200_comm_t *comm);
/* Collective.  Registers `out` (the BASE of an ib200_alloc / cudaMalloc allocation of `bytes` bytes, the same size on every
 * rank) as gather destination: the ranks exchange CUDA IPC handles and map each other's buffers.  From then on
 * ib200_allgatherv / ib200_allgather_slabs with this `out` push the local slab straight into the peers' buffers -- one
 * copy-engine transfer per peer over NVLink, no SM, no staging, no NCCL kernel -- and the arrival is published by a
 * one-double all-reduce (inside ib200_comm_join for side-stream calls).  Ordering contract: every such gather starts with
 * a one-double all-reduce on the calling stream, so no peer writes into this rank's `out` before this rank's stream has
 * reached the call -- earlier work queued on that stream (kernels reading the previous gather, own writes into `out`) is
 * safe, exactly as with an NCCL all-gather; work on OTHER streams of this rank that touches `out` must be ordered by the
 * caller.  IB200_ERROR_UNSUPPORTED, on every rank alike and without error text, when IPC mapping is unavailable or
 * $IB200_NO_P2P_GATHER is set: the NCCL path keeps serving.  IB200_ERROR_BAD_ARGUMENT (this rank; UNSUPPORTED on the
 * others) when `out` is not the base of its allocation. */
IB200_API int ib200_comm_register_gather(ib200_comm_t *comm, double *out, size_t bytes);
IB200_API int ib200_comm_unregister_gather(ib200_comm_t *comm);
/* in-place all-reduce(sum) of n doubles in device memory (projection norms) */
IB200_API int ib200_allreduce_sums(ib200_comm_t *comm, double *values, size_t n);

#ifdef __cplusplus
}
#endif
#endif /* INTERP_B200_H */
