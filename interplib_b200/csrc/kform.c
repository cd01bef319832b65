/* N4 -- k-form bookkeeping: the integer algebra that labels the components of a differential form and
 * lays their degrees of freedom out per element.  Host only, no floating point.
 *
 *   covector basis bundles   /root/reference/src/basis/covector_basis.{h,c}   (bit set of dx_i + sign)
 *   k-form DoF counts        /root/reference/src/basis/de_rham_sequence.h:66-110
 *   per-element DoF offsets  /root/reference/src/basis/de_rham_sequence.c:62-88 (prefix sums, discontinuous elements)
 *
 * The GPU side needs none of this to compute: a k-form mass matrix is the set of blocks M_ab =
 * compute_mass_matrix(space_b, space_a, ...) between the component spaces (ib200_mass_matrix_*_batch); these functions
 * say WHICH component spaces exist, in which order, and where element e's DoFs start. */
#include <string.h>

#include "../../include/interp_b200.h"

static unsigned popcount_u(unsigned x)
{
    unsigned c = 0;
    for (; x; x &= x - 1)
        ++c;
    return c;
}

/* covector_basis_make, covector_basis.c:68-87: indices ascending, unique, < dimension */
int ib200_covector_make(unsigned dimension, int sign, unsigned rank, const unsigned *indices, ib200_covector_t *out)
{
    if (!out || (rank && !indices) || dimension == 0 || dimension >= IB200_COVECTOR_MAX_DIM || rank > dimension)
        return IB200_ERROR_BAD_ARGUMENT;
    ib200_covector_t b = {dimension, sign < 0, 0u};
    for (unsigned i = 0; i < rank; ++i)
    {
        if (indices[i] >= dimension || (i > 0 && indices[i] <= indices[i - 1]))
            return IB200_ERROR_BAD_ARGUMENT;
        b.basis_bits |= 1u << indices[i];
    }
    *out = b;
    return INTERP_SUCCESS;
}

unsigned ib200_covector_rank(ib200_covector_t b)
{
    return popcount_u(b.basis_bits);
}

int ib200_covector_is_zero(ib200_covector_t b)
{
    return b.dimension == 0;
}

/* covector_basis_wedge, covector_basis.c:27-61: overlapping components give the zero bundle; every component of b2
 * passes the components of b1 that sit at or above it, each passage flips the sign */
int ib200_covector_wedge(ib200_covector_t b1, ib200_covector_t b2, ib200_covector_t *out)
{
    if (!out || b1.dimension != b2.dimension)
        return IB200_ERROR_BAD_ARGUMENT;
    memset(out, 0, sizeof *out);
    if (b1.basis_bits & b2.basis_bits)
        return INTERP_SUCCESS;
    unsigned sign = b1.sign ^ b2.sign;
    for (unsigned rest = b2.basis_bits; rest;)
    {
        const unsigned next = rest & (rest - 1), bit = rest ^ next;
        rest = next;
        if (popcount_u(b1.basis_bits & ~(bit - 1)) & 1u)
            sign = !sign;
    }
    out->dimension = b1.dimension;
    out->sign = sign;
    out->basis_bits = b1.basis_bits | b2.basis_bits;
    return INTERP_SUCCESS;
}

/* covector_basis_hodge, covector_basis.c:132-158 */
int ib200_covector_hodge(ib200_covector_t b, ib200_covector_t *out)
{
    if (!out)
        return IB200_ERROR_BAD_ARGUMENT;
    *out = b;
    if (b.dimension == 0)
        return INTERP_SUCCESS;
    unsigned bits = b.basis_bits ^ ((1u << b.dimension) - 1u);
    out->basis_bits = bits;
    unsigned flips = 0;
    while (bits)
    {
        const unsigned bit = bits & ~(bits - 1);
        bits ^= bit;
        const unsigned higher = b.basis_bits & ~(bit - 1);
        if (!higher)
            break;
        flips += popcount_u(higher);
    }
    if (flips & 1u)
        out->sign = !out->sign;
    return INTERP_SUCCESS;
}

/* covector_basis_apply_contra_basis, covector_basis.c:89-99 (the component `dim` must be present) */
int ib200_covector_apply_contra(ib200_covector_t b, unsigned dim, ib200_covector_t *out)
{
    if (!out || dim >= b.dimension || !(b.basis_bits & (1u << dim)))
        return IB200_ERROR_BAD_ARGUMENT;
    unsigned sign = b.sign;
    const unsigned bits = b.basis_bits ^ (1u << dim);
    const unsigned lower = bits & (bits - 1);
    if (popcount_u(lower) & 1u)
        sign = !sign;
    out->dimension = b.dimension;
    out->sign = sign;
    out->basis_bits = bits;
    return INTERP_SUCCESS;
}

/* covector_basis_determine_order, covector_basis.c:100-125: 0 before, 1 equal, 2 after, 3 not comparable */
int ib200_covector_order(ib200_covector_t b1, ib200_covector_t b2)
{
    if (b1.dimension != b2.dimension)
        return 3;
    const unsigned rank = popcount_u(b1.basis_bits);
    if (rank != popcount_u(b2.basis_bits))
        return 3;
    if (b1.basis_bits == b2.basis_bits)
        return 1;
    const unsigned u1 = b1.basis_bits & ~b2.basis_bits, u2 = b2.basis_bits & ~b1.basis_bits;
    const unsigned f1 = u1 & ~(u1 - 1), f2 = u2 & ~(u2 - 1);
    return ((f1 < f2) ^ (rank > b1.dimension / 2)) ? 0 : 2;
}

static unsigned long binomial(unsigned n, unsigned k)
{
    unsigned long r = 1;
    for (unsigned i = 1; i <= k; ++i)
        r = r * (n - k + i) / i;
    return r;
}

/* number of components of a k-form in ndim dimensions */
unsigned ib200_kform_component_count(unsigned ndim, unsigned k)
{
    return k > ndim ? 0u : (unsigned)binomial(ndim, k);
}

/* the components (positive sign) in ascending ib200_covector_order; out[ib200_kform_component_count(ndim, k)] */
int ib200_kform_components(unsigned ndim, unsigned k, ib200_covector_t *out)
{
    if (!out || ndim == 0 || ndim >= IB200_COVECTOR_MAX_DIM || k > ndim)
        return IB200_ERROR_BAD_ARGUMENT;
    unsigned cnt = 0;
    for (unsigned bits = 0; bits < (1u << ndim); ++bits)
        if (popcount_u(bits) == k)
        {
            const ib200_covector_t b = {ndim, 0u, bits};
            unsigned pos = cnt++;
            for (; pos > 0 && ib200_covector_order(b, out[pos - 1]) == 0; --pos)
                out[pos] = out[pos - 1];
            out[pos] = b;
        }
    return INTERP_SUCCESS;
}

/* function space of one component: the order drops by one along every dimension the component contains (dx_i carries
 * the edge/derivative direction); 2-D 1-form of order p: dx_0 -> (p-1, p), dx_1 -> (p, p-1), i.e. the two spaces whose
 * blocks BASELINE config 3 computes (de_rham_sequence.h:93-95: (p+1)p + p(p+1) DoFs) */
int ib200_kform_component_space(unsigned ndim, const basis_spec_t *basis, ib200_covector_t component, basis_spec_t *out)
{
    if (!basis || !out || component.dimension != ndim)
        return IB200_ERROR_BAD_ARGUMENT;
    for (unsigned d = 0; d < ndim; ++d)
    {
        out[d] = basis[d];
        if (component.basis_bits & (1u << d))
        {
            if (basis[d].order == 0)
                return IB200_ERROR_BAD_ARGUMENT;
            out[d].order = basis[d].order - 1;
        }
    }
    return INTERP_SUCCESS;
}

/* differential_form_degrees_of_freedom_count, de_rham_sequence.h:66-110, for any dimension */
int ib200_kform_dof_count(unsigned ndim, const basis_spec_t *basis, unsigned k, size_t *count)
{
    if (!basis || !count || ndim == 0 || ndim >= IB200_COVECTOR_MAX_DIM || k > ndim)
        return IB200_ERROR_BAD_ARGUMENT;
    size_t total = 0;
    for (unsigned bits = 0; bits < (1u << ndim); ++bits)
        if (popcount_u(bits) == k)
        {
            size_t n = 1;
            for (unsigned d = 0; d < ndim; ++d)
                n *= (bits & (1u << d)) ? basis[d].order : basis[d].order + 1;
            total += n;
        }
    *count = total;
    return INTERP_SUCCESS;
}

/* element_collection DoF offsets, de_rham_sequence.c:62-88: offsets[0] = 0, offsets[e+1] = offsets[e] + counts[e] */
int ib200_dof_offsets(size_t n_elements, const size_t *counts, size_t *offsets)
{
    if (!offsets || (n_elements && !counts))
        return IB200_ERROR_BAD_ARGUMENT;
    offsets[0] = 0;
    for (size_t e = 0; e < n_elements; ++e)
        offsets[e + 1] = offsets[e] + counts[e];
    return INTERP_SUCCESS;
}
