/* Stand-in for <cutl/iterators/multidim_iteration.h> of j4nr0th/cutl@46f63ed (absent
 * submodule).  TEST INFRASTRUCTURE ONLY.  Semantics are the ones the reference's call
 * sites require: the object lives in caller-provided
 * uninitialised memory, dims are declared in ascending order, the flat index is row-major
 * (last dimension fastest), advance(dim) carries towards dimension 0 and leaves higher
 * dimensions alone, and the end is reached when dimension 0 overflows. */
#ifndef CUTL_SHIM_MULTIDIM_ITERATION_H
#define CUTL_SHIM_MULTIDIM_ITERATION_H
#include <stddef.h>

typedef struct multidim_iterator_t
{
    size_t ndims;
    size_t slots[]; /* slots[2*k] = extent of dim k, slots[2*k+1] = current offset */
} multidim_iterator_t;

static inline size_t multidim_iterator_needed_memory(unsigned ndims)
{
    return sizeof(multidim_iterator_t) + 2 * (size_t)ndims * sizeof(size_t);
}
static inline void multidim_iterator_init_dim(multidim_iterator_t *it, unsigned dim, size_t size)
{
    it->ndims = (size_t)dim + 1;
    it->slots[2 * dim] = size;
    it->slots[2 * dim + 1] = 0;
}
static inline void multidim_iterator_set_to_start(multidim_iterator_t *it)
{
    for (size_t k = 0; k < it->ndims; ++k)
        it->slots[2 * k + 1] = 0;
}
static inline int multidim_iterator_is_at_end(const multidim_iterator_t *it)
{
    return it->slots[1] >= it->slots[0];
}
static inline void multidim_iterator_advance(multidim_iterator_t *it, unsigned dim, size_t n)
{
    size_t k = dim;
    it->slots[2 * k + 1] += n;
    while (k > 0 && it->slots[2 * k + 1] >= it->slots[2 * k])
    {
        const size_t carry = it->slots[2 * k + 1] / it->slots[2 * k];
        it->slots[2 * k + 1] %= it->slots[2 * k];
        k -= 1;
        it->slots[2 * k + 1] += carry;
    }
}
static inline size_t multidim_iterator_get_offset(const multidim_iterator_t *it, unsigned dim)
{
    return it->slots[2 * dim + 1];
}
static inline size_t multidim_iterator_get_flat_index(const multidim_iterator_t *it)
{
    size_t flat = 0;
    for (size_t k = 0; k < it->ndims; ++k)
        flat = flat * it->slots[2 * k] + it->slots[2 * k + 1];
    return flat;
}
static inline size_t multidim_iterator_total_size(const multidim_iterator_t *it)
{
    size_t total = 1;
    for (size_t k = 0; k < it->ndims; ++k)
        total *= it->slots[2 * k];
    return total;
}
#endif
