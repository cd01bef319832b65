/* Stand-in for <cutl/allocators.h> of j4nr0th/cutl@46f63ed (submodule absent from
 * /root/reference).  TEST INFRASTRUCTURE ONLY: used to compile
 * the unmodified reference into oracle/_ref.  Field names are fixed by the designated
 * initialisers at reference src/python/module.c:67-72 and test_c/common/common.c:29-34.
 * Contains no arithmetic. */
#ifndef CUTL_SHIM_ALLOCATORS_H
#define CUTL_SHIM_ALLOCATORS_H
#include <stddef.h>

typedef struct cutl_allocator_t
{
    void *(*allocate)(void *state, size_t size);
    void (*deallocate)(void *state, void *ptr);
    void *(*reallocate)(void *state, void *ptr, size_t new_size);
    void *state;
} cutl_allocator_t;

static inline void *cutl_alloc(const cutl_allocator_t *a, size_t size)
{
    return a->allocate(a->state, size);
}
static inline void cutl_dealloc(const cutl_allocator_t *a, void *ptr)
{
    a->deallocate(a->state, ptr);
}
static inline void *cutl_realloc(const cutl_allocator_t *a, void *ptr, size_t new_size)
{
    return a->reallocate(a->state, ptr, new_size);
}
#endif
