// C ABI of libinterp_b200.so (include/interp_b200.h): argument checking, the context with its device table
// cache, host<->device staging for host-pointer calls, and dispatch to the CUDA kernels.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>

#include <nvtx3/nvToolsExt.h>

#include "ctx.cuh"

NvtxRange::NvtxRange(const char *name)
{
    nvtxRangePushA(name);
}
NvtxRange::~NvtxRange()
{
    nvtxRangePop();
}

// ---- errors -----------------------------------------------------------------------------------------
static thread_local std::string g_last_error;

int ib200_fail(int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return code;
}

extern "C" const char *ib200_last_error(void)
{
    return g_last_error.c_str();
}

extern "C" const char *ib200_version(void)
{
    return "interp_b200 0.1 (sm_100a, FP64)";
}

// names follow the reference's interp_error_str (src/common/error.c:9-24), messages are our own wording
extern "C" const char *ib200_error_str(int code)
{
    switch (code)
    {
    case INTERP_SUCCESS: return "INTERP_SUCCESS";
    case INTERP_ERROR_NOT_IN_DOMAIN: return "INTERP_ERROR_NOT_IN_DOMAIN";
    case INTERP_ERROR_NOT_INCREASING: return "INTERP_ERROR_NOT_INCREASING";
    case INTERP_ERROR_FAILED_ALLOCATION: return "INTERP_ERROR_FAILED_ALLOCATION";
    case INTERP_ERROR_BAD_SYSTEM: return "INTERP_ERROR_BAD_SYSTEM";
    case INTERP_ERROR_INVALID_ENUM: return "INTERP_ERROR_INVALID_ENUM";
    case INTERP_ERROR_NOT_IN_REGISTRY: return "INTERP_ERROR_NOT_IN_REGISTRY";
    case INTERP_ERROR_GEOID_OUT_OF_RANGE: return "INTERP_ERROR_GEOID_OUT_OF_RANGE";
    case INTERP_ERROR_GEOID_NOT_VALID: return "INTERP_ERROR_GEOID_NOT_VALID";
    case INTERP_ERROR_SURFACE_NOT_CLOSED: return "INTERP_ERROR_SURFACE_NOT_CLOSED";
    case INTERP_ERROR_OBJECT_CONNECTED_TWICE: return "INTERP_ERROR_OBJECT_CONNECTED_TWICE";
    case INTERP_ERROR_MATRIX_DIMS_MISMATCH: return "INTERP_ERROR_MATRIX_DIMS_MISMATCH";
    case IB200_ERROR_CUDA: return "IB200_ERROR_CUDA";
    case IB200_ERROR_NO_DEVICE: return "IB200_ERROR_NO_DEVICE";
    case IB200_ERROR_BAD_ARGUMENT: return "IB200_ERROR_BAD_ARGUMENT";
    case IB200_ERROR_UNSUPPORTED: return "IB200_ERROR_UNSUPPORTED";
    case IB200_ERROR_NCCL: return "IB200_ERROR_NCCL";
    default: return "Unknown";
    }
}

extern "C" const char *ib200_error_msg(int code)
{
    switch (code)
    {
    case INTERP_SUCCESS: return "Success";
    case INTERP_ERROR_NOT_IN_DOMAIN: return "Argument was outside of the domain";
    case INTERP_ERROR_NOT_INCREASING: return "Input was not monotonically increasing";
    case INTERP_ERROR_FAILED_ALLOCATION: return "Memory allocation failed";
    case INTERP_ERROR_BAD_SYSTEM: return "System could not be solved";
    case INTERP_ERROR_INVALID_ENUM: return "Enum value was not valid";
    case INTERP_ERROR_NOT_IN_REGISTRY: return "Object was not in the registry";
    case INTERP_ERROR_GEOID_OUT_OF_RANGE: return "Geometry id was out of range";
    case INTERP_ERROR_GEOID_NOT_VALID: return "Geometry id was not valid";
    case INTERP_ERROR_SURFACE_NOT_CLOSED: return "Surface was not closed";
    case INTERP_ERROR_OBJECT_CONNECTED_TWICE: return "Object was connected twice";
    case INTERP_ERROR_MATRIX_DIMS_MISMATCH: return "Matrix dimensions did not match";
    case IB200_ERROR_CUDA: return "A CUDA call or kernel failed";
    case IB200_ERROR_NO_DEVICE: return "No usable CUDA device (the batched path has no CPU fallback)";
    case IB200_ERROR_BAD_ARGUMENT: return "Bad argument";
    case IB200_ERROR_UNSUPPORTED: return "Request is outside the implemented envelope";
    case IB200_ERROR_NCCL: return "An NCCL call failed or libnccl could not be loaded";
    default: return "Unknown error";
    }
}

// ---- host tables (thin checked wrappers around tables.c) ----------------------------------------------
static int check_basis_type(basis_set_type_t t)
{
    if (t < BASIS_LEGENDRE || t > BASIS_BERNSTEIN)
        return ib200_fail(INTERP_ERROR_INVALID_ENUM, "basis type %d is not valid", (int)t);
    return INTERP_SUCCESS;
}
static int check_rule_type(integration_rule_type_t t)
{
    if (t != INTEGRATION_RULE_TYPE_GAUSS_LEGENDRE && t != INTEGRATION_RULE_TYPE_GAUSS_LOBATTO)
        return ib200_fail(INTERP_ERROR_INVALID_ENUM, "integration rule type %d is not valid", (int)t);
    return INTERP_SUCCESS;
}

extern "C" int ib200_rule_1d(integration_spec_t spec, double *nodes, double *weights, unsigned *accuracy)
{
    if (!nodes || !weights)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "nodes/weights must not be NULL");
    IB_TRY(check_rule_type(spec.type));
    return ib200_tables_rule(spec, nodes, weights, accuracy);
}

extern "C" int ib200_rule_order_for_accuracy(integration_rule_type_t type, unsigned accuracy, unsigned *order)
{
    if (!order)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "order must not be NULL");
    IB_TRY(check_rule_type(type));
    return ib200_tables_order_for_accuracy(type, accuracy, order);
}

extern "C" int ib200_lagrange_roots(basis_spec_t spec, double *roots)
{
    if (!roots)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "roots must not be NULL");
    if (!ib200_tables_is_lagrange(spec.type))
        return ib200_fail(INTERP_ERROR_INVALID_ENUM, "basis type %d has no nodes", (int)spec.type);
    return ib200_tables_lagrange_roots(spec, roots);
}

extern "C" int ib200_basis_table_1d(basis_spec_t basis, integration_spec_t rule, double *values, double *derivatives)
{
    if (!values || !derivatives)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "values/derivatives must not be NULL");
    IB_TRY(check_basis_type(basis.type));
    IB_TRY(check_rule_type(rule.type));
    return ib200_tables_basis(basis, rule, values, derivatives);
}

extern "C" int ib200_integration_space_weights(unsigned ndim, const integration_spec_t *rules, double *weights)
{
    if (!rules || !weights || ndim == 0)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "rules/weights must not be NULL and ndim > 0");
    for (unsigned k = 0; k < ndim; ++k)
        IB_TRY(check_rule_type(rules[k].type));
    return ib200_tables_space_weights(ndim, rules, weights);
}

extern "C" int ib200_integration_space_nodes(unsigned ndim, const integration_spec_t *rules, double *nodes)
{
    if (!rules || !nodes || ndim == 0)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "rules/nodes must not be NULL and ndim > 0");
    for (unsigned k = 0; k < ndim; ++k)
        IB_TRY(check_rule_type(rules[k].type));
    return ib200_tables_space_nodes(ndim, rules, nodes);
}

extern "C" int ib200_values_at_integration_nodes(unsigned ndim, const basis_spec_t *basis,
                                                 const integration_spec_t *rules, double *out)
{
    if (!basis || !rules || !out || ndim == 0)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "basis/rules/out must not be NULL and ndim > 0");
    for (unsigned k = 0; k < ndim; ++k)
    {
        IB_TRY(check_basis_type(basis[k].type));
        IB_TRY(check_rule_type(rules[k].type));
    }
    return ib200_tables_values_at_nodes(ndim, basis, rules, out);
}

extern "C" int ib200_basis_at_points(basis_spec_t basis, size_t cnt, const double *x, double *values,
                                     double *derivatives)
{
    if (!x && cnt)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "x must not be NULL");
    IB_TRY(check_basis_type(basis.type));
    return ib200_tables_basis_at_points(basis, cnt, x, values, derivatives);
}

extern "C" int ib200_lagrange_samples(unsigned n_roots, const double *roots, size_t cnt, const double *x, unsigned derivative,
                                      double *out)
{
    if (derivative > 2)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "derivative order must be 0, 1 or 2, got %u", derivative);
    if ((n_roots && !roots) || (cnt && !x) || (cnt && n_roots && !out))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "roots/x/out must not be NULL");
    return ib200_tables_lagrange_samples(n_roots, roots, cnt, x, derivative, out);
}

extern "C" int ib200_gauss_lobatto(unsigned order, unsigned max_iter, double tol, double *nodes, double *weights,
                                   int *non_converged)
{
    if (!nodes || !weights)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "nodes/weights must not be NULL");
    if (!(tol >= 0.0))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "tolerance must not be negative");
    return ib200_tables_gauss_lobatto(order, max_iter, tol, nodes, weights, non_converged);
}

extern "C" int ib200_bernstein_samples(unsigned order, size_t cnt, const double *t, double *out)
{
    if (cnt && (!t || !out))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "t/out must not be NULL");
    return ib200_tables_bernstein_samples(order, cnt, t, out);
}

extern "C" int ib200_bernstein_from_power_series(unsigned n, double *coefficients)
{
    if (n && !coefficients)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "coefficients must not be NULL");
    return ib200_tables_bernstein_from_power(n, coefficients);
}

extern "C" int ib200_function_space_evaluate(unsigned ndim, const basis_spec_t *basis, size_t cnt,
                                             const double *const *x, double *out)
{
    if (!basis || !x || !out || ndim == 0)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "basis/x/out must not be NULL and ndim > 0");
    for (unsigned k = 0; k < ndim; ++k)
        IB_TRY(check_basis_type(basis[k].type));
    return ib200_tables_function_space_evaluate(ndim, basis, cnt, x, out);
}

// ---- context ---------------------------------------------------------------------------------------
extern "C" int ib200_device_count(int *count)
{
    if (!count)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "count must not be NULL");
    int n = 0;
    const cudaError_t err = cudaGetDeviceCount(&n);
    if (err != cudaSuccess)
    {
        *count = 0;
        return ib200_fail(IB200_ERROR_NO_DEVICE, "cudaGetDeviceCount: %s", cudaGetErrorString(err));
    }
    *count = n;
    return INTERP_SUCCESS;
}

extern "C" int ib200_ctx_create(int device, ib200_ctx_t **out)
{
    if (!out)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "out must not be NULL");
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0)
        return ib200_fail(IB200_ERROR_NO_DEVICE, "no CUDA device visible; the batched hot path has no CPU fallback");
    if (device < 0 || device >= n)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "device %d out of range (0..%d)", device, n - 1);
    ib200_ctx *ctx = new (std::nothrow) ib200_ctx;
    if (!ctx)
        return ib200_fail(INTERP_ERROR_FAILED_ALLOCATION, "out of host memory");
    ctx->device = device;
    if (const char *env = std::getenv("IB200_MMA_MIN_POINTS"))
        ctx->mma_min_points = std::atoi(env) <= 0 ? (1 << 30) : std::atoi(env);
    if (const char *env = std::getenv("IB200_NO_LO_KERNEL"))
        ctx->lo_kernel = std::atoi(env) == 0;
    if (const char *env = std::getenv("IB200_NO_FASTDIAG"))
        ctx->fastdiag = std::atoi(env) == 0;
    if (const char *env = std::getenv("IB200_NO_KRON"))
        ctx->kron = std::atoi(env) == 0;
    if (const char *env = std::getenv("IB200_KRON_GRADIENT"))
        ctx->kron_gradient = std::atoi(env) != 0;
    IB_CUDA(cudaSetDevice(device));
    IB_CUDA(cudaGetDeviceProperties(&ctx->prop, device));
    if (ctx->prop.major < 10)
    {
        const int major = ctx->prop.major, minor = ctx->prop.minor;
        delete ctx;
        return ib200_fail(IB200_ERROR_NO_DEVICE, "device %d is sm_%d%d; this library is built for sm_100a only", device,
                          major, minor);
    }
    IB_CUDA(cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
    ctx->stream = ctx->own_stream;
    for (int i = 0; i < 3; ++i)
    {
        IB_CUDA(cudaStreamCreateWithFlags(&ctx->io_stream[i], cudaStreamNonBlocking));
        IB_CUDA(cudaEventCreateWithFlags(&ctx->io_event[i], cudaEventDisableTiming));
    }
    for (int i = 0; i < 8; ++i)
        for (int j = 0; j < 2; ++j)
            IB_CUDA(cudaEventCreate(&ctx->timer[i][j]));
    IB_CUDA(cudaMalloc((void **)&ctx->sched_ring, sizeof(unsigned long long) * 2 * ib200_ctx::kSchedRing));
    IB_CUDA(cudaMemset(ctx->sched_ring, 0, sizeof(unsigned long long) * 2 * ib200_ctx::kSchedRing));
    IB_CUDA(cudaHostAlloc((void **)&ctx->status_host, sizeof(int), cudaHostAllocMapped));
    *ctx->status_host = 0;
    IB_CUDA(cudaHostGetDevicePointer((void **)&ctx->status_dev, ctx->status_host, 0));
    *out = ctx;
    return INTERP_SUCCESS;
}

static int ctx_check(ib200_ctx *ctx)
{
    if (!ctx)
        return ib200_fail(IB200_ERROR_NO_DEVICE, "context is NULL (no CUDA device?); there is no CPU fallback");
    const cudaError_t err = cudaSetDevice(ctx->device);
    if (err != cudaSuccess)
        return ib200_fail(IB200_ERROR_CUDA, "cudaSetDevice(%d): %s", ctx->device, cudaGetErrorString(err));
    return INTERP_SUCCESS;
}

extern "C" int ib200_ctx_clear_cache(ib200_ctx_t *ctx)
{
    IB_TRY(ctx_check(ctx));
    std::lock_guard<std::mutex> lock(ctx->mutex);
    IB_CUDA(cudaDeviceSynchronize());
    for (auto &kv : ctx->tables)
        cudaFree(kv.second.d_values);
    for (auto &kv : ctx->rules)
        cudaFree(kv.second.d_nodes);
    for (auto &kv : ctx->point_tabs)
        cudaFree(kv.second);
    ctx->point_tabs.clear();
    for (auto &kv : ctx->inc_tabs)
        cudaFree(kv.second);
    ctx->inc_tabs.clear();
    ctx->tables.clear();
    ctx->rules.clear();
    ctx->plans.clear();
    return INTERP_SUCCESS;
}

extern "C" int ib200_ctx_destroy(ib200_ctx_t *ctx)
{
    if (!ctx)
        return INTERP_SUCCESS;
    IB_TRY(ctx_check(ctx));
    cudaDeviceSynchronize();
    ib200_ctx_clear_cache(ctx);
    for (int i = 0; i < 3; ++i)
    {
        cudaFree(ctx->stage_dev[i]);
        cudaFreeHost(ctx->stage_pin[i]);
        cudaStreamDestroy(ctx->io_stream[i]);
        cudaEventDestroy(ctx->io_event[i]);
    }
    cudaFree(ctx->scratch);
    cudaFree(ctx->dot_partial);
    cudaFree(ctx->sched_ring);
    cudaFreeHost(ctx->status_host);
    for (int i = 0; i < 8; ++i)
        for (int j = 0; j < 2; ++j)
            cudaEventDestroy(ctx->timer[i][j]);
    cudaStreamDestroy(ctx->own_stream);
    delete ctx;
    return INTERP_SUCCESS;
}

int ctx_take_status(ib200_ctx *ctx)
{
    const int st = __atomic_exchange_n(ctx->status_host, 0, __ATOMIC_ACQ_REL);
    if (st & IB200_STATUS_BAD_SYSTEM)
        return ib200_fail(INTERP_ERROR_BAD_SYSTEM, "a batched solve met a matrix that is singular, indefinite or not finite");
    return INTERP_SUCCESS;
}

extern "C" int ib200_ctx_synchronize(ib200_ctx_t *ctx)
{
    IB_TRY(ctx_check(ctx));
    IB_CUDA(cudaStreamSynchronize(ctx->stream));
    return ctx_take_status(ctx);
}

extern "C" int ib200_ctx_set_stream(ib200_ctx_t *ctx, void *cuda_stream)
{
    IB_TRY(ctx_check(ctx));
    ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
    return INTERP_SUCCESS;
}

extern "C" int ib200_ctx_info(ib200_ctx_t *ctx, char *name, size_t len, int *sm_count, int *cc, size_t *mem_bytes)
{
    IB_TRY(ctx_check(ctx));
    if (name && len)
    {
        std::strncpy(name, ctx->prop.name, len - 1);
        name[len - 1] = 0;
    }
    if (sm_count)
        *sm_count = ctx->prop.multiProcessorCount;
    if (cc)
        *cc = ctx->prop.major * 10 + ctx->prop.minor;
    if (mem_bytes)
        *mem_bytes = ctx->prop.totalGlobalMem;
    return INTERP_SUCCESS;
}

extern "C" int ib200_ctx_stats(ib200_ctx_t *ctx, unsigned long long *kernel_launches, unsigned *cached_tables)
{
    IB_TRY(ctx_check(ctx));
    std::lock_guard<std::mutex> lock(ctx->mutex);
    if (kernel_launches)
        *kernel_launches = ctx->launches;
    if (cached_tables)
        *cached_tables = (unsigned)(ctx->tables.size() + ctx->rules.size());
    return INTERP_SUCCESS;
}

extern "C" int ib200_ctx_set_force_generic(ib200_ctx_t *ctx, int on)
{
    IB_TRY(ctx_check(ctx));
    ctx->force_generic = on != 0;
    return INTERP_SUCCESS;
}

extern "C" int ib200_ctx_set_mma_min_points(ib200_ctx_t *ctx, int points)
{
    IB_TRY(ctx_check(ctx));
    ctx->mma_min_points = points <= 0 ? (1 << 30) : points;
    return INTERP_SUCCESS;
}

extern "C" int ib200_ctx_set_lo_kernel(ib200_ctx_t *ctx, int on)
{
    IB_TRY(ctx_check(ctx));
    ctx->lo_kernel = on != 0;
    return INTERP_SUCCESS;
}

extern "C" int ib200_ctx_set_kron(ib200_ctx_t *ctx, int on)
{
    IB_TRY(ctx_check(ctx));
    ctx->kron = on != 0;
    return INTERP_SUCCESS;
}

extern "C" int ib200_ctx_set_kron_gradient(ib200_ctx_t *ctx, int on)
{
    IB_TRY(ctx_check(ctx));
    ctx->kron_gradient = on != 0;
    return INTERP_SUCCESS;
}

extern "C" int ib200_ctx_set_fastdiag(ib200_ctx_t *ctx, int on)
{
    IB_TRY(ctx_check(ctx));
    ctx->fastdiag = on != 0;
    return INTERP_SUCCESS;
}

extern "C" int ib200_ctx_set_chunk_elements(ib200_ctx_t *ctx, size_t elements)
{
    IB_TRY(ctx_check(ctx));
    ctx->chunk_elements = elements;
    return INTERP_SUCCESS;
}

extern "C" int ib200_device_alloc(ib200_ctx_t *ctx, size_t bytes, void **dptr)
{
    IB_TRY(ctx_check(ctx));
    if (!dptr)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "dptr must not be NULL");
    const cudaError_t err = cudaMalloc(dptr, bytes ? bytes : 1);
    if (err != cudaSuccess)
    {
        cudaGetLastError();
        return ib200_fail(INTERP_ERROR_FAILED_ALLOCATION, "cudaMalloc(%zu bytes): %s", bytes, cudaGetErrorString(err));
    }
    return INTERP_SUCCESS;
}

extern "C" int ib200_device_free(ib200_ctx_t *ctx, void *dptr)
{
    IB_TRY(ctx_check(ctx));
    IB_CUDA(cudaFree(dptr));
    return INTERP_SUCCESS;
}

extern "C" int ib200_pinned_alloc(ib200_ctx_t *ctx, size_t bytes, void **hptr)
{
    IB_TRY(ctx_check(ctx));
    if (!hptr)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "hptr must not be NULL");
    const cudaError_t err = cudaMallocHost(hptr, bytes ? bytes : 1);
    if (err != cudaSuccess)
    {
        cudaGetLastError();
        return ib200_fail(INTERP_ERROR_FAILED_ALLOCATION, "cudaMallocHost(%zu bytes): %s", bytes, cudaGetErrorString(err));
    }
    return INTERP_SUCCESS;
}

extern "C" int ib200_pinned_free(ib200_ctx_t *ctx, void *hptr)
{
    IB_TRY(ctx_check(ctx));
    IB_CUDA(cudaFreeHost(hptr));
    return INTERP_SUCCESS;
}

extern "C" int ib200_copy_h2d(ib200_ctx_t *ctx, void *dst_dev, const void *src_host, size_t bytes)
{
    IB_TRY(ctx_check(ctx));
    IB_CUDA(cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    IB_CUDA(cudaStreamSynchronize(ctx->stream));
    return INTERP_SUCCESS;
}

extern "C" int ib200_copy_d2h(ib200_ctx_t *ctx, void *dst_host, const void *src_dev, size_t bytes)
{
    IB_TRY(ctx_check(ctx));
    IB_CUDA(cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    IB_CUDA(cudaStreamSynchronize(ctx->stream));
    return INTERP_SUCCESS;
}

extern "C" int ib200_device_memset(ib200_ctx_t *ctx, void *dptr, int byte, size_t bytes)
{
    IB_TRY(ctx_check(ctx));
    IB_CUDA(cudaMemsetAsync(dptr, byte, bytes, ctx->stream));
    return INTERP_SUCCESS;
}

extern "C" int ib200_timer_start(ib200_ctx_t *ctx, unsigned slot)
{
    IB_TRY(ctx_check(ctx));
    if (slot >= 8)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "timer slot %u out of range", slot);
    IB_CUDA(cudaEventRecord(ctx->timer[slot][0], ctx->stream));
    return INTERP_SUCCESS;
}

extern "C" int ib200_timer_stop(ib200_ctx_t *ctx, unsigned slot)
{
    IB_TRY(ctx_check(ctx));
    if (slot >= 8)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "timer slot %u out of range", slot);
    IB_CUDA(cudaEventRecord(ctx->timer[slot][1], ctx->stream));
    return INTERP_SUCCESS;
}

extern "C" int ib200_timer_elapsed_ms(ib200_ctx_t *ctx, unsigned slot, float *ms)
{
    IB_TRY(ctx_check(ctx));
    if (slot >= 8 || !ms)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "bad timer slot or NULL output");
    IB_CUDA(cudaEventSynchronize(ctx->timer[slot][1]));
    IB_CUDA(cudaEventElapsedTime(ms, ctx->timer[slot][0], ctx->timer[slot][1]));
    return INTERP_SUCCESS;
}

// ---- device table cache (the registries' device-side mirror) ----------------------------------------------
int ctx_get_rule(ib200_ctx *ctx, integration_spec_t r, const DeviceRule **out)
{
    const auto key = std::make_tuple((int)r.type, r.order);
    auto it = ctx->rules.find(key);
    if (it == ctx->rules.end())
    {
        DeviceRule rule;
        rule.m = (int)r.order + 1;
        rule.h_nodes.resize(2 * (size_t)rule.m);
        IB_TRY(ib200_tables_rule(r, rule.h_nodes.data(), rule.h_nodes.data() + rule.m, nullptr));
        IB_CUDA(cudaMalloc(&rule.d_nodes, sizeof(double) * 2 * rule.m));
        IB_CUDA(cudaMemcpy(rule.d_nodes, rule.h_nodes.data(), sizeof(double) * 2 * rule.m, cudaMemcpyHostToDevice));
        it = ctx->rules.emplace(key, std::move(rule)).first;
    }
    *out = &it->second;
    return INTERP_SUCCESS;
}

int ctx_get_table(ib200_ctx *ctx, basis_spec_t b, integration_spec_t r, const DeviceTable **out)
{
    const auto key = std::make_tuple((int)b.type, b.order, (int)r.type, r.order);
    auto it = ctx->tables.find(key);
    if (it == ctx->tables.end())
    {
        DeviceTable tab;
        tab.n = (int)b.order + 1;
        tab.m = (int)r.order + 1;
        const size_t cnt = (size_t)tab.n * tab.m;
        tab.h_values.resize(2 * cnt);
        IB_TRY(ib200_tables_basis(b, r, tab.h_values.data(), tab.h_values.data() + cnt));
        IB_CUDA(cudaMalloc(&tab.d_values, sizeof(double) * 2 * cnt));
        IB_CUDA(cudaMemcpy(tab.d_values, tab.h_values.data(), sizeof(double) * 2 * cnt, cudaMemcpyHostToDevice));
        it = ctx->tables.emplace(key, std::move(tab)).first;
    }
    *out = &it->second;
    return INTERP_SUCCESS;
}

int ctx_make_op(ib200_ctx *ctx, unsigned ndim, const basis_spec_t *basis_in, const basis_spec_t *basis_out,
                const integration_spec_t *rules, OpDesc *op, OpHost *host)
{
    if (ndim == 0 || ndim > IB200_MAX_DIM)
        return ib200_fail(IB200_ERROR_UNSUPPORTED, "reference dimension %u not supported by the batched kernels (1..%d)", ndim,
                          (int)IB200_MAX_DIM);
    if (!basis_in || !rules)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "basis/rules must not be NULL");
    std::memset(op, 0, sizeof *op);
    std::memset(host, 0, sizeof *host);
    op->ndim = (int)ndim;
    long long n_in = 1, n_out = 1, q = 1;
    for (unsigned k = 0; k < ndim; ++k)
    {
        IB_TRY(check_basis_type(basis_in[k].type));
        IB_TRY(check_rule_type(rules[k].type));
        const basis_spec_t bo = basis_out ? basis_out[k] : basis_in[k];
        IB_TRY(check_basis_type(bo.type));
        const DeviceTable *ti, *to;
        const DeviceRule *r;
        IB_TRY(ctx_get_table(ctx, basis_in[k], rules[k], &ti));
        IB_TRY(ctx_get_table(ctx, bo, rules[k], &to));
        IB_TRY(ctx_get_rule(ctx, rules[k], &r));
        DimTab &d = op->dim[k];
        d.n_in = ti->n;
        d.n_out = to->n;
        d.m = ti->m;
        d.Bin = ti->d_values;
        d.Din = ti->d_values + (size_t)ti->n * ti->m;
        d.Bout = to->d_values;
        d.w = r->d_nodes + r->m;
        host->Bin[k] = ti->h_values.data();
        host->Din[k] = ti->h_values.data() + (size_t)ti->n * ti->m;
        host->Bout[k] = to->h_values.data();
        host->Dout[k] = to->h_values.data() + (size_t)to->n * to->m;
        host->w[k] = r->h_nodes.data() + r->m;
        n_in *= d.n_in;
        n_out *= d.n_out;
        q *= d.m;
    }
    for (unsigned k = ndim; k < IB200_MAX_DIM; ++k)
    {
        op->dim[k].n_in = op->dim[k].n_out = op->dim[k].m = 1;
    }
    if (n_in > (1 << 24) || n_out > (1 << 24) || q > (1 << 24) || n_in * n_out > (1ll << 30))
        return ib200_fail(IB200_ERROR_UNSUPPORTED, "element too large (N_in=%lld N_out=%lld Q=%lld)", n_in, n_out, q);
    op->N_in = (int)n_in;
    op->N_out = (int)n_out;
    op->Q = (int)q;
    return INTERP_SUCCESS;
}

int ctx_scratch(ib200_ctx *ctx, size_t bytes, void **out)
{
    if (ctx->scratch_bytes < bytes)
    {
        if (ctx->scratch)
        {
            IB_CUDA(cudaStreamSynchronize(ctx->stream));
            IB_CUDA(cudaFree(ctx->scratch));
            ctx->scratch = nullptr;
            ctx->scratch_bytes = 0;
        }
        const cudaError_t err = cudaMalloc(&ctx->scratch, bytes);
        if (err != cudaSuccess)
        {
            cudaGetLastError();
            return ib200_fail(INTERP_ERROR_FAILED_ALLOCATION, "cudaMalloc(%zu bytes of scratch): %s", bytes,
                              cudaGetErrorString(err));
        }
        ctx->scratch_bytes = bytes;
    }
    *out = ctx->scratch;
    return INTERP_SUCCESS;
}

// ---- batched execution: device pointers run in place, host pointers are streamed in chunks --------------------
namespace
{

enum ArrayKind
{
    ARR_IN,
    ARR_OUT,
    ARR_INOUT,
    ARR_SCRATCH
};

struct BatchArray
{
    ArrayKind kind;
    void *ptr;        // user pointer (host or device); NULL allowed for optional arrays and for scratch
    size_t elem_bytes; // bytes per element
};

// Copies between a caller's pageable array and a pinned staging slot: one host thread moves about 10 GB/s, a fifth of what
// PCIe then carries, so slabs of 2 MB and more are split over a few threads (IB200_STAGE_THREADS, default 4, 1 = off).
void stage_memcpy(void *dst, const void *src, size_t bytes)
{
    constexpr int kMaxParts = 16; // the caller + 15 workers
    static const int n_threads = [] {
        const char *env = getenv("IB200_STAGE_THREADS");
        const int want = env ? atoi(env) : 4;
        const int hw = (int)std::thread::hardware_concurrency();
        return std::max(1, std::min({want, hw > 1 ? hw / 2 : 1, kMaxParts}));
    }();
    if (n_threads == 1 || bytes < ((size_t)2 << 20))
    {
        std::memcpy(dst, src, bytes);
        return;
    }
    const size_t part = ((bytes / n_threads) + 4095) & ~(size_t)4095;
    std::thread workers[kMaxParts - 1];
    int n = 0;
    for (size_t off = part; off < bytes; off += part, ++n)
    {
        // the last worker takes whatever is left, so the parts always cover [0, bytes)
        const size_t len = (n == kMaxParts - 2) ? bytes - off : std::min(part, bytes - off);
        workers[n] = std::thread([=] { std::memcpy((char *)dst + off, (const char *)src + off, len); });
        if (n == kMaxParts - 2)
        {
            ++n;
            break;
        }
    }
    std::memcpy(dst, src, std::min(part, bytes));
    for (int i = 0; i < n; ++i)
        workers[i].join();
}

bool host_pointer_is_pinned(const void *p)
{
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, p) != cudaSuccess)
    {
        cudaGetLastError();
        return false;
    }
    return attr.type == cudaMemoryTypeHost;
}

int ensure_stage(ib200_ctx *ctx, int slot, size_t dev_bytes, size_t pin_bytes)
{
    if (ctx->stage_dev_bytes[slot] < dev_bytes)
    {
        if (ctx->stage_dev[slot])
            IB_CUDA(cudaFree(ctx->stage_dev[slot]));
        ctx->stage_dev[slot] = nullptr;
        ctx->stage_dev_bytes[slot] = 0;
        const cudaError_t err = cudaMalloc(&ctx->stage_dev[slot], dev_bytes);
        if (err != cudaSuccess)
        {
            cudaGetLastError();
            return ib200_fail(INTERP_ERROR_FAILED_ALLOCATION, "cudaMalloc(%zu staging bytes): %s", dev_bytes,
                              cudaGetErrorString(err));
        }
        ctx->stage_dev_bytes[slot] = dev_bytes;
    }
    if (ctx->stage_pin_bytes[slot] < pin_bytes)
    {
        if (ctx->stage_pin[slot])
            IB_CUDA(cudaFreeHost(ctx->stage_pin[slot]));
        ctx->stage_pin[slot] = nullptr;
        ctx->stage_pin_bytes[slot] = 0;
        const cudaError_t err = cudaMallocHost(&ctx->stage_pin[slot], pin_bytes);
        if (err != cudaSuccess)
        {
            cudaGetLastError();
            return ib200_fail(INTERP_ERROR_FAILED_ALLOCATION, "cudaMallocHost(%zu staging bytes): %s", pin_bytes,
                              cudaGetErrorString(err));
        }
        ctx->stage_pin_bytes[slot] = pin_bytes;
    }
    return INTERP_SUCCESS;
}

inline size_t align_up(size_t v, size_t a)
{
    return (v + a - 1) / a * a;
}

// launch(cfg, first_element, n_elements, device pointers in `arrays` order) -> int
template <class Launch>
int run_batched(ib200_ctx *ctx, size_t E, std::vector<BatchArray> arrays, int where, size_t max_device_chunk, Launch launch)
{
    const int n_arr = (int)arrays.size();
    if (where != IB200_HOST && where != IB200_DEVICE)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "`where` must be IB200_HOST or IB200_DEVICE");
    if (E == 0)
        return INTERP_SUCCESS; // an empty batch launches nothing, on either path
    if (where == IB200_DEVICE)
    {
        LaunchCfg cfg = make_launch_cfg(ctx, ctx->stream);
        // scratch arrays come from the context; bound them by running sub-chunks
        size_t scratch_per_elem = 0;
        for (auto &a : arrays)
            if (a.kind == ARR_SCRATCH)
                scratch_per_elem += align_up(a.elem_bytes, 8);
        size_t chunk = E;
        if (scratch_per_elem)
        {
            const size_t budget = (size_t)512 << 20;
            // a multiple of 1024 elements: sub-launches must start on the 16-byte (bulk copy) and tile boundaries the
            // specialised kernels need -- an odd chunk (512 MB / 288 B = 1 864 135 elements) sent every launch after
            // the first to the generic kernels (C5 shape: 2^21 elements took 8.5 ms instead of 1.6 ms)
            chunk = std::max<size_t>(1024, (budget / scratch_per_elem) & ~(size_t)1023);
            if (max_device_chunk)
                chunk = std::min(chunk, max_device_chunk);
            chunk = std::min(chunk, E);
        }
        char *scratch = nullptr;
        if (scratch_per_elem && E)
        {
            void *p;
            IB_TRY(ctx_scratch(ctx, chunk * scratch_per_elem + 256 * n_arr, &p));
            scratch = (char *)p;
        }
        std::vector<void *> dev(n_arr);
        for (size_t first = 0; first < E; first += chunk)
        {
            const size_t cnt = std::min(chunk, E - first);
            size_t off = 0;
            for (int i = 0; i < n_arr; ++i)
            {
                if (arrays[i].kind == ARR_SCRATCH)
                {
                    dev[i] = scratch + off;
                    off = align_up(off + chunk * arrays[i].elem_bytes, 256);
                }
                else
                {
                    dev[i] = arrays[i].ptr ? (char *)arrays[i].ptr + first * arrays[i].elem_bytes : nullptr;
                }
            }
            IB_TRY(launch(cfg, first, cnt, dev.data()));
        }
        return INTERP_SUCCESS;
    }

    // ---- host pointers: three slots, each with its own stream, device slab and pinned slab -----------------
    size_t per_elem = 0, dev_per_elem = 0; // bytes that cross PCIe / bytes of the device slab (scratch included)
    for (auto &a : arrays)
    {
        if (a.ptr)
            per_elem += a.elem_bytes;
        if (a.ptr || a.kind == ARR_SCRATCH)
            dev_per_elem += a.elem_bytes;
    }
    size_t chunk = ctx->chunk_elements;
    if (chunk == 0)
    {
        // bytes per chunk: large enough to saturate PCIe, small enough to pipeline; a multiple of 512 elements so that
        // every chunk starts on a 4 KB boundary of the host arrays (measured on C2: 48 MB chunks of an odd element count
        // 32.8 ms per 2^20 elements, 16 MB chunks of 8192 elements 30.4 ms = 53 GB/s device-to-host)
        const size_t target = (size_t)16 << 20;
        // scratch never crosses PCIe: it only bounds the chunk through the device slab (256 MB per slot)
        const size_t fit = std::min(target / std::max<size_t>(per_elem, 1), ((size_t)256 << 20) / std::max<size_t>(dev_per_elem, 1));
        chunk = fit >= 512 ? (fit & ~(size_t)511) : std::max<size_t>(256, fit); // big elements: launches of >= 256
    }
    chunk = std::min(chunk, E);
    std::vector<bool> pinned(n_arr, false);
    size_t dev_bytes = 0, pin_bytes = 0;
    std::vector<size_t> dev_off(n_arr, 0), pin_off(n_arr, 0);
    for (int i = 0; i < n_arr; ++i)
    {
        if (!arrays[i].ptr && arrays[i].kind != ARR_SCRATCH)
            continue;
        dev_off[i] = dev_bytes;
        dev_bytes = align_up(dev_bytes + chunk * arrays[i].elem_bytes, 256);
        if (arrays[i].kind == ARR_SCRATCH)
            continue;
        pinned[i] = host_pointer_is_pinned(arrays[i].ptr);
        if (!pinned[i])
        {
            pin_off[i] = pin_bytes;
            pin_bytes = align_up(pin_bytes + chunk * arrays[i].elem_bytes, 256);
        }
    }
    const int n_slots = 3;
    for (int s = 0; s < n_slots; ++s)
        IB_TRY(ensure_stage(ctx, s, dev_bytes, std::max<size_t>(pin_bytes, 1)));

    struct Pending
    {
        bool busy = false;
        size_t first = 0, cnt = 0;
    } pending[3];
    auto drain = [&](int s) -> int {
        if (!pending[s].busy)
            return INTERP_SUCCESS;
        IB_CUDA(cudaEventSynchronize(ctx->io_event[s]));
        for (int i = 0; i < n_arr; ++i)
        {
            const BatchArray &a = arrays[i];
            if (!a.ptr || pinned[i] || (a.kind != ARR_OUT && a.kind != ARR_INOUT))
                continue;
            stage_memcpy((char *)a.ptr + pending[s].first * a.elem_bytes, (char *)ctx->stage_pin[s] + pin_off[i],
                         pending[s].cnt * a.elem_bytes);
        }
        pending[s].busy = false;
        return INTERP_SUCCESS;
    };

    // one chunk: stage in, launch, stage out.  Any failure leaves earlier chunks' asynchronous copies into the caller's
    // arrays in flight, so the error path below waits for them before the call returns.
    std::vector<void *> dev(n_arr);
    auto submit = [&](int slot, size_t first, size_t cnt) -> int {
        IB_TRY(drain(slot));
        cudaStream_t st = ctx->io_stream[slot];
        for (int i = 0; i < n_arr; ++i)
        {
            const BatchArray &a = arrays[i];
            dev[i] = (a.ptr || a.kind == ARR_SCRATCH) ? (char *)ctx->stage_dev[slot] + dev_off[i] : nullptr;
            if (!a.ptr || (a.kind != ARR_IN && a.kind != ARR_INOUT))
                continue;
            const char *src = (const char *)a.ptr + first * a.elem_bytes;
            if (!pinned[i])
            {
                stage_memcpy((char *)ctx->stage_pin[slot] + pin_off[i], src, cnt * a.elem_bytes);
                src = (const char *)ctx->stage_pin[slot] + pin_off[i];
            }
            IB_CUDA(cudaMemcpyAsync(dev[i], src, cnt * a.elem_bytes, cudaMemcpyHostToDevice, st));
        }
        LaunchCfg cfg = make_launch_cfg(ctx, st);
        IB_TRY(launch(cfg, first, cnt, dev.data()));
        for (int i = 0; i < n_arr; ++i)
        {
            const BatchArray &a = arrays[i];
            if (!a.ptr || (a.kind != ARR_OUT && a.kind != ARR_INOUT))
                continue;
            char *dst = pinned[i] ? (char *)a.ptr + first * a.elem_bytes : (char *)ctx->stage_pin[slot] + pin_off[i];
            IB_CUDA(cudaMemcpyAsync(dst, dev[i], cnt * a.elem_bytes, cudaMemcpyDeviceToHost, st));
        }
        IB_CUDA(cudaEventRecord(ctx->io_event[slot], st));
        pending[slot].busy = true;
        pending[slot].first = first;
        pending[slot].cnt = cnt;
        return INTERP_SUCCESS;
    };
    int res = INTERP_SUCCESS, slot = 0;
    for (size_t first = 0; first < E && res == INTERP_SUCCESS; first += chunk, slot = (slot + 1) % n_slots)
        res = submit(slot, first, std::min(chunk, E - first));
    if (res != INTERP_SUCCESS)
    {
        // keep the first error's text; nothing may still be writing into caller memory when we return
        const std::string first_error = g_last_error;
        for (int s = 0; s < n_slots; ++s)
            cudaStreamSynchronize(ctx->io_stream[s]);
        cudaGetLastError();
        g_last_error = first_error;
        return res;
    }
    for (int s = 0; s < n_slots; ++s)
        IB_TRY(drain(s));
    return ctx_take_status(ctx); // everything this call queued has completed
}

bool spaces_match(unsigned ndim, const basis_spec_t *a, const basis_spec_t *b)
{
    // function_spaces_match, /root/reference/src/python/mass_matrices.c:142-159
    for (unsigned k = 0; k < ndim; ++k)
        if (a[k].order != b[k].order || a[k].type != b[k].type)
            return false;
    return true;
}

std::string plan_key(const char *tag, unsigned ndim, const basis_spec_t *bi, const basis_spec_t *bo, const integration_spec_t *r,
                     const basis_spec_t *gb, int a, int b, int c, int d)
{
    std::string k(tag);
    auto add = [&k](long v) { k += ':' + std::to_string(v); };
    add(ndim);
    for (unsigned i = 0; i < ndim; ++i)
    {
        add(bi[i].type), add(bi[i].order);
        add(bo ? bo[i].type : -1), add(bo ? bo[i].order : 0);
        add(r[i].type), add(r[i].order);
        add(gb ? gb[i].type : -1), add(gb ? gb[i].order : 0);
    }
    add(a), add(b), add(c), add(d);
    return k;
}

} // namespace

// ---- K1 ----------------------------------------------------------------------------------------------------
extern "C" int ib200_reconstruct_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis,
                                       const integration_spec_t *rules, size_t n_elements, const double *dofs,
                                       unsigned n_fields, const unsigned *deriv_masks, double *out, int where)
{
    IB_TRY(ctx_check(ctx));
    std::lock_guard<std::mutex> lock(ctx->mutex);
    NvtxRange nvtx("ib200_reconstruct_batch");
    if (!deriv_masks || n_fields == 0 || n_fields > 8)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "need 1..8 fields and their derivative masks");
    if (n_elements && (!dofs || !out))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "dofs/out must not be NULL");
    for (unsigned f = 0; f < n_fields; ++f)
        if (deriv_masks[f] >> ndim)
            return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "derivative mask 0x%x names a dimension >= %u", deriv_masks[f], ndim);
    OpDesc op;
    OpHost host;
    IB_TRY(ctx_make_op(ctx, ndim, basis, nullptr, rules, &op, &host));
    unsigned masks[8];
    std::copy(deriv_masks, deriv_masks + n_fields, masks);
    std::vector<BatchArray> arrays = {{ARR_IN, (void *)dofs, sizeof(double) * op.N_in},
                                      {ARR_OUT, out, sizeof(double) * n_fields * op.Q}};
    return run_batched(ctx, n_elements, arrays, where, 0,
                       [&](const LaunchCfg &cfg, size_t, size_t cnt, void **dev) -> int {
                           if (!cfg.force_generic)
                           {
                               const int res = launch_eval2d_fast(cfg, op, host, cnt, (const double *)dev[0], n_fields,
                                                                  masks, (double *)dev[1]);
                               if (res != IB200_ERROR_UNSUPPORTED)
                                   return res;
                           }
                           return launch_reconstruct_generic(cfg, op, cnt, (const double *)dev[0], n_fields, masks,
                                                             (double *)dev[1]);
                       });
}

// ---- N2: arbitrary reference points -------------------------------------------------------------------------
static int ctx_make_point_op(ib200_ctx *ctx, unsigned ndim, const basis_spec_t *basis, PointOp *op)
{
    if (ndim == 0 || ndim > IB200_POINT_MAX_DIM || !basis)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "arbitrary-point evaluation takes 1..%d dimensions (got %u)",
                          IB200_POINT_MAX_DIM, ndim);
    op->ndim = (int)ndim;
    op->N = 1;
    for (unsigned k = 0; k < ndim; ++k)
    {
        PointDim &d = op->dim[k];
        d.type = (int)basis[k].type;
        d.n = (int)basis[k].order + 1;
        d.roots = d.den = nullptr;
        op->N *= d.n;
        if (ib200_tables_is_lagrange(basis[k].type))
        {
            const auto key = std::make_tuple((int)basis[k].type, basis[k].order);
            auto it = ctx->point_tabs.find(key);
            if (it == ctx->point_tabs.end())
            {
                std::vector<double> h(3 * (size_t)d.n); // roots | denominators | reciprocal denominators
                IB_TRY(ib200_tables_lagrange_roots_denominators(basis[k], h.data(), h.data() + d.n));
                for (int j = 0; j < d.n; ++j)
                    h[2 * (size_t)d.n + j] = 1.0 / h[(size_t)d.n + j];
                double *dev = nullptr;
                IB_CUDA(cudaMalloc((void **)&dev, sizeof(double) * h.size()));
                IB_CUDA(cudaMemcpy(dev, h.data(), sizeof(double) * h.size(), cudaMemcpyHostToDevice));
                it = ctx->point_tabs.emplace(key, dev).first;
            }
            d.roots = it->second;
            d.den = it->second + d.n;
        }
        else if (basis[k].type != BASIS_LEGENDRE && basis[k].type != BASIS_BERNSTEIN)
            return ib200_fail(INTERP_ERROR_INVALID_ENUM, "unknown basis type %d", (int)basis[k].type);
    }
    return INTERP_SUCCESS;
}

extern "C" int ib200_evaluate_points_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis, size_t n_elements,
                                           size_t n_points, const double *x, double *out, int where)
{
    IB_TRY(ctx_check(ctx));
    std::lock_guard<std::mutex> lock(ctx->mutex);
    NvtxRange nvtx("ib200_evaluate_points_batch");
    if (n_elements && n_points && (!x || !out))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "x/out must not be NULL");
    PointOp op;
    IB_TRY(ctx_make_point_op(ctx, ndim, basis, &op));
    if (n_points == 0)
        return INTERP_SUCCESS;
    std::vector<BatchArray> arrays = {{ARR_IN, (void *)x, sizeof(double) * ndim * n_points},
                                      {ARR_OUT, out, sizeof(double) * n_points * op.N}};
    return run_batched(ctx, n_elements, arrays, where, 0, [&](const LaunchCfg &cfg, size_t, size_t cnt, void **dev) -> int {
        return launch_points_evaluate(cfg, op, cnt, n_points, (const double *)dev[0], (double *)dev[1]);
    });
}

extern "C" int ib200_reconstruct_points_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis, size_t n_elements,
                                              const double *dofs, size_t n_points, const double *x, unsigned n_fields,
                                              const unsigned *deriv_masks, double *out, int where)
{
    IB_TRY(ctx_check(ctx));
    std::lock_guard<std::mutex> lock(ctx->mutex);
    NvtxRange nvtx("ib200_reconstruct_points_batch");
    if (!deriv_masks || n_fields == 0 || n_fields > 8)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "need 1..8 fields and their derivative masks");
    if (n_elements && n_points && (!dofs || !x || !out))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "dofs/x/out must not be NULL");
    for (unsigned f = 0; f < n_fields; ++f)
        if (deriv_masks[f] >> ndim)
            return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "derivative mask 0x%x names a dimension >= %u", deriv_masks[f], ndim);
    PointOp op;
    IB_TRY(ctx_make_point_op(ctx, ndim, basis, &op));
    if (n_points == 0)
        return INTERP_SUCCESS;
    unsigned masks[8];
    std::copy(deriv_masks, deriv_masks + n_fields, masks);
    std::vector<BatchArray> arrays = {{ARR_IN, (void *)dofs, sizeof(double) * op.N},
                                      {ARR_IN, (void *)x, sizeof(double) * ndim * n_points},
                                      {ARR_OUT, out, sizeof(double) * n_fields * n_points}};
    return run_batched(ctx, n_elements, arrays, where, 0, [&](const LaunchCfg &cfg, size_t, size_t cnt, void **dev) -> int {
        return launch_points_reconstruct(cfg, op, cnt, n_points, (const double *)dev[0], (const double *)dev[1], n_fields, masks,
                                         (double *)dev[2]);
    });
}

// ---- N1: fast-diagonalisation tables (host) --------------------------------------------------------------------
extern "C" int ib200_fastdiag_tables(basis_spec_t basis, integration_spec_t rule, double *V, double *lambda, double *Vw)
{
    if (!V || !lambda || !Vw)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "V/lambda/Vw must not be NULL");
    const int res = ib200_tables_fastdiag(basis, rule, V, lambda, Vw);
    if (res == INTERP_ERROR_BAD_SYSTEM)
        return ib200_fail(res, "the 1-D mass matrix of basis (%d, %u) under rule (%d, %u) is not positive definite", (int)basis.type,
                          basis.order, (int)rule.type, rule.order);
    if (res != INTERP_SUCCESS)
        return ib200_fail(res, "fast-diagonalisation tables: invalid basis (%d) or rule (%d) type", (int)basis.type, (int)rule.type);
    return INTERP_SUCCESS;
}

extern "C" int ib200_pair_moments(basis_spec_t basis_out, basis_spec_t basis_in, integration_spec_t rule, unsigned kmax, int deriv_in,
                                  double *out)
{
    if (!out)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "out must not be NULL");
    if (kmax > 8 || (deriv_in && kmax > 1))
        return ib200_fail(IB200_ERROR_UNSUPPORTED, "moments up to x^8 (x^1 with the derivative of the input basis)");
    int res;
    if (deriv_in)
    {
        const size_t n = (size_t)(basis_out.order + 1) * (basis_in.order + 1);
        std::vector<double> X(n);
        res = ib200_tables_pair_moments_deriv(basis_out, basis_in, rule, out, X.data());
        if (res == INTERP_SUCCESS && kmax >= 1)
            std::memcpy(out + n, X.data(), n * sizeof(double));
    }
    else
        res = ib200_tables_pair_moments_k(basis_out, basis_in, rule, kmax, out);
    if (res != INTERP_SUCCESS)
        return ib200_fail(res, "pair moments: invalid basis (%d, %d) or rule (%d) type", (int)basis_out.type, (int)basis_in.type,
                          (int)rule.type);
    return INTERP_SUCCESS;
}

// ---- N3: incidence / derivative DoFs ------------------------------------------------------------------------
extern "C" int ib200_incidence_matrix(basis_spec_t basis, double *out)
{
    if (!out)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "out must not be NULL");
    if (basis.order == 0)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "Cannot compute the incidence matrix for a zero-dimensional basis.");
    const int res = ib200_tables_incidence_matrix(basis, out);
    if (res != INTERP_SUCCESS)
        return ib200_fail(res, "incidence matrix: unknown basis type %d", (int)basis.type);
    return INTERP_SUCCESS;
}

extern "C" int ib200_dof_derivative_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis, unsigned idim,
                                          size_t n_elements, const double *dofs, double *out, int where)
{
    IB_TRY(ctx_check(ctx));
    std::lock_guard<std::mutex> lock(ctx->mutex);
    NvtxRange nvtx("ib200_dof_derivative_batch");
    if (ndim == 0 || ndim > IB200_MAX_DIM || !basis)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "need 1..%d dimensions", (int)IB200_MAX_DIM);
    if (idim >= ndim)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "Expected an index between 0 and %u, but got %u.", ndim - 1, idim);
    if (basis[idim].order == 0)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT,
                          "Cannot compute the derivative of a function with order 0 in dimension %u.", idim);
    if (n_elements && (!dofs || !out))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "dofs/out must not be NULL");
    size_t pre = 1, post = 1;
    for (unsigned k = 0; k < idim; ++k)
        pre *= basis[k].order + 1;
    for (unsigned k = idim + 1; k < ndim; ++k)
        post *= basis[k].order + 1;
    const int n = (int)basis[idim].order;
    const double *d_inc = nullptr;
    if (ib200_tables_is_lagrange(basis[idim].type))
    {
        const auto key = std::make_tuple((int)basis[idim].type, basis[idim].order);
        auto it = ctx->inc_tabs.find(key);
        if (it == ctx->inc_tabs.end())
        {
            std::vector<double> h((size_t)n * (n + 1));
            IB_TRY(ib200_tables_incidence_matrix(basis[idim], h.data()));
            double *dev = nullptr;
            IB_CUDA(cudaMalloc((void **)&dev, sizeof(double) * h.size()));
            IB_CUDA(cudaMemcpy(dev, h.data(), sizeof(double) * h.size(), cudaMemcpyHostToDevice));
            it = ctx->inc_tabs.emplace(key, dev).first;
        }
        d_inc = it->second;
    }
    else if (basis[idim].type != BASIS_LEGENDRE && basis[idim].type != BASIS_BERNSTEIN)
        return ib200_fail(INTERP_ERROR_INVALID_ENUM, "Unsupported basis type %d.", (int)basis[idim].type);
    const int type = (int)basis[idim].type;
    std::vector<BatchArray> arrays = {{ARR_IN, (void *)dofs, sizeof(double) * pre * (n + 1) * post},
                                      {ARR_OUT, out, sizeof(double) * pre * n * post}};
    return run_batched(ctx, n_elements, arrays, where, 0, [&](const LaunchCfg &cfg, size_t, size_t cnt, void **dev) -> int {
        return launch_dof_derivative(cfg, type, n, pre, post, d_inc, cnt, (const double *)dev[0], (double *)dev[1]);
    });
}

// ---- K0 ----------------------------------------------------------------------------------------------------
static int check_geometry_dims(unsigned ndim, unsigned n_phys)
{
    if (n_phys < ndim)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT,
                          "can not map %u reference dimensions to %u physical dimensions (need n_phys >= ndim)", ndim, n_phys);
    if (n_phys > IB200_MAX_PHYS)
        return ib200_fail(IB200_ERROR_UNSUPPORTED, "at most %d physical dimensions (got %u)", (int)IB200_MAX_PHYS, n_phys);
    return INTERP_SUCCESS;
}

extern "C" int ib200_geometry_batch(ib200_ctx_t *ctx, unsigned ndim, unsigned n_phys, const basis_spec_t *geom_basis,
                                    const integration_spec_t *rules, size_t n_elements, const double *geom_dofs,
                                    double *coords, double *det, double *jinv, int where)
{
    IB_TRY(ctx_check(ctx));
    std::lock_guard<std::mutex> lock(ctx->mutex);
    NvtxRange nvtx("ib200_geometry_batch");
    IB_TRY(check_geometry_dims(ndim, n_phys));
    if (n_elements && !geom_dofs)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "geom_dofs must not be NULL");
    OpDesc op;
    OpHost host;
    IB_TRY(ctx_make_op(ctx, ndim, geom_basis, nullptr, rules, &op, &host));
    std::vector<BatchArray> arrays = {{ARR_IN, (void *)geom_dofs, sizeof(double) * n_phys * op.N_in},
                                      {ARR_OUT, coords, sizeof(double) * n_phys * (1 + ndim) * op.Q},
                                      {ARR_OUT, det, sizeof(double) * op.Q},
                                      {ARR_OUT, jinv, sizeof(double) * op.Q * ndim * n_phys}};
    return run_batched(ctx, n_elements, arrays, where, 0,
                       [&](const LaunchCfg &cfg, size_t, size_t cnt, void **dev) -> int {
                           return launch_geometry_generic(cfg, op, n_phys, cnt, (const double *)dev[0], (double *)dev[1],
                                                          (double *)dev[2], (double *)dev[3], nullptr, -1, 0);
                       });
}

// K2 dispatch: DMMA kernel for high orders (>= ctx->mma_min_points quadrature points per dimension), else the DFMA
// kernel; IB200_ERROR_UNSUPPORTED falls through to the next candidate and finally to the caller's generic kernel.
static int launch_mass2d_best(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, const OpDesc &op, const OpHost &host,
                              bool match0, bool match1, int deriv, int wsrc_kind, const OpHost *geom_host, int widx_in, int widx_out,
                              size_t cnt, const double *wsrc, double *out)
{
    if (deriv >= 0 && match0 && match1 && op.ndim == 2 && op.N_in == op.N_out && !getenv("IB200_NO_MIRROR_PASS"))
    {
        // Gradient mass matrix on matching spaces: the reference evaluates the lower triangle only and mirrors it
        // (mass_matrices.c:494-495, :551-560).  The sum-factorised kernels compute the true full matrix -- the pair table of
        // the dimension without the derivative is still symmetric, which is all they use of "matching" -- and a second,
        // in-place pass copies the transposed lower triangle over the upper one.  Same Jacobian, same quadrature sums as
        // the non-matching gradient path, hence the same per-matrix bound against the reference.
        // $IB200_MIRROR_CHUNK_MB > 0 runs both passes chunk by chunk (that many MiB of output each), so that a chunk still sits
        // in the 126 MB L2 when the mirror pass reads it back.  Measured SLOWER on B200 at every order (p = 4 / 8 / 16 with
        // 24, 48, 96 MiB chunks: 0.53-0.81 / 0.51-0.72 / 0.29-0.58 of the one-chunk rate, profiles/r03g_chunk_sweep.log --
        // the ramp and tail of many short persistent launches cost more than the saved DRAM reads), hence off by default and
        // kept as a knob only.  Chunks are multiples of 1024 elements where that fits (the specialised kernels' tiling),
        // else of 2 (16-byte alignment of `out` for odd N).
        static const long chunk_mb = getenv("IB200_MIRROR_CHUNK_MB") ? atol(getenv("IB200_MIRROR_CHUNK_MB")) : 0;
        const size_t per_elem = sizeof(double) * (size_t)op.N_in * op.N_out;
        size_t chunk = cnt;
        if (chunk_mb > 0)
        {
            chunk = ((size_t)chunk_mb << 20) / per_elem;
            chunk = chunk >= 1024 ? (chunk & ~(size_t)1023) : std::max<size_t>(2, chunk & ~(size_t)1);
        }
        // element stride of the weight source: fused geometry (kind 2) is bilinear quads in the plane only, geom[e][2][4]
        const size_t wstride = wsrc_kind == 2 ? 8 : (wsrc_kind == 1 ? (size_t)op.Q : 0);
        for (size_t done = 0; done < cnt; done += chunk)
        {
            const size_t n = std::min(chunk, cnt - done);
            double *o = out + done * (size_t)op.N_in * op.N_out;
            const int res = launch_mass2d_best(ctx, cfg, key, op, host, match0 && deriv != 0, match1 && deriv != 1, deriv, wsrc_kind,
                                               geom_host, widx_in, widx_out, n, wsrc ? wsrc + done * wstride : nullptr, o);
            if (res != INTERP_SUCCESS)
                return res; // UNSUPPORTED can only come from the first chunk: nothing has been written then
            IB_TRY(launch_mirror_lower(cfg, n, op.N_in, o));
        }
        return INTERP_SUCCESS;
    }
    if (op.ndim == 2 && op.dim[0].m >= ctx->mma_min_points)
    {
        const int res = launch_mass2d_mma(ctx, cfg, key, op, host, match0, match1, deriv, wsrc_kind, geom_host, widx_in, widx_out,
                                          cnt, wsrc, out);
        if (res != IB200_ERROR_UNSUPPORTED)
            return res;
    }
    if (ctx->lo_kernel)
    {
        const int res = launch_mass2d_lo(ctx, cfg, key, op, host, match0, match1, deriv, wsrc_kind, geom_host, widx_in, cnt, wsrc, out);
        if (res != IB200_ERROR_UNSUPPORTED)
            return res;
    }
    return launch_mass2d_fast(ctx, cfg, key, op, host, match0, match1, deriv, wsrc_kind, geom_host, widx_in, widx_out, cnt, wsrc, out);
}

// ---- K2 ----------------------------------------------------------------------------------------------------
extern "C" int ib200_mass_matrix_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis_in,
                                       const basis_spec_t *basis_out, const integration_spec_t *rules,
                                       size_t n_elements, const double *weight, int deriv_dim_in, double *out,
                                       int where)
{
    IB_TRY(ctx_check(ctx));
    std::lock_guard<std::mutex> lock(ctx->mutex);
    NvtxRange nvtx("ib200_mass_matrix_batch");
    if (!basis_in || !basis_out)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "basis_in/basis_out must not be NULL");
    if (n_elements && !out)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "out must not be NULL");
    if (deriv_dim_in >= (int)ndim)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "Index %d out of bounds for input space with %u dimensions.", deriv_dim_in,
                          ndim);
    OpDesc op;
    OpHost host;
    IB_TRY(ctx_make_op(ctx, ndim, basis_in, basis_out, rules, &op, &host));
    const bool symmetric = spaces_match(ndim, basis_in, basis_out);
    const int deriv = deriv_dim_in < 0 ? -1 : deriv_dim_in;
    const bool match0 = spaces_match(1, basis_in, basis_out);
    const bool match1 = ndim > 1 && spaces_match(1, basis_in + 1, basis_out + 1);
    const std::string key = plan_key("mass", ndim, basis_in, basis_out, rules, nullptr, deriv, weight ? 1 : 0, -1, 0);
    std::vector<BatchArray> arrays = {{ARR_IN, (void *)weight, sizeof(double) * op.Q},
                                      {ARR_OUT, out, sizeof(double) * op.N_out * op.N_in}};
    return run_batched(ctx, n_elements, arrays, where, 0,
                       [&](const LaunchCfg &cfg, size_t, size_t cnt, void **dev) -> int {
                           if (!cfg.force_generic && cnt)
                           {
                               const int res = launch_mass2d_best(ctx, cfg, key, op, host, match0, match1, deriv, weight ? 1 : 0,
                                                                  nullptr, -1, 0, cnt, (const double *)dev[0], (double *)dev[1]);
                               if (res != IB200_ERROR_UNSUPPORTED)
                                   return res;
                           }
                           return launch_mass_generic(cfg, op, symmetric, cnt, (const double *)dev[0], deriv, (double *)dev[1]);
                       });
}

// One mapped mass-matrix operation (spaces, rule, geometry basis, weight component) prepared for launching on device
// pointers: shared by ib200_mass_matrix_mapped_batch and the block-by-block path of ib200_kform_mass_mapped_batch.
struct MappedMassOp
{
    unsigned ndim = 0, n_phys = 0;
    const basis_spec_t *basis_in = nullptr, *basis_out = nullptr, *geom_basis = nullptr;
    const integration_spec_t *rules = nullptr;
    OpDesc op, gop;
    OpHost host, ghost;
    bool symmetric = false, match0 = false, match1 = false, fuse_geometry = false;
    int deriv = -1, idx_out = 0;
    std::string key_fused, key_w, key_kron;
};

static int mapped_mass_prepare(ib200_ctx *ctx, unsigned ndim, const basis_spec_t *basis_in, const basis_spec_t *basis_out,
                               const integration_spec_t *rules, unsigned n_phys, const basis_spec_t *geom_basis, int idx_in,
                               int idx_out, MappedMassOp *m)
{
    m->ndim = ndim;
    m->n_phys = n_phys;
    m->basis_in = basis_in;
    m->basis_out = basis_out;
    m->geom_basis = geom_basis;
    m->rules = rules;
    IB_TRY(ctx_make_op(ctx, ndim, basis_in, basis_out, rules, &m->op, &m->host));
    IB_TRY(ctx_make_op(ctx, ndim, geom_basis, nullptr, rules, &m->gop, &m->ghost));
    m->symmetric = spaces_match(ndim, basis_in, basis_out);
    m->deriv = idx_in < 0 ? -1 : idx_in;
    m->idx_out = idx_out;
    m->match0 = spaces_match(1, basis_in, basis_out);
    m->match1 = ndim > 1 && spaces_match(1, basis_in + 1, basis_out + 1);
    // bilinear quads in the plane: the geometry prologue is fused into the mass kernel (no [E][Q] weight array)
    m->fuse_geometry = ndim == 2 && n_phys == 2 && geom_basis[0].order == 1 && geom_basis[1].order == 1;
    m->key_fused = plan_key("massg", ndim, basis_in, basis_out, rules, geom_basis, m->deriv, 2, m->deriv, idx_out);
    m->key_w = plan_key("mass", ndim, basis_in, basis_out, rules, nullptr, m->deriv, 1, -1, 0);
    m->key_kron = plan_key("kron", ndim, basis_in, basis_out, rules, geom_basis, -1, 2, -1, 0);
    return INTERP_SUCCESS;
}

// Determinant-weighted mass matrices of bilinear quads: the compile-time specialised low-order kernel where it covers the
// shape (it sits on the write roofline), else the Kronecker kernel (two FP64 operations per entry, kernels_mass2d_kron.cu),
// else IB200_ERROR_UNSUPPORTED and the caller's sum-factorised candidates.
static int launch_mass2d_bilinear(ib200_ctx *ctx, const LaunchCfg &cfg, const MappedMassOp &m, size_t cnt, const double *geom,
                                  double *out)
{
    if (!m.fuse_geometry)
        return IB200_ERROR_UNSUPPORTED;
    if (m.deriv >= 0)
    {
        // gradient matrices: the cofactor weight of a bilinear map is linear in one reference coordinate -- the same Kronecker
        // form; matching spaces take the variant that reproduces the reference's lower-triangle mirroring
        // Opt-in: the cofactor J_ab of a nearly axis-aligned quad is a small difference of O(|x|) coordinates, which the
        // reference and this kernel round differently (absolute 1e-16 each); the default path restates the reference's
        // operation order for J and agrees with it to 1e-12 of every matrix, this one to 1e-12 of the LARGEST of the four.
        // Small non-matching shapes stay on the sum-factorised kernels, which are faster there.
        if (!ctx->kron || !ctx->kron_gradient || (!m.symmetric && m.op.N_in < 64))
            return IB200_ERROR_UNSUPPORTED;
        return launch_mass2d_kron(ctx, cfg, m.key_kron + ":d" + std::to_string(m.deriv) + std::to_string(m.idx_out), 1, m.basis_out, 1,
                                  m.basis_in, m.rules, m.geom_basis, cnt, geom, out, m.deriv, m.idx_out, m.symmetric);
    }
    if (ctx->lo_kernel)
    {
        const int res = launch_mass2d_lo(ctx, cfg, m.key_fused, m.op, m.host, m.match0, m.match1, -1, 2, &m.ghost, -1, cnt, geom, out);
        if (res != IB200_ERROR_UNSUPPORTED)
            return res;
    }
    if (ctx->kron)
        return launch_mass2d_kron(ctx, cfg, m.key_kron, 1, m.basis_out, 1, m.basis_in, m.rules, m.geom_basis, cnt, geom, out);
    return IB200_ERROR_UNSUPPORTED;
}

// geom[cnt][n_phys][N_g] -> out[cnt][N_out][N_in] on device pointers; wscratch[cnt][Q] is used by the two-pass paths only
static int mapped_mass_launch(ib200_ctx *ctx, const LaunchCfg &cfg, const MappedMassOp &m, size_t cnt, const double *geom, double *out,
                              double *wscratch)
{
    if (!cnt)
        return INTERP_SUCCESS;
    if (!cfg.force_generic && ctx->kron && m.ndim == 3 && m.n_phys == 3 && m.deriv < 0)
    {
        // trilinear hexahedra: det J is a 27-term polynomial, the mass matrix a 27-term Kronecker sum (kernels_mass3d_kron.cu)
        const int res = launch_mass3d_kron(ctx, cfg, m.key_kron, m.basis_out, m.basis_in, m.rules, m.geom_basis, cnt, geom, out);
        if (res != IB200_ERROR_UNSUPPORTED)
            return res;
    }
    if (!cfg.force_generic && m.fuse_geometry)
    {
        int res = launch_mass2d_bilinear(ctx, cfg, m, cnt, geom, out);
        if (res == IB200_ERROR_UNSUPPORTED)
            res = launch_mass2d_best(ctx, cfg, m.key_fused, m.op, m.host, m.match0, m.match1, m.deriv, 2, &m.ghost, m.deriv, m.idx_out,
                                     cnt, geom, out);
        if (res != IB200_ERROR_UNSUPPORTED)
            return res;
    }
    IB_TRY(launch_geometry_generic(cfg, m.gop, m.n_phys, cnt, geom, nullptr, nullptr, nullptr, wscratch, m.deriv, m.idx_out));
    if (!cfg.force_generic)
    {
        const int res = launch_mass2d_best(ctx, cfg, m.key_w, m.op, m.host, m.match0, m.match1, m.deriv, 1, nullptr, -1, 0, cnt, wscratch,
                                           out);
        if (res != IB200_ERROR_UNSUPPORTED)
            return res;
    }
    return launch_mass_generic(cfg, m.op, m.symmetric, cnt, wscratch, m.deriv, out);
}

extern "C" int ib200_mass_matrix_mapped_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis_in,
                                              const basis_spec_t *basis_out, const integration_spec_t *rules,
                                              unsigned n_phys, const basis_spec_t *geom_basis, size_t n_elements,
                                              const double *geom_dofs, int idx_in, int idx_out, double *out, int where)
{
    IB_TRY(ctx_check(ctx));
    std::lock_guard<std::mutex> lock(ctx->mutex);
    NvtxRange nvtx("ib200_mass_matrix_mapped_batch");
    IB_TRY(check_geometry_dims(ndim, n_phys));
    if (!basis_in || !basis_out || !geom_basis)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "basis_in/basis_out/geom_basis must not be NULL");
    if (n_elements && (!out || !geom_dofs))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "geom_dofs/out must not be NULL");
    if (idx_in >= (int)ndim)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "Index %d out of bounds for input space with %u dimensions.", idx_in, ndim);
    if (idx_in >= 0 && (idx_out < 0 || idx_out >= (int)n_phys))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "Index %d out of bounds for output space with %u dimensions.", idx_out,
                          n_phys);
    MappedMassOp m;
    IB_TRY(mapped_mass_prepare(ctx, ndim, basis_in, basis_out, rules, n_phys, geom_basis, idx_in, idx_out, &m));
    std::vector<BatchArray> arrays = {{ARR_IN, (void *)geom_dofs, sizeof(double) * n_phys * m.gop.N_in},
                                      {ARR_OUT, out, sizeof(double) * m.op.N_out * m.op.N_in},
                                      {ARR_SCRATCH, nullptr, sizeof(double) * m.op.Q}}; // per-point weights of the two-pass paths
    return run_batched(ctx, n_elements, arrays, where, 0, [&](const LaunchCfg &cfg, size_t, size_t cnt, void **dev) -> int {
        return mapped_mass_launch(ctx, cfg, m, cnt, (const double *)dev[0], (double *)dev[1], (double *)dev[2]);
    });
}

// ---- N4 + K2: the whole k-form element mass matrix -----------------------------------------------------------------
// out[e] = [M_ab] over the component spaces of the k-form in component order, M_ab = compute_mass_matrix(space_b, space_a,
// smap) with the determinant weight (the blocks the reference pins).  Bilinear quads: ONE launch of the Kronecker
// kernel writes the [N][N] matrix; everything else: one mapped mass launch per block into scratch + a strided device copy.
extern "C" int ib200_kform_mass_mapped_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis, unsigned k,
                                             const integration_spec_t *rules, unsigned n_phys, const basis_spec_t *geom_basis,
                                             size_t n_elements, const double *geom_dofs, double *out, int where)
{
    IB_TRY(ctx_check(ctx));
    std::lock_guard<std::mutex> lock(ctx->mutex);
    NvtxRange nvtx("ib200_kform_mass_mapped_batch");
    IB_TRY(check_geometry_dims(ndim, n_phys));
    if (!basis || !rules || !geom_basis)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "basis/rules/geom_basis must not be NULL");
    if (k > ndim)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "a %u-form does not exist in %u dimensions", k, ndim);
    if (n_elements && (!out || !geom_dofs))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "geom_dofs/out must not be NULL");
    const unsigned n_comp = ib200_kform_component_count(ndim, k);
    std::vector<ib200_covector_t> comps(n_comp);
    IB_TRY(ib200_kform_components(ndim, k, comps.data()));
    std::vector<basis_spec_t> spaces((size_t)n_comp * ndim);
    std::vector<size_t> n_dof(n_comp), offset(n_comp + 1, 0);
    for (unsigned c = 0; c < n_comp; ++c)
    {
        IB_TRY(ib200_kform_component_space(ndim, basis, comps[c], spaces.data() + (size_t)c * ndim));
        n_dof[c] = 1;
        for (unsigned d = 0; d < ndim; ++d)
            n_dof[c] *= spaces[(size_t)c * ndim + d].order + 1;
        offset[c + 1] = offset[c] + n_dof[c];
    }
    const size_t N = offset[n_comp];
    OpDesc gop;
    OpHost ghost;
    IB_TRY(ctx_make_op(ctx, ndim, geom_basis, nullptr, rules, &gop, &ghost));
    size_t Q = 1;
    for (unsigned d = 0; d < ndim; ++d)
        Q *= rules[d].order + 1;

    // one launch: bilinear quads in the plane, at most two components (every 2-D k-form)
    if (ndim == 2 && n_phys == 2 && n_comp <= 2 && ctx->kron && !ctx->force_generic)
    {
        std::string key = "kformkron:" + std::to_string(k);
        key += plan_key("", ndim, basis, nullptr, rules, geom_basis, -1, 2, -1, 0);
        const LaunchCfg probe_cfg = make_launch_cfg(ctx, ctx->stream);
        const int probe = launch_mass2d_kron(ctx, probe_cfg, key, (int)n_comp, spaces.data(), (int)n_comp, spaces.data(), rules, geom_basis,
                                             0, nullptr, nullptr);
        if (probe == INTERP_SUCCESS)
        {
            std::vector<BatchArray> arrays = {{ARR_IN, (void *)geom_dofs, sizeof(double) * n_phys * gop.N_in},
                                              {ARR_OUT, out, sizeof(double) * N * N}};
            return run_batched(ctx, n_elements, arrays, where, 0, [&](const LaunchCfg &cfg, size_t, size_t cnt, void **dev) -> int {
                if (!cnt)
                    return INTERP_SUCCESS;
                return launch_mass2d_kron(ctx, cfg, key, (int)n_comp, spaces.data(), (int)n_comp, spaces.data(), rules, geom_basis, cnt,
                                          (const double *)dev[0], (double *)dev[1]);
            });
        }
        if (probe != IB200_ERROR_UNSUPPORTED)
            return probe;
    }

    // block by block
    std::vector<MappedMassOp> ops((size_t)n_comp * n_comp);
    size_t max_block = 0;
    for (unsigned a = 0; a < n_comp; ++a)
        for (unsigned b = 0; b < n_comp; ++b)
        {
            IB_TRY(mapped_mass_prepare(ctx, ndim, spaces.data() + (size_t)b * ndim, spaces.data() + (size_t)a * ndim, rules, n_phys,
                                       geom_basis, -1, 0, &ops[(size_t)a * n_comp + b]));
            max_block = std::max(max_block, n_dof[a] * n_dof[b]);
        }
    std::vector<BatchArray> arrays = {{ARR_IN, (void *)geom_dofs, sizeof(double) * n_phys * gop.N_in},
                                      {ARR_OUT, out, sizeof(double) * N * N},
                                      {ARR_SCRATCH, nullptr, sizeof(double) * max_block},
                                      {ARR_SCRATCH, nullptr, sizeof(double) * Q}};
    return run_batched(ctx, n_elements, arrays, where, 0, [&](const LaunchCfg &cfg, size_t, size_t cnt, void **dev) -> int {
        if (!cnt)
            return INTERP_SUCCESS;
        for (unsigned a = 0; a < n_comp; ++a)
            for (unsigned b = 0; b < n_comp; ++b)
            {
                double *block = n_comp == 1 ? (double *)dev[1] : (double *)dev[2];
                IB_TRY(mapped_mass_launch(ctx, cfg, ops[(size_t)a * n_comp + b], cnt, (const double *)dev[0], block, (double *)dev[3]));
                if (n_comp == 1)
                    continue;
                // block[e][ja][ib] -> out[e][offset_a + ja][offset_b + ib]
                cudaMemcpy3DParms cp{};
                cp.srcPtr = make_cudaPitchedPtr(block, n_dof[b] * sizeof(double), n_dof[b] * sizeof(double), n_dof[a]);
                cp.dstPtr = make_cudaPitchedPtr((double *)dev[1] + offset[a] * N + offset[b], N * sizeof(double), N * sizeof(double), N);
                cp.extent = make_cudaExtent(n_dof[b] * sizeof(double), n_dof[a], cnt);
                cp.kind = cudaMemcpyDeviceToDevice;
                IB_CUDA(cudaMemcpy3DAsync(&cp, cfg.stream));
            }
        return INTERP_SUCCESS;
    });
}

// ---- K3 ----------------------------------------------------------------------------------------------------
extern "C" int ib200_l2_dual_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis,
                                   const integration_spec_t *rules, size_t n_elements, const double *fvals,
                                   const double *det, double *out, int where)
{
    IB_TRY(ctx_check(ctx));
    std::lock_guard<std::mutex> lock(ctx->mutex);
    NvtxRange nvtx("ib200_l2_dual_batch");
    if (n_elements && (!fvals || !out))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "fvals/out must not be NULL");
    OpDesc op;
    OpHost host;
    IB_TRY(ctx_make_op(ctx, ndim, basis, nullptr, rules, &op, &host));
    const std::string key = plan_key("l2", ndim, basis, nullptr, rules, nullptr, det ? 1 : 0, 0, 0, 0);
    std::vector<BatchArray> arrays = {{ARR_IN, (void *)fvals, sizeof(double) * op.Q},
                                      {ARR_IN, (void *)det, sizeof(double) * op.Q},
                                      {ARR_OUT, out, sizeof(double) * op.N_in}};
    return run_batched(ctx, n_elements, arrays, where, 0,
                       [&](const LaunchCfg &cfg, size_t, size_t cnt, void **dev) -> int {
                           if (!cfg.force_generic && cnt)
                           {
                               const int res = launch_l2dual2d_fast(ctx, cfg, key, op, host, det ? 1 : 0, nullptr, cnt,
                                                                    (const double *)dev[0], (const double *)dev[1], (double *)dev[2]);
                               if (res != IB200_ERROR_UNSUPPORTED)
                                   return res;
                           }
                           return launch_l2_generic(cfg, op, cnt, (const double *)dev[0], (const double *)dev[1],
                                                    (double *)dev[2]);
                       });
}

extern "C" int ib200_l2_dual_mapped_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis,
                                          const integration_spec_t *rules, unsigned n_phys,
                                          const basis_spec_t *geom_basis, size_t n_elements, const double *geom_dofs,
                                          const double *fvals, double *out, int where)
{
    IB_TRY(ctx_check(ctx));
    std::lock_guard<std::mutex> lock(ctx->mutex);
    NvtxRange nvtx("ib200_l2_dual_mapped_batch");
    IB_TRY(check_geometry_dims(ndim, n_phys));
    if (n_elements && (!fvals || !out || !geom_dofs))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "geom_dofs/fvals/out must not be NULL");
    OpDesc op, gop;
    OpHost host, ghost;
    IB_TRY(ctx_make_op(ctx, ndim, basis, nullptr, rules, &op, &host));
    IB_TRY(ctx_make_op(ctx, ndim, geom_basis, nullptr, rules, &gop, &ghost));
    const bool fuse_geometry = ndim == 2 && n_phys == 2 && geom_basis[0].order == 1 && geom_basis[1].order == 1;
    const std::string key = plan_key("l2g", ndim, basis, nullptr, rules, geom_basis, 2, 0, 0, 0);
    const std::string key_w = plan_key("l2", ndim, basis, nullptr, rules, nullptr, 1, 0, 0, 0);
    std::vector<BatchArray> arrays = {{ARR_IN, (void *)geom_dofs, sizeof(double) * n_phys * gop.N_in},
                                      {ARR_IN, (void *)fvals, sizeof(double) * op.Q},
                                      {ARR_OUT, out, sizeof(double) * op.N_in},
                                      {ARR_SCRATCH, nullptr, sizeof(double) * op.Q}};
    return run_batched(ctx, n_elements, arrays, where, 0,
                       [&](const LaunchCfg &cfg, size_t, size_t cnt, void **dev) -> int {
                           if (!cfg.force_generic && cnt && fuse_geometry)
                           {
                               const int res = launch_l2dual2d_fast(ctx, cfg, key, op, host, 2, &ghost, cnt, (const double *)dev[1],
                                                                    (const double *)dev[0], (double *)dev[2]);
                               if (res != IB200_ERROR_UNSUPPORTED)
                                   return res;
                           }
                           IB_TRY(launch_geometry_generic(cfg, gop, n_phys, cnt, (const double *)dev[0], nullptr, nullptr,
                                                          nullptr, (double *)dev[3], -1, 0));
                           if (!cfg.force_generic && cnt)
                           {
                               const int res = launch_l2dual2d_fast(ctx, cfg, key_w, op, host, 1, nullptr, cnt, (const double *)dev[1],
                                                                    (const double *)dev[3], (double *)dev[2]);
                               if (res != IB200_ERROR_UNSUPPORTED)
                                   return res;
                           }
                           return launch_l2_generic(cfg, op, cnt, (const double *)dev[1], (const double *)dev[3],
                                                    (double *)dev[2]);
                       });
}

// ---- K2 + K3 fused ------------------------------------------------------------------------------------------
// (M, b) per element, and with `solve` the primal DoFs u = M^-1 b: M then lives in the context's scratch only (mass ==
// NULL) and `rhs` receives u
static int l2_system_impl(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis, const integration_spec_t *rules,
                          unsigned n_phys, const basis_spec_t *geom_basis, size_t n_elements, const double *geom_dofs,
                          const double *fvals, double *mass, double *rhs, int where, bool solve)
{
    IB_TRY(ctx_check(ctx));
    std::lock_guard<std::mutex> lock(ctx->mutex);
    NvtxRange nvtx(solve ? "ib200_l2_project_mapped_batch" : "ib200_l2_system_mapped_batch");
    IB_TRY(check_geometry_dims(ndim, n_phys));
    if (!basis || !geom_basis)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "basis/geom_basis must not be NULL");
    if (n_elements && (!fvals || (!mass && !solve) || !rhs || !geom_dofs))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "geom_dofs/fvals/mass/rhs must not be NULL");
    OpDesc op, gop;
    OpHost host, ghost;
    IB_TRY(ctx_make_op(ctx, ndim, basis, basis, rules, &op, &host));
    IB_TRY(ctx_make_op(ctx, ndim, geom_basis, nullptr, rules, &gop, &ghost));
    const bool fuse_geometry = ndim == 2 && n_phys == 2 && geom_basis[0].order == 1 && geom_basis[1].order == 1;
    const std::string key_fused = plan_key("massg", ndim, basis, basis, rules, geom_basis, -1, 2, -1, 0);
    const std::string key_kron = plan_key("kron", ndim, basis, basis, rules, geom_basis, -1, 2, -1, 0);
    const std::string key_w = plan_key("mass", ndim, basis, basis, rules, nullptr, -1, 1, -1, 0);
    const std::string key_l2 = plan_key("l2g", ndim, basis, nullptr, rules, geom_basis, 2, 0, 0, 0);
    const std::string key_l2w = plan_key("l2", ndim, basis, nullptr, rules, nullptr, 1, 0, 0, 0);
    if (ndim == 1 && n_phys == 1 && geom_basis[0].order == 1 && ctx->fastdiag && !ctx->force_generic)
    {
        // straight 1-D segments: the determinant is one number per element, M = det A, b = det Bw f, u = P f -- one launch
        const std::string key_1d = plan_key("l2seg", ndim, basis, nullptr, rules, geom_basis, 2, 0, 0, 0);
        const LaunchCfg probe_cfg = make_launch_cfg(ctx, ctx->stream);
        const int probe = launch_l2_system_1d(ctx, probe_cfg, key_1d, basis[0], rules[0], geom_basis[0], 0, nullptr, nullptr, nullptr,
                                              nullptr, nullptr);
        if (probe == INTERP_SUCCESS)
        {
            std::vector<BatchArray> seg_arrays = {{ARR_IN, (void *)geom_dofs, sizeof(double) * 2},
                                                  {ARR_IN, (void *)fvals, sizeof(double) * op.Q},
                                                  {ARR_OUT, solve ? nullptr : mass, sizeof(double) * op.N_out * op.N_in},
                                                  {ARR_OUT, rhs, sizeof(double) * op.N_in}};
            return run_batched(ctx, n_elements, seg_arrays, where, 0, [&](const LaunchCfg &cfg, size_t, size_t cnt, void **dev) -> int {
                if (!cnt)
                    return INTERP_SUCCESS;
                return launch_l2_system_1d(ctx, cfg, key_1d, basis[0], rules[0], geom_basis[0], cnt, (const double *)dev[0],
                                           (const double *)dev[1], solve ? nullptr : (double *)dev[2], solve ? nullptr : (double *)dev[3],
                                           solve ? (double *)dev[3] : nullptr);
            });
        }
        if (probe != IB200_ERROR_UNSUPPORTED)
            return probe;
    }
    if (solve && fuse_geometry && ctx->fastdiag && !ctx->force_generic)
    {
        // bilinear quads: det J is linear in the reference coordinates, so M^-1 is two 1-D transforms and a diagonal
        // scaling (kernels_fastdiag.cu) -- no mass matrix, no factorisation, no scratch
        const std::string key_fd = plan_key("l2fd", ndim, basis, nullptr, rules, geom_basis, 2, 0, 0, 0);
        const LaunchCfg probe_cfg = make_launch_cfg(ctx, ctx->stream);
        const int probe = launch_l2proj_fd2d(ctx, probe_cfg, key_fd, basis, rules, geom_basis, 0, nullptr, nullptr, nullptr);
        if (probe == INTERP_SUCCESS)
        {
            std::vector<BatchArray> fd_arrays = {{ARR_IN, (void *)geom_dofs, sizeof(double) * n_phys * gop.N_in},
                                                 {ARR_IN, (void *)fvals, sizeof(double) * op.Q},
                                                 {ARR_OUT, rhs, sizeof(double) * op.N_in}};
            return run_batched(ctx, n_elements, fd_arrays, where, 0, [&](const LaunchCfg &cfg, size_t, size_t cnt, void **dev) -> int {
                if (!cnt)
                    return INTERP_SUCCESS;
                return launch_l2proj_fd2d(ctx, cfg, key_fd, basis, rules, geom_basis, cnt, (const double *)dev[1],
                                          (const double *)dev[0], (double *)dev[2]);
            });
        }
        if (probe != IB200_ERROR_UNSUPPORTED)
            return probe;
    }
    std::vector<BatchArray> arrays = {{ARR_IN, (void *)geom_dofs, sizeof(double) * n_phys * gop.N_in},
                                      {ARR_IN, (void *)fvals, sizeof(double) * op.Q},
                                      {solve ? ARR_SCRATCH : ARR_OUT, solve ? nullptr : mass, sizeof(double) * op.N_out * op.N_in},
                                      {ARR_OUT, rhs, sizeof(double) * op.N_in},
                                      {ARR_SCRATCH, nullptr, sizeof(double) * op.Q}};
    const auto assemble = [&](const LaunchCfg &cfg, size_t cnt, const double *g, const double *f, double *M, double *b, double *w) -> int {
                           if (!cfg.force_generic && fuse_geometry)
                           {
                               if (ctx->lo_kernel)
                               {
                                   // one launch: geometry + mass matrix + right-hand side
                                   const int res = launch_mass2d_lo(ctx, cfg, key_fused, op, host, true, true, -1, 2, &ghost, -1, cnt,
                                                                    g, M, f, b);
                                   if (res != IB200_ERROR_UNSUPPORTED)
                                       return res;
                               }
                               int res = ctx->kron ? launch_mass2d_kron(ctx, cfg, key_kron, 1, basis, 1, basis, rules, geom_basis, cnt, g, M)
                                                   : (int)IB200_ERROR_UNSUPPORTED;
                               if (res == IB200_ERROR_UNSUPPORTED)
                                   res = launch_mass2d_best(ctx, cfg, key_fused, op, host, true, true, -1, 2, &ghost, -1, 0, cnt, g, M);
                               if (res == INTERP_SUCCESS)
                               {
                                   // M is done; the right-hand side by the same chain as ib200_l2_dual_mapped_batch
                                   res = launch_l2dual2d_fast(ctx, cfg, key_l2, op, host, 2, &ghost, cnt, f, g, b);
                                   if (res != IB200_ERROR_UNSUPPORTED)
                                       return res;
                                   IB_TRY(launch_geometry_generic(cfg, gop, n_phys, cnt, g, nullptr, nullptr, nullptr, w, -1, 0));
                                   res = launch_l2dual2d_fast(ctx, cfg, key_l2w, op, host, 1, nullptr, cnt, f, w, b);
                                   if (res == IB200_ERROR_UNSUPPORTED)
                                       res = launch_l2_generic(cfg, op, cnt, f, w, b);
                                   return res;
                               }
                               if (res != IB200_ERROR_UNSUPPORTED)
                                   return res;
                           }
                           IB_TRY(launch_geometry_generic(cfg, gop, n_phys, cnt, g, nullptr, nullptr, nullptr, w, -1, 0));
                           int res = IB200_ERROR_UNSUPPORTED;
                           if (!cfg.force_generic && ctx->kron && ndim == 3 && n_phys == 3) // trilinear hexahedra: the same kernel as
                               res = launch_mass3d_kron(ctx, cfg, key_kron, basis, basis, rules, geom_basis, cnt, g, M); // the separate call
                           if (res == IB200_ERROR_UNSUPPORTED && !cfg.force_generic)
                               res = launch_mass2d_best(ctx, cfg, key_w, op, host, true, true, -1, 1, nullptr, -1, 0, cnt, w, M);
                           if (res == IB200_ERROR_UNSUPPORTED)
                               res = launch_mass_generic(cfg, op, true, cnt, w, -1, M);
                           IB_TRY(res);
                           res = IB200_ERROR_UNSUPPORTED;
                           if (!cfg.force_generic)
                               res = launch_l2dual2d_fast(ctx, cfg, key_l2w, op, host, 1, nullptr, cnt, f, w, b);
                           if (res == IB200_ERROR_UNSUPPORTED)
                               res = launch_l2_generic(cfg, op, cnt, f, w, b);
                           return res;
    };
    return run_batched(ctx, n_elements, arrays, where, 0, [&](const LaunchCfg &cfg, size_t, size_t cnt, void **dev) -> int {
        if (!cnt)
            return INTERP_SUCCESS;
        double *M = (double *)dev[2], *b = (double *)dev[3];
        IB_TRY(assemble(cfg, cnt, (const double *)dev[0], (const double *)dev[1], M, b, (double *)dev[4]));
        return solve ? launch_spd_solve(cfg, op.N_in, cnt, M, b) : INTERP_SUCCESS;
    });
}

extern "C" int ib200_l2_system_mapped_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis,
                                            const integration_spec_t *rules, unsigned n_phys,
                                            const basis_spec_t *geom_basis, size_t n_elements, const double *geom_dofs,
                                            const double *fvals, double *mass, double *rhs, int where)
{
    return l2_system_impl(ctx, ndim, basis, rules, n_phys, geom_basis, n_elements, geom_dofs, fvals, mass, rhs, where, false);
}

extern "C" int ib200_l2_project_mapped_batch(ib200_ctx_t *ctx, unsigned ndim, const basis_spec_t *basis,
                                             const integration_spec_t *rules, unsigned n_phys,
                                             const basis_spec_t *geom_basis, size_t n_elements, const double *geom_dofs,
                                             const double *fvals, double *primal, int where)
{
    return l2_system_impl(ctx, ndim, basis, rules, n_phys, geom_basis, n_elements, geom_dofs, fvals, nullptr, primal, where, true);
}

// ---- N1 ----------------------------------------------------------------------------------------------------
extern "C" int ib200_spd_solve_batch(ib200_ctx_t *ctx, unsigned n, size_t n_elements, double *mass, double *rhs,
                                     int where)
{
    IB_TRY(ctx_check(ctx));
    std::lock_guard<std::mutex> lock(ctx->mutex);
    NvtxRange nvtx("ib200_spd_solve_batch");
    if (n_elements && (!mass || !rhs))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "mass/rhs must not be NULL");
    if (n == 0 || n > 4096)
        return ib200_fail(IB200_ERROR_UNSUPPORTED, "system size %u outside 1..4096", n);
    // host matrices are not copied back: a host caller's `mass` is left unchanged
    std::vector<BatchArray> arrays = {{ARR_IN, mass, sizeof(double) * n * n}, {ARR_INOUT, rhs, sizeof(double) * n}};
    return run_batched(ctx, n_elements, arrays, where, 0,
                       [&](const LaunchCfg &cfg, size_t, size_t cnt, void **dev) -> int {
                           return launch_spd_solve(cfg, n, cnt, (double *)dev[0], (double *)dev[1]);
                       });
}

// ---- projection norms -------------------------------------------------------------------------------------------
extern "C" int ib200_dot(ib200_ctx_t *ctx, size_t n, const double *a, const double *b, double *out, int where)
{
    IB_TRY(ctx_check(ctx));
    std::lock_guard<std::mutex> lock(ctx->mutex);
    NvtxRange nvtx("ib200_dot");
    if (!out || (n && (!a || !b)))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "a/b/out must not be NULL");
    if (where != IB200_DEVICE && where != IB200_HOST)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "`where` must be IB200_HOST or IB200_DEVICE");
    LaunchCfg cfg = make_launch_cfg(ctx, ctx->stream);
    if (!ctx->dot_partial)
        IB_CUDA(cudaMalloc((void **)&ctx->dot_partial, sizeof(double) * (dot_scratch_doubles() + 1)));
    double *partial = ctx->dot_partial;
    double *result = partial + dot_scratch_doubles();
    if (where == IB200_DEVICE)
        return launch_dot(cfg, n, a, b, partial, out);
    // host vectors: stage through device scratch (norm vectors are small next to the matrices)
    void *stage = nullptr;
    IB_TRY(ctx_scratch(ctx, 2 * n * sizeof(double) + 256, &stage));
    double *da = (double *)stage, *db = da + n;
    IB_CUDA(cudaMemcpyAsync(da, a, n * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    IB_CUDA(cudaMemcpyAsync(db, b, n * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    IB_TRY(launch_dot(cfg, n, da, db, partial, result));
    IB_CUDA(cudaMemcpyAsync(out, result, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    IB_CUDA(cudaStreamSynchronize(ctx->stream));
    return INTERP_SUCCESS;
}
