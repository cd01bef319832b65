// Context, error plumbing and device-side table descriptors shared by the .cu files.
#pragma once

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <mutex>
#include <tuple>
#include <vector>

#include "../../include/interp_b200.h"
#include "tables.h"

// ---- error plumbing -----------------------------------------------------------------------------
int ib200_fail(int code, const char *fmt, ...);

#define IB_CUDA(call)                                                                                     \
    do                                                                                                    \
    {                                                                                                     \
        const cudaError_t ib_err__ = (call);                                                              \
        if (ib_err__ != cudaSuccess)                                                                      \
            return ib200_fail(IB200_ERROR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(ib_err__), \
                              __FILE__, __LINE__);                                                        \
    } while (0)

#define IB_TRY(call)                      \
    do                                    \
    {                                     \
        const int ib_res__ = (call);      \
        if (ib_res__ != INTERP_SUCCESS)   \
            return ib_res__;              \
    } while (0)

// ---- device-side descriptors ----------------------------------------------------------------------
// One tensor-product dimension of an operation.  Tables are basis-major [n][m] like the reference's
// basis_set_t payload (src/basis/basis_set.h:28-55).
struct DimTab
{
    int n_in;          // input-space basis count along this dimension
    int n_out;         // output-space basis count (== n_in for single-space operations)
    int m;             // quadrature points along this dimension
    const double *Bin; // [n_in][m]
    const double *Din; // [n_in][m]
    const double *Bout; // [n_out][m]
    const double *w;   // [m] rule weights
};

struct OpDesc
{
    int ndim;
    DimTab dim[IB200_MAX_DIM];
    int N_in, N_out, Q;
};

// Host copies of the same tables (the specialised kernels pass small tables through the kernel-parameter
// constant bank).
struct OpHost
{
    const double *Bin[IB200_MAX_DIM];
    const double *Din[IB200_MAX_DIM];
    const double *Bout[IB200_MAX_DIM];
    const double *Dout[IB200_MAX_DIM];
    const double *w[IB200_MAX_DIM];
};

// Arbitrary-point evaluation (kernels_points.cu): per dimension the basis family, its size and, for the Lagrange family,
// device pointers to the nodes and to the denominators prod_{k != j} (r_j - r_k).
#define IB200_POINT_MAX_DIM 3
struct PointDim
{
    int type; // basis_set_type_t
    int n;
    const double *roots;
    const double *den;
};
struct PointOp
{
    int ndim;
    PointDim dim[IB200_POINT_MAX_DIM];
    int N;
};

// ---- host-side context ----------------------------------------------------------------------------
struct DeviceTable
{
    double *d_values = nullptr; // [n][m] then derivatives [n][m]
    std::vector<double> h_values; // host copy (values then derivatives)
    int n = 0, m = 0;
};
struct DeviceRule
{
    double *d_nodes = nullptr; // [m] nodes then [m] weights
    std::vector<double> h_nodes;
    int m = 0;
};

struct ib200_ctx
{
    int device = 0;
    cudaDeviceProp prop{};
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr; // stream work is queued on (own_stream unless adopted)
    cudaStream_t io_stream[3] = {nullptr, nullptr, nullptr};
    cudaEvent_t io_event[3] = {nullptr, nullptr, nullptr};
    cudaEvent_t timer[8][2] = {};
    std::mutex mutex;
    std::map<std::tuple<int, unsigned, int, unsigned>, DeviceTable> tables;
    std::map<std::tuple<int, unsigned>, DeviceRule> rules;
    std::map<std::string, std::shared_ptr<void>> plans; // per-operation kernel plans (kernels_mass2d.cu)
    std::map<std::tuple<int, unsigned>, double *> point_tabs; // Lagrange roots[n] | denominators[n] per (type, order)
    std::map<std::tuple<int, unsigned>, double *> inc_tabs;   // Lagrange incidence matrices [n][n+1] per (type, order)
    unsigned long long launches = 0;
    bool force_generic = false;
    int mma_min_points = 13; // quadrature points per dimension from which the DMMA mass kernel is preferred
    bool lo_kernel = true;   // compile-time specialised low-order mass kernel (IB200_NO_LO_KERNEL=1 disables it)
    bool kron = true;        // Kronecker-sum mass kernel for bilinear quads (IB200_NO_KRON=1: sum-factorised kernels only)
    bool kron_gradient = false; // ... also for gradient matrices (IB200_KRON_GRADIENT=1; see ib200_ctx_set_kron_gradient)
    bool fastdiag = true;    // primal projections of bilinear quads by fast diagonalisation (IB200_NO_FASTDIAG=1: factorise M)
    size_t chunk_elements = 0;
    // staging (host-pointer calls) and scratch (fused generic path)
    void *stage_dev[3] = {nullptr, nullptr, nullptr};
    size_t stage_dev_bytes[3] = {0, 0, 0};
    void *stage_pin[3] = {nullptr, nullptr, nullptr};
    size_t stage_pin_bytes[3] = {0, 0, 0};
    void *scratch = nullptr;
    size_t scratch_bytes = 0;
    double *dot_partial = nullptr; // partial sums + result slot of ib200_dot
    // dynamic tile scheduling: ring of {next tile, finished CTAs} counter pairs, zero between launches (the last CTA
    // of a launch resets its pair); every LaunchCfg takes the next pair so launches in flight on different streams
    // never share one
    unsigned long long *sched_ring = nullptr;
    unsigned sched_next = 0;
    // status word the kernels raise (bit 0: a batched solve met a system it could not factorise): pinned host memory
    // mapped into the device, read by the host after a synchronisation without a copy
    int *status_host = nullptr;
    int *status_dev = nullptr;
    static constexpr unsigned kSchedRing = 256;
    unsigned long long *next_sched() { return sched_ring + 2 * (sched_next++ % kSchedRing); }
};

int ctx_get_table(ib200_ctx *ctx, basis_spec_t b, integration_spec_t r, const DeviceTable **out);
int ctx_get_rule(ib200_ctx *ctx, integration_spec_t r, const DeviceRule **out);
int ctx_scratch(ib200_ctx *ctx, size_t bytes, void **out);

// Build the device descriptor of an operation (uploads/caches tables as needed).
int ctx_make_op(ib200_ctx *ctx, unsigned ndim, const basis_spec_t *basis_in, const basis_spec_t *basis_out,
                const integration_spec_t *rules, OpDesc *op, OpHost *host);

// ---- kernel launchers (device pointers, queued on `stream`) -----------------------------------------
struct LaunchCfg
{
    cudaStream_t stream;
    int sm_count;
    bool force_generic;
    unsigned long long *launch_counter;
    unsigned long long *sched; // device counter pair {next tile, finished CTAs} for dynamically scheduled kernels
    int *status;               // device view of the context's status word (IB200_STATUS_* bits)
};
enum
{
    IB200_STATUS_BAD_SYSTEM = 1
};
inline LaunchCfg make_launch_cfg(ib200_ctx *ctx, cudaStream_t stream)
{
    return LaunchCfg{stream, ctx->prop.multiProcessorCount, ctx->force_generic, &ctx->launches, ctx->next_sched(), ctx->status_dev};
}
// Reads and clears the status word (call after the producing stream has been synchronised).
int ctx_take_status(ib200_ctx *ctx);

// NVTX range around every batched C-ABI call (visible in Nsight Systems / ncu --nvtx; a no-op without a tool attached).
struct NvtxRange
{
    explicit NvtxRange(const char *name);
    ~NvtxRange();
};

int launch_reconstruct_generic(const LaunchCfg &cfg, const OpDesc &op, size_t E, const double *dofs, unsigned n_fields,
                               const unsigned *masks, double *out);
int launch_geometry_generic(const LaunchCfg &cfg, const OpDesc &op, unsigned n_phys, size_t E, const double *geom_dofs,
                            double *coords, double *det, double *jinv, double *weight_out, int w_idx_in, int w_idx_out);
int launch_mass_generic(const LaunchCfg &cfg, const OpDesc &op, bool symmetric, size_t E, const double *weight,
                        int deriv_dim_in, double *out);
int launch_l2_generic(const LaunchCfg &cfg, const OpDesc &op, size_t E, const double *fvals, const double *det, double *out);
int launch_spd_solve(const LaunchCfg &cfg, unsigned n, size_t E, double *mass, double *rhs);
int launch_points_evaluate(const LaunchCfg &cfg, const PointOp &op, size_t E, size_t P, const double *x, double *out);
int launch_points_reconstruct(const LaunchCfg &cfg, const PointOp &op, size_t E, size_t P, const double *dofs, const double *x,
                              unsigned n_fields, const unsigned *masks, double *out);
int launch_dof_derivative(const LaunchCfg &cfg, int type, int n, size_t pre, size_t post, const double *d_inc, size_t E,
                          const double *in, double *out);
int launch_mirror_lower(const LaunchCfg &cfg, size_t E, int N, double *out);
int dot_scratch_doubles();
int launch_dot(const LaunchCfg &cfg, size_t n, const double *a, const double *b, double *partial, double *out);

// Specialised (sum-factorised, templated) 2-D paths: return IB200_ERROR_UNSUPPORTED (without setting the
// error text) when the shape is outside their instantiation list; the caller then uses the generic kernels.
int launch_eval2d_fast(const LaunchCfg &cfg, const OpDesc &op, const OpHost &host, size_t E, const double *dofs,
                       unsigned n_fields, const unsigned *masks, double *out);
int launch_mass2d_fast(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, const OpDesc &op, const OpHost &host,
                       bool match0, bool match1, int deriv_dim_in, int wsrc_kind, const OpHost *geom_host, int widx_in,
                       int widx_out, size_t E, const double *wsrc, double *out);
int launch_mass2d_mma(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, const OpDesc &op, const OpHost &host,
                      bool match0, bool match1, int deriv_dim_in, int wsrc_kind, const OpHost *geom_host, int widx_in,
                      int widx_out, size_t E, const double *wsrc, double *out);
int launch_mass2d_lo(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, const OpDesc &op, const OpHost &host,
                     bool match0, bool match1, int deriv_dim_in, int wsrc_kind, const OpHost *geom_host, int widx_in, size_t E,
                     const double *wsrc, double *out, const double *fvals = nullptr, double *rhs = nullptr);
int launch_l2dual2d_fast(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, const OpDesc &op, const OpHost &host,
                         int wsrc, const OpHost *geom_host, size_t E, const double *fvals, const double *aux, double *out);
int launch_mass3d_kron(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, const basis_spec_t *rows, const basis_spec_t *cols,
                       const integration_spec_t *rules, const basis_spec_t *geom_basis, size_t E, const double *geom, double *out);
int launch_l2_system_1d(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, basis_spec_t basis, integration_spec_t rule,
                        basis_spec_t geom_basis, size_t E, const double *geom, const double *fvals, double *mass, double *rhs,
                        double *primal);
int launch_mass2d_kron(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, int n_row_groups, const basis_spec_t *rows,
                       int n_col_groups, const basis_spec_t *cols, const integration_spec_t *rules, const basis_spec_t *geom_basis,
                       size_t E, const double *geom, double *out, int deriv_in = -1, int idx_out = 0, bool mirror = false);
int launch_l2proj_fd2d(ib200_ctx *ctx, const LaunchCfg &cfg, const std::string &key, const basis_spec_t *basis,
                       const integration_spec_t *rules, const basis_spec_t *geom_basis, size_t E, const double *fvals,
                       const double *geom, double *out);

// ---- dynamic tile scheduling (device side) ----------------------------------------------------------------------
// sched[0] = next unclaimed work item, sched[1] = participants that have finished.  SMs (and, for warp-autonomous
// kernels, SM sub-partitions hosting different numbers of warps) do not drain an HBM-bound stream at the same rate;
// with a static round-robin the fast ones idle at the end of the launch.
#ifdef __CUDACC__
// Closes a dynamically scheduled launch: the last participant (CTA or warp, `total` of them) to get here zeroes the
// pair for the next launch.  Every participant has made its last claim before it arrives, so nobody reads sched[0]
// after the reset.  Call from ONE thread per participant.
__device__ __forceinline__ void sched_finish_one(unsigned long long *sched, unsigned long long total)
{
    __threadfence();
    if (atomicAdd(sched + 1, 1ull) == total - 1)
    {
        sched[0] = 0;
        sched[1] = 0;
        __threadfence();
    }
}
__device__ __forceinline__ void sched_finish(unsigned long long *sched)
{
    if (threadIdx.x == 0)
        sched_finish_one(sched, gridDim.x);
}
#endif
