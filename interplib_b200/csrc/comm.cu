// the two collectives of the element-partitioned path behind the C ABI -- all-gather of per-rank DoF /
// right-hand-side slabs and all-reduce(sum) of projection norms -- over NCCL (NVLink 5 / NVSwitch on one box).
// The reference has no communication layer (single process, /root/reference has no MPI/NCCL call site); this file is
// new.  NCCL is bound at run time with dlopen so that libinterp_b200.so loads on hosts without it and shares the
// libnccl.so.2 a host framework (torch) may already have mapped.
#include <dlfcn.h>
#include <nccl.h> // types and enums only: every function is resolved with dlsym

#include <cstdint>
#include <cstdlib>
#include <mutex>
#include <vector>

#include "ctx.cuh"

namespace
{

struct NcclApi
{
    void *handle = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*CommAbort)(ncclComm_t) = nullptr;
    ncclResult_t (*CommGetAsyncError)(ncclComm_t, ncclResult_t *) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    const char *(*GetLastError)(ncclComm_t) = nullptr;
};

NcclApi g_nccl;
std::once_flag g_nccl_once;
std::string g_nccl_load_error;

void load_nccl()
{
    const char *env = std::getenv("INTERP_B200_NCCL");
    const char *names[] = {env, "libnccl.so.2", "libnccl.so"};
    for (const char *name : names)
    {
        if (!name || !*name)
            continue;
        g_nccl.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
        if (g_nccl.handle)
            break;
        g_nccl_load_error = dlerror();
    }
    if (!g_nccl.handle)
        return;
    bool ok = true;
    auto sym = [&](const char *name, bool required = true) -> void * {
        void *p = dlsym(g_nccl.handle, name);
        if (!p && required)
        {
            ok = false;
            g_nccl_load_error = std::string("symbol ") + name + " missing from libnccl";
        }
        return p;
    };
#define IB_SYM(field, name) g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(sym(name))
    IB_SYM(GetVersion, "ncclGetVersion");
    IB_SYM(GetUniqueId, "ncclGetUniqueId");
    IB_SYM(CommInitRank, "ncclCommInitRank");
    IB_SYM(CommDestroy, "ncclCommDestroy");
    IB_SYM(CommAbort, "ncclCommAbort");
    IB_SYM(CommGetAsyncError, "ncclCommGetAsyncError");
    IB_SYM(AllGather, "ncclAllGather");
    IB_SYM(AllReduce, "ncclAllReduce");
    IB_SYM(Broadcast, "ncclBroadcast");
    IB_SYM(GroupStart, "ncclGroupStart");
    IB_SYM(GroupEnd, "ncclGroupEnd");
    IB_SYM(GetErrorString, "ncclGetErrorString");
#undef IB_SYM
    g_nccl.GetLastError = reinterpret_cast<decltype(g_nccl.GetLastError)>(sym("ncclGetLastError", false));
    if (!ok)
    {
        dlclose(g_nccl.handle);
        g_nccl.handle = nullptr;
    }
}

int need_nccl()
{
    std::call_once(g_nccl_once, load_nccl);
    if (!g_nccl.handle)
        return ib200_fail(IB200_ERROR_NCCL, "libnccl.so.2 could not be loaded (%s); set $INTERP_B200_NCCL to its path",
                          g_nccl_load_error.c_str());
    return INTERP_SUCCESS;
}

} // namespace

struct ib200_comm
{
    ib200_ctx *ctx = nullptr;
    ncclComm_t nccl = nullptr;
    int n_ranks = 1, rank = 0;
    cudaStream_t side = nullptr; // stream of overlapped collectives
    cudaEvent_t fork = nullptr, join = nullptr;
    bool side_busy = false;
    // registered gather destination (ib200_comm_register_gather): the peers' buffers mapped through CUDA IPC, one push
    // stream + event per peer, a flag for the arrival barrier
    double *reg_out = nullptr;
    size_t reg_bytes = 0;
    std::vector<double *> peer;
    std::vector<cudaStream_t> push;
    std::vector<cudaEvent_t> pushed;
    double *flag = nullptr;
    bool p2p_pending_side = false;
};

namespace
{

int nccl_fail(ib200_comm *comm, ncclResult_t res, const char *what)
{
    const char *detail = (comm && comm->nccl && g_nccl.GetLastError) ? g_nccl.GetLastError(comm->nccl) : "";
    return ib200_fail(IB200_ERROR_NCCL, "%s failed: %s%s%s", what, g_nccl.GetErrorString(res), (detail && *detail) ? " -- " : "",
                      detail ? detail : "");
}

#define IB_NCCL(comm, call)                             \
    do                                                  \
    {                                                   \
        const ncclResult_t ib_nres__ = (call);          \
        if (ib_nres__ != ncclSuccess)                   \
            return nccl_fail(comm, ib_nres__, #call);   \
    } while (0)

int comm_check(ib200_comm *comm)
{
    if (!comm || !comm->ctx)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "communicator is NULL");
    IB_CUDA(cudaSetDevice(comm->ctx->device));
    return INTERP_SUCCESS;
}

} // namespace

extern "C" int ib200_partition_range(size_t n_elements, int n_ranks, int rank, size_t *first, size_t *count)
{
    if (n_ranks < 1 || rank < 0 || rank >= n_ranks || !first || !count)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "rank %d outside a world of %d (or NULL output)", rank, n_ranks);
    const size_t per = (n_elements + (size_t)n_ranks - 1) / (size_t)n_ranks;
    const size_t f = std::min(n_elements, (size_t)rank * per);
    *first = f;
    *count = std::min(n_elements, f + per) - f;
    return INTERP_SUCCESS;
}

extern "C" int ib200_comm_unique_id(void *id)
{
    if (!id)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "id must not be NULL");
    IB_TRY(need_nccl());
    static_assert(sizeof(ncclUniqueId) == IB200_COMM_ID_BYTES, "ncclUniqueId size");
    ncclUniqueId uid;
    IB_NCCL(nullptr, g_nccl.GetUniqueId(&uid));
    std::memcpy(id, &uid, sizeof uid);
    return INTERP_SUCCESS;
}

extern "C" int ib200_comm_create(ib200_ctx_t *ctx, int n_ranks, int rank, const void *id, ib200_comm_t **out)
{
    if (!ctx)
        return ib200_fail(IB200_ERROR_NO_DEVICE, "context is NULL (no CUDA device?)");
    if (!out || !id || n_ranks < 1 || rank < 0 || rank >= n_ranks)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "bad communicator arguments (rank %d of %d)", rank, n_ranks);
    *out = nullptr;
    IB_TRY(need_nccl());
    IB_CUDA(cudaSetDevice(ctx->device));
    ib200_comm *comm = new (std::nothrow) ib200_comm;
    if (!comm)
        return ib200_fail(INTERP_ERROR_FAILED_ALLOCATION, "out of host memory");
    comm->ctx = ctx;
    comm->n_ranks = n_ranks;
    comm->rank = rank;
    ncclUniqueId uid;
    std::memcpy(&uid, id, sizeof uid);
    const ncclResult_t res = g_nccl.CommInitRank(&comm->nccl, n_ranks, uid, rank);
    if (res != ncclSuccess)
    {
        delete comm;
        return nccl_fail(nullptr, res, "ncclCommInitRank");
    }
    // the side stream outranks the compute stream: the persistent compute kernels fill every SM, and a collective queued
    // behind them would otherwise only get CTAs once a whole compute launch has drained
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    if (cudaStreamCreateWithPriority(&comm->side, cudaStreamNonBlocking, prio_hi) != cudaSuccess ||
        cudaEventCreateWithFlags(&comm->fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&comm->join, cudaEventDisableTiming) != cudaSuccess)
    {
        ib200_comm_destroy(comm);
        return ib200_fail(IB200_ERROR_CUDA, "could not create the communicator's stream / events");
    }
    *out = comm;
    return INTERP_SUCCESS;
}

extern "C" int ib200_comm_destroy(ib200_comm_t *comm)
{
    if (!comm)
        return INTERP_SUCCESS;
    if (comm->ctx)
        cudaSetDevice(comm->ctx->device);
    ib200_comm_unregister_gather(comm);
    if (comm->flag)
        cudaFree(comm->flag);
    if (comm->side)
    {
        cudaStreamSynchronize(comm->side);
        cudaStreamDestroy(comm->side);
    }
    if (comm->fork)
        cudaEventDestroy(comm->fork);
    if (comm->join)
        cudaEventDestroy(comm->join);
    if (comm->nccl && g_nccl.handle)
    {
        if (comm->ctx)
            cudaStreamSynchronize(comm->ctx->stream);
        g_nccl.CommDestroy(comm->nccl);
    }
    delete comm;
    return INTERP_SUCCESS;
}

extern "C" int ib200_comm_info(ib200_comm_t *comm, int *n_ranks, int *rank, int *nccl_version)
{
    IB_TRY(comm_check(comm));
    if (n_ranks)
        *n_ranks = comm->n_ranks;
    if (rank)
        *rank = comm->rank;
    if (nccl_version)
        IB_NCCL(comm, g_nccl.GetVersion(nccl_version));
    return INTERP_SUCCESS;
}

extern "C" int ib200_comm_check(ib200_comm_t *comm)
{
    IB_TRY(comm_check(comm));
    ncclResult_t async = ncclSuccess;
    IB_NCCL(comm, g_nccl.CommGetAsyncError(comm->nccl, &async));
    if (async != ncclSuccess && async != ncclInProgress)
        return nccl_fail(comm, async, "asynchronous NCCL operation");
    return INTERP_SUCCESS;
}

extern "C" int ib200_allgatherv(ib200_comm_t *comm, const double *send, double *out, const size_t *counts,
                                const size_t *offsets, int on_side_stream)
{
    IB_TRY(comm_check(comm));
    if (!counts || !offsets)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "counts/offsets must not be NULL");
    size_t total = 0;
    bool even = true, packed = true;
    for (int r = 0; r < comm->n_ranks; ++r)
    {
        total += counts[r];
        even = even && counts[r] == counts[0];
        packed = packed && offsets[r] == (size_t)r * counts[0];
    }
    if (total == 0)
        return INTERP_SUCCESS;
    if (!out || (counts[comm->rank] && !send))
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "send/out must not be NULL");
    ib200_ctx *ctx = comm->ctx;
    std::lock_guard<std::mutex> lock(ctx->mutex);
    cudaStream_t st = ctx->stream;
    const bool p2p = comm->reg_out && out == comm->reg_out && comm->n_ranks > 1;
    if (p2p && (offsets[comm->rank] + counts[comm->rank]) * sizeof(double) > comm->reg_bytes)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "slab exceeds the registered gather buffer");
    if (on_side_stream)
    {
        IB_CUDA(cudaEventRecord(comm->fork, ctx->stream));
        IB_CUDA(cudaStreamWaitEvent(comm->side, comm->fork, 0));
        st = comm->side;
        comm->side_busy = true;
    }
    if (p2p)
    {
        // registered destination: push the local slab straight into every peer's buffer, one copy-engine transfer per peer
        // on its own stream (NVLink / NVSwitch), no SM and no staging involved; arrival is published by the barrier of
        // ib200_comm_join (side stream) or right here (compute stream)
        const size_t bytes = counts[comm->rank] * sizeof(double);
        // entry handshake: a one-double all-reduce on the calling stream before any push.  A peer's push into this
        // rank's `out` can then only start once THIS rank's stream has reached the call too, i.e. after everything this
        // rank queued earlier on that stream (kernels still reading the previous gather, its own writes into `out`) --
        // the stream-ordered entry an NCCL all-gather gives.  The pushes wait for it through `fork`, re-recorded here
        // (the side stream's wait on the earlier record is already queued and keeps that earlier state).
        IB_NCCL(comm, g_nccl.AllReduce(comm->flag, comm->flag, 1, ncclDouble, ncclSum, comm->nccl, st));
        IB_CUDA(cudaEventRecord(comm->fork, st));
        // peers in ring order starting after this rank: at any moment the copies in flight form a permutation, no GPU is
        // the target of all the others at once
        for (int k = 1; k < comm->n_ranks && bytes; ++k)
        {
            const int r = (comm->rank + k) % comm->n_ranks;
            IB_CUDA(cudaStreamWaitEvent(comm->push[r], comm->fork, 0));
            IB_CUDA(cudaMemcpyAsync(comm->peer[r] + offsets[comm->rank], send, bytes, cudaMemcpyDefault, comm->push[r]));
            IB_CUDA(cudaEventRecord(comm->pushed[r], comm->push[r]));
            IB_CUDA(cudaStreamWaitEvent(st, comm->pushed[r], 0));
        }
        if (bytes && send != out + offsets[comm->rank])
            IB_CUDA(cudaMemcpyAsync(out + offsets[comm->rank], send, bytes, cudaMemcpyDeviceToDevice, st));
        if (on_side_stream)
            comm->p2p_pending_side = true;
        else
            IB_NCCL(comm, g_nccl.AllReduce(comm->flag, comm->flag, 1, ncclDouble, ncclSum, comm->nccl, st));
        return INTERP_SUCCESS;
    }
    if (even && packed)
    {
        // equal, back-to-back slabs: the library's all-gather (in place when `send` already sits at its offset)
        IB_NCCL(comm, g_nccl.AllGather(send, out, counts[0], ncclDouble, comm->nccl, st));
        return INTERP_SUCCESS;
    }
    // ragged or strided destinations: one broadcast per rank, fused into one NCCL group (a single launch)
    IB_NCCL(comm, g_nccl.GroupStart());
    ncclResult_t res = ncclSuccess;
    for (int r = 0; r < comm->n_ranks && res == ncclSuccess; ++r)
        if (counts[r])
            res = g_nccl.Broadcast(r == comm->rank ? (const void *)send : (const void *)(out + offsets[r]), out + offsets[r],
                                   counts[r], ncclDouble, r, comm->nccl, st);
    const ncclResult_t end = g_nccl.GroupEnd();
    if (res != ncclSuccess)
        return nccl_fail(comm, res, "ncclBroadcast");
    IB_NCCL(comm, end);
    return INTERP_SUCCESS;
}

extern "C" int ib200_allgather_slabs(ib200_comm_t *comm, const double *local, size_t n_elements, size_t elem_doubles,
                                     double *out)
{
    IB_TRY(comm_check(comm));
    std::vector<size_t> counts(comm->n_ranks), offsets(comm->n_ranks);
    for (int r = 0; r < comm->n_ranks; ++r)
    {
        size_t first, cnt;
        IB_TRY(ib200_partition_range(n_elements, comm->n_ranks, r, &first, &cnt));
        counts[r] = cnt * elem_doubles;
        offsets[r] = first * elem_doubles;
    }
    return ib200_allgatherv(comm, local, out, counts.data(), offsets.data(), 0);
}

namespace
{
// push streams / events and IPC mappings of a (possibly half-finished) registration
void release_gather_mappings(ib200_comm *comm)
{
    for (size_t r = 0; r < comm->push.size(); ++r)
    {
        if (comm->push[r])
        {
            cudaStreamSynchronize(comm->push[r]);
            cudaStreamDestroy(comm->push[r]);
        }
        if (r < comm->pushed.size() && comm->pushed[r])
            cudaEventDestroy(comm->pushed[r]);
    }
    comm->push.clear();
    comm->pushed.clear();
    for (size_t r = 0; r < comm->peer.size(); ++r)
        if (comm->peer[r])
            cudaIpcCloseMemHandle(comm->peer[r]);
    comm->peer.clear();
    cudaGetLastError();
}

// true when `p` is the first byte of its device allocation (driver API through dlsym: the library does not link libcuda);
// when the driver entry point cannot be found the check is skipped
bool is_allocation_base(const void *p)
{
    typedef int (*get_range_fn)(unsigned long long *, size_t *, unsigned long long);
    static get_range_fn fn = []() -> get_range_fn {
        void *h = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
        return h ? (get_range_fn)dlsym(h, "cuMemGetAddressRange_v2") : nullptr;
    }();
    if (!fn)
        return true;
    unsigned long long base = 0;
    size_t size = 0;
    if (fn(&base, &size, (unsigned long long)(uintptr_t)p) != 0)
        return true;
    return base == (unsigned long long)(uintptr_t)p;
}
} // namespace

// Collective.  Maps every peer's gather destination into this process (CUDA IPC) so that ib200_allgatherv on `out` can push
// slabs with copy engines.  `out` must be the base of a cudaMalloc allocation (ib200_alloc) of `bytes` bytes, the same size
// on every rank.  IB200_ERROR_UNSUPPORTED (on every rank alike) when IPC mapping is not available: the caller keeps NCCL.
extern "C" int ib200_comm_register_gather(ib200_comm_t *comm, double *out, size_t bytes)
{
    IB_TRY(comm_check(comm));
    if (!out || !bytes)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "out must not be NULL");
    IB_TRY(ib200_comm_unregister_gather(comm));
    ib200_ctx *ctx = comm->ctx;
    std::lock_guard<std::mutex> lock(ctx->mutex);
    const int n = comm->n_ranks;
    if (!comm->flag)
    {
        IB_CUDA(cudaMalloc((void **)&comm->flag, sizeof(double)));
        IB_CUDA(cudaMemset(comm->flag, 0, sizeof(double)));
    }
    if (n == 1)
    {
        comm->reg_out = out;
        comm->reg_bytes = bytes;
        return INTERP_SUCCESS;
    }
    struct Slot
    {
        cudaIpcMemHandle_t handle;
        unsigned long long bytes;
        int ok;
        int pad;
    };
    static_assert(sizeof(Slot) % 8 == 0, "slot size");
    std::vector<Slot> slots(n);
    Slot mine{};
    bool not_base = false;
    mine.bytes = bytes;
    mine.ok = (std::getenv("IB200_NO_P2P_GATHER") == nullptr && cudaIpcGetMemHandle(&mine.handle, out) == cudaSuccess) ? 1 : 0;
    cudaGetLastError();
    // `out` must be the base of its allocation: cudaIpcOpenMemHandle hands the peers the BASE, so an interior pointer would
    // make every push land at the wrong offset without any error
    if (mine.ok && !is_allocation_base(out))
    {
        mine.ok = 0;
        not_base = true;
    }
    // from here on every rank must reach the agreeing all-reduce below: local failures only clear `ok`
    bool ok = true;
    Slot *d_slots = nullptr;
    ok = cudaMalloc((void **)&d_slots, sizeof(Slot) * n) == cudaSuccess;
    ok = ok && cudaMemcpyAsync(d_slots + comm->rank, &mine, sizeof mine, cudaMemcpyHostToDevice, ctx->stream) == cudaSuccess;
    if (!ok)
    {
        // without device memory this rank cannot even take part in the handle exchange: fail loudly (the peers see the
        // missing rank through ib200_comm_check / the NCCL watchdog)
        cudaFree(d_slots);
        cudaGetLastError();
        return ib200_fail(IB200_ERROR_CUDA, "ib200_comm_register_gather: no device memory for the handle exchange");
    }
    const ncclResult_t res = g_nccl.AllGather(d_slots + comm->rank, d_slots, sizeof(Slot), ncclChar, comm->nccl, ctx->stream);
    if (res != ncclSuccess)
    {
        cudaFree(d_slots);
        return nccl_fail(comm, res, "ncclAllGather of the IPC handles");
    }
    ok = cudaMemcpyAsync(slots.data(), d_slots, sizeof(Slot) * n, cudaMemcpyDeviceToHost, ctx->stream) == cudaSuccess;
    ok = cudaStreamSynchronize(ctx->stream) == cudaSuccess && ok;
    cudaFree(d_slots);
    for (int r = 0; r < n && ok; ++r)
        ok = ok && slots[r].ok && slots[r].bytes == bytes;
    comm->peer.assign(n, nullptr);
    for (int r = 0; r < n && ok; ++r)
    {
        if (r == comm->rank)
            continue;
        void *p = nullptr;
        if (cudaIpcOpenMemHandle(&p, slots[r].handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess)
        {
            ok = false;
            break;
        }
        comm->peer[r] = (double *)p;
    }
    // the push streams and events are created BEFORE the agreement, so that a failure here is agreed on as well
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    comm->push.assign(n, nullptr);
    comm->pushed.assign(n, nullptr);
    for (int r = 0; r < n && ok; ++r)
    {
        if (r == comm->rank)
            continue;
        ok = cudaStreamCreateWithPriority(&comm->push[r], cudaStreamNonBlocking, prio_hi) == cudaSuccess &&
             cudaEventCreateWithFlags(&comm->pushed[r], cudaEventDisableTiming) == cudaSuccess;
    }
    cudaGetLastError();
    // all ranks or none: agree on the outcome
    double agree = ok ? 0.0 : 1.0;
    IB_CUDA(cudaMemcpyAsync(comm->flag, &agree, sizeof agree, cudaMemcpyHostToDevice, ctx->stream));
    IB_NCCL(comm, g_nccl.AllReduce(comm->flag, comm->flag, 1, ncclDouble, ncclSum, comm->nccl, ctx->stream));
    IB_CUDA(cudaMemcpyAsync(&agree, comm->flag, sizeof agree, cudaMemcpyDeviceToHost, ctx->stream));
    IB_CUDA(cudaStreamSynchronize(ctx->stream));
    IB_CUDA(cudaMemsetAsync(comm->flag, 0, sizeof(double), ctx->stream));
    if (agree != 0.0)
    {
        release_gather_mappings(comm);
        if (not_base)
            return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "ib200_comm_register_gather: `out` is not the base of its allocation");
        return IB200_ERROR_UNSUPPORTED;
    }
    comm->reg_out = out;
    comm->reg_bytes = bytes;
    return INTERP_SUCCESS;
}

extern "C" int ib200_comm_unregister_gather(ib200_comm_t *comm)
{
    if (!comm || !comm->ctx)
        return INTERP_SUCCESS;
    cudaSetDevice(comm->ctx->device);
    release_gather_mappings(comm);
    comm->reg_out = nullptr;
    comm->reg_bytes = 0;
    comm->p2p_pending_side = false;
    return INTERP_SUCCESS;
}

extern "C" int ib200_comm_join(ib200_comm_t *comm)
{
    IB_TRY(comm_check(comm));
    std::lock_guard<std::mutex> lock(comm->ctx->mutex);
    if (!comm->side_busy)
        return INTERP_SUCCESS;
    if (comm->p2p_pending_side)
    {
        // every rank's pushes precede its entry into this all-reduce in stream order: when it completes here, the slabs
        // of all peers have landed in this rank's buffer
        IB_NCCL(comm, g_nccl.AllReduce(comm->flag, comm->flag, 1, ncclDouble, ncclSum, comm->nccl, comm->side));
        comm->p2p_pending_side = false;
    }
    IB_CUDA(cudaEventRecord(comm->join, comm->side));
    IB_CUDA(cudaStreamWaitEvent(comm->ctx->stream, comm->join, 0));
    comm->side_busy = false;
    return INTERP_SUCCESS;
}

extern "C" int ib200_allreduce_sums(ib200_comm_t *comm, double *values, size_t n)
{
    IB_TRY(comm_check(comm));
    if (n == 0)
        return INTERP_SUCCESS;
    if (!values)
        return ib200_fail(IB200_ERROR_BAD_ARGUMENT, "values must not be NULL");
    std::lock_guard<std::mutex> lock(comm->ctx->mutex);
    IB_NCCL(comm, g_nccl.AllReduce(values, values, n, ncclDouble, ncclSum, comm->nccl, comm->ctx->stream));
    return INTERP_SUCCESS;
}
