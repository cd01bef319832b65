// Write-bandwidth ceiling microbenchmark: the mass-matrix kernels (K2) are write-only streams, so the HBM roofline
// they can reach is the best a pure-write kernel achieves, which is below the copy number in MEASURED_PEAKS.json.
// Variants: plain st.global.cs (8 B and 16 B per lane), TMA bulk shared->global stores of 4/8/16/32 KB chunks
// (one buffer per warp, wait_group.read before refill), cudaMemsetAsync, and a read+write copy for comparison.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o write_peak write_peak.cu && ./write_peak
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x)                                                                          \
    do                                                                                 \
    {                                                                                  \
        cudaError_t e_ = (x);                                                          \
        if (e_ != cudaSuccess)                                                         \
        {                                                                              \
            fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); \
            return 1;                                                                  \
        }                                                                              \
    } while (0)

__global__ void __launch_bounds__(256) st8_kernel(double *out, size_t n)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        __stcs(out + i, 1.0);
}

__global__ void __launch_bounds__(256) st16_kernel(double2 *out, size_t n2)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride)
        __stcs(out + i, make_double2(1.0, 2.0));
}

__global__ void __launch_bounds__(256) copy16_kernel(const double2 *in, double2 *out, size_t n2)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride)
        __stcs(out + i, __ldcs(in + i));
}

// every warp owns NBUF chunk buffers; fills one (16 B per lane per pass), ships it with one bulk copy
template <int NBUF>
__global__ void __launch_bounds__(256) tma_kernel(double *out, size_t n_chunks, int chunk_doubles)
{
    extern __shared__ __align__(128) double smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double *buf0 = smem + (size_t)warp * NBUF * chunk_doubles;
    const size_t n_warps = (size_t)gridDim.x * (blockDim.x >> 5);
    int b = 0;
    for (size_t c = (size_t)blockIdx.x * (blockDim.x >> 5) + warp; c < n_chunks; c += n_warps)
    {
        double *buf = buf0 + (size_t)b * chunk_doubles;
        if (lane == 0)
        {
            if (NBUF == 1)
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            else
                asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        }
        __syncwarp();
        for (int i = lane * 2; i < chunk_doubles; i += 64)
            *reinterpret_cast<double2 *>(buf + i) = make_double2(1.0, (double)c);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0)
        {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out + c * (size_t)chunk_doubles),
                         "r"((unsigned)__cvta_generic_to_shared(buf)), "r"((unsigned)chunk_doubles * 8u)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        b = (b + 1) % NBUF;
    }
    if (lane == 0)
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <typename F>
static float time_ms(F &&launch, int reps = 10)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    for (int i = 0; i < 3; ++i)
        launch();
    float best = 1e30f;
    for (int r = 0; r < reps; ++r)
    {
        cudaEventRecord(a);
        launch();
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        best = ms < best ? ms : best;
    }
    return best;
}

int main()
{
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const size_t bytes = (size_t)4 << 30, n = bytes / 8;
    double *out, *in;
    CK(cudaMalloc(&out, bytes));
    CK(cudaMalloc(&in, bytes));
    CK(cudaMemset(in, 0, bytes));
    const int sms = prop.multiProcessorCount;
    printf("{\"gpu\": \"%s\", \"sm_count\": %d, \"bytes\": %zu", prop.name, sms, bytes);
    for (int per_sm : {2, 4, 8})
    {
        float ms = time_ms([&] { st8_kernel<<<sms * per_sm, 256>>>(out, n); });
        printf(", \"st8_x%d_gbs\": %.1f", per_sm, bytes / ms / 1e6);
        ms = time_ms([&] { st16_kernel<<<sms * per_sm, 256>>>((double2 *)out, n / 2); });
        printf(", \"st16_x%d_gbs\": %.1f", per_sm, bytes / ms / 1e6);
    }
    {
        float ms = time_ms([&] { copy16_kernel<<<sms * 8, 256>>>((const double2 *)in, (double2 *)out, n / 2); });
        printf(", \"copy16_rw_gbs\": %.1f", 2.0 * bytes / ms / 1e6);
        ms = time_ms([&] { cudaMemsetAsync(out, 0, bytes); });
        printf(", \"memset_gbs\": %.1f", bytes / ms / 1e6);
        ms = time_ms([&] { cudaMemcpyAsync(out, in, bytes, cudaMemcpyDeviceToDevice); });
        printf(", \"memcpy_d2d_rw_gbs\": %.1f", 2.0 * bytes / ms / 1e6);
    }
    for (int kb4 : {4, 8, 16, 32, 64, 128}) // quarter-KB units: 1, 2, 4, 8, 16, 32 KB
    {
        const int chunk_doubles = kb4 * 32;
        const size_t n_chunks = n / chunk_doubles;
        for (int nbuf : {1, 2})
        {
            const size_t smem = (size_t)8 * nbuf * chunk_doubles * 8; // 8 warps per CTA
            if (smem > 200 * 1024)
                continue;
            int per_sm = (int)((220 * 1024) / smem);
            per_sm = per_sm > 8 ? 8 : per_sm;
            float ms;
            if (nbuf == 1)
            {
                CK(cudaFuncSetAttribute(tma_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                ms = time_ms([&] { tma_kernel<1><<<sms * per_sm, 256, smem>>>(out, n_chunks, chunk_doubles); });
            }
            else
            {
                CK(cudaFuncSetAttribute(tma_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                ms = time_ms([&] { tma_kernel<2><<<sms * per_sm, 256, smem>>>(out, n_chunks, chunk_doubles); });
            }
            printf(", \"tma_%dB_nbuf%d_cta%d_gbs\": %.1f", chunk_doubles * 8, nbuf, per_sm, bytes / ms / 1e6);
        }
    }
    // chunks whose ends are 16-byte but not 32-byte (sector) aligned, like the row panels of odd-sized element matrices
    for (int chunk_doubles : {250, 626, 2500})
    {
        const size_t n_chunks = n / chunk_doubles;
        const size_t smem = (size_t)8 * chunk_doubles * 8;
        int per_sm = (int)((220 * 1024) / smem);
        per_sm = per_sm > 8 ? 8 : per_sm;
        CK(cudaFuncSetAttribute(tma_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        float ms = time_ms([&] { tma_kernel<1><<<sms * per_sm, 256, smem>>>(out, n_chunks, chunk_doubles); });
        printf(", \"tma_%dB_unaligned_nbuf1_cta%d_gbs\": %.1f", chunk_doubles * 8, per_sm, (double)n_chunks * chunk_doubles * 8 / ms / 1e6);
    }
    printf("}\n");
    CK(cudaDeviceSynchronize());
    return 0;
}
