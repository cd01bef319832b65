// Host <-> device copy ceiling of one box: what the end-to-end (host-buffer) path of bench.py can at best reach.
//
//   nvcc -O2 -std=c++17 -o tools/pcie_peak tools/pcie_peak.cu -lpthread && tools/pcie_peak > profiles/r02_pcie_peak.json
//
// For G = 1, 2, 4, 8 GPUs used at the same time (as many as the box has): one host thread per GPU, one 256 MB pinned
// buffer per direction, cudaMemcpyAsync on the GPU's own stream, timed with the host clock around a barrier-started
// burst of copies.  Three patterns -- host-to-device only, device-to-host only, both directions at once -- each with
// the threads unbound and bound to the CPUs next to their GPU (/sys/bus/pci/devices/<bus id>/local_cpulist) BEFORE
// the pinned buffers are allocated (first touch places them on that NUMA node).  Prints one JSON object.
#include <cuda_runtime.h>
#include <pthread.h>
#include <sched.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#define CK(call)                                                                                   \
    do                                                                                             \
    {                                                                                              \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
        {                                                                                          \
            fprintf(stderr, "%s: %s (%s:%d)\n", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
            exit(1);                                                                               \
        }                                                                                          \
    } while (0)

static const size_t kBytes = (size_t)256 << 20;
static const int kReps = 8;

static bool bind_near(int dev, std::string *desc)
{
    char bus[32];
    if (cudaDeviceGetPCIBusId(bus, sizeof bus, dev) != cudaSuccess)
        return false;
    for (char *p = bus; *p; ++p)
        *p = (char)tolower(*p);
    const std::string path = std::string("/sys/bus/pci/devices/") + bus + "/local_cpulist";
    FILE *f = fopen(path.c_str(), "r");
    if (!f)
        return false;
    char line[1024] = {0};
    const bool ok = fgets(line, sizeof line, f) != nullptr;
    fclose(f);
    if (!ok)
        return false;
    line[strcspn(line, "\n")] = 0;
    *desc = line;
    cpu_set_t set;
    CPU_ZERO(&set);
    int n_set = 0;
    for (char *tok = strtok(line, ","); tok; tok = strtok(nullptr, ","))
    {
        int a = 0, b = 0;
        if (sscanf(tok, "%d-%d", &a, &b) == 2)
            ;
        else if (sscanf(tok, "%d", &a) == 1)
            b = a;
        else
            continue;
        for (int c = a; c <= b && c < CPU_SETSIZE; ++c, ++n_set)
            CPU_SET(c, &set);
    }
    return n_set > 0 && sched_setaffinity(0, sizeof set, &set) == 0;
}

struct Result
{
    double h2d = 0, d2h = 0, bi_h2d = 0, bi_d2h = 0; // GB/s of this GPU
    std::string cpus;
};

static pthread_barrier_t g_bar;

static void worker(int dev, bool bind, Result *res)
{
    CK(cudaSetDevice(dev));
    if (bind)
        bind_near(dev, &res->cpus);
    void *h_in, *h_out, *d_in, *d_out;
    CK(cudaMallocHost(&h_in, kBytes));
    CK(cudaMallocHost(&h_out, kBytes));
    memset(h_in, 1, kBytes); // first touch on this thread's node
    memset(h_out, 0, kBytes);
    CK(cudaMalloc(&d_in, kBytes));
    CK(cudaMalloc(&d_out, kBytes));
    cudaStream_t s1, s2;
    CK(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
    auto burst = [&](bool up, bool down) {
        pthread_barrier_wait(&g_bar);
        const auto t0 = std::chrono::steady_clock::now();
        for (int r = 0; r < kReps; ++r)
        {
            if (up)
                CK(cudaMemcpyAsync(d_in, h_in, kBytes, cudaMemcpyHostToDevice, s1));
            if (down)
                CK(cudaMemcpyAsync(h_out, d_out, kBytes, cudaMemcpyDeviceToHost, s2));
        }
        CK(cudaStreamSynchronize(s1));
        CK(cudaStreamSynchronize(s2));
        const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        pthread_barrier_wait(&g_bar);
        return (double)kBytes * kReps / s / 1e9;
    };
    burst(true, true); // warm-up
    res->h2d = burst(true, false);
    res->d2h = burst(false, true);
    const double both = burst(true, true);
    res->bi_h2d = res->bi_d2h = both;
    CK(cudaFreeHost(h_in));
    CK(cudaFreeHost(h_out));
    CK(cudaFree(d_in));
    CK(cudaFree(d_out));
}

int main()
{
    int n_dev = 0;
    CK(cudaGetDeviceCount(&n_dev));
    printf("{\"bytes_per_copy\": %zu, \"copies_per_burst\": %d, \"gpus_on_box\": %d, \"host_threads\": %u, \"runs\": [", kBytes, kReps, n_dev,
           std::thread::hardware_concurrency());
    bool first = true;
    for (int g : {1, 2, 4, 8})
    {
        if (g > n_dev)
            break;
        for (int bind = 0; bind < 2; ++bind)
        {
            std::vector<Result> res(g);
            pthread_barrier_init(&g_bar, nullptr, g);
            std::vector<std::thread> th;
            for (int d = 0; d < g; ++d)
                th.emplace_back(worker, d, bind != 0, &res[d]);
            for (auto &t : th)
                t.join();
            pthread_barrier_destroy(&g_bar);
            double h2d = 0, d2h = 0, bi = 0, min_h2d = 1e30, min_d2h = 1e30;
            for (auto &r : res)
            {
                h2d += r.h2d, d2h += r.d2h, bi += r.bi_h2d;
                min_h2d = std::min(min_h2d, r.h2d), min_d2h = std::min(min_d2h, r.d2h);
            }
            printf("%s\n {\"gpus\": %d, \"numa_bound\": %s, \"cpus_of_gpu0\": \"%s\", \"h2d_gbs_total\": %.1f, \"d2h_gbs_total\": %.1f, "
                   "\"bidirectional_gbs_total_each_way\": %.1f, \"h2d_gbs_min_per_gpu\": %.1f, \"d2h_gbs_min_per_gpu\": %.1f}",
                   first ? "" : ",", g, bind ? "true" : "false", res[0].cpus.c_str(), h2d, d2h, bi, min_h2d, min_d2h);
            first = false;
        }
    }
    printf("\n]}\n");
    return 0;
}
