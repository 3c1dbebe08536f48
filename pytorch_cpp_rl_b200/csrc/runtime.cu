// runtime.cu — error state, device checks, launch counting and CUDA-event kernel timers.
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>

#include "common.cuh"

namespace cpprl_b200
{
static thread_local std::string t_error;
std::atomic<uint64_t> g_launches{0};

void set_error(const std::string &msg) { t_error = msg; }

int fail(int code, const std::string &msg)
{
    t_error = msg;
    return code;
}

int check_cuda(cudaError_t e, const char *what)
{
    if (e == cudaSuccess)
        return 0;
    t_error = std::string("CUDA error: ") + cudaGetErrorString(e) + " at " + what;
    return CPPRL_B200_ECUDA;
}

int require_device()
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess)
        return fail(CPPRL_B200_ECUDA, std::string("no CUDA device: ") + cudaGetErrorString(e) +
                                          " (cpprl_b200 has no CPU fallback)");
    static int cached_dev = -1, cached_major = 0;
    if (cached_dev != dev)
    {
        int major = 0;
        CPPRL_CUDA_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
        cached_dev = dev;
        cached_major = major;
    }
    if (cached_major != 10)
        return fail(CPPRL_B200_ECUDA, "device is not compute capability 10.x (kernels are built for sm_100a only)");
    return 0;
}

int sm_count()
{
    static int cached = 0;
    if (cached == 0)
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&cached, cudaDevAttrMultiProcessorCount, dev);
        if (cached <= 0)
            cached = 148;
    }
    return cached;
}

// ---- kernel timers ---------------------------------------------------------------------------
namespace
{
struct Family
{
    std::vector<cudaEvent_t> start, stop;
    size_t used = 0;
};
std::mutex g_timer_mu;
std::map<std::string, Family> g_families;
bool g_timing = false;
} // namespace

KernelTimer::KernelTimer(const char *fam, cudaStream_t s) : family(fam), stream(s), slot(-1)
{
    if (!g_timing)
        return;
    std::lock_guard<std::mutex> lock(g_timer_mu);
    Family &f = g_families[fam];
    if (f.used == f.start.size())
    {
        cudaEvent_t a, b;
        if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess)
            return;
        f.start.push_back(a);
        f.stop.push_back(b);
    }
    slot = static_cast<int>(f.used++);
    cudaEventRecord(f.start[slot], stream);
}

KernelTimer::~KernelTimer()
{
    if (slot < 0)
        return;
    std::lock_guard<std::mutex> lock(g_timer_mu);
    Family &f = g_families[family];
    cudaEventRecord(f.stop[slot], stream);
}
} // namespace cpprl_b200

using namespace cpprl_b200;

extern "C"
{
int cpprl_b200_abi_version(void) { return CPPRL_B200_ABI_VERSION; }

const char *cpprl_b200_last_error(void) { return t_error.c_str(); }

uint64_t cpprl_b200_launch_count(void) { return g_launches.load(); }

int cpprl_b200_device_info(int *sms, int *cc_major, int *cc_minor, char *name, int name_len)
{
    int dev = 0;
    CPPRL_CUDA_TRY(cudaGetDevice(&dev));
    cudaDeviceProp prop;
    CPPRL_CUDA_TRY(cudaGetDeviceProperties(&prop, dev));
    if (sms)
        *sms = prop.multiProcessorCount;
    if (cc_major)
        *cc_major = prop.major;
    if (cc_minor)
        *cc_minor = prop.minor;
    if (name && name_len > 0)
    {
        std::strncpy(name, prop.name, name_len - 1);
        name[name_len - 1] = 0;
    }
    return require_device();
}

void cpprl_b200_kernel_time_enable(int on)
{
    std::lock_guard<std::mutex> lock(g_timer_mu);
    g_timing = on != 0;
}

void cpprl_b200_kernel_time_reset(void)
{
    std::lock_guard<std::mutex> lock(g_timer_mu);
    for (auto &kv : g_families)
        kv.second.used = 0;
}

int cpprl_b200_kernel_time_ms(const char *family, double *total_ms, int64_t *launches)
{
    std::lock_guard<std::mutex> lock(g_timer_mu);
    double total = 0;
    int64_t n = 0;
    auto it = g_families.find(family ? family : "");
    if (it != g_families.end())
    {
        Family &f = it->second;
        for (size_t i = 0; i < f.used; ++i)
        {
            CPPRL_CUDA_TRY(cudaEventSynchronize(f.stop[i]));
            float ms = 0;
            CPPRL_CUDA_TRY(cudaEventElapsedTime(&ms, f.start[i], f.stop[i]));
            total += ms;
            ++n;
        }
    }
    if (total_ms)
        *total_ms = total;
    if (launches)
        *launches = n;
    return 0;
}
}
