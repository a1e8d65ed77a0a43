#include "common.cuh"
#include <cstring>

namespace gb200
{
static thread_local char t_err[512] = "";
std::atomic<uint64_t> g_launches{0};

int set_error(int code, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof t_err, fmt, ap);
    va_end(ap);
    return code;
}

int select_device(int device)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0)
        {
            cudaGetLastError();
            return set_error(GB200_ENODEVICE, "no CUDA device available (%s); this library has no CPU fallback", e != cudaSuccess ? cudaGetErrorString(e) : "0 devices");
        }
    if (device < 0 || device >= n) return set_error(GB200_EINVAL, "device %d out of range (0..%d)", device, n - 1);
    cudaDeviceProp p;
    GB_CUDA(cudaGetDeviceProperties(&p, device));
    if (p.major != 10) return set_error(GB200_ENODEVICE, "device %d is sm_%d%d; this library carries sm_100a code only", device, p.major, p.minor);
    GB_CUDA(cudaSetDevice(device));
    return GB200_OK;
}
}  // namespace gb200

extern "C" {
const char* gb200_last_error(void) { return gb200::t_err; }
uint64_t gb200_launch_count(void) { return gb200::g_launches.load(); }
int gb200_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess)
        {
            cudaGetLastError();
            return 0;
        }
    return n;
}
const char* gb200_version(void) { return "gnss_b200 0.1 (sm_100a)"; }
int gb200_host_alloc(void** out, size_t bytes)
{
    if (!out || bytes == 0) return gb200::set_error(GB200_EINVAL, "gb200_host_alloc: null argument or zero size");
    *out = nullptr;
    if (gb200_device_count() == 0) return gb200::set_error(GB200_ENODEVICE, "no CUDA device (no CPU fallback)");
    GB_CUDA(cudaHostAlloc(out, bytes, cudaHostAllocPortable));
    return GB200_OK;
}
int gb200_host_free(void* p)
{
    if (!p) return GB200_OK;
    GB_CUDA(cudaFreeHost(p));
    return GB200_OK;
}
}
