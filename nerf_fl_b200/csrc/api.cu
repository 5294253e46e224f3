// Error plumbing and ABI version of libnerfw_b200.so.
#include "common.cuh"
#include <algorithm>
#include <atomic>

namespace nwb {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};
void note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

char* err_buf() { return g_err; }

int set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int current_device_checked(int* dev_out) {
    static int sms[kMaxDevices] = {0};               // 0 = not queried yet
    int dev = 0;
    NWB_CUDA(cudaGetDevice(&dev));
    NWB_REQUIRE(dev >= 0 && dev < kMaxDevices, "device ordinal %d out of range", dev);
    if (sms[dev] == 0) {
        int n = 0;
        NWB_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
        sms[dev] = n;
    }
    NWB_REQUIRE(sms[dev] == kNumSMs, "device %d has %d SMs; the persistent grids of this build are sized for %d (B200)",
                dev, sms[dev], kNumSMs);
    *dev_out = dev;
    return NWB_OK;
}

// up to 8 small device buffers copied by ONE launch (a training step's batch -- rays, ts, rgbs -- into the static inputs of
// its CUDA graph: three copy launches in front of every replay otherwise)
struct CopySegs {
    const uint8_t* src[8];
    uint8_t* dst[8];
    long long bytes[8];
    int n;
};
__global__ void __launch_bounds__(256)
copy_segments_kernel(const CopySegs a) {
    const int s = blockIdx.y;
    const long long nb = a.bytes[s];
    const long long tid = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    const long long nth = static_cast<long long>(gridDim.x) * blockDim.x;
    if (((reinterpret_cast<uintptr_t>(a.src[s]) | reinterpret_cast<uintptr_t>(a.dst[s])) & 15u) == 0) {
        const uint4* src = reinterpret_cast<const uint4*>(a.src[s]);
        uint4* dst = reinterpret_cast<uint4*>(a.dst[s]);
        for (long long i = tid; i < (nb >> 4); i += nth) dst[i] = src[i];
        for (long long i = (nb & ~15ll) + tid; i < nb; i += nth) a.dst[s][i] = a.src[s][i];
    } else {
        for (long long i = tid; i < nb; i += nth) a.dst[s][i] = a.src[s][i];
    }
}

}  // namespace nwb

extern "C" int nwb_copy_segments(int n, const void* const* src, void* const* dst, const int64_t* bytes, void* stream) {
    NWB_REQUIRE(n >= 0 && n <= 8 && (n == 0 || (src && dst && bytes)), "nwb_copy_segments: 0..8 segments");
    if (n == 0) return NWB_OK;
    nwb::CopySegs a{};
    long long most = 0;
    for (int i = 0; i < n; ++i) {
        NWB_REQUIRE(bytes[i] >= 0 && (bytes[i] == 0 || (src[i] && dst[i])), "nwb_copy_segments: segment %d is null", i);
        a.src[i] = static_cast<const uint8_t*>(src[i]); a.dst[i] = static_cast<uint8_t*>(dst[i]); a.bytes[i] = bytes[i];
        most = bytes[i] > most ? bytes[i] : most;
    }
    a.n = n;
    if (most == 0) return NWB_OK;
    const int bx = static_cast<int>(std::min<long long>((most / 16 + 255) / 256 + 1, 64));
    nwb::copy_segments_kernel<<<dim3(bx, n), 256, 0, static_cast<cudaStream_t>(stream)>>>(a);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}

extern "C" int nwb_abi_version(void) { return 1; }
// kernels this library has launched in this process (every <<<>>> of ours passes NWB_LAUNCH_CHECK): what bench.py reports
// as gpu_launches; during a CUDA-graph capture it counts the kernels that end up as nodes of the graph
extern "C" long long nwb_kernel_launches(void) { return nwb::g_launches.load(std::memory_order_relaxed); }
extern "C" const char* nwb_last_error(void) { return nwb::err_buf(); }
