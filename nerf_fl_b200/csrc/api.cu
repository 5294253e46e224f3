// Error plumbing and ABI version of libnerfw_b200.so.
#include "common.cuh"
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

}  // namespace nwb

extern "C" int nwb_abi_version(void) { return 1; }
// kernels this library has launched in this process (every <<<>>> of ours passes NWB_LAUNCH_CHECK): what bench.py reports
// as gpu_launches; during a CUDA-graph capture it counts the kernels that end up as nodes of the graph
extern "C" long long nwb_kernel_launches(void) { return nwb::g_launches.load(std::memory_order_relaxed); }
extern "C" const char* nwb_last_error(void) { return nwb::err_buf(); }
