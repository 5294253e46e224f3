// Error plumbing and ABI version of libnerfw_b200.so.
#include "common.cuh"

namespace nwb {

static thread_local char g_err[512] = "";

char* err_buf() { return g_err; }

int set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

}  // namespace nwb

extern "C" int nwb_abi_version(void) { return 1; }
extern "C" const char* nwb_last_error(void) { return nwb::err_buf(); }
