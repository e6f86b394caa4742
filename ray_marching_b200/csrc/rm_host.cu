// Error reporting of the C ABI (include/rm_b200.h).
#include "rm_host.h"

namespace rm {

char* error_buffer() {
    static thread_local char buf[512] = {0};
    return buf;
}

int fail(int status, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(error_buffer(), 512, fmt, ap);
    va_end(ap);
    return status;
}

}  // namespace rm

extern "C" int rm_abi_version(void) { return RM_B200_ABI_VERSION; }
extern "C" const char* rm_last_error(void) { return rm::error_buffer(); }
