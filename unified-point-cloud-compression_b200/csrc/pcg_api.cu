// Error reporting, version and launch accounting for libpcg.so.
#include <stdarg.h>

#include "pcg_common.cuh"

thread_local char pcg_err_buf[512] = "";
std::atomic<long> pcg_launches{0};

int pcg_fail(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(pcg_err_buf, sizeof(pcg_err_buf), fmt, ap);
    va_end(ap);
    return 1;
}

extern "C" const char *pcg_last_error(void) { return pcg_err_buf; }
extern "C" int pcg_version(void) { return 100; }
extern "C" long pcg_launch_count(void) { return pcg_launches.load(std::memory_order_relaxed); }
