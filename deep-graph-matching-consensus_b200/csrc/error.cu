// dgmc_last_error(): thread-local message of the last failing entry point.
#include <stdarg.h>
#include <string.h>
#include "common.cuh"
#include "../../include/dgmc_b200.h"

namespace dgmc {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
static unsigned long long g_launches = 0;
void count_launches(int n) { __atomic_fetch_add(&g_launches, (unsigned long long)n, __ATOMIC_RELAXED); }
}  // namespace dgmc

extern "C" uint64_t dgmc_launch_count(void) {
    return __atomic_load_n(&dgmc::g_launches, __ATOMIC_RELAXED);
}

extern "C" const char* dgmc_last_error(void) { return dgmc::g_err; }
extern "C" int dgmc_abi_version(void) { return DGMC_B200_ABI_VERSION; }
