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
}  // namespace dgmc

extern "C" const char* dgmc_last_error(void) { return dgmc::g_err; }
extern "C" int dgmc_abi_version(void) { return DGMC_B200_ABI_VERSION; }
