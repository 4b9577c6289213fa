// Error reporting shared by all entry points of the C-ABI library.
#include "common.cuh"
#include "../../include/pinot_b200.h"

namespace pb {

static thread_local char g_last_error[512] = {0};
static unsigned long long g_launches = 0;

void count_launch() { __atomic_add_fetch(&g_launches, 1ull, __ATOMIC_RELAXED); }

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
    va_end(ap);
}

}  // namespace pb

extern "C" {

const char* pinot_b200_last_error(void) { return pb::g_last_error; }

int pinot_b200_abi_version(void) { return 4; }

int64_t pinot_b200_launch_count(void) { return (int64_t)__atomic_load_n(&pb::g_launches, __ATOMIC_RELAXED); }

}  // extern "C"
