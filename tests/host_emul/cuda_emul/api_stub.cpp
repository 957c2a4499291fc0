// api_stub.cpp — the pieces of csrc/api.cu the kernels' launchers need (error string, stage check, launch counter,
// profiling hooks as no-ops) for the CPU emulation build.  TEST INFRASTRUCTURE.
#include <stdarg.h>
#include <stdio.h>

#include "../../../gaussian_splatting_b200/csrc/common.cuh"

namespace gsr {
static thread_local char g_err[512] = "";
unsigned long long g_launches = 0;
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
int check_stage(const char*, bool, cudaStream_t) { return GSR_OK; }
void profile_begin(int, cudaStream_t) {}
void profile_end(int, cudaStream_t) {}
}  // namespace gsr

extern "C" const char* gsr_last_error(void) { return gsr::g_err; }
extern "C" uint64_t gsr_launch_count(void) { return gsr::g_launches; }
