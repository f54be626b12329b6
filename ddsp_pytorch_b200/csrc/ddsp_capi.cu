// Library-wide pieces of the C ABI: version, per-thread error text, FP32 peak probe.
#include <stdarg.h>
#include <atomic>
#include "ddsp_common.cuh"

namespace ddsp {
static thread_local char g_err[512] = "";

static std::atomic<unsigned long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// 8 independent FFMA chains per thread: enough ILP to saturate the FMA pipe with a few warps per scheduler.
__global__ void __launch_bounds__(1024) fma_probe_kernel(float* out, int iters) {
    float x0 = threadIdx.x * 1e-3f, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f;
    float x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f, x7 = x0 + 7.f;
    const float m = 0.999f, c = 1e-3f * (blockIdx.x + 1);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            x0 = fmaf(x0, m, c); x1 = fmaf(x1, m, c); x2 = fmaf(x2, m, c); x3 = fmaf(x3, m, c);
            x4 = fmaf(x4, m, c); x5 = fmaf(x5, m, c); x6 = fmaf(x6, m, c); x7 = fmaf(x7, m, c);
        }
    }
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}
}  // namespace ddsp

using namespace ddsp;

extern "C" int ddsp_abi_version(void) { return DDSP_B200_ABI_VERSION; }

extern "C" const char* ddsp_last_error(void) { return g_err; }

extern "C" unsigned long long ddsp_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" int ddsp_probe_fp32_fma(float* out, int blocks, int threads, int iters, void* stream) {
    DDSP_REQUIRE(out && blocks > 0 && threads > 0 && threads <= 1024 && iters > 0, "ddsp_probe_fp32_fma: bad arguments");
    fma_probe_kernel<<<blocks, threads, 0, static_cast<cudaStream_t>(stream)>>>(out, iters);
    DDSP_CHECK_LAUNCH("fma_probe_kernel");
    return 0;
}
