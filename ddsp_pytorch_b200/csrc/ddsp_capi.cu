// Library-wide pieces of the C ABI: version, per-thread error text, FP32 peak probe.
#include <stdarg.h>
#include <atomic>
#include <mutex>
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

#ifdef DDSP_HAVE_TUNED
#include <cstdio>
#include <cstdlib>
#include "ddsp_tuned.cuh"
namespace ddsp {
namespace {
struct DriverApi {
    CUresult (*moduleLoadData)(CUmodule*, const void*);
    CUresult (*moduleGetFunction)(CUfunction*, CUmodule, const char*);
    CUresult (*funcSetAttribute)(CUfunction, CUfunction_attribute, int);
    CUresult (*launchKernel)(CUfunction, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, CUstream, void**, void**);
    int state;                    // 0 not resolved, 1 ok, -1 missing
};
DriverApi g_drv = {};
std::mutex g_drv_lock;
std::atomic<int> g_tuned_on{1};

template <class F> bool drv_entry(const char* name, F& fn) {      // through the runtime: no link-time dependency on libcuda
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess || !p) return false;
    fn = reinterpret_cast<F>(p);
    return true;
}

bool driver_ready() {
    if (g_drv.state == 0) {
        std::lock_guard<std::mutex> lock(g_drv_lock);
        if (g_drv.state == 0)
            g_drv.state = (drv_entry("cuModuleLoadData", g_drv.moduleLoadData) && drv_entry("cuModuleGetFunction", g_drv.moduleGetFunction) &&
                           drv_entry("cuFuncSetAttribute", g_drv.funcSetAttribute) && drv_entry("cuLaunchKernel", g_drv.launchKernel)) ? 1 : -1;
    }
    return g_drv.state == 1;
}
}  // namespace

void tuned_enable(int on) { g_tuned_on.store(on ? 1 : 0, std::memory_order_relaxed); }
int tuned_enabled() { return g_tuned_on.load(std::memory_order_relaxed); }

CUfunction tuned_function(TunedImage& img, TunedKernel& k, cudaStream_t st, bool have_stream) {
    if (!tuned_enabled() || !k.name || !k.name[0]) return nullptr;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= TUNED_MAX_DEVICES) return nullptr;
    if (k.state[dev] == 0 || img.state[dev] == 0) {
        if (have_stream) {
            cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
            if (cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) {
                cudaGetLastError();
                return nullptr;
            }
        }
        std::lock_guard<std::mutex> lock(img.lock);
        if (img.state[dev] == 0) {
            int st = -1;
            const void* image = img.image;
            if (const char* path = img.env ? getenv(img.env) : nullptr) {
                if (FILE* fp = fopen(path, "rb")) {
                    fseek(fp, 0, SEEK_END);
                    const long n = ftell(fp);
                    fseek(fp, 0, SEEK_SET);
                    img.file.resize(n > 0 ? n : 0);
                    if (n > 0 && fread(img.file.data(), 1, n, fp) == (size_t)n) image = img.file.data();
                    fclose(fp);
                }
            }
            CUmodule mod = nullptr;
            if (driver_ready() && cudaFree(nullptr) == cudaSuccess && g_drv.moduleLoadData(&mod, image) == CUDA_SUCCESS) {
                img.module[dev] = mod;
                st = 1;
            }
            cudaGetLastError();
            img.state[dev] = st;
        }
        if (k.state[dev] == 0) {
            CUfunction f = nullptr;
            k.fn[dev] = (img.state[dev] == 1 && g_drv.moduleGetFunction(&f, img.module[dev], k.name) == CUDA_SUCCESS) ? f : nullptr;
            k.state[dev] = k.fn[dev] ? 1 : -1;
        }
    }
    return k.state[dev] == 1 ? k.fn[dev] : nullptr;
}

int tuned_launch(CUfunction f, dim3 grid, unsigned threads, size_t smem, cudaStream_t st, void** params, const char* what) {
    if (smem > 48 * 1024) {
        CUresult r = g_drv.funcSetAttribute(f, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, (int)smem);
        if (r == CUDA_ERROR_INVALID_CONTEXT) {
            cudaFree(nullptr);
            r = g_drv.funcSetAttribute(f, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, (int)smem);
        }
        DDSP_REQUIRE(r == CUDA_SUCCESS, "%s: cannot reserve %zu B of shared memory (driver error %d)", what, smem, (int)r);
    }
    CUresult r = g_drv.launchKernel(f, grid.x, grid.y, grid.z, threads, 1, 1, (unsigned)smem, st, params, nullptr);
    if (r == CUDA_ERROR_INVALID_CONTEXT) {
        // a thread that has made no runtime call yet (an autograd worker served from the caching allocator) has no current
        // context for the driver API: bind the device's primary context the way the runtime does, then launch again
        cudaFree(nullptr);
        r = g_drv.launchKernel(f, grid.x, grid.y, grid.z, threads, 1, 1, (unsigned)smem, st, params, nullptr);
    }
    count_launch();
    DDSP_REQUIRE(r == CUDA_SUCCESS, "%s: launch failed (driver error %d)", what, (int)r);
    return 0;
}
}  // namespace ddsp
#endif

namespace ddsp {
int osc_tuned_mask();        // ddsp_osc.cu: bit 0 synthesis, bit 1 gradient kernel available from the hand-scheduled image
}

using namespace ddsp;

extern "C" int ddsp_osc_tuned_sass(int mode) {
#ifdef DDSP_HAVE_TUNED
    if (mode == 0 || mode == 1) tuned_enable(mode);
    return tuned_enabled() ? osc_tuned_mask() : 0;
#else
    (void)mode;
    return 0;
#endif
}

extern "C" int ddsp_abi_version(void) { return DDSP_B200_ABI_VERSION; }

extern "C" const char* ddsp_last_error(void) { return g_err; }

extern "C" unsigned long long ddsp_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" int ddsp_probe_fp32_fma(float* out, int blocks, int threads, int iters, void* stream) {
    DDSP_REQUIRE(out && blocks > 0 && threads > 0 && threads <= 1024 && iters > 0, "ddsp_probe_fp32_fma: bad arguments");
    fma_probe_kernel<<<blocks, threads, 0, static_cast<cudaStream_t>(stream)>>>(out, iters);
    DDSP_CHECK_LAUNCH("fma_probe_kernel");
    return 0;
}
