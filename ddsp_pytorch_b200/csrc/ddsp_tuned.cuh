// Hand-scheduled kernel images (ddsp_pytorch_b200/sass_sched.py).  build.py compiles a source file once more to a cubin,
// lets the scheduler reorder the hot loops of the kernels listed there (same instructions, same registers, bit-identical
// results) and writes the image plus the kernels' mangled names into lib/<source>_tuned.inc, which the source includes.
// At run time the image is loaded through the driver API on the first use on a device (outside stream capture) and those
// kernels are launched from it; whenever something is missing -- no tuned image in this build, the scheduler declined a
// kernel, the driver refused the image, the process switched it off -- the compiled-in kernel runs instead.
#pragma once
#ifdef DDSP_HAVE_TUNED
#include <cuda.h>
#include <cuda_runtime.h>
#include <mutex>
#include <vector>

namespace ddsp {

constexpr int TUNED_MAX_DEVICES = 64;

struct TunedImage {
    const void* image;            // embedded cubin
    const char* env;              // environment variable naming a file that replaces it (scheduler experiments)
    std::mutex lock;
    int state[TUNED_MAX_DEVICES];          // 0 not tried, 1 loaded, -1 unavailable
    CUmodule module[TUNED_MAX_DEVICES];
    std::vector<char> file;
};

struct TunedKernel {
    const char* name;             // mangled name inside the image, "" when the scheduler declined the kernel
    int state[TUNED_MAX_DEVICES];
    CUfunction fn[TUNED_MAX_DEVICES];
};

// the kernel's function in the image on the current device, or nullptr (then launch the compiled-in kernel).  `st` is the
// stream the launch will go to: while it is capturing, an image that is not loaded yet stays unloaded (loading is not a
// capturable operation) and the compiled-in kernel is captured instead.
CUfunction tuned_function(TunedImage& img, TunedKernel& k, cudaStream_t st = nullptr, bool have_stream = false);
int tuned_launch(CUfunction f, dim3 grid, unsigned threads, size_t smem, cudaStream_t st, void** params, const char* what);
void tuned_enable(int on);
int tuned_enabled();

}  // namespace ddsp
#endif
