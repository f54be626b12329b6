// fft_convolve for sm_100a with one kernel PER ROW -- model/ddsp/filtered_noise.py:25-32 in its general form:
//   signal [..., n], kernel [..., n]  ->  irfft(rfft(pad(signal, right n)) * rfft(pad(kernel, left n)))[..., n:]
// which is the causal convolution truncated to the row:  y[r, m] = sum_{q <= m} kernel[r, q] * signal[r, m - q].
// The hot path never calls this (FilteredNoise fuses taps + noise + convolution in ddsp_noise.cu, Reverb's shared long
// kernel goes through the partitioned FFT convolution in ddsp_reverb.cu); it exists so that every call shape of the
// reference function has a CUDA path.  Direct form from shared memory, one CTA per row, n <= 4096:
// the O(n^2/2) triangle is cheaper than two padded FFTs for the row lengths the reference uses (n = hop <= 512).
#include "ddsp_common.cuh"

namespace ddsp {

constexpr int ROWCONV_THREADS = 256;
constexpr int ROWCONV_MAX_N = 4096;

// mode 0: out[m] = sum_{q=0..m}     b[q] * a[m - q]      (a = signal, b = kernel)
// mode 1: out[j] = sum_{p=0..n-1-j} b[p] * a[j + p]      (a = upstream gradient, b = kernel or signal: both adjoints)
template <int MODE>
__global__ void __launch_bounds__(ROWCONV_THREADS) rowconv_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                                  float* __restrict__ out, int n) {
    extern __shared__ __align__(16) float rsm[];
    float* as = rsm;                                    // [n + 4]
    float* bs = rsm + ((n + 4 + 3) & ~3);               // [n rounded up to 4], zero padded
    const size_t row = blockIdx.x;
    const int n4 = (n + 3) & ~3;
    for (int i = threadIdx.x; i < n4; i += ROWCONV_THREADS) {
        as[i] = i < n ? a[row * n + i] : 0.f;
        bs[i] = i < n ? b[row * n + i] : 0.f;
    }
    __syncthreads();
    for (int m = threadIdx.x; m < n; m += ROWCONV_THREADS) {
        float acc = 0.f;
        if (MODE == 0) {
            int q = 0;
            for (; q + 4 <= m + 1; q += 4) {
                const float4 k4 = *reinterpret_cast<const float4*>(bs + q);
                acc = fmaf(k4.x, as[m - q], acc);
                acc = fmaf(k4.y, as[m - q - 1], acc);
                acc = fmaf(k4.z, as[m - q - 2], acc);
                acc = fmaf(k4.w, as[m - q - 3], acc);
            }
            for (; q <= m; ++q) acc = fmaf(bs[q], as[m - q], acc);
        } else {
            const int len = n - m;
            int p = 0;
            for (; p + 4 <= len; p += 4) {
                const float4 k4 = *reinterpret_cast<const float4*>(bs + p);
                acc = fmaf(k4.x, as[m + p], acc);
                acc = fmaf(k4.y, as[m + p + 1], acc);
                acc = fmaf(k4.z, as[m + p + 2], acc);
                acc = fmaf(k4.w, as[m + p + 3], acc);
            }
            for (; p < len; ++p) acc = fmaf(bs[p], as[m + p], acc);
        }
        out[row * n + m] = acc;
    }
}

template <int MODE>
static int launch_rowconv(const float* a, const float* b, float* out, long long rows, int n, cudaStream_t st, const char* what) {
    const size_t smem = sizeof(float) * (((n + 4 + 3) & ~3) + ((n + 3) & ~3));
    rowconv_kernel<MODE><<<(unsigned)rows, ROWCONV_THREADS, smem, st>>>(a, b, out, n);
    DDSP_CHECK_LAUNCH(what);
    return 0;
}

static int check_rowconv(const char* fn, long long rows, int n) {
    DDSP_REQUIRE(rows >= 1 && rows <= 2147483647LL, "%s: rows=%lld out of range [1, 2^31)", fn, rows);
    DDSP_REQUIRE(n >= 1 && n <= ROWCONV_MAX_N, "%s: row length n=%d out of range [1,%d]", fn, n, ROWCONV_MAX_N);
    return 0;
}

}  // namespace ddsp

using namespace ddsp;

extern "C" int ddsp_rowconv_forward(const float* signal, const float* kernel, float* y, long long rows, int n, void* stream) {
    if (check_rowconv("ddsp_rowconv_forward", rows, n)) return 1;
    DDSP_REQUIRE(signal && kernel && y, "ddsp_rowconv_forward: null pointer argument");
    return launch_rowconv<0>(signal, kernel, y, rows, n, static_cast<cudaStream_t>(stream), "rowconv_kernel<forward>");
}

extern "C" int ddsp_rowconv_backward(const float* g, const float* signal, const float* kernel, float* gsignal, float* gkernel,
                                     long long rows, int n, void* stream) {
    if (check_rowconv("ddsp_rowconv_backward", rows, n)) return 1;
    DDSP_REQUIRE(g && (gsignal || gkernel), "ddsp_rowconv_backward: null pointer argument");
    DDSP_REQUIRE((!gsignal || kernel) && (!gkernel || signal), "ddsp_rowconv_backward: the other operand is needed for each gradient");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (gsignal && launch_rowconv<1>(g, kernel, gsignal, rows, n, st, "rowconv_kernel<gsignal>")) return 1;
    if (gkernel && launch_rowconv<1>(g, signal, gkernel, rows, n, st, "rowconv_kernel<gkernel>")) return 1;
    return 0;
}
