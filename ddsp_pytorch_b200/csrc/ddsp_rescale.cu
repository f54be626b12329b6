// OscillatorBank.rescale (harmonic_oscillator.py:52-55): F.interpolate(x.permute(0,2,1), scale_factor=hop, mode='linear')
// .permute(0,2,1) -- linear upsampling of frame-rate controls [B,T,C] to the sample rate [B,T*hop,C] with half-pixel centres
// (align_corners=False): output n reads position src = max((n + 1/2)/hop - 1/2, 0), left frame floor(src), right frame
// clamped to T-1.  The synthesis kernels never materialise this tensor (the oscillator folds the interpolation into its
// segments); this entry point exists for the reference's public helper methods, which the drop-in keeps.
// HBM-bound: 4 bytes written per output element, the input is read through L1/L2.
#include "ddsp_common.cuh"

namespace ddsp {

// same arithmetic as torch's upsample_linear1d (area_pixel_compute_source_index, scale = 1/scale_factor in fp32)
__device__ __forceinline__ void rescale_taps(long long n, int hop, int T, int& i0, int& i1, float& lam) {
    const float scale = 1.0f / (float)hop;
    float src = scale * ((float)n + 0.5f) - 0.5f;
    src = src < 0.f ? 0.f : src;
    i0 = (int)src;
    if (i0 > T - 1) i0 = T - 1;
    i1 = i0 + (i0 < T - 1 ? 1 : 0);
    lam = src - (float)i0;
}

__global__ void __launch_bounds__(256) rescale_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int T, int C, int hop,
                                                          long long total) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;       // over B * N * C, channel fastest
    if (e >= total) return;
    const long long N = (long long)T * hop;
    const int c = (int)(e % C);
    const long long bn = e / C;
    const long long n = bn % N, b = bn / N;
    int i0, i1;
    float lam;
    rescale_taps(n, hop, T, i0, i1, lam);
    const float* xb = x + (size_t)b * T * C + c;
    y[e] = (1.0f - lam) * xb[(size_t)i0 * C] + lam * xb[(size_t)i1 * C];
}

// adjoint: frame t collects every output whose left or right tap it is -- the samples of segments t and t+1 (and the clamped
// head / tail).  One thread per (b, t, c) walks the at most 2*hop + hop/2 candidates in order: a fixed summation order.
__global__ void __launch_bounds__(256) rescale_bwd_kernel(const float* __restrict__ gy, float* __restrict__ gx, int T, int C, int hop,
                                                          long long total) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;       // over B * T * C
    if (e >= total) return;
    const long long N = (long long)T * hop;
    const int c = (int)(e % C);
    const long long bt = e / C;
    const int t = (int)(bt % T);
    const long long b = bt / T;
    long long lo = (long long)(t - 1) * hop, hi = (long long)(t + 2) * hop;
    if (t == 0) lo = 0;
    if (t == T - 1) hi = N;
    lo = lo < 0 ? 0 : lo;
    hi = hi > N ? N : hi;
    const float* gb = gy + (size_t)b * N * C + c;
    float acc = 0.f;
    for (long long n = lo; n < hi; ++n) {
        int i0, i1;
        float lam;
        rescale_taps(n, hop, T, i0, i1, lam);
        float w = 0.f;
        if (i0 == t) w += 1.0f - lam;
        if (i1 == t) w += lam;
        if (w != 0.f) acc = fmaf(w, gb[(size_t)n * C], acc);
    }
    gx[e] = acc;
}

}  // namespace ddsp

using namespace ddsp;

static int check_rescale(const char* fn, int B, int T, int C, int hop) {
    DDSP_REQUIRE(B >= 1 && T >= 1 && C >= 1 && hop >= 1, "%s: B=%d, T=%d, C=%d, hop=%d must be positive", fn, B, T, C, hop);
    DDSP_REQUIRE((double)B * T * hop * C < 9.0e18 && ((long long)B * T * hop * C + 255) / 256 <= 2147483647LL,
                 "%s: %d x %d x %d x %d elements do not fit one launch", fn, B, T * hop, C, 1);
    return 0;
}

extern "C" int ddsp_rescale_forward(const float* x, float* y, int B, int T, int C, int hop, void* stream) {
    if (check_rescale("ddsp_rescale_forward", B, T, C, hop)) return 1;
    DDSP_REQUIRE(x && y, "ddsp_rescale_forward: null pointer argument");
    const long long total = (long long)B * T * hop * C;
    rescale_fwd_kernel<<<(unsigned)((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, y, T, C, hop, total);
    DDSP_CHECK_LAUNCH("rescale_fwd_kernel");
    return 0;
}

extern "C" int ddsp_rescale_backward(const float* gy, float* gx, int B, int T, int C, int hop, void* stream) {
    if (check_rescale("ddsp_rescale_backward", B, T, C, hop)) return 1;
    DDSP_REQUIRE(gy && gx, "ddsp_rescale_backward: null pointer argument");
    const long long total = (long long)B * T * C;
    rescale_bwd_kernel<<<(unsigned)((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(gy, gx, T, C, hop, total);
    DDSP_CHECK_LAUNCH("rescale_bwd_kernel");
    return 0;
}
