// Shared-memory complex FFT for power-of-two sizes (Stockham autosort, radix 4 with one radix-2 pass when
// log2(n) is odd).  Used by the long convolution of the reverb (reverb.py:31-38 via fft_convolve,
// filtered_noise.py:25-32) and by the STFT front end of the spectral loss (mss_loss.py:23-27), where the
// reference calls cuFFT through torch.fft / torch.stft.
//
//   * data: float2[n] ping-pong pair in shared memory, natural order in, natural order out;
//   * twiddles: float2 tw[n] in shared memory, tw[q] = (cos(2 pi q/n), -sin(2 pi q/n)), copied once per CTA
//     from a device table rounded from fp64 (fft_ensure_table must have run on the host side);
//   * a "group" of `gsize` threads (any divisor layout of the CTA) owns one transform; every pass ends
//     with a CTA-wide __syncthreads(), so all groups of a CTA must transform the same n together.
#pragma once
#include <cuda_runtime.h>

namespace ddsp {

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}
__device__ __forceinline__ float2 cmul_conj(float2 a, float2 b) {   // a * conj(b)
    return make_float2(fmaf(a.x, b.x, a.y * b.y), fmaf(a.y, b.x, -a.x * b.y));
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }

// Twiddle table for the largest transform (4096), rounded from fp64: tw[q] = (cos, -sin)(2 pi q / 4096).
// Exactly rounded twiddles keep the fp32 transform at the accuracy of a CPU fp32 FFT (rel. error 1.1e-7 at
// n = 2048; 1-ulp sincospif twiddles measured 2.0e-7), which the spectral-loss gradient is sensitive to.
// Each translation unit owns a copy, filled once per device by the first call that needs it.
constexpr int FFT_MAX = 4096;
static __device__ float2 g_fft_tw[FFT_MAX];
static __device__ double2 g_fft_tw64[FFT_MAX];          // same table in fp64, read straight from global/L1 by block_fft64

static __global__ void fft_init_table_kernel() {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q < FFT_MAX) {
        double sn, cs;
        sincospi(2.0 * (double)q / (double)FFT_MAX, &sn, &cs);
        g_fft_tw[q] = make_float2((float)cs, (float)-sn);
        g_fft_tw64[q] = make_double2(cs, -sn);
    }
}

static inline cudaError_t fft_ensure_table(cudaStream_t st) {
    static bool done[64] = {false};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64 || !done[dev]) {
        // first use on this device: the fill ends with a stream synchronise, which a capturing stream cannot take --
        // the FFT-using entry points must run once outside CUDA-graph capture (torch's graph warm-up does that)
        cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(st, &cap) == cudaSuccess && cap != cudaStreamCaptureStatusNone) return cudaErrorStreamCaptureUnsupported;
        fft_init_table_kernel<<<FFT_MAX / 256, 256, 0, st>>>();
        e = cudaStreamSynchronize(st);                  // one-time: later calls on other streams see the table
        if (e == cudaSuccess && dev >= 0 && dev < 64) done[dev] = true;
    }
    return e;
}

// copy the stride-(4096/n) subset of the table into shared memory: tw[q] = exp(-2 pi i q / n)
__device__ __forceinline__ void fft_fill_twiddles(float2* tw, int n) {
    const int stride = FFT_MAX / n;
    for (int q = threadIdx.x; q < n; q += blockDim.x) tw[q] = g_fft_tw[q * stride];
}

// One transform of size n by the `gsize` threads gtid = 0..gsize-1.  `a` holds the input; returns the buffer
// (a or b) that holds the output.  INV = true computes the unnormalised inverse (conjugated twiddles).
template <bool INV>
__device__ __forceinline__ float2* block_fft(float2* a, float2* b, const float2* __restrict__ tw, int n, int gtid, int gsize) {
    int p = 1;
    int log2n = 31 - __clz(n);
    if (log2n & 1) {                                   // radix-2 pass, p = 1 (no twiddle)
        const int t = n >> 1;
        for (int i = gtid; i < t; i += gsize) {
            const float2 u0 = a[i], u1 = a[i + t];
            b[2 * i] = cadd(u0, u1);
            b[2 * i + 1] = csub(u0, u1);
        }
        __syncthreads();
        float2* s = a; a = b; b = s;
        p = 2;
    }
    const int t = n >> 2;
    for (; p < n; p <<= 2) {
        const int tstep = n / (4 * p);                 // tw index step: w_{4p}^k = tw[k * n/(4p)]
        for (int i = gtid; i < t; i += gsize) {
            const int k = i & (p - 1);
            const int j = ((i - k) << 2) + k;
            float2 w1 = tw[k * tstep], w2 = tw[2 * k * tstep], w3 = tw[3 * k * tstep];
            if (INV) { w1.y = -w1.y; w2.y = -w2.y; w3.y = -w3.y; }
            const float2 u0 = a[i];
            const float2 u1 = cmul(a[i + t], w1);
            const float2 u2 = cmul(a[i + 2 * t], w2);
            const float2 u3 = cmul(a[i + 3 * t], w3);
            const float2 v0 = cadd(u0, u2), v1 = csub(u0, u2), v2 = cadd(u1, u3);
            const float2 d = csub(u1, u3);
            const float2 v3 = INV ? make_float2(-d.y, d.x) : make_float2(d.y, -d.x);   // (+/-)i * (u1-u3)
            b[j] = cadd(v0, v2);
            b[j + p] = cadd(v1, v3);
            b[j + 2 * p] = csub(v0, v2);
            b[j + 3 * p] = csub(v1, v3);
        }
        __syncthreads();
        float2* s = a; a = b; b = s;
    }
    return a;
}

// fp64 variant (same index algebra) for transforms whose small bins feed an ill-conditioned gradient (the
// log-spectral term of the loss).  Twiddles come from the global fp64 table, stride FFT_MAX/n.
__device__ __forceinline__ double2 cmul64(double2 a, double2 b) {
    return make_double2(fma(a.x, b.x, -a.y * b.y), fma(a.x, b.y, a.y * b.x));
}

__device__ __forceinline__ double2* block_fft64(double2* a, double2* b, int n, int gtid, int gsize) {
    int p = 1;
    const int log2n = 31 - __clz(n);
    const int tscale = FFT_MAX / n;
    if (log2n & 1) {
        const int t = n >> 1;
        for (int i = gtid; i < t; i += gsize) {
            const double2 u0 = a[i], u1 = a[i + t];
            b[2 * i] = make_double2(u0.x + u1.x, u0.y + u1.y);
            b[2 * i + 1] = make_double2(u0.x - u1.x, u0.y - u1.y);
        }
        __syncthreads();
        double2* s = a; a = b; b = s;
        p = 2;
    }
    const int t = n >> 2;
    for (; p < n; p <<= 2) {
        const int tstep = n / (4 * p) * tscale;
        for (int i = gtid; i < t; i += gsize) {
            const int k = i & (p - 1);
            const int j = ((i - k) << 2) + k;
            const double2 u0 = a[i];
            const double2 u1 = cmul64(a[i + t], g_fft_tw64[k * tstep]);
            const double2 u2 = cmul64(a[i + 2 * t], g_fft_tw64[2 * k * tstep]);
            const double2 u3 = cmul64(a[i + 3 * t], g_fft_tw64[3 * k * tstep]);
            const double2 v0 = make_double2(u0.x + u2.x, u0.y + u2.y), v1 = make_double2(u0.x - u2.x, u0.y - u2.y);
            const double2 v2 = make_double2(u1.x + u3.x, u1.y + u3.y);
            const double2 v3 = make_double2(u1.y - u3.y, -(u1.x - u3.x));      // -i * (u1-u3)
            b[j] = make_double2(v0.x + v2.x, v0.y + v2.y);
            b[j + p] = make_double2(v1.x + v3.x, v1.y + v3.y);
            b[j + 2 * p] = make_double2(v0.x - v2.x, v0.y - v2.y);
            b[j + 3 * p] = make_double2(v1.x - v3.x, v1.y - v3.y);
        }
        __syncthreads();
        double2* s = a; a = b; b = s;
    }
    return a;
}

}  // namespace ddsp
