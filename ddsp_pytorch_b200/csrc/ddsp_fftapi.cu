// Batched complex FFT entry point over the register-tiled transform of ddsp_fft8.cuh.  The reference reaches cuFFT
// through torch.fft (filtered_noise.py:10,29) and torch.stft (mss_loss.py:23-27); this is the transform underneath
// the reverb and spectral-loss kernels, exported on its own so that it can be held against a reference FFT.
#include "ddsp_common.cuh"
#include "ddsp_fft8.cuh"

namespace ddsp {

template <int N, typename T, bool INV>
__global__ void __launch_bounds__((N / 8 > 128 ? N / 8 : 128)) fft_c2c_kernel(const typename cplx<T>::type* __restrict__ in,
                                                                             typename cplx<T>::type* __restrict__ out,
                                                                             long long batch) {
    using C = typename cplx<T>::type;
    extern __shared__ __align__(16) unsigned char fraw[];
    constexpr int T8 = N / 8, THREADS = T8 > 128 ? T8 : 128, G = THREADS / T8;
    const int grp = threadIdx.x / T8, t = threadIdx.x % T8;
    C* sm = reinterpret_cast<C*>(fraw) + (size_t)grp * fft8_smem_elems(N);
    const long long xf = (long long)blockIdx.x * G + grp;
    const bool live = xf < batch;                       // dead groups still walk through the barriers
    C v[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        v[q].x = 0; v[q].y = 0;
        if (live) v[q] = in[xf * N + t + q * T8];
    }
    fft8_group<N, T, INV>(v, sm, t);
    if (live) {
#pragma unroll
        for (int q = 0; q < 8; ++q) out[xf * N + t + q * T8] = v[q];
    }
}

template <int N, typename T>
static int launch_c2c(const void* in, void* out, long long batch, int inverse, cudaStream_t st) {
    using C = typename cplx<T>::type;
    constexpr int T8 = N / 8, THREADS = T8 > 128 ? T8 : 128, G = THREADS / T8;
    const size_t smem = sizeof(C) * fft8_smem_elems(N) * G;
    const unsigned grid = (unsigned)((batch + G - 1) / G);
    if (inverse) {
        if (smem > 48 * 1024) cudaFuncSetAttribute(fft_c2c_kernel<N, T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        fft_c2c_kernel<N, T, true><<<grid, THREADS, smem, st>>>(static_cast<const C*>(in), static_cast<C*>(out), batch);
    } else {
        if (smem > 48 * 1024) cudaFuncSetAttribute(fft_c2c_kernel<N, T, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        fft_c2c_kernel<N, T, false><<<grid, THREADS, smem, st>>>(static_cast<const C*>(in), static_cast<C*>(out), batch);
    }
    DDSP_CHECK_LAUNCH("fft_c2c_kernel");
    return 0;
}

template <typename T>
static int dispatch_c2c(const void* in, void* out, int n, long long batch, int inverse, cudaStream_t st) {
    switch (n) {
        case 8: return launch_c2c<8, T>(in, out, batch, inverse, st);
        case 16: return launch_c2c<16, T>(in, out, batch, inverse, st);
        case 32: return launch_c2c<32, T>(in, out, batch, inverse, st);
        case 64: return launch_c2c<64, T>(in, out, batch, inverse, st);
        case 128: return launch_c2c<128, T>(in, out, batch, inverse, st);
        case 256: return launch_c2c<256, T>(in, out, batch, inverse, st);
        case 512: return launch_c2c<512, T>(in, out, batch, inverse, st);
        case 1024: return launch_c2c<1024, T>(in, out, batch, inverse, st);
        case 2048: return launch_c2c<2048, T>(in, out, batch, inverse, st);
        case 4096: return launch_c2c<4096, T>(in, out, batch, inverse, st);
    }
    set_error("ddsp_fft_c2c: n=%d must be a power of two in [8,4096]", n);
    return 1;
}

}  // namespace ddsp

using namespace ddsp;

extern "C" int ddsp_fft_c2c(const void* in, void* out, int n, long long batch, int inverse, int fp64, void* stream) {
    DDSP_REQUIRE(in && out && batch >= 1, "ddsp_fft_c2c: bad arguments");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    cudaError_t e = fft8_ensure_tables(st);
    DDSP_REQUIRE(e == cudaSuccess, "ddsp_fft_c2c: twiddle table initialisation failed: %s", cudaGetErrorString(e));
    return fp64 ? dispatch_c2c<double>(in, out, n, batch, inverse, st) : dispatch_c2c<float>(in, out, n, batch, inverse, st);
}
