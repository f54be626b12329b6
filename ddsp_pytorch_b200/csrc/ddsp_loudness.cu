// Loudness feature for sm_100a -- replaces LoudnessEncoder.forward (model/autoencoder/encoder.py:131-156):
//   torch.stft(signal, n_fft, hop, center=False) with the default rectangular window -> 20 log10(|X| + 1e-20)
//   + A-weighting per bin -> /90 + 1 -> mean over the n_fft/2+1 bins.
// One group of n_fft/8 threads transforms a PAIR of consecutive frames as one complex FFT (registers, ddsp_fft8.cuh),
// splits it by Hermitian symmetry and reduces the weighted log-magnitudes; the spectrogram never reaches HBM.
#include "ddsp_common.cuh"
#include "ddsp_fft8.cuh"

namespace ddsp {

template <int N>
struct LoudGeom {
    static constexpr int T8 = N / 8;
    static constexpr int THREADS = T8 > 128 ? T8 : 128;
    static constexpr int G = THREADS / T8;                        // frame pairs per CTA
    static constexpr size_t group_floats = 2 * (size_t)fft8_smem_elems(N) + 32;
};

template <int N>
__global__ void __launch_bounds__(LoudGeom<N>::THREADS) loudness_kernel(const float* __restrict__ x, const float* __restrict__ aw,
                                                                       float* __restrict__ out, int Nsig, int hop, int F) {
    using P = LoudGeom<N>;
    constexpr int T8 = P::T8, half = N / 2;
    extern __shared__ __align__(16) float lraw[];
    const int tid = threadIdx.x, grp = tid / T8, t = tid % T8;
    float2* sm = reinterpret_cast<float2*>(lraw + (size_t)grp * P::group_floats);
    float* red = reinterpret_cast<float*>(sm + fft8_smem_elems(N));        // [2][16] per-warp frame sums (groups wider than a warp)
    const int b = blockIdx.y;
    const float* xb = x + (size_t)b * Nsig;
    const int fa = 2 * (blockIdx.x * P::G + grp), fb = fa + 1;
    const bool live_a = fa < F, live_b = fb < F;      // dead groups still walk through every barrier

    float2 v[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const int m = t + q * T8;
        v[q] = make_float2(live_a ? xb[(size_t)fa * hop + m] : 0.f, live_b ? xb[(size_t)fb * hop + m] : 0.f);
    }
    fft8_group<N, float, false>(v, sm, t);
    fft8_sync<N>();
#pragma unroll
    for (int q = 0; q < 8; ++q) sm[fft8_pad<float>(t + q * T8)] = v[q];
    fft8_sync<N>();
    float sa = 0.f, sb = 0.f;
#pragma unroll
    for (int q = 0; q < 5; ++q) {
        if (q == 4 && t != 0) continue;               // bin n/2 belongs to thread 0
        const int k = q < 4 ? t + q * T8 : half;
        const float2 zk = q < 4 ? v[q] : sm[fft8_pad<float>(half)];
        const float2 zn = sm[fft8_pad<float>((N - k) & (N - 1))];
        // Z = FFT(xa + i xb):  Xa_k = (Z_k + conj Z_{n-k})/2,  Xb_k = (Z_k - conj Z_{n-k})/(2i)
        const float ar = 0.5f * (zk.x + zn.x), ai = 0.5f * (zk.y - zn.y);
        const float br = 0.5f * (zk.y + zn.y), bi = -0.5f * (zk.x - zn.x);
        const float w = aw[k];
        sa += (log10f(sqrtf(ar * ar + ai * ai) + 1e-20f) * 20.f + w) / 90.f + 1.f;      // encoder.py:148-151
        sb += (log10f(sqrtf(br * br + bi * bi) + 1e-20f) * 20.f + w) / 90.f + 1.f;
    }
    // sum over the group's threads
    constexpr int W = T8 < 32 ? T8 : 32;
#pragma unroll
    for (int o = W / 2; o > 0; o >>= 1) {
        sa += __shfl_xor_sync(0xffffffffu, sa, o);
        sb += __shfl_xor_sync(0xffffffffu, sb, o);
    }
    if (T8 > 32) {                                    // fixed-order sum of the warps' partials
        if ((t & 31) == 0) { red[t >> 5] = sa; red[16 + (t >> 5)] = sb; }
        __syncthreads();
        if (t == 0) {
            sa = sb = 0.f;
            for (int w = 0; w < T8 / 32; ++w) { sa += red[w]; sb += red[16 + w]; }
        }
    }
    if (t == 0) {
        const float inv = 1.0f / (float)(half + 1);
        if (live_a) out[(size_t)b * F + fa] = sa * inv;
        if (live_b) out[(size_t)b * F + fb] = sb * inv;
    }
}

template <int N>
static int launch_loudness(const float* x, const float* aw, float* out, int B, int Nsig, int hop, int F, cudaStream_t st) {
    using P = LoudGeom<N>;
    const size_t smem = sizeof(float) * P::group_floats * P::G;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(loudness_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        DDSP_REQUIRE(e == cudaSuccess, "loudness_kernel: cannot reserve %zu B of shared memory: %s", smem, cudaGetErrorString(e));
    }
    const int per_cta = 2 * P::G;
    loudness_kernel<N><<<dim3((F + per_cta - 1) / per_cta, B), P::THREADS, smem, st>>>(x, aw, out, Nsig, hop, F);
    DDSP_CHECK_LAUNCH("loudness_kernel");
    return 0;
}

}  // namespace ddsp

using namespace ddsp;

extern "C" int ddsp_loudness_frames(int N, int n_fft, int hop) {
    if (N < n_fft || n_fft < 1 || hop < 1) return 0;
    return 1 + (N - n_fft) / hop;                       // torch.stft, center=False
}

extern "C" int ddsp_loudness_forward(const float* x, const float* a_weight, float* out, int B, int N, int n_fft, int hop,
                                     void* stream) {
    DDSP_REQUIRE(x && a_weight && out, "ddsp_loudness_forward: null pointer argument");
    DDSP_REQUIRE(B >= 1 && B <= 65535 && hop >= 1, "ddsp_loudness_forward: bad B=%d / hop=%d", B, hop);
    DDSP_REQUIRE(n_fft >= 8 && n_fft <= 4096 && (n_fft & (n_fft - 1)) == 0,
                 "ddsp_loudness_forward: n_fft=%d must be a power of two in [8,4096]", n_fft);
    DDSP_REQUIRE(N >= n_fft, "ddsp_loudness_forward: signal of %d samples is shorter than n_fft=%d (center=False)", N, n_fft);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    cudaError_t e = fft8_ensure_tables(st);
    DDSP_REQUIRE(e == cudaSuccess, "ddsp_loudness_forward: twiddle table initialisation failed: %s", cudaGetErrorString(e));
    const int F = ddsp_loudness_frames(N, n_fft, hop);
    switch (n_fft) {
#define LOUD_CASE(NN) case NN: return launch_loudness<NN>(x, a_weight, out, B, N, hop, F, st)
        LOUD_CASE(8); LOUD_CASE(16); LOUD_CASE(32); LOUD_CASE(64); LOUD_CASE(128); LOUD_CASE(256);
        LOUD_CASE(512); LOUD_CASE(1024); LOUD_CASE(2048); LOUD_CASE(4096);
#undef LOUD_CASE
    }
    return 1;
}
