// Register-tiled complex FFT for power-of-two sizes 8..4096 (sm_100a).
//
// A group of n/8 threads owns one transform; every thread holds 8 elements in registers,
//     in : v[q] = x[t + q*n/8]          out : v[q] = X[t + q*n/8]            (t = thread index inside the group)
// so global loads/stores of a group are coalesced on both sides.  Stockham autosort, decimation in time: radix-8
// butterflies in registers, one shared-memory exchange between stages (n = 8^a * {1,2,4}; the radix-2 / radix-4
// stage, if any, comes last).  Compared with the radix-4 shared-memory passes of ddsp_fft.cuh this halves the
// number of passes and keeps all butterfly arithmetic in registers (profiles/summary_r01d.md).
//
//   * exchange buffer: n complex values per group with a padded index (fft8_pad) -- with that padding the scattered
//     writes of every stage and the unit-stride reads are bank-conflict free;
//   * twiddles: per-stage tables T[q-1][k] = exp(-+2 pi i q k / (r p)) (k < p, q < r) in global memory, rounded
//     from fp64, read through L1 with unit stride across the group (tables are shared by all transforms on the SM); a radix-8
//     stage reads w, w^2, w^4 and forms the other four as one product each (DDSP_FFT8_DERIVED_TWIDDLES);
//   * T = float or double (the spectral loss runs its predicted frames in fp64, see ddsp_mss.cu).
#pragma once
#include <cuda_runtime.h>

#ifndef DDSP_FFT8_DERIVED_TWIDDLES
#define DDSP_FFT8_DERIVED_TWIDDLES 1
#endif

namespace ddsp {

template <typename T> struct cplx;
template <> struct cplx<float> { using type = float2; };
template <> struct cplx<double> { using type = double2; };

// padded exchange index: float2 (8 B) elements skip 2 slots every 16, double2 (16 B) elements 1 slot every 8
template <typename T> __host__ __device__ constexpr int fft8_pad(int i) {
    return sizeof(T) == 4 ? i + ((i >> 4) << 1) : i + (i >> 3);
}
__host__ __device__ constexpr int fft8_smem_elems(int n) { return n + (n >> 3) + 2; }   // either element type

// ---- stage plan -----------------------------------------------------------------------------------------------
// slot(p, r): p in {8, 64, 512} -> 0..2, r in {2, 4, 8} -> 0..2; each slot holds (r-1)*p <= 3584 twiddles
constexpr int FFT8_SLOT = 3584;
__host__ __device__ constexpr int fft8_slot(int p, int r) {
    return ((p == 8 ? 0 : (p == 64 ? 1 : 2)) * 3 + (r == 2 ? 0 : (r == 4 ? 1 : 2))) * FFT8_SLOT;
}
static __device__ float2 g_fft8_tw32[9 * FFT8_SLOT];
static __device__ double2 g_fft8_tw64[9 * FFT8_SLOT];

static __global__ void fft8_init_tables_kernel() {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;          // over 9 slots x FFT8_SLOT
    if (idx >= 9 * FFT8_SLOT) return;
    const int slot = idx / FFT8_SLOT, e = idx % FFT8_SLOT;
    const int p = slot / 3 == 0 ? 8 : (slot / 3 == 1 ? 64 : 512);
    const int r = slot % 3 == 0 ? 2 : (slot % 3 == 1 ? 4 : 8);
    double2 w = make_double2(0.0, 0.0);
    if (e < (r - 1) * p) {
        const int q = e / p + 1, k = e % p;
        double sn, cs;
        sincospi(2.0 * (double)((long long)q * k) / (double)(r * p), &sn, &cs);
        w = make_double2(cs, -sn);
    }
    g_fft8_tw64[idx] = w;
    g_fft8_tw32[idx] = make_float2((float)w.x, (float)w.y);
}

static inline cudaError_t fft8_ensure_tables(cudaStream_t st) {
    static bool done[64] = {false};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64 || !done[dev]) {
        // first use on this device: the fill ends with a stream synchronise, which a capturing stream cannot take --
        // the FFT-using entry points must run once outside CUDA-graph capture (torch's graph warm-up does that)
        cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(st, &cap) == cudaSuccess && cap != cudaStreamCaptureStatusNone) return cudaErrorStreamCaptureUnsupported;
        fft8_init_tables_kernel<<<(9 * FFT8_SLOT + 255) / 256, 256, 0, st>>>();
        e = cudaStreamSynchronize(st);                  // one-time: later calls on other streams see the tables
        if (e == cudaSuccess && dev >= 0 && dev < 64) done[dev] = true;
    }
    return e;
}

template <typename T> __device__ __forceinline__ const typename cplx<T>::type* fft8_table(int p, int r);
template <> __device__ __forceinline__ const float2* fft8_table<float>(int p, int r) { return g_fft8_tw32 + fft8_slot(p, r); }
template <> __device__ __forceinline__ const double2* fft8_table<double>(int p, int r) { return g_fft8_tw64 + fft8_slot(p, r); }

// ---- complex helpers ---------------------------------------------------------------------------------------------
template <typename C> __device__ __forceinline__ C c_add(C a, C b) { C r; r.x = a.x + b.x; r.y = a.y + b.y; return r; }
template <typename C> __device__ __forceinline__ C c_sub(C a, C b) { C r; r.x = a.x - b.x; r.y = a.y - b.y; return r; }
// a * w (forward) or a * conj(w) (inverse)
template <bool INV, typename C> __device__ __forceinline__ C c_mul_tw(C a, C w) {
    C r;
    if (INV) { r.x = fma(a.x, w.x, a.y * w.y); r.y = fma(a.y, w.x, -(a.x * w.y)); }
    else     { r.x = fma(a.x, w.x, -(a.y * w.y)); r.y = fma(a.x, w.y, a.y * w.x); }
    return r;
}
// multiply by -i (forward) / +i (inverse)
template <bool INV, typename C> __device__ __forceinline__ C c_rot(C a) {
    C r;
    if (INV) { r.x = -a.y; r.y = a.x; } else { r.x = a.y; r.y = -a.x; }
    return r;
}
// multiply by w8 = exp(-+ i pi/4) and w8^3 = exp(-+ 3 i pi/4)
template <bool INV, typename C> __device__ __forceinline__ C c_w8(C a) {
    const decltype(a.x) h = (decltype(a.x))0.70710678118654752440;
    C r;
    if (INV) { r.x = (a.x - a.y) * h; r.y = (a.x + a.y) * h; } else { r.x = (a.x + a.y) * h; r.y = (a.y - a.x) * h; }
    return r;
}
template <bool INV, typename C> __device__ __forceinline__ C c_w83(C a) {
    const decltype(a.x) h = (decltype(a.x))0.70710678118654752440;
    C r;
    if (INV) { r.x = (-a.x - a.y) * h; r.y = (a.x - a.y) * h; } else { r.x = (a.y - a.x) * h; r.y = (-a.x - a.y) * h; }
    return r;
}

template <bool INV, typename C> __device__ __forceinline__ void radix2(C& u0, C& u1) {
    const C a = c_add(u0, u1), b = c_sub(u0, u1);
    u0 = a; u1 = b;
}
template <bool INV, typename C> __device__ __forceinline__ void radix4(C& u0, C& u1, C& u2, C& u3) {
    const C a0 = c_add(u0, u2), a1 = c_sub(u0, u2), a2 = c_add(u1, u3), a3 = c_rot<INV>(c_sub(u1, u3));
    u0 = c_add(a0, a2); u2 = c_sub(a0, a2); u1 = c_add(a1, a3); u3 = c_sub(a1, a3);
}
template <bool INV, typename C> __device__ __forceinline__ void radix8(C (&v)[8]) {
    const C a0 = c_add(v[0], v[4]), a1 = c_sub(v[0], v[4]), a2 = c_add(v[2], v[6]), a3 = c_rot<INV>(c_sub(v[2], v[6]));
    const C a4 = c_add(v[1], v[5]), a5 = c_sub(v[1], v[5]), a6 = c_add(v[3], v[7]), a7 = c_rot<INV>(c_sub(v[3], v[7]));
    const C b0 = c_add(a0, a2), b2 = c_sub(a0, a2), b1 = c_add(a1, a3), b3 = c_sub(a1, a3);
    const C b4 = c_add(a4, a6), b6 = c_rot<INV>(c_sub(a4, a6)), b5 = c_w8<INV>(c_add(a5, a7)), b7 = c_w83<INV>(c_sub(a5, a7));
    v[0] = c_add(b0, b4); v[4] = c_sub(b0, b4); v[1] = c_add(b1, b5); v[5] = c_sub(b1, b5);
    v[2] = c_add(b2, b6); v[6] = c_sub(b2, b6); v[3] = c_add(b3, b7); v[7] = c_sub(b3, b7);
}

// ---- the transform ---------------------------------------------------------------------------------------------------
template <int N> struct Fft8Plan {
    static_assert(N >= 8 && N <= 4096 && (N & (N - 1)) == 0, "fft8: n must be a power of two in [8, 4096]");
    static constexpr int A = N >= 4096 ? 4 : (N >= 512 ? 3 : (N >= 64 ? 2 : 1));      // radix-8 stages
    static constexpr int P8 = A == 4 ? 4096 : (A == 3 ? 512 : (A == 2 ? 64 : 8));     // 8^A
    static constexpr int RL = N / P8;                                                 // last radix: 1 (none), 2 or 4
    static constexpr int T8 = N / 8;                                                  // threads per transform
};

// groups of <= 32 threads are warp-aligned (power-of-two sizes); larger groups use the CTA barrier, so every thread
// of the CTA must run the same transform size in lockstep
template <int N> __device__ __forceinline__ void fft8_sync() {
    if (N / 8 > 32) __syncthreads(); else __syncwarp();
}

// In-register transform of the group's N values (layout above).  sm: the group's exchange buffer, fft8_smem_elems(N)
// elements; the caller makes sure nobody still uses it (the function synchronises between its own exchanges only).
// INV = true: unnormalised inverse.
template <int N, typename T, bool INV>
__device__ __forceinline__ void fft8_group(typename cplx<T>::type (&v)[8], typename cplx<T>::type* sm, int t) {
    using C = typename cplx<T>::type;
    using P = Fft8Plan<N>;
#pragma unroll
    for (int s = 0; s < P::A; ++s) {
        const int p = 1 << (3 * s);
        const int k = t & (p - 1);
        if (s > 0) {
            const C* tb = fft8_table<T>(p, 8);
#if DDSP_FFT8_DERIVED_TWIDDLES
            // w^1, w^2, w^4 from the table, the other four as one product each of those (1 ulp): every table read is a
            // wavefront (fp32) or two (fp64) on the L1 data pipe, which is what bounds the transform kernels
            // (profiles/summary_r02.md section 9).  Measured against all seven from the table: c2c n = 2048 0.375 -> 0.355 ms,
            // n = 4096 0.154 -> 0.150 ms, spectral loss cfg 5 0.631 -> 0.626 ms, reverb unchanged.
            C w[8];
            w[1] = __ldg(tb + k); w[2] = __ldg(tb + p + k); w[4] = __ldg(tb + 3 * p + k);
            w[3] = c_mul_tw<false>(w[1], w[2]); w[5] = c_mul_tw<false>(w[1], w[4]);
            w[6] = c_mul_tw<false>(w[2], w[4]); w[7] = c_mul_tw<false>(w[3], w[4]);
#pragma unroll
            for (int q = 1; q < 8; ++q) v[q] = c_mul_tw<INV>(v[q], w[q]);
#else
#pragma unroll
            for (int q = 1; q < 8; ++q) v[q] = c_mul_tw<INV>(v[q], __ldg(tb + (q - 1) * p + k));
#endif
        }
        radix8<INV>(v);
        const bool last = (s == P::A - 1) && P::RL == 1;
        if (!last) {
            if (s > 0) fft8_sync<N>();                   // the previous exchange's reads are done
            const int j = ((t - k) << 3) + k;
#pragma unroll
            for (int q = 0; q < 8; ++q) sm[fft8_pad<T>(j + q * p)] = v[q];
            fft8_sync<N>();
            const int rn = (s + 1 < P::A) ? 8 : P::RL;   // radix of the next stage
            const int tn = N / rn;
#pragma unroll
            for (int u = 0; u < 8 / rn; ++u)
#pragma unroll
                for (int q = 0; q < rn; ++q) v[u * rn + q] = sm[fft8_pad<T>(t + u * P::T8 + q * tn)];
        }
    }
    if (P::RL > 1) {
        constexpr int p = P::P8, RL = P::RL;
        const C* tb = fft8_table<T>(p, RL);
        C o[8];
#pragma unroll
        for (int u = 0; u < 8 / RL; ++u) {
            const int i = t + u * P::T8;                 // butterfly index, < N/RL = p, hence k = i
#pragma unroll
            for (int q = 1; q < RL; ++q) v[u * RL + q] = c_mul_tw<INV>(v[u * RL + q], __ldg(tb + (q - 1) * p + i));
            if (RL == 2) radix2<INV>(v[2 * u], v[2 * u + 1]);
            else radix4<INV>(v[(4 * u) & 7], v[(4 * u + 1) & 7], v[(4 * u + 2) & 7], v[(4 * u + 3) & 7]);
#pragma unroll
            for (int q = 0; q < RL; ++q) o[u + q * (8 / RL)] = v[u * RL + q];
        }
#pragma unroll
        for (int q = 0; q < 8; ++q) v[q] = o[q];
    }
}

}  // namespace ddsp
