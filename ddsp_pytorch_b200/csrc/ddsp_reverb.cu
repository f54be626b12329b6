// Learned reverb for sm_100a -- replaces Reverb (reverb.py:9-49) and the long-kernel use of fft_convolve
// (filtered_noise.py:25-32).
//
// The reference pads signal and kernel to 2N and runs three monolithic FFTs of size 2N = 176128..384000
// (non power of two) per call.  Here the causal, truncated convolution y[n] = sum_q ir[q] x[n-q] is a
// uniformly partitioned overlap-save convolution with P = 2048-sample blocks and 4096-point complex FFTs
// that live entirely in shared memory:
//   conv_spectra_kernel   X[b2,i,:] = FFT_4096(x[(i-1)P .. (i+1)P))            one CTA per block i (register-tiled FFT)
//   conv_mac_kernel       Y[b2,i,k] = sum_j H[j,k] X[b2,i-j,k]                  one thread per (bin, voice pair), register ring
//   conv_ifft_kernel      y block i = last P samples of IFFT(Y[b2,i])
// Two voices share one complex transform (z = x_a + i x_b): convolution with a REAL kernel is real-linear,
// so the inverse transform returns y_a + i y_b with no split/merge pass at all.
// Backward: gx is the same convolution run on the time-reversed gradient; gir is a frequency-domain
// cross-correlation sum over voices and blocks (conv_gir_kernel -> deterministic two-stage reduction ->
// conv_gir_finish_kernel); build_impulse has its own forward/backward kernels.
#include "ddsp_common.cuh"
#include "ddsp_fft8.cuh"

namespace ddsp {

constexpr int CONV_P = 2048;            // new samples per block
constexpr int CONV_N2 = 2 * CONV_P;     // FFT size
constexpr int CONV_THREADS = 512;
constexpr int GIR_CHUNKS = 32;          // partial sums over voice pairs kept apart for a fixed-order reduction
constexpr int GIR_NP = 24;             // IR partitions handled per pass of the gir kernel

struct ConvPlan {
    int B2, nblk, np;
    size_t off_H, off_S0, off_S1, off_part, total;
};

static ConvPlan conv_plan(int B, int N, int Lk) {
    ConvPlan p;
    p.B2 = (B + 1) / 2;
    p.nblk = (N + CONV_P - 1) / CONV_P;
    const int Le = Lk < N ? Lk : N;                     // taps at lags >= N never reach an output (reverb.py:34 crops)
    p.np = (Le + CONV_P - 1) / CONV_P;
    const size_t spec = sizeof(float2) * (size_t)p.B2 * p.nblk * CONV_N2;
    p.off_H = 0;
    p.off_S0 = align_up(sizeof(float2) * (size_t)p.np * CONV_N2, 256);
    p.off_S1 = p.off_S0 + align_up(spec, 256);
    p.off_part = p.off_S1 + align_up(spec, 256);
    p.total = p.off_part + align_up(sizeof(float2) * (size_t)GIR_CHUNKS * p.np * CONV_N2, 256);
    return p;
}

// ------------------------------------------------------------------------------------------------
// build_impulse (reverb.py:24-29) and its backward
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float softplus_f(float x) { return x > 20.f ? x : log1pf(expf(x)); }   // torch threshold
__device__ __forceinline__ float sigmoid_f(float x) { return 1.f / (1.f + expf(-x)); }

__global__ void reverb_impulse_kernel(const float* __restrict__ noise, const float* __restrict__ t, const float* __restrict__ decay,
                                      const float* __restrict__ wet, float* __restrict__ ir, int L) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L) return;
    const float sp = softplus_f(-decay[0]), sg = sigmoid_f(wet[0]);
    ir[i] = i == 0 ? 1.f : noise[i] * expf(-sp * t[i] * 500.f) * sg;
}

__global__ void __launch_bounds__(1024) reverb_impulse_bwd_kernel(const float* __restrict__ gir, const float* __restrict__ noise,
                                                                  const float* __restrict__ t, const float* __restrict__ decay,
                                                                  const float* __restrict__ wet, float* __restrict__ gnoise,
                                                                  float* __restrict__ gdecay, float* __restrict__ gwet, int L) {
    __shared__ double red[2][32];
    const float sp = softplus_f(-decay[0]), sg = sigmoid_f(wet[0]), sgd = sigmoid_f(-decay[0]);
    double aw = 0.0, ad = 0.0;
    constexpr int U = 8;                                  // loads of U strides are issued together (one CTA: latency bound)
    for (int i0 = threadIdx.x; i0 < L; i0 += U * blockDim.x) {
        float tv[U], gv[U], nv[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = i0 + u * blockDim.x;
            const bool in = i < L;
            tv[u] = in ? t[i] : 0.f;
            gv[u] = in ? gir[i] : 0.f;
            nv[u] = in ? noise[i] : 0.f;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = i0 + u * blockDim.x;
            if (i >= L) continue;
            float gn = 0.f;
            if (i > 0) {                                  // ir[0] is overwritten with 1 (:28): no gradient
                const float env = expf(-sp * tv[u] * 500.f), g = gv[u];
                gn = g * env * sg;
                const float gne = g * nv[u] * env;
                aw += (double)(gne * sg * (1.f - sg));
                ad += (double)(gne * sg * tv[u] * 500.f * sgd);
            }
            gnoise[i] = gn;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        aw += __shfl_xor_sync(0xffffffffu, aw, o);
        ad += __shfl_xor_sync(0xffffffffu, ad, o);
    }
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = aw; red[1][threadIdx.x >> 5] = ad; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double w = 0.0, d = 0.0;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) { w += red[0][i]; d += red[1][i]; }
        gwet[0] = (float)w;
        gdecay[0] = (float)d;
    }
}

// ------------------------------------------------------------------------------------------------
// block spectra.  Segment of block i holds samples n = (i-1)P + m, m = 0..2P-1.
//   halves = 3: both halves (overlap-save input);  2: second half only ([0, g_i], adjoint of "keep last P");
//   halves = 1 with block index i = j+1: first half only ([h_j, 0], an IR partition; single real signal).
// reverse != 0 reads the signals back to front (for gx).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(CONV_THREADS, 3) conv_spectra_kernel(const float* __restrict__ x, float2* __restrict__ S, int B,
                                                                    int N, int nblk, int halves, int reverse, int blk_shift) {
    extern __shared__ __align__(16) float2 fsm[];      // exchange buffer of the register-tiled FFT
    const int i = blockIdx.x, b2 = blockIdx.y, tid = threadIdx.x;
    const int v0 = 2 * b2, v1 = 2 * b2 + 1;
    const float* xa = x + (size_t)v0 * N;
    const float* xb = v1 < B ? x + (size_t)v1 * N : nullptr;
    const long long base = (long long)(i + blk_shift - 1) * CONV_P;
    float2 v[8];
    if (base >= 0 && base + CONV_N2 <= N) {
        // segment entirely inside the signal (all but the first and last block): plain strided loads, q < 4 is the first half
        const float* pa = reverse ? xa + (N - 1 - base) - tid : xa + base + tid;
        const float* pb = xb ? (reverse ? xb + (N - 1 - base) - tid : xb + base + tid) : nullptr;
        const int step = reverse ? -CONV_THREADS : CONV_THREADS;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const bool on = (q < 4) ? (halves & 1) : (halves & 2);
            v[q] = on ? make_float2(pa[q * step], pb ? pb[q * step] : 0.f) : make_float2(0.f, 0.f);
        }
    } else {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int m = tid + q * CONV_THREADS;
            const bool on = (m < CONV_P) ? (halves & 1) : (halves & 2);
            const long long n = base + m;
            float re = 0.f, im = 0.f;
            if (on && n >= 0 && n < N) {
                const long long idx = reverse ? (N - 1 - n) : n;
                re = xa[idx];
                if (xb) im = xb[idx];
            }
            v[q] = make_float2(re, im);
        }
    }
    fft8_group<CONV_N2, float, false>(v, fsm, tid);
    float2* dst = S + ((size_t)b2 * nblk + i) * CONV_N2;
#pragma unroll
    for (int q = 0; q < 8; ++q) dst[tid + q * CONV_THREADS] = v[q];
}

// Y[b2,i,k] (+)= sum_{j<MAC_NP} H[jg+j,k] X[b2,i-jg-j,k].  Two adjacent lanes share a (frequency bin, voice pair): lane
// parity s takes the partitions jg + s*MAC_NPT + [0, MAC_NPT) and walks the blocks in order with the last MAC_NPT input
// spectra of ITS lag range in a register ring (static indices: the walk is unrolled MAC_NPT blocks at a time); the two
// partial sums meet in one shuffle.  Every H value is read once, every X value twice (the second time MAC_NPT blocks
// later, from L2).  One thread per bin with all 24 partitions needed 199 registers (8 warps per SM) and ptxas serialised
// its accumulation chains: 42 % FMA-pipe utilisation, 38 % of warp samples in fixed-latency wait (profiles/summary_r01g.md).
constexpr int MAC_NP = 24;
constexpr int MAC_NPT = MAC_NP / 2;
constexpr int MAC_THREADS = 128;
constexpr int MAC_BINS = MAC_THREADS / 2;          // bins per CTA of conv_mac_kernel

// 8-byte asynchronous copy global -> shared; nbytes = 0 writes zeros without touching src (LDGSTS with zero fill)
__device__ __forceinline__ void cp_async8(void* dst_smem, const void* src, int nbytes) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(d), "l"(src), "r"(nbytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

constexpr int MAC_STAGES = 2;          // rounds staged in shared memory: the one being consumed + one in flight
constexpr size_t MAC_SMEM = sizeof(float2) * MAC_STAGES * MAC_NPT * MAC_THREADS;

// One round of the walk = MAC_NPT consecutive blocks i0 .. i0+MAC_NPT-1.  prev / cur hold this lane's input spectra of the
// previous / current round (cur[u] = X[i0 + u - jb], read from the thread's own staging slots), so output u needs cur[u],
// cur[u-1], .., cur[0], prev[NPT-1], .., prev[u+1] -- all static register indices.
// Packed arithmetic: with x = (xr, xi) as it lies in its register pair, A += x * (hr, hr) and B += x * (hi, hi) are one
// FFMA2 each and the product is (A.x - B.y, A.y + B.x): two issue slots per complex multiply-add instead of four (the FMA
// pipe does the same work either way; the kernel was issue-bound at 81 % of its slots with 58 % of the pipe).
__device__ __forceinline__ void mac_round(const float2 (&prev)[MAC_NPT], float2 (&cur)[MAC_NPT], const float2 (&hrr)[MAC_NPT],
                                          const float2 (&hii)[MAC_NPT], const float2* __restrict__ slot, float* __restrict__ yf,
                                          int i0, int nblk, int s, int accumulate) {
#pragma unroll
    for (int u = 0; u < MAC_NPT; ++u) cur[u] = slot[u * MAC_THREADS];
#pragma unroll
    for (int u = 0; u < MAC_NPT; ++u) {
        const int i = i0 + u;                          // warp-uniform
        if (i < nblk) {
            float2 A0 = make_float2(0.f, 0.f), A1 = A0, B0 = A0, B1 = A0;           // four independent packed chains
#pragma unroll
            for (int j = 0; j < MAC_NPT; j += 2) {
                const float2 x0 = (u - j >= 0) ? cur[(u - j + MAC_NPT) % MAC_NPT] : prev[(u - j + MAC_NPT) % MAC_NPT];
                const float2 x1 = (u - j - 1 >= 0) ? cur[(u - j - 1 + MAC_NPT) % MAC_NPT] : prev[(u - j - 1 + MAC_NPT) % MAC_NPT];
                A0 = __ffma2_rn(x0, hrr[j], A0);
                B0 = __ffma2_rn(x0, hii[j], B0);
                A1 = __ffma2_rn(x1, hrr[j + 1], A1);
                B1 = __ffma2_rn(x1, hii[j + 1], B1);
            }
            // lane s keeps component s of the sum over both lanes' partitions and stores that float: no divergent store
            const float re = (A0.x + A1.x) - (B0.y + B1.y), im = (A0.y + A1.y) + (B0.x + B1.x);
            float mine = (s ? im : re) + __shfl_xor_sync(0xffffffffu, s ? re : im, 1);
            float* dst = yf + (size_t)i * (2 * CONV_N2);
            if (accumulate) mine += *dst;
            *dst = mine;
        }
    }
}

__global__ void __launch_bounds__(MAC_THREADS, 5) conv_mac_kernel(const float2* __restrict__ S, const float2* __restrict__ Hf,
                                                                  float2* __restrict__ Y, int nblk, int np, int jg, int accumulate) {
    extern __shared__ __align__(16) float2 mac_sm[];   // [MAC_STAGES][MAC_NPT][MAC_THREADS]
    const int s = threadIdx.x & 1;
    const int k = blockIdx.x * MAC_BINS + (threadIdx.x >> 1), b2 = blockIdx.y;
    const int jb = jg + s * MAC_NPT;               // first partition of this lane
    const float2* xs = S + (size_t)b2 * nblk * CONV_N2 + k;
    float* yf = reinterpret_cast<float*>(Y + (size_t)b2 * nblk * CONV_N2 + k) + s;
    float2* mine = mac_sm + threadIdx.x;
    const int R = (nblk - jg + MAC_NPT - 1) / MAC_NPT;

    // the next round's spectra travel through the thread's own shared-memory slots (cp.async, zero fill at the edges) while
    // the current round is multiplied: no registers and no barrier for the loads in flight.  History: loads issued and
    // consumed in the same step 0.36 ms; register look-ahead with three rotating sets 0.265 ms (128 registers, 4 CTAs/SM)
    auto issue = [&](int r) {
        if (r < R) {
            const int i0 = jg + r * MAC_NPT;
            const float2* xr = xs + (long long)(i0 - jb) * CONV_N2;      // a zero-byte copy never reads its source
            float2* slot = mine + (size_t)(r % MAC_STAGES) * (MAC_NPT * MAC_THREADS);
#pragma unroll
            for (int u = 0; u < MAC_NPT; ++u) {
                const int in = i0 + u;
                cp_async8(slot + u * MAC_THREADS, xr + (size_t)u * CONV_N2, (in < nblk && in >= jb) ? 8 : 0);
            }
        }
        cp_async_commit();
    };
    issue(0);
    float2 hrr[MAC_NPT], hii[MAC_NPT], ra[MAC_NPT], rb[MAC_NPT];
#pragma unroll
    for (int j = 0; j < MAC_NPT; ++j) {
        const float2 h = (jb + j < np) ? Hf[(size_t)(jb + j) * CONV_N2 + k] : make_float2(0.f, 0.f);
        hrr[j] = make_float2(h.x, h.x);
        hii[j] = make_float2(h.y, h.y);
        ra[j] = rb[j] = make_float2(0.f, 0.f);
    }
    for (int r = 0; r < R; r += 2) {
        issue(r + 1);
        cp_async_wait<MAC_STAGES - 1>();
        mac_round(ra, rb, hrr, hii, mine + (size_t)(r % MAC_STAGES) * (MAC_NPT * MAC_THREADS), yf, jg + r * MAC_NPT, nblk, s, accumulate);
        if (r + 1 >= R) break;
        issue(r + 2);
        cp_async_wait<MAC_STAGES - 1>();
        mac_round(rb, ra, hrr, hii, mine + (size_t)((r + 1) % MAC_STAGES) * (MAC_NPT * MAC_THREADS), yf, jg + (r + 1) * MAC_NPT, nblk, s,
                  accumulate);
    }
    cp_async_wait<0>();
}

// y block i of voice pair b2 = last P samples of IFFT(Y[b2,i]) / 4096
__global__ void __launch_bounds__(CONV_THREADS) conv_ifft_kernel(const float2* __restrict__ Y, float* __restrict__ y, int B, int N,
                                                                 int nblk, int reverse) {
    extern __shared__ __align__(16) float2 fsm[];
    const int i = blockIdx.x, b2 = blockIdx.y, tid = threadIdx.x;
    const float2* src = Y + ((size_t)b2 * nblk + i) * CONV_N2;
    float2 v[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) v[q] = src[tid + q * CONV_THREADS];
    fft8_group<CONV_N2, float, true>(v, fsm, tid);
    const float scale = 1.0f / (float)CONV_N2;
    const int v0 = 2 * b2, v1 = 2 * b2 + 1;
#pragma unroll
    for (int q = 4; q < 8; ++q) {
        const long long n = (long long)i * CONV_P + (tid + (q - 4) * CONV_THREADS);
        if (n < N) {
            const long long idx = reverse ? (N - 1 - n) : n;
            y[(size_t)v0 * N + idx] = v[q].x * scale;
            if (v1 < B) y[(size_t)v1 * N + idx] = v[q].y * scale;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// gradient of the kernel:  gir[jP+q] = Re IFFT( sum_{b2,i} conj(X[b2,i-j]) * Gt[b2,i] )[q] / 4096,  q < P
// Two adjacent lanes share a frequency bin: lane parity s accumulates the partitions jg + s*GIR_NPT + [0, GIR_NPT).  A CTA
// (bin tile, voice-pair chunk) walks its voice pairs and blocks in rounds of GIR_NPT blocks; the lane's input spectra of the
// previous / current / next round sit in three register sets that take turns (static indices, no copies), the gradient
// spectra of the next round are requested a round ahead.  Chunk partials are summed in a fixed order by the finish kernel.
// ------------------------------------------------------------------------------------------------
constexpr int GIR_NPT = GIR_NP / 2;

constexpr int GIR_STAGES = 2;          // rounds staged in shared memory: the one being consumed + one in flight (48 KB: 4 CTAs per SM)
constexpr size_t GIR_SMEM = sizeof(float2) * GIR_STAGES * 2 * GIR_NPT * MAC_THREADS;

// One round: cur[u] = X[i0 + u - jb] and g[u] = G[i0 + u] come out of the thread's own staging slots;
// acc[j] += g[u] * conj(X[i0 + u - jb - j]) with X[.. < round start] in prev (the previous round's cur).
// Packed: with x = (xr, xi) in its register pair, A[j] += (gx, gx) * x and B[j] += (gy, gy) * x are one FFMA2 each and
// g * conj(x) accumulates as (A.x + B.y, B.x - A.y) -- combined once, at the end of the kernel.
__device__ __forceinline__ void gir_round(const float2 (&prev)[GIR_NPT], float2 (&cur)[GIR_NPT], float2 (&A)[GIR_NPT],
                                          float2 (&Bc)[GIR_NPT], const float2* __restrict__ slot) {
    float2 g[GIR_NPT];
#pragma unroll
    for (int u = 0; u < GIR_NPT; ++u) {
        cur[u] = slot[u * MAC_THREADS];
        g[u] = slot[(GIR_NPT + u) * MAC_THREADS];
    }
#pragma unroll
    for (int u = 0; u < GIR_NPT; ++u) {
        const float2 gxx = make_float2(g[u].x, g[u].x), gyy = make_float2(g[u].y, g[u].y);   // zero past the last block
#pragma unroll
        for (int j = 0; j < GIR_NPT; ++j) {
            const float2 xv = (u - j >= 0) ? cur[(u - j + GIR_NPT) % GIR_NPT] : prev[(u - j + GIR_NPT) % GIR_NPT];
            A[j] = __ffma2_rn(gxx, xv, A[j]);
            Bc[j] = __ffma2_rn(gyy, xv, Bc[j]);
        }
    }
}

// The walk over (voice pair, round of GIR_NPT blocks) is one flat sequence of rounds.  Every thread copies the 2 x GIR_NPT
// spectra values of a round into its own shared-memory slots with cp.async, GIR_STAGES - 1 rounds ahead of the one it
// multiplies: the loads in flight cost no registers and no barrier (a slot is only ever read by the thread that filled it).
// With the look-ahead held in registers (one round, 168 registers) the kernel spent 53 % of its warp samples on the long
// scoreboard (profiles/summary_r01g.md).
__global__ void __launch_bounds__(MAC_THREADS, 4) conv_gir_kernel(const float2* __restrict__ Xf, const float2* __restrict__ Gf,
                                                                  float2* __restrict__ part, int B2, int nblk, int np, int jg) {
    extern __shared__ __align__(16) float2 gir_sm[];   // [GIR_STAGES][2 * GIR_NPT][MAC_THREADS]
    const int s = threadIdx.x & 1;
    const int k = blockIdx.x * MAC_BINS + (threadIdx.x >> 1);
    const int jb = jg + s * GIR_NPT;
    const int c = blockIdx.y, nchunks = gridDim.y;
    const int per = (B2 + nchunks - 1) / nchunks;
    const int b_lo = c * per, b_hi = min(B2, b_lo + per);
    const int rpp = (nblk - jg + GIR_NPT - 1) / GIR_NPT;                 // rounds per voice pair
    const int R = (b_hi > b_lo ? b_hi - b_lo : 0) * rpp;                 // rounds of this CTA
    float2* mine = gir_sm + threadIdx.x;

    auto issue = [&](int r) {                          // request round r into stage r % GIR_STAGES (an empty group past the end)
        if (r < R) {
            const int b2 = b_lo + r / rpp, i0 = jg + (r % rpp) * GIR_NPT;
            // round base addresses; a zero-byte copy forms its source address but never reads it (lane 1's first rounds
            // point in front of the pair's first block)
            const long long row0 = (long long)b2 * nblk + i0;
            const float2* xr = Xf + (row0 - jb) * CONV_N2 + k;
            const float2* gr = Gf + row0 * CONV_N2 + k;
            float2* slot = mine + (size_t)(r % GIR_STAGES) * (2 * GIR_NPT * MAC_THREADS);
#pragma unroll
            for (int u = 0; u < GIR_NPT; ++u) {
                const int in = i0 + u;
                cp_async8(slot + u * MAC_THREADS, xr + (size_t)u * CONV_N2, (in < nblk && in >= jb) ? 8 : 0);
                cp_async8(slot + (GIR_NPT + u) * MAC_THREADS, gr + (size_t)u * CONV_N2, in < nblk ? 8 : 0);
            }
        }
        cp_async_commit();
    };

    float2 accA[GIR_NPT], accB[GIR_NPT], ra[GIR_NPT], rb[GIR_NPT];
#pragma unroll
    for (int j = 0; j < GIR_NPT; ++j) accA[j] = accB[j] = ra[j] = rb[j] = make_float2(0.f, 0.f);
#pragma unroll
    for (int r = 0; r < GIR_STAGES - 1; ++r) issue(r);
    for (int r = 0; r < R; r += 2) {
        // even round of the pair: previous spectra in ra, current into rb; odd round: the other way round
        issue(r + GIR_STAGES - 1);
        cp_async_wait<GIR_STAGES - 1>();
        if (r % rpp == 0) {
#pragma unroll
            for (int j = 0; j < GIR_NPT; ++j) ra[j] = make_float2(0.f, 0.f);      // a new voice pair starts with an empty history
        }
        gir_round(ra, rb, accA, accB, mine + (size_t)(r % GIR_STAGES) * (2 * GIR_NPT * MAC_THREADS));
        if (r + 1 >= R) break;
        issue(r + GIR_STAGES);
        cp_async_wait<GIR_STAGES - 1>();
        if ((r + 1) % rpp == 0) {
#pragma unroll
            for (int j = 0; j < GIR_NPT; ++j) rb[j] = make_float2(0.f, 0.f);
        }
        gir_round(rb, ra, accA, accB, mine + (size_t)((r + 1) % GIR_STAGES) * (2 * GIR_NPT * MAC_THREADS));
    }
    cp_async_wait<0>();
#pragma unroll
    for (int j = 0; j < GIR_NPT; ++j)
        if (jb + j < np) part[((size_t)c * np + jb + j) * CONV_N2 + k] = make_float2(accA[j].x + accB[j].y, accB[j].x - accA[j].y);
}

__global__ void __launch_bounds__(CONV_THREADS) conv_gir_finish_kernel(const float2* __restrict__ part, float* __restrict__ gir,
                                                                       int np, int Le, int nchunks) {
    extern __shared__ __align__(16) float2 fsm[];
    const int j = blockIdx.x, tid = threadIdx.x;
    float2 v[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const int k = tid + q * CONV_THREADS;
        float sx = 0.f, sy = 0.f;
#pragma unroll 4
        for (int c = 0; c < nchunks; ++c) {
            const float2 p = part[((size_t)c * np + j) * CONV_N2 + k];
            sx += p.x;
            sy += p.y;
        }
        v[q] = make_float2(sx, sy);
    }
    fft8_group<CONV_N2, float, true>(v, fsm, tid);
    const float scale = 1.0f / (float)CONV_N2;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int idx = j * CONV_P + tid + q * CONV_THREADS;
        if (idx < Le) gir[idx] = v[q].x * scale;
    }
}

// ------------------------------------------------------------------------------------------------
// live_forward (reverb.py:40-49): slide the buffer, then only the last n outputs of the convolution
// ------------------------------------------------------------------------------------------------
__global__ void copy_kernel(const float* __restrict__ src, float* __restrict__ dst, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[i];
}

__global__ void __launch_bounds__(256) reverb_live_dot_kernel(const float* __restrict__ buf, const float* __restrict__ ir,
                                                              float* __restrict__ y, int L, int n) {
    __shared__ float red[8];
    const int m = blockIdx.x;
    const int p = L - n + m;                           // output position inside the buffer
    float acc = 0.f;
    for (int q = threadIdx.x; q <= p; q += 256) acc = fmaf(ir[q], buf[p - q], acc);
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.f;
        for (int w = 0; w < 8; ++w) s += red[w];
        y[m] = s;
    }
}

static int conv_setup(cudaStream_t st) {
    cudaError_t e = fft8_ensure_tables(st);
    DDSP_REQUIRE(e == cudaSuccess, "conv: twiddle table initialisation failed: %s", cudaGetErrorString(e));
    return 0;
}

// voice-pair chunks of the kernel-gradient walk: as many as fit ONE wave of resident CTAs (the 32-chunk grid was 2.3 waves
// of 3 CTAs per SM, the last one a third full), evened out; GIR_CHUNKS bounds the partial-sum scratch
static int gir_chunk_count(int B2) {
    static int slots_cache[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    if (slots_cache[dev] == 0) {
        int sms = 148, per_sm = 3;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) sms = 148;
        cudaFuncSetAttribute(conv_gir_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GIR_SMEM);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, conv_gir_kernel, MAC_THREADS, GIR_SMEM) != cudaSuccess || per_sm < 1) per_sm = 3;
        const int s = sms * per_sm / (CONV_N2 / MAC_BINS);
        slots_cache[dev] = s < 1 ? 1 : s;
    }
    int n = slots_cache[dev];
    if (n > GIR_CHUNKS) n = GIR_CHUNKS;
    if (n > B2) n = B2;
    const int per = (B2 + n - 1) / n;
    return (B2 + per - 1) / per;
}

static int check_conv(const char* fn, int B, int N, int Lk) {
    DDSP_REQUIRE(B >= 1 && B <= 2 * 65535, "%s: B=%d out of range", fn, B);
    DDSP_REQUIRE(N >= 1 && Lk >= 1, "%s: N=%d, Lk=%d must be positive", fn, N, Lk);
    DDSP_REQUIRE((N + CONV_P - 1) / CONV_P <= 65535 * 32, "%s: N=%d too long", fn, N);
    return 0;
}

}  // namespace ddsp

using namespace ddsp;

extern "C" int ddsp_reverb_impulse(const float* noise, const float* t, const float* decay, const float* wet, float* ir, int L,
                                   void* stream) {
    DDSP_REQUIRE(noise && t && decay && wet && ir && L >= 1, "ddsp_reverb_impulse: bad arguments");
    reverb_impulse_kernel<<<(L + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(noise, t, decay, wet, ir, L);
    DDSP_CHECK_LAUNCH("reverb_impulse_kernel");
    return 0;
}

extern "C" int ddsp_reverb_impulse_backward(const float* gir, const float* noise, const float* t, const float* decay,
                                            const float* wet, float* gnoise, float* gdecay, float* gwet, int L, void* stream) {
    DDSP_REQUIRE(gir && noise && t && decay && wet && gnoise && gdecay && gwet && L >= 1,
                 "ddsp_reverb_impulse_backward: bad arguments");
    reverb_impulse_bwd_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(gir, noise, t, decay, wet, gnoise, gdecay, gwet, L);
    DDSP_CHECK_LAUNCH("reverb_impulse_bwd_kernel");
    return 0;
}

extern "C" size_t ddsp_conv_scratch_bytes(int B, int N, int Lk) {
    if (B < 1 || N < 1 || Lk < 1) return 0;
    return conv_plan(B, N, Lk).total;
}

extern "C" int ddsp_conv_forward(const float* x, const float* ir, float* y, void* scratch, int B, int N, int Lk, void* stream) {
    if (check_conv("ddsp_conv_forward", B, N, Lk)) return 1;
    DDSP_REQUIRE(x && ir && y && scratch, "ddsp_conv_forward: null pointer argument");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (conv_setup(st)) return 1;
    const ConvPlan p = conv_plan(B, N, Lk);
    char* base = reinterpret_cast<char*>(scratch);
    float2* Hf = reinterpret_cast<float2*>(base + p.off_H);
    float2* S0 = reinterpret_cast<float2*>(base + p.off_S0);
    const size_t smem = sizeof(float2) * fft8_smem_elems(CONV_N2);
    const int Le = Lk < N ? Lk : N;
    conv_spectra_kernel<<<dim3(p.np, 1), CONV_THREADS, smem, st>>>(ir, Hf, 1, Le, p.np, 1, 0, 1);
    DDSP_CHECK_LAUNCH("conv_spectra_kernel(ir)");
    conv_spectra_kernel<<<dim3(p.nblk, p.B2), CONV_THREADS, smem, st>>>(x, S0, B, N, p.nblk, 3, 0, 0);
    DDSP_CHECK_LAUNCH("conv_spectra_kernel(x)");
    float2* S1 = reinterpret_cast<float2*>(base + p.off_S1);
    for (int jg = 0; jg < p.np; jg += MAC_NP) {
        conv_mac_kernel<<<dim3(CONV_N2 / MAC_BINS, p.B2), MAC_THREADS, MAC_SMEM, st>>>(S0, Hf, S1, p.nblk, p.np, jg, jg > 0);
        DDSP_CHECK_LAUNCH("conv_mac_kernel");
    }
    conv_ifft_kernel<<<dim3(p.nblk, p.B2), CONV_THREADS, smem, st>>>(S1, y, B, N, p.nblk, 0);
    DDSP_CHECK_LAUNCH("conv_ifft_kernel");
    return 0;
}

extern "C" int ddsp_conv_backward(const float* g, const float* x, const float* ir, float* gx, float* gir, void* scratch, int B,
                                  int N, int Lk, int scratch_from_forward, void* stream) {
    if (check_conv("ddsp_conv_backward", B, N, Lk)) return 1;
    DDSP_REQUIRE(g && ir && scratch && (gx || gir) && (x || !gir), "ddsp_conv_backward: null pointer argument");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (conv_setup(st)) return 1;
    const ConvPlan p = conv_plan(B, N, Lk);
    char* base = reinterpret_cast<char*>(scratch);
    float2* Hf = reinterpret_cast<float2*>(base + p.off_H);
    float2* S0 = reinterpret_cast<float2*>(base + p.off_S0);
    float2* S1 = reinterpret_cast<float2*>(base + p.off_S1);
    float2* part = reinterpret_cast<float2*>(base + p.off_part);
    const size_t smem = sizeof(float2) * fft8_smem_elems(CONV_N2);
    const int Le = Lk < N ? Lk : N;
    // The kernel gradient goes first: with scratch_from_forward the block spectra of x (S0) and of the kernel (Hf) are
    // still the ones ddsp_conv_forward left there; the gx pass below then overwrites S0.
    if (gir) {
        if (!scratch_from_forward) {
            conv_spectra_kernel<<<dim3(p.nblk, p.B2), CONV_THREADS, smem, st>>>(x, S0, B, N, p.nblk, 3, 0, 0);
            DDSP_CHECK_LAUNCH("conv_spectra_kernel(x)");
        }
        conv_spectra_kernel<<<dim3(p.nblk, p.B2), CONV_THREADS, smem, st>>>(g, S1, B, N, p.nblk, 2, 0, 0);
        DDSP_CHECK_LAUNCH("conv_spectra_kernel(g tail)");
        const int nchunks = gir_chunk_count(p.B2);
        for (int jg = 0; jg < p.np; jg += GIR_NP) {
            conv_gir_kernel<<<dim3(CONV_N2 / MAC_BINS, nchunks), MAC_THREADS, GIR_SMEM, st>>>(S0, S1, part, p.B2, p.nblk, p.np, jg);
            DDSP_CHECK_LAUNCH("conv_gir_kernel");
        }
        cudaError_t e = cudaMemsetAsync(gir, 0, sizeof(float) * (size_t)Lk, st);
        DDSP_REQUIRE(e == cudaSuccess, "ddsp_conv_backward: memset failed: %s", cudaGetErrorString(e));
        conv_gir_finish_kernel<<<p.np, CONV_THREADS, smem, st>>>(part, gir, p.np, Le, nchunks);
        DDSP_CHECK_LAUNCH("conv_gir_finish_kernel");
    }
    if (gx) {      // gx = flip(conv(flip(g), ir))
        if (!scratch_from_forward) {
            conv_spectra_kernel<<<dim3(p.np, 1), CONV_THREADS, smem, st>>>(ir, Hf, 1, Le, p.np, 1, 0, 1);
            DDSP_CHECK_LAUNCH("conv_spectra_kernel(ir)");
        }
        conv_spectra_kernel<<<dim3(p.nblk, p.B2), CONV_THREADS, smem, st>>>(g, S0, B, N, p.nblk, 3, 1, 0);
        DDSP_CHECK_LAUNCH("conv_spectra_kernel(flip g)");
        for (int jg = 0; jg < p.np; jg += MAC_NP) {
            conv_mac_kernel<<<dim3(CONV_N2 / MAC_BINS, p.B2), MAC_THREADS, MAC_SMEM, st>>>(S0, Hf, S1, p.nblk, p.np, jg, jg > 0);
            DDSP_CHECK_LAUNCH("conv_mac_kernel(gx)");
        }
        conv_ifft_kernel<<<dim3(p.nblk, p.B2), CONV_THREADS, smem, st>>>(S1, gx, B, N, p.nblk, 1);
        DDSP_CHECK_LAUNCH("conv_ifft_kernel(gx)");
    }
    return 0;
}

extern "C" int ddsp_reverb_live(float* buffer, const float* x, const float* ir, float* y, void* scratch, int L, int n,
                                void* stream) {
    DDSP_REQUIRE(buffer && x && ir && y && scratch, "ddsp_reverb_live: null pointer argument");
    DDSP_REQUIRE(L >= 1 && n >= 1 && n <= L, "ddsp_reverb_live: need 1 <= n <= L (n=%d, L=%d)", n, L);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    float* tmp = reinterpret_cast<float*>(scratch);    // >= L floats
    if (L - n > 0) {
        copy_kernel<<<(L - n + 255) / 256, 256, 0, st>>>(buffer + n, tmp, L - n);
        copy_kernel<<<(L - n + 255) / 256, 256, 0, st>>>(tmp, buffer, L - n);
    }
    copy_kernel<<<(n + 255) / 256, 256, 0, st>>>(x, buffer + (L - n), n);
    reverb_live_dot_kernel<<<n, 256, 0, st>>>(buffer, ir, y, L, n);
    DDSP_CHECK_LAUNCH("reverb_live kernels");
    return 0;
}
