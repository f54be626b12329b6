// Filtered-noise synthesiser for sm_100a -- replaces FilteredNoise.forward (filtered_noise.py:40-53).
//
// The reference runs, per frame, irfft(H) -> roll/Hann/pad/roll (amp_to_impulse_response, :7-22) and
// then three FFTs of size 2*hop (fft_convolve, :25-32) on noise drawn by the CPU generator and copied
// to the GPU (:44-48).  Here one CTA owns one frame end to end:
//   * the M = 2(NB-1) tap zero-phase FIR is built in shared memory from the NB magnitudes by the
//     real inverse DFT written out as a cosine sum (r is even, so only r[0..M/2] is needed),
//     windowed with the periodic Hann;  after the reference's two rolls the filter is
//         causal taps      hc[q] = r[q]     * 0.5(1+cos(2 pi q/M)),  q = 0..M/2-1   at lag q
//         wrapped taps     ha[i] = r[M/2-i] * 0.5(1-cos(2 pi i/M)),  i = 0..M/2-1   at lag hop-M/2+i
//   * the uniform noise is generated in-kernel (Philox4x32-10) or read from the caller,
//   * the frame's causal, truncated convolution is done directly from shared memory with an
//     8-output x 8-tap register tile (64 FFMA per 4 LDS.128) -- no overlap-add, tails dropped (:50-51).
// The backward kernel regenerates the same noise, correlates it with the upstream gradient into the
// tap gradients and applies the transposed cosine sum.
#include "ddsp_common.cuh"

namespace ddsp {

struct NoiseSmem {
    float* Hs;    // [NB]        magnitudes, later reused for d/dH accumulation
    float* ct;    // [M]         cos(2 pi i / M)
    float* rr;    // [M/2+1]     r = irfft(H)[0..M/2]
    float* hc;    // [TP]        causal taps (zero padded to a multiple of 8)
    float* ha;    // [TP]        wrapped (anti-causal) taps
    float* u;     // [PADL + hop] noise with PADL leading zeros
    int TP, PADL;
};

__host__ __device__ inline int noise_tp(int NB) { return ((NB - 1) + 7) & ~7; }

__host__ __device__ inline size_t noise_smem_floats(int NB, int hop) {
    const int M = 2 * (NB - 1), TP = noise_tp(NB);
    size_t n = ((NB + 3) & ~3) + ((M + 3) & ~3) + ((M / 2 + 1 + 3) & ~3) + 2 * (size_t)TP + (TP + 8) + hop;
    return n;
}

__device__ __forceinline__ NoiseSmem carve_noise_smem(float* base, int NB, int hop) {
    const int M = 2 * (NB - 1);
    NoiseSmem s;
    s.TP = noise_tp(NB);
    s.PADL = s.TP + 8;
    s.Hs = base;          base += (NB + 3) & ~3;
    s.ct = base;          base += (M + 3) & ~3;
    s.rr = base;          base += (M / 2 + 1 + 3) & ~3;
    s.hc = base;          base += s.TP;
    s.ha = base;          base += s.TP;
    s.u = base;
    return s;
}

// Builds ct, rr, hc, ha from Hs (all in shared memory). Ends with a __syncthreads().
__device__ __forceinline__ void build_taps(const NoiseSmem& s, int NB) {
    const int M = 2 * (NB - 1), half = M / 2;
    const int tid = threadIdx.x, nt = blockDim.x;
    const float invM = 1.0f / (float)M;
    for (int i = tid; i < M; i += nt) s.ct[i] = cospif(2.0f * (float)i * invM);
    __syncthreads();
    // r[j] = (1/M) (H_0 + (-1)^j H_{NB-1} + 2 sum_{m=1}^{NB-2} H_m cos(2 pi m j / M))        (irfft, :8-10)
    for (int j = tid; j <= half; j += nt) {
        float acc = 0.f;
        int idx = j % M;
        for (int m = 1; m < NB - 1; ++m) {
            acc = fmaf(s.Hs[m], s.ct[idx], acc);
            idx += j;
            if (idx >= M) idx -= M;
        }
        const float edge = s.Hs[0] + ((j & 1) ? -s.Hs[NB - 1] : s.Hs[NB - 1]);
        s.rr[j] = (2.f * acc + edge) * invM;
    }
    __syncthreads();
    for (int q = tid; q < s.TP; q += nt) {
        float c = 0.f, a = 0.f;
        if (q < half) {
            c = s.rr[q] * (0.5f + 0.5f * s.ct[q]);                 // r[q]     * hann[M/2+q]
            a = s.rr[half - q] * (0.5f - 0.5f * s.ct[q]);          // r[M/2+q] * hann[q]
        }
        s.hc[q] = c;
        s.ha[q] = a;
    }
    __syncthreads();
}

// Loads (or generates) the frame's noise into s.u[PADL .. PADL+hop) and zeroes the left pad.
__device__ __forceinline__ void load_noise(const NoiseSmem& s, const float* __restrict__ noise_frame, int hop, uint32_t t,
                                           uint32_t voice, uint32_t offset, uint2 key) {
    const int tid = threadIdx.x, nt = blockDim.x;
    for (int i = tid; i < s.PADL; i += nt) s.u[i] = 0.f;
    float* u = s.u + s.PADL;
    for (int i = hop + tid; i < ((hop + 7) & ~7); i += nt) u[i] = 0.f;        // tail of the last 8-sample group (hop % 8 != 0)
    if (noise_frame) {
        for (int i = tid; i < hop; i += nt) u[i] = noise_frame[i];
    } else {
        const int quads = (hop + 3) >> 2;
        for (int q = tid; q < quads; q += nt) {
            const float4 v = philox_noise4((uint32_t)q, t, voice, offset, key);
            const int i = 4 * q;
            if (i + 3 < hop) {
                *reinterpret_cast<float4*>(u + i) = v;            // PADL and i are multiples of 4
            } else {
                if (i < hop) u[i] = v.x;
                if (i + 1 < hop) u[i + 1] = v.y;
                if (i + 2 < hop) u[i + 2] = v.z;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// forward: grid (T, B), ceil(hop/8) threads, thread owns outputs j0..j0+7 (masked past hop)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512) noise_fwd_kernel(const float* __restrict__ H, const float* __restrict__ noise,
                                                        float* __restrict__ y, int accumulate, uint2 key, uint32_t offset,
                                                        const uint32_t* __restrict__ offset_dev, uint32_t voice_offset, int T,
                                                        int NB, int hop) {
    extern __shared__ __align__(16) float smem[];
    if (offset_dev) offset += *offset_dev;
    const NoiseSmem s = carve_noise_smem(smem, NB, (hop + 7) & ~7);     // any hop: the last thread's group is partly masked
    const int t = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
    const size_t frame = (size_t)b * T + t;
    for (int m = tid; m < NB; m += blockDim.x) s.Hs[m] = H[frame * NB + m];
    load_noise(s, noise ? noise + frame * hop : nullptr, hop, (uint32_t)t, voice_offset + (uint32_t)b, offset, key);
    __syncthreads();
    build_taps(s, NB);

    const int half = NB - 1;
    const int j0 = 8 * tid;
    const float* u = s.u + s.PADL;
    float acc[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) acc[r] = 0.f;
    float win[16];
    {
        const float4 a = *reinterpret_cast<const float4*>(u + j0), c = *reinterpret_cast<const float4*>(u + j0 + 4);
        win[8] = a.x; win[9] = a.y; win[10] = a.z; win[11] = a.w;
        win[12] = c.x; win[13] = c.y; win[14] = c.z; win[15] = c.w;
    }
    const int qend = min(s.TP, j0 + 8);                 // lags beyond the frame start only see zeros
    for (int q0 = 0; q0 < qend; q0 += 8) {
        {
            const float4 a = *reinterpret_cast<const float4*>(u + j0 - q0 - 8);
            const float4 c = *reinterpret_cast<const float4*>(u + j0 - q0 - 4);
            win[0] = a.x; win[1] = a.y; win[2] = a.z; win[3] = a.w;
            win[4] = c.x; win[5] = c.y; win[6] = c.z; win[7] = c.w;
        }
        const float4 ta = *reinterpret_cast<const float4*>(s.hc + q0), tb = *reinterpret_cast<const float4*>(s.hc + q0 + 4);
        const float tap[8] = {ta.x, ta.y, ta.z, ta.w, tb.x, tb.y, tb.z, tb.w};
#pragma unroll
        for (int d = 0; d < 8; ++d)
#pragma unroll
            for (int r = 0; r < 8; ++r) acc[r] = fmaf(tap[d], win[8 + r - d], acc[r]);
#pragma unroll
        for (int i = 0; i < 8; ++i) win[8 + i] = win[i];
    }
    // wrapped taps: lag hop-half+i only reaches outputs j >= hop-half            (the roll by -M/2, :20)
    const int jw = hop - half;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const int jp = j0 + r - jw;
        if (jp >= 0 && j0 + r < hop) {
            float a2 = 0.f;
            for (int i = 0; i <= jp; ++i) a2 = fmaf(s.ha[i], u[jp - i], a2);
            acc[r] += a2;
        }
    }
    float* yp = y + frame * hop + j0;
    if ((hop & 3) == 0 && j0 + 8 <= hop) {
        float4 o0 = make_float4(acc[0], acc[1], acc[2], acc[3]), o1 = make_float4(acc[4], acc[5], acc[6], acc[7]);
        if (accumulate) {
            const float4 p0 = *reinterpret_cast<const float4*>(yp), p1 = *reinterpret_cast<const float4*>(yp + 4);
            o0.x += p0.x; o0.y += p0.y; o0.z += p0.z; o0.w += p0.w;
            o1.x += p1.x; o1.y += p1.y; o1.z += p1.z; o1.w += p1.w;
        }
        *reinterpret_cast<float4*>(yp) = o0;
        *reinterpret_cast<float4*>(yp + 4) = o1;
    } else {                                            // rows not 16-byte aligned, or the masked tail group
#pragma unroll
        for (int r = 0; r < 8; ++r)
            if (j0 + r < hop) yp[r] = accumulate ? yp[r] + acc[r] : acc[r];
    }
}

// ------------------------------------------------------------------------------------------------
// backward: grid (T, B), (TP/8)*nsl threads.  Thread (qg, sl) owns taps 8qg..8qg+7 and the sl-th slice
// of the frame's samples:   d hc[q] = sum_j g[j] u[j-q]        d ha[i] = sum_{j'<M/2} g[hop-M/2+j'] u[j'-i]
// then   d r[p] = d hc[p] hann[M/2+p] + d ha[M/2-p] hann[M/2-p]      (p = 0..M/2)
//        d H[m] = (kappa_m / M) sum_p d r[p] cos(2 pi m p / M),  kappa = 1 at the two edge bands, else 2.
// The gradient does not depend on H itself (the filter is linear in H).
// ------------------------------------------------------------------------------------------------
__host__ __device__ inline int noise_bwd_slices(int NB, int hop) {
    const int groups = noise_tp(NB) / 8;
    int nsl = 1;
    while (nsl * 2 * groups <= 128 && hop % (16 * nsl) == 0) nsl *= 2;
    return nsl;
}

__host__ __device__ inline size_t noise_bwd_smem_floats(int NB, int hop) {
    const int M = 2 * (NB - 1), TP = noise_tp(NB), nsl = noise_bwd_slices(NB, hop);
    return ((M + 3) & ~3) + ((M / 2 + 1 + 3) & ~3) + (size_t)hop + (TP + 8) + (TP + 8 + hop) + 2 * (size_t)TP * nsl;
}

__global__ void __launch_bounds__(512) noise_bwd_kernel(const float* __restrict__ g, const float* __restrict__ noise,
                                                        float* __restrict__ gH, uint2 key, uint32_t offset,
                                                        const uint32_t* __restrict__ offset_dev, uint32_t voice_offset, int T,
                                                        int NB, int hop, int nsl) {
    extern __shared__ __align__(16) float smem[];
    if (offset_dev) offset += *offset_dev;
    const int M = 2 * (NB - 1), half = NB - 1, TP = noise_tp(NB), PADL = TP + 8;
    const int hopp = (hop + 7) & ~7;                   // sample loops run in groups of 8; g and u are zero past hop
    float* ct = smem;
    float* dr = ct + ((M + 3) & ~3);
    float* gs = dr + ((half + 1 + 3) & ~3);            // [hopp]   upstream gradient of this frame
    float* g2 = gs + hopp;                             // [TP+8]   gradient seen by the wrapped taps, zero padded
    NoiseSmem s;
    s.TP = TP; s.PADL = PADL;
    s.u = g2 + TP + 8;                                 // [PADL+hopp]
    float* dpc = s.u + PADL + hopp;                    // [TP][nsl] partial d hc
    float* dpa = dpc + (size_t)TP * nsl;               // [TP][nsl] partial d ha
    const int t = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, nt = blockDim.x;
    const size_t frame = (size_t)b * T + t;
    const int jw = hop - half;
    const float invM = 1.0f / (float)M;

    for (int i = tid; i < M; i += nt) ct[i] = cospif(2.0f * (float)i * invM);
    for (int i = tid; i < hopp; i += nt) gs[i] = i < hop ? g[frame * hop + i] : 0.f;
    for (int i = tid; i < TP + 8; i += nt) g2[i] = (i < half) ? g[frame * hop + jw + i] : 0.f;
    load_noise(s, noise ? noise + frame * hop : nullptr, hop, (uint32_t)t, voice_offset + (uint32_t)b, offset, key);
    __syncthreads();

    const float* u = s.u + PADL;
    const int qg = tid / nsl, sl = tid % nsl;
    const int q0 = 8 * qg, SL = hopp / nsl;
    float acc[8], acc2[8];
#pragma unroll
    for (int d = 0; d < 8; ++d) acc[d] = acc2[d] = 0.f;
    {
        const int jb0 = sl * SL, jend = jb0 + SL;
        int jb = max(jb0, q0);
        float win[16];
        if (jb < jend) {
            const float4 a = *reinterpret_cast<const float4*>(u + jb - q0 - 8), c = *reinterpret_cast<const float4*>(u + jb - q0 - 4);
            win[8] = a.x; win[9] = a.y; win[10] = a.z; win[11] = a.w;
            win[12] = c.x; win[13] = c.y; win[14] = c.z; win[15] = c.w;
        }
        for (; jb < jend; jb += 8) {
#pragma unroll
            for (int i = 0; i < 8; ++i) win[i] = win[8 + i];
            const float4 a = *reinterpret_cast<const float4*>(u + jb - q0), c = *reinterpret_cast<const float4*>(u + jb - q0 + 4);
            win[8] = a.x; win[9] = a.y; win[10] = a.z; win[11] = a.w;
            win[12] = c.x; win[13] = c.y; win[14] = c.z; win[15] = c.w;
            const float4 ga = *reinterpret_cast<const float4*>(gs + jb), gb = *reinterpret_cast<const float4*>(gs + jb + 4);
            const float gv[8] = {ga.x, ga.y, ga.z, ga.w, gb.x, gb.y, gb.z, gb.w};
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int d = 0; d < 8; ++d) acc[d] = fmaf(gv[r], win[8 + r - d], acc[d]);
        }
        // wrapped taps: blocks of j' dealt round-robin to the slices
        for (int jp = 8 * sl; jp < TP; jp += 8 * nsl) {
            if (jp < q0) continue;
#pragma unroll
            for (int i = 0; i < 16; ++i) win[i] = u[jp - q0 - 8 + i];
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const float gr = g2[jp + r];
#pragma unroll
                for (int d = 0; d < 8; ++d) acc2[d] = fmaf(gr, win[8 + r - d], acc2[d]);
            }
        }
    }
#pragma unroll
    for (int d = 0; d < 8; ++d) {
        dpc[(size_t)(q0 + d) * nsl + sl] = acc[d];
        dpa[(size_t)(q0 + d) * nsl + sl] = acc2[d];
    }
    __syncthreads();
    for (int p = tid; p <= half; p += nt) {
        float v = 0.f;
        if (p < half) {
            float sc = 0.f;
            for (int i = 0; i < nsl; ++i) sc += dpc[(size_t)p * nsl + i];
            v = sc * (0.5f + 0.5f * ct[p]);
        }
        if (p >= 1) {
            const int i0 = half - p;
            float sa = 0.f;
            for (int i = 0; i < nsl; ++i) sa += dpa[(size_t)i0 * nsl + i];
            v = fmaf(sa, 0.5f - 0.5f * ct[i0], v);
        }
        dr[p] = v;
    }
    __syncthreads();
    for (int m = tid; m < NB; m += nt) {
        float a2 = 0.f;
        int idx = 0;
        for (int p = 0; p <= half; ++p) {
            a2 = fmaf(dr[p], ct[idx], a2);
            idx += m;
            if (idx >= M) idx -= M;
        }
        gH[frame * NB + m] = a2 * ((m == 0 || m == NB - 1) ? invM : 2.f * invM);
    }
}

// ================================================================================================
// Fast path (hop % 16 == 0): a TEAM of hop/16 threads owns one frame (one warp at hop = 512), four warps
// of teams per CTA.  Differences from the generic kernels above:
//   * noise window in shared memory is XOR-swizzled at float4 granularity (swz_f / swz_b below) so the strided float4
//     window loads of a team are bank-conflict free at every offset;
//   * 16 outputs x 8 taps register tile: 128 FFMA per 4 LDS.128, three rotating 8-float window chunks
//     (no register shuffling);
//   * r = irfft(H) by a cosine recurrence over m per output j (no table look-ups), and only for
//     j <= M/4: cos(2 pi m (M/2-j)/M) = (-1)^m cos(2 pi m j/M), so even/odd partial sums give r[j] and r[M/2-j];
//   * the wrapped taps' triangular corner is spread over the whole team (one output per thread per pass)
//     instead of serialising on the threads that own the frame's tail.
// ================================================================================================
constexpr int TEAM_BLOCK = 128;

// Shared-memory index swizzles of the noise window (float index in, float index out; a float4 stays a float4).
// An LDS.128 is served a quarter-warp at a time; its 8 lanes are conflict-free when their float4 indices differ mod 8.
//   forward: lane l reads float4 index 4*l + c (16 outputs per lane).  Unswizzled that is 2 distinct values mod 8.  XOR of
//     the index with its 32-float row number mod 4 makes the 8 lanes distinct for EVERY c (the lanes touch rows r..r+4 and
//     rows r, r+4 share a mask, so the misaligned case reduces to the aligned one).  The round-1 layout -- 4 floats of
//     padding per 32, i.e. a rotation that keeps growing with the row -- collided on rows r and r+4: ncu counted 18 % of
//     the forward kernel's shared-memory wavefronts as conflicts (profiles/summary_r01g.md).
//   backward: the 8 lanes of a quarter-warp read float4 index c - 2*tg (tap groups 8 taps apart).  XOR with the row
//     number mod 2 sends odd rows to the odd float4 slots; any 8 consecutive terms of a stride-2 progression then land in 8
//     distinct slots (33 % conflicts before).
// Both permute within an aligned block of 32 floats and leave 32-float-aligned contiguous float4 stores conflict-free.
__device__ __forceinline__ int swz_f(int i) { return i ^ (((i >> 5) & 3) << 2); }
__device__ __forceinline__ int swz_b(int i) { return i ^ (((i >> 5) & 1) << 2); }

// A team's shared memory is private to it: when the team is exactly one warp (hop 512) a warp barrier is enough, and the
// four frames of a CTA drift apart so one frame's FMA-bound convolution overlaps another's latency-bound tap build.
__device__ __forceinline__ void team_sync(int TS) {
    if (TS == 32) __syncwarp(); else __syncthreads();
}

struct TeamLayout {
    int half, TP, PADL, uc_off, n_H, n_rr, n_u, n_uc, n_cor, total;
};

__host__ __device__ inline TeamLayout team_layout(int NB, int hop) {
    TeamLayout L;
    L.half = NB - 1;
    L.TP = noise_tp(NB);
    L.PADL = (L.TP + 16 + 31) & ~31;
    L.n_H = (NB + 3) & ~3;
    L.n_rr = (L.half + 1 + 3) & ~3;
    const int logical = L.PADL + hop;
    L.n_u = (logical + 4 + 3) & ~3;                     // the XOR swizzles permute inside aligned 32-float blocks: no padding
    L.n_uc = (2 * L.half + 4 + 3) & ~3;                 // unswizzled copy of u[-half .. half) for the corner
    L.n_cor = (L.half + 3) & ~3;
    L.total = L.n_H + L.n_rr + 3 * L.TP + 4 + L.n_u + L.n_uc + L.n_cor;      // taps: two pair-swapped copies + wrapped + tap 0
    const int out_need = hop;                           // the output staging area reuses the u region
    if (L.n_u < out_need) L.total += out_need - L.n_u;
    return L;
}

// r[0..half] from H (team-cooperative); H, rr in shared memory (H 16-byte aligned, readable up to the next multiple of 4).
// Lane j handles r[j] and r[half-j] for j < half/2; the middle tap (theta = pi/2: cos(m pi/2) is 0, +-1) would cost the
// whole team a second pass for one lane, so it is summed directly instead.
__device__ __forceinline__ void team_build_r(const float* H, float* rr, int NB, int l, int TS) {
    const int half = NB - 1, M = 2 * half;
    const float invM = 1.0f / (float)M;
    const float h0 = H[0], hN = H[half];
    for (int j = l; 2 * j < half; j += TS) {
        const float sh = sinpif((float)j * invM);       // sin(theta/2), theta = 2 pi j / M < pi/2
        const float dl = -4.f * sh * sh;                // 2 cos(theta) - 2
        float e = 0.5f * dl, c = 1.f + e;               // cos(1*theta)
        float E = 0.f, O = 0.f;
#pragma unroll 4
        for (int m4 = 0; m4 < half; m4 += 4) {
            const float4 h = *reinterpret_cast<const float4*>(H + m4);
            const float hv[4] = {h.x, h.y, h.z, h.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int m = m4 + k;
                if (m == 0 || m >= half) continue;
                if (m & 1) O = fmaf(hv[k], c, O); else E = fmaf(hv[k], c, E);
                e = fmaf(dl, c, e); c += e;
            }
        }
        const float sj = (j & 1) ? -hN : hN, sk = ((half - j) & 1) ? -hN : hN;
        rr[j] = (h0 + sj + 2.f * (E + O)) * invM;
        rr[half - j] = (h0 + sk + 2.f * (E - O)) * invM;
    }
    if ((half & 1) == 0 && l == (half / 2) % TS) {      // j = half/2
        float E = 0.f;
        for (int m = 2; m < half; m += 4) E -= H[m];    // cos(m pi/2) = -1 at m = 2, 6, ..
        for (int m = 4; m < half; m += 4) E += H[m];    //              +1 at m = 4, 8, ..
        const float sj = ((half / 2) & 1) ? -hN : hN;
        rr[half / 2] = (h0 + sj + 2.f * E) * invM;
    }
}

// Packed tile: 16 outputs x 8 taps as 64 FFMA2.  Every output keeps TWO partial sums (one per lane of a float2) so that
// a tap pair (h[a], h[a+1]) meets an ascending, even-aligned window pair (w[x], w[x+1]):
//   odd  r : taps (q0+2k, q0+2k+1)   x window pair starting at 7+r-2k -> tap operand (h[2k+1], h[2k])   = copy A
//   even r : taps (q0+2k+1, q0+2k+2) x window pair starting at 6+r-2k -> tap operand (h[2k+2], h[2k+1]) = copy B
// i.e. even outputs run one tap behind: tap 0 is added before the loop and tap TP (zero) closes the last tile.
#define NOISE_TILE2(NW, M1, M2, Q0)                                                                             \
    {                                                                                                           \
        const float4 l0_ = *reinterpret_cast<const float4*>(u + swz_f(ub - (Q0) - 8));                            \
        const float4 l1_ = *reinterpret_cast<const float4*>(u + swz_f(ub - (Q0) - 4));                            \
        NW[0] = make_float2(l0_.x, l0_.y); NW[1] = make_float2(l0_.z, l0_.w);                                   \
        NW[2] = make_float2(l1_.x, l1_.y); NW[3] = make_float2(l1_.z, l1_.w);                                   \
        const float4 a0_ = *reinterpret_cast<const float4*>(hcA + (Q0));                                        \
        const float4 a1_ = *reinterpret_cast<const float4*>(hcA + (Q0) + 4);                                    \
        const float4 b0_ = *reinterpret_cast<const float4*>(hcB + (Q0));                                        \
        const float4 b1_ = *reinterpret_cast<const float4*>(hcB + (Q0) + 4);                                    \
        const float2 A2_[4] = {make_float2(a0_.x, a0_.y), make_float2(a0_.z, a0_.w), make_float2(a1_.x, a1_.y), \
                               make_float2(a1_.z, a1_.w)};                                                      \
        const float2 B2_[4] = {make_float2(b0_.x, b0_.y), make_float2(b0_.z, b0_.w), make_float2(b1_.x, b1_.y), \
                               make_float2(b1_.z, b1_.w)};                                                      \
        _Pragma("unroll") for (int k = 0; k < 4; ++k) {                                                         \
            _Pragma("unroll") for (int r = 0; r < 16; ++r) {                                                    \
                const int pi = ((r & 1) ? 7 + r - 2 * k : 6 + r - 2 * k) >> 1;                                  \
                const float2 wv = pi < 4 ? NW[pi] : (pi < 8 ? M1[pi - 4] : M2[pi - 8]);                         \
                acc2[r] = __ffma2_rn((r & 1) ? A2_[k] : B2_[k], wv, acc2[r]);                                   \
            }                                                                                                   \
        }                                                                                                       \
    }

// NBc / HOPc > 0: compile-time band count / hop of the common configurations (loops unroll, address arithmetic folds);
// 0: taken from the run-time arguments.
template <int NBc, int HOPc>
__global__ void __launch_bounds__(TEAM_BLOCK) noise_fwd_team_kernel(const float* __restrict__ H, const float* __restrict__ noise,
                                                                    float* __restrict__ y, int accumulate, uint2 key,
                                                                    uint32_t offset, const uint32_t* __restrict__ offset_dev,
                                                                    uint32_t voice_offset, int T, int NB_rt, int hop_rt,
                                                                    long long frames) {
    const int NB = NBc > 0 ? NBc : NB_rt, hop = HOPc > 0 ? HOPc : hop_rt;
    if (offset_dev) offset += *offset_dev;
    extern __shared__ __align__(16) float smem[];
    const int TS = hop >> 4, FPB = TEAM_BLOCK / TS;
    const int team = threadIdx.x / TS, l = threadIdx.x % TS;
    const TeamLayout L = team_layout(NB, hop);
    float* base = smem + (size_t)team * L.total;
    float* Hs = base;
    float* rr = Hs + L.n_H;
    float* hcA = rr + L.n_rr;                           // causal taps, pairs swapped: (h1,h0,h3,h2,..)
    float* hcB = hcA + L.TP;                            // causal taps, one behind:    (h2,h1,h4,h3,..,0,h[TP-1])
    float* ha = hcB + L.TP;                             // wrapped taps
    float* h0s = ha + L.TP;                             // h[0]
    float* u = h0s + 4;                                 // swizzled: logical index PADL + j
    float* uc = u + (L.n_u < hop ? hop : L.n_u);        // uc[half + x] = u[x], x in [-half, half)
    float* cor = uc + L.n_uc;
    const int half = L.half, TP = L.TP, PADL = L.PADL;
    const long long frame = (long long)blockIdx.x * FPB + team;
    const bool live = frame < frames;
    // (voice, frame-in-voice) of this team's frame: a 32-bit division when the frame index fits (it does unless B*T >= 2^32);
    // the 64-bit one is a ~100-instruction dependent sequence in front of the first Philox counter
    uint32_t b = 0u, t = 0u;
    if (live) {
        if (frame <= 0xffffffffLL) {
            b = (uint32_t)frame / (uint32_t)T;
            t = (uint32_t)frame - b * (uint32_t)T;
        } else {
            b = (uint32_t)(frame / T);
            t = (uint32_t)(frame % T);
        }
    }

    // ---- stage 1: H, noise (generated or read), zero pads -----------------------------------------
    // The frame's magnitudes are requested first and parked in registers while the noise is generated (as the gradient is in
    // the backward kernel): stored as they arrived, the load was the kernel's top stall line (13.6 % of warp samples).
    const bool h_early = NB <= 4 * TS;
    float hreg[4];
    if (live) {
        if (h_early) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int m = l + k * TS;
                hreg[k] = m < NB ? H[frame * NB + m] : 0.f;
            }
        } else {
            for (int m = l; m < NB; m += TS) Hs[m] = H[frame * NB + m];
        }
    }
    for (int i = l; i < PADL; i += TS) u[i] = 0.f;           // PADL is a multiple of 32: the swizzle stays inside
    for (int i = l; i < L.n_uc; i += TS) uc[i] = 0.f;
    team_sync(TS);
    if (live) {
        const float* nf = noise ? noise + frame * hop : nullptr;
        // hop/4 quads over hop/16 lanes: always 4 per lane.  Unrolled so the four Philox calls (ten dependent multiply rounds
        // each) interleave: run one after the other they were 32 % of the kernel's warp samples (profiles/summary_r02.md)
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int q = l + it * TS;
            const float4 v = nf ? *reinterpret_cast<const float4*>(nf + 4 * q)
                                : philox_noise4((uint32_t)q, t, voice_offset + b, offset, key);
            *reinterpret_cast<float4*>(u + swz_f(PADL + 4 * q)) = v;
            if (4 * q < half) {
                const float vv[4] = {v.x, v.y, v.z, v.w};
                for (int e = 0; e < 4; ++e)
                    if (4 * q + e < half) uc[half + 4 * q + e] = vv[e];
            }
        }
        if (h_early) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int m = l + k * TS;
                if (m < NB) Hs[m] = hreg[k];
            }
        }
    }
    if (h_early) team_sync(TS);
    if (live) team_build_r(Hs, rr, NB, l, TS);
    team_sync(TS);
    // ---- stage 2: windowed taps (filtered_noise.py:14-20) -----------------------------------------
    if (live) {
        const float invM = 0.5f / (float)half;
        for (int q = l; q < TP; q += TS) {
            float c = 0.f, a = 0.f;
            if (q < half) {
                const float cq = cospif(2.f * (float)q * invM);
                c = rr[q] * (0.5f + 0.5f * cq);
                a = rr[half - q] * (0.5f - 0.5f * cq);
            }
            hcA[q ^ 1] = c;
            if (q & 1) hcB[q] = c;
            else if (q >= 2) hcB[q - 2] = c;
            else { h0s[0] = c; hcB[TP - 2] = 0.f; }
            ha[q] = a;
        }
    }
    team_sync(TS);
    // ---- stage 3: causal taps, 16 outputs per thread ----------------------------------------------
    float acc[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) acc[r] = 0.f;
    if (live) {
        const int ub = PADL + 16 * l;                   // logical index of this thread's first output
        float2 acc2[16];
        float2 wa[4], wb[4], wc[4];
        {
            const float4 a0 = *reinterpret_cast<const float4*>(u + swz_f(ub)), a1 = *reinterpret_cast<const float4*>(u + swz_f(ub + 4));
            const float4 a2 = *reinterpret_cast<const float4*>(u + swz_f(ub + 8)), a3 = *reinterpret_cast<const float4*>(u + swz_f(ub + 12));
            wb[0] = make_float2(a0.x, a0.y); wb[1] = make_float2(a0.z, a0.w); wb[2] = make_float2(a1.x, a1.y); wb[3] = make_float2(a1.z, a1.w);
            wc[0] = make_float2(a2.x, a2.y); wc[1] = make_float2(a2.z, a2.w); wc[2] = make_float2(a3.x, a3.y); wc[3] = make_float2(a3.z, a3.w);
            const float h0 = h0s[0];                    // tap 0 of the even outputs (they run one tap behind)
#pragma unroll
            for (int r = 0; r < 16; r += 2) {
                const float2 own = (r < 8) ? wb[r >> 1] : wc[(r - 8) >> 1];
                acc2[r] = make_float2(h0 * own.x, 0.f);
                acc2[r + 1] = make_float2(0.f, 0.f);
            }
        }
        for (int q0 = 0; q0 < TP; q0 += 24) {
            NOISE_TILE2(wa, wb, wc, q0);
            if (q0 + 8 >= TP) break;
            NOISE_TILE2(wc, wa, wb, q0 + 8);
            if (q0 + 16 >= TP) break;
            NOISE_TILE2(wb, wc, wa, q0 + 16);
        }
#pragma unroll
        for (int r = 0; r < 16; ++r) acc[r] = acc2[r].x + acc2[r].y;
        // ---- stage 4: wrapped taps, one corner output per thread per pass -------------------------
        // u[x] = 0 for x < 0: output jp only sees taps i <= jp; the bound is rounded up to the pass so it stays team-uniform
        for (int jp = l; jp < half; jp += TS) {
            float a2 = 0.f;
            const float* ux = uc + half + jp;
            const int iend = min(half, (jp - l) + TS);
            int i = 0;
#pragma unroll 4
            for (; i + 4 <= iend; i += 4) {
                const float4 h4 = *reinterpret_cast<const float4*>(ha + i);
                a2 = fmaf(h4.x, ux[-i], a2);
                a2 = fmaf(h4.y, ux[-i - 1], a2);
                a2 = fmaf(h4.z, ux[-i - 2], a2);
                a2 = fmaf(h4.w, ux[-i - 3], a2);
            }
            for (; i < iend; ++i) a2 = fmaf(ha[i], ux[-i], a2);
            cor[jp] = a2;
        }
    }
    team_sync(TS);
    // ---- stage 5: add the corner, stage through shared memory for coalesced float4 stores ---------
    float* outs = u;                                    // window no longer needed
    if (live) {
        const int jw = hop - half;
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            const int j = 16 * l + r;
            if (j >= jw) acc[r] += cor[j - jw];
        }
    }
    team_sync(TS);
    if (live) {
#pragma unroll
        for (int r = 0; r < 16; r += 4)      // same stride-16 pattern as the window loads: swizzled, or it is a 4-way conflict
            *reinterpret_cast<float4*>(outs + swz_f(16 * l + r)) = make_float4(acc[r], acc[r + 1], acc[r + 2], acc[r + 3]);
    }
    team_sync(TS);
    if (live) {
        float* yp = y + frame * hop;
        for (int q = l; q < (hop >> 2); q += TS) {
            float4 o = *reinterpret_cast<const float4*>(outs + swz_f(4 * q));
            if (accumulate) {
                const float4 p = *reinterpret_cast<const float4*>(yp + 4 * q);
                o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w;
            }
            *reinterpret_cast<float4*>(yp + 4 * q) = o;
        }
    }
}

// ---- backward fast path -------------------------------------------------------------------------
// Work item (tap group tg of 8 taps, sample slice sl) = 8x8 register tiles over the slice; items are laid out
// so that the 8 threads of a quarter warp share a slice (g loads broadcast, window loads conflict-free).
struct TeamBwdLayout {
    int half, TP, PADL, ntg, nsl, n_u, n_uc, n_g2, n_part, n_dha, n_dr, total;
};

__host__ __device__ inline TeamBwdLayout team_bwd_layout(int NB, int hop) {
    TeamBwdLayout L;
    L.half = NB - 1;
    L.TP = noise_tp(NB);
    L.PADL = (L.TP + 16 + 31) & ~31;
    L.ntg = L.TP / 8;
    const int TS = hop >> 4;
    int nsl = 1;
    while (nsl * 2 * L.ntg <= TS && hop % (16 * nsl) == 0) nsl *= 2;
    L.nsl = nsl;
    const int logical = L.PADL + hop;
    L.n_u = (logical + 4 + 3) & ~3;                     // the XOR swizzles permute inside aligned 32-float blocks: no padding
    L.n_uc = (2 * L.half + 4 + 3) & ~3;
    L.n_g2 = (L.half + 3) & ~3;
    L.n_part = L.TP * nsl;
    L.n_dha = (L.half + 3) & ~3;
    L.n_dr = (L.half + 1 + 3) & ~3;
    L.total = hop + L.n_u + L.n_uc + L.n_g2 + L.n_part + L.n_dha + L.n_dr;
    return L;
}

#define NOISE_BTILE(LO, UP, JB)                                                                                 \
    {                                                                                                           \
        const float4 l0_ = *reinterpret_cast<const float4*>(u + swz_b(PADL + (JB) - q0));                         \
        const float4 l1_ = *reinterpret_cast<const float4*>(u + swz_b(PADL + (JB) - q0 + 4));                     \
        UP[0] = l0_.x; UP[1] = l0_.y; UP[2] = l0_.z; UP[3] = l0_.w;                                             \
        UP[4] = l1_.x; UP[5] = l1_.y; UP[6] = l1_.z; UP[7] = l1_.w;                                             \
        const float4 g0_ = *reinterpret_cast<const float4*>(gs + (JB));                                         \
        const float4 g1_ = *reinterpret_cast<const float4*>(gs + (JB) + 4);                                     \
        const float gv_[8] = {g0_.x, g0_.y, g0_.z, g0_.w, g1_.x, g1_.y, g1_.z, g1_.w};                          \
        _Pragma("unroll") for (int r = 0; r < 8; ++r) {                                                         \
            _Pragma("unroll") for (int d = 0; d < 8; ++d) {                                                     \
                const int ix = 8 + r - d;                                                                       \
                acc[d] = fmaf(gv_[r], ix < 8 ? LO[ix] : UP[ix - 8], acc[d]);                                    \
            }                                                                                                   \
        }                                                                                                       \
    }

template <int NBc, int HOPc>
__global__ void __launch_bounds__(TEAM_BLOCK) noise_bwd_team_kernel(const float* __restrict__ g, const float* __restrict__ noise,
                                                                    float* __restrict__ gH, uint2 key, uint32_t offset,
                                                                    const uint32_t* __restrict__ offset_dev, uint32_t voice_offset,
                                                                    int T, int NB_rt, int hop_rt, long long frames) {
    const int NB = NBc > 0 ? NBc : NB_rt, hop = HOPc > 0 ? HOPc : hop_rt;
    if (offset_dev) offset += *offset_dev;
    extern __shared__ __align__(16) float smem[];
    const int TS = hop >> 4, FPB = TEAM_BLOCK / TS;
    const int team = threadIdx.x / TS, l = threadIdx.x % TS;
    const TeamBwdLayout L = team_bwd_layout(NB, hop);
    float* gs = smem + (size_t)team * L.total;          // [hop] upstream gradient of the frame
    float* u = gs + hop;                                // swizzled noise, logical index PADL + j
    float* uc = u + L.n_u;                              // uc[half + x] = u[x], x in [-half, half)
    float* g2 = uc + L.n_uc;                            // gradient seen by the wrapped taps
    float* part = g2 + L.n_g2;                          // [nsl][TP], tap index swizzled (swz_b): float4 stores of a tap group
    float* dha = part + L.n_part;
    float* dr = dha + L.n_dha;
    const int half = L.half, PADL = L.PADL, nsl = L.nsl, ntg = L.ntg;
    const long long frame = (long long)blockIdx.x * FPB + team;
    const bool live = frame < frames;
    // (voice, frame-in-voice) of this team's frame: a 32-bit division when the frame index fits (it does unless B*T >= 2^32);
    // the 64-bit one is a ~100-instruction dependent sequence in front of the first Philox counter
    uint32_t b = 0u, t = 0u;
    if (live) {
        if (frame <= 0xffffffffLL) {
            b = (uint32_t)frame / (uint32_t)T;
            t = (uint32_t)frame - b * (uint32_t)T;
        } else {
            b = (uint32_t)(frame / T);
            t = (uint32_t)(frame % T);
        }
    }
    const int jw = hop - half;

    for (int i = l; i < PADL; i += TS) u[i] = 0.f;
    for (int i = l; i < L.n_uc; i += TS) uc[i] = 0.f;
    team_sync(TS);
    if (live) {
        const float* gf = g + frame * hop;
        const float* nf = noise ? noise + frame * hop : nullptr;
        // the frame's gradient is requested first (hop/4 float4 over hop/16 lanes: always 4 per lane) and parked in registers
        // while the noise is generated; copied load-by-load inside the loop it cost 13.6 % of the warp samples (ncu r01g)
        float4 gq[4];
#pragma unroll
        for (int it = 0; it < 4; ++it) gq[it] = *reinterpret_cast<const float4*>(gf + 4 * (l + it * TS));
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int q = l + it * TS;
            const float4 v = nf ? *reinterpret_cast<const float4*>(nf + 4 * q)
                                : philox_noise4((uint32_t)q, t, voice_offset + b, offset, key);
            *reinterpret_cast<float4*>(gs + 4 * q) = gq[it];
            *reinterpret_cast<float4*>(u + swz_b(PADL + 4 * q)) = v;
            if (4 * q < half) {
                const float vv[4] = {v.x, v.y, v.z, v.w};
                for (int e = 0; e < 4; ++e)
                    if (4 * q + e < half) uc[half + 4 * q + e] = vv[e];
            }
        }
        for (int i = l; i < L.n_g2; i += TS) g2[i] = i < half ? gf[jw + i] : 0.f;
    }
    team_sync(TS);
    if (live) {
        // ---- d hc[q] = sum_j g[j] u[j-q] ----------------------------------------------------------
        const int SL = hop / nsl;
        for (int w = l; w < ntg * nsl; w += TS) {
            const int sl = w / ntg, tg = w % ntg, q0 = 8 * tg;
            float acc[8];
#pragma unroll
            for (int d = 0; d < 8; ++d) acc[d] = 0.f;
            float wa[8], wb[8];
            const int jb0 = sl * SL;
            {
                const float4 a0 = *reinterpret_cast<const float4*>(u + swz_b(PADL + jb0 - q0 - 8));
                const float4 a1 = *reinterpret_cast<const float4*>(u + swz_b(PADL + jb0 - q0 - 4));
                wa[0] = a0.x; wa[1] = a0.y; wa[2] = a0.z; wa[3] = a0.w; wa[4] = a1.x; wa[5] = a1.y; wa[6] = a1.z; wa[7] = a1.w;
            }
            for (int jb = jb0; jb < jb0 + SL; jb += 16) {
                NOISE_BTILE(wa, wb, jb);
                if (jb + 8 >= jb0 + SL) break;
                NOISE_BTILE(wb, wa, jb + 8);
            }
            // two float4 per item; the quarter-warp's 8 tap groups land in 8 distinct float4 slots (the [TP][nsl] scalar layout
            // of round 1 put all 8 on one bank: 64 wavefronts for 8 stores)
            *reinterpret_cast<float4*>(part + sl * L.TP + swz_b(q0)) = make_float4(acc[0], acc[1], acc[2], acc[3]);
            *reinterpret_cast<float4*>(part + sl * L.TP + swz_b(q0 + 4)) = make_float4(acc[4], acc[5], acc[6], acc[7]);
        }
        // ---- d ha[i] = sum_{j'} g[hop-M/2+j'] u[j'-i] ---------------------------------------------
        // u[jp - i] = 0 for jp < i: start at the pass's first i (team-uniform, multiple of 4 when TS is)
        for (int i = l; i < half; i += TS) {
            float a2 = 0.f;
            const float* ux = uc + half - i;
            const int j0 = (i - l) & ~3;
            for (int jp = j0; jp < half; jp += 4) {
                const float4 g4 = *reinterpret_cast<const float4*>(g2 + jp);
                a2 = fmaf(g4.x, ux[jp], a2);
                a2 = fmaf(g4.y, ux[jp + 1], a2);
                a2 = fmaf(g4.z, ux[jp + 2], a2);
                a2 = fmaf(g4.w, ux[jp + 3], a2);
            }
            dha[i] = a2;
        }
    }
    team_sync(TS);
    if (live) {
        // ---- d r[p] = hann[M/2+p] * (d hc[p] + d ha[M/2-p]) ---------------------------------------
        const float invh = 1.0f / (float)half;
        for (int p = l; p <= half; p += TS) {
            float v = 0.f;
            if (p < half)
                for (int i = 0; i < nsl; ++i) v += part[i * L.TP + swz_b(p)];
            if (p >= 1) v += dha[half - p];
            dr[p] = v * (0.5f + 0.5f * cospif((float)p * invh));
        }
    }
    team_sync(TS);
    if (live) {
        // ---- d H[m] = kappa_m/M sum_p d r[p] cos(2 pi m p/M); m and M/2-m share even/odd partial sums
        const int M = 2 * half;
        const float invM = 1.0f / (float)M;
        for (int m = l; 2 * m < half; m += TS) {
            const float sh = sinpif((float)m * invM);
            const float dl = -4.f * sh * sh;
            float e = -0.5f * dl, c = 1.f;
            float E = 0.f, O = 0.f;
            for (int p4 = 0; p4 <= half; p4 += 4) {
                const float4 d4 = *reinterpret_cast<const float4*>(dr + p4);
                const float dv[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int p = p4 + k;
                    if (p > half) continue;
                    if (p & 1) O = fmaf(dv[k], c, O); else E = fmaf(dv[k], c, E);
                    e = fmaf(dl, c, e); c += e;
                }
            }
            const int m2 = half - m;
            gH[frame * NB + m] = (E + O) * ((m == 0 || m == half) ? invM : 2.f * invM);
            gH[frame * NB + m2] = (E - O) * ((m2 == 0 || m2 == half) ? invM : 2.f * invM);
        }
        if ((half & 1) == 0 && l == (half / 2) % TS) {   // m = half/2: cos(p pi/2) = 0, +-1
            float E = 0.f;
            for (int p = 0; p <= half; p += 4) E += dr[p];
            for (int p = 2; p <= half; p += 4) E -= dr[p];
            gH[frame * NB + half / 2] = E * (2.f * invM);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// small utility kernels
// ------------------------------------------------------------------------------------------------
__global__ void noise_uniform_kernel(float* __restrict__ out, uint2 key, uint32_t offset, uint32_t voice_offset, int T,
                                     int hop, long long total_quads) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total_quads) return;
    const int qpf = (hop + 3) >> 2;
    const long long frame = i / qpf;
    const int q = (int)(i % qpf);
    const uint32_t b = (uint32_t)(frame / T), t = (uint32_t)(frame % T);
    const float4 v = philox_noise4((uint32_t)q, t, voice_offset + b, offset, key);
    float* o = out + frame * hop + 4 * q;
    const float vals[4] = {v.x, v.y, v.z, v.w};
    for (int e = 0; e < 4; ++e)
        if (4 * q + e < hop) o[e] = vals[e];
}

// amp_to_impulse_response (:7-22) on its own: one CTA per row
__global__ void __launch_bounds__(256) amp_to_ir_kernel(const float* __restrict__ H, float* __restrict__ ir, int NB, int target) {
    extern __shared__ __align__(16) float smem[];
    const NoiseSmem s = carve_noise_smem(smem, NB, 0);
    const size_t row = blockIdx.x;
    const int half = NB - 1;
    for (int m = threadIdx.x; m < NB; m += blockDim.x) s.Hs[m] = H[row * NB + m];
    __syncthreads();
    build_taps(s, NB);
    float* o = ir + row * target;
    for (int i = threadIdx.x; i < target; i += blockDim.x) {
        float v = 0.f;
        if (i < half) v = s.hc[i];
        const int w = i - (target - half);
        if (w >= 0) v += s.ha[w];                        // M == target: both parts land on the same taps? no: ranges are disjoint
        o[i] = v;
    }
}

static int check_noise_shape(const char* fn, int B, int T, int NB, int hop) {
    DDSP_REQUIRE(B >= 1 && B <= 65535, "%s: B=%d out of range [1,65535]", fn, B);
    DDSP_REQUIRE(T >= 1, "%s: T=%d must be >= 1", fn, T);
    DDSP_REQUIRE(NB >= 2, "%s: NB=%d must be >= 2", fn, NB);
    DDSP_REQUIRE(hop >= 2 && hop <= 4096, "%s: hop=%d out of range [2,4096]", fn, hop);
    DDSP_REQUIRE(2 * (NB - 1) <= hop, "%s: FIR length 2(NB-1)=%d exceeds hop=%d (filtered_noise.py:19 pads by hop-M)", fn,
                 2 * (NB - 1), hop);
    return 0;
}

template <typename Kern>
static int reserve_smem(Kern kern, size_t bytes, const char* name) {
    DDSP_REQUIRE(bytes <= 227 * 1024, "%s: needs %zu B of shared memory (> 227 KB)", name, bytes);
    if (bytes > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        DDSP_REQUIRE(e == cudaSuccess, "%s: cannot reserve %zu B of shared memory: %s", name, bytes, cudaGetErrorString(e));
    }
    return 0;
}

}  // namespace ddsp

using namespace ddsp;

extern "C" int ddsp_noise_forward(const float* H, const float* noise, float* y, int accumulate, uint64_t seed,
                                  uint32_t offset, const uint32_t* offset_dev, uint32_t voice_offset, int B, int T, int NB,
                                  int hop, void* stream) {
    if (check_noise_shape("ddsp_noise_forward", B, T, NB, hop)) return 1;
    DDSP_REQUIRE(H && y, "ddsp_noise_forward: null pointer argument");
    const uint2 key = make_uint2((uint32_t)(seed & 0xffffffffu), (uint32_t)(seed >> 32));
    if (hop % 16 == 0 && hop / 16 <= TEAM_BLOCK && TEAM_BLOCK % (hop / 16) == 0) {
        const int FPB = TEAM_BLOCK / (hop / 16);
        const size_t tsm = sizeof(float) * (size_t)team_layout(NB, hop).total * FPB;
        if (tsm <= 200 * 1024) {
            const long long frames = (long long)B * T;
            const unsigned grid = (unsigned)((frames + FPB - 1) / FPB);
            cudaStream_t st = static_cast<cudaStream_t>(stream);
#define NOISE_FWD(NBc, HOPc)                                                                                         \
    do {                                                                                                             \
        if (reserve_smem(noise_fwd_team_kernel<NBc, HOPc>, tsm, "noise_fwd_team_kernel")) return 1;                  \
        noise_fwd_team_kernel<NBc, HOPc><<<grid, TEAM_BLOCK, tsm, st>>>(H, noise, y, accumulate, key, offset,        \
                                                                        offset_dev, voice_offset, T, NB, hop, frames);           \
    } while (0)
            if (NB == 65 && hop == 512) NOISE_FWD(65, 512);
            else if (NB == 195 && hop == 512) NOISE_FWD(195, 512);
            else NOISE_FWD(0, 0);
#undef NOISE_FWD
            DDSP_CHECK_LAUNCH("noise_fwd_team_kernel");
            return 0;
        }
    }
    const int hopp = (hop + 7) & ~7;
    const size_t smem = sizeof(float) * noise_smem_floats(NB, hopp);
    if (reserve_smem(noise_fwd_kernel, smem, "noise_fwd_kernel")) return 1;
    noise_fwd_kernel<<<dim3(T, B), hopp / 8, smem, static_cast<cudaStream_t>(stream)>>>(H, noise, y, accumulate, key, offset,
                                                                                      offset_dev, voice_offset, T, NB, hop);
    DDSP_CHECK_LAUNCH("noise_fwd_kernel");
    return 0;
}

extern "C" int ddsp_noise_backward(const float* g, const float* noise, float* gH, uint64_t seed, uint32_t offset,
                                   const uint32_t* offset_dev, uint32_t voice_offset, int B, int T, int NB, int hop,
                                   void* stream) {
    if (check_noise_shape("ddsp_noise_backward", B, T, NB, hop)) return 1;
    DDSP_REQUIRE(g && gH, "ddsp_noise_backward: null pointer argument");
    const uint2 key = make_uint2((uint32_t)(seed & 0xffffffffu), (uint32_t)(seed >> 32));
    if (hop % 16 == 0 && hop / 16 <= TEAM_BLOCK && TEAM_BLOCK % (hop / 16) == 0) {
        const int FPB = TEAM_BLOCK / (hop / 16);
        const size_t tsm = sizeof(float) * (size_t)team_bwd_layout(NB, hop).total * FPB;
        if (tsm <= 200 * 1024) {
            const long long frames = (long long)B * T;
            const unsigned grid = (unsigned)((frames + FPB - 1) / FPB);
            cudaStream_t st = static_cast<cudaStream_t>(stream);
#define NOISE_BWD(NBc, HOPc)                                                                                         \
    do {                                                                                                             \
        if (reserve_smem(noise_bwd_team_kernel<NBc, HOPc>, tsm, "noise_bwd_team_kernel")) return 1;                  \
        noise_bwd_team_kernel<NBc, HOPc><<<grid, TEAM_BLOCK, tsm, st>>>(g, noise, gH, key, offset, offset_dev, voice_offset, T,  \
                                                                        NB, hop, frames);                            \
    } while (0)
            if (NB == 65 && hop == 512) NOISE_BWD(65, 512);
            else if (NB == 195 && hop == 512) NOISE_BWD(195, 512);
            else NOISE_BWD(0, 0);
#undef NOISE_BWD
            DDSP_CHECK_LAUNCH("noise_bwd_team_kernel");
            return 0;
        }
    }
    const int hopp = (hop + 7) & ~7;
    const int nsl = noise_bwd_slices(NB, hopp);
    const int threads = noise_tp(NB) / 8 * nsl;
    DDSP_REQUIRE(threads <= 512, "ddsp_noise_backward: NB=%d needs %d threads per frame (max 512)", NB, threads);
    const size_t smem = sizeof(float) * noise_bwd_smem_floats(NB, hopp);
    if (reserve_smem(noise_bwd_kernel, smem, "noise_bwd_kernel")) return 1;
    noise_bwd_kernel<<<dim3(T, B), threads, smem, static_cast<cudaStream_t>(stream)>>>(g, noise, gH, key, offset, offset_dev,
                                                                                      voice_offset, T, NB, hop, nsl);
    DDSP_CHECK_LAUNCH("noise_bwd_kernel");
    return 0;
}

extern "C" int ddsp_noise_uniform(float* out, uint64_t seed, uint32_t offset, uint32_t voice_offset, int B, int T, int hop,
                                  void* stream) {
    DDSP_REQUIRE(out && B >= 1 && T >= 1 && hop >= 1, "ddsp_noise_uniform: bad arguments");
    const long long quads = (long long)B * T * ((hop + 3) >> 2);
    const uint2 key = make_uint2((uint32_t)(seed & 0xffffffffu), (uint32_t)(seed >> 32));
    noise_uniform_kernel<<<(unsigned)((quads + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        out, key, offset, voice_offset, T, hop, quads);
    DDSP_CHECK_LAUNCH("noise_uniform_kernel");
    return 0;
}

extern "C" int ddsp_amp_to_impulse_response(const float* H, float* ir, int R, int NB, int target, void* stream) {
    DDSP_REQUIRE(H && ir && R >= 1 && NB >= 2, "ddsp_amp_to_impulse_response: bad arguments");
    DDSP_REQUIRE(2 * (NB - 1) <= target, "ddsp_amp_to_impulse_response: FIR length %d exceeds target %d", 2 * (NB - 1), target);
    const size_t smem = sizeof(float) * noise_smem_floats(NB, 0);
    if (reserve_smem(amp_to_ir_kernel, smem, "amp_to_ir_kernel")) return 1;
    amp_to_ir_kernel<<<R, 256, smem, static_cast<cudaStream_t>(stream)>>>(H, ir, NB, target);
    DDSP_CHECK_LAUNCH("amp_to_ir_kernel");
    return 0;
}
