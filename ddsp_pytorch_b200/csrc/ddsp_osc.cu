// Harmonic oscillator bank for sm_100a -- replaces OscillatorBank (harmonic_oscillator.py:24-62).
//
// The reference materialises ~10 tensors of shape [B,N,K] (upsampled rad/sample, cumsum, sin, ...).
// Here nothing of that shape ever exists:
//   osc_prepare_kernel   frame rate: Nyquist mask + normalisation of c (:31-33), active-harmonic
//                        count, and an fp64 scan of the fundamental's phase over FRAMES (the
//                        cumsum over samples, :41, has a closed form inside each hop).
//   osc_synth_kernel     one CTA per (voice, hop segment): per sample one fp64 phase -> (sin, delta)
//                        seed, then all harmonics by the Reinsch sine recurrence in registers
//                        (2 FP32 ops per harmonic-sample, no MUFU), two FFMA dot products with the
//                        segment's amplitude pair held in shared memory, float4 store.
//   osc_grad_kernel      same recurrence, accumulates g*a_up*sin and g*a_up*w*sin per harmonic,
//                        reduced over the segment's samples through shared memory.
//   osc_finalize_kernel  frame rate: gathers the two segments touching each frame and runs the
//                        backward of mask+normalise.
// Harmonics masked in both frames of a segment have zero amplitude for the whole segment and are
// skipped (the loop runs to the segment's active count), exactly as multiplying by zero would.
#include <atomic>
#include "ddsp_common.cuh"

namespace ddsp {

// Segment s interpolates frames s-1 and s and holds the samples whose half-pixel position (n + 1/2)/hop - 1/2 lies in
// [s-1, s): n0 = s hop - ceil(hop/2) is its first sample and sample r sits at fraction w = (r + phi)/hop between the two
// frames, phi = 1/2 for even hops, 0 for odd hops (a sample then falls exactly on every frame centre).
__host__ __device__ inline int half_up(int hop) { return (hop + 1) >> 1; }
__host__ __device__ inline float seg_phi(int hop) { return (hop & 1) ? 0.f : 0.5f; }


// ------------------------------------------------------------------------------------------------
// frame-rate preparation: one CTA per voice
// ------------------------------------------------------------------------------------------------
constexpr int PREP_THREADS = 256;
constexpr int PREP_TILE = 32;          // frames normalised per CTA

__global__ void __launch_bounds__(PREP_THREADS) osc_prepare_kernel(
    const float* __restrict__ f0, const float* __restrict__ c, float* __restrict__ cn, float* __restrict__ csum,
    int* __restrict__ kact, double* __restrict__ phase0, SegFix* __restrict__ seg, int T, int K, int hop, int sample_rate) {
    const int b = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = PREP_THREADS / 32;
    __shared__ double warp_tot[NW];
    __shared__ double carry_s;
    const float* f0b = f0 + (size_t)b * T;

    // blockIdx.y >= 1: mask + normalise a tile of PREP_TILE frames, one warp per frame (harmonic_oscillator.py:31-33);
    // blockIdx.y == 0: the two phase scans of the voice
    if (blockIdx.y > 0) {
        const float nyq = (float)(sample_rate / 2);        // integer floor, compared as float (:31)
        const int t_lo = (blockIdx.y - 1) * PREP_TILE, t_hi = min(T, t_lo + PREP_TILE);
        for (int t = t_lo + warp; t < t_hi; t += NW) {
            const float f = f0b[t];
            const float* cr = c + ((size_t)b * T + t) * K;
            float* cnr = cn + ((size_t)b * T + t) * K;
            float sum = 0.f;
            int last = 0;
            float vreg[8];                                  // K <= 256: the frame's row stays in registers between the two sweeps
            if (K <= 256) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int k = lane + 32 * i;
                    const bool masked = (float)(k + 1) * f > nyq;
                    vreg[i] = (k < K && !masked) ? cr[k] : 0.f;
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int k = lane + 32 * i;
                    if (k < K) {
                        sum += vreg[i];
                        if (!((float)(k + 1) * f > nyq)) last = k + 1;
                    }
                }
            } else {
                for (int k = lane; k < K; k += 32) {
                    const bool masked = (float)(k + 1) * f > nyq;
                    const float v = masked ? 0.f : cr[k];
                    sum += v;
                    if (!masked) last = k + 1;
                }
            }
            sum = warp_sum(sum);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) last = max(last, __shfl_xor_sync(0xffffffffu, last, o));
            if (!(sum > 0.f)) last = K;                     // 0/0 or NaN frame: keep every harmonic so NaN propagates like :33
            if (K <= 256) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int k = lane + 32 * i;
                    if (k < K) cnr[k] = vreg[i] / sum;
                }
            } else {
                for (int k = lane; k < K; k += 32) {
                    const bool masked = (float)(k + 1) * f > nyq;
                    cnr[k] = (masked ? 0.f : cr[k]) / sum;
                }
            }
            if (lane == 0) {
                csum[(size_t)b * T + t] = sum;
                kact[(size_t)b * T + t] = last;
            }
        }
        return;
    }

    // ---- phase prefix over segments, fp64, in cycles -------------------------------------------
    // hu = ceil(hop/2) samples of segment 0 lie before n = 0:  E_0 = -hu u_0.  Summing u over a whole segment,
    // E_{s+1} = E_s + hu u_L + (hop - hu) u_R   (= hop (u_L + u_R)/2 for even hop)         (u = f0 / sample_rate)
    double* Eb = phase0 + (size_t)b * (T + 1);
    const double inv_sr = 1.0 / (double)sample_rate;
    if (tid == 0) {
        const double e0 = -(double)half_up(hop) * ((double)f0b[0] * inv_sr);
        carry_s = e0;
        Eb[0] = e0 - floor(e0);
    }
    __syncthreads();
    for (int base = 0; base < T; base += PREP_THREADS) {
        const int s = base + tid;
        double v = 0.0;
        if (s < T) {
            const double uL = (double)f0b[s > 0 ? s - 1 : 0] * inv_sr;
            const double uR = (double)f0b[s] * inv_sr;
            v = (double)half_up(hop) * uL + (double)(hop - half_up(hop)) * uR;
        }
        // inclusive scan inside the warp, then across warps
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const double n = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += n;
        }
        if (lane == 31) warp_tot[warp] = v;
        __syncthreads();
        double pre = carry_s;
        for (int w = 0; w < warp; ++w) pre += warp_tot[w];
        const double e = pre + v;                       // E_{s+1}
        if (s < T) Eb[s + 1] = e - floor(e);
        __syncthreads();
        if (tid == PREP_THREADS - 1) carry_s = e - floor(e);   // keep the carry small: only the fraction matters
        __syncthreads();
    }

    // ---- the same prefix in 0.64 fixed point (exact integer scan), with each segment's increment pair --------
    {
        __shared__ unsigned long long warp_tot_u[NW];
        __shared__ unsigned long long carry_u;
        SegFix* sb = seg + (size_t)b * (T + 1);
        const double two64 = 18446744073709551616.0;
        auto ufix = [&](int t) { return (unsigned long long)__double2ll_rn((double)f0b[t] * inv_sr * two64); };
        const unsigned long long hhL = (unsigned long long)half_up(hop), hhR = (unsigned long long)(hop - half_up(hop));
        if (tid == 0) carry_u = 0ull - hhL * ufix(0);
        __syncthreads();
        for (int base = 0; base <= T; base += PREP_THREADS) {
            const int s = base + tid;                      // segment index, E_s = E_0 + sum_{s'<s} (hu uL + (hop-hu) uR)
            unsigned long long uL = 0, uR = 0, v = 0;
            if (s <= T) {
                uL = ufix(s > 0 ? s - 1 : 0);
                uR = ufix(s < T ? s : T - 1);
                v = hhL * uL + hhR * uR;
            }
            unsigned long long inc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned long long n = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += n;
            }
            if (lane == 31) warp_tot_u[warp] = inc;
            __syncthreads();
            unsigned long long pre = carry_u;
            for (int w = 0; w < warp; ++w) pre += warp_tot_u[w];
            if (s <= T) {
                SegFix g;
                g.E = pre + inc - v;                       // exclusive prefix
                g.du = (long long)(uR - uL) / (2ll * hop);
                // Theta(r) = E + (r+1) uL + du (r+1)(r + 2 phi): phi = 1/2 (even hop) gives du (r+1)^2; phi = 0 (odd hop)
                // gives du (r+1)^2 - du (r+1), folded into the linear coefficient
                g.uL = (hop & 1) ? uL - (unsigned long long)g.du : uL;
                sb[s] = g;
            }
            __syncthreads();
            if (tid == PREP_THREADS - 1) carry_u = pre + inc;
            __syncthreads();
        }
    }
}

// Everything a thread needs about its segment.
struct Segment {
    int tL, tR;        // frames interpolated
    int r_lo, r_hi;    // valid sample range inside the segment
    long long n0;      // global sample index of r = 0 (may be negative for s = 0)
    double E, uL, du;  // phase prefix, left increment, (uR-uL)/(2*hop)    [cycles]
};

__device__ __forceinline__ Segment make_segment(int s, int T, int hop, const float* __restrict__ f0b,
                                                const double* __restrict__ Eb, double inv_sr) {
    Segment g;
    g.tL = s > 0 ? s - 1 : 0;
    g.tR = s < T ? s : T - 1;
    g.r_lo = s == 0 ? half_up(hop) : 0;
    g.r_hi = s == T ? half_up(hop) : hop;
    g.n0 = (long long)s * hop - half_up(hop);
    g.E = Eb[s];
    g.uL = (double)f0b[g.tL] * inv_sr;
    g.du = ((double)f0b[g.tR] * inv_sr - g.uL) / (2.0 * hop);
    if (hop & 1) g.uL -= g.du;                         // odd hop: see osc_prepare_kernel
    return g;
}

__device__ __forceinline__ PhaseSeed seed_at(const Segment& g, int r) {
    const double rr = (double)(r + 1);
    return phase_seed(wrap_cycles(fma(g.du * rr, rr, fma(rr, g.uL, g.E))));
}

// ------------------------------------------------------------------------------------------------
// Per-thread seeds for R consecutive samples r0..r0+R-1 of a segment, packed in pairs for FFMA2/FADD2
// (fma.rn.f32x2: one issue slot per two FMAs, which is what lets the FMA pipe fill -- profiles/summary_r01c.md).
// The phase runs in 0.64 fixed point by second differences: two 64-bit adds per sample.
// ------------------------------------------------------------------------------------------------
template <int RP>
__device__ __forceinline__ void seed_pairs(const SegFix& g, int r0, float2 (&sn)[RP], float2 (&dl)[RP], float2 (&sg)[RP]) {
    const unsigned long long R1 = (unsigned long long)(r0 + 1);
    const unsigned long long du = (unsigned long long)g.du;
    unsigned long long ph = g.E + R1 * g.uL + (R1 * R1) * du;
    unsigned long long D = g.uL + (2ull * R1 + 1ull) * du;         // Theta(r+1) - Theta(r)
    const unsigned long long D2 = 2ull * du;
#pragma unroll
    for (int j = 0; j < RP; ++j) {
        const PhaseSeed p0 = phase_seed_fix(ph);
        ph += D; D += D2;
        const PhaseSeed p1 = phase_seed_fix(ph);
        ph += D; D += D2;
        sn[j] = make_float2(p0.sin1, p1.sin1);
        dl[j] = make_float2(p0.delta, p1.delta);
        sg[j] = make_float2(p0.sigma, p1.sigma);
    }
}

struct SegGeom {
    int tL, tR;        // frames interpolated
    int r_lo, r_hi;    // valid sample range inside the segment
    long long n0;      // global sample index of r = 0 (negative for s = 0)
};

__device__ __forceinline__ SegGeom seg_geom(int s, int T, int hop) {
    SegGeom g;
    g.tL = s > 0 ? s - 1 : 0;
    g.tR = s < T ? s : T - 1;
    g.r_lo = s == 0 ? half_up(hop) : 0;
    g.r_hi = s == T ? half_up(hop) : hop;
    g.n0 = (long long)s * hop - half_up(hop);
    return g;
}

// ------------------------------------------------------------------------------------------------
// synthesis: grid (T+1, B), hop/R threads, R consecutive samples per thread (R = 2*RP)
// ------------------------------------------------------------------------------------------------
template <int RP>
// register caps instead of launch bounds (they exclude each other): 8 samples per thread (the shipped hop 512) measured
// fastest at 72 registers -- 76 uncapped, 3.32 -> 3.24 ms at cfg 4; 64 spills into the harmonic loop (3.51 ms)
__global__ void __maxnreg__(RP == 4 ? 72 : (RP >= 8 ? 255 : 128)) osc_synth_kernel(
    const float* __restrict__ cn, const float* __restrict__ a, const int* __restrict__ kact,
    const SegFix* __restrict__ seg, float* __restrict__ y, float* __restrict__ z_save, int accumulate, int T, int K, int hop) {
    // (c'_L, dc') per active harmonic.  Each value multiplies BOTH samples of a pair: written as (v, v) the compiler emits the
    // scalar-broadcast operand form of FFMA2 (R.F32), one register-file read instead of two (sass_sched.py has the arithmetic)
    extern __shared__ __align__(16) float2 amp2[];
    constexpr int R = 2 * RP;
    const int s = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
    const SegGeom g = seg_geom(s, T, hop);
    const size_t fL = (size_t)b * T + g.tL, fR = (size_t)b * T + g.tR;
    const int kmax = max(kact[fL], kact[fR]);
    const int kpad = (kmax + 1) & ~1;
    {
        const float* cl = cn + fL * K;
        const float* cr = cn + fR * K;
        for (int k = tid; k < kpad; k += blockDim.x) {
            float l = 0.f, d = 0.f;
            if (k < kmax) {
                l = cl[k];
                d = cr[k] - l;
            }
            amp2[k] = make_float2(l, d);
        }
    }
    __syncthreads();

    const int r0 = tid * R;
    if (r0 >= g.r_hi || r0 + R <= g.r_lo) return;      // the half-filled first / last segment

    float2 sn[RP], dn[RP], dl[RP], sg[RP], Pe[RP], Po[RP], Qe[RP], Qo[RP];
    seed_pairs<RP>(seg[(size_t)b * (T + 1) + s], r0, sn, dl, sg);
#pragma unroll
    for (int j = 0; j < RP; ++j) {
        dn[j] = sn[j];                                 // d_1 = s_1 - s_0, s_0 = 0
        Pe[j] = Po[j] = Qe[j] = Qo[j] = make_float2(0.f, 0.f);
    }

    // Reinsch recurrence over harmonics m = 1..kmax:  d_{m+1} = d_m + delta*s_m ; s_{m+1} = s_m + d_{m+1}
#pragma unroll 2
    for (int k = 0; k < kpad; k += 2) {
        const float4 AE = *reinterpret_cast<const float4*>(amp2 + k);   // odd harmonic k+1, even harmonic k+2
        const float2 Ac = make_float2(AE.x, AE.x), Ad = make_float2(AE.y, AE.y);
        const float2 Ec = make_float2(AE.z, AE.z), Ed = make_float2(AE.w, AE.w);
#pragma unroll
        for (int j = 0; j < RP; ++j) {
            Po[j] = __ffma2_rn(Ac, sn[j], Po[j]);
            Qo[j] = __ffma2_rn(Ad, sn[j], Qo[j]);
            dn[j] = __ffma2_rn(dl[j], sn[j], dn[j]);
            sn[j] = __fadd2_rn(sn[j], dn[j]);
        }
#pragma unroll
        for (int j = 0; j < RP; ++j) {
            Pe[j] = __ffma2_rn(Ec, sn[j], Pe[j]);
            Qe[j] = __ffma2_rn(Ed, sn[j], Qe[j]);
            dn[j] = __ffma2_rn(dl[j], sn[j], dn[j]);
            sn[j] = __fadd2_rn(sn[j], dn[j]);
        }
    }

    const float aL = a[fL], da = a[fR] - aL;
    const float inv_hop = 1.0f / (float)hop;
    const float wb = ((float)r0 + seg_phi(hop)) * inv_hop;        // fraction of the thread's first sample between the frames
    float out[R], zs[R];
#pragma unroll
    for (int j = 0; j < RP; ++j) {
        const float2 zP = __ffma2_rn(sg[j], Po[j], Pe[j]);     // sin(m theta) = sigma^m sin(m theta')
        const float2 zQ = __ffma2_rn(sg[j], Qo[j], Qe[j]);
        const float w0 = fmaf((float)(2 * j), inv_hop, wb), w1 = fmaf((float)(2 * j + 1), inv_hop, wb);
        zs[2 * j] = fmaf(w0, zQ.x, zP.x);
        zs[2 * j + 1] = fmaf(w1, zQ.y, zP.y);
        out[2 * j] = fmaf(w0, da, aL) * zs[2 * j];
        out[2 * j + 1] = fmaf(w1, da, aL) * zs[2 * j + 1];
    }
    const size_t N = (size_t)T * hop;
    float* yp = y + (size_t)b * N + (g.n0 + r0);
    float* zp = z_save ? z_save + (size_t)b * N + (g.n0 + r0) : nullptr;
    if (R % 4 == 0) {                                  // n0 + r0 is then a multiple of 4 (hop % (2R) == 0)
#pragma unroll
        for (int j = 0; j < R; j += 4) {
            float4 o = make_float4(out[j], out[j + 1], out[j + 2], out[j + 3]);
            if (accumulate) {
                const float4 p = *reinterpret_cast<const float4*>(yp + j);
                o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w;
            }
            *reinterpret_cast<float4*>(yp + j) = o;
            if (zp) *reinterpret_cast<float4*>(zp + j) = make_float4(zs[j], zs[j + 1], zs[j + 2], zs[j + 3]);
        }
    } else {
#pragma unroll
        for (int j = 0; j < R; ++j) {
            if (r0 + j < g.r_lo || r0 + j >= g.r_hi) continue;
            yp[j] = accumulate ? yp[j] + out[j] : out[j];
            if (zp) zp[j] = zs[j];
        }
    }
}

// ------------------------------------------------------------------------------------------------
// d/d f0 (harmonic_oscillator.py:26-43 under autograd: f0 -> k*f0 -> rad/sample -> upsample -> cumsum -> sin).
// With Theta_n the running phase of the fundamental in cycles, y_n = a_n sum_k c'_k,n sin(2 pi k Theta_n) and
// d Theta_n / d u_i = 1 for i <= n (u = upsampled f0 / sample_rate), so
//     dL/du_i = 2 pi * sum_{n >= i} q_n,   q_n = g_n a_n sum_k k c'_k,n cos(2 pi k Theta_n),
// and dL/df0 is the adjoint of the linear upsampling applied to that reverse running sum, over sample_rate.  The Nyquist
// mask and the amplitude normalisation do not depend on f0 differentiably (a comparison), exactly as in the reference.
// osc_dphase_kernel is the synthesis kernel with cosine seeds and amplitudes scaled by the harmonic number; it writes q.
// ------------------------------------------------------------------------------------------------
template <int RP>
__global__ void __maxnreg__(RP == 4 ? 80 : (RP >= 8 ? 255 : 128)) osc_dphase_kernel(
    const float* __restrict__ cn, const float* __restrict__ a, const int* __restrict__ kact, const SegFix* __restrict__ seg,
    const float* __restrict__ gup, float* __restrict__ q_out, int T, int K, int hop) {
    extern __shared__ __align__(16) float4 amp[];      // (k c'_L, k c'_L, k dc', k dc') per active harmonic k
    constexpr int R = 2 * RP;
    const int s = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
    const SegGeom g = seg_geom(s, T, hop);
    const size_t fL = (size_t)b * T + g.tL, fR = (size_t)b * T + g.tR;
    const int kmax = max(kact[fL], kact[fR]);
    const int kpad = (kmax + 1) & ~1;
    {
        const float* cl = cn + fL * K;
        const float* cr = cn + fR * K;
        for (int k = tid; k < kpad; k += blockDim.x) {
            float l = 0.f, d = 0.f;
            if (k < kmax) {
                const float hk = (float)(k + 1);
                l = hk * cl[k];
                d = hk * cr[k] - l;
            }
            amp[k] = make_float4(l, l, d, d);
        }
    }
    __syncthreads();

    const int r0 = tid * R;
    if (r0 >= g.r_hi || r0 + R <= g.r_lo) return;

    float2 cs[RP], dn[RP], dl[RP], sg[RP], Pe[RP], Po[RP], Qe[RP], Qo[RP];
    seed_pairs<RP>(seg[(size_t)b * (T + 1) + s], r0, cs, dl, sg);       // cs receives sin(theta') here; replaced below
#pragma unroll
    for (int j = 0; j < RP; ++j) {
        dn[j] = make_float2(0.5f * dl[j].x, 0.5f * dl[j].y);            // cos(theta') - cos(0) = delta/2
        cs[j] = make_float2(1.f + dn[j].x, 1.f + dn[j].y);              // cos(theta') = 1 - 2 sin^2(theta'/2)
        Pe[j] = Po[j] = Qe[j] = Qo[j] = make_float2(0.f, 0.f);
    }
    // the same Reinsch recurrence runs cosines: e_{m+1} = e_m + delta*c_m ; c_{m+1} = c_m + e_{m+1}
#pragma unroll 2
    for (int k = 0; k < kpad; k += 2) {
        const float4 A = amp[k], E = amp[k + 1];
        const float2 Ac = make_float2(A.x, A.y), Ad = make_float2(A.z, A.w);
        const float2 Ec = make_float2(E.x, E.y), Ed = make_float2(E.z, E.w);
#pragma unroll
        for (int j = 0; j < RP; ++j) {
            Po[j] = __ffma2_rn(Ac, cs[j], Po[j]);
            Qo[j] = __ffma2_rn(Ad, cs[j], Qo[j]);
            dn[j] = __ffma2_rn(dl[j], cs[j], dn[j]);
            cs[j] = __fadd2_rn(cs[j], dn[j]);
        }
#pragma unroll
        for (int j = 0; j < RP; ++j) {
            Pe[j] = __ffma2_rn(Ec, cs[j], Pe[j]);
            Qe[j] = __ffma2_rn(Ed, cs[j], Qe[j]);
            dn[j] = __ffma2_rn(dl[j], cs[j], dn[j]);
            cs[j] = __fadd2_rn(cs[j], dn[j]);
        }
    }
    const float aL = a[fL], da = a[fR] - aL;
    const float inv_hop = 1.0f / (float)hop;
    const float wb = ((float)r0 + seg_phi(hop)) * inv_hop;
    const size_t N = (size_t)T * hop;
    const float* gp = gup + (size_t)b * N + (g.n0 + r0);
    float* qp = q_out + (size_t)b * N + (g.n0 + r0);
#pragma unroll
    for (int j = 0; j < RP; ++j) {
        const float2 zP = __ffma2_rn(sg[j], Po[j], Pe[j]);              // cos(m theta) = sigma^m cos(m theta')
        const float2 zQ = __ffma2_rn(sg[j], Qo[j], Qe[j]);
        const float w0 = fmaf((float)(2 * j), inv_hop, wb), w1 = fmaf((float)(2 * j + 1), inv_hop, wb);
        const int ra = r0 + 2 * j, rb = ra + 1;
        if (ra >= g.r_lo && ra < g.r_hi) qp[2 * j] = gp[2 * j] * fmaf(w0, da, aL) * fmaf(w0, zQ.x, zP.x);
        if (rb >= g.r_lo && rb < g.r_hi) qp[2 * j + 1] = gp[2 * j + 1] * fmaf(w1, da, aL) * fmaf(w1, zQ.y, zP.y);
    }
}

// Q[n] = scale * sum_{m >= n} q[m] per voice (fp64 carry and partial sums: N reaches 10^5..10^6 terms), written over q.
__global__ void __launch_bounds__(256) osc_revscan_kernel(float* __restrict__ q, long long N, double scale) {
    constexpr int PER = 8, CHUNK = 256 * PER;
    __shared__ double wsum[8];
    __shared__ double carry_s;
    float* qb = q + (size_t)blockIdx.x * N;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) carry_s = 0.0;
    __syncthreads();
    for (long long hi = N; hi > 0; hi -= CHUNK) {
        // thread t owns the PER samples [hi - (t+1)*PER, hi - t*PER): later samples first
        const long long top = hi - (long long)tid * PER;
        double v[PER];
        double tot = 0.0;
#pragma unroll
        for (int e = 0; e < PER; ++e) {
            const long long idx = top - 1 - e;
            v[e] = idx >= 0 ? (double)qb[idx] : 0.0;
            tot += v[e];
        }
        double inc = tot;                                   // inclusive scan over threads (thread 0 = latest samples)
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const double up = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += up;
        }
        if (lane == 31) wsum[wid] = inc;
        __syncthreads();
        double before = carry_s;
        for (int w = 0; w < wid; ++w) before += wsum[w];
        double run = before + (inc - tot);                  // everything later than this thread's samples
#pragma unroll
        for (int e = 0; e < PER; ++e) {
            const long long idx = top - 1 - e;
            run += v[e];
            if (idx >= 0) qb[idx] = (float)(run * scale);
        }
        __syncthreads();
        if (tid == 255) carry_s = run;
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------
// backward over samples: grid (T+1, B), hop/R threads.
//   G0[b,s,k] = sum_r g*a_up*sin(k theta),  G1[b,s,k] = sum_r g*a_up*w*sin(k theta)     (segment s)
//   A0[b,s]   = sum_r g*z,                  A1[b,s]   = sum_r g*z*w
// Harmonics are processed in chunks of CH; every thread parks its R-sample partial sums in shared
// memory ([CH][NT] float2, rows padded by 4 floats so the float4 row sweeps are conflict-free) and
// one thread per harmonic sums its row.
// ------------------------------------------------------------------------------------------------
constexpr int GRAD_CH = 32;

template <int RP>
// 16 samples per thread: 144 registers (158 uncapped; 128 spills) -- 3.78 -> 3.76 ms at cfg 4
__global__ void __maxnreg__(RP >= 8 ? 144 : 128) osc_grad_kernel(
    const float* __restrict__ gup, const float* __restrict__ a, const float* __restrict__ z_save,
    const int* __restrict__ kact, const SegFix* __restrict__ seg, float2* __restrict__ gseg, float2* __restrict__ gaseg,
    int T, int K, int hop) {
    extern __shared__ __align__(16) float part[];      // [GRAD_CH][2*NT + 4]
    __shared__ float red[2][16];
    constexpr int R = 2 * RP;
    const int s = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, NT = blockDim.x;
    const int stride = ((2 * NT + 3) & ~3) + 4;          // multiple of 4 floats: the float4 row sweeps stay 16-byte aligned
    const SegGeom g = seg_geom(s, T, hop);
    const size_t fL = (size_t)b * T + g.tL, fR = (size_t)b * T + g.tR;
    const int kmax = max(kact[fL], kact[fR]);

    const int r0 = tid * R;
    const bool live = !(r0 >= g.r_hi || r0 + R <= g.r_lo);
    float2 sn[RP], dn[RP], dl[RP], sg[RP], ug[RP], uw[RP], ugo[RP], uwo[RP];
    float ga0 = 0.f, ga1 = 0.f;
    if (live) {
        seed_pairs<RP>(seg[(size_t)b * (T + 1) + s], r0, sn, dl, sg);
        const float aL = a[fL], da = a[fR] - aL;
        const float inv_hop = 1.0f / (float)hop;
        const size_t N = (size_t)T * hop;
        const float* gp = gup + (size_t)b * N + (g.n0 + r0);
        const float* zp = z_save + (size_t)b * N + (g.n0 + r0);
        float gj[R], zj[R];
        if (R % 4 == 0) {
#pragma unroll
            for (int j = 0; j < R; j += 4) {
                const float4 gv = *reinterpret_cast<const float4*>(gp + j), zv = *reinterpret_cast<const float4*>(zp + j);
                gj[j] = gv.x; gj[j + 1] = gv.y; gj[j + 2] = gv.z; gj[j + 3] = gv.w;
                zj[j] = zv.x; zj[j + 1] = zv.y; zj[j + 2] = zv.z; zj[j + 3] = zv.w;
            }
        } else {
#pragma unroll
            for (int j = 0; j < R; ++j) {
                const bool in = r0 + j >= g.r_lo && r0 + j < g.r_hi;
                gj[j] = in ? gp[j] : 0.f;
                zj[j] = in ? zp[j] : 0.f;
            }
        }
        float u1[R], u2[R];
#pragma unroll
        for (int j = 0; j < R; ++j) {
            const float w = ((float)(r0 + j) + seg_phi(hop)) * inv_hop;
            const float gz = gj[j] * zj[j];
            ga0 += gz;
            ga1 = fmaf(gz, w, ga1);
            u1[j] = gj[j] * fmaf(w, da, aL);
            u2[j] = u1[j] * w;
        }
#pragma unroll
        for (int j = 0; j < RP; ++j) {
            dn[j] = sn[j];
            ug[j] = make_float2(u1[2 * j], u1[2 * j + 1]);
            uw[j] = make_float2(u2[2 * j], u2[2 * j + 1]);
            ugo[j] = __fmul2_rn(ug[j], sg[j]);
            uwo[j] = __fmul2_rn(uw[j], sg[j]);
        }
    }

    float2* gout = gseg + ((size_t)b * (T + 1) + s) * K;
    for (int k0 = 0; k0 < kmax; k0 += GRAD_CH) {
        const int kn = min(GRAD_CH, kmax - k0);
        const int kn2 = (kn + 1) & ~1;
        if (live) {
            for (int kk = 0; kk < kn2; kk += 2) {
                float2 p0o = make_float2(0.f, 0.f), p1o = p0o, p0e = p0o, p1e = p0o;
#pragma unroll
                for (int j = 0; j < RP; ++j) {
                    p0o = __ffma2_rn(ugo[j], sn[j], p0o);
                    p1o = __ffma2_rn(uwo[j], sn[j], p1o);
                    dn[j] = __ffma2_rn(dl[j], sn[j], dn[j]);
                    sn[j] = __fadd2_rn(sn[j], dn[j]);
                }
#pragma unroll
                for (int j = 0; j < RP; ++j) {
                    p0e = __ffma2_rn(ug[j], sn[j], p0e);
                    p1e = __ffma2_rn(uw[j], sn[j], p1e);
                    dn[j] = __ffma2_rn(dl[j], sn[j], dn[j]);
                    sn[j] = __fadd2_rn(sn[j], dn[j]);
                }
                *reinterpret_cast<float2*>(&part[kk * stride + 2 * tid]) = make_float2(p0o.x + p0o.y, p1o.x + p1o.y);
                *reinterpret_cast<float2*>(&part[(kk + 1) * stride + 2 * tid]) = make_float2(p0e.x + p0e.y, p1e.x + p1e.y);
            }
        } else {
            for (int kk = 0; kk < kn2; ++kk)
                *reinterpret_cast<float2*>(&part[kk * stride + 2 * tid]) = make_float2(0.f, 0.f);
        }
        __syncthreads();
        for (int row = tid; row < kn; row += NT) {
            const float4* p4 = reinterpret_cast<const float4*>(&part[row * stride]);
#ifdef OSC_SWEEP_PACKED                               // measured: 3.757 -> 3.784 ms at cfg 4 (profiles/summary_r02.md); off
            float2 sa = make_float2(0.f, 0.f), sb = sa;
#pragma unroll 8
            for (int c4 = 0; c4 < NT / 2; ++c4) {
                const float4 v = p4[c4];
                sa = __fadd2_rn(sa, make_float2(v.x, v.y));
                sb = __fadd2_rn(sb, make_float2(v.z, v.w));
            }
            float s0 = sa.x, s1 = sa.y;
            const float t0 = sb.x, t1 = sb.y;
#else
            float s0 = 0.f, s1 = 0.f, t0 = 0.f, t1 = 0.f;
            for (int c4 = 0; c4 < NT / 2; ++c4) {
                const float4 v = p4[c4];
                s0 += v.x; s1 += v.y; t0 += v.z; t1 += v.w;
            }
#endif
            if (NT & 1) {                                   // odd thread count (tiny hops): last column
                s0 += part[row * stride + 2 * (NT - 1)];
                s1 += part[row * stride + 2 * (NT - 1) + 1];
            }
            gout[k0 + row] = make_float2(s0 + t0, s1 + t1);
        }
        __syncthreads();
    }

    // loudness partials: block reduction
    ga0 = warp_sum(ga0);
    ga1 = warp_sum(ga1);
    if ((tid & 31) == 0) { red[0][tid >> 5] = ga0; red[1][tid >> 5] = ga1; }
    __syncthreads();
    if (tid == 0) {
        float x0 = 0.f, x1 = 0.f;
        for (int w = 0; w < (NT + 31) / 32; ++w) { x0 += red[0][w]; x1 += red[1][w]; }
        gaseg[(size_t)b * (T + 1) + s] = make_float2(x0, x1);
    }
}

// ------------------------------------------------------------------------------------------------
// frame-rate backward: one warp per frame. Frame t is the LEFT frame of segment t+1 (and of segment 0
// when t = 0) with weight 1-w, and the RIGHT frame of segment t (and of segment T when t = T-1) with
// weight w.  Then the backward of c' = m*c / sum(m*c)   (harmonic_oscillator.py:32-33).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) osc_finalize_kernel(
    const float2* __restrict__ gseg, const float2* __restrict__ gaseg, const float* __restrict__ cn,
    const float* __restrict__ csum, const int* __restrict__ kact, const float* __restrict__ f0,
    float* __restrict__ gc, float* __restrict__ ga, int B, int T, int K, int sample_rate) {
    const int lane = threadIdx.x & 31;
    const long long frame = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (frame >= (long long)B * T) return;
    const int b = (int)(frame / T), t = (int)(frame % T);
    const float2* segR = gseg + ((size_t)b * (T + 1) + t) * K;        // segment t   : frame t is RIGHT
    const float2* segL = segR + K;                                     // segment t+1 : frame t is LEFT
    const float2* seg0 = gseg + ((size_t)b * (T + 1)) * K;            // segment 0   : frame 0 is also LEFT
    const float2* segT = gseg + ((size_t)b * (T + 1) + T) * K;        // segment T   : frame T-1 is also RIGHT
    const int kt = kact[frame];
    const float S = csum[frame];
    const float f = f0[frame];
    const float nyq = (float)(sample_rate / 2);
    const float* cnr = cn + (size_t)frame * K;
    float* gcr = gc + (size_t)frame * K;

    auto grad_cn = [&](int k) {
        const float2 l = segL[k], r = segR[k];
        float v = (l.x - l.y) + r.y;
        if (t == 0) { const float2 e = seg0[k]; v += e.x - e.y; }
        if (t == T - 1) { const float2 e = segT[k]; v += e.y; }
        return v;
    };
    float dot = 0.f;
    // (a one-pass variant with the frame's values parked in registers and all loads in flight measured slower: 0.33 -> 0.39 ms)
    for (int k = lane; k < kt; k += 32) dot = fmaf(grad_cn(k), cnr[k], dot);
    dot = warp_sum(dot);
    for (int k = lane; k < K; k += 32) {
        float v = 0.f;
        if (k < kt && !((float)(k + 1) * f > nyq)) v = (grad_cn(k) - dot) / S;
        gcr[k] = v;
    }
    if (lane == 0) {
        const float2* ab = gaseg + (size_t)b * (T + 1);
        float v = (ab[t + 1].x - ab[t + 1].y) + ab[t].y;
        if (t == 0) v += ab[0].x - ab[0].y;
        if (t == T - 1) v += ab[T].y;
        ga[frame] = v;
    }
}

// pick samples-per-thread (even: samples are processed in FFMA2 pairs).  R = 8 or 4 needs the half-filled first / last
// segment to split on a thread boundary (hop % 2R == 0); R = 2 masks per sample, so any even hop up to 1024 works.
static int pick_R(int hop, int prefer) {
    for (int R = prefer; R >= 4; R >>= 1)
        if (hop % (2 * R) == 0 && hop / R <= (R >= 16 ? 256 : 512) && hop / R >= 32) return R;
    if ((hop + 1) / 2 <= 512) return 2;
    for (int R = 4; R <= prefer; R <<= 1)
        if (hop % (2 * R) == 0 && hop / R <= 512) return R;
    return 0;
}

static int check_osc_shape(const char* fn, int B, int T, int K, int hop, int sample_rate) {
    DDSP_REQUIRE(B >= 1 && B <= 65535, "%s: B=%d out of range [1,65535]", fn, B);
    DDSP_REQUIRE(T >= 1, "%s: T=%d must be >= 1", fn, T);
    DDSP_REQUIRE(K >= 1 && K <= 4096, "%s: K=%d out of range [1,4096]", fn, K);
    DDSP_REQUIRE(hop >= 1, "%s: hop=%d must be >= 1", fn, hop);
    DDSP_REQUIRE(sample_rate >= 1, "%s: sample_rate=%d must be positive", fn, sample_rate);
    DDSP_REQUIRE(pick_R(hop, 8) > 0, "%s: hop=%d too large for one CTA per segment (odd or 2 mod 4: max 1024, else 4096)", fn, hop);
    return 0;
}

}  // namespace ddsp

using namespace ddsp;

// ------------------------------------------------------------------------------------------------
// Hand-scheduled harmonic loops.  build.py compiles this file once more to a cubin, lets sass_sched.py reorder the loop
// bodies of osc_synth_kernel<4> and osc_grad_kernel<8> for operand-cache reuse (same instructions, same registers,
// bit-identical results; sass_sched.py says why ptxas's own order leaves 12 % on the table) and embeds the result here.
// The image is loaded through the driver API on first use and its kernels are launched in place of the compiled-in
// ones; if anything is missing (no tuned image in this build, load refused) the stock kernels run.
// ------------------------------------------------------------------------------------------------
#ifdef DDSP_HAVE_TUNED
#include "ddsp_tuned.cuh"
#include "ddsp_osc_tuned.inc"      // kTunedCubin[], kTunedSynth4[], kTunedGrad8[] (mangled names, "" when not tuned)
namespace {
TunedImage g_osc_image = {kTunedCubin, "DDSP_B200_TUNED_CUBIN"};
TunedKernel g_tuned_synth4 = {kTunedSynth4}, g_tuned_grad8 = {kTunedGrad8}, g_tuned_dphase4 = {kTunedDphase4};
}  // namespace
#endif

namespace ddsp {
int osc_tuned_mask() {
#ifdef DDSP_HAVE_TUNED
    return (tuned_function(g_osc_image, g_tuned_synth4) ? 1 : 0) | (tuned_function(g_osc_image, g_tuned_grad8) ? 2 : 0) |
           (tuned_function(g_osc_image, g_tuned_dphase4) ? 4 : 0);
#else
    return 0;
#endif
}
}  // namespace ddsp

extern "C" size_t ddsp_osc_state_bytes(int B, int T, int K) {
    if (B < 1 || T < 1 || K < 1) return 0;
    return align_up(sizeof(double) * (size_t)B * (T + 1), 256) + align_up(sizeof(float) * (size_t)B * T * K, 256) +
           align_up(sizeof(float) * (size_t)B * T, 256) + align_up(sizeof(int) * (size_t)B * T, 256) +
           align_up(sizeof(SegFix) * (size_t)B * (T + 1), 256);
}

extern "C" size_t ddsp_osc_backward_scratch_bytes(int B, int T, int K) {
    if (B < 1 || T < 1 || K < 1) return 0;
    return align_up(sizeof(float2) * (size_t)B * (T + 1) * K, 256) + align_up(sizeof(float2) * (size_t)B * (T + 1), 256);
}

extern "C" int ddsp_osc_prepare(const float* f0, const float* c, void* state, int B, int T, int K, int hop, int sample_rate,
                                void* stream) {
    if (check_osc_shape("ddsp_osc_prepare", B, T, K, hop, sample_rate)) return 1;
    DDSP_REQUIRE(f0 && c && state, "ddsp_osc_prepare: null pointer argument");
    const OscState S = carve_osc_state(state, B, T, K);
    osc_prepare_kernel<<<dim3(B, 1 + (T + PREP_TILE - 1) / PREP_TILE), PREP_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(f0, c, S.cn, S.csum, S.kact, S.phase0, S.seg,
                                                                                 T, K, hop, sample_rate);
    DDSP_CHECK_LAUNCH("osc_prepare_kernel");
    return 0;
}

extern "C" int ddsp_osc_synthesize(const float* f0, const float* a, const void* state, float* y, float* z_save, int accumulate,
                                   int B, int T, int K, int hop, int sample_rate, void* stream) {
    if (check_osc_shape("ddsp_osc_synthesize", B, T, K, hop, sample_rate)) return 1;
    DDSP_REQUIRE(f0 && a && y && state, "ddsp_osc_synthesize: null pointer argument");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const OscState S = carve_osc_state(const_cast<void*>(state), B, T, K);
    const int R = pick_R(hop, 8);                      // 16 samples per thread measured 2 % slower here (one barrier only)
    const dim3 grid(T + 1, B);
    const size_t smem = sizeof(float2) * ((K + 1) & ~1);
#define LAUNCH_SYNTH(RP) \
    osc_synth_kernel<RP><<<grid, (hop + 2 * RP - 1) / (2 * RP), smem, st>>>(S.cn, a, S.kact, S.seg, y, z_save, accumulate, T, K, hop)
    if (smem > 48 * 1024) {
        cudaError_t e = cudaSuccess;
        switch (R) {
            case 16: e = cudaFuncSetAttribute(osc_synth_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); break;
            case 8: e = cudaFuncSetAttribute(osc_synth_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); break;
            case 4: e = cudaFuncSetAttribute(osc_synth_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); break;
            default: e = cudaFuncSetAttribute(osc_synth_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); break;
        }
        DDSP_REQUIRE(e == cudaSuccess, "osc_synth_kernel: cannot reserve %zu B of shared memory: %s", smem, cudaGetErrorString(e));
    }
#ifdef DDSP_HAVE_TUNED
    if (R == 8) {
        if (CUfunction fn = tuned_function(g_osc_image, g_tuned_synth4, st, true)) {
            const float* cn_ = S.cn; const int* kact_ = S.kact; const SegFix* seg_ = S.seg;
            void* params[] = {&cn_, &a, &kact_, &seg_, &y, &z_save, &accumulate, &T, &K, &hop};
            return tuned_launch(fn, grid, (unsigned)((hop + 7) / 8), smem, st, params, "osc_synth_kernel (tuned)");
        }
    }
#endif
    switch (R) {
        case 16: LAUNCH_SYNTH(8); break;
        case 8: LAUNCH_SYNTH(4); break;
        case 4: LAUNCH_SYNTH(2); break;
        default: LAUNCH_SYNTH(1); break;
    }
#undef LAUNCH_SYNTH
    DDSP_CHECK_LAUNCH("osc_synth_kernel");
    return 0;
}

extern "C" int ddsp_osc_forward(const float* f0, const float* c, const float* a, float* y, float* z_save, void* state,
                                int accumulate, int B, int T, int K, int hop, int sample_rate, void* stream) {
    int rc = ddsp_osc_prepare(f0, c, state, B, T, K, hop, sample_rate, stream);
    if (rc) return rc;
    return ddsp_osc_synthesize(f0, a, state, y, z_save, accumulate, B, T, K, hop, sample_rate, stream);
}

template <int RP>
static int launch_grad(const float* g, const float* a, const float* z, const OscState& S, float2* gseg, float2* gaseg, int B,
                       int T, int K, int hop, cudaStream_t st) {
    const int NT = (hop + 2 * RP - 1) / (2 * RP);
    const size_t smem = sizeof(float) * GRAD_CH * (((2 * NT + 3) & ~3) + 4);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(osc_grad_kernel<RP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        DDSP_REQUIRE(e == cudaSuccess, "osc_grad_kernel: cannot reserve %zu B of shared memory: %s", smem,
                     cudaGetErrorString(e));
    }
    osc_grad_kernel<RP><<<dim3(T + 1, B), NT, smem, st>>>(g, a, z, S.kact, S.seg, gseg, gaseg, T, K, hop);
    DDSP_CHECK_LAUNCH("osc_grad_kernel");
    return 0;
}

static float2* scratch_gaseg(void* scratch, int B, int T, int K) {
    return reinterpret_cast<float2*>(reinterpret_cast<char*>(scratch) + align_up(sizeof(float2) * (size_t)B * (T + 1) * K, 256));
}

extern "C" int ddsp_osc_grad_segments(const float* g, const float* f0, const float* a, const float* z_save, const void* state,
                                      void* scratch, int B, int T, int K, int hop, int sample_rate, void* stream) {
    if (check_osc_shape("ddsp_osc_grad_segments", B, T, K, hop, sample_rate)) return 1;
    DDSP_REQUIRE(g && f0 && a && z_save && state && scratch, "ddsp_osc_grad_segments: null pointer argument");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const OscState S = carve_osc_state(const_cast<void*>(state), B, T, K);
    float2* gseg = reinterpret_cast<float2*>(scratch);
    float2* gaseg = scratch_gaseg(scratch, B, T, K);
#ifdef DDSP_HAVE_TUNED
    if (pick_R(hop, 16) == 16) {
        if (CUfunction fn = tuned_function(g_osc_image, g_tuned_grad8, st, true)) {
            const int NT = (hop + 15) / 16;
            const size_t smem = sizeof(float) * GRAD_CH * (((2 * NT + 3) & ~3) + 4);
            const int* kact_ = S.kact; const SegFix* seg_ = S.seg;
            void* params[] = {&g, &a, &z_save, &kact_, &seg_, &gseg, &gaseg, &T, &K, &hop};
            return tuned_launch(fn, dim3(T + 1, B), (unsigned)NT, smem, st, params, "osc_grad_kernel (tuned)");
        }
    }
#endif
    switch (pick_R(hop, 16)) {
        case 16: return launch_grad<8>(g, a, z_save, S, gseg, gaseg, B, T, K, hop, st);
        case 8: return launch_grad<4>(g, a, z_save, S, gseg, gaseg, B, T, K, hop, st);
        case 4: return launch_grad<2>(g, a, z_save, S, gseg, gaseg, B, T, K, hop, st);
        default: return launch_grad<1>(g, a, z_save, S, gseg, gaseg, B, T, K, hop, st);
    }
}

extern "C" int ddsp_osc_grad_finalize(const float* f0, const void* state, const void* scratch, float* gc, float* ga, int B,
                                      int T, int K, int hop, int sample_rate, void* stream) {
    if (check_osc_shape("ddsp_osc_grad_finalize", B, T, K, hop, sample_rate)) return 1;
    DDSP_REQUIRE(f0 && state && scratch && gc && ga, "ddsp_osc_grad_finalize: null pointer argument");
    const OscState S = carve_osc_state(const_cast<void*>(state), B, T, K);
    const float2* gseg = reinterpret_cast<const float2*>(scratch);
    const float2* gaseg = scratch_gaseg(const_cast<void*>(scratch), B, T, K);
    const long long frames = (long long)B * T;
    const int wpb = 8;
    osc_finalize_kernel<<<(unsigned)((frames + wpb - 1) / wpb), wpb * 32, 0, static_cast<cudaStream_t>(stream)>>>(
        gseg, gaseg, S.cn, S.csum, S.kact, f0, gc, ga, B, T, K, sample_rate);
    DDSP_CHECK_LAUNCH("osc_finalize_kernel");
    return 0;
}

extern "C" int ddsp_osc_backward(const float* g, const float* f0, const float* a, const float* z_save, const void* state,
                                 void* scratch, float* gc, float* ga, int B, int T, int K, int hop, int sample_rate,
                                 void* stream) {
    int rc = ddsp_osc_grad_segments(g, f0, a, z_save, state, scratch, B, T, K, hop, sample_rate, stream);
    if (rc) return rc;
    return ddsp_osc_grad_finalize(f0, state, scratch, gc, ga, B, T, K, hop, sample_rate, stream);
}

// ------------------------------------------------------------------------------------------------
// streaming variant -- OscillatorBank.live (harmonic_oscillator.py:64-75)
// Voice 0 starts every harmonic from its own carried phase last[k] (radians), so harmonic k has phase
// last[k] + 2 pi k Theta and the single-seed recurrence does not apply; blocks are a few thousand samples,
// so each harmonic-sample is evaluated directly (fp64 phase, sinpi).  Launch-latency bound by nature.
// ------------------------------------------------------------------------------------------------
namespace ddsp {

__global__ void __launch_bounds__(256) osc_live_kernel(const float* __restrict__ cn, const float* __restrict__ a,
                                                       const float* __restrict__ f0, const int* __restrict__ kact,
                                                       const double* __restrict__ phase0, const float* __restrict__ last_in,
                                                       float* __restrict__ y, int T, int K, int hop, double inv_sr) {
    const int s = blockIdx.x, b = blockIdx.y;
    const Segment g = make_segment(s, T, hop, f0 + (size_t)b * T, phase0 + (size_t)b * (T + 1), inv_sr);
    const size_t fL = (size_t)b * T + g.tL, fR = (size_t)b * T + g.tR;
    const int kmax = max(kact[fL], kact[fR]);
    const float* cl = cn + fL * K;
    const float* cr = cn + fR * K;
    const float aL = a[fL], da = a[fR] - aL;
    const double inv_2pi = 0.15915494309189533577;
    for (int r = g.r_lo + threadIdx.x; r < g.r_hi; r += blockDim.x) {
        const double rr = (double)(r + 1);
        const double theta = fma(g.du * rr, rr, fma(rr, g.uL, g.E));        // cycles, not wrapped
        const float w = ((float)r + seg_phi(hop)) / (float)hop;
        float acc = 0.f;
        for (int k = 0; k < kmax; ++k) {
            double ph = theta * (double)(k + 1);
            if (b == 0) ph += (double)last_in[k] * inv_2pi;
            ph -= rint(ph);
            const float amp = fmaf(w, cr[k] - cl[k], cl[k]);
            acc = fmaf(amp, sinpif(2.0f * (float)ph), acc);
        }
        y[(size_t)b * T * hop + (g.n0 + r)] = fmaf(w, da, aL) * acc;
    }
}

// new carried phases: wrapped phase of voice 0's last sample (:72)
__global__ void osc_live_state_kernel(const float* __restrict__ f0, const double* __restrict__ phase0,
                                      const float* __restrict__ last_in, float* __restrict__ last_out, int T, int K, int hop,
                                      double inv_sr) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= K) return;
    const Segment g = make_segment(T, T, hop, f0, phase0, inv_sr);
    const double rr = (double)half_up(hop);                                  // r = ceil(hop/2) - 1 is the last sample
    const double theta = fma(g.du * rr, rr, fma(rr, g.uL, g.E));
    double ph = theta * (double)(k + 1) + (double)last_in[k] * 0.15915494309189533577;
    ph -= floor(ph);
    last_out[k] = (float)(ph * 6.283185307179586476925);
}

}  // namespace ddsp

extern "C" size_t ddsp_osc_backward_f0_scratch_bytes(int B, int T, int hop) {
    if (B < 1 || T < 1 || hop < 1) return 0;
    return sizeof(float) * (size_t)B * T * hop;
}

extern "C" int ddsp_osc_backward_f0(const float* g, const float* f0, const float* a, const void* state, void* scratch, float* gf0,
                                    int B, int T, int K, int hop, int sample_rate, void* stream) {
    if (check_osc_shape("ddsp_osc_backward_f0", B, T, K, hop, sample_rate)) return 1;
    DDSP_REQUIRE(g && f0 && a && state && scratch && gf0, "ddsp_osc_backward_f0: null pointer argument");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const OscState S = carve_osc_state(const_cast<void*>(state), B, T, K);
    float* q = static_cast<float*>(scratch);
    const int R = pick_R(hop, 8);
    const dim3 grid(T + 1, B);
    const size_t smem = sizeof(float4) * ((K + 1) & ~1);
#define LAUNCH_DPH(RP)                                                                                                  \
    do {                                                                                                                \
        if (smem > 48 * 1024) {                                                                                         \
            cudaError_t e = cudaFuncSetAttribute(osc_dphase_kernel<RP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
            DDSP_REQUIRE(e == cudaSuccess, "osc_dphase_kernel: cannot reserve %zu B of shared memory: %s", smem, cudaGetErrorString(e)); \
        }                                                                                                               \
        osc_dphase_kernel<RP><<<grid, (hop + 2 * RP - 1) / (2 * RP), smem, st>>>(S.cn, a, S.kact, S.seg, g, q, T, K, hop); \
    } while (0)
    bool launched = false;
#ifdef DDSP_HAVE_TUNED
    if (R == 8) {
        if (CUfunction fn = tuned_function(g_osc_image, g_tuned_dphase4, st, true)) {
            const float* cn_ = S.cn; const int* kact_ = S.kact; const SegFix* seg_ = S.seg;
            void* params[] = {&cn_, &a, &kact_, &seg_, &g, &q, &T, &K, &hop};
            if (tuned_launch(fn, grid, (unsigned)((hop + 7) / 8), smem, st, params, "osc_dphase_kernel (tuned)")) return 1;
            launched = true;
        }
    }
#endif
    if (!launched) switch (R) {
        case 16: LAUNCH_DPH(8); break;
        case 8: LAUNCH_DPH(4); break;
        case 4: LAUNCH_DPH(2); break;
        default: LAUNCH_DPH(1); break;
    }
#undef LAUNCH_DPH
    if (!launched) DDSP_CHECK_LAUNCH("osc_dphase_kernel");
    const long long N = (long long)T * hop;
    osc_revscan_kernel<<<B, 256, 0, st>>>(q, N, 6.283185307179586476925 / (double)sample_rate);
    DDSP_CHECK_LAUNCH("osc_revscan_kernel");
    return ddsp_rescale_backward(q, gf0, B, T, 1, hop, stream);       // adjoint of the half-pixel linear upsampling (:52-55)
}

extern "C" int ddsp_osc_live(const float* f0, const float* c, const float* a, const float* last_in, float* last_out, float* y,
                             void* state, int B, int T, int K, int hop, int sample_rate, void* stream) {
    if (check_osc_shape("ddsp_osc_live", B, T, K, hop, sample_rate)) return 1;
    DDSP_REQUIRE(f0 && c && a && last_in && last_out && y && state, "ddsp_osc_live: null pointer argument");
    DDSP_REQUIRE(last_in != last_out, "ddsp_osc_live: last_in and last_out must not alias");
    int rc = ddsp_osc_prepare(f0, c, state, B, T, K, hop, sample_rate, stream);
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const OscState S = carve_osc_state(state, B, T, K);
    const double inv_sr = 1.0 / (double)sample_rate;
    osc_live_kernel<<<dim3(T + 1, B), 256, 0, st>>>(S.cn, a, f0, S.kact, S.phase0, last_in, y, T, K, hop, inv_sr);
    DDSP_CHECK_LAUNCH("osc_live_kernel");
    osc_live_state_kernel<<<(K + 127) / 128, 128, 0, st>>>(f0, S.phase0, last_in, last_out, T, K, hop, inv_sr);
    DDSP_CHECK_LAUNCH("osc_live_state_kernel");
    return 0;
}

// ------------------------------------------------------------------------------------------------
// Allocation-free forward (no counterpart in the reference).
// Callers that cannot allocate -- code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc --
// get OscillatorBank.forward with the frame-rate state (ddsp_osc_state_bytes: normalised amplitudes, their sums, active
// counts, phase prefixes) in a file-scope device array instead of a caller-owned buffer.  Every GPU that loads the
// library holds the array until exit.  In HOST code `g_osc_noalloc_state` is only a stand-in for the symbol, not a device
// address: the address is asked for once per device, with that device current, through cudaGetSymbolAddress.
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_osc_noalloc_state[DDSP_OSC_NOALLOC_STATE_BYTES];

extern "C" int ddsp_osc_forward_noalloc(const float* f0, const float* c, const float* a, float* y, int accumulate, int B, int T,
                                        int K, int hop, int sample_rate, void* stream) {
    if (check_osc_shape("ddsp_osc_forward_noalloc", B, T, K, hop, sample_rate)) return 1;
    DDSP_REQUIRE(f0 && c && a && y, "ddsp_osc_forward_noalloc: null pointer argument");
    const size_t need = ddsp_osc_state_bytes(B, T, K);
    DDSP_REQUIRE(need <= (size_t)DDSP_OSC_NOALLOC_STATE_BYTES,
                 "ddsp_osc_forward_noalloc: B=%d T=%d K=%d needs %zu bytes of frame-rate state, the built-in scratch holds %zu "
                 "(DDSP_OSC_NOALLOC_STATE_BYTES); use ddsp_osc_forward with a caller-owned state buffer",
                 B, T, K, need, (size_t)DDSP_OSC_NOALLOC_STATE_BYTES);
    static void* dev_ptr[64] = {nullptr};               // per device: the array lives once in every GPU's copy of the module
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    DDSP_REQUIRE(e == cudaSuccess && dev >= 0 && dev < 64, "ddsp_osc_forward_noalloc: cudaGetDevice failed or device index %d >= 64: %s",
                 dev, cudaGetErrorString(e));
    void* state = dev_ptr[dev];
    if (!state) {
        e = cudaGetSymbolAddress(&state, g_osc_noalloc_state);          // `dev` is the current device
        DDSP_REQUIRE(e == cudaSuccess && state, "ddsp_osc_forward_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
        dev_ptr[dev] = state;
    }
    // prepare writes every byte of the state the synthesis kernel reads (nothing has to start at zero), so the previous
    // call's contents are simply overwritten -- which is also why two calls that overlap on the GPU would race
    int rc = ddsp_osc_prepare(f0, c, state, B, T, K, hop, sample_rate, stream);
    if (rc) return rc;
    return ddsp_osc_synthesize(f0, a, state, y, nullptr, accumulate, B, T, K, hop, sample_rate, stream);
}
