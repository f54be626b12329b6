// Shared helpers for the sm_100a DDSP kernels: error plumbing, phase math, Philox.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/ddsp_b200.h"

namespace ddsp {

void set_error(const char* fmt, ...);
void count_launch();    // every kernel this library launches (or records into a capturing stream) passes through here

#define DDSP_REQUIRE(cond, ...)                 \
    do {                                        \
        if (!(cond)) {                          \
            ::ddsp::set_error(__VA_ARGS__);     \
            return 1;                           \
        }                                       \
    } while (0)

#define DDSP_CHECK_LAUNCH(name)                                                          \
    do {                                                                                 \
        cudaError_t e__ = cudaGetLastError(); /* clears it: a failed launch must not poison later calls */ \
        ::ddsp::count_launch();                                                          \
        if (e__ != cudaSuccess) {                                                        \
            ::ddsp::set_error("%s: launch failed: %s", name, cudaGetErrorString(e__));   \
            return 2;                                                                    \
        }                                                                                \
    } while (0)

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- frame-rate state of the oscillator (device layout inside the caller's `state` buffer) ----
// Phase polynomial of one segment in 0.64 fixed-point CYCLES (integer arithmetic wraps mod 1 cycle for free and is
// exact):  Theta(r) = E + (r+1)*uL + (r+1)^2*du   (mod 2^64),  r = sample index inside the segment.
struct SegFix {
    unsigned long long E;    // phase accumulated before the segment
    unsigned long long uL;   // left frame's increment f0/sample_rate
    long long du;            // (uR - uL) / (2*hop)
};

struct OscState {
    float* cn;        // [B,T,K] masked + normalised harmonic amplitudes      (harmonic_oscillator.py:32-33)
    float* csum;      // [B,T]   sum of the masked amplitudes (backward of the normalisation)
    int* kact;        // [B,T]   1 + index of the last unmasked harmonic (K when the frame sum is not > 0)
    double* phase0;   // [B,T+1] fundamental phase in CYCLES accumulated before each segment, wrapped to [0,1) (live path)
    SegFix* seg;      // [B,T+1] the same prefix with the segment's increment pair, fixed point (forward/backward path)
};

inline OscState carve_osc_state(void* base, int B, int T, int K) {
    char* p = reinterpret_cast<char*>(base);
    OscState s;
    s.phase0 = reinterpret_cast<double*>(p);
    p += align_up(sizeof(double) * (size_t)B * (T + 1), 256);
    s.cn = reinterpret_cast<float*>(p);
    p += align_up(sizeof(float) * (size_t)B * T * K, 256);
    s.csum = reinterpret_cast<float*>(p);
    p += align_up(sizeof(float) * (size_t)B * T, 256);
    s.kact = reinterpret_cast<int*>(p);
    p += align_up(sizeof(int) * (size_t)B * T, 256);
    s.seg = reinterpret_cast<SegFix*>(p);
    return s;
}

// ---- per-sample phase of the fundamental ---------------------------------------------------------
// Segment s (0..T) covers samples [s*hop - hop/2, s*hop + hop/2) and interpolates frames
// L = max(s-1,0) and R = min(s,T-1) with weight w = (r+0.5)/hop, r = sample - segment start
// (F.interpolate linear, align_corners=False; harmonic_oscillator.py:52-55).  The reference's
// cumsum of the upsampled rad/sample (:41) has a closed form inside a segment:
//     Theta(r) = E_s + (r+1)*uL + (uR-uL)*(r+1)^2/(2*hop)            [cycles, u = f0/sample_rate]
// evaluated in fp64 and wrapped to [-0.5,0.5); harmonic k then has phase k*Theta (mod 1).
struct PhaseSeed {
    float sin1;    // sin(theta')            theta' = 2*pi*Theta folded into [-pi/2, pi/2]
    float delta;   // -4 sin^2(theta'/2)     Reinsch form of 2cos(theta') - 2
    float sigma;   // +1, or -1 when theta' = theta -+ pi (then sin(k theta) = sigma^k sin(k theta'))
};

__device__ __forceinline__ double wrap_cycles(double x) { return x - rint(x); }

__device__ __forceinline__ PhaseSeed phase_seed(double theta_cycles) {
    // theta_cycles in [-0.5, 0.5]
    PhaseSeed p;
    double t = theta_cycles;
    p.sigma = 1.0f;
    if (fabs(t) > 0.25) {
        t -= copysign(0.5, t);
        p.sigma = -1.0f;
    }
    // x = pi*t in [-pi/4, pi/4]; short fp64 Taylor polynomials (error < 1e-11) -- cheaper than sincospi(double)
    const double x = 3.14159265358979323846 * t;
    const double x2 = x * x;
    double s = -2.5052108385441718775e-8;            // -1/11!
    s = fma(s, x2, 2.7557319223985890653e-6);        //  1/9!
    s = fma(s, x2, -1.9841269841269841270e-4);       // -1/7!
    s = fma(s, x2, 8.3333333333333333333e-3);        //  1/5!
    s = fma(s, x2, -1.6666666666666666667e-1);       // -1/3!
    s = fma(s * x2, x, x);
    double c = 2.0876756987868098979e-9;             //  1/12!
    c = fma(c, x2, -2.7557319223985890653e-7);       // -1/10!
    c = fma(c, x2, 2.4801587301587301587e-5);        //  1/8!
    c = fma(c, x2, -1.3888888888888888889e-3);       // -1/6!
    c = fma(c, x2, 4.1666666666666666667e-2);        //  1/4!
    c = fma(c, x2, -0.5);
    c = fma(c, x2, 1.0);
    p.sin1 = (float)(2.0 * s * c);
    p.delta = (float)(-4.0 * s * s);
    return p;
}

// Seed from the fixed-point phase: the top 32 bits are the phase in [-0.5, 0.5) cycles as a signed integer; fold to
// [-0.25, 0.25) by flipping the top bit when bits 31 and 30 differ (theta' = theta -+ pi, sigma = -1); then
// sin/cos of theta'/2 = pi*t in [-pi/4, pi/4] by fp32 Taylor polynomials (truncation < 3e-9 relative).  No fp64, no MUFU.
__device__ __forceinline__ PhaseSeed phase_seed_fix(unsigned long long ph) {
    int h = (int)(ph >> 32);
    const int flip = (h ^ (h << 1)) & 0x80000000;
    h ^= flip;
    PhaseSeed p;
    p.sigma = __int_as_float(0x3f800000 | flip);
    const float hf = (float)h;
    const float x = fmaf(hf, 0x1.921fb6p-31f, hf * -0x1.777a5cp-56f);   // pi * 2^-32 split hi + lo: one rounding
    const float x2 = x * x;
    float sp = 2.7557319223985890653e-6f;                //  1/9!
    sp = fmaf(sp, x2, -1.9841269841269841270e-4f);       // -1/7!
    sp = fmaf(sp, x2, 8.3333333333333333333e-3f);        //  1/5!
    sp = fmaf(sp, x2, -1.6666666666666666667e-1f);       // -1/3!
    const float sv = fmaf(sp * x2, x, x);
    float cp = -2.7557319223985890653e-7f;               // -1/10!
    cp = fmaf(cp, x2, 2.4801587301587301587e-5f);        //  1/8!
    cp = fmaf(cp, x2, -1.3888888888888888889e-3f);       // -1/6!
    cp = fmaf(cp, x2, 4.1666666666666666667e-2f);        //  1/4!
    cp = fmaf(cp, x2, -0.5f);
    const float cv = fmaf(cp, x2, 1.0f);
    p.sin1 = 2.0f * sv * cv;
    p.delta = -4.0f * sv * sv;
    return p;
}

// ---- Philox4x32-10 (Salmon et al. 2011); counter layout documented in include/ddsp_b200.h --------
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
        const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += W0;
        key.y += W1;
    }
    return ctr;
}

__device__ __forceinline__ float u32_to_pm1(uint32_t w) {
    return (float)(w >> 8) * (2.0f * 5.9604644775390625e-8f) - 1.0f;   // 2*(w>>8)*2^-24 - 1, exact in fp32
}

__device__ __forceinline__ float4 philox_noise4(uint32_t quad, uint32_t t, uint32_t voice, uint32_t offset, uint2 key) {
    const uint4 r = philox4x32_10(make_uint4(quad, t, voice, offset), key);
    return make_float4(u32_to_pm1(r.x), u32_to_pm1(r.y), u32_to_pm1(r.z), u32_to_pm1(r.w));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace ddsp
