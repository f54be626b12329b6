// Experiment harness (not product code): FFMA vs FFMA2 peak probe and oscillator synth/grad inner-loop variants.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -o scratch/exp_osc scratch/exp_osc.cu \
//        -Lddsp_pytorch_b200/lib -lddsp_b200 -Xlinker -rpath -Xlinker '$ORIGIN/../ddsp_pytorch_b200/lib'
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../ddsp_pytorch_b200/csrc/ddsp_common.cuh"

using namespace ddsp;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1);} } while (0)

// ---------------------------------------------------------------- probes
__global__ void __launch_bounds__(1024) probe_ffma(float* out, int iters) {
    float x[8];
    for (int u = 0; u < 8; ++u) x[u] = threadIdx.x * 1e-3f + u;
    const float m = 0.999f, c = 1e-3f * (blockIdx.x + 1);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int u = 0; u < 8; ++u) x[u] = fmaf(x[u], m, c);
    }
    float s = 0.f;
    for (int u = 0; u < 8; ++u) s += x[u];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(1024) probe_ffma2(float* out, int iters) {
    float2 x[8];
    for (int u = 0; u < 8; ++u) x[u] = make_float2(threadIdx.x * 1e-3f + u, threadIdx.x * 2e-3f + u);
    const float2 m = make_float2(0.999f, 0.998f), c = make_float2(1e-3f * (blockIdx.x + 1), 2e-3f);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int u = 0; u < 8; ++u) x[u] = __ffma2_rn(x[u], m, c);
    }
    float s = 0.f;
    for (int u = 0; u < 8; ++u) s += x[u].x + x[u].y;
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// mixed: the synth loop's instruction mix (3 FFMA2 + 1 FADD2 per harmonic per sample pair), no memory
__global__ void __launch_bounds__(1024) probe_mix2(float* out, int iters) {
    float2 sn[4], dn[4], dl[4], P[4], Q[4];
    for (int u = 0; u < 4; ++u) {
        sn[u] = make_float2(threadIdx.x * 1e-3f + u, 0.5f);
        dn[u] = sn[u];
        dl[u] = make_float2(-1e-3f, -2e-3f);
        P[u] = Q[u] = make_float2(0.f, 0.f);
    }
    const float2 a = make_float2(0.25f, 0.25f), b = make_float2(0.5f, 0.5f);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                P[u] = __ffma2_rn(a, sn[u], P[u]);
                Q[u] = __ffma2_rn(b, sn[u], Q[u]);
                dn[u] = __ffma2_rn(dl[u], sn[u], dn[u]);
                sn[u] = __fadd2_rn(sn[u], dn[u]);
            }
    }
    float s = 0.f;
    for (int u = 0; u < 4; ++u) s += P[u].x + P[u].y + Q[u].x + Q[u].y;
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(1024) probe_mix1(float* out, int iters) {
    float sn[8], dn[8], dl[8], P[8], Q[8];
    for (int u = 0; u < 8; ++u) {
        sn[u] = threadIdx.x * 1e-3f + u;
        dn[u] = sn[u];
        dl[u] = -1e-3f;
        P[u] = Q[u] = 0.f;
    }
    const float a = 0.25f, b = 0.5f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                P[u] = fmaf(a, sn[u], P[u]);
                Q[u] = fmaf(b, sn[u], Q[u]);
                dn[u] = fmaf(dl[u], sn[u], dn[u]);
                sn[u] += dn[u];
            }
    }
    float s = 0.f;
    for (int u = 0; u < 8; ++u) s += P[u] + Q[u];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static float time_ms(F f, int reps = 5) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    f();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int i = 0; i < reps; ++i) {
        CK(cudaEventRecord(a));
        f();
        CK(cudaEventRecord(b));
        CK(cudaEventSynchronize(b));
        float ms;
        CK(cudaEventElapsedTime(&ms, a, b));
        best = fminf(best, ms);
    }
    CK(cudaGetLastError());
    return best;
}

// ---------------------------------------------------------------- segment helpers (as in ddsp_osc.cu)
struct Seg { int tL, tR, r_lo, r_hi; long long n0; double E, uL, du; };
__device__ __forceinline__ Seg mkseg(int s, int T, int hop, const float* f0b, const double* Eb, double inv_sr) {
    Seg g;
    g.tL = s > 0 ? s - 1 : 0; g.tR = s < T ? s : T - 1;
    g.r_lo = s == 0 ? hop / 2 : 0; g.r_hi = s == T ? hop / 2 : hop;
    g.n0 = (long long)s * hop - hop / 2;
    g.E = Eb[s]; g.uL = (double)f0b[g.tL] * inv_sr;
    g.du = ((double)f0b[g.tR] * inv_sr - g.uL) / (2.0 * hop);
    return g;
}
__device__ __forceinline__ PhaseSeed seedat(const Seg& g, int r) {
    const double rr = (double)(r + 1);
    return phase_seed(wrap_cycles(fma(g.du * rr, rr, fma(rr, g.uL, g.E))));
}

// fp64 only for the wrapped phase; sin/cos of theta'/2 by fp32 Taylor polynomials on [-pi/4, pi/4]
__device__ __forceinline__ PhaseSeed seed32(const Seg& g, int r) {
    const double rr = (double)(r + 1);
    double t = wrap_cycles(fma(g.du * rr, rr, fma(rr, g.uL, g.E)));
    PhaseSeed p;
    p.sigma = 1.0f;
    if (fabs(t) > 0.25) { t -= copysign(0.5, t); p.sigma = -1.0f; }
    const float x = (float)(3.14159265358979323846 * t);          // theta'/2 in [-pi/4, pi/4]
    const float x2 = x * x;
    float sp = 2.7557319223985890653e-6f;
    sp = fmaf(sp, x2, -1.9841269841269841270e-4f);
    sp = fmaf(sp, x2, 8.3333333333333333333e-3f);
    sp = fmaf(sp, x2, -1.6666666666666666667e-1f);
    const float sv = fmaf(sp * x2, x, x);
    float cp = -2.7557319223985890653e-7f;
    cp = fmaf(cp, x2, 2.4801587301587301587e-5f);
    cp = fmaf(cp, x2, -1.3888888888888888889e-3f);
    cp = fmaf(cp, x2, 4.1666666666666666667e-2f);
    cp = fmaf(cp, x2, -0.5f);
    const float cv = fmaf(cp, x2, 1.0f);
    p.sin1 = 2.0f * sv * cv;
    p.delta = -4.0f * sv * sv;
    return p;
}
__device__ __forceinline__ PhaseSeed seed_fake(const Seg& g, int r) {   // timing only: what if the seed were free
    PhaseSeed p;
    p.sin1 = (float)r * 1e-3f + (float)g.E; p.delta = -1e-3f * (float)g.uL; p.sigma = 1.f;
    return p;
}
template <int SM>
__device__ __forceinline__ PhaseSeed seed_sel(const Seg& g, int r) {
    if (SM == 1) return seed32(g, r);
    if (SM == 2) return seed_fake(g, r);
    return seedat(g, r);
}

// ---------------------------------------------------------------- synth variant: packed FFMA2, R samples (R/2 pairs)
template <int R, int SM>
__global__ void __launch_bounds__(512) synth_v2(const float* __restrict__ cn, const float* __restrict__ a,
                                                const float* __restrict__ f0, const int* __restrict__ kact,
                                                const double* __restrict__ phase0, float* __restrict__ y,
                                                float* __restrict__ z_save, int T, int K, int hop, double inv_sr) {
    extern __shared__ __align__(16) float4 amp4[];     // (cL, cL, dc, dc) per harmonic
    constexpr int RP = R / 2;
    const int s = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
    const Seg g = mkseg(s, T, hop, f0 + (size_t)b * T, phase0 + (size_t)b * (T + 1), inv_sr);
    const size_t fL = (size_t)b * T + g.tL, fR = (size_t)b * T + g.tR;
    const int kmax = max(kact[fL], kact[fR]);
    const int kpad = (kmax + 1) & ~1;
    {
        const float* cl = cn + fL * K;
        const float* cr = cn + fR * K;
        for (int k = tid; k < kpad; k += blockDim.x) {
            float l = 0.f, d = 0.f;
            if (k < kmax) { l = cl[k]; d = cr[k] - l; }
            amp4[k] = make_float4(l, l, d, d);
        }
    }
    __syncthreads();
    const int r0 = tid * R;
    if (r0 >= g.r_hi || r0 + R <= g.r_lo) return;
    float2 sn[RP], dn[RP], dl[RP], Pe[RP], Po[RP], Qe[RP], Qo[RP], sg[RP];
#pragma unroll
    for (int j = 0; j < RP; ++j) {
        const PhaseSeed p0 = seed_sel<SM>(g, r0 + 2 * j), p1 = seed_sel<SM>(g, r0 + 2 * j + 1);
        sn[j] = make_float2(p0.sin1, p1.sin1);
        dn[j] = sn[j];
        dl[j] = make_float2(p0.delta, p1.delta);
        sg[j] = make_float2(p0.sigma, p1.sigma);
        Pe[j] = Po[j] = Qe[j] = Qo[j] = make_float2(0.f, 0.f);
    }
#pragma unroll 2
    for (int k = 0; k < kpad; k += 2) {
        const float4 A = amp4[k], Bv = amp4[k + 1];
        const float2 Ac = make_float2(A.x, A.y), Ad = make_float2(A.z, A.w);
        const float2 Bc = make_float2(Bv.x, Bv.y), Bd = make_float2(Bv.z, Bv.w);
#pragma unroll
        for (int j = 0; j < RP; ++j) {
            Po[j] = __ffma2_rn(Ac, sn[j], Po[j]);
            Qo[j] = __ffma2_rn(Ad, sn[j], Qo[j]);
            dn[j] = __ffma2_rn(dl[j], sn[j], dn[j]);
            sn[j] = __fadd2_rn(sn[j], dn[j]);
        }
#pragma unroll
        for (int j = 0; j < RP; ++j) {
            Pe[j] = __ffma2_rn(Bc, sn[j], Pe[j]);
            Qe[j] = __ffma2_rn(Bd, sn[j], Qe[j]);
            dn[j] = __ffma2_rn(dl[j], sn[j], dn[j]);
            sn[j] = __fadd2_rn(sn[j], dn[j]);
        }
    }
    const float aL = a[fL], da = a[fR] - aL;
    const float inv_hop = 1.0f / (float)hop;
    float out[R], zs[R];
#pragma unroll
    for (int j = 0; j < RP; ++j) {
        const float2 zP = __ffma2_rn(sg[j], Po[j], Pe[j]), zQ = __ffma2_rn(sg[j], Qo[j], Qe[j]);
        const float w0 = ((float)(r0 + 2 * j) + 0.5f) * inv_hop, w1 = ((float)(r0 + 2 * j + 1) + 0.5f) * inv_hop;
        zs[2 * j] = fmaf(w0, zQ.x, zP.x);
        zs[2 * j + 1] = fmaf(w1, zQ.y, zP.y);
        out[2 * j] = fmaf(w0, da, aL) * zs[2 * j];
        out[2 * j + 1] = fmaf(w1, da, aL) * zs[2 * j + 1];
    }
    const size_t N = (size_t)T * hop;
    float* yp = y + (size_t)b * N + (g.n0 + r0);
    float* zp = z_save ? z_save + (size_t)b * N + (g.n0 + r0) : nullptr;
#pragma unroll
    for (int j = 0; j < R; j += 4) {
        *reinterpret_cast<float4*>(yp + j) = make_float4(out[j], out[j + 1], out[j + 2], out[j + 3]);
        if (zp) *reinterpret_cast<float4*>(zp + j) = make_float4(zs[j], zs[j + 1], zs[j + 2], zs[j + 3]);
    }
}

// ---------------------------------------------------------------- host
static double urand(uint64_t& s) { s = s * 6364136223846793005ull + 1442695040888963407ull; return (double)(s >> 11) / 9007199254740992.0; }

int main(int argc, char** argv) {
    const int B = argc > 1 ? atoi(argv[1]) : 1024, T = 375, K = 256, hop = 512, sr = 48000;
    const size_t N = (size_t)T * hop;
    const int only = argc > 2 ? atoi(argv[2]) : 0;      // run only this synth variant (for ncu)
    float* scratch_out;
    CK(cudaMalloc(&scratch_out, sizeof(float) * 148 * 8 * 1024));
    if (!only) {
        const int iters = 4096, blocks = 148 * 8, threads = 256;
        const double fl = 2.0 * 64 * iters * (double)blocks * threads;
        float t1 = time_ms([&] { probe_ffma<<<blocks, threads>>>(scratch_out, iters); });
        float t2 = time_ms([&] { probe_ffma2<<<blocks, threads>>>(scratch_out, iters); });
        printf("probe FFMA : %.3f ms  %.1f TFLOP/s\n", t1, fl / t1 * 1e-9);
        printf("probe FFMA2: %.3f ms  %.1f TFLOP/s (2 FMA per lane-instr)\n", t2, 2 * fl / t2 * 1e-9);
        // mix: per iteration 4 reps x 8 samples x 4 instr
        const double ops = 4.0 * 8 * 4 * iters * (double)blocks * threads;
        float t3 = time_ms([&] { probe_mix1<<<blocks, threads>>>(scratch_out, iters); });
        float t4 = time_ms([&] { probe_mix2<<<blocks, threads>>>(scratch_out, iters); });
        printf("mix scalar : %.3f ms  %.2f Tlane-op/s\n", t3, ops / t3 * 1e-9);
        printf("mix packed : %.3f ms  %.2f Tlane-op/s\n", t4, ops / t4 * 1e-9);
    }
    // synthetic controls like bench.py: log-uniform f0 in [80,1000] with vibrato, positive c, a
    std::vector<float> f0((size_t)B * T), c((size_t)B * T * K), a((size_t)B * T);
    uint64_t s = 1234;
    for (int b = 0; b < B; ++b) {
        const double base = 80.0 * pow(1000.0 / 80.0, urand(s));
        for (int t = 0; t < T; ++t) {
            f0[(size_t)b * T + t] = (float)(base * (1.0 + 0.01 * sin(2 * M_PI * 5.0 * t * hop / sr)) + (urand(s) - 0.5));
            a[(size_t)b * T + t] = (float)urand(s);
            for (int k = 0; k < K; ++k) c[((size_t)b * T + t) * K + k] = (float)(urand(s) + 1e-3);
        }
    }
    float *df0, *dc, *da, *y0, *y1, *z0, *z1;
    void* state;
    CK(cudaMalloc(&df0, f0.size() * 4)); CK(cudaMalloc(&dc, c.size() * 4)); CK(cudaMalloc(&da, a.size() * 4));
    CK(cudaMalloc(&y0, B * N * 4)); CK(cudaMalloc(&y1, B * N * 4)); CK(cudaMalloc(&z0, B * N * 4)); CK(cudaMalloc(&z1, B * N * 4));
    CK(cudaMalloc(&state, ddsp_osc_state_bytes(B, T, K)));
    CK(cudaMemcpy(df0, f0.data(), f0.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dc, c.data(), c.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(da, a.data(), a.size() * 4, cudaMemcpyHostToDevice));
    if (ddsp_osc_prepare(df0, dc, state, B, T, K, hop, sr, nullptr)) { printf("prepare: %s\n", ddsp_last_error()); return 1; }
    const OscState S = carve_osc_state(state, B, T, K);
    float t0 = time_ms([&] { ddsp_osc_synthesize(df0, da, state, y0, z0, 0, B, T, K, hop, sr, nullptr); });
    printf("synth V0 (library, R=4 scalar): %.3f ms\n", t0);
    const double inv_sr = 1.0 / sr;
    std::vector<float> h0(B * N), h1(B * N);
    CK(cudaMemcpy(h0.data(), y0, B * N * 4, cudaMemcpyDeviceToHost));
    auto check = [&](const char* name, float ms) {
        CK(cudaMemcpy(h1.data(), y1, B * N * 4, cudaMemcpyDeviceToHost));
        double md = 0;
        for (size_t i = 0; i < B * N; ++i) md = fmax(md, fabs((double)h0[i] - h1[i]));
        printf("%s: %.3f ms  (x%.2f vs V0)  max|dy| = %.2e\n", name, ms, t0 / ms, md);
    };
    if (!only || only == 4) {
        CK(cudaMemset(y1, 0, B * N * 4));
        float t = time_ms([&] { synth_v2<4, 0><<<dim3(T + 1, B), hop / 4, sizeof(float4) * (K + 2)>>>(S.cn, da, df0, S.kact, S.phase0, y1, z1, T, K, hop, inv_sr); });
        check("synth V2<4> packed", t);
    }
    if (!only || only == 8) {
        CK(cudaMemset(y1, 0, B * N * 4));
        float t = time_ms([&] { synth_v2<8, 0><<<dim3(T + 1, B), hop / 8, sizeof(float4) * (K + 2)>>>(S.cn, da, df0, S.kact, S.phase0, y1, z1, T, K, hop, inv_sr); });
        check("synth V2<8> packed", t);
    }
    if (!only || only == 81) {
        CK(cudaMemset(y1, 0, B * N * 4));
        float t = time_ms([&] { synth_v2<8, 1><<<dim3(T + 1, B), hop / 8, sizeof(float4) * (K + 2)>>>(S.cn, da, df0, S.kact, S.phase0, y1, z1, T, K, hop, inv_sr); });
        check("synth V2<8> packed, fp32 seed polys", t);
    }
    if (!only || only == 82) {
        float t = time_ms([&] { synth_v2<8, 2><<<dim3(T + 1, B), hop / 8, sizeof(float4) * (K + 2)>>>(S.cn, da, df0, S.kact, S.phase0, y1, z1, T, K, hop, inv_sr); });
        check("synth V2<8> packed, FAKE seeds (timing only)", t);
    }
    if (!only || only == 41) {
        CK(cudaMemset(y1, 0, B * N * 4));
        float t = time_ms([&] { synth_v2<4, 1><<<dim3(T + 1, B), hop / 4, sizeof(float4) * (K + 2)>>>(S.cn, da, df0, S.kact, S.phase0, y1, z1, T, K, hop, inv_sr); });
        check("synth V2<4> packed, fp32 seed polys", t);
    }
    return 0;
}
