// FFMA2 issue-rate probe: which operand patterns of fma.rn.f32x2 reach the FMA pipe's rate on sm_100a?
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o profiles/probes/ffma2_probe profiles/probes/ffma2_probe.cu
// Every variant runs ITERS iterations of an unrolled body of known packed-instruction count; the result is packed
// instructions per cycle per SM sub-partition (0.5 = the pipe's limit: 64 lane-FMAs per instruction, 32 lanes per cycle).
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

constexpr int RP = 4;

__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }

// A: scalar FFMA, constants as operands (the bench's peak probe)
__global__ void kA(float* out, int iters) {
    float x[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) x[j] = threadIdx.x * 1e-3f + j;
    const float m = 0.999f, c = 1e-3f * (blockIdx.x + 1);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int j = 0; j < 16; ++j) x[j] = fmaf(x[j], m, c);
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) s += x[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// B: FFMA2, loop-invariant multiplier and addend pairs
__global__ void kB(float* out, int iters) {
    float2 x[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] = f2(threadIdx.x * 1e-3f + j, j);
    const float2 m = f2(0.999f, 0.998f), c = f2(1e-3f * (blockIdx.x + 1), 2e-3f);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int j = 0; j < 8; ++j) x[j] = __ffma2_rn(x[j], m, c);
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += x[j].x + x[j].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// C: FFMA2, three distinct varying register pairs per instruction
__global__ void kC(float* out, int iters, float r) {
    float2 a[8], x[8], y[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { a[j] = f2(j, j); x[j] = f2(threadIdx.x * 1e-3f + j, 0.5f * r); y[j] = f2(0.999f * r + threadIdx.x, 1e-3f * j * r); }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int j = 0; j < 8; ++j) a[j] = __ffma2_rn(x[j], y[j], a[j]);
#pragma unroll
            for (int j = 0; j < 8; ++j) x[j] = __ffma2_rn(a[j], y[(j + 1) & 7], x[j]);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += a[j].x + a[j].y + x[j].x + x[j].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// D / E: the synthesis loop (RP sample pairs): amplitude pair from shared memory as a register PAIR (D) or as a
// scalar broadcast operand (E)
template <bool BCAST>
__global__ void kDE(float* out, int iters, int kpad, float r) {
    extern __shared__ __align__(16) float4 amp[];
    for (int k = threadIdx.x; k < kpad; k += blockDim.x) amp[k] = make_float4(1e-3f * k, 1e-3f * k, 2e-3f, 2e-3f);
    __syncthreads();
    float2 sn[RP], dn[RP], dl[RP], Po[RP], Qo[RP], Pe[RP], Qe[RP];
#pragma unroll
    for (int j = 0; j < RP; ++j) {
        sn[j] = f2(1e-3f * threadIdx.x, 2e-3f * j * r); dn[j] = sn[j]; dl[j] = f2(-1e-4f * r * (threadIdx.x + j), -2e-4f * r);
        Po[j] = Qo[j] = Pe[j] = Qe[j] = f2(0.f, 0.f);
    }
    const float2* amp2 = reinterpret_cast<const float2*>(amp);
    for (int i = 0; i < iters; ++i) {
#pragma unroll 2
        for (int k = 0; k < kpad; k += 2) {
            float2 Ac, Ad, Ec, Ed;
            if (BCAST) {
                const float4 v = *reinterpret_cast<const float4*>(amp2 + k);     // (l, d) of two harmonics: one LDS.128
                Ac = f2(v.x, v.x); Ad = f2(v.y, v.y); Ec = f2(v.z, v.z); Ed = f2(v.w, v.w);
            } else {
                const float4 A = amp[k], E = amp[k + 1];
                Ac = f2(A.x, A.y); Ad = f2(A.z, A.w); Ec = f2(E.x, E.y); Ed = f2(E.z, E.w);
            }
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
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < RP; ++j) s += Po[j].x + Po[j].y + Qo[j].x + Qo[j].y + Pe[j].x + Pe[j].y + Qe[j].x + Qe[j].y + sn[j].x;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// F: the gradient loop (8 sample pairs, accumulators chained over the pairs)
__global__ void __maxnreg__(144) kF(float* out, int iters, int kn, float r) {
    constexpr int R8 = 8;
    float2 sn[R8], dn[R8], dl[R8], ug[R8], uw[R8], ugo[R8], uwo[R8];
#pragma unroll
    for (int j = 0; j < R8; ++j) {
        const float q = r * (threadIdx.x + 1) * (j + 1);
        sn[j] = f2(1e-3f * q, 2e-3f * q); dn[j] = sn[j]; dl[j] = f2(-1e-4f * q, -2e-4f * q);
        ug[j] = f2(0.1f * q, 0.2f * q); uw[j] = f2(0.3f * q, 0.01f * q); ugo[j] = f2(-0.1f * q, 0.25f * q); uwo[j] = f2(0.35f * q, -0.01f * q);
    }
    float acc = 0.f;
    for (int i = 0; i < iters; ++i) {
        for (int kk = 0; kk < kn; kk += 2) {
            float2 p0o = f2(0.f, 0.f), p1o = p0o, p0e = p0o, p1e = p0o;
#pragma unroll
            for (int j = 0; j < R8; ++j) {
                p0o = __ffma2_rn(ugo[j], sn[j], p0o);
                p1o = __ffma2_rn(uwo[j], sn[j], p1o);
                dn[j] = __ffma2_rn(dl[j], sn[j], dn[j]);
                sn[j] = __fadd2_rn(sn[j], dn[j]);
            }
#pragma unroll
            for (int j = 0; j < R8; ++j) {
                p0e = __ffma2_rn(ug[j], sn[j], p0e);
                p1e = __ffma2_rn(uw[j], sn[j], p1e);
                dn[j] = __ffma2_rn(dl[j], sn[j], dn[j]);
                sn[j] = __fadd2_rn(sn[j], dn[j]);
            }
            acc += (p0o.x + p0o.y) + (p1o.x + p1o.y) + (p0e.x + p0e.y) + (p1e.x + p1e.y);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc + sn[0].x;
}

template <class L>
static void run(const char* name, double packed_per_thread_iter, int iters, int threads, int ctas_per_sm, L launch) {
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    int khz = 0;
    CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0));
    const int blocks = sms * ctas_per_sm;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    launch(blocks, threads, iters);                    // warm-up
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        CK(cudaEventRecord(e0));
        launch(blocks, threads, iters);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    const double warps = (double)blocks * threads / 32.0;
    const double instr = warps * packed_per_thread_iter * iters;
    const double cycles = best * 1e-3 * khz * 1e3;
    printf("%-34s thr %4d ctas/SM %2d warps/SM %2d : %.3f ms  %.4f packed instr / clk / SMSP  (%.1f TFLOP/s fma-equiv)\n", name, threads,
           ctas_per_sm, ctas_per_sm * threads / 32, best, instr / cycles / (sms * 4.0), instr * 128.0 / (best * 1e-3) / 1e12);
}

int main() {
    float* out = nullptr;
    CK(cudaMalloc(&out, sizeof(float) * 148 * 64 * 1024));
    const int it = 2000;
    for (int w : {3, 7, 12}) {      // warps per SMSP-ish via CTAs per SM of 128 threads
        run("A scalar FFMA (x2 = packed equiv)", 64 / 2.0, it, 128, w, [&](int b, int t, int i) { kA<<<b, t>>>(out, i); });
        run("B FFMA2 invariant operands", 64, it, 128, w, [&](int b, int t, int i) { kB<<<b, t>>>(out, i); });
        run("C FFMA2 three varying pairs", 64, it, 128, w, [&](int b, int t, int i) { kC<<<b, t>>>(out, i, 1.0f); });
    }
    const int kpad = 128;
    for (int c : {6, 10, 14}) {
        run("D synth loop, pair operands", 8.0 * RP * kpad / 2, 40, 64, c, [&](int b, int t, int i) { kDE<false><<<b, t, sizeof(float4) * kpad>>>(out, i, kpad, 1.0f); });
        run("E synth loop, scalar broadcast", 8.0 * RP * kpad / 2, 40, 64, c, [&](int b, int t, int i) { kDE<true><<<b, t, sizeof(float4) * kpad>>>(out, i, kpad, 1.0f); });
    }
    for (int c : {8, 12}) run("F grad loop", 8.0 * 8 * 128 / 2, 40, 32, c, [&](int b, int t, int i) { kF<<<b, t>>>(out, i, 128, 1.0f); });
    CK(cudaDeviceSynchronize());
    return 0;
}
