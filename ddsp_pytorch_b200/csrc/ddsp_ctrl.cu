// Controller tail for sm_100a -- replaces, as ONE forward launch and three backward launches,
//   harm_amps = modified_sigmoid(dense_harmonic(latent)); total_harm_amp = modified_sigmoid(dense_loudness(latent));
//   noise_distribution = modified_sigmoid(dense_filter(latent))           (model/autoencoder/decoder.py:74-76,96-100,110-116)
// The three Linear heads are one [R,U] x [U,J] product, J = K + 1 + NB, against the three weight matrices where they
// lie (no concatenated copy); bias, 2*sigmoid(z)^2.3026 + 1e-7 and the scatter into c / a / H -- the fp32, contiguous
// layouts the oscillator and noise kernels read -- happen in the GEMM epilogue.  Backward: dZ = g * msig'(z) with the
// bias-gradient partials (one pass), dX = dZ W (same GEMM kernel, other operand layout), dW = dZ^T X (split over the
// R dimension, partials summed in fixed order: deterministic).  Two implementations of the products: FP32 FFMA tiles (first
// half of this file; small R) and tcgen05 with bf16 x 3 operand splits (second half; R >= a few hundred) -- a plain tf32/bf16
// product cannot hold the 1e-3 waveform bound against the fp32 reference, the split product is fp32-accurate.
#include <algorithm>
#include <climits>
#include <cstdint>

#include "ddsp_umma.cuh"

namespace ddsp {

// A row-addressed matrix that may live in up to three separate allocations (rows [0,r0) / [r0,r1) / [r1,..)).
struct Rows3 {
    const float* p[3];
    int r0, r1;
    int ld;
    __device__ __forceinline__ const float* row(int j) const {
        return j < r0 ? p[0] + (size_t)j * ld : (j < r1 ? p[1] + (size_t)(j - r0) * ld : p[2] + (size_t)(j - r1) * ld);
    }
};
static Rows3 rows_plain(const float* p, int ld) { return Rows3{{p, p, p}, INT_MAX, INT_MAX, ld}; }

constexpr int G_BM = 128, G_BN = 64, G_BK = 16, G_THREADS = 256;
constexpr int G_LDA = G_BM + 4, G_LDB = G_BN + 4;
constexpr float MSIG_P = 2.3026f;                                      // decoder.py:113

__device__ __forceinline__ float softplus_neg(float z) {               // log(1 + exp(-z)), stable on both sides
    return fmaxf(-z, 0.f) + log1pf(expf(-fabsf(z)));
}
__device__ __forceinline__ float msig(float z) { return 2.0f * expf(-MSIG_P * softplus_neg(z)) + 1e-7f; }
__device__ __forceinline__ float msig_grad(float z) {                  // 2 p sigma^p (1 - sigma)
    const float one_minus = z >= 0.f ? expf(-z) / (1.f + expf(-z)) : 1.f / (1.f + expf(z));
    return 2.0f * MSIG_P * expf(-MSIG_P * softplus_neg(z)) * one_minus;
}

// The same two functions on the special-function unit (ex2 / lg2 / rcp approximations, a few 1e-6 relative: far inside the
// 1e-4 the outputs are held to) for the tensor-core path, whose epilogue has 256 threads per SM for 16 384 activations.
__device__ __forceinline__ float msig_fast(float z) {
    const float sp = fmaxf(-z, 0.f) + __logf(1.f + __expf(-fabsf(z)));
    return 2.0f * __expf(-MSIG_P * sp) + 1e-7f;
}
__device__ __forceinline__ float msig_grad_fast(float z) {
    const float e = __expf(-fabsf(z)), d = 1.f + e;
    const float one_minus = z >= 0.f ? __fdividef(e, d) : __fdividef(1.f, d);
    return 2.0f * MSIG_P * __expf(-MSIG_P * (fmaxf(-z, 0.f) + __logf(d))) * one_minus;
}

__device__ __forceinline__ void split3(float v, __nv_bfloat16 (&h)[3]) {
#pragma unroll
    for (int s = 0; s < 3; ++s) {
        h[s] = __float2bfloat16_rn(v);
        v -= __bfloat162float(h[s]);                        // exact: the remainder fits fp32
    }
}

enum { EPI_HEADS = 0, EPI_STORE = 1 };

struct HeadsOut {
    const float* bias[3];
    float *c, *a, *H, *z;
    int K, NB, ldz;
};

// C[M,N] = A[M,Kd] * B[Kd,N], fp32.  A_KMAJOR: A(m,k) = A.row(m)[k] (k contiguous) else A(m,k) = A.row(k)[m];
// B_KMAJOR: B(k,n) = B.row(n)[k] (k contiguous) else B(k,n) = B.row(k)[n].  blockIdx.z splits Kd (partials at C + z*M*ldc).
// Every row start and ld must be 16-byte aligned; ragged M, N, Kd edges are zero-filled.
template <bool A_KMAJOR, bool B_KMAJOR, int EPI>
__global__ void __launch_bounds__(G_THREADS, 2)
ctrl_gemm_kernel(Rows3 A, Rows3 B, int M, int N, int Kd, int ksplit, float* __restrict__ C, int ldc, HeadsOut ho) {
    __shared__ __align__(16) float As[2][G_BK][G_LDA];
    __shared__ __align__(16) float Bs[2][G_BK][G_LDB];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int m0 = blockIdx.x * G_BM, n0 = blockIdx.y * G_BN;
    const int kbeg = blockIdx.z * ksplit, kend = min(Kd, kbeg + ksplit);
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);

    float4 pa[2], pb;
    auto fetch = [&](int k0) {
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            const int f = tid + G_THREADS * s;
            if (A_KMAJOR) {
                const int r = m0 + (f >> 2), k = k0 + (f & 3) * 4;
                pa[s] = zero4;
                if (r < M && k < kend) {
                    const float* src = A.row(r) + k;
                    if (k + 4 <= kend) pa[s] = *reinterpret_cast<const float4*>(src);
                    else { pa[s].x = src[0]; if (k + 1 < kend) pa[s].y = src[1]; if (k + 2 < kend) pa[s].z = src[2]; }
                }
            } else {
                const int k = k0 + (f >> 5), m = m0 + (f & 31) * 4;
                pa[s] = zero4;
                if (k < kend && m < M) {
                    const float* src = A.row(k) + m;
                    if (m + 4 <= M) pa[s] = *reinterpret_cast<const float4*>(src);
                    else { pa[s].x = src[0]; if (m + 1 < M) pa[s].y = src[1]; if (m + 2 < M) pa[s].z = src[2]; }
                }
            }
        }
        if (B_KMAJOR) {
            const int n = n0 + (tid >> 2), k = k0 + (tid & 3) * 4;
            pb = zero4;
            if (n < N && k < kend) {
                const float* src = B.row(n) + k;
                if (k + 4 <= kend) pb = *reinterpret_cast<const float4*>(src);
                else { pb.x = src[0]; if (k + 1 < kend) pb.y = src[1]; if (k + 2 < kend) pb.z = src[2]; }
            }
        } else {
            const int k = k0 + (tid >> 4), n = n0 + (tid & 15) * 4;
            pb = zero4;
            if (k < kend && n < N) {
                const float* src = B.row(k) + n;
                if (n + 4 <= N) pb = *reinterpret_cast<const float4*>(src);
                else { pb.x = src[0]; if (n + 1 < N) pb.y = src[1]; if (n + 2 < N) pb.z = src[2]; }
            }
        }
    };
    auto stash = [&](int buf) {
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            const int f = tid + G_THREADS * s;
            if (A_KMAJOR) {
                const int r = f >> 2, kq = (f & 3) * 4;
                As[buf][kq + 0][r] = pa[s].x; As[buf][kq + 1][r] = pa[s].y;
                As[buf][kq + 2][r] = pa[s].z; As[buf][kq + 3][r] = pa[s].w;
            } else {
                *reinterpret_cast<float4*>(&As[buf][f >> 5][(f & 31) * 4]) = pa[s];
            }
        }
        if (B_KMAJOR) {
            const int n = tid >> 2, kq = (tid & 3) * 4;
            Bs[buf][kq + 0][n] = pb.x; Bs[buf][kq + 1][n] = pb.y; Bs[buf][kq + 2][n] = pb.z; Bs[buf][kq + 3][n] = pb.w;
        } else {
            *reinterpret_cast<float4*>(&Bs[buf][tid >> 4][(tid & 15) * 4]) = pb;
        }
    };

    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    const int ntiles = kend > kbeg ? (kend - kbeg + G_BK - 1) / G_BK : 0;
    if (ntiles > 0) { fetch(kbeg); stash(0); }
    __syncthreads();
    for (int t = 0; t < ntiles; ++t) {
        const int buf = t & 1;
        if (t + 1 < ntiles) fetch(kbeg + (t + 1) * G_BK);
#pragma unroll
        for (int k = 0; k < G_BK; ++k) {
            const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
            const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][k][64 + ty * 4]);
            const float4 b = *reinterpret_cast<const float4*>(&Bs[buf][k][tx * 4]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        if (t + 1 < ntiles) stash(buf ^ 1);
        __syncthreads();
    }

#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (m >= M) continue;
        const int n = n0 + tx * 4;
        if (EPI == EPI_STORE) {
            float* dst = C + (size_t)blockIdx.z * M * ldc + (size_t)m * ldc + n;
            if (n + 4 <= N) *reinterpret_cast<float4*>(dst) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
            else
#pragma unroll
                for (int j = 0; j < 4; ++j) if (n + j < N) dst[j] = acc[i][j];
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int col = n + j;
                if (col >= N) continue;
                float z;
                if (col < ho.K) {
                    z = acc[i][j] + ho.bias[0][col];
                    ho.c[(size_t)m * ho.K + col] = msig(z);
                } else if (col == ho.K) {
                    z = acc[i][j] + ho.bias[1][0];
                    ho.a[m] = msig(z);
                } else {
                    z = acc[i][j] + ho.bias[2][col - ho.K - 1];
                    ho.H[(size_t)m * ho.NB + (col - ho.K - 1)] = msig(z);
                }
                if (ho.z) ho.z[(size_t)m * ho.ldz + col] = z;
            }
        }
    }
}

// Forward for up to 255 rows (the real-time decoder: one block of 1-16 frames, rt/synth.py:21-55): the tiled product above
// puts 6-12 CTAs through 32 barrier-separated K steps (49 us for any R <= 128); here one warp owns one output column for a
// group of 16 rows -- lanes stride over U in float4, the group's x lives in shared memory, one shuffle tree per row: 4 us at
// R = 1, 6 at R = 4, 10 at R = 16-64, 14 at R = 128.
constexpr int SMALL_ROWS = 16, SMALL_WARPS = 8, SMALL_MAX_ROWS = 255;     // rows per CTA, columns per CTA, largest R served (the tensor-core path starts at 256)
__global__ void __launch_bounds__(32 * SMALL_WARPS)
ctrl_heads_small_kernel(const float* __restrict__ x, Rows3 W, int R, int U, int J, HeadsOut ho) {
    extern __shared__ __align__(16) float xs[];                     // [rows of this group][U]
    const int rbase = blockIdx.y * SMALL_ROWS;                      // blockIdx.y: group of SMALL_ROWS rows
    x += (size_t)rbase * U;
    R = min(SMALL_ROWS, R - rbase);
    for (int i = threadIdx.x * 4; i < R * U; i += 32 * SMALL_WARPS * 4)
        *reinterpret_cast<float4*>(xs + i) = *reinterpret_cast<const float4*>(x + i);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int col = blockIdx.x * SMALL_WARPS + (threadIdx.x >> 5);
    if (col >= J) return;
    const float* wr = W.row(col);
    float acc[SMALL_ROWS];
#pragma unroll
    for (int r = 0; r < SMALL_ROWS; ++r) acc[r] = 0.f;
    for (int u = lane * 4; u < U; u += 128) {
        const float4 w = *reinterpret_cast<const float4*>(wr + u);
#pragma unroll
        for (int r = 0; r < SMALL_ROWS; ++r) {
            if (r < R) {
                const float4 v = *reinterpret_cast<const float4*>(xs + (size_t)r * U + u);
                acc[r] = fmaf(v.w, w.w, fmaf(v.z, w.z, fmaf(v.y, w.y, fmaf(v.x, w.x, acc[r]))));
            }
        }
    }
#pragma unroll
    for (int r = 0; r < SMALL_ROWS; ++r) {
        if (r >= R) break;
        float a = acc[r];
        for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        if (lane == 0) {
            const size_t m = (size_t)rbase + r;
            float z;
            if (col < ho.K) {
                z = a + ho.bias[0][col];
                ho.c[m * ho.K + col] = msig(z);
            } else if (col == ho.K) {
                z = a + ho.bias[1][0];
                ho.a[m] = msig(z);
            } else {
                z = a + ho.bias[2][col - ho.K - 1];
                ho.H[m * ho.NB + (col - ho.K - 1)] = msig(z);
            }
            if (ho.z) ho.z[m * ho.ldz + col] = z;
        }
    }
}

constexpr int DZ_ROWS = 16, DZ_TX = 128, DZ_TY = 4;

// dZ = g * msig'(z) for the three heads at once, plus per-row-tile column sums (bias-gradient partials).  A CTA owns 16 rows:
// thread (tx, ty) takes columns tx, tx + 128, ... of rows ty, ty + 4, ...; the four row groups' sums meet in shared memory in
// a fixed order.  (First version: one thread per column walking the 32 rows, 4 warps per SM: 82 us at R = 5504, J = 376.)
template <bool FAST>
__global__ void __launch_bounds__(DZ_TX * DZ_TY)
ctrl_dz_kernel(const float* __restrict__ z, const float* __restrict__ gc, const float* __restrict__ ga,
               const float* __restrict__ gH, float* __restrict__ dz, float* __restrict__ bpart, int R, int K, int NB, int ldz,
               __nv_bfloat16* __restrict__ planes, int ld_planes, size_t plane) {
    __shared__ float red[DZ_TY][DZ_TX];
    const int J = K + 1 + NB;
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int r0 = blockIdx.x * DZ_ROWS, r1 = min(R, r0 + DZ_ROWS);
    for (int j0 = 0; j0 < ldz; j0 += DZ_TX) {
        const int j = j0 + tx;
        float sum = 0.f;
        if (j < J) {
#pragma unroll 4
            for (int r = r0 + ty; r < r1; r += DZ_TY) {
                float g;
                if (j < K) g = gc ? gc[(size_t)r * K + j] : 0.f;
                else if (j == K) g = ga ? ga[r] : 0.f;
                else g = gH ? gH[(size_t)r * NB + (j - K - 1)] : 0.f;
                const float zz = z[(size_t)r * ldz + j];
                const float d = g * (FAST ? msig_grad_fast(zz) : msig_grad(zz));
                dz[(size_t)r * ldz + j] = d;
                if (planes) {                          // the bf16 x 3 split the tensor-core product of gx reads (K-major in j)
                    __nv_bfloat16 h[3];
                    split3(d, h);
#pragma unroll
                    for (int s = 0; s < 3; ++s) planes[s * plane + (size_t)r * ld_planes + j] = h[s];
                }
                sum += d;
            }
        } else if (j < ldz) {
            for (int r = r0 + ty; r < r1; r += DZ_TY) dz[(size_t)r * ldz + j] = 0.f;
        }
        red[ty][tx] = sum;
        __syncthreads();
        if (ty == 0 && j < ldz) bpart[(size_t)blockIdx.x * ldz + j] = (red[0][tx] + red[1][tx]) + (red[2][tx] + red[3][tx]);
        __syncthreads();
    }
}

// Weight gradients: fixed-order sum of the split partials [S][J][U], scattered to the three matrices; bias gradients:
// fixed-order sum of the row-tile partials.
__global__ void __launch_bounds__(256)
ctrl_wgrad_finish_kernel(const float* __restrict__ part, int S, const float* __restrict__ bpart, int tiles, int ldz,
                         float* __restrict__ gwc, float* __restrict__ gwa, float* __restrict__ gwh,
                         float* __restrict__ gbc, float* __restrict__ gba, float* __restrict__ gbh, int U, int K, int NB) {
    const int J = K + 1 + NB, U4 = U / 4;
    const size_t total = (size_t)J * U4, plane = (size_t)J * U;
    const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (size_t e = gid; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int j = (int)(e / U4), u = (int)(e % U4) * 4;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
        for (int s = 0; s < S; ++s) {
            const float4 v = *reinterpret_cast<const float4*>(part + s * plane + (size_t)j * U + u);
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        float* dst = j < K ? gwc + (size_t)j * U : (j == K ? gwa : gwh + (size_t)(j - K - 1) * U);
        *reinterpret_cast<float4*>(dst + u) = acc;
    }
    // bias gradients: one warp per column, lane l sums tiles l, l + 32, ... and the lanes meet in a fixed tree (one thread per
    // column walked all tiles in sequence: 172 dependent additions at R = 5504, most of this kernel's 13 us)
    const size_t wid = gid >> 5;
    if (wid < (size_t)J) {
        const int j = (int)wid, lane = threadIdx.x & 31;
        float sum = 0.f;
        for (int t = lane; t < tiles; t += 32) sum += bpart[(size_t)t * ldz + j];
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        if (lane == 0) {
            if (j < K) gbc[j] = sum; else if (j == K) gba[0] = sum; else gbh[j - K - 1] = sum;
        }
    }
}

struct CtrlPlan {
    int J, ldz, tiles, S, ksplit;
    size_t part_floats, bpart_floats, dz_floats;
};
static CtrlPlan ctrl_plan(int R, int U, int K, int NB) {
    CtrlPlan p;
    p.J = K + 1 + NB;
    p.ldz = (p.J + 3) & ~3;
    p.tiles = (R + DZ_ROWS - 1) / DZ_ROWS;
    const int ctas = ((p.J + G_BM - 1) / G_BM) * ((U + G_BN - 1) / G_BN);
    int S = (2 * 148 + ctas - 1) / ctas;
    S = std::max(1, std::min(S, (R + 63) / 64));
    p.ksplit = (((R + S - 1) / S) + 15) & ~15;
    p.S = (R + p.ksplit - 1) / p.ksplit;
    p.part_floats = (size_t)p.S * p.J * U;
    p.bpart_floats = (size_t)p.tiles * p.ldz;
    p.dz_floats = (size_t)R * p.ldz;
    return p;
}


// ================================================================================================
// Tensor-core path (tcgen05 + TMEM + TMA) for R >= a few hundred rows: the same three products as C = A B^T with both
// operands K-major bf16 x 3 splits (v = v1 + v2 + v3, 8 mantissa bits each: 24 bits, i.e. the fp32 value exactly), six
// products a1b1 + (a1b2 + a2b1) + (a1b3 + a2b2 + a3b1) accumulated in fp32 in TMEM -- the dropped terms are below 2^-24 of the
// product, so the result is an fp32-accurate product and the 1e-3 waveform bound holds as with FFMA.  One 128 x 128 tile per
// CTA (256 threads: warp 0 TMA producer, warp 1 MMA issuer, warps 4-7 TMEM readers, all eight the epilogue), K in panels of 64 (one 128-byte swizzle
// atom) through a two-stage ring of 96 KB; blockIdx.z splits K for the weight gradient (partials summed in fixed order by
// ctrl_wgrad_finish_kernel: deterministic).  Pre-passes write the split planes, transposed where the product needs the
// other operand layout.  Ragged edges: the TMA unit zero-fills what lies outside a matrix.
// ================================================================================================
namespace ctc {
constexpr int BM = 128, BN = 128, BK = 64, NST = 2, THREADS = 256;
constexpr int PANEL = 128 * BK * 2;                    // bytes of one split of one operand: 128 rows x 128 B
constexpr int STAGE = 6 * PANEL;
constexpr int TMEM_COLS = 4 * BN;                      // four accumulators (two halves of K x {leading product, the rest})
constexpr int TILE_LD = BN + 4;                        // fp32 staging tile of the epilogue, conflict-free for float4 rows
static_assert((size_t)BM * TILE_LD * 4 <= (size_t)STAGE, "the epilogue tile lives in the first stage buffer");
constexpr size_t SMEM = 1024 + (size_t)NST * STAGE + 256;
constexpr uint32_t IDESC = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
struct Maps { CUtensorMap a[3], b[3]; };
}  // namespace ctc

template <int EPI>
__device__ __forceinline__ void ctrl_tc_tile(const ctc::Maps& maps, int M, int N, int Kd, int ksplit, float* __restrict__ C, int ldc,
                                             const HeadsOut& ho, int bx, int by, int bz) {
    using namespace ctc;
    extern __shared__ unsigned char ctc_raw[];
    unsigned char* base = reinterpret_cast<unsigned char*>(((uintptr_t)ctc_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + NST * STAGE);
    uint64_t* full = bars;                   // [NST] TMA -> MMA
    uint64_t* empty = bars + NST;            // [NST] MMA -> TMA
    uint64_t* tfull = bars + 2 * NST;        // [1]   MMA -> epilogue
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = bx * BM, n0 = by * BN;
    const int kbeg = bz * ksplit, kend = min(Kd, kbeg + ksplit);
    const int nk = (kend - kbeg + BK - 1) / BK;          // >= 1 (host)
    const int nk0 = (nk + 1) / 2;                        // panels of the first half of K

    if (threadIdx.x == 0) {
        for (int i = 0; i < NST; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        mbar_init(tfull, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            for (int it = 0; it < nk; ++it) {
                const uint32_t st = it % NST, ph = (it / NST) & 1u;
                mbar_wait(&empty[st], ph ^ 1u);
                mbar_expect_tx(&full[st], (uint32_t)STAGE);
                unsigned char* dst = base + st * STAGE;
                const int k0 = kbeg + it * BK;
                for (int s = 0; s < 3; ++s) tma_load_2d(dst + s * PANEL, &maps.a[s], k0, m0, &full[st]);
                for (int s = 0; s < 3; ++s) tma_load_2d(dst + (3 + s) * PANEL, &maps.b[s], k0, n0, &full[st]);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // The tensor core adds into its fp32 accumulator by truncation: one accumulator over all 6 x K/16 instructions
            // is off by up to 1e-4 at |z| = 20.  Four accumulators instead -- per half of K, the leading product a1 b1 apart from
            // the five small ones (2^-8 and 2^-16 of it) -- and one rounded sum of the four in the epilogue: 8e-6 in the same
            // case, below the 1e-5 of an FFMA chain (profiles/summary_r02.md section 10).
            const int ca[6] = {0, 0, 1, 0, 1, 2}, cb[6] = {0, 1, 0, 2, 1, 0};
            uint32_t started = 0;
            for (int it = 0; it < nk; ++it) {
                const uint32_t st = it % NST, ph = (it / NST) & 1u;
                mbar_wait(&full[st], ph);
                tc_fence_after();
                const uint32_t sb = smem_u32(base + st * STAGE);
                const uint32_t half = it >= nk0 ? 2u : 0u;
                for (int c = 0; c < 6; ++c) {
                    const uint32_t accn = half + (c ? 1u : 0u);
                    for (int ks = 0; ks < BK / 16; ++ks) {
                        const uint64_t ad = umma_desc(sb + ca[c] * PANEL + ks * 32, 128);
                        const uint64_t bd = umma_desc(sb + (3 + cb[c]) * PANEL + ks * 32, 128);
                        umma_bf16(tmem_base + accn * BN, ad, bd, IDESC, (started >> accn) & 1u);
                        started |= 1u << accn;
                    }
                }
                umma_commit(&empty[st]);
            }
            umma_commit(tfull);
        }
    } else if (warp >= 4) {
        // ===== TMEM -> shared memory: the rounded sum of the accumulators, row by row =====
        const int quad = warp & 3;                          // the TMEM lane quadrant this warp may read
        const int row = quad * 32 + lane;
        mbar_wait(tfull, 0);                                // every MMA has completed: the stage buffers are free
        tc_fence_after();
        float* tile = reinterpret_cast<float*>(base);       // [BM][TILE_LD] fp32 over the first stage
        const uint32_t t0 = tmem_base + ((uint32_t)(quad * 32) << 16);
#pragma unroll 1
        for (int j0 = 0; j0 < BN; j0 += 16) {
            float a0[16], b0[16], v[16];
            tmem_ld16(t0 + 0 * BN + j0, a0);
            tmem_ld16(t0 + 1 * BN + j0, b0);
            if (nk >= 2) {
                float a1[16], b1[16];
                tmem_ld16(t0 + 2 * BN + j0, a1);
                tmem_ld16(t0 + 3 * BN + j0, b1);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = (a0[i] + a1[i]) + (b0[i] + b1[i]);
            } else {
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = a0[i] + b0[i];
            }
#pragma unroll
            for (int q = 0; q < 4; ++q)
                *reinterpret_cast<float4*>(tile + (size_t)row * TILE_LD + j0 + 4 * q) = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(ctc::TMEM_COLS) : "memory");
    }
    // ===== all eight warps: bias, activation and stores, a row per warp and consecutive columns per lane =====
    const float* tile = reinterpret_cast<const float*>(base);
    for (int r = warp; r < BM; r += THREADS / 32) {
        const int m = m0 + r;
        if (m >= M) break;
#pragma unroll
        for (int i = 0; i < BN / 32; ++i) {
            const int col = n0 + lane + 32 * i;
            if (col >= N) continue;
            const float acc = tile[(size_t)r * TILE_LD + lane + 32 * i];
            if (EPI == EPI_STORE) {
                C[(size_t)bz * M * ldc + (size_t)m * ldc + col] = acc;
            } else {
                float z;
                if (col < ho.K) {
                    z = acc + ho.bias[0][col];
                    ho.c[(size_t)m * ho.K + col] = msig_fast(z);
                } else if (col == ho.K) {
                    z = acc + ho.bias[1][0];
                    ho.a[m] = msig_fast(z);
                } else {
                    z = acc + ho.bias[2][col - ho.K - 1];
                    ho.H[(size_t)m * ho.NB + (col - ho.K - 1)] = msig_fast(z);
                }
                if (ho.z) ho.z[(size_t)m * ho.ldz + col] = z;
            }
        }
    }
}

template <int EPI>
__global__ void __launch_bounds__(ctc::THREADS, 1)
ctrl_tc_gemm_kernel(const __grid_constant__ ctc::Maps maps, int M, int N, int Kd, int ksplit, float* __restrict__ C, int ldc, HeadsOut ho) {
    ctrl_tc_tile<EPI>(maps, M, N, Kd, ksplit, C, ldc, ho, blockIdx.x, blockIdx.y, blockIdx.z);
}

// Two independent products in one launch (the backward's dX and dW): the second one's CTAs fill what the first leaves of its
// last wave -- 172 + 120 CTAs are two waves of 148 SMs, launched apart they were three.
struct TcProblem {
    ctc::Maps maps;
    int M, N, Kd, ksplit, ldc;
    float* C;
    unsigned gx, gy, gz;
};
__global__ void __launch_bounds__(ctc::THREADS, 1)
ctrl_tc_gemm2_kernel(const __grid_constant__ TcProblem p0, const __grid_constant__ TcProblem p1) {
    const unsigned n0 = p0.gx * p0.gy * p0.gz;
    const bool first = blockIdx.x < n0;
    const TcProblem& p = first ? p0 : p1;
    const unsigned b = first ? blockIdx.x : blockIdx.x - n0;
    const HeadsOut none{};
    ctrl_tc_tile<EPI_STORE>(p.maps, p.M, p.N, p.Kd, p.ksplit, p.C, p.ldc, none, (int)(b % p.gx), (int)((b / p.gx) % p.gy), (int)(b / (p.gx * p.gy)));
}

// One launch splits up to three matrices (fewer launches matter: the path is host-bound in eager mode at the shapes it serves).
struct SplitJob {
    Rows3 src;
    int rows, cols, ld_out;
    __nv_bfloat16* planes;
    size_t plane;
    unsigned gx, gy;              // this job's grid
};
struct SplitJobs { SplitJob job[3]; };

// planes[s][r][c] (pitch ld_out) = split s of src(r, c); four columns per thread (cols % 4 == 0, rows 16-byte aligned)
__global__ void __launch_bounds__(256) ctrl_split_rows_kernel(const __grid_constant__ SplitJobs jobs) {
    const SplitJob& jb = jobs.job[blockIdx.y];
    if (blockIdx.x >= jb.gx) return;
    const int c4 = jb.cols >> 2;
    const size_t total = (size_t)jb.rows * c4;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)jb.gx * blockDim.x) {
        const int r = (int)(e / c4), c = (int)(e % c4) * 4;
        const float4 v = *reinterpret_cast<const float4*>(jb.src.row(r) + c);
        __nv_bfloat16 h[4][3];
        split3(v.x, h[0]); split3(v.y, h[1]); split3(v.z, h[2]); split3(v.w, h[3]);
#pragma unroll
        for (int s = 0; s < 3; ++s) {
            __nv_bfloat162 lo, hi;
            lo.x = h[0][s]; lo.y = h[1][s]; hi.x = h[2][s]; hi.y = h[3][s];
            __nv_bfloat162* dst = reinterpret_cast<__nv_bfloat162*>(jb.planes + s * jb.plane + (size_t)r * jb.ld_out + c);
            dst[0] = lo;
            dst[1] = hi;
        }
    }
}

// planes[s][c][r] (pitch ld_out) = split s of src(r, c): 64 x 32 tiles through shared memory, two rows per store
__global__ void __launch_bounds__(256) ctrl_split_transpose_kernel(const __grid_constant__ SplitJobs jobs) {
    __shared__ float tile[64][33];
    const SplitJob& jb = jobs.job[blockIdx.z];
    if (blockIdx.x >= jb.gx || blockIdx.y >= jb.gy) return;
    const int rows = jb.rows, cols = jb.cols;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;              // 32 x 8
    const int r0 = blockIdx.x * 64, c0 = blockIdx.y * 32;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int r = r0 + ty + 8 * i, c = c0 + tx;
        tile[ty + 8 * i][tx] = (r < rows && c < cols) ? jb.src.row(r)[c] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int c = c0 + ty + 8 * i, r = r0 + 2 * tx;
        if (c >= cols || r >= rows) continue;
        __nv_bfloat16 h0[3], h1[3];
        split3(tile[2 * tx][ty + 8 * i], h0);
        split3(tile[2 * tx + 1][ty + 8 * i], h1);
#pragma unroll
        for (int s = 0; s < 3; ++s) {
            __nv_bfloat16* dst = jb.planes + s * jb.plane + (size_t)c * jb.ld_out + r;
            if (r + 1 < rows) { __nv_bfloat162 p; p.x = h0[s]; p.y = h1[s]; *reinterpret_cast<__nv_bfloat162*>(dst) = p; }
            else dst[0] = h0[s];
        }
    }
}

static SplitJob split_rows_job(Rows3 src, int rows, int cols, __nv_bfloat16* planes, int ld_out) {
    const size_t items = (size_t)rows * (cols / 4);
    return SplitJob{src, rows, cols, ld_out, planes, (size_t)rows * ld_out, (unsigned)std::min<size_t>((items + 255) / 256, 148 * 16), 1};
}
static SplitJob split_transpose_job(Rows3 src, int rows, int cols, __nv_bfloat16* planes, int ld_out) {
    return SplitJob{src, rows, cols, ld_out, planes, (size_t)cols * ld_out, (unsigned)((rows + 63) / 64), (unsigned)((cols + 31) / 32)};
}
static int launch_split_rows(const SplitJob* jb, int n, cudaStream_t st) {
    SplitJobs jobs{};
    unsigned gx = 1;
    for (int i = 0; i < n; ++i) { jobs.job[i] = jb[i]; gx = std::max(gx, jb[i].gx); }
    ctrl_split_rows_kernel<<<dim3(gx, n), 256, 0, st>>>(jobs);
    DDSP_CHECK_LAUNCH("ctrl_split_rows_kernel");
    return 0;
}
static int launch_split_transpose(const SplitJob* jb, int n, cudaStream_t st) {
    SplitJobs jobs{};
    unsigned gx = 1, gy = 1;
    for (int i = 0; i < n; ++i) { jobs.job[i] = jb[i]; gx = std::max(gx, jb[i].gx); gy = std::max(gy, jb[i].gy); }
    DDSP_REQUIRE(gy <= 65535, "ctrl_split_transpose_kernel: matrix too wide");
    ctrl_split_transpose_kernel<<<dim3(gx, gy, n), 256, 0, st>>>(jobs);
    DDSP_CHECK_LAUNCH("ctrl_split_transpose_kernel");
    return 0;
}

struct CtrlTcPlan {
    int J, ldz, J8, R8, tiles, S, ksplit, S2, ksplit2;      // S2 / ksplit2: dW's K split when it shares a launch with dX
    // byte offsets into the workspace
    size_t off_xp, off_wp;                                      // forward: x and W planes, K-major in U
    size_t off_part, off_bpart, off_dz, off_dzp, off_wt, off_dzt, off_xt;     // backward
    size_t fwd_bytes, bwd_bytes;
};
static CtrlTcPlan ctrl_tc_plan(int R, int U, int K, int NB) {
    CtrlTcPlan p;
    p.J = K + 1 + NB;
    p.ldz = (p.J + 3) & ~3;
    p.J8 = (p.J + 7) & ~7;
    p.R8 = (R + 7) & ~7;
    p.tiles = (R + DZ_ROWS - 1) / DZ_ROWS;
    const int ctas = ((p.J + ctc::BM - 1) / ctc::BM) * ((U + ctc::BN - 1) / ctc::BN);
    auto split = [&](int s_want, int& ksplit, int& S) {
        s_want = std::max(1, std::min(s_want, (R + 127) / 128));
        ksplit = (((R + s_want - 1) / s_want) + ctc::BK - 1) / ctc::BK * ctc::BK;
        S = (R + ksplit - 1) / ksplit;
    };
    split(148 / std::max(ctas, 1), p.ksplit, p.S);                       // dW alone: one wave
    // dW launched with dX: as many K splits as fit the SMs dX leaves free in its last wave (one more wave if that is too few)
    const int n_dx = ((R + ctc::BM - 1) / ctc::BM) * ((U + ctc::BN - 1) / ctc::BN);
    int free_sms = (n_dx + 147) / 148 * 148 - n_dx;
    if (free_sms < ctas) free_sms += 148;
    split(free_sms / std::max(ctas, 1), p.ksplit2, p.S2);
    const size_t h = sizeof(__nv_bfloat16);
    size_t off = 0;
    p.off_xp = off; off += align_up(3 * h * (size_t)R * U, 1024);
    p.off_wp = off; off += align_up(3 * h * (size_t)p.J * U, 1024);
    p.fwd_bytes = off;
    off = 0;
    p.off_part = off;  off += align_up(sizeof(float) * (size_t)std::max(p.S, p.S2) * p.J * U, 1024);
    p.off_bpart = off; off += align_up(sizeof(float) * (size_t)p.tiles * p.ldz, 1024);
    p.off_dz = off;    off += align_up(sizeof(float) * (size_t)R * p.ldz, 1024);
    p.off_dzp = off;   off += align_up(3 * h * (size_t)R * p.J8, 1024);
    p.off_wt = off;    off += align_up(3 * h * (size_t)U * p.J8, 1024);
    p.off_dzt = off;   off += align_up(3 * h * (size_t)p.J * p.R8, 1024);
    p.off_xt = off;    off += align_up(3 * h * (size_t)U * p.R8, 1024);
    p.bwd_bytes = off;
    return p;
}

static int ctrl_tc_maps(ctc::Maps& maps, __nv_bfloat16* a, uint64_t a_rows, uint64_t a_ld, __nv_bfloat16* b, uint64_t b_rows, uint64_t b_ld,
                        uint64_t kd) {
    for (int s = 0; s < 3; ++s) {
        if (int rc = make_map_2d(&maps.a[s], a + (size_t)s * a_rows * a_ld, a_rows, kd, ctc::BM, ctc::BK, a_ld)) return rc;
        if (int rc = make_map_2d(&maps.b[s], b + (size_t)s * b_rows * b_ld, b_rows, kd, ctc::BN, ctc::BK, b_ld)) return rc;
    }
    return 0;
}

template <int EPI>
static int ctrl_tc_launch(const ctc::Maps& maps, int M, int N, int Kd, int ksplit, int S, float* C, int ldc, const HeadsOut& ho, cudaStream_t st,
                          const char* what) {
    cudaError_t e = cudaFuncSetAttribute(ctrl_tc_gemm_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ctc::SMEM);
    DDSP_REQUIRE(e == cudaSuccess, "%s: cannot reserve %zu B of shared memory: %s", what, (size_t)ctc::SMEM, cudaGetErrorString(e));
    dim3 grid((M + ctc::BM - 1) / ctc::BM, (N + ctc::BN - 1) / ctc::BN, S);
    DDSP_REQUIRE(grid.y <= 65535 && grid.z <= 65535, "%s: matrix too large", what);
    ctrl_tc_gemm_kernel<EPI><<<grid, ctc::THREADS, ctc::SMEM, st>>>(maps, M, N, Kd, ksplit, C, ldc, ho);
    DDSP_CHECK_LAUNCH(what);
    return 0;
}

}  // namespace ddsp

using namespace ddsp;

extern "C" int ddsp_ctrl_heads_ldz(int K, int NB) { return (K + 1 + NB + 3) & ~3; }

extern "C" size_t ddsp_ctrl_heads_workspace_bytes(int R, int U, int K, int NB) {
    if (R < 1 || U < 1 || K < 1 || NB < 1) return 0;
    const CtrlPlan p = ctrl_plan(R, U, K, NB);
    return sizeof(float) * (p.part_floats + p.bpart_floats + p.dz_floats);
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

extern "C" int ddsp_ctrl_heads_forward(const float* x, const float* w_c, const float* b_c, const float* w_a, const float* b_a,
                                       const float* w_h, const float* b_h, float* c, float* a, float* H, float* z, int R,
                                       int U, int K, int NB, void* stream) {
    DDSP_REQUIRE(x && w_c && b_c && w_a && b_a && w_h && b_h && c && a && H, "ddsp_ctrl_heads_forward: null pointer argument");
    DDSP_REQUIRE(R >= 1 && K >= 1 && NB >= 1, "ddsp_ctrl_heads_forward: bad R=%d K=%d NB=%d", R, K, NB);
    DDSP_REQUIRE(U >= 4 && U % 4 == 0, "ddsp_ctrl_heads_forward: U=%d must be a multiple of 4", U);
    DDSP_REQUIRE(aligned16(x) && aligned16(w_c) && aligned16(w_a) && aligned16(w_h),
                 "ddsp_ctrl_heads_forward: x and the weight matrices must be 16-byte aligned");
    const int J = K + 1 + NB;
    Rows3 A = rows_plain(x, U);
    Rows3 B{{w_c, w_a, w_h}, K, K + 1, U};
    HeadsOut ho{{b_c, b_a, b_h}, c, a, H, z, K, NB, ddsp_ctrl_heads_ldz(K, NB)};
    if (R <= SMALL_MAX_ROWS && (size_t)SMALL_ROWS * U * sizeof(float) <= 48 * 1024) {
        const int rows = std::min(R, SMALL_ROWS);
        ctrl_heads_small_kernel<<<dim3((J + SMALL_WARPS - 1) / SMALL_WARPS, (R + SMALL_ROWS - 1) / SMALL_ROWS), 32 * SMALL_WARPS,
                                  (size_t)rows * U * sizeof(float), static_cast<cudaStream_t>(stream)>>>(x, B, R, U, J, ho);
        DDSP_CHECK_LAUNCH("ctrl_heads_small_kernel");
        return 0;
    }
    dim3 grid((R + G_BM - 1) / G_BM, (J + G_BN - 1) / G_BN, 1);
    DDSP_REQUIRE(grid.y <= 65535, "ddsp_ctrl_heads_forward: K + NB too large");
    ctrl_gemm_kernel<true, true, EPI_HEADS><<<grid, G_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(A, B, R, J, U, U, nullptr, 0, ho);
    DDSP_CHECK_LAUNCH("ctrl_gemm_kernel<heads>");
    return 0;
}

extern "C" int ddsp_ctrl_heads_backward(const float* x, const float* w_c, const float* w_a, const float* w_h, const float* z,
                                        const float* g_c, const float* g_a, const float* g_h, float* gx, float* gw_c,
                                        float* gb_c, float* gw_a, float* gb_a, float* gw_h, float* gb_h, void* workspace,
                                        int R, int U, int K, int NB, void* stream) {
    DDSP_REQUIRE(x && w_c && w_a && w_h && z && workspace, "ddsp_ctrl_heads_backward: null pointer argument");
    DDSP_REQUIRE(gw_c && gb_c && gw_a && gb_a && gw_h && gb_h, "ddsp_ctrl_heads_backward: null gradient output");
    DDSP_REQUIRE(R >= 1 && K >= 1 && NB >= 1, "ddsp_ctrl_heads_backward: bad R=%d K=%d NB=%d", R, K, NB);
    DDSP_REQUIRE(U >= 4 && U % 4 == 0, "ddsp_ctrl_heads_backward: U=%d must be a multiple of 4", U);
    DDSP_REQUIRE(aligned16(x) && aligned16(w_c) && aligned16(w_a) && aligned16(w_h) && aligned16(z) && aligned16(workspace) &&
                 aligned16(gw_c) && aligned16(gw_a) && aligned16(gw_h) && (!gx || aligned16(gx)),
                 "ddsp_ctrl_heads_backward: matrices must be 16-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const CtrlPlan p = ctrl_plan(R, U, K, NB);
    float* part = static_cast<float*>(workspace);
    float* bpart = part + p.part_floats;
    float* dz = bpart + p.bpart_floats;                  // part_floats and bpart_floats are multiples of 4
    ctrl_dz_kernel<false><<<p.tiles, dim3(DZ_TX, DZ_TY), 0, st>>>(z, g_c, g_a, g_h, dz, bpart, R, K, NB, p.ldz, nullptr, 0, 0);
    DDSP_CHECK_LAUNCH("ctrl_dz_kernel");
    HeadsOut none{};
    if (gx) {       // gx[R,U] = dZ[R,J] * Wcat[J,U]
        Rows3 A = rows_plain(dz, p.ldz);
        Rows3 B{{w_c, w_a, w_h}, K, K + 1, U};
        dim3 grid((R + G_BM - 1) / G_BM, (U + G_BN - 1) / G_BN, 1);
        ctrl_gemm_kernel<true, false, EPI_STORE><<<grid, G_THREADS, 0, st>>>(A, B, R, U, p.J, (p.J + 15) & ~15, gx, U, none);
        DDSP_CHECK_LAUNCH("ctrl_gemm_kernel<dx>");
    }
    {               // partial[s][J,U] = dZ[rows of split s]^T * X[rows of split s]
        Rows3 A = rows_plain(dz, p.ldz);
        Rows3 B = rows_plain(x, U);
        dim3 grid((p.J + G_BM - 1) / G_BM, (U + G_BN - 1) / G_BN, p.S);
        ctrl_gemm_kernel<false, false, EPI_STORE><<<grid, G_THREADS, 0, st>>>(A, B, p.J, U, R, p.ksplit, part, U, none);
        DDSP_CHECK_LAUNCH("ctrl_gemm_kernel<dw>");
    }
    const size_t total4 = (size_t)p.J * (U / 4);
    const int blocks = (int)std::min<size_t>((total4 + 255) / 256, 148 * 8);
    ctrl_wgrad_finish_kernel<<<std::max(blocks, (p.J + 7) / 8), 256, 0, st>>>(part, p.S, bpart, p.tiles, p.ldz, gw_c, gw_a, gw_h,
                                                                                 gb_c, gb_a, gb_h, U, K, NB);
    DDSP_CHECK_LAUNCH("ctrl_wgrad_finish_kernel");
    return 0;
}

// ---- tensor-core variants (same arguments plus a workspace; U % 8 == 0) ---------------------------------------------------
extern "C" size_t ddsp_ctrl_heads_tc_workspace_bytes(int R, int U, int K, int NB, int backward) {
    if (R < 1 || U < 8 || U % 8 || K < 1 || NB < 1) return 0;
    const CtrlTcPlan p = ctrl_tc_plan(R, U, K, NB);
    return backward ? p.bwd_bytes : p.fwd_bytes;
}

extern "C" int ddsp_ctrl_heads_forward_tc(const float* x, const float* w_c, const float* b_c, const float* w_a, const float* b_a,
                                          const float* w_h, const float* b_h, float* c, float* a, float* H, float* z, void* workspace,
                                          int R, int U, int K, int NB, void* stream) {
    DDSP_REQUIRE(x && w_c && b_c && w_a && b_a && w_h && b_h && c && a && H && workspace, "ddsp_ctrl_heads_forward_tc: null pointer argument");
    DDSP_REQUIRE(R >= 1 && K >= 1 && NB >= 1, "ddsp_ctrl_heads_forward_tc: bad R=%d K=%d NB=%d", R, K, NB);
    DDSP_REQUIRE(U >= 8 && U % 8 == 0, "ddsp_ctrl_heads_forward_tc: U=%d must be a multiple of 8", U);
    DDSP_REQUIRE(aligned16(x) && aligned16(w_c) && aligned16(w_a) && aligned16(w_h) && aligned16(workspace),
                 "ddsp_ctrl_heads_forward_tc: x, the weight matrices and the workspace must be 16-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const CtrlTcPlan p = ctrl_tc_plan(R, U, K, NB);
    char* ws = static_cast<char*>(workspace);
    __nv_bfloat16* xp = reinterpret_cast<__nv_bfloat16*>(ws + p.off_xp);
    __nv_bfloat16* wp = reinterpret_cast<__nv_bfloat16*>(ws + p.off_wp);
    Rows3 W{{w_c, w_a, w_h}, K, K + 1, U};
    {
        const SplitJob jb[2] = {split_rows_job(rows_plain(x, U), R, U, xp, U), split_rows_job(W, p.J, U, wp, U)};
        if (int rc = launch_split_rows(jb, 2, st)) return rc;
    }
    ctc::Maps maps;
    if (int rc = ctrl_tc_maps(maps, xp, (uint64_t)R, (uint64_t)U, wp, (uint64_t)p.J, (uint64_t)U, (uint64_t)U)) return rc;
    HeadsOut ho{{b_c, b_a, b_h}, c, a, H, z, K, NB, p.ldz};
    return ctrl_tc_launch<EPI_HEADS>(maps, R, p.J, U, (U + ctc::BK - 1) / ctc::BK * ctc::BK, 1, nullptr, 0, ho, st, "ctrl_tc_gemm_kernel<heads>");
}

extern "C" int ddsp_ctrl_heads_backward_tc(const float* x, const float* w_c, const float* w_a, const float* w_h, const float* z,
                                           const float* g_c, const float* g_a, const float* g_h, float* gx, float* gw_c,
                                           float* gb_c, float* gw_a, float* gb_a, float* gw_h, float* gb_h, void* workspace,
                                           int R, int U, int K, int NB, void* stream) {
    DDSP_REQUIRE(x && w_c && w_a && w_h && z && workspace, "ddsp_ctrl_heads_backward_tc: null pointer argument");
    DDSP_REQUIRE(gw_c && gb_c && gw_a && gb_a && gw_h && gb_h, "ddsp_ctrl_heads_backward_tc: null gradient output");
    DDSP_REQUIRE(R >= 1 && K >= 1 && NB >= 1, "ddsp_ctrl_heads_backward_tc: bad R=%d K=%d NB=%d", R, K, NB);
    DDSP_REQUIRE(U >= 8 && U % 8 == 0, "ddsp_ctrl_heads_backward_tc: U=%d must be a multiple of 8", U);
    DDSP_REQUIRE(aligned16(x) && aligned16(w_c) && aligned16(w_a) && aligned16(w_h) && aligned16(z) && aligned16(workspace) &&
                 aligned16(gw_c) && aligned16(gw_a) && aligned16(gw_h) && (!gx || aligned16(gx)),
                 "ddsp_ctrl_heads_backward_tc: matrices must be 16-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const CtrlTcPlan p = ctrl_tc_plan(R, U, K, NB);
    char* ws = static_cast<char*>(workspace);
    float* part = reinterpret_cast<float*>(ws + p.off_part);
    float* bpart = reinterpret_cast<float*>(ws + p.off_bpart);
    float* dz = reinterpret_cast<float*>(ws + p.off_dz);
    __nv_bfloat16* dzp = reinterpret_cast<__nv_bfloat16*>(ws + p.off_dzp);
    __nv_bfloat16* wt = reinterpret_cast<__nv_bfloat16*>(ws + p.off_wt);
    __nv_bfloat16* dzt = reinterpret_cast<__nv_bfloat16*>(ws + p.off_dzt);
    __nv_bfloat16* xt = reinterpret_cast<__nv_bfloat16*>(ws + p.off_xt);
    ctrl_dz_kernel<true><<<p.tiles, dim3(DZ_TX, DZ_TY), 0, st>>>(z, g_c, g_a, g_h, dz, bpart, R, K, NB, p.ldz, gx ? dzp : nullptr, p.J8, (size_t)R * p.J8);
    DDSP_CHECK_LAUNCH("ctrl_dz_kernel");
    HeadsOut none{};
    ctc::Maps maps;
    {               // the transposed operands of both products in one launch: Wcat^T [U,J] (for gx), dZ^T [J,R] and X^T [U,R]
        Rows3 W{{w_c, w_a, w_h}, K, K + 1, U};
        const SplitJob jb[3] = {split_transpose_job(rows_plain(dz, p.ldz), R, p.J, dzt, p.R8), split_transpose_job(rows_plain(x, U), R, U, xt, p.R8),
                                split_transpose_job(W, p.J, U, wt, p.J8)};
        if (int rc = launch_split_transpose(jb, gx ? 3 : 2, st)) return rc;
    }
    int S_used = p.S;
    if (gx) {       // gx[R,U] = dZ[R,J] Wcat[J,U] (A = dZ, K-major in J: planes written by ctrl_dz_kernel; B = Wcat^T [U,J]) and, in
                    // the same launch, partial[s][J,U] = dZ[rows of split s]^T X[rows of split s] (A = dZ^T [J,R], B = X^T [U,R])
        TcProblem dx{}, dw{};
        if (int rc = ctrl_tc_maps(dx.maps, dzp, (uint64_t)R, (uint64_t)p.J8, wt, (uint64_t)U, (uint64_t)p.J8, (uint64_t)p.J)) return rc;
        if (int rc = ctrl_tc_maps(dw.maps, dzt, (uint64_t)p.J, (uint64_t)p.R8, xt, (uint64_t)U, (uint64_t)p.R8, (uint64_t)R)) return rc;
        dx.M = R; dx.N = U; dx.Kd = p.J; dx.ksplit = (p.J + ctc::BK - 1) / ctc::BK * ctc::BK; dx.ldc = U; dx.C = gx;
        dx.gx = (unsigned)((R + ctc::BM - 1) / ctc::BM); dx.gy = (unsigned)((U + ctc::BN - 1) / ctc::BN); dx.gz = 1;
        dw.M = p.J; dw.N = U; dw.Kd = R; dw.ksplit = p.ksplit2; dw.ldc = U; dw.C = part;
        dw.gx = (unsigned)((p.J + ctc::BM - 1) / ctc::BM); dw.gy = (unsigned)((U + ctc::BN - 1) / ctc::BN); dw.gz = (unsigned)p.S2;
        S_used = p.S2;
        cudaError_t e = cudaFuncSetAttribute(ctrl_tc_gemm2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ctc::SMEM);
        DDSP_REQUIRE(e == cudaSuccess, "ctrl_tc_gemm2_kernel: cannot reserve %zu B of shared memory: %s", (size_t)ctc::SMEM, cudaGetErrorString(e));
        const unsigned long long blocks2 = (unsigned long long)dx.gx * dx.gy + (unsigned long long)dw.gx * dw.gy * dw.gz;
        DDSP_REQUIRE(blocks2 <= 0x7fffffffull, "ctrl_tc_gemm2_kernel: matrix too large");
        ctrl_tc_gemm2_kernel<<<(unsigned)blocks2, ctc::THREADS, ctc::SMEM, st>>>(dx, dw);
        DDSP_CHECK_LAUNCH("ctrl_tc_gemm2_kernel");
    } else {        // partial[s][J,U] = dZ[rows of split s]^T * X[rows of split s]: A = dZ^T [J,R], B = X^T [U,R]
        if (int rc = ctrl_tc_maps(maps, dzt, (uint64_t)p.J, (uint64_t)p.R8, xt, (uint64_t)U, (uint64_t)p.R8, (uint64_t)R)) return rc;
        if (int rc = ctrl_tc_launch<EPI_STORE>(maps, p.J, U, R, p.ksplit, p.S, part, U, none, st, "ctrl_tc_gemm_kernel<dw>")) return rc;
    }
    const size_t total4 = (size_t)p.J * (U / 4);
    const int blocks = (int)std::min<size_t>((total4 + 255) / 256, 148 * 8);
    ctrl_wgrad_finish_kernel<<<std::max(blocks, (p.J + 7) / 8), 256, 0, st>>>(part, S_used, bpart, p.tiles, p.ldz, gw_c, gw_a, gw_h,
                                                                                 gb_c, gb_a, gb_h, U, K, NB);
    DDSP_CHECK_LAUNCH("ctrl_wgrad_finish_kernel");
    return 0;
}
