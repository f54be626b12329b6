// Controller tail for sm_100a -- replaces, as ONE forward launch and three backward launches,
//   harm_amps = modified_sigmoid(dense_harmonic(latent)); total_harm_amp = modified_sigmoid(dense_loudness(latent));
//   noise_distribution = modified_sigmoid(dense_filter(latent))           (model/autoencoder/decoder.py:74-76,96-100,110-116)
// The three Linear heads are one [R,U] x [U,J] product, J = K + 1 + NB, against the three weight matrices where they
// lie (no concatenated copy); bias, 2*sigmoid(z)^2.3026 + 1e-7 and the scatter into c / a / H -- the fp32, contiguous
// layouts the oscillator and noise kernels read -- happen in the GEMM epilogue.  Backward: dZ = g * msig'(z) with the
// bias-gradient partials (one pass), dX = dZ W (same GEMM kernel, other operand layout), dW = dZ^T X (split over the
// R dimension, partials summed in fixed order: deterministic).  FP32 FFMA throughout: the 1e-3 waveform bound against
// the fp32 reference leaves no room for a tf32/bf16 product here.
#include <algorithm>
#include <climits>
#include <cstdint>

#include "ddsp_common.cuh"

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

constexpr int DZ_ROWS = 32, DZ_THREADS = 128;

// dZ = g * msig'(z) for the three heads at once, plus per-row-tile column sums (bias-gradient partials).
__global__ void __launch_bounds__(DZ_THREADS)
ctrl_dz_kernel(const float* __restrict__ z, const float* __restrict__ gc, const float* __restrict__ ga,
               const float* __restrict__ gH, float* __restrict__ dz, float* __restrict__ bpart, int R, int K, int NB, int ldz) {
    const int J = K + 1 + NB;
    const int r0 = blockIdx.x * DZ_ROWS, r1 = min(R, r0 + DZ_ROWS);
    for (int j = threadIdx.x; j < ldz; j += DZ_THREADS) {
        float sum = 0.f;
        if (j < J) {
            for (int r = r0; r < r1; ++r) {
                float g;
                if (j < K) g = gc ? gc[(size_t)r * K + j] : 0.f;
                else if (j == K) g = ga ? ga[r] : 0.f;
                else g = gH ? gH[(size_t)r * NB + (j - K - 1)] : 0.f;
                const float d = g * msig_grad(z[(size_t)r * ldz + j]);
                dz[(size_t)r * ldz + j] = d;
                sum += d;
            }
        } else {
            for (int r = r0; r < r1; ++r) dz[(size_t)r * ldz + j] = 0.f;
        }
        bpart[(size_t)blockIdx.x * ldz + j] = sum;
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
        for (int s = 0; s < S; ++s) {
            const float4 v = *reinterpret_cast<const float4*>(part + s * plane + (size_t)j * U + u);
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        float* dst = j < K ? gwc + (size_t)j * U : (j == K ? gwa : gwh + (size_t)(j - K - 1) * U);
        *reinterpret_cast<float4*>(dst + u) = acc;
    }
    if (gid < (size_t)J) {
        const int j = (int)gid;
        float sum = 0.f;
        for (int t = 0; t < tiles; ++t) sum += bpart[(size_t)t * ldz + j];
        if (j < K) gbc[j] = sum; else if (j == K) gba[0] = sum; else gbh[j - K - 1] = sum;
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
    ctrl_dz_kernel<<<p.tiles, DZ_THREADS, 0, st>>>(z, g_c, g_a, g_h, dz, bpart, R, K, NB, p.ldz);
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
    ctrl_wgrad_finish_kernel<<<std::max(blocks, (p.J + 255) / 256), 256, 0, st>>>(part, p.S, bpart, p.tiles, p.ldz, gw_c, gw_a, gw_h,
                                                                                 gb_c, gb_a, gb_h, U, K, NB);
    DDSP_CHECK_LAUNCH("ctrl_wgrad_finish_kernel");
    return 0;
}
