// DFT-as-GEMM front end for one scale of the spectral loss on the 5th-generation tensor cores (tcgen05 + TMEM + TMA).
//
// north_star item 3 asks for "a tcgen05 DFT-as-GEMM ... only for the small FFT sizes where ncu shows it beats a
// shared-memory radix FFT".  This file is that experiment, complete enough to be measured against mss_fast_kernel<n>
// (csrc/ddsp_mss.cu) on the same frames: the forward of SSSLoss (mss_loss.py:25-33) for n_fft in {64, 128}, hop = n/4.
//
//   frames  : the reflect-padded signal is cut into rows of hop samples ("blocks").  Frame f is blocks f..f+3, so the
//             [128 frames x n] A tile of a GEMM is four K-panels, panel q = block rows f0+q .. f0+q+127 -- four plain 2-D
//             TMA loads of the same block matrix at row offsets 0..3 (no frame matrix is ever materialised);
//   DFT     : B operand = the real DFT matrix with the Hann window folded in, W[j][m] = w[m] cos(2 pi k m / n) for
//             j = k < n/2, w[m] sin(2 pi k m / n) for j = n/2 + k (slot j = n/2 carries the Nyquist bin), resident in
//             shared memory for the life of the CTA;
//   precision: bf16 x 3 split of both operands (x = x1 + x2 + x3, 8 mantissa bits each), six products
//             x1W1 + (x1W2 + x2W1) + (x1W3 + x2W2 + x3W1) accumulated in fp32 in TMEM; `splits` = 1 runs x1W1 only;
//   epilogue: four warps read the two signals' accumulators with tcgen05.ld, form |X|^2 and both L1 terms.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane), warps 2-5 = epilogue (TMEM lane
// quadrant = warp % 4).  Persistent: one CTA per SM strides over (voice, 128-frame tile) items.  Every mbarrier wait is
// bounded and traps instead of hanging.
#include "ddsp_umma.cuh"

namespace ddsp {

template <int NFFT>
struct DftGemm {
    static constexpr int HOP = NFFT / 4;                    // samples per block row = K-panel width
    static constexpr int ROW_BYTES = HOP * 2;               // bf16: 32 B (n = 64) or 64 B (n = 128) = the swizzle span
    static constexpr int KSTEPS = HOP / 16;                 // UMMA_K = 16 bf16 per instruction
    static constexpr int NST = NFFT == 64 ? 4 : 2;          // A stages in flight (one stage = one K-panel of one tile)
    static constexpr int A_PANEL = TILE_M * ROW_BYTES;      // one signal, one split
    static constexpr int W_PANEL = NFFT * ROW_BYTES;        // one split, one quarter of the DFT matrix
    static constexpr int STAGE_BYTES = 6 * A_PANEL;         // 2 signals x 3 splits
    static constexpr int W_BYTES = 12 * W_PANEL;            // 3 splits x 4 panels
    static constexpr int TMEM_COLS = 4 * NFFT;              // 2 signals x 2 accumulator buffers
    static constexpr size_t SMEM = 1024 + W_BYTES + NST * STAGE_BYTES + 256;
    static constexpr uint32_t IDESC = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(NFFT >> 3) << 17) | ((uint32_t)(TILE_M >> 4) << 24);
};

struct DftGemmMaps {
    CUtensorMap a[2][3];        // [signal: 0 true, 1 pred][split] : block rows [B*R, HOP] bf16
    CUtensorMap w[3];           // [split] : DFT matrix [n outputs, n samples] bf16
};

template <int NFFT>
__global__ void __launch_bounds__(192, 1) dftgemm_sss_kernel(const __grid_constant__ DftGemmMaps maps, double2* __restrict__ partials,
                                                             float* __restrict__ S_out, int B, int F, int R, int tiles_per_voice,
                                                             int splits, float eps) {
    using P = DftGemm<NFFT>;
    extern __shared__ unsigned char dg_raw[];
    unsigned char* base = reinterpret_cast<unsigned char*>(((uintptr_t)dg_raw + 1023) & ~(uintptr_t)1023);
    unsigned char* w_smem = base;
    unsigned char* a_smem = base + P::W_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(a_smem + P::NST * P::STAGE_BYTES);
    uint64_t* full = bars;                   // [NST]  TMA -> MMA
    uint64_t* empty = bars + P::NST;         // [NST]  MMA -> TMA
    uint64_t* tfull = bars + 2 * P::NST;     // [2]    MMA -> epilogue
    uint64_t* tempty = tfull + 2;            // [2]    epilogue -> MMA
    uint64_t* wfull = tempty + 2;            // [1]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wfull + 1);
    __shared__ double red[2][4];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_tiles = B * tiles_per_voice;

    if (threadIdx.x == 0) {
        for (int i = 0; i < P::NST; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tfull[i], 1); mbar_init(&tempty[i], 128); }
        mbar_init(wfull, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(P::TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            mbar_expect_tx(wfull, (uint32_t)(splits * 4 * P::W_PANEL));
            for (int s = 0; s < splits; ++s)
                for (int q = 0; q < 4; ++q) tma_load_2d(w_smem + (s * 4 + q) * P::W_PANEL, &maps.w[s], q * P::HOP, 0, wfull);
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                const int b = tile / tiles_per_voice, f0 = (tile % tiles_per_voice) * TILE_M;
                const int row0 = b * R + f0;
                for (int q = 0; q < 4; ++q, ++it) {
                    const uint32_t st = it % P::NST, ph = (it / P::NST) & 1u;
                    mbar_wait(&empty[st], ph ^ 1u);
                    mbar_expect_tx(&full[st], (uint32_t)(2 * splits * P::A_PANEL));
                    for (int sig = 0; sig < 2; ++sig)
                        for (int s = 0; s < splits; ++s)
                            tma_load_2d(a_smem + st * P::STAGE_BYTES + (sig * 3 + s) * P::A_PANEL, &maps.a[sig][s], 0, row0 + q, &full[st]);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: one lane issues every tcgen05.mma of the CTA =====
        if (lane == 0) {
            mbar_wait(wfull, 0);
            tc_fence_after();
            // (x split, W split) products in increasing order of magnitude class; splits == 1: the leading product only
            const int ca[6] = {0, 0, 1, 0, 1, 2}, cw[6] = {0, 1, 0, 2, 1, 0};
            const int n_combo = splits == 3 ? 6 : 1;
            uint32_t it = 0, lt = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++lt) {
                const uint32_t buf = lt & 1u;
                mbar_wait(&tempty[buf], ((lt >> 1) & 1u) ^ 1u);
                tc_fence_after();
                for (int q = 0; q < 4; ++q, ++it) {
                    const uint32_t st = it % P::NST, ph = (it / P::NST) & 1u;
                    mbar_wait(&full[st], ph);
                    tc_fence_after();
                    const uint32_t a_base = smem_u32(a_smem + st * P::STAGE_BYTES);
                    const uint32_t w_base = smem_u32(w_smem);
                    for (int sig = 0; sig < 2; ++sig) {
                        const uint32_t d = tmem_base + buf * 2 * NFFT + sig * NFFT;
                        for (int c = 0; c < n_combo; ++c)
                            for (int ks = 0; ks < P::KSTEPS; ++ks) {
                                const uint64_t ad = umma_desc(a_base + (sig * 3 + ca[c]) * P::A_PANEL + ks * 32, P::ROW_BYTES);
                                const uint64_t bd = umma_desc(w_base + (cw[c] * 4 + q) * P::W_PANEL + ks * 32, P::ROW_BYTES);
                                umma_bf16(d, ad, bd, P::IDESC, (q | c | ks) != 0);
                            }
                    }
                    umma_commit(&empty[st]);            // the stage's panels may be overwritten once these MMAs have read them
                }
                umma_commit(&tfull[buf]);               // the tile's accumulators are complete
            }
        }
    } else {
        // ===== epilogue: TMEM -> registers -> power spectra -> the two L1 sums =====
        const int quad = warp & 3;                      // TMEM lanes 32*quad .. 32*quad+31 belong to this warp
        const int row = quad * 32 + lane;
        double acc_lin = 0.0, acc_log = 0.0;
        uint32_t lt = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++lt) {
            const uint32_t buf = lt & 1u;
            const int b = tile / tiles_per_voice, f = (tile % tiles_per_voice) * TILE_M + row;
            const bool valid = f < F;
            mbar_wait(&tfull[buf], (lt >> 1) & 1u);
            tc_fence_after();
            const uint32_t t0 = tmem_base + ((uint32_t)(quad * 32) << 16) + buf * 2 * NFFT;
            float it_lin = 0.f, it_log = 0.f;
            auto bin = [&](float st, float sp, int k) {
                it_lin += fabsf(sp - st);
                it_log += fabsf(__log2f(st + eps) - __log2f(sp + eps));
                if (S_out && valid) {
                    const size_t o = ((size_t)b * F + f) * (NFFT / 2 + 1) + k;
                    S_out[o] = st;
                    S_out[(size_t)B * F * (NFFT / 2 + 1) + o] = sp;
                }
            };
#pragma unroll 1
            for (int k0 = 0; k0 < NFFT / 2; k0 += 16) {
                float tr[16], ti[16], pr[16], pi[16];
                tmem_ld16(t0 + k0, tr);
                tmem_ld16(t0 + NFFT / 2 + k0, ti);
                tmem_ld16(t0 + NFFT + k0, pr);
                tmem_ld16(t0 + NFFT + NFFT / 2 + k0, pi);
                tmem_ld_wait();
                if (valid) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const int k = k0 + i;
                        if (k == 0) {                   // slot n/2 carries the (real) Nyquist bin, bin 0 has no imaginary part
                            bin(tr[0] * tr[0], pr[0] * pr[0], 0);
                            bin(ti[0] * ti[0], pi[0] * pi[0], NFFT / 2);
                        } else {
                            bin(tr[i] * tr[i] + ti[i] * ti[i], pr[i] * pr[i] + pi[i] * pi[i], k);
                        }
                    }
                }
            }
            acc_lin += (double)it_lin;
            acc_log += (double)it_log;
            tc_fence_before();
            mbar_arrive(&tempty[buf]);
        }
        for (int o = 16; o > 0; o >>= 1) {
            acc_lin += __shfl_xor_sync(0xffffffffu, acc_lin, o);
            acc_log += __shfl_xor_sync(0xffffffffu, acc_log, o);
        }
        if (lane == 0) { red[0][quad] = acc_lin; red[1][quad] = acc_log; }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0)
        partials[blockIdx.x] = make_double2((red[0][0] + red[0][1]) + (red[0][2] + red[0][3]), (red[1][0] + red[1][1]) + (red[1][2] + red[1][3]));
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(P::TMEM_COLS) : "memory");
    }
}

// ---- pre-pass: reflect pad (torchaudio Spectrogram center=True, pad_mode="reflect") + bf16 x 3 split, block-row layout ----
// planes[sig][split][b][R*HOP] bf16, sample i of a voice's row space = signal position i - n/2; zero past the last frame
__global__ void dftgemm_split_kernel(const float* __restrict__ xt, const float* __restrict__ xp, __nv_bfloat16* __restrict__ planes,
                                     int B, int N, int n, long long row_elems) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;
    if (i >= row_elems) return;
    long long p = i - n / 2;
    const bool inside = p < (long long)N + n / 2;
    if (p < 0) p = -p;
    if (p >= N) p = 2LL * (N - 1) - p;
    const size_t plane = (size_t)B * row_elems;
    for (int sig = 0; sig < 2; ++sig) {
        float v = inside ? (sig == 0 ? xt : xp)[(size_t)b * N + p] : 0.f;
        for (int s = 0; s < 3; ++s) {
            const __nv_bfloat16 h = __float2bfloat16_rn(v);
            planes[(size_t)(sig * 3 + s) * plane + (size_t)b * row_elems + i] = h;
            v -= __bfloat162float(h);                       // exact: the remainder fits fp32
        }
    }
}

// W[s][j][m], j < n/2: w[m] cos(2 pi j m / n);  j >= n/2: w[m] sin(2 pi (j - n/2) m / n), except j = n/2: w[m] cos(pi m)
__global__ void dftgemm_matrix_kernel(const float* __restrict__ window, __nv_bfloat16* __restrict__ W, int n) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * n) return;
    const int j = idx / n, m = idx % n;
    const double w = window ? (double)window[m] : 0.5 - 0.5 * cospi(2.0 * (double)m / (double)n);
    double sn, cs;
    const int k = j < n / 2 ? j : j - n / 2;
    sincospi(2.0 * (double)(((long long)k * m) % n) / (double)n, &sn, &cs);
    double v = w * (j < n / 2 ? cs : sn);
    if (j == n / 2) v = w * ((m & 1) ? -1.0 : 1.0);
    float r = (float)v;                                      // the reference multiplies fp32 window by fp32 samples; the matrix
    for (int s = 0; s < 3; ++s) {                            // carries 24 bits of the product of window and twiddle
        const __nv_bfloat16 h = __float2bfloat16_rn(r);
        W[(size_t)s * n * n + idx] = h;
        r -= __bfloat162float(h);
    }
}

__global__ void dftgemm_finish_kernel(const double2* __restrict__ partials, int count, double* __restrict__ terms2) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        double l = 0.0, q = 0.0;
        for (int i = 0; i < count; ++i) { l += partials[i].x; q += partials[i].y; }
        terms2[0] = l;
        terms2[1] = q;
    }
}

struct DftGemmGeom {
    int F, tiles_per_voice, R, hop;
    long long row_elems;
    size_t off_planes, off_w, off_part, total;
};

static DftGemmGeom dftgemm_geom(int B, int N, int n) {
    DftGemmGeom g;
    g.hop = n / 4;
    g.F = 1 + N / g.hop;
    g.tiles_per_voice = (g.F + TILE_M - 1) / TILE_M;
    g.R = g.tiles_per_voice * TILE_M + 8;                    // block rows per voice (a tile reads rows f0 .. f0+127+3)
    g.row_elems = (long long)g.R * g.hop;
    size_t off = 0;
    g.off_planes = off; off += align_up(sizeof(__nv_bfloat16) * 6 * (size_t)B * g.row_elems, 1024);
    g.off_w = off;      off += align_up(sizeof(__nv_bfloat16) * 3 * (size_t)n * n, 1024);
    g.off_part = off;   off += align_up(sizeof(double2) * 1024, 1024);
    g.total = off;
    return g;
}

template <int NFFT>
static int launch_dftgemm(const DftGemmGeom& g, char* ws, double* terms2, float* S_out, int splits, float eps, int B, cudaStream_t st) {
    using P = DftGemm<NFFT>;
    DftGemmMaps maps;
    __nv_bfloat16* planes = reinterpret_cast<__nv_bfloat16*>(ws + g.off_planes);
    __nv_bfloat16* W = reinterpret_cast<__nv_bfloat16*>(ws + g.off_w);
    const size_t plane = (size_t)B * g.row_elems;
    for (int sig = 0; sig < 2; ++sig)
        for (int s = 0; s < 3; ++s)
            if (int rc = make_map_2d(&maps.a[sig][s], planes + (size_t)(sig * 3 + s) * plane, (uint64_t)B * g.R, P::HOP, TILE_M, P::HOP)) return rc;
    for (int s = 0; s < 3; ++s)
        if (int rc = make_map_2d(&maps.w[s], W + (size_t)s * NFFT * NFFT, NFFT, NFFT, NFFT, P::HOP)) return rc;
    cudaError_t e = cudaFuncSetAttribute(dftgemm_sss_kernel<NFFT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P::SMEM);
    DDSP_REQUIRE(e == cudaSuccess, "dftgemm_sss_kernel: cannot reserve %zu B of shared memory: %s", (size_t)P::SMEM, cudaGetErrorString(e));
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) sms = 148;
    const int n_tiles = B * g.tiles_per_voice;
    const int grid = n_tiles < sms ? n_tiles : sms;
    double2* partials = reinterpret_cast<double2*>(ws + g.off_part);
    dftgemm_sss_kernel<NFFT><<<grid, 192, P::SMEM, st>>>(maps, partials, S_out, B, g.F, g.R, g.tiles_per_voice, splits, eps);
    DDSP_CHECK_LAUNCH("dftgemm_sss_kernel");
    dftgemm_finish_kernel<<<1, 32, 0, st>>>(partials, grid, terms2);
    DDSP_CHECK_LAUNCH("dftgemm_finish_kernel");
    return 0;
}

}  // namespace ddsp

using namespace ddsp;

extern "C" size_t ddsp_dftgemm_workspace_bytes(int B, int N, int n_fft) {
    if (B < 1 || N < 2 || (n_fft != 64 && n_fft != 128) || n_fft / 2 >= N) return 0;
    return dftgemm_geom(B, N, n_fft).total;
}

extern "C" int ddsp_dftgemm_prepare(const float* x_pred, const float* x_true, const float* window, void* workspace, int n_fft,
                                    int B, int N, void* stream) {
    DDSP_REQUIRE(x_pred && x_true && workspace, "ddsp_dftgemm_prepare: null pointer argument");
    DDSP_REQUIRE(B >= 1 && B <= 65535 && (n_fft == 64 || n_fft == 128) && n_fft / 2 < N, "ddsp_dftgemm_prepare: n_fft must be 64 or 128, n_fft/2 < N");
    const DftGemmGeom g = dftgemm_geom(B, N, n_fft);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char* ws = static_cast<char*>(workspace);
    dftgemm_split_kernel<<<dim3((unsigned)((g.row_elems + 255) / 256), (unsigned)B), 256, 0, st>>>(
        x_true, x_pred, reinterpret_cast<__nv_bfloat16*>(ws + g.off_planes), B, N, n_fft, g.row_elems);
    DDSP_CHECK_LAUNCH("dftgemm_split_kernel");
    dftgemm_matrix_kernel<<<(n_fft * n_fft + 255) / 256, 256, 0, st>>>(window, reinterpret_cast<__nv_bfloat16*>(ws + g.off_w), n_fft);
    DDSP_CHECK_LAUNCH("dftgemm_matrix_kernel");
    return 0;
}

extern "C" int ddsp_dftgemm_sss_forward(void* workspace, double* terms2, float* S_out, int n_fft, int splits, float eps, int B, int N,
                                        void* stream) {
    DDSP_REQUIRE(workspace && terms2, "ddsp_dftgemm_sss_forward: null pointer argument");
    DDSP_REQUIRE(B >= 1 && B <= 65535 && (n_fft == 64 || n_fft == 128) && n_fft / 2 < N, "ddsp_dftgemm_sss_forward: n_fft must be 64 or 128, n_fft/2 < N");
    DDSP_REQUIRE(splits == 1 || splits == 3, "ddsp_dftgemm_sss_forward: splits must be 1 or 3");
    DDSP_REQUIRE(eps >= 1.17549435e-38f, "ddsp_dftgemm_sss_forward: eps must be a normal positive number");
    const DftGemmGeom g = dftgemm_geom(B, N, n_fft);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char* ws = static_cast<char*>(workspace);
    if (n_fft == 64) return launch_dftgemm<64>(g, ws, terms2, S_out, splits, eps, B, st);
    return launch_dftgemm<128>(g, ws, terms2, S_out, splits, eps, B, st);
}
