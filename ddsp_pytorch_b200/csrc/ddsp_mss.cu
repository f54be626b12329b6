// Multi-scale spectral loss for sm_100a -- replaces SSSLoss / MSSLoss (mss_loss.py:11-68).
//
// The reference runs, per scale and per signal, torchaudio Spectrogram (reflect pad + torch.stft through
// cuFFT + abs().pow(2)) and then ~8 elementwise / reduction passes over a 2-4x oversampled spectrogram,
// and autograd replays all of it backwards.  Here one kernel per scale does everything for a frame while
// it sits in shared memory:
//   reflect-pad gather + periodic Hann -> one complex FFT per PAIR of consecutive frames of the same signal
//   (z = w*frame_a + i w*frame_b, split by Hermitian symmetry; frames of one signal have like magnitudes, so
//   the packing costs sqrt(2) in rounding error instead of coupling pred to true) -> power spectra -> both L1
//   terms -> d loss / d S_pred -> ONE inverse FFT for the pair (Hermitian-extended gradient spectra as real /
//   imaginary part) -> window -> overlap-add -> d loss / d x_pred with the adjoint of the reflect pad.
// The spectrograms never reach HBM; the gradient is produced in the forward pass (the loss is a scalar, so
// backward is one gather + multiply by the upstream scalar).
//
// Bit-reproducible by construction (no floating-point atomics anywhere): the overlap-add is OWNER-COMPUTES.  With
// hop = n/4 a frame pair spans ten slices of n/8 samples; a pair owns ("home") the first four and spills the other six
// to the next two pairs.  Spills travel through shared memory between the pairs of one CTA item and through a side
// buffer across items; every home value is summed as (own + previous pair) + pair before that, wherever the spills
// came from, and written once, per scale, in a coordinate that still includes the reflect-padded margins.  The backward
// kernel gathers, for every sample, the scales in order and the (at most two) mirrored margin positions, and multiplies
// by the upstream scalar.  The two loss sums are per-CTA partials reduced in a fixed order.  Nothing depends on the
// batch size: a voice's gradient in a batch of B is bit for bit its gradient alone times the exact ratio of the means.
#include "ddsp_common.cuh"
#include "ddsp_fft.cuh"
#include "ddsp_fft8.cuh"

namespace ddsp {

constexpr int MSS_THREADS = 256;
constexpr int MSS_ITERS = 4;       // frame batches per CTA (amortises the twiddle / window tables)

__device__ __forceinline__ int reflect_index(int p, int N) {
    if (p < 0) p = -p;
    if (p >= N) p = 2 * (N - 1) - p;
    return p;
}

struct MssGeom {
    int gs;       // threads per frame PAIR (one complex transform carries two frames of the same signal)
    int fc;       // frame pairs processed concurrently by a CTA
};

__host__ __device__ inline MssGeom mss_geom(int n) {
    MssGeom g;
    g.gs = n / 4 < MSS_THREADS ? (n / 4 < 1 ? 1 : n / 4) : MSS_THREADS;
    g.fc = MSS_THREADS / g.gs;
    return g;
}

// per group: two double2[n] buffers (the fp32 transforms alias them) + S_true of both frames
__host__ __device__ inline size_t mss_group_floats(int n) { return 8 * (size_t)n + (size_t)n + 4; }

__host__ __device__ inline size_t mss_smem_bytes(int n) {
    const MssGeom g = mss_geom(n);
    return sizeof(float2) * n /*tw*/ + sizeof(float) * n /*window*/ + sizeof(float) * mss_group_floats(n) * g.fc;
}

// split of a packed transform Z = FFT(xa + i xb):  Xa_k = (Z_k + conj Z_{n-k})/2,  Xb_k = (Z_k - conj Z_{n-k})/(2i)
__device__ __forceinline__ void split_pair(float2 zk, float2 zn, float2& xa, float2& xb) {
    xa = make_float2(0.5f * (zk.x + zn.x), 0.5f * (zk.y - zn.y));
    xb = make_float2(0.5f * (zk.y + zn.y), -0.5f * (zk.x - zn.x));
}

__global__ void __launch_bounds__(MSS_THREADS) mss_scale_kernel(const float* __restrict__ xp, const float* __restrict__ xt,
                                                                double2* __restrict__ partials,
                                                                float* __restrict__ frames_out, const float* __restrict__ window,
                                                                int N, int n, int hop, int F, double inv_count, double alpha_d,
                                                                float eps) {
    extern __shared__ __align__(16) unsigned char mss_raw[];
    float2* tw = reinterpret_cast<float2*>(mss_raw);
    float* win = reinterpret_cast<float*>(tw + n);
    float* groups = win + n;
    __shared__ double red[2][MSS_THREADS / 32];

    const MssGeom geo = mss_geom(n);
    const int tid = threadIdx.x;
    const int grp = tid / geo.gs, gtid = tid % geo.gs;
    const int b = blockIdx.y;
    double2* Ad = reinterpret_cast<double2*>(groups + (size_t)grp * mss_group_floats(n));
    double2* Bd = Ad + n;
    float2* A = reinterpret_cast<float2*>(Ad);         // fp32 ping-pong pair inside the first fp64 buffer
    float2* Bf = A + n;
    float* St = reinterpret_cast<float*>(Bd + n);      // [2][half+1] power spectra of the two true frames
    const float* xpb = xp + (size_t)b * N;
    const float* xtb = xt + (size_t)b * N;
    float* gxb = frames_out ? frames_out + (size_t)b * F * n : nullptr;    // [F][n] windowed gradient frames of this voice
    const int half = n / 2;

    fft_fill_twiddles(tw, n);
    __syncthreads();
    // the reference's window is the fp32 buffer of its Spectrogram module (mss_loss.py:23); use the caller's copy of
    // it when given so both sides window with the same values, else the periodic Hann from the fp64-rounded table
    for (int m = tid; m < n; m += MSS_THREADS) win[m] = window ? window[m] : 0.5f - 0.5f * tw[m].x;
    __syncthreads();

    const float w_lin = (float)inv_count, w_log = (float)(alpha_d * inv_count);
    const float inv_ln2 = 1.4426950408889634f;
    double acc_lin = 0.0, acc_log = 0.0;
    for (int it = 0; it < MSS_ITERS; ++it) {
        const int fa = 2 * ((blockIdx.x * MSS_ITERS + it) * geo.fc + grp), fb = fa + 1;
        const bool live_a = fa < F, live_b = fb < F;   // dead groups still walk through every barrier
        const int pa = fa * hop - half, pb = fb * hop - half;

        // ---- true frames: power spectra only -----------------------------------------------------
        for (int m = gtid; m < n; m += geo.gs) {
            const float w = win[m];
            A[m] = make_float2(live_a ? w * xtb[reflect_index(pa + m, N)] : 0.f, live_b ? w * xtb[reflect_index(pb + m, N)] : 0.f);
        }
        __syncthreads();
        {
            const float2* Z = block_fft<false>(A, Bf, tw, n, gtid, geo.gs);
            for (int k = gtid; k <= half; k += geo.gs) {
                float2 xa, xb;
                split_pair(Z[k], Z[(n - k) & (n - 1)], xa, xb);
                St[k] = xa.x * xa.x + xa.y * xa.y;
                St[half + 1 + k] = xb.x * xb.x + xb.y * xb.y;
            }
        }
        __syncthreads();
        // ---- predicted frames in fp64: their small bins decide d log2(S+eps)/dx -----------------------
        for (int m = gtid; m < n; m += geo.gs) {
            const double w = (double)win[m];
            Ad[m] = make_double2(live_a ? w * (double)xpb[reflect_index(pa + m, N)] : 0.0,
                                 live_b ? w * (double)xpb[reflect_index(pb + m, N)] : 0.0);
        }
        __syncthreads();
        const double2* Z = block_fft64(Ad, Bd, n, gtid, geo.gs);
        float2* Y = reinterpret_cast<float2*>(Z == Ad ? Bd : Ad);      // the free fp64 buffer holds the fp32 inverse pair
        float2* Y2 = Y + n;
        for (int k = gtid; k <= half; k += geo.gs) {
            const double2 zk = Z[k], zn = Z[(n - k) & (n - 1)];
            const double xr[2] = {0.5 * (zk.x + zn.x), 0.5 * (zk.y + zn.y)};
            const double xi[2] = {0.5 * (zk.y - zn.y), -0.5 * (zk.x - zn.x)};
            float2 G[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                G[e] = make_float2(0.f, 0.f);
                if (e == 0 ? live_a : live_b) {
                    const float sp = (float)(xr[e] * xr[e] + xi[e] * xi[e]), st = St[e * (half + 1) + k];   // power=2 spectrogram
                    const float dl = sp - st;
                    const float dq = log2f(sp + eps) - log2f(st + eps);
                    acc_lin += (double)fabsf(dl);
                    acc_log += (double)fabsf(dq);
                    const float sl = (dl > 0.f) - (dl < 0.f), sq = (dq > 0.f) - (dq < 0.f);
                    const float gs_ = w_lin * sl + w_log * sq * inv_ln2 / (sp + eps);                      // d loss / d S_pred
                    G[e] = make_float2(2.f * (float)xr[e] * gs_, 2.f * (float)xi[e] * gs_);                 // d loss / d X_pred
                }
            }
            if (gxb) {
                // x-gradient of a frame = w * Re sum_{k<=n/2} G_k e^{+i 2 pi jk/n}: Hermitian-extend G (halved off the
                // edges) so the inverse transform is real, and carry the two frames as real / imaginary part
                if (k == 0 || k == half) {
                    Y[k] = make_float2(G[0].x, G[1].x);
                } else {
                    Y[k] = make_float2(0.5f * (G[0].x - G[1].y), 0.5f * (G[0].y + G[1].x));
                    Y[n - k] = make_float2(0.5f * (G[0].x + G[1].y), 0.5f * (G[1].x - G[0].y));
                }
            }
        }
        __syncthreads();
        if (gxb) {
            const float2* R = block_fft<true>(Y, Y2, tw, n, gtid, geo.gs);
            for (int m = gtid; m < n; m += geo.gs) {
                const float2 r = R[m];
                const float w = win[m];
                if (live_a) gxb[(size_t)fa * n + m] = r.x * w;          // overlap-added by the gather in mss_combine_kernel
                if (live_b) gxb[(size_t)fb * n + m] = r.y * w;
            }
            __syncthreads();
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        acc_lin += __shfl_xor_sync(0xffffffffu, acc_lin, o);
        acc_log += __shfl_xor_sync(0xffffffffu, acc_log, o);
    }
    if ((tid & 31) == 0) { red[0][tid >> 5] = acc_lin; red[1][tid >> 5] = acc_log; }
    __syncthreads();
    if (tid == 0) {
        double l = 0.0, q = 0.0;
        for (int w = 0; w < MSS_THREADS / 32; ++w) { l += red[0][w]; q += red[1][w]; }
        partials[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = make_double2(l, q);
    }
}

// ================================================================================================
// Fast path (8 <= n_fft <= 4096, hop = n_fft/4 -- the reference's overlap of 0.75): register-tiled FFTs (ddsp_fft8.cuh).
// A group of T8 = n/8 threads owns one PAIR of consecutive frames of a voice.  With hop = 2*T8 the pair spans ten slices
// of T8 samples starting at the first frame: thread t holds sample t of every slice (frame a = slices 0..7, frame b =
// slices 2..9, so the two frames share six of their eight loads), bins k = t + q*T8 (q = 0..3, plus n/2 on thread 0) of
// the spectra, and sample t of every slice of the gradient.  The window values a thread needs are the same for every
// frame, so they live in registers.
// ================================================================================================
template <int N>
struct MssFast {
    static constexpr int T8 = N / 8;                              // threads per frame pair
    static constexpr int THREADS = T8 > 128 ? T8 : 128;
    static constexpr int G = THREADS / T8;                        // frame pairs in flight per CTA (one work item)
    static constexpr int SL = G >= 2 ? 2 : 1;                     // pairs per item whose spills leave the CTA
    static constexpr size_t group_bytes = sizeof(double2) * fft8_smem_elems(N);
    static constexpr size_t smem = group_bytes * G;
};

template <int N>
__global__ void __launch_bounds__(MssFast<N>::THREADS, 512 / MssFast<N>::THREADS) mss_fast_kernel(const float* __restrict__ xp, const float* __restrict__ xt,
                                                                      double2* __restrict__ partials,
                                                                      float* __restrict__ gs, float* __restrict__ side,
                                                                      const float* __restrict__ window,
                                                                      int Nsig, int F, double inv_count, double alpha_d,
                                                                      float eps, int mufu, int items_per_voice, long long n_items) {
    using P = MssFast<N>;
    constexpr int T8 = P::T8, half = N / 2, hop = N / 4;
    extern __shared__ __align__(16) unsigned char mss_raw[];
    __shared__ double red[2][P::THREADS / 32];
    const int tid = threadIdx.x, grp = tid / T8, t = tid % T8;
    double2* smd = reinterpret_cast<double2*>(mss_raw + (size_t)grp * P::group_bytes);
    float2* smf = reinterpret_cast<float2*>(smd);

    // the reference's window is the fp32 buffer of its Spectrogram module (mss_loss.py:23); use the caller's copy when
    // given so both sides window with the same values, else the periodic Hann
    float win[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const int m = t + q * T8;
        win[q] = window ? window[m] : 0.5f - 0.5f * (float)cospi(2.0 * (double)m / (double)N);
    }
    const float w_lin = (float)inv_count, w_log = (float)(alpha_d * inv_count);
    const float inv_ln2 = 1.4426950408889634f;
    double acc_lin = 0.0, acc_log = 0.0;

    // d loss / d S_pred of one (bin, frame): both L1 terms accumulate, returns the factor that multiplies 2 X_pred.
    // mufu (eps is a normal positive number): sp + eps and st + eps are normal, so the two log2 are single MUFU.LG2
    // (absolute error 2^-22 on a difference whose sign alone enters the gradient) and the quotient a MUFU.RCP + multiply
    // (2 ulp) -- the library log2f / IEEE division were 23 % of the kernel's instructions (profiles/summary_r01g.md).
    // eps = 0 or denormal takes the library functions.
    const float c_log = w_log * inv_ln2;
    float it_lin = 0.f, it_log = 0.f;              // the <= 10 terms of one frame pair; the long sums are fp64
    auto bin_grad = [&](float sp, float st) -> float {
        const float dl = sp - st;
        const float dq = mufu ? __log2f(sp + eps) - __log2f(st + eps) : log2f(sp + eps) - log2f(st + eps);
        it_lin += fabsf(dl);
        it_log += fabsf(dq);
        const float sl = dl == 0.f ? 0.f : copysignf(w_lin, dl);
        const float sq = dq == 0.f ? 0.f : copysignf(c_log, dq);
        return sl + (mufu ? __fdividef(sq, sp + eps) : sq / (sp + eps));
    };

    // Work item = (voice, P::G consecutive frame pairs), one pair per thread group; item geometry depends on (n, N) only.
    // The grid is at most one wave of resident CTAs and every CTA strides over the items.
    for (long long item = blockIdx.x; item < n_items; item += gridDim.x) {
        const int b = (int)(item / items_per_voice), pblk = (int)(item % items_per_voice);
        const float* xpb = xp + (size_t)b * Nsig;
        const float* xtb = xt + (size_t)b * Nsig;
        const int pair = pblk * P::G + grp;
        const int fa = 2 * pair, fb = fa + 1;
        const bool live_a = fa < F, live_b = fb < F;   // dead groups still walk through every barrier (and own their slices)
        const int pa = fa * hop - half;
        // pairs that touch neither end of the signal (all but n/(2 hop) on each side) address it directly
        const bool interior = pa >= 0 && pa + hop + N <= Nsig;

        // ---- the ten slices of both signals (the predicted ones wait in registers) ---------------------
        float tr[10], pr[10];
        if (interior) {
            const float* st_ = xtb + pa + t;
            const float* sp_ = xpb + pa + t;
#pragma unroll
            for (int j = 0; j < 10; ++j) { tr[j] = st_[j * T8]; pr[j] = sp_[j * T8]; }
        } else {
#pragma unroll
            for (int j = 0; j < 10; ++j) {
                const bool need = (j < 8 && live_a) || (j >= 2 && live_b);
                const int i = need ? reflect_index(pa + t + j * T8, Nsig) : 0;
                tr[j] = need ? xtb[i] : 0.f;
                pr[j] = need ? xpb[i] : 0.f;
            }
        }
        const float ma = live_a ? 1.f : 0.f, mb = live_b ? 1.f : 0.f;

        // ---- true frames: power spectra of bins k = t + q*T8 (q < 4) and n/2, kept in registers ----------
        float2 v[8];
        float St[2][5];
        {
#pragma unroll
            for (int q = 0; q < 8; ++q) v[q] = make_float2(win[q] * tr[q] * ma, win[q] * tr[q + 2] * mb);
            fft8_group<N, float, false>(v, smf, t);
            fft8_sync<N>();
#pragma unroll
            for (int q = 0; q < 8; ++q) smf[fft8_pad<float>(t + q * T8)] = v[q];
            fft8_sync<N>();
#pragma unroll
            for (int q = 0; q < 5; ++q) {
                const int k = q < 4 ? t + q * T8 : half;
                const float2 zk = q < 4 ? v[q] : smf[fft8_pad<float>(half)];
                const float2 zn = smf[fft8_pad<float>((N - k) & (N - 1))];
                float2 xa, xb;
                split_pair(zk, zn, xa, xb);
                St[0][q] = xa.x * xa.x + xa.y * xa.y;
                St[1][q] = xb.x * xb.x + xb.y * xb.y;
            }
            fft8_sync<N>();
        }

        // ---- predicted frames in fp64: their small bins decide d log2(S+eps)/dx ---------------------------
        double2 vd[8];
#pragma unroll
        for (int q = 0; q < 8; ++q)
            vd[q] = make_double2(live_a ? (double)win[q] * (double)pr[q] : 0.0, live_b ? (double)win[q] * (double)pr[q + 2] : 0.0);
        fft8_group<N, double, false>(vd, smd, t);
        fft8_sync<N>();
#pragma unroll
        for (int q = 0; q < 8; ++q) smd[fft8_pad<double>(t + q * T8)] = vd[q];
        fft8_sync<N>();
        float2 Yk[5], Yn[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int k = t + q * T8;
            const double2 zk = vd[q];
            const double2 zn = smd[fft8_pad<double>((N - k) & (N - 1))];
            const double xr[2] = {0.5 * (zk.x + zn.x), 0.5 * (zk.y + zn.y)};
            const double xi[2] = {0.5 * (zk.y - zn.y), -0.5 * (zk.x - zn.x)};
            float2 G[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                G[e] = make_float2(0.f, 0.f);
                if (e == 0 ? live_a : live_b) {
                    const float gs_ = 2.f * bin_grad((float)(xr[e] * xr[e] + xi[e] * xi[e]), St[e][q]);   // power=2 spectrogram
                    G[e] = make_float2((float)xr[e] * gs_, (float)xi[e] * gs_);                           // d loss / d X_pred
                }
            }
            // x-gradient of a frame = w * Re sum_{k<=n/2} G_k e^{+i 2 pi jk/n}: Hermitian-extend G (halved off the
            // edges) so the inverse transform is real, and carry the two frames as real / imaginary part
            if (k == 0) {
                Yk[q] = make_float2(G[0].x, G[1].x);
                Yn[q] = make_float2(0.f, 0.f);
            } else {
                Yk[q] = make_float2(0.5f * (G[0].x - G[1].y), 0.5f * (G[0].y + G[1].x));
                Yn[q] = make_float2(0.5f * (G[0].x + G[1].y), 0.5f * (G[1].x - G[0].y));
            }
        }
        // bin n/2 (real for both frames): thread 0 of the group, one pass for the two frames' values of Z[n/2]
        Yk[4] = make_float2(0.f, 0.f);
        if (t == 0) {
            const double2 zh = smd[fft8_pad<double>(half)];          // Xa[n/2] = zh.x, Xb[n/2] = zh.y
            float ga = 0.f, gb = 0.f;
            if (live_a) ga = 2.f * (float)zh.x * bin_grad((float)(zh.x * zh.x), St[0][4]);
            if (live_b) gb = 2.f * (float)zh.y * bin_grad((float)(zh.y * zh.y), St[1][4]);
            Yk[4] = make_float2(ga, gb);
        }
        acc_lin += (double)it_lin;
        acc_log += (double)it_log;
        it_lin = it_log = 0.f;
        fft8_sync<N>();                                // everyone has read the fp64 spectra
        if (gs) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int k = t + q * T8;
                smf[fft8_pad<float>(k)] = Yk[q];
                if (k != 0) smf[fft8_pad<float>(N - k)] = Yn[q];
            }
            if (t == 0) smf[fft8_pad<float>(half)] = Yk[4];
            fft8_sync<N>();
#pragma unroll
            for (int q = 0; q < 8; ++q) v[q] = smf[fft8_pad<float>(t + q * T8)];
            fft8_sync<N>();
            fft8_group<N, float, true>(v, smf, t);
            // ---- owner-computes overlap-add ------------------------------------------------------------------
            // o[j]: this pair's windowed gradient on slice j (frame a on slices 0..7, frame b on 2..9); dead frames give 0
            float o[10];
#pragma unroll
            for (int j = 0; j < 10; ++j) {
                float x = j < 8 ? v[j].x * win[j] : 0.f;
                if (j >= 2) x += v[j - 2].y * win[j - 2];
                o[j] = x;
            }
            // slices 4..9 belong to the next two pairs: through the side buffer when they are in another item ...
            float* sd = nullptr;
            if (P::G == 1) sd = side + ((size_t)b * items_per_voice + pblk) * (6 * T8);
            else if (grp >= P::G - 2) sd = side + (((size_t)b * items_per_voice + pblk) * 2 + (grp - (P::G - 2))) * (6 * T8);
            if (sd) {
#pragma unroll
                for (int j = 0; j < 6; ++j) sd[j * T8 + t] = o[4 + j];
            }
            if (P::G >= 2) {
                // ... and through the groups' (now idle) exchange buffers inside the item; summed as (own + previous) + the
                // one before, the order mss_combine_kernel uses for the spills that crossed an item boundary
                fft8_sync<N>();                        // the inverse transform's last exchange has been read
                float* sp_ = reinterpret_cast<float*>(smd);
#pragma unroll
                for (int j = 0; j < 6; ++j) sp_[j * T8 + t] = o[4 + j];
                __syncthreads();
                if (grp >= 1) {
                    const float* pv = reinterpret_cast<const float*>(mss_raw + (size_t)(grp - 1) * P::group_bytes);
#pragma unroll
                    for (int j = 0; j < 4; ++j) o[j] += pv[j * T8 + t];
                }
                if (grp >= 2) {
                    const float* pv = reinterpret_cast<const float*>(mss_raw + (size_t)(grp - 2) * P::group_bytes);
#pragma unroll
                    for (int j = 0; j < 2; ++j) o[j] += pv[(4 + j) * T8 + t];
                }
                __syncthreads();                       // the next item's transforms may overwrite the buffers
            } else {
                fft8_sync<N>();                        // the inverse transform's last exchange is read
            }
            // home slices 0..3, in the margin-extended coordinate u = position + n/2 of this voice and scale
            float* g_ = gs + ((size_t)b * items_per_voice * P::G + pair) * half + t;
#pragma unroll
            for (int j = 0; j < 4; ++j) g_[j * T8] = o[j];
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        acc_lin += __shfl_xor_sync(0xffffffffu, acc_lin, o);
        acc_log += __shfl_xor_sync(0xffffffffu, acc_log, o);
    }
    if ((tid & 31) == 0) { red[0][tid >> 5] = acc_lin; red[1][tid >> 5] = acc_log; }
    __syncthreads();
    if (tid == 0) {
        double l = 0.0, q = 0.0;
        for (int w = 0; w < P::THREADS / 32; ++w) { l += red[0][w]; q += red[1][w]; }
        partials[blockIdx.x] = make_double2(l, q);
    }
}

// ---- plan: per-scale kernel choice, geometry and scratch layout (a function of the arguments and the device only) -------
struct MssScalePlan {
    int n, hop, F;
    int fast;                 // register-tiled kernel (n in 8..4096, hop == n/4); else the generic shared-memory kernel
    int T8, G, SL;            // fast: threads per pair, pairs per item, spilling pairs per item
    int items_per_voice;      // fast: items per voice (pairs incl. the two margin pairs, rounded up to G)
    long long U;              // fast: floats per voice of the margin-extended gradient = items_per_voice*G*n/2
    unsigned grid_x, grid_y;  // CTAs of the scale kernel; grid_x*grid_y loss partials
    size_t off_gs, off_side, off_part;     // byte offsets into the scratch buffer
    double inv_count;
};

template <int N>
static int mss_fast_slots(int* G, int* T8, int* SL) {
    using P = MssFast<N>;
    *G = P::G; *T8 = P::T8; *SL = P::SL;
    static int slots_cache[64] = {0};                  // resident CTAs of this instantiation per device
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    if (slots_cache[dev] == 0) {
        if (P::smem > 48 * 1024) cudaFuncSetAttribute(mss_fast_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P::smem);
        int sms = 148, per_sm = 1;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) sms = 148;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mss_fast_kernel<N>, P::THREADS, P::smem) != cudaSuccess || per_sm < 1)
            per_sm = 1;
        slots_cache[dev] = sms * per_sm;
    }
    return slots_cache[dev];
}

static int mss_plan(const int* n_ffts, const int* hops, int ns, int B, int N, MssScalePlan* pl, size_t* total) {
    size_t off = 0;
    for (int s = 0; s < ns; ++s) {
        MssScalePlan& p = pl[s];
        p.n = n_ffts[s]; p.hop = hops[s];
        DDSP_REQUIRE(p.n >= 4 && p.n <= 4096 && (p.n & (p.n - 1)) == 0, "ddsp_mss: n_fft=%d must be a power of two in [4,4096]", p.n);
        DDSP_REQUIRE(p.hop >= 1, "ddsp_mss: hop=%d must be positive", p.hop);
        DDSP_REQUIRE(p.n / 2 < N, "ddsp_mss: reflect padding n_fft/2=%d needs a signal longer than that (N=%d)", p.n / 2, N);
        p.F = 1 + N / p.hop;
        p.inv_count = 1.0 / (double)ddsp_mss_count(B, N, p.n, p.hop);
        p.fast = p.n >= 8 && p.hop * 4 == p.n;
        p.T8 = p.G = p.SL = p.items_per_voice = 0;
        p.U = 0;
        size_t gs_bytes, side_bytes;
        if (p.fast) {
            int slots = 0;
            switch (p.n) {
#define MSS_CASE(NN) case NN: slots = mss_fast_slots<NN>(&p.G, &p.T8, &p.SL); break
                MSS_CASE(8); MSS_CASE(16); MSS_CASE(32); MSS_CASE(64); MSS_CASE(128); MSS_CASE(256);
                MSS_CASE(512); MSS_CASE(1024); MSS_CASE(2048); MSS_CASE(4096);
#undef MSS_CASE
                default: break;
            }
            const int pairs = (p.F + 4) / 2;           // live pairs + the margin pairs that only collect spills
            p.items_per_voice = (pairs + p.G - 1) / p.G;
            p.U = (long long)p.items_per_voice * p.G * (p.n / 2);
            const long long n_items = (long long)B * p.items_per_voice;
            p.grid_x = (unsigned)(n_items < slots ? n_items : slots);
            p.grid_y = 1;
            gs_bytes = sizeof(float) * (size_t)B * (size_t)p.U;
            side_bytes = sizeof(float) * (size_t)B * p.items_per_voice * p.SL * 6 * p.T8;
        } else {
            const MssGeom g = mss_geom(p.n);
            const int per_cta = 2 * g.fc * MSS_ITERS;
            p.grid_x = (unsigned)((p.F + per_cta - 1) / per_cta);
            p.grid_y = (unsigned)B;
            gs_bytes = sizeof(float) * (size_t)B * p.F * p.n;      // windowed gradient frames [B,F,n]
            side_bytes = 0;
        }
        p.off_gs = off;   off += align_up(gs_bytes, 256);
        p.off_side = off; off += align_up(side_bytes, 256);
        p.off_part = off; off += align_up(sizeof(double2) * (size_t)p.grid_x * p.grid_y, 256);
    }
    *total = off;
    return 0;
}

template <int N>
static int launch_mss_fast(const MssScalePlan& p, const float* xp, const float* xt, double2* partials, float* gs, float* side,
                           const float* window, int B, int Nsig, double alpha_d, float eps, cudaStream_t st) {
    using P = MssFast<N>;
    if (P::smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(mss_fast_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P::smem);
        DDSP_REQUIRE(e == cudaSuccess, "mss_fast_kernel: cannot reserve %zu B of shared memory: %s", (size_t)P::smem, cudaGetErrorString(e));
    }
    const long long n_items = (long long)B * p.items_per_voice;
    mss_fast_kernel<N><<<p.grid_x, P::THREADS, P::smem, st>>>(xp, xt, partials, gs, side, window, Nsig, p.F, p.inv_count, alpha_d, eps,
                                                              eps >= 1.17549435e-38f ? 1 : 0, p.items_per_voice, n_items);
    DDSP_CHECK_LAUNCH("mss_fast_kernel");
    return 0;
}

// ---- fixed-order reduction of the per-CTA loss partials -------------------------------------------------------------
constexpr int MSS_MAX_SCALES = 16;
struct MssFinishArgs {
    int ns;
    const double2* part[MSS_MAX_SCALES];
    int count[MSS_MAX_SCALES];
    double inv_count[MSS_MAX_SCALES];
};

__global__ void __launch_bounds__(32 * MSS_MAX_SCALES) mss_finish_kernel(MssFinishArgs a, double alpha_d, double* __restrict__ terms,
                                                                          float* __restrict__ loss_f32) {
    // one warp per scale: lane-strided sums in a fixed order, xor-shuffle tree (every lane ends with the same bits), then
    // thread 0 adds the scales in order
    __shared__ double per_scale[MSS_MAX_SCALES];
    const int s = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (s < a.ns) {
        double l = 0.0, q = 0.0;
        for (int i = lane; i < a.count[s]; i += 32) { const double2 v = a.part[s][i]; l += v.x; q += v.y; }
        for (int o = 16; o > 0; o >>= 1) {
            l += __shfl_xor_sync(0xffffffffu, l, o);
            q += __shfl_xor_sync(0xffffffffu, q, o);
        }
        if (lane == 0) {
            terms[1 + 2 * s] = l;
            terms[2 + 2 * s] = q;
            per_scale[s] = (l + alpha_d * q) * a.inv_count[s];
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double loss = 0.0;
        for (int k = 0; k < a.ns; ++k) loss += per_scale[k];
        terms[0] = loss;
        if (loss_f32) loss_f32[0] = (float)loss;
    }
}

// ---- backward: owner-computes gather over scales and mirrored margins, times the upstream scalar -----------------------
struct MssCombineScale {
    const float* gs;
    const float* side;
    int fast, n, hop, F, T8, G, SL, items_per_voice;
    int lh, lt8, lg;          // log2 of n/2, T8 and G (all powers of two on the fast path): the gather divides by shifting
    long long U;
};
struct MssCombineArgs {
    int ns;
    MssCombineScale s[MSS_MAX_SCALES];
};

// complete gradient mass at extended coordinate u (position u - n/2) of voice b for a fast-path scale.  Index arithmetic
// by shifts and masks: with run-time divisors the six scales cost ~1200 integer instructions per sample and the kernel
// took 0.2 ms at cfg 5 (profiles/summary_r02.md).
__device__ __forceinline__ float mss_fast_value(const MssCombineScale& d, int b, long long u) {
    float x = d.gs[(size_t)b * d.U + u];
    const int p = (int)(u >> d.lh), r = (int)u & ((1 << d.lh) - 1), j = r >> d.lt8, t = r & (d.T8 - 1), g = p & (d.G - 1);
    if (g >= 2 || p == 0) return x;                     // both spill sources are pairs of this item: already summed in the CTA
    const size_t item_floats = (size_t)d.SL * 6 * d.T8;
    const float* voice_side = d.side + (size_t)b * d.items_per_voice * item_floats;
    auto side_of = [&](int q) {
        const int item = q >> d.lg, slot = d.G >= 2 ? (q & (d.G - 1)) - (d.G - 2) : 0;
        return voice_side + (size_t)item * item_floats + (size_t)slot * (6 * d.T8);
    };
    if (g == 0) x += side_of(p - 1)[(j << d.lt8) + t];
    if (j < 2 && p >= 2) x += side_of(p - 2)[((j + 4) << d.lt8) + t];
    return x;
}

// the same for a generic-path scale: frames in ascending order
__device__ __forceinline__ float mss_generic_value(const MssCombineScale& d, int b, long long q) {   // q: position, may be off the signal
    const int half = d.n >> 1;
    long long f_hi = (q + half) / d.hop;                 // frames f with f*hop - half <= q < f*hop + half
    if (q + half < 0) return 0.f;
    if (f_hi > d.F - 1) f_hi = d.F - 1;
    long long f_lo = q - half < 0 ? 0 : (q - half) / d.hop + 1;
    float x = 0.f;
    for (long long f = f_lo; f <= f_hi; ++f) x += d.gs[((size_t)b * d.F + f) * d.n + (size_t)(q - (f * d.hop - half))];
    return x;
}

// Four consecutive samples per thread when every scale took the fast path (n >= 32, so a group of four aligned samples lies in
// one slice of one pair: one set of index arithmetic and float4 loads for the four).  Groups that can receive mass reflected
// from the padded margins (within n_max/2 of either end) take the scalar route sample by sample.  Every sample's sum is formed
// in the order of mss_combine_kernel, so the two kernels give the same bits.
__global__ void __launch_bounds__(256) mss_combine4_kernel(MssCombineArgs a, const float* __restrict__ gloss, float* __restrict__ out,
                                                           int Nsig, int accumulate, int margin) {
    const int m0 = 4 * (blockIdx.x * blockDim.x + threadIdx.x), b = blockIdx.y;
    if (m0 >= Nsig) return;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (m0 > margin && m0 + 3 < Nsig - 2 - margin) {
        for (int s = 0; s < a.ns; ++s) {
            const MssCombineScale& d = a.s[s];
            const long long u = (long long)m0 + (d.n >> 1);
            float4 x = *reinterpret_cast<const float4*>(d.gs + (size_t)b * d.U + u);
            const int p = (int)(u >> d.lh), r = (int)u & ((1 << d.lh) - 1), j = r >> d.lt8, t = r & (d.T8 - 1), g = p & (d.G - 1);
            if (!(g >= 2 || p == 0)) {
                const size_t item_floats = (size_t)d.SL * 6 * d.T8;
                const float* voice_side = d.side + (size_t)b * d.items_per_voice * item_floats;
                auto side_of = [&](int q) {
                    const int item = q >> d.lg, slot = d.G >= 2 ? (q & (d.G - 1)) - (d.G - 2) : 0;
                    return voice_side + (size_t)item * item_floats + (size_t)slot * (6 * d.T8);
                };
                if (g == 0) {
                    const float4 y = *reinterpret_cast<const float4*>(side_of(p - 1) + (j << d.lt8) + t);
                    x.x += y.x; x.y += y.y; x.z += y.z; x.w += y.w;
                }
                if (j < 2 && p >= 2) {
                    const float4 y = *reinterpret_cast<const float4*>(side_of(p - 2) + ((j + 4) << d.lt8) + t);
                    x.x += y.x; x.y += y.y; x.z += y.z; x.w += y.w;
                }
            }
            acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
        }
    } else {
        float av[4] = {0.f, 0.f, 0.f, 0.f};
        for (int e = 0; e < 4; ++e) {
            const int m = m0 + e;
            if (m >= Nsig) break;
            const long long back = 2LL * (Nsig - 1) - m;
            for (int s = 0; s < a.ns; ++s) {
                const MssCombineScale& d = a.s[s];
                const int half = d.n >> 1;
                float v = mss_fast_value(d, b, (long long)m + half);
                if (m >= 1 && m <= half) v += mss_fast_value(d, b, (long long)half - m);
                if (m <= Nsig - 2 && back + half < d.U) v += mss_fast_value(d, b, back + half);
                av[e] += v;
            }
        }
        acc = make_float4(av[0], av[1], av[2], av[3]);
    }
    const float gl = gloss[0];
    float4 r = make_float4(acc.x * gl, acc.y * gl, acc.z * gl, acc.w * gl);
    float4* o = reinterpret_cast<float4*>(out + (size_t)b * Nsig + m0);
    if (accumulate) {
        const float4 q = *o;
        r.x += q.x; r.y += q.y; r.z += q.z; r.w += q.w;
    }
    *o = r;
}

__global__ void __launch_bounds__(256) mss_combine_kernel(MssCombineArgs a, const float* __restrict__ gloss, float* __restrict__ out,
                                                          int Nsig, int accumulate) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y;
    if (m >= Nsig) return;
    const long long back = 2LL * (Nsig - 1) - m;          // the other position that reflects onto m from beyond the end
    float acc = 0.f;
    for (int s = 0; s < a.ns; ++s) {
        const MssCombineScale& d = a.s[s];
        const int half = d.n >> 1;
        float v;
        if (d.fast) {
            v = mss_fast_value(d, b, (long long)m + half);
            if (m >= 1 && m <= half) v += mss_fast_value(d, b, (long long)half - m);
            if (m <= Nsig - 2 && back + half < d.U) v += mss_fast_value(d, b, back + half);
        } else {
            v = mss_generic_value(d, b, m);
            if (m >= 1 && m <= half) v += mss_generic_value(d, b, -(long long)m);
            if (m <= Nsig - 2) v += mss_generic_value(d, b, back);
        }
        acc += v;
    }
    const size_t i = (size_t)b * Nsig + m;
    const float r = acc * gloss[0];
    out[i] = accumulate ? out[i] + r : r;
}

}  // namespace ddsp

using namespace ddsp;

extern "C" int64_t ddsp_mss_count(int B, int N, int n_fft, int hop) {
    if (B < 1 || N < 1 || n_fft < 2 || hop < 1) return 0;
    return (int64_t)B * (n_fft / 2 + 1) * (1 + N / hop);
}

static int mss_check(const void* a, const void* b, const int* n_ffts_host, const int* hops_host, int n_scales, int B, int N, const char* who) {
    DDSP_REQUIRE(a && b && n_ffts_host && hops_host, "%s: null pointer argument", who);
    DDSP_REQUIRE(B >= 1 && B <= 65535 && N >= 2 && n_scales >= 1 && n_scales <= MSS_MAX_SCALES,
                 "%s: bad B/N/n_scales (1 <= B <= 65535, N >= 2, 1..%d scales)", who, MSS_MAX_SCALES);
    return 0;
}

extern "C" size_t ddsp_mss_scratch_bytes(const int* n_ffts_host, const int* hops_host, int n_scales, int B, int N) {
    if (!n_ffts_host || !hops_host || n_scales < 1 || n_scales > MSS_MAX_SCALES || B < 1 || B > 65535 || N < 2) return 0;
    MssScalePlan pl[MSS_MAX_SCALES];
    size_t total = 0;
    if (mss_plan(n_ffts_host, hops_host, n_scales, B, N, pl, &total) != 0) return 0;
    return total;
}

extern "C" int ddsp_mss_forward(const float* x_pred, const float* x_true, double* loss_terms, float* loss_f32, void* scratch,
                                int need_grad, const int* n_ffts_host, const int* hops_host, const float* const* windows_host,
                                int n_scales, float alpha, float eps, int B, int N, void* stream) {
    if (int rc = mss_check(x_pred, x_true, n_ffts_host, hops_host, n_scales, B, N, "ddsp_mss_forward")) return rc;
    DDSP_REQUIRE(loss_terms && scratch, "ddsp_mss_forward: null loss_terms / scratch");
    MssScalePlan pl[MSS_MAX_SCALES];
    size_t total = 0;
    if (int rc = mss_plan(n_ffts_host, hops_host, n_scales, B, N, pl, &total)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    cudaError_t e = fft_ensure_table(st);
    if (e == cudaSuccess) e = fft8_ensure_tables(st);
    DDSP_REQUIRE(e == cudaSuccess, "ddsp_mss_forward: twiddle table initialisation failed: %s", cudaGetErrorString(e));
    char* base = static_cast<char*>(scratch);
    for (int s = 0; s < n_scales; ++s) {
        const MssScalePlan& p = pl[s];
        const float* wptr = windows_host ? windows_host[s] : nullptr;
        double2* part = reinterpret_cast<double2*>(base + p.off_part);
        float* gs = need_grad ? reinterpret_cast<float*>(base + p.off_gs) : nullptr;
        float* side = reinterpret_cast<float*>(base + p.off_side);
        if (p.fast) {
            int rc = -1;
            switch (p.n) {
#define MSS_CASE(NN) case NN: rc = launch_mss_fast<NN>(p, x_pred, x_true, part, gs, side, wptr, B, N, (double)alpha, eps, st); break
                MSS_CASE(8); MSS_CASE(16); MSS_CASE(32); MSS_CASE(64); MSS_CASE(128); MSS_CASE(256);
                MSS_CASE(512); MSS_CASE(1024); MSS_CASE(2048); MSS_CASE(4096);
#undef MSS_CASE
                default: break;
            }
            if (rc != 0) { if (rc < 0) set_error("ddsp_mss_forward: no kernel for n_fft=%d", p.n); return rc < 0 ? 1 : rc; }
            continue;
        }
        // generic shared-memory kernel: n = 4, or a hop other than n/4 (overlap != 0.75)
        const size_t smem = mss_smem_bytes(p.n);
        if (smem > 48 * 1024) {
            e = cudaFuncSetAttribute(mss_scale_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            DDSP_REQUIRE(e == cudaSuccess, "mss_scale_kernel: cannot reserve %zu B of shared memory: %s", smem, cudaGetErrorString(e));
        }
        mss_scale_kernel<<<dim3(p.grid_x, p.grid_y), MSS_THREADS, smem, st>>>(x_pred, x_true, part, gs, wptr, N, p.n, p.hop, p.F,
                                                                             p.inv_count, (double)alpha, eps);
        DDSP_CHECK_LAUNCH("mss_scale_kernel");
    }
    MssFinishArgs fa;
    fa.ns = n_scales;
    for (int s = 0; s < n_scales; ++s) {
        fa.part[s] = reinterpret_cast<const double2*>(base + pl[s].off_part);
        fa.count[s] = (int)(pl[s].grid_x * pl[s].grid_y);
        fa.inv_count[s] = pl[s].inv_count;
    }
    mss_finish_kernel<<<1, 32 * n_scales, 0, st>>>(fa, (double)alpha, loss_terms, loss_f32);
    DDSP_CHECK_LAUNCH("mss_finish_kernel");
    return 0;
}

extern "C" int ddsp_mss_backward(const float* gloss, const void* scratch, float* gx_out, const int* n_ffts_host,
                                 const int* hops_host, int n_scales, int B, int N, void* stream) {
    if (int rc = mss_check(gloss, scratch, n_ffts_host, hops_host, n_scales, B, N, "ddsp_mss_backward")) return rc;
    DDSP_REQUIRE(gx_out, "ddsp_mss_backward: null output");
    MssScalePlan pl[MSS_MAX_SCALES];
    size_t total = 0;
    if (int rc = mss_plan(n_ffts_host, hops_host, n_scales, B, N, pl, &total)) return rc;
    const char* base = static_cast<const char*>(scratch);
    MssCombineArgs ca;
    ca.ns = n_scales;
    for (int s = 0; s < n_scales; ++s) {
        const MssScalePlan& p = pl[s];
        MssCombineScale& d = ca.s[s];
        d.gs = reinterpret_cast<const float*>(base + p.off_gs);
        d.side = reinterpret_cast<const float*>(base + p.off_side);
        d.fast = p.fast; d.n = p.n; d.hop = p.hop; d.F = p.F; d.T8 = p.T8 > 0 ? p.T8 : 1; d.G = p.G > 0 ? p.G : 1; d.SL = p.SL;
        d.items_per_voice = p.items_per_voice; d.U = p.U;
        auto lg2 = [](int v) { int k = 0; while ((1 << k) < v) ++k; return k; };
        d.lh = lg2(p.n / 2); d.lt8 = lg2(d.T8); d.lg = lg2(d.G);
    }
    bool vec4 = N % 4 == 0 && (reinterpret_cast<uintptr_t>(gx_out) & 15) == 0;
    int margin = 0;
    for (int s = 0; s < n_scales; ++s) {
        vec4 = vec4 && ca.s[s].fast && ca.s[s].T8 >= 4 && ca.s[s].U % 4 == 0;
        margin = ca.s[s].n / 2 > margin ? ca.s[s].n / 2 : margin;
    }
    if (vec4) {
        mss_combine4_kernel<<<dim3((unsigned)((N / 4 + 255) / 256), (unsigned)B), 256, 0, static_cast<cudaStream_t>(stream)>>>(ca, gloss, gx_out, N, 0, margin);
        DDSP_CHECK_LAUNCH("mss_combine4_kernel");
        return 0;
    }
    mss_combine_kernel<<<dim3((unsigned)((N + 255) / 256), (unsigned)B), 256, 0, static_cast<cudaStream_t>(stream)>>>(ca, gloss, gx_out, N, 0);
    DDSP_CHECK_LAUNCH("mss_combine_kernel");
    return 0;
}
