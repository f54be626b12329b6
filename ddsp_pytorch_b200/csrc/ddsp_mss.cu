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
//   imaginary part) -> window -> overlap-add into d loss / d x_pred with the adjoint of the reflect pad.
// The spectrograms never reach HBM; the gradient is produced in the forward pass (the loss is a scalar, so
// backward is one multiply by the upstream scalar).
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
                                                                double* __restrict__ loss, double* __restrict__ terms,
                                                                float* __restrict__ gx, const float* __restrict__ window,
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
    float* gxb = gx ? gx + (size_t)b * N : nullptr;
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
                if (live_a) atomicAdd(&gxb[reflect_index(pa + m, N)], r.x * w);
                if (live_b) atomicAdd(&gxb[reflect_index(pb + m, N)], r.y * w);
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
        atomicAdd(&terms[0], l);
        atomicAdd(&terms[1], q);
        atomicAdd(loss, (l + alpha_d * q) * inv_count);
    }
}

// ================================================================================================
// Fast path (n_fft >= 8): register-tiled FFTs (ddsp_fft8.cuh).  A group of n/8 threads owns one PAIR of
// consecutive frames of a voice; thread t holds elements m = t + q*n/8 (q = 0..7) of the windowed pair, bins
// k = t + q*n/8 (q = 0..3, plus n/2 on thread 0) of its spectra, and the same elements of the gradient frame.
// The window values a thread needs are the same for every frame, so they live in registers.
// ================================================================================================
template <int N>
struct MssFast {
    static constexpr int T8 = N / 8;                              // threads per frame pair
    static constexpr int THREADS = T8 > 128 ? T8 : 128;
    static constexpr int G = THREADS / T8;                        // frame pairs in flight per CTA
    static constexpr size_t group_bytes = sizeof(double2) * fft8_smem_elems(N);
    static constexpr size_t smem = group_bytes * G;
};

template <int N>
__global__ void __launch_bounds__(MssFast<N>::THREADS, 512 / MssFast<N>::THREADS) mss_fast_kernel(const float* __restrict__ xp, const float* __restrict__ xt,
                                                                      double* __restrict__ loss, double* __restrict__ terms,
                                                                      float* __restrict__ gx, const float* __restrict__ window,
                                                                      int Nsig, int hop, int F, double inv_count, double alpha_d,
                                                                      float eps, int items_per_voice, long long n_items) {
    using P = MssFast<N>;
    constexpr int T8 = P::T8, half = N / 2;
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
    // sp + eps and st + eps are normal positive numbers (eps > 0 is checked on the host), so the two log2 are single
    // MUFU.LG2 (absolute error 2^-22 on a difference whose sign alone enters the gradient) and the quotient a
    // MUFU.RCP + multiply (2 ulp) -- the library log2f / IEEE division were 23 % of the kernel's instructions
    // (profiles/summary_r01g.md).
    const float c_log = w_log * inv_ln2;
    float it_lin = 0.f, it_log = 0.f;              // the <= 10 terms of one frame pair; the long sums are fp64
    auto bin_grad = [&](float sp, float st) -> float {
        const float dl = sp - st;
        const float dq = __log2f(sp + eps) - __log2f(st + eps);
        it_lin += fabsf(dl);
        it_log += fabsf(dq);
        const float sl = dl == 0.f ? 0.f : copysignf(w_lin, dl);
        const float sq = dq == 0.f ? 0.f : copysignf(c_log, dq);
        return sl + __fdividef(sq, sp + eps);
    };

    // Work item = (voice, P::G consecutive frame pairs), one pair per thread group.  The grid is at most one wave of resident
    // CTAs and every CTA strides over the items: with a fixed 2-4 items per CTA the small shapes ran 1.2-2.4 waves, the last
    // one nearly empty (cfg 1': 704 CTAs on 592 slots, 40-50 us per scale for 25 us of work).
    for (long long item = blockIdx.x; item < n_items; item += gridDim.x) {
        const int b = (int)(item / items_per_voice), pblk = (int)(item % items_per_voice);
        const float* xpb = xp + (size_t)b * Nsig;
        const float* xtb = xt + (size_t)b * Nsig;
        float* gxb = gx ? gx + (size_t)b * Nsig : nullptr;
        const int fa = 2 * (pblk * P::G + grp), fb = fa + 1;
        const bool live_a = fa < F, live_b = fb < F;   // dead groups still walk through every barrier
        const int pa = fa * hop - half, pb = fb * hop - half;
        // frames that touch neither end of the signal (all but n/(2 hop) on each side) address it directly
        const bool interior = pa >= 0 && pb + N <= Nsig;

        // ---- gather both signals' samples of the two frames (the predicted ones wait in registers) ------------
        float2 v[8];
        float pra[8], prb[8];
        if (interior) {
            const float* st_ = xtb + pa + t;
            const float* sp_ = xpb + pa + t;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                v[q] = make_float2(st_[q * T8], st_[hop + q * T8]);
                pra[q] = sp_[q * T8];
                prb[q] = sp_[hop + q * T8];
            }
        } else {
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int ia = reflect_index(pa + t + q * T8, Nsig), ib = reflect_index(pb + t + q * T8, Nsig);
                v[q] = make_float2(live_a ? xtb[ia] : 0.f, live_b ? xtb[ib] : 0.f);
                pra[q] = live_a ? xpb[ia] : 0.f;
                prb[q] = live_b ? xpb[ib] : 0.f;
            }
        }

        // ---- true frames: power spectra of bins k = t + q*T8 (q < 4) and n/2, kept in registers ----------
        float St[2][5];
        {
#pragma unroll
            for (int q = 0; q < 8; ++q) v[q] = make_float2(win[q] * v[q].x, win[q] * v[q].y);
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
        for (int q = 0; q < 8; ++q) vd[q] = make_double2((double)win[q] * (double)pra[q], (double)win[q] * (double)prb[q]);
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
        if (gxb) {
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
            if (interior) {
                float* g_ = gxb + pa + t;
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    atomicAdd(g_ + q * T8, v[q].x * win[q]);
                    atomicAdd(g_ + hop + q * T8, v[q].y * win[q]);
                }
            } else {
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    if (live_a) atomicAdd(&gxb[reflect_index(pa + t + q * T8, Nsig)], v[q].x * win[q]);
                    if (live_b) atomicAdd(&gxb[reflect_index(pb + t + q * T8, Nsig)], v[q].y * win[q]);
                }
            }
            fft8_sync<N>();                            // the inverse transform's last exchange is read
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
        atomicAdd(&terms[0], l);
        atomicAdd(&terms[1], q);
        atomicAdd(loss, (l + alpha_d * q) * inv_count);
    }
}

template <int N>
static int launch_mss_fast(const float* xp, const float* xt, double* loss, double* terms, float* gx, const float* window, int B,
                           int Nsig, int hop, int F, double inv_count, double alpha_d, float eps, cudaStream_t st) {
    using P = MssFast<N>;
    if (P::smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(mss_fast_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P::smem);
        DDSP_REQUIRE(e == cudaSuccess, "mss_fast_kernel: cannot reserve %zu B of shared memory: %s", (size_t)P::smem, cudaGetErrorString(e));
    }
    static int slots_cache[64] = {0};                  // resident CTAs of this instantiation per device
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    if (slots_cache[dev] == 0) {
        int sms = 148, per_sm = 1;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) sms = 148;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mss_fast_kernel<N>, P::THREADS, P::smem) != cudaSuccess || per_sm < 1)
            per_sm = 1;
        slots_cache[dev] = sms * per_sm;
    }
    const int pairs = (F + 1) / 2;
    const int items_per_voice = (pairs + P::G - 1) / P::G;
    const long long n_items = (long long)B * items_per_voice;
    const unsigned grid = (unsigned)(n_items < slots_cache[dev] ? n_items : slots_cache[dev]);
    mss_fast_kernel<N><<<grid, P::THREADS, P::smem, st>>>(xp, xt, loss, terms, gx, window, Nsig, hop, F, inv_count, alpha_d, eps,
                                                          items_per_voice, n_items);
    DDSP_CHECK_LAUNCH("mss_fast_kernel");
    return 0;
}

__global__ void scale_by_scalar_kernel(const float* __restrict__ in, const float* __restrict__ s, float* __restrict__ out, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[i] * s[0];
}

__global__ void double_to_float_kernel(const double* __restrict__ in, float* __restrict__ out) { out[0] = (float)in[0]; }

}  // namespace ddsp

using namespace ddsp;

extern "C" int64_t ddsp_mss_count(int B, int N, int n_fft, int hop) {
    if (B < 1 || N < 1 || n_fft < 2 || hop < 1) return 0;
    return (int64_t)B * (n_fft / 2 + 1) * (1 + N / hop);
}

extern "C" int ddsp_mss_forward(const float* x_pred, const float* x_true, double* loss_terms, float* loss_f32, float* gx,
                                const int* n_ffts_host, const int* hops_host, const float* const* windows_host, int n_scales,
                                float alpha, float eps, int B, int N, void* stream) {
    DDSP_REQUIRE(x_pred && x_true && loss_terms && n_ffts_host && hops_host, "ddsp_mss_forward: null pointer argument");
    DDSP_REQUIRE(B >= 1 && B <= 65535 && N >= 2 && n_scales >= 1 && n_scales <= 64, "ddsp_mss_forward: bad B/N/n_scales");
    for (int s = 0; s < n_scales; ++s) {
        const int n = n_ffts_host[s], hop = hops_host[s];
        DDSP_REQUIRE(n >= 4 && n <= 4096 && (n & (n - 1)) == 0, "ddsp_mss_forward: n_fft=%d must be a power of two in [4,4096]", n);
        DDSP_REQUIRE(hop >= 1, "ddsp_mss_forward: hop=%d must be positive", hop);
        DDSP_REQUIRE(n / 2 < N, "ddsp_mss_forward: reflect padding n_fft/2=%d needs a signal longer than that (N=%d)", n / 2, N);
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    cudaError_t e = fft_ensure_table(st);
    if (e == cudaSuccess) e = fft8_ensure_tables(st);
    DDSP_REQUIRE(e == cudaSuccess, "ddsp_mss_forward: twiddle table initialisation failed: %s", cudaGetErrorString(e));
    e = cudaMemsetAsync(loss_terms, 0, sizeof(double) * (1 + 2 * (size_t)n_scales), st);
    if (e == cudaSuccess && gx) e = cudaMemsetAsync(gx, 0, sizeof(float) * (size_t)B * N, st);
    DDSP_REQUIRE(e == cudaSuccess, "ddsp_mss_forward: memset failed: %s", cudaGetErrorString(e));
    for (int s = 0; s < n_scales; ++s) {
        const int n = n_ffts_host[s], hop = hops_host[s];
        const int F = 1 + N / hop;
        const double cnt = (double)ddsp_mss_count(B, N, n, hop);
        const float* wptr = windows_host ? windows_host[s] : nullptr;
        double* terms = loss_terms + 1 + 2 * s;
        int rc = -1;
        // the register-tiled kernels take log2 / reciprocal of S + eps with MUFU: that needs a normal eps (the reference's
        // default is 1e-7); eps = 0 or denormal goes through the generic kernel with the library functions
        if (eps >= 1.17549435e-38f) switch (n) {
#define MSS_CASE(NN) case NN: rc = launch_mss_fast<NN>(x_pred, x_true, loss_terms, terms, gx, wptr, B, N, hop, F, 1.0 / cnt, (double)alpha, eps, st); break
            MSS_CASE(8); MSS_CASE(16); MSS_CASE(32); MSS_CASE(64); MSS_CASE(128); MSS_CASE(256);
            MSS_CASE(512); MSS_CASE(1024); MSS_CASE(2048); MSS_CASE(4096);
#undef MSS_CASE
            default: break;
        }
        if (rc == 0) continue;
        if (rc > 0) return rc;
        const MssGeom g = mss_geom(n);                  // n = 4: the generic shared-memory kernel
        const size_t smem = mss_smem_bytes(n);
        if (smem > 48 * 1024) {
            e = cudaFuncSetAttribute(mss_scale_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            DDSP_REQUIRE(e == cudaSuccess, "mss_scale_kernel: cannot reserve %zu B of shared memory: %s", smem, cudaGetErrorString(e));
        }
        const double count = (double)ddsp_mss_count(B, N, n, hop);
        const int per_cta = 2 * g.fc * MSS_ITERS;
        mss_scale_kernel<<<dim3((F + per_cta - 1) / per_cta, B), MSS_THREADS, smem, st>>>(
            x_pred, x_true, loss_terms, loss_terms + 1 + 2 * s, gx, windows_host ? windows_host[s] : nullptr, N, n, hop, F,
            1.0 / count, (double)alpha, eps);
        DDSP_CHECK_LAUNCH("mss_scale_kernel");
    }
    if (loss_f32) {
        double_to_float_kernel<<<1, 1, 0, st>>>(loss_terms, loss_f32);
        DDSP_CHECK_LAUNCH("double_to_float_kernel");
    }
    return 0;
}

extern "C" int ddsp_mss_backward(const float* gloss, const float* gx_saved, float* gx_out, int B, int N, void* stream) {
    DDSP_REQUIRE(gloss && gx_saved && gx_out && B >= 1 && N >= 1, "ddsp_mss_backward: bad arguments");
    const size_t n = (size_t)B * N;
    scale_by_scalar_kernel<<<(unsigned)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(gx_saved, gloss, gx_out, n);
    DDSP_CHECK_LAUNCH("scale_by_scalar_kernel");
    return 0;
}
