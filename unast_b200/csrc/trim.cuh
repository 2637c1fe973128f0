// K0 / K2 / K4: silence trimming (librosa.effects.trim, reference src/utils.py:248,304 and
// src/gl_vocoder.py:32), magnitude de-normalisation (utils.py:289-298) and de-emphasis
// (scipy.signal.lfilter([1],[1,-0.97]), utils.py:301).
#pragma once
#include <cfloat>
#include "common.cuh"
#include "frontend_pf.cuh"   // rfft_post_pairs (gl_init_f64_kernel)
#include "griffinlim.cuh"    // GL_F0_STRIDE

namespace unast {

__device__ __forceinline__ float ld_sample(const void* wave, int is_i16, long long idx) {
    if (is_i16) return (float)reinterpret_cast<const short*>(wave)[idx] * (1.0f / 32768.0f);
    return reinterpret_cast<const float*>(wave)[idx];
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

// energy of every 512-sample block of every utterance: be[b * nblk_max + j] = sum y^2 over block j
__global__ void block_energy_kernel(const void* wave, int is_i16, const long long* __restrict__ off, int B,
                                    int nblk_max, double* __restrict__ be) {
    const int b = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int j = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const long long base = off[b];
    const int N = (int)(off[b + 1] - base);
    if (j >= nblk_max || j * TRIM_HOP >= N) return;
    const int lo = j * TRIM_HOP, hi = min(N, lo + TRIM_HOP);
    double e = 0.0;
    for (int n = lo + lane; n < hi; n += 32) { double y = ld_sample(wave, is_i16, base + n); e += y * y; }
    e = warp_sum(e);
    if (lane == 0) be[(size_t)b * nblk_max + j] = e;
}

// One CTA per utterance: frame mean-square over reflect-padded 2048-sample frames (hop 512), threshold
// at -60 dB below the loudest frame, first / last non-silent frame -> [start, end).
__global__ void __launch_bounds__(128) trim_decide_kernel(const void* wave, int is_i16, const long long* __restrict__ off,
                                                          int nblk_max, const double* __restrict__ be,
                                                          int* __restrict__ bounds, int* err) {
    __shared__ double s_red[4];
    __shared__ double s_edge[8];
    __shared__ int s_first[4], s_last[4];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long base = off[b];
    const int N = (int)(off[b + 1] - base);
    if (N <= 0) {
        if (tid == 0) { bounds[2 * b] = 0; bounds[2 * b + 1] = 0; }     // (unast_frame_offsets reports the empty utterance)
        return;
    }
    const int nfr = 1 + N / TRIM_HOP;
    // Clips of <= 1024 samples: numpy reflects them more than once inside the 1024-sample padding; the block decomposition
    // below assumes a single reflection, so their (at most three) frames are summed directly.
    __shared__ double s_small[3];
    const bool small = N <= TRIM_FRAME / 2;
    if (small) {
        for (int i = 0; i < nfr; ++i) {
            double part = 0.0;
            for (int k = tid; k < TRIM_FRAME; k += 128) {
                const double y = ld_sample(wave, is_i16, base + reflect_index(TRIM_HOP * i - TRIM_FRAME / 2 + k, N));
                part += y * y;
            }
            part = warp_sum(part);
            if (lane == 0) s_red[warp] = part;
            __syncthreads();
            if (tid == 0) s_small[i] = s_red[0] + s_red[1] + s_red[2] + s_red[3];
            __syncthreads();
        }
    }
    const double* beb = be + (size_t)b * nblk_max;

    // reflected contributions exist only for frames 0, 1 (left) and the last <= 3 frames (right):
    // slots 0,1 = left part of frames 0,1; slots 2..4 = right part of frames nfr-3 .. nfr-1
    for (int e = 0; e < 5 && !small; ++e) {
        int i, lo, hi;                               // reflected source range [lo, hi] (inclusive) of frame i
        if (e < 2) { i = e; lo = 1; hi = TRIM_FRAME / 2 - TRIM_HOP * i; }
        else {
            i = nfr - 5 + e;
            const int last = TRIM_HOP * i + TRIM_FRAME / 2 - 1;      // last signal index the frame touches
            lo = 2 * (N - 1) - last; hi = N - 2;
        }
        double part = 0.0;
        if (i >= 0 && i < nfr && hi >= lo) {
            for (int n = lo + tid; n <= hi; n += 128) { double y = ld_sample(wave, is_i16, base + n); part += y * y; }
        }
        part = warp_sum(part);
        if (lane == 0) s_red[warp] = part;
        __syncthreads();
        if (tid == 0) s_edge[e] = s_red[0] + s_red[1] + s_red[2] + s_red[3];
        __syncthreads();
    }

    auto frame_energy = [&](int i) -> double {
        if (small) return s_small[i];
        const int lo = max(0, TRIM_HOP * i - TRIM_FRAME / 2), hi = min(N, TRIM_HOP * i + TRIM_FRAME / 2);
        double e = 0.0;
        for (int j = lo / TRIM_HOP; j * TRIM_HOP < hi; ++j) e += beb[j];
        if (i < 2) e += s_edge[i];
        if (i >= nfr - 3 && TRIM_HOP * i + TRIM_FRAME / 2 > N) e += s_edge[2 + (i - (nfr - 3))];
        return e;
    };

    double emax = 0.0;
    for (int i = tid; i < nfr; i += 128) emax = fmax(emax, frame_energy(i));
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) emax = fmax(emax, __shfl_xor_sync(0xffffffffu, emax, d));
    if (lane == 0) s_red[warp] = emax;
    __syncthreads();
    emax = fmax(fmax(s_red[0], s_red[1]), fmax(s_red[2], s_red[3]));
    const double ref_db = 10.0 * log10(fmax(1e-10, emax / TRIM_FRAME));
    int first = INT_MAX, last = -1;
    for (int i = tid; i < nfr; i += 128) {
        const double db = 10.0 * log10(fmax(1e-10, frame_energy(i) / TRIM_FRAME)) - ref_db;
        if (db > -60.0) { first = min(first, i); last = max(last, i); }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        first = min(first, __shfl_xor_sync(0xffffffffu, first, d));
        last = max(last, __shfl_xor_sync(0xffffffffu, last, d));
    }
    if (lane == 0) { s_first[warp] = first; s_last[warp] = last; }
    __syncthreads();
    if (tid == 0) {
        first = min(min(s_first[0], s_first[1]), min(s_first[2], s_first[3]));
        last = max(max(s_last[0], s_last[1]), max(s_last[2], s_last[3]));
        if (last < 0) { bounds[2 * b] = 0; bounds[2 * b + 1] = 0; }
        else { bounds[2 * b] = first * TRIM_HOP; bounds[2 * b + 1] = min(N, (last + 1) * TRIM_HOP); }
    }
}

// frame_off[b] = prefix sum of T_b = 1 + (end - start) / 275 (librosa.stft, center=True); single CTA.
__global__ void __launch_bounds__(1024) frame_offsets_kernel(const int* __restrict__ bounds, int B, int* __restrict__ frame_off,
                                                             int* err) {
    __shared__ int s_warp[32];
    __shared__ int s_total;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) { s_total = 0; frame_off[0] = 0; }
    __syncthreads();
    for (int base = 0; base < B; base += 1024) {
        const int b = base + tid;
        int T = 0;
        if (b < B) {
            const int N = bounds[2 * b + 1] - bounds[2 * b];
            if (N <= 0) atomicCAS(err, ERR_NONE, ERR_EMPTY);
            else T = 1 + N / HOP;
        }
        int x = T;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { int y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
        if (lane == 31) s_warp[warp] = x;
        __syncthreads();
        if (warp == 0) {
            int w = s_warp[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { int y = __shfl_up_sync(0xffffffffu, w, d); if (lane >= d) w += y; }
            s_warp[lane] = w;
        }
        __syncthreads();
        const int incl = x + (warp ? s_warp[warp - 1] : 0) + s_total;
        if (b < B) frame_off[b + 1] = incl;
        __syncthreads();
        if (tid == 1023) s_total = incl;
        __syncthreads();
    }
}

// K2: S = (10^((clip(m,0,1)*max_db - max_db + ref_db) * 0.05))^power, float32 (utils.py:289-298).
// power == 0 passes `mag` through unchanged (griffin_lim / invert_spectrogram called on a linear spectrogram).
__global__ void gl_prepare_kernel(const float* __restrict__ mag, long long total_frames, float power, float* __restrict__ S) {
    const long long n = total_frames * NBINS;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / NBINS;
        const int k = (int)(i - r * NBINS);
        if (power == 0.0f) { S[r * S_STRIDE + k] = mag[i]; continue; }   // input already is the linear target magnitude
        // numpy's float32 sequence, one rounding per operation (no FMA contraction); the two powf of the reference
        // (glibc: correctly rounded) as float64 pow rounded once -- the zero-phase start of utils.griffin_lim amplifies an
        // ulp of S to ~ -90 dB of the first estimate, so S has to match to the last bit, not to a few ulp
        const float c = fminf(fmaxf(mag[i], 0.0f), 1.0f);
        const float m = __fadd_rn(__fsub_rn(__fmul_rn(c, 100.0f), 100.0f), 20.0f);
        const float a = (float)exp10((double)__fmul_rn(m, 0.05f));
        S[r * S_STRIDE + k] = (float)exp2((double)power * log2((double)a));      // a^power; ~1e-15 before the rounding to float
    }
}

// K2b: the zero-phase start of utils.griffin_lim (utils.py:311-313: X_best = S, X_t = istft(S)) in float64.
// irfft of a real, smooth S piles the frame's energy up at sample 0 -- outside the window's support [473, 1575) -- so the
// taps that survive the window are the ~ -50 dB tail of a large cancellation: a float32 transform leaves them with ~3
// significant digits, and that one step decides how far the whole 100-iteration run ends from the reference (measured on
// the CPU: float32 everywhere 56-89 dB, this step alone in float64 101-108 dB; every later step is well conditioned).
// For real S the inverse is the forward transform of the even extension s[m] = S[min(m, 2048 - m)]:
//     irfft(S)[n] = X[n] / 2048,  X = rfft(s) real,  and irfft(S)[2048 - n] = irfft(S)[n],
// so the front end's float64 core (warp_fft1024_split + rfft_post_pairs) does it: one warp per frame, bins 473..1024 -> the
// windowed taps i = n - 473 = 0..551 (tap 1102 - i is the same value: the periodic Hann window is symmetric too), float32 out.
struct GlInitParams {
    const float* S;              // [F][S_STRIDE]
    long long F;
    const double2* tw;           // float64 inter-pass twiddles (front end's table)
    const double2* wp;           // float64 post-processing factors
    const double* hann;          // [1102] float64 periodic Hann
    float* frames0;              // [F][GL_F0_STRIDE]
};
constexpr int GL_INIT_WARPS = 4;
__global__ void __launch_bounds__(GL_INIT_WARPS * 32, 3) gl_init_f64_kernel(const GlInitParams P) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double* xs = reinterpret_cast<double*>(smem + warp * xs_bytes<double>());
    for (long long f = (long long)blockIdx.x * GL_INIT_WARPS + warp; f < P.F; f += (long long)gridDim.x * GL_INIT_WARPS) {
        const float* __restrict__ Sr = P.S + f * S_STRIDE;
        double2 v[32];                                                   // v[j] = (s[2 n], s[2 n + 1]) / 2, n = lane + 32 j
        static_for<0, 32>([&](auto jj) {
            constexpr int j = decltype(jj)::value;
            const int m = 2 * (lane + 32 * j);
            if constexpr (j < 16) {
                const float2 q = *reinterpret_cast<const float2*>(Sr + m);
                v[j] = mk(0.5 * (double)q.x, 0.5 * (double)q.y);
            } else {
                v[j] = mk(0.5 * (double)Sr[2048 - m], 0.5 * (double)Sr[2047 - m]);
            }
        });
        warp_fft1024_split<double, 32>(v, xs, TwMem<double>{P.tw, lane}, lane);
        float* dst = P.frames0 + f * GL_F0_STRIDE;
        auto put = [&](int n, double xr) {                               // bin n of X -> tap n - 473
            const int i = n - WIN_LPAD;
            if (i >= 0) dst[i] = (float)(xr * (1.0 / NFFT) * P.hann[i]);
        };
        rfft_post_pairs<double>(v, reinterpret_cast<double2*>(xs), WpLane<double>{P.wp[lane]}, lane, [&](auto jj, double2 xa, double2 xb) {
            constexpr int j = decltype(jj)::value;
            put(lane + 32 * j, xa.x);
            put(1024 - lane - 32 * j, xb.x);
        }, [&](double2 x) { put(512, x.x); });
        __syncwarp();                                                    // the exchange buffers are reused by the next frame
    }
}

// K4: y[n] = x[n] + 0.97 y[n-1] in float64 (scipy.signal.lfilter), x = ola / wss on the fly, float32 out.
// A warp owns a chunk; the recurrence is restarted `WARM` samples early (0.97^2048 ~ 1e-27, far below
// float64 resolution), so chunks are independent.  Within the chunk: 32-sample inclusive scans with carry.
struct DeemphParams {
    const float* ola;            // padded OLA sums, or nullptr when `x` is given
    const float* x;              // plain input signal (already normalised), or nullptr
    const int* frame_off;
    const long long* pad_off;
    const long long* wav_off;    // [B+1]
    float* out;
    const float* wss_per;
    const double* w2;
    double coef;                 // 0.97, or 0 to skip de-emphasis (copy / normalise only)
};

constexpr int DE_CHUNK = 4096;
constexpr int DE_WARM = 2048;

__global__ void deemph_kernel(const DeemphParams P) {
    const int b = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const int chunk = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int L = (int)(P.wav_off[b + 1] - P.wav_off[b]);
    const int c0 = chunk * DE_CHUNK;
    if (c0 >= L) return;
    const int c1 = min(L, c0 + DE_CHUNK);
    const int Tb = P.frame_off ? P.frame_off[b + 1] - P.frame_off[b] : 0;
    const float* src = P.ola ? P.ola + P.pad_off[b] + HALF : P.x + P.wav_off[b];
    float* dst = P.out + P.wav_off[b];
    const double a = P.coef;
    double apow[5];
    apow[0] = a;
#pragma unroll
    for (int i = 1; i < 5; ++i) apow[i] = apow[i - 1] * apow[i - 1];
    double alane = a;                                   // a^(lane+1)
    for (int i = 0; i < lane; ++i) alane *= a;
    double carry = 0.0;
    for (int n0 = max(0, c0 - DE_WARM) & ~31; n0 < c1; n0 += 32) {
        const int n = n0 + lane;
        double v = 0.0;
        if (n < c1) {
            float y = src[n];
            if (P.ola) { const float w = wss_at(n + HALF, Tb, P.wss_per, P.w2); if (w > FLT_MIN) y = y / w; }
            v = (double)y;
        }
#pragma unroll
        for (int s = 0; s < 5; ++s) {
            const double t = __shfl_up_sync(0xffffffffu, v, 1 << s);
            if (lane >= (1 << s)) v = fma(apow[s], t, v);
        }
        v = fma(alane, carry, v);
        carry = __shfl_sync(0xffffffffu, v, 31);
        if (n >= c0 && n < c1) dst[n] = (float)v;
    }
}

// The samples a fragment owns (sharded long-form Griffin-Lim): out = ola / window-sum envelope, no de-emphasis.
// Fragment b = frames [t0, t0 + n) of an utterance of Tg frames owns wave samples [275 t0 - 551, 275 (t0 + n) - 551),
// clipped to the utterance; its local overlap-add buffer starts at padded position 275 t0.
struct GlExtractParams {
    const float* ola;
    const int* frame_off;
    const long long* pad_off;
    const int* t0;
    const int* Tg;
    const long long* raw_off;    // [B+1] prefix sum of the fragments' sample counts
    float* out;
    const float* wss_per;
    const double* w2;
};

__global__ void gl_extract_kernel(const GlExtractParams P) {
    const int b = blockIdx.y;
    const int t0 = P.t0[b], Tg = P.Tg[b];
    const long long n_a = t0 == 0 ? 0 : (long long)HOP * t0 - (HALF - WIN_LPAD);
    const long long cnt = P.raw_off[b + 1] - P.raw_off[b];
    const float* src = P.ola + P.pad_off[b] - (long long)HOP * t0;         // indexed by the utterance's padded position
    float* dst = P.out + P.raw_off[b];
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < cnt; i += (long long)gridDim.x * blockDim.x) {
        const long long q = n_a + i + HALF;
        float y = src[q];
        const float w = wss_at((int)q, Tg, P.wss_per, P.w2);
        if (w > FLT_MIN) y = y / w;
        dst[i] = y;
    }
}

}  // namespace unast
