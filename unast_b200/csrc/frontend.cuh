// K1: fused mel / magnitude front end (replaces reference src/utils.py:251-276).
//
// One persistent CTA per SM walks a contiguous range of "steps"; a step is FPS consecutive frames of
// one utterance.  Per step:
//   1. all threads stage the step's samples into shared memory (int16/float load, reflect padding of
//      librosa.stft(center=True), pre-emphasis 0.97 in float32 exactly as numpy rounds it) as two
//      planes (even / odd sample) so that the packed-complex frame loads are conflict-free;
//   2. each warp transforms one frame (window -> 32x32 warp FFT -> real post-processing), writes the
//      dB-normalised magnitude row straight to HBM (coalesced) and leaves |X| in its own exchange
//      buffer, which doubles as the A operand row of the filterbank contraction;
//   3. the 1024x80 Slaney filterbank is applied on the tensor cores as a banded 3xTF32
//      mma.sync.m16n8k8 (only the (k-block, n-tile) pairs that hold non-zeros), then dB-normalised.
// T = float is the fast mode; T = double reproduces the reference's float64 FFT (numpy.fft) and is
// the mode that meets the 1e-4 parity gate on every bin (DESIGN.md "precision").
#pragma once
#include "fft_warp.cuh"

namespace unast {

struct MelPlan {                 // built once on the host (tables.cu), lives in device memory
    const float4* bfrag;         // [npairs][32] {b0_hi, b1_hi, b0_lo, b1_lo} in mma B-fragment order
    const int* pair_kb;          // [npairs] k-block (8 bins) of each pair
    int pair_begin[11];          // pairs of n-tile nt are [pair_begin[nt], pair_begin[nt+1])
};

struct FrontendParams {
    const void* wave;            // concatenated utterances, float32 or int16
    int wave_is_i16;
    const long long* wave_off;   // [B+1] first sample of each utterance in `wave`
    const int* bounds;           // [B][2] trim start / end relative to the utterance
    const int* frame_off;        // [B+1] prefix sum of T_b
    const int* step_off;         // [B+1] prefix sum of ceil(T_b / FPS)
    int B;
    float* mag_out;              // [sum T][1025] or nullptr
    float* mel_out;
    int mel_layout;              // 0: per-utterance mel-major [80][T_b] at 80*frame_off[b]; 1: [B][T_max][80]
    int T_max;
    const void* win2;            // cplx_t<T>[551]  (w[2n], w[2n+1])
    const void* tw1024;          // cplx_t<T>[32*32] W1024^(n1 k1) laid out [k1][n1]
    const void* wp;              // cplx_t<T>[1024] (-sin/2, -cos/2)
    MelPlan mel;
    int* err;
};

constexpr int A_STRIDE_PAD = 4;  // words of padding between warps' exchange regions (A-row bank spread)

template <class T, int WARPS>
struct FeLayout {
    using C = cplx_t<T>;
    static constexpr int FPS = WARPS;                                   // frames per step
    static constexpr int SPAN = HOP * (FPS - 1) + WIN;                  // samples a step touches
    static constexpr int PLANE = (SPAN + 1) / 2 + 1;
    static constexpr int XCH_BYTES = xch_elems<T>() * (int)sizeof(C) + A_STRIDE_PAD * 4;
    static constexpr int off_xch = 0;
    static constexpr int off_planes = off_xch + WARPS * XCH_BYTES;
    static constexpr int off_win = (off_planes + 2 * PLANE * 4 + 15) / 16 * 16;
    static constexpr int off_tw = off_win + (NLIVE_C * (int)sizeof(C) + 15) / 16 * 16;
    static constexpr int off_wp = off_tw + 1024 * (int)sizeof(C);
    static constexpr int bytes = off_wp + 1024 * (int)sizeof(C);
};

__device__ __forceinline__ uint32_t to_tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// (20*log10(max(1e-5, m)) - ref_db + max_db) / max_db clipped to [1e-8, 1]  (utils.py:267-272)
__device__ __forceinline__ float db_norm(float m) {
    float t = 20.0f * (0.30102999566398120f * __log2f(fmaxf(1e-5f, m)));
    t = (t - 20.0f + 100.0f) / 100.0f;
    return fminf(fmaxf(t, 1e-8f), 1.0f);
}

__device__ __forceinline__ float load_sample(const void* wave, int is_i16, long long idx) {
    if (is_i16) return (float)reinterpret_cast<const short*>(wave)[idx] * (1.0f / 32768.0f);
    return reinterpret_cast<const float*>(wave)[idx];
}

template <class T, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1) frontend_kernel(const FrontendParams P) {
    using C = cplx_t<T>;
    using L = FeLayout<T, WARPS>;
    constexpr int FPS = L::FPS;
    constexpr int NT = WARPS * 32;
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    C* xch = reinterpret_cast<C*>(smem + L::off_xch + warp * L::XCH_BYTES);
    float* PE = reinterpret_cast<float*>(smem + L::off_planes);
    float* PO = PE + L::PLANE;
    C* s_win = reinterpret_cast<C*>(smem + L::off_win);
    C* s_tw = reinterpret_cast<C*>(smem + L::off_tw);
    C* s_wp = reinterpret_cast<C*>(smem + L::off_wp);

    for (int i = tid; i < NLIVE_C; i += NT) s_win[i] = reinterpret_cast<const C*>(P.win2)[i];
    for (int i = tid; i < 1024; i += NT) {
        s_tw[i] = reinterpret_cast<const C*>(P.tw1024)[i];
        s_wp[i] = reinterpret_cast<const C*>(P.wp)[i];
    }

    const int total_steps = P.step_off[P.B];
    const int g0 = (int)(((long long)total_steps * blockIdx.x) / gridDim.x);
    const int g1 = (int)(((long long)total_steps * (blockIdx.x + 1)) / gridDim.x);
    if (g0 >= g1) return;
    int b = 0;
    {   // binary search: last b with step_off[b] <= g0
        int lo = 0, hi = P.B;
        while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (P.step_off[mid] <= g0) lo = mid; else hi = mid; }
        b = lo;
    }
    __syncthreads();

    for (int g = g0; g < g1; ++g) {
        while (g >= P.step_off[b + 1]) ++b;
        const int ta = (g - P.step_off[b]) * FPS;
        const int fbase = P.frame_off[b];
        const int Tb = P.frame_off[b + 1] - fbase;
        const int nf = min(FPS, Tb - ta);
        const int s0 = P.bounds[2 * b], N = P.bounds[2 * b + 1] - s0;
        const long long wbase = P.wave_off[b] + s0;
        const int p0 = HOP * ta + WIN_LPAD - HALF;       // signal index of the step's first staged sample

        // ---- 1. stage: reflect padding + pre-emphasis, split into even/odd planes ----
        const int span = HOP * (nf - 1) + WIN;
        for (int i = tid; i < span; i += NT) {
            int n = p0 + i;
            if (n < 0) n = -n; else if (n >= N) n = 2 * (N - 1) - n;
            float cur = load_sample(P.wave, P.wave_is_i16, wbase + n);
            float y = cur;
            if (n > 0) {
                float prev = load_sample(P.wave, P.wave_is_i16, wbase + n - 1);
                y = __fsub_rn(cur, __fmul_rn(0.97f, prev));              // utils.py:251, float32 like numpy
            }
            if (i & 1) PO[i >> 1] = y; else PE[i >> 1] = y;
        }
        __syncthreads();

        // ---- 2. one frame per warp ----
        float* arow = reinterpret_cast<float*>(xch);
        if (warp < nf) {
            const int o = HOP * warp;                    // frame offset inside the staged span
            const float* pa = (o & 1) ? PO + (o >> 1) : PE + (o >> 1);          // x'[2n]
            const float* pb = (o & 1) ? PE + ((o + 1) >> 1) : PO + (o >> 1);    // x'[2n+1]
            C v[32];
            static_for<0, 18>([&](auto jj) {
                constexpr int j = decltype(jj)::value;
                const int n = lane + 32 * j;
                if (j < 17 || n < NLIVE_C) {
                    C w = s_win[n];
                    v[j] = mk((T)pa[n] * w.x, (T)pb[n] * w.y);
                } else {
                    v[j] = mk((T)0, (T)0);
                }
            });
            warp_fft1024<T, 18>(v, xch, s_tw, lane);
            T nyq;
            rfft_post<T>(v, nyq, xch, s_wp, lane);

            float* mrow = P.mag_out ? P.mag_out + (size_t)(fbase + ta + warp) * NBINS : nullptr;
            static_for<0, 32>([&](auto jj) {
                constexpr int j = decltype(jj)::value;
                const int k = lane + 32 * j;
                float re = (float)v[j].x, im = (float)v[j].y;          // complex64 storage of librosa.stft
                float pw = fmaf(re, re, im * im);
                float m = pw > 0.0f ? pw * rsqrtf(pw) : 0.0f;           // np.abs
                arow[k] = m;
                if (mrow) mrow[k] = db_norm(m);
            });
            if (lane == 0 && mrow) mrow[1024] = db_norm(fabsf((float)nyq));
        }
        __syncthreads();

        // ---- 3. banded 3xTF32 filterbank contraction on the tensor cores (utils.py:263-264) ----
        for (int nt = warp; nt < NMELS / 8; nt += WARPS) {
            const int gq = lane >> 2, tq = lane & 3;
            const float* a_lo = reinterpret_cast<const float*>(smem + L::off_xch + gq * L::XCH_BYTES);
            const float* a_hi = reinterpret_cast<const float*>(smem + L::off_xch + (gq + 8) * L::XCH_BYTES);
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            for (int p = P.mel.pair_begin[nt]; p < P.mel.pair_begin[nt + 1]; ++p) {
                const int kb = P.mel.pair_kb[p];
                const float4 bq = P.mel.bfrag[p * 32 + lane];
                float a[4];
                a[0] = a_lo[8 * kb + tq];
                a[2] = a_lo[8 * kb + tq + 4];
                if (FPS > 8) { a[1] = a_hi[8 * kb + tq]; a[3] = a_hi[8 * kb + tq + 4]; } else { a[1] = 0.f; a[3] = 0.f; }
                uint32_t ah[4], al[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) { ah[q] = to_tf32(a[q]); al[q] = to_tf32(a[q] - __uint_as_float(ah[q])); }
                mma_tf32(acc, al, __float_as_uint(bq.x), __float_as_uint(bq.y));
                mma_tf32(acc, ah, __float_as_uint(bq.z), __float_as_uint(bq.w));
                mma_tf32(acc, ah, __float_as_uint(bq.x), __float_as_uint(bq.y));
            }
            const int m0 = 8 * nt + 2 * tq;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int f = gq + 8 * h;
                if (f < nf) {
                    const float v0 = db_norm(acc[2 * h]), v1 = db_norm(acc[2 * h + 1]);
                    if (P.mel_layout == 0) {
                        float* dst = P.mel_out + (size_t)NMELS * fbase + (ta + f);
                        dst[(size_t)m0 * Tb] = v0;
                        dst[(size_t)(m0 + 1) * Tb] = v1;
                    } else {
                        float* dst = P.mel_out + ((size_t)b * P.T_max + ta + f) * NMELS + m0;
                        dst[0] = v0;
                        dst[1] = v1;
                    }
                }
            }
        }
        __syncthreads();
    }
}

}  // namespace unast
