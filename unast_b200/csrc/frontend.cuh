// K1 building blocks shared by the fused mel / magnitude front end (replaces reference src/utils.py:251-276):
// the launch parameter block, the dB normalisation (utils.py:267-272), PCM16 -> float conversion, pre-emphasis with
// numpy's float32 roundings (utils.py:251) and the frame loader with librosa.stft's reflect padding (utils.py:254).
// The kernel itself is frontend_au_kernel (frontend_au.cuh).
//
// T = double reproduces the reference's float64 FFT (numpy.fft under librosa.stft) and is the mode that meets the
// 1e-4 parity gate on every bin; T = float is the fast mode (DESIGN.md "Precision").
#pragma once
#include "fft_warp.cuh"

namespace unast {

struct FrontendParams {
    const void* wave;            // concatenated utterances, float32 or int16 (16-byte aligned base for int16)
    int wave_is_i16;
    const long long* wave_off;   // [B+1] first sample of each utterance in `wave`
    const int* bounds;           // [B][2] trim start / end relative to the utterance
    const int* frame_off;        // [B+1] prefix sum of T_b
    int B;
    float* mag_out;              // [sum T][1025] or nullptr
    float* mel_out;
    int mel_layout;              // 0: per-utterance mel-major [80][T_b] at 80*frame_off[b]; 1: [B][T_max][80]
    int T_max;
    const void* win2;            // cplx_t<T>[551]  (w[2n], w[2n+1]) * 0.5 (* 2^-15 for PCM16 input)
    const void* tw1024;          // cplx_t<T>[32*32] W1024^(n1 k1) laid out [k1][n1]
    const void* wp2;             // cplx_t<T>[1024] (-sin th, -cos th), th = 2 pi k / 2048
    int* err;
};



// (20*log10(max(1e-5, m)) - ref_db + max_db) / max_db clipped to [1e-8, 1]  (utils.py:267-272), from m
__device__ __forceinline__ float db_norm(float m) {
    const float t = fmaf(0.060205999132796240f, lg2_approx(fmaxf(1e-5f, m)), 0.8f);
    return fminf(fmaxf(t, 1e-8f), 1.0f);
}
// the same from the power m^2:  20 log10(max(1e-5, m)) = 10 log10(max(1e-10, m^2))
__device__ __forceinline__ float db_norm_pw(float pw) {
    const float t = fmaf(0.030102999566398120f, lg2_approx(fmaxf(1e-10f, pw)), 0.8f);
    return fminf(fmaxf(t, 1e-8f), 1.0f);
}

__device__ __forceinline__ float load_sample(const void* wave, int is_i16, long long idx) {
    if (is_i16) return (float)reinterpret_cast<const short*>(wave)[idx] * (1.0f / 32768.0f);
    return reinterpret_cast<const float*>(wave)[idx];
}

// int16 halves of a 32-bit word -> float (exact), without the slow I2F path:
// bits 0x4B000000 | (s + 32768) is the float 8388608 + s + 32768.
__device__ __forceinline__ float s16lo_to_f(uint32_t w) {
    return __uint_as_float((w & 0xffffu) ^ 0x4B008000u) - 8421376.0f;
}
__device__ __forceinline__ float s16hi_to_f(uint32_t w) {
    return __uint_as_float((w >> 16) ^ 0x4B008000u) - 8421376.0f;
}

// y[n] = x[n] - 0.97 x[n-1] rounded like numpy float32 arithmetic (utils.py:251)
__device__ __forceinline__ float preemph(float cur, float prev) { return __fsub_rn(cur, __fmul_rn(0.97f, prev)); }

// Loads one frame (window taps only, circularly shifted to start at 0), pre-emphasised and windowed, into
// v[j] = (y[2n] w[2n], y[2n+1] w[2n+1]),  n = lane + 32 j.
// wbase: index of the utterance's first (trimmed) sample in `wave`; N: trimmed length; t: frame index.
// For PCM16 the samples are kept as integer-valued floats: the 2^-15 lives in the window table (exact).
// Edge taps of a clip that one reflection does not cover (<= 1024 samples; numpy.pad reflects repeatedly): out of line, so
// that the hot kernel carries a call it never takes for LJSpeech-length audio.
template <bool I16>
__device__ __noinline__ void edge_taps_general(float* scratch, const void* wave, long long wbase, int N, int n0, int lane) {
    for (int i = lane; i < WIN; i += 32) {
        const int m = reflect_index(n0 + i, N);
        float cur, prev = 0.0f;
        if constexpr (I16) {
            const short* s = reinterpret_cast<const short*>(wave) + wbase;
            cur = (float)s[m];
            if (m > 0) prev = (float)s[m - 1];
        } else {
            const float* s = reinterpret_cast<const float*>(wave) + wbase;
            cur = s[m];
            if (m > 0) prev = s[m - 1];
        }
        scratch[i] = m > 0 ? preemph(cur, prev) : cur;
    }
}

template <class T, bool I16>
__device__ __forceinline__ void load_frame(cplx_t<T> (&v)[32], const void* __restrict__ wave, long long wbase, int N, int t,
                                           const cplx_t<T>* __restrict__ s_win, float* scratch, int lane) {
    using C = cplx_t<T>;
    const int n0 = HOP * t + WIN_LPAD - HALF;                  // signal index of tap 0 (may be negative)
    const bool interior = (n0 >= 1) && (n0 + WIN + 2 <= N);    // no reflection, x[n0-1] exists, one word of slack
    if (interior) {
        if constexpr (I16) {
            const long long A = wbase + n0 - 1;                // absolute index of the "previous" sample of tap 0
            const uint32_t* __restrict__ w32 = reinterpret_cast<const uint32_t*>(wave) + (A >> 1) + lane;
            if (A & 1) {                                       // warp-uniform
                static_for<0, 18>([&](auto jj) {
                    constexpr int j = decltype(jj)::value;
                    const int n = lane + 32 * j;
                    if (j < 17 || n < NLIVE_C) {
                        const uint32_t a = w32[32 * j], b = w32[32 * j + 1];
                        const float pv = s16hi_to_f(a), e = s16lo_to_f(b), o = s16hi_to_f(b);
                        const C w = s_win[n];
                        v[j] = mk((T)preemph(e, pv) * w.x, (T)preemph(o, e) * w.y);
                    } else {
                        v[j] = mk((T)0, (T)0);
                    }
                });
            } else {
                static_for<0, 18>([&](auto jj) {
                    constexpr int j = decltype(jj)::value;
                    const int n = lane + 32 * j;
                    if (j < 17 || n < NLIVE_C) {
                        const uint32_t a = w32[32 * j], b = w32[32 * j + 1];
                        const float pv = s16lo_to_f(a), e = s16hi_to_f(a), o = s16lo_to_f(b);
                        const C w = s_win[n];
                        v[j] = mk((T)preemph(e, pv) * w.x, (T)preemph(o, e) * w.y);
                    } else {
                        v[j] = mk((T)0, (T)0);
                    }
                });
            }
        } else {
            const float* __restrict__ x = reinterpret_cast<const float*>(wave) + wbase + n0 + 2 * lane;
            static_for<0, 18>([&](auto jj) {
                constexpr int j = decltype(jj)::value;
                const int n = lane + 32 * j;
                if (j < 17 || n < NLIVE_C) {
                    const float pv = x[64 * j - 1], e = x[64 * j], o = x[64 * j + 1];
                    const C w = s_win[n];
                    v[j] = mk((T)preemph(e, pv) * w.x, (T)preemph(o, e) * w.y);
                } else {
                    v[j] = mk((T)0, (T)0);
                }
            });
        }
    } else {
        // edge frame: reflect padding of librosa.stft(center=True); y[0] = x[0] (np.append(y[0], ...), utils.py:251)
        auto sample_at = [&](int m) -> float {                 // pre-emphasised signal at index m in [0, N)
            float cur, prev = 0.0f;
            if constexpr (I16) {
                const short* s = reinterpret_cast<const short*>(wave) + wbase;
                cur = (float)s[m];
                if (m > 0) prev = (float)s[m - 1];
            } else {
                const float* s = reinterpret_cast<const float*>(wave) + wbase;
                cur = s[m];
                if (m > 0) prev = s[m - 1];
            }
            return m > 0 ? preemph(cur, prev) : cur;
        };
        // compact runtime loop (edge frames are rare; unrolling this path 36x bloats the instruction cache): stage the
        // taps in the warp's idle exchange buffer, then pick them up with static register indices.  The loop is bound by
        // the latency of its global loads, so its body stays branch-free (one reflection, what every clip of more than
        // 1024 samples needs) and unrollable; taps that one reflection does not bring home are redone below.
        bool again = false;
#pragma unroll 4
        for (int i = lane; i < WIN; i += 32) {
            int m = n0 + i;
            m = m < 0 ? -m : (m >= N ? 2 * (N - 1) - m : m);
            const bool ok = m >= 0 && m < N;
            again |= !ok;
            scratch[i] = sample_at(ok ? m : 0);
        }
        if (__any_sync(0xffffffffu, again)) edge_taps_general<I16>(scratch, wave, wbase, N, n0, lane);   // clips of <= 1024 samples
        __syncwarp();
        static_for<0, 18>([&](auto jj) {
            constexpr int j = decltype(jj)::value;
            const int n = lane + 32 * j;
            if (j < 17 || n < NLIVE_C) {
                const C w = s_win[n];
                const float2 y = reinterpret_cast<const float2*>(scratch)[n];
                v[j] = mk((T)y.x * w.x, (T)y.y * w.y);
            } else {
                v[j] = mk((T)0, (T)0);
            }
        });
        __syncwarp();
    }
    static_for<18, 32>([&](auto jj) { constexpr int j = decltype(jj)::value; v[j] = mk((T)0, (T)0); });
}

}  // namespace unast
