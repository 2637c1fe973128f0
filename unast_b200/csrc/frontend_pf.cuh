// K1 building blocks (continued): the mirror-pair real-FFT post-processing, the cp.async sample prefetch and the
// PCM16 word -> windowed complex input conversion used by frontend_au_kernel (reference src/utils.py:251-260).
#pragma once
#include "frontend.cuh"

namespace unast {

constexpr int PF_STAGE_BYTES = 2304;        // 140 16-byte chunks cover 1104 int16 samples at any alignment (+ pad)
constexpr int PF_SLACK = 24;                // samples of slack after the last tap so that chunk reads stay in bounds


// Real-FFT post-processing by pairs.  in: v[j] = Z[lane + 32 j] / 2.  Lane l computes, for j < 16, bin ka = l + 32 j
// and its mirror kb = 1024 - ka (lane 0, j = 0: DC and Nyquist) and hands both to emit(j, Xa, Xb); lane 0 also
// emits bin 512 through emit512(X).  With w = wp2[ka] = (-sin th, -cos th):
//   er = z.x + p.x, ei = z.y - p.y, or = z.x - p.x, oi = z.y + p.y   (p = Z[1024 - ka] / 2)
//   A = w.x or - w.y oi, B = w.x oi + w.y or;   X[ka] = (er + A, B + ei),  X[kb] = (er - A, B - ei).
// The mirror value of (lane l, register j) sits in lane 32 - l, register 31 - j (column 0 pairs with itself:
// register j <-> 32 - j).  Every lane parks its upper registers j' >= 16 in the idle transposition buffer, row
// 32 - j' (rows 1..16, 32 complex values each), lane 0 also its register 0 in row 0; then lane l >= 1 reads row
// j + 1, column 32 - l and lane 0 reads row j, column 0 -- one base pointer per lane, no selects, and only the 16
// lower registers stay live in the pair loop (room to batch the loads).  buf: 17 x 32 complex values.
template <class T, class WP, class F, class G>
__device__ __forceinline__ void rfft_post_pairs(cplx_t<T> (&v)[32], cplx_t<T>* buf, const WP& wp, int lane, F&& emit, G&& emit512) {
    using C = cplx_t<T>;
    static_for<16, 32>([&](auto jj) { constexpr int j = decltype(jj)::value; buf[(32 - j) * 32 + lane] = v[j]; });
    if (lane == 0) buf[0] = v[0];
    __syncwarp();
    const C* mine = lane ? buf + 32 + (32 - lane) : buf;
    static_for<0, 16>([&](auto jj) {
        constexpr int j = decltype(jj)::value;
        const C p = mine[32 * j];
        const C z = v[j], w = wp.template get<j>(lane);
        const T er = z.x + p.x, ei = z.y - p.y, orr = z.x - p.x, oi = z.y + p.y;
        const T A = fma_t(w.x, orr, -(w.y * oi)), B = fma_t(w.x, oi, w.y * orr);
        emit(jj, mk(er + A, B + ei), mk(er - A, B - ei));
    });
    if (lane == 0) {                                  // bin 512 (register 16 of lane 0) pairs with itself: X = 2 conj(z)
        const C z = buf[16 * 32];
        emit512(mk(z.x + z.x, -(z.y + z.y)));
    }
    __syncwarp();
}

// int16 halves of a word that was XORed with 0x80008000 -> float (exact): bytes {lo16, 0x00, 0x4B} are the float
// 8388608 + 32768 + s.
__device__ __forceinline__ float x16lo_to_f(uint32_t w) { return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7610)) - 8421376.0f; }
__device__ __forceinline__ float x16hi_to_f(uint32_t w) { return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7632)) - 8421376.0f; }

template <class T, bool ODD>
__device__ __forceinline__ void window_i16_words_a(cplx_t<T> (&v)[32], const uint32_t* w32, const cplx_t<T>* __restrict__ s_win, int lane) {
    using C = cplx_t<T>;
    static_for<0, 18>([&](auto jj) {
        constexpr int j = decltype(jj)::value;
        const int n = lane + 32 * j;
        if (j < 17 || n < NLIVE_C) {
            const uint32_t a = w32[32 * j] ^ 0x80008000u, b = w32[32 * j + 1] ^ 0x80008000u;
            // odd: prev = hi(a), even tap = lo(b), odd tap = hi(b);  else: prev = lo(a), even tap = hi(a), odd tap = lo(b)
            const float pv = ODD ? x16hi_to_f(a) : x16lo_to_f(a);
            const float e = ODD ? x16lo_to_f(b) : x16hi_to_f(a);
            const float o = ODD ? x16hi_to_f(b) : x16lo_to_f(b);
            const C w = s_win[n];
            v[j] = mk((T)preemph(e, pv) * w.x, (T)preemph(o, e) * w.y);
        } else {
            v[j] = mk((T)0, (T)0);
        }
    });
    static_for<18, 32>([&](auto jj) { constexpr int j = decltype(jj)::value; v[j] = mk((T)0, (T)0); });
}

// frame samples from 32-bit words holding int16 pairs (the staging buffer): w32 points at the word that holds the
// sample before tap 0 (in its high half when odd), already offset by the lane.
template <class T>
__device__ __forceinline__ void window_i16_words(cplx_t<T> (&v)[32], const uint32_t* w32, bool odd,
                                                 const cplx_t<T>* __restrict__ s_win, int lane) {
    if (odd) window_i16_words_a<T, true>(v, w32, s_win, lane);
    else window_i16_words_a<T, false>(v, w32, s_win, lane);
}

// db_norm_pw without the 1e-10 floor: any power below it (zero included: lg2 -> -inf) lands below 0 and is clipped to
// 1e-8 exactly like the floored value (-0.2).
__device__ __forceinline__ float db_norm_pw_nf(float pw) {
    return fminf(fmaxf(fmaf(0.030102999566398120f, lg2_approx(pw), 0.8f), 1e-8f), 1.0f);
}

// is frame t of an utterance with N (trimmed) samples served by the prefetch path?
__device__ __forceinline__ bool pf_interior(int t, int N) {
    const int n0 = HOP * t + WIN_LPAD - HALF;
    return (n0 >= 1) && (n0 + WIN + PF_SLACK <= N);
}

// issue the cp.async copies of frame t's samples (16-byte chunks around [A, A + 1104)), A = wbase + n0 - 1
__device__ __forceinline__ void pf_issue(const void* wave, long long wbase, int t, unsigned char* stage, int lane) {
    const long long A = wbase + (HOP * t + WIN_LPAD - HALF) - 1;
    const unsigned char* src = reinterpret_cast<const unsigned char*>(wave) + ((2 * A) & ~15LL);
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        const int c = lane + 32 * i;
        if (c < 140) cp_async16(stage + 16 * c, src + 16 * c);
    }
    cp_async_commit();
}

}  // namespace unast
