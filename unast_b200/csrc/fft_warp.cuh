// Warp-level 2048-point real FFT for sm_100a: one warp owns one frame.
//
// A real frame of 2048 samples (only the 1102 window taps are non-zero, circularly shifted so they
// start at sample 0 -- a shift only multiplies bin k by a unit phase, which |X| ignores and the
// Griffin-Lim forward/inverse pair cancels) is packed into 1024 complex points
// z[n] = x[2n] + i x[2n+1] and transformed as 32 x 32:
//   lane = n1, register = n2 (n = n1 + 32 n2)
//   pass 1: 32-point FFT over the registers            (fft32, all in registers)
//   twiddle W1024^(n1 k1) + transpose through shared memory (one exchange per FFT)
//   pass 2: 32-point FFT over the registers            -> Z[k1 + 32 k2] in lane k1 / register k2
//   real post-processing pairs bin k with 1024-k (mirrors parked in the idle transposition buffer).
// The index algebra is validated on the CPU by tools/fft_model.py.
//
// Replaces numpy.fft.rfft / irfft as called by librosa.stft / istft on the reference path
// (reference src/utils.py:254-257, :314, :328).
#pragma once
#include "common.cuh"

namespace unast {

// cos(2 pi j / 32), j = 0..8 (correctly rounded to double; float instantiations round once more).
constexpr double kCos32[9] = {
    1.0,
    0.980785280403230449126182236134,
    0.923879532511286756128183189397,
    0.831469612302545237078788377618,
    0.707106781186547524400844362105,
    0.555570233019602224742830813949,
    0.382683432365089771728459984030,
    0.195090322016128267848284868477,
    0.0};

__host__ __device__ constexpr double cos32(int j) { return j <= 8 ? kCos32[j] : -kCos32[16 - j]; }   // j in 0..16
__host__ __device__ constexpr double sin32(int j) { return j <= 8 ? kCos32[8 - j] : kCos32[j - 8]; }  // j in 0..16

// a * exp(-2 pi i J / 32), J in 0..15 (forward DIF twiddles)
template <int J, class C>
__device__ __forceinline__ C mul_w32(C a) {
    using T = decltype(a.x);
    if constexpr (J == 0) {
        return a;
    } else if constexpr (J == 8) {
        return mk(a.y, -a.x);
    } else if constexpr (J == 4) {
        constexpr T r = (T)kCos32[4];
        return mk((a.x + a.y) * r, (a.y - a.x) * r);
    } else if constexpr (J == 12) {
        constexpr T r = (T)kCos32[4];
        return mk((a.y - a.x) * r, -(a.x + a.y) * r);
    } else {
        constexpr T c = (T)cos32(J), s = (T)sin32(J);   // W = c - i s
        return mk(fma_t(a.x, c, a.y * s), fma_t(a.y, c, -(a.x * s)));
    }
}

__host__ __device__ constexpr int brev5(int i) {
    return ((i & 1) << 4) | ((i & 2) << 2) | (i & 4) | ((i & 8) >> 2) | ((i & 16) >> 4);
}

// One radix-2 decimation-in-time butterfly with the twiddle folded into the additions:
//   x = a + W b,  y = a - W b,   W = exp(-2 pi i J / 32), J in 0..15
// General W = c - i s: x in four fused multiply-adds, y = 2 a - x in two more (6 instructions against 8 for multiply-then-add);
// W = (+-1 - i) / sqrt 2: two additions and four fused multiply-adds; W = 1, -i: four additions.
template <int J, class C>
__device__ __forceinline__ void bfly_w32(C a, C b, C& x, C& y) {
    using T = decltype(a.x);
    if constexpr (J == 0) {
        x = cadd(a, b);
        y = csub(a, b);
    } else if constexpr (J == 8) {                       // W b = (b.y, -b.x)
        x = mk(a.x + b.y, a.y - b.x);
        y = mk(a.x - b.y, a.y + b.x);
    } else if constexpr (J == 4) {                       // W b = r (b.x + b.y, b.y - b.x)
        constexpr T r = (T)kCos32[4];
        const T sm = b.x + b.y, df = b.y - b.x;
        x = mk(fma_t(r, sm, a.x), fma_t(r, df, a.y));
        y = mk(fma_t(-r, sm, a.x), fma_t(-r, df, a.y));
    } else if constexpr (J == 12) {                      // W b = r (b.y - b.x, -(b.x + b.y))
        constexpr T r = (T)kCos32[4];
        const T sm = b.x + b.y, df = b.y - b.x;
        x = mk(fma_t(r, df, a.x), fma_t(-r, sm, a.y));
        y = mk(fma_t(-r, df, a.x), fma_t(r, sm, a.y));
    } else {                                             // W b = (c b.x + s b.y, c b.y - s b.x)
        constexpr T c = (T)cos32(J), sn = (T)sin32(J);
        x = mk(fma_t(c, b.x, fma_t(sn, b.y, a.x)), fma_t(c, b.y, fma_t(-sn, b.x, a.y)));
        y = mk(fma_t((T)2, a.x, -x.x), fma_t((T)2, a.y, -x.y));
    }
}

template <int SPAN, class C>
__device__ __forceinline__ void dit_stage(C (&v)[32]) {
    static_for<0, 16>([&](auto ii) {
        constexpr int i = decltype(ii)::value;
        constexpr int base = (i / SPAN) * 2 * SPAN, j = i % SPAN;
        bfly_w32<j*(16 / SPAN)>(v[base + j], v[base + j + SPAN], v[base + j], v[base + j + SPAN]);
    });
}

// In-register 32-point forward FFT (radix-2 decimation in time on the bit-reversed registers -- a compile-time renaming --,
// natural order in and out).
// NZ: only v[0..NZ) are non-zero on entry (16 <= NZ <= 32): the first stage pairs (v[r], v[r + 16]); where v[r + 16] is zero
// the butterfly is two copies.
template <int NZ, class C>
__device__ __forceinline__ void fft32(C (&v)[32]) {
    static_assert(NZ >= 16 && NZ <= 32, "pruning supports 16..32 live inputs");
    C t[32];
    static_for<0, 16>([&](auto ii) {
        constexpr int i = decltype(ii)::value;
        constexpr int r = brev5(2 * i);                  // brev5(2 i + 1) == r + 16
        if constexpr (r + 16 < NZ) {
            t[2 * i] = cadd(v[r], v[r + 16]);
            t[2 * i + 1] = csub(v[r], v[r + 16]);
        } else {
            t[2 * i] = v[r];
            t[2 * i + 1] = v[r];
        }
    });
    dit_stage<2>(t);
    dit_stage<4>(t);
    dit_stage<8>(t);
    dit_stage<16>(t);
    static_for<0, 32>([&](auto kk) { constexpr int k = decltype(kk)::value; v[k] = t[k]; });
}

// cos(2 pi j / 64), j = 0..16
constexpr double kCos64[17] = {
    1.0,
    0.99518472667219688624483695310948,
    0.980785280403230449126182236134239,
    0.95694033573220886493579788698027,
    0.923879532511286756128183189396788,
    0.881921264348355029712756863660388,
    0.831469612302545237078788377617906,
    0.77301045336273696081090660975847,
    0.707106781186547524400844362104849,
    0.634393284163645498215171613225493,
    0.555570233019602224742830813948533,
    0.471396736825997648556387625905254,
    0.382683432365089771728459984030399,
    0.290284677254462367636192375817395,
    0.195090322016128267848284868477022,
    0.0980171403295606019941955638886418,
    0.0};
__host__ __device__ constexpr double cos64(int j) { return j <= 16 ? kCos64[j] : -kCos64[32 - j]; }   // j in 0..32
__host__ __device__ constexpr double sin64(int j) { return j <= 16 ? kCos64[16 - j] : kCos64[j - 16]; }  // j in 0..32

// Source of the real-FFT post-processing factors  wp2[k] = -i W2048^k = (-sin th, -cos th), th = 2 pi k / 2048,
// for k = lane + 32 j.  Two policies:
//   WpTable: read the 1024-entry table (shared memory) -- one load per bin, no arithmetic (float32 kernels, which
//            are issue-bound);
//   WpLane : wp2[lane + 32 j] = wp2[lane] * W64^j with the per-lane factor in registers and W64^j a compile-time
//            constant -- four extra FMAs per bin, no memory traffic (float64 kernels, which are bound by the
//            shared-memory data path).
template <class T>
struct WpTable {
    const cplx_t<T>* tab;
    template <int J>
    __device__ __forceinline__ cplx_t<T> get(int lane) const { return tab[lane + 32 * J]; }
};
template <class T>
struct WpLane {
    cplx_t<T> b;   // wp2[lane]
    template <int J>
    __device__ __forceinline__ cplx_t<T> get(int) const {
        if constexpr (J == 0) {
            return b;
        } else if constexpr (J == 16) {                       // W64^16 = -i
            return mk(b.y, -b.x);
        } else {
            constexpr T c = (T)cos64(J), sn = (T)sin64(J);    // W64^J = c - i sn
            return mk(fma_t(b.x, c, b.y * sn), fma_t(b.y, c, -(b.x * sn)));
        }
    }
};

// ---- the 1024-point complex transform across one warp ---------------------------------------------------
// warp_fft1024_split transposes the real and the imaginary plane one after the other through a buffer of
// 32 x XS scalars (float: 4.5 KB, double: 8.5 KB per warp); the real-FFT post-processing on mirror pairs lives in
// frontend_pf.cuh (rfft_post_pairs) and fft_pk.cuh (pk_post_pairs / pk_pre_pairs).
template <class T> struct XsStride;
template <> struct XsStride<float> { static constexpr int value = 36; };    // 144 B rows: conflict-free 128-bit reads
template <> struct XsStride<double> { static constexpr int value = 34; };   // 272 B rows
template <class T> __host__ __device__ constexpr int xs_bytes() { return 32 * XsStride<T>::value * (int)sizeof(T); }

// mid(): called between the two planes, where only the 32 imaginary parts are live in registers -- the cheapest point
// of the transform for anything that needs registers of its own (e.g. a barrier wait).
// twiddle sources for warp_fft1024_split: a [k1][lane] table in shared / global memory, or the same table in tensor
// memory (row k1 = columns 4 k1 .. 4 k1 + 3 of the thread's lane for double2; float2 rows are not used from TMEM)
template <class T>
struct TwMem {
    const cplx_t<T>* tw;
    int lane;
    __device__ __forceinline__ void apply(cplx_t<T> (&v)[32]) const {
        static_for<1, 32>([&](auto kk) { constexpr int k1 = decltype(kk)::value; v[k1] = cmul(v[k1], tw[k1 * 32 + lane]); });
    }
};
struct TwTmemF64 {
    uint32_t base;                                   // lane quadrant | column of row 0
    // four twiddles per tcgen05.ld, the next four in flight while the current ones are used
    __device__ __forceinline__ void apply(double2 (&v)[32]) const {
        uint32_t a[16], b[16];
        tmem_ld16(base, a);
        static_for<0, 8>([&](auto cc) {
            constexpr int c = decltype(cc)::value;
            auto use = [&](uint32_t (&r)[16]) {
                static_for<0, 4>([&](auto ii) {
                    constexpr int k1 = 4 * c + decltype(ii)::value, i = decltype(ii)::value;
                    if constexpr (k1 > 0) {
                        const double wx = __hiloint2double((int)r[4 * i + 1], (int)r[4 * i]), wy = __hiloint2double((int)r[4 * i + 3], (int)r[4 * i + 2]);
                        v[k1] = cmul(v[k1], make_double2(wx, wy));
                    }
                });
            };
            if constexpr (c % 2 == 0) {
                tmem_wait16(a);
                if constexpr (c + 1 < 8) tmem_ld16(base + 16 * (c + 1), b);
                use(a);
            } else {
                tmem_wait16(b);
                if constexpr (c + 1 < 8) tmem_ld16(base + 16 * (c + 1), a);
                use(b);
            }
        });
    }
};

template <class T, int NZ, class TW, class MID>
__device__ __forceinline__ void warp_fft1024_split(cplx_t<T> (&v)[32], T* xs, const TW& tw, int lane, MID&& mid) {
    using C = cplx_t<T>;
    constexpr int XS = XsStride<T>::value;
    fft32<NZ>(v);
    tw.apply(v);
    auto read_row = [&](auto sel) {
        if constexpr (sizeof(T) == 4) {
            const float4* row = reinterpret_cast<const float4*>(xs + lane * XS);
            static_for<0, 8>([&](auto ii) {
                constexpr int i = decltype(ii)::value;
                const float4 q = row[i];
                sel(v[4 * i]) = q.x; sel(v[4 * i + 1]) = q.y; sel(v[4 * i + 2]) = q.z; sel(v[4 * i + 3]) = q.w;
            });
        } else {
            const double2* row = reinterpret_cast<const double2*>(xs + lane * XS);
            static_for<0, 16>([&](auto ii) {
                constexpr int i = decltype(ii)::value;
                const double2 q = row[i];
                sel(v[2 * i]) = q.x; sel(v[2 * i + 1]) = q.y;
            });
        }
    };
    static_for<0, 32>([&](auto kk) { constexpr int k1 = decltype(kk)::value; xs[k1 * XS + lane] = v[k1].x; });
    __syncwarp();
    mid();
    read_row([](C& c) -> T& { return c.x; });
    __syncwarp();
    static_for<0, 32>([&](auto kk) { constexpr int k1 = decltype(kk)::value; xs[k1 * XS + lane] = v[k1].y; });
    __syncwarp();
    read_row([](C& c) -> T& { return c.y; });
    __syncwarp();
    fft32<32>(v);
}

template <class T, int NZ, class TW>
__device__ __forceinline__ void warp_fft1024_split(cplx_t<T> (&v)[32], T* xs, const TW& tw, int lane) {
    warp_fft1024_split<T, NZ>(v, xs, tw, lane, []() {});
}

}  // namespace unast
