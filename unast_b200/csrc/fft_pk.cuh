// Packed-float32 variant of the warp-level 1024-point complex FFT (fft_warp.cuh): the same 32 x 32 decomposition, with
// the in-register 32-point transforms issued as f32x2 instructions (FADD2 / FMUL2 / FFMA2 of sm_100).
//
// A lane's 32 complex values live as 16 + 16 register pairs, real and imaginary parts apart (Pk).  Two pairings:
//   layout A: pair j holds elements (j, j + 16)      -- what a radix-2 DIF stage of span 16 leaves independent: after
//             that first (scalar) stage the two halves are two 16-point transforms with identical structure and
//             identical twiddles, so stages 2..5 run on pairs with duplicated (c, c) constants;
//   layout B: pair m holds elements (2 m, 2 m + 1)   -- where the bit-reversed outputs of those stages land
//             (pair p of layout A ends up holding outputs brev5(p), brev5(p) + 1; renamed at compile time).
// fft32_ab maps A -> B.  The twiddle between the two passes multiplies pairs of layout B by a table laid out to
// match, and the transposition writes element n1 of a row to column 2 (n1 % 16) + n1 / 16, so that the 128-bit
// row reads of the second pass come back as layout-A pairs without a single register move.
// Per 32-point transform: 175 packed + <= 128 scalar instructions (A -> B, decimation in frequency), 148 packed + 92 scalar
// (B -> A, decimation in time with the twiddles folded into fused multiply-adds), instead of 456 scalar ones.
//
// Replaces numpy.fft.rfft / irfft as called by librosa.stft / istft (reference src/utils.py:314, :328).
#pragma once
#include "fft_warp.cuh"

namespace unast {

struct Pk { float2 re[16], im[16]; };

// scalar views of element J in the two layouts
template <int J> __device__ __forceinline__ float& reA(Pk& p) { if constexpr (J < 16) return p.re[J].x; else return p.re[J - 16].y; }
template <int J> __device__ __forceinline__ float& imA(Pk& p) { if constexpr (J < 16) return p.im[J].x; else return p.im[J - 16].y; }
template <int J> __device__ __forceinline__ float& reB(Pk& p) { if constexpr (J % 2 == 0) return p.re[J / 2].x; else return p.re[J / 2].y; }
template <int J> __device__ __forceinline__ float& imB(Pk& p) { if constexpr (J % 2 == 0) return p.im[J / 2].x; else return p.im[J / 2].y; }

__device__ __forceinline__ float2 neg2(float2 a) { return make_float2(-a.x, -a.y); }
__device__ __forceinline__ float2 dup2(float c) { return make_float2(c, c); }

// ---- where a [rows][32 lanes] float4 table lives ---------------------------------------------------------------------
// Both policies hand the table to `use(chunk, r)` four rows (16 registers) at a time, in order.
//   TabMem : shared or global memory, one 128-bit load per row;
//   TabTmem: tensor memory (row m = columns 4 m .. 4 m + 3 of the thread's lane), one tcgen05.ld per chunk, the next chunk
//            in flight while the current one is used.
struct TabMem {
    const float4* t;
    int lane;
    template <int NCH, class F>
    __device__ __forceinline__ void stream(F&& use) const {
        static_for<0, NCH>([&](auto cc) {
            constexpr int c = decltype(cc)::value;
            uint32_t r[16];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float4 q = t[(4 * c + i) * 32 + lane];
                r[4 * i] = __float_as_uint(q.x); r[4 * i + 1] = __float_as_uint(q.y); r[4 * i + 2] = __float_as_uint(q.z); r[4 * i + 3] = __float_as_uint(q.w);
            }
            use(cc, r);
        });
    }
};
struct TabTmem {
    uint32_t base;                                   // lane quadrant | first column
    template <int NCH, class F>
    __device__ __forceinline__ void stream(F&& use) const {
        uint32_t a[16], b[16];
        tmem_ld16(base, a);
        static_for<0, NCH>([&](auto cc) {
            constexpr int c = decltype(cc)::value;
            if constexpr (c % 2 == 0) {
                tmem_wait16(a);
                if constexpr (c + 1 < NCH) tmem_ld16(base + 16 * (c + 1), b);
                use(cc, a);
            } else {
                tmem_wait16(b);
                if constexpr (c + 1 < NCH) tmem_ld16(base + 16 * (c + 1), a);
                use(cc, b);
            }
        });
    }
};

__host__ __device__ constexpr int brev4(int i) { return ((i & 1) << 3) | ((i & 2) << 1) | ((i & 4) >> 1) | ((i & 8) >> 3); }

// (re, im) * exp(-2 pi i J / 32) on pairs, J in 0..15
template <int J>
__device__ __forceinline__ void pk_mul_w32(float2& re, float2& im) {
    if constexpr (J == 0) {
    } else if constexpr (J == 8) {
        const float2 t = re;
        re = im;
        im = neg2(t);
    } else if constexpr (J == 4) {
        const float2 r = dup2((float)kCos32[4]);
        const float2 s = __fadd2_rn(re, im), d = __fadd2_rn(im, neg2(re));
        re = __fmul2_rn(s, r);
        im = __fmul2_rn(d, r);
    } else if constexpr (J == 12) {
        const float2 r = dup2((float)kCos32[4]);
        const float2 s = __fadd2_rn(re, im), d = __fadd2_rn(im, neg2(re));
        re = __fmul2_rn(d, r);
        im = __fmul2_rn(s, neg2(r));
    } else {
        const float2 c = dup2((float)cos32(J)), s = dup2((float)sin32(J));   // W = c - i s
        const float2 nr = __ffma2_rn(re, c, __fmul2_rn(im, s));
        const float2 ni = __ffma2_rn(im, c, __fmul2_rn(re, neg2(s)));
        re = nr;
        im = ni;
    }
}

// one DIF stage of the two 16-point halves at once (SPAN = 8, 4, 2, 1 inside a half)
template <int SPAN>
__device__ __forceinline__ void pk_stage(Pk& p) {
    static_for<0, 8>([&](auto ii) {
        constexpr int i = decltype(ii)::value;
        constexpr int lo = (i / SPAN) * 2 * SPAN + i % SPAN, hi = lo + SPAN, J = (i % SPAN) * (8 / SPAN) * 2;
        const float2 ar = p.re[lo], ai = p.im[lo], br = p.re[hi], bi = p.im[hi];
        p.re[lo] = __fadd2_rn(ar, br);
        p.im[lo] = __fadd2_rn(ai, bi);
        float2 dr = __fadd2_rn(ar, neg2(br)), di = __fadd2_rn(ai, neg2(bi));
        pk_mul_w32<J>(dr, di);
        p.re[hi] = dr;
        p.im[hi] = di;
    });
}

// In-register 32-point forward FFT, layout A in -> layout B out (natural order).
// NZ: only elements [0, NZ) are non-zero on entry (16 <= NZ <= 32) -- prunes the scalar first stage.
template <int NZ>
__device__ __forceinline__ void fft32_ab(Pk& p) {
    static_assert(NZ >= 16 && NZ <= 32, "pruning supports 16..32 live inputs");
    static_for<0, 16>([&](auto jj) {
        constexpr int j = decltype(jj)::value;
        const float2 a = make_float2(p.re[j].x, p.im[j].x);
        if constexpr (j + 16 < NZ) {
            const float2 b = make_float2(p.re[j].y, p.im[j].y);
            const float2 t = mul_w32<j>(make_float2(a.x - b.x, a.y - b.y));
            p.re[j] = make_float2(a.x + b.x, t.x);
            p.im[j] = make_float2(a.y + b.y, t.y);
        } else {
            const float2 t = mul_w32<j>(a);
            p.re[j].y = t.x;
            p.im[j].y = t.y;
        }
    });
    pk_stage<8>(p);
    pk_stage<4>(p);
    pk_stage<2>(p);
    pk_stage<1>(p);
    Pk t;
    static_for<0, 16>([&](auto mm) { constexpr int m = decltype(mm)::value; t.re[m] = p.re[brev4(m)]; t.im[m] = p.im[brev4(m)]; });
    p = t;
}

// column of element n1 in a row of the transposition buffer
__device__ __forceinline__ int pk_col(int n1) { return 2 * (n1 & 15) + (n1 >> 4); }

// 1024-point complex forward FFT across one warp, split-plane transposition (see warp_fft1024_split).
// in : layout A, element n2 = z[lane + 32 n2] (only the first NZ elements non-zero)
// out: layout B, element k2 = Z[lane + 32 k2]
// tw4: [16][32] float4 (c[2m], c[2m+1], s[2m], s[2m+1]) of W1024^(lane k1) = c + i s, k1 = 2m, 2m + 1
template <int NZ, class TAB>
__device__ __forceinline__ void warp_fft1024_pk(Pk& p, float* xs, const TAB& tw4, int lane) {
    constexpr int XS = XsStride<float>::value;
    fft32_ab<NZ>(p);
    tw4.template stream<4>([&](auto cc, const uint32_t (&r)[16]) {
        static_for<0, 4>([&](auto ii) {
            constexpr int m = 4 * decltype(cc)::value + decltype(ii)::value, i = decltype(ii)::value;
            const float2 c = make_float2(__uint_as_float(r[4 * i]), __uint_as_float(r[4 * i + 1]));
            const float2 s = make_float2(__uint_as_float(r[4 * i + 2]), __uint_as_float(r[4 * i + 3]));
            const float2 nr = __ffma2_rn(p.re[m], c, __fmul2_rn(p.im[m], neg2(s)));
            const float2 ni = __ffma2_rn(p.re[m], s, __fmul2_rn(p.im[m], c));
            p.re[m] = nr;
            p.im[m] = ni;
        });
    });
    float* col = xs + pk_col(lane);
    const float4* row = reinterpret_cast<const float4*>(xs + lane * XS);
    static_for<0, 32>([&](auto kk) { constexpr int k1 = decltype(kk)::value; col[k1 * XS] = reB<k1>(p); });
    __syncwarp();
    static_for<0, 8>([&](auto ii) {
        constexpr int i = decltype(ii)::value;
        const float4 q = row[i];
        p.re[2 * i] = make_float2(q.x, q.y);
        p.re[2 * i + 1] = make_float2(q.z, q.w);
    });
    __syncwarp();
    static_for<0, 32>([&](auto kk) { constexpr int k1 = decltype(kk)::value; col[k1 * XS] = imB<k1>(p); });
    __syncwarp();
    static_for<0, 8>([&](auto ii) {
        constexpr int i = decltype(ii)::value;
        const float4 q = row[i];
        p.im[2 * i] = make_float2(q.x, q.y);
        p.im[2 * i + 1] = make_float2(q.z, q.w);
    });
    __syncwarp();
    fft32_ab<32>(p);
}

// One decimation-in-time butterfly on two register pairs, the same twiddle W = exp(-2 pi i J / 32) in both halves, folded
// into the additions like bfly_w32 (fft_warp.cuh):  x = a + W b -> LO,  y = a - W b -> HI.
template <int LO, int HI, int J>
__device__ __forceinline__ void pk_bfly_dit(Pk& p) {
    const float2 ar = p.re[LO], ai = p.im[LO], br = p.re[HI], bi = p.im[HI];
    if constexpr (J == 0) {
        p.re[LO] = __fadd2_rn(ar, br); p.im[LO] = __fadd2_rn(ai, bi);
        p.re[HI] = __fadd2_rn(ar, neg2(br)); p.im[HI] = __fadd2_rn(ai, neg2(bi));
    } else if constexpr (J == 8) {                       // W b = (b.im, -b.re)
        p.re[LO] = __fadd2_rn(ar, bi); p.im[LO] = __fadd2_rn(ai, neg2(br));
        p.re[HI] = __fadd2_rn(ar, neg2(bi)); p.im[HI] = __fadd2_rn(ai, br);
    } else if constexpr (J == 4 || J == 12) {            // W b = r (sm, df) (J = 4), r (df, -sm) (J = 12)
        const float2 r = dup2((float)kCos32[4]);
        const float2 sm = __fadd2_rn(br, bi), df = __fadd2_rn(bi, neg2(br));
        if constexpr (J == 4) {
            p.re[LO] = __ffma2_rn(r, sm, ar); p.im[LO] = __ffma2_rn(r, df, ai);
            p.re[HI] = __ffma2_rn(neg2(r), sm, ar); p.im[HI] = __ffma2_rn(neg2(r), df, ai);
        } else {
            p.re[LO] = __ffma2_rn(r, df, ar); p.im[LO] = __ffma2_rn(neg2(r), sm, ai);
            p.re[HI] = __ffma2_rn(neg2(r), df, ar); p.im[HI] = __ffma2_rn(r, sm, ai);
        }
    } else {                                             // W = c - i s:  W b = (c b.re + s b.im, c b.im - s b.re)
        const float2 c = dup2((float)cos32(J)), s = dup2((float)sin32(J)), two = dup2(2.0f);
        const float2 xr = __ffma2_rn(c, br, __ffma2_rn(s, bi, ar)), xi = __ffma2_rn(c, bi, __ffma2_rn(neg2(s), br, ai));
        p.re[LO] = xr; p.im[LO] = xi;
        p.re[HI] = __ffma2_rn(two, ar, neg2(xr)); p.im[HI] = __ffma2_rn(two, ai, neg2(xi));
    }
}

// In-register 32-point forward FFT, layout B in -> layout A out (natural order), radix-2 decimation in time.  With
// t[k] = v[brev5(k)], register pair m holds (t[q], t[q + 16]), q = brev4(m): the stages of span 1, 2, 4, 8 act between
// register pairs with the SAME twiddle in both halves (trivial twiddles stay trivial), the last stage (span 16) inside each
// pair, scalar.  Pair m ends up holding outputs (q, q + 16): layout A after a compile-time renaming.
// 148 packed + 92 scalar instructions (the decimation-in-frequency form it replaces: 256 + 64).
__device__ __forceinline__ void fft32_ba(Pk& p) {
    static_for<0, 4>([&](auto ss) {
        constexpr int S = 1 << decltype(ss)::value;      // span in t-index
        static_for<0, 8>([&](auto ii) {
            constexpr int i = decltype(ii)::value;
            constexpr int k = (i / S) * 2 * S + i % S;   // t-index of the butterfly's first input, k + S the second
            pk_bfly_dit<brev4(k), brev4(k + S), (i % S) * (16 / S)>(p);
        });
    });
    static_for<0, 16>([&](auto mm) {
        constexpr int m = decltype(mm)::value, q = brev4(m);
        float2 x, y;
        bfly_w32<q>(make_float2(p.re[m].x, p.im[m].x), make_float2(p.re[m].y, p.im[m].y), x, y);
        p.re[m] = make_float2(x.x, y.x);
        p.im[m] = make_float2(x.y, y.y);
    });
    Pk t;
    static_for<0, 16>([&](auto aa) { constexpr int a = decltype(aa)::value; t.re[a] = p.re[brev4(a)]; t.im[a] = p.im[brev4(a)]; });
    p = t;
}

// The same transform for an input in layout B (first pass = fft32_ba, twiddle on layout-A pairs).
// tw4a: [16][32] float4 (c[j], c[j + 16], s[j], s[j + 16]) of W1024^(lane k1), k1 = j, j + 16
template <class TAB, class F>
__device__ __forceinline__ void warp_fft1024_pk_b(Pk& p, float* xs, const TAB& tw4a, int lane, F&& xs_idle) {
    constexpr int XS = XsStride<float>::value;
    fft32_ba(p);
    tw4a.template stream<4>([&](auto cc, const uint32_t (&r)[16]) {
        static_for<0, 4>([&](auto ii) {
            constexpr int j = 4 * decltype(cc)::value + decltype(ii)::value, i = decltype(ii)::value;
            const float2 c = make_float2(__uint_as_float(r[4 * i]), __uint_as_float(r[4 * i + 1]));
            const float2 s = make_float2(__uint_as_float(r[4 * i + 2]), __uint_as_float(r[4 * i + 3]));
            const float2 nr = __ffma2_rn(p.re[j], c, __fmul2_rn(p.im[j], neg2(s)));
            const float2 ni = __ffma2_rn(p.re[j], s, __fmul2_rn(p.im[j], c));
            p.re[j] = nr;
            p.im[j] = ni;
        });
    });
    float* col = xs + pk_col(lane);
    const float4* row = reinterpret_cast<const float4*>(xs + lane * XS);
    static_for<0, 32>([&](auto kk) { constexpr int k1 = decltype(kk)::value; col[k1 * XS] = reA<k1>(p); });
    __syncwarp();
    static_for<0, 8>([&](auto ii) {
        constexpr int i = decltype(ii)::value;
        const float4 q = row[i];
        p.re[2 * i] = make_float2(q.x, q.y);
        p.re[2 * i + 1] = make_float2(q.z, q.w);
    });
    __syncwarp();
    static_for<0, 32>([&](auto kk) { constexpr int k1 = decltype(kk)::value; col[k1 * XS] = imA<k1>(p); });
    __syncwarp();
    static_for<0, 8>([&](auto ii) {
        constexpr int i = decltype(ii)::value;
        const float4 q = row[i];
        p.im[2 * i] = make_float2(q.x, q.y);
        p.im[2 * i + 1] = make_float2(q.z, q.w);
    });
    __syncwarp();
    xs_idle();                                   // the buffer is free from here on (e.g. to prefetch into)
    fft32_ab<32>(p);
}

// ---- real-FFT post- / pre-processing by mirror pairs, packed ------------------------------------------------
// Layout B throughout.  Lane l keeps elements 0..15 (pairs m < 8: bins ka = l + 64 m, ka + 32) and fetches their
// mirrors kb = 1024 - ka, which are elements 31 - 2m, 30 - 2m of lane 32 - l (column 0 pairs with itself: element
// j <-> 32 - j, and bin 0 with the Nyquist bin), from a park buffer in shared memory: plane[row][slot], 33 rows of
// PK_PARK_S floats; lane l' parks element e >= 16 at row l', slot 31 - e; lane 0 instead at row 32, slot 32 - e, plus
// its element 0 at slot 0 -- so that every lane reads its two mirrors with one aligned 64-bit load from one base.
// After pk_post_pairs: pair m = X[ka] (both halves), pair m + 8 = X[kb] of the same two bins' mirrors.
// pk_pre_pairs takes that back to Z' (conjugated, times 2): own elements in place, the mirrors
// returned through the same rows (only lane l and its mirror lane ever touch row 32 - l), then read back as pairs
// 8..15.  Lane 0's element 16 (bin 512, its own mirror) is handled by the caller through pk_park_512 / row 0 slot 0.
constexpr int PK_PARK_S = 18;
constexpr int PK_PARK_PLANE = 33 * PK_PARK_S;             // floats per plane (re, im)
constexpr int PK_PARK_BYTES = 2 * PK_PARK_PLANE * 4;      // 4752

// wp4: [8][32] float4 (wx[ka0], wx[ka1], wy[ka0], wy[ka1]),  wp2[k] = (wx, wy) = (-sin th, -cos th), th = 2 pi k / 2048
template <class TAB>
__device__ __forceinline__ void pk_post_pairs(Pk& p, float* park, const TAB& wp4, int lane) {
    float* pre = park;
    float* pim = park + PK_PARK_PLANE;
    {
        const int wb = lane ? lane * PK_PARK_S : 32 * PK_PARK_S + 1;
        static_for<8, 16>([&](auto mm) {
            constexpr int m = decltype(mm)::value;
            pre[wb + 31 - 2 * m] = p.re[m].x; pre[wb + 30 - 2 * m] = p.re[m].y;
            pim[wb + 31 - 2 * m] = p.im[m].x; pim[wb + 30 - 2 * m] = p.im[m].y;
        });
        if (lane == 0) { pre[32 * PK_PARK_S] = p.re[0].x; pim[32 * PK_PARK_S] = p.im[0].x; }
    }
    __syncwarp();
    const int rb = lane ? (32 - lane) * PK_PARK_S : 32 * PK_PARK_S;
    wp4.template stream<2>([&](auto cc, const uint32_t (&r)[16]) {
        static_for<0, 4>([&](auto ii) {
            constexpr int m = 4 * decltype(cc)::value + decltype(ii)::value, i = decltype(ii)::value;
            const float2 qr = *reinterpret_cast<const float2*>(pre + rb + 2 * m), qi = *reinterpret_cast<const float2*>(pim + rb + 2 * m);
            const float2 wx = make_float2(__uint_as_float(r[4 * i]), __uint_as_float(r[4 * i + 1]));
            const float2 wy = make_float2(__uint_as_float(r[4 * i + 2]), __uint_as_float(r[4 * i + 3]));
            const float2 zr = p.re[m], zi = p.im[m];
            const float2 er = __fadd2_rn(zr, qr), ei = __fadd2_rn(zi, neg2(qi)), orr = __fadd2_rn(zr, neg2(qr)), oi = __fadd2_rn(zi, qi);
            // A = wx orr - wy oi, B = wx oi + wy orr:  X[ka] = (er + A, B + ei),  X[kb] = (er - A, B - ei), folded
            const float2 two = dup2(2.0f);
            const float2 xr = __ffma2_rn(wx, orr, __ffma2_rn(neg2(wy), oi, er)), xi = __ffma2_rn(wx, oi, __ffma2_rn(wy, orr, ei));
            p.re[m] = xr;
            p.im[m] = xi;
            p.re[m + 8] = __ffma2_rn(two, er, neg2(xr));
            p.im[m + 8] = __ffma2_rn(neg2(two), ei, xi);
        });
    });
}

template <class TAB>
__device__ __forceinline__ void pk_pre_pairs(Pk& p, float* park, const TAB& wp4, int lane) {
    float* pre = park;
    float* pim = park + PK_PARK_PLANE;
    const int wb = lane ? (32 - lane) * PK_PARK_S : 1;
    wp4.template stream<2>([&](auto cc, const uint32_t (&r)[16]) {
      static_for<0, 4>([&](auto ii) {
        constexpr int m = 4 * decltype(cc)::value + decltype(ii)::value, i = decltype(ii)::value;
        const float2 wx = make_float2(__uint_as_float(r[4 * i]), __uint_as_float(r[4 * i + 1]));
        const float2 wy = make_float2(__uint_as_float(r[4 * i + 2]), __uint_as_float(r[4 * i + 3]));
        const float2 xr = p.re[m], xi = p.im[m], rr = p.re[m + 8], ri = p.im[m + 8];
        const float2 er = __fadd2_rn(xr, rr), ei = __fadd2_rn(xi, neg2(ri)), dr = __fadd2_rn(xr, neg2(rr)), di = __fadd2_rn(xi, ri);
        // C = wx dr + wy di, D = wx di - wy dr:  own = (er + C, -(D + ei)),  mirror = (er - C, -(D - ei)), folded
        const float2 two = dup2(2.0f);
        const float2 yr = __ffma2_rn(wx, dr, __ffma2_rn(wy, di, er));
        const float2 yi = __ffma2_rn(wy, dr, __ffma2_rn(neg2(wx), di, neg2(ei)));
        p.re[m] = yr;
        p.im[m] = yi;
        const float2 br = __ffma2_rn(two, er, neg2(yr)), bi = __ffma2_rn(two, ei, yi);
        pre[wb + 15 - 2 * m] = br.x; pre[wb + 14 - 2 * m] = br.y;
        pim[wb + 15 - 2 * m] = bi.x; pim[wb + 14 - 2 * m] = bi.y;
      });
    });
}
// after the caller has dealt with bin 512 and synchronised the warp: elements 16..31 back from the own row
__device__ __forceinline__ void pk_pre_collect(Pk& p, const float* park, int lane) {
    const float* pre = park + lane * PK_PARK_S;
    const float* pim = pre + PK_PARK_PLANE;
    static_for<8, 16>([&](auto mm) {
        constexpr int m = decltype(mm)::value;
        p.re[m] = *reinterpret_cast<const float2*>(pre + 2 * (m - 8));
        p.im[m] = *reinterpret_cast<const float2*>(pim + 2 * (m - 8));
    });
}

}  // namespace unast
