// K1: fused mel / magnitude front end (replaces reference src/utils.py:251-276).
//
// One persistent CTA per SM walks a contiguous range of "tiles"; a tile is R = FW * RPW consecutive frames of one
// utterance.  The CTA is warp-specialised and has no CTA-wide barrier in its steady state:
//
//   FW "FFT warps"   each owns RPW frames of the tile and runs them end to end on its own:
//                    coalesced global loads of the frame's 1102 samples (PCM16 words or float32; the 4x overlap
//                    between consecutive frames is served by L1/L2), reflect padding of librosa.stft(center=True)
//                    on the few edge frames, pre-emphasis 0.97 in float32 exactly as numpy rounds it, periodic
//                    Hann window -> 32x32 warp FFT in registers (one private shared-memory transposition) ->
//                    real post-processing -> |X|^2; the dB-normalised magnitude row goes straight to HBM with
//                    coalesced stores, the raw |X| row goes to the tile's row buffer in shared memory.
//   MW "mel warps"   wait (mbarrier) until every row of the tile is there, apply the 1024x80 Slaney filterbank on
//                    the tensor cores as a banded 3xTF32 mma.sync.m16n8k8 (only the 149 (k-block, n-tile) pairs
//                    that hold non-zeros), dB-normalise and store mel, then hand the row buffer back.
//
// T = double reproduces the reference's float64 FFT (numpy.fft under librosa.stft) and is the mode that meets the
// 1e-4 parity gate on every bin; T = float is the fast mode (DESIGN.md "Precision").
#pragma once
#include "fft_warp.cuh"

namespace unast {

struct MelPlan {                 // built once on the host, lives in device memory
    const float2* bfrag;         // [npairs][32] {b0, b1} float32 weights in mma B-fragment order
    const int* pair_kb;          // [npairs] k-block (8 bins) of each pair
    int pair_begin[11];          // pairs of n-tile nt are [pair_begin[nt], pair_begin[nt+1])
};

struct FrontendParams {
    const void* wave;            // concatenated utterances, float32 or int16 (4-byte aligned base)
    int wave_is_i16;
    const long long* wave_off;   // [B+1] first sample of each utterance in `wave`
    const int* bounds;           // [B][2] trim start / end relative to the utterance
    const int* frame_off;        // [B+1] prefix sum of T_b
    const int* tile_off;         // [B+1] prefix sum of ceil(T_b / R)
    int B;
    float* mag_out;              // [sum T][1025] or nullptr
    float* mel_out;
    int mel_layout;              // 0: per-utterance mel-major [80][T_b] at 80*frame_off[b]; 1: [B][T_max][80]
    int T_max;
    const void* win2;            // cplx_t<T>[551]  (w[2n], w[2n+1]) * 0.5 (* 2^-15 for PCM16 input)
    const void* tw1024;          // cplx_t<T>[32*32] W1024^(n1 k1) laid out [k1][n1]
    const void* wp2;             // cplx_t<T>[1024] (-sin th, -cos th), th = 2 pi k / 2048
    MelPlan mel;
    int* err;
};

constexpr int ROW_STRIDE = 1028;         // floats per |X| row: 4 (mod 32) makes the mma A-fragment reads conflict-free
constexpr int MEL_MAX_PAIRS = 160;

template <class T, int FW, int RPW, int MW>
struct FeLayout {
    using C = cplx_t<T>;
    static constexpr int R = FW * RPW;                                  // frames per tile (<= 16)
    static constexpr bool WP_TABLE = sizeof(T) == 4;                    // float: table; double: per-lane factor
    static constexpr int XCH_BYTES = xch_elems<T>() * (int)sizeof(C);
    static constexpr int off_xch = 0;
    static constexpr int off_rows = off_xch + FW * XCH_BYTES;
    static constexpr int off_win = off_rows + 16 * ROW_STRIDE * 4;
    static constexpr int off_tw = off_win + (NLIVE_C * (int)sizeof(C) + 15) / 16 * 16;
    static constexpr int off_wp = off_tw + 1024 * (int)sizeof(C);
    static constexpr int off_kb = off_wp + (WP_TABLE ? 1024 * (int)sizeof(C) : 0);
    static constexpr int off_bar = off_kb + MEL_MAX_PAIRS * 4;
    static constexpr int off_tmem = off_bar + 2 * 8;
    static constexpr int bytes = off_tmem + 8;
    static_assert(R <= 16, "a tile is one m16 MMA tile of frames");
    static_assert(bytes <= 227 * 1024, "shared memory budget");
};

// TF32 operand split for 3xTF32: x = hi + lo with hi = x truncated to 10 mantissa bits (one LOP3) and lo = x - hi
// (exact).  The tensor core ignores the low 13 bits of a .tf32 operand, so lo can be passed as it is; the error of
// the three-product sum is ~2^-20 relative.  (cvt.rna.tf32.f32 is emulated with ~5 instructions on sm_100.)
__device__ __forceinline__ uint32_t tf32_hi(float x) { return __float_as_uint(x) & 0xffffe000u; }
__device__ __forceinline__ uint32_t tf32_lo(float x, uint32_t hi) { return __float_as_uint(x - __uint_as_float(hi)); }

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

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
        auto sample = [&](int m) -> float {                    // pre-emphasised signal at (reflected) index m
            if (m < 0) m = -m; else if (m >= N) m = 2 * (N - 1) - m;
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
        // taps in the warp's idle exchange buffer, then pick them up with static register indices
#pragma unroll 4
        for (int i = lane; i < WIN; i += 32) scratch[i] = sample(n0 + i);
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

struct TileInfo {
    int b, ta, nf, fbase, Tb, N;
    long long wbase;
};

// tile g (global tile index) -> utterance / first frame / sizes; `b` is a running lower bound (monotone in g)
__device__ __forceinline__ TileInfo tile_info(const FrontendParams& P, int g, int& b, int R) {
    while (g >= P.tile_off[b + 1]) ++b;
    TileInfo ti;
    ti.b = b;
    ti.ta = (g - P.tile_off[b]) * R;
    ti.fbase = P.frame_off[b];
    ti.Tb = P.frame_off[b + 1] - ti.fbase;
    ti.nf = min(R, ti.Tb - ti.ta);
    const int s0 = P.bounds[2 * b];
    ti.N = P.bounds[2 * b + 1] - s0;
    ti.wbase = P.wave_off[b] + s0;
    return ti;
}

template <class T, int FW, int RPW, int MW>
__global__ void __launch_bounds__((FW + MW) * 32, 1) frontend_kernel(const FrontendParams P) {
    using C = cplx_t<T>;
    using L = FeLayout<T, FW, RPW, MW>;
    constexpr int R = L::R;
    constexpr int NT = (FW + MW) * 32;
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float* rows = reinterpret_cast<float*>(smem + L::off_rows);
    C* s_win = reinterpret_cast<C*>(smem + L::off_win);
    C* s_tw = reinterpret_cast<C*>(smem + L::off_tw);
    int* s_kb = reinterpret_cast<int*>(smem + L::off_kb);
    uint64_t* bar_full = reinterpret_cast<uint64_t*>(smem + L::off_bar);
    uint64_t* bar_empty = bar_full + 1;
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + L::off_tmem);

    const int total_tiles = P.tile_off[P.B];
    const int g0 = (int)(((long long)total_tiles * blockIdx.x) / gridDim.x);
    const int g1 = (int)(((long long)total_tiles * (blockIdx.x + 1)) / gridDim.x);
    if (g0 >= g1) return;

    for (int i = tid; i < NLIVE_C; i += NT) s_win[i] = reinterpret_cast<const C*>(P.win2)[i];
    for (int i = tid; i < 1024; i += NT) s_tw[i] = reinterpret_cast<const C*>(P.tw1024)[i];
    if constexpr (L::WP_TABLE) {
        C* s_wp = reinterpret_cast<C*>(smem + L::off_wp);
        for (int i = tid; i < 1024; i += NT) s_wp[i] = reinterpret_cast<const C*>(P.wp2)[i];
    }
    const int npairs = P.mel.pair_begin[NMELS / 8];
    for (int i = tid; i < npairs; i += NT) s_kb[i] = P.mel.pair_kb[i];
    for (int i = tid; i < 16 * ROW_STRIDE; i += NT) rows[i] = 0.0f;      // rows >= R are never written
    if (tid == 0) {
        mbar_init(bar_full, FW);
        mbar_init(bar_empty, MW);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    int b;
    {   // binary search: last b with tile_off[b] <= g0
        int lo = 0, hi = P.B;
        while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (P.tile_off[mid] <= g0) lo = mid; else hi = mid; }
        b = lo;
    }
    if (warp == FW) tmem_alloc(s_tmem, 512);                              // whole tensor memory: one CTA per SM
    tmem_fence_before_sync();
    __syncthreads();                                                      // the only CTA-wide barrier
    tmem_fence_after_sync();
    const uint32_t tmem = *s_tmem;

    if (warp < FW) {
        // ================================ FFT warps ================================
        C* xch = reinterpret_cast<C*>(smem + L::off_xch + warp * L::XCH_BYTES);
        auto wp = [&]() {
            if constexpr (L::WP_TABLE) return WpTable<T>{reinterpret_cast<const C*>(smem + L::off_wp)};
            else return WpLane<T>{reinterpret_cast<const C*>(P.wp2)[lane]};
        }();
        for (int g = g0; g < g1; ++g) {
            const TileInfo ti = tile_info(P, g, b, R);
            const int it = g - g0;
#pragma unroll 1
            for (int r = 0; r < RPW; ++r) {
                const int f = warp * RPW + r;                            // row of the tile
                if (f < ti.nf) {
                    const int t = ti.ta + f;
                    C v[32];
                    if (P.wave_is_i16) load_frame<T, true>(v, P.wave, ti.wbase, ti.N, t, s_win, reinterpret_cast<float*>(xch), lane);
                    else load_frame<T, false>(v, P.wave, ti.wbase, ti.N, t, s_win, reinterpret_cast<float*>(xch), lane);
                    warp_fft1024<T, 18>(v, xch, s_tw, lane);
                    T nyq;
                    rfft_post<T>(v, nyq, xch, wp, lane);
                    if (r == 0 && it > 0) mbar_wait(bar_empty, (it - 1) & 1);   // mel warps are done with the previous tile
                    float* row = rows + f * ROW_STRIDE;
                    float* mrow = P.mag_out ? P.mag_out + (size_t)(ti.fbase + t) * NBINS : nullptr;
                    static_for<0, 32>([&](auto jj) {
                        constexpr int j = decltype(jj)::value;
                        const int k = lane + 32 * j;
                        const float re = (float)v[j].x, im = (float)v[j].y;      // complex64 storage of librosa.stft
                        const float pw = fmaf(re, re, im * im);
                        row[k] = sqrt_approx(pw);                                 // np.abs
                        if (mrow) mrow[k] = db_norm_pw(pw);
                    });
                    if (lane == 0 && mrow) mrow[1024] = db_norm(fabsf((float)nyq));
                } else if (r == 0 && it > 0) {
                    mbar_wait(bar_empty, (it - 1) & 1);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_full);
        }
    } else {
        // ================================ mel warps ================================
        // The filterbank (B operand) lives in tensor memory: this warp's pairs, in the order it visits them, at
        // columns 2 lp, 2 lp + 1 of its own 32-lane quadrant -- no shared-memory, L1 or L2 traffic in the tile loop.
        const int mw = warp - FW;
        const int gq = lane >> 2, tq = lane & 3;
        const uint32_t tbase = tmem + ((uint32_t)(32 * (warp & 3)) << 16);
        {
            int lp = 0;
            for (int nt = NMELS / 8 - 1 - mw; nt >= 0; nt -= MW)
                for (int p = P.mel.pair_begin[nt]; p < P.mel.pair_begin[nt + 1]; ++p, ++lp)
                    tmem_st2(tbase + 2 * lp, __ldg(P.mel.bfrag + p * 32 + lane));
            tmem_wait_st();
        }
        const float* a_lo = rows + gq * ROW_STRIDE + tq;
        const float* a_hi = rows + (gq + 8) * ROW_STRIDE + tq;
        // operands of one (k-block, n-tile) pair, split for 3xTF32: x = hi + lo with hi = tf32(x), lo = tf32(x - hi)
        struct PairOps { uint32_t ah[4], al[4], bh[2], bl[2]; };
        auto pair_ops = [&](int kb, float b0, float b1) {
            PairOps o;
            float a[4];
            a[0] = a_lo[8 * kb];
            a[1] = a_hi[8 * kb];
            a[2] = a_lo[8 * kb + 4];
            a[3] = a_hi[8 * kb + 4];
#pragma unroll
            for (int q = 0; q < 4; ++q) { o.ah[q] = tf32_hi(a[q]); o.al[q] = tf32_lo(a[q], o.ah[q]); }
            o.bh[0] = tf32_hi(b0); o.bh[1] = tf32_hi(b1);
            o.bl[0] = tf32_lo(b0, o.bh[0]); o.bl[1] = tf32_lo(b1, o.bh[1]);
            return o;
        };
        for (int g = g0; g < g1; ++g) {
            const TileInfo ti = tile_info(P, g, b, R);
            const int it = g - g0;
            mbar_wait(bar_full, it & 1);
            // n-tiles are dealt heaviest-first: with MW = 2 warp 0 takes {9, 7, 5, ...}, warp 1 {8, 6, 4, ...}
            int lp = 0;
            for (int nt = NMELS / 8 - 1 - mw; nt >= 0; nt -= MW) {
                // four independent accumulator sets: consecutive pairs go to different sets and the MMAs are issued
                // interleaved, so the tensor pipe never waits on its own previous result; summed in a fixed order.
                float acc[4][4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
#pragma unroll
                    for (int q = 0; q < 4; ++q) acc[u][q] = 0.f;
                const int p0 = P.mel.pair_begin[nt], p1 = P.mel.pair_begin[nt + 1];
                int p = p0;
                for (; p + 4 <= p1; p += 4, lp += 4) {
                    float4 q0 = tmem_ld4(tbase + 2 * lp), q1 = tmem_ld4(tbase + 2 * lp + 4);
                    tmem_wait_ld(q0, q1);
                    PairOps o[4] = {pair_ops(s_kb[p], q0.x, q0.y), pair_ops(s_kb[p + 1], q0.z, q0.w),
                                    pair_ops(s_kb[p + 2], q1.x, q1.y), pair_ops(s_kb[p + 3], q1.z, q1.w)};
#pragma unroll
                    for (int u = 0; u < 4; ++u) mma_tf32(acc[u], o[u].al, o[u].bh[0], o[u].bh[1]);
#pragma unroll
                    for (int u = 0; u < 4; ++u) mma_tf32(acc[u], o[u].ah, o[u].bl[0], o[u].bl[1]);
#pragma unroll
                    for (int u = 0; u < 4; ++u) mma_tf32(acc[u], o[u].ah, o[u].bh[0], o[u].bh[1]);
                }
                for (; p < p1; ++p, ++lp) {                               // tail: <= 3 pairs
                    float2 bw = tmem_ld2(tbase + 2 * lp), dummy = bw;
                    tmem_wait_ld(bw, dummy);
                    const PairOps o = pair_ops(s_kb[p], bw.x, bw.y);
                    mma_tf32(acc[0], o.al, o.bh[0], o.bh[1]);
                    mma_tf32(acc[0], o.ah, o.bl[0], o.bl[1]);
                    mma_tf32(acc[0], o.ah, o.bh[0], o.bh[1]);
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[0][q] = (acc[0][q] + acc[1][q]) + (acc[2][q] + acc[3][q]);
                const int m0 = 8 * nt + 2 * tq;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int f = gq + 8 * h;
                    if (f < ti.nf) {
                        const float v0 = db_norm(acc[0][2 * h]), v1 = db_norm(acc[0][2 * h + 1]);
                        if (P.mel_layout == 0) {
                            float* dst = P.mel_out + (size_t)NMELS * ti.fbase + (ti.ta + f);
                            dst[(size_t)m0 * ti.Tb] = v0;
                            dst[(size_t)(m0 + 1) * ti.Tb] = v1;
                        } else {
                            float* dst = P.mel_out + ((size_t)ti.b * P.T_max + ti.ta + f) * NMELS + m0;
                            dst[0] = v0;
                            dst[1] = v1;
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_empty);
        }
        if constexpr (MW > 1) asm volatile("bar.sync 1, %0;" ::"n"(MW * 32) : "memory");
        if (warp == FW) tmem_dealloc(tmem, 512);
    }
}

// dB-normalise and store one 16-frame x 8-mel accumulator tile (mma C-fragment layout) in the requested layout
__device__ __forceinline__ void store_mel_tile(const FrontendParams& P, const TileInfo& ti, int nt, const float (&s4)[4], int gq, int tq) {
    const int m0 = 8 * nt + 2 * tq;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int f = gq + 8 * h;
        if (f < ti.nf) {
            const float v0 = db_norm(s4[2 * h]), v1 = db_norm(s4[2 * h + 1]);
            if (P.mel_layout == 0) {
                float* dst = P.mel_out + (size_t)NMELS * ti.fbase + (ti.ta + f);
                dst[(size_t)m0 * ti.Tb] = v0;
                dst[(size_t)(m0 + 1) * ti.Tb] = v1;
            } else {
                float* dst = P.mel_out + ((size_t)ti.b * P.T_max + ti.ta + f) * NMELS + m0;
                dst[0] = v0;
                dst[1] = v1;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Cooperative variant: every warp is an FFT warp (balanced over the four SM sub-partitions, the whole FP64 pipe is
// used) and the tile's filterbank contraction is shared by all warps between two CTA barriers:
//     [each warp: RPW frames -> rows] -> barrier -> [each warp: its mel units on the tensor cores] -> barrier
// A mel unit is an n-tile (8 mels) or half of one; halves accumulate into a small shared tile with red.shared.add
// (exactly two contributions onto zero: order-independent) that a designated warp finalises after the second
// barrier.  B fragments live in tensor memory: warp w keeps the pairs of its units, in visiting order, at columns
// [128 (w / 4), ...) of quadrant w % 4.
struct MelUnit { int nt, p0, p1, split; };        // pairs [p0, p1) of n-tile nt; split = index of the shared tile or -1

struct MelCoop {                                   // built on the host for a given warp count
    const MelUnit* units;                          // [unit_begin[FW]]
    int unit_begin[17];                            // units of warp w are [unit_begin[w], unit_begin[w + 1])
    int nsplit;                                    // number of split n-tiles (<= 4)
    int split_nt[4];                               // their n-tile
};

constexpr int MEL_MAX_UNITS = 32;

template <class T, int FW, int RPW>
struct FeCoopLayout {
    using C = cplx_t<T>;
    static constexpr int R = FW * RPW;
    static constexpr bool WP_TABLE = sizeof(T) == 4;
    static constexpr int XCH_BYTES = xch_elems<T>() * (int)sizeof(C);
    static constexpr int off_xch = 0;
    static constexpr int off_rows = off_xch + FW * XCH_BYTES;
    static constexpr int off_win = off_rows + 16 * ROW_STRIDE * 4;
    static constexpr int off_tw = off_win + (NLIVE_C * (int)sizeof(C) + 15) / 16 * 16;
    static constexpr int off_wp = off_tw + 1024 * (int)sizeof(C);
    static constexpr int off_kb = off_wp + (WP_TABLE ? 1024 * (int)sizeof(C) : 0);
    static constexpr int off_units = off_kb + MEL_MAX_PAIRS * 4;
    static constexpr int off_part = off_units + MEL_MAX_UNITS * 16;
    static constexpr int off_tmem = off_part + 4 * 128 * 4;
    static constexpr int bytes = off_tmem + 16;
    static_assert(R == 16, "a tile is one m16 MMA tile of frames");
    static_assert(bytes <= 227 * 1024, "shared memory budget");
};

template <class T, int FW, int RPW>
__global__ void __launch_bounds__(FW * 32, 1) frontend_coop_kernel(const FrontendParams P, const MelCoop M) {
    using C = cplx_t<T>;
    using L = FeCoopLayout<T, FW, RPW>;
    constexpr int R = L::R;
    constexpr int NT = FW * 32;
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float* rows = reinterpret_cast<float*>(smem + L::off_rows);
    C* s_win = reinterpret_cast<C*>(smem + L::off_win);
    C* s_tw = reinterpret_cast<C*>(smem + L::off_tw);
    int* s_kb = reinterpret_cast<int*>(smem + L::off_kb);
    MelUnit* s_units = reinterpret_cast<MelUnit*>(smem + L::off_units);
    float* part = reinterpret_cast<float*>(smem + L::off_part);
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + L::off_tmem);

    const int total_tiles = P.tile_off[P.B];
    const int g0 = (int)(((long long)total_tiles * blockIdx.x) / gridDim.x);
    const int g1 = (int)(((long long)total_tiles * (blockIdx.x + 1)) / gridDim.x);
    if (g0 >= g1) return;

    for (int i = tid; i < NLIVE_C; i += NT) s_win[i] = reinterpret_cast<const C*>(P.win2)[i];
    for (int i = tid; i < 1024; i += NT) s_tw[i] = reinterpret_cast<const C*>(P.tw1024)[i];
    if constexpr (L::WP_TABLE) {
        C* s_wp = reinterpret_cast<C*>(smem + L::off_wp);
        for (int i = tid; i < 1024; i += NT) s_wp[i] = reinterpret_cast<const C*>(P.wp2)[i];
    }
    const int npairs = P.mel.pair_begin[NMELS / 8];
    for (int i = tid; i < npairs; i += NT) s_kb[i] = P.mel.pair_kb[i];
    for (int i = tid; i < M.unit_begin[FW]; i += NT) s_units[i] = M.units[i];
    for (int i = tid; i < 4 * 128; i += NT) part[i] = 0.0f;
    if (warp == 0) tmem_alloc(s_tmem, 512);                               // whole tensor memory: one CTA per SM
    int b;
    {   // binary search: last b with tile_off[b] <= g0
        int lo = 0, hi = P.B;
        while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (P.tile_off[mid] <= g0) lo = mid; else hi = mid; }
        b = lo;
    }
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    const uint32_t tbase = *s_tmem + ((uint32_t)(32 * (warp & 3)) << 16) + 128u * (uint32_t)(warp >> 2);   // <= 64 pairs per warp
    const int u0 = M.unit_begin[warp], u1 = M.unit_begin[warp + 1];
    {   // this warp's B fragments -> tensor memory, in visiting order
        int lp = 0;
        for (int u = u0; u < u1; ++u)
            for (int p = s_units[u].p0; p < s_units[u].p1; ++p, ++lp)
                tmem_st2(tbase + 2 * lp, __ldg(P.mel.bfrag + p * 32 + lane));
        tmem_wait_st();
    }

    C* xch = reinterpret_cast<C*>(smem + L::off_xch + warp * L::XCH_BYTES);
    auto wp = [&]() {
        if constexpr (L::WP_TABLE) return WpTable<T>{reinterpret_cast<const C*>(smem + L::off_wp)};
        else return WpLane<T>{reinterpret_cast<const C*>(P.wp2)[lane]};
    }();
    const int gq = lane >> 2, tq = lane & 3;
    const float* a_lo = rows + gq * ROW_STRIDE + tq;
    const float* a_hi = rows + (gq + 8) * ROW_STRIDE + tq;

    for (int g = g0; g < g1; ++g) {
        const TileInfo ti = tile_info(P, g, b, R);
        // ---------------- phase 1: this warp's frames ----------------
#pragma unroll 1
        for (int r = 0; r < RPW; ++r) {
            const int f = warp * RPW + r;
            if (f < ti.nf) {
                const int t = ti.ta + f;
                C v[32];
                if (P.wave_is_i16) load_frame<T, true>(v, P.wave, ti.wbase, ti.N, t, s_win, reinterpret_cast<float*>(xch), lane);
                else load_frame<T, false>(v, P.wave, ti.wbase, ti.N, t, s_win, reinterpret_cast<float*>(xch), lane);
                warp_fft1024<T, 18>(v, xch, s_tw, lane);
                T nyq;
                rfft_post<T>(v, nyq, xch, wp, lane);
                float* row = rows + f * ROW_STRIDE;
                float* mrow = P.mag_out ? P.mag_out + (size_t)(ti.fbase + t) * NBINS : nullptr;
                static_for<0, 32>([&](auto jj) {
                    constexpr int j = decltype(jj)::value;
                    const int k = lane + 32 * j;
                    const float re = (float)v[j].x, im = (float)v[j].y;      // complex64 storage of librosa.stft
                    const float pw = fmaf(re, re, im * im);
                    row[k] = sqrt_approx(pw);                                 // np.abs
                    if (mrow) mrow[k] = db_norm_pw(pw);
                });
                if (lane == 0 && mrow) mrow[1024] = db_norm(fabsf((float)nyq));
            }
        }
        __syncthreads();                                                  // rows of the tile complete
        // ---------------- phase 2: this warp's mel units (3xTF32 mma.sync, banded) ----------------
        int lp = 0;
        for (int u = u0; u < u1; ++u) {
            const MelUnit un = s_units[u];
            float acc[4][4];
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[a][q] = 0.f;
            auto pair_ops = [&](int kb, float b0, float b1, float (&d)[4], int pass) {
                float a[4] = {a_lo[8 * kb], a_hi[8 * kb], a_lo[8 * kb + 4], a_hi[8 * kb + 4]};
                uint32_t ah[4], al[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) { ah[q] = tf32_hi(a[q]); al[q] = tf32_lo(a[q], ah[q]); }
                const uint32_t b0h = tf32_hi(b0), b1h = tf32_hi(b1);
                (void)pass;
                mma_tf32(d, al, b0h, b1h);
                mma_tf32(d, ah, tf32_lo(b0, b0h), tf32_lo(b1, b1h));
                mma_tf32(d, ah, b0h, b1h);
            };
            int p = un.p0;
            for (; p + 4 <= un.p1; p += 4, lp += 4) {
                float4 q0 = tmem_ld4(tbase + 2 * lp), q1 = tmem_ld4(tbase + 2 * lp + 4);
                tmem_wait_ld(q0, q1);
                pair_ops(s_kb[p], q0.x, q0.y, acc[0], 0);
                pair_ops(s_kb[p + 1], q0.z, q0.w, acc[1], 0);
                pair_ops(s_kb[p + 2], q1.x, q1.y, acc[2], 0);
                pair_ops(s_kb[p + 3], q1.z, q1.w, acc[3], 0);
            }
            for (; p < un.p1; ++p, ++lp) {
                float2 bw = tmem_ld2(tbase + 2 * lp), dummy = bw;
                tmem_wait_ld(bw, dummy);
                pair_ops(s_kb[p], bw.x, bw.y, acc[0], 0);
            }
            float s4[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) s4[q] = (acc[0][q] + acc[1][q]) + (acc[2][q] + acc[3][q]);
            if (un.split >= 0) {
                float* dst = part + un.split * 128 + lane * 4;
#pragma unroll
                for (int q = 0; q < 4; ++q) atomicAdd(dst + q, s4[q]);
            } else {
                store_mel_tile(P, ti, un.nt, s4, gq, tq);
            }
        }
        __syncthreads();                                                  // rows free again, split tiles complete
        if (warp < M.nsplit) {
            float* src = part + warp * 128 + lane * 4;
            float s4[4] = {src[0], src[1], src[2], src[3]};
            src[0] = src[1] = src[2] = src[3] = 0.0f;
            store_mel_tile(P, ti, M.split_nt[warp], s4, gq, tq);
        }
    }
    __syncthreads();
    if (warp == 0) tmem_dealloc(*s_tmem, 512);
}

}  // namespace unast
