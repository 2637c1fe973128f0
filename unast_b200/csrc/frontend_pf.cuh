// K1 (v4): fused mel / magnitude front end with more resident warps (replaces reference src/utils.py:251-276).
//
// Same contract and tile structure as frontend_coop_kernel (frontend.cuh): one persistent CTA per SM walks a contiguous
// range of tiles of R = FW consecutive frames of one utterance, every warp transforms one frame of the tile, then all
// warps share the tile's filterbank contraction on the tensor cores between two CTA barriers.  What changed, driven by
// the ncu capture of the previous kernel (profiles/r1c_ncu_fe_f64.txt: FP64 pipe 36 % busy, shared-memory pipe 43 %,
// two warps per scheduler, a third of the time in load / epilogue phases that nothing overlapped):
//
//   * the transposition goes through a split-plane buffer (real plane, then imaginary plane: 8.5 KB per warp for
//     float64 instead of 16.5 KB), which is what lets FW = 12 warps of 168 registers fit one SM (three per scheduler
//     instead of two);
//   * the post-processing computes each (k, 1024 - k) pair once: lane l owns its registers j < 16 (bins l + 32 j) and
//     their mirrors, which lane 32 - l parks in the idle transposition buffer -- 12 instead of 16 flops per bin pair,
//     half the factor products, no shuffles, no lane-0 selects;
//   * the next frame's PCM16 samples are prefetched with cp.async into a warp-private staging buffer while the
//     current frame is transformed (the global-load latency was fully exposed before);
//   * tile bookkeeping is resolved once per CTA, two tiles ahead, into a ring in shared memory;
//   * the CTA barrier after the mel phase is replaced by an mbarrier whose wait sits in the middle of the next
//     transform; split n-tiles are finished by whichever of their two warps arrives second.
#pragma once
#include "frontend.cuh"

namespace unast {

constexpr int PF_STAGE_BYTES = 2304;        // 140 16-byte chunks cover 1104 int16 samples at any alignment (+ pad)
constexpr int PF_SLACK = 24;                // samples of slack after the last tap so that chunk reads stay in bounds

template <class T, int FW>
struct FePfLayout {
    using C = cplx_t<T>;
    static constexpr int R = FW;
    static constexpr int NR = R < 16 ? R + 1 : 16;                      // row R is an all-zero row (MMA rows >= R)
    static constexpr bool WP_TABLE = sizeof(T) == 4;
    static constexpr int XCH_BYTES = (xs_bytes<T>() > WIN * 4 ? xs_bytes<T>() : (WIN * 4 + 15) / 16 * 16);
    static constexpr int off_xch = 0;
    static constexpr int off_stage = off_xch + FW * XCH_BYTES;
    static constexpr int off_rows = off_stage + FW * PF_STAGE_BYTES;
    static constexpr int off_win = off_rows + NR * ROW_STRIDE * 4;
    static constexpr int off_tw = off_win + (NLIVE_C * (int)sizeof(C) + 15) / 16 * 16;
    static constexpr int off_wp = off_tw + 1024 * (int)sizeof(C);
    static constexpr int off_kb = off_wp + (WP_TABLE ? 1024 * (int)sizeof(C) : 0);
    static constexpr int off_units = off_kb + MEL_MAX_PAIRS * 4;
    static constexpr int off_part = off_units + MEL_MAX_UNITS * 16;
    static constexpr int off_ti = off_part + 4 * 128 * 4;                // TileInfo ring [4] + the current utterance record
    static constexpr int off_tmem = off_ti + 5 * 32;
    static constexpr int off_bar = off_tmem + 16;
    static constexpr int off_cnt = off_bar + 16;                        // arrival counters of the split n-tiles
    static constexpr int off_loop = off_cnt + 16;                       // per-warp tile counters (see the tile loop)
    static constexpr int bytes = off_loop + 16 * 4;
    static_assert(R <= 16, "a tile is one m16 MMA tile of frames");
    static_assert(bytes <= 227 * 1024, "shared memory budget");
};

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

// Tile bookkeeping is CTA-uniform: thread 0 resolves tile g + 2 (utterance, first frame, sizes) while tile g is being
// transformed and leaves it in a ring in shared memory, so no warp ever waits on the dependent global loads
// (tile_off -> frame_off / bounds / wave_off) that tile_info() chains.
struct UttRec { int b, tile_lo, tile_hi, fbase, Tb, N; long long wbase; };
static_assert(sizeof(UttRec) == 32 && sizeof(TileInfo) == 32, "ring slots are 32 bytes");

__device__ __forceinline__ void utt_load(const FrontendParams& P, int b, UttRec& u) {
    u.b = b;
    u.tile_lo = P.tile_off[b];
    u.tile_hi = P.tile_off[b + 1];
    u.fbase = P.frame_off[b];
    u.Tb = P.frame_off[b + 1] - u.fbase;
    const int s0 = P.bounds[2 * b];
    u.N = P.bounds[2 * b + 1] - s0;
    u.wbase = P.wave_off[b] + s0;
}
__device__ __forceinline__ TileInfo tile_of(const FrontendParams& P, UttRec& u, int g, int R) {
    while (g >= u.tile_hi) utt_load(P, u.b + 1, u);
    TileInfo ti;
    ti.b = u.b;
    ti.ta = (g - u.tile_lo) * R;
    ti.fbase = u.fbase;
    ti.Tb = u.Tb;
    ti.nf = min(R, u.Tb - ti.ta);
    ti.N = u.N;
    ti.wbase = u.wbase;
    return ti;
}

template <class T, int FW>
__global__ void __launch_bounds__(FW * 32, 1) frontend_pf_kernel(const FrontendParams P, const MelCoop M) {
    using C = cplx_t<T>;
    using L = FePfLayout<T, FW>;
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
    TileInfo* s_ti = reinterpret_cast<TileInfo*>(smem + L::off_ti);      // ring [4], slot g & 3
    UttRec* s_utt = reinterpret_cast<UttRec*>(smem + L::off_ti + 4 * 32);
    uint64_t* bar_empty = reinterpret_cast<uint64_t*>(smem + L::off_bar);
    int* s_cnt = reinterpret_cast<int*>(smem + L::off_cnt);
    volatile int* s_it = reinterpret_cast<volatile int*>(smem + L::off_loop) + warp;

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
    for (int i = tid; i < L::NR * ROW_STRIDE; i += NT) rows[i] = 0.0f;
    if (warp == 0) tmem_alloc(s_tmem, 512);                               // whole tensor memory: one CTA per SM
    if (tid < 4) s_cnt[tid] = 0;
    if (tid == 0) {
        mbar_init(bar_empty, FW);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid == NT - 1) {
        // binary search: last b with tile_off[b] <= g0; then the first two tiles of the ring
        int lo = 0, hi = P.B;
        while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (P.tile_off[mid] <= g0) lo = mid; else hi = mid; }
        UttRec u;
        utt_load(P, lo, u);
        s_ti[g0 & 3] = tile_of(P, u, g0, R);
        if (g0 + 1 < g1) s_ti[(g0 + 1) & 3] = tile_of(P, u, g0 + 1, R);
        *s_utt = u;
    }
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    const uint32_t tbase = *s_tmem + ((uint32_t)(32 * (warp & 3)) << 16) + 128u * (uint32_t)(warp >> 2);   // <= 31 pairs per warp
    const int u0 = M.unit_begin[warp], u1 = M.unit_begin[warp + 1];
    {   // this warp's B fragments -> tensor memory, in visiting order, already split for 3xTF32: 4 columns per pair
        int lp = 0;
        for (int u = u0; u < u1; ++u)
            for (int p = s_units[u].p0; p < s_units[u].p1; ++p, ++lp) {
                const float2 bw = __ldg(P.mel.bfrag + p * 32 + lane);
                const uint32_t h0 = tf32_hi(bw.x), h1 = tf32_hi(bw.y);
                tmem_st4(tbase + 4 * lp, make_float4(__uint_as_float(h0), __uint_as_float(h1),
                                                     bw.x - __uint_as_float(h0), bw.y - __uint_as_float(h1)));
            }
        tmem_wait_st();
    }

    T* xs = reinterpret_cast<T*>(smem + L::off_xch + warp * L::XCH_BYTES);
    unsigned char* stage = smem + L::off_stage + warp * PF_STAGE_BYTES;
    auto wp = [&]() {
        if constexpr (L::WP_TABLE) return WpTable<T>{reinterpret_cast<const C*>(smem + L::off_wp)};
        else return WpLane<T>{reinterpret_cast<const C*>(P.wp2)[lane]};
    }();
    const int gq = lane >> 2, tq = lane & 3;
    const float* a_lo = rows + gq * ROW_STRIDE + tq;
    const float* a_hi = rows + (gq + 8 < L::NR ? gq + 8 : L::NR - 1) * ROW_STRIDE + tq;   // rows >= R: the zero row

    const int f = warp;                                                   // this warp's row of every tile
    // is this warp's frame of tile `ti` served by the prefetch path?  (recomputed where needed: no flags live across
    // the transform)
    auto prefetched = [&](const TileInfo& ti) { return P.wave_is_i16 && f < ti.nf && pf_interior(ti.ta + f, ti.N); };
    {
        const TileInfo ti = s_ti[g0 & 3];
        if (prefetched(ti)) pf_issue(P.wave, ti.wbase, ti.ta + f, stage, lane);
    }

    // The loop state is CTA-uniform but, next to 128 registers of frame data, the register allocator spills it to
    // local memory and every reload stalls on the long scoreboard.  Keeping the warp's tile counter in shared memory
    // and re-reading it where it is needed (volatile) keeps it out of the long live ranges.
    if (lane == 0) *s_it = 0;
    __syncwarp();
    const int n_tiles = g1 - g0;
    for (int it = 0; it < n_tiles; it = *s_it) {
        const int g = g0 + it;
        if (tid == NT - 1 && g + 2 < g1) {                                // resolve tile g + 2 (read after two barriers)
            UttRec u = *s_utt;
            s_ti[(g + 2) & 3] = tile_of(P, u, g + 2, R);
            *s_utt = u;
        }
        // ---------------- phase 1: this warp's frame ----------------
        {
            const TileInfo cur = s_ti[g & 3];
            C v[32];
            const bool live = f < cur.nf;
            const int t = cur.ta + f;
            if (live) {
                if (prefetched(cur)) {
                    cp_async_wait_all();
                    __syncwarp();
                    const long long A = cur.wbase + (HOP * t + WIN_LPAD - HALF) - 1;
                    const int s0 = (int)(A & 7);                          // sample offset of A inside the first chunk
                    window_i16_words<T>(v, reinterpret_cast<const uint32_t*>(stage) + (s0 >> 1) + lane, (s0 & 1) != 0, s_win, lane);
                    __syncwarp();
                } else if (P.wave_is_i16) {
                    load_frame<T, true>(v, P.wave, cur.wbase, cur.N, t, s_win, reinterpret_cast<float*>(xs), lane);
                } else {
                    load_frame<T, false>(v, P.wave, cur.wbase, cur.N, t, s_win, reinterpret_cast<float*>(xs), lane);
                }
            }
            if (g + 1 < g1) {                                             // next tile's frame: prefetch while this one is transformed
                const TileInfo nxt = s_ti[(g + 1) & 3];
                if (prefetched(nxt)) pf_issue(P.wave, nxt.wbase, nxt.ta + f, stage, lane);
            }
            float* row = rows + f * ROW_STRIDE;
            float* mrow = P.mag_out ? P.mag_out + (size_t)(cur.fbase + t) * NBINS : nullptr;
            // Rows are about to be overwritten: every warp must have finished its mel share of the previous tile.  It
            // arrived on an mbarrier long ago, so this wait is normally free (unlike a second CTA barrier); it sits in
            // the middle of the transform, where the fewest registers are live.
            auto rows_free = [&]() { const int i2 = *s_it; if (i2 > 0) mbar_wait_inline(bar_empty, (uint32_t)(i2 - 1) & 1u); };
            if (live) warp_fft1024_split<T, 18>(v, xs, s_tw, lane, rows_free);
            else rows_free();
            if (live) {
                // power of every bin first (the float64 registers die pair by pair), then the float32 epilogue with 33
                // independent chains in flight
                float pa[16], pb[16], p512 = 0.0f;
                rfft_post_pairs<T>(v, reinterpret_cast<C*>(xs), wp, lane, [&](auto jj, C xa, C xb) {
                    constexpr int j = decltype(jj)::value;
                    const float ar = (float)xa.x, ai = (float)xa.y;      // complex64 storage of librosa.stft
                    const float br = (float)xb.x, bi = (float)xb.y;
                    pa[j] = fmaf(ar, ar, ai * ai);
                    pb[j] = fmaf(br, br, bi * bi);
                }, [&](C x) { const float r = (float)x.x, i = (float)x.y; p512 = fmaf(r, r, i * i); });
                float* rowa = row + lane;
                float* rowb = row + 1024 - lane;
                static_for<0, 16>([&](auto jj) {
                    constexpr int j = decltype(jj)::value;
                    rowa[32 * j] = sqrt_approx(pa[j]);                    // np.abs
                    rowb[-32 * j] = sqrt_approx(pb[j]);
                });
                if (lane == 0) row[512] = sqrt_approx(p512);
                if (mrow) {
                    float* ma = mrow + lane;
                    float* mb = mrow + 1024 - lane;
                    static_for<0, 16>([&](auto jj) {
                        constexpr int j = decltype(jj)::value;
                        ma[32 * j] = db_norm_pw_nf(pa[j]);
                        mb[-32 * j] = db_norm_pw_nf(pb[j]);
                    });
                    if (lane == 0) mrow[512] = db_norm_pw_nf(p512);
                }
            }
        }
        __syncthreads();                                                  // rows of the tile complete
        // ---------------- phase 2: this warp's mel units (3xTF32 mma.sync, banded) ----------------
        const int it2 = *s_it;
        const TileInfo cur = s_ti[(g0 + it2) & 3];
        int lp = 0;
        for (int u = u0; u < u1; ++u) {
            const MelUnit un = s_units[u];
            float acc[4][4];
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[a][q] = 0.f;
            // B operands (hi0, hi1, lo0, lo1 per pair) come from tensor memory a chunk of four pairs ahead of their use
            auto pair_ops = [&](int kb, const float4 bq, float (&d)[4]) {
                float a[4] = {a_lo[8 * kb], a_hi[8 * kb], a_lo[8 * kb + 4], a_hi[8 * kb + 4]};
                uint32_t ah[4], al[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) { ah[q] = tf32_hi(a[q]); al[q] = tf32_lo(a[q], ah[q]); }
                mma_tf32(d, al, __float_as_uint(bq.x), __float_as_uint(bq.y));
                mma_tf32(d, ah, __float_as_uint(bq.z), __float_as_uint(bq.w));
                mma_tf32(d, ah, __float_as_uint(bq.x), __float_as_uint(bq.y));
            };
            auto load4 = [&](float4 (&q)[4], int l) {
#pragma unroll
                for (int i = 0; i < 4; ++i) q[i] = tmem_ld4(tbase + 4 * (l + i));
            };
            float4 nb[4];
            load4(nb, lp);
            for (int p = un.p0; p < un.p1; p += 4, lp += 4) {
                tmem_wait_ld(nb[0], nb[1], nb[2], nb[3]);
                const float4 b0 = nb[0], b1 = nb[1], b2 = nb[2], b3 = nb[3];
                if (p + 4 < un.p1) load4(nb, lp + 4);
                pair_ops(s_kb[p], b0, acc[0]);
                if (p + 1 < un.p1) pair_ops(s_kb[p + 1], b1, acc[1]);
                if (p + 2 < un.p1) pair_ops(s_kb[p + 2], b2, acc[2]);
                if (p + 3 < un.p1) pair_ops(s_kb[p + 3], b3, acc[3]);
            }
            lp -= (4 - ((un.p1 - un.p0) & 3)) & 3;                        // the last chunk may have been partial
            float s4[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) s4[q] = (acc[0][q] + acc[1][q]) + (acc[2][q] + acc[3][q]);
            if (un.split >= 0) {
                // half of an n-tile: add into the shared tile (two contributions onto zero: order-independent); the
                // warp that arrives second turns the sum into mel and clears the tile for the next one
                float* dst = part + un.split * 128 + lane * 4;
#pragma unroll
                for (int q = 0; q < 4; ++q) atomicAdd(dst + q, s4[q]);
                __threadfence_block();
                __syncwarp();
                int old = 0;
                if (lane == 0) old = atomicAdd(s_cnt + un.split, 1);
                old = __shfl_sync(0xffffffffu, old, 0);
                if (old == 1) {
                    __threadfence_block();
                    float t4[4] = {dst[0], dst[1], dst[2], dst[3]};
                    dst[0] = dst[1] = dst[2] = dst[3] = 0.0f;
                    if (lane == 0) s_cnt[un.split] = 0;
                    store_mel_tile(P, cur, un.nt, t4, gq, tq);
                }
            } else {
                store_mel_tile(P, cur, un.nt, s4, gq, tq);
            }
        }
        __syncwarp();
        if (lane == 0) { mbar_arrive(bar_empty); *s_it = it2 + 1; }       // this warp is done with the tile's rows
        __syncwarp();
    }
    __syncthreads();
    __syncthreads();
    if (warp == 0) tmem_dealloc(*s_tmem, 512);
}

}  // namespace unast
