// K1 (experiment, not the default: -DUNAST_FE64_WS=1): warp-specialised fused mel / magnitude front end (replaces
// reference src/utils.py:251-276).  Parity-green on B200 but slower than frontend_pf_kernel (40.4 ms on C2) in every
// configuration tried: 12 + 4 warps with setmaxnreg 45.5 ms (profiles/r1n_ncu_fe_f64_ws.txt: the FFT warps get 152
// registers instead of 168 and spill, 91 local loads per frame, and the leading FFT warps still wait for the slowest
// warp plus the mel warps' latency), 60 ms with the wait moved to the end of the frame (more spills); 10 + 2 warps at
// 168 registers without setmaxnreg 50.4 ms (two mel warps cannot keep up with ten FFT warps: the 3xTF32 contraction
// is a latency chain of ~6,000 warp instructions per tile).
//
// frontend_pf_kernel (frontend_pf.cuh) stops all twelve FFT warps twice per tile: at the CTA barrier that waits for the
// slowest of them, and during the mel phase, when the FP64 pipe idles (about a quarter of its stall samples,
// profiles/r1j).  Here the filterbank contraction belongs to a fourth warp group:
//
//   warps 0..11  (three warp groups, setmaxnreg -> 152 registers): one frame of every 12-frame tile each, end to end,
//                exactly as in frontend_pf_kernel (prefetch, split-plane FFT, mirror-pair post-processing, epilogue);
//                a warp arrives on `full` when its |X| row is in shared memory and goes straight on to its next frame;
//   warps 12..15 (one warp group, setmaxnreg -> 56 registers): wait on `full`, contract the tile with the filterbank
//                on the tensor cores (3xTF32 mma.sync, operands pre-split in TMEM), store mel, arrive on `empty`;
//                one of their threads also resolves the tile bookkeeping three tiles ahead.
//
// The row tile is single-buffered: an FFT warp overwrites its row only at the very end of its next frame, and checks
// `empty` in the middle of that transform (where only 64 registers are live) -- by then the mel warps, which need a
// tenth of a tile time, have long finished.  There is no CTA-wide barrier in the steady state, and a fast FFT warp
// may run most of a tile ahead of a slow one.
#pragma once
#include "frontend_pf.cuh"

namespace unast {

#ifndef UNAST_WS_FW
#define UNAST_WS_FW 10
#endif
#ifndef UNAST_WS_MW
#define UNAST_WS_MW 2
#endif
constexpr int WS_FW = UNAST_WS_FW;          // FFT warps
constexpr int WS_MW = UNAST_WS_MW;          // mel warps
// With 12 + 4 warps the register file only works out with setmaxnreg (152 / 56 registers, warp-group granularity);
// with 10 + 2 every warp keeps the 168 registers of a 12-warp CTA and no reallocation is needed.
constexpr bool WS_SETREG = (WS_FW + WS_MW) * 32 * 168 > 65536;
constexpr int WS_REGS_FFT = 152, WS_REGS_MEL = 56;
static_assert(!WS_SETREG || (WS_FW % 4 == 0 && WS_MW % 4 == 0 && WS_FW * 32 * WS_REGS_FFT + WS_MW * 32 * WS_REGS_MEL <= 65536), "register file");

template <class T>
struct FeWsLayout : FePfLayout<T, WS_FW> {
    using Base = FePfLayout<T, WS_FW>;
    static constexpr int off_bar2 = Base::bytes;            // `full` barrier (Base::off_bar holds `empty`)
    static constexpr int bytes = off_bar2 + 16;
    static_assert(bytes <= 227 * 1024, "shared memory budget");
};

template <int N> __device__ __forceinline__ void reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> __device__ __forceinline__ void reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

template <class T>
__global__ void __launch_bounds__((WS_FW + WS_MW) * 32, 1) frontend_ws_kernel(const FrontendParams P, const MelCoop M) {
    using C = cplx_t<T>;
    using L = FeWsLayout<T>;
    constexpr int R = WS_FW;
    constexpr int NT = (WS_FW + WS_MW) * 32;
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
    uint64_t* bar_empty = reinterpret_cast<uint64_t*>(smem + L::off_bar);
    uint64_t* bar_full = reinterpret_cast<uint64_t*>(smem + L::off_bar2);
    int* s_cnt = reinterpret_cast<int*>(smem + L::off_cnt);

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
    for (int i = tid; i < M.unit_begin[WS_MW]; i += NT) s_units[i] = M.units[i];
    for (int i = tid; i < 4 * 128; i += NT) part[i] = 0.0f;
    for (int i = tid; i < L::NR * ROW_STRIDE; i += NT) rows[i] = 0.0f;
    if (warp == WS_FW) tmem_alloc(s_tmem, 512);                           // whole tensor memory: one CTA per SM
    if (tid < 4) s_cnt[tid] = 0;
    if (tid == 0) {
        mbar_init(bar_empty, WS_MW);
        mbar_init(bar_full, WS_FW);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    UttRec utt;                                                           // live in thread NT - 1 only
    if (tid == NT - 1) {
        // binary search: last b with tile_off[b] <= g0; then the first three tiles of the ring
        int lo = 0, hi = P.B;
        while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (P.tile_off[mid] <= g0) lo = mid; else hi = mid; }
        utt_load(P, lo, utt);
        for (int g = g0; g < g1 && g < g0 + 3; ++g) s_ti[g & 3] = tile_of(P, utt, g, R);
    }
    tmem_fence_before_sync();
    __syncthreads();                                                      // the only CTA-wide barrier
    tmem_fence_after_sync();
    const int n_tiles = g1 - g0;

    if (warp < WS_FW) {
        // ======================================= FFT warps =======================================
        if constexpr (WS_SETREG) reg_inc<WS_REGS_FFT>();
        T* xs = reinterpret_cast<T*>(smem + L::off_xch + warp * L::XCH_BYTES);
        unsigned char* stage = smem + L::off_stage + warp * PF_STAGE_BYTES;
        auto wp = [&]() {
            if constexpr (L::WP_TABLE) return WpTable<T>{reinterpret_cast<const C*>(smem + L::off_wp)};
            else return WpLane<T>{reinterpret_cast<const C*>(P.wp2)[lane]};
        }();
        const int f = warp;                                               // this warp's row of every tile
        auto prefetched = [&](const TileInfo& ti) { return P.wave_is_i16 && f < ti.nf && pf_interior(ti.ta + f, ti.N); };
        {
            const TileInfo ti = s_ti[g0 & 3];
            if (prefetched(ti)) pf_issue(P.wave, ti.wbase, ti.ta + f, stage, lane);
        }
        for (int it = 0; it < n_tiles; ++it) {
            const int g = g0 + it;
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
            if (it + 1 < n_tiles) {                                       // next tile's frame: prefetch while this one is transformed
                const TileInfo nxt = s_ti[(g + 1) & 3];
                if (prefetched(nxt)) pf_issue(P.wave, nxt.wbase, nxt.ta + f, stage, lane);
            }
            float* row = rows + f * ROW_STRIDE;
            float* mrow = P.mag_out ? P.mag_out + (size_t)(cur.fbase + t) * NBINS : nullptr;
            // this warp's row may be overwritten once the mel warps are done with the previous tile: checked in the
            // middle of the transform (fewest live registers), normally long satisfied
            if (live) warp_fft1024_split<T, 18>(v, xs, s_tw, lane);
            if (live) {
                float pa[16], pb[16], p512 = 0.0f;
                rfft_post_pairs<T>(v, reinterpret_cast<C*>(xs), wp, lane, [&](auto jj, C xa, C xb) {
                    constexpr int j = decltype(jj)::value;
                    const float ar = (float)xa.x, ai = (float)xa.y;      // complex64 storage of librosa.stft
                    const float br = (float)xb.x, bi = (float)xb.y;
                    pa[j] = fmaf(ar, ar, ai * ai);
                    pb[j] = fmaf(br, br, bi * bi);
                }, [&](C x) { const float r = (float)x.x, i = (float)x.y; p512 = fmaf(r, r, i * i); });
                // this warp's row may be overwritten once the mel warps are done with the previous tile: checked here,
                // at the very end of the frame, where only the 33 powers are live -- the mel warps have had almost a
                // whole tile time
                if (it > 0) mbar_wait_inline(bar_empty, (uint32_t)(it - 1) & 1u);
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
            // a warp without a frame in this tile still has to see `empty` of the previous tile before it arrives, or
            // it could arrive twice in one phase of `full`
            if (!live && it > 0) mbar_wait_inline(bar_empty, (uint32_t)(it - 1) & 1u);
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_full);                         // this warp's row of the tile is in place
        }
        return;
    }

    // ======================================= mel warps =======================================
    if constexpr (WS_SETREG) reg_dec<WS_REGS_MEL>();
    const int mw = warp - WS_FW;
    const int gq = lane >> 2, tq = lane & 3;
    const float* a_lo = rows + gq * ROW_STRIDE + tq;
    const float* a_hi = rows + (gq + 8 < L::NR ? gq + 8 : L::NR - 1) * ROW_STRIDE + tq;   // rows >= R: the zero row
    const uint32_t tbase = *s_tmem + ((uint32_t)(32 * (warp & 3)) << 16);                    // own quadrant, <= 127 pairs
    static_assert(WS_MW <= 4, "one TMEM lane quadrant per mel warp");
    const int u0 = M.unit_begin[mw], u1 = M.unit_begin[mw + 1];
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
    for (int it = 0; it < n_tiles; ++it) {
        const int g = g0 + it;
        mbar_wait(bar_full, (uint32_t)it & 1u);                           // every row of the tile is in place
        // Tile g + 3 goes into the slot of tile g - 1, which no FFT warp reads any more (all of them have finished
        // tile g); it is published by this warp's arrival on `empty`, which every FFT warp awaits before tile g + 2.
        if (tid == NT - 1 && g + 3 < g1) s_ti[(g + 3) & 3] = tile_of(P, utt, g + 3, R);
        const TileInfo cur = s_ti[g & 3];
        int lp = 0;
        for (int u = u0; u < u1; ++u) {
            const MelUnit un = s_units[u];
            float acc[4][4];
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[a][q] = 0.f;
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
        __threadfence_block();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_empty);                            // rows (and the new ring slot) are released
    }
    asm volatile("bar.sync 1, %0;" ::"n"(WS_MW * 32) : "memory");
    if (warp == WS_FW) tmem_dealloc(*s_tmem, 512);
}

}  // namespace unast
