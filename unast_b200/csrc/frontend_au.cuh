// K1 (autonomous): the fused front end of frontend_pf.cuh with NO CTA-wide phase (reference src/utils.py:251-276).
//
// frontend_pf_kernel ends every tile of 12 frames in a CTA barrier and a tensor-core mel phase during which the FP64
// pipe idles (profiles/r1m_ncu_fe_f64.txt: 32 % of the samples, 27 % of them barrier stalls, 693 warp instructions per
// frame of tile bookkeeping, fragment splitting and mma.sync).  Here every warp owns a contiguous run of frames (an
// equal share of all frames, located by one binary search) and finishes each frame alone:
//     samples (cp.async prefetch) -> window -> float64 FFT -> |X| row in its own shared-memory row -> mag row out
//     -> filterbank from that row in plain FP32 FMAs -> 80 mels out
// so warps drift apart and one warp's filterbank / epilogue overlaps the others' FP64 transforms.
//
// Filterbank without tensor cores: the 80 x 1025 Slaney matrix has 2,001 non-zeros in contiguous runs of 2..83 bins per
// mel.  The host cuts every run into ITEMS of 8 bins starting at a multiple of 4 (zero weights pad the ends), deals
// them to (round, lane) slots and lays their weights out so that one item is two 128-bit loads of |X|, two of weights
// and 8 FMAs.  A mel is the sum of its items' partial sums, added in item order (fixed: deterministic).  About 300
// items -> 10 rounds per frame.
#pragma once
#include "frontend_pf.cuh"

namespace unast {

#ifndef UNAST_FE_TMEM
#define UNAST_FE_TMEM 1                      // float64 mode: the [k1][lane] twiddle table lives in tensor memory
#endif

constexpr int AU_ROW = 1032;                 // floats per |X| row: bins 0..1024 + zero padding up to the last item's end
constexpr int AU_MEL_CAP0 = 2, AU_MEL_CAP1 = 6, AU_MEL_CAP2 = 12;   // most items of a mel in 0..31 / 32..63 / 64..79 (checked by the host)
constexpr int AU_BATCH = 5;                  // rounds per batch; the host pads the tables to a multiple of it
constexpr int AU_MAX_ROUNDS = 15;

struct MelItems {                            // device arrays, built once on the host (unast_spectral.cu)
    const float4* w;                         // [rounds][2][32]: weights of bins k0..k0+3 / k0+4..k0+7 of slot (round, lane)
    const int* k0;                           // [rounds][32]: first bin of the slot's item (multiple of 4); idle slots: 0, weights 0
    const int2* mel;                         // [80]: (first item id = round * 32 + lane, number of items)
    int rounds;                              // multiple of AU_BATCH (padding slots: k0 = 0, weights 0)
};

template <class T, int FW>
struct FeAuLayout {
    using C = cplx_t<T>;
    static constexpr bool WP_TABLE = sizeof(T) == 4;
    static constexpr int XCH_BYTES = (xs_bytes<T>() > WIN * 4 ? xs_bytes<T>() : (WIN * 4 + 15) / 16 * 16);
    static constexpr int off_xch = 0;
    static constexpr int off_stage = off_xch + FW * XCH_BYTES;
    static constexpr int off_rows = off_stage + FW * PF_STAGE_BYTES;
    static constexpr int off_win = off_rows + FW * AU_ROW * 4;
    static constexpr int off_tw = off_win + (NLIVE_C * (int)sizeof(C) + 15) / 16 * 16;
    static constexpr int off_wp = off_tw + 1024 * (int)sizeof(C);
    static constexpr int off_w = off_wp + (WP_TABLE ? 1024 * (int)sizeof(C) : 0);
    static constexpr int off_k0 = off_w + AU_MAX_ROUNDS * 64 * 16;
    static constexpr int off_mel = off_k0 + AU_MAX_ROUNDS * 32 * 4;
    static constexpr int off_state = off_mel + NMELS * 8;                // per-warp loop state, 8 words each
    static constexpr int bytes = off_state + FW * 32 + 16;               // + the TMEM base address word
    static_assert(XCH_BYTES >= AU_MAX_ROUNDS * 32 * 4, "partial sums live in the idle exchange buffer");
    static_assert(bytes <= 227 * 1024, "shared memory budget");
    static_assert(off_state % 16 == 0, "the loop state is read with 128-bit loads");
};

template <class T, int FW>
__global__ void __launch_bounds__(FW * 32, 1) frontend_au_kernel(const FrontendParams P, const MelItems MI) {
    using C = cplx_t<T>;
    using L = FeAuLayout<T, FW>;
    constexpr int NT = FW * 32;
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    C* s_win = reinterpret_cast<C*>(smem + L::off_win);
    C* s_tw = reinterpret_cast<C*>(smem + L::off_tw);
    float4* s_w = reinterpret_cast<float4*>(smem + L::off_w);
    int* s_k0 = reinterpret_cast<int*>(smem + L::off_k0);
    int2* s_mel = reinterpret_cast<int2*>(smem + L::off_mel);
    float* row = reinterpret_cast<float*>(smem + L::off_rows) + warp * AU_ROW;
    T* xs = reinterpret_cast<T*>(smem + L::off_xch + warp * L::XCH_BYTES);
    unsigned char* stage = smem + L::off_stage + warp * PF_STAGE_BYTES;

    for (int i = tid; i < NLIVE_C; i += NT) s_win[i] = reinterpret_cast<const C*>(P.win2)[i];
    constexpr bool TW_TMEM = UNAST_FE_TMEM && sizeof(T) == 8;
    uint32_t tq = 0;
    if constexpr (TW_TMEM) {
        // The inter-pass twiddle table W1024^(lane k1) goes to tensor memory (31 complex doubles per lane = columns [0, 128),
        // one copy per lane quadrant) and is read with tcgen05.ld in the middle of the float64 phase: 35.4 -> 32.3 ms on C2
        // (profiles/r3_fe_variants.txt).  Moving the window taps and the filterbank weights there as well changed nothing
        // (32.6 ms): it is the latency of the twiddle loads inside the FP64 critical path that mattered, not the bandwidth.
        uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + L::off_state + FW * 32);
        if (warp == 0) tmem_alloc(s_tmem, 128);
        tmem_fence_before_sync();
        __syncthreads();
        tmem_fence_after_sync();
        tq = *s_tmem + ((uint32_t)(32 * (warp & 3)) << 16);
        if (warp < 4) {
            for (int k1 = 0; k1 < 32; ++k1) {
                const double2 w = reinterpret_cast<const double2*>(P.tw1024)[k1 * 32 + lane];
                tmem_st4(tq + 4 * k1, make_float4(__int_as_float(__double2loint(w.x)), __int_as_float(__double2hiint(w.x)),
                                                  __int_as_float(__double2loint(w.y)), __int_as_float(__double2hiint(w.y))));
            }
            tmem_wait_st();
        }
        tmem_fence_before_sync();
    } else {
        for (int i = tid; i < 1024; i += NT) s_tw[i] = reinterpret_cast<const C*>(P.tw1024)[i];
    }
    if constexpr (L::WP_TABLE) {
        C* s_wp = reinterpret_cast<C*>(smem + L::off_wp);
        for (int i = tid; i < 1024; i += NT) s_wp[i] = reinterpret_cast<const C*>(P.wp2)[i];
    }
    const int rounds = MI.rounds;
    for (int i = tid; i < rounds * 64; i += NT) s_w[i] = MI.w[i];
    for (int i = tid; i < rounds * 32; i += NT) s_k0[i] = MI.k0[i];
    for (int i = tid; i < NMELS; i += NT) s_mel[i] = MI.mel[i];
    for (int i = tid; i < FW * AU_ROW; i += NT) reinterpret_cast<float*>(smem + L::off_rows)[i] = 0.0f;   // incl. the padding
    __syncthreads();                                                      // the only CTA-wide barrier of the steady state
    if constexpr (TW_TMEM) tmem_fence_after_sync();

    auto wp = [&]() {
        if constexpr (L::WP_TABLE) return WpTable<T>{reinterpret_cast<const C*>(smem + L::off_wp)};
        else return WpLane<T>{reinterpret_cast<const C*>(P.wp2)[lane]};
    }();

    // this warp's run of frames: an equal share of all frames of the batch
    const long long F = P.frame_off[P.B];
    const long long NW = (long long)gridDim.x * FW, gw = (long long)blockIdx.x * FW + warp;
    const int fa = (int)(F * gw / NW), fb = (int)(F * (gw + 1) / NW);
    auto release_tmem = [&]() {                                           // tensor memory is released when every warp is done
        if constexpr (TW_TMEM) {
            tmem_fence_before_sync();
            asm volatile("bar.sync 2, %0;" ::"n"(NT) : "memory");
            if (warp == 0) tmem_dealloc(*reinterpret_cast<uint32_t*>(smem + L::off_state + FW * 32), 128);
        }
    };
    if (fa >= fb) { release_tmem(); return; }
    int b;
    {   // last b with frame_off[b] <= fa
        int lo = 0, hi = P.B;
        while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (P.frame_off[mid] <= fa) lo = mid; else hi = mid; }
        b = lo;
    }
    // Loop state lives in shared memory, re-read (volatile) where it is needed: next to 128 registers of frame data the
    // register allocator would spill it to local memory, and every reload then stalls on the long scoreboard
    // (profiles/r2c_ncu_fe_f64.txt: ~9 % of the samples sat on such reloads).
    //   st[0] f, st[1] fb, st[2] fend, st[3] fbase, st[4] N, st[5] b, st[6..7] wbase
    volatile int* st = reinterpret_cast<volatile int*>(smem + L::off_state) + warp * 8;
    auto utt = [&](int bb) {                                              // lane 0 only
        st[3] = P.frame_off[bb];
        st[2] = P.frame_off[bb + 1];
        const int s0 = P.bounds[2 * bb];
        st[4] = P.bounds[2 * bb + 1] - s0;
        st[5] = bb;
        *reinterpret_cast<volatile long long*>(st + 6) = P.wave_off[bb] + s0;
    };
    if (lane == 0) { st[0] = fa; st[1] = fb; utt(b); }
    __syncwarp();
    auto wbase_of = [&]() { return *reinterpret_cast<volatile long long*>(st + 6); };
    auto prefetched = [&](int t) { return P.wave_is_i16 && pf_interior(t, st[4]); };
    if (prefetched(fa - st[3])) pf_issue(P.wave, wbase_of(), fa - st[3], stage, lane);

    float* partial = reinterpret_cast<float*>(xs);                        // item sums: the exchange buffer is idle then
    // the eight state words in two 128-bit loads: one shared-memory latency at the top of a frame instead of a chain of them
    auto ld_state = [&](int4& a, int4& b2) {
        const uint32_t sp = smem_u32(const_cast<int*>(st));
        asm volatile("ld.volatile.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w) : "r"(sp) : "memory");
        asm volatile("ld.volatile.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(b2.x), "=r"(b2.y), "=r"(b2.z), "=r"(b2.w) : "r"(sp + 16) : "memory");
    };
#pragma unroll 1
    for (;;) {
        int4 sa, sb;                                                      // (f, fb, fend, fbase), (N, b, wbase lo, wbase hi)
        ld_state(sa, sb);
        if (sa.x >= sa.y) break;
        if (sa.x >= sa.z) {                                               // the run crosses into the next utterance
            __syncwarp();
            if (lane == 0) { int bb = sb.y; do { ++bb; } while (sa.x >= P.frame_off[bb + 1]); utt(bb); }
            __syncwarp();
            ld_state(sa, sb);
        }
        C v[32];
        {
            const int t = sa.x - sa.w;
            const int N = sb.x;
            const long long wbase = (long long)(((unsigned long long)(unsigned)sb.w << 32) | (unsigned)sb.z);
            const bool pf_now = P.wave_is_i16 && pf_interior(t, N);
            const bool pf_next = sa.x + 1 < sa.y && sa.x + 1 < sa.z && P.wave_is_i16 && pf_interior(t + 1, N);
            if (pf_now) {
                cp_async_wait_all();
                __syncwarp();
                const long long A = wbase + (HOP * t + WIN_LPAD - HALF) - 1;
                const int s0 = (int)(A & 7);                              // sample offset of A inside the first chunk
                window_i16_words<T>(v, reinterpret_cast<const uint32_t*>(stage) + (s0 >> 1) + lane, (s0 & 1) != 0, s_win, lane);
                __syncwarp();
            } else if (P.wave_is_i16) {
                load_frame<T, true>(v, P.wave, wbase, N, t, s_win, reinterpret_cast<float*>(xs), lane);
            } else {
                load_frame<T, false>(v, P.wave, wbase, N, t, s_win, reinterpret_cast<float*>(xs), lane);
            }
            // next frame of the same utterance: its samples land while this one is transformed (the first frames of
            // an utterance are edge frames and never prefetched, so nothing is lost at a boundary)
            if (pf_next) pf_issue(P.wave, wbase, t + 1, stage, lane);
        }

        if constexpr (TW_TMEM) warp_fft1024_split<T, 18>(v, xs, TwTmemF64{tq}, lane);
        else warp_fft1024_split<T, 18>(v, xs, TwMem<T>{s_tw, lane}, lane);

        float* mrow = P.mag_out ? P.mag_out + (size_t)st[0] * NBINS : nullptr;
        float pa[16], pb[16], p512 = 0.0f;
        {
            rfft_post_pairs<T>(v, reinterpret_cast<C*>(xs), wp, lane, [&](auto jj, C xa, C xb) {
                constexpr int j = decltype(jj)::value;
                const float ar = (float)xa.x, ai = (float)xa.y;          // complex64 storage of librosa.stft
                const float br = (float)xb.x, bi = (float)xb.y;
                pa[j] = fmaf(ar, ar, ai * ai);
                pb[j] = fmaf(br, br, bi * bi);
            }, [&](C x) { const float r = (float)x.x, i = (float)x.y; p512 = fmaf(r, r, i * i); });
            float* rowa = row + lane;
            float* rowb = row + 1024 - lane;
            static_for<0, 16>([&](auto jj) {
                constexpr int j = decltype(jj)::value;
                rowa[32 * j] = sqrt_approx(pa[j]);                        // np.abs
                rowb[-32 * j] = sqrt_approx(pb[j]);
            });
            if (lane == 0) row[512] = sqrt_approx(p512);
        }
        // ---- filterbank: items of 8 bins, then the mels in item order ----
        // batches of AU_BATCH rounds: all addresses, then all 128-bit loads, then two FMA chains per item -- the
        // shared-memory latency of one batch is paid once, not once per item
        int k0n[AU_BATCH];                                                // the first bins of the next batch's items, one batch ahead
#pragma unroll
        for (int i = 0; i < AU_BATCH; ++i) k0n[i] = s_k0[i * 32 + lane];
        __syncwarp();                                                     // row complete; park buffer reads done
        float4 x0[AU_BATCH], x1[AU_BATCH], w0[AU_BATCH], w1[AU_BATCH];
        auto fb_loads = [&](int r0) {
            int k0[AU_BATCH];
#pragma unroll
            for (int i = 0; i < AU_BATCH; ++i) k0[i] = k0n[i];
            if (r0 + AU_BATCH < rounds) {
#pragma unroll
                for (int i = 0; i < AU_BATCH; ++i) k0n[i] = s_k0[(r0 + AU_BATCH + i) * 32 + lane];
            }
#pragma unroll
            for (int i = 0; i < AU_BATCH; ++i) {
                x0[i] = *reinterpret_cast<const float4*>(row + k0[i]);
                x1[i] = *reinterpret_cast<const float4*>(row + k0[i] + 4);
                w0[i] = s_w[(r0 + i) * 64 + lane];
                w1[i] = s_w[(r0 + i) * 64 + 32 + lane];
            }
        };
        auto fb_fmas = [&](int r0) {
#pragma unroll
            for (int i = 0; i < AU_BATCH; ++i) {
                float a = x0[i].x * w0[i].x, c = x1[i].x * w1[i].x;
                a = fmaf(x0[i].y, w0[i].y, a); c = fmaf(x1[i].y, w1[i].y, c);
                a = fmaf(x0[i].z, w0[i].z, a); c = fmaf(x1[i].z, w1[i].z, c);
                a = fmaf(x0[i].w, w0[i].w, a); c = fmaf(x1[i].w, w1[i].w, c);
                partial[(r0 + i) * 32 + lane] = a + c;
            }
        };
        fb_loads(0);
        // the normalised magnitudes go out while the first batch of the filterbank is on its way from shared memory
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
        fb_fmas(0);
#pragma unroll 1
        for (int r0 = AU_BATCH; r0 < rounds; r0 += AU_BATCH) {
            fb_loads(r0);
            fb_fmas(r0);
        }
        __syncwarp();
        {
            // The mels, in item order: lane l sums mels l, l + 32 and (l < 16) l + 64.  All item sums of the three are fetched
            // first (predicated 32-bit loads, at most AU_MEL_CAP* each: the host checks the plan against the caps), so the
            // shared-memory latency is paid once per frame instead of once per item; adding 0 for an absent item leaves the
            // sum unchanged, so the result is the sequential sum of the mel's items.
            const int2 i0 = s_mel[lane], i1 = s_mel[lane + 32], i2 = lane < NMELS - 64 ? s_mel[lane + 64] : make_int2(0, 0);
            float q0[AU_MEL_CAP0], q1[AU_MEL_CAP1], q2[AU_MEL_CAP2];
#pragma unroll
            for (int i = 0; i < AU_MEL_CAP0; ++i) q0[i] = i < i0.y ? partial[i0.x + i] : 0.0f;
#pragma unroll
            for (int i = 0; i < AU_MEL_CAP1; ++i) q1[i] = i < i1.y ? partial[i1.x + i] : 0.0f;
#pragma unroll
            for (int i = 0; i < AU_MEL_CAP2; ++i) q2[i] = i < i2.y ? partial[i2.x + i] : 0.0f;
            float s0 = q0[0], s1 = q1[0], s2 = q2[0];
#pragma unroll
            for (int i = 1; i < AU_MEL_CAP0; ++i) s0 += q0[i];
#pragma unroll
            for (int i = 1; i < AU_MEL_CAP1; ++i) s1 += q1[i];
#pragma unroll
            for (int i = 1; i < AU_MEL_CAP2; ++i) s2 += q2[i];
            const int fbase = st[3], t = st[0] - fbase, Tu = st[2] - fbase, bu = st[5];
            auto put = [&](int m, float sum) {
                const float val = db_norm(sum);
                if (P.mel_layout == 0) P.mel_out[(size_t)NMELS * fbase + (size_t)m * Tu + t] = val;
                else P.mel_out[((size_t)bu * P.T_max + t) * NMELS + m] = val;
            };
            put(lane, s0);
            put(lane + 32, s1);
            if (lane < NMELS - 64) put(lane + 64, s2);
        }
        __syncwarp();                                                     // partial sums read before xs is reused
        if (lane == 0) st[0] = st[0] + 1;
        __syncwarp();
    }
    release_tmem();
}

}  // namespace unast
