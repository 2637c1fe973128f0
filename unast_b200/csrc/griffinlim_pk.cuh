// K3: the persistent Griffin-Lim kernel (replaces reference src/utils.py:309-328 and librosa.griffinlim as called at
// src/gl_vocoder.py:26): ONE cooperative launch runs the initial inverse transform and all n_iter iterations.
//
// Every warp owns a contiguous run of frames for the whole call and, per iteration and frame, does
//   taps of the previous iteration's signal (x window / window-sum envelope) -> rFFT-2048 -> phase update against S
//   -> irFFT-2048 -> window -> overlap-add into a warp-private 827-sample carry, 275 final samples out
// on the packed f32x2 FFT core of fft_pk.cuh.  State between iterations is the time-domain overlap-add buffer (three
// rotating copies: read / write / seams being re-zeroed); the momentum variant also carries the previous STFT estimate.
//
// No grid-wide barrier: a warp starts iteration i as soon as its two NEIGHBOUR warps (the owners of the adjacent runs)
// have finished iteration i - 1 -- progress counters in global memory, release / acquire.  Seams between runs are
// float atomics onto zero (exactly two contributions per sample: order-independent, deterministic).  A seam can also be
// LINKED: its other side lives in another overlap-add buffer -- another GPU's memory, mapped over NVLink, when one long
// waveform is sharded by frame range, or another fragment of the same launch when that run is
// emulated on one GPU.  Both sides then add their partial sums into both copies (red.global.add over NVLink) and push
// their progress counter into the partner's mailbox (st.release.sys); no collective, no host round trip.
//
// The frame's target-magnitude row and the next frame's taps are fetched by TMA bulk copies (cp.async.bulk completing
// on a per-warp mbarrier) instead of 16-byte cp.async: the r3a ncu capture showed the latter costing 218 of 1,117
// shared-memory wavefronts per frame (a 512-byte LDGSTS lands sector by sector).
#pragma once
#include "griffinlim.cuh"
#include "fft_pk.cuh"

namespace unast {

// TMEM columns of the tables (row m of a [rows][32] float4 table = columns 4 m .. 4 m + 3 of the thread's lane)
constexpr int GLT_TWB = 0, GLT_TWA = 64, GLT_WP4 = 128, GLT_COLS = 256;

constexpr int GLP_XS_BYTES = PK_PARK_BYTES > GL_XS_BYTES ? PK_PARK_BYTES : GL_XS_BYTES;   // 4752: transposition / park buffer
constexpr int GLP_SROW_BYTES = S_STRIDE * 4;                                               // 4128
constexpr int GLP_WARP_BYTES = GLP_XS_BYTES + GLP_SROW_BYTES + GL_CARRY * 4;               // 12192
constexpr int GLP_X_BYTES = (WIN + 6 + 3) / 4 * 16;                                        // 4432: taps + alignment slack
constexpr int glp_off_warp = 0;
constexpr int glp_off_bar = glp_off_warp + GL_WARPS * GLP_WARP_BYTES;                      // two mbarriers per warp
constexpr int glp_off_wa = glp_off_bar + GL_WARPS * 16 + 16;                                 // + the TMEM base address word
constexpr int glp_off_win = glp_off_wa + (NLIVE_C * 8 + 15) / 16 * 16;
constexpr int glp_off_wini = glp_off_win + (NLIVE_C * 8 + 15) / 16 * 16;
constexpr int glp_smem_bytes = glp_off_wini + (NLIVE_C * 8 + 15) / 16 * 16;
static_assert(GLP_XS_BYTES % 16 == 0 && GLP_SROW_BYTES % 16 == 0 && GLP_WARP_BYTES % 16 == 0 && GLP_X_BYTES <= GLP_XS_BYTES &&
              glp_smem_bytes <= 227 * 1024, "shared memory layout");

// phase update of one bin (classic: utils.py:315-316; momentum: librosa.griffinlim), S from the row buffer
template <bool MOMENTUM>
__device__ __forceinline__ void gl_update_bin(float& xr, float& xi, int k, const float* __restrict__ srow, float2* tp,
                                              float alpha, bool first_iter) {
    const float s = srow[k];
    if (!MOMENTUM) {
        const float pw = fmaf(xr, xr, xi * xi);
        const float sc = s * rsqrtf(fmaxf(pw, 1e-16f));                  // == s / max(1e-8, |est|)
        xr *= sc;
        xi *= sc;
    } else {
        float ar = xr, ai = xi;
        if (!first_iter) { const float2 pr = tp[k]; ar = fmaf(-alpha, pr.x, ar); ai = fmaf(-alpha, pr.y, ai); }
        tp[k] = make_float2(xr, xi);
        const float sc = s * rsqrtf(fmaxf(fmaf(ar, ar, ai * ai), 1e-30f));  // librosa: angles / (|angles| + 1e-16)
        xr = ar * sc;
        xi = ai * sc;
    }
}

// classic update of two bins held as a register pair (same arithmetic per half as gl_update_bin<false>)
__device__ __forceinline__ void gl_update_pair(float2& xr, float2& xi, float sx, float sy) {
    const float2 pw = __ffma2_rn(xr, xr, __fmul2_rn(xi, xi));
    const float2 sc = __fmul2_rn(make_float2(sx, sy), make_float2(rsqrtf(fmaxf(pw.x, 1e-16f)), rsqrtf(fmaxf(pw.y, 1e-16f))));
    xr = __fmul2_rn(xr, sc);
    xi = __fmul2_rn(xi, sc);
}

// momentum update with the bin's previous estimate already in registers (zero on the first iteration)
__device__ __forceinline__ void gl_update_bin_m(float& xr, float& xi, int k, const float* __restrict__ srow, float2* tp,
                                                float alpha, float2 pr) {
    const float s = srow[k];
    const float ar = fmaf(-alpha, pr.x, xr), ai = fmaf(-alpha, pr.y, xi);
    tp[k] = make_float2(xr, xi);
    const float sc = s * rsqrtf(fmaxf(fmaf(ar, ar, ai * ai), 1e-30f));
    xr = ar * sc;
    xi = ai * sc;
}

// Wait until *p >= target (the neighbour has finished the previous iteration).  Lane 0 polls; a wait that lasts longer
// than GL_WAIT_NS (a peer rank that never arrived) raises ERR_SYNC and makes every warp of the grid leave.
constexpr unsigned long long GL_WAIT_NS = 4000000000ull;
template <bool SYS>
__device__ __forceinline__ bool gl_wait(const int* p, int target, int* abort_flag, int* err, int lane) {
    int ok = 1;
    if (lane == 0) {
        unsigned long long t0 = 0;
        int spins = 0;
        // relaxed polls (an acquire load invalidates the L1 on every poll), one acquire fence once the value is there
        while ((SYS ? ld_relaxed_sys(p) : ld_relaxed_gpu(p)) - target < 0) {
            if ((++spins & 63) == 0) {
                if (*reinterpret_cast<volatile int*>(abort_flag)) { ok = 0; break; }
                const unsigned long long now = global_timer_ns();
                if (t0 == 0) t0 = now;
                else if (now - t0 > GL_WAIT_NS) { atomicCAS(err, ERR_NONE, ERR_SYNC); atomicExch(abort_flag, 1); ok = 0; break; }
            }
            __nanosleep(20);
        }
        if (SYS) __threadfence_system(); else __threadfence();
    }
    ok = __shfl_sync(0xffffffffu, ok, 0);
    return ok != 0;
}

template <bool MOMENTUM, bool CANON>
__global__ void __launch_bounds__(GL_WARPS * 32, 1) gl_persist_kernel(const GlParams P) {
    constexpr int NT = GL_WARPS * 32;
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned char* wbase = smem + glp_off_warp + warp * GLP_WARP_BYTES;
    float* xs = reinterpret_cast<float*>(wbase);                         // transposition buffer / park buffer / edge scratch
    float* srow = reinterpret_cast<float*>(wbase + GLP_XS_BYTES);        // landing zone of the frame's target-magnitude row
    float* cb = reinterpret_cast<float*>(wbase + GLP_XS_BYTES + GLP_SROW_BYTES);
    uint64_t* bar_s = reinterpret_cast<uint64_t*>(smem + glp_off_bar) + 2 * warp;   // S row landed
    uint64_t* bar_x = bar_s + 1;                                                   // next frame's taps landed
    float2* s_wa = reinterpret_cast<float2*>(smem + glp_off_wa);
    float2* s_win = reinterpret_cast<float2*>(smem + glp_off_win);
    float2* s_wini = reinterpret_cast<float2*>(smem + glp_off_wini);

    for (int i = tid; i < NLIVE_C; i += NT) { s_wa[i] = P.wa2[i]; s_win[i] = P.win2[i]; s_wini[i] = P.wini2[i]; }
    // The three [rows][lane] float4 tables go to tensor memory, one copy per lane quadrant (a warp reaches the 32 lanes of
    // quadrant warp % 4 only): 192 of the ~890 shared-memory wavefronts per frame move to the tcgen05.ld data path.
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(smem + glp_off_bar + GL_WARPS * 16);
    if (warp == 0) tmem_alloc(s_tmem, GLT_COLS);
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    const uint32_t tq = *s_tmem + ((uint32_t)(32 * (warp & 3)) << 16);   // this warp's lane quadrant
    if (warp < 4) {
        for (int m = 0; m < 16; ++m) { tmem_st4(tq + GLT_TWB + 4 * m, P.tw4[m * 32 + lane]); tmem_st4(tq + GLT_TWA + 4 * m, P.tw4a[m * 32 + lane]); }
        for (int m = 0; m < 8; ++m) tmem_st4(tq + GLT_WP4 + 4 * m, P.wp4[m * 32 + lane]);
        tmem_wait_st();
    }
    tmem_fence_before_sync();
    const TabTmem t_twb{tq + GLT_TWB}, t_twa{tq + GLT_TWA}, t_wp4{tq + GLT_WP4};
    if (lane == 0) { mbar_init(bar_s, 1); mbar_init(bar_x, 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();                                                    // tables, barriers and tensor memory are set up
    tmem_fence_after_sync();

    // this warp's run of frames (equal shares of a cost that weights edge frames; resolved once on the host)
    const int gw = blockIdx.x * GL_WARPS + warp;
    if (gw >= P.n_runs) {
        asm volatile("bar.sync 2, %0;" ::"n"(NT) : "memory");            // tensor memory is released when every warp is done
        if (warp == 0) tmem_dealloc(*s_tmem, GLT_COLS);
        return;
    }
    const int cta_threads = 32 * min(GL_WARPS, P.n_runs - blockIdx.x * GL_WARPS);   // warps of this CTA that own a run
    bool dead = false;                                                   // a wait timed out: keep the barriers company, do nothing
    const int4 run = P.runs[gw];
    const int f0 = run.x, f1 = run.y;
    const int link_l = (run.w & 0xffff) - 1, link_r = (run.w >> 16) - 1;
    // whose progress this warp waits for, and whom it tells its own: the owners of the adjacent runs, or the partner
    // of a linked seam through the mailboxes
    // Inside a CTA the barrier at the start of every iteration already says that both neighbours have finished the previous one
    // (every warp publishes, fences and then arrives): only the first and the last warp of a CTA poll a counter -- their neighbour
    // lives in another CTA -- and only their counters are ever read.
    const bool cta_first = warp == 0, cta_last = 32 * (warp + 1) == cta_threads;
    const int* poll_l = link_l >= 0 ? P.links[link_l].poll : (gw > 0 && cta_first ? P.done + gw - 1 : nullptr);
    const int* poll_r = link_r >= 0 ? P.links[link_r].poll : (gw + 1 < P.n_runs && cta_last ? P.done + gw + 1 : nullptr);
    uint32_t ph_s = 0, ph_x = 0;                                         // mbarrier phases

    if (link_l >= 0 || link_r >= 0) {
        // linked seams: the partner's buffers must be initialised (its launch has started) before anything is added
        // into them, and it must not add into ours before this launch's memsets are done (stream order: they are)
        __syncwarp();
        if (lane == 0) {
            __threadfence_system();
            if (link_l >= 0) st_release_sys(P.links[link_l].push, P.epoch);
            if (link_r >= 0) st_release_sys(P.links[link_r].push, P.epoch);
        }
        bool ok = true;
        if (link_l >= 0) ok = gl_wait<true>(poll_l, P.epoch, P.abort, P.err, lane);
        if (ok && link_r >= 0) ok = gl_wait<true>(poll_r, P.epoch, P.abort, P.err, lane);
        dead = !ok;
    }

#pragma unroll 1
    for (int it = 0; it <= P.n_iter; ++it) {
        const bool iter_mode = it > 0;                                   // it == 0: inverse transform of S * angles0 only
        const bool first_iter = it == 1;
        const bool init64 = !iter_mode && P.frames0 != nullptr;          // zero-phase start: taps precomputed in float64
        const float* __restrict__ ola_r = P.ola[it % 3];
        float* __restrict__ ola_w = P.ola[(it + 1) % 3];
        float* __restrict__ ola_z = P.ola[(it + 2) % 3];
        float* peer_wl = link_l >= 0 ? P.links[link_l].peer[(it + 1) % 3] : nullptr;
        float* peer_wr = link_r >= 0 ? P.links[link_r].peer[(it + 1) % 3] : nullptr;

        // target-magnitude row of the first frame (every later row is fetched one frame ahead)
        __syncwarp();
        if (!dead && lane == 0) tma_load_1d(srow, P.S + (size_t)f0 * S_STRIDE, GLP_SROW_BYTES, bar_s);
        if (iter_mode && !dead) {
            // both neighbours have finished iteration it - 1: everything this iteration reads is complete, and
            // nobody still reads what it overwrites or re-zeroes
            bool ok = true;
            if (poll_l) ok = link_l >= 0 ? gl_wait<true>(poll_l, P.epoch + it, P.abort, P.err, lane)
                                         : gl_wait<false>(poll_l, P.epoch + it, P.abort, P.err, lane);
            if (ok && poll_r) ok = link_r >= 0 ? gl_wait<true>(poll_r, P.epoch + it, P.abort, P.err, lane)
                                               : gl_wait<false>(poll_r, P.epoch + it, P.abort, P.err, lane);
            dead = !ok;
        }
        // The CTA's warps (16 adjacent runs) start every iteration together: left alone they drift through the 98 KB loop
        // body and thrash the instruction cache (r3f capture: 1.5 warps per issue cycle stalled on instruction fetch against
        // 0.37 when a launch boundary realigned them every iteration).
        asm volatile("bar.sync 1, %0;" ::"r"(cta_threads) : "memory");
        if (dead) continue;
        // momentum: pull a frame's previous estimate (8,200 bytes) into L2 one frame before its bins are read
        auto prefetch_prev = [&](int f) {
            if (MOMENTUM && iter_mode && !first_iter) {
                const char* row = reinterpret_cast<const char*>(P.tprev + (size_t)f * NBINS);
                for (int c = lane; c < (NBINS * 8 + 127) / 128; c += 32) prefetch_l2(row + 128 * c);
            }
        };
        prefetch_prev(f0);

        int b = run.z;
        bool x_staged = false;          // this frame's input samples were prefetched into xs by the previous frame
        bool have_carry = false;        // cb holds the partial sums of this utterance's previous frames
        long long seam_end = -1;        // positions below this index of ola_w also receive the previous run's tail
        bool seam_linked = false;       // ... and that previous run lives behind a link
        // the fragment's frame range, global frame numbering and padded base stay in registers; reloaded only when the
        // run crosses into the next fragment
        int fbase = P.frame_off[b], fend = P.frame_off[b + 1], tofs = P.t0[b], Tb = P.Tg[b];
        long long ubase = P.pad_off[b] - (long long)HOP * tofs;         // local index of the utterance's padded sample 0
        for (int f = f0; f < f1; ++f) {
            if (f >= fend) {
                do { ++b; fend = P.frame_off[b + 1]; } while (f >= fend);
                fbase = P.frame_off[b];
                tofs = P.t0[b];
                Tb = P.Tg[b];
                ubase = P.pad_off[b] - (long long)HOP * tofs;
                have_carry = false;
            }
            const int t = f - fbase + tofs;                              // frame index inside the whole utterance
            const int L = HOP * (Tb - 1);
            const long long pos0 = ubase + HOP * t + WIN_LPAD;             // local padded index of the frame's tap 0
            const bool last_of_utt = t == Tb - 1;
            const bool run_end = f == f1 - 1;
            // the carry is flushed where the run ends -- and, with canonical blocks, at every block boundary inside it
            // (the warp then shares that seam with itself: tail added here, head added by the next frame)
            const bool last_of_run = run_end || (CANON && (t + 1) % P.block == 0);

            if (!have_carry) {
                for (int i = lane; i < GL_CARRY; i += 32) cb[i] = 0.0f;
                seam_end = -1;
                seam_linked = false;
                if (t > 0) {                                                 // head seam: shared with the previous run's tail
                    seam_end = pos0 + OVERLAP;
                    seam_linked = f == f0 && link_l >= 0;
                    for (int i = lane; i < OVERLAP; i += 32) ola_z[pos0 + i] = 0.0f;
                }
                __syncwarp();
            }

            if (f + 1 < f1) prefetch_prev(f + 1);
            Pk v;
            if (iter_mode) {
                // ---- 1. previous iteration's signal: normalise, reflect-pad, window (layout A) ----
                // (ld.global.cg everywhere: the buffers are rewritten by other SMs between iterations of this launch)
                if (x_staged) {
                    // interior frame whose taps the previous frame fetched (from the 16-byte boundary below pos0)
                    mbar_wait(bar_x, ph_x);
                    ph_x ^= 1;
                    const float* x = xs + (int)(pos0 & 3) + 2 * lane;
                    static_for<0, 18>([&](auto jj) {
                        constexpr int j = decltype(jj)::value;
                        const int n = lane + 32 * j;
                        if (j < 17 || n < NLIVE_C) {
                            const float2 w = s_wa[n];
                            reA<j>(v) = x[64 * j] * w.x;
                            imA<j>(v) = x[64 * j + 1] * w.y;
                        } else {
                            reA<j>(v) = 0.f;
                            imA<j>(v) = 0.f;
                        }
                    });
                    __syncwarp();
                } else if (t >= 4 && t <= Tb - 5) {
                    const float* __restrict__ x = ola_r + pos0 + 2 * lane;
                    static_for<0, 18>([&](auto jj) {
                        constexpr int j = decltype(jj)::value;
                        const int n = lane + 32 * j;
                        if (j < 17 || n < NLIVE_C) {
                            const float2 w = s_wa[n];
                            reA<j>(v) = __ldcg(x + 64 * j) * w.x;
                            imA<j>(v) = __ldcg(x + 64 * j + 1) * w.y;
                        } else {
                            reA<j>(v) = 0.f;
                            imA<j>(v) = 0.f;
                        }
                    });
                } else {
                    // Edge frame (rare): stage the normalised, reflect-padded taps in the idle transposition buffer
                    const float* __restrict__ src = ola_r + ubase;
                    const int nbase = HOP * t + WIN_LPAD - HALF;
                    if (Tb >= GL_TABLE_MIN_T) {
                        const float* __restrict__ rtab = (t < 4) ? P.rinv_l : P.rinv_r + (1100 - WIN_LPAD - HOP * (Tb - 1));
#pragma unroll 6
                        for (int i = lane; i < WIN; i += 32) {
                            int n = abs(nbase + i);                            // reflect at 0
                            n = min(n, 2 * (L - 1) - n);                       // reflect at L - 1
                            const int q = n + HALF;
                            xs[i] = __ldcg(src + q) * rtab[q];
                        }
                    } else {
#pragma unroll 1
                        for (int i = lane; i < WIN; i += 32) {
                            int n = abs(nbase + i);
                            n = min(n, 2 * (L - 1) - n);
                            const int q = n + HALF;
                            float y = __ldcg(src + q);
                            const float w = wss_at(q, Tb, P.wss_per, P.w2);
                            if (w > FLT_MIN) y = y / w;                         // librosa.istft: y[nz] /= wss[nz]
                            xs[i] = y;
                        }
                    }
                    __syncwarp();
                    static_for<0, 18>([&](auto jj) {
                        constexpr int j = decltype(jj)::value;
                        const int n = lane + 32 * j;
                        if (j < 17 || n < NLIVE_C) {
                            const float2 w = s_win[n];
                            const float2 y = reinterpret_cast<const float2*>(xs)[n];
                            reA<j>(v) = y.x * w.x;
                            imA<j>(v) = y.y * w.y;
                        } else {
                            reA<j>(v) = 0.f;
                            imA<j>(v) = 0.f;
                        }
                    });
                    __syncwarp();
                }
                static_for<18, 32>([&](auto jj) { constexpr int j = decltype(jj)::value; reA<j>(v) = 0.f; imA<j>(v) = 0.f; });
                // ---- 2. STFT column (layout B out), mirror pairs: pair m = X[ka], pair m + 8 = X[1024 - ka] ----
                warp_fft1024_pk<18>(v, xs, t_twb, lane);
                pk_post_pairs(v, xs, t_wp4, lane);
            }
            // ---- 3. phase update against S on the mirror pairs: ka = lane + 64 m (+ 32), kb = 1024 - ka ----
            mbar_wait(bar_s, ph_s);                                          // the frame's target-magnitude row has landed
            ph_s ^= 1;
            float2* tp = MOMENTUM ? P.tprev + (size_t)f * NBINS : nullptr;
            float x512r = 0.f, x512i = 0.f;                                  // lane 0: bin 512, its own mirror
            if (iter_mode && MOMENTUM) {
                // The previous estimates (8.2 KB per frame) stream through a four-slot ring in the idle transposition
                // buffer with cp.async, four groups of 4 x 32 bins ahead of the group being updated: register loads that
                // follow stores to the same array in program order cannot be hoisted, so they would put one L2 / HBM
                // round trip per group (per bin, unpipelined) on the critical path.  Every lane reads back only what it
                // copied itself, so cp.async.wait_group is all the synchronisation the ring needs.
                const bool prev = !first_iter;
                if (lane == 0) {                                             // bin 512 sits in the park rows the ring reuses
                    x512r = 2.0f * xs[32 * PK_PARK_S + 16];
                    x512i = -2.0f * xs[PK_PARK_PLANE + 32 * PK_PARK_S + 16];
                }
                float2 p512 = make_float2(0.f, 0.f);
                if (prev && lane == 0) p512 = tp[512];
                __syncwarp();
                float2* ring = reinterpret_cast<float2*>(xs) + lane;        // [4 slots][4 rows][32 lanes]
                auto issue = [&](int m) {
                    if (prev) {
                        float2* d = ring + (m & 3) * 128;
                        const int ka = lane + 64 * m;
                        cp_async8(d, tp + ka);
                        cp_async8(d + 32, tp + ka + 32);
                        cp_async8(d + 64, tp + 1024 - ka);
                        cp_async8(d + 96, tp + 992 - ka);
                    }
                    cp_async_commit();
                };
                issue(0); issue(1); issue(2); issue(3);
                static_for<0, 8>([&](auto mm) {
                    constexpr int m = decltype(mm)::value;
                    const int ka = lane + 64 * m;
                    cp_async_wait_group<(m < 5 ? 3 : 7 - m)>();
                    float2 pr[4] = {};
                    if (prev) {
                        const float2* d = ring + (m & 3) * 128;
                        pr[0] = d[0]; pr[1] = d[32]; pr[2] = d[64]; pr[3] = d[96];
                    }
                    gl_update_bin_m(v.re[m].x, v.im[m].x, ka, srow, tp, P.alpha, pr[0]);
                    gl_update_bin_m(v.re[m].y, v.im[m].y, ka + 32, srow, tp, P.alpha, pr[1]);
                    gl_update_bin_m(v.re[m + 8].x, v.im[m + 8].x, 1024 - ka, srow, tp, P.alpha, pr[2]);
                    gl_update_bin_m(v.re[m + 8].y, v.im[m + 8].y, 992 - ka, srow, tp, P.alpha, pr[3]);
                    if (m < 4) issue(m + 4);                                 // the slot's values are in registers by now
                });
                if (lane == 0) gl_update_bin_m(x512r, x512i, 512, srow, tp, P.alpha, p512);
                __syncwarp();                                                // ring reads done before the park rows are rewritten
            } else if (iter_mode) {
                static_for<0, 8>([&](auto mm) {
                    constexpr int m = decltype(mm)::value;
                    const int ka = lane + 64 * m;
                    gl_update_pair(v.re[m], v.im[m], srow[ka], srow[ka + 32]);
                    gl_update_pair(v.re[m + 8], v.im[m + 8], srow[1024 - ka], srow[992 - ka]);
                });
                if (lane == 0) {
                    x512r = 2.0f * xs[32 * PK_PARK_S + 16];                  // X[512] = 2 conj(z), z parked by pk_post_pairs
                    x512i = -2.0f * xs[PK_PARK_PLANE + 32 * PK_PARK_S + 16];
                    gl_update_bin<false>(x512r, x512i, 512, srow, tp, P.alpha, first_iter);
                }
            } else if (init64) {
                // zero-phase start: w * irfft(S) computed in float64 by gl_init_f64_kernel (tap i == tap 1102 - i), in the form
                // step 5 expects (real parts, negated imaginary parts, unit window)
                const float* __restrict__ f0 = P.frames0 + (size_t)f * GL_F0_STRIDE;
                static_for<0, 18>([&](auto jj) {
                    constexpr int j = decltype(jj)::value;
                    const int i0 = 2 * (lane + 32 * j);
                    if (j < 17 || i0 < WIN) {
                        reB<j>(v) = f0[i0 <= WIN / 2 ? i0 : WIN - i0];
                        imB<j>(v) = -f0[i0 + 1 <= WIN / 2 ? i0 + 1 : WIN - 1 - i0];
                    }
                });
            } else {
                // initial spectrum S * angles0 in the shifted frame
                const float2* a0 = P.angles0 ? P.angles0 + (size_t)f * NBINS : nullptr;
                auto init_bin = [&](float& xr, float& xi, int k) {
                    float2 ph = P.phi[k];
                    if (a0) ph = cmul(a0[k], ph);
                    const float s = srow[k];
                    xr = s * ph.x;
                    xi = s * ph.y;
                };
                static_for<0, 8>([&](auto mm) {
                    constexpr int m = decltype(mm)::value;
                    const int ka = lane + 64 * m;
                    init_bin(v.re[m].x, v.im[m].x, ka);
                    init_bin(v.re[m].y, v.im[m].y, ka + 32);
                    init_bin(v.re[m + 8].x, v.im[m + 8].x, 1024 - ka);
                    init_bin(v.re[m + 8].y, v.im[m + 8].y, 992 - ka);
                });
                if (lane == 0) {
                    v.im[0].x = 0.0f;                                        // irfft ignores Im of the DC bin
                    v.im[8].x = 0.0f;                                        // ... and of the Nyquist bin
                    init_bin(x512r, x512i, 512);
                }
            }
            // ---- 4. back to Z' (layout B, in place; mirrors through the park buffer), ISTFT column ----
            if (!init64) {
                pk_pre_pairs(v, xs, t_wp4, lane);
                if (lane == 0) { xs[0] = 2.0f * x512r; xs[PK_PARK_PLANE] = 2.0f * x512i; }   // element 16 of lane 0
            }
            __syncwarp();
            // next frame's S row (this one has been consumed); after the run's last frame, the first row for the next iteration
            // is fetched at the top of the iteration loop
            if (lane == 0 && f + 1 < f1) tma_load_1d(srow, P.S + (size_t)(f + 1) * S_STRIDE, GLP_SROW_BYTES, bar_s);
            // the next frame's taps go into the transposition buffer as soon as the inverse transform is done with it
            x_staged = iter_mode && f + 1 < f1 && f + 1 < fend && t + 1 >= 4 && t + 1 <= Tb - 5;
            if (!init64) {
                pk_pre_collect(v, xs, lane);
                __syncwarp();
                warp_fft1024_pk_b(v, xs, t_twa, lane, [&]() {
                    if (x_staged && lane == 0) tma_load_1d(xs, ola_r + ((pos0 + HOP) & ~3LL), GLP_X_BYTES, bar_x);
                });
            }

            // ---- 5. window, overlap-add with the carried partial sums, emit what became final ----
            // tap i of this frame sits at ola_w[pos0 + i]; taps [0, 275) are complete once this frame is added
            // (given the carry), taps [275, 1102) become the next frame's carry [0, 827).
            float* wdst = ola_w + pos0;
            float* pdst = seam_linked ? peer_wl + pos0 : nullptr;            // the linked partner's copy of the head seam
            const int nseam = (int)max(0LL, min((long long)HOP, seam_end - pos0));   // leading taps shared with the previous run
            static_for<0, 18>([&](auto jj) {
                constexpr int j = decltype(jj)::value;
                const int n = lane + 32 * j;
                if (j < 17 || n < NLIVE_C) {
                    const float2 w = init64 ? make_float2(1.0f, 1.0f) : s_wini[n];   // float64 start: v already holds the windowed taps
                    const int i0 = 2 * n;
                    float o0 = reB<j>(v) * w.x, o1 = -imB<j>(v) * w.y;
                    if (j < 12 || i0 < OVERLAP) {                            // carry covers taps [0, 827); cb[827] == 0
                        const float2 c = *reinterpret_cast<const float2*>(cb + i0);
                        o0 += c.x;
                        o1 += c.y;
                    }
                    auto put = [&](int i, float val) {
                        if (i < HOP) {
                            if (i < nseam) { atomicAdd(wdst + i, val); if (pdst) atomicAdd(pdst + i, val); }
                            else wdst[i] = val;
                        } else {
                            cb[i - HOP] = val;
                        }
                    };
                    if (j < 4) { put(i0, o0); put(i0 + 1, o1); }             // i0 + 1 <= 255 < 275: always emitted
                    else if (j == 4) { put(i0, o0); put(i0 + 1, o1); }       // straddles tap 275
                    else { cb[i0 - HOP] = o0; cb[i0 + 1 - HOP] = o1; }
                }
            });
            __syncwarp();
            have_carry = true;
            if (last_of_utt || last_of_run) {
                // flush the carry: final at the end of an utterance, otherwise the tail seam shared with the next run
                const long long q0 = pos0 + HOP;
                if (last_of_utt) {
                    for (int i = lane; i < OVERLAP; i += 32) {
                        const long long q = q0 + i;
                        if (q < seam_end) { atomicAdd(ola_w + q, cb[i]); if (pdst) atomicAdd(peer_wl + q, cb[i]); }
                        else ola_w[q] = cb[i];
                    }
                } else if (link_r >= 0 && run_end) {
                    // linked tail seam: both copies get this side's partial sums; this side re-zeroes its own copy
                    for (int i = lane; i < OVERLAP; i += 32) {
                        const float c = cb[i];
                        atomicAdd(ola_w + q0 + i, c);
                        atomicAdd(peer_wr + q0 + i, c);
                        if (q0 + i < seam_end) { if (pdst) atomicAdd(peer_wl + q0 + i, c); }   // a run shorter than the overlap
                        ola_z[q0 + i] = 0.0f;
                    }
                } else {
                    for (int i = lane; i < OVERLAP; i += 32) {
                        atomicAdd(ola_w + q0 + i, cb[i]);
                        if (pdst && q0 + i < seam_end) atomicAdd(peer_wl + q0 + i, cb[i]);
                    }
                }
                __syncwarp();
                have_carry = false;
            }
        }
        // ---- this warp has finished iteration `it`: publish (release) for the neighbours ----
        __syncwarp();
        if (lane == 0) {
            if (link_l >= 0 || link_r >= 0) {
                __threadfence_system();
                if (link_l >= 0) st_release_sys(P.links[link_l].push, P.epoch + it + 1);
                if (link_r >= 0) st_release_sys(P.links[link_r].push, P.epoch + it + 1);
            }
            __threadfence();                                             // (also what makes this warp's sums visible behind the CTA barrier)
            if (cta_first || cta_last) st_release_gpu(P.done + gw, P.epoch + it + 1);
        }
    }
    // Linked seams: the partner's partial sums of the LAST iteration land in this side's buffer from another launch (another
    // GPU): what follows this kernel on the stream (extraction, de-emphasis) may only run once they are there.  (Seams between
    // runs of this launch are complete when the kernel is.)
    if (!dead && (link_l >= 0 || link_r >= 0)) {
        bool ok = true;
        if (link_l >= 0) ok = gl_wait<true>(poll_l, P.epoch + P.n_iter + 1, P.abort, P.err, lane);
        if (ok && link_r >= 0) ok = gl_wait<true>(poll_r, P.epoch + P.n_iter + 1, P.abort, P.err, lane);
    }
    tmem_fence_before_sync();
    asm volatile("bar.sync 2, %0;" ::"n"(NT) : "memory");
    if (warp == 0) tmem_dealloc(*s_tmem, GLT_COLS);
}

}  // namespace unast
