// K3: Griffin-Lim iteration kernel (replaces reference src/utils.py:309-328 and librosa.griffinlim as
// called at src/gl_vocoder.py:26).
//
// State carried between iterations is the *time-domain* overlap-add buffer (raw window-weighted sums,
// one float per sample, padded coordinates of librosa.istft) -- classic Griffin-Lim needs no complex
// state at all; the fast (momentum) variant additionally carries the previous STFT estimate.
// One iteration = for every frame: [normalise by the window-sum envelope + reflect-pad + window] ->
// rFFT-2048 -> phase update against the target magnitude S -> irFFT-2048 -> window -> overlap-add.
// All transforms run in a circularly shifted frame (the 1102 window taps start at sample 0); the unit
// phase this puts on every bin cancels between the forward and the inverse transform.
//
// Every warp is autonomous (no CTA-wide barrier after the table set-up): it owns a contiguous run of frames and,
// per frame,
//   * loads the 1102 samples straight from the previous iteration's overlap-add buffer (L2-resident), multiplied
//     by a table that holds analysis window / window-sum envelope (edge frames: reflect padding + edge envelope
//     tables),
//   * runs rFFT -> phase update -> irFFT in registers; the transposition buffer in shared memory is private to
//     the warp and, between the two transforms, doubles as the landing zone of the frame's target-magnitude row,
//     which is prefetched with cp.async while the forward FFT is still running,
//   * overlap-adds into a warp-private 827-sample carry in shared memory and emits the 275 samples that became
//     final.  Only the seams between the runs of different warps go through float atomics in L2 (exactly two
//     contributions per sample, so the sum is order-independent and the result deterministic).  Three rotating
//     buffers (read / write / being re-zeroed at the seams) make that safe without an extra launch.
#pragma once
#include <cfloat>
#include "fft_warp.cuh"

namespace unast {

enum GlMode : int { GL_INIT = 0, GL_ITER = 1 };

constexpr int GL_EDGE_L = 2404;            // left-edge envelope table: padded positions [0, 2404)
constexpr int GL_EDGE_R = 2204;            // right-edge table: u + 1100, u = q - 473 - 275 (T - 1) in [-1100, 1104)
constexpr int GL_TABLE_MIN_T = 12;         // utterances shorter than this use the exact (slow) envelope on edge frames

struct GlParams {
    const float* S;              // [sum T][S_STRIDE] target magnitude (already raised to `power`)
    const int* frame_off;        // [B+1]
    const long long* pad_off;    // [B+1] prefix sum of Lp_b = 2048 + 275 (T_b - 1)
    int B;
    int total_frames;
    const int4* runs;            // [gridDim.x * GL_WARPS] (first frame, end frame, utterance of the first frame) per warp
    const long long* cost_off;   // [B+1] prefix sum of the utterances' cost (interior frame 4, edge frame 5): run partition
    const float* ola_r;          // read buffer  (unused in GL_INIT)
    float* ola_w;                // write buffer
    float* ola_z;                // buffer whose seams are zeroed for the next iteration
    float2* tprev;               // [sum T][1025] previous STFT estimate (momentum variant) or nullptr
    const float2* angles0;       // [sum T][1025] initial unit phases (GL_INIT) or nullptr = zero phase
    float alpha;                 // momentum / (1 + momentum); 0 selects classic Griffin-Lim
    int mode;
    int first_iter;              // momentum: treat tprev as zero (librosa starts from rebuilt = 0)
    const float2* wa2;           // interior frames: (w[2n], w[2n+1]) / 2 / window-sum envelope (periodic part)
    const float2* win2;          // (w[2n], w[2n+1]) / 2
    const float2* wini2;         // (w[2n], w[2n+1]) / 2048 (irfft scaling folded into the synthesis window)
    const float2* tw1024;
    const float4* tw4;           // packed layout of tw1024 for fft_pk.cuh: [16][32] (c[2m], c[2m+1], s[2m], s[2m+1])
    const float4* tw4a;          // ... for pairs (j, j + 16): [16][32] (c[j], c[j+16], s[j], s[j+16])
    const float4* wp4;           // packed wp2: [8][32] (wx[ka], wx[ka+32], wy[ka], wy[ka+32]), ka = lane + 64 m
    const float2* wp2;
    const float2* phi;           // [1025] exp(+2 pi i 473 k / 2048): shift phase applied to the initial spectrum
    const float* rinv_l;         // [GL_EDGE_L] 1 / envelope at padded position q (left edge)
    const float* rinv_r;         // [GL_EDGE_R] 1 / envelope at u + 1100 (right edge)
    const float* wss_per;        // [275] interior (periodic) window-sum envelope, float32 as librosa accumulates it
    const double* w2;            // [1102] hann^2 in float64 (exact edge envelope for very short utterances)
};

constexpr int GL_WARPS = 16;
constexpr int GL_XS_BYTES = xs_bytes<float>();                   // 4608: transposition buffer / S row landing zone
constexpr int GL_SIDE_OFF = 4352;                                 // 32 float2 of side buffer behind the 4128-byte S row
constexpr int GL_CARRY = 828;                                     // 827 carried samples + one zero
constexpr int GL_WARP_BYTES = GL_XS_BYTES + GL_CARRY * 4;         // 7920
constexpr int gl_off_warp = 0;
constexpr int gl_off_wa = gl_off_warp + GL_WARPS * GL_WARP_BYTES;
constexpr int gl_off_win = gl_off_wa + (NLIVE_C * 8 + 15) / 16 * 16;
constexpr int gl_off_wini = gl_off_win + (NLIVE_C * 8 + 15) / 16 * 16;
constexpr int gl_off_tw = gl_off_wini + (NLIVE_C * 8 + 15) / 16 * 16;
constexpr int gl_off_wp = gl_off_tw + 1024 * 8;
constexpr int gl_smem_bytes = gl_off_wp + 1024 * 8;
static_assert(S_STRIDE * 4 <= GL_SIDE_OFF && GL_SIDE_OFF + 256 <= GL_XS_BYTES, "S row and side buffer share the transposition buffer");

template <bool MOMENTUM>
__global__ void __launch_bounds__(GL_WARPS * 32, 1) gl_kernel(const GlParams P) {
    constexpr int NT = GL_WARPS * 32;
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float* xs = reinterpret_cast<float*>(smem + gl_off_warp + warp * GL_WARP_BYTES);
    float2* side = reinterpret_cast<float2*>(reinterpret_cast<unsigned char*>(xs) + GL_SIDE_OFF);
    float* cb = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(xs) + GL_XS_BYTES);
    float2* s_wa = reinterpret_cast<float2*>(smem + gl_off_wa);
    float2* s_win = reinterpret_cast<float2*>(smem + gl_off_win);
    float2* s_wini = reinterpret_cast<float2*>(smem + gl_off_wini);
    float2* s_tw = reinterpret_cast<float2*>(smem + gl_off_tw);
    float2* s_wp = reinterpret_cast<float2*>(smem + gl_off_wp);

    for (int i = tid; i < NLIVE_C; i += NT) { s_wa[i] = P.wa2[i]; s_win[i] = P.win2[i]; s_wini[i] = P.wini2[i]; }
    for (int i = tid; i < 1024; i += NT) { s_tw[i] = P.tw1024[i]; s_wp[i] = P.wp2[i]; }
    __syncthreads();                                                    // the only CTA-wide barrier

    // this warp's run of frames: equal shares of the total cost (edge frames are ~1.3x an interior frame)
    const long long W = (long long)gridDim.x * GL_WARPS, wid = (long long)blockIdx.x * GL_WARPS + warp;
    const long long total_cost = P.cost_off[P.B];
    auto frame_at = [&](long long c) -> int {                          // frame that contains cost point c
        if (c <= 0) return 0;
        if (c >= total_cost) return P.total_frames;
        int lo = 0, hi = P.B;
        while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (P.cost_off[mid] <= c) lo = mid; else hi = mid; }
        const int T = P.frame_off[lo + 1] - P.frame_off[lo];
        int ci = (int)(c - P.cost_off[lo]), t;
        const int head = 5 * min(4, T), inner = 4 * max(0, T - 8);
        if (ci < head) t = ci / 5;
        else if (ci < head + inner) t = 4 + (ci - head) / 4;
        else t = 4 + max(0, T - 8) + (ci - head - inner) / 5;
        return P.frame_off[lo] + min(t, T - 1);
    };
    const int f0 = frame_at((total_cost * wid) / W), f1 = frame_at((total_cost * (wid + 1)) / W);
    if (f0 >= f1) return;
    int b;
    {
        int lo = 0, hi = P.B;
        while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (P.frame_off[mid] <= f0) lo = mid; else hi = mid; }
        b = lo;
    }
    const WpTable<float> wp{s_wp};
    bool have_carry = false;        // cb holds the partial sums of this utterance's previous frames
    long long seam_end = -1;        // positions below this index of ola_w also receive the previous run's tail

    for (int f = f0; f < f1; ++f) {
        if (f >= P.frame_off[b + 1]) { while (f >= P.frame_off[b + 1]) ++b; have_carry = false; }
        const int fbase = P.frame_off[b];
        const int Tb = P.frame_off[b + 1] - fbase;
        const int t = f - fbase;
        const int L = HOP * (Tb - 1);
        const long long ubase = P.pad_off[b];
        const long long pos0 = ubase + HOP * t + WIN_LPAD;             // padded index of the frame's tap 0
        const bool last_of_utt = t == Tb - 1;
        const bool last_of_run = f == f1 - 1;
        const float* Srow_g = P.S + (size_t)f * S_STRIDE;

        if (!have_carry) {
            for (int i = lane; i < GL_CARRY; i += 32) cb[i] = 0.0f;
            seam_end = -1;
            if (t > 0) {                                                 // head seam: shared with the previous run's tail
                seam_end = pos0 + OVERLAP;
                for (int i = lane; i < OVERLAP; i += 32) P.ola_z[pos0 + i] = 0.0f;
            }
            __syncwarp();
        }

        float2 v[32];
        float nyq;
        if (P.mode == GL_ITER) {
            // ---- 1. previous iteration's signal: normalise, reflect-pad, window ----
            if (t >= 4 && t <= Tb - 5) {
                const float* __restrict__ x = P.ola_r + pos0 + 2 * lane;
                static_for<0, 18>([&](auto jj) {
                    constexpr int j = decltype(jj)::value;
                    const int n = lane + 32 * j;
                    if (j < 17 || n < NLIVE_C) {
                        const float2 w = s_wa[n];
                        v[j] = make_float2(x[64 * j] * w.x, x[64 * j + 1] * w.y);
                    } else {
                        v[j] = make_float2(0.f, 0.f);
                    }
                });
            } else {
                // Edge frame (rare): a compact runtime loop stages the normalised, reflect-padded taps in the idle
                // transposition buffer (unrolling this path 36x would bloat the instruction cache); the loop body is
                // branch-free so that the unrolled loads are issued back to back.
                const float* __restrict__ src = P.ola_r + ubase;
                const int nbase = HOP * t + WIN_LPAD - HALF;
                if (Tb >= GL_TABLE_MIN_T) {
                    const float* __restrict__ rtab = (t < 4) ? P.rinv_l : P.rinv_r + (1100 - WIN_LPAD - HOP * (Tb - 1));
#pragma unroll 6
                    for (int i = lane; i < WIN; i += 32) {
                        int n = abs(nbase + i);                            // reflect at 0
                        n = min(n, 2 * (L - 1) - n);                       // reflect at L - 1
                        const int q = n + HALF;
                        xs[i] = src[q] * rtab[q];
                    }
                } else {
#pragma unroll 1
                    for (int i = lane; i < WIN; i += 32) {
                        int n = abs(nbase + i);
                        n = min(n, 2 * (L - 1) - n);
                        const int q = n + HALF;
                        float y = src[q];
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
                        v[j] = make_float2(y.x * w.x, y.y * w.y);
                    } else {
                        v[j] = make_float2(0.f, 0.f);
                    }
                });
                __syncwarp();
            }
            static_for<18, 32>([&](auto jj) { constexpr int j = decltype(jj)::value; v[j] = make_float2(0.f, 0.f); });
            // ---- 2. STFT column ----
            warp_fft1024_split<float, 18>(v, xs, s_tw, lane);
        }
        // ---- 3. prefetch the frame's target-magnitude row into the (now idle) transposition buffer ----
        for (int c = lane; c < S_STRIDE / 4; c += 32) cp_async16(xs + 4 * c, Srow_g + 4 * c);
        cp_async_commit();
        if (P.mode == GL_ITER) rfft_post_shfl<float>(v, nyq, side, wp, lane);
        cp_async_wait_all();
        __syncwarp();

        // ---- 4. phase update against S ----
        if (P.mode == GL_ITER) {
            if (!MOMENTUM) {
                // utils.py:315-316: X = S * est / max(1e-8, |est|)
                static_for<0, 32>([&](auto jj) {
                    constexpr int j = decltype(jj)::value;
                    const float s = xs[lane + 32 * j];
                    const float pw = fmaf(v[j].x, v[j].x, v[j].y * v[j].y);
                    const float sc = s * rsqrtf(fmaxf(pw, 1e-16f));            // == s / max(1e-8, |est|)
                    v[j] = make_float2(v[j].x * sc, v[j].y * sc);
                });
                if (lane == 0) nyq = xs[1024] * (nyq / fmaxf(1e-8f, fabsf(nyq)));
            } else {
                // librosa.griffinlim: angles = rebuilt - alpha * tprev; angles /= |angles| + 1e-16
                float2* tp = P.tprev + (size_t)f * NBINS;
                static_for<0, 32>([&](auto jj) {
                    constexpr int j = decltype(jj)::value;
                    const int k = lane + 32 * j;
                    float2 a = v[j];
                    if (!P.first_iter) { const float2 q = tp[k]; a.x = fmaf(-P.alpha, q.x, a.x); a.y = fmaf(-P.alpha, q.y, a.y); }
                    tp[k] = v[j];
                    const float s = xs[k];
                    const float mag = sqrtf(fmaf(a.x, a.x, a.y * a.y)) + 1e-16f;
                    const float sc = s / mag;
                    v[j] = make_float2(a.x * sc, a.y * sc);
                });
                if (lane == 0) {
                    float an = nyq;
                    if (!P.first_iter) an = fmaf(-P.alpha, tp[1024].x, an);
                    tp[1024] = make_float2(nyq, 0.f);
                    nyq = xs[1024] * (an / (fabsf(an) + 1e-16f));
                }
            }
        } else {
            // initial spectrum S * angles0 in the shifted frame
            const float2* a0 = P.angles0 ? P.angles0 + (size_t)f * NBINS : nullptr;
            static_for<0, 32>([&](auto jj) {
                constexpr int j = decltype(jj)::value;
                const int k = lane + 32 * j;
                float2 ph = P.phi[k];
                if (a0) ph = cmul(a0[k], ph);
                const float s = xs[k];
                v[j] = make_float2(s * ph.x, s * ph.y);
            });
            if (lane == 0) {
                v[0].y = 0.0f;                                           // irfft ignores Im of the DC bin
                float2 pn = P.phi[1024];
                if (a0) pn = cmul(a0[1024], pn);
                nyq = xs[1024] * pn.x;                                   // ... and of the Nyquist bin
            }
        }
        // ---- 5. ISTFT column ----
        irfft_pre_shfl<float>(v, nyq, side, wp, lane);                   // (its first __syncwarp also fences the S reads)
        warp_fft1024_split<float, 32>(v, xs, s_tw, lane);

        // ---- 6. window, overlap-add with the carried partial sums, emit what became final ----
        // tap i of this frame sits at ola_w[pos0 + i]; taps [0, 275) are complete once this frame is added
        // (given the carry), taps [275, 1102) become the next frame's carry [0, 827).
        static_for<0, 18>([&](auto jj) {
            constexpr int j = decltype(jj)::value;
            const int n = lane + 32 * j;
            if (j < 17 || n < NLIVE_C) {
                const float2 w = s_wini[n];
                const int i0 = 2 * n;
                float o0 = v[j].x * w.x, o1 = -v[j].y * w.y;
                if (j < 12 || i0 < OVERLAP) {                            // carry covers taps [0, 827); cb[827] == 0
                    const float2 c = *reinterpret_cast<const float2*>(cb + i0);
                    o0 += c.x;
                    o1 += c.y;
                }
                auto put = [&](int i, float val) {
                    if (i < HOP) {
                        const long long q = pos0 + i;
                        if (q < seam_end) atomicAdd(P.ola_w + q, val);
                        else P.ola_w[q] = val;
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
                    if (q < seam_end) atomicAdd(P.ola_w + q, cb[i]); else P.ola_w[q] = cb[i];
                }
            } else {
                for (int i = lane; i < OVERLAP; i += 32) atomicAdd(P.ola_w + q0 + i, cb[i]);
            }
            __syncwarp();
            have_carry = false;
        }
    }
}

// wav[n] = ola[n + 1024] / wss  (the final librosa.istft of utils.py:317 / griffinlim's return)
struct GlFinalParams {
    const float* ola;
    const int* frame_off;
    const long long* pad_off;
    const long long* wav_off;    // [B+1] prefix sum of L_b = 275 (T_b - 1)
    int B;
    float* wav;
    const float* wss_per;
    const double* w2;
};

}  // namespace unast
