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
// A persistent CTA owns a contiguous run of 16-frame steps.  Inside a run the 827-sample overlap
// between consecutive steps is carried in shared memory; only the seams between runs of different
// CTAs go through float atomics in L2 (exactly two contributions per sample, so the sum is
// order-independent and the result deterministic).  Three rotating buffers (read / write / being
// re-zeroed at the seams) make that safe without an extra launch.
#pragma once
#include <cfloat>
#include "fft_warp.cuh"

namespace unast {

enum GlMode : int { GL_INIT = 0, GL_ITER = 1 };

struct GlParams {
    const float* S;              // [sum T][S_STRIDE] target magnitude (already raised to `power`)
    const int* frame_off;        // [B+1]
    const int* step_off;         // [B+1] prefix sum of ceil(T_b / 16)
    const long long* pad_off;    // [B+1] prefix sum of Lp_b = 2048 + 275 (T_b - 1)
    int B;
    const float* ola_r;          // read buffer  (unused in GL_INIT)
    float* ola_w;                // write buffer
    float* ola_z;                // buffer whose seams are zeroed for the next iteration
    float2* tprev;               // [sum T][1025] previous STFT estimate (momentum variant) or nullptr
    const float2* angles0;       // [sum T][1025] initial unit phases (GL_INIT) or nullptr = zero phase
    float alpha;                 // momentum / (1 + momentum); 0 selects classic Griffin-Lim
    int mode;
    int first_iter;              // momentum: treat tprev as zero (librosa starts from rebuilt = 0)
    const float2* win2;          // (w[2n], w[2n+1])
    const float2* wini2;         // same / 1024 (irfft scaling folded into the synthesis window)
    const float2* tw1024;
    const float2* wp;
    const float2* phi;           // [1025] exp(+2 pi i 473 k / 2048): shift phase applied to the initial spectrum
    const float* wss_per;        // [275] interior (periodic) window-sum envelope, float32 as librosa accumulates it
    const double* w2;            // [1102] hann^2 in float64 (edge envelope, computed on the fly)
};

constexpr int GL_WARPS = 16;
constexpr int GL_FPS = 16;
constexpr int GL_SPAN = HOP * (GL_FPS - 1) + WIN;      // 5227
constexpr int GL_EMIT = HOP * GL_FPS;                  // 4400 samples become final per full step

constexpr int GL_XCH_BYTES = 32 * 34 * 8;
constexpr int GL_PLANE = (GL_SPAN + 1) / 2 + 1;
constexpr int gl_off_xch = 0;
constexpr int gl_off_planes = gl_off_xch + GL_WARPS * GL_XCH_BYTES;
constexpr int gl_off_carry = gl_off_planes + 2 * GL_PLANE * 4;
constexpr int gl_off_win = (gl_off_carry + 2 * OVERLAP * 4 + 15) / 16 * 16;
constexpr int gl_off_wini = gl_off_win + (NLIVE_C * 8 + 15) / 16 * 16;
constexpr int gl_off_tw = gl_off_wini + (NLIVE_C * 8 + 15) / 16 * 16;
constexpr int gl_off_wp = gl_off_tw + 1024 * 8;
constexpr int gl_off_wss = gl_off_wp + 1024 * 8;
constexpr int gl_smem_bytes = gl_off_wss + HOP * 4;

__global__ void __launch_bounds__(GL_WARPS * 32, 1) gl_kernel(const GlParams P) {
    constexpr int NT = GL_WARPS * 32;
    extern __shared__ __align__(16) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float2* xch = reinterpret_cast<float2*>(smem + gl_off_xch + warp * GL_XCH_BYTES);
    float* PE = reinterpret_cast<float*>(smem + gl_off_planes);
    float* PO = PE + GL_PLANE;
    float* carry = reinterpret_cast<float*>(smem + gl_off_carry);
    float2* s_win = reinterpret_cast<float2*>(smem + gl_off_win);
    float2* s_wini = reinterpret_cast<float2*>(smem + gl_off_wini);
    float2* s_tw = reinterpret_cast<float2*>(smem + gl_off_tw);
    float2* s_wp = reinterpret_cast<float2*>(smem + gl_off_wp);
    float* s_wss = reinterpret_cast<float*>(smem + gl_off_wss);

    for (int i = tid; i < NLIVE_C; i += NT) { s_win[i] = P.win2[i]; s_wini[i] = P.wini2[i]; }
    for (int i = tid; i < 1024; i += NT) { s_tw[i] = P.tw1024[i]; s_wp[i] = P.wp[i]; }
    for (int i = tid; i < HOP; i += NT) s_wss[i] = P.wss_per[i];

    const int total_steps = P.step_off[P.B];
    const int g0 = (int)(((long long)total_steps * blockIdx.x) / gridDim.x);
    const int g1 = (int)(((long long)total_steps * (blockIdx.x + 1)) / gridDim.x);
    if (g0 >= g1) return;
    int b;
    {
        int lo = 0, hi = P.B;
        while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (P.step_off[mid] <= g0) lo = mid; else hi = mid; }
        b = lo;
    }
    __syncthreads();

    int cbuf = 0;               // carry double buffer selector
    bool have_carry = false;    // previous step of the same utterance was processed by this CTA
    for (int g = g0; g < g1; ++g) {
        if (g >= P.step_off[b + 1]) { while (g >= P.step_off[b + 1]) ++b; have_carry = false; }
        const int ta = (g - P.step_off[b]) * GL_FPS;
        const int fbase = P.frame_off[b];
        const int Tb = P.frame_off[b + 1] - fbase;
        const int nf = min(GL_FPS, Tb - ta);
        const bool last_of_utt = (ta + nf == Tb);
        const bool next_is_mine = !last_of_utt && (g + 1 < g1);
        const long long pbase = P.pad_off[b] + HOP * ta + WIN_LPAD;   // padded index of staged sample 0
        const int span = HOP * (nf - 1) + WIN;
        const int L = HOP * (Tb - 1);
        const bool head_seam = !have_carry && ta > 0;

        // ---- 1. stage the normalised, reflect-padded signal of the previous iteration ----
        if (P.mode == GL_ITER) {
            const float* src = P.ola_r + P.pad_off[b];
            for (int i = tid; i < span; i += NT) {
                int n = HOP * ta + WIN_LPAD + i - HALF;
                if (n < 0) n = -n; else if (n >= L) n = 2 * (L - 1) - n;
                const int q = n + HALF;
                float y = src[q];
                const float w = wss_at(q, Tb, s_wss, P.w2);
                if (w > FLT_MIN) y = y / w;                       // librosa.istft: y[nz] /= wss[nz]
                if (i & 1) PO[i >> 1] = y; else PE[i >> 1] = y;
            }
        }
        if (head_seam) {
            for (int i = tid; i < OVERLAP; i += NT) P.ola_z[pbase + i] = 0.0f;
        }
        __syncthreads();

        // ---- 2. one frame per warp: STFT column -> phase update -> ISTFT column ----
        if (warp < nf) {
            const size_t grow = (size_t)(fbase + ta + warp);
            const float* Srow = P.S + grow * S_STRIDE;
            float2 v[32];
            float nyq;
            if (P.mode == GL_ITER) {
                const int o = HOP * warp;
                const float* pa = (o & 1) ? PO + (o >> 1) : PE + (o >> 1);
                const float* pb = (o & 1) ? PE + ((o + 1) >> 1) : PO + (o >> 1);
                static_for<0, 18>([&](auto jj) {
                    constexpr int j = decltype(jj)::value;
                    const int n = lane + 32 * j;
                    if (j < 17 || n < NLIVE_C) {
                        float2 w = s_win[n];
                        v[j] = make_float2(pa[n] * w.x, pb[n] * w.y);
                    } else {
                        v[j] = make_float2(0.f, 0.f);
                    }
                });
                warp_fft1024<float, 18>(v, xch, s_tw, lane);
                rfft_post<float>(v, nyq, xch, WpTable<float>{s_wp}, lane);
                if (P.alpha == 0.0f) {
                    // utils.py:315-316: X = S * est / max(1e-8, |est|)
                    static_for<0, 32>([&](auto jj) {
                        constexpr int j = decltype(jj)::value;
                        const float s = Srow[lane + 32 * j];
                        const float pw = fmaf(v[j].x, v[j].x, v[j].y * v[j].y);
                        const float sc = s * (pw >= 1e-16f ? rsqrtf(pw) : 1e8f);
                        v[j] = make_float2(v[j].x * sc, v[j].y * sc);
                    });
                    nyq = Srow[1024] * (nyq / fmaxf(1e-8f, fabsf(nyq)));
                } else {
                    // librosa.griffinlim: angles = rebuilt - alpha * tprev; angles /= |angles| + 1e-16
                    float2* tp = P.tprev + grow * NBINS;
                    static_for<0, 32>([&](auto jj) {
                        constexpr int j = decltype(jj)::value;
                        const int k = lane + 32 * j;
                        float2 a = v[j];
                        if (!P.first_iter) { float2 q = tp[k]; a.x = fmaf(-P.alpha, q.x, a.x); a.y = fmaf(-P.alpha, q.y, a.y); }
                        tp[k] = v[j];
                        const float s = Srow[k];
                        const float mag = sqrtf(fmaf(a.x, a.x, a.y * a.y)) + 1e-16f;
                        const float sc = s / mag;
                        v[j] = make_float2(a.x * sc, a.y * sc);
                    });
                    float an = nyq;
                    if (lane == 0) {
                        if (!P.first_iter) an = fmaf(-P.alpha, tp[1024].x, an);
                        tp[1024] = make_float2(nyq, 0.f);
                    }
                    an = __shfl_sync(0xffffffffu, an, 0);
                    nyq = Srow[1024] * (an / (fabsf(an) + 1e-16f));
                }
            } else {
                // initial spectrum S * angles0 in the shifted frame
                const float2* a0 = P.angles0 ? P.angles0 + grow * NBINS : nullptr;
                static_for<0, 32>([&](auto jj) {
                    constexpr int j = decltype(jj)::value;
                    const int k = lane + 32 * j;
                    float2 ph = P.phi[k];
                    if (a0) ph = cmul(a0[k], ph);
                    const float s = Srow[k];
                    v[j] = make_float2(s * ph.x, s * ph.y);
                });
                if (lane == 0) v[0].y = 0.0f;                       // irfft ignores Im of the DC bin
                float2 pn = P.phi[1024];
                if (a0) pn = cmul(a0[1024], pn);
                nyq = Srow[1024] * pn.x;                            // ... and of the Nyquist bin
            }
            irfft_pre<float>(v, nyq, xch, WpTable<float>{s_wp}, lane);
            warp_fft1024<float, 32>(v, xch, s_tw, lane);
            float2* fb = xch;                                        // windowed frame, 1102 floats
            static_for<0, 18>([&](auto jj) {
                constexpr int j = decltype(jj)::value;
                const int n = lane + 32 * j;
                if (j < 17 || n < NLIVE_C) {
                    float2 w = s_wini[n];
                    fb[n] = make_float2(v[j].x * w.x, -v[j].y * w.y);
                }
            });
        }
        __syncthreads();

        // ---- 3. overlap-add the step's frames (+ carried overlap) and emit ----
        {
            const float* cr = carry + cbuf * OVERLAP;
            float* cw = carry + (cbuf ^ 1) * OVERLAP;
            const int n_emit = last_of_utt ? span : GL_EMIT;
            for (int i = tid; i < span; i += NT) {
                float tot = (have_carry && i < OVERLAP) ? cr[i] : 0.0f;
                const int f_hi = min(nf - 1, i / HOP);
                const int f_lo = i < WIN ? 0 : (i - WIN) / HOP + 1;
                for (int f = f_lo; f <= f_hi; ++f)
                    tot += reinterpret_cast<const float*>(smem + gl_off_xch + f * GL_XCH_BYTES)[i - HOP * f];
                if (i < n_emit) {
                    if (head_seam && i < OVERLAP) atomicAdd(P.ola_w + pbase + i, tot);
                    else P.ola_w[pbase + i] = tot;
                } else if (next_is_mine) {
                    cw[i - GL_EMIT] = tot;
                } else {
                    atomicAdd(P.ola_w + pbase + i, tot);              // spill into the next CTA's head seam
                }
            }
            cbuf ^= 1;
            have_carry = next_is_mine;
        }
        __syncthreads();
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

__global__ void gl_finalize_kernel(const GlFinalParams P) {
    const int b = blockIdx.y;
    const int Tb = P.frame_off[b + 1] - P.frame_off[b];
    const int L = HOP * (Tb - 1);
    const float* src = P.ola + P.pad_off[b] + HALF;
    float* dst = P.wav + P.wav_off[b];
    for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < L; n += gridDim.x * blockDim.x) {
        float y = src[n];
        const float w = wss_at(n + HALF, Tb, P.wss_per, P.w2);
        if (w > FLT_MIN) y = y / w;
        dst[n] = y;
    }
}

}  // namespace unast
