// K3 building blocks: parameter block and constants of the Griffin-Lim iteration kernel gl_persist_kernel (griffinlim_pk.cuh; replaces reference src/utils.py:309-328 and librosa.griffinlim as
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

constexpr int GL_EDGE_L = 2404;            // left-edge envelope table: padded positions [0, 2404)
constexpr int GL_EDGE_R = 2204;            // right-edge table: u + 1100, u = q - 473 - 275 (T - 1) in [-1100, 1104)
constexpr int GL_TABLE_MIN_T = 12;         // utterances shorter than this use the exact (slow) envelope on edge frames
constexpr int GL_MAX_LINKS = 16;

// A linked seam: the 827 samples shared by the last frames of one fragment and the first frames of the next one when
// the two live in different overlap-add buffers (another GPU's memory over NVLink, or -- in the single-GPU emulation of
// the multi-rank run -- another fragment of the same launch).  Both sides add their partial sums into both copies.
struct GlLink {
    float* peer[3];              // partner's copy of rotating buffer k, offset so that peer[k][q] pairs with local index q
    int* push;                   // partner's mailbox (peer memory): this side's progress is stored there
    const int* poll;             // this side's mailbox (local memory): the partner's progress
};

struct GlParams {
    const float* S;              // [sum T][S_STRIDE] target magnitude (already raised to `power`)
    const int* frame_off;        // [B+1] local prefix sum of the fragments' frame counts
    const long long* pad_off;    // [B+1] prefix sum of Lp_b = 2048 + 275 (n_b - 1)
    const int* t0;               // [B] index of the fragment's first frame inside its utterance (0 for whole utterances)
    const int* Tg;               // [B] frames of the whole utterance the fragment belongs to
    int B;
    int n_runs;                  // active warps (each owns a non-empty run of frames)
    const int4* runs;            // [n_runs] (first frame, end frame, fragment of the first frame, link_l + 1 | (link_r + 1) << 16)
    float* ola[3];               // rotating overlap-add buffers: iteration `it` reads [it % 3], writes [(it + 1) % 3], re-zeroes
                                 // the seams of [(it + 2) % 3]
    int n_iter;
    int block;                   // > 0: canonical summation blocks -- the carry restarts at every multiple of `block` frames of an
                                 // utterance, so the overlap-add result does not depend on how the frames are cut into runs / ranks
    int* done;                   // [n_runs] epoch + number of iterations warp w has completed (incl. the initial inverse)
    int epoch;                   // base value of this call's progress counters (counters are never reset)
    int* abort;                  // set when a wait timed out; every wait loop leaves when it is set
    int* err;
    GlLink links[GL_MAX_LINKS];
    float2* tprev;               // [sum T][1025] previous STFT estimate (momentum variant) or nullptr
    const float2* angles0;       // [sum T][1025] initial unit phases or nullptr = zero phase
    const float* frames0;        // zero-phase start only: [sum T][GL_F0_STRIDE] windowed taps 0..551 of w * irfft(S) computed in
                                 // float64 by gl_init_f64_kernel (tap i == tap 1102 - i), or nullptr
    float alpha;                 // momentum / (1 + momentum); 0 selects classic Griffin-Lim
    const float2* wa2;           // interior frames: (w[2n], w[2n+1]) / 2 / window-sum envelope (periodic part)
    const float2* win2;          // (w[2n], w[2n+1]) / 2
    const float2* wini2;         // (w[2n], w[2n+1]) / 2048 (irfft scaling folded into the synthesis window)
    const float4* tw4;           // packed W1024^(lane k1) for fft_pk.cuh: [16][32] (c[2m], c[2m+1], s[2m], s[2m+1])
    const float4* tw4a;          // ... for pairs (j, j + 16): [16][32] (c[j], c[j+16], s[j], s[j+16])
    const float4* wp4;           // packed wp2: [8][32] (wx[ka], wx[ka+32], wy[ka], wy[ka+32]), ka = lane + 64 m
    const float2* phi;           // [1025] exp(+2 pi i 473 k / 2048): shift phase applied to the initial spectrum
    const float* rinv_l;         // [GL_EDGE_L] 1 / envelope at padded position q (left edge)
    const float* rinv_r;         // [GL_EDGE_R] 1 / envelope at u + 1100 (right edge)
    const float* wss_per;        // [275] interior (periodic) window-sum envelope, float32 as librosa accumulates it
    const double* w2;            // [1102] hann^2 in float64 (exact edge envelope for very short utterances)
};

#ifndef GL_WARPS_N
#define GL_WARPS_N 16
#endif
constexpr int GL_WARPS = GL_WARPS_N;
constexpr int GL_F0_STRIDE = 552;  // floats per frame of GlParams::frames0 (taps 0..551; 2,208 bytes)
constexpr int GL_XS_BYTES = xs_bytes<float>();                   // 4608: transposition buffer
constexpr int GL_CARRY = 828;                                     // 827 carried samples + one zero

}  // namespace unast
