// Shared constants and small device helpers for the UNAST spectral kernels (sm_100a).
// Audio constants mirror reference src/audio_parameters.py:1-20.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <type_traits>
#include <cfloat>
#include <climits>

namespace unast {

constexpr int SR = 22050;
constexpr int NFFT = 2048;
constexpr int NBINS = 1025;
constexpr int HOP = 275;
constexpr int WIN = 1102;
constexpr int NMELS = 80;
constexpr int WIN_LPAD = (NFFT - WIN) / 2;      // 473: librosa pad_center(hann(1102), 2048)
constexpr int HALF = NFFT / 2;                  // 1024: centre padding of stft/istft
constexpr int NLIVE_C = WIN / 2;                // 551 packed complex inputs are non-zero
constexpr int OVERLAP = WIN - HOP;              // 827 samples shared with following frames
constexpr int TRIM_FRAME = 2048;
constexpr int TRIM_HOP = 512;

constexpr int S_STRIDE = 1032;                  // Griffin-Lim magnitude rows padded to a 32-byte multiple

// ---- compile-time loop -------------------------------------------------------------------------
template <int I, int N, class F>
__device__ __forceinline__ void static_for(F&& f) {
    if constexpr (I < N) {
        f(std::integral_constant<int, I>{});
        static_for<I + 1, N>(f);
    }
}

// ---- complex helpers (float2 / double2) --------------------------------------------------------
template <class T> struct cplx;
template <> struct cplx<float> { using type = float2; };
template <> struct cplx<double> { using type = double2; };
template <class T> using cplx_t = typename cplx<T>::type;

__device__ __forceinline__ float2 mk(float x, float y) { return make_float2(x, y); }
__device__ __forceinline__ double2 mk(double x, double y) { return make_double2(x, y); }

#ifndef UNAST_PACKED_F32
#define UNAST_PACKED_F32 0
#endif

__device__ __forceinline__ float2 cadd(float2 a, float2 b) {
#if UNAST_PACKED_F32
    return __fadd2_rn(a, b);
#else
    return make_float2(a.x + b.x, a.y + b.y);
#endif
}
__device__ __forceinline__ float2 csub(float2 a, float2 b) {
#if UNAST_PACKED_F32
    return __ffma2_rn(b, make_float2(-1.0f, -1.0f), a);      // a - b in one packed instruction (exact: b * -1 is exact)
#else
    return make_float2(a.x - b.x, a.y - b.y);
#endif
}
__device__ __forceinline__ double2 cadd(double2 a, double2 b) { return make_double2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ double2 csub(double2 a, double2 b) { return make_double2(a.x - b.x, a.y - b.y); }

// (a.x + i a.y) * (w.x + i w.y)
__device__ __forceinline__ float2 cmul(float2 a, float2 w) {
    return make_float2(fmaf(a.x, w.x, -(a.y * w.y)), fmaf(a.x, w.y, a.y * w.x));
}
__device__ __forceinline__ double2 cmul(double2 a, double2 w) {
    return make_double2(fma(a.x, w.x, -(a.y * w.y)), fma(a.x, w.y, a.y * w.x));
}

__device__ __forceinline__ float fma_t(float a, float b, float c) { return fmaf(a, b, c); }
__device__ __forceinline__ double fma_t(double a, double b, double c) { return fma(a, b, c); }

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- tensor memory (TMEM) as a lane-indexed table store ------------------------------------------------------------
// 512 columns x 128 lanes x 32 bit per SM.  With the 32x32b shape thread i of warp w reaches lane 32 (w % 4) + i, consecutive
// registers <-> consecutive columns: a table that every warp indexes as table[register][lane] (FFT twiddles, window taps)
// is stored once per lane quadrant and read with tcgen05.ld -- a data path of its own, next to the shared-memory pipe that
// the transposition and the row buffers keep busy.  Address: bits 31..16 lane, 15..0 column.
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result, int ncols /* power of two, 32..512 */) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, int ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_st4(uint32_t taddr, float4 v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(__float_as_uint(v.x)),
                 "r"(__float_as_uint(v.y)), "r"(__float_as_uint(v.z)), "r"(__float_as_uint(v.w))
                 : "memory");
}
__device__ __forceinline__ void tmem_st2(uint32_t taddr, float2 v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" ::"r"(taddr), "r"(__float_as_uint(v.x)), "r"(__float_as_uint(v.y)) : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// 16 columns -> 16 registers.  The registers may only be read after tmem_wait16 on the SAME array (its "+r" operands tie
// every later use to the wait); loads issued before a wait all complete at that wait.
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                   "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_wait16(uint32_t (&v)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]), "+r"(v[9]),
                   "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]));
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, uint32_t (&v)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(taddr));
}
__device__ __forceinline__ void tmem_wait4(uint32_t (&v)[4]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]));
}

// ---- TMA bulk copies (cp.async.bulk, SASS UBLKCP) completing on an mbarrier --------------------------------------
// One lane arms the barrier with the byte count and issues the copy; the copy engine writes shared memory in full
// lines (a 16-byte cp.async of 32 lanes lands sector by sector and costs ~4x the shared-memory wavefronts).
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra WAIT_DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "WAIT_DONE:\n\t}"
        ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}
// src and dst 16-byte aligned, bytes a multiple of 16.  The destination must not have outstanding generic-proxy
// accesses: callers __syncwarp() first; the proxy fence orders those accesses before the asynchronous write.
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// numpy.pad(..., mode="reflect") index: position m of the infinitely reflected signal of length N (period 2 (N - 1); N == 1: constant).
// A loop of single reflections: one pass for every clip of more than 1024 samples, no integer division, no call.
__device__ __forceinline__ int reflect_index(int m, int N) {
    if (N == 1) return 0;
    while (m < 0 || m >= N) m = m < 0 ? -m : 2 * (N - 1) - m;
    return m;
}

// ---- progress counters between warps (same GPU: .gpu scope; another GPU over NVLink: .sys scope) ------------------
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(int* p, int v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(int* p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_relaxed_gpu(const int* p) {
    int v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ int ld_relaxed_sys(const int* p) {
    int v;
    asm volatile("ld.relaxed.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// ---- cp.async (LDGSTS) helpers: 16-byte global -> shared copies that bypass registers ------------------
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gmem_src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
template <int N>
__device__ __forceinline__ void cp_async_wait_group() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// ---- fast float32 transcendental wrappers (single MUFU instruction each) -------------------------------
__device__ __forceinline__ float sqrt_approx(float x) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float lg2_approx(float x) {
    float r;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// Window-sum-square envelope at padded position q of an utterance with T frames, accumulated exactly
// like librosa.filters.window_sumsquare (float32 accumulator += float64 tap, ascending frame index).
__device__ __forceinline__ float wss_at(int q, int T, const float* __restrict__ wss_per, const double* __restrict__ w2) {
    const int c = q - WIN_LPAD;
    if (c >= OVERLAP && c < HOP * T) return wss_per[c % HOP];
    if (c < 0) return 0.0f;
    int t_hi = min(T - 1, c / HOP);
    int t_lo = c < WIN ? 0 : (c - WIN) / HOP + 1;
    float acc = 0.0f;
    for (int t = t_lo; t <= t_hi; ++t) acc = (float)((double)acc + w2[c - HOP * t]);
    return acc;
}

// Error word layout shared by all kernels (unast_ctx::d_err): first non-zero code wins.
enum DevErr : int {
    ERR_NONE = 0,
    ERR_TOO_SHORT = 1,     // (no longer raised by the front end: short clips are reflect-padded repeatedly, like numpy.pad)
    ERR_EMPTY = 2,         // trimmed signal is empty (reference: IndexError at utils.py:251)
    ERR_SYNC = 3,          // a neighbour wait of the persistent Griffin-Lim kernel timed out (a peer rank never arrived)
};

}  // namespace unast
