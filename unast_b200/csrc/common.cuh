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

// ---- mbarrier (shared-memory transaction barrier) helpers: producer/consumer hand-off between warps --------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!done);
}

// the same wait as one opaque statement (the retry loop lives inside the PTX block): no loop in the compiler's control
// flow graph, so a wait in the middle of a register-heavy section does not perturb register allocation
__device__ __forceinline__ void mbar_wait_inline(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "LAB_WAIT:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra DONE;\n\t"
        "bra LAB_WAIT;\n\t"
        "DONE:\n\t}"
        ::"r"(smem_u32(bar)), "r"(parity)
        : "memory");
}

// ---- tensor memory (TMEM) as a per-SM operand store ----------------------------------------------------
// 512 columns x 128 lanes x 32 bit per SM.  A warp reaches the 32 lanes of its own quadrant (warp id % 4) with the
// 32x32b shape: thread i <-> lane 32 (warp % 4) + i, consecutive registers <-> consecutive columns.
// Address: bits 31..16 lane, 15..0 column.
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
__device__ __forceinline__ float2 tmem_ld2(uint32_t taddr) {
    uint32_t a, b;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(a), "=r"(b) : "r"(taddr) : "memory");
    return make_float2(__uint_as_float(a), __uint_as_float(b));
}
__device__ __forceinline__ void tmem_wait_ld(float2& v, float2& w) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+f"(v.x), "+f"(v.y), "+f"(w.x), "+f"(w.y)::"memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float4 tmem_ld4(uint32_t taddr) {
    uint32_t a, b, c, d;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(taddr) : "memory");
    return make_float4(__uint_as_float(a), __uint_as_float(b), __uint_as_float(c), __uint_as_float(d));
}
// The loaded registers may only be read after the wait: the "+f" operands tie them to it so the compiler cannot
// hoist a use above the wait.
__device__ __forceinline__ void tmem_wait_ld(float4& v) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+f"(v.x), "+f"(v.y), "+f"(v.z), "+f"(v.w)::"memory");
}
__device__ __forceinline__ void tmem_wait_ld(float4& v, float4& w) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+f"(v.x), "+f"(v.y), "+f"(v.z), "+f"(v.w), "+f"(w.x), "+f"(w.y), "+f"(w.z), "+f"(w.w)::"memory");
}

__device__ __forceinline__ void tmem_wait_ld(float4& a, float4& b, float4& c, float4& d) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+f"(a.x), "+f"(a.y), "+f"(a.z), "+f"(a.w), "+f"(b.x), "+f"(b.y), "+f"(b.z), "+f"(b.w),
                   "+f"(c.x), "+f"(c.y), "+f"(c.z), "+f"(c.w), "+f"(d.x), "+f"(d.y), "+f"(d.z), "+f"(d.w)::"memory");
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
    ERR_TOO_SHORT = 1,     // an utterance is too short for reflect padding (librosa would raise too)
    ERR_EMPTY = 2,         // trimmed signal is empty (reference: IndexError at utils.py:251)
};

}  // namespace unast
