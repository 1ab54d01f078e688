// TEST INFRASTRUCTURE ONLY -- a SIMT emulator that lets the CPU test suite (-m "not gpu") execute the SAME kernel
// sources that nvcc compiles for sm_100a.  It is never part of the product: nothing under libadsb_b200/ includes,
// links or loads it; tests/emu/build_emu.py compiles libadsb_b200/csrc/* a second time with g++ against this header
// (which shadows <cuda_runtime.h>) into tests/emu/_build/libadsbcu_emu.so, and only tests/test_emu_*.py load that file.
//
// Model: one CTA at a time; every CUDA thread is a ucontext fiber on ONE OS thread (deterministic, race-free);
// fibers switch only at __syncthreads / named barriers / warp-collective intrinsics.  "Device" memory is host
// memory, streams are synchronous, events measure nothing.  What this cannot show: memory-ordering bugs, real
// concurrency between CTAs, performance.  Those are what the -m gpu tests and ncu are for.
#pragma once
#define ADSBCU_EMU 1

#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __shared__ static
#define __align__(n) __attribute__((aligned(n)))
#define __constant__ static
#define __launch_bounds__(...)
#define __restrict__

// ---- vector types ---------------------------------------------------------------------------------
struct uint3 { unsigned x, y, z; };
struct dim3 {
    unsigned x, y, z;
    dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct alignas(8) uint2 { unsigned x, y; };
struct alignas(16) uint4 { unsigned x, y, z, w; };
struct alignas(8) int2 { int x, y; };
struct alignas(16) int4 { int x, y, z, w; };
struct alignas(8) float2 { float x, y; };
struct alignas(4) ushort2 { unsigned short x, y; };
struct alignas(16) ulonglong2 { unsigned long long x, y; };
static inline uint2 make_uint2(unsigned x, unsigned y) { return uint2{x, y}; }
static inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return uint4{x, y, z, w}; }
static inline int2 make_int2(int x, int y) { return int2{x, y}; }
static inline int4 make_int4(int x, int y, int z, int w) { return int4{x, y, z, w}; }
static inline float2 make_float2(float x, float y) { return float2{x, y}; }
static inline ulonglong2 make_ulonglong2(unsigned long long x, unsigned long long y) { return ulonglong2{x, y}; }

extern uint3 threadIdx, blockIdx;
extern dim3 blockDim, gridDim;
constexpr int warpSize = 32;

// ---- scheduler hooks (tests/emu/emu_runtime.cpp) ----------------------------------------------------
namespace emu {
void launch(dim3 grid, dim3 block, size_t dyn_smem, const std::function<void()> &body);
void barrier(int id, int count);      // CTA barrier `id`; count = threads expected (0: every live thread of the CTA)
void warp_sync();                     // rendezvous of the live lanes of the calling fiber's warp
unsigned long long *warp_slots();     // 32 exchange slots of the calling fiber's warp
unsigned warp_live_mask();            // lanes of this warp that have not exited
int lane();
unsigned char *dyn_smem();            // dynamic shared memory of the running CTA (16-byte aligned, 232 KiB)
int sm_count();                       // ADSBCU_EMU_SMS (default 2): what cudaGetDeviceProperties reports
}  // namespace emu

static inline void __syncthreads() { emu::barrier(0, 0); }
static inline void __syncwarp(unsigned = 0xFFFFFFFFu) { emu::warp_sync(); }
static inline void __threadfence() {}
static inline void __threadfence_block() {}
static inline void __threadfence_system() {}
static inline void __nanosleep(unsigned) {}
static inline long long clock64() { return 0; }
static inline unsigned __activemask() { return emu::warp_live_mask(); }

namespace emu {
template <class T, class F>
inline T exchange(T v, F pick)          // pick(lane) -> source lane (or -1: keep own)
{
    static_assert(sizeof(T) <= 8, "shuffle of at most 64 bits");
    unsigned long long *s = warp_slots();
    const int l = lane();
    unsigned long long raw = 0;
    std::memcpy(&raw, &v, sizeof(T));
    s[l] = raw;
    warp_sync();
    const int src = pick(l);
    T r = v;
    if (src >= 0 && src < 32 && ((warp_live_mask() >> src) & 1u)) std::memcpy(&r, &s[src], sizeof(T));
    warp_sync();
    return r;
}
template <class T, class Op>
inline T reduce(T v, Op op)
{
    unsigned long long *s = warp_slots();
    const int l = lane();
    unsigned long long raw = 0;
    std::memcpy(&raw, &v, sizeof(T));
    s[l] = raw;
    warp_sync();
    const unsigned live = warp_live_mask();
    bool first = true;
    T acc{};
    for (int k = 0; k < 32; ++k) {
        if (!((live >> k) & 1u)) continue;
        T x;
        std::memcpy(&x, &s[k], sizeof(T));
        acc = first ? x : op(acc, x);
        first = false;
    }
    warp_sync();
    return acc;
}
}  // namespace emu

template <class T> static inline T __shfl_sync(unsigned, T v, int src, int = 32) { return emu::exchange(v, [=](int) { return src & 31; }); }
template <class T> static inline T __shfl_up_sync(unsigned, T v, unsigned d, int = 32) { return emu::exchange(v, [=](int l) { return l - int(d); }); }
template <class T> static inline T __shfl_down_sync(unsigned, T v, unsigned d, int = 32) { return emu::exchange(v, [=](int l) { return l + int(d) < 32 ? l + int(d) : -1; }); }
template <class T> static inline T __shfl_xor_sync(unsigned, T v, int m, int = 32) { return emu::exchange(v, [=](int l) { return l ^ m; }); }
static inline unsigned __ballot_sync(unsigned, int pred)
{
    unsigned long long *s = emu::warp_slots();
    s[emu::lane()] = pred ? 1ull : 0ull;
    emu::warp_sync();
    const unsigned live = emu::warp_live_mask();
    unsigned r = 0;
    for (int k = 0; k < 32; ++k) if (((live >> k) & 1u) && s[k]) r |= 1u << k;
    emu::warp_sync();
    return r;
}
static inline int __any_sync(unsigned m, int pred) { return __ballot_sync(m, pred) != 0; }
static inline int __all_sync(unsigned m, int pred) { return __ballot_sync(m, pred) == emu::warp_live_mask(); }
static inline unsigned __reduce_add_sync(unsigned, unsigned v) { return emu::reduce(v, [](unsigned a, unsigned b) { return a + b; }); }
static inline int __reduce_add_sync(unsigned, int v) { return emu::reduce(v, [](int a, int b) { return a + b; }); }
static inline unsigned __reduce_xor_sync(unsigned, unsigned v) { return emu::reduce(v, [](unsigned a, unsigned b) { return a ^ b; }); }
static inline unsigned __reduce_or_sync(unsigned, unsigned v) { return emu::reduce(v, [](unsigned a, unsigned b) { return a | b; }); }
static inline unsigned __reduce_and_sync(unsigned, unsigned v) { return emu::reduce(v, [](unsigned a, unsigned b) { return a & b; }); }
static inline unsigned __reduce_min_sync(unsigned, unsigned v) { return emu::reduce(v, [](unsigned a, unsigned b) { return a < b ? a : b; }); }
static inline unsigned __reduce_max_sync(unsigned, unsigned v) { return emu::reduce(v, [](unsigned a, unsigned b) { return a > b ? a : b; }); }

// ---- atomics (one OS thread: plain read-modify-write) ---------------------------------------------
template <class T, class U> static inline T atomicAdd(T *p, U v) { const T o = *p; *p = T(o + T(v)); return o; }
template <class T, class U> static inline T atomicSub(T *p, U v) { const T o = *p; *p = T(o - T(v)); return o; }
template <class T, class U> static inline T atomicOr(T *p, U v) { const T o = *p; *p = T(o | T(v)); return o; }
template <class T, class U> static inline T atomicAnd(T *p, U v) { const T o = *p; *p = T(o & T(v)); return o; }
template <class T, class U> static inline T atomicXor(T *p, U v) { const T o = *p; *p = T(o ^ T(v)); return o; }
template <class T, class U> static inline T atomicMin(T *p, U v) { const T o = *p; if (T(v) < o) *p = T(v); return o; }
template <class T, class U> static inline T atomicMax(T *p, U v) { const T o = *p; if (T(v) > o) *p = T(v); return o; }
template <class T, class U> static inline T atomicExch(T *p, U v) { const T o = *p; *p = T(v); return o; }
template <class T, class U, class V> static inline T atomicCAS(T *p, U cmp, V v) { const T o = *p; if (o == T(cmp)) *p = T(v); return o; }

// ---- integer / float intrinsics ----------------------------------------------------------------------
static inline unsigned __funnelshift_l(unsigned lo, unsigned hi, unsigned sh) { return unsigned((((unsigned long long)hi << 32 | lo) << (sh & 31u)) >> 32); }
static inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned sh) { return unsigned(((unsigned long long)hi << 32 | lo) >> (sh & 31u)); }
static inline unsigned __funnelshift_lc(unsigned lo, unsigned hi, unsigned sh) { sh = sh > 32 ? 32 : sh; return sh == 32 ? lo : __funnelshift_l(lo, hi, sh); }
static inline unsigned __funnelshift_rc(unsigned lo, unsigned hi, unsigned sh) { sh = sh > 32 ? 32 : sh; return sh == 32 ? hi : __funnelshift_r(lo, hi, sh); }
static inline unsigned __brev(unsigned x)
{
    x = (x >> 16) | (x << 16);
    x = ((x & 0xFF00FF00u) >> 8) | ((x & 0x00FF00FFu) << 8);
    x = ((x & 0xF0F0F0F0u) >> 4) | ((x & 0x0F0F0F0Fu) << 4);
    x = ((x & 0xCCCCCCCCu) >> 2) | ((x & 0x33333333u) << 2);
    return ((x & 0xAAAAAAAAu) >> 1) | ((x & 0x55555555u) << 1);
}
static inline int __ffs(int x) { return __builtin_ffs(x); }
static inline int __ffsll(long long x) { return __builtin_ffsll(x); }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int __popcll(unsigned long long x) { return __builtin_popcountll(x); }
static inline int __clz(int x) { return x ? __builtin_clz(unsigned(x)) : 32; }
static inline int __clzll(long long x) { return x ? __builtin_clzll((unsigned long long)x) : 64; }
static inline unsigned __byte_perm(unsigned a, unsigned b, unsigned s)
{
    const unsigned long long v = (unsigned long long)b << 32 | a;
    unsigned r = 0;
    for (int k = 0; k < 4; ++k) {
        const unsigned sel = (s >> (4 * k)) & 0xFu;
        unsigned byte = unsigned(v >> (8 * (sel & 7u))) & 0xFFu;
        if (sel & 8u) byte = (byte & 0x80u) ? 0xFFu : 0u;
        r |= byte << (8 * k);
    }
    return r;
}
static inline unsigned __umulhi(unsigned a, unsigned b) { return unsigned(((unsigned long long)a * b) >> 32); }
template <class T> static inline T __ldg(const T *p) { return *p; }
template <class T> static inline T __ldcs(const T *p) { return *p; }
template <class T> static inline T __ldcg(const T *p) { return *p; }
template <class T> static inline void __stcs(T *p, T v) { *p = v; }
// IEEE single/double operations with explicit rounding: the emulator is built with -ffp-contract=off on x86-64 (SSE2)
static inline float __fadd_rn(float a, float b) { volatile float r = a + b; return r; }
static inline float __fsub_rn(float a, float b) { volatile float r = a - b; return r; }
static inline float __fmul_rn(float a, float b) { volatile float r = a * b; return r; }
static inline float __fdiv_rn(float a, float b) { volatile float r = a / b; return r; }
static inline float __fsqrt_rn(float a) { return sqrtf(a); }
static inline float __fmaf_rn(float a, float b, float c) { return fmaf(a, b, c); }
static inline double __dadd_rn(double a, double b) { volatile double r = a + b; return r; }
static inline double __dsub_rn(double a, double b) { volatile double r = a - b; return r; }
static inline double __dmul_rn(double a, double b) { volatile double r = a * b; return r; }
static inline double __ddiv_rn(double a, double b) { volatile double r = a / b; return r; }
static inline double __dsqrt_rn(double a) { return sqrt(a); }
static inline float __double2float_rn(double a) { volatile float r = float(a); return r; }
static inline unsigned __float2uint_rz(float x) { return x <= 0.f ? 0u : x >= 4294967296.f ? 0xFFFFFFFFu : unsigned(x); }
static inline int __float2int_rz(float x) { return int(x); }
static inline int __float2int_rn(float x) { return int(nearbyintf(x)); }
static inline int __double2int_rz(double x) { return x != x ? 0 : x >= 2147483648.0 ? 2147483647 : x <= -2147483649.0 ? (-2147483647 - 1) : int(x); }
static inline int __double2int_rd(double x) { return int(floor(x)); }
static inline int __double2int_ru(double x) { return int(ceil(x)); }
static inline int __double2int_rn(double x) { return int(nearbyint(x)); }
static inline long long __double2ll_rz(double x) { return (long long)x; }
static inline double __int2double_rn(int x) { return double(x); }
static inline double __uint2double_rn(unsigned x) { return double(x); }
static inline float __int2float_rn(int x) { return float(x); }
static inline float __uint2float_rn(unsigned x) { return float(x); }
static inline unsigned __float_as_uint(float f) { unsigned u; std::memcpy(&u, &f, 4); return u; }
static inline float __uint_as_float(unsigned u) { float f; std::memcpy(&f, &u, 4); return f; }
static inline int __float_as_int(float f) { int u; std::memcpy(&u, &f, 4); return u; }
static inline long long __double_as_longlong(double d) { long long u; std::memcpy(&u, &d, 8); return u; }
static inline double __longlong_as_double(long long u) { double d; std::memcpy(&d, &u, 8); return d; }
static inline size_t __cvta_generic_to_shared(const void *p) { return size_t(p); }

// CUDA's overloaded min/max (mixed signedness promotes to unsigned, like the device headers)
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }
static inline unsigned min(unsigned a, unsigned b) { return a < b ? a : b; }
static inline unsigned max(unsigned a, unsigned b) { return a > b ? a : b; }
static inline unsigned min(unsigned a, int b) { return a < unsigned(b) ? a : unsigned(b); }
static inline unsigned min(int a, unsigned b) { return unsigned(a) < b ? unsigned(a) : b; }
static inline unsigned max(unsigned a, int b) { return a > unsigned(b) ? a : unsigned(b); }
static inline unsigned max(int a, unsigned b) { return unsigned(a) > b ? unsigned(a) : b; }
static inline long min(long a, long b) { return a < b ? a : b; }
static inline long max(long a, long b) { return a > b ? a : b; }
static inline unsigned long min(unsigned long a, unsigned long b) { return a < b ? a : b; }
static inline unsigned long max(unsigned long a, unsigned long b) { return a > b ? a : b; }
static inline long long min(long long a, long long b) { return a < b ? a : b; }
static inline long long max(long long a, long long b) { return a > b ? a : b; }
static inline unsigned long long min(unsigned long long a, unsigned long long b) { return a < b ? a : b; }
static inline unsigned long long max(unsigned long long a, unsigned long long b) { return a > b ? a : b; }
static inline float min(float a, float b) { return fminf(a, b); }
static inline float max(float a, float b) { return fmaxf(a, b); }
static inline double min(double a, double b) { return fmin(a, b); }
static inline double max(double a, double b) { return fmax(a, b); }

// ---- runtime API stubs ------------------------------------------------------------------------------
enum cudaError_t {
    cudaSuccess = 0, cudaErrorInvalidValue = 1, cudaErrorMemoryAllocation = 2, cudaErrorInvalidConfiguration = 9,
    cudaErrorNotReady = 600, cudaErrorUnknown = 999
};
typedef struct emuStream_st *cudaStream_t;
typedef struct emuEvent_st *cudaEvent_t;
enum cudaMemcpyKind { cudaMemcpyHostToHost = 0, cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
enum { cudaStreamNonBlocking = 1, cudaEventDisableTiming = 2, cudaEventDefault = 0, cudaHostAllocMapped = 2, cudaHostAllocDefault = 0,
       cudaHostAllocPortable = 1, cudaHostRegisterDefault = 0, cudaHostRegisterMapped = 2, cudaHostRegisterPortable = 1 };
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize = 8, cudaFuncAttributePreferredSharedMemoryCarveout = 9 };
enum cudaMemoryType { cudaMemoryTypeUnregistered = 0, cudaMemoryTypeHost = 1, cudaMemoryTypeDevice = 2, cudaMemoryTypeManaged = 3 };
struct cudaPointerAttributes { cudaMemoryType type; int device; void *devicePointer; void *hostPointer; };
struct cudaDeviceProp {
    char name[256];
    int major, minor, multiProcessorCount;
    size_t totalGlobalMem, sharedMemPerBlockOptin;
    int l2CacheSize, persistingL2CacheMaxSize;
};
static inline const char *cudaGetErrorString(cudaError_t e) { return e == cudaSuccess ? "no error" : "emulated CUDA error"; }
static inline cudaError_t cudaGetLastError() { return cudaSuccess; }
static inline cudaError_t cudaPeekAtLastError() { return cudaSuccess; }
static inline cudaError_t cudaSetDevice(int) { return cudaSuccess; }
static inline cudaError_t cudaGetDevice(int *d) { *d = 0; return cudaSuccess; }
static inline cudaError_t cudaGetDeviceCount(int *n) { *n = 8; return cudaSuccess; }     // several "devices" for the sharding tests
static inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
static inline cudaError_t cudaDeviceCanAccessPeer(int *can, int, int) { *can = 1; return cudaSuccess; }
static inline cudaError_t cudaDeviceEnablePeerAccess(int, unsigned) { return cudaSuccess; }
static inline cudaError_t cudaGetDeviceProperties(cudaDeviceProp *p, int)
{
    std::memset(p, 0, sizeof(*p));
    std::snprintf(p->name, sizeof(p->name), "SIMT emulator (tests/emu)");
    p->major = 10; p->minor = 0; p->multiProcessorCount = emu::sm_count();
    p->totalGlobalMem = size_t(180) << 30; p->sharedMemPerBlockOptin = 227 * 1024; p->l2CacheSize = 126 << 20;
    return cudaSuccess;
}
static inline cudaError_t cudaStreamCreate(cudaStream_t *s) { *s = reinterpret_cast<cudaStream_t>(new char); return cudaSuccess; }
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned) { return cudaStreamCreate(s); }
static inline cudaError_t cudaStreamCreateWithPriority(cudaStream_t *s, unsigned, int) { return cudaStreamCreate(s); }
static inline cudaError_t cudaDeviceGetStreamPriorityRange(int *lo, int *hi) { *lo = 0; *hi = -1; return cudaSuccess; }
static inline cudaError_t cudaStreamDestroy(cudaStream_t s) { delete reinterpret_cast<char *>(s); return cudaSuccess; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaStreamQuery(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned = 0) { return cudaSuccess; }
static inline cudaError_t cudaEventCreate(cudaEvent_t *e) { *e = reinterpret_cast<cudaEvent_t>(new char); return cudaSuccess; }
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t *e, unsigned) { return cudaEventCreate(e); }
static inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete reinterpret_cast<char *>(e); return cudaSuccess; }
static inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t = nullptr) { return cudaSuccess; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventQuery(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t, cudaEvent_t) { *ms = 0.001f; return cudaSuccess; }
template <class T> static inline cudaError_t cudaMalloc(T **p, size_t n)
{
    void *q = nullptr;
    if (posix_memalign(&q, 256, n ? n : 1)) return cudaErrorMemoryAllocation;
    std::memset(q, 0xA5, n);                 // device memory is not zeroed: make reliance on that visible
    *p = static_cast<T *>(q);
    return cudaSuccess;
}
static inline cudaError_t cudaFree(void *p) { std::free(p); return cudaSuccess; }
template <class T> static inline cudaError_t cudaHostAlloc(T **p, size_t n, unsigned)
{
    void *q = nullptr;
    if (posix_memalign(&q, 4096, n ? n : 1)) return cudaErrorMemoryAllocation;
    std::memset(q, 0, n);
    *p = static_cast<T *>(q);
    return cudaSuccess;
}
template <class T> static inline cudaError_t cudaMallocHost(T **p, size_t n) { return cudaHostAlloc(p, n, 0); }
static inline cudaError_t cudaFreeHost(void *p) { std::free(p); return cudaSuccess; }
template <class T> static inline cudaError_t cudaHostGetDevicePointer(T **d, void *h, unsigned) { *d = static_cast<T *>(h); return cudaSuccess; }
static inline cudaError_t cudaHostRegister(void *, size_t, unsigned) { return cudaSuccess; }
static inline cudaError_t cudaHostUnregister(void *) { return cudaSuccess; }
// tests mark a host range as "device memory" with adsbcu_emu_mark_device (emu_runtime.cpp); everything else is unregistered
extern "C" int adsbcu_emu_is_device(const void *p);
static inline cudaError_t cudaPointerGetAttributes(cudaPointerAttributes *a, const void *p)
{
    const bool dev = adsbcu_emu_is_device(p) != 0;
    a->type = dev ? cudaMemoryTypeDevice : cudaMemoryTypeUnregistered; a->device = 0;
    a->devicePointer = dev ? const_cast<void *>(p) : nullptr; a->hostPointer = dev ? nullptr : const_cast<void *>(p);
    return cudaSuccess;
}
static inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind) { if (n) std::memmove(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind, cudaStream_t = nullptr) { if (n) std::memmove(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemcpyPeerAsync(void *d, int, const void *s, int, size_t n, cudaStream_t = nullptr) { if (n) std::memmove(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaMemset(void *d, int v, size_t n) { if (n) std::memset(d, v, n); return cudaSuccess; }
static inline cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t = nullptr) { if (n) std::memset(d, v, n); return cudaSuccess; }
template <class T> static inline cudaError_t cudaGetSymbolAddress(void **p, T &sym) { *p = static_cast<void *>(&sym); return cudaSuccess; }
template <class T> static inline cudaError_t cudaMemcpyFromSymbol(void *d, const T &sym, size_t n) { std::memcpy(d, &sym, n); return cudaSuccess; }
template <class T> static inline cudaError_t cudaMemcpyToSymbol(T &sym, const void *s, size_t n) { std::memcpy(&sym, s, n); return cudaSuccess; }
template <class F> static inline cudaError_t cudaFuncSetAttribute(F, cudaFuncAttribute, int) { return cudaSuccess; }
