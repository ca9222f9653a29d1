// tests/emu/cuda_emu.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// A minimal single-OS-thread SIMT emulator so that the SAME kernel sources that nvcc
// compiles for sm_100a (acoss_b200/csrc/*.cu, *.cuh) can be compiled with g++ and stepped on
// the CPU-only build container: every CUDA thread of a block is a ucontext fiber, warps are
// scheduled lane-round-robin, __syncthreads / __shfl_*_sync / __ballot_sync are rendezvous
// points.  It exists to shake out indexing / synchronisation-logic bugs before a GPU call
// is spent (there is no GPU in the build container).  The resulting library
// (tests/emu/_build/libacoss_emu.so) is loaded ONLY by tests/test_emu_*.py; the product
// loader (acoss_b200/_native.py) never looks at it and fails loudly without the real CUDA
// library.  It models none of the hardware's memory-ordering hazards: racecheck /
// memcheck on the GPU box remain the authority.
#pragma once
#include <ucontext.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <vector>

#define ACOSS_EMU 1

// ---------------------------------------------------------------- qualifiers
#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __noinline__ inline
#define __restrict__
#define __launch_bounds__(...)
#define __shared__ static
#define __align__(n) __attribute__((aligned(n)))
#define __constant__ static

struct uint3 { unsigned x, y, z; };
struct dim3 {
    unsigned x, y, z;
    dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct int2 { int x, y; };
struct int4 { int x, y, z, w; };
struct uint2 { unsigned x, y; };
struct uint4 { unsigned x, y, z, w; };
struct float2 { float x, y; };
struct float4 { float x, y, z, w; };
struct double2 { double x, y; };
static inline int4 make_int4(int x, int y, int z, int w) { return int4{x, y, z, w}; }
static inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return uint4{x, y, z, w}; }

namespace emu {

struct Warp {
    uint64_t slot[32];
    int arrived = 0;
    unsigned gen = 0;
};

struct Block {
    int nthreads = 0;
    int bar_arrived = 0;
    unsigned bar_gen = 0;
    int named_arrived[16] = {0};
    unsigned named_gen[16] = {0};
    std::vector<Warp> warps;
};

struct Fiber {
    ucontext_t ctx;
    char* stack = nullptr;
    bool done = false;
    bool block_wait = false;
    const unsigned* wait_ptr = nullptr;  // parked while *wait_ptr == wait_val
    unsigned wait_val = 0;
    uint3 tid;
    int linear = 0;
};

struct State {
    ucontext_t main_ctx;
    Fiber* cur = nullptr;
    Block blk;
    uint3 bid;
    dim3 bdim, gdim;
    unsigned char* dyn_smem = nullptr;
    size_t dyn_smem_bytes = 0;
    const std::function<void()>* body = nullptr;
    std::vector<Fiber> fibers;
    long launches = 0;
};

inline State& st() {
    static State s;
    return s;
}

constexpr size_t kStackBytes = 256 * 1024;

inline void yield_() {
    State& s = st();
    Fiber* f = s.cur;
    swapcontext(&f->ctx, &s.main_ctx);
}

inline void fiber_entry() {
    State& s = st();
    (*s.body)();
    s.cur->done = true;
    // returning switches to uc_link (main_ctx)
}

inline void launch(dim3 grid, dim3 block, size_t smem, const std::function<void()>& body) {
    State& s = st();
    s.launches++;
    s.gdim = grid;
    s.bdim = block;
    s.body = &body;
    int nthreads = (int)(block.x * block.y * block.z);
    if (nthreads % 32 != 0) {
        fprintf(stderr, "emu: block size %d not a multiple of 32\n", nthreads);
        abort();
    }
    if ((int)s.fibers.size() < nthreads) {
        size_t old = s.fibers.size();
        s.fibers.resize(nthreads);
        for (size_t i = old; i < s.fibers.size(); ++i) s.fibers[i].stack = (char*)malloc(kStackBytes);
    }
    std::vector<unsigned char> smem_buf(smem + 64);
    // 16-byte align
    s.dyn_smem = (unsigned char*)(((uintptr_t)smem_buf.data() + 15) & ~(uintptr_t)15);
    s.dyn_smem_bytes = smem;
    int nwarps = nthreads / 32;
    for (unsigned bz = 0; bz < grid.z; ++bz)
        for (unsigned by = 0; by < grid.y; ++by)
            for (unsigned bx = 0; bx < grid.x; ++bx) {
                s.bid = uint3{bx, by, bz};
                s.blk.nthreads = nthreads;
                s.blk.bar_arrived = 0;
                s.blk.bar_gen = 0;
                for (int nb = 0; nb < 16; ++nb) s.blk.named_arrived[nb] = 0, s.blk.named_gen[nb] = 0;
                s.blk.warps.assign(nwarps, Warp());
                for (int t = 0; t < nthreads; ++t) {
                    Fiber& f = s.fibers[t];
                    f.done = false;
                    f.block_wait = false;
                    f.linear = t;
                    f.tid.x = t % block.x;
                    f.tid.y = (t / block.x) % block.y;
                    f.tid.z = t / (block.x * block.y);
                    getcontext(&f.ctx);
                    f.ctx.uc_stack.ss_sp = f.stack;
                    f.ctx.uc_stack.ss_size = kStackBytes;
                    f.ctx.uc_link = &s.main_ctx;
                    makecontext(&f.ctx, (void (*)())fiber_entry, 0);
                }
                int alive = nthreads;
                int guard = 0;
                while (alive > 0) {
                    bool any_progress = false;
                    for (int w = 0; w < nwarps; ++w) {
                        // run this warp until every lane is done or parked on the block barrier
                        bool progressed = true;
                        while (progressed) {
                            progressed = false;
                            for (int l = 0; l < 32; ++l) {
                                Fiber& f = s.fibers[w * 32 + l];
                                if (f.done) continue;
                                if (f.block_wait && *f.wait_ptr == f.wait_val) continue;
                                f.block_wait = false;
                                s.cur = &f;
                                swapcontext(&s.main_ctx, &f.ctx);
                                progressed = true;
                                any_progress = true;
                                if (f.done) {
                                    --alive;
                                }
                            }
                        }
                    }
                    if (!any_progress) {
                        if (++guard > 4) {
                            fprintf(stderr, "emu: deadlock in block (%u,%u,%u): %d threads alive, barrier arrived %d\n",
                                    bx, by, bz, alive, s.blk.bar_arrived);
                            abort();
                        }
                    } else {
                        guard = 0;
                    }
                }
            }
    s.cur = nullptr;
    s.body = nullptr;
    s.dyn_smem = nullptr;
}

inline void syncthreads() {
    State& s = st();
    Fiber* f = s.cur;
    unsigned gen = s.blk.bar_gen;
    if (++s.blk.bar_arrived == s.blk.nthreads) {
        s.blk.bar_arrived = 0;
        s.blk.bar_gen++;
        return;
    }
    while (s.blk.bar_gen == gen) {
        f->block_wait = true;
        f->wait_ptr = &s.blk.bar_gen;
        f->wait_val = gen;
        yield_();
    }
    f->block_wait = false;
}

// PTX bar.sync id, count / bar.arrive id, count (named barriers 1..15)
inline void bar_named(int id, int count, bool wait) {
    State& s = st();
    Fiber* f = s.cur;
    unsigned gen = s.blk.named_gen[id];
    if (++s.blk.named_arrived[id] == count) {
        s.blk.named_arrived[id] = 0;
        s.blk.named_gen[id]++;
        return;
    }
    if (!wait) return;
    while (s.blk.named_gen[id] == gen) {
        f->block_wait = true;
        f->wait_ptr = &s.blk.named_gen[id];
        f->wait_val = gen;
        yield_();
    }
    f->block_wait = false;
}

inline Warp& my_warp() {
    State& s = st();
    return s.blk.warps[s.cur->linear / 32];
}

inline void warp_barrier() {
    Warp& w = my_warp();
    unsigned gen = w.gen;
    if (++w.arrived == 32) {
        w.arrived = 0;
        w.gen++;
        return;
    }
    while (w.gen == gen) yield_();
}

inline int lane_id() { return st().cur->linear & 31; }

template <class T>
inline uint64_t to_bits(T v) {
    static_assert(sizeof(T) <= 8, "shuffle payload too large");
    uint64_t b = 0;
    memcpy(&b, &v, sizeof(T));
    return b;
}
template <class T>
inline T from_bits(uint64_t b) {
    T v;
    memcpy(&v, &b, sizeof(T));
    return v;
}

// all 32 lanes must call (full mask only)
template <class T>
inline T shfl_from(T v, int src, bool valid) {
    Warp& w = my_warp();
    w.slot[lane_id()] = to_bits(v);
    warp_barrier();
    T r = valid ? from_bits<T>(w.slot[src & 31]) : v;
    warp_barrier();
    return r;
}

}  // namespace emu

#define threadIdx (emu::st().cur->tid)
#define blockIdx (emu::st().bid)
#define blockDim (emu::st().bdim)
#define gridDim (emu::st().gdim)
#define warpSize 32

inline void __syncthreads() { emu::syncthreads(); }
inline void __syncwarp(unsigned = 0xffffffffu) { emu::warp_barrier(); }
inline void __threadfence() {}
inline void __threadfence_block() {}

template <class T>
inline T __shfl_sync(unsigned, T v, int src, int width = 32) {
    int lane = emu::lane_id();
    int base = lane & ~(width - 1);
    return emu::shfl_from(v, base + (src & (width - 1)), true);
}
template <class T>
inline T __shfl_up_sync(unsigned, T v, unsigned delta, int width = 32) {
    int lane = emu::lane_id();
    int base = lane & ~(width - 1);
    int src = lane - (int)delta;
    return emu::shfl_from(v, src, src >= base);
}
template <class T>
inline T __shfl_down_sync(unsigned, T v, unsigned delta, int width = 32) {
    int lane = emu::lane_id();
    int base = lane & ~(width - 1);
    int src = lane + (int)delta;
    return emu::shfl_from(v, src, src < base + width);
}
template <class T>
inline T __shfl_xor_sync(unsigned, T v, int mask, int width = 32) {
    int lane = emu::lane_id();
    int base = lane & ~(width - 1);
    int src = lane ^ mask;
    return emu::shfl_from(v, src, src >= base && src < base + width);
}
inline unsigned __ballot_sync(unsigned, int pred) {
    emu::Warp& w = emu::my_warp();
    w.slot[emu::lane_id()] = pred ? 1 : 0;
    emu::warp_barrier();
    unsigned r = 0;
    for (int l = 0; l < 32; ++l) r |= (unsigned)(w.slot[l] & 1) << l;
    emu::warp_barrier();
    return r;
}
inline int __reduce_add_sync(unsigned, int v) {
    emu::Warp& w = emu::my_warp();
    w.slot[emu::lane_id()] = (uint64_t)(uint32_t)v;
    emu::warp_barrier();
    int r = 0;
    for (int l = 0; l < 32; ++l) r += (int)(uint32_t)w.slot[l];
    emu::warp_barrier();
    return r;
}
inline int __reduce_max_sync(unsigned, int v) {
    emu::Warp& w = emu::my_warp();
    w.slot[emu::lane_id()] = (uint64_t)(uint32_t)v;
    emu::warp_barrier();
    int r = (int)(uint32_t)w.slot[0];
    for (int l = 1; l < 32; ++l) r = std::max(r, (int)(uint32_t)w.slot[l]);
    emu::warp_barrier();
    return r;
}
inline unsigned __reduce_min_sync(unsigned, unsigned v) {
    emu::Warp& w = emu::my_warp();
    w.slot[emu::lane_id()] = (uint64_t)v;
    emu::warp_barrier();
    unsigned r = (unsigned)w.slot[0];
    for (int l = 1; l < 32; ++l) r = std::min(r, (unsigned)w.slot[l]);
    emu::warp_barrier();
    return r;
}
inline unsigned __reduce_max_sync(unsigned, unsigned v) {
    emu::Warp& w = emu::my_warp();
    w.slot[emu::lane_id()] = (uint64_t)v;
    emu::warp_barrier();
    unsigned r = (unsigned)w.slot[0];
    for (int l = 1; l < 32; ++l) r = std::max(r, (unsigned)w.slot[l]);
    emu::warp_barrier();
    return r;
}
// SIMD-in-word intrinsics used by the packed counting pass
inline unsigned __byte_perm(unsigned x, unsigned y, unsigned sel) {
    const uint64_t w = ((uint64_t)y << 32) | x;
    unsigned r = 0;
    for (int i = 0; i < 4; ++i) r |= (unsigned)((w >> (8 * ((sel >> (4 * i)) & 7))) & 0xff) << (8 * i);
    return r;
}
inline unsigned __vneg2(unsigned a) {
    const unsigned lo = (unsigned)(-(int)(int16_t)(a & 0xffff)) & 0xffffu, hi = (unsigned)(-(int)(int16_t)(a >> 16)) & 0xffffu;
    return (hi << 16) | lo;
}
inline unsigned __vadd2(unsigned a, unsigned b) { return (((a >> 16) + (b >> 16)) << 16) | ((a + b) & 0xffffu); }
inline unsigned __viaddmin_s16x2_relu(unsigned a, unsigned b, unsigned c) {
    unsigned r = 0;
    for (int h = 0; h < 2; ++h) {
        int v = (int)(int16_t)((int16_t)(a >> (16 * h)) + (int16_t)(b >> (16 * h)));  // wraps like the hardware add
        const int cc = (int16_t)(c >> (16 * h));
        v = v < cc ? v : cc;
        v = v > 0 ? v : 0;
        r |= ((unsigned)v & 0xffffu) << (16 * h);
    }
    return r;
}
inline int __double2loint(double v) { return (int)(uint32_t)(emu::to_bits(v) & 0xffffffffu); }
inline int __double2hiint(double v) { return (int)(uint32_t)(emu::to_bits(v) >> 32); }
inline int __any_sync(unsigned m, int pred) { return __ballot_sync(m, pred) != 0; }
inline int __all_sync(unsigned m, int pred) { return __ballot_sync(m, pred) == 0xffffffffu; }

inline int __popc(unsigned v) { return __builtin_popcount(v); }
inline int __popcll(unsigned long long v) { return __builtin_popcountll(v); }
inline int __clz(int v) { return v == 0 ? 32 : __builtin_clz((unsigned)v); }
inline int __ffs(int v) { return __builtin_ffs(v); }
inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned shift) {
    uint64_t x = ((uint64_t)hi << 32) | lo;
    return (unsigned)(x >> (shift & 31));
}
inline unsigned __funnelshift_l(unsigned lo, unsigned hi, unsigned shift) {
    uint64_t x = ((uint64_t)hi << 32) | lo;
    return (unsigned)((x << (shift & 31)) >> 32);
}
template <class T>
inline T __ldg(const T* p) { return *p; }

inline double __dmul_rn(double a, double b) { volatile double r = a * b; return r; }
inline double __dadd_rn(double a, double b) { volatile double r = a + b; return r; }
inline double __fma_rn(double a, double b, double c) { return std::fma(a, b, c); }
inline float __fmaf_rn(float a, float b, float c) { return std::fmaf(a, b, c); }
inline float __fadd_rn(float a, float b) { volatile float r = a + b; return r; }
inline float __fdiv_rn(float a, float b) { volatile float r = a / b; return r; }
inline double __ddiv_rn(double a, double b) { volatile double r = a / b; return r; }
inline float __fmul_rn(float a, float b) { volatile float r = a * b; return r; }
inline float __fdividef(float a, float b) { return a / b; }
inline float __frcp_rn(float a) { return 1.0f / a; }
inline int __double2int_rn(double v) { return (int)std::nearbyint(v); }
inline double __longlong_as_double(long long v) { return emu::from_bits<double>((uint64_t)v); }
inline long long __double_as_longlong(double v) { return (long long)emu::to_bits(v); }
inline float __int_as_float(int v) { return emu::from_bits<float>((uint64_t)(uint32_t)v); }
inline int __float_as_int(float v) { return (int)emu::to_bits(v); }

template <class T>
inline T atomicAdd(T* p, T v) { T o = *p; *p = o + v; return o; }
template <class T>
inline T atomicMax(T* p, T v) { T o = *p; if (v > o) *p = v; return o; }
template <class T>
inline T atomicOr(T* p, T v) { T o = *p; *p = o | v; return o; }
template <class T>
inline T atomicAnd(T* p, T v) { T o = *p; *p = o & v; return o; }

using std::max;
using std::min;

// ---------------------------------------------------------------- runtime subset
typedef int cudaError_t;
#define cudaSuccess 0
#define cudaErrorInvalidValue 1
#define cudaErrorMemoryAllocation 2
typedef struct emuStream_* cudaStream_t;
typedef struct emuEvent_ { double t; }* cudaEvent_t;
enum cudaMemcpyKind { cudaMemcpyHostToHost, cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
enum cudaDeviceAttr { cudaDevAttrMultiProcessorCount = 16, cudaDevAttrMaxSharedMemoryPerBlockOptin = 97 };
#define cudaStreamNonBlocking 1
#define cudaEventDefault 0
#define cudaHostAllocDefault 0

inline const char* cudaGetErrorString(cudaError_t e) { return e == 0 ? "no error" : "emu error"; }
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
inline cudaError_t cudaPeekAtLastError() { return cudaSuccess; }
inline cudaError_t cudaSetDevice(int) { return cudaSuccess; }
inline cudaError_t cudaGetDeviceCount(int* n) { *n = 1; return cudaSuccess; }
inline cudaError_t cudaGetDevice(int* d) { *d = 0; return cudaSuccess; }
template <class T>
inline cudaError_t cudaGetSymbolAddress(void** p, T& sym) { *p = (void*)&sym; return cudaSuccess; }
inline cudaError_t cudaDeviceGetAttribute(int* v, cudaDeviceAttr a, int) {
    *v = (a == cudaDevAttrMultiProcessorCount) ? 2 : 227 * 1024;
    return cudaSuccess;
}
inline cudaError_t cudaMalloc(void** p, size_t n) { *p = calloc(1, n ? n : 1); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
template <class T>
inline cudaError_t cudaMalloc(T** p, size_t n) { return cudaMalloc((void**)p, n); }
inline cudaError_t cudaFree(void* p) { free(p); return cudaSuccess; }
inline cudaError_t cudaMallocHost(void** p, size_t n) { return cudaMalloc(p, n); }
template <class T>
inline cudaError_t cudaMallocHost(T** p, size_t n) { return cudaMalloc((void**)p, n); }
inline cudaError_t cudaFreeHost(void* p) { free(p); return cudaSuccess; }
inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { if (n) memcpy(d, s, n); return cudaSuccess; }
inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t = 0) { if (n) memcpy(d, s, n); return cudaSuccess; }
inline cudaError_t cudaMemset(void* d, int v, size_t n) { if (n) memset(d, v, n); return cudaSuccess; }
inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t = 0) { if (n) memset(d, v, n); return cudaSuccess; }
inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned) { *s = nullptr; return cudaSuccess; }
inline cudaError_t cudaStreamCreate(cudaStream_t* s) { *s = nullptr; return cudaSuccess; }
inline cudaError_t cudaStreamCreateWithPriority(cudaStream_t* s, unsigned, int) { *s = nullptr; return cudaSuccess; }
inline cudaError_t cudaDeviceGetStreamPriorityRange(int* lo, int* hi) { *lo = 0; *hi = 0; return cudaSuccess; }
inline cudaError_t cudaStreamWaitEvent(cudaStream_t, struct emuEvent_*, unsigned) { return cudaSuccess; }
inline cudaError_t cudaStreamDestroy(cudaStream_t) { return cudaSuccess; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = new emuEvent_{0.0}; return cudaSuccess; }
#define cudaEventDisableTiming 2
inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned) { *e = new emuEvent_{0.0}; return cudaSuccess; }
inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return cudaSuccess; }
inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t = 0) { return cudaSuccess; }
inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t, cudaEvent_t) { *ms = 0.f; return cudaSuccess; }
template <class F>
inline cudaError_t cudaFuncSetAttribute(F, cudaFuncAttribute, int) { return cudaSuccess; }
