// Shared helpers for libdlrmx (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>

#include "../../include/dlrmx.h"

namespace dlrmx {

void set_error(const char* fmt, ...);
extern std::atomic<int64_t> g_launches;

inline int check_launch(const char* what) {
    g_launches.fetch_add(1, std::memory_order_relaxed);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return DLRMX_E_CUDA;
    }
    return DLRMX_OK;
}

#define DLRMX_REQUIRE(cond, ...)           \
    do {                                   \
        if (!(cond)) {                     \
            dlrmx::set_error(__VA_ARGS__); \
            return DLRMX_E_ARG;            \
        }                                  \
    } while (0)

constexpr int kSMs = 148;  // B200

// Per-device one-time setup (cudaFuncSetAttribute opt-ins are per device): first() is true once per CUDA device.
struct DeviceOnce {
    std::atomic<bool> done[64];
    DeviceOnce() { for (auto& d : done) d.store(false); }
    bool first() {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return true;
        return !done[dev].exchange(true);
    }
};

// Test hooks (dlrmx_set_tunable); 0 = default.  2: force the CUDA-core interaction kernels; 3: sorted positions per
// group of the fused update (32/64/128/256); 10: cp.async staging instead of TMA bulk copies in the interaction kernels;
// 11: CTAs per SM of the interaction kernels' persistent grid.  They select between code paths that ship anyway (tests
// compare them bit for bit); the kernel-variant sweeps of round 1 are gone (numbers in DESIGN.md).
enum { DLRMX_TUNE_BWD_SEG = 3, DLRMX_TUNE_COUNT = 16 };
int get_tunable(int which);

__host__ __device__ inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- vector access ---------------------------------------------------------------------------
// Rows of the big tables are read exactly once per step: keep them out of L1 and mark them
// evict-first in L2 so the pooled outputs / gradients (re-read by the next kernel) stay resident.
template <int VEC>
struct RowVec;

template <>
struct RowVec<4> {
    float4 v;
    __device__ __forceinline__ void zero() { v = make_float4(0.f, 0.f, 0.f, 0.f); }
    __device__ __forceinline__ void load_stream(const float* p) {
        asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                     : "l"(p));
    }
    __device__ __forceinline__ void load(const float* p) { v = *reinterpret_cast<const float4*>(p); }
    // bf16 table rows (fp32 everywhere else): 4 elements = one 8-byte access, widened exactly / rounded to nearest even
    __device__ __forceinline__ void widen(uint2 r) {
        v.x = __uint_as_float(r.x << 16); v.y = __uint_as_float(r.x & 0xffff0000u);
        v.z = __uint_as_float(r.y << 16); v.w = __uint_as_float(r.y & 0xffff0000u);
    }
    __device__ __forceinline__ void load_stream(const __nv_bfloat16* p) {
        uint2 r;
        asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
        widen(r);
    }
    __device__ __forceinline__ void load(const __nv_bfloat16* p) { widen(*reinterpret_cast<const uint2*>(p)); }
    __device__ __forceinline__ void store(__nv_bfloat16* p) const {
        __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
        uint2 r;
        r.x = *reinterpret_cast<uint32_t*>(&lo); r.y = *reinterpret_cast<uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(p) = r;
    }
    __device__ __forceinline__ void load_ro(const float* p) { v = __ldg(reinterpret_cast<const float4*>(p)); }
    __device__ __forceinline__ void load_unaligned(const float* p) { v = make_float4(p[0], p[1], p[2], p[3]); }
    __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float4*>(p) = v; }
    __device__ __forceinline__ void add(const RowVec& o) {
        v.x = __fadd_rn(v.x, o.v.x); v.y = __fadd_rn(v.y, o.v.y);
        v.z = __fadd_rn(v.z, o.v.z); v.w = __fadd_rn(v.w, o.v.w);
    }
    __device__ __forceinline__ void add_scaled(const RowVec& o, float w) {  // acc + (row*w), two roundings
        v.x = __fadd_rn(v.x, __fmul_rn(o.v.x, w)); v.y = __fadd_rn(v.y, __fmul_rn(o.v.y, w));
        v.z = __fadd_rn(v.z, __fmul_rn(o.v.z, w)); v.w = __fadd_rn(v.w, __fmul_rn(o.v.w, w));
    }
    __device__ __forceinline__ float sumsq() const {
        return __fadd_rn(__fadd_rn(__fmul_rn(v.x, v.x), __fmul_rn(v.y, v.y)),
                         __fadd_rn(__fmul_rn(v.z, v.z), __fmul_rn(v.w, v.w)));
    }
    template <typename F>
    __device__ __forceinline__ void map(const RowVec& g, F f) {  // v = f(v, g) elementwise
        v.x = f(v.x, g.v.x); v.y = f(v.y, g.v.y); v.z = f(v.z, g.v.z); v.w = f(v.w, g.v.w);
    }
};

// 8 bf16 elements as loaded (4 registers): rows in flight stay packed and are widened when they are added
struct PackedRow8 {
    uint4 r;
    __device__ __forceinline__ void load_stream(const __nv_bfloat16* p) {
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    }
};

// 8 elements per lane: the lane geometry of bf16 table rows (one 16-byte access = 8 bf16), fp32 everywhere else (two
// 16-byte accesses).  Same arithmetic as RowVec<4>, element by element.
template <>
struct RowVec<8> {
    float4 a, b;
    __device__ __forceinline__ void zero() { a = make_float4(0.f, 0.f, 0.f, 0.f); b = a; }
    __device__ __forceinline__ void widen(uint4 r) {
        a.x = __uint_as_float(r.x << 16); a.y = __uint_as_float(r.x & 0xffff0000u);
        a.z = __uint_as_float(r.y << 16); a.w = __uint_as_float(r.y & 0xffff0000u);
        b.x = __uint_as_float(r.z << 16); b.y = __uint_as_float(r.z & 0xffff0000u);
        b.z = __uint_as_float(r.w << 16); b.w = __uint_as_float(r.w & 0xffff0000u);
    }
    __device__ __forceinline__ void load_stream(const __nv_bfloat16* p) {
        uint4 r;
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
        widen(r);
    }
    __device__ __forceinline__ void load(const __nv_bfloat16* p) { widen(*reinterpret_cast<const uint4*>(p)); }
    __device__ __forceinline__ void store(__nv_bfloat16* p) const {
        __nv_bfloat162 q0 = __floats2bfloat162_rn(a.x, a.y), q1 = __floats2bfloat162_rn(a.z, a.w);
        __nv_bfloat162 q2 = __floats2bfloat162_rn(b.x, b.y), q3 = __floats2bfloat162_rn(b.z, b.w);
        uint4 r;
        r.x = *reinterpret_cast<uint32_t*>(&q0); r.y = *reinterpret_cast<uint32_t*>(&q1);
        r.z = *reinterpret_cast<uint32_t*>(&q2); r.w = *reinterpret_cast<uint32_t*>(&q3);
        *reinterpret_cast<uint4*>(p) = r;
    }
    __device__ __forceinline__ void load_stream(const float* p) {
        asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "l"(p));
        asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p + 4));
    }
    __device__ __forceinline__ void load(const float* p) {
        a = *reinterpret_cast<const float4*>(p); b = *reinterpret_cast<const float4*>(p + 4);
    }
    __device__ __forceinline__ void load_ro(const float* p) {
        a = __ldg(reinterpret_cast<const float4*>(p)); b = __ldg(reinterpret_cast<const float4*>(p + 4));
    }
    __device__ __forceinline__ void store(float* p) const {
        *reinterpret_cast<float4*>(p) = a; *reinterpret_cast<float4*>(p + 4) = b;
    }
    __device__ __forceinline__ void add(const RowVec& o) {
        a.x = __fadd_rn(a.x, o.a.x); a.y = __fadd_rn(a.y, o.a.y); a.z = __fadd_rn(a.z, o.a.z); a.w = __fadd_rn(a.w, o.a.w);
        b.x = __fadd_rn(b.x, o.b.x); b.y = __fadd_rn(b.y, o.b.y); b.z = __fadd_rn(b.z, o.b.z); b.w = __fadd_rn(b.w, o.b.w);
    }
    __device__ __forceinline__ void add_scaled(const RowVec& o, float w) {  // acc + (row*w), two roundings
        a.x = __fadd_rn(a.x, __fmul_rn(o.a.x, w)); a.y = __fadd_rn(a.y, __fmul_rn(o.a.y, w));
        a.z = __fadd_rn(a.z, __fmul_rn(o.a.z, w)); a.w = __fadd_rn(a.w, __fmul_rn(o.a.w, w));
        b.x = __fadd_rn(b.x, __fmul_rn(o.b.x, w)); b.y = __fadd_rn(b.y, __fmul_rn(o.b.y, w));
        b.z = __fadd_rn(b.z, __fmul_rn(o.b.z, w)); b.w = __fadd_rn(b.w, __fmul_rn(o.b.w, w));
    }
    __device__ __forceinline__ float sumsq() const {
        return __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(a.x, a.x), __fmul_rn(a.y, a.y)), __fadd_rn(__fmul_rn(a.z, a.z), __fmul_rn(a.w, a.w))),
                         __fadd_rn(__fadd_rn(__fmul_rn(b.x, b.x), __fmul_rn(b.y, b.y)), __fadd_rn(__fmul_rn(b.z, b.z), __fmul_rn(b.w, b.w))));
    }
    template <typename F>
    __device__ __forceinline__ void map(const RowVec& g, F f) {  // v = f(v, g) elementwise
        a.x = f(a.x, g.a.x); a.y = f(a.y, g.a.y); a.z = f(a.z, g.a.z); a.w = f(a.w, g.a.w);
        b.x = f(b.x, g.b.x); b.y = f(b.y, g.b.y); b.z = f(b.z, g.b.z); b.w = f(b.w, g.b.w);
    }
    __device__ __forceinline__ void add(const PackedRow8& o) { RowVec t; t.widen(o.r); add(t); }
    __device__ __forceinline__ void add_scaled(const PackedRow8& o, float w) { RowVec t; t.widen(o.r); add_scaled(t, w); }
};

template <>
struct RowVec<1> {
    float v;
    __device__ __forceinline__ void zero() { v = 0.f; }
    __device__ __forceinline__ void load_stream(const float* p) { v = __ldg(p); }
    __device__ __forceinline__ void load(const float* p) { v = *p; }
    __device__ __forceinline__ void load_stream(const __nv_bfloat16* p) { v = __bfloat162float(*p); }
    __device__ __forceinline__ void load(const __nv_bfloat16* p) { v = __bfloat162float(*p); }
    __device__ __forceinline__ void store(__nv_bfloat16* p) const { *p = __float2bfloat16_rn(v); }
    __device__ __forceinline__ void load_ro(const float* p) { v = __ldg(p); }
    __device__ __forceinline__ void load_unaligned(const float* p) { v = *p; }
    __device__ __forceinline__ void store(float* p) const { *p = v; }
    __device__ __forceinline__ void add(const RowVec& o) { v = __fadd_rn(v, o.v); }
    __device__ __forceinline__ void add_scaled(const RowVec& o, float w) { v = __fadd_rn(v, __fmul_rn(o.v, w)); }
    __device__ __forceinline__ float sumsq() const { return __fmul_rn(v, v); }
    template <typename F>
    __device__ __forceinline__ void map(const RowVec& g, F f) { v = f(v, g.v); }
};

// Lane geometry for a row of D floats: groups of G lanes (power of two <= 32), each lane owning CH
// vectors of VEC floats at columns (lane + c*G)*VEC.
struct RowGeom {
    int vec, g, ch;
};
// max_vec: 4 (fp32 rows: 16 bytes per lane) or 8 (bf16 rows: 16 bytes per lane as well)
inline bool row_geometry(int D, RowGeom* out, int max_vec = 4) {
    if (D <= 0) return false;
    int vec = (max_vec == 8 && D % 8 == 0) ? 8 : ((D % 4 == 0 && max_vec != 8) ? 4 : 1);
    int nv = D / vec;
    int g = 1;
    while (g < nv && g < 32) g <<= 1;
    int ch = (nv + g - 1) / g;
    if (ch > 4) return false;
    if (ch == 3) ch = 4;
    out->vec = vec; out->g = g; out->ch = ch;
    return true;
}

}  // namespace dlrmx
