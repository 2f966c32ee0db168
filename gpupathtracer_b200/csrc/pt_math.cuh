// pt_math.cuh -- device vector math with a FIXED operation order.
//
// The parity contract (DESIGN.md "float mode") is IEEE-754 binary32 with every operation rounded
// separately: this library is compiled with -fmad=false -prec-div=true -prec-sqrt=true, so the
// expressions below evaluate exactly as written (left to right), which is how the reference's
// GLSL reads (common/*.glsl) and how the CPU checker in oracle/ evaluates them.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pt {

struct V3 { float x, y, z; };
__host__ __device__ __forceinline__ V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__host__ __device__ __forceinline__ V3 v3(float s) { return v3(s, s, s); }
__host__ __device__ __forceinline__ V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__host__ __device__ __forceinline__ V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__host__ __device__ __forceinline__ V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
__host__ __device__ __forceinline__ V3 operator/(V3 a, V3 b) { return v3(a.x / b.x, a.y / b.y, a.z / b.z); }
__host__ __device__ __forceinline__ V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
__host__ __device__ __forceinline__ V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
__host__ __device__ __forceinline__ V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
__host__ __device__ __forceinline__ V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
__host__ __device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// same operand order as glm::cross, which the GLSL cross() of the reference maps to
__host__ __device__ __forceinline__ V3 cross(V3 a, V3 b) {
    return v3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
__host__ __device__ __forceinline__ float length(V3 a) { return sqrtf(dot(a, a)); }
__host__ __device__ __forceinline__ float distance(V3 a, V3 b) { return length(a - b); }
__host__ __device__ __forceinline__ V3 normalize(V3 a) { return a * (1.0f / sqrtf(dot(a, a))); }
// GLSL min/max == IEEE minNum/maxNum == FMNMX
__host__ __device__ __forceinline__ float gmin(float a, float b) { return fminf(a, b); }
__host__ __device__ __forceinline__ float gmax(float a, float b) { return fmaxf(a, b); }
__host__ __device__ __forceinline__ float clampf(float x, float lo, float hi) { return gmin(gmax(x, lo), hi); }
__host__ __device__ __forceinline__ float mixf(float x, float y, float a) { return x * (1.0f - a) + y * a; }
__host__ __device__ __forceinline__ V3 mix(V3 x, V3 y, float a) {
    return v3(mixf(x.x, y.x, a), mixf(x.y, y.y, a), mixf(x.z, y.z, a));
}
__host__ __device__ __forceinline__ V3 reflect(V3 I, V3 N) { return I - N * dot(N, I) * 2.0f; }
__host__ __device__ __forceinline__ float nndot(V3 a, V3 b) { return gmax(dot(a, b), 0.0f); }
__host__ __device__ __forceinline__ float avdot(V3 a, V3 b) { return fabsf(dot(a, b)); }

#define PT_PI 3.141592653589793238462643383279f

__device__ __forceinline__ V3 xyz(float4 v) { return v3(v.x, v.y, v.z); }
__device__ __forceinline__ float4 f4(V3 v, float w) { return make_float4(v.x, v.y, v.z, w); }

// 128-bit read-only loads (LDG.E.128.CONSTANT)
__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }

// 256-bit read-only load (LDG.E.256.CONSTANT, new with sm_100): one request and one L1 data-pipe wavefront
// per lane for a 32-B half record instead of two. The traversal kernels are bound by exactly that pipe
// (l1tex__data_pipe_lsu_wavefronts ~ 89 % of peak with 128-bit loads, profiles/), so record loads use this.
struct F8 { float4 lo, hi; };
// PT_NODE_HINT (compile-time experiment): eviction priority of the BVH record loads. 0: none; 1: L1::evict_last;
// 2: L1::evict_last + L2::evict_last; 3: L2::evict_last (sm_100 accepts L2 priorities on 256-bit loads only).
#ifndef PT_NODE_HINT
#define PT_NODE_HINT 0
#endif
#if PT_NODE_HINT == 1
#define PT_LDG256 "ld.global.nc.L1::evict_last.v8.f32"
#elif PT_NODE_HINT == 2
#define PT_LDG256 "ld.global.nc.L1::evict_last.L2::evict_last.v8.f32"
#elif PT_NODE_HINT == 3
#define PT_LDG256 "ld.global.nc.L2::evict_last.v8.f32"
#else
#define PT_LDG256 "ld.global.nc.v8.f32"
#endif
__device__ __forceinline__ F8 ldg256(const void* p) {
    F8 r;
    asm(PT_LDG256 " {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(r.lo.x), "=f"(r.lo.y), "=f"(r.lo.z), "=f"(r.lo.w), "=f"(r.hi.x), "=f"(r.hi.y), "=f"(r.hi.z), "=f"(r.hi.w)
        : "l"(p));
    return r;
}

// Ray-queue / hit-record traffic of the traversal kernels is touched once per launch: with PT_STREAM_NA it bypasses
// L1 allocation (LDG.E.NA / STG.E.NA) so that it does not evict BVH records, which is what the L1 is for here.
#ifdef PT_STREAM_NA
__device__ __forceinline__ float4 ld_stream(const float4* p) {
    float4 a;
    asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "l"(p));
    return a;
}
__device__ __forceinline__ void st_stream(float4* p, float4 a) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w) : "memory");
}
#else
__device__ __forceinline__ float4 ld_stream(const float4* p) { return *p; }
__device__ __forceinline__ void st_stream(float4* p, float4 a) { *p = a; }
#endif

}  // namespace pt
