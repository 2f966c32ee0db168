// tools/l1_microbench.cu -- DEV TOOL: cost of fetching scattered 64-B records through L1 on sm_100a for the
// access patterns considered for BVH node fetches. Each warp chases pointers (next index comes from the loaded
// record) like a traversal does; 40 warps/SM, table sized to be L2-resident with ~75 % L1 hits not guaranteed.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

struct __align__(64) Rec { float4 q[4]; };

__device__ __forceinline__ void ld256(const void* p, float4& a, float4& b) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
}

template <int MODE>
__global__ void __launch_bounds__(128) k(const Rec* __restrict__ t, int n, int steps, float* out, int active_lanes) {
    const int lane = threadIdx.x & 31;
    unsigned idx = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u % n;
    float acc = 0.f;
    const bool act = lane < active_lanes;
    for (int s = 0; s < steps; s++) {
        float4 a, b, c, d;
        if (MODE == 0) {            // 4 x LDG.128 per lane
            if (act) { a = __ldg(&t[idx].q[0]); b = __ldg(&t[idx].q[1]); c = __ldg(&t[idx].q[2]); d = __ldg(&t[idx].q[3]); }
        } else if (MODE == 1) {     // 2 x LDG.256 per lane
            if (act) { ld256(&t[idx].q[0], a, b); ld256(&t[idx].q[2], c, d); }
        } else if (MODE == 2) {     // cooperative: 4 lanes fetch one record, 4 rounds, shuffle to owner
            float4 mine[4];
#pragma unroll
            for (int r = 0; r < 4; r++) {
                int owner = 8 * r + (lane >> 2);
                unsigned oi = __shfl_sync(0xffffffffu, idx, owner);
                float4 v = make_float4(0, 0, 0, 0);
                if (owner < active_lanes) v = __ldg(&t[oi].q[lane & 3]);
                // owner (lane == 8r + k) collects quads j = 0..3 from lanes 4k + j
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    int src = 4 * (lane & 7) + j;
                    float4 w;
                    w.x = __shfl_sync(0xffffffffu, v.x, src); w.y = __shfl_sync(0xffffffffu, v.y, src);
                    w.z = __shfl_sync(0xffffffffu, v.z, src); w.w = __shfl_sync(0xffffffffu, v.w, src);
                    if ((lane >> 3) == r) mine[j] = w;
                }
            }
            a = mine[0]; b = mine[1]; c = mine[2]; d = mine[3];
        } else if (MODE == 3) {     // quad: 4 lanes own one ray, each loads 16 B of the record
            unsigned oi = __shfl_sync(0xffffffffu, idx, lane & ~3);
            a = __ldg(&t[oi].q[lane & 3]); b = a; c = a; d = a;
            float s1 = __shfl_xor_sync(0xffffffffu, a.w, 1), s2 = __shfl_xor_sync(0xffffffffu, a.x, 2);
            a.w += s1 * 0.f + s2 * 0.f;
        } else if (MODE == 4) {     // 1 x LDG.128 per lane (16 B only)
            if (act) { a = __ldg(&t[idx].q[0]); b = a; c = a; d = a; }
        } else if (MODE == 5) {     // 1 x LDG.256 per lane (32 B only)
            if (act) { ld256(&t[idx].q[0], a, b); c = a; d = b; }
        } else if (MODE == 6) {     // lane pairs: two LDG.256, each fetching ONE record of the pair by halves (one line per pair per
                                    // instruction instead of two), then 8 SHFL.BFLY to hand the halves over
            const unsigned even_idx = __shfl_sync(0xffffffffu, idx, lane & ~1), odd_idx = __shfl_sync(0xffffffffu, idx, lane | 1);
            const bool even_act = (lane & ~1) < active_lanes, odd_act = (lane | 1) < active_lanes;
            const int half = lane & 1;
            float4 e0 = make_float4(0, 0, 0, 0), e1 = e0, o0 = e0, o1 = e0;
            if (even_act) ld256(&t[even_idx].q[2 * half], e0, e1);       // even lane's record: lane 2k its first half, lane 2k+1 its second
            if (odd_act) ld256(&t[odd_idx].q[2 * half], o0, o1);         // odd lane's record likewise
            // the even lane keeps e (first half) and needs the partner's e (second half); the odd lane keeps o (second half) and needs the partner's o
            float4 give0 = half ? e0 : o0, give1 = half ? e1 : o1, got0, got1;
            got0.x = __shfl_xor_sync(0xffffffffu, give0.x, 1); got0.y = __shfl_xor_sync(0xffffffffu, give0.y, 1);
            got0.z = __shfl_xor_sync(0xffffffffu, give0.z, 1); got0.w = __shfl_xor_sync(0xffffffffu, give0.w, 1);
            got1.x = __shfl_xor_sync(0xffffffffu, give1.x, 1); got1.y = __shfl_xor_sync(0xffffffffu, give1.y, 1);
            got1.z = __shfl_xor_sync(0xffffffffu, give1.z, 1); got1.w = __shfl_xor_sync(0xffffffffu, give1.w, 1);
            if (half) { a = got0; b = got1; c = o0; d = o1; } else { a = e0; b = e1; c = got0; d = got1; }
        }
        if (act || MODE == 2 || MODE == 3 || MODE == 6) {
            acc += a.x + b.y + c.z + d.x;
            idx = (__float_as_uint(a.w) + (unsigned)s * 7919u + lane * 104729u) % n;   // dependent next index
        }
    }
    if (acc == 12345.678f) out[0] = acc;
}

int main(int argc, char** argv) {
    int n = argc > 1 ? atoi(argv[1]) : (1 << 17);     // 128 Ki records = 8 MiB (L2 resident)
    int steps = 2000;
    std::vector<Rec> h(n);
    srand(1);
    for (int i = 0; i < n; i++) for (int k = 0; k < 4; k++) {
        unsigned r = (unsigned)rand() * 2654435761u;
        h[i].q[k] = make_float4(1.f, 2.f, 3.f, 0.f); ((unsigned*)&h[i].q[k])[3] = r;
    }
    Rec* d; float* out;
    cudaMalloc(&d, sizeof(Rec) * n); cudaMalloc(&out, 4);
    cudaMemcpy(d, h.data(), sizeof(Rec) * n, cudaMemcpyHostToDevice);
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int blocks = p.multiProcessorCount * 10;
    const char* names[] = {"4xLDG.128/lane", "2xLDG.256/lane", "coop 4 lanes/rec + 16 SHFL", "quad (8 rays/warp) 1xLDG.128", "1xLDG.128 (16B)", "1xLDG.256 (32B)", "lane pairs: 2xLDG.256 by halves + 8 SHFL"};
    for (int lanes : {32, 20, 10}) {
        for (int mode = 0; mode < 7; mode++) {
            cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
            for (int rep = 0; rep < 2; rep++) {
                cudaEventRecord(e0);
                switch (mode) {
                    case 0: k<0><<<blocks, 128>>>(d, n, steps, out, lanes); break;
                    case 1: k<1><<<blocks, 128>>>(d, n, steps, out, lanes); break;
                    case 2: k<2><<<blocks, 128>>>(d, n, steps, out, lanes); break;
                    case 3: k<3><<<blocks, 128>>>(d, n, steps, out, lanes); break;
                    case 4: k<4><<<blocks, 128>>>(d, n, steps, out, lanes); break;
                    case 5: k<5><<<blocks, 128>>>(d, n, steps, out, lanes); break;
                    case 6: k<6><<<blocks, 128>>>(d, n, steps, out, lanes); break;
                }
                cudaEventRecord(e1); cudaEventSynchronize(e1);
            }
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            double rays_per_warp = mode == 3 ? 8 : lanes;
            double recs = (double)blocks * 4 * rays_per_warp * steps;
            double cyc = ms * 1e-3 * 1.965e9 * p.multiProcessorCount;
            printf("lanes %2d  %-30s %8.3f ms  %6.2f SM-cycles per record fetch  (%5.1f Grec/s)\n", lanes, names[mode], ms, cyc / recs, recs / ms / 1e6);
        }
    }
    printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
