// probe.cuh -- micro-kernels that measure the two rates the MD kernel's roofline is stated against:
// non-tensor FP32 FMA throughput and shared-memory read bandwidth (SURVEY.md section 8(d) asks for both
// to be confirmed by a micro-benchmark; MEASURED_PEAKS.json holds only HBM copy and tensor numbers).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace igmk {

constexpr int PROBE_THREADS = 1024;
constexpr int PROBE_CHAINS = 8;          // independent accumulators per thread (hides the 4-cycle FMA latency at 8 warps/scheduler)

// iters x PROBE_CHAINS dependent-per-chain FFMA per thread; 2 flops each
__global__ void __launch_bounds__(PROBE_THREADS, 1) k_probe_ffma(int iters, float a, float b, float* out)
{
    float v[PROBE_CHAINS];
#pragma unroll
    for (int k = 0; k < PROBE_CHAINS; k++) v[k] = (float)(threadIdx.x + k);
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < PROBE_CHAINS; k++) v[k] = fmaf(v[k], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < PROBE_CHAINS; k++) s += v[k];
    if (s == 12345.678f) out[0] = s;      // never true: keeps the loop alive
}

// the same work issued as packed fma.rn.f32x2 (two lanes of arithmetic per instruction, sm_100)
__global__ void __launch_bounds__(PROBE_THREADS, 1) k_probe_ffma2(int iters, float a, float b, float* out)
{
    unsigned long long v[PROBE_CHAINS / 2], pa, pb;
    asm("mov.b64 %0, {%1, %1};" : "=l"(pa) : "f"(a));
    asm("mov.b64 %0, {%1, %1};" : "=l"(pb) : "f"(b));
#pragma unroll
    for (int k = 0; k < PROBE_CHAINS / 2; k++) {
        const float lo = (float)(threadIdx.x + 2 * k), hi = (float)(threadIdx.x + 2 * k + 1);
        asm("mov.b64 %0, {%1, %2};" : "=l"(v[k]) : "f"(lo), "f"(hi));
    }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < PROBE_CHAINS / 2; k++) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v[k]) : "l"(pa), "l"(pb));
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < PROBE_CHAINS / 2; k++) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v[k]));
        s += lo + hi;
    }
    if (s == 12345.678f) out[0] = s;
}

// iters x 4 conflict-free 16-byte shared-memory reads per thread (a warp covers 512 contiguous bytes)
__global__ void __launch_bounds__(PROBE_THREADS, 1) k_probe_lds(int iters, float* out)
{
    extern __shared__ __align__(16) unsigned char probe_smem[];
    float4* buf = (float4*)probe_smem;                   // 4 * PROBE_THREADS float4 = 64 KB
    for (int k = 0; k < 4; k++) buf[k * PROBE_THREADS + threadIdx.x] = make_float4(1.f, 2.f, 3.f, (float)threadIdx.x);
    __syncthreads();
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    unsigned idx = threadIdx.x;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const float4 q = buf[k * PROBE_THREADS + idx];
            acc.x += q.x; acc.y += q.y; acc.z += q.z; acc.w += q.w;
        }
        idx = (idx + 32) & (PROBE_THREADS - 1);          // another warp's segment next time: nothing stays in a register
    }
    if (acc.x + acc.y + acc.z + acc.w == 12345.678f) out[0] = acc.x;
}

}  // namespace igmk
