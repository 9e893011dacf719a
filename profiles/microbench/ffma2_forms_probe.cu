// ffma2_forms_probe.cu -- measurement tool (not product code).  The grid kernel's dense loop
// (raycast_kernels.cu, bound_group) rebuilt with one ingredient changed at a time, to find what
// each SASS form costs on sm_100a.  16 rays per thread (8 packed pairs), 4 primitives per group,
// coefficient records read as warp-uniform LDS.128/LDS.64, 3 CTAs x 256 threads per SM.
// Prints cycles per (thread, primitive) at the max SM clock for every variant.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o ffma2_forms_probe ffma2_forms_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define NPRIM 512
#define PAIRS 8
#define G 4

__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
    uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r;
}
__device__ __forceinline__ void unpack2u(uint64_t v, uint32_t& lo, uint32_t& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d;
}
__device__ __forceinline__ uint64_t ffma2v(uint64_t a, uint64_t b, uint64_t c) {  // never eliminated
    uint64_t d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d;
}
__device__ __forceinline__ uint64_t fmul2(uint64_t a, uint64_t b) {
    uint64_t d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d;
}

// V0 real loop: t = u*A + B; q = u*t + C; acc &= q.lo & q.hi
// V1 stage 1 only (pair, scalar, scalar) + LOP3
// V2 stage 2 only (pair, pair, scalar) + LOP3
// V3 both stages, volatile, no LOP3 at all
// V4 scalar FFMA 2-stage + LOP3 (the previous kernel)
// V5 stage 2 written t*u + C
// V6 stage 2 without addend: q = u*t (FMUL2 pair,pair), C folded into the bit test (wrong maths, timing only)
// V7 both stages + LOP3, but 2 rows of LOP3 replaced by one per 4 values (half the LOP3s; wrong maths, timing only)
// V8 stage 1 only, no LOP3 (volatile)
// V9 stage 2 only, no LOP3 (volatile)
template <int V>
__global__ void __launch_bounds__(256, 3) probe(const float4* __restrict__ recs, float* out, int iters, float x0) {
    __shared__ __align__(16) float4 sh[2 * NPRIM];
    for (int i = threadIdx.x; i < 2 * NPRIM; i += blockDim.x) sh[i] = recs[i];
    __syncthreads();
    uint64_t u2[PAIRS];
    float us[2 * PAIRS];
#pragma unroll
    for (int j = 0; j < PAIRS; ++j) {
        us[2 * j] = x0 + (threadIdx.x & 7) * 16 + 2 * j;
        us[2 * j + 1] = us[2 * j] + 1.0f;
        u2[j] = pack2(us[2 * j], us[2 * j + 1]);
    }
    const float v = (float)(threadIdx.x >> 3) + blockIdx.x;
    uint32_t hits = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll 1
        for (int p = 0; p < NPRIM; p += G) {
            uint32_t acc[G];
#pragma unroll
            for (int g = 0; g < G; ++g) {
                const float4 r0 = sh[2 * (p + g)];
                const float2 r1 = *reinterpret_cast<const float2*>(&sh[2 * (p + g) + 1]);
                const float rowB = fmaf(r0.y, v, r0.z), rowC = fmaf(fmaf(r0.w, v, r1.x), v, r1.y);
                const uint64_t A2 = pack2(r0.x, r0.x), B2 = pack2(rowB, rowB), C2 = pack2(rowC, rowC);
                uint32_t a = 0xFFFFFFFFu;
#pragma unroll
                for (int j = 0; j < PAIRS; ++j) {
                    uint32_t q0, q1;
                    if (V == 0) {
                        unpack2u(ffma2(u2[j], ffma2(u2[j], A2, B2), C2), q0, q1);
                        a &= q0 & q1;
                    } else if (V == 1) {
                        unpack2u(ffma2(u2[j], A2, B2), q0, q1);
                        a &= q0 & q1;
                    } else if (V == 2) {
                        unpack2u(ffma2(u2[j], u2[(j + 3) & 7], C2), q0, q1);
                        a &= q0 & q1;
                    } else if (V == 3) {
                        (void)ffma2v(u2[j], ffma2v(u2[j], A2, B2), C2);
                    } else if (V == 4) {
                        const float t0 = fmaf(fmaf(r0.x, us[2 * j], rowB), us[2 * j], rowC);
                        const float t1 = fmaf(fmaf(r0.x, us[2 * j + 1], rowB), us[2 * j + 1], rowC);
                        a &= __float_as_uint(t0) & __float_as_uint(t1);
                    } else if (V == 5) {
                        unpack2u(ffma2(ffma2(u2[j], A2, B2), u2[j], C2), q0, q1);
                        a &= q0 & q1;
                    } else if (V == 6) {
                        unpack2u(fmul2(u2[j], ffma2(u2[j], A2, B2)), q0, q1);
                        a &= q0 & q1;
                    } else if (V == 7) {
                        unpack2u(ffma2(u2[j], ffma2(u2[j], A2, B2), C2), q0, q1);
                        if (j & 1) a &= q0 & q1; else a &= q1;
                    } else if (V == 8) {
                        (void)ffma2v(u2[j], A2, B2);
                    } else if (V == 9) {
                        (void)ffma2v(u2[j], u2[(j + 3) & 7], C2);
                    }
                }
                acc[g] = a;
            }
            uint32_t all = acc[0];
#pragma unroll
            for (int g = 1; g < G; ++g) all &= acc[g];
            if (__any_sync(0xFFFFFFFFu, (all & 0x40000000u) == 0u)) hits += 1;
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = (float)hits;
}

template <int V>
void run(const char* what, const float4* d_rec, float* d_out, int sms, double clk_ghz) {
    const int blocks = sms * 3, iters = 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int w = 0; w < 2; ++w) probe<V><<<blocks, 256>>>(d_rec, d_out, iters, 100.0f);
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        probe<V><<<blocks, 256>>>(d_rec, d_out, iters, 100.0f);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    // per SMSP: 3 CTAs * 8 warps / 4 = 6 warps, each doing iters*NPRIM thread-primitives
    const double cyc = best * 1e-3 * clk_ghz * 1e9 / (6.0 * iters * NPRIM);
    std::printf("{\"variant\": %d, \"what\": \"%s\", \"ms\": %.4f, \"cycles_per_thread_prim\": %.2f, "
                "\"cycles_per_test\": %.3f, \"err\": \"%s\"}\n",
                V, what, best, cyc, cyc / 16.0, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const double clk = clk_khz * 1e-6;
    std::printf("{\"device\": \"%s\", \"sms\": %d, \"clock_ghz\": %.3f}\n", pr.name, pr.multiProcessorCount, clk);
    float4* h = new float4[2 * NPRIM];
    for (int i = 0; i < NPRIM; ++i) {  // far-away ellipses: q is large, the vote never fires
        h[2 * i] = make_float4(1e-3f + 1e-6f * i, 1e-5f, -9.0f - i, 1e-3f);
        h[2 * i + 1] = make_float4(-7.0f, 1e6f + i, 0.f, 0.f);
    }
    float4* d_rec; cudaMalloc(&d_rec, sizeof(float4) * 2 * NPRIM);
    cudaMemcpy(d_rec, h, sizeof(float4) * 2 * NPRIM, cudaMemcpyHostToDevice);
    float* d_out; cudaMalloc(&d_out, sizeof(float) * pr.multiProcessorCount * 3 * 256);
    const int sms = pr.multiProcessorCount;
    run<0>("real: 2 FFMA2 + LOP3 per pair", d_rec, d_out, sms, clk);
    run<1>("stage 1 only (pair,scalar,scalar) + LOP3", d_rec, d_out, sms, clk);
    run<2>("stage 2 only (pair,pair,scalar) + LOP3", d_rec, d_out, sms, clk);
    run<3>("both stages, no LOP3", d_rec, d_out, sms, clk);
    run<4>("scalar FFMA 2-stage + LOP3", d_rec, d_out, sms, clk);
    run<5>("stage 2 as t*u + C", d_rec, d_out, sms, clk);
    run<6>("stage 2 as FMUL2 (pair,pair)", d_rec, d_out, sms, clk);
    run<7>("both stages, LOP3 with 1.5 inputs avg", d_rec, d_out, sms, clk);
    run<8>("stage 1 only, no LOP3", d_rec, d_out, sms, clk);
    run<9>("stage 2 only, no LOP3", d_rec, d_out, sms, clk);
    return 0;
}
