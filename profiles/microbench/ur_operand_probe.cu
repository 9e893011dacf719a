// ur_operand_probe.cu -- measurement tool (not product code).  Does the dense loop get cheaper when
// the per-primitive coefficients are WARP-UNIFORM values held in uniform registers (UR operands of
// FFMA2) instead of vector registers?  The pixel tile of a warp is one row high here (v depends on
// blockIdx only), the bound records sit in __constant__ memory and are indexed by the loop counter,
// so everything but u is provably uniform.  Variant 0 = shipped form (table in shared memory,
// LDS.128, vector-register operands), for the same launch shape.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o ur_operand_probe ur_operand_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define NPRIM 1024
#define G 4
#define PAIRS 8

__constant__ float4 c_rec[2 * NPRIM];   // (sA, k1, k0, e), (cv, -, -, -)

__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
    uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r;
}
__device__ __forceinline__ void unpack2u(uint64_t v, uint32_t& lo, uint32_t& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d;
}

// V: 0 = shared-memory row table + LDS.128 (shipped form); 1 = constant-bank records, row terms computed
// from uniform values per primitive; 2 = constant-bank PRECOMPUTED (sA,rowW,rowC) (timing only: as if a
// per-row table could live in the constant bank)
template <int V>
__global__ void __launch_bounds__(128, 4) probe(const float4* __restrict__ recs, float* out, int iters, float x0) {
    __shared__ __align__(16) float4 tab[NPRIM];
    const float v = (float)blockIdx.x;
    if (V == 0) {
        for (int i = threadIdx.x; i < NPRIM; i += blockDim.x) {
            const float4 r0 = recs[2 * i];
            const float cv = recs[2 * i + 1].x;
            const float dv = v - cv;
            tab[i] = make_float4(r0.x, fmaf(r0.y, v, r0.z), (r0.w * dv) * dv, 0.f);
        }
        __syncthreads();
    }
    uint64_t u2[PAIRS];
#pragma unroll
    for (int j = 0; j < PAIRS; ++j) {
        const float u = x0 + threadIdx.x * (2 * PAIRS) + 2 * j;
        u2[j] = pack2(u, u + 1.0f);
    }
    uint32_t hits = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll 1
        for (int p = 0; p < NPRIM; p += G) {
            uint32_t acc[G];
#pragma unroll
            for (int g = 0; g < G; ++g) {
                float sA, rowW, rowC;
                if (V == 0) {
                    const float4 t = tab[p + g];
                    sA = t.x; rowW = t.y; rowC = t.z;
                } else if (V == 1) {
                    const float4 r0 = c_rec[2 * (p + g)];
                    const float cv = c_rec[2 * (p + g) + 1].x;
                    sA = r0.x;
                    rowW = fmaf(r0.y, v, r0.z);
                    const float dv = v - cv;
                    rowC = (r0.w * dv) * dv;
                } else {
                    const float4 t = c_rec[p + g];
                    sA = t.x; rowW = t.y; rowC = t.z;
                }
                const uint64_t S2 = pack2(sA, sA), W2 = pack2(rowW, rowW), C2 = pack2(rowC, rowC);
                uint32_t a = 0xFFFFFFFFu;
#pragma unroll
                for (int j = 0; j < PAIRS; ++j) {
                    const uint64_t w2 = ffma2(u2[j], S2, W2);
                    uint32_t q0, q1;
                    unpack2u(ffma2(w2, w2, C2), q0, q1);
                    a &= q0 & q1;
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
    const int blocks = sms * 4, iters = 4;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int w = 0; w < 2; ++w) probe<V><<<blocks, 128>>>(d_rec, d_out, iters, 100.0f);
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        probe<V><<<blocks, 128>>>(d_rec, d_out, iters, 100.0f);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double tp = 4.0 * iters * NPRIM;  // warp-level thread-primitives per scheduler (4 CTAs x 4 warps / 4)
    std::printf("{\"variant\": %d, \"what\": \"%s\", \"ms\": %.4f, \"cycles_per_thread_prim\": %.2f, \"err\": \"%s\"}\n", V, what, best,
                best * 1e-3 * clk_ghz * 1e9 / tp, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const double clk = clk_khz * 1e-6;
    std::printf("{\"device\": \"%s\", \"sms\": %d, \"clock_ghz\": %.3f}\n", pr.name, pr.multiProcessorCount, clk);
    float4* h = new float4[2 * NPRIM];
    for (int i = 0; i < NPRIM; ++i) {
        h[2 * i] = make_float4(1e-3f + 1e-6f * i, 1e-5f, -9.0f - i, 1e-3f);
        h[2 * i + 1] = make_float4(-7000.0f, 0.f, 0.f, 0.f);
    }
    float4* d_rec; cudaMalloc(&d_rec, sizeof(float4) * 2 * NPRIM);
    cudaMemcpy(d_rec, h, sizeof(float4) * 2 * NPRIM, cudaMemcpyHostToDevice);
    cudaMemcpyToSymbol(c_rec, h, sizeof(float4) * 2 * NPRIM);
    float* d_out; cudaMalloc(&d_out, sizeof(float) * pr.multiProcessorCount * 4 * 128);
    run<0>("shared-memory row table, LDS.128, vector-register coefficients (shipped form)", d_rec, d_out, pr.multiProcessorCount, clk);
    run<1>("constant-bank records, uniform row terms, uniform-register coefficients", d_rec, d_out, pr.multiProcessorCount, clk);
    run<2>("constant-bank precomputed coefficients (timing only)", d_rec, d_out, pr.multiProcessorCount, clk);
    return 0;
}
