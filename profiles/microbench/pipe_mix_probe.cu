// pipe_mix_probe.cu -- measurement tool (not product code).  Time per loop iteration of
// parameterised instruction mixes on sm_100a: NF scalar FFMA, NF2 packed FFMA2 (fma.rn.f32x2),
// NA FMNMX, N3 FMNMX3, NS FSETP(+predicate OR), NL warp-uniform LDS.128 per iteration, all on
// independent registers.  32 warps per SM (8 per scheduler), 148*4 blocks.  Output: JSON lines with
// ns per warp-iteration per scheduler; tools/loop_sass.py lists the SASS opcodes of each loop.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 2048

__device__ __forceinline__ uint64_t pack2(float a, float b) {
    uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r;
}
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d;
}
__device__ __forceinline__ float ffma(float a, float b, float c) {
    float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d;
}
__device__ __forceinline__ float fmax_(float a, float b) {
    float d; asm volatile("max.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b)); return d;
}
__device__ __forceinline__ float fmax3_(float a, float b, float c) {
    float d; asm volatile("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d;
}

template <int NF, int NF2, int NA, int N3, int NS, int NL>
__global__ void __launch_bounds__(256) mix(float* out, float s0, float s1, int iters) {
    __shared__ float4 sh[256];
    sh[threadIdx.x] = make_float4(s0, s1, s0, s1);
    __syncthreads();
    float x[16], y[8];
    uint64_t z[8];
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = s0 + 0.001f * (threadIdx.x + i);
#pragma unroll
    for (int i = 0; i < 8; ++i) { y[i] = s1 * i; z[i] = pack2(x[i], x[i + 8]); }
    const uint64_t k0 = pack2(s0, s0), k1 = pack2(s1, s1);
    int pred_count = 0;
    float4 l = make_float4(0, 0, 0, 0);
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
        bool any = false;
#pragma unroll
        for (int j = 0; j < NL; ++j) {
            const float4 v = sh[(it + j) & 255];  // warp-uniform address -> broadcast LDS.128
            l.x += v.x; l.y += v.y; l.z += v.z; l.w += v.w;  // (4 FADD per LDS; counted by the caller)
        }
#pragma unroll
        for (int j = 0; j < NF; ++j) x[j & 15] = ffma(x[j & 15], s0, s1);
#pragma unroll
        for (int j = 0; j < NF2; ++j) z[j & 7] = ffma2(z[j & 7], k0, k1);
#pragma unroll
        for (int j = 0; j < NA; ++j) y[j & 7] = fmax_(y[j & 7], x[(j + 3) & 15]);
#pragma unroll
        for (int j = 0; j < N3; ++j) y[j & 7] = fmax3_(y[j & 7], x[(j + 3) & 15], x[(j + 9) & 15]);
#pragma unroll
        for (int j = 0; j < NS; ++j) any |= (x[j & 15] <= y[j & 7]);
        if (NS > 0 && any) pred_count++;
    }
    float acc = l.x + l.y + l.z + l.w + pred_count;
#pragma unroll
    for (int i = 0; i < 16; ++i) acc += x[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float a, b; asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(z[i]));
        acc += y[i] + a + b;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int NF, int NF2, int NA, int N3, int NS, int NL>
void run(int sms) {
    const int blocks = sms * 4, threads = 256;
    float* out; cudaMalloc(&out, sizeof(float) * blocks * threads);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int w = 0; w < 3; ++w) mix<NF, NF2, NA, N3, NS, NL><<<blocks, threads>>>(out, 0.999f, 0.5f, ITERS);
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        mix<NF, NF2, NA, N3, NS, NL><<<blocks, threads>>>(out, 0.999f, 0.5f, ITERS);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    // 8 warps per scheduler run concurrently: ns per warp-iteration per scheduler
    const double ns = best * 1e6 / ((double)ITERS * 8.0);
    std::printf("{\"ffma\": %d, \"ffma2\": %d, \"fmnmx\": %d, \"fmnmx3\": %d, \"fsetp\": %d, \"lds128\": %d, "
                "\"ms\": %.4f, \"ns_per_warp_iter\": %.3f, \"err\": \"%s\"}\n",
                NF, NF2, NA, N3, NS, NL, best, ns, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    std::printf("{\"device\": \"%s\", \"sms\": %d}\n", p.name, sms);
    //   FFMA FFMA2 FMNMX FMNMX3 FSETP LDS
    run<64, 0, 0, 0, 0, 0>(sms);   // calibration: issue-bound FFMA
    run<128, 0, 0, 0, 0, 0>(sms);
    run<0, 32, 0, 0, 0, 0>(sms);   // 64 FMAs in 32 issue slots
    run<0, 64, 0, 0, 0, 0>(sms);
    run<0, 0, 64, 0, 0, 0>(sms);   // ALU pipe alone
    run<0, 0, 0, 64, 0, 0>(sms);
    run<16, 0, 0, 0, 64, 0>(sms);  // FSETP rate (16 FFMA keep x moving)
    run<64, 0, 16, 0, 0, 0>(sms);  // FFMA : FMNMX = 4 : 1
    run<64, 0, 32, 0, 0, 0>(sms);  // 2 : 1
    run<64, 0, 64, 0, 0, 0>(sms);  // 1 : 1
    run<0, 32, 16, 0, 0, 0>(sms);  // FFMA2-packed equivalents
    run<0, 32, 32, 0, 0, 0>(sms);
    run<0, 32, 64, 0, 0, 0>(sms);
    run<48, 0, 0, 16, 8, 0>(sms);  // scalar cube-test mix for 8 px: 48 FMA-pipe, 16 FMNMX3, 8 FSETP
    run<0, 24, 0, 16, 8, 0>(sms);  // packed
    run<24, 0, 0, 8, 8, 0>(sms);   // footprint-only mix for 8 px: 24 FFMA, 8 FMNMX3, 8 FSETP
    run<0, 12, 0, 8, 8, 0>(sms);
    run<24, 0, 0, 8, 8, 3>(sms);   // + 3 LDS.128 (one cube record) (+12 FADD)
    run<0, 12, 0, 8, 8, 3>(sms);
    run<16, 0, 0, 0, 8, 1>(sms);   // sphere mix 8 px: 16 FMA-pipe + 8 FSETP + 1 LDS
    run<0, 8, 0, 0, 8, 1>(sms);
    run<64, 0, 0, 0, 0, 4>(sms);   // LDS cost on top of FFMA
    run<64, 0, 0, 0, 0, 16>(sms);
    return 0;
}
