// fp32_issue_probe.cu -- measurement tool (not product code): what the sm_100a SM sustains per
// clock for the instruction mixes the ray-cast kernel is built from.  Prints one JSON line per mix:
// lane-ops per clock per SM (128 = one FP32 op per CUDA core per clock).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp32_issue_probe fp32_issue_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 4096

__device__ __forceinline__ void ffma2(float2& d, const float2 a, const float2 b, const float2 c) {
    uint64_t ra, rb, rc, rd;
    asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
}
__device__ __forceinline__ void fadd2(float2& d, const float2 a, const float2 b) {
    uint64_t ra, rb, rd;
    asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
}

// mix 0: 8 independent FFMA chains
// mix 1: 8 independent FFMA2 chains (16 FMAs per "step")
// mix 2: 6 FFMA + 2 FMNMX3 + 1 FSETP  (scalar cube test shape)
// mix 3: 3 FFMA2 + 2 FMNMX3 + 1 FSETP  (cube test, 2 pixels packed on near/far)
// mix 4: 8 FFMA + 4 FMNMX (can ALU-pipe ops hide under FMA issue?)
// mix 5: 4 FFMA2 + 4 FMNMX
// mix 6: 4 FFMA2 + 8 FMNMX
// mix 7: 8 FADD chains
template <int MIX>
__global__ void __launch_bounds__(256) probe(float* out, long long* cyc, float s0, float s1) {
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = s0 + 0.001f * (threadIdx.x + i);
    float m = s1, acc = 0.f;
    bool any = false;
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
        if (MIX == 0) {
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], s0, s1);
        } else if (MIX == 1) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float2 d;
                ffma2(d, make_float2(a[2 * i], a[2 * i + 1]), make_float2(s0, s0), make_float2(s1, s1));
                a[2 * i] = d.x; a[2 * i + 1] = d.y;
            }
        } else if (MIX == 2) {
            // two pixels of the scalar cube test per iteration: 12 FFMA-pipe + 4 FMNMX3 + 2 FSETP
#pragma unroll
            for (int p = 0; p < 2; ++p) {
                const float u = a[p];
                const float n0 = fmaf(s0, u, a[2]), n1 = fmaf(s1, u, a[3]), n2 = fmaf(s0, u, a[4]);
                const float f0 = n0 + a[5], f1 = n1 + a[6], f2 = n2 + a[7];
                const float tn = fmaxf(fmaxf(n0, n1), n2), tf = fminf(fminf(f0, f1), f2);
                any |= (tn <= tf);
                a[p] = u + 1.0f;  // 1 more FADD to keep the chain live
            }
        } else if (MIX == 3) {
            const float2 u = make_float2(a[0], a[1]);
            float2 n0, n1, n2, f0, f1, f2;
            ffma2(n0, make_float2(s0, s0), u, make_float2(a[2], a[2]));
            ffma2(n1, make_float2(s1, s1), u, make_float2(a[3], a[3]));
            ffma2(n2, make_float2(s0, s0), u, make_float2(a[4], a[4]));
            fadd2(f0, n0, make_float2(a[5], a[5]));
            fadd2(f1, n1, make_float2(a[6], a[6]));
            fadd2(f2, n2, make_float2(a[7], a[7]));
            const float tnx = fmaxf(fmaxf(n0.x, n1.x), n2.x), tfx = fminf(fminf(f0.x, f1.x), f2.x);
            const float tny = fmaxf(fmaxf(n0.y, n1.y), n2.y), tfy = fminf(fminf(f0.y, f1.y), f2.y);
            any |= (tnx <= tfx);
            any |= (tny <= tfy);
            float2 un;
            fadd2(un, u, make_float2(1.0f, 1.0f));
            a[0] = un.x; a[1] = un.y;
        } else if (MIX == 4) {
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], s0, s1);
#pragma unroll
            for (int i = 0; i < 4; ++i) a[8 + i] = fmaxf(a[8 + i], a[i]);
        } else if (MIX == 5 || MIX == 6) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float2 d;
                ffma2(d, make_float2(a[2 * i], a[2 * i + 1]), make_float2(s0, s0), make_float2(s1, s1));
                a[2 * i] = d.x; a[2 * i + 1] = d.y;
            }
#pragma unroll
            for (int i = 0; i < (MIX == 5 ? 4 : 8); ++i) a[8 + i] = fmaxf(a[8 + i], a[i]);
        } else if (MIX == 7) {
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = a[i] + s1;
        }
    }
    const long long t1 = clock64();
#pragma unroll
    for (int i = 0; i < 16; ++i) acc += a[i];
    if (any) acc += m;
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MIX>
void run(const char* name, double fma_lane_ops_per_iter, double alu_lane_ops_per_iter, int sms) {
    const int blocks_per_sm = 4, threads = 256;
    const int blocks = sms * blocks_per_sm;
    float* out; long long* cyc;
    cudaMalloc(&out, sizeof(float) * blocks * threads);
    cudaMalloc(&cyc, sizeof(long long) * blocks);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<MIX><<<blocks, threads>>>(out, cyc, 0.999f, 0.5f);
    cudaEventRecord(e0);
    probe<MIX><<<blocks, threads>>>(out, cyc, 0.999f, 0.5f);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long* h = new long long[blocks];
    cudaMemcpy(h, cyc, sizeof(long long) * blocks, cudaMemcpyDeviceToHost);
    double mean = 0; for (int i = 0; i < blocks; ++i) mean += (double)h[i]; mean /= blocks;
    const double thr_per_sm = (double)blocks_per_sm * threads;
    const double fma_rate = fma_lane_ops_per_iter * ITERS * thr_per_sm / mean;
    const double alu_rate = alu_lane_ops_per_iter * ITERS * thr_per_sm / mean;
    const double warp_instr_per_clk_smsp = 0;  (void)warp_instr_per_clk_smsp;
    std::printf("{\"mix\": \"%s\", \"fma_pipe_lane_ops_per_clk_per_sm\": %.1f, \"alu_pipe_lane_ops_per_clk_per_sm\": %.1f, "
                "\"cycles\": %.0f, \"ms\": %.4f, \"eff_clock_mhz\": %.0f, \"cuda_error\": \"%s\"}\n",
                name, fma_rate, alu_rate, mean, ms, mean / (ms * 1e3), cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(cyc); delete[] h;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    std::printf("{\"device\": \"%s\", \"sms\": %d, \"cc\": \"%d.%d\"}\n", p.name, p.multiProcessorCount, p.major, p.minor);
    const int sms = p.multiProcessorCount;
    // lane-op counts per thread per iteration: (FMA-pipe ops counting an f32x2 op as 2), (ALU-pipe ops)
    run<0>("8xFFMA", 8, 0, sms);
    run<1>("8xFFMA2", 16, 0, sms);
    run<7>("8xFADD", 8, 0, sms);
    run<2>("scalar cube test x2px: 14 FMA-pipe + 4 FMNMX3 + 2 FSETP", 14, 6, sms);
    run<3>("packed cube test x2px: 7 f32x2 (=14) + 4 FMNMX3 + 2 FSETP", 14, 6, sms);
    run<4>("8xFFMA + 4xFMNMX", 8, 4, sms);
    run<5>("4xFFMA2 + 4xFMNMX", 8, 4, sms);
    run<6>("4xFFMA2 + 8xFMNMX", 8, 8, sms);
    return 0;
}
