// pair_rate_probe.cu -- measurement tool (not product code): what one sm_100a scheduler sustains on the two
// instructions the dense ray-cast loop is made of -- FFMA2 (w' = sAn*u + rWn for a ray pair) and the 3-input FMNMX
// with |.| on two operands (running minimum of |w'|) -- alone, paired as in the kernel, and with the loop's other
// instructions (LDS.128, integer address arithmetic) mixed in.  Occupancy as in raycast_dense_kernel: 4 CTAs of 4
// warps per SM = 4 warps per scheduler.  One JSON line per mix: scheduler clocks per warp instruction, and for the
// paired mixes clocks per (FFMA2 + FMNMX3) pair = the per-ray-pair, per-primitive cost that bounds the kernel.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pair_rate_probe pair_rate_probe.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 2048

__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

// MIX 0: 8 FFMA2                    1: 8 FMNMX3(|.|,|.|)            2: 8 x (FFMA2 -> FMNMX3), the dense loop body
//     3: 8 FMNMX (2 inputs)         4: 8 x (FFMA2 -> 2 FMNMX)       5: mix 2 + 1 LDS.128 + 3 integer ops per 8 pairs
//     6: 8 x (FFMA2 -> LOP3)        7: 16 FFMA (scalar)
template <int MIX>
__global__ void __launch_bounds__(128, 4) probe(float* out, long long* cyc, float s0, float s1, const float4* tab) {
    __shared__ float4 sm[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) sm[i] = tab[i];
    __syncthreads();
    uint64_t u[8];
    float m[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        u[i] = pack2(s0 + 0.001f * (threadIdx.x + i), s0 - 0.002f * (threadIdx.x + i));
        m[i] = 1e30f;
    }
    uint64_t S2 = pack2(s0, s0), W2 = pack2(s1, s1);
    unsigned acc_i = threadIdx.x;
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
        if (MIX == 5) {
            const float4 e = sm[(acc_i + it) & 255];
            S2 = pack2(e.x, e.x);
            W2 = pack2(e.y, e.y);
            acc_i = acc_i * 5u + 1u;
            acc_i ^= (unsigned)it;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float lo, hi;
            if (MIX == 0) {
                u[i] = ffma2(u[i], S2, W2);
            } else if (MIX == 1) {
                unpack2(u[i], lo, hi);
                m[i] = fminf(fminf(m[i], fabsf(lo)), fabsf(hi));
                u[i] = pack2(m[i], hi);  // keeps the operands changing without another instruction: a register rename
            } else if (MIX == 2 || MIX == 5) {
                const uint64_t w = ffma2(u[i], S2, W2);
                unpack2(w, lo, hi);
                m[i] = fminf(fminf(m[i], fabsf(lo)), fabsf(hi));
            } else if (MIX == 3) {
                unpack2(u[i], lo, hi);
                m[i] = fminf(m[i], lo);
                u[i] = pack2(m[i], hi);
            } else if (MIX == 4) {
                const uint64_t w = ffma2(u[i], S2, W2);
                unpack2(w, lo, hi);
                m[i] = fminf(m[i], fabsf(lo));
                m[(i + 1) & 7] = fminf(m[(i + 1) & 7], fabsf(hi));
            } else if (MIX == 6) {
                const uint64_t w = ffma2(u[i], S2, W2);
                unpack2(w, lo, hi);
                acc_i = acc_i & (__float_as_uint(lo) | __float_as_uint(hi));
            } else if (MIX == 7) {
                unpack2(u[i], lo, hi);
                lo = fmaf(lo, s0, s1);
                hi = fmaf(hi, s0, s1);
                u[i] = pack2(lo, hi);
            }
        }
        if (MIX == 2 || MIX == 4 || MIX == 5 || MIX == 6) {  // feed the results back so that nothing is loop-invariant
            W2 = pack2(s1 + m[0] * 1e-30f, s1 + (float)(acc_i & 1u) * 1e-30f);
        }
    }
    const long long t1 = clock64();
    float acc = (float)acc_i;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float lo, hi;
        unpack2(u[i], lo, hi);
        acc += m[i] + lo + hi;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MIX>
void run(const char* name, int two_ray_instr, int other_instr, int pairs, int sms, const float4* tab) {
    const int blocks = sms * 4, threads = 128;
    float* out;
    long long* cyc;
    cudaMalloc(&out, sizeof(float) * blocks * threads);
    cudaMalloc(&cyc, sizeof(long long) * blocks);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    probe<MIX><<<blocks, threads>>>(out, cyc, 0.999f, 0.5f, tab);
    cudaEventRecord(e0);
    probe<MIX><<<blocks, threads>>>(out, cyc, 0.999f, 0.5f, tab);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    long long* h = new long long[blocks];
    cudaMemcpy(h, cyc, sizeof(long long) * blocks, cudaMemcpyDeviceToHost);
    double mean = 0;
    for (int i = 0; i < blocks; ++i) mean += (double)h[i];
    mean /= blocks;
    // 4 warps per scheduler, each runs ITERS iterations: scheduler clocks per warp iteration
    const double clk_per_iter = mean / ITERS / 4.0;
    std::printf("{\"mix\": \"%s\", \"instr_per_iter\": %d, \"clocks_per_warp_iter_per_scheduler\": %.2f, \"clocks_per_instr\": %.3f",
                name, two_ray_instr + other_instr, clk_per_iter, clk_per_iter / (two_ray_instr + other_instr));
    if (pairs) std::printf(", \"clocks_per_ffma2_fmnmx3_pair\": %.3f", clk_per_iter / pairs);
    std::printf(", \"cycles\": %.0f, \"ms\": %.4f, \"eff_clock_mhz\": %.0f, \"cuda_error\": \"%s\"}\n", mean, ms, mean / (ms * 1e3),
                cudaGetErrorString(cudaGetLastError()));
    cudaFree(out);
    cudaFree(cyc);
    delete[] h;
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    std::printf("{\"device\": \"%s\", \"sms\": %d, \"cc\": \"%d.%d\", \"occupancy\": \"4 CTAs x 4 warps per SM\"}\n", p.name,
                p.multiProcessorCount, p.major, p.minor);
    const int sms = p.multiProcessorCount;
    float4* tab;
    cudaMalloc(&tab, sizeof(float4) * 256);
    float4 h[256];
    for (int i = 0; i < 256; ++i) h[i] = make_float4(0.999f + i * 1e-6f, 0.5f, 0.f, 0.f);
    cudaMemcpy(tab, h, sizeof h, cudaMemcpyHostToDevice);
    run<0>("8xFFMA2", 8, 0, 0, sms, tab);
    run<7>("16xFFMA", 16, 0, 0, sms, tab);
    run<1>("8xFMNMX3(|a|,|b|)", 8, 0, 0, sms, tab);
    run<3>("8xFMNMX", 8, 0, 0, sms, tab);
    run<2>("8x(FFMA2 -> FMNMX3): the dense loop body", 16, 0, 8, sms, tab);
    run<5>("8x(FFMA2 -> FMNMX3) + LDS.128 + integer ops", 16, 6, 8, sms, tab);
    run<4>("8x(FFMA2 -> 2 FMNMX)", 24, 0, 0, sms, tab);
    run<6>("8x(FFMA2 -> LOP3)", 16, 0, 0, sms, tab);
    cudaFree(tab);
    return 0;
}
