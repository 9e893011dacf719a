// half_rate_probe.cu -- measurement tool (not product code): issue cost on one sm_100a scheduler of the packed
// half-precision forms of the dense loop's two instructions -- HFMA2 (w' for a ray pair in one 32-bit register) and
// VHMNMX (3-input packed minimum with |.| on two operands: four rays per instruction) -- alone and paired, beside the
// FP32 pair (FFMA2 + FMNMX3) of pair_rate_probe.cu.  Occupancy as in raycast_dense_kernel (4 CTAs x 4 warps per SM).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o half_rate_probe half_rate_probe.cu
#include <cstdint>
#include <cstdio>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define ITERS 2048

// MIX 0: 16 HFMA2            1: 8 VHMNMX(m, |a|, |b|)        2: 8 x (2 HFMA2 -> VHMNMX): 32 rays
//     3: 16 HFMA2.BF16       4: 8 x (2 HFMA2.BF16 -> VHMNMX.BF16)   5: mix 2 + per-"primitive" FFMA + F2FP pack + LDS.64
//     6: 16 x HMNMX2 (2-input)
template <int MIX>
__global__ void __launch_bounds__(128, 4) probe(float* out, long long* cyc, float s0, float s1, const float4* tab) {
    __shared__ float4 sm[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) sm[i] = tab[i];
    __syncthreads();
    __half2 u[16], m[8];
    __nv_bfloat162 ub[16], mb[8];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        u[i] = __floats2half2_rn(s0 + 0.001f * (threadIdx.x + i), s0 - 0.002f * (threadIdx.x + i));
        ub[i] = __floats2bfloat162_rn(s0 + 0.001f * (threadIdx.x + i), s0 - 0.002f * (threadIdx.x + i));
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        m[i] = __floats2half2_rn(60000.f, 60000.f);
        mb[i] = __floats2bfloat162_rn(1e30f, 1e30f);
    }
    __half2 S2 = __floats2half2_rn(s0, s0), W2 = __floats2half2_rn(s1, s1);
    __nv_bfloat162 S2b = __floats2bfloat162_rn(s0, s0), W2b = __floats2bfloat162_rn(s1, s1);
    unsigned acc_i = threadIdx.x;
    float uc = s0 * threadIdx.x;
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
        if (MIX == 5) {
            const float2 e = reinterpret_cast<const float2*>(sm)[(acc_i + it) & 511];
            const float c = fmaf(e.x, uc, e.y);                 // c_t = sAn * u_c + rWn
            W2 = __floats2half2_rn(c, c);                         // F2FP pack
            S2 = *reinterpret_cast<const __half2*>(&e.x);        // sAn kept as half2 in the table
            acc_i = acc_i * 5u + 1u;
            acc_i ^= (unsigned)it;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MIX == 0) {
                u[2 * i] = __hfma2(u[2 * i], S2, W2);
                u[2 * i + 1] = __hfma2(u[2 * i + 1], S2, W2);
            } else if (MIX == 1) {
                m[i] = __hmin2(__hmin2(m[i], __habs2(u[2 * i])), __habs2(u[2 * i + 1]));
                u[2 * i] = m[i];
            } else if (MIX == 2 || MIX == 5) {
                const __half2 a = __hfma2(u[2 * i], S2, W2), b = __hfma2(u[2 * i + 1], S2, W2);
                m[i] = __hmin2(__hmin2(m[i], __habs2(a)), __habs2(b));
            } else if (MIX == 3) {
                ub[2 * i] = __hfma2(ub[2 * i], S2b, W2b);
                ub[2 * i + 1] = __hfma2(ub[2 * i + 1], S2b, W2b);
            } else if (MIX == 4) {
                const __nv_bfloat162 a = __hfma2(ub[2 * i], S2b, W2b), b = __hfma2(ub[2 * i + 1], S2b, W2b);
                mb[i] = __hmin2(__hmin2(mb[i], __habs2(a)), __habs2(b));
            } else if (MIX == 6) {
                m[i] = __hmin2(m[i], u[2 * i]);
                u[2 * i] = __hmin2(u[2 * i + 1], m[i]);
            }
        }
        if (MIX == 2 || MIX == 4) {  // feed the results back so that nothing is loop-invariant
            W2 = __hadd2(W2, __hmul2(m[0], __floats2half2_rn(1e-7f, 1e-7f)));
            W2b = __hadd2(W2b, __hmul2(mb[0], __floats2bfloat162_rn(1e-30f, 1e-30f)));
        }
    }
    const long long t1 = clock64();
    float acc = (float)acc_i;
#pragma unroll
    for (int i = 0; i < 8; ++i)
        acc += __low2float(m[i]) + __high2float(m[i]) + __low2float(mb[i]) + __high2float(mb[i]);
#pragma unroll
    for (int i = 0; i < 16; ++i) acc += __low2float(u[i]) + __high2float(u[i]) + __low2float(ub[i]) + __high2float(ub[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MIX>
void run(const char* name, int instr, int rays, int sms, const float4* tab) {
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
    const double clk_per_iter = mean / ITERS / 4.0;  // 4 warps per scheduler
    std::printf("{\"mix\": \"%s\", \"instr_per_iter\": %d, \"clocks_per_warp_iter_per_scheduler\": %.2f, \"clocks_per_instr\": %.3f", name, instr,
                clk_per_iter, clk_per_iter / instr);
    if (rays) std::printf(", \"rays_per_iter\": %d, \"clocks_per_ray_pair\": %.3f", rays, clk_per_iter / (rays / 2));
    std::printf(", \"cycles\": %.0f, \"ms\": %.4f, \"cuda_error\": \"%s\"}\n", mean, ms, cudaGetErrorString(cudaGetLastError()));
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
    run<0>("16xHFMA2", 16, 0, sms, tab);
    run<1>("8xVHMNMX(m,|a|,|b|)", 8, 0, sms, tab);
    run<6>("16xHMNMX2", 16, 0, sms, tab);
    run<2>("8x(2 HFMA2 -> VHMNMX): 32 rays", 24, 32, sms, tab);
    run<5>("8x(2 HFMA2 -> VHMNMX) + LDS.64 + FFMA + F2FP + integer ops: 32 rays", 30, 32, sms, tab);
    run<3>("16xHFMA2.BF16", 16, 0, sms, tab);
    run<4>("8x(2 HFMA2.BF16 -> VHMNMX.BF16): 32 rays", 24, 32, sms, tab);
    cudaFree(tab);
    return 0;
}
