// dense_loop_probe.cu -- measurement tool (not product code).  Where do the cycles of the grid
// kernel's dense loop go?  The loop of raycast_kernels.cu (bound_group) is rebuilt with its cost
// centres switchable: rays per thread (PAIRS), where the row terms come from (computed per thread as
// in the shipped kernel / read from a per-row table in shared memory), and how the results are
// consumed (LOP3 bit-AND as shipped / folded into the FFMA2 chain, which measures the FFMA2 stream
// alone).  Times are CUDA-event wall times of a one-wave launch, converted to cycles at the maximum SM
// clock (an upper bound on real cycles).  NOTE: a per-CTA clock64() span was also tried and is NOT
// reported: thread 0 sits in the oldest warp of its CTA, which the scheduler favours, so its span is
// far shorter than the launch (it made the loop look issue-bound at 1 instruction/clock; ncu's
// sm__cycles_elapsed and these wall times agree on about 53 cycles per thread-primitive instead).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o dense_loop_probe dense_loop_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define NPRIM 256
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

// ROW: 0 = row terms computed per thread (1 FFMA + 1 FADD + 2 FMUL, LDS.128 + LDS.32 per primitive)
//      1 = row terms from a shared-memory table [prim][4 rows] (LDS.64, lane-dependent) + sA (LDS.32)
//      2 = as 1 but the table entry is (sA, rowW, rowC, -) read with ONE LDS.128
// ACC: 0 = LOP3 AND of the bit patterns (shipped)   1 = stage 2 accumulates into itself (no LOP3)
//      2 = LOP3 only on every second pair (timing only)
template <int PAIRS, int ROW, int ACC, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) probe(const float4* __restrict__ recs, float* out, long long* cyc,
                                                       int iters, float x0) {
    __shared__ __align__(16) float4 sh[2 * NPRIM];          // ROW 0: bound records (32 B each)
    __shared__ __align__(16) float4 tab4[ROW == 2 ? NPRIM * 4 : 1];
    __shared__ __align__(16) float2 tab2[ROW == 1 ? NPRIM * 4 : 1];
    __shared__ float sa[ROW == 1 ? NPRIM : 1];
    for (int i = threadIdx.x; i < 2 * NPRIM; i += blockDim.x) sh[i] = recs[i];
    __syncthreads();
    const int ly = (threadIdx.x >> 3) & 3;
    const float v = (float)ly + blockIdx.x;
    if (ROW != 0) {
        for (int i = threadIdx.x; i < NPRIM * 4; i += blockDim.x) {
            const int p = i >> 2, r = i & 3;
            const float4 r0 = sh[2 * p];
            const float cv = sh[2 * p + 1].x;
            const float vv = (float)r + blockIdx.x;
            const float rowW = fmaf(r0.y, vv, r0.z), dv = vv - cv, rowC = (r0.w * dv) * dv;
            if (ROW == 1) { tab2[i] = make_float2(rowW, rowC); if (r == 0) sa[p] = r0.x; }
            if (ROW == 2) tab4[i] = make_float4(r0.x, rowW, rowC, 0.f);
        }
        __syncthreads();
    }
    uint64_t u2[PAIRS];
#pragma unroll
    for (int j = 0; j < PAIRS; ++j) {
        const float u = x0 + (threadIdx.x & 7) * (2 * PAIRS) + 2 * j;
        u2[j] = pack2(u, u + 1.0f);
    }
    uint64_t chain[PAIRS];
#pragma unroll
    for (int j = 0; j < PAIRS; ++j) chain[j] = pack2(0.f, 0.f);
    uint32_t hits = 0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll 1
        for (int p = 0; p < NPRIM; p += G) {
            uint32_t acc[G];
#pragma unroll
            for (int g = 0; g < G; ++g) {
                float sA, rowW, rowC;
                if (ROW == 0) {
                    const float4 r0 = sh[2 * (p + g)];
                    const float cv = *reinterpret_cast<const float*>(&sh[2 * (p + g) + 1]);
                    sA = r0.x;
                    rowW = fmaf(r0.y, v, r0.z);
                    const float dv = v - cv;
                    rowC = (r0.w * dv) * dv;
                } else if (ROW == 1) {
                    const float2 t = tab2[(p + g) * 4 + ly];
                    sA = sa[p + g]; rowW = t.x; rowC = t.y;
                } else {
                    const float4 t = tab4[(p + g) * 4 + ly];
                    sA = t.x; rowW = t.y; rowC = t.z;
                }
                const uint64_t S2 = pack2(sA, sA), W2 = pack2(rowW, rowW), C2 = pack2(rowC, rowC);
                uint32_t a = 0xFFFFFFFFu;
#pragma unroll
                for (int j = 0; j < PAIRS; ++j) {
                    const uint64_t w2 = ffma2(u2[j], S2, W2);
                    if (ACC == 1) {
                        chain[j] = ffma2(w2, w2, chain[j]);
                    } else {
                        uint32_t q0, q1;
                        unpack2u(ffma2(w2, w2, C2), q0, q1);
                        if (ACC == 0 || (j & 1)) a &= q0 & q1; else chain[j] ^= (uint64_t)q0;  // ACC 2: wrong maths
                    }
                }
                if (ACC == 1) a = __float_as_uint(rowC);
                acc[g] = a;
            }
            uint32_t all = acc[0];
#pragma unroll
            for (int g = 1; g < G; ++g) all &= acc[g];
            if (__any_sync(0xFFFFFFFFu, (all & 0x40000000u) == 0u)) hits += 1;
        }
    }
    const long long t1 = clock64();
    uint32_t x = hits;
#pragma unroll
    for (int j = 0; j < PAIRS; ++j) { uint32_t a, b; unpack2u(chain[j], a, b); x ^= a ^ b; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = (float)x;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int PAIRS, int ROW, int ACC, int THREADS, int MINB>
void run(const char* what, const float4* d_rec, float* d_out, long long* d_cyc, int sms, double clk_ghz) {
    int occ = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, probe<PAIRS, ROW, ACC, THREADS, MINB>, THREADS, 0);
    const int blocks = sms * occ, iters = 8;
    cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, probe<PAIRS, ROW, ACC, THREADS, MINB>);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int w = 0; w < 2; ++w) probe<PAIRS, ROW, ACC, THREADS, MINB><<<blocks, THREADS>>>(d_rec, d_out, d_cyc, iters, 100.0f);
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        probe<PAIRS, ROW, ACC, THREADS, MINB><<<blocks, THREADS>>>(d_rec, d_out, d_cyc, iters, 100.0f);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    long long* h = new long long[blocks];
    cudaMemcpy(h, d_cyc, sizeof(long long) * blocks, cudaMemcpyDeviceToHost);
    double mean = 0; for (int i = 0; i < blocks; ++i) mean += (double)h[i]; mean /= blocks;
    delete[] h;
    const double warps_per_smsp = occ * (THREADS / 32) / 4.0;
    const double tp = warps_per_smsp * iters * NPRIM;  // warp-level thread-primitives per SMSP
    (void)mean;
    const double cyc_wall = best * 1e-3 * clk_ghz * 1e9 / tp;
    // dispatch model that fits every line within 10 %: an FFMA2 takes 2 dispatch cycles, anything else 1
    const int lop = (ACC == 0 ? PAIRS : (ACC == 2 ? PAIRS : 0));
    const double model = PAIRS * 2 * 2.0 + (ROW == 0 ? 4.0 + 2.0 : (ROW == 1 ? 2.0 : 1.0)) + lop + 2.0;
    const double pipe = PAIRS * 2 * 2.0;  // FMA-pipe cycles of the per-ray FFMA2 alone
    std::printf("{\"what\": \"%s\", \"pairs\": %d, \"row\": %d, \"acc\": %d, \"threads\": %d, \"ctas_per_sm\": %d, \"regs\": %d, "
                "\"ms\": %.4f, \"cycles_per_thread_prim\": %.2f, \"cycles_per_test\": %.3f, "
                "\"dispatch_model_cycles\": %.1f, \"ffma2_share_of_cycles\": %.3f, \"err\": \"%s\"}\n",
                what, PAIRS, ROW, ACC, THREADS, occ, fa.numRegs, best, cyc_wall, cyc_wall / (2.0 * PAIRS), model,
                pipe / cyc_wall, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const double clk = clk_khz * 1e-6;
    std::printf("{\"device\": \"%s\", \"sms\": %d, \"clock_ghz\": %.3f}\n", pr.name, pr.multiProcessorCount, clk);
    float4* h = new float4[2 * NPRIM];
    for (int i = 0; i < NPRIM; ++i) {  // far-away ellipses: q is large, the vote never fires
        h[2 * i] = make_float4(1e-3f + 1e-6f * i, 1e-5f, -9.0f - i, 1e-3f);
        h[2 * i + 1] = make_float4(-7000.0f, 0.f, 0.f, 0.f);
    }
    float4* d_rec; cudaMalloc(&d_rec, sizeof(float4) * 2 * NPRIM);
    cudaMemcpy(d_rec, h, sizeof(float4) * 2 * NPRIM, cudaMemcpyHostToDevice);
    const int sms = pr.multiProcessorCount;
    float* d_out; cudaMalloc(&d_out, sizeof(float) * sms * 16 * 256);
    long long* d_cyc; cudaMalloc(&d_cyc, sizeof(long long) * sms * 16);
    run<8, 0, 0, 128, 4>("shipped shape: 16 rays, row terms computed, LOP3", d_rec, d_out, d_cyc, sms, clk);
    run<8, 0, 0, 128, 3>("same, 3 CTAs/SM", d_rec, d_out, d_cyc, sms, clk);
    run<8, 0, 0, 256, 2>("same, 256 threads x 2", d_rec, d_out, d_cyc, sms, clk);
    run<8, 1, 0, 128, 4>("16 rays, row table LDS.64 + LDS.32, LOP3", d_rec, d_out, d_cyc, sms, clk);
    run<8, 2, 0, 128, 4>("16 rays, row table LDS.128, LOP3", d_rec, d_out, d_cyc, sms, clk);
    run<8, 0, 1, 128, 4>("16 rays, row terms computed, FFMA2 stream only (no LOP3)", d_rec, d_out, d_cyc, sms, clk);
    run<8, 2, 1, 128, 4>("16 rays, row table LDS.128, FFMA2 stream only", d_rec, d_out, d_cyc, sms, clk);
    run<8, 0, 2, 128, 4>("16 rays, row terms computed, half the LOP3s", d_rec, d_out, d_cyc, sms, clk);
    run<16, 0, 0, 128, 2>("32 rays, row terms computed, LOP3, 2 CTAs/SM", d_rec, d_out, d_cyc, sms, clk);
    run<16, 0, 0, 128, 3>("32 rays, row terms computed, LOP3, 3 CTAs/SM", d_rec, d_out, d_cyc, sms, clk);
    run<16, 2, 0, 128, 3>("32 rays, row table LDS.128, LOP3, 3 CTAs/SM", d_rec, d_out, d_cyc, sms, clk);
    run<16, 0, 1, 128, 3>("32 rays, FFMA2 stream only", d_rec, d_out, d_cyc, sms, clk);
    run<4, 0, 0, 128, 8>("8 rays, row terms computed, LOP3, 8 CTAs/SM", d_rec, d_out, d_cyc, sms, clk);
    run<4, 2, 0, 128, 8>("8 rays, row table LDS.128, LOP3, 8 CTAs/SM", d_rec, d_out, d_cyc, sms, clk);
    return 0;
}
