// Integer issue-rate microbenchmark for the int-ALU roofline: dependency-free streams of
// IADD3/LOP3 (alu pipe), IMAD (fma pipe) and a 1:1 mix, 8 independent chains per thread, every SM filled.
// Prints one JSON object: thread-instructions per second for each mix, and the SM clock it implies.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gpurun_out/int_peak tools/int_peak.cu && gpurun_out/int_peak
#include <cstdio>
#include <cuda_runtime.h>

constexpr int CHAINS = 8, ITERS = 4096;

template <int MIX> __global__ void __launch_bounds__(256) k_issue(int *out, int a, int b)
{
    int x[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; c++) x[c] = threadIdx.x + c;
    for (int i = 0; i < ITERS; i++) {
#pragma unroll
        for (int c = 0; c < CHAINS; c++) {
            if (MIX == 0) {                 // 2 alu-pipe instructions
                x[c] = (x[c] + a) ^ b;
            } else if (MIX == 1) {          // 2 fma-pipe instructions
                x[c] = x[c] * a + b;
                x[c] = x[c] * b + a;
            } else {                        // 1 + 1
                x[c] = (x[c] * a + b) ^ a;  // IMAD + LOP3
            }
        }
    }
    int s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; c++) s += x[c];
    if (s == 0x7fffffff) out[0] = s;
}

template <int MIX> static double run(int *d, int sms, double per_iter)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = sms * 8;
    k_issue<MIX><<<blocks, 256>>>(d, 3, 5);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0);
        k_issue<MIX><<<blocks, 256>>>(d, 3 + r, 5);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return (double)blocks * 256 * ITERS * CHAINS * per_iter / (best * 1e-3);
}

int main()
{
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, 0) != cudaSuccess) { printf("{\"error\": \"no device\"}\n"); return 1; }
    int *d;
    cudaMalloc(&d, 4);
    const double alu = run<0>(d, p.multiProcessorCount, 2), fma = run<1>(d, p.multiProcessorCount, 2),
                 mix = run<2>(d, p.multiProcessorCount, 2);
    int clk = 0;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"sm_clock_khz_max\": %d, \"alu_thread_instr_per_s\": %.4g, "
           "\"imad_thread_instr_per_s\": %.4g, \"mixed_thread_instr_per_s\": %.4g, "
           "\"mixed_lanes_per_clk_per_sm_at_max_clock\": %.1f, "
           "\"how\": \"8 independent chains/thread, 256 thr x 8 blocks/SM, best of 5, CUDA events; SASS-checked: 2 "
           "instructions per chain-iteration in each mix (VIADD+LOP3, IMAD+IMAD, IMAD+LOP3)\"}\n",
           p.name, p.multiProcessorCount, clk, alu, fma, mix, mix / ((double)clk * 1e3) / p.multiProcessorCount);
    return 0;
}
