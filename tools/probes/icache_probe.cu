// Instruction-supply probe for sm_100a: straight-line FFMA code of N instructions (no loop reuse inside one
// pass), executed by W warps per SM; prints issued instructions per clock per SM.
// Usage: icache_probe   (sweeps code size and warps per SM)
#include <cstdio>
#include <cuda_runtime.h>

template <int N>
__global__ void __launch_bounds__(1024, 1) straight(float* out, int passes, float a, float b) {
    float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int p = 0; p < passes; p++) {
#pragma unroll
        for (int i = 0; i < N / 8; i++) {
            x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
            x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

template <int N>
void run(float* d, int sms, float clk_ghz) {
    const int warps_list[] = {1, 2, 4, 8, 16, 32};
    for (int w : warps_list) {
        int passes = 64;
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        straight<N><<<sms, 32 * w>>>(d, 2, 1.0001f, 0.5f);
        cudaEventRecord(e0);
        straight<N><<<sms, 32 * w>>>(d, passes, 1.0001f, 0.5f);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        double instr_per_sm = double(N) * passes * w;
        double clk = ms * 1e-3 * clk_ghz * 1e9;
        printf("N=%6d (%4d KB) warps/SM=%2d  %.3f ms  %.2f instr/clk/SM\n", N, N * 16 / 1024, w, ms, instr_per_sm / clk);
    }
}

int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    float ghz = khz * 1e-6f;
    printf("SMs %d clock %.3f GHz\n", sms, ghz);
    float* d; cudaMalloc(&d, sms * 1024 * sizeof(float));
    run<256>(d, sms, ghz);
    run<2048>(d, sms, ghz);
    run<8192>(d, sms, ghz);
    run<16384>(d, sms, ghz);
    return 0;
}
