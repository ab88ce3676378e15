// micro-benchmark: issue rate of legacy tensor / integer instructions on sm_100a (per SM per clock)
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#define ITERS 4096
__global__ void k_imma(int * out) {
    int c[4][4] = {}; int a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
    for (int i = 0; i < ITERS; i++) {
#pragma unroll
        for (int u = 0; u < 4; u++)
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                         : "+r"(c[u][0]), "+r"(c[u][1]), "+r"(c[u][2]), "+r"(c[u][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = c[0][0] + c[1][1] + c[2][2] + c[3][3];
}
__global__ void k_hmma(float * out) {
    float c[4][4] = {}; unsigned a0 = threadIdx.x | 0x3c003c00, a1 = a0, a2 = a0, a3 = a0, b0 = a0, b1 = a0;
    for (int i = 0; i < ITERS; i++) {
#pragma unroll
        for (int u = 0; u < 4; u++)
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                         : "+f"(c[u][0]), "+f"(c[u][1]), "+f"(c[u][2]), "+f"(c[u][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = c[0][0] + c[1][1] + c[2][2] + c[3][3];
}
__global__ void k_dp4a(int * out) {
    int c[8] = {}; int a = threadIdx.x * 0x01010101, b = a * 7;
    for (int i = 0; i < ITERS; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++) c[u] = __dp4a(a + u, b, c[u]);
    }
    int s = 0; for (int u = 0; u < 8; u++) s += c[u];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_imad(int * out) {
    int c[8] = {}; int a = threadIdx.x, b = a * 7 + 1;
    for (int i = 0; i < ITERS; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++) c[u] = c[u] * b + a;
    }
    int s = 0; for (int u = 0; u < 8; u++) s += c[u];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <typename F> float timeit(F f) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}
int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    void * buf; cudaMalloc(&buf, 1 << 24);
    const int ctas = sms * 4, thr = 128;   // 16 warps per SM
    const double warps = (double) ctas * thr / 32;
    float ms;
    ms = timeit([&] { k_imma<<<ctas, thr>>>((int *) buf); });
    printf("IMMA.16832.S8 : %.3f ms  -> %.1f cycles per warp-instr per SMSP, %.0f int8 MAC/clk/SM (clock %d MHz)\n", ms, ms * 1e-3 * khz * 1e3 / (warps * ITERS * 4 / (sms * 4)), warps * ITERS * 4 * 4096.0 / (ms * 1e-3 * khz * 1e3) / sms, khz / 1000);
    ms = timeit([&] { k_hmma<<<ctas, thr>>>((float *) buf); });
    printf("HMMA.16816.F32: %.3f ms  -> %.1f cycles per warp-instr per SMSP, %.0f fp16 MAC/clk/SM\n", ms, ms * 1e-3 * khz * 1e3 / (warps * ITERS * 4 / (sms * 4)), warps * ITERS * 4 * 2048.0 / (ms * 1e-3 * khz * 1e3) / sms);
    ms = timeit([&] { k_dp4a<<<ctas, thr>>>((int *) buf); });
    printf("IDP.4A        : %.3f ms  -> %.1f cycles per warp-instr per SMSP, %.0f int8 MAC/clk/SM\n", ms, ms * 1e-3 * khz * 1e3 / (warps * ITERS * 8 / (sms * 4)), warps * ITERS * 8 * 128.0 / (ms * 1e-3 * khz * 1e3) / sms);
    ms = timeit([&] { k_imad<<<ctas, thr>>>((int *) buf); });
    printf("IMAD          : %.3f ms  -> %.1f cycles per warp-instr per SMSP\n", ms, ms * 1e-3 * khz * 1e3 / (warps * ITERS * 8 / (sms * 4)));
    return 0;
}
