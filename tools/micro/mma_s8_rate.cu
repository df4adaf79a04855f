// Micro-benchmark: issue rate of legacy mma.sync.m16n8k32 s8 on sm_100a (is an int8 tensor-core DCT worth building?)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_s8_rate mma_s8_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void mma_s8(int (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int CHAINS>
__global__ void __launch_bounds__(192) k_mma(int iters, int* out, uint32_t seed) {
    uint32_t a[4] = {seed, seed * 3, seed * 5, seed * 7}, b[2] = {seed * 11 + threadIdx.x, seed * 13};
    int c[CHAINS][4] = {};
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < CHAINS; j++) mma_s8(c[j], a, b);
    }
    int s = 0;
#pragma unroll
    for (int j = 0; j < CHAINS; j++) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
    if (s == 0x12345678) out[0] = s;
}

// mixed: per mma, K fp32 adds on the ALU pipe (does the tensor pipe overlap with ALU issue?)
template <int CHAINS, int ALU>
__global__ void __launch_bounds__(192) k_mix(int iters, int* out, uint32_t seed) {
    uint32_t a[4] = {seed, seed * 3, seed * 5, seed * 7}, b[2] = {seed * 11 + threadIdx.x, seed * 13};
    int c[CHAINS][4] = {};
    float f[8];
    for (int j = 0; j < 8; j++) f[j] = (float) (seed + j);
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < CHAINS; j++) {
            mma_s8(c[j], a, b);
#pragma unroll
            for (int k = 0; k < ALU; k++) f[k & 7] = __fadd_rn(f[k & 7], f[(k + 1) & 7]);
        }
    }
    int s = 0;
#pragma unroll
    for (int j = 0; j < CHAINS; j++) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
    float fs = 0;
    for (int j = 0; j < 8; j++) fs += f[j];
    if (s == 0x12345678 || fs == 1.2345f) out[0] = s;
}

template <typename F>
static float time_ms(F f) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}

int main() {
    int* out; cudaMalloc(&out, 4);
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount, iters = 4096;
    const double clk = p.clockRate * 1e3;
    printf("%s, %d SMs, %.0f MHz\n", p.name, sms, clk / 1e6);
    for (int ctas = 1; ctas <= 4; ctas *= 2) {
        float ms = time_ms([&] { k_mma<8><<<sms * ctas, 192>>>(iters, out, 1); });
        double n = (double) sms * ctas * 6 * iters * 8;
        printf("mma only, %d CTA/SM x 6 warps: %.3f ms, %.2f mma/clk/SM (at max clock), %.1f Tops\n", ctas, ms, n / (ms * 1e-3) / clk / sms,
               n * 16 * 8 * 32 * 2 / (ms * 1e-3) / 1e12);
    }
    {
        float ms = time_ms([&] { k_mix<8, 0><<<sms * 4, 192>>>(iters, out, 1); });
        printf("mix ALU=0: %.3f ms\n", ms);
        ms = time_ms([&] { k_mix<8, 4><<<sms * 4, 192>>>(iters, out, 1); });
        printf("mix ALU=4: %.3f ms\n", ms);
        ms = time_ms([&] { k_mix<8, 8><<<sms * 4, 192>>>(iters, out, 1); });
        printf("mix ALU=8: %.3f ms\n", ms);
        ms = time_ms([&] { k_mix<8, 16><<<sms * 4, 192>>>(iters, out, 1); });
        printf("mix ALU=16: %.3f ms\n", ms);
    }
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return 1; }
    return 0;
}
