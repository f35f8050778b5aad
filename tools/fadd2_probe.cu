// tools/fadd2_probe.cu — issue rate of packed fp32 adds (add.rn.f32x2, sm_100+) against scalar FADD:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fadd2_probe tools/fadd2_probe.cu && ./fadd2_probe
// 16 independent chains per thread, 512 threads per SM; prints adds per clock per SM.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned long long add2(unsigned long long a, unsigned long long b) {
    unsigned long long d;
    asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

template <bool PACKED>
__global__ void __launch_bounds__(512, 1) chains(float* out, float seed, int iters, long long* clocks) {
    const long long t0 = clock64();
    if (PACKED) {
        unsigned long long acc[16], v;
        float2 s = make_float2(seed, seed * 0.5f);
        v = *reinterpret_cast<unsigned long long*>(&s);
        for (int i = 0; i < 16; ++i) acc[i] = v + (unsigned long long)(threadIdx.x + i);
#pragma unroll 1
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int i = 0; i < 16; ++i) acc[i] = add2(acc[i], v);
        unsigned long long x = 0;
        for (int i = 0; i < 16; ++i) x ^= acc[i];
        float2 f = *reinterpret_cast<float2*>(&x);
        out[blockIdx.x * blockDim.x + threadIdx.x] = f.x + f.y;
    } else {
        float acc[32];
        for (int i = 0; i < 32; ++i) acc[i] = seed + threadIdx.x + i;
#pragma unroll 1
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int i = 0; i < 32; ++i) acc[i] = __fadd_rn(acc[i], seed);
        float x = 0;
        for (int i = 0; i < 32; ++i) x += acc[i];
        out[blockIdx.x * blockDim.x + threadIdx.x] = x;
    }
    if (threadIdx.x == 0) clocks[blockIdx.x] = clock64() - t0;
}

int main() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float* out;
    long long* clk;
    cudaMalloc(&out, sizeof(float) * 512 * sms);
    cudaMallocManaged(&clk, sizeof(long long) * sms);
    const int iters = 20000;
    for (int packed = 0; packed < 2; ++packed) {
        for (int rep = 0; rep < 2; ++rep) {
            if (packed) chains<true><<<sms, 512>>>(out, 1.0f, iters, clk);
            else chains<false><<<sms, 512>>>(out, 1.0f, iters, clk);
            cudaDeviceSynchronize();
        }
        const double adds = 512.0 * iters * 8 * 32;  // fp32 adds per SM (32 scalar chains = 16 packed chains)
        printf("%s: %.1f fp32 adds / clock / SM (%lld clocks)  %s\n", packed ? "add.rn.f32x2" : "FADD        ", adds / clk[0],
               clk[0], cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
