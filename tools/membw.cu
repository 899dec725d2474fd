// Calibration only (not part of the product): pure-read HBM bandwidth with the access shapes the engine uses.
//   ./membw [GiB]     1) grid-stride float4 sum, L1 no-allocate    2) same with default caching   3) copy
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ float4 ld_na(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

template <int U, bool NA>
__global__ void read_kernel(const float4* __restrict__ p, size_t n, float* out) {
    size_t i = (size_t)blockIdx.x * blockDim.x * U + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x * U;
    float acc = 0.f;
    for (; i + (U - 1) * blockDim.x < n; i += stride) {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = NA ? ld_na(p + i + u * blockDim.x) : __ldg(p + i + u * blockDim.x);
#pragma unroll
        for (int u = 0; u < U; ++u) acc += v[u].x + v[u].y + v[u].z + v[u].w;
    }
    if (acc == 12345.678f) *out = acc;
}

__global__ void copy_kernel(const float4* __restrict__ a, float4* __restrict__ b, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) b[i] = a[i];
}

template <typename F>
float time_it(F f, int reps = 5) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    f();
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        cudaEventRecord(a);
        f();
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    return best;
}

int main(int argc, char** argv) {
    const double gib = argc > 1 ? atof(argv[1]) : 16.0;
    const size_t bytes = (size_t)(gib * (1ull << 30));
    const size_t n = bytes / 16;
    float4 *a, *b;
    float* out;
    if (cudaMalloc(&a, bytes) != cudaSuccess || cudaMalloc(&b, bytes) != cudaSuccess) { printf("alloc failed\n"); return 1; }
    cudaMalloc(&out, 4);
    cudaMemset(a, 1, bytes);
    cudaMemset(b, 0, bytes);
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    for (int mult : {2, 4, 8, 16}) {
        float t;
        t = time_it([&] { read_kernel<8, true><<<sms * mult, 256>>>(a, n, out); });
        printf("read  U8 na  grid=%4d x256: %8.1f GB/s\n", sms * mult, bytes / t / 1e6);
        t = time_it([&] { read_kernel<16, true><<<sms * mult, 256>>>(a, n, out); });
        printf("read  U16 na grid=%4d x256: %8.1f GB/s\n", sms * mult, bytes / t / 1e6);
        t = time_it([&] { read_kernel<8, false><<<sms * mult, 256>>>(a, n, out); });
        printf("read  U8 ldg grid=%4d x256: %8.1f GB/s\n", sms * mult, bytes / t / 1e6);
    }
    float t = time_it([&] { copy_kernel<<<sms * 8, 256>>>(a, b, n); });
    printf("copy kernel              : %8.1f GB/s (read+write)\n", 2.0 * bytes / t / 1e6);
    t = time_it([&] { cudaMemcpyAsync(b, a, bytes, cudaMemcpyDeviceToDevice); });
    printf("cudaMemcpy D2D           : %8.1f GB/s (read+write)\n", 2.0 * bytes / t / 1e6);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
