// L2 / HBM read-bandwidth microbenchmark (the extend kernel's node and primitive fetches are
// bounded by L2, for which MEASURED_PEAKS.json has no figure).  A grid of 148 x 8 blocks streams a buffer with
// 128-bit loads `reps` times; a buffer that fits L2 (32 MB) measures the L2 read stream, one far beyond it (2 GB)
// measures HBM.  Prints one JSON line.     nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2bw l2bw.cu
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

__global__ void __launch_bounds__(256) k_read(const float4* __restrict__ p, size_t n4, int reps, float* sink) {
    float acc = 0.f;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int r = 0; r < reps; ++r) {
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
            float4 v;
            asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p + i));  // .cg: bypass L1
            acc += v.x + v.y + v.z + v.w;
        }
    }
    if (acc == 123.456f) *sink = acc;
}

static double run(size_t bytes, int reps) {
    float4* p;
    float* sink;
    cudaMalloc(&p, bytes);
    cudaMalloc(&sink, 4);
    cudaMemset(p, 0, bytes);
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0;
    for (int t = 0; t < 5; ++t) {
        k_read<<<sms * 8, 256>>>(p, bytes / 16, t == 0 ? 1 : reps, sink);  // t = 0 warms the cache
        cudaEventRecord(e0);
        k_read<<<sms * 8, 256>>>(p, bytes / 16, reps, sink);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double gbs = (double)bytes * reps / (ms * 1e-3) / 1e9;
        if (gbs > best) best = gbs;
    }
    cudaFree(p);
    cudaFree(sink);
    return best;
}

int main() {
    int l2 = 0;
    cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, 0);
    const double l2_16 = run((size_t)16 << 20, 200), l2_32 = run((size_t)32 << 20, 100), l2_64 = run((size_t)64 << 20, 50);
    const double hbm = run((size_t)2 << 30, 4);
    if (cudaGetLastError() != cudaSuccess) { fprintf(stderr, "cuda error\n"); return 1; }
    printf("{\"l2_bytes\": %d, \"l2_read_gbs_16MB\": %.1f, \"l2_read_gbs_32MB\": %.1f, \"l2_read_gbs_64MB\": %.1f, \"hbm_read_gbs_2GB\": %.1f}\n", l2, l2_16, l2_32,
           l2_64, hbm);
    return 0;
}
