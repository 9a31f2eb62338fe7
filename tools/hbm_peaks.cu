// Micro-benchmark: what can this B200 sustain for PURE WRITE, pure read and copy streams?
// (MEASURED_PEAKS.json's hbm_gbs is a copy: half reads, half writes.  The LBS kernel is ~99 % writes.)
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__global__ void fill(float4* p, size_t n, float v)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x, s = (size_t)gridDim.x * blockDim.x;
    float4 x = make_float4(v, v, v, v);
    for (; i < n; i += s) p[i] = x;
}
__global__ void fill_cs(float4* p, size_t n, float v) // streaming (evict-first) stores
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x, s = (size_t)gridDim.x * blockDim.x;
    float4 x = make_float4(v, v, v, v);
    for (; i < n; i += s) __stcs(p + i, x);
}
__global__ void rd(const float4* p, size_t n, float* out)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x, s = (size_t)gridDim.x * blockDim.x;
    float a = 0;
    for (; i < n; i += s) { float4 x = __ldg(p + i); a += x.x + x.y + x.z + x.w; }
    if (a == 123.456f) *out = a;
}
__global__ void cp(const float4* a, float4* b, size_t n)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x, s = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += s) b[i] = __ldg(a + i);
}
int main()
{
    const size_t bytes = (size_t)8 << 30, n = bytes / 16;
    float4 *a, *b; float* o;
    CK(cudaMalloc(&a, bytes)); CK(cudaMalloc(&b, bytes)); CK(cudaMalloc(&o, 4));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float ms;
    for (int grid : {148 * 4, 148 * 8, 148 * 16, 148 * 32})
    {
        for (int rep = 0; rep < 2; rep++) fill<<<grid, 512>>>(a, n, 1.f);
        cudaEventRecord(e0); for (int r = 0; r < 5; r++) fill<<<grid, 512>>>(a, n, 2.f); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
        cudaEventElapsedTime(&ms, e0, e1); printf("fill     grid %5d: %7.1f GB/s\n", grid, 5.0 * bytes / ms / 1e6);
        cudaEventRecord(e0); for (int r = 0; r < 5; r++) fill_cs<<<grid, 512>>>(a, n, 2.f); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
        cudaEventElapsedTime(&ms, e0, e1); printf("fill.cs  grid %5d: %7.1f GB/s\n", grid, 5.0 * bytes / ms / 1e6);
        cudaEventRecord(e0); for (int r = 0; r < 5; r++) rd<<<grid, 512>>>(a, n, o); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
        cudaEventElapsedTime(&ms, e0, e1); printf("read     grid %5d: %7.1f GB/s\n", grid, 5.0 * bytes / ms / 1e6);
        cudaEventRecord(e0); for (int r = 0; r < 5; r++) cp<<<grid, 512>>>(a, b, n); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
        cudaEventElapsedTime(&ms, e0, e1); printf("copy     grid %5d: %7.1f GB/s (read+write)\n", grid, 5.0 * 2 * bytes / ms / 1e6);
    }
    cudaEventRecord(e0); for (int r = 0; r < 5; r++) cudaMemsetAsync(a, 0, bytes); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
    cudaEventElapsedTime(&ms, e0, e1); printf("cudaMemset        : %7.1f GB/s\n", 5.0 * bytes / ms / 1e6);
    cudaEventRecord(e0); for (int r = 0; r < 5; r++) cudaMemcpyAsync(b, a, bytes, cudaMemcpyDeviceToDevice); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
    cudaEventElapsedTime(&ms, e0, e1); printf("cudaMemcpy D2D    : %7.1f GB/s (read+write)\n", 5.0 * 2 * bytes / ms / 1e6);
    return 0;
}
