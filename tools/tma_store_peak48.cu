// Micro-benchmark (companion of tma_store_peak.cu): the same output streams written by PER-WARP tensor stores of 32 instances x 48 bytes
// (4 vertices of positions, or 2 vertices of normal+tangent) - the shape a barrier-free, warp-private variant of lbs_cta_kernel would use.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/tma_store_peak48 tools/tma_store_peak48.cu -lcuda
#include <cstdint>
#include <cstdio>
#include <cuda.h>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tma_store_2d(const void* tmap, int c0, int c1, const void* src)
{
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(src)) : "memory");
}
// 8 warps per CTA; warp w owns vertices 4w..4w+3 of every 32-vertex superchunk; DEPTH mini-tiles (3 x 1536 B each) in flight per warp
template<int DEPTH>
__global__ void __launch_bounds__(256, 1) store_only48(const __grid_constant__ CUtensorMap tm_pos, const __grid_constant__ CUtensorMap tm_nt, int n_sc, int vsplit)
{
    extern __shared__ __align__(1024) unsigned char sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char* mine = sm + (size_t)warp * DEPTH * 4608;
    for (int i = lane; i < DEPTH * 4608 / 4; i += 32)
        reinterpret_cast<float*>(mine)[i] = (float)i;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    const int tile = blockIdx.x / vsplit, split = blockIdx.x - tile * vsplit;
    const int per = (n_sc + vsplit - 1) / vsplit, s_lo = split * per, s_hi = min(n_sc, s_lo + per);
    const int row0 = tile * 32;
    for (int s = s_lo; s < s_hi; s++)
    {
        const unsigned char* tb = mine + (size_t)((s - s_lo) % DEPTH) * 4608;
        const int            v0 = s * 32 + warp * 4;
        if (lane == 0)
            tma_store_2d(&tm_pos, v0 * 3, row0, tb);
        else if (lane == 1)
            tma_store_2d(&tm_nt, v0 * 6, row0, tb + 1536);
        else if (lane == 2)
            tma_store_2d(&tm_nt, v0 * 6 + 12, row0, tb + 3072);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(DEPTH - 1) : "memory");
        __syncwarp();
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                              CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static bool make_map(encode_fn enc, CUtensorMap* tm, float* base, long long rows, long long row_floats)
{
    const cuuint64_t dims[2] = {(cuuint64_t)row_floats, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)row_floats * 4};
    const cuuint32_t box[2] = {12, 32}, estr[2] = {1, 1};
    return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
               CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
template<int DEPTH>
static int run(const CUtensorMap& tp, const CUtensorMap& tn, long long rows, int Vp, int vsplit, double bytes)
{
    const int smem = 200 * 1024;
    CK(cudaFuncSetAttribute(store_only48<DEPTH>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    const unsigned grid = (unsigned)(rows / 32 * vsplit);
    for (int r = 0; r < 2; r++)
        store_only48<DEPTH><<<grid, 256, smem>>>(tp, tn, Vp / 32, vsplit);
    cudaEventRecord(e0);
    for (int r = 0; r < 5; r++)
        store_only48<DEPTH><<<grid, 256, smem>>>(tp, tn, Vp / 32, vsplit);
    cudaEventRecord(e1);
    CK(cudaEventSynchronize(e1));
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    printf("per-warp 32 x 48 B tensor stores: depth %d, vsplit %2d: %.3f ms per launch, %7.1f GB/s\n", DEPTH, vsplit, ms / 5, 5.0 * bytes / ms / 1e6);
    return 0;
}
int main()
{
    const long long rows = 25024;
    const int       Vp   = 16384;
    float *         dp, *dn;
    CK(cudaMalloc(&dp, (size_t)rows * Vp * 12));
    CK(cudaMalloc(&dn, (size_t)rows * Vp * 24));
    void*                           fn = nullptr;
    cudaDriverEntryPointQueryResult qr;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
    CUtensorMap tp, tn;
    if (!make_map((encode_fn)fn, &tp, dp, rows, (long long)Vp * 3) || !make_map((encode_fn)fn, &tn, dn, rows, (long long)Vp * 6))
    {
        printf("cuTensorMapEncodeTiled failed\n");
        return 1;
    }
    const double bytes = (double)rows * Vp * 36;
    for (int vsplit : {1, 3})
        if (run<2>(tp, tn, rows, Vp, vsplit, bytes) || run<3>(tp, tn, rows, Vp, vsplit, bytes) || run<4>(tp, tn, rows, Vp, vsplit, bytes))
            return 1;
    return 0;
}
