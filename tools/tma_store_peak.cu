// Micro-benchmark: what does the STORE SIDE of lbs_cta_kernel sustain on its own?
// Same output geometry as bench.py's LBS launch (25 024 ring slots x 16 384 padded vertices, 12 B + 24 B per vertex),
// same tensor maps (32 instances x 128 B boxes, 128-byte swizzle), same 9 tensor stores per 32-vertex superchunk with
// 2 or 3 staging tiles in flight - but nothing is computed: the tiles are filled once.  The number is the ceiling the
// compute side of the kernel is measured against (profiles/README.md).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tma_store_peak tools/tma_store_peak.cu && /tmp/tma_store_peak
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

template<int TILES, int LANES>
__global__ void __launch_bounds__(32, 1) store_only(const __grid_constant__ CUtensorMap tm_pos, const __grid_constant__ CUtensorMap tm_nt, int n_sc, int vsplit, int stagger)
{
    extern __shared__ __align__(1024) unsigned char sm[];
    const int lane = threadIdx.x;
    for (int i = lane; i < TILES * 36864 / 4; i += 32)
        reinterpret_cast<float*>(sm)[i] = (float)i;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    const int tile = blockIdx.x / vsplit, split = blockIdx.x - tile * vsplit;
    const int per = (n_sc + vsplit - 1) / vsplit, s_lo = split * per, s_hi = min(n_sc, s_lo + per);
    const int row0 = tile * 32;
    const int n = s_hi - s_lo;
    const int rot = stagger ? (int)((tile * 2654435761u >> 8) % (unsigned)max(n, 1)) : 0; // every CTA starts its run somewhere else
    for (int it = 0; it < n; it++)
    {
        const int s = s_lo + (it + rot) % n;
        const unsigned char* tb = sm + (size_t)(it % TILES) * 36864;
        const int            v0 = s * 32;
        if (LANES == 1)
        {
            if (lane == 0)
            {
                for (int k = 0; k < 3; k++)
                    tma_store_2d(&tm_pos, v0 * 3 + k * 32, row0, tb + k * 4096);
                for (int k = 0; k < 6; k++)
                    tma_store_2d(&tm_nt, v0 * 6 + k * 32, row0, tb + (3 + k) * 4096);
            }
        }
        else if (lane < 9)
        {
            if (lane < 3)
                tma_store_2d(&tm_pos, v0 * 3 + lane * 32, row0, tb + lane * 4096);
            else
                tma_store_2d(&tm_nt, v0 * 6 + (lane - 3) * 32, row0, tb + lane * 4096);
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(TILES - 1) : "memory");
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
    const cuuint32_t box[2] = {32, 32}, estr[2] = {1, 1};
    return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
               CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template<int TILES, int LANES>
static int run(const CUtensorMap& tp, const CUtensorMap& tn, long long rows, int Vp, int vsplit, double bytes, int stagger, int pitch)
{
    const int n_sc = Vp / 32;
    const int smem = 200 * 1024; // like the real kernel: one CTA per SM, whatever the number of tiles
    CK(cudaFuncSetAttribute(store_only<TILES, LANES>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    const unsigned grid = (unsigned)(rows / 32 * vsplit);
    for (int r = 0; r < 2; r++)
        store_only<TILES, LANES><<<grid, 32, smem>>>(tp, tn, n_sc, vsplit, stagger);
    cudaEventRecord(e0);
    for (int r = 0; r < 5; r++)
        store_only<TILES, LANES><<<grid, 32, smem>>>(tp, tn, n_sc, vsplit, stagger);
    cudaEventRecord(e1);
    CK(cudaEventSynchronize(e1));
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    printf("tensor-map stores only: row pitch %5d vertices, stagger %d, tiles %d, vsplit %2d, grid %5u: %.3f ms per launch, %7.1f GB/s\n", pitch, stagger, TILES,
           vsplit, grid, ms / 5, 5.0 * bytes / ms / 1e6);
    return 0;
}

int main()
{
    const long long rows = 25024;
    const int       Vp   = 16384;
    float *         dp, *dn;
    const int       max_pitch = 16384 + 512;
    CK(cudaMalloc(&dp, (size_t)rows * max_pitch * 12));
    CK(cudaMalloc(&dn, (size_t)rows * max_pitch * 24));
    void*                           fn = nullptr;
    cudaDriverEntryPointQueryResult qr;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
    const double bytes = (double)rows * Vp * 36;
    for (int pitch : {16384, 16384 + 32, 16384 + 96, 16384 + 160, 16384 + 352})
    {
        CUtensorMap tp, tn;
        if (!make_map((encode_fn)fn, &tp, dp, rows, (long long)pitch * 3) || !make_map((encode_fn)fn, &tn, dn, rows, (long long)pitch * 6))
        {
            printf("cuTensorMapEncodeTiled failed\n");
            return 1;
        }
        for (int stagger = 0; stagger < 2; stagger++)
            for (int vsplit : {1, 3, 6})
                if (run<3, 9>(tp, tn, rows, Vp, vsplit, bytes, stagger, pitch))
                    return 1;
    }
    // sustained: the same stream for ~3 s (does the 1 kW power cap, which pulls the SM clock down under the full kernel, touch it?)
    {
        CUtensorMap tp, tn;
        make_map((encode_fn)fn, &tp, dp, rows, (long long)16384 * 3), make_map((encode_fn)fn, &tn, dn, rows, (long long)16384 * 6);
        const int smem = 200 * 1024;
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0), cudaEventCreate(&e1);
        for (int r = 0; r < 1200; r++)
            store_only<3, 9><<<2346, 32, smem>>>(tp, tn, Vp / 32, 3, 0);
        cudaEventRecord(e0);
        for (int r = 0; r < 100; r++)
            store_only<3, 9><<<2346, 32, smem>>>(tp, tn, Vp / 32, 3, 0);
        cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1));
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        printf("sustained (after ~2.7 s of the same launches): %.3f ms per launch, %7.1f GB/s\n", ms / 100, 100.0 * bytes / ms / 1e6);
    }
    return 0;
}
