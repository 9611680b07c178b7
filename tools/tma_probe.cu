// tma_probe.cu — isolates the TMA tiled load used by k2_fused_scale_kernel (debug tool, not product).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cudaTypedefs.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cstdint>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int RANK>
__global__ void probe(const __grid_constant__ CUtensorMap map, uint32_t *out, int bw, int bh, int cx, int cy)
{
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    uint32_t *tile = reinterpret_cast<uint32_t *>(smem);
    if (threadIdx.x == 0) {
        uint32_t b = smem_u32(&bar);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        uint32_t bytes = (uint32_t)bw * bh * 4;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
        if (RANK == 3)
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(smem_u32(tile)), "l"(reinterpret_cast<uint64_t>(&map)), "r"(cx), "r"(cy), "r"(0), "r"(b) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(smem_u32(tile)), "l"(reinterpret_cast<uint64_t>(&map)), "r"(cx), "r"(cy), "r"(b) : "memory");
    }
    __syncthreads();
    uint32_t b = smem_u32(&bar), done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred q;\n\tmbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2;\n\tselp.u32 %0, 1, 0, q;\n\t}"
                     : "=r"(done) : "r"(b), "r"(0u) : "memory");
    for (int i = threadIdx.x; i < bw * bh; i += blockDim.x) out[i] = tile[i];
}

int main(int argc, char **argv)
{
    int W = 1920, H = 1080;
    int bw = argc > 1 ? atoi(argv[1]) : 132, bh = argc > 2 ? atoi(argv[2]) : 32;
    int rank = argc > 3 ? atoi(argv[3]) : 3, cx = argc > 4 ? atoi(argv[4]) : 6, cy = argc > 5 ? atoi(argv[5]) : 4;
    size_t pitch = (size_t)W * 4;
    std::vector<uint32_t> h((size_t)W * H);
    for (size_t i = 0; i < h.size(); i++) h[i] = (uint32_t)i;
    uint32_t *d_in, *d_out;
    cudaMalloc(&d_in, h.size() * 4);
    cudaMalloc(&d_out, (size_t)bw * bh * 4);
    cudaMemcpy(d_in, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    auto enc = (PFN_cuTensorMapEncodeTiled_v12000)fn;
    CUtensorMap map;
    cuuint64_t dims[3] = { (cuuint64_t)W, (cuuint64_t)H, 1 };
    cuuint64_t strides[2] = { pitch, pitch * H };
    cuuint32_t box[3] = { (cuuint32_t)bw, (cuuint32_t)bh, 1 };
    cuuint32_t es[3] = { 1, 1, 1 };
    CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_UINT32, rank, d_in, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rank %d box %dx%d -> %d\n", rank, bw, bh, (int)r);
    size_t smem = (size_t)bw * bh * 4;
    cudaFuncSetAttribute(probe<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(probe<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (rank == 3) probe<3><<<1, 256, smem>>>(map, d_out, bw, bh, cx, cy);
    else probe<2><<<1, 256, smem>>>(map, d_out, bw, bh, cx, cy);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    if (e == cudaSuccess) {
        std::vector<uint32_t> o((size_t)bw * bh);
        cudaMemcpy(o.data(), d_out, o.size() * 4, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int y = 0; y < bh; y++)
            for (int x = 0; x < bw; x++) {
                int gx = cx + x, gy = cy + y;
                uint32_t want = (gx < 0 || gy < 0 || gx >= W || gy >= H) ? 0u : (uint32_t)((size_t)gy * W + gx);
                if (o[(size_t)y * bw + x] != want) bad++;
            }
        printf("mismatches: %d\n", bad);
    }
    return 0;
}
