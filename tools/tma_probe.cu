// tma_probe.cu -- stand-alone check of the TMA plumbing me_refine uses: one BW x BH byte box of a padded 8-bit plane into
// shared memory, box origin at any byte (x, y).  usage: tma_probe <mode 1|2> <x> <box_w> <box_h>   (mode 2: descriptor in global memory)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o tma_probe tma_probe.cu   (run on a B200)
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__global__ void probe(const __grid_constant__ CUtensorMap map, const CUtensorMap *gmap, int x, int y, unsigned *out, int mode, int bytes) {
    __shared__ __align__(128) unsigned win[512];
    __shared__ __align__(8) unsigned long long bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bytes) : "memory");
        const CUtensorMap *m = mode == 2 ? gmap : &map;
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     ::"r"(smem_u32(win)), "l"(reinterpret_cast<unsigned long long>(m)), "r"(smem_u32(&bar)), "r"(x), "r"(y) : "memory");
    }
    __syncwarp();
    asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(smem_u32(&bar)), "r"(0) : "memory");
    for (int i = threadIdx.x; i < 512; i += 32) out[i] = win[i];
}
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s -> %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
int main(int argc, char **argv) {
    const int mode = argc > 1 ? atoi(argv[1]) : 1, X = argc > 2 ? atoi(argv[2]) : 13, BW = argc > 3 ? atoi(argv[3]) : 32, BH = argc > 4 ? atoi(argv[4]) : 22;
    const int stride = 256, rows = 200, Y = 5;
    std::vector<unsigned char> h(stride * rows);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (unsigned char)(i * 7 + (i >> 8));
    unsigned char *d; unsigned *o;
    CK(cudaMalloc(&d, h.size())); CK(cudaMalloc(&o, 512 * 4));
    CK(cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice));
    void *fn = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    typedef CUresult (*F)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    alignas(64) CUtensorMap m; memset(&m, 0, sizeof(m));
    cuuint64_t dims[2] = {(cuuint64_t)stride, (cuuint64_t)rows}, strides[1] = {(cuuint64_t)stride};
    cuuint32_t box[2] = {(cuuint32_t)BW, (cuuint32_t)BH}, es[2] = {1, 1};
    CUresult r = ((F)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("mode %d x %d box %dx%d: encode -> %d; ", mode, X, BW, BH, (int)r);
    CUtensorMap *gm; CK(cudaMalloc(&gm, sizeof(m))); CK(cudaMemcpy(gm, &m, sizeof(m), cudaMemcpyHostToDevice));
    probe<<<1, 32>>>(m, gm, X, Y, o, mode, BW * BH);
    cudaError_t e = cudaDeviceSynchronize();
    printf("run: %s; ", cudaGetErrorString(e));
    if (e != cudaSuccess) { printf("\n"); return 1; }
    std::vector<unsigned> got(512);
    CK(cudaMemcpy(got.data(), o, 512 * 4, cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int r2 = 0; r2 < BH; ++r2)
        for (int c = 0; c < BW; ++c) {
            const int i = r2 * BW + c;
            unsigned char v = (unsigned char)(got[i / 4] >> (8 * (i & 3)));
            if (v != h[(Y + r2) * stride + X + c]) ++bad;
        }
    printf("mismatches: %d\n", bad);
    return bad != 0;
}
