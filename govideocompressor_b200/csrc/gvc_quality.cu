// gvc_quality.cu -- K9: PSNR / SSIM measurement of a coded plane against its source on the device (include/gvc.h "quality").
//
// Tier 2 of north_star compares whole segments with libx264's output by PSNR, SSIM and bitrate; this is the measuring
// device for it.  Arithmetic follows x264's common/pixel.c: pixel_ssd_wxh (sum of squared differences
// over every sample) and pixel_ssim_wxh -- ssim_4x4x2_core's four integer sums per 4x4 block (s1 = sum a, s2 = sum b,
// ss = sum a^2 + b^2, s12 = sum a b), combined over the 8x8 windows that step by four samples (four neighbouring blocks) by
// ssim_end1's float formula.  The integer sums are exact; every window's value is computed with separately rounded float
// operations (no fused multiply-add), i.e. the same floats as the C code gives; they are summed in double, per CTA on the
// device in a fixed order and across CTAs on the host (x264 itself accumulates in float).
// Oracle: oracle/quality.c; parity test: tests/test_gpu_quality.py.
#include <cstdint>
#include <vector>

#include "gvc_common.h"

namespace gvc {
namespace {

constexpr int kQTile = 16;                 // 4x4 blocks per tile side (64 x 64 samples); one more row / column of blocks as halo
struct QPartial { unsigned long long ssd; double ssim; };

__device__ __forceinline__ unsigned ld4(const uint8_t *row, int x, int w) {     // samples x..x+3 of a row, 0 beyond w
    unsigned v = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (x + i < w) v |= (unsigned)row[x + i] << (8 * i);
    return v;
}
__device__ __forceinline__ float ssim_end1(int s1, int s2, int ss, int s12) {    // x264 ssim_end1, 8-bit depth
    const int   c1 = 416, c2 = 235963;      // (.01*255)^2 * 64 and (.03*255)^2 * 64 * 63, rounded as x264 does
    const int   vars = ss * 64 - s1 * s1 - s2 * s2, covar = s12 * 64 - s1 * s2;
    const float num = __fmul_rn((float)(2 * s1 * s2 + c1), (float)(2 * covar + c2));
    const float den = __fmul_rn((float)(s1 * s1 + s2 * s2 + c1), (float)(vars + c2));
    return __fdiv_rn(num, den);
}

__global__ void __launch_bounds__(256) psnr_ssim_kernel(const uint8_t *a, int sa, const uint8_t *b, int sb, int w, int h, QPartial *out) {
    __shared__ int4 blk[kQTile + 1][kQTile + 1];          // s1, s2, ss, s12 of every complete 4x4 block of the tile (+ halo)
    __shared__ unsigned long long ssd_part[8];
    __shared__ double ssim_part[8];
    const int bx0 = blockIdx.x * kQTile, by0 = blockIdx.y * kQTile;      // first block of the tile
    const int bw = w >> 2, bh = h >> 2;                                    // complete blocks in the plane
    unsigned long long ssd = 0;
    for (int i = threadIdx.x; i < (kQTile + 1) * (kQTile + 1); i += 256) {
        const int ly = i / (kQTile + 1), lx = i - ly * (kQTile + 1), bx = bx0 + lx, by = by0 + ly;
        int s1 = 0, s2 = 0, ss = 0, s12 = 0;
        unsigned sq = 0;
        for (int r = 0; r < 4; ++r) {
            const int y = 4 * by + r;
            if (y >= h || 4 * bx >= w) break;
            const unsigned pa = ld4(a + (size_t)y * sa, 4 * bx, w), pb = ld4(b + (size_t)y * sb, 4 * bx, w);
            s1 = (int)__dp4a(pa, 0x01010101u, (unsigned)s1); s2 = (int)__dp4a(pb, 0x01010101u, (unsigned)s2);
            ss = (int)__dp4a(pa, pa, __dp4a(pb, pb, (unsigned)ss)); s12 = (int)__dp4a(pa, pb, (unsigned)s12);
            const unsigned d = __vabsdiffu4(pa, pb);
            sq = __dp4a(d, d, sq);
        }
        blk[ly][lx] = make_int4(s1, s2, ss, s12);
        if (lx < kQTile && ly < kQTile) ssd += sq;         // every sample belongs to exactly one tile's own blocks (partial blocks included)
    }
    __syncthreads();
    double ssim = 0.0;
    {
        const int ly = threadIdx.x / kQTile, lx = threadIdx.x - ly * kQTile, bx = bx0 + lx, by = by0 + ly;
        if (bx + 1 < bw && by + 1 < bh) {                  // the 8x8 window whose top left block is (bx, by)
            const int4 p = blk[ly][lx], q = blk[ly][lx + 1], r = blk[ly + 1][lx], t = blk[ly + 1][lx + 1];
            ssim = (double)ssim_end1(p.x + q.x + r.x + t.x, p.y + q.y + r.y + t.y, p.z + q.z + r.z + t.z, p.w + q.w + r.w + t.w);
        }
    }
    for (int d = 16; d; d >>= 1) { ssd += __shfl_down_sync(0xffffffffu, ssd, d); ssim += __shfl_down_sync(0xffffffffu, ssim, d); }
    if ((threadIdx.x & 31) == 0) { ssd_part[threadIdx.x >> 5] = ssd; ssim_part[threadIdx.x >> 5] = ssim; }
    __syncthreads();
    if (threadIdx.x == 0) {
        QPartial o{0, 0.0};
        for (int k = 0; k < 8; ++k) { o.ssd += ssd_part[k]; o.ssim += ssim_part[k]; }
        out[blockIdx.y * gridDim.x + blockIdx.x] = o;
    }
}

}  // namespace
}  // namespace gvc

using namespace gvc;

extern "C" int gvc_psnr_ssim(const uint8_t *a, int stride_a, const uint8_t *b, int stride_b, int w, int h, unsigned long long *ssd,
                             double *ssim, void *stream) {
    int rc = enter();
    if (rc) return rc;
    if (!a || !b || w < 1 || h < 1 || stride_a < w || stride_b < w || w > 16384 || h > 16384) { set_error("gvc_psnr_ssim: bad arguments"); return GVC_ERR_INVALID_ARG; }
    const int tx = ((w + 3) / 4 + kQTile - 1) / kQTile, ty = ((h + 3) / 4 + kQTile - 1) / kQTile;
    QPartial *d = nullptr;
    GVC_CUDA(cudaMalloc((void **)&d, sizeof(QPartial) * tx * ty));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    psnr_ssim_kernel<<<dim3(tx, ty), 256, 0, st>>>(a, stride_a, b, stride_b, w, h, d);
    std::vector<QPartial> host(size_t(tx) * ty);
    cudaError_t ce = cudaGetLastError();
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(host.data(), d, sizeof(QPartial) * host.size(), cudaMemcpyDeviceToHost, st);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(st);
    cudaFree(d);
    if (ce != cudaSuccess) { set_error("gvc_psnr_ssim: %s", cudaGetErrorString(ce)); return GVC_ERR_CUDA; }
    unsigned long long s = 0;
    double             m = 0.0;
    for (const QPartial &p : host) { s += p.ssd; m += p.ssim; }
    const long long windows = (long long)((w >> 2) - 1) * ((h >> 2) - 1);
    if (ssd) *ssd = s;
    if (ssim) *ssim = windows > 0 ? m / (double)windows : 1.0;
    return GVC_OK;
}
