// gvc_entropy.cu -- kernels and launcher of the GPU CAVLC writer (entropy_kernels.cuh).
#include <cuda_runtime.h>

#include "enc_launch.h"
#include "entropy_kernels.cuh"
#include "gvc_common.h"

namespace gvc {
namespace {

constexpr int kEntWarps = 4;
// every kernel of the pass: grid.y = segment of the batch (its EntropyPicture record)
__global__ void __launch_bounds__(32 * kEntWarps) ent_mb_kernel(const __grid_constant__ Batch<EntropyPicture> B) {
    __shared__ EntSmem sms[kEntWarps];
    const EntropyArgs &a = B.a[blockIdx.y].ent;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int mb = blockIdx.x * kEntWarps + warp;
    if (mb >= a.mbw * a.mbh) return;
    EntSmem &sm = sms[warp];
    if (lane == 0) { sm.kind = 0; sm.overflow = 0; }
    __syncwarp();
    const int bits = ent_lane_encode(lane, a, mb, sm);
    if (!__any_sync(0xffffffffu, bits != 0)) {       // P_Skip: nothing to lay out
        __syncwarp();
        if (lane == 0) {
            a.mb_bits[mb] = 0;
            a.mb_kind[mb] = (uint8_t)(sm.kind | (sm.overflow ? 128 : 0));
        }
        return;
    }
    // bit offsets in coding order: thread p takes the length of the lane that sits at position p, scans, hands it back
    int v = __shfl_sync(0xffffffffu, bits, ent_lane_at(lane));
    const int mine = v;
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += o;
    }
    const int total = __shfl_sync(0xffffffffu, v, 31);
    const int off = __shfl_sync(0xffffffffu, v - mine, lane < kEntLanes ? ent_lane_order(lane) : lane);
    const int words = (total + 31) >> 5;
    for (int j = lane; j < words; j += 32) sm.mb[j] = 0;
    __syncwarp();
    if (lane < kEntLanes && bits) ent_or_bits(sm.mb, kMbBufWords, (unsigned long long)off, sm.lane_buf[lane], bits);
    __syncwarp();
    uint32_t *dst = a.mb_buf + (size_t)mb * kMbBufWords;
    for (int j = lane; j < words; j += 32) dst[j] = sm.mb[j];
    if (lane == 0) {
        a.mb_bits[mb] = (uint32_t)total;
        a.mb_kind[mb] = (uint8_t)(sm.kind | (sm.overflow ? 128 : 0));
    }
}

constexpr int kScanThreads = 1024;

// inclusive scan over the CTA's threads; op: 0 = sum, 1 = max (signed).  `total` = the value of the last thread.
template <int OP> __device__ __forceinline__ int block_scan(int v, int *warp_part, int &total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v = OP ? max(v, o) : v + o;
    }
    if (lane == 31) warp_part[warp] = v;
    __syncthreads();
    if (warp == 0) {
        int w = warp_part[lane];
        for (int d = 1; d < 32; d <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, w, d);
            if (lane >= d) w = OP ? max(w, o) : w + o;
        }
        warp_part[lane] = w;
    }
    __syncthreads();
    if (warp > 0) v = OP ? max(v, warp_part[warp - 1]) : v + warp_part[warp - 1];
    total = warp_part[31];
    __syncthreads();
    return v;
}

__global__ void __launch_bounds__(kScanThreads) ent_scan_kernel(const __grid_constant__ Batch<EntropyPicture> B) {
    const SliceArgs &a = B.a[blockIdx.x].slice;
    __shared__ int      part[32];
    __shared__ unsigned counts[4];
    __shared__ int      any_overflow;
    const int t = threadIdx.x, per = (a.nmb + kScanThreads - 1) / kScanThreads;
    const int lo = min(t * per, a.nmb), hi = min(lo + per, a.nmb);
    if (t < 4) counts[t] = 0;
    if (t == 0) any_overflow = 0;
    __syncthreads();
    // last coded macroblock at or before each position
    int      last = -1;
    unsigned c[4] = {0, 0, 0, 0};
    bool     ovf = false;
    for (int i = lo; i < hi; ++i) {
        const int k = a.mb_kind[i];
        if (k & 127) last = i;
        ++c[k & 3];
        ovf |= (k & 128) != 0;
    }
    for (int k = 0; k < 4; ++k)
        if (c[k]) atomicAdd(&counts[k], c[k]);
    if (ovf) any_overflow = 1;
    int last_all;
    const int incl_last = block_scan<1>(last, part, last_all);
    int prev = __shfl_up_sync(0xffffffffu, incl_last, 1);          // exclusive: the previous thread's inclusive value
    if ((t & 31) == 0) prev = t ? part[(t >> 5) - 1] : -1;         // (part[] still holds the per-warp inclusive maxima)
    __syncthreads();                                               // part[] is reused by the next scan
    // bits each thread's macroblocks take, mb_skip_run included
    int sum = 0;
    {
        int p = prev;
        for (int i = lo; i < hi; ++i)
            if (a.mb_kind[i] & 127) {
                const unsigned run = (unsigned)(i - 1 - p);
                a.mb_run[i] = run;
                sum += (a.p_slice ? ue_len(run) : 0) + (int)a.mb_bits[i];
                p = i;
            }
    }
    int data_bits;
    const int incl_sum = block_scan<0>(sum, part, data_bits);
    {
        unsigned off = (unsigned)(a.header_bits + incl_sum - sum);
        int      p = prev;
        for (int i = lo; i < hi; ++i)
            if (a.mb_kind[i] & 127) {
                a.mb_off[i] = off;
                off += (unsigned)((a.p_slice ? ue_len((unsigned)(i - 1 - p)) : 0) + (int)a.mb_bits[i]);
                p = i;
            }
    }
    // clear what the slice will occupy (ent_place ORs into it)
    const long long need = min((long long)a.slice_words, ((long long)a.header_bits + data_bits + 64 + 1 + 31) / 32 + 4);
    for (long long j = t; j < need; j += kScanThreads) a.slice[j] = 0;
    __syncthreads();
    if (t == 0) {
        const unsigned cc[4] = {counts[0], counts[1], counts[2], counts[3]};
        ent_finish_slice(a, (unsigned long long)data_bits, (unsigned)(a.nmb - 1 - last_all), cc, any_overflow != 0);
    }
}

__global__ void __launch_bounds__(128) ent_place_kernel(const __grid_constant__ Batch<EntropyPicture> B) {
    const SliceArgs &a = B.a[blockIdx.y].slice;
    const int mb = blockIdx.x * 128 + threadIdx.x;
    if (mb < a.nmb) ent_place(a, mb);
}

// slice bytes -> the segment's buffer in mapped host memory, appended after the previous pictures
__global__ void __launch_bounds__(256) ent_copy_kernel(const __grid_constant__ Batch<EntropyPicture> B) {
    const EntropyPicture &p = B.a[blockIdx.y];
    const uint32_t *slice = p.slice.slice, *info = p.slice.info;
    uint8_t *host_buf = p.host_buf;
    const unsigned host_cap = p.host_cap;
    unsigned *used = p.used;        // [f], [f+1]
    uint32_t *pos = p.pos;          // host: base, bytes, bits, skipped, p16, p8, intra, error
    const unsigned bytes = info[1], base = used[0], padded = (bytes + 15u) & ~15u;
    const bool     fits = info[6] == 0 && base + padded <= host_cap;
    if (fits) {
        uint4       *dst = reinterpret_cast<uint4 *>(host_buf + base);
        const uint4 *src = reinterpret_cast<const uint4 *>(slice);
        for (unsigned j = blockIdx.x * 256 + threadIdx.x; j < padded / 16; j += gridDim.x * 256) {
            uint4 v = src[j];
            v.x = __byte_perm(v.x, 0, 0x0123); v.y = __byte_perm(v.y, 0, 0x0123); v.z = __byte_perm(v.z, 0, 0x0123); v.w = __byte_perm(v.w, 0, 0x0123);
            dst[j] = v;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        used[1] = fits ? base + padded : base;
        pos[0] = base; pos[1] = bytes; pos[2] = info[0]; pos[3] = info[2]; pos[4] = info[3]; pos[5] = info[4]; pos[6] = info[5];
        pos[7] = fits ? 0u : 1u;
    }
}

}  // namespace

int launch_entropy_picture(const EntropyPicture *p, int count, cudaStream_t st) {
    if (count < 1 || count > kMaxBatch) { set_error("batch of %d segments outside 1..%d", count, kMaxBatch); return GVC_ERR_INVALID_ARG; }
    Batch<EntropyPicture> B;
    for (int i = 0; i < kMaxBatch; ++i) B.a[i] = p[i < count ? i : 0];
    const int nmb = p[0].ent.mbw * p[0].ent.mbh;
    ent_mb_kernel<<<dim3((nmb + kEntWarps - 1) / kEntWarps, count), 32 * kEntWarps, 0, st>>>(B);
    GVC_CUDA(cudaGetLastError());
    ent_scan_kernel<<<count, kScanThreads, 0, st>>>(B);
    GVC_CUDA(cudaGetLastError());
    ent_place_kernel<<<dim3((nmb + 127) / 128, count), 128, 0, st>>>(B);
    GVC_CUDA(cudaGetLastError());
    ent_copy_kernel<<<dim3(16, count), 256, 0, st>>>(B);
    GVC_CUDA(cudaGetLastError());
    return GVC_OK;
}

}  // namespace gvc
