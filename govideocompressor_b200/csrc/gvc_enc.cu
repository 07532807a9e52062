// gvc_enc.cu -- launchers and C ABI of the encoder-side kernels (include/gvc.h "encoder stages").
// Kernel bodies: enc_kernels.cuh.  These entry points take DEVICE pointers and are asynchronous on `stream`.
#include <cuda.h>

#include <cstring>

#include "enc_host.h"
#include "enc_kernels.cuh"
#include "enc_launch.h"
#include "gvc_common.h"
#include "tma.cuh"

namespace gvc {

__global__ void __launch_bounds__(256) extend_border_kernel(uint8_t *p, int stride, int w, int h, int x_lo, int x_hi, int y_lo, int y_hi) {
    const int x = x_lo + blockIdx.x * 256 + threadIdx.x, y = y_lo + blockIdx.y;
    if (x < x_hi && y < y_hi) extend_border_px(p, stride, w, h, x, y);
}

__global__ void __launch_bounds__(256) lowres_kernel(const uint8_t *src, int sstride, int x_lo, int x_hi, int y_lo, int y_hi,
                                                     uint8_t *d0, uint8_t *dh, uint8_t *dv, uint8_t *dc, int dstride,
                                                     int lx_lo, int lx_hi, int ly_lo, int ly_hi) {
    const int lx = lx_lo + blockIdx.x * 256 + threadIdx.x, ly = ly_lo + blockIdx.y;
    if (lx < lx_hi && ly < ly_hi) lowres_px(src, sstride, x_lo, x_hi, y_lo, y_hi, d0, dh, dv, dc, dstride, lx, ly);
}

__global__ void __launch_bounds__(256) extend_border_batch_kernel(const __grid_constant__ Batch<BorderBatchArgs> B) {
    const BorderBatchArgs &a = B.a[blockIdx.z];
    const int f = blockIdx.y / 3, k = blockIdx.y - 3 * f;
    int x, y;
    if (f < a.n_frames && border_piece_xy(a, k, blockIdx.x, threadIdx.x, x, y))
        extend_border_px(a.p[k] + (size_t)f * a.frame_stride[k], a.stride[k], a.w[k], a.h[k], x, y);
}

__global__ void __launch_bounds__(256) extend_border16_batch_kernel(const __grid_constant__ Batch<BorderBatchArgs> B) {
    const BorderBatchArgs &a = B.a[blockIdx.z];
    const int f = blockIdx.y / 3, k = blockIdx.y - 3 * f;
    int x, y;
    if (f < a.n_frames && border16_piece_xy(a, k, blockIdx.x * 8 + (threadIdx.x >> 5), threadIdx.x & 31, x, y))
        extend_border_16(a.p[k] + (size_t)f * a.frame_stride[k], a.stride[k], a.w[k], a.h[k], x, y);
}

// grid.z = segment * frames_max + picture
__global__ void __launch_bounds__(256) lowres_batch_kernel(const __grid_constant__ Batch<LowresBatchArgs> B, int frames_max) {
    const int seg = blockIdx.z / frames_max, f = blockIdx.z - seg * frames_max;
    const LowresBatchArgs &a = B.a[seg];
    const int lx = a.lx_lo + 4 * (blockIdx.x * 256 + threadIdx.x), ly = a.ly_lo + blockIdx.y;      // four samples per thread
    if (f < a.n_frames && lx < a.lx_hi && ly < a.ly_hi)
        lowres_quad(a, lowres_aligned(a), a.src + (size_t)f * a.sframe, a.dst + (size_t)f * a.dframe, lx, ly);
}

__global__ void __launch_bounds__(kHpThreads) hpel_kernel(const __grid_constant__ Batch<HpelArgs> B) {
    __shared__ HpelSmem sm;
    const HpelArgs &a = B.a[blockIdx.z];
    const int x0 = a.x_lo + blockIdx.x * kHpTileW, y0 = a.y_lo + blockIdx.y * kHpTileH;
    const bool al = hpel_aligned(a);
    hpel_phase_a(threadIdx.x, a, al, x0, y0, sm);
    __syncthreads();
    hpel_phase_b(threadIdx.x, a, al, x0, y0, sm);
    __syncthreads();
    hpel_phase_c(threadIdx.x, a, al, x0, y0, sm);
}

// The same three planes for word-aligned arguments, without shared memory: a lane owns one 4-sample column and walks down
// kHpColRows rows with the six source rows of the vertical filter in registers (packed as in hpel_phase_b); the
// horizontal neighbours' words and vertical sums come from the adjacent lanes by shuffle.  Lanes 0 and 31 only feed their
// neighbours, so a warp writes 30 words (120 samples) per row.
constexpr int kHpColRows = 16, kHpColWarps = 4, kHpColSpan = 120;
__global__ void __launch_bounds__(32 * kHpColWarps) hpel_cols_kernel(const __grid_constant__ Batch<HpelArgs> B) {
    const HpelArgs &a = B.a[blockIdx.z];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x = a.x_lo - 4 + kHpColSpan * blockIdx.x + 4 * lane;
    const int y0 = a.y_lo + kHpColRows * (blockIdx.y * kHpColWarps + warp);
    if (y0 >= a.y_hi) return;
    // edge replication: a word outside the readable columns is the edge sample four times
    const int      xc = min(max(x, a.cx_lo), a.cx_hi - 4);
    const unsigned sel = x < a.cx_lo ? 0x0000u : (x >= a.cx_hi ? 0x3333u : 0x3210u);
    const bool     ok = lane >= 1 && lane <= 30 && x < a.x_hi;
    const uint8_t *sp = a.src + xc;
    const int      stride = a.stride, cy_lo = a.cy_lo, cy_top = a.cy_hi - 1, y_hi = a.y_hi;
    unsigned e[6], o[6];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const int      yy = min(max(y0 - 2 + k, cy_lo), cy_top);
        const unsigned w = __byte_perm(__ldg(reinterpret_cast<const unsigned *>(sp + (long long)yy * stride)), 0, sel);
        e[k] = pack_even(w);
        o[k] = pack_odd(w);
    }
    long long ofs = (long long)y0 * stride + x;
#pragma unroll
    for (int r = 0; r < kHpColRows; ++r) {
        const int y = y0 + r;
        {
            const int      yy = min(max(y + 3, cy_lo), cy_top);
            const unsigned w = __byte_perm(__ldg(reinterpret_cast<const unsigned *>(sp + (long long)yy * stride)), 0, sel);
            e[(r + 5) % 6] = pack_even(w);
            o[(r + 5) % 6] = pack_odd(w);
        }
        const unsigned e0 = e[r % 6], e1 = e[(r + 1) % 6], e2 = e[(r + 2) % 6], e3 = e[(r + 3) % 6], e4 = e[(r + 4) % 6], e5 = e[(r + 5) % 6];
        const unsigned o0 = o[r % 6], o1 = o[(r + 1) % 6], o2 = o[(r + 2) % 6], o3 = o[(r + 3) % 6], o4 = o[(r + 4) % 6], o5 = o[(r + 5) % 6];
        unsigned ve = (e0 + e5) + 20u * (e2 + e3) - 5u * (e1 + e4);
        unsigned vo = (o0 + o5) + 20u * (o2 + o3) - 5u * (o1 + o4);
        ve += (ve & 0x8000u) << 1;
        vo += (vo & 0x8000u) << 1;
        // V plane
        const unsigned be = (bias_clamp_s16x2(ve, 0x00100010u, 0x1fff1fffu) >> 5) & 0x00ff00ffu;
        const unsigned bo = (bias_clamp_s16x2(vo, 0x00100010u, 0x1fff1fffu) >> 5) & 0x00ff00ffu;
        const unsigned out_v = __byte_perm(be, bo, 0x6240);
        // H plane: samples x-2 .. x+6 of row y
        const unsigned w1 = __byte_perm(e2, o2, 0x6240);
        const unsigned w0 = __shfl_up_sync(0xffffffffu, w1, 1), w2 = __shfl_down_sync(0xffffffffu, w1, 1);
        const int h0 = dp4a_us(__funnelshift_r(w1, w2, 16), kTapB, dp4a_us(__funnelshift_r(w0, w1, 16), kTapA, 16));
        const int h1 = dp4a_us(__funnelshift_r(w1, w2, 24), kTapB, dp4a_us(__funnelshift_r(w0, w1, 24), kTapA, 16));
        const int h2 = dp4a_us(w2, kTapB, dp4a_us(w1, kTapA, 16));
        const int h3 = dp4a_us(w2 >> 8, kTapB, dp4a_us(__funnelshift_r(w1, w2, 8), kTapA, 16));
        const unsigned out_h = pack_sat_u8x4(h0 >> 5, h1 >> 5, h2 >> 5, h3 >> 5);
        // C plane: the taps over the vertical sums of the same samples
        const unsigned p01 = __byte_perm(ve, vo, 0x5410), p23 = __byte_perm(ve, vo, 0x7632);
        const unsigned l23 = __shfl_up_sync(0xffffffffu, p23, 1);
        const unsigned r01 = __shfl_down_sync(0xffffffffu, p01, 1), r23 = __shfl_down_sync(0xffffffffu, p23, 1);
        const unsigned s0 = __funnelshift_r(l23, p01, 16), s1 = __funnelshift_r(p01, p23, 16), s2 = __funnelshift_r(p23, r01, 16),
                       s3 = __funnelshift_r(r01, r23, 16);
        const int q0 = dp2a_ss(p23, kTapB, dp2a_ss(p01, kTapA, dp2a_ss(l23, kTapA, 512, false), true), false);
        const int q1 = dp2a_ss(s2, kTapB, dp2a_ss(s1, kTapA, dp2a_ss(s0, kTapA, 512, false), true), false);
        const int q2 = dp2a_ss(r01, kTapB, dp2a_ss(p23, kTapA, dp2a_ss(p01, kTapA, 512, false), true), false);
        const int q3 = dp2a_ss(s3, kTapB, dp2a_ss(s2, kTapA, dp2a_ss(s1, kTapA, 512, false), true), false);
        const unsigned out_c = pack_sat_u8x4(q0 >> 10, q1 >> 10, q2 >> 10, q3 >> 10);
        if (ok && y < y_hi) {
            *reinterpret_cast<unsigned *>(a.dv + ofs) = out_v;
            *reinterpret_cast<unsigned *>(a.dh + ofs) = out_h;
            *reinterpret_cast<unsigned *>(a.dc + ofs) = out_c;
        }
        ofs += stride;
    }
}

__global__ void __launch_bounds__(128) transform8x8_kernel(const __grid_constant__ Transform8Args a) {
    const int b = blockIdx.x * 128 + threadIdx.x;
    if (b < a.bw * a.bh) transform8x8_block(a, b % a.bw, b / a.bw);
}

constexpr int kMbWarps = 4;   // macroblocks (warps) per CTA in the per-MB kernels

// per-macroblock kernels: a CTA takes kMbWarps neighbouring macroblocks of one row (grid.x), grid.y = macroblock row, grid.z = segment
__global__ void __launch_bounds__(32 * kMbWarps) me_lowres_kernel(const __grid_constant__ Batch<MeLowArgs> B) {
    __shared__ MeLowSmem sms[kMbWarps];
    const MeLowArgs &a = B.a[blockIdx.z];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int mx = blockIdx.x * kMbWarps + warp, my = blockIdx.y;
    if (mx >= a.mbw) return;
    if (me_lowres_zero_wins(a, __reduce_add_sync(0xffffffffu, me_lowres_zero_sad(lane, a, mx, my)))) {
        me_lowres_out(lane, a, mx, my, (unsigned)(a.r * (2 * a.r + 1) + a.r));      // index of displacement (0, 0)
        return;
    }
    me_lowres_load(lane, a, mx, my, sms[warp]);
    __syncwarp();
    const unsigned mine = me_lowres_chunk(a.r) == kLowChunk ? me_lowres_lane<kLowChunk>(lane, a, mx, my, sms[warp])
                                                            : me_lowres_lane<kLowChunkWide>(lane, a, mx, my, sms[warp]);
    const unsigned best = __reduce_min_sync(0xffffffffu, mine);
    me_lowres_out(lane, a, mx, my, best);
}

// ---- TMA plumbing of me_refine (see me_refine.cuh): four tensor maps, one per sample plane of the reference picture, each
// describing the whole padded plane as a 2-D byte tensor (row pitch = the plane's stride); a window is one box of it.
struct MeRefTma {
    CUtensorMap map[4];          // TMA takes the descriptor's ADDRESS, which must stay in parameter space (never a local copy): the
    int on;                      // kernel only forms &T.a[segment].map[p] of its __grid_constant__ parameter.  on = 0: planes not
};                               // addressable by TMA (alignment); the windows are then staged with plain loads
constexpr int kMeWarps = 4;   // macroblocks (warps) per CTA of me_refine: 24.6 KB of windows per CTA
__global__ void __launch_bounds__(32 * kMeWarps, 8)
me_refine_kernel(const __grid_constant__ Batch<MeRefArgs> B, const __grid_constant__ Batch<MeRefTma> T) {
    __shared__ MeRefSmem sms[kMeWarps];
    const MeRefArgs &a = B.a[blockIdx.z];
    const CUtensorMap *maps = T.a[blockIdx.z].map;
    const int tma_on = T.a[blockIdx.z].on;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int mx = blockIdx.x * kMeWarps + warp, my = blockIdx.y;      // a CTA takes kMeWarps neighbouring macroblocks of one row
    if (mx >= a.mbw) return;
    MeRefSmem &sm = sms[warp];
    me_ref_load(lane, a, mx, my, sm);
    __syncwarp();
    me_mvp(sm, lane);
    __syncwarp();
    // zero-vector SAD: per 8x8 quadrant (lanes of one quadrant: same bit 4, same bit 0), then the whole macroblock
    unsigned zq[4];
    {
        unsigned v = me_ref_zero_part(lane, a, mx, my, sm);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
#pragma unroll
        for (int q = 0; q < 4; ++q) zq[q] = __shfl_sync(0xffffffffu, v, ((q >> 1) << 4) | (q & 1));
    }
    const int zero_sad = (int)(zq[0] + zq[1] + zq[2] + zq[3]);
    if (me_ref_early(a, zero_sad)) { me_ref_early_out(lane, a, mx, my, sm, zero_sad); return; }
    const MeGeom g = me_geom(a, sm, mx, my);
    if (tma_on) {
        if (lane == 0) {
            mbar_init(&sm.bar[0], 1);
            mbar_init(&sm.bar[1], 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");      // the barriers exist before the copy engine signals them
            constexpr unsigned kBox = kMeWinPitch * kMeWinRows;
            mbar_expect_tx(&sm.bar[0], 2 * kBox);                                    // tensor coordinates: the plane's padding is part of the tensor
            tma_box_2d(sm.win[0], &maps[0], g.wox + kEncPad, g.woy + kEncPad, &sm.bar[0]);
            tma_box_2d(sm.pred, &maps[0], g.pox + kEncPad, g.poy + kEncPad, &sm.bar[0]);
            mbar_expect_tx(&sm.bar[1], 3 * kBox);
            tma_box_2d(sm.win[1], &maps[1], g.wox + kEncPad, g.woy + kEncPad, &sm.bar[1]);
            tma_box_2d(sm.win[2], &maps[2], g.wox + kEncPad, g.woy + kEncPad, &sm.bar[1]);
            tma_box_2d(sm.win[3], &maps[3], g.wox + kEncPad, g.woy + kEncPad, &sm.bar[1]);
        }
        __syncwarp();
        mbar_wait(&sm.bar[0], 0);
    } else {
        me_stage_plain(lane, a, g, sm);
        __syncwarp();
    }
    // integer candidates, one per lane; the best per partition by a key-packed warp minimum, its vector from the winning lane
    int ccx[5], ccy[5];
    {
        unsigned key[5];
        me_ref_int(lane, a, sm, g, zq, key);
        int cx, cy;
        me_int_cand(sm, g, lane < 27 ? lane : 26, cx, cy);
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const int w = (int)(__reduce_min_sync(0xffffffffu, key[k]) & 31u);
            ccx[k] = __shfl_sync(0xffffffffu, cx, w);
            ccy[k] = __shfl_sync(0xffffffffu, cy, w);
        }
    }
    if (tma_on) mbar_wait(&sm.bar[1], 0);
    MeLane L;
    const bool one = ccx[1] == ccx[0] && ccx[2] == ccx[0] && ccx[3] == ccx[0] && ccx[4] == ccx[0] &&
                     ccy[1] == ccy[0] && ccy[2] == ccy[0] && ccy[3] == ccy[0] && ccy[4] == ccy[0];
    if (one) {      // warp-uniform: both half-warps on the 16x16 partition, two candidates per round
        me_lane_init(lane & 15, a, mx, my, sm, g, ccx[0], ccy[0], L);
        MeBest b;
        b.key = 0xffffffffu; b.x = L.cx; b.y = L.cy;
        for (int step = 2; step >= 1; --step) {
            const MeStepCost sc = me_step_cost(a, L, step);
#pragma unroll
            for (int r = 0; r < 5; ++r) {
                if (step == 1 && r == 4) break;
                const int ci = me_fast_cand(lane, r, step);
                int v = ci < 8 ? me_fast_satd(a, sm, L, ci, step) : 0;
                v += __shfl_xor_sync(0xffffffffu, v, 1);
                v += __shfl_xor_sync(0xffffffffu, v, 2);
                v += __shfl_xor_sync(0xffffffffu, v, 4);
                v += __shfl_xor_sync(0xffffffffu, v, 8);
                if (ci < 8) me_take_desc(L, sc, v, me_desc(L.cx, L.cy, ci, step), ci, step, b);
            }
            // the better half-warp wins (order bits break ties as the sequential search would); it is the next centre
            const unsigned ok = __shfl_xor_sync(0xffffffffu, b.key, 16);
            const int      ox = __shfl_xor_sync(0xffffffffu, b.x, 16), oy = __shfl_xor_sync(0xffffffffu, b.y, 16);
            if (ok < b.key) { b.key = ok; b.x = ox; b.y = oy; }
            me_next_step(L, b);
        }
        me_fast_out(lane, a, mx, my, b);
        return;
    }
    int P, ox, oy;
    me_lane_block(lane, P, ox, oy);
    const int pcx = P == 0 ? ccx[0] : P == 1 ? ccx[1] : P == 2 ? ccx[2] : P == 3 ? ccx[3] : ccx[4];
    const int pcy = P == 0 ? ccy[0] : P == 1 ? ccy[1] : P == 2 ? ccy[2] : P == 3 ? ccy[3] : ccy[4];
    me_lane_init(lane, a, mx, my, sm, g, pcx, pcy, L);
    MeBest b;
    b.key = 0xffffffffu; b.x = L.cx; b.y = L.cy;
    {
        int v[9];
        me_half_all(a, sm, L, v);
        const MeStepCost sc = me_step_cost(a, L, 2);
#pragma unroll
        for (int ci = -1; ci < 8; ++ci) me_take(L, sc, me_group_sum(lane, v[ci + 1]), ci, 2, b);
        me_next_step(L, b);
    }
    {
        int v[8];
        me_quarter_all(a, sm, L, v);
        const MeStepCost sc = me_step_cost(a, L, 1);
#pragma unroll
        for (int ci = 0; ci < 8; ++ci) me_take(L, sc, me_group_sum(lane, v[ci]), ci, 1, b);
    }
    MeBest all[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const int src = k == 0 ? 0 : 12 + 4 * k;
        all[k].key = __shfl_sync(0xffffffffu, b.key, src);
        all[k].x = __shfl_sync(0xffffffffu, b.x, src);
        all[k].y = __shfl_sync(0xffffffffu, b.y, src);
    }
    me_ref_out(lane, a, mx, my, all);
}

__global__ void __launch_bounds__(32 * kMbWarps) encode_inter_kernel(const __grid_constant__ Batch<EncMbArgs> B) {
    __shared__ EncMbSmem sms[kMbWarps];
    const EncMbArgs &a = B.a[blockIdx.z];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int mx = blockIdx.x * kMbWarps + warp, my = blockIdx.y;
    if (mx >= a.mbw) return;
    EncMbSmem &sm = sms[warp];
    const bool al = enc_mb_aligned(a);
    EncLane L;
    enc_mb_p0(lane, a, mx, my, sm);
    __syncwarp();
    enc_mb_p1(lane, a, al, mx, my, sm, L);
    __syncwarp();
    if (__any_sync(0xffffffffu, enc_mb_any_level(lane, a, sm, L))) enc_mb_p2(lane, a, al, mx, my, sm, L);
    else enc_mb_p2_zero(lane, a, al, mx, my, sm, L);      // warp-uniform: nothing of this macroblock survives the quantiser
    __syncwarp();
    enc_mb_p3(lane, a, mx, my, sm);
    __syncwarp();
    enc_mb_p4(lane, a, mx, my, sm, L);
}

// One warp per macroblock row; rows advance as a wavefront (see enc_kernels.cuh K7).
__global__ void __launch_bounds__(32) intra_frame_kernel(const __grid_constant__ Batch<IntraArgs> B) {
    __shared__ IntraSmem sm;
    const IntraArgs &a = B.a[blockIdx.y];
    const int my = blockIdx.x, lane = threadIdx.x;
    if (my >= a.mbh) return;
    const bool al = intra_aligned(a);
    // The CTA is one warp.  __syncthreads() (bar.sync.aligned) instead of __syncwarp() between the phases: an aligned barrier
    // requires the warp to be converged, a warp-level sync does not bring the lanes back together -- with __syncwarp() the
    // two half-warps (luma lanes / chroma + idle lanes) ran the whole loop body separately, 1.8x the instructions (ncu).
    IntraLane L;
    unsigned  next[4] = {0, 0, 0, 0};
    int       row_count = 0;
    intra_load_cur(lane, a, al, 0, my, next);
    for (int mx = 0; mx < a.mbw; ++mx) {
        for (int j = 0; j < 4; ++j) L.cur[j] = next[j];
        if (mx + 1 < a.mbw) intra_load_cur(lane, a, al, mx + 1, my, next);   // in flight while this macroblock is coded
        intra_p0(lane, a, mx, my, sm);
        __syncthreads();
        intra_p1(lane, a, mx, my, sm, L);
        __syncthreads();
        intra_p2(lane, a, mx, my, sm, L);
        __syncthreads();
        intra_p3a(lane, sm);
        __syncthreads();
        intra_p3b(lane, a, sm);
        __syncthreads();
        intra_p3c(lane, sm);
        __syncthreads();
        intra_p3d(lane, a, al, mx, my, sm, L);
        __syncthreads();
        intra_p4(lane, a, mx, my, sm, row_count);
        __syncthreads();
        intra_p5(lane, a, mx, my, sm, L);
        __syncthreads();
    }
    intra_row_done(lane, a, row_count);
}

// ---- launchers shared with the segment encoder -------------------------------------------------
int launch_extend_border(uint8_t *p, int stride, int w, int h, int x_lo, int x_hi, int y_lo, int y_hi, cudaStream_t st) {
    if (x_hi <= x_lo || y_hi <= y_lo) return GVC_OK;
    extend_border_kernel<<<dim3((x_hi - x_lo + 255) / 256, y_hi - y_lo), 256, 0, st>>>(p, stride, w, h, x_lo, x_hi, y_lo, y_hi);
    GVC_CUDA(cudaGetLastError());
    return GVC_OK;
}
int launch_lowres(const uint8_t *src, int sstride, int x_lo, int x_hi, int y_lo, int y_hi, uint8_t *d0, uint8_t *dh, uint8_t *dv,
                  uint8_t *dc, int dstride, int lx_lo, int lx_hi, int ly_lo, int ly_hi, cudaStream_t st) {
    if (lx_hi <= lx_lo || ly_hi <= ly_lo) return GVC_OK;
    lowres_kernel<<<dim3((lx_hi - lx_lo + 255) / 256, ly_hi - ly_lo), 256, 0, st>>>(src, sstride, x_lo, x_hi, y_lo, y_hi, d0, dh, dv, dc,
                                                                                    dstride, lx_lo, lx_hi, ly_lo, ly_hi);
    GVC_CUDA(cudaGetLastError());
    return GVC_OK;
}
// ---- batched launchers: `count` (<= kMaxBatch) argument records of equal geometry, one per segment; one launch covers them all
template <class A> static void fill_batch(Batch<A> &B, const A *a, int count) {
    for (int i = 0; i < count; ++i) B.a[i] = a[i];
    for (int i = count; i < kMaxBatch; ++i) B.a[i] = a[0];      // never indexed by the grid; keeps the parameter block defined
}
#define GVC_CHECK_BATCH(count) do { if ((count) < 1 || (count) > kMaxBatch) { set_error("batch of %d segments outside 1..%d", (count), kMaxBatch); return GVC_ERR_INVALID_ARG; } } while (0)

int launch_extend_border_batch(const BorderBatchArgs *a, int count, cudaStream_t st) {
    GVC_CHECK_BATCH(count);
    int frames = 0;
    bool vec = true;
    for (int i = 0; i < count; ++i) { frames = a[i].n_frames > frames ? a[i].n_frames : frames; vec = vec && border16_ok(a[i]); }
    if (frames < 1) return GVC_OK;
    int pieces = 1;                                   // the largest plane decides the grid; smaller planes' extra blocks exit
    for (int k = 0; k < 3; ++k) {
        const int n = vec ? (border16_pieces(a[0], k) + 7) / 8 : border_pieces(a[0], k);
        pieces = n > pieces ? n : pieces;
    }
    for (int f0 = 0; f0 < frames; f0 += 16384) {     // gridDim.y <= 65535
        Batch<BorderBatchArgs> B;
        fill_batch(B, a, count);
        int nf = 0;
        for (int i = 0; i < kMaxBatch; ++i) {
            BorderBatchArgs &b = B.a[i];
            const int left = b.n_frames - f0;
            b.n_frames = left < 0 ? 0 : left < 16384 ? left : 16384;
            for (int k = 0; k < 3; ++k) b.p[k] += (size_t)f0 * b.frame_stride[k];
            if (i < count) nf = b.n_frames > nf ? b.n_frames : nf;
        }
        if (vec) extend_border16_batch_kernel<<<dim3(pieces, nf * 3, count), 256, 0, st>>>(B);
        else     extend_border_batch_kernel<<<dim3(pieces, nf * 3, count), 256, 0, st>>>(B);
        GVC_CUDA(cudaGetLastError());
    }
    return GVC_OK;
}
int launch_lowres_batch(const LowresBatchArgs *a, int count, cudaStream_t st) {
    GVC_CHECK_BATCH(count);
    int frames = 0;
    for (int i = 0; i < count; ++i) frames = a[i].n_frames > frames ? a[i].n_frames : frames;
    if (frames < 1 || a[0].lx_hi <= a[0].lx_lo || a[0].ly_hi <= a[0].ly_lo) return GVC_OK;
    const int per = 65535 / count;                   // gridDim.z <= 65535
    for (int f0 = 0; f0 < frames; f0 += per) {
        Batch<LowresBatchArgs> B;
        fill_batch(B, a, count);
        int nf = 0;
        for (int i = 0; i < kMaxBatch; ++i) {
            LowresBatchArgs &b = B.a[i];
            const int left = b.n_frames - f0;
            b.n_frames = left < 0 ? 0 : left < per ? left : per;
            b.src += (size_t)f0 * b.sframe; b.dst += (size_t)f0 * b.dframe;
            if (i < count) nf = b.n_frames > nf ? b.n_frames : nf;
        }
        lowres_batch_kernel<<<dim3((a[0].lx_hi - a[0].lx_lo + 1023) / 1024, a[0].ly_hi - a[0].ly_lo, nf * count), 256, 0, st>>>(B, nf);
        GVC_CUDA(cudaGetLastError());
    }
    return GVC_OK;
}
int launch_hpel(const HpelArgs *a, int count, cudaStream_t st) {
    GVC_CHECK_BATCH(count);
    if (a[0].x_hi <= a[0].x_lo || a[0].y_hi <= a[0].y_lo) return GVC_OK;
    Batch<HpelArgs> B;
    fill_batch(B, a, count);
    bool cols = a[0].cx_hi - a[0].cx_lo >= 4;
    for (int i = 0; i < count; ++i) cols = cols && hpel_aligned(a[i]);
    if (cols) {
        const int band = kHpColRows * kHpColWarps;
        hpel_cols_kernel<<<dim3((a[0].x_hi - a[0].x_lo + kHpColSpan - 1) / kHpColSpan, (a[0].y_hi - a[0].y_lo + band - 1) / band, count),
                           32 * kHpColWarps, 0, st>>>(B);
    } else {
        hpel_kernel<<<dim3((a[0].x_hi - a[0].x_lo + kHpTileW - 1) / kHpTileW, (a[0].y_hi - a[0].y_lo + kHpTileH - 1) / kHpTileH, count), kHpThreads, 0, st>>>(B);
    }
    GVC_CUDA(cudaGetLastError());
    return GVC_OK;
}
int launch_me_lowres(const MeLowArgs *a, int count, cudaStream_t st) {
    GVC_CHECK_BATCH(count);
    Batch<MeLowArgs> B;
    fill_batch(B, a, count);
    me_lowres_kernel<<<dim3((a[0].mbw + kMbWarps - 1) / kMbWarps, a[0].mbh, count), 32 * kMbWarps, 0, st>>>(B);
    GVC_CUDA(cudaGetLastError());
    return GVC_OK;
}
__global__ void signal_picture_kernel(const __grid_constant__ Batch<SignalArgs> B) {
    const SignalArgs &a = B.a[threadIdx.x];
    *a.count_out = *a.count_in;
    __threadfence_system();
    *reinterpret_cast<volatile int *>(a.done) = 1;
}
int launch_signal_picture(const SignalArgs *a, int count, cudaStream_t st) {
    GVC_CHECK_BATCH(count);
    Batch<SignalArgs> B;
    fill_batch(B, a, count);
    signal_picture_kernel<<<1, count, 0, st>>>(B);
    GVC_CUDA(cudaGetLastError());
    return GVC_OK;
}
// Describes the padded planes of a.ref to TMA; tma.on = 0 when they cannot be (base or pitch not a multiple of 16 bytes).
static void me_refine_tma(const MeRefArgs &a, MeRefTma &tma) {
    memset(&tma, 0, sizeof(tma));
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc || (a.stride & 15) || a.stride < a.w16 + 2 * kEncPad) return;
    const uint8_t *pl[4] = {a.ref.y, a.ref.hh, a.ref.hv, a.ref.hc};
    for (int p = 0; p < 4; ++p) {
        const uint8_t *base = pl[p] - ((size_t)kEncPad * a.stride + kEncPad);        // first byte of the padded plane
        if (reinterpret_cast<uintptr_t>(base) & 15) return;
        const cuuint64_t dims[2] = {(cuuint64_t)a.stride, (cuuint64_t)(a.h16 + 2 * kEncPad)};
        const cuuint64_t strides[1] = {(cuuint64_t)a.stride};
        const cuuint32_t box[2] = {(cuuint32_t)kMeWinPitch, (cuuint32_t)kMeWinRows}, es[2] = {1, 1};
        if (enc(&tma.map[p], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<uint8_t *>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
            return;
    }
    tma.on = 1;
}
int launch_me_refine(const MeRefArgs *a, int count, cudaStream_t st) {
    GVC_CHECK_BATCH(count);
    Batch<MeRefArgs> B;
    Batch<MeRefTma>  T;
    fill_batch(B, a, count);
    for (int i = 0; i < kMaxBatch; ++i) me_refine_tma(B.a[i], T.a[i]);
    static const bool carve = [] { return cudaFuncSetAttribute(me_refine_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100) == cudaSuccess; }();
    (void)carve;
    me_refine_kernel<<<dim3((a[0].mbw + kMeWarps - 1) / kMeWarps, a[0].mbh, count), 32 * kMeWarps, 0, st>>>(B, T);
    GVC_CUDA(cudaGetLastError());
    return GVC_OK;
}
int launch_encode_inter(const EncMbArgs *a, int count, cudaStream_t st) {
    GVC_CHECK_BATCH(count);
    Batch<EncMbArgs> B;
    fill_batch(B, a, count);
    encode_inter_kernel<<<dim3((a[0].mbw + kMbWarps - 1) / kMbWarps, a[0].mbh, count), 32 * kMbWarps, 0, st>>>(B);
    GVC_CUDA(cudaGetLastError());
    return GVC_OK;
}

int launch_encode_intra(const IntraArgs *a, int count, cudaStream_t st) {
    GVC_CHECK_BATCH(count);
    for (int i = 0; i < count; ++i) GVC_CUDA(cudaMemsetAsync(a[i].mail, 0, sizeof(unsigned long long) * 8 * a[i].mbw * a[i].mbh, st));
    Batch<IntraArgs> B;
    fill_batch(B, a, count);
    intra_frame_kernel<<<dim3(a[0].mbh, count), 32, 0, st>>>(B);
    GVC_CUDA(cudaGetLastError());
    return GVC_OK;
}

}  // namespace gvc

using namespace gvc;

static_assert(sizeof(gvc_mb_motion) == sizeof(MbMotion), "public and internal motion records must match");

extern "C" {

int gvc_extend_border(uint8_t *plane, int stride, int w, int h, int x_lo, int x_hi, int y_lo, int y_hi, void *stream) {
    int rc = enter();
    if (rc) return rc;
    if (!plane || w < 1 || h < 1 || stride < 1) { set_error("gvc_extend_border: bad arguments"); return GVC_ERR_INVALID_ARG; }
    return launch_extend_border(plane, stride, w, h, x_lo, x_hi, y_lo, y_hi, static_cast<cudaStream_t>(stream));
}

int gvc_lowres(const uint8_t *src, int sstride, int x_lo, int x_hi, int y_lo, int y_hi, uint8_t *d0, uint8_t *dh, uint8_t *dv,
               uint8_t *dc, int dstride, int lx_lo, int lx_hi, int ly_lo, int ly_hi, void *stream) {
    int rc = enter();
    if (rc) return rc;
    if (!src || !d0) { set_error("gvc_lowres: null plane"); return GVC_ERR_INVALID_ARG; }
    return launch_lowres(src, sstride, x_lo, x_hi - 1, y_lo, y_hi - 1, d0, dh, dv, dc, dstride, lx_lo, lx_hi, ly_lo, ly_hi,
                         static_cast<cudaStream_t>(stream));
}

int gvc_hpel_filter(const uint8_t *src, uint8_t *dh, uint8_t *dv, uint8_t *dc, int stride, int x_lo, int x_hi, int y_lo, int y_hi,
                    void *stream) {
    int rc = enter();
    if (rc) return rc;
    if (!src || !dh || !dv || !dc) { set_error("gvc_hpel_filter: null plane"); return GVC_ERR_INVALID_ARG; }
    HpelArgs a;
    a.src = src; a.dh = dh; a.dv = dv; a.dc = dc; a.stride = stride;
    a.x_lo = x_lo; a.x_hi = x_hi; a.y_lo = y_lo; a.y_hi = y_hi;
    a.cx_lo = x_lo; a.cx_hi = x_hi; a.cy_lo = y_lo; a.cy_hi = y_hi;      // the C ABI's region is also the readable part of src
    return launch_hpel(&a, 1, static_cast<cudaStream_t>(stream));
}

int gvc_me_lowres(const uint8_t *cur_low, const uint8_t *ref_low, int stride, int mbw, int mbh, int range, int lambda,
                  int16_t *out_mv, void *stream) {
    int rc = enter();
    if (rc) return rc;
    if (!cur_low || !ref_low || !out_mv || mbw < 1 || mbh < 1) { set_error("gvc_me_lowres: bad arguments"); return GVC_ERR_INVALID_ARG; }
    if (range < 0 || range > 22) { set_error("gvc_me_lowres: lowres range %d outside 0..22", range); return GVC_ERR_INVALID_ARG; }
    MeLowArgs a;
    a.cur = cur_low; a.ref = ref_low; a.stride = stride; a.mbw = mbw; a.mbh = mbh; a.r = range; a.lambda = lambda; a.out_mv = out_mv;
    return launch_me_lowres(&a, 1, static_cast<cudaStream_t>(stream));
}

static RefPlanes to_ref(const gvc_ref_planes *r) {
    RefPlanes o;
    o.y = r->y; o.hh = r->h; o.hv = r->v; o.hc = r->c; o.u = r->cb; o.v = r->cr; o.low = r->low;
    return o;
}

int gvc_me_refine(const uint8_t *cur_y, const gvc_ref_planes *ref, int stride, int w16, int h16, const int16_t *lowres_mv,
                  int lambda, gvc_mb_motion *out, void *stream) {
    int rc = enter();
    if (rc) return rc;
    if (!cur_y || !ref || !ref->y || !ref->h || !ref->v || !ref->c || !lowres_mv || !out || (w16 & 15) || (h16 & 15) || w16 < 16 || h16 < 16) {
        set_error("gvc_me_refine: bad arguments");
        return GVC_ERR_INVALID_ARG;
    }
    MeRefArgs a;
    a.cur = cur_y; a.ref = to_ref(ref); a.stride = stride; a.w16 = w16; a.h16 = h16; a.mbw = w16 / 16; a.mbh = h16 / 16;
    a.lambda = lambda; a.lowres_mv = lowres_mv; a.out = reinterpret_cast<MbMotion *>(out); a.out_host = nullptr;
    return launch_me_refine(&a, 1, static_cast<cudaStream_t>(stream));
}

int gvc_encode_inter(const uint8_t *cur_y, const uint8_t *cur_u, const uint8_t *cur_v, const gvc_ref_planes *ref, uint8_t *rec_y,
                     uint8_t *rec_u, uint8_t *rec_v, int sy, int sc, int w16, int h16, const gvc_mb_motion *motion, int qp,
                     int16_t *coefs, uint32_t *nzmask, void *stream) {
    int rc = enter();
    if (rc) return rc;
    if (!cur_y || !cur_u || !cur_v || !ref || !rec_y || !rec_u || !rec_v || !motion || !coefs || !nzmask || (w16 & 15) || (h16 & 15)) {
        set_error("gvc_encode_inter: bad arguments");
        return GVC_ERR_INVALID_ARG;
    }
    if (qp < 0 || qp > 51) { set_error("gvc_encode_inter: qp %d outside 0..51", qp); return GVC_ERR_INVALID_ARG; }
    EncMbArgs a;
    a.cur_y = cur_y; a.cur_u = cur_u; a.cur_v = cur_v; a.ref = to_ref(ref);
    a.rec_y = rec_y; a.rec_u = rec_u; a.rec_v = rec_v; a.sy = sy; a.sc = sc; a.mbw = w16 / 16; a.mbh = h16 / 16;
    fill_quant_params(a.q, qp, false);
    a.motion = reinterpret_cast<const MbMotion *>(motion); a.coefs = coefs; a.nzmask = nzmask;
    a.pack.blocks = nullptr; a.pack.offsets = nullptr; a.pack.counter = nullptr; a.nnz_y = a.nnz_u = a.nnz_v = nullptr;
    return launch_encode_inter(&a, 1, static_cast<cudaStream_t>(stream));
}

int gvc_encode_intra(const uint8_t *cur_y, const uint8_t *cur_u, const uint8_t *cur_v, uint8_t *rec_y, uint8_t *rec_u, uint8_t *rec_v,
                     int sy, int sc, int w16, int h16, int qp, int16_t *coefs, uint32_t *nzmask, uint8_t *modes, void *scratch,
                     void *stream) {
    int rc = enter();
    if (rc) return rc;
    if (!cur_y || !cur_u || !cur_v || !rec_y || !rec_u || !rec_v || !coefs || !nzmask || !modes || !scratch || (w16 & 15) || (h16 & 15) ||
        w16 < 16 || h16 < 16) {
        set_error("gvc_encode_intra: bad arguments");
        return GVC_ERR_INVALID_ARG;
    }
    if (qp < 0 || qp > 51) { set_error("gvc_encode_intra: qp %d outside 0..51", qp); return GVC_ERR_INVALID_ARG; }
    if (h16 / 16 > 1024) { set_error("gvc_encode_intra: more than 1024 macroblock rows"); return GVC_ERR_TOO_LARGE; }
    IntraArgs a;
    a.cur_y = cur_y; a.cur_u = cur_u; a.cur_v = cur_v; a.rec_y = rec_y; a.rec_u = rec_u; a.rec_v = rec_v;
    a.sy = sy; a.sc = sc; a.mbw = w16 / 16; a.mbh = h16 / 16;
    fill_quant_params(a.q, qp, true);
    a.coefs = coefs; a.nzmask = nzmask; a.modes = modes; a.mail = static_cast<unsigned long long *>(scratch);
    a.pack.blocks = nullptr; a.pack.offsets = nullptr; a.pack.counter = nullptr; a.nnz_y = a.nnz_u = a.nnz_v = nullptr;
    return launch_encode_intra(&a, 1, static_cast<cudaStream_t>(stream));
}

int gvc_transform8x8(const uint8_t *src, int sstride, const uint8_t *pred, int pstride, uint8_t *rec, int rstride, int w, int h, int qp,
                     int intra, int16_t *coefs, uint8_t *nz, void *stream) {
    int rc = enter();
    if (rc) return rc;
    if (!src || !pred || !rec || !coefs || !nz || w < 8 || h < 8 || (w & 7) || (h & 7)) { set_error("gvc_transform8x8: bad arguments"); return GVC_ERR_INVALID_ARG; }
    if (qp < 0 || qp > 51) { set_error("gvc_transform8x8: qp %d outside 0..51", qp); return GVC_ERR_INVALID_ARG; }
    Transform8Args a;
    a.src = src; a.pred = pred; a.rec = rec; a.sstride = sstride; a.pstride = pstride; a.rstride = rstride; a.bw = w / 8; a.bh = h / 8;
    fill_quant8_params(a.q, qp, intra != 0);
    a.coefs = coefs; a.nz = nz;
    transform8x8_kernel<<<(a.bw * a.bh + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(a);
    GVC_CUDA(cudaGetLastError());
    return GVC_OK;
}

int gvc_lambda_for_qp(int qp) { return lambda_for_qp(qp); }

}  // extern "C"
