// gvc_resize.cu -- K1 resize: plan construction, kernel launches, C ABI (include/gvc.h "resize").
//
// Replaces libswscale on the reference's path (client.go:105-115: args such as `-s 1280x720` are
// spliced into the ffmpeg command line, which inserts the scaler).  See resize_kernels.cuh for the
// arithmetic contract and the kernel design.
#include <algorithm>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include "filter_bank.h"
#include "gvc_common.h"
#include "resize_kernels.cuh"
#include "resize_plan.h"
#include "tma.cuh"

namespace gvc {

// ------------------------------------------------------------------------------- kernels ------
// The three planes' tensor maps (3-D: bytes of a row as 16-bit elements, rows, frames) travel as kernel parameters; on = 0 for a
// plane TMA cannot address (NV12 chroma, unaligned base / pitch): its tiles are then filled with plain loads.
struct ResizeTma { CUtensorMap map[3]; int on[3]; };

template <int NG, int VT>
__global__ void __launch_bounds__(kThreads) resize_fused_kernel(const __grid_constant__ ResizeParams P, const __grid_constant__ ResizeTma M, int s_chunks_max) {
    extern __shared__ __align__(128) uint4 smem[];
    __shared__ unsigned long long bar;
    const int            tile = blockIdx.x;
    const int            j    = find_job(P, tile);
    const PlaneJob      &pj   = P.job[j];
    const PlaneClassDev &pc   = P.cls[pj.cls];
    const int            t    = tile - pj.tile_base;
    const int            ty = pc.tiles_x_magic ? (int)(((unsigned)t * (unsigned)pc.tiles_x_magic) >> 20) : t / pc.tiles_x, tx = t - ty * pc.tiles_x;
    const TileGeom       g   = tile_geom(pc, tx, ty);
    const uint8_t       *src = pj.src + (size_t)blockIdx.y * pj.sframe;
    uint8_t             *dst = pj.dst + (size_t)blockIdx.y * pj.dframe;
    unsigned            *S   = reinterpret_cast<unsigned *>(smem);
    int4                *T   = reinterpret_cast<int4 *>(smem + kTileRows * s_chunks_max);
    const int            pitch_w = 4 * s_chunks_max;

    if (M.on[j]) {      // the whole source tile is one box: 16 * s_chunks_max bytes x kTileRows rows of frame blockIdx.y
        if (threadIdx.x == 0) {
            mbar_init(&bar, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            mbar_expect_tx(&bar, (unsigned)(16 * s_chunks_max * kTileRows));
            tma_box_3d(S, &M.map[j], 2 * g.qa0, g.r0, (int)blockIdx.y, &bar);      // x in 16-bit elements: byte 4 * qa0
        }
        __syncthreads();
        mbar_wait(&bar, 0);
    } else {
        resize_phase0(threadIdx.x, pc, pj, src, g, S, pitch_w);
        __syncthreads();
    }
    resize_phase1<NG>(threadIdx.x, pc, g, S, pitch_w, T);
    __syncthreads();
    resize_phase2<VT>(threadIdx.x, pc, pj, dst, g, T);
}

struct GenericJob {
    const uint8_t *src; int sstride; int spix; unsigned long long sframe;
    uint8_t       *dst; int dstride; int dpix; unsigned long long dframe;
    int cls; unsigned long long scratch_off;  // int16 elements, within one frame's scratch
};
struct GenericParams {
    GenericClassDev cls[2];
    GenericJob      job[3];
    int             njobs;
    unsigned long long scratch_per_frame;      // int16 elements
};

// grid: (ceil(dw/256), sh, frames*njobs)
__global__ void __launch_bounds__(256) resize_hpass_kernel(const __grid_constant__ GenericParams P, int16_t *scratch) {
    const int frame = blockIdx.z / P.njobs, j = blockIdx.z - frame * P.njobs;
    const GenericJob      &pj = P.job[j];
    const GenericClassDev &pc = P.cls[pj.cls];
    const int x = blockIdx.x * 256 + threadIdx.x, r = blockIdx.y;
    if (x >= pc.dw || r >= pc.sh) return;
    const uint8_t *row = pj.src + (size_t)frame * pj.sframe + (size_t)r * pj.sstride;
    int16_t *tmp = scratch + (size_t)frame * P.scratch_per_frame + pj.scratch_off;
    tmp[(size_t)r * pc.dw + x] = (int16_t)generic_h_sample(pc, row, pj.spix, x);
}

// grid: (ceil(dw/256), dh, frames*njobs)
__global__ void __launch_bounds__(256) resize_vpass_kernel(const __grid_constant__ GenericParams P, const int16_t *scratch) {
    const int frame = blockIdx.z / P.njobs, j = blockIdx.z - frame * P.njobs;
    const GenericJob      &pj = P.job[j];
    const GenericClassDev &pc = P.cls[pj.cls];
    const int x = blockIdx.x * 256 + threadIdx.x, y = blockIdx.y;
    if (x >= pc.dw || y >= pc.dh) return;
    const int16_t *tmp = scratch + (size_t)frame * P.scratch_per_frame + pj.scratch_off;
    uint8_t *d = pj.dst + (size_t)frame * pj.dframe + (size_t)y * pj.dstride;
    d[(size_t)x * pj.dpix] = (uint8_t)generic_v_sample(pc, tmp, y, x);
}

}  // namespace gvc

// ------------------------------------------------------------------------------- plan ---------
using namespace gvc;

struct gvc_resize_plan {
    int fmt = 0;
    ResizeTables t;
    bool force_two_pass = false;
    std::vector<void *> dev_allocs;        // device tables
    // host-path staging: kSlots independent pipelines
    static constexpr int kSlots = 3;
    cudaStream_t stream[kSlots] = {};
    uint8_t *stage_src[kSlots] = {}, *stage_dst[kSlots] = {};
    size_t   stage_src_cap[kSlots] = {}, stage_dst_cap[kSlots] = {};
    // two-pass intermediate (int16): one per host-path slot, last entry for gvc_resize_frames_device callers
    int16_t *scratch[kSlots + 1] = {}; size_t scratch_elems[kSlots + 1] = {};
};

namespace {

int plane_w(int w, int k) { return k == 0 ? w : chroma_dim(w); }
int plane_h(int h, int k) { return k == 0 ? h : chroma_dim(h); }

template <typename T>
int upload(gvc_resize_plan *p, const std::vector<T> &v, const T **out) {
    void *d = nullptr;
    GVC_CUDA(cudaMalloc(&d, std::max<size_t>(v.size() * sizeof(T), 16)));
    p->dev_allocs.push_back(d);
    GVC_CUDA(cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    *out = static_cast<const T *>(d);
    return GVC_OK;
}

int upload_tables(gvc_resize_plan *p) {
    ResizeTables &t = p->t;
    int rc;
    for (int c = 0; c < 2; ++c) {
        GenericClassDev &g = t.gcls[c];
        if ((rc = upload(p, t.hb[c].coef, &g.hc))) return rc;
        if ((rc = upload(p, t.hb[c].pos, &g.hp))) return rc;
        if ((rc = upload(p, t.vb[c].coef, &g.vc))) return rc;
        if ((rc = upload(p, t.vb[c].pos, &g.vp))) return rc;
        if (t.fused_ok) {
            if ((rc = upload(p, t.htab[c], &t.cls[c].htab))) return rc;
            if ((rc = upload(p, t.vtab[c], &t.cls[c].vtab))) return rc;
        }
    }
    return GVC_OK;
}

typedef void (*fused_fn)(const ResizeParams, const ResizeTma, int);
template <int NG, int VT>
struct PickVt {
    static fused_fn get(int vt) { return vt >= VT ? static_cast<fused_fn>(resize_fused_kernel<NG, VT>) : PickVt<NG, VT - 1>::get(vt); }
};
template <int NG>
struct PickVt<NG, 1> {
    static fused_fn get(int) { return resize_fused_kernel<NG, 1>; }
};
fused_fn pick_fused(int nt, int vt) {
    switch (nt) {
        case 1: return PickVt<1, kMaxVPairs>::get(vt);
        case 2: return PickVt<2, kMaxVPairs>::get(vt);
        case 3: return PickVt<3, kMaxVPairs>::get(vt);
        case 4: return PickVt<4, kMaxVPairs>::get(vt);
        case 5: return PickVt<5, kMaxVPairs>::get(vt);
        default: return PickVt<6, kMaxVPairs>::get(vt);
    }
}

int check_frames(const gvc_resize_plan *p, const gvc_frames *f, int w, int h, const char *what) {
    if (!f) { set_error("%s is NULL", what); return GVC_ERR_INVALID_ARG; }
    const int np = p->fmt == GVC_PIX_NV12 ? 2 : 3;
    for (int k = 0; k < np; ++k) {
        const int pw = p->fmt == GVC_PIX_NV12 && k == 1 ? 2 * plane_w(w, 1) : plane_w(w, k);
        if (!f->plane[k] || f->stride[k] < pw) {
            set_error("%s plane %d: null pointer or stride %d < row bytes %d", what, k, f->stride[k], pw);
            return GVC_ERR_INVALID_ARG;
        }
        (void)h;
    }
    return GVC_OK;
}

// Tensor maps of the source planes of `nf` frames: plane k as a 3-D tensor of 16-bit elements (a row's bytes in pairs, so a
// box may be up to 512 bytes wide), rows, frames.  A box = one tile's source rows; what lies outside the plane reads as zeros.
void resize_tma(const ResizeParams &P, int nf, int s_chunks_max, ResizeTma &M) {
    memset(&M, 0, sizeof(M));
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc || 16 * s_chunks_max > 512) return;
    for (int k = 0; k < P.njobs; ++k) {
        const PlaneJob      &j  = P.job[k];
        const PlaneClassDev &pc = P.cls[j.cls];
        unsigned long long fstride = j.sframe;
        if (nf == 1 && (fstride & 15)) fstride = ((unsigned long long)j.sstride * pc.sh + 15) & ~15ull;      // never used as an offset
        if (j.spix != 1 || (reinterpret_cast<uintptr_t>(j.src) & 15) || (j.sstride & 15) || (fstride & 15) || fstride == 0) continue;
        const cuuint64_t dims[3] = {(cuuint64_t)((pc.sw + 1) / 2), (cuuint64_t)pc.sh, (cuuint64_t)nf};
        const cuuint64_t strides[2] = {(cuuint64_t)j.sstride, (cuuint64_t)fstride};
        const cuuint32_t box[3] = {(cuuint32_t)(8 * s_chunks_max), (cuuint32_t)kTileRows, 1}, es[3] = {1, 1, 1};
        if (enc(&M.map[k], CU_TENSOR_MAP_DATA_TYPE_UINT16, 3, const_cast<uint8_t *>(j.src), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS)
            M.on[k] = 1;
    }
}

int launch_device(const gvc_resize_plan *p, const gvc_frames *src, const gvc_frames *dst, int n_frames, cudaStream_t st, int sslot) {
    const bool nv12 = p->fmt == GVC_PIX_NV12;
    if (p->t.fused_ok && !p->force_two_pass) {
        ResizeParams P;
        memset(&P, 0, sizeof(P));
        P.cls[0] = p->t.cls[0]; P.cls[1] = p->t.cls[1];
        P.njobs = 3;
        for (int k = 0; k < 3; ++k) {
            PlaneJob &j = P.job[k];
            const int sk = nv12 && k == 2 ? 1 : k;       // NV12: jobs 1,2 read/write the interleaved plane
            j.src = src->plane[sk] + (nv12 && k == 2 ? 1 : 0);
            j.dst = dst->plane[sk] + (nv12 && k == 2 ? 1 : 0);
            j.sstride = src->stride[sk]; j.dstride = dst->stride[sk];
            j.sframe = src->frame_stride[sk]; j.dframe = dst->frame_stride[sk];
            j.spix = j.dpix = nv12 && k > 0 ? 2 : 1;
            j.cls = k == 0 ? 0 : 1;
        }
        // tile ids: luma first (heaviest), then the two chroma planes
        const int lt = p->t.cls[0].tiles_x * p->t.cls[0].tiles_y, ct = p->t.cls[1].tiles_x * p->t.cls[1].tiles_y;
        P.job[0].tile_base = 0; P.job[1].tile_base = lt; P.job[2].tile_base = lt + ct;
        P.tiles_per_frame = lt + 2 * ct;
        fused_fn fn = pick_fused(p->t.nt, p->t.vt);
        GVC_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p->t.smem_bytes));
        for (int f0 = 0; f0 < n_frames; f0 += 32768) {
            const int nf = std::min(32768, n_frames - f0);
            ResizeParams Q = P;
            for (int k = 0; k < 3; ++k) { Q.job[k].src += (size_t)f0 * Q.job[k].sframe; Q.job[k].dst += (size_t)f0 * Q.job[k].dframe; }
            ResizeTma M;
            resize_tma(Q, nf, p->t.s_chunks_max, M);
            fn<<<dim3(P.tiles_per_frame, nf), kThreads, p->t.smem_bytes, st>>>(Q, M, p->t.s_chunks_max);
            GVC_CUDA(cudaGetLastError());
        }
        return GVC_OK;
    }
    // ---- two-pass path
    GenericParams G;
    memset(&G, 0, sizeof(G));
    G.cls[0] = p->t.gcls[0]; G.cls[1] = p->t.gcls[1]; G.njobs = 3;
    unsigned long long off = 0;
    for (int k = 0; k < 3; ++k) {
        GenericJob &j = G.job[k];
        const int sk = nv12 && k == 2 ? 1 : k;
        j.src = src->plane[sk] + (nv12 && k == 2 ? 1 : 0);
        j.dst = dst->plane[sk] + (nv12 && k == 2 ? 1 : 0);
        j.sstride = src->stride[sk]; j.dstride = dst->stride[sk];
        j.sframe = src->frame_stride[sk]; j.dframe = dst->frame_stride[sk];
        j.spix = j.dpix = nv12 && k > 0 ? 2 : 1;
        j.cls = k == 0 ? 0 : 1;
        j.scratch_off = off;
        off += (unsigned long long)G.cls[j.cls].sh * G.cls[j.cls].dw;
    }
    G.scratch_per_frame = off;
    gvc_resize_plan *mp = const_cast<gvc_resize_plan *>(p);   // scratch is a cache, not plan state
    const size_t budget = size_t(256) << 20;                  // bytes of scratch per launch group
    int chunk = (int)std::max<size_t>(1, std::min<size_t>(n_frames, budget / (off * 2)));
    chunk = std::min(chunk, 65535 / 3);
    if (mp->scratch_elems[sslot] < (size_t)chunk * off) {
        if (mp->scratch[sslot]) cudaFree(mp->scratch[sslot]);
        mp->scratch[sslot] = nullptr; mp->scratch_elems[sslot] = 0;
        GVC_CUDA(cudaMalloc((void **)&mp->scratch[sslot], (size_t)chunk * off * 2));
        mp->scratch_elems[sslot] = (size_t)chunk * off;
    }
    for (int f0 = 0; f0 < n_frames; f0 += chunk) {
        const int nf = std::min(chunk, n_frames - f0);
        GenericParams Q = G;
        for (int k = 0; k < 3; ++k) { Q.job[k].src += (size_t)f0 * Q.job[k].sframe; Q.job[k].dst += (size_t)f0 * Q.job[k].dframe; }
        const int mdw = G.cls[0].dw, msh = G.cls[0].sh, mdh = G.cls[0].dh;
        resize_hpass_kernel<<<dim3((mdw + 255) / 256, msh, nf * 3), 256, 0, st>>>(Q, mp->scratch[sslot]);
        GVC_CUDA(cudaGetLastError());
        resize_vpass_kernel<<<dim3((mdw + 255) / 256, mdh, nf * 3), 256, 0, st>>>(Q, mp->scratch[sslot]);
        GVC_CUDA(cudaGetLastError());
    }
    return GVC_OK;
}

struct Range { const uint8_t *lo, *hi; };

// Byte ranges the frames [f0, f0+nf) of `f` occupy, overlapping/adjacent ranges merged.
int frame_ranges(const gvc_resize_plan *p, const gvc_frames *f, int w, int h, int f0, int nf, Range out[3]) {
    const int np = p->fmt == GVC_PIX_NV12 ? 2 : 3;
    Range r[3];
    for (int k = 0; k < np; ++k) {
        const int pw = p->fmt == GVC_PIX_NV12 && k == 1 ? 2 * plane_w(w, 1) : plane_w(w, k);
        const int ph = plane_h(h, k);
        r[k].lo = f->plane[k] + (size_t)f0 * f->frame_stride[k];
        r[k].hi = r[k].lo + (size_t)(nf - 1) * f->frame_stride[k] + (size_t)(ph - 1) * f->stride[k] + pw;
    }
    std::sort(r, r + np, [](const Range &a, const Range &b) { return a.lo < b.lo; });
    int n = 0;
    for (int k = 0; k < np; ++k) {
        if (n && r[k].lo <= out[n - 1].hi) out[n - 1].hi = std::max(out[n - 1].hi, r[k].hi);
        else out[n++] = r[k];
    }
    return n;
}

}  // namespace

// ------------------------------------------------------------------------------- C ABI --------
extern "C" {

int gvc_resize_plan_create(gvc_resize_plan **plan, int sw, int sh, int dw, int dh, int pix_fmt, int algo) {
    if (!plan) { set_error("plan is NULL"); return GVC_ERR_INVALID_ARG; }
    *plan = nullptr;
    int rc = enter();
    if (rc) return rc;
    if (sw < 1 || sh < 1 || dw < 1 || dh < 1 || sw > 16384 || sh > 16384 || dw > 16384 || dh > 16384) {
        set_error("bad geometry %dx%d -> %dx%d", sw, sh, dw, dh);
        return GVC_ERR_INVALID_ARG;
    }
    if (pix_fmt != GVC_PIX_YUV420P && pix_fmt != GVC_PIX_NV12) { set_error("pix_fmt %d unsupported", pix_fmt); return GVC_ERR_UNSUPPORTED; }
    if (algo != GVC_SCALE_BICUBIC && algo != GVC_SCALE_BILINEAR) { set_error("scaler %d unsupported", algo); return GVC_ERR_UNSUPPORTED; }
    gvc_resize_plan *p = new (std::nothrow) gvc_resize_plan;
    if (!p) return GVC_ERR_NOMEM;
    try {
        p->fmt = pix_fmt;
        if (!build_resize_tables(p->t, sw, sh, dw, dh, algo)) {
            set_error("filter construction failed for %dx%d -> %dx%d", sw, sh, dw, dh);
            rc = GVC_ERR_INVALID_ARG;
        } else {
            rc = upload_tables(p);
        }
    } catch (const std::bad_alloc &) {
        rc = GVC_ERR_NOMEM;
    }
    if (rc) { gvc_resize_plan_destroy(p); return rc; }
    *plan = p;
    return GVC_OK;
}

void gvc_resize_plan_destroy(gvc_resize_plan *p) {
    if (!p) return;
    if (gvc::initialised()) cudaSetDevice(gvc::current_device());
    for (void *d : p->dev_allocs) cudaFree(d);
    for (int s = 0; s <= gvc_resize_plan::kSlots; ++s)
        if (p->scratch[s]) cudaFree(p->scratch[s]);
    for (int s = 0; s < gvc_resize_plan::kSlots; ++s) {
        if (p->stage_src[s]) cudaFree(p->stage_src[s]);
        if (p->stage_dst[s]) cudaFree(p->stage_dst[s]);
        if (p->stream[s]) cudaStreamDestroy(p->stream[s]);
    }
    delete p;
}

int gvc_resize_plan_info(const gvc_resize_plan *p, int *fused, int *h_taps_luma, int *v_taps_luma) {
    if (!p) { set_error("plan is NULL"); return GVC_ERR_INVALID_ARG; }
    if (fused) *fused = p->t.fused_ok && !p->force_two_pass;
    if (h_taps_luma) *h_taps_luma = p->t.hb[0].taps;
    if (v_taps_luma) *v_taps_luma = p->t.vb[0].taps;
    return GVC_OK;
}

int gvc_resize_plan_force_two_pass(gvc_resize_plan *p, int on) {
    if (!p) { set_error("plan is NULL"); return GVC_ERR_INVALID_ARG; }
    p->force_two_pass = on != 0;
    return GVC_OK;
}

int gvc_resize_frames_device(const gvc_resize_plan *p, const gvc_frames *src, const gvc_frames *dst, int n_frames, void *stream) {
    int rc = enter();
    if (rc) return rc;
    if (!p) { set_error("plan is NULL"); return GVC_ERR_INVALID_ARG; }
    if (n_frames < 0) { set_error("n_frames < 0"); return GVC_ERR_INVALID_ARG; }
    if (n_frames == 0) return GVC_OK;
    if ((rc = check_frames(p, src, p->t.sw, p->t.sh, "src"))) return rc;
    if ((rc = check_frames(p, dst, p->t.dw, p->t.dh, "dst"))) return rc;
    return launch_device(p, src, dst, n_frames, static_cast<cudaStream_t>(stream), gvc_resize_plan::kSlots);
}

int gvc_resize_frames_host(gvc_resize_plan *p, const gvc_frames *src, const gvc_frames *dst, int n_frames) {
    int rc = enter();
    if (rc) return rc;
    if (!p) { set_error("plan is NULL"); return GVC_ERR_INVALID_ARG; }
    if (n_frames < 0) { set_error("n_frames < 0"); return GVC_ERR_INVALID_ARG; }
    if (n_frames == 0) return GVC_OK;
    if ((rc = check_frames(p, src, p->t.sw, p->t.sh, "src"))) return rc;
    if ((rc = check_frames(p, dst, p->t.dw, p->t.dh, "dst"))) return rc;
    constexpr int kSlots = gvc_resize_plan::kSlots;
    for (int s = 0; s < kSlots; ++s)
        if (!p->stream[s]) GVC_CUDA(cudaStreamCreateWithFlags(&p->stream[s], cudaStreamNonBlocking));
    const int np = p->fmt == GVC_PIX_NV12 ? 2 : 3;
    const int chunk = std::max(1, std::min(n_frames, 16));
    int slot = 0;
    for (int f0 = 0; f0 < n_frames; f0 += chunk, slot = (slot + 1) % kSlots) {
        const int nf = std::min(chunk, n_frames - f0);
        cudaStream_t st = p->stream[slot];
        Range sr[3], dr[3];
        const int nsr = frame_ranges(p, src, p->t.sw, p->t.sh, f0, nf, sr);
        const int ndr = frame_ranges(p, dst, p->t.dw, p->t.dh, f0, nf, dr);
        size_t sbytes = 0, dbytes = 0, soff[3], doff[3];
        for (int i = 0; i < nsr; ++i) { soff[i] = sbytes; sbytes += ((size_t)(sr[i].hi - sr[i].lo) + 255) & ~size_t(255); }
        for (int i = 0; i < ndr; ++i) { doff[i] = dbytes; dbytes += ((size_t)(dr[i].hi - dr[i].lo) + 255) & ~size_t(255); }
        GVC_CUDA(cudaStreamSynchronize(st));   // slot reuse: previous chunk on this slot must have drained
        if (p->stage_src_cap[slot] < sbytes) {
            if (p->stage_src[slot]) cudaFree(p->stage_src[slot]);
            p->stage_src[slot] = nullptr; p->stage_src_cap[slot] = 0;
            GVC_CUDA(cudaMalloc((void **)&p->stage_src[slot], sbytes));
            p->stage_src_cap[slot] = sbytes;
        }
        if (p->stage_dst_cap[slot] < dbytes) {
            if (p->stage_dst[slot]) cudaFree(p->stage_dst[slot]);
            p->stage_dst[slot] = nullptr; p->stage_dst_cap[slot] = 0;
            GVC_CUDA(cudaMalloc((void **)&p->stage_dst[slot], dbytes));
            p->stage_dst_cap[slot] = dbytes;
        }
        for (int i = 0; i < nsr; ++i)
            GVC_CUDA(cudaMemcpyAsync(p->stage_src[slot] + soff[i], sr[i].lo, sr[i].hi - sr[i].lo, cudaMemcpyHostToDevice, st));
        gvc_frames ds = *src, dd = *dst;
        for (int k = 0; k < np; ++k) {
            const uint8_t *hp = src->plane[k] + (size_t)f0 * src->frame_stride[k];
            for (int i = 0; i < nsr; ++i)
                if (hp >= sr[i].lo && hp < sr[i].hi) ds.plane[k] = p->stage_src[slot] + soff[i] + (hp - sr[i].lo);
            const uint8_t *hq = dst->plane[k] + (size_t)f0 * dst->frame_stride[k];
            for (int i = 0; i < ndr; ++i)
                if (hq >= dr[i].lo && hq < dr[i].hi) dd.plane[k] = p->stage_dst[slot] + doff[i] + (hq - dr[i].lo);
        }
        // A merged destination range may contain bytes that are not ours (row padding, gaps between
        // planes): upload the range first so the copy back leaves them as the caller had them.
        size_t payload = 0, span = 0;
        for (int k = 0; k < np; ++k) {
            const int pw = p->fmt == GVC_PIX_NV12 && k == 1 ? 2 * plane_w(p->t.dw, 1) : plane_w(p->t.dw, k);
            payload += (size_t)pw * plane_h(p->t.dh, k) * nf;
        }
        for (int i = 0; i < ndr; ++i) span += dr[i].hi - dr[i].lo;
        if (span != payload)
            for (int i = 0; i < ndr; ++i)
                GVC_CUDA(cudaMemcpyAsync(p->stage_dst[slot] + doff[i], dr[i].lo, dr[i].hi - dr[i].lo, cudaMemcpyHostToDevice, st));
        if ((rc = launch_device(p, &ds, &dd, nf, st, slot))) return rc;
        for (int i = 0; i < ndr; ++i)
            GVC_CUDA(cudaMemcpyAsync(const_cast<uint8_t *>(dr[i].lo), p->stage_dst[slot] + doff[i], dr[i].hi - dr[i].lo, cudaMemcpyDeviceToHost, st));
    }
    for (int s = 0; s < kSlots; ++s) GVC_CUDA(cudaStreamSynchronize(p->stream[s]));
    return GVC_OK;
}

}  // extern "C"

// ------------------------------------------------------------------ allocation-free variant ----
// gvc_resize_frames_device_static(): the two-pass resize with its int16 intermediate plane held in a
// fixed array that is part of the library image instead of a buffer obtained from cudaMalloc.
//
// Who this is for.  gvc_resize_frames_device() may call cudaMalloc the first time a plan takes the
// two-pass path (or when a bigger batch arrives).  Two kinds of callers cannot live with that:
//   * a worker in its steady state, with copy streams of the previous and next segment in flight:
//     cudaMalloc/cudaFree synchronise the whole device, so a lazy allocation stalls every stream;
//   * a deployment where the GPU's memory is budgeted up front (several workers per GPU under a
//     memory cap, or frame pools sized to fill HBM): a late allocation can fail after the job was
//     accepted, and the reference dispatcher never re-queues a failed piece (server.go:276-283).
// Such callers create their plans and buffers during start-up and then use this entry point, which
// performs no allocation of any kind: no cudaMalloc/cudaFree, no allocating library call.
//
// The price is a fixed ceiling.  The array below occupies kStaticScratchBytes of device memory on
// the GPU gvc_init() selects, from gvc_init() until the process exits, whether or not this entry
// point is ever called.  (CUDA loads a module's device image lazily, at the first use of anything
// in it; gvc_init() asks for the array's address so that this happens -- and the 16 MiB are
// reserved, or found to be unavailable -- during start-up and not inside the first steady-state
// call.)  A frame needs 2 * (src_h * dst_w + 2 * ceil(src_h/2) * ceil(dst_w/2)) bytes of it (4.1 MB
// for 1080p->720p, 12.4 MB for 2160p->1080p); frames are processed in groups that fit, and a plan
// whose single frame does not fit is rejected with GVC_ERR_TOO_LARGE.
namespace {

constexpr size_t kStaticScratchBytes = size_t(16) << 20;   // 16 MiB, the stated maximum
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

// In host code `g_static_scratch` is only a handle for the symbol, not an address the GPU can use.
// The device address is looked up once per device (with that device current) and remembered here.
constexpr int kMaxDevices = 64;
int16_t   *g_static_scratch_addr[kMaxDevices] = {};
std::mutex g_static_scratch_mutex;

int static_scratch_address(int dev, int16_t **out) {     // `dev` must be the current device
    if (dev < 0 || dev >= kMaxDevices) { set_error("device ordinal %d outside the supported range 0..%d", dev, kMaxDevices - 1); return GVC_ERR_INVALID_ARG; }
    std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
    if (!g_static_scratch_addr[dev]) {
        void *p = nullptr;
        GVC_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
        // the two kernels the entry point launches are loaded now as well (lazy loading would otherwise do it, with
        // whatever bookkeeping memory that takes, inside the first call)
        cudaFuncAttributes fa;
        GVC_CUDA(cudaFuncGetAttributes(&fa, gvc::resize_hpass_kernel));
        GVC_CUDA(cudaFuncGetAttributes(&fa, gvc::resize_vpass_kernel));
        g_static_scratch_addr[dev] = static_cast<int16_t *>(p);
    }
    if (out) *out = g_static_scratch_addr[dev];
    return GVC_OK;
}

}  // namespace

// Called by gvc_init() with the selected device current: loads this file's device image (which reserves the array)
// and remembers the array's address, so gvc_resize_frames_device_static() itself never triggers either.
int gvc::resize_static_prepare(int dev) { return static_scratch_address(dev, nullptr); }

extern "C" int gvc_resize_frames_device_static(const gvc_resize_plan *p, const gvc_frames *src, const gvc_frames *dst,
                                               int n_frames, void *stream) {
    int rc = enter();
    if (rc) return rc;
    if (!p) { set_error("plan is NULL"); return GVC_ERR_INVALID_ARG; }
    if (n_frames < 0) { set_error("n_frames < 0"); return GVC_ERR_INVALID_ARG; }
    if ((rc = check_frames(p, src, p->t.sw, p->t.sh, "src"))) return rc;
    if ((rc = check_frames(p, dst, p->t.dw, p->t.dh, "dst"))) return rc;

    const bool nv12 = p->fmt == GVC_PIX_NV12;
    GenericParams G;
    memset(&G, 0, sizeof(G));
    G.cls[0] = p->t.gcls[0]; G.cls[1] = p->t.gcls[1]; G.njobs = 3;
    unsigned long long elems = 0;   // int16 elements of intermediate per frame
    for (int k = 0; k < 3; ++k) {
        GenericJob &j = G.job[k];
        const int sk = nv12 && k == 2 ? 1 : k;
        j.src = src->plane[sk] + (nv12 && k == 2 ? 1 : 0);
        j.dst = dst->plane[sk] + (nv12 && k == 2 ? 1 : 0);
        j.sstride = src->stride[sk]; j.dstride = dst->stride[sk];
        j.sframe = src->frame_stride[sk]; j.dframe = dst->frame_stride[sk];
        j.spix = j.dpix = nv12 && k > 0 ? 2 : 1;
        j.cls = k == 0 ? 0 : 1;
        j.scratch_off = elems;
        elems += (unsigned long long)G.cls[j.cls].sh * G.cls[j.cls].dw;
    }
    G.scratch_per_frame = elems;
    const size_t capacity = kStaticScratchBytes / sizeof(int16_t);
    if (elems > capacity) {
        set_error("one %dx%d -> %dx%d frame needs %llu bytes of intermediate; the built-in scratch of "
                  "gvc_resize_frames_device_static holds %zu bytes (use gvc_resize_frames_device instead)",
                  p->t.sw, p->t.sh, p->t.dw, p->t.dh, elems * 2ull, kStaticScratchBytes);
        return GVC_ERR_TOO_LARGE;
    }
    if (n_frames == 0) return GVC_OK;
    int16_t *scratch = nullptr;      // resolved by gvc_init(); the lookup below only reads the remembered address
    if ((rc = static_scratch_address(gvc::current_device(), &scratch))) return rc;

    // Nothing in the scratch has to start at zero: the horizontal pass writes every element the
    // vertical pass reads (all src_h x dst_w samples of each plane), so whatever an earlier call left
    // behind is overwritten before it can be observed.
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int group = (int)std::min<unsigned long long>({(unsigned long long)n_frames, capacity / elems, 65535ull / 3});
    for (int f0 = 0; f0 < n_frames; f0 += group) {
        const int nf = std::min(group, n_frames - f0);
        GenericParams Q = G;
        for (int k = 0; k < 3; ++k) { Q.job[k].src += (size_t)f0 * Q.job[k].sframe; Q.job[k].dst += (size_t)f0 * Q.job[k].dframe; }
        const int mdw = G.cls[0].dw, msh = G.cls[0].sh, mdh = G.cls[0].dh;
        resize_hpass_kernel<<<dim3((mdw + 255) / 256, msh, nf * 3), 256, 0, st>>>(Q, scratch);
        GVC_CUDA(cudaGetLastError());
        resize_vpass_kernel<<<dim3((mdw + 255) / 256, mdh, nf * 3), 256, 0, st>>>(Q, scratch);
        GVC_CUDA(cudaGetLastError());
    }
    return GVC_OK;
}
