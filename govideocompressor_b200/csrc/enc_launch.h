// enc_launch.h -- stage launchers defined in gvc_enc.cu, shared with the segment encoder (gvc_segment.cu).
#pragma once
#include <cuda_runtime.h>

#include "enc_kernels.cuh"
#include "entropy_kernels.cuh"

namespace gvc {
int launch_extend_border(uint8_t *p, int stride, int w, int h, int x_lo, int x_hi, int y_lo, int y_hi, cudaStream_t st);
int launch_lowres(const uint8_t *src, int sstride, int x_lo, int x_hi, int y_lo, int y_hi, uint8_t *d0, uint8_t *dh, uint8_t *dv,
                  uint8_t *dc, int dstride, int lx_lo, int lx_hi, int ly_lo, int ly_hi, cudaStream_t st);
// batched launchers: `count` (1..kMaxBatch) argument records of equal geometry, one per segment, in ONE launch per kernel
int launch_extend_border_batch(const BorderBatchArgs *a, int count, cudaStream_t st);
int launch_lowres_batch(const LowresBatchArgs *a, int count, cudaStream_t st);
int launch_hpel(const HpelArgs *a, int count, cudaStream_t st);
int launch_me_lowres(const MeLowArgs *a, int count, cudaStream_t st);
int launch_me_refine(const MeRefArgs *a, int count, cudaStream_t st);
int launch_signal_picture(const SignalArgs *a, int count, cudaStream_t st);
int launch_encode_inter(const EncMbArgs *a, int count, cudaStream_t st);
int launch_encode_intra(const IntraArgs *a, int count, cudaStream_t st);
int launch_entropy_picture(const EntropyPicture *p, int count, cudaStream_t st);
}  // namespace gvc
