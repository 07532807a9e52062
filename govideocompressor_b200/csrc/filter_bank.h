// filter_bank.h -- host-side polyphase filter banks for the resize kernels (product code).
//
// Same numerical contract as libswscale's utils.c initFilter() for the configuration ffmpeg uses on
// the reference's path (client.go:115 -> `-s WxH` / `-vf scale=`): bicubic (B=0, C=0.6) or bilinear,
// no src/dst blur vectors, SWS_BITEXACT tap trimming.  The resulting (position, taps) pairs are what
// the CUDA kernels consume; tests compare them tap-for-tap with the oracle's bank.
#pragma once
#include <cstdint>
#include <vector>

namespace gvc {

struct FilterBank {
    int                  taps = 0;     // taps per output sample
    int                  dst_len = 0;
    std::vector<int16_t> coef;         // dst_len * taps, each row sums to `one`
    std::vector<int32_t> pos;          // first source sample of each row, 0 <= pos, pos+taps <= max(src_len,taps)
};

enum : int { kBilinear = 2, kBicubic = 4 };

// phase128: sample phase in 1/128 source-sample units (128 = co-sited centres; what swscale's
// get_local_pos() yields for luma and for default-sited 4:2:0 chroma).
bool build_filter_bank(FilterBank &out, int src_len, int dst_len, int algo, int one, int phase128 = 128);

}  // namespace gvc
