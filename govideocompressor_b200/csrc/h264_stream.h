// h264_stream.h -- host-side H.264 bitstream writer (product code, C++): SPS / PPS / slice headers and CAVLC
// slice data from the per-macroblock records the GPU kernels produce.
//
// north_star keeps entropy coding and muxing on the host; in the reference they happen inside the spawned
// ffmpeg (libx264 CAVLC/CABAC + libavformat, client.go:115).  Syntax follows ITU-T H.264 clauses 7.3 (syntax),
// 8.4.1.3 (motion vector prediction), 9.1 (Exp-Golomb), 9.2 (CAVLC).  Stream shape: Constrained Baseline,
// one slice per picture, IDR + P pictures with one reference, Intra16x16 / P_L0_16x16 / P_8x8 / P_Skip,
// 4x4 transform, in-loop deblocking disabled in the slice header (so the kernels' reconstruction is exactly what
// a decoder produces).
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

#include "enc_types.h"

namespace gvc {

class BitWriter {
public:
    BitWriter() { buf_.resize(4096); }
    void put(uint32_t value, int bits) {         // MSB first, 0 <= bits <= 32
        if (bits < 32) value &= (1u << bits) - 1u;
        acc_ = (acc_ << bits) | value;
        n_ += bits;
        if (n_ >= 32) {
            if (len_ + 4 > buf_.size()) buf_.resize(buf_.size() * 2);
            const uint32_t w = uint32_t(acc_ >> (n_ - 32));
            uint8_t *p = buf_.data() + len_;
            p[0] = uint8_t(w >> 24); p[1] = uint8_t(w >> 16); p[2] = uint8_t(w >> 8); p[3] = uint8_t(w);
            len_ += 4;
            n_ -= 32;
        }
    }
    void ue(uint32_t v) {
        const uint64_t code = uint64_t(v) + 1;
        const int len = 63 - __builtin_clzll(code);     // floor(log2(code))
        put(0, len);
        put(uint32_t(code), len + 1);
    }
    void se(int32_t v) { ue(v > 0 ? uint32_t(2 * v - 1) : uint32_t(-2 * int64_t(v))); }
    void rbsp_trailing();                        // stop bit + zero alignment; finalises bytes()
    size_t bit_count() const { return len_ * 8 + n_; }
    std::vector<uint8_t> peek() const {           // the bits written so far, the last byte padded with zeros
        std::vector<uint8_t> v(buf_.begin(), buf_.begin() + len_);
        for (int left = n_; left > 0; left -= 8) v.push_back(uint8_t(left >= 8 ? acc_ >> (left - 8) : acc_ << (8 - left)));
        return v;
    }
    const std::vector<uint8_t> &bytes() const { return out_; }   // valid after rbsp_trailing()
private:
    std::vector<uint8_t> buf_, out_;
    size_t   len_ = 0;
    uint64_t acc_ = 0;
    int      n_   = 0;
};

struct StreamParams {
    int w = 0, h = 0;            // display size
    int w16 = 0, h16 = 0;        // coded size
    int fps_num = 30, fps_den = 1;
    int level_idc = 31;
};

int  pick_level(int mbw, int mbh, int fps_num, int fps_den);
void write_sps(BitWriter &bw, const StreamParams &p);
void write_pps(BitWriter &bw);

// Wraps an RBSP into a NAL unit (header byte + emulation prevention).  annexb: prefix 00 00 00 01, else a
// 4-byte big-endian length (the MP4 sample format).
void append_nal(std::vector<uint8_t> &out, int nal_ref_idc, int nal_unit_type, const std::vector<uint8_t> &rbsp, bool annexb);

// Everything the slice writer needs about one coded picture (host copies of the kernels' outputs).
struct PictureData {
    int             mbw = 0, mbh = 0;
    bool            idr = false;
    int             frame_num = 0, idr_pic_id = 0, qp = 26;
    const int16_t  *coefs = nullptr;    // kMbCoefs per MB (layout: enc_kernels.cuh), or, with compact = true, only the blocks
                                        // whose compact_mask bit is set, 16 coefficients each, packed MB after MB
    bool            compact = false;
    const uint32_t *block_offsets = nullptr;   // compact: per MB, index (in blocks) of its first packed block
    const uint32_t *nzmask = nullptr;   // per MB
    const MbMotion *motion = nullptr;   // P pictures
    const uint8_t  *modes = nullptr;    // IDR pictures: Intra16x16 prediction mode
};

struct SliceStats { int skipped = 0, p16 = 0, p8 = 0, intra = 0; size_t bits = 0; };

// Writes slice_layer_without_partitioning_rbsp() for the whole picture.
void write_slice(BitWriter &bw, const PictureData &pic, SliceStats *stats);
// slice_header() alone (uses idr, frame_num, idr_pic_id, qp): its length in bits, and its bits as bytes (zero padded) --
// for slices whose slice_data() was written by the GPU behind room for the header.
void write_slice_header(BitWriter &bw, const PictureData &pic);
size_t slice_header_bits(const PictureData &pic);
std::vector<uint8_t> slice_header_bytes(const PictureData &pic);

}  // namespace gvc
