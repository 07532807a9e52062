// mp4_mux.h -- minimal ISO BMFF (MP4) writer for one H.264 video track and, optionally, one AAC track copied from the input
// (product code, host).
// In the reference the container is written by libavformat inside the spawned ffmpeg (client.go:115); the
// segments it produces are later joined by `ffmpeg -f concat -c copy` (server.go:357), which needs every
// segment to carry the same codec parameters: avcC here always holds the one SPS/PPS pair of the stream.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace gvc {

struct Mp4Sample { size_t offset, size; bool sync; };   // offset/size inside the sample payload blob

// An AAC track carried over unchanged from the input piece (the reference's splitter keeps it with `-acodec copy -map 0:1`,
// server.go:199-200): AudioSpecificConfig + the raw access units, 1024 samples each.
struct Mp4Audio {
    std::vector<uint8_t>              asc;       // AudioSpecificConfig (esds DecoderSpecificInfo)
    int                               sample_rate = 0, channels = 0;
    std::vector<std::vector<uint8_t>> packets;   // one raw AAC frame each
};
// Fills sample_rate / channels from the AudioSpecificConfig; false when it is not plain AAC (object types 1..4, no explicit rate).
bool parse_aac_config(Mp4Audio &a);

// samples: length-prefixed NAL units (4-byte big-endian), one access unit per sample.
// Returns false and fills err on I/O failure.  The file is written to `path` directly (callers use a temp name).
bool write_mp4(const std::string &path, int width, int height, int fps_num, int fps_den, const std::vector<uint8_t> &sps_nal,
               const std::vector<uint8_t> &pps_nal, const std::vector<uint8_t> &payload, const std::vector<Mp4Sample> &samples,
               std::string &err, const Mp4Audio *audio = nullptr);

}  // namespace gvc
