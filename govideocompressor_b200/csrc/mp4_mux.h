// mp4_mux.h -- minimal ISO BMFF (MP4) writer for one H.264 video track (product code, host).
// In the reference the container is written by libavformat inside the spawned ffmpeg (client.go:115); the
// segments it produces are later joined by `ffmpeg -f concat -c copy` (server.go:357), which needs every
// segment to carry the same codec parameters: avcC here always holds the one SPS/PPS pair of the stream.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace gvc {

struct Mp4Sample { size_t offset, size; bool sync; };   // offset/size inside the sample payload blob

// samples: length-prefixed NAL units (4-byte big-endian), one access unit per sample.
// Returns false and fills err on I/O failure.  The file is written to `path` directly (callers use a temp name).
bool write_mp4(const std::string &path, int width, int height, int fps_num, int fps_den, const std::vector<uint8_t> &sps_nal,
               const std::vector<uint8_t> &pps_nal, const std::vector<uint8_t> &payload, const std::vector<Mp4Sample> &samples,
               std::string &err);

}  // namespace gvc
