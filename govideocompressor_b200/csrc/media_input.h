// media_input.h -- host-side readers for the segment handed to gvc_transcode_file (product code).
//
// The reference's pieces are `<id>.mp4` files holding H.264 video (server.go:199 splits with `-vcodec copy`), and the
// spawned ffmpeg demuxes and decodes them itself (client.go:115).  north_star keeps muxing/demuxing and entropy
// (de)coding on the host "in Go/ffmpeg": here the MP4 demuxer and H.264 decoder are FFmpeg's own libavformat /
// libavcodec, loaded at run time with dlopen when they can be found (env GVC_FFMPEG_LIBDIR, or the default library
// path); raw YUV4MPEG2 input needs no external library.  Only functions and the long-stable leading fields of
// AVFrame / AVPacket / AVFormatContext / AVStream are used -- no FFmpeg headers are required to build libgvc.
#pragma once
#include <cstdint>
#include <cstdio>
#include <memory>
#include <string>

#include "mp4_mux.h"

namespace gvc {

class FrameSource {
public:
    virtual ~FrameSource() {}
    int w = 0, h = 0, fps_num = 30, fps_den = 1;
    std::string format_name;
    // The piece's audio stream (the splitter maps one: server.go:199-200).  AAC is carried into the output unchanged: its
    // packets are collected while the pictures are read (complete once read() has returned 0); any other codec is only flagged.
    bool     has_audio = false, audio_copyable = false;
    int      audio_codec_id = 0;
    Mp4Audio audio;
    // Reads up to max_frames pictures as tight I420 into dst (w*h*3/2 bytes each).  Returns the number read
    // (0 at end of stream) or a negative value on a damaged input (err explains).
    virtual int read(uint8_t *dst, int max_frames, std::string &err) = 0;
};

// Opens path as YUV4MPEG2, else through libavformat/libavcodec.  Returns null and fills err on failure.
std::unique_ptr<FrameSource> open_frame_source(const std::string &path, std::string &err);

}  // namespace gvc
