// ff_args.h -- parser for the ffmpeg-style argument string of the reference protocol.
// The reference forwards a free-form string: server flag --ff/-f (server.go:30), shorthand expansion
// "264" -> "-threads 4 -vcodec libx264" (server.go:69-70), sent verbatim (server.go:179), split on single spaces
// with no quoting and spliced between `-i <in>` and the output name (client.go:105-112).  This parser accepts the
// subset that selects this path and rejects the rest (there is no CPU fallback to hand other codecs to).
#pragma once
#include <string>

namespace gvc {

struct TranscodeOptions {
    int  dst_w = 0, dst_h = 0;        // 0 = keep source size
    int  crf = 23;                    // libx264 default
    int  scale_algo = 4;              // GVC_SCALE_BICUBIC (ffmpeg's default scaler)
    int  gop = 250;                   // -g (a segment always starts with an IDR)
    int  merange = 16;                // -me_range
    int  fps_num = 0, fps_den = 0;    // -r (0 = input rate)
    std::string preset = "medium";
    bool drop_audio = false;          // -an
    long bitrate = 0;                 // -b:v in bits per second (0 = constant quality from -crf)
};

// Returns 0 or a negative gvc_status; err describes the offending token.
int parse_ffmpeg_args(const std::string &args, TranscodeOptions &opt, std::string &err);

}  // namespace gvc
