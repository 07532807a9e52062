// ff_args.cpp -- see ff_args.h.
#include "ff_args.h"

#include <cstdlib>
#include <sstream>
#include <vector>

#include "../../include/gvc.h"

namespace gvc {
namespace {

bool parse_size(const std::string &s, char sep, int &w, int &h) {
    const size_t p = s.find(sep);
    if (p == std::string::npos) return false;
    char *e1 = nullptr, *e2 = nullptr;
    const long a = strtol(s.c_str(), &e1, 10), b = strtol(s.c_str() + p + 1, &e2, 10);
    if (e1 != s.c_str() + p || a < 2 || b < 2 || a > 16384 || b > 16384) return false;
    w = int(a); h = int(b);
    return *e2 == 0 || *e2 == ':';
}

bool parse_flags(const std::string &f, int &algo) {
    if (f == "bicubic") { algo = GVC_SCALE_BICUBIC; return true; }
    if (f == "bilinear") { algo = GVC_SCALE_BILINEAR; return true; }
    return false;
}

}  // namespace

int parse_ffmpeg_args(const std::string &args, TranscodeOptions &opt, std::string &err) {
    std::vector<std::string> tok;
    {   // client.go:105 uses strings.Split(args, " "): empty tokens from double spaces are possible; skip them
        std::stringstream ss(args);
        std::string       t;
        while (std::getline(ss, t, ' '))
            if (!t.empty()) tok.push_back(t);
    }
    bool have_codec = false;
    for (size_t i = 0; i < tok.size(); ++i) {
        const std::string &k = tok[i];
        auto need = [&](std::string &v) {
            if (i + 1 >= tok.size()) { err = "option " + k + " needs a value"; return false; }
            v = tok[++i];
            return true;
        };
        std::string v;
        if (k == "-threads" || k == "-pix_fmt" || k == "-profile:v" || k == "-level" || k == "-tune" || k == "-movflags") {
            if (!need(v)) return GVC_ERR_INVALID_ARG;
            if (k == "-pix_fmt" && v != "yuv420p") { err = "only -pix_fmt yuv420p is supported"; return GVC_ERR_UNSUPPORTED; }
        } else if (k == "-vcodec" || k == "-c:v" || k == "-codec:v") {
            if (!need(v)) return GVC_ERR_INVALID_ARG;
            if (v != "libx264" && v != "h264") { err = "video codec '" + v + "' is not available on this path (no CPU fallback)"; return GVC_ERR_UNSUPPORTED; }
            have_codec = true;
        } else if (k == "-crf") {
            if (!need(v)) return GVC_ERR_INVALID_ARG;
            opt.crf = atoi(v.c_str());
            if (opt.crf < 0 || opt.crf > 51) { err = "-crf out of range 0..51"; return GVC_ERR_INVALID_ARG; }
        } else if (k == "-preset") {
            if (!need(opt.preset)) return GVC_ERR_INVALID_ARG;
        } else if (k == "-s") {
            if (!need(v)) return GVC_ERR_INVALID_ARG;
            if (!parse_size(v, 'x', opt.dst_w, opt.dst_h)) { err = "bad -s value '" + v + "'"; return GVC_ERR_INVALID_ARG; }
        } else if (k == "-vf" || k == "-filter:v") {
            if (!need(v)) return GVC_ERR_INVALID_ARG;
            if (v.compare(0, 6, "scale=") != 0) { err = "only -vf scale=W:H[:flags=...] is supported"; return GVC_ERR_UNSUPPORTED; }
            const std::string body = v.substr(6);
            if (!parse_size(body, ':', opt.dst_w, opt.dst_h)) { err = "bad scale filter '" + v + "'"; return GVC_ERR_INVALID_ARG; }
            const size_t fp = body.find("flags=");
            if (fp != std::string::npos && !parse_flags(body.substr(fp + 6), opt.scale_algo)) { err = "unsupported scaler flags in '" + v + "'"; return GVC_ERR_UNSUPPORTED; }
        } else if (k == "-sws_flags") {
            if (!need(v)) return GVC_ERR_INVALID_ARG;
            if (!parse_flags(v, opt.scale_algo)) { err = "unsupported -sws_flags '" + v + "'"; return GVC_ERR_UNSUPPORTED; }
        } else if (k == "-g") {
            if (!need(v)) return GVC_ERR_INVALID_ARG;
            opt.gop = atoi(v.c_str());
            if (opt.gop < 1) { err = "bad -g"; return GVC_ERR_INVALID_ARG; }
        } else if (k == "-me_range") {
            if (!need(v)) return GVC_ERR_INVALID_ARG;
            opt.merange = atoi(v.c_str());
            if (opt.merange < 4 || opt.merange > 44) { err = "-me_range outside 4..44"; return GVC_ERR_INVALID_ARG; }
        } else if (k == "-r") {
            if (!need(v)) return GVC_ERR_INVALID_ARG;
            int a = 0, b = 1;
            if (!parse_size(v, '/', a, b)) { a = atoi(v.c_str()); b = 1; }
            if (a < 1) { err = "bad -r"; return GVC_ERR_INVALID_ARG; }
            opt.fps_num = a; opt.fps_den = b;
        } else if (k == "-an") {
            opt.drop_audio = true;
        } else if (k == "-y") {
        } else if (k == "-acodec" || k == "-c:a" || k == "-codec:a") {
            if (!need(v)) return GVC_ERR_INVALID_ARG;
            // the audio stream is always stream-copied (what the reference's own splitter does, server.go:199-200); a request to
            // re-encode it cannot be honoured on this path
            if (v != "copy" && v != "aac") { err = "audio codec '" + v + "' is not available on this path (audio is stream-copied)"; return GVC_ERR_UNSUPPORTED; }
        } else if (k == "-b:v" || k == "-b") {
            if (!need(v)) return GVC_ERR_INVALID_ARG;
            char *end = nullptr;
            double b = strtod(v.c_str(), &end);
            if (end && (*end == 'k' || *end == 'K')) { b *= 1e3; ++end; }
            else if (end && (*end == 'M' || *end == 'm')) { b *= 1e6; ++end; }
            if (end == v.c_str() || (end && *end) || b < 1000 || b > 2e9) { err = "bad " + k + " value '" + v + "'"; return GVC_ERR_INVALID_ARG; }
            opt.bitrate = long(b);
        } else {
            err = "unrecognised option '" + k + "'";
            return GVC_ERR_INVALID_ARG;
        }
    }
    if (!have_codec) { err = "no -vcodec libx264 in the argument string: this path only encodes H.264"; return GVC_ERR_UNSUPPORTED; }
    return GVC_OK;
}

}  // namespace gvc
