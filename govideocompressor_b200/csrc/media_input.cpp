// media_input.cpp -- see media_input.h.
#include "media_input.h"

#include <dirent.h>
#include <dlfcn.h>

#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

namespace gvc {
namespace {

// ------------------------------------------------------------------------------- YUV4MPEG2 -----
class Y4mSource : public FrameSource {
public:
    explicit Y4mSource(FILE *f) : f_(f) {}
    ~Y4mSource() override { if (f_) fclose(f_); }
    bool parse_header(std::string &err) {
        char line[256];
        if (!fgets(line, sizeof(line), f_) || strncmp(line, "YUV4MPEG2", 9) != 0) { err = "not a YUV4MPEG2 stream"; return false; }
        bool c420 = true;
        for (char *t = strtok(line + 9, " \n"); t; t = strtok(nullptr, " \n")) {
            if (t[0] == 'W') w = atoi(t + 1);
            else if (t[0] == 'H') h = atoi(t + 1);
            else if (t[0] == 'F') { if (sscanf(t + 1, "%d:%d", &fps_num, &fps_den) != 2 || fps_num < 1 || fps_den < 1) { fps_num = 30; fps_den = 1; } }
            else if (t[0] == 'C') c420 = strncmp(t + 1, "420", 3) == 0 && strstr(t, "p1") == nullptr;
            else if (t[0] == 'I' && t[1] != 'p' && t[1] != '?') { err = "interlaced input is not supported"; return false; }
        }
        if (!c420) { err = "only 8-bit 4:2:0 input is supported"; return false; }
        if (w < 16 || h < 16) { err = "bad picture size in the Y4M header"; return false; }
        format_name = "yuv4mpegpipe";
        return true;
    }
    int read(uint8_t *dst, int max_frames, std::string &err) override {
        const size_t fb = size_t(w) * h + 2 * size_t((w + 1) / 2) * ((h + 1) / 2);
        int n = 0;
        while (n < max_frames) {
            char fl[128];
            if (!fgets(fl, sizeof(fl), f_)) break;
            if (strncmp(fl, "FRAME", 5) != 0) { err = "damaged Y4M frame header"; return -1; }
            if (fread(dst + fb * n, 1, fb, f_) != fb) { err = "truncated Y4M frame"; return -1; }
            ++n;
        }
        return n;
    }
private:
    FILE *f_;
};

// ------------------------------------------------------------------------------- libav* via dlopen
struct AvFrameHead { uint8_t *data[8]; int linesize[8]; uint8_t **extended_data; int width, height, nb_samples, format; };
struct AvPacketHead { void *buf; int64_t pts, dts; uint8_t *data; int size; int stream_index; };
struct AvRational { int num, den; };

struct LavApi {
    void *h_util = nullptr, *h_codec = nullptr, *h_format = nullptr;
    int (*avformat_open_input)(void **, const char *, void *, void **) = nullptr;
    int (*avformat_find_stream_info)(void *, void **) = nullptr;
    int (*av_find_best_stream)(void *, int, int, int, const void **, int) = nullptr;
    void (*avformat_close_input)(void **) = nullptr;
    int (*av_read_frame)(void *, void *) = nullptr;
    AvRational (*av_guess_frame_rate)(void *, void *, void *) = nullptr;
    void *(*avcodec_alloc_context3)(const void *) = nullptr;
    int (*avcodec_parameters_to_context)(void *, const void *) = nullptr;
    int (*avcodec_open2)(void *, const void *, void **) = nullptr;
    void (*avcodec_free_context)(void **) = nullptr;
    int (*avcodec_send_packet)(void *, const void *) = nullptr;
    int (*avcodec_receive_frame)(void *, void *) = nullptr;
    void *(*av_packet_alloc)() = nullptr;
    void (*av_packet_free)(void **) = nullptr;
    void (*av_packet_unref)(void *) = nullptr;
    void *(*av_frame_alloc)() = nullptr;
    void (*av_frame_free)(void **) = nullptr;
    void (*av_frame_unref)(void *) = nullptr;
    unsigned (*avformat_version)() = nullptr;
    int (*av_opt_set_int)(void *, const char *, int64_t, int) = nullptr;      // optional: decoder threads
    bool ok = false;
    std::string where;
};

void *open_lib(const std::string &dir, const char *stem) {
    if (!dir.empty()) {
        if (DIR *d = opendir(dir.c_str())) {
            std::string best;
            while (dirent *e = readdir(d)) {
                const std::string n = e->d_name;
                if (n.compare(0, strlen(stem), stem) == 0 && n.find(".so") != std::string::npos) best = dir + "/" + n;
            }
            closedir(d);
            if (!best.empty())
                if (void *h = dlopen(best.c_str(), RTLD_NOW | RTLD_GLOBAL)) return h;
        }
    }
    const std::string plain = std::string(stem) + ".so";
    return dlopen(plain.c_str(), RTLD_NOW | RTLD_GLOBAL);
}

LavApi &lav() {
    static LavApi api;
    static bool   tried = false;
    if (tried) return api;
    tried = true;
    const char *env = getenv("GVC_FFMPEG_LIBDIR");
    const std::string dir = env ? env : "";
    api.h_util = open_lib(dir, "libavutil");
    api.h_codec = open_lib(dir, "libavcodec");
    api.h_format = open_lib(dir, "libavformat");
    if (!api.h_util || !api.h_codec || !api.h_format) {
        const char *why = dlerror();
        api.where = std::string("libavformat/libavcodec/libavutil could not be loaded (set GVC_FFMPEG_LIBDIR)") + (why ? std::string(": ") + why : std::string());
        return api;
    }
    bool ok = true;
#define LOAD(handle, name) ok = ok && (*(void **)(&api.name) = dlsym(handle, #name)) != nullptr
    LOAD(api.h_format, avformat_open_input); LOAD(api.h_format, avformat_find_stream_info); LOAD(api.h_format, av_find_best_stream);
    LOAD(api.h_format, avformat_close_input); LOAD(api.h_format, av_read_frame); LOAD(api.h_format, av_guess_frame_rate);
    LOAD(api.h_format, avformat_version);
    LOAD(api.h_codec, avcodec_alloc_context3); LOAD(api.h_codec, avcodec_parameters_to_context); LOAD(api.h_codec, avcodec_open2);
    LOAD(api.h_codec, avcodec_free_context); LOAD(api.h_codec, avcodec_send_packet); LOAD(api.h_codec, avcodec_receive_frame);
    LOAD(api.h_codec, av_packet_alloc); LOAD(api.h_codec, av_packet_free); LOAD(api.h_codec, av_packet_unref);
    LOAD(api.h_util, av_frame_alloc); LOAD(api.h_util, av_frame_free); LOAD(api.h_util, av_frame_unref);
#undef LOAD
    if (!ok) { api.where = "libav* symbols missing"; return api; }
    *(void **)(&api.av_opt_set_int) = dlsym(api.h_util, "av_opt_set_int");
    const unsigned major = api.avformat_version() >> 16;
    if (major < 59 || major > 63) { api.where = "libavformat major version " + std::to_string(major) + " is outside the range this reader was written for (59..63)"; return api; }
    api.ok = true;
    return api;
}

class LavSource : public FrameSource {
public:
    ~LavSource() override {
        LavApi &a = lav();
        {
            std::lock_guard<std::mutex> g(mu_);
            stop_ = true;
        }
        cv_job_.notify_all();
        for (auto &t : copiers_) t.join();
        for (void *&f : ring_)
            if (f) a.av_frame_free(&f);
        frame_ = nullptr;
        if (pkt_) a.av_packet_free(&pkt_);
        if (cctx_) a.avcodec_free_context(&cctx_);
        if (fmt_) a.avformat_close_input(&fmt_);
    }
    bool open(const std::string &path, std::string &err) {
        LavApi &a = lav();
        if (!a.ok) { err = a.where; return false; }
        if (a.avformat_open_input(&fmt_, path.c_str(), nullptr, nullptr) < 0) { fmt_ = nullptr; err = "cannot open / unknown container"; return false; }
        // AVFormatContext: av_class, iformat, ...; AVInputFormat: name, ...  An MP4 / MOV header already carries what this reader
        // needs (stream types, decoder configuration, frame rate); avformat_find_stream_info would decode the first pictures on
        // one thread just to learn the pixel format, which the real decoder reports anyway -- a quarter of a short piece's time.
        const char *const *ifmt = *reinterpret_cast<const char *const **>(static_cast<char *>(fmt_) + 8);
        const bool mov = ifmt && *ifmt && strncmp(*ifmt, "mov,", 4) == 0;
        if (!mov && a.avformat_find_stream_info(fmt_, nullptr) < 0) { err = "no stream information"; return false; }
        const void *dec = nullptr;
        stream_ = a.av_find_best_stream(fmt_, /*AVMEDIA_TYPE_VIDEO*/ 0, -1, -1, &dec, 0);
        if (stream_ < 0 || !dec) { err = "no decodable video stream"; return false; }
        // AVFormatContext: av_class, iformat, oformat, priv_data, pb (5 pointers), ctx_flags, nb_streams, streams
        const unsigned nb = *reinterpret_cast<unsigned *>(static_cast<char *>(fmt_) + 44);
        void **streams = *reinterpret_cast<void ***>(static_cast<char *>(fmt_) + 48);
        if (nb == 0 || nb > 64 || unsigned(stream_) >= nb || !streams) { err = "unexpected AVFormatContext layout"; return false; }
        void *st = streams[stream_];
        // AVStream: av_class, index, id, codecpar
        void *codecpar = *reinterpret_cast<void **>(static_cast<char *>(st) + 16);
        if (*reinterpret_cast<int *>(static_cast<char *>(st) + 8) != stream_ || !codecpar) { err = "unexpected AVStream layout"; return false; }
        cctx_ = a.avcodec_alloc_context3(dec);
        if (!cctx_ || a.avcodec_parameters_to_context(cctx_, codecpar) < 0) { err = "cannot open the video decoder"; return false; }
        // host decode is what bounds a worker once the kernels take milliseconds: frame threads, but not one per core as the
        // spawned ffmpeg's decoder would take -- a piece is a few dozen pictures, and every frame thread adds a picture of
        // start-up latency (measured on 38-picture 1080p pieces, 16 cores: 747 / 1,564 / 1,543 / 1,229 pictures/s with 1 / 4 / 8 /
        // 16 threads, tools/exp_decode_threads.py).  GVC_DECODE_THREADS overrides (1 = single-threaded, 0 = libavcodec's choice).
        if (a.av_opt_set_int) {
            const char *env = getenv("GVC_DECODE_THREADS");
            const long  cores = long(std::thread::hardware_concurrency());
            const long  n = env ? atol(env) : std::max(1L, std::min(8L, cores));
            a.av_opt_set_int(cctx_, "threads", n < 0 ? 0 : n, 0);
        }
        if (a.avcodec_open2(cctx_, dec, nullptr) < 0) { err = "cannot open the video decoder"; return false; }
        // audio: AVCodecParameters starts with codec_type, codec_id, codec_tag, extradata, extradata_size in every libavcodec
        // this reader accepts; AV_CODEC_ID_AAC = 0x15002
        astream_ = a.av_find_best_stream(fmt_, /*AVMEDIA_TYPE_AUDIO*/ 1, -1, -1, nullptr, 0);
        if (astream_ >= 0 && unsigned(astream_) < nb) {
            const char *ap = *reinterpret_cast<char **>(static_cast<char *>(streams[astream_]) + 16);
            has_audio = true;
            if (ap) {
                audio_codec_id = *reinterpret_cast<const int *>(ap + 4);
                const uint8_t *xd = *reinterpret_cast<uint8_t *const *>(ap + 16);
                const int xn = *reinterpret_cast<const int *>(ap + 24);
                if (audio_codec_id == 0x15002 && xd && xn >= 2 && xn <= 64) {
                    audio.asc.assign(xd, xd + xn);
                    audio_copyable = parse_aac_config(audio);
                }
            }
        } else {
            astream_ = -1;
        }
        AvRational r = a.av_guess_frame_rate(fmt_, st, nullptr);
        if (mov && (r.num <= 0 || r.den <= 0) && a.avformat_find_stream_info(fmt_, nullptr) >= 0)    // a header without a usable rate: probe after all
            r = a.av_guess_frame_rate(fmt_, st, nullptr);
        if (r.num > 0 && r.den > 0) { fps_num = r.num; fps_den = r.den; }
        pkt_ = a.av_packet_alloc();
        for (void *&f : ring_) f = a.av_frame_alloc();
        frame_ = ring_[0];
        if (!pkt_ || std::find(ring_, ring_ + kRing, nullptr) != ring_ + kRing) { err = "out of memory"; return false; }
        format_name = "lavf";
        // decode the first picture now: the caller needs the size before it reads
        if (!pump(err)) return false;
        if (!have_) { err = "no pictures in the video stream"; return false; }
        return true;
    }
    int read(uint8_t *dst, int max_frames, std::string &err) override {
        const size_t fb = size_t(w) * h + 2 * size_t((w + 1) / 2) * ((h + 1) / 2);
        int n = 0;
        while (n < max_frames) {
            if (!have_ && !pump(err)) { drain(); return -1; }
            if (!have_) break;
            hand_to_copier(dst + fb * n);                     // the picture's 3 MB are moved by a helper while the decoder is fed
            have_ = false;
            ++n;
        }
        drain();                                               // the caller reads dst next
        return n;
    }
private:
    // Advances the decoder until a picture is available in frame_ (have_) or the stream is exhausted.
    bool pump(std::string &err) {
        LavApi &a = lav();
        for (;;) {
            const int rc = a.avcodec_receive_frame(cctx_, frame_);
            if (rc == 0) {
                const AvFrameHead *fh = static_cast<const AvFrameHead *>(frame_);
                if (fh->format != 0 && fh->format != 12) { err = "decoded pixel format is not 8-bit 4:2:0"; return false; }
                if (w == 0) { w = fh->width; h = fh->height; }
                if (fh->width != w || fh->height != h) { err = "picture size changes inside the segment"; return false; }
                have_ = true;
                return true;
            }
            if (flushed_) return true;                        // drained
            int rr;
            for (;;) {                                         // next packet of our stream; audio packets are kept on the way
                rr = a.av_read_frame(fmt_, pkt_);
                const AvPacketHead *ph = static_cast<AvPacketHead *>(pkt_);
                if (rr < 0 || ph->stream_index == stream_) break;
                if (audio_copyable && ph->stream_index == astream_ && ph->data && ph->size > 0) audio.packets.emplace_back(ph->data, ph->data + ph->size);
                a.av_packet_unref(pkt_);
            }
            if (rr < 0) { a.avcodec_send_packet(cctx_, nullptr); flushed_ = true; continue; }
            const int sr = a.avcodec_send_packet(cctx_, pkt_);
            a.av_packet_unref(pkt_);
            if (sr < 0) { err = "the decoder rejected a packet"; return false; }
        }
    }
    // Decoded pictures leave through a ring of AVFrames: the reading thread keeps pumping packets into the decoder while
    // kCopiers helpers move finished pictures into the caller's (pinned) buffer -- at 1080p the row-wise copy of a picture
    // costs about as much host time as the frame-threaded decoder needs to produce it.
    static constexpr int kRing = 6, kCopiers = 2;
    struct CopyJob { int slot; uint8_t *dst; };
    void hand_to_copier(uint8_t *dst) {
        std::unique_lock<std::mutex> lk(mu_);
        if (copiers_.empty())
            for (int i = 0; i < kCopiers; ++i) copiers_.emplace_back([this] { copier(); });
        int slot = 0;
        while (ring_[slot] != frame_) ++slot;
        busy_[slot] = true;
        ++pending_;
        jobs_.push_back(CopyJob{slot, dst});
        cv_job_.notify_one();
        // next picture goes into a free frame
        cv_done_.wait(lk, [&] { return std::find(busy_, busy_ + kRing, false) != busy_ + kRing; });
        frame_ = ring_[std::find(busy_, busy_ + kRing, false) - busy_];
    }
    void drain() {
        std::unique_lock<std::mutex> lk(mu_);
        cv_done_.wait(lk, [&] { return pending_ == 0; });
    }
    void copier() {
        std::unique_lock<std::mutex> lk(mu_);
        for (;;) {
            cv_job_.wait(lk, [&] { return stop_ || !jobs_.empty(); });
            if (jobs_.empty()) return;                         // stop_
            const CopyJob j = jobs_.front();
            jobs_.pop_front();
            lk.unlock();
            copy_out(ring_[j.slot], j.dst);
            lav().av_frame_unref(ring_[j.slot]);
            lk.lock();
            busy_[j.slot] = false;
            --pending_;
            cv_done_.notify_all();
        }
    }
    void copy_out(const void *frame, uint8_t *dst) const {
        const AvFrameHead *fh = static_cast<const AvFrameHead *>(frame);
        for (int k = 0; k < 3; ++k) {
            const int pw = k ? (w + 1) / 2 : w, ph = k ? (h + 1) / 2 : h;
            for (int y = 0; y < ph; ++y) memcpy(dst + size_t(y) * pw, fh->data[k] + size_t(y) * fh->linesize[k], pw);
            dst += size_t(pw) * ph;
        }
    }
    void *fmt_ = nullptr, *cctx_ = nullptr, *pkt_ = nullptr, *frame_ = nullptr;      // frame_: the ring entry the decoder fills next
    void *ring_[kRing] = {};
    bool  busy_[kRing] = {};
    std::mutex mu_;
    std::condition_variable cv_job_, cv_done_;
    std::deque<CopyJob> jobs_;
    std::vector<std::thread> copiers_;
    int   pending_ = 0;
    bool  stop_ = false;
    int   stream_ = -1, astream_ = -1;
    bool  have_ = false, flushed_ = false;
};

}  // namespace

std::unique_ptr<FrameSource> open_frame_source(const std::string &path, std::string &err) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) { err = "No such file or directory"; return nullptr; }
    char magic[10] = {0};
    const size_t got = fread(magic, 1, 9, f);
    if (got == 9 && memcmp(magic, "YUV4MPEG2", 9) == 0) {
        rewind(f);
        std::unique_ptr<Y4mSource> y(new Y4mSource(f));
        if (!y->parse_header(err)) return nullptr;
        return std::unique_ptr<FrameSource>(y.release());
    }
    fclose(f);
    std::unique_ptr<LavSource> l(new LavSource);
    if (!l->open(path, err)) return nullptr;
    return std::unique_ptr<FrameSource>(l.release());
}

}  // namespace gvc
