// mp4_mux.cpp -- see mp4_mux.h.  Box layout: ftyp, moov(mvhd, trak(tkhd, mdia(mdhd, hdlr, minf(vmhd, dinf(dref),
// [second trak: the copied AAC track, mp4a/esds],
// stbl(stsd(avc1(avcC)), stts, stss, stsc, stsz, stco))))), mdat.  ISO/IEC 14496-12 / 14496-15.
#include "mp4_mux.h"

#include <cstdio>
#include <cstring>

namespace gvc {
namespace {

struct Box {
    std::vector<uint8_t> d;
    void u8(uint32_t v) { d.push_back(uint8_t(v)); }
    void u16(uint32_t v) { u8(v >> 8); u8(v); }
    void u32(uint32_t v) { u16(v >> 16); u16(v & 0xffff); }
    void tag(const char *t) { d.insert(d.end(), t, t + 4); }
    void zeros(int n) { d.insert(d.end(), size_t(n), 0); }
    void bytes(const std::vector<uint8_t> &b) { d.insert(d.end(), b.begin(), b.end()); }
    size_t begin(const char *t) { const size_t at = d.size(); u32(0); tag(t); return at; }
    void   end(size_t at) { const uint32_t n = uint32_t(d.size() - at); d[at] = uint8_t(n >> 24); d[at + 1] = uint8_t(n >> 16); d[at + 2] = uint8_t(n >> 8); d[at + 3] = uint8_t(n); }
    void   full(uint32_t version_flags) { u32(version_flags); }
};

void matrix(Box &b) {
    const uint32_t m[9] = {0x10000, 0, 0, 0, 0x10000, 0, 0, 0, 0x40000000};
    for (uint32_t v : m) b.u32(v);
}

}  // namespace

bool parse_aac_config(Mp4Audio &a) {
    if (a.asc.size() < 2) return false;
    static const int rates[13] = {96000, 88200, 64000, 48000, 44100, 32000, 24000, 22050, 16000, 12000, 11025, 8000, 7350};
    const int obj = a.asc[0] >> 3, fi = ((a.asc[0] & 7) << 1) | (a.asc[1] >> 7), ch = (a.asc[1] >> 3) & 15;
    if (obj < 1 || obj > 4 || fi > 12 || ch < 1 || ch > 7) return false;
    a.sample_rate = rates[fi];
    a.channels = ch == 7 ? 8 : ch;
    return true;
}

namespace {
// esds descriptor length: one byte is enough for the sizes written here (< 128)
void descr(Box &b, int tag, const std::vector<uint8_t> &body) { b.u8(uint32_t(tag)); b.u8(uint32_t(body.size())); b.bytes(body); }

// trak of the AAC track: tkhd, mdia(mdhd, hdlr soun, minf(smhd, dinf, stbl(stsd(mp4a(esds)), stts, stsc, stsz, stco)))
void audio_trak(Box &b, const Mp4Audio &a, uint32_t movie_timescale, size_t &stco_value_at) {
    const uint32_t n = uint32_t(a.packets.size()), dur = n * 1024u;
    const uint32_t movie_dur = uint32_t(uint64_t(dur) * movie_timescale / uint32_t(a.sample_rate));
    const size_t trak = b.begin("trak");
    size_t at = b.begin("tkhd"); b.full(3);
    b.u32(0); b.u32(0); b.u32(2); b.u32(0); b.u32(movie_dur); b.zeros(8); b.u16(0); b.u16(1); b.u16(0x100); b.u16(0); matrix(b);
    b.u32(0); b.u32(0);
    b.end(at);
    const size_t mdia = b.begin("mdia");
    at = b.begin("mdhd"); b.full(0); b.u32(0); b.u32(0); b.u32(uint32_t(a.sample_rate)); b.u32(dur); b.u16(0x55c4); b.u16(0); b.end(at);
    at = b.begin("hdlr"); b.full(0); b.u32(0); b.tag("soun"); b.zeros(12); b.d.insert(b.d.end(), {'g', 'v', 'c', 0}); b.end(at);
    const size_t minf = b.begin("minf");
    at = b.begin("smhd"); b.full(0); b.u16(0); b.u16(0); b.end(at);
    const size_t dinf = b.begin("dinf");
    at = b.begin("dref"); b.full(0); b.u32(1);
    { const size_t u = b.begin("url "); b.full(1); b.end(u); }
    b.end(at);
    b.end(dinf);
    const size_t stbl = b.begin("stbl");
    at = b.begin("stsd"); b.full(0); b.u32(1);
    {
        const size_t mp4a = b.begin("mp4a");
        b.zeros(6); b.u16(1); b.zeros(8); b.u16(uint32_t(a.channels)); b.u16(16); b.u16(0); b.u16(0); b.u32(uint32_t(a.sample_rate) << 16);
        const size_t esds = b.begin("esds"); b.full(0);
        Box dsi, dcd, es;
        descr(dsi, 5, a.asc);                                          // DecoderSpecificInfo = AudioSpecificConfig
        dcd.u8(0x40); dcd.u8(0x15); dcd.u8(0); dcd.u16(0); dcd.u32(0); dcd.u32(0);   // MPEG-4 audio, audio stream, buffer size / bit rates unknown
        dcd.bytes(dsi.d);
        es.u16(0); es.u8(0);                                           // ES_ID, flags
        descr(es, 4, dcd.d);
        descr(es, 6, std::vector<uint8_t>{2});                          // SLConfigDescriptor: MP4 predefined
        descr(b, 3, es.d);
        b.end(esds);
        b.end(mp4a);
    }
    b.end(at);
    at = b.begin("stts"); b.full(0); b.u32(1); b.u32(n); b.u32(1024); b.end(at);
    at = b.begin("stsc"); b.full(0); b.u32(1); b.u32(1); b.u32(n); b.u32(1); b.end(at);
    at = b.begin("stsz"); b.full(0); b.u32(0); b.u32(n);
    for (const std::vector<uint8_t> &p : a.packets) b.u32(uint32_t(p.size()));
    b.end(at);
    at = b.begin("stco"); b.full(0); b.u32(1);
    stco_value_at = b.d.size();
    b.u32(0);
    b.end(at);
    b.end(stbl); b.end(minf); b.end(mdia); b.end(trak);
}
}  // namespace

bool write_mp4(const std::string &path, int width, int height, int fps_num, int fps_den, const std::vector<uint8_t> &sps_nal,
               const std::vector<uint8_t> &pps_nal, const std::vector<uint8_t> &payload, const std::vector<Mp4Sample> &samples,
               std::string &err, const Mp4Audio *audio) {
    if (audio && (audio->packets.empty() || audio->sample_rate <= 0 || audio->asc.empty() || audio->asc.size() > 64)) audio = nullptr;
    const uint32_t timescale = uint32_t(fps_num), delta = uint32_t(fps_den), n = uint32_t(samples.size());
    const uint32_t duration = n * delta;
    Box b;
    size_t at = b.begin("ftyp");
    b.tag("isom"); b.u32(0x200); b.tag("isom"); b.tag("iso2"); b.tag("avc1"); b.tag("mp41");
    b.end(at);

    const size_t moov = b.begin("moov");
    at = b.begin("mvhd"); b.full(0);
    b.u32(0); b.u32(0); b.u32(timescale); b.u32(duration); b.u32(0x10000); b.u16(0x100); b.zeros(10); matrix(b); b.zeros(24); b.u32(audio ? 3 : 2);
    b.end(at);
    const size_t trak = b.begin("trak");
    at = b.begin("tkhd"); b.full(3);
    b.u32(0); b.u32(0); b.u32(1); b.u32(0); b.u32(duration); b.zeros(8); b.u16(0); b.u16(0); b.u16(0); b.u16(0); matrix(b);
    b.u32(uint32_t(width) << 16); b.u32(uint32_t(height) << 16);
    b.end(at);
    const size_t mdia = b.begin("mdia");
    at = b.begin("mdhd"); b.full(0); b.u32(0); b.u32(0); b.u32(timescale); b.u32(duration); b.u16(0x55c4); b.u16(0); b.end(at);
    at = b.begin("hdlr"); b.full(0); b.u32(0); b.tag("vide"); b.zeros(12); b.d.insert(b.d.end(), {'g', 'v', 'c', 0}); b.end(at);
    const size_t minf = b.begin("minf");
    at = b.begin("vmhd"); b.full(1); b.zeros(8); b.end(at);
    const size_t dinf = b.begin("dinf");
    at = b.begin("dref"); b.full(0); b.u32(1);
    { const size_t u = b.begin("url "); b.full(1); b.end(u); }
    b.end(at);
    b.end(dinf);
    const size_t stbl = b.begin("stbl");
    at = b.begin("stsd"); b.full(0); b.u32(1);
    {
        const size_t avc1 = b.begin("avc1");
        b.zeros(6); b.u16(1); b.zeros(16); b.u16(uint32_t(width)); b.u16(uint32_t(height)); b.u32(0x480000); b.u32(0x480000); b.u32(0);
        b.u16(1); b.zeros(32); b.u16(0x18); b.u16(0xffff);
        const size_t avcc = b.begin("avcC");
        b.u8(1); b.u8(sps_nal[1]); b.u8(sps_nal[2]); b.u8(sps_nal[3]); b.u8(0xff);   // 4-byte NAL lengths
        b.u8(0xe1); b.u16(uint32_t(sps_nal.size())); b.bytes(sps_nal);
        b.u8(1); b.u16(uint32_t(pps_nal.size())); b.bytes(pps_nal);
        b.end(avcc);
        b.end(avc1);
    }
    b.end(at);
    at = b.begin("stts"); b.full(0); b.u32(1); b.u32(n); b.u32(delta); b.end(at);
    at = b.begin("stss"); b.full(0);
    {
        uint32_t ns = 0;
        for (const Mp4Sample &s : samples) ns += s.sync;
        b.u32(ns);
        for (uint32_t i = 0; i < n; ++i)
            if (samples[i].sync) b.u32(i + 1);
    }
    b.end(at);
    at = b.begin("stsc"); b.full(0); b.u32(1); b.u32(1); b.u32(n); b.u32(1); b.end(at);
    at = b.begin("stsz"); b.full(0); b.u32(0); b.u32(n);
    for (const Mp4Sample &s : samples) b.u32(uint32_t(s.size));
    b.end(at);
    at = b.begin("stco"); b.full(0); b.u32(1);
    const size_t stco_value_at = b.d.size();
    b.u32(0);
    b.end(at);
    b.end(stbl); b.end(minf); b.end(mdia); b.end(trak);
    size_t audio_stco_at = 0, audio_bytes = 0;
    if (audio) {
        audio_trak(b, *audio, timescale, audio_stco_at);
        for (const std::vector<uint8_t> &p : audio->packets) audio_bytes += p.size();
    }
    b.end(moov);

    const uint32_t mdat_payload_at = uint32_t(b.d.size() + 8);
    if (audio) {      // the audio chunk follows the video chunk inside mdat
        const uint64_t aat = uint64_t(mdat_payload_at) + payload.size();
        if (aat + audio_bytes > 0xffffffffull) { err = "segment larger than 4 GiB"; return false; }
        b.d[audio_stco_at] = uint8_t(aat >> 24); b.d[audio_stco_at + 1] = uint8_t(aat >> 16); b.d[audio_stco_at + 2] = uint8_t(aat >> 8); b.d[audio_stco_at + 3] = uint8_t(aat);
    }
    b.d[stco_value_at] = uint8_t(mdat_payload_at >> 24); b.d[stco_value_at + 1] = uint8_t(mdat_payload_at >> 16);
    b.d[stco_value_at + 2] = uint8_t(mdat_payload_at >> 8); b.d[stco_value_at + 3] = uint8_t(mdat_payload_at);
    if (payload.size() + audio_bytes + 8 > 0xffffffffull) { err = "segment larger than 4 GiB"; return false; }
    b.u32(uint32_t(payload.size() + audio_bytes + 8)); b.tag("mdat");

    FILE *f = fopen(path.c_str(), "wb");
    if (!f) { err = "cannot open " + path; return false; }
    bool ok = fwrite(b.d.data(), 1, b.d.size(), f) == b.d.size();
    ok = ok && (payload.empty() || fwrite(payload.data(), 1, payload.size(), f) == payload.size());
    if (audio)
        for (const std::vector<uint8_t> &p : audio->packets) ok = ok && (p.empty() || fwrite(p.data(), 1, p.size(), f) == p.size());
    ok = (fclose(f) == 0) && ok;
    if (!ok) { err = "write failed: " + path; remove(path.c_str()); }
    return ok;
}

}  // namespace gvc
