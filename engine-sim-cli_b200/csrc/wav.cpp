// wav.cpp — RIFF/WAVE 16-bit PCM reader and mono writer.
// Reader stands in for dr_wav as used by engine-sim-bridge/include/common/wav_loader.h:41-63
// (drwav_read_pcm_frames_s16: all channels interleaved); the writer follows the
// 44-byte header the reference's spike emits (RealEngineAudioSpike.cpp:22-47) but with a
// consistent mono header.
#include "host.h"

#include <cstdio>
#include <cstring>

namespace esb {

static uint32_t rd32(const uint8_t *p) { return p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24); }
static uint16_t rd16(const uint8_t *p) { return (uint16_t)(p[0] | (p[1] << 8)); }

bool readWav(const std::string &path, WavData &out, std::string &err) {
    FILE *f = std::fopen(path.c_str(), "rb");
    if (!f) { err = "Cannot open WAV file: " + path; return false; }
    std::vector<uint8_t> buf;
    uint8_t tmp[65536]; size_t n;
    while ((n = std::fread(tmp, 1, sizeof(tmp), f)) > 0) buf.insert(buf.end(), tmp, tmp + n);
    std::fclose(f);
    if (buf.size() < 12 || std::memcmp(buf.data(), "RIFF", 4) != 0 || std::memcmp(buf.data() + 8, "WAVE", 4) != 0) {
        err = "Not a RIFF/WAVE file: " + path; return false;
    }
    int fmtTag = 0, bits = 0; bool haveFmt = false;
    size_t pos = 12;
    while (pos + 8 <= buf.size()) {
        uint32_t sz = rd32(&buf[pos + 4]);
        const uint8_t *body = &buf[pos + 8];
        size_t avail = buf.size() - (pos + 8);
        if (std::memcmp(&buf[pos], "fmt ", 4) == 0 && sz >= 16 && avail >= 16) {
            fmtTag = rd16(body); out.channels = rd16(body + 2); out.sampleRate = (int)rd32(body + 4); bits = rd16(body + 14);
            if (fmtTag == 0xFFFE && sz >= 26 && avail >= 26) fmtTag = rd16(body + 24); // WAVE_FORMAT_EXTENSIBLE
            haveFmt = true;
        } else if (std::memcmp(&buf[pos], "data", 4) == 0) {
            if (!haveFmt) { err = "WAV data chunk before fmt chunk: " + path; return false; }
            if (fmtTag != 1 || bits != 16 || out.channels < 1) { err = "Only 16-bit PCM WAV is supported: " + path; return false; }
            size_t bytes = sz < avail ? sz : avail;
            size_t count = bytes / 2;
            count -= count % (size_t)out.channels;
            out.samples.resize(count);
            for (size_t i = 0; i < count; ++i) out.samples[i] = (int16_t)rd16(body + 2 * i);
            return true;
        }
        pos += 8 + (size_t)sz + (sz & 1);
    }
    err = "WAV file has no data chunk: " + path;
    return false;
}

bool writeWavMono16(const std::string &path, const int16_t *samples, size_t n, int sampleRate, std::string &err) {
    FILE *f = std::fopen(path.c_str(), "wb");
    if (!f) { err = "Cannot create WAV file: " + path; return false; }
    uint8_t h[44];
    auto w32 = [&](int o, uint32_t v) { h[o] = v & 255; h[o + 1] = (v >> 8) & 255; h[o + 2] = (v >> 16) & 255; h[o + 3] = (v >> 24) & 255; };
    auto w16 = [&](int o, uint16_t v) { h[o] = v & 255; h[o + 1] = (v >> 8) & 255; };
    uint32_t dataBytes = (uint32_t)(n * 2);
    std::memcpy(h, "RIFF", 4); w32(4, 36 + dataBytes); std::memcpy(h + 8, "WAVEfmt ", 8);
    w32(16, 16); w16(20, 1); w16(22, 1); w32(24, (uint32_t)sampleRate); w32(28, (uint32_t)sampleRate * 2); w16(32, 2); w16(34, 16);
    std::memcpy(h + 36, "data", 4); w32(40, dataBytes);
    bool ok = std::fwrite(h, 1, 44, f) == 44 && std::fwrite(samples, 2, n, f) == n;
    std::fclose(f);
    if (!ok) err = "Short write to WAV file: " + path;
    return ok;
}

} // namespace esb
