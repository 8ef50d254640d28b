// Minimal OpenEXR scanline I/O for the three image files on the path (util/image.cpp subset):
//   read : the image-space guiding buffer (tinyexr, uncompressed FLOAT, multi-layer, layer VSP.Y --
//          openpgl/imagespace/ImageSpaceGuidingBuffer.h:290-293) and the transmittance buffer (OpenEXR, ZIP, FLOAT,
//          channels Transmittance.{R,G,B} -- cpu/trbuffer.h:75-98)
//   write: the film (R,G,B; HALF unless savefp16=false; ZIP) -- film.cpp RGBFilm::WriteImage -> util/image.cpp
// Supported: single-part scanline files, compression NONE / ZIPS / ZIP, pixel types HALF / FLOAT / UINT(read as float).
#include <zlib.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/vspg.h"

namespace {

float half_to_float(uint16_t h) {
    uint32_t s = (h >> 15) & 1, e = (h >> 10) & 0x1f, m = h & 0x3ff, out;
    if (e == 0) {
        if (m == 0) out = s << 31;
        else {  // subnormal
            int sh = 0;
            while (!(m & 0x400)) { m <<= 1; sh++; }
            m &= 0x3ff;
            out = (s << 31) | ((uint32_t)(127 - 15 - sh + 1) << 23) | (m << 13);
        }
    } else if (e == 31) out = (s << 31) | 0x7f800000u | (m << 13);
    else out = (s << 31) | ((e - 15 + 127) << 23) | (m << 13);
    float f; memcpy(&f, &out, 4); return f;
}
uint16_t float_to_half(float f) {  // round to nearest even (util/float.h Half ctor semantics)
    uint32_t x; memcpy(&x, &f, 4);
    uint32_t s = (x >> 16) & 0x8000, em = x & 0x7fffffff;
    if (em >= 0x7f800000) return (uint16_t)(s | 0x7c00 | (em > 0x7f800000 ? 0x200 : 0));   // inf / nan
    if (em >= 0x477ff000) return (uint16_t)(s | 0x7c00);                                    // overflow -> inf
    if (em < 0x33000001) return (uint16_t)s;                                                // underflow -> 0
    int e = (int)(em >> 23) - 127 + 15;
    uint32_t m = em & 0x7fffff;
    if (e <= 0) {  // subnormal half
        m |= 0x800000;
        int shift = 14 - e;
        uint32_t r = m >> shift, rem = m & ((1u << shift) - 1), half = 1u << (shift - 1);
        if (rem > half || (rem == half && (r & 1))) r++;
        return (uint16_t)(s | r);
    }
    uint32_t r = ((uint32_t)e << 10) | (m >> 13), rem = m & 0x1fff;
    if (rem > 0x1000 || (rem == 0x1000 && (r & 1))) r++;
    return (uint16_t)(s | r);
}

struct Channel { std::string name; int type; };

struct Cursor {
    const unsigned char *p, *end; bool ok = true;
    bool need(size_t n) { if ((size_t)(end - p) < n) { ok = false; return false; } return true; }
    int32_t i32() { int32_t v = 0; if (need(4)) { memcpy(&v, p, 4); p += 4; } return v; }
    uint64_t u64() { uint64_t v = 0; if (need(8)) { memcpy(&v, p, 8); p += 8; } return v; }
    std::string str() { std::string s; while (p < end && *p) s.push_back((char)*p++); if (p < end) p++; else ok = false; return s; }
};

void set_err(char *err, size_t n, const std::string &s) { if (err && n) snprintf(err, n, "%s", s.c_str()); }

// ZIP block post-processing (ImfZip.cpp): undo the byte-delta predictor, then the even/odd byte split
void unpredict_and_interleave(std::vector<unsigned char> &buf) {
    size_t n = buf.size();
    for (size_t i = 1; i < n; i++) buf[i] = (unsigned char)(buf[i - 1] + buf[i] - 128);
    std::vector<unsigned char> out(n);
    const unsigned char *t1 = buf.data(), *t2 = buf.data() + (n + 1) / 2;
    for (size_t i = 0, a = 0, b = 0; i < n; i++) out[i] = (i & 1) ? t2[b++] : t1[a++];
    buf.swap(out);
}
void split_and_predict(std::vector<unsigned char> &buf) {
    size_t n = buf.size();
    std::vector<unsigned char> t(n);
    unsigned char *t1 = t.data(), *t2 = t.data() + (n + 1) / 2;
    for (size_t i = 0; i < n; i++) { if (i & 1) *t2++ = buf[i]; else *t1++ = buf[i]; }
    unsigned char prev = t[0];
    for (size_t i = 1; i < n; i++) { unsigned char cur = t[i]; t[i] = (unsigned char)(cur - prev + 128 + 256); prev = cur; }
    buf.swap(t);
}

}  // namespace

extern "C" {

void vspg_free(void *p) { free(p); }

int vspg_exr_read_channel(const char *path, const char *channel, float **out, int *w, int *h, char *err, size_t err_len) {
    if (!path || !channel || !out || !w || !h) return VSPG_ERR_INVALID;
    *out = nullptr;
    FILE *fp = fopen(path, "rb");
    if (!fp) { set_err(err, err_len, std::string("cannot open '") + path + "'"); return VSPG_ERR_IO; }
    std::vector<unsigned char> file;
    fseek(fp, 0, SEEK_END); long sz = ftell(fp); fseek(fp, 0, SEEK_SET);
    file.resize(sz > 0 ? (size_t)sz : 0);
    size_t got = fread(file.data(), 1, file.size(), fp);
    fclose(fp);
    if (got != file.size() || file.size() < 8) { set_err(err, err_len, "short read"); return VSPG_ERR_IO; }
    Cursor c{file.data(), file.data() + file.size()};
    if (c.i32() != 20000630) { set_err(err, err_len, std::string("'") + path + "' is not an OpenEXR file"); return VSPG_ERR_IO; }
    int version = c.i32();
    if (version & 0x1a00) { set_err(err, err_len, "tiled / multi-part / deep EXR files are not supported"); return VSPG_ERR_UNSUPPORTED; }
    std::vector<Channel> chans;
    int compression = -1, dw[4] = {0, 0, -1, -1};
    while (c.ok) {
        std::string name = c.str();
        if (name.empty()) break;
        std::string type = c.str();
        int size = c.i32();
        if (!c.need((size_t)std::max(size, 0))) break;
        const unsigned char *v = c.p;
        if (name == "channels") {
            Cursor cc{v, v + size};
            while (cc.ok) {
                std::string cn = cc.str();
                if (cn.empty()) break;
                Channel ch; ch.name = cn; ch.type = cc.i32();
                cc.need(4); cc.p += 4;                       // pLinear + reserved
                int xs = cc.i32(), ys = cc.i32();
                if (xs != 1 || ys != 1) { set_err(err, err_len, "subsampled channels are not supported"); return VSPG_ERR_UNSUPPORTED; }
                chans.push_back(ch);
            }
        } else if (name == "compression") compression = v[0];
        else if (name == "dataWindow") memcpy(dw, v, 16);
        c.p += size;
    }
    if (!c.ok || chans.empty() || compression < 0 || dw[2] < dw[0] || dw[3] < dw[1]) { set_err(err, err_len, "malformed EXR header"); return VSPG_ERR_IO; }
    if (compression != 0 && compression != 2 && compression != 3) { set_err(err, err_len, "EXR compression other than NONE/ZIPS/ZIP is not supported"); return VSPG_ERR_UNSUPPORTED; }
    const int W = dw[2] - dw[0] + 1, H = dw[3] - dw[1] + 1;
    int target = -1; size_t lineBytes = 0, targetOff = 0;
    for (size_t i = 0; i < chans.size(); i++) {
        size_t b = chans[i].type == 1 ? 2 : 4;
        if (chans[i].name == channel) { target = (int)i; targetOff = lineBytes; }
        lineBytes += b * W;
    }
    if (target < 0) { set_err(err, err_len, std::string("channel '") + channel + "' not found in '" + path + "'"); return VSPG_ERR_IO; }
    const int linesPerBlock = compression == 3 ? 16 : 1;
    const int nBlocks = (H + linesPerBlock - 1) / linesPerBlock;
    std::vector<uint64_t> offsets(nBlocks);
    for (int i = 0; i < nBlocks; i++) offsets[i] = c.u64();
    if (!c.ok) { set_err(err, err_len, "truncated offset table"); return VSPG_ERR_IO; }
    float *img = (float *)malloc(sizeof(float) * (size_t)W * H);
    if (!img) return VSPG_ERR_INVALID;
    std::vector<unsigned char> raw;
    for (int b = 0; b < nBlocks; b++) {
        if (offsets[b] + 8 > file.size()) { free(img); set_err(err, err_len, "chunk offset outside the file"); return VSPG_ERR_IO; }
        Cursor k{file.data() + offsets[b], file.data() + file.size()};
        int y = k.i32(), dataSize = k.i32();
        int lines = std::min(linesPerBlock, dw[3] - y + 1);
        if (!k.ok || dataSize < 0 || !k.need((size_t)dataSize) || y < dw[1] || lines <= 0) { free(img); set_err(err, err_len, "malformed chunk"); return VSPG_ERR_IO; }
        size_t rawSize = lineBytes * (size_t)lines;
        raw.assign(rawSize, 0);
        if (compression == 0 || (size_t)dataSize == rawSize) memcpy(raw.data(), k.p, std::min(rawSize, (size_t)dataSize));
        else {
            uLongf dst = (uLongf)rawSize;
            if (uncompress(raw.data(), &dst, k.p, (uLong)dataSize) != Z_OK || dst != rawSize) { free(img); set_err(err, err_len, "zlib error in chunk"); return VSPG_ERR_IO; }
            unpredict_and_interleave(raw);
        }
        for (int l = 0; l < lines; l++) {
            const unsigned char *src = raw.data() + lineBytes * l + targetOff;
            float *dst = img + (size_t)(y - dw[1] + l) * W;
            int t = chans[target].type;
            for (int x = 0; x < W; x++) {
                if (t == 1) { uint16_t hv; memcpy(&hv, src + 2 * x, 2); dst[x] = half_to_float(hv); }
                else if (t == 2) memcpy(&dst[x], src + 4 * x, 4);
                else { uint32_t u; memcpy(&u, src + 4 * x, 4); dst[x] = (float)u; }
            }
        }
    }
    *out = img; *w = W; *h = H;
    return VSPG_OK;
}

namespace {
struct ExrChan { std::string name; const float *base; size_t stride; };   // value(x, y) = base[stride * (y * w + x)]
}
static int write_exr_channels(const char *path, std::vector<ExrChan> chans, int w, int h, int as_half);

int vspg_exr_write_rgb(const char *path, const float *rgb, int w, int h, int as_half) {
    if (!path || !rgb || w <= 0 || h <= 0) return VSPG_ERR_INVALID;
    return write_exr_channels(path, {{"R", rgb, 3}, {"G", rgb + 1, 3}, {"B", rgb + 2, 3}}, w, h, as_half);
}

int vspg_exr_write_channels(const char *path, int w, int h, int n, const char *const *names, const float *const *planes) {
    if (!path || w <= 0 || h <= 0 || n <= 0 || !names || !planes) return VSPG_ERR_INVALID;
    std::vector<ExrChan> chans;
    for (int i = 0; i < n; i++) {
        if (!names[i] || !names[i][0] || !planes[i]) return VSPG_ERR_INVALID;
        chans.push_back({names[i], planes[i], 1});
    }
    return write_exr_channels(path, chans, w, h, 0);
}

}  // extern "C"

static int write_exr_channels(const char *path, std::vector<ExrChan> chans, int w, int h, int as_half) {
    std::sort(chans.begin(), chans.end(), [](const ExrChan &a, const ExrChan &b) { return a.name < b.name; });   // chlist is alphabetical
    const int nch = (int)chans.size();
    std::vector<unsigned char> hdr;
    auto put = [&](const void *p, size_t n) { const unsigned char *q = (const unsigned char *)p; hdr.insert(hdr.end(), q, q + n); };
    auto puts0 = [&](const char *s) { put(s, strlen(s) + 1); };
    auto puti = [&](int32_t v) { put(&v, 4); };
    auto putf = [&](float v) { put(&v, 4); };
    puti(20000630); puti(2);
    const int ptype = as_half ? 1 : 2;
    {
        std::vector<unsigned char> cl;
        for (const ExrChan &c : chans) {
            cl.insert(cl.end(), c.name.begin(), c.name.end()); cl.push_back(0);
            int32_t v = ptype; cl.insert(cl.end(), (unsigned char *)&v, (unsigned char *)&v + 4);
            cl.insert(cl.end(), {0, 0, 0, 0});
            v = 1; cl.insert(cl.end(), (unsigned char *)&v, (unsigned char *)&v + 4); cl.insert(cl.end(), (unsigned char *)&v, (unsigned char *)&v + 4);
        }
        cl.push_back(0);
        puts0("channels"); puts0("chlist"); puti((int)cl.size()); put(cl.data(), cl.size());
    }
    puts0("compression"); puts0("compression"); puti(1); { unsigned char z = 3; put(&z, 1); }   // ZIP
    int box[4] = {0, 0, w - 1, h - 1};
    puts0("dataWindow"); puts0("box2i"); puti(16); put(box, 16);
    puts0("displayWindow"); puts0("box2i"); puti(16); put(box, 16);
    puts0("lineOrder"); puts0("lineOrder"); puti(1); { unsigned char z = 0; put(&z, 1); }
    puts0("pixelAspectRatio"); puts0("float"); puti(4); putf(1.f);
    puts0("screenWindowCenter"); puts0("v2f"); puti(8); putf(0.f); putf(0.f);
    puts0("screenWindowWidth"); puts0("float"); puti(4); putf(1.f);
    { unsigned char z = 0; put(&z, 1); }
    const int nBlocks = (h + 15) / 16;
    const size_t bpp = as_half ? 2 : 4, lineBytes = nch * bpp * w;
    std::vector<std::vector<unsigned char>> chunks(nBlocks);
    for (int b = 0; b < nBlocks; b++) {
        int y0 = b * 16, lines = std::min(16, h - y0);
        std::vector<unsigned char> raw(lineBytes * lines);
        for (int l = 0; l < lines; l++)
            for (int ch = 0; ch < nch; ch++) {
                unsigned char *dst = raw.data() + lineBytes * l + (size_t)ch * bpp * w;
                for (int x = 0; x < w; x++) {
                    float v = chans[ch].base[chans[ch].stride * ((size_t)(y0 + l) * w + x)];
                    if (as_half) { uint16_t hv = float_to_half(v); memcpy(dst + 2 * x, &hv, 2); }
                    else memcpy(dst + 4 * x, &v, 4);
                }
            }
        std::vector<unsigned char> pre = raw;
        split_and_predict(pre);
        uLongf bound = compressBound((uLong)pre.size());
        std::vector<unsigned char> z(bound);
        if (compress2(z.data(), &bound, pre.data(), (uLong)pre.size(), Z_DEFAULT_COMPRESSION) != Z_OK) return VSPG_ERR_IO;
        std::vector<unsigned char> &c = chunks[b];
        int32_t y = y0, n = bound < raw.size() ? (int32_t)bound : (int32_t)raw.size();
        c.insert(c.end(), (unsigned char *)&y, (unsigned char *)&y + 4);
        c.insert(c.end(), (unsigned char *)&n, (unsigned char *)&n + 4);
        if (bound < raw.size()) c.insert(c.end(), z.begin(), z.begin() + bound); else c.insert(c.end(), raw.begin(), raw.end());
    }
    FILE *fp = fopen(path, "wb");
    if (!fp) return VSPG_ERR_IO;
    bool ok = fwrite(hdr.data(), 1, hdr.size(), fp) == hdr.size();
    uint64_t off = hdr.size() + 8ull * nBlocks;
    for (int b = 0; b < nBlocks && ok; b++) { ok = fwrite(&off, 8, 1, fp) == 1; off += chunks[b].size(); }
    for (int b = 0; b < nBlocks && ok; b++) ok = fwrite(chunks[b].data(), 1, chunks[b].size(), fp) == chunks[b].size();
    ok = (fclose(fp) == 0) && ok;
    return ok ? VSPG_OK : VSPG_ERR_IO;
}
