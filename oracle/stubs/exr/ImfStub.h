// Test-infrastructure stub (NOT product code).
// A minimal stand-in for the slice of the OpenEXR C++ API that the reference's util/image.cpp uses
// (ReadEXR / Image::WriteEXR, image.cpp:1023-1255).  The reference vendors OpenEXR but not Imath (its cmake fetches
// Imath from the network), so the real library cannot be built here.  This stub reads single-part scanline files with
// NONE / ZIPS / ZIP compression and HALF / FLOAT channels, and writes uncompressed scanline files: enough for
// TrBuffer::Load/Store (cpu/trbuffer.h:52-98) and the film output of the trimmed reference build
// (oracle/build_ref_pbrt.sh).  Own code, written against the public EXR file layout.
#pragma once
#include <zlib.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace Imath {
struct V2i { int x = 0, y = 0; V2i() {} V2i(int a, int b) : x(a), y(b) {} };
struct V2f { float x = 0, y = 0; V2f() {} V2f(float a, float b) : x(a), y(b) {} };
struct Box2i { V2i min, max; Box2i() {} Box2i(V2i a, V2i b) : min(a), max(b) {} };
}  // namespace Imath

namespace Imf {
enum PixelType { UINT = 0, HALF = 1, FLOAT = 2 };

struct Slice {
    PixelType type = HALF; char *base = nullptr; size_t xStride = 0, yStride = 0;
    Slice() {}
    Slice(PixelType t, char *b, size_t xs, size_t ys) : type(t), base(b), xStride(xs), yStride(ys) {}
};
struct Channel { PixelType type = HALF; Channel() {} Channel(PixelType t) : type(t) {} };

template <class V> struct NamedMap {
    std::map<std::string, V> m;
    struct Iter {
        typename std::map<std::string, V>::iterator it;
        const char *name() const { return it->first.c_str(); }
        V &slice() const { return it->second; }
        V &channel() const { return it->second; }
        Iter &operator++() { ++it; return *this; }
        bool operator!=(const Iter &o) const { return it != o.it; }
        bool operator==(const Iter &o) const { return it == o.it; }
    };
    Iter begin() const { return Iter{const_cast<std::map<std::string, V> &>(m).begin()}; }
    Iter end() const { return Iter{const_cast<std::map<std::string, V> &>(m).end()}; }
    void insert(const std::string &n, const V &v) { m[n] = v; }
};
using FrameBuffer = NamedMap<Slice>;
using ChannelList = NamedMap<Channel>;

struct Attribute {
    virtual ~Attribute() {}
    virtual const char *typeName() const = 0;
    virtual Attribute *clone() const = 0;
    virtual std::string bytes() const = 0;
};
template <class T> inline std::string podBytes(const T &v) { return std::string((const char *)&v, sizeof(T)); }
#define IMF_STUB_ATTR(Name, T, tname, BYTES)                                        \
    struct Name : Attribute {                                                       \
        T v;                                                                        \
        Name() : v() {}                                                             \
        Name(const T &x) : v(x) {}                                                  \
        const T &value() const { return v; }                                        \
        T &value() { return v; }                                                    \
        static const char *staticTypeName() { return tname; }                       \
        const char *typeName() const override { return tname; }                     \
        Attribute *clone() const override { return new Name(v); }                   \
        std::string bytes() const override { BYTES; }                               \
    };
struct M44f { float m[4][4]; M44f() { std::memset(m, 0, sizeof(m)); } M44f(const float a[4][4]) { std::memcpy(m, a, sizeof(m)); }
              float *operator[](int i) { return m[i]; } const float *operator[](int i) const { return m[i]; }
              const float *getValue() const { return &m[0][0]; } };
struct Chromaticities {
    Imath::V2f red{0.64f, 0.33f}, green{0.3f, 0.6f}, blue{0.15f, 0.06f}, white{0.3127f, 0.329f};
    Chromaticities() {}
    Chromaticities(Imath::V2f r, Imath::V2f g, Imath::V2f b, Imath::V2f w) : red(r), green(g), blue(b), white(w) {}
};
using StringVector = std::vector<std::string>;
IMF_STUB_ATTR(FloatAttribute, float, "float", return podBytes(v))
IMF_STUB_ATTR(IntAttribute, int, "int", return podBytes(v))
IMF_STUB_ATTR(M44fAttribute, M44f, "m44f", return podBytes(v))
IMF_STUB_ATTR(ChromaticitiesAttribute, Chromaticities, "chromaticities", return podBytes(v))
IMF_STUB_ATTR(StringAttribute, std::string, "string", return v)
IMF_STUB_ATTR(StringVectorAttribute, StringVector, "stringvector",
              std::string s; for (auto &e : v) { int n = (int)e.size(); s += podBytes(n); s += e; } return s)

class Header {
  public:
    Header() {}
    Header(const Imath::Box2i &disp, const Imath::Box2i &data) : disp_(disp), data_(data) {}
    Header(const Header &o) : disp_(o.disp_), data_(o.data_), ch_(o.ch_), compression(o.compression) {
        for (auto &a : o.attrs_) attrs_[a.first] = a.second->clone();
    }
    Header &operator=(const Header &) = delete;
    ~Header() { for (auto &a : attrs_) delete a.second; }
    const Imath::Box2i &dataWindow() const { return data_; }
    const Imath::Box2i &displayWindow() const { return disp_; }
    Imath::Box2i &dataWindow() { return data_; }
    Imath::Box2i &displayWindow() { return disp_; }
    ChannelList &channels() { return ch_; }
    const ChannelList &channels() const { return ch_; }
    void insert(const std::string &name, const Attribute &a) {
        auto it = attrs_.find(name);
        if (it != attrs_.end()) delete it->second;
        attrs_[name] = a.clone();
    }
    template <class T> const T *findTypedAttribute(const char *name) const {
        auto it = attrs_.find(name);
        return it == attrs_.end() ? nullptr : dynamic_cast<const T *>(it->second);
    }
    struct ConstIterator {
        std::map<std::string, Attribute *>::const_iterator it;
        const char *name() const { return it->first.c_str(); }
        Attribute &attribute() const { return *it->second; }
        ConstIterator &operator++() { ++it; return *this; }
        bool operator!=(const ConstIterator &o) const { return it != o.it; }
    };
    ConstIterator begin() const { return ConstIterator{attrs_.begin()}; }
    ConstIterator end() const { return ConstIterator{attrs_.end()}; }
    int compression = 0;

  private:
    Imath::Box2i disp_, data_;
    ChannelList ch_;
    std::map<std::string, Attribute *> attrs_;
};

namespace stub {
inline int pixelBytes(PixelType t) { return t == HALF ? 2 : 4; }
struct Reader {
    const std::vector<unsigned char> &b; size_t p = 0;
    explicit Reader(const std::vector<unsigned char> &bb) : b(bb) {}
    void need(size_t n) const { if (p + n > b.size()) throw std::runtime_error("EXR stub: truncated file"); }
    template <class T> T get() { need(sizeof(T)); T v; std::memcpy(&v, &b[p], sizeof(T)); p += sizeof(T); return v; }
    std::string str() { std::string s; while (true) { need(1); char c = (char)b[p++]; if (!c) break; s += c; } return s; }
};
}  // namespace stub

class InputFile {
  public:
    explicit InputFile(const char *name) {
        FILE *f = std::fopen(name, "rb");
        if (!f) throw std::runtime_error(std::string("EXR stub: cannot open ") + name);
        std::fseek(f, 0, SEEK_END); long n = std::ftell(f); std::fseek(f, 0, SEEK_SET);
        buf_.resize(n > 0 ? (size_t)n : 0);
        size_t got = buf_.empty() ? 0 : std::fread(buf_.data(), 1, buf_.size(), f);
        std::fclose(f);
        if (got != buf_.size()) throw std::runtime_error("EXR stub: short read");
        stub::Reader r(buf_);
        if (r.get<uint32_t>() != 20000630u) throw std::runtime_error("EXR stub: bad magic");
        uint32_t ver = r.get<uint32_t>();
        if (ver & 0x1a00u) throw std::runtime_error("EXR stub: tiled / deep / multi-part files are not supported");
        while (true) {
            std::string an = r.str();
            if (an.empty()) break;
            std::string ty = r.str();
            int sz = r.get<int>();
            r.need((size_t)sz);
            size_t start = r.p;
            if (an == "channels") {
                while (true) {
                    std::string cn = r.str();
                    if (cn.empty()) break;
                    int pt = r.get<int>(); r.p += 4; int xs = r.get<int>(), ys = r.get<int>();
                    if (xs != 1 || ys != 1) throw std::runtime_error("EXR stub: subsampled channels are not supported");
                    if (pt != HALF && pt != FLOAT) throw std::runtime_error("EXR stub: UINT channels are not supported");
                    hdr_.channels().insert(cn, Channel((PixelType)pt));
                }
            } else if (an == "compression") hdr_.compression = r.get<unsigned char>();
            else if (an == "dataWindow") { Imath::Box2i b; b.min.x = r.get<int>(); b.min.y = r.get<int>(); b.max.x = r.get<int>(); b.max.y = r.get<int>(); hdr_.dataWindow() = b; }
            else if (an == "displayWindow") { Imath::Box2i b; b.min.x = r.get<int>(); b.min.y = r.get<int>(); b.max.x = r.get<int>(); b.max.y = r.get<int>(); hdr_.displayWindow() = b; }
            else if (ty == "float" && sz == 4) hdr_.insert(an, FloatAttribute(r.get<float>()));
            else if (ty == "int" && sz == 4) hdr_.insert(an, IntAttribute(r.get<int>()));
            else if (ty == "string") hdr_.insert(an, StringAttribute(std::string((const char *)&buf_[start], (size_t)sz)));
            else if (ty == "m44f" && sz == 64) { M44f m; std::memcpy(m.m, &buf_[start], 64); hdr_.insert(an, M44fAttribute(m)); }
            else if (ty == "chromaticities" && sz == 32) { Chromaticities c; std::memcpy(&c, &buf_[start], 32); hdr_.insert(an, ChromaticitiesAttribute(c)); }
            r.p = start + (size_t)sz;
        }
        if (hdr_.compression != 0 && hdr_.compression != 2 && hdr_.compression != 3)
            throw std::runtime_error("EXR stub: only NONE / ZIPS / ZIP compression is supported");
        tablePos_ = r.p;
    }
    const Header &header() const { return hdr_; }
    void setFrameBuffer(const FrameBuffer &fb) { fb_ = fb; }
    void readPixels(int y0, int y1) {
        const Imath::Box2i dw = hdr_.dataWindow();
        const int W = dw.max.x - dw.min.x + 1, H = dw.max.y - dw.min.y + 1;
        const int lpb = hdr_.compression == 3 ? 16 : 1, nblk = (H + lpb - 1) / lpb;
        size_t rowBytes = 0;
        for (auto &c : hdr_.channels().m) rowBytes += (size_t)W * stub::pixelBytes(c.second.type);
        stub::Reader r(buf_); r.p = tablePos_;
        std::vector<uint64_t> off(nblk);
        for (int i = 0; i < nblk; ++i) off[i] = r.get<uint64_t>();
        std::vector<unsigned char> raw, tmp;
        for (int bi = 0; bi < nblk; ++bi) {
            r.p = (size_t)off[bi];
            int by = r.get<int>(), sz = r.get<int>();
            r.need((size_t)sz);
            int rows = std::min(lpb, dw.max.y - by + 1);
            if (by + rows - 1 < y0 || by > y1) continue;
            size_t want = rowBytes * rows;
            raw.resize(want);
            if (hdr_.compression == 0 || (size_t)sz == want) std::memcpy(raw.data(), &buf_[r.p], want);
            else {
                tmp.resize(want); uLongf dl = (uLongf)want;
                if (uncompress(tmp.data(), &dl, &buf_[r.p], (uLong)sz) != Z_OK || dl != want) throw std::runtime_error("EXR stub: zlib error");
                for (size_t i = 1; i < want; ++i) tmp[i] = (unsigned char)(tmp[i - 1] + tmp[i] - 128);
                size_t half = (want + 1) / 2;
                for (size_t i = 0; i < want; ++i) raw[i] = (i & 1) ? tmp[half + i / 2] : tmp[i / 2];
            }
            const unsigned char *src = raw.data();
            for (int yy = 0; yy < rows; ++yy) {
                int y = by + yy;
                for (auto &c : hdr_.channels().m) {
                    int pb = stub::pixelBytes(c.second.type);
                    auto it = fb_.m.find(c.first);
                    if (it != fb_.m.end() && y >= y0 && y <= y1) {
                        const Slice &s = it->second;
                        if (s.type != c.second.type) throw std::runtime_error("EXR stub: slice / channel type mismatch");
                        for (int x = 0; x < W; ++x)
                            std::memcpy(s.base + (size_t)(dw.min.x + x) * s.xStride + (size_t)y * s.yStride, src + (size_t)x * pb, pb);
                    }
                    src += (size_t)W * pb;
                }
            }
        }
    }

  private:
    std::vector<unsigned char> buf_;
    Header hdr_;
    FrameBuffer fb_;
    size_t tablePos_ = 0;
};

class OutputFile {
  public:
    OutputFile(const char *name, const Header &h) : name_(name), hdr_(h) {}
    void setFrameBuffer(const FrameBuffer &fb) { fb_ = fb; }
    void writePixels(int nLines) {
        const Imath::Box2i dw = hdr_.dataWindow();
        const int W = dw.max.x - dw.min.x + 1, H = dw.max.y - dw.min.y + 1;
        if (nLines != H) throw std::runtime_error("EXR stub: writePixels expects the whole image");
        std::string out;
        auto put = [&](const void *p, size_t n) { out.append((const char *)p, n); };
        auto attr = [&](const std::string &n, const std::string &t, const std::string &data) {
            put(n.c_str(), n.size() + 1); put(t.c_str(), t.size() + 1); int sz = (int)data.size(); put(&sz, 4); out += data; };
        uint32_t magic = 20000630u, ver = 2; put(&magic, 4); put(&ver, 4);
        std::string ch;
        for (auto &c : hdr_.channels().m) {
            ch.append(c.first.c_str(), c.first.size() + 1);
            int pt = c.second.type, one = 1, zero = 0; ch.append((char *)&pt, 4); ch.append((char *)&zero, 4); ch.append((char *)&one, 4); ch.append((char *)&one, 4);
        }
        ch.push_back('\0');
        attr("channels", "chlist", ch);
        attr("compression", "compression", std::string(1, '\0'));
        int dwv[4] = {dw.min.x, dw.min.y, dw.max.x, dw.max.y};
        const Imath::Box2i ds = hdr_.displayWindow(); int dsv[4] = {ds.min.x, ds.min.y, ds.max.x, ds.max.y};
        attr("dataWindow", "box2i", std::string((char *)dwv, 16));
        attr("displayWindow", "box2i", std::string((char *)dsv, 16));
        attr("lineOrder", "lineOrder", std::string(1, '\0'));
        float par = 1.f, swc[2] = {0.f, 0.f}, sww = 1.f;
        attr("pixelAspectRatio", "float", podBytes(par));
        attr("screenWindowCenter", "v2f", std::string((char *)swc, 8));
        attr("screenWindowWidth", "float", podBytes(sww));
        for (auto it = hdr_.begin(); it != hdr_.end(); ++it) attr(it.name(), it.attribute().typeName(), it.attribute().bytes());
        out.push_back('\0');
        size_t rowBytes = 0;
        for (auto &c : hdr_.channels().m) rowBytes += (size_t)W * stub::pixelBytes(c.second.type);
        uint64_t pos = out.size() + (uint64_t)H * 8;
        for (int y = 0; y < H; ++y) { put(&pos, 8); pos += 8 + rowBytes; }
        for (int yy = 0; yy < H; ++yy) {
            int y = dw.min.y + yy, sz = (int)rowBytes; put(&y, 4); put(&sz, 4);
            for (auto &c : hdr_.channels().m) {
                int pb = stub::pixelBytes(c.second.type);
                const Slice &s = fb_.m.at(c.first);
                for (int x = 0; x < W; ++x) put(s.base + (size_t)(dw.min.x + x) * s.xStride + (size_t)y * s.yStride, pb);
            }
        }
        FILE *f = std::fopen(name_.c_str(), "wb");
        if (!f) throw std::runtime_error("EXR stub: cannot open " + name_ + " for writing");
        bool ok = std::fwrite(out.data(), 1, out.size(), f) == out.size();
        ok = (std::fclose(f) == 0) && ok;
        if (!ok) throw std::runtime_error("EXR stub: write failed");
    }

  private:
    std::string name_;
    Header hdr_;
    FrameBuffer fb_;
};
}  // namespace Imf
