// .pbrt front end for the guidedvolpath(VSPG) render pass: tokenizer + directive interpreter for the subset of the
// scene-description language the hot path consumes, producing the POD descs of include/vspg.h.
//
// Follows the *behaviour* of the reference's parser / scene builder, not its structure:
//   directive grammar, parameter lists, "type name" declarations       src/pbrt/parser.cpp:200-1100, paramdict.cpp:30-420
//   unused-parameter => ErrorExit                                      src/pbrt/paramdict.cpp:642-664
//   graphics state, CTM, named coordinate systems, attribute stack    src/pbrt/scene.cpp:93-620 (BasicSceneBuilder)
//   render space "cameraworld" = world translated by -cameraPosition   src/pbrt/cameras.cpp:29-66 (CameraTransform)
//   perspective camera, screen window, raster transforms              src/pbrt/cameras.h:180-260, cameras.cpp:340-410
//   film / filter / sampler parameters                                src/pbrt/film.cpp:200-260,  filters.cpp, samplers.cpp
//   media creation parameters                                         src/pbrt/media.cpp:160-216, 272-361, 411-484
//   lights (RGB build: SpectrumToPhotometric == 1)                     src/pbrt/lights.cpp:281-345, 1595-1760; util/spectrum.cpp:37-51
//   integrator parameters                                             src/pbrt/cpu/guidedvolpathvspgintegrator.cpp:1260-1322
// One process-wide difference, stated in the header: a requested store*/training run, or a load* whose file is
// missing, is a hard error here (the reference would fall back to training a fresh cache on the CPU).
//
// Transforms are float (m, mInv) pairs composed with the reference's own compensated arithmetic (host_xform.h), so
// camera matrices, render-space vertices and medium transforms equal the reference's bit for bit
// (tests/test_ref_pbrt.py pins them against the reference's own objects).
#include <cctype>
#include <cerrno>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/vspg.h"
#include "host_nvdb.h"
#include "host_ply.h"
#include "host_xform.h"

namespace {

struct ParseError : std::runtime_error {
    int code;
    ParseError(int c, const std::string &m) : std::runtime_error(m), code(c) {}
};

// ------------------------------------------------------------------------------------------------ transforms
using xf::XF;
using xf::M4f;
vspg_transform to_abi(const XF &t) {
    vspg_transform r;
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) { r.m[i][j] = t.m.m[i][j]; r.mInv[i][j] = t.inv.m[i][j]; }
    return r;
}

// ------------------------------------------------------------------------------------------------ tokens
struct Tok { enum K { END, WORD, STR, LB, RB } kind = END; std::string s; int line = 0; };

struct Lexer {
    std::string text, file;
    size_t pos = 0;
    int line = 1;
    Tok peeked; bool has_peek = false;
    Tok next() {
        if (has_peek) { has_peek = false; return peeked; }
        for (;;) {
            while (pos < text.size() && isspace((unsigned char)text[pos])) { if (text[pos] == '\n') line++; pos++; }
            if (pos < text.size() && text[pos] == '#') { while (pos < text.size() && text[pos] != '\n') pos++; continue; }
            break;
        }
        Tok t; t.line = line;
        if (pos >= text.size()) return t;
        char c = text[pos];
        if (c == '[') { pos++; t.kind = Tok::LB; return t; }
        if (c == ']') { pos++; t.kind = Tok::RB; return t; }
        if (c == '"') {
            size_t e = ++pos;
            while (e < text.size() && text[e] != '"') { if (text[e] == '\n') fail("unterminated string"); e++; }
            if (e >= text.size()) fail("unterminated string");
            t.kind = Tok::STR; t.s = text.substr(pos, e - pos); pos = e + 1; return t;
        }
        size_t e = pos;
        while (e < text.size() && !isspace((unsigned char)text[e]) && text[e] != '[' && text[e] != ']' && text[e] != '"' && text[e] != '#') e++;
        t.kind = Tok::WORD; t.s = text.substr(pos, e - pos); pos = e; return t;
    }
    const Tok &peek() { if (!has_peek) { peeked = next(); has_peek = true; } return peeked; }
    [[noreturn]] void fail(const std::string &m, int code = VSPG_ERR_INVALID) const {
        throw ParseError(code, file + ":" + std::to_string(line) + ": " + m);
    }
};

// ------------------------------------------------------------------------------------------------ parameters
struct Param {
    std::string type, name;
    std::vector<double> nums;          // numeric payload (floats, ints, points, rgb)
    std::vector<std::string> strs;     // string / bool / texture payload
    mutable bool used = false;
    int line = 0;
};

struct Params {
    std::vector<Param> list;
    std::string where;  // "Integrator \"guidedvolpath\"" for messages
    const Param *find(const std::string &name, std::initializer_list<const char *> types) const {
        for (const Param &p : list)
            if (p.name == name)
                for (const char *t : types)
                    if (p.type == t) { p.used = true; return &p; }
        return nullptr;
    }
    double one_float(const std::string &n, double def) const {
        const Param *p = find(n, {"float"});
        if (!p) return def;
        if (p->nums.size() != 1) throw ParseError(VSPG_ERR_INVALID, where + ": \"" + n + "\" expects a single value");
        return p->nums[0];
    }
    int one_int(const std::string &n, int def) const {
        const Param *p = find(n, {"integer"});
        if (!p) return def;
        if (p->nums.size() != 1) throw ParseError(VSPG_ERR_INVALID, where + ": \"" + n + "\" expects a single value");
        return (int)p->nums[0];
    }
    bool one_bool(const std::string &n, bool def) const {
        const Param *p = find(n, {"bool"});
        if (!p) return def;
        if (p->strs.size() != 1) throw ParseError(VSPG_ERR_INVALID, where + ": \"" + n + "\" expects a single value");
        return p->strs[0] == "true";
    }
    std::string one_string(const std::string &n, const std::string &def) const {
        const Param *p = find(n, {"string"});
        if (!p) return def;
        if (p->strs.size() != 1) throw ParseError(VSPG_ERR_INVALID, where + ": \"" + n + "\" expects a single value");
        return p->strs[0];
    }
    bool point3(const std::string &n, double out[3]) const {
        const Param *p = find(n, {"point3", "point", "vector3", "vector", "normal", "normal3"});
        if (!p) return false;
        if (p->nums.size() != 3) throw ParseError(VSPG_ERR_INVALID, where + ": \"" + n + "\" expects 3 values");
        for (int i = 0; i < 3; i++) out[i] = p->nums[i];
        return true;
    }
    // GetOneSpectrum in the RGB build: "rgb" triple, or a single "float" (constant spectrum).  Named / sampled
    // spectra and blackbody need the spectral tables and are outside the RGB hot path.
    bool rgb(const std::string &n, float out[3]) const {
        if (const Param *p = find(n, {"rgb"})) {
            if (p->nums.size() != 3) throw ParseError(VSPG_ERR_INVALID, where + ": rgb \"" + n + "\" expects 3 values");
            for (int i = 0; i < 3; i++) out[i] = (float)p->nums[i];
            return true;
        }
        if (const Param *p = find(n, {"float"})) {
            if (p->nums.size() != 1) throw ParseError(VSPG_ERR_INVALID, where + ": \"" + n + "\" expects a single value");
            out[0] = out[1] = out[2] = (float)p->nums[0];
            return true;
        }
        if (find(n, {"spectrum", "blackbody"}))
            throw ParseError(VSPG_ERR_UNSUPPORTED, where + ": \"" + n + "\": sampled/named/blackbody spectra are outside the RGB render path; give an rgb value");
        if (find(n, {"texture"}))
            throw ParseError(VSPG_ERR_UNSUPPORTED, where + ": \"" + n + "\": textures are outside the hot-path scope (SURVEY 8f3)");
        return false;
    }
    const std::vector<double> *floats(const std::string &n) const { const Param *p = find(n, {"float"}); return p ? &p->nums : nullptr; }
    const std::vector<double> *ints(const std::string &n) const { const Param *p = find(n, {"integer"}); return p ? &p->nums : nullptr; }
    const std::vector<double> *points(const std::string &n) const { const Param *p = find(n, {"point3", "point"}); return p ? &p->nums : nullptr; }
    const std::vector<double> *rgbs(const std::string &n) const { const Param *p = find(n, {"rgb"}); return p ? &p->nums : nullptr; }
    void report_unused() const {  // paramdict.cpp:642-664
        for (const Param &p : list)
            if (!p.used) throw ParseError(VSPG_ERR_INVALID, where + ": parameter \"" + p.name + "\" is unused (line " + std::to_string(p.line) + ")");
    }
};

Params parse_params(Lexer &lx, const std::string &where) {
    Params ps; ps.where = where;
    while (lx.peek().kind == Tok::STR) {
        Tok decl = lx.next();
        Param p; p.line = decl.line;
        std::istringstream ss(decl.s);
        if (!(ss >> p.type >> p.name)) lx.fail("parameter declaration \"" + decl.s + "\" must be \"type name\"");
        static const char *known[] = {"integer", "float", "point2", "vector2", "point3", "point", "vector3", "vector", "normal", "normal3",
                                      "rgb", "spectrum", "blackbody", "bool", "string", "texture"};
        bool ok = false;
        for (const char *k : known) ok |= p.type == k;
        if (!ok) lx.fail("unknown parameter type \"" + p.type + "\"");
        bool stringy = p.type == "bool" || p.type == "string" || p.type == "texture";
        auto take = [&](const Tok &t) {
            if (t.kind == Tok::STR) { if (!stringy && p.type != "spectrum") lx.fail("\"" + p.name + "\": string value for a numeric parameter"); p.strs.push_back(t.s); }
            else if (t.kind == Tok::WORD) {
                if (p.type == "bool") { if (t.s != "true" && t.s != "false") lx.fail("\"" + p.name + "\": bool must be true or false"); p.strs.push_back(t.s); }
                else {
                    if (stringy) lx.fail("\"" + p.name + "\": expected a quoted string");
                    errno = 0; char *end = nullptr;
                    double v = strtod(t.s.c_str(), &end);
                    if (end == t.s.c_str() || *end) lx.fail("\"" + p.name + "\": \"" + t.s + "\" is not a number");
                    p.nums.push_back(v);
                }
            } else lx.fail("\"" + p.name + "\": missing value");
        };
        if (lx.peek().kind == Tok::LB) {
            lx.next();
            if (!stringy && p.type != "spectrum") {  // fast path for big numeric arrays (density grids)
                for (;;) {
                    // scan numbers straight from the text
                    while (lx.pos < lx.text.size()) {
                        char c = lx.text[lx.pos];
                        if (c == '\n') { lx.line++; lx.pos++; }
                        else if (isspace((unsigned char)c)) lx.pos++;
                        else if (c == '#') { while (lx.pos < lx.text.size() && lx.text[lx.pos] != '\n') lx.pos++; }
                        else break;
                    }
                    if (lx.pos >= lx.text.size()) lx.fail("\"" + p.name + "\": missing ]");
                    if (lx.text[lx.pos] == ']') { lx.pos++; break; }
                    const char *b = lx.text.c_str() + lx.pos; char *end = nullptr;
                    double v = strtod(b, &end);
                    if (end == b) lx.fail("\"" + p.name + "\": expected a number");
                    p.nums.push_back(v);
                    lx.pos += end - b;
                }
            } else {
                for (;;) {
                    Tok t = lx.next();
                    if (t.kind == Tok::RB) break;
                    if (t.kind == Tok::END) lx.fail("\"" + p.name + "\": missing ]");
                    take(t);
                }
            }
        } else take(lx.next());
        if (p.type == "integer") for (double v : p.nums) if (v != std::floor(v)) lx.fail("\"" + p.name + "\": integer parameter has a fractional value");
        ps.list.push_back(std::move(p));
    }
    return ps;
}

// ------------------------------------------------------------------------------------------------ scene state
struct GState {
    XF ctm;
    bool reverse = false;
    int material = 0;          // index into materials; -1 interface
    bool area = false; float area_L[3] = {1, 1, 1}; float area_scale = 1; bool area_two = false; double area_power = -1;
    std::string inside, outside;
    bool has_mi = false;       // a MediumInterface directive is in effect
};

std::string dirname_of(const std::string &p) { size_t s = p.find_last_of('/'); return s == std::string::npos ? "." : p.substr(0, s); }
std::string read_file(const std::string &p) {
    std::ifstream f(p, std::ios::binary);
    if (!f) throw ParseError(VSPG_ERR_IO, "cannot open '" + p + "'");
    std::ostringstream ss; ss << f.rdbuf(); return ss.str();
}

}  // namespace

struct vspg_host_scene {
    vspg_scene_desc desc{};
    vspg_guiding_desc guiding{};
    bool has_guiding = false;
    vspg_integrator_params ip{};
    int spp = 16, seed = 0;
    std::string outfile = "pbrt.exr", base_dir = ".";
    std::vector<vspg_medium> media; std::vector<vspg_quad> quads; std::vector<vspg_triangle> tris; std::vector<vspg_sphere> spheres;
    std::vector<vspg_material> materials; std::vector<vspg_light> lights;
    std::vector<std::unique_ptr<std::vector<float>>> arrays;   // grids owned by the handle
    vspg_host_field *field = nullptr; float *isg = nullptr, *tr = nullptr;
    // builder state
    std::map<std::string, int> medium_ids, named_materials;
    std::map<std::string, XF> named_cs;
    std::vector<GState> stack; GState gs;
    XF world_from_camera, render_from_world; float eye[3] = {0, 0, 0};
    std::string camera_medium;
    bool have_camera = false, world = false, have_integrator = false;
    float fov = 90, lens_radius = 0, focal_distance = 1e6f, frame_aspect = -1; std::vector<double> screen_window;
    int xres = 1280, yres = 720; float iso = 100, exposure = 1, max_comp = INFINITY;
    int filter_type = 1; float filter_r[2] = {1.5f, 1.5f}, filter_sigma = 0.5f;   // pbrt's default filter is "gaussian"
    std::string guiding_cache_file, isg_file, tr_file; bool load_cache = false, load_isg = false, load_tr = false, store_tr = false, store_isg = false, train_isg = false;
    std::string store_tr_path, store_isg_path;
    ~vspg_host_scene() { if (field) vspg_host_field_free(field); if (isg) vspg_free(isg); if (tr) vspg_free(tr); }
};

namespace {

struct Builder {
    vspg_host_scene &S;
    explicit Builder(vspg_host_scene &s) : S(s) {}

    std::string resolve(const std::string &f) const { return (!f.empty() && f[0] == '/') ? f : S.base_dir + "/" + f; }
    // scene.h:477-479: Transform((renderFromWorld * ctm).GetMatrix()) -- the inverse is re-derived from the product
    XF render_from_object() const { return XF(xf::mul(S.render_from_world.m, S.gs.ctm.m)); }
    int medium_id(const std::string &name, const Lexer &lx) const {
        if (name.empty()) return -1;
        auto it = S.medium_ids.find(name);
        if (it == S.medium_ids.end()) lx.fail("medium \"" + name + "\" is not defined");
        return it->second;
    }
    void to_render(const XF &rfo, const double p[3], float out[3]) const {
        float pf[3] = {(float)p[0], (float)p[1], (float)p[2]};   // the parser hands Float values to the shapes (parser.cpp)
        xf::apply_point(rfo.m, pf, out);
    }

    float *own(std::vector<float> &&v) { S.arrays.emplace_back(new std::vector<float>(std::move(v))); return S.arrays.back()->data(); }

    // ---------------------------------------------------------------- directives
    void camera(const std::string &type, const Params &ps, Lexer &lx) {
        if (type != "perspective") lx.fail("Camera \"" + type + "\": only \"perspective\" is on the supported path", VSPG_ERR_UNSUPPORTED);
        S.fov = (float)ps.one_float("fov", 90);
        S.lens_radius = (float)ps.one_float("lensradius", 0);
        S.focal_distance = (float)ps.one_float("focaldistance", 1e6);
        S.frame_aspect = (float)ps.one_float("frameaspectratio", -1);
        if (const auto *sw = ps.floats("screenwindow")) { if (sw->size() != 4) lx.fail("\"screenwindow\" expects 4 values"); S.screen_window = *sw; }
        ps.one_float("shutteropen", 0); ps.one_float("shutterclose", 1);   // no motion on this path
        ps.report_unused();
        // scene.cpp:139-156 + CameraTransform (cameras.cpp:26-57), rendering space "cameraworld" (the default):
        // worldFromRender = Translate(worldFromCamera(0,0,0)), renderFromWorld = its inverse
        S.world_from_camera = xf::inverse(S.gs.ctm);
        const float o[3] = {0, 0, 0};
        xf::apply_point(S.world_from_camera.m, o, S.eye);
        S.render_from_world = xf::inverse(xf::translate(S.eye[0], S.eye[1], S.eye[2]));
        S.named_cs["camera"] = xf::inverse(S.gs.ctm);
        S.camera_medium = S.gs.outside;
        S.have_camera = true;
    }
    void film(const std::string &type, const Params &ps, Lexer &lx) {
        if (type != "rgb") lx.fail("Film \"" + type + "\": only \"rgb\" is supported", VSPG_ERR_UNSUPPORTED);
        S.xres = ps.one_int("xresolution", 1280); S.yres = ps.one_int("yresolution", 720);
        S.outfile = ps.one_string("filename", "pbrt.exr");
        S.iso = (float)ps.one_float("iso", 100); S.exposure = (float)ps.one_float("exposuretime", 1);
        S.max_comp = (float)ps.one_float("maxcomponentvalue", INFINITY);
        if (ps.one_float("whitebalance", 0) != 0) lx.fail("Film: \"whitebalance\" is not supported", VSPG_ERR_UNSUPPORTED);
        std::string sensor = ps.one_string("sensor", "cie1931");
        if (sensor != "cie1931") lx.fail("Film: only the \"cie1931\" sensor is supported", VSPG_ERR_UNSUPPORTED);
        ps.one_float("diagonal", 35); ps.one_bool("savefp16", true);
        if (ps.find("cropwindow", {"float"}) || ps.find("pixelbounds", {"integer"}))
            lx.fail("Film: crop windows are expressed through vspg_render's pixel window, not the scene file", VSPG_ERR_UNSUPPORTED);
        ps.report_unused();
        if (S.xres <= 0 || S.yres <= 0) lx.fail("Film: resolution must be positive");
    }
    void sampler(const std::string &type, const Params &ps, Lexer &lx) {
        // The render pass draws from IndependentSampler streams (samplers.h:442-476); other samplers are SURVEY 8f3.
        if (type != "independent") lx.fail("Sampler \"" + type + "\": only \"independent\" is on the supported path", VSPG_ERR_UNSUPPORTED);
        S.spp = ps.one_int("pixelsamples", 16);
        S.seed = ps.one_int("seed", S.seed);
        ps.report_unused();
    }
    void pixel_filter(const std::string &type, const Params &ps, Lexer &lx) {
        if (type == "box") {
            S.filter_type = 0;
            S.filter_r[0] = (float)ps.one_float("xradius", 0.5); S.filter_r[1] = (float)ps.one_float("yradius", 0.5);
        } else if (type == "gaussian") {
            S.filter_type = 1;
            S.filter_r[0] = (float)ps.one_float("xradius", 1.5); S.filter_r[1] = (float)ps.one_float("yradius", 1.5);
            S.filter_sigma = (float)ps.one_float("sigma", 0.5);
        } else lx.fail("PixelFilter \"" + type + "\": only \"box\" and \"gaussian\" are supported", VSPG_ERR_UNSUPPORTED);
        ps.report_unused();
    }
    void integrator(const std::string &type, const Params &ps, Lexer &lx) {
        // integrators.cpp:3739-3744: "guidedvolpathvspg" is the VSPG class; "guidedvolpath" (the name README and
        // BASELINE.json use) is routed to the same path.
        if (type != "guidedvolpathvspg" && type != "guidedvolpath")
            lx.fail("Integrator \"" + type + "\": this library implements \"guidedvolpath\" / \"guidedvolpathvspg\" only", VSPG_ERR_UNSUPPORTED);
        vspg_integrator_params &ip = S.ip;
        ip.max_depth = ps.one_int("maxdepth", 5);
        ip.min_rr_depth = ps.one_int("minrrdepth", 1);
        ip.use_nee = ps.one_bool("usenee", true);
        ip.surface_guiding = ps.one_bool("surfaceguiding", true);
        ip.volume_guiding = ps.one_bool("volumeguiding", true);
        ip.surface_guiding_ris = ps.one_string("surfaceguidingtype", "ris") == "mis" ? 0 : 1;
        ip.volume_guiding_ris = ps.one_string("volumeguidingtype", "mis") == "mis" ? 0 : 1;
        bool store_cache = ps.one_bool("storeGuidingCache", false);
        S.load_cache = ps.one_bool("loadGuidingCache", false);
        S.guiding_cache_file = ps.one_string("guidingCacheFileName", "");
        ip.vsp_guiding = ps.one_bool("vspguiding", true);
        ip.vsp_primary_guiding = ps.one_bool("vspprimaryguiding", true);
        ip.vsp_secondary_guiding = ps.one_bool("vspsecondaryguiding", true);
        ip.vsp_mis_ratio = (float)ps.one_float("vspmisratio", 0.5);
        std::string crit = ps.one_string("vspcriterion", "variance");
        if (crit == "Contribution" || crit == "contribution") ip.vsp_criterion_variance = 0;
        else if (crit == "Variance" || crit == "variance") ip.vsp_criterion_variance = 1;
        else lx.fail("Integrator: vspcriterion must be \"variance\" or \"contribution\"");
        std::string meth = ps.one_string("vspsamplingmethod", "resampling");
        if (meth == "Resampling" || meth == "resampling") ip.vsp_sampling_nds = 0;
        else if (meth == "NDS" || meth == "nds") ip.vsp_sampling_nds = 1;
        else lx.fail("Integrator: vspsamplingmethod must be \"resampling\" or \"nds\"");
        ip.collision_probability_bias = ps.one_bool("collisionProbabilityBias", false);
        S.store_isg = ps.one_bool("storeISGBuffer", false);
        S.load_isg = ps.one_bool("loadISGBuffer", false);
        S.isg_file = ps.one_string("isgBufferFileName", "");
        S.store_tr = ps.one_bool("storeTrBuffer", false);
        S.load_tr = ps.one_bool("loadTrBuffer", false);
        S.tr_file = ps.one_string("trBufferFileName", "");
        // guided RR needs OPENPGL_RADIANCE_CACHES, which the reference builds OFF: the flags are inert.
        const bool rrGuiding = ps.one_bool("rrguiding", false);   // surfacerrguiding / volumerrguiding only matter when it is on (vspg.cpp:279-290)
        ps.one_bool("surfacerrguiding", true); ps.one_bool("volumerrguiding", true);
        std::string ls = ps.one_string("lightsampler", "bvh");
        if (ls == "uniform") ip.light_sampler = VSPG_LIGHTSAMPLER_UNIFORM;
        else if (ls == "power") ip.light_sampler = VSPG_LIGHTSAMPLER_POWER;
        else if (ls == "bvh") ip.light_sampler = VSPG_LIGHTSAMPLER_BVH;
        else lx.fail("Integrator: lightsampler \"" + ls + "\" unknown");
        ip.regularize = ps.one_bool("regularize", false);
        ps.report_unused();
        if (rrGuiding)
            lx.fail("Integrator: rrguiding (adjoint-driven Russian roulette from the denoised pixel-contribution estimate, vspg.cpp:279-290) is "
                    "outside the hot-path scope; the throughput-based roulette of the default configuration is built", VSPG_ERR_UNSUPPORTED);
        if (store_cache)
            lx.fail("Integrator: storeGuidingCache requests a training pass of the guiding cache, which is out of scope (the cache is a "
                    "loaded input; storeTrBuffer and storeISGBuffer are supported)", VSPG_ERR_UNSUPPORTED);
        if (S.store_isg) {   // vspg.cpp:161-178, 215-217: the buffer is trained by the render itself and stored afterwards
            if (S.isg_file.empty()) lx.fail("Integrator: storeISGBuffer without isgBufferFileName");
            if (S.load_isg) lx.fail("Integrator: storeISGBuffer together with loadISGBuffer (the reference would only re-write the loaded file)", VSPG_ERR_UNSUPPORTED);
            if (!(ip.vsp_guiding && ip.vsp_primary_guiding))
                lx.fail("Integrator: storeISGBuffer without vspguiding + vspprimaryguiding stores no VSP layer (vspg.cpp:169-176)", VSPG_ERR_UNSUPPORTED);
        }
        if (S.store_tr) {   // vspg.cpp:190-193, 219-221: the Tr buffer is accumulated by the render itself and written afterwards
            if (S.tr_file.empty()) lx.fail("Integrator: storeTrBuffer without trBufferFileName");
            if (S.load_tr) lx.fail("Integrator: storeTrBuffer together with loadTrBuffer (the reference would only re-write the loaded file)", VSPG_ERR_UNSUPPORTED);
        }
        S.have_integrator = true;
    }

    void make_named_medium(const std::string &name, const Params &ps, Lexer &lx) {
        std::string type = ps.one_string("type", "");
        if (type.empty()) lx.fail("MakeNamedMedium \"" + name + "\": no \"string type\" given");
        vspg_medium m{};
        m.bounds_min[0] = m.bounds_min[1] = m.bounds_min[2] = 0; m.bounds_max[0] = m.bounds_max[1] = m.bounds_max[2] = 1;
        m.render_from_medium = to_abi(render_from_object());
        m.maj_res[0] = m.maj_res[1] = m.maj_res[2] = 16;
        m.sigma_scale = 1;
        float scale_ = (float)ps.one_float("scale", 1);
        m.g = (float)ps.one_float("g", 0);
        float Le[3] = {0, 0, 0};
        const bool emits = ps.rgb("Le", Le) && std::max(std::max(Le[0], Le[1]), Le[2]) != 0;   // media.cpp:191, 311: "Le" absent or MaxValue() == 0 -> no emission
        if (emits && type != "homogeneous" && type != "uniformgrid")
            lx.fail("medium emission is built for homogeneous and uniformgrid media only (SURVEY 8f3)", VSPG_ERR_UNSUPPORTED);
        if (type != "uniformgrid" && type != "nanovdb") {
            float LeScale = (float)ps.one_float("Lescale", 1);
            if (emits) for (int i = 0; i < 3; i++) m.Le[i] = Le[i] * LeScale;   // media.h:243 Le_spec.Scale(LeScale) (photometric == 1)
        }
        auto grid_dims = [&](size_t n) {
            m.nx = ps.one_int("nx", 1); m.ny = ps.one_int("ny", 1); m.nz = ps.one_int("nz", 1);
            if ((size_t)m.nx * m.ny * m.nz != n)
                lx.fail("medium \"" + name + "\" has " + std::to_string(n) + " density values; expected nx*ny*nz = " + std::to_string((size_t)m.nx * m.ny * m.nz));
            double p0[3] = {0, 0, 0}, p1[3] = {1, 1, 1};
            ps.point3("p0", p0); ps.point3("p1", p1);
            for (int i = 0; i < 3; i++) { m.bounds_min[i] = (float)p0[i]; m.bounds_max[i] = (float)p1[i]; }
        };
        if (type == "homogeneous") {
            m.type = VSPG_MEDIUM_HOMOGENEOUS;
            if (!ps.one_string("preset", "").empty()) lx.fail("medium presets need the spectral tables; give sigma_a / sigma_s as rgb", VSPG_ERR_UNSUPPORTED);
            float sa[3] = {.0011f, .0024f, .014f}, ss[3] = {2.55f, 3.21f, 3.77f};   // media.cpp:178-186 defaults
            ps.rgb("sigma_a", sa); ps.rgb("sigma_s", ss);
            for (int i = 0; i < 3; i++) { m.sigma_a[i] = sa[i] * scale_; m.sigma_s[i] = ss[i] * scale_; }   // media.h:240-242
        } else if (type == "uniformgrid") {
            m.type = VSPG_MEDIUM_GRID;
            std::vector<float> d;
            std::string file = ps.one_string("densityfile", "");   // extension: raw little-endian float32, x fastest
            if (const auto *v = ps.floats("density")) d.assign(v->begin(), v->end());
            else if (!file.empty()) {
                std::string raw = read_file(resolve(file));
                d.resize(raw.size() / 4);
                memcpy(d.data(), raw.data(), d.size() * 4);
            } else lx.fail("No \"density\" value provided for grid medium.");
            if (ps.floats("temperature")) lx.fail("grid medium \"temperature\" (emission) is outside the hot-path scope", VSPG_ERR_UNSUPPORTED);
            grid_dims(d.size());
            {   // media.cpp:304-329: "Lescale" is a grid here (nx*ny*nz values); without one, a 1x1x1 grid holding LeNorm = 1
                const auto *ls = ps.floats("Lescale");
                if (ls && !ls->empty() && ls->size() != d.size())
                    lx.fail("Expected " + std::to_string(m.nx) + " x " + std::to_string(m.ny) + " " + std::to_string(m.nz) + " = " + std::to_string(d.size()) +
                            " values for \"Lescale\" but were given " + std::to_string(ls->size()) + ".");
                if (emits) {
                    for (int i = 0; i < 3; i++) m.Le[i] = Le[i];
                    if (ls && !ls->empty()) { m.le_scale = own(std::vector<float>(ls->begin(), ls->end())); m.le_nx = m.nx; m.le_ny = m.ny; m.le_nz = m.nz; }
                    else { m.le_scale = own(std::vector<float>(1, 1.f)); m.le_nx = m.le_ny = m.le_nz = 1; }
                }
            }
            float sa[3] = {1, 1, 1}, ss[3] = {1, 1, 1};
            ps.rgb("sigma_a", sa); ps.rgb("sigma_s", ss);
            for (int i = 0; i < 3; i++) { m.sigma_a[i] = sa[i] * scale_; m.sigma_s[i] = ss[i] * scale_; }   // media.cpp:249-250
            m.density = own(std::move(d));
        } else if (type == "rgbgrid") {
            m.type = VSPG_MEDIUM_RGBGRID;
            m.sigma_scale = scale_;
            for (int i = 0; i < 3; i++) m.sigma_a[i] = m.sigma_s[i] = 1;
            const auto *sa = ps.rgbs("sigma_a"), *ss = ps.rgbs("sigma_s");
            if (!sa && !ss) lx.fail("rgbgrid medium needs \"sigma_a\" and/or \"sigma_s\" rgb arrays");
            if (sa && ss && sa->size() != ss->size()) lx.fail("rgbgrid: different number of \"sigma_a\" and \"sigma_s\" values");
            size_t n = (sa ? sa->size() : ss->size()) / 3;
            grid_dims(n);
            if (sa) m.sigma_a_rgb = own(std::vector<float>(sa->begin(), sa->end()));
            if (ss) m.sigma_s_rgb = own(std::vector<float>(ss->begin(), ss->end()));
        } else if (type == "nanovdb") {
            make_nanovdb_medium(m, name, scale_, ps, lx);
        } else lx.fail("medium type \"" + type + "\" is outside the hot-path scope (homogeneous, uniformgrid, rgbgrid, nanovdb)", VSPG_ERR_UNSUPPORTED);
        ps.report_unused();
        if (S.medium_ids.count(name)) lx.fail("medium \"" + name + "\" redefined");
        S.medium_ids[name] = (int)S.media.size();
        S.media.push_back(m);
    }

    // media.cpp:683-734 NanoVDBMedium::Create + :549-670 the constructor (bounds, 64^3 majorant grid with the +-1 voxel
    // filter slop).  The sparse tree is flattened over its index bounding box (host_nvdb.cpp); the CUDA path turns that
    // into cell records like any other density grid.
    void make_nanovdb_medium(vspg_medium &m, const std::string &name, float scale_, const Params &ps, Lexer &lx) {
        m.type = VSPG_MEDIUM_NANOVDB;
        std::string filename = ps.one_string("filename", "");
        if (filename.empty()) lx.fail("Must supply \"filename\" to \"nanovdb\" medium.");
        filename = resolve(filename);
        std::string gridname = ps.one_string("gridname", "density");
        std::string temperaturename = ps.one_string("temperaturename", "temperature");
        float LeScale = (float)ps.one_float("Lescale", 1);
        ps.one_float("temperatureoffset", ps.one_float("temperaturecutoff", 0)); ps.one_float("temperaturescale", 1);
        const float densityOffset = (float)ps.one_float("densityoffset", 0), majorantScale = (float)ps.one_float("majorantscale", 1);
        float sa[3] = {1, 1, 1}, ss[3] = {1, 1, 1};
        ps.rgb("sigma_a", sa); ps.rgb("sigma_s", ss);
        for (int i = 0; i < 3; i++) { m.sigma_a[i] = sa[i] * scale_; m.sigma_s[i] = ss[i] * scale_; }   // media.cpp:582-583
        vspg_host::NvdbGrid g;
        std::string err;
        if (!vspg_host::read_nvdb_float_grid(filename, gridname, &g, &err)) lx.fail("nanovdb medium \"" + name + "\": " + err, VSPG_ERR_IO);
        if (LeScale > 0 && vspg_host::nvdb_has_grid(filename, temperaturename))   // media.h:687 IsEmissive
            lx.fail("nanovdb medium \"" + name + "\": the file has a \"" + temperaturename + "\" grid (blackbody emission), which is outside the hot-path scope; pass \"float Lescale\" 0", VSPG_ERR_UNSUPPORTED);
        if (g.nx == 0) lx.fail("nanovdb medium \"" + name + "\": grid \"" + gridname + "\" is empty");
        for (int i = 0; i < 3; i++) {   // media.cpp:585-587: Bounds3f from the double world bounding box
            m.bounds_min[i] = (float)g.world_min[i]; m.bounds_max[i] = (float)g.world_max[i];
            m.index_min[i] = g.bbox_min[i]; m.index_origin[i] = g.vec_f[i];
        }
        for (int i = 0; i < 9; i++) m.index_from_medium[i] = g.inv_mat_f[i];
        m.nx = g.nx; m.ny = g.ny; m.nz = g.nz;
        m.density_offset = densityOffset; m.background = g.background; m.majorant_scale = majorantScale;
        m.maj_res[0] = m.maj_res[1] = m.maj_res[2] = 64;
        // worldToIndexF on a Vec3R (media.cpp:629-632): Map::applyInverseMapF subtracts the float translation in double, and
        // matMult(const float*, Vec3T) (NanoVDB.h:1408-1414) hands the doubles to fmaf(float, float, float)
        auto worldToIndexF = [&](const float w[3], double out[3]) {
            double d[3] = {(double)w[0] - (double)g.vec_f[0], (double)w[1] - (double)g.vec_f[1], (double)w[2] - (double)g.vec_f[2]};
            for (int r = 0; r < 3; r++) {
                const float *mt = g.inv_mat_f + 3 * r;
                out[r] = (double)fmaf((float)d[0], mt[0], fmaf((float)d[1], mt[1], (float)(d[2] * (double)mt[2])));
            }
        };
        auto value = [&](int i, int j, int k) { return g.dense[((size_t)(k - g.bbox_min[2]) * g.ny + (j - g.bbox_min[1])) * g.nx + (i - g.bbox_min[0])]; };
        auto lerp = [](float t, float a, float b) { return (1 - t) * a + t * b; };   // util/math.h:210, Bounds3::Lerp
        std::vector<float> maj((size_t)64 * 64 * 64);
        for (int z = 0; z < 64; z++)
            for (int y = 0; y < 64; y++)
                for (int x = 0; x < 64; x++) {
                    const int c[3] = {x, y, z};
                    float lo[3], hi[3];
                    for (int a = 0; a < 3; a++) {
                        lo[a] = lerp(float(c[a]) / 64, m.bounds_min[a], m.bounds_max[a]);
                        hi[a] = lerp(float(c[a] + 1) / 64, m.bounds_min[a], m.bounds_max[a]);
                    }
                    double i0[3], i1[3];
                    worldToIndexF(lo, i0); worldToIndexF(hi, i1);
                    const float delta = 1.f;   // filter slop
                    int n0[3], n1[3];
                    for (int a = 0; a < 3; a++) {
                        n0[a] = std::max(int(i0[a] - delta), g.bbox_min[a]);
                        n1[a] = std::min(int(i1[a] + delta), g.bbox_max[a]);
                    }
                    float maxValue = 0;
                    for (int nz = n0[2]; nz <= n1[2]; ++nz)
                        for (int ny = n0[1]; ny <= n1[1]; ++ny)
                            for (int nx = n0[0]; nx <= n1[0]; ++nx) maxValue = std::max(maxValue, value(nx, ny, nz));
                    maj[x + 64 * (y + 64 * (size_t)z)] = (maxValue + densityOffset) * majorantScale;
                }
        m.density = own(std::move(g.dense));
        m.majorant = own(std::move(maj));
    }

    int make_material(const std::string &type, const Params &ps, Lexer &lx) {
        if (type == "interface" || type.empty() || type == "none") { ps.report_unused(); return -1; }
        if (type != "diffuse") lx.fail("Material \"" + type + "\": only \"diffuse\" and \"interface\" are on the supported path (SURVEY 8f3)", VSPG_ERR_UNSUPPORTED);
        vspg_material m{};
        m.type = VSPG_MATERIAL_DIFFUSE;
        float r[3] = {0.5f, 0.5f, 0.5f};
        ps.rgb("reflectance", r);
        if (ps.find("displacement", {"texture"}) || ps.find("normalmap", {"string"})) lx.fail("Material: displacement / normal maps are not supported", VSPG_ERR_UNSUPPORTED);
        for (int i = 0; i < 3; i++) m.reflectance[i] = std::min(std::max(r[i], 0.f), 1.f);   // materials.h DiffuseMaterial: Clamp(reflectance, 0, 1)
        ps.report_unused();
        S.materials.push_back(m);
        return (int)S.materials.size() - 1;
    }

    void light_source(const std::string &type, const Params &ps, Lexer &lx) {
        vspg_light l{};
        l.shape_index = -1; l.medium = -1;
        float L[3] = {1, 1, 1};   // RGB build: the colour space's illuminant samples to (1,1,1)
        float sc = (float)ps.one_float("scale", 1);
        XF rfl = render_from_object();
        if (type == "point") {
            l.type = VSPG_LIGHT_POINT;
            ps.rgb("I", L);
            double from[3] = {0, 0, 0};
            ps.point3("from", from);
            float pw = (float)ps.one_float("power", -1);
            if (pw > 0) sc *= pw / (4 * 3.14159265358979323846f);   // lights.cpp:234-238 PointLight::Create: k_e = 4 Pi (photometric == 1)
            // lights.cpp:240-242: finalRenderFromLight = renderFromLight * Translate(from); the light sits at its origin
            const double origin[3] = {0, 0, 0};
            to_render(rfl * xf::translate((float)from[0], (float)from[1], (float)from[2]), origin, l.pos);
            l.medium = medium_id(S.gs.outside, lx);
        } else if (type == "spot") {
            l.type = VSPG_LIGHT_SPOT;
            ps.rgb("I", L);
            float coneangle = (float)ps.one_float("coneangle", 30), conedelta = (float)ps.one_float("conedeltaangle", 5);
            double from[3] = {0, 0, 0}, to[3] = {0, 0, 1};
            ps.point3("from", from); ps.point3("to", to);
            if (conedelta < 0) lx.fail("spot light: conedeltaangle must not be negative");   // CHECK_LE(falloffStart, totalWidth), lights.cpp:1440
            // lights.cpp:1518-1520: dirToZ = Transform(Frame::FromZ(Normalize(to - from))), whose inverse comes from the
            // numeric matrix inversion of Transform(SquareMatrix); t = Translate(from) * Inverse(dirToZ)
            float w[3] = {(float)to[0] - (float)from[0], (float)to[1] - (float)from[1], (float)to[2] - (float)from[2]};
            if (xf::length3(w[0], w[1], w[2]) == 0) lx.fail("spot light: from == to");
            xf::normalize3(w);
            float sign = std::copysign(1.f, w[2]), ca = -1 / (sign + w[2]), cb = w[0] * w[1] * ca;   // CoordinateSystem (util/vecmath.h:1007-1013)
            float v1[3] = {1 + sign * (w[0] * w[0]) * ca, sign * cb, -sign * w[0]}, v2[3] = {cb, sign + (w[1] * w[1]) * ca, -w[1]};
            M4f fm = M4f::identity();
            for (int i = 0; i < 3; i++) { fm.m[0][i] = v1[i]; fm.m[1][i] = v2[i]; fm.m[2][i] = w[i]; }
            XF fin = rfl * (xf::translate((float)from[0], (float)from[1], (float)from[2]) * xf::inverse(XF(fm)));
            l.cos_falloff_end = vspg_cosf(xf::radians(coneangle));
            l.cos_falloff_start = vspg_cosf(xf::radians(coneangle - conedelta));
            float pw = (float)ps.one_float("power", -1);
            if (pw > 0) {   // lights.cpp:1524-1531 (photometric == 1)
                float k_e = 2 * 3.14159265358979323846f * ((1 - l.cos_falloff_start) + (l.cos_falloff_start - l.cos_falloff_end) / 2);
                sc *= pw / k_e;
            }
            const double origin[3] = {0, 0, 0};
            to_render(fin, origin, l.pos);
            for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) l.light_from_render[3 * i + j] = fin.inv.m[i][j];
            const float z[3] = {0, 0, 1};
            float ax[3];
            xf::apply_vector(fin.m, z, ax);
            xf::normalize3(ax);
            for (int i = 0; i < 3; i++) l.axis[i] = ax[i];
            l.medium = medium_id(S.gs.outside, lx);
        } else if (type == "distant") {
            l.type = VSPG_LIGHT_DISTANT;
            ps.rgb("L", L);
            double from[3] = {0, 0, 0}, to[3] = {0, 0, 1};
            ps.point3("from", from); ps.point3("to", to);
            float ev = (float)ps.one_float("illuminance", -1);
            if (ev > 0) sc *= ev;
            // lights.cpp:289-298: w = Normalize(from - to) becomes the z column of a frame t; finalRenderFromLight =
            // renderFromLight * t; the light direction is Normalize(finalRenderFromLight(Vector3f(0,0,1))) (lights.cpp:259)
            float w[3] = {(float)from[0] - (float)to[0], (float)from[1] - (float)to[1], (float)from[2] - (float)to[2]};
            if (xf::length3(w[0], w[1], w[2]) == 0) lx.fail("distant light: from == to");
            xf::normalize3(w);
            M4f t = M4f::identity();
            float sign = std::copysign(1.f, w[2]), ca = -1 / (sign + w[2]), cb = w[0] * w[1] * ca;   // CoordinateSystem (util/vecmath.h:1007-1013)
            float v1[3] = {1 + sign * (w[0] * w[0]) * ca, sign * cb, -sign * w[0]}, v2[3] = {cb, sign + (w[1] * w[1]) * ca, -w[1]};
            for (int i = 0; i < 3; i++) { t.m[i][0] = v1[i]; t.m[i][1] = v2[i]; t.m[i][2] = w[i]; }
            M4f fin = xf::mul(rfl.m, t);
            const float z[3] = {0, 0, 1};
            float wr[3];
            xf::apply_vector(fin, z, wr);
            xf::normalize3(wr);
            for (int i = 0; i < 3; i++) l.pos[i] = wr[i];
        } else if (type == "infinite") {
            l.type = VSPG_LIGHT_UNIFORM_INFINITE;
            ps.rgb("L", L);
            if (!ps.one_string("filename", "").empty() || ps.find("portal", {"point3", "point"}))
                lx.fail("image / portal infinite lights are outside the hot-path scope (SURVEY 8f3)", VSPG_ERR_UNSUPPORTED);
            float ev = (float)ps.one_float("illuminance", -1);
            if (ev > 0) sc *= ev / 3.14159265358979323846f;
        } else lx.fail("LightSource \"" + type + "\": supported types are point, spot, distant, infinite (uniform)", VSPG_ERR_UNSUPPORTED);
        ps.report_unused();
        for (int i = 0; i < 3; i++) l.L[i] = L[i];
        l.scale = sc;
        S.lights.push_back(l);
    }

    void area_light(const std::string &type, const Params &ps, Lexer &lx) {
        if (type != "diffuse") lx.fail("AreaLightSource \"" + type + "\": only \"diffuse\" exists");
        GState &g = S.gs;
        g.area = true;
        g.area_L[0] = g.area_L[1] = g.area_L[2] = 1;
        ps.rgb("L", g.area_L);
        g.area_scale = (float)ps.one_float("scale", 1);
        g.area_two = ps.one_bool("twosided", false);
        g.area_power = ps.one_float("power", -1);
        if (!ps.one_string("filename", "").empty()) lx.fail("textured area lights are not supported", VSPG_ERR_UNSUPPORTED);
        ps.report_unused();
    }

    // shared per-shape attributes (scene.cpp BasicSceneBuilder::Shape + cpu/render.cpp primitive creation)
    struct ShapeAttr { int material, area_light, mi, mo, has_interface, flip; };
    ShapeAttr shape_attr(Lexer &lx, const XF &rfo) {
        ShapeAttr a;
        a.material = S.gs.material;
        a.area_light = -1;
        a.mi = medium_id(S.gs.inside, lx); a.mo = medium_id(S.gs.outside, lx);
        // interaction.h:117-121 / primitive.h GeometricPrimitive: a MediumInterface is attached only when it
        // marks a transition (inside != outside); otherwise rays keep their current medium.
        a.has_interface = (a.mi != a.mo) ? 1 : 0;
        a.flip = (S.gs.reverse ? 1 : 0) ^ (xf::det3(rfo.m) < 0 ? 1 : 0);   // Transform::SwapsHandedness
        return a;
    }
    int attach_area_light(int kind, int index, float area, Lexer &lx) {
        const GState &g = S.gs;
        if (!g.area) return -1;
        vspg_light l{};
        l.type = VSPG_LIGHT_DIFFUSE_AREA;
        float sc = g.area_scale;
        if (g.area_power > 0) sc *= (float)(g.area_power / ((g.area_two ? 2 : 1) * (double)area * M_PI));   // lights.cpp DiffuseAreaLight::Create
        for (int i = 0; i < 3; i++) l.L[i] = g.area_L[i];
        l.scale = sc;
        l.shape_kind = kind; l.shape_index = index; l.two_sided = g.area_two; l.medium = -1;
        S.lights.push_back(l);
        return (int)S.lights.size() - 1;
    }

    void shape(const std::string &type, const Params &ps, Lexer &lx);
};

float tri_area(const float *a, const float *b, const float *c) {
    double u[3] = {(double)b[0] - a[0], (double)b[1] - a[1], (double)b[2] - a[2]}, v[3] = {(double)c[0] - a[0], (double)c[1] - a[1], (double)c[2] - a[2]};
    double x = u[1] * v[2] - u[2] * v[1], y = u[2] * v[0] - u[0] * v[2], z = u[0] * v[1] - u[1] * v[0];
    return (float)(0.5 * std::sqrt(x * x + y * y + z * z));
}

void Builder::shape(const std::string &type, const Params &ps, Lexer &lx) {
    XF rfo = render_from_object();
    ShapeAttr a = shape_attr(lx, rfo);
    if (ps.find("alpha", {"texture", "float"})) lx.fail("Shape: alpha textures are not supported", VSPG_ERR_UNSUPPORTED);
    auto emit_quads = [&](const std::vector<double> &P, size_t nv, const std::vector<double> &idx, const char *what) {
        for (size_t q = 0; q < idx.size(); q += 4) {
            vspg_quad d{};
            float *dst[4] = {d.p00, d.p10, d.p01, d.p11};
            for (int k = 0; k < 4; k++) {
                size_t vi = (size_t)idx[q + k];
                if (idx[q + k] < 0 || vi >= nv) lx.fail(std::string(what) + ": vertex index out of range");
                to_render(rfo, &P[3 * vi], dst[k]);
            }
            d.material = a.material; d.medium_inside = a.mi; d.medium_outside = a.mo; d.has_interface = a.has_interface; d.flip_normal = a.flip;
            float area = tri_area(d.p00, d.p10, d.p11) + tri_area(d.p00, d.p11, d.p01);   // only used for "power"; exact area is computed by set_scene
            d.area_light = attach_area_light(0, (int)S.quads.size(), area, lx);
            S.quads.push_back(d);
        }
    };
    auto emit_tris = [&](const std::vector<double> &P, size_t nv, const std::vector<double> &idx, const char *what) {
        for (size_t t = 0; t < idx.size(); t += 3) {
            vspg_triangle d{};
            float *dst[3] = {d.p0, d.p1, d.p2};
            for (int k = 0; k < 3; k++) {
                size_t vi = (size_t)idx[t + k];
                if (idx[t + k] < 0 || vi >= nv) lx.fail(std::string(what) + ": vertex index out of range");
                to_render(rfo, &P[3 * vi], dst[k]);
            }
            d.material = a.material; d.medium_inside = a.mi; d.medium_outside = a.mo; d.has_interface = a.has_interface; d.flip_normal = a.flip;
            d.area_light = attach_area_light(1, (int)S.tris.size(), tri_area(d.p0, d.p1, d.p2), lx);
            S.tris.push_back(d);
        }
    };
    if (type == "bilinearmesh") {
        const auto *P = ps.points("P");
        if (!P || P->size() % 3) lx.fail("bilinearmesh: vertex positions \"P\" not provided / malformed");
        size_t nv = P->size() / 3;
        std::vector<double> idx;
        if (const auto *I = ps.ints("indices")) idx = *I;
        else if (nv == 4) idx = {0, 1, 2, 3};
        else lx.fail("bilinearmesh: vertex indices \"indices\" must be provided with more than 4 vertices");
        if (idx.size() % 4) lx.fail("bilinearmesh: number of indices must be a multiple of 4");
        if (ps.find("uv", {"point2"}) || ps.find("N", {"normal", "normal3"}) || !ps.one_string("emissionfilename", "").empty())
            lx.fail("bilinearmesh: per-vertex uv / N / emission images are not supported", VSPG_ERR_UNSUPPORTED);
        ps.report_unused();
        emit_quads(*P, nv, idx, "bilinearmesh");
    } else if (type == "trianglemesh") {
        const auto *P = ps.points("P");
        if (!P || P->size() % 3) lx.fail("trianglemesh: vertex positions \"P\" not provided / malformed");
        size_t nv = P->size() / 3;
        std::vector<double> idx;
        if (const auto *I = ps.ints("indices")) idx = *I;
        else if (nv == 3) idx = {0, 1, 2};
        else lx.fail("trianglemesh: vertex indices \"indices\" must be provided with more than 3 vertices");
        if (idx.size() % 3) lx.fail("trianglemesh: number of indices must be a multiple of 3");
        if (ps.find("N", {"normal", "normal3"}) || ps.find("S", {"vector3", "vector"}))
            lx.fail("trianglemesh: shading normals / tangents are not supported (geometric normals only)", VSPG_ERR_UNSUPPORTED);
        ps.find("uv", {"point2"});   // uv only feeds textures, which this path does not have
        ps.report_unused();
        emit_tris(*P, nv, idx, "trianglemesh");
    } else if (type == "plymesh") {   // shapes.cpp:1415-1472: triangles first, then the quads as bilinear patches
        std::string filename = ps.one_string("filename", "");
        if (filename.empty()) lx.fail("plymesh: no \"filename\" given");
        ps.one_float("edgelength", 1);
        if (ps.find("displacement", {"texture"})) lx.fail("plymesh: displacement is not supported", VSPG_ERR_UNSUPPORTED);
        ps.report_unused();
        vspg_host::PlyMesh mesh;
        std::string err;
        if (!vspg_host::read_ply(resolve(filename), &mesh, &err)) lx.fail(err, VSPG_ERR_IO);
        if (mesh.has_normals) lx.fail("plymesh: \"" + filename + "\" carries per-vertex normals; shading normals are not supported (geometric normals only)", VSPG_ERR_UNSUPPORTED);
        std::vector<double> P(mesh.p.begin(), mesh.p.end());   // (uv only feeds textures, which this path does not have)
        const size_t nv = P.size() / 3;
        emit_tris(P, nv, std::vector<double>(mesh.tri.begin(), mesh.tri.end()), "plymesh");
        emit_quads(P, nv, std::vector<double>(mesh.quad.begin(), mesh.quad.end()), "plymesh");
    } else if (type == "sphere") {
        double r = ps.one_float("radius", 1);
        double zmin = ps.one_float("zmin", -r), zmax = ps.one_float("zmax", r), phimax = ps.one_float("phimax", 360);
        ps.report_unused();
        if (zmin > -r || zmax < r || phimax < 360) lx.fail("sphere: partial spheres are not supported", VSPG_ERR_UNSUPPORTED);
        // object space must be translate(center) * uniform scale: the sphere stays a sphere in render space
        double sx = std::sqrt((double)rfo.m.m[0][0] * (double)rfo.m.m[0][0] + (double)rfo.m.m[1][0] * (double)rfo.m.m[1][0] + (double)rfo.m.m[2][0] * (double)rfo.m.m[2][0]);
        double sy = std::sqrt((double)rfo.m.m[0][1] * (double)rfo.m.m[0][1] + (double)rfo.m.m[1][1] * (double)rfo.m.m[1][1] + (double)rfo.m.m[2][1] * (double)rfo.m.m[2][1]);
        double sz = std::sqrt((double)rfo.m.m[0][2] * (double)rfo.m.m[0][2] + (double)rfo.m.m[1][2] * (double)rfo.m.m[1][2] + (double)rfo.m.m[2][2] * (double)rfo.m.m[2][2]);
        if (std::fabs(sx - sy) > 1e-6 * sx || std::fabs(sx - sz) > 1e-6 * sx) lx.fail("sphere: non-uniform scaling is not supported", VSPG_ERR_UNSUPPORTED);
        vspg_sphere d{};
        const double origin[3] = {0, 0, 0};
        to_render(rfo, origin, d.center);
        d.radius = (float)(r * sx);
        d.material = a.material; d.medium_inside = a.mi; d.medium_outside = a.mo; d.has_interface = a.has_interface; d.flip_normal = a.flip;
        d.area_light = attach_area_light(2, (int)S.spheres.size(), (float)(4 * M_PI * d.radius * d.radius), lx);
        S.spheres.push_back(d);
    } else
        lx.fail("Shape \"" + type + "\": supported shapes are bilinearmesh, trianglemesh, plymesh, sphere (curves / cylinders / disks / subdivision surfaces are SURVEY 8f3)", VSPG_ERR_UNSUPPORTED);
}

double num(Lexer &lx, const char *what) {
    Tok t = lx.next();
    if (t.kind != Tok::WORD) lx.fail(std::string(what) + ": expected a number");
    char *end = nullptr;
    double v = strtod(t.s.c_str(), &end);
    if (end == t.s.c_str() || *end) lx.fail(std::string(what) + ": \"" + t.s + "\" is not a number");
    return v;
}
std::string str(Lexer &lx, const char *what) {
    Tok t = lx.next();
    if (t.kind != Tok::STR) lx.fail(std::string(what) + ": expected a quoted string");
    return t.s;
}
XF matrix16(Lexer &lx, const char *what) {
    bool br = lx.peek().kind == Tok::LB;
    if (br) lx.next();
    double v[16];
    for (double &x : v) x = num(lx, what);
    if (br && lx.next().kind != Tok::RB) lx.fail(std::string(what) + ": expected ]");
    // scene.cpp:583-593: Transpose(Transform(SquareMatrix<4>(tr))) -- the inverse is taken of the matrix as given
    // (row-major), then both are transposed (pbrt files list matrices column-major)
    M4f m;
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) m.m[i][j] = (float)v[4 * i + j];
    XF t(m);
    if (std::isnan(t.inv.m[0][0])) lx.fail(std::string(what) + ": singular transformation matrix");
    return xf::transpose(t);
}

void run(Lexer &lx, Builder &B, int depth);

void include_file(const std::string &rel, Builder &B, int depth) {
    if (depth > 16) throw ParseError(VSPG_ERR_INVALID, "Include nesting too deep");
    Lexer sub; sub.file = B.resolve(rel); sub.text = read_file(sub.file);
    run(sub, B, depth + 1);
}

void directive(Lexer &lx, Builder &B, int depth, const Tok &t);

void run(Lexer &lx, Builder &B, int depth) {
    for (;;) {
        Tok t = lx.next();
        if (t.kind == Tok::END) return;
        if (t.kind != Tok::WORD) lx.fail("expected a directive");
        try { directive(lx, B, depth, t); }
        catch (const ParseError &e) {   // every message carries file:line, like the reference's FileLoc
            std::string m = e.what();
            if (m.find(".pbrt:") != std::string::npos || m.find("<string>:") != std::string::npos || m.rfind(lx.file + ":", 0) == 0) throw;
            throw ParseError(e.code, lx.file + ":" + std::to_string(t.line) + ": " + m);
        }
    }
}

void directive(Lexer &lx, Builder &B, int depth, const Tok &t) {
    vspg_host_scene &S = B.S;
    {
        const std::string &d = t.s;
        auto pre_world = [&] { if (S.world) lx.fail(d + " is not allowed after WorldBegin"); };
        auto in_world = [&] { if (!S.world) lx.fail(d + " is not allowed before WorldBegin"); };
        if (d == "Identity") S.gs.ctm = XF();
        else if (d == "Translate") { float x = (float)num(lx, "Translate"), y = (float)num(lx, "Translate"), z = (float)num(lx, "Translate"); S.gs.ctm = S.gs.ctm * xf::translate(x, y, z); }
        else if (d == "Scale") { float x = (float)num(lx, "Scale"), y = (float)num(lx, "Scale"), z = (float)num(lx, "Scale"); if (x == 0 || y == 0 || z == 0) lx.fail("Scale: zero factor"); S.gs.ctm = S.gs.ctm * xf::scale(x, y, z); }
        else if (d == "Rotate") {
            float a = (float)num(lx, "Rotate"), x = (float)num(lx, "Rotate"), y = (float)num(lx, "Rotate"), z = (float)num(lx, "Rotate");
            XF r;
            if (!xf::rotate(a, x, y, z, &r)) lx.fail("Rotate: zero-length axis");
            S.gs.ctm = S.gs.ctm * r;
        }
        else if (d == "Transform") S.gs.ctm = matrix16(lx, "Transform");
        else if (d == "ConcatTransform") S.gs.ctm = S.gs.ctm * matrix16(lx, "ConcatTransform");
        else if (d == "LookAt") {
            float v[9];
            for (float &x : v) x = (float)num(lx, "LookAt");
            if (v[0] == v[3] && v[1] == v[4] && v[2] == v[5]) lx.fail("LookAt: eye and look-at point coincide");
            XF la;
            if (!xf::look_at(v, v + 3, v + 6, &la)) lx.fail("LookAt: up vector and viewing direction are collinear");
            S.gs.ctm = S.gs.ctm * la;
        }
        else if (d == "CoordinateSystem") S.named_cs[str(lx, "CoordinateSystem")] = S.gs.ctm;
        else if (d == "CoordSysTransform") {
            std::string n = str(lx, "CoordSysTransform");
            auto it = S.named_cs.find(n);
            if (it == S.named_cs.end()) lx.fail("CoordSysTransform: unknown coordinate system \"" + n + "\"");
            S.gs.ctm = it->second;
        }
        else if (d == "ActiveTransform") { Tok w = lx.next(); if (w.s != "All") lx.fail("ActiveTransform: motion blur is outside the hot-path scope", VSPG_ERR_UNSUPPORTED); }
        else if (d == "TransformTimes") { num(lx, d.c_str()); num(lx, d.c_str()); }
        else if (d == "ReverseOrientation") S.gs.reverse = !S.gs.reverse;
        else if (d == "AttributeBegin" || d == "TransformBegin") S.stack.push_back(S.gs);
        else if (d == "AttributeEnd" || d == "TransformEnd") {
            if (S.stack.empty()) lx.fail("unmatched " + d);
            if (d == "TransformEnd") { XF c = S.stack.back().ctm; S.stack.pop_back(); S.gs.ctm = c; }
            else { S.gs = S.stack.back(); S.stack.pop_back(); }
        }
        else if (d == "Option") {
            Params ps = parse_params(lx, "Option");
            S.seed = ps.one_int("seed", S.seed);
            for (const Param &p : ps.list) p.used = true;   // other options (threads, quiet, ...) do not affect the render pass
        }
        else if (d == "ColorSpace") { if (str(lx, "ColorSpace") != "srgb") lx.fail("only the srgb colour space is supported", VSPG_ERR_UNSUPPORTED); }
        else if (d == "Camera") { pre_world(); std::string ty = str(lx, "Camera"); B.camera(ty, parse_params(lx, "Camera \"" + ty + "\""), lx); }
        else if (d == "Film") { pre_world(); std::string ty = str(lx, "Film"); B.film(ty, parse_params(lx, "Film \"" + ty + "\""), lx); }
        else if (d == "Sampler") { pre_world(); std::string ty = str(lx, "Sampler"); B.sampler(ty, parse_params(lx, "Sampler \"" + ty + "\""), lx); }
        else if (d == "PixelFilter") { pre_world(); std::string ty = str(lx, "PixelFilter"); B.pixel_filter(ty, parse_params(lx, "PixelFilter \"" + ty + "\""), lx); }
        else if (d == "Integrator") { pre_world(); std::string ty = str(lx, "Integrator"); B.integrator(ty, parse_params(lx, "Integrator \"" + ty + "\""), lx); }
        else if (d == "Accelerator") { pre_world(); str(lx, "Accelerator"); Params ps = parse_params(lx, "Accelerator"); (void)ps; }   // the library builds its own threaded BVH
        else if (d == "MakeNamedMedium") { std::string n = str(lx, d.c_str()); B.make_named_medium(n, parse_params(lx, "MakeNamedMedium \"" + n + "\""), lx); }
        else if (d == "MediumInterface") {
            S.gs.inside = str(lx, d.c_str());
            S.gs.outside = lx.peek().kind == Tok::STR ? lx.next().s : S.gs.inside;   // parser.cpp: one name => both sides
            S.gs.has_mi = true;
        }
        else if (d == "WorldBegin") {
            pre_world();
            if (!S.have_camera) lx.fail("WorldBegin before a Camera directive (a perspective camera is required)");
            S.world = true;
            S.gs.ctm = XF();
            S.named_cs["world"] = S.gs.ctm;
        }
        else if (d == "Material") { in_world(); std::string ty = str(lx, d.c_str()); S.gs.material = B.make_material(ty, parse_params(lx, "Material \"" + ty + "\""), lx); }
        else if (d == "MakeNamedMaterial") {
            in_world();
            std::string n = str(lx, d.c_str());
            Params ps = parse_params(lx, "MakeNamedMaterial \"" + n + "\"");
            std::string ty = ps.one_string("type", "");
            if (ty.empty()) lx.fail("MakeNamedMaterial \"" + n + "\": no \"string type\" given");
            S.named_materials[n] = B.make_material(ty, ps, lx);
        }
        else if (d == "NamedMaterial") {
            in_world();
            std::string n = str(lx, d.c_str());
            auto it = S.named_materials.find(n);
            if (it == S.named_materials.end()) lx.fail("NamedMaterial \"" + n + "\" is not defined");
            S.gs.material = it->second;
        }
        else if (d == "LightSource") { in_world(); std::string ty = str(lx, d.c_str()); B.light_source(ty, parse_params(lx, "LightSource \"" + ty + "\""), lx); }
        else if (d == "AreaLightSource") { in_world(); std::string ty = str(lx, d.c_str()); B.area_light(ty, parse_params(lx, "AreaLightSource \"" + ty + "\""), lx); }
        else if (d == "Shape") { in_world(); std::string ty = str(lx, d.c_str()); B.shape(ty, parse_params(lx, "Shape \"" + ty + "\""), lx); }
        else if (d == "Include") include_file(str(lx, "Include"), B, depth);
        else if (d == "Import") include_file(str(lx, "Import"), B, depth);
        else if (d == "Texture" || d == "ObjectBegin" || d == "ObjectEnd" || d == "ObjectInstance" || d == "Attribute")
            lx.fail(d + " is outside the hot-path scope (SURVEY 8f3)", VSPG_ERR_UNSUPPORTED);
        else lx.fail("unknown directive \"" + d + "\"");
    }
}

void finish(vspg_host_scene &S) {
    if (!S.world) throw ParseError(VSPG_ERR_INVALID, "scene has no WorldBegin");
    if (!S.stack.empty()) throw ParseError(VSPG_ERR_INVALID, "unmatched AttributeBegin at end of file");
    if (!S.have_integrator) {
        // pbrt's default integrator is "volpath"; this library is the guidedvolpath render pass only.
        throw ParseError(VSPG_ERR_UNSUPPORTED, "scene has no Integrator \"guidedvolpath\" directive");
    }
    // ---- camera (cameras.h ProjectiveCamera ctor / cameras.cpp PerspectiveCamera::Create)
    vspg_camera &c = S.desc.camera;
    // cameras.cpp:499-533 PerspectiveCamera::Create, cameras.h:259-278 ProjectiveCamera, :348-351 Perspective(fov, 1e-2, 1000)
    float frame = S.frame_aspect > 0 ? S.frame_aspect : float(S.xres) / float(S.yres);
    float sw[4];
    if (frame > 1.f) { sw[0] = -frame; sw[1] = frame; sw[2] = -1.f; sw[3] = 1.f; }
    else { sw[0] = -1.f; sw[1] = 1.f; sw[2] = -1.f / frame; sw[3] = 1.f / frame; }
    if (!S.screen_window.empty()) for (int i = 0; i < 4; i++) sw[i] = (float)S.screen_window[i];
    XF screen_from_camera = xf::perspective(S.fov, 1e-2f, 1000.f);
    XF ndc_from_screen = xf::scale(1 / (sw[1] - sw[0]), 1 / (sw[3] - sw[2]), 1) * xf::translate(-sw[0], -sw[3], 0);
    XF raster_from_ndc = xf::scale((float)S.xres, -(float)S.yres, 1);
    XF raster_from_screen = raster_from_ndc * ndc_from_screen;
    XF camera_from_raster = xf::inverse(screen_from_camera) * xf::inverse(raster_from_screen);
    c.camera_from_raster = to_abi(camera_from_raster);
    c.render_from_camera = to_abi(S.render_from_world * S.world_from_camera);   // cameras.cpp:52-56
    c.lens_radius = S.lens_radius; c.focal_distance = S.focal_distance;
    auto mid = S.medium_ids.find(S.camera_medium);
    if (!S.camera_medium.empty() && mid == S.medium_ids.end()) throw ParseError(VSPG_ERR_INVALID, "camera medium \"" + S.camera_medium + "\" is not defined");
    c.medium = S.camera_medium.empty() ? -1 : mid->second;
    c.width = S.xres; c.height = S.yres;
    c.filter_type = S.filter_type; c.filter_radius[0] = S.filter_r[0]; c.filter_radius[1] = S.filter_r[1]; c.filter_sigma = S.filter_type == 1 ? S.filter_sigma : 0.f;
    c.imaging_ratio = S.exposure * S.iso / 100.f;   // film.cpp:233
    c.max_component_value = S.max_comp;
    // ---- light order: BasicScene::CreateLights (scene.cpp:1320-1376) appends the area lights first, in shape order, and
    // the LightSource lights after them.  Light samplers index this array (uniform pick, alias table, light BVH), so the
    // order is part of the result.
    {
        std::vector<int> order;
        for (int pass = 0; pass < 2; pass++)
            for (int i = 0; i < (int)S.lights.size(); i++)
                if ((S.lights[i].type == VSPG_LIGHT_DIFFUSE_AREA) == (pass == 0)) order.push_back(i);
        std::vector<vspg_light> sorted;
        std::vector<int> new_index(S.lights.size());
        for (int k = 0; k < (int)order.size(); k++) { sorted.push_back(S.lights[order[k]]); new_index[order[k]] = k; }
        S.lights.swap(sorted);
        for (auto &q : S.quads) if (q.area_light >= 0) q.area_light = new_index[q.area_light];
        for (auto &q : S.tris) if (q.area_light >= 0) q.area_light = new_index[q.area_light];
        for (auto &q : S.spheres) if (q.area_light >= 0) q.area_light = new_index[q.area_light];
    }
    // ---- arrays
    vspg_scene_desc &d = S.desc;
    d.n_media = (int)S.media.size(); d.media = S.media.data();
    d.n_quads = (int)S.quads.size(); d.quads = S.quads.data();
    d.n_triangles = (int)S.tris.size(); d.triangles = S.tris.data();
    d.n_spheres = (int)S.spheres.size(); d.spheres = S.spheres.data();
    d.n_materials = (int)S.materials.size(); d.materials = S.materials.data();
    d.n_lights = (int)S.lights.size(); d.lights = S.lights.data();
    // ---- guiding inputs (vspg.cpp:111-193).  Guiding requested without a loadable file => hard error: the reference
    // would start training a fresh field here, which this library does not do.
    const vspg_integrator_params &ip = S.ip;
    char err[512];
    bool need_field = ip.surface_guiding || ip.volume_guiding || (ip.vsp_guiding && ip.vsp_secondary_guiding);
    if (need_field) {
        if (!S.load_cache || S.guiding_cache_file.empty())
            throw ParseError(VSPG_ERR_UNSUPPORTED, "Integrator: guiding is enabled but \"bool loadGuidingCache\" true + \"string guidingCacheFileName\" are not given; "
                                                   "this library renders with a frozen, loaded cache only (no training pass)");
        std::string p = S.guiding_cache_file[0] == '/' ? S.guiding_cache_file : S.base_dir + "/" + S.guiding_cache_file;
        int rc = vspg_host_field_load(p.c_str(), &S.field, err, sizeof err);
        if (rc != VSPG_OK) throw ParseError(rc, std::string("loadGuidingCache: ") + err);
        S.guiding.surface = vspg_host_field_surface(S.field);
        S.guiding.volume = vspg_host_field_volume(S.field);
        S.has_guiding = true;
    }
    if (ip.vsp_guiding && ip.vsp_primary_guiding) {
        std::string p = S.isg_file.empty() ? "" : (S.isg_file[0] == '/' ? S.isg_file : S.base_dir + "/" + S.isg_file);
        if (!S.load_isg || S.isg_file.empty()) {
            // vspg.cpp:161-178: no loaded buffer -> the image-space buffer is trained on the fly by the render's waves
            // (vspg_isg_store_*; the driver renders one sample index per wave, see vspg_pbrt_main.cpp / Context.render_training_isg)
            S.train_isg = true;
            if (S.store_isg) S.store_isg_path = p;
        } else {
        int w = 0, h = 0;
        int rc = vspg_exr_read_channel(p.c_str(), "VSP.Y", &S.isg, &w, &h, err, sizeof err);
        if (rc != VSPG_OK) throw ParseError(rc, std::string("loadISGBuffer: ") + err);
        S.guiding.isg_vsp = S.isg; S.guiding.isg_width = w; S.guiding.isg_height = h;
        S.has_guiding = true;
        }
    }
    if (S.store_tr) S.store_tr_path = S.tr_file[0] == '/' ? S.tr_file : S.base_dir + "/" + S.tr_file;
    if (S.load_tr) {
        if (S.tr_file.empty()) throw ParseError(VSPG_ERR_INVALID, "Integrator: loadTrBuffer without trBufferFileName");
        std::string p = S.tr_file[0] == '/' ? S.tr_file : S.base_dir + "/" + S.tr_file;
        float *ch[3] = {nullptr, nullptr, nullptr};
        int w = 0, h = 0;
        const char *names[3] = {"Transmittance.R", "Transmittance.G", "Transmittance.B"};   // cpu/trbuffer.h:52-98
        for (int k = 0; k < 3; k++) {
            int rc = vspg_exr_read_channel(p.c_str(), names[k], &ch[k], &w, &h, err, sizeof err);
            if (rc != VSPG_OK) { for (float *q : ch) if (q) vspg_free(q); throw ParseError(rc, std::string("loadTrBuffer: ") + err); }
        }
        S.tr = (float *)malloc((size_t)w * h * 3 * sizeof(float));
        for (size_t i = 0; i < (size_t)w * h; i++) for (int k = 0; k < 3; k++) S.tr[3 * i + k] = ch[k][i];
        for (float *q : ch) vspg_free(q);
        S.guiding.tr_rgb = S.tr; S.guiding.tr_width = w; S.guiding.tr_height = h;
        S.has_guiding = true;
    } else if (ip.vsp_guiding && ip.vsp_sampling_nds && ip.collision_probability_bias)
        throw ParseError(VSPG_ERR_UNSUPPORTED, "Integrator: collisionProbabilityBias (NDS+) needs \"bool loadTrBuffer\" true + \"string trBufferFileName\"");
}

int parse_common(Lexer &lx, const std::string &base_dir, vspg_host_scene **out, char *err, size_t err_len) {
    std::unique_ptr<vspg_host_scene> S(new vspg_host_scene);
    S->base_dir = base_dir;
    vspg_integrator_defaults(&S->ip);
    try {
        Builder B(*S);
        run(lx, B, 0);
        finish(*S);
    } catch (const ParseError &e) {
        if (err && err_len) snprintf(err, err_len, "%s", e.what());
        return e.code;
    } catch (const std::exception &e) {
        if (err && err_len) snprintf(err, err_len, "%s", e.what());
        return VSPG_ERR_INVALID;
    }
    *out = S.release();
    return VSPG_OK;
}

}  // namespace

extern "C" {

int vspg_host_scene_parse(const char *pbrt_path, vspg_host_scene **out, char *err, size_t err_len) {
    if (!pbrt_path || !out) return VSPG_ERR_INVALID;
    *out = nullptr;
    Lexer lx; lx.file = pbrt_path;
    try { lx.text = read_file(pbrt_path); }
    catch (const ParseError &e) { if (err && err_len) snprintf(err, err_len, "%s", e.what()); return e.code; }
    return parse_common(lx, dirname_of(pbrt_path), out, err, err_len);
}

int vspg_host_scene_parse_string(const char *pbrt_text, const char *base_dir, vspg_host_scene **out, char *err, size_t err_len) {
    if (!pbrt_text || !out) return VSPG_ERR_INVALID;
    *out = nullptr;
    Lexer lx; lx.file = "<string>"; lx.text = pbrt_text;
    return parse_common(lx, base_dir && *base_dir ? base_dir : ".", out, err, err_len);
}

void vspg_host_scene_free(vspg_host_scene *s) { delete s; }
const vspg_scene_desc *vspg_host_scene_desc(const vspg_host_scene *s) { return s ? &s->desc : nullptr; }
const vspg_guiding_desc *vspg_host_scene_guiding(const vspg_host_scene *s) { return s && s->has_guiding ? &s->guiding : nullptr; }
const vspg_integrator_params *vspg_host_scene_integrator(const vspg_host_scene *s) { return s ? &s->ip : nullptr; }
int vspg_host_scene_spp(const vspg_host_scene *s) { return s ? s->spp : 0; }
int vspg_host_scene_seed(const vspg_host_scene *s) { return s ? s->seed : 0; }
const char *vspg_host_scene_outfile(const vspg_host_scene *s) { return s ? s->outfile.c_str() : ""; }
const char *vspg_host_scene_store_tr_file(const vspg_host_scene *s) { return s ? s->store_tr_path.c_str() : ""; }
const char *vspg_host_scene_store_isg_file(const vspg_host_scene *s) { return s ? s->store_isg_path.c_str() : ""; }
int vspg_host_scene_train_isg(const vspg_host_scene *s) { return s && s->train_isg ? 1 : 0; }

}  // extern "C"
