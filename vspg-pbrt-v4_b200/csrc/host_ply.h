// Minimal PLY reader for Shape "plymesh" (reference: util/mesh.cpp:322-420 TriQuadMesh::ReadPLY, which drives the rply
// library): ascii and binary_little_endian files, element "vertex" with x / y / z (any scalar type, stored as float like
// rply_vertex_callback does) and element "face" with the list "vertex_indices"; triangles and quads,
// the quad index order re-arranged to (v0, v1, v3, v2) as ReadPLY does (:301-305).  Other elements and properties are
// skipped.  Reports whether per-vertex normals are present (the caller rejects those: geometric normals only).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <sstream>
#include <string>
#include <vector>

namespace vspg_host {

struct PlyMesh {
    std::vector<float> p;            // 3 per vertex
    std::vector<int> tri, quad;      // 3 / 4 indices per face
    bool has_normals = false, has_uv = false;
};

namespace ply_detail {
inline int type_size(const std::string &t) {
    if (t == "char" || t == "uchar" || t == "int8" || t == "uint8") return 1;
    if (t == "short" || t == "ushort" || t == "int16" || t == "uint16") return 2;
    if (t == "int" || t == "uint" || t == "int32" || t == "uint32" || t == "float" || t == "float32") return 4;
    if (t == "double" || t == "float64") return 8;
    return 0;
}
inline double decode(const unsigned char *b, const std::string &t) {
    if (t == "char" || t == "int8") return (double)*(const int8_t *)b;
    if (t == "uchar" || t == "uint8") return (double)*b;
    if (t == "short" || t == "int16") { int16_t v; memcpy(&v, b, 2); return v; }
    if (t == "ushort" || t == "uint16") { uint16_t v; memcpy(&v, b, 2); return v; }
    if (t == "int" || t == "int32") { int32_t v; memcpy(&v, b, 4); return v; }
    if (t == "uint" || t == "uint32") { uint32_t v; memcpy(&v, b, 4); return v; }
    if (t == "float" || t == "float32") { float v; memcpy(&v, b, 4); return v; }
    double v; memcpy(&v, b, 8); return v;
}
struct Prop { std::string name, type, countType; bool list = false; };
struct Elem { std::string name; long count = 0; std::vector<Prop> props; };
}  // namespace ply_detail

inline bool read_ply(const std::string &path, PlyMesh *out, std::string *err) {
    using namespace ply_detail;
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) { *err = "Couldn't open PLY file \"" + path + "\""; return false; }
    struct Closer { FILE *f; ~Closer() { fclose(f); } } closer{f};
    auto line = [&](std::string &s) {
        s.clear();
        int c;
        while ((c = fgetc(f)) != EOF && c != '\n') if (c != '\r') s.push_back((char)c);
        return c != EOF || !s.empty();
    };
    std::string l;
    if (!line(l) || l != "ply") { *err = "Unable to read the header of PLY file \"" + path + "\""; return false; }
    bool ascii = false, haveFormat = false;
    std::vector<Elem> elems;
    while (true) {
        if (!line(l)) { *err = "Unable to read the header of PLY file \"" + path + "\""; return false; }
        std::istringstream is(l);
        std::string kw; is >> kw;
        if (kw == "end_header") break;
        if (kw == "format") {
            std::string fmt; is >> fmt;
            if (fmt == "ascii") ascii = true;
            else if (fmt != "binary_little_endian") { *err = path + ": PLY format \"" + fmt + "\" is not supported (ascii, binary_little_endian)"; return false; }
            haveFormat = true;
        } else if (kw == "element") { Elem e; is >> e.name >> e.count; elems.push_back(e); }
        else if (kw == "property") {
            if (elems.empty()) { *err = path + ": property before any element"; return false; }
            Prop p; std::string t; is >> t;
            if (t == "list") { p.list = true; is >> p.countType >> p.type >> p.name; } else { p.type = t; is >> p.name; }
            if (!type_size(p.type) || (p.list && !type_size(p.countType))) { *err = path + ": unknown PLY property type in \"" + l + "\""; return false; }
            elems.back().props.push_back(p);
        }   // comment / obj_info: skipped
    }
    if (!haveFormat) { *err = path + ": PLY header without a format line"; return false; }
    long nv = 0, nf = 0;
    for (auto &e : elems) { if (e.name == "vertex") nv = e.count; if (e.name == "face") nf = e.count; }
    if (nv == 0 || nf == 0) { *err = path + ": PLY file is invalid! No face/vertex elements found!"; return false; }
    auto next = [&](const std::string &t, double *v) {
        if (ascii) return fscanf(f, "%lf", v) == 1;
        unsigned char b[8];
        int n = type_size(t);
        if (fread(b, 1, n, f) != (size_t)n) return false;
        *v = decode(b, t);
        return true;
    };
    bool gotXYZ[3] = {false, false, false}, gotIdx = false;
    for (auto &e : elems) {
        const bool isV = e.name == "vertex", isF = e.name == "face";
        if (isV) out->p.assign(3 * (size_t)e.count, 0.f);
        for (long i = 0; i < e.count; i++)
            for (auto &p : e.props) {
                if (!p.list) {
                    double v;
                    if (!next(p.type, &v)) { *err = path + ": unable to read the contents of PLY file"; return false; }
                    if (isV) {
                        int a = p.name == "x" ? 0 : p.name == "y" ? 1 : p.name == "z" ? 2 : -1;
                        if (a >= 0) { out->p[3 * (size_t)i + a] = (float)v; gotXYZ[a] = true; }
                        if (p.name == "nx" || p.name == "ny" || p.name == "nz") out->has_normals = true;
                        if (p.name == "u" || p.name == "s" || p.name == "texture_u" || p.name == "texture_s") out->has_uv = true;
                    }
                    continue;
                }
                double cnt;
                if (!next(p.countType, &cnt)) { *err = path + ": unable to read the contents of PLY file"; return false; }
                const int n = (int)cnt;
                int face[4] = {0, 0, 0, 0};
                for (int k = 0; k < n; k++) {
                    double v;
                    if (!next(p.type, &v)) { *err = path + ": unable to read the contents of PLY file"; return false; }
                    if (k < 4) face[k] = (int)v;
                }
                if (isF && p.name == "vertex_indices") {   // (the reference looks for exactly this name, mesh.cpp:395-397)
                    gotIdx = true;
                    if (n == 3) out->tri.insert(out->tri.end(), {face[0], face[1], face[2]});
                    else if (n == 4) out->quad.insert(out->quad.end(), {face[0], face[1], face[3], face[2]});   // mesh.cpp:301-305
                    else { *err = "plymesh: Only triangles and quads are supported! Face with " + std::to_string(n) + " vertices in \"" + path + "\""; return false; }
                }
            }
    }
    if (!gotXYZ[0] || !gotXYZ[1] || !gotXYZ[2]) { *err = path + ": Vertex coordinate property not found!"; return false; }
    if (!gotIdx) { *err = path + ": vertex indices not found in PLY file"; return false; }
    for (int idx : out->tri) if (idx < 0 || idx >= nv) { *err = "plymesh: Vertex index " + std::to_string(idx) + " is out of bounds! Valid range is [0.." + std::to_string(nv) + ")"; return false; }
    for (int idx : out->quad) if (idx < 0 || idx >= nv) { *err = "plymesh: Vertex index " + std::to_string(idx) + " is out of bounds! Valid range is [0.." + std::to_string(nv) + ")"; return false; }
    return true;
}

}  // namespace vspg_host
