// prt_loader.h — loader -> device path: the host-side ingestion the reference does in Mesh::loadObj
// (mesh.cpp:151-298), Material::load (material.cpp:45-85) and Texture::load (texture.cpp:212-254), restated without the
// reference's third-party parsers (tinyobjloader, stb_image).  The outputs are the buffers the C ABI uploads
// (prt_cuda_scene_add_mesh / prt_cuda_scene_add_texture), so a scene on disk reaches the GPU through
//     Mesh m; m.loadObj("scene.obj"); Bvh* b = new Bvh; b->build(std::move(m)); scene.add(b);
// exactly as in the reference's main.cpp:58-105.
//
// Reference behaviours kept on purpose (they shape the device buffers):
//   * positions are ALL `v` records of the file; indices are the faces' position indices (mesh.cpp:266-275);
//   * texcoords are stored per POSITION index: the last face corner that references a position decides its uv, v is
//     flipped (1 - vt.y), corners without vt give (0, 0) (mesh.cpp:268-274);
//   * quads are split along the shorter diagonal, 0-2 if |v2-v0|^2 < |v3-v1|^2 else 1-3 (tiny_obj_loader.h:1476-1523);
//     polygons with more than 4 corners (tinyobj's ear clipping) are rejected here;
//   * a face before any `usemtl` gets material id -1 = 0xffffffff in the uint32 table (mesh.cpp:277);
//   * Kd defaults to 0.6 when map_Kd is given and NO material of the file has set Kd so far — tinyobj never resets its
//     has_kd flag at newmtl (tiny_obj_loader.h:1941, 2024, 2180-2184);
//   * textures: 1-component files stay 1 component, everything else becomes RGBA8; a 3-component `map_bump` is a normal
//     map and is converted to a height map with convertNormalToBump (texture.cpp:185-200, 236-244);
//     alphaTest = any alpha < 255 in the diffuse map (material.cpp:77, texture.cpp:338-350) is computed by the C ABI.
// Formats: OBJ / MTL text; TGA images (uncompressed and RLE, 8 / 24 / 32 bit, either origin) decoded as stb_image does.
// PNG / JPEG need an inflate / DCT decoder and are not restated: Texture::load throws for them.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cctype>
#include <cmath>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace prt {
namespace loader {

inline std::string dirOf(const std::string& path) {
    size_t p = path.find_last_of("/\\");
    return p == std::string::npos ? std::string(".") : path.substr(0, p);
}

inline bool readFile(const std::string& path, std::vector<uint8_t>& out) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    out.resize(n > 0 ? (size_t)n : 0);
    size_t got = out.empty() ? 0 : std::fread(out.data(), 1, out.size(), f);
    std::fclose(f);
    return got == out.size();
}

// TGA -> top-down RGB(A) / grey, as stb_image's tga loader: BGR(A) swapped to RGB(A), bottom-origin files flipped.
// comp = components in the file (1, 3 or 4).
inline bool decodeTga(const std::vector<uint8_t>& d, int& w, int& h, int& comp, std::vector<uint8_t>& px) {
    if (d.size() < 18) return false;
    const int idLen = d[0], cmapType = d[1], type = d[2], bpp = d[16], desc = d[17];
    w = d[12] | (d[13] << 8); h = d[14] | (d[15] << 8);
    if (cmapType != 0 || w <= 0 || h <= 0) return false;
    const bool rle = type == 10 || type == 11, grey = type == 3 || type == 11;
    if (!(type == 2 || type == 3 || rle)) return false;
    if (grey ? bpp != 8 : (bpp != 24 && bpp != 32)) return false;
    comp = bpp / 8;
    px.assign((size_t)w * h * comp, 0);
    size_t p = 18 + (size_t)idLen;
    const size_t n = (size_t)w * h;
    uint8_t raw[4] = {0, 0, 0, 0};
    int rleCount = 0; bool rleRepeat = false, readNext = true;
    for (size_t i = 0; i < n; i++) {
        if (rle) {
            if (rleCount == 0) {
                if (p >= d.size()) return false;
                int c = d[p++];
                rleCount = 1 + (c & 127); rleRepeat = (c >> 7) != 0; readNext = true;
            } else if (!rleRepeat) readNext = true;
        } else readNext = true;
        if (readNext) {
            if (p + comp > d.size()) return false;
            for (int k = 0; k < comp; k++) raw[k] = d[p++];
            readNext = false;
        }
        for (int k = 0; k < comp; k++) px[i * comp + k] = raw[k];
        rleCount--;
    }
    if (!((desc >> 5) & 1)) {                                     // bottom-left origin: flip rows
        const size_t row = (size_t)w * comp;
        std::vector<uint8_t> tmp(row);
        for (int y = 0; y * 2 < h; y++) {
            uint8_t* a = &px[(size_t)y * row]; uint8_t* b = &px[(size_t)(h - 1 - y) * row];
            std::memcpy(tmp.data(), a, row); std::memcpy(a, b, row); std::memcpy(b, tmp.data(), row);
        }
    }
    if (comp >= 3) for (size_t i = 0; i < n; i++) { uint8_t t = px[i * comp]; px[i * comp] = px[i * comp + 2]; px[i * comp + 2] = t; }
    return true;
}

// convertNormalToBump (texture.cpp:185-200): RGBA8 normal map -> 8-bit height, 0xff * clamp(n.z^4, 0, 1), n normalised.
inline void convertNormalToBump(uint8_t* bump, const uint8_t* normal, uint32_t width, uint32_t height) {
    const float kChannelScale = 1.0f / 255.0f;
    for (uint32_t y = 0; y < height; y++)
        for (uint32_t x = 0; x < width; x++) {
            const uint32_t index = 4 * (x + y * width);
            float nx = 2.0f * (normal[index + 0] * kChannelScale - 0.5f);
            float ny = 2.0f * (normal[index + 1] * kChannelScale - 0.5f);
            float nz = 2.0f * (normal[index + 2] * kChannelScale - 0.5f);
            float xx = nx * nx, yy = ny * ny, zz = nz * nz;
            float inv = 1.0f / std::sqrt(xx + yy + zz);           // normalize (vecmath.h:1200-1204)
            float z = inv * nz;
            float c = std::fmin(std::fmax(z * z * z * z, 0.0f), 1.0f);
            bump[x + y * width] = static_cast<uint8_t>(0xff * c);
        }
}

// ---- MTL / OBJ text (the subset Mesh::loadObj consumes) ----
struct MtlData {
    std::string name;
    float diffuse[3] = {0, 0, 0}, ambient[3] = {0, 0, 0}, specular[3] = {0, 0, 0}, emission[3] = {0, 0, 0};
    std::string diffuseTex, bumpTex;
};
struct ObjData {
    std::vector<float> positions;           // 3 per `v`
    std::vector<float> texcoords;           // 2 per position (see header comment)
    std::vector<uint32_t> indices;          // 3 per triangle, position indices
    std::vector<uint32_t> primMaterial;     // per triangle, 0xffffffff = none
    std::vector<MtlData> materials;
};

inline const char* skipWs(const char* p) { while (*p == ' ' || *p == '\t') p++; return p; }
inline bool isWs(char c) { return c == ' ' || c == '\t'; }
inline float parseFloat(const char*& p) { p = skipWs(p); char* e = nullptr; double v = std::strtod(p, &e); if (e == p) { while (*p && !isWs(*p)) p++; return 0.0f; } p = e; return (float)v; }
inline std::string parseToken(const char*& p) { p = skipWs(p); const char* b = p; while (*p && !isWs(*p) && *p != '\r' && *p != '\n') p++; return std::string(b, p); }
// texture statement: options (-bm 1.0, -o u v w, -s ..., -clamp on, ...) precede the file name (tiny_obj_loader.h ParseTextureNameAndOption)
inline std::string parseTexName(const char* p) {
    std::string name;
    for (;;) {
        std::string t = parseToken(p);
        if (t.empty()) break;
        if (t[0] == '-') {
            int nargs = (t == "-o" || t == "-s" || t == "-t") ? 3 : (t == "-mm" ? 2 : 1);
            for (int i = 0; i < nargs; i++) {
                const char* save = p; std::string a = parseToken(p);
                if (i > 0 && !a.empty() && !(std::isdigit((unsigned char)a[0]) || a[0] == '-' || a[0] == '+' || a[0] == '.')) { p = save; break; }   // optional v / w
            }
            continue;
        }
        name = t;
        const char* q = skipWs(p);                                // file names may contain blanks: take the rest of the line
        std::string rest(q);
        while (!rest.empty() && (rest.back() == '\r' || rest.back() == '\n' || isWs(rest.back()))) rest.pop_back();
        if (!rest.empty()) name += " " + rest;
        break;
    }
    return name;
}

inline void parseMtl(const std::string& path, std::vector<MtlData>& out, std::map<std::string, int>& byName) {
    std::vector<uint8_t> buf;
    if (!readFile(path, buf)) return;                             // tinyobj only warns when the .mtl is missing
    buf.push_back(0);
    MtlData cur; bool have = false, hasKd = false;
    auto flush = [&]() { if (have) { byName[cur.name] = (int)out.size(); out.push_back(cur); } };
    const char* p = (const char*)buf.data();
    while (*p) {
        const char* eol = p; while (*eol && *eol != '\n') eol++;
        std::string line(p, eol); p = *eol ? eol + 1 : eol;
        while (!line.empty() && (line.back() == '\r' || isWs(line.back()))) line.pop_back();
        const char* t = skipWs(line.c_str());
        if (!*t || *t == '#') continue;
        auto key = [&](const char* k) { size_t n = std::strlen(k); return std::strncmp(t, k, n) == 0 && isWs(t[n]); };
        auto rgb = [&](float* d, size_t skip) { const char* q = t + skip; d[0] = parseFloat(q); d[1] = parseFloat(q); d[2] = parseFloat(q); };
        if (key("newmtl")) { flush(); cur = MtlData(); const char* q = t + 6; cur.name = parseToken(q); have = true; }   // hasKd is NOT reset: tinyobj keeps it across materials
        else if (key("Ka")) rgb(cur.ambient, 2);
        else if (key("Kd")) { rgb(cur.diffuse, 2); hasKd = true; }
        else if (key("Ks")) rgb(cur.specular, 2);
        else if (key("Ke")) rgb(cur.emission, 2);
        else if (key("map_Kd")) {
            cur.diffuseTex = parseTexName(t + 6);
            if (!hasKd) { cur.diffuse[0] = cur.diffuse[1] = cur.diffuse[2] = 0.6f; }      // tiny_obj_loader.h:2180-2184
        }
        else if (key("map_bump") || key("map_Bump")) cur.bumpTex = parseTexName(t + 8);
        else if (key("bump")) cur.bumpTex = parseTexName(t + 4);
    }
    flush();
}

inline void parseObj(const std::string& path, ObjData& o) {
    std::vector<uint8_t> buf;
    if (!readFile(path, buf)) throw std::runtime_error("Failed to load " + path);
    buf.push_back(0);
    std::vector<float> vt;                                        // 2 per `vt`
    std::map<std::string, int> byName;
    struct Corner { int v, t; };
    std::vector<std::vector<Corner>> faces; std::vector<uint32_t> faceMat;
    uint32_t curMat = 0xffffffffu;
    const char* p = (const char*)buf.data();
    while (*p) {
        const char* eol = p; while (*eol && *eol != '\n') eol++;
        std::string line(p, eol); p = *eol ? eol + 1 : eol;
        const char* t = skipWs(line.c_str());
        if (t[0] == 'v' && isWs(t[1])) { const char* q = t + 1; for (int k = 0; k < 3; k++) o.positions.push_back(parseFloat(q)); }
        else if (t[0] == 'v' && t[1] == 't' && isWs(t[2])) { const char* q = t + 2; vt.push_back(parseFloat(q)); vt.push_back(parseFloat(q)); }
        else if (t[0] == 'f' && isWs(t[1])) {
            const char* q = t + 1; std::vector<Corner> f;
            for (;;) {
                std::string tok = parseToken(q);
                if (tok.empty()) break;
                Corner c{0, 0}; bool hasT = false;                // v, v/vt, v//vn, v/vt/vn; 1-based, negative = relative to the end
                const char* s = tok.c_str(); char* e = nullptr;
                long vi = std::strtol(s, &e, 10);
                long ti = 0;
                if (*e == '/') { s = e + 1; if (*s != '/' && *s) { ti = std::strtol(s, &e, 10); hasT = true; } }
                c.v = vi > 0 ? (int)vi - 1 : (int)(o.positions.size() / 3) + (int)vi;
                c.t = hasT ? (ti > 0 ? (int)ti - 1 : (int)(vt.size() / 2) + (int)ti) : -1;
                f.push_back(c);
            }
            if (f.size() >= 3) { faces.push_back(f); faceMat.push_back(curMat); }
        }
        else if (std::strncmp(t, "usemtl", 6) == 0 && isWs(t[6])) {
            const char* q = t + 6; std::string name = parseToken(q);
            auto it = byName.find(name);
            curMat = it == byName.end() ? 0xffffffffu : (uint32_t)it->second;
        }
        else if (std::strncmp(t, "mtllib", 6) == 0 && isWs(t[6])) {
            const char* q = t + 6;
            for (;;) { std::string name = parseToken(q); if (name.empty()) break; parseMtl(dirOf(path) + "/" + name, o.materials, byName); }
        }
    }
    // faces are triangulated once every `v` is known: like tinyobj, which triangulates when a shape is flushed
    const uint32_t nv = (uint32_t)(o.positions.size() / 3);
    o.texcoords.assign(2 * (size_t)nv, 0.0f);
    auto emit = [&](const Corner& c) {
        o.indices.push_back((uint32_t)c.v);
        if ((uint32_t)c.v >= nv) return;
        if (c.t >= 0) { o.texcoords[2 * (size_t)c.v] = vt[2 * (size_t)c.t]; o.texcoords[2 * (size_t)c.v + 1] = 1.0f - vt[2 * (size_t)c.t + 1]; }
        else { o.texcoords[2 * (size_t)c.v] = 0.0f; o.texcoords[2 * (size_t)c.v + 1] = 0.0f; }
    };
    for (size_t fi = 0; fi < faces.size(); fi++) {
        const auto& f = faces[fi];
        if (f.size() == 3) { emit(f[0]); emit(f[1]); emit(f[2]); o.primMaterial.push_back(faceMat[fi]); }
        else if (f.size() == 4) {
            for (const auto& c : f) if ((uint32_t)c.v >= nv) throw std::runtime_error("loadObj: face with invalid vertex index");
            const float* v0 = &o.positions[3 * (size_t)f[0].v]; const float* v1 = &o.positions[3 * (size_t)f[1].v];
            const float* v2 = &o.positions[3 * (size_t)f[2].v]; const float* v3 = &o.positions[3 * (size_t)f[3].v];
            float e02x = v2[0] - v0[0], e02y = v2[1] - v0[1], e02z = v2[2] - v0[2];
            float e13x = v3[0] - v1[0], e13y = v3[1] - v1[1], e13z = v3[2] - v1[2];
            float sqr02 = e02x * e02x + e02y * e02y + e02z * e02z, sqr13 = e13x * e13x + e13y * e13y + e13z * e13z;
            if (sqr02 < sqr13) { emit(f[0]); emit(f[1]); emit(f[2]); emit(f[0]); emit(f[2]); emit(f[3]); }
            else { emit(f[0]); emit(f[1]); emit(f[3]); emit(f[1]); emit(f[2]); emit(f[3]); }
            o.primMaterial.push_back(faceMat[fi]); o.primMaterial.push_back(faceMat[fi]);
        }
        else throw std::runtime_error("loadObj: polygons with more than 4 corners are not supported (tinyobj ear clipping is not restated)");
    }
}

}  // namespace loader
}  // namespace prt
