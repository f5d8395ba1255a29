// prt_loader.h — loader -> device path: the host-side ingestion the reference does in Mesh::loadObj
// (mesh.cpp:151-298), Material::load (material.cpp:45-85) and Texture::load (texture.cpp:212-254), restated without the
// reference's third-party parsers (tinyobjloader, stb_image).  The outputs are the buffers the C ABI uploads
// (prt_cuda_scene_add_mesh / prt_cuda_scene_add_texture), so a scene on disk reaches the GPU through
//     Mesh m; m.loadObj("scene.obj"); Bvh* b = new Bvh; b->build(std::move(m)); scene.add(b);
// exactly as in the reference's main.cpp:58-105.
//
// Reference behaviours kept on purpose (they shape the device buffers):
//   * numbers are read with tinyobj's own decimal reader, whose double arithmetic is not correctly rounded (parseFloat below);
//   * positions are ALL `v` records of the file; indices are the faces' position indices (mesh.cpp:266-275);
//   * texcoords are stored per POSITION index: the last face corner that references a position decides its uv, v is
//     flipped (1 - vt.y), corners without vt give (0, 0) (mesh.cpp:268-274);
//   * quads are split along the shorter diagonal, 0-2 if |v2-v0|^2 < |v3-v1|^2 else 1-3 (tiny_obj_loader.h:1476-1523);
//     polygons with more than 4 corners go through tinyobj's ear clipping, restated in fp32 (tiny_obj_loader.h:1536-1820);
//   * a face before any `usemtl` gets material id -1 = 0xffffffff in the uint32 table (mesh.cpp:277);
//   * Kd defaults to 0.6 when map_Kd is given and NO material of the file has set Kd so far — tinyobj never resets its
//     has_kd flag at newmtl (tiny_obj_loader.h:1941, 2024, 2180-2184);
//   * textures: 1-component files stay 1 component, everything else becomes RGBA8; a 3-component `map_bump` is a normal
//     map and is converted to a height map with convertNormalToBump (texture.cpp:185-200, 236-244);
//     alphaTest = any alpha < 255 in the diffuse map (material.cpp:77, texture.cpp:338-350) is computed by the C ABI.
// Formats: OBJ / MTL text; TGA images (uncompressed and RLE, 8 / 24 / 32 bit, either origin) and PNG images (grey, grey+alpha,
// RGB, RGBA and palette at every legal bit depth, tRNS transparency, all five row filters, Adam7; own inflate) decoded to what stb_image hands
// Texture::load; scan-line OpenEXR (HALF / FLOAT, NONE / RLE / ZIPS / ZIP / PIZ) decoded to what tinyexr hands Texture::loadExr.
// sequential and progressive JPEG decoded to stb_image's bytes (its IDCT, up-sampling and colour conversion).  Tiled / multi-part EXRs and
// arithmetic-coded JPEGs are not restated: the loaders throw for them.
#pragma once
#include <algorithm>
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

// ---- PNG (RFC 2083) ----
// RFC 1951 inflate of a zlib stream (RFC 1950 header skipped, Adler-32 not checked): stored, fixed and dynamic Huffman blocks.
class Inflater {
public:
    Inflater(const uint8_t* d, size_t n) : m_d(d), m_n(n) {}
    bool run(std::vector<uint8_t>& out, size_t expected) {
        out.clear(); out.reserve(expected);
        if (m_n < 2 || (m_d[0] & 15) != 8 || ((m_d[0] << 8) | m_d[1]) % 31 != 0 || (m_d[1] & 32)) return false;
        m_p = 2;
        for (;;) {
            const uint32_t last = bits(1), type = bits(2);
            if (m_fail) return false;
            if (type == 0) {
                m_acc = 0; m_cnt = 0;                                  // to the next byte boundary
                if (m_p + 4 > m_n) return false;
                const uint32_t len = m_d[m_p] | (m_d[m_p + 1] << 8), nlen = m_d[m_p + 2] | (m_d[m_p + 3] << 8);
                m_p += 4;
                if ((len ^ 0xffffu) != nlen || m_p + len > m_n) return false;
                out.insert(out.end(), m_d + m_p, m_d + m_p + len);
                m_p += len;
            } else if (type == 1 || type == 2) {
                Table lit, dist;
                if (type == 1) {
                    uint8_t l[288], dd[30];
                    for (int i = 0; i < 288; i++) l[i] = i < 144 ? 8 : (i < 256 ? 9 : (i < 280 ? 7 : 8));
                    for (int i = 0; i < 30; i++) dd[i] = 5;
                    build(lit, l, 288); build(dist, dd, 30);
                } else if (!dynamicTables(lit, dist)) return false;
                if (!block(lit, dist, out)) return false;
            } else return false;
            if (last) break;
        }
        return !m_fail;
    }
private:
    struct Table { uint16_t count[16]; uint16_t symbol[288]; };
    uint32_t bits(int n) {
        while (m_cnt < n) {
            if (m_p >= m_n) { m_fail = true; return 0; }
            m_acc |= (uint32_t)m_d[m_p++] << m_cnt; m_cnt += 8;
        }
        const uint32_t v = m_acc & ((1u << n) - 1u);
        m_acc >>= n; m_cnt -= n;
        return v;
    }
    static void build(Table& t, const uint8_t* len, int n) {            // canonical code: symbols ordered by (length, value)
        std::memset(t.count, 0, sizeof(t.count));
        for (int i = 0; i < n; i++) t.count[len[i]]++;
        t.count[0] = 0;
        uint16_t offs[16]; offs[1] = 0;
        for (int l = 1; l < 15; l++) offs[l + 1] = (uint16_t)(offs[l] + t.count[l]);
        for (int i = 0; i < n; i++) if (len[i]) t.symbol[offs[len[i]]++] = (uint16_t)i;
    }
    int decode(const Table& t) {                                        // one bit at a time along the canonical code
        int code = 0, first = 0, index = 0;
        for (int l = 1; l <= 15; l++) {
            code |= (int)bits(1);
            if (m_fail) return -1;
            const int c = t.count[l];
            if (code - c < first) return t.symbol[index + (code - first)];
            index += c; first += c; first <<= 1; code <<= 1;
        }
        return -1;
    }
    bool dynamicTables(Table& lit, Table& dist) {
        static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
        const int nlit = (int)bits(5) + 257, ndist = (int)bits(5) + 1, ncode = (int)bits(4) + 4;
        if (m_fail || nlit > 286 || ndist > 30) return false;
        uint8_t cl[19] = {0};
        for (int i = 0; i < ncode; i++) cl[order[i]] = (uint8_t)bits(3);
        Table ct; build(ct, cl, 19);
        uint8_t len[320] = {0};
        for (int i = 0; i < nlit + ndist;) {
            const int sym = decode(ct);
            if (sym < 0) return false;
            if (sym < 16) { len[i++] = (uint8_t)sym; continue; }
            int rep, val = 0;
            if (sym == 16) { if (i == 0) return false; val = len[i - 1]; rep = 3 + (int)bits(2); }
            else if (sym == 17) rep = 3 + (int)bits(3);
            else rep = 11 + (int)bits(7);
            if (m_fail || i + rep > nlit + ndist) return false;
            while (rep--) len[i++] = (uint8_t)val;
        }
        if (len[256] == 0) return false;
        build(lit, len, nlit); build(dist, len + nlit, ndist);
        return true;
    }
    bool block(const Table& lit, const Table& dist, std::vector<uint8_t>& out) {
        static const uint16_t lbase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
        static const uint8_t lext[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
        static const uint16_t dbase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
        static const uint8_t dext[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
        for (;;) {
            const int sym = decode(lit);
            if (sym < 0) return false;
            if (sym < 256) { out.push_back((uint8_t)sym); continue; }
            if (sym == 256) return true;
            if (sym > 285) return false;
            const size_t len = lbase[sym - 257] + bits(lext[sym - 257]);
            const int ds = decode(dist);
            if (ds < 0 || ds > 29) return false;
            const size_t d = dbase[ds] + bits(dext[ds]);
            if (m_fail || d > out.size()) return false;
            for (size_t k = 0; k < len; k++) out.push_back(out[out.size() - d]);
        }
    }
    const uint8_t* m_d; size_t m_n, m_p = 0;
    uint32_t m_acc = 0; int m_cnt = 0; bool m_fail = false;
};

// PNG -> top-down pixels.  comp = the component count Texture::load works with (texture.cpp:218-236): grey 1, grey+alpha 2,
// RGB 3, RGBA 4, palette 3 (4 with a tRNS chunk), RGB with a tRNS colour key 4 (stbi_load counts the key's alpha, stb_image.h
// "++s->img_n", so such a file is never taken for a normal map) — but grey with a colour key stays 1: the reference sizes its
// request from stbi_info, whose header scan stops at IHDR for non-palette files, asks for 1 component and the key's alpha is
// dropped.  pxComp = channels stored in px.  Bit depths 1 / 2 / 4 (grey, palette), 8 and 16 (reduced to the high byte);
// Adam7-interlaced files too.
inline bool decodePng(const std::vector<uint8_t>& d, int& w, int& h, int& comp, int& pxComp, std::vector<uint8_t>& px) {
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (d.size() < 8 + 25 || std::memcmp(d.data(), sig, 8) != 0) return false;
    auto be32 = [&](size_t p) { return ((uint32_t)d[p] << 24) | ((uint32_t)d[p + 1] << 16) | ((uint32_t)d[p + 2] << 8) | d[p + 3]; };
    int depth = 0, ctype = -1, interlace = 0;
    std::vector<uint8_t> idat, plte, trns;
    bool haveTrns = false;
    for (size_t p = 8; p + 12 <= d.size();) {
        const uint32_t len = be32(p);
        const char* t = (const char*)&d[p + 4];
        if (p + 12 + (size_t)len > d.size()) return false;
        const uint8_t* body = &d[p + 8];
        if (!std::memcmp(t, "IHDR", 4)) {
            if (len != 13) return false;
            w = (int)be32(p + 8); h = (int)be32(p + 12); depth = body[8]; ctype = body[9]; interlace = body[12];
            if (body[10] != 0 || body[11] != 0) return false;
        } else if (!std::memcmp(t, "PLTE", 4)) plte.assign(body, body + len);
        else if (!std::memcmp(t, "tRNS", 4)) { trns.assign(body, body + len); haveTrns = true; }
        else if (!std::memcmp(t, "IDAT", 4)) idat.insert(idat.end(), body, body + len);
        else if (!std::memcmp(t, "IEND", 4)) break;
        p += 12 + (size_t)len;
    }
    if (w <= 0 || h <= 0 || interlace > 1) return false;
    int fileComp;                                                   // samples per pixel in the file
    switch (ctype) { case 0: fileComp = 1; break; case 2: fileComp = 3; break; case 3: fileComp = 1; break; case 4: fileComp = 2; break; case 6: fileComp = 4; break; default: return false; }
    if (!(depth == 8 || depth == 16 || ((ctype == 0 || ctype == 3) && (depth == 1 || depth == 2 || depth == 4))) || (ctype == 3 && depth == 16)) return false;
    if (ctype == 3 && (plte.empty() || plte.size() % 3 != 0)) return false;
    // Sub-images: the whole picture, or the seven Adam7 passes (RFC 2083 section 2.6), each filtered on its own.
    static const int xo[7] = {0, 4, 0, 2, 0, 1, 0}, yo[7] = {0, 0, 4, 0, 2, 0, 1}, xs[7] = {8, 8, 4, 4, 2, 2, 1}, ys[7] = {8, 8, 8, 4, 4, 2, 2};
    const int passes = interlace ? 7 : 1;
    size_t total = 0;
    for (int ps = 0; ps < passes; ps++) {
        const int pw = interlace ? (w - xo[ps] + xs[ps] - 1) / xs[ps] : w, ph = interlace ? (h - yo[ps] + ys[ps] - 1) / ys[ps] : h;
        if (pw > 0 && ph > 0) total += ((((size_t)pw * fileComp * depth) + 7) / 8 + 1) * ph;
    }
    std::vector<uint8_t> raw;
    if (!Inflater(idat.data(), idat.size()).run(raw, total) || raw.size() < total) return false;
    const int fb = depth < 8 ? 1 : fileComp * depth / 8;            // filter distance in bytes
    std::vector<uint16_t> samp((size_t)w * h * fileComp);           // raw sample values, full picture
    std::vector<uint8_t> prev, cur;
    size_t rp = 0;
    for (int ps = 0; ps < passes; ps++) {
        const int pw = interlace ? (w - xo[ps] + xs[ps] - 1) / xs[ps] : w, ph = interlace ? (h - yo[ps] + ys[ps] - 1) / ys[ps] : h;
        if (pw <= 0 || ph <= 0) continue;
        const size_t stride = (((size_t)pw * fileComp * depth) + 7) / 8;
        prev.assign(stride, 0); cur.assign(stride, 0);
        for (int y = 0; y < ph; y++) {
            const uint8_t* in = &raw[rp]; rp += stride + 1;
            const int f = in[0];
            if (f > 4) return false;
            for (size_t x = 0; x < stride; x++) {                   // undo the row filter: a = left, b = above, c = above left
                const int a = x >= (size_t)fb ? cur[x - fb] : 0, b = prev[x], c = x >= (size_t)fb ? prev[x - fb] : 0;
                int pred = 0;
                if (f == 1) pred = a;
                else if (f == 2) pred = b;
                else if (f == 3) pred = (a + b) >> 1;
                else if (f == 4) { const int pp = a + b - c, pa = std::abs(pp - a), pb = std::abs(pp - b), pc = std::abs(pp - c); pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); }
                cur[x] = (uint8_t)(in[1 + x] + pred);
            }
            const int Y = interlace ? yo[ps] + y * ys[ps] : y;
            for (int x = 0; x < pw; x++) {
                const int X = interlace ? xo[ps] + x * xs[ps] : x;
                uint16_t* o = &samp[((size_t)Y * w + X) * fileComp];
                for (int k = 0; k < fileComp; k++) {
                    const size_t si = (size_t)x * fileComp + k;     // sample index within the row
                    if (depth == 8) o[k] = cur[si];
                    else if (depth == 16) o[k] = (uint16_t)((cur[2 * si] << 8) | cur[2 * si + 1]);
                    else { const size_t bit = si * depth; o[k] = (uint16_t)((cur[bit >> 3] >> (8 - depth - (bit & 7))) & ((1 << depth) - 1)); }   // packed, leftmost pixel in the high bits
                }
            }
            prev.swap(cur);
        }
    }
    // to 8 bits per sample as stb_image does: 16-bit samples keep their high byte, 1 / 2 / 4-bit grey is scaled to 0..255, palette indices stay
    const size_t nsamp = samp.size();
    std::vector<uint8_t> img(nsamp);
    const int scale = (ctype == 0 && depth < 8) ? (depth == 1 ? 0xff : (depth == 2 ? 0x55 : 0x11)) : 1;
    for (size_t i = 0; i < nsamp; i++) img[i] = depth == 16 ? (uint8_t)(samp[i] >> 8) : (uint8_t)(samp[i] * scale);
    const size_t n = (size_t)w * h;
    if (ctype == 3) {                                               // palette -> RGB, or RGBA when the file has a tRNS chunk
        comp = pxComp = haveTrns ? 4 : 3;
        px.resize(n * comp);
        for (size_t i = 0; i < n; i++) {
            const size_t k = img[i];
            if (3 * k + 2 >= plte.size()) return false;
            px[comp * i] = plte[3 * k]; px[comp * i + 1] = plte[3 * k + 1]; px[comp * i + 2] = plte[3 * k + 2];
            if (haveTrns) px[comp * i + 3] = k < trns.size() ? trns[k] : 255;
        }
    } else if (haveTrns && ctype == 2) {                            // RGB colour key: alpha 0 where the pixel equals the key
        if (trns.size() < 6) return false;
        comp = pxComp = 4;
        px.resize(n * 4);
        for (size_t i = 0; i < n; i++) {
            bool key = true;                                       // the key is compared at the file's depth (all 16 bits of a 16-bit sample)
            for (int k = 0; k < 3; k++) {
                px[4 * i + k] = img[3 * i + k];
                const uint16_t kv = depth == 16 ? (uint16_t)((trns[2 * k] << 8) | trns[2 * k + 1]) : trns[2 * k + 1];
                key = key && samp[3 * i + k] == kv;
            }
            px[4 * i + 3] = key ? 0 : 255;
        }
    } else { comp = pxComp = fileComp; px.swap(img); }
    return true;
}

// ---- OpenEXR (what Texture::loadExr gets from tinyexr's LoadEXRFromMemory, texture.cpp:256-336): single-part scan-line
// files, HALF or FLOAT channels, compression NONE / RLE / ZIPS / ZIP / PIZ -> RGBA32F, row 0 = top of the data window.  Channel
// rules as tinyexr's LoadEXR: one channel is replicated into all four components; otherwise R, G and B are required and a
// missing A reads 1.0.  Tiled, multi-part, deep, PXR24 / B44 / DWA files are refused. ----
inline float halfToFloat(uint16_t h) {
    const uint32_t s = (uint32_t)(h >> 15) << 31, e = (h >> 10) & 31u, m = h & 1023u;
    uint32_t u;
    if (e == 0) {
        if (m == 0) u = s;
        else { int k = 0; uint32_t mm = m; while (!(mm & 1024u)) { mm <<= 1; k++; } u = s | ((uint32_t)(113 - k) << 23) | ((mm & 1023u) << 13); }   // subnormal half
    } else if (e == 31) u = s | 0x7f800000u | (m << 13);
    else u = s | ((e + 112u) << 23) | (m << 13);
    float f; std::memcpy(&f, &u, 4); return f;
}
// PIZ block (OpenEXR's wavelet + Huffman scheme): a bitmap of the 16-bit values in use maps them to a dense range, every channel's
// 16-bit planes go through a 2-D Haar-like wavelet (14-bit arithmetic when the dense range allows, modulo-16-bit otherwise), and the
// coefficients are Huffman coded (canonical codes of up to 58 bits, a run-length symbol).  words: per channel 1 (HALF) or 2 (FLOAT).
inline bool decodePizBlock(const uint8_t* src, size_t n, int width, int lines, const std::vector<int>& words, std::vector<uint8_t>& out) {
    auto u16 = [&](size_t p) { return (uint32_t)src[p] | ((uint32_t)src[p + 1] << 8); };
    auto u32 = [&](size_t p) { return u16(p) | (u16(p + 2) << 16); };
    if (n < 8) return false;
    const uint32_t minNz = u16(0), maxNz = u16(2);
    if (maxNz >= 8192) return false;
    std::vector<uint8_t> bitmap(8192, 0);
    size_t p = 4;
    if (minNz <= maxNz) { if (p + (maxNz - minNz + 1) > n) return false; std::memcpy(&bitmap[minNz], src + p, maxNz - minNz + 1); p += maxNz - minNz + 1; }
    std::vector<uint16_t> lut(65536, 0);
    int k = 0;
    for (int i = 0; i < 65536; i++) if (i == 0 || (bitmap[i >> 3] & (1 << (i & 7)))) lut[k++] = (uint16_t)i;
    const int maxValue = k - 1;
    if (p + 4 > n) return false;
    const size_t hlen = u32(p); p += 4;
    if (p + hlen > n || hlen < 20) return false;
    const uint8_t* hp = src + p;
    auto h32 = [&](size_t q) { return (uint32_t)hp[q] | ((uint32_t)hp[q + 1] << 8) | ((uint32_t)hp[q + 2] << 16) | ((uint32_t)hp[q + 3] << 24); };
    const uint32_t im = h32(0), iM = h32(4), nBits = h32(12);
    if (im > 65536 || iM > 65536 || im > iM) return false;
    // code lengths: 6 bits each, MSB first; 63 = long zero run (8 more bits, + 6), 59..62 = short zero run (2..5)
    std::vector<uint8_t> len(65537, 0);
    size_t q = 20; uint64_t acc = 0; int cnt = 0;
    auto bits = [&](int nb) -> int { while (cnt < nb) { acc = (acc << 8) | (q < hlen ? hp[q] : 0); q++; cnt += 8; } cnt -= nb; return (int)((acc >> cnt) & ((1u << nb) - 1u)); };
    for (uint32_t i = im; i <= iM; i++) {
        if (q > hlen) return false;
        const int l = bits(6);
        if (l == 63) { int z = bits(8) + 6; if (i + z > iM + 1) return false; i += z - 1; }
        else if (l >= 59) { int z = l - 59 + 2; if (i + z > iM + 1) return false; i += z - 1; }
        else len[i] = (uint8_t)l;
    }
    if (q > hlen || (size_t)nBits > 8 * (hlen - q)) return false;
    // canonical codes: the longest length starts at 0; within a length, codes ascend with the symbol index
    uint64_t first[60] = {0}, count[60] = {0};
    for (int i = 0; i <= 65536; i++) count[len[i]]++;
    count[0] = 0;
    { uint64_t c = 0; for (int l = 58; l > 0; l--) { const uint64_t nc = (c + count[l]) >> 1; first[l] = c; c = nc; } }
    std::vector<uint32_t> order; order.reserve(65537);
    size_t start[60] = {0};
    for (int l = 1; l <= 58; l++) { start[l] = order.size(); for (int i = 0; i <= 65536; i++) if (len[i] == l) order.push_back((uint32_t)i); }
    size_t total = 0;
    for (int wd : words) total += (size_t)wd * width * lines;
    std::vector<uint16_t> buf(total, 0);
    {
        const uint8_t* in = hp + q; size_t bitPos = 0, o = 0;
        auto bit1 = [&]() -> int { const int b = (in[bitPos >> 3] >> (7 - (bitPos & 7))) & 1; bitPos++; return b; };
        while (bitPos < nBits) {
            uint64_t code = 0; int sym = -1;
            for (int l = 1; l <= 58 && bitPos < nBits; l++) {
                code = (code << 1) | (uint64_t)bit1();
                if (count[l] && code >= first[l] && code - first[l] < count[l]) { sym = (int)order[start[l] + (size_t)(code - first[l])]; break; }
            }
            if (sym < 0) return false;
            if ((uint32_t)sym == iM) {                                          // run: 8-bit count of copies of the previous value
                if (bitPos + 8 > nBits || o == 0) return false;
                int c = 0; for (int i = 0; i < 8; i++) c = (c << 1) | bit1();
                if (o + c > total) return false;
                for (int i = 0; i < c; i++) { buf[o] = buf[o - 1]; o++; }
            } else { if (o >= total) return false; buf[o++] = (uint16_t)sym; }
        }
        if (o != total) return false;
    }
    // inverse wavelet per channel and 16-bit plane
    const bool w14 = maxValue < (1 << 14);
    auto dec = [&](uint16_t l, uint16_t hh, uint16_t& a, uint16_t& b) {
        if (w14) { const int hi = (int16_t)hh; const int ai = (int16_t)l + (hi & 1) + (hi >> 1); a = (uint16_t)(int16_t)ai; b = (uint16_t)(int16_t)(ai - hi); }
        else { const int bb = ((int)l - ((int)hh >> 1)) & 0xffff; const int aa = ((int)hh + bb - 0x8000) & 0xffff; b = (uint16_t)bb; a = (uint16_t)aa; }
    };
    size_t base = 0;
    for (int wd : words) {
        for (int j = 0; j < wd; j++) {
            uint16_t* in = &buf[base + j];
            const int nx = width, ox = wd, ny = lines, oy = width * wd;
            const int nmin = nx > ny ? ny : nx;
            int pw = 1; while (pw <= nmin) pw <<= 1;
            pw >>= 1; int p2 = pw; pw >>= 1;
            while (pw >= 1) {
                const int oy1 = oy * pw, oy2 = oy * p2, ox1 = ox * pw, ox2 = ox * p2;
                int y = 0;
                for (; y <= ny - p2; y += p2) {
                    uint16_t* py = in + (size_t)oy * y;
                    int x = 0;
                    for (; x <= nx - p2; x += p2) {
                        uint16_t* px = py + (size_t)ox * x;
                        uint16_t *p01 = px + ox1, *p10 = px + oy1, *p11 = p10 + ox1, i00, i01, i10, i11;
                        dec(*px, *p10, i00, i10); dec(*p01, *p11, i01, i11);
                        dec(i00, i01, *px, *p01); dec(i10, i11, *p10, *p11);
                    }
                    if (nx & pw) { uint16_t* px = py + (size_t)ox * x; uint16_t* p10 = px + oy1; uint16_t i00; dec(*px, *p10, i00, *p10); *px = i00; }
                }
                if (ny & pw) {
                    uint16_t* py = in + (size_t)oy * y;
                    for (int x = 0; x <= nx - p2; x += p2) { uint16_t* px = py + (size_t)ox * x; uint16_t* p01 = px + ox1; uint16_t i00; dec(*px, *p01, i00, *p01); *px = i00; }
                }
                (void)oy2; (void)ox2;
                p2 = pw; pw >>= 1;
            }
        }
        base += (size_t)wd * width * lines;
    }
    for (auto& v : buf) v = lut[v];
    // channel-major planes -> the scan-line layout of an uncompressed block (per line: channel after channel)
    out.resize(total * 2);
    size_t o = 0;
    std::vector<size_t> cpos(words.size());
    { size_t b2 = 0; for (size_t c = 0; c < words.size(); c++) { cpos[c] = b2; b2 += (size_t)words[c] * width * lines; } }
    for (int y = 0; y < lines; y++)
        for (size_t c = 0; c < words.size(); c++) {
            const size_t cnt2 = (size_t)words[c] * width;
            for (size_t i = 0; i < cnt2; i++) { const uint16_t v = buf[cpos[c] + i]; out[o++] = (uint8_t)(v & 255); out[o++] = (uint8_t)(v >> 8); }
            cpos[c] += cnt2;
        }
    return true;
}

inline bool decodeExr(const std::vector<uint8_t>& d, int& w, int& h, std::vector<float>& rgba) {
    auto u32 = [&](size_t p) { return (uint32_t)d[p] | ((uint32_t)d[p + 1] << 8) | ((uint32_t)d[p + 2] << 16) | ((uint32_t)d[p + 3] << 24); };
    if (d.size() < 8 || u32(0) != 20000630u) return false;
    const uint32_t ver = u32(4);
    if ((ver & 0xffu) != 2u || (ver & 0x1a00u)) return false;                 // tiled / deep / multi-part
    struct Chan { std::string name; int type; };
    std::vector<Chan> chans;
    int compression = -1, xmin = 0, ymin = 0, xmax = -1, ymax = -1;
    size_t p = 8;
    for (;;) {                                                                // attributes: name\0 type\0 size value
        if (p >= d.size()) return false;
        if (d[p] == 0) { p++; break; }
        std::string name, type;
        while (p < d.size() && d[p]) name += (char)d[p++];
        p++;
        while (p < d.size() && d[p]) type += (char)d[p++];
        p++;
        if (p + 4 > d.size()) return false;
        const size_t size = u32(p); p += 4;
        if (p + size > d.size()) return false;
        if (name == "channels" && type == "chlist") {
            size_t q = p;
            while (q < p + size && d[q]) {
                Chan c;
                while (q < p + size && d[q]) c.name += (char)d[q++];
                q++;
                if (q + 16 > p + size) return false;
                c.type = (int)u32(q);
                if (u32(q + 8) != 1u || u32(q + 12) != 1u) return false;      // sub-sampled channels
                q += 16;
                chans.push_back(c);
            }
        } else if (name == "compression" && size == 1) compression = d[p];
        else if (name == "dataWindow" && size == 16) { xmin = (int)u32(p); ymin = (int)u32(p + 4); xmax = (int)u32(p + 8); ymax = (int)u32(p + 12); }
        p += size;
    }
    w = xmax - xmin + 1; h = ymax - ymin + 1;
    if (w <= 0 || h <= 0 || chans.empty() || compression < 0 || compression > 4) return false;
    size_t lineBytes = 0;
    std::vector<size_t> chanOff(chans.size());
    for (size_t c = 0; c < chans.size(); c++) {
        if (chans[c].type != 1 && chans[c].type != 2) return false;           // UINT channels
        chanOff[c] = lineBytes; lineBytes += (size_t)w * (chans[c].type == 1 ? 2 : 4);
    }
    int iR = -1, iG = -1, iB = -1, iA = -1;
    for (size_t c = 0; c < chans.size() && c < 4; c++) {
        if (chans[c].name == "R") iR = (int)c; else if (chans[c].name == "G") iG = (int)c;
        else if (chans[c].name == "B") iB = (int)c; else if (chans[c].name == "A") iA = (int)c;
    }
    const bool single = chans.size() == 1;
    if (!single && (iR < 0 || iG < 0 || iB < 0)) return false;
    const int linesPerBlock = compression == 3 ? 16 : (compression == 4 ? 32 : 1);
    const size_t blocks = ((size_t)h + linesPerBlock - 1) / linesPerBlock;
    if (p + 8 * blocks > d.size()) return false;
    rgba.assign(4 * (size_t)w * h, 0.0f);
    std::vector<uint8_t> tmp, pix;
    for (size_t b = 0; b < blocks; b++) {
        const uint64_t off = (uint64_t)u32(p + 8 * b) | ((uint64_t)u32(p + 8 * b + 4) << 32);
        if (off + 8 > d.size()) return false;
        const int y0 = (int)u32((size_t)off) - ymin;
        const size_t dataSize = u32((size_t)off + 4);
        if (y0 < 0 || y0 >= h || off + 8 + dataSize > d.size()) return false;
        const int lines = std::min(linesPerBlock, h - y0);
        const size_t want = lineBytes * lines;
        const uint8_t* src = &d[(size_t)off + 8];
        if (compression == 0 || dataSize == want) pix.assign(src, src + std::min(dataSize, want));   // stored as is when compression did not help
        else if (compression == 4) {
            std::vector<int> words;
            for (const auto& c : chans) words.push_back(c.type == 1 ? 1 : 2);
            if (!decodePizBlock(src, dataSize, w, lines, words, pix)) return false;
        } else {
            if (compression == 1) {                                           // RLE: count < 0: -count literals, else count + 1 copies of the next byte
                tmp.clear();
                for (size_t q = 0; q < dataSize;) {
                    const int c = (int8_t)src[q++];
                    if (c < 0) { if (q + (size_t)(-c) > dataSize) return false; tmp.insert(tmp.end(), src + q, src + q + (-c)); q += (size_t)(-c); }
                    else { if (q >= dataSize) return false; tmp.insert(tmp.end(), (size_t)c + 1, src[q++]); }
                }
            } else if (!Inflater(src, dataSize).run(tmp, want)) return false;
            if (tmp.size() != want) return false;
            for (size_t i = 1; i < want; i++) tmp[i] = (uint8_t)(tmp[i - 1] + tmp[i] - 128);       // predictor
            pix.resize(want);
            const size_t half = (want + 1) / 2;                               // de-interleave: first half = even bytes, second half = odd bytes
            for (size_t i = 0; i < want; i++) pix[i] = (i & 1) ? tmp[half + i / 2] : tmp[i / 2];
        }
        if (pix.size() != want) return false;
        for (int l = 0; l < lines; l++) {
            const uint8_t* line = &pix[lineBytes * l];
            float* out = &rgba[4 * (size_t)w * (y0 + l)];
            auto value = [&](int c, int x) {
                const uint8_t* q = line + chanOff[c];
                if (chans[c].type == 1) { uint16_t v; std::memcpy(&v, q + 2 * (size_t)x, 2); return halfToFloat(v); }
                float f; std::memcpy(&f, q + 4 * (size_t)x, 4); return f;
            };
            for (int x = 0; x < w; x++) {
                if (single) { const float v = value(0, x); out[4 * x] = out[4 * x + 1] = out[4 * x + 2] = out[4 * x + 3] = v; continue; }
                out[4 * x] = value(iR, x); out[4 * x + 1] = value(iG, x); out[4 * x + 2] = value(iB, x);
                out[4 * x + 3] = iA >= 0 ? value(iA, x) : 1.0f;
            }
        }
    }
    return true;
}

// ---- JPEG: sequential (SOF0 / SOF1) and progressive (SOF2) DCT, 8-bit, Huffman, 1 or 3 components, any sampling factors, restart intervals —
// decoded to the bytes stb_image produces: its integer "islow" IDCT (12-bit constants, 2 extra bits kept between the passes), its
// chroma up-sampling filters ((3 near + far) / 4 per axis, nearest for other ratios) and its 20-bit fixed-point YCbCr -> RGB.
// comp = 1 (grey) or 3 (YCbCr, Adobe RGB, and CMYK / YCCK reduced to RGB as stb_image does).  Arithmetic-coded, lossless, 12-bit files are refused. ----
class JpegDecoder {
public:
    bool run(const std::vector<uint8_t>& d, int& w, int& h, int& comp, std::vector<uint8_t>& px) {
        m_d = d.data(); m_n = d.size();
        if (m_n < 4 || m_d[0] != 0xff || m_d[1] != 0xd8) return false;
        size_t p = 2;
        bool haveFrame = false, done = false;
        while (!done) {
            while (p < m_n && m_d[p] != 0xff) p++;
            while (p < m_n && m_d[p] == 0xff) p++;
            if (p >= m_n) return false;
            const int m = m_d[p++];
            if (m == 0xd9) break;
            if (m == 0x00 || m == 0x01 || (m >= 0xd0 && m <= 0xd7)) continue;     // stuffed byte of left-over entropy data, TEM, stray RSTn
            if (p + 2 > m_n) return false;
            const size_t len = ((size_t)m_d[p] << 8) | m_d[p + 1];
            if (len < 2 || p + len > m_n) return false;
            const uint8_t* b = m_d + p + 2; const size_t bl = len - 2;
            if (m == 0xdb) {                                                   // DQT
                for (size_t q = 0; q < bl;) {
                    const int pq = b[q] >> 4, tq = b[q] & 15; q++;
                    if (pq > 1 || tq > 3 || q + (pq ? 128u : 64u) > bl) return false;
                    for (int i = 0; i < 64; i++) { m_q[tq][kZig[i]] = pq ? (uint16_t)((b[q] << 8) | b[q + 1]) : b[q]; q += pq ? 2 : 1; }
                }
            } else if (m == 0xc4) {                                            // DHT
                for (size_t q = 0; q < bl;) {
                    if (q + 17 > bl) return false;
                    const int tc = b[q] >> 4, th = b[q] & 15; q++;
                    if (tc > 1 || th > 3) return false;
                    Huff& t = m_h[tc][th];
                    int total = 0;
                    for (int i = 0; i < 16; i++) { t.count[i + 1] = b[q + i]; total += b[q + i]; }
                    q += 16;
                    if (total > 256 || q + total > bl) return false;
                    for (int i = 0; i < total; i++) t.sym[i] = b[q + i];
                    q += total;
                    int code = 0, k = 0;                                       // canonical codes: first code / first index per length
                    for (int l = 1; l <= 16; l++) { t.first[l] = code; t.index[l] = k; code = (code + t.count[l]) << 1; k += t.count[l]; }
                }
            } else if (m == 0xc0 || m == 0xc1 || m == 0xc2) {                  // SOF0 / SOF1 (Huffman, sequential), SOF2 (progressive)
                if (bl < 6 || b[0] != 8) return false;
                m_prog = m == 0xc2;
                m_h_ = (b[1] << 8) | b[2]; m_w_ = (b[3] << 8) | b[4]; m_nc = b[5];
                if (m_w_ <= 0 || m_h_ <= 0 || (m_nc != 1 && m_nc != 3 && m_nc != 4) || bl < 6 + 3 * (size_t)m_nc) return false;
                m_hmax = m_vmax = 1;
                for (int c = 0; c < m_nc; c++) {
                    Comp& C = m_c[c];
                    C.id = b[6 + 3 * c]; C.hs = b[7 + 3 * c] >> 4; C.vs = b[7 + 3 * c] & 15; C.tq = b[8 + 3 * c];
                    if (C.hs < 1 || C.hs > 4 || C.vs < 1 || C.vs > 4 || C.tq > 3) return false;
                    m_hmax = std::max(m_hmax, C.hs); m_vmax = std::max(m_vmax, C.vs);
                }
                for (int c = 0; c < m_nc; c++) if (m_hmax % m_c[c].hs || m_vmax % m_c[c].vs) return false;
                m_rgbIds = m_nc == 3 && m_c[0].id == 'R' && m_c[1].id == 'G' && m_c[2].id == 'B';
                m_mcux = (m_w_ + 8 * m_hmax - 1) / (8 * m_hmax); m_mcuy = (m_h_ + 8 * m_vmax - 1) / (8 * m_vmax);
                for (int c = 0; c < m_nc; c++) {
                    Comp& C = m_c[c];
                    C.x = (m_w_ * C.hs + m_hmax - 1) / m_hmax; C.y = (m_h_ * C.vs + m_vmax - 1) / m_vmax;
                    C.w2 = m_mcux * C.hs * 8; C.h2 = m_mcuy * C.vs * 8;
                    C.data.assign((size_t)C.w2 * C.h2, 0);
                    if (m_prog) { C.cw = C.w2 / 8; C.coef.assign((size_t)C.cw * (C.h2 / 8) * 64, 0); }
                }
                haveFrame = true;
            } else if (m >= 0xc3 && m <= 0xcf && m != 0xc4 && m != 0xc8 && m != 0xcc) return false;   // lossless, differential, arithmetic
            else if (m == 0xdd) { if (bl < 2) return false; m_restart = (b[0] << 8) | b[1]; }
            else if (m == 0xe0) { if (bl >= 5 && !std::memcmp(b, "JFIF", 5)) m_jfif = true; }
            else if (m == 0xee) { if (bl >= 12 && !std::memcmp(b, "Adobe", 6)) m_adobe = b[11]; }
            else if (m == 0xda) {                                              // SOS + entropy-coded data
                if (!haveFrame || bl < 1) return false;
                const int ns = b[0];
                if (ns < 1 || ns > m_nc || bl < 1 + 2 * (size_t)ns + 3) return false;
                int order[4];
                for (int i = 0; i < ns; i++) {
                    int c = 0;
                    while (c < m_nc && m_c[c].id != b[1 + 2 * i]) c++;
                    if (c == m_nc) return false;
                    m_c[c].td = b[2 + 2 * i] >> 4; m_c[c].ta = b[2 + 2 * i] & 15;
                    if (m_c[c].td > 3 || m_c[c].ta > 3) return false;
                    order[i] = c;
                }
                m_ss = b[1 + 2 * ns]; m_se = b[2 + 2 * ns]; m_ah = b[3 + 2 * ns] >> 4; m_al = b[3 + 2 * ns] & 15;
                if (m_prog) { if (m_ss > 63 || m_se > 63 || m_ss > m_se || m_ah > 13 || m_al > 13) return false; }
                else if (m_ss != 0 || m_ah != 0 || m_al != 0) return false;
                p += len;
                if (!scan(p, order, ns)) return false;
                continue;
            }
            p += len;
        }
        if (!haveFrame) return false;
        return finish(w, h, comp, px);
    }
private:
    struct Huff { int count[17] = {0}, first[17] = {0}, index[17] = {0}; uint8_t sym[256] = {0}; };
    struct Comp { int id = 0, hs = 1, vs = 1, tq = 0, td = 0, ta = 0, x = 0, y = 0, w2 = 0, h2 = 0, pred = 0, cw = 0; std::vector<uint8_t> data; std::vector<short> coef; };
    static constexpr uint8_t kZig[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
                                         35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
    // entropy-coded segment reader: 0xff 0x00 is a stuffed 0xff, any other marker ends the data (zeros are fed from there on)
    int bit() {
        if (m_cnt == 0) {
            int v = 0;
            if (!m_hitMarker && m_p < m_n) {
                v = m_d[m_p];
                if (v == 0xff) {
                    if (m_p + 1 < m_n && m_d[m_p + 1] == 0) m_p += 2;
                    else { m_hitMarker = true; v = 0; }
                } else m_p++;
            }
            m_acc = v; m_cnt = 8;
        }
        return (m_acc >> --m_cnt) & 1;
    }
    int receiveExtend(int n) {                                                 // n-bit magnitude, JPEG sign convention
        if (n == 0) return 0;
        int v = 0;
        for (int i = 0; i < n; i++) v = (v << 1) | bit();
        return v < (1 << (n - 1)) ? v - (1 << n) + 1 : v;
    }
    int symbol(const Huff& t) {
        int code = 0;
        for (int l = 1; l <= 16; l++) {
            code = (code << 1) | bit();
            if (code - t.first[l] < t.count[l]) return t.sym[t.index[l] + code - t.first[l]];
        }
        m_bad = true; return 0;
    }
    // one 1-D pass of the 8-point inverse DCT (Loeffler-Ligtenberg-Moschytz as in IJG's jidctint), inputs s[0..7], 12 fractional bits:
    // e[] = even part sums, o[] = odd part; outputs are e[k] +- o[3 - k]
    static void idct1d(const int* s, int* e, int* o) {
        auto f = [](double x) { return (int)(x * 4096 + 0.5); };
        static const int c0 = f(0.5411961f), c1 = f(-1.847759065f), c2 = f(0.765366865f), c3 = f(1.175875602f), c4 = f(0.298631336f), c5 = f(2.053119869f),
                         c6 = f(3.072711026f), c7 = f(1.501321110f), c8 = f(-0.899976223f), c9 = f(-2.562915447f), c10 = f(-1.961570560f), c11 = f(-0.390180644f);
        const int z = (s[2] + s[6]) * c0, a2 = z + s[6] * c1, a3 = z + s[2] * c2;
        const int a0 = (s[0] + s[4]) * 4096, a1 = (s[0] - s[4]) * 4096;
        e[0] = a0 + a3; e[3] = a0 - a3; e[1] = a1 + a2; e[2] = a1 - a2;
        int t0 = s[7], t1 = s[5], t2 = s[3], t3 = s[1];
        int p3 = t0 + t2, p4 = t1 + t3, p1 = t0 + t3, p2 = t1 + t2;
        const int p5 = (p3 + p4) * c3;
        t0 *= c4; t1 *= c5; t2 *= c6; t3 *= c7;
        p1 = p5 + p1 * c8; p2 = p5 + p2 * c9; p3 *= c10; p4 *= c11;
        o[3] = t3 + p1 + p4; o[2] = t2 + p2 + p3; o[1] = t1 + p2 + p4; o[0] = t0 + p1 + p3;
    }
    static uint8_t clamp255(int x) { return (uint8_t)(x < 0 ? 0 : (x > 255 ? 255 : x)); }
    static void idctBlock(uint8_t* out, int stride, const short* d) {
        int v[64];
        for (int i = 0; i < 8; i++) {                                           // columns; 2 extra bits kept
            if (!d[8 + i] && !d[16 + i] && !d[24 + i] && !d[32 + i] && !d[40 + i] && !d[48 + i] && !d[56 + i]) {
                for (int k = 0; k < 8; k++) v[8 * k + i] = d[i] * 4;
                continue;
            }
            const int s[8] = {d[i], d[8 + i], d[16 + i], d[24 + i], d[32 + i], d[40 + i], d[48 + i], d[56 + i]};
            int e[4], o[4];
            idct1d(s, e, o);
            for (int k = 0; k < 4; k++) { v[8 * k + i] = (e[k] + 512 + o[3 - k]) >> 10; v[8 * (7 - k) + i] = (e[k] + 512 - o[3 - k]) >> 10; }
        }
        for (int r = 0; r < 8; r++) {                                           // rows; +128 level shift and rounding folded into the bias
            int e[4], o[4];
            idct1d(v + 8 * r, e, o);
            const int bias = 65536 + (128 << 17);
            for (int k = 0; k < 4; k++) {
                out[r * stride + k] = clamp255((e[k] + bias + o[3 - k]) >> 17);
                out[r * stride + 7 - k] = clamp255((e[k] + bias - o[3 - k]) >> 17);
            }
        }
    }
    bool block(Comp& C, int bx, int by) {
        short d[64] = {0};
        const Huff& hd = m_h[0][C.td]; const Huff& ha = m_h[1][C.ta];
        const uint16_t* q = m_q[C.tq];
        const int t = symbol(hd);
        if (m_bad || t > 15) return false;
        C.pred += receiveExtend(t);
        d[0] = (short)(C.pred * q[0]);
        for (int k = 1; k < 64;) {
            const int rs = symbol(ha);
            if (m_bad) return false;
            const int r = rs >> 4, sz = rs & 15;
            if (sz == 0) { if (rs != 0xf0) break; k += 16; continue; }
            k += r;
            if (k > 63) return false;
            const int z = kZig[k++];
            d[z] = (short)(receiveExtend(sz) * q[z]);
        }
        idctBlock(&C.data[(size_t)by * 8 * C.w2 + (size_t)bx * 8], C.w2, d);
        return true;
    }
    // Progressive scans (ITU T.81 annex G): DC first / refinement, AC first (with end-of-band runs) / refinement, on the stored coefficients.
    bool blockProg(Comp& C, int bx, int by) {
        short* d = &C.coef[64 * ((size_t)bx + (size_t)by * C.cw)];
        if (m_ss == 0) {                                                        // DC scan
            if (m_se != 0) return false;
            if (m_ah == 0) {
                std::memset(d, 0, 64 * sizeof(short));
                const int t = symbol(m_h[0][C.td]);
                if (m_bad || t > 15) return false;
                C.pred += receiveExtend(t);
                d[0] = (short)(C.pred * (1 << m_al));
            } else if (bit()) d[0] = (short)(d[0] + (short)(1 << m_al));
            return true;
        }
        const Huff& ha = m_h[1][C.ta];
        if (m_ah == 0) {                                                        // AC first pass
            if (m_eobRun) { m_eobRun--; return true; }
            int k = m_ss;
            do {
                const int rs = symbol(ha);
                if (m_bad) return false;
                const int r = rs >> 4, sz = rs & 15;
                if (sz == 0) {
                    if (r < 15) {
                        m_eobRun = 1 << r;
                        for (int i = 0, v = 0; i < r; i++) { v = (v << 1) | bit(); if (i == r - 1) m_eobRun += v; }
                        m_eobRun--;
                        break;
                    }
                    k += 16;
                } else {
                    k += r;
                    if (k > 63) return false;
                    d[kZig[k++]] = (short)(receiveExtend(sz) * (1 << m_al));
                }
            } while (k <= m_se);
            return true;
        }
        const short one = (short)(1 << m_al);                                   // AC refinement
        auto refine = [&](short* c) { if (bit() && (*c & one) == 0) *c = (short)(*c > 0 ? *c + one : *c - one); };
        if (m_eobRun) {
            m_eobRun--;
            for (int k = m_ss; k <= m_se; k++) { short* c = &d[kZig[k]]; if (*c != 0) refine(c); }
            return true;
        }
        int k = m_ss;
        do {
            const int rs = symbol(ha);
            if (m_bad) return false;
            int r = rs >> 4, sz = rs & 15;
            if (sz == 0) {
                if (r < 15) {
                    m_eobRun = (1 << r) - 1;
                    for (int i = 0, v = 0; i < r; i++) { v = (v << 1) | bit(); if (i == r - 1) m_eobRun += v; }
                    r = 64;                                                     // the rest of the band only gets correction bits
                }
            } else {
                if (sz != 1) return false;
                sz = bit() ? one : -one;
            }
            while (k <= m_se) {
                short* c = &d[kZig[k++]];
                if (*c != 0) refine(c);
                else { if (r == 0) { *c = (short)sz; break; } r--; }
            }
        } while (k <= m_se);
        return true;
    }
    void restartInterval() {                                                    // skip to just behind the RSTn marker, reset predictors
        m_cnt = 0; m_hitMarker = false;
        while (m_p + 1 < m_n && !(m_d[m_p] == 0xff && m_d[m_p + 1] >= 0xd0 && m_d[m_p + 1] <= 0xd7)) {
            if (m_d[m_p] == 0xff && m_d[m_p + 1] != 0 && m_d[m_p + 1] != 0xff) return;   // some other marker: leave it to the outer loop
            m_p++;
        }
        if (m_p + 1 < m_n) m_p += 2;
        for (auto& c : m_c) c.pred = 0;
        m_eobRun = 0;
    }
    bool scan(size_t& p, const int* order, int ns) {
        m_p = p; m_cnt = 0; m_hitMarker = false; m_eobRun = 0;
        for (auto& c : m_c) c.pred = 0;
        int todo = m_restart ? m_restart : 0x7fffffff;
        if (m_prog && ns > 1 && m_ss != 0) return false;                         // AC bands are never interleaved
        if (ns == 1) {                                                          // non-interleaved: the component's own block grid
            Comp& C = m_c[order[0]];
            const int bw = (C.x + 7) >> 3, bh = (C.y + 7) >> 3;
            for (int by = 0; by < bh; by++) for (int bx = 0; bx < bw; bx++) {
                if (!(m_prog ? blockProg(C, bx, by) : block(C, bx, by))) return false;
                if (--todo <= 0) { restartInterval(); todo = m_restart; }
            }
        } else {
            for (int my = 0; my < m_mcuy; my++) for (int mx = 0; mx < m_mcux; mx++) {
                for (int i = 0; i < ns; i++) {
                    Comp& C = m_c[order[i]];
                    for (int y = 0; y < C.vs; y++) for (int x = 0; x < C.hs; x++)
                        if (!(m_prog ? blockProg(C, mx * C.hs + x, my * C.vs + y) : block(C, mx * C.hs + x, my * C.vs + y))) return false;
                }
                if (--todo <= 0) { restartInterval(); todo = m_restart; }
            }
        }
        p = m_p;
        return true;
    }
    bool finish(int& w, int& h, int& comp, std::vector<uint8_t>& px) {
        if (m_prog)                                                              // de-quantise and transform the accumulated coefficients
            for (int c = 0; c < m_nc; c++) {
                Comp& C = m_c[c];
                const int bw = (C.x + 7) >> 3, bh = (C.y + 7) >> 3;
                for (int by = 0; by < bh; by++) for (int bx = 0; bx < bw; bx++) {
                    short* d = &C.coef[64 * ((size_t)bx + (size_t)by * C.cw)];
                    for (int k = 0; k < 64; k++) d[k] = (short)(d[k] * m_q[C.tq][k]);
                    idctBlock(&C.data[(size_t)by * 8 * C.w2 + (size_t)bx * 8], C.w2, d);
                }
            }
        w = m_w_; h = m_h_; comp = m_nc >= 3 ? 3 : 1;                           // stb_image reports "3" for CMYK / YCCK files too
        px.assign((size_t)w * h * comp, 0);
        std::vector<std::vector<uint8_t>> line(m_nc, std::vector<uint8_t>((size_t)w + 8 * 4 + 3));
        struct Rs { int hs, vs, ystep, wl, ypos; const uint8_t *l0, *l1; } rs[4];
        const bool isRgb = m_nc == 3 && (m_rgbIds || (m_adobe == 0 && !m_jfif));  // components stored as R, G, B: no colour conversion
        auto blinn = [](int x, int y) { const unsigned t = (unsigned)(x * y + 128); return (uint8_t)((t + (t >> 8)) >> 8); };
        for (int c = 0; c < m_nc; c++) {
            rs[c].hs = m_hmax / m_c[c].hs; rs[c].vs = m_vmax / m_c[c].vs; rs[c].ystep = rs[c].vs >> 1;
            rs[c].wl = (w + rs[c].hs - 1) / rs[c].hs; rs[c].ypos = 0; rs[c].l0 = rs[c].l1 = m_c[c].data.data();
        }
        for (int j = 0; j < h; j++) {
            const uint8_t* row[4] = {nullptr, nullptr, nullptr, nullptr};
            for (int c = 0; c < m_nc; c++) {
                Rs& r = rs[c];
                const bool bot = r.ystep >= (r.vs >> 1);
                const uint8_t* nr = bot ? r.l1 : r.l0; const uint8_t* fr = bot ? r.l0 : r.l1;
                uint8_t* o = line[c].data();
                const int n = r.wl;
                if (r.hs == 1 && r.vs == 1) row[c] = nr;
                else {
                    if (r.hs == 1 && r.vs == 2) for (int i = 0; i < n; i++) o[i] = (uint8_t)((3 * nr[i] + fr[i] + 2) >> 2);
                    else if (r.hs == 2 && r.vs == 1) {
                        if (n == 1) o[0] = o[1] = nr[0];
                        else {
                            o[0] = nr[0]; o[1] = (uint8_t)((nr[0] * 3 + nr[1] + 2) >> 2);
                            for (int i = 1; i < n - 1; i++) { const int t = 3 * nr[i] + 2; o[2 * i] = (uint8_t)((t + nr[i - 1]) >> 2); o[2 * i + 1] = (uint8_t)((t + nr[i + 1]) >> 2); }
                            o[2 * (n - 1)] = (uint8_t)((nr[n - 2] * 3 + nr[n - 1] + 2) >> 2); o[2 * (n - 1) + 1] = nr[n - 1];
                        }
                    } else if (r.hs == 2 && r.vs == 2) {
                        if (n == 1) o[0] = o[1] = (uint8_t)((3 * nr[0] + fr[0] + 2) >> 2);
                        else {
                            int t1 = 3 * nr[0] + fr[0];
                            o[0] = (uint8_t)((t1 + 2) >> 2);
                            for (int i = 1; i < n; i++) { const int t0 = t1; t1 = 3 * nr[i] + fr[i]; o[2 * i - 1] = (uint8_t)((3 * t0 + t1 + 8) >> 4); o[2 * i] = (uint8_t)((3 * t1 + t0 + 8) >> 4); }
                            o[2 * n - 1] = (uint8_t)((t1 + 2) >> 2);
                        }
                    } else for (int i = 0; i < n; i++) for (int k = 0; k < r.hs; k++) o[i * r.hs + k] = nr[i];
                    row[c] = o;
                }
                if (++r.ystep >= r.vs) { r.ystep = 0; r.l0 = r.l1; if (++r.ypos < m_c[c].y) r.l1 += m_c[c].w2; }
            }
            uint8_t* out = &px[(size_t)j * w * comp];
            if (m_nc == 1) std::memcpy(out, row[0], (size_t)w);
            else if (isRgb) for (int i = 0; i < w; i++) { out[3 * i] = row[0][i]; out[3 * i + 1] = row[1][i]; out[3 * i + 2] = row[2][i]; }
            else if (m_nc == 4 && m_adobe == 0) for (int i = 0; i < w; i++) {   // CMYK (Adobe, stored inverted): colour x black
                const int k4 = row[3][i];
                out[3 * i] = blinn(row[0][i], k4); out[3 * i + 1] = blinn(row[1][i], k4); out[3 * i + 2] = blinn(row[2][i], k4);
            }
            else for (int i = 0; i < w; i++) {                                  // 20-bit fixed point, the green Cb term truncated to 16 bits first
                auto fx = [](float x) { return ((int)(x * 4096.0f + 0.5f)) << 8; };
                const int yf = (row[0][i] << 20) + (1 << 19), cr = row[2][i] - 128, cb = row[1][i] - 128;
                const int r = (yf + cr * fx(1.40200f)) >> 20;
                const int g = (int)(yf + (cr * -fx(0.71414f)) + (int)(((unsigned)(cb * -fx(0.34414f))) & 0xffff0000u)) >> 20;
                const int b = (yf + cb * fx(1.77200f)) >> 20;
                out[3 * i] = clamp255(r); out[3 * i + 1] = clamp255(g); out[3 * i + 2] = clamp255(b);
                if (m_nc == 4 && m_adobe == 2) {                                // YCCK: the converted colour is inverted CMY, times black
                    const int k4 = row[3][i];
                    out[3 * i] = blinn(255 - out[3 * i], k4); out[3 * i + 1] = blinn(255 - out[3 * i + 1], k4); out[3 * i + 2] = blinn(255 - out[3 * i + 2], k4);
                }
            }
        }
        return true;
    }
    const uint8_t* m_d = nullptr; size_t m_n = 0, m_p = 0;
    int m_acc = 0, m_cnt = 0; bool m_hitMarker = false, m_bad = false, m_jfif = false, m_rgbIds = false; int m_adobe = -1;
    uint16_t m_q[4][64] = {{0}};
    Huff m_h[2][4];
    Comp m_c[4];
    int m_w_ = 0, m_h_ = 0, m_nc = 0, m_hmax = 1, m_vmax = 1, m_mcux = 0, m_mcuy = 0, m_restart = 0;
    bool m_prog = false; int m_ss = 0, m_se = 63, m_ah = 0, m_al = 0, m_eobRun = 0;
};
inline bool decodeJpeg(const std::vector<uint8_t>& d, int& w, int& h, int& comp, std::vector<uint8_t>& px) { return JpegDecoder().run(d, w, h, comp, px); }

inline bool hasSuffix(const std::string& s, const char* suf) {
    const size_t n = std::strlen(suf);
    if (s.size() < n) return false;
    for (size_t i = 0; i < n; i++) if (std::tolower((unsigned char)s[s.size() - n + i]) != suf[i]) return false;
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
// One real number, read the way tinyobj reads it (tiny_obj_loader.h:866-1000, parseReal) — NOT strtod: the token runs to the next
// blank; digits before the point accumulate as m = 10 m + d, digits behind it as m += d * 10^-k (a table for k < 8, pow() beyond), an
// exponent e is applied as ldexp(m * 5^e, e); every step rounds in double, so the result can sit a few double ulps away from the
// correctly rounded value and, on a float tie, land on the other float ("-66.5789875e5" does).  Anything unparsable reads 0.
inline float parseFloat(const char*& p) {
    p = skipWs(p);
    const char* c = p;
    const char* end = p;
    while (*end && *end != ' ' && *end != '\t' && *end != '\r') end++;
    p = end;
    if (c >= end) return 0.0f;
    double m = 0.0; int exponent = 0, read = 0;
    char sign = '+', esign = '+';
    bool leadingDot = false;
    auto digit = [](char ch) { return ch >= '0' && ch <= '9'; };
    if (*c == '+' || *c == '-') { sign = *c++; if (c != end && *c == '.') leadingDot = true; }
    else if (digit(*c)) {}
    else if (*c == '.') leadingDot = true;
    else return 0.0f;
    if (!leadingDot) {
        while (c != end && digit(*c)) { m *= 10; m += (int)(*c - '0'); c++; read++; }
        if (read == 0) return 0.0f;
    }
    bool haveValue = true;
    if (c != end) {
        if (*c == '.') {
            static const double lut[] = {1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001};
            c++; read = 1;
            while (c != end && digit(*c)) { m += (int)(*c - '0') * (read < 8 ? lut[read] : std::pow(10.0, -read)); read++; c++; }
        } else if (*c != 'e' && *c != 'E') c = end;                 // trailing junk: the value so far stands
        if (c != end && (*c == 'e' || *c == 'E')) {
            c++;
            if (c != end && (*c == '+' || *c == '-')) esign = *c++;
            else if (c != end && digit(*c)) {}
            else haveValue = false;
            if (haveValue) {
                read = 0;
                while (c != end && digit(*c)) { if (exponent > 2147483647 / 10) { haveValue = false; break; } exponent *= 10; exponent += (int)(*c - '0'); c++; read++; }
                exponent *= (esign == '+' ? 1 : -1);
                if (read == 0) haveValue = false;
            }
        }
    }
    if (!haveValue) return 0.0f;
    const double v = (sign == '+' ? 1 : -1) * (exponent ? std::ldexp(m * std::pow(5.0, exponent), exponent) : m);
    return (float)v;
}
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
        else {
            // Polygons with more than 4 corners: ear clipping as tinyobj does it (tiny_obj_loader.h:1536-1820), all in fp32.
            // The polygon is projected on two axes picked at the first corner whose edge cross product is not ~0 (the plane
            // facing the cross product's largest component, ties resolved as there); then, walking a cursor round the remaining
            // corners, a corner (a, b, c) is an ear when it turns the same way as sign(ax*by - ay*bx) and no other remaining corner
            // lies inside triangle abc (crossing-number test); an ear emits (a, b, c) and drops b.  The walk gives up after a full
            // round without progress, dropping what is left unless exactly a triangle remains.
            for (const auto& c : f) if ((uint32_t)c.v >= nv) throw std::runtime_error("loadObj: face with invalid vertex index");
            const size_t n0 = f.size();
            int ax0 = 1, ax1 = 2;
            for (size_t k = 0; k < n0; k++) {
                const float* a = &o.positions[3 * (size_t)f[k % n0].v]; const float* b = &o.positions[3 * (size_t)f[(k + 1) % n0].v];
                const float* c = &o.positions[3 * (size_t)f[(k + 2) % n0].v];
                const float e0x = b[0] - a[0], e0y = b[1] - a[1], e0z = b[2] - a[2], e1x = c[0] - b[0], e1y = c[1] - b[1], e1z = c[2] - b[2];
                const float cx = std::fabs(e0y * e1z - e0z * e1y), cy = std::fabs(e0z * e1x - e0x * e1z), cz = std::fabs(e0x * e1y - e0y * e1x);
                const float eps = 1.1920928955078125e-07f;
                if (cx > eps || cy > eps || cz > eps) {
                    if (!(cx > cy && cx > cz)) { ax0 = 0; if (cz > cx && cz > cy) ax1 = 1; }
                    break;
                }
            }
            std::vector<Corner> rem(f.begin(), f.end());
            size_t guess = 0, left = n0, prevCount = n0;
            while (rem.size() > 3 && left > 0) {
                const size_t np = rem.size();
                if (guess >= np) guess -= np;
                if (prevCount != np) { prevCount = np; left = np; } else left--;
                Corner tri[3]; float vx[3], vy[3];
                for (int k = 0; k < 3; k++) {
                    tri[k] = rem[(guess + k) % np];
                    vx[k] = o.positions[3 * (size_t)tri[k].v + ax0]; vy[k] = o.positions[3 * (size_t)tri[k].v + ax1];
                }
                const float e0x = vx[1] - vx[0], e0y = vy[1] - vy[0], e1x = vx[2] - vx[1], e1y = vy[2] - vy[1];
                const float cross = e0x * e1y - e0y * e1x;
                const float area = (vx[0] * vy[1] - vy[0] * vx[1]) * 0.5f;
                if (cross * area < 0.0f) { guess++; continue; }
                bool overlap = false;
                for (size_t other = 3; other < np && !overlap; other++) {
                    const Corner& q = rem[(guess + other) % np];
                    const float tx = o.positions[3 * (size_t)q.v + ax0], ty = o.positions[3 * (size_t)q.v + ax1];
                    bool in = false;                                  // crossing number against the 3 edges
                    for (int i = 0, j = 2; i < 3; j = i++)
                        if (((vy[i] > ty) != (vy[j] > ty)) && (tx < (vx[j] - vx[i]) * (ty - vy[i]) / (vy[j] - vy[i]) + vx[i])) in = !in;
                    overlap = in;
                }
                if (overlap) { guess++; continue; }
                emit(tri[0]); emit(tri[1]); emit(tri[2]); o.primMaterial.push_back(faceMat[fi]);
                rem.erase(rem.begin() + (std::ptrdiff_t)((guess + 1) % np));
            }
            if (rem.size() == 3) { emit(rem[0]); emit(rem[1]); emit(rem[2]); o.primMaterial.push_back(faceMat[fi]); }
        }
    }
}

}  // namespace loader
}  // namespace prt
