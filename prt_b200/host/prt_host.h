// prt_host.h — C++ host API with the reference's class names and method meanings (Scene / Mesh / Bvh / Camera /
// PathTracer / Image), implemented as thin shims over the C ABI in include/prt_cuda.h.
//
// A program written against amada/PRT's classes (src/main.cpp:22-55, 107-190) keeps its shape:
//     Mesh m; m.create(...); fill buffers; Bvh* b = new Bvh; b->build(std::move(m)); scene.add(b);
//     camera.create(pos, dir, w, h); PathTracer t; t.TraceBlock(image, x0, y0, x1, y1, scene, camera, spp);
// What changes: one ray per call would waste a GPU, so Scene::intersect / Scene::occluded also exist in batch form;
// TraceBlock renders the whole block as a wavefront on the device.  Errors: the reference traps (PRT_ASSERT ->
// __builtin_trap, prt.h:7-13); here every failing C-ABI call throws std::runtime_error with prt_cuda_last_error().
//
// Out of scope here, as in SURVEY §2: OBJ/MTL/PNG/EXR loaders (host file I/O — load with any loader and fill the Mesh
// buffers), thread pool, stats.  SURVEY §8f-2/§8f-3 are in: GbufferVisualizer, Image::savePpm / Image::saveExr (the
// pixel conversions run on the device; the files are written here).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../../include/prt_cuda.h"
#include "prt_loader.h"

namespace prt {

struct Vector2f { float x = 0, y = 0; Vector2f() = default; Vector2f(float a, float b) : x(a), y(b) {} };
struct Vector3f {
    float x = 0, y = 0, z = 0;
    Vector3f() = default;
    Vector3f(float f) : x(f), y(f), z(f) {}
    Vector3f(float a, float b, float c) : x(a), y(b), z(c) {}
};

enum class ReflectionType : uint32_t { kDiffuse = 0, kSpecular, kRefraction };   // material.h:24-28
enum class LightType : uint32_t { kDirectional, kInfiniteArea };                 // light.h:9-12

// RGBA8 / R8 / RGBA32F texel block (texture.h:13-22); the Scene uploads it when the owning mesh is added.
struct Texture {
    uint16_t width = 0, height = 0;
    uint8_t component = 0;
    bool isFloat = false;
    std::vector<uint8_t> texels;
    bool isValid() const { return !texels.empty(); }
    void init() { *this = Texture(); }
    // Texture::load (texture.cpp:212-254): 1-component images stay 1 component, everything else becomes RGBA8; a
    // 3-component bump texture is a normal map and is converted to a height map.  TGA and 8-bit PNG (prt_loader.h).
    void load(const char* path, bool bumpTexture = false) {
        std::vector<uint8_t> file, px;
        int w = 0, h = 0, comp = 0, pc = 0;                           // comp: what stb_image reports; pc: channels stored in px
        const bool isPng = loader::hasSuffix(path, ".png"), isJpg = loader::hasSuffix(path, ".jpg") || loader::hasSuffix(path, ".jpeg");
        bool ok = loader::readFile(path, file);
        if (ok) {
            if (isPng) ok = loader::decodePng(file, w, h, comp, pc, px);
            else if (isJpg) { ok = loader::decodeJpeg(file, w, h, comp, px); pc = comp; }
            else { ok = loader::decodeTga(file, w, h, comp, px); pc = comp; }
        }
        if (!ok) throw std::runtime_error(std::string("Texture::load: cannot decode '") + path + "' (TGA 8/24/32-bit, PNG, baseline JPEG)");
        width = (uint16_t)w; height = (uint16_t)h; isFloat = false;
        const size_t n = (size_t)w * h;
        if (comp == 1) { component = 1; texels = px; return; }
        std::vector<uint8_t> rgba(4 * n);
        for (size_t i = 0; i < n; i++) {                              // stb_image's conversion to 4 components: grey(+alpha) replicates the luminance
            if (pc == 2) { rgba[4 * i] = rgba[4 * i + 1] = rgba[4 * i + 2] = px[2 * i]; rgba[4 * i + 3] = px[2 * i + 1]; continue; }
            rgba[4 * i] = px[pc * i]; rgba[4 * i + 1] = px[pc * i + 1]; rgba[4 * i + 2] = px[pc * i + 2];
            rgba[4 * i + 3] = pc == 4 ? px[4 * i + 3] : 255;
        }
        if (bumpTexture && comp == 3) {
            component = 1; texels.assign(n, 0);
            loader::convertNormalToBump(texels.data(), rgba.data(), (uint32_t)w, (uint32_t)h);
        } else { component = 4; texels.swap(rgba); }
    }
    // Texture::loadExr (texture.cpp:256-336): RGBA32F, 4 components (scan-line EXR, see prt_loader.h)
    void loadExr(const char* path) {
        std::vector<uint8_t> file; std::vector<float> rgba;
        int w = 0, h = 0;
        if (!loader::readFile(path, file) || !loader::decodeExr(file, w, h, rgba))
            throw std::runtime_error(std::string("Texture::loadExr: cannot decode '") + path + "' (single-part scan-line EXR, HALF / FLOAT, NONE / RLE / ZIPS / ZIP / PIZ only)");
        width = (uint16_t)w; height = (uint16_t)h; component = 4; isFloat = true;
        texels.resize(rgba.size() * sizeof(float));
        std::memcpy(texels.data(), rgba.data(), texels.size());
    }
};

struct Material {                                                                  // material.h:30-46
    Vector3f diffuse, ambient, specular, emissive;
    Texture diffuseMap, bumpMap;
    ReflectionType reflectionType = ReflectionType::kDiffuse;
    bool alphaTest = false;                                                        // filled in by Scene::add (material.cpp:77)
    void init() { *this = Material(); }
};

struct Ray { Vector3f org, dir; float maxT = 0; };                                 // ray.h:14-41 (prepare() happens on the device)
struct RayHit {                                                                    // ray.h:182-198
    static constexpr float kMissT = -1.0f;
    float t = kMissT, i = 0, j = 0, k = 0;
    uint32_t primId = 0, meshId = 0;
    bool isHit() const { return t != kMissT; }
};
static_assert(sizeof(RayHit) == sizeof(prt_hit), "RayHit must match prt_hit");

struct SurfaceProperties {                                                         // mesh.h:19-28 (SurfacePropertiesT<Vector3f, const Material*>)
    Vector3f normal;
    const Material* material = nullptr;
    Vector2f uv, duv01, duv02;
    Vector3f dp01, dp02;
};

class Mesh {                                                                       // mesh.h:34-103
public:
    static const uint32_t kVertexCountPerPrim = 3;
    void create(uint32_t primCount, uint32_t vertexCount, uint32_t materialCount, bool hasVertexNormal) {   // mesh.cpp:87-105
        m_indices.assign(3 * (size_t)primCount, 0); m_positions.assign(vertexCount, Vector3f());
        m_texcoords.assign(vertexCount, Vector2f()); m_normals.assign(hasVertexNormal ? vertexCount : 0, Vector3f());
        m_primMaterial.assign(primCount, 0); m_materials.assign(materialCount, Material());
        m_hasVertexNormal = hasVertexNormal;
    }
    uint32_t getPrimCount() const { return (uint32_t)(m_indices.size() / 3); }
    uint32_t getVertexCount() const { return (uint32_t)m_positions.size(); }
    uint32_t getMaterialCount() const { return (uint32_t)m_materials.size(); }
    uint32_t* getIndexBuffer() { return m_indices.data(); }
    Vector3f* getPositionBuffer() { return m_positions.data(); }
    Vector3f* getNormalBuffer() { return m_normals.data(); }
    Vector2f* getTexcoordBuffer() { m_hasTexcoord = true; return m_texcoords.data(); }
    uint32_t* getPrimMateialBuffer() { return m_primMaterial.data(); }              // (sic) mesh.h:77
    Material* getMaterialBuffer() { return m_materials.data(); }
    bool hasVertexNormal() const { return m_hasVertexNormal; }

    // Mesh::calculateVertexNormals (mesh.cpp:108-149): faces visited last to first; an accumulation that would make
    // the running normal zero-length is skipped; NaN face normals count as zero.
    void calculateVertexNormals() {
        auto sub = [](Vector3f a, Vector3f b) { return Vector3f(a.x - b.x, a.y - b.y, a.z - b.z); };
        auto cross = [](Vector3f a, Vector3f b) { return Vector3f(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); };
        auto length = [](Vector3f a) { float x = a.x * a.x, y = a.y * a.y, z = a.z * a.z; return std::sqrt(x + y + z); };
        auto normalize = [&](Vector3f a) { float inv = 1.0f / length(a); return Vector3f(inv * a.x, inv * a.y, inv * a.z); };
        m_normals.assign(m_positions.size(), Vector3f(0.0f));
        for (int32_t i = (int32_t)getPrimCount() - 1; i >= 0; i--) {
            uint32_t v[3] = {m_indices[3 * i], m_indices[3 * i + 1], m_indices[3 * i + 2]};
            Vector3f n = normalize(cross(sub(m_positions[v[1]], m_positions[v[0]]), sub(m_positions[v[2]], m_positions[v[0]])));
            if (std::isnan(n.x) || std::isnan(n.y) || std::isnan(n.z)) n = Vector3f(0.0f);
            for (int j = 0; j < 3; j++) {
                Vector3f s(m_normals[v[j]].x + n.x, m_normals[v[j]].y + n.y, m_normals[v[j]].z + n.z);
                if (length(s) > 0.0f) m_normals[v[j]] = s;
            }
        }
        for (auto& n : m_normals) n = normalize(n);
        m_hasVertexNormal = true;
    }
    void calculateBounds() {}                                                       // done by the C ABI at Scene::add (mesh.cpp:300-309)

    // Mesh::loadObj(path) (mesh.cpp:212-298): OBJ + MTL + textures -> the buffers Scene::add uploads.  See prt_loader.h
    // for the reference behaviours that are kept (uv per position index, quad split rule, material id -1, ...).
    void loadObj(const char* path) {
        loader::ObjData o;
        loader::parseObj(path, o);
        const uint32_t primCount = (uint32_t)(o.indices.size() / 3), vertexCount = (uint32_t)(o.positions.size() / 3);
        create(primCount, vertexCount, (uint32_t)o.materials.size(), false);
        m_indices = o.indices;
        std::memcpy((void*)m_positions.data(), o.positions.data(), o.positions.size() * sizeof(float));
        std::memcpy((void*)m_texcoords.data(), o.texcoords.data(), o.texcoords.size() * sizeof(float));
        m_primMaterial = o.primMaterial;
        const std::string dir = loader::dirOf(path);
        for (size_t i = 0; i < o.materials.size(); i++) {                           // Material::load (material.cpp:45-85)
            const loader::MtlData& s = o.materials[i];
            Material& d = m_materials[i];
            d.diffuse = Vector3f(s.diffuse[0], s.diffuse[1], s.diffuse[2]); d.ambient = Vector3f(s.ambient[0], s.ambient[1], s.ambient[2]);
            d.specular = Vector3f(s.specular[0], s.specular[1], s.specular[2]); d.emissive = Vector3f(s.emission[0], s.emission[1], s.emission[2]);
            auto fix = [](std::string p) { for (auto& c : p) if (c == '\\') c = '/'; return p; };
            if (!s.diffuseTex.empty()) d.diffuseMap.load(fix(dir + "/" + s.diffuseTex).c_str());
            if (!s.bumpTex.empty()) d.bumpMap.load(fix(dir + "/" + s.bumpTex).c_str(), true);
            d.alphaTest = false;                                                    // Texture::isAlphaTestRequired (texture.cpp:338-350)
            if (d.diffuseMap.isValid() && d.diffuseMap.component == 4)
                for (size_t k = 3; k < d.diffuseMap.texels.size(); k += 4) if (d.diffuseMap.texels[k] < 255) { d.alphaTest = true; break; }
        }
        m_hasTexcoord = true;
    }
    // Mesh::loadObj(path, mat) (mesh.cpp:151-210): geometry only, one caller-supplied material for every primitive.
    void loadObj(const char* path, const Material& mat) {
        loader::ObjData o;
        loader::parseObj(path, o);
        create((uint32_t)(o.indices.size() / 3), (uint32_t)(o.positions.size() / 3), 1, false);
        m_indices = o.indices;
        std::memcpy((void*)m_positions.data(), o.positions.data(), o.positions.size() * sizeof(float));
        std::memcpy((void*)m_texcoords.data(), o.texcoords.data(), o.texcoords.size() * sizeof(float));
        m_materials[0] = mat;
        m_hasTexcoord = true;
    }
    const uint32_t* indexData() const { return m_indices.data(); }
    const float* positionData() const { return &m_positions[0].x; }
    const float* texcoordData() const { return &m_texcoords[0].x; }
    const uint32_t* primMaterialData() const { return m_primMaterial.data(); }
    const Material* materialData() const { return m_materials.data(); }

private:
    friend class Scene;
    std::vector<uint32_t> m_indices, m_primMaterial;
    std::vector<Vector3f> m_positions, m_normals;
    std::vector<Vector2f> m_texcoords;
    std::vector<Material> m_materials;
    bool m_hasVertexNormal = false, m_hasTexcoord = false;
};

// Bvh::build(Mesh&&) (bvh.h:107): takes ownership of the mesh; the tree itself is built by the C ABI (host-side
// binned SAH identical to the reference's, then flattened to the device layout) when the Bvh is added to a Scene.
class Bvh {
public:
    void build(Mesh&& mesh) { m_mesh = std::move(mesh); }
    const Mesh& mesh() const { return m_mesh; }
private:
    friend class Scene;
    Mesh m_mesh;
    int m_id = -1;
};

class Camera {                                                                      // camera.h:14-53
public:
    void create(const Vector3f& pos, const Vector3f& dir, uint32_t width, uint32_t height) { m_pos = pos; m_dir = dir; m_width = width; m_height = height; }
    const Vector3f& getPosition() const { return m_pos; }
    uint32_t getWidth() const { return m_width; }
    uint32_t getHeight() const { return m_height; }
private:
    friend class PathTracer;
    friend class GbufferVisualizer;
    friend class Scene;
    Vector3f m_pos, m_dir;
    uint32_t m_width = 0, m_height = 0;
};

class Image {                                                                       // image.h:8-30
public:
    Image(uint32_t width, uint32_t height, bool tonemap = true, float exposure = 1.0f)
        : m_pixels(3 * (size_t)width * height, 0.0f), m_width(width), m_height(height), m_tonemap(tonemap), m_exposure(exposure) {}
    void writePixel(uint32_t x, uint32_t y, const Vector3f& c) {                    // image.cpp:44-50
        float* o = &m_pixels[3 * ((size_t)y * m_width + x)];
        o[0] = m_exposure * c.x; o[1] = m_exposure * c.y; o[2] = m_exposure * c.z;
    }
    uint32_t getWidth() const { return m_width; }
    uint32_t getHeigit() const { return m_height; }                                 // (sic) image.h:19
    float* data() { return m_pixels.data(); }
    const float* data() const { return m_pixels.data(); }
    float getExposure() const { return m_exposure; }

    // Image::savePpm (image.cpp:52-80): same header, same payload bytes (tone-map, clamp, gamma, truncation on the device).
    void savePpm(const char* path) const {
        std::vector<uint8_t> bytes(m_pixels.size());
        check(prt_cuda_image_to_ppm(ctx(), m_pixels.data(), m_width, m_height, m_tonemap ? 1 : 0, bytes.data()));
        FILE* fp = std::fopen(path, "wb");
        if (!fp) throw std::runtime_error(std::string("Error saving ") + path);
        std::fprintf(fp, "P6\n%d %d\n255\n", (int)m_width, (int)m_height);
        std::fwrite(bytes.data(), bytes.size(), 1, fp);
        std::fclose(fp);
    }
    // Image::saveExr (image.cpp:82-139): channels B, G, R stored as HALF with tinyexr's FP32->half rounding (done on
    // the device).  The reference lets tinyexr ZIP-compress the scanlines; this writer stores them uncompressed —
    // any EXR reader decodes both files to the same pixels.
    void saveExr(const char* path) const {
        const size_t n = (size_t)m_width * m_height;
        std::vector<uint16_t> planes(3 * n);
        check(prt_cuda_image_to_half_bgr(ctx(), m_pixels.data(), m_width, m_height, planes.data()));
        std::vector<uint8_t> f;
        auto put = [&](const void* p, size_t k) { const uint8_t* b = (const uint8_t*)p; f.insert(f.end(), b, b + k); };
        auto str = [&](const char* s) { put(s, std::strlen(s) + 1); };
        auto i32 = [&](int32_t v) { put(&v, 4); };
        auto f32 = [&](float v) { put(&v, 4); };
        auto attr = [&](const char* name, const char* type, int32_t size) { str(name); str(type); i32(size); };
        const uint8_t magic[8] = {0x76, 0x2f, 0x31, 0x01, 2, 0, 0, 0};
        put(magic, 8);
        attr("channels", "chlist", 3 * 18 + 1);
        for (const char* c : {"B", "G", "R"}) { str(c); i32(1 /*HALF*/); const uint8_t lin[4] = {0, 0, 0, 0}; put(lin, 4); i32(1); i32(1); }
        f.push_back(0);
        attr("compression", "compression", 1); f.push_back(0);
        attr("dataWindow", "box2i", 16); i32(0); i32(0); i32((int32_t)m_width - 1); i32((int32_t)m_height - 1);
        attr("displayWindow", "box2i", 16); i32(0); i32(0); i32((int32_t)m_width - 1); i32((int32_t)m_height - 1);
        attr("lineOrder", "lineOrder", 1); f.push_back(0);
        attr("pixelAspectRatio", "float", 4); f32(1.0f);
        attr("screenWindowCenter", "v2f", 8); f32(0.0f); f32(0.0f);
        attr("screenWindowWidth", "float", 4); f32(1.0f);
        f.push_back(0);
        const uint64_t lineBytes = 8 + 3 * 2 * (uint64_t)m_width;
        uint64_t off = f.size() + 8 * (uint64_t)m_height;
        for (uint32_t y = 0; y < m_height; y++, off += lineBytes) put(&off, 8);
        for (uint32_t y = 0; y < m_height; y++) {
            i32((int32_t)y); i32((int32_t)(3 * 2 * m_width));
            for (int c = 0; c < 3; c++) put(&planes[c * n + (size_t)y * m_width], 2 * (size_t)m_width);
        }
        FILE* fp = std::fopen(path, "wb");
        if (!fp) throw std::runtime_error(std::string("Error saving ") + path);
        std::fwrite(f.data(), f.size(), 1, fp);
        std::fclose(fp);
    }
    // The conversions run on the GPU of the Scene that rendered into this image (set by TraceBlock), or bind one here.
    void bind(prt_ctx* ctx) { m_ctx = ctx; }
private:
    prt_ctx* ctx() const { if (!m_ctx) throw std::runtime_error("Image: not bound to a Scene (render into it or call bind)"); return m_ctx; }
    void check(int r) const { if (r < 0) throw std::runtime_error(std::string("prt_cuda: ") + prt_cuda_last_error(m_ctx)); }
    std::vector<float> m_pixels;
    uint32_t m_width, m_height;
    bool m_tonemap;
    float m_exposure;
    prt_ctx* m_ctx = nullptr;
};

class Scene {                                                                       // scene.h:11-72
public:
    explicit Scene(int device = 0) { if (prt_cuda_create(device, &m_ctx) != PRT_OK) throw std::runtime_error(prt_cuda_last_error(nullptr)); }
    // One Scene on several GPUs: the scene is replicated, PathTracer::TraceBlock deals the block's rows to the
    // GPUs in bands and gathers the framebuffer with one NCCL all-gather, ray batches are split across the GPUs' PCIe links.
    // This replaces the reference's tile loop over its thread pool (main.cpp:121-158): one TraceBlock call per frame.
    explicit Scene(const std::vector<int>& devices) {
        if (prt_cuda_create_multi(devices.data(), (int)devices.size(), &m_ctx) != PRT_OK) throw std::runtime_error(prt_cuda_last_error(nullptr));
    }
    static std::vector<int> allDevices() { std::vector<int> d(prt_cuda_device_count()); for (size_t i = 0; i < d.size(); i++) d[i] = (int)i; return d; }
    int deviceCount() const { int n = 1; prt_cuda_comm_info(m_ctx, nullptr, nullptr, &n); return n; }
    ~Scene() { prt_cuda_destroy(m_ctx); }
    Scene(const Scene&) = delete;
    Scene& operator=(const Scene&) = delete;
    void init() {}                                                                  // scene.cpp:10-17: a fresh context is already empty
    // No counterpart in the reference (it has one traversal): the opt-in fast traversal mode (include/prt_cuda.h, options).
    // buildFastTraversal() before the first add(); setFastTraversal(true / false) switches at any time afterwards.
    void buildFastTraversal(bool on = true) { check(prt_cuda_set_option(m_ctx, PRT_OPTION_BUILD_FAST_TRAVERSAL, on ? 1 : 0)); }
    void setFastTraversal(bool on) { check(prt_cuda_set_option(m_ctx, PRT_OPTION_TRAVERSAL, on ? PRT_TRAVERSAL_FAST : PRT_TRAVERSAL_EXACT)); }
    bool fastTraversal() const { int v = 0; prt_cuda_get_option(m_ctx, PRT_OPTION_TRAVERSAL, &v); return v == PRT_TRAVERSAL_FAST; }

    void add(Bvh* bvh) {                                                            // scene.cpp:19-27
        Mesh& m = bvh->m_mesh;
        std::vector<prt_material> mats(m.m_materials.size());
        for (size_t i = 0; i < mats.size(); i++) {
            const Material& s = m.m_materials[i];
            prt_material& d = mats[i];
            d.diffuse[0] = s.diffuse.x; d.diffuse[1] = s.diffuse.y; d.diffuse[2] = s.diffuse.z;
            d.emissive[0] = s.emissive.x; d.emissive[1] = s.emissive.y; d.emissive[2] = s.emissive.z;
            d.reflType = (uint32_t)s.reflectionType;
            d.diffuseTex = upload(s.diffuseMap); d.bumpTex = upload(s.bumpMap); d.alphaTest = 0;
        }
        prt_mesh_desc md;
        md.primCount = m.getPrimCount(); md.vertexCount = m.getVertexCount(); md.materialCount = (uint32_t)mats.size();
        md.indices = m.m_indices.data(); md.positions = &m.m_positions[0].x;
        md.normals = m.m_hasVertexNormal ? &m.m_normals[0].x : nullptr;
        md.texcoords = m.m_hasTexcoord ? &m.m_texcoords[0].x : nullptr;
        md.primMaterial = m.m_primMaterial.data(); md.materials = mats.data();
        bvh->m_id = check(prt_cuda_scene_add_mesh(m_ctx, &md));
        for (size_t i = 0; i < mats.size(); i++) m.m_materials[i].alphaTest = mats[i].alphaTest != 0;
        m_bvh.push_back(bvh);
    }
    // Scene::getSurfaceProperties<SurfaceProperties, RayHit> (scene.cpp:30-45): one hit, and the batch form a device needs.
    void getSurfaceProperties(SurfaceProperties& properties, const RayHit& hit) const { getSurfaceProperties(&properties, &hit, 1); }
    void getSurfaceProperties(SurfaceProperties* properties, const RayHit* hits, uint32_t n) const {
        std::vector<prt_surface> s(n);
        check(prt_cuda_surface_properties(m_ctx, n, reinterpret_cast<const prt_hit*>(hits), s.data()));
        for (uint32_t i = 0; i < n; i++) {
            SurfaceProperties& p = properties[i];
            p = SurfaceProperties();
            if (!hits[i].isHit()) continue;
            p.normal = Vector3f(s[i].normal[0], s[i].normal[1], s[i].normal[2]);
            p.material = &m_bvh[hits[i].meshId]->m_mesh.m_materials[s[i].material];
            p.uv = Vector2f(s[i].uv[0], s[i].uv[1]); p.duv01 = Vector2f(s[i].duv01[0], s[i].duv01[1]); p.duv02 = Vector2f(s[i].duv02[0], s[i].duv02[1]);
            p.dp01 = Vector3f(s[i].dp01[0], s[i].dp01[1], s[i].dp01[2]); p.dp02 = Vector3f(s[i].dp02[0], s[i].dp02[1], s[i].dp02[2]);
        }
    }
    void setDirectionalLight(const Vector3f& dir, const Vector3f& intensity) {     // scene.h:30-35
        check(prt_cuda_scene_set_directional_light(m_ctx, &dir.x, &intensity.x));
        m_lights = 1u << (uint32_t)LightType::kDirectional;
    }
    // The decoded RGBA32F lat-long image (what Texture::loadExr produces) ...
    void setInfiniteAreaLight(const float* rgba, int width, int height) {
        check(prt_cuda_scene_set_infinite_light(m_ctx, width, height, rgba));
        m_lights = 1u << (uint32_t)LightType::kInfiniteArea;
    }
    // ... or the reference's form: an .exr path (scene.h:41-45 -> InfiniteAreaLight::create -> Texture::loadExr, light.cpp:30-32)
    void setInfiniteAreaLight(const char* path) {
        Texture t; t.loadExr(path);
        setInfiniteAreaLight(reinterpret_cast<const float*>(t.texels.data()), t.width, t.height);
    }
    bool isLightAvailable(LightType t) const { return (m_lights & (1u << (uint32_t)t)) != 0; }
    float getRadius() const { float r; check(prt_cuda_scene_get_radius(m_ctx, &r)); return r; }

    // One-ray forms of Scene::intersect / Scene::occluded (scene.cpp:47-94), kept for API compatibility.
    void intersect(RayHit& hit, const Ray& ray) const { intersect(&hit, &ray, 1); }
    bool occluded(const Ray& ray) const { uint8_t o; occluded(&o, &ray, 1); return o != 0; }
    // Batch forms: single-ray rules (what every secondary ray uses).
    void intersect(RayHit* hits, const Ray* rays, uint32_t n) const {
        std::vector<prt_ray> r = pack(rays, n);
        check(prt_cuda_intersect(m_ctx, n, r.data(), nullptr, PRT_MODE_SINGLE, reinterpret_cast<prt_hit*>(hits)));
    }
    void occluded(uint8_t* out, const Ray* rays, uint32_t n) const {
        std::vector<prt_ray> r = pack(rays, n);
        check(prt_cuda_occluded(m_ctx, n, r.data(), PRT_MODE_SINGLE, out));
    }
    // Packet forms: rays in groups of 8 sharing avgDir (what TracePacket uses for primary rays).
    void intersectPackets(RayHit* hits, const Ray* rays, const Vector3f* avgDir, uint32_t nPackets) const {
        std::vector<prt_ray> r = pack(rays, nPackets * 8);
        check(prt_cuda_intersect(m_ctx, nPackets * 8, r.data(), &avgDir[0].x, PRT_MODE_PACKET, reinterpret_cast<prt_hit*>(hits)));
    }
    // A closest-hit batch and an any-hit batch in ONE library call (prt_cuda_trace_jobs): the upload of the second batch runs
    // under the traversal and download of the first.  Either count may be 0.  Same results as the two calls above.
    void intersectAndOccluded(RayHit* hits, const Ray* rays, uint32_t n, uint8_t* out, const Ray* shadowRays, uint32_t nShadow) const {
        std::vector<prt_ray> r = pack(rays, n), s = pack(shadowRays, nShadow);
        prt_trace_job jobs[2] = {};
        jobs[0].kind = PRT_JOB_INTERSECT; jobs[0].n = n; jobs[0].mode = PRT_MODE_SINGLE; jobs[0].rays = r.data(); jobs[0].out = hits;
        jobs[1].kind = PRT_JOB_OCCLUDED; jobs[1].n = nShadow; jobs[1].mode = PRT_MODE_SINGLE; jobs[1].rays = s.data(); jobs[1].out = out;
        check(prt_cuda_trace_jobs(m_ctx, 2, jobs));
    }
    prt_ctx* context() const { return m_ctx; }

private:
    friend class PathTracer;
    friend class GbufferVisualizer;
    int check(int r) const { if (r < 0) throw std::runtime_error(std::string("prt_cuda: ") + prt_cuda_last_error(m_ctx)); return r; }
    int upload(const Texture& t) { return t.isValid() ? check(prt_cuda_scene_add_texture(m_ctx, t.width, t.height, t.component, t.isFloat ? 1 : 0, t.texels.data())) : -1; }
    static std::vector<prt_ray> pack(const Ray* rays, uint32_t n) {
        std::vector<prt_ray> r(n);
        for (uint32_t i = 0; i < n; i++) {
            r[i].org[0] = rays[i].org.x; r[i].org[1] = rays[i].org.y; r[i].org[2] = rays[i].org.z;
            r[i].dir[0] = rays[i].dir.x; r[i].dir[1] = rays[i].dir.y; r[i].dir[2] = rays[i].dir.z;
            r[i].maxT = rays[i].maxT; r[i].pad = 0.0f;
        }
        return r;
    }
    prt_ctx* m_ctx = nullptr;
    uint32_t m_lights = 0;
    std::vector<Bvh*> m_bvh;                                                        // non-owning, as the reference (scene.h:70)
};

class PathTracer {                                                                  // path_tracer.h:16-38
public:
    // The reference seeds its xorshift32 with rand() per tile task (random.h:15-17), which is not reproducible; here the
    // stream is reseeded per packet from (seed, pixel, packet) — SURVEY §8c P2 — so any tiling gives the same image.
    void setSeed(uint32_t seed) { m_seed = seed; }
    // x1 / y1 are INCLUSIVE, as in the reference (path_tracer.cpp:23-24).  Call it once per frame or per large block, not per
    // 16x16 tile: every call is a wavefront of a few dozen kernel launches (on a multi-GPU Scene: on every GPU, plus the
    // all-gather).  The image must have the camera's dimensions (the reference indexes Image with camera pixel coordinates).
    void TraceBlock(Image& image, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, const Scene& scene, const Camera& camera, uint32_t samples) {
        if (image.getWidth() != camera.m_width || image.getHeigit() != camera.m_height) throw std::runtime_error("PathTracer::TraceBlock: image and camera dimensions differ");
        scene.check(prt_cuda_set_camera(scene.m_ctx, &camera.m_pos.x, &camera.m_dir.x, camera.m_width, camera.m_height));
        scene.check(prt_cuda_render(scene.m_ctx, x0, y0, x1 + 1, y1 + 1, samples, m_seed, image.getExposure(), image.data()));
        image.bind(scene.m_ctx);
    }
private:
    uint32_t m_seed = 0;
};

class GbufferVisualizer {                                                           // gbuffer_visualizer.h:14-34
public:
    enum class Type { kDiffuse, kMeshNormal, kNormal };
    GbufferVisualizer(Type type = Type::kDiffuse) : m_type(type) {}
    void setSeed(uint32_t seed) { m_seed = seed; }                                  // per-pixel reseeding, as PathTracer::setSeed
    // x1 / y1 INCLUSIVE (gbuffer_visualizer.cpp:19-20)
    void TraceBlock(Image& image, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, const Scene& scene, const Camera& camera) {
        if (image.getWidth() != camera.m_width || image.getHeigit() != camera.m_height) throw std::runtime_error("GbufferVisualizer::TraceBlock: image and camera dimensions differ");
        scene.check(prt_cuda_set_camera(scene.m_ctx, &camera.m_pos.x, &camera.m_dir.x, camera.m_width, camera.m_height));
        scene.check(prt_cuda_gbuffer(scene.m_ctx, x0, y0, x1 + 1, y1 + 1, (int)m_type, m_seed, image.getExposure(), image.data()));
        image.bind(scene.m_ctx);
    }
private:
    Type m_type;
    uint32_t m_seed = 0;
};

}  // namespace prt
