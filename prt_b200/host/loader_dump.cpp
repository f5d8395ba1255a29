// loader_dump.cpp — loads an OBJ (+MTL +TGA) with the reference-shaped host API (prt_host.h / prt_loader.h), dumps every buffer in the layout of
// oracle/ref_harness.cpp:ref_load_obj_dump, so tests/test_loader.py can compare byte for byte with the reference loader.
//   loader_dump <scene.obj> <out.bin>
//   loader_dump --exr <image.exr> <out.bin>      (Texture::loadExr against tinyexr's LoadEXR)
#include <cstdio>
#include "prt_host.h"

using namespace prt;

int main(int argc, char** argv) {
    if (argc < 3) { std::fprintf(stderr, "usage: loader_dump scene.obj out.bin | loader_dump --exr image.exr out.bin\n"); return 2; }
    if (argc >= 4 && std::string(argv[1]) == "--exr") {                        // Texture::loadExr: width, height, then RGBA32F pixels
        try {
            Texture t; t.loadExr(argv[2]);
            FILE* f = std::fopen(argv[3], "wb");
            if (!f) return 2;
            uint32_t hdr[2] = {t.width, t.height};
            std::fwrite(hdr, 4, 2, f); std::fwrite(t.texels.data(), 1, t.texels.size(), f);
            std::fclose(f);
            return 0;
        } catch (const std::exception& e) { std::fprintf(stderr, "%s\n", e.what()); return 1; }
    }
    try {
        Mesh mesh;
        mesh.loadObj(argv[1]);
        const Mesh& m = mesh;
        FILE* f = std::fopen(argv[2], "wb");
        if (!f) return 2;
        uint32_t hdr[3] = {m.getPrimCount(), m.getVertexCount(), m.getMaterialCount()};
        std::fwrite(hdr, 4, 3, f);
        std::fwrite(m.indexData(), 4, 3 * (size_t)hdr[0], f);
        std::fwrite(m.positionData(), 4, 3 * (size_t)hdr[1], f);
        std::fwrite(m.texcoordData(), 4, 2 * (size_t)hdr[1], f);
        std::fwrite(m.primMaterialData(), 4, hdr[0], f);
        for (uint32_t i = 0; i < hdr[2]; i++) {
            const Material& mt = m.materialData()[i];
            float v[6] = {mt.diffuse.x, mt.diffuse.y, mt.diffuse.z, mt.emissive.x, mt.emissive.y, mt.emissive.z};
            uint32_t u[2] = {(uint32_t)mt.reflectionType, mt.alphaTest ? 1u : 0u};
            std::fwrite(v, 4, 6, f); std::fwrite(u, 4, 2, f);
            for (const Texture* t : {&mt.diffuseMap, &mt.bumpMap}) {
                uint32_t d[3] = {0, 0, 0};
                if (t->isValid()) { d[0] = t->width; d[1] = t->height; d[2] = t->component; }
                std::fwrite(d, 4, 3, f);
                if (t->isValid()) std::fwrite(t->texels.data(), 1, t->texels.size(), f);
            }
        }
        std::fclose(f);
        std::printf("loaded %u triangles, %u vertices, %u materials\n", hdr[0], hdr[1], hdr[2]);
        // hand the buffers to the C ABI (host-only context: BVH build, texture registration, validation — no GPU needed)
        try {
            Scene scene(-1);
            Bvh* bvh = new Bvh;
            bvh->build(std::move(mesh));
            scene.add(bvh);
            std::printf("scene.add: ok, radius %.9g\n", scene.getRadius());
        } catch (const std::exception& e) { std::printf("scene.add: %s\n", e.what()); }
    } catch (const std::exception& e) { std::fprintf(stderr, "error: %s\n", e.what()); return 1; }
    return 0;
}
