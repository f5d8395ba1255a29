// demo.cpp — a driver written against the reference-shaped C++ host API (prt_host.h), in the style of the reference's
// raytrace_scene() (main.cpp:107-190).  Reads a scene dump (written by tests/test_host_cpp.py), builds Mesh/Bvh/Scene/
// Camera, renders tiles with PathTracer::TraceBlock and writes the RGB32F image.
//   demo <scene.bin> <out.raw> <width> <height> <spp> <seed> <device> [tile]
// <device>: one CUDA device index, -1 (host-only), "all", or a comma list ("0,1,2,3"): several GPUs behind ONE Scene — the
// block's rows are dealt to the GPUs in bands and the framebuffer is gathered with NCCL inside PathTracer::TraceBlock.
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <memory>
#include "prt_host.h"

using namespace prt;

static bool rd(FILE* f, void* p, size_t n) { return fread(p, 1, n, f) == n; }

int main(int argc, char** argv) {
    if (argc < 8) { fprintf(stderr, "usage: demo scene.bin out.raw W H spp seed device [tile]\n"); return 2; }
    uint32_t W = atoi(argv[3]), H = atoi(argv[4]), spp = atoi(argv[5]), seed = atoi(argv[6]);
    int device = atoi(argv[7]);
    std::vector<int> devices;
    if (!strcmp(argv[7], "all")) devices = Scene::allDevices();
    else if (strchr(argv[7], ',')) for (const char* p = argv[7]; *p; ) { devices.push_back(atoi(p)); p = strchr(p, ','); if (!p) break; p++; }
    if (!devices.empty()) device = devices[0];
    uint32_t tile = argc > 8 ? atoi(argv[8]) : 0;
    FILE* f = fopen(argv[1], "rb");
    if (!f) { perror("scene"); return 2; }
    try {
        std::unique_ptr<Scene> scenePtr(devices.size() > 1 ? new Scene(devices) : new Scene(device));
        Scene& scene = *scenePtr;
        scene.init();
        printf("gpus %d\n", scene.deviceCount());
        const bool fast = getenv("PRT_DEMO_FAST") && atoi(getenv("PRT_DEMO_FAST")) != 0;     // the opt-in fast traversal mode
        if (fast) scene.buildFastTraversal();
        uint32_t nmesh = 0;
        rd(f, &nmesh, 4);
        for (uint32_t m = 0; m < nmesh; m++) {
            uint32_t hdr[4];                                     // nprim, nvert, nmat, hasNormals
            rd(f, hdr, 16);
            Mesh mesh;
            mesh.create(hdr[0], hdr[1], hdr[2], hdr[3] != 0);
            rd(f, mesh.getIndexBuffer(), 12 * (size_t)hdr[0]);
            rd(f, mesh.getPositionBuffer(), 12 * (size_t)hdr[1]);
            if (hdr[3]) rd(f, mesh.getNormalBuffer(), 12 * (size_t)hdr[1]);
            rd(f, mesh.getPrimMateialBuffer(), 4 * (size_t)hdr[0]);
            for (uint32_t i = 0; i < hdr[2]; i++) {
                float v[6]; uint32_t refl;
                rd(f, v, 24); rd(f, &refl, 4);
                Material& mat = mesh.getMaterialBuffer()[i];
                mat.init();
                mat.diffuse = Vector3f(v[0], v[1], v[2]); mat.emissive = Vector3f(v[3], v[4], v[5]);
                mat.reflectionType = (ReflectionType)refl;
            }
            if (hdr[3] == 2) mesh.calculateVertexNormals();      // 2: recompute normals with the reference's algorithm
            Bvh* bvh = new Bvh;
            bvh->build(std::move(mesh));
            scene.add(bvh);
        }
        uint32_t lightType = 0;
        rd(f, &lightType, 4);
        if (lightType == 1) { float v[6]; rd(f, v, 24); scene.setDirectionalLight(Vector3f(v[0], v[1], v[2]), Vector3f(v[3], v[4], v[5])); }
        float cam[6];
        rd(f, cam, 24);
        fclose(f);
        Camera camera;
        camera.create(Vector3f(cam[0], cam[1], cam[2]), Vector3f(cam[3], cam[4], cam[5]), W, H);
        printf("radius %.9g\n", scene.getRadius());
        if (fast) { scene.setFastTraversal(true); printf("traversal %s\n", scene.fastTraversal() ? "fast" : "exact"); }
        if (device < 0) { printf("host-only context: scene built, not rendering\n"); return 0; }
        Image image(W, H, true, 1.0f);
        PathTracer tracer;
        tracer.setSeed(seed);
        if (tile == 0) tracer.TraceBlock(image, 0, 0, W - 1, H - 1, scene, camera, spp);
        else                                                      // non-overlapping tiles (the reference's 17x17 overlap is a quirk)
            for (uint32_t y = 0; y < H; y += tile)
                for (uint32_t x = 0; x < W; x += tile)
                    tracer.TraceBlock(image, x, y, std::min(x + tile, W) - 1, std::min(y + tile, H) - 1, scene, camera, spp);
        // one-ray API forms
        Ray r; r.org = camera.getPosition(); r.dir = Vector3f(cam[3], cam[4], cam[5]); r.maxT = 100000.0f;
        RayHit hit; scene.intersect(hit, r);
        printf("centre ray: t=%.9g prim=%u mesh=%u occluded=%d\n", hit.t, hit.primId, hit.meshId, (int)scene.occluded(r));
        if (hit.isHit()) {                                        // Scene::getSurfaceProperties (scene.cpp:30-45)
            SurfaceProperties sp; scene.getSurfaceProperties(sp, hit);
            printf("centre surface: n=%.9g %.9g %.9g uv=%.9g %.9g diffuse=%.9g %.9g %.9g\n", sp.normal.x, sp.normal.y, sp.normal.z, sp.uv.x, sp.uv.y,
                   sp.material->diffuse.x, sp.material->diffuse.y, sp.material->diffuse.z);
        }
        FILE* o = fopen(argv[2], "wb");
        fwrite(image.data(), 4, 3 * (size_t)W * H, o);
        fclose(o);
        // output stage + G-buffer pass: <out>.ppm, <out>.exr, <out>.normal.raw
        std::string base(argv[2]);
        image.savePpm((base + ".ppm").c_str());
        image.saveExr((base + ".exr").c_str());
        Image aov(W, H, false, 1.0f);
        GbufferVisualizer vis(GbufferVisualizer::Type::kNormal);
        vis.setSeed(seed);
        vis.TraceBlock(aov, 0, 0, W - 1, H - 1, scene, camera);
        o = fopen((base + ".normal.raw").c_str(), "wb");
        fwrite(aov.data(), 4, 3 * (size_t)W * H, o);
        fclose(o);
    } catch (const std::exception& e) { fprintf(stderr, "error: %s\n", e.what()); return 1; }
    return 0;
}
