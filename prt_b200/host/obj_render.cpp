// obj_render.cpp — disk -> device path with the reference-shaped host API, in the style of the reference's Sponza setup
// (main.cpp:58-105): Mesh::loadObj (OBJ + MTL + TGA / PNG), Bvh::build, Scene::add, directional light, Camera, TraceBlock.
//   obj_render <scene.obj> <out.raw> W H spp seed device  px py pz  dx dy dz  lx ly lz  ir ig ib  [sky.exr]
#include <cstdio>
#include <cstdlib>
#include "prt_host.h"

using namespace prt;

int main(int argc, char** argv) {
    if (argc < 20) { std::fprintf(stderr, "usage: obj_render scene.obj out.raw W H spp seed device pos(3) dir(3) lightDir(3) lightIntensity(3)\n"); return 2; }
    const uint32_t W = std::atoi(argv[3]), H = std::atoi(argv[4]), spp = std::atoi(argv[5]), seed = std::atoi(argv[6]);
    const int device = std::atoi(argv[7]);
    float a[12];
    for (int i = 0; i < 12; i++) a[i] = (float)std::atof(argv[8 + i]);
    try {
        Scene scene(device);
        scene.init();
        Mesh mesh;
        mesh.loadObj(argv[1]);
        Bvh* bvh = new Bvh;
        bvh->build(std::move(mesh));
        scene.add(bvh);
        if (argc > 20) scene.setInfiniteAreaLight(argv[20]);                     // an .exr lat-long sky instead (scene.h:41-45)
        else scene.setDirectionalLight(Vector3f(a[6], a[7], a[8]), Vector3f(a[9], a[10], a[11]));
        Camera camera;
        camera.create(Vector3f(a[0], a[1], a[2]), Vector3f(a[3], a[4], a[5]), W, H);
        Image image(W, H, true, 1.0f);
        PathTracer tracer;
        tracer.setSeed(seed);
        tracer.TraceBlock(image, 0, 0, W - 1, H - 1, scene, camera, spp);
        FILE* o = std::fopen(argv[2], "wb");
        std::fwrite(image.data(), 4, 3 * (size_t)W * H, o);
        std::fclose(o);
        std::printf("rendered %ux%u at %u spp, radius %.9g\n", W, H, spp, scene.getRadius());
    } catch (const std::exception& e) { std::fprintf(stderr, "error: %s\n", e.what()); return 1; }
    return 0;
}
