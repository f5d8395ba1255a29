/* include/prt_cuda.h — C ABI of the B200-native ray-tracing hot path (libprt_cuda.so).
 *
 * The reference (amada/PRT) has no FFI layer; its seam is the C++ class surface of Scene / Bvh / Camera /
 * PathTracer.  Those classes work one ray or one 8-ray packet per call, which is far too fine
 * for a device launch, so this ABI is batch-shaped: the C++ host classes in prt_b200/host/prt/ (same names
 * and methods as the reference) and the reference-side shim shown in INTEGRATION.md forward to it.
 *
 * Conventions: plain C, opaque context, every function returns 0 on success or a negative prt_status and
 * records a message retrievable with prt_cuda_last_error().  Host buffers are caller-owned.  "_dev" variants
 * take device pointers that live on the context's device and run on the context's stream.  One context per
 * GPU, not re-entrant.  There is NO CPU fallback: without a usable CUDA device every compute entry point
 * fails with PRT_ERR_CUDA.
 */
#ifndef PRT_CUDA_H
#define PRT_CUDA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct prt_ctx prt_ctx;

enum prt_status { PRT_OK = 0, PRT_ERR_CUDA = -1, PRT_ERR_ARG = -2, PRT_ERR_STATE = -3 };

/* Ray record, 32 bytes.  Replaces Ray{org,dir,maxT} (ray.h:14-41) / one lane of SoaRay (ray.h:43-72);
 * invDir and the dominant-axis swap masks of Ray::prepare / SoaRay::prepare are recomputed on the device. */
typedef struct prt_ray { float org[3]; float dir[3]; float maxT; float pad; } prt_ray;

/* Hit record, 24 bytes == RayHit = RayHitT<float,uint32_t> (ray.h:182-198).  Miss: t == -1.0f. */
typedef struct prt_hit { float t, i, j, k; uint32_t primId, meshId; } prt_hit;

/* Traversal rules.  The reference runs two code paths whose results can differ in
 * tie / near-edge cases; the caller says which one a batch belongs to. */
enum prt_ray_mode {
    /* Bvh::intersect<SingleRayHitPacket,SingleRayPacket> (bvh.cpp:429-570) / Bvh::occluded<bool,SingleRayPacket>
     * (bvh.cpp:576-657): Ray::prepare signed dominant axis (ray.h:26-40), strict bool box test at the root /
     * at every node for occlusion (vecmath.h:1449-1466), near-child-first by entry distance (vecmath.h:1402-1424). */
    PRT_MODE_SINGLE = 0,
    /* Bvh::intersect<RayHitPacket,RayPacket> / Bvh::occluded<RayPacketMask,RayPacket>, one lane of the 8-ray
     * packet: SoaRay::prepare abs dominant axis (ray.h:58-71), non-strict SoA box test at pop
     * (vecmath.h:1504-1518), child order fixed by the sign bits of the packet's avgDir (bvh.cpp:444-445,523-529). */
    PRT_MODE_PACKET = 1
};

/* Material, mirrors the fields of Material (material.h:30-46) the hot path reads. */
typedef struct prt_material {
    float diffuse[3];
    float emissive[3];
    uint32_t reflType;    /* ReflectionType: 0 diffuse, 1 specular, 2 refraction (material.h:24-28) */
    int32_t diffuseTex;   /* texture id from prt_cuda_scene_add_texture, or -1 */
    int32_t bumpTex;      /* texture id (single-channel height map, texture.cpp:185-200), or -1 */
    uint32_t alphaTest;   /* out: set by add_mesh as Material::load does (material.cpp:77, texture.cpp:338-350) */
} prt_material;

/* Mesh buffers, mirrors Mesh (mesh.h:83-98) after loading.  texcoords are indexed by position index
 * (mesh.cpp:261).  normals / texcoords may be NULL (geometric normal, mesh.cpp:334; default uv, mesh.cpp:352-354). */
typedef struct prt_mesh_desc {
    uint32_t primCount, vertexCount, materialCount;
    const uint32_t* indices;       /* 3*primCount */
    const float* positions;        /* 3*vertexCount */
    const float* normals;          /* 3*vertexCount or NULL */
    const float* texcoords;        /* 2*vertexCount or NULL */
    const uint32_t* primMaterial;  /* primCount */
    prt_material* materials;       /* materialCount (alphaTest written back) */
} prt_mesh_desc;

typedef struct prt_counters {
    uint64_t extendRays, shadowRays, primaryRays, kernelLaunches;
    float lastKernelMs;            /* CUDA-event time of the most recent traversal / render call, device side */
    uint64_t shadowRaysSkipped;    /* renders: shadow rays not traced because their radiance is exactly zero (not in shadowRays) */
} prt_counters;

/* ---- lifetime ---- */
int prt_cuda_create(int device, prt_ctx** out);
void prt_cuda_destroy(prt_ctx* ctx);
const char* prt_cuda_last_error(const prt_ctx* ctx);   /* ctx may be NULL: last error of a failed create */
int prt_cuda_device_count(void);

/* ---- scene: replaces Scene::init/add/set*Light (scene.h:14-45) + Bvh::build (bvh.cpp:173-228) ---- */
/* channel: 0 = 8-bit unorm, 1 = fp32 (texture.h:8-11); returns texture id >= 0 */
int prt_cuda_scene_add_texture(prt_ctx* ctx, int width, int height, int components, int channel, const void* texels);
/* Builds the BVH on the host with the reference's binned-SAH algorithm (bvh.cpp:31-171), flattens it to the
 * device layout and uploads.  Returns mesh id >= 0 (== Mesh::m_id, scene.cpp:21). */
int prt_cuda_scene_add_mesh(prt_ctx* ctx, prt_mesh_desc* mesh);
int prt_cuda_scene_set_directional_light(prt_ctx* ctx, const float dir[3], const float intensity[3]);
/* rgba: height*width*4 floats, lat-long (what Texture::loadExr produces, texture.cpp:256-336); builds the
 * marginal/conditional CDF tables of InfiniteAreaLight::create (light.cpp:30-82) on the host. */
int prt_cuda_scene_set_infinite_light(prt_ctx* ctx, int width, int height, const float* rgba);
int prt_cuda_scene_get_radius(const prt_ctx* ctx, float* radius);           /* Scene::getRadius (scene.h:58) */
/* Camera::create (camera.h:17-36) */
int prt_cuda_set_camera(prt_ctx* ctx, const float pos[3], const float dir[3], uint32_t width, uint32_t height);

/* ---- host-only introspection (no GPU needed): builder parity ---- */
/* Reference layout of the flattened BVH of mesh `mesh` (LinearBvhNode, bvh.h:49-62): nodes32 = nodeCount x 8
 * words {lower, upper, primOrSecondNodeIndex, triVectorIndex:26|primCount:4|splitAxis:2}; remap = primCount. */
int prt_cuda_bvh_node_count(const prt_ctx* ctx, int mesh, uint32_t* nodeCount);
int prt_cuda_bvh_get(const prt_ctx* ctx, int mesh, uint32_t* nodes32, uint32_t* remap);
/* InfiniteAreaLight tables as built (light.cpp:46-81): vertP[height], horiP[width*height]. */
int prt_cuda_light_get_tables(const prt_ctx* ctx, float* vertP, float* horiP);
/* Camera basis as built: pos, dir, up, right (3 floats each), invWidth, invHeight. */
int prt_cuda_get_camera(const prt_ctx* ctx, float out[14]);

/* ---- hot path 1: closest hit.  Replaces Scene::intersect (scene.cpp:47-63) over a batch. ----
 * mode PRT_MODE_PACKET: rays are consecutive groups of 8 (n % 8 == 0), avgDir has 3 floats per group. */
int prt_cuda_intersect(prt_ctx* ctx, uint32_t n, const prt_ray* rays, const float* avgDir, int mode, prt_hit* hits);
int prt_cuda_intersect_dev(prt_ctx* ctx, uint32_t n, const prt_ray* d_rays, const float* d_avgDir, int mode, prt_hit* d_hits);

/* ---- hot path 2: any hit.  Replaces Scene::occluded (scene.cpp:69-94) over a batch; out[i] = 0/1. ---- */
int prt_cuda_occluded(prt_ctx* ctx, uint32_t n, const prt_ray* rays, int mode, uint8_t* out);
int prt_cuda_occluded_dev(prt_ctx* ctx, uint32_t n, const prt_ray* d_rays, int mode, uint8_t* d_out);

/* ---- hot paths 1 + 2 batched: several closest-hit / any-hit batches in ONE host-buffer call.  The per-ray work of a frame
 * reaches Scene::intersect and Scene::occluded as separate queries (path_tracer.cpp:57-75 primary, :209-238 shadow); handing
 * the library all batches that are ready at once lets it keep the PCIe link busy across batch boundaries (the upload of batch
 * j+1 runs under the traversal and download of batch j) instead of filling and draining the copy pipeline once per call.
 * Results are exactly those of prt_cuda_intersect / prt_cuda_occluded on each job (rays are independent). ---- */
enum prt_job_kind { PRT_JOB_INTERSECT = 0, PRT_JOB_OCCLUDED = 1 };
typedef struct prt_trace_job {
    uint32_t kind;          /* prt_job_kind */
    uint32_t n;             /* rays in this batch (may be 0) */
    int32_t mode;           /* prt_trace_mode */
    uint32_t reserved;
    const prt_ray* rays;    /* HOST, n */
    const float* avgDir;    /* HOST, 3 floats per group of 8 rays; packet-mode closest hit only, else NULL */
    void* out;              /* HOST: prt_hit[n] (PRT_JOB_INTERSECT) or uint8_t[n] (PRT_JOB_OCCLUDED) */
} prt_trace_job;
int prt_cuda_trace_jobs(prt_ctx* ctx, uint32_t count, const prt_trace_job* jobs);

/* ---- hot path 3: primary rays.  Camera::GenerateJitteredRayPacket (camera.cpp:35-73) + packet closest hit for
 * pixels [y0,y1) x [x0,x1), packetsPerPixel packets of 8 samples each, RNG reseeded per packet with
 * prt_packet_seed().  Outputs are DEVICE pointers and may be NULL (rays: 8 per packet, avgDir: 3 floats per
 * packet, hits: 8 per packet).  mode PRT_MODE_SINGLE generates Camera::GenerateJitteredRay rays instead
 * (camera.cpp:12-33; packetsPerPixel then counts single samples). */
int prt_cuda_primary_dev(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t packetsPerPixel,
                         uint32_t seed, int mode, prt_ray* d_rays, float* d_avgDir, prt_hit* d_hits);

/* ---- hot path 4: full path tracing.  Replaces PathTracer::TraceBlock (path_tracer.cpp:17-33) over image rows
 * [y0,y1) x [x0,x1) as a wavefront of kernels; writes exposure*colour/spp (image.cpp:44-50) into the RGB32F
 * row-major image (Image layout, 3 floats per pixel, full-image stride).  spp/8 packets per pixel (integer
 * division as path_tracer.cpp:65).  rgb is a HOST pointer in prt_cuda_render, a device pointer in _dev. */
int prt_cuda_render(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t spp, uint32_t seed,
                    float exposure, float* rgb);
int prt_cuda_render_dev(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t spp, uint32_t seed,
                        float exposure, float* d_rgb);

/* Multi-GPU form of hot path 4: the image is cut into bands of `bandRows` rows dealt round-robin to `world`
 * ranks (band b belongs to rank b % world); this call renders rank `rank`'s bands and writes them consecutively
 * (tile-major) into d_local, a DEVICE buffer of ceil(ceil(H/bandRows)/world)*bandRows rows x width x 3 floats — exactly
 * the send buffer of the framebuffer all-gather, so no repack precedes the collective.  Rows past the image end are
 * written as zeros.  Per-packet seeding makes the pixels independent of the partition (1/2/4/8-GPU images are identical). */
int prt_cuda_render_bands_dev(prt_ctx* ctx, uint32_t bandRows, uint32_t rank, uint32_t world, uint32_t spp, uint32_t seed,
                              float exposure, float* d_local);

/* ---- G-buffer pass.  Replaces GbufferVisualizer::TraceBlock (gbuffer_visualizer.cpp:17-51) over
 * pixels [y0,y1) x [x0,x1): one jittered single-ray primary per pixel (the ray of prt_cuda_primary_dev in
 * PRT_MODE_SINGLE at 1 spp with the same seed), closest hit, surface properties, then the selected quantity times
 * exposure into the RGB32F row-major image.  Misses write 0. ---- */
enum prt_gbuffer_type {
    PRT_GBUFFER_DIFFUSE = 0,       /* Material::sampleDiffuse(uv)            (gbuffer_visualizer.cpp:43) */
    PRT_GBUFFER_MESH_NORMAL = 1,   /* Material::sampleBump(prop)*0.5 + 0.5   (gbuffer_visualizer.cpp:44) */
    PRT_GBUFFER_NORMAL = 2         /* same expression                        (gbuffer_visualizer.cpp:45) */
};
int prt_cuda_gbuffer(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, int type, uint32_t seed,
                     float exposure, float* rgb);
int prt_cuda_gbuffer_dev(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, int type, uint32_t seed,
                         float exposure, float* d_rgb);

/* ---- output stage ----
 * image_to_ppm: the 8-bit payload Image::savePpm writes after its "P6\nW H\n255\n" header (image.cpp:60-69):
 * c/(c+1) when tonemap, clamp to [0,1], powf(c, 1/2.2f)*255, truncation; out = 3*width*height bytes, RGB interleaved.
 * image_to_half_bgr: the half-precision channel planes Image::saveExr hands to tinyexr (image.cpp:82-139), converted
 * with tinyexr's float_to_half_full rounding (tinyexr.h:899-934); planes = 3 x width*height uint16 in B, G, R order. */
int prt_cuda_image_to_ppm(prt_ctx* ctx, const float* rgb, uint32_t width, uint32_t height, int tonemap, uint8_t* out);
int prt_cuda_image_to_ppm_dev(prt_ctx* ctx, const float* d_rgb, uint32_t width, uint32_t height, int tonemap, uint8_t* d_out);
int prt_cuda_image_to_half_bgr(prt_ctx* ctx, const float* rgb, uint32_t width, uint32_t height, uint16_t* planes);
int prt_cuda_image_to_half_bgr_dev(prt_ctx* ctx, const float* d_rgb, uint32_t width, uint32_t height, uint16_t* d_planes);

/* ---- plumbing ---- */
int prt_cuda_synchronize(prt_ctx* ctx);
int prt_cuda_get_counters(prt_ctx* ctx, prt_counters* out);
int prt_cuda_reset_counters(prt_ctx* ctx);
/* cudaStream_t of the context, as an integer handle (for CUDA-event timing on the launching stream) */
uint64_t prt_cuda_stream(prt_ctx* ctx);
/* Measurement aid: L2 read bandwidth (GB/s, 32 MiB buffer, ld.global.cg) and FP32 warp-instruction issue rate
 * (10^9 warp instructions / s) of the context's device, from two micro-benchmark kernels.  No reference counterpart. */
int prt_cuda_probe_peaks(prt_ctx* ctx, double* l2ReadGBs, double* fp32WarpGinstS);
/* Per-packet xorshift32 seed shared with the oracle: fmix32(seed*0x9E3779B9 + pixel*packetsPerPixel + packet) | 1 */
uint32_t prt_packet_seed(uint32_t seed, uint32_t pixel, uint32_t packetsPerPixel, uint32_t packet);

#ifdef __cplusplus
}
#endif
#endif /* PRT_CUDA_H */
