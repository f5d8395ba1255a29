/* include/prt_cuda.h — C ABI of the B200-native ray-tracing hot path (libprt_cuda.so).
 *
 * The reference (amada/PRT) has no FFI layer; its seam is the C++ class surface of Scene / Bvh / Camera /
 * PathTracer.  Those classes work one ray or one 8-ray packet per call, which is far too fine
 * for a device launch, so this ABI is batch-shaped: the C++ host classes in prt_b200/host/prt_host.h (same names
 * and methods as the reference) and the reference-side shim shown in INTEGRATION.md forward to it.
 *
 * Conventions: plain C, opaque context, every function returns 0 on success or a negative prt_status and
 * records a message retrievable with prt_cuda_last_error().  Host buffers are caller-owned.  "_dev" variants
 * take device pointers that live on the context's device and run on the context's stream.  A context is bound
 * to its device: every entry point switches to it for the duration of the call and restores the caller's current
 * device, so contexts on different GPUs can be interleaved from one thread and a context can be used from any
 * thread (one call at a time per context: not re-entrant).  There is NO CPU fallback: without a usable CUDA
 * device every compute entry point fails with PRT_ERR_CUDA.
 *
 * Multi-GPU comes in two forms with the same data path — the scene replicated per GPU, the rows
 * of a render call dealt to the GPUs in bands, ONE collective (ncclAllGather of the band buffers over NVLink /
 * NVSwitch) and a de-tile kernel:
 *   - one process, several GPUs: prt_cuda_create_multi(devices, n) returns ONE context that drives all of them
 *     (one host thread + stream per GPU, ncclCommInitAll); every scene call is replicated, prt_cuda_render[_dev]
 *     shards the region, host-buffer ray batches are split across the GPUs' PCIe links;
 *   - one process per GPU (torchrun / MPI): every rank creates its own context and joins a communicator with
 *     prt_cuda_comm_unique_id / prt_cuda_comm_init_rank; prt_cuda_render[_dev] is then a collective call that leaves
 *     the full region on every rank.
 */
#ifndef PRT_CUDA_H
#define PRT_CUDA_H

#include <stddef.h>
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

/* SurfaceProperties (mesh.h:19-28) of one hit, same field order; `material` is the index into the hit mesh's own
 * material table (what Mesh::getPrimToMaterial(primId) returns, mesh.cpp:362) instead of a host pointer. */
typedef struct prt_surface {
    float normal[3];
    uint32_t material;
    float uv[2], duv01[2], duv02[2];
    float dp01[3], dp02[3];
} prt_surface;

/* Ray classes of the traversal counters below. */
enum prt_ray_class { PRT_CLASS_CLOSEST_SINGLE = 0, PRT_CLASS_CLOSEST_PACKET = 1, PRT_CLASS_OCCLUDE = 2, PRT_CLASS_COUNT = 3 };

typedef struct prt_counters {
    uint64_t extendRays, shadowRays, primaryRays, kernelLaunches;
    float lastKernelMs;            /* CUDA-event time of the most recent traversal / render call, device side */
    uint64_t shadowRaysSkipped;    /* renders: shadow rays not traced because their radiance is exactly zero (not in shadowRays) */
    float lastGatherMs;            /* multi-GPU renders: the framebuffer all-gather + de-tile of the most recent call, device side */
    /* Traversal work as the kernels did it, per ray class — what Stats counts inside the reference's hit packet
     * (stats.h:10-33, bvh.cpp:464-466,555-557).  Filled ONLY by the counting build of the library (libprt_cuda_stats.so,
     * -DPRT_STATS=1; the reference's PRT_ENABLE_STATS, prt.h:5, is a compile-time switch too); zero otherwise.
     *   statRays      rays traced by the dynamic-fetch kernels
     *   statNodeSteps 64-byte two-child node records fetched (= interior nodes the reference pops)
     *   statLeafVisits leaves entered (reference nodesTraversed = statNodeSteps + statLeafVisits + root-only cases)
     *   statTriTests  48-byte triangle records fetched and tested (reference primsTraversed counts the leaf's primCount)
     *   statUvLoads   24-byte uv side records fetched for alpha tests */
    uint64_t statRays[3], statNodeSteps[3], statLeafVisits[3], statTriTests[3], statUvLoads[3];
} prt_counters;

/* ---- lifetime ---- */
int prt_cuda_create(int device, prt_ctx** out);
void prt_cuda_destroy(prt_ctx* ctx);
const char* prt_cuda_last_error(const prt_ctx* ctx);   /* ctx may be NULL: last error of a failed create */
int prt_cuda_device_count(void);

/* ---- multi-GPU (replaces the tile loop over the thread pool, main.cpp:121-158) ----
 * One process, ndev GPUs: one context driving all of them.  Scene calls are replicated (the BVH is built once on the host
 * and uploaded to every GPU), renders are sharded by row bands + all-gathered, host-buffer batches (prt_cuda_intersect /
 * prt_cuda_occluded / prt_cuda_trace_jobs) are split evenly across the GPUs.  "_dev" pointers live on devices[0]. */
int prt_cuda_create_multi(const int* devices, int ndev, prt_ctx** out);
/* One process per GPU: rank 0 obtains an id (ncclGetUniqueId), hands the PRT_COMM_ID_BYTES bytes to every rank by any
 * means (torch.distributed / MPI broadcast, a file), and all ranks call init_rank collectively on single-GPU contexts.
 * From then on prt_cuda_render / prt_cuda_render_dev are collective over the communicator. */
#define PRT_COMM_ID_BYTES 128
int prt_cuda_comm_unique_id(void* id);
int prt_cuda_comm_init_rank(prt_ctx* ctx, const void* id, int rank, int world);
/* rank / world of the context's communicator (0 / 1 without one) and the number of GPUs this context drives itself. */
int prt_cuda_comm_info(const prt_ctx* ctx, int* rank, int* world, int* localDevices);

/* ---- page-locked host memory for the host-buffer entry points.  The reference keeps its buffers in new[] / std::vector
 * memory (image.cpp:30-36, mesh.cpp:87-105); a binding either allocates the ray / hit buffers here or registers the
 * ones it has.  Pageable buffers work too, at the staging-copy rate of the driver (measured in profiles/). ---- */
int prt_cuda_host_alloc(size_t bytes, void** out);
int prt_cuda_host_free(void* p);
int prt_cuda_host_register(void* p, size_t bytes);
int prt_cuda_host_unregister(void* p);

/* ---- options ----
 * PRT_OPTION_BUILD_FAST_TRAVERSAL (0 / 1, default 0; set BEFORE the first prt_cuda_scene_add_mesh): every mesh added afterwards
 *   also gets the structure of the fast traversal mode — a second tree with leaves of <= 4 triangles collapsed to 8-wide nodes
 *   with 8-bit quantised child boxes (SURVEY §8f-1 "flatten-time re-encoding"; prt_b200/csrc/prt_wide.cuh).
 * PRT_OPTION_TRAVERSAL (prt_traversal; needs the option above): which structure rays walk from now on — the secondary and
 *   shadow rays of prt_cuda_render* and the batches of prt_cuda_intersect / occluded / trace_jobs in both modes (the mode still
 *   decides how a ray is prepared: dominant-axis rule, root-box test).  The renderer's camera rays and prt_cuda_primary_dev with
 *   ray outputs stay on the exact path (coherent rays gain nothing from the wide nodes).
 *   PRT_TRAVERSAL_EXACT (default) reproduces the reference bit for bit.  PRT_TRAVERSAL_FAST finds the same closest t with the
 *   same triangle arithmetic, but may report the other triangle where two are hit at exactly the same t, and sees hits the
 *   reference's box tests reject at grazing incidence; renders then differ in the few pixels whose paths meet such a case. */
enum prt_option { PRT_OPTION_BUILD_FAST_TRAVERSAL = 1, PRT_OPTION_TRAVERSAL = 2 };
enum prt_traversal { PRT_TRAVERSAL_EXACT = 0, PRT_TRAVERSAL_FAST = 1 };
int prt_cuda_set_option(prt_ctx* ctx, int option, int value);
int prt_cuda_get_option(const prt_ctx* ctx, int option, int* value);

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
/* The fast traversal mode's structure of a mesh (parity / invariant checks, tests/test_fast_build_cpu.py): Wide8 records as
 * uploaded (24 words each, mesh-local references; prt_b200/csrc/bvh_build.h) and the 48-byte triangle records in that tree's leaf
 * order (12 floats each; vertex 0's w = slot of the triangle in the exact array, bit 31 of vertex 1's w = never seen by the
 * reference's single-ray any-hit traversal).  Any output may be NULL.  PRT_ERR_STATE when the mesh was added without
 * PRT_OPTION_BUILD_FAST_TRAVERSAL, or — for nodes / tris — on device contexts, which drop the host copy after the upload
 * (host-only contexts, device < 0, keep it). */
int prt_cuda_bvh_get_fast(const prt_ctx* ctx, int mesh, uint32_t* nodeCount, uint32_t* rootRef, uint32_t* nodes, float* tris);
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
/* How a job's rays reach the device.  A frame's batches are far from arbitrary — camera rays share their origin
 * (camera.cpp:12-33), directional shadow rays share direction and maxT (path_tracer.cpp:236-238), and both can be built on
 * the device from a few bytes — so the host can send 12 or 0 bytes per ray instead of 32.  Results are bit-identical to
 * the same rays sent as PRT_RAYS_FULL (tests/test_gpu_compact_jobs.py). */
enum prt_ray_layout {
    PRT_RAYS_FULL = 0,           /* rays = prt_ray[n] (32 B per ray) */
    PRT_RAYS_SHARED_ORIGIN = 1,  /* rays = float[3*n] directions; origin = shared[0..2], maxT = shared[3]  (12 B per ray) */
    PRT_RAYS_SHARED_DIR = 2,     /* rays = float[3*n] origins; direction = shared[0..2], maxT = shared[3]  (12 B per ray) */
    /* No upload: the n = (x1-x0)*(y1-y0)*perPixel*(8 in packet mode) camera rays of region[] = {x0,y0,x1,y1}, generated on the
     * device exactly as prt_cuda_primary_dev does (Camera::GenerateJitteredRay / GenerateJitteredRayPacket, per-packet seed);
     * closest-hit jobs only.  avgDir is not needed (computed with the rays). */
    PRT_RAYS_PRIMARY = 3,
    /* No upload: shadow rays of an earlier closest-hit job of the SAME call, jobs[fromJob], toward the light direction
     * shared[0..2]: for every ray of that job that hit, pos = t*dir + org, then org = pos + kFar*L, dir = -L,
     * maxT = kFar - 0.0008 with kFar = 2*Scene::getRadius() (path_tracer.cpp:192-193, 209-212, 236-238).  n must equal that
     * job's n; out[i] = 0 / 1 for a ray that hit, 2 where the source ray missed (no shadow ray).  Any-hit jobs only. */
    PRT_RAYS_FROM_HITS = 4
};
typedef struct prt_trace_job {
    uint32_t kind;          /* prt_job_kind */
    uint32_t n;             /* rays in this batch (may be 0) */
    int32_t mode;           /* prt_ray_mode */
    uint32_t layout;        /* prt_ray_layout (0 = PRT_RAYS_FULL, the round-1 ABI) */
    const void* rays;       /* HOST, per layout; NULL for PRT_RAYS_PRIMARY / PRT_RAYS_FROM_HITS */
    const float* avgDir;    /* HOST, 3 floats per group of 8 rays; packet-mode closest hit with uploaded rays only, else NULL */
    void* out;              /* HOST: prt_hit[n] (PRT_JOB_INTERSECT) or uint8_t[n] (PRT_JOB_OCCLUDED) */
    float shared[4];        /* PRT_RAYS_SHARED_ORIGIN / SHARED_DIR / FROM_HITS, see above */
    uint32_t region[4];     /* PRT_RAYS_PRIMARY: x0, y0, x1, y1 (exclusive) */
    uint32_t perPixel;      /* PRT_RAYS_PRIMARY: samples (single mode) or packets (packet mode) per pixel */
    uint32_t seed;          /* PRT_RAYS_PRIMARY */
    int32_t fromJob;        /* PRT_RAYS_FROM_HITS: index of the source closest-hit job, < this job's index */
    uint32_t reserved;
} prt_trace_job;
int prt_cuda_trace_jobs(prt_ctx* ctx, uint32_t count, const prt_trace_job* jobs);

/* ---- hot path 3: primary rays.  Camera::GenerateJitteredRayPacket (camera.cpp:35-73) + packet closest hit for
 * pixels [y0,y1) x [x0,x1), packetsPerPixel packets of 8 samples each, RNG reseeded per packet with
 * prt_packet_seed().  Outputs are DEVICE pointers and may be NULL (rays: 8 per packet, avgDir: 3 floats per
 * packet, hits: 8 per packet).  mode PRT_MODE_SINGLE generates Camera::GenerateJitteredRay rays instead
 * (camera.cpp:12-33; packetsPerPixel then counts single samples). */
int prt_cuda_primary_dev(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t packetsPerPixel,
                         uint32_t seed, int mode, prt_ray* d_rays, float* d_avgDir, prt_hit* d_hits);

/* ---- No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Allocates nothing — no cudaMalloc /
 * cudaFree / allocating library routine: the single-rule camera rays of pixels [y0,y1) x [x0,x1), perPixel samples each,
 * are traced for their closest hit exactly as prt_cuda_primary_dev(..., PRT_MODE_SINGLE, ...) does, and d_out[i] (DEVICE,
 * one byte per ray) receives what a PRT_RAYS_FROM_HITS job of prt_cuda_trace_jobs would: 1 if the shadow ray from ray i's
 * hit toward lightDir (path_tracer.cpp:236-238) is occluded, 0 if not, 2 if ray i hit nothing.  The intermediate hit
 * records and the work counters sit in a static __device__ array of the library (256 B + 24 B x PRT_NOALLOC_MAX_RAYS
 * = 6 MiB per GPU that loads the library, held until exit); more rays than that -> PRT_ERR_ARG, split the region.
 * The scene must already be committed (any earlier tracing call does it; a changed scene -> PRT_ERR_STATE, because the
 * upload allocates).  Always the exact traversal (PRT_OPTION_TRAVERSAL is ignored).  Asynchronous on the context's
 * stream like the other _dev calls.  All calls on one GPU share the array, whatever the context: calls that overlap on
 * the GPU (two contexts' streams, two threads) race; calls on one stream, or separated by a synchronize, do not. ---- */
#define PRT_NOALLOC_MAX_RAYS 262144u
int prt_cuda_primary_shadow_noalloc(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t perPixel, uint32_t seed,
                                    const float lightDir[3], uint8_t* d_out);

/* ---- hot path 4: full path tracing.  Replaces PathTracer::TraceBlock (path_tracer.cpp:17-33) over image rows
 * [y0,y1) x [x0,x1) as a wavefront of kernels; writes exposure*colour/spp (image.cpp:44-50) into the RGB32F
 * row-major image (Image layout, 3 floats per pixel, full-image stride).  spp/8 packets per pixel (integer
 * division as path_tracer.cpp:65).  rgb is a HOST pointer in prt_cuda_render, a device pointer in _dev. */
int prt_cuda_render(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t spp, uint32_t seed,
                    float exposure, float* rgb);
int prt_cuda_render_dev(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t spp, uint32_t seed,
                        float exposure, float* d_rgb);
/* On a multi-GPU context (prt_cuda_create_multi, or after prt_cuda_comm_init_rank — then collectively on every rank) both
 * calls shard the region's rows over the communicator's GPUs in bands of 8 rows dealt round-robin, all-gather the band
 * buffers with NCCL and de-tile into rgb / d_rgb; per-packet seeding makes the pixels independent of the partition, so the
 * image is bit-identical to the single-GPU one.  prt_cuda_render copies back only the region's pixels: call it once per
 * frame (or per large block), not per 16x16 tile — every call costs a few dozen kernel launches. */

/* Multi-GPU form of hot path 4: the image is cut into bands of `bandRows` rows dealt round-robin to `world`
 * ranks (band b belongs to rank b % world); this call renders rank `rank`'s bands and writes them consecutively
 * (tile-major) into d_local, a DEVICE buffer of ceil(ceil(H/bandRows)/world)*bandRows rows x width x 3 floats — exactly
 * the send buffer of the framebuffer all-gather, so no repack precedes the collective.  Rows past the image end are
 * written as zeros.  Per-packet seeding makes the pixels independent of the partition (1/2/4/8-GPU images are identical). */
int prt_cuda_render_bands_dev(prt_ctx* ctx, uint32_t bandRows, uint32_t rank, uint32_t world, uint32_t spp, uint32_t seed,
                              float exposure, float* d_local);

/* The collective half alone: all-gathers the band buffers the last sharded render of this region left on the GPUs and de-tiles
 * them into d_rgb again (same result).  Measurement aid — it times the framebuffer exchange without a render in front of it
 * (prt_counters.lastGatherMs after a render also contains the wait for the slowest GPU).  Collective like the render itself. */
int prt_cuda_gather_bands_dev(prt_ctx* ctx, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, float* d_rgb);

/* ---- Scene::getSurfaceProperties (scene.cpp:30-45 -> Mesh::getSurfaceProperties, mesh.cpp:311-372) over a batch of hit
 * records as prt_cuda_intersect returns them (t, i, j, k, primId, meshId).  Misses (t == -1) produce an all-zero record.
 * A host integrator that calls Scene::intersect and shades itself gets normals / uv / material / bump frame from here. ---- */
int prt_cuda_surface_properties(prt_ctx* ctx, uint32_t n, const prt_hit* hits, prt_surface* out);
int prt_cuda_surface_properties_dev(prt_ctx* ctx, uint32_t n, const prt_hit* d_hits, prt_surface* d_out);

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
