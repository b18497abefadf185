/* rtb200.h — C ABI of the B200-native ray-traversal engine (librtb200.so).
 *
 * Drop-in boundary for the reference's hot path. The reference (wtzhang23/gpu-ray-tracer) has no
 * FFI layer; its seam is the C++ free-function API in include/raytracer.h:18-28 fed by
 * SceneBuilder::build_gpu_scene (include/scene_builder.h:115). Every entry point below names the
 * reference interface it replaces. Plain pointers and sizes only; all functions return 0 on
 * success or an RtStatus error code (the reference's assert(rv == 0) / unchecked calls become
 * return codes, src/raytracer.cu:83,115; include/gputils/alloc.h:18-19) and never abort.
 *
 * All arithmetic types follow the reference: fp32 geometry/shading, RGBA8 canvas packed as
 * R<<24 | G<<16 | B<<8 | A in an int32 at buffer[y*W + x] (src/rayenv/color.cu:24-26,
 * src/rayenv/canvas.cu:9-12).
 */
#ifndef RTB200_H
#define RTB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum RtStatus {
    RT_OK = 0,
    RT_ERR_INVALID_ARG = 1,
    RT_ERR_CUDA = 2,
    RT_ERR_NO_DEVICE = 3,
    RT_ERR_OUT_OF_MEMORY = 4,
    RT_ERR_NO_BVH = 5,
    RT_ERR_LIMIT = 6
} RtStatus;

/* rprimitives::Material (include/rayprimitives/material.h:14-27): 6 RGBA vectors + alpha + eta. */
typedef struct RtMaterial {
    float Ke[4], Ka[4], Kd[4], Ks[4], Kt[4], Kr[4];
    float alpha, eta;
} RtMaterial;

/* rprimitives::{gpu,cpu}::TriInner (include/rayprimitives/gpu/trimesh.cuh:25-53): vertex index triple, per-triangle
 * material and rprimitives::TextureCoords (include/rayprimitives/texture_coords.h:12-29): atlas origin (texels) and the
 * atlas-space directions of the two barycentric axes. The reference never samples its atlas (`// TODO: implement
 * texture mapping`, src/rayprimitives/phong.cu:19-23), so with tex_degenerate != 0 — every shipped world — the tex_*
 * fields are ignored. With tex_degenerate == 0 and an atlas in the scene (no reference counterpart)
 * the diffuse colour of a hit with Isect::uv = (b1, b2) is the atlas texel at
 *   (tex_x + b1 * tex_u[0] + b2 * tex_v[0],  tex_y + b1 * tex_u[1] + b2 * tex_v[1]),
 * point-sampled, clamped, unnormalised coordinates (the reference's texture descriptor, include/gputils/alloc.h:52-56),
 * and replaces Kd in phong()'s diffuse term. */
typedef struct RtTriangle {
    int32_t v[3];
    int32_t tex_degenerate;
    RtMaterial mat;
    int32_t tex_x, tex_y;
    float tex_u[2], tex_v[2];
} RtTriangle;

/* rprimitives::Entity pose (include/rayprimitives/entity.h:16-30): quaternion (i,j,k,r) + position. */
typedef struct RtPose {
    float rot[4];
    float pos[3];
} RtPose;

/* One hitable (rprimitives::gpu::Hitable, include/rayprimitives/gpu/hitable.cuh:27-47): its own Entity pose and
 *  - RT_HITABLE_TRIMESH: a [tri_begin, tri_end) range into the triangle array (Trimesh, MeshBuilder,
 *    include/scene_builder.h:29-49) — the reference's only Hitable subclass;
 *  - RT_HITABLE_SPHERE (no reference counterpart): an analytic sphere of `radius` centred at the
 *    hitable's local origin. It owns exactly ONE record of the triangle array, [tri_begin, tri_begin + 1), which
 *    carries its Material (and texture flag); that record's vertex indices are ignored. */
enum { RT_HITABLE_TRIMESH = 0, RT_HITABLE_SPHERE = 1 };
typedef struct RtMesh {
    RtPose pose;
    int32_t tri_begin, tri_end;
    int32_t kind;
    float radius;
} RtMesh;

/* renv::Transformation (include/rayenv/transformation.h:13-23): pose + hitable index. */
typedef struct RtInstance {
    RtPose pose;
    int32_t mesh;
} RtInstance;

enum { RT_LIGHT_POINT = 0, RT_LIGHT_DIRECTIONAL = 1 };
/* PointLight / DirLight (include/rayprimitives/gpu/light.cuh:17-67). `v` is the position for a
 * point light and the NORMALISED shine direction for a directional one (set_shine_dir normalises).
 * Order matters for float summation: point lights first, then directional
 * (src/scene_builder.cu:61-81). */
typedef struct RtLight {
    int32_t kind;
    float color[4];
    float v[3];
} RtLight;

/* renv::Camera (include/rayenv/camera.h:15-36, src/rayenv/camera.cu:6-9,32-42). */
typedef struct RtCamera {
    RtPose pose;
    float global_near;
    float unit_to_pixels;
    float canvas_width, canvas_height;
} RtCamera;

/* renv::Environment scalars (include/rayenv/environment.h:19-49). depth <= 9 (RayFrame[10],
 * src/rayenv/scene.cu:25,96). */
typedef struct RtEnv {
    float ambience[4];
    float dist_atten[3];
    int32_t depth;
} RtEnv;

typedef struct RtSceneDesc {
    const float* vertices; /* n_vertices * 3 */
    const float* normals;  /* n_vertices * 3 (SceneBuilder::generate_normals, src/scene_builder.cc:11-29) */
    int32_t n_vertices;
    const RtTriangle* triangles;
    int32_t n_triangles;
    const RtMesh* meshes;
    int32_t n_meshes;
    const RtInstance* instances;
    int32_t n_instances;
    const RtLight* lights;
    int32_t n_lights;
    RtCamera camera;
    RtEnv env;
    int32_t width, height;
    /* texture atlas (rprimitives::gpu::Texture, a float4 2-D texture in the reference): RGBA floats, row-major,
     * atlas_width * atlas_height * 4 values; NULL = no atlas, every triangle shades with its Kd */
    const float* atlas;
    int32_t atlas_width, atlas_height;
} RtSceneDesc;

typedef struct RtScene RtScene; /* opaque; owns all device memory */

/* Which bounce semantics propagate_ray follows. The reference's CUDA and serial paths differ for
 * reflective / refractive materials; RT_SEMANTICS_GPU restates
 * renv::gpu::propagate_ray (src/rayenv/scene.cu:92-188), RT_SEMANTICS_CPU restates
 * renv::cpu::Scene::propagate_helper (src/rayenv/scene.cu:222-268). */
enum { RT_SEMANTICS_GPU = 0, RT_SEMANTICS_CPU = 1 };

enum {
    RT_RENDER_REBUILD_BVH = 1,   /* rebuild the BVH inside the call (reference timing scope) */
    RT_RENDER_NO_BVH = 2,        /* `-r/--unoptimize`: brute force over instances, src/rayenv/scene.cu:48-52 */
    RT_RENDER_CANVAS_DEVICE = 4, /* `canvas` is a device pointer (no D2H copy) */
    RT_RENDER_SORT_RAYS = 8,     /* bounce waves as one launch per phase with the live paths sorted between the waves by the
                                    Morton cell of the pending ray's origin and its direction octant. Same frame; measured
                                    slower than the default (all waves in one cooperative launch, queue kept in tile order):
                                    world8_stress at 4K 2.18 ms of waves against 1.46 ms */
    RT_RENDER_TILED_OUTPUT = 16, /* write only this rank's tiles, compacted (multi-GPU) */
    RT_RENDER_STATS = 32,        /* count nodes / triangle tests per frame (slower kernels) */
    RT_RENDER_FUSED_SHADE = 64,  /* fused shading / bounce kernels instead of the split gen / shadow / resolve pipeline */
    RT_RENDER_ASYNC = 256,       /* rt_render / rt_untile only enqueue on the scene's stream; rt_frame_finish() waits for
                                    the frame and returns its statistics (a host canvas is valid only after that) */
    RT_RENDER_REFIT_BVH = 512,   /* refit the existing BVH inside the call (rt_bvh_refit) instead of rebuilding it; falls back
                                    to a build when there is no tree to refit. RT_RENDER_REBUILD_BVH wins when both are set */
    RT_RENDER_HOST_BANDS = 1024, /* multi-GPU frames straight into ONE host canvas: with RT_RENDER_TILED_OUTPUT and
                                    tile_w >= width (full-width bands of tile_h rows dealt round-robin to the ranks), `canvas` is
                                    the full W x H row-major HOST canvas (e.g. pinned shared memory mapped by every rank) and
                                    the call copies only this rank's bands into it, each GPU over its own PCIe link */
    RT_RENDER_HOST_DIRECT = 2048, /* host canvas in page-locked, device-addressable memory (rt_host_canvas_open, cudaHostAlloc,
                                    cudaHostRegister): the kernels write it directly over PCIe while the frame renders instead
                                    of rendering into a device canvas that is copied afterwards. Falls back to the staged copy
                                    for pageable memory. Opt-in: slower than the staged, chunked copy on B200 */
    RT_RENDER_SPLIT_SHADE = 4096, /* wave 0 as round 1's split pipeline: k_primary with the per-lane BVH8 walk, then k_shadow_gen ->
                                    k_shadow -> k_resolve. Default: the camera rays' closest hits come from the warp-coherent
                                    k_tile<1> (rt_beam.cuh: one walk per 8x4 tile) in front of the same shading kernels. Same frame */
    RT_RENDER_TILE_KERNEL = 16384, /* wave 0 entirely in the tile kernels: closest hits, then shadow rays + Phong + pixel per tile with the
                                    warp-coherent traversal — no shadow-ray queues, 2 launches (1 with RT_TILE_PHASES=1). Same
                                    frame; 26 % fewer instructions but lower issue rate (instruction cache, 128 registers), so it
                                    is opt-in: equal to the default on 1 GPU, slower on 8. Scenes with a refractive
                                    material always shade through the split kernels (closest-hit shadow walks, light.cu:34-60) */
    RT_RENDER_PEER_STORE = 8192, /* multi-GPU frame gathered by the kernels that produce the pixels: this rank's tiles (rank / world
                                    of RtRenderParams) are stored straight into the row-major peer canvases registered with
                                    rt_scene_set_peers over NVLink, and the frame's completion is signalled through their
                                    mailboxes; `canvas` is ignored (may be NULL) */
    RT_RENDER_GRAPH = 32768,     /* the frame as ONE CUDA graph: a frame whose key (these parameters, canvas, camera, environment, instance
                                    and buffer generation, BVH state, exchange parity) equals the previous frame's is captured from the
                                    scene's stream — build kernel, frame kernels, counters copy, the exchange — instantiated and from then
                                    on replayed with one cudaGraphLaunch; a changed key (camera moved, instances set) falls back to plain
                                    launches and a new capture. Device or peer canvas, BVH traversal, RT_SEMANTICS_GPU only; frames whose
                                    enqueue needs the host fall back silently. Per-kernel times stay available (external event nodes) */
    RT_RENDER_FAST_ACCEPT = 65536, /* OPT-IN TOLERANCE MODE (the default is bit-exact): a hit that lies clearly inside a triangle — inside
                                    all edges by twice the filter margin, |d.n| >= 4e-5, t >= 1e-4 — is accepted with the conservative
                                    filter's own t / uv instead of the reference-exact test (40 % of the traversal kernels' instructions);
                                    everything near an edge, the 1e-5 threshold or the light distance still takes the exact path.
                                    Within north_star's tolerances (ids >= 99.99 %, t 1e-4 relative, RMSE 1e-3): measured parity and speed
                                    in profiles/r2_fast_accept_parity.md. Ignored for scenes with a rotated instance or mesh */
    RT_RENDER_TRACE_UNLIT = 128  /* also trace shadow rays whose Phong weight is exactly zero (the reference does; the
                                    split pipeline skips them by default: they cannot change a pixel) */
};

typedef struct RtRenderParams {
    int32_t flags;
    int32_t semantics;
    /* Multi-GPU screen-space sharding: this process renders tiles t with t % world == rank. */
    int32_t rank, world;
    int32_t tile_w, tile_h;
    /* Host canvas only: number of chunks of tile rows the frame is cut into so that chunk c is copied to the host while
     * chunk c + 1 renders (1..8; chunk sizes fall geometrically so that little is left to copy after the last one).
     * 0 = automatic: 2 from 16 MiB of canvas per rank on when no material spawns secondary rays, else 1. */
    int32_t host_chunks;
} RtRenderParams;

typedef struct RtFrameStats {
    uint64_t rays_total;      /* closest-hit queries = the reference's cast_ray calls */
    uint64_t rays_primary;
    uint64_t rays_shadow;     /* Light::attenuate casts the reference performs for this frame (traced + rays_culled) */
    uint64_t rays_secondary;
    uint64_t nodes_visited;   /* only filled by the stats build of the kernels */
    uint64_t tris_tested;     /* exact (reference-arithmetic) triangle tests */
    uint64_t prims_prefiltered; /* conservative candidate tests */
    float ms_build;           /* CUDA-event time of the BVH build inside this call (0 if none) */
    float ms_trace;           /* CUDA-event time of ray generation + traversal + shading */
    float ms_total;           /* CUDA-event time of everything incl. canvas copy */
    int32_t waves;            /* wavefront iterations (bounce levels) executed */
    int32_t kernel_launches;  /* kernels launched by this call */
    float ms_primary;         /* CUDA-event time of k_primary (ray generation + closest-hit traversal) */
    float ms_shade;           /* CUDA-event time of k_shade (shadow-ray traversal + Phong) */
    float ms_bounce;          /* CUDA-event time of all k_bounce waves (incl. the per-wave read-backs) */
    int32_t n_hits;           /* primary rays that hit something */
    float ms_shadow;          /* CUDA-event time of k_shadow (any-hit traversal of the shadow-ray queue; part of ms_shade) */
    float ms_wave_phase[4];   /* RT_RENDER_STATS: time all bounce waves spent in their four phases (path trace, shadow-ray
                                 generation, shadow traversal, shade + advance), clocked by block 0 of k_bounce_waves */
    uint64_t rays_culled;     /* shadow queries counted in rays_shadow / rays_total that were answered without a
                                 traversal (zero Phong weight, see RT_RENDER_TRACE_UNLIT) */
    int32_t host_chunks;      /* chunks the frame was cut into for the host copy (RtRenderParams.host_chunks); the per-kernel
                                 times above are sums over the chunks. 0 = no copy: the kernels wrote the host canvas directly */
    int32_t tile_kernel;      /* form of wave 0: 0 = split pipeline with the per-lane k_primary, 2 = warp-coherent closest hits
                                 (k_tile<1>) + split shading (default), 1 = tile kernels only (ms_primary is their time,
                                 ms_shade / ms_shadow are 0) */
    uint64_t node_fetches;    /* RT_RENDER_STATS, k_tile: 80-byte node fetches after the warp broadcast (one per warp-level
                                 visit); nodes_visited counts the same visits once per active ray */
    int32_t graph;            /* RT_RENDER_GRAPH: 0 = plain launches, 1 = this frame was captured and launched as a graph, 2 = replayed */
} RtFrameStats;

typedef struct RtBuildStats {
    int32_t n_prims;          /* world-space triangles (instances x mesh triangles) */
    int32_t n_bvh2_nodes;     /* 2*n_prims - 1 */
    int32_t n_bvh8_nodes;
    int32_t kernel_launches;
    float ms_build;
    int32_t n_levels8;        /* depth of the BVH8 (levels of the breadth-first layout). The traversal stack holds one entry
                                 per level; a tree deeper than the stack (32) is refused with RT_ERR_LIMIT at build time */
} RtBuildStats;

typedef struct RtHit {
    int32_t instance; /* renv::Transformation index, -1 on miss */
    int32_t triangle; /* index within the instance's mesh */
    float t;
    float normal[3];
    float color[4];   /* propagate_ray result for rt_debug_cast */
} RtHit;

/* Device selection for this process (one process per GPU). */
int rt_init(int device_ordinal);
int rt_device_count(int* out);

/* Replaces the upload half of SceneBuilder::build_gpu_scene (src/scene_builder.cu:83-179):
 * flat arrays in, no device-side `new`, no managed memory. */
int rt_scene_create(const RtSceneDesc* desc, RtScene** out);
/* Replaces Scene::free (include/rayenv/gpu/scene.h:55-69). */
int rt_scene_destroy(RtScene* scene);

/* Host writes to the managed Scene in the reference (camera motion src/main.cc:140-179,
 * Environment setters include/rayenv/environment.h:36-53). */
int rt_scene_set_camera(RtScene* scene, const RtCamera* cam);
int rt_scene_set_env(RtScene* scene, const RtEnv* env);
int rt_scene_set_instances(RtScene* scene, const RtInstance* instances, int32_t n_instances);
int rt_scene_set_canvas_size(RtScene* scene, int32_t width, int32_t height);

/* Replaces rtracer::gpu::create_bvh + ropt::gpu::BVH::BVH (src/raytracer.cu:76-89,
 * src/rayopt/bvh.cu:74-91) and BVH::free (src/rayopt/bvh.cu:93-96). */
int rt_bvh_build(RtScene* scene, RtBuildStats* stats);
int rt_bvh_free(RtScene* scene);
/* No counterpart in the reference, which rebuilds its BVH from scratch every frame because the scene is meant
 * to be interactive (src/raytracer.cu:102-107, src/main.cc:140-179): after rt_scene_set_instances moved
 * instances (same count, same meshes), keeps the tree built by rt_bvh_build and recomputes the world-space
 * triangles and every BVH2 / BVH8 box. Frames are identical to those of a rebuilt tree (traversal is a
 * conservative filter in front of the reference-exact test). RT_ERR_NO_BVH when there is nothing to refit.
 * rt_render does this on its own, without RT_RENDER_REBUILD_BVH, when instances moved since the last build. */
int rt_bvh_refit(RtScene* scene, RtBuildStats* stats);

/* Replaces rtracer::gpu::update_scene(scene, kernel_dim, optimize) (src/raytracer.cu:102-120):
 * (re)build the BVH when optimize != 0, trace the frame, synchronise, leave the RGBA8 frame in
 * host_canvas (width*height int32). kernel_dim is accepted for compatibility and advisory. */
int rt_update_scene(RtScene* scene, int32_t kernel_dim, int32_t optimize, uint32_t* host_canvas);

/* Full-control render entry (tile sharding, device-resident canvas, stats). */
int rt_render(RtScene* scene, const RtRenderParams* params, uint32_t* canvas, RtFrameStats* stats);

/* Stream-ordered use (replaces the reference's default-stream + cudaDeviceSynchronize per frame,
 * src/raytracer.cu:108-114): rt_scene_set_stream makes every later call of this scene enqueue on the
 * caller's CUDA stream (external != 0; cudaStream_t as void*, where a null handle is the legacy default
 * stream) or on the scene's own stream again (external == 0), so a frame
 * can be ordered against the caller's work (an NCCL all-gather of the tiles, a copy) without host
 * synchronisation. With RT_RENDER_ASYNC, rt_render returns once the frame is enqueued and `stats`
 * is not written; rt_frame_finish waits for that frame and fills the statistics. One frame can be
 * pending per scene; starting another finishes the pending one first. */
int rt_scene_set_stream(RtScene* scene, void* cuda_stream, int32_t external);
int rt_frame_finish(RtScene* scene, RtFrameStats* stats);
/* Concurrent lanes (no counterpart in the reference, whose trace is one kernel, src/raytracer.cu:108-113): a device-resident
 * frame without secondary rays is cut into `lanes` (1..8) contiguous ranges of its screen tiles, and every range runs the
 * frame's kernels on a stream of its own, forked from and joined back into the scene's stream, with its own slice of the
 * queues — the tail of one range's traversal kernel (a few SMs finishing their longest work units) is filled by the next
 * range's kernel. Same kernels, same frame, bit for bit. 0 = the library default (RT_FRAME_LANES in the environment, else 1).
 * Frames that cannot use lanes (host canvas in chunks, secondary rays, graph capture, statistics, the opt-in kernel
 * variants) run as before; RT_DEBUG_FRAME_LANES tells how many the last frame used. */
int rt_scene_set_lanes(RtScene* scene, int32_t lanes);

/* Multi-GPU helper: number of uint32 pixels rt_render writes with RT_RENDER_TILED_OUTPUT, and the
 * de-interleave of `world` gathered slabs back into a row-major canvas (both device pointers). */
int rt_tiled_slab_pixels(RtScene* scene, const RtRenderParams* params, int64_t* out_pixels);
int rt_untile(RtScene* scene, const RtRenderParams* params, const uint32_t* gathered_dev, uint32_t* canvas_dev);

/* Multi-GPU frame exchange over NVLink peer memory (one process per GPU), instead of all-gather +
 * rt_untile: every rank owns a canvas other ranks can map (rt_peer_canvas_alloc returns the device
 * pointer and a 64-byte CUDA IPC handle to send to the peers; rt_peer_canvas_open maps a peer's
 * handle). After rt_render with RT_RENDER_TILED_OUTPUT, rt_push_tiles stores this rank's tiles into
 * the row-major canvases of all `n_peers` ranks (peer_canvases[rank] is the own canvas). A barrier
 * across the ranks on the same stream (any collective) then tells every rank its canvas is complete. */
#define RT_MAX_PEERS 16
int rt_peer_canvas_alloc(int64_t pixels, uint32_t** dev_ptr, unsigned char ipc_handle[64]);
int rt_peer_canvas_free(uint32_t* dev_ptr);
int rt_peer_canvas_open(const unsigned char ipc_handle[64], uint32_t** dev_ptr);
int rt_peer_canvas_close(uint32_t* dev_ptr);
int rt_push_tiles(RtScene* scene, const RtRenderParams* params, const uint32_t* slab_dev, uint32_t* const* peer_canvases,
                  int32_t n_peers);
/* rt_push_tiles followed by the exchange barrier, both over peer memory (no collective call on the frame's path):
 * every canvas from rt_peer_canvas_alloc carries a mailbox of RT_MAX_PEERS words behind its `canvas_pixels` pixels.
 * After its last tile store the rank writes `epoch` (grow it by one per frame, start at 1) into its slot of every
 * rank's mailbox; a second kernel on the same stream then waits until all `n_peers` slots of the rank's OWN mailbox
 * carry the epoch, i.e. until every rank's tiles have landed in this rank's canvas. peer_canvases[rank] must be the
 * rank's own canvas and n_peers == world. The consumers of frame i must be done before any rank pushes frame i + 1. */
int rt_push_tiles_signal(RtScene* scene, const RtRenderParams* params, const uint32_t* slab_dev, uint32_t* const* peer_canvases,
                         int32_t n_peers, int64_t canvas_pixels, uint32_t epoch);
/* The gather fused into the frame (RT_RENDER_PEER_STORE): instead of rendering into a slab that rt_push_tiles then copies,
 * the frame's kernels store every finished pixel into the peer canvases themselves and the last block of the frame signals.
 *   canvases     2 * n_ranks device pointers from rt_peer_canvas_alloc / rt_peer_canvas_open: [parity][rank]. Frame e
 *                (e = 1, 2, ... counted per scene, the same on every rank) uses the set e & 1, so a rank that is still
 *                reading frame e (display, copy — enqueued on the scene's stream after the frame) is never overwritten by
 *                ranks already rendering frame e + 1: they cannot start frame e + 2 before every rank has produced e + 1,
 *                which in stream order is after its consumers of frame e.
 *   gather_root  r >= 0: only rank r receives the frame (north_star's "framebuffer gathered": 4 bytes per pixel cross
 *                NVLink once); -1: every rank receives it (all-gather: every pixel is stored n_ranks times).
 *                With one receiver the senders wait for nobody's signal, so rank r acknowledges every gathered epoch in
 *                each sender's own mailbox (slot r of its parity-0 canvas) and a sender stores frame e only after the
 *                acknowledgement of frame e - 1: it is never more than one frame ahead of the gathering rank, and a
 *                gathering rank that stopped rendering makes the sender's frame RT_ERR_CUDA after the time-out.
 *                The consumer's side of the contract: reads of frame e are enqueued on the scene's stream (or complete)
 *                before rt_render is called for frame e + 1.
 * A receiving rank's rt_render ends with a one-warp wait on its own mailbox (all ranks' epochs), so when the call — or
 * rt_frame_finish after RT_RENDER_ASYNC — returns RT_OK the frame in rt_peer_frame_canvas() is complete; a peer that never
 * signals makes it RT_ERR_CUDA after the time-out instead. Frames with secondary rays or refractive materials render into
 * the rank's own canvas and forward their tiles with one extra kernel. n_ranks == 0 clears the registration. */
int rt_scene_set_peers(RtScene* scene, uint32_t* const* canvases, int32_t n_ranks, int64_t canvas_pixels, int32_t gather_root);
/* The canvas of `rank` that holds (or will hold) the frame of the last rt_render with RT_RENDER_PEER_STORE, and its epoch. */
int rt_peer_frame_canvas(RtScene* scene, int32_t rank, uint32_t** canvas, uint32_t* epoch);
/* Blocking device-to-host copy of `pixels` uint32 on the scene's stream. */
int rt_copy_to_host(RtScene* scene, const uint32_t* dev_ptr, uint32_t* host_ptr, int64_t pixels);
/* The host canvas of RT_RENDER_HOST_BANDS: page-locked memory every rank of a box can map. The reference keeps its
 * canvas in managed memory read by the host after each frame (src/rayenv/canvas.cu:7, src/main.cc:181-208); with one
 * process per GPU the equivalent is one POSIX shared-memory canvas (`shm_name`, e.g. "/rtb200_canvas"; rank 0 passes
 * create = 1, the others open it after a barrier), registered with CUDA in every process so that each GPU copies its
 * bands into it over its own PCIe link. shm_name == NULL gives private pinned memory (single process). */
int rt_host_canvas_open(const char* shm_name, int64_t pixels, int32_t create, uint32_t** host_ptr);
int rt_host_canvas_close(uint32_t* host_ptr, int64_t pixels, const char* shm_name, int32_t unlink_name);

/* Host-only (no GPU needed): the tile map the kernels use. For every pixel (row-major) writes the
 * index rank `rank` stores it at inside its slab, or -1 if another rank owns the pixel. */
int rt_tile_map(int32_t width, int32_t height, int32_t tile_w, int32_t tile_h, int32_t rank, int32_t world,
                int64_t* slab_index, int64_t* slab_pixels);

/* Parity hooks. Primary-ray closest hit per pixel (row-major y*W+x), host outputs. The reference's
 * Isect carries no ids (include/rayprimitives/isect.h:15-24); these are derived the same way the
 * oracle derives them. use_bvh: 1 = BVH8 traversal with the reference's leaf-box rule (the BVH
 * path), 0 = brute force over all instances (`-r`), 2 = BVH8 traversal without the leaf-box rule,
 * 3 = brute force with the leaf-box rule. */
int rt_trace_primary(RtScene* scene, int32_t use_bvh, int32_t* instance, int32_t* triangle, float* t);
/* renv::gpu::cast_ray on caller-provided rays (src/rayenv/scene.cu:42-73). origins/dirs: n*3 floats,
 * directions are normalised like the rmath::Ray constructor does (include/raymath/geometry.h:216). */
int rt_cast_rays(RtScene* scene, int32_t use_bvh, const float* origins, const float* dirs, int32_t n,
                 int32_t* instance, int32_t* triangle, float* t);
/* rt_cast_rays for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it). It has NO counterpart in the reference.
 * Same arguments, same results (bit for bit) as rt_cast_rays; the device staging of the rays and their results lives
 * in a static device array inside the library instead of a grow-only cudaMalloc buffer:
 *   - at most RT_CAST_STATIC_MAX_RAYS rays per call (36 bytes each = 2.25 MiB, held by every GPU that loads the library
 *     until the process exits); a larger n is refused with RT_ERR_LIMIT and a message naming the limit;
 *   - the call allocates and frees nothing and builds nothing: for use_bvh 1 / 2 a BVH must already exist (rt_bvh_build
 *     or a rendered frame), else RT_ERR_NO_BVH; a pending asynchronous frame must be finished first;
 *   - all calls on one GPU share the one array, which keeps the last call's contents (nothing in it has to start at zero:
 *     a call reads only what it wrote). Calls that overlap on the GPU race; the entry point serialises its callers on the
 *     host and synchronises the scene's stream before it returns, so calls in a row are safe. */
#define RT_CAST_STATIC_MAX_RAYS 65536
int rt_cast_rays_static(RtScene* scene, int32_t use_bvh, const float* origins, const float* dirs, int32_t n,
                        int32_t* instance, int32_t* triangle, float* t);
/* Replaces rtracer::gpu::debug_cast (src/raytracer.cu:91-100) without device printf. */
int rt_debug_cast(RtScene* scene, int32_t x, int32_t y, int32_t semantics, RtHit* out);

/* LBVH inspection for the bit-exact host comparison: any pointer may be NULL. Sizes: codes/perm
 * n_prims; parent/left/right/boxes for 2*n_prims-1 BVH2 nodes (boxes: 6 floats each). */
int rt_bvh_download(RtScene* scene, uint64_t* morton, int32_t* perm, int32_t* left, int32_t* right,
                    int32_t* parent, float* boxes, float* scene_bounds6);
/* BVH8: nodes80 = 80 bytes per node, prims48 = 48 bytes per triangle in leaf order, prim_order =
 * sorted (Morton) position of each leaf-order triangle. */
int rt_bvh8_download(RtScene* scene, void* nodes80, int32_t max_nodes, int32_t* n_nodes, void* prims48,
                     int32_t* prim_order, int32_t max_prims, int32_t* n_prims);

/* Introspection for the tests (no reference counterpart): counters of the scene's BVH builder. */
enum {
    RT_DEBUG_PREFIX_UPLOADS = 0, /* builds that had to upload the per-instance triangle prefix (unchanged meshes: once) */
    RT_DEBUG_BVH8_LEVELS = 1,    /* depth of the BVH8 of the last build */
    RT_DEBUG_BUILD_PATH = 2,     /* 1 = multi-kernel pipeline, 2 = one-launch thread-block-cluster build */
    RT_DEBUG_BUILD_LAUNCHES = 3, /* kernels the last build / refit launched */
    RT_DEBUG_GRAPH_REPLAYS = 5,  /* frames of this scene replayed from a captured graph */
    RT_DEBUG_GRAPH_CAPTURES = 6, /* frames captured into a graph */
    RT_DEBUG_FRAME_LANES = 7,    /* concurrent lanes the last frame was cut into (RT_FRAME_LANES; 1 = none) */
    RT_DEBUG_BUILD_CLUSTER = 4,  /* CTAs in the cluster of the last one-launch build (0: multi-kernel pipeline) */
    RT_DEBUG_BUILD_PHASE_NS = 16, /* + i (0..6): ns per phase of the last cluster build, with RT_BUILD_TIMERS=1 in the environment */
    RT_DEBUG_BUILD_LEVEL_NS = 32 /* + i (0..15): raw globaltimer stamp after phase A (even i) / phase B (odd i) of collapse level i / 2 */
};
int rt_debug_get(RtScene* scene, int32_t what, int64_t* out);

/* Measured memory peaks used by bench.py for the L2 roofline (resident-working-set read). */
int rt_measure_l2_bandwidth(int64_t bytes, int32_t iters, float* gbps);

const char* rt_last_error_string(void);
const char* rt_version(void);

#ifdef __cplusplus
}
#endif
#endif
