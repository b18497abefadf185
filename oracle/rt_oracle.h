/* rt_oracle.h — CPU ORACLE. TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the reference's hot path (closest hit, shading, bounce logic, camera,
 * colour packing), each function citing the /root/reference file:line it follows. Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this; the product library
 * (librtb200.so) never links or calls it.
 *
 * Parity pin: validated bit-for-bit (canvas hashes, per-pixel ids and t) against the UNMODIFIED
 * reference compiled from its own sources (oracle/_ref/ref_cpu, recipe in oracle/ref_build/) —
 * see tests/test_oracle_vs_ref.py and tests/golden/. Compile with -ffp-contract=off: the
 * reference's host build has no FMA contraction and every + - * / sqrt below is a single
 * correctly rounded fp32 operation, in the reference's order of evaluation.
 */
#ifndef RT_ORACLE_H
#define RT_ORACLE_H

#include <stdint.h>
#include "rtb200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct OrScene OrScene;

typedef struct OrStats {
    uint64_t rays_total;     /* cast_ray calls */
    uint64_t rays_primary;
    uint64_t rays_shadow;
    uint64_t rays_secondary;
    uint64_t box_tests;      /* leaf-box tests (one per instance per ray with use_bvh) */
    uint64_t leaf_tests;     /* cast_local calls */
    uint64_t tri_tests;
} OrStats;

OrScene* or_scene_create(const RtSceneDesc* desc);
void or_scene_destroy(OrScene* s);
void or_scene_set_camera(OrScene* s, const RtCamera* cam);
void or_scene_set_env(OrScene* s, const RtEnv* env);
void or_scene_set_canvas_size(OrScene* s, int32_t w, int32_t h);

/* World AABB per instance as rtracer::cpu::create_bvh computes it (6 floats: min xyz, max xyz). */
void or_instance_boxes(const OrScene* s, float* boxes6);

/* Renders rows [y0, y1) of the frame into canvas (full W*H buffer, row-major). semantics:
 * RT_SEMANTICS_GPU / RT_SEMANTICS_CPU. use_bvh: 1 = leaf-AABB-filtered (BVH path), 0 = brute force.
 * threads <= 1: scalar, else rows are split over pthreads (pixels are independent). */
void or_render(const OrScene* s, int semantics, int use_bvh, int y0, int y1, uint32_t* canvas, OrStats* stats,
               int threads);

void or_trace_primary(const OrScene* s, int use_bvh, int32_t* instance, int32_t* triangle, float* t, int threads);
void or_cast_rays(const OrScene* s, int use_bvh, const float* origins, const float* dirs, int32_t n,
                  int32_t* instance, int32_t* triangle, float* t);
/* propagate_ray for one pixel: colour (unclamped RGBA), plus the primary hit record. */
void or_debug_cast(const OrScene* s, int semantics, int use_bvh, int x, int y, RtHit* out);
void or_primary_ray(const OrScene* s, int x, int y, float* origin3, float* dir3);

uint64_t or_canvas_hash(const uint32_t* canvas, int64_t n_pixels, int32_t* nonzero);

#ifdef __cplusplus
}
#endif
#endif
