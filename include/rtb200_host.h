/* rtb200_host.h — C ABI of the host-side scene loader (librtb200_host.so, no CUDA dependency).
 *
 * Wraps the C++ mirror of the reference's loader, procedural::{gpu,cpu}::generate
 * (/root/reference/include/procedural/cube_world.h:20-27), for FFI callers (tests, bench.py):
 * world*.json in, a flat RtSceneDesc (include/rtb200.h) out. */
#ifndef RTB200_HOST_H
#define RTB200_HOST_H

#include <stdint.h>
#include "rtb200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct RthWorld RthWorld; /* owns the arrays the RtSceneDesc points into */

/* width/height/grid_size <= 0 and depth < 0 keep the file's values. Returns 0 or non-zero
 * (message via rth_last_error_string). */
int rth_world_load(const char* json_path, int32_t width, int32_t height, int32_t depth, int32_t grid_size,
                   RthWorld** out);
int rth_world_load_text(const char* json_text, int32_t width, int32_t height, int32_t depth, int32_t grid_size,
                        RthWorld** out);
const RtSceneDesc* rth_world_desc(const RthWorld* w);
void rth_world_free(RthWorld* w);
const char* rth_last_error_string(void);
/* Writes the scene in the line format of oracle/ref_build/ref_driver.cu's --scene dump. */
int rth_world_dump(const RthWorld* w, const char* path);

/* Per-frame motion, the reference's Entity / event-loop arithmetic restated on the
 * host so that a moved RtCamera / RtInstance carries the bits the reference's objects would:
 *   rth_pose_translate_global  Entity::translate_global(dp): p += dp      (include/rayprimitives/entity.h:49-52)
 *   rth_pose_translate         Entity::translate(dp): p += o * dp          (entity.h:54-57, geometry.h:176-181)
 *   rth_pose_rotate            Entity::rotate(dr): o = dr * o              (entity.h:64-67)
 *   rth_quat_axis_angle        rmath::Quat(axis, theta), theta in radians  (include/raymath/geometry.h:34-41)
 *   rth_camera_key             the w/s/a/d cases of src/main.cc:140-165 (translate by +-MOVE_SPEED = 0.2 along local z / x)
 *   rth_camera_mouse           SDL_MOUSEMOTION, src/main.cc:166-177 (ROT_SPEED = 0.01)
 * After moving, hand the camera to rt_scene_set_camera and the instances to rt_scene_set_instances. */
void rth_pose_translate_global(RtPose* pose, const float dp[3]);
void rth_pose_translate(RtPose* pose, const float dp[3]);
void rth_pose_rotate(RtPose* pose, const float quat_ijkr[4]);
void rth_quat_axis_angle(const float axis[3], float theta, float quat_ijkr[4]);
int rth_camera_key(RtCamera* cam, char key);
void rth_camera_mouse(RtCamera* cam, float xrel, float yrel);

#ifdef __cplusplus
}
#endif
#endif
