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

#ifdef __cplusplus
}
#endif
#endif
