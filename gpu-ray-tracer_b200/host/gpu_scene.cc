// gpu_scene.cc — SceneBuilder::build_gpu_scene: flatten, then hand the arrays to the C ABI
// (replaces the upload half of /root/reference/src/scene_builder.cu:83-179). Linked only into
// programs that also link librtb200.so (the CUDA-free loader library leaves it out).
#include "rtb200.h"
#include "scene_builder.h"

namespace rtracer {

int SceneBuilder::build_gpu_scene(int width, int height, const Camera& camera, RtScene** out, FlatScene* keep) const {
    FlatScene local;
    FlatScene& fs = keep ? *keep : local;
    fs = build_flat_scene(width, height, camera);
    RtSceneDesc d = fs.desc();
    return rt_scene_create(&d, out);
}

}  // namespace rtracer
