#include <cmath>
#include <cstdio>
#include <cstring>
#include <exception>
#include <string>

#include "cube_world.h"
#include "hmath.h"
#include "rtb200_host.h"

struct RthWorld {
    rtracer::FlatScene scene;
    RtSceneDesc desc;
};

static thread_local std::string g_err;

static int load_impl(bool is_text, const char* s, int32_t w, int32_t h, int32_t depth, int32_t grid, RthWorld** out) {
    if (!s || !out) { g_err = "null argument"; return 1; }
    try {
        procedural::WorldOverrides ov;
        ov.width = w; ov.height = h; ov.depth = depth; ov.grid_size = grid;
        RthWorld* world = new RthWorld();
        world->scene = is_text ? procedural::generate_from_text(s, ov) : procedural::generate(s, ov);
        world->desc = world->scene.desc();
        *out = world;
        return 0;
    } catch (const std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

extern "C" {

int rth_world_load(const char* p, int32_t w, int32_t h, int32_t d, int32_t g, RthWorld** out) { return load_impl(false, p, w, h, d, g, out); }
int rth_world_load_text(const char* t, int32_t w, int32_t h, int32_t d, int32_t g, RthWorld** out) { return load_impl(true, t, w, h, d, g, out); }
const RtSceneDesc* rth_world_desc(const RthWorld* w) { return w ? &w->desc : nullptr; }
void rth_world_free(RthWorld* w) { delete w; }
const char* rth_last_error_string(void) { return g_err.c_str(); }

static void pvec(FILE* f, const char* key, const float* v, int n) {
    fprintf(f, "%s %d", key, n);
    for (int i = 0; i < n; i++) { uint32_t u; memcpy(&u, &v[i], 4); fprintf(f, " %08x", u); }
    fprintf(f, "\n");
}

int rth_world_dump(const RthWorld* w, const char* path) {
    if (!w || !path) return 1;
    FILE* f = fopen(path, "w");
    if (!f) return 1;
    const RtSceneDesc& d = w->desc;
    fprintf(f, "canvas %d %d\n", d.width, d.height);
    float camv[11] = {d.camera.pose.rot[0], d.camera.pose.rot[1], d.camera.pose.rot[2], d.camera.pose.rot[3],
                      d.camera.pose.pos[0], d.camera.pose.pos[1], d.camera.pose.pos[2], d.camera.global_near,
                      d.camera.unit_to_pixels, d.camera.canvas_width, d.camera.canvas_height};
    pvec(f, "camera", camv, 11);
    pvec(f, "dist_atten", d.env.dist_atten, 3);
    pvec(f, "ambience", d.env.ambience, 4);
    fprintf(f, "depth %d\n", d.env.depth);
    fprintf(f, "n_vertices %d\n", d.n_vertices);
    for (int i = 0; i < d.n_vertices; i++) {
        float v[6] = {d.vertices[3 * i], d.vertices[3 * i + 1], d.vertices[3 * i + 2],
                      d.normals[3 * i], d.normals[3 * i + 1], d.normals[3 * i + 2]};
        pvec(f, "v", v, 6);
    }
    fprintf(f, "n_meshes %d\n", d.n_meshes);
    for (int m = 0; m < d.n_meshes; m++) {
        const RtMesh& mesh = d.meshes[m];
        float pose[7] = {mesh.pose.rot[0], mesh.pose.rot[1], mesh.pose.rot[2], mesh.pose.rot[3],
                         mesh.pose.pos[0], mesh.pose.pos[1], mesh.pose.pos[2]};
        pvec(f, "mesh_pose", pose, 7);
        fprintf(f, "mesh_tris %d\n", mesh.tri_end - mesh.tri_begin);
        for (int t = mesh.tri_begin; t < mesh.tri_end; t++) {
            const RtTriangle& tri = d.triangles[t];
            fprintf(f, "tri %d %d %d %d\n", tri.v[0], tri.v[1], tri.v[2], tri.tex_degenerate);
            pvec(f, "mat", (const float*)&tri.mat, 26);
        }
    }
    fprintf(f, "n_trans %d\n", d.n_instances);
    for (int i = 0; i < d.n_instances; i++) {
        const RtInstance& t = d.instances[i];
        float tv[7] = {t.pose.rot[0], t.pose.rot[1], t.pose.rot[2], t.pose.rot[3], t.pose.pos[0], t.pose.pos[1], t.pose.pos[2]};
        fprintf(f, "trans_mesh %d\n", t.mesh);
        pvec(f, "trans", tv, 7);
    }
    fprintf(f, "n_lights %d\n", d.n_lights);
    for (int i = 0; i < d.n_lights; i++) {
        const RtLight& l = d.lights[i];
        float lv[7] = {l.color[0], l.color[1], l.color[2], l.color[3], l.v[0], l.v[1], l.v[2]};
        pvec(f, l.kind == RT_LIGHT_POINT ? "point_light" : "dir_light", lv, 7);
    }
    fclose(f);
    return 0;
}

// ---- motion (declared in rtb200_host.h; reference anchors there) ----
static rtb::Quat pose_q(const RtPose* p) { return rtb::Quat{p->rot[0], p->rot[1], p->rot[2], p->rot[3]}; }
static void pose_set_q(RtPose* p, rtb::Quat q) { p->rot[0] = q.i; p->rot[1] = q.j; p->rot[2] = q.k; p->rot[3] = q.r; }

void rth_pose_translate_global(RtPose* pose, const float dp[3]) {
    rtb::Float3 p = rtb::Float3{pose->pos[0], pose->pos[1], pose->pos[2]} + rtb::Float3{dp[0], dp[1], dp[2]};
    pose->pos[0] = p.x; pose->pos[1] = p.y; pose->pos[2] = p.z;
}
void rth_pose_translate(RtPose* pose, const float dp[3]) {
    rtb::Float3 l = pose_q(pose) * rtb::Float3{dp[0], dp[1], dp[2]};  // Entity::vec_to_local = o * v
    const float g[3] = {l.x, l.y, l.z};
    rth_pose_translate_global(pose, g);
}
void rth_pose_rotate(RtPose* pose, const float q[4]) { pose_set_q(pose, rtb::Quat{q[0], q[1], q[2], q[3]} * pose_q(pose)); }
void rth_quat_axis_angle(const float axis[3], float theta, float out[4]) {
    rtb::Quat q = rtb::quat_axis_angle(rtb::Float3{axis[0], axis[1], axis[2]}, theta);
    out[0] = q.i; out[1] = q.j; out[2] = q.k; out[3] = q.r;
}
int rth_camera_key(RtCamera* cam, char key) {
    const float kMoveSpeed = 0.2f;  // MOVE_SPEED, src/main.cc:19
    float dp[3] = {0.f, 0.f, 0.f};
    switch (key) {
        case 'w': dp[2] = kMoveSpeed; break;
        case 's': dp[2] = -kMoveSpeed; break;
        case 'a': dp[0] = -kMoveSpeed; break;
        case 'd': dp[0] = kMoveSpeed; break;
        default: return 1;
    }
    rth_pose_translate(&cam->pose, dp);
    return 0;
}
void rth_camera_mouse(RtCamera* cam, float xrel, float yrel) {
    const float kRotSpeed = 0.01f;  // ROT_SPEED, src/main.cc:20
    // Vec<float, 2>::normalized (linear.h:160-168)
    float s = 0;
    s += xrel * xrel;
    s += yrel * yrel;
    float len = std::sqrt(s);
    float rx = 0.f, ry = 0.f;
    if (len > rtb::kThreshold) { float inv = 1 / len; rx = xrel * inv; ry = yrel * inv; }
    rtb::Basis b = rtb::quat_basis(pose_q(&cam->pose));
    rtb::Quat rot = rtb::quat_axis_angle(b.up, kRotSpeed * rx) * rtb::quat_axis_angle(b.right, kRotSpeed * ry);
    pose_set_q(&cam->pose, rot * pose_q(&cam->pose));
}

}  // extern "C"
