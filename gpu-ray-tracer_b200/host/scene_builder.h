// scene_builder.h — host-side mirror of the reference's rtracer::SceneBuilder / MeshBuilder API
// (/root/reference/include/scene_builder.h:29-117): same method names and argument meaning, but
// the product of a build is a flat RtSceneDesc (include/rtb200.h) handed to the C-ABI library
// instead of device-side `new`-ed polymorphic objects (src/scene_builder.cu:29-81).
#ifndef RTB200_SCENE_BUILDER_H
#define RTB200_SCENE_BUILDER_H

#include <string>
#include <vector>

#include "hmath.h"
#include "rtb200.h"

namespace rtracer {

// rprimitives::TextureCoords (include/rayprimitives/texture_coords.h:12-29): atlas origin + the atlas-space directions
// of the two barycentric axes. The reference only ever reads `degenerate` (src/rayprimitives/phong.cu:19-23).
struct TextureCoords {
    int texture_x = 0, texture_y = 0;
    float u[2] = {0, 0}, v[2] = {0, 0};
    bool degenerate = true;
    TextureCoords() {}
    TextureCoords(int tx, int ty, float u0, float u1, float v0, float v1)
        : texture_x(tx), texture_y(ty), u{u0, u1}, v{v0, v1}, degenerate(false) {}
};

// rprimitives::Material (include/rayprimitives/material.h:14-115) as a POD with the same defaults
struct Material {
    RtMaterial m;
    Material() {
        for (int i = 0; i < 4; i++) m.Ke[i] = m.Ka[i] = m.Kd[i] = m.Ks[i] = m.Kt[i] = m.Kr[i] = 0.f;
        m.alpha = 0.f;
        m.eta = 1.f;
    }
    void set_Ke(rtb::Float4 v) { put(m.Ke, v); }
    void set_Ka(rtb::Float4 v) { put(m.Ka, v); }
    void set_Kd(rtb::Float4 v) { put(m.Kd, v); }
    void set_Ks(rtb::Float4 v) { put(m.Ks, v); }
    void set_Kt(rtb::Float4 v) { put(m.Kt, v); }
    void set_Kr(rtb::Float4 v) { put(m.Kr, v); }
    void set_alpha(float a) { m.alpha = a; }
    void set_eta(float e) { m.eta = e; }

private:
    static void put(float* d, rtb::Float4 v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w; }
};

class SceneBuilder;

class MeshBuilder {
public:
    void add_triangle(int v0, int v1, int v2, const TextureCoords& coords, const Material& mat) {
        RtTriangle t;
        t.v[0] = v0; t.v[1] = v1; t.v[2] = v2;
        t.tex_degenerate = coords.degenerate ? 1 : 0;
        t.tex_x = coords.texture_x; t.tex_y = coords.texture_y;
        t.tex_u[0] = coords.u[0]; t.tex_u[1] = coords.u[1]; t.tex_v[0] = coords.v[0]; t.tex_v[1] = coords.v[1];
        t.mat = mat.m;
        triangles_.push_back(t);
    }
    int hitable_idx() const { return hitable_idx_; }

private:
    friend class SceneBuilder;
    MeshBuilder(int hi, rtb::Float3 pos, rtb::Quat rot) : hitable_idx_(hi), pos_(pos), rot_(rot) {}
    int hitable_idx_;
    int kind_ = RT_HITABLE_TRIMESH;  // RT_HITABLE_SPHERE: triangles_ holds the one material record, radius_ the radius
    float radius_ = 0.f;
    std::vector<RtTriangle> triangles_;
    rtb::Float3 pos_;
    rtb::Quat rot_;
};

// renv::Transformation (include/rayenv/transformation.h:13-23) + the Entity setters used by
// the loader (include/rayprimitives/entity.h:49-76)
struct Transformation {
    rtb::Quat o = rtb::Quat::identity();
    rtb::Float3 p = {0.f, 0.f, 0.f};
    int hitable_idx = -1;
    void set_position(rtb::Float3 np) { p = np; }
    void set_orientation(rtb::Quat no) { o = no; }
    void translate_global(rtb::Float3 dp) { p = p + dp; }
    void rotate(rtb::Quat dr) { o = dr * o; }
};

// renv::Camera (include/rayenv/camera.h:15-36; ctor src/rayenv/camera.cu:6-9)
struct Camera {
    rtb::Quat o = rtb::Quat::identity();
    rtb::Float3 p = {0.f, 0.f, 0.f};
    float global_near = 0, unit_to_pixels = 0, canvas_width = 0, canvas_height = 0;
    Camera() {}
    Camera(float fov, float unit_to_pixels, int width, int height);
    void set_position(rtb::Float3 np) { p = np; }
    void set_orientation(rtb::Quat no) { o = no; }
    void translate_global(rtb::Float3 dp) { p = p + dp; }
    void rotate(rtb::Quat dr) { o = dr * o; }
    RtCamera to_abi() const;
};

// The flat scene a SceneBuilder produces; owns the arrays an RtSceneDesc points into.
struct FlatScene {
    std::vector<float> vertices, normals;
    std::vector<RtTriangle> triangles;
    std::vector<RtMesh> meshes;
    std::vector<RtInstance> instances;
    std::vector<RtLight> lights;
    std::vector<float> atlas;  // RGBA floats (rprimitives::gpu::Texture), empty = none
    int atlas_width = 0, atlas_height = 0;
    RtCamera camera;
    RtEnv env;
    int width = 0, height = 0;
    RtSceneDesc desc() const;
};

class SceneBuilder {
public:
    explicit SceneBuilder(std::string atlas_path = "") : atlas_path_(atlas_path) {}

    int add_vertex(rtb::Float3 v) {
        int idx = (int)vertices_.size();
        vertices_.push_back(v);
        return idx;
    }
    int add_vertex(float x, float y, float z) { return add_vertex(rtb::Float3{x, y, z}); }
    MeshBuilder& get_mesh_builder(int idx) { return meshes_[idx]; }
    Transformation& get_transformation(int idx) { return trans_[idx]; }
    int n_meshes() const { return (int)meshes_.size(); }
    int n_trans() const { return (int)trans_.size(); }
    int create_mesh(rtb::Float3 pos, rtb::Quat rot) {
        int hi = (int)meshes_.size();
        meshes_.push_back(MeshBuilder(hi, pos, rot));
        return hi;
    }
    int create_mesh() { return create_mesh(rtb::Float3{0.f, 0.f, 0.f}, rtb::Quat::identity()); }
    int add_trans(const MeshBuilder& builder) {
        int idx = (int)trans_.size();
        Transformation t;
        t.hitable_idx = builder.hitable_idx_;
        trans_.push_back(t);
        return idx;
    }
    // 12-triangle, 36-vertex cube (src/scene_builder.cu:181-239)
    int build_cube(float scale, const TextureCoords& coords, const Material& mat);
    // build_cube whose faces each show one square tile of the atlas (texture mapping is a TODO in
    // the reference, src/procedural/cube_world.cc:79-81)
    int build_textured_cube(float scale, int tile_x, int tile_y, int tile_size, const Material& mat);
    // Analytic sphere hitable centred at its own origin (the reference has no second Hitable
    // subclass). Returns the hitable index, to be instanced with add_trans(get_mesh_builder(idx)) like a cube.
    int build_sphere(float radius, const TextureCoords& coords, const Material& mat) {
        int hi = create_mesh();
        MeshBuilder& mb = meshes_[hi];
        mb.kind_ = RT_HITABLE_SPHERE;
        mb.radius_ = radius;
        mb.add_triangle(0, 0, 0, coords, mat);  // the material record; its vertex indices are ignored
        return hi;
    }
    void add_directional_light(rtb::Float3 dir, rtb::Float4 col) {
        dir_light_dir_.push_back(dir);
        dir_light_col_.push_back(col);
    }
    // The atlas the reference reads with assets::read_png (src/assets.cc, libpng — absent here): RGBA floats in
    // [0, 1] (Texture::color_to_vec, src/rayprimitives/texture.cu:14-16), row-major.
    void set_atlas(std::vector<float> rgba, int width, int height) {
        atlas_ = std::move(rgba); atlas_w_ = width; atlas_h_ = height;
    }
    void add_point_light(rtb::Float3 pos, rtb::Float4 col) {
        point_light_pos_.push_back(pos);
        point_light_col_.push_back(col);
    }
    // per-vertex normals: sum of adjacent normalised face normals, renormalised
    // (src/scene_builder.cc:11-29)
    std::vector<rtb::Float3> generate_normals() const;

    // Flatten. The Environment defaults follow renv::Environment's ctor
    // (include/rayenv/environment.h:31-32): dist_atten 0, ambience 0, depth 0.
    FlatScene build_flat_scene(int width, int height, const Camera& camera) const;
    // build_gpu_scene(canvas, camera) equivalent: flatten + rt_scene_create
    // (replaces src/scene_builder.cu:83-179). Returns an RtStatus.
    int build_gpu_scene(int width, int height, const Camera& camera, RtScene** out, FlatScene* keep = nullptr) const;

private:
    std::vector<rtb::Float3> vertices_;
    std::vector<MeshBuilder> meshes_;
    std::vector<rtb::Float3> point_light_pos_, dir_light_dir_;
    std::vector<rtb::Float4> point_light_col_, dir_light_col_;
    std::vector<Transformation> trans_;
    std::string atlas_path_;
    std::vector<float> atlas_;
    int atlas_w_ = 0, atlas_h_ = 0;
};

}  // namespace rtracer

#endif
