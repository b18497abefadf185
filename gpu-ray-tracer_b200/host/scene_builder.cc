// scene_builder.cc — flattening half of the SceneBuilder mirror (no CUDA dependency).
#include "scene_builder.h"

#include <cmath>

namespace rtracer {

Camera::Camera(float fov, float unit, int width, int height) {
    // ref: src/rayenv/camera.cu:6-9 — near = 0.5f * W / unit / tan(fov), evaluated in fp32
    canvas_width = (float)width;
    canvas_height = (float)height;
    unit_to_pixels = unit;
    global_near = 0.5f * width / unit / std::tan(fov);
}

RtCamera Camera::to_abi() const {
    RtCamera c;
    c.pose.rot[0] = o.i; c.pose.rot[1] = o.j; c.pose.rot[2] = o.k; c.pose.rot[3] = o.r;
    c.pose.pos[0] = p.x; c.pose.pos[1] = p.y; c.pose.pos[2] = p.z;
    c.global_near = global_near;
    c.unit_to_pixels = unit_to_pixels;
    c.canvas_width = canvas_width;
    c.canvas_height = canvas_height;
    return c;
}

RtSceneDesc FlatScene::desc() const {
    RtSceneDesc d;
    d.vertices = vertices.data();
    d.normals = normals.data();
    d.n_vertices = (int32_t)(vertices.size() / 3);
    d.triangles = triangles.data();
    d.n_triangles = (int32_t)triangles.size();
    d.meshes = meshes.data();
    d.n_meshes = (int32_t)meshes.size();
    d.instances = instances.data();
    d.n_instances = (int32_t)instances.size();
    d.lights = lights.data();
    d.n_lights = (int32_t)lights.size();
    d.camera = camera;
    d.env = env;
    d.width = width;
    d.height = height;
    d.atlas = atlas.empty() ? nullptr : atlas.data();
    d.atlas_width = atlas_width;
    d.atlas_height = atlas_height;
    return d;
}

int SceneBuilder::build_cube(float scale, const TextureCoords& coords, const Material& mat) {
    // Corner naming follows the reference's sketch (src/scene_builder.cu:182-197):
    // a,b,c,d = front face (z = -0.5), e,f,g,h = back face (z = +0.5).
    const rtb::Float3 corner[8] = {
        {-0.5f, 0.5f, -0.5f},  /* a */ {0.5f, 0.5f, -0.5f},  /* b */
        {-0.5f, -0.5f, -0.5f}, /* c */ {0.5f, -0.5f, -0.5f}, /* d */
        {-0.5f, 0.5f, 0.5f},   /* e */ {0.5f, 0.5f, 0.5f},   /* f */
        {-0.5f, -0.5f, 0.5f},  /* g */ {0.5f, -0.5f, 0.5f},  /* h */
    };
    enum { A, B, C, D, E, F, G, H };
    // 12 triangles in the reference's order: front, top, right, left, back, bottom
    static const int faces[12][3] = {
        {D, A, B}, {C, A, D}, {A, E, B}, {E, F, B}, {D, B, H}, {B, F, H},
        {C, G, A}, {A, G, E}, {G, H, E}, {E, H, F}, {G, C, D}, {D, H, G},
    };
    rtb::Float3 scaled[8];
    for (int i = 0; i < 8; i++) scaled[i] = scale * corner[i];
    int builder_idx = create_mesh();
    for (int t = 0; t < 12; t++) {
        // vertices are NOT shared between triangles (36 per cube) => flat face normals
        int v0 = add_vertex(scaled[faces[t][0]]);
        int v1 = add_vertex(scaled[faces[t][1]]);
        int v2 = add_vertex(scaled[faces[t][2]]);
        get_mesh_builder(builder_idx).add_triangle(v0, v1, v2, coords, mat);
    }
    return builder_idx;
}

// build_cube with every face mapped onto one square atlas tile (origin tile_x, tile_y; edge tile_size texels): a
// corner's tile position comes from its two in-face coordinates, TextureCoords = atlas position of the triangle's first
// vertex + the atlas-space edges to the second and third (Isect::uv weighs exactly those edges).
int SceneBuilder::build_textured_cube(float scale, int tile_x, int tile_y, int tile_size, const Material& mat) {
    int idx = build_cube(scale, TextureCoords(), mat);
    MeshBuilder& mb = get_mesh_builder(idx);
    for (RtTriangle& t : mb.triangles_) {
        rtb::Float3 c[3] = {vertices_[t.v[0]], vertices_[t.v[1]], vertices_[t.v[2]]};
        // the face's normal axis: the coordinate all three corners share
        int axis = (c[0].x == c[1].x && c[1].x == c[2].x) ? 0 : ((c[0].y == c[1].y && c[1].y == c[2].y) ? 1 : 2);
        float a[3][2];
        for (int k = 0; k < 3; k++) {
            float p[3] = {c[k].x / scale, c[k].y / scale, c[k].z / scale};  // back to +-0.5
            float s0 = p[(axis + 1) % 3], s1 = p[(axis + 2) % 3];
            a[k][0] = (s0 + 0.5f) * (float)tile_size;
            a[k][1] = (0.5f - s1) * (float)tile_size;
        }
        t.tex_degenerate = 0;
        // integer origin (TextureCoords::texture_x/y are ints): the first vertex sits on a tile corner
        t.tex_x = tile_x + (int)a[0][0];
        t.tex_y = tile_y + (int)a[0][1];
        t.tex_u[0] = a[1][0] - a[0][0]; t.tex_u[1] = a[1][1] - a[0][1];
        t.tex_v[0] = a[2][0] - a[0][0]; t.tex_v[1] = a[2][1] - a[0][1];
    }
    return idx;
}

std::vector<rtb::Float3> SceneBuilder::generate_normals() const {
    std::vector<rtb::Float3> normals(vertices_.size(), rtb::Float3{0.f, 0.f, 0.f});
    for (const MeshBuilder& mesh : meshes_) {
        if (mesh.kind_ != RT_HITABLE_TRIMESH) continue;  // a sphere's one record is its material, not a triangle
        for (const RtTriangle& tri : mesh.triangles_) {
            rtb::Float3 a = vertices_[tri.v[1]] - vertices_[tri.v[0]];
            rtb::Float3 b = vertices_[tri.v[2]] - vertices_[tri.v[0]];
            rtb::Float3 n = rtb::normalized(rtb::cross(a, b));
            for (int k = 0; k < 3; k++) normals[tri.v[k]] = normals[tri.v[k]] + n;
        }
    }
    for (rtb::Float3& n : normals) n = rtb::normalized(n);
    return normals;
}

FlatScene SceneBuilder::build_flat_scene(int width, int height, const Camera& camera) const {
    FlatScene fs;
    fs.width = width;
    fs.height = height;
    fs.camera = camera.to_abi();
    for (int i = 0; i < 4; i++) fs.env.ambience[i] = 0.f;
    for (int i = 0; i < 3; i++) fs.env.dist_atten[i] = 0.f;
    fs.env.depth = 0;

    std::vector<rtb::Float3> normals = generate_normals();
    fs.vertices.reserve(vertices_.size() * 3);
    fs.normals.reserve(vertices_.size() * 3);
    for (size_t i = 0; i < vertices_.size(); i++) {
        fs.vertices.push_back(vertices_[i].x); fs.vertices.push_back(vertices_[i].y); fs.vertices.push_back(vertices_[i].z);
        fs.normals.push_back(normals[i].x); fs.normals.push_back(normals[i].y); fs.normals.push_back(normals[i].z);
    }
    for (const MeshBuilder& mb : meshes_) {
        RtMesh m;
        m.pose.rot[0] = mb.rot_.i; m.pose.rot[1] = mb.rot_.j; m.pose.rot[2] = mb.rot_.k; m.pose.rot[3] = mb.rot_.r;
        m.pose.pos[0] = mb.pos_.x; m.pose.pos[1] = mb.pos_.y; m.pose.pos[2] = mb.pos_.z;
        m.tri_begin = (int32_t)fs.triangles.size();
        fs.triangles.insert(fs.triangles.end(), mb.triangles_.begin(), mb.triangles_.end());
        m.tri_end = (int32_t)fs.triangles.size();
        m.kind = mb.kind_;
        m.radius = mb.radius_;
        fs.meshes.push_back(m);
    }
    fs.atlas = atlas_;
    fs.atlas_width = atlas_w_;
    fs.atlas_height = atlas_h_;
    for (const Transformation& t : trans_) {
        RtInstance inst;
        inst.pose.rot[0] = t.o.i; inst.pose.rot[1] = t.o.j; inst.pose.rot[2] = t.o.k; inst.pose.rot[3] = t.o.r;
        inst.pose.pos[0] = t.p.x; inst.pose.pos[1] = t.p.y; inst.pose.pos[2] = t.p.z;
        inst.mesh = t.hitable_idx;
        fs.instances.push_back(inst);
    }
    // point lights first, then directional (src/scene_builder.cu:61-81, src/scene_builder.cc:56-71)
    for (size_t i = 0; i < point_light_pos_.size(); i++) {
        RtLight l;
        l.kind = RT_LIGHT_POINT;
        l.color[0] = point_light_col_[i].x; l.color[1] = point_light_col_[i].y;
        l.color[2] = point_light_col_[i].z; l.color[3] = point_light_col_[i].w;
        l.v[0] = point_light_pos_[i].x; l.v[1] = point_light_pos_[i].y; l.v[2] = point_light_pos_[i].z;
        fs.lights.push_back(l);
    }
    for (size_t i = 0; i < dir_light_dir_.size(); i++) {
        RtLight l;
        l.kind = RT_LIGHT_DIRECTIONAL;
        l.color[0] = dir_light_col_[i].x; l.color[1] = dir_light_col_[i].y;
        l.color[2] = dir_light_col_[i].z; l.color[3] = dir_light_col_[i].w;
        // DirLight::set_shine_dir normalises (include/rayprimitives/gpu/light.cuh:58-60)
        rtb::Float3 d = rtb::normalized(dir_light_dir_[i]);
        l.v[0] = d.x; l.v[1] = d.y; l.v[2] = d.z;
        fs.lights.push_back(l);
    }
    return fs;
}

}  // namespace rtracer
