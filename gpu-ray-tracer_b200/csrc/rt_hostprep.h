// rt_hostprep.h — host-side flattening of an RtSceneDesc into the POD arrays the kernels read
// (the upload half of SceneBuilder::build_gpu_scene, src/scene_builder.cu:83-179, without
// device-side `new`). Header-only; used by rt_engine.cu and by the host-side unit harness.
#ifndef RTB200_HOSTPREP_H
#define RTB200_HOSTPREP_H

#include <cmath>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "rtb200.h"
#include "rt_scene.cuh"

namespace rth {

struct HostScene {
    std::vector<float4> verts, norms;
    std::vector<rtd::TriRec> tris;
    std::vector<rtd::TexRec> texc;   // per triangle record; empty unless the scene has an atlas and a textured triangle
    std::vector<float4> atlas;
    int atlas_w = 0, atlas_h = 0;
    std::vector<rtd::MatRec> mats;
    std::vector<rtd::MeshRec> meshes;
    std::vector<float> mesh_box;  // 8 per mesh: lo[3], hi[3], non_degenerate, pad
    std::vector<rtd::LightRec> lights;
    std::vector<int> mesh_tri_count;
    int any_refractive = 0, any_secondary = 0;
    int cull_ok = 1;  // every colour that can enter a light's `incoming` is finite (rt_shade.cuh light_unlit)
};

inline bool is_identity_rot(const float* q) {
    const float id[4] = {0.f, 0.f, 0.f, 1.f};
    return std::memcmp(q, id, sizeof id) == 0;  // bit pattern +0,+0,+0,1
}
inline rtx::Pose to_pose(const RtPose& p) {
    rtx::Pose o;
    o.q.i = p.rot[0]; o.q.j = p.rot[1]; o.q.k = p.rot[2]; o.q.r = p.rot[3];
    o.p = make_float3(p.pos[0], p.pos[1], p.pos[2]);
    o.identity_rot = is_identity_rot(p.rot) ? 1 : 0;
    return o;
}

inline void prepare_host_scene(const RtSceneDesc* d, HostScene& hs) {
    hs.verts.resize(d->n_vertices);
    hs.norms.resize(d->n_vertices);
    for (int i = 0; i < d->n_vertices; i++) {
        hs.verts[i] = make_float4(d->vertices[3 * i], d->vertices[3 * i + 1], d->vertices[3 * i + 2], 1.f);
        hs.norms[i] = make_float4(d->normals[3 * i], d->normals[3 * i + 1], d->normals[3 * i + 2], 0.f);
    }
    // de-duplicate per-triangle materials (the reference stores a 104-byte Material in every TriInner)
    hs.tris.resize(d->n_triangles);
    std::map<std::string, int> seen;
    for (int i = 0; i < d->n_triangles; i++) {
        const RtMaterial& m = d->triangles[i].mat;
        std::string key((const char*)&m, sizeof m);
        auto it = seen.find(key);
        int id;
        if (it == seen.end()) {
            rtd::MatRec r;
            r.Ke = make_float4(m.Ke[0], m.Ke[1], m.Ke[2], m.Ke[3]); r.Ka = make_float4(m.Ka[0], m.Ka[1], m.Ka[2], m.Ka[3]);
            r.Kd = make_float4(m.Kd[0], m.Kd[1], m.Kd[2], m.Kd[3]); r.Ks = make_float4(m.Ks[0], m.Ks[1], m.Ks[2], m.Ks[3]);
            r.Kt = make_float4(m.Kt[0], m.Kt[1], m.Kt[2], m.Kt[3]); r.Kr = make_float4(m.Kr[0], m.Kr[1], m.Kr[2], m.Kr[3]);
            r.alpha = m.alpha; r.eta = m.eta;
            r.reflective = (m.Kr[0] > 0.f || m.Kr[1] > 0.f || m.Kr[2] > 0.f || m.Kr[3] > 0.f) ? 1 : 0;
            r.refractive = (m.Kt[0] > 0.f || m.Kt[1] > 0.f || m.Kt[2] > 0.f || m.Kt[3] > 0.f) ? 1 : 0;
            for (int k = 0; k < 4; k++)
                if (!std::isfinite(m.Kd[k]) || !std::isfinite(m.Ks[k]) || !(m.Kt[k] <= 1.0f)) hs.cull_ok = 0;
            if (r.refractive) hs.any_refractive = 1;
            if (r.refractive || r.reflective) hs.any_secondary = 1;
            id = (int)hs.mats.size();
            hs.mats.push_back(r);
            seen.emplace(key, id);
        } else {
            id = it->second;
        }
        hs.tris[i].v0 = d->triangles[i].v[0]; hs.tris[i].v1 = d->triangles[i].v[1]; hs.tris[i].v2 = d->triangles[i].v[2];
        hs.tris[i].mat = id;
    }
    // texture atlas + TextureCoords, only when something can be sampled
    bool any_tex = false;
    if (d->atlas && d->atlas_width > 0 && d->atlas_height > 0)
        for (int i = 0; i < d->n_triangles && !any_tex; i++) any_tex = d->triangles[i].tex_degenerate == 0;
    if (any_tex) {
        hs.atlas_w = d->atlas_width; hs.atlas_h = d->atlas_height;
        hs.atlas.resize((size_t)hs.atlas_w * hs.atlas_h);
        for (size_t i = 0; i < hs.atlas.size(); i++) {
            hs.atlas[i] = make_float4(d->atlas[4 * i], d->atlas[4 * i + 1], d->atlas[4 * i + 2], d->atlas[4 * i + 3]);
            for (int k = 0; k < 4; k++)
                if (!std::isfinite(d->atlas[4 * i + k])) hs.cull_ok = 0;
        }
        hs.texc.resize(d->n_triangles);
        for (int i = 0; i < d->n_triangles; i++) {
            const RtTriangle& t = d->triangles[i];
            rtd::TexRec r;
            r.x = t.tex_x; r.y = t.tex_y; r.u0 = t.tex_u[0]; r.u1 = t.tex_u[1]; r.v0 = t.tex_v[0]; r.v1 = t.tex_v[1];
            r.degenerate = t.tex_degenerate ? 1 : 0; r.pad = 0;
            hs.texc[i] = r;
        }
    }
    // meshes: pose + local box over the vertices of their triangles (Trimesh::compute_bounding_box,
    // src/rayprimitives/trimesh.cu:21-32: plain min/max, exact wherever it is evaluated)
    hs.meshes.resize(d->n_meshes);
    hs.mesh_box.assign(8 * (size_t)(d->n_meshes > 0 ? d->n_meshes : 1), 0.f);
    hs.mesh_tri_count.resize(d->n_meshes);
    for (int i = 0; i < d->n_meshes; i++) {
        hs.meshes[i].pose = to_pose(d->meshes[i].pose);
        hs.meshes[i].tri_begin = d->meshes[i].tri_begin;
        hs.meshes[i].tri_end = d->meshes[i].tri_end;
        hs.mesh_tri_count[i] = d->meshes[i].tri_end - d->meshes[i].tri_begin;
        hs.meshes[i].kind = d->meshes[i].kind;
        hs.meshes[i].radius = d->meshes[i].radius;
        float* b = &hs.mesh_box[8 * (size_t)i];
        if (d->meshes[i].kind == RT_HITABLE_SPHERE) {  // Sphere::compute_bounding_box: the local cube [-r, r]^3
            for (int a = 0; a < 3; a++) { b[a] = -d->meshes[i].radius; b[3 + a] = d->meshes[i].radius; }
            b[6] = 1.f;
            continue;
        }
        bool ok = false;
        for (int t = d->meshes[i].tri_begin; t < d->meshes[i].tri_end; t++)
            for (int k = 0; k < 3; k++) {
                const float* v = d->vertices + 3 * (size_t)d->triangles[t].v[k];
                if (!ok) { for (int a = 0; a < 3; a++) b[a] = b[3 + a] = v[a]; ok = true; continue; }
                for (int a = 0; a < 3; a++) { if (v[a] < b[a]) b[a] = v[a]; if (v[a] > b[3 + a]) b[3 + a] = v[a]; }
            }
        b[6] = ok ? 1.f : 0.f;
    }
    hs.lights.resize(d->n_lights);
    for (int i = 0; i < d->n_lights; i++) {
        hs.lights[i].kind = d->lights[i].kind;
        hs.lights[i].color = make_float4(d->lights[i].color[0], d->lights[i].color[1], d->lights[i].color[2], d->lights[i].color[3]);
        hs.lights[i].v = make_float3(d->lights[i].v[0], d->lights[i].v[1], d->lights[i].v[2]);
        for (int k = 0; k < 4; k++)
            if (!std::isfinite(d->lights[i].color[k])) hs.cull_ok = 0;
    }
}

inline void fill_instances(const RtInstance* inst, int n, rtd::InstRec* recs) {
    for (int i = 0; i < n; i++) {
        std::memset(&recs[i], 0, sizeof(rtd::InstRec));
        recs[i].pose = to_pose(inst[i].pose);
        recs[i].mesh = inst[i].mesh;
    }
}

inline std::vector<rtd::InstRec> prepare_instances(const RtInstance* inst, int n) {
    std::vector<rtd::InstRec> recs((size_t)n);
    fill_instances(inst, n, recs.data());
    return recs;
}

}  // namespace rth
#endif
