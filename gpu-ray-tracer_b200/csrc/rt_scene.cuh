// rt_scene.cuh — device-side view of a scene (flat POD arrays, no vtables, no managed memory)
// and the reference-exact candidate test.
//
// Layout in HBM (all arrays owned by RtScene, rt_engine.cu):
//   verts/norms  float4[n_vertices]   mesh-local vertices / generated normals (w unused)
//   tris         TriRec[n_triangles]  vertex ids + material id          (TriInner, trimesh.cuh:25-53)
//   meshes       MeshRec[n_meshes]    pose + triangle range             (Trimesh + Entity); or pose + radius + one
//                                     material record for an analytic sphere hitable
//   insts        InstRec[n_instances] pose + mesh + reference leaf box  (Transformation + create_boxes)
//   mats         MatRec[n_materials]  de-duplicated rprimitives::Material
//   lights       LightRec[n_lights]
//   nodes/prims  compressed BVH8 + world-space triangles in leaf order (rt_bvh.cuh)
#ifndef RTB200_SCENE_CUH
#define RTB200_SCENE_CUH

#include "rt_exact.cuh"
#include "rtb200.h"

namespace rtd {

struct TriRec { int v0, v1, v2, mat; };
struct TexRec { int x, y; float u0, u1, v0, v1; int degenerate, pad; };  // TextureCoords; only present in scenes with an atlas
struct MeshRec { rtx::Pose pose; int tri_begin, tri_end; int kind; float radius; };  // kind: RT_HITABLE_TRIMESH / RT_HITABLE_SPHERE
struct InstRec {
    rtx::Pose pose;
    int mesh;
    int box_ok;      // BoundingBox::non_degenerate
    float box[6];    // world AABB exactly as rtracer::gpu::create_boxes computes it (src/raytracer.cu:54-74)
};
struct MatRec {
    float4 Ke, Ka, Kd, Ks, Kt, Kr;
    float alpha, eta;
    int reflective, refractive;  // Material::reflective()/refractive() (material.h:103-111)
};
struct LightRec { int kind; float4 color; float3 v; };

struct Hit {
    float t;       // world time after both rescalings (cast_local / Hitable::hit)
    float t_local; // time in mesh-local units, what Trimesh::hit_local compares inside one instance
    float u, v;
    int inst;      // renv::Transformation index, -1 = miss
    int tri;       // global triangle index
};

struct SceneView {
    const float4* verts;
    const float4* norms;
    const TriRec* tris;
    const MeshRec* meshes;
    const InstRec* insts;
    const MatRec* mats;
    const LightRec* lights;
    const float4* tri_pre; // per triangle record: unit plane normal + area (rtx::ref_triangle_pre); null = computed per candidate
    const TexRec* texc;    // per triangle record, null when the scene has no atlas or no textured triangle
    const float4* atlas;   // RGBA texels, row-major
    int atlas_w, atlas_h;
    int n_insts, n_lights, n_prims;
    // BVH8 (null when no BVH is built)
    const uint4* nodes;
    const float4* prims;
    // environment
    rtx::CameraBasis cam;
    float4 ambience;
    float dist_c0, dist_c1, dist_c2;
    int depth;
    int width, height;
    int any_refractive;  // some material has Kt > 0: shadow rays need the closest-hit loop
    int any_secondary;   // some material has Kr > 0 or Kt > 0
    int cull_unlit;      // skip shadow rays of lights with zero Phong weight (rt_shade.cuh light_unlit)
    unsigned one_bits;   // 0x3f800000, kept opaque to the compiler (rt_bvh.cuh qfloat)
    int fast_accept;     // host-side only: the frame runs the tolerance-mode kernel variants (rt_bvh.cuh tri_filter_clear)
};

// create_boxes (src/raytracer.cu:54-74): world box of an instance = from_local(from_local(mesh box,
// mesh pose), instance pose). BoundingBox::from_local (src/rayopt/bounding_box.cu:51-60) refits
// only the transformed min and max corners — reproduced as is, because the reference culls
// instances by exactly this box. mesh_box: 8 floats per mesh = lo[3], hi[3], non_degenerate, pad.
RT_HD void box_fit(float* b, int& ok, float3 v) {  // BoundingBox::fit_vertex (bounding_box.cu:6-22)
    if (!ok) { b[0] = b[3] = v.x; b[1] = b[4] = v.y; b[2] = b[5] = v.z; ok = 1; return; }
    if (v.x < b[0]) b[0] = v.x;
    if (v.x > b[3]) b[3] = v.x;
    if (v.y < b[1]) b[1] = v.y;
    if (v.y > b[4]) b[4] = v.y;
    if (v.z < b[2]) b[2] = v.z;
    if (v.z > b[5]) b[5] = v.z;
}
RT_HD void instance_box(InstRec& I, const MeshRec* meshes, const float* mesh_box) {
    const float* mb = mesh_box + 8 * (size_t)I.mesh;
    float b1[6] = {0, 0, 0, 0, 0, 0}, b2[6] = {0, 0, 0, 0, 0, 0};
    int ok1 = 0, ok2 = 0;
    if (meshes[I.mesh].kind == RT_HITABLE_SPHERE) {
        // a sphere's box does not depend on the rotations: world centre +- radius (the reference's two-corner refit would
        // cut a rotated sphere)
        const float3 c = rtx::point_from_local(I.pose, rtx::point_from_local(meshes[I.mesh].pose, rtx::mk3(0.f, 0.f, 0.f)));
        const float rad = meshes[I.mesh].radius;
        b2[0] = c.x - rad; b2[1] = c.y - rad; b2[2] = c.z - rad; b2[3] = c.x + rad; b2[4] = c.y + rad; b2[5] = c.z + rad;
        ok2 = 1;
    } else if (mb[6] != 0.0f) {
        const rtx::Pose mp = meshes[I.mesh].pose;
        box_fit(b1, ok1, rtx::point_from_local(mp, rtx::mk3(mb[0], mb[1], mb[2])));
        box_fit(b1, ok1, rtx::point_from_local(mp, rtx::mk3(mb[3], mb[4], mb[5])));
        box_fit(b2, ok2, rtx::point_from_local(I.pose, rtx::mk3(b1[0], b1[1], b1[2])));
        box_fit(b2, ok2, rtx::point_from_local(I.pose, rtx::mk3(b1[3], b1[4], b1[5])));
    }
    for (int k = 0; k < 6; k++) I.box[k] = b2[k];
    I.box_ok = ok2;
}

RT_HD void hit_reset(Hit& h) {
    h.t = INFINITY; h.t_local = INFINITY; h.u = 0.f; h.v = 0.f; h.inst = -1; h.tri = -1;
}

// Ray in the frame of one instance's mesh: cast_local (src/rayenv/scene.cu:27-40) followed by
// Hitable::hit's HitHandle (src/rayprimitives/hitable.cu:7-38). Cached per instance so that
// consecutive candidates of the same instance share it.
struct InstCtx {
    int inst;
    rtx::Ray lray;     // mesh-local ray
    float scale;       // |mesh-local direction| (HitHandle::scale)
    float dir_len;     // |instance-local direction| (cast_local)
    int box_pass;      // BoundingBox::intersects on the instance's leaf box
    float radius;      // >= 0: the instance's hitable is a sphere of this radius; < 0: a triangle mesh
};

// A ray into the frame of one Entity: the direction through vec_to_local (its length is what cast_local / HitHandle
// rescale the hit time by), the origin through point_to_local, the Ray constructor's normalisation. One out-of-line copy
// serves both levels (instance, then mesh).
RT_HD_NOINLINE float ray_into_pose(const rtx::Pose& pose, const rtx::Ray& in, rtx::Ray& out) {
    float3 local_dir = rtx::vec_to_local(pose, in.d);
    float len = rtx::len3(local_dir);
    out = rtx::ray_make(rtx::point_to_local(pose, in.o), local_dir);
    return len;
}
RT_HD_NOINLINE void inst_ctx_load(const SceneView& s, const rtx::Ray& r, int inst, bool use_box, InstCtx& c) {
    const InstRec& I = s.insts[inst];
    c.inst = inst;
    c.box_pass = use_box ? (rtx::ref_box_intersects(I.box, I.box_ok != 0, r) ? 1 : 0) : 1;
    if (!c.box_pass) return;
    rtx::Ray l1;
    c.dir_len = ray_into_pose(I.pose, r, l1);
    const MeshRec& M = s.meshes[I.mesh];
    c.scale = ray_into_pose(M.pose, l1, c.lray);
    c.radius = M.kind == RT_HITABLE_SPHERE ? M.radius : -1.0f;
}

// Sphere::hit_local — the analytic sphere hitable (no reference counterpart; written in the reference's style: unit
// local direction, rmath::dot's summation order, THRESHOLD comparisons). Centre at the local origin.
//   b = o.d, c = o.o - r^2, disc = b^2 - c; roots -b -+ sqrt(disc); the near root unless it is below THRESHOLD
//   (origin inside the sphere or on its surface), then the far one.
RT_HD bool ref_sphere_hit(float radius, const rtx::Ray& lray, float& time) {
    float b = rtx::dot3(lray.o, lray.d);
    float c = rtx::dot3(lray.o, lray.o) - radius * radius;
    float disc = b * b - c;
    if (!(disc >= 0.0f)) return false;
    float sq = sqrtf(disc);
    float t = -b - sq;
    if (!(t >= RT_THR)) t = -b + sq;
    time = t;
    return true;
}

// One (instance, triangle) candidate through the reference's own arithmetic: TriInner::tri_hit
// (src/rayprimitives/trimesh.cu:46-68) with `time >= THRESHOLD && time < isect.time`. Within one
// instance times are compared in mesh-local units like Trimesh::hit_local does; across instances
// after rescaling. Exact ties go to the lower (instance, triangle) index, i.e. the one the
// reference's index-ordered loops reach first.
RT_HD_NOINLINE bool exact_candidate(const SceneView& s, const rtx::Ray& r, int inst, int tri, bool use_box, InstCtx& ctx,
                           Hit& best) {
    if (ctx.inst != inst) inst_ctx_load(s, r, inst, use_box, ctx);
    if (!ctx.box_pass) return false;
    float time, u, v;
    if (ctx.radius >= 0.0f) {
        if (!ref_sphere_hit(ctx.radius, ctx.lray, time)) return false;
        u = v = 0.0f;
    } else {
        const TriRec T = s.tris[tri];
        float4 a4 = s.verts[T.v0], b4 = s.verts[T.v1], c4 = s.verts[T.v2];
        if (s.tri_pre) {
            const float4 pn = s.tri_pre[tri];
            if (!rtx::ref_triangle_hit_pre(rtx::mk3(a4.x, a4.y, a4.z), rtx::mk3(b4.x, b4.y, b4.z), rtx::mk3(c4.x, c4.y, c4.z),
                                           rtx::mk3(pn.x, pn.y, pn.z), pn.w, ctx.lray, time, u, v))
                return false;
        } else if (!rtx::ref_triangle_hit(rtx::mk3(a4.x, a4.y, a4.z), rtx::mk3(b4.x, b4.y, b4.z), rtx::mk3(c4.x, c4.y, c4.z),
                                          ctx.lray, time, u, v)) {
            return false;
        }
    }
    if (!(time >= RT_THR)) return false;
    float t_world = (time * ctx.scale) * ctx.dir_len;
    bool better;
    if (best.inst == inst) better = time < best.t_local || (time == best.t_local && tri < best.tri);
    else better = t_world < best.t || (t_world == best.t && inst < best.inst);
    if (!better) return false;
    best.t = t_world; best.t_local = time; best.u = u; best.v = v; best.inst = inst; best.tri = tri;
    return true;
}

// Tolerance mode (SceneView::fast_accept): a clear hit of the conservative filter, taken with the filter's own t / uv instead of
// the reference-exact test. World-space t stands in for the mesh-local time too (the transforms are rigid: they differ by
// rounding). Same ordering rules as exact_candidate.
RT_HD bool fast_candidate(int inst, int tri, float t, float u, float v, Hit& best) {
    bool better;
    if (best.inst == inst) better = t < best.t_local || (t == best.t_local && tri < best.tri);
    else better = t < best.t || (t == best.t && inst < best.inst);
    if (!better) return false;
    best.t = t; best.t_local = t; best.u = u; best.v = v; best.inst = inst; best.tri = tri;
    return true;
}

// Shading normal of a recorded hit: TriInner::tri_hit's interpolation (trimesh.cu:57-63), then
// back through the mesh pose (HitHandle dtor, hitable.cu:20-26) and the instance pose
// (cast_local, scene.cu:35-38). The reference recomputes it on every accepted candidate; only
// the last one survives, so it is evaluated once per closest hit here.
// `r` is the ray the hit was found with: a sphere's normal is the local hit point over the radius direction,
// normalized(lray.at(t_local)), which needs the local ray again.
RT_HD_NOINLINE float3 hit_normal(const SceneView& s, const Hit& h, const rtx::Ray& r) {
    const InstRec& I = s.insts[h.inst];
    const MeshRec& M = s.meshes[I.mesh];
    float3 n;
    if (M.kind == RT_HITABLE_SPHERE) {
        InstCtx c;
        inst_ctx_load(s, r, h.inst, false, c);
        n = rtx::normalized3(rtx::ray_at(c.lray, h.t_local));
    } else {
        const TriRec T = s.tris[h.tri];
        float4 n0 = s.norms[T.v0], n1 = s.norms[T.v1], n2 = s.norms[T.v2];
        float bary0 = 1.0f - h.u - h.v;
        n = rtx::normalized3(rtx::add3(rtx::add3(rtx::scale3(bary0, rtx::mk3(n0.x, n0.y, n0.z)),
                                                  rtx::scale3(h.u, rtx::mk3(n1.x, n1.y, n1.z))),
                                        rtx::scale3(h.v, rtx::mk3(n2.x, n2.y, n2.z))));
    }
    n = rtx::normalized3(rtx::vec_from_local(M.pose, n));
    return rtx::vec_from_local(I.pose, n);
}

// The diffuse colour phong() uses for a hit: the material's Kd, or — triangle with non-degenerate TextureCoords in a
// scene with an atlas — the texel at origin + b1 * u + b2 * v (point sampling, clamped, unnormalised coordinates like
// the reference's texture descriptor, include/gputils/alloc.h:52-56). No reference counterpart (phong.cu:19-23 is a TODO).
RT_HD float4 diffuse_colour(const SceneView& s, const MatRec& m, int tri, float u, float v) {
    if (!s.texc) return m.Kd;
    const TexRec t = s.texc[tri];
    if (t.degenerate) return m.Kd;
    float tx = ((float)t.x + u * t.u0) + v * t.v0;
    float ty = ((float)t.y + u * t.u1) + v * t.v1;
    int ix = (int)floorf(tx), iy = (int)floorf(ty);
    ix = ix < 0 ? 0 : (ix >= s.atlas_w ? s.atlas_w - 1 : ix);
    iy = iy < 0 ? 0 : (iy >= s.atlas_h ? s.atlas_h - 1 : iy);
    return s.atlas[(size_t)iy * s.atlas_w + ix];
}

// cast_ray's `-r/--unoptimize` branch (src/rayenv/scene.cu:48-52): every instance in index order.
// With use_box the instance's leaf box must pass BoundingBox::intersects first, which is what the
// reference's BVH walk reduces to (an inner box is the union of its leaves).
struct BruteTracer {
    const SceneView* s;
    bool use_box;
    RT_HD bool closest(const rtx::Ray& r, Hit& best) const {
        hit_reset(best);
        InstCtx ctx;
        ctx.inst = -1;
        for (int i = 0; i < s->n_insts; i++) {
            inst_ctx_load(*s, r, i, use_box, ctx);
            if (!ctx.box_pass) continue;
            const MeshRec& M = s->meshes[s->insts[i].mesh];
            for (int k = M.tri_begin; k < M.tri_end; k++) exact_candidate(*s, r, i, k, use_box, ctx, best);
        }
        return best.inst >= 0;
    }
    RT_HD bool occluded(const rtx::Ray& r, float max_t) const {
        Hit h;
        return closest(r, h) && !(h.t > max_t);
    }
};

}  // namespace rtd
#endif
