// rt_shade.cuh — shading and the bounce state machine, following the reference's CUDA semantics.
//
//   illuminate / phong / org_light        src/rayprimitives/phong.cu:14-53
//   Light::attenuate / shine / calc_*      src/rayprimitives/light.cu:11-77
//   propagate_ray (RayFrame stack)         src/rayenv/scene.cu:75-188
//   serial propagate_helper (optional)     src/rayenv/scene.cu:222-268
//
// Templated on a Tracer with closest(ray, Hit&) and occluded(ray, max_t) so the same code runs
// over the BVH8 traversal and over the `-r` brute-force path.
#ifndef RTB200_SHADE_CUH
#define RTB200_SHADE_CUH

#include "rt_scene.cuh"

namespace rtd {

struct RayCounts { unsigned primary, shadow, secondary; };

// calc_dist_atten (light.cu:11-17)
RT_HD float dist_atten(const SceneView& s, float dist) {
    float quad = s.dist_c0 + s.dist_c1 * dist + s.dist_c2 * dist * dist;
    return quad < 1.0f ? 1.0f : 1.0f / quad;
}

// The closest-hit loop of Light::attenuate (light.cu:34-60) on its own (k_shadow<LOOP>).
template <class Tracer>
RT_HD float4 light_attenuate_loop(const SceneView& s, const Tracer& tr, const LightRec& L, rtx::Ray cur, float max_t,
                                  RayCounts& rc) {
    float4 rv = L.color;
    for (;;) {
        Hit sh;
        rc.shadow++;
        if (!tr.closest(cur, sh)) return rv;
        if (sh.t > max_t) return rv;
        const MatRec& m = s.mats[s.tris[sh.tri].mat];
        if (!m.refractive) return rtx::mk4(0.f, 0.f, 0.f, 0.f);
        float3 n = hit_normal(s, sh, cur);
        if (rtx::dot3(n, cur.d) > 0.0f) {  // leaving the solid: calc_shadow_atten = pow(Kt, dist)
            rv = rtx::mul4(rv, rtx::mk4(rtx::ref_powf(m.Kt.x, sh.t), rtx::ref_powf(m.Kt.y, sh.t),
                                        rtx::ref_powf(m.Kt.z, sh.t), rtx::ref_powf(m.Kt.w, sh.t)));
        }
        cur = rtx::ray_make(rtx::ray_at(cur, sh.t), cur.d);
        max_t -= sh.t;
    }
}

// Light::attenuate (light.cu:29-61): `cur` is the shadow ray (see shadow_ray()), then a loop of
// CLOSEST-hit casts that walks through refractive occluders. When no material in the scene is
// refractive the loop body runs once and only asks "is the closest hit within max_t", which an
// any-hit query answers exactly, so that case terminates on the first accepted occluder.
// OPAQUE_ONLY (compile time): the caller guarantees !s.any_refractive, so the closest-hit loop is
// not even instantiated (code size and registers).
template <bool OPAQUE_ONLY, class Tracer>
RT_HD float4 light_attenuate(const SceneView& s, const Tracer& tr, const LightRec& L, rtx::Ray cur, float max_t,
                             RayCounts& rc) {
    float4 rv = L.color;
    if (OPAQUE_ONLY || !s.any_refractive) {
        rc.shadow++;
        return tr.occluded(cur, max_t) ? rtx::mk4(0.f, 0.f, 0.f, 0.f) : rv;
    }
    if (OPAQUE_ONLY) return rv;
    return light_attenuate_loop(s, tr, L, cur, max_t, rc);
}

// PointLight::shine / DirLight::shine (light.cu:63-77), the part before attenuate(): direction to
// the light, distance limit and distance attenuation.
RT_HD void light_setup(const SceneView& s, const LightRec& L, float3 hit_pos, float3& dir_to_light, float& max_t, float& da) {
    max_t = INFINITY;
    da = 1.0f;
    if (L.kind == 0) {
        float3 disp = rtx::sub3(L.v, hit_pos);
        float dist = rtx::len3(disp);
        da = dist_atten(s, dist);
        dir_to_light = rtx::normalized3(disp);
        max_t = dist;
    } else {
        dir_to_light = rtx::neg3(L.v);
    }
}
// The Ray handed to attenuate() (Ray ctor normalises) and the shadow ray attenuate() starts from
// it: origin moved THRESHOLD along the direction, direction normalised once more (light.cu:31-33).
RT_HD rtx::Ray shadow_ray(float3 hit_pos, float3 dir_to_light) {
    rtx::Ray to_light = rtx::ray_make(hit_pos, dir_to_light);
    return rtx::ray_make(rtx::ray_at(to_light, RT_THR), to_light.d);
}

// One call site of light_attenuate for both kinds (the traversal is inlined there).
template <bool OPAQUE_ONLY, class Tracer>
RT_HD float4 light_shine(const SceneView& s, const Tracer& tr, const LightRec& L, float3 hit_pos, float3& dir_to_light,
                         RayCounts& rc) {
    float max_t, da;
    light_setup(s, L, hit_pos, dir_to_light, max_t, da);
    float4 a = light_attenuate<OPAQUE_ONLY>(s, tr, L, shadow_ray(hit_pos, dir_to_light), max_t, rc);
    return L.kind == 0 ? rtx::scale4(da, a) : a;
}

// phong (phong.cu:14-33). `kd` is the diffuse colour: the material's Kd, or the atlas texel of a textured triangle
// (diffuse_colour(); the texture branch is an empty TODO in the reference).
RT_HD float4 phong(const MatRec& m, float4 kd, float3 norm, float4 incoming, float3 ray_dir, float3 dir_to_light) {
    float norm_dot = fmaxf(rtx::dot3(dir_to_light, norm), 0.0f);
    float4 diffuse = rtx::scale4(norm_dot, kd);
    float4 specular;
    if (m.Ks.x == 0.0f && m.Ks.y == 0.0f && m.Ks.z == 0.0f && m.Ks.w == 0.0f && m.alpha >= 0.0f) {
        // Ks = +-0 in every channel (all shipped worlds): pow(max(.,0), alpha >= 0) is finite and carries a plus sign, so
        // pow * Ks is Ks itself, bit for bit, and the reflection vector and the double-precision pow (half of k_resolve's
        // instructions) need not be evaluated.
        specular = m.Ks;
    } else {
        float3 reflected = rtx::ref_reflect(rtx::neg3(dir_to_light), norm);
        float reflect_dot = rtx::dot3(rtx::neg3(reflected), ray_dir);
        specular = rtx::scale4(rtx::ref_powf(fmaxf(reflect_dot, 0.0f), m.alpha), m.Ks);
    }
    return rtx::mul4(rtx::add4(diffuse, specular), incoming);
}

// org_light + illuminate (phong.cu:35-53)
template <bool OPAQUE_ONLY = false, class Tracer>
RT_HD float4 illuminate(const SceneView& s, const Tracer& tr, const rtx::Ray& ray, float t, float3 norm,
                        const MatRec& m, float4 kd, RayCounts& rc) {
    float4 sum = rtx::add4(m.Ke, rtx::mul4(m.Ka, s.ambience));
    float3 hit_pos = rtx::ray_at(ray, t);
    for (int i = 0; i < s.n_lights; i++) {
        float3 dir_to_light;
        float4 incoming = light_shine<OPAQUE_ONLY>(s, tr, s.lights[i], hit_pos, dir_to_light, rc);
        sum = rtx::add4(sum, phong(m, kd, norm, incoming, ray.d, dir_to_light));
    }
    return sum;
}

// A light whose Phong weight is exactly zero: max(L.N, 0) == 0 and Ks == 0 (with alpha >= 0 the
// specular power is finite, so specular = pow * 0 = 0). phong() then returns (0 + 0) * incoming = 0
// for every finite `incoming`, lit or shadowed, and sum + 0 == sum: the shadow ray of such a light
// cannot change the pixel and the split pipelines do not trace it (SceneView::cull_unlit, which
// the host only sets when every colour that enters `incoming` is finite).
RT_HD bool light_unlit(const MatRec& m, float3 norm, float3 dir_to_light) {
    return fmaxf(rtx::dot3(dir_to_light, norm), 0.0f) == 0.0f && m.Ks.x == 0.0f && m.Ks.y == 0.0f && m.Ks.z == 0.0f &&
           m.Ks.w == 0.0f && m.alpha >= 0.0f;
}

// illuminate for the split pipelines: the shadow rays of this hit were traced by k_shadow;
// atten(i) is what Light::attenuate returned for light i (the light colour, zero when blocked,
// or the colour filtered by the refractive occluders). Same operations as illuminate().
template <class Atten>
RT_HD float4 illuminate_resolved(const SceneView& s, const rtx::Ray& ray, float t, float3 norm, const MatRec& m,
                                 float4 kd, Atten atten) {
    float4 sum = rtx::add4(m.Ke, rtx::mul4(m.Ka, s.ambience));
    float3 hit_pos = rtx::ray_at(ray, t);
    for (int i = 0; i < s.n_lights; i++) {
        const LightRec& L = s.lights[i];
        float3 dir_to_light;
        float max_t, da;
        light_setup(s, L, hit_pos, dir_to_light, max_t, da);
        if (s.cull_unlit && light_unlit(m, norm, dir_to_light)) continue;  // its shadow ray was not traced
        float4 a = atten(i, L);
        float4 incoming = L.kind == 0 ? rtx::scale4(da, a) : a;
        sum = rtx::add4(sum, phong(m, kd, norm, incoming, ray.d, dir_to_light));
    }
    return sum;
}

// ---- renv::gpu::propagate_ray (scene.cu:75-188) as a resumable machine ----
enum { FR_NORMAL = 0, FR_REFLECT = 1, FR_REFRACT = 2 };
struct Frame {       // RayFrame (scene.cu:75-90); 96 bytes
    rtx::Ray ray;
    float3 hit_pt, norm;
    float4 atten;
    int last_mat;    // material index standing for RayFrame::last_mat
    int type, depth, in_obj;
    int pad0, pad1;
};
#define RT_MAX_FRAMES 10  // MAX_DEPTH, scene.cu:25

struct PathHead {    // 64 bytes in front of a path's frames
    float4 acc;
    int out_index, stack_top, isect_mat;
    int pending;     // 1: `hit` holds the closest hit of the top frame's ray, not yet shaded
    Hit hit;
    int pad0, pad1;
};

RT_HD void frame_init_root(Frame& f, const rtx::Ray& r, int depth) {
    f.ray = r;
    f.hit_pt = rtx::mk3(0.f, 0.f, 0.f);
    f.norm = rtx::mk3(0.f, 0.f, 0.f);
    f.atten = rtx::mk4(1.f, 1.f, 1.f, 1.f);
    f.last_mat = -1;
    f.type = FR_NORMAL;
    f.depth = depth;
    f.in_obj = 0;
    f.pad0 = f.pad1 = 0;
}

// Consumes the result of the cast for the NORMAL frame on top of the stack (the body of
// `case NORMAL`), then runs REFLECT / REFRACT transitions until the next NORMAL frame is on top
// (return true: cast frames[stack_top].ray next) or the stack is empty (return false).
// isect_mat is the material of the most recent successful closest hit of this pixel: the
// reference keeps one Isect for the whole ray tree and reads isect.mat in the REFLECT/REFRACT
// cases after deeper bounces may have overwritten it.
// `frames` may point to local or global memory (one path per thread either way).
// `illum(ray, t, normal, material)` returns illuminate() of the hit: traced in place (path_advance)
// or assembled from shadow rays a previous kernel traced (the split bounce pipeline).
// The frame on top of the stack lives in registers (`top`) while the transitions run and is
// stored only when another frame goes on top of it or the function returns: the stack is in
// global memory in the wavefront kernels, and a store followed by dependent loads of the same
// frame cost k_resolve / k_path_advance most of their time.
// have_top: `top` already holds frames[stack_top] (not yet stored) on entry.
template <class Illum>
RT_HD bool path_advance_core(const SceneView& s, Illum illum, Frame* frames, int& stack_top, int& isect_mat,
                             float4& acc, bool hit, const Hit& h, Frame& top, bool have_top) {
    bool dirty = have_top;
    if (hit) {
        if (!have_top) { top = frames[stack_top]; have_top = true; }
        int mat_id = s.tris[h.tri].mat;
        const MatRec& m = s.mats[mat_id];
        float3 n = hit_normal(s, h, top.ray);
        rtx::Ray ray = top.ray;
        float4 atten = top.atten;
        if (top.depth > 0) {
            if (top.in_obj) {  // trans_atten = pow(time, Kt) (scene.cu:14-22)
                atten = rtx::mul4(atten, rtx::mk4(rtx::ref_powf(h.t, m.Kt.x), rtx::ref_powf(h.t, m.Kt.y),
                                                 rtx::ref_powf(h.t, m.Kt.z), rtx::ref_powf(h.t, m.Kt.w)));
                top.atten = atten;
            }
            top.type = FR_REFLECT;
            top.hit_pt = rtx::ray_at(ray, h.t);
            top.last_mat = mat_id;
            top.norm = n;
            dirty = true;
        } else {
            stack_top--;
            have_top = false;
        }
        isect_mat = mat_id;
        acc = rtx::add4(acc, rtx::mul4(atten, illum(ray, h.t, n, m, diffuse_colour(s, m, h.tri, h.u, h.v))));
    } else {
        stack_top--;
        have_top = false;
    }
    while (stack_top >= 0) {
        if (!have_top) { top = frames[stack_top]; have_top = true; dirty = false; }
        if (top.type == FR_NORMAL) {
            if (dirty) frames[stack_top] = top;
            return true;
        }
        const MatRec& im = s.mats[isect_mat];
        if (top.type == FR_REFLECT) {
            top.type = FR_REFRACT;
            dirty = true;
            if (im.reflective) {
                Frame nt;
                nt.type = FR_NORMAL;
                nt.last_mat = top.last_mat;
                nt.in_obj = top.in_obj;
                nt.atten = rtx::mul4(top.atten, im.Kr);
                nt.depth = top.depth - 1;
                nt.ray = rtx::ray_make(top.hit_pt, rtx::ref_reflect(top.ray.d, rtx::normalized3(top.norm)));
                nt.hit_pt = rtx::mk3(0.f, 0.f, 0.f);
                nt.norm = rtx::mk3(0.f, 0.f, 0.f);
                nt.pad0 = nt.pad1 = 0;
                frames[stack_top] = top;  // the parent stays below the new frame
                stack_top++;
                top = nt;
            }
        } else {  // FR_REFRACT
            if (im.refractive) {
                top.type = FR_NORMAL;
                dirty = true;
                float eta = s.mats[top.last_mat].eta;
                float n1, n2;
                if (top.in_obj) { n1 = eta; n2 = 1.0f; } else { n1 = 1.0f; n2 = eta; }
                bool tir;
                float3 rd = rtx::ref_refract(top.ray.d, rtx::normalized3(top.norm), n1, n2, tir);
                if (tir) {
                    stack_top--;
                    have_top = false;
                } else {
                    top.ray = rtx::ray_make(top.hit_pt, rd);
                    top.in_obj = !top.in_obj;
                    top.depth--;
                }
            } else {
                stack_top--;
                have_top = false;
            }
        }
    }
    return false;
}

template <class Illum>
RT_HD bool path_advance_with(const SceneView& s, Illum illum, Frame* frames, int& stack_top, int& isect_mat,
                             float4& acc, bool hit, const Hit& h) {
    Frame top;
    return path_advance_core(s, illum, frames, stack_top, isect_mat, acc, hit, h, top, false);
}

template <class Tracer>
RT_HD bool path_advance(const SceneView& s, const Tracer& tr, Frame* frames, int& stack_top, int& isect_mat,
                        float4& acc, bool hit, const Hit& h, RayCounts& rc) {
    return path_advance_with(
        s, [&](const rtx::Ray& ray, float t, float3 n, const MatRec& m, float4 kd) { return illuminate(s, tr, ray, t, n, m, kd, rc); },
        frames, stack_top, isect_mat, acc, hit, h);
}

// Whole pixel in one thread (rt_debug_cast and the host-side harness): propagate_ray verbatim.
template <class Tracer>
RT_HD float4 propagate_gpu(const SceneView& s, const Tracer& tr, const rtx::Ray& r, RayCounts& rc) {
    Frame frames[RT_MAX_FRAMES];
    frame_init_root(frames[0], r, s.depth);
    int stack_top = 0, isect_mat = 0;
    float4 acc = rtx::mk4(0.f, 0.f, 0.f, 0.f);
    bool first = true;
    for (;;) {
        Hit h;
        if (first) { rc.primary++; first = false; } else rc.secondary++;
        bool hit = tr.closest(frames[stack_top].ray, h);
        if (!path_advance(s, tr, frames, stack_top, isect_mat, acc, hit, h, rc)) break;
    }
    return acc;
}

// renv::cpu::Scene::propagate_helper (scene.cu:222-268): the serial path's bounce logic (Kr never
// applied, in_obj never toggles, the depth guard never fires, stale Isect) — RT_SEMANTICS_CPU.
// Recursion unrolled by hand is not possible (two children share one Isect), so this walks the
// tree with an explicit stack of continuation records.
struct CpuIsect { float time; float3 norm; int mat; };
// The reference's serial recursion is bounded only by misses; rays caught
// in the 0.001-wide gap between two reflective cubes recurse thousands of levels there, adding
// light to pixels that are already clamped to 1. Device recursion is cut here.
#ifndef RT_CPU_SEM_MAX_LEVEL
#define RT_CPU_SEM_MAX_LEVEL 16
#endif

template <class Tracer>
__host__ __device__ float4 propagate_cpu_rec(const SceneView& s, const Tracer& tr, const rtx::Ray& r, CpuIsect& isect,
                               float& last_time, int in_obj, int depth, int level, RayCounts& rc) {
    float4 zero = rtx::mk4(0.f, 0.f, 0.f, 0.f);
    if (depth > s.depth) return zero;
    if (level > RT_CPU_SEM_MAX_LEVEL) return zero;
    isect.time = INFINITY;
    if (level == 0) rc.primary++; else rc.secondary++;
    Hit h;
    if (!tr.closest(r, h)) return zero;
    isect.time = h.t;
    isect.norm = hit_normal(s, h, r);
    isect.mat = s.tris[h.tri].mat;
    float4 acc = illuminate(s, tr, r, isect.time, isect.norm, s.mats[isect.mat],
                            diffuse_colour(s, s.mats[isect.mat], h.tri, h.u, h.v), rc);
    last_time = isect.time;
    float next_time = 0.0f;
    if (s.mats[isect.mat].reflective) {
        float3 rd = rtx::ref_reflect(r.d, isect.norm);
        rtx::Ray rr = rtx::ray_make(rtx::ray_at(r, isect.time), rd);
        acc = rtx::add4(acc, propagate_cpu_rec(s, tr, rr, isect, next_time, in_obj, depth - 1, level + 1, rc));
    }
    if (s.mats[isect.mat].refractive) {
        const MatRec& m = s.mats[isect.mat];
        float n1, n2;
        if (in_obj) { n1 = m.eta; n2 = 1.0f; } else { n1 = 1.0f; n2 = m.eta; }
        bool tir;
        float3 rd = rtx::ref_refract(r.d, isect.norm, n1, n2, tir);
        if (!tir) {
            next_time = 0.0f;
            rtx::Ray rr = rtx::ray_make(rtx::ray_at(r, isect.time), rd);
            float4 un = propagate_cpu_rec(s, tr, rr, isect, next_time, in_obj, depth - 1, level + 1, rc);
            const MatRec& m2 = s.mats[isect.mat];
            float4 ta = rtx::mk4(rtx::ref_powf(next_time, m2.Kt.x), rtx::ref_powf(next_time, m2.Kt.y),
                                 rtx::ref_powf(next_time, m2.Kt.z), rtx::ref_powf(next_time, m2.Kt.w));
            acc = rtx::add4(acc, rtx::mul4(ta, un));
        }
    }
    return acc;
}

}  // namespace rtd
#endif
