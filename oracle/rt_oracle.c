/* rt_oracle.c — CPU ORACLE. TEST INFRASTRUCTURE ONLY (see rt_oracle.h for the contract).
 *
 * Every function restates one reference function; "ref:" comments give /root/reference paths.
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -fPIC -shared -pthread (oracle/Makefile).
 */
#include "rt_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ref: include/raymath/linear.h:15 — `constexpr double THRESHOLD = 1E-5f;` every comparison
 * against it promotes the fp32 operand to double; both sides are exactly representable in fp32,
 * so comparing against the fp32 constant is identical. */
#define THR 1E-5f

typedef struct { float x, y, z; } V3;
typedef struct { float c[4]; } V4;
typedef struct { float i, j, k, r; } Q4;
typedef struct { V3 o, d; } Ray;

/* ---- rmath::Vec (include/raymath/linear.h:24-185) ---- */
static V3 v3(float x, float y, float z) { V3 v = {x, y, z}; return v; }
static V3 v3add(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static V3 v3sub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static V3 v3scale(float c, V3 v) { return v3(v.x * c, v.y * c, v.z * c); } /* operator*(T coef, Vec): coords[i] *= coef */
static V3 v3neg(V3 v) { return v3scale(-1.0f, v); }                        /* operator-: (T)-1 * vec */
/* ref: linear.h:197-204 dot — sum starts at 0 and accumulates in index order */
static float v3dot(V3 a, V3 b) {
    float s = 0;
    s += a.x * b.x;
    s += a.y * b.y;
    s += a.z * b.z;
    return s;
}
static float v3len(V3 v) { return sqrtf(v3dot(v, v)); }
/* ref: linear.h:160-168 normalized — zero vector at or below the threshold */
static V3 v3normalized(V3 v) {
    float len = v3len(v);
    if (len > THR) return v3scale(1 / len, v);
    return v3(0, 0, 0);
}
/* ref: linear.h:206-213 cross */
static V3 v3cross(V3 a, V3 b) {
    float x = a.y * b.z - a.z * b.y;
    float y = a.z * b.x - a.x * b.z;
    float z = a.x * b.y - a.y * b.x;
    return v3(x, y, z);
}
static V4 v4mul(V4 a, V4 b) { V4 r; for (int i = 0; i < 4; i++) r.c[i] = a.c[i] * b.c[i]; return r; }
static V4 v4add(V4 a, V4 b) { V4 r; for (int i = 0; i < 4; i++) r.c[i] = a.c[i] + b.c[i]; return r; }
static V4 v4scale(float c, V4 a) { V4 r; for (int i = 0; i < 4; i++) r.c[i] = a.c[i] * c; return r; }
static V4 v4zero(void) { V4 r = {{0, 0, 0, 0}}; return r; }
static V4 v4from(const float* p) { V4 r = {{p[0], p[1], p[2], p[3]}}; return r; }

/* ---- rmath::Quat (include/raymath/geometry.h:17-204) ---- */
static float q4dot(Q4 a, Q4 b) {
    float s = 0;
    s += a.i * b.i;
    s += a.j * b.j;
    s += a.k * b.k;
    s += a.r * b.r;
    return s;
}
/* ref: geometry.h:158-161 Quat::normalized -> Vec4::normalized */
static Q4 q4normalized(Q4 q) {
    float len = sqrtf(q4dot(q, q));
    Q4 z = {0, 0, 0, 0};
    if (len > THR) {
        float inv = 1 / len;
        Q4 r = {q.i * inv, q.j * inv, q.k * inv, q.r * inv};
        return r;
    }
    return z;
}
/* ref: geometry.h:145-156 Quat::inverse */
static Q4 q4inverse(Q4 q) {
    float sq_norm = q4dot(q, q);
    Q4 z = {0, 0, 0, 0};
    if (sq_norm < THR) return z;
    float inv = 1 / sq_norm;
    Q4 r = {q.i * -inv, q.j * -inv, q.k * -inv, q.r * inv};
    return r;
}
/* ref: geometry.h:163-170 Quat::operator*= (Hamilton product, this evaluation order) */
static Q4 q4mul(Q4 a, Q4 o) {
    Q4 r;
    r.i = a.i * o.r + a.r * o.i + a.j * o.k - a.k * o.j;
    r.j = a.j * o.r + a.r * o.j + a.k * o.i - a.i * o.k;
    r.k = a.k * o.r + a.r * o.k + a.i * o.j - a.j * o.i;
    r.r = a.r * o.r - a.i * o.i - a.j * o.j - a.k * o.k;
    return r;
}
/* ref: geometry.h:176-181 operator*(Quat, Vec3): rotate, then restore the input length */
static V3 q4rotate(Q4 q, V3 v) {
    float length = v3len(v);
    Q4 qv = {v.x, v.y, v.z, 0};
    Q4 prod = q4mul(q4mul(q4normalized(q), qv), q4inverse(q));
    return v3scale(length, v3normalized(v3(prod.i, prod.j, prod.k)));
}

/* ---- rprimitives::Entity (src/rayprimitives/entity.cu:5-37) ---- */
typedef struct { Q4 o; V3 p; } Entity;
static Entity entity_from_pose(const RtPose* p) {
    Entity e;
    e.o.i = p->rot[0]; e.o.j = p->rot[1]; e.o.k = p->rot[2]; e.o.r = p->rot[3];
    e.p = v3(p->pos[0], p->pos[1], p->pos[2]);
    return e;
}
static V3 point_to_local(const Entity* e, V3 v) { return q4rotate(e->o, v3sub(v, e->p)); }
static V3 point_from_local(const Entity* e, V3 v) { return v3add(q4rotate(q4inverse(e->o), v), e->p); }
static V3 vec_to_local(const Entity* e, V3 v) { return q4rotate(e->o, v); }
static V3 vec_from_local(const Entity* e, V3 v) { return q4rotate(q4inverse(e->o), v); }
/* ref: geometry.h:216 Ray ctor normalises the direction */
static Ray ray_make(V3 o, V3 d) { Ray r; r.o = o; r.d = v3normalized(d); return r; }
static V3 ray_at(const Ray* r, float t) { return v3add(r->o, v3scale(t, r->d)); }
static Ray ray_to_local(const Entity* e, const Ray* r) {
    V3 ld = vec_to_local(e, r->d);
    V3 lp = point_to_local(e, r->o);
    return ray_make(lp, ld);
}

/* ---- ropt::BoundingBox (src/rayopt/bounding_box.cu:6-104) ---- */
typedef struct { V3 mn, mx; int nondegenerate; } Box;
static void box_fit(Box* b, V3 v) {
    if (!b->nondegenerate) { b->mn = v; b->mx = v; b->nondegenerate = 1; return; }
    if (v.x < b->mn.x) b->mn.x = v.x;
    if (v.x > b->mx.x) b->mx.x = v.x;
    if (v.y < b->mn.y) b->mn.y = v.y;
    if (v.y > b->mx.y) b->mx.y = v.y;
    if (v.z < b->mn.z) b->mn.z = v.z;
    if (v.z > b->mx.z) b->mx.z = v.z;
}
/* ref: bounding_box.cu:51-60 from_local — only the min and max corners are transformed */
static Box box_from_local(const Box* a, const Entity* e) {
    Box rv; memset(&rv, 0, sizeof rv);
    if (!a->nondegenerate) return rv;
    box_fit(&rv, point_from_local(e, a->mn));
    box_fit(&rv, point_from_local(e, a->mx));
    return rv;
}
/* ref: bounding_box.cu:62-104 intersects — Kay/Kajiya slabs, real divides, no test against the
 * current closest hit */
static int box_intersects(const Box* b, const Ray* r) {
    if (!b->nondegenerate) return 0;
    const float dir[3] = {r->d.x, r->d.y, r->d.z};
    const float org[3] = {r->o.x, r->o.y, r->o.z};
    const float mn[3] = {b->mn.x, b->mn.y, b->mn.z};
    const float mx[3] = {b->mx.x, b->mx.y, b->mx.z};
    float time_min = -INFINITY, time_max = INFINITY;
    for (int axis = 0; axis < 3; axis++) {
        if (dir[axis] == 0) continue;
        float time_near = (mn[axis] - org[axis]) / dir[axis];
        float time_far = (mx[axis] - org[axis]) / dir[axis];
        if (time_near > time_far) { float tmp = time_near; time_near = time_far; time_far = tmp; }
        if (time_near > time_min) time_min = time_near;
        if (time_far < time_max) time_max = time_far;
        if (time_min > time_max || time_max < THR) return 0;
    }
    return 1;
}

/* ---- scene ---- */
struct OrScene {
    int n_vertices, n_triangles, n_meshes, n_instances, n_lights;
    V3* vertices;
    V3* normals;
    RtTriangle* tris;
    RtMesh* meshes;
    RtInstance* instances;
    RtLight* lights;
    Box* inst_boxes;
    RtCamera cam;
    RtEnv env;
    int width, height;
    float* atlas; /* RGBA floats, NULL = none */
    int atlas_w, atlas_h;
};

typedef struct {
    float time;     /* Isect::time (reference to caller storage in the reference, isect.h:16) */
    V3 norm;
    float u, v;
    const RtTriangle* tri; /* stands for Isect::mat / Isect::coords (pointers into the TriInner) */
    int inst, tri_idx;     /* derived ids (the reference's Isect carries none) */
} Isect;

typedef struct { OrStats st; } Ctx;

/* ref: include/raymath/geometry.h:253-261 Plane::hit and :275-290 Triangle::hit */
static int triangle_hit(V3 a, V3 b, V3 c, const Ray* r, float* time, float* u, float* v) {
    V3 plane_norm = v3cross(v3sub(b, a), v3sub(c, a));
    V3 n = v3normalized(plane_norm); /* Plane ctor */
    float denom = v3dot(r->d, n);
    if (fabsf(denom) < THR) return 0;
    *time = (1.0f / denom) * v3dot(v3sub(a, r->o), n);
    V3 p = ray_at(r, *time);
    float tri_area = v3len(plane_norm);
    float bary0 = v3len(v3cross(v3sub(c, p), v3sub(b, p))) / tri_area;
    float bary1 = v3len(v3cross(v3sub(c, p), v3sub(a, p))) / tri_area;
    float bary2 = v3len(v3cross(v3sub(a, p), v3sub(b, p))) / tri_area;
    if (fabsf(bary0 + bary1 + bary2 - 1.0f) <= THR) {
        *u = bary1;
        *v = bary2;
        return 1;
    }
    return 0;
}

/* ref: src/rayprimitives/trimesh.cu:46-68 (gpu) / :94-118 (cpu) TriInner::tri_hit */
static int tri_hit(const OrScene* s, const RtTriangle* tri, const Ray* local_ray, Isect* isect, Ctx* ctx) {
    V3 a = s->vertices[tri->v[0]], b = s->vertices[tri->v[1]], c = s->vertices[tri->v[2]];
    float time, u, v;
    ctx->st.tri_tests++;
    if (triangle_hit(a, b, c, local_ray, &time, &u, &v) && time >= THR && time < isect->time) {
        isect->tri = tri;
        isect->u = u;
        isect->v = v;
        V3 n0 = s->normals[tri->v[0]], n1 = s->normals[tri->v[1]], n2 = s->normals[tri->v[2]];
        float bary0 = 1.0f - u - v;
        isect->norm = v3normalized(v3add(v3add(v3scale(bary0, n0), v3scale(u, n1)), v3scale(v, n2)));
        isect->time = time;
        return 1;
    }
    return 0;
}

/* Sphere::hit_local — the analytic sphere hitable. NO reference counterpart: the reference has one
 * Hitable subclass (include/rayprimitives/gpu/hitable.cuh:27-47). Written in the reference's conventions: unit local
 * direction (Ray ctor), rmath::dot's summation order, `time >= THRESHOLD && time < isect.time` like TriInner::tri_hit,
 * normal = normalized(hit point) in the hitable's frame. Centre at the local origin, b = o.d, c = o.o - r^2,
 * roots -b -+ sqrt(b^2 - c): the near one unless it is below THRESHOLD (origin inside or on the surface). */
static int sphere_hit(const OrScene* s, const RtMesh* m, const Ray* local_ray, Isect* isect, Ctx* ctx) {
    float b = v3dot(local_ray->o, local_ray->d);
    float c = v3dot(local_ray->o, local_ray->o) - m->radius * m->radius;
    float disc = b * b - c;
    ctx->st.tri_tests++;
    if (!(disc >= 0.0f)) return 0;
    float sq = sqrtf(disc);
    float time = -b - sq;
    if (!(time >= THR)) time = -b + sq;
    if (time >= THR && time < isect->time) {
        isect->tri = &s->tris[m->tri_begin];
        isect->u = 0.0f;
        isect->v = 0.0f;
        isect->norm = v3normalized(ray_at(local_ray, time));
        isect->time = time;
        return 1;
    }
    return 0;
}

/* ref: src/rayprimitives/hitable.cu:7-38 HitHandle + Hitable::hit, trimesh.cu:11-19 hit_local */
static int hitable_hit(const OrScene* s, int mesh_idx, const Ray* ray, Isect* isect, Ctx* ctx) {
    const RtMesh* m = &s->meshes[mesh_idx];
    Entity e = entity_from_pose(&m->pose);
    float scale = v3len(vec_to_local(&e, ray->d));
    Ray local_ray = ray_to_local(&e, ray);
    int hit = 0;
    if (m->kind == RT_HITABLE_SPHERE) {
        if (sphere_hit(s, m, &local_ray, isect, ctx)) {
            hit = 1;
            isect->tri_idx = 0;
        }
    } else
    for (int i = m->tri_begin; i < m->tri_end; i++) {
        if (tri_hit(s, &s->tris[i], &local_ray, isect, ctx)) {
            hit = 1;
            isect->tri_idx = i - m->tri_begin;
        }
    }
    if (hit) {
        isect->norm = v3normalized(vec_from_local(&e, isect->norm));
        isect->time *= scale;
    }
    return hit;
}

/* ref: src/rayenv/scene.cu:27-40 (gpu) / :192-203 (cpu) cast_local */
static int cast_local(const OrScene* s, const Ray* r, Isect* isect, int inst, Ctx* ctx) {
    const RtInstance* t = &s->instances[inst];
    Entity e = entity_from_pose(&t->pose);
    V3 local_dir = vec_to_local(&e, r->d);
    float dir_len = v3len(local_dir);
    Ray local_ray = ray_make(point_to_local(&e, r->o), local_dir);
    ctx->st.leaf_tests++;
    int rv = hitable_hit(s, t->mesh, &local_ray, isect, ctx);
    if (rv) {
        isect->norm = vec_from_local(&e, isect->norm);
        isect->time *= dir_len;
        isect->inst = inst;
    }
    return rv;
}

/* ref: src/rayenv/scene.cu:42-73 (gpu) / :205-220 (cpu) cast_ray.
 * use_bvh: both the serial tree walk (src/rayopt/bvh.cc:48-67) and the CUDA iterator
 * (src/rayopt/bvh.cu:98-156) call cast_local exactly for the instances whose LEAF box passes
 * BoundingBox::intersects — an inner box is the union of its leaves and every operation of the
 * slab test is monotone in the box bounds, so a leaf hit implies all its ancestors hit. The tree
 * shape therefore only changes the visiting order, which matters only for exact-t ties. */
static int cast_ray(const OrScene* s, const Ray* r, Isect* isect, int use_bvh, Ctx* ctx) {
    int hit = 0;
    ctx->st.rays_total++;
    for (int i = 0; i < s->n_instances; i++) {
        if (use_bvh) {
            ctx->st.box_tests++;
            if (!box_intersects(&s->inst_boxes[i], r)) continue;
        }
        hit |= cast_local(s, r, isect, i, ctx);
    }
    return hit;
}

/* ---- rprimitives::Material predicates (include/rayprimitives/material.h:103-111) ---- */
static int mat_reflective(const RtMaterial* m) { return m->Kr[0] > 0.0f || m->Kr[1] > 0.0f || m->Kr[2] > 0.0f || m->Kr[3] > 0.0f; }
static int mat_refractive(const RtMaterial* m) { return m->Kt[0] > 0.0f || m->Kt[1] > 0.0f || m->Kt[2] > 0.0f || m->Kt[3] > 0.0f; }

/* ref: include/raymath/linear.h:215-225 reflect */
static V3 reflect3(V3 dir, V3 norm) {
    float d_len = v3len(dir);
    V3 d_norm = v3normalized(dir);
    V3 n_norm = v3normalized(norm);
    V3 projection = v3scale(v3dot(d_norm, n_norm), n_norm);
    V3 r_norm = v3normalized(v3sub(d_norm, v3scale(2, projection)));
    return v3scale(d_len, r_norm);
}
/* ref: include/raymath/linear.h:227-242 refract */
static V3 refract3(V3 dir, V3 norm, float index_from, float index_to, int* tir) {
    float d_len = v3len(dir);
    V3 d_norm = v3normalized(dir);
    V3 n_norm = v3normalized(norm);
    float index_ratio = index_from / index_to;
    float cosi = v3dot(d_norm, n_norm);
    float sint_2 = index_ratio * index_ratio * (1 - cosi * cosi);
    if (sint_2 > 1) {
        *tir = 1;
        return v3scale(d_len, reflect3(d_norm, n_norm));
    }
    *tir = 0;
    return v3scale(d_len, v3add(v3scale(index_ratio, d_norm), v3scale(index_ratio * cosi - sqrtf(1 - sint_2), n_norm)));
}

/* ref: src/rayenv/scene.cu:14-22 trans_atten: pow(time, Kt) */
static V4 trans_atten(const RtMaterial* m, float time) {
    V4 r = {{powf(time, m->Kt[0]), powf(time, m->Kt[1]), powf(time, m->Kt[2]), powf(time, m->Kt[3])}};
    return r;
}
/* ref: src/rayprimitives/light.cu:11-17 calc_dist_atten */
static float calc_dist_atten(const OrScene* s, float dist) {
    const float* c = s->env.dist_atten;
    float quad = c[0] + c[1] * dist + c[2] * dist * dist;
    return quad < 1.0f ? 1.0f : 1.0f / quad;
}
/* ref: src/rayprimitives/light.cu:19-26 calc_shadow_atten: pow(Kt, dist) */
static V4 calc_shadow_atten(const float* kt, float dist) {
    V4 r = {{powf(kt[0], dist), powf(kt[1], dist), powf(kt[2], dist), powf(kt[3], dist)}};
    return r;
}

/* ref: src/rayprimitives/light.cu:29-61 (gpu) / :81-114 (cpu) Light::attenuate */
static V4 light_attenuate(const OrScene* s, const RtLight* l, const Ray* to_light, float max_t, int use_bvh, Ctx* ctx) {
    V4 rv = v4from(l->color);
    Ray cur = ray_make(ray_at(to_light, THR), to_light->d);
    for (;;) {
        Isect sh;
        memset(&sh, 0, sizeof sh);
        sh.time = INFINITY;
        sh.inst = -1;
        ctx->st.rays_shadow++;
        if (cast_ray(s, &cur, &sh, use_bvh, ctx)) {
            if (sh.time > max_t) return rv;
            if (!mat_refractive(&sh.tri->mat)) return v4zero();
            if (v3dot(sh.norm, cur.d) > 0) rv = v4mul(rv, calc_shadow_atten(sh.tri->mat.Kt, sh.time));
            cur = ray_make(ray_at(&cur, sh.time), cur.d);
            max_t -= sh.time;
        } else {
            return rv;
        }
    }
}
/* ref: src/rayprimitives/light.cu:63-77 PointLight::shine / DirLight::shine */
static V4 light_shine(const OrScene* s, const RtLight* l, V3 hit_pos, V3* dir_to_light, int use_bvh, Ctx* ctx) {
    V3 lv = v3(l->v[0], l->v[1], l->v[2]);
    if (l->kind == RT_LIGHT_POINT) {
        V3 disp = v3sub(lv, hit_pos);
        float dist = v3len(disp);
        float dist_atten = calc_dist_atten(s, dist);
        *dir_to_light = v3normalized(disp);
        Ray to_light = ray_make(hit_pos, *dir_to_light);
        return v4scale(dist_atten, light_attenuate(s, l, &to_light, dist, use_bvh, ctx));
    }
    *dir_to_light = v3neg(lv);
    Ray to_light = ray_make(hit_pos, *dir_to_light);
    return light_attenuate(s, l, &to_light, INFINITY, use_bvh, ctx);
}
/* The texture branch phong.cu:19-23 leaves as `// TODO: implement texture mapping` (NO reference
 * counterpart): TextureCoords (include/rayprimitives/texture_coords.h:12-29) = atlas origin + the atlas-space directions
 * of the barycentric axes, Isect::uv = (b1, b2); texel fetch like get_color_from_texture (src/rayprimitives/texture.cu:8-12)
 * through the reference's descriptor: point filter, clamp, unnormalised coordinates (include/gputils/alloc.h:52-56). */
static V4 diffuse_colour(const OrScene* s, const Isect* isect) {
    const RtTriangle* t = isect->tri;
    if (t->tex_degenerate || !s->atlas) return v4from(t->mat.Kd);
    float tx = ((float)t->tex_x + isect->u * t->tex_u[0]) + isect->v * t->tex_v[0];
    float ty = ((float)t->tex_y + isect->u * t->tex_u[1]) + isect->v * t->tex_v[1];
    int ix = (int)floorf(tx), iy = (int)floorf(ty);
    ix = ix < 0 ? 0 : (ix >= s->atlas_w ? s->atlas_w - 1 : ix);
    iy = iy < 0 ? 0 : (iy >= s->atlas_h ? s->atlas_h - 1 : iy);
    return v4from(s->atlas + 4 * ((size_t)iy * s->atlas_w + ix));
}

/* ref: src/rayprimitives/phong.cu:14-33 phong. The reference computes diffuse_col and then multiplies Kd regardless; with
 * a degenerate TextureCoords (every shipped world) both are Kd. */
static V4 phong(const OrScene* s, const Isect* isect, V4 incoming, V3 ray_dir, V3 dir_to_light) {
    const RtMaterial* m = &isect->tri->mat;
    float norm_dot = fmaxf(v3dot(dir_to_light, isect->norm), 0.0f);
    V4 diffuse = v4scale(norm_dot, diffuse_colour(s, isect));
    V3 reflected = reflect3(v3neg(dir_to_light), isect->norm);
    float reflect_dot = v3dot(v3neg(reflected), ray_dir);
    V4 specular = v4scale(powf(fmaxf(reflect_dot, 0.0f), m->alpha), v4from(m->Ks));
    return v4mul(v4add(diffuse, specular), incoming);
}
/* ref: src/rayprimitives/phong.cu:35-53 org_light + illuminate */
static V4 illuminate(const OrScene* s, const Ray* org_ray, const Isect* isect, int use_bvh, Ctx* ctx) {
    const RtMaterial* m = &isect->tri->mat;
    V4 sum = v4add(v4from(m->Ke), v4mul(v4from(m->Ka), v4from(s->env.ambience)));
    for (int i = 0; i < s->n_lights; i++) {
        V3 dir_to_light;
        V4 incoming = light_shine(s, &s->lights[i], ray_at(org_ray, isect->time), &dir_to_light, use_bvh, ctx);
        sum = v4add(sum, phong(s, isect, incoming, org_ray->d, dir_to_light));
    }
    return sum;
}

/* ref: src/rayenv/scene.cu:222-273 renv::cpu::Scene::propagate_helper / propagate_ray (serial
 * semantics: Kr never applied, in_obj never toggles, depth guard never fires — SURVEY.md §3.4) */
static V4 propagate_cpu(const OrScene* s, const Ray* r, Isect* isect, float* last_time, int in_obj, int depth,
                        int use_bvh, Ctx* ctx, int level) {
    if (depth > s->env.depth) return v4zero();
    isect->time = INFINITY;
    if (level == 0) ctx->st.rays_primary++; else ctx->st.rays_secondary++;
    if (cast_ray(s, r, isect, use_bvh, ctx)) {
        V4 acc = illuminate(s, r, isect, use_bvh, ctx);
        *last_time = isect->time;
        float next_time = 0;
        if (mat_reflective(&isect->tri->mat)) {
            V3 rd = reflect3(r->d, isect->norm);
            Ray rr = ray_make(ray_at(r, isect->time), rd);
            acc = v4add(acc, propagate_cpu(s, &rr, isect, &next_time, in_obj, depth - 1, use_bvh, ctx, level + 1));
        }
        if (mat_refractive(&isect->tri->mat)) {
            float n1, n2;
            int tir;
            if (in_obj) { n1 = isect->tri->mat.eta; n2 = 1.0f; } else { n1 = 1.0f; n2 = isect->tri->mat.eta; }
            V3 rd = refract3(r->d, isect->norm, n1, n2, &tir);
            if (!tir) {
                next_time = 0;
                Ray rr = ray_make(ray_at(r, isect->time), rd);
                V4 un = propagate_cpu(s, &rr, isect, &next_time, in_obj, depth - 1, use_bvh, ctx, level + 1);
                acc = v4add(acc, v4mul(trans_atten(&isect->tri->mat, next_time), un));
            }
        }
        return acc;
    }
    return v4zero();
}

/* ref: src/rayenv/scene.cu:75-188 renv::gpu::propagate_ray — explicit RayFrame[10] stack */
enum { FR_NORMAL, FR_REFLECT, FR_REFRACT };
typedef struct {
    Ray ray;
    V3 hit_pt, norm;
    V4 atten;
    const RtMaterial* last_mat;
    int type, depth, in_obj;
} RayFrame;
#define OR_MAX_DEPTH 10

static V4 propagate_gpu(const OrScene* s, const Ray* r, Isect* isect, int use_bvh, Ctx* ctx) {
    RayFrame frames[OR_MAX_DEPTH];
    memset(frames, 0, sizeof frames);
    frames[0].ray = *r;
    frames[0].atten.c[0] = frames[0].atten.c[1] = frames[0].atten.c[2] = frames[0].atten.c[3] = 1.0f;
    frames[0].last_mat = NULL;
    frames[0].type = FR_NORMAL;
    frames[0].depth = s->env.depth;
    frames[0].in_obj = 0;
    int stack_top = 0;
    V4 acc = v4zero();
    int first = 1;
    while (stack_top >= 0) {
        RayFrame* top = &frames[stack_top];
        switch (top->type) {
        case FR_NORMAL: {
            isect->time = INFINITY;
            if (first) { ctx->st.rays_primary++; first = 0; } else ctx->st.rays_secondary++;
            if (cast_ray(s, &top->ray, isect, use_bvh, ctx)) {
                if (top->depth > 0) {
                    if (top->in_obj) top->atten = v4mul(top->atten, trans_atten(&isect->tri->mat, isect->time));
                    top->type = FR_REFLECT;
                    top->hit_pt = ray_at(&top->ray, isect->time);
                    top->last_mat = &isect->tri->mat;
                    top->norm = isect->norm;
                } else {
                    stack_top--;
                }
                acc = v4add(acc, v4mul(top->atten, illuminate(s, &top->ray, isect, use_bvh, ctx)));
            } else {
                stack_top--;
            }
            break;
        }
        case FR_REFLECT: {
            /* isect still holds the most recent successful closest hit (possibly from a deeper
             * bounce — the reference's "stale isect") */
            V4 kr = v4from(isect->tri->mat.Kr);
            top->type = FR_REFRACT;
            if (mat_reflective(&isect->tri->mat)) {
                stack_top++;
                RayFrame* nt = &frames[stack_top];
                nt->type = FR_NORMAL;
                nt->last_mat = top->last_mat;
                nt->in_obj = top->in_obj;
                nt->atten = v4mul(top->atten, kr);
                nt->depth = top->depth - 1;
                V3 rd = reflect3(top->ray.d, v3normalized(top->norm));
                nt->ray = ray_make(top->hit_pt, rd);
            }
            break;
        }
        case FR_REFRACT: {
            if (mat_refractive(&isect->tri->mat)) {
                top->type = FR_NORMAL;
                float n1, n2;
                int tir;
                if (top->in_obj) { n1 = top->last_mat->eta; n2 = 1.0f; } else { n1 = 1.0f; n2 = top->last_mat->eta; }
                V3 rd = refract3(top->ray.d, v3normalized(top->norm), n1, n2, &tir);
                if (tir) {
                    stack_top--;
                } else {
                    top->ray = ray_make(top->hit_pt, rd);
                    top->in_obj = !top->in_obj;
                    top->depth--;
                }
            } else {
                stack_top--;
            }
            break;
        }
        }
    }
    return acc;
}

/* ref: include/raymath/geometry.h:183-198 Quat::to_Mat3; src/rayenv/camera.cu:32-42 Camera::at */
static Ray camera_at(const OrScene* s, float canvas_x, float canvas_y) {
    const RtCamera* c = &s->cam;
    float qi = c->pose.rot[0], qj = c->pose.rot[1], qk = c->pose.rot[2], qr = c->pose.rot[3];
    Q4 q = {qi, qj, qk, qr};
    float length = sqrtf(q4dot(q, q));
    float ni = qi / length, nj = qj / length, nk = qk / length, nr = qr / length;
    float ii = 2.0f * ni * ni, jj = 2.0f * nj * nj, kk = 2.0f * nk * nk;
    float ri = 2.0f * nr * ni, rj = 2.0f * nr * nj, rk = 2.0f * nr * nk, ij = 2.0f * ni * nj, ik = 2.0f * ni * nk,
          jk = 2.0f * nj * nk;
    float m[3][3] = {{1 - (jj + kk), ij - rk, ik + rj}, {ij + rk, 1 - (ii + kk), jk - ri}, {ik - rj, jk + ri, 1 - (ii + jj)}};
    V3 r = v3normalized(v3(m[0][0], m[1][0], m[2][0]));
    V3 u = v3normalized(v3(m[0][1], m[1][1], m[2][1]));
    V3 f = v3normalized(v3(m[0][2], m[1][2], m[2][2]));
    float global_x = (canvas_x - (0.5f * c->canvas_width)) / c->unit_to_pixels;
    float global_y = (0.5f * c->canvas_height - canvas_y) / c->unit_to_pixels;
    V3 dir = v3add(v3add(v3scale(c->global_near, f), v3scale(global_x, r)), v3scale(global_y, u));
    return ray_make(v3(c->pose.pos[0], c->pose.pos[1], c->pose.pos[2]), dir);
}

/* ref: src/raytracer.cu:37-40 clamp; include/rayenv/color.h:41-42 float->u8 truncation;
 * src/rayenv/color.cu:24-26 to_encoding */
static uint32_t pack_color(V4 c) {
    uint32_t ch[4];
    for (int i = 0; i < 4; i++) {
        float v = c.c[i] > 1.0f ? 1.0f : c.c[i];
        ch[i] = (uint32_t)(uint8_t)(int)(255.0f * v);
    }
    return (uint32_t)((int)(ch[0] << 24) + (int)(ch[1] << 16) + (int)(ch[2] << 8) + (int)ch[3]);
}

static V4 propagate(const OrScene* s, int semantics, const Ray* r, Isect* isect, int use_bvh, Ctx* ctx) {
    if (semantics == RT_SEMANTICS_CPU) {
        float t;
        return propagate_cpu(s, r, isect, &t, 0, s->env.depth, use_bvh, ctx, 0);
    }
    return propagate_gpu(s, r, isect, use_bvh, ctx);
}

static void isect_reset(Isect* i) {
    memset(i, 0, sizeof *i);
    i->time = INFINITY;
    i->inst = -1;
    i->tri_idx = -1;
}

/* ---- scene construction ---- */
/* ref: src/raytracer.cc:10-27 create_bvh box computation; src/rayprimitives/trimesh.cu:80-92
 * Trimesh::compute_bounding_box */
static void compute_instance_boxes(OrScene* s) {
    for (int i = 0; i < s->n_instances; i++) {
        const RtInstance* t = &s->instances[i];
        const RtMesh* m = &s->meshes[t->mesh];
        Box local;
        memset(&local, 0, sizeof local);
        if (m->kind == RT_HITABLE_SPHERE) { /* rotation-independent: world centre +- radius */
            Entity me = entity_from_pose(&m->pose), te = entity_from_pose(&t->pose);
            V3 c = point_from_local(&te, point_from_local(&me, v3(0, 0, 0)));
            s->inst_boxes[i].mn = v3(c.x - m->radius, c.y - m->radius, c.z - m->radius);
            s->inst_boxes[i].mx = v3(c.x + m->radius, c.y + m->radius, c.z + m->radius);
            s->inst_boxes[i].nondegenerate = 1;
            continue;
        }
        for (int k = m->tri_begin; k < m->tri_end; k++)
            for (int j = 0; j < 3; j++) box_fit(&local, s->vertices[s->tris[k].v[j]]);
        Entity me = entity_from_pose(&m->pose);
        Box mesh_box = box_from_local(&local, &me);
        Entity te = entity_from_pose(&t->pose);
        s->inst_boxes[i] = box_from_local(&mesh_box, &te);
    }
}

static void* dup_mem(const void* p, size_t n) {
    void* q = malloc(n ? n : 1);
    if (n) memcpy(q, p, n);
    return q;
}

OrScene* or_scene_create(const RtSceneDesc* d) {
    OrScene* s = (OrScene*)calloc(1, sizeof(OrScene));
    s->n_vertices = d->n_vertices;
    s->n_triangles = d->n_triangles;
    s->n_meshes = d->n_meshes;
    s->n_instances = d->n_instances;
    s->n_lights = d->n_lights;
    s->vertices = (V3*)dup_mem(d->vertices, sizeof(float) * 3 * (size_t)d->n_vertices);
    s->normals = (V3*)dup_mem(d->normals, sizeof(float) * 3 * (size_t)d->n_vertices);
    s->tris = (RtTriangle*)dup_mem(d->triangles, sizeof(RtTriangle) * (size_t)d->n_triangles);
    s->meshes = (RtMesh*)dup_mem(d->meshes, sizeof(RtMesh) * (size_t)d->n_meshes);
    s->instances = (RtInstance*)dup_mem(d->instances, sizeof(RtInstance) * (size_t)d->n_instances);
    s->lights = (RtLight*)dup_mem(d->lights, sizeof(RtLight) * (size_t)d->n_lights);
    s->inst_boxes = (Box*)calloc((size_t)(d->n_instances ? d->n_instances : 1), sizeof(Box));
    s->cam = d->camera;
    s->env = d->env;
    s->width = d->width;
    s->height = d->height;
    if (d->atlas && d->atlas_width > 0 && d->atlas_height > 0) {
        s->atlas = (float*)dup_mem(d->atlas, sizeof(float) * 4 * (size_t)d->atlas_width * (size_t)d->atlas_height);
        s->atlas_w = d->atlas_width;
        s->atlas_h = d->atlas_height;
    }
    compute_instance_boxes(s);
    return s;
}

void or_scene_destroy(OrScene* s) {
    if (!s) return;
    free(s->vertices); free(s->normals); free(s->tris); free(s->meshes);
    free(s->instances); free(s->lights); free(s->inst_boxes); free(s->atlas); free(s);
}
void or_scene_set_camera(OrScene* s, const RtCamera* cam) { s->cam = *cam; }
void or_scene_set_env(OrScene* s, const RtEnv* env) { s->env = *env; }
void or_scene_set_canvas_size(OrScene* s, int32_t w, int32_t h) {
    s->width = w; s->height = h;
    s->cam.canvas_width = (float)w; s->cam.canvas_height = (float)h;
}

void or_instance_boxes(const OrScene* s, float* b) {
    for (int i = 0; i < s->n_instances; i++) {
        const Box* x = &s->inst_boxes[i];
        b[6 * i + 0] = x->mn.x; b[6 * i + 1] = x->mn.y; b[6 * i + 2] = x->mn.z;
        b[6 * i + 3] = x->mx.x; b[6 * i + 4] = x->mx.y; b[6 * i + 5] = x->mx.z;
    }
}

/* ---- frame drivers (ref: src/raytracer.cc:41-62 update_scene; pixels are independent) ---- */
typedef struct {
    const OrScene* s;
    int semantics, use_bvh, y0, y1, mode; /* mode 0 = render, 1 = primary ids */
    uint32_t* canvas;
    int32_t *inst, *tri;
    float* t;
    OrStats st;
} Job;

static void stats_add(OrStats* a, const OrStats* b) {
    a->rays_total += b->rays_total; a->rays_primary += b->rays_primary; a->rays_shadow += b->rays_shadow;
    a->rays_secondary += b->rays_secondary; a->box_tests += b->box_tests; a->leaf_tests += b->leaf_tests;
    a->tri_tests += b->tri_tests;
}

static void* job_run(void* arg) {
    Job* j = (Job*)arg;
    const OrScene* s = j->s;
    Ctx ctx;
    memset(&ctx, 0, sizeof ctx);
    for (int y = j->y0; y < j->y1; y++) {
        for (int x = 0; x < s->width; x++) {
            Ray r = camera_at(s, (float)x, (float)y);
            Isect isect;
            isect_reset(&isect);
            size_t p = (size_t)y * s->width + x;
            if (j->mode == 0) {
                V4 c = propagate(s, j->semantics, &r, &isect, j->use_bvh, &ctx);
                j->canvas[p] = pack_color(c);
            } else {
                ctx.st.rays_primary++;
                int hit = cast_ray(s, &r, &isect, j->use_bvh, &ctx);
                j->inst[p] = hit ? isect.inst : -1;
                j->tri[p] = hit ? isect.tri_idx : -1;
                j->t[p] = isect.time;
            }
        }
    }
    j->st = ctx.st;
    return NULL;
}

static void run_jobs(Job proto, int y0, int y1, int threads, OrStats* stats) {
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    int rows = y1 - y0;
    if (rows <= 0) return;
    /* interleave small row bands so sky rows and geometry rows mix across threads */
    int band = 4;
    int nb = (rows + band - 1) / band;
    if (threads > nb) threads = nb;
    Job* jobs = (Job*)calloc((size_t)nb, sizeof(Job));
    for (int b = 0; b < nb; b++) {
        jobs[b] = proto;
        jobs[b].y0 = y0 + b * band;
        jobs[b].y1 = jobs[b].y0 + band < y1 ? jobs[b].y0 + band : y1;
    }
    if (threads == 1) {
        for (int b = 0; b < nb; b++) job_run(&jobs[b]);
    } else {
        /* static round-robin of bands over threads */
        typedef struct { Job* jobs; int nb, first, stride; } Pack;
        Pack* packs = (Pack*)calloc((size_t)threads, sizeof(Pack));
        pthread_t* th = (pthread_t*)calloc((size_t)threads, sizeof(pthread_t));
        extern void* or_pack_run(void*);
        for (int t = 0; t < threads; t++) {
            packs[t].jobs = jobs; packs[t].nb = nb; packs[t].first = t; packs[t].stride = threads;
            pthread_create(&th[t], NULL, or_pack_run, &packs[t]);
        }
        for (int t = 0; t < threads; t++) pthread_join(th[t], NULL);
        free(packs);
        free(th);
    }
    if (stats) for (int b = 0; b < nb; b++) stats_add(stats, &jobs[b].st);
    free(jobs);
}

void* or_pack_run(void* arg) {
    struct { Job* jobs; int nb, first, stride; }* p = arg;
    for (int b = p->first; b < p->nb; b += p->stride) job_run(&p->jobs[b]);
    return NULL;
}

void or_render(const OrScene* s, int semantics, int use_bvh, int y0, int y1, uint32_t* canvas, OrStats* stats,
               int threads) {
    Job j;
    memset(&j, 0, sizeof j);
    j.s = s; j.semantics = semantics; j.use_bvh = use_bvh; j.mode = 0; j.canvas = canvas;
    if (stats) memset(stats, 0, sizeof *stats);
    run_jobs(j, y0, y1, threads, stats);
}

void or_trace_primary(const OrScene* s, int use_bvh, int32_t* instance, int32_t* triangle, float* t, int threads) {
    Job j;
    memset(&j, 0, sizeof j);
    j.s = s; j.use_bvh = use_bvh; j.mode = 1; j.inst = instance; j.tri = triangle; j.t = t;
    run_jobs(j, 0, s->height, threads, NULL);
}

void or_cast_rays(const OrScene* s, int use_bvh, const float* origins, const float* dirs, int32_t n,
                  int32_t* instance, int32_t* triangle, float* t) {
    Ctx ctx;
    memset(&ctx, 0, sizeof ctx);
    for (int i = 0; i < n; i++) {
        Ray r = ray_make(v3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]),
                         v3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]));
        Isect isect;
        isect_reset(&isect);
        int hit = cast_ray(s, &r, &isect, use_bvh, &ctx);
        instance[i] = hit ? isect.inst : -1;
        triangle[i] = hit ? isect.tri_idx : -1;
        t[i] = isect.time;
    }
}

void or_primary_ray(const OrScene* s, int x, int y, float* o, float* d) {
    Ray r = camera_at(s, (float)x, (float)y);
    o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z;
    d[0] = r.d.x; d[1] = r.d.y; d[2] = r.d.z;
}

void or_debug_cast(const OrScene* s, int semantics, int use_bvh, int x, int y, RtHit* out) {
    Ctx ctx;
    memset(&ctx, 0, sizeof ctx);
    Ray r = camera_at(s, (float)x, (float)y);
    Isect isect;
    isect_reset(&isect);
    int hit = cast_ray(s, &r, &isect, use_bvh, &ctx);
    out->instance = hit ? isect.inst : -1;
    out->triangle = hit ? isect.tri_idx : -1;
    out->t = isect.time;
    out->normal[0] = isect.norm.x; out->normal[1] = isect.norm.y; out->normal[2] = isect.norm.z;
    isect_reset(&isect);
    V4 c = propagate(s, semantics, &r, &isect, use_bvh, &ctx);
    for (int i = 0; i < 4; i++) out->color[i] = c.c[i];
}

/* the survey's canvas hash: h = h*1315423911 + pixel, row-major, 64-bit wrap */
uint64_t or_canvas_hash(const uint32_t* canvas, int64_t n, int32_t* nonzero) {
    uint64_t h = 0;
    int32_t nz = 0;
    for (int64_t i = 0; i < n; i++) {
        h = h * 1315423911ull + canvas[i];
        if (canvas[i]) nz++;
    }
    if (nonzero) *nonzero = nz;
    return h;
}
