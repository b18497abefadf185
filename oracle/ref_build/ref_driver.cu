// ref_driver.cu — TEST INFRASTRUCTURE ONLY (never linked into the product library).
//
// Headless driver around the UNMODIFIED reference sources under /root/reference (compiled where
// they lie by oracle/ref_build/Makefile, outputs only into oracle/_ref/). It calls the reference's
// own public entry points:
//   procedural::{cpu,gpu}::generate      (/root/reference/include/procedural/cube_world.h:20-27)
//   rtracer::{cpu,gpu}::update_scene     (/root/reference/include/raytracer.h:18-28)
// exactly like the reference's bench() does (/root/reference/src/main.cc:210-216), and dumps
//   * the RGBA8 canvas + the survey's canvas hash,
//   * per-pixel primary-ray closest hit (instance idx, triangle idx, t), with and without the
//     leaf-AABB filter the BVH path applies (src/rayopt/bounding_box.cu:62-104),
//   * the flattened scene the loader produced (to validate our own host loader against it).
//
// Built twice: ref_cpu (-DREF_INTERPOSE: cudaMallocManaged/cudaFree -> calloc/free so the serial
// path runs on a box without a GPU; Canvas and Perlin allocate managed memory even in serial
// mode, src/rayenv/canvas.cu:7, src/procedural/perlin.cu:87-88) and ref_gpu (real CUDA runtime).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <functional>
#include <iostream>
#include <string>
#include <vector>
#include <cuda_runtime.h>

#define private public
#define protected public
#include "raytracer.h"
#include "procedural/cube_world.h"
#include "rayenv/cpu/scene.h"
#include "rayenv/gpu/scene.h"
#include "rayenv/transformation.h"
#include "rayprimitives/cpu/hitable.h"
#include "rayprimitives/cpu/trimesh.h"
#include "rayprimitives/cpu/light.h"
#include "rayprimitives/cpu/vertex_buffer.h"
#include "rayprimitives/isect.h"
#include "rayopt/bounding_box.h"
#undef private
#undef protected

// (the allocator interposer for ref_cpu lives in interpose.c, compiled by gcc and linked first)

static uint32_t fbits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

static uint64_t canvas_hash(renv::Canvas& canvas, int& nonzero) {
    uint64_t h = 0;
    nonzero = 0;
    for (int y = 0; y < canvas.get_height(); y++) {
        for (int x = 0; x < canvas.get_width(); x++) {
            uint32_t px = (uint32_t)canvas.buffer[y * canvas.get_width() + x];
            h = h * 1315423911ull + px;
            if (px) nonzero++;
        }
    }
    return h;
}

static void dump_canvas(renv::Canvas& canvas, const std::string& path) {
    FILE* f = fopen(path.c_str(), "wb");
    int w = canvas.get_width(), h = canvas.get_height();
    fwrite(&w, 4, 1, f);
    fwrite(&h, 4, 1, f);
    fwrite(canvas.buffer, 4, (size_t)w * h, f);
    fclose(f);
}

static void pvec(FILE* f, const char* key, const float* v, int n) {
    fprintf(f, "%s %d", key, n);
    for (int i = 0; i < n; i++) fprintf(f, " %08x", fbits(v[i]));
    fprintf(f, "\n");
}

static void dump_scene(renv::cpu::Scene* scene, const std::string& path) {
    FILE* f = fopen(path.c_str(), "w");
    renv::Environment& env = scene->get_environment();
    renv::Camera& cam = env.get_camera();
    fprintf(f, "canvas %d %d\n", env.get_canvas().get_width(), env.get_canvas().get_height());
    float camv[11] = {cam.o.i(), cam.o.j(), cam.o.k(), cam.o.r(), cam.p[0], cam.p[1], cam.p[2],
                      cam.global_near, cam.unit_to_pixels, cam.canvas_width, cam.canvas_height};
    pvec(f, "camera", camv, 11);
    float da[3] = {env.dist_atten[0], env.dist_atten[1], env.dist_atten[2]};
    pvec(f, "dist_atten", da, 3);
    float amb[4] = {env.ambience[0], env.ambience[1], env.ambience[2], env.ambience[3]};
    pvec(f, "ambience", amb, 4);
    fprintf(f, "depth %d\n", env.recurse_depth);
    rprimitives::cpu::VertexBuffer& vb = scene->get_vertex_buffer();
    int nv = (int)vb.v.size();
    fprintf(f, "n_vertices %d\n", nv);
    for (int i = 0; i < nv; i++) {
        float v[6] = {vb.v[i][0], vb.v[i][1], vb.v[i][2], vb.n[i][0], vb.n[i][1], vb.n[i][2]};
        pvec(f, "v", v, 6);
    }
    auto& hitables = scene->get_hitables();
    fprintf(f, "n_meshes %d\n", (int)hitables.size());
    for (size_t m = 0; m < hitables.size(); m++) {
        rprimitives::cpu::Trimesh* mesh = (rprimitives::cpu::Trimesh*)hitables[m];
        float pose[7] = {mesh->o.i(), mesh->o.j(), mesh->o.k(), mesh->o.r(), mesh->p[0], mesh->p[1], mesh->p[2]};
        pvec(f, "mesh_pose", pose, 7);
        fprintf(f, "mesh_tris %d\n", (int)mesh->triangles.size());
        for (auto& t : mesh->triangles) {
            fprintf(f, "tri %d %d %d %d\n", t.indices[0], t.indices[1], t.indices[2], t.coords.degenerate ? 1 : 0);
            const rprimitives::Material& mt = t.mat;
            float mv[26];
            for (int c = 0; c < 4; c++) {
                mv[c] = mt.Ke[c]; mv[4 + c] = mt.Ka[c]; mv[8 + c] = mt.Kd[c];
                mv[12 + c] = mt.Ks[c]; mv[16 + c] = mt.Kt[c]; mv[20 + c] = mt.Kr[c];
            }
            mv[24] = mt.alpha; mv[25] = mt.eta;
            pvec(f, "mat", mv, 26);
        }
    }
    fprintf(f, "n_trans %d\n", env.n_trans());
    for (int i = 0; i < env.n_trans(); i++) {
        renv::Transformation& t = env.get_trans()[i];
        float tv[7] = {t.o.i(), t.o.j(), t.o.k(), t.o.r(), t.p[0], t.p[1], t.p[2]};
        fprintf(f, "trans_mesh %d\n", t.get_hitable_idx());
        pvec(f, "trans", tv, 7);
    }
    auto& lights = scene->get_lights();
    fprintf(f, "n_lights %d\n", (int)lights.size());
    for (auto* l : lights) {
        if (auto* pl = dynamic_cast<rprimitives::cpu::PointLight*>(l)) {
            float lv[7] = {pl->color[0], pl->color[1], pl->color[2], pl->color[3], pl->pos[0], pl->pos[1], pl->pos[2]};
            pvec(f, "point_light", lv, 7);
        } else if (auto* dl = dynamic_cast<rprimitives::cpu::DirLight*>(l)) {
            float lv[7] = {dl->color[0], dl->color[1], dl->color[2], dl->color[3], dl->dir[0], dl->dir[1], dl->dir[2]};
            pvec(f, "dir_light", lv, 7);
        }
    }
    fclose(f);
}

struct IdRec {
    int32_t trans, tri;
    float t;
    int32_t trans_bf, tri_bf;
    float t_bf;
};

// Primary-ray closest hit ids. Follows the brute-force branch of Scene::cast_ray
// (src/rayenv/scene.cu:209-212); the "bvh" variant additionally requires the instance's leaf box
// (src/raytracer.cc:17-21) to pass BoundingBox::intersects, which is what BVH traversal reduces to.
static void dump_ids(renv::cpu::Scene* scene, const std::string& path, bool bvh_only) {
    renv::Environment& env = scene->get_environment();
    renv::Camera& cam = env.get_camera();
    int w = env.get_canvas().get_width(), h = env.get_canvas().get_height();
    int n = env.n_trans();
    renv::Transformation* trans = env.get_trans();
    auto& hitables = scene->get_hitables();
    std::vector<ropt::BoundingBox> boxes;
    for (int i = 0; i < n; i++) {
        rprimitives::cpu::Hitable& hh = *hitables[trans[i].get_hitable_idx()];
        boxes.push_back(ropt::from_local(hh.compute_bounding_box(scene), trans[i]));
    }
    std::vector<IdRec> recs((size_t)w * h);
    for (int y = 0; y < h; y++) {
        for (int x = 0; x < w; x++) {
            rmath::Ray<float> r = cam.at(x, y);
            IdRec rec{-1, -1, INFINITY, -1, -1, INFINITY};
            for (int pass = 0; pass < (bvh_only ? 1 : 2); pass++) {
                float time = INFINITY;
                rprimitives::Isect isect{time};
                int best = -1, best_tri = -1;
                for (int i = 0; i < n; i++) {
                    if (pass == 0) {
                        float bt;
                        if (!boxes[i].intersects(r, bt)) continue;
                    }
                    if (scene->cast_local(r, isect, trans[i])) {
                        best = i;
                        rprimitives::cpu::Trimesh* mesh = (rprimitives::cpu::Trimesh*)hitables[trans[i].get_hitable_idx()];
                        const char* base = (const char*)&mesh->triangles[0].mat;
                        best_tri = (int)(((const char*)isect.mat - base) / (ptrdiff_t)sizeof(rprimitives::cpu::TriInner));
                    }
                }
                if (pass == 0) { rec.trans = best; rec.tri = best_tri; rec.t = time; }
                else { rec.trans_bf = best; rec.tri_bf = best_tri; rec.t_bf = time; }
            }
            recs[(size_t)y * w + x] = rec;
        }
    }
    FILE* f = fopen(path.c_str(), "wb");
    fwrite(&w, 4, 1, f);
    fwrite(&h, 4, 1, f);
    fwrite(recs.data(), sizeof(IdRec), recs.size(), f);
    fclose(f);
}

// --moves: per-frame motion through the reference's own Entity API before the frame is traced, in the
// arithmetic of its event loop (src/main.cc:19-20,140-177). Comma-separated operations:
//   w s a d                  camera key presses: cam.translate(+-MOVE_SPEED along local z / x)
//   m:xrel:yrel              SDL_MOUSEMOTION: cam.rotate(Quat(up, ROT_SPEED*rx) * Quat(right, ROT_SPEED*ry))
//   g:idx:dx:dy:dz           trans[idx].translate_global(dp)      (include/rayprimitives/entity.h:49-52)
//   l:idx:dx:dy:dz           trans[idx].translate(dp)             (entity.h:54-57)
//   r:idx:ax:ay:az:theta     trans[idx].rotate(Quat(axis, theta)) (entity.h:64-67)
static void apply_moves(renv::Environment& env, const std::string& spec) {
    const float MOVE_SPEED = 0.2f, ROT_SPEED = 0.01f;
    size_t pos = 0;
    while (pos < spec.size()) {
        size_t end = spec.find(',', pos);
        if (end == std::string::npos) end = spec.size();
        std::string op = spec.substr(pos, end - pos);
        pos = end + 1;
        if (op.empty()) continue;
        std::vector<std::string> f;
        size_t q = 0;
        while (q <= op.size()) {
            size_t e = op.find(':', q);
            if (e == std::string::npos) e = op.size();
            f.push_back(op.substr(q, e - q));
            q = e + 1;
        }
        renv::Camera& cam = env.get_camera();
        auto num = [&](size_t i) { return (float)atof(f[i].c_str()); };
        if (f[0] == "w") cam.translate(rmath::Vec3<float>{0, 0, MOVE_SPEED});
        else if (f[0] == "s") cam.translate(rmath::Vec3<float>{0, 0, -MOVE_SPEED});
        else if (f[0] == "a") cam.translate(rmath::Vec3<float>{-MOVE_SPEED, 0, 0});
        else if (f[0] == "d") cam.translate(rmath::Vec3<float>{MOVE_SPEED, 0, 0});
        else if (f[0] == "m" && f.size() == 3) {
            rmath::Vec<float, 2> rel_mot = rmath::Vec<float, 2>({num(1), num(2)}).normalized();
            rmath::Quat<float> rot = rmath::Quat<float>(cam.up().direction(), ROT_SPEED * rel_mot[0])
                                   * rmath::Quat<float>(cam.right().direction(), ROT_SPEED * rel_mot[1]);
            cam.rotate(rot);
        } else if ((f[0] == "g" || f[0] == "l") && f.size() == 5) {
            renv::Transformation& t = env.get_trans()[atoi(f[1].c_str())];
            rmath::Vec3<float> dp({num(2), num(3), num(4)});
            if (f[0] == "g") t.translate_global(dp); else t.translate(dp);
        } else if (f[0] == "r" && f.size() == 6) {
            renv::Transformation& t = env.get_trans()[atoi(f[1].c_str())];
            t.rotate(rmath::Quat<float>(rmath::Vec3<float>({num(2), num(3), num(4)}), num(5)));
        } else {
            fprintf(stderr, "bad --moves operation '%s'\n", op.c_str());
            exit(2);
        }
    }
}

int main(int argc, char** argv) {
    if (argc < 2) {
        fprintf(stderr, "usage: %s world.json [--gpu] [--no-bvh] [--frames N] [--kd D] [--canvas f] [--ids f] [--scene f] [--moves ops]\n", argv[0]);
        return 2;
    }
    std::string json = argv[1], canvas_out, ids_out, scene_out, moves;
    bool gpu = false, optimize = true, ids_bvh_only = false;
    int frames = 1, kd = 16;
    for (int i = 2; i < argc; i++) {
        std::string a = argv[i];
        if (a == "--gpu") gpu = true;
        else if (a == "--no-bvh") optimize = false;
        else if (a == "--ids-bvh-only") ids_bvh_only = true;
        else if (a == "--frames" && i + 1 < argc) frames = atoi(argv[++i]);
        else if (a == "--kd" && i + 1 < argc) kd = atoi(argv[++i]);
        else if (a == "--canvas" && i + 1 < argc) canvas_out = argv[++i];
        else if (a == "--ids" && i + 1 < argc) ids_out = argv[++i];
        else if (a == "--scene" && i + 1 < argc) scene_out = argv[++i];
        else if (a == "--moves" && i + 1 < argc) moves = argv[++i];
    }
    std::vector<double> ms;
    int nonzero = 0;
    uint64_t hash = 0;
    int n_trans = 0, w = 0, h = 0;
    if (!gpu) {
        renv::cpu::Scene* scene = procedural::cpu::generate(json);
        n_trans = scene->get_environment().n_trans();
        w = scene->get_environment().get_canvas().get_width();
        h = scene->get_environment().get_canvas().get_height();
        if (!moves.empty()) apply_moves(scene->get_environment(), moves);
        if (!scene_out.empty()) dump_scene(scene, scene_out);
        for (int f = 0; f < frames; f++) {
            auto t0 = std::chrono::high_resolution_clock::now();
            rtracer::cpu::update_scene(scene, kd, optimize);
            auto t1 = std::chrono::high_resolution_clock::now();
            ms.push_back(std::chrono::duration<double, std::milli>(t1 - t0).count());
        }
        hash = canvas_hash(scene->get_environment().get_canvas(), nonzero);
        if (!canvas_out.empty()) dump_canvas(scene->get_environment().get_canvas(), canvas_out);
        if (!ids_out.empty()) dump_ids(scene, ids_out, ids_bvh_only);
    } else {
#ifdef REF_INTERPOSE
        fprintf(stderr, "this binary interposes the CUDA allocator; use ref_gpu for --gpu\n");
        return 2;
#else
        renv::gpu::Scene* scene = procedural::gpu::generate(json);
        cudaDeviceSynchronize();
        n_trans = scene->get_environment().n_trans();
        w = scene->get_environment().get_canvas().get_width();
        h = scene->get_environment().get_canvas().get_height();
        if (!moves.empty()) apply_moves(scene->get_environment(), moves);
        for (int f = 0; f < frames; f++) {
            auto t0 = std::chrono::high_resolution_clock::now();
            rtracer::gpu::update_scene(scene, kd, optimize);
            auto t1 = std::chrono::high_resolution_clock::now();
            ms.push_back(std::chrono::duration<double, std::milli>(t1 - t0).count());
        }
        cudaError_t err = cudaDeviceSynchronize();
        if (err != cudaSuccess) {
            fprintf(stderr, "cuda error: %s\n", cudaGetErrorString(err));
            return 1;
        }
        hash = canvas_hash(scene->get_environment().get_canvas(), nonzero);
        if (!canvas_out.empty()) dump_canvas(scene->get_environment().get_canvas(), canvas_out);
#endif
    }
    if (ms.empty()) ms.push_back(0.0);
    std::vector<double> sorted = ms;
    std::sort(sorted.begin(), sorted.end());
    printf("{\"impl\": \"%s\", \"scene\": \"%s\", \"width\": %d, \"height\": %d, \"n_trans\": %d, \"bvh\": %s, "
           "\"frames\": %d, \"ms_first\": %.4f, \"ms_min\": %.4f, \"ms_median\": %.4f, \"hash\": \"%016llx\", \"nonzero\": %d}\n",
           gpu ? "ref_gpu" : "ref_cpu", json.c_str(), w, h, n_trans, optimize ? "true" : "false", frames, ms[0],
           sorted[0], sorted[sorted.size() / 2], (unsigned long long)hash, nonzero);
    return 0;
}
