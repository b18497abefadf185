// main.cc — rtb200_cli: the reference's command line (/root/reference/src/main.cc:31-79) over the
// B200 engine. Same flags: -c/--config <json> (required), -b/--bench, -r/--unoptimize,
// -s/--serial, -d/--dim <int>. Differences, all forced by the environment:
//   * there is no window (no display stack): without -b the program renders one frame too;
//     -o <file.ppm> writes the canvas, --hash prints the canvas hash used by the parity fixtures;
//   * -s selects the SERIAL PATH'S BOUNCE SEMANTICS (renv::cpu::Scene::propagate_helper) but still
//     runs on the GPU — this program has no CPU renderer;
//   * --frames N repeats the frame (the reference times exactly one, cold).
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "cube_world.h"
#include "rtb200.h"

static void usage() {
    std::cerr << "usage: rtb200_cli -c world.json [-b] [-r] [-s] [-d dim] [--frames N] [--width W --height H --depth D]"
                 " [-o out.ppm] [--hash]\n";
}

int main(int argc, char** argv) {
    std::string config, out_ppm;
    bool bench = false, unoptimize = false, serial = false, hash = false;
    int dim = 16, frames = 1;
    procedural::WorldOverrides ov;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto next = [&]() -> const char* { return i + 1 < argc ? argv[++i] : nullptr; };
        if (a == "-c" || a == "--config") { const char* v = next(); if (v) config = v; }
        else if (a == "-b" || a == "--bench") bench = true;
        else if (a == "-r" || a == "--unoptimize") unoptimize = true;
        else if (a == "-s" || a == "--serial") serial = true;
        else if (a == "-d" || a == "--dim") { const char* v = next(); if (v) dim = atoi(v); }
        else if (a == "--frames") { const char* v = next(); if (v) frames = atoi(v); }
        else if (a == "--width") { const char* v = next(); if (v) ov.width = atoi(v); }
        else if (a == "--height") { const char* v = next(); if (v) ov.height = atoi(v); }
        else if (a == "--depth") { const char* v = next(); if (v) ov.depth = atoi(v); }
        else if (a == "-o") { const char* v = next(); if (v) out_ppm = v; }
        else if (a == "--hash") hash = true;
        else { usage(); return 2; }
    }
    if (config.empty()) { usage(); return 2; }
    rtracer::FlatScene fs;
    try {
        fs = procedural::generate(config, ov);
    } catch (const std::exception& e) {
        std::cerr << "cannot load " << config << ": " << e.what() << "\n";
        return 1;
    }
    RtSceneDesc desc = fs.desc();
    RtScene* scene = nullptr;
    if (rt_scene_create(&desc, &scene) != RT_OK) {
        std::cerr << "rt_scene_create: " << rt_last_error_string() << "\n";
        return 1;
    }
    std::cout << "Loaded scene" << std::endl;
    std::vector<uint32_t> canvas((size_t)fs.width * fs.height);
    RtRenderParams p;
    std::memset(&p, 0, sizeof p);
    p.flags = unoptimize ? RT_RENDER_NO_BVH : RT_RENDER_REBUILD_BVH;  // update_scene(scene, dim, !r)
    p.semantics = serial ? RT_SEMANTICS_CPU : RT_SEMANTICS_GPU;
    (void)dim;
    for (int f = 0; f < frames; f++) {
        RtFrameStats st;
        auto from = std::chrono::high_resolution_clock::now();
        int rc = rt_render(scene, &p, canvas.data(), &st);
        auto to = std::chrono::high_resolution_clock::now();
        if (rc != RT_OK) {
            std::cerr << "rt_render: " << rt_last_error_string() << "\n";
            return 1;
        }
        std::chrono::duration<double, std::milli> elapsed = to - from;
        if (bench || frames > 1)
            std::cout << "Time: " << elapsed.count() << " ms" << "  (device: build " << st.ms_build << " ms, trace "
                      << st.ms_trace << " ms, " << st.rays_total << " rays)" << std::endl;
    }
    if (hash) {
        uint64_t h = 0;
        int nz = 0;
        for (uint32_t px : canvas) { h = h * 1315423911ull + px; if (px) nz++; }
        printf("hash %016llx nonzero %d\n", (unsigned long long)h, nz);
    }
    if (!out_ppm.empty()) {
        FILE* f = fopen(out_ppm.c_str(), "wb");
        if (!f) { std::cerr << "cannot write " << out_ppm << "\n"; return 1; }
        fprintf(f, "P6\n%d %d\n255\n", fs.width, fs.height);
        for (uint32_t px : canvas) {
            unsigned char rgb[3] = {(unsigned char)(px >> 24), (unsigned char)(px >> 16), (unsigned char)(px >> 8)};
            fwrite(rgb, 1, 3, f);
        }
        fclose(f);
    }
    rt_scene_destroy(scene);
    return 0;
}
