#include "cube_world.h"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <fstream>
#include <sstream>
#include <stdexcept>

#include "mini_json.h"
#include "perlin.h"

namespace procedural {

namespace {
// defaults: src/procedural/cube_world.cc:15-21
const int kDefaultSeed = 42;
const int kDefaultGridSize = 8;
const int kDefaultWidth = 640;
const int kDefaultHeight = 480;
const float kDefaultFov = M_PI / 4;
const float kDefaultUnitLen = 200;
const float kDefaultAmplitude = 1.0f;

std::vector<float> read_ppm(const std::string& path, int& w, int& h) {
    std::ifstream f(path.c_str(), std::ios::binary);
    if (!f) throw std::runtime_error("cannot open atlas: " + path);
    std::string magic;
    int maxv = 0;
    f >> magic >> w >> h >> maxv;
    if (magic != "P6" || w <= 0 || h <= 0 || maxv != 255) throw std::runtime_error("atlas is not a binary PPM with maxval 255: " + path);
    f.get();  // the single whitespace after the header
    std::vector<unsigned char> raw((size_t)w * h * 3);
    f.read((char*)raw.data(), (std::streamsize)raw.size());
    if ((size_t)f.gcount() != raw.size()) throw std::runtime_error("atlas is truncated: " + path);
    std::vector<float> rgba((size_t)w * h * 4);
    const float inv255 = 1.0f / UINT8_MAX;  // Texture::color_to_vec (src/rayprimitives/texture.cu:14-16)
    for (size_t i = 0; i < (size_t)w * h; i++) {
        for (int k = 0; k < 3; k++) rgba[4 * i + k] = inv255 * (float)raw[3 * i + k];
        rgba[4 * i + 3] = inv255 * 255.0f;
    }
    return rgba;
}

rtb::Float3 read_vec3(const mini_json::Value& v) { return {v.at(0).as_float(), v.at(1).as_float(), v.at(2).as_float()}; }
rtb::Float4 read_vec4(const mini_json::Value& v) {
    return {v.at(0).as_float(), v.at(1).as_float(), v.at(2).as_float(), v.at(3).as_float()};
}

const float kInv255 = 1.0f / UINT8_MAX;

// Top-level scalars of a world file (src/procedural/cube_world.cc:39-61,141-144), then the caller's overrides.
struct WorldSettings {
    int seed = kDefaultSeed, grid_size = kDefaultGridSize, width = kDefaultWidth, height = kDefaultHeight;
    float fov = kDefaultFov, unit_length = kDefaultUnitLen, amplitude = kDefaultAmplitude;
    WorldSettings(const mini_json::Value& doc, const WorldOverrides& ov) {
        struct { const char* key; int* dst; } ints[] = {{"seed", &seed}, {"grid_size", &grid_size}, {"width", &width}, {"height", &height}};
        for (auto& e : ints) if (doc.has(e.key)) *e.dst = doc.at(e.key).as_int();
        if (doc.has("fov")) fov = (float)(doc.at("fov").as_double() * M_PI) / 180;  // degrees -> radians (:57)
        if (doc.has("unit_length")) unit_length = (float)doc.at("unit_length").as_double();
        if (doc.has("amplitude")) amplitude = (float)doc.at("amplitude").as_double();
        if (ov.width > 0) width = ov.width;
        if (ov.height > 0) height = ov.height;
        if (ov.grid_size > 0) grid_size = ov.grid_size;
    }
};

// Material keys of a "cubes" (or "spheres") entry (:82-105): the four colour terms are 0-255 in the file, Kt / Kr are not.
rtracer::Material read_material(const mini_json::Value& e) {
    rtracer::Material mat;
    if (e.has("Ke")) mat.set_Ke(kInv255 * read_vec4(e.at("Ke")));
    if (e.has("Ka")) mat.set_Ka(kInv255 * read_vec4(e.at("Ka")));
    if (e.has("Kd")) mat.set_Kd(kInv255 * read_vec4(e.at("Kd")));
    if (e.has("Ks")) mat.set_Ks(kInv255 * read_vec4(e.at("Ks")));
    if (e.has("Kt")) mat.set_Kt(read_vec4(e.at("Kt")));
    if (e.has("Kr")) mat.set_Kr(read_vec4(e.at("Kr")));
    if (e.has("alpha")) mat.set_alpha((float)e.at("alpha").as_double());
    if (e.has("eta")) mat.set_eta((float)e.at("eta").as_double());
    return mat;
}

// The terrain (:145-170): for every cube mesh a fresh height field with the SAME seed (so every layer has the same relief),
// sampled on the integer grid; floor(0.5 (noise + amplitude)) + 1 cubes of that mesh are stacked on the column, on top of
// what the meshes before it left there. Instances are created in mesh / row / column / bottom-up order — the order is
// the instance id. Returns the highest column.
float stack_terrain(rtracer::SceneBuilder& builder, const WorldSettings& ws, int n_meshes) {
    const int g = ws.grid_size;
    std::vector<float> column_top((size_t)g * g, 0.0f);
    float highest = 0.0f;
    for (int mesh = 0; mesh < n_meshes; mesh++) {
        Perlin relief(ws.seed, (g + 4) / 5);
        relief.set_amplitude(ws.amplitude);
        relief.set_period((float)g);
        for (int row = 0; row < g; row++) {
            for (int col = 0; col < g; col++) {
                float& top = column_top[(size_t)row * g + col];
                const float x = row - g / 2.0f, z = col - g / 2.0f;
                const float layers = std::floor(0.5f * (relief.sample((float)row, (float)col, 0.0f) + ws.amplitude)) + 1;
                for (int k = 0; k < layers; k++) {
                    const int id = builder.add_trans(builder.get_mesh_builder(mesh));
                    builder.get_transformation(id).set_position({x, top + k, z});
                }
                top += layers;
                highest = std::max(highest, top);
            }
        }
    }
    return highest;
}
}  // namespace

rtracer::FlatScene generate_from_text(const std::string& json_text, const WorldOverrides& ov) {
    std::shared_ptr<mini_json::Value> root = mini_json::parse_text(json_text);
    const mini_json::Value& doc = *root;
    if (!doc.is_object()) throw std::runtime_error("world json: top level is not an object");

    const WorldSettings ws(doc, ov);
    const int grid_size = ws.grid_size, width = ws.width, height = ws.height;

    (void)doc.at("atlas").as_string();  // required key in the reference (:64); the atlas is never sampled
    rtracer::Camera cam(ws.fov, ws.unit_length, width, height);
    rtracer::SceneBuilder builder;

    // cube meshes, one per "cubes" entry (:70-114)
    int n_cubes = 0;
    if (doc.has("cubes")) {
        const mini_json::Value& cubes = doc.at("cubes");
        n_cubes = (int)cubes.size();
        for (int i = 0; i < n_cubes; i++) {
            const mini_json::Value& cube = cubes.at((size_t)i);
            const rtracer::Material mat = read_material(cube);
            // "texture": [tile_x, tile_y, tile_size] — schema extension: every face shows that atlas tile
            if (cube.has("texture")) {
                const mini_json::Value& t = cube.at("texture");
                builder.build_textured_cube(.999f, t.at(0).as_int(), t.at(1).as_int(), t.at(2).as_int(), mat);
            } else {
                builder.build_cube(.999f, rtracer::TextureCoords(), mat);  // degenerate coordinates: texture mapping is a TODO in the reference (:79-81)
            }
        }
    }
    // "atlas_ppm": binary PPM (P6, maxval 255) next to the scene file or at an absolute path — the reference reads a PNG
    // through libpng (src/assets.cc), which this image does not have; texels become RGBA floats c / 255, alpha 1
    if (doc.has("atlas_ppm")) {
        std::string path = doc.at("atlas_ppm").as_string();
        if (!path.empty() && path[0] != '/' && !ov.base_dir.empty()) path = ov.base_dir + "/" + path;
        int aw = 0, ah = 0;
        std::vector<float> rgba = read_ppm(path, aw, ah);
        builder.set_atlas(std::move(rgba), aw, ah);
    }

    // "spheres": schema extension with no reference counterpart: analytic sphere hitables, each
    // {"radius": r, "pos": [x, y, z], material keys as for cubes}, placed at `pos` after the cubes' transformations.
    struct SphereAt { int hitable; rtb::Float3 pos; };
    std::vector<SphereAt> spheres;
    if (doc.has("spheres")) {
        const mini_json::Value& sp = doc.at("spheres");
        for (size_t i = 0; i < sp.size(); i++) {
            const mini_json::Value& e = sp.at(i);
            const rtracer::Material mat = read_material(e);
            rtracer::TextureCoords coords;
            spheres.push_back({builder.build_sphere((float)e.at("radius").as_double(), coords, mat), read_vec3(e.at("pos"))});
        }
    }

    // lights (:116-138)
    if (doc.has("lights")) {
        const mini_json::Value& lights = doc.at("lights");
        if (lights.has("directional")) {
            const mini_json::Value& d = lights.at("directional");
            for (size_t i = 0; i < d.size(); i++)
                builder.add_directional_light(read_vec3(d.at(i).at("dir")), kInv255 * read_vec4(d.at(i).at("col")));
        }
        if (lights.has("point")) {
            const mini_json::Value& p = lights.at("point");
            for (size_t i = 0; i < p.size(); i++)
                builder.add_point_light(read_vec3(p.at(i).at("pos")), kInv255 * read_vec4(p.at(i).at("col")));
        }
    }

    const float max_height = stack_terrain(builder, ws, n_cubes);

    for (const SphereAt& sa : spheres) {
        int tid = builder.add_trans(builder.get_mesh_builder(sa.hitable));
        builder.get_transformation(tid).set_position(sa.pos);
    }

    // camera pose (:172-173): note the angle is 45 RADIANS about +x
    cam.set_position({0.0f, max_height + 10.0f, -(float)grid_size / 2});
    cam.set_orientation(rtb::quat_axis_angle({1.0f, 0.0f, 0.0f}, 45));

    rtracer::FlatScene fs = builder.build_flat_scene(width, height, cam);

    // finish_env (:177-191)
    if (doc.has("ambience")) {
        rtb::Float4 a = read_vec4(doc.at("ambience"));
        fs.env.ambience[0] = a.x; fs.env.ambience[1] = a.y; fs.env.ambience[2] = a.z; fs.env.ambience[3] = a.w;
    }
    if (doc.has("depth")) fs.env.depth = doc.at("depth").as_int();
    if (ov.depth >= 0) fs.env.depth = ov.depth;
    if (doc.has("distance_attenuation")) {
        const mini_json::Value& a = doc.at("distance_attenuation");
        fs.env.dist_atten[0] = (float)a.at("constant_term").as_double();
        fs.env.dist_atten[1] = (float)a.at("linear_term").as_double();
        fs.env.dist_atten[2] = (float)a.at("quadratic_term").as_double();
    }
    return fs;
}

rtracer::FlatScene generate(const std::string& config_path, const WorldOverrides& ov) {
    std::ifstream ifs(config_path.c_str());
    if (!ifs) throw std::runtime_error("cannot open scene file: " + config_path);
    std::stringstream ss;
    ss << ifs.rdbuf();
    WorldOverrides o2 = ov;
    if (o2.base_dir.empty()) {
        size_t slash = config_path.find_last_of('/');
        o2.base_dir = slash == std::string::npos ? "." : config_path.substr(0, slash);
    }
    return generate_from_text(ss.str(), o2);
}

}  // namespace procedural
