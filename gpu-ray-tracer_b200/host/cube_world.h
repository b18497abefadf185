// cube_world.h — world*.json loader: mirror of the reference's procedural::{gpu,cpu}::generate
// (/root/reference/include/procedural/cube_world.h:20-27, src/procedural/cube_world.cc:38-224).
// Same JSON schema and defaults; output is a flat scene instead of a renv::Scene.
#ifndef RTB200_CUBE_WORLD_H
#define RTB200_CUBE_WORLD_H

#include <string>

#include "scene_builder.h"

namespace procedural {

// Optional key overrides, equivalent to editing the JSON file (<= 0 / < 0 means "keep").
struct WorldOverrides {
    int width = 0, height = 0, depth = -1, grid_size = 0;
    std::string base_dir;  // where relative "atlas_ppm" paths are looked up; generate() fills in the scene file's directory
};

// Throws std::runtime_error on unreadable / malformed input.
rtracer::FlatScene generate_from_text(const std::string& json_text, const WorldOverrides& ov = WorldOverrides());
rtracer::FlatScene generate(const std::string& config_path, const WorldOverrides& ov = WorldOverrides());

}  // namespace procedural
#endif
