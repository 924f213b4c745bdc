// SceneFile.h -- `.igdb`: the loader's OUTPUT (SceneDatabase blobs + the POD shading tables of include/ignis_b200.h)
// in one file, so the frontends of this directory run where the reference's loader (JSON parser, mesh readers,
// Artic generators: out of scope) does not exist.  Written by ignis-masterthesis_b200/fixtures.py
// (`save_igdb`) from the tables `sceneio.load_scene` builds off the reference's own scene files.
//
// Layout: "IGB200DB" | u32 version (1) | u32 section_count | section_count x { char name[24]; u64 bytes; u64 offset }
// | payloads at 16-byte aligned offsets.  Sections: lookups, shape_data, entities, leaves0 (trimesh EntityLeaf1),
// leaves1 (sphere), bbox (6 f32), material_count (u64), entity_per_material (i32[]), camera, technique, bsdfs,
// materials, infinite, finite, textures, blob, selector (2 x u32), film (u32 width, height).
#pragma once

#include "Device.h"

#include <string>
#include <vector>

namespace IG {
struct ShadingTables {
    igb200_camera                camera{};
    igb200_technique             technique{};
    std::vector<igb200_bsdf>     bsdfs;
    std::vector<igb200_material> materials;
    std::vector<igb200_light>    infinite, finite;
    std::vector<igb200_texture>  textures;
    std::vector<uint8_t>         blob;
    uint32_t                     selector = 0, hierarchy = 0;
    // view over the vectors above (valid while this object is alive and unmodified)
    igb200_shading view() const;
};

struct SceneFile {
    SceneDatabase      database;
    ShadingTables      shading;
    std::vector<int32> entity_per_material;
    size_t             width = 0, height = 0;
};

// false + message on a malformed file (the reference's loader logs and returns false, Runtime.cpp:157-166)
bool loadSceneFile(const std::string& path, SceneFile& out, std::string& error);
} // namespace IG
