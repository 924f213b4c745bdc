// SceneLoader.h -- scene ingest for the C++ frontends: the reference's scene JSON -> SceneDatabase blobs + POD shading
// tables, i.e. what `.igdb` files carry, built directly from the scene description.
//
// Follows, for the supported subset, src/runtime/loader/Parser.cpp (JSON dialect, transforms, externals),
// src/runtime/shape/TriMeshProvider.cpp:19-299,480-617 + src/runtime/mesh/TriMesh.cpp (procedural shapes, OBJ / PLY,
// normals, subdivision, DynTable layout), src/runtime/shape/SphereProvider.cpp, src/runtime/loader/LoaderEntity.cpp:32-205
// (material groups, entity rows, EntityLeaf1), src/runtime/loader/LoaderLight.cpp + src/runtime/light/*.cpp (light
// tables, light selectors), src/runtime/bsdf/*.cpp, src/runtime/pattern/*.cpp, src/runtime/camera/*.cpp and
// src/runtime/technique/*.cpp (the descriptor side: what the reference bakes into generated shader source travels here
// as the igb200_* tables of include/ignis_b200.h).
// Subset: everything BASELINE.json's scenes and the evaluation scenes pinned in tests/golden use; anything else is
// refused with a message naming the element (the reference logs an error and skips or aborts likewise).
#pragma once

#include "SceneFile.h"

namespace IG {
namespace json {
struct Value;
}
/// The loader proper, on an already parsed document (host/Json.h) whose externals have been merged; `base_dir` resolves
/// relative file names.  This is the entry the reference-side bridge (host/ReferenceBridge.cpp) uses.
bool loadSceneDocument(const json::Value& scene, const std::string& base_dir, SceneFile& out, std::string& error, size_t width = 0, size_t height = 0);
/// Loads `path` (scene JSON).  width / height override the film size when non-zero (igcli --width / --height).
bool loadSceneJson(const std::string& path, SceneFile& out, std::string& error, size_t width = 0, size_t height = 0);
/// Writes the container host/SceneFile.h describes.
bool saveSceneFile(const std::string& path, const SceneFile& scene, std::string& error);
/// `.json` -> loadSceneJson, anything else -> loadSceneFile.
bool loadAnyScene(const std::string& path, SceneFile& out, std::string& error, size_t width = 0, size_t height = 0);
} // namespace IG
