// ReferenceBridge.h -- the reference-side half of the descriptor bridge.
//
// The reference turns BSDFs, lights, camera and technique into generated Artic source (src/runtime/shader/ShaderUtils.cpp:95-122,
// src/runtime/loader/LoaderLight.cpp:106-240, src/runtime/technique/PathTechnique.cpp:36-79); this backend takes them as the POD
// tables of include/ignis_b200.h.  Inside the reference tree (IG_B200_IN_REFERENCE_TREE) this file's function produces those
// tables from the reference's own parsed `IG::Scene` and `IG::LoaderContext`, in the REFERENCE'S numbering of entities,
// materials and lights, by re-expressing the scene in the document model of host/Json.h and running the table builders of
// host/SceneLoader.cpp on it.  Outside the reference tree the file compiles to nothing: the frontends of this directory
// load scene JSON through SceneLoader.cpp directly.
#pragma once
#ifdef IG_B200_IN_REFERENCE_TREE

#include "SceneFile.h"

namespace IG {
class LoaderContext;
/// Call after Loader::load has filled `ctx` (Runtime::load, src/runtime/Runtime.cpp:168-292), before Device::assignScene:
///     ShadingTables tables; if (!describeSceneForB200(ctx, tables, err)) ...; mDevice->assignShading(tables.view());
/// Fails (false + message) when the scene uses something outside the backend's subset or when the numbering derived here
/// disagrees with ctx (entity / material counts), so a mismatch can never render silently.
bool describeSceneForB200(const LoaderContext& ctx, ShadingTables& out, std::string& error);
} // namespace IG

#endif
