// ReferenceBridge.cpp -- see ReferenceBridge.h.  Compiled only inside the reference tree.
#ifdef IG_B200_IN_REFERENCE_TREE
#include "ReferenceBridge.h"

#include "Json.h"
#include "SceneLoader.h"

#include "Scene.h"
#include "loader/LoaderContext.h"
#include "loader/LoaderEntity.h"
#include "loader/LoaderLight.h"

namespace IG {
namespace {
using json::Value;

Value number(double x)
{
    Value v;
    v.type   = Value::Number;
    v.number = x;
    return v;
}
Value text(const std::string& s)
{
    Value v;
    v.type   = Value::String;
    v.string = s;
    return v;
}

// One SceneProperty (src/runtime/SceneProperty.h:17-27) as a document value.  Transforms travel as the 16 numbers of the
// row-major matrix, which Parser.cpp:124-243 and SceneLoader.cpp's parseTransform read alike.
Value fromProperty(const SceneProperty& p)
{
    Value v;
    switch (p.type()) {
    case SceneProperty::PT_BOOL:
        v.type    = Value::Bool;
        v.boolean = p.getBool();
        break;
    case SceneProperty::PT_INTEGER: v = number((double)p.getInteger()); break;
    case SceneProperty::PT_NUMBER: v = number((double)p.getNumber()); break;
    case SceneProperty::PT_STRING: v = text(p.getString()); break;
    case SceneProperty::PT_VECTOR2: {
        v.type = Value::Array;
        for (int i = 0; i < 2; ++i) v.array.push_back(number((double)p.getVector2()(i)));
        break;
    }
    case SceneProperty::PT_VECTOR3: {
        v.type = Value::Array;
        for (int i = 0; i < 3; ++i) v.array.push_back(number((double)p.getVector3()(i)));
        break;
    }
    case SceneProperty::PT_TRANSFORM: {
        v.type        = Value::Array;
        const auto& m = p.getTransform().matrix();
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) v.array.push_back(number((double)m(r, c)));
        break;
    }
    case SceneProperty::PT_INTEGER_ARRAY: {
        v.type = Value::Array;
        for (auto x : p.getIntegerArray()) v.array.push_back(number((double)x));
        break;
    }
    case SceneProperty::PT_NUMBER_ARRAY: {
        v.type = Value::Array;
        for (auto x : p.getNumberArray()) v.array.push_back(number((double)x));
        break;
    }
    default: break; // PT_NONE -> null
    }
    return v;
}

Value fromObject(const std::string& name, const SceneObject& obj)
{
    Value v;
    v.type = Value::Object;
    v.members.emplace_back("type", text(obj.pluginType()));
    if (!name.empty()) v.members.emplace_back("name", text(name));
    for (const auto& kv : obj.properties()) {
        // file names are relative to the object's base directory (LoaderContext::handlePath): make them absolute here
        if (kv.first == "filename" && kv.second.type() == SceneProperty::PT_STRING) {
            const Path file = kv.second.getString();
            v.members.emplace_back(kv.first, text(file.is_absolute() ? file.u8string() : (obj.baseDir() / file).u8string()));
        } else {
            v.members.emplace_back(kv.first, fromProperty(kv.second));
        }
    }
    return v;
}

template <typename Map>
Value fromMap(const Map& objects)
{
    // ITERATION ORDER of the reference's containers, on purpose: LoaderEntity::load (LoaderEntity.cpp:42-103) and
    // LoaderLight::prepare walk these same unordered_maps, and entity / material / light ids follow from that walk.
    Value v;
    v.type = Value::Array;
    for (const auto& kv : objects) v.array.push_back(fromObject(kv.first, *kv.second));
    return v;
}
} // namespace

bool describeSceneForB200(const LoaderContext& ctx, ShadingTables& out, std::string& error)
{
    const Scene* scene = ctx.Options.Scene;
    if (!scene) {
        error = "no scene";
        return false;
    }
    Value doc;
    doc.type = Value::Object;
    if (scene->technique()) doc.members.emplace_back("technique", fromObject("", *scene->technique()));
    if (scene->camera()) doc.members.emplace_back("camera", fromObject("", *scene->camera()));
    {
        Value film, size;
        film.type = Value::Object;
        size.type = Value::Array;
        size.array.push_back(number((double)ctx.Options.FilmWidth));
        size.array.push_back(number((double)ctx.Options.FilmHeight));
        film.members.emplace_back("size", size);
        doc.members.emplace_back("film", film);
    }
    doc.members.emplace_back("textures", fromMap(scene->textures()));
    doc.members.emplace_back("bsdfs", fromMap(scene->bsdfs()));
    doc.members.emplace_back("shapes", fromMap(scene->shapes()));
    doc.members.emplace_back("entities", fromMap(scene->entities()));
    doc.members.emplace_back("lights", fromMap(scene->lights()));

    // file names were made absolute in fromObject, so the base directory handed over here is never consulted
    SceneFile built;
    if (!loadSceneDocument(doc, ".", built, error, ctx.Options.FilmWidth, ctx.Options.FilmHeight)) return false;

    // The numbering derived here must be the reference's: same walk, same grouping rule.  Check what can be checked.
    const size_t entities = built.database.FixTables.at("entities").entryCount();
    if (entities != ctx.Entities->entityCount() || built.database.MaterialCount != ctx.Materials.size()) {
        error = "descriptor bridge: entity / material numbering differs from the loader's (" + std::to_string(entities) + " / "
                + std::to_string(built.database.MaterialCount) + " against " + std::to_string(ctx.Entities->entityCount()) + " / "
                + std::to_string(ctx.Materials.size()) + ")";
        return false;
    }
    for (size_t i = 0; i < ctx.Materials.size(); ++i) {
        const bool emissive = built.shading.materials[i].light >= 0;
        if (emissive != ctx.Materials[i].hasEmission()) {
            error = "descriptor bridge: material " + std::to_string(i) + " disagrees on emission";
            return false;
        }
    }
    if (built.shading.finite.size() != ctx.Lights->finiteLightCount() || built.shading.infinite.size() != ctx.Lights->infiniteLightCount()) {
        error = "descriptor bridge: light tables differ in size from the loader's";
        return false;
    }
    out = std::move(built.shading);
    return true;
}
} // namespace IG
#endif // IG_B200_IN_REFERENCE_TREE
