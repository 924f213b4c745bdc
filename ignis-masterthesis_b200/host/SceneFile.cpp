#include "SceneFile.h"

#include <cstdio>
#include <cstring>
#include <map>
#include <new>
#include <sys/stat.h>

namespace IG {
igb200_shading ShadingTables::view() const
{
    igb200_shading sh{};
    sh.camera               = camera;
    sh.technique            = technique;
    sh.bsdfs                = bsdfs.data();
    sh.bsdf_count           = bsdfs.size();
    sh.materials            = materials.data();
    sh.material_count       = materials.size();
    sh.infinite_lights      = infinite.data();
    sh.infinite_light_count = infinite.size();
    sh.finite_lights        = finite.data();
    sh.finite_light_count   = finite.size();
    sh.textures             = textures.data();
    sh.texture_count        = textures.size();
    sh.blob                 = blob.data();
    sh.blob_bytes           = blob.size();
    sh.selector             = selector;
    sh.hierarchy            = hierarchy;
    return sh;
}

namespace {
struct TocEntry {
    char     name[24];
    uint64_t bytes;
    uint64_t offset;
};
static_assert(sizeof(TocEntry) == 40, "toc entry layout");

template <typename T>
bool typed(const std::vector<uint8_t>& raw, std::vector<T>& out)
{
    if (raw.size() % sizeof(T) != 0) return false;
    out.resize(raw.size() / sizeof(T));
    if (!raw.empty()) std::memcpy(out.data(), raw.data(), raw.size());
    return true;
}
} // namespace

bool loadSceneFile(const std::string& path, SceneFile& out, std::string& error)
{
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) {
        error = "cannot open " + path;
        return false;
    }
    std::map<std::string, std::vector<uint8_t>> sec;
    bool ok = false;
    do {
        char     magic[8];
        uint32_t head[2];
        if (std::fread(magic, 1, 8, f) != 8 || std::memcmp(magic, "IGB200DB", 8) != 0) { error = "not an .igdb file"; break; }
        if (std::fread(head, 4, 2, f) != 2 || head[0] != 1u || head[1] > 64u) { error = "unsupported .igdb version"; break; }
        std::vector<TocEntry> toc(head[1]);
        if (std::fread(toc.data(), sizeof(TocEntry), toc.size(), f) != toc.size()) { error = "truncated table of contents"; break; }
        // the table of contents is untrusted input: every section must lie inside the file before anything is allocated
        struct stat st {};
        if (fstat(fileno(f), &st) != 0) { error = "cannot stat " + path; break; }
        const uint64_t file_size = (uint64_t)st.st_size;
        bool bad = false;
        for (const TocEntry& e : toc) {
            if (e.offset > file_size || e.bytes > file_size - e.offset) {
                error = std::string("section ") + std::string(e.name, strnlen(e.name, 24)) + " lies outside the file";
                bad   = true;
                break;
            }
            std::vector<uint8_t> data;
            try {
                data.resize(e.bytes);
            } catch (const std::bad_alloc&) {
                error = "out of memory reading " + path;
                bad   = true;
                break;
            }
            if (fseeko(f, (off_t)e.offset, SEEK_SET) != 0 || (e.bytes && std::fread(data.data(), 1, e.bytes, f) != e.bytes)) {
                error = std::string("truncated section ") + std::string(e.name, strnlen(e.name, 24));
                bad   = true;
                break;
            }
            sec[std::string(e.name, strnlen(e.name, 24))] = std::move(data);
        }
        if (bad) break;
        ok = true;
    } while (false);
    std::fclose(f);
    if (!ok) return false;

    auto need = [&](const char* name, size_t bytes) -> const std::vector<uint8_t>* {
        auto it = sec.find(name);
        if (it == sec.end() || (bytes && it->second.size() != bytes)) {
            error = std::string("missing or malformed section ") + name;
            return nullptr;
        }
        return &it->second;
    };

    SceneDatabase& db = out.database;
    std::vector<LookupEntry> lookups;
    const auto* s_look = need("lookups", 0);
    const auto* s_data = need("shape_data", 0);
    const auto* s_ent  = need("entities", 0);
    const auto* s_l0   = need("leaves0", 0);
    const auto* s_l1   = need("leaves1", 0);
    const auto* s_bbox = need("bbox", 24);
    const auto* s_mat  = need("material_count", 8);
    if (!s_look || !s_data || !s_ent || !s_l0 || !s_l1 || !s_bbox || !s_mat) return false;
    if (!typed(*s_look, lookups) || s_ent->size() % (36 * 4) != 0 || s_l0->size() % 96 != 0 || s_l1->size() % 96 != 0) {
        error = "table size is not a multiple of its record size";
        return false;
    }
    db.DynTables["shapes"].assign(std::move(lookups), *s_data);
    db.FixTables["entities"].assign(s_ent->size() / (36 * 4), *s_ent);
    db.SceneBVHs["trimesh"].Leaves = *s_l0;
    db.SceneBVHs["sphere"].Leaves  = *s_l1;
    std::memcpy(db.SceneBBox.min.data(), s_bbox->data(), 12);
    std::memcpy(db.SceneBBox.max.data(), s_bbox->data() + 12, 12);
    uint64_t mc = 0;
    std::memcpy(&mc, s_mat->data(), 8);
    db.MaterialCount = (size_t)mc;
    if (auto it = sec.find("entity_per_material"); it != sec.end()) typed(it->second, out.entity_per_material);

    ShadingTables& sh = out.shading;
    const auto* s_cam = need("camera", sizeof(igb200_camera));
    const auto* s_tec = need("technique", sizeof(igb200_technique));
    const auto* s_sel = need("selector", 8);
    const auto* s_flm = need("film", 8);
    if (!s_cam || !s_tec || !s_sel || !s_flm) return false;
    std::memcpy(&sh.camera, s_cam->data(), sizeof(igb200_camera));
    std::memcpy(&sh.technique, s_tec->data(), sizeof(igb200_technique));
    uint32_t two[2];
    std::memcpy(two, s_sel->data(), 8);
    sh.selector  = two[0];
    sh.hierarchy = two[1];
    std::memcpy(two, s_flm->data(), 8);
    out.width  = two[0];
    out.height = two[1];
    for (const char* name : { "bsdfs", "materials", "infinite", "finite", "textures", "blob" })
        if (!need(name, 0)) return false; // a missing shading section is an error, not an empty table
    const bool tables_ok = typed(sec["bsdfs"], sh.bsdfs) && typed(sec["materials"], sh.materials) && typed(sec["infinite"], sh.infinite)
                           && typed(sec["finite"], sh.finite) && typed(sec["textures"], sh.textures);
    if (!tables_ok) {
        error = "shading table size is not a multiple of its record size";
        return false;
    }
    sh.blob = sec["blob"];
    return true;
}
} // namespace IG
