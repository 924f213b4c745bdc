// igdbtool -- host-only checks of the frontends' file formats (no device needed):
//   igdbtool inspect scene.igdb        prints the table sizes the .igdb reader sees, one `key value` per line
//   igdbtool testimage out.(exr|pfm) W H   writes the image rgb(x, y) = (x, y, x * y + 0.5) through ImageIO
//   igdbtool json2igdb scene.json out.igdb [W H]   runs the scene loader (SceneLoader.cpp) and stores its tables
#include "ImageIO.h"
#include "SceneFile.h"
#include "SceneLoader.h"

#include <cstdlib>
#include <iostream>

using namespace IG;

int main(int argc, char** argv)
{
    const std::string cmd = argc > 1 ? argv[1] : "";
    if (cmd == "inspect" && argc == 3) {
        SceneFile   scene;
        std::string err;
        if (!loadSceneFile(argv[2], scene, err)) {
            std::cerr << "[ERROR] " << err << std::endl;
            return EXIT_FAILURE;
        }
        const SceneDatabase& db = scene.database;
        std::cout << "shapes " << db.DynTables.at("shapes").entryCount() << "\nshape_bytes " << db.DynTables.at("shapes").data().size()
                  << "\nentities " << db.FixTables.at("entities").entryCount() << "\ntrimesh_leaves " << db.SceneBVHs.at("trimesh").Leaves.size() / 96
                  << "\nsphere_leaves " << db.SceneBVHs.at("sphere").Leaves.size() / 96 << "\nmaterials " << db.MaterialCount
                  << "\nbsdfs " << scene.shading.bsdfs.size() << "\ninfinite_lights " << scene.shading.infinite.size()
                  << "\nfinite_lights " << scene.shading.finite.size() << "\ntextures " << scene.shading.textures.size()
                  << "\nblob_bytes " << scene.shading.blob.size() << "\nselector " << scene.shading.selector
                  << "\ntechnique " << scene.shading.technique.type << "\nmax_depth " << scene.shading.technique.max_depth
                  << "\nwidth " << scene.width << "\nheight " << scene.height << "\nbbox";
        for (int k = 0; k < 3; ++k) std::cout << " " << db.SceneBBox.min(k);
        for (int k = 0; k < 3; ++k) std::cout << " " << db.SceneBBox.max(k);
        std::cout << std::endl;
        return EXIT_SUCCESS;
    }
    if (cmd == "testimage" && argc == 5) {
        const size_t w = std::strtoull(argv[3], nullptr, 10), h = std::strtoull(argv[4], nullptr, 10);
        std::vector<float> rgb(3 * w * h);
        for (size_t y = 0; y < h; ++y)
            for (size_t x = 0; x < w; ++x) {
                float* p = rgb.data() + 3 * (y * w + x);
                p[0]     = (float)x;
                p[1]     = (float)y;
                p[2]     = (float)(x * y) + 0.5f;
            }
        return saveImageRGB(argv[2], rgb.data(), w, h, 1.0f) ? EXIT_SUCCESS : EXIT_FAILURE;
    }
    if (cmd == "json2igdb" && (argc == 4 || argc == 6)) {
        SceneFile   scene;
        std::string err;
        const size_t w = argc == 6 ? std::strtoull(argv[4], nullptr, 10) : 0, h = argc == 6 ? std::strtoull(argv[5], nullptr, 10) : 0;
        if (!loadSceneJson(argv[2], scene, err, w, h) || !saveSceneFile(argv[3], scene, err)) {
            std::cerr << "[ERROR] " << err << std::endl;
            return EXIT_FAILURE;
        }
        return EXIT_SUCCESS;
    }
    std::cerr << "usage: igdbtool json2igdb scene.json out.igdb [W H] | igdbtool inspect scene.igdb | igdbtool testimage out.(exr|pfm) W H" << std::endl;
    return EXIT_FAILURE;
}
