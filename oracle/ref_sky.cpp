// TEST INFRASTRUCTURE ONLY.  Checker for the product's sky bake (ignis-masterthesis_b200/host/SkyModel.cpp).
// Restates the bake loop of SkyModel::SkyModel (src/runtime/skysun/SkyModel.cpp:9-55) around the REFERENCE's own
// Hosek-Wilkie model, ArHosekSkyModel.cpp, which `make -C oracle ref` compiles from /root/reference into the same
// oracle/_ref/libhosek.so.  Output rows are in the bake's own order (row y = elevation index), alpha 0 as SkyModel fills it.
#include <algorithm>
#include <cmath>
#include <cstddef>

struct ArHosekSkyModelState;
ArHosekSkyModelState* arhosek_rgb_skymodelstate_alloc_init(const double turbidity, const double albedo, const double elevation);
void                  arhosekskymodelstate_free(ArHosekSkyModelState* state);
double                arhosek_tristim_skymodel_radiance(ArHosekSkyModelState* state, double theta, double gamma, int channel);

extern "C" void igo_ref_sky_bake(const float* ground, float sun_elevation, float sun_azimuth, float turbidity, int res_az, int res_el,
                                 float* out)
{
    const float Pi = 3.14159265358979323846f, Pi2 = 1.57079632679489661923f, Pi4 = 0.78539816339744830961f;
    const float solar_elevation = Pi2 - sun_elevation;                  // SkyModel.cpp:13
    const float sun_se = std::sin(solar_elevation), sun_ce = std::cos(solar_elevation);
    ArHosekSkyModelState* states[3];
    for (int k = 0; k < 3; ++k) states[k] = arhosek_rgb_skymodelstate_alloc_init(turbidity, ground[k], solar_elevation);
    for (size_t y = 0; y < (size_t)res_el; ++y) {
        const float theta = Pi2 * y / (float)res_el;                      // ELEVATION_RANGE * y / count, :33
        const float st = std::sin(theta), ct = std::cos(theta);
        for (size_t x = 0; x < (size_t)res_az; ++x) {
            float azimuth = 2 * Pi * x / (float)res_az - Pi4;             // :37
            if (azimuth < 0) azimuth += 2 * Pi;
            const float cosGamma = ct * sun_ce + st * sun_se * std::cos(azimuth - sun_azimuth);
            const float gamma    = std::acos(std::min(1.0f, std::max(-1.0f, cosGamma)));
            for (int k = 0; k < 3; ++k) {
                const float CIEYSum  = 106.856980f;
                const float radiance = (float)arhosek_tristim_skymodel_radiance(states[k], theta, gamma, k) / CIEYSum;
                out[(y * res_az + x) * 4 + k] = std::max(0.0f, radiance);
            }
            out[(y * res_az + x) * 4 + 3] = 0.0f;
        }
    }
    for (int k = 0; k < 3; ++k) arhosekskymodelstate_free(states[k]);
}
