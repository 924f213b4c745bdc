// SkyModel.h -- Hosek-Wilkie sky bake of the `sky` light.
//
// Replaces, on the loader side of the boundary, src/runtime/skysun/SkyModel.cpp:9-55 (the 512 x 256 azimuth x elevation
// bake), src/runtime/light/SkyLight.cpp:33-50 (bake -> EXR -> image resource) and the RGB evaluation of
// src/runtime/skysun/model/ArHosekSkyModel.cpp:147-261,514-553.  The evaluation here is this repo's own statement of the
// published model (Hosek & Wilkie, "An Analytic Model for Full Spectral Sky-Dome Radiance", SIGGRAPH 2012, RGB fit);
// only the fitted coefficient tables are taken over (HosekRGBData.inc, scripts/make_hosek_tables.py).
// The reference's own object code is the CHECKER in tests/test_sky_bake.py, never linked or loaded from here.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace IG {

constexpr size_t RES_AZ = 512; // src/runtime/skysun/SkySunConfig.h
constexpr size_t RES_EL = 256;

/// One colour channel of the RGB fit for a fixed (turbidity, ground albedo, solar elevation).
struct HosekChannel {
    double coeff[9];   // A..I of the extended Perez formula
    double radiance;   // mean radiance scale of the channel
    /// theta: angle from the zenith, gamma: angle to the sun (radians).
    double eval(double theta, double gamma) const;
};

/// Interpolates the fitted tables: quintic Bezier in cbrt(elevation / (pi/2)), linear in albedo and turbidity.
HosekChannel hosekCookChannel(int channel, double turbidity, double albedo, double solarElevation);

class SkyModel {
public:
    /// sunElevation / sunAzimuth: ElevationAzimuth of LoaderUtils::getEA (radians, azimuth west of south).
    SkyModel(const float groundAlbedo[3], float sunElevation, float sunAzimuth, float turbidity = 3.0f,
             size_t resAzimuth = RES_AZ, size_t resElevation = RES_EL);

    size_t azimuthCount() const { return mAzimuthCount; }
    size_t elevationCount() const { return mElevationCount; }
    /// RGBA rows as SkyModel.cpp fills them: row y = elevation index (theta = pi/2 * y / rows), alpha 0.
    const std::vector<float>& data() const { return mData; }
    /// SkyModel::computeTotal (SkyModel.cpp:62-71)
    void computeTotal(float rgb[3]) const;
    /// The image as the device sees it after the EXR round trip of SkyLight.cpp:33-50: rows flipped
    /// (Image.cpp:559-560), alpha 1.  `out` holds 4 * azimuthCount * elevationCount floats.
    void toDeviceImage(float* out) const;

private:
    std::vector<float> mData;
    size_t             mAzimuthCount, mElevationCount;
};

} // namespace IG

extern "C" {
/// C entry for bindings (Python loader, tests): bakes the device-side sky image (RGBA f32, rows flipped, alpha 1).
/// Returns 0 on success, -1 on bad arguments.
int ig_b200_bake_sky(const float ground_albedo[3], float sun_elevation, float sun_azimuth, float turbidity,
                     uint32_t res_azimuth, uint32_t res_elevation, float* out_rgba);
/// One Hosek RGB radiance sample (for the parity test): channel 0..2.
double ig_b200_hosek_radiance(int channel, double turbidity, double albedo, double solar_elevation, double theta, double gamma);
}
