// SkyModel.cpp -- see SkyModel.h.  Own statement of the Hosek-Wilkie RGB sky-dome model + the reference's bake loop.
#include "SkyModel.h"

#include <algorithm>
#include <cmath>

namespace IG {
namespace {
#include "HosekRGBData.inc"

constexpr double kHalfPi = 1.57079632679489661923;

// Bernstein basis of degree 5 at x: the fitted control points are blended along the (cube-root warped) solar elevation.
void bernstein5(double x, double w[6])
{
    const double y  = 1.0 - x;
    const double x2 = x * x, y2 = y * y;
    w[0] = y2 * y2 * y;
    w[1] = 5.0 * y2 * y2 * x;
    w[2] = 10.0 * y2 * y * x2;
    w[3] = 10.0 * y2 * x2 * x;
    w[4] = 5.0 * y * x2 * x2;
    w[5] = x2 * x2 * x;
}
} // namespace

HosekChannel hosekCookChannel(int channel, double turbidity, double albedo, double solarElevation)
{
    // Table axes: albedo {0, 1}, integer turbidity 1..10; both interpolated linearly (ArHosekSkyModel.cpp:147-234 does
    // the same four-corner blend; turbidity 10 has no upper neighbour).
    const int    t_lo = (int)turbidity;
    const double t_f  = turbidity - (double)t_lo;
    double       w[6];
    bernstein5(std::cbrt(solarElevation / kHalfPi), w);

    HosekChannel out{};
    for (int corner = 0; corner < 4; ++corner) {
        const int    a  = corner & 1;
        const int    hi = corner >> 1;
        const int    t  = t_lo - 1 + hi; // table row (turbidity 1 -> row 0)
        if (hi && t_lo == 10) continue;
        if (t < 0 || t > 9) continue;    // outside the fitted range (the reference reads out of bounds here)
        const double weight = (a ? albedo : 1.0 - albedo) * (hi ? t_f : 1.0 - t_f);
        for (int k = 0; k < 6; ++k) {
            const double wk = weight * w[k];
            for (int i = 0; i < 9; ++i) out.coeff[i] += wk * kHosekRGBConfig[channel][a][t][k][i];
            out.radiance += wk * kHosekRGBRadiance[channel][a][t][k];
        }
    }
    return out;
}

double HosekChannel::eval(double theta, double gamma) const
{
    // F(theta, gamma) = (1 + A e^{B / (cos theta + 0.01)}) (C + D e^{E gamma} + F cos^2 gamma + G chi(H, gamma) + I sqrt(cos theta))
    // chi = (1 + cos^2 gamma) / (1 + H^2 - 2 H cos gamma)^{3/2}   (the anisotropic Mie term)
    const double* c  = coeff;
    const double  cg = std::cos(gamma), ct = std::cos(theta);
    const double  chi = (1.0 + cg * cg) / std::pow(1.0 + c[8] * c[8] - 2.0 * c[8] * cg, 1.5);
    const double  horizon = 1.0 + c[0] * std::exp(c[1] / (ct + 0.01));
    const double  lobes   = c[2] + c[3] * std::exp(c[4] * gamma) + c[5] * cg * cg + c[6] * chi + c[7] * std::sqrt(ct);
    return horizon * lobes * radiance;
}

SkyModel::SkyModel(const float groundAlbedo[3], float sunElevation, float sunAzimuth, float turbidity, size_t resAzimuth,
                   size_t resElevation)
    : mData(resAzimuth * resElevation * 4, 0.0f)
    , mAzimuthCount(resAzimuth)
    , mElevationCount(resElevation)
{
    constexpr float Pi = 3.14159265358979323846f, Pi2 = Pi / 2, Pi4 = Pi / 4;
    // The reference hands `pi/2 - elevation` to the model as its solar elevation (SkyModel.cpp:13) and uses the same
    // angle for the sun's sine / cosine below; kept as is -- the bake is the contract, not the astronomy.
    const float solar_elevation = Pi2 - sunElevation;
    const float sun_se = std::sin(solar_elevation), sun_ce = std::cos(solar_elevation);

    HosekChannel ch[3];
    for (int k = 0; k < 3; ++k) ch[k] = hosekCookChannel(k, (double)turbidity, (double)groundAlbedo[k], (double)solar_elevation);

    constexpr float CIEYSum = 106.856980f;
    for (size_t y = 0; y < mElevationCount; ++y) {
        const float theta = Pi2 * y / (float)mElevationCount;
        const float st = std::sin(theta), ct = std::cos(theta);
        for (size_t x = 0; x < mAzimuthCount; ++x) {
            float azimuth = 2 * Pi * x / (float)mAzimuthCount - Pi4;
            if (azimuth < 0) azimuth += 2 * Pi;
            const float cosGamma = ct * sun_ce + st * sun_se * std::cos(azimuth - sunAzimuth);
            const float gamma    = std::acos(std::min(1.0f, std::max(-1.0f, cosGamma)));
            float*      px       = &mData[(y * mAzimuthCount + x) * 4];
            for (int k = 0; k < 3; ++k) px[k] = std::max(0.0f, (float)ch[k].eval((double)theta, (double)gamma) / CIEYSum);
        }
    }
}

void SkyModel::computeTotal(float rgb[3]) const
{
    rgb[0] = rgb[1] = rgb[2] = 0.0f;
    for (size_t i = 0; i < mElevationCount * mAzimuthCount; ++i)
        for (int k = 0; k < 3; ++k) rgb[k] += mData[i * 4 + k];
}

void SkyModel::toDeviceImage(float* out) const
{
    const size_t row = mAzimuthCount * 4;
    for (size_t y = 0; y < mElevationCount; ++y) {
        const float* src = &mData[(mElevationCount - 1 - y) * row];
        float*       dst = out + y * row;
        for (size_t x = 0; x < mAzimuthCount; ++x) {
            dst[4 * x + 0] = src[4 * x + 0];
            dst[4 * x + 1] = src[4 * x + 1];
            dst[4 * x + 2] = src[4 * x + 2];
            dst[4 * x + 3] = 1.0f;
        }
    }
}

} // namespace IG

extern "C" int ig_b200_bake_sky(const float ground_albedo[3], float sun_elevation, float sun_azimuth, float turbidity,
                                uint32_t res_azimuth, uint32_t res_elevation, float* out_rgba)
{
    if (!ground_albedo || !out_rgba || res_azimuth == 0 || res_elevation == 0) return -1;
    IG::SkyModel model(ground_albedo, sun_elevation, sun_azimuth, turbidity, res_azimuth, res_elevation);
    model.toDeviceImage(out_rgba);
    return 0;
}

extern "C" double ig_b200_hosek_radiance(int channel, double turbidity, double albedo, double solar_elevation, double theta, double gamma)
{
    if (channel < 0 || channel > 2) return 0.0;
    return IG::hosekCookChannel(channel, turbidity, albedo, solar_elevation).eval(theta, gamma);
}
