// Device::tonemap (src/runtime/device/Device.cpp:1662-1687) -> ig_tonemap_pipeline (src/artic/entrypoints/tonemap.art):
// film or AOV sums -> packed 8-bit ARGB, one thread per pixel.  HBM-bound: 12 B read + 4 B written per pixel.
#include "wavefront.cuh"

namespace igb {
namespace {

constexpr float kFltEps = 1.1920928955078125e-7f;

__device__ __forceinline__ float safe_div(float a, float b) { return fabsf(b) <= kFltEps ? 0.0f : __fdiv_rn(a, b); } // core/common.art

__device__ __forceinline__ float srgb_gamma(float x) // core/color.art:136-142
{
    return x <= 0.0031308f ? 12.92f * x : 1.055f * powf(x, 0.416666667f) - 0.055f;
}

__device__ __forceinline__ float tone_curve(uint32_t method, float L) // mod tonemapping, core/color.art:101-134
{
    switch (method) {
    case 0: return L;
    case 1: return safe_div(L, 1.0f + L);                                   // reinhard
    case 2: return safe_div(L * (1.0f + __fdiv_rn(L, 4.0f * 4.0f)), 1.0f + L); // reinhard_modified, white point 4
    case 3: return safe_div(L * (2.51f * L + 0.03f), L * (2.43f * L + 0.59f) + 0.14f); // aces (Narkowicz 2015)
    default: {                                                              // uncharted2
        auto f = [](float x) {
            const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
            return __fdiv_rn(x * (A * x + C * B) + D * E, x * (A * x + B) + D * F) - __fdiv_rn(E, F);
        };
        return __fdiv_rn(f(L), f(11.2f));
    }
    }
}

__device__ __forceinline__ uint32_t pack(uint32_t r, uint32_t g, uint32_t b, uint32_t a) { return (a << 24) | (r << 16) | (g << 8) | b; }
__device__ __forceinline__ uint32_t to_byte(float c) { return (uint32_t)(fminf(fmaxf(c, 0.0f), 1.0f) * 255.0f) & 0xFFu; } // `as u8` truncates

__global__ void tonemap_kernel(const float* __restrict__ in, uint32_t* __restrict__ out, uint64_t n, TonemapParams p)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float r = in[3 * i + 0] * p.scale, g = in[3 * i + 1] * p.scale, b = in[3 * i + 2] * p.scale;
    // srgb_to_xyY, core/color.art:84-99
    const float X = 0.4124564f * r + 0.3575761f * g + 0.1804375f * b;
    const float Y = 0.2126729f * r + 0.7151522f * g + 0.0721750f * b;
    const float Z = 0.0193339f * r + 0.1191920f * g + 0.9503041f * b;
    const float s = X + Y + Z;
    float       cx = 0.0f, cy = 0.0f, cY = 0.0f;
    if (!(s <= kFltEps)) { cx = __fdiv_rn(X, s); cy = __fdiv_rn(Y, s); cY = Y; }
    uint32_t q;
    if (isnan(cY)) q = pack(0, 255, 255, 255);                       // cyan
    else if (!isfinite(cY)) q = pack(255, 0, 150, 255);              // pink
    else if (cx < 0.0f || cy < 0.0f || cY < 0.0f) q = pack(255, 255, 0, 255); // "orange"
    else {
        const float L  = p.exposure_factor * cY + p.exposure_offset;
        const float nL = tone_curve(p.method, L);
        float       R = 0.0f, G = 0.0f, B = 0.0f; // xyY_to_srgb, core/color.art:77-93
        if (!(cy <= kFltEps)) {
            const float x2 = __fdiv_rn(cx * nL, cy), y2 = nL, z2 = __fdiv_rn((1.0f - cx - cy) * nL, cy);
            R = 3.2404542f * x2 - 1.5371385f * y2 - 0.4985314f * z2;
            G = -0.9692660f * x2 + 1.8760108f * y2 + 0.0415560f * z2;
            B = 0.0556434f * x2 - 0.2040259f * y2 + 1.0572252f * z2;
        }
        if (p.use_gamma) { R = srgb_gamma(R); G = srgb_gamma(G); B = srgb_gamma(B); }
        q = pack(to_byte(R), to_byte(G), to_byte(B), 255u); // alpha is 1 on both branches of xyY_to_srgb
    }
    out[i] = q;
}

} // namespace

cudaError_t launch_tonemap(const float* in, uint32_t* out, uint64_t n, const TonemapParams& p, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    tonemap_kernel<<<(uint32_t)((n + 255) / 256), 256, 0, stream>>>(in, out, n, p);
    return cudaGetLastError();
}

} // namespace igb
