// IgTypes.h -- the handful of runtime types the device boundary of the reference is written against.
//
// The reference's `IG::Device` (src/runtime/device/Device.h:14-74) takes `IG::Ray`
// (src/runtime/RuntimeStructs.h:71-75), `IG::SceneDatabase` (src/runtime/table/SceneDatabase.h:8-20) with its
// `DynTable` / `FixTable` (src/runtime/table/DynTable.h:6-36, FixTable.h) and `TonemapSettings`
// (RuntimeStructs.h).  When the shim of this directory is compiled INTO the reference, those headers are used
// and this file is not; it exists so that the same Device.cpp / Runtime.cpp / frontends build and run in a tree
// that has neither Eigen nor the reference (the GPU box).  Only the members the boundary touches are declared,
// with the reference's names; layouts that cross the C ABI (Ray, LookupEntry) are byte-identical.
#pragma once

#include <cstddef>
#include <cstdint>
#include <string>
#include <string_view>
#include <unordered_map>
#include <vector>

namespace IG {
using uint8  = uint8_t;
using int32  = int32_t;
using uint32 = uint32_t;
using uint64 = uint64_t;

// Stand-in for the Eigen vectors the reference uses: `v(i)` element access is all the boundary needs.
template <int N>
struct VectorNf {
    float  e[N] = {};
    float& operator()(int i) { return e[i]; }
    float  operator()(int i) const { return e[i]; }
    const float* data() const { return e; }
    float*       data() { return e; }
};
using Vector2f = VectorNf<2>;
using Vector3f = VectorNf<3>;

// RuntimeStructs.h:71-75 -- 32 bytes, the igtrace input record (StreamRay, src/artic/traversal/ray.art:2-7).
struct Ray {
    Vector3f Origin;
    Vector3f Direction;
    Vector2f Range;
};
static_assert(sizeof(Ray) == 32, "IG::Ray must stay byte-identical to igb200_ray");

// table/DynTable.h:6-10
struct LookupEntry {
    uint32 TypeID;
    uint32 Flags;
    uint64 Offset;
};
static_assert(sizeof(LookupEntry) == 16, "LookupEntry must stay byte-identical to igb200_lookup_entry");

class DynTable {
public:
    size_t entryCount() const { return mLookups.size(); }
    const std::vector<LookupEntry>& lookups() const { return mLookups; }
    const std::vector<uint8>&       data() const { return mData; }
    // the .igdb reader fills the table wholesale; the reference's loader appends entry by entry (addLookup)
    void assign(std::vector<LookupEntry> lookups, std::vector<uint8> data)
    {
        mLookups = std::move(lookups);
        mData    = std::move(data);
    }

private:
    std::vector<LookupEntry> mLookups;
    std::vector<uint8>       mData;
};

class FixTable {
public:
    size_t entryCount() const { return mCount; }
    const std::vector<uint8>& data() const { return mData; }
    void assign(size_t count, std::vector<uint8> data)
    {
        mCount = count;
        mData  = std::move(data);
    }

private:
    size_t             mCount = 0;
    std::vector<uint8> mData;
};

struct BoundingBox {
    Vector3f min, max;
};

// table/SceneDatabase.h:8-20
struct SceneBVH {
    std::vector<uint8> Nodes;  // the reference's Node2/4/8 array: not consumed by this backend
    std::vector<uint8> Leaves; // EntityLeaf1 records, 96 bytes each
};
struct SceneDatabase {
    std::unordered_map<std::string_view, SceneBVH> SceneBVHs;
    std::unordered_map<std::string, DynTable>      DynTables;
    std::unordered_map<std::string, FixTable>      FixTables;
    float       SceneRadius   = 0;
    BoundingBox SceneBBox;
    size_t      MaterialCount = 0;
};

// RuntimeStructs.h (TonemapSettings): what Device::tonemap consumes (Device.cpp:1662-1687)
struct TonemapSettings {
    const char* AOV            = "";
    size_t      Method         = 0;
    bool        UseGamma       = true;
    float       Scale          = 1;
    float       ExposureFactor = 1;
    float       ExposureOffset = 0;
};

// technique/TechniqueInfo.h: the slice of TechniqueVariantInfo the device reads (AOV names, film lock, spi override)
struct TechniqueVariantInfo {
    std::vector<std::string> EnabledAOVs;
    bool   LockFramebuffer = false;
    size_t OverrideSPI     = 0;
    size_t GetSPI(size_t spi) const { return OverrideSPI ? OverrideSPI : spi; }
};
// The reference hands JIT-compiled function pointers here (technique/TechniqueVariant.h:13-36); this backend has no
// JIT and ignores them, so the stand-ins are empty.
struct TechniqueVariantShaderSet {};
struct ParameterSet {};
class Statistics;

// RuntimeStructs.h:15-54 -- utility entry points of the device (out of the hot path; Device.h:70-73)
struct GlareSettings {
    const char* AOV                 = nullptr;
    float       Scale               = 1;
    float       LuminanceMax        = 0;
    float       LuminanceAverage    = 0;
    float       LuminanceMultiplier = 0;
    float       VerticalIlluminance = 0;
};
struct GlareOutput {
    float DGP                 = 0;
    float VerticalIlluminance = 0;
    int   NumPixels           = 0;
    float AvgLum              = 0;
    float AvgOmega            = 0;
};
struct ImageInfoSettings {
    const char* AOV   = nullptr;
    float       Scale = 1;
    size_t      Bins  = 0;
    int*        HistogramR = nullptr;
    int*        HistogramG = nullptr;
    int*        HistogramB = nullptr;
    int*        HistogramL = nullptr;
    bool        AcquireErrorStats = false;
    bool        AcquireHistogram  = false;
};
struct ImageInfoOutput {
    float Min = 0, Max = 0, Average = 0, SoftMin = 0, SoftMax = 0, Median = 0;
    int   InfCount = 0, NaNCount = 0, NegCount = 0;
};
// device/ShaderOutput.h: the JIT'd entry point of a shader; this backend never executes it
template <typename T>
struct ShaderOutput {
    T Exec{};
};

// Target.h: only the device ordinal matters to this backend.
class Target {
public:
    explicit Target(int device = 0) : mDevice(device) {}
    int device() const { return mDevice; }
    std::string toString() const { return "b200:" + std::to_string(mDevice); }

private:
    int mDevice;
};
} // namespace IG
