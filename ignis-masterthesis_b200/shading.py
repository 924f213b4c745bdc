"""POD shading tables of include/ignis_b200.h (camera, BSDFs, materials, lights, technique).

In the reference these parameters are baked into per-scene Artic source by the loader
(src/runtime/shader/ShaderUtils.cpp:95-122, src/runtime/loader/LoaderLight.cpp:106-250,
src/runtime/technique/PathTechnique.cpp:36-79, src/runtime/camera/PerspectiveCamera.cpp:27-67); here they
travel as plain tables.  The helpers reproduce the loader-side arithmetic (fov -> scale, plane detection for
area lights, light table split into infinite / finite).
"""
import ctypes as C
import math

import numpy as np

BSDF_DIFFUSE, BSDF_DIELECTRIC, BSDF_MIRROR, BSDF_CONDUCTOR, BSDF_BUMPMAP, BSDF_THIN_DIELECTRIC = 0, 1, 2, 3, 4, 5
CAMERA_PERSPECTIVE, CAMERA_FISHLENS = 0, 1
LIGHT_POINT, LIGHT_ENV, LIGHT_AREA_SHAPE, LIGHT_AREA_PLANE, LIGHT_SPOT, LIGHT_DIRECTIONAL, LIGHT_ENV_TEXTURED = 0, 1, 2, 3, 4, 5, 6
LIGHT_AREA_SPHERE, LIGHT_ENV_FUNCTION, LIGHT_SUN = 7, 8, 9
SKY_CIE, SKY_CIE_SUNNY, SKY_PEREZ = 0, 1, 2
SKYFLAG_VARIANT, SKYFLAG_HAS_GROUND = 1, 2
TEX_CHECKERBOARD, TEX_IMAGE = 0, 1
FILTER_NEAREST, FILTER_BILINEAR, FILTER_BICUBIC = 0, 1, 2
WRAP_REPEAT, WRAP_MIRROR, WRAP_CLAMP = 0, 1, 2
IMAGE_RGBA8, IMAGE_RGBA32F = 0, 1
SELECTOR_UNIFORM, SELECTOR_HIERARCHY, SELECTOR_CDF = 0, 1, 2
NO_CDF = 0xFFFFFFFF
TECH_PATH, TECH_DEBUG, TECH_INFOBUFFER = 0, 1, 2
TECHFLAG_ONE_SIDED_DIFFUSE = 1
TECHFLAG_AOV_MIS = 2
TECHFLAG_IB_HANDLE_SPECULAR = 4
LIGHTFLAG_SCALED_NEE = 1


class Camera(C.Structure):
    _fields_ = [("eye", C.c_float * 3), ("dir", C.c_float * 3), ("up", C.c_float * 3), ("scale", C.c_float * 2),
                ("tmin", C.c_float), ("tmax", C.c_float), ("type", C.c_uint32)]


class Word(C.Union):
    _fields_ = [("f", C.c_float), ("u", C.c_uint32)]


class Bsdf(C.Structure):
    _fields_ = [("type", C.c_uint32), ("color", C.c_float * 3), ("color2", C.c_float * 3), ("p", C.c_float * 3),
                ("q", Word * 2)]


class Texture(C.Structure):
    _fields_ = [("type", C.c_uint32), ("filter", C.c_uint32), ("wrap_u", C.c_uint32), ("wrap_v", C.c_uint32),
                ("transform", C.c_float * 9), ("color0", C.c_float * 3), ("color1", C.c_float * 3), ("scale", C.c_float * 2),
                ("width", C.c_uint32), ("height", C.c_uint32), ("format", C.c_uint32), ("opaque", C.c_uint32),
                ("offset", C.c_uint64), ("_pad", C.c_uint32 * 2)]


class Material(C.Structure):
    _fields_ = [("bsdf", C.c_int32), ("light", C.c_int32)]


class Light(C.Structure):
    _fields_ = [("type", C.c_uint32), ("entity", C.c_int32), ("color", C.c_float * 3), ("v", C.c_float * 16),
                ("tex", C.c_uint32), ("cdf", C.c_uint32), ("flags", C.c_uint32)]


class Technique(C.Structure):
    _fields_ = [("type", C.c_uint32), ("max_depth", C.c_int32), ("min_depth", C.c_int32), ("clamp", C.c_float),
                ("enable_nee", C.c_int32), ("debug_mode", C.c_int32), ("flags", C.c_uint32), ("_pad", C.c_int32)]


class ShadingC(C.Structure):
    _fields_ = [("camera", Camera), ("technique", Technique),
                ("bsdfs", C.c_void_p), ("bsdf_count", C.c_uint64),
                ("materials", C.c_void_p), ("material_count", C.c_uint64),
                ("infinite_lights", C.c_void_p), ("infinite_light_count", C.c_uint64),
                ("finite_lights", C.c_void_p), ("finite_light_count", C.c_uint64),
                ("textures", C.c_void_p), ("texture_count", C.c_uint64), ("blob", C.c_void_p), ("blob_bytes", C.c_uint64),
                ("selector", C.c_uint32), ("hierarchy", C.c_uint32)]


class RenderSettings(C.Structure):
    _fields_ = [("rays", C.c_void_p), ("ray_count", C.c_uint64), ("spi", C.c_uint32), ("width", C.c_uint32),
                ("height", C.c_uint32), ("iteration", C.c_uint32), ("frame", C.c_uint32), ("user_seed", C.c_uint32)]


class TonemapSettings(C.Structure):
    _fields_ = [("aov", C.c_char_p), ("method", C.c_uint32), ("use_gamma", C.c_int32), ("scale", C.c_float),
                ("exposure_factor", C.c_float), ("exposure_offset", C.c_float)]


class RenderStats(C.Structure):
    _fields_ = [("camera_rays", C.c_uint64), ("bounce_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("rounds", C.c_uint64), ("kernel_launches", C.c_uint64), ("device_ms", C.c_float), ("_pad", C.c_float)]


assert C.sizeof(Bsdf) == 48 and C.sizeof(Light) == 96 and C.sizeof(Camera) == 56 and C.sizeof(Technique) == 32
assert C.sizeof(Texture) == 120


def _f3(dst, src):
    s = np.asarray(src, np.float32).reshape(-1)
    if s.size == 1:
        s = np.repeat(s, 3)
    for k in range(3):
        dst[k] = float(s[k])


def as_plane(vertices, faces):
    """TriMesh::getAsPlane (src/runtime/mesh/TriMesh.cpp:520-628) for the 4-vertex / 2-face case.
    Returns (origin, x_axis, y_axis) in object space or None."""
    v = np.asarray(vertices, np.float32).reshape(-1, 3)
    f = np.asarray(faces).reshape(-1, 3)
    if f.shape[0] != 2 or v.shape[0] != 4:
        return None
    fn = lambda t: np.cross(v[t[1]] - v[t[0]], v[t[2]] - v[t[0]])
    n0, n1 = fn(f[0]), fn(f[1])
    n0, n1 = n0 / np.linalg.norm(n0), n1 / np.linalg.norm(n1)
    if not np.allclose(n0, n1, atol=1e-5):
        return None
    origin = v[0]

    def angle(start):
        a = v[(start + 0) % 3 + 1] - origin
        b = v[(start + 1) % 3 + 1] - origin
        return abs(math.acos(float(np.clip(np.dot(a / np.linalg.norm(a), b / np.linalg.norm(b)), -1, 1))))

    a12, a23, a31 = angle(0), angle(1), angle(2)
    if a12 >= a23 and a12 >= a31:
        sel = 0
    elif a23 >= a31 and a23 >= a12:
        sel = 1
    else:
        sel = 2
    x_axis = v[(sel + 0) % 3 + 1] - origin
    y_axis = v[(sel + 1) % 3 + 1] - origin
    if float(np.dot(n0, np.cross(x_axis, y_axis))) < 0:
        x_axis, y_axis = y_axis, x_axis
    return origin.astype(np.float32), x_axis.astype(np.float32), y_axis.astype(np.float32)


class Shading:
    """Builder + owner of the POD shading tables; `c_shading()` gives the igb200_shading view."""

    def __init__(self):
        self.camera = Camera()
        self.technique = Technique(type=TECH_PATH, max_depth=64, min_depth=2, clamp=0.0, enable_nee=1, debug_mode=0)
        self.bsdfs, self.materials, self.infinite, self.finite = [], [], [], []
        self.textures, self.blob = [], bytearray()
        self.selector, self.hierarchy = SELECTOR_UNIFORM, 0
        self._keep = None

    # ---- camera: PerspectiveCamera.cpp:27-67 + compute_scale_from_hfov/vfov (camera/perspective.art:1-11) ----
    def set_perspective(self, eye, direction, up, fov_deg, width, height, near=0.01, far=100.0, vertical=False):
        _f3(self.camera.eye, eye)
        _f3(self.camera.dir, direction)
        _f3(self.camera.up, up)
        aspect = np.float32(width) / np.float32(height)
        t = np.float32(math.tan(np.float32(np.float32(fov_deg) * np.float32(math.pi / 180.0)) / 2))
        if vertical:
            sh = t
            sw = np.float32(sh * aspect)
        else:
            sw = t
            sh = np.float32(sw / aspect)
        self.camera.scale[0], self.camera.scale[1] = float(sw), float(sh)
        if far < near:
            near, far = far, near
        self.camera.tmin, self.camera.tmax = float(near), float(far)
        self.camera.type = CAMERA_PERSPECTIVE
        return self

    def set_fishlens_from_transform(self, m, width, height, near=0.0, far=float(np.finfo(np.float32).max), mode="circular"):
        """make_fishlens_camera (camera/fishlens.art:8-36): scale = (xasp, yasp) of the aspect mode; orientation as for the
        perspective camera (FishLensCamera.cpp:59-66)."""
        m = np.asarray(m, np.float64).reshape(-1, 4)
        _f3(self.camera.eye, m[:3, 3])
        _f3(self.camera.dir, m[:3, 2])
        _f3(self.camera.up, m[:3, 1])
        asp = np.float32(width) / np.float32(height)
        one = np.float32(1)
        if mode == "circular":
            xasp, yasp = (one if asp < 1 else asp), (one if asp > 1 else asp)
        elif mode == "cropped":
            xasp, yasp = (one / asp if asp < 1 else one), (one / asp if asp > 1 else one)
        else:  # full
            f = np.float32(np.sqrt(asp * asp + one) * np.float32(height)) / np.float32(min(width, height))
            xasp, yasp = (f if asp < 1 else f / asp), (f if asp > 1 else f * asp)
        self.camera.scale[0], self.camera.scale[1] = float(xasp), float(yasp)
        if far < near:
            near, far = far, near
        self.camera.tmin, self.camera.tmax = float(near), float(far)
        self.camera.type = CAMERA_FISHLENS
        return self

    def set_perspective_from_transform(self, m, fov_deg, width, height, near=0.01, far=100.0, vertical=False):
        """Camera orientation from a 4x4 (or 3x4) transform: eye = T*0, dir = col 2, up = col 1
        (PerspectiveCamera::getOrientation, PerspectiveCamera.cpp:69-78)."""
        m = np.asarray(m, np.float64).reshape(-1, 4)
        return self.set_perspective(m[:3, 3], m[:3, 2], m[:3, 1], fov_deg, width, height, near, far, vertical)

    def set_path(self, max_depth=64, min_depth=2, clamp=0.0, nee=True, one_sided_diffuse=False, aov_mis=False):
        """PathTechnique.cpp:8-17; aov_mis adds the AOVs "Direct Weights" and "NEE Weights" (:23-27)."""
        self.technique = Technique(type=TECH_PATH, max_depth=max_depth, min_depth=min_depth, clamp=clamp,
                                   enable_nee=1 if nee else 0, debug_mode=0,
                                   flags=(TECHFLAG_ONE_SIDED_DIFFUSE if one_sided_diffuse else 0) | (TECHFLAG_AOV_MIS if aov_mis else 0))
        return self

    def set_infobuffer(self, max_depth=64, handle_specular=False):
        """InfoBufferTechnique.cpp:6-12: AOVs "Normals", "Albedo", "Depth"; the film is locked."""
        self.technique = Technique(type=TECH_INFOBUFFER, max_depth=max_depth, min_depth=1, clamp=0.0, enable_nee=0, debug_mode=0,
                                   flags=TECHFLAG_IB_HANDLE_SPECULAR if handle_specular else 0)
        return self

    def set_debug(self, mode):
        self.technique = Technique(type=TECH_DEBUG, max_depth=1, min_depth=1, clamp=0.0, enable_nee=0, debug_mode=mode)
        return self

    # ---- BSDFs ----
    def _bsdf(self, type_, color=(0, 0, 0), color2=(0, 0, 0), p=()):
        b = Bsdf(type=type_)
        _f3(b.color, color)
        _f3(b.color2, color2)
        for i, x in enumerate(p):
            b.p[i] = float(x)
        self.bsdfs.append(b)
        return len(self.bsdfs) - 1

    def add_diffuse(self, reflectance=(0.8, 0.8, 0.8), texture=None):
        """texture: id from add_checkerboard / add_image_texture replacing the constant reflectance."""
        i = self._bsdf(BSDF_DIFFUSE, reflectance)
        if texture is not None:
            self.bsdfs[i].q[0].u = texture + 1
        return i

    def add_dielectric(self, int_ior=1.5046, ext_ior=1.0, specular_reflectance=(1, 1, 1), specular_transmittance=(1, 1, 1), thin=False):
        # n1 = ext_ior (default vacuum), n2 = int_ior (default bk7): runtime/bsdf/DielectricBSDF.cpp:14-35, BSDF.cpp:8-9;
        # "thin": make_thin_dielectric_bsdf (bsdf/dielectric.art:27-47)
        return self._bsdf(BSDF_THIN_DIELECTRIC if thin else BSDF_DIELECTRIC, specular_reflectance, specular_transmittance, (ext_ior, int_ior))

    def add_mirror(self, specular_reflectance=(1, 1, 1)):
        return self._bsdf(BSDF_MIRROR, specular_reflectance)

    def add_conductor(self, eta, k, specular_reflectance=(1, 1, 1), roughness=0.0, anisotropic=0.0, alpha=None):
        """roughness/anisotropic -> (alpha_u, alpha_v) by microfacet::compute_explicit (core/microfacet.art:409-414);
        alpha = (u, v) gives them explicitly.  0 = smooth."""
        kk = np.asarray(k, np.float32).reshape(-1)
        kk = np.repeat(kk, 3) if kk.size == 1 else kk
        i = self._bsdf(BSDF_CONDUCTOR, specular_reflectance, eta, tuple(kk[:3]))
        if alpha is None:
            r = np.float32(roughness)
            aspect = np.float32(1) if anisotropic == 0 else np.sqrt(np.float32(1) - np.float32(min(max(anisotropic, 0), 1)) * np.float32(0.99))
            alpha = (r / aspect, r * aspect)
        self.bsdfs[i].q[0].f, self.bsdfs[i].q[1].f = float(alpha[0]), float(alpha[1])
        return i

    def add_bumpmap(self, inner, texture, strength=1.0):
        i = self._bsdf(BSDF_BUMPMAP, p=(strength,))
        self.bsdfs[i].q[0].u, self.bsdfs[i].q[1].u = texture + 1, inner + 1
        return i

    # ---- textures and the blob ----
    def _blob_append(self, raw):
        while len(self.blob) % 16:
            self.blob.append(0)
        off = len(self.blob)
        self.blob += bytes(raw)
        return off

    def add_checkerboard(self, scale=(2, 2), color0=(0, 0, 0), color1=(1, 1, 1), transform=None):
        t = Texture(type=TEX_CHECKERBOARD)
        _f3(t.color0, color0)
        _f3(t.color1, color1)
        t.scale[0], t.scale[1] = float(scale[0]), float(scale[1])
        self._tex_transform(t, transform)
        self.textures.append(t)
        return len(self.textures) - 1

    @staticmethod
    def _tex_transform(t, transform):
        m = np.eye(3, dtype=np.float32) if transform is None else np.asarray(transform, np.float32).reshape(3, 3)
        for c in range(3):
            for r in range(3):
                t.transform[3 * c + r] = float(m[r, c])

    def add_image_texture(self, pixels, filter=FILTER_BICUBIC, wrap=(WRAP_REPEAT, WRAP_REPEAT), transform=None, opaque=False):
        """pixels: (h, w) uint32 packed RGBA8 (already mapped to linear like Image::loadAsPacked) or (h, w, 4) float32,
        rows in the reference's in-memory order (bottom row first)."""
        px = np.ascontiguousarray(pixels)
        t = Texture(type=TEX_IMAGE, filter=filter, wrap_u=wrap[0], wrap_v=wrap[1], opaque=1 if opaque else 0)
        if px.dtype == np.uint32 and px.ndim == 2:
            t.format = IMAGE_RGBA8
        elif px.dtype == np.float32 and px.ndim == 3 and px.shape[2] == 4:
            t.format = IMAGE_RGBA32F
        else:
            raise ValueError("image textures are (h, w) uint32 RGBA8 or (h, w, 4) float32")
        t.height, t.width = px.shape[0], px.shape[1]
        t.offset = self._blob_append(px.tobytes())
        self._tex_transform(t, transform)
        self.textures.append(t)
        return len(self.textures) - 1

    def add_cdf2d(self, marginal, conditional):
        """The buffer CDF::computeForImage writes (marginal then conditional rows, runtime/CDF.cpp:42-133)."""
        m, c = np.asarray(marginal, np.float32), np.asarray(conditional, np.float32)
        head = np.array([c.shape[1], c.shape[0], 0, 0], np.uint32)
        off = self._blob_append(head.tobytes() + m.tobytes() + c.tobytes())
        return off // 4

    def set_hierarchy(self, codes, nodes):
        """codes: u32 per finite light, nodes: (n, 8) float32 rows (LightHierarchy::setup)."""
        codes = np.asarray(codes, np.uint32)
        pad = (-codes.size) % 4
        raw = np.concatenate([codes, np.zeros(pad, np.uint32)]).tobytes() + np.asarray(nodes, np.float32).tobytes()
        self.hierarchy = self._blob_append(raw) // 4
        self.selector = SELECTOR_HIERARCHY
        return self

    def set_light_cdf(self, fluxes):
        """The "simple" selector: CDF::computeForArray (runtime/CDF.cpp:11-40) over the finite lights' estimated flux."""
        v = np.asarray(fluxes, np.float32)
        cdf = np.empty_like(v)
        acc = np.float32(0)
        for i, x in enumerate(v):  # f32 running sum, like the reference
            acc = np.float32(acc + x)
            cdf[i] = acc
        if acc > np.float32(1e-5):
            cdf = (cdf * (np.float32(1) / acc)).astype(np.float32)
        else:
            cdf = (np.arange(v.size, dtype=np.float32) * (np.float32(1) / np.float32(v.size))).astype(np.float32)
        cdf[-1] = 1
        self.hierarchy = self._blob_append(cdf.tobytes()) // 4
        self.selector = SELECTOR_CDF
        return self

    def add_material(self, bsdf, light=-1):
        self.materials.append(Material(bsdf=bsdf, light=light))
        return len(self.materials) - 1

    # ---- lights ----
    def _light(self, table, type_, color, v=(), entity=-1):
        l = Light(type=type_, entity=entity)
        _f3(l.color, color)
        for i, x in enumerate(v):
            l.v[i] = float(x)
        table.append(l)
        return len(table) - 1

    def add_point_light(self, position, intensity=(1, 1, 1)):
        return self._light(self.finite, LIGHT_POINT, intensity, tuple(np.asarray(position, np.float32)))

    def add_spot_light(self, position, direction, cutoff_deg=30.0, falloff_deg=20.0, intensity=(1, 1, 1)):
        d = np.asarray(direction, np.float64)
        d = (d / np.linalg.norm(d)).astype(np.float32)
        v = list(np.asarray(position, np.float32)) + list(d) + [np.float32(math.radians(cutoff_deg)),
                                                                  np.float32(math.radians(falloff_deg))]
        return self._light(self.finite, LIGHT_SPOT, intensity, v)

    def add_directional_light(self, direction, irradiance=(1, 1, 1)):
        d = np.asarray(direction, np.float64)
        d = (d / np.linalg.norm(d)).astype(np.float32)
        return self._light(self.infinite, LIGHT_DIRECTIONAL, irradiance, tuple(d))

    def add_sun_light(self, direction, irradiance=(1, 1, 1), angle_deg=11.4, radius=None):
        """make_sun_light (light/sun.art:4-28); cos_sun_angle as SunLight.cpp:37-40 derives it from `angle` or `radius`."""
        d = np.asarray(direction, np.float64)
        d = (d / np.linalg.norm(d)).astype(np.float32)
        cos_angle = np.float32(1 / math.sqrt(radius * radius + 1)) if radius is not None else np.float32(math.cos(math.radians(np.float32(angle_deg) / 2)))
        return self._light(self.infinite, LIGHT_SUN, irradiance, list(d) + [cos_angle])

    def add_env_light(self, radiance=(1, 1, 1), scale=(1, 1, 1)):
        r = np.asarray(radiance, np.float32).reshape(-1)
        s = np.asarray(scale, np.float32).reshape(-1)
        c = (np.repeat(r, 3) if r.size == 1 else r) * (np.repeat(s, 3) if s.size == 1 else s)
        return self._light(self.infinite, LIGHT_ENV, c, (1, 0, 0, 0, 1, 0, 0, 0, 1))

    def add_env_textured(self, texture, scale=(1, 1, 1), cdf=None, transform=None, scaled_nee=False):
        """transform: 3x3 (the loader passes transform.linear().transpose().inverse(), EnvironmentLight.cpp:44).
        scaled_nee sets IGB200_LIGHTFLAG_SCALED_NEE (the reference omits the scale in sample_direct, light/env.art:139)."""
        m = np.eye(3, dtype=np.float32) if transform is None else np.asarray(transform, np.float32).reshape(3, 3)
        i = self._light(self.infinite, LIGHT_ENV_TEXTURED, scale, tuple(m.T.reshape(-1)))
        self.infinite[i].tex = texture
        self.infinite[i].cdf = NO_CDF if cdf is None else cdf
        self.infinite[i].flags = LIGHTFLAG_SCALED_NEE if scaled_nee else 0
        return i

    def add_sky_function(self, kind, variant, has_ground, scale=(1, 1, 1), zenith=(1, 1, 1), ground=(1, 1, 1), zenith_brightness=1.0,
                         ground_brightness=0.2, c2=1.0, sun_dir=(0, 1, 0), transform=None):
        """make_cie_sky_light (kind SKY_CIE, variant = cloudy) / make_cie_sunny_light (kind SKY_CIE_SUNNY, variant = is_clear),
        light/cie.art:9-42, sampled by make_environment_light_function (hemisphere when there is no ground, light/env.art:24-108)."""
        m = np.eye(3, dtype=np.float32) if transform is None else np.asarray(transform, np.float32).reshape(3, 3)
        params = np.zeros(16, np.float32)
        params[0:3], params[3:6], params[6:9] = scale, zenith, ground
        params[9], params[10], params[11] = zenith_brightness, ground_brightness, c2
        params[12:15] = sun_dir
        i = self._light(self.infinite, LIGHT_ENV_FUNCTION, (1, 1, 1), tuple(m.T.reshape(-1)))
        self.infinite[i].tex = kind
        self.infinite[i].cdf = self._blob_append(params.tobytes()) // 4
        self.infinite[i].flags = (SKYFLAG_VARIANT if variant else 0) | (SKYFLAG_HAS_GROUND if has_ground else 0)
        return i

    def add_perez_sky(self, sun_dir, lum, ground, abcde, has_ground=True, transform=None):
        """make_perez_light (light/cie.art:52-57); make_environment_light_function is called with half = false there."""
        a, b, c, d, e = [float(x) for x in abcde]
        i = self.add_sky_function(SKY_PEREZ, False, True, scale=(e, 0, 0), zenith=lum, ground=ground, zenith_brightness=a, ground_brightness=b,
                                  c2=c, sun_dir=sun_dir, transform=transform)
        blob = np.frombuffer(self.blob, np.float32)  # writable view on the bytearray
        blob[self.infinite[i].cdf + 15] = d
        if has_ground:
            self.infinite[i].flags |= SKYFLAG_HAS_GROUND
        else:
            self.infinite[i].flags &= ~SKYFLAG_HAS_GROUND
        return i

    def add_sphere_area_light(self, entity, radiance, origin, radius):
        """make_sphere_area_emitter (light/area.art:241-296): `entity` is an analytic sphere entity; origin / radius in object space."""
        return self._light(self.finite, LIGHT_AREA_SPHERE, radiance, list(np.asarray(origin, np.float32)) + [np.float32(radius)], entity)

    def add_area_light(self, entity, radiance, plane=None):
        """plane = (origin, x_axis, y_axis) in WORLD space selects make_plane_area_emitter
        (AreaLight.cpp:56-70,143-160), otherwise make_shape_area_emitter."""
        if plane is None:
            return self._light(self.finite, LIGHT_AREA_SHAPE, radiance, (), entity)
        o, x, y = [np.asarray(a, np.float32) for a in plane]
        n = np.cross(x, y).astype(np.float32)
        area = np.float32(np.linalg.norm(n))
        n = (n / area).astype(np.float32)
        return self._light(self.finite, LIGHT_AREA_PLANE, radiance, list(o) + list(x) + list(y) + list(n) + [area], entity)

    def c_shading(self):
        arr = lambda cls, items: (cls * max(1, len(items)))(*items)
        b, m, i, f = arr(Bsdf, self.bsdfs), arr(Material, self.materials), arr(Light, self.infinite), arr(Light, self.finite)
        t = arr(Texture, self.textures)
        blob = (C.c_uint8 * max(16, len(self.blob))).from_buffer_copy(bytes(self.blob) + b"\0" * max(0, 16 - len(self.blob)))
        self._keep = (b, m, i, f, t, blob)
        sh = ShadingC(camera=self.camera, technique=self.technique)
        sh.textures, sh.texture_count = C.addressof(t), len(self.textures)
        sh.blob, sh.blob_bytes = C.addressof(blob), len(self.blob)
        sh.selector, sh.hierarchy = self.selector, self.hierarchy
        sh.bsdfs, sh.bsdf_count = C.addressof(b), len(self.bsdfs)
        sh.materials, sh.material_count = C.addressof(m), len(self.materials)
        sh.infinite_lights, sh.infinite_light_count = C.addressof(i), len(self.infinite)
        sh.finite_lights, sh.finite_light_count = C.addressof(f), len(self.finite)
        return sh
