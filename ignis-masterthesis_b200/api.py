"""The reference's Python API (`import ignis`, src/frontend/python/runtime.cpp:112-265) over the B200 backend.

Same names and behaviour for the members on the hot path: `loadFromString` / `loadFromFile` return a `RuntimeWrap`
(context manager, one runtime per process at a time -- runtime.cpp:76-101), the runtime has `step`, `trace`, `reset`,
`getFramebufferForHost(aov="")` (RGB SUMS over iterations, shape (h, w, 3)), `getFramebufferForDevice`, `tonemap`,
`clearFramebuffer`, and the properties `IterationCount`, `SampleCount`, `FrameCount`, `FramebufferWidth/Height`, `SPI`,
`Technique`, `Target`.  The reference's own integrator tests (src/tests/integrator) run against this module
unchanged (tests/test_reference_integrator_gpu.py).

Scene loading is `sceneio.load_scene` (the JSON subset of the five BASELINE configs); the reference's loader proper is
out of scope.  Everything after loading goes through the C ABI; there is no CPU fallback.
"""
import json
import math
import os

import numpy as np

from . import sceneio
from .device import Device
from .tables import RAY_DTYPE

_FLT_MAX = float(np.finfo(np.float32).max)


class Target:
    """Target.h: only the GPU ordinal matters to this backend."""

    def __init__(self, device=0):
        self.Device = int(device)

    IsValid, IsCPU, IsGPU = True, False, True

    def toString(self):
        return f"b200:{self.Device}"

    __str__ = toString

    @staticmethod
    def pickGPU(device=0):
        return Target(device)

    pickBest = makeGPU = pickGPU


class RuntimeOptions:
    """RuntimeSettings.h: the options of runtime.cpp:151-165 that reach the device."""

    def __init__(self):
        self.Target = Target(0)
        self.AcquireStats = False
        self.SPI = 0                   # 0 = recommendSPI (Runtime.cpp:61-69)
        self.Seed = 0
        self.OverrideTechnique = ""
        self.OverrideCamera = ""
        self.OverrideFilmSize = None   # (width, height)
        self.EnableTonemapping = True
        self.IsTracer = False
        self.DumpShader = self.DumpShaderFull = False
        self.WarnUnused = True
        # not in the reference: selects the one-sided lambertian eval its own goldens assume (note 2)
        self.OneSidedDiffuse = False

    @staticmethod
    def makeDefault(trace=False):
        o = RuntimeOptions()
        o.IsTracer = bool(trace)
        return o


class Ray:
    """RuntimeStructs.h:71-75; runtime.cpp:167-172."""

    def __init__(self, org, dir, tmin=0.0, tmax=_FLT_MAX):
        self.Origin = np.asarray(org, np.float32)
        self.Direction = np.asarray(dir, np.float32)
        self.Range = np.asarray([tmin, tmax], np.float32)


def _recommend_spi(width, height, interactive=False):
    spi_f = 8 // (2 if interactive else 1)
    return max(1, min(64, int(math.ceil(spi_f / ((width / 1000.0) * (height / 1000.0))))))


class Runtime:
    """IG::Runtime (src/runtime/Runtime.cpp:121-448) as the Python API exposes it."""

    def __init__(self, opts, scene, base):
        self._opts = opts
        size = opts.OverrideFilmSize
        tables, sh, w, h, _ = sceneio.load_scene(scene, base, *(size if size else (None, None)))
        if opts.OverrideTechnique:
            t = dict(scene.get("technique", {})) if isinstance(scene, dict) else {}
            t["type"] = opts.OverrideTechnique
            if opts.OverrideTechnique == "debug":
                sh.set_debug(t.get("mode", 0))
            elif opts.OverrideTechnique == "infobuffer":
                sh.set_infobuffer(t.get("max_depth", 64), bool(t.get("handle_specular", False)))
            elif opts.OverrideTechnique != "path":
                raise ValueError(f"technique '{opts.OverrideTechnique}' is outside the supported subset")
        if opts.OneSidedDiffuse:
            from .shading import TECHFLAG_ONE_SIDED_DIFFUSE
            sh.technique.flags |= TECHFLAG_ONE_SIDED_DIFFUSE
        self._sh = sh
        self._w, self._h = (1, 1) if opts.IsTracer else (w, h)
        from .shading import TECH_INFOBUFFER
        self._lock_film = sh.technique.type == TECH_INFOBUFFER      # InfoBufferTechnique.cpp:36
        self._spi = 1 if self._lock_film else (opts.SPI or (1 if opts.IsTracer else _recommend_spi(w, h)))
        self._iteration = self._samples = self._frame = 0
        self._dev = Device(opts.Target.Device, acquire_stats=opts.AcquireStats)
        try:
            self._dev.assign_scene(tables)
            self._dev.assign_shading(sh)
            self._dev.resize(self._w, self._h)
            self._dev.clear_framebuffer()
        except Exception:
            self._dev.close()
            raise

    # ---- Runtime.cpp:294-349
    def step(self, ignoreDenoiser=False):
        if self._opts.IsTracer:
            raise RuntimeError("Trying to use step() in a trace driver!")
        # pipelined step: returns once the iteration's rays are generated; every framebuffer access completes the film first
        self._dev.render(self._w, self._h, spi=self._spi, iteration=self._iteration, frame=self._frame, seed=self._opts.Seed,
                         pipelined=True)
        self._iteration += 1
        if not self._lock_film:
            self._samples += self._spi

    # ---- Runtime.cpp:351-383, runtime.cpp:181-185: the film of an N x 1 render, (N, 3), sums over iterations
    def trace(self, rays):
        if not self._opts.IsTracer:
            raise RuntimeError("Trying to use trace() in a camera driver!")
        if isinstance(rays, np.ndarray) and rays.dtype == RAY_DTYPE:
            packed = rays
        else:
            packed = np.zeros(len(rays), RAY_DTYPE)
            for i, r in enumerate(rays):
                packed[i] = (r.Origin, r.Direction, r.Range[0], r.Range[1])
        self._dev.render(spi=self._spi, rays=packed, iteration=self._iteration, frame=self._frame, seed=self._opts.Seed)
        self._w, self._h = packed.size, 1
        self._iteration += 1
        self._samples += self._spi
        return self.getFramebufferForHost().reshape(-1, 3)

    def reset(self):  # Runtime.cpp:437-443
        self._dev.clear_framebuffer()
        self._iteration = self._samples = 0

    def getFramebufferForHost(self, aov=""):
        film, _ = self._dev.framebuffer(copy=False, name=aov or None)
        return film

    def getFramebufferForDevice(self, aov=""):
        """Device pointer + shape (no copy); wrap with `torch.as_tensor` through __cuda_array_interface__."""
        if aov:
            raise NotImplementedError("device access to named AOVs goes through igb200_framebuffer(name, device=1)")
        ptr, _ = self._dev.framebuffer_device()
        return _DeviceImage(ptr, (self._h, self._w, 3))

    def tonemap(self, output):  # runtime.cpp:227-242: method 0, gamma, scale = 1 / iterations
        if output.shape != (self._h, self._w):
            raise BufferError("Incompatible buffer: Buffer has not the correct shape matching the framebuffer size")
        output[...] = self._dev.tonemap(None, 0, True, 1.0 / self._iteration if self._iteration else 1.0, 1.0, 0.0)

    def clearFramebuffer(self, aov=None):
        self._dev.clear_framebuffer(aov)

    def incFrameCount(self):
        self._frame += 1

    IterationCount = property(lambda self: self._iteration)
    SampleCount = property(lambda self: self._samples)
    FrameCount = property(lambda self: self._frame)
    FramebufferWidth = property(lambda self: self._w)
    FramebufferHeight = property(lambda self: self._h)
    SPI = property(lambda self: self._spi)
    Target = property(lambda self: self._opts.Target)
    Technique = property(lambda self: {0: "path", 1: "debug", 2: "infobuffer"}[int(self._sh.technique.type)])
    Camera = property(lambda self: "perspective")
    AvailableCameraTypes = ["perspective"]
    AvailableTechniqueTypes = ["path", "debug", "infobuffer"]

    def renderStatistics(self):
        """Ray counters of Statistics.h:56-63 for the last step (ours: the reference dumps them as text)."""
        return self._dev.render_stats()

    def _shutdown(self):
        self._dev.close()


class _DeviceImage:
    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (ptr, False), "version": 3, "strides": None}


class RuntimeWrap:
    """runtime.cpp:27-107: creates the runtime on __enter__ / .instance, destroys it on __exit__ / shutdown();
    a second live runtime is refused ('Trying to create multiple runtime instances!')."""
    _live = None

    def __init__(self, opts, scene, base):
        self._opts, self._scene, self._base = opts, scene, base
        self._mine = None

    def __enter__(self):
        return self._create()

    def __exit__(self, *exc):
        self.shutdown()
        return False

    @property
    def instance(self):
        return self._mine if self._mine is not None and RuntimeWrap._live is self._mine else self._create()

    def shutdown(self):
        if RuntimeWrap._live is not None and RuntimeWrap._live is self._mine:
            RuntimeWrap._live._shutdown()
            RuntimeWrap._live = None
        self._mine = None

    def _create(self):
        if RuntimeWrap._live is not None:
            print("[ERROR] Trying to create multiple runtime instances!")
            return None
        self._mine = RuntimeWrap._live = Runtime(self._opts, self._scene, self._base)
        return self._mine


def loadFromString(source, opts=None, base="."):
    """runtime.cpp: `loadFromString(str[, opts])`; an empty scene `{}` is valid (test_init.py:9-12)."""
    return RuntimeWrap(opts or RuntimeOptions.makeDefault(), json.loads(source), base)


def loadFromFile(path, opts=None):
    with open(path) as fh:
        scene = json.load(fh)
    return RuntimeWrap(opts or RuntimeOptions.makeDefault(), scene, os.path.dirname(os.path.abspath(path)))


def flushLog():
    import sys
    sys.stdout.flush()
    sys.stderr.flush()
