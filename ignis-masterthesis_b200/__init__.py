"""ignis-b200: B200-native ray-traversal backend behind Ignis' device boundary.

The directory name contains a hyphen (it is the repo's required package name), so import it with
``importlib.import_module("ignis-masterthesis_b200")`` -- see ``tests/conftest.py``.
"""
from . import _lib, fixtures, shading, tables, workloads  # noqa: F401
from .device import Comm, Device, DeviceGroup, partition_share  # noqa: F401
from .shading import Shading  # noqa: F401
from .tables import SceneBuilder, SceneTables, make_rays  # noqa: F401

__all__ = ["Comm", "Device", "DeviceGroup", "partition_share", "SceneBuilder", "SceneTables", "Shading", "make_rays", "shading", "tables", "workloads"]
