"""hiprt-path-tracer_b200: B200-native (sm_100a) implementation of HIPRT-Path-Tracer's render hot path
(CameraRays -> [ReSTIR DI passes] -> FullPathTracer) behind the reference renderer's own data contract.

The product is libb200pt.so (C ABI, include/b200pt.h) + libb200pt_host.so (C++ GPURenderer re-hosting); this Python
package only binds them (ctypes) and carries scene inputs. Import with importlib.import_module("hiprt-path-tracer_b200")."""
from . import structs  # noqa: F401
from ._capi import B200ptError, lib, host  # noqa: F401
from .renderer import Context, GPURenderer  # noqa: F401
from .scene import Scene, load_gltf, make_camera, scene_camera, default_material, make_safe  # noqa: F401
from .luts import load_brdf_luts  # noqa: F401
