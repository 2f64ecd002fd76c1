"""cudapathtracer_b200 — B200-native path-tracing hot path behind the scene/camera/BVH API of
CulDeVu/CUDAPathTracer.  The product is libcupt.so (C-ABI in include/cupt.h, CUDA kernels for sm_100a);
this package is its ctypes face plus the procedural scenes used by tests and bench."""
from . import api, scenes  # noqa: F401
from .api import (BvhHandle, Context, CuptError, SceneHandle, load_library, make_camera)  # noqa: F401
