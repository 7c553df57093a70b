"""mobileraytracer_b200: B200-native (sm_100a CUDA) drop-in for MobileRT's per-pixel ray-tracing hot path.
The package is a thin host-side mirror of the reference interface over libmobilert_b200.so; see include/mobilert_b200.h."""
from . import capi
from .renderer import Config, Renderer, RayTrace, stopRender, build_renderer, round_down_to_multiple_of
from .scene import Scene, Camera

__all__ = ["capi", "Config", "Renderer", "RayTrace", "stopRender", "build_renderer", "Scene", "Camera",
           "round_down_to_multiple_of"]
