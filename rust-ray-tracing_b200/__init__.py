"""rtb200 — B200-native (sm_100a) path-tracing backend behind the Renderer boundary of
semicolonTransistor/rust-ray-tracing.  Importing this package loads librtb200.so; there is no CPU fallback."""
from . import _abi
from ._abi import Rtb200Error
from .image_io import write_png, write_ppm
from .renderer import CudaRenderer, MultiGpuRenderer, Renderer, RenderStat
from .scene import Camera, Material, Scene, Sphere, get_materials, load_object_from_toml, to_float
from . import scenes

__all__ = ["CudaRenderer", "MultiGpuRenderer", "Renderer", "RenderStat", "Camera", "Material", "Scene", "Sphere", "get_materials",
           "load_object_from_toml", "to_float", "scenes", "write_png", "write_ppm", "Rtb200Error"]
