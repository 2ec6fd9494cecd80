"""B200-native render hot path of LevKruglyak/RayTracingRust behind the crate's scene API.

`scene` mirrors the reference's scene-construction API; `render` is the GPU replacement of
`render(&mut RenderTarget, &Scene)`.  Importing the package does not load CUDA; the first
render/loader call loads `librtr_b200.so` and fails loudly if it is missing.
"""
from . import _abi
from .scene import (Camera, Color, Dielectric, Emission, GradientBackground, Isotropic, Lambertian,
                    MaterialHandle, Mesh, Metal, MixMaterial, ObjectHandle, RenderMode, RenderSettings,
                    RenderTarget, Scene, SkyMap, Sphere, UniformBackground, Vec3, Volume)
from .render import Context, DeviceScene, Group, GroupScene, RtrError, make_params, render, load_obj, load_scene_json, lib, write_png, write_exr, load_image

__all__ = [n for n in dir() if not n.startswith("_")]
