"""Host-side mirror of the reference crate's scene API (same names, argument meaning and
error behaviour), lowering to the plain `rtr_scene_desc` of include/rtr.h.

Mirrors (reference file:line):
  Scene / RenderSettings / RenderMode / handles   lib/core/scene.rs:15-114
  Camera                                         lib/core/camera.rs:27-34
  Sphere                                         lib/objects.rs:15-20
  Mesh.from_file / from_buffers / build_bvh      lib/core/mesh.rs:63-121
  Volume, Isotropic                              lib/volume.rs:19-31,75-87
  Lambertian, Metal, Emission, Dielectric, Mix   lib/materials.rs:18-156
  UniformBackground, GradientBackground, SkyMap  lib/backgrounds.rs:14-92
  RenderTarget                                   lib/core/render.rs:14-30

This module only describes scenes; all rendering happens in the CUDA library (render.py).
Where the reference panics (unwrap/expect/index out of range) these raise.
"""
import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

from . import _abi as A
from ._abi import Vec3  # re-exported: cgmath Vector3<f32>


@dataclass
class Color:
    """lib/utils/color.rs:6-11"""
    r: float = 0.0
    g: float = 1.0  # Default is green (color.rs:43-47)
    b: float = 0.0

    def vec3(self):
        return Vec3(self.r, self.g, self.b)


class RenderMode:
    Full, Clay, Random, Normal = A.MODE_FULL, A.MODE_CLAY, A.MODE_RANDOM, A.MODE_NORMAL
    _names = {"Full": 0, "Clay": 1, "Random": 2, "Normal": 3}


@dataclass
class RenderSettings:
    samples_per_pixel: int = 5
    max_ray_depth: int = 6
    clamp_indirect: float = 10.0
    enable_multithreading: bool = True
    enable_bvh_tree: bool = True
    mode: int = RenderMode.Full

    def to_c(self):
        if not 0 <= self.max_ray_depth <= 255:
            raise ValueError("max_ray_depth is a u8 in the reference")
        return A.SettingsS(self.samples_per_pixel, self.max_ray_depth, self.clamp_indirect,
                           int(self.enable_multithreading), int(self.enable_bvh_tree), self.mode)

    @staticmethod
    def from_c(s):
        return RenderSettings(s.samples_per_pixel, s.max_ray_depth, s.clamp_indirect,
                              bool(s.enable_multithreading), bool(s.enable_bvh_tree), s.mode)


@dataclass
class Camera:
    lookfrom: Vec3 = field(default_factory=lambda: Vec3(0, 0, 0))
    lookat: Vec3 = field(default_factory=lambda: Vec3(0, 0, -1))
    vertical: Vec3 = field(default_factory=lambda: Vec3(0, 1, 0))
    vertical_fov: float = 90.0
    aspect_ratio: float = 1.0

    def to_c(self):
        return A.CameraS(self.lookfrom, self.lookat, self.vertical, self.vertical_fov, self.aspect_ratio)

    @staticmethod
    def from_c(c):
        return Camera(Vec3(*c.lookfrom.tuple()), Vec3(*c.lookat.tuple()), Vec3(*c.vertical.tuple()),
                      c.vertical_fov, c.aspect_ratio)


@dataclass(frozen=True)
class MaterialHandle:
    index: int


@dataclass(frozen=True)
class ObjectHandle:
    index: int


# ------------------------------------------------------------------ materials
class Material:
    pass


@dataclass
class Lambertian(Material):
    albedo: Color


@dataclass
class Metal(Material):
    albedo: Color
    fuzz: float


class Emission(Material):
    def __init__(self, color: Color, strength: float = 1.0):
        # Emission::new pre-multiplies (materials.rs:79-83); f32 product
        s = np.float32(strength)
        self.color = Color(float(np.float32(color.r) * s), float(np.float32(color.g) * s),
                           float(np.float32(color.b) * s))


@dataclass
class Dielectric(Material):
    ir: float


@dataclass
class MixMaterial(Material):
    first: Material
    second: Material
    factor: float


@dataclass
class Isotropic(Material):
    color: Color


# ------------------------------------------------------------------ objects
class Object:
    pass


@dataclass
class Sphere(Object):
    center: Vec3
    radius: float
    material: MaterialHandle


class Mesh(Object):
    def __init__(self, positions, normals, indices, material: MaterialHandle):
        self.positions = np.ascontiguousarray(positions, dtype=np.float32).reshape(-1, 3)
        self.normals = np.ascontiguousarray(normals, dtype=np.float32).reshape(-1, 3)
        self.indices = np.ascontiguousarray(indices, dtype=np.uint32).reshape(-1)
        if self.positions.shape != self.normals.shape:
            raise ValueError("positions/normals mismatch")
        if self.indices.size and int(self.indices.max()) >= len(self.positions):
            raise IndexError("vertex index out of range")  # the reference would panic on index
        self.material = material
        self._bvh_built = False

    @staticmethod
    def from_buffers(positions, normals, indices, material):
        """Mesh::from_buffers (mesh.rs:78-117)"""
        return Mesh(positions, normals, indices, material)

    @staticmethod
    def from_file(path: str, material: MaterialHandle):
        """Mesh::from_file (mesh.rs:63-76) via the library's OBJ loader."""
        from .render import load_obj
        pos, nrm, idx = load_obj(path)
        return Mesh(pos, nrm, idx, material)

    def build_bvh(self):
        """mesh.rs:119-121 — without it bvh_root stays None and the mesh is never hit."""
        self._bvh_built = True

    @property
    def n_triangles(self):
        return self.indices.size // 3


class Volume(Object):
    def __init__(self, boundary: Object, density: float):
        """Volume::new (volume.rs:26-31): stores -1.0 / density (f32)."""
        self.boundary = boundary
        self.neg_inv_density = float(np.float32(-1.0) / np.float32(density))


# ------------------------------------------------------------------ backgrounds
class Background:
    pass


@dataclass
class UniformBackground(Background):
    color: Color


@dataclass
class GradientBackground(Background):
    top: Color
    bottom: Color


class SkyMap(Background):
    """SkyMap (backgrounds.rs:51-92): row-major RGB f32 texels, image[x + y*width]."""

    def __init__(self, image, width: Optional[int] = None, height: Optional[int] = None):
        if isinstance(image, str):
            image = _read_exr(image)  # SkyMap::new(path) (backgrounds.rs:59-81)
        img = np.ascontiguousarray(image, dtype=np.float32)
        if img.ndim == 3:
            height, width = img.shape[:2]
            img = img[:, :, :3]
        self.width, self.height = int(width), int(height)
        self.image = np.ascontiguousarray(img).reshape(self.height * self.width, 3)


def _read_exr(path):
    from .render import load_image  # the host library's own OpenEXR / Radiance reader (rtr_image_load)
    try:
        return load_image(path)
    except Exception as e:
        raise IOError(f"could not read image! ({e})")  # .expect (backgrounds.rs:72)


# ------------------------------------------------------------------ render target
class RenderTarget:
    """RenderTarget::new (render.rs:21-30): zero-initialised RGBA8."""

    def __init__(self, width: int, height: int):
        self.width, self.height = width, height
        self.data = np.zeros(width * height * 4, dtype=np.uint8)
        self.request_redraw = False
        self.accum = None   # optional f32 accumulation buffer (extension)
        self.stats = None

    def image(self):
        return self.data.reshape(self.height, self.width, 4)

    def save_png(self, path):
        """Headless replacement of the window blit (src/main.rs:92-97)."""
        from .render import write_png
        write_png(path, self.image())


# ------------------------------------------------------------------ scene
class LoweredScene:
    """Owns the ctypes arrays behind an rtr_scene_desc (keeps the numpy buffers alive)."""

    def __init__(self):
        self.desc = A.SceneDesc()
        self._keep = []


class Scene:
    def __init__(self, settings: RenderSettings, camera: Camera, background: Background):
        """Scene::new (scene.rs:78-86)"""
        self.camera = camera
        self.settings = settings
        self.background = background
        self.objects: List[Object] = []
        self._materials: List[Material] = []

    @staticmethod
    def default():
        """Scene::default (scene.rs:42-65)"""
        return Scene(RenderSettings(), Camera(), UniformBackground(Color(0.8, 0.8, 0.8)))

    @staticmethod
    def from_file(path: str):
        """Scene::from_file (scene.rs:74-76) via the library's JSON loader."""
        from .render import load_scene_json
        return load_scene_json(path)

    def add_object(self, obj: Object) -> ObjectHandle:
        self.objects.append(obj)
        return ObjectHandle(len(self.objects) - 1)

    def object(self, h: ObjectHandle) -> Object:
        return self.objects[h.index]

    def add_material(self, m: Material) -> MaterialHandle:
        self._materials.append(m)
        return MaterialHandle(len(self._materials) - 1)

    def material(self, h: MaterialHandle) -> Material:
        return self._materials[h.index]

    # -------------------------------------------------------------- lowering
    def lower(self) -> LoweredScene:
        out = LoweredScene()
        mats: List[A.MaterialS] = []

        def lower_material(m, slot=None):
            if slot is None:
                slot = len(mats)
                mats.append(A.MaterialS())
            rec = A.MaterialS()
            if isinstance(m, Lambertian):
                rec.kind, rec.color = A.MAT_LAMBERTIAN, m.albedo.vec3()
            elif isinstance(m, Metal):
                rec.kind, rec.color, rec.param = A.MAT_METAL, m.albedo.vec3(), m.fuzz
            elif isinstance(m, Emission):
                rec.kind, rec.color = A.MAT_EMISSION, m.color.vec3()
            elif isinstance(m, Dielectric):
                rec.kind, rec.param = A.MAT_DIELECTRIC, m.ir
            elif isinstance(m, Isotropic):
                rec.kind, rec.color = A.MAT_ISOTROPIC, m.color.vec3()
            elif isinstance(m, MixMaterial):
                rec.kind, rec.param = A.MAT_MIX, m.factor
                rec.first = lower_material(m.first)
                rec.second = lower_material(m.second)
            else:
                raise TypeError(f"unknown material {m!r}")
            mats[slot] = rec
            return slot

        n_handles = len(self._materials)
        mats.extend(A.MaterialS() for _ in range(n_handles))  # handles keep their indices
        for i, m in enumerate(self._materials):
            lower_material(m, i)

        sph_chunks, meshes, volumes, obj_chunks = [], [], [], []
        n_sph = [0]
        SPH = np.dtype([("c", np.float32, 3), ("r", np.float32), ("m", np.uint32)])

        def add_spheres(centers, radii, mat):
            n = len(radii)
            a = np.zeros(n, SPH)
            a["c"], a["r"], a["m"] = centers, radii, mat
            sph_chunks.append(a)
            n_sph[0] += n
            return n_sph[0] - n

        def check_mat(h):
            idx = h.index if isinstance(h, MaterialHandle) else int(h)
            if not 0 <= idx < n_handles:
                raise IndexError("material handle out of range")
            return idx

        def lower_object(o, allow_volume=True):
            if isinstance(o, Sphere):
                first = add_spheres([o.center.tuple()], [o.radius], check_mat(o.material))
                return (A.OBJ_SPHERE, first)
            if isinstance(o, Mesh):
                ms = A.MeshS()
                ms.positions = o.positions.ctypes.data_as(C.POINTER(A.f32))
                ms.normals = o.normals.ctypes.data_as(C.POINTER(A.f32))
                ms.indices = o.indices.ctypes.data_as(C.POINTER(A.u32))
                ms.n_vertices = len(o.positions)
                # no build_bvh() -> BvhNode::None -> never hit; bounds still come from the vertices
                ms.n_indices = (o.indices.size // 3) * 3 if o._bvh_built else 0
                ms.material = check_mat(o.material)
                out._keep += [o.positions, o.normals, o.indices]
                meshes.append(ms)
                return (A.OBJ_MESH, len(meshes) - 1)
            if isinstance(o, Volume) and allow_volume:
                # boundary: Box<dyn Object> (volume.rs:20-23): a Sphere, a Mesh or again a Volume
                b = lower_object(o.boundary, allow_volume=isinstance(o.boundary, Volume))
                volumes.append(A.VolumeS(A.ObjectRef(*b), o.neg_inv_density))
                return (A.OBJ_VOLUME, len(volumes) - 1)
            raise TypeError(f"unknown object {o!r}")

        for o in self.objects:
            if isinstance(o, _SphereBlock):
                first = add_spheres(o.centers, o.radii, check_mat(o.material))
                refs = np.empty((len(o.radii), 2), np.uint32)
                refs[:, 0], refs[:, 1] = A.OBJ_SPHERE, np.arange(first, first + len(o.radii))
                obj_chunks.append(refs)
            else:
                obj_chunks.append(np.array([lower_object(o)], np.uint32).reshape(1, 2))
        spheres = np.concatenate(sph_chunks) if sph_chunks else np.zeros(0, SPH)
        objects = np.concatenate(obj_chunks) if obj_chunks else np.zeros((0, 2), np.uint32)
        spheres = np.ascontiguousarray(spheres)
        objects = np.ascontiguousarray(objects)
        out._keep += [spheres, objects]

        def arr(ctype, items):
            a = (ctype * max(len(items), 1))(*items)
            out._keep.append(a)
            return a

        d = out.desc
        assert SPH.itemsize == C.sizeof(A.SphereS)
        d.objects, d.n_objects = objects.ctypes.data_as(C.POINTER(A.ObjectRef)), len(objects)
        d.spheres, d.n_spheres = spheres.ctypes.data_as(C.POINTER(A.SphereS)), len(spheres)
        d.meshes, d.n_meshes = arr(A.MeshS, meshes), len(meshes)
        d.volumes, d.n_volumes = arr(A.VolumeS, volumes), len(volumes)
        d.materials, d.n_materials = arr(A.MaterialS, mats), len(mats)

        bg = self.background
        if isinstance(bg, UniformBackground):
            d.background.kind, d.background.color = A.BG_UNIFORM, bg.color.vec3()
        elif isinstance(bg, GradientBackground):
            d.background.kind = A.BG_GRADIENT
            d.background.color, d.background.bottom = bg.top.vec3(), bg.bottom.vec3()
        elif isinstance(bg, SkyMap):
            d.background.kind = A.BG_SKYMAP
            d.background.texels = bg.image.ctypes.data_as(C.POINTER(A.f32))
            d.background.width, d.background.height = bg.width, bg.height
            out._keep.append(bg.image)
        else:
            raise TypeError(f"unknown background {bg!r}")
        return out

    def add_spheres_bulk(self, centers, radii, material: MaterialHandle):
        """Convenience for the 8 000 / 1 M sphere stress scenes (benches/bvh_tree_stress.rs:18-28)."""
        self.objects.append(_SphereBlock(np.ascontiguousarray(centers, np.float32).reshape(-1, 3),
                                         np.ascontiguousarray(radii, np.float32).reshape(-1), material))


class _SphereBlock(Object):
    """N spheres added in order; lowered to N consecutive Sphere objects."""

    def __init__(self, centers, radii, material):
        self.centers, self.radii, self.material = centers, radii, material
