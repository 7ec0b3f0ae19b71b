"""Scene description types — Python mirror of the reference's L1 layer.

Mirrors (paths relative to the reference root): `Vec3::from_toml` src/geometry.rs:190-211, `Color::from_toml`
src/color.rs:99-123, `to_float` src/toml_utils.rs:2-11, `get_materials`/`load_material_from_toml`
src/materials.rs:12-33, the three `from_table` loaders (:65-68, :99-106, :149-154), `get_object_list` /
`load_object_from_toml` / `Sphere::from_table` src/objects.rs:38-52,292-299, `Scene::from_list`
src/ray_tracing.rs:222-224 and `Camera::new` / `Camera::from_toml` (:27-75).  Error behaviour follows the
reference: what panics there raises here (KeyError / TypeError / ValueError with the reference's message where it has one).
"""
from __future__ import annotations

import ctypes as C
import tomllib
from dataclasses import dataclass, field
from typing import Dict, List, Sequence

from . import _abi_types as _abi     # structs and constants only: the library itself is loaded where it is called (Camera)


def to_float(value) -> float:
    """src/toml_utils.rs:2-11 — a TOML float or integer; anything else is an error (the reference unwraps None)."""
    if isinstance(value, bool) or not isinstance(value, (int, float)):
        raise TypeError(f"expected a TOML float or integer, got {value!r}")
    return float(value)


def vec3_from_toml(value) -> tuple:
    """Vec3::from_toml, src/geometry.rs:190-211: table {x,y,z} or array of >= 3 numbers."""
    if isinstance(value, dict):
        return (to_float(value["x"]), to_float(value["y"]), to_float(value["z"]))
    if isinstance(value, list):
        if len(value) < 3:
            raise ValueError("assertion failed: array.len() >= 3")
        return (to_float(value[0]), to_float(value[1]), to_float(value[2]))
    raise TypeError(f"expected a table or an array for a vector, got {value!r}")


def color_from_toml(value) -> tuple:
    """Color::from_toml, src/color.rs:99-123: table {red,green,blue} (a missing key is 0.0) or array of >= 3."""
    if isinstance(value, dict):
        def get(k):
            v = value.get(k)
            if v is None or isinstance(v, bool) or not isinstance(v, (int, float)):
                return 0.0   # to_float(...).unwrap_or(0.0) — also reached for a wrongly typed value
            return float(v)
        if not all(k in value for k in ("red", "green", "blue")):
            # indexing a missing key panics in the reference (table["red"]); only a wrongly *typed* value falls back
            missing = [k for k in ("red", "green", "blue") if k not in value]
            raise KeyError(missing[0])
        return (get("red"), get("green"), get("blue"))
    if isinstance(value, list):
        if len(value) < 3:
            raise ValueError("assertion failed: array.len() >= 3")
        return (to_float(value[0]), to_float(value[1]), to_float(value[2]))
    raise TypeError(f"expected a table or an array for a colour, got {value!r}")


@dataclass
class Material:
    """One of the reference's three `impl Material` types (src/materials.rs:43-147)."""
    kind: int
    albedo: tuple = (0.0, 0.0, 0.0)
    param: float = 0.0       # Metal::fuzzy_factor | Dielectric::index_of_refraction
    hollow: bool = False

    @staticmethod
    def lambertian(albedo) -> "Material":
        return Material(_abi.LAMBERTIAN, tuple(float(c) for c in albedo))

    @staticmethod
    def metal(albedo, fuzzy_factor: float) -> "Material":
        f = float(fuzzy_factor)
        return Material(_abi.METAL, tuple(float(c) for c in albedo), f if f < 1.0 else 1.0)   # Metal::new, :79-88

    @staticmethod
    def dielectric(index_of_refraction: float, hollow: bool = False) -> "Material":
        return Material(_abi.DIELECTRIC, (0.0, 0.0, 0.0), float(index_of_refraction), bool(hollow))

    @staticmethod
    def from_table(table: dict) -> "Material":
        """load_material_from_toml, src/materials.rs:21-33."""
        kind = table["type"]
        if not isinstance(kind, str):
            raise TypeError("material type must be a string")
        kind = kind.lower()   # to_ascii_lowercase
        if kind == "lambertian":
            return Material.lambertian(color_from_toml(table["albedo"]))
        if kind == "metal":
            return Material.metal(color_from_toml(table["albedo"]), to_float(table["fuzzy_factor"]))
        if kind == "dielectric":
            hollow = table["hollow"]
            if not isinstance(hollow, bool):
                raise TypeError("hollow must be a boolean")
            return Material.dielectric(to_float(table["index_of_refraction"]), hollow)
        raise ValueError(f"Unknown material type {kind}!")

    def to_toml(self, name: str) -> str:
        if self.kind == _abi.LAMBERTIAN:
            return f'[materials.{name}]\ntype = "lambertian"\nalbedo = [{self.albedo[0]!r}, {self.albedo[1]!r}, {self.albedo[2]!r}]\n'
        if self.kind == _abi.METAL:
            return (f'[materials.{name}]\ntype = "metal"\nalbedo = [{self.albedo[0]!r}, {self.albedo[1]!r}, {self.albedo[2]!r}]\n'
                    f'fuzzy_factor = {self.param!r}\n')
        return (f'[materials.{name}]\ntype = "dielectric"\nindex_of_refraction = {self.param!r}\n'
                f'hollow = {"true" if self.hollow else "false"}\n')


def get_materials(table: dict) -> Dict[str, Material]:
    """get_materials, src/materials.rs:12-19."""
    return {key: Material.from_table(value) for key, value in table.items()}


@dataclass
class Sphere:
    """Sphere { center, radius, material } — src/objects.rs:203-214."""
    center: tuple
    radius: float
    material: str

    @staticmethod
    def from_table(table: dict, material_table: Dict[str, Material]) -> "Sphere":
        """Sphere::from_table, src/objects.rs:292-299."""
        center = vec3_from_toml(table["center"])
        radius = to_float(table["radius"])
        name = table["material"]
        if not isinstance(name, str):
            raise TypeError("material must be a string")
        if name not in material_table:
            raise KeyError(name)   # material_table.get(name).unwrap()
        return Sphere(center, radius, name)


def load_object_from_toml(table: dict, material_table: Dict[str, Material]) -> Sphere:
    """load_object_from_toml, src/objects.rs:44-52."""
    kind = table["type"]
    if not isinstance(kind, str):
        raise TypeError("object type must be a string")
    kind = kind.lower()
    if kind == "sphere":
        return Sphere.from_table(table, material_table)
    raise ValueError(f"Unknown object type {kind}")


@dataclass
class Scene:
    """Scene { objects: Vec<Object> } — src/ray_tracing.rs:100-104.  Object order is significant."""
    materials: Dict[str, Material] = field(default_factory=dict)
    objects: List[Sphere] = field(default_factory=list)

    @staticmethod
    def from_list(objects: Sequence[Sphere], materials: Dict[str, Material]) -> "Scene":
        return Scene(dict(materials), list(objects))

    @staticmethod
    def from_toml(text: str) -> "Scene":
        """What main does with scene.toml: src/main.rs:36-43."""
        data = tomllib.loads(text)
        materials = get_materials(data["materials"])
        hitables = data["hitables"]
        if not isinstance(hitables, list):
            raise TypeError("hitables must be an array of tables")
        objects = [load_object_from_toml(t, materials) for t in hitables]
        return Scene(materials, objects)

    @staticmethod
    def from_file(path: str) -> "Scene":
        with open(path, "rb") as f:
            return Scene.from_toml(f.read().decode("utf-8"))

    def to_toml(self) -> str:
        out = [m.to_toml(name) for name, m in self.materials.items()]
        for s in self.objects:
            out.append(f'[[hitables]]\ntype = "sphere"\ncenter = [{s.center[0]!r}, {s.center[1]!r}, {s.center[2]!r}]\n'
                       f'radius = {s.radius!r}\nmaterial = "{s.material}"\n')
        return "\n".join(out)

    def __len__(self) -> int:   # Scene::len, src/ray_tracing.rs:308-310
        return len(self.objects)

    # ---- flattening to the C ABI -----------------------------------------------------------------------
    def to_abi(self):
        """Returns (rtb200_scene, keepalive) — AoS arrays in object order, materials in table order.  The flattening is
        done once per Scene (the reference builds its Vec<Object> once in Scene::from_list, src/ray_tracing.rs:299-306);
        a Scene whose object or material count changed is flattened again."""
        key = (len(self.objects), len(self.materials))
        cached = self.__dict__.get("_abi_cache")
        if cached is not None and cached[0] == key:
            return cached[1], cached[2]
        names = list(self.materials.keys())
        index = {n: i for i, n in enumerate(names)}
        mats = (_abi.Material * max(1, len(names)))()
        for i, n in enumerate(names):
            m = self.materials[n]
            mats[i].kind = m.kind
            mats[i].hollow = 1 if m.hollow else 0
            mats[i].albedo[0], mats[i].albedo[1], mats[i].albedo[2] = m.albedo
            mats[i].param = m.param
        sph = (_abi.Sphere * max(1, len(self.objects)))()
        for i, s in enumerate(self.objects):
            sph[i].center[0], sph[i].center[1], sph[i].center[2] = s.center
            sph[i].radius = s.radius
            sph[i].material = index[s.material]
        sc = _abi.Scene(C.cast(sph, C.POINTER(_abi.Sphere)), len(self.objects), C.cast(mats, C.POINTER(_abi.Material)), len(names))
        self.__dict__["_abi_cache"] = (key, sc, (sph, mats))
        return sc, (sph, mats)

    def arrays(self):
        """Plain numpy view of the scene (for test harnesses): spheres[n,4] f64, material index[n] i32,
        kind[m] i32, hollow[m] i32, albedo[m,3] f64, param[m] f64."""
        import numpy as np
        names = list(self.materials.keys())
        index = {n: i for i, n in enumerate(names)}
        spheres = np.array([[*s.center, s.radius] for s in self.objects], dtype=np.float64).reshape(-1, 4)
        mat_index = np.array([index[s.material] for s in self.objects], dtype=np.int32)
        kind = np.array([self.materials[n].kind for n in names], dtype=np.int32)
        hollow = np.array([1 if self.materials[n].hollow else 0 for n in names], dtype=np.int32)
        albedo = np.array([self.materials[n].albedo for n in names], dtype=np.float64).reshape(-1, 3)
        param = np.array([self.materials[n].param for n in names], dtype=np.float64)
        return spheres, mat_index, kind, hollow, albedo, param


class Camera:
    """Camera — src/ray_tracing.rs:15-98.  Holds the eight derived f64 quantities (rtb200_camera)."""

    def __init__(self, image_width: int, image_height: int, focal_length: float, view_angle: float,
                 center, look_at, up, defocus_angle: float):
        """Camera::new, src/ray_tracing.rs:27-62 (evaluated by rtb200_camera_init in f64)."""
        from . import _abi as _lib     # Camera::new runs inside the library (rtb200_camera_init): loads librtb200.so
        self.abi = _abi.Camera()
        d3 = C.c_double * 3
        _lib.check(_lib.lib.rtb200_camera_init(C.byref(self.abi), int(image_width), int(image_height), float(focal_length),
                                               float(view_angle), d3(*center), d3(*look_at), d3(*up), float(defocus_angle)))

    @staticmethod
    def from_toml(table: dict) -> "Camera":
        """Camera::from_toml, src/ray_tracing.rs:64-75.  focal_length / view_angle must be TOML floats (as_float).
        The reference reads defocus_angle from the "view_angle" key (:72, a slip: SURVEY F12); here it has its own
        key, `defocus_angle`, default 0.0."""
        for k in ("image_width", "image_height"):
            if isinstance(table[k], bool) or not isinstance(table[k], int):
                raise TypeError(f"{k} must be an integer")
        for k in ("focal_length", "view_angle"):
            if not isinstance(table[k], float):
                raise TypeError(f"{k} must be a float")
        defocus = table.get("defocus_angle", 0.0)
        return Camera(table["image_width"], table["image_height"], table["focal_length"], table["view_angle"],
                      vec3_from_toml(table["center"]), vec3_from_toml(table["look_at"]), vec3_from_toml(table["up"]),
                      to_float(defocus))

    @staticmethod
    def default(image_width: int = 3840, image_height: int = 2160) -> "Camera":
        """The camera main hard-codes, src/main.rs:45-58."""
        return Camera(image_width, image_height, 10.0, 30.0, (16.0, 2.0, 18.5), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 0.0)

    def image_width(self) -> int:    # src/ray_tracing.rs:91-93
        return self.abi.image_width

    def image_height(self) -> int:   # :95-97
        return self.abi.image_height

    def vectors(self):
        """center, ulc, viewport_u, viewport_v, defocus_u, defocus_v as a flat list of 18 floats."""
        a = self.abi
        out = []
        for v in (a.center, a.viewport_upper_left_corner, a.viewport_u, a.viewport_v, a.defocus_u, a.defocus_v):
            out.extend([v[0], v[1], v[2]])
        return out
