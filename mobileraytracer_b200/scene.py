"""Host-side scene containers and the reference's four built-in scenes as input data.

Mirrors MobileRT::Scene (app/MobileRT/Scene.hpp:20-27) and app/Scenes/Scenes.cpp:63-302. All arithmetic that feeds
the device (triangle edges and face normals) is done in float32 in the reference's operation order
(Triangle::Builder, app/MobileRT/Shapes/Triangle.cpp:328-340)."""
import numpy as np
from . import capi

f32 = np.float32


def _v(x):
    return np.asarray(x, dtype=f32)


def _dot(a, b):
    return f32(f32(f32(a[0] * b[0]) + f32(a[1] * b[1])) + f32(a[2] * b[2]))


def _cross(x, y):
    return np.array([f32(x[1] * y[2]) - f32(y[1] * x[2]), f32(x[2] * y[0]) - f32(y[2] * x[0]),
                     f32(x[0] * y[1]) - f32(y[0] * x[1])], dtype=f32)


def _normalize(a):
    inv = f32(1.0) / np.sqrt(_dot(a, a), dtype=f32)
    return (a * inv).astype(f32)


def normalize_rows(a):
    """glm::normalize over the rows of an (n, 3) float32 array, in glm's operation order."""
    a = np.ascontiguousarray(a, f32)
    d = ((a[:, 0] * a[:, 0] + a[:, 1] * a[:, 1]).astype(f32) + a[:, 2] * a[:, 2]).astype(f32)
    with np.errstate(divide="ignore", invalid="ignore"):
        return (a * (f32(1.0) / np.sqrt(d, dtype=f32))[:, None]).astype(f32)


def make_triangle(A, B, Cc, material=-1, normals=None, texcoords=None):
    """Triangle::Builder(pointA, pointB, pointC)[.withNormals][.withTexCoords].withMaterialIndex().build()"""
    A, B, Cc = _v(A), _v(B), _v(Cc)
    t = np.zeros((), capi.TRIANGLE)
    t["AC"] = Cc - A
    t["AB"] = B - A
    t["A"] = A
    if normals is None:
        n = _normalize(_cross(t["AC"], t["AB"]))
        normals = (n, n, n)
    # the Triangle ctor normalises whatever the builder holds once more (Triangle.cpp:18-20)
    t["nA"], t["nB"], t["nC"] = _normalize(_v(normals[0])), _normalize(_v(normals[1])), _normalize(_v(normals[2]))
    tc = texcoords if texcoords is not None else ((-1, -1), (-1, -1), (-1, -1))
    t["tA"], t["tB"], t["tC"] = _v(tc[0]), _v(tc[1]), _v(tc[2])
    t["material"] = material
    return t


def make_material(Kd=(0, 0, 0), Ks=(0, 0, 0), Kt=(0, 0, 0), ior=1.0, Le=(0, 0, 0)):
    """Material(Kd, Ks, Kt, refractiveIndice, Le) (app/MobileRT/Material.hpp)"""
    m = np.zeros((), capi.MATERIAL)
    m["Kd"], m["Ks"], m["Kt"], m["ior"], m["Le"] = _v(Kd), _v(Ks), _v(Kt), ior, _v(Le)
    return m


class Scene:
    """MobileRT::Scene: triangles_, spheres_, planes_, lights_, materials_."""

    def __init__(self):
        self.triangles_ = []
        self.spheres_ = []
        self.planes_ = []
        self.lights_ = []
        self.materials_ = []
        self.textures_ = []            # decoded images, uint8 (h, w, channels) as stb_image returns them (Texture.cpp:58-118)
        self.material_texture_ = []    # per material: index into textures_ or -1 (Material::texture_)

    # -- builders used by the built-in scenes and the loaders
    def add_material(self, m):
        self.materials_.append(m)
        return len(self.materials_) - 1

    def add_sphere(self, center, radius, material):
        s = np.zeros((), capi.SPHERE)
        s["center"] = _v(center)
        s["sq_radius"] = f32(radius) * f32(radius)      # Sphere ctor, Sphere.cpp:15-19
        s["material"] = material
        self.spheres_.append(s)

    def add_plane(self, point, normal, material):
        p = np.zeros((), capi.PLANE)
        p["normal"] = _normalize(_v(normal))            # Plane ctor, Plane.cpp:15-19
        p["point"] = _v(point)
        p["material"] = material
        self.planes_.append(p)

    def add_point_light(self, Le, position):
        l = np.zeros((), capi.LIGHT)
        l["kind"] = capi.LIGHT_POINT
        l["position"] = _v(position)
        l["Le"] = _v(Le)
        self.lights_.append(l)

    def add_area_light(self, Le, triangle):
        l = np.zeros((), capi.LIGHT)
        l["kind"] = capi.LIGHT_AREA
        l["triangle"] = triangle
        l["Le"] = _v(Le)
        self.lights_.append(l)

    def arrays(self):
        def pack(lst, dt):
            a = np.zeros(len(lst), dt)
            for i, x in enumerate(lst):
                a[i] = x
            return a
        tri = self.triangles_ if isinstance(self.triangles_, np.ndarray) else pack(self.triangles_, capi.TRIANGLE)
        out = {"triangles": np.ascontiguousarray(tri), "spheres": pack(self.spheres_, capi.SPHERE),
               "planes": pack(self.planes_, capi.PLANE), "lights": pack(self.lights_, capi.LIGHT),
               "materials": pack(self.materials_, capi.MATERIAL) if not isinstance(self.materials_, np.ndarray) else self.materials_}
        if self.textures_:
            mt = list(self.material_texture_) + [-1] * (len(out["materials"]) - len(self.material_texture_))
            out["textures"], out["material_texture"] = list(self.textures_), np.asarray(mt, np.int32)
        return out


class Camera:
    """Arguments of Components::Perspective / Components::Orthographic."""

    def __init__(self, kind, position, look_at, up, a, b):
        self.kind, self.position, self.look_at, self.up, self.a, self.b = kind, _v(position), _v(look_at), _v(up), f32(a), f32(b)

    def desc(self):
        d = capi.CameraDesc()
        d.kind = self.kind
        for k in range(3):
            d.position[k], d.look_at[k], d.up[k] = float(self.position[k]), float(self.look_at[k]), float(self.up[k])
        d.a, d.b = float(self.a), float(self.b)
        return d


# ---- app/Scenes/Scenes.cpp ------------------------------------------------------------------------------------------
_lightMat = dict(Le=(0.9, 0.9, 0.9), ior=1.0)
_mirrorMat = dict(Ks=(0.9, 0.9, 0.9), ior=1.0)
_transmissionMat = dict(Kt=(0.9, 0.9, 0.9), ior=1.9)
_lightGray, _red, _yellow = (0.7, 0.7, 0.7), (0.9, 0.0, 0.0), (0.9, 0.9, 0.0)
_green, _blue, _sand, _lightBlue = (0.0, 0.9, 0.0), (0.0, 0.0, 0.9), (0.914, 0.723, 0.531), (0.0, 0.9, 0.9)
_tri = ((0.5, -0.5, 0.99), (0.5, 0.5, 1.001), (-0.5, -0.5, 0.99))


def _diffuse(scene, kd):
    # Material(Kd) leaves refractiveIndice_ at its default 1.0 (Material.hpp ctor defaults)
    return scene.add_material(make_material(Kd=kd, ior=1.0))


def _cornell_box(scene):
    """cornellBox() (Scenes.cpp:47-94): six axis-aligned planes."""
    for point, normal, kd in (((0, 0, 1), (0, 0, -1), _lightGray), ((0, 0, -3.5), (0, 0, 1), _lightBlue),
                              ((0, -1, 0), (0, 1, 0), _lightGray), ((0, 1, 0), (0, -1, 0), _lightGray),
                              ((-1, 0, 0), (1, 0, 0), _red), ((1, 0, 0), (-1, 0, 0), _blue)):
        idx = len(scene.materials_)
        scene.add_plane(point, normal, idx)
        _diffuse(scene, kd)
    return scene


def cornellBox_Scene(scene=None):
    scene = scene or Scene()
    scene.add_point_light(_lightMat["Le"], (0.0, 0.99, 0.0))
    scene.triangles_.append(make_triangle(*_tri, material=len(scene.materials_)))
    _diffuse(scene, _yellow)
    scene.add_sphere((0.45, -0.65, 0.4), 0.35, len(scene.materials_))
    scene.add_material(make_material(**_mirrorMat))
    scene.add_sphere((-0.45, -0.1, 0.0), 0.35, len(scene.materials_))
    _diffuse(scene, _green)
    return _cornell_box(scene)


def cornellBox_Cam(ratio):
    return Camera(capi.CAMERA_PERSPECTIVE, (0, 0, -3.4), (0, 0, 1), (0, 1, 0), f32(45.0) * f32(ratio), 45.0)


def cornellBox2_Scene(scene=None):
    scene = scene or Scene()
    scene.add_area_light(_lightMat["Le"], make_triangle((-0.25, 0.99, -0.25), (0.25, 0.99, -0.25), (0.25, 0.99, 0.25)))
    scene.add_area_light(_lightMat["Le"], make_triangle((0.25, 0.99, 0.25), (-0.25, 0.99, 0.25), (-0.25, 0.99, -0.25)))
    scene.triangles_.append(make_triangle(*_tri, material=len(scene.materials_)))
    _diffuse(scene, _yellow)
    scene.triangles_.append(make_triangle((-0.5, 0.5, 0.99), (-0.5, -0.5, 0.99), (0.5, 0.5, 0.99), material=len(scene.materials_)))
    _diffuse(scene, _green)
    scene.add_sphere((0.45, -0.65, 0.4), 0.35, len(scene.materials_))
    scene.add_material(make_material(**_mirrorMat))
    scene.add_sphere((-0.4, -0.3, 0.0), 0.35, len(scene.materials_))
    scene.add_material(make_material(**_transmissionMat))
    return _cornell_box(scene)


def spheres_Scene(scene=None):
    scene = scene or Scene()
    scene.add_sphere((4.0, 4.0, 4.0), 4.0, len(scene.materials_))
    _diffuse(scene, _red)
    scene.triangles_.append(make_triangle((0, 10, 10), (0, 0, 10), (10, 0, 10), material=len(scene.materials_)))
    _diffuse(scene, _sand)
    return scene


def spheres_Cam(ratio):
    return Camera(capi.CAMERA_ORTHOGRAPHIC, (0, 1, -10), (0, 1, 7), (0, 1, 0), f32(10.0) * f32(ratio), 10.0)


def spheres2_Scene(scene=None):
    scene = scene or Scene()
    scene.add_point_light(_lightMat["Le"], (0.0, 15.0, 4.0))
    for center, radius, mat in (((-1, 1, 6), 1.0, _red), ((-0.5, 2, 5), 0.3, _blue), ((0, 2, 7), 1.0, None),
                                ((0.5, 0.5, 5), 0.2, _yellow), ((1, 0.5, 4.5), 0.5, _green)):
        scene.add_sphere(center, radius, len(scene.materials_))
        if mat is None:
            scene.add_material(make_material(**_mirrorMat))
        else:
            _diffuse(scene, mat)
    scene.add_plane((0, 0, 0), (0, 1, 0), len(scene.materials_))
    _diffuse(scene, _sand)
    return scene


def spheres2_Cam(ratio):
    return Camera(capi.CAMERA_PERSPECTIVE, (0, 0.5, 1), (0, 0, 7), (0, 1, 0), f32(60.0) * f32(ratio), 60.0)


BUILTIN = {0: (cornellBox_Scene, cornellBox_Cam, (1, 1, 1)), 1: (spheres_Scene, spheres_Cam, (8, 8, 8)),
           2: (cornellBox2_Scene, cornellBox_Cam, (1, 1, 1)), 3: (spheres2_Scene, spheres2_Cam, (8, 8, 8))}
