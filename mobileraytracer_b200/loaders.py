"""Host-side scene ingest: Wavefront OBJ/MTL and the reference's .cam files -> the flat arrays the C-ABI takes.

Mirrors what the reference does around tinyobjloader v1.0.7 (app/Components/Loaders/OBJLoader.cpp:116-184,276-497):
  * x is negated for positions and normals (OBJLoader.cpp:139-141,167-171)
  * faces without normals get the face normal normalize(cross(AC, AB)) (OBJLoader.cpp:173-181)
  * a face whose material has a positive (colour-normalised) emission becomes an AreaLight INSTEAD of a scene
    triangle (OBJLoader.cpp:367-387); Ke is divided by its largest component when that exceeds 1 (Utils.cpp:189-196)
  * transmittance = Tf * (1 - dissolve) (OBJLoader.cpp:330)
  * materials are de-duplicated by value in first-use order (OBJLoader.cpp:406-418)
  * faces without a material take the first vertex's colour as Kd (OBJLoader.cpp:423-455)
and Components::PerspectiveLoader / OrthographicLoader / CameraFactory for `.cam` (PerspectiveLoader.cpp:18-64).
Triangle order is file order here (the reference's order depends on its loader threads, OBJLoader.cpp:83-97).
Textures: a material with map_Kd in a file that has `vt` lines gets the decoded image (PIL here, stb_image in the reference:
both return the PNG's own 8-bit channels) and its faces the fract() of their texture coordinates (OBJLoader.cpp:328-364,
193-212, Utils.cpp:177-180); the texture is part of the material's identity for de-duplication (Material.cpp:36-45)."""
import os

import numpy as np

from . import capi
from .scene import Scene, Camera, normalize_rows

f32 = np.float32


def _parse_mtl(path):
    """tinyobj::LoadMtl subset: name -> dict(Kd, Ks, Kt, Ke, ior, dissolve, map_Kd) with tinyobj's defaults."""
    mats, names, cur = [], {}, None
    has_d = False
    try:
        fh = open(path, "r", errors="replace")
    except OSError:
        return mats, names
    with fh:
        for raw in fh:
            line = raw.strip()
            if not line or line[0] == "#":
                continue
            parts = line.split()
            key, val = parts[0], parts[1:]
            if key == "newmtl":
                cur = {"name": " ".join(val), "Kd": [0, 0, 0], "Ks": [0, 0, 0], "Kt": [0, 0, 0], "Ke": [0, 0, 0], "ior": 1.0,
                       "dissolve": 1.0, "map_Kd": ""}
                names[cur["name"]] = len(mats)   # a repeated name maps to the latest definition, as in tinyobj
                mats.append(cur)
                has_d = False
            elif cur is None:
                continue
            elif key in ("Kd", "Ks", "Ke"):
                cur[key] = [float(x) for x in val[:3]]
            elif key in ("Kt", "Tf"):
                cur["Kt"] = [float(x) for x in val[:3]]
            elif key == "Ni":
                cur["ior"] = float(val[0])
            elif key == "d":
                cur["dissolve"] = float(val[0])
                has_d = True
            elif key == "Tr" and not has_d:
                cur["dissolve"] = 1.0 - float(val[0])
            elif key == "map_Kd":
                cur["map_Kd"] = val[-1] if val else ""
    return mats, names


def _color_normalize(c):
    """MobileRT::normalize(vec3) (Utils.cpp:189-196)."""
    c = np.asarray(c, f32)
    m = max(max(c[0], c[1]), c[2])
    return (c / f32(m)).astype(f32) if m > 1.0 else c


def load_obj(obj_path, mtl_path=None):
    """OBJLoader(obj, mtl) + fillScene -> Scene (triangles as one structured numpy array)."""
    mats, mat_names = _parse_mtl(mtl_path) if mtl_path else ([], {})
    v_txt, vn_txt, vt_txt, f_lines = [], [], [], []
    cur_mat = -1
    with open(obj_path, "r", errors="replace") as fh:
        for line in fh:
            if len(line) < 3:
                continue
            c0, c1 = line[0], line[1]
            if c0 == "v":
                if c1 == " " or c1 == "\t":
                    v_txt.append(line[2:])
                elif c1 == "n":
                    vn_txt.append(line[3:])
                elif c1 == "t":
                    vt_txt.append(line[3:])
            elif c0 == "f" and (c1 == " " or c1 == "\t"):
                f_lines.append((cur_mat, line[2:]))
            elif c0 == "u" and line.startswith("usemtl"):
                cur_mat = mat_names.get(line[6:].strip(), -1)
    nv = len(v_txt)
    vals = np.array(" ".join(v_txt).split(), dtype=np.float64)
    per = vals.size // nv if nv else 3
    vals = vals.reshape(nv, per) if nv else np.zeros((0, 3))
    V = vals[:, :3].astype(f32)
    colors = vals[:, 3:6].astype(f32) if per >= 6 else np.ones((nv, 3), f32)   # tinyobj: white when the file has no colours
    VN = np.array(" ".join(vn_txt).split(), dtype=np.float64).reshape(-1, 3).astype(f32) if vn_txt else np.zeros((0, 3), f32)
    V[:, 0] = -V[:, 0]
    if len(VN):
        VN[:, 0] = -VN[:, 0]

    # faces -> fan-triangulated index triples (tinyobj triangulates polygons on load)
    vi, ni, ti, fm = [], [], [], []
    n_vn = len(VN)
    VT = np.array([x.split()[:2] for x in vt_txt], dtype=np.float64).astype(f32).reshape(-1, 2) if vt_txt else np.zeros((0, 2), f32)
    n_vt = len(VT)
    for m, txt in f_lines:
        toks = txt.split()
        if len(toks) < 3:
            continue
        pv, pn, pt = [], [], []
        for t in toks:
            a = t.split("/")
            i = int(a[0])
            pv.append(i - 1 if i > 0 else nv + i)
            if len(a) > 1 and a[1]:
                j = int(a[1])
                pt.append(j - 1 if j > 0 else n_vt + j)
            else:
                pt.append(-1)
            if len(a) > 2 and a[2]:
                j = int(a[2])
                pn.append(j - 1 if j > 0 else n_vn + j)
            else:
                pn.append(-1)
        for k in range(1, len(pv) - 1):
            vi.append((pv[0], pv[k], pv[k + 1]))
            ni.append((pn[0], pn[k], pn[k + 1]))
            ti.append((pt[0], pt[k], pt[k + 1]))
            fm.append(m)
    vi = np.asarray(vi, np.int64).reshape(-1, 3)
    ni = np.asarray(ni, np.int64).reshape(-1, 3)
    ti = np.asarray(ti, np.int64).reshape(-1, 3)
    fm = np.asarray(fm, np.int64)
    n = len(vi)
    A, B, Cc = V[vi[:, 0]], V[vi[:, 1]], V[vi[:, 2]]
    tri = np.zeros(n, capi.TRIANGLE)
    tri["A"] = A
    tri["AB"] = B - A                     # Triangle ctor: AB = B - A, AC = C - A (Triangle.cpp:328-340)
    tri["AC"] = Cc - A
    if n_vn:
        # the reference indexes attrib.normals with normal_index as soon as the file has ANY normals (OBJLoader.cpp:162-171)
        nA, nB, nC = VN[ni[:, 0]], VN[ni[:, 1]], VN[ni[:, 2]]
    else:
        ac, ab = tri["AC"], tri["AB"]
        cx = np.stack([ac[:, 1] * ab[:, 2] - ab[:, 1] * ac[:, 2], ac[:, 2] * ab[:, 0] - ab[:, 2] * ac[:, 0],
                       ac[:, 0] * ab[:, 1] - ab[:, 0] * ac[:, 1]], axis=1).astype(f32)
        nA = nB = nC = normalize_rows(cx)
    # the Triangle ctor normalises the builder's normals once more (Triangle.cpp:18-20)
    tri["nA"], tri["nB"], tri["nC"] = normalize_rows(nA), normalize_rows(nB), normalize_rows(nC)
    tri["tA"] = tri["tB"] = tri["tC"] = -1.0

    scene = Scene()
    # per-file-material records, then value de-duplication in first-use order
    recs = []
    for m in mats:
        Kd = np.asarray(m["Kd"], f32)
        Ks = np.asarray(m["Ks"], f32)
        Kt = (np.asarray(m["Kt"], f32) * f32(f32(1.0) - f32(m["dissolve"]))).astype(f32)
        Le = _color_normalize(m["Ke"])
        recs.append((Kd, Ks, Kt, f32(m["ior"]), Le))
    dedup, out_mats = {}, []
    tex_cache = {}      # texture file name -> index into scene.textures_ or -1 when it cannot be decoded (OBJLoader.cpp:253-270)

    def texture_index(name):
        if name not in tex_cache:
            tex_cache[name] = -1
            try:
                from PIL import Image
                img = Image.open(os.path.join(os.path.dirname(obj_path), name))
                if img.mode in ("P", "1") or img.mode.startswith("I") or img.mode == "F":
                    img = img.convert("RGBA" if "transparency" in img.info else "RGB")   # stb_image expands palettes / 16 bit to 8-bit channels
                px = np.asarray(img, np.uint8)
                scene.textures_.append(px if px.ndim == 3 else px[:, :, None])
                tex_cache[name] = len(scene.textures_) - 1
            except Exception:
                pass
        return tex_cache[name]

    def mat_index(Kd, Ks, Kt, ior, Le, tex=-1):
        key = (Kd.tobytes(), Ks.tobytes(), Kt.tobytes(), np.float32(ior).tobytes(), Le.tobytes(), tex)
        if key not in dedup:
            dedup[key] = len(out_mats)
            r = np.zeros((), capi.MATERIAL)
            r["Kd"], r["Ks"], r["Kt"], r["ior"], r["Le"] = Kd, Ks, Kt, ior, Le
            out_mats.append(r)
            scene.material_texture_.append(tex)
        return dedup[key]

    is_light = np.zeros(n, bool)
    mat_of_file = {}
    order_first_use = []
    # first-use order over faces (vectorised: first occurrence of every file material, faces without material by colour)
    if n:
        uniq, first = np.unique(fm, return_index=True)
        for m, fi in sorted(zip(uniq.tolist(), first.tolist()), key=lambda x: x[1]):
            order_first_use.append(m)
    tri_mat = np.full(n, -1, np.int32)
    for m in order_first_use:
        sel = fm == m
        if m >= 0:
            Kd, Ks, Kt, ior, Le = recs[m]
            if (Le > 0).any():
                is_light |= sel
                mat_of_file[m] = None
            else:
                tex = -1
                if mats[m]["map_Kd"] and n_vt:            # hasTexture && hasCoordTex (OBJLoader.cpp:332-336)
                    tex = texture_index(mats[m]["map_Kd"])
                    if tex >= 0:
                        for name, col in (("tA", 0), ("tB", 1), ("tC", 2)):
                            tc = VT[ti[sel, col]]
                            tri[name][sel] = (tc - np.floor(tc)).astype(f32)     # glm::fract
                tri_mat[sel] = mat_index(Kd, Ks, Kt, ior, Le, tex)
        else:
            cols = colors[vi[sel, 0]]
            idxs = np.nonzero(sel)[0]
            z = np.zeros(3, f32)
            for i, c in zip(idxs, cols):
                tri_mat[i] = mat_index(c.astype(f32), z, z, f32(1.0), z)
    tri["material"] = tri_mat
    for i in np.nonzero(is_light)[0]:
        t = tri[i].copy()
        t["material"] = -1
        scene.add_area_light(recs[fm[i]][4], t)
    scene.triangles_ = np.ascontiguousarray(tri[~is_light])
    scene.materials_ = out_mats
    return scene


def load_camera(cam_path, aspect_ratio):
    """CameraFactory::loadFromFile (CameraFactory.cpp:6-22) -> PerspectiveLoader / OrthographicLoader."""
    kind, pos, look, up, fov = "perspective", (0, 0, 0), (0, 0, 0), (0, 0, 0), (0, 0)
    with open(cam_path, "r") as fh:
        for line in fh:
            if not line.strip():
                continue
            key, rest = line[0], line[1:].split("#")[0].split()
            if key == "t":
                kind = rest[0]
            elif key == "p":
                pos = [float(x) for x in rest[:3]]
            elif key == "l":
                look = [float(x) for x in rest[:3]]
            elif key == "u":
                up = [float(x) for x in rest[:3]]
            elif key == "f":
                fov = [float(x) for x in rest[:2]]
    pos = [-pos[0], pos[1], pos[2]]   # "Invert X axis" (PerspectiveLoader.cpp:51)
    if kind.startswith("o"):
        return Camera(capi.CAMERA_ORTHOGRAPHIC, pos, look, up, f32(fov[0]) * f32(aspect_ratio), fov[1])
    return Camera(capi.CAMERA_PERSPECTIVE, pos, look, up, f32(fov[0]) * f32(aspect_ratio), fov[1])


def replicate_scene(arrays, nx=6, nz=5, pitch_x=4200.0, pitch_z=5200.0):
    """BASELINE config 5's synthetic scene: the scene's triangles and lights copied nx x nz times on a grid (flattened, the
    reference has no instancing); the copy at grid position (nx // 3, nz // 2) stays where it is, under the camera."""
    tri, lights = arrays["triangles"], arrays["lights"]
    tris, ls = [], []
    for ix in range(nx):
        for iz in range(nz):
            off = np.array([(ix - nx // 3) * pitch_x, 0.0, (iz - nz // 2) * pitch_z], np.float32)
            t = tri.copy()
            t["A"] += off
            tris.append(t)
            lg = lights.copy()
            lg["triangle"]["A"] += off
            lg["position"] += off
            ls.append(lg)
    out = dict(arrays)
    out["triangles"] = np.ascontiguousarray(np.concatenate(tris))
    out["lights"] = np.ascontiguousarray(np.concatenate(ls))
    return out
