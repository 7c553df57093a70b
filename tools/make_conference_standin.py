"""Procedural stand-in for WavefrontOBJs/conference/conference.obj, which is missing from the reference snapshot
(/root/reference/.MISSING_LARGE_BLOBS:1). Emits a Wavefront OBJ + MTL + .cam triple that both the
reference's own loader (tinyobjloader) and mobileraytracer_b200.loaders read:

  a closed room with panelled walls, a long table, rows of chairs, table-top objects and 4 emissive ceiling panels
  (material `light`, Ke 0.9 like conference.mtl:360-361), 328 120 triangles at the default detail, conference.mtl's own
  materials (Kd by file position, the three specular ones included),
  camera = WavefrontOBJs/conference/conference.cam (p -460 500 -1000, l 0 400 0, f 45 45).

Everything is generated from a fixed seed; the same arguments always give the same bytes. Numbers measured on this scene
are labelled "conference stand-in" everywhere. If a real conference.obj is supplied it drops in unchanged."""
import argparse
import os

import numpy as np

N_MATERIALS = 36
DEFAULT_DETAIL = 1.095   # gives 328 120 faces (the original conference.obj has about 331 k)


class Mesh:
    def __init__(self):
        self.v, self.f, self.m = [], [], []
        self.nv = 0

    def add(self, verts, faces, mat):
        verts = np.asarray(verts, np.float64).reshape(-1, 3)
        faces = np.asarray(faces, np.int64).reshape(-1, 3)
        self.v.append(verts)
        self.f.append(faces + self.nv)
        self.m.append(np.full(len(faces), mat, np.int64))
        self.nv += len(verts)

    def arrays(self):
        v, f, m = np.concatenate(self.v), np.concatenate(self.f), np.concatenate(self.m)
        # drop zero-area faces (poles of the lathes and of the polar table top): they can never be hit and have no normal
        a, b, c = v[f[:, 0]], v[f[:, 1]], v[f[:, 2]]
        keep = np.linalg.norm(np.cross(b - a, c - a), axis=1) > 1e-9
        return v, f[keep], m[keep]


def grid_faces(nu, nv, flip=False):
    """Two triangles per cell of an (nu+1) x (nv+1) vertex grid."""
    i, j = np.meshgrid(np.arange(nu), np.arange(nv), indexing="ij")
    a = (i * (nv + 1) + j).ravel()
    b, c, d = a + (nv + 1), a + (nv + 1) + 1, a + 1
    f = np.concatenate([np.stack([a, b, c], 1), np.stack([a, c, d], 1)])
    return f[:, ::-1] if flip else f


def surface(fn, nu, nv, flip=False):
    u, v = np.meshgrid(np.linspace(0, 1, nu + 1), np.linspace(0, 1, nv + 1), indexing="ij")
    p = fn(u.ravel(), v.ravel())
    return np.stack(p, 1), grid_faces(nu, nv, flip)


def add_box(mesh, lo, hi, mat, n=(1, 1, 1)):
    lo, hi = np.asarray(lo, float), np.asarray(hi, float)
    for axis in range(3):
        a1, a2 = (axis + 1) % 3, (axis + 2) % 3
        for side in (0, 1):
            def fn(u, v, axis=axis, a1=a1, a2=a2, side=side):
                p = [None, None, None]
                p[axis] = np.full_like(u, hi[axis] if side else lo[axis])
                p[a1] = lo[a1] + u * (hi[a1] - lo[a1])
                p[a2] = lo[a2] + v * (hi[a2] - lo[a2])
                return p
            vs, fs = surface(fn, n[a1], n[a2], flip=not side)
            mesh.add(vs, fs, mat)


def add_cylinder(mesh, base, radius, height, mat, seg=12, rings=4, axis=1):
    base = np.asarray(base, float)

    def fn(u, v):
        ang = 2 * np.pi * u
        p = [None, None, None]
        p[axis] = base[axis] + v * height
        p[(axis + 1) % 3] = base[(axis + 1) % 3] + radius * np.cos(ang)
        p[(axis + 2) % 3] = base[(axis + 2) % 3] + radius * np.sin(ang)
        return p
    vs, fs = surface(fn, seg, rings)
    mesh.add(vs, fs, mat)


def add_lathe(mesh, center, profile_r, profile_y, mat, seg=24):
    """Surface of revolution around +y through `center`; profile arrays give radius and height along the profile."""
    center = np.asarray(center, float)
    pr, py = np.asarray(profile_r, float), np.asarray(profile_y, float)
    rings = len(pr) - 1

    def fn(u, v):
        k = np.clip(np.rint(v * rings).astype(int), 0, rings)
        ang = 2 * np.pi * u
        return [center[0] + pr[k] * np.cos(ang), center[1] + py[k], center[2] + pr[k] * np.sin(ang)]
    vs, fs = surface(fn, seg, rings)
    mesh.add(vs, fs, mat)


def add_chair(mesh, pos, yaw, mats, d):
    """Seat + curved back + 4 legs + 2 arm rests, rotated by yaw around y, feet on the floor at pos."""
    sub = Mesh()
    s = max(2, int(round(10 * d)))
    add_box(sub, (-28, 42, -28), (28, 50, 28), mats[0], n=(s, 2, s))
    bn = max(4, int(round(26 * d)))

    def back(u, v):
        x = -30 + 60 * u
        return [x, 52 + 62 * v, -30 - 9 * np.cos((u - 0.5) * np.pi) * (0.3 + 0.7 * v) + 2 * np.sin(6 * np.pi * v)]
    for off, flip in ((0.0, False), (-4.0, True)):
        vs, fs = surface(back, bn, bn, flip)
        vs[:, 2] += off
        sub.add(vs, fs, mats[1])
    seg = max(6, int(round(14 * d)))
    for lx in (-24, 24):
        for lz in (-24, 24):
            add_cylinder(sub, (lx, 0, lz), 2.2, 42, mats[2], seg=seg, rings=max(2, int(round(6 * d))))
    for ax in (-30, 30):
        add_cylinder(sub, (ax, 68, -26), 2.5, 50, mats[2], seg=seg, rings=max(2, int(round(8 * d))), axis=2)
        add_cylinder(sub, (ax, 50, 20), 1.8, 18, mats[2], seg=seg, rings=2)
    v, f, m = sub.arrays()
    c, sn = np.cos(yaw), np.sin(yaw)
    x, z = v[:, 0] * c + v[:, 2] * sn, -v[:, 0] * sn + v[:, 2] * c
    v = np.stack([x + pos[0], v[:, 1] + pos[1], z + pos[2]], 1)
    for mat in np.unique(m):
        sel = m == mat
        # re-index the selected faces into a compact vertex list
        used = np.unique(f[sel])
        remap = np.full(len(v), -1, np.int64)
        remap[used] = np.arange(len(used))
        mesh.add(v[used], remap[f[sel]], int(mat))


def generate(detail=DEFAULT_DETAIL, seed=7):
    """Returns (vertices float64 (n,3) in WORLD coordinates, faces int64 (m,3), face material ids; id N_MATERIALS = light)."""
    rng = np.random.default_rng(seed)
    d = float(detail)
    mesh = Mesh()
    LIGHT = N_MATERIALS
    X, Y, Z = 900.0, 800.0, 1400.0
    wn = max(4, int(round(56 * d)))

    # floor, ceiling and four walls: relief panels so that the big surfaces are made of many triangles like the original's
    def floor(u, v):
        return [-X + 2 * X * u, 0 * u, -Z + 2 * Z * v]
    vs, fs = surface(floor, wn, int(wn * 1.5), flip=True)
    mesh.add(vs, fs, 0)

    def ceil(u, v):
        return [-X + 2 * X * u, Y + 6 * np.sin(8 * np.pi * u) * np.sin(12 * np.pi * v), -Z + 2 * Z * v]
    vs, fs = surface(ceil, wn, int(wn * 1.5))
    mesh.add(vs, fs, 1)
    for sx in (-1, 1):
        def wall(u, v, sx=sx):
            return [sx * (X - 5 * np.abs(np.sin(10 * np.pi * u))), Y * v, -Z + 2 * Z * u]
        vs, fs = surface(wall, int(wn * 1.5), wn, flip=sx < 0)
        mesh.add(vs, fs, 2 if sx < 0 else 3)
    for sz in (-1, 1):
        def wall(u, v, sz=sz):
            return [-X + 2 * X * u, Y * v, sz * (Z - 5 * np.abs(np.sin(7 * np.pi * u)))]
        vs, fs = surface(wall, wn, wn, flip=sz > 0)
        mesh.add(vs, fs, 4 if sz < 0 else 5)

    # ceiling light panels (emissive): 4 quads = 8 area lights
    for lx in (-300.0, 300.0):
        for lz in (-500.0, 500.0):
            vsq = np.array([[lx - 90, Y - 12, lz - 160], [lx + 90, Y - 12, lz - 160], [lx + 90, Y - 12, lz + 160], [lx - 90, Y - 12, lz + 160]])
            mesh.add(vsq, np.array([[0, 1, 2], [0, 2, 3]]), LIGHT)

    # the table: rounded slab + thick edge + legs
    tn = max(4, int(round(70 * d)))

    def top(u, v, y=300.0):
        ang = 2 * np.pi * u
        r = v
        return [330 * r * np.cos(ang) * (1 + 0.15 * np.cos(2 * ang)), y + 0 * u, 820 * r * np.sin(ang)]
    vs, fs = surface(top, tn * 2, tn // 2)
    mesh.add(vs, fs, 6)
    vs, fs = surface(lambda u, v: top(u, v, 284.0), tn * 2, tn // 2, flip=True)
    mesh.add(vs, fs, 6)

    def rim(u, v):
        ang = 2 * np.pi * u
        return [330 * np.cos(ang) * (1 + 0.15 * np.cos(2 * ang)), 284 + 16 * v, 820 * np.sin(ang)]
    vs, fs = surface(rim, tn * 2, 3, flip=True)
    mesh.add(vs, fs, 7)
    for lx in (-180, 180):
        for lz in (-560, 0, 560):
            add_cylinder(mesh, (lx, 0, lz), 14, 284, 8, seg=max(8, int(round(20 * d))), rings=max(2, int(round(10 * d))))

    # chairs around the table and along the walls
    chairs = []
    for k in range(9):
        z = -720 + 180 * k
        chairs.append(((-430, 0, z), np.pi / 2))
        chairs.append(((430, 0, z), -np.pi / 2))
    chairs.append(((0, 0, -960), 0.0))
    chairs.append(((0, 0, 960), np.pi))
    for k in range(8):
        z = -1050 + 300 * k
        chairs.append(((-800, 0, z), np.pi / 2))
        chairs.append(((800, 0, z), -np.pi / 2))
    for i, (pos, yaw) in enumerate(chairs):
        jitter = rng.uniform(-0.25, 0.25)
        shift = rng.uniform(-12, 12, 3) * np.array([1, 0, 1])
        mats = (9 + i % 5, 14 + i % 7, 21 + i % 3)
        add_chair(mesh, np.asarray(pos) + shift, yaw + jitter, mats, d)

    # objects on the table: cups, bottles, stacks of paper, laptops
    seg = max(8, int(round(28 * d)))
    for i in range(22):
        cx, cz = rng.uniform(-200, 200), rng.uniform(-700, 700)
        if i % 3 == 0:
            r = np.array([0.0, 9, 10, 10.5, 11, 10, 9.5, 9, 0.0])
            y = np.array([0.0, 0, 4, 12, 22, 22, 12, 3, 3.0])
        elif i % 3 == 1:
            r = np.array([0.0, 8, 8.5, 8.5, 7, 3.5, 3.5, 4, 0.0])
            y = np.array([0.0, 0, 6, 40, 52, 58, 70, 72, 72.0])
        else:
            r = np.array([0.0, 16, 17, 15, 9, 4, 0.0])
            y = np.array([0.0, 0, 8, 20, 28, 31, 32.0])
        k = max(1, int(round(4 * d)))
        rr = np.interp(np.linspace(0, len(r) - 1, (len(r) - 1) * k + 1), np.arange(len(r)), r)
        yy = np.interp(np.linspace(0, len(y) - 1, (len(y) - 1) * k + 1), np.arange(len(y)), y)
        add_lathe(mesh, (cx, 300.0, cz), rr, yy, 24 + i % 6, seg=seg)
    for i in range(12):
        cx, cz = rng.uniform(-220, 220), rng.uniform(-720, 720)
        w, h = rng.uniform(18, 30), rng.uniform(25, 40)
        add_box(mesh, (cx - w, 300, cz - h), (cx + w, 300 + rng.uniform(1.5, 6), cz + h), 30 + i % 3, n=(3, 1, 3))
    for i in range(8):
        cx, cz = (-1) ** i * 200.0, -630 + 180 * i
        add_box(mesh, (cx - 32, 300, cz - 24), (cx + 32, 303, cz + 24), 33, n=(6, 1, 6))

        def lid(u, v, cx=cx, cz=cz, s=(-1) ** i):
            return [cx + s * (24 + 14 * v), 303 + 44 * v, cz - 24 + 48 * u]
        vs, fs = surface(lid, max(2, int(8 * d)), max(2, int(8 * d)))
        mesh.add(vs, fs, 34)
        mesh.add(vs + np.array([(-1) ** i * 1.5, 0, 0]), fs[:, ::-1], 33)

    # a projector screen, door frame and wall cabinets
    add_box(mesh, (-420, 260, Z - 40), (420, 700, Z - 30), 35, n=(max(2, int(30 * d)), max(2, int(20 * d)), 1))
    add_box(mesh, (X - 60, 0, -300), (X - 10, 520, 300), 20, n=(2, max(2, int(24 * d)), max(2, int(24 * d))))
    add_box(mesh, (-X + 10, 0, 700), (-X + 70, 420, 1100), 20, n=(2, max(2, int(20 * d)), max(2, int(20 * d))))
    return mesh.arrays()


# Kd of the 35 surface materials of the reference's WavefrontOBJs/conference/conference.mtl, in file order (mesh0_SG ...
# mesh27_SG); the three with a specular term (Ks 0.1 0.1 0.1: mesh11_SG "green boards", mesh12_SG, mesh13_SG, conference.mtl:104-138)
# are file positions 10, 11, 12. A constant table of the reference's data file: the mesh is a stand-in, the materials are not.
CONFERENCE_KD = [
    (0.639216, 0.6, 0.607843), (0.639216, 0.4, 0.129412), (0.737255, 0.4, 0.156863), (0.176471, 0.058824, 0.058824),
    (0.898039, 0.898039, 0.898039), (0.066667, 0.066667, 0.066667), (0.698039, 0.698039, 0.698039), (0.686275, 0.427451, 0.168627),
    (0.709804, 0.709804, 0.709804), (0.6, 0.6, 0.6), (0.117647, 0.258824, 0.14902), (0.427451, 0.427451, 0.427451),
    (0.517647, 0.447059, 0.337255), (0.447059, 0.447059, 0.447059), (1.0, 1.0, 1.0), (0.498039, 0.498039, 0.498039),
    (0.4, 0.4, 0.4), (0.878431, 0.717647, 0.576471), (0.827451, 0.698039, 0.427451), (0.286275, 0.14902, 0.129412),
    (0.078431, 0.078431, 0.078431), (0.819608, 0.086275, 0.066667), (0.698039, 0.698039, 0.698039), (0.6, 0.6, 0.6),
    (0.639216, 0.439216, 0.317647), (1.0, 1.0, 1.0), (0.058824, 0.058824, 0.058824), (0.976471, 0.176471, 0.168627),
    (0.6, 0.6, 0.6), (0.478431, 0.239216, 0.117647), (0.047059, 0.047059, 0.047059), (0.6, 0.6, 0.6),
    (0.098039, 0.098039, 0.098039), (0.047059, 0.047059, 0.047059), (0.466667, 0.447059, 0.427451)]
CONFERENCE_KS = {10: (0.1, 0.1, 0.1), 11: (0.1, 0.1, 0.1), 12: (0.1, 0.1, 0.1)}


def material_table(seed=7):
    """(Kd, Ks) of the stand-in's 36 material ids: conference.mtl's own values by file position (id 35 wraps to 0)."""
    kd = np.array([CONFERENCE_KD[i % len(CONFERENCE_KD)] for i in range(N_MATERIALS)])
    ks = np.array([CONFERENCE_KS.get(i % len(CONFERENCE_KD), (0.0, 0.0, 0.0)) for i in range(N_MATERIALS)])
    return kd, ks


def write(out_dir, detail=DEFAULT_DETAIL, seed=7, name="conference_standin"):
    """Writes <name>.obj/.mtl/.cam into out_dir; returns their paths and the triangle count."""
    os.makedirs(out_dir, exist_ok=True)
    v, f, m = generate(detail, seed)
    obj, mtl, cam = (os.path.join(out_dir, name + ext) for ext in (".obj", ".mtl", ".cam"))
    kd, ks = material_table(seed)
    with open(mtl, "w") as fh:
        for i in range(N_MATERIALS):
            spec = "  Ks %.6f %.6f %.6f\n" % tuple(ks[i]) if ks[i].any() else ""
            fh.write("newmtl mesh%d_SG\n  Ns 40.0\n  Ni 1.5\n  d 1.0\n  Tf 1.0 1.0 1.0\n  illum 2\n  Ka 0.0 0.0 0.0\n  Kd %.6f %.6f %.6f\n%s  Ke 0.0 0.0 0.0\n\n" % (i, *kd[i], spec))
        fh.write("newmtl light\n  Ke 0.9 0.9 0.9\n")
    with open(obj, "w") as fh:
        fh.write("# procedural stand-in for the missing conference.obj (see tools/make_conference_standin.py)\nmtllib %s.mtl\n" % name)
        vf = v.copy()
        vf[:, 0] = -vf[:, 0]      # the loaders negate x (OBJLoader.cpp:139-141): store file coordinates
        fh.write("".join("v %.4f %.4f %.4f\n" % tuple(r) for r in vf))
        order = np.argsort(m, kind="stable")
        last = -1
        chunks = []
        # negating x mirrors the mesh: swap two indices so that face normals computed by the loaders keep pointing inside
        for idx in order:
            if m[idx] != last:
                last = m[idx]
                chunks.append("usemtl %s\n" % ("light" if last == N_MATERIALS else "mesh%d_SG" % last))
            a, b, c = f[idx] + 1
            chunks.append("f %d %d %d\n" % (a, c, b))
        fh.write("".join(chunks))
    with open(cam, "w") as fh:
        fh.write("t perspective\np -460.0 500.0 -1000.0\nl 0.0 400.0 0.0\nu 0.0 1.0 0.0\nf 45 45\n")
    return {"obj": obj, "mtl": mtl, "cam": cam, "triangles": int(len(f))}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/scenes")
    ap.add_argument("--detail", type=float, default=DEFAULT_DETAIL)
    ap.add_argument("--seed", type=int, default=7)
    a = ap.parse_args()
    print(write(a.out, a.detail, a.seed))
