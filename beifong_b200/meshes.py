"""Mesh ingestion and transforms for the host side of the radar path.

- `load_ply` / `load_obj`: the vertex / face buffers the reference's loaders produce
  (src/shapes/ply.cpp:195-322, src/shapes/obj.cpp): positions transformed by `to_world` in fp32,
  faces as u32 triples, primitive numbering == file order.
- `vertex_normals`: Mesh::recompute_vertex_normals (src/librender/mesh.cpp:201-278), the
  angle-weighted scheme of Thuermer & Wuethrich the loaders run when the file has no normals and
  `face_normals` is false.
- `Transform`: the subset of ScalarTransform4f used by the radar scenes
  (include/mitsuba/core/transform.h: translate, scale, rotate, look_at, perspective).
- procedural stand-ins for the meshes missing from the reference tree (.MISSING_LARGE_BLOBS).
"""
import numpy as np


# ---- transforms (row-major 4x4, column vectors: p' = M p) ----------------------------------------------
class Transform:
    def __init__(self, m=None):
        self.m = np.eye(4, dtype=np.float64) if m is None else np.asarray(m, dtype=np.float64).reshape(4, 4)

    def __mul__(self, other):
        return Transform(self.m @ other.m)

    def matrix32(self):
        return self.m.astype(np.float32)

    def inverse(self):
        return Transform(np.linalg.inv(self.m))

    @staticmethod
    def translate(v):
        t = np.eye(4)
        t[:3, 3] = v
        return Transform(t)

    @staticmethod
    def scale(v):
        v = np.broadcast_to(np.asarray(v, dtype=np.float64), (3,))
        return Transform(np.diag([v[0], v[1], v[2], 1.0]))

    @staticmethod
    def rotate(axis, angle_deg):
        a = np.asarray(axis, dtype=np.float64)
        a = a / np.linalg.norm(a)
        th = np.deg2rad(angle_deg)
        s, c = np.sin(th), np.cos(th)
        x, y, z = a
        r = np.array([[c + x * x * (1 - c), x * y * (1 - c) - z * s, x * z * (1 - c) + y * s, 0],
                      [y * x * (1 - c) + z * s, c + y * y * (1 - c), y * z * (1 - c) - x * s, 0],
                      [z * x * (1 - c) - y * s, z * y * (1 - c) + x * s, c + z * z * (1 - c), 0],
                      [0, 0, 0, 1]])
        return Transform(r)

    @staticmethod
    def look_at(origin, target, up):
        o, t, u = (np.asarray(v, dtype=np.float64) for v in (origin, target, up))
        d = (t - o) / np.linalg.norm(t - o)
        left = np.cross(u, d)
        left /= np.linalg.norm(left)
        new_up = np.cross(d, left)
        m = np.eye(4)
        m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = left, new_up, d, o
        return Transform(m)

    @staticmethod
    def perspective(fov_deg, near, far):
        recip = 1.0 / (far - near)
        cot = 1.0 / np.tan(np.deg2rad(fov_deg * 0.5))
        m = np.zeros((4, 4))
        m[0, 0], m[1, 1], m[2, 2], m[2, 3], m[3, 2] = cot, cot, far * recip, -near * far * recip, 1.0
        return Transform(m)


def perspective_sample_to_camera(fov_deg, near, far, film_w=1, film_h=1):
    """inverse of perspective_projection (include/mitsuba/render/sensor.h:196-231), full-film crop."""
    aspect = film_w / film_h
    cam_to_sample = (Transform.scale([-0.5, -0.5 * aspect, 1.0]) * Transform.translate([-1.0, -1.0 / aspect, 0.0]) *
                     Transform.perspective(fov_deg, near, far))
    return cam_to_sample.inverse().matrix32()


def transform_points(m, pts):
    """Transform4f::transform_affine(Point) in fp32 with the reference's FMA order."""
    m = np.asarray(m, dtype=np.float32)
    p = np.asarray(pts, dtype=np.float32).reshape(-1, 3)
    p64 = p.astype(np.float64)
    out = np.empty_like(p)
    for j in range(3):
        # fmadd(m[j][2], z, fmadd(m[j][1], y, fmadd(m[j][0], x, m[j][3]))): each fma rounds once
        acc = np.float32(1) * (np.float64(m[j, 0]) * p64[:, 0] + np.float64(m[j, 3])).astype(np.float32)
        acc = (np.float64(m[j, 1]) * p64[:, 1] + acc.astype(np.float64)).astype(np.float32)
        acc = (np.float64(m[j, 2]) * p64[:, 2] + acc.astype(np.float64)).astype(np.float32)
        out[:, j] = acc
    return out


# ---- loaders ----------------------------------------------------------------------------------------------
_PLY_TYPES = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2", "ushort": "u2",
              "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4", "float": "f4", "float32": "f4",
              "double": "f8", "float64": "f8"}


def load_ply(path, to_world=None):
    """Returns (positions float32[V,3], faces uint32[F,3], normals float32[V,3] or None)."""
    with open(path, "rb") as f:
        data = f.read()
    end = data.index(b"end_header")
    end = data.index(b"\n", end) + 1
    header = data[:end].decode("ascii", "replace").splitlines()
    if not header or header[0].strip() != "ply":
        raise RuntimeError(f"{path}: invalid PLY header")
    fmt, elements = None, []
    for line in header[1:]:
        tok = line.split()
        if not tok:
            continue
        if tok[0] == "format":
            fmt = tok[1]
        elif tok[0] == "element":
            elements.append({"name": tok[1], "count": int(tok[2]), "props": []})
        elif tok[0] == "property":
            if tok[1] == "list":
                elements[-1]["props"].append(("list", tok[2], tok[3], tok[4]))
            else:
                elements[-1]["props"].append(("scalar", tok[1], tok[2]))
    if fmt not in ("binary_little_endian", "binary_big_endian", "ascii"):
        raise RuntimeError(f"{path}: unsupported PLY format {fmt}")
    pos = nrm = faces = None
    body = data[end:]
    if fmt == "ascii":
        tokens = body.split()
        cur = 0
    else:
        bo = "<" if fmt == "binary_little_endian" else ">"
        off = 0
    for el in elements:
        n = el["count"]
        if el["name"] == "vertex":
            names = [p[2] for p in el["props"]]
            if fmt == "ascii":
                arr = np.array(tokens[cur:cur + n * len(names)], dtype=np.float64).reshape(n, len(names))
                cur += n * len(names)
                cols = {nm: arr[:, i] for i, nm in enumerate(names)}
            else:
                dt = np.dtype([(p[2], bo + _PLY_TYPES[p[1]]) for p in el["props"]])
                arr = np.frombuffer(body, dtype=dt, count=n, offset=off)
                off += n * dt.itemsize
                cols = {nm: arr[nm] for nm in names}
            pos = np.stack([cols["x"], cols["y"], cols["z"]], axis=1).astype(np.float32)
            if all(k in cols for k in ("nx", "ny", "nz")):
                nrm = np.stack([cols["nx"], cols["ny"], cols["nz"]], axis=1).astype(np.float32)
        elif el["name"] == "face":
            tris = []
            lp = [p for p in el["props"] if p[0] == "list"]
            if len(el["props"]) != 1 or not lp:
                raise RuntimeError(f"{path}: face element with extra properties is not supported")
            if fmt == "ascii":
                for _ in range(n):
                    k = int(tokens[cur])
                    idx = [int(t) for t in tokens[cur + 1:cur + 1 + k]]
                    cur += 1 + k
                    tris.extend([idx[0], idx[i], idx[i + 1]] for i in range(1, k - 1))   # ply.cpp fan for quads
                faces = np.array(tris, dtype=np.uint32).reshape(-1, 3)
            else:
                ct, it = np.dtype(bo + _PLY_TYPES[lp[0][1]]), np.dtype(bo + _PLY_TYPES[lp[0][2]])
                # fast path: all triangles
                rec = np.dtype([("n", ct), ("i", it, (3,))])
                if off + n * rec.itemsize <= len(body):
                    arr = np.frombuffer(body, dtype=rec, count=n, offset=off)
                    if n == 0 or np.all(arr["n"] == 3):
                        faces = arr["i"].astype(np.uint32)
                        off += n * rec.itemsize
                        continue
                for _ in range(n):
                    k = int(np.frombuffer(body, dtype=ct, count=1, offset=off)[0])
                    off += ct.itemsize
                    idx = np.frombuffer(body, dtype=it, count=k, offset=off).astype(np.int64)
                    off += k * it.itemsize
                    tris.extend([idx[0], idx[i], idx[i + 1]] for i in range(1, k - 1))
                faces = np.array(tris, dtype=np.uint32).reshape(-1, 3)
        else:
            if fmt == "ascii":
                cur += n * len(el["props"])
            else:
                if any(p[0] == "list" for p in el["props"]):
                    raise RuntimeError(f"{path}: cannot skip list element {el['name']}")
                off += n * sum(np.dtype(_PLY_TYPES[p[1]]).itemsize for p in el["props"])
    if pos is None or faces is None:
        raise RuntimeError(f"{path}: PLY without vertex / face elements")
    if to_world is not None:
        m = np.asarray(to_world, dtype=np.float32)
        pos = transform_points(m, pos)
        if nrm is not None:
            inv_t = np.linalg.inv(m.astype(np.float64)).T[:3, :3]
            nrm = (nrm.astype(np.float64) @ inv_t.T)
            nrm = (nrm / np.linalg.norm(nrm, axis=1, keepdims=True)).astype(np.float32)
    return pos, faces, nrm


def load_obj(path, to_world=None):
    """Positions / triangle faces of a Wavefront OBJ (src/shapes/obj.cpp); no vertex de-duplication by
    normal / uv index (faces index positions only)."""
    pos, faces = [], []
    with open(path, "r") as f:
        for line in f:
            tok = line.split()
            if not tok:
                continue
            if tok[0] == "v":
                pos.append([float(tok[1]), float(tok[2]), float(tok[3])])
            elif tok[0] == "f":
                idx = []
                for t in tok[1:]:
                    i = int(t.split("/")[0])
                    idx.append(i - 1 if i > 0 else len(pos) + i)
                for i in range(1, len(idx) - 1):
                    faces.append([idx[0], idx[i], idx[i + 1]])
    p = np.array(pos, dtype=np.float32).reshape(-1, 3)
    if to_world is not None:
        p = transform_points(np.asarray(to_world, dtype=np.float32), p)
    return p, np.array(faces, dtype=np.uint32).reshape(-1, 3), None


def load_serialized(path, to_world=None, shape_index=0):
    """Mitsuba `.serialized` triangle mesh (src/shapes/serialized.cpp:196-300): little-endian header (0x041C, version
    3 or 4), one zlib stream per sub-shape (flags, [v4: name], vertex / face counts as u64, positions, optional normals /
    texcoords / colours in single or double precision, u32 face indices), sub-shape offsets in a dictionary at the
    end of the file (u64 offsets in v4, u32 in v3).  Returns (positions, faces, normals or None)."""
    import struct
    import zlib
    with open(path, "rb") as f:
        data = f.read()
    if len(data) < 4:
        raise RuntimeError(f"{path}: encountered an invalid file format!")
    fmt, version = struct.unpack_from("<HH", data, 0)
    if fmt != 0x041C:
        raise RuntimeError(f"{path}: encountered an invalid file format!")
    if version not in (3, 4):
        raise RuntimeError(f"{path}: encountered an incompatible file version!")
    start = 0
    if shape_index != 0:
        if shape_index < 0:
            raise RuntimeError("shape index must be nonnegative!")
        count = struct.unpack_from("<I", data, len(data) - 4)[0]
        if shape_index >= count:
            raise RuntimeError(f"Unable to unserialize mesh, shape index is out of range! (requested {shape_index} out of 0..{count - 1})")
        if version == 4:
            start = struct.unpack_from("<Q", data, len(data) - 8 * (count - shape_index) - 4)[0]
        else:
            start = struct.unpack_from("<I", data, len(data) - 4 * (count - shape_index + 1))[0]
    body = zlib.decompressobj().decompress(data[start + 4:])
    off = 0
    flags = struct.unpack_from("<I", body, off)[0]
    off += 4
    if version == 4:
        off = body.index(b"\0", off) + 1                     # utf-8 name, null-terminated
    n_vertices, n_faces = struct.unpack_from("<QQ", body, off)
    off += 16
    ft = np.dtype("<f8") if flags & 0x2000 else np.dtype("<f4")

    def take(n_comp):
        nonlocal off
        a = np.frombuffer(body, dtype=ft, count=n_vertices * n_comp, offset=off).reshape(n_vertices, n_comp)
        off += a.nbytes
        return a.astype(np.float32)
    pos = take(3)
    nrm = take(3) if flags & 0x0001 else None
    if flags & 0x0002:
        take(2)
    if flags & 0x0008:
        take(3)
    faces = np.frombuffer(body, dtype="<u4", count=n_faces * 3, offset=off).reshape(n_faces, 3).astype(np.uint32)
    if faces.size and int(faces.max()) >= n_vertices:
        raise RuntimeError(f"{path}: face index out of range")
    if to_world is not None:
        m = np.asarray(to_world, dtype=np.float32)
        pos = transform_points(m, pos)
        if nrm is not None:
            inv_t = np.linalg.inv(m.astype(np.float64)).T[:3, :3]
            nrm = nrm.astype(np.float64) @ inv_t.T
            nrm = (nrm / np.linalg.norm(nrm, axis=1, keepdims=True)).astype(np.float32)
    return pos, faces, nrm


def save_serialized(path, meshes_list, version=4, double_precision=False):
    """Writer for the same format (tests, scene conversion): `meshes_list` = [(positions, faces, normals or None), ...]."""
    import struct
    import zlib
    ft = "<f8" if double_precision else "<f4"
    out, offsets = bytearray(), []
    for pos, faces, nrm in meshes_list:
        offsets.append(len(out))
        flags = (0x2000 if double_precision else 0x1000) | (0x0001 if nrm is not None else 0)
        body = struct.pack("<I", flags) + (b"mesh\0" if version == 4 else b"") + struct.pack("<QQ", len(pos), len(faces))
        body += np.asarray(pos, dtype=ft).tobytes()
        if nrm is not None:
            body += np.asarray(nrm, dtype=ft).tobytes()
        body += np.asarray(faces, dtype="<u4").tobytes()
        out += struct.pack("<HH", 0x041C, version) + zlib.compress(body)
    for o in offsets:
        out += struct.pack("<Q" if version == 4 else "<I", o)
    out += struct.pack("<I", len(offsets))
    with open(path, "wb") as f:
        f.write(bytes(out))


def vertex_normals(pos, faces):
    """Angle-weighted vertex normals (Mesh::recompute_vertex_normals, src/librender/mesh.cpp:201-278)."""
    p = np.asarray(pos, dtype=np.float32).astype(np.float64)
    f = np.asarray(faces, dtype=np.int64)
    v0, v1, v2 = p[f[:, 0]], p[f[:, 1]], p[f[:, 2]]
    n = np.cross(v1 - v0, v2 - v0)
    ln = np.linalg.norm(n, axis=1)
    ok = ln > 0
    n[ok] /= ln[ok, None]

    def ang(a, b):
        a = a / np.maximum(np.linalg.norm(a, axis=1, keepdims=True), 1e-300)
        b = b / np.maximum(np.linalg.norm(b, axis=1, keepdims=True), 1e-300)
        return np.arccos(np.clip(np.sum(a * b, axis=1), -1.0, 1.0))

    angles = [ang(v1 - v0, v2 - v0), ang(v2 - v1, v0 - v1), ang(v0 - v2, v1 - v2)]
    out = np.zeros_like(p)
    for j in range(3):
        w = np.where(ok, angles[j], 0.0)
        np.add.at(out, f[:, j], n * w[:, None])
    l = np.linalg.norm(out, axis=1)
    bad = l == 0
    out[~bad] /= l[~bad, None]
    out[bad] = [1.0, 0.0, 0.0]
    return out.astype(np.float32)


# ---- procedural geometry ------------------------------------------------------------------------------------
def grid_faces(nu, nv, wrap_u):
    """Two triangles per cell of an nu x nv vertex grid (row-major, u fastest)."""
    cu = nu if wrap_u else nu - 1
    u, v = np.meshgrid(np.arange(cu), np.arange(nv - 1), indexing="xy")
    u, v = u.ravel(), v.ravel()
    u1 = (u + 1) % nu
    a, b, c, d = v * nu + u, v * nu + u1, (v + 1) * nu + u1, (v + 1) * nu + u
    return np.stack([np.stack([a, b, c], 1), np.stack([a, c, d], 1)], 1).reshape(-1, 3).astype(np.uint32)


def car_body(n_u=264, n_v=133, length=4.5, width=1.8, height=1.4):
    """Procedural stand-in for python_scripts/Car-body.ply (listed in .MISSING_LARGE_BLOBS): a closed
    super-ellipsoid hull with a cabin, wheel arches and panel ripple.  Default 2*264*132 = 69 696
    triangles, the size class of the reference's bunny.ply stand-in (69 451).  Object space: x forward,
    y left, z up, centred on the origin."""
    u = np.linspace(0.0, 2.0 * np.pi, n_u, endpoint=False)
    v = np.linspace(0.0, np.pi, n_v)
    uu, vv = np.meshgrid(u, v, indexing="xy")

    def spow(x, e):
        return np.sign(x) * np.abs(x) ** e

    e1, e2 = 0.45, 0.6
    x = 0.5 * length * spow(np.sin(vv), e1) * spow(np.cos(uu), e2)
    y = 0.5 * width * spow(np.sin(vv), e1) * spow(np.sin(uu), e2)
    z = 0.5 * height * 0.62 * spow(np.cos(vv), e1)
    # cabin: raise the roof over the middle of the car
    cab = np.exp(-((x / (0.22 * length)) ** 2)) * np.exp(-((y / (0.42 * width)) ** 4))
    z = z + np.where(z > 0, 0.5 * height * 0.75 * cab * (z / (0.5 * height * 0.62)), 0.0)
    # wheel arches: four indentations on the lower flanks
    for wx in (-0.3 * length, 0.3 * length):
        for wy in (-0.5 * width, 0.5 * width):
            dsq = ((x - wx) / 0.38) ** 2 + ((z + 0.25 * height) / 0.34) ** 2
            arch = np.exp(-dsq * 2.0) * (np.abs(y - wy) < 0.35 * width)
            y = y - np.sign(wy) * 0.18 * arch
    # panel ripple (keeps the BVH honest: no large flat regions)
    z = z + 0.004 * np.sin(37.0 * uu) * np.sin(29.0 * vv)
    pos = np.stack([x, y, z], -1).reshape(-1, 3).astype(np.float32)
    return pos, grid_faces(n_u, n_v, True)


def stairs(n_steps):
    """The stairs fixture of src/librender/tests/mesh_generation.py:30-60 (same vertex and face numbering):
    n_steps treads of unit width along y, tread i at height z = i / n_steps, joined by vertical risers.
    Known answer (test_kdtrees.py:25-57): rays from z = 2 going down hit at t = 2 - floor(y n) / n."""
    h = 1.0 / n_steps
    pos = np.zeros((4 * n_steps, 3), np.float32)
    faces = []
    for i in range(n_steps):
        k, y0, y1, z = 4 * i, i * h, (i + 1) * h, i * h
        pos[k:k + 4] = [[0, y0, z], [1, y0, z], [0, y1, z], [1, y1, z]]
        faces += [[k, k + 1, k + 2], [k + 1, k + 3, k + 2]]
        if i < n_steps - 1:
            faces += [[k + 2, k + 3, k + 5], [k + 5, k + 4, k + 2]]
    return pos, np.array(faces, np.uint32)


def _pcg32_floats(n, seed=42, seq=54):
    """n uniform floats in [0, 1) from one PCG32 stream (the generator of the procedural city)."""
    mask = (1 << 64) - 1
    mult, inc = 0x5851f42d4c957f2d, ((seq << 1) | 1) & mask
    state = 0
    state = (state * mult + inc) & mask
    state = (state + seed) & mask
    state = (state * mult + inc) & mask
    out = np.empty(n, dtype=np.float32)
    for i in range(n):
        old = state
        state = (old * mult + inc) & mask
        xs = (((old >> 18) ^ old) >> 27) & 0xffffffff
        rot = old >> 59
        u = ((xs >> rot) | (xs << ((-rot) & 31))) & 0xffffffff
        out[i] = np.uint32((u >> 9) | 0x3f800000).view(np.float32) - np.float32(1.0)
    return out


def city_blocks(n_side=100, pitch=10.0, footprint=6.0, h_min=5.0, h_max=60.0, face_grid=10, seed=42):
    """Procedural urban scene (BASELINE config 5 stand-in): n_side x n_side axis-aligned boxes on a square lattice of
    `pitch` metres centred on the origin (streets run along the multiples of `pitch`, so the origin is a crossing), footprint
    `footprint` x `footprint`, heights ~ U(h_min, h_max) from PCG32(seed); the 5 exposed faces of a box are tessellated into
    face_grid x face_grid x 2 triangles each (default 1000 triangles per block: 100 x 100 blocks = 1.0e7 triangles)."""
    heights = h_min + (h_max - h_min) * _pcg32_floats(n_side * n_side, seed)
    g = face_grid
    a, b = np.meshgrid(np.linspace(0.0, 1.0, g + 1), np.linspace(0.0, 1.0, g + 1), indexing="xy")
    a, b = a.reshape(-1), b.reshape(-1)
    quad = grid_faces(g + 1, g + 1, False)                    # (2 g^2, 3) over a (g+1) x (g+1) vertex grid
    nv_face = (g + 1) * (g + 1)
    # unit box [0,1]^3 faces as (origin, du, dv): four walls + roof, outward orientation
    faces = [((0, 0, 0), (1, 0, 0), (0, 0, 1)), ((1, 0, 0), (0, 1, 0), (0, 0, 1)), ((1, 1, 0), (-1, 0, 0), (0, 0, 1)),
             ((0, 1, 0), (0, -1, 0), (0, 0, 1)), ((0, 0, 1), (1, 0, 0), (0, 1, 0))]
    unit = np.concatenate([np.asarray(o, np.float64)[None] + a[:, None] * np.asarray(du, np.float64)[None] + b[:, None] * np.asarray(dv, np.float64)[None]
                           for o, du, dv in faces])            # (5 nv_face, 3)
    unit_faces = np.concatenate([quad + k * nv_face for k in range(5)])
    n_blocks = n_side * n_side
    ii, jj = np.meshgrid(np.arange(n_side), np.arange(n_side), indexing="ij")
    cx = (ii.reshape(-1) + 0.5 - n_side / 2.0) * pitch
    cy = (jj.reshape(-1) + 0.5 - n_side / 2.0) * pitch
    scale = np.stack([np.full(n_blocks, footprint), np.full(n_blocks, footprint), heights.astype(np.float64)], -1)
    origin = np.stack([cx - footprint / 2.0, cy - footprint / 2.0, np.zeros(n_blocks)], -1)
    pos = (unit[None] * scale[:, None] + origin[:, None]).reshape(-1, 3).astype(np.float32)
    idx = (unit_faces[None].astype(np.int64) + (np.arange(n_blocks) * unit.shape[0])[:, None, None]).reshape(-1, 3).astype(np.uint32)
    return pos, idx
