"""Host-side scene loading for the B200 volpath: Mitsuba-style scene dictionaries
(the shape `mitsuba.core.xml.load_dict` takes, src/libcore/python/xml_v.cpp:100-240)
-> the flat B2vpSceneDesc of include/b2vp.h.

Each plugin parser below reads the same properties, with the same defaults, as the
reference plugin of that name (file:line cited per parser) and evaluates derived
quantities in the variant's scalar type (float32 for scalar_mono, float64 for
scalar_mono_double).  Content outside SURVEY.md section 8 raises, like an unknown
plugin name does in the reference.
"""
import ctypes as C
import math
import os
import struct
import zlib

import numpy as np

from . import abi
from .transform import Transform, coordinate_system

VARIANTS = {"scalar_mono": (np.float32, abi.PREC_F32), "scalar_mono_double": (np.float64, abi.PREC_F64)}


def read_vol(path):
    """.vol reader (src/textures/volume_data.h:42-102): returns (data[z,y,x,c] float32, bbox 6 floats)."""
    with open(path, "rb") as f:
        raw = f.read()
    if raw[:3] != b"VOL" or raw[3] != 3:
        raise RuntimeError("%s: not a version-3 VOL file" % path)
    typ, nx, ny, nz, ch = struct.unpack_from("<5i", raw, 4)
    if typ != 1:
        raise RuntimeError("%s: only float32 volumes are supported (type %d)" % (path, typ))
    if nx * ny * nz < 8:
        raise RuntimeError("%s: invalid grid dimensions: %d x %d x %d < 8 (must have at least one value at each corner)" %
                           (path, nx, ny, nz))
    bbox = struct.unpack_from("<6f", raw, 24)
    data = np.frombuffer(raw, dtype="<f4", count=nx * ny * nz * ch, offset=48).reshape(nz, ny, nx, ch).copy()
    return data, bbox


def write_vol(path, data, bbox=(0, 0, 0, 1, 1, 1)):
    """Writes data[z,y,x(,c)] as a version-3 float32 VOL file."""
    data = np.asarray(data, dtype="<f4")
    if data.ndim == 3:
        data = data[..., None]
    nz, ny, nx, ch = data.shape
    with open(path, "wb") as f:
        f.write(b"VOL\x03")
        f.write(struct.pack("<5i", 1, nx, ny, nz, ch))
        f.write(struct.pack("<6f", *bbox))
        f.write(data.tobytes())


_SER_HEADER, _SER_NORMALS, _SER_TEXCOORDS, _SER_COLORS, _SER_DOUBLE = 0x041C, 0x0001, 0x0002, 0x0008, 0x2000


def read_serialized(path, shape_index=0):
    """One mesh of a `.serialized` container as src/shapes/serialized.cpp:174-340 reads it.  Little endian; every
    sub-mesh is [u16 0x041C][u16 version 3|4] followed by a zlib stream of: u32 flags, (v4) a NUL-terminated name,
    u64 vertex count, u64 face count, positions, [normals], [texcoords], [colours, skipped], u32 face indices;
    reals are float32 or (flag 0x2000) float64 narrowed to float32.  For shape_index > 0 the offset comes from the
    dictionary at the end of the file (v4: u64 offsets, v3: u32 offsets, then a u32 count).  Returns (positions,
    faces, normals or None, texcoords or None), positions and normals still in object space."""
    def fail(msg):
        raise RuntimeError('Error while loading serialized file "%s": %s!' % (os.path.basename(path), msg))
    if not os.path.exists(path):
        fail("file not found")
    if shape_index < 0:
        fail("shape index must be nonnegative!")
    with open(path, "rb") as f:
        raw = f.read()
    fmt, version = struct.unpack_from("<HH", raw, 0)
    if fmt != _SER_HEADER:
        fail("encountered an invalid file format!")
    if version not in (3, 4):
        fail("encountered an incompatible file version!")
    start = 0
    if shape_index != 0:
        (count,) = struct.unpack_from("<I", raw, len(raw) - 4)
        if shape_index > count:
            fail("Unable to unserialize mesh, shape index is out of range! (requested %d out of 0..%d)"
                 % (shape_index, count - 1))
        if version == 4:
            (start,) = struct.unpack_from("<Q", raw, len(raw) - 8 * (count - shape_index) - 4)
        else:
            (start,) = struct.unpack_from("<I", raw, len(raw) - 4 * (count - shape_index + 1))
    try:
        data = zlib.decompressobj().decompress(raw[start + 4:])
    except zlib.error as e:
        fail("inflate(): %s" % e)
    (flags,) = struct.unpack_from("<I", data, 0)
    off = 4
    if version == 4:
        off = data.index(b"\0", off) + 1
    n_vert, n_face = struct.unpack_from("<QQ", data, off)
    off += 16
    real = "<f8" if flags & _SER_DOUBLE else "<f4"

    def reals(dim):
        nonlocal off
        a = np.frombuffer(data, real, n_vert * dim, off).reshape(n_vert, dim)
        off += a.nbytes
        return a.astype(np.float32)

    pos = reals(3)
    nrm = reals(3) if flags & _SER_NORMALS else None
    uv = reals(2) if flags & _SER_TEXCOORDS else None
    if flags & _SER_COLORS:
        reals(3)
    faces = np.frombuffer(data, "<u4", n_face * 3, off).reshape(n_face, 3).astype(np.uint32)
    return pos, faces, nrm, uv


def write_serialized(path, meshes, version=4, double=False):
    """Writes meshes = [(name, positions, faces, normals or None, texcoords or None), ...] as one `.serialized`
    container (the layout read_serialized documents), with the offset dictionary at the end."""
    offsets, out = [], bytearray()
    for name, pos, faces, nrm, uv in meshes:
        offsets.append(len(out))
        flags = (_SER_DOUBLE if double else 0x1000) | (_SER_NORMALS if nrm is not None else 0) \
            | (_SER_TEXCOORDS if uv is not None else 0)
        real = "<f8" if double else "<f4"
        body = struct.pack("<I", flags) + (name.encode() + b"\0" if version == 4 else b"")
        body += struct.pack("<QQ", len(pos), len(faces)) + np.asarray(pos, real).tobytes()
        for a in (nrm, uv):
            if a is not None:
                body += np.asarray(a, real).tobytes()
        body += np.asarray(faces, "<u4").tobytes()
        out += struct.pack("<HH", _SER_HEADER, version) + zlib.compress(body)
    out += b"".join(struct.pack("<Q" if version == 4 else "<I", o) for o in offsets) + struct.pack("<I", len(offsets))
    with open(path, "wb") as f:
        f.write(bytes(out))


def read_obj(path, flip_tex_coords=True):
    """Wavefront OBJ as src/shapes/obj.cpp:140-325 reads it: `v`, `vn`, `vt`, `f`; one output vertex per distinct
    (v, vt, vn) index triple, numbered in order of first use; polygons as triangle fans.  Returns
    (positions[n, 3], faces[m, 3], normals[n, 3] or None, texcoords[n, 2] or None), all float32 / uint32,
    positions and normals still in object space."""
    vs, vns, vts, tris = [], [], [], []
    ids, order = {}, []
    with open(path, "r") as f:
        for line in f:
            tok = line.split()
            if not tok:
                continue
            if tok[0] == "v":
                vs.append([np.float32(x) for x in tok[1:4]])
            elif tok[0] == "vn":
                vns.append([np.float32(x) for x in tok[1:4]])
            elif tok[0] == "vt":
                u, v = np.float32(tok[1]), np.float32(tok[2])
                vts.append([u, np.float32(1) - v if flip_tex_coords else v])
            elif tok[0] == "f":
                corner = []
                for item in tok[1:]:
                    parts = (item.split("/") + ["", ""])[:3]
                    key = tuple(int(x) if x else 0 for x in parts)
                    if not 1 <= key[0] <= len(vs):
                        raise RuntimeError("Error while loading OBJ file %r: reference to invalid vertex %d!" % (path, key[0]))
                    if key not in ids:
                        ids[key] = len(order)
                        order.append(key)
                    corner.append(ids[key])
                for k in range(1, len(corner) - 1):
                    tris.append([corner[0], corner[k], corner[k + 1]])
    pos = np.array([vs[k[0] - 1] for k in order], dtype=np.float32).reshape(-1, 3)
    nrm = uv = None
    if vns:
        nrm = np.array([vns[k[2] - 1] if k[2] else [0, 0, 0] for k in order], dtype=np.float32).reshape(-1, 3)
    if vts:
        uv = np.array([vts[k[1] - 1] if k[1] else [0, 0] for k in order], dtype=np.float32).reshape(-1, 2)
    return pos, np.array(tris, dtype=np.uint32).reshape(-1, 3), nrm, uv


_PLY_TYPES = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2",
              "ushort": "u2", "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4",
              "float": "f4", "float32": "f4", "double": "f8", "float64": "f8"}


def read_ply(path):
    """Stanford PLY as src/shapes/ply.cpp:102-380 reads it: ascii or binary (either byte order); element `vertex`
    with x y z, optional nx ny nz and texture coordinates (u v / texture_u texture_v / s t); element `face` with a
    `vertex_index(es)` list of exactly three indices per face (anything else is "incompatible contents -- is this a
    triangle mesh?"); other elements and properties are skipped.  Returns (positions, faces, normals or None,
    texcoords or None) as float32 / uint32, positions and normals still in object space."""
    def fail(msg):
        raise RuntimeError('Error while loading PLY file "%s": %s!' % (os.path.basename(path), msg))
    with open(path, "rb") as f:
        raw = f.read()
    end = raw.find(b"end_header")
    if not raw.startswith(b"ply") or end < 0:
        fail("invalid PLY header")
    nl = raw.index(b"\n", end) + 1
    fmt, elements = None, []
    for line in raw[:end].decode("ascii", "replace").splitlines()[1:]:
        tok = line.split()
        if not tok or tok[0] in ("comment", "obj_info"):
            continue
        if tok[0] == "format":
            fmt = tok[1]
        elif tok[0] == "element":
            elements.append({"name": tok[1], "count": int(tok[2]), "props": []})
        elif tok[0] == "property":
            if tok[1] == "list":
                elements[-1]["props"].append((tok[4], "list", _PLY_TYPES[tok[2]], _PLY_TYPES[tok[3]]))
            else:
                elements[-1]["props"].append((tok[2], _PLY_TYPES[tok[1]]))
    if fmt not in ("ascii", "binary_little_endian", "binary_big_endian"):
        fail("invalid PLY format")
    order = ">" if fmt == "binary_big_endian" else "<"
    pos = faces = nrm = uv = None
    off = nl
    words = raw[nl:].split() if fmt == "ascii" else None
    wi = 0
    for el in elements:
        n, props = el["count"], el["props"]
        has_list = any(p[1] == "list" for p in props)
        if has_list and not (el["name"] == "face" and len([p for p in props if p[1] == "list"]) == 1):
            fail("unsupported list property in element '%s'" % el["name"])
        if has_list:   # every list must hold exactly three entries: then the records are fixed-size
            fields = []
            for p in props:
                if p[1] == "list":
                    fields += [(p[0] + ".count", p[2]), (p[0], p[3], (3,))]
                else:
                    fields.append((p[0], p[1]))
        else:
            fields = [(p[0], p[1]) for p in props]
        if fmt == "ascii":
            per = sum(int(np.prod(fl[2])) if len(fl) > 2 else 1 for fl in fields)
            vals = np.array(words[wi:wi + n * per], dtype=np.float64).reshape(n, per) if n else np.zeros((0, per))
            wi += n * per
            rec, c = {}, 0
            for fl in fields:
                k = int(np.prod(fl[2])) if len(fl) > 2 else 1
                rec[fl[0]] = vals[:, c] if k == 1 else vals[:, c:c + k]
                c += k
        else:
            dt = np.dtype([(fl[0], order + fl[1]) + ((fl[2],) if len(fl) > 2 else ()) for fl in fields])
            if off + n * dt.itemsize > len(raw):
                fail("file is truncated")
            arr = np.frombuffer(raw, dtype=dt, count=n, offset=off)
            off += n * dt.itemsize
            rec = {name: arr[name] for name in dt.names}
        if el["name"] == "vertex":
            if not all(k in rec for k in "xyz"):
                fail("vertex positions not found")
            pos = np.stack([rec["x"], rec["y"], rec["z"]], axis=1).astype(np.float32)
            if all(k in rec for k in ("nx", "ny", "nz")):
                nrm = np.stack([rec["nx"], rec["ny"], rec["nz"]], axis=1).astype(np.float32)
            for a, b in (("u", "v"), ("texture_u", "texture_v"), ("s", "t")):
                if a in rec and b in rec:
                    uv = np.stack([rec[a], rec[b]], axis=1).astype(np.float32)
                    break
        elif el["name"] == "face":
            name = "vertex_index" if "vertex_index" in rec else "vertex_indices" if "vertex_indices" in rec else None
            if name is None:
                fail("vertex_index/vertex_indices property not found")
            if n and not (np.asarray(rec[name + ".count"]) == 3).all():
                fail("incompatible contents -- is this a triangle mesh?")
            faces = np.asarray(rec[name]).astype(np.uint32).reshape(-1, 3)
    if fmt != "ascii" and off != len(raw):
        fail("invalid file -- trailing content")
    if pos is None or faces is None:
        fail("vertex / face element not found")
    return pos, faces, nrm, uv



def rfilter_table(node, dtype):
    """The film's reconstruction filter as ImageBlock::put reads it: radius, the 32-entry table of
    ReconstructionFilter::init_discretization (src/libcore/rfilter.cpp:9-21) with its scale factor and border, and the
    footprint width of imageblock.cpp:122.  The filters are src/rfilters/{box,tent,gaussian,mitchell,catmullrom,
    lanczos}.cpp, evaluated in the variant's scalar type `dtype`; exp and sin are taken in double and rounded, which
    is where a single-precision table can differ from the reference's expf/sinf by one unit in the last place."""
    f = dtype
    t = node.get("type", "box")
    eps = f(np.finfo(dtype).eps) / f(2)
    ray_eps = eps * f(1500)                                    # math.h:36-37
    known = {"box": ("radius",), "tent": (), "gaussian": ("stddev",), "mitchell": ("B", "C"), "catmullrom": (),
             "lanczos": ("lobes",)}
    if t not in known:
        raise RuntimeError("reconstruction filter plugin %r is not supported by the B200 volpath" % t)
    for key in node:
        if key not in ("type", "id") + known[t]:
            raise RuntimeError("Unreferenced property %r in plugin of type %r" % (key, t))
    if t == "box":
        radius = f(np.float32(node.get("radius", 0.5))) + ray_eps   # box.cpp:28; Properties keeps floats in single precision
        ev = lambda x: f(1) if abs(x) <= radius else f(0)
    elif t == "tent":
        radius = f(1)
        inv = f(1) / radius
        ev = lambda x: max(f(0), f(1) - abs(x * inv))
    elif t == "gaussian":
        stddev = f(np.float32(node.get("stddev", 0.5)))
        radius = f(4) * stddev                                 # gaussian.cpp:31-36
        alpha = f(-1) / (f(2) * stddev * stddev)
        bias = f(math.exp(float(alpha * (radius * radius))))
        ev = lambda x: max(f(0), f(math.exp(float(alpha * (x * x)))) - bias)
    elif t in ("mitchell", "catmullrom"):
        radius = f(2)
        third = np.float32(1) / np.float32(3)                  # the defaults and 1/6 are float literals in every variant
        B = f(np.float32(node.get("B", third))) if t == "mitchell" else f(0)
        Cc = f(np.float32(node.get("C", third))) if t == "mitchell" else f(0.5)
        sixth = f(np.float32(1) / np.float32(6))

        def ev(x):                                             # mitchell.cpp:40-53, catmullrom.cpp:27-42
            x = abs(x)
            x2 = x * x
            x3 = x2 * x
            if x < 1:
                r = (f(12) - f(9) * B - f(6) * Cc) * x3 + (f(-18) + f(12) * B + f(6) * Cc) * x2 + (f(6) - f(2) * B)
            else:
                r = ((-B - f(6) * Cc) * x3 + (f(6) * B + f(30) * Cc) * x2 + (f(-12) * B - f(48) * Cc) * x
                     + (f(8) * B + f(24) * Cc))
            return sixth * r if x < 2 else f(0)
    else:                                                      # lanczos.cpp:34-48
        radius = f(int(node.get("lobes", 3)))
        pi = f(math.pi)

        def ev(x):
            x = abs(x)
            if x < eps:
                return f(1)
            if x > radius:
                return f(0)
            x1 = pi * x
            x2 = x1 / radius
            return f(f(math.sin(float(x1))) * f(math.sin(float(x2)))) / (x1 * x2)
    if not radius > 0:
        raise RuntimeError("reconstruction filter: the radius must be positive")
    res = abi.RFILTER_RESOLUTION
    with np.errstate(all="ignore"):
        values = [float(f(ev((radius * f(i)) / f(res)))) for i in range(res)] + [0.0]
        scale = f(res) / radius
        border = int(math.ceil(float(radius - f(0.5) - f(2) * ray_eps)))
        taps = int(math.ceil(float((radius - f(2) * ray_eps) * f(2))))
    if radius > f(0.5) + ray_eps and taps > abi.RFILTER_MAX_TAPS:
        raise RuntimeError("reconstruction filter: a radius of %g pixels is more than the B200 volpath carries (%g)"
                           % (float(radius), abi.RFILTER_MAX_TAPS / 2))
    return {"radius": float(radius), "scale": float(scale), "taps": taps, "border": max(border, 0), "values": values}


class _Builder:
    def __init__(self, variant):
        if variant not in VARIANTS:
            raise RuntimeError("variant %r is not supported by the B200 volpath (mono only)" % (variant,))
        self.variant = variant
        self.dtype, self.precision = VARIANTS[variant]
        self.dt = np.dtype(self.dtype).type
        self.shapes, self.bsdfs, self.media, self.volumes, self.phases, self.emitters = [], [], [], [], [], []
        self.volume_arrays = []
        self.vpos, self.vnrm, self.vuv, self.faces = [], [], [], []
        self.n_vertices = 0
        self.n_faces = 0
        self.ids = {}        # id -> ("medium"|"bsdf"|"shape", index)
        self.bbox_min = np.full(3, np.inf, dtype=self.dtype)
        self.bbox_max = np.full(3, -np.inf, dtype=self.dtype)
        self.sensor = None
        self.shape_sensor = None   # (sensor node, index of its parent shape)
        self.integrator = None
        self.keep = []

    # ------------------------------------------------------------ helpers
    def prop(self, v):
        """A float property as a plugin reads it: Properties stores floats, 3-vectors and transforms in single
        precision in every variant (include/mitsuba/core/properties.h:55-56), the plugin then widens it."""
        return self.dt(np.float32(v))

    def prop3(self, v):
        return np.asarray(v, dtype=np.float32).astype(self.dtype)

    def xf(self, v):
        """props.transform(): a Transform4f in single precision (matrix and inverse transpose), widened."""
        if v is None:
            return Transform(dtype=self.dtype)
        if isinstance(v, Transform):
            return v.astype(np.float32).astype(self.dtype)
        return Transform(np.asarray(v, dtype=np.float64).reshape(4, 4), dtype=np.float32).astype(self.dtype)

    def mat(self, t):
        m = abi.Mat4()
        m.m[:] = t.colmajor()
        return m

    def expand(self, p):
        p = np.asarray(p, dtype=self.dtype)
        self.bbox_min = np.minimum(self.bbox_min, p)
        self.bbox_max = np.maximum(self.bbox_max, p)

    def resolve(self, node, kind, parser):
        if isinstance(node, dict) and node.get("type") == "ref":
            k, idx = self.ids[node["id"]]
            if k != kind:
                raise RuntimeError("reference %r is a %s, expected a %s" % (node["id"], k, kind))
            return idx
        return parser(node)

    # ------------------------------------------------------------ plugins
    def volume(self, node, default):
        """props.volume(): a float becomes a constvolume (src/libcore/properties.cpp); gridvolume =
        src/textures/grid3d.cpp:43-160; constvolume = src/textures/constant3d.cpp:13-55."""
        v = abi.Volume()
        if node is None:
            node = default
        if isinstance(node, (int, float)):
            node = {"type": "constvolume", "color": float(node)}
        t = node["type"]
        to_world = self.xf(node.get("to_world"))
        w2l = to_world.inverse()
        arr = None
        if t == "constvolume":
            v.type = abi.VOLUME_CONST
            if "value" in node:   # the plugin's property is `color` (constant3d.cpp:14); the reference rejects others
                raise RuntimeError('Unreferenced property "value" in plugin of type "constvolume" (the property is "color")')
            color = node.get("color", 1.0)
            if isinstance(color, dict):
                color = color.get("value", 1.0)   # a nested `uniform` spectrum
            v.value = float(self.prop(color))
            v.max_value = v.value
            v.channels = 1
        elif t == "gridvolume":
            v.type = abi.VOLUME_GRID
            if "filename" in node:
                data, file_bbox = read_vol(node["filename"])
            else:  # in-memory grid: data[z, y, x(, c)]
                data = np.asarray(node["data"], dtype=np.float32)
                if data.ndim == 3:
                    data = data[..., None]
                file_bbox = node.get("bbox", (0, 0, 0, 1, 1, 1))
            nz, ny, nx, ch = data.shape
            if ch not in (1, 3):
                raise RuntimeError("Unsupported channel count: %d (expected 1 or 3)" % ch)
            if node.get("raw", False):
                raise RuntimeError("gridvolume raw=true is not supported in mono variants")
            ft = node.get("filter_type", "trilinear")
            wm = node.get("wrap_mode", "clamp")
            try:
                v.filter = {"nearest": abi.FILTER_NEAREST, "trilinear": abi.FILTER_TRILINEAR}[ft]
                v.wrap = {"repeat": abi.WRAP_REPEAT, "mirror": abi.WRAP_MIRROR, "clamp": abi.WRAP_CLAMP}[wm]
            except KeyError:
                raise RuntimeError("Invalid filter type / wrap mode %r / %r" % (ft, wm))
            v.channels = ch
            v.res[:] = [nx, ny, nz]
            arr = np.ascontiguousarray(data, dtype=np.float32)
            v.max_value = float(self.prop(node["max_value"])) if "max_value" in node else float(arr.max())
            if node.get("use_grid_bbox", False):  # grid3d.cpp:152-155
                mn = np.asarray(file_bbox[:3], dtype=self.dtype)
                mx = np.asarray(file_bbox[3:], dtype=self.dtype)
                meta = (Transform.scale(self.dt(1) / (mx - mn), self.dtype) @ Transform.translate(-mn, self.dtype))
                w2l = meta @ w2l
            v.data = arr.ctypes.data_as(C.POINTER(C.c_float))
        else:
            raise RuntimeError("volume plugin %r is not supported by the B200 volpath" % t)
        v.world_to_local = self.mat(w2l)
        l2w = w2l.inverse()  # texture.h:265-272
        a, b = l2w.point([0, 0, 0]), l2w.point([1, 1, 1])
        v.bbox_min[:] = [float(x) for x in np.minimum(a, b)]
        v.bbox_max[:] = [float(x) for x in np.maximum(a, b)]
        self.volumes.append(v)
        self.volume_arrays.append(arr)
        return len(self.volumes) - 1

    def phase(self, node):
        """src/phase/{isotropic,rayleigh,hg}.cpp"""
        p = abi.Phase()
        t = "isotropic" if node is None else node["type"]
        if t == "isotropic":
            p.type = abi.PHASE_ISOTROPIC
        elif t == "rayleigh":
            p.type = abi.PHASE_RAYLEIGH
        elif t == "hg":
            p.type = abi.PHASE_HG
            g = self.prop(node.get("g", 0.8))  # hg.cpp:41
            if not (-1 < g < 1):
                raise RuntimeError("The asymmetry parameter must lie in the interval (-1, 1)!")
            p.g = float(g)
        elif t == "blendphase":  # blendphase.cpp:42-66: two nested phase functions in key order + a weight volume
            kids = [v for k, v in sorted(node.items()) if isinstance(v, dict) and k != "weight"
                    and v.get("type") in ("isotropic", "rayleigh", "hg", "tabphase")]
            if len(kids) != 2:
                raise RuntimeError("BlendPhase: Two child phase functions must be specified!")
            if "weight" not in node:
                raise RuntimeError('Property "weight" has not been specified!')
            c0, c1 = self.phase(kids[0]), self.phase(kids[1])
            p.type = abi.PHASE_BLEND
            p.child[0], p.child[1] = c0, c1
            p.weight_volume = self.volume(node["weight"], None)
        elif t == "tabphase":  # tabphase.cpp:43-66: `values` is a string (or here also a sequence) of samples
            p.type = abi.PHASE_TABULATED
            vals = node["values"]
            if isinstance(vals, str):
                vals = [float(x) for x in vals.replace(",", " ").split()]
            vals = np.asarray(vals, dtype=np.float64).astype(self.dtype).astype(np.float64)  # (ScalarFloat) std::stod
            if len(vals) < 2:
                raise RuntimeError("ContinuousDistribution: needs at least two entries!")
            if (vals < 0).any():
                raise RuntimeError("ContinuousDistribution: entries must be non-negative!")
            if not ((vals[:-1] + vals[1:]) > 0).any():
                raise RuntimeError("ContinuousDistribution: no probability mass found!")
            vals = np.ascontiguousarray(vals)
            self.keep.append(vals)
            p.n_values = len(vals)
            p.values = vals.ctypes.data_as(C.POINTER(C.c_double))
        else:
            raise RuntimeError("phase function %r is not supported by the B200 volpath" % t)
        self.phases.append(p)
        return len(self.phases) - 1

    def medium(self, node):
        """src/media/homogeneous.cpp:21-54, heterogeneous.cpp:21-54, src/librender/medium.cpp:12-30"""
        m = abi.Medium()
        t = node["type"]
        if t == "homogeneous":
            m.type = abi.MEDIUM_HOMOGENEOUS
        elif t == "heterogeneous":
            m.type = abi.MEDIUM_HETEROGENEOUS
        else:
            raise RuntimeError("medium plugin %r is not supported by the B200 volpath" % t)
        m.albedo = self.volume(node.get("albedo"), 0.75)
        m.sigma_t = self.volume(node.get("sigma_t"), 1.0)
        m.scale = float(self.prop(node.get("scale", 1.0)))
        m.has_spectral_extinction = int(bool(node.get("has_spectral_extinction", True)))
        m.sample_emitters = int(bool(node.get("sample_emitters", True)))
        m.phase = self.phase(node.get("phase"))
        if m.type == abi.MEDIUM_HETEROGENEOUS and self.volumes[m.sigma_t].type != abi.VOLUME_GRID:
            raise RuntimeError("heterogeneous medium needs a gridvolume sigma_t (ConstVolume::max is not implemented)")
        self.media.append(m)
        idx = len(self.media) - 1
        if "id" in node:
            self.ids[node["id"]] = ("medium", idx)
        return idx

    def bsdf(self, node):
        """src/bsdfs/diffuse.cpp:70-76, rpv.cpp:66-83, null.cpp:30-39, bilambertian.cpp:50-61"""
        b = abi.Bsdf()
        if node is None:
            node = {"type": "diffuse"}  # shape.cpp: default BSDF
        t = node["type"]

        def scalar(key, default):
            val = node.get(key, default)
            if isinstance(val, dict):
                if val.get("type") != "uniform":
                    raise RuntimeError("only uniform (constant) textures are supported for %r" % key)
                val = val.get("value", default)
            return float(self.prop(val))
        if t == "diffuse":
            b.type = abi.BSDF_DIFFUSE
            b.p[0] = scalar("reflectance", 0.5)
        elif t == "null":
            b.type = abi.BSDF_NULL
        elif t == "rpv":
            b.type = abi.BSDF_RPV
            b.p[0] = scalar("rho_0", 0.1)
            b.p[1] = scalar("k", 0.1)
            b.p[2] = scalar("g", 0.0)
            b.p[3] = scalar("rho_c", b.p[0]) if "rho_c" in node else b.p[0]  # rpv.cpp:72-80
        elif t == "bilambertian":  # bilambertian.cpp:51-53
            b.type = abi.BSDF_BILAMBERTIAN
            b.p[0] = scalar("reflectance", 0.5)
            b.p[1] = scalar("transmittance", 0.5)
        else:
            raise RuntimeError("BSDF plugin %r is not supported by the B200 volpath" % t)
        self.bsdfs.append(b)
        idx = len(self.bsdfs) - 1
        if "id" in node:
            self.ids[node["id"]] = ("bsdf", idx)
        return idx

    def add_mesh(self, s, positions, faces, normals=None, texcoords=None):
        positions = np.ascontiguousarray(positions, dtype=np.float32).reshape(-1, 3)
        faces = np.ascontiguousarray(faces, dtype=np.uint32).reshape(-1, 3)
        s.type = abi.SHAPE_MESH
        s.vertex_offset, s.vertex_count = self.n_vertices, len(positions)
        s.face_offset, s.face_count = self.n_faces, len(faces)
        s.has_vertex_normals = int(normals is not None)
        s.has_vertex_texcoords = int(texcoords is not None)
        self.vpos.append(positions)
        self.vnrm.append(np.zeros_like(positions) if normals is None else
                         np.ascontiguousarray(normals, dtype=np.float32).reshape(-1, 3))
        self.vuv.append(np.zeros((len(positions), 2), np.float32) if texcoords is None else
                        np.ascontiguousarray(texcoords, dtype=np.float32).reshape(-1, 2))
        self.faces.append(faces)
        self.n_vertices += len(positions)
        self.n_faces += len(faces)
        if len(positions):
            self.expand(positions.min(axis=0))
            self.expand(positions.max(axis=0))

    def shape(self, node):
        s = abi.Shape()
        t = node["type"]
        dt = self.dt
        to_world = self.xf(node.get("to_world"))
        s.bsdf = self.resolve(node.get("bsdf"), "bsdf", self.bsdf)
        s.interior_medium = self.resolve(node["interior"], "medium", self.medium) if "interior" in node else -1
        s.exterior_medium = self.resolve(node["exterior"], "medium", self.medium) if "exterior" in node else -1
        if "emitter" in node:
            raise RuntimeError("area emitters are not supported by the B200 volpath")
        if "sensor" in node:  # a sensor attached to the shape (shape.cpp:47-54): irradiancemeter on a rectangle
            if node["sensor"].get("type") != "irradiancemeter" or t != "rectangle":
                raise RuntimeError("only an irradiancemeter on a rectangle is supported as a shape sensor")
            if self.shape_sensor is not None:
                raise RuntimeError("the B200 volpath renders one sensor per scene")
            self.shape_sensor = (node["sensor"], len(self.shapes))
        if t == "rectangle":  # src/shapes/rectangle.cpp:75-104
            s.type = abi.SHAPE_RECTANGLE
            if node.get("flip_normals", False):
                to_world = to_world @ Transform.scale([1, 1, -1], self.dtype)
            dp_du, dp_dv = to_world.vector([2, 0, 0]), to_world.vector([0, 2, 0])
            n = to_world.normal([0, 0, 1])
            n = (n / np.sqrt(np.dot(n, n))).astype(self.dtype)
            s.frame_s[:] = [float(x) for x in dp_du]
            s.frame_t[:] = [float(x) for x in dp_dv]
            s.frame_n[:] = [float(x) for x in n]
            area = np.sqrt(np.sum(np.cross(dp_du, dp_dv).astype(self.dtype) ** 2, dtype=self.dtype))
            s.inv_surface_area = float(dt(1) / dt(area))
            for c in ([-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0]):
                self.expand(to_world.point_affine(c))
        elif t == "sphere":  # src/shapes/sphere.cpp:95-160
            s.type = abi.SHAPE_SPHERE
            to_world = to_world @ Transform.translate(self.prop3(node.get("center", [0, 0, 0])), self.dtype)
            to_world = to_world @ Transform.scale(self.prop(node.get("radius", 1.0)), self.dtype)
            m = to_world.matrix.astype(np.float64)
            # polar decomposition of the upper 3x3 (transform_decompose): S symmetric, Q rotation
            u, sv, vt = np.linalg.svd(m[:3, :3])
            q = u @ vt
            smat = vt.T @ np.diag(sv) @ vt
            if np.linalg.det(q) < 0:
                q, smat = -q, -smat
            radius = dt(smat[0, 0])
            flip = bool(node.get("flip_normals", False))
            if radius <= 0:
                radius, flip = dt(abs(radius)), not flip
            center = m[:3, 3].astype(self.dtype)
            m2 = np.eye(4)
            m2[:3, :3] = q * float(radius)
            m2[:3, 3] = center
            to_world = Transform(m2, dtype=self.dtype)
            s.center[:] = [float(x) for x in center]
            s.radius = float(radius)
            s.flip_normals = int(flip)
            s.inv_surface_area = float(dt(1) / (dt(4) * dt(np.pi) * radius * radius))
            self.expand(center - radius)
            self.expand(center + radius)
        elif t == "cube":  # src/shapes/cube.cpp:54-116
            verts = np.array([[1, -1, -1], [1, -1, 1], [-1, -1, 1], [-1, -1, -1], [1, 1, -1], [-1, 1, -1], [-1, 1, 1],
                              [1, 1, 1], [1, -1, -1], [1, 1, -1], [1, 1, 1], [1, -1, 1], [1, -1, 1], [1, 1, 1],
                              [-1, 1, 1], [-1, -1, 1], [-1, -1, 1], [-1, 1, 1], [-1, 1, -1], [-1, -1, -1], [1, 1, -1],
                              [1, -1, -1], [-1, -1, -1], [-1, 1, -1]], dtype=np.float32)
            nrms = np.repeat(np.array([[0, -1, 0], [0, 1, 0], [1, 0, 0], [0, 0, 1], [-1, 0, 0], [0, 0, -1]],
                                      dtype=np.float32), 4, axis=0)
            uvs = np.tile(np.array([[0, 1], [1, 1], [1, 0], [0, 0]], dtype=np.float32), (6, 1))
            tris = np.array([[0, 1, 2], [3, 0, 2], [4, 5, 6], [7, 4, 6], [8, 9, 10], [11, 8, 10], [12, 13, 14],
                             [15, 12, 14], [16, 17, 18], [19, 16, 18], [20, 21, 22], [23, 20, 22]], dtype=np.uint32)
            # Mesh buffers are always float32 (mesh.h:22-28): transform in float32
            # cube.cpp:104-107: transformed and normalised in the variant's type, stored as float32
            p = to_world.point_affine(verts.astype(self.dtype)).astype(np.float32)
            n = to_world.normal(nrms.astype(self.dtype))
            n = (n / np.sqrt(np.sum(n * n, axis=1, dtype=self.dtype))[:, None]).astype(np.float32)
            if node.get("flip_normals", False):
                n = -n
            self.add_mesh(s, p, tris, n, uvs)
        elif t in ("obj", "ply", "serialized"):  # src/shapes/obj.cpp, ply.cpp, serialized.cpp
            if t == "serialized":
                pos, faces, nrm, uv = read_serialized(node["filename"], int(node.get("shape_index", 0)))
            elif t == "obj":
                pos, faces, nrm, uv = read_obj(node["filename"], bool(node.get("flip_tex_coords", True)))
            else:
                pos, faces, nrm, uv = read_ply(node["filename"])
            face_normals = bool(node.get("face_normals", False))  # mesh.cpp:30-31: no vertex normals at all
            if nrm is None and not face_normals:
                raise RuntimeError("mesh files without normals need face_normals=true (smooth vertex normals, "
                                   "Mesh::recompute_vertex_normals, are not built in the B200 volpath)")
            if face_normals:
                nrm = None
            pos = to_world.point_affine(pos.astype(self.dtype)).astype(np.float32)
            if nrm is not None:
                nrm = to_world.normal(nrm.astype(self.dtype))
                nrm = (nrm / np.sqrt(np.sum(nrm * nrm, axis=1, dtype=self.dtype))[:, None]).astype(np.float32)
            self.add_mesh(s, pos, faces, nrm, uv)
        elif t == "mesh":  # in-memory triangle mesh (what obj/ply/serialized loaders produce, src/librender/mesh.cpp)
            # obj.cpp:172,186: float32 input, transformed in the variant's type, stored as float32
            pos = np.asarray(node["vertex_positions"], dtype=np.float32).reshape(-1, 3)
            pos = to_world.point_affine(pos.astype(self.dtype)).astype(np.float32)
            nrm = node.get("vertex_normals")
            if nrm is not None:
                nrm = to_world.normal(np.asarray(nrm, dtype=np.float32).reshape(-1, 3).astype(self.dtype))
                nrm = (nrm / np.sqrt(np.sum(nrm * nrm, axis=1, dtype=self.dtype))[:, None]).astype(np.float32)
            self.add_mesh(s, pos, node["faces"], nrm, node.get("vertex_texcoords"))
        else:
            raise RuntimeError("shape plugin %r is not supported by the B200 volpath" % t)
        s.to_world = self.mat(to_world)
        s.to_object = self.mat(to_world.inverse())
        self.shapes.append(s)
        idx = len(self.shapes) - 1
        if "id" in node:
            self.ids[node["id"]] = ("shape", idx)
        return idx

    def emitter(self, node):
        """src/emitters/directional.cpp:45-73"""
        if node["type"] == "constant":  # src/emitters/constant.cpp:35-40
            e = abi.Emitter()
            e.type = abi.EMITTER_CONSTANT
            rad = node.get("radiance", 1.0)
            if isinstance(rad, dict):
                rad = rad.get("value", 1.0)
            e.irradiance = float(self.prop(rad))
            self.emitters.append(e)
            return
        if node["type"] != "directional":
            raise RuntimeError("emitter plugin %r is not supported by the B200 volpath" % node["type"])
        e = abi.Emitter()
        e.type = abi.EMITTER_DIRECTIONAL
        if "direction" in node:
            if "to_world" in node:
                raise RuntimeError("Only one of the parameters 'direction' and 'to_world' can be specified at the same time!")
            d = np.asarray(node["direction"], dtype=np.float32)  # normalize(props.vector3f(..)): single precision
            d = (d / np.sqrt(np.dot(d, d))).astype(np.float32).astype(self.dtype)
            up, _ = coordinate_system(d, self.dtype)
            tw = Transform.look_at([0, 0, 0], d, up, self.dtype)
        else:
            tw = self.xf(node.get("to_world"))
        # Endpoint::m_world_transform is an AnimatedTransform: single precision in every variant (transform.h:366)
        tw = tw.astype(np.float32).astype(self.dtype)
        e.direction[:] = [float(x) for x in tw.vector([0, 0, 1])]
        irr = node.get("irradiance", 1.0)
        if isinstance(irr, dict):
            irr = irr.get("value", 1.0)
        e.irradiance = float(self.prop(irr))
        self.emitters.append(e)

    def parse_sensor(self, node):
        """src/sensors/{distant,mdistant,radiancemeter,mradiancemeter}.cpp ctors, src/librender/sensor.cpp:15-55,
        src/librender/film.cpp, src/films/hdrfilm.cpp:100-183, src/samplers/independent.cpp"""
        sn = abi.Sensor()
        t = node["type"]
        dtp, dt = self.dtype, self.dt
        film = node.get("film", {"type": "hdrfilm"})
        if film.get("type", "hdrfilm") != "hdrfilm":
            raise RuntimeError("film plugin %r is not supported by the B200 volpath" % film["type"])
        rf = rfilter_table(film.get("rfilter", {"type": "box"}), dtp)
        sn.rfilter_radius, sn.rfilter_scale = rf["radius"], rf["scale"]
        sn.rfilter_taps, sn.rfilter_border = rf["taps"], rf["border"]
        sn.rfilter_values[:] = rf["values"]
        w, h = int(film.get("width", 768)), int(film.get("height", 576))
        sn.film_size[:] = [w, h]
        sn.crop_offset[:] = [int(film.get("crop_offset_x", 0)), int(film.get("crop_offset_y", 0))]
        sn.crop_size[:] = [int(film.get("crop_width", w)), int(film.get("crop_height", h))]
        # Film::set_crop_window (src/librender/film.cpp:54-63): the crop size does not adjust itself to the offset
        for k, full in enumerate((w, h)):
            if sn.crop_offset[k] < 0 or sn.crop_size[k] <= 0 or sn.crop_offset[k] + sn.crop_size[k] > full:
                raise RuntimeError("Invalid crop window specification!\noffset %s + crop size %s vs full size %s" % (
                    list(sn.crop_offset), list(sn.crop_size), [w, h]))
        # HDRFilm (src/films/hdrfilm.cpp:100-155): the develop() parameters are validated at construction
        if film.get("file_format", "openexr") not in ("openexr", "exr", "rgbe", "pfm"):
            raise RuntimeError("The \"file_format\" parameter must either be equal to \"openexr\", \"pfm\", or \"rgbe\","
                               " found %s instead." % film["file_format"])
        if film.get("pixel_format", "luminance") not in ("luminance", "luminance_alpha", "rgb", "rgba", "xyz", "xyza"):
            raise RuntimeError("The \"pixel_format\" parameter must either be equal to \"luminance\", \"luminance_alpha\", "
                               "\"rgb\", \"rgba\",  \"xyz\" or \"xyza\". Found %s." % film["pixel_format"])
        if film.get("component_format", "float16") not in ("float16", "float32", "uint32"):
            raise RuntimeError("The \"component_format\" parameter must either be equal to \"float16\", \"float32\", or "
                               "\"uint32\". Found %s instead." % film["component_format"])
        sampler = node.get("sampler", {"type": "independent", "sample_count": 4})
        if sampler.get("type", "independent") != "independent":
            raise RuntimeError("sampler plugin %r is not supported by the B200 volpath" % sampler["type"])
        sn.sample_count = int(sampler.get("sample_count", 4))
        sn.sampler_seed = int(sampler.get("seed", 0))
        sn.shutter_open = float(node.get("shutter_open", 0.0))
        sn.shutter_open_time = float(node.get("shutter_close", 0.0)) - sn.shutter_open
        if sn.shutter_open_time < 0:
            raise RuntimeError("Shutter opening time must be less than or equal to the shutter closing time!")
        sn.medium = self.resolve(node["medium"], "medium", self.medium) if "medium" in node else -1
        sn.target_shape = sn.origin_shape = -1
        to_world = self.xf(node.get("to_world"))
        xfs = None

        def own_shape(sub):
            """A shape given to the sensor: a reference to a shape of the scene, or a shape of the sensor's own
            (distant.cpp:270-290) -- used by the sensor, but not a child of the scene: never intersected by paths and
            not part of the scene's bounding box."""
            if sub.get("type") == "ref":
                return self.resolve(sub, "shape", self.shape)
            box = (self.bbox_min.copy(), self.bbox_max.copy())
            idx = self.shape(sub)
            self.bbox_min, self.bbox_max = box
            self.shapes[idx].sample_only = 1
            return idx

        def origin(key):   # distant.cpp:281-290 `ray_origin`, distantflux.cpp:171-180 `origin`
            if key not in node:
                sn.origin_type = abi.ORIGIN_BSPHERE
                return
            if not isinstance(node[key], dict):
                raise RuntimeError("Invalid parameter %s, must be a Shape." % key)
            sn.origin_type = abi.ORIGIN_SHAPE
            sn.origin_shape = own_shape(node[key])
            if self.shapes[sn.origin_shape].type != abi.SHAPE_RECTANGLE:
                raise RuntimeError("sensor origin shape must be a rectangle in the B200 volpath")

        def target(key):
            tgt = node.get(key)
            if tgt is None:
                sn.target_type = abi.TARGET_NONE
            elif isinstance(tgt, dict):
                sn.target_type = abi.TARGET_SHAPE
                sn.target_shape = own_shape(tgt)
                if self.shapes[sn.target_shape].type != abi.SHAPE_RECTANGLE:
                    raise RuntimeError("sensor target shape must be a rectangle in the B200 volpath")
            else:
                sn.target_type = abi.TARGET_POINT
                sn.target_point[:] = [float(self.prop(x)) for x in tgt]

        if t == "distant":  # distant.cpp:225-290
            sn.type = abi.SENSOR_DISTANT
            sn.flip_directions = int(bool(node.get("flip_directions", False)))
            sn.direction_type = (abi.DIR_SINGLE if (w, h) == (1, 1) else
                                 abi.DIR_SAMPLE_WIDTH if h == 1 else abi.DIR_SAMPLE_ALL)
            if "direction" in node:
                if "to_world" in node:
                    raise RuntimeError("Only one of the parameters 'direction' and 'to_world' can be specified at the same time!")
                d = np.asarray(node["direction"], dtype=np.float32)  # distant.cpp:253: normalized in single precision
                d = (d / np.sqrt(np.dot(d, d))).astype(np.float32).astype(dtp)
                if "orientation" in node:
                    up = np.cross(d, self.prop3(node["orientation"])).astype(dtp)
                    up = (up / np.sqrt(np.dot(up, up))).astype(dtp)
                else:
                    _, up = coordinate_system(d, dtp)
                to_world = Transform.look_at([0, 0, 0], d, up, dtp)
            target("ray_target")
            origin("ray_origin")
            sn.needs_aperture_sample = 1
        elif t == "irradiancemeter":  # irradiancemeter.cpp:54-80
            sn.type = abi.SENSOR_IRRADIANCEMETER
            if "to_world" in node:
                raise RuntimeError("Found a 'to_world' transformation -- this is not allowed. The irradiance meter "
                                   "inherits this transformation from its parent shape.")
            if (w, h) != (1, 1):
                raise RuntimeError("This sensor only supports films of size 1x1 Pixels!")
            sn.origin_shape = node["_parent_shape"]
            sn.needs_aperture_sample = 1
        elif t == "distantflux":  # distantflux.cpp:59-176
            sn.type = abi.SENSOR_DISTANTFLUX
            target("target")
            origin("origin")
            sn.needs_aperture_sample = 1
            # m_reference_normal: props.transform("to_world") (single precision, widened) times (0, 0, 1)
            sn.reference_normal[:] = [float(x) for x in to_world.vector([0, 0, 1])]
        elif t == "mdistant":  # mdistant.cpp:146-207
            sn.type = abi.SENSOR_MDISTANT
            dirs = np.asarray(node["directions"], dtype=np.float32).reshape(-1, 3)  # std::stof
            if (w, h) != (len(dirs), 1):
                raise RuntimeError("Film size must be [sensor_count, 1]. Expected [%d, 1], got [%d, %d]" % (len(dirs), w, h))
            xfs = []
            for d in dirs.astype(dtp):
                _, up = coordinate_system(d, dtp)
                xfs.append(Transform.look_at([0, 0, 0], d, up, dtp))
            target("target")
            sn.needs_aperture_sample = 1
        elif t == "radiancemeter":  # radiancemeter.cpp:61-104
            sn.type = abi.SENSOR_RADIANCEMETER
            if "to_world" not in node and ("direction" in node) != ("origin" in node):
                raise RuntimeError("If the sensor is specified through origin and direction both values must be set!")
            if "to_world" not in node and "direction" in node:
                o = self.prop3(node["origin"])
                d = self.prop3(node["direction"])
                up, _ = coordinate_system(d, dtp)
                to_world = Transform.look_at(o, o + d, up, dtp)
            if (w, h) != (1, 1):
                raise RuntimeError("This sensor only supports films of size 1x1 Pixels!")
            sn.needs_aperture_sample = 0
        elif t == "mradiancemeter":  # mradiancemeter.cpp:75-131
            sn.type = abi.SENSOR_MRADIANCEMETER
            if "to_world" in node:
                raise RuntimeError("This sensor is specified through a set of origin and direction values and "
                                   "cannot use the to_world transform.")
            origins = np.asarray(node["origins"], dtype=np.float32).reshape(-1, 3).astype(dtp)
            dirs = np.asarray(node["directions"], dtype=np.float32).reshape(-1, 3).astype(dtp)
            if origins.shape != dirs.shape:
                raise RuntimeError("Invalid specification! Number of parameters for origins and directions are not equal.")
            if (w, h) != (len(dirs), 1):
                raise RuntimeError("Film size must be [sensor_count, 1]. Expected [%d, 1], got [%d, %d]" % (len(dirs), w, h))
            xfs = []
            for o, d in zip(origins, dirs):
                up, _ = coordinate_system(d, dtp)
                xfs.append(Transform.look_at(o, o + d, up, dtp))
            sn.needs_aperture_sample = 0
        else:
            raise RuntimeError("sensor plugin %r is not supported by the B200 volpath" % t)
        # distant / radiancemeter keep to_world in Endpoint::m_world_transform (AnimatedTransform, single precision
        # in every variant, transform.h:366); mdistant / mradiancemeter store ScalarFloat matrices of their own
        to_world = to_world.astype(np.float32).astype(self.dtype)
        sn.to_world = self.mat(to_world)
        if xfs is not None:
            arr = np.ascontiguousarray(np.stack([x.matrix.T.reshape(-1) for x in xfs]), dtype=np.float64)
            self.keep.append(arr)
            sn.n_transforms = len(xfs)
            sn.transforms = arr.ctypes.data_as(C.POINTER(C.c_double))
        self.sensor = sn

    def parse_integrator(self, node):
        """src/librender/integrator.cpp:23-39,302-315"""
        outer = node
        moment = 0
        if node.get("type", "volpath") == "moment":  # src/integrators/moment.cpp:39-60: one nested volpath
            nested = [v for v in node.values() if isinstance(v, dict)]
            if len(nested) != 1 or nested[0].get("type", "volpath") not in ("volpath", "volpathmis"):
                raise RuntimeError("the B200 moment integrator nests exactly one volpath / volpathmis")
            node, moment = nested[0], 1
        if node.get("type", "volpath") not in ("volpath", "volpathmis"):
            raise RuntimeError("integrator plugin %r is not supported by the B200 volpath" % node["type"])
        it = abi.Integrator()
        it.moment = moment
        if node.get("type") == "volpathmis":  # volpathmis.cpp:35-52: use_spectral_mis selects the implementation
            it.mis = 1 if node.get("use_spectral_mis", True) else 2
        it.max_depth = int(node.get("max_depth", -1))
        if it.max_depth < 0 and it.max_depth != -1:
            raise RuntimeError("\"max_depth\" must be set to -1 (infinite) or a value >= 0")
        it.rr_depth = int(node.get("rr_depth", 5))
        if it.rr_depth <= 0:
            raise RuntimeError("\"rr_depth\" must be set to a value greater than zero!")
        it.hide_emitters = int(bool(node.get("hide_emitters", False)))
        # block_size / samples_per_pass are read by the SamplingIntegrator that render() runs on: the outer one
        it.block_size = int(outer.get("block_size", 0))
        it.samples_per_pass = int(outer.get("samples_per_pass", 0xffffffff)) & 0xffffffff
        it.arg_order_ltr = int(node.get("arg_order_ltr", 0))
        self.integrator = it


class Scene:
    """A loaded scene: owns the B2vpSceneDesc and every array it points to."""

    def __init__(self, desc, variant, precision, keep):
        self.desc, self.variant, self.precision, self._keep = desc, variant, precision, keep
        self.film_size = (desc.sensor.crop_size[0], desc.sensor.crop_size[1])
        self.film_channels = 11 if desc.integrator.moment else 5   # B2VP_FILM_CHANNELS(_MOMENT)


def load_dict(d, variant="scalar_mono"):
    """Mirror of mitsuba.core.xml.load_dict for the scoped plugin set."""
    if d.get("type", "scene") != "scene":
        raise RuntimeError("load_dict: the top-level dictionary must be a scene")
    b = _Builder(variant)
    sensor_node = integrator_node = None
    emitters = []
    shape_kinds = ("rectangle", "sphere", "cube", "mesh", "obj", "ply", "serialized")
    # children are visited in key order, as Properties::objects() does (std::map)
    items = [(k, v) for k, v in sorted(d.items()) if isinstance(v, dict)]
    # first pass: named media / bsdfs so that references resolve regardless of order
    for k, v in items:
        v = dict(v)
        v.setdefault("id", k)
        if v["type"] in ("homogeneous", "heterogeneous"):
            b.medium(v)
        elif v["type"] in ("diffuse", "rpv", "null", "bilambertian"):
            b.bsdf(v)
    for k, v in items:
        t = v["type"]
        if t in shape_kinds:
            v = dict(v)
            v.setdefault("id", k)
            b.shape(v)
        elif t in ("directional", "constant"):
            emitters.append(v)
        elif t in ("distant", "mdistant", "radiancemeter", "mradiancemeter", "distantflux"):
            if sensor_node is not None:
                raise RuntimeError("the B200 volpath renders one sensor per scene")
            sensor_node = v
        elif t in ("volpath", "volpathmis", "moment"):
            integrator_node = v
        elif t in ("homogeneous", "heterogeneous", "diffuse", "rpv", "null", "bilambertian"):
            pass
        else:
            raise RuntimeError("plugin %r is not supported by the B200 volpath" % t)
    for e in emitters:
        b.emitter(e)
    if b.shape_sensor is not None:
        if sensor_node is not None:
            raise RuntimeError("the B200 volpath renders one sensor per scene")
        sensor_node = dict(b.shape_sensor[0], _parent_shape=b.shape_sensor[1])
    if sensor_node is None:
        raise RuntimeError("scene has no sensor")
    b.parse_sensor(sensor_node)
    b.parse_integrator(integrator_node or {"type": "volpath"})

    desc = abi.SceneDesc()
    desc.abi_version, desc.struct_size, desc.precision = abi.ABI_VERSION, C.sizeof(abi.SceneDesc), b.precision
    keep = [b.keep, b.volume_arrays]

    def arr(cls, items):
        a = (cls * max(len(items), 1))(*items)
        keep.append(a)
        return a
    desc.n_shapes, desc.shapes = len(b.shapes), arr(abi.Shape, b.shapes)
    desc.n_bsdfs, desc.bsdfs = len(b.bsdfs), arr(abi.Bsdf, b.bsdfs)
    desc.n_media, desc.media = len(b.media), arr(abi.Medium, b.media)
    desc.n_volumes, desc.volumes = len(b.volumes), arr(abi.Volume, b.volumes)
    desc.n_phases, desc.phases = len(b.phases), arr(abi.Phase, b.phases)
    desc.n_emitters, desc.emitters = len(b.emitters), arr(abi.Emitter, b.emitters)
    if b.n_vertices:
        vpos, vnrm, vuv = np.concatenate(b.vpos), np.concatenate(b.vnrm), np.concatenate(b.vuv)
        faces = np.concatenate(b.faces)
        keep += [vpos, vnrm, vuv, faces]
        desc.n_vertices, desc.n_faces = b.n_vertices, b.n_faces
        desc.vertex_positions = vpos.ctypes.data_as(C.POINTER(C.c_float))
        desc.vertex_normals = vnrm.ctypes.data_as(C.POINTER(C.c_float))
        desc.vertex_texcoords = vuv.ctypes.data_as(C.POINTER(C.c_float))
        desc.faces = faces.ctypes.data_as(C.POINTER(C.c_uint32))
    dt = b.dt
    if not np.all(np.isfinite(b.bbox_min)):
        b.bbox_min[:] = 0
        b.bbox_max[:] = 0
    desc.bbox_min[:] = [float(x) for x in b.bbox_min]
    desc.bbox_max[:] = [float(x) for x in b.bbox_max]
    # BoundingBox::bounding_sphere (bbox.h:329-332) + the padding of distant.cpp:292-297 / directional.cpp:68-73
    center = ((b.bbox_max + b.bbox_min) * dt(0.5)).astype(b.dtype)
    diff = (center - b.bbox_max).astype(b.dtype)
    radius = dt(np.sqrt(dt(diff[0] * diff[0] + diff[1] * diff[1]) + diff[2] * diff[2]))
    ray_eps = dt(np.finfo(b.dtype).eps) / dt(2) * dt(1500)
    radius = max(ray_eps, dt(radius * (dt(1) + ray_eps)))
    desc.bsphere_center[:] = [float(x) for x in center]
    desc.bsphere_radius = float(radius)
    desc.sensor = b.sensor
    desc.integrator = b.integrator
    return Scene(desc, variant, b.precision, keep)
