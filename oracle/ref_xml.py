"""TEST INFRASTRUCTURE: runs the UNMODIFIED reference (oracle/_ref, built by oracle/Makefile.ref) on the same
scene dictionaries the product's load_dict takes.

`to_xml` writes the dictionary as a reference XML scene (src/libcore/xml.cpp) plus the .vol / .obj files it
refers to; `render` runs oracle/_ref/refprobe on it and returns the raw film storage (X, Y, Z, A, W) and the
wall time of Integrator::render.  `round_f32` rounds every number of a dictionary to float32: xml.cpp parses
all numbers with std::stof (xml.cpp:100-103) even in the double variant, so a dictionary that is to give
identical parameters on both sides must hold float32-representable values."""
import os
import struct
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
PROBE = os.path.join(REF_DIR, "refprobe")
# the timing build (Makefile.ref `release`): the reference's Release flags; never used for parity
RELEASE_PROBE = os.path.join(REF_DIR, "release", "refprobe")
RELEASE_FLAGS = "-O3 -march=x86-64-v3 -ffp-contract=fast -fno-math-errno (grid3d.cpp at -O1)"
PARITY_FLAGS = "-O3 -march=x86-64-v3 -ffp-contract=off (grid3d.cpp at -O1)"
REFERENCE_SRC = "/root/reference"


def available(build="parity"):
    return os.path.exists(RELEASE_PROBE if build == "release" else PROBE)


def build():
    """Compiles the reference where it lies (only possible where /root/reference exists)."""
    if not os.path.isdir(REFERENCE_SRC):
        return False
    subprocess.check_call(["make", "-s", "-f", "Makefile.ref", "-C", HERE, "-j", str(os.cpu_count() or 4)])
    subprocess.check_call(["make", "-s", "-f", "Makefile.ref", "-C", HERE, "-j", str(os.cpu_count() or 4), "release"])
    return True


def round_f32(node):
    if isinstance(node, dict):
        return {k: round_f32(v) for k, v in node.items()}
    if isinstance(node, (list, tuple)) and node and all(isinstance(x, (int, float)) for x in node):
        return [float(np.float32(x)) if isinstance(x, float) else x for x in node]
    if isinstance(node, bool) or isinstance(node, int):
        return node
    if isinstance(node, float):
        return float(np.float32(node))
    if hasattr(node, "matrix") and hasattr(node, "astype"):  # Transform
        return node.astype(np.float32)
    return node


def _num(x):
    return repr(float(x))


def _matrix(t):
    m = np.asarray(t.matrix if hasattr(t, "matrix") else t, dtype=np.float64).reshape(4, 4)
    return '<transform name="to_world"><matrix value="%s"/></transform>' % " ".join(_num(v) for v in m.reshape(-1))


def _xyz(tag, name, v):
    return '<%s name="%s" x="%s" y="%s" z="%s"/>' % (tag, name, _num(v[0]), _num(v[1]), _num(v[2]))


def _write_obj(path, positions, faces, normals=None):
    """Wavefront OBJ (src/shapes/obj.cpp parses it with strtof: float32 values written with repr round-trip).
    The .ply loader is not used: it converts records through the asmjit struct converter, which is not built."""
    positions = np.asarray(positions, dtype=np.float32).reshape(-1, 3)
    faces = np.asarray(faces, dtype=np.int64).reshape(-1, 3) + 1
    with open(path, "w") as f:
        for p in positions:
            f.write("v %s %s %s\n" % (_num(p[0]), _num(p[1]), _num(p[2])))
        if normals is not None:
            for n in np.asarray(normals, dtype=np.float32).reshape(-1, 3):
                f.write("vn %s %s %s\n" % (_num(n[0]), _num(n[1]), _num(n[2])))
            for a_, b_, c_ in faces:
                f.write("f %d//%d %d//%d %d//%d\n" % (a_, a_, b_, b_, c_, c_))
        else:
            for a_, b_, c_ in faces:
                f.write("f %d %d %d\n" % (a_, b_, c_))


class _Writer:
    def __init__(self, workdir, write_vol):
        self.dir, self.write_vol, self.n_files = workdir, write_vol, 0

    def scalar_or_texture(self, name, v):
        if isinstance(v, dict):  # {"type": "uniform", "value": x}
            return '<spectrum name="%s" value="%s"/>' % (name, _num(v.get("value", 1.0)))
        return '<float name="%s" value="%s"/>' % (name, _num(v))

    def volume(self, name, v):
        if isinstance(v, (int, float)):
            return '<float name="%s" value="%s"/>' % (name, _num(v))
        out = ['<volume type="%s" name="%s">' % (v["type"], name)]
        if v["type"] == "gridvolume":
            if "filename" in v:
                path = v["filename"]
            else:
                path = os.path.join(self.dir, "grid%d.vol" % self.n_files)
                self.n_files += 1
                self.write_vol(path, v["data"], v.get("bbox", (0, 0, 0, 1, 1, 1)))
            out.append('<string name="filename" value="%s"/>' % path)
            for key in ("filter_type", "wrap_mode"):
                if key in v:
                    out.append('<string name="%s" value="%s"/>' % (key, v[key]))
            for key in ("use_grid_bbox", "raw"):
                if key in v:
                    out.append('<boolean name="%s" value="%s"/>' % (key, "true" if v[key] else "false"))
            if "max_value" in v:
                out.append('<float name="max_value" value="%s"/>' % _num(v["max_value"]))
        else:
            out.append('<float name="color" value="%s"/>' % _num(v.get("color", 1.0)))   # constant3d.cpp:14
        if "to_world" in v:
            out.append(_matrix(v["to_world"]))
        out.append("</volume>")
        return "".join(out)

    def phase(self, v, name=None):
        out = '<phase type="%s"%s>' % (v["type"], ' name="%s"' % name if name else "")
        if "g" in v:
            out += '<float name="g" value="%s"/>' % _num(v["g"])
        if "values" in v:
            vals = v["values"]
            if not isinstance(vals, str):
                vals = ", ".join(repr(float(x)) for x in vals)
            out += '<string name="values" value="%s"/>' % vals
        if v["type"] == "blendphase":   # nested phase functions are named so that they sort like the dictionary keys
            for k, child in sorted(v.items()):
                if isinstance(child, dict) and k != "weight":
                    out += self.phase(child, name=k)
            out += self.volume("weight", v["weight"])
        return out + "</phase>"

    def medium(self, v, ident=None, name=None):
        out = ['<medium type="%s"%s%s>' % (v["type"], ' id="%s"' % ident if ident else "",
                                           ' name="%s"' % name if name else "")]
        for key in ("sigma_t", "albedo"):
            if key in v:
                out.append(self.volume(key, v[key]))
        if "scale" in v:
            out.append('<float name="scale" value="%s"/>' % _num(v["scale"]))
        for key in ("has_spectral_extinction", "sample_emitters"):
            if key in v:
                out.append('<boolean name="%s" value="%s"/>' % (key, "true" if v[key] else "false"))
        if "phase" in v:
            out.append(self.phase(v["phase"]))
        out.append("</medium>")
        return "".join(out)

    def bsdf(self, v, ident=None):
        out = ['<bsdf type="%s"%s>' % (v["type"], ' id="%s"' % ident if ident else "")]
        for key in ("reflectance", "transmittance", "rho_0", "k", "g", "rho_c"):
            if key in v:
                out.append(self.scalar_or_texture(key, v[key]))
        out.append("</bsdf>")
        return "".join(out)

    def child(self, name, v, kind):
        if v.get("type") == "ref":
            return '<ref name="%s" id="%s"/>' % (name, v["id"])
        return self.medium(v, name=name) if kind == "medium" else self.bsdf(v)

    def shape(self, v, ident):
        t = v["type"]
        out = []
        if t in ("obj", "ply", "serialized"):
            out.append('<shape type="%s" id="%s"><string name="filename" value="%s"/>' % (t, ident, v["filename"]))
            if "shape_index" in v:
                out.append('<integer name="shape_index" value="%d"/>' % v["shape_index"])
            if "face_normals" in v:
                out.append('<boolean name="face_normals" value="%s"/>' % ("true" if v["face_normals"] else "false"))
        elif t == "mesh":
            path = os.path.join(self.dir, "mesh%d.obj" % self.n_files)
            self.n_files += 1
            _write_obj(path, v["vertex_positions"], v["faces"], v.get("vertex_normals"))
            out.append('<shape type="obj" id="%s"><string name="filename" value="%s"/>' % (ident, path))
            if v.get("vertex_normals") is None:
                out.append('<boolean name="face_normals" value="true"/>')
        else:
            out.append('<shape type="%s" id="%s">' % (t, ident))
        if "to_world" in v:
            out.append(_matrix(v["to_world"]))
        if "flip_normals" in v:
            out.append('<boolean name="flip_normals" value="%s"/>' % ("true" if v["flip_normals"] else "false"))
        if "center" in v:
            out.append(_xyz("point", "center", v["center"]))
        if "radius" in v:
            out.append('<float name="radius" value="%s"/>' % _num(v["radius"]))
        if "sensor" in v:
            out.append(self.sensor(v["sensor"]))
        if "bsdf" in v:
            out.append(self.child("bsdf", v["bsdf"], "bsdf"))
        for key in ("interior", "exterior"):
            if key in v:
                out.append(self.child(key, v[key], "medium"))
        out.append("</shape>")
        return "".join(out)

    def emitter(self, v):
        out = ['<emitter type="%s">' % v["type"]]
        if "direction" in v:
            out.append(_xyz("vector", "direction", v["direction"]))
        if "to_world" in v:
            out.append(_matrix(v["to_world"]))
        for key in ("irradiance", "radiance"):
            if key in v:
                out.append(self.scalar_or_texture(key, v[key]))
        out.append("</emitter>")
        return "".join(out)

    def sensor(self, v):
        out = ['<sensor type="%s">' % v["type"]]
        for key in ("direction", "orientation"):
            if key in v:
                out.append(_xyz("vector", key, v[key]))
        for key in ("origin",):
            if key in v and not isinstance(v[key], dict):
                out.append(_xyz("point", key, v[key]))
        for key in ("ray_target", "target", "ray_origin", "origin"):
            if key in v and (isinstance(v[key], dict) or key in ("ray_target", "target")):
                if isinstance(v[key], dict) and v[key].get("type") != "ref":   # a shape of the sensor's own, not in the scene
                    self.n_files += 1
                    out.append(self.shape(v[key], "target%d" % self.n_files).replace("<shape ", '<shape name="%s" ' % key, 1))
                elif isinstance(v[key], dict):
                    out.append('<ref name="%s" id="%s"/>' % (key, v[key]["id"]))
                else:
                    out.append(_xyz("point", key, v[key]))
        for key in ("directions", "origins"):
            if key in v:
                out.append('<string name="%s" value="%s"/>' % (key, ", ".join(_num(np.float32(x)) for x in v[key])))
        if "flip_directions" in v:
            out.append('<boolean name="flip_directions" value="%s"/>' % ("true" if v["flip_directions"] else "false"))
        for key in ("shutter_open", "shutter_close"):
            if key in v:
                out.append('<float name="%s" value="%s"/>' % (key, _num(v[key])))
        if "to_world" in v:
            out.append(_matrix(v["to_world"]))
        if "medium" in v:
            out.append(self.child("medium", v["medium"], "medium"))
        s = v.get("sampler", {"type": "independent", "sample_count": 4})
        out.append('<sampler type="%s"><integer name="sample_count" value="%d"/>' % (s.get("type", "independent"),
                                                                                    s.get("sample_count", 4)))
        if "seed" in s:
            out.append('<integer name="seed" value="%d"/>' % s["seed"])
        out.append("</sampler>")
        f = v.get("film", {"type": "hdrfilm"})
        out.append('<film type="hdrfilm">')
        for key in ("width", "height", "crop_offset_x", "crop_offset_y", "crop_width", "crop_height"):
            if key in f:
                out.append('<integer name="%s" value="%d"/>' % (key, f[key]))
        rf = f.get("rfilter", {"type": "box"})
        out.append('<rfilter type="%s">' % rf.get("type", "box"))
        for key in ("radius", "stddev", "B", "C"):
            if key in rf:
                out.append('<float name="%s" value="%r"/>' % (key, float(rf[key])))
        if "lobes" in rf:
            out.append('<integer name="lobes" value="%d"/>' % rf["lobes"])
        out.append("</rfilter>")
        out.append("</film></sensor>")
        return "".join(out)

    def integrator(self, v):
        out = ['<integrator type="%s">' % v.get("type", "volpath")]
        for key in ("max_depth", "rr_depth", "block_size", "samples_per_pass"):
            if key in v:
                out.append('<integer name="%s" value="%d"/>' % (key, v[key]))
        if "use_spectral_mis" in v:
            out.append('<boolean name="use_spectral_mis" value="%s"/>' % ("true" if v["use_spectral_mis"] else "false"))
        if v.get("moment") is True:
            out.append('<boolean name="moment" value="true"/>')
        if "algorithm" in v:
            out.append('<string name="algorithm" value="%s"/>' % v["algorithm"])
        if "seeding" in v:
            out.append('<string name="seeding" value="%s"/>' % v["seeding"])
        if "math" in v:
            out.append('<string name="math" value="%s"/>' % v["math"])
        for key in ("majorant_grid", "n_devices", "device"):
            if key in v:
                out.append('<integer name="%s" value="%d"/>' % (key, v[key]))
        if "timeout" in v:
            out.append('<float name="timeout" value="%r"/>' % float(v["timeout"]))
        if "hide_emitters" in v:
            out.append('<boolean name="hide_emitters" value="%s"/>' % ("true" if v["hide_emitters"] else "false"))
        if "nested" in v:  # moment integrator: one nested integrator
            out.append(self.integrator(v["nested"]))
        out.append("</integrator>")
        return "".join(out)


SHAPES = ("rectangle", "sphere", "cube", "mesh", "obj", "ply", "serialized")
MEDIA = ("homogeneous", "heterogeneous")
BSDFS = ("diffuse", "rpv", "null", "bilambertian")
SENSORS = ("distant", "mdistant", "radiancemeter", "mradiancemeter", "distantflux")
INTEGRATORS = ("volpath", "volpath_b200", "volpathmis", "moment", "path")


def to_xml(scene, workdir, write_vol):
    """Children are written in key order: that is the order load_dict (std::map of named children) and the
    product's loader give them, and the XML loader keeps the order of appearance (`_arg_<i>` names) for up to
    ten unnamed children."""
    w = _Writer(workdir, write_vol)
    out = ['<scene version="2.0.0">']
    items = [(k, v) for k, v in sorted(scene.items()) if isinstance(v, dict)]
    # referenced objects first: the XML loader resolves ids in document order independent of use
    for k, v in items:
        if v["type"] in MEDIA:
            out.append(w.medium(v, ident=v.get("id", k)))
        elif v["type"] in BSDFS:
            out.append(w.bsdf(v, ident=v.get("id", k)))
    for k, v in items:
        t = v["type"]
        if t in SHAPES:
            out.append(w.shape(v, v.get("id", k)))
        elif t in ("directional", "constant"):
            out.append(w.emitter(v))
        elif t in SENSORS:
            out.append(w.sensor(v))
        elif t in INTEGRATORS:
            out.append(w.integrator(v))
    out.append("</scene>")
    path = os.path.join(workdir, "scene.xml")
    with open(path, "w") as f:
        f.write("\n".join(out))
    return path


PLUGIN = os.path.join(REF_DIR, "plugins", "volpath_b200.so")
ORACLE_LIB = os.path.join(HERE, "libb2vp_oracle.so")


def with_plugin(scene, seeding="scalar", type_name="volpath_b200", **props):
    """The same scene with <integrator type="volpath_b200">: the product's plugin inside the reference binary.
    seeding="scalar" asks for the scalar variants' stream order (films comparable sample by sample with the
    reference's CPU volpath), "wavefront" for the plugin's default.  type_name="volpath" + render(plugin_root=True):
    the plugin under the reference's own name.  props: math, majorant_grid, n_devices, timeout."""
    d = dict(scene)
    for k, v in scene.items():
        if isinstance(v, dict) and v.get("type") in ("volpath", "volpathmis"):
            # under the integrator's own name (the drop-in .so of that name) the algorithm is the plugin's default
            alg = {} if type_name == v["type"] else {"algorithm": v["type"]}
            d[k] = dict(v, type=type_name, seeding=seeding, **alg, **props)
        elif isinstance(v, dict) and v.get("type") == "moment":   # moment around a volpath -> the plugin's own `moment`
            inner = [c for c in v.values() if isinstance(c, dict)][0]
            outer = {kk: vv for kk, vv in v.items() if not isinstance(vv, dict) and kk != "type"}
            # block_size / samples_per_pass belong to the integrator render() runs on, the outer one; what the nested
            # integrator says about them is never read (integrator.cpp:51-110)
            inner = {kk: vv for kk, vv in inner.items() if kk not in ("block_size", "samples_per_pass")}
            d[k] = dict(inner, type=type_name, seeding=seeding, algorithm=inner["type"], moment=True, **dict(outer, **props))
    return d


def render(scene, variant="scalar_mono", threads=1, repeats=1, write_vol=None, keep_dir=None,
           describe_oracle=False, build="parity", plugin_root=False, record=False):
    """Runs the reference binary on `scene`.  Returns (film[h, w, channels], seconds of the last Integrator::render).
    describe_oracle (scenes from with_plugin only): do not render on a GPU; run the plugin's describe() and render
    its descriptor with the CPU oracle (refprobe --describe-oracle)."""
    probe = RELEASE_PROBE if build == "release" else PROBE
    if not available(build):
        raise RuntimeError("oracle/_ref/refprobe is not built (run make -f oracle/Makefile.ref where /root/reference exists)")
    tmp = None
    if keep_dir is None:
        tmp = tempfile.TemporaryDirectory(prefix="b2vp_ref_")
        keep_dir = tmp.name
    try:
        xml = to_xml(scene, keep_dir, write_vol)
        out = os.path.join(keep_dir, "film.bin")
        env = dict(os.environ)
        rec_path = os.path.join(keep_dir, "records.bin")
        if record:   # the scene's sampler must be {"type": "recording"} (oracle/ref_recording_sampler.cpp)
            env["B2VP_RECORD_PATH"] = rec_path
        cmd = [probe, variant, xml, out, "-t", str(threads), "-r", str(repeats)]
        if describe_oracle:
            cmd += ["--describe-oracle", ORACLE_LIB]
        if plugin_root:   # <integrator type="volpath"> resolves to the product's plugin (oracle/_ref/b200/plugins/volpath.so)
            cmd += ["--plugin-root", os.path.join(REF_DIR, "b200")]
        p = subprocess.run(cmd, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True, env=env)
        if p.returncode != 0:
            raise RuntimeError("refprobe failed (%d): %s" % (p.returncode, p.stderr[-2000:]))
        with open(out, "rb") as f:
            raw = f.read()
        w, h, c, is_double = struct.unpack_from("<4i", raw, 0)
        seconds, = struct.unpack_from("<d", raw, 16)
        film = np.frombuffer(raw, dtype="<f8" if is_double else "<f4", offset=24).reshape(h, w, c).copy()
        if record:
            # (seed value, sample index within the seed, PCG32 state after the sample, draws since the seed) per advance()
            rec = np.fromfile(rec_path, dtype="<u8").reshape(-1, 4)
            return film, seconds, rec
        return film, seconds
    finally:
        if tmp is not None:
            tmp.cleanup()
