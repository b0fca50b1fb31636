"""Integrator-level entry points: the Python counterpart of `scene.integrator().render(scene, sensor)`
(src/librender/python/integrator_v.cpp:124-153) on top of the C-ABI in include/b2vp.h."""
import ctypes as C

import numpy as np

from . import abi


# Defaults of DeviceScene.render's `kernel` and `math` arguments.  "auto" + "exact" is the bit-exact product path;
# the parity tests patch "kernel" to run every test on every device kernel, bench.py asks for math="fast".
render_defaults = {"kernel": "auto", "math": "exact"}


def _set_launch_options(a, kernel, math, tuning, debug, majorant_grid=0):
    kernel = render_defaults["kernel"] if kernel is None else kernel
    math = render_defaults["math"] if math is None else math
    a.kernel = abi.KERNELS[kernel] if isinstance(kernel, str) else int(kernel)
    a.math_mode = {"exact": abi.MATH_EXACT, "fast": abi.MATH_FAST}[math] if isinstance(math, str) else int(math)
    if tuning:   # dict: slots_per_thread, stay, iter, cta_phase (B2vpRenderArgs::tuning)
        a.tuning[0] = int(round(16 * tuning.get("slots_per_thread", 0)))
        a.tuning[1] = int(tuning.get("stay", 0))
        a.tuning[2] = int(tuning.get("iter", 0))
        a.tuning[3] = int(tuning["cta_phase"]) + 1 if "cta_phase" in tuning else 0
    a.debug_flags = int(debug)
    a.majorant_grid = int(majorant_grid)


class RenderResult:
    def __init__(self, film, totals, state=None, counters=None, radiance=None):
        self.film = film            # (H, W, 5) XYZAW sums, like the hdrfilm storage ImageBlock
        self.totals = totals        # abi.Counters
        self.state, self.counters, self.radiance = state, counters, radiance

    def bitmap(self):
        """HDRFilm::bitmap for pixel_format=luminance: Y / W (src/films/hdrfilm.cpp:251-321)."""
        w = self.film[..., 4]
        with np.errstate(divide="ignore", invalid="ignore"):
            return np.where(w > 0, self.film[..., 1] / w, 0).astype(self.film.dtype)

    def write_pfm(self, path):
        """film->develop() for file_format=pfm, pixel_format=luminance, component_format=float32 (the output the
        reference CLI is driven with): `Pf`, size, `-1` (little endian), then the scanlines
        bottom to top (Bitmap::write_pfm, src/libcore/bitmap.cpp:2206-2230)."""
        img = np.ascontiguousarray(self.bitmap(), dtype="<f4")
        h, w = img.shape
        with open(path, "wb") as f:
            f.write(("Pf\n%u %u\n-1\n" % (w, h)).encode())
            f.write(img[::-1].tobytes())

    def write_exr(self, path, component_format="float16", pixel_format="luminance"):
        """film->develop() for file_format=openexr (the hdrfilm default, src/films/hdrfilm.cpp:251-343 ->
        Bitmap::write_openexr, src/libcore/bitmap.cpp): a scan-line OpenEXR 2 file, channels in alphabetical order,
        increasing y, data and display window = the crop window.  pixel_format "luminance" writes the one channel `Y`
        (= Y / W, what mono variants develop), "xyzaw" the raw accumulation buffer (channels A, W, X, Y, Z) as
        `develop(raw=true)` would keep it.  component_format float16 (hdrfilm's default) | float32.  Written without
        compression (the reference's default is PIZ; compression is a container detail, the pixels are the same) by
        this file alone: no OpenEXR library is needed."""
        import struct
        ctype, dtype = {"float16": (1, "<f2"), "float32": (2, "<f4")}[component_format]
        if pixel_format == "luminance":
            planes = {"Y": self.bitmap()}
        elif pixel_format == "xyzaw":
            planes = {n: self.film[..., i] for i, n in enumerate("XYZAW")}
        else:
            raise ValueError("pixel_format must be 'luminance' or 'xyzaw'")
        names = sorted(planes)
        h, w = planes[names[0]].shape
        rows = np.stack([np.ascontiguousarray(planes[n], dtype=dtype) for n in names], axis=1)   # (h, channels, w)

        def attr(name, typ, payload):
            return name.encode() + b"\0" + typ.encode() + b"\0" + struct.pack("<i", len(payload)) + payload
        chlist = b"".join(n.encode() + b"\0" + struct.pack("<iB3xii", ctype, 0, 1, 1) for n in names) + b"\0"
        box = struct.pack("<4i", 0, 0, w - 1, h - 1)
        header = (struct.pack("<II", 20000630, 2)
                  + attr("channels", "chlist", chlist)
                  + attr("compression", "compression", b"\0")
                  + attr("dataWindow", "box2i", box) + attr("displayWindow", "box2i", box)
                  + attr("lineOrder", "lineOrder", b"\0")
                  + attr("pixelAspectRatio", "float", struct.pack("<f", 1.0))
                  + attr("screenWindowCenter", "v2f", struct.pack("<2f", 0.0, 0.0))
                  + attr("screenWindowWidth", "float", struct.pack("<f", 1.0))
                  + b"\0")
        row_bytes = rows.shape[1] * w * rows.dtype.itemsize
        first = len(header) + 8 * h
        with open(path, "wb") as f:
            f.write(header)
            f.write(np.arange(h, dtype="<u8").__mul__(row_bytes + 8).__add__(first).tobytes())   # scan-line offset table
            for y in range(h):
                f.write(struct.pack("<ii", y, row_bytes))
                f.write(rows[y].tobytes())

    @property
    def segments(self):
        return self.totals.n_medium_steps + self.totals.n_intersect


class DeviceScene:
    """A scene resident on one CUDA device (b2vp_scene_create .. b2vp_scene_destroy)."""

    def __init__(self, scene, device=0):
        self.scene = scene
        self.device = device
        self._h = C.c_void_p()
        abi.check(abi.lib().b2vp_scene_create(C.byref(scene.desc), device, C.byref(self._h)))

    def close(self):
        if self._h:
            abi.lib().b2vp_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def stream_count(self, seed_mode):
        n, per = C.c_uint64(), C.c_uint32()
        abi.check(abi.lib().b2vp_stream_count(self._h, seed_mode, C.byref(n), C.byref(per)))
        return n.value, per.value

    def update(self, medium=-1, medium_scale=0.0, bsdf=-1, bsdf_p=(0, 0, 0, 0), emitter=-1, emitter_irradiance=0.0,
               volume=-1, volume_value=0.0):
        p = abi.ParamPatch()
        p.abi_version, p.struct_size = abi.ABI_VERSION, C.sizeof(abi.ParamPatch)
        p.medium, p.medium_scale, p.bsdf, p.emitter, p.emitter_irradiance = medium, medium_scale, bsdf, emitter, emitter_irradiance
        p.bsdf_p[:] = list(bsdf_p)
        p.volume, p.volume_value = volume, volume_value
        abi.check(abi.lib().b2vp_scene_update(self._h, C.byref(p)))

    def update_grid(self, volume, data, keep_max=False):
        """New values data[z, y, x(, c)] for grid volume `volume` (index into Scene.volumes) of unchanged resolution:
        what the reference does through traverse() -> "data" -> GridVolume::parameters_changed
        (src/textures/grid3d.cpp:366-391).  The grid is overwritten on the device, nothing else is re-uploaded."""
        arr = np.ascontiguousarray(data, dtype=np.float32).ravel()
        abi.check(abi.lib().b2vp_scene_update_grid(self._h, int(volume), arr.ctypes.data_as(C.POINTER(C.c_float)),
                                                   arr.size, int(bool(keep_max))))

    def cancel(self):
        abi.lib().b2vp_cancel(self._h)

    def info(self):
        """b2vp_scene_info: BVH figures and the kernel instance the last render launched."""
        i = abi.SceneInfo()
        i.abi_version, i.struct_size = abi.ABI_VERSION, C.sizeof(abi.SceneInfo)
        abi.check(abi.lib().b2vp_scene_info(self._h, C.byref(i)))
        return i

    def render(self, seed_mode=abi.SEED_WAVEFRONT, precision=None, first_stream=0, n_streams=0, explicit=None, interleave=None,
               want_paths=False, film=None, stream=None, device_film=None, clear=True, kernel=None, math=None,
               tuning=None, debug=0, majorant_grid=0, timeout=0.0, devices=None):
        """Renders a stream range.  `explicit` = (state u64[], inc u64[], pixel u32[], samples_per_stream).
        `film`: optional (H, W, 5) array to accumulate into (otherwise a fresh zeroed film).
        `kernel`: "auto" | "generic" | "q256" | "lanes" | "mega"; `math`: "exact" (bit-exact against the oracle) |
        "fast" (fp32 tolerance mode, B2VP_MATH_FAST); `tuning`: queue-kernel scheduler parameters; `majorant_grid`:
        non-zero = opt-in free-flight sampling against a grid of local majorants (B2vpRenderArgs::majorant_grid: same
        estimator, other random sequence -- statistically, not per-path, equal to the reference's global majorant);
        `timeout`: seconds after which no new streams are started (the integrator's `timeout` property; the result
        holds the samples taken so far and totals.stopped_early = 1); `devices`: list of CUDA devices for a multi-GPU
        render inside the library (devices[0] = this scene's device; films merged by one NCCL reduce)."""
        prec = self.scene.precision if precision is None else precision
        dtype = np.float32 if prec == abi.PREC_F32 else np.float64
        w, h = self.scene.film_size
        a = abi.RenderArgs()
        a.abi_version, a.struct_size = abi.ABI_VERSION, C.sizeof(abi.RenderArgs)
        a.precision, a.seed_mode, a.device = prec, seed_mode, self.device
        a.cuda_stream = stream
        a.first_stream, a.n_streams = first_stream, n_streams
        _set_launch_options(a, kernel, math, tuning, debug, majorant_grid)
        a.timeout_s = float(timeout)
        if devices is not None:
            dev = (C.c_int32 * len(devices))(*devices)
            a.devices, a.n_devices = dev, len(devices)
        if interleave is not None:   # (block, count, index): B2vpRenderArgs::interleave_*
            a.interleave_block, a.interleave_count, a.interleave_index = interleave
        keep = []
        if seed_mode == abi.SEED_EXPLICIT:
            st, inc, pix, per = explicit
            st = np.ascontiguousarray(st, dtype=np.uint64); inc = np.ascontiguousarray(inc, dtype=np.uint64)
            pix = np.ascontiguousarray(pix, dtype=np.uint32)
            keep += [st, inc, pix]
            a.explicit_state = st.ctypes.data_as(C.POINTER(C.c_uint64))
            a.explicit_inc = inc.ctypes.data_as(C.POINTER(C.c_uint64))
            a.explicit_pixel = pix.ctypes.data_as(C.POINTER(C.c_uint32))
            a.explicit_samples_per_stream = per
            a.n_streams = len(st)
            count, sps = len(st), per
        else:
            total, sps = self.stream_count(seed_mode)
            count = n_streams if n_streams else total - first_stream
        if device_film is not None:
            # film stays resident in HBM: `device_film` is a device pointer to H*W*5 reals (e.g. tensor.data_ptr())
            a.film_on_device, a.clear_film = 1, int(clear)
        elif film is None:
            film = np.zeros((h, w, self.scene.film_channels), dtype=dtype)
            a.clear_film = 1
        else:
            assert film.dtype == dtype and film.shape == (h, w, self.scene.film_channels) and film.flags.c_contiguous
            a.clear_film = 0
        tot = abi.Counters()
        a.totals = C.pointer(tot)
        state = counters = radiance = None
        if want_paths:
            state = np.zeros(count, dtype=np.uint64)
            counters = np.zeros((count, 6), dtype=np.uint32)
            radiance = np.zeros((count, sps), dtype=dtype)
            a.out_state = state.ctypes.data_as(C.POINTER(C.c_uint64))
            a.out_counters = C.cast(counters.ctypes.data, C.POINTER(abi.PathCounters))
            a.out_radiance = radiance.ctypes.data
        abi.check(abi.lib().b2vp_render(self._h, C.byref(a), device_film if device_film is not None
                                        else film.ctypes.data))
        return RenderResult(film, tot, state, counters, radiance)


    def render_noalloc(self, device_film, seed_mode=abi.SEED_WAVEFRONT, precision=None, first_stream=0, n_streams=0,
                       interleave=None, stream=None, clear=True, kernel=None, math=None, tuning=None, majorant_grid=0):
        """b2vp_render_noalloc: the same render with the kernels' scratch in the library's static device array -- no
        device allocation in the call (for callers under CUDA graph capture or in loops that must not cudaMalloc;
        no counterpart in the reference).  `device_film` is a device pointer to H*W*channels reals owned by the
        caller; the scene must have been rendered once with `render` before (that call uploads it).  Returns the
        launch totals.  Calls on one device share the scratch: do not overlap them."""
        a = abi.RenderArgs()
        a.abi_version, a.struct_size = abi.ABI_VERSION, C.sizeof(abi.RenderArgs)
        a.precision = self.scene.precision if precision is None else precision
        a.seed_mode, a.device = seed_mode, self.device
        a.cuda_stream = stream
        a.first_stream, a.n_streams = first_stream, n_streams
        _set_launch_options(a, kernel, math, tuning, 0, majorant_grid)
        if interleave is not None:
            a.interleave_block, a.interleave_count, a.interleave_index = interleave
        a.film_on_device, a.clear_film = 1, int(clear)
        tot = abi.Counters()
        a.totals = C.pointer(tot)
        abi.check(abi.lib().b2vp_render_noalloc(self._h, C.byref(a), device_film))
        return tot


def read_exr(path):
    """Reads back what RenderResult.write_exr writes (scan lines, no compression, HALF / FLOAT channels):
    {channel name: (h, w) array}.  For tests and for callers without an OpenEXR library."""
    import struct
    b = open(path, "rb").read()
    magic, version = struct.unpack_from("<II", b, 0)
    if magic != 20000630 or (version & 0xff) != 2 or version & 0x200:
        raise ValueError("not a scan-line OpenEXR 2 file")
    pos, attrs = 8, {}
    while b[pos] != 0:
        e = b.index(b"\0", pos); name = b[pos:e].decode(); pos = e + 1
        e = b.index(b"\0", pos); typ = b[pos:e].decode(); pos = e + 1
        (n,) = struct.unpack_from("<i", b, pos); pos += 4
        attrs[name] = (typ, b[pos:pos + n]); pos += n
    pos += 1
    if attrs["compression"][1] != b"\0":
        raise ValueError("compressed OpenEXR files are not supported by this reader")
    x0, y0, x1, y1 = struct.unpack("<4i", attrs["dataWindow"][1])
    w, h = x1 - x0 + 1, y1 - y0 + 1
    ch, c, p = [], attrs["channels"][1], 0
    while c[p] != 0:
        e = c.index(b"\0", p); name = c[p:e].decode(); p = e + 1
        (ctype,) = struct.unpack_from("<i", c, p); p += 16
        ch.append((name, {1: "<f2", 2: "<f4", 0: "<u4"}[ctype]))
    offsets = np.frombuffer(b, dtype="<u8", count=h, offset=pos)
    out = {name: np.zeros((h, w), dtype=dt) for name, dt in ch}
    for off in offsets:
        y, size = struct.unpack_from("<ii", b, int(off))
        q = int(off) + 8
        for name, dt in ch:
            out[name][y - y0] = np.frombuffer(b, dtype=dt, count=w, offset=q)
            q += w * np.dtype(dt).itemsize
    return out


def shard_streams(count, n, index):
    """b2vp_shard_streams: (block, count_i) of device `index` of `n` for a render of `count` streams -- the partition a
    multi-device b2vp_render uses; with first_stream it gives interleave=(block, n, index), n_streams=count_i."""
    block, ci = C.c_uint64(), C.c_uint64()
    abi.check(abi.lib().b2vp_shard_streams(int(count), int(n), int(index), C.byref(block), C.byref(ci)))
    return block.value, ci.value


def render(scene, device=0, **kw):
    """One-shot `integrator.render(scene, sensor)`."""
    with DeviceScene(scene, device) as ds:
        return ds.render(**kw)
