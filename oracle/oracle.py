"""TEST INFRASTRUCTURE: ctypes wrapper of the CPU oracle (oracle/volpath_oracle.cpp).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm may
import this module; the product package never does."""
import ctypes as C
import os
import subprocess
import threading

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb2vp_oracle.so")
_lib = None


def build():
    subprocess.check_call(["make", "-C", HERE, "-s"])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.b2vp_oracle_last_error.restype = C.c_char_p
        L.b2vp_oracle_tea32.restype = C.c_uint32
        L.b2vp_oracle_tea32.argtypes = [C.c_uint32, C.c_uint32, C.c_int]
        L.b2vp_oracle_tea64.restype = C.c_uint64
        L.b2vp_oracle_tea64.argtypes = [C.c_uint32, C.c_uint32, C.c_int]
        L.b2vp_oracle_tea64_u64.restype = C.c_uint64
        L.b2vp_oracle_tea64_u64.argtypes = [C.c_uint64, C.c_uint64]
        L.b2vp_oracle_tea_float32.restype = C.c_float
        L.b2vp_oracle_tea_float32.argtypes = [C.c_uint32, C.c_uint32, C.c_int]
        L.b2vp_oracle_tea_float64.restype = C.c_double
        L.b2vp_oracle_tea_float64.argtypes = [C.c_uint32, C.c_uint32, C.c_int]
        L.b2vp_oracle_rpv_eval.restype = C.c_double
        L.b2vp_oracle_rpv_eval.argtypes = [C.c_double] * 4 + [C.POINTER(C.c_double)] * 2
        L.b2vp_oracle_pcg32.argtypes = [C.c_uint64, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.b2vp_oracle_pcg32.restype = None
        L.b2vp_oracle_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def render(pkg, scene, seed_mode, precision=None, first_stream=0, n_streams=0, explicit=None, want_paths=False,
           interleave=None):
    """Same contract as eradiate_kernel_b200.DeviceScene.render, computed by the CPU oracle.
    `pkg` is the loaded product package (for its ctypes struct definitions only)."""
    abi = pkg.abi
    prec = scene.precision if precision is None else precision
    dtype = np.float32 if prec == abi.PREC_F32 else np.float64
    w, h = scene.film_size
    a = abi.RenderArgs()
    a.abi_version, a.struct_size = abi.ABI_VERSION, C.sizeof(abi.RenderArgs)
    a.precision, a.seed_mode = prec, seed_mode
    a.first_stream, a.n_streams = first_stream, n_streams
    if interleave is not None:
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
        n, per = C.c_uint64(), C.c_uint32()
        rc = lib().b2vp_oracle_stream_count(C.byref(scene.desc), seed_mode, C.byref(n), C.byref(per))
        if rc:
            raise RuntimeError(lib().b2vp_oracle_last_error().decode())
        count = n_streams if n_streams else n.value - first_stream
        sps = per.value
    film = np.zeros((h, w, scene.film_channels), dtype=dtype)
    a.clear_film = 1
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
    rc = lib().b2vp_oracle_render(C.byref(scene.desc), C.byref(a), film.ctypes.data)
    if rc:
        raise RuntimeError("oracle error %d: %s" % (rc, lib().b2vp_oracle_last_error().decode()))
    return pkg.RenderResult(film, tot, state, counters, radiance)


def render_parallel(pkg, scene, seed_mode, ranges, n_threads, precision=None):
    """Renders a list of (first_stream, n_streams) ranges on `n_threads` host threads (ctypes releases the
    GIL).  Returns (summed film in float64, list of totals)."""
    results, lock, todo = [], threading.Lock(), list(ranges)

    def work():
        while True:
            with lock:
                if not todo:
                    return
                lo, n = todo.pop()
            r = render(pkg, scene, seed_mode, precision, lo, n)
            with lock:
                results.append(r)
    ts = [threading.Thread(target=work) for _ in range(n_threads)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    film = sum(r.film.astype(np.float64) for r in results)
    return film, [r.totals for r in results]


def sensor_ray(scene, fx, fy, ax=0.5, ay=0.5):
    """One Sensor::sample_ray of the scene's sensor: film position (fx, fy) in [0,1)^2, aperture sample
    (ax, ay).  Returns (origin, direction, weight)."""
    o, d, w = (C.c_double * 3)(), (C.c_double * 3)(), C.c_double()
    f = lib().b2vp_oracle_probe_sensor_ray
    f.argtypes = [C.c_void_p, C.c_int] + [C.c_double] * 4 + [C.c_void_p] * 3
    rc = f(C.byref(scene.desc), scene.precision, fx, fy, ax, ay, o, d, C.byref(w))
    if rc:
        raise RuntimeError("oracle error %d" % rc)
    return np.array(o[:]), np.array(d[:]), w.value


def ray_intersect(scene, o, d, mint=-1.0, maxt=-1.0):
    """One Scene::ray_intersect.  Returns dict(t (None if no hit), p, n, wi, shape)."""
    oo, dd = (C.c_double * 3)(*o), (C.c_double * 3)(*d)
    t, p, n, wi, shape = C.c_double(), (C.c_double * 3)(), (C.c_double * 3)(), (C.c_double * 3)(), C.c_int()
    f = lib().b2vp_oracle_probe_ray_intersect
    f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_double] + [C.c_void_p] * 5
    rc = f(C.byref(scene.desc), scene.precision, oo, dd, mint, maxt, C.byref(t), p, n, C.byref(shape), wi)
    if rc:
        raise RuntimeError("oracle error %d" % rc)
    return {"t": None if t.value < 0 else t.value, "p": np.array(p[:]), "n": np.array(n[:]),
            "wi": np.array(wi[:]), "shape": shape.value}


def _vec(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(C.c_void_p)


def phase_samples(pkg, phase, wi, samples):
    """n samples of a phase function (abi.Phase) for incident direction wi: (wo[n, 3], value[n]) -- the value of the
    phase function in each sampled direction, which is also its pdf (phase functions are normalised)."""
    samples, sp = _vec(samples)
    n = len(samples)
    wi, wp = _vec(wi)
    wo, value = np.zeros((n, 3)), np.zeros(n)
    lib().b2vp_oracle_probe_phase(C.byref(phase), wp, C.c_int(n), sp, wo.ctypes.data_as(C.c_void_p), value.ctypes.data_as(C.c_void_p))
    return wo, value


def phase_eval(pkg, phase, wi, wo):
    f = lib().b2vp_oracle_probe_phase_eval
    f.restype = C.c_double
    wi, wp = _vec(wi)
    wo, op = _vec(wo)
    return f(C.byref(phase), wp, op)


def bsdf_samples(pkg, bsdf, wi, samples):
    """n samples of a BSDF (abi.Bsdf) in the local frame: (wo[n, 3], value[n], pdf[n], weight[n])."""
    samples, sp = _vec(samples)
    n = len(samples)
    wi, wp = _vec(wi)
    wo, value, pdf, weight = np.zeros((n, 3)), np.zeros(n), np.zeros(n), np.zeros(n)
    lib().b2vp_oracle_probe_bsdf(C.byref(bsdf), wp, C.c_int(n), sp, *[a.ctypes.data_as(C.c_void_p) for a in (wo, value, pdf, weight)])
    return wo, value, pdf, weight


def bsdf_eval(pkg, bsdf, wi, wo):
    wo, op = _vec(wo)
    wi, wp = _vec(wi)
    n = len(wo)
    value, pdf = np.zeros(n), np.zeros(n)
    lib().b2vp_oracle_probe_bsdf_eval(C.byref(bsdf), wp, C.c_int(n), op, value.ctypes.data_as(C.c_void_p), pdf.ctypes.data_as(C.c_void_p))
    return value, pdf
