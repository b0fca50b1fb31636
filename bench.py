#!/usr/bin/env python
"""bench.py -- volpath (mono) throughput on BASELINE.json configs[1]: the 1D plane-parallel heterogeneous
Rayleigh atmosphere (gridvolume 1x1x1000, rpv surface, distant 256x1 film), 2^14 spp = 4 194 304 paths
per GPU per step.  One JSON line on stdout (see the task's bench contract).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--spp S]

A "step" is one full render of the workload (one pass of the hot path over one batch of paths).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

FLOP_PER_SEGMENT = 150  # FP32 flop-equivalents per tentative collision
WIDTH = 256


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def host_info():
    """CPU model, physical cores and hardware threads of the box (for the cpu_baseline record)."""
    model, cores = "unknown", set()
    try:
        phys = core = None
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name") and model == "unknown":
                model = line.split(":", 1)[1].strip()
            elif line.startswith("physical id"):
                phys = line.split(":", 1)[1].strip()
            elif line.startswith("core id"):
                core = line.split(":", 1)[1].strip()
                cores.add((phys, core))
    except OSError:
        pass
    return {"cpu_model": model, "physical_cores": len(cores) or None, "hardware_threads": os.cpu_count()}


REF_BLOCK_SIZE = 2   # 256 x 1 film -> 128 image blocks: several per thread, so TBB can balance pixels of different cost


def workload(spp):
    import scenes
    return scenes.c2(spp=spp, width=WIDTH, tau=0.5, nz=1000)


def config(spp, n_gpus, extra=None, math="exact"):
    c = {"workload": "BASELINE configs[1]: 1D plane-parallel heterogeneous atmosphere, gridvolume sigma_t "
                     "1x1x1000 (trilinear, exp profile, tau=0.5), rayleigh, rpv, distant %dx1 film, volpath "
                     "max_depth=-1 rr_depth=5, scalar_mono (fp32)" % WIDTH,
         "spp_per_gpu": spp, "paths_per_step": WIDTH * spp * n_gpus, "seeding": "wavefront (sampler.cpp:89-92)",
         "sharding": "interleaved stream blocks (every GPU renders spp of the spp x N samples of every pixel) + one NCCL "
                     "film reduce" if n_gpus > 1 else "single GPU",
         "l2": "flushed between steps (256 MiB write); the 4 KB grid is staged in shared memory",
         "math": "fast: fp32 tolerance mode (B2VP_MATH_FAST: FMA contraction, hardware lg2/ex2/rcp/rsq; PCG32 exact; "
                 "gated against the reference films, tests/test_gpu_fastmath.py)" if math == "fast" else
                 "exact: every fp32 operation in the oracle's IEEE order (bit-exact against the CPU oracle)"}
    if extra:
        c.update(extra)
    return c


# ----------------------------------------------------------------------------- CPU (oracle) timing
def cpu_time_sample(pkg, scene, budget_s, threads):
    """Times the CPU oracle on a bounded prefix of the workload's stream space with `threads` host threads."""
    from oracle import oracle
    abi = pkg.abi
    t0 = time.perf_counter()
    pilot = oracle.render(pkg, scene, abi.SEED_WAVEFRONT, first_stream=0, n_streams=4096)
    rate1 = 4096 / max(time.perf_counter() - t0, 1e-6)      # paths/s on one thread
    total = scene.film_size[0] * scene.film_size[1] * scene.desc.sensor.sample_count
    n = int(min(total, max(4096, rate1 * threads * budget_s)))
    # spread the sample over the stream space so that every sensor direction is represented
    chunks = max(threads * 4, 64)
    per = max(1, n // chunks)
    ranges = [(int(i * (total - per) / max(chunks - 1, 1)), per) for i in range(chunks)]
    n = per * chunks
    t0 = time.perf_counter()
    film, tots = oracle.render_parallel(pkg, scene, abi.SEED_WAVEFRONT, ranges, threads)
    dt = time.perf_counter() - t0
    segs = sum(t.n_medium_steps + t.n_intersect for t in tots)
    return {"paths": n, "seconds": dt, "mpaths_s": n / dt / 1e6, "segments_s": segs / dt,
            "pilot_mpaths_s_1thread": rate1 / 1e6, "pilot": pilot}


def ref_time_sample(pkg, budget_s, threads):
    """Times the UNMODIFIED reference (oracle/_ref/refprobe, compiled from /root/reference by oracle/Makefile.ref;
    scalar_mono, TBB with `threads` threads) on the workload at a reduced sample count chosen to take about
    `budget_s` seconds.  The time is that of Integrator::render (the reference's "Processing finished (took ..)"
    interval): scene loading and kd-tree build are outside it.  Returns None where the reference is not built."""
    from oracle import ref_xml
    build = "release" if ref_xml.available("release") else "parity"
    if not ref_xml.available(build):
        return None
    import tempfile
    with tempfile.TemporaryDirectory(prefix="b2vp_refbench_") as tmp:
        def run(spp):
            d = ref_xml.round_f32(workload(spp))
            d["integrator"]["block_size"] = REF_BLOCK_SIZE
            _, sec = ref_xml.render(d, "scalar_mono", threads=threads, write_vol=pkg.scene.write_vol, keep_dir=tmp, build=build)
            return sec
        pilot_spp = 64
        t = max(run(pilot_spp), 1e-4)
        spp = int(min(16384, max(pilot_spp, pilot_spp * budget_s / t)))
        spp = 1 << (spp.bit_length() - 1)   # a power of two like the workload's 2^14
        sec = run(spp)
    n = WIDTH * spp
    return {"paths": n, "seconds": sec, "mpaths_s": n / sec / 1e6, "spp": spp,
            "build": "the reference's Release flags: " + ref_xml.RELEASE_FLAGS if build == "release"
                     else "parity build: " + ref_xml.PARITY_FLAGS,
            "block_size": REF_BLOCK_SIZE}


def plugin_time(pkg, spp):
    """The drop-in path end to end: the UNMODIFIED reference binary (oracle/_ref/refprobe) loads the workload's XML
    scene with <integrator type="volpath_b200"> and calls Integrator::render three times; the time of the last call
    (object graph -> descriptor, scene upload, kernel, film read-back, film->put) is reported.  None where
    oracle/_ref is not built."""
    from oracle import ref_xml
    if not (ref_xml.available() and os.path.exists(ref_xml.PLUGIN)):
        return None
    d = ref_xml.with_plugin(ref_xml.round_f32(workload(spp)), seeding="wavefront")
    _, sec = ref_xml.render(d, "scalar_mono", threads=1, repeats=3, write_vol=pkg.scene.write_vol)
    return {"value": WIDTH * spp / sec / 1e6, "unit": "Mpaths/s", "ms_per_render": sec * 1e3,
            "what": "reference binary + volpath_b200 plugin (wavefront seeding), time of Integrator::render "
                    "(third call in the process): object graph -> descriptor, scene upload, kernel, film read-back"}


def other_configs(pkg, math, fp32_peak, hbm_gbs):
    """BASELINE configs[0], [2], [4] at full size on one GPU, each with the roofline that bounds it:
    kernel time from CUDA events around the launch (B2vpCounters::gpu_ms), best of two after a warm-up render."""
    import scenes
    abi = pkg.abi
    out = []

    def run(name, what, d):
        sc = pkg.load_dict(d)
        with pkg.DeviceScene(sc, device=0) as ds:
            ds.render(seed_mode=abi.SEED_WAVEFRONT, math=math)
            best = None
            for _ in range(2):
                r = ds.render(seed_mode=abi.SEED_WAVEFRONT, math=math)
                if best is None or r.totals.gpu_ms < best.totals.gpu_ms:
                    best = r
            kernel = ds.info().last_kernel.decode()
        t, k = best.totals, best.totals.gpu_ms * 1e-3
        return {"config": name, "workload": what, "kernel": kernel, "paths": t.n_paths, "kernel_ms": t.gpu_ms,
                "mpaths_s": t.n_paths / k / 1e6, "segments_per_s": (t.n_medium_steps + t.n_intersect) / k}, t, k

    e, t, k = run("C1", "homogeneous Rayleigh cube over a diffuse rectangle, distant 1x1, 10^6 spp", scenes.c1(spp=1000000))
    ach = t.n_medium_steps * FLOP_PER_SEGMENT / k / 1e12
    e["roofline"] = {"bound": "fp32", "achieved": ach, "peak": fp32_peak, "unit": "TFLOP/s", "frac": ach / fp32_peak,
                     "note": "2.2 ms launch with 7 paths per slot: too short to fill the machine"}
    out.append(e)
    e, t, k = run("C3", "cloud field: gridvolume 256^3 sigma_t and albedo, hg g=0.85, distant 512x512, 64 spp",
                  scenes.c3(spp=64, n=256, size=512))
    alg = t.n_medium_steps * 32 * 2     # one 32-byte cell record per lookup and field (sigma_t, albedo)
    e["roofline"] = {"bound": "hbm", "achieved": alg / k / 1e9, "peak": hbm_gbs, "unit": "GB/s",
                     "frac": alg / k / 1e9 / hbm_gbs if hbm_gbs else None,
                     "algorithmic": "%d lookups x 32 B x 2 fields" % t.n_medium_steps,
                     "note": "gather-latency bound at 16 warps per SM, not bandwidth bound (DESIGN.md section 5)"}
    out.append(e)
    e, t, k = run("C5", "canopy: 10^6 triangles (BVH), homogeneous Rayleigh slab, diffuse leaves, mradiancemeter 4096, 256 spp",
                  scenes.c5(spp=256, n_leaves=500000, n_sensors=4096))
    alg = t.n_bvh_nodes * 32 + t.n_bvh_tris * 48
    e["roofline"] = {"bound": "hbm", "achieved": alg / k / 1e9, "peak": hbm_gbs, "unit": "GB/s",
                     "frac": alg / k / 1e9 / hbm_gbs if hbm_gbs else None,
                     "algorithmic": "%d BVH nodes x 32 B + %d leaf triangles x 48 B (kernel counters)" % (t.n_bvh_nodes, t.n_bvh_tris),
                     "note": "traversal is latency bound: 31 MB of nodes + 48 MB of leaf triangles sit in the 126 MB L2"}
    out.append(e)
    return out


class ClockSampler:
    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = [int(r[0]) for r in self.rows if r and r[0].isdigit()]
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        reasons = set()
        for r in self.rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------- arms
def run_reference(args):
    """Reference arm: the reference's own CPU implementation of the path on all host threads.  Where oracle/_ref
    is built (kind 'reference') that is the unmodified reference binary; otherwise the CPU restatement (oracle/,
    kind 'port')."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import b2vp_import
    pkg = b2vp_import.load()
    threads = host_threads()
    scene = pkg.load_dict(workload(args.spp))
    per_step = max(2.0, min(10.0, 120.0 / (args.steps + args.warmup)))
    res, kind = [], "port"
    for i in range(args.warmup + args.steps):
        r = ref_time_sample(pkg, per_step, threads)
        if r is not None:
            kind = "reference"
        else:
            r = cpu_time_sample(pkg, scene, per_step, threads)
        if i >= args.warmup:
            res.append(r)
    paths = sum(r["paths"] for r in res)
    secs = sum(r["seconds"] for r in res)
    v = paths / secs / 1e6
    # ray segments per path of this workload, counted by the instrumented restatement on a small sample
    from oracle import oracle
    pilot = oracle.render(pkg, scene, pkg.abi.SEED_WAVEFRONT, first_stream=0, n_streams=8192).totals
    seg = v * 1e6 * (pilot.n_medium_steps + pilot.n_intersect) / pilot.n_paths
    if kind == "reference":
        sample = ("the workload at %d spp (%d paths) per step, unmodified reference binary (refprobe, scalar_mono, TBB %d "
                  "threads, block_size %d = 128 blocks; %s), time of Integrator::render" % (
                      res[0]["spp"], res[0]["paths"], threads, res[0]["block_size"], res[0]["build"]))
    else:
        sample = "%d of %d streams of the workload per step (64+ chunks spread evenly over the film), %d threads" % (
            res[0]["paths"], WIDTH * args.spp, threads)
    print(json.dumps({
        "impl": "reference", "metric": "Mpaths/s", "value": v, "unit": "Mpaths/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / len(res) * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "segments_per_s": seg, "config": config(args.spp, args.gpus),
        "cpu_baseline": dict({"value": v, "unit": "Mpaths/s", "cores": threads, "kind": kind, "sample": sample,
                              "per_thread_mpaths_s": v / threads}, **host_info()),
        "e2e": {"value": v, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_ours(args):
    import torch
    import torch.distributed as dist
    import b2vp_import
    pkg = b2vp_import.load()
    abi = pkg.abi
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the B200 volpath has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"   # the version banner goes to stdout, where ONE JSON line is expected
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        # a CPU-side group for the one wait that must not touch the GPUs (strong scaling below): a rank that waits in
        # an NCCL barrier keeps a spinning kernel on its GPU, and the render kernel is one persistent CTA per SM
        host_group = dist.new_group(backend="gloo")
    n_gpus = world

    # weak scaling: every GPU renders `spp` samples per pixel; the global stream space has spp * N
    spp_total = args.spp * n_gpus
    scene = pkg.load_dict(workload(spp_total))
    per_gpu = WIDTH * args.spp
    # every rank renders `spp` of the spp * N samples of EVERY pixel (interleaved blocks, dist.shard_interleaved):
    # contiguous ranges would give each rank its own pixels, i.e. its own mix of path lengths (10 % imbalance at N=8)
    shard = pkg.shard_interleaved(spp_total, WIDTH, rank, world) if world > 1 else {"first_stream": 0, "n_streams": per_gpu}
    ds = pkg.DeviceScene(scene, device=local)
    film = torch.zeros((1, WIDTH, 5), dtype=torch.float32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        r = ds.render(seed_mode=abi.SEED_WAVEFRONT, device_film=film.data_ptr(), clear=True, stream=stream,
                      math=args.math, **shard)
        if world > 1:
            dist.reduce(film, dst=0, op=dist.ReduceOp.SUM)
        return r.totals

    for _ in range(args.warmup):
        flush.fill_(1)
        step()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    barrier()
    step_ms, kern_ms, launches, tot_last = [], [], 0, None
    for _ in range(args.steps):
        flush.fill_(1)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        tot = step()
        e1.record()
        barrier()
        step_ms.append(e0.elapsed_time(e1))
        kern_ms.append(tot.gpu_ms)
        launches += tot.n_launches
        tot_last = tot
    clk = clocks.stop() if rank == 0 else None
    t = torch.tensor([sum(step_ms)], dtype=torch.float64, device="cuda")
    segs = torch.tensor([float(tot_last.n_medium_steps + tot_last.n_intersect), float(tot_last.n_medium_steps),
                         float(sum(kern_ms) / len(kern_ms)), float(launches)], dtype=torch.float64, device="cuda")
    seg_sum = segs.clone()
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(seg_sum, op=dist.ReduceOp.SUM)
    total_ms = float(t.item())
    launches = int(seg_sum[3].item())   # kernels launched by all ranks inside the timed region
    ms_per_step = total_ms / args.steps
    paths_per_step = per_gpu * n_gpus
    value = paths_per_step / (ms_per_step * 1e-3) / 1e6
    seg_per_s = float(seg_sum[0].item()) / (ms_per_step * 1e-3)

    # ---- e2e: through the public API with HOST buffers: scene upload (H2D) + render + film read-back (D2H)
    host_film = np.zeros((1, WIDTH, 5), dtype=np.float32)
    grid_bytes = sum(a.nbytes for a in scene._keep[1] if a is not None)
    h2d = grid_bytes + 4096
    e2e_ms = []
    for i in range(0 if args.profile else 2 + min(args.steps, 5)):
        barrier()
        t0 = time.perf_counter()
        with pkg.DeviceScene(scene, device=local) as ds2:
            r = ds2.render(seed_mode=abi.SEED_WAVEFRONT, film=host_film * 0, math=args.math, **shard)
            if world > 1:
                f = torch.from_numpy(r.film).cuda()
                dist.reduce(f, dst=0, op=dist.ReduceOp.SUM)
                r.film[...] = f.cpu().numpy()
        barrier()
        if i >= 2:
            e2e_ms.append((time.perf_counter() - t0) * 1e3)
    e2e_t = torch.tensor([sum(e2e_ms) / max(len(e2e_ms), 1) if e2e_ms else float("nan")], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_val = paths_per_step / (float(e2e_t.item()) * 1e-3) / 1e6

    # ---- the bit-exact mode timed beside the product (fast) mode: same workload, kernel-only
    exact = None
    if world == 1 and args.math == "fast" and not args.profile:
        ms = []
        for i in range(5):
            flush.fill_(1)
            torch.cuda.synchronize()
            r = ds.render(seed_mode=abi.SEED_WAVEFRONT, device_film=film.data_ptr(), clear=True, stream=stream, math="exact", **shard)
            if i >= 2:
                ms.append(r.totals.gpu_ms)
        exact = {"value": per_gpu / (sum(ms) / len(ms) * 1e-3) / 1e6, "unit": "Mpaths/s", "kernel_ms": sum(ms) / len(ms),
                 "what": "B2VP_MATH_EXACT: every fp32 operation in the oracle's IEEE order; films, PCG32 states and event "
                         "counters bit-identical to the CPU oracle"}

    # ---- strong scaling: the FIXED single-GPU workload (spp per pixel, not spp x N) split over the N GPUs, one process,
    # through b2vp_render's device list (host thread per GPU, one NCCL film reduce); rank 0 drives, the others wait
    strong = None
    if world > 1 and not args.profile:
        barrier()
        if rank == 0:
            try:
                sc1 = pkg.load_dict(workload(args.spp))
                with pkg.DeviceScene(sc1, device=local) as dsm:
                    devs = [local] + [d for d in range(world) if d != local]
                    dsm.render(math=args.math, devices=devs)
                    ts = []
                    for _ in range(5):
                        t0 = time.perf_counter()
                        r = dsm.render(math=args.math, devices=devs)
                        ts.append(time.perf_counter() - t0)
                    t1 = min(ts)
                strong = {"gpus": world, "paths": int(r.totals.n_paths), "wall_ms": t1 * 1e3, "mpaths_s": r.totals.n_paths / t1 / 1e6,
                          "slowest_kernel_ms": r.totals.gpu_ms,
                          "what": "the fixed %d-path workload of N=1 split over %d GPUs inside b2vp_render (device list), "
                                  "wall time of the call incl. the NCCL film reduce and the film read-back" % (r.totals.n_paths, world)}
            except Exception as e:   # reported, never fatal for the bench line
                strong = {"error": str(e)[-300:]}
        dist.barrier(group=host_group)   # the other ranks wait here on the CPU, their GPUs idle for rank 0
        barrier()

    if rank == 0:
        # ---- roofline of the dominant kernel (the queue kernel, vp_queue.cuh): FP32 issue bound on this workload
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        fp32 = abi.c_d()
        abi.check(abi.lib().b2vp_measure_fp32_peak(local, abi.C.byref(fp32)))
        k_ms = sum(kern_ms) / len(kern_ms)
        flops = tot_last.n_medium_steps * FLOP_PER_SEGMENT
        achieved = flops / (k_ms * 1e-3) / 1e12
        # DRAM traffic of one launch of this kernel on this workload, from the committed `ncu --set full`
        # capture (profiles/ncu_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum); None if the
        # capture is for another spp
        traffic, traffic_source = None, "not measured at N > 1"
        if world == 1:
            try:
                tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))[args.math]
                if int(tr.get("spp", -1)) == args.spp:
                    traffic = tr["dram_bytes_per_launch"]
                    traffic_source = "dram__bytes_read.sum + dram__bytes_write.sum of this kernel on this workload from the " \
                                     "committed capture %s (N = 1; ncu cannot run inside the timed region)" % tr["capture"]
            except (OSError, ValueError, KeyError):
                traffic_source = "no committed capture for this mode"
        roofline = {"bound": "fp32", "achieved": achieved, "peak": fp32.value, "unit": "TFLOP/s",
                    "frac": achieved / fp32.value if fp32.value else None, "traffic": traffic, "traffic_source": traffic_source,
                    "peak_is": "the FMA peak (2 flop per instruction); the exact mode issues no contracted FMAs, its "
                               "ceiling is half of it" if args.math == "exact" else "the FMA peak (2 flop per instruction)",
                    "kernel": "vp::volpath_q_kernel [%s]" % ds.info().last_kernel.decode(), "kernel_ms": k_ms,
                    "algorithmic": "%d medium steps x %d flop" % (tot_last.n_medium_steps, FLOP_PER_SEGMENT),
                    "peak_source": "b2vp_measure_fp32_peak (FMA microbenchmark, measured in this run); "
                                   "HBM peak %s GB/s (MEASURED_PEAKS.json) is not the bound here: the grid "
                                   "lives in shared memory, algorithmic HBM bytes ~ 0" % peaks.get("hbm_gbs")}
        # ---- CPU baseline beside it (bounded sample, all host threads)
        threads = host_threads()
        # (timed at N=1 only, as the bench contract asks: at N>1 the other ranks must not wait for host work)
        host_legs = world == 1 and not args.profile
        cb = cpu_time_sample(pkg, pkg.load_dict(workload(args.spp)), 8.0 if host_legs else 0.2, threads)
        rb = ref_time_sample(pkg, 8.0, threads) if host_legs else None
        configs = None
        if host_legs:
            try:
                configs = other_configs(pkg, args.math, fp32.value, peaks.get("hbm_gbs"))
            except Exception as e:   # reported, never fatal for the bench line
                configs = [{"error": str(e)[-300:]}]
        out = {
            "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": n_gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "segments_per_s": seg_per_s, "config": config(args.spp, n_gpus, math=args.math),
            "clocks": clk, "gpu_launches": launches,
            "e2e": {"value": e2e_val, "unit": "Mpaths/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(host_film.nbytes), "ms_per_step": float(e2e_t.item())},
            "roofline": roofline,
            "cpu_baseline": {"value": cb["mpaths_s"], "unit": "Mpaths/s", "cores": threads, "kind": "port",
                             "sample": "%d streams of the workload spread evenly over the film, %.1f s" % (cb["paths"], cb["seconds"]),
                             "segments_per_s": cb["segments_s"],
                             "one_thread_mpaths_s": cb["pilot_mpaths_s_1thread"]},
        }
        if host_legs:
            try:
                pl = plugin_time(pkg, args.spp)
            except RuntimeError as e:   # reported, never fatal for the bench line
                pl = {"error": str(e)[-300:]}
            if pl is not None:
                out["e2e_plugin"] = pl
        if exact is not None:
            out["exact_mode"] = exact
        if strong is not None:
            out["strong_scaling"] = strong
        if configs is not None:
            out["configs"] = configs
        if not host_legs:
            out["cpu_baseline"]["sample"] += " (a token sample: the CPU legs are timed in the N=1 run)"
        if rb is not None:  # the unmodified reference binary is built on this box: it is the baseline, the port is kept beside it
            out["cpu_baseline"] = {"value": rb["mpaths_s"], "unit": "Mpaths/s", "cores": threads, "kind": "reference",
                                   "sample": "the workload at %d spp (%d paths), unmodified reference binary "
                                             "(refprobe, scalar_mono, TBB %d threads, block_size %d = 128 image blocks), "
                                             "%.1f s of Integrator::render" % (rb["spp"], rb["paths"], threads,
                                                                                rb["block_size"], rb["seconds"]),
                                   "build": rb["build"], "per_thread_mpaths_s": rb["mpaths_s"] / threads,
                                   "port_mpaths_s": cb["mpaths_s"], "port_segments_per_s": cb["segments_s"],
                                   "port_one_thread_mpaths_s": cb["pilot_mpaths_s_1thread"]}
        out["cpu_baseline"].update(host_info())
        print(json.dumps(out))
    ds.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--profile", action="store_true", help="kernel-only run for ncu: no e2e leg, no CPU baseline")
    ap.add_argument("--spp", type=int, default=16384, help="samples per pixel per GPU (BASELINE: 2^14)")
    ap.add_argument("--math", default="fast", choices=["exact", "fast"],
                    help="fp32 mode of the kernels: fast (the product mode for scalar_mono renders: tolerance mode, gated "
                         "against the reference's films) or exact (bit-exact against the oracle; timed beside it at N=1)")
    args = ap.parse_args()
    # ONE JSON line on stdout: native libraries that write to file descriptor 1 themselves (NCCL prints its version
    # banner there at N>1) are sent to stderr; Python's own stdout keeps the real one
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w")
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
