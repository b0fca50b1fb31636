/*
 * vp_kernels.cuh -- the render kernel: persistent CTAs pull PCG32 streams from
 * a global work counter; each thread renders its stream's samples
 * (render_sample, src/librender/integrator.cpp:233-288), splats them into the
 * film (ImageBlock::put with the box filter, src/librender/imageblock.cpp:79-172)
 * and optionally writes per-stream parity outputs.
 *
 * Shared memory layout (dynamic): [scene tables blob][staged grids][CTA film].
 * Grids small enough to fit are staged once per CTA with a bulk asynchronous
 * copy (TMA, cp.async.bulk) that completes on an mbarrier.
 */
#pragma once
#include "vp_integrator.cuh"

namespace VP_NS {

enum { SEED_SCALAR_BLOCKS = 0, SEED_WAVEFRONT = 1, SEED_EXPLICIT = 2 };
enum { TOT_PATHS = 0, TOT_DRAWS, TOT_STEPS, TOT_NULL, TOT_REAL, TOT_ISECT, TOT_SURF, TOT_INVALID, TOT_BVH_NODES, TOT_BVH_TRIS, TOT_N };

template <typename F> struct RenderOut {
    F *film;                        /* crop_w * crop_h * 5, XYZAW interleaved */
    uint64_t *out_state;            /* per stream or NULL */
    PathCounters *out_counters;     /* per stream or NULL */
    F *out_radiance;                /* per stream * samples or NULL */
    unsigned long long *totals;     /* TOT_N */
    unsigned long long *work_counter;
    uint32_t film_in_smem;          /* CTA-private film accumulated in shared memory */
    uint32_t film_smem_offset;
    uint32_t grid_smem_offset;
    /* queue kernel (vp_queue.cuh): path-slot pool and phase queues in shared memory */
    uint32_t q_slots_offset, q_ctl_offset, q_n_slots, q_ring_cap;
    uint32_t q_cta_phase;           /* 1: the sticky phase is shared by the CTA's warps, 0: per warp */
    uint32_t q_stay_min, q_max_iter; /* STEP keeps its batch while >= q_stay_min lanes stay, <= q_max_iter steps */
    unsigned long long *q_stats;    /* optional scheduler statistics (B2VP_QSTATS): pops[8], slots[8], switches, idle polls */
    void *q_cold;                   /* grid * q_n_slots PathCold records */
    /* majorant grids of the scene's media (floats, staged in shared memory at mg_smem_offset), or null */
    const float *mg;
    uint32_t mg_bytes, mg_smem_offset;
    /* raised by the host (cancel / timeout) while the kernel runs: no new stream is handed out once it is non-zero */
    const int *stop_flag;
};

template <typename F> VP_D bool stop_requested(const RenderOut<F> &out) {
    return out.stop_flag != nullptr && *reinterpret_cast<const volatile int *>(out.stop_flag) != 0;
}

struct StreamInfo {
    uint64_t state, inc;
    int px, py, box, boy, bsx, bsy;
    uint32_t n_samples;
    bool empty;
};

/* stream k -> generator state + pixel (integrator.cpp:147-153,197-200, sampler.cpp:83-96,
   spiral.cpp:27-72) */
VP_D StreamInfo make_stream(const DStreams &sp, uint64_t k) {
    StreamInfo st;
    st.n_samples = sp.samples_per_stream;
    st.empty = false;
    Pcg32 rng;
    if (sp.mode == SEED_SCALAR_BLOCKS) {
        uint64_t bs2 = (uint64_t) sp.bs * sp.bs;
        uint64_t per_pass = (uint64_t) sp.n_blocks * bs2;
        uint64_t pass = k / per_pass, r = k % per_pass;
        uint64_t counter = r / bs2;
        uint32_t i = (uint32_t) (r % bs2);
        /* spiral.cpp:42: block_id = counter + (remaining_passes - 1) * block_count */
        uint64_t block_id = counter + (uint64_t) (sp.n_passes - 1 - pass) * sp.n_blocks;
        SpiralBlockD b = sp.blocks[counter];
        rng.seed(sp.seed + block_id * bs2 + i, VP_PCG32_DEFAULT_STREAM);
        uint32_t x = compact_bits(i), y = compact_bits(i >> 1);
        st.empty = (int) x >= b.sx || (int) y >= b.sy;
        st.px = b.ox + (int) x; st.py = b.oy + (int) y;
        st.box = b.ox; st.boy = b.oy; st.bsx = b.sx; st.bsy = b.sy;
    } else if (sp.mode == SEED_WAVEFRONT) {
        rng.seed(tea64_u64(sp.seed, k), tea64_u64(k, sp.seed));
        uint64_t pixel = k / sp.spp_pass;
        st.px = (int) (pixel % (uint64_t) sp.w) + sp.ox;
        st.py = (int) (pixel / (uint64_t) sp.w) + sp.oy;
        st.box = sp.ox; st.boy = sp.oy; st.bsx = sp.w; st.bsy = sp.h;
    } else {
        rng.state = sp.ex_state[k];
        rng.inc = sp.ex_inc[k];
        uint32_t pixel = sp.ex_pixel[k];
        st.px = (int) (pixel % (uint32_t) sp.w) + sp.ox;
        st.py = (int) (pixel / (uint32_t) sp.w) + sp.oy;
        st.box = sp.ox; st.boy = sp.oy; st.bsx = sp.w; st.bsy = sp.h;
    }
    st.state = rng.state;
    st.inc = rng.inc;
    return st;
}

VP_D uint32_t smem_addr(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }

VP_D void film_add(float *p, float v) { atomicAdd(p, v); }
VP_D void film_add(double *p, double v) { atomicAdd(p, v); }

/* ImageBlock::put for a reconstruction filter wider than a pixel (imageblock.cpp:117-152), followed by the
   film's put(block) (hdrfilm.cpp:207-211, imageblock.cpp:52-76), which keeps what falls inside the crop window.
   The sample at film position (posx, posy) was drawn in the block at (box, boy) of bsx x bsy pixels; the block
   carries a border of rf_border pixels on every side, so its pixel (0, 0) is film pixel (box - border, boy -
   border).  Weights come from the filter's 32-entry table as eval_discretized() reads it (rfilter.h:68-71).
   Out of line: the box filter of every BASELINE configuration never comes here. */
template <typename F>
__device__ __noinline__ void film_put_wide(F *film, const DSensor<F> &sn, int NCH, F posx, F posy, int box, int boy,
                                           int bsx, int bsy, int ox, int oy, F value, F alpha, F raw) {
    const int b = sn.rf_border, n = sn.rf_taps;
    const int W = sn.crop_size[0], H = sn.crop_size[1];
    const F radius = sn.rf_radius;
    const F px = posx - ((F) (box - b) + (F) 0.5), py = posy - ((F) (boy - b) + (F) 0.5);
    const int lox = max((int) ceil_(px - radius), 0), loy = max((int) ceil_(py - radius), 0);
    const int hix = min((int) floor_(px + radius), bsx + 2 * b - 1), hiy = min((int) floor_(py + radius), bsy + 2 * b - 1);
    const F basex = (F) lox - px, basey = (F) loy - py;
    F wx[B2VP_RFILTER_MAX_TAPS], wy[B2VP_RFILTER_MAX_TAPS];
    for (int i = 0; i < n; ++i) {
        F qx = (basex + (F) i) * sn.rf_scale, qy = (basey + (F) i) * sn.rf_scale;
        if (qx < 0) qx = -qx;
        if (qy < 0) qy = -qy;
        /* the comparison in F first: int(x) of a huge x is not what min() should see */
        wx[i] = sn.rf_values[qx < (F) B2VP_RFILTER_RESOLUTION ? (int) qx : B2VP_RFILTER_RESOLUTION];
        wy[i] = sn.rf_values[qy < (F) B2VP_RFILTER_RESOLUTION ? (int) qy : B2VP_RFILTER_RESOLUTION];
    }
    const F raw2 = raw * raw;
    for (int yr = 0; yr < n; ++yr) {
        const int y = loy + yr;
        if (y > hiy) break;
        const int fy = y - b + boy - oy;
        for (int xr = 0; xr < n; ++xr) {
            const int x = lox + xr;
            if (x > hix) break;
            const int fx = x - b + box - ox;
            if (fx < 0 || fy < 0 || fx >= W || fy >= H) continue;
            const F weight = wy[yr] * wx[xr];
            F *p = film + ((size_t) fy * W + fx) * NCH;
            const F v = value * weight;
            film_add(p, v); film_add(p + 1, v); film_add(p + 2, v);
            film_add(p + 3, alpha * weight); film_add(p + 4, weight);
            if (NCH > 5) { /* moment.cpp:78-83 */
                const F a = raw * weight, a2 = raw2 * weight;
                film_add(p + 5, a); film_add(p + 6, a); film_add(p + 7, a);
                film_add(p + 8, a2); film_add(p + 9, a2); film_add(p + 10, a2);
            }
        }
    }
}

/* CTA prologue: stage the scene tables and the small grids in shared memory. */
template <typename F>
VP_D void stage_scene(const DScene<F> &g, const RenderOut<F> &out, unsigned char *smem, SceneView<F> &sv) {
    __shared__ __align__(8) unsigned long long bar;
    /* tables: plain cooperative copy (a few KB) */
    const uint32_t n16 = g.blob_bytes / 16;
    const uint4 *src = reinterpret_cast<const uint4 *>(g.blob);
    uint4 *dst = reinterpret_cast<uint4 *>(smem);
    for (uint32_t i = threadIdx.x; i < n16; i += blockDim.x) dst[i] = src[i];
    if (out.film_in_smem) {
        F *f = reinterpret_cast<F *>(smem + out.film_smem_offset);
        uint32_t n = (uint32_t) g.sensor.crop_size[0] * g.sensor.crop_size[1] * g.film_ch;
        for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) f[i] = (F) 0;
    }
    __syncthreads();
    sv.g = &g;
    sv.shapes = reinterpret_cast<const DShape<F> *>(smem + g.ofs_shapes);
    sv.analytic = reinterpret_cast<const int *>(smem + g.ofs_analytic);
    sv.bsdfs = reinterpret_cast<const DBsdf<F> *>(smem + g.ofs_bsdfs);
    sv.media = reinterpret_cast<const DMedium<F> *>(smem + g.ofs_media);
    sv.volumes = reinterpret_cast<const DVolume<F> *>(smem + g.ofs_volumes);
    sv.phases = reinterpret_cast<const DPhase<F> *>(smem + g.ofs_phases);
    sv.emitters = reinterpret_cast<const DEmitter<F> *>(smem + g.ofs_emitters);
    sv.grid_smem = smem + out.grid_smem_offset;
    sv.grid_saddr = smem_addr(sv.grid_smem);
    sv.bvh_saddr = smem_addr(smem + g.ofs_bvh);
    sv.tri_saddr = smem_addr(smem + g.ofs_bvh_tris);
    sv.mg_saddr = 0;
    if (out.mg) {   /* majorant grids: a few KB, plain cooperative copy */
        float *dstm = reinterpret_cast<float *>(smem + out.mg_smem_offset);
        for (uint32_t i = threadIdx.x; i < out.mg_bytes / 4; i += blockDim.x) dstm[i] = out.mg[i];
        sv.mg_saddr = smem_addr(dstm);
        __syncthreads();
    }

    if (g.grid_smem_bytes > 0) {
        /* TMA bulk copies global -> shared, completion counted on an mbarrier */
        const uint32_t bar_a = smem_addr(&bar);
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a),
                         "r"(g.grid_smem_bytes) : "memory");
            for (int i = 0; i < g.n_volumes; ++i) {
                const DVolume<F> &v = sv.volumes[i];
                if (v.type != 1 || v.smem_offset < 0) continue;
                uint32_t d = smem_addr(sv.grid_smem + v.smem_offset);
                uint32_t bytes = (uint32_t) v.data_bytes;
                asm volatile(
                    "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                    ::"r"(d), "l"(v.data), "r"(bytes), "r"(bar_a) : "memory");
            }
        }
        __syncthreads();
        uint32_t done = 0;
        while (!done) {
            asm volatile(
                "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n selp.u32 %0, 1, 0, p;\n}\n"
                : "=r"(done) : "r"(bar_a) : "memory");
        }
    }
}

template <typename F>
__global__ void __launch_bounds__(256)
volpath_kernel(const __grid_constant__ DScene<F> g, const __grid_constant__ DStreams sp,
               const __grid_constant__ RenderOut<F> out) {
    extern __shared__ __align__(128) unsigned char smem[];
    SceneView<F> sv;
    stage_scene(g, out, smem, sv);

    const DSensor<F> &sn = g.sensor;
    const int W = sn.crop_size[0], C = g.film_ch;
    F *film = out.film_in_smem ? reinterpret_cast<F *>(smem + out.film_smem_offset) : out.film;

    unsigned long long tot[TOT_N];
#pragma unroll
    for (int i = 0; i < TOT_N; ++i) tot[i] = 0;

    for (;;) {
        unsigned long long k = atomicAdd(out.work_counter, 1ull);
        if (k >= sp.count || stop_requested(out)) break;
        StreamInfo st = make_stream(sp, stream_index(sp, k));
        Pcg32 rng;
        rng.state = st.state; rng.inc = st.inc; rng.draws = 0;
        PathCounters cnt;
        cnt.draws = 0; cnt.medium_steps = 0; cnt.n_null = 0; cnt.n_real = 0; cnt.n_intersect = 0; cnt.n_surface = 0;
        F acc0 = 0, acc3 = 0, acc4 = 0;
        bool acc_used = false;
        for (uint32_t j = 0; j < st.n_samples && !st.empty; ++j) {
            /* render_sample (integrator.cpp:233-288) */
            F posx = (F) (uint32_t) st.px + rng.next(F(), true);
            F posy = (F) (uint32_t) st.py + rng.next(F(), true);
            F apx = (F) 0.5, apy = (F) 0.5;
            if (sn.needs_aperture_sample) { apx = rng.next(F(), true); apy = rng.next(F(), true); }
            if (sn.has_shutter) (void) rng.next(F(), true);
            (void) rng.next(F(), true); /* wavelength sample, drawn even in mono (integrator.cpp:252) */
            F ax = (posx - (F) sn.crop_offset[0]) / (F) sn.crop_size[0];
            F ay = (posy - (F) sn.crop_offset[1]) / (F) sn.crop_size[1];
            Ray<F> ray;
            F ray_weight = sensor_sample_ray(sv, ax, ay, apx, apy, ray);

            PathCtx<F> ctx{ sv, rng, cnt, g.max_depth, g.rr_depth, g.hide_emitters != 0, g.ltr != 0, g.mis != 2 };
            bool valid;
            const F raw_value = g.mis ? ctx.sample_mis(ray, sn.medium, valid) : ctx.sample(ray, sn.medium, valid);
            F value = ray_weight * raw_value;
            tot[TOT_PATHS]++;
            if (out.out_radiance)
                out.out_radiance[k * st.n_samples + j] = value;

            /* ImageBlock::put, box filter (imageblock.cpp:85-109,162-169) */
            F alpha = valid ? (F) 1 : (F) 0;
            bool ok = C > 5 || (value >= (F) -1e-5f && finite_(value)); /* no check with AOVs (integrator.cpp:113-115) */
            if (!ok) { tot[TOT_INVALID]++; continue; }
            F bx = posx - ((F) st.box + (F) 0.5), by = posy - ((F) st.boy + (F) 0.5);
            int lx = (int) ceil_(bx - (F) 0.5), ly = (int) ceil_(by - (F) 0.5);
            if (sn.rf_taps > 0) {
                film_put_wide(film, sn, C, posx, posy, st.box, st.boy, st.bsx, st.bsy, sp.ox, sp.oy, value, alpha, raw_value);
                continue;
            }
            if (lx < 0 || ly < 0 || lx >= st.bsx || ly >= st.bsy) continue;
            int fx = lx + st.box - sp.ox, fy = ly + st.boy - sp.oy;
            if (C > 5) { /* moment.cpp:78-83: the nested integrator's X Y Z (without the ray weight) and squares */
                F *q = film + ((size_t) fy * W + fx) * C + 5;
                const F raw = raw_value, raw2 = raw * raw;
                film_add(q, raw); film_add(q + 1, raw); film_add(q + 2, raw);
                film_add(q + 3, raw2); film_add(q + 4, raw2); film_add(q + 5, raw2);
            }
            if (fx == st.px - sp.ox && fy == st.py - sp.oy) {
                acc0 += value; acc3 += alpha; acc4 += (F) 1;
                acc_used = true;
            } else { /* position sample rounded onto a neighbouring pixel */
                F *p = film + ((size_t) fy * W + fx) * C;
                film_add(p, value); film_add(p + 1, value); film_add(p + 2, value);
                film_add(p + 3, alpha); film_add(p + 4, (F) 1);
            }
        }
        if (acc_used) {
            F *p = film + ((size_t) (st.py - sp.oy) * W + (st.px - sp.ox)) * C;
            film_add(p, acc0); film_add(p + 1, acc0); film_add(p + 2, acc0);
            film_add(p + 3, acc3); film_add(p + 4, acc4);
        }
        if (out.out_state) out.out_state[k] = rng.state;
        if (out.out_counters) {
            PathCounters o = cnt;
            o.draws = rng.draws;
            out.out_counters[k] = o;
        }
        tot[TOT_DRAWS] += rng.draws; tot[TOT_STEPS] += cnt.medium_steps; tot[TOT_NULL] += cnt.n_null;
        tot[TOT_REAL] += cnt.n_real; tot[TOT_ISECT] += cnt.n_intersect; tot[TOT_SURF] += cnt.n_surface;
    }

    /* totals: warp reduce, one atomic per warp and counter */
#pragma unroll
    for (int i = 0; i < TOT_N; ++i) {
        unsigned long long v = tot[i];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(out.totals + i, v);
    }
    if (out.film_in_smem) {
        __syncthreads();
        const F *f = reinterpret_cast<const F *>(smem + out.film_smem_offset);
        uint32_t n = (uint32_t) W * sn.crop_size[1] * C;
        for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
            F v = f[i];
            if (v != (F) 0) film_add(out.film + i, v);
        }
    }
}

} // namespace VP_NS
