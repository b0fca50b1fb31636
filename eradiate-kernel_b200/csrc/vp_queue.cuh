/*
 * vp_queue.cuh -- the production render kernel: a CTA-resident wavefront with
 * shared-memory phase queues.
 *
 * A CTA owns a pool of path slots (about 2.4 per thread) whose hot state lives in
 * shared memory.  A path is always in exactly one of five phase queues (also
 * shared memory) or in flight in a warp:
 *
 *   STEP     one tentative-collision step (loop top, Medium::sample_interaction,
 *            null / real / surface decision) -- shared by the main path, the
 *            emitter-sampling march (sample_emitter) and the direct-light march
 *            (evaluate_direct_light);
 *   SCATTER  real medium scattering: emitter sample + phase eval, phase sample;
 *   SURFACE  surface event: emitter sample + BSDF eval, BSDF sample;
 *   MARCH    a march crosses a surface or ends;
 *   END      film splat of a finished path and path regeneration.
 *
 * Every phase but STEP makes rays; Scene::ray_intersect runs at the end of the
 * visit that made the ray (one shared call site), so a path never queues for it.
 *
 * Each warp repeatedly pops up to 32 slots of ONE phase (the phase the whole CTA
 * is working on while that queue holds a full batch, else the fullest queue),
 * loads the fields that phase needs into registers, runs the phase for all 32
 * lanes together, stores the fields back and pushes every slot onto the queue of
 * its next phase.  This is the compaction / sorting by event type: warps execute
 * full, converged batches, and an SM's instruction cache sees one phase's code at
 * a time (the kernel is instruction-fetch bound otherwise, see profiles/).
 *
 * Arithmetic and random-number order are exactly those of vp_integrator.cuh (the
 * one-thread-per-path twin, reference src/integrators/volpath.cpp:38-483); only
 * the interleaving between paths changes.
 */
#pragma once
#include "vp_kernels.cuh"

namespace VP_NS {

enum QState : int {
    Q_STEP = 0,     /* one tentative-collision step                            */
    Q_SCATTER,      /* real medium scattering: albedo, start emitter march     */
    Q_SCATTER_B,    /* phase function sampling                                 */
    Q_SURFACE,      /* surface hit: start emitter march                        */
    Q_SURF_B,       /* BSDF sampling, second intersection                      */
    Q_MARCH_SURF,   /* a march crosses a surface (null-BSDF boundary or blocker) */
    Q_MARCH_END,    /* a march is over: hand back its result, restore the path */
    Q_PATH_END,
    Q_NEW,          /* next sample of the stream / next stream                 */
    Q_DEAD
};
enum QPhase : int { PH_STEP = 0, PH_SCATTER, PH_SURFACE, PH_MARCH, PH_END, PH_N };
enum QMode : int { QM_MAIN = 0, QM_NEE = 1, QM_DL = 2 };
enum QCont : int { QC_STEP_B = 0, QC_SCATTER_B, QC_SURF_B, QC_START_DL, QC_AFTER_DL };

VP_D int phase_of(int state) {
    return state == Q_STEP ? PH_STEP : state <= Q_SCATTER_B ? PH_SCATTER
           : state <= Q_SURF_B ? PH_SURFACE : state <= Q_MARCH_END ? PH_MARCH : PH_END;
}

/* flag word of a slot */
enum : uint32_t {
    QF_STATE = 0xfu, QF_MODE_SHIFT = 4, QF_MODE = 0x3u << 4, QF_CONT_SHIFT = 6, QF_CONT = 0x7u << 6,
    QF_TOP_DONE = 1u << 10,  /* the part of this loop iteration that precedes the step (roulette / remaining
                                distance, intersection) was done where the ray was made */
    QF_VALID_RAY = 1u << 11,
    QF_SPEC_CHAIN = 1u << 12, /* `specular_chain` (volpath.cpp:53, volpathmis.cpp:104): kept only with an environment */
    QF_HAVE_STREAM = 1u << 13,
    QF_ACC_USED = 1u << 14
};

/* A ray / hit / medium / weight "context": the thing a STEP marches.  In main mode
   it is the path's own ray and throughput; during a march it is the shadow / BSDF
   ray and the transmittance, and the main path's context sits in `saved`. */
template <typename F> struct MarchCtx {
    F o[3], d[3], d_rcp[3], mint, maxt;   /* Ray3f                                  */
    F hit_t, hit_p[3], hit_n[3];          /* closest surface along the ray          */
    int hit_shape;
    int medium;
    F w;                                  /* throughput / transmittance             */
};

/* Per-path state is split by how often it is touched.  PathSlot (shared memory) holds what the
   tentative-collision step reads and writes every iteration; PathCold (global memory, L2-resident, one
   array per CTA) holds what only the rarer phases need.  The smaller the shared-memory slot, the more paths
   a CTA keeps in flight and the fuller its batches. */
/* PathSlot is 35 words in fp32: an ODD word stride, so that one field of the 32 slots of a batch spreads over the 32
   shared-memory banks (slot ids within a batch are arbitrary; with the even stride of an 8-byte aligned record one
   field of 32 slots could reach 16 banks at most -- 61 % of all shared wavefronts were conflict replays in the round-1
   capture).  That is why the two PCG32 words are stored as 32-bit halves (4-byte alignment). */
template <typename F> struct PathSlot {
    uint32_t rng_state_lo, rng_state_hi, rng_inc_lo, rng_inc_hi;
    uint32_t rng_draws;
    uint32_t c_steps, c_null, c_isect;
    uint32_t flags;
    MarchCtx<F> cur;
    F eta;
    uint32_t depth;
    /* emitter march: distance covered, distance to the emitter.  (The remaining distance of the current march
       iteration is not stored: whoever makes or continues a march ray stores it as the ray's maxt.) */
    F total_dist, ds_dist;
    F w2;   /* volpathmis: the second weight of the current context -- p_over_f_uni of an emitter march (volpathmis.cpp:330-444) */
};
template <typename F> struct PathCold {
    unsigned long long k;
    uint32_t j;
    uint32_t c_real, c_surf;
    MarchCtx<F> saved;                    /* the main path's context during a march */
    uint32_t prim; F bu, bv;              /* mesh face and barycentrics of the main path's hit (dead once the surface
                                             frame is complete; volpathmis keeps the caller's pdf of an emitter march in
                                             `bu` and its phase / BSDF value in `nee_w` while the march runs) */
    F wi[3], sh_s[3], sh_t[3], sh_n[3];   /* its shading frame, completed at the surface event  */
    F result;
    int medium_next;
    F emitter_val, nee_w;
    F acc0, acc3, acc4, posx, posy, ray_weight;
};

/* Lock-free phase queues: one ring of slot ids per phase.  A pusher reserves positions at `tail`, writes the
   ids, then publishes them by raising `avail`; a popper first takes entries out of `avail` (compare-and-swap),
   then reserves positions at `head` and waits for each id to appear (the pusher that reserved a position is
   already writing it).  Entries are reset to Q_EMPTY when read, so a ring position is never read twice. */
struct QueueCtl {
    int head[PH_N], tail[PH_N], avail[PH_N];
    int n_live;
    int cur;   /* the phase the CTA is working on (see the pop) */
};
#define Q_EMPTY 0xffffu

template <typename F> VP_D void ld_rng(const PathSlot<F> &s, Pcg32 &rng) {
    rng.state = ((uint64_t) s.rng_state_hi << 32) | s.rng_state_lo;
    rng.inc = ((uint64_t) s.rng_inc_hi << 32) | s.rng_inc_lo;
    rng.draws = s.rng_draws;
}
template <typename F> VP_D void st_rng(PathSlot<F> &s, const Pcg32 &rng) {   /* the increment never changes */
    s.rng_state_lo = (uint32_t) rng.state; s.rng_state_hi = (uint32_t) (rng.state >> 32);
    s.rng_draws = rng.draws;
}
template <typename F> VP_D void ld_ray(const MarchCtx<F> &c, Ray<F> &r) {
    r.o = ld3(c.o); r.d = ld3(c.d); r.d_rcp = ld3(c.d_rcp); r.mint = c.mint; r.maxt = c.maxt;
}
template <typename F> VP_D void st3(F *p, V3<F> v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }
template <typename F> VP_D void st_ray(MarchCtx<F> &c, const Ray<F> &r) {
    st3(c.o, r.o); st3(c.d, r.d); st3(c.d_rcp, r.d_rcp); c.mint = r.mint; c.maxt = r.maxt;
}
template <typename F> VP_D void copy_ctx(MarchCtx<F> &dst, const MarchCtx<F> &src) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        dst.o[i] = src.o[i]; dst.d[i] = src.d[i]; dst.d_rcp[i] = src.d_rcp[i];
        dst.hit_p[i] = src.hit_p[i]; dst.hit_n[i] = src.hit_n[i];
    }
    dst.mint = src.mint; dst.maxt = src.maxt; dst.hit_t = src.hit_t; dst.hit_shape = src.hit_shape;
    dst.medium = src.medium; dst.w = src.w;
}

/* Top of the main loop: termination and Russian roulette (volpath.cpp:72-87).  False: the path ends. */
template <typename F>
VP_D bool loop_top_rr(Pcg32 &rng, F &w, F eta, uint32_t depth, uint32_t rr_depth, uint32_t max_depth) {
    bool active = w != 0;
    F q = min_(w * sqr(eta), (F) 0.95f);
    bool perform_rr = depth > rr_depth;
    F rr_sample = rng.next(F(), active);
    active = active && (rr_sample < q || !perform_rr);
    if (perform_rr) w *= rcp_(q);
    return active && depth < max_depth;
}

/* ---- volpathmis (src/integrators/volpathmis.cpp), mono: the 1 x 1 weight "matrices" are scalars.  A path carries
   p_over_f (pdf over contribution) where volpath carries a throughput; every event multiplies a ratio into it. */
template <typename F> VP_D void mis_update(F &p_over_f, F p, F f) {   /* update_weights, volpathmis.cpp:447-465 */
    F ratio = p / f;
    ratio = finite_(ratio) ? ratio : (F) 0;
    ratio *= p_over_f;
    p_over_f = (ratio != ratio) ? (F) 0 : ratio;
}
template <typename F> VP_D F mis_w1(F p_over_f) { return p_over_f == 0 ? (F) 0 : (F) 1 / p_over_f; }          /* :467-482 */
template <typename F> VP_D F mis_w2(F a, F b) { F sum = a + b; return sum == 0 ? (F) 0 : (F) 1 / sum; }       /* :485-500 */
/* Top of the main loop (volpathmis.cpp:122-133).  False: the path ends. */
template <typename F>
VP_D bool mis_loop_top(Pcg32 &rng, F &p_over_f, F eta, uint32_t depth, uint32_t rr_depth, uint32_t max_depth) {
    const F q = min_(mis_w1(p_over_f) * sqr(eta), (F) 0.95f);
    const bool perform_rr = depth > rr_depth;
    const F rr_sample = rng.next(F(), true);
    bool active = !(rr_sample >= q && perform_rr);
    if (perform_rr) mis_update(p_over_f, q, (F) 1);
    active = active && depth < max_depth;
    return active && mis_w1(p_over_f) != 0;
}

/* Scene::sample_emitter_direction as the integrators call it: returns the sample and the emitter VALUE the caller
   works with -- the importance weight for volpath (volpath.cpp:266-269), weight * pdf for volpathmis
   (volpathmis.cpp:341-343); 0 when the pdf is 0. */
template <typename F, uint32_t FEAT>
VP_D DirSample<F> queue_sample_emitter(const SceneView<F> &sv, Pcg32 &rng, F &ev) {
    const F sx = rng.next(F(), true), sy = rng.next(F(), true);
    DirSample<F> ds = (FEAT & FT_ENV) ? sample_emitter_direction_env(sv, sx, sy, ev) : sample_emitter_direction(sv, sx, ev);
    if (FEAT & FT_MIS) ev = ev * ds.pdf;
    if (ds.pdf == 0) ev = 0;
    return ds;
}
/* Scene::pdf_emitter_direction for the environment (scene.cpp:221-235 over constant.cpp:106-110) */
template <typename F> VP_D F env_pdf(const SceneView<F> &sv) {
    F pdf = K<F>::InvFourPi;
    if (sv.g->n_emitters > 1) pdf = pdf * (F) (1.f / (float) sv.g->n_emitters);   /* scene.cpp:233: a float quotient in every variant */
    return pdf;
}
/* The main path's ray left the scene: the environment is the emitter it "hits" (volpath.cpp:186-191,
   volpathmis.cpp:252-268).  w = throughput (volpath) or p_over_f (volpathmis), w_nee = p_over_f_nee. */
template <typename F, uint32_t FEAT>
VP_D void env_escape_main(const SceneView<F> &sv, PathCold<F> &c, uint32_t fl, F w, F w_nee, uint32_t depth) {
    const DScene<F> &g = *sv.g;
    if (g.environment < 0) return;
    const F emitted = sv.emitters[g.environment].irradiance;
    if (FEAT & FT_MIS) {
        if (depth == 0 && g.hide_emitters) return;
        const bool count_direct = depth == 0 || (fl & QF_SPEC_CHAIN);
        if (!count_direct) mis_update(w_nee, env_pdf(sv), (F) 1);
        c.result += (count_direct ? mis_w1(w) : mis_w2(w, w_nee)) * emitted;
    } else if (fl & QF_SPEC_CHAIN) {
        c.result += w * emitted;
    }
}

/* Scene::ray_intersect for a ray made in this visit, and what follows from the hit without any further draw.
   start_dl: the ray is intersect2 (volpath.cpp:238-252) and the direct-light march starts from the new hit.
   Called from ONE place in the kernel (after the phase bodies), so that ray_intersect is inlined there once: its
   Ray / SurfaceHit / counters then live in registers instead of crossing a call boundary through local memory
   (ncu: what was left of the long_scoreboard stalls after the BVH moved to shared memory). */
template <typename F, uint32_t FEAT>
VP_D void intersect_new_ray(const SceneView<F> &sv, PathSlot<F> &s, PathCold<F> &c, const Ray<F> &ir, bool isect_main,
                            bool start_dl, uint32_t &fl, int &state, BvhWork &work) {
    SurfaceHit<F> h;
    PathCounters cnt;
    cnt.n_intersect = 0;
    /* BVH work is counted for meshes whose BVH lives in HBM (its memory roofline, bench.py); the few-triangle
       meshes staged in shared memory are not worth the registers: counting them cost the C2 instance 8 % */
    ray_intersect_impl<F, FEAT, false>(sv, ir, h, cnt, (FEAT & FT_MESH) ? &work : nullptr);
    s.c_isect += 1;
    s.cur.hit_t = h.t; st3(s.cur.hit_p, h.p); st3(s.cur.hit_n, h.n); s.cur.hit_shape = h.shape;
    if (isect_main) { /* the main path will want the whole SurfaceInteraction if it scatters here: keep what
                         surface_frame() needs to finish it */
        c.prim = h.prim; c.bu = h.bu; c.bv = h.bv;
    }
    if (start_dl) {
        /* evaluate_direct_light prologue (volpath.cpp:370-383): march along the BSDF ray from the new hit; the
           main path (same ray, new hit, throughput) waits in `saved` */
        copy_ctx(c.saved, s.cur);
        s.cur.w = 1;
        c.emitter_val = 0;
        fl = (fl & ~(QF_MODE | QF_CONT | QF_TOP_DONE)) | ((uint32_t) QM_DL << QF_MODE_SHIFT) |
             ((uint32_t) QC_AFTER_DL << QF_CONT_SHIFT);
    }
    if (s.cur.medium < 0) {
        /* outside every medium the step that follows is trivial -- no draw, no collision, just "did the ray hit
           something" (volpath.cpp:178-190, 336-345, 417-425) -- so take it here */
        const int mode = (int) ((fl & QF_MODE) >> QF_MODE_SHIFT);
        const bool valid = h.t != inf_<F>();
        if (mode == QM_MAIN) {
            state = valid ? Q_SURFACE : Q_PATH_END;
            if ((FEAT & FT_ENV) && !valid) env_escape_main<F, FEAT>(sv, c, fl, s.cur.w, s.w2, s.depth);
        } else {
            if (mode == QM_NEE) s.total_dist += h.t;
            state = valid ? Q_MARCH_SURF : Q_MARCH_END;
            /* the direct-light march reached the environment (volpath.cpp:429-438) */
            if ((FEAT & FT_ENV) && mode == QM_DL && !valid && sv.g->environment >= 0)
                c.emitter_val = sv.emitters[sv.g->environment].irradiance;
        }
        fl &= ~QF_TOP_DONE;
    }
}

/* pixel and image block of stream k, without seeding a generator (see make_stream) */
VP_D void stream_pixel(const DStreams &sp, uint64_t k, int &px, int &py, int &box, int &boy, int &bsx, int &bsy,
                       bool &empty) {
    empty = false;
    if (sp.mode == SEED_SCALAR_BLOCKS) {
        uint64_t bs2 = (uint64_t) sp.bs * sp.bs;
        uint64_t r = k % ((uint64_t) sp.n_blocks * bs2);
        uint64_t counter = r / bs2;
        uint32_t i = (uint32_t) (r % bs2);
        SpiralBlockD b = sp.blocks[counter];
        uint32_t x = compact_bits(i), y = compact_bits(i >> 1);
        empty = (int) x >= b.sx || (int) y >= b.sy;
        px = b.ox + (int) x; py = b.oy + (int) y;
        box = b.ox; boy = b.oy; bsx = b.sx; bsy = b.sy;
    } else {
        uint64_t pixel = sp.mode == SEED_WAVEFRONT ? k / sp.spp_pass : (uint64_t) sp.ex_pixel[k];
        px = (int) (pixel % (uint64_t) sp.w) + sp.ox;
        py = (int) (pixel / (uint64_t) sp.w) + sp.oy;
        box = sp.ox; boy = sp.oy; bsx = sp.w; bsy = sp.h;
    }
}

template <typename F, int MAXT, uint32_t FEAT>
__global__ void __launch_bounds__(MAXT, 1)
volpath_q_kernel(const __grid_constant__ DScene<F> g, const __grid_constant__ DStreams sp,
                 const __grid_constant__ RenderOut<F> out) {
    extern __shared__ __align__(128) unsigned char smem[];
    SceneView<F> sv;
    stage_scene(g, out, smem, sv);

    const DSensor<F> &sn = g.sensor;
    const int W = sn.crop_size[0], NCH = g.film_ch;
    F *film = out.film_in_smem ? reinterpret_cast<F *>(smem + out.film_smem_offset) : out.film;
    const uint32_t max_depth = g.max_depth, rr_depth = g.rr_depth;
    const bool ltr = g.ltr != 0;
    constexpr bool MIS = (FEAT & FT_MIS) != 0;   /* this instance runs volpathmis' state machine */
    constexpr bool ENV = (FEAT & FT_ENV) != 0;   /* ... for scenes with an environment emitter */
    const F INF = inf_<F>();
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;

    PathSlot<F> *S = reinterpret_cast<PathSlot<F> *>(smem + out.q_slots_offset);
    PathCold<F> *C = reinterpret_cast<PathCold<F> *>(out.q_cold) + (size_t) blockIdx.x * out.q_n_slots;
    __builtin_assume(__isGlobal(C));   /* LDG / STG instead of generic loads and stores */
    __builtin_assume(__isGlobal(out.film)); __builtin_assume(__isGlobal(out.totals));
    QueueCtl *ctl = reinterpret_cast<QueueCtl *>(smem + out.q_ctl_offset);
    unsigned short *Q = reinterpret_cast<unsigned short *>(ctl + 1); /* [PH_N][cap] rings of slot ids */
    const int P = (int) out.q_n_slots;

    const int cap = (int) out.q_ring_cap, cap_mask = cap - 1;  /* power of two >= P */
    for (int i = threadIdx.x; i < PH_N * cap; i += blockDim.x) Q[i] = (unsigned short) Q_EMPTY;
    __syncthreads();
    for (int i = threadIdx.x; i < P; i += blockDim.x) {
        S[i].flags = Q_NEW;
        Q[PH_END * cap + i] = (unsigned short) i;
    }
    if (threadIdx.x == 0) {
        for (int p = 0; p < PH_N; ++p) { ctl->head[p] = 0; ctl->tail[p] = 0; ctl->avail[p] = 0; }
        ctl->tail[PH_END] = P;
        ctl->avail[PH_END] = P;
        ctl->n_live = P;
        ctl->cur = PH_END;
    }
    __syncthreads();

    /* Per-thread launch totals.  The BVH work counters are kept apart and exist only in the instances that count
       them: as elements 8 and 9 of this array they cost the C2 instance 8 % (221 against 240 Mpaths/s; A/B over four
       builds on one box, profiles/r2_history.md).  Tried instead: shared-memory totals fed by atomics when a stream
       ends -- with wavefront seeding a stream is one path, and six contended 64-bit shared atomics per path cost
       more (196). */
    constexpr int TOT_Q = TOT_BVH_NODES;
    unsigned long long tot[TOT_Q];
#pragma unroll
    for (int i = 0; i < TOT_Q; ++i) tot[i] = 0;
    unsigned long long bvh_nodes = 0, bvh_tris = 0;   /* FT_MESH instances only (dead code elsewhere) */

    int last_phase = PH_END;
    for (;;) {
        /* ---------------- pop a batch: prefer the phase run last, else the fullest queue ---------------- */
        int phase = -1, n = 0, h = 0, live = 1;
        for (int attempt = 0; attempt < 4 && phase < 0; ++attempt) {
            /* lanes 0..4 look at one queue each; a max-reduction picks the fullest (or the sticky one) */
            const int cur_phase = out.q_cta_phase ? *((volatile int *) &ctl->cur) : last_phase;
            const int c = lane < PH_N ? ((volatile int *) ctl->avail)[lane] : 0;
            int key = lane < PH_N ? ((c << 3) | lane) : 0;
            if (lane == cur_phase && c >= 32) key |= 1 << 29;
            /* when the CTA has to move on, full batches of the ray-making phases go before STEP: they run back
               to back (they share ray_intersect's code) and STEP then runs long on everything they fed it */
            if (out.q_cta_phase == 2 && lane != PH_STEP && lane < PH_N && c >= 32) key |= 1 << 28;
            const int top = __reduce_max_sync(FULL, key);
            const int ph = top & 7, cnt = (top & ~(3 << 28)) >> 3;
            if (cnt == 0) break;
            const int take = cnt < 32 ? cnt : 32;
            int ok = 0;
            if (lane == 0 && atomicCAS(&ctl->avail[ph], cnt, cnt - take) == cnt) {
                ok = 1;
                h = atomicAdd(&ctl->head[ph], take);
            }
            ok = __shfl_sync(FULL, ok, 0);
            if (ok) {
                phase = ph; n = take; h = __shfl_sync(FULL, h, 0);
                if (lane == 0 && ph != cur_phase) {
                    *((volatile int *) &ctl->cur) = ph;
                    if (out.q_stats) atomicAdd(out.q_stats + 16, 1ull);
                }
                if (lane == 0 && out.q_stats) {
                    atomicAdd(out.q_stats + ph, 1ull);
                    atomicAdd(out.q_stats + 8 + ph, (unsigned long long) take);
                }
            }
        }
        if (phase < 0) live = *((volatile int *) &ctl->n_live);
        if (phase < 0) {
            if (live == 0) break;
            if (lane == 0 && out.q_stats) atomicAdd(out.q_stats + 17, 1ull);
            __nanosleep(100);
            continue;
        }
        int slot = -1;
        if (lane < n) {
            volatile unsigned short *e = Q + phase * cap + ((h + lane) & cap_mask);
            unsigned v;
            while ((v = *e) == Q_EMPTY) { }
            *e = (unsigned short) Q_EMPTY;
            slot = (int) v;
        }
        __threadfence_block();
        __syncwarp();
        last_phase = phase;
        const bool has = slot >= 0;
        PathSlot<F> &s = S[has ? slot : 0];
        PathCold<F> &c = C[has ? slot : 0];
        int state = Q_DEAD;
        PathCounters cnt;
        cnt.draws = 0; cnt.medium_steps = 0; cnt.n_null = 0; cnt.n_real = 0; cnt.n_intersect = 0; cnt.n_surface = 0;

        /* Rays are intersected where they are made (scatter, surface, march, end phases all end in one call of the
           shared, out-of-line ray_intersect), together with the part of the next loop iteration that precedes the
           intersection in the reference -- Russian roulette for the main path (volpath.cpp:72-87), the
           remaining-distance test for an emitter march (volpath.cpp:285-291).  QF_TOP_DONE tells STEP that this
           has happened.  Random numbers are drawn in the reference's order (ray_intersect draws none), and it is
           called exactly when the reference calls it, so the per-path counters agree. */
        Ray<F> ir;                 /* ray to intersect at the end of this visit */
        uint32_t fl = 0;           /* flag word of the slot (phases other than STEP) */
        bool do_isect = false;
        bool isect_main = false;   /* it is the main path's ray: keep (prim, u, v) for surface_frame() */
        bool start_dl = false;     /* it is intersect2: the direct-light march starts from the new hit */
        ir.o = mk3<F>(0, 0, 0); ir.d = ir.o; ir.d_rcp = ir.o; ir.mint = 0; ir.maxt = 0;

        if (phase == PH_STEP) {
            /* ================================ STEP ================================ */
            /* A warp keeps its batch for several steps while most of its lanes stay in this phase (null
               collisions come in runs): the pop / push cost is paid once per run, not once per step. */
            /* The lanes that stay keep their path in REGISTERS from one step to the next: the slot is read once when the
               batch is taken and written once when a lane leaves the phase (or the warp gives the batch back) -- not
               once per step (slot ids within a batch are arbitrary, so every slot access is a ~3-way bank conflict). */
            bool mine = has;
            uint32_t fl = mine ? s.flags : (uint32_t) Q_DEAD;
            int st = (int) (fl & QF_STATE);
            const int mode = (int) ((fl & QF_MODE) >> QF_MODE_SHIFT);
            const bool main_mode = mode == QM_MAIN, nee = mode == QM_NEE;   /* a path does not change mode inside STEP */
            Pcg32 rng;
            Ray<F> r;
            int med = -1;
            F w = 0, hit_t = INF;
            F w2 = 0;          /* volpathmis: p_over_f_uni of an emitter march */
            F total_dist = 0;  /* emitter march: distance covered */
            uint32_t acc_steps = 0, acc_null = 0;
            r.o = mk3<F>(0, 0, 0); r.d = r.o; r.d_rcp = r.o; r.mint = 0; r.maxt = 0;
            rng.state = 0; rng.inc = 1; rng.draws = 0;
            if (mine) {
                ld_rng(s, rng);
                ld_ray(s.cur, r);
                med = s.cur.medium;
                w = s.cur.w;
                hit_t = s.cur.hit_t;
                if (MIS && (nee || ENV)) w2 = s.w2;   /* main mode: p_over_f_nee, read where a path meets the environment */
                if (nee) total_dist = s.total_dist;
            }
            for (int it = 1;; ++it) {
            cnt.medium_steps = 0; cnt.n_null = 0;
            F remaining = 0;
            MediumHit<F> mi;
            mi.t = INF; mi.mint = 0; mi.p = mk3<F>(0, 0, 0); mi.sigma_s = 0; mi.sigma_n = 0; mi.sigma_t = 0; mi.combined = 0;
            bool go = mine;
            if (go) {
                if (fl & QF_TOP_DONE) {
                    if (nee) remaining = r.maxt;   /* stored by the visit that made or continued the march ray */
                } else if (main_mode) {
                    /* termination / Russian roulette (volpath.cpp:72-87; volpathmis.cpp:122-133) */
                    const bool on = MIS ? mis_loop_top(rng, w, s.eta, s.depth, rr_depth, max_depth)
                                        : loop_top_rr(rng, w, s.eta, s.depth, rr_depth, max_depth);
                    if (!on) { st = Q_PATH_END; go = false; }
                } else if (nee) {
                    remaining = s.ds_dist * ((F) 1 - K<F>::ShadowEpsilon) - total_dist;
                    r.maxt = remaining;
                    if (!(remaining > 0)) { st = Q_MARCH_END; go = false; }
                }
                fl &= ~QF_TOP_DONE;
            }
            __syncwarp();
            /* one free-flight sample for whichever ray this path is marching (medium.cpp:36-75).  For a
               homogeneous medium the reference shortens ray.maxt to mi.t before it intersects; the hit found
               here over the full range differs from that only when it lies beyond a valid mi.t, and then it is
               never looked at (a homogeneous medium has no null collisions: the path scatters, a march ends). */
            const bool in_medium = go && med >= 0;
            bool mg = false;   /* this step samples against the medium's majorant grid (opt-in fast mode) */
            if (in_medium) {
                const DMedium<F> &m = sv.media[med];
                mg = (FEAT & FT_MAJORANT_GRID) && sv.mg_saddr != 0 && m.mg_offset >= 0;
                if (mg) {
                    sample_interaction_mg<F, FEAT>(sv, m, r, rng.next(F(), true), nee ? min_(hit_t, remaining) : hit_t, mi, cnt);
                } else {
                    sample_interaction<F, FEAT>(sv, m, r, rng.next(F(), true), mi, cnt);
                    if (m.type == 0 && mi.valid())
                        r.maxt = nee ? min_(mi.t, remaining) : mi.t;
                }
            }
            __syncwarp();
            /* surface-vs-collision ordering and the transmittance / pdf ratio
               (Medium::eval_tr_and_pdf, medium.cpp:77-89; volpath.cpp:113-117,300-311,402-404) */
            bool spectral = false;
            F ratio = 1;
            F ff_p = 1, ff_f = 1;   /* volpathmis: free-flight pdf and transmittance, applied as update_weights(.., p, f) */
            if (in_medium) {
                spectral = sv.media[med].spectral != 0;
                if (hit_t < mi.t) mi.t = INF;
                if (spectral && mg) {
                    /* tr / pdf of Medium::eval_tr_and_pdf with a piecewise-constant majorant: the transmittance cancels,
                       1 / mu_cell is left for a collision and 1 when the surface (or the end of the march) comes first */
                    const bool surface_first = hit_t < mi.t || (nee && mi.t > remaining);
                    ratio = surface_first ? (F) 1 : (F) 1 / mi.combined;
                    ff_p = surface_first ? (F) 1 : mi.combined;
                } else if (spectral) {
                    F tt = min_(mi.t, hit_t);
                    if (nee) tt = min_(remaining, tt);
                    F t = tt - mi.mint;
                    F tr = exp_(-t * mi.combined);
                    bool surface_first = hit_t < mi.t || (nee && mi.t > remaining);
                    F pdf = surface_first ? tr : tr * mi.combined;
                    if (MIS) { ff_p = pdf; ff_f = tr; }
                    else ratio = pdf > 0 ? tr / pdf : (F) 0;
                }
            }
            __syncwarp();
            if (MIS && go && main_mode) {
                /* volpathmis.cpp:135-250: free-flight weights, null / real decision */
                bool act_null = false, act_real = false, escaped = false;
                uint32_t depth = s.depth;
                if (in_medium) {
                    if (spectral) mis_update(w, ff_p, ff_f);
                    if (ENV && spectral) mis_update(w2, ff_p, ff_f);
                    escaped = !mi.valid();
                    const bool active_medium = mi.valid();
                    F u = rng.next(F(), active_medium);
                    bool null_scatter = u >= mi.sigma_t / mi.combined;
                    act_null = null_scatter && active_medium;
                    act_real = !act_null && active_medium;
                    if (act_real) { depth += 1; s.depth = depth; }
                }
                const bool real_beyond_depth = act_real && !(depth < max_depth);
                act_real = act_real && depth < max_depth;
                if (act_null) {
                    cnt.n_null++;
                    if (spectral) mis_update(w, mi.sigma_n / mi.combined, mi.sigma_n);
                    else mis_update(w, mi.sigma_n, mi.sigma_n);
                    if (ENV) {
                        if (spectral) mis_update(w2, (F) 1, mi.sigma_n);
                        else mis_update(w2, (F) 1, mi.sigma_n / mi.combined);
                    }
                    r.o = mi.p;
                    r.mint = 0;
                    hit_t = hit_t - mi.t;
                } else if (act_real) {
                    if (spectral) mis_update(w, mi.sigma_t / mi.combined, mi.sigma_s);
                    else mis_update(w, mi.sigma_t, mi.sigma_s);
                    r.o = mi.p;
                    st = Q_SCATTER;
                } else if (!real_beyond_depth && (!in_medium || escaped) && hit_t != INF) {
                    st = Q_SURFACE;
                } else {
                    st = Q_PATH_END;
                    if (ENV && (!in_medium || escaped) && hit_t == INF) env_escape_main<F, FEAT>(sv, c, fl, w, w2, depth);
                }
            } else if (MIS && go) {
                /* sample_emitter of volpathmis (volpathmis.cpp:362-442): two weights along the march */
                bool escaped = false, active_medium = false;
                if (in_medium) {
                    if (spectral) { mis_update(w, ff_p, ff_f); mis_update(w2, ff_p, ff_f); }
                    if (mi.t > remaining && mi.valid()) total_dist = s.ds_dist;
                    if (mi.t > remaining) mi.t = INF;
                    escaped = !mi.valid();
                    active_medium = mi.valid();
                    if (active_medium) {
                        total_dist += mi.t;
                        cnt.n_null++;
                        r.o = mi.p;
                        r.mint = 0;
                        hit_t = hit_t - mi.t;
                        if (spectral) { mis_update(w, (F) 1, mi.sigma_n); mis_update(w2, mi.sigma_n / mi.combined, mi.sigma_n); }
                        else { mis_update(w, (F) 1, mi.sigma_n / mi.combined); mis_update(w2, mi.sigma_n, mi.sigma_n); }
                    }
                }
                bool active_surface = !in_medium || escaped;
                if (active_surface) total_dist += hit_t;
                active_surface = active_surface && hit_t != INF && !active_medium;
                r.maxt = remaining;
                if (active_surface) st = Q_MARCH_SURF;
                else if (active_medium && (g.mis != 2 ? mis_w1(w2) != 0 : (w2 != 0 || w != 0))) st = Q_STEP;
                else st = Q_MARCH_END;
            } else if (go && main_mode) {
                bool act_null = false, act_real = false, escaped = false;
                uint32_t depth = s.depth;
                if (in_medium) {
                    if (spectral) w *= ratio;
                    escaped = !mi.valid();
                    const bool active_medium = mi.valid();
                    F u = rng.next(F(), active_medium);
                    bool null_scatter = u >= mi.sigma_t / mi.combined;
                    act_null = null_scatter && active_medium;
                    act_real = !act_null && active_medium;
                    if (spectral && act_null) w *= mi.sigma_n * mi.combined / mi.sigma_n;
                    if (act_real) { depth += 1; s.depth = depth; }
                }
                const bool active = depth < max_depth;
                act_real = act_real && active;
                if (act_null) { /* volpath.cpp:140-144 */
                    cnt.n_null++;
                    r.o = mi.p;
                    r.mint = 0;
                    hit_t = hit_t - mi.t;
                } else if (act_real) {
                    /* real collision: the albedo weight (volpath.cpp:148-152); the scattering event itself
                       needs only the position, which becomes the ray origin */
                    if (spectral) w *= mi.sigma_s * mi.combined / mi.sigma_t;
                    else w *= mi.sigma_s / mi.sigma_t;
                    r.o = mi.p;
                    st = Q_SCATTER;
                } else if ((!in_medium || escaped) && hit_t != INF) {
                    st = Q_SURFACE;
                } else {
                    st = Q_PATH_END;
                    if (ENV && (!in_medium || escaped) && hit_t == INF) env_escape_main<F, FEAT>(sv, c, fl, w, w, depth);
                }
            } else if (go) {
                /* sample_emitter (volpath.cpp:285-365) / evaluate_direct_light (volpath.cpp:384-463) */
                bool escaped = false, active_medium = false;
                if (in_medium) {
                    if (spectral) w *= ratio;
                    if (nee) {
                        if (mi.t > remaining && mi.valid()) total_dist = s.ds_dist;
                        if (mi.t > remaining) mi.t = INF;
                    }
                    escaped = !mi.valid();
                    active_medium = mi.valid();
                    if (active_medium) {
                        if (nee) total_dist += mi.t;
                        cnt.n_null++;
                        r.o = mi.p;
                        r.mint = 0;
                        hit_t = hit_t - mi.t;
                        if (spectral) w *= mi.sigma_n;
                        else w *= mi.sigma_n / mi.combined;
                    }
                }
                bool active_surface = !in_medium || escaped;
                if (nee && active_surface) total_dist += hit_t;
                active_surface = active_surface && hit_t != INF && !active_medium;
                if (nee) r.maxt = remaining;
                /* a surface on the march (null.cpp:69-72 lets it through, anything else blocks it) and the
                   end of the march are rare: they run as their own phase, in full batches */
                if (active_surface) st = Q_MARCH_SURF;
                else if (active_medium && w != 0) st = Q_STEP;
                else {
                    st = Q_MARCH_END;
                    /* the direct-light march reached the environment (volpath.cpp:429-438) */
                    if (ENV && !nee && !active_medium && hit_t == INF && g.environment >= 0)
                        c.emitter_val = sv.emitters[g.environment].irradiance;
                }
            }
            acc_steps += cnt.medium_steps; acc_null += cnt.n_null;
            const bool staying = mine && st == Q_STEP;
            const int stay = __popc(__ballot_sync(FULL, staying));
            const bool give_back = stay < (int) out.q_stay_min || it >= (int) out.q_max_iter;
            if (mine && (!staying || give_back)) {   /* this lane's path leaves the warp: back into its slot */
                st_rng(s, rng);
                s.c_steps += acc_steps; s.c_null += acc_null;
                st3(s.cur.o, r.o); s.cur.mint = r.mint; s.cur.maxt = r.maxt;
                s.cur.w = w;
                if (MIS && (nee || ENV)) s.w2 = w2;
                if (nee) s.total_dist = total_dist;
                s.cur.hit_t = hit_t;
                s.flags = (fl & ~QF_STATE) | (uint32_t) st;
                state = st;
            }
            mine = staying;
            if (give_back) break;
            }
        } else if (phase == PH_SCATTER) {
            /* =============================== SCATTER ============================== */
            if (has) {
                fl = s.flags;
                state = (int) (fl & QF_STATE);
                Pcg32 rng;
                ld_rng(s, rng);
                const int med = s.cur.medium;
                const DMedium<F> &m = sv.media[med];
                const V3<F> d = ld3(s.cur.d);
                const V3<F> p = ld3(s.cur.o); /* the scattering position (STEP left it as the ray origin) */
                if (MIS && state == Q_SCATTER) { /* volpathmis.cpp:201-218, sample_emitter prologue :330-349 */
                    c.c_real += 1;
                    fl |= QF_VALID_RAY;
                    state = Q_SCATTER_B;
                    if (ENV && m.sample_emitters) fl &= ~QF_SPEC_CHAIN;   /* volpathmis.cpp:171 */
                    if (m.sample_emitters) {
                        F ev;
                        DirSample<F> ds = queue_sample_emitter<F, FEAT>(sv, rng, ev);
                        F nee = s.cur.w, uni = s.cur.w;
                        if (ds.pdf != 0) mis_update(nee, ds.pdf, (F) 1);
                        /* what the caller applies to the march's two weights when it is over (volpathmis.cpp:211-216) */
                        const F phase_val = phase_eval_at<F, FEAT>(sv, sv.phases[m.phase], p, -d, ds.d);
                        const F phase_p = ds.delta ? (F) 0 : phase_val;
                        if (ds.pdf != 0) {
                            copy_ctx(c.saved, s.cur);
                            Ray<F> r2 = spawn_ray(p, ds.d);
                            r2.mint = 0;
                            s.total_dist = 0; s.ds_dist = ds.dist; c.emitter_val = ev; c.nee_w = phase_val; c.bu = phase_p;
                            fl = (fl & ~(QF_MODE | QF_CONT)) | ((uint32_t) QM_NEE << QF_MODE_SHIFT) |
                                 ((uint32_t) QC_SCATTER_B << QF_CONT_SHIFT);
                            const F remaining = ds.dist * ((F) 1 - K<F>::ShadowEpsilon) - (F) 0;
                            r2.maxt = remaining;
                            st_ray(s.cur, r2);
                            s.cur.hit_t = INF; s.cur.hit_shape = -1;
                            s.cur.w = nee; s.w2 = uni;
                            if (remaining > 0) { state = Q_STEP; fl |= QF_TOP_DONE; do_isect = true; ir = r2; }
                            else state = Q_MARCH_END;
                        } else {
                            mis_update(nee, (F) 1, phase_val);
                            mis_update(uni, phase_p, phase_val);
                            c.result += mis_w2(nee, uni) * ev;
                        }
                    }
                }
                if (MIS && state == Q_SCATTER_B) { /* phase function sampling (volpathmis.cpp:220-236) */
                    const DPhase<F> &phase_fn = sv.phases[m.phase];
                    F da = rng.next(F(), true), db = rng.next(F(), true), dc = rng.next(F(), true);
                    F s1 = ltr ? da : dc, s2x = ltr ? db : da, s2y = ltr ? dc : db;
                    F phase_pdf;   /* what phase->sample returns with the direction (volpathmis.cpp:229) */
                    V3<F> wo = phase_sample_at<F, FEAT>(sv, phase_fn, p, d, s1, s2x, s2y, &phase_pdf);
                    Ray<F> r = spawn_ray(p, wo);
                    r.mint = 0;
                    st_ray(s.cur, r);
                    F w = s.cur.w;
                    if (ENV) {   /* p_over_f_nee restarts from the scattering event (volpathmis.cpp:219, 234) */
                        F w_nee = w;
                        mis_update(w_nee, (F) 1, phase_pdf);
                        s.w2 = w_nee;
                    }
                    mis_update(w, phase_pdf, phase_pdf);
                    /* top of the next loop iteration */
                    if (mis_loop_top(rng, w, s.eta, s.depth, rr_depth, max_depth)) {
                        state = Q_STEP; fl |= QF_TOP_DONE; do_isect = true; isect_main = true; ir = r;
                    } else {
                        state = Q_PATH_END;
                    }
                    s.cur.w = w;
                }
                if (!MIS && state == Q_SCATTER) { /* volpath.cpp:146-168 */
                    c.c_real += 1;
                    const F throughput = s.cur.w;
                    fl |= QF_VALID_RAY;
                    state = Q_SCATTER_B;
                    if (ENV) fl = m.sample_emitters ? (fl & ~QF_SPEC_CHAIN) : (fl | QF_SPEC_CHAIN);   /* volpath.cpp:158-159 */
                    if (m.sample_emitters) { /* sample_emitter prologue (volpath.cpp:261-283) */
                        F ev;
                        DirSample<F> ds = queue_sample_emitter<F, FEAT>(sv, rng, ev);
                        F nee_w = throughput * phase_eval_at<F, FEAT>(sv, sv.phases[m.phase], p, -d, ds.d);
                        if (ds.pdf != 0) {
                            copy_ctx(c.saved, s.cur);
                            Ray<F> r2 = spawn_ray(p, ds.d);
                            r2.mint = 0;
                            s.total_dist = 0; s.ds_dist = ds.dist; c.emitter_val = ev; c.nee_w = nee_w;
                            fl = (fl & ~(QF_MODE | QF_CONT)) | ((uint32_t) QM_NEE << QF_MODE_SHIFT) |
                                 ((uint32_t) QC_SCATTER_B << QF_CONT_SHIFT);
                            /* top of the march's first iteration (volpath.cpp:285-291) */
                            const F remaining = ds.dist * ((F) 1 - K<F>::ShadowEpsilon) - (F) 0;
                            r2.maxt = remaining;
                            st_ray(s.cur, r2);
                            s.cur.hit_t = INF; s.cur.hit_shape = -1;
                            s.cur.w = 1;
                            if (remaining > 0) { state = Q_STEP; fl |= QF_TOP_DONE; do_isect = true; ir = r2; }
                            else state = Q_MARCH_END;
                        } else {
                            c.result += nee_w * ev;
                        }
                    }
                }
                /* no emitter march was started: sample the phase function in the same visit */
                if (!MIS && state == Q_SCATTER_B) { /* phase function sampling (volpath.cpp:170-176) */
                    const DPhase<F> &phase_fn = sv.phases[m.phase];
                    F da = rng.next(F(), true), db = rng.next(F(), true), dc = rng.next(F(), true);
                    F s1 = ltr ? da : dc, s2x = ltr ? db : da, s2y = ltr ? dc : db;
                    V3<F> wo = phase_sample_at<F, FEAT>(sv, phase_fn, p, d, s1, s2x, s2y);
                    Ray<F> r = spawn_ray(p, wo);
                    r.mint = 0;
                    st_ray(s.cur, r);
                    /* top of the next loop iteration */
                    F w = s.cur.w;
                    if (loop_top_rr(rng, w, s.eta, s.depth, rr_depth, max_depth)) {
                        state = Q_STEP; fl |= QF_TOP_DONE; do_isect = true; isect_main = true; ir = r;
                    } else {
                        state = Q_PATH_END;
                    }
                    s.cur.w = w;
                }
                st_rng(s, rng);
            }
        } else if (phase == PH_SURFACE) {
            /* =============================== SURFACE ============================== */
            if (has) {
                fl = s.flags;
                state = (int) (fl & QF_STATE);
                Pcg32 rng;
                ld_rng(s, rng);
                const DShape<F> &sh = sv.shapes[s.cur.hit_shape];
                const DBsdf<F> &bsdf = sv.bsdfs[sh.bsdf];
                const V3<F> sp_ = ld3(s.cur.hit_p);
                V3<F> wi;
                Frame<F> fr;
                if (state == Q_SURFACE) {
                    /* finish the SurfaceInteraction of the hit (interaction.h:571-596) */
                    SurfaceHit<F> h;
                    h.shape = s.cur.hit_shape; h.prim = c.prim; h.bu = c.bu; h.bv = c.bv;
                    h.p = sp_; h.n = ld3(s.cur.hit_n);
                    surface_frame<F, FEAT>(sv, ld3(s.cur.d), h);
                    wi = h.wi; fr = h.sh;
                    st3(c.wi, wi); st3(c.sh_s, fr.s); st3(c.sh_t, fr.t); st3(c.sh_n, fr.n);
                } else {
                    wi = ld3(c.wi);
                    fr.s = ld3(c.sh_s); fr.t = ld3(c.sh_t); fr.n = ld3(c.sh_n);
                }
                F throughput = s.cur.w;
                uint32_t depth = s.depth;
                if (MIS && state == Q_SURFACE) { /* volpathmis.cpp:270-290 */
                    c.c_surf += 1;
                    const bool active_e = (bsdf_flags(bsdf.type) & BF_Smooth) != 0 && (depth + 1 < max_depth);
                    state = Q_SURF_B;
                    if (active_e) {
                        F ev;
                        DirSample<F> ds = queue_sample_emitter<F, FEAT>(sv, rng, ev);
                        F nee = throughput, uni = throughput;   /* p_over_f */
                        if (ds.pdf != 0) mis_update(nee, ds.pdf, (F) 1);
                        const V3<F> wo = fr.to_local(ds.d);
                        const F bsdf_val = bsdf_eval(bsdf, wi, wo);
                        const F bsdf_p = ds.delta ? (F) 0 : bsdf_pdf(bsdf, wi, wo);
                        if (ds.pdf != 0) {
                            copy_ctx(c.saved, s.cur);
                            Ray<F> r2 = spawn_ray(sp_, ds.d);
                            s.total_dist = 0; s.ds_dist = ds.dist; c.emitter_val = ev; c.nee_w = bsdf_val; c.bu = bsdf_p;
                            fl = (fl & ~(QF_MODE | QF_CONT)) | ((uint32_t) QM_NEE << QF_MODE_SHIFT) |
                                 ((uint32_t) QC_SURF_B << QF_CONT_SHIFT);
                            const F remaining = ds.dist * ((F) 1 - K<F>::ShadowEpsilon) - (F) 0;
                            r2.maxt = remaining;
                            st_ray(s.cur, r2);
                            s.cur.hit_t = INF; s.cur.hit_shape = -1;
                            s.cur.w = nee; s.w2 = uni;
                            if (remaining > 0) { state = Q_STEP; fl |= QF_TOP_DONE; do_isect = true; ir = r2; }
                            else state = Q_MARCH_END;
                        } else {
                            mis_update(nee, (F) 1, bsdf_val);
                            mis_update(uni, bsdf_p, bsdf_val);
                            c.result += mis_w2(nee, uni) * ev;
                        }
                    }
                }
                if (MIS && state == Q_SURF_B) { /* BSDF sampling (volpathmis.cpp:292-322) */
                    F da = rng.next(F(), true), db = rng.next(F(), true), dc = rng.next(F(), true);
                    F s1 = ltr ? da : dc, s2x = ltr ? db : da, s2y = ltr ? dc : db;
                    V3<F> bwo;
                    F bs_pdf, bs_eta, bsdf_weight;
                    uint32_t sampled_type;
                    bsdf_sample(bsdf, wi, s1, s2x, s2y, bwo, bs_pdf, bs_eta, sampled_type, bsdf_weight);
                    const bool invalid_sample = bs_pdf == 0;
                    const bool alive = bs_pdf > 0;
                    const bool non_null = alive && !(sampled_type & BF_Null);
                    if (non_null || invalid_sample) fl |= QF_VALID_RAY;
                    if (ENV) {   /* volpathmis.cpp:311-312 */
                        if (non_null && (sampled_type & BF_Delta)) fl |= QF_SPEC_CHAIN;
                        if (alive && (sampled_type & BF_Smooth)) fl &= ~QF_SPEC_CHAIN;
                    }
                    state = Q_PATH_END;
                    if (alive) {
                        s.eta *= bs_eta;
                        Ray<F> r = spawn_ray(sp_, fr.to_world(bwo));
                        st_ray(s.cur, r);
                        if (non_null) { depth += 1; s.depth = depth; }
                        if (ENV) {   /* p_over_f_nee (volpathmis.cpp:314-317); a null BSDF leaves it as it is */
                            F w_nee = non_null ? throughput : s.w2;
                            if (non_null) mis_update(w_nee, (F) 1, bsdf_weight * bs_pdf);
                            s.w2 = w_nee;
                        }
                        mis_update(throughput, bs_pdf, bsdf_weight * bs_pdf);
                        if (sh.interior >= 0 || sh.exterior >= 0)
                            s.cur.medium = dot(r.d, ld3(s.cur.hit_n)) > 0 ? sh.exterior : sh.interior;
                        /* top of the next loop iteration */
                        if (mis_loop_top(rng, throughput, s.eta, depth, rr_depth, max_depth)) {
                            state = Q_STEP; fl |= QF_TOP_DONE; do_isect = true; isect_main = true; ir = r;
                        }
                    }
                    s.cur.w = throughput;
                }
                if (!MIS && state == Q_SURFACE) { /* volpath.cpp:192-212 */
                    c.c_surf += 1;
                    bool active_e = (bsdf_flags(bsdf.type) & BF_Smooth) != 0 && (depth + 1 < max_depth);
                    state = Q_SURF_B;
                    if (active_e) { /* sample_emitter prologue (volpath.cpp:261-283) */
                        F ev;
                        DirSample<F> ds = queue_sample_emitter<F, FEAT>(sv, rng, ev);
                        V3<F> wo = fr.to_local(ds.d);
                        F bsdf_val = bsdf_eval(bsdf, wi, wo);
                        F bpdf = bsdf_pdf(bsdf, wi, wo);
                        F pa = ds.pdf, pb = ds.delta ? (F) 0 : bpdf;
                        pa *= pa; pb *= pb;
                        F mis = pa > 0 ? pa / (pa + pb) : (F) 0; /* mis_weight (volpath.cpp:479-483) */
                        F nee_w = throughput * bsdf_val * mis;
                        if (ds.pdf != 0) {
                            copy_ctx(c.saved, s.cur);
                            Ray<F> r2 = spawn_ray(sp_, ds.d);
                            s.total_dist = 0; s.ds_dist = ds.dist; c.emitter_val = ev; c.nee_w = nee_w;
                            fl = (fl & ~(QF_MODE | QF_CONT)) | ((uint32_t) QM_NEE << QF_MODE_SHIFT) |
                                 ((uint32_t) QC_SURF_B << QF_CONT_SHIFT);
                            const F remaining = ds.dist * ((F) 1 - K<F>::ShadowEpsilon) - (F) 0;
                            r2.maxt = remaining;
                            st_ray(s.cur, r2);
                            s.cur.hit_t = INF; s.cur.hit_shape = -1;
                            s.cur.w = 1;
                            if (remaining > 0) { state = Q_STEP; fl |= QF_TOP_DONE; do_isect = true; ir = r2; }
                            else state = Q_MARCH_END;
                        } else {
                            c.result += nee_w * ev;
                        }
                    }
                }
                /* no emitter march was started (null BSDF, or nothing to sample): go on in the same visit */
                if (!MIS && state == Q_SURF_B) { /* BSDF sampling (volpath.cpp:214-241) */
                    F da = rng.next(F(), true), db = rng.next(F(), true), dc = rng.next(F(), true);
                    F s1 = ltr ? da : dc, s2x = ltr ? db : da, s2y = ltr ? dc : db;
                    V3<F> bwo;
                    F bs_pdf, bs_eta, bsdf_val;
                    uint32_t sampled_type;
                    bsdf_sample(bsdf, wi, s1, s2x, s2y, bwo, bs_pdf, bs_eta, sampled_type, bsdf_val);
                    throughput *= bsdf_val;
                    s.eta *= bs_eta;
                    Ray<F> r = spawn_ray(sp_, fr.to_world(bwo));
                    st_ray(s.cur, r);
                    bool non_null = !(sampled_type & BF_Null);
                    if (non_null) { depth += 1; s.depth = depth; fl |= QF_VALID_RAY; }
                    if (ENV) {   /* volpath.cpp:228-229 */
                        if (non_null && (sampled_type & BF_Delta)) fl |= QF_SPEC_CHAIN;
                        if (sampled_type & BF_Smooth) fl &= ~QF_SPEC_CHAIN;
                    }
                    bool add_emitter = !(sampled_type & BF_Delta) && throughput != 0 && (depth < max_depth);
                    /* medium switch (volpath.cpp:249-250), applied after the direct-light march */
                    const int medium = s.cur.medium;
                    const int medium_next = (sh.interior >= 0 || sh.exterior >= 0)
                                                ? (dot(r.d, ld3(s.cur.hit_n)) > 0 ? sh.exterior : sh.interior) : medium;
                    if (add_emitter) { /* intersect2: the new hit replaces si (volpath.cpp:238-252) */
                        c.medium_next = medium_next;
                        if (ENV) {   /* mis_weight(bs.pdf, emitter_pdf) for the emitter the march can reach (volpath.cpp:244) */
                            F pa = bs_pdf, pb = env_pdf(sv);
                            pa *= pa; pb *= pb;
                            c.nee_w = pa > 0 ? pa / (pa + pb) : (F) 0;
                        }
                        do_isect = true; isect_main = true; start_dl = true; ir = r;
                        state = Q_STEP;
                    } else {
                        s.cur.medium = medium_next;
                        /* top of the next loop iteration */
                        if (loop_top_rr(rng, throughput, s.eta, depth, rr_depth, max_depth)) {
                            state = Q_STEP; fl |= QF_TOP_DONE; do_isect = true; isect_main = true; ir = r;
                        } else {
                            state = Q_PATH_END;
                        }
                    }
                    s.cur.w = throughput;
                }
                st_rng(s, rng);
            }
        } else if (phase == PH_MARCH) {
            /* ================================ MARCH =============================== */
            /* first half: a march crosses a surface; the second half (a march is over) follows the intersection below */
            if (has) {
                fl = s.flags;
                state = (int) (fl & QF_STATE);
                if (state == Q_MARCH_SURF) { /* volpath.cpp:329-352, 427-450 */
                    const bool nee = (int) ((fl & QF_MODE) >> QF_MODE_SHIFT) == QM_NEE;
                    F w = s.cur.w;
                    const DShape<F> &sh = sv.shapes[s.cur.hit_shape];
                    /* eval_null_transmission: 1 for `null`, 0 otherwise (null.cpp:69-72, bsdf.cpp:11-14) */
                    const F null_tr = sv.bsdfs[sh.bsdf].type == 1 ? (F) 1 : (F) 0;
                    bool on;
                    if (MIS) {   /* volpathmis.cpp:415-440: both weights; the march goes on while they say so */
                        F w2 = s.w2;
                        mis_update(w, (F) 1, null_tr);
                        mis_update(w2, (F) 1, null_tr);
                        s.w2 = w2;
                        on = g.mis != 2 ? mis_w1(w2) != 0 : (w2 != 0 || w != 0);
                    } else {
                        w *= null_tr;
                        on = w != 0;
                    }
                    Ray<F> r = spawn_ray(ld3(s.cur.hit_p), ld3(s.cur.d));
                    if (sh.interior >= 0 || sh.exterior >= 0)
                        s.cur.medium = dot(r.d, ld3(s.cur.hit_n)) > 0 ? sh.exterior : sh.interior;
                    s.cur.w = w;
                    state = Q_MARCH_END;
                    if (on) {
                        /* top of the march's next iteration */
                        bool cont = true;
                        if (nee) {
                            const F remaining = s.ds_dist * ((F) 1 - K<F>::ShadowEpsilon) - s.total_dist;
                            r.maxt = remaining;
                            cont = remaining > 0;
                        }
                        /* the new ray is intersected at once (below); outside the medium that settles what the march
                           does next, often that it is over (the ray left the atmosphere towards the sun) */
                        if (cont) { state = Q_STEP; fl |= QF_TOP_DONE; do_isect = true; ir = r; }
                    }
                    st_ray(s.cur, r);
                }
            }
        } else {
            /* ================================= END ================================ */
            /* the accumulator of a finished stream, splatted after the per-lane work by the whole warp together */
            int splat_pix = -1;
            F splat0 = 0, splat3 = 0, splat4 = 0;
            if (has) {
                fl = s.flags;
                state = (int) (fl & QF_STATE);
                const uint32_t n_samples = sp.samples_per_stream;
                unsigned long long k = c.k;
                uint32_t j = c.j;
                int px = 0, py = 0, box = 0, boy = 0, bsx = 0, bsy = 0;
                bool empty = true;
                if (fl & QF_HAVE_STREAM) stream_pixel(sp, stream_index(sp, k), px, py, box, boy, bsx, bsy, empty);
                if (state == Q_PATH_END) {
                    /* render_sample tail (integrator.cpp:267-287) + ImageBlock::put (imageblock.cpp:85-169) */
                    F value = c.ray_weight * c.result;
                    tot[TOT_PATHS]++;
                    if (out.out_radiance) out.out_radiance[k * n_samples + j] = value;
                    F alpha = (fl & QF_VALID_RAY) ? (F) 1 : (F) 0;
                    bool ok = NCH > 5 || (value >= (F) -1e-5f && finite_(value)); /* no check with AOVs */
                    if (!ok) {
                        tot[TOT_INVALID]++;
                    } else {
                        F bx = c.posx - ((F) box + (F) 0.5), by = c.posy - ((F) boy + (F) 0.5);
                        int lx = (int) ceil_(bx - (F) 0.5), ly = (int) ceil_(by - (F) 0.5);
                        if ((FEAT & FT_RFILTER) && sn.rf_taps > 0) {
                            film_put_wide(film, sn, NCH, c.posx, c.posy, box, boy, bsx, bsy, sp.ox, sp.oy, value, alpha, c.result);
                        } else if (!(lx < 0 || ly < 0 || lx >= bsx || ly >= bsy)) {
                            int fx = lx + box - sp.ox, fy = ly + boy - sp.oy;
                            if (NCH > 5) { /* moment.cpp:78-83: the nested integrator's X Y Z (without the ray weight) and squares */
                                F *q = film + ((size_t) fy * W + fx) * NCH + 5;
                                const F raw = c.result, raw2 = raw * raw;
                                film_add(q, raw); film_add(q + 1, raw); film_add(q + 2, raw);
                                film_add(q + 3, raw2); film_add(q + 4, raw2); film_add(q + 5, raw2);
                            }
                            if (fx == px - sp.ox && fy == py - sp.oy) {
                                c.acc0 += value; c.acc3 += alpha; c.acc4 += (F) 1;
                                fl |= QF_ACC_USED;
                            } else {
                                F *p = film + ((size_t) fy * W + fx) * NCH;
                                film_add(p, value); film_add(p + 1, value); film_add(p + 2, value);
                                film_add(p + 3, alpha); film_add(p + 4, (F) 1);
                            }
                        }
                    }
                    j++;
                    state = Q_NEW;
                }
                /* Q_NEW */
                if ((fl & QF_HAVE_STREAM) && (j >= n_samples || empty)) {
                    /* stream finished: splat its accumulator, write the parity outputs */
                    if (fl & QF_ACC_USED) {
                        splat_pix = (py - sp.oy) * W + (px - sp.ox);
                        splat0 = c.acc0; splat3 = c.acc3; splat4 = c.acc4;
                    }
                    if (out.out_state) out.out_state[k] = ((uint64_t) s.rng_state_hi << 32) | s.rng_state_lo;
                    if (out.out_counters) {
                        PathCounters o;
                        o.draws = s.rng_draws; o.medium_steps = s.c_steps; o.n_null = s.c_null; o.n_real = c.c_real;
                        o.n_intersect = s.c_isect; o.n_surface = c.c_surf;
                        out.out_counters[k] = o;
                    }
                    tot[TOT_DRAWS] += s.rng_draws; tot[TOT_STEPS] += s.c_steps; tot[TOT_NULL] += s.c_null;
                    tot[TOT_REAL] += c.c_real; tot[TOT_ISECT] += s.c_isect; tot[TOT_SURF] += c.c_surf;
                    fl &= ~(QF_HAVE_STREAM | QF_ACC_USED);
                }
                if (!(fl & QF_HAVE_STREAM)) {
                    k = atomicAdd(out.work_counter, 1ull);
                    if (k >= sp.count || stop_requested(out)) {   /* should_stop() (integrator.h:143-146) */
                        state = Q_DEAD;
                    } else {
                        StreamInfo st = make_stream(sp, stream_index(sp, k));
                        c.k = k;
                        s.rng_state_lo = (uint32_t) st.state; s.rng_state_hi = (uint32_t) (st.state >> 32);
                        s.rng_inc_lo = (uint32_t) st.inc; s.rng_inc_hi = (uint32_t) (st.inc >> 32);
                        s.rng_draws = 0;
                        s.c_steps = 0; s.c_null = 0; s.c_isect = 0; c.c_real = 0; c.c_surf = 0;
                        c.acc0 = 0; c.acc3 = 0; c.acc4 = 0;
                        j = 0;
                        px = st.px; py = st.py; empty = st.empty;
                        fl |= QF_HAVE_STREAM;
                    }
                }
                c.j = j;
                if (state == Q_NEW && (fl & QF_HAVE_STREAM) && !(j >= n_samples || empty)) {
                    /* render_sample prologue (integrator.cpp:233-264); an empty stream (pixel outside its
                       block) is finished on the next visit */
                    Pcg32 rng;
                    ld_rng(s, rng);
                    F posx = (F) (uint32_t) px + rng.next(F(), true);
                    F posy = (F) (uint32_t) py + rng.next(F(), true);
                    F apx = (F) 0.5, apy = (F) 0.5;
                    if (sn.needs_aperture_sample) { apx = rng.next(F(), true); apy = rng.next(F(), true); }
                    if (sn.has_shutter) (void) rng.next(F(), true);
                    (void) rng.next(F(), true); /* wavelength sample (integrator.cpp:252) */
                    F ax = (posx - (F) sn.crop_offset[0]) / (F) sn.crop_size[0];
                    F ay = (posy - (F) sn.crop_offset[1]) / (F) sn.crop_size[1];
                    Ray<F> ray;
                    c.ray_weight = sensor_sample_ray<F, FEAT>(sv, ax, ay, apx, apy, ray);
                    c.posx = posx; c.posy = posy;
                    st_rng(s, rng);
                    /* VolumetricPathIntegrator::sample prologue (volpath.cpp:45-70) */
                    st_ray(s.cur, ray);
                    s.cur.hit_t = INF; s.cur.hit_shape = -1;
                    s.cur.medium = sn.medium;
                    s.cur.w = 1;
                    c.result = 0; s.eta = 1; s.depth = 0;
                    fl &= ~(QF_MODE | QF_CONT | QF_VALID_RAY | QF_TOP_DONE | QF_SPEC_CHAIN);
                    if (ENV) {   /* volpath.cpp:51-53, volpathmis.cpp:100-104 */
                        if (!g.hide_emitters) fl |= QF_SPEC_CHAIN;
                        if (!g.hide_emitters && g.environment >= 0) fl |= QF_VALID_RAY;
                        if (MIS) s.w2 = 1;
                    }
                    /* top of the first loop iteration */
                    F w0 = 1;
                    if (MIS ? mis_loop_top(rng, w0, (F) 1, 0u, rr_depth, max_depth)
                            : loop_top_rr(rng, w0, (F) 1, 0u, rr_depth, max_depth)) {
                        state = Q_STEP; fl |= QF_TOP_DONE; do_isect = true; isect_main = true; ir = ray;
                    } else {
                        state = Q_PATH_END;
                    }
                    s.cur.w = w0;
                    st_rng(s, rng);
                }
            }
            /* Warp-aggregated film accumulation (north star (5)).  In wavefront seeding consecutive streams share a
               pixel, so the lanes of a batch that finish together mostly splat onto ONE pixel: 32 x 5 atomics on the
               same five addresses.  When every finishing lane of the warp has the same pixel the accumulators are
               summed by a butterfly and one lane adds five numbers; otherwise each lane adds its own. */
            const unsigned splat_mask = __ballot_sync(FULL, splat_pix >= 0);
            if (splat_mask) {
                const int first = __ffs(splat_mask) - 1;
                const int pix0 = __shfl_sync(FULL, splat_pix, first);
                if (__all_sync(FULL, splat_pix < 0 || splat_pix == pix0)) {
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) {
                        splat0 += __shfl_xor_sync(FULL, splat0, o);
                        splat3 += __shfl_xor_sync(FULL, splat3, o);
                        splat4 += __shfl_xor_sync(FULL, splat4, o);
                    }
                    if (lane == first) {
                        F *p = film + (size_t) pix0 * NCH;
                        film_add(p, splat0); film_add(p + 1, splat0); film_add(p + 2, splat0);
                        film_add(p + 3, splat3); film_add(p + 4, splat4);
                    }
                } else if (splat_pix >= 0) {
                    F *p = film + (size_t) splat_pix * NCH;
                    film_add(p, splat0); film_add(p + 1, splat0); film_add(p + 2, splat0);
                    film_add(p + 3, splat3); film_add(p + 4, splat4);
                }
            }
        }

        /* ---------------- the one place where rays are intersected ---------------- */
        if (phase != PH_STEP) {
            if (has && do_isect) {
                BvhWork work;
                work.nodes = 0; work.tris = 0;
                intersect_new_ray<F, FEAT>(sv, s, c, ir, isect_main, start_dl, fl, state, work);
                if (FEAT & FT_MESH) { bvh_nodes += work.nodes; bvh_tris += work.tris; }
            }
            if (phase == PH_MARCH && has && state == Q_MARCH_END) {
                /* a march is over: hand the result back and restore the main path's context */
                const bool nee = (int) ((fl & QF_MODE) >> QF_MODE_SHIFT) == QM_NEE;
                if (MIS) {   /* the caller's half of the emitter sample (volpathmis.cpp:211-216, 282-289) */
                    F nee_end = s.cur.w, uni_end = s.w2;
                    mis_update(nee_end, (F) 1, c.nee_w);
                    mis_update(uni_end, c.bu, c.nee_w);
                    c.result += mis_w2(nee_end, uni_end) * c.emitter_val;
                } else if (nee) c.result += c.nee_w * (s.cur.w * c.emitter_val);
                /* QM_DL: emitter_val = 0 and emitter_pdf = 0 for delta emitters: nothing is added.  With an environment
                   the march may have reached it: emitter_val is then its radiance (0 otherwise: adds nothing) */
                else if (ENV && c.emitter_val != 0) c.result += c.nee_w * c.saved.w * (s.cur.w * c.emitter_val);
                copy_ctx(s.cur, c.saved);
                const int cont = (int) ((fl & QF_CONT) >> QF_CONT_SHIFT);
                if (cont == QC_SCATTER_B) state = Q_SCATTER_B;
                else if (cont == QC_SURF_B) state = Q_SURF_B;
                else { /* QC_AFTER_DL: medium switch (volpath.cpp:249-250); the hit of intersect2 is current */
                    s.cur.medium = c.medium_next;
                    state = Q_STEP;
                }
                fl &= ~(QF_MODE | QF_TOP_DONE);
            }
            if (has) s.flags = (fl & ~QF_STATE) | (uint32_t) state;
        }

        /* ---------------- push every slot onto the queue of its next phase ---------------- */
        const int dest = (has && state != Q_DEAD) ? phase_of(state) : -1;
        const unsigned peers = __match_any_sync(FULL, dest);      /* lanes going to the same queue */
        const int leader = __ffs(peers) - 1;
        const unsigned dead = __ballot_sync(FULL, has && state == Q_DEAD);
        __threadfence_block(); /* slot state before the id becomes visible */
        int t = 0;
        if (lane == leader && dest >= 0) t = atomicAdd(&ctl->tail[dest], __popc(peers));
        t = __shfl_sync(FULL, t, leader);
        if (dest >= 0) {
            volatile unsigned short *e = Q + dest * cap + ((t + __popc(peers & lt_mask)) & cap_mask);
            while (*e != Q_EMPTY) { }
            *e = (unsigned short) slot;
        }
        __threadfence_block();
        __syncwarp();
        if (lane == leader && dest >= 0) atomicAdd(&ctl->avail[dest], __popc(peers));
        if (lane == 0 && dead) atomicSub(&ctl->n_live, __popc(dead));
    }

    /* totals: warp reduce, one atomic per warp and counter */
#pragma unroll
    for (int i = 0; i < TOT_Q; ++i) {
        unsigned long long v = tot[i];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(FULL, v, o);
        if (lane == 0 && v) atomicAdd(out.totals + i, v);
    }
    if (FEAT & FT_MESH) {
        if (bvh_nodes) atomicAdd(out.totals + TOT_BVH_NODES, bvh_nodes);
        if (bvh_tris) atomicAdd(out.totals + TOT_BVH_TRIS, bvh_tris);
    }
    if (out.film_in_smem) {
        __syncthreads();
        const F *f = reinterpret_cast<const F *>(smem + out.film_smem_offset);
        uint32_t n = (uint32_t) W * sn.crop_size[1] * NCH;
        for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
            F v = f[i];
            if (v != (F) 0) film_add(out.film + i, v);
        }
    }
}

} // namespace VP_NS
