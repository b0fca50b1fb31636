/*
 * vp_queue.cuh -- the production render kernel: a CTA-resident wavefront with
 * shared-memory phase queues.
 *
 * A CTA owns a pool of path slots (about twice its thread count) whose state
 * lives in shared memory.  A path is always in exactly one of five phase queues
 * (also shared memory) or in flight in a warp:
 *
 *   STEP     one tentative-collision step (Medium::sample_interaction and its
 *            bookkeeping) -- shared by the main path, the emitter-sampling march
 *            (sample_emitter) and the direct-light march (evaluate_direct_light);
 *   ISECT    Scene::ray_intersect;
 *   SCATTER  real medium scattering: albedo weight, emitter sample, phase sample;
 *   SURFACE  surface event: emitter sample + BSDF eval, BSDF sample;
 *   END      film splat of a finished path and path regeneration.
 *
 * Each warp repeatedly pops up to 32 slots of ONE phase (it prefers the phase it
 * ran last, then the fullest queue), loads the fields that phase needs into
 * registers, runs the phase for all 32 lanes together, stores the fields back
 * and pushes every slot onto the queue of its next phase.  This is the
 * compaction / sorting by event type: warps always execute full, converged
 * batches, and a warp tends to stay inside one phase's code (the kernel is
 * instruction-fetch bound otherwise, see profiles/).
 *
 * Arithmetic and random-number order are exactly those of vp_integrator.cuh (the
 * one-thread-per-path twin, reference src/integrators/volpath.cpp:38-483); only
 * the interleaving between paths changes.
 */
#pragma once
#include "vp_kernels.cuh"

namespace vp {

enum QState : int {
    Q_STEP_A = 0,   /* loop top + free-flight sampling                         */
    Q_STEP_B,       /* rest of the tentative-collision step                    */
    Q_ISECT,        /* Scene::ray_intersect                                    */
    Q_SCATTER,      /* real medium scattering: albedo, start emitter march     */
    Q_SCATTER_B,    /* phase function sampling                                 */
    Q_SURFACE,      /* surface hit: start emitter march                        */
    Q_SURF_B,       /* BSDF sampling, second intersection                      */
    Q_PATH_END,
    Q_NEW,          /* next sample of the stream / next stream                 */
    Q_DEAD
};
enum QPhase : int { PH_STEP = 0, PH_ISECT, PH_SCATTER, PH_SURFACE, PH_END, PH_N };
enum QMode : int { QM_MAIN = 0, QM_NEE = 1, QM_DL = 2 };
enum QCont : int { QC_STEP_B = 0, QC_SCATTER_B, QC_SURF_B, QC_START_DL, QC_AFTER_DL };

VP_D int phase_of(int state) {
    return state <= Q_STEP_B ? PH_STEP : state == Q_ISECT ? PH_ISECT : state <= Q_SCATTER_B ? PH_SCATTER
           : state <= Q_SURF_B ? PH_SURFACE : PH_END;
}

/* flag word of a slot */
enum : uint32_t {
    QF_STATE = 0xfu, QF_MODE_SHIFT = 4, QF_MODE = 0x3u << 4, QF_CONT_SHIFT = 6, QF_CONT = 0x7u << 6,
    QF_ISECT_DL = 1u << 9,   /* the pending intersection is intersect2: start the direct-light march after it */
    QF_NEEDS_ISECT = 1u << 10, QF_VALID_RAY = 1u << 11, QF_IN_MEDIUM = 1u << 12, QF_HAVE_STREAM = 1u << 13,
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

template <typename F> struct PathSlot {
    unsigned long long rng_state, rng_inc, k;
    uint32_t rng_draws, j;
    uint32_t c_steps, c_null, c_real, c_isect, c_surf;
    uint32_t flags;
    MarchCtx<F> cur, saved;
    F wi[3], sh_s[3], sh_t[3], sh_n[3];   /* shading frame of the main path's hit   */
    F result, eta;
    uint32_t depth;
    int medium_next;
    F total_dist, ds_dist, emitter_val, nee_w, remaining;
    F mi_t, mi_mint, mi_p[3], mi_ss, mi_sn, mi_st, mi_comb; /* MediumInteraction carried from A to B */
    F acc0, acc3, acc4, posx, posy, ray_weight;
    uint32_t pad_to_odd;                  /* odd word stride: consecutive slots fall in distinct banks */
};

struct QueueCtl {
    int lock;
    int count[PH_N];
    int n_live;
    int pad;
};

VP_D void q_lock(int *lock) {
    while (atomicCAS(lock, 0, 1) != 0) __nanosleep(20);
    __threadfence_block();
}
VP_D void q_unlock(int *lock) {
    __threadfence_block();
    atomicExch(lock, 0);
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
