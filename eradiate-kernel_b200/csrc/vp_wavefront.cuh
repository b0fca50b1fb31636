/*
 * vp_wavefront.cuh -- the production render kernel: a register-resident
 * wavefront.  Every lane owns one path and runs it as an explicit state machine;
 * each warp schedules itself so that its lanes execute the SAME piece of code
 * together.  Work is cut into five phases:
 *
 *   STEP     one tentative-collision step (Medium::sample_interaction and its
 *            bookkeeping).  The main path, the emitter-sampling march
 *            (sample_emitter) and the direct-light march (evaluate_direct_light)
 *            all share this code, so lanes in different marches stay converged;
 *   ISECT    Scene::ray_intersect;
 *   SCATTER  real medium scattering: albedo weight, emitter sample, phase sample;
 *   SURFACE  surface event: emitter sample + BSDF eval, BSDF sample;
 *   END      film splat of a finished path and path regeneration.
 *
 * Per iteration the warp counts (ballot) how many lanes wait for each phase and
 * runs the one with the most lanes; the others keep their state in registers
 * and wait.  Running one phase per iteration keeps the instruction stream of
 * an iteration compact (the kernel is instruction-fetch bound otherwise, see
 * profiles/) and gives every phase a large active mask.
 *
 * Arithmetic and random-number order are exactly those of vp_integrator.cuh (the
 * one-thread-per-path twin, reference src/integrators/volpath.cpp:38-483); only
 * the interleaving between lanes changes.
 */
#pragma once
#include "vp_kernels.cuh"

namespace VP_NS {

enum LaneState : int {
    L_STEP_A = 0,   /* loop top + free-flight sampling                         */
    L_STEP_B,       /* rest of the tentative-collision step                    */
    L_ISECT,        /* Scene::ray_intersect                                    */
    L_SCATTER,      /* real medium scattering: albedo, start emitter march     */
    L_SCATTER_B,    /* phase function sampling                                 */
    L_SURFACE,      /* surface hit: start emitter march                        */
    L_SURF_B,       /* BSDF sampling, second intersection                      */
    L_PATH_END,
    L_NEW,          /* next sample of the stream / next stream                 */
    L_DONE
};
enum MarchMode : int { M_MAIN = 0, M_NEE = 1, M_DL = 2 };
/* what follows a march / an intersection */
enum Continuation : int { C_STEP_B = 0, C_SCATTER_B, C_SURF_B, C_START_DL, C_AFTER_DL };

/* what a march needs from a surface hit (spawn point, normal, shape) */
template <typename F> struct MarchHit {
    F t;
    V3<F> p, n;
    int shape;
};

template <typename F, int MIN_BLOCKS>
__global__ void __launch_bounds__(256, MIN_BLOCKS)
volpath_wf_kernel(const __grid_constant__ DScene<F> g, const __grid_constant__ DStreams sp,
                  const __grid_constant__ RenderOut<F> out) {
    extern __shared__ __align__(128) unsigned char smem[];
    SceneView<F> sv;
    stage_scene(g, out, smem, sv);

    const DSensor<F> &sn = g.sensor;
    const int W = sn.crop_size[0], C = g.film_ch;
    F *film = out.film_in_smem ? reinterpret_cast<F *>(smem + out.film_smem_offset) : out.film;
    const uint32_t max_depth = g.max_depth, rr_depth = g.rr_depth;
    const bool ltr = g.ltr != 0;
    const F INF = inf_<F>();
    const unsigned FULL = 0xffffffffu;

    unsigned long long tot[TOT_N];
#pragma unroll
    for (int i = 0; i < TOT_N; ++i) tot[i] = 0;

    /* ---- lane state ------------------------------------------------------ */
    int state = L_NEW, mode = M_MAIN, cont = C_STEP_B, isect_ret = C_STEP_B;
    bool isect_march = false;
    /* stream */
    unsigned long long k = 0;
    StreamInfo st;
    st.n_samples = 0; st.empty = true;
    uint32_t j = 0;
    bool have_stream = false, acc_used = false;
    F acc0 = 0, acc3 = 0, acc4 = 0, posx = 0, posy = 0, ray_weight = 0;
    Pcg32 rng;
    rng.state = 0; rng.inc = 1; rng.draws = 0;
    PathCounters cnt;
    cnt.draws = 0; cnt.medium_steps = 0; cnt.n_null = 0; cnt.n_real = 0; cnt.n_intersect = 0; cnt.n_surface = 0;
    /* main path */
    Ray<F> ray;
    ray.o = mk3<F>(0, 0, 0); ray.d = ray.o; ray.d_rcp = ray.o; ray.mint = 0; ray.maxt = 0;
    SurfaceHit<F> si;
    si.t = INF; si.shape = -1; si.p = ray.o; si.n = ray.o; si.wi = ray.o; si.sh.s = ray.o; si.sh.t = ray.o; si.sh.n = ray.o;
    int medium = -1, medium_next = -1;
    F throughput = 1, result = 0, eta = 1;
    uint32_t depth = 0;
    bool needs_isect = true, valid_ray = false, a_in_medium = false;
    /* march (sample_emitter / evaluate_direct_light) */
    Ray<F> ray2 = ray;
    MarchHit<F> si2;
    si2.t = INF; si2.p = ray.o; si2.n = ray.o; si2.shape = -1;
    int medium2 = -1;
    bool needs_isect2 = false;
    F transmittance = 1, total_dist = 0, ds_dist = 0, emitter_val = 0, nee_w = 0, remaining = 0;
    /* free-flight sample shared by step A and step B */
    MediumHit<F> mi;
    mi.t = INF; mi.mint = 0; mi.p = ray.o; mi.sigma_s = 0; mi.sigma_n = 0; mi.sigma_t = 0; mi.combined = 0;

    auto next_1d = [&](bool m) { return rng.next(F(), m); };
    auto mis_weight = [&](F pdf_a, F pdf_b) -> F {
        pdf_a *= pdf_a; pdf_b *= pdf_b;
        return pdf_a > 0 ? pdf_a / (pdf_a + pdf_b) : (F) 0;
    };
    /* evaluate_direct_light prologue (volpath.cpp:370-383): march along the BSDF ray from the new hit */
    auto start_dl = [&]() {
        ray2 = ray;
        medium2 = medium;
        si2.t = si.t; si2.p = si.p; si2.n = si.n; si2.shape = si.shape;
        needs_isect2 = false;
        transmittance = 1;
        emitter_val = 0;
        mode = M_DL;
        cont = C_AFTER_DL;
        state = L_STEP_A;
    };
    /* end of a march: hand the result back to the main path */
    auto march_done = [&]() {
        if (mode == M_NEE) result += nee_w * (transmittance * emitter_val);
        /* M_DL: emitter_val = 0 and emitter_pdf = 0 for delta emitters: nothing is added */
        mode = M_MAIN;
        if (cont == C_SCATTER_B) state = L_SCATTER_B;
        else if (cont == C_SURF_B) state = L_SURF_B;
        else { medium = medium_next; state = L_STEP_A; } /* C_AFTER_DL: medium switch (volpath.cpp:249-250) */
    };
    /* sample_emitter prologue (volpath.cpp:261-283): returns false when no march is needed */
    auto begin_nee = [&](V3<F> ref_p, bool is_medium_interaction, DirSample<F> &ds) -> bool {
        F sx = next_1d(true);
        (void) next_1d(true);
        F ev;
        ds = sample_emitter_direction(sv, sx, ev);
        if (ds.pdf == 0) ev = 0;
        emitter_val = ev;
        transmittance = 1;
        if (ds.pdf == 0) return false;
        ray2 = spawn_ray(ref_p, ds.d);
        if (is_medium_interaction) ray2.mint = 0;
        total_dist = 0;
        ds_dist = ds.dist;
        si2.t = INF; si2.shape = -1;
        needs_isect2 = true;
        medium2 = medium;
        return true;
    };

    for (;;) {
        /* ---- warp scheduler: run the phase most lanes are waiting for ---- */
        const int n_step = __popc(__ballot_sync(FULL, state <= L_STEP_B));
        const int n_isect = __popc(__ballot_sync(FULL, state == L_ISECT));
        const int n_scat = __popc(__ballot_sync(FULL, state == L_SCATTER || state == L_SCATTER_B));
        const int n_surf = __popc(__ballot_sync(FULL, state == L_SURFACE || state == L_SURF_B));
        const int n_end = __popc(__ballot_sync(FULL, state == L_PATH_END || state == L_NEW));
        if ((n_step | n_isect | n_scat | n_surf | n_end) == 0) break;
        int phase = 0, best = n_step;
        if (n_isect > best) { best = n_isect; phase = 1; }
        if (n_surf > best) { best = n_surf; phase = 3; }
        if (n_scat > best) { best = n_scat; phase = 2; }
        if (n_end > best) { best = n_end; phase = 4; }

        if (phase == 0) {
            /* ================================ STEP ================================ */
            const bool do_a = state == L_STEP_A;
            bool go = do_a;
            if (do_a) {
                if (mode == M_MAIN) {
                    /* termination / Russian roulette (volpath.cpp:72-87) */
                    bool active = throughput != 0;
                    F q = min_(throughput * sqr(eta), (F) 0.95f);
                    bool perform_rr = depth > rr_depth;
                    F rr_sample = next_1d(active);
                    active = active && (rr_sample < q || !perform_rr);
                    if (perform_rr) throughput *= rcp_(q);
                    if (!active || depth >= max_depth) { state = L_PATH_END; go = false; }
                } else if (mode == M_NEE) {
                    remaining = ds_dist * ((F) 1 - K<F>::ShadowEpsilon) - total_dist;
                    ray2.maxt = remaining;
                    if (!(remaining > 0)) { march_done(); go = false; }
                }
            }
            __syncwarp();
            /* one free-flight sample for whichever ray this lane is marching */
            const bool main_mode = mode == M_MAIN;
            const int med = main_mode ? medium : medium2;
            if (go) a_in_medium = med >= 0;
            if (go && a_in_medium) {
                const DMedium<F> &m = sv.media[med];
                Ray<F> r;
                r.o = main_mode ? ray.o : ray2.o;
                r.d = main_mode ? ray.d : ray2.d;
                r.d_rcp = main_mode ? ray.d_rcp : ray2.d_rcp;
                r.mint = main_mode ? ray.mint : ray2.mint;
                r.maxt = main_mode ? ray.maxt : ray2.maxt;
                sample_interaction(sv, m, r, next_1d(true), mi, cnt);
                if (m.type == 0 && mi.valid()) {
                    if (main_mode) ray.maxt = mi.t;
                    else ray2.maxt = mode == M_NEE ? min_(mi.t, remaining) : mi.t;
                }
            }
            __syncwarp();
            if (go) {
                const bool need = main_mode ? needs_isect : needs_isect2;
                if (need) { state = L_ISECT; isect_ret = C_STEP_B; isect_march = !main_mode; }
                else state = L_STEP_B;
            }
            const bool do_b = state == L_STEP_B;
            /* common part of step B: surface-vs-collision ordering and the transmittance / pdf ratio
               (Medium::eval_tr_and_pdf, medium.cpp:77-89; volpath.cpp:113-117,300-311,402-404) */
            const bool nee = mode == M_NEE;
            bool spectral = false;
            F ratio = 1;
            if (do_b && a_in_medium) {
                const F si_t = main_mode ? si.t : si2.t;
                spectral = sv.media[med].spectral != 0;
                if (si_t < mi.t) mi.t = INF;
                if (spectral) {
                    F tt = min_(mi.t, si_t);
                    if (nee) tt = min_(remaining, tt);
                    F t = tt - mi.mint;
                    F tr = exp_(-t * mi.combined);
                    bool surface_first = si_t < mi.t || (nee && mi.t > remaining);
                    F pdf = surface_first ? tr : tr * mi.combined;
                    ratio = pdf > 0 ? tr / pdf : (F) 0;
                }
            }
            __syncwarp();
            if (do_b && main_mode) {
                bool act_null = false, act_real = false, escaped = false;
                if (a_in_medium) {
                    if (spectral) throughput *= ratio;
                    escaped = !mi.valid();
                    const bool active_medium = mi.valid();
                    F u = next_1d(active_medium);
                    bool null_scatter = u >= mi.sigma_t / mi.combined;
                    act_null = null_scatter && active_medium;
                    act_real = !act_null && active_medium;
                    if (spectral && act_null) throughput *= mi.sigma_n * mi.combined / mi.sigma_n;
                    if (act_real) depth += 1;
                }
                const bool active = depth < max_depth;
                act_real = act_real && active;
                if (act_null) { /* volpath.cpp:140-144 */
                    cnt.n_null++;
                    ray.o = mi.p;
                    ray.mint = 0;
                    si.t = si.t - mi.t;
                    state = L_STEP_A;
                } else if (act_real) {
                    state = L_SCATTER;
                } else if ((!a_in_medium || escaped) && si.valid()) {
                    state = L_SURFACE;
                } else {
                    state = L_PATH_END;
                }
            } else if (do_b) {
                /* sample_emitter (volpath.cpp:285-365) / evaluate_direct_light (volpath.cpp:384-463) */
                bool escaped = false, active_medium = false;
                if (a_in_medium) {
                    if (spectral) transmittance *= ratio;
                    if (nee) {
                        if (mi.t > remaining && mi.valid()) total_dist = ds_dist;
                        if (mi.t > remaining) mi.t = INF;
                    }
                    escaped = !mi.valid();
                    active_medium = mi.valid();
                    if (active_medium) {
                        if (nee) total_dist += mi.t;
                        cnt.n_null++;
                        ray2.o = mi.p;
                        ray2.mint = 0;
                        si2.t = si2.t - mi.t;
                        if (spectral) transmittance *= mi.sigma_n;
                        else transmittance *= mi.sigma_n / mi.combined;
                    }
                }
                bool active_surface = !a_in_medium || escaped;
                if (nee && active_surface) total_dist += si2.t;
                active_surface = active_surface && si2.t != INF && !active_medium;
                if (active_surface) {
                    const DShape<F> &sh = sv.shapes[si2.shape];
                    /* eval_null_transmission: 1 for `null`, 0 otherwise (null.cpp:69-72, bsdf.cpp:11-14) */
                    transmittance *= sv.bsdfs[sh.bsdf].type == 1 ? (F) 1 : (F) 0;
                    ray2 = spawn_ray(si2.p, ray2.d);
                    if (sh.interior >= 0 || sh.exterior >= 0)
                        medium2 = dot(ray2.d, si2.n) > 0 ? sh.exterior : sh.interior;
                }
                if (nee) ray2.maxt = remaining;
                needs_isect2 = needs_isect2 || active_surface;
                const bool active = (active_medium || active_surface) && transmittance != 0;
                if (active) state = L_STEP_A;
                else march_done();
            }
        } else if (phase == 1) {
            /* ================================ ISECT =============================== */
            if (state == L_ISECT) {
                Ray<F> r;
                r.o = isect_march ? ray2.o : ray.o;
                r.d = isect_march ? ray2.d : ray.d;
                r.d_rcp = isect_march ? ray2.d_rcp : ray.d_rcp;
                r.mint = isect_march ? ray2.mint : ray.mint;
                r.maxt = isect_march ? ray2.maxt : ray.maxt;
                SurfaceHit<F> h;
                ray_intersect(sv, r, h, cnt);
                if (isect_march) {
                    si2.t = h.t; si2.p = h.p; si2.n = h.n; si2.shape = h.shape;
                    needs_isect2 = false;
                } else {
                    si = h;
                    needs_isect = false;
                }
                if (isect_ret == C_START_DL) start_dl();
                else state = L_STEP_B;
            }
        } else if (phase == 2) {
            /* =============================== SCATTER ============================== */
            if (state == L_SCATTER) { /* volpath.cpp:146-168 */
                cnt.n_real++;
                const DMedium<F> &m = sv.media[medium];
                if (m.spectral) throughput *= mi.sigma_s * mi.combined / mi.sigma_t;
                else throughput *= mi.sigma_s / mi.sigma_t;
                valid_ray = true;
                ray.o = mi.p; /* the old origin is dead: keep the scattering position here */
                state = L_SCATTER_B;
                if (m.sample_emitters) {
                    DirSample<F> ds;
                    bool march = begin_nee(mi.p, true, ds);
                    nee_w = throughput * phase_eval_at(sv, sv.phases[m.phase], mi.p, -ray.d, ds.d);
                    cont = C_SCATTER_B;
                    if (march) { mode = M_NEE; state = L_STEP_A; }
                    else result += nee_w * emitter_val;
                }
            } else if (state == L_SCATTER_B) { /* phase function sampling (volpath.cpp:170-176) */
                const DPhase<F> &phase_fn = sv.phases[sv.media[medium].phase];
                F da = next_1d(true), db = next_1d(true), dc = next_1d(true);
                F s1 = ltr ? da : dc, s2x = ltr ? db : da, s2y = ltr ? dc : db;
                V3<F> wo = phase_sample_at(sv, phase_fn, ray.o, ray.d, s1, s2x, s2y);
                ray = spawn_ray(ray.o, wo);
                ray.mint = 0;
                needs_isect = true;
                state = L_STEP_A;
            }
        } else if (phase == 3) {
            /* =============================== SURFACE ============================== */
            if (state == L_SURFACE) { /* volpath.cpp:192-212 */
                cnt.n_surface++;
                const DBsdf<F> &bsdf = sv.bsdfs[sv.shapes[si.shape].bsdf];
                bool active_e = (bsdf_flags(bsdf.type) & BF_Smooth) != 0 && (depth + 1 < max_depth);
                state = L_SURF_B;
                if (active_e) {
                    DirSample<F> ds;
                    bool march = begin_nee(si.p, false, ds);
                    V3<F> wo = si.sh.to_local(ds.d);
                    F bsdf_val = bsdf_eval(bsdf, si.wi, wo);
                    F bpdf = bsdf_pdf(bsdf, si.wi, wo);
                    nee_w = throughput * bsdf_val * mis_weight(ds.pdf, ds.delta ? (F) 0 : bpdf);
                    cont = C_SURF_B;
                    if (march) { mode = M_NEE; state = L_STEP_A; }
                    else result += nee_w * emitter_val;
                }
            } else if (state == L_SURF_B) { /* BSDF sampling (volpath.cpp:214-241) */
                const DShape<F> &sh = sv.shapes[si.shape];
                const DBsdf<F> &bsdf = sv.bsdfs[sh.bsdf];
                F da = next_1d(true), db = next_1d(true), dc = next_1d(true);
                F s1 = ltr ? da : dc, s2x = ltr ? db : da, s2y = ltr ? dc : db;
                V3<F> bwo;
                F bs_pdf, bs_eta, bsdf_val;
                uint32_t sampled_type;
                bsdf_sample(bsdf, si.wi, s1, s2x, s2y, bwo, bs_pdf, bs_eta, sampled_type, bsdf_val);
                throughput *= bsdf_val;
                eta *= bs_eta;
                ray = spawn_ray(si.p, si.sh.to_world(bwo));
                needs_isect = true;
                bool non_null = !(sampled_type & BF_Null);
                if (non_null) depth += 1;
                valid_ray = valid_ray || non_null;
                bool add_emitter = !(sampled_type & BF_Delta) && throughput != 0 && (depth < max_depth);
                /* medium switch (volpath.cpp:249-250), applied after the direct-light march */
                medium_next = (sh.interior >= 0 || sh.exterior >= 0) ? (dot(ray.d, si.n) > 0 ? sh.exterior : sh.interior)
                                                                     : medium;
                if (add_emitter) { /* intersect2: the new hit replaces si (volpath.cpp:238-252) */
                    state = L_ISECT; isect_ret = C_START_DL; isect_march = false;
                } else {
                    medium = medium_next;
                    state = L_STEP_A;
                }
            }
        } else {
            /* ================================= END ================================ */
            if (state == L_PATH_END) {
                /* render_sample tail (integrator.cpp:267-287) + ImageBlock::put (imageblock.cpp:85-169) */
                F value = ray_weight * result;
                tot[TOT_PATHS]++;
                if (out.out_radiance) out.out_radiance[k * st.n_samples + j] = value;
                F alpha = valid_ray ? (F) 1 : (F) 0;
                bool ok = C > 5 || (value >= (F) -1e-5f && finite_(value)); /* no check with AOVs */
                if (!ok) {
                    tot[TOT_INVALID]++;
                } else {
                    F bx = posx - ((F) st.box + (F) 0.5), by = posy - ((F) st.boy + (F) 0.5);
                    int lx = (int) ceil_(bx - (F) 0.5), ly = (int) ceil_(by - (F) 0.5);
                    if (sn.rf_taps > 0) {
                        film_put_wide(film, sn, C, posx, posy, st.box, st.boy, st.bsx, st.bsy, sp.ox, sp.oy, value, alpha, result);
                    } else if (!(lx < 0 || ly < 0 || lx >= st.bsx || ly >= st.bsy)) {
                        int fx = lx + st.box - sp.ox, fy = ly + st.boy - sp.oy;
                        if (C > 5) { /* moment.cpp:78-83: the nested integrator's X Y Z (without the ray weight) and squares */
                            F *q = film + ((size_t) fy * W + fx) * C + 5;
                            const F raw = result, raw2 = raw * raw;
                            film_add(q, raw); film_add(q + 1, raw); film_add(q + 2, raw);
                            film_add(q + 3, raw2); film_add(q + 4, raw2); film_add(q + 5, raw2);
                        }
                        if (fx == st.px - sp.ox && fy == st.py - sp.oy) {
                            acc0 += value; acc3 += alpha; acc4 += (F) 1;
                            acc_used = true;
                        } else {
                            F *p = film + ((size_t) fy * W + fx) * C;
                            film_add(p, value); film_add(p + 1, value); film_add(p + 2, value);
                            film_add(p + 3, alpha); film_add(p + 4, (F) 1);
                        }
                    }
                }
                j++;
                state = L_NEW;
            }
            if (state == L_NEW) {
                if (have_stream && (j >= st.n_samples || st.empty)) {
                    /* stream finished: splat its accumulator, write the parity outputs */
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
                    have_stream = false;
                }
                if (!have_stream) {
                    k = atomicAdd(out.work_counter, 1ull);
                    if (k >= sp.count || stop_requested(out)) {
                        state = L_DONE;
                    } else {
                        st = make_stream(sp, stream_index(sp, k));
                        rng.state = st.state; rng.inc = st.inc; rng.draws = 0;
                        cnt.draws = 0; cnt.medium_steps = 0; cnt.n_null = 0; cnt.n_real = 0;
                        cnt.n_intersect = 0; cnt.n_surface = 0;
                        acc0 = 0; acc3 = 0; acc4 = 0; acc_used = false;
                        j = 0;
                        have_stream = true;
                    }
                }
                if (state == L_NEW && have_stream && !(j >= st.n_samples || st.empty)) {
                    /* render_sample prologue (integrator.cpp:233-264); an empty stream (pixel outside
                       its block) is finished on the next visit */
                    posx = (F) (uint32_t) st.px + rng.next(F(), true);
                    posy = (F) (uint32_t) st.py + rng.next(F(), true);
                    F apx = (F) 0.5, apy = (F) 0.5;
                    if (sn.needs_aperture_sample) { apx = rng.next(F(), true); apy = rng.next(F(), true); }
                    if (sn.has_shutter) (void) rng.next(F(), true);
                    (void) rng.next(F(), true); /* wavelength sample (integrator.cpp:252) */
                    F ax = (posx - (F) sn.crop_offset[0]) / (F) sn.crop_size[0];
                    F ay = (posy - (F) sn.crop_offset[1]) / (F) sn.crop_size[1];
                    ray_weight = sensor_sample_ray(sv, ax, ay, apx, apy, ray);
                    /* VolumetricPathIntegrator::sample prologue (volpath.cpp:45-70) */
                    valid_ray = false;
                    eta = 1; throughput = 1; result = 0;
                    medium = sn.medium;
                    depth = 0;
                    si.t = INF; si.shape = -1;
                    needs_isect = true;
                    mode = M_MAIN;
                    state = L_STEP_A;
                }
            }
        }
    }

    /* totals: warp reduce, one atomic per warp and counter */
#pragma unroll
    for (int i = 0; i < TOT_N; ++i) {
        unsigned long long v = tot[i];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(FULL, v, o);
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
