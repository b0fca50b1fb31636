/*
 * vp_integrator.cuh -- VolumetricPathIntegrator::sample and its two nested
 * marches, one path per thread (reference: src/integrators/volpath.cpp:38-483).
 *
 * The random-number draw order is part of the contract:
 * masked draws advance the generator only when their mask is true, and the
 * phase/BSDF sample arguments are drawn in the order the reference's compiler
 * chose (`ltr`: 0 = GCC, right to left).
 */
#pragma once
#include "vp_scene.cuh"

namespace VP_NS {

template <typename F> struct PathCtx {
    const SceneView<F> &sv;
    Pcg32 &rng;
    PathCounters &cnt;
    uint32_t max_depth, rr_depth;
    bool hide_emitters, ltr;
    bool spectral_mis;   /* volpathmis `use_spectral_mis` (volpathmis.cpp:36) */

    VP_D F next_1d(bool active) { return rng.next(F(), active); }

    /* SurfaceInteraction::target_medium (interaction.h:186-200) */
    VP_D int target_medium(const SurfaceHit<F> &si, V3<F> d) const {
        const DShape<F> &sh = sv.shapes[si.shape];
        return dot(d, si.n) > 0 ? sh.exterior : sh.interior;
    }
    /* Shape::is_medium_transition (include/mitsuba/render/shape.h:341) */
    VP_D bool is_medium_transition(const SurfaceHit<F> &si) const {
        const DShape<F> &sh = sv.shapes[si.shape];
        return sh.interior >= 0 || sh.exterior >= 0;
    }
    VP_D F null_transmission(const SurfaceHit<F> &si) const { /* null.cpp:69-72, bsdf.cpp:11-14 */
        return sv.bsdfs[sv.shapes[si.shape].bsdf].type == 1 ? (F) 1 : (F) 0;
    }
    /* volpath.cpp:479-483 */
    VP_D F mis_weight(F pdf_a, F pdf_b) const {
        pdf_a *= pdf_a;
        pdf_b *= pdf_b;
        return pdf_a > 0 ? pdf_a / (pdf_a + pdf_b) : (F) 0;
    }

    /* sample_emitter (volpath.cpp:261-367): emitter sample + ratio-tracking
       transmittance towards it through media and null interfaces */
    __device__ __noinline__ F sample_emitter(V3<F> ref_p, bool is_medium_interaction, int medium, bool active,
                                             DirSample<F> &ds) {
        F transmittance = 1;
        F sx = next_1d(active);
        F sy = next_1d(active);
        F emitter_val;
        ds = sv.g->environment >= 0 ? sample_emitter_direction_env(sv, sx, sy, emitter_val)
                                    : sample_emitter_direction(sv, sx, emitter_val);
        if (ds.pdf == 0) emitter_val = 0;
        active = active && ds.pdf != 0;
        if (!active)
            return emitter_val;

        Ray<F> ray = spawn_ray(ref_p, ds.d);
        if (is_medium_interaction) ray.mint = 0;

        F total_dist = 0;
        SurfaceHit<F> si;
        si.t = inf_<F>(); si.shape = -1;
        bool needs_intersection = true;
        while (active) {
            F remaining_dist = ds.dist * ((F) 1 - K<F>::ShadowEpsilon) - total_dist;
            ray.maxt = remaining_dist;
            active = active && remaining_dist > 0;
            if (!active) break;

            bool escaped_medium = false;
            bool active_medium = active && medium >= 0;
            bool active_surface = active && !active_medium;

            if (active_medium) {
                const DMedium<F> &m = sv.media[medium];
                MediumHit<F> mi;
                sample_interaction(sv, m, ray, next_1d(active_medium), mi, cnt);
                if (m.type == 0 && mi.valid())
                    ray.maxt = min_(mi.t, remaining_dist);
                if (needs_intersection)
                    ray_intersect(sv, ray, si, cnt);
                if (si.t < mi.t) mi.t = inf_<F>();
                needs_intersection = false;

                bool is_spectral = m.spectral != 0, not_spectral = !is_spectral;
                if (is_spectral) {
                    F t = min_(remaining_dist, min_(mi.t, si.t)) - mi.mint;
                    F tr = exp_(-t * mi.combined);
                    F free_flight_pdf = (si.t < mi.t || mi.t > remaining_dist) ? tr : tr * mi.combined;
                    transmittance *= free_flight_pdf > 0 ? tr / free_flight_pdf : (F) 0;
                }
                if (mi.t > remaining_dist && mi.valid()) total_dist = ds.dist;
                if (mi.t > remaining_dist) mi.t = inf_<F>();

                escaped_medium = !mi.valid();
                active_medium = mi.valid();
                is_spectral = is_spectral && active_medium;
                not_spectral = not_spectral && active_medium;

                if (active_medium) {
                    total_dist += mi.t;
                    cnt.n_null++;
                    ray.o = mi.p;
                    ray.mint = 0;
                    si.t = si.t - mi.t;
                    if (is_spectral) transmittance *= mi.sigma_n;
                    if (not_spectral) transmittance *= mi.sigma_n / mi.combined;
                }
            }

            bool intersect = active_surface && needs_intersection;
            if (intersect) ray_intersect(sv, ray, si, cnt);
            needs_intersection = needs_intersection && !intersect;
            active_surface = active_surface || escaped_medium;
            if (active_surface) total_dist += si.t;

            active_surface = active_surface && si.valid() && active && !active_medium;
            if (active_surface) {
                transmittance *= null_transmission(si);
                ray = spawn_ray(si.p, ray.d);
            }
            ray.maxt = remaining_dist;
            needs_intersection = needs_intersection || active_surface;

            active = active && (active_medium || active_surface) && transmittance != 0;

            if (active_surface && is_medium_transition(si))
                medium = target_medium(si, ray.d);
        }
        return transmittance * emitter_val;
    }

    /* evaluate_direct_light (volpath.cpp:370-465).  Only delta emitters are in
       scope, so nothing can be hit and the value is 0 -- but the march consumes
       random numbers and counts as ray segments. */
    __device__ __noinline__ void evaluate_direct_light(int medium, Ray<F> ray, SurfaceHit<F> &si, bool active,
                                                       F &emitted, F &emitter_pdf) {
        F emitter_val = 0, transmittance = 1;
        emitter_pdf = 0;
        bool needs_intersection = false;
        while (active) {
            bool escaped_medium = false;
            bool active_medium = active && medium >= 0;
            bool active_surface = active && !active_medium;
            if (active_medium) {
                const DMedium<F> &m = sv.media[medium];
                MediumHit<F> mi;
                sample_interaction(sv, m, ray, next_1d(active_medium), mi, cnt);
                if (m.type == 0 && mi.valid()) ray.maxt = mi.t;
                if (needs_intersection) ray_intersect(sv, ray, si, cnt);
                if (si.t < mi.t) mi.t = inf_<F>();
                if (m.spectral) transmittance *= tr_over_pdf(mi, si.t);
                needs_intersection = false;
                escaped_medium = !mi.valid();
                active_medium = mi.valid();
                if (active_medium) {
                    cnt.n_null++;
                    ray.o = mi.p;
                    ray.mint = 0;
                    si.t = si.t - mi.t;
                    if (m.spectral) transmittance *= mi.sigma_n;
                    else transmittance *= mi.sigma_n / mi.combined;
                }
            }
            bool intersect = active_surface && needs_intersection;
            if (intersect) ray_intersect(sv, ray, si, cnt);
            needs_intersection = needs_intersection && !intersect;
            active_surface = active_surface || escaped_medium;

            /* si.emitter(scene) (interaction.h:571-596): the environment when nothing was hit (volpath.cpp:429-438;
               Scene::pdf_emitter_direction, scene.cpp:221-235, over constant.cpp:106-110) */
            if (sv.g->environment >= 0 && !si.valid() && active_surface) {
                emitter_val = sv.emitters[sv.g->environment].irradiance;
                emitter_pdf = K<F>::InvFourPi;
                if (sv.g->n_emitters > 1) emitter_pdf = emitter_pdf * (F) (1.f / (float) sv.g->n_emitters);   /* scene.cpp:233 */
                active = false; active_surface = false; active_medium = false;
            }

            active_surface = active_surface && si.valid() && !active_medium;
            if (active_surface) {
                transmittance *= null_transmission(si);
                ray = spawn_ray(si.p, ray.d);
            }
            needs_intersection = needs_intersection || active_surface;
            active = active && (active_medium || active_surface) && transmittance != 0;
            if (active_surface && is_medium_transition(si))
                medium = target_medium(si, ray.d);
        }
        emitted = transmittance * emitter_val;
    }

    /* ================= VolpathMisIntegratorImpl (src/integrators/volpathmis.cpp), mono ==================
       WeightMatrix is 1 x 1 in mono.  p_over_f_nee of the main loop is read when a path hits an emitter
       (volpathmis.cpp:253-268): here, when it escapes to the environment emitter. */
    VP_D static void update_weights(F &p_over_f, F p, F f, bool active) { /* volpathmis.cpp:447-465 */
        F ratio = p / f;
        ratio = finite_(ratio) ? ratio : (F) 0;
        ratio *= p_over_f;
        if (active) p_over_f = (ratio != ratio) ? (F) 0 : ratio;
    }
    VP_D static F mis_weight1(F p_over_f) { return p_over_f == 0 ? (F) 0 : (F) 1 / p_over_f; }            /* :467-482 */
    VP_D static F mis_weight2(F a, F b) { F sum = a + b; return sum == 0 ? (F) 0 : (F) 1 / sum; }           /* :485-500 */

    /* volpathmis.cpp:330-444: returns the emitter value; p_over_f_nee / p_over_f_uni through the references */
    __device__ __noinline__ F sample_emitter_mis(V3<F> ref_p, bool is_medium_interaction, int medium, F p_over_f,
                                                 bool active, DirSample<F> &ds, F &p_over_f_nee, F &p_over_f_uni) {
        p_over_f_nee = p_over_f; p_over_f_uni = p_over_f;
        F sx = next_1d(active);
        F sy = next_1d(active);
        F weight;
        ds = sv.g->environment >= 0 ? sample_emitter_direction_env(sv, sx, sy, weight)
                                    : sample_emitter_direction(sv, sx, weight);
        F emitter_val = weight * ds.pdf;
        if (ds.pdf == 0) emitter_val = 0;
        active = active && ds.pdf != 0;
        update_weights(p_over_f_nee, ds.pdf, (F) 1, active);
        if (!active) return emitter_val;

        Ray<F> ray = spawn_ray(ref_p, ds.d);
        if (is_medium_interaction) ray.mint = 0;
        F total_dist = 0;
        SurfaceHit<F> si;
        si.t = inf_<F>(); si.shape = -1;
        bool needs_intersection = true;
        while (active) {
            F remaining_dist = ds.dist * ((F) 1 - K<F>::ShadowEpsilon) - total_dist;
            ray.maxt = remaining_dist;
            active = active && remaining_dist > 0;
            if (!active) break;
            bool escaped_medium = false;
            bool active_medium = active && medium >= 0;
            bool active_surface = active && !active_medium;
            if (active_medium) {
                const DMedium<F> &m = sv.media[medium];
                MediumHit<F> mi;
                sample_interaction(sv, m, ray, next_1d(active_medium), mi, cnt);
                if (m.type == 0 && mi.valid()) ray.maxt = min_(mi.t, remaining_dist);
                if (needs_intersection) ray_intersect(sv, ray, si, cnt);
                if (si.t < mi.t) mi.t = inf_<F>();
                needs_intersection = false;
                bool is_spectral = m.spectral != 0, not_spectral = !is_spectral;
                if (is_spectral) {
                    F t = min_(remaining_dist, min_(mi.t, si.t)) - mi.mint;
                    F tr = exp_(-t * mi.combined);
                    F free_flight_pdf = (si.t < mi.t || mi.t > remaining_dist) ? tr : tr * mi.combined;
                    update_weights(p_over_f_nee, free_flight_pdf, tr, true);
                    update_weights(p_over_f_uni, free_flight_pdf, tr, true);
                }
                if (mi.t > remaining_dist && mi.valid()) total_dist = ds.dist;
                if (mi.t > remaining_dist) mi.t = inf_<F>();
                escaped_medium = !mi.valid();
                active_medium = mi.valid();
                is_spectral = is_spectral && active_medium;
                not_spectral = not_spectral && active_medium;
                if (active_medium) {
                    total_dist += mi.t;
                    cnt.n_null++;
                    ray.o = mi.p; ray.mint = 0;
                    si.t = si.t - mi.t;
                    if (is_spectral) {
                        update_weights(p_over_f_nee, (F) 1, mi.sigma_n, true);
                        update_weights(p_over_f_uni, mi.sigma_n / mi.combined, mi.sigma_n, true);
                    }
                    if (not_spectral) {
                        update_weights(p_over_f_nee, (F) 1, mi.sigma_n / mi.combined, true);
                        update_weights(p_over_f_uni, mi.sigma_n, mi.sigma_n, true);
                    }
                }
            }
            bool intersect = active_surface && needs_intersection;
            if (intersect) ray_intersect(sv, ray, si, cnt);
            active_surface = active_surface || escaped_medium;
            if (active_surface) total_dist += si.t;
            active_surface = active_surface && si.valid() && active && !active_medium;
            if (active_surface) {
                F bsdf_val = null_transmission(si);
                update_weights(p_over_f_nee, (F) 1, bsdf_val, true);
                update_weights(p_over_f_uni, (F) 1, bsdf_val, true);
            }
            if (active_surface) ray = spawn_ray(si.p, ray.d);
            ray.maxt = remaining_dist;
            needs_intersection = needs_intersection || active_surface;
            if (spectral_mis)
                active = active && (active_medium || active_surface) && mis_weight1(p_over_f_uni) != 0;
            else
                active = active && (active_medium || active_surface) && (p_over_f_uni != 0 || p_over_f_nee != 0);
            if (active_surface && is_medium_transition(si))
                medium = target_medium(si, ray.d);
        }
        return emitter_val;
    }

    /* volpathmis.cpp:91-329 */
    VP_D F sample_mis(Ray<F> ray, int initial_medium, bool &valid_ray_out) {
        bool active = true;
        bool valid_ray = !hide_emitters && sv.g->environment >= 0; /* volpathmis.cpp:100 */
        F eta = 1, result = 0;
        int medium = initial_medium;
        MediumHit<F> mi;
        mi.t = inf_<F>();
        bool specular_chain = active && !hide_emitters;
        uint32_t depth = 0;
        F p_over_f = 1, p_over_f_nee = 1;
        SurfaceHit<F> si;
        si.t = inf_<F>(); si.shape = -1;
        bool needs_intersection = true;
        for (;;) {
            /* Russian roulette (122-128); last_event_was_null is never set in this version of the integrator */
            F mis_throughput = mis_weight1(p_over_f);
            F q = min_(mis_throughput * sqr(eta), (F) 0.95f);
            bool perform_rr = active && depth > rr_depth;
            F rr_sample = next_1d(active);
            active = active && !(rr_sample >= q && perform_rr);
            update_weights(p_over_f, q, (F) 1, perform_rr);
            bool exceeded_max_depth = depth >= max_depth;
            active = active && !exceeded_max_depth;
            active = active && mis_weight1(p_over_f) != 0;
            if (!active) break;

            bool active_medium = active && medium >= 0;
            bool active_surface = active && !active_medium;
            bool act_null_scatter = false, act_medium_scatter = false, escaped_medium = false;
            bool is_spectral = active_medium, not_spectral = false;
            if (active_medium) {
                is_spectral = is_spectral && sv.media[medium].spectral != 0;
                not_spectral = !is_spectral && active_medium;
            }
            if (active_medium) {
                const DMedium<F> &m = sv.media[medium];
                sample_interaction(sv, m, ray, next_1d(active_medium), mi, cnt);
                if (m.type == 0 && mi.valid()) ray.maxt = mi.t;
                if (needs_intersection) ray_intersect(sv, ray, si, cnt);
                needs_intersection = false;
                if (si.t < mi.t) mi.t = inf_<F>();
                if (is_spectral) { /* Medium::eval_tr_and_pdf (medium.cpp:77-89) */
                    F t = min_(mi.t, si.t) - mi.mint;
                    F tr = exp_(-t * mi.combined);
                    F free_flight_pdf = si.t < mi.t ? tr : tr * mi.combined;
                    update_weights(p_over_f, free_flight_pdf, tr, true);
                    update_weights(p_over_f_nee, free_flight_pdf, tr, true);
                }
                escaped_medium = !mi.valid();
                active_medium = mi.valid();
                is_spectral = is_spectral && active_medium;
                not_spectral = not_spectral && active_medium;
            }
            if (active_medium) {
                const DMedium<F> &m = sv.media[medium];
                F u = next_1d(active_medium);
                bool null_scatter = u >= mi.sigma_t / mi.combined;
                act_null_scatter = null_scatter && active_medium;
                act_medium_scatter = !act_null_scatter && active_medium;
                if (act_medium_scatter) depth += 1;
                bool sample_emitters = m.sample_emitters != 0;
                active = active && depth < max_depth;
                act_medium_scatter = act_medium_scatter && active;
                specular_chain = specular_chain && !(act_medium_scatter && sample_emitters);
                if (act_null_scatter) {
                    cnt.n_null++;
                    if (is_spectral) {
                        update_weights(p_over_f, mi.sigma_n / mi.combined, mi.sigma_n, true);
                        update_weights(p_over_f_nee, (F) 1, mi.sigma_n, true);
                    }
                    if (not_spectral) {
                        update_weights(p_over_f, mi.sigma_n, mi.sigma_n, true);
                        update_weights(p_over_f_nee, (F) 1, mi.sigma_n / mi.combined, true);
                    }
                    ray.o = mi.p; ray.mint = 0;
                    si.t = si.t - mi.t;
                }
                if (act_medium_scatter) {
                    cnt.n_real++;
                    if (is_spectral) update_weights(p_over_f, mi.sigma_t / mi.combined, mi.sigma_s, true);
                    if (not_spectral) update_weights(p_over_f, mi.sigma_t, mi.sigma_s, true);
                    valid_ray = true;
                    if (sample_emitters) {
                        DirSample<F> ds;
                        F nee_end, uni_end;
                        F emitted = sample_emitter_mis(mi.p, true, medium, p_over_f, true, ds, nee_end, uni_end);
                        F phase_val = phase_eval_at(sv, sv.phases[m.phase], mi.p, -ray.d, ds.d);
                        update_weights(nee_end, (F) 1, phase_val, true);
                        update_weights(uni_end, ds.delta ? (F) 0 : phase_val, phase_val, true);
                        result += mis_weight2(nee_end, uni_end) * emitted;
                    }
                    p_over_f_nee = p_over_f;
                    F da = next_1d(true), db = next_1d(true), dc = next_1d(true);
                    F s1 = ltr ? da : dc, s2x = ltr ? db : da, s2y = ltr ? dc : db;
                    F phase_pdf;   /* what phase->sample returns with the direction (volpathmis.cpp:229) */
                    V3<F> wo = phase_sample_at(sv, sv.phases[m.phase], mi.p, ray.d, s1, s2x, s2y, &phase_pdf);
                    ray = spawn_ray(mi.p, wo);
                    ray.mint = 0;
                    needs_intersection = true;
                    update_weights(p_over_f, phase_pdf, phase_pdf, true);
                    update_weights(p_over_f_nee, (F) 1, phase_pdf, true);
                }
            }

            active_surface = active_surface || escaped_medium;
            bool intersect = active_surface && needs_intersection;
            if (intersect) ray_intersect(sv, ray, si, cnt);
            /* emitter hit (252-268): the environment when the ray escaped; shapes carry no emitters */
            if (active_surface && sv.g->environment >= 0 && !si.valid() && !(depth == 0 && hide_emitters)) {
                bool count_direct = depth == 0 || specular_chain;
                if (!count_direct) { /* Scene::pdf_emitter_direction (scene.cpp:221-235) over constant.cpp:106-110 */
                    F emitter_pdf = K<F>::InvFourPi;
                    if (sv.g->n_emitters > 1) emitter_pdf = emitter_pdf * (F) (1.f / (float) sv.g->n_emitters);   /* scene.cpp:233 */
                    update_weights(p_over_f_nee, emitter_pdf, (F) 1, true);
                }
                F emitted = sv.emitters[sv.g->environment].irradiance;
                result += (count_direct ? mis_weight1(p_over_f) : mis_weight2(p_over_f, p_over_f_nee)) * emitted;
            }
            active_surface = active_surface && si.valid();
            if (active_surface) {
                cnt.n_surface++;
                const DShape<F> &sh = sv.shapes[si.shape];
                const DBsdf<F> &bsdf = sv.bsdfs[sh.bsdf];
                bool active_e = (bsdf_flags(bsdf.type) & BF_Smooth) != 0 && (depth + 1 < max_depth);
                if (active_e) {
                    DirSample<F> ds;
                    F nee_end, uni_end;
                    F emitted = sample_emitter_mis(si.p, false, medium, p_over_f, true, ds, nee_end, uni_end);
                    V3<F> wo_local = si.sh.to_local(ds.d);
                    F bsdf_val = bsdf_eval(bsdf, si.wi, wo_local);
                    F bpdf = bsdf_pdf(bsdf, si.wi, wo_local);
                    update_weights(nee_end, (F) 1, bsdf_val, true);
                    update_weights(uni_end, ds.delta ? (F) 0 : bpdf, bsdf_val, true);
                    result += mis_weight2(nee_end, uni_end) * emitted;
                }
                F da = next_1d(true), db = next_1d(true), dc = next_1d(true);
                F s1 = ltr ? da : dc, s2x = ltr ? db : da, s2y = ltr ? dc : db;
                V3<F> bwo;
                F bs_pdf, bs_eta, bsdf_weight;
                uint32_t sampled_type;
                bsdf_sample(bsdf, si.wi, s1, s2x, s2y, bwo, bs_pdf, bs_eta, sampled_type, bsdf_weight);
                bool invalid_bsdf_sample = bs_pdf == 0;
                active_surface = active_surface && bs_pdf > 0;
                if (active_surface) {
                    eta *= bs_eta;
                    ray = spawn_ray(si.p, si.sh.to_world(bwo));
                }
                needs_intersection = needs_intersection || active_surface;
                bool non_null_bsdf = active_surface && !(sampled_type & BF_Null);
                valid_ray = valid_ray || non_null_bsdf || invalid_bsdf_sample;
                specular_chain = specular_chain || (non_null_bsdf && (sampled_type & BF_Delta) != 0);
                specular_chain = specular_chain && !(active_surface && (sampled_type & BF_Smooth) != 0);
                if (non_null_bsdf) depth += 1;
                if (non_null_bsdf) p_over_f_nee = p_over_f;
                update_weights(p_over_f, bs_pdf, bsdf_weight * bs_pdf, active_surface);
                update_weights(p_over_f_nee, (F) 1, bsdf_weight * bs_pdf, non_null_bsdf);
                if (active_surface && is_medium_transition(si)) medium = target_medium(si, ray.d);
            }
            active = active && (active_surface || active_medium);
        }
        valid_ray_out = valid_ray;
        return result;
    }

    /* VolumetricPathIntegrator::sample (volpath.cpp:38-257) */
    __device__ F sample(Ray<F> ray, int initial_medium, bool &valid_ray_out) {
        bool active = true;
        bool valid_ray = !hide_emitters && sv.g->environment >= 0; /* volpath.cpp:51 */
        F eta = 1, throughput = 1, result = 0;
        int medium = initial_medium;
        MediumHit<F> mi;
        mi.t = inf_<F>(); mi.mint = 0; mi.p = mk3<F>(0, 0, 0);
        mi.sigma_s = 0; mi.sigma_n = 0; mi.sigma_t = 0; mi.combined = 0;
        bool specular_chain = active && !hide_emitters;
        uint32_t depth = 0;
        SurfaceHit<F> si;
        si.t = inf_<F>(); si.shape = -1;
        bool needs_intersection = true;

        for (;;) {
            /* termination / Russian roulette (volpath.cpp:72-87) */
            active = active && throughput != 0;
            F q = min_(throughput * sqr(eta), (F) 0.95f);
            bool perform_rr = depth > rr_depth;
            F rr_sample = next_1d(active);
            active = active && (rr_sample < q || !perform_rr);
            if (perform_rr) throughput *= rcp_(q);
            bool exceeded_max_depth = depth >= max_depth;
            if (!active || exceeded_max_depth) break;

            /* sampling the RTE (volpath.cpp:90-134) */
            bool active_medium = active && medium >= 0;
            bool active_surface = active && !active_medium;
            bool act_null_scatter = false, act_medium_scatter = false, escaped_medium = false;
            bool is_spectral = false, not_spectral = false;
            if (active_medium) {
                const DMedium<F> &m = sv.media[medium];
                is_spectral = m.spectral != 0;
                not_spectral = !is_spectral;
                sample_interaction(sv, m, ray, next_1d(active_medium), mi, cnt);
                if (m.type == 0 && mi.valid()) ray.maxt = mi.t;
                if (needs_intersection) ray_intersect(sv, ray, si, cnt);
                needs_intersection = false;
                if (si.t < mi.t) mi.t = inf_<F>();
                if (is_spectral) throughput *= tr_over_pdf(mi, si.t);

                escaped_medium = !mi.valid();
                active_medium = mi.valid();

                F u = next_1d(active_medium);
                bool null_scatter = u >= mi.sigma_t / mi.combined;
                act_null_scatter = null_scatter && active_medium;
                act_medium_scatter = !act_null_scatter && active_medium;

                if (is_spectral && act_null_scatter)
                    throughput *= mi.sigma_n * mi.combined / mi.sigma_n;

                if (act_medium_scatter) depth += 1;
            }

            active = active && depth < max_depth;
            act_medium_scatter = act_medium_scatter && active;

            if (act_null_scatter) { /* volpath.cpp:140-144 */
                cnt.n_null++;
                ray.o = mi.p;
                ray.mint = 0;
                si.t = si.t - mi.t;
            }

            if (act_medium_scatter) { /* volpath.cpp:146-176 */
                cnt.n_real++;
                if (is_spectral) throughput *= mi.sigma_s * mi.combined / mi.sigma_t;
                if (not_spectral) throughput *= mi.sigma_s / mi.sigma_t;

                const DMedium<F> &m = sv.media[medium];
                const DPhase<F> &phase = sv.phases[m.phase];
                bool sample_emitters = m.sample_emitters != 0;
                valid_ray = true;
                specular_chain = specular_chain && !act_medium_scatter;
                specular_chain = specular_chain || (act_medium_scatter && !sample_emitters);

                if (sample_emitters) {
                    DirSample<F> ds;
                    F emitted = sample_emitter(mi.p, true, medium, true, ds);
                    F phase_val = phase_eval_at(sv, phase, mi.p, -ray.d, ds.d);
                    result += throughput * phase_val * emitted;
                }
                /* phase->sample(ctx, mi, next_1d(), next_2d()) (volpath.cpp:171) */
                F da = next_1d(true), db = next_1d(true), dc = next_1d(true);
                F s1 = ltr ? da : dc, s2x = ltr ? db : da, s2y = ltr ? dc : db;
                V3<F> wo = phase_sample_at(sv, phase, mi.p, ray.d, s1, s2x, s2y);
                ray = spawn_ray(mi.p, wo);
                ray.mint = 0;
                needs_intersection = true;
            }

            /* surface interactions (volpath.cpp:178-253) */
            active_surface = active_surface || escaped_medium;
            bool intersect = active_surface && needs_intersection;
            if (intersect) ray_intersect(sv, ray, si, cnt);

            /* emitter hit (volpath.cpp:186-191): the environment when the ray escaped; shapes carry no emitters */
            if (active_surface && specular_chain && sv.g->environment >= 0 && !si.valid())
                result += throughput * sv.emitters[sv.g->environment].irradiance;

            active_surface = active_surface && si.valid();
            if (active_surface) {
                cnt.n_surface++;
                const DShape<F> &sh = sv.shapes[si.shape];
                const DBsdf<F> &bsdf = sv.bsdfs[sh.bsdf];
                bool active_e = (bsdf_flags(bsdf.type) & BF_Smooth) != 0 && (depth + 1 < max_depth);
                if (active_e) {
                    DirSample<F> ds;
                    F emitted = sample_emitter(si.p, false, medium, true, ds);
                    V3<F> wo = si.sh.to_local(ds.d);
                    F bsdf_val = bsdf_eval(bsdf, si.wi, wo);
                    F bpdf = bsdf_pdf(bsdf, si.wi, wo);
                    result += throughput * bsdf_val * mis_weight(ds.pdf, ds.delta ? (F) 0 : bpdf) * emitted;
                }

                /* bsdf->sample(ctx, si, next_1d(), next_2d()) (volpath.cpp:215-216) */
                F da = next_1d(true), db = next_1d(true), dc = next_1d(true);
                F s1 = ltr ? da : dc, s2x = ltr ? db : da, s2y = ltr ? dc : db;
                V3<F> bwo;
                F bs_pdf, bs_eta, bsdf_val;
                uint32_t sampled_type;
                bsdf_sample(bsdf, si.wi, s1, s2x, s2y, bwo, bs_pdf, bs_eta, sampled_type, bsdf_val);

                throughput *= bsdf_val;
                eta *= bs_eta;

                ray = spawn_ray(si.p, si.sh.to_world(bwo));
                needs_intersection = true;

                bool non_null_bsdf = !(sampled_type & BF_Null);
                if (non_null_bsdf) depth += 1;
                valid_ray = valid_ray || non_null_bsdf;
                specular_chain = specular_chain || (non_null_bsdf && (sampled_type & BF_Delta) != 0);
                specular_chain = specular_chain && !((sampled_type & BF_Smooth) != 0);

                bool add_emitter = !(sampled_type & BF_Delta) && throughput != 0 && (depth < max_depth);
                act_null_scatter = act_null_scatter || (sampled_type & BF_Null) != 0;

                bool intersect2 = needs_intersection && add_emitter;
                SurfaceHit<F> si_new = si;
                if (intersect2) ray_intersect(sv, ray, si_new, cnt);
                needs_intersection = needs_intersection && !intersect2;

                F emitted, emitter_pdf;
                SurfaceHit<F> si_march = si_new;
                evaluate_direct_light(medium, ray, si_march, add_emitter, emitted, emitter_pdf);
                if (add_emitter && emitter_pdf != 0)
                    result += mis_weight(bs_pdf, emitter_pdf) * throughput * emitted;

                if (is_medium_transition(si)) medium = target_medium(si, ray.d);

                if (intersect2) si = si_new;
            }
            active = active && (active_surface || active_medium);
        }
        (void) specular_chain;
        valid_ray_out = valid_ray;
        return result;
    }
};

} // namespace VP_NS
