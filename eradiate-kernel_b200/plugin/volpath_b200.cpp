// volpath_b200 -- the B200 volpath as a plugin of the reference itself.
//
// This is the host side of the drop-in boundary in the reference's own language and plugin API: a Mitsuba
// integrator plugin (MTS_EXPORT_PLUGIN, loaded by the reference's PluginManager from `plugins/volpath_b200.so`,
// src/libcore/plugin.cpp:27-36,94-111) whose Integrator::render (include/mitsuba/render/integrator.h:42) reads
// the object graph the reference's XML / load_dict loaders built, flattens it into the POD tables of
// include/b2vp.h and calls the C-ABI (libb2vp.so: hand-written CUDA for sm_100a).  A scene file switches to the
// GPU path by writing <integrator type="volpath_b200"> instead of <integrator type="volpath">; everything else
// (sensors, films, samplers, media, shapes, the `mitsuba` driver, film->develop()) stays the reference's.
//
// There is no CPU fallback: without a usable device b2vp_scene_create fails and render() throws.
//
// describe() reads plugin parameters where they live.  The reference exposes only part of them through
// accessors / traverse() (INTEGRATION.md section 2 lists which); the rest are private members of the plugin
// classes (rectangle frame, gridvolume filter / wrap mode, distant-sensor target, ...).  A maintainer adopting
// this plugin would add accessors; here the class definitions are taken from the reference's own plugin sources
// with member access opened, and objects are recognised by Class::name() (plugins are dlopen'ed RTLD_LOCAL, so
// dynamic_cast across plugin boundaries is not available anyway).  Nothing of those sources is instantiated in
// this library except member offsets.
//
// Built by oracle/Makefile.ref (needs the reference's headers, so only where /root/reference exists) into
// oracle/_ref/plugins/volpath_b200.so, next to the reference's own plugins.
#ifndef B2VP_DEFAULT_ALGORITHM
#define B2VP_DEFAULT_ALGORITHM "volpath"
#endif
#include <algorithm>
#include <array>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <functional>
#include <iostream>
#include <cstdio>
#include <cstdlib>
#include <unistd.h>
#include <map>
#include <memory>
#include <mutex>
#include <sstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#define private public
#define protected public
#include <mitsuba/core/fresolver.h>
#include <mitsuba/core/plugin.h>
#include <mitsuba/core/properties.h>
#include <mitsuba/core/rfilter.h>
#include <mitsuba/core/transform.h>
#include <mitsuba/core/warp.h>
#include <mitsuba/render/bsdf.h>
#include <mitsuba/render/emitter.h>
#include <mitsuba/render/film.h>
#include <mitsuba/render/imageblock.h>
#include <mitsuba/render/integrator.h>
#include <mitsuba/render/interaction.h>
#include <mitsuba/render/medium.h>
#include <mitsuba/render/mesh.h>
#include <mitsuba/render/phase.h>
#include <mitsuba/render/sampler.h>
#include <mitsuba/render/scene.h>
#include <mitsuba/render/sensor.h>
#include <mitsuba/render/shape.h>
#include <mitsuba/render/texture.h>

// class definitions of the plugins whose parameters describe() reads (no plugin entry points from them)
#undef MTS_EXPORT_PLUGIN
#define MTS_EXPORT_PLUGIN(...)
#include <shapes/rectangle.cpp>
#include <shapes/sphere.cpp>
#include <bsdfs/diffuse.cpp>
#include <bsdfs/rpv.cpp>
#include <bsdfs/bilambertian.cpp>
#include <media/homogeneous.cpp>
#include <media/heterogeneous.cpp>
#include <textures/grid3d.cpp>
#include <phase/hg.cpp>
#include <phase/tabphase.cpp>
#include <phase/blendphase.cpp>
#include <emitters/directional.cpp>
#include <emitters/constant.cpp>
#include <sensors/distant.cpp>
#define RayTargetType MultiDistantRayTargetType   /* distant.cpp and mdistant.cpp both define this enum */
#include <sensors/mdistant.cpp>
#undef RayTargetType
#define RayTargetType DistantFluxRayTargetType
#include <sensors/distantflux.cpp>
#undef RayTargetType
#include <sensors/mradiancemeter.cpp>
#undef private
#undef protected
#undef MTS_EXPORT_PLUGIN
#define MTS_EXPORT_PLUGIN(Name, Descr)                                                                       \
    extern "C" {                                                                                             \
        MTS_EXPORT const char *plugin_name() { return #Name; }                                               \
        MTS_EXPORT const char *plugin_descr() { return Descr; }                                              \
    }                                                                                                        \
    MTS_INSTANTIATE_CLASS(Name)

#include "b2vp.h"

NAMESPACE_BEGIN(mitsuba)

/* The flattened scene: the descriptor and every array it points to. */
struct B200SceneTables {
    B2vpSceneDesc desc{};
    std::vector<B2vpShape> shapes; std::vector<B2vpBsdf> bsdfs; std::vector<B2vpMedium> media;
    std::vector<B2vpVolume> volumes; std::vector<B2vpPhase> phases; std::vector<B2vpEmitter> emitters;
    std::vector<float> vpos, vnrm, vuv; std::vector<uint32_t> faces;
    std::vector<std::vector<float>> grids;
    std::vector<double> sensor_xf;
    std::vector<std::vector<double>> phase_values;   /* tabphase tables, parallel to `phases` */
    std::map<const void *, int> bsdf_ids, medium_ids, shape_ids;
    /* Fingerprint of everything the device scene is built from.  Taken BEFORE finish() fills the pointer fields, so
       it depends on content only; render() keeps the resident device scene while it is unchanged. */
    uint64_t fingerprint() const {
        uint64_t h = 0xcbf29ce484222325ull;
        auto mix = [&h](const void *p, size_t n) {
            const unsigned char *b = (const unsigned char *) p;
            size_t i = 0;
            for (; i + 8 <= n; i += 8) { uint64_t w; std::memcpy(&w, b + i, 8); h = (h ^ w) * 0x100000001b3ull; h ^= h >> 29; }
            for (; i < n; ++i) h = (h ^ b[i]) * 0x100000001b3ull;
            h = (h ^ (uint64_t) n) * 0x100000001b3ull;
        };
        auto vec = [&mix](const auto &v) { mix(v.data(), v.size() * sizeof(v[0])); };
        mix(&desc, sizeof(desc));
        vec(shapes); vec(bsdfs); vec(media); vec(volumes); vec(phases); vec(emitters);
        vec(vpos); vec(vnrm); vec(vuv); vec(faces); vec(sensor_xf);
        for (const auto &g : grids) vec(g);
        for (const auto &t : phase_values) vec(t);
        return h;
    }
    uint64_t hash = 0;
    void finish() {
        desc.abi_version = B2VP_ABI_VERSION; desc.struct_size = (uint32_t) sizeof(B2vpSceneDesc);
        desc.n_shapes = (uint32_t) shapes.size(); desc.shapes = shapes.data();
        desc.n_bsdfs = (uint32_t) bsdfs.size(); desc.bsdfs = bsdfs.data();
        desc.n_media = (uint32_t) media.size(); desc.media = media.data();
        desc.n_volumes = (uint32_t) volumes.size(); desc.volumes = volumes.data();
        for (size_t i = 0, g = 0; i < volumes.size(); ++i)
            if (volumes[i].type == B2VP_VOLUME_GRID) volumes[i].data = grids[g++].data();
        for (size_t i = 0; i < phases.size(); ++i)
            if (phases[i].type == B2VP_PHASE_TABULATED) {
                phases[i].n_values = (uint32_t) phase_values[i].size(); phases[i].values = phase_values[i].data();
            }
        desc.n_phases = (uint32_t) phases.size(); desc.phases = phases.data();
        desc.n_emitters = (uint32_t) emitters.size(); desc.emitters = emitters.data();
        desc.n_vertices = (uint32_t) (vpos.size() / 3); desc.n_faces = (uint32_t) (faces.size() / 3);
        desc.vertex_positions = vpos.data(); desc.vertex_normals = vnrm.data(); desc.vertex_texcoords = vuv.data();
        desc.faces = faces.data();
        if (!sensor_xf.empty()) desc.sensor.transforms = sensor_xf.data();
    }
};

/* Reads the object graph of a loaded scene (the read-out include/b2vp.h documents per field). */
template <typename Float, typename Spectrum> struct B200Describe {
    MTS_IMPORT_TYPES(Scene, Sensor, Shape, Mesh, BSDF, Medium, Volume, PhaseFunction, Emitter, Texture, Film, Sampler)
    using ScalarTransform = ScalarTransform4f;
    B200SceneTables &t;
    explicit B200Describe(B200SceneTables &tables) : t(tables) {}

    static std::string cname(const Object *o) { return o->class_()->name(); }
    [[noreturn]] static void unsupported(const std::string &what) {
        Throw("volpath_b200: %s is outside the scope of the B200 volpath (SURVEY.md section 8)", what);
    }
    template <typename M> static B2vpMat4 mat(const M &m) { /* column-major, m[c*4+r] */
        B2vpMat4 o;
        for (int c = 0; c < 4; ++c) for (int r = 0; r < 4; ++r) o.m[c * 4 + r] = (double) m(r, c);
        return o;
    }
    /* a constant texture / spectrum of a mono variant: its value anywhere (src/spectra/uniform.cpp:40-50) */
    static double texture_value(const Texture *tex) {
        SurfaceInteraction3f si = zero<SurfaceInteraction3f>();
        if (cname(tex) != "UniformSpectrum")
            unsupported("texture class " + cname(tex) + " (only constant reflectances)");
        return (double) tex->eval(si, true)[0];
    }

    int phase(const PhaseFunction *p) {
        B2vpPhase o{};
        std::vector<double> values;
        std::string n = cname(p);
        if (n == "IsotropicPhaseFunction") o.type = B2VP_PHASE_ISOTROPIC;
        else if (n == "RayleighPhaseFunction") o.type = B2VP_PHASE_RAYLEIGH;
        else if (n == "HGPhaseFunction") {
            o.type = B2VP_PHASE_HG;
            o.g = (double) static_cast<const HGPhaseFunction<Float, Spectrum> *>(p)->m_g;
        } else if (n == "BlendPhaseFunction") {
            auto *b = static_cast<const BlendPhaseFunction<Float, Spectrum> *>(p);
            o.type = B2VP_PHASE_BLEND;
            for (int k = 0; k < 2; ++k) {
                if (cname(b->m_nested_phase[k].get()) == "BlendPhaseFunction") unsupported("nested blendphase");
                o.child[k] = phase(b->m_nested_phase[k].get());
            }
            o.weight_volume = volume(b->m_weight.get());
        } else if (n == "TabulatedPhaseFunction") {
            o.type = B2VP_PHASE_TABULATED;   /* the values as the plugin holds them (tabphase.cpp:43-60) */
            const auto &pdf = static_cast<const TabulatedPhaseFunction<Float, Spectrum> *>(p)->m_distr.m_pdf;
            for (size_t i = 0; i < pdf.size(); ++i) values.push_back((double) pdf[i]);
        } else unsupported("phase function " + n);
        t.phases.push_back(o);
        t.phase_values.push_back(std::move(values));
        return (int) t.phases.size() - 1;
    }

    int volume(const Volume *v) {
        B2vpVolume o{};
        std::string n = cname(v);
        Interaction3f it = zero<Interaction3f>();
        if (n == "ConstVolume") {
            o.type = B2VP_VOLUME_CONST; o.channels = 1;
            o.value = (double) v->eval(it, true)[0];
            o.max_value = o.value;
        } else if (n.rfind("GridVolumeImpl_", 0) == 0) {   /* GridVolumeImpl_<channels>_<raw> (grid3d.cpp:420-440) */
            /* the members read here do not depend on the Channels / Raw template arguments */
            using G = GridVolumeImpl<Float, Spectrum, 1, false>;
            const G *g = static_cast<const G *>(v);
            o.type = B2VP_VOLUME_GRID;
            o.channels = (int32_t) g->m_metadata.channel_count;
            o.res[0] = g->m_metadata.shape.x(); o.res[1] = g->m_metadata.shape.y(); o.res[2] = g->m_metadata.shape.z();
            o.filter = g->m_filter_type == FilterType::Nearest ? B2VP_FILTER_NEAREST : B2VP_FILTER_TRILINEAR;
            o.wrap = g->m_wrap_mode == WrapMode::Repeat ? B2VP_WRAP_REPEAT
                     : g->m_wrap_mode == WrapMode::Mirror ? B2VP_WRAP_MIRROR : B2VP_WRAP_CLAMP;
            o.max_value = (double) g->m_metadata.max;
            size_t count = (size_t) o.res[0] * o.res[1] * o.res[2] * o.channels;
            std::vector<float> data(count);
            for (size_t i = 0; i < count; ++i) data[i] = (float) g->m_data[i];   /* float32 on disk, widened at load */
            t.grids.push_back(std::move(data));
        } else unsupported("volume class " + n);
        o.world_to_local = mat(v->m_world_to_local.matrix);
        for (int k = 0; k < 3; ++k) { o.bbox_min[k] = (double) v->m_bbox.min[k]; o.bbox_max[k] = (double) v->m_bbox.max[k]; }
        t.volumes.push_back(o);
        return (int) t.volumes.size() - 1;
    }

    int medium(const Medium *m) {
        if (!m) return -1;
        auto it = t.medium_ids.find(m);
        if (it != t.medium_ids.end()) return it->second;
        B2vpMedium o{};
        std::string n = cname(m);
        const Volume *sigma_t, *albedo;
        if (n == "HomogeneousMedium") {
            auto *h = static_cast<const HomogeneousMedium<Float, Spectrum> *>(m);
            o.type = B2VP_MEDIUM_HOMOGENEOUS; sigma_t = h->m_sigmat.get(); albedo = h->m_albedo.get(); o.scale = (double) h->m_scale;
        } else if (n == "HeterogeneousMedium") {
            auto *h = static_cast<const HeterogeneousMedium<Float, Spectrum> *>(m);
            o.type = B2VP_MEDIUM_HETEROGENEOUS; sigma_t = h->m_sigmat.get(); albedo = h->m_albedo.get(); o.scale = (double) h->m_scale;
        } else unsupported("medium " + n);
        o.albedo = volume(albedo);
        o.sigma_t = volume(sigma_t);
        o.has_spectral_extinction = m->has_spectral_extinction();
        o.sample_emitters = m->use_emitter_sampling();
        o.phase = phase(m->phase_function());
        t.media.push_back(o);
        return t.medium_ids[m] = (int) t.media.size() - 1;
    }

    int bsdf(const BSDF *b) {
        auto it = t.bsdf_ids.find(b);
        if (it != t.bsdf_ids.end()) return it->second;
        B2vpBsdf o{};
        std::string n = cname(b);
        if (n == "SmoothDiffuse") {
            o.type = B2VP_BSDF_DIFFUSE;
            o.p[0] = texture_value(static_cast<const SmoothDiffuse<Float, Spectrum> *>(b)->m_reflectance.get());
        } else if (n == "Null") {
            o.type = B2VP_BSDF_NULL;
        } else if (n == "RPV") {
            auto *r = static_cast<const RPV<Float, Spectrum> *>(b);
            o.type = B2VP_BSDF_RPV;
            o.p[0] = texture_value(r->m_rho_0.get()); o.p[1] = texture_value(r->m_k.get());
            o.p[2] = texture_value(r->m_g.get()); o.p[3] = texture_value(r->m_rho_c.get());
        } else if (n == "BiLambertian") {
            auto *r = static_cast<const BiLambertian<Float, Spectrum> *>(b);
            o.type = B2VP_BSDF_BILAMBERTIAN;
            o.p[0] = texture_value(r->m_reflectance.get()); o.p[1] = texture_value(r->m_transmittance.get());
        } else unsupported("BSDF " + n);
        t.bsdfs.push_back(o);
        return t.bsdf_ids[b] = (int) t.bsdfs.size() - 1;
    }

    /* scene->shapes(): meshes arrive in world space (cube.cpp:104-107, obj.cpp:172); rectangles keep to_world */
    int shape(const Shape *s) {
        B2vpShape o{};
        std::string n = cname(s);
        o.bsdf = bsdf(s->bsdf());
        o.interior_medium = medium(s->interior_medium());
        o.exterior_medium = medium(s->exterior_medium());
        if (s->is_emitter()) unsupported("area emitters");
        if (s->is_mesh()) {
            const Mesh *m = static_cast<const Mesh *>(s);
            o.type = B2VP_SHAPE_MESH;
            const uint32_t nv = (uint32_t) m->vertex_count(), nf = (uint32_t) m->face_count();
            o.vertex_offset = (uint32_t) (t.vpos.size() / 3); o.vertex_count = nv;
            o.face_offset = (uint32_t) (t.faces.size() / 3); o.face_count = nf;
            o.has_vertex_normals = m->has_vertex_normals(); o.has_vertex_texcoords = m->has_vertex_texcoords();
            const float *p = m->m_vertex_positions_buf.data();
            t.vpos.insert(t.vpos.end(), p, p + (size_t) nv * 3);
            if (o.has_vertex_normals) { const float *q = m->m_vertex_normals_buf.data(); t.vnrm.insert(t.vnrm.end(), q, q + (size_t) nv * 3); }
            else t.vnrm.insert(t.vnrm.end(), (size_t) nv * 3, 0.f);
            if (o.has_vertex_texcoords) { const float *q = m->m_vertex_texcoords_buf.data(); t.vuv.insert(t.vuv.end(), q, q + (size_t) nv * 2); }
            else t.vuv.insert(t.vuv.end(), (size_t) nv * 2, 0.f);
            const uint32_t *f = m->m_faces_buf.data();
            t.faces.insert(t.faces.end(), f, f + (size_t) nf * 3);
            ScalarTransform id;
            o.to_world = mat(id.matrix); o.to_object = mat(id.matrix);
        } else if (n == "Rectangle") {
            auto *r = static_cast<const Rectangle<Float, Spectrum> *>(s);
            o.type = B2VP_SHAPE_RECTANGLE;
            o.to_world = mat(r->m_to_world.matrix); o.to_object = mat(r->m_to_object.matrix);
            for (int k = 0; k < 3; ++k) {
                o.frame_s[k] = (double) r->m_frame.s[k]; o.frame_t[k] = (double) r->m_frame.t[k];
                o.frame_n[k] = (double) r->m_frame.n[k];
            }
            o.inv_surface_area = (double) r->m_inv_surface_area;
        } else if (n == "Sphere") {
            auto *r = static_cast<const Sphere<Float, Spectrum> *>(s);
            o.type = B2VP_SHAPE_SPHERE;
            o.to_world = mat(r->m_to_world.matrix); o.to_object = mat(r->m_to_object.matrix);
            for (int k = 0; k < 3; ++k) o.center[k] = (double) r->m_center[k];
            o.radius = (double) r->m_radius;
            o.flip_normals = r->m_flip_normals;
            o.inv_surface_area = (double) r->m_inv_surface_area;
        } else unsupported("shape " + n);
        t.shapes.push_back(o);
        return t.shape_ids[s] = (int) t.shapes.size() - 1;
    }

    void emitter(const Emitter *e) {
        if (cname(e) == "ConstantBackgroundEmitter") { /* the environment (constant.cpp:35-40) */
            B2vpEmitter o{};
            o.type = B2VP_EMITTER_CONSTANT;
            o.irradiance = texture_value(static_cast<const ConstantBackgroundEmitter<Float, Spectrum> *>(e)->m_radiance.get());
            t.emitters.push_back(o);
            return;
        }
        if (cname(e) != "DirectionalEmitter") unsupported("emitter " + cname(e));
        auto *d = static_cast<const DirectionalEmitter<Float, Spectrum> *>(e);
        B2vpEmitter o{};
        o.type = B2VP_EMITTER_DIRECTIONAL;
        /* directional.cpp:112: the direction is to_world * (0, 0, 1), to_world in single precision (transform.h:366) */
        auto trafo = d->m_world_transform->eval(0.f);
        ScalarVector3f dir = trafo.transform_affine(ScalarVector3f(0.f, 0.f, 1.f));
        for (int k = 0; k < 3; ++k) o.direction[k] = (double) dir[k];
        o.irradiance = texture_value(d->m_irradiance.get());
        t.emitters.push_back(o);
    }

    void sensor(const Sensor *sn) {
        B2vpSensor o{};
        std::string n = cname(sn);
        const Film *film = sn->film();
        const Sampler *sampler = sn->m_sampler.get();
        if (cname(film) != "HDRFilm") unsupported("film " + cname(film));
        if (cname(sampler) != "IndependentSampler") unsupported("sampler " + cname(sampler));
        {   /* the film's reconstruction filter, through its public interface only (rfilter.h:55-71): the table is
               read back entry by entry with eval_discretized() at the centre of each of its cells */
            const ReconstructionFilter<Float, Spectrum> *rf = film->reconstruction_filter();
            const ScalarFloat radius = rf->radius(), scale = ScalarFloat(MTS_FILTER_RESOLUTION) / radius;   /* rfilter.cpp:19 */
            o.rfilter_radius = (double) radius; o.rfilter_scale = (double) scale;
            o.rfilter_border = rf->border_size();
            o.rfilter_taps = enoki::ceil2int<uint32_t>((radius - 2.f * math::RayEpsilon<ScalarFloat>) * 2.f);   /* imageblock.cpp:122 */
            for (int i = 0; i <= MTS_FILTER_RESOLUTION; ++i)
                o.rfilter_values[i] = (double) rf->eval_discretized((ScalarFloat(i) + ScalarFloat(0.5)) / scale);
            static_assert(MTS_FILTER_RESOLUTION == B2VP_RFILTER_RESOLUTION, "filter table size");
            if (radius > 0.5f + math::RayEpsilon<Float> && o.rfilter_taps > B2VP_RFILTER_MAX_TAPS)
                unsupported("reconstruction filters wider than " + std::to_string(B2VP_RFILTER_MAX_TAPS / 2) + " pixels");
        }
        o.film_size[0] = film->size().x(); o.film_size[1] = film->size().y();
        o.crop_offset[0] = film->crop_offset().x(); o.crop_offset[1] = film->crop_offset().y();
        o.crop_size[0] = film->crop_size().x(); o.crop_size[1] = film->crop_size().y();
        o.sample_count = sampler->sample_count();
        o.sampler_seed = sampler->m_base_seed;
        o.shutter_open = (double) sn->shutter_open(); o.shutter_open_time = (double) sn->shutter_open_time();
        o.medium = medium(sn->medium());
        o.target_shape = o.origin_shape = -1;
        o.needs_aperture_sample = sn->needs_aperture_sample();
        if (sn->m_world_transform) {   /* AnimatedTransform: single precision; shape sensors have none of their own */
            auto world = sn->m_world_transform->eval(0.f);
            ScalarTransform4f world_s(world);
            o.to_world = mat(world_s.matrix);
        } else {
            ScalarTransform4f id;
            o.to_world = mat(id.matrix);
        }
        auto target_point = [&](const auto &p) { for (int k = 0; k < 3; ++k) o.target_point[k] = (double) p[k]; };
        /* a shape given to the sensor: one of the scene's, or a shape nested in the sensor, which is not a child of the
           scene (scene.cpp:23-40): used by the sensor, never hit by a path.  The device handles rectangles */
        auto own_shape = [&](const Shape *shape, const char *role) {
            auto it = t.shape_ids.find(shape);
            int index;
            if (it != t.shape_ids.end()) index = it->second;
            else {
                index = this->shape(shape);
                t.shapes[index].sample_only = 1;
            }
            if (t.shapes[index].type != B2VP_SHAPE_RECTANGLE) unsupported(std::string("a sensor ") + role + " shape other than a rectangle");
            return index;
        };
        /* a shape as target: area sampling of the shape (distant.cpp:345-354) */
        auto target_shape = [&](const Shape *shape) { o.target_type = B2VP_TARGET_SHAPE; o.target_shape = own_shape(shape, "target"); };
        /* a shape as ray origin: the target projected onto it along the ray (distant.cpp:366-375) */
        auto origin_shape = [&](const Shape *shape) { o.origin_type = B2VP_ORIGIN_SHAPE; o.origin_shape = own_shape(shape, "origin"); };
        if (n.rfind("DistantSensor_", 0) == 0) {
            /* the members read do not depend on the target / origin template arguments */
            using D = DistantSensorImpl<Float, Spectrum, RayTargetType::Point, RayOriginType::BoundingSphere>;
            const D *d = static_cast<const D *>(sn);
            o.type = B2VP_SENSOR_DISTANT;
            o.flip_directions = d->m_flip_directions;
            o.direction_type = d->m_direction_type == RayDirectionType::Single ? B2VP_DIR_SINGLE
                               : d->m_direction_type == RayDirectionType::SampleWidth ? B2VP_DIR_SAMPLE_WIDTH : B2VP_DIR_SAMPLE_ALL;
            o.origin_type = B2VP_ORIGIN_BSPHERE;
            if (n.rfind("DistantSensor_Point_", 0) == 0) { o.target_type = B2VP_TARGET_POINT; target_point(d->m_ray_target_point); }
            else if (n.rfind("DistantSensor_None_", 0) == 0) o.target_type = B2VP_TARGET_NONE;
            else if (n.rfind("DistantSensor_Shape_", 0) == 0) target_shape(d->m_ray_target_shape.get());
            else unsupported("distant sensor " + n);
            if (n.size() > 6 && n.compare(n.size() - 6, 6, "_Shape") == 0) origin_shape(d->m_ray_origin_shape.get());
            else if (n.size() < 15 || n.compare(n.size() - 15, 15, "_BoundingSphere") != 0) unsupported("distant sensor " + n);
        } else if (n.rfind("DistantFluxSensor_", 0) == 0) {
            using D = DistantFluxSensorImpl<Float, Spectrum, DistantFluxRayTargetType::Point>;
            const D *d = static_cast<const D *>(sn);
            o.type = B2VP_SENSOR_DISTANTFLUX;
            if (n == "DistantFluxSensor_Point") { o.target_type = B2VP_TARGET_POINT; target_point(d->m_target_point); }
            else if (n == "DistantFluxSensor_NoTarget") o.target_type = B2VP_TARGET_NONE;
            else if (n == "DistantFluxSensor_Shape") target_shape(d->m_target_shape.get());
            else unsupported("distantflux sensor " + n);
            o.origin_type = B2VP_ORIGIN_BSPHERE;
            if (d->m_origin_shape) origin_shape(d->m_origin_shape.get());
            for (int k = 0; k < 3; ++k) o.reference_normal[k] = (double) d->m_reference_normal[k];
        } else if (n == "IrradianceMeter") {   /* attached to a shape (shape.cpp:47-54): a rectangle here */
            o.type = B2VP_SENSOR_IRRADIANCEMETER;
            auto it = t.shape_ids.find(sn->shape());
            if (it == t.shape_ids.end() || t.shapes[it->second].type != B2VP_SHAPE_RECTANGLE)
                unsupported("an irradiancemeter on a shape other than a rectangle");
            o.origin_shape = it->second;
        } else if (n == "RadianceMeter") {
            o.type = B2VP_SENSOR_RADIANCEMETER;
        } else if (n.rfind("MultiDistantSensor_", 0) == 0) {
            using D = MultiDistantSensorImpl<Float, Spectrum, MultiDistantRayTargetType::Point>;
            const D *d = static_cast<const D *>(sn);
            o.type = B2VP_SENSOR_MDISTANT;
            if (n == "MultiDistantSensor_Point") { o.target_type = B2VP_TARGET_POINT; target_point(d->m_target_point); }
            else if (n == "MultiDistantSensor_NoTarget") o.target_type = B2VP_TARGET_NONE;
            else if (n == "MultiDistantSensor_Shape") target_shape(d->m_target_shape.get());
            else unsupported("mdistant sensor " + n);
            o.n_transforms = (int32_t) d->m_sensor_count;
            for (size_t i = 0; i < d->m_sensor_count * 16; ++i) t.sensor_xf.push_back((double) d->m_transforms[i]);
        } else if (n == "MultiRadianceMeter") {
            auto *d = static_cast<const MultiRadianceMeter<Float, Spectrum> *>(sn);
            o.type = B2VP_SENSOR_MRADIANCEMETER;
            o.n_transforms = (int32_t) d->m_sensor_count;
            for (size_t i = 0; i < d->m_sensor_count * 16; ++i) t.sensor_xf.push_back((double) d->m_transforms[i]);
        } else unsupported("sensor " + n);
        t.desc.sensor = o;
    }

    /* ---- load-time self-check ------------------------------------------------------------------------------
       describe() reads private members of classes that live in OTHER shared objects (the reference's plugins), through
       class definitions compiled into this one.  That is only right while both sides were compiled from the same
       sources with the same flags; a drift would be silent memory corruption.  So before the first render, canary
       objects of every such class are created through the reference's own PluginManager with known parameters and read
       back through the very code paths describe() uses: any disagreement throws, naming the class.  (A plugin that is
       not installed is skipped: scenes cannot contain it either.) */
    template <typename T> static ref<T> canary(const Properties &props) {
        try {
            ref<T> o = PluginManager::instance()->create_object<T>(props);
            /* wrapper plugins (distant, mdistant, gridvolume, ...) stand for the object expand() returns, as in the
               XML loader (src/libcore/xml.cpp) */
            std::vector<ref<Object>> ex = o->expand();
            if (!ex.empty()) o = static_cast<T *>(ex[0].get());
            return o;
        } catch (const std::exception &) {
            return nullptr;
        }
    }
    static void expect(bool ok, const char *what) {
        if (!ok)
            Throw("volpath_b200: self-check failed for %s: this plugin and the reference's plugins were not built from "
                  "the same sources / flags (class layouts differ); refusing to read scene objects", what);
    }
    static bool near(double a, double b) { return std::fabs(a - b) <= 1e-6 * (1.0 + std::fabs(b)); }
    static void self_check() {
        B200SceneTables scratch;
        B200Describe d(scratch);
        const bool trace_on = std::getenv("B2VP_SELFCHECK_TRACE") != nullptr;   /* diagnostics: which canary is being read */
        auto trace = [trace_on, &scratch](const char *name) {
            if (trace_on) std::fprintf(stderr, "[volpath_b200 self-check] %s\n", name);
            /* the tables remember objects by address, and a canary may be allocated where the previous one was freed */
            scratch.bsdf_ids.clear(); scratch.medium_ids.clear(); scratch.shape_ids.clear();
        };
        { trace("hg"); Properties p("hg"); p.set_float("g", 0.3125f);
          if (auto o = canary<PhaseFunction>(p)) { int i = d.phase(o.get()); expect(scratch.phases[i].type == B2VP_PHASE_HG && near(scratch.phases[i].g, 0.3125), "HGPhaseFunction"); } }
        { trace("tabphase"); Properties p("tabphase"); p.set_string("values", "0.5, 1.0, 2.0");
          if (auto o = canary<PhaseFunction>(p)) { int i = d.phase(o.get()); const auto &v = scratch.phase_values[i];
              expect(v.size() == 3 && near(v[0], 0.5) && near(v[1], 1.0) && near(v[2], 2.0), "TabulatedPhaseFunction"); } }
        { trace("diffuse"); Properties p("diffuse"); p.set_float("reflectance", 0.375f);
          if (auto o = canary<BSDF>(p)) { int i = d.bsdf(o.get()); expect(scratch.bsdfs[i].type == B2VP_BSDF_DIFFUSE && near(scratch.bsdfs[i].p[0], 0.375), "SmoothDiffuse"); } }
        { trace("rpv"); Properties p("rpv"); p.set_float("rho_0", 0.125f); p.set_float("k", 0.75f); p.set_float("g", -0.25f); p.set_float("rho_c", 0.0625f);
          if (auto o = canary<BSDF>(p)) { int i = d.bsdf(o.get()); const auto &b = scratch.bsdfs[i];
              expect(b.type == B2VP_BSDF_RPV && near(b.p[0], 0.125) && near(b.p[1], 0.75) && near(b.p[2], -0.25) && near(b.p[3], 0.0625), "RPV"); } }
        { trace("bilambertian"); Properties p("bilambertian"); p.set_float("reflectance", 0.375f); p.set_float("transmittance", 0.25f);
          if (auto o = canary<BSDF>(p)) { int i = d.bsdf(o.get()); const auto &b = scratch.bsdfs[i];
              expect(b.type == B2VP_BSDF_BILAMBERTIAN && near(b.p[0], 0.375) && near(b.p[1], 0.25), "BiLambertian"); } }
        { trace("rectangle"); Properties p("rectangle");
          p.set_transform("to_world", ScalarTransform4f::translate(ScalarVector3f(1.f, 2.f, 3.f)) * ScalarTransform4f::scale(ScalarVector3f(2.f, 4.f, 1.f)));
          if (auto o = canary<Shape>(p)) { int i = d.shape(o.get()); const auto &r = scratch.shapes[i];
              expect(r.type == B2VP_SHAPE_RECTANGLE && near(r.to_world.m[12], 1.0) && near(r.to_world.m[13], 2.0) && near(r.to_world.m[14], 3.0) &&
                     near(r.frame_s[0], 4.0) && near(r.frame_t[1], 8.0) && near(r.frame_n[2], 1.0) && near(r.inv_surface_area, 1.0 / 32.0), "Rectangle"); } }
        { trace("sphere"); Properties p("sphere"); p.set_array3f("center", Properties::Array3f(1.f, 2.f, 3.f)); p.set_float("radius", 0.5f);
          if (auto o = canary<Shape>(p)) { int i = d.shape(o.get()); const auto &r = scratch.shapes[i];
              expect(r.type == B2VP_SHAPE_SPHERE && near(r.center[0], 1.0) && near(r.center[1], 2.0) && near(r.center[2], 3.0) && near(r.radius, 0.5) &&
                     r.flip_normals == 0 && near(r.inv_surface_area, 1.0 / (4.0 * 3.14159265358979323846 * 0.25)), "Sphere"); } }
        { Properties ph("isotropic");
          auto pf = canary<PhaseFunction>(ph);
          trace("homogeneous"); Properties p("homogeneous"); p.set_float("sigma_t", 0.625f); p.set_float("albedo", 0.875f); p.set_float("scale", 2.f);
          if (pf) p.set_object("phase", pf.get());
          if (auto o = canary<Medium>(p)) { int i = d.medium(o.get()); const auto &m = scratch.media[i];
              expect(m.type == B2VP_MEDIUM_HOMOGENEOUS && near(m.scale, 2.0) && near(scratch.volumes[m.sigma_t].value, 0.625) &&
                     near(scratch.volumes[m.albedo].value, 0.875) && scratch.phases[m.phase].type == B2VP_PHASE_ISOTROPIC, "HomogeneousMedium"); } }
        { trace("directional"); Properties p("directional"); p.set_array3f("direction", Properties::Array3f(0.f, 0.f, -1.f)); p.set_float("irradiance", 1.5f);
          if (auto o = canary<Emitter>(p)) { d.emitter(o.get()); const auto &e = scratch.emitters.back();
              expect(e.type == B2VP_EMITTER_DIRECTIONAL && near(e.direction[0], 0.0) && near(e.direction[1], 0.0) && near(e.direction[2], -1.0) && near(e.irradiance, 1.5), "DirectionalEmitter"); } }
        { trace("constant"); Properties p("constant"); p.set_float("radiance", 0.25f);
          if (auto o = canary<Emitter>(p)) { d.emitter(o.get()); const auto &e = scratch.emitters.back();
              expect(e.type == B2VP_EMITTER_CONSTANT && near(e.irradiance, 0.25), "ConstantBackgroundEmitter"); } }
        { trace("gridvolume");   /* a 2 x 2 x 2 version-3 .vol file (volume_data.h:42-102) with the values 1 .. 8 */
          char path[] = "/tmp/b2vp_canary_XXXXXX";
          const int fd = mkstemp(path);
          if (fd >= 0) {
              unsigned char head[48] = { 'V', 'O', 'L', 3 };
              const int32_t meta[5] = { 1, 2, 2, 2, 1 };
              const float box[6] = { 0, 0, 0, 1, 1, 1 }, vals[8] = { 1, 2, 3, 4, 5, 6, 7, 8 };
              std::memcpy(head + 4, meta, 20); std::memcpy(head + 24, box, 24);
              const bool written = write(fd, head, 48) == 48 && write(fd, vals, 32) == 32;
              close(fd);
              Properties p("gridvolume"); p.set_string("filename", path); p.set_string("filter_type", "nearest"); p.set_string("wrap_mode", "mirror");
              if (written) if (auto o = canary<Volume>(p)) { int i = d.volume(o.get()); const auto &v = scratch.volumes[i];
                  expect(v.type == B2VP_VOLUME_GRID && v.channels == 1 && v.res[0] == 2 && v.res[1] == 2 && v.res[2] == 2 && v.filter == B2VP_FILTER_NEAREST &&
                         v.wrap == B2VP_WRAP_MIRROR && near(v.max_value, 8.0) && scratch.grids.back().size() == 8 && scratch.grids.back()[0] == 1.f &&
                         scratch.grids.back()[7] == 8.f && near(v.bbox_max[2], 1.0), "GridVolumeImpl"); }
              unlink(path);
          } }
        /* sensors: film and sampler children, then the members of the sensor classes themselves */
        auto film_of = [&trace](int w, int h) {
            trace("box"); Properties rf("box"); auto filter = canary<mitsuba::ReconstructionFilter<Float, Spectrum>>(rf);
            trace("hdrfilm"); Properties p("hdrfilm"); p.set_int("width", w); p.set_int("height", h);
            if (filter) p.set_object("rfilter", filter.get());
            return canary<Film>(p);
        };
        auto sampler_of = [&trace]() { trace("independent"); Properties p("independent"); p.set_int("sample_count", 7); p.set_int("seed", 11); return canary<Sampler>(p); };
        { auto film = film_of(3, 1); auto smp = sampler_of();
          trace("distant"); Properties p("distant"); p.set_array3f("direction", Properties::Array3f(0.f, 0.f, 1.f)); p.set_array3f("ray_target", Properties::Array3f(0.5f, -0.25f, 0.125f));
          if (film && smp) { p.set_object("film", film.get()); p.set_object("sampler", smp.get());
              if (auto o = canary<Sensor>(p)) { d.sensor(o.get()); const auto &sn = scratch.desc.sensor;
                  expect(sn.type == B2VP_SENSOR_DISTANT && sn.direction_type == B2VP_DIR_SAMPLE_WIDTH && sn.target_type == B2VP_TARGET_POINT &&
                         near(sn.target_point[0], 0.5) && near(sn.target_point[1], -0.25) && near(sn.target_point[2], 0.125) &&
                         sn.film_size[0] == 3 && sn.film_size[1] == 1 && sn.sample_count == 7 && sn.sampler_seed == 11 && sn.flip_directions == 0,
                         "DistantSensor / HDRFilm / IndependentSampler"); } } }
        { auto film = film_of(2, 1); auto smp = sampler_of();
          trace("mdistant"); Properties p("mdistant"); p.set_string("directions", "0, 0, -1, 1, 0, -1"); p.set_array3f("target", Properties::Array3f(0.5f, -0.25f, 0.125f));
          if (film && smp) { p.set_object("film", film.get()); p.set_object("sampler", smp.get());
              scratch.sensor_xf.clear();
              if (auto o = canary<Sensor>(p)) { d.sensor(o.get()); const auto &sn = scratch.desc.sensor;
                  expect(sn.type == B2VP_SENSOR_MDISTANT && sn.n_transforms == 2 && scratch.sensor_xf.size() == 32 && sn.target_type == B2VP_TARGET_POINT &&
                         near(sn.target_point[0], 0.5) && near(scratch.sensor_xf[10], -1.0) /* first look-at: z axis = (0, 0, -1) */, "MultiDistantSensor"); } } }
        { auto film = film_of(2, 1); auto smp = sampler_of();
          trace("mradiancemeter"); Properties p("mradiancemeter"); p.set_string("origins", "1, 2, 3, 4, 5, 6"); p.set_string("directions", "0, 0, -1, 0, 0, 1");
          if (film && smp) { p.set_object("film", film.get()); p.set_object("sampler", smp.get());
              scratch.sensor_xf.clear();
              if (auto o = canary<Sensor>(p)) { d.sensor(o.get()); const auto &sn = scratch.desc.sensor;
                  expect(sn.type == B2VP_SENSOR_MRADIANCEMETER && sn.n_transforms == 2 && scratch.sensor_xf.size() == 32 &&
                         near(scratch.sensor_xf[12], 1.0) && near(scratch.sensor_xf[13], 2.0) && near(scratch.sensor_xf[14], 3.0) &&
                         near(scratch.sensor_xf[16 + 12], 4.0) && near(scratch.sensor_xf[16 + 14], 6.0), "MultiRadianceMeter"); } } }
    }

    void scene(const Scene *sc, const Sensor *sn) {
        t.desc.precision = std::is_same_v<ScalarFloat, double> ? B2VP_PREC_F64 : B2VP_PREC_F32;
        for (auto &s : sc->shapes()) shape(s.get());
        for (auto &e : sc->emitters()) emitter(e.get());
        sensor(sn);
        /* Scene::bbox() and the padded bounding sphere of distant.cpp:292-297 / directional.cpp:68-73 */
        ScalarBoundingBox3f bbox = sc->bbox();
        auto bs = bbox.bounding_sphere();
        bs.radius = enoki::max(math::RayEpsilon<Float>, bs.radius * (1.f + math::RayEpsilon<Float>));
        for (int k = 0; k < 3; ++k) {
            t.desc.bbox_min[k] = (double) bbox.min[k]; t.desc.bbox_max[k] = (double) bbox.max[k];
            t.desc.bsphere_center[k] = (double) bs.center[k];
        }
        t.desc.bsphere_radius = (double) bs.radius;
    }
};

/* The integrator plugin.  Properties are those of `volpath` (src/integrators/volpath.cpp:19-25 over
   MonteCarloIntegrator / SamplingIntegrator, src/librender/integrator.cpp:23-39,302-315): max_depth, rr_depth,
   hide_emitters, block_size, samples_per_pass, timeout; plus
     device   CUDA ordinal (default 0)
     seeding  "wavefront" (default): one PCG32 stream per sample, (state, stream) = (tea64(seed, i), tea64(i, seed)),
              the seeding of the reference's own wavefront variants (sampler.cpp:89-92, integrator.cpp:147-153):
              every sample is an independent unit of work, which is what fills a GPU.
              "scalar": the scalar variants' seeding (one stream per pixel, seed = block_id * bs^2 + morton index,
              all samples of a pixel drawn serially, integrator.cpp:89-130,198): the film then equals the one
              scalar_mono / scalar_mono_double render on the CPU (to the parity tolerance); parallelism is limited to
              the number of pixels, so it is for verification, not for speed. */
template <typename Float, typename Spectrum>
class B200VolumetricPathIntegrator final : public MonteCarloIntegrator<Float, Spectrum> {
public:
    MTS_IMPORT_BASE(MonteCarloIntegrator, m_max_depth, m_rr_depth, m_hide_emitters, m_stop, m_block_size,
                    m_samples_per_pass, m_render_timer, m_timeout)
    MTS_IMPORT_TYPES(Scene, Sensor, Film, ImageBlock, Sampler, Medium)

    B200VolumetricPathIntegrator(const Properties &props) : Base(props) {
        if constexpr (!is_monochromatic_v<Spectrum> || is_array_v<Float>)
            Throw("volpath_b200 supports the scalar_mono and scalar_mono_double variants only");
        m_device = props.int_("device", 0);
        std::string seeding = props.string("seeding", "wavefront");
        if (seeding == "wavefront") m_seed_mode = B2VP_SEED_WAVEFRONT;
        else if (seeding == "scalar") m_seed_mode = B2VP_SEED_SCALAR_BLOCKS;
        else Throw("volpath_b200: seeding must be \"wavefront\" or \"scalar\"");
        /* moment=true: the film also gets the channels a `moment` integrator around this one would add
           (src/integrators/moment.cpp:39-60): nested.X/Y/Z and m2_nested.X/Y/Z, for per-pixel variance estimates.
           (A `moment` plugin cannot wrap this integrator itself: it calls sample(), and the batch boundary here is
           render().) */
        m_moment = props.bool_("moment", false);
        /* algorithm: "volpath" (default) or "volpathmis" (src/integrators/volpathmis.cpp; use_spectral_mis as there) */
        /* B2VP_DEFAULT_ALGORITHM: the same source is also built as the drop-in `volpathmis.so` (oracle/Makefile.ref) */
        std::string algorithm = props.string("algorithm", B2VP_DEFAULT_ALGORITHM);
        if (algorithm == "volpathmis") m_mis = props.bool_("use_spectral_mis", true) ? 1 : 2;
        else if (algorithm != "volpath") Throw("volpath_b200: algorithm must be \"volpath\" or \"volpathmis\"");
        /* math: "fast" (default in scalar_mono: the fp32 tolerance mode, B2VP_MATH_FAST) or "exact" (every fp32
           operation in the oracle's order; the only mode in scalar_mono_double) */
        std::string math = props.string("math", std::is_same_v<scalar_t<Float>, double> ? "exact" : "fast");
        if (math == "fast") m_math = B2VP_MATH_FAST;
        else if (math == "exact") m_math = B2VP_MATH_EXACT;
        else Throw("volpath_b200: math must be \"fast\" or \"exact\"");
        if (m_math == B2VP_MATH_FAST && std::is_same_v<scalar_t<Float>, double>)
            m_math = B2VP_MATH_EXACT;   /* no fast instance in double precision */
        /* majorant_grid: 0 (default) = the reference's global majorant; see B2vpRenderArgs::majorant_grid */
        m_majorant_grid = (uint32_t) props.int_("majorant_grid", 0);
        /* n_devices: render on this many GPUs (device, then the others in ascending order), one NCCL film reduce */
        int n_dev = props.int_("n_devices", 1);
        if (n_dev < 1 || (n_dev > 1 && n_dev > b2vp_device_count())) Throw("volpath_b200: n_devices = %d, %d CUDA devices present", n_dev, b2vp_device_count());
        m_devices.push_back(m_device);
        for (int d = 0; (int) m_devices.size() < n_dev; ++d)
            if (d != m_device) m_devices.push_back(d);
    }
    ~B200VolumetricPathIntegrator() { release(); }

    /* Integrator::render (integrator.h:42); replaces SamplingIntegrator::render (integrator.cpp:51-179) */
    bool render(Scene *scene, Sensor *sensor) override {
        m_stop = false;
        B200SceneTables tables;
        describe(scene, sensor, tables);
        {
            /* The device scene (grids, BVH, tables in HBM) stays resident between render() calls while the flattened
               scene is unchanged -- the object graph is re-read every time (cheap, host only), so any edit through
               traverse() / parameters_changed, a new sensor or other integrator properties is seen without the
               plugin having to be told. */
            std::lock_guard<std::mutex> lock(m_mutex);
            if (!m_scene || tables.hash != m_scene_hash) {
                release_locked();
                check(b2vp_scene_create(&tables.desc, m_device, &m_scene));
                m_scene_hash = tables.hash;
            }
        }

        ref<Film> film = sensor->film();
        std::vector<std::string> channels = { "X", "Y", "Z", "A", "W" };   /* integrator.cpp:67-73 */
        if (m_moment)
            for (const char *name : { "nested.X", "nested.Y", "nested.Z", "m2_nested.X", "m2_nested.Y", "m2_nested.Z" })
                channels.push_back(name);
        film->prepare(channels);
        ScalarVector2i size = film->crop_size();
        /* b2vp_render returns the film of the crop window with the reconstruction filter applied and the part of the
           footprints outside the window already dropped, which is what the film's put(block) would do to a bordered
           block (imageblock.cpp:52-76): so this block has no border */
        ref<ImageBlock> block = new ImageBlock(size, channels.size(), film->reconstruction_filter(), false, true, false);
        block->set_offset(film->crop_offset());
        block->clear();

        m_render_timer.reset();
        B2vpRenderArgs args{};
        args.abi_version = B2VP_ABI_VERSION; args.struct_size = (uint32_t) sizeof(args);
        args.device = m_device;
        args.precision = tables.desc.precision;
        args.seed_mode = m_seed_mode;
        args.clear_film = 1;
        args.totals = &m_totals;
        args.math_mode = m_math;
        args.majorant_grid = m_majorant_grid;
        args.timeout_s = (double) m_timeout;           /* integrator.h:143-146: <= 0 = none */
        if (m_devices.size() > 1) { args.devices = m_devices.data(); args.n_devices = (uint32_t) m_devices.size(); }
        int rc = b2vp_render(m_scene, &args, block->data().data());   /* host XYZAW buffer, crop_size x 5 */
        if (rc == B2VP_ERR_CANCELLED) return false;
        check(rc);
        film->put(block);                              /* as the reference's wavefront branch, integrator.cpp:155-171 */
        if (!m_stop)
            Log(Info, "Processing finished. (took %s; %d paths, %d kernel launches, %.2f ms on the GPU)",
                util::time_string(m_render_timer.value(), true), (uint64_t) m_totals.n_paths,
                (uint32_t) m_totals.n_launches, m_totals.gpu_ms);
        return !m_stop;
    }

    /* integrator.h:44-51: may come from another thread while render() runs; m_scene is only replaced under the lock */
    void cancel() override {
        m_stop = true;
        std::lock_guard<std::mutex> lock(m_mutex);
        if (m_scene) b2vp_cancel(m_scene);
    }

    /* the batch boundary is render(): a single-sample entry point does not exist on the GPU path */
    std::pair<Spectrum, Mask> sample(const Scene *, Sampler *, const RayDifferential3f &, const Medium *, Float *,
                                     Mask) const override {
        NotImplementedError("sample");
    }

    /* object graph -> POD tables of include/b2vp.h */
    void describe(const Scene *scene, const Sensor *sensor, B200SceneTables &tables) const {
        static std::once_flag checked;   /* per variant (this is a class template): canaries before any real object */
        std::call_once(checked, [] { B200Describe<Float, Spectrum>::self_check(); });
        B200Describe<Float, Spectrum> d(tables);
        d.scene(scene, sensor);
        B2vpIntegrator &it = tables.desc.integrator;
        it.max_depth = m_max_depth; it.rr_depth = m_rr_depth; it.hide_emitters = m_hide_emitters;
        it.block_size = m_block_size;
        it.samples_per_pass = (uint32_t) m_samples_per_pass;
        it.arg_order_ltr = 0;   /* GCC evaluates the sample arguments right to left (include/b2vp.h) */
        it.moment = m_moment ? 1u : 0u;
        it.mis = (uint32_t) m_mis;
        tables.hash = tables.fingerprint();
        tables.finish();
    }

    std::string to_string() const override {
        return tfm::format("B200VolumetricPathIntegrator[\n  max_depth = %i,\n  rr_depth = %i,\n  device = %i\n]",
                           m_max_depth, m_rr_depth, m_device);
    }
    MTS_DECLARE_CLASS()
private:
    static void check(int rc) { if (rc != B2VP_OK) Throw("volpath_b200: %s", b2vp_last_error()); }
    void release_locked() { if (m_scene) b2vp_scene_destroy(m_scene); m_scene = nullptr; }
    void release() { std::lock_guard<std::mutex> lock(m_mutex); release_locked(); }
    std::mutex m_mutex;
    b2vp_scene_t *m_scene = nullptr;
    uint64_t m_scene_hash = 0;
    int m_math = B2VP_MATH_EXACT;
    uint32_t m_majorant_grid = 0;
    std::vector<int32_t> m_devices;
    B2vpCounters m_totals{};
    int m_device = 0;
    int m_seed_mode = B2VP_SEED_WAVEFRONT;
    int m_mis = 0;
    bool m_moment = false;
};

MTS_IMPLEMENT_CLASS_VARIANT(B200VolumetricPathIntegrator, MonteCarloIntegrator)
MTS_EXPORT_PLUGIN(B200VolumetricPathIntegrator, "B200 volumetric path tracer (CUDA, sm_100a)")
NAMESPACE_END(mitsuba)

/* Test hook (tests/ only): flattens a loaded scene and hands the descriptor to a callback, so that the CPU test
   suite can check describe() without a GPU.  `integrator` must be a volpath_b200 instance of the variant. */
extern "C" MTS_EXPORT int b2vp_plugin_describe(void *scene, void *sensor, void *integrator, int is_double,
                                               void (*cb)(const B2vpSceneDesc *, void *), void *user) {
    using namespace mitsuba;
    B200SceneTables tables;
    if (is_double) {
        using F = double; using S = Color<double, 1>;
        static_cast<B200VolumetricPathIntegrator<F, S> *>((Object *) integrator)
            ->describe((const Scene<F, S> *) scene, (const Sensor<F, S> *) sensor, tables);
    } else {
        using F = float; using S = Color<float, 1>;
        static_cast<B200VolumetricPathIntegrator<F, S> *>((Object *) integrator)
            ->describe((const Scene<F, S> *) scene, (const Sensor<F, S> *) sensor, tables);
    }
    cb(&tables.desc, user);
    return 0;
}
