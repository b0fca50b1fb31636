/*
 * b2vp.hpp -- C++17 host layer above the C-ABI of b2vp.h (header only).
 *
 * The reference's host language is C++: scenes are trees of plugins configured through
 * `Properties` (include/mitsuba/core/properties.h) and rendered with
 * `Integrator::render(Scene*, Sensor*)` (include/mitsuba/render/integrator.h:42).  This header gives
 * C++ callers the same vocabulary without the reference's build: one small struct per scoped plugin
 * with the plugin's property names and defaults (file:line per struct), a `SceneBuilder<Float>` that
 * flattens them into the POD `B2vpSceneDesc`, and RAII wrappers of the device scene and the render
 * call that turn error codes into exceptions, as `Throw` does in the reference.
 *
 *   b2vp::SceneBuilder<float> sb;                       // scalar_mono   (double: scalar_mono_double)
 *   int atm = sb.add_medium(b2vp::Homogeneous{.5, 1.0}, b2vp::Phase::rayleigh());
 *   sb.add_cube(b2vp::Mat4<float>::translate(0, 0, .45f) * b2vp::Mat4<float>::scale(1, 1, .55f),
 *               sb.add_bsdf(b2vp::Null{}), atm);
 *   ...
 *   b2vp::DeviceScene dev(sb.finish(), 0);              // b2vp_scene_create
 *   b2vp::Film film = dev.render();                     // b2vp_render
 *
 * Derived quantities (frames, areas, look-at matrices, the bounding sphere) are computed in `Float`,
 * like the reference computes them in the variant's scalar type.
 */
#pragma once
#include "b2vp.h"

#include <algorithm>
#include <array>
#ifdef B2VP_WITH_ZLIB
#include <zlib.h>
#endif
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <limits>
#include <map>
#include <sstream>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace b2vp {

/* `Throw(...)` of the reference (include/mitsuba/core/logger.h) becomes this exception. */
class Error : public std::runtime_error {
public:
    Error(int code, const std::string &what) : std::runtime_error(what), code_(code) {}
    int code() const { return code_; }
private:
    int code_;
};
inline void check(int rc) {
    if (rc != B2VP_OK) throw Error(rc, std::string("b2vp: ") + b2vp_last_error());
}

template <typename F> using Vec3 = std::array<F, 3>;
template <typename F> inline Vec3<F> cross(const Vec3<F> &a, const Vec3<F> &b) {
    return { a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0] };
}
template <typename F> inline F dot(const Vec3<F> &a, const Vec3<F> &b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
template <typename F> inline Vec3<F> normalize(const Vec3<F> &a) {
    F n = std::sqrt(dot(a, a));
    return { a[0] / n, a[1] / n, a[2] / n };
}
/* coordinate_system (include/mitsuba/core/vector.h:116-136) */
template <typename F> inline void coordinate_system(const Vec3<F> &n, Vec3<F> &s, Vec3<F> &t) {
    F sign = std::copysign((F) 1, n[2]);
    F a = -(F) 1 / (sign + n[2]);
    F b = n[0] * n[1] * a;
    s = { n[0] * n[0] * a * sign + (F) 1, b * sign, -n[0] * sign };
    t = { b, n[1] * n[1] * a + sign, -n[1] };
}

/* Column-major 4x4 (m[c*4+r]), the layout of enoki::Matrix and of B2vpMat4. */
template <typename F> struct Mat4 {
    F m[16];
    Mat4() { for (int i = 0; i < 16; ++i) m[i] = (i % 5 == 0) ? (F) 1 : (F) 0; }
    F &operator()(int r, int c) { return m[c * 4 + r]; }
    F operator()(int r, int c) const { return m[c * 4 + r]; }
    static Mat4 translate(F x, F y, F z) { Mat4 t; t(0, 3) = x; t(1, 3) = y; t(2, 3) = z; return t; }
    static Mat4 scale(F x, F y, F z) { Mat4 t; t(0, 0) = x; t(1, 1) = y; t(2, 2) = z; return t; }
    /* Transform::look_at (include/mitsuba/core/transform.h:262-283) */
    static Mat4 look_at(const Vec3<F> &origin, const Vec3<F> &target, const Vec3<F> &up) {
        Vec3<F> dir = normalize(Vec3<F>{ target[0] - origin[0], target[1] - origin[1], target[2] - origin[2] });
        Vec3<F> left = normalize(cross(up, dir));
        Vec3<F> new_up = cross(dir, left);
        Mat4 t;
        for (int r = 0; r < 3; ++r) { t(r, 0) = left[r]; t(r, 1) = new_up[r]; t(r, 2) = dir[r]; t(r, 3) = origin[r]; }
        return t;
    }
    Mat4 operator*(const Mat4 &o) const {
        Mat4 r;
        for (int i = 0; i < 4; ++i)
            for (int j = 0; j < 4; ++j) {
                F s = 0;
                for (int k = 0; k < 4; ++k) s += (*this)(i, k) * o(k, j);
                r(i, j) = s;
            }
        return r;
    }
    Vec3<F> point(const Vec3<F> &p) const {   /* affine */
        return { (*this)(0, 0) * p[0] + (*this)(0, 1) * p[1] + (*this)(0, 2) * p[2] + (*this)(0, 3),
                 (*this)(1, 0) * p[0] + (*this)(1, 1) * p[1] + (*this)(1, 2) * p[2] + (*this)(1, 3),
                 (*this)(2, 0) * p[0] + (*this)(2, 1) * p[1] + (*this)(2, 2) * p[2] + (*this)(2, 3) };
    }
    Vec3<F> vector(const Vec3<F> &v) const {
        return { (*this)(0, 0) * v[0] + (*this)(0, 1) * v[1] + (*this)(0, 2) * v[2],
                 (*this)(1, 0) * v[0] + (*this)(1, 1) * v[1] + (*this)(1, 2) * v[2],
                 (*this)(2, 0) * v[0] + (*this)(2, 1) * v[1] + (*this)(2, 2) * v[2] };
    }
    /* general inverse by Gauss-Jordan with partial pivoting, evaluated in double */
    Mat4 inverse() const {
        double a[4][8];
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) { a[r][c] = (double) (*this)(r, c); a[r][4 + c] = r == c ? 1.0 : 0.0; }
        for (int c = 0; c < 4; ++c) {
            int piv = c;
            for (int r = c + 1; r < 4; ++r) if (std::fabs(a[r][c]) > std::fabs(a[piv][c])) piv = r;
            if (a[piv][c] == 0.0) throw Error(B2VP_ERR_INVALID, "b2vp: singular transform");
            if (piv != c) for (int k = 0; k < 8; ++k) std::swap(a[piv][k], a[c][k]);
            double inv = 1.0 / a[c][c];
            for (int k = 0; k < 8; ++k) a[c][k] *= inv;
            for (int r = 0; r < 4; ++r)
                if (r != c) { double f = a[r][c]; for (int k = 0; k < 8; ++k) a[r][k] -= f * a[c][k]; }
        }
        Mat4 o;
        for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) o(r, c) = (F) a[r][4 + c];
        return o;
    }
    /* The inverse as a Transform4f read from Properties carries it: single precision in every variant
       (properties.h:55-56 -- matrix AND inverse transpose are float members, widened afterwards) */
    Mat4 inverse_f32() const {
        Mat4 o = inverse();
        for (int i = 0; i < 16; ++i) o.m[i] = (F) (float) o.m[i];
        return o;
    }
    /* normals transform with the inverse transpose (of a transform that came through Properties: inverse_f32) */
    Vec3<F> normal(const Vec3<F> &n) const {
        Mat4 it = inverse_f32();
        return { it(0, 0) * n[0] + it(1, 0) * n[1] + it(2, 0) * n[2], it(0, 1) * n[0] + it(1, 1) * n[1] + it(2, 1) * n[2],
                 it(0, 2) * n[0] + it(1, 2) * n[1] + it(2, 2) * n[2] };
    }
    B2vpMat4 pod() const { B2vpMat4 o; for (int i = 0; i < 16; ++i) o.m[i] = (double) m[i]; return o; }
};

/* ---------------------------------------------------------------- plugins ----
   One struct per scoped plugin: members are the plugin's properties with the reference's defaults. */

/* src/phase/isotropic.cpp, rayleigh.cpp, hg.cpp:41 (g defaults to 0.8) */
struct Phase {
    int type = B2VP_PHASE_ISOTROPIC;
    double g = 0.8;
    static Phase isotropic() { return {}; }
    static Phase rayleigh() { Phase p; p.type = B2VP_PHASE_RAYLEIGH; return p; }
    static Phase hg(double g) {
        if (!(g > -1 && g < 1)) throw Error(B2VP_ERR_INVALID, "The asymmetry parameter must lie in the interval (-1, 1)!");
        Phase p; p.type = B2VP_PHASE_HG; p.g = g; return p;
    }
    /* src/phase/tabphase.cpp:43-66: the phase function at regularly spaced cosines on [-1, 1] */
    std::vector<double> values;
    static Phase tabulated(std::vector<double> v) {
        if (v.size() < 2) throw Error(B2VP_ERR_INVALID, "ContinuousDistribution: needs at least two entries!");
        for (double x : v) if (x < 0) throw Error(B2VP_ERR_INVALID, "ContinuousDistribution: entries must be non-negative!");
        Phase p; p.type = B2VP_PHASE_TABULATED; p.values = std::move(v); return p;
    }
    /* src/phase/blendphase.cpp:42-140: (1 - w) a + w b with the weight volume w evaluated where the path scatters.
       The weight is a constant (`weight_value`) or a grid (`weight_grid`: data[((z ny + y) nx + x)], resolution,
       filter / wrap mode and a row-major 4 x 4 to_world, the members of GridVolume below) */
    std::vector<Phase> children;
    double weight_value = 0.5;
    std::vector<float> weight_grid; int weight_res[3] = { 0, 0, 0 };
    std::string weight_filter = "trilinear", weight_wrap = "clamp";
    std::array<double, 16> weight_to_world{ 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1 };
    static Phase blend(const Phase &a, const Phase &b, double weight) {
        if (a.type == B2VP_PHASE_BLEND || b.type == B2VP_PHASE_BLEND)
            throw Error(B2VP_ERR_UNSUPPORTED, "blendphase: nested blends are not supported by the B200 volpath");
        Phase p; p.type = B2VP_PHASE_BLEND; p.children = { a, b }; p.weight_value = weight; return p;
    }
};
/* src/textures/constant3d.cpp:13-55 */
struct ConstVolume { double value = 1.0; };
/* src/textures/grid3d.cpp:43-160: in-memory grid data[((z*ny+y)*nx+x)*channels+c], float32 as on disk */
struct GridVolume {
    std::vector<float> data;
    int nx = 0, ny = 0, nz = 0, channels = 1;
    std::string filter_type = "trilinear", wrap_mode = "clamp";
    double max_value = -1;   /* < 0: the maximum of the data (grid3d.cpp:157-160) */
    /* use_grid_bbox (grid3d.cpp:152-155): the file's bounding box (min xyz, max xyz; read_vol returns it) is mapped onto
       the unit cube before to_world applies */
    bool use_grid_bbox = false; std::array<float, 6> bbox{ 0, 0, 0, 1, 1, 1 };
};
/* A version-3 float32 `.vol` file (src/textures/volume_data.h:42-102) as a GridVolume.  `bbox` (if given) receives the
   file's bounding box: min xyz, max xyz (what `use_grid_bbox` applies, grid3d.cpp:152-155). */
inline GridVolume read_vol(const std::string &path, std::array<float, 6> *bbox = nullptr);
/* src/media/homogeneous.cpp:21-54 / heterogeneous.cpp:21-54, src/librender/medium.cpp:12-30 */
struct Homogeneous { double sigma_t = 1.0, albedo = 0.75, scale = 1.0; bool has_spectral_extinction = true, sample_emitters = true; };
struct Heterogeneous { double scale = 1.0; bool has_spectral_extinction = true, sample_emitters = true; };
/* src/bsdfs/diffuse.cpp:70-76, null.cpp:30-39, rpv.cpp:66-83 (rho_c defaults to rho_0) */
struct Diffuse { double reflectance = 0.5; };
struct Null {};
struct Rpv { double rho_0 = 0.1, k = 0.1, g = 0.0, rho_c = -1.0; };
/* src/bsdfs/bilambertian.cpp:50-61 */
struct Bilambertian { double reflectance = 0.5, transmittance = 0.5; };
/* src/emitters/directional.cpp:45-73 */
struct Directional {
    std::array<double, 3> direction{ 0, 0, 1 }; double irradiance = 1.0;
    /* instead of `direction` (directional.cpp:50-60): the emitter's to_world, row-major; the light travels along its z axis */
    bool has_to_world = false; std::array<double, 16> to_world{ 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1 };
};
/* src/emitters/constant.cpp:35-40: the environment emitter */
struct Constant { double radiance = 1.0; };
/* src/films/hdrfilm.cpp:100-183, src/samplers/independent.cpp */
/* src/rfilters/{box,tent,gaussian,mitchell,catmullrom,lanczos}.cpp: the film's reconstruction filter */
struct RFilter {
    enum Type { Box, Tent, Gaussian, Mitchell, CatmullRom, Lanczos } type = Box;
    double radius = 0.5;                       /* box.cpp:28 (RayEpsilon is added to it) */
    double stddev = 0.5;                       /* gaussian.cpp:31 */
    double B = 1.f / 3.f, C = 1.f / 3.f;       /* mitchell.cpp:34-36: float literals in every variant */
    int lobes = 3;                             /* lanczos.cpp:34 */
};
struct HdrFilm {
    int width = 768, height = 576;
    RFilter rfilter{};
    int crop_offset_x = 0, crop_offset_y = 0, crop_width = -1, crop_height = -1;   /* film.cpp:30-63; -1 = the whole film */
};
struct IndependentSampler { uint32_t sample_count = 4; uint64_t seed = 0; };
/* src/sensors/distant.cpp:225-290 */
struct Distant {
    std::array<double, 3> direction{ 0, 0, 1 };
    bool has_orientation = false; std::array<double, 3> orientation{ 1, 0, 0 };
    bool has_ray_target = false; std::array<double, 3> ray_target{ 0, 0, 0 };
    bool flip_directions = false;
    int medium = -1;
    /* a rectangle as ray_target / ray_origin (distant.cpp:270-290): the index add_rectangle returned -- for a shape of
       the sensor's own, one added with sample_only = true.  -1: the point target above / the bounding sphere */
    int ray_target_shape = -1, ray_origin_shape = -1;
    /* instead of `direction` / `orientation` (distant.cpp:246-262): the sensor's to_world, row-major */
    bool has_to_world = false; std::array<double, 16> to_world{ 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1 };
};
/* src/sensors/distantflux.cpp:140-182: a film over the hemisphere around to_world * (0, 0, 1); target point, target
   rectangle (an index add_rectangle returned) or none; rays start on the bounding sphere or on an origin rectangle */
struct DistantFlux {
    std::array<double, 16> to_world{ 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1 };   /* row-major */
    bool has_target = false; std::array<double, 3> target{ 0, 0, 0 };
    int target_shape = -1, origin_shape = -1;
    int medium = -1;
};
/* src/sensors/irradiancemeter.cpp:54-80: attached to a rectangle of the scene (`shape` = the index add_rectangle returned) */
struct IrradianceMeter { int shape = -1; int medium = -1; };
/* src/sensors/radiancemeter.cpp:61-104 */
struct RadianceMeter { std::array<double, 3> origin{ 0, 0, 0 }, direction{ 0, 0, 1 }; int medium = -1; };
/* src/shapes/sphere.cpp:95-160: `center`, `radius`, `flip_normals` (+ a to_world given to add_sphere) */
struct Sphere { std::array<double, 3> center{ 0, 0, 0 }; double radius = 1.0; bool flip_normals = false; };
/* src/sensors/mdistant.cpp:146-207: one direction per film column (film size [n, 1]), optional target point */
struct MDistant {
    std::vector<std::array<double, 3>> directions;
    bool has_target = false; std::array<double, 3> target{ 0, 0, 0 };
    int medium = -1;
    int target_shape = -1;   /* a rectangle as target (mdistant.cpp:193-200), see Distant::ray_target_shape */
};
/* src/sensors/mradiancemeter.cpp:75-131: one (origin, direction) pair per film column */
struct MRadianceMeter { std::vector<std::array<double, 3>> origins, directions; int medium = -1; };
/* A triangle mesh as the reference's loaders leave it in memory (src/librender/mesh.cpp): object-space positions,
   optional vertex normals and texture coordinates */
struct MeshData { std::vector<float> positions, normals, texcoords; std::vector<uint32_t> faces; };
/* Wavefront OBJ as src/shapes/obj.cpp:140-325 reads it: v / vn / vt / f, one output vertex per distinct (v, vt, vn)
   index triple in order of first use, polygons as triangle fans, v flipped to 1 - v unless flip_tex_coords = false */
inline MeshData load_obj(const std::string &path, bool flip_tex_coords = true);
/* Stanford PLY as src/shapes/ply.cpp:102-380 reads it: ascii or binary of either byte order; element `vertex` with
   x y z, optional nx ny nz and texture coordinates (u v / texture_u texture_v / s t); element `face` with one
   vertex_index(es) list of exactly three indices per face; other elements and scalar properties are skipped */
inline MeshData load_ply(const std::string &path);
#ifdef B2VP_WITH_ZLIB
/* One mesh of a `.serialized` container (src/shapes/serialized.cpp:174-340): [u16 0x041C][u16 version 3 | 4], a zlib
   stream of flags, (v4) name, vertex and face counts, positions, [normals], [texcoords], [colours], u32 indices; reals
   float32 or (flag 0x2000) float64; for shape_index > 0 the offset dictionary at the end of the file.  Needs zlib:
   compile with -DB2VP_WITH_ZLIB and link -lz */
inline MeshData load_serialized(const std::string &path, int shape_index = 0);
#endif
/* src/integrators/volpath.cpp:19-25, src/librender/integrator.cpp:23-39,302-315 */
struct Volpath {
    int max_depth = -1, rr_depth = 5; bool hide_emitters = false; uint32_t block_size = 0, samples_per_pass = 0xffffffffu;
    /* true: nested in a `moment` integrator (src/integrators/moment.cpp): the film gets 6 more channels per pixel,
       the unweighted X Y Z and their squares */
    bool moment = false;
    /* 1: `volpathmis` (src/integrators/volpathmis.cpp) instead of `volpath`; 2: with use_spectral_mis=false */
    int mis = 0;
};

/* A finished scene: owns the descriptor and every array it points to. */
struct Scene {
    B2vpSceneDesc desc{};
    std::vector<B2vpShape> shapes; std::vector<B2vpBsdf> bsdfs; std::vector<B2vpMedium> media;
    std::vector<B2vpVolume> volumes; std::vector<B2vpPhase> phases; std::vector<B2vpEmitter> emitters;
    std::vector<float> vpos, vnrm, vuv; std::vector<uint32_t> faces;
    std::vector<std::vector<float>> grids;
    std::vector<std::vector<double>> phase_values;   /* tabphase tables, parallel to `phases` */
    std::vector<double> sensor_transforms;           /* mdistant / mradiancemeter: 16 doubles per film column */
    int width() const { return desc.sensor.crop_size[0]; }
    int height() const { return desc.sensor.crop_size[1]; }
};

template <typename F> class SceneBuilder {
public:
    SceneBuilder() : sc_(new Scene()) {
        for (int k = 0; k < 3; ++k) { bmin_[k] = std::numeric_limits<F>::infinity(); bmax_[k] = -bmin_[k]; }
    }
    /* A float property as a plugin reads it: Properties keeps floats, 3-vectors and transforms in single precision
       in every variant (include/mitsuba/core/properties.h:55-56); the plugin widens the value to Float. */
    static F prop(double v) { return (F) (float) v; }
    static Vec3<F> prop3(const std::array<double, 3> &v) { return { prop(v[0]), prop(v[1]), prop(v[2]) }; }
    static Mat4<F> prop_xf(const Mat4<F> &t) { Mat4<F> r; for (int i = 0; i < 16; ++i) r.m[i] = (F) (float) t.m[i]; return r; }
    /* normalize(props.vector3f(..)): the normalisation itself runs in single precision (distant.cpp:253, directional.cpp:55) */
    static Vec3<F> prop_dir(const std::array<double, 3> &v) {
        Vec3<float> d = normalize(Vec3<float>{ (float) v[0], (float) v[1], (float) v[2] });
        return { (F) d[0], (F) d[1], (F) d[2] };
    }
    int add_bsdf(const Diffuse &b) { return push_bsdf(B2VP_BSDF_DIFFUSE, prop(b.reflectance), 0, 0, 0); }
    int add_bsdf(const Null &) { return push_bsdf(B2VP_BSDF_NULL, 0, 0, 0, 0); }
    int add_bsdf(const Bilambertian &b) { return push_bsdf(B2VP_BSDF_BILAMBERTIAN, prop(b.reflectance), prop(b.transmittance), 0, 0); }
    int add_bsdf(const Rpv &b) { return push_bsdf(B2VP_BSDF_RPV, prop(b.rho_0), prop(b.k), prop(b.g), prop(b.rho_c < 0 ? b.rho_0 : b.rho_c)); }

    int add_medium(const Homogeneous &m, const Phase &ph) {
        B2vpMedium o{};
        o.type = B2VP_MEDIUM_HOMOGENEOUS;
        o.albedo = add_volume(ConstVolume{ m.albedo }); o.sigma_t = add_volume(ConstVolume{ m.sigma_t });
        return push_medium(o, prop(m.scale), m.has_spectral_extinction, m.sample_emitters, ph);
    }
    /* sigma_t must be a grid (ConstVolume::max is not implemented in the reference, heterogeneous.cpp:29) */
    int add_medium(const Heterogeneous &m, const GridVolume &sigma_t, const Mat4<F> &sigma_t_to_world, double albedo,
                   const Phase &ph) {
        B2vpMedium o{};
        o.type = B2VP_MEDIUM_HETEROGENEOUS;
        o.albedo = add_volume(ConstVolume{ albedo }); o.sigma_t = add_volume(sigma_t, sigma_t_to_world);
        return push_medium(o, prop(m.scale), m.has_spectral_extinction, m.sample_emitters, ph);
    }
    /* the same with an albedo grid (C3: sigma_t and albedo fields of one cloud scene) */
    int add_medium(const Heterogeneous &m, const GridVolume &sigma_t, const Mat4<F> &sigma_t_to_world, const GridVolume &albedo,
                   const Mat4<F> &albedo_to_world, const Phase &ph) {
        B2vpMedium o{};
        o.type = B2VP_MEDIUM_HETEROGENEOUS;
        o.albedo = add_volume(albedo, albedo_to_world); o.sigma_t = add_volume(sigma_t, sigma_t_to_world);
        return push_medium(o, prop(m.scale), m.has_spectral_extinction, m.sample_emitters, ph);
    }
    int add_volume(const ConstVolume &v) {
        B2vpVolume o{};
        o.type = B2VP_VOLUME_CONST; o.channels = 1; o.value = (double) prop(v.value); o.max_value = o.value;
        finish_volume(o, Mat4<F>());
        return (int) sc_->volumes.size() - 1;
    }
    int add_volume(const GridVolume &v, const Mat4<F> &to_world) {
        if ((size_t) v.nx * v.ny * v.nz * v.channels != v.data.size() || v.data.empty())
            throw Error(B2VP_ERR_INVALID, "gridvolume: data size does not match the resolution");
        if (v.channels != 1 && v.channels != 3) throw Error(B2VP_ERR_UNSUPPORTED, "Unsupported channel count (expected 1 or 3)");
        B2vpVolume o{};
        o.type = B2VP_VOLUME_GRID; o.channels = v.channels; o.res[0] = v.nx; o.res[1] = v.ny; o.res[2] = v.nz;
        if (v.filter_type == "nearest") o.filter = B2VP_FILTER_NEAREST;
        else if (v.filter_type == "trilinear") o.filter = B2VP_FILTER_TRILINEAR;
        else throw Error(B2VP_ERR_INVALID, "Invalid filter type \"" + v.filter_type + "\"");
        if (v.wrap_mode == "repeat") o.wrap = B2VP_WRAP_REPEAT;
        else if (v.wrap_mode == "mirror") o.wrap = B2VP_WRAP_MIRROR;
        else if (v.wrap_mode == "clamp") o.wrap = B2VP_WRAP_CLAMP;
        else throw Error(B2VP_ERR_INVALID, "Invalid wrap mode \"" + v.wrap_mode + "\"");
        sc_->grids.push_back(v.data);
        o.data = sc_->grids.back().data();
        o.max_value = v.max_value >= 0 ? (double) prop(v.max_value) : (double) *std::max_element(v.data.begin(), v.data.end());
        finish_volume(o, prop_xf(to_world), v.use_grid_bbox ? &v.bbox : nullptr);
        return (int) sc_->volumes.size() - 1;
    }

    /* src/shapes/rectangle.cpp:75-104 */
    /* sample_only: a rectangle of a sensor's own (its ray_target / ray_origin), not a child of the scene: never
       intersected by paths and outside the scene's bounding box (B2vpShape::sample_only) */
    int add_rectangle(const Mat4<F> &to_world_, int bsdf, int interior = -1, int exterior = -1, bool sample_only = false,
                      bool flip_normals = false) {
        /* flip_normals: to_world * scale(1, 1, -1) (rectangle.cpp:80-83) */
        const Mat4<F> to_world = flip_normals ? prop_xf(to_world_) * Mat4<F>::scale(1, 1, -1) : prop_xf(to_world_);
        B2vpShape s{};
        s.type = B2VP_SHAPE_RECTANGLE; s.bsdf = bsdf; s.interior_medium = interior; s.exterior_medium = exterior;
        Vec3<F> dp_du = to_world.vector({ 2, 0, 0 }), dp_dv = to_world.vector({ 0, 2, 0 });
        Vec3<F> n = normalize(to_world.normal({ 0, 0, 1 }));
        Vec3<F> cr = cross(dp_du, dp_dv);
        F area = std::sqrt(dot(cr, cr));
        for (int k = 0; k < 3; ++k) { s.frame_s[k] = dp_du[k]; s.frame_t[k] = dp_dv[k]; s.frame_n[k] = n[k]; }
        s.inv_surface_area = (double) ((F) 1 / area);
        const F corners[4][3] = { { -1, -1, 0 }, { 1, -1, 0 }, { 1, 1, 0 }, { -1, 1, 0 } };
        if (!sample_only)
            for (auto &c : corners) expand(to_world.point({ c[0], c[1], c[2] }));
        s.sample_only = sample_only ? 1 : 0;
        s.to_world = to_world.pod(); s.to_object = to_world.inverse_f32().pod();
        sc_->shapes.push_back(s);
        return (int) sc_->shapes.size() - 1;
    }
    /* in-memory triangle mesh in world space (what the obj / ply / serialized loaders produce,
       src/librender/mesh.cpp): positions 3 floats per vertex, faces 3 indices per triangle */
    int add_mesh(const std::vector<float> &positions, const std::vector<uint32_t> &faces, int bsdf, int interior = -1,
                 int exterior = -1, const std::vector<float> *normals = nullptr, const std::vector<float> *texcoords = nullptr) {
        if (positions.size() % 3 || faces.size() % 3) throw Error(B2VP_ERR_INVALID, "mesh: positions / faces must come in triples");
        B2vpShape s{};
        s.type = B2VP_SHAPE_MESH; s.bsdf = bsdf; s.interior_medium = interior; s.exterior_medium = exterior;
        const uint32_t nv = (uint32_t) (positions.size() / 3), nf = (uint32_t) (faces.size() / 3);
        for (uint32_t i : faces) if (i >= nv) throw Error(B2VP_ERR_INVALID, "mesh: face index out of range");
        s.vertex_offset = (uint32_t) (sc_->vpos.size() / 3); s.vertex_count = nv;
        s.face_offset = (uint32_t) (sc_->faces.size() / 3); s.face_count = nf;
        s.has_vertex_normals = normals != nullptr; s.has_vertex_texcoords = texcoords != nullptr;
        sc_->vpos.insert(sc_->vpos.end(), positions.begin(), positions.end());
        if (normals) sc_->vnrm.insert(sc_->vnrm.end(), normals->begin(), normals->end());
        else sc_->vnrm.insert(sc_->vnrm.end(), positions.size(), 0.f);
        if (texcoords) sc_->vuv.insert(sc_->vuv.end(), texcoords->begin(), texcoords->end());
        else sc_->vuv.insert(sc_->vuv.end(), (size_t) nv * 2, 0.f);
        sc_->faces.insert(sc_->faces.end(), faces.begin(), faces.end());
        for (uint32_t v = 0; v < nv; ++v) expand({ (F) positions[3 * v], (F) positions[3 * v + 1], (F) positions[3 * v + 2] });
        Mat4<F> id;
        s.to_world = id.pod(); s.to_object = id.pod();
        sc_->shapes.push_back(s);
        return (int) sc_->shapes.size() - 1;
    }
    /* A loaded mesh (load_obj, or the caller's own loader) under `to_world`: positions by the fmadd chain of
       Transform::transform_affine in Float, normals by the inverse transpose and renormalised, both rounded to
       float32 like the reference's mesh buffers (obj.cpp:172-190) */
    int add_mesh(const MeshData &m, const Mat4<F> &to_world_, int bsdf, int interior = -1, int exterior = -1) {
        const Mat4<F> tw = prop_xf(to_world_);
        std::vector<float> p(m.positions.size()), n(m.normals.size());
        for (size_t v = 0; v + 2 < m.positions.size(); v += 3) {
            Vec3<F> q = tw.point({ (F) m.positions[v], (F) m.positions[v + 1], (F) m.positions[v + 2] });
            for (int k = 0; k < 3; ++k) p[v + k] = (float) q[k];
        }
        for (size_t v = 0; v + 2 < m.normals.size(); v += 3) {
            Vec3<F> q = normalize(tw.normal({ (F) m.normals[v], (F) m.normals[v + 1], (F) m.normals[v + 2] }));
            for (int k = 0; k < 3; ++k) n[v + k] = (float) q[k];
        }
        return add_mesh(p, m.faces, bsdf, interior, exterior, m.normals.empty() ? nullptr : &n,
                        m.texcoords.empty() ? nullptr : &m.texcoords);
    }
    /* src/shapes/sphere.cpp:95-160.  to_world * translate(center) * scale(radius) must be a similarity (rotation,
       uniform scale, translation): the reference takes the radius from a polar decomposition and warns otherwise;
       here anything else is refused */
    int add_sphere(const Sphere &sp, const Mat4<F> &to_world_, int bsdf, int interior = -1, int exterior = -1) {
        const Vec3<F> c0 = prop3(sp.center);
        const F r0 = prop(sp.radius);
        Mat4<F> m = prop_xf(to_world_) * Mat4<F>::translate(c0[0], c0[1], c0[2]) * Mat4<F>::scale(r0, r0, r0);
        Vec3<F> col[3];
        for (int c = 0; c < 3; ++c) col[c] = { m(0, c), m(1, c), m(2, c) };
        const bool diagonal = m(0, 1) == 0 && m(0, 2) == 0 && m(1, 0) == 0 && m(1, 2) == 0 && m(2, 0) == 0 && m(2, 1) == 0;
        F radius = diagonal ? m(0, 0) : std::sqrt(dot(col[0], col[0]));
        const F tol = (F) 1e-4 * std::fabs(radius);
        for (int c = 0; c < 3; ++c) {
            if (std::fabs(std::sqrt(dot(col[c], col[c])) - std::fabs(radius)) > tol ||
                std::fabs(dot(col[c], col[(c + 1) % 3])) > tol * std::fabs(radius))
                throw Error(B2VP_ERR_UNSUPPORTED, "sphere: 'to_world' transform shouldn't contain any scale along individual axes");
        }
        const Vec3<F> cr = cross(col[0], col[1]);
        if (!diagonal && dot(cr, col[2]) < 0) radius = -radius;   /* a reflection: det(Q) > 0 puts the sign into the scale */
        bool flip = sp.flip_normals;
        if (radius <= 0) { radius = std::fabs(radius); flip = !flip; }
        if (radius == 0) throw Error(B2VP_ERR_INVALID, "sphere: radius must be non-zero");
        B2vpShape s{};
        s.type = B2VP_SHAPE_SPHERE; s.bsdf = bsdf; s.interior_medium = interior; s.exterior_medium = exterior;
        Mat4<F> tw;   /* rotation * radius + centre */
        const F sign = (diagonal ? m(0, 0) : (dot(cr, col[2]) < 0 ? (F) -1 : (F) 1)) < 0 ? (F) -1 : (F) 1;
        for (int c = 0; c < 3; ++c)
            for (int r = 0; r < 3; ++r) tw(r, c) = diagonal ? (r == c ? radius : (F) 0) : sign * col[c][r];
        const Vec3<F> center = { m(0, 3), m(1, 3), m(2, 3) };
        for (int k = 0; k < 3; ++k) { tw(k, 3) = center[k]; s.center[k] = (double) center[k]; }
        s.radius = (double) radius; s.flip_normals = flip ? 1 : 0;
        const F pi = (F) 3.14159265358979323846;
        s.inv_surface_area = (double) ((F) 1 / ((F) 4 * pi * radius * radius));
        s.to_world = tw.pod(); s.to_object = tw.inverse().pod();   /* recomposed and inverted in the variant's type (sphere.cpp:126-127) */
        expand({ center[0] - radius, center[1] - radius, center[2] - radius });
        expand({ center[0] + radius, center[1] + radius, center[2] + radius });
        sc_->shapes.push_back(s);
        return (int) sc_->shapes.size() - 1;
    }
    /* src/shapes/cube.cpp:54-116: 24 vertices, 12 triangles, face normals and texcoords; transformed in f32 */
    int add_cube(const Mat4<F> &to_world, int bsdf, int interior = -1, int exterior = -1) {
        static const float V[24][3] = { { 1, -1, -1 }, { 1, -1, 1 }, { -1, -1, 1 }, { -1, -1, -1 }, { 1, 1, -1 }, { -1, 1, -1 },
            { -1, 1, 1 }, { 1, 1, 1 }, { 1, -1, -1 }, { 1, 1, -1 }, { 1, 1, 1 }, { 1, -1, 1 }, { 1, -1, 1 }, { 1, 1, 1 },
            { -1, 1, 1 }, { -1, -1, 1 }, { -1, -1, 1 }, { -1, 1, 1 }, { -1, 1, -1 }, { -1, -1, -1 }, { 1, 1, -1 },
            { 1, -1, -1 }, { -1, -1, -1 }, { -1, 1, -1 } };
        static const float N[6][3] = { { 0, -1, 0 }, { 0, 1, 0 }, { 1, 0, 0 }, { 0, 0, 1 }, { -1, 0, 0 }, { 0, 0, -1 } };
        static const float UV[4][2] = { { 0, 1 }, { 1, 1 }, { 1, 0 }, { 0, 0 } };
        static const uint32_t T[12][3] = { { 0, 1, 2 }, { 3, 0, 2 }, { 4, 5, 6 }, { 7, 4, 6 }, { 8, 9, 10 }, { 11, 8, 10 },
            { 12, 13, 14 }, { 15, 12, 14 }, { 16, 17, 18 }, { 19, 16, 18 }, { 20, 21, 22 }, { 23, 20, 22 } };
        /* cube.cpp:104-107: transformed and normalised in Float, stored as float32 */
        const Mat4<F> tw = prop_xf(to_world);
        std::vector<float> p, n, uv;
        std::vector<uint32_t> f;
        for (int v = 0; v < 24; ++v) {
            Vec3<F> q = tw.point({ (F) V[v][0], (F) V[v][1], (F) V[v][2] });
            Vec3<F> m = normalize(tw.normal({ (F) N[v / 4][0], (F) N[v / 4][1], (F) N[v / 4][2] }));
            for (int k = 0; k < 3; ++k) { p.push_back((float) q[k]); n.push_back((float) m[k]); }
            uv.push_back(UV[v % 4][0]); uv.push_back(UV[v % 4][1]);
        }
        for (auto &t : T) f.insert(f.end(), t, t + 3);
        return add_mesh(p, f, bsdf, interior, exterior, &n, &uv);
    }
    void add_emitter(const Constant &e) {
        B2vpEmitter o{};
        o.type = B2VP_EMITTER_CONSTANT; o.irradiance = (double) prop(e.radiance);
        sc_->emitters.push_back(o);
    }
    void add_emitter(const Directional &e) {
        Vec3<F> d = prop_dir(e.direction);
        Vec3<F> up, t;
        coordinate_system(d, up, t);
        /* Endpoint::m_world_transform is an AnimatedTransform: single precision in every variant (transform.h:366) */
        Mat4<F> tw = prop_xf(Mat4<F>::look_at({ 0, 0, 0 }, d, up));
        if (e.has_to_world) {
            for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) tw(r, c) = (F) e.to_world[(size_t) (r * 4 + c)];
            tw = prop_xf(tw);
        }
        Vec3<F> z = tw.vector({ 0, 0, 1 });
        B2vpEmitter o{};
        o.type = B2VP_EMITTER_DIRECTIONAL; o.irradiance = (double) prop(e.irradiance);
        for (int k = 0; k < 3; ++k) o.direction[k] = z[k];
        sc_->emitters.push_back(o);
    }
    void set_sensor(const Distant &sn, const HdrFilm &film, const IndependentSampler &sampler) {
        B2vpSensor o = base_sensor(film, sampler, sn.medium);
        o.type = B2VP_SENSOR_DISTANT; o.flip_directions = sn.flip_directions; o.needs_aperture_sample = 1;
        o.direction_type = (film.width == 1 && film.height == 1) ? B2VP_DIR_SINGLE
                           : film.height == 1 ? B2VP_DIR_SAMPLE_WIDTH : B2VP_DIR_SAMPLE_ALL;
        Vec3<F> d = prop_dir(sn.direction);
        Vec3<F> up, tmp;
        if (sn.has_orientation)
            up = normalize(cross(d, prop3(sn.orientation)));
        else
            coordinate_system(d, tmp, up);
        o.to_world = prop_xf(Mat4<F>::look_at({ 0, 0, 0 }, d, up)).pod();   /* AnimatedTransform: single precision */
        if (sn.has_to_world) {
            Mat4<F> tw;
            for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) tw(r, c) = (F) sn.to_world[(size_t) (r * 4 + c)];
            o.to_world = prop_xf(tw).pod();
        }
        o.target_type = sn.has_ray_target ? B2VP_TARGET_POINT : B2VP_TARGET_NONE;
        for (int k = 0; k < 3; ++k) o.target_point[k] = (double) prop(sn.ray_target[k]);
        if (sn.ray_target_shape >= 0) { o.target_type = B2VP_TARGET_SHAPE; o.target_shape = rectangle_index(sn.ray_target_shape); }
        o.origin_type = B2VP_ORIGIN_BSPHERE;
        if (sn.ray_origin_shape >= 0) { o.origin_type = B2VP_ORIGIN_SHAPE; o.origin_shape = rectangle_index(sn.ray_origin_shape); }
        sc_->desc.sensor = o;
        have_sensor_ = true;
    }
    void set_sensor(const DistantFlux &sn, const HdrFilm &film, const IndependentSampler &sampler) {
        B2vpSensor o = base_sensor(film, sampler, sn.medium);
        o.type = B2VP_SENSOR_DISTANTFLUX; o.needs_aperture_sample = 1;
        Mat4<F> tw;
        for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) tw(r, c) = (F) sn.to_world[(size_t) (r * 4 + c)];
        tw = prop_xf(tw);                           /* AnimatedTransform / props.transform(): single precision */
        o.to_world = tw.pod();
        const Vec3<F> n = tw.vector({ 0, 0, 1 });   /* m_reference_normal (distantflux.cpp:173-175) */
        for (int k = 0; k < 3; ++k) o.reference_normal[k] = (double) n[k];
        o.target_type = sn.has_target ? B2VP_TARGET_POINT : B2VP_TARGET_NONE;
        for (int k = 0; k < 3; ++k) o.target_point[k] = (double) prop(sn.target[k]);
        if (sn.target_shape >= 0) { o.target_type = B2VP_TARGET_SHAPE; o.target_shape = rectangle_index(sn.target_shape); }
        o.origin_type = B2VP_ORIGIN_BSPHERE;
        if (sn.origin_shape >= 0) { o.origin_type = B2VP_ORIGIN_SHAPE; o.origin_shape = rectangle_index(sn.origin_shape); }
        sc_->desc.sensor = o;
        have_sensor_ = true;
    }
    void set_sensor(const IrradianceMeter &sn, const IndependentSampler &sampler, const HdrFilm *film = nullptr) {
        B2vpSensor o = base_sensor(fixed_film(film, 1), sampler, sn.medium);
        o.type = B2VP_SENSOR_IRRADIANCEMETER; o.needs_aperture_sample = 1;
        if (sn.shape < 0) throw Error(B2VP_ERR_INVALID, "irradiancemeter: no parent shape");
        o.origin_shape = rectangle_index(sn.shape);
        sc_->desc.sensor = o;
        have_sensor_ = true;
    }
    void set_sensor(const RadianceMeter &sn, const IndependentSampler &sampler, const HdrFilm *film = nullptr) {
        B2vpSensor o = base_sensor(fixed_film(film, 1), sampler, sn.medium);
        o.type = B2VP_SENSOR_RADIANCEMETER; o.needs_aperture_sample = 0;
        Vec3<F> org = prop3(sn.origin), d = prop3(sn.direction);
        Vec3<F> up, tmp;
        coordinate_system(d, up, tmp);
        o.to_world = prop_xf(Mat4<F>::look_at(org, { org[0] + d[0], org[1] + d[1], org[2] + d[2] }, up)).pod();
        sc_->desc.sensor = o;
        have_sensor_ = true;
    }
    /* src/sensors/mdistant.cpp:146-207: a look-at matrix per direction, ScalarFloat matrices of the plugin's own */
    void set_sensor(const MDistant &sn, const IndependentSampler &sampler, const HdrFilm *film = nullptr) {
        const int n = (int) sn.directions.size();
        if (n == 0) throw Error(B2VP_ERR_INVALID, "mdistant: no directions");
        B2vpSensor o = base_sensor(fixed_film(film, n), sampler, sn.medium);
        o.type = B2VP_SENSOR_MDISTANT; o.needs_aperture_sample = 1;
        o.target_type = sn.has_target ? B2VP_TARGET_POINT : B2VP_TARGET_NONE;
        for (int k = 0; k < 3; ++k) o.target_point[k] = (double) prop(sn.target[k]);
        if (sn.target_shape >= 0) { o.target_type = B2VP_TARGET_SHAPE; o.target_shape = rectangle_index(sn.target_shape); }
        o.origin_type = B2VP_ORIGIN_BSPHERE;
        sc_->sensor_transforms.clear();
        for (const auto &dir : sn.directions) {
            const Vec3<F> d = prop3(dir);   /* std::stof values widened; not normalised by the plugin */
            Vec3<F> tmp, up;
            coordinate_system(d, tmp, up);
            const B2vpMat4 t = Mat4<F>::look_at({ 0, 0, 0 }, d, up).pod();
            sc_->sensor_transforms.insert(sc_->sensor_transforms.end(), t.m, t.m + 16);
        }
        o.n_transforms = n;
        sc_->desc.sensor = o;
        have_sensor_ = true;
    }
    /* src/sensors/mradiancemeter.cpp:75-131 */
    void set_sensor(const MRadianceMeter &sn, const IndependentSampler &sampler, const HdrFilm *film = nullptr) {
        const int n = (int) sn.directions.size();
        if (n == 0 || sn.origins.size() != sn.directions.size())
            throw Error(B2VP_ERR_INVALID, "Invalid specification! Number of parameters for origins and directions are not equal.");
        B2vpSensor o = base_sensor(fixed_film(film, n), sampler, sn.medium);
        o.type = B2VP_SENSOR_MRADIANCEMETER; o.needs_aperture_sample = 0;
        sc_->sensor_transforms.clear();
        for (int i = 0; i < n; ++i) {
            const Vec3<F> org = prop3(sn.origins[i]), d = prop3(sn.directions[i]);
            Vec3<F> up, tmp;
            coordinate_system(d, up, tmp);
            const B2vpMat4 t = Mat4<F>::look_at(org, { org[0] + d[0], org[1] + d[1], org[2] + d[2] }, up).pod();
            sc_->sensor_transforms.insert(sc_->sensor_transforms.end(), t.m, t.m + 16);
        }
        o.n_transforms = n;
        sc_->desc.sensor = o;
        have_sensor_ = true;
    }
    void set_integrator(const Volpath &v) {
        if (v.max_depth < -1) throw Error(B2VP_ERR_INVALID, "\"max_depth\" must be set to -1 (infinite) or a value >= 0");
        if (v.rr_depth <= 0) throw Error(B2VP_ERR_INVALID, "\"rr_depth\" must be set to a value greater than zero!");
        B2vpIntegrator &it = sc_->desc.integrator;
        it.max_depth = v.max_depth; it.rr_depth = v.rr_depth; it.hide_emitters = v.hide_emitters;
        it.block_size = v.block_size; it.samples_per_pass = v.samples_per_pass; it.arg_order_ltr = 0;
        it.moment = v.moment ? 1u : 0u;
        it.mis = (uint32_t) v.mis;
        have_integrator_ = true;
    }

    /* Scene::bbox() and its bounding sphere with the padding of distant.cpp:292-297 / directional.cpp:68-73 */
    std::unique_ptr<Scene> finish() {
        if (!have_sensor_) throw Error(B2VP_ERR_INVALID, "scene has no sensor");
        if (!have_integrator_) set_integrator(Volpath{});
        Scene &s = *sc_;
        B2vpSceneDesc &d = s.desc;
        d.abi_version = B2VP_ABI_VERSION; d.struct_size = (uint32_t) sizeof(B2vpSceneDesc);
        d.precision = std::is_same<F, float>::value ? B2VP_PREC_F32 : B2VP_PREC_F64;
        d.n_shapes = (uint32_t) s.shapes.size(); d.shapes = s.shapes.data();
        d.n_bsdfs = (uint32_t) s.bsdfs.size(); d.bsdfs = s.bsdfs.data();
        d.n_media = (uint32_t) s.media.size(); d.media = s.media.data();
        d.n_volumes = (uint32_t) s.volumes.size(); d.volumes = s.volumes.data();
        for (size_t i = 0; i < s.phases.size(); ++i)
            if (s.phases[i].type == B2VP_PHASE_TABULATED) {
                s.phases[i].n_values = (uint32_t) s.phase_values[i].size(); s.phases[i].values = s.phase_values[i].data();
            }
        d.n_phases = (uint32_t) s.phases.size(); d.phases = s.phases.data();
        d.n_emitters = (uint32_t) s.emitters.size(); d.emitters = s.emitters.data();
        if (d.sensor.n_transforms) d.sensor.transforms = s.sensor_transforms.data();
        d.n_vertices = (uint32_t) (s.vpos.size() / 3); d.n_faces = (uint32_t) (s.faces.size() / 3);
        if (d.n_vertices) {
            d.vertex_positions = s.vpos.data(); d.vertex_normals = s.vnrm.data();
            d.vertex_texcoords = s.vuv.data(); d.faces = s.faces.data();
        }
        /* grid pointers: the vectors may have moved while volumes were added */
        size_t gi = 0;
        for (auto &v : s.volumes) if (v.type == B2VP_VOLUME_GRID) v.data = s.grids[gi++].data();
        if (!(bmin_[0] <= bmax_[0])) for (int k = 0; k < 3; ++k) bmin_[k] = bmax_[k] = 0;
        Vec3<F> c, df;
        for (int k = 0; k < 3; ++k) { c[k] = (bmax_[k] + bmin_[k]) * (F) 0.5; df[k] = c[k] - bmax_[k]; }
        F radius = std::sqrt(df[0] * df[0] + df[1] * df[1] + df[2] * df[2]);
        const F ray_eps = std::numeric_limits<F>::epsilon() / (F) 2 * (F) 1500;
        radius = std::max(ray_eps, radius * ((F) 1 + ray_eps));
        for (int k = 0; k < 3; ++k) { d.bbox_min[k] = bmin_[k]; d.bbox_max[k] = bmax_[k]; d.bsphere_center[k] = c[k]; }
        d.bsphere_radius = radius;
        return std::move(sc_);
    }

private:
    int push_bsdf(int type, F p0, F p1, F p2, F p3) {
        B2vpBsdf b{};
        b.type = type; b.p[0] = p0; b.p[1] = p1; b.p[2] = p2; b.p[3] = p3;
        sc_->bsdfs.push_back(b);
        return (int) sc_->bsdfs.size() - 1;
    }
    int push_phase(const Phase &ph) {
        B2vpPhase p{};
        p.type = ph.type; p.g = (double) prop(ph.g);
        if (ph.type == B2VP_PHASE_BLEND) {
            if (ph.children.size() != 2) throw Error(B2VP_ERR_INVALID, "BlendPhase: Two child phase functions must be specified!");
            p.child[0] = push_phase(ph.children[0]); p.child[1] = push_phase(ph.children[1]);
            if (ph.weight_grid.empty()) p.weight_volume = add_volume(ConstVolume{ ph.weight_value });
            else {
                GridVolume g;
                g.data = ph.weight_grid; g.nx = ph.weight_res[0]; g.ny = ph.weight_res[1]; g.nz = ph.weight_res[2];
                g.filter_type = ph.weight_filter; g.wrap_mode = ph.weight_wrap;
                Mat4<F> tw;
                for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) tw(r, c) = (F) ph.weight_to_world[(size_t) (r * 4 + c)];
                p.weight_volume = add_volume(g, tw);
            }
        }
        std::vector<double> vals;
        for (double x : ph.values) vals.push_back((double) (F) x);   /* (ScalarFloat) std::stod(..) */
        sc_->phases.push_back(p);
        sc_->phase_values.push_back(std::move(vals));
        return (int) sc_->phases.size() - 1;
    }
    int push_medium(B2vpMedium o, F scale, bool spectral, bool sample_emitters, const Phase &ph) {
        o.phase = push_phase(ph);
        o.scale = scale; o.has_spectral_extinction = spectral; o.sample_emitters = sample_emitters;
        sc_->media.push_back(o);
        return (int) sc_->media.size() - 1;
    }
    /* Volume base class: world_to_local = to_world^-1, bbox = image of the unit cube (texture.h:265-272) */
    void finish_volume(B2vpVolume &o, const Mat4<F> &to_world, const std::array<float, 6> *grid_bbox = nullptr) {
        Mat4<F> w2l = to_world.inverse_f32();
        Mat4<F> l2w = to_world;
        if (grid_bbox) {   /* grid3d.cpp:152-155: world_to_local = scale(1 / extents) * translate(-min) * world_to_local */
            const auto &bb = *grid_bbox;
            const F mn[3] = { (F) bb[0], (F) bb[1], (F) bb[2] }, ex[3] = { (F) bb[3] - (F) bb[0], (F) bb[4] - (F) bb[1], (F) bb[5] - (F) bb[2] };
            w2l = Mat4<F>::scale((F) 1 / ex[0], (F) 1 / ex[1], (F) 1 / ex[2]) * Mat4<F>::translate(-mn[0], -mn[1], -mn[2]) * w2l;
            l2w = to_world * Mat4<F>::translate(mn[0], mn[1], mn[2]) * Mat4<F>::scale(ex[0], ex[1], ex[2]);
        }
        o.world_to_local = w2l.pod();
        Vec3<F> a = l2w.point({ 0, 0, 0 }), b = l2w.point({ 1, 1, 1 });
        for (int k = 0; k < 3; ++k) { o.bbox_min[k] = std::min(a[k], b[k]); o.bbox_max[k] = std::max(a[k], b[k]); }
        sc_->volumes.push_back(o);
    }
    /* sensors whose film size is given by the sensor ([n, 1]): the caller's film, if any, brings its reconstruction
       filter and crop window; its size must be the sensor's (mdistant.cpp:178-183, radiancemeter.cpp:94-95) */
    static HdrFilm fixed_film(const HdrFilm *film, int n) {
        if (!film) return HdrFilm{ n, 1 };
        if (film->width != n || film->height != 1)
            throw Error(B2VP_ERR_INVALID, "Film size must be [" + std::to_string(n) + ", 1], got [" + std::to_string(film->width) + ", " +
                                          std::to_string(film->height) + "]");
        return *film;
    }
    int rectangle_index(int shape) const {
        if (shape >= (int) sc_->shapes.size() || sc_->shapes[(size_t) shape].type != B2VP_SHAPE_RECTANGLE)
            throw Error(B2VP_ERR_UNSUPPORTED, "sensor target / origin shape must be a rectangle");
        return shape;
    }
    B2vpSensor base_sensor(const HdrFilm &film, const IndependentSampler &sampler, int medium) {
        B2vpSensor o{};
        o.film_size[0] = film.width; o.film_size[1] = film.height;
        o.crop_offset[0] = film.crop_offset_x; o.crop_offset[1] = film.crop_offset_y;
        o.crop_size[0] = film.crop_width < 0 ? film.width : film.crop_width;
        o.crop_size[1] = film.crop_height < 0 ? film.height : film.crop_height;
        for (int k = 0; k < 2; ++k)   /* Film::set_crop_window (film.cpp:54-63) */
            if (o.crop_offset[k] < 0 || o.crop_size[k] <= 0 || o.crop_offset[k] + o.crop_size[k] > o.film_size[k])
                throw Error(B2VP_ERR_INVALID, "Invalid crop window specification!");
        rfilter_table(film.rfilter, o);
        o.sample_count = sampler.sample_count; o.sampler_seed = sampler.seed;
        o.medium = medium; o.target_shape = o.origin_shape = -1;
        o.to_world = Mat4<F>().pod();
        return o;
    }
    void expand(const Vec3<F> &p) { for (int k = 0; k < 3; ++k) { bmin_[k] = std::min(bmin_[k], p[k]); bmax_[k] = std::max(bmax_[k], p[k]); } }

    /* ReconstructionFilter::init_discretization (src/libcore/rfilter.cpp:9-21) over the eval() of the six filter
       plugins, in the variant's scalar type, plus the footprint width ImageBlock::put derives (imageblock.cpp:122) */
    static void rfilter_table(const RFilter &rf, B2vpSensor &o) {
        const F eps = std::numeric_limits<F>::epsilon() / 2, ray_eps = eps * 1500;   /* math.h:36-37 */
        F radius = 0;
        const F stddev = (F) rf.stddev, alpha = (F) -1 / ((F) 2 * stddev * stddev);
        const F B = rf.type == RFilter::Mitchell ? (F) rf.B : (F) 0, C = rf.type == RFilter::Mitchell ? (F) rf.C : (F) 0.5;
        switch (rf.type) {
            case RFilter::Box: radius = (F) rf.radius + ray_eps; break;
            case RFilter::Tent: radius = 1; break;
            case RFilter::Gaussian: radius = (F) 4 * stddev; break;
            case RFilter::Mitchell: case RFilter::CatmullRom: radius = 2; break;
            case RFilter::Lanczos: radius = (F) rf.lobes; break;
        }
        if (!(radius > 0)) throw Error(B2VP_ERR_INVALID, "reconstruction filter: the radius must be positive");
        const F bias = std::exp(alpha * (radius * radius));
        auto eval = [&](F x) -> F {
            x = std::abs(x);
            switch (rf.type) {
                case RFilter::Box: return x <= radius ? (F) 1 : (F) 0;
                case RFilter::Tent: return std::max((F) 0, (F) 1 - std::abs(x * ((F) 1 / radius)));
                case RFilter::Gaussian: return std::max((F) 0, std::exp(alpha * (x * x)) - bias);
                case RFilter::Mitchell: case RFilter::CatmullRom: {
                    const F x2 = x * x, x3 = x2 * x;
                    const F r = x < 1 ? ((F) 12 - (F) 9 * B - (F) 6 * C) * x3 + ((F) -18 + (F) 12 * B + (F) 6 * C) * x2 + ((F) 6 - (F) 2 * B)
                                      : (-B - (F) 6 * C) * x3 + ((F) 6 * B + (F) 30 * C) * x2 + ((F) -12 * B - (F) 48 * C) * x + ((F) 8 * B + (F) 24 * C);
                    return x < 2 ? (F) (1.f / 6.f) * r : (F) 0;
                }
                default: {
                    if (x < eps) return 1;
                    if (x > radius) return 0;
                    const F x1 = (F) 3.14159265358979323846 * x, x2 = x1 / radius;
                    return (std::sin(x1) * std::sin(x2)) / (x1 * x2);
                }
            }
        };
        for (int i = 0; i < B2VP_RFILTER_RESOLUTION; ++i) o.rfilter_values[i] = (double) eval((radius * (F) i) / (F) B2VP_RFILTER_RESOLUTION);
        o.rfilter_values[B2VP_RFILTER_RESOLUTION] = 0;
        o.rfilter_radius = (double) radius;
        o.rfilter_scale = (double) ((F) B2VP_RFILTER_RESOLUTION / radius);
        o.rfilter_border = (uint32_t) std::max(0, (int) std::ceil(radius - (F) 0.5 - (F) 2 * ray_eps));
        o.rfilter_taps = (uint32_t) std::ceil((radius - (F) 2 * ray_eps) * (F) 2);
        if (radius > (F) 0.5 + ray_eps && o.rfilter_taps > B2VP_RFILTER_MAX_TAPS)
            throw Error(B2VP_ERR_UNSUPPORTED, "reconstruction filter: a radius of " + std::to_string((double) radius) +
                                              " pixels is more than the B200 volpath carries");
    }

    std::unique_ptr<Scene> sc_;
    F bmin_[3], bmax_[3];
    bool have_sensor_ = false, have_integrator_ = false;
};

/* The film of one render: crop_size x 5 channels (X, Y, Z, A, W), ImageBlock layout (imageblock.cpp); 11 channels
   under a moment integrator (+ nested X Y Z and their squares). */
struct Film {
    int width = 0, height = 0, channels = B2VP_FILM_CHANNELS, precision = B2VP_PREC_F32;
    std::vector<float> f32; std::vector<double> f64;
    B2vpCounters totals{};
    double at(int x, int y, int ch) const {
        size_t i = ((size_t) y * width + x) * channels + ch;
        return precision == B2VP_PREC_F32 ? (double) f32[i] : f64[i];
    }
    /* HDRFilm::bitmap for pixel_format=luminance: Y / W (hdrfilm.cpp:251-321) */
    double luminance(int x, int y) const { double w = at(x, y, 4); return w > 0 ? at(x, y, 1) / w : 0.0; }
    /* film->develop() for file_format=pfm, pixel_format=luminance (Bitmap::write_pfm, src/libcore/bitmap.cpp:2206-2230):
       `Pf`, size, `-1` (little endian), scan lines bottom to top */
    void write_pfm(const std::string &path) const {
        std::ofstream f(path, std::ios::binary);
        if (!f) throw Error(B2VP_ERR_INVALID, "cannot write " + path);
        f << "Pf\n" << width << " " << height << "\n-1\n";
        std::vector<float> row((size_t) width);
        for (int y = height - 1; y >= 0; --y) {
            for (int x = 0; x < width; ++x) row[(size_t) x] = (float) luminance(x, y);
            f.write(reinterpret_cast<const char *>(row.data()), (std::streamsize) (row.size() * sizeof(float)));
        }
    }
    /* film->develop() for file_format=openexr, pixel_format=luminance, component_format=float32 (hdrfilm.cpp:251-343):
       a scan-line OpenEXR 2 file with the one channel `Y`, uncompressed, written without the OpenEXR library */
    void write_exr(const std::string &path) const {
        std::string h;
        auto put32 = [&h](int32_t v) { h.append(reinterpret_cast<const char *>(&v), 4); };
        auto putf = [&h](float v) { h.append(reinterpret_cast<const char *>(&v), 4); };
        auto attr = [&](const char *name, const char *type, const std::string &payload) {
            h.append(name); h.push_back('\0'); h.append(type); h.push_back('\0');
            put32((int32_t) payload.size()); h.append(payload);
        };
        put32(20000630); put32(2);
        { std::string c("Y"); c.push_back('\0'); int32_t v[4] = { 2 /* FLOAT */, 0, 1, 1 };
          c.append(reinterpret_cast<const char *>(v), 16); c.push_back('\0'); attr("channels", "chlist", c); }
        attr("compression", "compression", std::string(1, '\0'));
        { int32_t b[4] = { 0, 0, width - 1, height - 1 }; std::string w(reinterpret_cast<const char *>(b), 16);
          attr("dataWindow", "box2i", w); attr("displayWindow", "box2i", w); }
        attr("lineOrder", "lineOrder", std::string(1, '\0'));
        { std::string one; float v = 1.0f; one.append(reinterpret_cast<const char *>(&v), 4);
          attr("pixelAspectRatio", "float", one);
          std::string z(8, '\0'); attr("screenWindowCenter", "v2f", z);
          attr("screenWindowWidth", "float", one); }
        (void) putf;
        h.push_back('\0');
        std::ofstream f(path, std::ios::binary);
        if (!f) throw Error(B2VP_ERR_INVALID, "cannot write " + path);
        f.write(h.data(), (std::streamsize) h.size());
        const uint64_t row_bytes = (uint64_t) width * 4, first = h.size() + 8ull * (uint64_t) height;
        for (int y = 0; y < height; ++y) { uint64_t o = first + (uint64_t) y * (row_bytes + 8); f.write(reinterpret_cast<const char *>(&o), 8); }
        std::vector<float> row((size_t) width);
        for (int y = 0; y < height; ++y) {
            int32_t head[2] = { y, (int32_t) row_bytes };
            f.write(reinterpret_cast<const char *>(head), 8);
            for (int x = 0; x < width; ++x) row[(size_t) x] = (float) luminance(x, y);
            f.write(reinterpret_cast<const char *>(row.data()), (std::streamsize) row_bytes);
        }
    }
};

/* What Integrator::render is given besides the scene: the stream range, and the knobs of B2vpRenderArgs that the
   reference reads from integrator properties (`timeout`) or has no counterpart for (math mode, devices, majorant grid) */
struct RenderOptions {
    int seed_mode = B2VP_SEED_WAVEFRONT;
    uint64_t first_stream = 0, n_streams = 0;
    int math = B2VP_MATH_EXACT;        /* B2VP_MATH_FAST: the fp32 tolerance mode */
    uint32_t majorant_grid = 0;        /* opt-in local majorants (statistically, not per path, the reference's estimator) */
    double timeout_s = 0;              /* integrator.h:143-146: <= 0 = none */
    std::vector<int32_t> devices;      /* empty: the scene's device; else devices[0] must be it (multi-GPU render) */
};

/* A scene resident on one CUDA device: b2vp_scene_create .. b2vp_scene_destroy. */
class DeviceScene {
public:
    DeviceScene(std::unique_ptr<Scene> scene, int device = 0) : scene_(std::move(scene)), device_(device) {
        check(b2vp_scene_create(&scene_->desc, device, &h_));
    }
    ~DeviceScene() { if (h_) b2vp_scene_destroy(h_); }
    DeviceScene(const DeviceScene &) = delete;
    DeviceScene &operator=(const DeviceScene &) = delete;
    const Scene &scene() const { return *scene_; }
    /* traverse() -> "data" -> GridVolume::parameters_changed (src/textures/grid3d.cpp:366-391): new values for a
       grid of unchanged resolution, written in place on the device */
    void update_grid(int volume, const std::vector<float> &data, bool keep_max = false) {
        check(b2vp_scene_update_grid(h_, volume, data.data(), data.size(), keep_max ? 1 : 0));
    }
    /* Integrator::render: all streams of the seeding mode, or the range [first, first + n) */
    Film render(int seed_mode = B2VP_SEED_WAVEFRONT, uint64_t first_stream = 0, uint64_t n_streams = 0) {
        RenderOptions o;
        o.seed_mode = seed_mode; o.first_stream = first_stream; o.n_streams = n_streams;
        return render(o);
    }
    /* kd-tree statistics of the reference's log, here the BVH's; and which kernel instance the last render launched */
    B2vpSceneInfo info() const {
        B2vpSceneInfo i{};
        i.abi_version = B2VP_ABI_VERSION; i.struct_size = (uint32_t) sizeof(i);
        check(b2vp_scene_info(h_, &i));
        return i;
    }
    Film render(const RenderOptions &opt) {
        const int seed_mode = opt.seed_mode;
        const uint64_t first_stream = opt.first_stream, n_streams = opt.n_streams;
        Film f;
        f.width = scene_->width(); f.height = scene_->height(); f.precision = scene_->desc.precision;
        f.channels = scene_->desc.integrator.moment ? B2VP_FILM_CHANNELS_MOMENT : B2VP_FILM_CHANNELS;
        size_t n = (size_t) f.width * f.height * f.channels;
        void *buf;
        if (f.precision == B2VP_PREC_F32) { f.f32.assign(n, 0.f); buf = f.f32.data(); }
        else { f.f64.assign(n, 0.0); buf = f.f64.data(); }
        B2vpRenderArgs a{};
        a.abi_version = B2VP_ABI_VERSION; a.struct_size = (uint32_t) sizeof(a);
        a.precision = f.precision; a.seed_mode = seed_mode; a.device = device_;
        a.first_stream = first_stream; a.n_streams = n_streams; a.clear_film = 1; a.totals = &f.totals;
        a.math_mode = opt.math; a.majorant_grid = opt.majorant_grid; a.timeout_s = opt.timeout_s;
        if (!opt.devices.empty()) { a.devices = opt.devices.data(); a.n_devices = (uint32_t) opt.devices.size(); }
        check(b2vp_render(h_, &a, buf));
        return f;
    }
    void cancel() { b2vp_cancel(h_); }   /* Integrator::cancel (integrator.h:51) */
private:
    std::unique_ptr<Scene> scene_;
    int device_ = 0;
    b2vp_scene_t *h_ = nullptr;
};

inline GridVolume read_vol(const std::string &path, std::array<float, 6> *bbox) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw Error(B2VP_ERR_INVALID, path + ": file does not exist!");
    char head[48];
    f.read(head, 48);
    if (!f || std::memcmp(head, "VOL", 3) != 0 || head[3] != 3) throw Error(B2VP_ERR_INVALID, path + ": not a version-3 VOL file");
    int32_t v[5];
    std::memcpy(v, head + 4, 20);
    if (v[0] != 1) throw Error(B2VP_ERR_UNSUPPORTED, path + ": only float32 volumes are supported");
    GridVolume g;
    g.nx = v[1]; g.ny = v[2]; g.nz = v[3]; g.channels = v[4];
    if ((int64_t) g.nx * g.ny * g.nz < 8)
        throw Error(B2VP_ERR_INVALID, path + ": invalid grid dimensions (must have at least one value at each corner)");
    if (bbox) std::memcpy(bbox->data(), head + 24, 24);
    g.data.resize((size_t) g.nx * g.ny * g.nz * g.channels);
    f.read(reinterpret_cast<char *>(g.data.data()), (std::streamsize) (g.data.size() * sizeof(float)));
    if (!f) throw Error(B2VP_ERR_INVALID, path + ": file is truncated");
    return g;
}

inline MeshData load_obj(const std::string &path, bool flip_tex_coords) {
    std::ifstream f(path);
    if (!f) throw Error(B2VP_ERR_INVALID, path + ": file does not exist!");
    std::vector<std::array<float, 3>> vs, vns;
    std::vector<std::array<float, 2>> vts;
    std::map<std::array<int, 3>, uint32_t> ids;
    std::vector<std::array<int, 3>> order;
    MeshData m;
    std::string line, tok;
    while (std::getline(f, line)) {
        std::istringstream in(line);
        if (!(in >> tok)) continue;
        if (tok == "v" || tok == "vn") {
            std::string a, b, c;
            in >> a >> b >> c;
            (tok == "v" ? vs : vns).push_back({ std::stof(a), std::stof(b), std::stof(c) });   /* the reference parses with stof */
        } else if (tok == "vt") {
            std::string a, b;
            in >> a >> b;
            const float u = std::stof(a), v = std::stof(b);
            vts.push_back({ u, flip_tex_coords ? 1.0f - v : v });
        } else if (tok == "f") {
            std::vector<uint32_t> corner;
            std::string item;
            while (in >> item) {
                std::array<int, 3> key{ 0, 0, 0 };
                size_t p0 = 0;
                for (int k = 0; k < 3 && p0 <= item.size(); ++k) {
                    size_t p1 = item.find('/', p0);
                    std::string part = item.substr(p0, p1 == std::string::npos ? std::string::npos : p1 - p0);
                    if (!part.empty()) key[k] = std::stoi(part);
                    if (p1 == std::string::npos) break;
                    p0 = p1 + 1;
                }
                if (key[0] < 1 || key[0] > (int) vs.size())
                    throw Error(B2VP_ERR_INVALID, "Error while loading OBJ file \"" + path + "\": reference to invalid vertex " + std::to_string(key[0]) + "!");
                auto it = ids.find(key);
                if (it == ids.end()) { it = ids.emplace(key, (uint32_t) order.size()).first; order.push_back(key); }
                corner.push_back(it->second);
            }
            for (size_t k = 1; k + 1 < corner.size(); ++k) { m.faces.push_back(corner[0]); m.faces.push_back(corner[k]); m.faces.push_back(corner[k + 1]); }
        }
    }
    for (const auto &k : order) {
        const auto &p = vs[(size_t) k[0] - 1];
        m.positions.insert(m.positions.end(), p.begin(), p.end());
        if (!vns.empty()) {
            std::array<float, 3> n{ 0, 0, 0 };
            if (k[2] >= 1 && k[2] <= (int) vns.size()) n = vns[(size_t) k[2] - 1];
            m.normals.insert(m.normals.end(), n.begin(), n.end());
        }
        if (!vts.empty()) {
            std::array<float, 2> t{ 0, 0 };
            if (k[1] >= 1 && k[1] <= (int) vts.size()) t = vts[(size_t) k[1] - 1];
            m.texcoords.insert(m.texcoords.end(), t.begin(), t.end());
        }
    }
    return m;
}

namespace detail {
inline std::string read_file(const std::string &path, const std::string &what) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw Error(B2VP_ERR_INVALID, "Error while loading " + what + " file \"" + path + "\": file not found!");
    std::ostringstream ss;
    ss << f.rdbuf();
    return ss.str();
}
/* a scalar of the given PLY type at p (byte order `big`), as double; returns its size */
inline size_t ply_scalar(const std::string &type, const unsigned char *p, bool big, double &out) {
    auto load = [&](size_t n, unsigned char *dst) { for (size_t i = 0; i < n; ++i) dst[i] = p[big ? n - 1 - i : i]; };
    if (type == "char" || type == "int8") { out = (double) (int8_t) p[0]; return 1; }
    if (type == "uchar" || type == "uint8") { out = (double) p[0]; return 1; }
    if (type == "short" || type == "int16") { int16_t v; load(2, (unsigned char *) &v); out = v; return 2; }
    if (type == "ushort" || type == "uint16") { uint16_t v; load(2, (unsigned char *) &v); out = v; return 2; }
    if (type == "int" || type == "int32") { int32_t v; load(4, (unsigned char *) &v); out = v; return 4; }
    if (type == "uint" || type == "uint32") { uint32_t v; load(4, (unsigned char *) &v); out = v; return 4; }
    if (type == "float" || type == "float32") { float v; load(4, (unsigned char *) &v); out = v; return 4; }
    if (type == "double" || type == "float64") { double v; load(8, (unsigned char *) &v); out = v; return 8; }
    throw Error(B2VP_ERR_INVALID, "PLY: unknown property type " + type);
}
} // namespace detail

inline MeshData load_ply(const std::string &path) {
    const std::string raw = detail::read_file(path, "PLY");
    auto fail = [&](const std::string &msg) -> void { throw Error(B2VP_ERR_INVALID, "Error while loading PLY file \"" + path + "\": " + msg + "!"); };
    const size_t end = raw.find("end_header");
    if (raw.compare(0, 3, "ply") != 0 || end == std::string::npos) fail("invalid PLY header");
    const size_t body = raw.find('\n', end) + 1;
    struct Prop { std::string name, type, count_type; bool list = false; };
    struct Element { std::string name; size_t count = 0; std::vector<Prop> props; };
    std::vector<Element> elements;
    std::string format;
    {
        std::istringstream header(raw.substr(0, end));
        std::string line;
        std::getline(header, line);   /* "ply" */
        while (std::getline(header, line)) {
            std::istringstream in(line);
            std::string tok;
            if (!(in >> tok) || tok == "comment" || tok == "obj_info") continue;
            if (tok == "format") in >> format;
            else if (tok == "element") { Element e; in >> e.name >> e.count; elements.push_back(e); }
            else if (tok == "property" && !elements.empty()) {
                Prop p; std::string t;
                in >> t;
                if (t == "list") { p.list = true; in >> p.count_type >> p.type >> p.name; }
                else { p.type = t; in >> p.name; }
                elements.back().props.push_back(p);
            }
        }
    }
    if (format != "ascii" && format != "binary_little_endian" && format != "binary_big_endian") fail("invalid PLY format");
    const bool ascii = format == "ascii", big = format == "binary_big_endian";
    std::istringstream words(ascii ? raw.substr(body) : std::string());
    size_t off = body;
    auto next = [&](const std::string &type) -> double {
        double v = 0;
        if (ascii) { if (!(words >> v)) fail("file is truncated"); return v; }
        unsigned char buf[8] = { 0 };   /* a copy, so that a truncated file is never read past its end */
        const size_t left = off < raw.size() ? raw.size() - off : 0;
        std::memcpy(buf, raw.data() + std::min(off, raw.size()), std::min<size_t>(8, left));
        const size_t n = detail::ply_scalar(type, buf, big, v);
        if (n > left) fail("file is truncated");
        off += n;
        return v;
    };
    MeshData m;
    bool have_vertex = false, have_face = false;
    for (const Element &el : elements) {
        std::map<std::string, int> col;
        int n_lists = 0;
        for (size_t i = 0; i < el.props.size(); ++i) { col[el.props[i].name] = (int) i; n_lists += el.props[i].list; }
        if (n_lists && !(el.name == "face" && n_lists == 1)) fail("unsupported list property in element '" + el.name + "'");
        const bool vertex = el.name == "vertex", face = el.name == "face";
        int cx = -1, cy = -1, cz = -1, cnx = -1, cny = -1, cnz = -1, cu = -1, cv = -1;
        auto find = [&](const char *k) { auto it = col.find(k); return it == col.end() ? -1 : it->second; };
        if (vertex) {
            cx = find("x"); cy = find("y"); cz = find("z");
            if (cx < 0 || cy < 0 || cz < 0) fail("vertex positions not found");
            cnx = find("nx"); cny = find("ny"); cnz = find("nz");
            if (cnx < 0 || cny < 0 || cnz < 0) cnx = cny = cnz = -1;
            const char *names[3][2] = { { "u", "v" }, { "texture_u", "texture_v" }, { "s", "t" } };
            for (auto &pr : names) if (find(pr[0]) >= 0 && find(pr[1]) >= 0) { cu = find(pr[0]); cv = find(pr[1]); break; }
        }
        if (face && find("vertex_index") < 0 && find("vertex_indices") < 0) fail("vertex_index/vertex_indices property not found");
        std::vector<double> rec(el.props.size());
        for (size_t r = 0; r < el.count; ++r) {
            for (size_t i = 0; i < el.props.size(); ++i) {
                const Prop &p = el.props[i];
                if (!p.list) { rec[i] = next(p.type); continue; }
                const int cnt = (int) next(p.count_type);
                if (cnt != 3) fail("incompatible contents -- is this a triangle mesh?");
                for (int k = 0; k < 3; ++k) m.faces.push_back((uint32_t) next(p.type));
            }
            if (vertex) {
                for (int c : { cx, cy, cz }) m.positions.push_back((float) rec[(size_t) c]);
                if (cnx >= 0) for (int c : { cnx, cny, cnz }) m.normals.push_back((float) rec[(size_t) c]);
                if (cu >= 0) for (int c : { cu, cv }) m.texcoords.push_back((float) rec[(size_t) c]);
            }
        }
        have_vertex |= vertex; have_face |= face;
        if (!face && !vertex) continue;
    }
    if (!ascii && off != raw.size()) fail("invalid file -- trailing content");
    if (!have_vertex || !have_face) fail("vertex / face element not found");
    return m;
}

#ifdef B2VP_WITH_ZLIB
inline MeshData load_serialized(const std::string &path, int shape_index) {
    auto fail = [&](const std::string &msg) -> void { throw Error(B2VP_ERR_INVALID, "Error while loading serialized file \"" + path + "\": " + msg + "!"); };
    if (shape_index < 0) fail("shape index must be nonnegative");
    const std::string raw = detail::read_file(path, "serialized");
    auto rd = [&](size_t at, size_t n, void *dst) { if (at + n > raw.size()) fail("file is truncated"); std::memcpy(dst, raw.data() + at, n); };
    uint16_t fmt = 0, version = 0;
    rd(0, 2, &fmt); rd(2, 2, &version);
    if (fmt != 0x041C) fail("encountered an invalid file format");
    if (version != 3 && version != 4) fail("encountered an incompatible file version");
    size_t start = 0;
    if (shape_index != 0) {   /* the offset dictionary at the end of the file (serialized.cpp:215-246) */
        uint32_t count = 0;
        rd(raw.size() - 4, 4, &count);
        if ((uint32_t) shape_index > count)
            fail("Unable to unserialize mesh, shape index is out of range! (requested " + std::to_string(shape_index) + " out of 0.." + std::to_string((int) count - 1) + ")");
        if (version == 4) { uint64_t o; rd(raw.size() - 8 * (size_t) (count - (uint32_t) shape_index) - 4, 8, &o); start = (size_t) o; }
        else { uint32_t o; rd(raw.size() - 4 * (size_t) (count - (uint32_t) shape_index + 1), 4, &o); start = o; }
    }
    if (start + 4 > raw.size()) fail("file is truncated");
    std::string data;
    {
        z_stream zs{};
        if (inflateInit(&zs) != Z_OK) fail("inflateInit()");
        zs.next_in = (Bytef *) (raw.data() + start + 4); zs.avail_in = (uInt) (raw.size() - start - 4);
        std::vector<char> buf(1 << 16);
        int rc = Z_OK;
        while (rc == Z_OK) {
            zs.next_out = (Bytef *) buf.data(); zs.avail_out = (uInt) buf.size();
            rc = inflate(&zs, Z_NO_FLUSH);
            data.append(buf.data(), buf.size() - zs.avail_out);
        }
        inflateEnd(&zs);
        if (rc != Z_STREAM_END) fail("inflate(): stream error");
    }
    size_t off = 0;
    auto take = [&](size_t n, void *dst) { if (off + n > data.size()) fail("stream is truncated"); std::memcpy(dst, data.data() + off, n); off += n; };
    uint32_t flags = 0;
    take(4, &flags);
    if (version == 4) { const size_t z = data.find('\0', off); if (z == std::string::npos) fail("stream is truncated"); off = z + 1; }
    uint64_t n_vert = 0, n_face = 0;
    take(8, &n_vert); take(8, &n_face);
    const bool dbl = (flags & 0x2000u) != 0;
    auto reals = [&](size_t dim, std::vector<float> *out) {
        for (uint64_t i = 0; i < n_vert * dim; ++i) {
            float v;
            if (dbl) { double d; take(8, &d); v = (float) d; } else take(4, &v);
            if (out) out->push_back(v);
        }
    };
    MeshData m;
    reals(3, &m.positions);
    if (flags & 0x0001u) reals(3, &m.normals);
    if (flags & 0x0002u) reals(2, &m.texcoords);
    if (flags & 0x0008u) reals(3, nullptr);
    m.faces.resize((size_t) n_face * 3);
    take(m.faces.size() * 4, m.faces.data());
    return m;
}
#endif

} // namespace b2vp
