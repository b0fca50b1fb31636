/*
 * volpath_oracle.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the reference's `volpath` hot path in mono mode
 * (scalar_mono = float, scalar_mono_double = double), written as plain scalar
 * C++ that follows the reference's control flow statement by statement so that
 * RNG draw order, event counts and arithmetic order can be audited against it.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
 * arm may load this library; the product (eradiate-kernel_b200/) never does.
 *
 * Parity status: pinned against the reference's own golden vectors where they
 * exist (TEA, PCG32, RPV eval, analytic sensor/shape/emitter results, the
 * rho/pi end-to-end test) and against reference outputs recorded in
 * SURVEY.md/BASELINE.md (statistical).  `VolumetricPathIntegrator::sample`
 * itself has no test in the reference and the reference cannot
 * be built here without its cmake build system, so per-path bit parity with
 * the reference binary is UNPINNED; see DESIGN.md.
 *
 * Floating point: IEEE operations in the order written here, explicit fma only
 * where the reference writes fmadd/fmsub/fnmadd, libm replaced by
 * oracle_math.h.  Compile with -ffp-contract=off.
 *
 * Reference files followed (relative to the reference tree):
 *   src/integrators/volpath.cpp:38-483     src/librender/integrator.cpp:181-288
 *   src/librender/medium.cpp:36-89         src/media/{heterogeneous,homogeneous}.cpp
 *   src/textures/grid3d.cpp:220-360        src/phase/{rayleigh,hg,isotropic}.cpp
 *   src/bsdfs/{rpv,diffuse,null}.cpp       src/sensors/{distant,mdistant,radiancemeter,mradiancemeter}.cpp
 *   src/emitters/directional.cpp:108-135   src/librender/scene.cpp:168-235
 *   src/shapes/{rectangle,sphere}.cpp      include/mitsuba/render/mesh.h:195-226
 *   src/librender/mesh.cpp:450-545         include/mitsuba/render/kdtree.h:2078-2171
 *   include/mitsuba/render/interaction.h   include/mitsuba/core/{ray,bbox,vector,frame,warp,math,transform,random}.h
 *   ext/enoki/include/enoki/random.h       src/librender/{sampler,spiral,imageblock}.cpp
 */
#include "../include/b2vp.h"
#include "oracle_math.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <limits>
#include <string>
#include <vector>

namespace {

using namespace omath;

thread_local std::string g_error;

/* ======================================================================
 *  PCG32 + TEA  (ext/enoki/include/enoki/random.h:33-150,
 *                include/mitsuba/core/random.h:105-116)
 * ====================================================================== */
constexpr uint64_t PCG32_DEFAULT_STATE  = 0x853c49e6748fea9bULL;
constexpr uint64_t PCG32_DEFAULT_STREAM = 0xda3e39cb94b95bdbULL;
constexpr uint64_t PCG32_MULT           = 0x5851f42d4c957f2dULL;

struct PCG32 {
    uint64_t state = 0, inc = 0;
    uint32_t draws = 0; /* instrumentation: number of advances */

    void seed(uint64_t initstate, uint64_t initseq) { /* random.h:62-68 */
        state = 0;
        inc = (initseq << 1) | 1u;
        next_uint32(true);
        state += initstate;
        next_uint32(true);
        draws = 0;
    }
    /* masked variant: does not advance when the mask is false (random.h:79-85) */
    uint32_t next_uint32(bool mask) {
        uint64_t oldstate = state;
        if (mask) { state = oldstate * PCG32_MULT + inc; draws++; }
        uint32_t xorshifted = (uint32_t) (((oldstate >> 18u) ^ oldstate) >> 27u);
        uint32_t rot = (uint32_t) (oldstate >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((-rot) & 31));
    }
    float next_float(float, bool mask) { /* random.h:113-120 */
        return from_bits<float>((next_uint32(mask) >> 9) | 0x3f800000u) - 1.f;
    }
    double next_float(double, bool mask) { /* random.h:129-141: ONE u32 per double */
        return from_bits<double>(((uint64_t) next_uint32(mask) << 20) | 0x3ff0000000000000ull) - 1.0;
    }
};

/* sample_tea_64 as instantiated by PCG32Sampler::seed in wavefront mode: the
   arguments are UInt64 arrays, so the rounds run in 64-bit arithmetic
   (src/librender/sampler.cpp:89-92, include/mitsuba/core/random.h:105-116). */
uint64_t tea64_u64(uint64_t v0, uint64_t v1, int rounds = 4) {
    uint64_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9;
        v0 += ((v1 << 4) + 0xa341316c) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4);
        v1 += ((v0 << 4) + 0xad90777d) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761e);
    }
    return v0 + (v1 << 32);
}

/* The 32-bit flavours the reference's test_random.py pins. */
uint32_t tea32(uint32_t v0, uint32_t v1, int rounds) {
    uint32_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9;
        v0 += ((v1 << 4) + 0xa341316c) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4);
        v1 += ((v0 << 4) + 0xad90777d) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761e);
    }
    return v1;
}
uint64_t tea64(uint32_t v0, uint32_t v1, int rounds) {
    uint32_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9;
        v0 += ((v1 << 4) + 0xa341316c) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4);
        v1 += ((v0 << 4) + 0xad90777d) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761e);
    }
    return (uint64_t) v0 + ((uint64_t) v1 << 32);
}

/* enoki::morton_decode<Point2u> (ext/enoki/include/enoki/morton.h): even bits -> x */
inline uint32_t compact_bits(uint32_t x) {
    x &= 0x55555555u;
    x = (x ^ (x >> 1)) & 0x33333333u;
    x = (x ^ (x >> 2)) & 0x0f0f0f0fu;
    x = (x ^ (x >> 4)) & 0x00ff00ffu;
    x = (x ^ (x >> 8)) & 0x0000ffffu;
    return x;
}
inline void morton_decode(uint32_t m, uint32_t &x, uint32_t &y) {
    x = compact_bits(m);
    y = compact_bits(m >> 1);
}

/* Spiral block order (src/librender/spiral.cpp:8-72). */
struct SpiralBlock { int ox, oy, sx, sy; };
std::vector<SpiralBlock> spiral_blocks(int w, int h, int offx, int offy, int bs) {
    int bx = (int) std::ceil((float) w / (float) bs), by = (int) std::ceil((float) h / (float) bs);
    int count = bx * by, counter = 0;
    int px = bx / 2, py = by / 2, dir = 0 /*Right,Down,Left,Up*/, steps_left = 1, steps = 1;
    std::vector<SpiralBlock> out;
    while (counter < count) {
        SpiralBlock b;
        b.ox = px * bs; b.oy = py * bs;
        b.sx = std::min(bs, w - b.ox); b.sy = std::min(bs, h - b.oy);
        b.ox += offx; b.oy += offy;
        out.push_back(b);
        ++counter;
        if (counter != count) {
            do {
                switch (dir) {
                    case 0: ++px; break; case 1: ++py; break;
                    case 2: --px; break; case 3: --py; break;
                }
                if (--steps_left == 0) {
                    dir = (dir + 1) % 4;
                    if (dir == 2 || dir == 0) ++steps;
                    steps_left = steps;
                }
            } while (px < 0 || py < 0 || px >= bx || py >= by);
        }
    }
    return out;
}

/* ======================================================================
 *  Small vector algebra in the operation order fixed by oracle_math.h
 * ====================================================================== */
template <typename F> struct V3 { F x, y, z; };
template <typename F> inline V3<F> operator+(V3<F> a, V3<F> b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
template <typename F> inline V3<F> operator-(V3<F> a, V3<F> b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
template <typename F> inline V3<F> operator-(V3<F> a) { return { -a.x, -a.y, -a.z }; }
template <typename F> inline V3<F> operator*(V3<F> a, F s) { return { a.x * s, a.y * s, a.z * s }; }
template <typename F> inline V3<F> operator*(F s, V3<F> a) { return { s * a.x, s * a.y, s * a.z }; }
template <typename F> inline F dot(V3<F> a, V3<F> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
/* enoki::cross: fmsub(a.yzx, b.zxy, a.zxy * b.yzx) (array_router.h:684) */
template <typename F> inline V3<F> cross(V3<F> a, V3<F> b) {
    return { fma_(a.y, b.z, -(a.z * b.y)), fma_(a.z, b.x, -(a.x * b.z)), fma_(a.x, b.y, -(a.y * b.x)) };
}
template <typename F> inline F rcp_(F x) { return (F) 1 / x; }
template <typename F> inline F rsqrt_(F x) { return (F) 1 / sqrt_(x); }
template <typename F> inline V3<F> normalize(V3<F> a) { return a * rsqrt_(dot(a, a)); }
template <typename F> inline F norm(V3<F> a) { return sqrt_(dot(a, a)); }
template <typename F> inline F sqr(F x) { return x * x; }
template <typename F> inline F abs_(F x) { return x < 0 ? -x : (x == 0 ? (F) 0 : x); }
template <typename F> inline F max_(F a, F b) { return a > b ? a : b; } /* maxps-like: returns b on NaN a */
template <typename F> inline F min_(F a, F b) { return a < b ? a : b; }
template <typename F> inline F hmax_abs(V3<F> p) { return max_(max_(abs_(p.x), abs_(p.y)), abs_(p.z)); }
template <typename F> inline bool signbit_(F x) { return std::signbit(x); }
template <typename F> inline F mulsign(F a, F b) { return signbit_(b) ? -a : a; }
template <typename F> inline F mulsign_neg(F a, F b) { return signbit_(b) ? a : -a; }
template <typename F> inline V3<F> fmadd3(V3<F> a, F s, V3<F> b) { /* fmadd(a, s, b) */
    return { fma_(a.x, s, b.x), fma_(a.y, s, b.y), fma_(a.z, s, b.z) };
}

template <typename F> struct Consts {
    static constexpr F Pi = (F) 3.14159265358979323846;
    static constexpr F InvPi = (F) 0.31830988618379067154;
    static constexpr F InvFourPi = (F) 0.07957747154594766788;
    static constexpr F InvTwoPi = (F) 0.15915494309189533577;
    static constexpr F Epsilon = std::numeric_limits<F>::epsilon() / 2;  /* math.h:35 */
    static constexpr F RayEpsilon = Epsilon * 1500;                      /* math.h:36 */
    static constexpr F ShadowEpsilon = RayEpsilon * 10;                  /* math.h:37 */
    static constexpr F Inf = std::numeric_limits<F>::infinity();
};

/* coordinate_system (include/mitsuba/core/vector.h:116-136) */
template <typename F> inline void coordinate_system(V3<F> n, V3<F> &s, V3<F> &t) {
    F sign = signbit_(n.z) ? (F) -1 : (F) 1;
    F a = -rcp_(sign + n.z);
    F b = n.x * n.y * a;
    s = { mulsign(sqr(n.x) * a, n.z) + (F) 1, mulsign(b, n.z), mulsign_neg(n.x, n.z) };
    t = { b, sign + sqr(n.y) * a, -n.y };
}

template <typename F> struct Frame {
    V3<F> s, t, n;
    V3<F> to_local(V3<F> v) const { return { dot(v, s), dot(v, t), dot(v, n) }; }  /* frame.h:30 */
    V3<F> to_world(V3<F> v) const { return s * v.x + t * v.y + n * v.z; }         /* frame.h:35 */
};
template <typename F> inline Frame<F> make_frame(V3<F> v) {
    Frame<F> f; f.n = v; coordinate_system(v, f.s, f.t); return f;
}

/* Column-major 4x4 (enoki::Matrix); operators from transform.h:92-128. */
template <typename F> struct M4 {
    F m[16];
    F c(int col, int row) const { return m[col * 4 + row]; }
};
template <typename F> inline M4<F> load_mat(const B2vpMat4 &a) {
    M4<F> r; for (int i = 0; i < 16; ++i) r.m[i] = (F) a.m[i]; return r;
}
/* Transform * Point with the w divide (transform.h:107-115) */
template <typename F> inline V3<F> xf_point(const M4<F> &M, V3<F> p) {
    F r[4];
    for (int k = 0; k < 4; ++k) {
        F v = M.c(3, k);
        v = fma_(M.c(0, k), p.x, v);
        v = fma_(M.c(1, k), p.y, v);
        v = fma_(M.c(2, k), p.z, v);
        r[k] = v;
    }
    return { r[0] / r[3], r[1] / r[3], r[2] / r[3] };
}
/* transform_affine(Point) (transform.h:93-101) */
template <typename F> inline V3<F> xf_point_affine(const M4<F> &M, V3<F> p) {
    F r[3];
    for (int k = 0; k < 3; ++k) {
        F v = M.c(3, k);
        v = fma_(M.c(0, k), p.x, v);
        v = fma_(M.c(1, k), p.y, v);
        v = fma_(M.c(2, k), p.z, v);
        r[k] = v;
    }
    return { r[0], r[1], r[2] };
}
/* Transform * Vector (transform.h:121-129) */
template <typename F> inline V3<F> xf_vector(const M4<F> &M, V3<F> d) {
    F r[3];
    for (int k = 0; k < 3; ++k) {
        F v = M.c(0, k) * d.x;
        v = fma_(M.c(1, k), d.y, v);
        v = fma_(M.c(2, k), d.z, v);
        r[k] = v;
    }
    return { r[0], r[1], r[2] };
}

/* Ray (include/mitsuba/core/ray.h:30-65) */
template <typename F> struct Ray {
    V3<F> o, d, d_rcp;
    F mint = Consts<F>::RayEpsilon, maxt = Consts<F>::Inf;
    void update() { d_rcp = { rcp_(d.x), rcp_(d.y), rcp_(d.z) }; }
    V3<F> at(F t) const { return fmadd3(d, t, o); }
};
template <typename F> inline Ray<F> make_ray(V3<F> o, V3<F> d, F mint, F maxt) {
    Ray<F> r; r.o = o; r.d = d; r.mint = mint; r.maxt = maxt; r.update(); return r;
}

/* BoundingBox::ray_intersect (include/mitsuba/core/bbox.h:302-325) */
template <typename F> struct BBox { V3<F> min, max; };
template <typename F>
inline bool bbox_ray_intersect(const BBox<F> &b, const Ray<F> &ray, F &mint, F &maxt) {
    bool active = (ray.d.x != 0 || ray.o.x > b.min.x || ray.o.x < b.max.x) &&
                  (ray.d.y != 0 || ray.o.y > b.min.y || ray.o.y < b.max.y) &&
                  (ray.d.z != 0 || ray.o.z > b.min.z || ray.o.z < b.max.z);
    V3<F> t1 = { (b.min.x - ray.o.x) * ray.d_rcp.x, (b.min.y - ray.o.y) * ray.d_rcp.y,
                 (b.min.z - ray.o.z) * ray.d_rcp.z };
    V3<F> t2 = { (b.max.x - ray.o.x) * ray.d_rcp.x, (b.max.y - ray.o.y) * ray.d_rcp.y,
                 (b.max.z - ray.o.z) * ray.d_rcp.z };
    V3<F> t1p = { min_(t1.x, t2.x), min_(t1.y, t2.y), min_(t1.z, t2.z) };
    V3<F> t2p = { max_(t1.x, t2.x), max_(t1.y, t2.y), max_(t1.z, t2.z) };
    mint = max_(max_(t1p.x, t1p.y), t1p.z);
    maxt = min_(min_(t2p.x, t2p.y), t2p.z);
    return active && maxt >= mint;
}

/* ======================================================================
 *  Scene, typed in the variant's Float
 * ====================================================================== */
template <typename F> struct OShape {
    int type, bsdf, interior, exterior;
    M4<F> to_world, to_object;
    Frame<F> frame;      /* rectangle: s = dp_du, t = dp_dv, n */
    V3<F> translation;   /* to_world.translation() */
    F inv_surface_area;
    V3<F> center; F radius; bool flip_normals;
    bool sample_only = false;   /* B2vpShape::sample_only */
    uint32_t vofs, vcount, fofs, fcount; bool has_normals, has_uv;
};
template <typename F> struct OVolume {
    int type, filter, wrap, channels, nx, ny, nz;
    F value, max_value;
    BBox<F> bbox;
    M4<F> world_to_local;
    std::vector<F> data; /* widened to Float at load (volume_data.h:85-90) */
};
template <typename F> struct OMedium {
    int type, sigma_t, albedo, phase; bool spectral, sample_emitters; F scale;
    F max_density; BBox<F> aabb; /* heterogeneous.cpp:29-30 */
};
/* ContinuousDistribution<Float> over [-1, 1] (include/mitsuba/core/distr_1d.h:258-470) as tabphase holds it */
template <typename F> struct TabDistr {
    std::vector<F> pdf, cdf;
    F integral = 0, normalization = 0, interval_size = 0, inv_interval_size = 0;
    uint32_t valid_x = 0, valid_y = 0;
    /* update(), distr_1d.h:291-336: the running integral is accumulated in double */
    void build(const double *values, uint32_t n) {
        pdf.resize(n); cdf.resize(n - 1);
        for (uint32_t i = 0; i < n; ++i) pdf[i] = (F) values[i];
        valid_x = valid_y = (uint32_t) -1;
        double range = 1.0 - (-1.0), isz = range / (n - 1), integ = 0.;
        for (uint32_t i = 0; i + 1 < n; ++i) {
            double y0 = (double) pdf[i], y1 = (double) pdf[i + 1];
            double value = 0.5 * isz * (y0 + y1);
            integ += value;
            cdf[i] = (F) integ;
            if (value > 0.) { if (valid_x == (uint32_t) -1) valid_x = i; valid_y = i; }
        }
        integral = (F) integ; normalization = (F) (1. / integ);
        interval_size = (F) isz; inv_interval_size = (F) (1. / isz);
    }
    /* eval_pdf (distr_1d.h:365-380); outside [-1, 1] the gathers are masked: y0 = y1 = 0 */
    F eval_pdf(F x) const {
        bool active = x >= (F) -1 && x <= (F) 1;
        x = (x - (F) -1) * inv_interval_size;
        /* Index(x): float -> uint32 conversion, clamp to [0, n - 2] */
        long long xi = (long long) x;
        uint32_t index = (uint32_t) std::min<long long>(std::max<long long>(xi, 0), (long long) pdf.size() - 2);
        F y0 = active ? pdf[index] : (F) 0, y1 = active ? pdf[index + 1] : (F) 0;
        F w1 = x - (F) index, w0 = (F) 1 - w1;
        return fma_(w0, y0, w1 * y1);
    }
    /* sample (distr_1d.h:416-439) with enoki::binary_search (array_utils.h:150-171) */
    F sample(F value) const {
        value *= integral;
        uint32_t start = valid_x, end = valid_y;
        uint32_t iterations = 0;
        if (start < end) { uint32_t d = end - start; while (d) { iterations++; d >>= 1; } } /* log2i(d) + 1 */
        for (uint32_t i = 0; i < iterations; ++i) {
            uint32_t middle = (start + end) >> 1;
            bool cond = cdf[middle] < value;
            if (cond) start = std::min(middle + 1, end); else end = middle;
        }
        uint32_t index = start;
        F y0 = pdf[index], y1 = pdf[index + 1], c0 = index > 0 ? cdf[index - 1] : (F) 0;
        value = (value - c0) * inv_interval_size;
        F t_linear = (y0 - safe_sqrt_(sqr(y0) + (F) 2 * value * (y1 - y0))) / (y0 - y1), t_const = value / y0;
        F t = y0 == y1 ? t_const : t_linear;
        return fma_((F) index + t, interval_size, (F) -1);
    }
};

template <typename F> struct OScene {
    std::vector<TabDistr<F>> tabs;   /* one per phase function (empty unless tabulated) */
    int environment = -1;            /* index of the environment emitter (Scene::environment()), or -1 */
    std::vector<OShape<F>> shapes;
    std::vector<B2vpBsdf> bsdfs;
    std::vector<OMedium<F>> media;
    std::vector<OVolume<F>> volumes;
    std::vector<B2vpPhase> phases;
    std::vector<B2vpEmitter> emitters;
    const float *vpos = nullptr, *vnrm = nullptr, *vuv = nullptr;
    const uint32_t *faces = nullptr;
    BBox<F> kd_bbox;      /* Scene bbox padded as in kdtree.h:1811-1813 */
    V3<F> bs_center; F bs_radius;
    B2vpSensor sensor; M4<F> sensor_to_world; std::vector<M4<F>> sensor_xf;
    B2vpIntegrator integ;
};

template <typename F> inline V3<F> ld3(const double *p) { return { (F) p[0], (F) p[1], (F) p[2] }; }

template <typename F> bool build_scene(const B2vpSceneDesc *d, OScene<F> &s) {
    for (uint32_t i = 0; i < d->n_shapes; ++i) {
        const B2vpShape &a = d->shapes[i];
        OShape<F> o;
        o.type = a.type; o.bsdf = a.bsdf; o.interior = a.interior_medium; o.exterior = a.exterior_medium;
        o.sample_only = a.sample_only != 0;
        o.to_world = load_mat<F>(a.to_world); o.to_object = load_mat<F>(a.to_object);
        o.frame.s = ld3<F>(a.frame_s); o.frame.t = ld3<F>(a.frame_t); o.frame.n = ld3<F>(a.frame_n);
        o.translation = { o.to_world.c(3, 0), o.to_world.c(3, 1), o.to_world.c(3, 2) };
        o.inv_surface_area = (F) a.inv_surface_area;
        o.center = ld3<F>(a.center); o.radius = (F) a.radius; o.flip_normals = a.flip_normals != 0;
        o.vofs = a.vertex_offset; o.vcount = a.vertex_count; o.fofs = a.face_offset; o.fcount = a.face_count;
        o.has_normals = a.has_vertex_normals != 0; o.has_uv = a.has_vertex_texcoords != 0;
        s.shapes.push_back(o);
    }
    s.bsdfs.assign(d->bsdfs, d->bsdfs + d->n_bsdfs);
    s.phases.assign(d->phases, d->phases + d->n_phases);
    s.tabs.resize(d->n_phases);
    for (uint32_t i = 0; i < d->n_phases; ++i)
        if (d->phases[i].type == B2VP_PHASE_TABULATED) {
            if (d->phases[i].n_values < 2 || !d->phases[i].values) { g_error = "tabphase needs at least two values"; return false; }
            s.tabs[i].build(d->phases[i].values, d->phases[i].n_values);
        }
    s.emitters.assign(d->emitters, d->emitters + d->n_emitters);
    for (uint32_t i = 0; i < d->n_emitters; ++i)
        if (d->emitters[i].type == B2VP_EMITTER_CONSTANT) {
            if (s.environment >= 0) { g_error = "Only one environment emitter can be specified per scene."; return false; }
            s.environment = (int) i;
        }

    for (uint32_t i = 0; i < d->n_volumes; ++i) {
        const B2vpVolume &a = d->volumes[i];
        OVolume<F> o;
        o.type = a.type; o.filter = a.filter; o.wrap = a.wrap; o.channels = a.channels;
        o.nx = a.res[0]; o.ny = a.res[1]; o.nz = a.res[2];
        o.value = (F) a.value; o.max_value = (F) a.max_value;
        o.bbox.min = ld3<F>(a.bbox_min); o.bbox.max = ld3<F>(a.bbox_max);
        o.world_to_local = load_mat<F>(a.world_to_local);
        if (a.type == B2VP_VOLUME_GRID) {
            size_t n = (size_t) o.nx * o.ny * o.nz * o.channels;
            o.data.resize(n);
            for (size_t k = 0; k < n; ++k) o.data[k] = (F) a.data[k];
        }
        s.volumes.push_back(std::move(o));
    }
    for (uint32_t i = 0; i < d->n_media; ++i) {
        const B2vpMedium &a = d->media[i];
        OMedium<F> o;
        o.type = a.type; o.sigma_t = a.sigma_t; o.albedo = a.albedo; o.phase = a.phase;
        o.spectral = a.has_spectral_extinction != 0; o.sample_emitters = a.sample_emitters != 0;
        o.scale = (F) a.scale;
        o.max_density = 0; o.aabb = { { 0, 0, 0 }, { 0, 0, 0 } };
        if (a.type == B2VP_MEDIUM_HETEROGENEOUS) {
            const OVolume<F> &v = s.volumes[a.sigma_t];
            if (v.type != B2VP_VOLUME_GRID) {
                g_error = "heterogeneous medium needs a gridvolume sigma_t (ConstVolume::max is not implemented, constant3d.cpp:39)";
                return false;
            }
            o.max_density = o.scale * v.max_value;
            o.aabb = v.bbox;
        }
        s.media.push_back(o);
    }
    s.vpos = d->vertex_positions; s.vnrm = d->vertex_normals; s.vuv = d->vertex_texcoords;
    s.faces = d->faces;
    /* kd-tree bounding box: scene bbox padded by (extents + 1) * eps (kdtree.h:1811-1813) */
    V3<F> bmin = ld3<F>(d->bbox_min), bmax = ld3<F>(d->bbox_max);
    V3<F> extra = { ((bmax.x - bmin.x) + (F) 1) * Consts<F>::Epsilon,
                    ((bmax.y - bmin.y) + (F) 1) * Consts<F>::Epsilon,
                    ((bmax.z - bmin.z) + (F) 1) * Consts<F>::Epsilon };
    s.kd_bbox.min = bmin - extra; s.kd_bbox.max = bmax + extra;
    s.bs_center = ld3<F>(d->bsphere_center); s.bs_radius = (F) d->bsphere_radius;
    s.sensor = d->sensor; s.sensor_to_world = load_mat<F>(d->sensor.to_world);
    for (int i = 0; i < d->sensor.n_transforms; ++i) {
        M4<F> m; for (int k = 0; k < 16; ++k) m.m[k] = (F) d->sensor.transforms[i * 16 + k];
        s.sensor_xf.push_back(m);
    }
    s.integ = d->integrator;
    return true;
}

/* ======================================================================
 *  Surface interaction + shapes
 * ====================================================================== */
template <typename F> struct SurfaceInteraction {
    F t = Consts<F>::Inf;
    V3<F> p{ 0, 0, 0 }, n{ 0, 0, 0 }, wi{ 0, 0, 0 }, dp_du{ 0, 0, 0 };
    Frame<F> sh_frame{ { 0, 0, 0 }, { 0, 0, 0 }, { 0, 0, 0 } };
    int shape = -1; uint32_t prim = 0;
    bool is_valid() const { return t != Consts<F>::Inf; }
};

struct Counters { uint32_t medium_steps = 0, n_null = 0, n_real = 0, n_intersect = 0, n_surface = 0; };

/* Rectangle::ray_intersect_preliminary (src/shapes/rectangle.cpp:139-159) */
template <typename F> inline F rectangle_intersect(const OShape<F> &sh, const Ray<F> &ray_) {
    Ray<F> ray = make_ray(xf_point_affine(sh.to_object, ray_.o), xf_vector(sh.to_object, ray_.d),
                          ray_.mint, ray_.maxt);
    F t = -ray.o.z * ray.d_rcp.z;
    V3<F> local = ray.at(t);
    bool active = t >= ray.mint && t <= ray.maxt && abs_(local.x) <= (F) 1 && abs_(local.y) <= (F) 1;
    return active ? t : Consts<F>::Inf;
}

/* math::solve_quadratic (include/mitsuba/core/math.h) in double, as Sphere uses it */
inline bool solve_quadratic(double a, double b, double c, double &x0, double &x1) {
    bool linear_case = a == 0.0;
    bool valid_linear = linear_case && b != 0.0;
    x0 = x1 = -c / b;
    double discrim = fma_(b, b, -(4.0 * a) * c);
    bool valid_quadratic = !linear_case && discrim >= 0.0;
    if (valid_quadratic) {
        double root = std::sqrt(discrim);
        double temp = -0.5 * (b + (std::signbit(b) ? -root : root));
        double x0p = temp / a, x1p = c / temp;
        double x0m = min_(x0p, x1p), x1m = max_(x0p, x1p);
        x0 = x0m; x1 = x1m;
    }
    return valid_linear || valid_quadratic;
}

/* Sphere::ray_intersect_preliminary (src/shapes/sphere.cpp:272-306): double precision */
template <typename F> inline F sphere_intersect(const OShape<F> &sh, const Ray<F> &ray) {
    double mint = (double) ray.mint, maxt = (double) ray.maxt;
    V3<double> o = { (double) ray.o.x - (double) sh.center.x, (double) ray.o.y - (double) sh.center.y,
                     (double) ray.o.z - (double) sh.center.z };
    V3<double> d = { (double) ray.d.x, (double) ray.d.y, (double) ray.d.z };
    double A = dot(d, d), B = 2.0 * dot(o, d), C = dot(o, o) - sqr((double) sh.radius);
    double near_t, far_t;
    bool found = solve_quadratic(A, B, C, near_t, far_t);
    bool out_bounds = !(near_t <= maxt && far_t >= mint);
    bool in_bounds = near_t < mint && far_t > maxt;
    bool active = found && !out_bounds && !in_bounds;
    return active ? (near_t < mint ? (F) far_t : (F) near_t) : Consts<F>::Inf;
}

template <typename F> inline V3<F> vtx(const OScene<F> &s, const OShape<F> &sh, uint32_t i) {
    const float *p = s.vpos + 3 * (size_t) (sh.vofs + i);
    return { (F) p[0], (F) p[1], (F) p[2] };
}

/* Mesh::ray_intersect_triangle (include/mitsuba/render/mesh.h:195-226) */
template <typename F>
inline F triangle_intersect(const OScene<F> &s, const OShape<F> &sh, uint32_t f, const Ray<F> &ray,
                            F &u_out, F &v_out) {
    const uint32_t *fi = s.faces + 3 * (size_t) (sh.fofs + f);
    V3<F> p0 = vtx(s, sh, fi[0]), p1 = vtx(s, sh, fi[1]), p2 = vtx(s, sh, fi[2]);
    V3<F> e1 = p1 - p0, e2 = p2 - p0;
    V3<F> pvec = cross(ray.d, e2);
    F inv_det = rcp_(dot(e1, pvec));
    V3<F> tvec = ray.o - p0;
    F u = dot(tvec, pvec) * inv_det;
    bool active = u >= 0 && u <= (F) 1;
    V3<F> qvec = cross(tvec, e1);
    F v = dot(ray.d, qvec) * inv_det;
    active = active && v >= 0 && u + v <= (F) 1;
    F t = dot(e2, qvec) * inv_det;
    active = active && t >= ray.mint && t <= ray.maxt;
    u_out = u; v_out = v;
    return active ? t : Consts<F>::Inf;
}

/* SurfaceInteraction::initialize_sh_frame (interaction.h:153-156) */
template <typename F> inline void initialize_sh_frame(SurfaceInteraction<F> &si) {
    F d = dot(si.sh_frame.n, si.dp_du);
    V3<F> v = { fma_(-si.sh_frame.n.x, d, si.dp_du.x), fma_(-si.sh_frame.n.y, d, si.dp_du.y),
                fma_(-si.sh_frame.n.z, d, si.dp_du.z) };
    si.sh_frame.s = normalize(v);
    si.sh_frame.t = cross(si.sh_frame.n, si.sh_frame.s);
}

/* Scene::ray_intersect: closest hit over all primitives.  The reference walks a
   kd-tree (kdtree.h:2078-2171); its result is the closest hit with
   ray.mint <= t <= ray.maxt after clipping against the padded scene box, a later
   primitive winning exact ties (`t <= maxt`).  The oracle tests primitives in
   shape order, which has the same semantics up to the order of exact ties.   */
template <typename F>
SurfaceInteraction<F> ray_intersect(const OScene<F> &s, const Ray<F> &ray_in, Counters &cnt) {
    cnt.n_intersect++;
    Ray<F> ray = ray_in;
    SurfaceInteraction<F> si;
    si.wi = -ray.d; /* scene_native.inl:36-38 */
    F bmin, bmax;
    bbox_ray_intersect(s.kd_bbox, ray, bmin, bmax);
    F mint = max_(ray.mint, bmin), maxt = min_(ray.maxt, bmax);
    if (!(mint <= maxt))
        return si;
    int best_shape = -1; uint32_t best_prim = 0; F best_u = 0, best_v = 0;
    for (size_t i = 0; i < s.shapes.size(); ++i) {
        const OShape<F> &sh = s.shapes[i];
        if (sh.sample_only) continue;   /* a sensor's own target shape is not a child of the scene (scene.cpp:23-40) */
        if (sh.type == B2VP_SHAPE_RECTANGLE) {
            F t = rectangle_intersect(sh, ray);
            if (t != Consts<F>::Inf) { best_shape = (int) i; ray.maxt = t; }
        } else if (sh.type == B2VP_SHAPE_SPHERE) {
            F t = sphere_intersect(sh, ray);
            if (t != Consts<F>::Inf) { best_shape = (int) i; ray.maxt = t; }
        } else {
            for (uint32_t f = 0; f < sh.fcount; ++f) {
                F u, v;
                F t = triangle_intersect(s, sh, f, ray, u, v);
                if (t != Consts<F>::Inf) { best_shape = (int) i; best_prim = f; best_u = u; best_v = v; ray.maxt = t; }
            }
        }
    }
    if (best_shape < 0)
        return si;

    /* PreliminaryIntersection::compute_surface_interaction (interaction.h:571-596) */
    const OShape<F> &sh = s.shapes[best_shape];
    F t = ray.maxt;
    si.t = t; si.shape = best_shape; si.prim = best_prim;
    if (sh.type == B2VP_SHAPE_RECTANGLE) { /* rectangle.cpp:175-212 */
        V3<F> p = ray_in.at(t);
        F dist = dot(sh.translation - p, sh.frame.n);
        si.p = fmadd3(sh.frame.n, dist, p);
        si.n = sh.frame.n; si.sh_frame.n = sh.frame.n;
        si.dp_du = sh.frame.s;
    } else if (sh.type == B2VP_SHAPE_SPHERE) { /* sphere.cpp:335-400 */
        si.sh_frame.n = normalize(ray_in.at(t) - sh.center);
        si.p = fmadd3(si.sh_frame.n, sh.radius, sh.center);
        V3<F> local = xf_point_affine(sh.to_object, si.p);
        si.dp_du = { -local.y, local.x, (F) 0 };
        si.dp_du = xf_vector(sh.to_world, si.dp_du) * ((F) 2 * Consts<F>::Pi);
        if (sh.flip_normals) si.sh_frame.n = -si.sh_frame.n;
        si.n = si.sh_frame.n;
    } else { /* mesh.cpp:450-545 */
        F b1 = best_u, b2 = best_v, b0 = (F) 1 - b1 - b2;
        const uint32_t *fi = s.faces + 3 * (size_t) (sh.fofs + best_prim);
        V3<F> p0 = vtx(s, sh, fi[0]), p1 = vtx(s, sh, fi[1]), p2 = vtx(s, sh, fi[2]);
        V3<F> dp0 = p1 - p0, dp1 = p2 - p0;
        si.p = p0 * b0 + p1 * b1 + p2 * b2;
        si.n = normalize(cross(dp0, dp1));
        V3<F> dp_dv;
        coordinate_system(si.n, si.dp_du, dp_dv);
        if (sh.has_uv) {
            const float *t0 = s.vuv + 2 * (size_t) (sh.vofs + fi[0]), *t1 = s.vuv + 2 * (size_t) (sh.vofs + fi[1]),
                        *t2 = s.vuv + 2 * (size_t) (sh.vofs + fi[2]);
            F du0 = (F) t1[0] - (F) t0[0], dv0 = (F) t1[1] - (F) t0[1];
            F du1 = (F) t2[0] - (F) t0[0], dv1 = (F) t2[1] - (F) t0[1];
            F det = fma_(du0, dv1, -(dv0 * du1)), inv_det = rcp_(det);
            if (det != 0) {
                /* fmsub(duv1.y, dp0, duv0.y * dp1) * inv_det */
                si.dp_du = { fma_(dv1, dp0.x, -(dv0 * dp1.x)) * inv_det,
                             fma_(dv1, dp0.y, -(dv0 * dp1.y)) * inv_det,
                             fma_(dv1, dp0.z, -(dv0 * dp1.z)) * inv_det };
            }
        }
        if (sh.has_normals) {
            auto nrm = [&](uint32_t i) {
                const float *q = s.vnrm + 3 * (size_t) (sh.vofs + i);
                return V3<F>{ (F) q[0], (F) q[1], (F) q[2] };
            };
            V3<F> n0 = nrm(fi[0]), n1 = nrm(fi[1]), n2 = nrm(fi[2]);
            si.sh_frame.n = normalize(n0 * b0 + n1 * b1 + n2 * b2);
        } else {
            si.sh_frame.n = si.n;
        }
    }
    initialize_sh_frame(si);
    si.wi = si.sh_frame.to_local(-ray_in.d);
    return si;
}

/* Interaction::spawn_ray (interaction.h:58-61) */
template <typename F> inline Ray<F> spawn_ray(V3<F> p, V3<F> d) {
    return make_ray(p, d, ((F) 1 + hmax_abs(p)) * Consts<F>::RayEpsilon, Consts<F>::Inf);
}

/* ======================================================================
 *  Volumes and media
 * ====================================================================== */
inline int floor2int(float x) { return (int) std::floor(x); }
inline int floor2int(double x) { return (int) std::floor(x); }

/* GridVolumeImpl::wrap (src/textures/grid3d.cpp:234-250) for one coordinate */
inline int wrap_coord(int v, int res, int mode) {
    if (mode == B2VP_WRAP_CLAMP)
        return std::min(std::max(v, 0), res - 1);
    /* enoki::divisor<int32_t> == truncating integer division */
    int div = v / res, mod = v - div * res;
    if (mod < 0) mod += res;
    if (mode == B2VP_WRAP_MIRROR) {
        bool keep = ((div & 1) == 0) != (v < 0);
        mod = keep ? mod : res - 1 - mod;
    }
    return mod;
}

template <typename F> inline F luminance3(F r, F g, F b) { /* mitsuba::luminance(Color3f) */
    return r * (F) 0.212671f + g * (F) 0.715160f + b * (F) 0.072169f; /* float literals, spectrum.h:247 */
}

/* GridVolumeImpl::eval (grid3d.cpp:163-185,220-360) / ConstVolume::eval */
template <typename F> F volume_eval(const OVolume<F> &v, V3<F> p_world) {
    if (v.type == B2VP_VOLUME_CONST)
        return v.value;
    V3<F> p = xf_point(v.world_to_local, p_world); /* grid3d.cpp:227: full 4x4 with w divide */
    const int nx = v.nx, ny = v.ny, nz = v.nz, ch = v.channels;
    auto fetch = [&](int x, int y, int z, int c) -> F {
        size_t idx = ((size_t) z * ny + y) * nx + x;
        return v.data[idx * ch + c];
    };
    F res[3];
    if (v.filter == B2VP_FILTER_TRILINEAR) {
        p = { fma_(p.x, (F) nx, (F) -0.5), fma_(p.y, (F) ny, (F) -0.5), fma_(p.z, (F) nz, (F) -0.5) };
        int ix = floor2int(p.x), iy = floor2int(p.y), iz = floor2int(p.z);
        F w1x = p.x - (F) ix, w1y = p.y - (F) iy, w1z = p.z - (F) iz;
        F w0x = (F) 1 - w1x, w0y = (F) 1 - w1y, w0z = (F) 1 - w1z;
        int x0 = wrap_coord(ix, nx, v.wrap), x1 = wrap_coord(ix + 1, nx, v.wrap);
        int y0 = wrap_coord(iy, ny, v.wrap), y1 = wrap_coord(iy + 1, ny, v.wrap);
        int z0 = wrap_coord(iz, nz, v.wrap), z1 = wrap_coord(iz + 1, nz, v.wrap);
        for (int c = 0; c < ch; ++c) {
            F d000 = fetch(x0, y0, z0, c), d100 = fetch(x1, y0, z0, c), d010 = fetch(x0, y1, z0, c),
              d110 = fetch(x1, y1, z0, c), d001 = fetch(x0, y0, z1, c), d101 = fetch(x1, y0, z1, c),
              d011 = fetch(x0, y1, z1, c), d111 = fetch(x1, y1, z1, c);
            F v00 = fma_(w0x, d000, w1x * d100), v01 = fma_(w0x, d001, w1x * d101),
              v10 = fma_(w0x, d010, w1x * d110), v11 = fma_(w0x, d011, w1x * d111);
            F v0 = fma_(w0y, v00, w1y * v10), v1 = fma_(w0y, v01, w1y * v11);
            res[c] = fma_(w0z, v0, w1z * v1);
        }
    } else {
        p = { p.x * (F) nx, p.y * (F) ny, p.z * (F) nz };
        int x = wrap_coord(floor2int(p.x), nx, v.wrap), y = wrap_coord(floor2int(p.y), ny, v.wrap),
            z = wrap_coord(floor2int(p.z), nz, v.wrap);
        for (int c = 0; c < ch; ++c) res[c] = fetch(x, y, z, c);
    }
    return ch == 1 ? res[0] : luminance3(res[0], res[1], res[2]);
}

template <typename F> struct MediumInteraction {
    F t = Consts<F>::Inf, mint = 0;
    V3<F> p{ 0, 0, 0 }, wi{ 0, 0, 0 };
    Frame<F> sh_frame{ { 0, 0, 0 }, { 0, 0, 0 }, { 0, 0, 0 } };
    F sigma_s = 0, sigma_n = 0, sigma_t = 0, combined_extinction = 0;
    int medium = -1;
    bool is_valid() const { return t != Consts<F>::Inf; }
};

/* Medium::sample_interaction (src/librender/medium.cpp:36-75) */
template <typename F>
MediumInteraction<F> sample_interaction(const OScene<F> &s, int medium_id, const Ray<F> &ray, F sample,
                                        Counters &cnt) {
    cnt.medium_steps++;
    const OMedium<F> &m = s.media[medium_id];
    MediumInteraction<F> mi;
    mi.sh_frame = make_frame(ray.d);
    mi.wi = -ray.d;
    bool aabb_its; F mint, maxt;
    if (m.type == B2VP_MEDIUM_HOMOGENEOUS) { /* homogeneous.cpp:51-54 */
        aabb_its = true; mint = 0; maxt = Consts<F>::Inf;
    } else {                                  /* heterogeneous.cpp:50-53 */
        aabb_its = bbox_ray_intersect(m.aabb, ray, mint, maxt);
    }
    aabb_its = aabb_its && (std::isfinite(mint) || std::isfinite(maxt));
    bool active = aabb_its;
    if (!active) { mint = 0; maxt = Consts<F>::Inf; }
    mint = max_(ray.mint, mint);
    maxt = min_(ray.maxt, maxt);
    F combined;
    if (m.type == B2VP_MEDIUM_HOMOGENEOUS)
        combined = volume_eval(s.volumes[m.sigma_t], ray.o) * m.scale; /* homogeneous.cpp:30-38 (const volume) */
    else
        combined = m.max_density;
    F sampled_t = mint + (-log_((F) 1 - sample) / combined);
    bool valid_mi = active && (sampled_t <= maxt);
    mi.t = valid_mi ? sampled_t : Consts<F>::Inf;
    mi.p = ray.at(sampled_t);
    mi.medium = medium_id;
    mi.mint = mint;
    if (valid_mi) { /* coefficients of an invalid interaction are never used by volpath */
        F sigmat, albedo = volume_eval(s.volumes[m.albedo], mi.p);
        if (m.type == B2VP_MEDIUM_HOMOGENEOUS) {
            sigmat = volume_eval(s.volumes[m.sigma_t], mi.p) * m.scale;
            mi.sigma_n = 0;
        } else {
            sigmat = m.scale * volume_eval(s.volumes[m.sigma_t], mi.p);
            mi.sigma_n = combined - sigmat;
        }
        mi.sigma_t = sigmat;
        mi.sigma_s = sigmat * albedo;
    }
    mi.combined_extinction = combined;
    return mi;
}

/* Medium::eval_tr_and_pdf (medium.cpp:77-89) followed by the
   `tr / tr_pdf` pattern of volpath.cpp:113-117,402-404 */
template <typename F> inline F tr_over_pdf(const MediumInteraction<F> &mi, F si_t) {
    F t = min_(mi.t, si_t) - mi.mint;
    F tr = exp_(-t * mi.combined_extinction);
    F pdf = si_t < mi.t ? tr : tr * mi.combined_extinction;
    return pdf > 0 ? tr / pdf : (F) 0;
}

/* ======================================================================
 *  Warps, phase functions, BSDFs
 * ====================================================================== */
/* warp::square_to_uniform_disk_concentric (include/mitsuba/core/warp.h:54-90) */
template <typename F> inline void square_to_uniform_disk_concentric(F sx, F sy, F &ox, F &oy) {
    F x = fma_((F) 2, sx, (F) -1), y = fma_((F) 2, sy, (F) -1);
    bool is_zero = x == 0 && y == 0, quadrant_1_or_3 = abs_(x) < abs_(y);
    F r = quadrant_1_or_3 ? y : x, rp = quadrant_1_or_3 ? x : y;
    F phi = (F) 0.25 * Consts<F>::Pi * rp / r;
    if (quadrant_1_or_3) phi = (F) 0.5 * Consts<F>::Pi - phi;
    if (is_zero) phi = 0;
    auto sc = sincos_(phi);
    ox = r * sc.second; oy = r * sc.first;
}
/* warp::square_to_cosine_hemisphere (warp.h:325-334) */
template <typename F> inline V3<F> square_to_cosine_hemisphere(F sx, F sy) {
    F px, py; square_to_uniform_disk_concentric(sx, sy, px, py);
    F z = safe_sqrt_((F) 1 - (px * px + py * py));
    return { px, py, z };
}
/* warp::square_to_uniform_hemisphere (warp.h:298-303) */
template <typename F> inline V3<F> square_to_uniform_hemisphere(F sx, F sy) {
    F px, py; square_to_uniform_disk_concentric(sx, sy, px, py);
    F z = (F) 1 - (px * px + py * py);
    F k = sqrt_(z + (F) 1);
    return { px * k, py * k, z };
}
/* warp::square_to_uniform_sphere (warp.h:255-260) */
template <typename F> inline V3<F> square_to_uniform_sphere(F sx, F sy) {
    F z = fma_((F) -2, sy, (F) 1);
    F r = safe_sqrt_(fma_(-z, z, (F) 1)); /* circ(z) = sqrt(fnmadd(z, z, 1)) */
    auto sc = sincos_((F) 2 * Consts<F>::Pi * sx);
    return { r * sc.second, r * sc.first, z };
}

/* PhaseFunction::eval (rayleigh.cpp:42-45, hg.cpp:47-50, isotropic.cpp) */
template <typename F> inline F phase_eval_cos(const B2vpPhase &ph, const TabDistr<F> &tab, F cos_theta) {
    if (ph.type == B2VP_PHASE_TABULATED) /* tabphase.cpp:91-99: eval_pdf_normalized(-cos_theta) / (2 pi) */
        return tab.eval_pdf(-cos_theta) * tab.normalization * Consts<F>::InvTwoPi;
    if (ph.type == B2VP_PHASE_RAYLEIGH)
        return ((F) 3 / (F) 16) * Consts<F>::InvPi * ((F) 1 + sqr(cos_theta));
    if (ph.type == B2VP_PHASE_HG) {
        F g = (F) ph.g;
        F temp = (F) 1 + g * g + (F) 2 * g * cos_theta;
        return Consts<F>::InvFourPi * ((F) 1 - g * g) / (temp * sqrt_(temp));
    }
    return Consts<F>::InvFourPi;
}
template <typename F>
inline F phase_eval(const B2vpPhase &ph, const TabDistr<F> &tab, const MediumInteraction<F> &mi, V3<F> wo) {
    if (ph.type == B2VP_PHASE_ISOTROPIC) return Consts<F>::InvFourPi;
    return phase_eval_cos(ph, tab, dot(wo, mi.wi));
}
/* PhaseFunction::sample (rayleigh.cpp:47-66, hg.cpp:52-76, isotropic.cpp:31-39) */
template <typename F>
inline V3<F> phase_sample(const B2vpPhase &ph, const TabDistr<F> &tab, const MediumInteraction<F> &mi, F s2x, F s2y,
                          F *pdf = nullptr) {
    if (ph.type == B2VP_PHASE_ISOTROPIC) {
        if (pdf) *pdf = Consts<F>::InvFourPi;   /* square_to_uniform_sphere_pdf (isotropic.cpp:37) */
        return square_to_uniform_sphere(s2x, s2y);
    }
    F cos_theta;
    if (ph.type == B2VP_PHASE_TABULATED) { /* tabphase.cpp:68-89 */
        cos_theta = tab.sample(s2x);
    } else if (ph.type == B2VP_PHASE_RAYLEIGH) {
        F z = (F) 2 * ((F) 2 * s2x - (F) 1);
        F tmp = sqrt_(sqr(z) + (F) 1);
        F A = cbrt_(z + tmp), B = cbrt_(z - tmp);
        cos_theta = A + B;
    } else {
        F g = (F) ph.g;
        if (abs_(g) < Consts<F>::Epsilon) {
            cos_theta = (F) 1 - (F) 2 * s2x;
        } else {
            F sqr_term = ((F) 1 - g * g) / ((F) 1 - g + (F) 2 * g * s2x);
            cos_theta = ((F) 1 + g * g - sqr_term * sqr_term) / ((F) 2 * g);
        }
    }
    F sin_theta = safe_sqrt_((F) 1 - cos_theta * cos_theta);
    auto sc = sincos_((F) 2 * Consts<F>::Pi * s2y);
    V3<F> wo = { sin_theta * sc.second, sin_theta * sc.first, cos_theta };
    /* the pdf returned with the sample is the phase function at the LOCAL sampled cosine, eval_*(-cos_theta)
       (rayleigh.cpp:63, hg.cpp:73, tabphase.cpp:79-80) -- not eval(wo): with a ray direction that is not exactly of
       unit length (a sensor frame rounded to single precision) dot(to_world(wo), wi) is another number */
    /* tabphase: sample() returns m_distr.eval_pdf_normalized(-cos_theta) (tabphase.cpp:79) where eval() of the same
       direction, whose dot(wo, wi) is -cos_theta, looks up eval_pdf_normalized(+cos_theta) (tabphase.cpp:94-96): the
       reference's sampled pdf is the table read at the mirrored cosine.  Mirrored here, not corrected */
    if (pdf) *pdf = phase_eval_cos(ph, tab, ph.type == B2VP_PHASE_TABULATED ? cos_theta : -cos_theta);
    return mi.sh_frame.to_world(wo);
}

/* BlendPhaseFunction (src/phase/blendphase.cpp:68-140, default context): the weight volume is evaluated at the
   scattering position and clamped to [0, 1] */
template <typename F> inline F blend_weight(const OScene<F> &s, const B2vpPhase &ph, const MediumInteraction<F> &mi) {
    return min_(max_(volume_eval(s.volumes[ph.weight_volume], mi.p), (F) 0), (F) 1);
}
template <typename F> inline F phase_eval_at(const OScene<F> &s, int idx, const MediumInteraction<F> &mi, V3<F> wo) {
    const B2vpPhase &ph = s.phases[idx];
    if (ph.type != B2VP_PHASE_BLEND) return phase_eval(ph, s.tabs[idx], mi, wo);
    F weight = blend_weight(s, ph, mi);
    return phase_eval(s.phases[ph.child[0]], s.tabs[ph.child[0]], mi, wo) * ((F) 1 - weight) +
           phase_eval(s.phases[ph.child[1]], s.tabs[ph.child[1]], mi, wo) * weight;
}
template <typename F>
inline V3<F> phase_sample_at(const OScene<F> &s, int idx, const MediumInteraction<F> &mi, F s1, F s2x, F s2y,
                             F *pdf = nullptr) {
    const B2vpPhase &ph = s.phases[idx];
    if (ph.type != B2VP_PHASE_BLEND) return phase_sample(ph, s.tabs[idx], mi, s2x, s2y, pdf);
    F weight = blend_weight(s, ph, mi);
    /* the rescaled sample1 is passed on, none of the nested phase functions in scope uses it; the pdf is that of the
       nested function that was picked, unweighted (blendphase.cpp:97-112) */
    int c = s1 > weight ? ph.child[0] : ph.child[1];
    return phase_sample(s.phases[c], s.tabs[c], mi, s2x, s2y, pdf);
}

enum : uint32_t { /* BSDFFlags (include/mitsuba/render/bsdf.h:44-123) */
    F_Null = 0x1, F_DiffuseReflection = 0x2, F_DiffuseTransmission = 0x4, F_GlossyReflection = 0x8,
    F_Smooth = 0x2 | 0x4 | 0x8 | 0x10, F_Delta = 0x1 | 0x20 | 0x40
};
template <typename F> struct BSDFSample { V3<F> wo{ 0, 0, 0 }; F pdf = 0, eta = 0; uint32_t sampled_type = 0; };

inline uint32_t bsdf_flags(const B2vpBsdf &b) {
    if (b.type == B2VP_BSDF_BILAMBERTIAN) return F_DiffuseReflection | F_DiffuseTransmission; /* bilambertian.cpp:55-60 */
    return b.type == B2VP_BSDF_NULL ? F_Null : (b.type == B2VP_BSDF_DIFFUSE ? F_DiffuseReflection : F_GlossyReflection);
}
/* BiLambertian lobe weights (src/bsdfs/bilambertian.cpp:80-87): r / (r + t), 1 - that, NaN (r = t = 0) -> 0 */
template <typename F> inline void bilambertian_weights(const B2vpBsdf &b, F &r, F &t, F &rw, F &tw) {
    r = (F) b.p[0]; t = (F) b.p[1];
    rw = r / (r + t); tw = (F) 1 - rw;
    if (rw != rw) rw = 0;
    if (tw != tw) tw = 0;
}

/* Frame::sincos_phi (frame.h:107-118) */
template <typename F> inline void sincos_phi(V3<F> v, F &s, F &c) {
    F sin_theta_2 = fma_(v.x, v.x, sqr(v.y));
    F inv_sin_theta = rsqrt_(sin_theta_2);
    F rx = v.x * inv_sin_theta, ry = v.y * inv_sin_theta;
    if (abs_(sin_theta_2) <= (F) 4 * Consts<F>::Epsilon) { rx = 1; ry = 0; }
    else { rx = min_(max_(rx, (F) -1), (F) 1); ry = min_(max_(ry, (F) -1), (F) 1); }
    s = ry; c = rx;
}
template <typename F> inline F frame_sin_theta(V3<F> v) { return safe_sqrt_(fma_(v.x, v.x, sqr(v.y))); }
template <typename F> inline F frame_tan_theta(V3<F> v) { return safe_sqrt_(fma_(-v.z, v.z, (F) 1)) / v.z; }

/* RPV::eval_rpv (src/bsdfs/rpv.cpp:107-141); pow is enoki's exp(log(x)*y) */
template <typename F> F eval_rpv(const B2vpBsdf &b, V3<F> wi, V3<F> wo) {
    F rho_0 = (F) b.p[0], k = (F) b.p[1], g = (F) b.p[2], rho_c = (F) b.p[3];
    F sin_phi1, cos_phi1, sin_phi2, cos_phi2;
    sincos_phi(wi, sin_phi1, cos_phi1);
    sincos_phi(wo, sin_phi2, cos_phi2);
    F cos_phi1_minus_phi2 = cos_phi1 * cos_phi2 + sin_phi1 * sin_phi2;
    F sin_theta1 = frame_sin_theta(wi), cos_theta1 = wi.z, tan_theta1 = frame_tan_theta(wi);
    F sin_theta2 = frame_sin_theta(wo), cos_theta2 = wo.z, tan_theta2 = frame_tan_theta(wo);
    F G = safe_sqrt_(sqr(tan_theta1) + sqr(tan_theta2) - (F) 2 * tan_theta1 * tan_theta2 * cos_phi1_minus_phi2);
    F cos_g = cos_theta1 * cos_theta2 + sin_theta1 * sin_theta2 * cos_phi1_minus_phi2;
    F Fg = ((F) 1 - sqr(g)) / pow_(((F) 1 + sqr(g) + (F) 2 * g * cos_g), (F) 1.5);
    F value = rho_0 * (pow_(cos_theta1 * cos_theta2 * (cos_theta1 + cos_theta2), k - (F) 1) * Fg *
                       ((F) 1 + ((F) 1 - rho_c) / ((F) 1 + G))) * Consts<F>::InvPi;
    return value;
}

/* BSDF::sample (diffuse.cpp:78-102, rpv.cpp:85-105, null.cpp:41-57) */
template <typename F> BSDFSample<F> bsdf_sample(const B2vpBsdf &b, V3<F> wi, F s1, F s2x, F s2y, F &weight) {
    BSDFSample<F> bs;
    weight = 0;
    if (b.type == B2VP_BSDF_NULL) {
        bs.wo = -wi; bs.sampled_type = F_Null; bs.eta = 1; bs.pdf = 1; weight = 1;
        return bs;
    }
    if (b.type == B2VP_BSDF_BILAMBERTIAN) { /* bilambertian.cpp:63-119 */
        V3<F> wo = square_to_cosine_hemisphere(s2x, s2y);
        F r, t, rw, tw; bilambertian_weights(b, r, t, rw, tw);
        bool selected_r = s1 < rw;
        F value = 1;
        if (selected_r) value *= r / rw; else value *= t / tw;
        bs.pdf = Consts<F>::InvPi * wo.z;
        bs.pdf = selected_r ? bs.pdf * rw : bs.pdf * tw;
        bs.eta = 1;
        bs.sampled_type = selected_r ? F_DiffuseReflection : F_DiffuseTransmission;
        if (!(wi.z > 0)) wo.z = -wo.z;   /* the incoming direction comes from "behind" */
        if (!selected_r) wo.z = -wo.z;   /* transmission */
        bs.wo = wo;
        weight = bs.pdf > 0 ? value : (F) 0;
        return bs;
    }
    F cos_theta_i = wi.z;
    bool active = cos_theta_i > 0;
    if (b.type == B2VP_BSDF_DIFFUSE) {
        if (!active) return bs; /* zero-initialised sample, weight 0 (diffuse.cpp:90-92) */
        bs.wo = square_to_cosine_hemisphere(s2x, s2y);
        bs.pdf = Consts<F>::InvPi * bs.wo.z;
        bs.eta = 1; bs.sampled_type = F_DiffuseReflection;
        weight = (active && bs.pdf > 0) ? (F) b.p[0] : (F) 0;
        return bs;
    }
    /* RPV: the sample is always generated; the weight is the BRDF value f
       (NOT f*cos/pdf) -- reference behaviour, SURVEY.md F8 */
    bs.wo = square_to_cosine_hemisphere(s2x, s2y);
    bs.pdf = Consts<F>::InvPi * bs.wo.z;
    bs.eta = 1; bs.sampled_type = F_GlossyReflection;
    F value = eval_rpv(b, wi, bs.wo);
    weight = (active && bs.pdf > 0) ? value : (F) 0;
    return bs;
}
/* BSDF::eval (diffuse.cpp:104-120, rpv.cpp:143-155, null.cpp:59-62) */
template <typename F> F bsdf_eval(const B2vpBsdf &b, V3<F> wi, V3<F> wo) {
    if (b.type == B2VP_BSDF_NULL) return 0;
    if (b.type == B2VP_BSDF_BILAMBERTIAN) { /* bilambertian.cpp:121-151; sign() is copysign(1, x) */
        F result = signbit_(wi.z) == signbit_(wo.z) ? (F) b.p[0] : (F) b.p[1];
        return result * (Consts<F>::InvPi * abs_(wo.z));
    }
    bool active = wi.z > 0 && wo.z > 0;
    if (b.type == B2VP_BSDF_DIFFUSE) {
        F value = (F) b.p[0] * Consts<F>::InvPi * wo.z;
        return active ? value : (F) 0;
    }
    F value = eval_rpv(b, wi, wo);
    return active ? value * abs_(wo.z) : (F) 0;
}
/* BSDF::pdf (diffuse.cpp:122-135, rpv.cpp:157-168) */
template <typename F> F bsdf_pdf(const B2vpBsdf &b, V3<F> wi, V3<F> wo) {
    if (b.type == B2VP_BSDF_NULL) return 0;
    if (b.type == B2VP_BSDF_BILAMBERTIAN) { /* bilambertian.cpp:153-193 */
        F r, t, rw, tw; bilambertian_weights(b, r, t, rw, tw);
        F result = Consts<F>::InvPi * abs_(wo.z);
        return result * (signbit_(wi.z) == signbit_(wo.z) ? rw : tw);
    }
    F pdf = Consts<F>::InvPi * wo.z;
    return (wi.z > 0 && wo.z > 0) ? pdf : (F) 0;
}
template <typename F> inline F bsdf_eval_null_transmission(const B2vpBsdf &b) {
    return b.type == B2VP_BSDF_NULL ? (F) 1 : (F) 0; /* null.cpp:69-72, bsdf.cpp:11-14 */
}

/* ======================================================================
 *  Emitter and sensor
 * ====================================================================== */
template <typename F> struct DirectionSample { V3<F> d{ 0, 0, 0 }; F dist = 0, pdf = 0; bool delta = false; };

/* Scene::sample_emitter_direction (scene.cpp:168-218) over
   DirectionalEmitter::sample_direction (directional.cpp:108-135) */
template <typename F>
DirectionSample<F> sample_emitter_direction(const OScene<F> &s, F sx, F sy, F &spec) {
    DirectionSample<F> ds;
    spec = 0;
    size_t n = s.emitters.size();
    if (n == 0) return ds;
    size_t index = 0;
    F emitter_pdf = 1;
    if (n > 1) {
        emitter_pdf = (F) (1.f / (float) n);   /* `ScalarFloat emitter_pdf = 1.f / m_emitters.size()` (scene.cpp:183): a float quotient in every variant */
        index = std::min((size_t) (uint32_t) (sx * (F) n), n - 1);
        sx = (sx - (F) index * emitter_pdf) * (F) n; /* rescale sample.x() to [0, 1) again (scene.cpp:190) */
    }
    const B2vpEmitter &e = s.emitters[index];
    if (e.type == B2VP_EMITTER_CONSTANT) { /* ConstantBackgroundEmitter::sample_direction (constant.cpp:81-104) */
        ds.d = square_to_uniform_sphere(sx, sy);
        ds.dist = (F) 2 * s.bs_radius;
        ds.pdf = Consts<F>::InvFourPi;  /* square_to_uniform_sphere_pdf (warp.h:272-279, no domain test) */
        ds.delta = false;
        spec = (F) e.irradiance / ds.pdf;
    } else {
        V3<F> d = ld3<F>(e.direction);
        ds.d = -d; ds.dist = (F) 2 * s.bs_radius; ds.pdf = 1; ds.delta = true;
        spec = (F) e.irradiance;
    }
    if (n > 1) { ds.pdf *= emitter_pdf; spec *= rcp_(emitter_pdf); }
    return ds;
}

template <typename F> inline F rect_area(const OShape<F> &sh) { return norm(cross(sh.frame.s, sh.frame.t)); }

/* Sensor::sample_ray (distant.cpp:299-386, mdistant.cpp:216-266,
   radiancemeter.cpp:106-139, mradiancemeter.cpp:134-159); returns the weight */
/* The ray origin on a `ray_origin` / `origin` shape (distant.cpp:366-375, distantflux.cpp:245-252): the target is
   projected onto the shape along -d with Shape::ray_intersect(.., Minimal) -- for a rectangle ray_intersect_preliminary
   (rectangle.cpp:139-159) and the re-projected hit point of compute_surface_interaction (rectangle.cpp:194-197).  The
   temporary ray is Ray3f(o, d, time): mint = RayEpsilon, maxt = inf (ray.h:33-34,44-47).  A miss leaves pi.t = inf, so
   the origin is not finite, and the caller zeroes the ray weight (`ray_weight && active`). */
template <typename F>
V3<F> origin_on_shape(const OShape<F> &sh, const V3<F> &target, const V3<F> &d, bool &valid) {
    Ray<F> tmp = make_ray(target, -d, Consts<F>::RayEpsilon, Consts<F>::Inf);
    const F t = rectangle_intersect(sh, tmp);
    valid = t != Consts<F>::Inf;
    V3<F> p = tmp.at(t);
    F dist = dot(sh.translation - p, sh.frame.n);
    return fmadd3(sh.frame.n, dist, p);
}

template <typename F>
F sensor_sample_ray(const OScene<F> &s, F fx, F fy, F ax, F ay, Ray<F> &ray, bool &ok) {
    const B2vpSensor &sn = s.sensor;
    ok = true;
    ray = Ray<F>();
    F weight = 0;
    if (sn.type == B2VP_SENSOR_DISTANT) {
        const M4<F> &trafo = s.sensor_to_world;
        V3<F> v0 = { 0, 0, 1 };
        if (sn.direction_type == B2VP_DIR_SAMPLE_ALL)
            v0 = square_to_uniform_hemisphere(fx, fy);
        else if (sn.direction_type == B2VP_DIR_SAMPLE_WIDTH) {
            auto sc = sincos_(Consts<F>::Pi * fx);
            v0.x = sc.second; v0.z = sc.first;
        }
        ray.d = sn.flip_directions ? xf_vector(trafo, v0) : xf_vector(trafo, -v0);
        V3<F> ray_target = ld3<F>(sn.target_point);
        if (sn.target_type == B2VP_TARGET_POINT) {
            weight = 1;
        } else if (sn.target_type == B2VP_TARGET_SHAPE) {
            const OShape<F> &sh = s.shapes[sn.target_shape]; /* rectangle.cpp:111-125 */
            ray_target = xf_point_affine(sh.to_world, V3<F>{ ax * (F) 2 - (F) 1, ay * (F) 2 - (F) 1, (F) 0 });
            weight = (F) 1 / sh.inv_surface_area / rect_area(sh);
        } else {
            F ox, oy; square_to_uniform_disk_concentric(ax, ay, ox, oy);
            V3<F> perp = xf_vector(trafo, V3<F>{ ox, oy, (F) 0 });
            ray_target = s.bs_center + perp * s.bs_radius;
            weight = (F) 1;
            weight /= dot(-ray.d, V3<F>{ 0, 0, 1 });
        }
        if (sn.origin_type == B2VP_ORIGIN_SHAPE) {
            bool valid;
            ray.o = origin_on_shape(s.shapes[sn.origin_shape], ray_target, ray.d, valid);
            if (!valid) weight = 0;
        } else if (sn.target_type == B2VP_TARGET_NONE)
            ray.o = ray_target - ray.d * s.bs_radius;
        else
            ray.o = ray_target - ray.d * (F) 2 * s.bs_radius;
    } else if (sn.type == B2VP_SENSOR_MDISTANT) {
        int idx = (int) (fx * (F) sn.n_transforms);
        const M4<F> &trafo = s.sensor_xf[idx];
        ray.d = xf_vector(trafo, V3<F>{ 0, 0, 1 });
        if (sn.target_type == B2VP_TARGET_POINT) {
            ray.o = ld3<F>(sn.target_point) - (F) 2 * ray.d * s.bs_radius;
            weight = 1;
        } else if (sn.target_type == B2VP_TARGET_SHAPE) {
            const OShape<F> &sh = s.shapes[sn.target_shape];
            V3<F> p = xf_point_affine(sh.to_world, V3<F>{ ax * (F) 2 - (F) 1, ay * (F) 2 - (F) 1, (F) 0 });
            ray.o = p - (F) 2 * ray.d * s.bs_radius;
            weight = (F) 1 / (sh.inv_surface_area * rect_area(sh));
        } else {
            F ox, oy; square_to_uniform_disk_concentric(ax, ay, ox, oy);
            V3<F> perp = xf_vector(trafo, V3<F>{ ox, oy, (F) 0 });
            ray.o = s.bs_center + perp * s.bs_radius - ray.d * s.bs_radius;
            weight = 1;
        }
    } else if (sn.type == B2VP_SENSOR_DISTANTFLUX) { /* distantflux.cpp:184-251 */
        const M4<F> &trafo = s.sensor_to_world;
        ray.d = -xf_vector(trafo, square_to_uniform_hemisphere(fx, fy));
        const F npixels = (F) ((size_t) sn.film_size[0] * (size_t) sn.film_size[1]);
        /* square_to_uniform_hemisphere_pdf is the constant 1 / (2 pi) (warp.h:305-312, no domain test) */
        weight = dot(-ray.d, ld3<F>(sn.reference_normal)) / (Consts<F>::InvTwoPi * npixels);
        V3<F> ray_target = ld3<F>(sn.target_point);
        if (sn.target_type == B2VP_TARGET_SHAPE) {
            const OShape<F> &sh = s.shapes[sn.target_shape];
            ray_target = xf_point_affine(sh.to_world, V3<F>{ ax * (F) 2 - (F) 1, ay * (F) 2 - (F) 1, (F) 0 });
            weight *= (F) 1 / (sh.inv_surface_area * rect_area(sh));
        } else if (sn.target_type == B2VP_TARGET_NONE) {
            F ox, oy; square_to_uniform_disk_concentric(ax, ay, ox, oy);
            V3<F> perp = xf_vector(trafo, V3<F>{ ox, oy, (F) 0 });
            ray_target = s.bs_center + perp * s.bs_radius;
        }
        if (sn.origin_type == B2VP_ORIGIN_SHAPE) {
            bool valid;
            ray.o = origin_on_shape(s.shapes[sn.origin_shape], ray_target, ray.d, valid);
            if (!valid) weight = 0;
        } else
            ray.o = ray_target - ray.d * (F) 2 * s.bs_radius;
    } else if (sn.type == B2VP_SENSOR_IRRADIANCEMETER) { /* irradiancemeter.cpp:88-118 over rectangle.cpp:111-125 */
        const OShape<F> &sh = s.shapes[sn.origin_shape];
        ray.o = xf_point_affine(sh.to_world, V3<F>{ fx * (F) 2 - (F) 1, fy * (F) 2 - (F) 1, (F) 0 });
        V3<F> local = square_to_cosine_hemisphere(ax, ay);
        ray.d = make_frame(sh.frame.n).to_world(local);
        weight = Consts<F>::Pi;
    } else if (sn.type == B2VP_SENSOR_RADIANCEMETER) {
        ray.o = xf_point_affine(s.sensor_to_world, V3<F>{ 0, 0, 0 });
        ray.d = xf_vector(s.sensor_to_world, V3<F>{ 0, 0, 1 });
        weight = 1;
    } else {
        int idx = (int) (fx * (F) sn.n_transforms);
        const M4<F> &trafo = s.sensor_xf[idx];
        ray.o = xf_point_affine(trafo, V3<F>{ 0, 0, 0 });
        ray.d = xf_vector(trafo, V3<F>{ 0, 0, 1 });
        weight = 1;
    }
    ray.update();
    return weight;
}

/* ======================================================================
 *  VolumetricPathIntegrator (src/integrators/volpath.cpp)
 * ====================================================================== */
template <typename F> struct Integrator {
    const OScene<F> &scene;
    PCG32 &rng;
    Counters &cnt;
    uint32_t max_depth, rr_depth;
    bool hide_emitters;
    bool ltr; /* argument evaluation order of the compiler that built the reference */
    bool spectral_mis = true; /* volpathmis `use_spectral_mis` (volpathmis.cpp:36) */

    F next_1d(bool active) { return rng.next_float(F(), active); }

    int target_medium(const SurfaceInteraction<F> &si, V3<F> d) const { /* interaction.h:186-200 */
        const OShape<F> &sh = scene.shapes[si.shape];
        return dot(d, si.n) > 0 ? sh.exterior : sh.interior;
    }
    bool is_medium_transition(const SurfaceInteraction<F> &si) const { /* shape.h:341 */
        const OShape<F> &sh = scene.shapes[si.shape];
        return sh.interior >= 0 || sh.exterior >= 0;
    }
    F mis_weight(F pdf_a, F pdf_b) const { /* volpath.cpp:479-483 */
        pdf_a *= pdf_a; pdf_b *= pdf_b;
        return pdf_a > 0 ? pdf_a / (pdf_a + pdf_b) : (F) 0;
    }

    /* volpath.cpp:261-367 */
    F sample_emitter(V3<F> ref_p, bool is_medium_interaction, int medium, bool active, DirectionSample<F> &ds) {
        F transmittance = 1;
        F sx = next_1d(active), sy = next_1d(active);
        F emitter_val;
        ds = sample_emitter_direction(scene, sx, sy, emitter_val);
        if (ds.pdf == 0) emitter_val = 0;
        active = active && ds.pdf != 0;
        if (!active)
            return emitter_val;

        Ray<F> ray = spawn_ray(ref_p, ds.d);
        if (is_medium_interaction) ray.mint = 0;

        F total_dist = 0;
        SurfaceInteraction<F> si;
        bool needs_intersection = true;
        while (active) {
            F remaining_dist = ds.dist * ((F) 1 - Consts<F>::ShadowEpsilon) - total_dist;
            ray.maxt = remaining_dist;
            active = active && remaining_dist > 0;
            if (!active) break;

            bool escaped_medium = false;
            bool active_medium = active && medium >= 0;
            bool active_surface = active && !active_medium;

            if (active_medium) {
                const OMedium<F> &m = scene.media[medium];
                MediumInteraction<F> mi = sample_interaction(scene, medium, ray, next_1d(active_medium), cnt);
                if (m.type == B2VP_MEDIUM_HOMOGENEOUS && mi.is_valid())
                    ray.maxt = min_(mi.t, remaining_dist);
                if (needs_intersection)
                    si = ray_intersect(scene, ray, cnt);
                if (si.t < mi.t) mi.t = Consts<F>::Inf;
                needs_intersection = false;

                bool is_spectral = m.spectral, not_spectral = !m.spectral;
                if (is_spectral) {
                    F t = min_(remaining_dist, min_(mi.t, si.t)) - mi.mint;
                    F tr = exp_(-t * mi.combined_extinction);
                    F free_flight_pdf = (si.t < mi.t || mi.t > remaining_dist) ? tr : tr * mi.combined_extinction;
                    transmittance *= free_flight_pdf > 0 ? tr / free_flight_pdf : (F) 0;
                }
                if (mi.t > remaining_dist && mi.is_valid()) total_dist = ds.dist;
                if (mi.t > remaining_dist) mi.t = Consts<F>::Inf;

                escaped_medium = !mi.is_valid();
                active_medium = mi.is_valid();
                is_spectral = is_spectral && active_medium;
                not_spectral = not_spectral && active_medium;

                if (active_medium) {
                    total_dist += mi.t;
                    cnt.n_null++;
                    ray.o = mi.p; ray.mint = 0;
                    si.t = si.t - mi.t;
                    if (is_spectral) transmittance *= mi.sigma_n;
                    if (not_spectral) transmittance *= mi.sigma_n / mi.combined_extinction;
                }
            }

            bool intersect = active_surface && needs_intersection;
            if (intersect) si = ray_intersect(scene, ray, cnt);
            needs_intersection = needs_intersection && !intersect;
            active_surface = active_surface || escaped_medium;
            if (active_surface) total_dist += si.t;

            active_surface = active_surface && si.is_valid() && active && !active_medium;
            if (active_surface)
                transmittance *= bsdf_eval_null_transmission<F>(scene.bsdfs[scene.shapes[si.shape].bsdf]);

            if (active_surface) ray = spawn_ray(si.p, ray.d);
            ray.maxt = remaining_dist;
            needs_intersection = needs_intersection || active_surface;

            active = active && (active_medium || active_surface) && transmittance != 0;

            if (active_surface && is_medium_transition(si))
                medium = target_medium(si, ray.d);
        }
        return transmittance * emitter_val;
    }

    /* volpath.cpp:370-465.  With only delta/directional emitters nothing can be
       hit, so the value is always 0 -- but the march draws random numbers. */
    void evaluate_direct_light(int medium, Ray<F> ray, const SurfaceInteraction<F> &si_ray, bool active,
                               F &emitted, F &emitter_pdf) {
        F emitter_val = 0, transmittance = 1;
        emitter_pdf = 0;
        bool needs_intersection = false;
        SurfaceInteraction<F> si = si_ray;
        while (active) {
            bool escaped_medium = false;
            bool active_medium = active && medium >= 0;
            bool active_surface = active && !active_medium;
            if (active_medium) {
                const OMedium<F> &m = scene.media[medium];
                MediumInteraction<F> mi = sample_interaction(scene, medium, ray, next_1d(active_medium), cnt);
                if (m.type == B2VP_MEDIUM_HOMOGENEOUS && mi.is_valid()) ray.maxt = mi.t;
                if (needs_intersection) si = ray_intersect(scene, ray, cnt);
                if (si.t < mi.t) mi.t = Consts<F>::Inf;
                if (m.spectral) transmittance *= tr_over_pdf(mi, si.t);
                needs_intersection = false;
                escaped_medium = !mi.is_valid();
                active_medium = mi.is_valid();
                if (active_medium) {
                    cnt.n_null++;
                    ray.o = mi.p; ray.mint = 0;
                    si.t = si.t - mi.t;
                    if (m.spectral) transmittance *= mi.sigma_n;
                    else transmittance *= mi.sigma_n / mi.combined_extinction;
                }
            }
            bool intersect = active_surface && needs_intersection;
            if (intersect) si = ray_intersect(scene, ray, cnt);
            needs_intersection = needs_intersection && !intersect;
            active_surface = active_surface || escaped_medium;

            /* si.emitter(scene) (interaction.h:571-596): the environment when nothing was hit; no area emitters */
            bool emitter_hit = scene.environment >= 0 && !si.is_valid() && active_surface;
            if (emitter_hit) { /* volpath.cpp:429-438 */
                emitter_val = (F) scene.emitters[scene.environment].irradiance;
                /* Scene::pdf_emitter_direction (scene.cpp:221-235) over constant.cpp:106-110 */
                emitter_pdf = Consts<F>::InvFourPi;
                if (scene.emitters.size() > 1) emitter_pdf = emitter_pdf * (F) (1.f / (float) scene.emitters.size());   /* scene.cpp:233: (1.f / m_emitters.size()) */
                active = false; active_surface = false; active_medium = false;
            }

            active_surface = active_surface && si.is_valid() && !active_medium;
            if (active_surface)
                transmittance *= bsdf_eval_null_transmission<F>(scene.bsdfs[scene.shapes[si.shape].bsdf]);
            if (active_surface) ray = spawn_ray(si.p, ray.d);
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
    static void update_weights(F &p_over_f, F p, F f, bool active) { /* volpathmis.cpp:447-465 */
        F ratio = p / f;
        ratio = std::isfinite(ratio) ? ratio : (F) 0;
        ratio *= p_over_f;
        if (active) p_over_f = (ratio != ratio) ? (F) 0 : ratio;
    }
    static F mis_weight1(F p_over_f) { return p_over_f == 0 ? (F) 0 : (F) 1 / p_over_f; }            /* :467-482 */
    static F mis_weight2(F a, F b) { F sum = a + b; return sum == 0 ? (F) 0 : (F) 1 / sum; }           /* :485-500 */

    /* volpathmis.cpp:330-444: returns the emitter value; p_over_f_nee / p_over_f_uni through the references */
    F sample_emitter_mis(V3<F> ref_p, bool is_medium_interaction, int medium, F p_over_f, bool active,
                         DirectionSample<F> &ds, F &p_over_f_nee, F &p_over_f_uni) {
        p_over_f_nee = p_over_f; p_over_f_uni = p_over_f;
        F sx = next_1d(active), sy = next_1d(active);
        F weight;
        ds = sample_emitter_direction(scene, sx, sy, weight);
        F emitter_val = weight * ds.pdf;
        if (ds.pdf == 0) emitter_val = 0;
        active = active && ds.pdf != 0;
        update_weights(p_over_f_nee, ds.pdf, (F) 1, active);
        if (!active) return emitter_val;

        Ray<F> ray = spawn_ray(ref_p, ds.d);
        if (is_medium_interaction) ray.mint = 0;
        F total_dist = 0;
        SurfaceInteraction<F> si;
        bool needs_intersection = true;
        while (active) {
            F remaining_dist = ds.dist * ((F) 1 - Consts<F>::ShadowEpsilon) - total_dist;
            ray.maxt = remaining_dist;
            active = active && remaining_dist > 0;
            if (!active) break;
            bool escaped_medium = false;
            bool active_medium = active && medium >= 0;
            bool active_surface = active && !active_medium;
            if (active_medium) {
                const OMedium<F> &m = scene.media[medium];
                MediumInteraction<F> mi = sample_interaction(scene, medium, ray, next_1d(active_medium), cnt);
                if (m.type == B2VP_MEDIUM_HOMOGENEOUS && mi.is_valid()) ray.maxt = min_(mi.t, remaining_dist);
                if (needs_intersection) si = ray_intersect(scene, ray, cnt);
                if (si.t < mi.t) mi.t = Consts<F>::Inf;
                needs_intersection = false;
                bool is_spectral = m.spectral, not_spectral = !m.spectral;
                if (is_spectral) {
                    F t = min_(remaining_dist, min_(mi.t, si.t)) - mi.mint;
                    F tr = exp_(-t * mi.combined_extinction);
                    F free_flight_pdf = (si.t < mi.t || mi.t > remaining_dist) ? tr : tr * mi.combined_extinction;
                    update_weights(p_over_f_nee, free_flight_pdf, tr, true);
                    update_weights(p_over_f_uni, free_flight_pdf, tr, true);
                }
                if (mi.t > remaining_dist && mi.is_valid()) total_dist = ds.dist;
                if (mi.t > remaining_dist) mi.t = Consts<F>::Inf;
                escaped_medium = !mi.is_valid();
                active_medium = mi.is_valid();
                is_spectral = is_spectral && active_medium;
                not_spectral = not_spectral && active_medium;
                if (active_medium) {
                    total_dist += mi.t;
                    cnt.n_null++;
                    ray.o = mi.p; ray.mint = 0;
                    si.t = si.t - mi.t;
                    if (is_spectral) {
                        update_weights(p_over_f_nee, (F) 1, mi.sigma_n, true);
                        update_weights(p_over_f_uni, mi.sigma_n / mi.combined_extinction, mi.sigma_n, true);
                    }
                    if (not_spectral) {
                        update_weights(p_over_f_nee, (F) 1, mi.sigma_n / mi.combined_extinction, true);
                        update_weights(p_over_f_uni, mi.sigma_n, mi.sigma_n, true);
                    }
                }
            }
            bool intersect = active_surface && needs_intersection;
            if (intersect) si = ray_intersect(scene, ray, cnt);
            active_surface = active_surface || escaped_medium;
            if (active_surface) total_dist += si.t;
            active_surface = active_surface && si.is_valid() && active && !active_medium;
            if (active_surface) {
                F bsdf_val = bsdf_eval_null_transmission<F>(scene.bsdfs[scene.shapes[si.shape].bsdf]);
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
    F sample_mis(Ray<F> ray, int initial_medium, bool active, bool &valid_ray_out) {
        bool valid_ray = !hide_emitters && scene.environment >= 0; /* volpathmis.cpp:100 */
        F eta = 1, result = 0;
        int medium = initial_medium;
        MediumInteraction<F> mi;
        bool specular_chain = active && !hide_emitters;
        uint32_t depth = 0;
        F p_over_f = 1, p_over_f_nee = 1;
        SurfaceInteraction<F> si;
        bool needs_intersection = true;
        for (int bounce = 0;; ++bounce) {
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
                is_spectral = is_spectral && scene.media[medium].spectral;
                not_spectral = !is_spectral && active_medium;
            }
            if (active_medium) {
                const OMedium<F> &m = scene.media[medium];
                mi = sample_interaction(scene, medium, ray, next_1d(active_medium), cnt);
                if (m.type == B2VP_MEDIUM_HOMOGENEOUS && mi.is_valid()) ray.maxt = mi.t;
                if (needs_intersection) si = ray_intersect(scene, ray, cnt);
                needs_intersection = false;
                if (si.t < mi.t) mi.t = Consts<F>::Inf;
                if (is_spectral) { /* Medium::eval_tr_and_pdf (medium.cpp:77-89) */
                    F t = min_(mi.t, si.t) - mi.mint;
                    F tr = exp_(-t * mi.combined_extinction);
                    F free_flight_pdf = si.t < mi.t ? tr : tr * mi.combined_extinction;
                    update_weights(p_over_f, free_flight_pdf, tr, true);
                    update_weights(p_over_f_nee, free_flight_pdf, tr, true);
                }
                escaped_medium = !mi.is_valid();
                active_medium = mi.is_valid();
                is_spectral = is_spectral && active_medium;
                not_spectral = not_spectral && active_medium;
            }
            if (active_medium) {
                const OMedium<F> &m = scene.media[medium];
                F u = next_1d(active_medium);
                bool null_scatter = u >= mi.sigma_t / mi.combined_extinction;
                act_null_scatter = null_scatter && active_medium;
                act_medium_scatter = !act_null_scatter && active_medium;
                if (act_medium_scatter) depth += 1;
                bool sample_emitters = m.sample_emitters;
                active = active && depth < max_depth;
                act_medium_scatter = act_medium_scatter && active;
                specular_chain = specular_chain && !(act_medium_scatter && sample_emitters);
                if (act_null_scatter) {
                    cnt.n_null++;
                    if (is_spectral) {
                        update_weights(p_over_f, mi.sigma_n / mi.combined_extinction, mi.sigma_n, true);
                        update_weights(p_over_f_nee, (F) 1, mi.sigma_n, true);
                    }
                    if (not_spectral) {
                        update_weights(p_over_f, mi.sigma_n, mi.sigma_n, true);
                        update_weights(p_over_f_nee, (F) 1, mi.sigma_n / mi.combined_extinction, true);
                    }
                    ray.o = mi.p; ray.mint = 0;
                    si.t = si.t - mi.t;
                }
                if (act_medium_scatter) {
                    cnt.n_real++;
                    if (is_spectral) update_weights(p_over_f, mi.sigma_t / mi.combined_extinction, mi.sigma_s, true);
                    if (not_spectral) update_weights(p_over_f, mi.sigma_t, mi.sigma_s, true);
                    valid_ray = true;
                    if (sample_emitters) {
                        DirectionSample<F> ds;
                        F nee_end, uni_end;
                        F emitted = sample_emitter_mis(mi.p, true, medium, p_over_f, true, ds, nee_end, uni_end);
                        F phase_val = phase_eval_at(scene, m.phase, mi, ds.d);
                        update_weights(nee_end, (F) 1, phase_val, true);
                        update_weights(uni_end, ds.delta ? (F) 0 : phase_val, phase_val, true);
                        result += mis_weight2(nee_end, uni_end) * emitted;
                    }
                    p_over_f_nee = p_over_f;
                    F da = next_1d(true), db = next_1d(true), dc = next_1d(true);
                    F s1 = ltr ? da : dc, s2x = ltr ? db : da, s2y = ltr ? dc : db;
                    F phase_pdf;   /* what phase->sample returns with the direction (volpathmis.cpp:229) */
                    V3<F> wo = phase_sample_at(scene, m.phase, mi, s1, s2x, s2y, &phase_pdf);
                    ray = spawn_ray(mi.p, wo);
                    ray.mint = 0;
                    needs_intersection = true;
                    update_weights(p_over_f, phase_pdf, phase_pdf, true);
                    update_weights(p_over_f_nee, (F) 1, phase_pdf, true);
                }
            }

            active_surface = active_surface || escaped_medium;
            bool intersect = active_surface && needs_intersection;
            if (intersect) si = ray_intersect(scene, ray, cnt);
            /* emitter hit (252-268): the environment when the ray escaped; shapes carry no emitters */
            if (active_surface && scene.environment >= 0 && !si.is_valid() && !(depth == 0 && hide_emitters)) {
                bool count_direct = depth == 0 || specular_chain;
                if (!count_direct) { /* Scene::pdf_emitter_direction (scene.cpp:221-235) over constant.cpp:106-110 */
                    F emitter_pdf = Consts<F>::InvFourPi;
                    if (scene.emitters.size() > 1) emitter_pdf = emitter_pdf * (F) (1.f / (float) scene.emitters.size());   /* scene.cpp:233: (1.f / m_emitters.size()) */
                    update_weights(p_over_f_nee, emitter_pdf, (F) 1, true);
                }
                F emitted = (F) scene.emitters[scene.environment].irradiance;
                result += (count_direct ? mis_weight1(p_over_f) : mis_weight2(p_over_f, p_over_f_nee)) * emitted;
            }
            active_surface = active_surface && si.is_valid();
            if (active_surface) {
                cnt.n_surface++;
                const OShape<F> &sh = scene.shapes[si.shape];
                const B2vpBsdf &bsdf = scene.bsdfs[sh.bsdf];
                bool active_e = (bsdf_flags(bsdf) & F_Smooth) != 0 && (depth + 1 < max_depth);
                if (active_e) {
                    DirectionSample<F> ds;
                    F nee_end, uni_end;
                    F emitted = sample_emitter_mis(si.p, false, medium, p_over_f, true, ds, nee_end, uni_end);
                    V3<F> wo_local = si.sh_frame.to_local(ds.d);
                    F bsdf_val = bsdf_eval(bsdf, si.wi, wo_local);
                    F bsdf_pdf = bsdf_pdf_(bsdf, si.wi, wo_local);
                    update_weights(nee_end, (F) 1, bsdf_val, true);
                    update_weights(uni_end, ds.delta ? (F) 0 : bsdf_pdf, bsdf_val, true);
                    result += mis_weight2(nee_end, uni_end) * emitted;
                }
                F da = next_1d(true), db = next_1d(true), dc = next_1d(true);
                F s1 = ltr ? da : dc, s2x = ltr ? db : da, s2y = ltr ? dc : db;
                F bsdf_weight;
                BSDFSample<F> bs = bsdf_sample(bsdf, si.wi, s1, s2x, s2y, bsdf_weight);
                bool invalid_bsdf_sample = bs.pdf == 0;
                active_surface = active_surface && bs.pdf > 0;
                if (active_surface) {
                    eta *= bs.eta;
                    ray = spawn_ray(si.p, si.sh_frame.to_world(bs.wo));
                }
                needs_intersection = needs_intersection || active_surface;
                bool non_null_bsdf = active_surface && !(bs.sampled_type & F_Null);
                valid_ray = valid_ray || non_null_bsdf || invalid_bsdf_sample;
                specular_chain = specular_chain || (non_null_bsdf && (bs.sampled_type & F_Delta) != 0);
                specular_chain = specular_chain && !(active_surface && (bs.sampled_type & F_Smooth) != 0);
                if (non_null_bsdf) depth += 1;
                if (non_null_bsdf) p_over_f_nee = p_over_f;
                update_weights(p_over_f, bs.pdf, bsdf_weight * bs.pdf, active_surface);
                update_weights(p_over_f_nee, (F) 1, bsdf_weight * bs.pdf, non_null_bsdf);
                if (active_surface && is_medium_transition(si)) medium = target_medium(si, ray.d);
            }
            active = active && (active_surface || active_medium);
        }
        valid_ray_out = valid_ray;
        return result;
    }

    /* volpath.cpp:38-257 */
    F sample(Ray<F> ray, int initial_medium, bool active, bool &valid_ray_out) {
        bool valid_ray = !hide_emitters && scene.environment >= 0; /* volpath.cpp:51 */
        F eta = 1, throughput = 1, result = 0;
        int medium = initial_medium;
        MediumInteraction<F> mi;
        bool specular_chain = active && !hide_emitters;
        uint32_t depth = 0;
        SurfaceInteraction<F> si;
        bool needs_intersection = true;

        for (int bounce = 0;; ++bounce) {
            /* ----------------- termination / Russian roulette (72-87) */
            active = active && throughput != 0;
            F q = min_(throughput * sqr(eta), (F) 0.95f);
            bool perform_rr = depth > rr_depth;
            F rr_sample = next_1d(active);
            active = active && (rr_sample < q || !perform_rr);
            if (perform_rr) throughput *= rcp_(q);
            bool exceeded_max_depth = depth >= max_depth;
            if (!active || exceeded_max_depth) break;

            /* ----------------- sampling the RTE (90-134) */
            bool active_medium = active && medium >= 0;
            bool active_surface = active && !active_medium;
            bool act_null_scatter = false, act_medium_scatter = false, escaped_medium = false;
            bool is_spectral = active_medium, not_spectral = false;
            if (active_medium) {
                is_spectral = is_spectral && scene.media[medium].spectral;
                not_spectral = !is_spectral && active_medium;
            }
            if (active_medium) {
                const OMedium<F> &m = scene.media[medium];
                mi = sample_interaction(scene, medium, ray, next_1d(active_medium), cnt);
                if (m.type == B2VP_MEDIUM_HOMOGENEOUS && mi.is_valid()) ray.maxt = mi.t;
                if (needs_intersection) si = ray_intersect(scene, ray, cnt);
                needs_intersection = false;
                if (si.t < mi.t) mi.t = Consts<F>::Inf;
                if (is_spectral) throughput *= tr_over_pdf(mi, si.t);

                escaped_medium = !mi.is_valid();
                active_medium = mi.is_valid();

                F u = next_1d(active_medium);
                bool null_scatter = u >= mi.sigma_t / mi.combined_extinction;
                act_null_scatter = null_scatter && active_medium;
                act_medium_scatter = !act_null_scatter && active_medium;

                if (is_spectral && act_null_scatter)
                    throughput *= mi.sigma_n * mi.combined_extinction / mi.sigma_n;

                if (act_medium_scatter) depth += 1;
            }

            active = active && depth < max_depth;
            act_medium_scatter = act_medium_scatter && active;

            if (act_null_scatter) { /* 140-144 */
                cnt.n_null++;
                ray.o = mi.p; ray.mint = 0;
                si.t = si.t - mi.t;
            }

            if (act_medium_scatter) { /* 146-176 */
                cnt.n_real++;
                if (is_spectral) throughput *= mi.sigma_s * mi.combined_extinction / mi.sigma_t;
                if (not_spectral) throughput *= mi.sigma_s / mi.sigma_t;

                const OMedium<F> &m = scene.media[mi.medium];
                bool sample_emitters = m.sample_emitters;
                valid_ray = true;
                specular_chain = specular_chain && !act_medium_scatter;
                specular_chain = specular_chain || (act_medium_scatter && !sample_emitters);

                bool active_e = act_medium_scatter && sample_emitters;
                if (active_e) {
                    DirectionSample<F> ds;
                    F emitted = sample_emitter(mi.p, true, medium, active_e, ds);
                    F phase_val = phase_eval_at(scene, m.phase, mi, ds.d);
                    result += throughput * phase_val * emitted;
                }
                /* phase->sample(ctx, mi, next_1d(), next_2d(), ..) (volpath.cpp:171): the
                   order in which the two argument expressions draw is the compiler's
                   choice -- GCC evaluates right to left (next_2d first), clang left to
                   right.  `ltr` selects the clang order; default is GCC's. */
                F da = next_1d(act_medium_scatter), db = next_1d(act_medium_scatter),
                  dc = next_1d(act_medium_scatter);
                F s1 = ltr ? da : dc, s2x = ltr ? db : da, s2y = ltr ? dc : db;
                V3<F> wo = phase_sample_at(scene, m.phase, mi, s1, s2x, s2y);
                Ray<F> new_ray = spawn_ray(mi.p, wo);
                new_ray.mint = 0;
                ray = new_ray;
                needs_intersection = true;
            }

            /* ----------------- surface interactions (178-253) */
            active_surface = active_surface || escaped_medium;
            bool intersect = active_surface && needs_intersection;
            if (intersect) si = ray_intersect(scene, ray, cnt);

            /* emitter hit (186-191): the environment when the ray escaped; shapes carry no emitters */
            if (active_surface && specular_chain && scene.environment >= 0 && !si.is_valid())
                result += throughput * (F) scene.emitters[scene.environment].irradiance;

            active_surface = active_surface && si.is_valid();
            if (active_surface) {
                cnt.n_surface++;
                const OShape<F> &sh = scene.shapes[si.shape];
                const B2vpBsdf &bsdf = scene.bsdfs[sh.bsdf];
                bool active_e = active_surface && (bsdf_flags(bsdf) & F_Smooth) != 0 && (depth + 1 < max_depth);
                if (active_e) {
                    DirectionSample<F> ds;
                    F emitted = sample_emitter(si.p, false, medium, active_e, ds);
                    V3<F> wo = si.sh_frame.to_local(ds.d);
                    F bsdf_val = bsdf_eval(bsdf, si.wi, wo);
                    F bsdf_pdf = bsdf_pdf_(bsdf, si.wi, wo);
                    result += throughput * bsdf_val * mis_weight(ds.pdf, ds.delta ? (F) 0 : bsdf_pdf) * emitted;
                }

                /* bsdf->sample(ctx, si, next_1d(), next_2d(), ..) (volpath.cpp:215-216):
                   same argument-order caveat as for the phase function */
                F da = next_1d(active_surface), db = next_1d(active_surface), dc = next_1d(active_surface);
                F s1 = ltr ? da : dc, s2x = ltr ? db : da, s2y = ltr ? dc : db;
                F bsdf_val;
                BSDFSample<F> bs = bsdf_sample(bsdf, si.wi, s1, s2x, s2y, bsdf_val);

                throughput *= bsdf_val;
                eta *= bs.eta;

                Ray<F> bsdf_ray = spawn_ray(si.p, si.sh_frame.to_world(bs.wo));
                ray = bsdf_ray;
                needs_intersection = true;

                bool non_null_bsdf = active_surface && !(bs.sampled_type & F_Null);
                if (non_null_bsdf) depth += 1;
                valid_ray = valid_ray || non_null_bsdf;
                specular_chain = specular_chain || (non_null_bsdf && (bs.sampled_type & F_Delta) != 0);
                specular_chain = specular_chain && !(active_surface && (bs.sampled_type & F_Smooth) != 0);

                bool add_emitter = active_surface && !(bs.sampled_type & F_Delta) && throughput != 0 &&
                                   (depth < max_depth);
                act_null_scatter = act_null_scatter || (active_surface && (bs.sampled_type & F_Null) != 0);

                bool intersect2 = active_surface && needs_intersection && add_emitter;
                SurfaceInteraction<F> si_new = si;
                if (intersect2) si_new = ray_intersect(scene, ray, cnt);
                needs_intersection = needs_intersection && !intersect2;

                F emitted, emitter_pdf;
                evaluate_direct_light(medium, ray, si_new, add_emitter, emitted, emitter_pdf);
                if (add_emitter && emitter_pdf != 0)
                    result += mis_weight(bs.pdf, emitter_pdf) * throughput * emitted;

                bool has_medium_trans = active_surface && is_medium_transition(si);
                if (has_medium_trans) medium = target_medium(si, ray.d);

                if (intersect2) si = si_new;
            }
            active = active && (active_surface || active_medium);
        }
        valid_ray_out = valid_ray;
        return result;
    }
    F bsdf_pdf_(const B2vpBsdf &b, V3<F> wi, V3<F> wo) const { return bsdf_pdf(b, wi, wo); }
};

/* ======================================================================
 *  Driver: streams, render_sample, film (integrator.cpp:51-288,
 *  imageblock.cpp:79-172, hdrfilm.cpp:190-211)
 * ====================================================================== */
struct Stream {
    uint64_t state, inc;
    int px, py;                 /* absolute pixel (incl. crop offset) */
    int box, boy, bsx, bsy;     /* image block this stream renders into */
    uint32_t n_samples;
    bool empty;
};

uint32_t choose_block_size(const B2vpSceneDesc *d, unsigned n_threads) { /* integrator.cpp:89-97 */
    if (d->integrator.block_size != 0) return d->integrator.block_size;
    uint32_t bs = 32;
    int w = d->sensor.crop_size[0], h = d->sensor.crop_size[1];
    while (true) {
        if (bs == 1 || (uint64_t) ((w + bs - 1) / bs) * ((h + bs - 1) / bs) >= n_threads) break;
        bs /= 2;
    }
    return bs;
}

struct StreamSpace {
    int mode; uint64_t n_streams; uint32_t samples_per_stream;
    uint32_t spp, spp_pass, n_passes, bs;
    std::vector<SpiralBlock> blocks;
    int w, h, ox, oy; uint64_t seed;
};

bool make_stream_space(const B2vpSceneDesc *d, const B2vpRenderArgs *a, StreamSpace &sp) {
    sp.mode = a->seed_mode;
    sp.w = d->sensor.crop_size[0]; sp.h = d->sensor.crop_size[1];
    sp.ox = d->sensor.crop_offset[0]; sp.oy = d->sensor.crop_offset[1];
    sp.seed = d->sensor.sampler_seed;
    sp.spp = d->sensor.sample_count;
    sp.spp_pass = d->integrator.samples_per_pass == 0xffffffffu
                      ? sp.spp : std::min(d->integrator.samples_per_pass, sp.spp);
    if (sp.spp_pass == 0 || sp.spp % sp.spp_pass != 0) {
        g_error = "sample_count must be a multiple of samples_per_pass"; /* integrator.cpp:61-63 */
        return false;
    }
    sp.n_passes = sp.spp / sp.spp_pass;
    sp.bs = 0;
    if (sp.mode == B2VP_SEED_SCALAR_BLOCKS) {
        sp.bs = choose_block_size(d, 1);
        sp.blocks = spiral_blocks(sp.w, sp.h, sp.ox, sp.oy, (int) sp.bs);
        sp.n_streams = (uint64_t) sp.n_passes * sp.blocks.size() * sp.bs * sp.bs;
        sp.samples_per_stream = sp.spp_pass;
    } else if (sp.mode == B2VP_SEED_WAVEFRONT) {
        sp.n_streams = (uint64_t) sp.w * sp.h * sp.spp_pass;
        sp.samples_per_stream = sp.n_passes;
    } else if (sp.mode == B2VP_SEED_EXPLICIT) {
        sp.n_streams = a->n_streams;
        sp.samples_per_stream = a->explicit_samples_per_stream;
    } else {
        g_error = "unknown seed mode";
        return false;
    }
    return true;
}

Stream make_stream(const StreamSpace &sp, const B2vpRenderArgs *a, uint64_t k) {
    Stream st{};
    st.n_samples = sp.samples_per_stream;
    st.empty = false;
    PCG32 rng;
    if (sp.mode == B2VP_SEED_SCALAR_BLOCKS) {
        uint64_t per_pass = (uint64_t) sp.blocks.size() * sp.bs * sp.bs;
        uint64_t pass = k / per_pass, r = k % per_pass;
        uint64_t counter = r / ((uint64_t) sp.bs * sp.bs);
        uint32_t i = (uint32_t) (r % ((uint64_t) sp.bs * sp.bs));
        /* spiral.cpp:42: block_id = counter + (remaining_passes - 1) * block_count */
        uint64_t block_id = counter + (uint64_t) (sp.n_passes - 1 - pass) * sp.blocks.size();
        const SpiralBlock &b = sp.blocks[counter];
        /* sampler->seed(block_id * pixel_count + i) (integrator.cpp:198, sampler.cpp:83-96) */
        rng.seed(sp.seed + block_id * ((uint64_t) sp.bs * sp.bs) + i, PCG32_DEFAULT_STREAM);
        uint32_t x, y; morton_decode(i, x, y);
        st.empty = (int) x >= b.sx || (int) y >= b.sy;
        st.px = b.ox + (int) x; st.py = b.oy + (int) y;
        st.box = b.ox; st.boy = b.oy; st.bsx = b.sx; st.bsy = b.sy;
    } else if (sp.mode == B2VP_SEED_WAVEFRONT) {
        /* sampler->seed(0, wavefront_size) (integrator.cpp:147-148, sampler.cpp:89-92) */
        rng.seed(tea64_u64(sp.seed, k), tea64_u64(k, sp.seed));
        uint64_t pixel = k / sp.spp_pass;
        st.px = (int) (pixel % (uint64_t) sp.w) + sp.ox; st.py = (int) (pixel / (uint64_t) sp.w) + sp.oy;
        st.box = sp.ox; st.boy = sp.oy; st.bsx = sp.w; st.bsy = sp.h;
    } else {
        rng.state = a->explicit_state[k]; rng.inc = a->explicit_inc[k];
        uint32_t pixel = a->explicit_pixel[k];
        st.px = (int) (pixel % (uint32_t) sp.w) + sp.ox; st.py = (int) (pixel / (uint32_t) sp.w) + sp.oy;
        st.box = sp.ox; st.boy = sp.oy; st.bsx = sp.w; st.bsy = sp.h;
    }
    st.state = rng.state; st.inc = rng.inc;
    return st;
}

inline int ceil2int(float x) { return (int) std::ceil(x); }
inline int ceil2int(double x) { return (int) std::ceil(x); }

template <typename F>
int render_typed(const B2vpSceneDesc *d, const B2vpRenderArgs *a, F *film) {
    OScene<F> scene;
    if (!build_scene(d, scene)) return B2VP_ERR_UNSUPPORTED;
    StreamSpace sp;
    if (!make_stream_space(d, a, sp)) return B2VP_ERR_INVALID;
    uint64_t first = a->first_stream;
    uint64_t count = a->n_streams ? a->n_streams : sp.n_streams - first;
    if (sp.mode == B2VP_SEED_EXPLICIT) { first = 0; count = a->n_streams; }
    const uint64_t il_block = a->interleave_block;
    auto stream_index = [&](uint64_t k) -> uint64_t { /* B2vpRenderArgs::interleave_* */
        if (!il_block) return first + k;
        return first + ((k / il_block) * a->interleave_count + a->interleave_index) * il_block + k % il_block;
    };
    if (il_block && (!a->n_streams || !a->interleave_count || a->interleave_index >= a->interleave_count ||
                     sp.mode == B2VP_SEED_EXPLICIT)) { g_error = "invalid interleaved stream range"; return B2VP_ERR_INVALID; }
    if (count && stream_index(count - 1) >= sp.n_streams) { g_error = "stream range out of bounds"; return B2VP_ERR_INVALID; }
    if (!il_block && first + count > sp.n_streams) { g_error = "stream range out of bounds"; return B2VP_ERR_INVALID; }

    const int C = d->integrator.moment ? B2VP_FILM_CHANNELS_MOMENT : B2VP_FILM_CHANNELS;
    const size_t film_n = (size_t) sp.w * sp.h * C;
    if (a->clear_film) std::fill(film, film + film_n, (F) 0);

    B2vpCounters tot{};
    const B2vpSensor &sn = d->sensor;
    uint32_t max_depth = d->integrator.max_depth < 0 ? 0xffffffffu : (uint32_t) d->integrator.max_depth;

    for (uint64_t k = 0; k < count; ++k) {
        Stream st = make_stream(sp, a, stream_index(k));
        PCG32 rng; rng.state = st.state; rng.inc = st.inc;
        Counters cnt;
        F acc[B2VP_FILM_CHANNELS_MOMENT] = { 0 };
        bool acc_used = false;
        for (uint32_t j = 0; j < st.n_samples && !st.empty; ++j) {
            /* render_sample (integrator.cpp:233-288) */
            F posx = (F) (uint32_t) st.px + rng.next_float(F(), true);
            F posy = (F) (uint32_t) st.py + rng.next_float(F(), true);
            F apx = (F) 0.5, apy = (F) 0.5;
            if (sn.needs_aperture_sample) { apx = rng.next_float(F(), true); apy = rng.next_float(F(), true); }
            if (sn.shutter_open_time > 0) (void) rng.next_float(F(), true);
            (void) rng.next_float(F(), true); /* wavelength sample, drawn even in mono */
            F ax = (posx - (F) sn.crop_offset[0]) / (F) sn.crop_size[0];
            F ay = (posy - (F) sn.crop_offset[1]) / (F) sn.crop_size[1];
            Ray<F> ray; bool ok;
            F ray_weight = sensor_sample_ray(scene, ax, ay, apx, apy, ray, ok);
            if (!ok) { g_error = "sensor configuration not supported (ray_origin shape)"; return B2VP_ERR_UNSUPPORTED; }

            Integrator<F> integ{ scene, rng, cnt, max_depth, (uint32_t) d->integrator.rr_depth,
                                 d->integrator.hide_emitters != 0, d->integrator.arg_order_ltr != 0,
                                 d->integrator.mis != 2 };
            bool valid;
            const F raw_value = d->integrator.mis ? integ.sample_mis(ray, sn.medium, true, valid)
                                                  : integ.sample(ray, sn.medium, true, valid);
            F value = ray_weight * raw_value;
            tot.n_paths++;
            if (a->out_radiance)
                ((F *) a->out_radiance)[k * st.n_samples + j] = value;

            /* ImageBlock::put, box filter (imageblock.cpp:85-109,162-169) */
            /* MomentIntegrator::sample (moment.cpp:63-91): the nested integrator's X Y Z without the ray weight,
               then their squares; with AOVs the block does not check the sample (integrator.cpp:113-115) */
            F aovs[B2VP_FILM_CHANNELS_MOMENT] = { value, value, value, valid ? (F) 1 : (F) 0, (F) 1,
                                                  raw_value, raw_value, raw_value,
                                                  raw_value * raw_value, raw_value * raw_value, raw_value * raw_value };
            bool is_valid = true;
            for (int c = 0; c < 5 && C == 5; ++c) is_valid = is_valid && aovs[c] >= (F) -1e-5f && std::isfinite(aovs[c]);
            if (!is_valid) { tot.n_invalid_samples++; continue; }
            if (sn.rfilter_radius > (double) ((F) 0.5 + Consts<F>::RayEpsilon)) {
                /* ImageBlock::put with a filter wider than a pixel (imageblock.cpp:117-152).  The block of this
                   stream has a border of m_border_size pixels (integrator.cpp:114-116, rfilter.cpp:20); the film's
                   own put(block) then adds the block to the storage, which has no border, and so drops what lies
                   outside the crop window (hdrfilm.cpp:201-211, imageblock.cpp:52-76). */
                const int border = (int) sn.rfilter_border;
                const F radius = (F) sn.rfilter_radius, scale = (F) sn.rfilter_scale;
                const int size_x = st.bsx + 2 * border, size_y = st.bsy + 2 * border;
                const F qx = posx - ((F) (st.box - border) + (F) 0.5), qy = posy - ((F) (st.boy - border) + (F) 0.5);
                const int lo_x = std::max(ceil2int(qx - radius), 0), lo_y = std::max(ceil2int(qy - radius), 0);
                const int hi_x = std::min((int) std::floor(qx + radius), size_x - 1),
                          hi_y = std::min((int) std::floor(qy + radius), size_y - 1);
                const uint32_t n = sn.rfilter_taps;
                std::vector<F> weights_x(n), weights_y(n);
                auto eval_discretized = [&](F x) { /* rfilter.h:68-71 */
                    const int index = std::min((int) std::abs(x * scale), B2VP_RFILTER_RESOLUTION);
                    return (F) sn.rfilter_values[index];
                };
                for (uint32_t i = 0; i < n; ++i) {
                    weights_x[i] = eval_discretized(((F) lo_x - qx) + (F) i);
                    weights_y[i] = eval_discretized(((F) lo_y - qy) + (F) i);
                }
                for (uint32_t yr = 0; yr < n; ++yr) {
                    const int y = lo_y + (int) yr;
                    bool enabled = y <= hi_y;
                    for (uint32_t xr = 0; xr < n; ++xr) {
                        const int x = lo_x + (int) xr;
                        const F weight = weights_y[yr] * weights_x[xr];
                        enabled = enabled && x <= hi_x;
                        if (!enabled) continue;
                        const int fx = x - border + st.box - sp.ox, fy = y - border + st.boy - sp.oy;
                        if (fx < 0 || fy < 0 || fx >= (int) sp.w || fy >= (int) sp.h) continue; /* the film's put */
                        for (int c = 0; c < C; ++c) film[((size_t) fy * sp.w + fx) * C + c] += aovs[c] * weight;
                    }
                }
                continue;
            }
            F bx = posx - ((F) st.box + (F) 0.5), by = posy - ((F) st.boy + (F) 0.5);
            int lx = ceil2int(bx - (F) 0.5), ly = ceil2int(by - (F) 0.5);
            if (lx < 0 || ly < 0 || lx >= st.bsx || ly >= st.bsy) continue;
            int fx = lx + st.box - sp.ox, fy = ly + st.boy - sp.oy;
            if (fx == st.px - sp.ox && fy == st.py - sp.oy) {
                for (int c = 0; c < C; ++c) acc[c] += aovs[c];
                acc_used = true;
            } else { /* sample landed on a neighbour pixel (position sample == 0) */
                for (int c = 0; c < C; ++c) film[((size_t) fy * sp.w + fx) * C + c] += aovs[c];
            }
        }
        if (acc_used) {
            size_t base = ((size_t) (st.py - sp.oy) * sp.w + (st.px - sp.ox)) * C;
            for (int c = 0; c < C; ++c) film[base + c] += acc[c];
        }
        if (a->out_state) a->out_state[k] = rng.state;
        if (a->out_counters) {
            B2vpPathCounters &o = a->out_counters[k];
            o.n_rng_draws = rng.draws; o.n_medium_steps = cnt.medium_steps; o.n_null = cnt.n_null;
            o.n_real = cnt.n_real; o.n_intersect = cnt.n_intersect; o.n_surface = cnt.n_surface;
        }
        tot.n_rng_draws += rng.draws; tot.n_medium_steps += cnt.medium_steps; tot.n_null += cnt.n_null;
        tot.n_real += cnt.n_real; tot.n_intersect += cnt.n_intersect; tot.n_surface += cnt.n_surface;
    }
    if (a->totals) *a->totals = tot;
    return B2VP_OK;
}

} // namespace

/* ======================================================================
 *  C entry points (loaded with ctypes by tests/ and bench.py only)
 * ====================================================================== */
/* Probes for the known-answer tests (tests/test_oracle_golden.py): one sensor ray / one Scene::ray_intersect
   of a described scene, in the given precision.  Outputs are doubles. */
template <typename F>
static int probe_sensor_ray_t(const B2vpSceneDesc *desc, double fx, double fy, double ax, double ay, double *o,
                              double *d, double *weight) {
    OScene<F> sc;
    if (!build_scene<F>(desc, sc)) return B2VP_ERR_INVALID;
    Ray<F> ray;
    bool ok = true;
    F w = sensor_sample_ray<F>(sc, (F) fx, (F) fy, (F) ax, (F) ay, ray, ok);
    o[0] = ray.o.x; o[1] = ray.o.y; o[2] = ray.o.z;
    d[0] = ray.d.x; d[1] = ray.d.y; d[2] = ray.d.z;
    *weight = ok ? (double) w : 0.0;
    return B2VP_OK;
}
template <typename F>
static int probe_ray_intersect_t(const B2vpSceneDesc *desc, const double *o, const double *d, double mint,
                                 double maxt, double *t, double *p, double *n, int *shape, double *wi) {
    OScene<F> sc;
    if (!build_scene<F>(desc, sc)) return B2VP_ERR_INVALID;
    Ray<F> ray = make_ray<F>(V3<F>{ (F) o[0], (F) o[1], (F) o[2] }, V3<F>{ (F) d[0], (F) d[1], (F) d[2] },
                             mint < 0 ? Consts<F>::RayEpsilon : (F) mint, maxt < 0 ? Consts<F>::Inf : (F) maxt);
    Counters cnt;
    SurfaceInteraction<F> si = ray_intersect<F>(sc, ray, cnt);
    *t = si.is_valid() ? (double) si.t : -1.0;
    p[0] = si.p.x; p[1] = si.p.y; p[2] = si.p.z;
    n[0] = si.n.x; n[1] = si.n.y; n[2] = si.n.z;
    wi[0] = si.wi.x; wi[1] = si.wi.y; wi[2] = si.wi.z;
    *shape = si.shape;
    return B2VP_OK;
}

extern "C" {

int b2vp_oracle_render(const B2vpSceneDesc *desc, const B2vpRenderArgs *args, void *film) {
    if (!desc || !args || !film || desc->abi_version != B2VP_ABI_VERSION ||
        desc->struct_size != sizeof(B2vpSceneDesc) || args->struct_size != sizeof(B2vpRenderArgs)) {
        g_error = "bad descriptor";
        return B2VP_ERR_INVALID;
    }
    if (args->precision == B2VP_PREC_F32) return render_typed<float>(desc, args, (float *) film);
    return render_typed<double>(desc, args, (double *) film);
}

int b2vp_oracle_stream_count(const B2vpSceneDesc *desc, int32_t seed_mode, uint64_t *n, uint32_t *per) {
    B2vpRenderArgs a{}; a.seed_mode = seed_mode;
    StreamSpace sp;
    if (!make_stream_space(desc, &a, sp)) return B2VP_ERR_INVALID;
    *n = sp.n_streams; *per = sp.samples_per_stream;
    return B2VP_OK;
}

const char *b2vp_oracle_last_error(void) { return g_error.c_str(); }

/* --- unit-level probes used to pin the oracle to the reference's golden vectors */
uint32_t b2vp_oracle_tea32(uint32_t v0, uint32_t v1, int rounds) { return tea32(v0, v1, rounds); }
uint64_t b2vp_oracle_tea64(uint32_t v0, uint32_t v1, int rounds) { return tea64(v0, v1, rounds); }
uint64_t b2vp_oracle_tea64_u64(uint64_t v0, uint64_t v1) { return tea64_u64(v0, v1); }
/* include/mitsuba/core/random.h:137-160 sample_tea_float32/64 */
float b2vp_oracle_tea_float32(uint32_t v0, uint32_t v1, int rounds) {
    return from_bits<float>((tea32(v0, v1, rounds) >> 9) | 0x3f800000u) - 1.f;
}
double b2vp_oracle_tea_float64(uint32_t v0, uint32_t v1, int rounds) {
    return from_bits<double>((tea64(v0, v1, rounds) >> 12) | 0x3ff0000000000000ull) - 1.0;
}
void b2vp_oracle_pcg32(uint64_t initstate, uint64_t initseq, int n, uint32_t *out_u32, float *out_f32,
                       double *out_f64) {
    PCG32 a, b, c;
    a.seed(initstate, initseq); b = a; c = a;
    for (int i = 0; i < n; ++i) {
        if (out_u32) out_u32[i] = a.next_uint32(true);
        if (out_f32) out_f32[i] = b.next_float(0.f, true);
        if (out_f64) out_f64[i] = c.next_float(0.0, true);
    }
}
void b2vp_oracle_math_f32(int fn, int n, const float *x, const float *y, float *out, float *out2) {
    for (int i = 0; i < n; ++i) {
        switch (fn) {
            case 0: out[i] = log_(x[i]); break;
            case 1: out[i] = exp_(x[i]); break;
            case 2: { auto sc = sincos_(x[i]); out[i] = sc.first; out2[i] = sc.second; } break;
            case 3: out[i] = cbrt_(x[i]); break;
            case 4: out[i] = pow_(x[i], y[i]); break;
        }
    }
}
void b2vp_oracle_math_f64(int fn, int n, const double *x, const double *y, double *out, double *out2) {
    for (int i = 0; i < n; ++i) {
        switch (fn) {
            case 0: out[i] = log_(x[i]); break;
            case 1: out[i] = exp_(x[i]); break;
            case 2: { auto sc = sincos_(x[i]); out[i] = sc.first; out2[i] = sc.second; } break;
            case 3: out[i] = cbrt_(x[i]); break;
            case 4: out[i] = pow_(x[i], y[i]); break;
        }
    }
}
/* RPV eval through the BSDF interface: value = f * |cos_o| (rpv.cpp:143-155) */
double b2vp_oracle_rpv_eval(double rho_0, double k, double g, double rho_c, const double *wi, const double *wo) {
    B2vpBsdf b{}; b.type = B2VP_BSDF_RPV; b.p[0] = rho_0; b.p[1] = k; b.p[2] = g; b.p[3] = rho_c;
    return bsdf_eval<double>(b, V3<double>{ wi[0], wi[1], wi[2] }, V3<double>{ wo[0], wo[1], wo[2] });
}

/* Probes for the statistical tests of the reference (src/phase/tests/test_*.py, src/bsdfs/tests/test_*.py):
   n samples of a phase function / BSDF for the incident direction wi, with the value and pdf of each sampled direction.
   samples: n x 3 (sample1, sample2.x, sample2.y); wo: n x 3; value / pdf / weight: n. */
void b2vp_oracle_probe_phase(const B2vpPhase *ph, const double *wi, int n, const double *samples, double *wo, double *value) {
    TabDistr<double> tab;
    if (ph->type == B2VP_PHASE_TABULATED) tab.build(ph->values, ph->n_values);
    MediumInteraction<double> mi;
    mi.wi = V3<double>{ wi[0], wi[1], wi[2] };
    mi.sh_frame = make_frame(-mi.wi);   /* the frame sample_interaction gives a collision: around the ray direction = -wi (medium.cpp:60-62) */
    for (int i = 0; i < n; ++i) {
        V3<double> w = phase_sample(*ph, tab, mi, samples[3 * i + 1], samples[3 * i + 2]);
        wo[3 * i] = w.x; wo[3 * i + 1] = w.y; wo[3 * i + 2] = w.z;
        value[i] = phase_eval(*ph, tab, mi, w);
    }
}
double b2vp_oracle_probe_phase_eval(const B2vpPhase *ph, const double *wi, const double *wo) {
    TabDistr<double> tab;
    if (ph->type == B2VP_PHASE_TABULATED) tab.build(ph->values, ph->n_values);
    MediumInteraction<double> mi;
    mi.wi = V3<double>{ wi[0], wi[1], wi[2] };
    mi.sh_frame = make_frame(-mi.wi);
    return phase_eval(*ph, tab, mi, V3<double>{ wo[0], wo[1], wo[2] });
}
void b2vp_oracle_probe_bsdf(const B2vpBsdf *b, const double *wi, int n, const double *samples, double *wo, double *value,
                            double *pdf, double *weight) {
    const V3<double> w_i{ wi[0], wi[1], wi[2] };
    for (int i = 0; i < n; ++i) {
        double wgt;
        BSDFSample<double> bs = bsdf_sample<double>(*b, w_i, samples[3 * i], samples[3 * i + 1], samples[3 * i + 2], wgt);
        wo[3 * i] = bs.wo.x; wo[3 * i + 1] = bs.wo.y; wo[3 * i + 2] = bs.wo.z;
        value[i] = bsdf_eval<double>(*b, w_i, bs.wo);
        pdf[i] = bsdf_pdf<double>(*b, w_i, bs.wo);
        weight[i] = wgt;
    }
}
void b2vp_oracle_probe_bsdf_eval(const B2vpBsdf *b, const double *wi, int n, const double *wo, double *value, double *pdf) {
    const V3<double> w_i{ wi[0], wi[1], wi[2] };
    for (int i = 0; i < n; ++i) {
        const V3<double> w{ wo[3 * i], wo[3 * i + 1], wo[3 * i + 2] };
        value[i] = bsdf_eval<double>(*b, w_i, w);
        pdf[i] = bsdf_pdf<double>(*b, w_i, w);
    }
}

int b2vp_oracle_probe_sensor_ray(const B2vpSceneDesc *desc, int precision, double fx, double fy, double ax, double ay,
                                 double *o, double *d, double *weight) {
    return precision == B2VP_PREC_F32 ? probe_sensor_ray_t<float>(desc, fx, fy, ax, ay, o, d, weight)
                                      : probe_sensor_ray_t<double>(desc, fx, fy, ax, ay, o, d, weight);
}
int b2vp_oracle_probe_ray_intersect(const B2vpSceneDesc *desc, int precision, const double *o, const double *d,
                                    double mint, double maxt, double *t, double *p, double *n, int *shape, double *wi) {
    return precision == B2VP_PREC_F32 ? probe_ray_intersect_t<float>(desc, o, d, mint, maxt, t, p, n, shape, wi)
                                      : probe_ray_intersect_t<double>(desc, o, d, mint, maxt, t, p, n, shape, wi);
}
}
