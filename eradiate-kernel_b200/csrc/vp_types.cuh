/*
 * vp_types.cuh -- vector algebra, PCG32/TEA and the device-side scene tables.
 *
 * Arithmetic follows the reference's operation order (file:line cited per
 * function, relative to the reference tree) with every fused multiply-add
 * written explicitly; the translation unit is compiled with -fmad=false.
 */
#pragma once
#include "vp_math.cuh"
#include "../../include/b2vp.h"   /* B2VP_RFILTER_RESOLUTION, B2VP_RFILTER_MAX_TAPS */

namespace VP_NS {

/* ------------------------------------------------------------------ vectors */
template <typename F> struct V3 { F x, y, z; };
template <typename F> VP_D V3<F> mk3(F x, F y, F z) { V3<F> r; r.x = x; r.y = y; r.z = z; return r; }
template <typename F> VP_D V3<F> operator+(V3<F> a, V3<F> b) { return mk3<F>(a.x + b.x, a.y + b.y, a.z + b.z); }
template <typename F> VP_D V3<F> operator-(V3<F> a, V3<F> b) { return mk3<F>(a.x - b.x, a.y - b.y, a.z - b.z); }
template <typename F> VP_D V3<F> operator-(V3<F> a) { return mk3<F>(-a.x, -a.y, -a.z); }
template <typename F> VP_D V3<F> operator*(V3<F> a, F s) { return mk3<F>(a.x * s, a.y * s, a.z * s); }
template <typename F> VP_D V3<F> operator*(F s, V3<F> a) { return mk3<F>(s * a.x, s * a.y, s * a.z); }
template <typename F> VP_D F dot(V3<F> a, V3<F> b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
/* enoki::cross = fmsub(a.yzx, b.zxy, a.zxy * b.yzx) (ext/enoki/include/enoki/array_router.h:684) */
template <typename F> VP_D V3<F> cross(V3<F> a, V3<F> b) {
    return mk3<F>(fma_(a.y, b.z, -(a.z * b.y)), fma_(a.z, b.x, -(a.x * b.z)), fma_(a.x, b.y, -(a.y * b.x)));
}
template <typename F> VP_D F rcp_(F x) { return (F) 1 / x; }
template <typename F> VP_D F rsqrt_(F x) { return (F) 1 / sqrt_(x); }
template <typename F> VP_D V3<F> normalize(V3<F> a) { return a * rsqrt_(dot(a, a)); }
template <typename F> VP_D F norm(V3<F> a) { return sqrt_(dot(a, a)); }
template <typename F> VP_D F sqr(F x) { return x * x; }
template <typename F> VP_D F abs_(F x) { return x < 0 ? -x : (x == 0 ? (F) 0 : x); }
template <typename F> VP_D F max_(F a, F b) { return a > b ? a : b; }
template <typename F> VP_D F min_(F a, F b) { return a < b ? a : b; }
template <typename F> VP_D F hmax_abs(V3<F> p) { return max_(max_(abs_(p.x), abs_(p.y)), abs_(p.z)); }
VP_D bool signbit_(float x) { return (__float_as_uint(x) >> 31) != 0; }
VP_D bool signbit_(double x) { return (__double_as_longlong(x) < 0); }
template <typename F> VP_D F mulsign(F a, F b) { return signbit_(b) ? -a : a; }
template <typename F> VP_D F mulsign_neg(F a, F b) { return signbit_(b) ? a : -a; }
template <typename F> VP_D V3<F> fmadd3(V3<F> a, F s, V3<F> b) {
    return mk3<F>(fma_(a.x, s, b.x), fma_(a.y, s, b.y), fma_(a.z, s, b.z));
}
template <typename F> VP_D V3<F> ld3(const F *p) { return mk3<F>(p[0], p[1], p[2]); }

template <typename F> struct K;
template <> struct K<float> {
    static constexpr float Pi = 3.14159265358979323846f;
    static constexpr float InvPi = 0.31830988618379067154f;
    static constexpr float InvFourPi = 0.07957747154594766788f;
    static constexpr float InvTwoPi = 0.15915494309189533577f;
    static constexpr float Epsilon = 1.1920928955078125e-7f / 2;     /* include/mitsuba/core/math.h:35 */
    static constexpr float RayEpsilon = Epsilon * 1500;              /* math.h:36 */
    static constexpr float ShadowEpsilon = RayEpsilon * 10;          /* math.h:37 */
};
template <> struct K<double> {
    static constexpr double Pi = 3.14159265358979323846;
    static constexpr double InvPi = 0.31830988618379067154;
    static constexpr double InvFourPi = 0.07957747154594766788;
    static constexpr double InvTwoPi = 0.15915494309189533577;
    static constexpr double Epsilon = 2.220446049250313e-16 / 2;
    static constexpr double RayEpsilon = Epsilon * 1500;
    static constexpr double ShadowEpsilon = RayEpsilon * 10;
};
template <typename F> VP_D F inf_() { return Lim<F>::inf(); }

/* coordinate_system (include/mitsuba/core/vector.h:116-136) */
template <typename F> VP_D void coordinate_system(V3<F> n, V3<F> &s, V3<F> &t) {
    F sign = signbit_(n.z) ? (F) -1 : (F) 1;
    F a = -rcp_(sign + n.z);
    F b = n.x * n.y * a;
    s = mk3<F>(mulsign(sqr(n.x) * a, n.z) + (F) 1, mulsign(b, n.z), mulsign_neg(n.x, n.z));
    t = mk3<F>(b, sign + sqr(n.y) * a, -n.y);
}

/* Frame3f (include/mitsuba/core/frame.h:30-35) */
template <typename F> struct Frame {
    V3<F> s, t, n;
    VP_D V3<F> to_local(V3<F> v) const { return mk3<F>(dot(v, s), dot(v, t), dot(v, n)); }
    VP_D V3<F> to_world(V3<F> v) const { return s * v.x + t * v.y + n * v.z; }
};
template <typename F> VP_D Frame<F> make_frame(V3<F> v) {
    Frame<F> f;
    f.n = v;
    coordinate_system(v, f.s, f.t);
    return f;
}

/* column-major 4x4 (enoki::Matrix), m[c*4+r]; include/mitsuba/core/transform.h:92-129 */
template <typename F> VP_D V3<F> xf_point(const F *M, V3<F> p) { /* with the w divide */
    F r[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        F v = M[12 + k];
        v = fma_(M[k], p.x, v);
        v = fma_(M[4 + k], p.y, v);
        v = fma_(M[8 + k], p.z, v);
        r[k] = v;
    }
    return mk3<F>(r[0] / r[3], r[1] / r[3], r[2] / r[3]);
}
template <typename F> VP_D V3<F> xf_point_affine(const F *M, V3<F> p) {
    F r[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        F v = M[12 + k];
        v = fma_(M[k], p.x, v);
        v = fma_(M[4 + k], p.y, v);
        v = fma_(M[8 + k], p.z, v);
        r[k] = v;
    }
    return mk3<F>(r[0], r[1], r[2]);
}
template <typename F> VP_D V3<F> xf_vector(const F *M, V3<F> d) {
    F r[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        F v = M[k] * d.x;
        v = fma_(M[4 + k], d.y, v);
        v = fma_(M[8 + k], d.z, v);
        r[k] = v;
    }
    return mk3<F>(r[0], r[1], r[2]);
}

/* Ray3f (include/mitsuba/core/ray.h:30-65) */
template <typename F> struct Ray {
    V3<F> o, d, d_rcp;
    F mint, maxt;
    VP_D void update() { d_rcp = mk3<F>(rcp_(d.x), rcp_(d.y), rcp_(d.z)); }
    VP_D V3<F> at(F t) const { return fmadd3(d, t, o); }
};
template <typename F> VP_D Ray<F> make_ray(V3<F> o, V3<F> d, F mint, F maxt) {
    Ray<F> r;
    r.o = o; r.d = d; r.mint = mint; r.maxt = maxt;
    r.update();
    return r;
}
/* Interaction::spawn_ray (include/mitsuba/render/interaction.h:58-61) */
template <typename F> VP_D Ray<F> spawn_ray(V3<F> p, V3<F> d) {
    return make_ray(p, d, ((F) 1 + hmax_abs(p)) * K<F>::RayEpsilon, inf_<F>());
}

/* BoundingBox::ray_intersect (include/mitsuba/core/bbox.h:302-325) */
template <typename F>
VP_D bool bbox_ray_intersect(const F *bmin, const F *bmax, const Ray<F> &ray, F &mint, F &maxt) {
    bool active = (ray.d.x != 0 || ray.o.x > bmin[0] || ray.o.x < bmax[0]) &&
                  (ray.d.y != 0 || ray.o.y > bmin[1] || ray.o.y < bmax[1]) &&
                  (ray.d.z != 0 || ray.o.z > bmin[2] || ray.o.z < bmax[2]);
    F t1x = (bmin[0] - ray.o.x) * ray.d_rcp.x, t1y = (bmin[1] - ray.o.y) * ray.d_rcp.y,
      t1z = (bmin[2] - ray.o.z) * ray.d_rcp.z;
    F t2x = (bmax[0] - ray.o.x) * ray.d_rcp.x, t2y = (bmax[1] - ray.o.y) * ray.d_rcp.y,
      t2z = (bmax[2] - ray.o.z) * ray.d_rcp.z;
    mint = max_(max_(min_(t1x, t2x), min_(t1y, t2y)), min_(t1z, t2z));
    maxt = min_(min_(max_(t1x, t2x), max_(t1y, t2y)), max_(t1z, t2z));
    return active && maxt >= mint;
}

/* ------------------------------------------------------------------ PCG32 */
/* enoki::PCG32 (ext/enoki/include/enoki/random.h:33-150) */
#define VP_PCG32_MULT 0x5851f42d4c957f2dULL
#define VP_PCG32_DEFAULT_STREAM 0xda3e39cb94b95bdbULL

struct Pcg32 {
    uint64_t state, inc;
    uint32_t draws;
    VP_D uint32_t next_u32(bool mask) {
        uint64_t old = state;
        if (mask) { state = old * VP_PCG32_MULT + inc; draws++; }
        uint32_t xs = (uint32_t) (((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t) (old >> 59u);
        return __funnelshift_r(xs, xs, rot);
    }
    VP_D void seed(uint64_t initstate, uint64_t initseq) {
        state = 0;
        inc = (initseq << 1) | 1u;
        next_u32(true);
        state += initstate;
        next_u32(true);
        draws = 0;
    }
    VP_D float next(float, bool mask) { return __uint_as_float((next_u32(mask) >> 9) | 0x3f800000u) - 1.f; }
    /* random.h:129-141: one 32-bit draw per double */
    VP_D double next(double, bool mask) {
        return __longlong_as_double((long long) (((uint64_t) next_u32(mask) << 20) | 0x3ff0000000000000ull)) - 1.0;
    }
};

/* sample_tea_64 on UInt64 operands, as PCG32Sampler::seed instantiates it in
   wavefront mode (src/librender/sampler.cpp:89-92, include/mitsuba/core/random.h:105-116) */
VP_D uint64_t tea64_u64(uint64_t v0, uint64_t v1) {
    uint64_t sum = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        sum += 0x9e3779b9;
        v0 += ((v1 << 4) + 0xa341316c) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4);
        v1 += ((v0 << 4) + 0xad90777d) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761e);
    }
    return v0 + (v1 << 32);
}

/* enoki::morton_decode (ext/enoki/include/enoki/morton.h): even bits -> x */
VP_D uint32_t compact_bits(uint32_t x) {
    x &= 0x55555555u;
    x = (x ^ (x >> 1)) & 0x33333333u;
    x = (x ^ (x >> 2)) & 0x0f0f0f0fu;
    x = (x ^ (x >> 4)) & 0x00ff00ffu;
    x = (x ^ (x >> 8)) & 0x0000ffffu;
    return x;
}

/* ------------------------------------------------------------ scene tables */
template <typename F> struct DShape {
    int type, bsdf, interior, exterior;
    F to_world[16], to_object[16];
    F fs[3], ft[3], fn[3];      /* rectangle frame: dp_du, dp_dv, n (rectangle.cpp:86-90) */
    F translation[3];
    F inv_surface_area;
    F center[3];
    F radius;
    int flip_normals;
    uint32_t vofs, fofs, fcount;
    int has_normals, has_uv;
};
template <typename F> struct DBsdf { int type, pad; F p[4]; };
/* tabphase: the table (pdf[n] then cdf[n - 1], ContinuousDistribution of distr_1d.h:258-470) lies in the scene blob
   `tab_rel` bytes after this record, so it is staged in shared memory with the other tables. */
template <typename F> struct DPhase {
    int type, n;
    F g;
    int tab_rel;
    uint32_t valid_x, valid_y;
    int weight_volume;               /* blendphase: index into volumes */
    int child[2];                    /* blendphase: indices of the nested phase functions */
    F integral, normalization, interval_size, inv_interval_size;
};
template <typename F> struct DEmitter { F dir[3]; F irradiance; int type, pad[3]; };   /* type 1: constant (environment), `irradiance` = radiance */
template <typename F> struct DVolume {
    int type, filter, wrap, channels;
    int nx, ny, nz, smem_offset;     /* smem_offset: byte offset of the staged grid, or -1 */
    F value, max_value;
    F bmin[3], bmax[3];
    F w2l[16];
    const F *data;                   /* [((z*ny+y)*nx+x)*channels+c] (volume_data.h:65-95) */
    unsigned long long data_bytes;
    /* Trilinear cell records (large single-channel clamp-mode grids only, else null): for every cell
       (cx, cy, cz) in [0, nx] x [0, ny] x [0, nz] the eight corner values d000, d100, d010, d110, d001, d101,
       d011, d111 of the lookup whose floor index is (cx-1, cy-1, cz-1), contiguous and 8*sizeof(F)-aligned:
       one lookup reads one 32-byte sector (fp32) instead of eight words from four rows. */
    const F *cells;
};
template <typename F> struct DMedium {
    int type, sigma_t, albedo, phase;
    int spectral, sample_emitters;
    int paired, pad0;                /* sigma_t and albedo are single-channel trilinear grids of the same geometry */
    F scale, max_density;            /* heterogeneous.cpp:29 */
    F amin[3], amax[3];              /* heterogeneous.cpp:30 */
    /* Majorant grid (opt-in fast mode, no counterpart in the reference, which has one global majorant:
       heterogeneous.cpp:29,38): mg_res cells per axis over the sigma_t volume's unit cube, each holding the maximum
       of UNSCALED sigma_t over everything a lookup inside the cell can touch; mg_offset = index of the first cell in
       the staged array, or -1 when the medium has none. */
    int mg_res[3], mg_offset;
};
template <typename F> struct DSensor {
    int type, direction_type, target_type, origin_type;
    int target_shape, flip_directions, needs_aperture_sample, medium;
    int n_transforms, has_shutter;
    F to_world[16];
    F target_point[3];
    F reference_normal[3];           /* distantflux */
    int crop_offset[2], crop_size[2];
    int n_pixels, origin_shape;      /* distantflux: film width * height; irradiancemeter: the parent rectangle */
    /* reconstruction filter wider than a pixel (rf_taps == 0: the box filter, one pixel per sample): radius,
       table scale, footprint taps, block border and the discretised table (rfilter.h:55-98, rfilter.cpp:9-21) */
    int rf_taps, rf_border;
    F rf_radius, rf_scale;
    F rf_values[B2VP_RFILTER_RESOLUTION + 1];
};

/* 32-byte BVH node over the scene-wide triangle pool.  count == 0: inner node
   with children at child, child+1; count > 0: leaf over prims[child .. child+count). */
struct BvhNode {
    float bmin[3];
    uint32_t child;
    float bmax[3];
    uint32_t count;
};
struct BvhPrim { uint32_t shape, face; };
struct alignas(16) BvhTri { float p0[3], p1[3], p2[3]; uint32_t shape, face, pad; };  /* 48 B */

struct SpiralBlockD { int ox, oy, sx, sy; };

template <typename F> struct DScene {
    int n_shapes, n_analytic, n_media, n_volumes, n_emitters, n_bvh_nodes;
    const DShape<F> *shapes;        /* all shapes; analytic ones (rect/sphere) are listed in analytic[] */
    const int *analytic;
    const DBsdf<F> *bsdfs;
    const DMedium<F> *media;
    const DVolume<F> *volumes;
    const DPhase<F> *phases;
    const DEmitter<F> *emitters;
    const float *vpos, *vnrm, *vuv;
    const uint32_t *faces;
    const BvhNode *bvh;
    const BvhPrim *bvh_prims;
    const BvhTri *bvh_tris;         /* leaf triangles by value, in leaf order */
    const F *sensor_xf;
    F kd_min[3], kd_max[3];         /* scene bbox padded as in kdtree.h:1811-1813 */
    F bs_center[3], bs_radius;
    DSensor<F> sensor;
    uint32_t max_depth, rr_depth;
    int hide_emitters, ltr;
    int film_ch, mis;               /* film channels per pixel: 5, or 11 under a `moment` integrator; volpathmis */
    int environment, pad_env;       /* index of the environment emitter (Scene::environment()) or -1 */
    /* small tables packed in one blob so a CTA can stage them in shared memory */
    const unsigned char *blob;
    uint32_t blob_bytes;
    uint32_t ofs_shapes, ofs_analytic, ofs_bsdfs, ofs_media, ofs_volumes, ofs_phases, ofs_emitters;
    uint32_t grid_smem_bytes;       /* total bytes of grids staged in shared memory */
    uint32_t features;              /* FT_* bits the scene uses (vp_scene.cuh) */
    /* small meshes: BVH nodes and leaf triangles are part of the blob (staged in shared memory with the tables) */
    uint32_t bvh_in_blob, ofs_bvh, ofs_bvh_tris, pad_bvh;
};

/* stream space (integrator.cpp:51-231, sampler.cpp:83-96, spiral.cpp:8-72) */
struct DStreams {
    int mode;
    uint64_t first, count;          /* range rendered by this launch */
    uint64_t il_block;              /* > 0: interleaved blocks (B2vpRenderArgs::interleave_*) */
    uint32_t il_count, il_index;
    int il_shift;                   /* log2(il_block), or -1: no interleaving */
    uint32_t samples_per_stream, spp_pass, n_passes, bs;
    uint32_t n_blocks;
    const SpiralBlockD *blocks;
    int w, h, ox, oy;
    uint64_t seed;
    const uint64_t *ex_state, *ex_inc;
    const uint32_t *ex_pixel;
};

/* global stream index of the k-th stream of a launch.  Interleaved sharding (B2vpRenderArgs::interleave_*): blocks are
   a power of two (il_shift = log2), so the map is a shift, a multiply and a mask -- cheap enough to be part of every
   kernel instance (round 1 kept it out of the single-GPU instances because the general map needed two 64-bit divisions). */
VP_D uint64_t stream_index(const DStreams &sp, uint64_t k) {
    if (sp.il_shift < 0) return sp.first + k;
    return sp.first + ((((k >> sp.il_shift) * sp.il_count + sp.il_index) << sp.il_shift) | (k & (sp.il_block - 1)));
}

struct PathCounters { uint32_t draws, medium_steps, n_null, n_real, n_intersect, n_surface; };
/* BVH work of one ray_intersect call (launch totals only: the memory roofline of the mesh configuration) */
struct BvhWork { uint32_t nodes, tris; };

} // namespace VP_NS
