/*
 * vp_scene.cuh -- device-side shapes, BVH traversal, volumes, media, phase
 * functions, BSDFs, emitter and sensors of the volpath hot path (mono).
 *
 * Each function cites the reference code whose result it reproduces (paths
 * relative to the reference tree).  The reference's bugs are part of the
 * contract: they are reproduced, not fixed.
 */
#pragma once
#include "vp_types.cuh"

namespace VP_NS {

/* Per-thread view of the scene: table pointers resolved to shared memory when
   the CTA staged them there. */
template <typename F> struct SceneView {
    const DScene<F> *g;             /* kernel parameter (constant bank) */
    const DShape<F> *shapes;
    const int *analytic;
    const DBsdf<F> *bsdfs;
    const DMedium<F> *media;
    const DVolume<F> *volumes;
    const DPhase<F> *phases;
    const DEmitter<F> *emitters;
    const unsigned char *grid_smem; /* base of the staged grids */
    /* 32-bit shared-memory addresses (ld.shared) of what the CTA staged: the grids, and -- for scenes whose meshes
       are small (the 12-triangle cube of the atmosphere scenes) -- the BVH nodes and leaf triangles, which are part
       of the table blob then.  Read with explicit ld.shared: a pointer that may be shared OR global is a generic
       pointer to the compiler, and generic loads cost a trip through the global-memory pipeline. */
    uint32_t grid_saddr, bvh_saddr, tri_saddr;
    uint32_t mg_saddr;              /* staged majorant grids, 0 when the launch does not use them */
};

VP_D float4 lds_f4(uint32_t a) {
    float4 v;
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
VP_D float lds_real(uint32_t a, float) { float v; asm("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a)); return v; }
VP_D double lds_real(uint32_t a, double) { double v; asm("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a)); return v; }

template <typename F> struct SurfaceHit {
    F t;
    V3<F> p, n, wi;
    Frame<F> sh;
    int shape;
    uint32_t prim;   /* mesh face */
    F bu, bv;        /* barycentric coordinates on it */
    VP_D bool valid() const { return t != inf_<F>(); }
};

/* ------------------------------------------------------------------ shapes */

/* Scene features a kernel instance is compiled for.  The generic instance (FT_ALL) handles everything;
   the host picks the smallest preset that covers the scene (b2vp_api.cu) so that the hot
   phases -- the kernel is instruction-fetch bound -- carry no code for features the scene does not use.
   Every preset computes bit-identical results to FT_ALL on the scenes it accepts. */
enum : uint32_t {
    FT_HOMOG = 1u << 0,        /* a homogeneous medium is present                         */
    FT_HETERO = 1u << 1,       /* a heterogeneous medium is present                       */
    FT_GRID_ALBEDO = 1u << 2,  /* some medium's albedo is a grid                          */
    FT_GRID_HOMOG = 1u << 3,   /* a homogeneous medium reads a grid volume                */
    FT_NEAREST = 1u << 4,      /* a grid uses the nearest filter                          */
    FT_TRILINEAR = 1u << 5,    /* a grid uses the trilinear filter                        */
    FT_WRAP = 1u << 6,         /* a grid uses repeat / mirror wrapping                    */
    FT_CH3 = 1u << 7,          /* a grid has 3 channels                                   */
    FT_PROJECTIVE = 1u << 8,   /* a grid's world_to_local has a non-trivial last row      */
    FT_RECT = 1u << 9,         /* a rectangle is present                                  */
    FT_SPHERE = 1u << 10,      /* a sphere is present                                     */
    FT_MESH = 1u << 11,        /* triangle meshes whose BVH is read from global memory    */
    FT_MESH_ATTR = 1u << 12,   /* some mesh has vertex normals or texture coordinates     */
    FT_TABPHASE = 1u << 13,    /* a tabulated phase function is present                   */
    FT_BLENDPHASE = 1u << 14,  /* a blended phase function (weight volume) is present     */
    FT_DISTANTFLUX = 1u << 15, /* the sensor is a distantflux or an irradiancemeter, or takes its ray origins from a shape */
    FT_RFILTER = 1u << 16,     /* the film's reconstruction filter is wider than a pixel (no preset: generic instances) */
    FT_GRID_SMEM = 1u << 17,   /* some grid is staged in shared memory                    */
    FT_GRID_GLOBAL = 1u << 18, /* some grid is read from global memory                    */
    FT_MESH_SMEM = 1u << 19,   /* triangle meshes whose BVH and triangles are staged in shared memory (small meshes) */
    FT_MAJORANT_GRID = 1u << 20, /* the launch samples free flights against majorant grids (opt-in fast mode) */
    FT_ALL = 0x1fffffu,
    /* the integrator is volpathmis (src/integrators/volpathmis.cpp) instead of volpath: not a scene feature but another
       state machine in the same phases, so it is outside FT_ALL -- an instance is compiled for one or the other */
    FT_MIS = 1u << 21,
    /* the scene has an environment emitter (src/emitters/constant.cpp): non-delta emitter sampling with both samples,
       radiance where a ray escapes, the specular-chain flag.  Like FT_MIS it selects other phase bodies: outside
       FT_ALL, own instances */
    FT_ENV = 1u << 22
};
constexpr uint32_t FT_ANY_MESH = FT_MESH | FT_MESH_SMEM;


/* Rectangle::ray_intersect_preliminary (src/shapes/rectangle.cpp:139-159) */
template <typename F> VP_D F rectangle_intersect(const DShape<F> &sh, const Ray<F> &r) {
    V3<F> o = xf_point_affine(sh.to_object, r.o), d = xf_vector(sh.to_object, r.d);
    F drz = rcp_(d.z);
    F t = -o.z * drz;
    F lx = fma_(d.x, t, o.x), ly = fma_(d.y, t, o.y);
    bool active = t >= r.mint && t <= r.maxt && abs_(lx) <= (F) 1 && abs_(ly) <= (F) 1;
    return active ? t : inf_<F>();
}

/* The ray origin on a `ray_origin` / `origin` rectangle (distant.cpp:366-375, distantflux.cpp:245-252): the target
   projected onto the shape along -d by Shape::ray_intersect(.., Minimal), i.e. ray_intersect_preliminary above and the
   re-projected hit point of rectangle.cpp:194-197, for the ray Ray3f(target, -d, time) with mint = RayEpsilon.  On a
   miss pi.t is inf, the origin is not finite and the caller zeroes the ray weight. */
template <typename F> VP_D V3<F> origin_on_shape(const DShape<F> &sh, V3<F> target, V3<F> d, bool &valid) {
    Ray<F> tmp = make_ray(target, -d, K<F>::RayEpsilon, inf_<F>());
    const F t = rectangle_intersect(sh, tmp);
    valid = t != inf_<F>();
    V3<F> p = tmp.at(t);
    V3<F> fn = ld3(sh.fn);
    F dist = dot(ld3(sh.translation) - p, fn);
    return fmadd3(fn, dist, p);
}

/* math::solve_quadratic (include/mitsuba/core/math.h) in double */
VP_D bool solve_quadratic(double a, double b, double c, double &x0, double &x1) {
    bool linear_case = a == 0.0;
    bool valid_linear = linear_case && b != 0.0;
    x0 = x1 = -c / b;
    double discrim = fma_(b, b, -(4.0 * a) * c);
    bool valid_quadratic = !linear_case && discrim >= 0.0;
    if (valid_quadratic) {
        double root = sqrt_(discrim);
        double temp = -0.5 * (b + (signbit_(b) ? -root : root));
        double x0p = temp / a, x1p = c / temp;
        x0 = min_(x0p, x1p);
        x1 = max_(x0p, x1p);
    }
    return valid_linear || valid_quadratic;
}

/* Sphere::ray_intersect_preliminary (src/shapes/sphere.cpp:272-306): always double */
template <typename F> VP_D F sphere_intersect(const DShape<F> &sh, const Ray<F> &r) {
    double mint = (double) r.mint, maxt = (double) r.maxt;
    V3<double> o = mk3<double>((double) r.o.x - (double) sh.center[0], (double) r.o.y - (double) sh.center[1],
                               (double) r.o.z - (double) sh.center[2]);
    V3<double> d = mk3<double>((double) r.d.x, (double) r.d.y, (double) r.d.z);
    double A = dot(d, d), B = 2.0 * dot(o, d), C = dot(o, o) - sqr((double) sh.radius);
    double near_t, far_t;
    bool found = solve_quadratic(A, B, C, near_t, far_t);
    bool out_bounds = !(near_t <= maxt && far_t >= mint);
    bool in_bounds = near_t < mint && far_t > maxt;
    bool active = found && !out_bounds && !in_bounds;
    return active ? (near_t < mint ? (F) far_t : (F) near_t) : inf_<F>();
}

template <typename F> VP_D V3<F> load_vertex(const DScene<F> &g, const DShape<F> &sh, uint32_t i) {
    const float *p = g.vpos + 3 * (size_t) (sh.vofs + i);
    return mk3<F>((F) __ldg(p), (F) __ldg(p + 1), (F) __ldg(p + 2));
}

/* Where the BVH lives for this kernel instance: staged in shared memory (FT_MESH_SMEM), in global memory (FT_MESH), or
   decided per scene at run time when the instance is compiled for both (the all-features instance). */
template <typename F, uint32_t FEAT> struct BvhMem {
    const float4 *nodes, *tris;
    uint32_t nodes_s, tris_s;
    bool smem;
    VP_D BvhMem(const SceneView<F> &sv)
        : nodes(reinterpret_cast<const float4 *>(sv.g->bvh)), tris(reinterpret_cast<const float4 *>(sv.g->bvh_tris)),
          nodes_s(sv.bvh_saddr), tris_s(sv.tri_saddr),
          smem((FEAT & FT_MESH_SMEM) && (!(FEAT & FT_MESH) || sv.g->bvh_in_blob)) {}
    /* k-th 16-byte quarter after node i (a node is 32 bytes; two sibling nodes are adjacent) */
    VP_D float4 node(uint32_t i, int k) const { return smem ? lds_f4(nodes_s + i * 32u + (uint32_t) k * 16u) : __ldg(nodes + 2 * (size_t) i + k); }
    VP_D float4 tri(uint32_t i, int k) const { return smem ? lds_f4(tris_s + i * 48u + (uint32_t) k * 16u) : __ldg(tris + 3 * (size_t) i + k); }
};

/* One leaf triangle of the BVH: the three vertex positions (always f32, mesh.h:22-28) and its (shape, face). */
template <typename F, uint32_t FEAT>
VP_D void load_bvh_tri(const BvhMem<F, FEAT> &bm, uint32_t i, V3<F> &p0, V3<F> &p1, V3<F> &p2, BvhPrim &pr) {
    const float4 a = bm.tri(i, 0), b = bm.tri(i, 1), c = bm.tri(i, 2);
    p0 = mk3<F>((F) a.x, (F) a.y, (F) a.z);
    p1 = mk3<F>((F) a.w, (F) b.x, (F) b.y);
    p2 = mk3<F>((F) b.z, (F) b.w, (F) c.x);
    pr.shape = __float_as_uint(c.y);
    pr.face = __float_as_uint(c.z);
}

/* Mesh::ray_intersect_triangle (include/mitsuba/render/mesh.h:195-226), Moeller-Trumbore */
template <typename F>
VP_D F triangle_intersect(V3<F> p0, V3<F> p1, V3<F> p2, const Ray<F> &r, F &u_out, F &v_out) {
    V3<F> e1 = p1 - p0, e2 = p2 - p0;
    V3<F> pvec = cross(r.d, e2);
    F inv_det = rcp_(dot(e1, pvec));
    V3<F> tvec = r.o - p0;
    F u = dot(tvec, pvec) * inv_det;
    bool active = u >= 0 && u <= (F) 1;
    V3<F> qvec = cross(tvec, e1);
    F v = dot(r.d, qvec) * inv_det;
    active = active && v >= 0 && u + v <= (F) 1;
    F t = dot(e2, qvec) * inv_det;
    active = active && t >= r.mint && t <= r.maxt;
    u_out = u;
    v_out = v;
    return active ? t : inf_<F>();
}

/* SurfaceInteraction::initialize_sh_frame (include/mitsuba/render/interaction.h:153-156) */
template <typename F> VP_D void initialize_sh_frame(SurfaceHit<F> &si, V3<F> dp_du) {
    F d = dot(si.sh.n, dp_du);
    V3<F> v = mk3<F>(fma_(-si.sh.n.x, d, dp_du.x), fma_(-si.sh.n.y, d, dp_du.y), fma_(-si.sh.n.z, d, dp_du.z));
    si.sh.s = normalize(v);
    si.sh.t = cross(si.sh.n, si.sh.s);
}

/* Scene::ray_intersect (src/librender/scene.cpp:117-125): closest hit with
   ray.mint <= t <= ray.maxt after clipping against the padded scene box
   (kdtree.h:2078-2171).  Exact ties go to the primitive that comes later in
   (shape, face) order -- the outcome of the reference's `t <= maxt` update rule
   when primitives are visited in that order.  Triangles go through a BVH; its
   slab test is conservative so the candidate set always contains the winner.  */
/* SHADING = false stops after position and geometric normal (all a march needs, and all the main path needs
   until it actually scatters on the surface); surface_frame() below completes the interaction later from
   (shape, prim, bu, bv). */
template <typename F, uint32_t FEAT = FT_ALL, bool SHADING = true>
VP_D void ray_intersect_impl(const SceneView<F> &sv, const Ray<F> &ray_in, SurfaceHit<F> &si, PathCounters &cnt,
                             BvhWork *work = nullptr) {
    const DScene<F> &g = *sv.g;
    cnt.n_intersect++;
    si.t = inf_<F>();
    si.shape = -1;
    si.prim = 0; si.bu = 0; si.bv = 0;
    si.wi = -ray_in.d; /* scene_native.inl:36-38 */
    si.p = mk3<F>(0, 0, 0); si.n = si.p; si.sh.s = si.p; si.sh.t = si.p; si.sh.n = si.p;
    Ray<F> ray = ray_in;
    F bmin, bmax;
    bbox_ray_intersect(g.kd_min, g.kd_max, ray, bmin, bmax);
    F mint = max_(ray.mint, bmin), maxt = min_(ray.maxt, bmax);
    if (!(mint <= maxt))
        return;

    int best_shape = -1;
    uint32_t best_prim = 0, best_tri = 0;
    F best_u = 0, best_v = 0;

    /* analytic shapes, in shape order */
    for (int k = 0; (FEAT & (FT_RECT | FT_SPHERE)) && k < g.n_analytic; ++k) {
        int i = sv.analytic[k];
        const DShape<F> &sh = sv.shapes[i];
        F t;
        if ((FEAT & FT_RECT) && (!(FEAT & FT_SPHERE) || sh.type == 0)) t = rectangle_intersect(sh, ray);
        else t = sphere_intersect(sh, ray);
        if (t != inf_<F>() && (t < ray.maxt || i > best_shape)) {
            best_shape = i;
            ray.maxt = t;
        }
    }

    /* triangles: near-child-first traversal.  An inner node's two children are adjacent (64 bytes, four 16-byte
       loads); both boxes are tested, the nearer one is entered and the farther one pushed with its entry
       distance, so that it can be skipped once a closer hit is known.  Visiting order does not change the
       answer: the winner is the smallest t, ties going to the later (shape, face), and boxes that touch the
       current maxt are still visited. */
    const BvhMem<F, FEAT> bm(sv);
    if ((FEAT & FT_ANY_MESH) && g.n_bvh_nodes > 0) {
        uint32_t stack_n[64];
        float stack_t[64];
        int sp = 0;
        /* conservative slab test in float: widen by a few ulps */
        const float ox = (float) ray.o.x, oy = (float) ray.o.y, oz = (float) ray.o.z;
        const float ix = (float) ray.d_rcp.x, iy = (float) ray.d_rcp.y, iz = (float) ray.d_rcp.z;
        const float lo_s = 1.f - 8.f * 1.1920929e-7f, hi_s = 1.f + 8.f * 1.1920929e-7f;
        const float rmint_lo = (float) ray.mint * lo_s;
        auto inflate_max = [&](float m) { return m > 0.f ? m * hi_s : m * lo_s; };
        auto slab = [&](const float4 a, const float4 b, float cur_max, float &tn) {
            float t1x = (a.x - ox) * ix, t2x = (b.x - ox) * ix;
            float t1y = (a.y - oy) * iy, t2y = (b.y - oy) * iy;
            float t1z = (a.z - oz) * iz, t2z = (b.z - oz) * iz;
            /* NaN (0 * inf) never rejects: fminf/fmaxf drop NaN operands */
            tn = fmaxf(fmaxf(fminf(t1x, t2x), fminf(t1y, t2y)), fminf(t1z, t2z));
            float tf = fminf(fminf(fmaxf(t1x, t2x), fmaxf(t1y, t2y)), fmaxf(t1z, t2z));
            tn = tn > 0.f ? tn * lo_s : tn * hi_s;
            tf = tf > 0.f ? tf * hi_s : tf * lo_s;
            return tn <= tf && tf >= rmint_lo && tn <= cur_max;
        };
        /* a node reference: first child / first triangle in the low 29 bits, triangle count above */
        auto ref_of = [](const float4 a, const float4 b) { return __float_as_uint(a.w) | (__float_as_uint(b.w) << 29); };
        uint32_t cur;
        bool have;
        {
            const float4 a = bm.node(0, 0), b = bm.node(0, 1);
            float tn;
            have = slab(a, b, inflate_max((float) ray.maxt), tn);
            cur = ref_of(a, b);
        }
        while (have) {
            const uint32_t count = cur >> 29, first = cur & 0x1fffffffu;
            if (count == 0) {
                const float4 a0 = bm.node(first, 0), b0 = bm.node(first, 1), a1 = bm.node(first, 2), b1 = bm.node(first, 3);
                if (work) work->nodes += 2;
                const float cur_max = inflate_max((float) ray.maxt);
                float t0, t1;
                const bool h0 = slab(a0, b0, cur_max, t0), h1 = slab(a1, b1, cur_max, t1);
                const uint32_t r0 = ref_of(a0, b0), r1 = ref_of(a1, b1);
                if (h0 && h1) {
                    const bool swap = t1 < t0;
                    stack_n[sp] = swap ? r0 : r1;
                    stack_t[sp] = swap ? t0 : t1;
                    sp++;
                    cur = swap ? r1 : r0;
                    continue;
                }
                if (h0 || h1) { cur = h0 ? r0 : r1; continue; }
            } else {
                for (uint32_t k = 0; k < count; ++k) {
                    /* leaf triangles are stored by value in leaf order: three 16-byte loads per test */
                    V3<F> p0, p1, p2;
                    BvhPrim pr;
                    load_bvh_tri(bm, first + k, p0, p1, p2, pr);
                    if (work) work->tris += 1;
                    F u, v;
                    F t = triangle_intersect(p0, p1, p2, ray, u, v);
                    if (t != inf_<F>()) {
                        bool later = (int) pr.shape > best_shape ||
                                     ((int) pr.shape == best_shape && pr.face > best_prim);
                        if (t < ray.maxt || later) {
                            best_shape = (int) pr.shape; best_prim = pr.face; best_u = u; best_v = v;
                            best_tri = first + k;
                            ray.maxt = t;
                        }
                    }
                }
            }
            /* next: the nearest pending node that can still hold a hit */
            have = false;
            const float cur_max = inflate_max((float) ray.maxt);
            while (sp > 0) {
                --sp;
                if (stack_t[sp] <= cur_max) { cur = stack_n[sp]; have = true; break; }
            }
        }
    }
    if (best_shape < 0)
        return;

    /* PreliminaryIntersection::compute_surface_interaction (interaction.h:571-596) */
    const DShape<F> &sh = sv.shapes[best_shape];
    F t = ray.maxt;
    si.t = t;
    si.shape = best_shape;
    si.prim = best_prim; si.bu = best_u; si.bv = best_v;
    V3<F> dp_du;
    if ((FEAT & FT_RECT) && sh.type == 0) { /* rectangle.cpp:175-212 */
        V3<F> p = ray_in.at(t);
        V3<F> fn = ld3(sh.fn);
        F dist = dot(ld3(sh.translation) - p, fn);
        si.p = fmadd3(fn, dist, p);
        si.n = fn;
        si.sh.n = fn;
        if (SHADING) dp_du = ld3(sh.fs);
    } else if ((FEAT & FT_SPHERE) && sh.type == 1) { /* sphere.cpp:335-400 */
        V3<F> c = ld3(sh.center);
        si.sh.n = normalize(ray_in.at(t) - c);
        si.p = fmadd3(si.sh.n, sh.radius, c);
        if (SHADING) {
            V3<F> local = xf_point_affine(sh.to_object, si.p);
            dp_du = mk3<F>(-local.y, local.x, (F) 0);
            dp_du = xf_vector(sh.to_world, dp_du) * ((F) 2 * K<F>::Pi);
        }
        if (sh.flip_normals) si.sh.n = -si.sh.n;
        si.n = si.sh.n;
    } else if (FEAT & FT_ANY_MESH) { /* src/librender/mesh.cpp:450-545 */
        F b1 = best_u, b2 = best_v, b0 = (F) 1 - b1 - b2;
        V3<F> p0, p1, p2;
        BvhPrim pr;
        load_bvh_tri(bm, best_tri, p0, p1, p2, pr);
        V3<F> dp0 = p1 - p0, dp1 = p2 - p0;
        si.p = p0 * b0 + p1 * b1 + p2 * b2;
        si.n = normalize(cross(dp0, dp1));
        if (!SHADING) return;
        const uint32_t *fi = g.faces + 3 * (size_t) (sh.fofs + best_prim);
        const uint32_t i0 = __ldg(fi), i1 = __ldg(fi + 1), i2 = __ldg(fi + 2);
        V3<F> dp_dv;
        coordinate_system(si.n, dp_du, dp_dv);
        if ((FEAT & FT_MESH_ATTR) && sh.has_uv) {
            const float *t0 = g.vuv + 2 * (size_t) (sh.vofs + i0), *t1 = g.vuv + 2 * (size_t) (sh.vofs + i1),
                        *t2 = g.vuv + 2 * (size_t) (sh.vofs + i2);
            F du0 = (F) t1[0] - (F) t0[0], dv0 = (F) t1[1] - (F) t0[1];
            F du1 = (F) t2[0] - (F) t0[0], dv1 = (F) t2[1] - (F) t0[1];
            F det = fma_(du0, dv1, -(dv0 * du1)), inv_det = rcp_(det);
            if (det != 0) {
                dp_du = mk3<F>(fma_(dv1, dp0.x, -(dv0 * dp1.x)) * inv_det, fma_(dv1, dp0.y, -(dv0 * dp1.y)) * inv_det,
                               fma_(dv1, dp0.z, -(dv0 * dp1.z)) * inv_det);
            }
        }
        if ((FEAT & FT_MESH_ATTR) && sh.has_normals) {
            const float *q0 = g.vnrm + 3 * (size_t) (sh.vofs + i0), *q1 = g.vnrm + 3 * (size_t) (sh.vofs + i1),
                        *q2 = g.vnrm + 3 * (size_t) (sh.vofs + i2);
            V3<F> n0 = mk3<F>((F) q0[0], (F) q0[1], (F) q0[2]), n1 = mk3<F>((F) q1[0], (F) q1[1], (F) q1[2]),
                  n2 = mk3<F>((F) q2[0], (F) q2[1], (F) q2[2]);
            si.sh.n = normalize(n0 * b0 + n1 * b1 + n2 * b2);
        } else {
            si.sh.n = si.n;
        }
    }
    if (!SHADING) return;
    initialize_sh_frame(si, dp_du);
    si.wi = si.sh.to_local(-ray_in.d);
}

/* Out-of-line entry for the kernels that intersect from many places (the one-thread-per-stream and lanes twins); the
   queue kernel inlines ray_intersect_impl at its single call site. */
template <typename F, uint32_t FEAT = FT_ALL, bool SHADING = true>
__device__ __noinline__ void ray_intersect(const SceneView<F> &sv, const Ray<F> &ray_in, SurfaceHit<F> &si,
                                           PathCounters &cnt) {
    ray_intersect_impl<F, FEAT, SHADING>(sv, ray_in, si, cnt);
}

/* The rest of PreliminaryIntersection::compute_surface_interaction (interaction.h:571-596) for a hit found by
   ray_intersect<.., false>: shading normal, dp_du, shading frame and wi -- the same operations on the same
   inputs, so the same bits.  `si` carries t-independent inputs: shape, prim, bu, bv, p, n; `d` is the
   direction of the ray that hit. */
template <typename F, uint32_t FEAT = FT_ALL>
__device__ __noinline__ void surface_frame(const SceneView<F> &sv, V3<F> d, SurfaceHit<F> &si) {
    const DScene<F> &g = *sv.g;
    const DShape<F> &sh = sv.shapes[si.shape];
    V3<F> dp_du;
    if ((FEAT & FT_RECT) && sh.type == 0) {
        si.sh.n = si.n;
        dp_du = ld3(sh.fs);
    } else if ((FEAT & FT_SPHERE) && sh.type == 1) {
        si.sh.n = si.n;
        V3<F> local = xf_point_affine(sh.to_object, si.p);
        dp_du = mk3<F>(-local.y, local.x, (F) 0);
        dp_du = xf_vector(sh.to_world, dp_du) * ((F) 2 * K<F>::Pi);
    } else if (FEAT & FT_ANY_MESH) {
        F b1 = si.bu, b2 = si.bv, b0 = (F) 1 - b1 - b2;
        const uint32_t *fi = g.faces + 3 * (size_t) (sh.fofs + si.prim);
        uint32_t i0 = __ldg(fi), i1 = __ldg(fi + 1), i2 = __ldg(fi + 2);
        V3<F> dp_dv;
        coordinate_system(si.n, dp_du, dp_dv);
        if ((FEAT & FT_MESH_ATTR) && sh.has_uv) {
            V3<F> p0 = load_vertex(g, sh, i0), p1 = load_vertex(g, sh, i1), p2 = load_vertex(g, sh, i2);
            V3<F> dp0 = p1 - p0, dp1 = p2 - p0;
            const float *t0 = g.vuv + 2 * (size_t) (sh.vofs + i0), *t1 = g.vuv + 2 * (size_t) (sh.vofs + i1),
                        *t2 = g.vuv + 2 * (size_t) (sh.vofs + i2);
            F du0 = (F) t1[0] - (F) t0[0], dv0 = (F) t1[1] - (F) t0[1];
            F du1 = (F) t2[0] - (F) t0[0], dv1 = (F) t2[1] - (F) t0[1];
            F det = fma_(du0, dv1, -(dv0 * du1)), inv_det = rcp_(det);
            if (det != 0) {
                dp_du = mk3<F>(fma_(dv1, dp0.x, -(dv0 * dp1.x)) * inv_det, fma_(dv1, dp0.y, -(dv0 * dp1.y)) * inv_det,
                               fma_(dv1, dp0.z, -(dv0 * dp1.z)) * inv_det);
            }
        }
        if ((FEAT & FT_MESH_ATTR) && sh.has_normals) {
            const float *q0 = g.vnrm + 3 * (size_t) (sh.vofs + i0), *q1 = g.vnrm + 3 * (size_t) (sh.vofs + i1),
                        *q2 = g.vnrm + 3 * (size_t) (sh.vofs + i2);
            V3<F> n0 = mk3<F>((F) q0[0], (F) q0[1], (F) q0[2]), n1 = mk3<F>((F) q1[0], (F) q1[1], (F) q1[2]),
                  n2 = mk3<F>((F) q2[0], (F) q2[1], (F) q2[2]);
            si.sh.n = normalize(n0 * b0 + n1 * b1 + n2 * b2);
        } else {
            si.sh.n = si.n;
        }
    }
    initialize_sh_frame(si, dp_du);
    si.wi = si.sh.to_local(-d);
}

/* ----------------------------------------------------------------- volumes */

/* GridVolumeImpl::wrap (src/textures/grid3d.cpp:234-250), one coordinate */
VP_D int wrap_coord(int v, int res, int mode) {
    if (mode == 2 /* clamp */)
        return min(max(v, 0), res - 1);
    int dv = v / res, md = v - dv * res; /* enoki::divisor<int32_t> truncates */
    if (md < 0) md += res;
    if (mode == 1 /* mirror */) {
        bool keep = ((dv & 1) == 0) != (v < 0);
        md = keep ? md : res - 1 - md;
    }
    return md;
}

template <typename F> VP_D int floor2int(F x) { return (int) floor_(x); }

/* One trilinear cell record (DVolume::cells): eight corner values in one aligned, vectorised, read-only load. */
VP_D void load_cell(const float *c, float (&d)[8]) {
    const float4 a = __ldg(reinterpret_cast<const float4 *>(c)), b = __ldg(reinterpret_cast<const float4 *>(c) + 1);
    d[0] = a.x; d[1] = a.y; d[2] = a.z; d[3] = a.w; d[4] = b.x; d[5] = b.y; d[6] = b.z; d[7] = b.w;
}
VP_D void load_cell(const double *c, double (&d)[8]) {
    const double2 *q = reinterpret_cast<const double2 *>(c);
    const double2 a = __ldg(q), b = __ldg(q + 1), e = __ldg(q + 2), f = __ldg(q + 3);
    d[0] = a.x; d[1] = a.y; d[2] = b.x; d[3] = b.y; d[4] = e.x; d[5] = e.y; d[6] = f.x; d[7] = f.y;
}
template <typename F> VP_D size_t cell_index(const DVolume<F> &v, int ix, int iy, int iz) {
    /* clamp mode: every floor index <= -1 reads (0, 0), every index >= n-1 reads (n-1, n-1): same record */
    const int cx = min(max(ix, -1), v.nx - 1) + 1, cy = min(max(iy, -1), v.ny - 1) + 1,
              cz = min(max(iz, -1), v.nz - 1) + 1;
    return (((size_t) cz * (v.ny + 1) + cy) * (v.nx + 1) + cx) * 8;
}
/* Fills the cell records from the plain grid (scene upload, once per grid). */
template <typename F>
__global__ void build_cells_kernel(const F *__restrict__ data, F *__restrict__ cells, int nx, int ny, int nz) {
    const size_t n = (size_t) (nx + 1) * (ny + 1) * (nz + 1);
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x) {
        const int cx = (int) (i % (nx + 1)), cy = (int) ((i / (nx + 1)) % (ny + 1)), cz = (int) (i / ((size_t) (nx + 1) * (ny + 1)));
        const int x0 = min(max(cx - 1, 0), nx - 1), x1 = min(cx, nx - 1);
        const int y0 = min(max(cy - 1, 0), ny - 1), y1 = min(cy, ny - 1);
        const int z0 = min(max(cz - 1, 0), nz - 1), z1 = min(cz, nz - 1);
        const size_t r00 = ((size_t) z0 * ny + y0) * nx, r10 = ((size_t) z0 * ny + y1) * nx,
                     r01 = ((size_t) z1 * ny + y0) * nx, r11 = ((size_t) z1 * ny + y1) * nx;
        F *o = cells + i * 8;
        o[0] = data[r00 + x0]; o[1] = data[r00 + x1]; o[2] = data[r10 + x0]; o[3] = data[r10 + x1];
        o[4] = data[r01 + x0]; o[5] = data[r01 + x1]; o[6] = data[r11 + x0]; o[7] = data[r11 + x1];
    }
}


/* GridVolumeImpl::eval (src/textures/grid3d.cpp:163-185,220-360) and
   ConstVolume::eval (src/textures/constant3d.cpp:31-37) */
template <typename F, uint32_t FEAT, bool MAY_BE_GRID = true>
VP_D F volume_eval(const SceneView<F> &sv, const DVolume<F> &v, V3<F> p_world) {
    if (!MAY_BE_GRID || v.type == 0)
        return v.value;
    /* grid3d.cpp:227: full 4x4 with the w divide.  With a last row of (0,0,0,1) the divisor is exactly 1
       and x / 1 == x, so the affine product gives the same bits. */
    V3<F> p = (FEAT & FT_PROJECTIVE) ? xf_point(v.w2l, p_world) : xf_point_affine(v.w2l, p_world);
    const int nx = v.nx, ny = v.ny, nz = v.nz;
    const int ch = (FEAT & FT_CH3) ? v.channels : 1;
    const int wrap = (FEAT & FT_WRAP) ? v.wrap : 2;
    /* a staged grid is read with ld.shared, a grid in HBM with global loads; instances compiled for scenes that have
       only one kind (FT_GRID_SMEM / FT_GRID_GLOBAL) carry only that path */
    const bool staged = (FEAT & FT_GRID_SMEM) && (!(FEAT & FT_GRID_GLOBAL) || v.smem_offset >= 0);
    const uint32_t gs = sv.grid_saddr + (uint32_t) v.smem_offset;
    auto data = [&](size_t i) -> F { return staged ? lds_real(gs + (uint32_t) i * (uint32_t) sizeof(F), F()) : v.data[i]; };
    F res[3];
    const bool trilinear = (FEAT & FT_TRILINEAR) && (!(FEAT & FT_NEAREST) || v.filter == 1);
    if (trilinear) { /* trilinear, grid3d.cpp:271-341 */
        p = mk3<F>(fma_(p.x, (F) nx, (F) -0.5), fma_(p.y, (F) ny, (F) -0.5), fma_(p.z, (F) nz, (F) -0.5));
        int ix = floor2int(p.x), iy = floor2int(p.y), iz = floor2int(p.z);
        F w1x = p.x - (F) ix, w1y = p.y - (F) iy, w1z = p.z - (F) iz;
        F w0x = (F) 1 - w1x, w0y = (F) 1 - w1y, w0z = (F) 1 - w1z;
        if ((FEAT & FT_GRID_GLOBAL) && v.cells) { /* large grid: one aligned record holds the eight corners (same values, same arithmetic) */
            F d[8];
            load_cell(v.cells + cell_index(v, ix, iy, iz), d);
            F v00 = fma_(w0x, d[0], w1x * d[1]), v01 = fma_(w0x, d[4], w1x * d[5]);
            F v10 = fma_(w0x, d[2], w1x * d[3]), v11 = fma_(w0x, d[6], w1x * d[7]);
            F v0 = fma_(w0y, v00, w1y * v10), v1 = fma_(w0y, v01, w1y * v11);
            return fma_(w0z, v0, w1z * v1);
        }
        int z0 = wrap_coord(iz, nz, wrap), z1 = wrap_coord(iz + 1, nz, wrap);
        if (nx == 1 && ny == 1 && ch == 1) {
            /* a 1 x 1 x N column (the plane-parallel atmospheres of C2 / C4): every wrap mode maps both x and both y
               indices to 0, so four of the eight corners are data[z0] and four data[z1].  Two loads instead of eight;
               the interpolation is evaluated on the same operands in the same order, so the bits are the same (the
               x- and y-lerps of equal values are NOT skipped: w0 * a + w1 * a need not round to a). */
            const F a = data((size_t) z0), b = data((size_t) z1);
            const F v00 = fma_(w0x, a, w1x * a), v01 = fma_(w0x, b, w1x * b);
            const F v0 = fma_(w0y, v00, w1y * v00), v1 = fma_(w0y, v01, w1y * v01);
            return fma_(w0z, v0, w1z * v1);
        }
        int x0 = wrap_coord(ix, nx, wrap), x1 = wrap_coord(ix + 1, nx, wrap);
        int y0 = wrap_coord(iy, ny, wrap), y1 = wrap_coord(iy + 1, ny, wrap);
        size_t r00 = ((size_t) z0 * ny + y0) * nx, r10 = ((size_t) z0 * ny + y1) * nx,
               r01 = ((size_t) z1 * ny + y0) * nx, r11 = ((size_t) z1 * ny + y1) * nx;
#pragma unroll 1
        for (int c = 0; c < ch; ++c) {
            F d000 = data((r00 + x0) * ch + c), d100 = data((r00 + x1) * ch + c);
            F d010 = data((r10 + x0) * ch + c), d110 = data((r10 + x1) * ch + c);
            F d001 = data((r01 + x0) * ch + c), d101 = data((r01 + x1) * ch + c);
            F d011 = data((r11 + x0) * ch + c), d111 = data((r11 + x1) * ch + c);
            F v00 = fma_(w0x, d000, w1x * d100), v01 = fma_(w0x, d001, w1x * d101);
            F v10 = fma_(w0x, d010, w1x * d110), v11 = fma_(w0x, d011, w1x * d111);
            F v0 = fma_(w0y, v00, w1y * v10), v1 = fma_(w0y, v01, w1y * v11);
            res[c] = fma_(w0z, v0, w1z * v1);
        }
    } else { /* nearest, grid3d.cpp:342-356 */
        p = mk3<F>(p.x * (F) nx, p.y * (F) ny, p.z * (F) nz);
        int x = wrap_coord(floor2int(p.x), nx, wrap), y = wrap_coord(floor2int(p.y), ny, wrap),
            z = wrap_coord(floor2int(p.z), nz, wrap);
        size_t idx = (((size_t) z * ny + y) * nx + x) * ch;
#pragma unroll 1
        for (int c = 0; c < ch; ++c) res[c] = data(idx + c);
    }
    if (ch == 1) return res[0];
    /* mitsuba::luminance(Color3f) with float literals (include/mitsuba/core/spectrum.h:247) */
    return res[0] * (F) 0.212671f + res[1] * (F) 0.715160f + res[2] * (F) 0.072169f;
}

/* Two single-channel trilinear grids with the same geometry (resolution, world_to_local, wrap mode) evaluated at
   one point: the local position, the eight indices and the weights are the same numbers for both, so they are
   computed once, and the sixteen gathers are issued together (they are the latency that bounds large grids).
   Each value goes through exactly the operations of volume_eval, so the bits are the same. */
template <typename F, uint32_t FEAT>
VP_D void volume_eval_pair(const SceneView<F> &sv, const DVolume<F> &va, const DVolume<F> &vb, V3<F> p_world, F &ra,
                           F &rb) {
    V3<F> p = (FEAT & FT_PROJECTIVE) ? xf_point(va.w2l, p_world) : xf_point_affine(va.w2l, p_world);
    const int nx = va.nx, ny = va.ny, nz = va.nz;
    const int wrap = (FEAT & FT_WRAP) ? va.wrap : 2;
    const bool sa = (FEAT & FT_GRID_SMEM) && (!(FEAT & FT_GRID_GLOBAL) || va.smem_offset >= 0);
    const bool sb = (FEAT & FT_GRID_SMEM) && (!(FEAT & FT_GRID_GLOBAL) || vb.smem_offset >= 0);
    const uint32_t ga = sv.grid_saddr + (uint32_t) va.smem_offset, gb = sv.grid_saddr + (uint32_t) vb.smem_offset;
    auto da = [&](size_t i) -> F { return sa ? lds_real(ga + (uint32_t) i * (uint32_t) sizeof(F), F()) : va.data[i]; };
    auto db = [&](size_t i) -> F { return sb ? lds_real(gb + (uint32_t) i * (uint32_t) sizeof(F), F()) : vb.data[i]; };
    p = mk3<F>(fma_(p.x, (F) nx, (F) -0.5), fma_(p.y, (F) ny, (F) -0.5), fma_(p.z, (F) nz, (F) -0.5));
    int ix = floor2int(p.x), iy = floor2int(p.y), iz = floor2int(p.z);
    F w1x = p.x - (F) ix, w1y = p.y - (F) iy, w1z = p.z - (F) iz;
    F w0x = (F) 1 - w1x, w0y = (F) 1 - w1y, w0z = (F) 1 - w1z;
    if ((FEAT & FT_GRID_GLOBAL) && va.cells && vb.cells) { /* both grids have cell records: two sector loads instead of sixteen gathers */
        F a[8], b[8];
        const size_t ci = cell_index(va, ix, iy, iz);
        load_cell(va.cells + ci, a);
        load_cell(vb.cells + ci, b);
        {
            F v00 = fma_(w0x, a[0], w1x * a[1]), v01 = fma_(w0x, a[4], w1x * a[5]);
            F v10 = fma_(w0x, a[2], w1x * a[3]), v11 = fma_(w0x, a[6], w1x * a[7]);
            F v0 = fma_(w0y, v00, w1y * v10), v1 = fma_(w0y, v01, w1y * v11);
            ra = fma_(w0z, v0, w1z * v1);
        }
        {
            F v00 = fma_(w0x, b[0], w1x * b[1]), v01 = fma_(w0x, b[4], w1x * b[5]);
            F v10 = fma_(w0x, b[2], w1x * b[3]), v11 = fma_(w0x, b[6], w1x * b[7]);
            F v0 = fma_(w0y, v00, w1y * v10), v1 = fma_(w0y, v01, w1y * v11);
            rb = fma_(w0z, v0, w1z * v1);
        }
        return;
    }
    int x0 = wrap_coord(ix, nx, wrap), x1 = wrap_coord(ix + 1, nx, wrap);
    int y0 = wrap_coord(iy, ny, wrap), y1 = wrap_coord(iy + 1, ny, wrap);
    int z0 = wrap_coord(iz, nz, wrap), z1 = wrap_coord(iz + 1, nz, wrap);
    size_t r00 = ((size_t) z0 * ny + y0) * nx, r10 = ((size_t) z0 * ny + y1) * nx,
           r01 = ((size_t) z1 * ny + y0) * nx, r11 = ((size_t) z1 * ny + y1) * nx;
    F a000 = da(r00 + x0), a100 = da(r00 + x1), a010 = da(r10 + x0), a110 = da(r10 + x1);
    F a001 = da(r01 + x0), a101 = da(r01 + x1), a011 = da(r11 + x0), a111 = da(r11 + x1);
    F b000 = db(r00 + x0), b100 = db(r00 + x1), b010 = db(r10 + x0), b110 = db(r10 + x1);
    F b001 = db(r01 + x0), b101 = db(r01 + x1), b011 = db(r11 + x0), b111 = db(r11 + x1);
    {
        F v00 = fma_(w0x, a000, w1x * a100), v01 = fma_(w0x, a001, w1x * a101);
        F v10 = fma_(w0x, a010, w1x * a110), v11 = fma_(w0x, a011, w1x * a111);
        F v0 = fma_(w0y, v00, w1y * v10), v1 = fma_(w0y, v01, w1y * v11);
        ra = fma_(w0z, v0, w1z * v1);
    }
    {
        F v00 = fma_(w0x, b000, w1x * b100), v01 = fma_(w0x, b001, w1x * b101);
        F v10 = fma_(w0x, b010, w1x * b110), v11 = fma_(w0x, b011, w1x * b111);
        F v0 = fma_(w0y, v00, w1y * v10), v1 = fma_(w0y, v01, w1y * v11);
        rb = fma_(w0z, v0, w1z * v1);
    }
}

/* ------------------------------------------------------------------- media */
template <typename F> struct MediumHit {
    F t, mint;
    V3<F> p;
    F sigma_s, sigma_n, sigma_t, combined;
    VP_D bool valid() const { return t != inf_<F>(); }
};

/* Medium::sample_interaction (src/librender/medium.cpp:36-75) over
   HeterogeneousMedium / HomogeneousMedium (src/media/*.cpp).  The shading frame
   Frame3f(ray.d) and wi = -ray.d that the reference stores per step are pure
   functions of ray.d and are rebuilt where a scattering event needs them. */
template <typename F, uint32_t FEAT = FT_ALL>
VP_D void sample_interaction(const SceneView<F> &sv, const DMedium<F> &m, const Ray<F> &ray, F sample,
                             MediumHit<F> &mi, PathCounters &cnt) {
    cnt.medium_steps++;
    /* a preset without FT_HOMOG / FT_HETERO knows the medium type at compile time */
    const bool homog = (FEAT & FT_HOMOG) && (!(FEAT & FT_HETERO) || m.type == 0);
    constexpr bool HGRID = (FEAT & FT_GRID_HOMOG) != 0;
    bool aabb_its;
    F mint, maxt;
    if (homog) { /* homogeneous.cpp:51-54 */
        aabb_its = true; mint = 0; maxt = inf_<F>();
    } else {     /* heterogeneous.cpp:50-53 */
        aabb_its = bbox_ray_intersect(m.amin, m.amax, ray, mint, maxt);
    }
    aabb_its = aabb_its && (finite_(mint) || finite_(maxt));
    bool active = aabb_its;
    if (!active) { mint = 0; maxt = inf_<F>(); }
    mint = max_(ray.mint, mint);
    maxt = min_(ray.maxt, maxt);
    F combined;
    if (homog)
        combined = volume_eval<F, FEAT, HGRID>(sv, sv.volumes[m.sigma_t], ray.o) * m.scale; /* homogeneous.cpp:30-38 */
    else
        combined = m.max_density;
    F sampled_t = mint + (-log_((F) 1 - sample) / combined);
    bool valid_mi = active && (sampled_t <= maxt);
    mi.t = valid_mi ? sampled_t : inf_<F>();
    mi.p = ray.at(sampled_t);
    mi.mint = mint;
    mi.sigma_s = 0; mi.sigma_n = 0; mi.sigma_t = 0;
    if (valid_mi) {
        F albedo, sigmat;
        if ((FEAT & FT_GRID_ALBEDO) && (FEAT & FT_TRILINEAR) && !homog && m.paired) {
            F st;
            volume_eval_pair<F, FEAT>(sv, sv.volumes[m.albedo], sv.volumes[m.sigma_t], mi.p, albedo, st);
            sigmat = m.scale * st;
            mi.sigma_n = combined - sigmat;
        } else {
        albedo = volume_eval<F, FEAT, (FEAT & FT_GRID_ALBEDO) != 0>(sv, sv.volumes[m.albedo], mi.p);
        if (homog) {
            sigmat = volume_eval<F, FEAT, HGRID>(sv, sv.volumes[m.sigma_t], mi.p) * m.scale;
            mi.sigma_n = 0;
        } else {
            sigmat = m.scale * volume_eval<F, FEAT, true>(sv, sv.volumes[m.sigma_t], mi.p);
            mi.sigma_n = combined - sigmat;
        }
        }
        mi.sigma_t = sigmat;
        mi.sigma_s = sigmat * albedo;
    }
    mi.combined = combined;
}

/* Free-flight sampling against a MAJORANT GRID instead of the medium's one global majorant.
   Not in the reference (heterogeneous.cpp:29,38 has only the global majorant): an opt-in mode whose
   tentative collisions are a different -- much sparser -- random sequence with the same distribution of real
   collisions, so it changes how many random numbers a path draws and is validated statistically, never bit for bit.
   One uniform variate gives the optical depth tau = -log(1 - u) to the next tentative collision; the ray is walked
   through the coarse cells (3D-DDA in the volume's local coordinates), accumulating mu_cell * dt until tau is used up.
   mi.combined is the majorant of the cell the collision falls in: every weight of the integrator that refers to the
   majorant (null probability sigma_n / mu, ratio-tracking factor, tr / pdf = 1 / mu) uses it unchanged.
   t_limit: nothing beyond it is looked at by the caller (the closest surface, the remaining distance of a march). */
template <typename F, uint32_t FEAT>
VP_D void sample_interaction_mg(const SceneView<F> &sv, const DMedium<F> &m, const Ray<F> &ray, F sample, F t_limit,
                                MediumHit<F> &mi, PathCounters &cnt) {
    cnt.medium_steps++;
    F mint, maxt;
    bool active = bbox_ray_intersect(m.amin, m.amax, ray, mint, maxt);
    active = active && (finite_(mint) || finite_(maxt));
    if (!active) { mint = 0; maxt = inf_<F>(); }
    mint = max_(ray.mint, mint);
    maxt = min_(min_(ray.maxt, maxt), t_limit);
    mi.t = inf_<F>(); mi.mint = mint; mi.p = ray.o;
    mi.sigma_s = 0; mi.sigma_n = 0; mi.sigma_t = 0; mi.combined = m.max_density;
    if (!active || !(mint < maxt)) return;
    const DVolume<F> &vol = sv.volumes[m.sigma_t];
    const int Rx = m.mg_res[0], Ry = m.mg_res[1], Rz = m.mg_res[2];
    /* local coordinates scaled to cells: q(t) = qo + t * qd.  Everything per axis lives in scalars (an array indexed
       by the axis of the next crossing would live in local memory). */
    const V3<F> ol = xf_point_affine(vol.w2l, ray.o), dl = xf_vector(vol.w2l, ray.d);
    const F qox = ol.x * (F) Rx, qoy = ol.y * (F) Ry, qoz = ol.z * (F) Rz;
    const F qdx = dl.x * (F) Rx, qdy = dl.y * (F) Ry, qdz = dl.z * (F) Rz;
    F t = mint;
    int cx, cy, cz, sx, sy, sz;
    F tnx, tny, tnz, tdx, tdy, tdz;
    auto setup = [&](F qo, F qd, int R, int &c, int &st, F &tn, F &td) {
        c = (int) floor_(fma_(qd, t, qo));
        c = c < 0 ? 0 : (c > R - 1 ? R - 1 : c);
        if (qd > 0) { st = 1; td = (F) 1 / qd; tn = ((F) (c + 1) - qo) * td; }
        else if (qd < 0) { st = -1; td = (F) -1 / qd; tn = ((F) c - qo) / qd; }
        else { st = 0; td = inf_<F>(); tn = inf_<F>(); }
    };
    setup(qox, qdx, Rx, cx, sx, tnx, tdx);
    setup(qoy, qdy, Ry, cy, sy, tny, tdy);
    setup(qoz, qdz, Rz, cz, sz, tnz, tdz);
    const F tau = -log_((F) 1 - sample);
    const uint32_t base = sv.mg_saddr + (uint32_t) m.mg_offset * 4u;
    F acc = 0;
    /* At most MG_WALK cells per call: the number of cells to the next collision differs wildly between the lanes of a
       warp (1 .. 3 * 16), and a warp pays for its longest walk.  A walk that is not over after MG_WALK cells returns a
       VIRTUAL collision at the cell face it has reached: sigma_t = 0 there, so it is a null collision with
       probability 1 and weight exactly 1, and the caller draws a fresh optical depth from that point on -- which the
       exponential distribution allows (it has no memory).  Same estimator, bounded work per visit. */
    constexpr int MG_WALK = 8;
#pragma unroll 1
    for (int guard = 0; guard < MG_WALK; ++guard) {
        const F tn = min_(tnx, min_(tny, tnz));
        const F t_exit = min_(max_(tn, t), maxt);
        const F mu = m.scale * (F) lds_real(base + (uint32_t) ((cz * Ry + cy) * Rx + cx) * 4u, float());
        const F dtau = mu * (t_exit - t);
        if (acc + dtau >= tau && mu > 0) {   /* the collision falls in this cell */
            const F tc = t + (tau - acc) / mu;
            if (tc <= maxt) {
                mi.t = tc;
                mi.p = ray.at(tc);
                mi.combined = mu;
                F albedo, sigmat;
                if ((FEAT & FT_GRID_ALBEDO) && (FEAT & FT_TRILINEAR) && m.paired) {
                    F st;
                    volume_eval_pair<F, FEAT>(sv, sv.volumes[m.albedo], vol, mi.p, albedo, st);
                    sigmat = m.scale * st;
                } else {
                    albedo = volume_eval<F, FEAT, (FEAT & FT_GRID_ALBEDO) != 0>(sv, sv.volumes[m.albedo], mi.p);
                    sigmat = m.scale * volume_eval<F, FEAT, true>(sv, vol, mi.p);
                }
                mi.sigma_t = sigmat;
                mi.sigma_n = mu - sigmat;
                mi.sigma_s = sigmat * albedo;
            }
            return;
        }
        acc += dtau;
        t = t_exit;
        if (!(t < maxt)) return;
        /* cross into the next cell along the axis whose boundary comes first */
        if (tnx <= tny && tnx <= tnz) { cx += sx; tnx += tdx; if (cx < 0 || cx >= Rx) return; }
        else if (tny <= tnz) { cy += sy; tny += tdy; if (cy < 0 || cy >= Ry) return; }
        else { cz += sz; tnz += tdz; if (cz < 0 || cz >= Rz) return; }
    }
    /* virtual collision at the face reached (t < maxt here): null with certainty, weight sigma_n / combined = 1 */
    mi.t = t;
    mi.p = ray.at(t);
    mi.combined = m.max_density;
    mi.sigma_n = m.max_density;
    mi.sigma_t = 0; mi.sigma_s = 0;
}

/* Medium::eval_tr_and_pdf (medium.cpp:77-89) followed by tr / tr_pdf
   (src/integrators/volpath.cpp:113-117,402-404) */
template <typename F> VP_D F tr_over_pdf(const MediumHit<F> &mi, F si_t) {
    F t = min_(mi.t, si_t) - mi.mint;
    F tr = exp_(-t * mi.combined);
    F pdf = si_t < mi.t ? tr : tr * mi.combined;
    return pdf > 0 ? tr / pdf : (F) 0;
}

/* ------------------------------------------------------------------- warps */
/* warp::square_to_uniform_disk_concentric (include/mitsuba/core/warp.h:54-90) */
template <typename F> VP_D void square_to_uniform_disk_concentric(F sx, F sy, F &ox, F &oy) {
    F x = fma_((F) 2, sx, (F) -1), y = fma_((F) 2, sy, (F) -1);
    bool is_zero = x == 0 && y == 0, quadrant_1_or_3 = abs_(x) < abs_(y);
    F r = quadrant_1_or_3 ? y : x, rp = quadrant_1_or_3 ? x : y;
    F phi = (F) 0.25 * K<F>::Pi * rp / r;
    if (quadrant_1_or_3) phi = (F) 0.5 * K<F>::Pi - phi;
    if (is_zero) phi = 0;
    F s, c;
    sincos_(phi, s, c);
    ox = r * c;
    oy = r * s;
}
/* warp::square_to_cosine_hemisphere (warp.h:325-334) */
template <typename F> VP_D V3<F> square_to_cosine_hemisphere(F sx, F sy) {
    F px, py;
    square_to_uniform_disk_concentric(sx, sy, px, py);
    F z = safe_sqrt_((F) 1 - (px * px + py * py));
    return mk3<F>(px, py, z);
}
/* warp::square_to_uniform_hemisphere (warp.h:298-303) */
template <typename F> VP_D V3<F> square_to_uniform_hemisphere(F sx, F sy) {
    F px, py;
    square_to_uniform_disk_concentric(sx, sy, px, py);
    F z = (F) 1 - (px * px + py * py);
    F k = sqrt_(z + (F) 1);
    return mk3<F>(px * k, py * k, z);
}
/* warp::square_to_uniform_sphere (warp.h:255-260) */
template <typename F> VP_D V3<F> square_to_uniform_sphere(F sx, F sy) {
    F z = fma_((F) -2, sy, (F) 1);
    F r = safe_sqrt_(fma_(-z, z, (F) 1));
    F s, c;
    sincos_((F) 2 * K<F>::Pi * sx, s, c);
    return mk3<F>(r * c, r * s, z);
}

/* ------------------------------------------------------------------- phase */
/* PhaseFunction::eval (src/phase/rayleigh.cpp:42-45,68-74, hg.cpp:47-50,78-84, isotropic.cpp:41) */
template <typename F> VP_D const F *phase_table(const DPhase<F> &ph) {
    return reinterpret_cast<const F *>(reinterpret_cast<const unsigned char *>(&ph) + ph.tab_rel);
}
/* ContinuousDistribution::eval_pdf (distr_1d.h:365-380) on [-1, 1]; masked gathers read 0 outside the range */
template <typename F> VP_D F tab_eval_pdf(const DPhase<F> &ph, F x) {
    const F *pdf = phase_table(ph);
    const bool active = x >= (F) -1 && x <= (F) 1;
    x = (x - (F) -1) * ph.inv_interval_size;
    long long xi = (long long) x;
    uint32_t index = (uint32_t) min(max(xi, 0ll), (long long) ph.n - 2);
    F y0 = active ? pdf[index] : (F) 0, y1 = active ? pdf[index + 1] : (F) 0;
    F w1 = x - (F) index, w0 = (F) 1 - w1;
    return fma_(w0, y0, w1 * y1);
}
/* ContinuousDistribution::sample (distr_1d.h:416-439) with enoki::binary_search (array_utils.h:150-171) */
template <typename F> VP_D F tab_sample(const DPhase<F> &ph, F value) {
    const F *pdf = phase_table(ph), *cdf = pdf + ph.n;
    value *= ph.integral;
    uint32_t start = ph.valid_x, end = ph.valid_y;
    uint32_t iterations = start < end ? (32u - (uint32_t) __clz(end - start)) : 0u;   /* log2i(d) + 1 */
    for (uint32_t i = 0; i < iterations; ++i) {
        uint32_t middle = (start + end) >> 1;
        if (cdf[middle] < value) start = min(middle + 1, end); else end = middle;
    }
    const uint32_t index = start;
    F y0 = pdf[index], y1 = pdf[index + 1], c0 = index > 0 ? cdf[index - 1] : (F) 0;
    value = (value - c0) * ph.inv_interval_size;
    F t_linear = (y0 - safe_sqrt_(sqr(y0) + (F) 2 * value * (y1 - y0))) / (y0 - y1), t_const = value / y0;
    F t = y0 == y1 ? t_const : t_linear;
    return fma_((F) index + t, ph.interval_size, (F) -1);
}
/* the phase function as a function of the cosine its eval() computes, dot(wo, wi) (isotropic: a constant) */
template <typename F, uint32_t FEAT = FT_ALL> VP_D F phase_eval_cos(const DPhase<F> &ph, F cos_theta) {
    if (ph.type == 0) return K<F>::InvFourPi;
    if ((FEAT & FT_TABPHASE) && ph.type == 3) /* tabphase.cpp:91-99 */
        return tab_eval_pdf(ph, -cos_theta) * ph.normalization * K<F>::InvTwoPi;
    if (ph.type == 1)
        return ((F) 3 / (F) 16) * K<F>::InvPi * ((F) 1 + sqr(cos_theta));
    F g = ph.g;
    F temp = (F) 1 + g * g + (F) 2 * g * cos_theta;
    return K<F>::InvFourPi * ((F) 1 - g * g) / (temp * sqrt_(temp));
}
template <typename F, uint32_t FEAT = FT_ALL> VP_D F phase_eval(const DPhase<F> &ph, V3<F> wi, V3<F> wo) {
    if (ph.type == 0) return K<F>::InvFourPi;
    return phase_eval_cos<F, FEAT>(ph, dot(wo, wi));
}
/* PhaseFunction::sample (rayleigh.cpp:47-66, hg.cpp:52-76, isotropic.cpp:31-39);
   `d` is the incoming ray direction, the local frame is Frame3f(d) (medium.cpp:42).  `pdf` (volpathmis only): what
   sample() returns with the direction -- the phase function at the LOCAL sampled cosine, eval_*(-cos_theta)
   (rayleigh.cpp:63, hg.cpp:73, tabphase.cpp:79-80), not eval(wo): for a ray direction that is not exactly of unit
   length (a sensor frame rounded to single precision) dot(to_world(wo), wi) is another number */
template <typename F, uint32_t FEAT = FT_ALL> VP_D V3<F> phase_sample(const DPhase<F> &ph, V3<F> d, F s2x, F s2y, F *pdf = nullptr) {
    if (ph.type == 0) {
        if (pdf) *pdf = K<F>::InvFourPi;
        return square_to_uniform_sphere(s2x, s2y);
    }
    F cos_theta;
    if ((FEAT & FT_TABPHASE) && ph.type == 3) { /* tabphase.cpp:68-89 */
        cos_theta = tab_sample(ph, s2x);
    } else if (ph.type == 1) {
        F z = (F) 2 * ((F) 2 * s2x - (F) 1);
        F tmp = sqrt_(sqr(z) + (F) 1);
        F A = cbrt_(z + tmp), B = cbrt_(z - tmp);
        cos_theta = A + B;
    } else {
        F g = ph.g;
        if (abs_(g) < K<F>::Epsilon) {
            cos_theta = (F) 1 - (F) 2 * s2x;
        } else {
            F sqr_term = ((F) 1 - g * g) / ((F) 1 - g + (F) 2 * g * s2x);
            cos_theta = ((F) 1 + g * g - sqr_term * sqr_term) / ((F) 2 * g);
        }
    }
    F sin_theta = safe_sqrt_((F) 1 - cos_theta * cos_theta);
    F s, c;
    sincos_((F) 2 * K<F>::Pi * s2y, s, c);
    V3<F> wo = mk3<F>(sin_theta * c, sin_theta * s, cos_theta);
    /* tabphase: sample() returns eval_pdf_normalized(-cos_theta) (tabphase.cpp:79) where eval() of the same direction,
       whose dot(wo, wi) is -cos_theta, reads eval_pdf_normalized(+cos_theta) (tabphase.cpp:94-96): the reference's
       sampled pdf is the table at the mirrored cosine.  Mirrored here, not corrected */
    if (pdf) *pdf = phase_eval_cos<F, FEAT>(ph, ((FEAT & FT_TABPHASE) && ph.type == 3) ? cos_theta : -cos_theta);
    return make_frame(d).to_world(wo);
}

/* BlendPhaseFunction (src/phase/blendphase.cpp:68-140, default context): (1 - w) * child[0] + w * child[1] with
   w = clamp(weight volume at the scattering position, 0, 1); sampling picks child[0] when sample1 > w.
   Other phase functions pass straight through. */
template <typename F, uint32_t FEAT = FT_ALL>
VP_D F phase_eval_at(const SceneView<F> &sv, const DPhase<F> &ph, V3<F> p, V3<F> wi, V3<F> wo) {
    if (!(FEAT & FT_BLENDPHASE) || ph.type != 4) return phase_eval<F, FEAT>(ph, wi, wo);
    F w = min_(max_(volume_eval<F, FT_ALL, true>(sv, sv.volumes[ph.weight_volume], p), (F) 0), (F) 1);
    return phase_eval<F, FEAT>(sv.phases[ph.child[0]], wi, wo) * ((F) 1 - w) +
           phase_eval<F, FEAT>(sv.phases[ph.child[1]], wi, wo) * w;
}
template <typename F, uint32_t FEAT = FT_ALL>
VP_D V3<F> phase_sample_at(const SceneView<F> &sv, const DPhase<F> &ph, V3<F> p, V3<F> d, F s1, F s2x, F s2y, F *pdf = nullptr) {
    if (!(FEAT & FT_BLENDPHASE) || ph.type != 4) return phase_sample<F, FEAT>(ph, d, s2x, s2y, pdf);
    F w = min_(max_(volume_eval<F, FT_ALL, true>(sv, sv.volumes[ph.weight_volume], p), (F) 0), (F) 1);
    /* the pdf of a blend is that of the nested function that was picked, unweighted (blendphase.cpp:97-112) */
    return phase_sample<F, FEAT>(sv.phases[s1 > w ? ph.child[0] : ph.child[1]], d, s2x, s2y, pdf);
}

/* ------------------------------------------------------------------- BSDFs */
enum : uint32_t { /* BSDFFlags (include/mitsuba/render/bsdf.h:44-123) */
    BF_Null = 0x1, BF_DiffuseReflection = 0x2, BF_DiffuseTransmission = 0x4, BF_GlossyReflection = 0x8,
    BF_Smooth = 0x2 | 0x4 | 0x8 | 0x10, BF_Delta = 0x1 | 0x20 | 0x40
};
VP_D uint32_t bsdf_flags(int type) {
    if (type == 3) return BF_DiffuseReflection | BF_DiffuseTransmission; /* bilambertian.cpp:55-60 */
    return type == 1 ? BF_Null : (type == 0 ? BF_DiffuseReflection : BF_GlossyReflection);
}
/* BiLambertian lobe weights (src/bsdfs/bilambertian.cpp:80-87): r / (r + t), 1 - that, NaN (r = t = 0) -> 0 */
template <typename F> VP_D void bilambertian_weights(const DBsdf<F> &b, F &r, F &t, F &rw, F &tw) {
    r = b.p[0]; t = b.p[1];
    rw = r / (r + t); tw = (F) 1 - rw;
    if (rw != rw) rw = 0;
    if (tw != tw) tw = 0;
}

/* Frame::sincos_phi (include/mitsuba/core/frame.h:107-118) */
template <typename F> VP_D void sincos_phi(V3<F> v, F &s, F &c) {
    F sin_theta_2 = fma_(v.x, v.x, sqr(v.y));
    F inv_sin_theta = rsqrt_(sin_theta_2);
    F rx = v.x * inv_sin_theta, ry = v.y * inv_sin_theta;
    if (abs_(sin_theta_2) <= (F) 4 * K<F>::Epsilon) { rx = 1; ry = 0; }
    else { rx = min_(max_(rx, (F) -1), (F) 1); ry = min_(max_(ry, (F) -1), (F) 1); }
    s = ry;
    c = rx;
}
template <typename F> VP_D F frame_sin_theta(V3<F> v) { return safe_sqrt_(fma_(v.x, v.x, sqr(v.y))); }
template <typename F> VP_D F frame_tan_theta(V3<F> v) { return safe_sqrt_(fma_(-v.z, v.z, (F) 1)) / v.z; }

/* RPV::eval_rpv (src/bsdfs/rpv.cpp:107-141); pow = exp(log(x)*y) as enoki evaluates it */
template <typename F> __device__ __noinline__ F eval_rpv(const DBsdf<F> &b, V3<F> wi, V3<F> wo) {
    F rho_0 = b.p[0], k = b.p[1], g = b.p[2], rho_c = b.p[3];
    F sin_phi1, cos_phi1, sin_phi2, cos_phi2;
    sincos_phi(wi, sin_phi1, cos_phi1);
    sincos_phi(wo, sin_phi2, cos_phi2);
    F cos_dphi = cos_phi1 * cos_phi2 + sin_phi1 * sin_phi2;
    F sin_theta1 = frame_sin_theta(wi), cos_theta1 = wi.z, tan_theta1 = frame_tan_theta(wi);
    F sin_theta2 = frame_sin_theta(wo), cos_theta2 = wo.z, tan_theta2 = frame_tan_theta(wo);
    F G = safe_sqrt_(sqr(tan_theta1) + sqr(tan_theta2) - (F) 2 * tan_theta1 * tan_theta2 * cos_dphi);
    F cos_g = cos_theta1 * cos_theta2 + sin_theta1 * sin_theta2 * cos_dphi;
    F Fg = ((F) 1 - sqr(g)) / pow_(((F) 1 + sqr(g) + (F) 2 * g * cos_g), (F) 1.5);
    return rho_0 * (pow_(cos_theta1 * cos_theta2 * (cos_theta1 + cos_theta2), k - (F) 1) * Fg *
                    ((F) 1 + ((F) 1 - rho_c) / ((F) 1 + G))) * K<F>::InvPi;
}

/* BSDF::sample (src/bsdfs/diffuse.cpp:78-102, rpv.cpp:85-105 [weight = f, F8], null.cpp:41-57) */
template <typename F>
VP_D void bsdf_sample(const DBsdf<F> &b, V3<F> wi, F s1, F s2x, F s2y, V3<F> &wo, F &pdf, F &eta,
                      uint32_t &sampled_type, F &weight) {
    wo = mk3<F>(0, 0, 0); pdf = 0; eta = 0; sampled_type = 0; weight = 0;
    if (b.type == 1) {
        wo = -wi; sampled_type = BF_Null; eta = 1; pdf = 1; weight = 1;
        return;
    }
    if (b.type == 3) { /* bilambertian.cpp:63-119 */
        wo = square_to_cosine_hemisphere(s2x, s2y);
        F r, t, rw, tw; bilambertian_weights(b, r, t, rw, tw);
        bool selected_r = s1 < rw;
        F value = 1;
        if (selected_r) value *= r / rw; else value *= t / tw;
        pdf = K<F>::InvPi * wo.z;
        pdf = selected_r ? pdf * rw : pdf * tw;
        eta = 1;
        sampled_type = selected_r ? BF_DiffuseReflection : BF_DiffuseTransmission;
        if (!(wi.z > 0)) wo.z = -wo.z;   /* the incoming direction comes from "behind" */
        if (!selected_r) wo.z = -wo.z;   /* transmission */
        weight = pdf > 0 ? value : (F) 0;
        return;
    }
    bool active = wi.z > 0;
    if (b.type == 0) {
        if (!active) return; /* zero-initialised sample (diffuse.cpp:90-92) */
        wo = square_to_cosine_hemisphere(s2x, s2y);
        pdf = K<F>::InvPi * wo.z;
        eta = 1; sampled_type = BF_DiffuseReflection;
        weight = (active && pdf > 0) ? b.p[0] : (F) 0;
        return;
    }
    wo = square_to_cosine_hemisphere(s2x, s2y);
    pdf = K<F>::InvPi * wo.z;
    eta = 1; sampled_type = BF_GlossyReflection;
    F value = eval_rpv(b, wi, wo);
    weight = (active && pdf > 0) ? value : (F) 0;
}
/* BSDF::eval (diffuse.cpp:104-120, rpv.cpp:143-155, null.cpp:59-62) */
template <typename F> VP_D F bsdf_eval(const DBsdf<F> &b, V3<F> wi, V3<F> wo) {
    if (b.type == 1) return 0;
    if (b.type == 3) { /* bilambertian.cpp:121-151; sign() is copysign(1, x) */
        F result = signbit_(wi.z) == signbit_(wo.z) ? b.p[0] : b.p[1];
        return result * (K<F>::InvPi * abs_(wo.z));
    }
    bool active = wi.z > 0 && wo.z > 0;
    if (b.type == 0) {
        F value = b.p[0] * K<F>::InvPi * wo.z;
        return active ? value : (F) 0;
    }
    F value = eval_rpv(b, wi, wo);
    return active ? value * abs_(wo.z) : (F) 0;
}
/* BSDF::pdf (diffuse.cpp:122-135, rpv.cpp:157-168) */
template <typename F> VP_D F bsdf_pdf(const DBsdf<F> &b, V3<F> wi, V3<F> wo) {
    if (b.type == 1) return 0;
    if (b.type == 3) { /* bilambertian.cpp:153-193 */
        F r, t, rw, tw; bilambertian_weights(b, r, t, rw, tw);
        F result = K<F>::InvPi * abs_(wo.z);
        return result * (signbit_(wi.z) == signbit_(wo.z) ? rw : tw);
    }
    F pdf = K<F>::InvPi * wo.z;
    return (wi.z > 0 && wo.z > 0) ? pdf : (F) 0;
}

/* ----------------------------------------------------------------- emitter */
template <typename F> struct DirSample { V3<F> d; F dist, pdf; bool delta; };

/* Scene::sample_emitter_direction (src/librender/scene.cpp:168-218) over
   DirectionalEmitter::sample_direction (src/emitters/directional.cpp:108-135) */
template <typename F> VP_D DirSample<F> sample_emitter_direction(const SceneView<F> &sv, F sx, F &spec) {
    DirSample<F> ds;
    ds.d = mk3<F>(0, 0, 0); ds.dist = 0; ds.pdf = 0; ds.delta = false;
    spec = 0;
    int n = sv.g->n_emitters;
    if (n == 0) return ds;
    int index = 0;
    F emitter_pdf = 1;
    if (n > 1) {
        emitter_pdf = (F) (1.f / (float) n);   /* scene.cpp:183: `1.f / m_emitters.size()`, a float quotient in every variant */
        uint32_t idx = (uint32_t) (sx * (F) n);
        index = (int) (idx < (uint32_t) (n - 1) ? idx : (uint32_t) (n - 1));
    }
    const DEmitter<F> &e = sv.emitters[index];
    ds.d = -ld3(e.dir);
    ds.dist = (F) 2 * sv.g->bs_radius;
    ds.pdf = 1;
    ds.delta = true;
    spec = e.irradiance;
    if (n > 1) { ds.pdf *= emitter_pdf; spec *= rcp_(emitter_pdf); }
    return ds;
}

/* The same with an environment emitter in the scene (ConstantBackgroundEmitter::sample_direction,
   src/emitters/constant.cpp:81-104): needs the second sample and the rescaling of the first (scene.cpp:190).
   Used by the one-thread-per-stream kernel, the only one that renders scenes with an environment. */
template <typename F> VP_D DirSample<F> sample_emitter_direction_env(const SceneView<F> &sv, F sx, F sy, F &spec) {
    DirSample<F> ds;
    ds.d = mk3<F>(0, 0, 0); ds.dist = 0; ds.pdf = 0; ds.delta = false;
    spec = 0;
    int n = sv.g->n_emitters;
    if (n == 0) return ds;
    int index = 0;
    F emitter_pdf = 1;
    if (n > 1) {
        emitter_pdf = (F) (1.f / (float) n);   /* scene.cpp:183: `1.f / m_emitters.size()`, a float quotient in every variant */
        uint32_t idx = (uint32_t) (sx * (F) n);
        index = (int) (idx < (uint32_t) (n - 1) ? idx : (uint32_t) (n - 1));
        sx = (sx - (F) index * emitter_pdf) * (F) n;
    }
    const DEmitter<F> &e = sv.emitters[index];
    ds.dist = (F) 2 * sv.g->bs_radius;
    if (e.type == 1) {
        ds.d = square_to_uniform_sphere(sx, sy);
        ds.pdf = K<F>::InvFourPi;
        ds.delta = false;
        spec = e.irradiance / ds.pdf;
    } else {
        ds.d = -ld3(e.dir);
        ds.pdf = 1;
        ds.delta = true;
        spec = e.irradiance;
    }
    if (n > 1) { ds.pdf *= emitter_pdf; spec *= rcp_(emitter_pdf); }
    return ds;
}

/* ------------------------------------------------------------------ sensors */
template <typename F> VP_D F rect_area(const DShape<F> &sh) { return norm(cross(ld3(sh.fs), ld3(sh.ft))); }

/* Sensor::sample_ray_differential (src/sensors/distant.cpp:299-386, mdistant.cpp:216-266,
   radiancemeter.cpp:106-139, mradiancemeter.cpp:134-159); returns the ray weight */
template <typename F, uint32_t FEAT = FT_ALL>
VP_D F sensor_sample_ray(const SceneView<F> &sv, F fx, F fy, F ax, F ay, Ray<F> &ray) {
    const DScene<F> &g = *sv.g;
    const DSensor<F> &sn = g.sensor;
    ray.mint = K<F>::RayEpsilon;
    ray.maxt = inf_<F>();
    F weight = 0;
    V3<F> bs_c = ld3(g.bs_center);
    F bs_r = g.bs_radius;
    if (sn.type == 0) {
        V3<F> v0 = mk3<F>(0, 0, 1);
        if (sn.direction_type == 2)
            v0 = square_to_uniform_hemisphere(fx, fy);
        else if (sn.direction_type == 1) {
            F s, c;
            sincos_(K<F>::Pi * fx, s, c);
            v0.x = c; v0.z = s;
        }
        ray.d = sn.flip_directions ? xf_vector(sn.to_world, v0) : xf_vector(sn.to_world, -v0);
        V3<F> ray_target = ld3(sn.target_point);
        if (sn.target_type == 1) {
            weight = 1;
        } else if (sn.target_type == 2) {
            const DShape<F> &sh = sv.shapes[sn.target_shape]; /* rectangle.cpp:111-125 */
            ray_target = xf_point_affine(sh.to_world, mk3<F>(ax * (F) 2 - (F) 1, ay * (F) 2 - (F) 1, (F) 0));
            weight = (F) 1 / sh.inv_surface_area / rect_area(sh);
        } else {
            F ox, oy;
            square_to_uniform_disk_concentric(ax, ay, ox, oy);
            V3<F> perp = xf_vector(sn.to_world, mk3<F>(ox, oy, (F) 0));
            ray_target = bs_c + perp * bs_r;
            weight = (F) 1;
            weight /= dot(-ray.d, mk3<F>(0, 0, 1));
        }
        if ((FEAT & FT_DISTANTFLUX) && sn.origin_type == 1) {
            bool valid;
            ray.o = origin_on_shape(sv.shapes[sn.origin_shape], ray_target, ray.d, valid);
            if (!valid) weight = 0;
        } else if (sn.target_type == 0)
            ray.o = ray_target - ray.d * bs_r;
        else
            ray.o = ray_target - ray.d * (F) 2 * bs_r;
    } else if (sn.type == 1) {
        int idx = (int) (fx * (F) sn.n_transforms);
        const F *trafo = g.sensor_xf + 16 * (size_t) idx;
        ray.d = xf_vector(trafo, mk3<F>(0, 0, 1));
        if (sn.target_type == 1) {
            ray.o = ld3(sn.target_point) - (F) 2 * ray.d * bs_r;
            weight = 1;
        } else if (sn.target_type == 2) {
            const DShape<F> &sh = sv.shapes[sn.target_shape];
            V3<F> p = xf_point_affine(sh.to_world, mk3<F>(ax * (F) 2 - (F) 1, ay * (F) 2 - (F) 1, (F) 0));
            ray.o = p - (F) 2 * ray.d * bs_r;
            weight = (F) 1 / (sh.inv_surface_area * rect_area(sh));
        } else {
            F ox, oy;
            square_to_uniform_disk_concentric(ax, ay, ox, oy);
            V3<F> perp = xf_vector(trafo, mk3<F>(ox, oy, (F) 0));
            ray.o = bs_c + perp * bs_r - ray.d * bs_r;
            weight = 1;
        }
    } else if ((FEAT & FT_DISTANTFLUX) && sn.type == 4) { /* distantflux.cpp:184-251 */
        ray.d = -xf_vector(sn.to_world, square_to_uniform_hemisphere(fx, fy));
        /* square_to_uniform_hemisphere_pdf is the constant 1 / (2 pi) (warp.h:305-312) */
        weight = dot(-ray.d, ld3(sn.reference_normal)) / (K<F>::InvTwoPi * (F) sn.n_pixels);
        V3<F> ray_target = ld3(sn.target_point);
        if (sn.target_type == 2) {
            const DShape<F> &sh = sv.shapes[sn.target_shape];
            ray_target = xf_point_affine(sh.to_world, mk3<F>(ax * (F) 2 - (F) 1, ay * (F) 2 - (F) 1, (F) 0));
            weight *= (F) 1 / (sh.inv_surface_area * rect_area(sh));
        } else if (sn.target_type == 0) {
            F ox, oy;
            square_to_uniform_disk_concentric(ax, ay, ox, oy);
            V3<F> perp = xf_vector(sn.to_world, mk3<F>(ox, oy, (F) 0));
            ray_target = bs_c + perp * bs_r;
        }
        if (sn.origin_type == 1) {
            bool valid;
            ray.o = origin_on_shape(sv.shapes[sn.origin_shape], ray_target, ray.d, valid);
            if (!valid) weight = 0;
        } else
            ray.o = ray_target - ray.d * (F) 2 * bs_r;
    } else if ((FEAT & FT_DISTANTFLUX) && sn.type == 5) { /* irradiancemeter.cpp:88-118 over rectangle.cpp:111-125 */
        const DShape<F> &sh = sv.shapes[sn.origin_shape];
        ray.o = xf_point_affine(sh.to_world, mk3<F>(fx * (F) 2 - (F) 1, fy * (F) 2 - (F) 1, (F) 0));
        V3<F> local = square_to_cosine_hemisphere(ax, ay);
        ray.d = make_frame(ld3(sh.fn)).to_world(local);
        weight = K<F>::Pi;
    } else if (sn.type == 2) {
        ray.o = xf_point_affine(sn.to_world, mk3<F>(0, 0, 0));
        ray.d = xf_vector(sn.to_world, mk3<F>(0, 0, 1));
        weight = 1;
    } else {
        int idx = (int) (fx * (F) sn.n_transforms);
        const F *trafo = g.sensor_xf + 16 * (size_t) idx;
        ray.o = xf_point_affine(trafo, mk3<F>(0, 0, 0));
        ray.d = xf_vector(trafo, mk3<F>(0, 0, 1));
        weight = 1;
    }
    ray.update();
    return weight;
}

} // namespace VP_NS
