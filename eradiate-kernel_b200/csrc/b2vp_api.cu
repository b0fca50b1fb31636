/*
 * b2vp_api.cu -- the C-ABI of include/b2vp.h on top of the CUDA kernels.
 *
 * There is no CPU fallback: every entry point that computes needs an sm_100
 * device and fails with B2VP_ERR_NO_DEVICE / B2VP_ERR_CUDA otherwise.
 */
#include "../../include/b2vp.h"
#include "bvh_build.h"
#include "vp_queue.cuh"
#include "vp_launch.h"

#include <cstdlib>
#include <dlfcn.h>
#include <nccl.h>     /* types and prototypes only: the library is loaded with dlopen at the first multi-device render */

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <chrono>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

using namespace vp;

thread_local std::string g_error;

int fail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_error = buf;
    return code;
}

#define CU_CHECK(expr)                                                                                       \
    do {                                                                                                     \
        cudaError_t e_ = (expr);                                                                             \
        if (e_ != cudaSuccess)                                                                               \
            return fail(e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver ? B2VP_ERR_NO_DEVICE    \
                                                                                     : B2VP_ERR_CUDA,        \
                        "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__);         \
    } while (0)

/* RAII device buffer */
/* Freed device blocks are kept for the next scene instead of going back to the driver: cudaFree
   synchronises the device and costs 1-30 ms, which is most of the end-to-end overhead of a short render.
   Blocks are reused only for requests of exactly the same (256-byte rounded) size on the same device; every
   user synchronises its stream before releasing, so a cached block is idle.  At most CACHE_MAX bytes stay
   cached per process. */
class DeviceCache {
public:
    void *take(size_t n) {
        int dev = 0;
        cudaGetDevice(&dev);
        std::lock_guard<std::mutex> lock(m_);
        for (size_t i = 0; i < blocks_.size(); ++i)
            if (blocks_[i].bytes == n && blocks_[i].device == dev) {
                void *p = blocks_[i].p;
                total_ -= n;
                blocks_.erase(blocks_.begin() + (long) i);
                return p;
            }
        return nullptr;
    }
    void put(void *p, size_t n) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, p) == cudaSuccess) dev = at.device; else cudaGetLastError();
        std::lock_guard<std::mutex> lock(m_);
        if (n > CACHE_MAX / 2 || total_ + n > CACHE_MAX) { cudaFree(p); return; }
        blocks_.push_back({p, n, dev});
        total_ += n;
    }
    void trim() {
        std::lock_guard<std::mutex> lock(m_);
        for (auto &b : blocks_) cudaFree(b.p);
        blocks_.clear();
        total_ = 0;
    }
private:
    static constexpr size_t CACHE_MAX = (size_t) 1 << 30;
    struct Block { void *p; size_t bytes; int device; };
    std::mutex m_;
    std::vector<Block> blocks_;
    size_t total_ = 0;
};
inline DeviceCache &device_cache() { static DeviceCache *c = new DeviceCache(); return *c; } /* never destroyed: outlives the CUDA context teardown order */

struct DevBuf {
    void *p = nullptr;
    size_t bytes = 0;
    DevBuf() = default;
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    ~DevBuf() { release(); }
    void release() { if (p) device_cache().put(p, bytes); p = nullptr; bytes = 0; }
    cudaError_t alloc(size_t n) {
        release();
        if (n == 0) n = 16;
        n = (n + 255) & ~(size_t) 255;
        p = device_cache().take(n);
        if (p) { bytes = n; return cudaSuccess; }
        cudaError_t e = cudaMalloc(&p, n);
        if (e != cudaSuccess) { /* give cached blocks back to the driver and retry once */
            cudaGetLastError();
            device_cache().trim();
            e = cudaMalloc(&p, n);
        }
        if (e == cudaSuccess) bytes = n; else p = nullptr;
        return e;
    }
    cudaError_t upload(const void *src, size_t n, size_t padded = 0) {
        cudaError_t e = alloc(padded > n ? padded : n);
        if (e != cudaSuccess) return e;
        if (padded > n) { e = cudaMemset(p, 0, padded); if (e != cudaSuccess) return e; }
        return n ? cudaMemcpy(p, src, n, cudaMemcpyHostToDevice) : cudaSuccess;
    }
    template <typename T> T *as() const { return (T *) p; }
};

/* Device copy of a scene in one arithmetic precision. */
template <typename F> struct DeviceScene {
    DScene<F> g{};
    DevBuf blob, vpos, vnrm, vuv, faces, bvh, bvh_prims, bvh_tris, sensor_xf;
    std::vector<std::unique_ptr<DevBuf>> grids;
    std::vector<std::unique_ptr<DevBuf>> cells;   /* trilinear cell records of the grids gathered from HBM */
    size_t cell_bytes = 0;
    std::vector<unsigned char> host_blob;  /* kept for scene_update */
    uint32_t smem_static = 0;              /* blob + staged grids */
    size_t global_grid_bytes = 0;          /* grids gathered from HBM / L2 (not staged in shared memory) */
    DevBuf mg;                             /* majorant grids of all eligible media, floats (B2vpRenderArgs::majorant_grid) */
    uint32_t mg_bytes = 0, mg_request = 0; /* size, and the request (resolution) it was built for; 0 = not built */
    bool ready = false;
};

} // namespace

/* Everything that lives on ONE GPU: the device copies of the scene (one per arithmetic precision, made on first use)
   and the buffers of a render.  A scene has one context per device it has rendered on; contexts[0] is the device the
   scene was created for and the root of multi-device renders. */
struct DeviceCtx {
    int device = 0;
    int sm_count = 0;
    size_t smem_optin = 0;
    size_t l2_window_max = 0;   /* largest access-policy window (0: no persisting L2 on this device) */
    size_t l2_persist_max = 0;   /* largest persisting L2 carve-out the device allows */
    size_t l2_persist_bytes = 0; /* the carve-out this library has set (cudaLimitPersistingL2CacheSize); 0 = none yet */
    DeviceScene<float> dev32;
    DeviceScene<double> dev64;
    DevBuf film_buf, totals_buf, counter_buf, blocks_buf, stop_flag;
    DevBuf ex_state, ex_inc, ex_pixel, o_state, o_counters, o_radiance, q_cold;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_done = nullptr;
    cudaStream_t ctl_stream = nullptr;    /* raises stop_flag while a render kernel runs on another stream */
    cudaStream_t work_stream = nullptr;   /* render stream of a non-root device in a multi-device render */
};

struct b2vp_scene {
    int device = 0;
    B2vpSceneDesc desc{};
    std::vector<B2vpShape> shapes;
    std::vector<B2vpBsdf> bsdfs;
    std::vector<B2vpMedium> media;
    std::vector<B2vpVolume> volumes;
    std::vector<std::vector<float>> volume_data;
    std::vector<char> volume_pinned;                 /* volume_data[i] is page-locked (b2vp_scene_update_grid) */
    std::vector<B2vpPhase> phases;
    std::vector<std::vector<double>> phase_values;   /* tabphase tables, copied from the descriptor */
    std::vector<B2vpEmitter> emitters;
    std::vector<float> vpos, vnrm, vuv;
    std::vector<uint32_t> faces;
    std::vector<double> sensor_xf;
    std::vector<BvhNode> bvh_nodes;
    std::vector<BvhPrim> bvh_prims;
    BvhStats bvh_stats;
    std::string last_kernel;                         /* kernel instance launched by the last render (b2vp_scene_info) */
    std::vector<std::unique_ptr<DeviceCtx>> contexts;
    std::mutex mutex;                                /* guards contexts / last_kernel between the threads of a multi-device render */
    std::atomic<int> cancelled{ 0 };
};

namespace {

/* ---- validation + deep copy --------------------------------------------- */
int copy_desc(const B2vpSceneDesc *d, b2vp_scene *s) {
    s->desc = *d;
    s->shapes.assign(d->shapes, d->shapes + d->n_shapes);
    s->bsdfs.assign(d->bsdfs, d->bsdfs + d->n_bsdfs);
    s->media.assign(d->media, d->media + d->n_media);
    s->volumes.assign(d->volumes, d->volumes + d->n_volumes);
    s->phases.assign(d->phases, d->phases + d->n_phases);
    s->phase_values.resize(d->n_phases);
    for (uint32_t i = 0; i < d->n_phases; ++i)
        if (d->phases[i].type == B2VP_PHASE_TABULATED && d->phases[i].values) {
            s->phase_values[i].assign(d->phases[i].values, d->phases[i].values + d->phases[i].n_values);
            s->phases[i].values = nullptr;
        }
    s->emitters.assign(d->emitters, d->emitters + d->n_emitters);
    s->volume_data.resize(d->n_volumes);
    for (uint32_t i = 0; i < d->n_volumes; ++i) {
        const B2vpVolume &v = d->volumes[i];
        if (v.type == B2VP_VOLUME_GRID) {
            if (!v.data || v.res[0] <= 0 || v.res[1] <= 0 || v.res[2] <= 0 || (v.channels != 1 && v.channels != 3))
                return fail(B2VP_ERR_INVALID, "volume %u: bad grid (res/channels/data)", i);
            size_t n = (size_t) v.res[0] * v.res[1] * v.res[2] * v.channels;
            s->volume_data[i].assign(v.data, v.data + n);
        } else if (v.type != B2VP_VOLUME_CONST)
            return fail(B2VP_ERR_UNSUPPORTED, "volume %u: type %d not supported by the B200 volpath", i, v.type);
        s->volumes[i].data = nullptr;
    }
    for (uint32_t i = 0; i < d->n_media; ++i) {
        const B2vpMedium &m = d->media[i];
        if (m.sigma_t < 0 || m.sigma_t >= (int) d->n_volumes || m.albedo < 0 || m.albedo >= (int) d->n_volumes ||
            m.phase < 0 || m.phase >= (int) d->n_phases)
            return fail(B2VP_ERR_INVALID, "medium %u: volume/phase index out of range", i);
        if (m.type == B2VP_MEDIUM_HETEROGENEOUS && d->volumes[m.sigma_t].type != B2VP_VOLUME_GRID)
            return fail(B2VP_ERR_UNSUPPORTED,
                        "heterogeneous medium needs a gridvolume sigma_t (ConstVolume::max is not implemented, "
                        "constant3d.cpp:39)");
        if (m.type != B2VP_MEDIUM_HOMOGENEOUS && m.type != B2VP_MEDIUM_HETEROGENEOUS)
            return fail(B2VP_ERR_UNSUPPORTED, "medium %u: type %d not supported by the B200 volpath", i, m.type);
    }
    for (uint32_t i = 0; i < d->n_shapes; ++i) {
        const B2vpShape &a = d->shapes[i];
        if (a.type < 0 || a.type > B2VP_SHAPE_MESH)
            return fail(B2VP_ERR_UNSUPPORTED, "shape %u: type %d not supported by the B200 volpath", i, a.type);
        if (a.bsdf < 0 || a.bsdf >= (int) d->n_bsdfs || a.interior_medium >= (int) d->n_media ||
            a.exterior_medium >= (int) d->n_media)
            return fail(B2VP_ERR_INVALID, "shape %u: bsdf/medium index out of range", i);
        const bool is_target = d->sensor.target_type == B2VP_TARGET_SHAPE && d->sensor.target_shape == (int) i;
        const bool is_origin = d->sensor.origin_type == B2VP_ORIGIN_SHAPE && d->sensor.origin_shape == (int) i;
        if (a.sample_only && (a.type != B2VP_SHAPE_RECTANGLE || !(is_target || is_origin)))
            return fail(B2VP_ERR_INVALID, "shape %u: sample_only is for the rectangle the sensor targets or starts its rays on", i);
        if (a.type == B2VP_SHAPE_MESH &&
            ((uint64_t) a.vertex_offset + a.vertex_count > d->n_vertices ||
             (uint64_t) a.face_offset + a.face_count > d->n_faces))
            return fail(B2VP_ERR_INVALID, "shape %u: vertex/face range out of the pools", i);
    }
    if (d->n_faces >= (1ull << 28))
        return fail(B2VP_ERR_UNSUPPORTED, "more than 2^28 triangles (the BVH node reference holds 29 bits)");
    for (uint32_t i = 0; i < d->n_bsdfs; ++i)
        if (d->bsdfs[i].type < 0 || d->bsdfs[i].type > B2VP_BSDF_BILAMBERTIAN)
            return fail(B2VP_ERR_UNSUPPORTED, "bsdf %u: type %d not supported by the B200 volpath", i, d->bsdfs[i].type);
    for (uint32_t i = 0; i < d->n_phases; ++i)
        if (d->phases[i].type < 0 || d->phases[i].type > B2VP_PHASE_BLEND)
            return fail(B2VP_ERR_UNSUPPORTED, "phase %u: type %d not supported by the B200 volpath", i, d->phases[i].type);
    for (uint32_t i = 0; i < d->n_phases; ++i)
        if (d->phases[i].type == B2VP_PHASE_TABULATED && s->phase_values[i].size() < 2)
            return fail(B2VP_ERR_INVALID, "phase %u: ContinuousDistribution: needs at least two entries!", i);
    for (uint32_t i = 0; i < d->n_phases; ++i) {
        const B2vpPhase &p = d->phases[i];
        if (p.type != B2VP_PHASE_BLEND) continue;
        for (int k = 0; k < 2; ++k)
            if (p.child[k] < 0 || p.child[k] >= (int) d->n_phases || d->phases[p.child[k]].type == B2VP_PHASE_BLEND)
                return fail(B2VP_ERR_INVALID, "phase %u: BlendPhase: Two child phase functions must be specified!", i);
        if (p.weight_volume < 0 || p.weight_volume >= (int) d->n_volumes)
            return fail(B2VP_ERR_INVALID, "phase %u: blendphase weight volume out of range", i);
    }
    for (uint32_t i = 0; i < d->n_emitters; ++i)
        if (d->emitters[i].type != B2VP_EMITTER_DIRECTIONAL && d->emitters[i].type != B2VP_EMITTER_CONSTANT)
            return fail(B2VP_ERR_UNSUPPORTED, "emitter %u: only `directional` and `constant` are supported by the B200 volpath", i);
    {
        int n_env = 0;
        for (uint32_t i = 0; i < d->n_emitters; ++i) n_env += d->emitters[i].type == B2VP_EMITTER_CONSTANT;
        if (n_env > 1) return fail(B2VP_ERR_INVALID, "Only one environment emitter can be specified per scene.");
    }
    const B2vpSensor &sn = d->sensor;
    if (sn.type < 0 || sn.type > B2VP_SENSOR_IRRADIANCEMETER)
        return fail(B2VP_ERR_UNSUPPORTED, "sensor type %d not supported by the B200 volpath", sn.type);
    if (sn.origin_type == B2VP_ORIGIN_SHAPE) {   /* distant.cpp:281-290, distantflux.cpp:171-180 */
        if (sn.type != B2VP_SENSOR_DISTANT && sn.type != B2VP_SENSOR_DISTANTFLUX)
            return fail(B2VP_ERR_INVALID, "only distant and distantflux sensors take their ray origins from a shape");
        if (sn.origin_shape < 0 || sn.origin_shape >= (int) d->n_shapes || d->shapes[sn.origin_shape].type != B2VP_SHAPE_RECTANGLE)
            return fail(B2VP_ERR_UNSUPPORTED, "sensor ray_origin shape must be a rectangle");
    }
    if ((sn.type == B2VP_SENSOR_DISTANT || sn.type == B2VP_SENSOR_MDISTANT || sn.type == B2VP_SENSOR_DISTANTFLUX) &&
        sn.target_type == B2VP_TARGET_SHAPE &&
        (sn.target_shape < 0 || sn.target_shape >= (int) d->n_shapes ||
         d->shapes[sn.target_shape].type != B2VP_SHAPE_RECTANGLE))
        return fail(B2VP_ERR_UNSUPPORTED, "sensor ray_target shape must be a rectangle");
    if (sn.type == B2VP_SENSOR_IRRADIANCEMETER &&
        (sn.origin_shape < 0 || sn.origin_shape >= (int) d->n_shapes || d->shapes[sn.origin_shape].type != B2VP_SHAPE_RECTANGLE))
        return fail(B2VP_ERR_UNSUPPORTED, "irradiancemeter: the parent shape must be a rectangle");
    if ((sn.type == B2VP_SENSOR_MDISTANT || sn.type == B2VP_SENSOR_MRADIANCEMETER) &&
        (sn.n_transforms <= 0 || !sn.transforms))
        return fail(B2VP_ERR_INVALID, "multi-sensor without transforms");
    if (sn.medium >= (int) d->n_media)
        return fail(B2VP_ERR_INVALID, "sensor medium index out of range");
    if (sn.crop_size[0] <= 0 || sn.crop_size[1] <= 0)
        return fail(B2VP_ERR_INVALID, "empty film");
    if (sn.rfilter_radius > 0.5) { /* possibly wider than a pixel (the device decides in the scene's precision): the table may be read */
        if (!(sn.rfilter_radius <= 0.5 * B2VP_RFILTER_MAX_TAPS) || sn.rfilter_taps == 0 ||
            sn.rfilter_taps > B2VP_RFILTER_MAX_TAPS || !(sn.rfilter_scale > 0))
            return fail(B2VP_ERR_INVALID, "reconstruction filter: radius %g with %u taps (at most %d taps)",
                        sn.rfilter_radius, sn.rfilter_taps, B2VP_RFILTER_MAX_TAPS);
        if ((double) sn.rfilter_border + 0.5 < sn.rfilter_radius - 1e-3)
            return fail(B2VP_ERR_INVALID, "reconstruction filter: border %u too small for radius %g", sn.rfilter_border,
                        sn.rfilter_radius);
        for (int k = 0; k <= B2VP_RFILTER_RESOLUTION; ++k)
            if (!std::isfinite(sn.rfilter_values[k]))
                return fail(B2VP_ERR_INVALID, "reconstruction filter: table entry %d is not finite", k);
    }
    if (sn.n_transforms > 0 && sn.transforms)
        s->sensor_xf.assign(sn.transforms, sn.transforms + (size_t) sn.n_transforms * 16);
    s->desc.sensor.transforms = nullptr;
    if (d->n_vertices) {
        s->vpos.assign(d->vertex_positions, d->vertex_positions + (size_t) d->n_vertices * 3);
        if (d->vertex_normals) s->vnrm.assign(d->vertex_normals, d->vertex_normals + (size_t) d->n_vertices * 3);
        if (d->vertex_texcoords) s->vuv.assign(d->vertex_texcoords, d->vertex_texcoords + (size_t) d->n_vertices * 2);
    }
    if (d->n_faces) s->faces.assign(d->faces, d->faces + (size_t) d->n_faces * 3);
    for (uint32_t i = 0; i < d->n_shapes; ++i) {
        const B2vpShape &a = d->shapes[i];
        if (a.type != B2VP_SHAPE_MESH) continue;
        if ((a.has_vertex_normals && s->vnrm.empty()) || (a.has_vertex_texcoords && s->vuv.empty()))
            return fail(B2VP_ERR_INVALID, "shape %u declares vertex normals/texcoords but the pool is NULL", i);
        for (uint32_t f = 0; f < a.face_count; ++f)
            for (int k = 0; k < 3; ++k)
                if (s->faces[3 * (size_t) (a.face_offset + f) + k] >= a.vertex_count)
                    return fail(B2VP_ERR_INVALID, "shape %u face %u: vertex index out of range", i, f);
    }
    s->desc.shapes = nullptr; s->desc.bsdfs = nullptr; s->desc.media = nullptr; s->desc.volumes = nullptr;
    s->desc.phases = nullptr; s->desc.emitters = nullptr;
    s->desc.vertex_positions = s->desc.vertex_normals = s->desc.vertex_texcoords = nullptr;
    s->desc.faces = nullptr;
    return B2VP_OK;
}

int build_bvh(b2vp_scene *s) {
    std::vector<BuildTri> tris;
    for (size_t i = 0; i < s->shapes.size(); ++i) {
        const B2vpShape &a = s->shapes[i];
        if (a.type != B2VP_SHAPE_MESH) continue;
        for (uint32_t f = 0; f < a.face_count; ++f) {
            BuildTri t;
            t.shape = (uint32_t) i; t.face = f;
            for (int k = 0; k < 3; ++k) { t.bmin[k] = INFINITY; t.bmax[k] = -INFINITY; }
            for (int v = 0; v < 3; ++v) {
                const float *p = &s->vpos[3 * (size_t) (a.vertex_offset + s->faces[3 * (size_t) (a.face_offset + f) + v])];
                for (int k = 0; k < 3; ++k) { t.bmin[k] = std::min(t.bmin[k], p[k]); t.bmax[k] = std::max(t.bmax[k], p[k]); }
            }
            for (int k = 0; k < 3; ++k) t.c[k] = 0.5f * (t.bmin[k] + t.bmax[k]);
            tris.push_back(t);
        }
    }
    /* pad: ~16 float ulps of the largest coordinate a ray origin can have
       (sensors and emitters sit up to 2 bounding-sphere radii from the centre) */
    double scale = 0;
    for (int k = 0; k < 3; ++k)
        scale = std::max(scale, std::max(std::fabs(s->desc.bbox_min[k]), std::fabs(s->desc.bbox_max[k])));
    scale += 3.0 * s->desc.bsphere_radius;
    float pad = (float) (scale * 2e-6);
    int max_leaf = 2;   /* measured: 2-3 triangles per leaf beat 4 and 1 on both the 12-triangle cube and the 1e6-triangle canopy */
    if (const uint32_t v = (s->desc.options >> B2VP_OPT_BVH_LEAF_SHIFT) & 7u) max_leaf = (int) v;
    bvh_build(tris, pad, s->bvh_nodes, s->bvh_prims, max_leaf, &s->bvh_stats);
    /* the device traversal pushes at most one node per level onto a BVH_STACK_CAPACITY-entry stack; the builder
       bounds the depth by 24 + log2(n) (index-median splits from level 24 on), n < 2^28 */
    if (s->bvh_stats.max_depth >= BVH_STACK_CAPACITY)
        return fail(B2VP_ERR_UNSUPPORTED, "BVH depth %u exceeds the device traversal stack (%u entries)",
                    s->bvh_stats.max_depth, BVH_STACK_CAPACITY);
    return B2VP_OK;
}

} // namespace

namespace {

/* The context of `device`, created on first use (device attributes, events, control stream). */
int noalloc_scratch(int device, unsigned char **out);   /* defined with b2vp_render_noalloc below */

int context_for(b2vp_scene *s, int device, DeviceCtx **out) {
    std::lock_guard<std::mutex> lock(s->mutex);
    for (auto &c : s->contexts)
        if (c->device == device) { *out = c.get(); return B2VP_OK; }
    int n_dev = 0;
    CU_CHECK(cudaGetDeviceCount(&n_dev));
    if (device < 0 || device >= n_dev) return fail(B2VP_ERR_INVALID, "device ordinal %d out of range", device);
    /* single attributes, not cudaGetDeviceProperties: that call costs milliseconds on every scene */
    int cc_major = 0, cc_minor = 0, sm_count = 0, smem_optin = 0;
    CU_CHECK(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, device));
    CU_CHECK(cudaDeviceGetAttribute(&cc_minor, cudaDevAttrComputeCapabilityMinor, device));
    CU_CHECK(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device));
    CU_CHECK(cudaDeviceGetAttribute(&smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    if (cc_major != 10)
        return fail(B2VP_ERR_NO_DEVICE, "device %d is sm_%d%d; this library is built for sm_100a only", device,
                    cc_major, cc_minor);
    std::unique_ptr<DeviceCtx> c(new DeviceCtx());
    c->device = device; c->sm_count = sm_count; c->smem_optin = (size_t) smem_optin;
    CU_CHECK(cudaSetDevice(device));
    {
        int l2_persist_max = 0, window_max = 0;
        CU_CHECK(cudaDeviceGetAttribute(&l2_persist_max, cudaDevAttrMaxPersistingL2CacheSize, device));
        CU_CHECK(cudaDeviceGetAttribute(&window_max, cudaDevAttrMaxAccessPolicyWindowSize, device));
        if (l2_persist_max > 0 && window_max > 0) { c->l2_window_max = (size_t) window_max; c->l2_persist_max = (size_t) l2_persist_max; }
        (void) cudaGetLastError();
    }
    CU_CHECK(cudaEventCreate(&c->ev0));
    CU_CHECK(cudaEventCreate(&c->ev1));
    CU_CHECK(cudaEventCreateWithFlags(&c->ev_done, cudaEventDisableTiming));
    CU_CHECK(cudaStreamCreateWithFlags(&c->ctl_stream, cudaStreamNonBlocking));
    CU_CHECK(cudaStreamCreateWithFlags(&c->work_stream, cudaStreamNonBlocking));
    CU_CHECK(c->stop_flag.alloc(256));
    CU_CHECK(cudaMemset(c->stop_flag.p, 0, 256));
    /* The static scratch array of b2vp_render_noalloc is part of this library's device image; with CUDA's lazy module
       loading it would be instantiated (64 MiB) by the first call that asks for its address.  Ask here, at set-up, so
       that b2vp_render_noalloc itself never makes the driver allocate (tests/test_zz_no_alloc.py). */
    {
        unsigned char *scratch = nullptr;
        int rc = noalloc_scratch(device, &scratch);
        if (rc != B2VP_OK) return rc;
    }
    *out = c.get();
    s->contexts.push_back(std::move(c));
    return B2VP_OK;
}

/* ---- device upload in precision F ---------------------------------------- */
template <typename F> void cvt(const double *src, F *dst, int n) { for (int i = 0; i < n; ++i) dst[i] = (F) src[i]; }

inline uint32_t align16(uint32_t x) { return (x + 15u) & ~15u; }

/* shared-memory budget that still lets two CTAs share an SM */
constexpr uint32_t SMEM_BUDGET = 100 * 1024;
constexpr uint32_t FILM_SMEM_MAX = 16 * 1024;
constexpr size_t BVH_SMEM_MAX = 16 * 1024;     /* BVH nodes + leaf triangles staged with the tables when they fit */

template <typename F> int upload_scene(b2vp_scene *s, DeviceCtx *dc, DeviceScene<F> &ds) {
    const B2vpSceneDesc &d = s->desc;
    DScene<F> &g = ds.g;
    std::memset(&g, 0, sizeof(g));
    const uint32_t n_shapes = (uint32_t) s->shapes.size();

    std::vector<DShape<F>> shapes(n_shapes);
    std::vector<int> analytic;
    for (uint32_t i = 0; i < n_shapes; ++i) {
        const B2vpShape &a = s->shapes[i];
        DShape<F> &o = shapes[i];
        std::memset(&o, 0, sizeof(o));
        o.type = a.type; o.bsdf = a.bsdf; o.interior = a.interior_medium; o.exterior = a.exterior_medium;
        cvt(a.to_world.m, o.to_world, 16); cvt(a.to_object.m, o.to_object, 16);
        cvt(a.frame_s, o.fs, 3); cvt(a.frame_t, o.ft, 3); cvt(a.frame_n, o.fn, 3);
        for (int k = 0; k < 3; ++k) o.translation[k] = o.to_world[12 + k];
        o.inv_surface_area = (F) a.inv_surface_area;
        cvt(a.center, o.center, 3);
        o.radius = (F) a.radius; o.flip_normals = a.flip_normals;
        o.vofs = a.vertex_offset; o.fofs = a.face_offset; o.fcount = a.face_count;
        o.has_normals = a.has_vertex_normals; o.has_uv = a.has_vertex_texcoords;
        if (a.type != B2VP_SHAPE_MESH && !a.sample_only) analytic.push_back((int) i);   /* a sensor's own target shape is sampled, never hit */
    }
    std::vector<DBsdf<F>> bsdfs(s->bsdfs.size());
    for (size_t i = 0; i < bsdfs.size(); ++i) {
        bsdfs[i].type = s->bsdfs[i].type; bsdfs[i].pad = 0;
        cvt(s->bsdfs[i].p, bsdfs[i].p, 4);
    }
    /* phase functions; tabulated ones get their ContinuousDistribution (distr_1d.h:291-336: running integral in
       double) appended to the phase table: records first, then pdf[n] + cdf[n - 1] per table */
    std::vector<DPhase<F>> phases(s->phases.size());
    std::vector<F> phase_tabs;
    bool have_tabphase = false, have_blendphase = false;
    for (size_t i = 0; i < phases.size(); ++i) {
        std::memset(&phases[i], 0, sizeof(DPhase<F>));
        phases[i].type = s->phases[i].type; phases[i].g = (F) s->phases[i].g;
        if (s->phases[i].type == B2VP_PHASE_BLEND) {
            have_blendphase = true;
            phases[i].child[0] = s->phases[i].child[0]; phases[i].child[1] = s->phases[i].child[1];
            phases[i].weight_volume = s->phases[i].weight_volume;
        }
        if (s->phases[i].type != B2VP_PHASE_TABULATED) continue;
        have_tabphase = true;
        const std::vector<double> &vals = s->phase_values[i];
        const uint32_t n = (uint32_t) vals.size();
        DPhase<F> &o = phases[i];
        o.n = (int) n;
        o.tab_rel = (int) ((phases.size() - i) * sizeof(DPhase<F>) + phase_tabs.size() * sizeof(F));
        o.valid_x = o.valid_y = 0xffffffffu;
        const size_t base = phase_tabs.size();
        phase_tabs.resize(base + 2 * (size_t) n - 1);
        for (uint32_t k = 0; k < n; ++k) phase_tabs[base + k] = (F) vals[k];
        double isz = 2.0 / (n - 1), integ = 0.;
        for (uint32_t k = 0; k + 1 < n; ++k) {
            double y0 = (double) phase_tabs[base + k], y1 = (double) phase_tabs[base + k + 1];
            double value = 0.5 * isz * (y0 + y1);
            integ += value;
            phase_tabs[base + n + k] = (F) integ;
            if (value > 0.) { if (o.valid_x == 0xffffffffu) o.valid_x = k; o.valid_y = k; }
        }
        if (o.valid_x == 0xffffffffu) return fail(B2VP_ERR_INVALID, "ContinuousDistribution: no probability mass found!");
        o.integral = (F) integ; o.normalization = (F) (1. / integ);
        o.interval_size = (F) isz; o.inv_interval_size = (F) (1. / isz);
    }
    std::vector<DEmitter<F>> emitters(s->emitters.size());
    for (size_t i = 0; i < emitters.size(); ++i) {
        cvt(s->emitters[i].direction, emitters[i].dir, 3);
        emitters[i].irradiance = (F) s->emitters[i].irradiance;
        emitters[i].type = s->emitters[i].type; emitters[i].pad[0] = emitters[i].pad[1] = emitters[i].pad[2] = 0;
    }

    /* volumes: grids are widened to F at load like the reference (volume_data.h:85-90) */
    std::vector<DVolume<F>> volumes(s->volumes.size());
    ds.grids.clear();
    ds.cells.clear();
    ds.cell_bytes = 0;
    ds.mg_request = 0; ds.mg_bytes = 0;
    for (size_t i = 0; i < volumes.size(); ++i) {
        const B2vpVolume &a = s->volumes[i];
        DVolume<F> &o = volumes[i];
        std::memset(&o, 0, sizeof(o));
        o.type = a.type; o.filter = a.filter; o.wrap = a.wrap; o.channels = a.channels;
        o.nx = a.res[0]; o.ny = a.res[1]; o.nz = a.res[2]; o.smem_offset = -1;
        o.value = (F) a.value; o.max_value = (F) a.max_value;
        cvt(a.bbox_min, o.bmin, 3); cvt(a.bbox_max, o.bmax, 3);
        cvt(a.world_to_local.m, o.w2l, 16);
        ds.grids.emplace_back(new DevBuf());
        if (a.type == B2VP_VOLUME_GRID) {
            const std::vector<float> &src = s->volume_data[i];
            size_t bytes = src.size() * sizeof(F), padded = (bytes + 15) & ~(size_t) 15;
            cudaError_t e;
            if (sizeof(F) == sizeof(float)) {
                e = ds.grids.back()->upload(src.data(), bytes, padded);
            } else {
                std::vector<F> wide(src.begin(), src.end());
                e = ds.grids.back()->upload(wide.data(), bytes, padded);
            }
            if (e != cudaSuccess) return fail(B2VP_ERR_CUDA, "grid upload failed: %s", cudaGetErrorString(e));
            o.data = ds.grids.back()->template as<F>();
            o.data_bytes = padded;
        }
    }
    std::vector<DMedium<F>> media(s->media.size());
    for (size_t i = 0; i < media.size(); ++i) {
        const B2vpMedium &a = s->media[i];
        DMedium<F> &o = media[i];
        std::memset(&o, 0, sizeof(o));
        o.type = a.type; o.sigma_t = a.sigma_t; o.albedo = a.albedo; o.phase = a.phase;
        o.spectral = a.has_spectral_extinction; o.sample_emitters = a.sample_emitters;
        o.scale = (F) a.scale;
        o.mg_offset = -1;
        if (a.type == B2VP_MEDIUM_HETEROGENEOUS) {
            const DVolume<F> &v = volumes[a.sigma_t];
            o.max_density = o.scale * v.max_value; /* heterogeneous.cpp:29 */
            for (int k = 0; k < 3; ++k) { o.amin[k] = v.bmin[k]; o.amax[k] = v.bmax[k]; }
            /* same grid geometry for sigma_t and albedo: one index computation serves both (volume_eval_pair) */
            const DVolume<F> &b = volumes[a.albedo];
            o.paired = v.type == 1 && b.type == 1 && v.filter == B2VP_FILTER_TRILINEAR && b.filter == B2VP_FILTER_TRILINEAR &&
                       v.channels == 1 && b.channels == 1 && v.wrap == b.wrap && v.nx == b.nx && v.ny == b.ny &&
                       v.nz == b.nz && std::memcmp(v.w2l, b.w2l, sizeof(v.w2l)) == 0;
        }
    }

    /* scene features: lets render() pick a kernel instance without code for unused features */
    uint32_t ft = 0;
    for (size_t i = 0; i < media.size(); ++i) {
        const DMedium<F> &m = media[i];
        if (m.type == B2VP_MEDIUM_HOMOGENEOUS) {
            ft |= FT_HOMOG;
            if (volumes[m.sigma_t].type == 1) ft |= FT_GRID_HOMOG;
        } else {
            ft |= FT_HETERO;
        }
        if (volumes[m.albedo].type == 1) ft |= FT_GRID_ALBEDO;
    }
    for (size_t i = 0; i < volumes.size(); ++i) {
        const DVolume<F> &v = volumes[i];
        if (v.type != 1) continue;
        ft |= v.filter == B2VP_FILTER_TRILINEAR ? FT_TRILINEAR : FT_NEAREST;
        if (v.wrap != B2VP_WRAP_CLAMP) ft |= FT_WRAP;
        if (v.channels != 1) ft |= FT_CH3;
        if (!(v.w2l[3] == 0 && v.w2l[7] == 0 && v.w2l[11] == 0 && v.w2l[15] == 1)) ft |= FT_PROJECTIVE;
    }
    /* leaf triangles by value, in leaf order (three 16-byte loads per test) */
    std::vector<BvhTri> soup(s->bvh_prims.size());
    for (size_t i = 0; i < soup.size(); ++i) {
        const BvhPrim &pr = s->bvh_prims[i];
        const B2vpShape &a = s->shapes[pr.shape];
        const uint32_t *fi = &s->faces[3 * (size_t) (a.face_offset + pr.face)];
        float *dst[3] = {soup[i].p0, soup[i].p1, soup[i].p2};
        for (int v = 0; v < 3; ++v)
            std::memcpy(dst[v], &s->vpos[3 * (size_t) (a.vertex_offset + fi[v])], 3 * sizeof(float));
        soup[i].shape = pr.shape; soup[i].face = pr.face; soup[i].pad = 0;
    }
    /* A small BVH (the 12-triangle cube that bounds the atmosphere in C1-C4: 11 nodes + 12 triangles = 928 bytes) is
       appended to the table blob, i.e. staged in shared memory by every CTA: Scene::ray_intersect runs ~12 times per
       path, each traversal is a chain of dependent loads, and the L1 left beside 222 KB of shared memory does not keep
       even 1 KB resident against the stream of cold path records (ncu: 29 % of all stall samples were
       long_scoreboard, led by the slab tests of this cube). */
    const size_t bvh_bytes = s->bvh_nodes.size() * sizeof(BvhNode) + soup.size() * sizeof(BvhTri);
    const bool bvh_small = !s->bvh_nodes.empty() && bvh_bytes <= BVH_SMEM_MAX;
    for (uint32_t i = 0; i < n_shapes; ++i) {
        const B2vpShape &a = s->shapes[i];
        if (a.sample_only) continue;   /* not intersected: no intersection code needed for it */
        ft |= a.type == B2VP_SHAPE_MESH ? (bvh_small ? FT_MESH_SMEM : FT_MESH) : a.type == B2VP_SHAPE_SPHERE ? FT_SPHERE : FT_RECT;
        if (a.type == B2VP_SHAPE_MESH && (a.has_vertex_normals || a.has_vertex_texcoords)) ft |= FT_MESH_ATTR;
    }
    if (have_tabphase) ft |= FT_TABPHASE;
    if (have_blendphase) ft |= FT_BLENDPHASE;
    if (d.sensor.type == B2VP_SENSOR_DISTANTFLUX || d.sensor.type == B2VP_SENSOR_IRRADIANCEMETER ||
        d.sensor.origin_type == B2VP_ORIGIN_SHAPE) ft |= FT_DISTANTFLUX;
    if (d.sensor.rfilter_radius > (double) ((F) 0.5 + K<F>::RayEpsilon)) ft |= FT_RFILTER;   /* imageblock.cpp:117; the box filter's radius is 0.5 + RayEpsilon */

    /* blob layout */
    uint32_t off = 0;
    auto place = [&](size_t bytes) { uint32_t o = off; off = align16(off + (uint32_t) bytes); return o; };
    g.ofs_shapes = place(shapes.size() * sizeof(DShape<F>));
    g.ofs_analytic = place(analytic.size() * sizeof(int));
    g.ofs_bsdfs = place(bsdfs.size() * sizeof(DBsdf<F>));
    g.ofs_media = place(media.size() * sizeof(DMedium<F>));
    g.ofs_volumes = place(volumes.size() * sizeof(DVolume<F>));
    g.ofs_phases = place(phases.size() * sizeof(DPhase<F>) + phase_tabs.size() * sizeof(F));
    g.ofs_emitters = place(emitters.size() * sizeof(DEmitter<F>));
    g.bvh_in_blob = bvh_small ? 1u : 0u;
    g.ofs_bvh = g.ofs_bvh_tris = 0;
    if (bvh_small) {
        g.ofs_bvh = place(s->bvh_nodes.size() * sizeof(BvhNode));
        g.ofs_bvh_tris = place(soup.size() * sizeof(BvhTri));
    }
    g.blob_bytes = align16(off ? off : 16);
    if (g.blob_bytes > 64 * 1024)
        return fail(B2VP_ERR_UNSUPPORTED, "scene tables (%u bytes) exceed the shared-memory table budget", g.blob_bytes);

    /* decide which grids are staged in shared memory */
    uint32_t grid_off = 0;
    ds.global_grid_bytes = 0;
    for (size_t i = 0; i < volumes.size(); ++i) {
        DVolume<F> &v = volumes[i];
        if (v.type != 1) continue;
        if (g.blob_bytes + FILM_SMEM_MAX + grid_off + v.data_bytes <= SMEM_BUDGET) {
            v.smem_offset = (int) grid_off;
            grid_off += (uint32_t) v.data_bytes;
        } else {
            ds.global_grid_bytes += v.data_bytes;
            /* A grid gathered from HBM / L2 gets trilinear cell records (vp_types.cuh, DVolume::cells): 8x the
               memory (2 x 257^3 x 32 B = 1.1 GB for C3, of 180 GB), one 32-byte sector per lookup instead of
               eight words in four rows.  B2VP_OPT_NO_CELL_RECORDS keeps the plain layout (A/B measurements). */
            const bool want_cells = !(d.options & B2VP_OPT_NO_CELL_RECORDS);
            const size_t n_cells = (size_t) (v.nx + 1) * (v.ny + 1) * (v.nz + 1), cbytes = n_cells * 8 * sizeof(F);
            size_t free_b = 0, total_b = 0;
            cudaMemGetInfo(&free_b, &total_b);
            if (want_cells && v.filter == B2VP_FILTER_TRILINEAR && v.channels == 1 && v.wrap == B2VP_WRAP_CLAMP &&
                cbytes < free_b / 4) {
                ds.cells.emplace_back(new DevBuf());
                cudaError_t e = ds.cells.back()->alloc(cbytes);
                if (e != cudaSuccess) return fail(B2VP_ERR_CUDA, "cell-record allocation failed: %s", cudaGetErrorString(e));
                F *cp = ds.cells.back()->template as<F>();
                const int blocks = (int) std::min<size_t>((n_cells + 255) / 256, (size_t) dc->sm_count * 16);
                build_cells_kernel<F><<<blocks, 256>>>(v.data, cp, v.nx, v.ny, v.nz);
                CU_CHECK(cudaGetLastError());
                v.cells = cp;
                ds.cell_bytes += cbytes;
            }
        }
    }
    CU_CHECK(cudaDeviceSynchronize());
    g.grid_smem_bytes = grid_off;
    for (size_t i = 0; i < volumes.size(); ++i)
        if (volumes[i].type == 1) ft |= volumes[i].smem_offset >= 0 ? FT_GRID_SMEM : FT_GRID_GLOBAL;
    g.features = ft;

    ds.host_blob.assign(g.blob_bytes, 0);
    auto put = [&](uint32_t o, const void *p, size_t n) { if (n) std::memcpy(ds.host_blob.data() + o, p, n); };
    put(g.ofs_shapes, shapes.data(), shapes.size() * sizeof(DShape<F>));
    put(g.ofs_analytic, analytic.data(), analytic.size() * sizeof(int));
    put(g.ofs_bsdfs, bsdfs.data(), bsdfs.size() * sizeof(DBsdf<F>));
    put(g.ofs_media, media.data(), media.size() * sizeof(DMedium<F>));
    put(g.ofs_volumes, volumes.data(), volumes.size() * sizeof(DVolume<F>));
    put(g.ofs_phases, phases.data(), phases.size() * sizeof(DPhase<F>));
    put(g.ofs_phases + (uint32_t) (phases.size() * sizeof(DPhase<F>)), phase_tabs.data(), phase_tabs.size() * sizeof(F));
    put(g.ofs_emitters, emitters.data(), emitters.size() * sizeof(DEmitter<F>));
    if (bvh_small) {
        put(g.ofs_bvh, s->bvh_nodes.data(), s->bvh_nodes.size() * sizeof(BvhNode));
        put(g.ofs_bvh_tris, soup.data(), soup.size() * sizeof(BvhTri));
    }
    CU_CHECK(ds.blob.upload(ds.host_blob.data(), ds.host_blob.size()));
    g.blob = ds.blob.template as<unsigned char>();

    g.n_shapes = (int) n_shapes; g.n_analytic = (int) analytic.size(); g.n_media = (int) media.size();
    g.n_volumes = (int) volumes.size(); g.n_emitters = (int) emitters.size();
    g.n_bvh_nodes = (int) s->bvh_nodes.size();
    if (!s->vpos.empty()) { CU_CHECK(ds.vpos.upload(s->vpos.data(), s->vpos.size() * 4)); g.vpos = ds.vpos.template as<float>(); }
    if (!s->vnrm.empty()) { CU_CHECK(ds.vnrm.upload(s->vnrm.data(), s->vnrm.size() * 4)); g.vnrm = ds.vnrm.template as<float>(); }
    if (!s->vuv.empty()) { CU_CHECK(ds.vuv.upload(s->vuv.data(), s->vuv.size() * 4)); g.vuv = ds.vuv.template as<float>(); }
    if (!s->faces.empty()) { CU_CHECK(ds.faces.upload(s->faces.data(), s->faces.size() * 4)); g.faces = ds.faces.template as<uint32_t>(); }
    if (!s->bvh_nodes.empty()) {
        CU_CHECK(ds.bvh.upload(s->bvh_nodes.data(), s->bvh_nodes.size() * sizeof(BvhNode)));
        CU_CHECK(ds.bvh_prims.upload(s->bvh_prims.data(), s->bvh_prims.size() * sizeof(BvhPrim)));
        g.bvh = ds.bvh.template as<BvhNode>(); g.bvh_prims = ds.bvh_prims.template as<BvhPrim>();
        CU_CHECK(ds.bvh_tris.upload(soup.data(), soup.size() * sizeof(BvhTri)));
        g.bvh_tris = ds.bvh_tris.template as<BvhTri>();
    }
    if (!s->sensor_xf.empty()) {
        std::vector<F> xf(s->sensor_xf.begin(), s->sensor_xf.end());
        CU_CHECK(ds.sensor_xf.upload(xf.data(), xf.size() * sizeof(F)));
        g.sensor_xf = ds.sensor_xf.template as<F>();
    }

    /* kd-tree box: scene bbox padded by (extents + 1) * eps (kdtree.h:1811-1813) */
    for (int k = 0; k < 3; ++k) {
        F bmin = (F) d.bbox_min[k], bmax = (F) d.bbox_max[k];
        F extra = ((bmax - bmin) + (F) 1) * K<F>::Epsilon;
        g.kd_min[k] = bmin - extra; g.kd_max[k] = bmax + extra;
        g.bs_center[k] = (F) d.bsphere_center[k];
    }
    g.bs_radius = (F) d.bsphere_radius;

    const B2vpSensor &sn = d.sensor;
    DSensor<F> &o = g.sensor;
    o.type = sn.type; o.direction_type = sn.direction_type; o.target_type = sn.target_type;
    o.origin_type = sn.origin_type; o.target_shape = sn.target_shape; o.flip_directions = sn.flip_directions;
    o.needs_aperture_sample = sn.needs_aperture_sample; o.medium = sn.medium; o.n_transforms = sn.n_transforms;
    o.has_shutter = sn.shutter_open_time > 0 ? 1 : 0;
    cvt(sn.to_world.m, o.to_world, 16); cvt(sn.target_point, o.target_point, 3);
    cvt(sn.reference_normal, o.reference_normal, 3);
    o.n_pixels = sn.film_size[0] * sn.film_size[1]; o.origin_shape = sn.origin_shape;
    for (int k = 0; k < 2; ++k) { o.crop_offset[k] = sn.crop_offset[k]; o.crop_size[k] = sn.crop_size[k]; }
    /* imageblock.cpp:117: the footprint loop only for a radius above 0.5 + RayEpsilon */
    const bool wide = sn.rfilter_radius > (double) ((F) 0.5 + K<F>::RayEpsilon);
    o.rf_taps = wide ? (int) sn.rfilter_taps : 0; o.rf_border = wide ? (int) sn.rfilter_border : 0;
    o.rf_radius = (F) sn.rfilter_radius; o.rf_scale = (F) sn.rfilter_scale;
    for (int k = 0; k <= B2VP_RFILTER_RESOLUTION; ++k) o.rf_values[k] = wide ? (F) sn.rfilter_values[k] : (F) 0;

    g.max_depth = d.integrator.max_depth < 0 ? 0xffffffffu : (uint32_t) d.integrator.max_depth;
    g.rr_depth = (uint32_t) d.integrator.rr_depth;
    g.hide_emitters = d.integrator.hide_emitters; g.ltr = (int) d.integrator.arg_order_ltr;
    g.film_ch = d.integrator.moment ? B2VP_FILM_CHANNELS_MOMENT : B2VP_FILM_CHANNELS; g.mis = (int) d.integrator.mis;
    g.environment = -1; g.pad_env = 0;
    for (uint32_t i = 0; i < d.n_emitters; ++i) if (s->emitters[i].type == B2VP_EMITTER_CONSTANT) g.environment = (int) i;
    ds.ready = true;
    return B2VP_OK;
}

/* ---- stream space (host mirror of integrator.cpp:51-110, spiral.cpp:8-72) -- */
std::vector<SpiralBlockD> spiral_blocks(int w, int h, int offx, int offy, int bs) {
    int bx = (int) std::ceil((float) w / (float) bs), by = (int) std::ceil((float) h / (float) bs);
    int count = bx * by, counter = 0;
    int px = bx / 2, py = by / 2, dir = 0, steps_left = 1, steps = 1;
    std::vector<SpiralBlockD> out;
    while (counter < count) {
        SpiralBlockD b;
        b.ox = px * bs; b.oy = py * bs;
        b.sx = std::min(bs, w - b.ox); b.sy = std::min(bs, h - b.oy);
        b.ox += offx; b.oy += offy;
        out.push_back(b);
        if (++counter == count) break;
        do {
            switch (dir) { case 0: ++px; break; case 1: ++py; break; case 2: --px; break; default: --py; break; }
            if (--steps_left == 0) {
                dir = (dir + 1) % 4;
                if (dir == 2 || dir == 0) ++steps;
                steps_left = steps;
            }
        } while (px < 0 || py < 0 || px >= bx || py >= by);
    }
    return out;
}

struct StreamSpace {
    DStreams d{};
    std::vector<SpiralBlockD> blocks;
    uint64_t n_streams = 0;
};

int make_stream_space(const b2vp_scene *s, int32_t seed_mode, uint64_t explicit_n, uint32_t explicit_per,
                      StreamSpace &sp) {
    const B2vpSceneDesc &d = s->desc;
    DStreams &o = sp.d;
    o.mode = seed_mode;
    o.w = d.sensor.crop_size[0]; o.h = d.sensor.crop_size[1];
    o.ox = d.sensor.crop_offset[0]; o.oy = d.sensor.crop_offset[1];
    o.seed = d.sensor.sampler_seed;
    uint32_t spp = d.sensor.sample_count;
    o.spp_pass = d.integrator.samples_per_pass == 0xffffffffu ? spp : std::min(d.integrator.samples_per_pass, spp);
    if (o.spp_pass == 0 || spp % o.spp_pass != 0) /* integrator.cpp:61-63 */
        return fail(B2VP_ERR_INVALID, "sample_count must be a multiple of samples_per_pass");
    o.n_passes = spp / o.spp_pass;
    o.bs = 0;
    if (seed_mode == B2VP_SEED_SCALAR_BLOCKS) {
        uint32_t bs = d.integrator.block_size;
        if (bs == 0) { /* integrator.cpp:89-97 with one thread */
            bs = 32;
            while (!(bs == 1 || (uint64_t) ((o.w + bs - 1) / bs) * ((o.h + bs - 1) / bs) >= 1)) bs /= 2;
        }
        o.bs = bs;
        sp.blocks = spiral_blocks(o.w, o.h, o.ox, o.oy, (int) bs);
        o.n_blocks = (uint32_t) sp.blocks.size();
        sp.n_streams = (uint64_t) o.n_passes * sp.blocks.size() * bs * bs;
        o.samples_per_stream = o.spp_pass;
    } else if (seed_mode == B2VP_SEED_WAVEFRONT) {
        sp.n_streams = (uint64_t) o.w * o.h * o.spp_pass;
        o.samples_per_stream = o.n_passes;
    } else if (seed_mode == B2VP_SEED_EXPLICIT) {
        sp.n_streams = explicit_n;
        o.samples_per_stream = explicit_per;
    } else
        return fail(B2VP_ERR_INVALID, "unknown seed mode %d", seed_mode);
    return B2VP_OK;
}

} // namespace

namespace {

/* 16 independent FMA chains per thread: the non-tensor FP32 roofline denominator */
__global__ void fma_peak_kernel(float *sink, int iters, float m) {
    float a[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) a[k] = (float) (threadIdx.x + k);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 16; ++k) a[k] = __fmaf_rn(a[k], m, 0.5f);
    }
    float r = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) r += a[k];
    if (r == 123.456f) sink[threadIdx.x] = r;
}

template <typename F> DeviceScene<F> &dev_of(DeviceCtx *dc);
template <> DeviceScene<float> &dev_of<float>(DeviceCtx *dc) { return dc->dev32; }
template <> DeviceScene<double> &dev_of<double>(DeviceCtx *dc) { return dc->dev64; }

inline uint32_t align128(uint32_t x) { return (x + 127u) & ~127u; }
inline size_t align256(size_t x) { return (x + 255) & ~(size_t) 255; }

/* ---- NCCL, loaded on demand ------------------------------------------------------------------------------
   Only a multi-device render needs it, and a process may already hold a copy (torch bundles its own libnccl.so.2
   under the same SONAME): dlopen by SONAME returns whichever is loaded.  Communicators are made once per device set
   and kept for the life of the process (ncclCommInitAll costs about a second). */
struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Reduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
std::mutex g_nccl_mutex;
NcclApi g_nccl;
std::map<std::vector<int>, std::vector<ncclComm_t>> g_nccl_comms;

int nccl_load() {
    if (g_nccl.lib) return B2VP_OK;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
    if (!h) return fail(B2VP_ERR_UNSUPPORTED, "multi-device render needs NCCL: dlopen(libnccl.so.2) failed: %s", dlerror());
    NcclApi a;
    a.lib = h;
    a.CommInitAll = (decltype(a.CommInitAll)) dlsym(h, "ncclCommInitAll");
    a.GroupStart = (decltype(a.GroupStart)) dlsym(h, "ncclGroupStart");
    a.GroupEnd = (decltype(a.GroupEnd)) dlsym(h, "ncclGroupEnd");
    a.Reduce = (decltype(a.Reduce)) dlsym(h, "ncclReduce");
    a.GetErrorString = (decltype(a.GetErrorString)) dlsym(h, "ncclGetErrorString");
    if (!a.CommInitAll || !a.GroupStart || !a.GroupEnd || !a.Reduce || !a.GetErrorString)
        return fail(B2VP_ERR_UNSUPPORTED, "libnccl.so.2 lacks an entry point this library needs");
    g_nccl = a;
    return B2VP_OK;
}
#define NCCL_CHECK(expr)                                                                                          \
    do {                                                                                                          \
        ncclResult_t r_ = (expr);                                                                                 \
        if (r_ != ncclSuccess)                                                                                    \
            return fail(B2VP_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, g_nccl.GetErrorString(r_), __FILE__, __LINE__); \
    } while (0)

int nccl_comms_for(const std::vector<int> &devices, std::vector<ncclComm_t> **out) {
    std::lock_guard<std::mutex> lock(g_nccl_mutex);
    int rc = nccl_load();
    if (rc != B2VP_OK) return rc;
    auto it = g_nccl_comms.find(devices);
    if (it == g_nccl_comms.end()) {
        std::vector<ncclComm_t> comms(devices.size());
        NCCL_CHECK(g_nccl.CommInitAll(comms.data(), (int) devices.size(), devices.data()));
        it = g_nccl_comms.emplace(devices, std::move(comms)).first;
    }
    *out = &it->second;
    return B2VP_OK;
}

/* Interleaved sharding of `count` streams over n devices (b2vp_shard_streams): block = a power of two small enough
   that every device gets many blocks (balance: pixels differ in cost, every device should see every pixel) and large
   enough that a block is a few warps' worth of streams. */
void shard_streams(uint64_t count, uint32_t n, uint32_t index, uint64_t *block, uint64_t *count_i) {
    uint64_t B = 1024;
    while (B > 32 && count / B < (uint64_t) n * 8) B >>= 1;
    const uint64_t n_blocks = (count + B - 1) / B;
    uint64_t c = 0;
    if (n && index < n && n_blocks > index) {
        const uint64_t mine = (n_blocks - index + n - 1) / n;          /* blocks index, index + n, ... */
        c = mine * B;
        const uint64_t last_block = index + (mine - 1) * (uint64_t) n;
        if (last_block == n_blocks - 1 && count % B) c -= B - count % B;   /* the ragged last block is mine */
    }
    *block = B;
    *count_i = c;
}

/* ---- majorant grids (B2vpRenderArgs::majorant_grid) ---------------------------------------------------------
   Built on first use, per precision instance.  For every heterogeneous medium whose sigma_t is a single-channel
   clamp-mode grid with an affine world_to_local: R[a] cells per axis over the volume's unit cube, each holding the
   maximum of the UNSCALED grid over all voxels a lookup at a position inside the cell can read -- the trilinear
   support floor(p n - 0.5) + {0, 1} of the cell's range plus one voxel of margin on each side, which also covers a
   position that rounding puts just across a cell face.  The medium's scale is applied by the kernel, so
   b2vp_scene_update of the scale costs nothing; b2vp_scene_update_grid drops the grids (rebuilt at the next use). */
constexpr uint32_t MG_MAX_CELLS = 4096, MG_MAX_AXIS = 64;

template <typename F> int ensure_majorant_grids(b2vp_scene *s, DeviceScene<F> &ds, uint32_t request, bool *any) {
    *any = false;
    if (ds.mg_request == request) { *any = ds.mg_bytes > 0; return B2VP_OK; }
    unsigned char *b = ds.host_blob.data();
    auto *media = reinterpret_cast<DMedium<F> *>(b + ds.g.ofs_media);
    std::vector<float> all;
    for (size_t i = 0; i < s->media.size(); ++i) {
        DMedium<F> &m = media[i];
        m.mg_offset = -1; m.mg_res[0] = m.mg_res[1] = m.mg_res[2] = 0;
        if (m.type != B2VP_MEDIUM_HETEROGENEOUS) continue;
        const B2vpVolume &v = s->volumes[m.sigma_t];
        const double *w = v.world_to_local.m;
        if (v.type != B2VP_VOLUME_GRID || v.channels != 1 || v.wrap != B2VP_WRAP_CLAMP ||
            !(w[3] == 0 && w[7] == 0 && w[11] == 0 && w[15] == 1))
            continue;
        const int n[3] = { v.res[0], v.res[1], v.res[2] };
        int R[3];
        const int longest = std::max(n[0], std::max(n[1], n[2]));
        for (int a = 0; a < 3; ++a) {
            if (request > 1)   /* n cells along the longest axis, the others in proportion */
                R[a] = (int) std::max<int64_t>(1, std::min<int64_t>((int64_t) request * n[a] / longest, std::min<int>(n[a], MG_MAX_AXIS)));
            else
                R[a] = std::min(n[a], 16);
        }
        /* 16 cells along an axis is the measured optimum on the 1D atmosphere (C2, 1 x 1 x 1000 voxels: 8-16 layers
           422-429 Mpaths/s, 32 layers 391, 64 layers 315 -- every cell crossed costs a DDA step, and the exponential
           profile is bounded within a factor e^0.8 per layer at 16) */
        while ((uint64_t) R[0] * R[1] * R[2] > MG_MAX_CELLS) {
            int a = R[0] >= R[1] ? (R[0] >= R[2] ? 0 : 2) : (R[1] >= R[2] ? 1 : 2);
            R[a] = (R[a] + 1) / 2;
        }
        const std::vector<float> &data = s->volume_data[m.sigma_t];
        int lo[3][MG_MAX_AXIS], hi[3][MG_MAX_AXIS];
        for (int a = 0; a < 3; ++a)
            for (int c = 0; c < R[a]; ++c) {
                const double q0 = (double) c / R[a] * n[a] - 0.5, q1 = (double) (c + 1) / R[a] * n[a] - 0.5;
                lo[a][c] = std::max(0, (int) std::floor(q0) - 1);
                hi[a][c] = std::min(n[a] - 1, (int) std::floor(q1) + 2);
            }
        m.mg_offset = (int) all.size();
        m.mg_res[0] = R[0]; m.mg_res[1] = R[1]; m.mg_res[2] = R[2];
        for (int cz = 0; cz < R[2]; ++cz)
            for (int cy = 0; cy < R[1]; ++cy)
                for (int cx = 0; cx < R[0]; ++cx) {
                    float mx = 0.f;
                    for (int z = lo[2][cz]; z <= hi[2][cz]; ++z)
                        for (int y = lo[1][cy]; y <= hi[1][cy]; ++y) {
                            const float *row = &data[((size_t) z * n[1] + y) * n[0]];
                            for (int x = lo[0][cx]; x <= hi[0][cx]; ++x) mx = row[x] > mx ? row[x] : mx;
                        }
                    /* the exact maximum, not rounded up: where sigma_t equals it (the inside of a uniform cloud) the null
                       coefficient must be exactly 0 as under the global majorant, so that an emitter march ends at its
                       first collision there instead of dragging a 1e-7, 1e-14, ... transmittance through the cloud */
                    all.push_back(mx);
                }
    }
    ds.mg_request = request;
    ds.mg_bytes = (uint32_t) (all.size() * sizeof(float));
    CU_CHECK(cudaMemcpy(ds.blob.p, b, ds.host_blob.size(), cudaMemcpyHostToDevice));   /* the patched medium records */
    if (!all.empty()) {
        CU_CHECK(ds.mg.upload(all.data(), all.size() * sizeof(float)));
        *any = true;
    }
    return B2VP_OK;
}

/* ---- launch plan: which kernel, with which shape and shared-memory layout ------------------------------ */
struct LaunchPlan {
    const void *kern = nullptr;
    int threads = 0, grid = 0;
    size_t smem = 0;
    bool queue_kernel = false;
    size_t cold_bytes = 0;          /* queue kernel: grid * q_n_slots cold path records */
    std::string name;
};

template <typename F> const void *queue_entry(uint32_t ft, int maxt, bool fast, bool mis, bool env, const char **preset, int *threads);
template <> const void *queue_entry<float>(uint32_t ft, int maxt, bool fast, bool mis, bool env, const char **preset, int *threads) {
    *threads = maxt;
    if (env) return fast ? vp_launch::queue_f32_fast_env(mis, preset) : vp_launch::queue_f32_env(mis, preset);
    if (mis) return fast ? vp_launch::queue_f32_fast_mis(ft, preset, threads) : vp_launch::queue_f32_mis(ft, preset, threads);
    return fast ? vp_launch::queue_f32_fast(ft, preset, threads) : vp_launch::queue_f32(ft, maxt, preset, threads);
}
template <> const void *queue_entry<double>(uint32_t ft, int maxt, bool, bool mis, bool env, const char **preset, int *threads) {
    *threads = maxt;
    if (env) return vp_launch::queue_f64_env(mis, preset);
    return mis ? vp_launch::queue_f64_mis(ft, preset, threads) : vp_launch::queue_f64(ft, maxt, preset, threads);
}
template <typename F> const void *stream_entry() { return sizeof(F) == 4 ? vp_launch::stream_f32() : vp_launch::stream_f64(); }
template <typename F> const void *lanes_entry() { return sizeof(F) == 4 ? vp_launch::lanes_f32() : vp_launch::lanes_f64(); }

/* Fills the shared-memory layout fields of `out` and the plan.  `count` = streams of this launch. */
template <typename F>
int plan_launch(b2vp_scene *s, DeviceCtx *dc, DeviceScene<F> &ds, const B2vpRenderArgs *a, const DStreams &spd, size_t film_bytes,
                RenderOut<F> &out, LaunchPlan &pl, bool may_allocate = true) {
    /* shared memory: [tables][grids][film] */
    out.grid_smem_offset = align128(ds.g.blob_bytes);
    out.film_smem_offset = align128(out.grid_smem_offset + ds.g.grid_smem_bytes);
    out.film_in_smem = film_bytes <= FILM_SMEM_MAX ? 1u : 0u;
    size_t smem = out.film_smem_offset + (out.film_in_smem ? film_bytes : 0);
    bool fast = a->math_mode == B2VP_MATH_FAST;
    if (a->math_mode != B2VP_MATH_EXACT && !fast) return fail(B2VP_ERR_INVALID, "unknown math_mode %d", a->math_mode);
    if (fast && sizeof(F) != 4) return fail(B2VP_ERR_INVALID, "B2VP_MATH_FAST is an fp32 mode (precision F32)");
    /* Kernel selection.  Product: the CTA-resident wavefront with shared-memory phase queues (vp_queue.cuh), for
       volpath and -- as other bodies of the same phases -- volpathmis (FT_MIS) and scenes with an
       environment emitter (FT_ENV; constant.cpp: non-delta emitter sampling, light on escape).  Those two exist as
       512-thread instances only: when another kernel is asked for (tests, comparisons) they run on the
       one-thread-per-stream kernel, an order of magnitude slower; B2vpSceneInfo::last_kernel says which. */
    const bool mis = s->desc.integrator.mis != 0;
    bool env = false;
    for (const B2vpEmitter &e : s->emitters) env = env || e.type == B2VP_EMITTER_CONSTANT;
    int kernel = a->kernel;
    if (kernel < B2VP_KERNEL_AUTO || kernel > B2VP_KERNEL_STREAM) return fail(B2VP_ERR_INVALID, "unknown kernel %d", kernel);
    const bool needs_stream = (mis || env) && kernel != B2VP_KERNEL_AUTO && kernel != B2VP_KERNEL_QUEUE_GENERIC;
    if ((mis || env) && a->majorant_grid)
        return fail(B2VP_ERR_UNSUPPORTED, "majorant_grid is a mode of volpath without an environment emitter, not of volpathmis");
    if (needs_stream) {
        if (fast)
            return fail(B2VP_ERR_UNSUPPORTED, "B2VP_MATH_FAST: volpathmis / environment-emitter scenes have fast instances of the 512-thread queue kernel only");
        kernel = B2VP_KERNEL_STREAM;
    }
    if (fast && (kernel == B2VP_KERNEL_LANES || kernel == B2VP_KERNEL_STREAM || kernel == B2VP_KERNEL_QUEUE_256))
        return fail(B2VP_ERR_INVALID, "B2VP_MATH_FAST exists for the 512-thread queue kernel only");
    uint32_t launch_features = ds.g.features;
    const char *preset = "";
    pl.queue_kernel = kernel == B2VP_KERNEL_AUTO || kernel == B2VP_KERNEL_QUEUE_GENERIC || kernel == B2VP_KERNEL_QUEUE_256;
    out.mg = nullptr; out.mg_bytes = 0; out.mg_smem_offset = 0;
    if (a->majorant_grid) {
        if (!pl.queue_kernel) return fail(B2VP_ERR_INVALID, "majorant_grid is a mode of the queue kernel");
        if (!may_allocate && ds.mg_request != a->majorant_grid)
            return fail(B2VP_ERR_INVALID, "b2vp_render_noalloc: the majorant grids for this request are not built yet; "
                                          "call b2vp_render once with the same majorant_grid at set-up");
        bool any = false;
        int rc = ensure_majorant_grids<F>(s, ds, a->majorant_grid, &any);
        if (rc != B2VP_OK) return rc;
        if (any) {   /* media without an eligible sigma_t grid keep the global majorant */
            launch_features |= FT_MAJORANT_GRID;
            out.mg = ds.mg.template as<float>();
            out.mg_bytes = ds.mg_bytes;
            out.mg_smem_offset = (uint32_t) ((smem + 127) & ~(size_t) 127);
            smem = out.mg_smem_offset + ds.mg_bytes;
        }
    }
    if (pl.queue_kernel) {
        pl.threads = kernel == B2VP_KERNEL_QUEUE_256 ? 256 : 512;
        pl.kern = queue_entry<F>(launch_features | (kernel == B2VP_KERNEL_QUEUE_GENERIC ? vp_launch::FORCE_GENERIC : 0u),
                                 pl.threads, fast, mis, env, &preset, &pl.threads);
        pl.name = std::string("queue/") + preset + (kernel == B2VP_KERNEL_QUEUE_256 ? "/256" : "") + (fast ? " fast" : "");
    } else if (kernel == B2VP_KERNEL_LANES) {
        pl.kern = lanes_entry<F>(); pl.threads = 256; pl.name = "lanes";
    } else {
        pl.kern = stream_entry<F>(); pl.threads = 256; pl.name = "stream (one thread per stream)";
    }
    pl.name += sizeof(F) == 4 ? " f32" : " f64";
    if (pl.queue_kernel) {
        /* path slots per thread.  With everything hot on chip the shared memory left sets the pool size (the more
           paths, the longer a CTA stays in one phase).  When large grids are gathered through L2 the cold path
           records compete with them for it: measured on 2 x 256^3 grids, 2 slots per thread give 90 Mpaths/s,
           3 and more 56. */
        const bool big_grids = ds.global_grid_bytes > ((size_t) 16 << 20);
        double oversub = big_grids ? 2.0 : 4.0;
        /* the large-mesh preset (768-thread CTAs): its BVH is read through L1 / L2, and every KB of shared memory the
           pool does not take is L1 -- C5 fast 153.7 Mpaths/s with the pool that fills shared memory (1.8 slots per
           thread), 165 at 1.5, **172 at 1.0-1.25** */
        if (pl.threads > 512) oversub = 1.125;
        if (a->tuning[0]) oversub = std::min(8.0, std::max(1.0, a->tuning[0] / 16.0));
        /* shared memory: [tables][grids][film][queue control + rings][path slots] */
        size_t base = (smem + 127) & ~(size_t) 127;
        size_t avail = dc->smem_optin > base + 1024 ? dc->smem_optin - base - 1024 : 0;
        size_t per_slot = sizeof(PathSlot<F>) + 2 * PH_N * sizeof(unsigned short);
        size_t max_slots = avail > 256 ? (avail - sizeof(QueueCtl) - 64) / per_slot : 0;
        size_t n_slots = std::min<size_t>((size_t) (pl.threads * oversub), max_slots) & ~(size_t) 31;
        if (n_slots < 64)
            return fail(B2VP_ERR_CUDA, "not enough shared memory for the path-slot pool (%zu bytes free)", avail);
        n_slots = std::min<size_t>(n_slots, 32768);
        size_t cap = 64;
        while (cap < n_slots) cap *= 2;
        out.q_ctl_offset = (uint32_t) base;
        size_t qbytes = (sizeof(QueueCtl) + PH_N * cap * sizeof(unsigned short) + 15) & ~(size_t) 15;
        out.q_slots_offset = (uint32_t) (base + qbytes);
        out.q_n_slots = (uint32_t) n_slots;
        out.q_ring_cap = (uint32_t) cap;
        /* a warp keeps its STEP batch while at least this many lanes stay (sweep after STEP went register-resident:
           16-24 equal within 0.5 % in fast mode, 24 best in exact mode, 26 costs 2.5 %; profiles/r2_history.md) */
        out.q_stay_min = a->tuning[1] ? std::min(33u, a->tuning[1]) : 24u;
        out.q_max_iter = a->tuning[2] ? std::min(1024u, a->tuning[2]) : 8u;
        /* 2: when the CTA moves on, the ray-making phases are served before STEP (+4 % with everything on chip);
           1: plain fullest-queue order, better when STEP is bound by gather latency (large grids: 93 vs 84 Mpaths/s) */
        out.q_cta_phase = a->tuning[3] ? std::min(2u, a->tuning[3] - 1u) : (big_grids ? 1u : 2u);
        smem = out.q_slots_offset + n_slots * sizeof(PathSlot<F>);
    }
    int per_sm = 0;
    CU_CHECK(cudaFuncSetAttribute(pl.kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pl.kern, pl.threads, smem));
    if (per_sm < 1) return fail(B2VP_ERR_CUDA, "kernel does not fit on an SM (smem %zu)", smem);
    const uint64_t want = (spd.count + pl.threads - 1) / pl.threads;
    pl.grid = (int) std::min<uint64_t>((uint64_t) dc->sm_count * per_sm, std::max<uint64_t>(want, 1));
    pl.smem = smem;
    pl.cold_bytes = pl.queue_kernel ? (size_t) pl.grid * out.q_n_slots * sizeof(PathCold<F>) : 0;
    return B2VP_OK;
}

/* stream range of a launch: [first, first + count) or interleaved blocks (B2vpRenderArgs::interleave_*) */
int set_stream_range(const B2vpRenderArgs *a, StreamSpace &sp) {
    uint64_t first = a->first_stream;
    uint64_t count = a->n_streams ? a->n_streams : (first <= sp.n_streams ? sp.n_streams - first : 0);
    if (a->seed_mode == B2VP_SEED_EXPLICIT) { first = 0; count = a->n_streams; }
    sp.d.il_block = 0; sp.d.il_count = 1; sp.d.il_index = 0; sp.d.il_shift = -1;
    if (a->interleave_block) {
        if (a->seed_mode == B2VP_SEED_EXPLICIT || !a->n_streams || !a->interleave_count ||
            a->interleave_index >= a->interleave_count)
            return fail(B2VP_ERR_INVALID, "invalid interleaved stream range");
        const uint64_t B = a->interleave_block, kmax = count - 1;
        if (B & (B - 1)) return fail(B2VP_ERR_INVALID, "interleave_block must be a power of two (b2vp_shard_streams gives one)");
        const uint64_t last = first + ((kmax / B) * a->interleave_count + a->interleave_index) * B + kmax % B;
        if (last >= sp.n_streams) return fail(B2VP_ERR_INVALID, "stream range out of bounds");
        sp.d.il_block = B; sp.d.il_count = a->interleave_count; sp.d.il_index = a->interleave_index;
        { int sh = 0; while ((B >> sh) != 1) ++sh; sp.d.il_shift = sh; }
    } else if (first + count > sp.n_streams) return fail(B2VP_ERR_INVALID, "stream range out of bounds");
    sp.d.first = first;
    sp.d.count = count;
    return B2VP_OK;
}

template <typename F>
cudaError_t launch_kernel(const LaunchPlan &pl, const DeviceScene<F> &ds, const DStreams &spd, const RenderOut<F> &out,
                          cudaStream_t stream, DeviceCtx *dc = nullptr, bool pin_cold = false) {
    void *kargs[3] = { (void *) &ds.g, (void *) &spd, (void *) &out };
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) pl.grid); cfg.blockDim = dim3((unsigned) pl.threads);
    cfg.dynamicSmemBytes = pl.smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    if (pin_cold && dc && dc->l2_window_max && out.q_cold && pl.cold_bytes) {
        if (!dc->l2_persist_bytes) {   /* the carve-out takes capacity from everything else: set it only when it is used */
            const size_t want = std::min(dc->l2_persist_max, (pl.cold_bytes + ((size_t) 1 << 20)) & ~(((size_t) 1 << 20) - 1));
            if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) == cudaSuccess) dc->l2_persist_bytes = want;
            (void) cudaGetLastError();
        }
        if (!dc->l2_persist_bytes) pin_cold = false;
    }
    if (pin_cold && dc && dc->l2_window_max && out.q_cold && pl.cold_bytes) {
        /* EXPERIMENT (debug_flags bit 1), off by default: a persisting L2 window over the cold path records.  Measured
           (profiles/r2_history.md): no change in kernel time on C2 or C3, and the carve-out it needs takes capacity
           from everything else -- DRAM writes per C2 launch 0.40 GB without it, 1.08 GB with a 41 MB carve-out,
           5.3 GB with the device maximum of 79 MB */
        attr[0].id = cudaLaunchAttributeAccessPolicyWindow;
        attr[0].val.accessPolicyWindow.base_ptr = out.q_cold;
        const size_t window = std::min(pl.cold_bytes, dc->l2_window_max);
        attr[0].val.accessPolicyWindow.num_bytes = window;
        /* a window larger than the persisting carve-out thrashes inside it: mark only the share that fits */
        attr[0].val.accessPolicyWindow.hitRatio = (float) std::min(1.0, (double) dc->l2_persist_bytes / (double) window);
        attr[0].val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        attr[0].val.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        cfg.attrs = attr; cfg.numAttrs = 1;
    }
    return cudaLaunchKernelExC(&cfg, pl.kern, kargs);
}

void fill_totals(B2vpCounters &t, const unsigned long long *tot, float ms, uint32_t launches) {
    std::memset(&t, 0, sizeof(t));
    t.n_paths = tot[TOT_PATHS]; t.n_rng_draws = tot[TOT_DRAWS]; t.n_medium_steps = tot[TOT_STEPS];
    t.n_null = tot[TOT_NULL]; t.n_real = tot[TOT_REAL]; t.n_intersect = tot[TOT_ISECT];
    t.n_surface = tot[TOT_SURF]; t.n_invalid_samples = tot[TOT_INVALID];
    t.n_bvh_nodes = tot[TOT_BVH_NODES]; t.n_bvh_tris = tot[TOT_BVH_TRIS];
    t.gpu_ms = ms;
    t.n_launches = launches;
}

/* Waits for everything queued on `stream` while watching for b2vp_cancel and the integrator's timeout
   (should_stop(), include/mitsuba/render/integrator.h:143-146).  Either raises the device stop flag through the
   context's control stream -- a 4-byte copy that overtakes the running kernel -- after which the kernel hands out no
   new streams and drains.  *stopped: 0 ran to the end, 1 timeout, 2 cancelled. */
int wait_for_render(b2vp_scene *s, DeviceCtx *dc, cudaStream_t stream, double timeout_s,
                    std::chrono::steady_clock::time_point t_start, uint32_t *stopped) {
    static const int one = 1;
    *stopped = 0;
    CU_CHECK(cudaEventRecord(dc->ev_done, stream));
    for (;;) {
        cudaError_t e = cudaEventQuery(dc->ev_done);
        if (e == cudaSuccess) break;
        if (e != cudaErrorNotReady) return fail(B2VP_ERR_CUDA, "render failed: %s", cudaGetErrorString(e));
        if (!*stopped) {
            uint32_t why = 0;
            if (s->cancelled.load(std::memory_order_relaxed)) why = 2;
            else if (timeout_s > 0 &&
                     std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count() > timeout_s) why = 1;
            if (why) {
                CU_CHECK(cudaMemcpyAsync(dc->stop_flag.p, &one, sizeof(int), cudaMemcpyHostToDevice, dc->ctl_stream));
                *stopped = why;
            }
        }
        std::this_thread::yield();
    }
    return B2VP_OK;
}

template <typename F> int render_typed(b2vp_scene *s, DeviceCtx *dc, const B2vpRenderArgs *a, void *film_user) {
    CU_CHECK(cudaSetDevice(dc->device));
    DeviceScene<F> &ds = dev_of<F>(dc);
    if (!ds.ready) {
        int rc = upload_scene<F>(s, dc, ds);
        if (rc != B2VP_OK) return rc;
    }
    StreamSpace sp;
    int rc = make_stream_space(s, a->seed_mode, a->n_streams, a->explicit_samples_per_stream, sp);
    if (rc != B2VP_OK) return rc;
    rc = set_stream_range(a, sp);
    if (rc != B2VP_OK) return rc;
    const uint64_t count = sp.d.count;
    if (a->seed_mode == B2VP_SEED_EXPLICIT && count &&
        (!a->explicit_state || !a->explicit_inc || !a->explicit_pixel))
        return fail(B2VP_ERR_INVALID, "explicit seeding needs state, inc and pixel arrays");
    cudaStream_t stream = (cudaStream_t) a->cuda_stream;

    const int W = sp.d.w, H = sp.d.h;
    const size_t film_n = (size_t) W * H * (s->desc.integrator.moment ? B2VP_FILM_CHANNELS_MOMENT : B2VP_FILM_CHANNELS), film_bytes = film_n * sizeof(F);
    RenderOut<F> out{};
    LaunchPlan pl;
    rc = plan_launch<F>(s, dc, ds, a, sp.d, film_bytes, out, pl);
    if (rc != B2VP_OK) return rc;

    F *d_film;
    if (a->film_on_device) {
        d_film = (F *) film_user;
        if (a->clear_film) CU_CHECK(cudaMemsetAsync(d_film, 0, film_bytes, stream));
    } else {
        if (dc->film_buf.bytes < film_bytes) CU_CHECK(dc->film_buf.alloc(film_bytes));
        d_film = dc->film_buf.as<F>();
        if (a->clear_film) CU_CHECK(cudaMemsetAsync(d_film, 0, film_bytes, stream));
        else CU_CHECK(cudaMemcpyAsync(d_film, film_user, film_bytes, cudaMemcpyHostToDevice, stream));
    }

    if (!dc->totals_buf.p) CU_CHECK(dc->totals_buf.alloc((TOT_N + 32) * sizeof(unsigned long long)));
    if (!dc->counter_buf.p) CU_CHECK(dc->counter_buf.alloc(sizeof(unsigned long long)));
    CU_CHECK(cudaMemsetAsync(dc->totals_buf.p, 0, (TOT_N + 32) * sizeof(unsigned long long), stream));
    CU_CHECK(cudaMemsetAsync(dc->counter_buf.p, 0, sizeof(unsigned long long), stream));

    if (a->seed_mode == B2VP_SEED_SCALAR_BLOCKS) {
        CU_CHECK(dc->blocks_buf.upload(sp.blocks.data(), sp.blocks.size() * sizeof(SpiralBlockD)));
        sp.d.blocks = dc->blocks_buf.as<SpiralBlockD>();
    }
    if (a->seed_mode == B2VP_SEED_EXPLICIT && count) {
        for (uint64_t k = 0; k < count; ++k)
            if (a->explicit_pixel[k] >= (uint32_t) (W * H))
                return fail(B2VP_ERR_INVALID, "explicit_pixel[%llu] outside the film", (unsigned long long) k);
        CU_CHECK(dc->ex_state.upload(a->explicit_state, count * 8));
        CU_CHECK(dc->ex_inc.upload(a->explicit_inc, count * 8));
        CU_CHECK(dc->ex_pixel.upload(a->explicit_pixel, count * 4));
        sp.d.ex_state = dc->ex_state.as<uint64_t>(); sp.d.ex_inc = dc->ex_inc.as<uint64_t>();
        sp.d.ex_pixel = dc->ex_pixel.as<uint32_t>();
    }

    out.film = d_film;
    out.totals = dc->totals_buf.as<unsigned long long>();
    const bool want_qstats = (a->debug_flags & 1u) != 0;
    if (want_qstats) out.q_stats = out.totals + TOT_N;
    out.work_counter = dc->counter_buf.as<unsigned long long>();
    const uint32_t sps = sp.d.samples_per_stream;
    if (a->out_state && count) { CU_CHECK(dc->o_state.alloc(count * 8)); out.out_state = dc->o_state.as<uint64_t>(); }
    if (a->out_counters && count) {
        static_assert(sizeof(PathCounters) == sizeof(B2vpPathCounters), "counter layout");
        CU_CHECK(dc->o_counters.alloc(count * sizeof(PathCounters)));
        out.out_counters = dc->o_counters.as<PathCounters>();
    }
    if (a->out_radiance && count) {
        CU_CHECK(dc->o_radiance.alloc(count * sps * sizeof(F)));
        CU_CHECK(cudaMemsetAsync(dc->o_radiance.p, 0, count * sps * sizeof(F), stream));
        out.out_radiance = dc->o_radiance.as<F>();
    }
    if (pl.queue_kernel) {
        /* cold per-path state: one PathCold array per CTA in global memory (L2-resident) */
        if (dc->q_cold.bytes < pl.cold_bytes) CU_CHECK(dc->q_cold.alloc(pl.cold_bytes));
        out.q_cold = dc->q_cold.p;
    }
    { std::lock_guard<std::mutex> lock(s->mutex); s->last_kernel = pl.name; }
    out.stop_flag = dc->stop_flag.template as<int>();
    CU_CHECK(cudaMemsetAsync(dc->stop_flag.p, 0, sizeof(int), stream));

    uint32_t launches = 0;
    const auto t_start = std::chrono::steady_clock::now();
    CU_CHECK(cudaEventRecord(dc->ev0, stream));
    if (count) {
        CU_CHECK(launch_kernel<F>(pl, ds, sp.d, out, stream, dc, (a->debug_flags & 2u) != 0));
        launches++;
    }
    CU_CHECK(cudaEventRecord(dc->ev1, stream));
    CU_CHECK(cudaGetLastError());
    /* watch for cancel / timeout while the kernel runs -- BEFORE any read-back is queued: a device-to-host copy into
       pageable memory blocks the host until the kernel is done */
    uint32_t stopped = 0;
    rc = wait_for_render(s, dc, stream, a->timeout_s, t_start, &stopped);
    if (rc != B2VP_OK) return rc;

    unsigned long long tot[TOT_N + 32];
    CU_CHECK(cudaMemcpyAsync(tot, dc->totals_buf.p, sizeof(tot), cudaMemcpyDeviceToHost, stream));
    if (!a->film_on_device)
        CU_CHECK(cudaMemcpyAsync(film_user, d_film, film_bytes, cudaMemcpyDeviceToHost, stream));
    if (out.out_state) CU_CHECK(cudaMemcpyAsync(a->out_state, out.out_state, count * 8, cudaMemcpyDeviceToHost, stream));
    if (out.out_counters)
        CU_CHECK(cudaMemcpyAsync(a->out_counters, out.out_counters, count * sizeof(PathCounters),
                                 cudaMemcpyDeviceToHost, stream));
    if (out.out_radiance)
        CU_CHECK(cudaMemcpyAsync(a->out_radiance, out.out_radiance, count * sps * sizeof(F), cudaMemcpyDeviceToHost, stream));
    CU_CHECK(cudaStreamSynchronize(stream));
    if (want_qstats) {
        const unsigned long long *q = tot + TOT_N;
        static const char *names[PH_N] = {"STEP", "SCATTER", "SURFACE", "MARCH", "END"};
        std::fprintf(stderr, "[b2vp qstats] phase switches %llu, idle polls %llu\n", q[16], q[17]);
        for (int p = 0; p < PH_N; ++p)
            std::fprintf(stderr, "[b2vp qstats]   %-8s batches %10llu  slots %12llu  fill %.1f\n", names[p], q[p], q[8 + p],
                         q[p] ? (double) q[8 + p] / (double) q[p] : 0.0);
    }
    if (a->totals) {
        float ms = 0;
        CU_CHECK(cudaEventElapsedTime(&ms, dc->ev0, dc->ev1));
        fill_totals(*a->totals, tot, ms, launches);
        a->totals->stopped_early = stopped;
    }
    if (stopped == 2)
        return fail(B2VP_ERR_CANCELLED, "render cancelled (%llu of %llu paths were traced)", (unsigned long long) tot[TOT_PATHS],
                    (unsigned long long) count * std::max<uint32_t>(sps, 1));
    return B2VP_OK;
}

/* ---- multi-device render (B2vpRenderArgs::devices) ---------------------------------------------------------
   Paths are independent and the only shared state is the film: the stream range is dealt out in
   interleaved blocks (shard_streams), every device renders its share into a private film from its own host thread
   (scene replica uploaded on first use), and one ncclReduce(sum) over NVLink merges the films on the root. */
template <typename F> int render_multi(b2vp_scene *s, const B2vpRenderArgs *a, void *film_user) {
    const uint32_t n = a->n_devices;
    if (!a->devices || n > 64) return fail(B2VP_ERR_INVALID, "bad device list");
    if (a->devices[0] != s->device)
        return fail(B2VP_ERR_INVALID, "devices[0] must be the scene's device %d (the root of the film reduce)", s->device);
    for (uint32_t i = 0; i < n; ++i)
        for (uint32_t j = 0; j < i; ++j)
            if (a->devices[i] == a->devices[j]) return fail(B2VP_ERR_INVALID, "device %d listed twice", a->devices[i]);
    if (a->seed_mode == B2VP_SEED_EXPLICIT || a->out_state || a->out_counters || a->out_radiance || a->interleave_block)
        return fail(B2VP_ERR_UNSUPPORTED, "multi-device render: explicit seeding, per-stream outputs and the caller's own "
                                          "interleave_* are single-device features");
    std::vector<DeviceCtx *> ctx(n);
    std::vector<int> devs(a->devices, a->devices + n);
    for (uint32_t i = 0; i < n; ++i) {
        int rc = context_for(s, devs[i], &ctx[i]);
        if (rc != B2VP_OK) return rc;
    }
    StreamSpace sp;
    int rc = make_stream_space(s, a->seed_mode, 0, 0, sp);
    if (rc != B2VP_OK) return rc;
    const uint64_t first = a->first_stream;
    if (first > sp.n_streams || (a->n_streams && first + a->n_streams > sp.n_streams))
        return fail(B2VP_ERR_INVALID, "stream range out of bounds");
    const uint64_t count = a->n_streams ? a->n_streams : sp.n_streams - first;
    const size_t film_n = (size_t) sp.d.w * sp.d.h * (s->desc.integrator.moment ? B2VP_FILM_CHANNELS_MOMENT : B2VP_FILM_CHANNELS);
    const size_t film_bytes = film_n * sizeof(F);

    /* films: the root's is the caller's device film, or a staging buffer for a host film */
    std::vector<F *> film(n);
    std::vector<cudaStream_t> stream(n);
    for (uint32_t i = 0; i < n; ++i) {
        CU_CHECK(cudaSetDevice(devs[i]));
        stream[i] = i == 0 ? (cudaStream_t) a->cuda_stream : ctx[i]->work_stream;
        if (i == 0 && a->film_on_device) { film[0] = (F *) film_user; continue; }
        if (ctx[i]->film_buf.bytes < film_bytes) CU_CHECK(ctx[i]->film_buf.alloc(film_bytes));
        film[i] = ctx[i]->film_buf.template as<F>();
        if (i == 0 && !a->clear_film)
            CU_CHECK(cudaMemcpyAsync(film[0], film_user, film_bytes, cudaMemcpyHostToDevice, stream[0]));
    }
    std::vector<ncclComm_t> *comms = nullptr;
    rc = nccl_comms_for(devs, &comms);
    if (rc != B2VP_OK) return rc;

    std::vector<B2vpCounters> tot(n);
    std::vector<int> rcs(n, B2VP_OK);
    std::vector<std::string> errs(n);
    auto work = [&](uint32_t i) {
        B2vpRenderArgs ai = *a;
        uint64_t B = 0, ci = 0;
        shard_streams(count, n, i, &B, &ci);
        ai.devices = nullptr; ai.n_devices = 0; ai.device = devs[i];
        ai.film_on_device = 1;
        ai.clear_film = i == 0 ? a->clear_film : 1u;
        ai.first_stream = first; ai.n_streams = ci;
        ai.interleave_block = B; ai.interleave_count = n; ai.interleave_index = i;
        ai.totals = &tot[i];
        ai.cuda_stream = (void *) stream[i];
        std::memset(&tot[i], 0, sizeof(B2vpCounters));
        if (ci == 0) {   /* nothing to render here: the film only has to be zero for the reduce */
            if (cudaSetDevice(devs[i]) != cudaSuccess || (ai.clear_film && cudaMemsetAsync(film[i], 0, film_bytes, stream[i]) != cudaSuccess)) {
                rcs[i] = B2VP_ERR_CUDA; errs[i] = "clearing the film of an idle device failed";
            }
            return;
        }
        rcs[i] = render_typed<F>(s, ctx[i], &ai, film[i]);
        if (rcs[i] != B2VP_OK) errs[i] = g_error;   /* thread-local: carry it to the caller's thread */
    };
    std::vector<std::thread> pool;
    for (uint32_t i = 1; i < n; ++i) pool.emplace_back(work, i);
    work(0);
    for (auto &t : pool) t.join();
    bool cancelled = false;
    for (uint32_t i = 0; i < n; ++i) {
        if (rcs[i] == B2VP_ERR_CANCELLED) { cancelled = true; continue; }
        if (rcs[i] != B2VP_OK) return fail(rcs[i], "device %d: %s", devs[i], errs[i].c_str());
    }

    /* the single collective of this path: sum the films onto the root, in place */
    const ncclDataType_t dt = sizeof(F) == 4 ? ncclFloat32 : ncclFloat64;
    NCCL_CHECK(g_nccl.GroupStart());
    for (uint32_t i = 0; i < n; ++i) {
        CU_CHECK(cudaSetDevice(devs[i]));
        NCCL_CHECK(g_nccl.Reduce(film[i], film[i], film_n, dt, ncclSum, 0, (*comms)[i], stream[i]));
    }
    NCCL_CHECK(g_nccl.GroupEnd());
    for (uint32_t i = n; i-- > 0;) {
        CU_CHECK(cudaSetDevice(devs[i]));
        if (i == 0 && !a->film_on_device)
            CU_CHECK(cudaMemcpyAsync(film_user, film[0], film_bytes, cudaMemcpyDeviceToHost, stream[0]));
        CU_CHECK(cudaStreamSynchronize(stream[i]));
    }
    if (a->totals) {
        B2vpCounters &t = *a->totals;
        std::memset(&t, 0, sizeof(t));
        for (uint32_t i = 0; i < n; ++i) {
            t.n_paths += tot[i].n_paths; t.n_rng_draws += tot[i].n_rng_draws; t.n_medium_steps += tot[i].n_medium_steps;
            t.n_null += tot[i].n_null; t.n_real += tot[i].n_real; t.n_intersect += tot[i].n_intersect;
            t.n_surface += tot[i].n_surface; t.n_invalid_samples += tot[i].n_invalid_samples;
            t.n_bvh_nodes += tot[i].n_bvh_nodes; t.n_bvh_tris += tot[i].n_bvh_tris;
            t.gpu_ms = std::max(t.gpu_ms, tot[i].gpu_ms);     /* the devices run side by side */
            t.n_launches += tot[i].n_launches;
            t.stopped_early = std::max(t.stopped_early, tot[i].stopped_early);
        }
    }
    if (cancelled) return fail(B2VP_ERR_CANCELLED, "render cancelled");
    return B2VP_OK;
}

} // namespace

/* ---- render without device allocation (b2vp_render_noalloc) ---------------------------------------------
   Scratch of the render kernels (launch totals, stream work counter, spiral block table, cold path records of
   the queue kernel) in ONE file-scope device array instead of cudaMalloc'ed buffers.  Every GPU that loads the
   library holds B2VP_NOALLOC_SCRATCH_BYTES of it until exit.  In host code the array's name is only a stand-in
   for the symbol: the device address comes from cudaGetSymbolAddress, once per device with that device current
   (noalloc_scratch).  All calls on one device share the array and it keeps the last call's contents: what must
   start at zero (totals, work counter) is cleared on the launch stream; calls that overlap on the GPU race. */
static __device__ __align__(256) unsigned char g_noalloc_scratch[B2VP_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mutex;
unsigned char *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

int noalloc_scratch(int device, unsigned char **out) {
    if (device < 0 || device >= NOALLOC_MAX_DEVICES)
        return fail(B2VP_ERR_INVALID, "b2vp_render_noalloc: device ordinal %d out of range", device);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        CU_CHECK(cudaSetDevice(device));
        void *p = nullptr;
        CU_CHECK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) return fail(B2VP_ERR_CUDA, "cudaGetSymbolAddress returned a null scratch address");
        g_noalloc_ptr[device] = (unsigned char *) p;
    }
    *out = g_noalloc_ptr[device];
    return B2VP_OK;
}

template <typename F> int render_noalloc_typed(b2vp_scene *s, DeviceCtx *dc, const B2vpRenderArgs *a, void *film_device) {
    CU_CHECK(cudaSetDevice(dc->device));
    DeviceScene<F> &ds = dev_of<F>(dc);
    if (!ds.ready)
        return fail(B2VP_ERR_INVALID, "b2vp_render_noalloc: the scene is not resident on the device in this precision "
                                      "yet; call b2vp_render once at set-up (it allocates), then this entry point");
    if (!a->film_on_device)
        return fail(B2VP_ERR_INVALID, "b2vp_render_noalloc: the film must be a device buffer of the caller (film_on_device = 1)");
    if (a->seed_mode == B2VP_SEED_EXPLICIT || a->out_state || a->out_counters || a->out_radiance)
        return fail(B2VP_ERR_UNSUPPORTED, "b2vp_render_noalloc: explicit seeding and per-stream outputs need device "
                                          "staging buffers; use b2vp_render");
    StreamSpace sp;
    int rc = make_stream_space(s, a->seed_mode, a->n_streams, a->explicit_samples_per_stream, sp);
    if (rc != B2VP_OK) return rc;
    rc = set_stream_range(a, sp);
    if (rc != B2VP_OK) return rc;
    const uint64_t count = sp.d.count;
    cudaStream_t stream = (cudaStream_t) a->cuda_stream;
    const size_t film_n = (size_t) sp.d.w * sp.d.h * (s->desc.integrator.moment ? B2VP_FILM_CHANNELS_MOMENT : B2VP_FILM_CHANNELS);
    const size_t film_bytes = film_n * sizeof(F);
    F *d_film = (F *) film_device;
    RenderOut<F> out{};
    out.film = d_film;
    LaunchPlan pl;
    rc = plan_launch<F>(s, dc, ds, a, sp.d, film_bytes, out, pl, false);
    if (rc != B2VP_OK) return rc;

    /* scratch layout inside the static array: [totals + work counter][spiral blocks][cold path records] */
    const size_t head_bytes = align256((TOT_N + 32 + 1) * sizeof(unsigned long long));
    const size_t blocks_bytes = a->seed_mode == B2VP_SEED_SCALAR_BLOCKS ? align256(sp.blocks.size() * sizeof(SpiralBlockD)) : 0;
    const size_t need = head_bytes + blocks_bytes + pl.cold_bytes;
    if (need > (size_t) B2VP_NOALLOC_SCRATCH_BYTES)
        return fail(B2VP_ERR_UNSUPPORTED,
                    "b2vp_render_noalloc: this launch needs %zu bytes of scratch (%zu header, %zu spiral blocks, %zu path "
                    "records), the static scratch array holds B2VP_NOALLOC_SCRATCH_BYTES = %zu; use b2vp_render",
                    need, head_bytes, blocks_bytes, pl.cold_bytes, (size_t) B2VP_NOALLOC_SCRATCH_BYTES);
    unsigned char *scr = nullptr;
    rc = noalloc_scratch(dc->device, &scr);
    if (rc != B2VP_OK) return rc;
    out.totals = reinterpret_cast<unsigned long long *>(scr);
    out.work_counter = out.totals + TOT_N + 32;
    CU_CHECK(cudaMemsetAsync(scr, 0, head_bytes, stream));   /* the shared array keeps the last call's contents */
    if (blocks_bytes) {
        CU_CHECK(cudaMemcpyAsync(scr + head_bytes, sp.blocks.data(), sp.blocks.size() * sizeof(SpiralBlockD),
                                 cudaMemcpyHostToDevice, stream));
        sp.d.blocks = reinterpret_cast<const SpiralBlockD *>(scr + head_bytes);
    }
    if (pl.queue_kernel) out.q_cold = scr + head_bytes + blocks_bytes;   /* every field is written before it is read */
    if (a->clear_film) CU_CHECK(cudaMemsetAsync(d_film, 0, film_bytes, stream));
    { std::lock_guard<std::mutex> lock(s->mutex); s->last_kernel = pl.name; }

    uint32_t launches = 0;
    CU_CHECK(cudaEventRecord(dc->ev0, stream));
    if (count) {
        CU_CHECK(launch_kernel<F>(pl, ds, sp.d, out, stream));
        launches++;
    }
    CU_CHECK(cudaEventRecord(dc->ev1, stream));
    CU_CHECK(cudaGetLastError());
    unsigned long long tot[TOT_N];
    CU_CHECK(cudaMemcpyAsync(tot, out.totals, sizeof(tot), cudaMemcpyDeviceToHost, stream));
    CU_CHECK(cudaStreamSynchronize(stream));   /* also keeps sp.blocks alive until the copy above has run */
    if (a->totals) {
        float ms = 0;
        CU_CHECK(cudaEventElapsedTime(&ms, dc->ev0, dc->ev1));
        fill_totals(*a->totals, tot, ms, launches);
    }
    return B2VP_OK;
}

} // namespace

/* ======================================================================== */
extern "C" {

uint32_t b2vp_abi_version(void) { return B2VP_ABI_VERSION; }

int b2vp_scene_info(const b2vp_scene_t *scene, B2vpSceneInfo *info) {
    if (!scene || !info) return fail(B2VP_ERR_INVALID, "null argument");
    if (info->abi_version != B2VP_ABI_VERSION || info->struct_size != sizeof(B2vpSceneInfo))
        return fail(B2VP_ERR_INVALID, "B2vpSceneInfo ABI mismatch");
    const BvhStats &b = scene->bvh_stats;
    info->n_triangles = scene->faces.size() / 3;
    info->bvh_nodes = b.n_nodes; info->bvh_leaves = b.n_leaves; info->bvh_max_depth = b.max_depth;
    info->bvh_forced_median_splits = b.n_forced_median; info->bvh_degenerate_splits = b.n_degenerate_median;
    info->bvh_stack_capacity = BVH_STACK_CAPACITY;
    std::memset(info->last_kernel, 0, sizeof(info->last_kernel));
    std::strncpy(info->last_kernel, scene->last_kernel.c_str(), sizeof(info->last_kernel) - 1);
    return B2VP_OK;
}

int b2vp_render_noalloc(b2vp_scene_t *scene, const B2vpRenderArgs *args, void *film_device) {
    if (!scene || !args || !film_device) return fail(B2VP_ERR_INVALID, "null argument");
    if (args->abi_version != B2VP_ABI_VERSION || args->struct_size != sizeof(B2vpRenderArgs))
        return fail(B2VP_ERR_INVALID, "B2vpRenderArgs ABI mismatch");
    if (args->device != scene->device)
        return fail(B2VP_ERR_INVALID, "scene lives on device %d, render asked for %d", scene->device, args->device);
    if (args->n_devices > 1) return fail(B2VP_ERR_UNSUPPORTED, "b2vp_render_noalloc renders on the scene's device only");
    DeviceCtx *dc = scene->contexts[0].get();
    if (args->precision == B2VP_PREC_F32) return render_noalloc_typed<float>(scene, dc, args, film_device);
    if (args->precision == B2VP_PREC_F64) return render_noalloc_typed<double>(scene, dc, args, film_device);
    return fail(B2VP_ERR_INVALID, "unknown precision %d", args->precision);
}

const char *b2vp_last_error(void) { return g_error.c_str(); }

int b2vp_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int b2vp_scene_create(const B2vpSceneDesc *desc, int32_t device, b2vp_scene_t **out) {
    if (!desc || !out) return fail(B2VP_ERR_INVALID, "null argument");
    *out = nullptr;
    if (desc->abi_version != B2VP_ABI_VERSION || desc->struct_size != sizeof(B2vpSceneDesc))
        return fail(B2VP_ERR_INVALID, "B2vpSceneDesc ABI mismatch (version %u size %u, library %u/%zu)",
                    desc->abi_version, desc->struct_size, B2VP_ABI_VERSION, sizeof(B2vpSceneDesc));
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(B2VP_ERR_NO_DEVICE, "no CUDA device available (%s); the B200 volpath has no CPU fallback",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
    }
    if (device < 0 || device >= n) return fail(B2VP_ERR_INVALID, "device ordinal %d out of range", device);
    std::unique_ptr<b2vp_scene> s(new b2vp_scene());
    s->device = device;
    int rc = copy_desc(desc, s.get());
    if (rc != B2VP_OK) return rc;
    rc = build_bvh(s.get());
    if (rc != B2VP_OK) return rc;
    DeviceCtx *dc = nullptr;
    rc = context_for(s.get(), device, &dc);   /* contexts[0]: the scene's own device */
    if (rc != B2VP_OK) return rc;
    *out = s.release();
    return B2VP_OK;
}

void b2vp_scene_destroy(b2vp_scene_t *scene) {
    if (!scene) return;
    for (auto &c : scene->contexts) {
        cudaSetDevice(c->device);
        if (c->ev0) cudaEventDestroy(c->ev0);
        if (c->ev1) cudaEventDestroy(c->ev1);
        if (c->ev_done) cudaEventDestroy(c->ev_done);
        if (c->ctl_stream) cudaStreamDestroy(c->ctl_stream);
        if (c->work_stream) cudaStreamDestroy(c->work_stream);
    }
    cudaSetDevice(scene->device);
    for (size_t i = 0; i < scene->volume_pinned.size(); ++i)
        if (scene->volume_pinned[i]) cudaHostUnregister(scene->volume_data[i].data());
    delete scene;
}

int b2vp_stream_count(const b2vp_scene_t *scene, int32_t seed_mode, uint64_t *n_streams, uint32_t *samples_per_stream) {
    if (!scene || !n_streams || !samples_per_stream) return fail(B2VP_ERR_INVALID, "null argument");
    StreamSpace sp;
    int rc = make_stream_space(scene, seed_mode, 0, 0, sp);
    if (rc != B2VP_OK) return rc;
    *n_streams = sp.n_streams;
    *samples_per_stream = sp.d.samples_per_stream;
    return B2VP_OK;
}

int b2vp_scene_update(b2vp_scene_t *scene, const B2vpParamPatch *patch) {
    if (!scene || !patch) return fail(B2VP_ERR_INVALID, "null argument");
    if (patch->abi_version != B2VP_ABI_VERSION || patch->struct_size != sizeof(B2vpParamPatch))
        return fail(B2VP_ERR_INVALID, "B2vpParamPatch ABI mismatch");
    if (patch->medium >= (int) scene->media.size() || patch->bsdf >= (int) scene->bsdfs.size() ||
        patch->emitter >= (int) scene->emitters.size() || patch->volume >= (int) scene->volumes.size())
        return fail(B2VP_ERR_INVALID, "patch index out of range");
    if (patch->medium >= 0) scene->media[patch->medium].scale = patch->medium_scale;
    if (patch->bsdf >= 0) std::memcpy(scene->bsdfs[patch->bsdf].p, patch->bsdf_p, sizeof(patch->bsdf_p));
    if (patch->emitter >= 0) scene->emitters[patch->emitter].irradiance = patch->emitter_irradiance;
    if (patch->volume >= 0) {
        if (scene->volumes[patch->volume].type != B2VP_VOLUME_CONST)
            return fail(B2VP_ERR_INVALID, "volume patch targets a grid volume");
        scene->volumes[patch->volume].value = patch->volume_value;
    }
    /* only the small tables change: rebuild and re-upload the blob, grids and BVH stay resident (on every device the
       scene has been rendered on) */
    auto patch_dev = [&](auto &ds) -> int {
        using DS = std::remove_reference_t<decltype(ds)>;
        using F = decltype(ds.g.bs_radius);
        (void) sizeof(DS);
        if (!ds.ready) return B2VP_OK;
        unsigned char *b = ds.host_blob.data();
        auto *media = reinterpret_cast<DMedium<F> *>(b + ds.g.ofs_media);
        auto *volumes = reinterpret_cast<DVolume<F> *>(b + ds.g.ofs_volumes);
        auto *bsdfs = reinterpret_cast<DBsdf<F> *>(b + ds.g.ofs_bsdfs);
        auto *emitters = reinterpret_cast<DEmitter<F> *>(b + ds.g.ofs_emitters);
        if (patch->medium >= 0) {
            DMedium<F> &m = media[patch->medium];
            m.scale = (F) patch->medium_scale;
            if (m.type == B2VP_MEDIUM_HETEROGENEOUS) m.max_density = m.scale * volumes[m.sigma_t].max_value;
        }
        if (patch->bsdf >= 0) for (int k = 0; k < 4; ++k) bsdfs[patch->bsdf].p[k] = (F) patch->bsdf_p[k];
        if (patch->emitter >= 0) emitters[patch->emitter].irradiance = (F) patch->emitter_irradiance;
        if (patch->volume >= 0) volumes[patch->volume].value = (F) patch->volume_value;
        CU_CHECK(cudaMemcpy(ds.blob.p, b, ds.host_blob.size(), cudaMemcpyHostToDevice));
        return B2VP_OK;
    };
    for (auto &c : scene->contexts) {
        CU_CHECK(cudaSetDevice(c->device));
        int rc = patch_dev(c->dev32);
        if (rc != B2VP_OK) return rc;
        rc = patch_dev(c->dev64);
        if (rc != B2VP_OK) return rc;
    }
    return B2VP_OK;
}

int b2vp_scene_update_grid(b2vp_scene_t *scene, int32_t volume, const float *data, uint64_t n_values, int32_t keep_max) {
    if (!scene || !data) return fail(B2VP_ERR_INVALID, "null argument");
    if (volume < 0 || volume >= (int) scene->volumes.size())
        return fail(B2VP_ERR_INVALID, "volume index out of range");
    B2vpVolume &hv = scene->volumes[volume];
    if (hv.type != B2VP_VOLUME_GRID) return fail(B2VP_ERR_INVALID, "grid update targets a constant volume");
    std::vector<float> &held = scene->volume_data[volume];
    if (n_values != held.size())   /* grid3d.cpp:373-383 also takes a doubled resolution; this path does not */
        return fail(B2VP_ERR_UNSUPPORTED, "Unsupported GridVolume data size update: %llu -> %llu",
                    (unsigned long long) held.size(), (unsigned long long) n_values);
    CU_CHECK(cudaSetDevice(scene->device));
    /* The host copy of the grid (kept for a device instance of the other precision built later) is page-locked at
       the first update, so that sweeps upload at PCIe rate instead of through the driver's pageable staging; the
       copy into it and the maximum (grid3d.cpp:387-390: follows the data unless max_value was given) are one
       pass over the caller's array on a few host threads. */
    if (scene->volume_pinned.size() != scene->volume_data.size()) scene->volume_pinned.assign(scene->volume_data.size(), 0);
    if (!scene->volume_pinned[volume] && n_values >= (1u << 18)) {
        if (cudaHostRegister(held.data(), held.size() * sizeof(float), cudaHostRegisterDefault) == cudaSuccess)
            scene->volume_pinned[volume] = 1;
        else
            cudaGetLastError();   /* stays pageable: slower, not wrong */
    }
    {
        const unsigned hw = std::thread::hardware_concurrency();
        const size_t n_thr = n_values < (1u << 20) ? 1 : std::min<size_t>(8, hw ? hw : 1);
        std::vector<float> part(n_thr, data[0]);
        auto work = [&](size_t t) {
            const size_t lo = n_values * t / n_thr, hi = n_values * (t + 1) / n_thr;
            std::memcpy(held.data() + lo, data + lo, (hi - lo) * sizeof(float));
            float m[8];
            for (int q = 0; q < 8; ++q) m[q] = data[lo];
            size_t i = lo;
            for (; i + 8 <= hi; i += 8)
                for (int q = 0; q < 8; ++q) m[q] = data[i + q] > m[q] ? data[i + q] : m[q];
            for (; i < hi; ++i) m[0] = data[i] > m[0] ? data[i] : m[0];
            for (int q = 1; q < 8; ++q) m[0] = m[q] > m[0] ? m[q] : m[0];
            part[t] = m[0];
        };
        std::vector<std::thread> pool;
        for (size_t t = 1; t < n_thr; ++t) pool.emplace_back(work, t);
        work(0);
        for (auto &th : pool) th.join();
        if (!keep_max) {
            float mx = part[0];
            for (float x : part) mx = x > mx ? x : mx;
            hv.max_value = mx;
        }
    }
    int cur_sm_count = 0;
    auto patch_dev = [&](auto &ds) -> int {
        using F = decltype(ds.g.bs_radius);
        if (!ds.ready) return B2VP_OK;
        unsigned char *b = ds.host_blob.data();
        auto *media = reinterpret_cast<DMedium<F> *>(b + ds.g.ofs_media);
        auto *volumes = reinterpret_cast<DVolume<F> *>(b + ds.g.ofs_volumes);
        DVolume<F> &v = volumes[volume];
        if (sizeof(F) == sizeof(float)) {
            CU_CHECK(cudaMemcpy((void *) v.data, held.data(), held.size() * sizeof(float), cudaMemcpyHostToDevice));
        } else {
            std::vector<F> wide(held.begin(), held.end());
            CU_CHECK(cudaMemcpy((void *) v.data, wide.data(), wide.size() * sizeof(F), cudaMemcpyHostToDevice));
        }
        if (v.cells) {   /* the trilinear cell records are derived from the grid: rebuild them in place */
            const size_t n_cells = (size_t) (v.nx + 1) * (v.ny + 1) * (v.nz + 1);
            const int blocks = (int) std::min<size_t>((n_cells + 255) / 256, (size_t) cur_sm_count * 16);
            build_cells_kernel<F><<<blocks, 256>>>(v.data, const_cast<F *>(v.cells), v.nx, v.ny, v.nz);
            CU_CHECK(cudaGetLastError());
        }
        v.max_value = (F) hv.max_value;
        ds.mg_request = 0; ds.mg_bytes = 0;   /* majorant grids follow the data: rebuilt at their next use */
        for (size_t i = 0; i < scene->media.size(); ++i) {
            DMedium<F> &m = media[i];
            if (m.type == B2VP_MEDIUM_HETEROGENEOUS && m.sigma_t == volume) m.max_density = m.scale * v.max_value;
        }
        CU_CHECK(cudaMemcpy(ds.blob.p, b, ds.host_blob.size(), cudaMemcpyHostToDevice));
        CU_CHECK(cudaDeviceSynchronize());
        return B2VP_OK;
    };
    for (auto &c : scene->contexts) {
        CU_CHECK(cudaSetDevice(c->device));
        cur_sm_count = c->sm_count;
        int rc = patch_dev(c->dev32);
        if (rc != B2VP_OK) return rc;
        rc = patch_dev(c->dev64);
        if (rc != B2VP_OK) return rc;
    }
    return B2VP_OK;
}

int b2vp_render(b2vp_scene_t *scene, const B2vpRenderArgs *args, void *film) {
    if (!scene || !args || !film) return fail(B2VP_ERR_INVALID, "null argument");
    if (args->abi_version != B2VP_ABI_VERSION || args->struct_size != sizeof(B2vpRenderArgs))
        return fail(B2VP_ERR_INVALID, "B2vpRenderArgs ABI mismatch");
    if (args->device != scene->device)
        return fail(B2VP_ERR_INVALID, "scene lives on device %d, render asked for %d", scene->device, args->device);
    /* like m_stop = false at the top of SamplingIntegrator::render (integrator.cpp:59): a cancel only stops the render
       that is running when it arrives, and the handle stays usable afterwards */
    scene->cancelled.store(0);
    if (args->n_devices > 1) {
        if (args->precision == B2VP_PREC_F32) return render_multi<float>(scene, args, film);
        if (args->precision == B2VP_PREC_F64) return render_multi<double>(scene, args, film);
        return fail(B2VP_ERR_INVALID, "unknown precision %d", args->precision);
    }
    if (args->n_devices == 1 && args->devices && args->devices[0] != scene->device)
        return fail(B2VP_ERR_INVALID, "devices[0] must be the scene's device %d", scene->device);
    DeviceCtx *dc = scene->contexts[0].get();
    if (args->precision == B2VP_PREC_F32) return render_typed<float>(scene, dc, args, film);
    if (args->precision == B2VP_PREC_F64) return render_typed<double>(scene, dc, args, film);
    return fail(B2VP_ERR_INVALID, "unknown precision %d", args->precision);
}

int b2vp_shard_streams(uint64_t count, uint32_t n, uint32_t index, uint64_t *block, uint64_t *count_i) {
    if (!block || !count_i || n == 0 || index >= n) return fail(B2VP_ERR_INVALID, "bad sharding request");
    shard_streams(count, n, index, block, count_i);
    return B2VP_OK;
}

int b2vp_measure_fp32_peak(int32_t device, double *tflops) {
    if (!tflops) return fail(B2VP_ERR_INVALID, "null argument");
    CU_CHECK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU_CHECK(cudaGetDeviceProperties(&prop, device));
    DevBuf sink;
    CU_CHECK(sink.alloc(sizeof(float) * 1024));
    cudaEvent_t e0, e1;
    CU_CHECK(cudaEventCreate(&e0));
    CU_CHECK(cudaEventCreate(&e1));
    const int iters = 1 << 15, blocks = prop.multiProcessorCount * 8, threads = 256;
    double best = 0;
    for (int rep = 0; rep < 4; ++rep) {
        CU_CHECK(cudaEventRecord(e0));
        fma_peak_kernel<<<blocks, threads>>>(sink.as<float>(), iters, 1.0001f);
        CU_CHECK(cudaEventRecord(e1));
        CU_CHECK(cudaEventSynchronize(e1));
        float ms = 0;
        CU_CHECK(cudaEventElapsedTime(&ms, e0, e1));
        double flop = 2.0 * 16 * (double) iters * blocks * threads;
        if (rep > 0) best = std::max(best, flop / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *tflops = best;
    return B2VP_OK;
}

void b2vp_cancel(b2vp_scene_t *scene) {
    if (scene) scene->cancelled.store(1);
}

} /* extern "C" */
