// TEST INFRASTRUCTURE -- driver for the UNMODIFIED reference (eradiate-kernel @ /root/reference), built by
// oracle/Makefile.ref into oracle/_ref/refprobe.  Nothing here is reference code: it is the ~100 lines a user
// of the reference's C++ API writes to load an XML scene and call Integrator::render (what
// src/mitsuba/mitsuba.cpp:80-110,123-230 does), plus a raw dump of the film storage instead of film->develop()
// (Bitmap needs libpng / libjpeg / OpenEXR, which are not built here).
//
//   refprobe <variant> <scene.xml> <out.bin> [-t threads] [-r repeats] [-Dkey=value ...] [--describe-oracle lib]
//
// --describe-oracle <libb2vp_oracle.so>: the scene's integrator must be the product's plugin `volpath_b200`.
// Instead of rendering on a GPU, the plugin's describe() (object graph -> B2vpSceneDesc) is called through its
// test hook and the descriptor is rendered by the CPU oracle: a GPU-less check that the plugin reads the
// reference's object graph correctly (the film must equal the one the reference's own volpath gives).
//
// out.bin: int32 width, height, channels, is_double; float64 seconds of the last Integrator::render call;
// then width*height*channels film values (X, Y, Z, A, W; float or double) exactly as HDRFilm::m_storage holds
// them after SamplingIntegrator::render (src/librender/integrator.cpp:51-179, src/films/hdrfilm.cpp:190-211).
#include <chrono>
#include <mitsuba/core/argparser.h>
#include <mitsuba/core/filesystem.h>
#include <mitsuba/core/fresolver.h>
#include <mitsuba/core/jit.h>
#include <mitsuba/core/logger.h>
#include <mitsuba/core/thread.h>
#include <mitsuba/core/timer.h>
#include <mitsuba/core/util.h>
#include <mitsuba/core/xml.h>
#include <mitsuba/render/film.h>
#include <mitsuba/render/imageblock.h>
#include <mitsuba/render/integrator.h>
#include <mitsuba/render/scene.h>
#include <mitsuba/render/sensor.h>
#include <tbb/task_scheduler_init.h>
#include <cstdio>
#include <cstring>
#include <vector>
#include <dlfcn.h>
#include "b2vp.h"

// The film's accumulation buffer is a private member of the plugin class; the driver reads it in place.
// Only the class definition is needed here (nothing of it is instantiated in this executable except the
// member access): the object itself is created by plugins/hdrfilm.so, built from the same source.
#undef MTS_EXPORT_PLUGIN
#define MTS_EXPORT_PLUGIN(...)
#define private public
#define protected public
#include <films/hdrfilm.cpp>
#undef private
#undef protected

using namespace mitsuba;

struct OracleJob { void *oracle; const char *out_path; int rc; };

static void oracle_callback(const B2vpSceneDesc *desc, void *user) {
    OracleJob *job = (OracleJob *) user;
    auto render = (int (*)(const B2vpSceneDesc *, const B2vpRenderArgs *, void *)) dlsym(job->oracle, "b2vp_oracle_render");
    auto last_error = (const char *(*)()) dlsym(job->oracle, "b2vp_oracle_last_error");
    const bool f64 = desc->precision == B2VP_PREC_F64;
    int32_t hdr[4] = { desc->sensor.crop_size[0], desc->sensor.crop_size[1],
                       desc->integrator.moment ? B2VP_FILM_CHANNELS_MOMENT : B2VP_FILM_CHANNELS, (int32_t) f64 };
    size_t n = (size_t) hdr[0] * hdr[1] * hdr[2];
    std::vector<unsigned char> film(n * (f64 ? 8 : 4));
    B2vpRenderArgs a{};
    a.abi_version = B2VP_ABI_VERSION; a.struct_size = (uint32_t) sizeof(a);
    a.precision = desc->precision; a.seed_mode = B2VP_SEED_SCALAR_BLOCKS; a.clear_film = 1;
    job->rc = render(desc, &a, film.data());
    if (job->rc) { fprintf(stderr, "refprobe: oracle: %s\n", last_error ? last_error() : "?"); return; }
    FILE *f = fopen(job->out_path, "wb");
    double seconds = 0;
    fwrite(hdr, sizeof(hdr), 1, f); fwrite(&seconds, sizeof(double), 1, f); fwrite(film.data(), 1, film.size(), f);
    fclose(f);
}

template <typename Float, typename Spectrum>
int run(Object *parsed, const char *out_path, int repeats, const char *oracle_path) {
    using Scene_ = Scene<Float, Spectrum>;
    using HDR = HDRFilm<Float, Spectrum>;
    auto *scene = dynamic_cast<Scene_ *>(parsed);
    if (!scene)
        Throw("Root element of the input file must be a <scene> tag!");
    if (scene->sensors().empty())
        Throw("No sensor specified for scene");
    auto *sensor = scene->sensors()[0].get();
    auto *integrator = scene->integrator();
    if (!integrator)
        Throw("No integrator specified for scene");
    if (oracle_path) {
        if (std::string(integrator->class_()->name()) != "B200VolumetricPathIntegrator")
            Throw("--describe-oracle needs <integrator type=\"volpath_b200\"> (or type=\"volpath\" with --plugin-root)");
        fs::path plug = util::library_path().parent_path() / "plugins" / "volpath_b200.so";
        void *h = dlopen(plug.native().c_str(), RTLD_LAZY | RTLD_LOCAL);   /* already loaded: same handle */
        void *orc = dlopen(oracle_path, RTLD_NOW | RTLD_LOCAL);
        if (!h || !orc) Throw("dlopen failed: %s", dlerror());
        auto describe = (int (*)(void *, void *, void *, int, void (*)(const B2vpSceneDesc *, void *), void *))
            dlsym(h, "b2vp_plugin_describe");
        if (!describe) Throw("b2vp_plugin_describe not found");
        OracleJob job{ orc, out_path, 0 };
        describe((void *) scene, (void *) sensor, (void *) (Object *) integrator, sizeof(scalar_t<Float>) == 8,
                 oracle_callback, &job);
        return job.rc ? 4 : 0;
    }
    double seconds = 0;
    for (int r = 0; r < repeats; ++r) {
        /* the reference's Timer truncates to whole milliseconds: too coarse for a 12 ms render on the GPU */
        const auto t0 = std::chrono::steady_clock::now();
        if (!integrator->render(scene, sensor))
            return 3;
        seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    }
    auto *film = static_cast<HDR *>(sensor->film());
    const ImageBlock<Float, Spectrum> *st = film->m_storage.get();
    int32_t hdr[4] = { (int32_t) st->width(), (int32_t) st->height(), (int32_t) st->channel_count(),
                       (int32_t) (sizeof(scalar_t<Float>) == 8) };
    FILE *f = fopen(out_path, "wb");
    if (!f)
        Throw("cannot write %s", out_path);
    fwrite(hdr, sizeof(hdr), 1, f);
    fwrite(&seconds, sizeof(double), 1, f);
    size_t n = (size_t) hdr[0] * hdr[1] * hdr[2];
    // the block has a border of `border_size` pixels (0 for the box filter); copy the interior rows
    int b = st->border_size();
    int sw = hdr[0] + 2 * b;
    const scalar_t<Float> *data = st->data().data();
    for (int y = 0; y < hdr[1]; ++y)
        fwrite(data + ((size_t) (y + b) * sw + b) * hdr[2], sizeof(scalar_t<Float>), (size_t) hdr[0] * hdr[2], f);
    fclose(f);
    (void) n;
    return 0;
}

int main(int argc, char **argv) {
    if (argc < 4) {
        fprintf(stderr, "usage: refprobe <variant> <scene.xml> <out.bin> [-t threads] [-r repeats] [-Dk=v ...]\n");
        return 2;
    }
    Jit::static_initialization();
    Class::static_initialization();
    Thread::static_initialization();
    Logger::static_initialization();
    librender_nop();
    int rc = 0;
    try {
        std::string mode = argv[1];
        int threads = 1, repeats = 1;
        const char *oracle_path = nullptr, *plugin_root = nullptr;
        xml::ParameterList params;
        for (int i = 4; i < argc; ++i) {
            if (!strcmp(argv[i], "-t") && i + 1 < argc)
                threads = atoi(argv[++i]);
            else if (!strcmp(argv[i], "-r") && i + 1 < argc)
                repeats = atoi(argv[++i]);
            else if (!strcmp(argv[i], "--describe-oracle") && i + 1 < argc)
                oracle_path = argv[++i];
            else if (!strcmp(argv[i], "--plugin-root") && i + 1 < argc)   /* a directory searched BEFORE the reference's own */
                plugin_root = argv[++i];
            else if (!strncmp(argv[i], "-D", 2)) {
                std::string kv = argv[i] + 2;
                auto sep = kv.find('=');
                params.push_back({ kv.substr(0, sep), kv.substr(sep + 1) });
            }
        }
        __global_thread_count = (size_t) threads;
        tbb::task_scheduler_init init(threads);
        Thread::thread()->logger()->set_log_level(Warn);
        ref<Thread> thread = Thread::thread();
        ref<FileResolver> fr = thread->file_resolver();
        fs::path base = util::library_path().parent_path();   // plugins/ lives next to libmitsuba-core.so
        if (plugin_root)   /* <root>/plugins/volpath.so shadows the stock volpath.so (plugin.cpp:94-105 resolves in order) */
            fr->prepend(fs::path(plugin_root));
        if (!fr->contains(base))
            fr->append(base);
        fs::path scene_dir = fs::path(argv[2]).parent_path();
        if (!fr->contains(scene_dir))
            fr->append(scene_dir);
        ref<Object> parsed = xml::load_file(argv[2], mode, params, false);
        if (mode == "scalar_mono")
            rc = run<float, Color<float, 1>>(parsed.get(), argv[3], repeats, oracle_path);
        else if (mode == "scalar_mono_double")
            rc = run<double, Color<double, 1>>(parsed.get(), argv[3], repeats, oracle_path);
        else
            Throw("refprobe: variant %s is not built", mode);
    } catch (const std::exception &e) {
        fprintf(stderr, "refprobe: %s\n", e.what());
        rc = 1;
    }
    Logger::static_shutdown();
    Thread::static_shutdown();
    Class::static_shutdown();
    Jit::static_shutdown();
    return rc;
}
