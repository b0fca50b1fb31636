// TEST INFRASTRUCTURE -- driver for the UNMODIFIED reference (eradiate-kernel @ /root/reference), built by
// oracle/Makefile.ref into oracle/_ref/refprobe.  Nothing here is reference code: it is the ~100 lines a user
// of the reference's C++ API writes to load an XML scene and call Integrator::render (what
// src/mitsuba/mitsuba.cpp:80-110,123-230 does), plus a raw dump of the film storage instead of film->develop()
// (Bitmap needs libpng / libjpeg / OpenEXR, which are not built here).
//
//   refprobe <variant> <scene.xml> <out.bin> [-t threads] [-r repeats] [-Dkey=value ...]
//
// out.bin: int32 width, height, channels, is_double; float64 seconds of the last Integrator::render call;
// then width*height*channels film values (X, Y, Z, A, W; float or double) exactly as HDRFilm::m_storage holds
// them after SamplingIntegrator::render (src/librender/integrator.cpp:51-179, src/films/hdrfilm.cpp:190-211).
#include <mitsuba/core/argparser.h>
#include <mitsuba/core/filesystem.h>
#include <mitsuba/core/fresolver.h>
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

template <typename Float, typename Spectrum>
int run(Object *parsed, const char *out_path, int repeats) {
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
    double seconds = 0;
    for (int r = 0; r < repeats; ++r) {
        Timer timer;
        if (!integrator->render(scene, sensor))
            return 3;
        seconds = timer.value() * 1e-3;
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
    Class::static_initialization();
    Thread::static_initialization();
    Logger::static_initialization();
    librender_nop();
    int rc = 0;
    try {
        std::string mode = argv[1];
        int threads = 1, repeats = 1;
        xml::ParameterList params;
        for (int i = 4; i < argc; ++i) {
            if (!strcmp(argv[i], "-t") && i + 1 < argc)
                threads = atoi(argv[++i]);
            else if (!strcmp(argv[i], "-r") && i + 1 < argc)
                repeats = atoi(argv[++i]);
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
        if (!fr->contains(base))
            fr->append(base);
        fs::path scene_dir = fs::path(argv[2]).parent_path();
        if (!fr->contains(scene_dir))
            fr->append(scene_dir);
        ref<Object> parsed = xml::load_file(argv[2], mode, params, false);
        if (mode == "scalar_mono")
            rc = run<float, Color<float, 1>>(parsed.get(), argv[3], repeats);
        else if (mode == "scalar_mono_double")
            rc = run<double, Color<double, 1>>(parsed.get(), argv[3], repeats);
        else
            Throw("refprobe: variant %s is not built", mode);
    } catch (const std::exception &e) {
        fprintf(stderr, "refprobe: %s\n", e.what());
        rc = 1;
    }
    Logger::static_shutdown();
    Thread::static_shutdown();
    Class::static_shutdown();
    return rc;
}
