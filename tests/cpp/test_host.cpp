// TEST CODE: builds two scenes with the C++ host layer (include/b2vp.hpp), renders them through the C-ABI
// and checks the film against the CPU oracle (oracle/libb2vp_oracle.so) bit for bit.
//   test_host cpu   -> no GPU needed: the product must refuse to run (B2VP_ERR_NO_DEVICE), the oracle renders
//                      the C++-built descriptors and the C1 radiance is checked against its known value
//   test_host gpu   -> product vs oracle, every film byte and every counter
#include "b2vp.hpp"

#include <cstdio>
#include <cstdlib>
#include <string>

extern "C" int b2vp_oracle_render(const B2vpSceneDesc *, const B2vpRenderArgs *, void *);

template <typename F> std::unique_ptr<b2vp::Scene> scene_c1(uint32_t spp) {
    using M = b2vp::Mat4<F>;
    b2vp::SceneBuilder<F> sb;
    int atm = sb.add_medium(b2vp::Homogeneous{ 0.5, 1.0 }, b2vp::Phase::rayleigh());
    sb.add_cube(M::translate(0, 0, (F) 0.45) * M::scale(1, 1, (F) 0.55), sb.add_bsdf(b2vp::Null{}), atm);
    sb.add_rectangle(M(), sb.add_bsdf(b2vp::Diffuse{ 0.5 }));
    sb.add_emitter(b2vp::Directional{ { 0, 0, -1 }, 1.0 });
    b2vp::Distant sn;
    sn.has_ray_target = true;
    sb.set_sensor(sn, b2vp::HdrFilm{ 1, 1 }, b2vp::IndependentSampler{ spp, 0 });
    return sb.finish();
}

template <typename F> std::unique_ptr<b2vp::Scene> scene_c2(uint32_t spp, int width) {
    using M = b2vp::Mat4<F>;
    b2vp::SceneBuilder<F> sb;
    b2vp::GridVolume g;
    g.nx = 1; g.ny = 1; g.nz = 200;
    for (int z = 0; z < g.nz; ++z) g.data.push_back((float) std::exp(-((z + 0.5) / g.nz) * 100.0 / 8.0));
    M box = M::translate(-10, -10, 0) * M::scale(20, 20, 1);
    int atm = sb.add_medium(b2vp::Heterogeneous{ 4.0 }, g, box, 1.0, b2vp::Phase::rayleigh());
    sb.add_cube(M::translate(0, 0, (F) 0.45) * M::scale(10, 10, (F) 0.55), sb.add_bsdf(b2vp::Null{}), atm);
    sb.add_rectangle(M::scale(10, 10, 1), sb.add_bsdf(b2vp::Rpv{ 0.1, 0.7, -0.1 }));
    sb.add_emitter(b2vp::Directional{ { 0.5, 0, -0.8660254037844386 }, 1.0 });
    b2vp::Distant sn;
    sn.has_orientation = true; sn.has_ray_target = true;
    sb.set_sensor(sn, b2vp::HdrFilm{ width, 1 }, b2vp::IndependentSampler{ spp, 0 });
    return sb.finish();
}

/* A scene through the loaders and plugins the builder gained in round 2: a `.vol` grid, an OBJ mesh under a
   transform, an analytic sphere, and an mdistant / mradiancemeter sensor.  tests/test_cpp_host.py builds the same
   scene with the Python host (`load_dict`) and compares the films. */
template <typename F> std::unique_ptr<b2vp::Scene> scene_ext(const std::string &dir, const std::string &sensor, uint32_t spp) {
    using M = b2vp::Mat4<F>;
    b2vp::SceneBuilder<F> sb;
    std::array<float, 6> bbox;
    b2vp::GridVolume g = b2vp::read_vol(dir + "/col.vol", &bbox);
    if (g.nx != 1 || g.ny != 1 || g.nz != 64 || bbox[5] != 1.0f) throw b2vp::Error(B2VP_ERR_INVALID, "unexpected col.vol");
    M box = M::translate(-10, -10, 0) * M::scale(20, 20, 1);
    int atm = sb.add_medium(b2vp::Heterogeneous{ 4.0 }, g, box, 1.0, b2vp::Phase::rayleigh());
    sb.add_cube(M::translate(0, 0, (F) 0.45) * M::scale(10, 10, (F) 0.55), sb.add_bsdf(b2vp::Null{}), atm);
    sb.add_rectangle(M::scale(10, 10, 1), sb.add_bsdf(b2vp::Rpv{ 0.1, 0.7, -0.1 }));
    sb.add_sphere(b2vp::Sphere{ { 0.5, -0.25, 0.5 }, 0.25 }, M(), sb.add_bsdf(b2vp::Diffuse{ 0.75 }));
    sb.add_mesh(b2vp::load_obj(dir + "/roof.obj"), M::translate((F) 0.25, (F) -0.5, (F) 0.125) * M::scale(2, (F) 1.5, 1),
                sb.add_bsdf(b2vp::Diffuse{ 0.75 }));
    sb.add_emitter(b2vp::Directional{ { 0.5, 0, -0.8660254037844386 }, 1.0 });
    const int n = 8;
    if (sensor == "mdistant") {
        b2vp::MDistant sn;
        for (int i = 0; i < n; ++i) {
            const double th = (i + 0.5) / n * 3.14159265358979323846;
            sn.directions.push_back({ -std::cos(th), 0.0, -std::sin(th) });
        }
        sn.has_target = true;
        sb.set_sensor(sn, b2vp::IndependentSampler{ spp, 0 });
    } else {
        b2vp::MRadianceMeter sn;
        for (int i = 0; i < n; ++i) {
            sn.origins.push_back({ -1.0 + 0.25 * i, 0.125 * i - 0.5, 2.0 });
            sn.directions.push_back({ 0.0625 * i, -0.03125 * i, -1.0 });
        }
        sb.set_sensor(sn, b2vp::IndependentSampler{ spp, 0 });
    }
    return sb.finish();
}

static b2vp::Film oracle_render(const b2vp::Scene &s, int seed_mode = B2VP_SEED_WAVEFRONT) {
    b2vp::Film f;
    f.width = s.width(); f.height = s.height(); f.precision = s.desc.precision;
    size_t n = (size_t) f.width * f.height * 5;
    void *buf;
    if (f.precision == B2VP_PREC_F32) { f.f32.assign(n, 0.f); buf = f.f32.data(); } else { f.f64.assign(n, 0.0); buf = f.f64.data(); }
    B2vpRenderArgs a{};
    a.abi_version = B2VP_ABI_VERSION; a.struct_size = sizeof(a); a.precision = f.precision;
    a.seed_mode = seed_mode; a.clear_film = 1; a.totals = &f.totals;
    if (b2vp_oracle_render(&s.desc, &a, buf) != 0) { std::fprintf(stderr, "oracle failed\n"); std::exit(2); }
    return f;
}

static int compare(const char *name, const b2vp::Film &a, const b2vp::Film &b) {
    /* same addends in a different order: compare the sums to summation accuracy, the event counters exactly */
    int bad = 0;
    for (int x = 0; x < a.width; ++x)
        for (int ch = 0; ch < 5; ++ch) {
            double u = a.at(x, 0, ch), v = b.at(x, 0, ch), tol = (a.precision == B2VP_PREC_F32 ? 2e-4 : 1e-12) * std::fabs(v) + 1e-30;
            if (std::fabs(u - v) > tol) bad++;
        }
    bad += a.totals.n_rng_draws != b.totals.n_rng_draws || a.totals.n_medium_steps != b.totals.n_medium_steps ||
           a.totals.n_null != b.totals.n_null || a.totals.n_real != b.totals.n_real ||
           a.totals.n_intersect != b.totals.n_intersect || a.totals.n_surface != b.totals.n_surface;
    std::printf("%s: draws %llu steps %llu isect %llu  -> %s\n", name, (unsigned long long) a.totals.n_rng_draws,
                (unsigned long long) a.totals.n_medium_steps, (unsigned long long) a.totals.n_intersect, bad ? "MISMATCH" : "ok");
    return bad;
}

int main(int argc, char **argv) {
    const std::string mode = argc > 1 ? argv[1] : "cpu";
    if (mode == "cpu") {
        auto s = scene_c1<double>(20000);
        b2vp::Film ref = oracle_render(*s);
        double L = ref.luminance(0, 0);
        std::printf("C1 (C++ host, oracle): L = %.5f\n", L);
        if (!(std::fabs(L - 0.1432) < 0.004)) { std::printf("unexpected C1 radiance\n"); return 1; }
        if (b2vp_device_count() == 0) {
            try {
                b2vp::DeviceScene dev(std::move(s), 0);
                std::printf("scene_create succeeded without a device?\n");
                return 1;
            } catch (const b2vp::Error &e) {
                std::printf("no device: %s (code %d)\n", e.what(), e.code());
                return e.code() == B2VP_ERR_NO_DEVICE ? 0 : 1;
            }
        }
        return 0;
    }
    if (mode == "dump") {
        /* test_host dump <f32|f64> <out.bin>: the C2-shaped scene of scene_c2 (1 sample per pixel, 256 pixels) in
           the reference's scalar seeding order, rendered by the oracle; tests/test_cpp_host.py compares the film
           with the one the reference binary renders from the same parameters (tests/golden/ref_golden.npz). */
        const bool f64 = argc > 2 && std::string(argv[2]) == "f64";
        auto s = f64 ? scene_c2<double>(1, 256) : scene_c2<float>(1, 256);
        b2vp::Film f = oracle_render(*s, B2VP_SEED_SCALAR_BLOCKS);
        FILE *out = std::fopen(argc > 3 ? argv[3] : "film.bin", "wb");
        if (!out) return 2;
        if (f64) std::fwrite(f.f64.data(), sizeof(double), f.f64.size(), out);
        else std::fwrite(f.f32.data(), sizeof(float), f.f32.size(), out);
        std::fclose(out);
        return 0;
    }
    if (mode == "ext") {
        /* test_host ext <dir> <f32|f64> <mdistant|mradiancemeter> <out.bin>: oracle film of scene_ext, plus
           <dir>/film.pfm and <dir>/film.exr developed by the C++ host */
        const std::string dir = argv[2], sensor = argv[4];
        const bool f64 = std::string(argv[3]) == "f64";
        auto s = f64 ? scene_ext<double>(dir, sensor, 64) : scene_ext<float>(dir, sensor, 64);
        b2vp::Film f = oracle_render(*s);
        FILE *out = std::fopen(argv[5], "wb");
        if (!out) return 2;
        if (f64) std::fwrite(f.f64.data(), sizeof(double), f.f64.size(), out);
        else std::fwrite(f.f32.data(), sizeof(float), f.f32.size(), out);
        std::fclose(out);
        f.write_pfm(dir + "/film.pfm");
        f.write_exr(dir + "/film.exr");
        return 0;
    }
    int bad = 0;
    if (argc > 2) {   /* gpu <dir>: the round-2 plugins and loaders on the device, both sensors, both precisions */
        const std::string dir = argv[2];
        for (const char *sensor : { "mdistant", "mradiancemeter" }) {
            { auto s = scene_ext<float>(dir, sensor, 256); b2vp::Film ref = oracle_render(*s); b2vp::DeviceScene dev(std::move(s));
              bad += compare((std::string("ext ") + sensor + " f32").c_str(), dev.render(), ref); }
            { auto s = scene_ext<double>(dir, sensor, 64); b2vp::Film ref = oracle_render(*s); b2vp::DeviceScene dev(std::move(s));
              bad += compare((std::string("ext ") + sensor + " f64").c_str(), dev.render(), ref); }
        }
        /* render options: the tolerance mode stays within 1 % of the exact film mean; BVH info is filled */
        auto s = scene_ext<float>(dir, "mdistant", 1024);
        b2vp::DeviceScene dev(std::move(s));
        b2vp::Film exact = dev.render();
        b2vp::RenderOptions opt; opt.math = B2VP_MATH_FAST;
        b2vp::Film fast = dev.render(opt);
        double me = 0, mf = 0;
        for (int x = 0; x < exact.width; ++x) { me += exact.luminance(x, 0); mf += fast.luminance(x, 0); }
        const B2vpSceneInfo info = dev.info();
        const bool ok = std::fabs(me - mf) < 0.01 * me && fast.totals.n_paths == exact.totals.n_paths && info.bvh_nodes > 0;
        std::printf("render options: exact mean %.6f fast mean %.6f, last kernel '%s', %u BVH nodes -> %s\n", me / exact.width,
                    mf / exact.width, info.last_kernel, (unsigned) info.bvh_nodes, ok ? "ok" : "MISMATCH");
        bad += !ok;
    }
    { auto s = scene_c1<float>(4096); b2vp::Film ref = oracle_render(*s); b2vp::DeviceScene dev(std::move(s)); bad += compare("C1 f32", dev.render(), ref); }
    { auto s = scene_c1<double>(2048); b2vp::Film ref = oracle_render(*s); b2vp::DeviceScene dev(std::move(s)); bad += compare("C1 f64", dev.render(), ref); }
    { auto s = scene_c2<float>(256, 16); b2vp::Film ref = oracle_render(*s); b2vp::DeviceScene dev(std::move(s)); bad += compare("C2 f32", dev.render(), ref); }
    { auto s = scene_c2<double>(128, 16); b2vp::Film ref = oracle_render(*s); b2vp::DeviceScene dev(std::move(s)); bad += compare("C2 f64", dev.render(), ref); }
    return bad ? 1 : 0;
}
