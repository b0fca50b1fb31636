/*
 * py_b2vp.cpp -- pybind11 module `_b2vp`: `load_dict` and `render` in C++ over include/b2vp.hpp.
 *
 * The reference's Python front end is a pybind11 module: `mitsuba.core.xml.load_dict` turns nested dictionaries into
 * plugin instances (src/libcore/python/xml_v.cpp:100-240) and `Integrator.render(scene, sensor)` releases the GIL,
 * routes SIGINT to `Integrator::cancel` and calls the C++ render (src/librender/python/integrator_v.cpp:124-153).
 * This module is the same pair for the B200 path: `load_dict` walks the dictionary in C++ and drives the
 * `SceneBuilder` of b2vp.hpp (so it covers the plugin types that header covers; the ctypes host in scene.py covers
 * the rest: ply / serialized meshes, blendphase, distantflux, irradiancemeter, crop windows), `DeviceScene.render`
 * calls `b2vp_render` with the GIL released and Ctrl-C wired to `b2vp_cancel`.  Built by __graft_entry__.build()
 * with g++ against the installed pybind11 3; no CUDA in this file.
 */
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <atomic>
#include <csignal>
#include <functional>
#include <map>

#define B2VP_WITH_ZLIB 1   /* the .serialized reader (zlib) */
#include "b2vp.hpp"

namespace py = pybind11;
using namespace pybind11::literals;

namespace {

[[noreturn]] void unsupported(const std::string &kind, const std::string &type) {
    throw std::runtime_error(kind + " plugin \"" + type + "\" is not supported by the C++ host of the B200 volpath "
                             "(include/b2vp.hpp); the ctypes host (load_dict of the package) may cover it");
}

std::string type_of(const py::dict &d) {
    if (!d.contains("type")) throw std::runtime_error("load_dict: missing key 'type'");
    return py::cast<std::string>(d["type"]);
}
template <typename T> T get(const py::dict &d, const char *key, T def) {
    return d.contains(key) ? py::cast<T>(d[key]) : def;
}
std::array<double, 3> vec3(const py::handle &h) {
    auto v = py::cast<std::vector<double>>(h);
    if (v.size() != 3) throw std::runtime_error("load_dict: expected a 3-vector");
    return { v[0], v[1], v[2] };
}

template <typename F> struct Loader {
    b2vp::SceneBuilder<F> sb;
    std::map<std::string, int> media;

    /* a Transform of the package (attribute `matrix`) or anything numpy turns into a 4 x 4 array */
    b2vp::Mat4<F> xf(const py::dict &d, const char *key = "to_world") {
        b2vp::Mat4<F> m;
        if (!d.contains(key)) return m;
        py::object o = d[key];
        if (py::hasattr(o, "matrix")) o = o.attr("matrix");
        auto a = py::array_t<double, py::array::c_style | py::array::forcecast>::ensure(o);
        if (!a || a.ndim() != 2 || a.shape(0) != 4 || a.shape(1) != 4) throw std::runtime_error("load_dict: to_world must be 4 x 4");
        for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) m(r, c) = (F) a.at(r, c);
        return m;
    }
    b2vp::Phase phase(const py::dict &d) {
        const std::string t = type_of(d);
        if (t == "isotropic") return b2vp::Phase::isotropic();
        if (t == "rayleigh") return b2vp::Phase::rayleigh();
        if (t == "hg") return b2vp::Phase::hg(get<double>(d, "g", 0.8));
        if (t == "tabphase") return b2vp::Phase::tabulated(py::cast<std::vector<double>>(d["values"]));
        if (t == "blendphase") {   /* blendphase.cpp:42-66: the two nested phase functions in key order + `weight` */
            std::map<std::string, py::dict> kids;
            for (auto kv : d) {
                const std::string k = py::cast<std::string>(kv.first);
                if (k != "weight" && py::isinstance<py::dict>(kv.second)) kids.emplace(k, py::cast<py::dict>(kv.second));
            }
            if (kids.size() != 2) throw std::runtime_error("BlendPhase: Two child phase functions must be specified!");
            if (!d.contains("weight")) throw std::runtime_error("Property \"weight\" has not been specified!");
            auto it = kids.begin();
            const b2vp::Phase a = phase(it->second), b = phase((++it)->second);
            if (!py::isinstance<py::dict>(d["weight"])) return b2vp::Phase::blend(a, b, py::cast<double>(d["weight"]));
            const py::dict w = py::cast<py::dict>(d["weight"]);
            if (type_of(w) == "constvolume") return b2vp::Phase::blend(a, b, get<double>(w, "color", 1.0));   /* constant3d.cpp:14 */
            if (type_of(w) != "gridvolume") unsupported("volume", type_of(w));
            b2vp::Phase p = b2vp::Phase::blend(a, b, 0.0);
            const b2vp::GridVolume g = grid(w);
            if (g.channels != 1) throw std::runtime_error("blendphase: the weight grid must have one channel");
            p.weight_grid = g.data; p.weight_res[0] = g.nx; p.weight_res[1] = g.ny; p.weight_res[2] = g.nz;
            p.weight_filter = g.filter_type; p.weight_wrap = g.wrap_mode;
            const b2vp::Mat4<F> tw = xf(w);
            for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) p.weight_to_world[(size_t) (r * 4 + c)] = (double) tw(r, c);
            return p;
        }
        unsupported("phase function", t);
    }
    b2vp::GridVolume grid(const py::dict &d) {
        b2vp::GridVolume g;
        if (d.contains("filename")) {
            std::array<float, 6> file_bbox{ 0, 0, 0, 1, 1, 1 };
            g = b2vp::read_vol(py::cast<std::string>(d["filename"]), &file_bbox);
            g.bbox = file_bbox;
        } else {   /* in-memory data[z, y, x(, c)] */
            auto a = py::array_t<float, py::array::c_style | py::array::forcecast>::ensure(d["data"]);
            if (!a || (a.ndim() != 3 && a.ndim() != 4)) throw std::runtime_error("gridvolume: data must be [z, y, x] or [z, y, x, c]");
            g.nz = (int) a.shape(0); g.ny = (int) a.shape(1); g.nx = (int) a.shape(2); g.channels = a.ndim() == 4 ? (int) a.shape(3) : 1;
            g.data.assign(a.data(), a.data() + a.size());
        }
        g.filter_type = get<std::string>(d, "filter_type", "trilinear");
        g.wrap_mode = get<std::string>(d, "wrap_mode", "clamp");
        if (d.contains("max_value")) g.max_value = py::cast<double>(d["max_value"]);
        if (get<bool>(d, "raw", false)) throw std::runtime_error("gridvolume raw=true is not supported in mono variants");
        g.use_grid_bbox = get<bool>(d, "use_grid_bbox", false);
        if (!d.contains("filename") && d.contains("bbox")) {   /* in-memory data: the bounding box a file would carry */
            const auto bb = py::cast<std::vector<float>>(d["bbox"]);
            if (bb.size() != 6) throw std::runtime_error("gridvolume: bbox must be (min xyz, max xyz)");
            for (size_t i = 0; i < 6; ++i) g.bbox[i] = bb[i];
        }
        return g;
    }
    void medium(const std::string &id, const py::dict &d) {
        const std::string t = type_of(d);
        if (!d.contains("phase")) throw std::runtime_error("medium \"" + id + "\": no phase function");
        const b2vp::Phase ph = phase(py::cast<py::dict>(d["phase"]));
        const bool spectral = get<bool>(d, "has_spectral_extinction", true), se = get<bool>(d, "sample_emitters", true);
        const double scale = get<double>(d, "scale", 1.0);
        /* props.volume(): a number, or a `constvolume` (its `color`, constant3d.cpp:14); grids are for heterogeneous media */
        auto constant = [&](const char *key, double fallback) -> double {
            if (!d.contains(key) || !py::isinstance<py::dict>(d[key])) return get<double>(d, key, fallback);
            const py::dict v = py::cast<py::dict>(d[key]);
            if (type_of(v) != "constvolume") unsupported("volume of a homogeneous medium", type_of(v));
            return get<double>(v, "color", 1.0);
        };
        if (t == "homogeneous") {
            media[id] = sb.add_medium(b2vp::Homogeneous{ constant("sigma_t", 1.0), constant("albedo", 0.75), scale, spectral, se }, ph);
        } else if (t == "heterogeneous") {
            if (!py::isinstance<py::dict>(d["sigma_t"])) throw std::runtime_error("heterogeneous: sigma_t must be a gridvolume");
            const py::dict st = py::cast<py::dict>(d["sigma_t"]);
            if (type_of(st) != "gridvolume") unsupported("volume", type_of(st));
            const b2vp::Heterogeneous h{ scale, spectral, se };
            if (d.contains("albedo") && py::isinstance<py::dict>(d["albedo"])) {
                const py::dict al = py::cast<py::dict>(d["albedo"]);
                if (type_of(al) == "gridvolume")
                    media[id] = sb.add_medium(h, grid(st), xf(st), grid(al), xf(al), ph);
                else if (type_of(al) == "constvolume")
                    media[id] = sb.add_medium(h, grid(st), xf(st), get<double>(al, "color", 1.0), ph);   /* constant3d.cpp:14 */
                else unsupported("volume", type_of(al));
            } else {
                media[id] = sb.add_medium(h, grid(st), xf(st), get<double>(d, "albedo", 0.75), ph);
            }
        } else unsupported("medium", t);
    }
    int medium_ref(const py::dict &d, const char *key) {
        if (!d.contains(key)) return -1;
        const py::dict r = py::cast<py::dict>(d[key]);
        if (type_of(r) != "ref") throw std::runtime_error("load_dict: nested media are not supported by the C++ host; use {\"type\": \"ref\", \"id\": ...}");
        const std::string id = py::cast<std::string>(r["id"]);
        auto it = media.find(id);
        if (it == media.end()) throw std::runtime_error("load_dict: reference to unknown medium \"" + id + "\"");
        return it->second;
    }
    int bsdf(const py::dict &shape) {
        if (!shape.contains("bsdf")) return sb.add_bsdf(b2vp::Diffuse{});   /* the reference's default BSDF */
        const py::dict d = py::cast<py::dict>(shape["bsdf"]);
        const std::string t = type_of(d);
        if (t == "diffuse") return sb.add_bsdf(b2vp::Diffuse{ get<double>(d, "reflectance", 0.5) });
        if (t == "null") return sb.add_bsdf(b2vp::Null{});
        if (t == "rpv") return sb.add_bsdf(b2vp::Rpv{ get<double>(d, "rho_0", 0.1), get<double>(d, "k", 0.1), get<double>(d, "g", 0.0), get<double>(d, "rho_c", -1.0) });
        if (t == "bilambertian") return sb.add_bsdf(b2vp::Bilambertian{ get<double>(d, "reflectance", 0.5), get<double>(d, "transmittance", 0.5) });
        unsupported("BSDF", t);
    }
    std::map<std::string, int> shapes;   /* id (or key) -> shape index, for sensors that refer to a shape */
    py::dict shape_sensor; int shape_sensor_index = -1; bool have_shape_sensor = false;
    /* a shape given to a sensor (distant.cpp:270-290): a reference to a rectangle of the scene, or a rectangle of the
       sensor's own, which is used by the sensor only (sample_only) */
    int sensor_shape(const py::dict &d, const char *key) {
        const py::dict r = py::cast<py::dict>(d[key]);
        if (type_of(r) == "ref") {
            auto it = shapes.find(py::cast<std::string>(r["id"]));
            if (it == shapes.end()) throw std::runtime_error(std::string("load_dict: ") + key + " refers to an unknown shape (rectangles only)");
            return it->second;
        }
        if (type_of(r) != "rectangle") throw std::runtime_error("sensor target shape must be a rectangle in the B200 volpath");
        return sb.add_rectangle(xf(r), bsdf(r), medium_ref(r, "interior"), medium_ref(r, "exterior"), true, get<bool>(r, "flip_normals", false));
    }
    bool shape(const std::string &key, const py::dict &d) {
        const std::string t = type_of(d);
        if (t != "cube" && t != "rectangle" && t != "sphere" && t != "obj" && t != "ply" && t != "serialized" && t != "mesh") return false;
        const int b = bsdf(d), in = medium_ref(d, "interior"), ex = medium_ref(d, "exterior");
        if (t == "cube") sb.add_cube(xf(d), b, in, ex);
        else if (t == "rectangle") {
            const int index = sb.add_rectangle(xf(d), b, in, ex, false, get<bool>(d, "flip_normals", false));
            shapes[get<std::string>(d, "id", key)] = index;
            if (d.contains("sensor")) {   /* a sensor attached to the shape (shape.cpp:47-54): the irradiancemeter */
                if (have_shape_sensor) throw std::runtime_error("the B200 volpath renders one sensor per scene");
                shape_sensor = py::cast<py::dict>(d["sensor"]); shape_sensor_index = index; have_shape_sensor = true;
                if (type_of(shape_sensor) != "irradiancemeter") throw std::runtime_error("only an irradiancemeter on a rectangle is supported as a shape sensor");
            }
        }
        else if (t == "sphere") {
            b2vp::Sphere s;
            if (d.contains("center")) s.center = vec3(d["center"]);
            s.radius = get<double>(d, "radius", 1.0); s.flip_normals = get<bool>(d, "flip_normals", false);
            sb.add_sphere(s, xf(d), b, in, ex);
        } else if (t == "mesh") {   /* an in-memory triangle mesh: what the file loaders produce (src/librender/mesh.cpp) */
            b2vp::MeshData m;
            auto p = py::array_t<float, py::array::c_style | py::array::forcecast>::ensure(d["vertex_positions"]);
            auto f = py::array_t<uint32_t, py::array::c_style | py::array::forcecast>::ensure(d["faces"]);
            if (!p || !f || p.size() % 3 || f.size() % 3) throw std::runtime_error("mesh: vertex_positions / faces must be (n, 3) arrays");
            m.positions.assign(p.data(), p.data() + p.size());
            m.faces.assign(f.data(), f.data() + f.size());
            if (d.contains("vertex_normals")) {
                auto n = py::array_t<float, py::array::c_style | py::array::forcecast>::ensure(d["vertex_normals"]);
                if (n) m.normals.assign(n.data(), n.data() + n.size());
            }
            if (d.contains("vertex_texcoords")) {
                auto uv = py::array_t<float, py::array::c_style | py::array::forcecast>::ensure(d["vertex_texcoords"]);
                if (uv) m.texcoords.assign(uv.data(), uv.data() + uv.size());
            }
            sb.add_mesh(m, xf(d), b, in, ex);
        } else {   /* src/shapes/obj.cpp, ply.cpp, serialized.cpp */
            const std::string file = py::cast<std::string>(d["filename"]);
            b2vp::MeshData m = t == "obj" ? b2vp::load_obj(file, get<bool>(d, "flip_tex_coords", true))
                               : t == "ply" ? b2vp::load_ply(file) : b2vp::load_serialized(file, get<int>(d, "shape_index", 0));
            const bool face_normals = get<bool>(d, "face_normals", false);   /* mesh.cpp:30-31: no vertex normals at all */
            if (m.normals.empty() && !face_normals)
                throw std::runtime_error("mesh files without normals need face_normals=true (smooth vertex normals, "
                                         "Mesh::recompute_vertex_normals, are not built in the B200 volpath)");
            if (face_normals) m.normals.clear();
            sb.add_mesh(m, xf(d), b, in, ex);
        }
        return true;
    }
    void sensor(const py::dict &d) {
        const std::string t = type_of(d);
        b2vp::IndependentSampler smp;
        if (d.contains("sampler")) {
            const py::dict s = py::cast<py::dict>(d["sampler"]);
            if (type_of(s) != "independent") unsupported("sampler", type_of(s));
            smp.sample_count = get<uint32_t>(s, "sample_count", 4); smp.seed = get<uint64_t>(s, "seed", 0);
        }
        b2vp::HdrFilm film;
        if (d.contains("film")) {
            const py::dict f = py::cast<py::dict>(d["film"]);
            if (type_of(f) != "hdrfilm") unsupported("film", type_of(f));
            film.width = get<int>(f, "width", 768); film.height = get<int>(f, "height", 576);
            if (f.contains("rfilter")) {   /* src/rfilters/*.cpp */
                const py::dict r = py::cast<py::dict>(f["rfilter"]);
                const std::string rt = type_of(r);
                b2vp::RFilter &rf = film.rfilter;
                if (rt == "box") { rf.type = b2vp::RFilter::Box; rf.radius = (double) (float) get<double>(r, "radius", 0.5); }
                else if (rt == "tent") rf.type = b2vp::RFilter::Tent;
                else if (rt == "gaussian") { rf.type = b2vp::RFilter::Gaussian; rf.stddev = (double) (float) get<double>(r, "stddev", 0.5); }
                else if (rt == "mitchell") {
                    rf.type = b2vp::RFilter::Mitchell;
                    rf.B = (double) (float) get<double>(r, "B", 1.f / 3.f); rf.C = (double) (float) get<double>(r, "C", 1.f / 3.f);
                } else if (rt == "catmullrom") rf.type = b2vp::RFilter::CatmullRom;
                else if (rt == "lanczos") { rf.type = b2vp::RFilter::Lanczos; rf.lobes = get<int>(r, "lobes", 3); }
                else unsupported("reconstruction filter", rt);
            }
            film.crop_offset_x = get<int>(f, "crop_offset_x", 0); film.crop_offset_y = get<int>(f, "crop_offset_y", 0);
            film.crop_width = get<int>(f, "crop_width", film.width); film.crop_height = get<int>(f, "crop_height", film.height);
        }
        const int med = medium_ref(d, "medium");
        if (t == "distant") {
            b2vp::Distant s;
            if (d.contains("direction")) s.direction = vec3(d["direction"]);
            if (d.contains("orientation")) { s.has_orientation = true; s.orientation = vec3(d["orientation"]); }
            if (d.contains("to_world")) {
                if (d.contains("direction")) throw std::runtime_error("Only one of the parameters 'direction' and 'to_world' can be specified at the same time!");
                const b2vp::Mat4<F> tw = xf(d);
                s.has_to_world = true;
                for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) s.to_world[(size_t) (r * 4 + c)] = (double) tw(r, c);
            }
            if (d.contains("ray_target")) {
                if (py::isinstance<py::dict>(d["ray_target"])) s.ray_target_shape = sensor_shape(d, "ray_target");
                else { s.has_ray_target = true; s.ray_target = vec3(d["ray_target"]); }
            }
            if (d.contains("ray_origin")) s.ray_origin_shape = sensor_shape(d, "ray_origin");
            s.flip_directions = get<bool>(d, "flip_directions", false); s.medium = med;
            sb.set_sensor(s, film, smp);
        } else if (t == "distantflux") {   /* distantflux.cpp:140-182 */
            b2vp::DistantFlux s;
            const b2vp::Mat4<F> tw = xf(d);
            for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) s.to_world[(size_t) (r * 4 + c)] = (double) tw(r, c);
            if (d.contains("target")) {
                if (py::isinstance<py::dict>(d["target"])) s.target_shape = sensor_shape(d, "target");
                else { s.has_target = true; s.target = vec3(d["target"]); }
            }
            if (d.contains("origin")) s.origin_shape = sensor_shape(d, "origin");
            s.medium = med;
            sb.set_sensor(s, film, smp);
        } else if (t == "irradiancemeter") {   /* irradiancemeter.cpp:54-80 */
            if (d.contains("to_world"))
                throw std::runtime_error("Found a 'to_world' transformation -- this is not allowed. The irradiance meter inherits this transformation from its parent shape.");
            if (d.contains("film") && (film.width != 1 || film.height != 1)) throw std::runtime_error("This sensor only supports films of size 1x1 Pixels!");
            b2vp::IrradianceMeter s;
            s.shape = shape_sensor_index; s.medium = med;
            sb.set_sensor(s, smp, d.contains("film") ? &film : nullptr);
        } else if (t == "radiancemeter") {
            b2vp::RadianceMeter s;
            s.origin = vec3(d["origin"]); s.direction = vec3(d["direction"]); s.medium = med;
            sb.set_sensor(s, smp, d.contains("film") ? &film : nullptr);
        } else if (t == "mdistant" || t == "mradiancemeter") {
            auto triples = [](const py::handle &h) {
                auto v = py::cast<std::vector<double>>(h);
                if (v.size() % 3) throw std::runtime_error("load_dict: directions / origins must come in triples");
                std::vector<std::array<double, 3>> o;
                for (size_t i = 0; i < v.size(); i += 3) o.push_back({ v[i], v[i + 1], v[i + 2] });
                return o;
            };
            const auto dirs = triples(d["directions"]);
            if (d.contains("film") && (film.width != (int) dirs.size() || film.height != 1))
                throw std::runtime_error("Film size must be [sensor_count, 1]. Expected [" + std::to_string(dirs.size()) + ", 1], got [" +
                                         std::to_string(film.width) + ", " + std::to_string(film.height) + "]");
            if (t == "mdistant") {
                b2vp::MDistant s;
                s.directions = dirs; s.medium = med;
                if (d.contains("target")) {
                    if (py::isinstance<py::dict>(d["target"])) s.target_shape = sensor_shape(d, "target");
                    else { s.has_target = true; s.target = vec3(d["target"]); }
                }
                sb.set_sensor(s, smp, d.contains("film") ? &film : nullptr);
            } else {
                b2vp::MRadianceMeter s;
                s.directions = dirs; s.origins = triples(d["origins"]); s.medium = med;
                sb.set_sensor(s, smp, d.contains("film") ? &film : nullptr);
            }
        } else unsupported("sensor", t);
    }
    void integrator(py::dict d) {
        b2vp::Volpath v;
        const py::dict outer = d;   /* block_size / samples_per_pass: of the integrator render() runs on (integrator.cpp:51-110) */
        if (type_of(d) == "moment") {
            v.moment = true;
            for (auto kv : d)
                if (py::isinstance<py::dict>(kv.second)) { d = py::cast<py::dict>(kv.second); break; }
        }
        const std::string t = type_of(d);
        if (t == "volpathmis") v.mis = get<bool>(d, "use_spectral_mis", true) ? 1 : 2;
        else if (t != "volpath") unsupported("integrator", t);
        v.max_depth = get<int>(d, "max_depth", -1); v.rr_depth = get<int>(d, "rr_depth", 5);
        v.hide_emitters = get<bool>(d, "hide_emitters", false);
        v.block_size = get<uint32_t>(outer, "block_size", 0);
        if (outer.contains("samples_per_pass")) v.samples_per_pass = py::cast<uint32_t>(outer["samples_per_pass"]);
        sb.set_integrator(v);
    }

    std::unique_ptr<b2vp::Scene> load(const py::dict &root) {
        if (get<std::string>(root, "type", "scene") != "scene") throw std::runtime_error("load_dict: the top-level dictionary must be a scene");
        /* children are visited in KEY order, as Properties::objects() hands them to the Scene constructor (a
           std::map; src/libcore/properties.cpp): the order of shapes and emitters is part of the result (ties between
           coincident hits go to the later shape, emitters are picked by index).  Media first: shapes and sensors
           refer to them by id. */
        std::map<std::string, py::dict> items;
        for (auto kv : root)
            if (py::isinstance<py::dict>(kv.second)) items.emplace(py::cast<std::string>(kv.first), py::cast<py::dict>(kv.second));
        for (auto &kv : items) {
            const std::string t = type_of(kv.second);
            if (t == "homogeneous" || t == "heterogeneous") medium(kv.first, kv.second);
        }
        for (auto &kv : items) {
            const py::dict &d = kv.second;
            const std::string t = type_of(d);
            if (t == "homogeneous" || t == "heterogeneous") continue;
            if (shape(kv.first, d)) continue;
            if (t == "directional") {
                b2vp::Directional e;
                if (d.contains("to_world")) {
                    if (d.contains("direction")) throw std::runtime_error("Only one of the parameters 'direction' and 'to_world' can be specified at the same time!");
                    const b2vp::Mat4<F> tw = xf(d);
                    e.has_to_world = true;
                    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) e.to_world[(size_t) (r * 4 + c)] = (double) tw(r, c);
                }
                if (d.contains("direction")) e.direction = vec3(d["direction"]);
                e.irradiance = get<double>(d, "irradiance", 1.0);
                sb.add_emitter(e);
            } else if (t == "constant") sb.add_emitter(b2vp::Constant{ get<double>(d, "radiance", 1.0) });
            else if (t == "volpath" || t == "volpathmis" || t == "moment") integrator(d);
            else if (t == "distant" || t == "radiancemeter" || t == "mdistant" || t == "mradiancemeter" || t == "distantflux") continue;   /* below */
            else unsupported("scene", t);
        }
        /* the sensor last: it may refer to a shape whose key sorts after its own */
        for (auto &kv : items) {
            const std::string t = type_of(kv.second);
            if (t == "distant" || t == "radiancemeter" || t == "mdistant" || t == "mradiancemeter" || t == "distantflux") sensor(kv.second);
        }
        if (have_shape_sensor) sensor(shape_sensor);
        return sb.finish();
    }
};

/* integrator_v.cpp:124-153: while a render runs, SIGINT cancels it instead of killing the interpreter */
std::function<void()> g_sigint_handler;
void (*g_sigint_prev)(int) = nullptr;

struct PyScene {
    std::unique_ptr<b2vp::Scene> scene;
    std::string variant;
};

struct PyDeviceScene {
    std::unique_ptr<b2vp::DeviceScene> dev;
    int channels = B2VP_FILM_CHANNELS;
    bool f64 = false;

    py::tuple render(const std::string &seeding, const std::string &math, const std::vector<int32_t> &devices, double timeout,
                     uint32_t majorant_grid, uint64_t first_stream, uint64_t n_streams) {
        b2vp::RenderOptions opt;
        if (seeding == "wavefront") opt.seed_mode = B2VP_SEED_WAVEFRONT;
        else if (seeding == "scalar") opt.seed_mode = B2VP_SEED_SCALAR_BLOCKS;
        else throw std::runtime_error("seeding must be \"wavefront\" or \"scalar\"");
        if (math == "fast") opt.math = B2VP_MATH_FAST;
        else if (math != "exact") throw std::runtime_error("math must be \"exact\" or \"fast\"");
        opt.devices = devices; opt.timeout_s = timeout; opt.majorant_grid = majorant_grid;
        opt.first_stream = first_stream; opt.n_streams = n_streams;
        b2vp::Film film;
        {
            py::gil_scoped_release release;
            b2vp::DeviceScene *d = dev.get();
            g_sigint_handler = [d]() { d->cancel(); };
            g_sigint_prev = std::signal(SIGINT, [](int) {
                if (g_sigint_handler) { g_sigint_handler(); g_sigint_handler = std::function<void()>(); }
            });
            try {
                film = dev->render(opt);
            } catch (...) {
                std::signal(SIGINT, g_sigint_prev);
                g_sigint_handler = std::function<void()>();
                throw;
            }
            std::signal(SIGINT, g_sigint_prev);
            g_sigint_handler = std::function<void()>();
        }
        py::object arr;
        const std::vector<py::ssize_t> shape{ film.height, film.width, film.channels };
        if (film.precision == B2VP_PREC_F32) { py::array_t<float> a(shape); std::memcpy(a.mutable_data(), film.f32.data(), film.f32.size() * 4); arr = a; }
        else { py::array_t<double> a(shape); std::memcpy(a.mutable_data(), film.f64.data(), film.f64.size() * 8); arr = a; }
        const B2vpCounters &t = film.totals;
        py::dict tot("n_paths"_a = t.n_paths, "n_rng_draws"_a = t.n_rng_draws, "n_medium_steps"_a = t.n_medium_steps,
                     "n_null"_a = t.n_null, "n_real"_a = t.n_real, "n_intersect"_a = t.n_intersect, "n_surface"_a = t.n_surface,
                     "n_invalid_samples"_a = t.n_invalid_samples, "gpu_ms"_a = t.gpu_ms, "n_launches"_a = t.n_launches,
                     "stopped_early"_a = t.stopped_early);
        return py::make_tuple(arr, tot);
    }
};

} // namespace

PYBIND11_MODULE(_b2vp, m) {
    m.doc() = "pybind11 front end of the B200 volpath: load_dict / render over include/b2vp.hpp (libb2vp.so)";
    m.attr("abi_version") = b2vp_abi_version();
    py::register_exception<b2vp::Error>(m, "B2vpError", PyExc_RuntimeError);

    py::class_<PyScene>(m, "Scene", "A loaded scene (B2vpSceneDesc + the arrays it points to)")
        .def_property_readonly("variant", [](const PyScene &s) { return s.variant; })
        .def_property_readonly("film_size", [](const PyScene &s) { return py::make_tuple(s.scene->width(), s.scene->height()); })
        .def_property_readonly("desc_address", [](const PyScene &s) { return (uintptr_t) &s.scene->desc; },
                               "address of the B2vpSceneDesc (for ctypes callers; valid while this object lives)");

    m.def("load_dict", [](const py::dict &d, const std::string &variant) {
        PyScene s;
        s.variant = variant;
        if (variant == "scalar_mono") s.scene = Loader<float>().load(d);
        else if (variant == "scalar_mono_double") s.scene = Loader<double>().load(d);
        else throw std::runtime_error("variant must be scalar_mono or scalar_mono_double");
        return s;
    }, "dict"_a, "variant"_a = "scalar_mono",
    "mitsuba.core.xml.load_dict for the plugin types of include/b2vp.hpp (src/libcore/python/xml_v.cpp:100-240)");

    py::class_<PyDeviceScene>(m, "DeviceScene", "A scene resident on a CUDA device (b2vp_scene_create .. b2vp_scene_destroy)")
        .def(py::init([](PyScene &s, int device) {
            if (!s.scene) throw std::runtime_error("this Scene was already moved onto a device");
            PyDeviceScene d;
            d.channels = s.scene->desc.integrator.moment ? B2VP_FILM_CHANNELS_MOMENT : B2VP_FILM_CHANNELS;
            d.f64 = s.scene->desc.precision == B2VP_PREC_F64;
            d.dev.reset(new b2vp::DeviceScene(std::move(s.scene), device));
            return d;
        }), "scene"_a, "device"_a = 0)
        .def("render", &PyDeviceScene::render, "seeding"_a = "wavefront", "math"_a = "exact", "devices"_a = std::vector<int32_t>(),
             "timeout"_a = 0.0, "majorant_grid"_a = 0u, "first_stream"_a = 0ull, "n_streams"_a = 0ull,
             "Integrator.render (src/librender/python/integrator_v.cpp:124-153): GIL released, Ctrl-C cancels the render. "
             "Returns (film[h, w, channels], totals).")
        .def("cancel", [](PyDeviceScene &d) { d.dev->cancel(); }, "Integrator.cancel")
        .def("last_kernel", [](PyDeviceScene &d) { return std::string(d.dev->info().last_kernel); });
}
