// rtr_cli — the crate's callers, rewritten against include/rtr.hpp (C++ host side over the C ABI).
//
//   rtr_cli render SCENE.json W H OUT.png [--seed S] [--spp N] [--exr OUT.exr] [--kernel auto|megakernel|wavefront]
//        what src/gui.rs:326 does with a loaded scene: Scene::from_file + render(&mut target, &scene), headless
//   rtr_cli lower simple|mesh|bvh_stress|SCENE.json [--assets DIR]     the lowered scene's counts and checksums (no GPU)
//   rtr_cli bench simple|mesh|bvh_stress [--assets DIR] [--iters N]
//        benches/simple_render_bench.rs:8-11, benches/mesh.rs:10-21, benches/bvh_tree_stress.rs:11-29 (criterion's
//        b.iter loop replaced by a plain timing loop; each iteration is one full render(), BVH build included, as
//        in the crate, plus one that reuses the uploaded scene)
// Every run prints one JSON line with the device time, ray count and an FNV-1a checksum of the RGBA8 image, so the
// tests can check that this path and the Python path produce the same bytes.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../include/rtr.hpp"

using namespace ray_tracing_rust;
using core::render::RenderTarget;
using core::scene::Scene;
using utils::Color;
using utils::Vec3;

static uint64_t fnv1a(const std::vector<uint8_t>& d) {
    uint64_t h = 1469598103934665603ull;
    for (uint8_t b : d) { h ^= b; h *= 1099511628211ull; }
    return h;
}
static double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
static const char* opt(int argc, char** argv, const char* name, const char* dflt) {
    for (int i = 0; i + 1 < argc; ++i) if (!std::strcmp(argv[i], name)) return argv[i + 1];
    return dflt;
}
static void report(const char* what, const RenderTarget& t, const rtr_stats& st, double wall_ms, double reuse_ms) {
    std::printf("{\"run\": \"%s\", \"width\": %zu, \"height\": %zu, \"device_ms\": %.4f, \"render_call_ms\": %.3f, "
                "\"rerender_ms\": %.3f, \"rays\": %llu, \"samples\": %llu, \"kernel\": %u, \"fnv1a\": \"%016llx\"}\n",
                what, t.width, t.height, st.device_ms, wall_ms, reuse_ms, (unsigned long long)st.rays,
                (unsigned long long)st.samples, st.kernel_used, (unsigned long long)fnv1a(t.data));
}

static int cmd_render(int argc, char** argv) {
    if (argc < 6) { std::fprintf(stderr, "usage: rtr_cli render SCENE.json W H OUT.png [--seed S] [--spp N] [--exr OUT.exr]\n"); return 2; }
    Scene scene = Scene::from_file(argv[2]);
    RenderTarget target((size_t)std::atol(argv[3]), (size_t)std::atol(argv[4]));
    const uint64_t seed = std::strtoull(opt(argc, argv, "--seed", "0"), nullptr, 10);
    if (const char* spp = opt(argc, argv, "--spp", nullptr)) scene.settings.samples_per_pixel = (uint32_t)std::atol(spp);
    const std::string k = opt(argc, argv, "--kernel", "auto");
    const uint32_t kernel = k == "megakernel" ? RTR_KERNEL_MEGAKERNEL : k == "wavefront" ? RTR_KERNEL_WAVEFRONT : RTR_KERNEL_AUTO;
    const char* exr = opt(argc, argv, "--exr", nullptr);
    core::render::Device dev(0);
    std::vector<float> accum;
    const double t0 = now_ms();
    const rtr_stats st = dev.render(target, scene, seed, exr ? &accum : nullptr, kernel);
    const double t1 = now_ms();
    target.save_png(argv[5]);
    if (exr) {  // linear radiance: sum / spp
        const float inv = 1.0f / (float)scene.settings.samples_per_pixel;
        for (size_t i = 0; i < accum.size(); i += 4) { accum[i] *= inv; accum[i + 1] *= inv; accum[i + 2] *= inv; accum[i + 3] = 1.0f; }
        check(rtr_write_exr(exr, accum.data(), (uint32_t)target.width, (uint32_t)target.height, 4));
    }
    report("render", target, st, t1 - t0, 0.0);
    return 0;
}

// the scenes of the crate's three benches; returns false for an unknown name
static bool bench_scene(const std::string& which, const std::string& assets, Scene& scene, size_t& W, size_t& H) {
    W = H = 100;
    if (which == "simple") {  // benches/simple_render_bench.rs:8-9
        scene = Scene::from_file(assets + "/scenes/simple.json");
    } else if (which == "mesh") {  // benches/mesh.rs:10-19
        scene.settings.max_ray_depth = 50;
        scene.settings.samples_per_pixel = 10;
        const auto m = scene.add_material(std::make_unique<materials::Dielectric>(1.5f));
        auto mesh = core::mesh::Mesh::from_file(assets + "/assets/monkey.obj", m);
        mesh->build_bvh();
        scene.add_object(std::move(mesh));
        W = H = 40;
    } else if (which == "bvh_stress") {  // benches/bvh_tree_stress.rs:11-26 (that bench times build_bvh only)
        scene.camera.lookfrom.y = -20.0f;
        scene.settings.max_ray_depth = 50;
        const auto m = scene.add_material(std::make_unique<materials::Dielectric>(1.5f));
        for (int x = -10; x < 10; ++x)
            for (int y = -10; y < 10; ++y)
                for (int z = -10; z < 10; ++z)
                    scene.add_object(std::make_unique<objects::Sphere>(Vec3{(float)x, (float)y, (float)z}, 0.5f, m));
    } else {
        return false;
    }
    return true;
}

template <typename T>
static uint64_t fnv_of(const std::vector<T>& v) {
    const uint8_t* p = reinterpret_cast<const uint8_t*>(v.data());
    return fnv1a(std::vector<uint8_t>(p, p + v.size() * sizeof(T)));
}

// rtr_cli lower NAME|SCENE.json: what lower() hands to rtr_scene_upload, no GPU involved
static int cmd_lower(int argc, char** argv) {
    if (argc < 3) { std::fprintf(stderr, "usage: rtr_cli lower simple|mesh|bvh_stress|SCENE.json [--assets DIR]\n"); return 2; }
    const std::string which = argv[2], assets = opt(argc, argv, "--assets", ".");
    Scene scene;
    size_t W, H;
    if (!bench_scene(which, assets, scene, W, H)) scene = Scene::from_file(which);
    const core::scene::Lowered l = scene.lower();
    uint64_t n_tri = 0;
    for (const rtr_mesh& m : l.b.meshes) n_tri += m.n_indices / 3;
    std::printf("{\"objects\": %zu, \"spheres\": %zu, \"meshes\": %zu, \"volumes\": %zu, \"materials\": %zu, \"triangles\": %llu, "
                "\"background\": %u, \"spp\": %u, \"depth\": %u, \"fnv_objects\": \"%016llx\", \"fnv_spheres\": \"%016llx\", "
                "\"fnv_materials\": \"%016llx\"}\n",
                l.b.objects.size(), l.b.spheres.size(), l.b.meshes.size(), l.b.volumes.size(), l.b.materials.size(),
                (unsigned long long)n_tri, l.desc.background.kind, scene.settings.samples_per_pixel, (unsigned)scene.settings.max_ray_depth,
                (unsigned long long)fnv_of(l.b.objects), (unsigned long long)fnv_of(l.b.spheres), (unsigned long long)fnv_of(l.b.materials));
    return 0;
}

static int cmd_bench(int argc, char** argv) {
    if (argc < 3) { std::fprintf(stderr, "usage: rtr_cli bench simple|mesh|bvh_stress [--assets DIR] [--iters N]\n"); return 2; }
    const std::string which = argv[2], assets = opt(argc, argv, "--assets", ".");
    const int iters = std::atoi(opt(argc, argv, "--iters", "10"));
    Scene scene;
    size_t W = 100, H = 100;
    if (!bench_scene(which, assets, scene, W, H)) {
        std::fprintf(stderr, "unknown bench '%s'\n", which.c_str());
        return 2;
    }
    RenderTarget target(W, H);
    core::render::Device dev(0);
    rtr_stats st{};
    dev.render(target, scene, 0);  // warm-up (context, module load)
    double full = 0.0, reuse = 0.0;
    for (int i = 0; i < iters; ++i) {
        const double t0 = now_ms();
        dev.upload(scene);  // the crate's render() rebuilds the BVH every call (render.rs:67)
        st = dev.render(target, scene, 0);
        const double t1 = now_ms();
        st = dev.render(target, scene, 0);
        const double t2 = now_ms();
        full += t1 - t0;
        reuse += t2 - t1;
    }
    report(which.c_str(), target, st, full / iters, reuse / iters);
    return 0;
}

int main(int argc, char** argv) {
    try {
        if (argc >= 2 && !std::strcmp(argv[1], "render")) return cmd_render(argc, argv);
        if (argc >= 2 && !std::strcmp(argv[1], "bench")) return cmd_bench(argc, argv);
        if (argc >= 2 && !std::strcmp(argv[1], "lower")) return cmd_lower(argc, argv);
        std::fprintf(stderr, "usage: rtr_cli render|bench|lower ...\n");
        return 2;
    } catch (const std::exception& e) {  // the crate would panic here
        std::fprintf(stderr, "panic: %s\n", e.what());
        return 101;  // Rust's panic exit code
    }
}
