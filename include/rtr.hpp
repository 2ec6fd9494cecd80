// rtr.hpp — C++17 mirror of the RayTracingRust crate's public scene/render API over the C ABI of rtr.h.
//
// The reference is compiled code (Rust) with no plugin interface: its callers (`benches/*.rs`, `src/gui.rs`) build a
// `Scene` out of boxed trait objects and call `render(&mut target, &scene)`.  No Rust toolchain exists on the build
// or GPU boxes, so this header is the compiled-language host side: same names, argument meaning and error behaviour
// (where the crate panics, these throw), each class lowering itself into the plain `rtr_scene_desc` exactly as the
// `lower()` trait method of INTEGRATION.md section 3 would inside the crate.  Header-only; link librtr_b200.so.
//
//   crate item (file:line)                                   here
//   Scene, RenderSettings, RenderMode  lib/core/scene.rs:15-114   ray_tracing_rust::core::scene::*
//   Camera                             lib/core/camera.rs:27-34   core::camera::Camera
//   render, RenderTarget               lib/core/render.rs:14-127  core::render::*
//   Mesh                               lib/core/mesh.rs:63-121    core::mesh::Mesh
//   Sphere                             lib/objects.rs:15-20       objects::Sphere
//   Volume, Isotropic                  lib/volume.rs:19-31,75-87  volume::*
//   Lambertian .. MixMaterial          lib/materials.rs:18-156    materials::*
//   UniformBackground .. SkyMap        lib/backgrounds.rs:14-92   backgrounds::*
#ifndef RTR_HPP_
#define RTR_HPP_

#include <algorithm>
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "rtr.h"

namespace ray_tracing_rust {

// the crate panics; the mirror throws
struct Panic : std::runtime_error {
    rtr_status status;
    Panic(rtr_status s, const std::string& what) : std::runtime_error(what), status(s) {}
};
inline void check(rtr_status s, const rtr_context* ctx = nullptr) {
    if (s != RTR_OK) {
        const char* m = rtr_last_error(ctx);
        throw Panic(s, m ? m : "rtr error");
    }
}

namespace utils {
using Float = float;                       // lib/utils/types.rs:4
struct Vec3 { Float x = 0, y = 0, z = 0; };  // cgmath::Vector3<f32>
struct Color {                             // lib/utils/color.rs:6-11; Default is green (:43-47)
    Float r = 0, g = 1, b = 0;
    static Color new_(Float r, Float g, Float b) { return Color{r, g, b}; }
};
inline rtr_vec3 c(const Vec3& v) { return rtr_vec3{v.x, v.y, v.z}; }
inline rtr_vec3 c(const Color& v) { return rtr_vec3{v.r, v.g, v.b}; }
}  // namespace utils

namespace core {

namespace scene {
struct MaterialHandle { size_t index; };  // scene.rs:68
struct ObjectHandle { size_t index; };    // scene.rs:71
}  // namespace scene

// What `lower()` appends to: typed arrays + the object list
struct SceneBuilder {
    std::vector<rtr_object_ref> objects;
    std::vector<rtr_sphere> spheres;
    std::vector<rtr_mesh> meshes;
    std::vector<rtr_volume> volumes;
    std::vector<rtr_material> materials;
    rtr_background background{};
    size_t n_handles = 0;
    uint32_t material(scene::MaterialHandle h) const {
        if (h.index >= n_handles) throw std::out_of_range("material handle out of range");  // Vec index panic
        return (uint32_t)h.index;
    }
};

namespace traits {
struct Material {  // lib/core/traits.rs:22-24
    virtual ~Material() = default;
    // fills materials[slot]; a MixMaterial appends its children behind the handles
    virtual void lower(SceneBuilder& b, uint32_t slot) const = 0;
};
struct Object {  // lib/core/traits.rs:17-20
    virtual ~Object() = default;
    virtual rtr_object_ref lower(SceneBuilder& b) const = 0;
    virtual bool is_volume() const { return false; }
};
struct Background {  // lib/core/traits.rs:26-28
    virtual ~Background() = default;
    virtual void lower(SceneBuilder& b) const = 0;
};
}  // namespace traits

namespace camera {
struct Camera {  // camera.rs:27-34
    utils::Vec3 lookfrom{0, 0, 0}, lookat{0, 0, -1}, vertical{0, 1, 0};
    utils::Float vertical_fov = 90.0f, aspect_ratio = 1.0f;
};
}  // namespace camera

}  // namespace core

// ---------------------------------------------------------------------------------- materials (lib/materials.rs)
namespace materials {
using core::SceneBuilder;
using utils::Color;
using utils::Float;
inline rtr_material rec(uint32_t kind, Color c, Float param) { return rtr_material{kind, utils::c(c), param, 0u, 0u}; }

struct Lambertian : core::traits::Material {  // :18-41
    Color albedo;
    explicit Lambertian(Color a) : albedo(a) {}
    void lower(SceneBuilder& b, uint32_t slot) const override { b.materials[slot] = rec(RTR_MAT_LAMBERTIAN, albedo, 0); }
};
struct Metal : core::traits::Material {  // :44-70
    Color albedo;
    Float fuzz;
    Metal(Color a, Float f) : albedo(a), fuzz(f) {}
    void lower(SceneBuilder& b, uint32_t slot) const override { b.materials[slot] = rec(RTR_MAT_METAL, albedo, fuzz); }
};
struct Emission : core::traits::Material {  // :73-90; new() pre-multiplies the strength (:79-83)
    Color color;
    Emission(Color c, Float strength) : color{c.r * strength, c.g * strength, c.b * strength} {}
    void lower(SceneBuilder& b, uint32_t slot) const override { b.materials[slot] = rec(RTR_MAT_EMISSION, color, 0); }
};
struct Dielectric : core::traits::Material {  // :93-138
    Float ir;
    explicit Dielectric(Float i) : ir(i) {}
    void lower(SceneBuilder& b, uint32_t slot) const override { b.materials[slot] = rec(RTR_MAT_DIELECTRIC, Color{0, 0, 0}, ir); }
};
struct MixMaterial : core::traits::Material {  // :141-156, arbitrarily nested
    std::unique_ptr<core::traits::Material> first, second;
    Float factor;
    MixMaterial(std::unique_ptr<core::traits::Material> a, std::unique_ptr<core::traits::Material> c, Float f)
        : first(std::move(a)), second(std::move(c)), factor(f) {}
    void lower(SceneBuilder& b, uint32_t slot) const override {
        rtr_material m = rec(RTR_MAT_MIX, Color{0, 0, 0}, factor);
        m.first = (uint32_t)b.materials.size();
        b.materials.push_back(rtr_material{});
        first->lower(b, m.first);
        m.second = (uint32_t)b.materials.size();
        b.materials.push_back(rtr_material{});
        second->lower(b, m.second);
        b.materials[slot] = m;
    }
};
}  // namespace materials

// ---------------------------------------------------------------------------------- objects
namespace objects {
struct Sphere : core::traits::Object {  // lib/objects.rs:15-20 (derive_new: center, radius, material)
    utils::Vec3 center;
    utils::Float radius;
    core::scene::MaterialHandle material;
    Sphere(utils::Vec3 c, utils::Float r, core::scene::MaterialHandle m) : center(c), radius(r), material(m) {}
    rtr_object_ref lower(core::SceneBuilder& b) const override {
        b.spheres.push_back(rtr_sphere{utils::c(center), radius, b.material(material)});
        return rtr_object_ref{RTR_OBJ_SPHERE, (uint32_t)b.spheres.size() - 1u};
    }
};
}  // namespace objects

namespace core {
namespace mesh {
class Mesh : public traits::Object {  // lib/core/mesh.rs:31-38
public:
    // Mesh::from_file (mesh.rs:63-76): obj-rs load of a triangulated OBJ; a missing file panics (:65)
    static std::unique_ptr<Mesh> from_file(const std::string& path, scene::MaterialHandle material) {
        rtr_host_mesh* hm = nullptr;
        check(rtr_mesh_load_obj(path.c_str(), (uint32_t)material.index, &hm));
        const rtr_mesh* m = rtr_host_mesh_get(hm);
        auto out = from_buffers(std::vector<float>(m->positions, m->positions + (size_t)m->n_vertices * 3),
                                std::vector<float>(m->normals, m->normals + (size_t)m->n_vertices * 3),
                                std::vector<uint32_t>(m->indices, m->indices + m->n_indices), material);
        rtr_host_mesh_destroy(hm);
        return out;
    }
    // Mesh::from_buffers (mesh.rs:78-117): xyz triples, one normal per vertex, 3 indices per triangle
    static std::unique_ptr<Mesh> from_buffers(std::vector<float> positions, std::vector<float> normals,
                                              std::vector<uint32_t> indices, scene::MaterialHandle material) {
        if (positions.size() != normals.size() || positions.size() % 3 != 0) throw std::invalid_argument("positions/normals mismatch");
        for (uint32_t i : indices)
            if ((size_t)i >= positions.size() / 3) throw std::out_of_range("vertex index out of range");
        std::unique_ptr<Mesh> m(new Mesh());
        m->positions_ = std::move(positions);
        m->normals_ = std::move(normals);
        m->indices_ = std::move(indices);
        m->material_ = material;
        return m;
    }
    // mesh.rs:119-121: without it bvh_root stays BvhNode::None and the mesh is never hit
    void build_bvh() { bvh_built_ = true; }
    size_t n_triangles() const { return indices_.size() / 3; }
    rtr_object_ref lower(SceneBuilder& b) const override {
        rtr_mesh m{};
        m.positions = positions_.data();
        m.normals = normals_.data();
        m.indices = indices_.data();
        m.n_vertices = (uint32_t)(positions_.size() / 3);
        m.n_indices = bvh_built_ ? (uint32_t)(indices_.size() / 3 * 3) : 0u;
        m.material = b.material(material_);
        b.meshes.push_back(m);
        return rtr_object_ref{RTR_OBJ_MESH, (uint32_t)b.meshes.size() - 1u};
    }

private:
    Mesh() = default;
    std::vector<float> positions_, normals_;
    std::vector<uint32_t> indices_;
    scene::MaterialHandle material_{0};
    bool bvh_built_ = false;
};
}  // namespace mesh
}  // namespace core

namespace volume {
struct Isotropic : core::traits::Material {  // lib/volume.rs:75-87
    utils::Color color;
    explicit Isotropic(utils::Color c) : color(c) {}
    void lower(core::SceneBuilder& b, uint32_t slot) const override { b.materials[slot] = materials::rec(RTR_MAT_ISOTROPIC, color, 0); }
};
struct Volume : core::traits::Object {  // lib/volume.rs:19-31: stores -1/density
    std::unique_ptr<core::traits::Object> boundary;
    utils::Float neg_inv_density;
    Volume(std::unique_ptr<core::traits::Object> b, utils::Float density) : boundary(std::move(b)), neg_inv_density(-1.0f / density) {}
    bool is_volume() const override { return true; }
    rtr_object_ref lower(core::SceneBuilder& b) const override {
        const rtr_object_ref inner = boundary->lower(b);  // Box<dyn Object>: a Sphere, a Mesh or again a Volume (volume.rs:20-23)
        b.volumes.push_back(rtr_volume{inner, neg_inv_density});
        return rtr_object_ref{RTR_OBJ_VOLUME, (uint32_t)b.volumes.size() - 1u};
    }
};
}  // namespace volume

// ---------------------------------------------------------------------------------- backgrounds (lib/backgrounds.rs)
namespace backgrounds {
struct UniformBackground : core::traits::Background {  // :14-30
    utils::Color color;
    explicit UniformBackground(utils::Color c) : color(c) {}
    void lower(core::SceneBuilder& b) const override { b.background = rtr_background{RTR_BG_UNIFORM, utils::c(color), rtr_vec3{0, 0, 0}, nullptr, 0, 0}; }
};
struct GradientBackground : core::traits::Background {  // :33-48
    utils::Color top, bottom;
    GradientBackground(utils::Color t, utils::Color b) : top(t), bottom(b) {}
    void lower(core::SceneBuilder& b) const override { b.background = rtr_background{RTR_BG_GRADIENT, utils::c(top), utils::c(bottom), nullptr, 0, 0}; }
};
struct SkyMap : core::traits::Background {  // :51-92
    std::vector<float> image;  // RGB, image[x + y*width]
    uint32_t width = 0, height = 0;
    // SkyMap::new(path) (:58-80): first RGBA layer of an EXR file; "could not read image!" panics (:72)
    explicit SkyMap(const std::string& path) {
        rtr_host_image* img = nullptr;
        const rtr_status s = rtr_image_load(path.c_str(), &img);
        if (s != RTR_OK) throw Panic(s, std::string("could not read image! ") + rtr_last_error(nullptr));
        width = rtr_host_image_width(img);
        height = rtr_host_image_height(img);
        const float* p = rtr_host_image_rgb(img);
        image.assign(p, p + (size_t)width * height * 3);
        rtr_host_image_destroy(img);
    }
    SkyMap(std::vector<float> rgb, uint32_t w, uint32_t h) : image(std::move(rgb)), width(w), height(h) {
        if (image.size() != (size_t)w * h * 3) throw std::invalid_argument("SkyMap: texel count mismatch");
    }
    void lower(core::SceneBuilder& b) const override {
        b.background = rtr_background{RTR_BG_SKYMAP, rtr_vec3{0, 0, 0}, rtr_vec3{0, 0, 0}, image.data(), width, height};
    }
};
}  // namespace backgrounds

// ---------------------------------------------------------------------------------- Scene (lib/core/scene.rs)
namespace core {
namespace scene {
enum class RenderMode : uint32_t { Full = RTR_MODE_FULL, Clay = RTR_MODE_CLAY, Random = RTR_MODE_RANDOM, Normal = RTR_MODE_NORMAL };  // :14-19

struct RenderSettings {  // :22-29, defaults :52-59
    uint32_t samples_per_pixel = 5;
    uint8_t max_ray_depth = 6;
    float clamp_indirect = 10.0f;
    bool enable_multithreading = true;
    bool enable_bvh_tree = true;
    RenderMode mode = RenderMode::Full;
};

// the lowered scene: owns the arrays an rtr_scene_desc points into (borrowed from the Scene's meshes / sky map)
struct Lowered {
    SceneBuilder b;
    rtr_scene_desc desc{};
};

class Scene {
public:
    camera::Camera camera;                            // pub fields (:33-40)
    RenderSettings settings;
    std::unique_ptr<traits::Background> background;
    std::vector<std::unique_ptr<traits::Object>> objects;

    // Scene::default (:42-65)
    Scene() : background(new backgrounds::UniformBackground(utils::Color{0.8f, 0.8f, 0.8f})) {}
    // Scene::new (:78-86)
    Scene(RenderSettings s, camera::Camera c, std::unique_ptr<traits::Background> bg) : camera(c), settings(s), background(std::move(bg)) {}
    Scene(Scene&&) = default;
    Scene& operator=(Scene&&) = default;

    ObjectHandle add_object(std::unique_ptr<traits::Object> o) { objects.push_back(std::move(o)); return ObjectHandle{objects.size() - 1}; }   // :88-91
    const traits::Object& object(ObjectHandle h) const { return *objects.at(h.index); }                                                        // :93-98
    MaterialHandle add_material(std::unique_ptr<traits::Material> m) { materials_.push_back(std::move(m)); return MaterialHandle{materials_.size() - 1}; }  // :100-103
    const traits::Material& material(MaterialHandle h) const { return *materials_.at(h.index); }                                               // :105-110
    // Scene::build_bvh (:112-114): the tree is built by rtr_scene_upload (host or device), nothing to do here
    void build_bvh() const {}

    // Scene::from_file (:74-76): the serde/typetag JSON the crate writes.  unwrap() panics -> throws.
    static Scene from_file(const std::string& path);

    Lowered lower() const {
        Lowered out;
        SceneBuilder& b = out.b;
        b.n_handles = materials_.size();
        b.materials.resize(materials_.size());  // handles keep their indices, Mix children go behind them
        for (size_t i = 0; i < materials_.size(); ++i) materials_[i]->lower(b, (uint32_t)i);
        for (const auto& o : objects) b.objects.push_back(o->lower(b));
        background->lower(b);
        rtr_scene_desc& d = out.desc;
        d.objects = b.objects.data(); d.n_objects = (uint32_t)b.objects.size();
        d.spheres = b.spheres.data(); d.n_spheres = (uint32_t)b.spheres.size();
        d.meshes = b.meshes.data(); d.n_meshes = (uint32_t)b.meshes.size();
        d.volumes = b.volumes.data(); d.n_volumes = (uint32_t)b.volumes.size();
        d.materials = b.materials.data(); d.n_materials = (uint32_t)b.materials.size();
        d.background = b.background;
        return out;
    }
    rtr_camera c_camera() const { return rtr_camera{utils::c(camera.lookfrom), utils::c(camera.lookat), utils::c(camera.vertical), camera.vertical_fov, camera.aspect_ratio}; }
    rtr_settings c_settings() const {
        return rtr_settings{settings.samples_per_pixel, settings.max_ray_depth, settings.clamp_indirect,
                            settings.enable_multithreading ? 1u : 0u, settings.enable_bvh_tree ? 1u : 0u, (uint32_t)settings.mode};
    }

private:
    std::vector<std::unique_ptr<traits::Material>> materials_;  // private in the crate too (:39)
};

namespace detail {
inline std::unique_ptr<traits::Material> material_from(const rtr_scene_desc& d, uint32_t i) {
    if (i >= d.n_materials) throw std::out_of_range("material index out of range");
    const rtr_material& m = d.materials[i];
    const utils::Color c{m.color.x, m.color.y, m.color.z};
    switch (m.kind) {
        case RTR_MAT_LAMBERTIAN: return std::unique_ptr<traits::Material>(new materials::Lambertian(c));
        case RTR_MAT_METAL: return std::unique_ptr<traits::Material>(new materials::Metal(c, m.param));
        case RTR_MAT_EMISSION: return std::unique_ptr<traits::Material>(new materials::Emission(c, 1.0f));  // JSON stores the product (Q13)
        case RTR_MAT_DIELECTRIC: return std::unique_ptr<traits::Material>(new materials::Dielectric(m.param));
        case RTR_MAT_ISOTROPIC: return std::unique_ptr<traits::Material>(new volume::Isotropic(c));
        case RTR_MAT_MIX: return std::unique_ptr<traits::Material>(new materials::MixMaterial(material_from(d, m.first), material_from(d, m.second), m.param));
        default: throw std::invalid_argument("unknown material kind");
    }
}
inline std::unique_ptr<traits::Object> object_from(const rtr_scene_desc& d, rtr_object_ref r) {
    if (r.kind == RTR_OBJ_SPHERE) {
        const rtr_sphere& s = d.spheres[r.index];
        return std::unique_ptr<traits::Object>(new objects::Sphere(utils::Vec3{s.center.x, s.center.y, s.center.z}, s.radius, MaterialHandle{s.material}));
    }
    if (r.kind == RTR_OBJ_MESH) {
        const rtr_mesh& m = d.meshes[r.index];
        auto mesh = mesh::Mesh::from_buffers(std::vector<float>(m.positions, m.positions + (size_t)m.n_vertices * 3),
                                             std::vector<float>(m.normals, m.normals + (size_t)m.n_vertices * 3),
                                             std::vector<uint32_t>(m.indices, m.indices + m.n_indices), MaterialHandle{m.material});
        if (m.n_indices) mesh->build_bvh();
        return mesh;
    }
    const rtr_volume& v = d.volumes[r.index];
    std::unique_ptr<volume::Volume> vol(new volume::Volume(object_from(d, v.boundary), 1.0f));
    vol->neg_inv_density = v.neg_inv_density;
    return vol;
}
}  // namespace detail

inline Scene Scene::from_file(const std::string& path) {
    rtr_host_scene* hs = nullptr;
    check(rtr_scene_load_json(path.c_str(), &hs));
    struct Guard { rtr_host_scene* h; ~Guard() { rtr_host_scene_destroy(h); } } guard{hs};
    const rtr_scene_desc& d = *rtr_host_scene_desc(hs);
    const rtr_camera& c = *rtr_host_scene_camera(hs);
    const rtr_settings& s = *rtr_host_scene_settings(hs);
    Scene out;
    out.camera = camera::Camera{{c.lookfrom.x, c.lookfrom.y, c.lookfrom.z}, {c.lookat.x, c.lookat.y, c.lookat.z},
                                {c.vertical.x, c.vertical.y, c.vertical.z}, c.vertical_fov, c.aspect_ratio};
    out.settings = RenderSettings{s.samples_per_pixel, (uint8_t)s.max_ray_depth, s.clamp_indirect, s.enable_multithreading != 0,
                                  s.enable_bvh_tree != 0, (RenderMode)s.mode};
    // Scene.materials are the leading entries; children of a MixMaterial were appended behind them
    uint32_t n_handles = d.n_materials;
    for (uint32_t i = 0; i < d.n_materials; ++i)
        if (d.materials[i].kind == RTR_MAT_MIX) n_handles = std::min(n_handles, std::min(d.materials[i].first, d.materials[i].second));
    for (uint32_t i = 0; i < n_handles; ++i) out.add_material(detail::material_from(d, i));
    for (uint32_t i = 0; i < d.n_objects; ++i) out.add_object(detail::object_from(d, d.objects[i]));
    const rtr_background& bg = d.background;
    if (bg.kind == RTR_BG_UNIFORM) out.background.reset(new backgrounds::UniformBackground(utils::Color{bg.color.x, bg.color.y, bg.color.z}));
    else if (bg.kind == RTR_BG_GRADIENT)
        out.background.reset(new backgrounds::GradientBackground(utils::Color{bg.color.x, bg.color.y, bg.color.z}, utils::Color{bg.bottom.x, bg.bottom.y, bg.bottom.z}));
    else out.background.reset(new backgrounds::SkyMap(std::vector<float>(bg.texels, bg.texels + (size_t)bg.width * bg.height * 3), bg.width, bg.height));
    return out;
}
}  // namespace scene

// ---------------------------------------------------------------------------------- render (lib/core/render.rs)
namespace render {
struct RenderTarget {  // :14-30
    size_t width, height;
    std::vector<uint8_t> data;  // RGBA8, row 0 = top
    bool request_redraw = false;
    static RenderTarget new_(size_t w, size_t h) { return RenderTarget(w, h); }
    RenderTarget(size_t w, size_t h) : width(w), height(h), data(w * h * 4, 0) {}
    void save_png(const std::string& path) const { check(rtr_write_png(path.c_str(), data.data(), (uint32_t)width, (uint32_t)height)); }
};

// One GPU: context + the uploaded scene, for callers that render the same scene repeatedly (the crate rebuilds its
// BVH on every render(), render.rs:67; here the upload is the build and render() below does both).
class Device {
public:
    explicit Device(int device = 0) { check(rtr_context_create(device, nullptr, &ctx_)); }
    ~Device() { if (scene_) rtr_scene_destroy(scene_); if (ctx_) rtr_context_destroy(ctx_); }
    Device(const Device&) = delete;
    Device& operator=(const Device&) = delete;
    void upload(const scene::Scene& s) {
        if (scene_) { rtr_scene_destroy(scene_); scene_ = nullptr; }
        const scene::Lowered l = s.lower();
        check(rtr_scene_upload(ctx_, &l.desc, &scene_), ctx_);
    }
    // fp32 accumulation buffer (W*H*4: sum r,g,b + sample count) is optional
    rtr_stats render(RenderTarget& target, const scene::Scene& s, uint64_t seed = 0, std::vector<float>* accum = nullptr,
                     uint32_t kernel = RTR_KERNEL_AUTO, uint32_t flags = 0) {
        if (!scene_) upload(s);
        rtr_render_params p{};
        p.width = (uint32_t)target.width; p.height = (uint32_t)target.height;
        p.camera = s.c_camera(); p.settings = s.c_settings();
        p.seed = seed; p.kernel = kernel; p.flags = flags;
        if (accum) accum->assign(target.width * target.height * 4, 0.0f);
        rtr_stats st{};
        check(rtr_render(ctx_, scene_, &p, target.data.data(), accum ? accum->data() : nullptr, &st), ctx_);
        target.request_redraw = true;  // render.rs:62
        return st;
    }
    rtr_context* context() const { return ctx_; }

private:
    rtr_context* ctx_ = nullptr;
    rtr_scene* scene_ = nullptr;
};

// pub fn render(target: &mut RenderTarget, scene: &Scene) (render.rs:61): builds the BVH, renders, fills target.data
inline rtr_stats render(RenderTarget& target, const scene::Scene& s, uint64_t seed = 0) {
    Device dev(0);
    return dev.render(target, s, seed);
}
}  // namespace render
}  // namespace core
}  // namespace ray_tracing_rust
#endif  // RTR_HPP_
