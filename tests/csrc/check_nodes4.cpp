// Host-only check of the 4-wide traversal tree (flat.h): below every entry point it must reach exactly the leaf slots the
// 2-wide traversal tree reaches, each carrying the same box, and every interior slot's box must be one the 2-wide tree holds.
// Built and run by tests/test_nodes4.py (g++ on this file + csrc/host_flatten.cpp; no GPU).
#include <cstdio>
#include <cstring>
#include <map>
#include <random>
#include <set>
#include <string>
#include <vector>
#include "host_flatten.h"
using namespace rtr;
struct B6 { float v[6]; bool operator<(const B6& o) const { return std::memcmp(v, o.v, 24) < 0; } bool operator==(const B6& o) const { return std::memcmp(v, o.v, 24) == 0; } };
static uint32_t bitsof(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static void leaves2(const HostFlat& f, uint32_t ref, std::multimap<uint32_t, B6>& out, std::set<B6>& interior, const B6* box) {
    if (ref == REF_EMPTY_VALUE) return;
    if (RTR_REF_KIND(ref) != REF_NODE) { out.insert({ref, box ? *box : B6{}}); return; }
    if (box) interior.insert(*box);
    const F4* q = &f.nodes[(size_t)RTR_REF_INDEX(ref) * 4];
    B6 l{{q[0].x, q[0].y, q[0].z, q[1].x, q[1].y, q[1].z}}, r{{q[2].x, q[2].y, q[2].z, q[3].x, q[3].y, q[3].z}};
    leaves2(f, bitsof(q[0].w), out, interior, &l); leaves2(f, bitsof(q[1].w), out, interior, &r);
}
static int depth4 = 0;
static void leaves4(const HostFlat& f, uint32_t ref, std::multimap<uint32_t, B6>& out, std::set<B6>& interior, const B6* box, int d) {
    if (ref == REF_EMPTY_VALUE) return;
    if (d > depth4) depth4 = d;
    if (RTR_REF_KIND(ref) != REF_NODE) { out.insert({ref, box ? *box : B6{}}); return; }
    if (box) interior.insert(*box);
    const F4* q = &f.nodes4[(size_t)RTR_REF_INDEX(ref) * 8];
    bool seen_empty = false;
    for (int i = 0; i < 4; ++i) {
        const uint32_t c = bitsof(q[i].w);
        if (c == REF_EMPTY_VALUE) { seen_empty = true; continue; }
        if (seen_empty) { std::printf("FAIL: hole in a 4-wide node\n"); std::exit(1); }
        B6 b{{q[2 * i].x, q[2 * i].y, q[2 * i].z, q[2 * i + 1].x, q[2 * i + 1].y, q[2 * i + 1].z}};
        leaves4(f, c, out, interior, &b, d + 1);
    }
}
static int compare(const HostFlat& f, uint32_t e2, uint32_t e4, const char* what) {
    std::multimap<uint32_t, B6> a, b; std::set<B6> ia, ib;
    leaves2(f, e2, a, ia, nullptr); leaves4(f, e4, b, ib, nullptr, 0);
    if (a != b) { std::printf("FAIL %s: leaf sets differ (%zu vs %zu)\n", what, a.size(), b.size()); return 1; }
    for (const B6& x : ib) if (!ia.count(x)) { std::printf("FAIL %s: interior box not in the 2-wide tree\n", what); return 1; }
    return 0;
}
int main() {
    std::mt19937 g(7); std::uniform_real_distribution<float> U(-10.f, 10.f), R(0.05f, 0.6f);
    std::vector<rtr_sphere> sph; std::vector<rtr_object_ref> obj; std::vector<rtr_mesh> meshes; std::vector<rtr_volume> vols;
    for (int i = 0; i < 1500; ++i) { sph.push_back({{U(g), U(g), U(g)}, R(g), 0}); obj.push_back({RTR_OBJ_SPHERE, (uint32_t)i}); }
    std::vector<float> pos, nrm; std::vector<uint32_t> idx;
    for (int t = 0; t < 4001; ++t) { float c[3] = {U(g), U(g), U(g)}; for (int k = 0; k < 3; ++k) { for (int a = 0; a < 3; ++a) { pos.push_back(c[a] + R(g)); nrm.push_back(a == 1); } idx.push_back(t * 3 + k); } }
    meshes.push_back({pos.data(), nrm.data(), idx.data(), (uint32_t)(pos.size() / 3), (uint32_t)idx.size(), 0, 0});
    std::vector<uint32_t> idx2(idx.begin(), idx.begin() + 3 * 77);
    meshes.push_back({pos.data(), nrm.data(), idx2.data(), (uint32_t)(pos.size() / 3), (uint32_t)idx2.size(), 0, 0});
    obj.push_back({RTR_OBJ_MESH, 0});
    sph.push_back({{0, 0, 0}, 3.f, 1});
    vols.push_back({{RTR_OBJ_SPHERE, (uint32_t)sph.size() - 1}, -2.f}); vols.push_back({{RTR_OBJ_MESH, 1}, -1.f});
    obj.push_back({RTR_OBJ_VOLUME, 0}); obj.push_back({RTR_OBJ_VOLUME, 1});
    rtr_material mats[2]; std::memset(mats, 0, sizeof(mats)); mats[1].kind = RTR_MAT_ISOTROPIC;
    rtr_scene_desc d; std::memset(&d, 0, sizeof(d));
    d.objects = obj.data(); d.n_objects = (uint32_t)obj.size(); d.spheres = sph.data(); d.n_spheres = (uint32_t)sph.size();
    d.meshes = meshes.data(); d.n_meshes = 2; d.volumes = vols.data(); d.n_volumes = 2; d.materials = mats; d.n_materials = 2;
    HostFlat f; std::string err;
    if (flatten_scene(d, f, err, 8) != 0) { std::printf("flatten failed: %s\n", err.c_str()); return 1; }
    int bad = compare(f, f.root_entry, f.root_entry4, "scene");
    const int root_depth = depth4;
    for (size_t i = 0; i < f.volumes.size(); ++i) bad += compare(f, f.volumes[i].entry, f.volumes[i].entry4, "volume boundary");
    for (size_t i = 0; i < f.objects.size(); ++i) bad += compare(f, f.objects[i].entry, f.objects[i].entry4, "object");
    std::printf("%s: %zu 2-wide nodes (depth %u), %zu 4-wide nodes (depth %u reported, %d walked from the root)\n", bad ? "FAIL" : "OK",
                f.nodes.size() / 4, f.max_depth, f.nodes4.size() / 8, f.max_depth4, root_depth);
    return bad != 0;
}
