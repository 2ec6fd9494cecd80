// Host-only check of the 4-wide traversal tree (flat.h): below every entry point it must reach exactly the leaf slots the
// 2-wide traversal tree reaches, each carrying the same box, and every interior slot's box must be one the 2-wide tree holds.
// Built and run by tests/test_nodes4.py (g++ on this file + csrc/host_flatten.cpp; no GPU).
#include <cmath>
#include <cstdio>
#include <cstdlib>
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
// brute-force counterparts of count_nodes4 / count_gated (host_flatten.h): walk the median-split shape node by node
static uint32_t brute_nodes4(uint32_t n, uint32_t depth) {
    if (n < 2) return 0;
    return (((depth & 1u) == 0u && (n >= 3 || depth == 0)) ? 1u : 0u) + brute_nodes4(n / 2, depth + 1) + brute_nodes4(n - n / 2, depth + 1);
}
static uint32_t brute_gated(uint32_t n) { return n <= 2 ? (n ? 1u : 0u) : brute_gated(n / 2) + brute_gated(n - n / 2); }
static int check_counts() {
    int bad = 0;
    for (uint32_t n = 1; n <= 70000; n = n < 3000 ? n + 1 : n + n / 7 + 1) {
        if (n >= 2 && count_nodes4(n) != brute_nodes4(n, 0)) { std::printf("FAIL count_nodes4(%u) = %u, expected %u\n", n, count_nodes4(n), brute_nodes4(n, 0)); ++bad; }
        if (count_gated(n) != brute_gated(n)) { std::printf("FAIL count_gated(%u) = %u, expected %u\n", n, count_gated(n), brute_gated(n)); ++bad; }
    }
    return bad;
}

// the hybrid build's host half: a SAH tree over gated entries reaches every entry exactly once
static int check_over_gates() {
    std::mt19937 g(11); std::uniform_real_distribution<float> U(-5.f, 5.f), R(0.01f, 0.5f);
    const uint32_t n = 5003, base = 777;
    std::vector<F4> gated(2 * (size_t)n);
    std::multiset<uint32_t> want;
    for (uint32_t p = 0; p < n; ++p) {
        const bool second_of_pair = p > 0 && (p % 5u) == 1u;  // position p-1 holds a paired entry
        uint32_t ref = second_of_pair ? REF_EMPTY_VALUE : RTR_MAKE_REF((p % 5u) == 0u && p + 1 < n ? REF_TRI2 : REF_TRI, p);
        if ((p % 5u) == 0u && p + 1 >= n) ref = RTR_MAKE_REF(REF_TRI, p);
        const float c[3] = {U(g), U(g), U(g)}, r = R(g);
        float w; std::memcpy(&w, &ref, 4);
        gated[2 * (size_t)p] = F4{c[0] - r, c[1] - r, c[2] - r, w};
        gated[2 * (size_t)p + 1] = F4{c[0] + r, c[1] + r, c[2] + r, 0.0f};
        if (ref != REF_EMPTY_VALUE) want.insert(ref);
    }
    std::vector<F4> nodes;
    uint32_t depth = 0;
    build_traversal_over_gates(gated.data(), n, base, 8, nodes, &depth);
    int bad = 0;
    if (nodes.size() / 4 != want.size() - 1) { std::printf("FAIL over gates: %zu nodes for %zu entries\n", nodes.size() / 4, want.size()); ++bad; }
    std::multiset<uint32_t> got2;
    std::vector<uint32_t> stack{RTR_MAKE_REF(REF_NODE, base)};
    while (!stack.empty()) {
        const uint32_t ref = stack.back(); stack.pop_back();
        if (RTR_REF_KIND(ref) != REF_NODE) { got2.insert(ref); continue; }
        const F4* q = &nodes[(size_t)(RTR_REF_INDEX(ref) - base) * 4];
        for (int k = 0; k < 2; ++k) { const uint32_t c = bitsof(q[k].w); if (c != REF_EMPTY_VALUE) stack.push_back(c); }
    }
    if (got2 != want) { std::printf("FAIL over gates: the 2-wide tree does not reach every entry once\n"); ++bad; }
    if (depth > 13 + 8) { std::printf("FAIL over gates: depth %u\n", depth); ++bad; }
    return bad;
}

// the host build's helper threads (RTR_HOST_THREADS) must not change a byte of the flattened scene
static int check_threads() {
    std::mt19937 g(3); std::uniform_real_distribution<float> U(-10.f, 10.f), R(0.05f, 0.3f);
    const int ntri = 70001;
    std::vector<float> pos, nrm; std::vector<uint32_t> idx;
    for (int t = 0; t < ntri; ++t) { float c[3] = {U(g), U(g), std::round(U(g))}; for (int k = 0; k < 3; ++k) { for (int a = 0; a < 3; ++a) { pos.push_back(c[a] + (a == 2 ? 0.f : R(g))); nrm.push_back(a == 1); } idx.push_back(t * 3 + k); } }
    rtr_mesh m{pos.data(), nrm.data(), idx.data(), (uint32_t)(pos.size() / 3), (uint32_t)idx.size(), 0, 0};
    rtr_object_ref o{RTR_OBJ_MESH, 0};
    rtr_material mat; std::memset(&mat, 0, sizeof(mat));
    rtr_scene_desc d; std::memset(&d, 0, sizeof(d));
    d.objects = &o; d.n_objects = 1; d.meshes = &m; d.n_meshes = 1; d.materials = &mat; d.n_materials = 1;
    HostFlat f[2];
    const char* threads[2] = {"1", "8"};
    for (int k = 0; k < 2; ++k) {
        setenv("RTR_HOST_THREADS", threads[k], 1);
        std::string err;
        if (flatten_scene(d, f[k], err, 8) != 0) { std::printf("FAIL threads: %s\n", err.c_str()); return 1; }
    }
    unsetenv("RTR_HOST_THREADS");
    auto same = [](const std::vector<F4>& a, const std::vector<F4>& b) { return a.size() == b.size() && std::memcmp(a.data(), b.data(), a.size() * sizeof(F4)) == 0; };
    if (!same(f[0].nodes, f[1].nodes) || !same(f[0].nodes4, f[1].nodes4) || !same(f[0].tris, f[1].tris) || !same(f[0].tri_vn, f[1].tri_vn) ||
        f[0].root_entry != f[1].root_entry || f[0].root_entry4 != f[1].root_entry4 || f[0].max_depth != f[1].max_depth) {
        std::printf("FAIL threads: the flattened scene depends on the thread count\n");
        return 1;
    }
    return 0;
}

int main() {
    if (check_counts() || check_over_gates() || check_threads()) { std::printf("FAIL\n"); return 1; }
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
