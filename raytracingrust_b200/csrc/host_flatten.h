// host_flatten.h — host image of the flattened scene (see flat.h, host_flatten.cpp)
#ifndef RTR_HOST_FLATTEN_H_
#define RTR_HOST_FLATTEN_H_

#include <string>
#include <vector>

#include "../../include/rtr.h"
#include "flat.h"

namespace rtr {

// A mesh whose BVH is built on the device (bvh_build.cu) instead of by the host flattener.  The host only reserves its
// index ranges — device-built records occupy the FRONT of the device arrays, the host-built ones follow — and refers
// to it through its root Split, node_base (4-wide tree: node4_base).
struct DeviceMesh {
    uint32_t mesh_index, object, material, n_tris;
    uint32_t slot_base;    // first triangle slot (tris / tri_vn)
    uint32_t node_base;    // first 2-wide node; the mesh owns n_tris of them
    uint32_t node4_base;   // first 4-wide node; the mesh owns count_nodes4(n_tris)
    uint32_t rank_base;    // DFS leaf rank of its first triangle (bvh.rs:183-187 tie rule)
    // hybrid: the device builds the reference tree (gates, ranks, records), the host then builds the SAH traversal tree over
    // the gates it reads back (build_traversal_over_gates): n_gated - 1 nodes from sah_base
    uint32_t hybrid, sah_base, n_gated, slack, depth4_estimate;
};
uint32_t count_nodes4(uint32_t n_leaves);  // 4-wide nodes the device collapse emits for a median-split tree of n leaves
uint32_t count_gated(uint32_t n_leaves);   // gated entries of that tree: single leaves + 2-leaf Splits (paired leaves)
// SAH traversal tree over gated entries read back from the device (2 x F4 per leaf position: {gate.min, bits(ref)}
// {gate.max, 0}; ref EMPTY = the second leaf of a pair).  Node indices start at node_base; n_gated - 1 nodes come out.
void build_traversal_over_gates(const F4* gated, uint32_t n_positions, uint32_t node_base, uint32_t slack, std::vector<F4>& nodes,
                                uint32_t* depth);
struct HostFlat;
// Hybrid scenes: the 4-wide fold of the WHOLE traversal tree (scene level + every mesh's SAH tree, mesh_sah[k] for
// device_meshes[k]) once those exist, so that a mesh's top folds into the scene's nodes as in a host-built scene.
// Returns the node array (indices from 0) and the entry refs (root, VolumeRec::entry4, ObjectRec::entry4 in the copies).
void collapse_scene4(const HostFlat& flat, const std::vector<const std::vector<F4>*>& mesh_sah, std::vector<F4>& nodes4,
                     uint32_t* root_entry4, std::vector<VolumeRec>& volumes, std::vector<ObjectRec>& objects, uint32_t* max_depth4);

struct HostFlat {
    std::vector<DeviceMesh> device_meshes;
    uint32_t device_nodes = 0, device_nodes4 = 0, device_tri_slots = 0;  // sizes of the device-built prefixes; host indices start there
    std::vector<F4> nodes, spheres, tris, tri_vn, env;
    std::vector<F4> nodes4;                       // 4-wide traversal tree (flat.h), 8 per node; empty when not built
    uint32_t root_entry4 = 0xFFFFFFFFu;
    uint32_t max_depth4 = 0;                      // deepest root-to-leaf path of any 4-wide tree (stack need: 3 per level)
    std::vector<PrimInfo> sphere_info, tri_info;
    std::vector<VolumeRec> volumes;
    std::vector<ObjectRec> objects;
    std::vector<MaterialRec> materials;
    BackgroundRec background;
    uint32_t root_entry = 0xFFFFFFFFu;            // traversal tree
    uint32_t root_entry_reference = 0xFFFFFFFFu;  // reference topology
    uint32_t max_depth = 0;           // levels of the fused scene tree incl. the pseudo parent
    uint32_t max_boundary_depth = 0;  // deepest volume-boundary tree
};

// returns an rtr_status; on failure `err` says why.  `slack` = levels a collection's SAH traversal tree may exceed its
// balanced depth ceil(log2 n) by (0: every split is the median, the minimal depth)
// `device_mesh_min_tris`: meshes with at least this many triangles are left to the device builder (0 = none);
// `hybrid`: ... and get a host-built SAH traversal tree over the device-built gates at upload time
int flatten_scene(const rtr_scene_desc& desc, HostFlat& out, std::string& err, uint32_t slack = 8, uint32_t device_mesh_min_tris = 0,
                  bool hybrid = false);

}  // namespace rtr
#endif
