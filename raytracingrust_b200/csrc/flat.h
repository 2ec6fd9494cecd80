// flat.h — the flattened, device-resident scene layout (host + device view).
//
// Produced once per rtr_scene_upload by host_flatten.cpp from the reference's pointer BVH
// topology (lib/core/bvh.rs:59-144 for the scene and, nested, for every mesh
// lib/core/mesh.rs:119-121).  The two levels are fused into ONE linear node array: the
// top-level leaf of a mesh object is simply that mesh's root.
//
// Node (64 B, 64-B aligned, 4 x float4) — a Split node stores the boxes its CHILDREN would
// test on entry (bvh.rs:161), so one fetch yields both slab tests and the near/far order:
//   q0 = { left.min.xyz , bits(left_ref)  }
//   q1 = { left.max.xyz , bits(right_ref) }
//   q2 = { right.min.xyz, 0 }
//   q3 = { right.max.xyz, 0 }
// Leaves carry no box (bvh.rs:157-159, quirk Q3): a leaf child is tested unconditionally once
// its parent's box was pierced.  A tree root has no parent, so every root that is a Split gets
// a pseudo parent {left = root + its own box, right = EMPTY}; queries start at that "entry".
//
// Child reference (u32): kind in bits 31..29, index in bits 28..0.
#ifndef RTR_FLAT_H_
#define RTR_FLAT_H_

#include <stdint.h>

namespace rtr {

enum : uint32_t {
    REF_NODE = 0u,    // index of a 64-B node
    REF_SPHERE = 1u,  // slot in spheres[] / sphere_info[]
    REF_TRI = 2u,     // slot in tris[] (x4) / tri_vn[] (x3)
    REF_VOLUME = 3u,  // index in volumes[]
    REF_TRI2 = 4u,    // traversal tree only: triangle slots i and i+1, siblings under one reference Split (one gate)
    REF_SPHERE2 = 5u, // traversal tree only: sphere slots i and i+1, likewise
    REF_EMPTY = 7u    // BvhNode::None
};
#define RTR_REF_KIND(r) ((r) >> 29)
#define RTR_REF_INDEX(r) ((r)&0x1FFFFFFFu)
#define RTR_MAKE_REF(kind, idx) (((uint32_t)(kind) << 29) | (uint32_t)(idx))
static const uint32_t REF_EMPTY_VALUE = 0xFFFFFFFFu;  // kind 7

// 4-wide traversal node (128 B, 128-B aligned, 8 x float4): the traversal tree with every other level folded away
// (host_flatten.cpp collapse4 for host-built trees, bvh_build.cu k_emit_nodes4 for device-built ones).  Child k's box
// is the SAME box its slot carries in the 2-wide tree (an interior box is a union of gates, a leaf slot carries its
// gate), so the candidate set — every leaf whose gate is pierced — is unchanged:
//   q0 = { c0.min, bits(ref0) }  q1 = { c0.max, bits(ref1) }  q2 = { c1.min, bits(ref2) }  q3 = { c1.max, bits(ref3) }
//   q4 = { c2.min, 0 }           q5 = { c2.max, 0 }           q6 = { c3.min, 0 }           q7 = { c3.max, 0 }
// REF_NODE references inside nodes4 index nodes4; unused slots hold REF_EMPTY.  Only the 4-wide instantiations of
// wf_extend walk it (SceneView::root_entry4, VolumeRec::entry4, ObjectRec::entry4); every other walk uses the 2-wide arrays.
struct F4 { float x, y, z, w; };
struct U4 { uint32_t x, y, z, w; };

// per-primitive side record, fetched once per ray for the winning hit only
//   x = Scene.objects index, y = triangle index inside its mesh (original order),
//   z = material index, w = DFS leaf rank (tie-break: among equal t the LAST leaf in the
//       reference's left-to-right DFS wins, bvh.rs:183-187, quirk Q2)
typedef U4 PrimInfo;

struct VolumeRec {        // Volume (lib/volume.rs:19-31)
    uint32_t entry;       // ref of the boundary: SPHERE slot | NODE (pseudo parent) | TRI | EMPTY
    float neg_inv_density;
    uint32_t material;    // boundary.material()
    uint32_t object;      // Scene.objects index
    uint32_t rank;        // DFS leaf rank of the volume leaf
    uint32_t index;       // index in rtr_scene_desc.volumes (RNG iteration slot)
    uint32_t entry_reference;  // the boundary in the reference topology (RTR_FLAG_REFERENCE_TREE)
    uint32_t entry4;      // the boundary in the 4-wide traversal tree (REF_NODE -> nodes4), or the same leaf ref
};

struct ObjectRec {        // Scene.objects[i], for the BVH-off linear scan (scene.rs:141-156)
    uint32_t entry;       // SPHERE slot | NODE (pseudo parent) | TRI | VOLUME | EMPTY
    uint32_t entry_reference;  // the same in the reference topology
    uint32_t entry4;      // the same in the 4-wide traversal tree
    uint32_t _pad;
};

struct MaterialRec {      // Box<dyn Material>
    uint32_t kind;
    float r, g, b;
    float param;
    uint32_t first, second;
    uint32_t _pad;
};

struct BackgroundRec {
    uint32_t kind;
    float r, g, b;        // uniform | gradient top
    float br, bg, bb;     // gradient bottom
    uint32_t width, height;
};

// Device view: raw pointers into one allocation each.
struct SceneView {
    const F4* nodes;          // 4 per node
    const F4* nodes4;         // 8 per 4-wide node (null: no 4-wide tree — RTR_WF_WIDE=0, or its stack need exceeds the limit)
    uint32_t root_entry4;     // entry ref of the 4-wide traversal tree
    uint32_t _pad4;
    const F4* spheres;        // (cx, cy, cz, r) per slot
    const PrimInfo* sphere_info;
    const F4* tris;           // 4 per slot (64 B, one aligned record = two 256-bit loads): (v0.xyz, fn.x) (e1.xyz, fn.y)
                              // (e2.xyz, fn.z) bits(PrimInfo)
    const F4* tri_vn;         // 3 per slot: vertex normals n0, n1, n2 (w unused)
    const VolumeRec* volumes;
    const ObjectRec* objects;
    const MaterialRec* materials;
    const F4* env;            // SkyMap texels as float4 (rgb, 0)
    const F4* linear_spheres; // Scene.objects order copy of the spheres when EVERY object is a sphere (else null):
                              // the BVH-off scan (scene.rs:141-156) then reads consecutive records, no dispatch
    BackgroundRec background;
    uint32_t root_entry;      // entry ref of the scene tree being walked (traversal tree, or the reference topology)
    uint32_t reference_tree;  // 1: walk the reference topology (leaf slots carry no box) — counting / validation
    uint32_t n_objects;
    uint32_t n_materials;
    uint32_t n_volumes;
    uint32_t n_nodes, n_sphere_slots, n_tri_slots;  // array extents (debug bounds checks)
    uint32_t linear_needs_inv;  // some object is a mesh or a volume (the linear scan then needs 1/d)
};

}  // namespace rtr
#endif
