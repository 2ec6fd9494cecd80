// bvh_build.h — device BVH build for sphere scenes (bvh_build.cu)
#ifndef RTR_BVH_BUILD_H_
#define RTR_BVH_BUILD_H_

#include <cuda_runtime_api.h>
#include <stdint.h>

#include <vector>

#include "../../include/rtr.h"
#include "flat.h"
#include "host_flatten.h"

namespace rtr {

struct DeviceBvh {
    F4* nodes = nullptr;        // n nodes of 64 B (n-1 Splits in pre-order + the root's pseudo parent)
    F4* spheres = nullptr;      // n slots in DFS leaf order
    PrimInfo* sphere_info = nullptr;
    ObjectRec* objects = nullptr;
    F4* linear_spheres = nullptr;  // the spheres in Scene.objects order (BVH-off scan)
    F4* nodes4 = nullptr;       // the 4-wide traversal tree (flat.h), when asked for: every other level of `nodes` folded away
    uint32_t n_nodes4 = 0, root_entry4 = 0xFFFFFFFFu, max_depth4 = 0;
    uint32_t n_nodes = 0, root_entry = 0xFFFFFFFFu, max_depth = 0;
    uint64_t device_bytes = 0, h2d_bytes = 0;
    bool nan_centroid = false;
};

// true when every object of the scene is a Sphere (the case the device builder covers)
bool device_build_applies(const rtr_scene_desc& d);

// Builds the reference tree (bvh.rs:59-144) for Scene.objects on `st`.  The four result arrays are cudaMalloc'ed
// and appended to `allocs` (owned by the caller); scratch is freed before returning.  Synchronises `st`.
cudaError_t build_sphere_scene_on_device(const rtr_scene_desc& d, cudaStream_t st, cudaMemPool_t pool, DeviceBvh* out, std::vector<void*>& allocs,
                                         bool want_nodes4 = false);

// mesh.build_bvh() (mesh.rs:119-121) of one mesh on `st`: writes the mesh's node, 4-wide node (nodes4 may be null) and
// triangle records into the ranges `dm` reserves in the scene's device arrays.  Scratch and the device copies of the vertex
// buffers come from `pool` and are freed before returning.  Synchronises `st`.
cudaError_t build_mesh_on_device(const rtr_mesh& mesh, const DeviceMesh& dm, cudaStream_t st, cudaMemPool_t pool, F4* nodes, F4* nodes4,
                                 F4* tris, F4* tri_vn, uint64_t* h2d_bytes, bool* nan_centroid, bool* count_mismatch,
                                 std::vector<F4>* gated_out = nullptr);  // hybrid: the gated entries (host_flatten.h), after the sync

}  // namespace rtr
#endif
