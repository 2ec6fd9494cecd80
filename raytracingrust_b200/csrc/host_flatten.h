// host_flatten.h — host image of the flattened scene (see flat.h, host_flatten.cpp)
#ifndef RTR_HOST_FLATTEN_H_
#define RTR_HOST_FLATTEN_H_

#include <string>
#include <vector>

#include "../../include/rtr.h"
#include "flat.h"

namespace rtr {

struct HostFlat {
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
int flatten_scene(const rtr_scene_desc& desc, HostFlat& out, std::string& err, uint32_t slack = 8);

}  // namespace rtr
#endif
