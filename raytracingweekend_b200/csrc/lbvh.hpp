// lbvh.hpp — device-side linear BVH build (lbvh.cu), opt-in alternative to the host SAH build of api.cu.
#pragma once
#include <vector>

#include <cuda_runtime.h>

namespace rtw {

// boxes: n x (lo.xyz, hi.xyz).  Output: n - 1 internal nodes, node 0 = root; left/right >= 0: internal node, < 0: ~(object index);
// node_box: (n - 1) x 6.  n < 2: empty output.  The current device is used.
cudaError_t lbvh_build(const std::vector<float>& boxes, std::vector<int>& left, std::vector<int>& right, std::vector<float>& node_box);

}  // namespace rtw
