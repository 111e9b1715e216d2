// lbvh.cu — device-side tree build (opt-in: RTW_DEVICE_BVH=1 at rtw_scene_create).
//
// The reference builds its bvh_node on the host with a random split axis and std::sort (hittable.h:45-95); this backend keeps
// that topology for rtw_intersect_batch and, for RENDERING, builds its own trees at upload — by default a binned-SAH build on
// the host (api.cu build_tree).  This file is the device-side alternative: a linear BVH (Lauterbach et al. 2009 / Karras 2012)
//   1. 30-bit Morton code of every object's box centre inside the scene bounds, made unique with the object index (64-bit key);
//   2. bitonic sort of the keys in global memory (log^2 n passes; the scenes here have 10^2..10^4 objects);
//   3. one thread per internal node finds its key range and split from the longest common prefixes (Karras 2012, sec. 3);
//   4. one thread per leaf walks to the root and fits the boxes bottom-up (the second arrival at a node does the union).
// Output: n - 1 internal nodes as (left, right, box); a child >= 0 is an internal node, < 0 is ~(object index).  The caller
// (api.cu) turns them into DevNode records, so traversal, box padding and the 4-wide collapse are shared with the SAH build.
// Results are independent of the tree: the closest hit of a ray is the same whichever tree finds it, which the parity test
// `test_upload_time_trees_do_not_change_hits_or_images` checks bit for bit on the images.
#include <cstdint>
#include <vector>

#include <cuda_runtime.h>

#include "lbvh.hpp"

namespace rtw {
namespace {

__device__ __forceinline__ uint32_t expand_bits10(uint32_t v) {  // 10 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__global__ void lbvh_keys(const float* __restrict__ boxes, int n, int n_pad, float3 lo, float3 inv_ext, unsigned long long* __restrict__ keys) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pad) return;
    if (i >= n) {
        keys[i] = ~0ull;  // padding sorts to the end
        return;
    }
    const float* b = boxes + 6 * (size_t)i;
    const float cx = (0.5f * (b[0] + b[3]) - lo.x) * inv_ext.x, cy = (0.5f * (b[1] + b[4]) - lo.y) * inv_ext.y,
                cz = (0.5f * (b[2] + b[5]) - lo.z) * inv_ext.z;
    const uint32_t x = (uint32_t)fminf(fmaxf(cx * 1024.0f, 0.0f), 1023.0f), y = (uint32_t)fminf(fmaxf(cy * 1024.0f, 0.0f), 1023.0f),
                   z = (uint32_t)fminf(fmaxf(cz * 1024.0f, 0.0f), 1023.0f);
    const uint32_t code = (expand_bits10(x) << 2) | (expand_bits10(y) << 1) | expand_bits10(z);
    keys[i] = ((unsigned long long)code << 32) | (uint32_t)i;
}

__global__ void lbvh_bitonic_step(unsigned long long* __restrict__ keys, int n_pad, int j, int k) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pad) return;
    const int p = i ^ j;
    if (p > i) {
        const unsigned long long a = keys[i], b = keys[p];
        const bool up = (i & k) == 0;
        if ((a > b) == up) keys[i] = b, keys[p] = a;
    }
}

// longest common prefix of keys i and j (-1 outside the range); the keys are unique, so it is < 64
__device__ __forceinline__ int lcp(const unsigned long long* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    return __clzll((long long)(keys[i] ^ keys[j]));
}

__global__ void lbvh_nodes(const unsigned long long* __restrict__ keys, int n, int* __restrict__ left, int* __restrict__ right,
                           int* __restrict__ parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    // direction of the range and its other end (Karras 2012, fig. 4)
    const int d = (lcp(keys, n, i, i + 1) - lcp(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = lcp(keys, n, i, i - d);
    int lmax = 2;
    while (lcp(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (lcp(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    // the split: highest differing bit inside [i, j]
    const int dnode = lcp(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) / 2;; t = (t + 1) / 2) {
        if (lcp(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int lc = (lo == gamma) ? ~(int)(uint32_t)keys[gamma] : gamma;          // a leaf child = ~(object index)
    const int rc = (hi == gamma + 1) ? ~(int)(uint32_t)keys[gamma + 1] : gamma + 1;
    left[i] = lc, right[i] = rc;
    if (lo == gamma) parent[n - 1 + gamma] = i;  // leaf slots of `parent` follow the n - 1 internal ones, in sorted order
    else parent[gamma] = i;
    if (hi == gamma + 1) parent[n - 1 + gamma + 1] = i;
    else parent[gamma + 1] = i;
    if (i == 0) parent[0] = -1;
}

__global__ void lbvh_fit(const unsigned long long* __restrict__ keys, const float* __restrict__ boxes, int n, const int* __restrict__ left,
                         const int* __restrict__ right, const int* __restrict__ parent, float* __restrict__ node_box, int* __restrict__ visits) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;  // sorted leaf position
    if (s >= n) return;
    int node = parent[n - 1 + s];
    while (node >= 0) {
        if (atomicAdd(&visits[node], 1) == 0) return;  // the first child to arrive stops; the second has both boxes at hand
        __threadfence();
        float b[6] = {3.0e38f, 3.0e38f, 3.0e38f, -3.0e38f, -3.0e38f, -3.0e38f};
        const int c[2] = {left[node], right[node]};
        for (int k = 0; k < 2; k++) {
            const volatile float* cb = c[k] < 0 ? boxes + 6 * (size_t)(~c[k]) : node_box + 6 * (size_t)c[k];
            b[0] = fminf(b[0], cb[0]), b[1] = fminf(b[1], cb[1]), b[2] = fminf(b[2], cb[2]);
            b[3] = fmaxf(b[3], cb[3]), b[4] = fmaxf(b[4], cb[4]), b[5] = fmaxf(b[5], cb[5]);
        }
        for (int k = 0; k < 6; k++) node_box[6 * (size_t)node + k] = b[k];
        __threadfence();
        node = parent[node];
    }
    (void)keys;
}

}  // namespace

cudaError_t lbvh_build(const std::vector<float>& boxes, std::vector<int>& left, std::vector<int>& right, std::vector<float>& node_box) {
    const int n = (int)(boxes.size() / 6);
    left.clear(), right.clear(), node_box.clear();
    if (n < 2) return cudaSuccess;
    int n_pad = 1;
    while (n_pad < n) n_pad *= 2;
    float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    for (int i = 0; i < n; i++)
        for (int k = 0; k < 3; k++) {
            const float c = 0.5f * (boxes[6 * i + k] + boxes[6 * i + 3 + k]);
            lo[k] = c < lo[k] ? c : lo[k], hi[k] = c > hi[k] ? c : hi[k];
        }
    float3 l3 = make_float3(lo[0], lo[1], lo[2]);
    float3 inv = make_float3(hi[0] > lo[0] ? 1.0f / (hi[0] - lo[0]) : 0.0f, hi[1] > lo[1] ? 1.0f / (hi[1] - lo[1]) : 0.0f,
                             hi[2] > lo[2] ? 1.0f / (hi[2] - lo[2]) : 0.0f);
    float *d_boxes = nullptr, *d_nbox = nullptr;
    unsigned long long* d_keys = nullptr;
    int *d_left = nullptr, *d_right = nullptr, *d_parent = nullptr, *d_visits = nullptr;
    cudaError_t e = cudaSuccess;
    auto chk = [&](cudaError_t x) {
        if (e == cudaSuccess) e = x;
    };
    chk(cudaMalloc(&d_boxes, boxes.size() * sizeof(float)));
    chk(cudaMalloc(&d_nbox, (size_t)(n - 1) * 6 * sizeof(float)));
    chk(cudaMalloc(&d_keys, (size_t)n_pad * sizeof(unsigned long long)));
    chk(cudaMalloc(&d_left, (size_t)(n - 1) * sizeof(int)));
    chk(cudaMalloc(&d_right, (size_t)(n - 1) * sizeof(int)));
    chk(cudaMalloc(&d_parent, (size_t)(2 * n - 1) * sizeof(int)));
    chk(cudaMalloc(&d_visits, (size_t)(n - 1) * sizeof(int)));
    if (e == cudaSuccess) {
        chk(cudaMemcpy(d_boxes, boxes.data(), boxes.size() * sizeof(float), cudaMemcpyHostToDevice));
        chk(cudaMemset(d_visits, 0, (size_t)(n - 1) * sizeof(int)));
        const int T = 256;
        lbvh_keys<<<(n_pad + T - 1) / T, T>>>(d_boxes, n, n_pad, l3, inv, d_keys);
        for (int k = 2; k <= n_pad; k *= 2)
            for (int j = k / 2; j > 0; j /= 2) lbvh_bitonic_step<<<(n_pad + T - 1) / T, T>>>(d_keys, n_pad, j, k);
        lbvh_nodes<<<(n - 1 + T - 1) / T, T>>>(d_keys, n, d_left, d_right, d_parent);
        lbvh_fit<<<(n + T - 1) / T, T>>>(d_keys, d_boxes, n, d_left, d_right, d_parent, d_nbox, d_visits);
        chk(cudaGetLastError());
        left.resize(n - 1), right.resize(n - 1), node_box.resize((size_t)(n - 1) * 6);
        chk(cudaMemcpy(left.data(), d_left, (size_t)(n - 1) * sizeof(int), cudaMemcpyDeviceToHost));
        chk(cudaMemcpy(right.data(), d_right, (size_t)(n - 1) * sizeof(int), cudaMemcpyDeviceToHost));
        chk(cudaMemcpy(node_box.data(), d_nbox, (size_t)(n - 1) * 6 * sizeof(float), cudaMemcpyDeviceToHost));
    }
    cudaFree(d_boxes), cudaFree(d_nbox), cudaFree(d_keys), cudaFree(d_left), cudaFree(d_right), cudaFree(d_parent), cudaFree(d_visits);
    return e;
}

}  // namespace rtw
