// flat.hpp — owning storage behind the public `rtw_flat_scene` view (include/rtw_b200.h).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/rtw_b200.h"

namespace rtw {

struct FlatStorage {
    std::vector<rtw_flat_item> items;
    std::vector<rtw_flat_prim> prims;
    std::vector<rtw_flat_xf> xfs;
    std::vector<rtw_flat_node> nodes;
    std::vector<rtw_flat_leaf> leaves;
    std::vector<rtw_flat_material> materials;
    std::vector<rtw_flat_texture> textures;
    std::vector<rtw_flat_light> lights;
    std::vector<uint8_t> texels;
    std::vector<double> perlin_ranvec;  // 768 or empty
    std::vector<int32_t> perlin_perm;   // 768 or empty
    int32_t n_world_items = 0;
    int32_t render_type = RTW_RENDER_SHADED;
    int32_t background_type = RTW_BG_GRADIENT;
    rtw_flat_camera camera{};
    rtw_flat_scene view{};

    void refresh_view();
};

void set_error(const std::string& msg);
void clear_error();

}  // namespace rtw
