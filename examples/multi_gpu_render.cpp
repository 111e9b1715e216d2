// multi_gpu_render.cpp — what a maintainer of the reference writes to render on several B200s: the UNMODIFIED reference's
// scene class (Scene/scene.h) is described to the backend by the adapter, every rank (one process per GPU) calls
// rtw_b200::render with the same arguments, and rank 0 alone gets the canvas and runs the reference's own PPM writer
// loop (RayTracingWeekend.cpp:252-275) — everything else of main() stays as it is.
//
//   multi_gpu_render <rank> <world> <id-file> <nx> <ny> <spp> <out.ppm>
//
// Rank 0 draws the NCCL id (rtw_comm_unique_id) and publishes it through <id-file>; the other ranks wait for the file.
// Built by oracle/Makefile (`make example`) against /root/reference, like the oracle harness, into oracle/_ref/ (it
// includes reference headers, so it is built where the reference lies and is not part of the product library).
#include <cfloat>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>
#define __int64 long long

#define main reference_main
#include "RayTracingWeekend.cpp"  // the reference translation unit: scene classes, vec3, camera (its color() is not called here)
#undef main

#include "rtw_b200_adapter.hpp"

int main(int argc, char** argv) {
    if (argc < 8) {
        fprintf(stderr, "usage: %s <rank> <world> <id-file> <nx> <ny> <spp> <out.ppm>\n", argv[0]);
        return 2;
    }
    const int rank = atoi(argv[1]), world = atoi(argv[2]);
    const std::string id_file = argv[3];
    const int nx = atoi(argv[4]), ny = atoi(argv[5]), spp = atoi(argv[6]);
    uint8_t id[RTW_COMM_ID_BYTES];
    if (rank == 0) {
        if (rtw_comm_unique_id(id) != RTW_OK) {
            fprintf(stderr, "rtw_comm_unique_id: %s\n", rtw_last_error());
            return 1;
        }
        FILE* f = fopen((id_file + ".tmp").c_str(), "wb");
        if (!f || fwrite(id, 1, sizeof id, f) != sizeof id) return 1;
        fclose(f);
        rename((id_file + ".tmp").c_str(), id_file.c_str());  // atomic publish
    } else {
        for (int tries = 0;; tries++) {
            FILE* f = fopen(id_file.c_str(), "rb");
            if (f) {
                const size_t n = fread(id, 1, sizeof id, f);
                fclose(f);
                if (n == sizeof id) break;
            }
            if (tries > 3000) {
                fprintf(stderr, "rank %d: no NCCL id in %s\n", rank, id_file.c_str());
                return 1;
            }
            std::this_thread::sleep_for(std::chrono::milliseconds(10));
        }
    }
    rtw_comm* comm = rtw_comm_create(id, rank, world, /*device=*/rank);
    if (!comm) {
        fprintf(stderr, "rank %d: rtw_comm_create: %s\n", rank, rtw_last_error());
        return 1;
    }

    cornell_box_scene scene(nx * 1.0 / ny);  // RayTracingWeekend.cpp:199-204, the scene typedef'd there
    std::vector<vec3> canvas;                // :210
    rtw_b200::render_options opt;
    opt.device = rank, opt.comm = comm, opt.root = 0;
    rtw_render_stats st{};
    opt.stats = &st;
    try {
        rtw_b200::render(scene, nx, ny, spp, /*max_depth=*/100, canvas, opt);  // replaces :211-250 on every rank
    } catch (const std::exception& e) {
        fprintf(stderr, "rank %d: %s\n", rank, e.what());
        return 1;
    }
    printf("rank %d of %d: %llu paths, %llu rays, %.2f ms\n", rank, world, (unsigned long long)st.paths, (unsigned long long)st.rays, st.kernel_ms);
    int rc = 0;
    if (rank == 0) {  // :252-275, the reference's writer loop on the root's canvas
        FILE* f = fopen(argv[7], "w");
        if (!f) return 1;
        fprintf(f, "P3\n%d %d\n255\n", nx, ny);
        for (int j = ny - 1; j >= 0; j--)
            for (int i = 0; i < nx; i++) {
                const vec3& c = canvas[(size_t)j * nx + i];
                fprintf(f, "%d %d %d\n", int(255.99f * c.r), int(255.99f * c.g), int(255.99f * c.b));
            }
        fclose(f);
    } else if (!canvas.empty()) {
        rc = 1;  // only the root owns an image
    }
    rtw_comm_destroy(comm);
    return rc;
}
