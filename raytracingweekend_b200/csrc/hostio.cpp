// hostio.cpp — host-side tail of the path: tonemap + PPM writer (rank 0).
// RayTracingWeekend.cpp:241-244 (mean, gamma 2, clamp) and :252-275 (ASCII P3, top row first).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <string>

#include "flat.hpp"

extern "C" {

int rtw_tonemap(const float* accum, int nx, int ny, double* canvas) {
    if (!accum || !canvas || nx <= 0 || ny <= 0) {
        rtw::set_error("tonemap: bad arguments");
        return RTW_ERR_INVALID;
    }
    for (size_t i = 0; i < (size_t)nx * ny; i++) {
        double n = (double)accum[4 * i + 3];  // sample count (subPixelCount, :241)
        for (int c = 0; c < 3; c++) {
            double col = (double)accum[4 * i + c] / n;
            canvas[3 * i + c] = std::min(std::sqrt(col), 1.0);  // :244 (NaN propagates like the reference)
        }
    }
    return RTW_OK;
}

int rtw_write_ppm(const char* path, const double* canvas, int nx, int ny) {
    if (!path || !canvas || nx <= 0 || ny <= 0) {
        rtw::set_error("write_ppm: bad arguments");
        return RTW_ERR_INVALID;
    }
    FILE* f = fopen(path, "w");
    if (!f) {
        rtw::set_error(std::string("write_ppm: cannot open ") + path);
        return RTW_ERR_IO;
    }
    fprintf(f, "P3\n%d %d\n255\n", nx, ny);  // :257
    for (int j = ny - 1; j >= 0; j--)       // :261
        for (int i = 0; i < nx; i++) {
            const double* col = &canvas[3 * ((size_t)j * nx + i)];
            int ir = int(255.99f * col[0]);  // :268-270 (float literal promoted to double)
            int ig = int(255.99f * col[1]);
            int ib = int(255.99f * col[2]);
            fprintf(f, "%d %d %d\n", ir, ig, ib);
        }
    bool ok = fclose(f) == 0;
    if (!ok) {
        rtw::set_error("write_ppm: write failed");
        return RTW_ERR_IO;
    }
    return RTW_OK;
}

}  // extern "C"
