// imageio.cpp — the data formats either side of the hot path, host only, no third-party code:
//   * rtw_decode_jpeg : baseline JPEG -> RGB8 rows, top row first — the bytes image_texture expects from its caller
//     (texture.h:73-99; the reference ships RayTracingWeekend/earth.jpg but no loader).  The arithmetic follows the
//     published libjpeg (IJG v6b) algorithms — the 13-bit "islow" LL&M inverse DCT, triangle-filter ("fancy") h2v1/h2v2
//     chroma upsampling and the 16-bit fixed-point YCbCr->RGB tables — so the output is byte-identical to what
//     Pillow/libjpeg-turbo produce (pinned by SHA-256 against tests/golden/earth_256x256_rgb8.bin).
//   * rtw_write_ppm_binary : "P6" twin of the reference's ASCII writer (RayTracingWeekend.cpp:252-275), same quantisation.
//   * rtw_write_png : 8-bit RGB PNG, replacing the reference's `system("magick 1.ppm 1.png")` (RayTracingWeekend.cpp:285);
//     zlib stream of stored (uncompressed) deflate blocks, CRC-32 and Adler-32 computed here.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "flat.hpp"

namespace {

// ------------------------------------------------------------------------------------------------ JPEG
struct Huff {
    // canonical decoding tables (ITU T.81 Annex F.2.2.3): per code length, min/max code and first value index
    int mincode[17], maxcode[18], valptr[17];
    uint8_t vals[256];
    bool present = false;
};
struct Comp {
    int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0;
    int bw = 0, bh = 0;  // size in blocks (padded to whole MCUs)
    int dw = 0, dh = 0;  // downsampled_width / height in samples
    std::vector<uint8_t> plane;  // bw*8 x bh*8
    int pred = 0;
};
struct BitReader {
    const uint8_t* p;
    const uint8_t* end;
    uint32_t acc = 0;
    int cnt = 0;
    bool marker = false;
    int bit() {
        if (cnt == 0) {
            uint8_t b = 0;
            if (!marker && p < end) {
                b = *p++;
                if (b == 0xFF) {
                    uint8_t b2 = p < end ? *p : 0;
                    if (b2 == 0) p++;            // stuffed zero
                    else { marker = true; p--; b = 0; }  // a marker: feed zeros, leave the pointer on it
                }
            }
            acc = b;
            cnt = 8;
        }
        cnt--;
        return (acc >> cnt) & 1;
    }
    int bits(int n) {
        int v = 0;
        while (n--) v = (v << 1) | bit();
        return v;
    }
    void reset() { acc = 0, cnt = 0, marker = false; }
};
const int kZigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48,
                         41, 34, 27, 20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22,
                         15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

int decode_symbol(BitReader& br, const Huff& h) {
    int code = 0;
    for (int l = 1; l <= 16; l++) {
        code = (code << 1) | br.bit();
        if (h.maxcode[l] >= 0 && code <= h.maxcode[l]) return h.vals[h.valptr[l] + code - h.mincode[l]];
    }
    return -1;
}
inline int extend(int v, int t) { return v < (1 << (t - 1)) ? v - (1 << t) + 1 : v; }  // T.81 F.2.2.1
inline uint8_t clamp8(int x) { return (uint8_t)(x < 0 ? 0 : (x > 255 ? 255 : x)); }

// jpeg_idct_islow (IJG jidctint.c): Loeffler-Ligtenberg-Moschytz, CONST_BITS 13, PASS1_BITS 2
void idct_islow(const int* coef, const uint16_t* q, uint8_t* out, int stride) {
    const int CB = 13, P1 = 2;
    const int F_0_298 = 2446, F_0_390 = 3196, F_0_541 = 4433, F_0_765 = 6270, F_0_899 = 7373, F_1_175 = 9633, F_1_501 = 12299,
              F_1_847 = 15137, F_1_961 = 16069, F_2_053 = 16819, F_2_562 = 20995, F_3_072 = 25172;
    auto descale = [](int64_t x, int n) { return (int)((x + ((int64_t)1 << (n - 1))) >> n); };
    int ws[64];
    for (int c = 0; c < 8; c++) {
        int in[8];
        for (int r = 0; r < 8; r++) in[r] = coef[r * 8 + c] * (int)q[r * 8 + c];
        int64_t z2 = in[2], z3 = in[6];
        int64_t z1 = (z2 + z3) * F_0_541;
        int64_t tmp2 = z1 + z3 * (-F_1_847), tmp3 = z1 + z2 * F_0_765;
        z2 = in[0], z3 = in[4];
        int64_t tmp0 = (z2 + z3) << CB, tmp1 = (z2 - z3) << CB;
        int64_t tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
        tmp0 = in[7], tmp1 = in[5], tmp2 = in[3], tmp3 = in[1];
        z1 = tmp0 + tmp3, z2 = tmp1 + tmp2, z3 = tmp0 + tmp2;
        int64_t z4 = tmp1 + tmp3, z5 = (z3 + z4) * F_1_175;
        tmp0 *= F_0_298, tmp1 *= F_2_053, tmp2 *= F_3_072, tmp3 *= F_1_501;
        z1 *= -F_0_899, z2 *= -F_2_562, z3 *= -F_1_961, z4 *= -F_0_390;
        z3 += z5, z4 += z5;
        tmp0 += z1 + z3, tmp1 += z2 + z4, tmp2 += z2 + z3, tmp3 += z1 + z4;
        ws[0 * 8 + c] = descale(tmp10 + tmp3, CB - P1), ws[7 * 8 + c] = descale(tmp10 - tmp3, CB - P1);
        ws[1 * 8 + c] = descale(tmp11 + tmp2, CB - P1), ws[6 * 8 + c] = descale(tmp11 - tmp2, CB - P1);
        ws[2 * 8 + c] = descale(tmp12 + tmp1, CB - P1), ws[5 * 8 + c] = descale(tmp12 - tmp1, CB - P1);
        ws[3 * 8 + c] = descale(tmp13 + tmp0, CB - P1), ws[4 * 8 + c] = descale(tmp13 - tmp0, CB - P1);
    }
    for (int r = 0; r < 8; r++) {
        const int* w = ws + r * 8;
        int64_t z2 = w[2], z3 = w[6];
        int64_t z1 = (z2 + z3) * F_0_541;
        int64_t tmp2 = z1 + z3 * (-F_1_847), tmp3 = z1 + z2 * F_0_765;
        int64_t tmp0 = ((int64_t)w[0] + w[4]) << CB, tmp1 = ((int64_t)w[0] - w[4]) << CB;
        int64_t tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
        tmp0 = w[7], tmp1 = w[5], tmp2 = w[3], tmp3 = w[1];
        z1 = tmp0 + tmp3, z2 = tmp1 + tmp2, z3 = tmp0 + tmp2;
        int64_t z4 = tmp1 + tmp3, z5 = (z3 + z4) * F_1_175;
        tmp0 *= F_0_298, tmp1 *= F_2_053, tmp2 *= F_3_072, tmp3 *= F_1_501;
        z1 *= -F_0_899, z2 *= -F_2_562, z3 *= -F_1_961, z4 *= -F_0_390;
        z3 += z5, z4 += z5;
        tmp0 += z1 + z3, tmp1 += z2 + z4, tmp2 += z2 + z3, tmp3 += z1 + z4;
        const int S = CB + P1 + 3;
        uint8_t* o = out + r * stride;
        o[0] = clamp8(descale(tmp10 + tmp3, S) + 128), o[7] = clamp8(descale(tmp10 - tmp3, S) + 128);
        o[1] = clamp8(descale(tmp11 + tmp2, S) + 128), o[6] = clamp8(descale(tmp11 - tmp2, S) + 128);
        o[2] = clamp8(descale(tmp12 + tmp1, S) + 128), o[5] = clamp8(descale(tmp12 - tmp1, S) + 128);
        o[3] = clamp8(descale(tmp13 + tmp0, S) + 128), o[4] = clamp8(descale(tmp13 - tmp0, S) + 128);
    }
}

// Upsample one component plane to full resolution (IJG jdsample.c: fullsize copy, h2v1_fancy, h2v2_fancy).
bool upsample(const Comp& c, int hmax, int vmax, int width, int height, std::vector<uint8_t>& out) {
    out.assign((size_t)width * height, 0);
    const int stride = c.bw * 8;
    auto in_row = [&](int r) { return c.plane.data() + (size_t)(r < 0 ? 0 : (r >= c.dh ? c.dh - 1 : r)) * stride; };
    if (c.h == hmax && c.v == vmax) {
        for (int y = 0; y < height; y++) memcpy(&out[(size_t)y * width], in_row(y), width);
        return true;
    }
    const int dw = c.dw;
    std::vector<uint8_t> row((size_t)dw * 2 + 2);
    if (c.h * 2 == hmax && c.v == vmax) {  // h2v1_fancy_upsample
        for (int y = 0; y < height; y++) {
            const uint8_t* in = in_row(y);
            if (dw == 1) {
                row[0] = row[1] = in[0];
            } else {
                row[0] = in[0];
                row[1] = (uint8_t)((in[0] * 3 + in[1] + 2) >> 2);
                for (int x = 1; x < dw - 1; x++) {
                    int v = in[x] * 3;
                    row[2 * x] = (uint8_t)((v + in[x - 1] + 1) >> 2);
                    row[2 * x + 1] = (uint8_t)((v + in[x + 1] + 2) >> 2);
                }
                row[2 * dw - 2] = (uint8_t)((in[dw - 1] * 3 + in[dw - 2] + 1) >> 2);
                row[2 * dw - 1] = in[dw - 1];
            }
            memcpy(&out[(size_t)y * width], row.data(), width);
        }
        return true;
    }
    if (c.h * 2 == hmax && c.v * 2 == vmax) {  // h2v2_fancy_upsample
        for (int y = 0; y < height; y++) {
            const int r = y >> 1;
            const uint8_t* in0 = in_row(r);
            const uint8_t* in1 = in_row((y & 1) ? r + 1 : r - 1);
            if (dw == 1) {
                int s = in0[0] * 3 + in1[0];
                row[0] = (uint8_t)((s * 4 + 8) >> 4), row[1] = (uint8_t)((s * 4 + 7) >> 4);
            } else {
                int thiscol = in0[0] * 3 + in1[0], nextcol = in0[1] * 3 + in1[1], lastcol;
                row[0] = (uint8_t)((thiscol * 4 + 8) >> 4);
                row[1] = (uint8_t)((thiscol * 3 + nextcol + 7) >> 4);
                lastcol = thiscol, thiscol = nextcol;
                for (int x = 1; x < dw - 1; x++) {
                    nextcol = in0[x + 1] * 3 + in1[x + 1];
                    row[2 * x] = (uint8_t)((thiscol * 3 + lastcol + 8) >> 4);
                    row[2 * x + 1] = (uint8_t)((thiscol * 3 + nextcol + 7) >> 4);
                    lastcol = thiscol, thiscol = nextcol;
                }
                row[2 * dw - 2] = (uint8_t)((thiscol * 3 + lastcol + 8) >> 4);
                row[2 * dw - 1] = (uint8_t)((thiscol * 4 + 7) >> 4);
            }
            memcpy(&out[(size_t)y * width], row.data(), width);
        }
        return true;
    }
    return false;
}

int fail(const char* msg, int code = RTW_ERR_UNSUPPORTED) {
    rtw::set_error(std::string("decode_jpeg: ") + msg);
    return code;
}

// ------------------------------------------------------------------------------------------------ PNG helpers
uint32_t crc32_update(uint32_t crc, const uint8_t* p, size_t n) {
    static uint32_t table[256];
    static bool init = false;
    if (!init) {
        for (uint32_t i = 0; i < 256; i++) {
            uint32_t c = i;
            for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            table[i] = c;
        }
        init = true;
    }
    for (size_t i = 0; i < n; i++) crc = table[(crc ^ p[i]) & 0xFF] ^ (crc >> 8);
    return crc;
}
void put_be32(std::vector<uint8_t>& v, uint32_t x) {
    v.push_back((uint8_t)(x >> 24)), v.push_back((uint8_t)(x >> 16)), v.push_back((uint8_t)(x >> 8)), v.push_back((uint8_t)x);
}
void put_chunk(std::vector<uint8_t>& file, const char type[4], const std::vector<uint8_t>& data) {
    put_be32(file, (uint32_t)data.size());
    size_t start = file.size();
    file.insert(file.end(), type, type + 4);
    file.insert(file.end(), data.begin(), data.end());
    uint32_t crc = crc32_update(0xFFFFFFFFu, file.data() + start, file.size() - start) ^ 0xFFFFFFFFu;
    put_be32(file, crc);
}
// :268-270 quantisation shared by every writer: int(255.99f * c), top row first (:261)
void quantise_rows(const double* canvas, int nx, int ny, std::vector<uint8_t>& rgb) {
    rgb.resize((size_t)nx * ny * 3);
    size_t o = 0;
    for (int j = ny - 1; j >= 0; j--)
        for (int i = 0; i < nx; i++) {
            const double* col = &canvas[3 * ((size_t)j * nx + i)];
            for (int c = 0; c < 3; c++) {
                int v = int(255.99f * col[c]);
                rgb[o++] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
            }
        }
}
int write_file(const char* path, const std::vector<uint8_t>& bytes, const char* what) {
    FILE* f = fopen(path, "wb");
    if (!f) {
        rtw::set_error(std::string(what) + ": cannot open " + path);
        return RTW_ERR_IO;
    }
    bool ok = fwrite(bytes.data(), 1, bytes.size(), f) == bytes.size();
    ok = (fclose(f) == 0) && ok;
    if (!ok) {
        rtw::set_error(std::string(what) + ": write failed");
        return RTW_ERR_IO;
    }
    return RTW_OK;
}

}  // namespace

extern "C" {

int rtw_decode_jpeg(const uint8_t* data, size_t size, uint8_t* rgb8, size_t rgb8_capacity, int* out_nx, int* out_ny) {
    if (!data || size < 4 || !out_nx || !out_ny) return fail("bad arguments", RTW_ERR_INVALID);
    if (data[0] != 0xFF || data[1] != 0xD8) return fail("not a JPEG (no SOI)", RTW_ERR_INVALID);
    uint16_t qt[4][64] = {};
    Huff dc[4], ac[4];
    std::vector<Comp> comps;
    int width = 0, height = 0, restart = 0;
    int adobe_transform = -1;
    size_t i = 2;
    bool have_sof = false;
    while (i + 4 <= size) {
        if (data[i] != 0xFF) return fail("marker expected", RTW_ERR_INVALID);
        while (i < size && data[i] == 0xFF) i++;  // fill bytes
        const int m = data[i++];
        if (m == 0xD9) break;
        if (i + 2 > size) return fail("truncated", RTW_ERR_INVALID);
        const size_t len = ((size_t)data[i] << 8) | data[i + 1];
        if (len < 2 || i + len > size) return fail("bad segment length", RTW_ERR_INVALID);
        const uint8_t* s = data + i + 2;
        const size_t n = len - 2;
        if (m == 0xDB) {  // DQT
            size_t k = 0;
            while (k < n) {
                int pq = s[k] >> 4, tq = s[k] & 15;
                k++;
                if (tq > 3 || k + (pq ? 128 : 64) > n) return fail("bad DQT", RTW_ERR_INVALID);
                for (int z = 0; z < 64; z++) {
                    qt[tq][kZigzag[z]] = pq ? (uint16_t)((s[k] << 8) | s[k + 1]) : s[k];
                    k += pq ? 2 : 1;
                }
            }
        } else if (m == 0xC0 || m == 0xC1) {  // SOF0/1: baseline / extended sequential, Huffman
            if (n < 6 || s[0] != 8) return fail("only 8-bit precision");
            height = (s[1] << 8) | s[2], width = (s[3] << 8) | s[4];
            int nc = s[5];
            if ((nc != 1 && nc != 3) || n < 6 + 3 * (size_t)nc || width <= 0 || height <= 0) return fail("only 1 or 3 components");
            comps.resize(nc);
            for (int c = 0; c < nc; c++) {
                comps[c].id = s[6 + 3 * c], comps[c].h = s[7 + 3 * c] >> 4, comps[c].v = s[7 + 3 * c] & 15, comps[c].tq = s[8 + 3 * c] & 3;
                if (comps[c].h < 1 || comps[c].h > 2 || comps[c].v < 1 || comps[c].v > 2) return fail("sampling factors beyond 2 are not supported");
            }
            have_sof = true;
        } else if (m == 0xC2 || (m >= 0xC3 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC)) {
            return fail("progressive / lossless / arithmetic JPEG is not supported (baseline only)");
        } else if (m == 0xC4) {  // DHT
            size_t k = 0;
            while (k + 17 <= n) {
                int tc = s[k] >> 4, th = s[k] & 15;
                if (tc > 1 || th > 3) return fail("bad DHT", RTW_ERR_INVALID);
                Huff& h = tc ? ac[th] : dc[th];
                int counts[17] = {}, total = 0;
                for (int l = 1; l <= 16; l++) counts[l] = s[k + l], total += counts[l];
                k += 17;
                if (total > 256 || k + total > n) return fail("bad DHT", RTW_ERR_INVALID);
                memcpy(h.vals, s + k, total);
                k += total;
                int code = 0, idx = 0;
                for (int l = 1; l <= 16; l++) {
                    h.valptr[l] = idx, h.mincode[l] = code;
                    code += counts[l], idx += counts[l];
                    h.maxcode[l] = counts[l] ? code - 1 : -1;
                    code <<= 1;
                }
                h.present = true;
            }
        } else if (m == 0xDD) {
            if (n >= 2) restart = (s[0] << 8) | s[1];
        } else if (m == 0xEE) {  // Adobe APP14: transform flag
            if (n >= 12 && memcmp(s, "Adobe", 5) == 0) adobe_transform = s[11];
        } else if (m == 0xDA) {  // SOS: the single scan of a baseline file
            if (!have_sof) return fail("SOS before SOF", RTW_ERR_INVALID);
            if (n < 1) return fail("truncated SOS", RTW_ERR_INVALID);
            int ns = s[0];
            if (ns != (int)comps.size() || n < 1 + 2 * (size_t)ns + 3) return fail("non-interleaved multi-scan files are not supported");
            for (int c = 0; c < ns; c++) {
                int cid = s[1 + 2 * c];
                for (auto& cp : comps)
                    if (cp.id == cid) cp.td = s[2 + 2 * c] >> 4, cp.ta = s[2 + 2 * c] & 15;
            }
            for (auto& cp : comps)  // four tables of each class exist (DHT above rejects th > 3): a selector beyond them is corrupt
                if (cp.td > 3 || cp.ta > 3) return fail("bad Huffman table selector in SOS", RTW_ERR_INVALID);
            int hmax = 1, vmax = 1;
            for (auto& c : comps) hmax = c.h > hmax ? c.h : hmax, vmax = c.v > vmax ? c.v : vmax;
            const int mcux = (width + 8 * hmax - 1) / (8 * hmax), mcuy = (height + 8 * vmax - 1) / (8 * vmax);
            for (auto& c : comps) {
                c.bw = mcux * c.h, c.bh = mcuy * c.v;
                c.dw = (width * c.h + hmax - 1) / hmax, c.dh = (height * c.v + vmax - 1) / vmax;
                c.plane.assign((size_t)c.bw * 8 * c.bh * 8, 0);
                c.pred = 0;
                if (!dc[c.td].present || !ac[c.ta].present) return fail("missing Huffman table", RTW_ERR_INVALID);
            }
            BitReader br{data + i + len, data + size};
            int until_restart = restart;
            for (int my = 0; my < mcuy; my++)
                for (int mx = 0; mx < mcux; mx++) {
                    if (restart && until_restart == 0) {  // RSTn: byte-align, skip the marker, reset predictors
                        br.reset();
                        while (br.p + 1 < br.end && !(br.p[0] == 0xFF && br.p[1] >= 0xD0 && br.p[1] <= 0xD7)) br.p++;
                        br.p += 2;
                        for (auto& c : comps) c.pred = 0;
                        until_restart = restart;
                    }
                    until_restart--;
                    for (auto& c : comps)
                        for (int by = 0; by < c.v; by++)
                            for (int bx = 0; bx < c.h; bx++) {
                                int coef[64] = {};
                                int t = decode_symbol(br, dc[c.td]);
                                if (t < 0 || t > 11) return fail("corrupt DC code", RTW_ERR_INVALID);
                                int diff = t ? extend(br.bits(t), t) : 0;
                                c.pred += diff;
                                coef[0] = c.pred;
                                for (int k = 1; k < 64;) {
                                    int rs = decode_symbol(br, ac[c.ta]);
                                    if (rs < 0) return fail("corrupt AC code", RTW_ERR_INVALID);
                                    int r = rs >> 4, sz = rs & 15;
                                    if (sz == 0) {
                                        if (r != 15) break;  // EOB
                                        k += 16;
                                        continue;
                                    }
                                    k += r;
                                    if (k > 63) return fail("corrupt AC run", RTW_ERR_INVALID);
                                    coef[kZigzag[k]] = extend(br.bits(sz), sz);
                                    k++;
                                }
                                const int px = (mx * c.h + bx) * 8, py = (my * c.v + by) * 8;
                                idct_islow(coef, qt[c.tq], &c.plane[(size_t)py * c.bw * 8 + px], c.bw * 8);
                            }
                }
            *out_nx = width, *out_ny = height;
            const size_t need = (size_t)width * height * 3;
            if (!rgb8) return RTW_OK;  // size query
            if (rgb8_capacity < need) return fail("output buffer too small", RTW_ERR_INVALID);
            std::vector<uint8_t> full[3];
            for (size_t c = 0; c < comps.size(); c++)
                if (!upsample(comps[c], hmax, vmax, width, height, full[c])) return fail("unsupported chroma subsampling");
            if (comps.size() == 1) {
                for (size_t k = 0; k < (size_t)width * height; k++) rgb8[3 * k] = rgb8[3 * k + 1] = rgb8[3 * k + 2] = full[0][k];
                return RTW_OK;
            }
            if (adobe_transform == 0) {  // stored as RGB
                for (size_t k = 0; k < (size_t)width * height; k++) rgb8[3 * k] = full[0][k], rgb8[3 * k + 1] = full[1][k], rgb8[3 * k + 2] = full[2][k];
                return RTW_OK;
            }
            // IJG jdcolor.c build_ycc_rgb_table / ycc_rgb_convert: SCALEBITS 16
            const int SB = 16, HALF = 1 << 15;
            auto FIX = [](double x) { return (int)(x * 65536.0 + 0.5); };
            int cr_r[256], cb_b[256], cr_g[256], cb_g[256];
            for (int v = 0; v < 256; v++) {
                int x = v - 128;
                cr_r[v] = (FIX(1.40200) * x + HALF) >> SB;
                cb_b[v] = (FIX(1.77200) * x + HALF) >> SB;
                cr_g[v] = -FIX(0.71414) * x;
                cb_g[v] = -FIX(0.34414) * x + HALF;
            }
            for (size_t k = 0; k < (size_t)width * height; k++) {
                int y = full[0][k], cb = full[1][k], cr = full[2][k];
                rgb8[3 * k] = clamp8(y + cr_r[cr]);
                rgb8[3 * k + 1] = clamp8(y + ((cb_g[cb] + cr_g[cr]) >> SB));
                rgb8[3 * k + 2] = clamp8(y + cb_b[cb]);
            }
            return RTW_OK;
        }
        i += len;
    }
    return fail("no scan found", RTW_ERR_INVALID);
}

int rtw_write_ppm_binary(const char* path, const double* canvas, int nx, int ny) {
    if (!path || !canvas || nx <= 0 || ny <= 0) {
        rtw::set_error("write_ppm_binary: bad arguments");
        return RTW_ERR_INVALID;
    }
    std::vector<uint8_t> rgb, file;
    quantise_rows(canvas, nx, ny, rgb);
    char head[64];
    int hl = snprintf(head, sizeof head, "P6\n%d %d\n255\n", nx, ny);
    file.insert(file.end(), head, head + hl);
    file.insert(file.end(), rgb.begin(), rgb.end());
    return write_file(path, file, "write_ppm_binary");
}

int rtw_write_png(const char* path, const double* canvas, int nx, int ny) {
    if (!path || !canvas || nx <= 0 || ny <= 0) {
        rtw::set_error("write_png: bad arguments");
        return RTW_ERR_INVALID;
    }
    std::vector<uint8_t> rgb;
    quantise_rows(canvas, nx, ny, rgb);
    // raw scanlines: filter byte 0 + RGB
    std::vector<uint8_t> raw;
    raw.reserve((size_t)ny * (1 + 3 * (size_t)nx));
    for (int y = 0; y < ny; y++) {
        raw.push_back(0);
        raw.insert(raw.end(), rgb.begin() + (size_t)y * nx * 3, rgb.begin() + (size_t)(y + 1) * nx * 3);
    }
    // zlib container, stored deflate blocks (<= 65535 bytes each), Adler-32
    std::vector<uint8_t> z;
    z.push_back(0x78), z.push_back(0x01);
    size_t pos = 0;
    do {
        size_t n = raw.size() - pos;
        if (n > 65535) n = 65535;
        z.push_back(pos + n == raw.size() ? 1 : 0);
        z.push_back((uint8_t)(n & 0xFF)), z.push_back((uint8_t)(n >> 8));
        z.push_back((uint8_t)(~n & 0xFF)), z.push_back((uint8_t)((~n >> 8) & 0xFF));
        z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + n);
        pos += n;
    } while (pos < raw.size());
    uint32_t a = 1, b = 0;
    for (uint8_t v : raw) a = (a + v) % 65521u, b = (b + a) % 65521u;
    put_be32(z, (b << 16) | a);
    std::vector<uint8_t> file = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    std::vector<uint8_t> ihdr;
    put_be32(ihdr, (uint32_t)nx), put_be32(ihdr, (uint32_t)ny);
    ihdr.push_back(8), ihdr.push_back(2), ihdr.push_back(0), ihdr.push_back(0), ihdr.push_back(0);  // 8-bit RGB
    put_chunk(file, "IHDR", ihdr);
    put_chunk(file, "IDAT", z);
    put_chunk(file, "IEND", {});
    return write_file(path, file, "write_png");
}

}  // extern "C"
