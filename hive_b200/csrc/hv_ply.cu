// PLY writers (host code): the on-disk format either side of the mesh kernels.
//
// Replace meshwrite / pcwrite of third_party/tsdf_fusion_python/fusion.py:378-438, which format one vertex per Python
// `%` operation (several seconds for the ~10^6 vertices of a 512^3 extraction -- longer than integrating the frames).
// The ASCII output is byte-identical to the reference's ("%f %f %f %f %f %f %d %d %d\n" per vertex, "3 %d %d %d\n" per
// face, same header); rows are formatted by all host threads into per-chunk buffers and written in order.
// `binary = 1` writes binary_little_endian 1.0 with the same elements and properties instead (not a reference format).
#include "hv_common.cuh"

#include <string.h>
#include <string>
#include <thread>
#include <vector>

namespace hv {

static int host_threads(int64_t rows) {
    unsigned int n = std::thread::hardware_concurrency();
    if (n == 0) n = 4;
    if (n > 32) n = 32;
    const int64_t by_rows = rows / 4096 + 1;  // not worth a thread below a few thousand rows
    return (int)(by_rows < (int64_t)n ? by_rows : n);
}

// Runs fmt(begin, end, out) over [0, rows) in contiguous chunks on `host_threads` threads; chunks[i] holds chunk i's text.
template <typename F>
static void format_rows(int64_t rows, std::vector<std::string>& chunks, F fmt) {
    const int nt = host_threads(rows);
    chunks.assign(nt, std::string());
    std::vector<std::thread> pool;
    for (int t = 0; t < nt; ++t) {
        const int64_t b = rows * t / nt, e = rows * (t + 1) / nt;
        pool.emplace_back([&, t, b, e] { fmt(b, e, chunks[t]); });
    }
    for (auto& th : pool) th.join();
}

static bool write_chunks(FILE* f, const std::vector<std::string>& chunks) {
    for (const auto& c : chunks)
        if (!c.empty() && fwrite(c.data(), 1, c.size(), f) != c.size()) return false;
    return true;
}

}  // namespace hv

extern "C" int hv_meshwrite_ply(const char* filename, const double* verts, const int32_t* faces, const double* norms,
                                const uint8_t* colors, int64_t n_verts, int64_t n_faces, int binary) {
    HV_REQUIRE(filename && n_verts >= 0 && n_faces >= 0, "hv_meshwrite_ply: bad arguments");
    HV_REQUIRE((n_verts == 0 || (verts && norms && colors)) && (n_faces == 0 || faces), "hv_meshwrite_ply: null pointer");
    FILE* f = fopen(filename, binary ? "wb" : "w");
    HV_REQUIRE(f != nullptr, "hv_meshwrite_ply: cannot open %s", filename);
    fprintf(f,
            "ply\nformat %s 1.0\nelement vertex %lld\nproperty float x\nproperty float y\nproperty float z\n"
            "property float nx\nproperty float ny\nproperty float nz\nproperty uchar red\nproperty uchar green\n"
            "property uchar blue\nelement face %lld\nproperty list uchar int vertex_index\nend_header\n",
            binary ? "binary_little_endian" : "ascii", (long long)n_verts, (long long)n_faces);
    std::vector<std::string> chunks;
    bool ok = true;
    if (!binary) {
        hv::format_rows(n_verts, chunks, [&](int64_t b, int64_t e, std::string& out) {
            out.reserve((size_t)(e - b) * 72);
            char line[512];
            for (int64_t i = b; i < e; ++i) {
                const int n = snprintf(line, sizeof(line), "%f %f %f %f %f %f %d %d %d\n", verts[3 * i], verts[3 * i + 1],
                                       verts[3 * i + 2], norms[3 * i], norms[3 * i + 1], norms[3 * i + 2],
                                       (int)colors[3 * i], (int)colors[3 * i + 1], (int)colors[3 * i + 2]);
                out.append(line, (size_t)(n < (int)sizeof(line) ? n : (int)sizeof(line) - 1));
            }
        });
        ok = hv::write_chunks(f, chunks);
        hv::format_rows(n_faces, chunks, [&](int64_t b, int64_t e, std::string& out) {
            out.reserve((size_t)(e - b) * 28);
            char line[64];
            for (int64_t i = b; i < e; ++i) {
                const int n = snprintf(line, sizeof(line), "3 %d %d %d\n", faces[3 * i], faces[3 * i + 1], faces[3 * i + 2]);
                out.append(line, (size_t)n);
            }
        });
        ok = ok && hv::write_chunks(f, chunks);
    } else {
        hv::format_rows(n_verts, chunks, [&](int64_t b, int64_t e, std::string& out) {
            out.resize((size_t)(e - b) * 27);
            char* p = &out[0];
            for (int64_t i = b; i < e; ++i, p += 27) {
                float v[6] = {(float)verts[3 * i], (float)verts[3 * i + 1], (float)verts[3 * i + 2],
                              (float)norms[3 * i], (float)norms[3 * i + 1], (float)norms[3 * i + 2]};
                memcpy(p, v, 24);
                memcpy(p + 24, colors + 3 * i, 3);
            }
        });
        ok = hv::write_chunks(f, chunks);
        hv::format_rows(n_faces, chunks, [&](int64_t b, int64_t e, std::string& out) {
            out.resize((size_t)(e - b) * 13);
            char* p = &out[0];
            for (int64_t i = b; i < e; ++i, p += 13) {
                p[0] = 3;
                memcpy(p + 1, faces + 3 * i, 12);
            }
        });
        ok = ok && hv::write_chunks(f, chunks);
    }
    ok = (fclose(f) == 0) && ok;
    HV_REQUIRE(ok, "hv_meshwrite_ply: short write to %s", filename);
    return HV_OK;
}

extern "C" int hv_pcwrite_ply(const char* filename, const double* xyz, const uint8_t* rgb, int64_t n, int binary) {
    HV_REQUIRE(filename && n >= 0 && (n == 0 || (xyz && rgb)), "hv_pcwrite_ply: bad arguments");
    FILE* f = fopen(filename, binary ? "wb" : "w");
    HV_REQUIRE(f != nullptr, "hv_pcwrite_ply: cannot open %s", filename);
    fprintf(f,
            "ply\nformat %s 1.0\nelement vertex %lld\nproperty float x\nproperty float y\nproperty float z\n"
            "property uchar red\nproperty uchar green\nproperty uchar blue\nend_header\n",
            binary ? "binary_little_endian" : "ascii", (long long)n);
    std::vector<std::string> chunks;
    if (!binary) {
        hv::format_rows(n, chunks, [&](int64_t b, int64_t e, std::string& out) {
            out.reserve((size_t)(e - b) * 40);
            char line[1024];
            for (int64_t i = b; i < e; ++i) {
                const int k = snprintf(line, sizeof(line), "%f %f %f %d %d %d\n", xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2],
                                       (int)rgb[3 * i], (int)rgb[3 * i + 1], (int)rgb[3 * i + 2]);
                out.append(line, (size_t)(k < (int)sizeof(line) ? k : (int)sizeof(line) - 1));
            }
        });
    } else {
        hv::format_rows(n, chunks, [&](int64_t b, int64_t e, std::string& out) {
            out.resize((size_t)(e - b) * 15);
            char* p = &out[0];
            for (int64_t i = b; i < e; ++i, p += 15) {
                float v[3] = {(float)xyz[3 * i], (float)xyz[3 * i + 1], (float)xyz[3 * i + 2]};
                memcpy(p, v, 12);
                memcpy(p + 12, rgb + 3 * i, 3);
            }
        });
    }
    bool ok = hv::write_chunks(f, chunks);
    ok = (fclose(f) == 0) && ok;
    HV_REQUIRE(ok, "hv_pcwrite_ply: short write to %s", filename);
    return HV_OK;
}
