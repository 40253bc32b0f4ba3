// ply_io.cu -- host-side codec of the Gaussian PLY file (include/b200splat_next.h; the wire format of
// GaussianModel.save_ply / load_ply, /root/reference/scene/gaussian_model.py:218-259,267-320).  Host memory
// only; compiled with the library so the format lives next to the kernels that consume it.
//
// File: "ply / format binary_little_endian 1.0 / element vertex P / property float <name> ... / end_header",
// then P rows of little-endian fp32 in header order.  Every worker thread (de)interleaves its own contiguous range
// of rows through a small staging buffer and writes / reads it at its own file offset (pwrite / pread); the SH coefficients change layout between memory ([P, K, 3]) and file (channel-major).
#include <atomic>
#include <cerrno>
#include <cstdio>
#include <cstring>
#include <functional>
#include <string>
#include <thread>
#include <vector>

#include <fcntl.h>
#include <sys/types.h>
#include <unistd.h>

#include "../../include/b200splat_next.h"
#include "splat_internal.cuh"

using namespace b200splat;

namespace {

struct Prop { std::string name; int size; int offset; bool is_f32; };
struct Header { int64_t P = -1; std::vector<Prop> props; int row_bytes = 0; long data_start = 0; };

int type_size(const std::string& t, bool* f32) {
    *f32 = (t == "float" || t == "float32");
    if (t == "char" || t == "uchar" || t == "int8" || t == "uint8") return 1;
    if (t == "short" || t == "ushort" || t == "int16" || t == "uint16") return 2;
    if (t == "int" || t == "uint" || t == "int32" || t == "uint32" || *f32) return 4;
    if (t == "double" || t == "float64") return 8;
    return 0;
}

int parse_header(FILE* f, Header& h, std::string& err) {
    char line[512];
    bool first = true, in_vertex = false, fmt_ok = false, seen_vertex = false;
    while (fgets(line, sizeof(line), f)) {
        std::string s(line);
        while (!s.empty() && (s.back() == '\n' || s.back() == '\r' || s.back() == ' ')) s.pop_back();
        if (first) {
            if (s != "ply") { err = "not a PLY file"; return B200SPLAT_ERR_INVALID_ARG; }
            first = false;
            continue;
        }
        if (s == "end_header") {
            h.data_start = ftell(f);
            if (!fmt_ok) { err = "PLY: only format binary_little_endian 1.0 is supported"; return B200SPLAT_ERR_INVALID_ARG; }
            if (h.P < 0) { err = "PLY: no vertex element"; return B200SPLAT_ERR_INVALID_ARG; }
            return B200SPLAT_OK;
        }
        char a[64], b[64], c[256];
        int n = sscanf(s.c_str(), "%63s %63s %255s", a, b, c);
        if (n < 1) continue;
        std::string key(a);
        if (key == "comment" || key == "obj_info") continue;
        if (key == "format") {
            fmt_ok = (n >= 3 && std::string(b) == "binary_little_endian");
        } else if (key == "element") {
            if (n < 3) { err = "PLY: malformed element line"; return B200SPLAT_ERR_INVALID_ARG; }
            in_vertex = (std::string(b) == "vertex");
            if (in_vertex) {
                if (seen_vertex) { err = "PLY: two vertex elements"; return B200SPLAT_ERR_INVALID_ARG; }
                seen_vertex = true;
                h.P = strtoll(c, nullptr, 10);
            } else if (!seen_vertex && strtoll(c, nullptr, 10) != 0) {
                err = "PLY: a non-empty element precedes the vertex element";
                return B200SPLAT_ERR_INVALID_ARG;
            }
        } else if (key == "property" && in_vertex) {
            if (n < 3 || std::string(b) == "list") { err = "PLY: unsupported vertex property"; return B200SPLAT_ERR_INVALID_ARG; }
            bool f32 = false;
            int sz = type_size(b, &f32);
            if (sz == 0) { err = std::string("PLY: unknown property type ") + b; return B200SPLAT_ERR_INVALID_ARG; }
            h.props.push_back({c, sz, h.row_bytes, f32});
            h.row_bytes += sz;
        }
    }
    err = "PLY: header not terminated";
    return B200SPLAT_ERR_INVALID_ARG;
}

int find_f32(const Header& h, const std::string& name) {
    for (const Prop& p : h.props)
        if (p.name == name) return p.is_f32 ? p.offset : -2;
    return -1;
}

int count_rest(const Header& h) {
    int n = 0;
    for (const Prop& p : h.props)
        if (p.name.rfind("f_rest_", 0) == 0) ++n;
    return n;
}

// Rows [0, P) split into one contiguous range per worker thread; fn(lo, hi) returns false on an I/O error.
bool parallel_ranges(int64_t P, const std::function<bool(int64_t, int64_t)>& fn) {
    unsigned hw = std::thread::hardware_concurrency();
    int64_t nt = hw ? (int64_t)hw : 4;
    if (nt > 16) nt = 16;
    if (P < 65536 || nt < 2) return fn(0, P);
    std::vector<std::thread> th;
    std::atomic<bool> ok{true};
    int64_t per = (P + nt - 1) / nt;
    for (int64_t t = 0; t < nt; ++t) {
        int64_t lo = t * per, hi = lo + per < P ? lo + per : P;
        if (lo >= hi) break;
        th.emplace_back([&fn, &ok, lo, hi] { if (!fn(lo, hi)) ok = false; });
    }
    for (auto& t : th) t.join();
    return ok;
}

bool pwrite_all(int fd, const void* buf, size_t n, off_t off) {
    const char* p = (const char*)buf;
    while (n > 0) {
        ssize_t w = pwrite(fd, p, n, off);
        if (w <= 0) { if (w < 0 && errno == EINTR) continue; return false; }
        p += w; n -= (size_t)w; off += w;
    }
    return true;
}

bool pread_all(int fd, void* buf, size_t n, off_t off) {
    char* p = (char*)buf;
    while (n > 0) {
        ssize_t r = pread(fd, p, n, off);
        if (r <= 0) { if (r < 0 && errno == EINTR) continue; return false; }
        p += r; n -= (size_t)r; off += r;
    }
    return true;
}

constexpr int64_t SUB_ROWS = 8192;   // rows per thread-local staging buffer (2 MB at SH3)

}  // namespace

extern "C" {

int b200splat_ply_write(const char* path, int64_t P, int32_t n_rest, const float* xyz, const float* f_dc,
                        const float* f_rest, const float* opacity, const float* scale, const float* rot) {
    if (!path || P < 0 || n_rest < 0 || n_rest % 3 != 0) return report_error(B200SPLAT_ERR_INVALID_ARG, "ply_write: bad arguments");
    if (P > 0 && (!xyz || !f_dc || !opacity || !scale || !rot || (n_rest > 0 && !f_rest)))
        return report_error(B200SPLAT_ERR_INVALID_ARG, "ply_write: null array");
    const int K = n_rest / 3;                 // coefficients per channel in f_rest
    const int cols = 3 + 3 + 3 + n_rest + 1 + 3 + 4;
    std::string hdr = "ply\nformat binary_little_endian 1.0\nelement vertex " + std::to_string((long long)P) + "\n";
    const char* head[] = {"x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"};
    for (const char* n : head) hdr += std::string("property float ") + n + "\n";
    for (int i = 0; i < n_rest; ++i) hdr += "property float f_rest_" + std::to_string(i) + "\n";
    hdr += "property float opacity\n";
    for (int i = 0; i < 3; ++i) hdr += "property float scale_" + std::to_string(i) + "\n";
    for (int i = 0; i < 4; ++i) hdr += "property float rot_" + std::to_string(i) + "\n";
    hdr += "end_header\n";

    int fd = open(path, O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) return report_error(B200SPLAT_ERR_INVALID_ARG, (std::string("ply_write: cannot open ") + path + ": " + strerror(errno)).c_str());
    bool ok = pwrite_all(fd, hdr.data(), hdr.size(), 0);
    const off_t data0 = (off_t)hdr.size();
    const size_t row_bytes = sizeof(float) * (size_t)cols;
    // every worker interleaves its own contiguous range of rows and writes it at its own file offset
    ok = ok && parallel_ranges(P, [&](int64_t lo, int64_t hi) {
        std::vector<float> buf((size_t)SUB_ROWS * cols);
        for (int64_t r0 = lo; r0 < hi; r0 += SUB_ROWS) {
            int64_t n = hi - r0 < SUB_ROWS ? hi - r0 : SUB_ROWS;
            for (int64_t j = 0; j < n; ++j) {
                int64_t i = r0 + j;
                float* o = buf.data() + (size_t)j * cols;
                o[0] = xyz[3 * i]; o[1] = xyz[3 * i + 1]; o[2] = xyz[3 * i + 2];
                o[3] = o[4] = o[5] = 0.0f;
                o[6] = f_dc[3 * i]; o[7] = f_dc[3 * i + 1]; o[8] = f_dc[3 * i + 2];
                const float* fr = f_rest + (size_t)i * n_rest;     // [K,3] in memory -> [3,K] in the file
                for (int c = 0; c < 3; ++c)
                    for (int k = 0; k < K; ++k) o[9 + c * K + k] = fr[k * 3 + c];
                float* t = o + 9 + n_rest;
                t[0] = opacity[i];
                t[1] = scale[3 * i]; t[2] = scale[3 * i + 1]; t[3] = scale[3 * i + 2];
                t[4] = rot[4 * i]; t[5] = rot[4 * i + 1]; t[6] = rot[4 * i + 2]; t[7] = rot[4 * i + 3];
            }
            if (!pwrite_all(fd, buf.data(), row_bytes * (size_t)n, data0 + (off_t)(row_bytes * (size_t)r0))) return false;
        }
        return true;
    });
    ok = (close(fd) == 0) && ok;
    if (!ok) return report_error(B200SPLAT_ERR_INVALID_ARG, (std::string("ply_write: write failed: ") + path).c_str());
    return B200SPLAT_OK;
}

int b200splat_ply_read_header(const char* path, int64_t* P, int32_t* n_rest) {
    if (!path || !P || !n_rest) return report_error(B200SPLAT_ERR_INVALID_ARG, "ply_read_header: null argument");
    FILE* f = fopen(path, "rb");
    if (!f) return report_error(B200SPLAT_ERR_INVALID_ARG, (std::string("ply_read: cannot open ") + path + ": " + strerror(errno)).c_str());
    Header h;
    std::string err;
    int rc = parse_header(f, h, err);
    fclose(f);
    if (rc != B200SPLAT_OK) return report_error(rc, err.c_str());
    *P = h.P;
    *n_rest = count_rest(h);
    return B200SPLAT_OK;
}

int b200splat_ply_read(const char* path, int64_t P, int32_t n_rest, float* xyz, float* f_dc, float* f_rest,
                       float* opacity, float* scale, float* rot) {
    if (!path || P < 0 || n_rest < 0 || n_rest % 3 != 0) return report_error(B200SPLAT_ERR_INVALID_ARG, "ply_read: bad arguments");
    FILE* f = fopen(path, "rb");
    if (!f) return report_error(B200SPLAT_ERR_INVALID_ARG, (std::string("ply_read: cannot open ") + path + ": " + strerror(errno)).c_str());
    Header h;
    std::string err;
    int rc = parse_header(f, h, err);
    if (rc != B200SPLAT_OK) { fclose(f); return report_error(rc, err.c_str()); }
    if (h.P != P || count_rest(h) != n_rest) {
        fclose(f);
        return report_error(B200SPLAT_ERR_INVALID_ARG, "ply_read: P / n_rest differ from the file's header");
    }
    if (P > 0 && (!xyz || !f_dc || !opacity || !scale || !rot || (n_rest > 0 && !f_rest))) {
        fclose(f);
        return report_error(B200SPLAT_ERR_INVALID_ARG, "ply_read: null array");
    }
    const int K = n_rest / 3;
    // byte offsets of the wanted properties inside a row
    std::vector<int> off;
    std::vector<std::string> names = {"x", "y", "z", "f_dc_0", "f_dc_1", "f_dc_2"};
    for (int i = 0; i < n_rest; ++i) names.push_back("f_rest_" + std::to_string(i));
    names.push_back("opacity");
    for (int i = 0; i < 3; ++i) names.push_back("scale_" + std::to_string(i));
    for (int i = 0; i < 4; ++i) names.push_back("rot_" + std::to_string(i));
    for (const std::string& n : names) {
        int o = find_f32(h, n);
        if (o < 0) {
            fclose(f);
            return report_error(B200SPLAT_ERR_INVALID_ARG,
                                (std::string("ply_read: property ") + n + (o == -1 ? " missing" : " is not float")).c_str());
        }
        off.push_back(o);
    }
    fclose(f);
    int fd = open(path, O_RDONLY);
    if (fd < 0) return report_error(B200SPLAT_ERR_INVALID_ARG, (std::string("ply_read: cannot open ") + path + ": " + strerror(errno)).c_str());
    const size_t row_bytes = (size_t)h.row_bytes;
    const off_t data0 = (off_t)h.data_start;
    bool ok = parallel_ranges(P, [&](int64_t lo, int64_t hi) {
        std::vector<unsigned char> buf((size_t)SUB_ROWS * row_bytes);
        for (int64_t r0 = lo; r0 < hi; r0 += SUB_ROWS) {
            int64_t n = hi - r0 < SUB_ROWS ? hi - r0 : SUB_ROWS;
            if (!pread_all(fd, buf.data(), row_bytes * (size_t)n, data0 + (off_t)(row_bytes * (size_t)r0))) return false;
            for (int64_t j = 0; j < n; ++j) {
                int64_t i = r0 + j;
                const unsigned char* row = buf.data() + (size_t)j * row_bytes;
                auto get = [&](int idx) { float v; memcpy(&v, row + off[idx], 4); return v; };
                xyz[3 * i] = get(0); xyz[3 * i + 1] = get(1); xyz[3 * i + 2] = get(2);
                f_dc[3 * i] = get(3); f_dc[3 * i + 1] = get(4); f_dc[3 * i + 2] = get(5);
                float* fr = f_rest + (size_t)i * n_rest;           // file [3,K] -> memory [K,3]
                for (int c = 0; c < 3; ++c)
                    for (int k = 0; k < K; ++k) fr[k * 3 + c] = get(6 + c * K + k);
                int b = 6 + n_rest;
                opacity[i] = get(b);
                scale[3 * i] = get(b + 1); scale[3 * i + 1] = get(b + 2); scale[3 * i + 2] = get(b + 3);
                rot[4 * i] = get(b + 4); rot[4 * i + 1] = get(b + 5); rot[4 * i + 2] = get(b + 6); rot[4 * i + 3] = get(b + 7);
            }
        }
        return true;
    });
    close(fd);
    if (!ok) return report_error(B200SPLAT_ERR_INVALID_ARG, "ply_read: file shorter than its header says");
    return B200SPLAT_OK;
}

}  // extern "C"
