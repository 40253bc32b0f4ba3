// ply_io.cu -- host-side codec of the Gaussian PLY file (include/b200splat_next.h; the wire format of
// GaussianModel.save_ply / load_ply, /root/reference/scene/gaussian_model.py:218-259,267-320).  Host memory
// only; compiled with the library so the format lives next to the kernels that consume it.
//
// File: "ply / format binary_little_endian 1.0 / element vertex P / property float <name> ... / end_header",
// then P rows of little-endian fp32 in header order.  Rows are (de)interleaved by worker threads over
// contiguous row ranges; the SH coefficients change layout between memory ([P, K, 3]) and file (channel-major).
#include <cerrno>
#include <cstdio>
#include <cstring>
#include <functional>
#include <string>
#include <thread>
#include <vector>

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

void parallel_rows(int64_t P, const std::function<void(int64_t, int64_t)>& fn) {
    unsigned hw = std::thread::hardware_concurrency();
    int64_t nt = hw ? (int64_t)hw : 4;
    if (nt > 32) nt = 32;
    if (P < 65536 || nt < 2) { fn(0, P); return; }
    std::vector<std::thread> th;
    int64_t per = (P + nt - 1) / nt;
    for (int64_t t = 0; t < nt; ++t) {
        int64_t lo = t * per, hi = lo + per < P ? lo + per : P;
        if (lo >= hi) break;
        th.emplace_back(fn, lo, hi);
    }
    for (auto& t : th) t.join();
}

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

    FILE* f = fopen(path, "wb");
    if (!f) return report_error(B200SPLAT_ERR_INVALID_ARG, (std::string("ply_write: cannot open ") + path + ": " + strerror(errno)).c_str());
    bool ok = fwrite(hdr.data(), 1, hdr.size(), f) == hdr.size();
    const int64_t CH = 1 << 16;               // rows per chunk
    std::vector<float> buf((size_t)CH * cols);
    for (int64_t r0 = 0; ok && r0 < P; r0 += CH) {
        int64_t n = P - r0 < CH ? P - r0 : CH;
        parallel_rows(n, [&](int64_t lo, int64_t hi) {
            for (int64_t j = lo; j < hi; ++j) {
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
        });
        ok = fwrite(buf.data(), sizeof(float) * cols, (size_t)n, f) == (size_t)n;
    }
    ok = (fclose(f) == 0) && ok;
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
    const int64_t CH = 1 << 16;
    std::vector<unsigned char> buf((size_t)CH * h.row_bytes);
    bool ok = fseek(f, h.data_start, SEEK_SET) == 0;
    for (int64_t r0 = 0; ok && r0 < P; r0 += CH) {
        int64_t n = P - r0 < CH ? P - r0 : CH;
        ok = fread(buf.data(), (size_t)h.row_bytes, (size_t)n, f) == (size_t)n;
        if (!ok) break;
        parallel_rows(n, [&](int64_t lo, int64_t hi) {
            for (int64_t j = lo; j < hi; ++j) {
                int64_t i = r0 + j;
                const unsigned char* row = buf.data() + (size_t)j * h.row_bytes;
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
        });
    }
    fclose(f);
    if (!ok) return report_error(B200SPLAT_ERR_INVALID_ARG, "ply_read: file shorter than its header says");
    return B200SPLAT_OK;
}

}  // extern "C"
