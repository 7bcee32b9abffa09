// Host-side loader of MS-GS model files: binary little-endian PLY as GaussianModel.save_ply writes it
// (/root/reference/scene/gaussian_model.py:293-344: x y z nx ny nz f_dc_* f_rest_* opacity occ_multiplier_* dc_delta_*_*
// scale_* rot_* base_gaussian_mask (uchar) max_pixel_sizes min_pixel_sizes), read straight into the layout the
// rasterizer takes — what GaussianModel.load_ply (:358-417) + get_features' torch.cat (:127-153) produce:
//   means3D (P,3) | shs (P,M,3): shs[p][0][c] = f_dc_c, shs[p][k][c] = f_rest_{c*(M-1) + k-1} (the reshape (P,3,M-1) +
//   transpose(1,2) of :380-381, :404) | opacity (P,1), scales (P,3), rotations (P,4): RAW (pre-activation) values |
//   occ_multiplier (P,L,1) | dc_delta (P,3D,1) | base_mask (P) | max / min pixel sizes (P).
// The file is memory-mapped; rows are converted by `threads` host threads into caller-owned (ideally pinned) buffers,
// so the model goes file -> pinned staging -> device with one pass over the bytes and no intermediate numpy arrays
// (the reference builds ~15 float64 temporaries).  Plain 3DGS files (no MS-GS attributes) load with occ_multiplier = 1,
// dc_delta = 0, base_mask = 0, max / min pixel sizes = -1 (the "no level" values of the reference's create_from_pcd).
// Pure host code: no CUDA, built by g++ into libmsgs_ply.so.
#include "../../include/msgs_ply.h"

#include <algorithm>
#include <cstdint>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace {

thread_local std::string g_ply_err;

enum PType { T_F32, T_F64, T_U8, T_I8, T_U16, T_I16, T_U32, T_I32 };
struct Prop { std::string name; PType type; size_t offset; };

size_t type_size(PType t) {
    switch (t) {
        case T_F64: return 8;
        case T_F32: case T_U32: case T_I32: return 4;
        case T_U16: case T_I16: return 2;
        default: return 1;
    }
}
bool parse_type(const std::string& s, PType& t) {
    if (s == "float" || s == "float32") t = T_F32;
    else if (s == "double" || s == "float64") t = T_F64;
    else if (s == "uchar" || s == "uint8") t = T_U8;
    else if (s == "char" || s == "int8") t = T_I8;
    else if (s == "ushort" || s == "uint16") t = T_U16;
    else if (s == "short" || s == "int16") t = T_I16;
    else if (s == "uint" || s == "uint32") t = T_U32;
    else if (s == "int" || s == "int32") t = T_I32;
    else return false;
    return true;
}
inline float load_as_float(const unsigned char* p, PType t) {
    switch (t) {
        case T_F32: { float v; memcpy(&v, p, 4); return v; }
        case T_F64: { double v; memcpy(&v, p, 8); return (float)v; }
        case T_U8: return (float)*p;
        case T_I8: return (float)*(const signed char*)p;
        case T_U16: { uint16_t v; memcpy(&v, p, 2); return (float)v; }
        case T_I16: { int16_t v; memcpy(&v, p, 2); return (float)v; }
        case T_U32: { uint32_t v; memcpy(&v, p, 4); return (float)v; }
        default: { int32_t v; memcpy(&v, p, 4); return (float)v; }
    }
}

struct Ply {
    int fd = -1;
    const unsigned char* map = nullptr;
    size_t map_len = 0, data_off = 0, row = 0;
    int64_t P = 0;
    std::vector<Prop> props;
    // resolved columns (index into props, -1 = absent)
    int xyz[3] = {-1, -1, -1}, opacity = -1, mask = -1, maxpx = -1, minpx = -1;
    std::vector<int> fdc, frest, scale, rot, occ, dcd;
    int M = 0;      // SH coefficients per channel (1 + rest/3)
};

int find(const Ply& p, const std::string& n) {
    for (size_t i = 0; i < p.props.size(); ++i) if (p.props[i].name == n) return (int)i;
    return -1;
}
// columns named prefix + "0", prefix + "1", ... in index order, as far as they exist
std::vector<int> find_seq(const Ply& p, const std::string& prefix) {
    std::vector<int> out;
    for (int i = 0;; ++i) {
        const int k = find(p, prefix + std::to_string(i));
        if (k < 0) break;
        out.push_back(k);
    }
    return out;
}

}  // namespace

extern "C" {

const char* msgs_ply_last_error(void) { return g_ply_err.c_str(); }

void msgs_ply_close(void* handle) {
    Ply* p = (Ply*)handle;
    if (!p) return;
    if (p->map && p->map_len) munmap((void*)p->map, p->map_len);
    if (p->fd >= 0) close(p->fd);
    delete p;
}

void* msgs_ply_open(const char* path, msgs_ply_info* info) {
    if (!path || !info) { g_ply_err = "msgs_ply_open: null argument"; return nullptr; }
    Ply* p = new Ply();
    auto fail = [&](const std::string& why) -> void* { g_ply_err = "msgs_ply_open(" + std::string(path) + "): " + why; msgs_ply_close(p); return nullptr; };
    p->fd = open(path, O_RDONLY);
    if (p->fd < 0) return fail("cannot open");
    struct stat st;
    if (fstat(p->fd, &st) != 0) return fail("cannot stat");
    p->map_len = (size_t)st.st_size;
    if (p->map_len < 16) return fail("not a PLY file");
    void* m = mmap(nullptr, p->map_len, PROT_READ, MAP_PRIVATE, p->fd, 0);
    if (m == MAP_FAILED) { p->map_len = 0; return fail("mmap failed"); }
    p->map = (const unsigned char*)m;

    // ---- header: ascii lines up to "end_header\n" ----
    const char* c = (const char*)p->map;
    const size_t hmax = std::min(p->map_len, (size_t)1 << 20);
    size_t pos = 0;
    auto next_line = [&](std::string& line) -> bool {
        if (pos >= hmax) return false;
        const void* nl = memchr(c + pos, '\n', hmax - pos);
        if (!nl) return false;
        const size_t end = (const char*)nl - c;
        line.assign(c + pos, end - pos);
        if (!line.empty() && line.back() == '\r') line.pop_back();
        pos = end + 1;
        return true;
    };
    std::string line;
    if (!next_line(line) || line != "ply") return fail("missing 'ply' magic");
    bool fmt_ok = false, in_vertex = false, seen_vertex = false, done = false;
    size_t off = 0;
    while (next_line(line)) {
        if (line == "end_header") { done = true; break; }
        char a[64] = {0}, b[64] = {0}, d[64] = {0};
        const int n = sscanf(line.c_str(), "%63s %63s %63s", a, b, d);
        if (n < 1) continue;
        const std::string kw = a;
        if (kw == "format") {
            if (std::string(b) != "binary_little_endian") return fail("only binary_little_endian files are supported (got '" + std::string(b) + "')");
            fmt_ok = true;
        } else if (kw == "element") {
            if (seen_vertex && in_vertex) { in_vertex = false; }           // a later element: its data follows the vertices
            else if (std::string(b) == "vertex") {
                if (seen_vertex) return fail("two vertex elements");
                in_vertex = seen_vertex = true;
                char* endp = nullptr;
                errno = 0;
                p->P = (int64_t)strtoll(d, &endp, 10);
                if (endp == d || *endp != '\0' || errno == ERANGE) return fail("malformed vertex count '" + std::string(d) + "'");
                if (p->P < 0) return fail("negative vertex count");
            } else if (!seen_vertex) {
                return fail("an element precedes 'vertex' (unsupported)");
            }
        } else if (kw == "property" && in_vertex) {
            if (std::string(b) == "list") return fail("list property inside the vertex element");
            PType t;
            if (n < 3 || !parse_type(b, t)) return fail("unknown property type '" + std::string(b) + "'");
            p->props.push_back({d, t, off});
            off += type_size(t);
        }
    }
    if (!done) return fail("no end_header");
    if (!fmt_ok) return fail("no format line");
    if (!seen_vertex) return fail("no vertex element");
    p->data_off = pos;
    p->row = off;
    if (p->row == 0 && p->P > 0) return fail("vertex element has no properties");
    // (no product: a huge count from an untrusted header must not wrap size_t)
    if (p->data_off > p->map_len || (p->row && (uint64_t)p->P > (uint64_t)(p->map_len - p->data_off) / (uint64_t)p->row))
        return fail("file is truncated");

    // ---- columns (names of construct_list_of_attributes, gaussian_model.py:293-312) ----
    const char* xyzn[3] = {"x", "y", "z"};
    for (int k = 0; k < 3; ++k) if ((p->xyz[k] = find(*p, xyzn[k])) < 0) return fail("missing x / y / z");
    p->fdc = find_seq(*p, "f_dc_");
    p->frest = find_seq(*p, "f_rest_");
    p->scale = find_seq(*p, "scale_");
    p->rot = find_seq(*p, "rot_");
    p->occ = find_seq(*p, "occ_multiplier_");
    for (int i = 0;; ++i) {                                   // dc_delta_{i}_{j}, j = 0..2
        int k[3];
        bool all = true;
        for (int j = 0; j < 3; ++j) all = all && (k[j] = find(*p, "dc_delta_" + std::to_string(i) + "_" + std::to_string(j))) >= 0;
        if (!all) break;
        for (int j = 0; j < 3; ++j) p->dcd.push_back(k[j]);
    }
    p->opacity = find(*p, "opacity");
    p->mask = find(*p, "base_gaussian_mask");
    p->maxpx = find(*p, "max_pixel_sizes");
    p->minpx = find(*p, "min_pixel_sizes");
    if (p->fdc.size() != 3) return fail("need f_dc_0..2");
    if (p->frest.size() % 3 != 0) return fail("number of f_rest_* columns is not a multiple of 3");
    if (p->scale.size() != 3 || p->rot.size() != 4 || p->opacity < 0) return fail("need opacity, scale_0..2, rot_0..3");
    p->M = 1 + (int)p->frest.size() / 3;

    info->num_points = p->P;
    info->sh_coeffs = p->M;
    info->n_lvl_occ = (int)p->occ.size();
    info->n_lvl_dc = (int)p->dcd.size() / 3;
    info->has_ms_attributes = (p->mask >= 0 && p->maxpx >= 0 && p->minpx >= 0) ? 1 : 0;
    info->row_bytes = (int64_t)p->row;
    info->data_bytes = (int64_t)p->row * p->P;
    return p;
}

int msgs_ply_read(void* handle, float* means3D, float* shs, float* opacity, float* scales, float* rotations,
                  float* occ_multiplier, float* dc_delta, uint8_t* base_mask, float* max_pixel_sizes,
                  float* min_pixel_sizes, int threads) {
    Ply* p = (Ply*)handle;
    if (!p) { g_ply_err = "msgs_ply_read: null handle"; return MSGS_PLY_ERR_ARG; }
    const int64_t P = p->P;
    if (P == 0) return MSGS_PLY_OK;
    if (threads <= 0) threads = (int)std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
    threads = (int)std::min<int64_t>(threads, std::max<int64_t>(1, P / 4096));
    const int M = p->M, L = (int)p->occ.size(), D3 = (int)p->dcd.size();
    const unsigned char* base = p->map + p->data_off;
    const size_t row = p->row;
    // column tables resolved once: (byte offset in the row, type) per output element, in output order
    struct Col { uint32_t off; PType type; };
    auto cols_of = [&](const std::vector<int>& idx) {
        std::vector<Col> c;
        for (int k : idx) c.push_back({(uint32_t)p->props[(size_t)k].offset, p->props[(size_t)k].type});
        return c;
    };
    std::vector<int> sh_idx;                                   // (k, channel) order of the (P,M,3) tensor
    for (int ch = 0; ch < 3; ++ch) sh_idx.push_back(p->fdc[ch]);
    for (int k = 1; k < M; ++k)
        for (int ch = 0; ch < 3; ++ch) sh_idx.push_back(p->frest[(size_t)ch * (M - 1) + (k - 1)]);
    const std::vector<Col> c_xyz = cols_of({p->xyz[0], p->xyz[1], p->xyz[2]}), c_sh = cols_of(sh_idx),
                           c_op = cols_of({p->opacity}), c_sc = cols_of(p->scale), c_rot = cols_of(p->rot),
                           c_occ = cols_of(p->occ), c_dcd = cols_of(p->dcd);
    std::vector<Col> c_tail;                                   // mask, max, min (absent in a plain 3DGS file)
    if (p->mask >= 0) c_tail = cols_of({p->mask});
    const std::vector<Col> c_max = p->maxpx >= 0 ? cols_of({p->maxpx}) : std::vector<Col>(),
                           c_min = p->minpx >= 0 ? cols_of({p->minpx}) : std::vector<Col>();
    bool all_f32 = true;                                       // save_ply writes float32 columns: plain 4-byte moves
    for (const auto* v : {&c_xyz, &c_sh, &c_op, &c_sc, &c_rot, &c_occ, &c_dcd, &c_max, &c_min})
        for (const Col& c : *v) all_f32 = all_f32 && c.type == T_F32;

    auto gather = [&](const unsigned char* r, const std::vector<Col>& c, float* out, bool fast) {
        const size_t n = c.size();
        if (fast) for (size_t k = 0; k < n; ++k) memcpy(out + k, r + c[k].off, 4);
        else for (size_t k = 0; k < n; ++k) out[k] = load_as_float(r + c[k].off, c[k].type);
    };
    auto work = [&](int64_t i0, int64_t i1) {
        const bool fast = all_f32;
        for (int64_t i = i0; i < i1; ++i) {
            const unsigned char* r = base + (size_t)i * row;
            if (means3D) gather(r, c_xyz, means3D + 3 * i, fast);
            if (shs) gather(r, c_sh, shs + (size_t)i * M * 3, fast);
            if (opacity) gather(r, c_op, opacity + i, fast);
            if (scales) gather(r, c_sc, scales + 3 * i, fast);
            if (rotations) gather(r, c_rot, rotations + 4 * i, fast);
            if (occ_multiplier && L) gather(r, c_occ, occ_multiplier + (size_t)i * L, fast);
            if (dc_delta && D3) gather(r, c_dcd, dc_delta + (size_t)i * D3, fast);
            if (base_mask) base_mask[i] = c_tail.empty() ? (uint8_t)0 : (uint8_t)(load_as_float(r + c_tail[0].off, c_tail[0].type) != 0.0f);
            if (max_pixel_sizes) { if (c_max.empty()) max_pixel_sizes[i] = -1.0f; else gather(r, c_max, max_pixel_sizes + i, fast); }
            if (min_pixel_sizes) { if (c_min.empty()) min_pixel_sizes[i] = -1.0f; else gather(r, c_min, min_pixel_sizes + i, fast); }
        }
    };
    if (threads == 1) { work(0, P); return MSGS_PLY_OK; }
    std::vector<std::thread> pool;
    const int64_t per = (P + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
        const int64_t a = t * per, b = std::min(P, a + per);
        if (a < b) pool.emplace_back(work, a, b);
    }
    for (auto& th : pool) th.join();
    return MSGS_PLY_OK;
}

}  // extern "C"
