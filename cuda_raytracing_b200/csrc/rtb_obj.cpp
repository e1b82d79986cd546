// rtb_obj.cpp -- Wavefront OBJ ingestion for the drop-in OBJLoader::load (reference: OBJLoader.hpp:15-179).
// Host I/O, not a kernel.  Same accepted grammar and the same output triangles as the reference loader:
//   `v x y z`, `vt u v`  (vn and every other record ignored, 55-62), faces `f` with tokens `v`, `v/vt`,
//   `v/vt/vn`; polygons fan-triangulated 0,i,i+1 (140); all vertices are collected before any face is
//   resolved (two passes, 36-77); face normal = normalize(cross(v_i - v_0, v_{i+1} - v_0)) with the
//   Q_rsqrt normalize (143-147).  Faces without vt get zero UVs (the reference leaves them uninitialised).
// The reference reads with getline + istringstream + vector<string> per line on one thread (seconds per million faces).
// Here: the file is mapped, cut into one chunk per host thread at line boundaries, every chunk is scanned once with
// pointer arithmetic (numbers by an exact fast path, strtof for anything unusual), and the triangles are written straight
// into the output array at the chunk's prefix-summed offset.  An index that names no vertex fails the load (the reference
// reads out of bounds).
#include "rtb_hostmath.h"
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

using namespace rtb;

namespace {

inline bool is_blank(char ch) { return ch == ' ' || ch == '\t' || ch == '\r' || ch == '\v' || ch == '\f'; }

// Decimal -> binary32 with strtof's result.  Fast path (Clinger): <= 19 significant digits that fit 2^53 and a power of
// ten <= 22 give the correctly rounded double in one multiplication or division; narrowing that double to float is again
// correctly rounded unless the double sits exactly on a float rounding boundary, which is detected and sent, like every
// other unusual spelling (inf, nan, hex, long mantissas, huge exponents), to strtof itself.  `p` stops after the number;
// no number at `p` (end of line, a letter) yields 0 with `p` unchanged, as `iss >> float` leaves 0 behind.
float parse_float(const char*& p, const char* end)
{
    static const double P10[23] = { 1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16,
                                    1e17, 1e18, 1e19, 1e20, 1e21, 1e22 };
    while (p < end && is_blank(*p)) p++;
    const char* s = p;
    bool neg = false;
    if (s < end && (*s == '-' || *s == '+')) { neg = *s == '-'; s++; }
    uint64_t mant = 0;
    int digits = 0, exp10 = 0;
    bool any = false, slow = false;
    while (s < end && *s >= '0' && *s <= '9') { any = true; if (mant || *s != '0') { if (digits < 19) { mant = mant * 10 + (uint64_t)(*s - '0'); digits++; } else slow = true; } s++; }
    if (s < end && *s == '.') {
        s++;
        while (s < end && *s >= '0' && *s <= '9') { any = true; if (mant || *s != '0') { if (digits < 19) { mant = mant * 10 + (uint64_t)(*s - '0'); digits++; exp10--; } else slow = true; } else exp10--; s++; }
    }
    if (!any) {                                         // not a plain decimal: let strtof decide (inf, nan, or no number at all)
        if (s < end && ((*s | 32) == 'i' || (*s | 32) == 'n')) { char* e; float v = strtof(p, &e); p = e; return v; }
        return 0.0f;
    }
    if (s < end && (*s | 32) == 'e') {
        const char* t = s + 1;
        bool eneg = false;
        if (t < end && (*t == '-' || *t == '+')) { eneg = *t == '-'; t++; }
        if (t < end && *t >= '0' && *t <= '9') {
            int ev = 0;
            while (t < end && *t >= '0' && *t <= '9') { if (ev < 10000) ev = ev * 10 + (*t - '0'); t++; }
            exp10 += eneg ? -ev : ev;
            s = t;
        }
    }
    if (s < end && ((*s | 32) == 'x' || (*s | 32) == 'p')) slow = true;     // hex float
    if (!slow && mant <= (1ull << 53) && exp10 >= -22 && exp10 <= 22) {
        double d = (double)mant;
        d = exp10 < 0 ? d / P10[-exp10] : d * P10[exp10];
        uint64_t bits;
        memcpy(&bits, &d, 8);
        if ((bits & 0x1fffffffull) != 0x10000000ull) {                      // not on a float rounding boundary
            p = s;
            const float f = (float)d;
            return neg ? -f : f;
        }
    }
    char* e;
    const float v = strtof(p, &e);
    p = e;
    return v;
}

// strtol(…, 10) on a token: optional sign, digits; no digits -> 0 with `p` unchanged
inline long parse_long(const char*& p, const char* end)
{
    const char* s = p;
    bool neg = false;
    if (s < end && (*s == '-' || *s == '+')) { neg = *s == '-'; s++; }
    if (!(s < end && *s >= '0' && *s <= '9')) return 0;
    long v = 0;
    while (s < end && *s >= '0' && *s <= '9') { v = v * 10 + (*s - '0'); s++; }
    p = s;
    return neg ? -v : v;
}

inline bool is_tag(const char* p, const char* end, const char* tag, int len)
{
    if (end - p < len) return false;
    for (int i = 0; i < len; i++) if (p[i] != tag[i]) return false;
    if (p + len >= end) return true;
    const char t = p[len];
    return is_blank(t) || t == '\n';
}

struct Chunk {
    const char* begin; const char* end;
    std::vector<hm::V3> pos;
    std::vector<float> tex;                 // u, v pairs
    std::vector<int> corners;               // per face corner: vertex index, texture index (-1: none), 0-based
    std::vector<int> face_n;                // corners per face
    size_t n_tris = 0, pos0 = 0, tex0 = 0, tri0 = 0;
    bool bad = false;
};

void scan_chunk(Chunk& c)
{
    const char* p = c.begin; const char* end = c.end;
    while (p < end) {
        while (p < end && is_blank(*p)) p++;
        if (is_tag(p, end, "v", 1)) {
            const char* e = p + 1;
            const float x = parse_float(e, end), y = parse_float(e, end), z = parse_float(e, end);
            c.pos.push_back(hm::V3{ x, y, z });
        } else if (is_tag(p, end, "vt", 2)) {
            const char* e = p + 2;
            const float u = parse_float(e, end), v = parse_float(e, end);
            c.tex.push_back(u); c.tex.push_back(v);
        } else if (is_tag(p, end, "f", 1)) {
            const char* e = p + 1;
            int n = 0;
            for (;;) {
                while (e < end && is_blank(*e)) e++;
                if (e >= end || *e == '\n') break;
                const char* t0 = e;
                const long v = parse_long(e, end);
                if (e == t0) break;                                         // not a number: stop this face
                long t = 0;
                if (e < end && *e == '/') {                                 // OBJLoader.hpp:106-110: text after the first slash
                    e++;
                    t = parse_long(e, end);                                 // `v//vn`: no vt, t stays 0 -> -1 below
                    if (e < end && *e == '/') { e++; parse_long(e, end); }  // vn index parsed and ignored (113-117)
                }
                c.corners.push_back((int)(v - 1)); c.corners.push_back((int)(t - 1));
                n++;
                while (e < end && *e != '\n' && !is_blank(*e)) e++;
            }
            c.face_n.push_back(n);
            if (n >= 3) c.n_tris += (size_t)(n - 2);
        }
        while (p < end && *p != '\n') p++;
        if (p < end) p++;
    }
}

void emit_chunk(Chunk& c, const hm::V3* pos, size_t n_pos, const float* tex, size_t n_tex, rtb_triangle* out)
{
    rtb_triangle* o = out + c.tri0;
    const int* k = c.corners.data();
    for (int n : c.face_n) {
        bool has_t = n > 0;
        for (int i = 0; i < n; i++) {
            if (k[2 * i] < 0 || (size_t)k[2 * i] >= n_pos) { c.bad = true; return; }
            if (k[2 * i + 1] < 0) has_t = false;
            else if ((size_t)k[2 * i + 1] >= n_tex) { c.bad = true; return; }
        }
        for (int i = 1; i + 1 < n; i++) {
            rtb_triangle t;
            memset(&t, 0, sizeof t);
            const hm::V3 a = pos[k[0]], b = pos[k[2 * i]], d = pos[k[2 * i + 2]];
            t.vertices[0][0] = a.x; t.vertices[0][1] = a.y; t.vertices[0][2] = a.z;
            t.vertices[1][0] = b.x; t.vertices[1][1] = b.y; t.vertices[1][2] = b.z;
            t.vertices[2][0] = d.x; t.vertices[2][1] = d.y; t.vertices[2][2] = d.z;
            const hm::V3 nr = hm::face_normal(a, b, d);
            t.normal[0] = nr.x; t.normal[1] = nr.y; t.normal[2] = nr.z;
            if (has_t) {
                const int q[3] = { k[1], k[2 * i + 1], k[2 * i + 3] };
                for (int j = 0; j < 3; j++) { t.uv_coords[j][0] = tex[2 * (size_t)q[j]]; t.uv_coords[j][1] = tex[2 * (size_t)q[j] + 1]; }
            }
            *o++ = t;
        }
        k += 2 * n;
    }
}

template <class F>
void run_parallel(std::vector<Chunk>& chunks, F f)
{
    std::vector<std::thread> th;
    for (size_t i = 1; i < chunks.size(); i++) th.emplace_back([&, i] { f(chunks[i]); });
    f(chunks[0]);
    for (auto& t : th) t.join();
}

}  // namespace

extern "C" int rtb_obj_load(const char* path, rtb_triangle** tris_out, int* n_out)
{
    if (!path || !tris_out || !n_out) return RTB_ERR_INVALID;
    const int fd = open(path, O_RDONLY);
    if (fd < 0) return RTB_ERR_INVALID;
    struct stat sb;
    if (fstat(fd, &sb) != 0 || !S_ISREG(sb.st_mode)) { close(fd); return RTB_ERR_INVALID; }
    const size_t size = (size_t)sb.st_size;
    // one spare zero byte behind the text keeps strtof (the slow path) inside the buffer when the file does not end in '\n'
    std::vector<char> tail;
    const char* data = nullptr;
    void* map = nullptr;
    if (size > 0) {
        map = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
        if (map == MAP_FAILED) { close(fd); return RTB_ERR_INVALID; }
        madvise(map, size, MADV_SEQUENTIAL);
        data = (const char*)map;
    }
    close(fd);
    if (size > 0 && data[size - 1] != '\n' && size % (size_t)sysconf(_SC_PAGESIZE) == 0) {
        tail.assign(data, data + size);                 // the rare file whose last page is full and unterminated: private copy + NUL
        tail.push_back(0);
        munmap(map, size); map = nullptr;
        data = tail.data();
    }

    unsigned hw = std::thread::hardware_concurrency();
    size_t n_chunks = std::max<size_t>(1, std::min<size_t>({ hw ? hw : 1, 32, size / (256 << 10) + 1 }));
    std::vector<Chunk> chunks(n_chunks);
    const char* end = data + size;
    const char* cur = data;
    for (size_t i = 0; i < n_chunks; i++) {
        const char* stop = i + 1 == n_chunks ? end : data + size * (i + 1) / n_chunks;
        if (stop < cur) stop = cur;
        while (stop < end && stop > data && stop[-1] != '\n') stop++;        // cut at a line boundary
        chunks[i].begin = cur; chunks[i].end = stop;
        cur = stop;
    }
    run_parallel(chunks, scan_chunk);

    size_t n_pos = 0, n_tex = 0, n_tris = 0;
    for (auto& c : chunks) { c.pos0 = n_pos; c.tex0 = n_tex; c.tri0 = n_tris; n_pos += c.pos.size(); n_tex += c.tex.size() / 2; n_tris += c.n_tris; }
    std::vector<hm::V3> pos(n_pos);
    std::vector<float> tex(2 * n_tex);
    run_parallel(chunks, [&](Chunk& c) {
        if (!c.pos.empty()) memcpy(pos.data() + c.pos0, c.pos.data(), sizeof(hm::V3) * c.pos.size());
        if (!c.tex.empty()) memcpy(tex.data() + 2 * c.tex0, c.tex.data(), sizeof(float) * c.tex.size());
    });
    if (map) munmap(map, size);
    if (n_tris > (size_t)INT32_MAX) return RTB_ERR_INVALID;
    rtb_triangle* res = (rtb_triangle*)malloc(sizeof(rtb_triangle) * (n_tris ? n_tris : 1));
    if (!res) return RTB_ERR_INVALID;
    run_parallel(chunks, [&](Chunk& c) { emit_chunk(c, pos.data(), n_pos, tex.data(), n_tex, res); });
    for (auto& c : chunks) if (c.bad) { free(res); return RTB_ERR_INVALID; }
    *tris_out = res;
    *n_out = (int)n_tris;
    return RTB_OK;
}
