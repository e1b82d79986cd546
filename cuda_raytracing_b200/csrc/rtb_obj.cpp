// rtb_obj.cpp -- Wavefront OBJ ingestion for the drop-in OBJLoader::load (reference: OBJLoader.hpp:15-179).
// Host I/O, not a kernel.  Same accepted grammar and the same output triangles as the reference loader:
//   `v x y z`, `vt u v`  (vn and every other record ignored, 55-62), faces `f` with tokens `v`, `v/vt`,
//   `v/vt/vn`; polygons fan-triangulated 0,i,i+1 (140); all vertices are collected before any face is
//   resolved (two passes, 36-77); face normal = normalize(cross(v_i - v_0, v_{i+1} - v_0)) with the
//   Q_rsqrt normalize (143-147).  Faces without vt get zero UVs (the reference leaves them uninitialised).
// Whole-file read + pointer scanning instead of getline/istringstream/vector<string> per line.
#include "rtb_hostmath.h"
#include <cstdio>
#include <cstdlib>
#include <vector>

using namespace rtb;

namespace {
struct Cursor {
    const char* p; const char* end;
    bool eof() const { return p >= end; }
    void skip_blank() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\r' || *p == '\v' || *p == '\f')) p++; }
    void next_line() { while (p < end && *p != '\n') p++; if (p < end) p++; }
    bool at_eol() const { return p >= end || *p == '\n'; }
};
inline bool is_tag(const Cursor& c, const char* tag, int len)
{
    if (c.end - c.p < len) return false;
    for (int i = 0; i < len; i++) if (c.p[i] != tag[i]) return false;
    if (c.p + len >= c.end) return true;
    char t = c.p[len];
    return t == ' ' || t == '\t' || t == '\r' || t == '\n' || t == '\v' || t == '\f';
}
}  // namespace

extern "C" int rtb_obj_load(const char* path, rtb_triangle** tris_out, int* n_out)
{
    if (!path || !tris_out || !n_out) return RTB_ERR_INVALID;
    FILE* f = fopen(path, "rb");
    if (!f) return RTB_ERR_INVALID;
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    std::vector<char> buf((size_t)sz + 1);
    size_t got = sz > 0 ? fread(buf.data(), 1, (size_t)sz, f) : 0;
    fclose(f);
    buf[got] = 0;

    std::vector<hm::V3> pos;
    std::vector<float> tex;            // u,v pairs
    Cursor c{ buf.data(), buf.data() + got };
    while (!c.eof()) {                 // pass 1: positions and texture coordinates
        c.skip_blank();
        if (is_tag(c, "v", 1)) {
            char* e = const_cast<char*>(c.p + 1);
            float x = strtof(e, &e), y = strtof(e, &e), z = strtof(e, &e);
            pos.push_back(hm::V3{ x, y, z });
        } else if (is_tag(c, "vt", 2)) {
            char* e = const_cast<char*>(c.p + 2);
            float u = strtof(e, &e), v = strtof(e, &e);
            tex.push_back(u); tex.push_back(v);
        }
        c.next_line();
    }
    std::vector<rtb_triangle> out;
    std::vector<long> vi, ti;
    c = Cursor{ buf.data(), buf.data() + got };
    while (!c.eof()) {                 // pass 2: faces
        c.skip_blank();
        if (is_tag(c, "f", 1)) {
            c.p += 1;
            vi.clear(); ti.clear();
            for (;;) {
                c.skip_blank();
                if (c.at_eol()) break;
                char* e = const_cast<char*>(c.p);
                long v = strtol(e, &e, 10);
                if (e == c.p) break;                                    // not a number: stop this face
                vi.push_back(v - 1);
                if (*e == '/') {                                        // OBJLoader.hpp:106-110: text after the first slash
                    char* e2 = e + 1;
                    long t = strtol(e2, &e2, 10);
                    ti.push_back(t - 1);
                    e = e2;
                    if (*e == '/') { e++; strtol(e, &e, 10); }          // vn index parsed and ignored (113-117)
                }
                c.p = e;
                while (!c.at_eol() && *c.p != ' ' && *c.p != '\t' && *c.p != '\r') c.p++;
            }
            for (size_t i = 1; i + 1 < vi.size(); i++) {
                rtb_triangle t;
                memset(&t, 0, sizeof t);
                const hm::V3 a = pos[(size_t)vi[0]], b = pos[(size_t)vi[i]], d = pos[(size_t)vi[i + 1]];
                t.vertices[0][0] = a.x; t.vertices[0][1] = a.y; t.vertices[0][2] = a.z;
                t.vertices[1][0] = b.x; t.vertices[1][1] = b.y; t.vertices[1][2] = b.z;
                t.vertices[2][0] = d.x; t.vertices[2][1] = d.y; t.vertices[2][2] = d.z;
                const hm::V3 n = hm::face_normal(a, b, d);
                t.normal[0] = n.x; t.normal[1] = n.y; t.normal[2] = n.z;
                if (!ti.empty()) {
                    const long k[3] = { ti[0], ti[i], ti[i + 1] };
                    for (int q = 0; q < 3; q++) { t.uv_coords[q][0] = tex[2 * (size_t)k[q]]; t.uv_coords[q][1] = tex[2 * (size_t)k[q] + 1]; }
                }
                out.push_back(t);
            }
        }
        c.next_line();
    }
    rtb_triangle* res = (rtb_triangle*)malloc(sizeof(rtb_triangle) * (out.size() ? out.size() : 1));
    if (!res) return RTB_ERR_INVALID;
    if (!out.empty()) memcpy(res, out.data(), sizeof(rtb_triangle) * out.size());
    *tris_out = res;
    *n_out = (int)out.size();
    return RTB_OK;
}
