// rtb_api.cu -- the C-ABI (include/rtb200.h): context, BLAS/TLAS management, scene upload, render.
// Host code only orchestrates; every pixel and every tree node is produced by the CUDA kernels in
// rtb_build.cu / rtb_render.cu.  There is no CPU fallback anywhere in this library.
#include "rtb_build.cuh"
#include "rtb_render.cuh"
#include "rtb_hostmath.h"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

using namespace rtb;

static thread_local std::string g_err;
static int fail(int code, const std::string& msg) { g_err = msg; return code; }
static int cuda_fail(cudaError_t e, const char* where)
{
    g_err = std::string(where) + ": " + cudaGetErrorString(e);
    return RTB_ERR_CUDA;
}
#define CU(x, where) do { cudaError_t _e = (x); if (_e != cudaSuccess) return cuda_fail(_e, where); } while (0)

namespace rtb { void wave_release(void* state); }

struct BlasBack {                   // the buffer an asynchronous rebuild writes while frames still read the front one
    float4* tri_rec = nullptr; float2* uvs = nullptr;
    int tri_cap = 0, n = 0;
    BvhResult bvh;
    cudaEvent_t built = nullptr;    // recorded after the build's last kernel and its counter read-back
    void* counters_h = nullptr;     // pinned
    bool pending = false;           // a build has been enqueued and not committed yet
    float build_ms = 0.f;
};

struct MeshRec {
    bool alive = false;
    int n = 0;
    float4* tri_rec = nullptr; float2* uvs = nullptr;
    int tri_cap = 0;
    BvhResult bvh;
    BlasBack back;                  // allocated by the first rtb_mesh_rebuild_async
    rtb_mesh_info info{};
    // what the TLAS knows about this mesh (see mesh_cut / pack_instance)
    bool dynamic = false;           // rebuilt at least once: its world boxes come from the (slightly inflated) root bounds
    std::vector<float> cut;         // mesh-space boxes (min xyz, max xyz) that together cover every triangle
    bool cut_valid = false;         // the cut covers the current geometry
    int cut_budget = 0;             // box budget the cut was made for
    int cut_serial = 0;             // bumped whenever the cut changes (instances remember the one they were packed against)
};

namespace rtb {
cudaError_t mega_static_setup(int device);                            // rtb_mega.cu: the static scratch of rtb_render_noalloc
size_t mega_static_bytes();
bool mega_static_fits(const Tune& tune, const FrameParams& P, long long* slots, long long* pixels);
cudaError_t launch_mega_static(const Tune& tune, LaunchInfo& info, int device, const SceneDev& sc, const FrameParams& P, unsigned char* img, size_t pitch,
                               const rtb_aovs& aov, Counters* ctr, bool count, cudaStream_t st);
}

// Morton key width: the 32-bit keys (11 + 11 + 10 bits, four radix passes) sort 10 M triangles into a tree the tracer finds as good
// as the one from 39-bit keys (8.32 nodes and 1.50 triangles per C5 ray against 8.32 and 1.48, profiles/r04i_*) and the build is
// 0.17 ms shorter; beyond 2^24 primitives the cells start to fill up and the wide keys take over.
static int auto_key_bits(int n) { return n > (1 << 24) ? 64 : 32; }

struct rtb_context {
    int device = 0;
    std::vector<MeshRec> meshes;
    std::vector<rtb_material> materials_h;
    std::vector<rtb_instance> instances_h;
    rtb_material* materials_d = nullptr; int materials_cap = 0;
    float4* inst_d = nullptr; int inst_cap = 0;
    MeshDev* meshes_d = nullptr; int meshes_cap = 0;
    bool mesh_table_dirty = true;
    BvhResult tlas;
    float4* ibox_min = nullptr; float4* ibox_max = nullptr; int ibox_cap = 0;
    std::vector<float4> inst_rec_h, ibox_lo_h, ibox_hi_h;   // packed instance records and world boxes (host copies)
    std::vector<int> inst_serial;  // cut_serial of the mesh each instance was packed against
    bool tlas_dirty = false;       // a mesh rebuild outgrew the boxes the TLAS was built from: rebuilt before the next frame
    int opt_inst_cut = 256;        // mesh-space boxes per static mesh used to bound its instances (1 = the root box)
    Counters* counters_d = nullptr;
    Arena ws;                      // build workspace (grow-only)
    Arena stage;                   // triangle upload staging + prim boxes
    std::vector<unsigned char*> textures;
    // internal frame for rtb_render_host
    unsigned char* frame_d = nullptr; size_t frame_pitch = 0; int frame_w = 0, frame_h = 0;
    int32_t* aov_prim_d = nullptr; int32_t* aov_inst_d = nullptr; float* aov_t_d = nullptr; size_t aov_cap = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t frame_done = nullptr;   // recorded after every rtb_render on its stream: BLAS / TLAS / material updates (stream 0) wait for it
    cudaStream_t build_stream = nullptr;    // asynchronous rebuilds (rtb_mesh_rebuild_async) run here, next to the frame being rendered
    Arena ws_async, stage_async;        // their workspace (the synchronous path keeps its own: the two may overlap in time)
    cudaEvent_t ev_b0 = nullptr;
    float last_ms = 0.f;
    int last_launches = 0;
    int opt_counters = 0, opt_leaf = 8, opt_key_bits = 0;
    int opt_build_fused = 1;       // option "build_fused": fit + chunk-local collapse in one kernel (rtb_build.cu k_fit_collapse)
    rtb::Tune tune;                // kernel tuning knobs of THIS context (rtb_set_option)
    float sah_node_cost = 1.0f;    // option "sah_node" (permille)
    rtb::LaunchInfo last_info;
    void* wf = nullptr;            // wavefront state (rtb_wavefront.cu)
};

extern "C" {

const char* rtb_last_error(void) { return g_err.c_str(); }
int rtb_version(void) { return 100; }

int rtb_context_create(int device, rtb_context** out)
{
    if (!out) return fail(RTB_ERR_INVALID, "rtb_context_create: out is NULL");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0 || device < 0 || device >= count)
        return fail(RTB_ERR_NO_DEVICE, std::string("rtb_context_create: no usable CUDA device (") +
                                           (e != cudaSuccess ? cudaGetErrorString(e) : "index out of range") + "); there is no CPU fallback");
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device), "cudaGetDeviceProperties");
    if (prop.major != 10)
        return fail(RTB_ERR_NO_DEVICE, "rtb_context_create: device is not sm_100 (kernels are built for sm_100a only)");
    CU(cudaSetDevice(device), "cudaSetDevice");
    rtb_context* c = new rtb_context();
    c->device = device;
    CU(cudaMalloc((void**)&c->counters_d, sizeof(Counters)), "cudaMalloc counters");
    CU(cudaMemset(c->counters_d, 0, sizeof(Counters)), "cudaMemset counters");
    CU(cudaEventCreate(&c->ev0), "cudaEventCreate");
    CU(cudaEventCreate(&c->ev1), "cudaEventCreate");
    CU(cudaEventCreateWithFlags(&c->frame_done, cudaEventDisableTiming), "cudaEventCreate");
    rtb::mega_static_setup(device);     // address of rtb_render_noalloc's static scratch on this device; a failure is reported by that call
    *out = c;
    return RTB_OK;
}

static void free_mesh(MeshRec& m)
{
    if (m.back.tri_rec) cudaFree(m.back.tri_rec);
    if (m.back.uvs) cudaFree(m.back.uvs);
    if (m.back.bvh.nodes) cudaFree(m.back.bvh.nodes);
    if (m.back.bvh.order) cudaFree(m.back.bvh.order);
    if (m.back.built) cudaEventDestroy(m.back.built);
    if (m.back.counters_h) cudaFreeHost(m.back.counters_h);
    if (m.tri_rec) cudaFree(m.tri_rec);
    if (m.uvs) cudaFree(m.uvs);
    if (m.bvh.nodes) cudaFree(m.bvh.nodes);
    if (m.bvh.order) cudaFree(m.bvh.order);
    m = MeshRec();
}


void rtb_context_destroy(rtb_context* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    for (auto& m : c->meshes) if (m.alive) free_mesh(m);
    for (auto* t : c->textures) cudaFree(t);
    if (c->materials_d) cudaFree(c->materials_d);
    if (c->inst_d) cudaFree(c->inst_d);
    if (c->meshes_d) cudaFree(c->meshes_d);
    if (c->tlas.nodes) cudaFree(c->tlas.nodes);
    if (c->tlas.order) cudaFree(c->tlas.order);
    if (c->ibox_min) cudaFree(c->ibox_min);
    if (c->ibox_max) cudaFree(c->ibox_max);
    if (c->counters_d) cudaFree(c->counters_d);
    if (c->frame_d) cudaFree(c->frame_d);
    if (c->aov_prim_d) cudaFree(c->aov_prim_d);
    if (c->aov_inst_d) cudaFree(c->aov_inst_d);
    if (c->aov_t_d) cudaFree(c->aov_t_d);
    wave_release(c->wf);
    c->ws.release(); c->stage.release(); c->ws_async.release(); c->stage_async.release();
    if (c->build_stream) cudaStreamDestroy(c->build_stream);
    if (c->ev_b0) cudaEventDestroy(c->ev_b0);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->frame_done) cudaEventDestroy(c->frame_done);
    delete c;
}

int rtb_set_option(rtb_context* c, const char* name, int value)
{
    if (!c || !name) return fail(RTB_ERR_INVALID, "rtb_set_option: NULL argument");
    std::string n(name);
    if (n == "counters") c->opt_counters = value;
    else if (n == "leaf_size") c->opt_leaf = std::max(1, std::min(value, (int)MAX_LEAF));
    else if (n == "key_bits") c->opt_key_bits = value;
    else if (n == "sah_node") c->sah_node_cost = 0.001f * (float)std::max(1, std::min(value, 100000));
    else if (n == "build_fused") c->opt_build_fused = value != 0;
    else if (n == "inst_cut") { c->opt_inst_cut = std::max(1, std::min(value, 1024)); c->tlas_dirty = !c->instances_h.empty(); }
    else if (n == "fused") c->tune.fused = value;
    else if (n == "cost_order") c->tune.cost_order = value;
    else if (n == "ring_split") c->tune.ring_split = value ? 1 : 0;
    else if (n == "cost_block") c->tune.cost_block = std::max(0, std::min(value, 3));
    else if (n == "queue_keep") c->tune.queue_keep = std::max(1, std::min(value, 33));
    else if (n == "tail_idle") c->tune.tail_idle = std::max(1, std::min(value, 32));
    else if (n == "spec_thr") c->tune.spec_thr = std::max(0, std::min(value, 31));
    else if (n == "smem_top") c->tune.smem_top = std::max(0, std::min(value, 341));
    else if (n == "trace_minb") c->tune.trace_minb = value;
    else if (n == "direct_minb") c->tune.direct_minb = value;
    else if (n == "refill_idle") c->tune.refill_idle = std::max(1, std::min(value, 32));
    else return fail(RTB_ERR_INVALID, "rtb_set_option: unknown option " + n);
    return RTB_OK;
}

// ------------------------------------------------------------------ BLAS
static int build_mesh(rtb_context* c, MeshRec& m, const rtb_triangle* tris, int n, int on_device)
{
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    cudaStream_t st = 0;
    CU(cudaStreamWaitEvent(st, c->frame_done, 0), "wait for the frame in flight");   // a frame on a non-blocking stream may still read this BLAS
    const size_t nn = (size_t)(n > 0 ? n : 1);
    c->stage.reset();
    CU(c->stage.reserve(nn * (on_device ? 32 : 32 + sizeof(rtb_triangle)) + 4096), "stage reserve");
    float4* bmin = c->stage.take<float4>(nn);
    float4* bmax = c->stage.take<float4>(nn);
    const rtb_triangle* tris_d = tris;
    if (!on_device) {
        rtb_triangle* up = c->stage.take<rtb_triangle>(nn);
        if (n > 0) CU(cudaMemcpyAsync(up, tris, sizeof(rtb_triangle) * (size_t)n, cudaMemcpyHostToDevice, st), "upload triangles");
        tris_d = up;
    }
    if (m.tri_cap < n || !m.tri_rec) {
        if (m.tri_rec) cudaFree(m.tri_rec);
        if (m.uvs) cudaFree(m.uvs);
        m.tri_rec = nullptr; m.uvs = nullptr;
        CU(cudaMalloc((void**)&m.tri_rec, sizeof(float4) * TRI_F4 * nn), "cudaMalloc triangle records");
        CU(cudaMalloc((void**)&m.uvs, sizeof(float2) * 3 * nn), "cudaMalloc uvs");
        m.tri_cap = (int)nn;
    }
    CU(cudaEventRecord(c->ev0, st), "event");
    int key_bits = c->opt_key_bits ? c->opt_key_bits : auto_key_bits(n);
    CU(build_bvh(tris_d, bmin, bmax, n, c->opt_leaf, key_bits, c->sah_node_cost, m.bvh, c->ws, st, m.tri_rec, m.uvs, c->ev1, nullptr, c->opt_build_fused), "build_bvh");
    CU(cudaEventSynchronize(c->ev1), "build sync");
    float ms = 0.f;
    cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    m.n = n; m.alive = true;
    rtb_mesh_info& I = m.info;
    I.num_triangles = n; I.num_nodes = m.bvh.num_nodes; I.num_leaves = m.bvh.num_leaves; I.max_leaf_size = m.bvh.max_leaf;
    I.depth = m.bvh.depth; I.stack_need = 3 * m.bvh.depth + 2; I.build_ms = ms; I.sah_cost = m.bvh.sah;
    for (int k = 0; k < 3; k++) { I.bounds_min[k] = m.bvh.bounds[k]; I.bounds_max[k] = m.bvh.bounds[k + 3]; }
    I.device_bytes = sizeof(float4) * TRI_F4 * nn + sizeof(float2) * 3 * nn + sizeof(float4) * NODE_F4 * (size_t)m.bvh.node_cap +
                     sizeof(unsigned) * (size_t)m.bvh.order_cap;
    c->mesh_table_dirty = true;
    return RTB_OK;
}

static void after_rebuild(rtb_context* c, MeshRec& m, int id, bool had_dynamic_box, int n);

int rtb_mesh_create(rtb_context* c, const rtb_triangle* tris, int n, int on_device, int* mesh_id)
{
    if (!c || n < 0 || (n > 0 && !tris) || !mesh_id) return fail(RTB_ERR_INVALID, "rtb_mesh_create: bad argument");
    c->meshes.emplace_back();
    int rc = build_mesh(c, c->meshes.back(), tris, n, on_device);
    if (rc != RTB_OK) { c->meshes.pop_back(); return rc; }
    *mesh_id = (int)c->meshes.size() - 1;
    return RTB_OK;
}

int rtb_mesh_rebuild(rtb_context* c, int id, const rtb_triangle* tris, int n, int on_device)
{
    if (!c || id < 0 || id >= (int)c->meshes.size() || !c->meshes[id].alive || n < 0 || (n > 0 && !tris))
        return fail(RTB_ERR_INVALID, "rtb_mesh_rebuild: bad argument");
    MeshRec& m = c->meshes[id];
    const bool had_dynamic_box = m.dynamic && m.cut_valid && m.cut.size() == 6;
    m.cut_valid = false;
    int rc = build_mesh(c, m, tris, n, on_device);
    if (rc != RTB_OK) return rc;
    after_rebuild(c, m, id, had_dynamic_box, n);
    return RTB_OK;
}

// The instance world boxes in the TLAS were made from the OLD geometry.  A deforming mesh keeps them as long as its new bounds
// stay inside the (inflated) box they were made from -- no TLAS work in the steady state of an animation -- otherwise the TLAS
// is rebuilt before the next frame (make_scene).
static void after_rebuild(rtb_context* c, MeshRec& m, int id, bool had_dynamic_box, int n)
{
    m.dynamic = true;
    bool covered = had_dynamic_box && n > 0;
    for (int k = 0; k < 3 && covered; k++) covered = m.bvh.bounds[k] >= m.cut[k] && m.bvh.bounds[k + 3] <= m.cut[k + 3];
    if (covered) m.cut_valid = true;
    else for (const auto& in : c->instances_h) if (in.mesh_index == id) { c->tlas_dirty = true; break; }
}

int rtb_mesh_rebuild_async(rtb_context* c, int id, const rtb_triangle* tris, int n, int on_device)
{
    if (!c || id < 0 || id >= (int)c->meshes.size() || !c->meshes[id].alive || n < 0 || (n > 0 && !tris))
        return fail(RTB_ERR_INVALID, "rtb_mesh_rebuild_async: bad argument");
    if (!on_device) return fail(RTB_ERR_INVALID, "rtb_mesh_rebuild_async: the triangles must be resident on the device (a pageable host copy would synchronise)");
    MeshRec& m = c->meshes[id];
    BlasBack& b = m.back;
    if (b.pending) return fail(RTB_ERR_STATE, "rtb_mesh_rebuild_async: the previous asynchronous rebuild of this mesh was not committed");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    if (!c->build_stream) {
        CU(cudaStreamCreateWithFlags(&c->build_stream, cudaStreamNonBlocking), "build stream");
        CU(cudaEventCreate(&c->ev_b0), "cudaEventCreate");
    }
    if (!b.built) {
        CU(cudaEventCreate(&b.built), "cudaEventCreate");
        CU(cudaMallocHost(&b.counters_h, build_counters_bytes()), "cudaMallocHost");
    }
    cudaStream_t st = c->build_stream;
    const size_t nn = (size_t)(n > 0 ? n : 1);
    if (b.tri_cap < n || !b.tri_rec) {
        if (b.tri_rec) cudaFree(b.tri_rec);
        if (b.uvs) cudaFree(b.uvs);
        b.tri_rec = nullptr; b.uvs = nullptr;
        CU(cudaMalloc((void**)&b.tri_rec, sizeof(float4) * TRI_F4 * nn), "cudaMalloc triangle records");
        CU(cudaMalloc((void**)&b.uvs, sizeof(float2) * 3 * nn), "cudaMalloc uvs");
        b.tri_cap = (int)nn;
    }
    // (this buffer was the front buffer until the last commit: that commit made the build stream wait for its last reader)
    c->stage_async.reset();
    CU(c->stage_async.reserve(nn * 32 + 4096), "stage reserve");
    float4* bmin = c->stage_async.take<float4>(nn);
    float4* bmax = c->stage_async.take<float4>(nn);
    CU(cudaEventRecord(c->ev_b0, st), "event");
    const int key_bits = c->opt_key_bits ? c->opt_key_bits : auto_key_bits(n);
    CU(build_bvh(tris, bmin, bmax, n, c->opt_leaf, key_bits, c->sah_node_cost, b.bvh, c->ws_async, st, b.tri_rec, b.uvs, b.built, b.counters_h, c->opt_build_fused), "build_bvh");
    b.n = n; b.pending = true;
    return RTB_OK;
}

int rtb_mesh_commit(rtb_context* c, int id)
{
    if (!c || id < 0 || id >= (int)c->meshes.size() || !c->meshes[id].alive) return fail(RTB_ERR_INVALID, "rtb_mesh_commit: bad argument");
    MeshRec& m = c->meshes[id];
    BlasBack& b = m.back;
    if (!b.pending) return fail(RTB_ERR_STATE, "rtb_mesh_commit: no asynchronous rebuild pending for this mesh");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    CU(cudaEventSynchronize(b.built), "build sync");          // the HOST waits for the build only; frames in flight keep running
    b.pending = false;
    CU(build_bvh_finish(b.bvh, b.counters_h), "build_bvh (asynchronous)");
    cudaEventElapsedTime(&b.build_ms, c->ev_b0, b.built);
    // The frames enqueued so far traverse the old front buffer, which becomes the target of the NEXT asynchronous build: the
    // build stream waits here for the last of them (frame_done as recorded by that frame); later frames read the new buffer.
    CU(cudaStreamWaitEvent(c->build_stream, c->frame_done, 0), "order the next build after the frames in flight");
    const bool had_dynamic_box = m.dynamic && m.cut_valid && m.cut.size() == 6;
    m.cut_valid = false;
    std::swap(m.tri_rec, b.tri_rec); std::swap(m.uvs, b.uvs); std::swap(m.tri_cap, b.tri_cap); std::swap(m.n, b.n);
    std::swap(m.bvh, b.bvh);
    const size_t nn = (size_t)(m.n > 0 ? m.n : 1);
    rtb_mesh_info& I = m.info;
    I.num_triangles = m.n; I.num_nodes = m.bvh.num_nodes; I.num_leaves = m.bvh.num_leaves; I.max_leaf_size = m.bvh.max_leaf;
    I.depth = m.bvh.depth; I.stack_need = 3 * m.bvh.depth + 2; I.build_ms = b.build_ms; I.sah_cost = m.bvh.sah;
    for (int k = 0; k < 3; k++) { I.bounds_min[k] = m.bvh.bounds[k]; I.bounds_max[k] = m.bvh.bounds[k + 3]; }
    I.device_bytes = sizeof(float4) * TRI_F4 * nn + sizeof(float2) * 3 * nn + sizeof(float4) * NODE_F4 * (size_t)m.bvh.node_cap +
                     sizeof(unsigned) * (size_t)m.bvh.order_cap;
    c->mesh_table_dirty = true;
    after_rebuild(c, m, id, had_dynamic_box, m.n);
    return RTB_OK;
}

int rtb_mesh_info_get(rtb_context* c, int id, rtb_mesh_info* out)
{
    if (!c || !out || id < 0 || id >= (int)c->meshes.size() || !c->meshes[id].alive)
        return fail(RTB_ERR_INVALID, "rtb_mesh_info_get: bad argument");
    *out = c->meshes[id].info;
    return RTB_OK;
}

int rtb_mesh_read(rtb_context* c, int id, int what, void* out, size_t bytes)
{
    if (!c || !out || id < 0 || id >= (int)c->meshes.size() || !c->meshes[id].alive) return fail(RTB_ERR_INVALID, "rtb_mesh_read: bad argument");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    const MeshRec& m = c->meshes[id];
    const void* src = nullptr; size_t have = 0;
    if (what == 0) { src = m.bvh.nodes; have = sizeof(float4) * NODE_F4 * (size_t)m.bvh.num_nodes; }
    else if (what == 1) { src = m.tri_rec; have = sizeof(float4) * TRI_F4 * (size_t)m.n; }
    else if (what == 2) { src = m.bvh.order; have = sizeof(unsigned) * (size_t)m.n; }
    else return fail(RTB_ERR_INVALID, "rtb_mesh_read: unknown array");
    if (bytes < have) return fail(RTB_ERR_INVALID, "rtb_mesh_read: output buffer too small");
    if (have) CU(cudaMemcpy(out, src, have, cudaMemcpyDeviceToHost), "rtb_mesh_read copy");
    return RTB_OK;
}

int rtb_mesh_destroy(rtb_context* c, int id)
{
    if (!c || id < 0 || id >= (int)c->meshes.size() || !c->meshes[id].alive) return fail(RTB_ERR_INVALID, "rtb_mesh_destroy: bad argument");
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    free_mesh(c->meshes[id]);
    c->mesh_table_dirty = true;
    return RTB_OK;
}

}  // extern "C"

// ------------------------------------------------------------------ scene
namespace rtb { cudaError_t launch_wavefront(const Tune& tune, LaunchInfo& info, void** wf_state, const SceneDev& sc, const FrameParams& P, unsigned char* img,
                                             size_t pitch, const rtb_aovs& aov, Counters* ctr, bool count, cudaStream_t st); }

extern "C" {

int rtb_texture_upload(rtb_context* c, const unsigned char* bgr, int w, int h, unsigned char** dev, size_t* pitch)
{
    if (!c || !bgr || w <= 0 || h <= 0 || !dev || !pitch) return fail(RTB_ERR_INVALID, "rtb_texture_upload: bad argument");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    unsigned char* d = nullptr;
    CU(cudaMallocPitch((void**)&d, pitch, (size_t)w * 3, (size_t)h), "cudaMallocPitch texture");   // as Material.hpp:37
    CU(cudaMemcpy2D(d, *pitch, bgr, (size_t)w * 3, (size_t)w * 3, (size_t)h, cudaMemcpyHostToDevice), "texture upload");
    c->textures.push_back(d);
    *dev = d;
    return RTB_OK;
}

int rtb_set_materials(rtb_context* c, const rtb_material* mats, int n)
{
    if (!c || n < 0 || (n > 0 && !mats)) return fail(RTB_ERR_INVALID, "rtb_set_materials: bad argument");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    CU(cudaStreamWaitEvent(0, c->frame_done, 0), "wait for the frame in flight");
    c->materials_h.assign(mats, mats + n);
    if (c->materials_cap < n || !c->materials_d) {
        if (c->materials_d) cudaFree(c->materials_d);
        c->materials_d = nullptr;
        CU(cudaMalloc((void**)&c->materials_d, sizeof(rtb_material) * (size_t)std::max(n, 1)), "cudaMalloc materials");
        c->materials_cap = std::max(n, 1);
    }
    if (n > 0) CU(cudaMemcpy(c->materials_d, mats, sizeof(rtb_material) * (size_t)n, cudaMemcpyHostToDevice), "materials upload");
    return RTB_OK;
}

// Mesh-space boxes that together cover every triangle of a mesh: the child boxes of a cut through the top levels of its
// BLAS (read back once per build, <= 85 nodes), opened breadth-first up to `budget` boxes.  An instance's world box is the
// union of the transformed cut boxes instead of the transformed ROOT box: for a rotated instance the latter is up to
// sqrt(2) too wide per axis, so neighbouring instances overlap and rays enter BLASes they cannot hit.  A mesh that is being
// rebuilt per frame (C5) uses its root bounds, inflated by 2 %, so that no read-back or TLAS rebuild sits in the frame loop.
static int mesh_cut(rtb_context* c, MeshRec& m, int budget)
{
    if (m.cut_valid && (m.dynamic || m.cut_budget == budget)) return RTB_OK;
    m.cut.clear();
    const float* b = m.bvh.bounds;
    bool finite = true;
    for (int k = 0; k < 6; k++) finite = finite && std::isfinite(b[k]);
    if (m.n > 0 && m.dynamic && finite) {
        for (int k = 0; k < 3; k++) {
            const float pad = 0.02f * (b[k + 3] - b[k]) + 1e-6f * std::max(std::fabs(b[k]), std::fabs(b[k + 3]));
            m.cut.push_back(b[k] - pad);
        }
        for (int k = 0; k < 3; k++) m.cut.push_back(b[k + 3] + (b[k] - m.cut[k]));
    } else if (m.n > 0 && finite && budget > 1 && m.bvh.num_nodes >= 1) {
        // the nodes of the top levels are not contiguous (fused build: only node 0 is at a known place), so they are fetched one
        // by one as the cut opens them: at most 85 small copies, once per build of a static mesh
        const int K = m.bvh.num_nodes;
        float4 nd[NODE_F4];
        int fetched = 0;
        cudaError_t fetch_err = cudaSuccess;
        struct Box { float lo[3], hi[3]; int ref; };
        std::vector<Box> open, done;                      // open: FIFO (head index), so the cut is opened level by level
        auto children = [&](int node) {
            cudaError_t e = cudaMemcpy(nd, m.bvh.nodes + (size_t)node * NODE_F4, sizeof nd, cudaMemcpyDeviceToHost);
            if (e != cudaSuccess) { fetch_err = e; return; }
            fetched++;
            const float* f = (const float*)nd;
            for (int k = 0; k < 4; k++) {
                int ref; memcpy(&ref, f + 24 + k, 4);
                if (ref == EMPTY_REF) continue;
                Box x;
                for (int a = 0; a < 3; a++) { x.lo[a] = f[8 * a + k]; x.hi[a] = f[8 * a + 4 + k]; }
                x.ref = ref;
                open.push_back(x);
            }
        };
        children(0);
        size_t head = 0;
        while (head < open.size()) {
            const Box x = open[head++];
            if (x.ref >= 0 && x.ref < K && fetched < 85 && done.size() + (open.size() - head) + 4 <= (size_t)budget) children(x.ref);
            else done.push_back(x);
        }
        CU(fetch_err, "read top BLAS levels");
        for (const Box& x : done)
            for (int a = 0; a < 3; a++) finite = finite && std::isfinite(x.lo[a]) && std::isfinite(x.hi[a]) && x.lo[a] <= x.hi[a];
        if (finite && !done.empty()) {
            for (const Box& x : done) { m.cut.insert(m.cut.end(), x.lo, x.lo + 3); m.cut.insert(m.cut.end(), x.hi, x.hi + 3); }
        }
    }
    if (m.cut.empty() && m.n > 0) m.cut.assign(b, b + 6);              // the root box (tiny meshes, budget 1, non-finite bounds)
    m.cut_valid = true; m.cut_budget = budget; m.cut_serial++;
    return RTB_OK;
}

// instance record (see rtb_internal.cuh) + conservative world box
static int pack_instance(rtb_context* c, const rtb_instance& in, float4 rec[INST_F4], float lo[3], float hi[3], int* serial)
{
    if (in.mesh_index < 0 || in.mesh_index >= (int)c->meshes.size() || !c->meshes[in.mesh_index].alive)
        return fail(RTB_ERR_INVALID, "instance refers to a mesh that does not exist");
    if (in.material_index < 0 || in.material_index >= (int)c->materials_h.size())
        return fail(RTB_ERR_INVALID, "instance refers to a material that does not exist (call rtb_set_materials first)");
    hm::Quat qp = hm::euler_to_quat(in.rotation[0], in.rotation[1], in.rotation[2]);             // raycast.cu:33
    hm::Quat qi = hm::euler_to_quat(in.inv_rotation[0], in.inv_rotation[1], in.inv_rotation[2]); // raycast.cu:115
    // apply_lre(pose, .) rotates by pose.yaw/pitch/roll (transforms.hpp:223-226); MeshInstance::build_inv makes
    // rotation == pose angles, but a caller may have filled the record by hand, so honour each field separately.
    hm::Quat qpl = hm::euler_to_quat(in.pose.yaw, in.pose.pitch, in.pose.roll);
    hm::Quat qil = hm::euler_to_quat(in.inv_pose.yaw, in.inv_pose.pitch, in.inv_pose.roll);
    if (memcmp(&qp, &qpl, sizeof qp) != 0 || memcmp(&qi, &qil, sizeof qi) != 0)
        return fail(RTB_ERR_INVALID, "instance rotation/inv_rotation differ from pose/inv_pose angles (call build_inv)");
    auto asf = [](int i) { float f; memcpy(&f, &i, 4); return f; };
    rec[0] = make_float4(qp.x, qp.y, qp.z, qp.w);
    rec[1] = make_float4(in.pose.x, in.pose.y, in.pose.z, asf(in.mesh_index));
    rec[2] = make_float4(in.inv_scale[0], in.inv_scale[1], in.inv_scale[2], asf(in.material_index));
    rec[3] = make_float4(qi.x, qi.y, qi.z, qi.w);
    rec[4] = make_float4(in.inv_pose.x, in.inv_pose.y, in.inv_pose.z, 0.f);
    rec[5] = make_float4(in.scale[0], in.scale[1], in.scale[2], 0.f);
    // world box: the ray enters mesh space as m = (R (w - t)) * inv_scale, so w = R^T (m / inv_scale) + t = M m + t.  The image
    // of a box (centre c, half extent h) lies in the box (M c + t) +- |M| h.  Double precision + padding: the box only culls,
    // it must merely contain every world point the forward map sends inside.
    MeshRec& mesh = c->meshes[in.mesh_index];
    int rc = mesh_cut(c, mesh, c->opt_inst_cut);
    if (rc != RTB_OK) return rc;
    *serial = mesh.cut_serial;
    float Rf[3][3];
    hm::euler_to_matrix(in.pose.yaw, in.pose.pitch, in.pose.roll, Rf);
    double M[3][3];
    for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) M[a][b] = (double)Rf[b][a] / (double)in.inv_scale[b];
    double wlo[3] = { 1e300, 1e300, 1e300 }, whi[3] = { -1e300, -1e300, -1e300 };
    for (size_t o = 0; o + 6 <= mesh.cut.size(); o += 6) {
        const float* x = &mesh.cut[o];
        double ctr[3], half[3];
        for (int b = 0; b < 3; b++) { ctr[b] = 0.5 * ((double)x[b] + (double)x[b + 3]); half[b] = 0.5 * ((double)x[b + 3] - (double)x[b]); }
        for (int a = 0; a < 3; a++) {
            const double wc = M[a][0] * ctr[0] + M[a][1] * ctr[1] + M[a][2] * ctr[2] + (double)(&in.pose.x)[a];
            const double wh = std::fabs(M[a][0]) * half[0] + std::fabs(M[a][1]) * half[1] + std::fabs(M[a][2]) * half[2];
            wlo[a] = std::min(wlo[a], wc - wh); whi[a] = std::max(whi[a], wc + wh);
        }
    }
    double mag = 0;
    for (int a = 0; a < 3; a++) mag = std::max(mag, std::max(std::fabs(wlo[a]), std::fabs(whi[a])));
    if (mesh.cut.empty()) { for (int a = 0; a < 3; a++) wlo[a] = whi[a] = (&in.pose.x)[a]; mag = 0; }   // empty BLAS: a point
    const double pad = 1e-4 * mag + 1e-5;
    for (int a = 0; a < 3; a++) { lo[a] = (float)(wlo[a] - pad); hi[a] = (float)(whi[a] + pad); }
    return RTB_OK;
}

// Packs the instances whose record or mesh cut changed (all of them when `only` < 0 or the count changed), uploads, and
// rebuilds the TLAS on the device.
static int rebuild_tlas(rtb_context* c, int only)
{
    CU(cudaStreamWaitEvent(0, c->frame_done, 0), "wait for the frame in flight");   // the uploads below are ordered on stream 0
    const int n = (int)c->instances_h.size();
    const int cap = std::max(n, 1);
    if (c->inst_cap < cap) {
        if (c->inst_d) cudaFree(c->inst_d);
        if (c->ibox_min) cudaFree(c->ibox_min);
        if (c->ibox_max) cudaFree(c->ibox_max);
        c->inst_d = c->ibox_min = c->ibox_max = nullptr;
        CU(cudaMalloc((void**)&c->inst_d, sizeof(float4) * INST_F4 * (size_t)cap), "cudaMalloc instances");
        CU(cudaMalloc((void**)&c->ibox_min, sizeof(float4) * (size_t)cap), "cudaMalloc instance boxes");
        CU(cudaMalloc((void**)&c->ibox_max, sizeof(float4) * (size_t)cap), "cudaMalloc instance boxes");
        c->inst_cap = cap;
    }
    const bool all = only < 0 || (int)c->inst_serial.size() != n;
    if (all) {
        c->inst_rec_h.assign((size_t)cap * INST_F4, make_float4(0.f, 0.f, 0.f, 0.f));
        c->ibox_lo_h.assign(cap, make_float4(0.f, 0.f, 0.f, 0.f)); c->ibox_hi_h.assign(cap, make_float4(0.f, 0.f, 0.f, 0.f));
        c->inst_serial.assign(n, -1);
    }
    for (int i = 0; i < n; i++) {
        const int mi = c->instances_h[i].mesh_index;
        const bool fresh = !all && i != only && mi >= 0 && mi < (int)c->meshes.size() && c->meshes[mi].alive && c->meshes[mi].cut_valid &&
                           (c->meshes[mi].dynamic || c->meshes[mi].cut_budget == c->opt_inst_cut) && c->meshes[mi].cut_serial == c->inst_serial[i];
        if (fresh) continue;
        float l[3], h[3];
        int rc = pack_instance(c, c->instances_h[i], &c->inst_rec_h[(size_t)i * INST_F4], l, h, &c->inst_serial[i]);
        if (rc != RTB_OK) { c->inst_serial.clear(); return rc; }
        c->ibox_lo_h[i] = make_float4(l[0], l[1], l[2], 0.f); c->ibox_hi_h[i] = make_float4(h[0], h[1], h[2], 0.f);
    }
    if (n > 0) {
        CU(cudaMemcpy(c->inst_d, c->inst_rec_h.data(), sizeof(float4) * INST_F4 * (size_t)n, cudaMemcpyHostToDevice), "instances upload");
        CU(cudaMemcpy(c->ibox_min, c->ibox_lo_h.data(), sizeof(float4) * (size_t)n, cudaMemcpyHostToDevice), "boxes upload");
        CU(cudaMemcpy(c->ibox_max, c->ibox_hi_h.data(), sizeof(float4) * (size_t)n, cudaMemcpyHostToDevice), "boxes upload");
    }
    CU(build_bvh(nullptr, c->ibox_min, c->ibox_max, n, 1, 32, c->sah_node_cost, c->tlas, c->ws, 0), "build TLAS");
    c->tlas_dirty = false;
    return RTB_OK;
}

int rtb_set_instances(rtb_context* c, const rtb_instance* insts, int n)
{
    if (!c || n < 0 || (n > 0 && !insts)) return fail(RTB_ERR_INVALID, "rtb_set_instances: bad argument");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    c->instances_h.assign(insts, insts + n);
    return rebuild_tlas(c, -1);
}

int rtb_update_instance(rtb_context* c, int index, const rtb_instance* inst)
{
    if (!c || !inst || index < 0 || index >= (int)c->instances_h.size()) return fail(RTB_ERR_INVALID, "rtb_update_instance: bad argument");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    c->instances_h[index] = *inst;
    return rebuild_tlas(c, index);
}

// ------------------------------------------------------------------ render
static float bounds_extent(const float b[6], int n)
{
    float e = 0.f;
    if (n > 0) for (int k = 0; k < 6; k++) e = std::max(e, std::fabs(b[k]));
    return e;
}

static int make_scene(rtb_context* c, SceneDev& sc)
{
    if (!c->tlas.nodes) return fail(RTB_ERR_STATE, "render: no instances set (call rtb_set_instances)");
    for (const auto& in : c->instances_h)
        if (!c->meshes[in.mesh_index].alive) return fail(RTB_ERR_STATE, "render: an instance refers to a destroyed mesh");
    if (c->tlas_dirty) { int rc = rebuild_tlas(c, (int)c->instances_h.size()); if (rc != RTB_OK) return rc; }   // stale boxes only
    // Worst-case traversal stack (rtb_trace_state.cuh pushes up to 3 children per wide-node visit): 3 entries per TLAS level, the
    // sentinel of the instance entered, 3 per level of the deepest BLAS in use.  A scene that needs more than the kernels' fixed
    // stack is refused here with a clear error -- never rendered wrong, and not only reported by the synchronous render path.
    {
        int deepest = 0;
        for (const auto& in : c->instances_h) deepest = std::max(deepest, c->meshes[in.mesh_index].bvh.depth);
        const int need = (c->instances_h.size() > 1 ? 3 * c->tlas.depth + 1 : 0) + 3 * deepest + 2;
        if (need > STACK_CAP)
            return fail(RTB_ERR_OVERFLOW, "render: the scene's trees need a traversal stack of " + std::to_string(need) + " entries, the kernels have " +
                                              std::to_string((int)STACK_CAP) + " (BLAS depth " + std::to_string(deepest) + ", TLAS depth " +
                                              std::to_string(c->tlas.depth) + ": a degenerate mesh -- many primitives with one centroid?)");
    }
    if (c->mesh_table_dirty) {
        const int n = (int)c->meshes.size(), cap = std::max(n, 1);
        if (c->meshes_cap < cap) {
            if (c->meshes_d) cudaFree(c->meshes_d);
            c->meshes_d = nullptr;
            CU(cudaMalloc((void**)&c->meshes_d, sizeof(MeshDev) * (size_t)cap), "cudaMalloc mesh table");
            c->meshes_cap = cap;
        }
        std::vector<MeshDev> t(cap);
        for (int i = 0; i < n; i++) {
            const MeshRec& m = c->meshes[i];
            t[i].nodes = m.bvh.nodes; t[i].tris = m.tri_rec; t[i].uvs = m.uvs; t[i].num_triangles = m.n; t[i].extent = bounds_extent(m.bvh.bounds, m.n);
        }
        CU(cudaMemcpy(c->meshes_d, t.data(), sizeof(MeshDev) * (size_t)cap, cudaMemcpyHostToDevice), "mesh table upload");
        c->mesh_table_dirty = false;
    }
    sc.tlas_nodes = c->tlas.nodes; sc.tlas_index = (const int*)c->tlas.order; sc.inst = c->inst_d; sc.meshes = c->meshes_d;
    sc.materials = c->materials_d; sc.num_instances = (int)c->instances_h.size(); sc.num_materials = (int)c->materials_h.size();
    sc.tlas_extent = bounds_extent(c->tlas.bounds, sc.num_instances);
    // top tree levels to stage in shared memory (rtb_tma.cuh): the TLAS prefix and the prefix of the BLAS most instances use
    sc.top_blas = nullptr; sc.top_tlas_n = sc.top_blas_n = 0;
    sc.spec_thr = c->tune.spec_thr;
    if (c->tune.smem_top > 0 && !c->instances_h.empty()) {
        std::vector<int> uses(c->meshes.size(), 0);
        for (const auto& in : c->instances_h) uses[in.mesh_index]++;
        const int best = (int)(std::max_element(uses.begin(), uses.end()) - uses.begin());
        const MeshRec& m = c->meshes[best];
        if (m.bvh.nodes && m.bvh.num_nodes > 0) { sc.top_blas = m.bvh.nodes; sc.top_blas_n = std::min(m.bvh.num_nodes, c->tune.smem_top); }
        if (sc.num_instances > 1) sc.top_tlas_n = std::min(c->tlas.num_nodes, 48);
    }
    return RTB_OK;
}

static int make_frame(const rtb_render_params* p, FrameParams& P)
{
    if (p->width <= 0 || p->height <= 0) return fail(RTB_ERR_INVALID, "render: width/height must be positive");
    if (p->mode < RTB_MODE_REFERENCE || p->mode > RTB_MODE_PATH) return fail(RTB_ERR_INVALID, "render: unknown mode");
    P.width = p->width; P.height = p->height;
    memcpy(P.K_inv, p->K_inv, sizeof P.K_inv); memcpy(P.D, p->D, sizeof P.D);
    P.ox = p->pose.x; P.oy = p->pose.y; P.oz = p->pose.z;
    hm::Quat q = hm::euler_to_quat(p->inv_pose.yaw, p->inv_pose.pitch, p->inv_pose.roll);       // raycast.cu:185
    P.q_cam = make_float4(q.x, q.y, q.z, q.w);
    P.mode = p->mode; P.max_depth = std::max(0, p->max_depth); P.spp = std::max(1, p->spp); P.seed = p->seed;
    P.tile_count = std::max(1, p->tile_count); P.tile_rank = p->tile_rank;
    if (P.tile_rank < 0 || P.tile_rank >= P.tile_count) return fail(RTB_ERR_INVALID, "render: tile_rank out of range");
    hm::V3 L = hm::normalize(hm::V3{ (float)-0.2, 0.0f, 1.0f });                                 // raycast.cu:249-250
    P.light_dir = make_float3(L.x, L.y, L.z);
    return RTB_OK;
}

static int render_impl(rtb_context* c, const rtb_render_params* p, unsigned char* img, size_t pitch, const rtb_aovs* aovs, cudaStream_t st)
{
    SceneDev sc; FrameParams P;
    int rc = make_scene(c, sc); if (rc != RTB_OK) return rc;
    rc = make_frame(p, P); if (rc != RTB_OK) return rc;
    if (pitch < (size_t)p->width * 3) return fail(RTB_ERR_INVALID, "render: pitch smaller than a row");
    rtb_aovs a = { nullptr, nullptr, nullptr };
    if (aovs) a = *aovs;
    CU(cudaMemsetAsync(c->counters_d, 0, sizeof(Counters), st), "counters reset");
    if (p->mode == RTB_MODE_REFERENCE || p->mode == RTB_MODE_SHADOW)
    {
        CU(launch_render(c->tune, sc, P, img, pitch, a, c->counters_d, c->opt_counters != 0, st), "render launch");
        c->last_info.launches = 1; c->last_info.variant = -1;
    } else {
        CU(launch_wavefront(c->tune, c->last_info, &c->wf, sc, P, img, pitch, a, c->counters_d, c->opt_counters != 0, st), "wavefront launch");
    }
    c->last_launches = c->last_info.launches;
    CU(cudaEventRecord(c->frame_done, st), "frame event");
    return RTB_OK;
}

int rtb_render(rtb_context* c, const rtb_render_params* p, unsigned char* img, size_t pitch, const rtb_aovs* aovs, void* stream)
{
    if (!c || !p || !img) return fail(RTB_ERR_INVALID, "rtb_render: NULL argument");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    return render_impl(c, p, img, pitch, aovs, (cudaStream_t)stream);
}

// No allocation on this path: no cudaMalloc / cudaFree, no table upload, no std::vector -- a scene that still needs any of that
// is refused (RTB_ERR_STATE) and the caller renders one frame through rtb_render first.
int rtb_render_noalloc(rtb_context* c, const rtb_render_params* p, unsigned char* img, size_t pitch, const rtb_aovs* aovs, void* stream)
{
    if (!c || !p || !img) return fail(RTB_ERR_INVALID, "rtb_render_noalloc: NULL argument");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    if (p->mode != RTB_MODE_REFLECT && p->mode != RTB_MODE_PATH)
        return fail(RTB_ERR_INVALID, "rtb_render_noalloc: only RTB_MODE_REFLECT / RTB_MODE_PATH use scratch memory; modes 0 / 1 allocate nothing in rtb_render");
    FrameParams P;
    int rc = make_frame(p, P); if (rc != RTB_OK) return rc;
    if (pitch < (size_t)p->width * 3) return fail(RTB_ERR_INVALID, "rtb_render_noalloc: pitch smaller than a row");
    long long slots = 0, pixels = 0;
    if (!rtb::mega_static_fits(c->tune, P, &slots, &pixels))
        return fail(RTB_ERR_INVALID, "rtb_render_noalloc: frame of " + std::to_string(p->width) + " x " + std::to_string(p->height) + " needs " +
                                         std::to_string(pixels) + " pixels and " + std::to_string(slots) + " pixel slots, the static scratch holds " +
                                         std::to_string((long long)RTB_NOALLOC_MAX_SLOTS) + " (RTB_NOALLOC_MAX_SLOTS); use rtb_render");
    if (!c->tlas.nodes) return fail(RTB_ERR_STATE, "rtb_render_noalloc: no instances set (call rtb_set_instances)");
    if (c->tlas_dirty || c->mesh_table_dirty || !c->meshes_d)
        return fail(RTB_ERR_STATE, "rtb_render_noalloc: the scene changed since the last frame and its device tables are stale; "
                                   "render one frame with rtb_render first (that call may allocate)");
    int deepest = 0;
    for (const auto& in : c->instances_h) {
        if (!c->meshes[in.mesh_index].alive) return fail(RTB_ERR_STATE, "rtb_render_noalloc: an instance refers to a destroyed mesh");
        deepest = std::max(deepest, c->meshes[in.mesh_index].bvh.depth);
    }
    if ((c->instances_h.size() > 1 ? 3 * c->tlas.depth + 1 : 0) + 3 * deepest + 2 > STACK_CAP)
        return fail(RTB_ERR_OVERFLOW, "rtb_render_noalloc: the scene's trees need a deeper traversal stack than the kernels have");
    SceneDev sc;
    sc.tlas_nodes = c->tlas.nodes; sc.tlas_index = (const int*)c->tlas.order; sc.inst = c->inst_d; sc.meshes = c->meshes_d;
    sc.materials = c->materials_d; sc.num_instances = (int)c->instances_h.size(); sc.num_materials = (int)c->materials_h.size();
    sc.tlas_extent = bounds_extent(c->tlas.bounds, sc.num_instances);
    sc.top_blas = nullptr; sc.top_tlas_n = sc.top_blas_n = 0;           // no top levels staged in shared memory (only a placement choice)
    sc.spec_thr = c->tune.spec_thr;
    rtb_aovs a = { nullptr, nullptr, nullptr };
    if (aovs) a = *aovs;
    cudaStream_t st = (cudaStream_t)stream;
    CU(cudaMemsetAsync(c->counters_d, 0, sizeof(Counters), st), "counters reset");
    cudaError_t e = rtb::launch_mega_static(c->tune, c->last_info, c->device, sc, P, img, pitch, a, c->counters_d, c->opt_counters != 0, st);
    if (e != cudaSuccess)
        return fail(RTB_ERR_CUDA, std::string("rtb_render_noalloc: ") + cudaGetErrorString(e) +
                                      (e == cudaErrorNotReady ? " (static scratch address not set up for this device)" : ""));
    c->last_launches = c->last_info.launches;
    CU(cudaEventRecord(c->frame_done, st), "frame event");
    return RTB_OK;
}

int rtb_render_host(rtb_context* c, const rtb_render_params* p, unsigned char* img_host, size_t pitch_host, const rtb_aovs* aovs_host)
{
    if (!c || !p || !img_host) return fail(RTB_ERR_INVALID, "rtb_render_host: NULL argument");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    if (p->width <= 0 || p->height <= 0) return fail(RTB_ERR_INVALID, "render: width/height must be positive");
    if (pitch_host < (size_t)p->width * 3) return fail(RTB_ERR_INVALID, "render: pitch smaller than a row");
    if (c->frame_w != p->width || c->frame_h != p->height) {
        if (c->frame_d) cudaFree(c->frame_d);
        c->frame_d = nullptr; c->frame_w = c->frame_h = 0;
        CU(cudaMallocPitch((void**)&c->frame_d, &c->frame_pitch, (size_t)p->width * 3, (size_t)p->height), "cudaMallocPitch frame");
        c->frame_w = p->width; c->frame_h = p->height;
    }
    const size_t npx = (size_t)p->width * p->height;
    rtb_aovs dev = { nullptr, nullptr, nullptr };
    if (aovs_host && (aovs_host->prim || aovs_host->inst || aovs_host->t)) {
        if (c->aov_cap < npx) {
            if (c->aov_prim_d) cudaFree(c->aov_prim_d);
            if (c->aov_inst_d) cudaFree(c->aov_inst_d);
            if (c->aov_t_d) cudaFree(c->aov_t_d);
            c->aov_prim_d = c->aov_inst_d = nullptr; c->aov_t_d = nullptr; c->aov_cap = 0;
            CU(cudaMalloc((void**)&c->aov_prim_d, 4 * npx), "cudaMalloc aov");
            CU(cudaMalloc((void**)&c->aov_inst_d, 4 * npx), "cudaMalloc aov");
            CU(cudaMalloc((void**)&c->aov_t_d, 4 * npx), "cudaMalloc aov");
            c->aov_cap = npx;
        }
        dev.prim = c->aov_prim_d; dev.inst = c->aov_inst_d; dev.t = c->aov_t_d;
    }
    cudaStream_t st = 0;
    CU(cudaEventRecord(c->ev0, st), "event");
    int rc = render_impl(c, p, c->frame_d, c->frame_pitch, &dev, st);
    if (rc != RTB_OK) return rc;
    CU(cudaEventRecord(c->ev1, st), "event");
    // only the rows this rank rendered are meaningful when tile_count > 1; copy the whole frame anyway
    CU(cudaMemcpy2DAsync(img_host, pitch_host, c->frame_d, c->frame_pitch, (size_t)p->width * 3, (size_t)p->height, cudaMemcpyDeviceToHost, st), "frame download");
    if (dev.prim) {
        if (aovs_host->prim) CU(cudaMemcpyAsync(aovs_host->prim, dev.prim, 4 * npx, cudaMemcpyDeviceToHost, st), "aov download");
        if (aovs_host->inst) CU(cudaMemcpyAsync(aovs_host->inst, dev.inst, 4 * npx, cudaMemcpyDeviceToHost, st), "aov download");
        if (aovs_host->t) CU(cudaMemcpyAsync(aovs_host->t, dev.t, 4 * npx, cudaMemcpyDeviceToHost, st), "aov download");
    }
    CU(cudaStreamSynchronize(st), "render sync");
    cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1);
    Counters h;
    CU(cudaMemcpy(&h, c->counters_d, sizeof h, cudaMemcpyDeviceToHost), "counters");
    if (h.overflow) return fail(RTB_ERR_OVERFLOW, "render: traversal stack / ray queue overflow (result discarded)");
    return RTB_OK;
}

// ------------------------------------------------------------------ multi-GPU frame exchange over peer memory
int rtb_frame_alloc_shared(rtb_context* c, size_t bytes, unsigned char** dev, unsigned char handle[RTB_IPC_HANDLE_BYTES])
{
    static_assert(sizeof(cudaIpcMemHandle_t) == RTB_IPC_HANDLE_BYTES, "handle size");
    if (!c || !dev || !handle || !bytes) return fail(RTB_ERR_INVALID, "rtb_frame_alloc_shared: bad argument");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    void* p = nullptr;
    CU(cudaMalloc(&p, bytes), "cudaMalloc shared frame");
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); return fail(RTB_ERR_CUDA, std::string("cudaIpcGetMemHandle: ") + cudaGetErrorString(e)); }
    CU(cudaMemset(p, 0, bytes), "cudaMemset shared frame");
    memcpy(handle, &h, sizeof h);
    *dev = (unsigned char*)p;
    return RTB_OK;
}

int rtb_frame_free_shared(rtb_context* c, unsigned char* dev)
{
    if (!c) return fail(RTB_ERR_INVALID, "rtb_frame_free_shared: NULL context");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    if (dev) CU(cudaFree(dev), "cudaFree shared frame");
    return RTB_OK;
}

int rtb_frame_open_shared(rtb_context* c, const unsigned char handle[RTB_IPC_HANDLE_BYTES], unsigned char** dev)
{
    if (!c || !dev || !handle) return fail(RTB_ERR_INVALID, "rtb_frame_open_shared: NULL argument");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof h);
    void* p = nullptr;
    CU(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle");
    *dev = (unsigned char*)p;
    return RTB_OK;
}

int rtb_frame_close_shared(rtb_context* c, unsigned char* dev)
{
    if (!c) return fail(RTB_ERR_INVALID, "rtb_frame_close_shared: NULL context");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    if (dev) CU(cudaIpcCloseMemHandle(dev), "cudaIpcCloseMemHandle");
    return RTB_OK;
}

int rtb_push_bands(rtb_context* c, const unsigned char* src, unsigned char* dst, size_t pitch, int height, int tile_rank, int tile_count, void* stream)
{
    if (!c || !src || !dst || !pitch || height <= 0 || tile_count <= 0 || tile_rank < 0 || tile_rank >= tile_count)
        return fail(RTB_ERR_INVALID, "rtb_push_bands: bad argument");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    cudaStream_t st = (cudaStream_t)stream;
    const int bands = (height + rtb::BAND_ROWS - 1) / rtb::BAND_ROWS;
    const int mine = (bands - tile_rank + tile_count - 1) / tile_count;            // bands tile_rank, tile_rank + tile_count, ...
    if (mine <= 0) return RTB_OK;
    const int last = tile_rank + (mine - 1) * tile_count;                           // my last band: may be cut by the frame's end
    const int last_rows = std::min(rtb::BAND_ROWS, height - last * rtb::BAND_ROWS);
    const int full = last_rows == rtb::BAND_ROWS ? mine : mine - 1;
    const size_t band_bytes = (size_t)rtb::BAND_ROWS * pitch, stride = band_bytes * (size_t)tile_count, off = band_bytes * (size_t)tile_rank;
    if (full > 0) CU(cudaMemcpy2DAsync(dst + off, stride, src + off, stride, band_bytes, (size_t)full, cudaMemcpyDefault, st), "band push");
    if (full < mine) {
        const size_t o = band_bytes * (size_t)last;
        CU(cudaMemcpyAsync(dst + o, src + o, (size_t)last_rows * pitch, cudaMemcpyDefault, st), "band push (last band)");
    }
    return RTB_OK;
}

// ------------------------------------------------------------------ frame-complete signal over peer memory
}  // extern "C"
// One thread publishes `value` in the owner's flag word after everything this stream wrote before it is visible system-wide.
__global__ void k_frame_signal(unsigned* flags, int slot, unsigned value)
{
    __threadfence_system();
    *((volatile unsigned*)flags + slot) = value;
    __threadfence_system();
}
// Thread i waits until flags[first + i] has reached `value` (wrap-safe compare).  Bounded: ~2 s of spinning sets the
// frame's overflow flag instead of hanging the device (a missing peer must not cost a GPU).
__global__ void k_frame_wait(const unsigned* flags, int first, int count, unsigned value, rtb::Counters* ctr)
{
    const int i = threadIdx.x;
    if (i >= count) return;
    const volatile unsigned* f = (const volatile unsigned*)flags + first + i;
    const long long t0 = clock64();
    while ((int)(*f - value) < 0) {
        __nanosleep(200);
        if (clock64() - t0 > 4000000000ll) { ctr->overflow = 1; break; }
    }
    __threadfence_system();
}
extern "C" {
int rtb_frame_signal(rtb_context* c, unsigned char* flags, int slot, unsigned value, void* stream)
{
    if (!c || !flags || slot < 0) return fail(RTB_ERR_INVALID, "rtb_frame_signal: bad argument");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    k_frame_signal<<<1, 1, 0, (cudaStream_t)stream>>>((unsigned*)flags, slot, value);
    CU(cudaGetLastError(), "k_frame_signal");
    return RTB_OK;
}

int rtb_frame_wait(rtb_context* c, const unsigned char* flags, int first, int count, unsigned value, void* stream)
{
    if (!c || !flags || first < 0 || count < 0 || count > 1024) return fail(RTB_ERR_INVALID, "rtb_frame_wait: bad argument");
    if (count == 0) return RTB_OK;
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    k_frame_wait<<<1, ((count + 31) / 32) * 32, 0, (cudaStream_t)stream>>>((const unsigned*)flags, first, count, value, c->counters_d);
    CU(cudaGetLastError(), "k_frame_wait");
    return RTB_OK;
}

int rtb_synchronize(rtb_context* c)
{
    if (!c) return fail(RTB_ERR_INVALID, "rtb_synchronize: NULL context");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    CU(cudaDeviceSynchronize(), "cudaDeviceSynchronize");
    return RTB_OK;
}

int rtb_frame_stats_get(rtb_context* c, rtb_frame_stats* out)
{
    if (!c || !out) return fail(RTB_ERR_INVALID, "rtb_frame_stats_get: NULL argument");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    Counters h;
    CU(cudaMemcpy(&h, c->counters_d, sizeof h, cudaMemcpyDeviceToHost), "counters");
    out->rays = h.rays; out->nodes_visited = h.nodes; out->tris_tested = h.tris; out->device_ms = c->last_ms; out->overflow = h.overflow;
    out->launches = c->last_launches; out->kernel_variant = c->last_info.variant;
    if (h.overflow) return fail(RTB_ERR_OVERFLOW, "render: traversal stack / ray queue overflow");
    return RTB_OK;
}

// ------------------------------------------------------------------ device-math inspection (tests)
}  // extern "C"
__global__ void k_debug_math(int what, const float* __restrict__ in, float* __restrict__ out, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (what == 0) out[i] = rtb::xatanf(in[i]);
    else if (what == 1) out[i] = rtb::q_rsqrt(in[i]);
}
extern "C" {
int rtb_debug_math(rtb_context* c, int what, const float* in_host, float* out_host, int n)
{
    if (!c || !in_host || !out_host || n <= 0 || what < 0 || what > 1) return fail(RTB_ERR_INVALID, "rtb_debug_math: bad argument");
    CU(cudaSetDevice(c->device), "cudaSetDevice");
    float *di = nullptr, *dout = nullptr;
    CU(cudaMalloc((void**)&di, sizeof(float) * (size_t)n), "cudaMalloc");
    CU(cudaMalloc((void**)&dout, sizeof(float) * (size_t)n), "cudaMalloc");
    CU(cudaMemcpy(di, in_host, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice), "copy");
    k_debug_math<<<(n + 255) / 256, 256>>>(what, di, dout, n);
    CU(cudaMemcpy(out_host, dout, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost), "copy");
    cudaFree(di); cudaFree(dout);
    return RTB_OK;
}

// ------------------------------------------------------------------ host helpers
void rtb_invert_intrinsic(const float K[9], float Ki[9]) { hm::invert_intrinsic(K, Ki); }
void rtb_invert_lre(const rtb_lre* in, rtb_lre* out) { *out = hm::invert_pose(*in); }
void rtb_instance_build_inv(rtb_instance* inst) { hm::instance_build_inv(*inst); }
void rtb_triangle_init(rtb_triangle* t, const float a[3], const float b[3], const float c[3])
{
    memset(t, 0, sizeof *t);
    for (int k = 0; k < 3; k++) { t->vertices[0][k] = a[k]; t->vertices[1][k] = b[k]; t->vertices[2][k] = c[k]; }
    hm::V3 n = hm::face_normal(hm::V3{ a[0], a[1], a[2] }, hm::V3{ b[0], b[1], b[2] }, hm::V3{ c[0], c[1], c[2] });
    t->normal[0] = n.x; t->normal[1] = n.y; t->normal[2] = n.z;
}
void rtb_free(void* p) { free(p); }

}  // extern "C"
