/*
 * rtb200.h -- C-ABI of the B200-native ray-casting core (librtb200.so).
 *
 * Drop-in boundary for the render path of AFIDclan/cuda-raytracing.  The reference has no FFI layer:
 * its boundary is the public C++ surface consumed by kernel.cu:166-299.  Every entry point below
 * names the reference interface it replaces (paths under /root/reference/CudaRaytracer/).  The C++
 * classes with the reference's names (cuda_raytracing_b200/cpp/*.h) and the ctypes binding
 * (cuda_raytracing_b200/api.py) are thin callers of exactly these functions.
 *
 * Conventions: plain pointers and sizes, no C++/torch types; every function returns 0 on success and
 * a negative rtb_status otherwise (rtb_last_error() gives the text).  There is NO CPU fallback: if no
 * CUDA device is usable rtb_context_create fails.  POD layouts are byte-identical to the
 * reference's structs so existing arrays can be passed through unchanged.
 */
#ifndef RTB200_H
#define RTB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTB_API __attribute__((visibility("default")))

typedef enum {
    RTB_OK = 0,
    RTB_ERR_INVALID = -1,      /* bad argument / handle / index                  */
    RTB_ERR_CUDA = -2,         /* a CUDA runtime call failed                     */
    RTB_ERR_NO_DEVICE = -3,    /* no usable sm_100 device (there is no fallback) */
    RTB_ERR_STATE = -4,        /* scene not committed, nothing to render, ...    */
    RTB_ERR_OVERFLOW = -5      /* traversal stack / queue capacity exceeded      */
} rtb_status;

/* TrianglePrimitive (TrianglePrimitive.hpp:8-11), 72 bytes */
typedef struct { float vertices[3][3]; float normal[3]; float uv_coords[3][2]; } rtb_triangle;
/* transforms::lre (transforms.hpp:10-14), 24 bytes */
typedef struct { float x, y, z, yaw, pitch, roll; } rtb_lre;
/* MeshInstance (MeshInstance.hpp:6-18), 104 bytes; inv_* as produced by build_inv() (39-46) */
typedef struct {
    int mesh_index, material_index;
    rtb_lre pose, inv_pose;
    float rotation[3], inv_rotation[3], scale[3], inv_scale[3];
} rtb_instance;
/* Material (Material.hpp:6-19), 48 bytes; `texture` is a DEVICE pointer to pitched BGR texels */
typedef struct {
    float roughness; float albedo[3]; float metallic; float illumination;
    const unsigned char* texture; int texture_width, texture_height; size_t texture_pitch;
} rtb_material;

typedef struct rtb_context rtb_context;   /* owns every device allocation (replaces the leaked cudaMallocs
                                             of BVHTree.hpp:97-114,364-383 / MeshPrimitive.cpp:17-36 / Scene.cpp:25-65) */

enum {                       /* rtb_render_params.mode */
    RTB_MODE_REFERENCE = 0,  /* raycast.cu:146-297 as shipped: primary closest hit, unlit albedo/texture     */
    RTB_MODE_SHADOW    = 1,  /* + Lambert term and shadow ray of the commented block raycast.cu:249-290       */
    RTB_MODE_REFLECT   = 2,  /* + mirror reflections to max_depth, weight Material::metallic (wavefront)     */
    RTB_MODE_PATH      = 3   /* diffuse path tracing, spp samples, max_depth bounces (wavefront)             */
};

/* Arguments of the reference's render<<<>>> launch (raycast.cu:146, marshalled in Camera.cu:18-36). */
typedef struct {
    int width, height;
    float K_inv[9];            /* Camera::K_inv = invert_intrinsic(K) (utils.hpp:142-160), row-major  */
    float D[4];                /* Camera::D fisheye coefficients (raycast.cu:172)                      */
    rtb_lre pose;              /* Camera::pose                                                         */
    rtb_lre inv_pose;          /* invert_lre(pose) (Camera.cu:21)                                      */
    int mode, max_depth, spp;
    uint32_t seed;
    int tile_rank, tile_count; /* multi-GPU image partition: this call renders the 8-row bands b with
                                  b % tile_count == tile_rank (0,1 = whole image)                      */
} rtb_render_params;

/* Optional per-pixel outputs (dense width*height DEVICE arrays, any may be NULL): what the reference's
 * HitInfo (raycast.cu:10-18) would need to carry to expose ids -- SURVEY 8c patch P3. */
typedef struct { int32_t* prim; int32_t* inst; float* t; } rtb_aovs;

typedef struct {
    int num_triangles, num_nodes, num_leaves, max_leaf_size, depth, stack_need;
    float build_ms;            /* device time of the last (re)build: Morton -> sort -> hierarchy -> fit -> collapse -> emit */
    float sah_cost;
    float bounds_min[3], bounds_max[3];
    size_t device_bytes;
} rtb_mesh_info;

typedef struct {
    uint64_t rays;             /* ray segments traced by the last rtb_render (primary + shadow + secondary) */
    uint64_t nodes_visited;    /* wide nodes fetched   (only when the context has counters enabled)         */
    uint64_t tris_tested;      /* triangle records fetched                                                  */
    float device_ms;           /* cudaEvent time of the last rtb_render_host / rtb_render_timed             */
    int overflow;              /* traversal stack or queue overflow flag (render fails loudly)              */
    int launches;              /* kernels of this library launched by the last rtb_render                   */
    int kernel_variant;        /* modes 2/3: 2 = persistent kernel with in-kernel ray queues, 1 = persistent kernel,
                                  0 = staged wavefront pipeline; -1 = the direct kernel (modes 0/1)           */
} rtb_frame_stats;

/* ---- context ------------------------------------------------------------------------------ */
RTB_API int rtb_context_create(int device, rtb_context** out);
RTB_API void rtb_context_destroy(rtb_context* ctx);
RTB_API const char* rtb_last_error(void);
RTB_API int rtb_version(void);
RTB_API int rtb_set_option(rtb_context* ctx, const char* name, int value);   /* names and defaults: INTEGRATION.md, "Run-time options" */

/* ---- BLAS: replaces MeshPrimitive::MeshPrimitive/build_bvh (MeshPrimitive.cpp:5-15,38-56),
 *      BVHTree::fill (BVHTree.hpp:203-292) and MeshPrimitive::to_device / BVHTree::compile_tree
 *      (MeshPrimitive.cpp:17-36, BVHTree.hpp:364-383) with an on-device LBVH build. ------------- */
RTB_API int rtb_mesh_create(rtb_context* ctx, const rtb_triangle* tris, int n, int tris_on_device, int* mesh_id);
/* Rebuilds a BLAS in place from new triangles (deforming geometry; the reference has no such call).  Instances of the mesh
 * keep working: when the new bounds leave the world boxes the TLAS was built from, the TLAS is rebuilt before the next
 * frame.  Like every scene-changing call it runs on the default stream and returns when the device has finished; a frame
 * still in flight on a NON-blocking stream must be waited for first (the reference has one stream and the same rule). */
RTB_API int rtb_mesh_rebuild(rtb_context* ctx, int mesh_id, const rtb_triangle* tris, int n, int tris_on_device);
/* The same rebuild off the frame's critical path (config C5: build frame f+1 while frame f renders).  _async enqueues the whole
 * build on the context's own build stream into the mesh's BACK buffer (the triangles must be on the device and stay untouched
 * until the commit) and returns at once; frames keep traversing the front buffer.  _commit waits -- the host, for that build
 * only -- swaps the buffers and refreshes rtb_mesh_info; frames enqueued afterwards see the new BLAS.  One build per mesh may be
 * pending.  No counterpart in the reference (its BVH is built once on the host, BVHTree.hpp:203-292). */
RTB_API int rtb_mesh_rebuild_async(rtb_context* ctx, int mesh_id, const rtb_triangle* tris_on_device, int n, int tris_on_device_flag);
RTB_API int rtb_mesh_commit(rtb_context* ctx, int mesh_id);
RTB_API int rtb_mesh_info_get(rtb_context* ctx, int mesh_id, rtb_mesh_info* out);
RTB_API int rtb_mesh_destroy(rtb_context* ctx, int mesh_id);
/* Inspection (tools/tests; the analogue of BVHTree::print_stats, BVHTree.hpp:117-172): copies device arrays to the host.
 * what = 0: wide nodes (num_nodes * 112 B; node 0 is the root, the order of the others is the builder's business), 1: triangle
 * records (num_triangles * 64 B), 2: leaf order (num_triangles * 4 B) */
RTB_API int rtb_mesh_read(rtb_context* ctx, int mesh_id, int what, void* out, size_t bytes);

/* ---- scene: replaces Scene::upload_to_device (Scene.cpp:25-65) / update_mesh_instance (67-74),
 *      Material::to_device (Material.hpp:21-27) and the texel half of upload_texture (29-43). ------ */
RTB_API int rtb_texture_upload(rtb_context* ctx, const unsigned char* bgr_host, int width, int height,
                               unsigned char** dev_ptr, size_t* pitch);
RTB_API int rtb_set_materials(rtb_context* ctx, const rtb_material* mats, int n);
RTB_API int rtb_set_instances(rtb_context* ctx, const rtb_instance* insts, int n);      /* builds the TLAS */
RTB_API int rtb_update_instance(rtb_context* ctx, int index, const rtb_instance* inst);  /* re-packs this instance, rebuilds the TLAS */

/* ---- render: replaces Camera::render_scene + the render kernel (Camera.cu:18-41, raycast.cu:146-297).
 *      img is a caller-owned DEVICE pointer, BGR uchar3 rows `pitch` bytes apart (kernel.cu:247-253).
 *      Returns after enqueue on `stream` (a cudaStream_t, NULL = default stream), like the reference. */
RTB_API int rtb_render(rtb_context* ctx, const rtb_render_params* p, unsigned char* img, size_t pitch,
                       const rtb_aovs* aovs, void* stream);
/* Host-buffer form (end-to-end): renders into an internal device frame, copies the frame (and AOVs, if
 * given as HOST pointers) to the caller's host memory and synchronises. */
RTB_API int rtb_render_host(rtb_context* ctx, const rtb_render_params* p, unsigned char* img_host, size_t pitch_host,
                            const rtb_aovs* aovs_host);
/* The same frame as rtb_render for RTB_MODE_REFLECT / RTB_MODE_PATH, for callers that cannot allocate (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  This entry point has no counterpart in the reference.
 * It calls no cudaMalloc, cudaFree or allocating library routine: the fused
 * kernel's per-pixel scratch lives in a static device array of RTB_NOALLOC_MAX_SLOTS pixel slots (17 318 144 bytes, about 16.5 MiB, that every GPU
 * which loads the library holds until the process exits).  Limits: width * height <= RTB_NOALLOC_MAX_SLOTS, and the rank's own
 * bands padded to claim blocks (64 x 32 pixels at the default "cost_block" 3) <= RTB_NOALLOC_MAX_SLOTS -- e.g. 320 x 240,
 * 448 x 256, 512 x 256; larger frames are rejected with RTB_ERR_INVALID and a message naming both numbers.  Modes 0 / 1 need
 * no scratch: use rtb_render.  The scene must be ready: at least one rtb_render / rtb_render_host after the last
 * rtb_set_instances / rtb_update_instance / mesh call (those upload tables and may allocate), otherwise RTB_ERR_STATE.
 * Pixels and AOVs are bit-identical to rtb_render's.  Every call shares the one array, which keeps the last call's contents
 * (the call clears the words that must start at zero); calls of one process that overlap on one GPU -- two streams, two
 * contexts on the same device -- race on it and produce garbage: serialise them (one stream, or events between streams). */
#define RTB_NOALLOC_MAX_SLOTS 131072
RTB_API int rtb_render_noalloc(rtb_context* ctx, const rtb_render_params* p, unsigned char* img, size_t pitch,
                               const rtb_aovs* aovs, void* stream);
RTB_API int rtb_synchronize(rtb_context* ctx);                                            /* cudaDeviceSynchronize, kernel.cu:279 */
RTB_API int rtb_frame_stats_get(rtb_context* ctx, rtb_frame_stats* out);
/* ---- multi-GPU frame exchange over peer memory (one process per GPU on one NVSwitch node; no counterpart in the
 *      reference, which renders on one GPU -- this is the (e) row of SURVEY.md 8).  The rank that owns the complete
 *      frame allocates it with rtb_frame_alloc_shared and passes the 64-byte handle to the other processes (any
 *      channel: torch.distributed object broadcast in tiles.py); they map it with rtb_frame_open_shared and either
 *      render their bands straight into the mapped pointer (rtb_render: the pixel stores travel over NVLink) or
 *      render locally and enqueue rtb_push_bands (one strided peer copy of the bands they own).  Both are stream
 *      ordered; the caller signals completion to the owner (a 4-byte all-reduce in tiles.py). */
#define RTB_IPC_HANDLE_BYTES 64
RTB_API int rtb_frame_alloc_shared(rtb_context* ctx, size_t bytes, unsigned char** dev_ptr, unsigned char handle[RTB_IPC_HANDLE_BYTES]);
RTB_API int rtb_frame_free_shared(rtb_context* ctx, unsigned char* dev_ptr);
RTB_API int rtb_frame_open_shared(rtb_context* ctx, const unsigned char handle[RTB_IPC_HANDLE_BYTES], unsigned char** dev_ptr);
RTB_API int rtb_frame_close_shared(rtb_context* ctx, unsigned char* dev_ptr);
/* copies rows [8b, 8b+8) for every band b with b % tile_count == tile_rank from src to dst (same pitch and height) */
RTB_API int rtb_push_bands(rtb_context* ctx, const unsigned char* src, unsigned char* dst, size_t pitch, int height,
                           int tile_rank, int tile_count, void* stream);
/* Completion signal over the same peer memory, so that no collective sits on the frame's path: `flags` points at 32-bit words
 * in memory every rank can reach (tiles.py keeps them behind the shared frame).  rtb_frame_signal enqueues, behind the caller's
 * render / push on `stream`, a system-scope fence and the store flags[slot] = value; rtb_frame_wait enqueues on the owner's
 * stream a kernel that returns once flags[first .. first+count) have all reached `value` (wrap-safe, so `value` is simply
 * the frame number).  The wait is bounded (about two seconds): a peer that never signals sets rtb_frame_stats.overflow
 * instead of hanging the device.  Ranks must be separate GPUs (a spinning kernel and its producer cannot share one). */
RTB_API int rtb_frame_signal(rtb_context* ctx, unsigned char* flags, int slot, unsigned value, void* stream);
RTB_API int rtb_frame_wait(rtb_context* ctx, const unsigned char* flags, int first, int count, unsigned value, void* stream);
/* Evaluates a device math helper on n host floats (tests): what = 0 atanf with host-libm bits (raycast.cu:170),
 * 1 Q_rsqrt (utils.hpp:12-27). */
RTB_API int rtb_debug_math(rtb_context* ctx, int what, const float* in_host, float* out_host, int n);

/* ---- host helpers with reference-exact arithmetic (kernel inputs must be bit-compatible) ------- */
RTB_API void rtb_invert_intrinsic(const float K[9], float K_inv[9]);      /* utils.hpp:142-160      */
RTB_API void rtb_invert_lre(const rtb_lre* in, rtb_lre* out);             /* transforms.hpp:232-235 */
RTB_API void rtb_instance_build_inv(rtb_instance* inst);                  /* MeshInstance.hpp:39-46 */
RTB_API void rtb_triangle_init(rtb_triangle* t, const float a[3], const float b[3], const float c[3]); /* TrianglePrimitive.hpp:15-24 */
RTB_API int rtb_obj_load(const char* path, rtb_triangle** tris, int* n);  /* OBJLoader.hpp:15-179; caller frees with rtb_free */
RTB_API void rtb_free(void* p);

#ifdef __cplusplus
}
#endif
#endif
