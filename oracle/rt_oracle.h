/*
 * rt_oracle.h -- CPU restatement of the AFIDclan/cuda-raytracing render path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 * The product (cuda_raytracing_b200/) never links, imports or falls back to this code.
 *
 * Parity status: the reference ships no tests, golden vectors or fixtures so
 * the reference itself leaves parity unpinned.  This restatement is pinned instead against
 *   (1) the reference's own sources compiled as host C++ (oracle/_ref/libref_host.so, built by
 *       oracle/build_ref.sh from the files where they lie under /root/reference) -- bit-identical
 *       colour / primitive-id / distance images and bit-identical BVH node arrays, and
 *   (2) the survey known-answer values (Q_rsqrt bit patterns, struct sizes, BVH node counts),
 *       frozen in tests/golden/.
 *
 * Every function cites the reference file:line (under /root/reference/CudaRaytracer/) it follows.
 * Arithmetic is IEEE binary32 with no FMA contraction (build with -ffp-contract=off), which is what
 * the reference's host path does on x86-64.  Q_rsqrt uses the 32-bit pun (the reference author's
 * Windows/MSVC `long`; SURVEY.md 8a Q1 -- on LP64 the as-is code is undefined behaviour).
 */
#ifndef RT_ORACLE_H
#define RT_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct { float x, y, z; } o_f3;
typedef struct { float x, y; } o_f2;
typedef struct { float x, y, z, w; } o_f4;

/* TrianglePrimitive.hpp:8-11  (72 bytes) */
typedef struct { o_f3 vertices[3]; o_f3 normal; o_f2 uv_coords[3]; } o_tri;
/* transforms.hpp:10-14  (24 bytes) */
typedef struct { float x, y, z, yaw, pitch, roll; } o_lre;
/* MeshInstance.hpp:6-18  (104 bytes) */
typedef struct {
    int mesh_index, material_index;
    o_lre pose, inv_pose;
    o_f3 rotation, inv_rotation, scale, inv_scale;
} o_instance;
/* Material.hpp:6-19  (48 bytes) */
typedef struct {
    float roughness; o_f3 albedo; float metallic; float illumination;
    const uint8_t* texture; int texture_width, texture_height; size_t texture_pitch;
} o_material;

/* Flattened BVHTree (BVHTree.hpp:59-74): node i of master_list_trees. */
typedef struct {
    o_f3 min, max;
    int tri_offset, tri_count;      /* leaf: slice of o_mesh.indices; interior: 0,0 (BVHTree.hpp:97-114) */
    int child_a, child_b;           /* -1 for leaves (BVHTree.hpp:66-67) */
} o_node;

typedef struct {
    int num_triangles; o_tri* triangles;
    int num_nodes; o_node* nodes;
    int num_indices; int* indices;
} o_mesh;

typedef struct {
    int num_meshes; o_mesh** meshes;
    int num_materials; o_material* materials;
    int num_instances; o_instance* instances;
} o_scene;

enum {
    O_MODE_REFERENCE = 0,   /* raycast.cu:146-297 exactly as shipped (unlit)                       */
    O_MODE_SHADOW    = 1,   /* + the commented-out Lambert / shadow-ray block raycast.cu:249-290   */
    O_MODE_REFLECT   = 2,   /* + mirror reflections (extension; spec in DESIGN.md)                 */
    O_MODE_PATH      = 3    /* diffuse path tracing (extension; spec in DESIGN.md)                 */
};

typedef struct {
    int width, height;
    float K_inv[9];           /* row-major float3x3 (utils.hpp:142-160) */
    float D[4];
    o_lre pose, inv_pose;     /* Camera.cu:21 */
    int mode, max_depth, spp;
    uint32_t seed;
    int y0, y1;               /* rows [y0, y1) are rendered */
} o_render_params;

typedef struct {            /* per-frame work counters (for the bounded cpu_baseline sample) */
    uint64_t rays, nodes_popped, boxes_tested, tris_tested;
} o_stats;

/* ---- math (utils.hpp / transforms.hpp) ---- */
float o_Q_rsqrt(float number);
void o_libm_atanf(const float* in, float* out, int n);   /* libm atanf as called at raycast.cu:170 */
o_f3 o_normalize(o_f3 v);
o_f3 o_apply_euler(o_f3 euler, o_f3 v);
o_f3 o_apply_lre(o_lre l, o_f3 v);
o_f4 o_euler2quat(o_f3 euler);
o_f3 o_apply_quat(o_f4 q, o_f3 v);
o_lre o_invert_lre(o_lre l);
void o_invert_intrinsic(const float K[9], float K_inv[9]);
void o_instance_build_inv(o_instance* inst);

/* ---- geometry ---- */
o_tri o_make_triangle(o_f3 a, o_f3 b, o_f3 c);             /* OBJLoader.hpp:143-168 normal convention */
o_mesh* o_mesh_create(const o_tri* tris, int n);           /* MeshPrimitive.cpp:5-15,38-56 + BVHTree::fill */
void o_mesh_destroy(o_mesh* m);
void o_mesh_stats(const o_mesh* m, int out[6]);            /* BVHTree.hpp:117-172: nodes,max,min,depth,leaves,ntris */
int o_obj_load(const char* path, o_tri** out_tris);        /* OBJLoader.hpp:15-179; returns count or -1 */
void o_free(void* p);

/* ---- render (raycast.cu) ---- */
void o_render(const o_scene* scene, const o_render_params* p, uint8_t* img, size_t pitch,
              int32_t* aov_prim, int32_t* aov_inst, float* aov_t, int nthreads, o_stats* stats);

/* single ray closest hit, for unit tests: returns 1 on hit */
int o_cast_ray(const o_scene* scene, o_f3 origin, o_f3 direction, int* prim, int* inst, float* t,
               o_f3* location, o_f3* normal, o_f2* uv);
/* parity-report support (see rt_oracle.c): one candidate's verdict, the BVH-free closest hit, a pixel's camera ray */
int o_candidate(const o_scene* scene, o_f3 origin, o_f3 direction, int inst, int prim, float* distance);
int o_cast_ray_exhaustive(const o_scene* scene, o_f3 origin, o_f3 direction, int* prim, int* inst, float* t);
void o_pixel_ray(const o_render_params* p, int x, int y, o_f3* origin, o_f3* direction);
void o_set_exhaustive(int on);            /* every cast_ray becomes a BVH-free search over all triangles (process-wide) */
void o_render_pixel(const o_scene* scene, const o_render_params* p, int x, int y, uint8_t bgr[3], int32_t* prim, int32_t* inst, float* t);

#ifdef __cplusplus
}
#endif
#endif
