// raytracer.hpp -- C++ drop-in surface of the B200 ray-casting core.
//
// One header that provides the class and function names an application written against
// AFIDclan/cuda-raytracing uses (kernel.cu:141-302 is the model caller): Scene, Camera, MeshInstance, Material,
// MeshPrimitive, BVHTree::print_stats, OBJLoader::load, TrianglePrimitive, transforms::*, and the free vector
// helpers.  The small forwarding headers next to this file (Scene.h, Camera.h, ...) keep the reference's
// #include lines working.  Every device-side action goes through the C-ABI in include/rtb200.h; this file
// contains no kernels and no CPU implementation of the render path.
//
// Differences that are visible to a caller (all listed in INTEGRATION.md):
//   * the BVH is built on the GPU when Scene::upload_to_device() runs, not in the MeshPrimitive constructor;
//   * Material::upload_texture() reads binary PPM (P6) or takes raw BGR texels -- image decoding via OpenCV is out of scope;
//   * Scene::d_* are opaque non-null tokens after upload (the device arrays live behind the context handle);
//   * Camera has three extra public fields (mode, max_depth, spp) selecting the north-star extensions; defaults
//     reproduce the reference's output.
#pragma once
#include <cuda_runtime.h>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>
#include <string>
#include <vector>
#include "../../include/rtb200.h"

// ----------------------------------------------------------------------------------------------- transforms
namespace transforms {

struct lre {                       // transforms.hpp:10-14 -- x,y,z + yaw(z) pitch(x) roll(y)
    float x, y, z, yaw, pitch, roll;
    lre() : x(0.f), y(0.f), z(0.f), yaw(0.f), pitch(0.f), roll(0.f) {}
};
struct float3x3 { float m[3][3]; };
struct float4x4 { float m[4][4]; };

namespace detail {
inline rtb_lre to_c(const lre& p) { rtb_lre r = { p.x, p.y, p.z, p.yaw, p.pitch, p.roll }; return r; }
inline lre from_c(const rtb_lre& r) { lre p; p.x = r.x; p.y = r.y; p.z = r.z; p.yaw = r.yaw; p.pitch = r.pitch; p.roll = r.roll; return p; }
}  // namespace detail

inline float3x3 invert_rotmat(const float3x3& r)
{
    float3x3 t;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) t.m[i][j] = r.m[j][i];
    return t;
}
inline float3 apply_rotmat(const float3x3& r, const float3& v)
{
    return make_float3(r.m[0][0] * v.x + r.m[0][1] * v.y + r.m[0][2] * v.z, r.m[1][0] * v.x + r.m[1][1] * v.y + r.m[1][2] * v.z,
                       r.m[2][0] * v.x + r.m[2][1] * v.y + r.m[2][2] * v.z);
}
inline float3x3 euler2rotmat(float3 e)            // transforms.hpp:129-144
{
    const float sy = sinf(e.x), cy = cosf(e.x), sp = sinf(e.y), cp = cosf(e.y), sr = sinf(e.z), cr = cosf(e.z);
    float3x3 r;
    r.m[0][0] = cr * cy + sr * sp * sy; r.m[0][1] = -cr * sy + sr * sp * cy; r.m[0][2] = -sr * cp;
    r.m[1][0] = cp * sy;                r.m[1][1] = cp * cy;                 r.m[1][2] = sp;
    r.m[2][0] = sr * cy - cr * sp * sy; r.m[2][1] = -sr * sy - cr * sp * cy; r.m[2][2] = cr * cp;
    return r;
}
inline float3 rotmat2euler(float3x3 r)            // transforms.hpp:119-126
{
    float a = r.m[1][2];
    a = a > 1 ? 1 : (a < -1 ? -1 : a);
    return make_float3(atan2f(r.m[1][0], r.m[1][1]), asinf(a), atan2f(-r.m[0][2], r.m[2][2]));
}
inline float4 euler2quat(float3 e)                // transforms.hpp:148-163
{
    const float hy = (float)(e.x * 0.5), hp = (float)(e.y * 0.5), hr = (float)(e.z * 0.5);
    const float sy = sinf(hy), cy = cosf(hy), sp = sinf(hp), cp = cosf(hp), sr = sinf(hr), cr = cosf(hr);
    return make_float4(sy * sp * sr + cy * cp * cr, cy * sp * cr + sy * cp * sr, -sy * sp * cr + cy * cp * sr, cy * sp * sr - sy * cp * cr);
}
inline float3 apply_quat(float4 q, float3 v)      // transforms.hpp:165-176
{
    const float a = -v.x * q.y - v.y * q.z - v.z * q.w, b = v.x * q.x + v.y * q.w - v.z * q.z;
    const float c = v.y * q.x + v.z * q.y - v.x * q.w, d = v.z * q.x + v.x * q.z - v.y * q.y;
    return make_float3(q.x * b - q.y * a - q.z * d + q.w * c, q.x * c - q.z * a - q.w * b + q.y * d, q.x * d - q.w * a - q.y * c + q.z * b);
}
inline float3 apply_euler(float3 e, float3 v) { return apply_quat(euler2quat(e), v); }
inline float3 apply_lre(lre p, float3 v) { return apply_euler(make_float3(p.yaw, p.pitch, p.roll), make_float3(v.x - p.x, v.y - p.y, v.z - p.z)); }
inline float4x4 matmul(float4x4 a, float4x4 b)
{
    float4x4 r;
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) { r.m[i][j] = 0.0f; for (int k = 0; k < 4; k++) r.m[i][j] += a.m[i][k] * b.m[k][j]; }
    return r;
}
inline float4x4 compose_homo(float4x4 h1, float4x4 h2) { return matmul(h2, h1); }
inline float4x4 lre2homo(lre p)                   // transforms.hpp:178-193
{
    const float3x3 R = euler2rotmat(make_float3(p.yaw, p.pitch, p.roll));
    const float3 t = apply_rotmat(R, make_float3(-p.x, -p.y, -p.z));
    float4x4 h;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) h.m[i][j] = R.m[i][j];
    h.m[0][3] = t.x; h.m[1][3] = t.y; h.m[2][3] = t.z;
    h.m[3][0] = h.m[3][1] = h.m[3][2] = 0.0f; h.m[3][3] = 1.0f;
    return h;
}
inline float4x4 invert_homo(const float4x4& H)    // transforms.hpp:76-100
{
    float3x3 R;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) R.m[i][j] = H.m[i][j];
    const float3x3 Ri = invert_rotmat(R);
    const float3 t = apply_rotmat(Ri, make_float3(-H.m[0][3], -H.m[1][3], -H.m[2][3]));
    float4x4 o;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) o.m[i][j] = Ri.m[i][j];
    o.m[0][3] = t.x; o.m[1][3] = t.y; o.m[2][3] = t.z;
    o.m[3][0] = o.m[3][1] = o.m[3][2] = 0.0f; o.m[3][3] = 1.0f;
    return o;
}
inline lre homo2lre(float4x4 H)                   // transforms.hpp:195-216
{
    float3x3 R;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) R.m[i][j] = H.m[i][j];
    const float3 e = rotmat2euler(R);
    const float3 s = apply_rotmat(invert_rotmat(R), make_float3(H.m[0][3], H.m[1][3], H.m[2][3]));
    lre o;
    o.x = -s.x; o.y = -s.y; o.z = -s.z; o.yaw = e.x; o.pitch = e.y; o.roll = e.z;
    return o;
}
inline lre compose_lre(lre a, lre b) { return homo2lre(compose_homo(lre2homo(a), lre2homo(b))); }
// kernel inputs go through the library's own host arithmetic so they carry exactly the bits the kernels expect
inline lre invert_lre(lre p) { rtb_lre in = detail::to_c(p), out; rtb_invert_lre(&in, &out); return detail::from_c(out); }

inline void print(const lre& p) { std::cout << p.x << ", " << p.y << ", " << p.z << ", " << p.yaw << ", " << p.pitch << ", " << p.roll << std::endl; }
inline void print(const float3 v) { std::cout << v.x << ", " << v.y << ", " << v.z << std::endl; }
inline void print(const float4 v) { std::cout << v.x << ", " << v.y << ", " << v.z << ", " << v.w << std::endl; }
inline void print(const float3x3& m) { for (int i = 0; i < 3; i++) { for (int j = 0; j < 3; j++) std::cout << m.m[i][j] << ' '; std::cout << std::endl; } }
inline void print(const float4x4& m) { for (int i = 0; i < 4; i++) { for (int j = 0; j < 4; j++) std::cout << m.m[i][j] << ' '; std::cout << std::endl; } }

}  // namespace transforms
using namespace transforms;

// ----------------------------------------------------------------------------------------------- vector helpers (utils.hpp)
static inline float Q_rsqrt(float number)        // utils.hpp:12-27, 32-bit pun
{
    float y = number;
    int32_t i;
    std::memcpy(&i, &y, 4);
    i = 0x5f3759df - (i >> 1);
    std::memcpy(&y, &i, 4);
    return y * (1.5F - ((number * 0.5F) * y * y));
}
static inline float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline float3 cross(float3 a, float3 b) { return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
static inline float magnitude(float3 v) { return sqrtf(v.x * v.x + v.y * v.y + v.z * v.z); }
static inline float magnitude(float2 v) { return sqrtf(v.x * v.x + v.y * v.y); }
static inline float inv_magnitude(float3 v) { return Q_rsqrt(v.x * v.x + v.y * v.y + v.z * v.z); }
static inline float3 normalize(float3 v) { const float s = inv_magnitude(v); return make_float3(v.x * s, v.y * s, v.z * s); }
static inline float3 operator+(float3 a, float3 b) { return make_float3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline float3 operator-(float3 a, float3 b) { return make_float3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline float3 operator-(float3 a, float b) { return make_float3(a.x - b, a.y - b, a.z - b); }
static inline float3 operator*(float3 a, float3 b) { return make_float3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline float3 operator*(float3 a, float b) { return make_float3(a.x * b, a.y * b, a.z * b); }
static inline float3 operator*(float b, float3 a) { return make_float3(a.x * b, a.y * b, a.z * b); }
static inline float3 operator/(float3 a, float b) { return make_float3(a.x / b, a.y / b, a.z / b); }
static inline float3 f3_min(float3 a, float3 b) { return make_float3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
static inline float3 f3_max(float3 a, float3 b) { return make_float3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
static inline float2 operator*(float2 a, float b) { return make_float2(a.x * b, a.y * b); }
static inline float2 operator*(float a, float2 b) { return make_float2(a * b.x, a * b.y); }
static inline float2 operator*(float2 a, float2 b) { return make_float2(a.x * b.x, a.y * b.y); }
static inline float2 operator+(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
static inline float2 operator-(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
static inline float2 operator/(float2 a, float b) { return make_float2(a.x / b, a.y / b); }
template <typename T> void cu_swap(T& a, T& b) { T t = a; a = b; b = t; }
static inline float3 apply_matrix(const float3x3& m, const float3& v) { return transforms::apply_rotmat(m, v); }
static inline float4 apply_matrix(const float4x4& m, const float4& v)
{
    float o[4];
    for (int i = 0; i < 4; i++) o[i] = m.m[i][0] * v.x + m.m[i][1] * v.y + m.m[i][2] * v.z + m.m[i][3] * v.w;
    return make_float4(o[0], o[1], o[2], o[3]);
}
static inline float3x3 invert_intrinsic(const float3x3& K)      // utils.hpp:142-160, evaluated by the library
{
    float3x3 r;
    rtb_invert_intrinsic(&K.m[0][0], &r.m[0][0]);
    return r;
}

// ----------------------------------------------------------------------------------------------- geometry
struct TrianglePrimitive {          // TrianglePrimitive.hpp:8-60; layout == rtb_triangle (72 bytes)
    float3 vertices[3];
    float3 normal;
    float2 uv_coords[3];
    TrianglePrimitive() { std::memset(this, 0, sizeof *this); }
    TrianglePrimitive(float3 a, float3 b, float3 c)
    {
        rtb_triangle_init(reinterpret_cast<rtb_triangle*>(this), &a.x, &b.x, &c.x);     // normal = normalize(cross(b-a, c-a))
    }
    TrianglePrimitive(float3 a, float3 b, float3 c, float3 n)
    {
        std::memset(this, 0, sizeof *this);
        vertices[0] = a; vertices[1] = b; vertices[2] = c; normal = n;
    }
    TrianglePrimitive(float3 a, float3 b, float3 c, float3 n, float2 ta, float2 tb, float2 tc) : TrianglePrimitive(a, b, c, n)
    {
        uv_coords[0] = ta; uv_coords[1] = tb; uv_coords[2] = tc;
    }
    float3 center() const { return (vertices[0] + vertices[1] + vertices[2]) / 3.0f; }
};
static_assert(sizeof(TrianglePrimitive) == sizeof(rtb_triangle), "TrianglePrimitive must stay 72 bytes");

namespace rtb_detail {
struct Context {                    // one C-ABI context per Scene
    rtb_context* ctx = nullptr;
    ~Context() { if (ctx) rtb_context_destroy(ctx); }
};
struct MeshState {                  // shared by every copy of a MeshPrimitive (Scene::add_mesh copies by value)
    std::vector<TrianglePrimitive> triangles;
    std::shared_ptr<Context> owner;
    int mesh_id = -1;
    int replicas = 0;               // devices the BLAS is resident on (Scene::use_devices)
};
struct TextureState { std::vector<unsigned char> bgr; int w = 0, h = 0; bool uploaded = false; std::vector<unsigned char*> dev; size_t pitch = 0; };
inline void die(const char* what)
{
    std::cerr << what << ": " << rtb_last_error() << std::endl;
    std::exit(1);
}
}  // namespace rtb_detail

// BVHTree.hpp:117-172 -- only the inspection entry point survives; the tree itself is built and kept on the GPU.
struct BVHTree {
    std::shared_ptr<rtb_detail::MeshState> state;
    void print_stats()
    {
        std::cout << "BVH Stats: " << std::endl;
        if (!state || state->mesh_id < 0 || !state->owner) {
            std::cout << "Triangles: " << (state ? state->triangles.size() : 0) << " (BVH is built on the GPU by Scene::upload_to_device)" << std::endl;
            return;
        }
        rtb_mesh_info info;
        if (rtb_mesh_info_get(state->owner->ctx, state->mesh_id, &info) != RTB_OK) rtb_detail::die("print_stats");
        std::cout << "Number of nodes: " << info.num_nodes << " (4-wide)" << std::endl;
        std::cout << "Max triangles per node: " << info.max_leaf_size << std::endl;
        std::cout << "Max depth: " << info.depth << std::endl;
        std::cout << "Number of leaves: " << info.num_leaves << std::endl;
        std::cout << "Average triangles per leaf: " << (info.num_leaves ? (float)info.num_triangles / (float)info.num_leaves : 0.f) << std::endl;
        std::cout << "Device build time (ms): " << info.build_ms << std::endl;
    }
};

struct d_MeshPrimitive { int num_triangles; int mesh_id; };      // MeshPrimitive.h:13-24 stand-in (opaque to callers)

class MeshPrimitive {               // MeshPrimitive.h:26-43
public:
    MeshPrimitive(std::vector<TrianglePrimitive> triangles)
    {
        bvh_top.state = std::make_shared<rtb_detail::MeshState>();
        bvh_top.state->triangles = std::move(triangles);
        num_triangles = (int)bvh_top.state->triangles.size();
    }
    int num_triangles;
    BVHTree bvh_top;
    d_MeshPrimitive* to_device() { handle.num_triangles = num_triangles; handle.mesh_id = bvh_top.state->mesh_id; return &handle; }
    const std::vector<TrianglePrimitive>& host_triangles() const { return bvh_top.state->triangles; }
private:
    d_MeshPrimitive handle = { 0, -1 };
};

struct Material {                   // Material.hpp:6-44; first 48 bytes == rtb_material
    float roughness;
    float3 albedo;
    float metallic;
    float illumination;
    uchar3* texture;
    int texture_width;
    int texture_height;
    size_t texture_pitch;
    std::shared_ptr<rtb_detail::TextureState> texels;            // host copy until the Scene uploads it
    Material() : roughness(0.0f), albedo(make_float3(1.0f, 1.0f, 1.0f)), metallic(0.0f), illumination(0.0f), texture(nullptr),
                 texture_width(0), texture_height(0), texture_pitch(0) {}
    Material* to_device() { return this; }
    void upload_texture(const unsigned char* bgr, int width, int height)
    {
        texels = std::make_shared<rtb_detail::TextureState>();
        texels->bgr.assign(bgr, bgr + (size_t)width * height * 3);
        texels->w = width; texels->h = height;
        texture_width = width; texture_height = height;
    }
    void upload_texture(std::string fp)                          // binary PPM (P6, maxval 255); RGB is stored as BGR like cv::imread
    {
        std::ifstream f(fp, std::ios::binary);
        std::string magic; int w = 0, h = 0, maxv = 0;
        if (!(f >> magic >> w >> h >> maxv) || magic != "P6" || maxv != 255 || w <= 0 || h <= 0) {
            std::cerr << "Could not read texture " << fp << " (binary PPM expected)" << std::endl;
            texture_width = texture_height = 0;
            return;
        }
        f.get();
        std::vector<unsigned char> rgb((size_t)w * h * 3);
        f.read((char*)rgb.data(), (std::streamsize)rgb.size());
        for (size_t i = 0; i + 2 < rgb.size(); i += 3) std::swap(rgb[i], rgb[i + 2]);
        upload_texture(rgb.data(), w, h);
    }
};

struct MeshInstance {               // MeshInstance.hpp:6-48; layout == rtb_instance (104 bytes)
    int mesh_index;
    int material_index;
    lre pose;
    lre inv_pose;
    float3 rotation;
    float3 inv_rotation;
    float3 scale;
    float3 inv_scale;
    MeshInstance() { std::memset(this, 0, sizeof *this); mesh_index = -1; scale = make_float3(1.0f, 1.0f, 1.0f); }
    MeshInstance(int mesh_index, int material_index) : MeshInstance()
    {
        this->mesh_index = mesh_index; this->material_index = material_index;
        build_inv();
    }
    MeshInstance(int mesh_index, int material_index, lre pose, float3 scale) : MeshInstance()
    {
        this->mesh_index = mesh_index; this->material_index = material_index; this->pose = pose; this->scale = scale;
        build_inv();
    }
    void build_inv() { rtb_instance_build_inv(reinterpret_cast<rtb_instance*>(this)); }      // MeshInstance.hpp:39-46
};
static_assert(sizeof(MeshInstance) == sizeof(rtb_instance), "MeshInstance must stay 104 bytes");

#include "raytracer_scene.hpp"
