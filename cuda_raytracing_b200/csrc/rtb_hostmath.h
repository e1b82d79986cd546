// rtb_hostmath.h -- host-side pose / intrinsics arithmetic whose results are kernel inputs.
//
// These values (inverse poses, quaternions, K_inv, face normals, the light direction) are consumed by
// the exact predicates on the device, so they must carry the same bits the reference's host code
// produces: binary32, reference operand order, no FMA (build with -ffp-contract=off).  Cited lines are
// /root/reference/CudaRaytracer/{transforms.hpp,utils.hpp,MeshInstance.hpp}.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include "../../include/rtb200.h"

namespace rtb { namespace hm {

struct Quat { float x, y, z, w; };
struct V3 { float x, y, z; };

// utils.hpp:12-27 with a 32-bit pun
inline float q_rsqrt(float number)
{
    const float half = number * 0.5F;
    int32_t bits;
    std::memcpy(&bits, &number, 4);
    bits = 0x5f3759df - (bits >> 1);
    float y;
    std::memcpy(&y, &bits, 4);
    return y * (1.5F - (half * y * y));
}
// utils.hpp:37-47
inline V3 normalize(V3 v)
{
    const float s = q_rsqrt(v.x * v.x + v.y * v.y + v.z * v.z);
    return V3{ v.x * s, v.y * s, v.z * s };
}
// transforms.hpp:148-163 (half angles: float * 0.5 in double, narrowed -- exact)
inline Quat euler_to_quat(float yaw, float pitch, float roll)
{
    const float hy = (float)(yaw * 0.5), hp = (float)(pitch * 0.5), hr = (float)(roll * 0.5);
    const float sy = sinf(hy), cy = cosf(hy), sp = sinf(hp), cp = cosf(hp), sr = sinf(hr), cr = cosf(hr);
    Quat q;
    q.x = sy * sp * sr + cy * cp * cr;
    q.y = cy * sp * cr + sy * cp * sr;
    q.z = -sy * sp * cr + cy * cp * sr;
    q.w = cy * sp * sr - sy * cp * cr;
    return q;
}

// 3x3 rotation from yaw/pitch/roll, transforms.hpp:129-144
inline void euler_to_matrix(float yaw, float pitch, float roll, float R[3][3])
{
    const float sy = sinf(yaw), cy = cosf(yaw), sp = sinf(pitch), cp = cosf(pitch), sr = sinf(roll), cr = cosf(roll);
    R[0][0] = cr * cy + sr * sp * sy;  R[0][1] = -cr * sy + sr * sp * cy;  R[0][2] = -sr * cp;
    R[1][0] = cp * sy;                 R[1][1] = cp * cy;                  R[1][2] = sp;
    R[2][0] = sr * cy - cr * sp * sy;  R[2][1] = -sr * sy - cr * sp * cy;  R[2][2] = cr * cp;
}
inline V3 mat_apply(const float R[3][3], V3 v)
{
    return V3{ R[0][0] * v.x + R[0][1] * v.y + R[0][2] * v.z, R[1][0] * v.x + R[1][1] * v.y + R[1][2] * v.z,
               R[2][0] * v.x + R[2][1] * v.y + R[2][2] * v.z };
}
inline V3 mat_apply_transposed(const float R[3][3], V3 v)
{
    return V3{ R[0][0] * v.x + R[1][0] * v.y + R[2][0] * v.z, R[0][1] * v.x + R[1][1] * v.y + R[2][1] * v.z,
               R[0][2] * v.x + R[1][2] * v.y + R[2][2] * v.z };
}

// transforms.hpp:232-235 invert_lre = homo2lre(invert_homo(lre2homo(p))), written out on the 3x3 + column:
//   lre2homo   (178-193): H = [R | R*(-p)]
//   invert_homo (76-100): Hi = [R^T | R^T*(-(R*(-p)))]
//   homo2lre  (195-216): euler = rotmat2euler(R^T) (119-126); xyz = -( (R^T)^T * Hi_t )
inline rtb_lre invert_pose(const rtb_lre& p)
{
    float R[3][3];
    euler_to_matrix(p.yaw, p.pitch, p.roll, R);
    const V3 ht = mat_apply(R, V3{ -p.x, -p.y, -p.z });
    const V3 it = mat_apply_transposed(R, V3{ -ht.x, -ht.y, -ht.z });
    // Ri = R^T ; rotmat2euler reads Ri[1][2], Ri[1][0], Ri[1][1], Ri[0][2], Ri[2][2]
    float a = R[2][1];
    if (a > 1) a = 1; else if (a < -1) a = -1;
    rtb_lre o;
    o.yaw = atan2f(R[0][1], R[1][1]);
    o.pitch = asinf(a);
    o.roll = atan2f(-R[2][0], R[2][2]);
    const V3 back = mat_apply(R, it);           // (Ri)^T = R applied to the inverse translation column
    o.x = -back.x; o.y = -back.y; o.z = -back.z;
    return o;
}

// utils.hpp:142-160
inline void invert_intrinsic(const float K[9], float Ki[9])
{
    const float fxi = 1.0f / K[0], fyi = 1.0f / K[4];
    Ki[0] = fxi;  Ki[1] = 0.0f; Ki[2] = -K[2] * fxi;
    Ki[3] = 0.0f; Ki[4] = fyi;  Ki[5] = -K[5] * fyi;
    Ki[6] = 0.0f; Ki[7] = 0.0f; Ki[8] = 1.0f;
}

// MeshInstance.hpp:39-46
inline void instance_build_inv(rtb_instance& in)
{
    in.inv_pose = invert_pose(in.pose);
    for (int k = 0; k < 3; k++) in.inv_scale[k] = 1 / in.scale[k];
    in.rotation[0] = in.pose.yaw; in.rotation[1] = in.pose.pitch; in.rotation[2] = in.pose.roll;
    in.inv_rotation[0] = in.inv_pose.yaw; in.inv_rotation[1] = in.inv_pose.pitch; in.inv_rotation[2] = in.inv_pose.roll;
}

// face normal as OBJLoader.hpp:143-147 / TrianglePrimitive.hpp:15-24: normalize(cross(b-a, c-a))
inline V3 face_normal(V3 a, V3 b, V3 c)
{
    const V3 p{ b.x - a.x, b.y - a.y, b.z - a.z }, q{ c.x - a.x, c.y - a.y, c.z - a.z };
    return normalize(V3{ p.y * q.z - p.z * q.y, p.z * q.x - p.x * q.z, p.x * q.y - p.y * q.x });
}

}}  // namespace rtb::hm
