// Device restatement of the subset of Piccolo's core/math the animation path uses.
// fp32 throughout, same operation order as the reference so that the only numerical difference is
// FMA contraction (the reference's x86-64 build has none; nvcc contracts a*b+c).
// Citations are relative to /root/reference/engine/source/runtime.
#pragma once
#include <cuda_runtime.h>

namespace pcl
{
    struct Quat // (w,x,y,z), core/math/quaternion.h:19
    {
        float w, x, y, z;
    };
    struct TRS
    {
        float3 p;
        Quat   q;
        float3 s;
    };

    __device__ __forceinline__ float3 operator+(float3 a, float3 b) { return make_float3(a.x + b.x, a.y + b.y, a.z + b.z); }
    __device__ __forceinline__ float3 operator-(float3 a, float3 b) { return make_float3(a.x - b.x, a.y - b.y, a.z - b.z); }
    __device__ __forceinline__ float3 operator*(float3 a, float3 b) { return make_float3(a.x * b.x, a.y * b.y, a.z * b.z); }
    __device__ __forceinline__ float3 operator*(float s, float3 a) { return make_float3(s * a.x, s * a.y, s * a.z); }

    // Vector3::lerp, core/math/vector3.h:433
    __device__ __forceinline__ float3 lerp3(float3 a, float3 b, float t) { return a + t * (b - a); }
    // Vector3::crossProduct, core/math/vector3.h:284-287
    __device__ __forceinline__ float3 cross3(float3 a, float3 b)
    {
        return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
    }
    // Quaternion::operator*(const Quaternion&), core/math/quaternion.cpp:13-19
    __device__ __forceinline__ Quat qmul(Quat a, Quat b)
    {
        Quat r;
        r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
        r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
        r.y = a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z;
        r.z = a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x;
        return r;
    }
    __device__ __forceinline__ Quat  qscale(Quat a, float s) { return Quat {a.w * s, a.x * s, a.y * s, a.z * s}; }
    __device__ __forceinline__ float qdot(Quat a, Quat b) { return a.w * b.w + a.x * b.x + a.y * b.y + a.z * b.z; }
    // Quaternion::normalise, core/math/quaternion.h:131-135: factor = 1 / length; q = q * factor.
    // The reference's factor carries two roundings (sqrt, then the division).  IEEE sqrtf and division each compile to a fast
    // path plus a branch to a slow-path subroutine; three normalisations per bone put six such branches into every bone of the
    // pose kernels and cut their instruction streams into blocks ptxas cannot schedule across.  PCL_EXACT_NORMALISE = 0 (default)
    // takes MUFU.RSQ and one Newton step instead: branch-free, within 1 ulp of 1/sqrt(x), i.e. within 2 ulp of the reference's
    // factor - the same size as the FMA-contraction differences the parity tolerance (1e-5 scaled) already covers.
#ifndef PCL_EXACT_NORMALISE
#define PCL_EXACT_NORMALISE 0
#endif
    __device__ __forceinline__ float rsqrt_newton(float x)
    {
        float y;
        asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
        const float e = fmaf(-x * y, 0.5f * y, 0.5f); // 0.5 (1 - x y^2)
        return fmaf(y, e, y);
    }
    // a / b without the slow-path branch of the IEEE division: reciprocal estimate, one Newton step, one residual correction
    // (correctly rounded for the blend weights it is used on: normal range, b >= 1e-4).
    __device__ __forceinline__ float fdiv_nobranch(float a, float b)
    {
        float r;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(b));
        r             = fmaf(fmaf(-b, r, 1.0f), r, r);
        const float q = a * r;
        return fmaf(fmaf(-b, q, a), r, q);
    }
    __device__ __forceinline__ Quat qnormalise(Quat a)
    {
        const float len2 = a.w * a.w + a.x * a.x + a.y * a.y + a.z * a.z;
#if PCL_EXACT_NORMALISE
        const float factor = 1.0f / sqrtf(len2);
#else
        const float factor = rsqrt_newton(len2);
#endif
        return qscale(a, factor);
    }
    // Quaternion::nLerp(t, kp, kq, shortest_path = true), core/math/quaternion.cpp:391-405
    __device__ __forceinline__ Quat qnlerp(float t, Quat kp, Quat kq)
    {
        float c = qdot(kp, kq);
        if (c < 0.0f)
            kq = Quat {-kq.w, -kq.x, -kq.y, -kq.z};
        Quat r {kp.w + (kq.w - kp.w) * t, kp.x + (kq.x - kp.x) * t, kp.y + (kq.y - kp.y) * t, kp.z + (kq.z - kp.z) * t};
        return qnormalise(r);
    }
    // Quaternion::operator*(const Vector3&), core/math/quaternion.cpp:269-280
    __device__ __forceinline__ float3 qrotate(Quat q, float3 v)
    {
        float3 qv  = make_float3(q.x, q.y, q.z);
        float3 uv  = cross3(qv, v);
        float3 uuv = cross3(qv, uv);
        uv         = (2.0f * q.w) * uv;
        uuv        = 2.0f * uuv;
        return (v + uv) + uuv;
    }

    // Node::updateDerivedTransform, function/animation/node.cpp:32-65 (TRS composition, no shear)
    __device__ __forceinline__ TRS compose(const TRS& parent, const TRS& local)
    {
        TRS d;
        d.q = qnormalise(qmul(parent.q, local.q));
        d.s = parent.s * local.s;
        d.p = qrotate(parent.q, parent.s * local.p) + parent.p;
        return d;
    }

    // Transform::getMatrix -> Matrix4x4::makeTransform (core/math/matrix4.cpp:82-111) with
    // Quaternion::toRotationMatrix (core/math/quaternion.cpp:62-86), then
    // objMat * inverseTpose (Matrix4x4::concatenate, core/math/matrix4.h:289-329) restricted to the three
    // affine rows: ib row 3 is (0,0,0,1) (checked at upload), so the dropped products are exact zeros.
    __device__ __forceinline__ void palette_rows(const TRS& d, float4 ib0, float4 ib1, float4 ib2, float4& r0, float4& r1,
                                                 float4& r2)
    {
        const Quat& q   = d.q;
        float       fTx = q.x + q.x, fTy = q.y + q.y, fTz = q.z + q.z;
        float       fTwx = fTx * q.w, fTwy = fTy * q.w, fTwz = fTz * q.w;
        float       fTxx = fTx * q.x, fTxy = fTy * q.x, fTxz = fTz * q.x;
        float       fTyy = fTy * q.y, fTyz = fTz * q.y, fTzz = fTz * q.z;
        float       m00 = d.s.x * (1.0f - (fTyy + fTzz)), m01 = d.s.y * (fTxy - fTwz), m02 = d.s.z * (fTxz + fTwy);
        float       m10 = d.s.x * (fTxy + fTwz), m11 = d.s.y * (1.0f - (fTxx + fTzz)), m12 = d.s.z * (fTyz - fTwx);
        float       m20 = d.s.x * (fTxz - fTwy), m21 = d.s.y * (fTyz + fTwx), m22 = d.s.z * (1.0f - (fTxx + fTyy));
        r0 = make_float4(m00 * ib0.x + m01 * ib1.x + m02 * ib2.x, m00 * ib0.y + m01 * ib1.y + m02 * ib2.y,
                         m00 * ib0.z + m01 * ib1.z + m02 * ib2.z, m00 * ib0.w + m01 * ib1.w + m02 * ib2.w + d.p.x);
        r1 = make_float4(m10 * ib0.x + m11 * ib1.x + m12 * ib2.x, m10 * ib0.y + m11 * ib1.y + m12 * ib2.y,
                         m10 * ib0.z + m11 * ib1.z + m12 * ib2.z, m10 * ib0.w + m11 * ib1.w + m12 * ib2.w + d.p.y);
        r2 = make_float4(m20 * ib0.x + m21 * ib1.x + m22 * ib2.x, m20 * ib0.y + m21 * ib1.y + m22 * ib2.y,
                         m20 * ib0.z + m21 * ib1.z + m22 * ib2.z, m20 * ib0.w + m21 * ib1.w + m22 * ib2.w + d.p.z);
    }
} // namespace pcl
