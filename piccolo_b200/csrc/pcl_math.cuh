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

    // -------------------------------------------------------------------------------------------
    // Packed fp32 (sm_100: add/mul/fma.rn.f32x2 -> SASS FADD2 / FMUL2 / FFMA2): the pose kernels' arithmetic on PAIRS of
    // components of ONE instance.  Each lane of a packed instruction is an ordinary IEEE fp32 operation, so a packed chain
    // rounds exactly like the scalar chain it replaces.  The hardware operand forms do the shuffling for free: a pair
    // operand can be a register pair as it lies, the pair swapped, one 32-bit register broadcast to both lanes, and any of
    // them negated - ptxas picks those forms when it sees pk(x, x), pk(hi, lo) and pk(-lo, -hi) (checked in SASS: `.F32`,
    // `.F32x2.LO_HI`, a leading `-`), so none of the pk() calls below costs an instruction as long as the two halves
    // already sit in one aligned register pair (vector loads and packed results do).
    //
    // Layout of a transform in pairs (chosen so that every product of the path is lane-aligned):
    //   pxy = (p.x, p.y)   sxy = (s.x, s.y)   pzsz = (p.z, s.z)   q.wx = (q.w, q.x)   q.yz = (q.y, q.z)
    // -------------------------------------------------------------------------------------------
    typedef unsigned long long f2;
    __device__ __forceinline__ f2 pk(float a, float b)
    {
        f2 r;
        asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
        return r;
    }
    __device__ __forceinline__ float lo(f2 v)
    {
        float a;
        asm("{ .reg .b32 t; mov.b64 {%0, t}, %1; }" : "=f"(a) : "l"(v));
        return a;
    }
    __device__ __forceinline__ float hi(f2 v)
    {
        float b;
        asm("{ .reg .b32 t; mov.b64 {t, %0}, %1; }" : "=f"(b) : "l"(v));
        return b;
    }
    __device__ __forceinline__ f2 bc(float x) { return pk(x, x); }            // broadcast operand
    __device__ __forceinline__ f2 sw(f2 v) { return pk(hi(v), lo(v)); }       // swapped operand
    __device__ __forceinline__ f2 ng(f2 v) { return pk(-lo(v), -hi(v)); }     // negated operand
    __device__ __forceinline__ f2 fma2p(f2 a, f2 b, f2 c)
    {
        f2 d;
        asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
        return d;
    }
    __device__ __forceinline__ f2 mul2p(f2 a, f2 b)
    {
        f2 d;
        asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
        return d;
    }
    __device__ __forceinline__ f2 add2p(f2 a, f2 b)
    {
        f2 d;
        asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
        return d;
    }
    __device__ __forceinline__ f2 sub2p(f2 a, f2 b)
    {
        f2 d;
        asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
        return d;
    }

    struct QuatP
    {
        f2 wx, yz;
    };
    struct TRSP
    {
        f2    pxy, sxy, pzsz;
        QuatP q;
    };

    // Vector3::lerp on a pair, core/math/vector3.h:433: a + t * (b - a)
    __device__ __forceinline__ f2 lerp2(f2 a, f2 b, float t) { return fma2p(sub2p(b, a), bc(t), a); }

    __device__ __forceinline__ float qdotp(const QuatP& a, const QuatP& b)
    {
        const f2 t = fma2p(a.yz, b.yz, mul2p(a.wx, b.wx)); // (w w' + y y', x x' + z z')
        return lo(t) + hi(t);
    }
    // Quaternion::normalise, core/math/quaternion.h:131-135 (factor as in qnormalise above)
    // `skip`: leave the quaternion as it is (factor 1) - how the branch-free hierarchy step reproduces the reference's
    // "a root's derived transform IS its local transform" (node.cpp:61-63) while composing it with an identity parent
    __device__ __forceinline__ QuatP qnormalisep(const QuatP& a, bool skip = false)
    {
        const f2    t    = fma2p(a.yz, a.yz, mul2p(a.wx, a.wx));
        const float len2 = lo(t) + hi(t);
#if PCL_EXACT_NORMALISE
        float factor = 1.0f / sqrtf(len2);
#else
        float factor = rsqrt_newton(len2);
#endif
        factor = skip ? 1.0f : factor;
        return QuatP {mul2p(a.wx, bc(factor)), mul2p(a.yz, bc(factor))};
    }
    // Quaternion::nLerp(t, kp, kq, shortest_path = true), core/math/quaternion.cpp:391-405.  -kq for a negative dot product is
    // folded into the difference: kq * (+-1) is exact, so fma(kq, sg, -kp) rounds like (+-kq) - kp.
    __device__ __forceinline__ QuatP qnlerpp(float t, const QuatP& kp, const QuatP& kq)
    {
        const float sg = qdotp(kp, kq) < 0.0f ? -1.0f : 1.0f;
        const f2    d0 = fma2p(kq.wx, bc(sg), ng(kp.wx)), d1 = fma2p(kq.yz, bc(sg), ng(kp.yz));
        return qnormalisep(QuatP {fma2p(d0, bc(t), kp.wx), fma2p(d1, bc(t), kp.yz)});
    }
    // Quaternion::operator*(const Quaternion&), core/math/quaternion.cpp:13-19, a's components broadcast:
    //   (r.w, r.x) = a.w (b.w, b.x) + a.x (-b.x, b.w) + a.y (-b.y, b.z) + a.z (-b.z, -b.y)
    //   (r.y, r.z) = a.w (b.y, b.z) + a.x (-b.z, b.y) + a.y (b.w, -b.x) + a.z (b.x, b.w)
    // every component accumulates its four products in the order a.w, a.x, a.y, a.z
    __device__ __forceinline__ QuatP qmulp(const QuatP& a, const QuatP& b)
    {
        const float aw = lo(a.wx), ax = hi(a.wx), ay = lo(a.yz), az = hi(a.yz);
        const f2    b0n = pk(lo(b.wx), -hi(b.wx)); // (b.w, -b.x)
        const f2    b1n = pk(-lo(b.yz), hi(b.yz)); // (-b.y, b.z)
        QuatP       r;
        r.wx = mul2p(b.wx, bc(aw));
        r.wx = fma2p(sw(b0n), bc(ax), r.wx);
        r.wx = fma2p(b1n, bc(ay), r.wx);
        r.wx = fma2p(ng(sw(b.yz)), bc(az), r.wx);
        r.yz = mul2p(b.yz, bc(aw));
        r.yz = fma2p(ng(sw(b1n)), bc(ax), r.yz);
        r.yz = fma2p(b0n, bc(ay), r.yz);
        r.yz = fma2p(sw(b.wx), bc(az), r.yz);
        return r;
    }
    // the same product with a CONSTANT left factor whose helper pairs were laid out at upload (StepRec::q01, q23):
    //   c0 = (a.w, a.x)  c1 = (a.y, a.z)  c2 = (-a.x, a.w)  c3 = (-a.z, a.y);  b's components broadcast:
    //   (r.w, r.x) = b.w c0 + b.x c2 - b.y c1 + b.z c3        (r.y, r.z) = b.w c1 - b.x c3 + b.y c0 + b.z c2
    __device__ __forceinline__ QuatP qmul_constp(f2 c0, f2 c1, f2 c2, f2 c3, const QuatP& b)
    {
        const float bw = lo(b.wx), bx = hi(b.wx), by = lo(b.yz), bz = hi(b.yz);
        QuatP       r;
        r.wx = mul2p(c0, bc(bw));
        r.wx = fma2p(c2, bc(bx), r.wx);
        r.wx = fma2p(ng(c1), bc(by), r.wx);
        r.wx = fma2p(c3, bc(bz), r.wx);
        r.yz = mul2p(c1, bc(bw));
        r.yz = fma2p(ng(c3), bc(bx), r.yz);
        r.yz = fma2p(c0, bc(by), r.yz);
        r.yz = fma2p(c2, bc(bz), r.yz);
        return r;
    }
    // Node::updateDerivedTransform, function/animation/node.cpp:32-65, on pairs
    // `root`: the parent is the identity transform and the result must be `local` exactly: every product below is then exact
    // (1 * x, 0 * x, x + 0), only the normalisation would touch the quaternion, so it is switched off
    __device__ __forceinline__ TRSP composep(const TRSP& parent, const TRSP& local, bool root = false)
    {
        TRSP d;
        d.q = qnormalisep(qmulp(parent.q, local.q), root);
        d.sxy = mul2p(parent.sxy, local.sxy);
        const f2 vxy = mul2p(parent.sxy, local.pxy);               // parent.s * local.p
        const f2 zz  = mul2p(local.pzsz, bc(hi(parent.pzsz)));     // (parent.s.z * local.p.z, parent.s.z * local.s.z)
        const float vx = lo(vxy), vy = hi(vxy), vz = lo(zz);
        // Quaternion::operator*(const Vector3&), core/math/quaternion.cpp:269-280: v + 2w (q x v) + 2 (q x (q x v))
        const float qw = lo(parent.q.wx), qx = hi(parent.q.wx), qy = lo(parent.q.yz), qz = hi(parent.q.yz);
        const float ux = qy * vz - qz * vy, uy = qz * vx - qx * vz, uz = qx * vy - qy * vx;
        const float wx_ = qy * uz - qz * uy, wy_ = qz * ux - qx * uz, wz_ = qx * uy - qy * ux;
        const float tw = 2.0f * qw;
        const f2    rxy = fma2p(pk(wx_, wy_), bc(2.0f), fma2p(pk(ux, uy), bc(tw), vxy));
        const float rz  = fmaf(wz_, 2.0f, fmaf(uz, tw, vz));
        d.pxy  = add2p(rxy, parent.pxy);
        d.pzsz = pk(rz + lo(parent.pzsz), hi(zz));
        return d;
    }
    // Transform::getMatrix x inverse bind, as palette_rows above.  Every entry m_rc of the scaled rotation is a scalar that
    // multiplies one row of the inverse bind, so the three products of an output row are packed chains over the constant
    // pairs (ib.x, ib.y), (ib.z, ib.w) with m_rc broadcast: 18 packed instead of 36 scalar multiply-adds, and the results
    // land as the (x, y), (z, w) pairs the 128-bit staging stores want.
    __device__ __forceinline__ void palette_rowsp(const TRSP& d, float4 ib0, float4 ib1, float4 ib2, float4& r0, float4& r1, float4& r2)
    {
        const float qw = lo(d.q.wx), qx = hi(d.q.wx), qy = lo(d.q.yz), qz = hi(d.q.yz);
        const float sx = lo(d.sxy), sy = hi(d.sxy), sz = hi(d.pzsz);
        const float fTx = qx + qx;
        const f2    fTyz = add2p(d.q.yz, d.q.yz);
        const float fTwx = fTx * qw, fTxx = fTx * qx;
        const f2    tw = mul2p(fTyz, bc(qw)), tx = mul2p(fTyz, bc(qx)), ty = mul2p(fTyz, bc(qy)); // (fTwy,fTwz) (fTxy,fTxz) (fTyy,fTyz)
        const float fTwy = lo(tw), fTwz = hi(tw), fTxy = lo(tx), fTxz = hi(tx), fTyy = lo(ty), fTyz_ = hi(ty);
        const float fTzz = hi(fTyz) * qz;
        const float m00 = sx * (1.0f - (fTyy + fTzz)), m01 = sy * (fTxy - fTwz), m02 = sz * (fTxz + fTwy);
        const float m10 = sx * (fTxy + fTwz), m11 = sy * (1.0f - (fTxx + fTzz)), m12 = sz * (fTyz_ - fTwx);
        const float m20 = sx * (fTxz - fTwy), m21 = sy * (fTyz_ + fTwx), m22 = sz * (1.0f - (fTxx + fTyy));
        const f2    a0 = pk(ib0.x, ib0.y), b0 = pk(ib0.z, ib0.w), a1 = pk(ib1.x, ib1.y), b1 = pk(ib1.z, ib1.w);
        const f2    a2 = pk(ib2.x, ib2.y), b2 = pk(ib2.z, ib2.w);
        f2 xy, zw;
        xy = fma2p(a2, bc(m02), fma2p(a1, bc(m01), mul2p(a0, bc(m00))));
        zw = fma2p(b2, bc(m02), fma2p(b1, bc(m01), mul2p(b0, bc(m00))));
        r0 = make_float4(lo(xy), hi(xy), lo(zw), hi(zw) + lo(d.pxy));
        xy = fma2p(a2, bc(m12), fma2p(a1, bc(m11), mul2p(a0, bc(m10))));
        zw = fma2p(b2, bc(m12), fma2p(b1, bc(m11), mul2p(b0, bc(m10))));
        r1 = make_float4(lo(xy), hi(xy), lo(zw), hi(zw) + hi(d.pxy));
        xy = fma2p(a2, bc(m22), fma2p(a1, bc(m21), mul2p(a0, bc(m20))));
        zw = fma2p(b2, bc(m22), fma2p(b1, bc(m21), mul2p(b0, bc(m20))));
        r2 = make_float4(lo(xy), hi(xy), lo(zw), hi(zw) + lo(d.pzsz));
    }
} // namespace pcl
