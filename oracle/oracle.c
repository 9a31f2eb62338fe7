/*
 * TEST INFRASTRUCTURE ONLY.  CPU restatement (plain C, fp32, no FMA contraction) of the reference's
 * runtime animation path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this; the product library never does.
 *
 * PARITY STATUS: pinned.  tests/test_oracle.py checks this file against (a) golden palettes produced
 * by the reference's own Skeleton code (oracle/_ref, built from /root/reference by oracle/Makefile;
 * vectors committed under tests/golden/ with tools/make_golden.py), including the key-count
 * carry-over quirk on a doctored clip, and (b) oracle/_ref live when that library is present.
 * Single-clip results are BIT-EXACT against oracle/_ref on x86-64 (same operation order, no FMA).
 * The multi-clip blend (K > 1) has no executed counterpart in the reference (skeleton.cpp:46 loops
 * "clip_index < 1", weight forced to 1 at :63-64): for K > 1 parity is UNPINNED by construction and
 * the semantics below are this project's documented extension ("A4").
 *
 * All file:line citations are relative to /root/reference/engine/source/runtime.
 */
#include <limits.h>
#include <math.h>
#include <pthread.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { float x, y, z; } v3;
typedef struct { float w, x, y, z; } quat; /* member order of core/math/quaternion.h:19 */

/* ---- core/math restated ------------------------------------------------------------------ */
static v3 v3_add(v3 a, v3 b) { v3 r = {a.x + b.x, a.y + b.y, a.z + b.z}; return r; }
static v3 v3_sub(v3 a, v3 b) { v3 r = {a.x - b.x, a.y - b.y, a.z - b.z}; return r; }
static v3 v3_mul(v3 a, v3 b) { v3 r = {a.x * b.x, a.y * b.y, a.z * b.z}; return r; }
static v3 v3_scale(float s, v3 a) { v3 r = {s * a.x, s * a.y, s * a.z}; return r; }
/* Vector3::lerp, core/math/vector3.h:433 : lhs + alpha * (rhs - lhs) */
static v3 v3_lerp(v3 a, v3 b, float t) { return v3_add(a, v3_scale(t, v3_sub(b, a))); }
/* Vector3::crossProduct, core/math/vector3.h:284-287 */
static v3 v3_cross(v3 a, v3 b)
{
    v3 r = {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
    return r;
}
/* Quaternion::operator*(Quaternion), core/math/quaternion.cpp:13-19 */
static quat q_mul(quat a, quat b)
{
    quat r = {a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
              a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z, a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x};
    return r;
}
static quat q_scale(quat a, float s) { quat r = {a.w * s, a.x * s, a.y * s, a.z * s}; return r; }
static quat q_add(quat a, quat b) { quat r = {a.w + b.w, a.x + b.x, a.y + b.y, a.z + b.z}; return r; }
static quat q_sub(quat a, quat b) { quat r = {a.w - b.w, a.x - b.x, a.y - b.y, a.z - b.z}; return r; }
static quat q_neg(quat a) { quat r = {-a.w, -a.x, -a.y, -a.z}; return r; }
static float q_dot(quat a, quat b) { return a.w * b.w + a.x * b.x + a.y * b.y + a.z * b.z; }
/* Quaternion::normalise, core/math/quaternion.h:131-135 : factor = 1/length; *this = *this * factor */
static quat q_normalise(quat a)
{
    float factor = 1.0f / sqrtf(a.w * a.w + a.x * a.x + a.y * a.y + a.z * a.z);
    return q_scale(a, factor);
}
/* Quaternion::nLerp(t, kp, kq, shortest_path), core/math/quaternion.cpp:391-405 */
static quat q_nlerp(float t, quat kp, quat kq, int shortest)
{
    quat  r;
    float c = q_dot(kp, kq);
    if (c < 0.0f && shortest)
        r = q_add(kp, q_scale(q_sub(q_neg(kq), kp), t));
    else
        r = q_add(kp, q_scale(q_sub(kq, kp), t));
    return q_normalise(r);
}
/* Quaternion::operator*(Vector3), core/math/quaternion.cpp:269-280 ("nVidia SDK implementation") */
static v3 q_rotate(quat q, v3 v)
{
    v3 qv = {q.x, q.y, q.z};
    v3 uv = v3_cross(qv, v), uuv = v3_cross(qv, uv);
    float tw = 2.0f * q.w;
    uv.x *= tw, uv.y *= tw, uv.z *= tw;
    uuv.x *= 2.0f, uuv.y *= 2.0f, uuv.z *= 2.0f;
    return v3_add(v3_add(v, uv), uuv);
}
static int q_isnan(quat q) { return isnan(q.x) || isnan(q.y) || isnan(q.z) || isnan(q.w); }

/* Quaternion::toRotationMatrix(Matrix3x3&) core/math/quaternion.cpp:62-86 then
 * Matrix4x4::makeTransform core/math/matrix4.cpp:82-111 (M[r][c] = scale[c]*R[r][c], M[r][3] = pos[r]) */
static void make_transform(v3 p, v3 s, quat q, float m[16])
{
    float fTx = q.x + q.x, fTy = q.y + q.y, fTz = q.z + q.z;
    float fTwx = fTx * q.w, fTwy = fTy * q.w, fTwz = fTz * q.w;
    float fTxx = fTx * q.x, fTxy = fTy * q.x, fTxz = fTz * q.x;
    float fTyy = fTy * q.y, fTyz = fTz * q.y, fTzz = fTz * q.z;
    float r00 = 1.0f - (fTyy + fTzz), r01 = fTxy - fTwz, r02 = fTxz + fTwy;
    float r10 = fTxy + fTwz, r11 = 1.0f - (fTxx + fTzz), r12 = fTyz - fTwx;
    float r20 = fTxz - fTwy, r21 = fTyz + fTwx, r22 = 1.0f - (fTxx + fTyy);
    m[0] = s.x * r00, m[1] = s.y * r01, m[2] = s.z * r02, m[3] = p.x;
    m[4] = s.x * r10, m[5] = s.y * r11, m[6] = s.z * r12, m[7] = p.y;
    m[8] = s.x * r20, m[9] = s.y * r21, m[10] = s.z * r22, m[11] = p.z;
    m[12] = 0, m[13] = 0, m[14] = 0, m[15] = 1;
}
/* Matrix4x4::concatenate core/math/matrix4.h:289-329 (full 4x4, left-to-right sums) */
static void mat_concat(const float a[16], const float b[16], float r[16])
{
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++)
            r[4 * i + j] = a[4 * i + 0] * b[0 + j] + a[4 * i + 1] * b[4 + j] + a[4 * i + 2] * b[8 + j] +
                           a[4 * i + 3] * b[12 + j];
}

/* ---- asset views (filled from Python via ctypes; plain pointers, caller-owned) ------------ */
typedef struct
{
    int32_t        J;
    const int32_t* index;    /* [J]   RawBone.index                                   */
    const int32_t* parent;   /* [J]   RawBone.parent_index (raw, INT_MAX = none)      */
    const float*   bind_pos; /* [3J]                                                  */
    const float*   bind_rot; /* [4J]  (w,x,y,z) raw                                   */
    const float*   bind_scl; /* [3J]                                                  */
    const float*   inv_bind; /* [16J] RawBone.tpose_matrix, row-major                 */
    int32_t        is_flat, in_topological_order;
} orc_skeleton;

typedef struct
{
    int32_t        total_frame, node_count;
    int32_t        C;           /* node_channels.size()                                */
    int32_t        convert_len; /* AnimSkelMap.convert.size()                          */
    const int32_t* key_counts;  /* [3C] position, rotation, scaling key counts         */
    const int32_t* key_offsets; /* [3C] first key of each track inside pos/rot/scl     */
    const float*   pos;         /* ragged [sum][3]                                     */
    const float*   rot;         /* ragged [sum][4] (w,x,y,z)                           */
    const float*   scl;         /* ragged [sum][3]                                     */
    const int32_t* convert;     /* [convert_len] channel -> bone                       */
} orc_clip;

/* Skeleton::buildSkeleton function/animation/skeleton.cpp:17-37 + find_by_index utilities.cpp:7-27
 * (flat case: key == INT_MAX or key >= i  =>  no parent) + Bone::initialize node.cpp:220-234
 * (setOrientation normalises, NaN -> identity, node.cpp:71-85).  Returns 0, or -1 when the reference
 * silently builds nothing (skeleton.cpp:24-28). */
int orc_build(const orc_skeleton* s, int32_t* parent_out, float* rot_out)
{
    if (!s->is_flat || !s->in_topological_order)
        return -1;
    for (int i = 0; i < s->J; i++)
    {
        int key       = s->parent[i];
        parent_out[i] = (key == INT_MAX || key >= i) ? -1 : key;
        quat q        = {s->bind_rot[4 * i], s->bind_rot[4 * i + 1], s->bind_rot[4 * i + 2], s->bind_rot[4 * i + 3]};
        if (q_isnan(q))
        {
            quat id = {1, 0, 0, 0};
            q       = id;
        }
        else
            q = q_normalise(q);
        rot_out[4 * i] = q.w, rot_out[4 * i + 1] = q.x, rot_out[4 * i + 2] = q.y, rot_out[4 * i + 3] = q.z;
    }
    return 0;
}

/* AnimationManager::getBlendStateWithClipData weights, function/animation/animation_system.cpp:117-149:
 * w[c][b] = enabled[c][b] ? weight[c] / sum_{c' enabled} weight[c'] : 0.   enabled: [K][J] */
void orc_blend_weights(int K, int J, const float* weight, const int32_t* enabled, float* out /*[K][J]*/)
{
    for (int b = 0; b < J; b++)
    {
        float sum = 0;
        for (int c = 0; c < K; c++)
            if (enabled[c * J + b])
                sum += weight[c];
        for (int c = 0; c < K; c++)
            out[c * J + b] = enabled[c * J + b] ? weight[c] / sum : 0.f;
    }
}

/* AnimationComponent::tick phase update, function/framework/component/animation/animation_component.cpp:19-21 */
float orc_advance(float ratio, float dt, float clip_len)
{
    ratio += (dt / clip_len);
    ratio -= floorf(ratio);
    return ratio;
}

typedef struct { v3 p; quat q; v3 s; } trs;

/* One channel sample, skeleton.cpp:75-95.  *hi / *lo carry over between channels (declared outside
 * the channel loop at skeleton.cpp:52-55) -- that is the key-count clamp quirk. */
static void sample_channel(const orc_clip* c, int n, int* lo, int* hi, float r, v3* P, v3* S, quat* Q)
{
    int np = c->key_counts[3 * n], nr = c->key_counts[3 * n + 1], ns = c->key_counts[3 * n + 2];
    if (np <= *hi) *hi = np - 1;
    if (ns <= *hi) *hi = ns - 1;
    if (nr <= *hi) *hi = nr - 1;
    *lo = (*lo < *hi) ? *lo : *hi;
    const float* pk = c->pos + 3 * (size_t)c->key_offsets[3 * n];
    const float* rk = c->rot + 4 * (size_t)c->key_offsets[3 * n + 1];
    const float* sk = c->scl + 3 * (size_t)c->key_offsets[3 * n + 2];
    v3   p0 = {pk[3 * *lo], pk[3 * *lo + 1], pk[3 * *lo + 2]}, p1 = {pk[3 * *hi], pk[3 * *hi + 1], pk[3 * *hi + 2]};
    v3   s0 = {sk[3 * *lo], sk[3 * *lo + 1], sk[3 * *lo + 2]}, s1 = {sk[3 * *hi], sk[3 * *hi + 1], sk[3 * *hi + 2]};
    quat q0 = {rk[4 * *lo], rk[4 * *lo + 1], rk[4 * *lo + 2], rk[4 * *lo + 3]};
    quat q1 = {rk[4 * *hi], rk[4 * *hi + 1], rk[4 * *hi + 2], rk[4 * *hi + 3]};
    *P = v3_lerp(p0, p1, r);
    *S = v3_lerp(s0, s1, r);
    *Q = q_nlerp(r, q0, q1, 1);
}

/*
 * Skeleton::applyAnimation + Skeleton::outputAnimationResult for ONE instance.
 *   K == 1 : exactly function/animation/skeleton.cpp:39-117 (weights/masks ignored: weight forced to 1).
 *   K  > 1 : documented extension A4 -- per bone, blend the clips' sampled deltas with the normalised
 *            per-bone weights wn[c][b] of animation_system.cpp:117-149 (positions/scales: weighted sum;
 *            rotations: sign-aligned to the first contributing clip, weighted sum), then apply the
 *            blended delta with rotate/scale/translate exactly like a single-clip sample (rotate()
 *            normalises).  Bones whose enabled weight sum is < 1e-4, and bones no clip has a channel
 *            for, keep the bind pose.  A clip without a channel for a bone that another clip does
 *            animate contributes the identity delta.
 * Outputs: mats [J][16] row-major (AnimationResultElement.transform), index [J] (= RawBone.index + 1),
 *          derived [J][10] = P.xyz, Q.wxyz, S.xyz (may be NULL).
 * parent / rot_norm come from orc_build.  Returns 0.
 */
int orc_pose(const orc_skeleton* s, const int32_t* parent, const float* rot_norm, int K, const orc_clip* const* clips,
             const float* phase, const float* weight, const int32_t* enabled /*[K][J] or NULL*/, float* mats,
             int32_t* index, float* derived)
{
    int  J   = s->J;
    trs* loc = (trs*)malloc(sizeof(trs) * (size_t)J);
    trs* der = (trs*)malloc(sizeof(trs) * (size_t)J);
    /* resetSkeleton -> Node::resetToInitialPose, skeleton.cpp:45, node.cpp:197-206 */
    for (int i = 0; i < J; i++)
    {
        v3   p = {s->bind_pos[3 * i], s->bind_pos[3 * i + 1], s->bind_pos[3 * i + 2]};
        v3   c = {s->bind_scl[3 * i], s->bind_scl[3 * i + 1], s->bind_scl[3 * i + 2]};
        quat q = {rot_norm[4 * i], rot_norm[4 * i + 1], rot_norm[4 * i + 2], rot_norm[4 * i + 3]};
        loc[i].p = p, loc[i].q = q, loc[i].s = c;
    }
    if (K == 1)
    {
        const orc_clip* c     = clips[0];
        float           exact = phase[0] * (c->total_frame - 1);
        int             lo = (int)floorf(exact), hi = (int)ceilf(exact);
        float           r = exact - lo;
        for (int n = 0; n < c->node_count && n < c->convert_len; n++)
        {
            int bone = c->convert[n];
            if (bone == -1) /* size_t(-1) == numeric_limits<size_t>::max(), skeleton.cpp:69-73 */
                continue;
            v3 P, S;
            quat Q;
            sample_channel(c, n, &lo, &hi, r, &P, &S, &Q);
            /* bone->rotate(LOCAL) node.cpp:135-157 ; scale :182-186 ; translate(AREN) :107-132 */
            loc[bone].q = q_mul(loc[bone].q, q_normalise(Q));
            loc[bone].s = v3_mul(loc[bone].s, S);
            loc[bone].p = v3_add(loc[bone].p, P);
        }
    }
    else
    {
        float* wn   = (float*)malloc(sizeof(float) * (size_t)K * J);
        float* wsum = (float*)calloc((size_t)J, sizeof(float));
        trs*   acc  = (trs*)calloc((size_t)J, sizeof(trs));
        quat*  qref = (quat*)calloc((size_t)J, sizeof(quat));
        char*  any  = (char*)calloc((size_t)J, 1);
        char*  tch  = (char*)calloc((size_t)J, 1); /* some clip has a channel for this bone */
        trs*   dl   = (trs*)malloc(sizeof(trs) * (size_t)J);
        char*  has  = (char*)malloc((size_t)J);
        for (int b = 0; b < J; b++)
        {
            float sum = 0;
            for (int c = 0; c < K; c++)
                if (!enabled || enabled[c * J + b])
                    sum += weight[c];
            wsum[b] = sum;
            for (int c = 0; c < K; c++)
                wn[c * J + b] = (!enabled || enabled[c * J + b]) ? weight[c] / sum : 0.f;
        }
        for (int ci = 0; ci < K; ci++)
        {
            const orc_clip* c     = clips[ci];
            float           exact = phase[ci] * (c->total_frame - 1);
            int             lo = (int)floorf(exact), hi = (int)ceilf(exact);
            float           r = exact - lo;
            memset(has, 0, (size_t)J);
            for (int n = 0; n < c->node_count && n < c->convert_len; n++)
            {
                int bone = c->convert[n];
                if (bone == -1)
                    continue;
                v3 P, S;
                quat Q;
                sample_channel(c, n, &lo, &hi, r, &P, &S, &Q);
                if (!has[bone])
                    dl[bone].p = P, dl[bone].s = S, dl[bone].q = Q, has[bone] = 1, tch[bone] = 1;
                else
                    dl[bone].p = v3_add(dl[bone].p, P), dl[bone].s = v3_mul(dl[bone].s, S), dl[bone].q = q_mul(dl[bone].q, Q);
            }
            for (int b = 0; b < J; b++)
            {
                float w = wn[ci * J + b];
                if (!(fabsf(wsum[b]) >= 0.0001f) || w == 0.f)
                    continue;
                trs d;
                if (has[b])
                    d = dl[b];
                else
                {
                    v3 z = {0, 0, 0}, o = {1, 1, 1};
                    quat id = {1, 0, 0, 0};
                    d.p = z, d.s = o, d.q = id;
                }
                if (!any[b])
                {
                    qref[b]  = d.q;
                    acc[b].p = v3_scale(w, d.p), acc[b].s = v3_scale(w, d.s), acc[b].q = q_scale(d.q, w);
                    any[b]   = 1;
                }
                else
                {
                    float sg = (q_dot(d.q, qref[b]) < 0.0f) ? -w : w;
                    acc[b].p = v3_add(acc[b].p, v3_scale(w, d.p));
                    acc[b].s = v3_add(acc[b].s, v3_scale(w, d.s));
                    acc[b].q = q_add(acc[b].q, q_scale(d.q, sg));
                }
            }
        }
        for (int b = 0; b < J; b++)
            if (any[b] && tch[b]) /* bones no clip animates keep the bind pose exactly */
            {
                loc[b].q = q_mul(loc[b].q, q_normalise(acc[b].q));
                loc[b].s = v3_mul(loc[b].s, acc[b].s);
                loc[b].p = v3_add(loc[b].p, acc[b].p);
            }
        free(wn), free(wsum), free(acc), free(qref), free(any), free(tch), free(dl), free(has);
    }
    /* m_bones[i].update() in index order, skeleton.cpp:110-113 -> Node::updateDerivedTransform node.cpp:32-65 */
    for (int i = 0; i < J; i++)
    {
        int p = parent[i];
        if (p >= 0)
        {
            der[i].q = q_normalise(q_mul(der[p].q, loc[i].q));
            der[i].s = v3_mul(der[p].s, loc[i].s);
            der[i].p = v3_add(q_rotate(der[p].q, v3_mul(der[p].s, loc[i].p)), der[p].p);
        }
        else
            der[i] = loc[i];
    }
    /* outputAnimationResult skeleton.cpp:119-156 */
    for (int i = 0; i < J; i++)
    {
        float obj[16];
        make_transform(der[i].p, der[i].s, der[i].q, obj);
        if (mats)
            mat_concat(obj, s->inv_bind + 16 * (size_t)i, mats + 16 * (size_t)i);
        if (index)
            index[i] = s->index[i] + 1; /* getID()+1, skeleton.cpp:127, node.cpp:237-244 */
        if (derived)
        {
            float* o = derived + 10 * (size_t)i;
            o[0] = der[i].p.x, o[1] = der[i].p.y, o[2] = der[i].p.z;
            o[3] = der[i].q.w, o[4] = der[i].q.x, o[5] = der[i].q.y, o[6] = der[i].q.z;
            o[7] = der[i].s.x, o[8] = der[i].s.y, o[9] = der[i].s.z;
        }
    }
    free(loc), free(der);
    return 0;
}

/* N instances split over n_threads pthreads.  clip_ids/phase/weight: [N][K]; mask_ids: [N][K] rows of
 * `masks` ([n_masks][J]) or NULL (all enabled).  mats: [N][J][16]; derived: [N][J][10] or NULL. */
typedef struct
{
    const orc_skeleton*    s;
    const orc_clip* const* all_clips;
    const int32_t *        masks, *clip_ids, *mask_ids, *parent;
    const float *          phase, *weight, *rotn;
    float *                mats, *derived;
    int                    K, i0, i1;
} batch_job;

static void* batch_worker(void* arg)
{
    batch_job* j = (batch_job*)arg;
    int        J = j->s->J, K = j->K;
    int32_t*   en = (j->mask_ids && j->masks) ? (int32_t*)malloc(sizeof(int32_t) * (size_t)K * J) : NULL;
    for (int i = j->i0; i < j->i1; i++)
    {
        const orc_clip* cl[16];
        for (int k = 0; k < K && k < 16; k++)
            cl[k] = j->all_clips[j->clip_ids[(size_t)i * K + k]];
        if (en)
            for (int k = 0; k < K; k++)
                memcpy(en + (size_t)k * J, j->masks + (size_t)j->mask_ids[(size_t)i * K + k] * J, sizeof(int32_t) * (size_t)J);
        orc_pose(j->s, j->parent, j->rotn, K, cl, j->phase + (size_t)i * K, j->weight ? j->weight + (size_t)i * K : NULL,
                 en, j->mats + (size_t)i * J * 16, NULL, j->derived ? j->derived + (size_t)i * J * 10 : NULL);
    }
    free(en);
    return NULL;
}

int orc_pose_batch(const orc_skeleton* s, int n_clips, const orc_clip* const* all_clips, int n_masks, const int32_t* masks,
                   int N, int K, const int32_t* clip_ids, const float* phase, const float* weight,
                   const int32_t* mask_ids, float* mats, float* derived, int n_threads)
{
    int      J      = s->J;
    int32_t* parent = (int32_t*)malloc(sizeof(int32_t) * (size_t)J);
    float*   rotn   = (float*)malloc(sizeof(float) * 4 * (size_t)J);
    if (K > 16 || orc_build(s, parent, rotn) != 0)
    {
        free(parent), free(rotn);
        return -1;
    }
    (void)n_clips, (void)n_masks;
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    pthread_t th[256];
    batch_job jobs[256];
    for (int t = 0; t < n_threads; t++)
    {
        batch_job jb = {s, all_clips, masks, clip_ids, mask_ids, parent, phase, weight, rotn, mats, derived, K,
                        (int)((long long)N * t / n_threads), (int)((long long)N * (t + 1) / n_threads)};
        jobs[t] = jb;
        pthread_create(&th[t], NULL, batch_worker, &jobs[t]);
    }
    for (int t = 0; t < n_threads; t++)
        pthread_join(th[t], NULL);
    free(parent), free(rotn);
    return 0;
}

/* ---- skinning: engine/shader/glsl/mesh.vert:72-111 ---------------------------------------- */
/* Binding weights renormalised as RenderResource::updateVertexBuffer does,
 * function/render/render_resource.cpp:714-725 (sum == 0 -> factor 1). */
void orc_normalise_weights(int V, const float* w_raw, float* w_out)
{
    for (int v = 0; v < V; v++)
    {
        const float* w   = w_raw + 4 * (size_t)v;
        float        inv = w[0] + w[1] + w[2] + w[3];
        inv              = (inv != 0.0f) ? 1 / inv : 1.0f;
        for (int k = 0; k < 4; k++)
            w_out[4 * (size_t)v + k] = w[k] * inv;
    }
}

/* palette16: [n_slots][16] row-major with slot 0 = identity (mesh_component.cpp:72-79); w4 already
 * renormalised.  B accumulates from a zero matrix in x,y,z,w order, each influence gated by
 * (w > 0 && idx > 0) (mesh.vert:77-102); p' = (B*[p,1]).xyz (:104); n' = normalize(B3x3*n) (:106-110). */
void orc_lbs(int V, const float* pos, const float* nrm, const float* tan, const int32_t* idx4, const float* w4,
             const float* palette16, float* out_pos, float* out_nrm, float* out_tan)
{
    for (int v = 0; v < V; v++)
    {
        float B[16];
        memset(B, 0, sizeof(B));
        for (int k = 0; k < 4; k++)
        {
            int   id = idx4[4 * (size_t)v + k];
            float w  = w4[4 * (size_t)v + k];
            if (w > 0.0f && id > 0)
            {
                const float* m = palette16 + 16 * (size_t)id;
                for (int e = 0; e < 16; e++)
                    B[e] = B[e] + m[e] * w;
            }
        }
        const float* p = pos + 3 * (size_t)v;
        for (int r = 0; r < 3; r++)
            out_pos[3 * (size_t)v + r] = B[4 * r] * p[0] + B[4 * r + 1] * p[1] + B[4 * r + 2] * p[2] + B[4 * r + 3] * 1.0f;
        for (int pass = 0; pass < 2; pass++)
        {
            const float* in  = pass ? tan : nrm;
            float*       out = pass ? out_tan : out_nrm;
            if (!in || !out)
                continue;
            const float* a = in + 3 * (size_t)v;
            float        t[3];
            for (int r = 0; r < 3; r++)
                t[r] = B[4 * r] * a[0] + B[4 * r + 1] * a[1] + B[4 * r + 2] * a[2];
            float l = sqrtf(t[0] * t[0] + t[1] * t[1] + t[2] * t[2]);
            out[3 * (size_t)v] = t[0] / l, out[3 * (size_t)v + 1] = t[1] / l, out[3 * (size_t)v + 2] = t[2] / l;
        }
    }
}

/* Slot layout the renderer consumes: slot 0 = identity, slot index[b] = mats[b]
 * (MeshComponent::tick mesh_component.cpp:72-79 pushes results in order, i.e. slot b+1 == index[b]
 * for the flat skeletons the reference accepts). */
void orc_palette_slots(int J, const float* mats, float* palette16 /*[J+1][16]*/)
{
    static const float I[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    memcpy(palette16, I, sizeof(I));
    memcpy(palette16 + 16, mats, sizeof(float) * 16 * (size_t)J);
}
