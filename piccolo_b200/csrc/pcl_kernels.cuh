// Hand-written sm_100a kernels of the animation + skinning hot path.
//   pose_kernel<K>   K1-K4 fused: keyframe sampling, (blend,) bind-relative local pose, hierarchy, palette
//   lbs_kernel<..>   K5: 4-weight linear-blend skinning, TMA-staged streams
//   export_kernel    compact 3x4 palette -> AnimationResult 4x4 (readback path)
//   advance_kernel   AnimationComponent::tick phase update
// Citations are relative to /root/reference/engine/source/runtime.
#pragma once
#include "pcl_math.cuh"

#include <cuda.h> // CUtensorMap (type only; the encoder is fetched through cudaGetDriverEntryPoint)

#include <cstdint>
#include <type_traits>

namespace pcl
{
    // -------------------------------------------------------------------------------------------
    // device-side asset records
    // -------------------------------------------------------------------------------------------
    // One record per processing step (= one bone), in dependency order.  Everything in it is uniform
    // across a warp, so it is fetched with broadcast loads.
    struct alignas(16) StepRec
    {
        int32_t joint; // bone position in bones_map; palette slot = joint + 1
        int32_t psrc;  // parent's derived TRS: -1 root, -2 result of the previous step (still in registers), >= 0 smem slot
        int32_t store; // smem slot to keep this bone's derived TRS in for later children, or -1
        int32_t pstep; // processing step of the parent (-1 root): pose_level_kernel's parent lookup.  In the records of a
                       // half (pose_simple_kernel<SPLIT>): this bone's step in the WHOLE skeleton = its row in the key tables
        // bind TRS (Node::m_initial_*) laid out for the packed arithmetic of pcl_math.cuh:
        float4  bA;    // (p.x, p.y, s.x, s.y)
        float4  bC;    // (1, s.z, p.z, 0): local (p.z, s.z) = sampled (p.z, s.z) * (1, s.z) + (p.z, 0) in one packed fma
        float4  q01;   // bind rotation (w, x, y, z), normalised as Node::setOrientation does
        float4  q23;   // (-x, w, -z, y): helper pairs of the quaternion product with a constant left factor
        float4  ib0, ib1, ib2;     // inverse bind rows 0..2 (Node::m_inverse_Tpose)
    };

    // One uploaded clip+map.  keys: [C][F][3] float4 = {pos.xyz_, rot.wxyz, scl.xyz_} per key, tracks
    // zero-padded to F keys (the clamp below never reads the padding).
    struct alignas(16) ClipRec
    {
        const float4*  keys;
        const int32_t* first_chan; // [J]  first played channel driving this bone, or -1
        const int32_t* next_chan;  // [C]  next played channel driving the same bone, or -1
        const int32_t* hi_clamp;   // [C]  prefix-min over played channels 0..c of (shortest track length - 1):
                                   //      the data-parallel form of the carry-over clamp at skeleton.cpp:75-87
        int32_t F;
        int32_t total_frame_m1;    // total_frame - 1
        int32_t C;
        int32_t simple;            // every bone is driven by at most one played channel: the step-ordered form below exists
        // step-ordered form (pose_simple_kernel): one track of exactly F keys per processing step - the identity track for
        // bones no channel drives, and every track CONTINUED WITH COPIES OF ITS KEY hi_clamp up to F keys.  Reading key
        // min(frame, F - 1) of such a track returns what the clamp of skeleton.cpp:75-87 returns (beyond the clamp lo and hi
        // name the same key, and interpolating a key with its copy gives the key back exactly), so the two key indices
        // are computed ONCE per instance and a key's address depends on (step, instance) only - no clamp table on the hot path
        const float4*   keys_step;  // three tables, [Jp][F] each: a (16 B) | q (16 B) | c (8 B), see KeyPair
        const uint32_t* anim_step;  // [ceil(Jp / 32)] bit s: a played channel drives step s (blend path: bones no clip animates keep the bind pose)
        int32_t        bucket_base; // first (clip, frame) bucket of this clip in the context-wide numbering (order_* kernels)
        int32_t        pad_[3];
    };

    struct PoseParams
    {
        const StepRec* steps;
        const ClipRec* clips;
        const int32_t* masks;    // [n_masks][J]
        const int32_t* clip_ids; // [K][N]
        float*         phase;    // [K][N]
        const float*   weight;   // [K][N]
        const float*   clip_len; // [K][N]
        const int32_t* mask_ids; // [K][N]  (-1 = all enabled)
        float4*        palette;  // [N][pitch][3]
        float*         derived;  // [N][J][10] or nullptr
        float4*        out12;    // [N][J][3] rows 0..2 in bones_map order, or nullptr: pose_level_kernel writes the AnimationResult
                                 // rows here as well (a pinned host block for pcl_batch_frame: no export kernel, no DMA)
        const int32_t* sched;    // [n_rounds][POSE_L_WARPS] step index or -1 (pose_level_kernel)
        const int32_t* order;    // [N] visiting order (frame-coherent, see order_* kernels) or nullptr = 0, 1, 2, ...
        long long      N;
        int32_t        J, pitch, nslot, n_steps_padded, n_rounds;
        int32_t        split_base[2], split_n[2]; // pose_simple_kernel<SPLIT>: where each half's step records start in c_steps, how many
        float*         phase_out;                 // pose_simple_kernel<SPLIT> with a phase advance: where the advanced phases go
        float          dt;
        int32_t        advance; // 0, PCL_ADVANCE_FIRST_ONLY, PCL_ADVANCE_ALL
    };

    constexpr int POSE_G = 4; // bones staged in shared memory between two coalesced palette flushes

    // per-warp dynamic shared memory, in floats
    constexpr int POSE_SLOT_FLOATS = 10 * 32; // one parked derived TRS: 5 pairs x 32 lanes (pose_simple / pose_level: [pair][lane] of 8 bytes)
    // nslot parked parents + one slot holding the identity transform (the "parent" of a root in pose_simple_kernel)
    __host__ __device__ inline size_t pose_smem_floats_per_warp(int nslot) { return (size_t)(nslot + 1) * POSE_SLOT_FLOATS + 32 * (POSE_G * 12 + 4); }

    // frame-coherent one-clip launches: a warp whose 32 instances fall into ONE (clip, frame) bucket - 98 % of the warps of a large
    // sorted crowd - copies that bucket's two keys of every bone into shared memory once (J x 96 bytes) and reads them back as
    // broadcasts: no global load is left in its bone loop
    __host__ __device__ inline size_t pose_wkeys_floats(int J) { return (size_t)J * 24; }

    struct SlotState // one clip slot of one instance, resident in registers
    {
        const float4*  keys;
        const int32_t* first;
        const int32_t* next;
        const int32_t* hic;
        int32_t        F, lo, hi, mask_off;
        float          r, w;
    };

    // skeleton.cpp:75-95 for one channel of one instance
    __device__ __forceinline__ void sample_channel(const SlotState& st, int chan, float3& P, float3& S, Quat& Q)
    {
        int hc = __ldg(st.hic + chan);
        int h  = max(min(st.hi, hc), 0);
        int l  = max(min(st.lo, h), 0);
        const float4* kl = st.keys + ((size_t)chan * st.F + l) * 3;
        const float4* kh = st.keys + ((size_t)chan * st.F + h) * 3;
        float4 p0 = __ldg(kl), q0 = __ldg(kl + 1), s0 = __ldg(kl + 2);
        float4 p1 = __ldg(kh), q1 = __ldg(kh + 1), s1 = __ldg(kh + 2);
        P = lerp3(make_float3(p0.x, p0.y, p0.z), make_float3(p1.x, p1.y, p1.z), st.r);
        S = lerp3(make_float3(s0.x, s0.y, s0.z), make_float3(s1.x, s1.y, s1.z), st.r);
        Q = qnlerp(st.r, Quat {q0.x, q0.y, q0.z, q0.w}, Quat {q1.x, q1.y, q1.z, q1.w});
    }

    // -------------------------------------------------------------------------------------------
    // pose kernel: lane = instance, the warp walks the bones in dependency order.
    //   - clip keys: 128-bit read-only loads (the whole clip is L1/L2 resident, shared by all instances)
    //   - parent lookups: registers when the parent was the previous step, else a per-warp shared-memory
    //     slot [slot][component][lane] (conflict-free); slots are recycled once a bone's last child is done
    //   - palette rows are staged [lane][bone] in shared memory and flushed as coalesced 128-bit stores
    //     (a run of POSE_G consecutive slots of one instance is contiguous in HBM)
    // -------------------------------------------------------------------------------------------
    template<int K>
    __global__ void __launch_bounds__(256) pose_kernel(const PoseParams par)
    {
        extern __shared__ __align__(16) float smem[];
        const int       lane  = threadIdx.x & 31;
        const int       wib   = threadIdx.x >> 5;
        const long long wglob = (long long)blockIdx.x * (blockDim.x >> 5) + wib;
        const long long inst0 = wglob * 32;
        if (inst0 >= par.N)
            return;
        const long long inst   = min(inst0 + lane, par.N - 1); // idle lanes shadow the last instance, never store
        const bool      active = (inst0 + lane) < par.N;
        float*          slots  = smem + (size_t)wib * pose_smem_floats_per_warp(par.nslot);
        float*          stage  = slots + (size_t)(par.nslot + 1) * POSE_SLOT_FLOATS;
        constexpr int   SROW   = POSE_G * 12 + 4; // floats per staged instance row (+4: conflict-free float4 columns)

        // ---- per-instance clip slot state (A1 + head of A3) -----------------------------------
        SlotState st[K];
#pragma unroll
        for (int k = 0; k < K; k++)
        {
            const size_t  o  = (size_t)k * par.N + inst;
            const ClipRec cr = par.clips[__ldg(par.clip_ids + o)];
            float         ph = par.phase[o];
            if (par.advance == 2 || (par.advance == 1 && k == 0))
            {
                // AnimationComponent::tick, function/framework/component/animation/animation_component.cpp:19-21
                ph += par.dt / __ldg(par.clip_len + o);
                ph -= floorf(ph);
                if (active)
                    par.phase[o] = ph;
            }
            // skeleton.cpp:52-55 (computed once per clip, before any clamping)
            float exact = ph * (float)cr.total_frame_m1;
            st[k].lo    = (int)floorf(exact);
            st[k].hi    = (int)ceilf(exact);
            st[k].r     = exact - (float)st[k].lo;
            st[k].keys = cr.keys, st[k].first = cr.first_chan, st[k].next = cr.next_chan, st[k].hic = cr.hi_clamp;
            st[k].F = cr.F;
            st[k].w = (K > 1) ? __ldg(par.weight + o) : 1.0f;
            int mid = (K > 1 && par.mask_ids) ? __ldg(par.mask_ids + o) : -1;
            st[k].mask_off = mid < 0 ? -1 : mid * par.J;
        }

        // slot 0 of the palette is the identity (mesh_component.cpp:72-79)
        if (active)
        {
            float4* p0 = par.palette + (size_t)inst * par.pitch * 3;
            p0[0] = make_float4(1, 0, 0, 0), p0[1] = make_float4(0, 1, 0, 0), p0[2] = make_float4(0, 0, 1, 0);
        }

        TRS prev; // derived TRS of the previous step
        prev.p = make_float3(0, 0, 0), prev.q = Quat {1, 0, 0, 0}, prev.s = make_float3(1, 1, 1);

        for (int s0 = 0; s0 < par.n_steps_padded; s0 += POSE_G)
        {
#pragma unroll
            for (int g = 0; g < POSE_G; g++)
            {
                const int s = s0 + g;
                if (s >= par.J)
                    break;
                const StepRec* rec  = par.steps + s;
                const int4     hdr  = __ldg(reinterpret_cast<const int4*>(rec));
                const int      bone = hdr.x;
                const float4   bA = __ldg(&rec->bA), bC = __ldg(&rec->bC), br = __ldg(&rec->q01);
                // resetSkeleton -> Node::resetToInitialPose, skeleton.cpp:45, node.cpp:197-206
                TRS loc;
                loc.p = make_float3(bA.x, bA.y, bC.z), loc.q = Quat {br.x, br.y, br.z, br.w}, loc.s = make_float3(bA.z, bA.w, bC.y);

                if (K == 1)
                {
                    // the reference's executed path: every played channel of this bone, in channel order
                    // (skeleton.cpp:57-105); rotate(LOCAL) node.cpp:135-157, scale :182-186, translate(AREN) :107-132
                    int chan = __ldg(st[0].first + bone);
                    while (chan >= 0)
                    {
                        float3 P, S;
                        Quat   Q;
                        sample_channel(st[0], chan, P, S, Q);
                        loc.q = qmul(loc.q, qnormalise(Q));
                        loc.s = loc.s * S;
                        loc.p = loc.p + P;
                        chan  = __ldg(st[0].next + chan);
                    }
                }
                else
                {
                    // documented blend extension; weights as animation_system.cpp:117-149
                    float wsum = 0.f;
                    bool  en[K];
#pragma unroll
                    for (int k = 0; k < K; k++)
                    {
                        en[k] = st[k].mask_off < 0 ? true : (__ldg(par.masks + st[k].mask_off + bone) != 0);
                        if (en[k])
                            wsum += st[k].w;
                    }
                    TRS  acc;
                    Quat qref {1, 0, 0, 0};
                    bool any = false, touched = false;
                    acc.p = make_float3(0, 0, 0), acc.q = Quat {0, 0, 0, 0}, acc.s = make_float3(0, 0, 0);
#pragma unroll
                    for (int k = 0; k < K; k++)
                    {
                        float wn   = en[k] ? st[k].w / wsum : 0.f;
                        int   chan = __ldg(st[k].first + bone);
                        bool  has  = false;
                        TRS   d;
                        d.p = make_float3(0, 0, 0), d.q = Quat {1, 0, 0, 0}, d.s = make_float3(1, 1, 1);
                        while (chan >= 0)
                        {
                            float3 P, S;
                            Quat   Q;
                            sample_channel(st[k], chan, P, S, Q);
                            if (!has)
                                d.p = P, d.s = S, d.q = Q, has = true;
                            else
                                d.p = d.p + P, d.s = d.s * S, d.q = qmul(d.q, Q);
                            chan = __ldg(st[k].next + chan);
                        }
                        touched |= has;
                        if (fabsf(wsum) >= 0.0001f && wn != 0.f)
                        {
                            if (!any)
                            {
                                qref  = d.q;
                                acc.p = wn * d.p, acc.s = wn * d.s, acc.q = qscale(d.q, wn);
                                any   = true;
                            }
                            else
                            {
                                float sg = (qdot(d.q, qref) < 0.0f) ? -wn : wn;
                                acc.p = acc.p + wn * d.p, acc.s = acc.s + wn * d.s;
                                acc.q = Quat {acc.q.w + d.q.w * sg, acc.q.x + d.q.x * sg, acc.q.y + d.q.y * sg, acc.q.z + d.q.z * sg};
                            }
                        }
                    }
                    if (any && touched)
                    {
                        loc.q = qmul(loc.q, qnormalise(acc.q));
                        loc.s = loc.s * acc.s;
                        loc.p = loc.p + acc.p;
                    }
                }

                // m_bones[i].update(), skeleton.cpp:110-113 -> Node::updateDerivedTransform node.cpp:32-65
                TRS der;
                if (hdr.y == -1)
                    der = loc;
                else
                {
                    TRS par_trs = prev;
                    if (hdr.y >= 0)
                    {
                        const float* sp = slots + (size_t)hdr.y * POSE_SLOT_FLOATS + lane;
                        par_trs.p = make_float3(sp[0], sp[32], sp[64]);
                        par_trs.q = Quat {sp[96], sp[128], sp[160], sp[192]};
                        par_trs.s = make_float3(sp[224], sp[256], sp[288]);
                    }
                    der = compose(par_trs, loc);
                }
                if (hdr.z >= 0)
                {
                    float* sp = slots + (size_t)hdr.z * POSE_SLOT_FLOATS + lane;
                    sp[0] = der.p.x, sp[32] = der.p.y, sp[64] = der.p.z;
                    sp[96] = der.q.w, sp[128] = der.q.x, sp[160] = der.q.y, sp[192] = der.q.z;
                    sp[224] = der.s.x, sp[256] = der.s.y, sp[288] = der.s.z;
                }
                prev = der;
                if (par.derived && active)
                {
                    float* o = par.derived + ((size_t)inst * par.J + bone) * 10;
                    o[0] = der.p.x, o[1] = der.p.y, o[2] = der.p.z, o[3] = der.q.w, o[4] = der.q.x;
                    o[5] = der.q.y, o[6] = der.q.z, o[7] = der.s.x, o[8] = der.s.y, o[9] = der.s.z;
                }

                // outputAnimationResult, skeleton.cpp:119-156
                float4 r0, r1, r2;
                palette_rows(der, __ldg(&rec->ib0), __ldg(&rec->ib1), __ldg(&rec->ib2), r0, r1, r2);
                float4* row = reinterpret_cast<float4*>(stage + (size_t)lane * SROW + g * 12);
                row[0] = r0, row[1] = r1, row[2] = r2;
            }
            __syncwarp();
            // coalesced flush: 32 instances x POSE_G bones x 3 float4
#pragma unroll
            for (int it = 0; it < POSE_G * 3; it++)
            {
                const int f  = it * 32 + lane;
                const int il = f / (POSE_G * 3);
                const int rm = f % (POSE_G * 3);
                const int g  = rm / 3;
                const int s  = s0 + g;
                if (s < par.J && inst0 + il < par.N)
                {
                    const int    bone = __ldg(&par.steps[s].joint);
                    const float4 v    = *reinterpret_cast<const float4*>(stage + (size_t)il * SROW + rm * 4);
                    par.palette[((size_t)(inst0 + il) * par.pitch + bone + 1) * 3 + (rm - g * 3)] = v;
                }
            }
            __syncwarp();
        }
    }

    // -------------------------------------------------------------------------------------------
    // pose kernel, step-ordered form: same mapping and arithmetic as pose_kernel, for the common case in
    // which every clip of the launch drives each bone with at most one channel (ClipRec::simple) and the
    // skeleton's step records fit the constant bank.
    //   - step records (bind TRS, inverse bind, parent/slot plan) come from __constant__ memory: uniform
    //     loads through the constant cache instead of 7 global loads per bone per lane
    //   - keys are read from the clip's step-ordered copy: the address depends on (step, frame) only, so the
    //     six 128-bit key loads of bone s+1 are issued before bone s is computed (L2 latency hidden);
    //     the clamp values of a whole group of POSE_G bones arrive as ONE 128-bit load
    //   - bones without a channel sample the identity track, which reproduces the bind pose exactly
    //     (q_bind * (1,0,0,0), s_bind * 1, p_bind + 0), so there is no per-bone branching
    //   - the palette flush uses the conflict-free row mapping of the LBS kernel with hoisted addressing
    // -------------------------------------------------------------------------------------------
    constexpr int POSE_CMAX = 320; // bones whose step records fit the constant bank (320 * 128 B = 40 KB)
    __constant__ StepRec c_steps[POSE_CMAX];

    // Step-ordered keys, one table per component (ClipRec::keys_step): {p.x p.y s.x s.y} [Jp][F] | {q.w q.x q.y q.z} [Jp][F] |
    // {p.z s.z} [Jp][F] - the pair layout of TRSP, so that the registers of the vector loads ARE the packed operands, and no
    // loaded register is dead (a dead lane of a vector load gets re-used by ptxas as a scratch register at once, and the
    // write-after-write hazard then stalls the warp for the full load latency right behind the prefetch - seen as
    // long_scoreboard on the first FMUL after the LDGs in profiles/r1_pose_v2_*).
    // Why per component: the 32 lanes of a warp sit at 32 different frames of the SAME step, and a load instruction costs the
    // L1 one pass per distinct 128-byte line.  A step's row of one component is F x 16 bytes - three or four lines for the
    // player's 23-frame clip - where the interleaved 48-byte keys spread the same 32 lanes over ~18 lines per instruction
    // (l1tex at 73 % of peak, warps waiting on it half of the time: profiles/r2_pose_v7_k2_summary.txt).
    __host__ __device__ inline int pose_jp(int J) { return (J + POSE_G - 1) / POSE_G * POSE_G + POSE_G; } // rows per table
    struct KeyPair
    {
        float4 a0, q0, a1, q1;
        float2 c0, c1;
    };
    struct SimpleSlot
    {
        const float4*   keys;
        const uint32_t* anim;
        uint32_t        F, lo, hi; // keys per track, index of the lower / upper key of this instance
        int32_t         mask_off;
        float           r, w;
    };
    // skeleton.cpp:52-55 and :75-87 for the step-ordered tracks (see ClipRec::keys_step)
    __device__ __forceinline__ void slot_frames(SimpleSlot& st, float ph, int total_frame_m1, int F)
    {
        const float exact = ph * (float)total_frame_m1;
        const int   fl = (int)floorf(exact), ce = (int)ceilf(exact);
        st.r           = exact - (float)fl;
        const int h    = max(min(ce, F - 1), 0);
        const int l    = max(min(fl, h), 0);
        st.F = (uint32_t)F, st.lo = (uint32_t)l, st.hi = (uint32_t)h;
    }
    // `after` is an opaque zero derived from a value the caller has just computed.  Adding it to the address
    // pins the loads BEHIND that computation, i.e. behind the consumption of the keys requested one bone ago.
    __device__ __forceinline__ int opaque_zero(float v)
    {
        int z;
        asm volatile("and.b32 %0, %1, 0;" : "=r"(z) : "r"(__float_as_int(v)));
        return z;
    }
    __device__ __forceinline__ KeyPair fetch_keys(const SimpleSlot& st, int s, int jp, int after = 0)
    {
        // a clip's step-ordered copy is far below 2^32 units
        const uint32_t ia = (uint32_t)s * st.F + (uint32_t)after, iq = (uint32_t)(s + jp) * st.F + (uint32_t)after;
        const uint32_t ic = (uint32_t)(s + 4 * jp) * st.F + (uint32_t)after; // (32-bit sums first: one IMAD.WIDE per address)
        const float2*  kc = reinterpret_cast<const float2*>(st.keys);
        KeyPair        k;
        k.a0 = __ldg(st.keys + (ia + st.lo)), k.q0 = __ldg(st.keys + (iq + st.lo)), k.c0 = __ldg(kc + (ic + st.lo));
        k.a1 = __ldg(st.keys + (ia + st.hi)), k.q1 = __ldg(st.keys + (iq + st.hi)), k.c1 = __ldg(kc + (ic + st.hi));
        return k;
    }
    // (Tried and rejected: staging the keys with cp.async - LDGSTS, one 80-byte shared-memory cell per lane and clip, four bones
    // in flight, completion per copy group instead of per scoreboard.  1 M instances 1.18 -> 2.04 ms: in the frame-coherent
    // order the 32 lanes of a plain load share one or two addresses and coalesce into one request, the per-lane copies do not.)
    // skeleton.cpp:88-95 on pairs: Vector3::lerp for position and scale, Quaternion::nLerp (shortest path) for the rotation
    struct SampleP
    {
        f2    pxy, sxy, pzsz;
        QuatP q;
    };
    __device__ __forceinline__ SampleP sample_keys(const KeyPair& k, float r)
    {
        SampleP s;
        s.pxy  = lerp2(pk(k.a0.x, k.a0.y), pk(k.a1.x, k.a1.y), r);
        s.sxy  = lerp2(pk(k.a0.z, k.a0.w), pk(k.a1.z, k.a1.w), r);
        s.pzsz = lerp2(pk(k.c0.x, k.c0.y), pk(k.c1.x, k.c1.y), r);
        s.q    = qnlerpp(r, QuatP {pk(k.q0.x, k.q0.y), pk(k.q0.z, k.q0.w)}, QuatP {pk(k.q1.x, k.q1.y), pk(k.q1.z, k.q1.w)});
        return s;
    }
    // Node::resetToInitialPose + rotate(LOCAL) / scale / translate(AREN) (node.cpp:197-206,135-157,182-186,107-132) of one
    // sampled (or blended) delta: q_l = q_bind * normalise(q), s_l = s_bind * s, p_l = p_bind + p
    __device__ __forceinline__ TRSP local_pose(const StepRec& rec, f2 pxy, f2 sxy, f2 pzsz, const QuatP& q)
    {
        TRSP loc;
        loc.q = qmul_constp(pk(rec.q01.x, rec.q01.y), pk(rec.q01.z, rec.q01.w), pk(rec.q23.x, rec.q23.y), pk(rec.q23.z, rec.q23.w), qnormalisep(q));
        loc.pxy  = add2p(pk(rec.bA.x, rec.bA.y), pxy);
        loc.sxy  = mul2p(pk(rec.bA.z, rec.bA.w), sxy);
        loc.pzsz = fma2p(pzsz, pk(rec.bC.x, rec.bC.y), pk(rec.bC.z, rec.bC.w));
        return loc;
    }
    __device__ __forceinline__ TRSP bind_pose(const StepRec& rec)
    {
        TRSP loc;
        loc.q    = QuatP {pk(rec.q01.x, rec.q01.y), pk(rec.q01.z, rec.q01.w)};
        loc.pxy  = pk(rec.bA.x, rec.bA.y);
        loc.sxy  = pk(rec.bA.z, rec.bA.w);
        loc.pzsz = pk(rec.bC.z, rec.bC.y);
        return loc;
    }
    // parked derived transforms: five 8-byte pairs, [pair][lane] (consecutive lanes on consecutive 8-byte units: conflict-free
    // 64-bit accesses that read and write the packed registers as they are)
    __device__ __forceinline__ TRSP load_trs(const float* slot, int lane)
    {
        const f2* s2 = reinterpret_cast<const f2*>(slot) + lane;
        TRSP      t;
        t.pxy = s2[0], t.sxy = s2[32], t.pzsz = s2[64], t.q.wx = s2[96], t.q.yz = s2[128];
        return t;
    }
    __device__ __forceinline__ void store_trs(float* slot, int lane, const TRSP& t)
    {
        f2* s2 = reinterpret_cast<f2*>(slot) + lane;
        s2[0] = t.pxy, s2[32] = t.sxy, s2[64] = t.pzsz, s2[96] = t.q.wx, s2[128] = t.q.yz;
    }
    // the same, predicated instead of branched (the transform keeps its value when `p` is false): a uniform branch around five
    // loads ends a basic block, and with it what ptxas can overlap
    __device__ __forceinline__ void load_trs_if(bool p, const float* slot, int lane, TRSP& t)
    {
        const uint32_t a = (uint32_t)__cvta_generic_to_shared(slot) + (uint32_t)lane * 8u;
        asm volatile("{\n.reg .pred q;\nsetp.ne.b32 q, %6, 0;\n"
                     "@q ld.shared.b64 %0, [%5];\n@q ld.shared.b64 %1, [%5+256];\n@q ld.shared.b64 %2, [%5+512];\n"
                     "@q ld.shared.b64 %3, [%5+768];\n@q ld.shared.b64 %4, [%5+1024];\n}"
                     : "+l"(t.pxy), "+l"(t.sxy), "+l"(t.pzsz), "+l"(t.q.wx), "+l"(t.q.yz)
                     : "r"(a), "r"((int)p)
                     : "memory");
    }
    __device__ __forceinline__ void store_trs_if(bool p, float* slot, int lane, const TRSP& t)
    {
        const uint32_t a = (uint32_t)__cvta_generic_to_shared(slot) + (uint32_t)lane * 8u;
        asm volatile("{\n.reg .pred q;\nsetp.ne.b32 q, %6, 0;\n"
                     "@q st.shared.b64 [%5], %0;\n@q st.shared.b64 [%5+256], %1;\n@q st.shared.b64 [%5+512], %2;\n"
                     "@q st.shared.b64 [%5+768], %3;\n@q st.shared.b64 [%5+1024], %4;\n}" ::"l"(t.pxy),
                     "l"(t.sxy), "l"(t.pzsz), "l"(t.q.wx), "l"(t.q.yz), "r"(a), "r"((int)p)
                     : "memory");
    }
    __device__ __forceinline__ TRSP identity_trs()
    {
        TRSP t;
        t.pxy = pk(0.f, 0.f), t.sxy = pk(1.f, 1.f), t.pzsz = pk(0.f, 1.f), t.q = QuatP {pk(1.f, 0.f), pk(0.f, 0.f)};
        return t;
    }
    __device__ __forceinline__ void store_derived(float* o, const TRSP& t)
    {
        o[0] = lo(t.pxy), o[1] = hi(t.pxy), o[2] = lo(t.pzsz), o[3] = lo(t.q.wx), o[4] = hi(t.q.wx);
        o[5] = lo(t.q.yz), o[6] = hi(t.q.yz), o[7] = lo(t.sxy), o[8] = hi(t.sxy), o[9] = hi(t.pzsz);
    }

    // MB = resident blocks per SM the register allocation is made for (__launch_bounds__): 2 -> 255 registers, 3 -> 168, 4 -> 128.
    // With IEEE sqrt / division in the normalisations (slow-path branches after every one of them) the kernel hid its latencies
    // only through resident warps: at N = 1 M, K = 1, 8 / 12 / 16 warps per SM ran 2.38 / 2.10 / 1.78 ms.  With the branch-free
    // normalise (pcl_math.cuh) ptxas schedules across whole bones and the registers are worth more than the warps:
    // 1.36 / 1.40 / 1.44 ms, and 33 us instead of 50 us for 10 k instances.  Two clips per instance still prefer 12 warps for the
    // largest crowds (2.04 vs 2.23 ms at 1 M) and 8 below (0.236 vs 0.254 ms at 100 k); the launcher picks.
    // palette flush stores: 0 plain, 1 streaming (st.global.cs), 2 cache-global (st.global.cg).  Measured (tools/pose_bench.py,
    // profiles/README.md): streaming stores -13 % at 100 k instances (K = 1 and K = 2), -4 % at 1 M; .cg within noise.
#ifndef PCL_POSE_ST_MODE
#define PCL_POSE_ST_MODE 1
#endif
#if PCL_POSE_ST_MODE == 1
#define PCL_PST(p, v) __stcs((p), (v))
#elif PCL_POSE_ST_MODE == 2
#define PCL_PST(p, v) __stcg((p), (v))
#else
#define PCL_PST(p, v) (*(p) = (v))
#endif
    constexpr int POSE_S_THREADS = 128;
    // SORTED: lane l of warp w works on instance order[32 w + l] instead of 32 w + l.  The order groups instances by the
    // (clip, frame) their phase falls in, so the 32 key gathers of a warp collapse onto one or two addresses (broadcast)
    // instead of up to 32 cache lines; every instance is still computed by exactly the same instruction sequence and
    // written to its own palette row, so the result does not depend on the order.
    // where the step records are read from: 0 = the constant bank, 1 = a copy in shared memory (made by the block, coalesced,
    // at its start - what pose_level_kernel does).  Measured (tools/pose_bench.py, K = 1 / K = 2): shared memory 29.0 / 41.2 us
    // against 30.0 / 41.3 at 10 k instances, but 147 / 207 us against 134 / 191 at 100 k and level at 1 M: the constant bank stays.
#ifndef PCL_POSE_STEPS_SMEM
#define PCL_POSE_STEPS_SMEM 0
#endif
    // BULK (palettes that live in ANOTHER GPU's memory, pcl_batch_attach_pool): the staged rows of a group leave as one bulk copy
    // per instance (cp.async.bulk shared -> global, 192 contiguous bytes) instead of 16-byte stores from the lanes.  Stores over
    // NVLink keep the issuing warp's store queue busy far longer than local ones, and the pose kernel - latency-bound as it is -
    // stalled behind them (376 GB/s per producer); the bulk copies are asynchronous, the warp goes on with the next group and only
    // has to find its staging tile read before it overwrites it.  (A plain store-side bulk copy: nothing is multicast, nothing is
    // loaded from the peer.)
    __device__ __forceinline__ void pose_bulk_store(void* dst_global, const void* src_smem, uint32_t bytes)
    {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // this lane's staging stores, for the async proxy
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_global), "r"((uint32_t)__cvta_generic_to_shared(src_smem)),
                     "r"(bytes)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    __device__ __forceinline__ void pose_bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
    __device__ __forceinline__ void pose_bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

    // DER: the derived transforms are written out too (LevelDebugger-style read-back); a template parameter so that the common
    // form carries no branch for it.  A group of POSE_G bones is ONE straight-line block there (no bounds checks in whole groups,
    // parent fetch and parking predicated instead of branched): ptxas may then run bone s + 1's sampling and local pose - which do
    // not depend on the hierarchy - inside bone s's composition chain.
    // SPLIT (small one-clip crowds, where a warp is alone on its scheduler and the 58-step walk IS the run time): the skeleton
    // is cut into two halves - a set of whole subtrees plus their ancestors, and everything outside those subtrees - and
    // two warps (blockIdx.y) take one half each of the same 32 instances.  The shared ancestors are computed by both, from the
    // same inputs with the same instructions, so the rows both write are bit-identical.  Each half has its own step records
    // (parking plan of its own, upload time); the key tables stay those of the whole skeleton (StepRec::pstep = the row).
    // MASKS = false (no bone mask uploaded for the skeleton: par.mask_ids == nullptr): every clip contributes to every bone, so
    // the normalised blend weights (animation_system.cpp:117-149) are the same for all bones of an instance - computed once,
    // before the walk, by the same expression; with masks they depend on the bone and are computed per bone.
    template<int K, bool SORTED, int MB, bool DER, bool BULK = false, bool SPLIT = false, bool MASKS = true>
    __global__ void __launch_bounds__(POSE_S_THREADS, MB) pose_simple_kernel(const PoseParams par)
    {
        static_assert(!SPLIT || (K == 1 && !SORTED && !DER && !BULK), "the two-half form exists for plain one-clip crowds");
        const int sbase = SPLIT ? par.split_base[blockIdx.y] : 0;  // first step record of this warp's list
        const int nst   = SPLIT ? par.split_n[blockIdx.y] : par.J; // steps this warp walks
        extern __shared__ __align__(16) float smem[];
        const int       lane  = threadIdx.x & 31;
        const int       wib   = threadIdx.x >> 5;
        const long long wglob = (long long)blockIdx.x * (blockDim.x >> 5) + wib;
        const long long inst0 = wglob * 32;
#if PCL_POSE_STEPS_SMEM
        const StepRec* s_steps = reinterpret_cast<const StepRec*>(smem + (size_t)(blockDim.x >> 5) * pose_smem_floats_per_warp(par.nslot));
        {
            const int4* src = reinterpret_cast<const int4*>(par.steps);
            int4*       dst = reinterpret_cast<int4*>(const_cast<StepRec*>(s_steps));
            for (int i = threadIdx.x; i < par.J * (int)(sizeof(StepRec) / 16); i += blockDim.x)
                dst[i] = __ldg(src + i);
            __syncthreads();
        }
#define PCL_STEP(s_) s_steps[sbase + (s_)]
#else
#define PCL_STEP(s_) c_steps[sbase + (s_)]
#endif
#define PCL_KROW(s_) (SPLIT ? PCL_STEP(s_).pstep : (s_)) /* the step's row in the key tables */
        if (inst0 >= par.N)
            return;
        const bool      active = (inst0 + lane) < par.N;
        const long long visit  = min(inst0 + lane, par.N - 1);
        const long long inst   = SORTED ? (long long)__ldg(par.order + visit) : visit;
        const int       inst32 = (int)inst;
        const int       ni     = (int)min(32LL, par.N - inst0);
        constexpr bool  WKEYS  = K == 1 && SORTED; // see pose_wkeys_floats
        float*          slots  = smem + (size_t)wib * (pose_smem_floats_per_warp(par.nslot) + (WKEYS ? pose_wkeys_floats(par.J) : 0));
        float*          stage  = slots + (size_t)(par.nslot + 1) * POSE_SLOT_FLOATS;
        constexpr int   SROW   = POSE_G * 12 + 4;
        float4*         wkeys  = reinterpret_cast<float4*>(stage + 32 * SROW); // [J][6] float4: lower key, upper key of every step
        store_trs(slots + (size_t)par.nslot * POSE_SLOT_FLOATS, lane, identity_trs()); // slot nslot: what a root composes with

        SimpleSlot st[K];
#pragma unroll
        for (int k = 0; k < K; k++)
        {
            const size_t  o  = (size_t)k * par.N + inst;
            const ClipRec cr = par.clips[__ldg(par.clip_ids + o)];
            float         ph = par.phase[o];
            if (par.advance == 2 || (par.advance == 1 && k == 0))
            {
                ph += par.dt / __ldg(par.clip_len + o); // animation_component.cpp:19-21
                ph -= floorf(ph);
                // (SPLIT: both halves read the old phase, so the new one goes to a second array; the host swaps the two)
                if (active && (!SPLIT || blockIdx.y == 0))
                    (SPLIT ? par.phase_out : par.phase)[o] = ph;
            }
            slot_frames(st[k], ph, cr.total_frame_m1, cr.F);
            st[k].keys = cr.keys_step, st[k].anim = cr.anim_step;
            st[k].w = (K > 1) ? __ldg(par.weight + o) : 1.0f;
            const int mid = (K > 1 && par.mask_ids) ? __ldg(par.mask_ids + o) : -1;
            st[k].mask_off = mid < 0 ? -1 : mid * par.J;
        }
        float we0[K]; // MASKS = false: each clip's effective weight, for every bone
        if (K > 1 && !MASKS)
        {
            float wsum = 0.f;
#pragma unroll
            for (int k = 0; k < K; k++)
                wsum += st[k].w;
            const bool live = fabsf(wsum) >= 0.0001f;
#pragma unroll
            for (int k = 0; k < K; k++)
            {
                const float wn = PCL_EXACT_NORMALISE ? st[k].w / wsum : fdiv_nobranch(st[k].w, wsum);
                we0[k]         = (live && wn != 0.f) ? wn : 0.f;
            }
        }
        // (slot 0 = identity, mesh_component.cpp:72-79: it leaves with the first group's flush, see `group` below)
        const int jp = pose_jp(par.J); // rows per key table (fetch_keys)
        // lane-dependent parts of the flush addresses (rows = instances, 12 units = POSE_G bones x 3 rows)
        const int fa = lane >> 3, fb = lane & 7, fc4 = ((lane >> 2) & 1) * 4, fd = lane & 3;
        const int s_f0 = fa * SROW + fb * 4, s_f2 = (fa + fc4) * SROW + (8 + fd) * 4;
        float4*   g_f0 = par.palette + (size_t)(inst0 + fa) * par.pitch * 3 + fb;
        float4*   g_f2 = par.palette + (size_t)(inst0 + fa + fc4) * par.pitch * 3 + 8 + fd;
        const size_t rowstep = (size_t)par.pitch * 3 * 4; // float4 per 4 instances
        const size_t rowsz   = (size_t)par.pitch * 3;     // float4 per instance
        // SORTED: staging row r belongs to the instance lane r works on.  The twelve rows this lane flushes are fixed for the
        // whole kernel, so their instance numbers are fetched once, here, instead of with twelve shuffles per group
        constexpr bool HOIST = SORTED && MB == 2; // (the 168- and 128-register forms have no room for twelve more registers)
        uint32_t       frow[HOIST ? 12 : 1];
        if (HOIST)
        {
#pragma unroll
            for (int gq = 0; gq < 4; gq++)
            {
#ifdef PCL_POSE_WRAP_ROWS /* experiment: all palette rows land in a small, L2-resident window - is the kernel bound by where it writes? */
#define PCL_WRAP(x) ((x) & (PCL_POSE_WRAP_ROWS - 1))
#else
#define PCL_WRAP(x) (x)
#endif
                frow[gq * 3 + 0] = (uint32_t)PCL_WRAP(__shfl_sync(0xffffffffu, inst32, fa + gq * 8));
                frow[gq * 3 + 1] = (uint32_t)PCL_WRAP(__shfl_sync(0xffffffffu, inst32, fa + gq * 8 + 4));
                frow[gq * 3 + 2] = (uint32_t)PCL_WRAP(__shfl_sync(0xffffffffu, inst32, fa + fc4 + gq * 8));
            }
        }

        TRSP prev = identity_trs(); // derived transform of the previous step = the parent of a chain's next bone

        // one bucket for the whole warp?  Then its keys come from shared memory (copied here, coalesced, in one burst)
        bool uni = false;
        if (WKEYS)
        {
            // (all three exchanges before any comparison: a short-circuited && would keep lanes out of a full-mask shuffle)
            const unsigned long long kp  = reinterpret_cast<unsigned long long>(st[0].keys);
            const unsigned long long kp0 = __shfl_sync(0xffffffffu, kp, 0);
            const uint32_t           lo0 = __shfl_sync(0xffffffffu, st[0].lo, 0), hi0 = __shfl_sync(0xffffffffu, st[0].hi, 0);
            const bool               same = (kp == kp0) & (st[0].lo == lo0) & (st[0].hi == hi0);
            uni = __all_sync(0xffffffffu, same) != 0;
            if (uni)
            {
                for (int i = lane; i < par.J * 6; i += 32)
                {
                    const int sI = i / 6, u = i - sI * 6;
                    const uint32_t f = u < 3 ? st[0].lo : st[0].hi, c = u < 3 ? u : u - 3; // component: 0 a, 1 q, 2 c
                    if (c < 2)
                        wkeys[i] = __ldg(st[0].keys + ((uint32_t)(sI + c * jp) * st[0].F + f));
                    else
                    {
                        const float2 t = __ldg(reinterpret_cast<const float2*>(st[0].keys) + ((uint32_t)(sI + 4 * jp) * st[0].F + f));
                        wkeys[i]       = make_float4(t.x, t.y, 0.f, 0.f);
                    }
                }
                __syncwarp();
            }
        }
        // software pipeline: keys one bone ahead (register budget allows it for K <= 2)
        constexpr bool PREFETCH = K <= 2;
        // One clip per instance and the 255-register form: keys TWO bones ahead (two rotating buffers; POSE_G is even, so a
        // bone's buffer is its position in the group modulo 2) - one bone's worth of work did not cover a miss to L2
        constexpr int  PD = (K == 1 && MB == 2) ? 2 : 1;
        KeyPair  nk[PREFETCH ? K * PD : 1];
        uint32_t amask[K]; // K > 1: which of the current 32 steps each clip animates
#pragma unroll
        for (int k = 0; k < K; k++)
        {
            amask[k] = 0xffffffffu;
            if (PREFETCH && !uni)
            {
                // (step s sits at position (s + 1) & 3 of its group: step 0 uses buffer 1, step 1 buffer 0)
                nk[PD == 2 ? K + k : k] = fetch_keys(st[k], PCL_KROW(0), jp);
                if (PD == 2)
                    nk[k] = fetch_keys(st[k], PCL_KROW(1), jp);
            }
        }

        auto group = [&](const int s0, auto full_tag, auto uni_tag) {
            constexpr bool FULL = decltype(full_tag)::value; // all POSE_G bones of the group exist
            constexpr bool UNI  = decltype(uni_tag)::value;  // the warp's keys come from its shared-memory copy
            // A group is four consecutive palette SLOTS starting at a multiple of four - 192 bytes on a 64-byte boundary of the
            // instance's row (the row pitch is a multiple of 4 slots) - so its flush writes whole 32-byte sectors only.  Slot 0 is
            // the identity, slot j + 1 bone j: the first group is the identity + steps 0..2 (s0 = -1), the next one steps 3..6, ...
            // (Groups of four STEPS put every flush 48 bytes off the boundary: both end sectors of each instance's 192 bytes
            // were written in halves by two different flushes.  1 M instances, one clip: 1.13 -> 0.95 ms, tools/pose_bench.py.)
            const int slot0 = s0 < 0 ? 0 : PCL_STEP(s0).joint + 1;
            bool contiguous = true; // the group's palette slots are consecutive (always, unless the upload re-ordered the bones)
#pragma unroll
            for (int g = 0; g < POSE_G; g++)
            {
                const int s = s0 + g;
                if (g == 0 && s0 < 0)
                {
                    float4* row = reinterpret_cast<float4*>(stage + (size_t)lane * SROW);
                    row[0] = make_float4(1, 0, 0, 0), row[1] = make_float4(0, 1, 0, 0), row[2] = make_float4(0, 0, 1, 0);
                    continue;
                }
                if (!FULL && s >= nst)
                    break;
                if (K > 1 && g == 1 && ((s0 + 1) & 31) == 0) // (step s0 itself is the last bit of the previous word)
                {
#pragma unroll
                    for (int k = 0; k < K; k++)
                        amask[k] = __ldg(st[k].anim + ((s0 + 1) >> 5));
                }
                const StepRec& rec  = PCL_STEP(s);
                const int      bone = rec.joint;
                contiguous          = contiguous && (bone + 1 == slot0 + g);
                KeyPair ck[K];
#pragma unroll
                for (int k = 0; k < K; k++)
                    ck[k] = PREFETCH ? nk[(PD == 2 ? (g & 1) * K : 0) + k] : fetch_keys(st[k], PCL_KROW(s), jp);
                if (WKEYS && UNI)
                {
                    const float4* wk = wkeys + s * 6; // the same address for all lanes: broadcast reads
                    const float4  t0 = wk[2], t1 = wk[5];
                    ck[0].a0 = wk[0], ck[0].q0 = wk[1], ck[0].c0 = make_float2(t0.x, t0.y);
                    ck[0].a1 = wk[3], ck[0].q1 = wk[4], ck[0].c1 = make_float2(t1.x, t1.y);
                }
                // this bone's samples first (they consume the keys requested one bone ago) ...
                SampleP sm[K];
#pragma unroll
                for (int k = 0; k < K; k++)
                    sm[k] = sample_keys(ck[k], st[k].r);
                // ... then the next bone's keys are requested; they land while the rest of this bone is computed
                if (PREFETCH && !UNI && (FULL || s + PD < nst)) // (whole groups: the key tracks are padded by a group, §2)
                {
#pragma unroll
                    for (int k = 0; k < K; k++)
                        nk[(PD == 2 ? (g & 1) * K : 0) + k] = fetch_keys(st[k], PCL_KROW(s + PD), jp, opaque_zero(lo(sm[k].q.wx)));
                }
                TRSP loc;
                if (K == 1)
                {
                    // skeleton.cpp:88-100 on top of Node::resetToInitialPose (skeleton.cpp:45)
                    loc = local_pose(rec, sm[0].pxy, sm[0].sxy, sm[0].pzsz, sm[0].q);
                }
                else
                {
                    // documented blend extension; weights as animation_system.cpp:117-149
                    float wsum = 0.f;
                    bool  en[K], touched = false;
#pragma unroll
                    for (int k = 0; k < K; k++)
                    {
                        en[k] = (!MASKS || st[k].mask_off < 0) ? true : (__ldg(par.masks + st[k].mask_off + bone) != 0);
                        if (MASKS && en[k])
                            wsum += st[k].w;
                        touched |= ((amask[k] >> (s & 31)) & 1u) != 0;
                    }
                    // One instruction stream for every mix of contributing and gated clips: a clip that does not contribute enters
                    // with weight 0 (x + 0 * d = x exactly), the first contributing one becomes the sign reference (selects)
                    f2    accA = pk(0.f, 0.f), accB = accA, accC = accA;
                    QuatP accq {accA, accA}, qref {pk(1.f, 0.f), accA};
                    bool  any = false;
                    const bool live = fabsf(wsum) >= 0.0001f;
#pragma unroll
                    for (int k = 0; k < K; k++)
                    {
                        // (an identity track samples to the identity delta, exactly)
                        bool  ok;
                        float we;
                        if (MASKS)
                        {
                            const float wn = en[k] ? (PCL_EXACT_NORMALISE ? st[k].w / wsum : fdiv_nobranch(st[k].w, wsum)) : 0.f; // animation_system.cpp:117-149
                            ok = live && wn != 0.f, we = ok ? wn : 0.f;
                        }
                        else
                            we = we0[k], ok = we != 0.f;
                        const float sg = (any && qdotp(sm[k].q, qref) < 0.0f) ? -we : we;
                        accA = fma2p(sm[k].pxy, bc(we), accA), accB = fma2p(sm[k].sxy, bc(we), accB), accC = fma2p(sm[k].pzsz, bc(we), accC);
                        accq = QuatP {fma2p(sm[k].q.wx, bc(sg), accq.wx), fma2p(sm[k].q.yz, bc(sg), accq.yz)};
                        const bool first = ok && !any;
                        qref.wx = first ? sm[k].q.wx : qref.wx, qref.yz = first ? sm[k].q.yz : qref.yz;
                        any |= ok;
                    }
                    // bones no clip animates or none contributes to keep the bind pose: the identity delta reproduces it exactly
                    // (q_bind * (1,0,0,0), s_bind * 1, p_bind + 0), selected here instead of branched to
                    const bool keep = !(any && touched);
                    accA = keep ? pk(0.f, 0.f) : accA, accB = keep ? pk(1.f, 1.f) : accB, accC = keep ? pk(0.f, 1.f) : accC;
                    accq.wx = keep ? pk(1.f, 0.f) : accq.wx, accq.yz = keep ? pk(0.f, 0.f) : accq.yz;
                    loc = local_pose(rec, accA, accB, accC, accq);
                }
                // m_bones[i].update(), skeleton.cpp:110-113 -> Node::updateDerivedTransform node.cpp:32-65
                // (one code path on purpose: a parent fetched from its slot - for a root, the identity in slot nslot - REPLACES
                // `prev` in place, and a root goes through the same composition with the normalisation switched off; selecting
                // between register sets cost ~30 copies per bone)
                load_trs_if(rec.psrc != -2, slots + (size_t)(rec.psrc == -1 ? par.nslot : max(rec.psrc, 0)) * POSE_SLOT_FLOATS, lane, prev);
                prev = composep(prev, loc, rec.psrc == -1);
                const TRSP& der = prev;
                store_trs_if(rec.store >= 0, slots + (size_t)max(rec.store, 0) * POSE_SLOT_FLOATS, lane, der);
                if (DER && active)
                    store_derived(par.derived + ((size_t)inst * par.J + bone) * 10, der);
                float4 r0, r1, r2; // outputAnimationResult, skeleton.cpp:119-156
                palette_rowsp(der, rec.ib0, rec.ib1, rec.ib2, r0, r1, r2);
                if (BULK && g == 0)
                    pose_bulk_wait_read(); // the previous group's bulk copy has read this lane's staging row
                float4* row = reinterpret_cast<float4*>(stage + (size_t)lane * SROW + g * 12);
                row[0] = r0, row[1] = r1, row[2] = r2;
            }
            __syncwarp();
            if (BULK && contiguous && FULL && active)
            {
                // this lane's own row: 4 slots x 48 bytes, contiguous in the (remote) palette
                pose_bulk_store(par.palette + (size_t)inst * rowsz + (size_t)slot0 * 3, stage + (size_t)lane * SROW, POSE_G * 48);
            }
            else if (BULK && contiguous && FULL)
            {
            }
            else if (contiguous && ni == 32 && FULL)
            {
                // 32 rows x 12 units, conflict-free (see lbs_i_flush12); the group's 4 slots are 12 consecutive units
                const size_t off = (size_t)slot0 * 3;
#pragma unroll
                for (int gq = 0; gq < 4; gq++)
                {
                    if (SORTED && !HOIST)
                    {
                        const size_t i0 = (size_t)__shfl_sync(0xffffffffu, inst32, fa + gq * 8);
                        const size_t i1 = (size_t)__shfl_sync(0xffffffffu, inst32, fa + gq * 8 + 4);
                        const size_t i2 = (size_t)__shfl_sync(0xffffffffu, inst32, fa + fc4 + gq * 8);
                        PCL_PST(par.palette + i0 * rowsz + off + fb, *reinterpret_cast<const float4*>(stage + s_f0 + gq * 8 * SROW));
                        PCL_PST(par.palette + i1 * rowsz + off + fb, *reinterpret_cast<const float4*>(stage + s_f0 + (gq * 8 + 4) * SROW));
                        PCL_PST(par.palette + i2 * rowsz + off + 8 + fd, *reinterpret_cast<const float4*>(stage + s_f2 + gq * 8 * SROW));
                    }
                    else if (SORTED)
                    {
                        float4* gb0 = par.palette + off + fb;
                        float4* gb2 = par.palette + off + 8 + fd;
                        PCL_PST(gb0 + (size_t)frow[gq * 3 + 0] * rowsz, *reinterpret_cast<const float4*>(stage + s_f0 + gq * 8 * SROW));
                        PCL_PST(gb0 + (size_t)frow[gq * 3 + 1] * rowsz, *reinterpret_cast<const float4*>(stage + s_f0 + (gq * 8 + 4) * SROW));
                        PCL_PST(gb2 + (size_t)frow[gq * 3 + 2] * rowsz, *reinterpret_cast<const float4*>(stage + s_f2 + gq * 8 * SROW));
                    }
                    else
                    {
                        PCL_PST(g_f0 + off + (2 * gq) * rowstep, *reinterpret_cast<const float4*>(stage + s_f0 + gq * 8 * SROW));
                        PCL_PST(g_f0 + off + (2 * gq + 1) * rowstep, *reinterpret_cast<const float4*>(stage + s_f0 + (gq * 8 + 4) * SROW));
                        PCL_PST(g_f2 + off + (2 * gq) * rowstep, *reinterpret_cast<const float4*>(stage + s_f2 + gq * 8 * SROW));
                    }
                }
            }
            else
            {
#pragma unroll
                for (int it = 0; it < POSE_G * 3; it++)
                {
                    const int f  = it * 32 + lane;
                    const int il = f / (POSE_G * 3);
                    const int rm = f % (POSE_G * 3);
                    const int g  = rm / 3;
                    const int s  = s0 + g;
                    const size_t irow = SORTED ? (size_t)__shfl_sync(0xffffffffu, inst32, il & 31) : (size_t)(inst0 + il);
                    if (s < nst && il < ni)
                        par.palette[(irow * par.pitch + (s < 0 ? 0 : PCL_STEP(s).joint + 1)) * 3 + (rm - g * 3)] =
                            *reinterpret_cast<const float4*>(stage + (size_t)il * SROW + rm * 4);
                }
            }
            __syncwarp();
        };
        int s0 = -1; // (the identity slot takes the first place of the first group)
        if (WKEYS && uni)
        {
            for (; s0 + POSE_G <= nst; s0 += POSE_G)
                group(s0, std::true_type {}, std::true_type {});
            if (s0 < nst)
                group(s0, std::false_type {}, std::true_type {});
        }
        else
        {
            for (; s0 + POSE_G <= nst; s0 += POSE_G)
                group(s0, std::true_type {}, std::false_type {});
            if (s0 < nst)
                group(s0, std::false_type {}, std::false_type {});
        }
        if (BULK)
            pose_bulk_wait_all();
    }

    // -------------------------------------------------------------------------------------------
    // pose kernel for SMALL crowds (one clip per instance): bones of one hierarchy level in parallel.
    //
    // pose_simple_kernel gives every instance one lane, which walks all J bones: 58 x ~1000 cycles = 30 us however few
    // instances there are - the time of ONE AnimationComponent::tick on the host.  Bones of the same depth do not depend on each
    // other, so here a CTA of POSE_L_WARPS warps shares 32 instances: in every round each warp takes one bone of the current
    // level (the schedule is built at upload: level by level, ceil(width / warps) rounds per level), derived transforms are
    // exchanged through shared memory ([step][component][lane]) and a __syncthreads() closes the round.  The player's 58 bones
    // in 13 levels (widths 1,1,1,3,3,3,5,5,4,2,10,10,10) take 16 rounds instead of 58 steps.  Same device functions, same
    // operation order per bone as the other pose kernels - the results are bit-identical; only worth it while the crowd leaves
    // SMs idle (the launcher uses it up to ~3 CTAs per SM), because at narrow levels most warps wait.
    // -------------------------------------------------------------------------------------------
    // POSE_L_WARPS "derive" warps take one bone of the current level each per round (sample, local pose, composition with
    // the parent, result parked in shared memory for the children); as many "palette" warps trail ONE ROUND behind and turn
    // the parked transforms into palette rows (Transform::getMatrix x inverse bind, the stores to global memory, the derived
    // read-back) - a third of a bone's instructions, none of which a child waits for, off the chain of rounds.
    // 12 + 12 warps: the player's widest levels (10 bones) take one round each, 13 rounds in all instead of 16.
    constexpr int POSE_L_WARPS = 12;
    // shared memory: parked transforms [J] | the skeleton's step records [J] | the schedule [n_rounds][POSE_L_WARPS]
    __host__ __device__ inline size_t pose_l_smem_bytes(int J, int n_rounds)
    {
        return (size_t)J * POSE_SLOT_FLOATS * sizeof(float) + (size_t)J * sizeof(StepRec) + (((size_t)n_rounds * POSE_L_WARPS * 4 + 15) & ~(size_t)15);
    }

    __global__ void __launch_bounds__(POSE_L_WARPS * 64) pose_level_kernel(const PoseParams par)
    {
        extern __shared__ __align__(16) float smem[];
        const int       lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        const long long inst0 = (long long)blockIdx.x * 32;
        const bool      active = (inst0 + lane) < par.N;
        const long long inst   = min(inst0 + lane, par.N - 1); // idle lanes shadow the last instance, never store
        float4*         pal    = par.palette + (size_t)inst * par.pitch * 3;
        // The step records and the schedule are copied into shared memory first (coalesced, once): a round's look-ups are then
        // shared-memory broadcasts.  Read through the constant bank, a CTA that is alone on its SM - one character, the
        // reference's everyday case - paid a constant-cache miss per round, most of a round's time.
        StepRec* s_steps = reinterpret_cast<StepRec*>(smem + (size_t)par.J * POSE_SLOT_FLOATS);
        int32_t* s_sched = reinterpret_cast<int32_t*>(s_steps + par.J);
        {
            const int4* src = reinterpret_cast<const int4*>(par.steps);
            int4*       dst = reinterpret_cast<int4*>(s_steps);
            for (int i = threadIdx.x; i < par.J * (int)(sizeof(StepRec) / 16); i += blockDim.x)
                dst[i] = __ldg(src + i);
            for (int i = threadIdx.x; i < par.n_rounds * POSE_L_WARPS; i += blockDim.x)
                s_sched[i] = __ldg(par.sched + i);
        }

        if (warp >= POSE_L_WARPS)
        {
            // ---------------- palette warps ----------------
            const int32_t* my = s_sched + (warp - POSE_L_WARPS);
            if (warp == POSE_L_WARPS && active)
            {
                pal[0] = make_float4(1, 0, 0, 0), pal[1] = make_float4(0, 1, 0, 0), pal[2] = make_float4(0, 0, 1, 0); // slot 0 = identity
            }
            __syncthreads(); // (the phase hand-over of the derive warps)
            int s_prev = -1;
            for (int r = 0; r <= par.n_rounds; r++)
            {
                const int s = s_prev;
                s_prev      = r < par.n_rounds ? my[r * POSE_L_WARPS] : -1;
                if (s >= 0 && active)
                {
                    const StepRec& rec  = s_steps[s];
                    const int      bone = rec.joint;
                    const TRSP     der  = load_trs(smem + (size_t)s * POSE_SLOT_FLOATS, lane);
                    if (par.derived)
                        store_derived(par.derived + ((size_t)inst * par.J + bone) * 10, der);
                    float4 r0, r1, r2; // outputAnimationResult, skeleton.cpp:119-156
                    palette_rowsp(der, rec.ib0, rec.ib1, rec.ib2, r0, r1, r2);
                    if (par.out12)
                    {
                        float4* o = par.out12 + ((size_t)inst * par.J + bone) * 3;
                        o[0] = r0, o[1] = r1, o[2] = r2;
                    }
                    float4* row = pal + (size_t)(bone + 1) * 3;
                    row[0] = r0, row[1] = r1, row[2] = r2;
                }
                __syncthreads();
            }
            return;
        }

        // ---------------- derive warps ----------------
        const ClipRec cr = par.clips[__ldg(par.clip_ids + inst)];
        float         ph = par.phase[inst];
        if (par.advance != 0)
        {
            ph += par.dt / __ldg(par.clip_len + inst); // animation_component.cpp:19-21
            ph -= floorf(ph);
        }
        SimpleSlot st;
        slot_frames(st, ph, cr.total_frame_m1, cr.F);
        st.keys = cr.keys_step, st.anim = cr.anim_step, st.w = 1.0f, st.mask_off = -1;
        __syncthreads(); // every warp has read the old phase before warp 0 replaces it
        if (warp == 0 && active && par.advance != 0)
            par.phase[inst] = ph;

        // software pipeline over the warp's own rounds: keys one round ahead
        const int32_t* my = s_sched + warp;
        int            s_cur = my[0], s_nxt = par.n_rounds > 1 ? my[POSE_L_WARPS] : -1;
        const int      jp    = pose_jp(par.J);
        KeyPair        k_cur = fetch_keys(st, max(s_cur, 0), jp);
        for (int r = 0; r <= par.n_rounds; r++)
        {
            const int     s = s_cur;
            const KeyPair k = k_cur;
            const int s_nn = r + 2 < par.n_rounds ? my[(r + 2) * POSE_L_WARPS] : -1;
            if (s_nxt >= 0)
                k_cur = fetch_keys(st, s_nxt, jp);
            s_cur = s_nxt, s_nxt = s_nn;
            if (s >= 0)
            {
                const StepRec& rec = s_steps[s];
                const SampleP  sm  = sample_keys(k, st.r);
                // Node::resetToInitialPose + rotate(LOCAL) / scale / translate(AREN), as in pose_simple_kernel
                const TRSP loc = local_pose(rec, sm.pxy, sm.sxy, sm.pzsz, sm.q);
                TRSP par_trs = identity_trs();
                if (rec.pstep >= 0)
                    par_trs = load_trs(smem + (size_t)rec.pstep * POSE_SLOT_FLOATS, lane);
                store_trs(smem + (size_t)s * POSE_SLOT_FLOATS, lane, composep(par_trs, loc, rec.pstep < 0));
            }
            __syncthreads();
        }
    }

    // -------------------------------------------------------------------------------------------
    // frame-coherent visiting order (counting sort, three launches): histogram of the (clip, frame) buckets the
    // instances' phases fall in -> exclusive scan -> scatter of the instance numbers.  Which instance lands where
    // inside a bucket is decided by atomics and may differ from run to run; the pose does not depend on it.
    // With two or more clips per instance and few enough buckets the key is the pair (bucket of clip 0, bucket of clip 1).
    // -------------------------------------------------------------------------------------------
    constexpr int ORDER_MAX_BUCKETS = 8192; // 32 KB of shared-memory counters per block
    constexpr int ORDER_THREADS     = 512;
    // items per thread: 2 048 per block.  Measured (tools/pose_bench.py, one clip, order + pose): 16 / 8 / 4 / 2 per thread:
    // 125 k instances 0.153 / 0.143 / 0.139 / 0.136 ms, 1 M 0.934 / 0.923 / 0.927 / 0.926 - with 16 a 1 M crowd is 123 blocks,
    // fewer than SMs, each thread walking a chain of dependent loads per item; 4 keeps the per-block tables of the two-clip
    // form (up to 8 192 pair buckets) a fraction of the block's work
#ifndef PCL_ORDER_PER_THREAD
#define PCL_ORDER_PER_THREAD 4
#endif
    constexpr int ORDER_PER_THREAD  = PCL_ORDER_PER_THREAD;
    struct OrderParams
    {
        const ClipRec* clips;
        const int32_t* clip_ids;
        const float*   phase;
        const float*   clip_len;
        int32_t*       hist;   // [nbk] zero on entry of order_hist_kernel; order_scan_kernel re-zeroes it
        int32_t*       cursor; // [nbk]
        int32_t*       order;  // [N]
        long long      N;
        int32_t        nb, nbk, levels, advance;
        float          dt;
    };
    __device__ __forceinline__ int order_key(const OrderParams& p, long long i)
    {
        int key = 0;
        for (int k = 0; k < p.levels; k++)
        {
            const size_t   o  = (size_t)k * p.N + i;
            const ClipRec* cr = p.clips + __ldg(p.clip_ids + o);
            const int      tf = __ldg(&cr->total_frame_m1);
            float          ph = __ldg(p.phase + o);
            if (p.advance == 2 || (p.advance == 1 && k == 0))
            {
                ph += p.dt / __ldg(p.clip_len + o); // what the pose kernel is about to do (animation_component.cpp:19-21)
                ph -= floorf(ph);
            }
            const int lo = min(max((int)floorf(ph * (float)tf), 0), tf);
            key          = key * p.nb + __ldg(&cr->bucket_base) + lo;
        }
        return key;
    }
    __global__ void __launch_bounds__(ORDER_THREADS) order_hist_kernel(const OrderParams p)
    {
        extern __shared__ int32_t osm[];
        for (int b = threadIdx.x; b < p.nbk; b += ORDER_THREADS)
            osm[b] = 0;
        __syncthreads();
        const long long base = (long long)blockIdx.x * (ORDER_THREADS * ORDER_PER_THREAD);
#pragma unroll 4
        for (int j = 0; j < ORDER_PER_THREAD; j++)
        {
            const long long i = base + (long long)j * ORDER_THREADS + threadIdx.x;
            if (i < p.N)
                atomicAdd(&osm[order_key(p, i)], 1);
        }
        __syncthreads();
        for (int b = threadIdx.x; b < p.nbk; b += ORDER_THREADS)
            if (osm[b])
                atomicAdd(&p.hist[b], osm[b]);
    }
    __global__ void __launch_bounds__(1024) order_scan_kernel(const OrderParams p)
    {
        __shared__ int32_t part[1024];
        const int per = (p.nbk + 1023) / 1024;
        const int lo = threadIdx.x * per, hi = min(p.nbk, lo + per);
        int sum = 0;
        for (int b = lo; b < hi; b++)
            sum += p.hist[b];
        part[threadIdx.x] = sum;
        __syncthreads();
        for (int d = 1; d < 1024; d <<= 1) // Hillis-Steele inclusive scan of the 1024 partial sums
        {
            const int v = threadIdx.x >= d ? part[threadIdx.x - d] : 0;
            __syncthreads();
            part[threadIdx.x] += v;
            __syncthreads();
        }
        int run = part[threadIdx.x] - sum;
        for (int b = lo; b < hi; b++)
        {
            const int h = p.hist[b];
            p.cursor[b] = run, p.hist[b] = 0;
            run += h;
        }
    }
    __global__ void __launch_bounds__(ORDER_THREADS) order_scatter_kernel(const OrderParams p)
    {
        extern __shared__ int32_t osm[]; // [nbk] counters, then [nbk] this block's first position in each bucket
        int32_t* cnt = osm;
        int32_t* pos = osm + p.nbk;
        for (int b = threadIdx.x; b < p.nbk; b += ORDER_THREADS)
            cnt[b] = 0;
        __syncthreads();
        const long long base = (long long)blockIdx.x * (ORDER_THREADS * ORDER_PER_THREAD);
        int keys[ORDER_PER_THREAD], rank[ORDER_PER_THREAD];
#pragma unroll
        for (int j = 0; j < ORDER_PER_THREAD; j++)
        {
            const long long i = base + (long long)j * ORDER_THREADS + threadIdx.x;
            keys[j] = i < p.N ? order_key(p, i) : -1;
            rank[j] = keys[j] >= 0 ? atomicAdd(&cnt[keys[j]], 1) : 0;
        }
        __syncthreads();
        for (int b = threadIdx.x; b < p.nbk; b += ORDER_THREADS)
            if (cnt[b])
                pos[b] = atomicAdd(&p.cursor[b], cnt[b]);
        __syncthreads();
#pragma unroll
        for (int j = 0; j < ORDER_PER_THREAD; j++)
        {
            const long long i = base + (long long)j * ORDER_THREADS + threadIdx.x;
            if (keys[j] >= 0)
                p.order[pos[keys[j]] + rank[j]] = (int32_t)i;
        }
    }

    // -------------------------------------------------------------------------------------------
    // phase update alone (pcl_batch_advance), animation_component.cpp:19-21
    // -------------------------------------------------------------------------------------------
    __global__ void advance_kernel(float* phase, const float* clip_len, long long n_first, long long n_total, float dt)
    {
        long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= n_total)
            return;
        (void)n_first;
        float ph = phase[i];
        ph += dt / clip_len[i];
        ph -= floorf(ph);
        phase[i] = ph;
    }

    // -------------------------------------------------------------------------------------------
    // readback helper: compact palette [N][pitch][3] float4 -> AnimationResult [count][J][16]
    // (Matrix4x4::toMatrix4x4_, core/math/matrix4.h:98-119; row 3 = 0 0 0 1)
    // -------------------------------------------------------------------------------------------
    __global__ void export_kernel(const float4* __restrict__ palette, float4* __restrict__ out, long long first,
                                  long long count, int J, int pitch)
    {
        long long t     = (long long)blockIdx.x * blockDim.x + threadIdx.x; // one thread per output float4
        long long total = count * J * 4;
        if (t >= total)
            return;
        int       r = (int)(t & 3);
        long long e = t >> 2;
        long long i = e / J;
        int       b = (int)(e - i * J);
        out[t] = (r == 3) ? make_float4(0, 0, 0, 1) : palette[((size_t)(first + i) * pitch + b + 1) * 3 + r];
    }

    // compact read-back: rows 0..2 only, [count][J][3] float4 (row 3 of an AnimationResult matrix is the constant 0 0 0 1 and is
    // filled in on the host).  Per instance this is ONE contiguous run of J * 48 bytes on both sides (slots 1..J of the
    // palette row), so consecutive threads read and write consecutive 16-byte units: whole 128-byte lines when `out` is a
    // pinned host buffer written over PCIe.
    __global__ void export3x4_kernel(const float4* __restrict__ palette, float4* __restrict__ out, long long first, long long count,
                                     int J, int pitch)
    {
        const long long t     = (long long)blockIdx.x * blockDim.x + threadIdx.x; // one thread per float4
        const int       per   = J * 3;
        const long long total = count * per;
        if (t >= total)
            return;
        const long long i = t / per;
        const int       u = (int)(t - i * per);
        out[t]            = __ldg(palette + (size_t)(first + i) * pitch * 3 + 3 + u);
    }

    // -------------------------------------------------------------------------------------------
    // render hand-off: compact palette -> the per-draw-call joint SSBO of the reference's mesh passes,
    // joint_matrices[1024 * instance_in_draw + slot] (4x4 row-major, <= 64 instances per draw):
    // function/render/passes/main_camera_pass.cpp:2284-2298, function/render/render_common.h:19-20,70-73,
    // engine/shader/include/constants.h:3-7.  Only the J+1 live slots of an instance are written, like the
    // reference's memcpy loop; one thread per 16-byte row.
    // -------------------------------------------------------------------------------------------
    constexpr int RENDER_MAX_INSTANCES = 64;   // m_mesh_per_drawcall_max_instance_count
    constexpr int RENDER_MAX_JOINTS    = 1024; // m_mesh_vertex_blending_max_joint_count
    __global__ void render_export_kernel(const float4* __restrict__ palette, float4* __restrict__ out, long long first, long long count,
                                         int slots, int pitch, long long draw_first)
    {
        long long t     = (long long)blockIdx.x * blockDim.x + threadIdx.x;
        long long total = count * slots * 4;
        if (t >= total)
            return;
        const int       r    = (int)(t & 3);
        const long long e    = t >> 2;
        const long long i    = e / slots;
        const int       slot = (int)(e - i * slots);
        const long long draw = draw_first + i / RENDER_MAX_INSTANCES;
        const int       in   = (int)(i % RENDER_MAX_INSTANCES);
        const float4    v    = (r == 3) ? make_float4(0, 0, 0, 1) : palette[((size_t)(first + i) * pitch + slot) * 3 + r];
        out[(((size_t)draw * RENDER_MAX_INSTANCES + in) * RENDER_MAX_JOINTS + slot) * 4 + r] = v;
    }

    // -------------------------------------------------------------------------------------------
    // K5: linear-blend skinning, engine/shader/glsl/mesh.vert:72-111
    // -------------------------------------------------------------------------------------------
    struct LbsParams
    {
        const float*  pos;  // [Vt][3]   Vt = V rounded up to LBS_VT (zero padded)
        const float*  nrm;  // [Vt][3]
        const float*  tan;  // [Vt][3]
        const int4*   idx;  // [Vt]      palette slots, gated to 0 where the shader's test fails
        const float4* w;    // [Vt]      renormalised weights, 0 where gated
        const float4* palette;
        float*        out_pos; // [cap][Vp][3]
        float*        out_nt;  // [cap][Vp][6]
        long long     first, count, slot_first;
        int32_t       V, Vp, n_vtiles, pitch, inst_tile, vsplit;
    };

    constexpr int LBS_VT      = 256; // vertices per tile = threads per CTA (one vertex per thread)
    constexpr int LBS_THREADS = 256;

    __host__ __device__ inline size_t lbs_smem_bytes(int inst_tile, int pitch)
    {
        return (size_t)inst_tile * pitch * 48      // palette tile
               + LBS_VT * (36 + 32)                // vertex streams: pos, nrm, tan, idx4, w4
               + (LBS_THREADS / 32) * 2 * 1152     // per-warp double-buffered output staging
               + 64;                               // mbarriers
    }

    __device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
    __device__ __forceinline__ void     mbar_init(uint64_t* bar, int count)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    }
    __device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
    {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
    }
    __device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
    {
        asm volatile("{\n"
                     ".reg .pred p;\n"
                     "W_%=:\n"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                     "@p bra D_%=;\n"
                     "bra W_%=;\n"
                     "D_%=:\n"
                     "}" ::"r"(smem_u32(bar)),
                     "r"(parity)
                     : "memory");
    }
    // TMA bulk copy global -> shared (1-D, no tensor map); completes on the mbarrier.  SASS: UBLKCP
    __device__ __forceinline__ void tma_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
    {
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                     "l"(src), "r"(bytes), "r"(smem_u32(bar))
                     : "memory");
    }
    // TMA bulk copy shared -> global
    __device__ __forceinline__ void tma_store(void* dst, const void* src, uint32_t bytes)
    {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
    }
    __device__ __forceinline__ void tma_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
    template<int N>
    __device__ __forceinline__ void tma_wait_read()
    {
        asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
    }
    __device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

    // CTA = (instance tile, vertex split).  The instance tile's palettes (one contiguous block of the
    // compact palette) are brought into shared memory by ONE TMA bulk copy; each vertex tile's five
    // streams by five more, prefetched one tile ahead (the tile is consumed into registers at once, so a
    // single shared buffer double-buffers against registers).  thread = vertex, loop over the tile's
    // instances: palette rows are gathered from shared memory with 128-bit loads (lanes bound to the same
    // bone broadcast), results are staged per warp and leave either as coalesced 128-bit stores
    // (STORE_TMA = false) or as TMA bulk stores of the warp's contiguous 384 B + 768 B runs (true).
    template<bool NT, bool STORE_TMA>
    __global__ void __launch_bounds__(LBS_THREADS, 2) lbs_kernel(const LbsParams par)
    {
        extern __shared__ __align__(1024) unsigned char lsm[];
        const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
        const int tile  = blockIdx.x / par.vsplit;
        const int split = blockIdx.x - tile * par.vsplit;
        const long long i0 = (long long)tile * par.inst_tile; // first instance of this tile, relative to par.first
        const int       ni = (int)min((long long)par.inst_tile, par.count - i0);

        float4*        pal_s  = reinterpret_cast<float4*>(lsm);
        unsigned char* vbuf   = lsm + (size_t)par.inst_tile * par.pitch * 48;
        float*         vpos   = reinterpret_cast<float*>(vbuf);
        float*         vnrm   = vpos + LBS_VT * 3;
        float*         vtan   = vnrm + LBS_VT * 3;
        int4*          vidx   = reinterpret_cast<int4*>(vtan + LBS_VT * 3);
        float4*        vw     = reinterpret_cast<float4*>(vidx + LBS_VT);
        float*         ostage = reinterpret_cast<float*>(vbuf + LBS_VT * 68) + (size_t)warp * 2 * 288;
        uint64_t*      bars   = reinterpret_cast<uint64_t*>(vbuf + LBS_VT * 68 + (LBS_THREADS / 32) * 2 * 1152);

        // vertex tiles of this split
        const int per   = (par.n_vtiles + par.vsplit - 1) / par.vsplit;
        const int vt_lo = split * per, vt_hi = min(par.n_vtiles, vt_lo + per);
        if (vt_lo >= vt_hi)
            return;

        auto issue_vtile = [&](int vt) {
            const size_t v0 = (size_t)vt * LBS_VT;
            mbar_expect_tx(&bars[1], LBS_VT * 68);
            tma_load(vpos, par.pos + v0 * 3, LBS_VT * 12, &bars[1]);
            tma_load(vnrm, par.nrm + v0 * 3, LBS_VT * 12, &bars[1]);
            tma_load(vtan, par.tan + v0 * 3, LBS_VT * 12, &bars[1]);
            tma_load(vidx, par.idx + v0, LBS_VT * 16, &bars[1]);
            tma_load(vw, par.w + v0, LBS_VT * 16, &bars[1]);
        };

        if (tid == 0)
        {
            mbar_init(&bars[0], 1);
            mbar_init(&bars[1], 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        if (tid == 0)
        {
            const uint32_t pbytes = (uint32_t)ni * par.pitch * 48;
            mbar_expect_tx(&bars[0], pbytes);
            tma_load(pal_s, par.palette + (size_t)(par.first + i0) * par.pitch * 3, pbytes, &bars[0]);
            issue_vtile(vt_lo);
        }
        mbar_wait(&bars[0], 0);

        uint32_t vphase = 0;
        int      obuf   = 0;
        for (int vt = vt_lo; vt < vt_hi; vt++)
        {
            mbar_wait(&bars[1], vphase);
            vphase ^= 1;
            const float3 p  = make_float3(vpos[tid * 3], vpos[tid * 3 + 1], vpos[tid * 3 + 2]);
            float3       n  = make_float3(0, 0, 0), t = make_float3(0, 0, 0);
            if (NT)
            {
                n = make_float3(vnrm[tid * 3], vnrm[tid * 3 + 1], vnrm[tid * 3 + 2]);
                t = make_float3(vtan[tid * 3], vtan[tid * 3 + 1], vtan[tid * 3 + 2]);
            }
            const int4   id = vidx[tid];
            const float4 w  = vw[tid];
            __syncthreads(); // everyone holds its vertex in registers: the buffer is free again
            if (tid == 0 && vt + 1 < vt_hi)
                issue_vtile(vt + 1);

            // influences nobody in the warp uses are skipped (warp-uniform)
            const bool a1 = __any_sync(0xffffffffu, w.y != 0.f), a2 = __any_sync(0xffffffffu, w.z != 0.f),
                       a3 = __any_sync(0xffffffffu, w.w != 0.f);
            const int  o0 = id.x * 3, o1 = id.y * 3, o2 = id.z * 3, o3 = id.w * 3;
            const int  wv0    = vt * LBS_VT + warp * 32;            // first vertex of this warp
            const int  nvalid = min(32, par.Vp - wv0);              // vertices (incl. <= 3 pad) this warp stores
            if (nvalid <= 0)
                continue;

            for (int i = 0; i < ni; i++)
            {
                const float4* P = pal_s + (size_t)i * par.pitch * 3;
                // vertex_blending_matrix = sum_k joint_matrices[idx_k] * w_k (mesh.vert:77-102); a gated
                // influence has idx = 0, w = 0 and adds an exact zero
                float4 m0 = P[o0], m1 = P[o0 + 1], m2 = P[o0 + 2];
                float4 b0 = make_float4(m0.x * w.x, m0.y * w.x, m0.z * w.x, m0.w * w.x);
                float4 b1 = make_float4(m1.x * w.x, m1.y * w.x, m1.z * w.x, m1.w * w.x);
                float4 b2 = make_float4(m2.x * w.x, m2.y * w.x, m2.z * w.x, m2.w * w.x);
#define PCL_ACC(o, ww)                                                                               \
    {                                                                                                \
        m0 = P[o], m1 = P[o + 1], m2 = P[o + 2];                                                     \
        b0.x += m0.x * ww, b0.y += m0.y * ww, b0.z += m0.z * ww, b0.w += m0.w * ww;                  \
        b1.x += m1.x * ww, b1.y += m1.y * ww, b1.z += m1.z * ww, b1.w += m1.w * ww;                  \
        b2.x += m2.x * ww, b2.y += m2.y * ww, b2.z += m2.z * ww, b2.w += m2.w * ww;                  \
    }
                if (a1) PCL_ACC(o1, w.y)
                if (a2) PCL_ACC(o2, w.z)
                if (a3) PCL_ACC(o3, w.w)
#undef PCL_ACC
                // model_position = (B * vec4(p, 1)).xyz (mesh.vert:104)
                const float px = b0.x * p.x + b0.y * p.y + b0.z * p.z + b0.w;
                const float py = b1.x * p.x + b1.y * p.y + b1.z * p.z + b1.w;
                const float pz = b2.x * p.x + b2.y * p.y + b2.z * p.z + b2.w;
                float*      os = ostage + obuf * 288;
                if (STORE_TMA)
                {
                    if (lane == 0)
                        tma_wait_read<1>(); // the bulk store issued from this buffer two instances ago has read it
                    __syncwarp();
                }
                os[lane * 3] = px, os[lane * 3 + 1] = py, os[lane * 3 + 2] = pz;
                if (NT)
                {
                    // normalize(mat3(B) * n), normalize(mat3(B) * t) (mesh.vert:106-110)
                    float nx = b0.x * n.x + b0.y * n.y + b0.z * n.z, ny = b1.x * n.x + b1.y * n.y + b1.z * n.z,
                          nz = b2.x * n.x + b2.y * n.y + b2.z * n.z;
                    float tx = b0.x * t.x + b0.y * t.y + b0.z * t.z, ty = b1.x * t.x + b1.y * t.y + b1.z * t.z,
                          tz = b2.x * t.x + b2.y * t.y + b2.z * t.z;
                    const float rn = 1.0f / sqrtf(nx * nx + ny * ny + nz * nz);
                    const float rt = 1.0f / sqrtf(tx * tx + ty * ty + tz * tz);
                    float2*     on = reinterpret_cast<float2*>(os + 96 + lane * 6);
                    on[0] = make_float2(nx * rn, ny * rn), on[1] = make_float2(nz * rn, tx * rt), on[2] = make_float2(ty * rt, tz * rt);
                }
                const size_t slot = (size_t)(par.slot_first + i0 + i);
                float*       gp   = par.out_pos + (slot * par.Vp + wv0) * 3;
                float*       gn   = NT ? par.out_nt + (slot * par.Vp + wv0) * 6 : nullptr;
                if (STORE_TMA)
                {
                    fence_async_smem();
                    __syncwarp();
                    if (lane == 0)
                    {
                        tma_store(gp, os, nvalid * 12);
                        if (NT)
                            tma_store(gn, os + 96, nvalid * 24);
                        tma_commit();
                    }
                }
                else
                {
                    __syncwarp();
                    const float4* s4 = reinterpret_cast<const float4*>(os);
                    if (lane * 4 < nvalid * 3)
                        reinterpret_cast<float4*>(gp)[lane] = s4[lane];
                    if (NT)
                    {
                        if (lane * 4 < nvalid * 6)
                            reinterpret_cast<float4*>(gn)[lane] = s4[24 + lane];
                        if ((lane + 32) * 4 < nvalid * 6)
                            reinterpret_cast<float4*>(gn)[lane + 32] = s4[24 + 32 + lane];
                    }
                }
                obuf ^= 1;
            }
        }
        if (STORE_TMA)
        {
            if (lane == 0)
                tma_wait_read<0>();
            __syncwarp();
        }
    }

    // -------------------------------------------------------------------------------------------
    // K5, crowd form: lane = instance.
    //
    // The lane = vertex kernel above gathers 4 x 48 B of palette per vertex-instance from shared memory and
    // is bound by the shared-memory pipe (ncu: L1/TEX 97 %, DRAM 42 %; profiles/r1_lbs_v0_*).  A bone's
    // matrix is reused by every vertex bound to it, so here a warp takes 32 INSTANCES and walks a run of
    // consecutive vertices: every lane keeps the four matrices of "its" instance in registers and reloads
    // a slot only when the vertex stream says that slot's bone changes (a warp-uniform decision; the upload
    // packs the influences into slots so that neighbouring vertices - which share bones in any real,
    // fully-expanded mesh - keep their slots).  Vertex data is warp-uniform and read with broadcast loads.
    //   - palettes: 32 instances x pitch x 48 B, one TMA bulk copy per instance into rows whose stride is an
    //     odd number of 16-byte units (conflict-free 128-bit loads with lane = instance)
    //   - vertex records (64 B, VRec): per-warp TMA bulk copies, LBS_I_VSTAGES chunks ahead
    //   - results: 4 vertices are held in registers, written to a per-warp staging tile with 128-bit stores,
    //     and every 8 vertices the tile is flushed with coalesced 128-bit stores: per instance a run of
    //     8 vertices is 96 B of positions + 192 B of normal/tangent, i.e. whole 32-byte sectors.
    // -------------------------------------------------------------------------------------------
    struct alignas(16) VRec // one vertex, warp-uniform
    {
        float px, py, pz, w0;
        float nx, tx, ny, ty; // normal and tangent interleaved: (n.k, t.k) pairs feed the packed fp32 transforms of lbs_cta_kernel
        float nz, tz, w1, w2;
        float    w3;
        uint32_t i01;  // 3 * palette slot of influence 0 | 3 * slot of influence 1 << 16 (16-byte units into a palette row)
        uint32_t i23;
        uint32_t mask; // bit k: influence k's bone differs from the previous vertex's influence k
    };

    struct LbsIParams
    {
        const VRec*   vrec;    // [Vt] zero padded to a multiple of LBS_I_CHUNK * anything the launch needs
        const float4* palette; // [N][pitch][3]
        float*        out_pos; // [cap][Vp][3]
        float*        out_nt;  // [cap][Vp][6]
        long long     first, count, slot_first;
        int32_t       Vp, n_chunks, pitch, prow, vsplit; // prow: shared-memory palette row stride in float4 (odd)
    };

    constexpr int LBS_I_CHUNK   = 8;  // vertices per position flush
    constexpr int LBS_I_VSTAGES = 3;  // vertex-record stages in flight per warp
    constexpr int LBS_I_SCH     = 2;  // chunks per stage: one mbarrier wait (~100 cycles even when complete) per 16 vertices
    constexpr int LBS_I_POSROW  = LBS_I_CHUNK * 3 + 4; // floats; /4 odd -> conflict-free 128-bit column stores
    // normal+tangent staging row: a whole chunk (12 units) or, in the HALF form that trades a little flush
    // efficiency for 50 % more resident warps, 4 vertices (6 units)
    __host__ __device__ constexpr int lbs_i_ntrow(bool half) { return half ? 4 * 6 + 4 : LBS_I_CHUNK * 6 + 4; }

    __host__ __device__ inline size_t lbs_i_smem_bytes(int prow, int warps, bool half)
    {
        return (size_t)32 * prow * 16                                             // palettes
               + (size_t)warps * 32 * (LBS_I_POSROW + lbs_i_ntrow(half)) * 4       // staging tiles
               + (size_t)warps * LBS_I_VSTAGES * LBS_I_SCH * LBS_I_CHUNK * sizeof(VRec) // vertex records
               + 8 * (1 + warps * LBS_I_VSTAGES) + 64;                            // mbarriers
    }

#ifndef PCL_LBS_FFMA2
#define PCL_LBS_FFMA2 1
#endif
    // packed fp32 pairs (sm_100: fma.rn.f32x2, mul.rn.f32x2)
    __device__ __forceinline__ unsigned long long pack2(float a, float b)
    {
        unsigned long long r;
        asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
        return r;
    }
    __device__ __forceinline__ void unpack2(unsigned long long v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
    __device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c)
    {
        unsigned long long d;
        asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
        return d;
    }
    __device__ __forceinline__ unsigned long long mul2(unsigned long long a, unsigned long long b)
    {
        unsigned long long d;
        asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
        return d;
    }

    // one MUFU.RSQ, no denormal rescue sequence (a squared normal below 1e-38 has no direction worth keeping)
    __device__ __forceinline__ float rsqrt_ftz(float x)
    {
        float r;
        asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
        return r;
    }

    // one bone matrix (3 rows) from this lane's palette row; off3 = slot * 3 (16-byte units)
    __device__ __forceinline__ void lbs_i_load(float4 (&m)[3], const float4* prow, uint32_t off3)
    {
        const float4* s = prow + off3;
        m[0] = s[0], m[1] = s[1], m[2] = s[2];
    }

    // skinned streams are written once and never re-read by this engine: streaming (evict-first) stores
#ifndef PCL_ST_MODE
#define PCL_ST_MODE 0 /* measured: plain, .cg and .cs stores are within noise of each other */
#endif
#if PCL_ST_MODE == 1
#define PCL_ST(p, v) __stcs((p), (v))
#elif PCL_ST_MODE == 2
#define PCL_ST(p, v) __stcg((p), (v))
#else
#define PCL_ST(p, v) (*(p) = (v))
#endif
    // Flushes with coalesced 128-bit stores.  Lane -> (staging row, 16-byte unit) maps every quarter-warp
    // onto ONE row (6-unit rows: lanes 6,7 of each quarter idle; 12-unit rows: units 0-7) or onto rows r and
    // r+4 (units 8-11 of the 12-unit rows), which keeps the reads conflict-free with the odd row strides the
    // 128-bit column stores need.  All lane-dependent address parts are hoisted by the caller.
    template<bool FULL, int ROW>
    __device__ __forceinline__ void lbs_i_flush6(const float* tile, int s0, float4* g0, size_t rowstep, size_t off, int lane, int ni)
    {
        const int a = lane >> 3, b = lane & 7;
        if (b < 6)
        {
#pragma unroll
            for (int it = 0; it < 8; it++)
                if (FULL || it * 4 + a < ni)
                    PCL_ST(&g0[off + it * rowstep], *reinterpret_cast<const float4*>(tile + s0 + it * 4 * ROW));
        }
    }
    template<bool FULL, int ROW>
    __device__ __forceinline__ void lbs_i_flush12(const float* snt, int s_nt0, int s_nt2, float4* g_nt0, float4* g_nt2, size_t rowstep_nt,
                                                 size_t chunk_nt, int lane, int ni)
    {
        const int a = lane >> 3, c4 = ((lane >> 2) & 1) * 4;
#pragma unroll
        for (int g = 0; g < 4; g++)
        {
            if (FULL || g * 8 + a < ni)
                PCL_ST(&g_nt0[chunk_nt + (2 * g) * rowstep_nt], *reinterpret_cast<const float4*>(snt + s_nt0 + g * 8 * ROW));
            if (FULL || g * 8 + 4 + a < ni)
                PCL_ST(&g_nt0[chunk_nt + (2 * g + 1) * rowstep_nt], *reinterpret_cast<const float4*>(snt + s_nt0 + (g * 8 + 4) * ROW));
            if (FULL || g * 8 + a + c4 < ni)
                PCL_ST(&g_nt2[chunk_nt + (2 * g) * rowstep_nt], *reinterpret_cast<const float4*>(snt + s_nt2 + g * 8 * ROW));
        }
    }

    template<bool NT, int WARPS, bool HALF>
    __global__ void __launch_bounds__(WARPS * 32, 1) lbs_inst_kernel(const LbsIParams par)
    {
        constexpr int LBS_I_WARPS = WARPS;
        constexpr int LBS_I_NTROW = lbs_i_ntrow(HALF);
        extern __shared__ __align__(1024) unsigned char lsm[];
        const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
        const int tile  = blockIdx.x / par.vsplit;
        const int split = blockIdx.x - tile * par.vsplit;
        const long long i0 = (long long)tile * 32;
        const int       ni = (int)min(32LL, par.count - i0);

        float4*   pal_s = reinterpret_cast<float4*>(lsm);
        float*    spos  = reinterpret_cast<float*>(lsm + (size_t)32 * par.prow * 16) + (size_t)warp * 32 * (LBS_I_POSROW + LBS_I_NTROW);
        float*    snt   = spos + 32 * LBS_I_POSROW;
        VRec*     vring = reinterpret_cast<VRec*>(lsm + (size_t)32 * par.prow * 16 + (size_t)LBS_I_WARPS * 32 * (LBS_I_POSROW + LBS_I_NTROW) * 4) +
                      (size_t)warp * LBS_I_VSTAGES * LBS_I_SCH * LBS_I_CHUNK;
        uint64_t* bars  = reinterpret_cast<uint64_t*>(lsm + (size_t)32 * par.prow * 16 + (size_t)LBS_I_WARPS * 32 * (LBS_I_POSROW + LBS_I_NTROW) * 4 +
                                                     (size_t)LBS_I_WARPS * LBS_I_VSTAGES * LBS_I_SCH * LBS_I_CHUNK * sizeof(VRec));
        uint64_t* vbar  = bars + 1 + warp * LBS_I_VSTAGES;

        // this warp's run of vertex chunks
        const int per_split = (par.n_chunks + par.vsplit - 1) / par.vsplit;
        const int s_lo = split * per_split, s_hi = min(par.n_chunks, s_lo + per_split);
        const int per_warp = (max(s_hi - s_lo, 0) + LBS_I_WARPS - 1) / LBS_I_WARPS;
        const int c_lo = s_lo + warp * per_warp, c_hi = min(s_hi, c_lo + per_warp);

        if (tid == 0)
        {
            mbar_init(&bars[0], 1);
            for (int i = 0; i < LBS_I_WARPS * LBS_I_VSTAGES; i++)
                mbar_init(&bars[1 + i], 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        if (tid == 0)
        {
            const uint32_t row_bytes = (uint32_t)par.pitch * 48;
            mbar_expect_tx(&bars[0], row_bytes * ni);
            for (int i = 0; i < ni; i++)
                tma_load(pal_s + (size_t)i * par.prow, par.palette + (size_t)(par.first + i0 + i) * par.pitch * 3, row_bytes, &bars[0]);
        }
        constexpr uint32_t STAGE_BYTES = LBS_I_SCH * LBS_I_CHUNK * sizeof(VRec); // the record array is padded, a stage may overrun c_hi
        if (lane == 0)
            for (int st = 0; st < LBS_I_VSTAGES && c_lo + st * LBS_I_SCH < c_hi; st++)
            {
                mbar_expect_tx(&vbar[st], STAGE_BYTES);
                tma_load(vring + st * LBS_I_SCH * LBS_I_CHUNK, par.vrec + (size_t)(c_lo + st * LBS_I_SCH) * LBS_I_CHUNK, STAGE_BYTES, &vbar[st]);
            }
        mbar_wait(&bars[0], 0);
        if (c_lo >= c_hi)
            return;

        const int     li   = min(lane, ni - 1); // idle lanes shadow the tile's last instance and never store
        const float4* prow = pal_s + (size_t)li * par.prow;
        // the run starts cold: bring in the four bones of its first vertex
        float4 m[4][3];
        {
            const uint4 d = __ldg(reinterpret_cast<const uint4*>(&par.vrec[(size_t)c_lo * LBS_I_CHUNK].w3));
            lbs_i_load(m[0], prow, d.y & 0xffffu), lbs_i_load(m[1], prow, d.y >> 16);
            lbs_i_load(m[2], prow, d.z & 0xffffu), lbs_i_load(m[3], prow, d.z >> 16);
        }
        // lane-dependent parts of the flush addresses (see lbs_i_flush)
        const int    fa = lane >> 3, fb = lane & 7, fc4 = ((lane >> 2) & 1) * 4, fd = lane & 3;
        const int    s_pos0 = fa * LBS_I_POSROW + fb * 4, s_nt0 = fa * LBS_I_NTROW + fb * 4, s_nt2 = (fa + fc4) * LBS_I_NTROW + (8 + fd) * 4;
        const size_t S0 = (size_t)(par.slot_first + i0);
        float4*      g_pos0 = reinterpret_cast<float4*>(par.out_pos + (S0 + fa) * par.Vp * 3) + fb;
        float4*      g_nt0  = NT ? reinterpret_cast<float4*>(par.out_nt + (S0 + fa) * par.Vp * 6) + fb : nullptr;
        float4*      g_nt2  = NT ? reinterpret_cast<float4*>(par.out_nt + (S0 + fa + fc4) * par.Vp * 6) + 8 + fd : nullptr;
        const size_t rowstep_pos = (size_t)par.Vp * 3, rowstep_nt = (size_t)par.Vp * 6; // float4 units per 4 instances

        for (int c = c_lo; c < c_hi; c++)
        {
            const int sidx = (c - c_lo) / LBS_I_SCH, sub = (c - c_lo) - sidx * LBS_I_SCH;
            const int st   = sidx % LBS_I_VSTAGES;
            if (sub == 0)
                mbar_wait(&vbar[st], (uint32_t)(sidx / LBS_I_VSTAGES) & 1);
            const VRec* vr = vring + (st * LBS_I_SCH + sub) * LBS_I_CHUNK;
            // records are fetched one vertex ahead so that the broadcast loads overlap the previous vertex's math
            float4 na = *reinterpret_cast<const float4*>(&vr[0].px), nb = *reinterpret_cast<const float4*>(&vr[0].nx);
            float4 nt_ = *reinterpret_cast<const float4*>(&vr[0].nz);
            uint4  nd = *reinterpret_cast<const uint4*>(&vr[0].w3);
#pragma unroll
            for (int half = 0; half < 2; half++)
            {
                float4 op[3], on[6];
                float* opf = reinterpret_cast<float*>(op);
                float* onf = reinterpret_cast<float*>(on);
#pragma unroll
                for (int j = 0; j < 4; j++)
                {
                    const float4 a = na;
                    const float4 b = make_float4(nb.x, nb.z, nt_.x, nt_.z), t = make_float4(nb.y, nb.w, nt_.y, nt_.w); // {n, w1}, {t, w2}
                    const uint4  d = nd;
                    if (half * 4 + j + 1 < LBS_I_CHUNK)
                    {
                        na  = *reinterpret_cast<const float4*>(&vr[half * 4 + j + 1].px);
                        nb  = *reinterpret_cast<const float4*>(&vr[half * 4 + j + 1].nx);
                        nt_ = *reinterpret_cast<const float4*>(&vr[half * 4 + j + 1].nz);
                        nd  = *reinterpret_cast<const uint4*>(&vr[half * 4 + j + 1].w3);
                    }
                    if (d.w) // some slot changes bone at this vertex (warp-uniform, rare in a coherent mesh)
                    {
                        if (d.w & 1) lbs_i_load(m[0], prow, d.y & 0xffffu);
                        if (d.w & 2) lbs_i_load(m[1], prow, d.y >> 16);
                        if (d.w & 4) lbs_i_load(m[2], prow, d.z & 0xffffu);
                        if (d.w & 8) lbs_i_load(m[3], prow, d.z >> 16);
                    }
                    const float w0 = a.w, w1 = b.w, w2 = t.w, w3 = __uint_as_float(d.x);
                    // vertex_blending_matrix = sum_k joint_matrices[idx_k] * w_k (mesh.vert:77-102), in slot order
#if PCL_LBS_FFMA2
                    // Blackwell packed fp32 (fma.rn.f32x2 / SASS FFMA2): two matrix entries per instruction, same
                    // rounding as two scalar FMAs
                    float4 B[3];
                    {
                        const unsigned long long ww0 = pack2(w0, w0), ww1 = pack2(w1, w1), ww2 = pack2(w2, w2), ww3 = pack2(w3, w3);
#pragma unroll
                        for (int r = 0; r < 3; r++)
                        {
                            unsigned long long lo = mul2(pack2(m[0][r].x, m[0][r].y), ww0), hi = mul2(pack2(m[0][r].z, m[0][r].w), ww0);
                            lo = fma2(pack2(m[1][r].x, m[1][r].y), ww1, lo), hi = fma2(pack2(m[1][r].z, m[1][r].w), ww1, hi);
                            lo = fma2(pack2(m[2][r].x, m[2][r].y), ww2, lo), hi = fma2(pack2(m[2][r].z, m[2][r].w), ww2, hi);
                            lo = fma2(pack2(m[3][r].x, m[3][r].y), ww3, lo), hi = fma2(pack2(m[3][r].z, m[3][r].w), ww3, hi);
                            unpack2(lo, B[r].x, B[r].y), unpack2(hi, B[r].z, B[r].w);
                        }
                    }
#else
                    float4 B[3];
#pragma unroll
                    for (int r = 0; r < 3; r++)
                    {
                        B[r].x = m[0][r].x * w0, B[r].y = m[0][r].y * w0, B[r].z = m[0][r].z * w0, B[r].w = m[0][r].w * w0;
                        B[r].x += m[1][r].x * w1, B[r].y += m[1][r].y * w1, B[r].z += m[1][r].z * w1, B[r].w += m[1][r].w * w1;
                        B[r].x += m[2][r].x * w2, B[r].y += m[2][r].y * w2, B[r].z += m[2][r].z * w2, B[r].w += m[2][r].w * w2;
                        B[r].x += m[3][r].x * w3, B[r].y += m[3][r].y * w3, B[r].z += m[3][r].z * w3, B[r].w += m[3][r].w * w3;
                    }
#endif
                    // model_position = (B * vec4(p, 1)).xyz (mesh.vert:104)
                    opf[j * 3 + 0] = B[0].x * a.x + B[0].y * a.y + B[0].z * a.z + B[0].w;
                    opf[j * 3 + 1] = B[1].x * a.x + B[1].y * a.y + B[1].z * a.z + B[1].w;
                    opf[j * 3 + 2] = B[2].x * a.x + B[2].y * a.y + B[2].z * a.z + B[2].w;
                    if (NT)
                    {
                        // normalize(mat3(B) * n), normalize(mat3(B) * t) (mesh.vert:106-110)
                        const float nx = B[0].x * b.x + B[0].y * b.y + B[0].z * b.z, ny = B[1].x * b.x + B[1].y * b.y + B[1].z * b.z,
                                    nz = B[2].x * b.x + B[2].y * b.y + B[2].z * b.z;
                        const float tx = B[0].x * t.x + B[0].y * t.y + B[0].z * t.z, ty = B[1].x * t.x + B[1].y * t.y + B[1].z * t.z,
                                    tz = B[2].x * t.x + B[2].y * t.y + B[2].z * t.z;
                        const float rn = rsqrt_ftz(nx * nx + ny * ny + nz * nz);
                        const float rt = rsqrt_ftz(tx * tx + ty * ty + tz * tz);
                        onf[j * 6 + 0] = nx * rn, onf[j * 6 + 1] = ny * rn, onf[j * 6 + 2] = nz * rn;
                        onf[j * 6 + 3] = tx * rt, onf[j * 6 + 4] = ty * rt, onf[j * 6 + 5] = tz * rt;
                    }
                }
                float4* sp = reinterpret_cast<float4*>(spos + lane * LBS_I_POSROW + half * 12);
                sp[0] = op[0], sp[1] = op[1], sp[2] = op[2];
                if (NT)
                {
                    float4* sn = reinterpret_cast<float4*>(snt + lane * LBS_I_NTROW + (HALF ? 0 : half * 24));
#pragma unroll
                    for (int q = 0; q < 6; q++)
                        sn[q] = on[q];
                    if (HALF)
                    {
                        __syncwarp();
                        const size_t off = (size_t)c * (LBS_I_CHUNK * 6 / 4) + half * 6;
                        if (ni == 32)
                            lbs_i_flush6<true, LBS_I_NTROW>(snt, s_nt0, g_nt0, rowstep_nt, off, lane, ni);
                        else
                            lbs_i_flush6<false, LBS_I_NTROW>(snt, s_nt0, g_nt0, rowstep_nt, off, lane, ni);
                        __syncwarp();
                    }
                }
            }
            __syncwarp();
            // the records of this stage are consumed: refill it with the stage LBS_I_VSTAGES ahead
            if (lane == 0 && sub == LBS_I_SCH - 1 && c_lo + (sidx + LBS_I_VSTAGES) * LBS_I_SCH < c_hi)
            {
                mbar_expect_tx(&vbar[st], STAGE_BYTES);
                tma_load(vring + st * LBS_I_SCH * LBS_I_CHUNK, par.vrec + (size_t)(c_lo + (sidx + LBS_I_VSTAGES) * LBS_I_SCH) * LBS_I_CHUNK, STAGE_BYTES,
                         &vbar[st]);
            }
            const size_t chunk_pos = (size_t)c * (LBS_I_CHUNK * 3 / 4), chunk_nt = (size_t)c * (LBS_I_CHUNK * 6 / 4);
            if (ni == 32)
            {
                lbs_i_flush6<true, LBS_I_POSROW>(spos, s_pos0, g_pos0, rowstep_pos, chunk_pos, lane, ni);
                if (NT && !HALF)
                    lbs_i_flush12<true, LBS_I_NTROW>(snt, s_nt0, s_nt2, g_nt0, g_nt2, rowstep_nt, chunk_nt, lane, ni);
            }
            else
            {
                lbs_i_flush6<false, LBS_I_POSROW>(spos, s_pos0, g_pos0, rowstep_pos, chunk_pos, lane, ni);
                if (NT && !HALF)
                    lbs_i_flush12<false, LBS_I_NTROW>(snt, s_nt0, s_nt2, g_nt0, g_nt2, rowstep_nt, chunk_nt, lane, ni);
            }
            __syncwarp();
        }
    }

    // -------------------------------------------------------------------------------------------
    // K5, crowd form with a TMA store pipeline (warp-specialised).
    //
    // lbs_inst_kernel moves every result through shared memory twice with LSU instructions (column stores,
    // row loads, global stores: ~29 of its ~40 shared/global wavefronts per warp-vertex) and ends up bound by
    // the L1/TEX pipe.  Here the read-back half is handed to the TMA engine: the CTA's 8 compute warps work on
    // the SAME 32 instances and on the same 32-vertex "superchunk" (warp w owns vertices 4w..4w+3 of it) and
    // write their results into one CTA-wide staging tile; a ninth warp then issues NINE tensor-map bulk stores
    // per superchunk (cp.async.bulk.tensor.2d, SASS UTMASTG: boxes of 32 instances x 128 bytes; 3 for the
    // positions, 6 for normal+tangent).  The tile uses the TMA 128-byte swizzle, which is exactly what makes the
    // lane = instance 128-bit column stores bank-conflict free (16-byte unit u of row r lives at unit u ^ (r & 7)).
    // (1-D bulk stores of the per-instance runs were tried first: 64 per superchunk, 2.0 TB/s - the non-tensor
    // bulk path moves only ~14 GB/s per SM.)  Two tiles and named barriers (full/empty) decouple the two sides;
    // the store warp also refills the CTA-wide vertex-record ring (TMA bulk loads, complete_tx on mbarriers).
    // Bone matrices stay in registers exactly as in lbs_inst_kernel; because a warp now visits vertices
    // 4w..4w+3 of every superchunk, the upload packs the influence slots along that visiting order (`vrec_cta`).
    // A ragged last tile is clipped by the TMA itself: for such launches the launcher encodes tensor maps whose row count
    // is slot_first + count, and the store boxes' rows beyond it are dropped.
    // -------------------------------------------------------------------------------------------
    constexpr int LBS_C_WARPS   = 8;                       // compute warps
    constexpr int LBS_C_THREADS = (LBS_C_WARPS + 1) * 32;  // + 1 store / load warp
    constexpr int LBS_C_VPW     = 4;                       // consecutive vertices a (warp, stream) owns in every superchunk
    constexpr int LBS_C_TILE_BYTES_POS = 3 * 32 * 128;     // per staging tile: 32 lanes x 4 vertices x 8 warps x 12 B
    constexpr int LBS_C_TILE_BYTES_NT  = 6 * 32 * 128;
    constexpr int LBS_C_RSTAGE_BYTES   = 8192;             // one vertex-record stage (128 records)
    constexpr int LBS_C_BLK_VERTS      = 1024;             // compact form: vertices per block (32 / 16 / 8 superchunks for T = 32 / 16 / 8)
    // A tile holds T = 32, 16 or 8 INSTANCES; a warp then carries S = 32 / T vertex streams side by side (lanes
    // [st*T, st*T+T) = the T instances of stream st), which is what lets skeletons whose 32 palettes do not fit in shared
    // memory (more than 82 bones) still use this kernel: 16 palettes fit up to ~165 bones, 8 up to ~340.
    // Superchunk = 32 S vertices; "virtual warp" vw = warp * S + st owns vertices 4 vw .. 4 vw + 3 of it.
    __host__ __device__ constexpr int lbs_c_scv(int T) { return 32 * (32 / T); }                  // vertices per superchunk
    __host__ __device__ constexpr int lbs_c_rsc(int T) { return LBS_C_RSTAGE_BYTES / (lbs_c_scv(T) * 64); } // superchunks per record stage
    // staging tiles in flight between the compute warps and the TMA stores: 3 when the palettes leave room, else 2;
    // the vertex-record ring gets the stage the third tile gives up
    __host__ __device__ constexpr int lbs_c_rstages(int tiles) { return tiles >= 3 ? 2 : 3; }

    __host__ __device__ inline size_t lbs_c_tile_bytes(bool nt) { return (size_t)LBS_C_TILE_BYTES_POS + (nt ? LBS_C_TILE_BYTES_NT : 0); }
    __host__ __device__ inline size_t lbs_c_smem_bytes(int prow, bool nt, int tiles, int T)
    {
        return tiles * lbs_c_tile_bytes(nt)                                            // staging tiles (1024-byte aligned, first)
               + (size_t)T * prow * 16                                                 // palettes
               + (size_t)lbs_c_rstages(tiles) * LBS_C_RSTAGE_BYTES                     // vertex-record ring
               + 8 * (1 + lbs_c_rstages(tiles)) + 64;                                  // mbarriers
    }

    __device__ __forceinline__ void named_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
    __device__ __forceinline__ void named_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }
    // 2-D tensor-map store shared -> global (SASS UTMASTG); c0 = innermost coordinate (floats), c1 = instance slot
    __device__ __forceinline__ void tma_store_2d(const void* tmap, int c0, int c1, const void* src)
    {
        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(src))
                     : "memory");
    }

    struct LbsCParams
    {
        const VRec*   vrec;    // packed along the (warp, stream) visiting order for this T, padded to whole record stages
        const float4* palette; // [N][pitch][3]
        long long     first, slot_first, count; // count need not be a multiple of T: see the last tile below
        int32_t       n_tiles, Vp, n_sc, pitch, prow, vsplit;
        int32_t       per_sc;  // superchunks per CTA (a multiple of blk_sc in the compact form)
        // compact form (big skeletons, coherent meshes): instead of whole palettes the CTA keeps, per instance, only the rows
        // the current BLOCK of LBS_C_BLK_VERTS vertices references; the records carry block-local row numbers
        int32_t         blk_sc;    // superchunks per block, 0 = whole palettes
        int32_t         blk_cap;   // row stride of blk_slots
        const uint16_t* blk_slots; // [n_blocks][blk_cap] palette slots of each block, ascending
        const int32_t*  blk_cnt;   // [n_blocks]
    };

    // COMPACT (block-local palette rows, see LbsCParams::blk_sc) is a template parameter: the whole-palette form - the headline's
    // kernel, which runs at the board's power cap where every instruction counts - carries none of its code.
    template<bool NT, int LBS_C_TILES, int T, bool COMPACT>
    __global__ void __launch_bounds__(LBS_C_THREADS, 1)
        lbs_cta_kernel(const LbsCParams par, const __grid_constant__ CUtensorMap tm_pos, const __grid_constant__ CUtensorMap tm_nt)
    {
        constexpr int S             = 32 / T;               // vertex streams per warp
        constexpr int SCV           = lbs_c_scv(T);         // vertices per superchunk
        constexpr int RSC           = lbs_c_rsc(T);         // superchunks per record stage
        constexpr int LBS_C_RSTAGES = lbs_c_rstages(LBS_C_TILES);
        constexpr int SUB           = T * 128;              // one swizzled sub-tile: T instances x 128 bytes
        constexpr int POS_SUBS = 3 * S, NT_SUBS = 6 * S;    // 384 S and 768 S bytes per instance and superchunk
        extern __shared__ __align__(1024) unsigned char lsm[];
        const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
        const int tile  = blockIdx.x / par.vsplit;
        const int split = blockIdx.x - tile * par.vsplit;
        const long long i0 = (long long)tile * T;

        const size_t   TILE  = lbs_c_tile_bytes(NT);
        unsigned char* tiles = lsm;
        float4*        pal_s = reinterpret_cast<float4*>(lsm + LBS_C_TILES * TILE);
        VRec*          vring = reinterpret_cast<VRec*>(lsm + LBS_C_TILES * TILE + (size_t)T * par.prow * 16);
        uint64_t*      bars  = reinterpret_cast<uint64_t*>(reinterpret_cast<unsigned char*>(vring) + LBS_C_RSTAGES * LBS_C_RSTAGE_BYTES);
        uint64_t*      rbar  = bars + 1;

        // this CTA's run of superchunks
        const int per  = par.per_sc;
        const int s_lo = split * per, s_hi = min(par.n_sc, s_lo + per);
        if (s_lo >= s_hi)
            return;
        const int      n_steps = s_hi - s_lo;
        constexpr bool compact = COMPACT;
        // named barriers: 1 + b = tile b full (compute arrives, store warp syncs); 1 + TILES + b = tile b empty (store arrives, compute syncs)

        if (tid == 0)
        {
            mbar_init(&bars[0], 1);
            for (int i = 0; i < LBS_C_RSTAGES; i++)
                mbar_init(&rbar[i], 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();

        if (warp == LBS_C_WARPS)
        {
            // ---------------- store / load warp ----------------
            if (lane == 0)
            {
                if (!compact)
                {
                    const uint32_t row_bytes = (uint32_t)par.pitch * 48;
                    mbar_expect_tx(&bars[0], row_bytes * T);
                    // a ragged last tile: the missing instances shadow the last real one; their rows lie beyond the row count of
                    // the tensor maps the launcher built for this launch, so the stores drop them
                    for (int i = 0; i < T; i++)
                        tma_load(pal_s + (size_t)i * par.prow, par.palette + (size_t)(par.first + min(i0 + i, par.count - 1)) * par.pitch * 3,
                                 row_bytes, &bars[0]);
                }
                for (int st = 0; st < LBS_C_RSTAGES && st * RSC < n_steps; st++)
                {
                    mbar_expect_tx(&rbar[st], LBS_C_RSTAGE_BYTES);
                    tma_load(reinterpret_cast<unsigned char*>(vring) + st * LBS_C_RSTAGE_BYTES, par.vrec + (size_t)(s_lo + st * RSC) * SCV,
                             LBS_C_RSTAGE_BYTES, &rbar[st]);
                }
            }
            const int row0 = (int)(par.slot_first + i0);
            for (int s = 0; s < n_steps; s++)
            {
                const int b = s % LBS_C_TILES;
                named_sync(1 + b, LBS_C_THREADS); // all compute warps have written (and fenced) tile b
                const int v0 = (s_lo + s) * SCV;
                for (int op = lane; op < POS_SUBS + (NT ? NT_SUBS : 0); op += 32)
                {
                    if (op < POS_SUBS)
                        tma_store_2d(&tm_pos, v0 * 3 + op * 32, row0, tiles + b * TILE + op * SUB);
                    else
                        tma_store_2d(&tm_nt, v0 * 6 + (op - POS_SUBS) * 32, row0, tiles + b * TILE + op * SUB);
                }
                tma_commit();
                // the records of superchunk s are consumed by everybody: refill the stage when its last superchunk is done
                if (lane == 0 && (s % RSC) == RSC - 1)
                {
                    const int nxt = s / RSC + LBS_C_RSTAGES; // stage index to fetch into the slot just freed
                    if (nxt * RSC < n_steps)
                    {
                        const int st = (s / RSC) % LBS_C_RSTAGES;
                        mbar_expect_tx(&rbar[st], LBS_C_RSTAGE_BYTES);
                        tma_load(reinterpret_cast<unsigned char*>(vring) + st * LBS_C_RSTAGE_BYTES, par.vrec + (size_t)(s_lo + nxt * RSC) * SCV,
                                 LBS_C_RSTAGE_BYTES, &rbar[st]);
                    }
                }
                if (s >= LBS_C_TILES - 1)
                {
                    tma_wait_read<LBS_C_TILES - 1>(); // the stores issued for superchunk s-(TILES-1) have read their tile ...
                    __syncwarp();
                    named_arrive(1 + LBS_C_TILES + (s + 1) % LBS_C_TILES, LBS_C_THREADS); // ... which superchunk s+1 may now overwrite
                }
            }
            asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); // the stores themselves, not only their reads of the tiles
            return;
        }

        // ---------------- compute warps ----------------
        const int     row  = lane % T;          // instance within the tile
        const int     vw   = warp * S + lane / T; // virtual warp: which 4 vertices of every superchunk
        const float4* prow = pal_s + (size_t)row * par.prow;
        float4 m[4][3];
        if (!compact)
        {
            mbar_wait(&bars[0], 0);
            const uint4 d = __ldg(reinterpret_cast<const uint4*>(&par.vrec[(size_t)s_lo * SCV + vw * LBS_C_VPW].w3));
            lbs_i_load(m[0], prow, d.y & 0xffffu), lbs_i_load(m[1], prow, d.y >> 16);
            lbs_i_load(m[2], prow, d.z & 0xffffu), lbs_i_load(m[3], prow, d.z >> 16);
        }
        else
        {
#pragma unroll
            for (int k = 0; k < 4; k++)
                m[k][0] = m[k][1] = m[k][2] = make_float4(0.f, 0.f, 0.f, 0.f); // every stream's first record of a block reloads all four
        }
        // swizzled staging offsets of this lane for the 16-byte units its virtual warp owns: positions: units 3 vw .. 3 vw + 2,
        // normal+tangent: units 6 vw .. 6 vw + 5; unit U -> sub-tile U >> 3, chunk (U & 7) ^ (row & 7)
        uint32_t off_p[3], off_n[6];
#pragma unroll
        for (int q = 0; q < 3; q++)
        {
            const int U = vw * 3 + q;
            off_p[q]    = (U >> 3) * SUB + row * 128 + (((U & 7) ^ (row & 7)) << 4);
        }
#pragma unroll
        for (int q = 0; q < 6; q++)
        {
            const int U = vw * 6 + q;
            off_n[q]    = (POS_SUBS + (U >> 3)) * SUB + row * 128 + (((U & 7) ^ (row & 7)) << 4);
        }
        for (int s = 0; s < n_steps; s++)
        {
            const int b = s % LBS_C_TILES;
            if (compact && ((s_lo + s) % par.blk_sc) == 0)
            {
                // A new block of vertices: bring in, for each of the tile's instances, the palette rows this block references
                // (a few dozen of a big skeleton's hundreds: what lets 32 instances share a CTA where 32 whole palettes would
                // not fit).  The compute warps copy them - 48-byte rows from all over each instance's palette, far too many and
                // too small for bulk copies - between two rendezvous: nobody still reads the old rows, nobody reads half-written
                // new ones.  ~60 KB per ~1.2 MB of output: a bubble of a few percent.
                const int       blk  = (s_lo + s) / par.blk_sc;
                const uint16_t* list = par.blk_slots + (size_t)blk * par.blk_cap;
                const int       nu   = __ldg(par.blk_cnt + blk) * 3; // 16-byte units per instance
                // where this lane's units come from is the same for every instance: looked up once (up to 8 units per lane:
                // 82 rows x 3 units is the most a tile of 32 holds), then every row copy is a burst of independent loads
                constexpr int GU = 8;
                int           goff[GU];
#pragma unroll
                for (int k = 0; k < GU; k++)
                {
                    const int u = lane + 32 * k, q = u / 3;
                    goff[k]     = u < nu ? (int)__ldg(list + q) * 3 + (u - q * 3) : -1;
                }
                named_sync(1 + 2 * LBS_C_TILES, LBS_C_WARPS * 32);
                for (int inst = warp; inst < T; inst += LBS_C_WARPS)
                {
                    const float4* src = par.palette + (size_t)(par.first + min(i0 + inst, par.count - 1)) * par.pitch * 3;
                    float4*       dst = pal_s + (size_t)inst * par.prow;
                    float4        v[GU];
#pragma unroll
                    for (int k = 0; k < GU; k++)
                        if (goff[k] >= 0)
                            v[k] = __ldg(src + goff[k]);
#pragma unroll
                    for (int k = 0; k < GU; k++)
                        if (goff[k] >= 0)
                            dst[lane + 32 * k] = v[k];
                    for (int u = lane + 32 * GU; u < nu; u += 32) // (tiles of 16 / 8 instances can hold more rows)
                    {
                        const int q = u / 3;
                        dst[u]      = __ldg(src + (int)__ldg(list + q) * 3 + (u - q * 3));
                    }
                }
                named_sync(1 + 2 * LBS_C_TILES, LBS_C_WARPS * 32);
            }
            const int ridx = s / RSC, rsub = s - ridx * RSC, rst = ridx % LBS_C_RSTAGES;
            if (rsub == 0)
                mbar_wait(&rbar[rst], (uint32_t)(ridx / LBS_C_RSTAGES) & 1);
            const VRec* vr = reinterpret_cast<const VRec*>(reinterpret_cast<const unsigned char*>(vring) + rst * LBS_C_RSTAGE_BYTES) + rsub * SCV + vw * LBS_C_VPW;
            float4 op[3], on[6];
            float* opf = reinterpret_cast<float*>(op);
            float* onf = reinterpret_cast<float*>(on);
            float4 na = *reinterpret_cast<const float4*>(&vr[0].px), nb = *reinterpret_cast<const float4*>(&vr[0].nx);
            float4 nt_ = *reinterpret_cast<const float4*>(&vr[0].nz);
            uint4  nd = *reinterpret_cast<const uint4*>(&vr[0].w3);
            // The previous superchunk's tile is handed to the store warp HERE, not right behind its column stores: the proxy
            // fence (MEMBAR + FENCE.VIEW.ASYNC) and bar.arrive wait for those stores to drain - 11 % of a compute warp's time
            // when issued at once (profiles/r1_lbs_v6_*).  Behind the record loads of the next superchunk the drain has had a few
            // dozen cycles: +1 %.  Any later (behind the first vertex's matrix reloads -1.3 %, behind its blend -1 %, behind the whole
            // vertex -3 %) the tensor stores start too late.
            // (Only with whole-warp tiles: with 2 or 4 vertex streams per warp the late hand-over costs 3 % / 20 %.)
            constexpr bool LATE_HANDOVER = T == 32;
            if (LATE_HANDOVER && s > 0)
            {
                fence_async_smem();
                named_arrive(1 + (s - 1) % LBS_C_TILES, LBS_C_THREADS);
            }
#pragma unroll
            for (int j = 0; j < LBS_C_VPW; j++)
            {
                const float4 a = na, q1 = nb, q2 = nt_; // q1 = {n.x, t.x, n.y, t.y}, q2 = {n.z, t.z, w1, w2}
                const uint4  d = nd;
                if (j + 1 < LBS_C_VPW)
                {
                    na  = *reinterpret_cast<const float4*>(&vr[j + 1].px);
                    nb  = *reinterpret_cast<const float4*>(&vr[j + 1].nx);
                    nt_ = *reinterpret_cast<const float4*>(&vr[j + 1].nz);
                    nd  = *reinterpret_cast<const uint4*>(&vr[j + 1].w3);
                }
                if (d.w)
                {
                    if (d.w & 1) lbs_i_load(m[0], prow, d.y & 0xffffu);
                    if (d.w & 2) lbs_i_load(m[1], prow, d.y >> 16);
                    if (d.w & 4) lbs_i_load(m[2], prow, d.z & 0xffffu);
                    if (d.w & 8) lbs_i_load(m[3], prow, d.z >> 16);
                }
                const float w0 = a.w, w1 = q2.z, w2 = q2.w, w3 = __uint_as_float(d.x);
                float4 B[3];
                {
                    const unsigned long long ww0 = pack2(w0, w0), ww1 = pack2(w1, w1), ww2 = pack2(w2, w2), ww3 = pack2(w3, w3);
#pragma unroll
                    for (int r = 0; r < 3; r++)
                    {
                        unsigned long long lo = mul2(pack2(m[0][r].x, m[0][r].y), ww0), hi = mul2(pack2(m[0][r].z, m[0][r].w), ww0);
                        lo = fma2(pack2(m[1][r].x, m[1][r].y), ww1, lo), hi = fma2(pack2(m[1][r].z, m[1][r].w), ww1, hi);
                        lo = fma2(pack2(m[2][r].x, m[2][r].y), ww2, lo), hi = fma2(pack2(m[2][r].z, m[2][r].w), ww2, hi);
                        lo = fma2(pack2(m[3][r].x, m[3][r].y), ww3, lo), hi = fma2(pack2(m[3][r].z, m[3][r].w), ww3, hi);
                        unpack2(lo, B[r].x, B[r].y), unpack2(hi, B[r].z, B[r].w);
                    }
                }
                // mesh.vert:104-110.  Position: one FMA chain per row.  Normal and tangent go through the same 3x3 block, so
                // row r of both is ONE packed chain on the (n.k, t.k) pairs the record delivers: 9 FFMA2 instead of 18 FFMA, and
                // both squared lengths are one more packed chain (fewer instructions = less energy: the kernel runs at the
                // board's power cap, where time follows joules rather than latencies).
                opf[j * 3 + 0] = fmaf(B[0].x, a.x, fmaf(B[0].y, a.y, fmaf(B[0].z, a.z, B[0].w)));
                opf[j * 3 + 1] = fmaf(B[1].x, a.x, fmaf(B[1].y, a.y, fmaf(B[1].z, a.z, B[1].w)));
                opf[j * 3 + 2] = fmaf(B[2].x, a.x, fmaf(B[2].y, a.y, fmaf(B[2].z, a.z, B[2].w)));
                if (NT)
                {
                    const unsigned long long vx = pack2(q1.x, q1.y), vy = pack2(q1.z, q1.w), vz = pack2(q2.x, q2.y);
                    unsigned long long       o[3];
#pragma unroll
                    for (int r = 0; r < 3; r++)
                    {
                        o[r] = mul2(vx, pack2(B[r].x, B[r].x));
                        o[r] = fma2(vy, pack2(B[r].y, B[r].y), o[r]);
                        o[r] = fma2(vz, pack2(B[r].z, B[r].z), o[r]);
                    }
                    unsigned long long l2 = mul2(o[0], o[0]);
                    l2                    = fma2(o[1], o[1], l2);
                    l2                    = fma2(o[2], o[2], l2);
                    float ln, lt;
                    unpack2(l2, ln, lt);
                    const float rn = rsqrt_ftz(ln), rt = rsqrt_ftz(lt);
#pragma unroll
                    for (int r = 0; r < 3; r++)
                    {
                        float nr, tr;
                        unpack2(o[r], nr, tr);
                        onf[j * 6 + r] = nr * rn, onf[j * 6 + 3 + r] = tr * rt;
                    }
                }
            }
            if (s >= LBS_C_TILES)
                named_sync(1 + LBS_C_TILES + b, LBS_C_THREADS); // the tensor stores of superchunk s-TILES have read tile b
            unsigned char* tb = tiles + b * TILE;
#pragma unroll
            for (int q = 0; q < 3; q++)
                *reinterpret_cast<float4*>(tb + off_p[q]) = op[q];
            if (NT)
            {
#pragma unroll
                for (int q = 0; q < 6; q++)
                    *reinterpret_cast<float4*>(tb + off_n[q]) = on[q];
            }
            if (!LATE_HANDOVER)
            {
                fence_async_smem();
                named_arrive(1 + b, LBS_C_THREADS);
            }
        }
        if (T == 32)
        {
            fence_async_smem();
            named_arrive(1 + (n_steps - 1) % LBS_C_TILES, LBS_C_THREADS);
        }
    }
} // namespace pcl
