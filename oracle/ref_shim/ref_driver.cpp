// TEST INFRASTRUCTURE ONLY -- part of the oracle, never linked into the product library.
//
// C driver around the UNMODIFIED reference animation code (compiled in place from
// $PICCOLO_REF by oracle/Makefile).  It reproduces what AnimationComponent does
// (engine/source/runtime/function/framework/component/animation/animation_component.cpp:8-25)
// without the GObject/Component framework:
//     postLoadResource : AnimationManager::tryLoadSkeleton + Skeleton::buildSkeleton
//     tick             : ratio += dt/len; ratio -= floor(ratio);
//                        Skeleton::applyAnimation(AnimationManager::getBlendStateWithClipData(bs));
//                        Skeleton::outputAnimationResult()
// and restates the skinning block of engine/shader/glsl/mesh.vert:72-111 with the reference's
// own Matrix4x4 / Vector4 operators (the GLSL itself cannot run: no graphics stack).
#include "runtime/function/animation/animation_system.h"
#include "runtime/function/animation/skeleton.h"

#include "runtime/core/math/matrix4.h"
#include "runtime/core/math/vector4.h"

#include <atomic>
#include <chrono>
#include <cmath>
#include <cstring>
#include <memory>
#include <thread>
#include <vector>

using namespace Piccolo;

namespace
{
    struct Session
    {
        BlendState                    bs;
        std::shared_ptr<SkeletonData> skel_data;
        Skeleton                      skeleton; // never copied (raw owning pointer, no copy ctor)
    };
    std::vector<std::unique_ptr<Session>> g_sessions;

    void dumpResult(const AnimationResult& r, float* out_mats, int* out_index)
    {
        for (size_t b = 0; b < r.node.size(); b++)
        {
            if (out_index)
                out_index[b] = r.node[b].index;
            if (out_mats)
                std::memcpy(out_mats + 16 * b, &r.node[b].transform.v0, 16 * sizeof(float));
        }
    }

    void dumpDerived(const Skeleton& s, float* out) // per bone: P.xyz, Q.wxyz, S.xyz
    {
        const Bone* bones = s.getBones();
        for (int b = 0; b < s.getBonesCount(); b++)
        {
            const Vector3&    p = bones[b]._getDerivedPosition();
            const Quaternion& q = bones[b]._getDerivedOrientation();
            const Vector3&    c = bones[b]._getDerivedScale();
            float*            o = out + 10 * b;
            o[0] = p.x, o[1] = p.y, o[2] = p.z;
            o[3] = q.w, o[4] = q.x, o[5] = q.y, o[6] = q.z;
            o[7] = c.x, o[8] = c.y, o[9] = c.z;
        }
    }
} // namespace

extern "C"
{
    // Mirrors player.object.json: clip_count=1, one clip, one map, TWO mask paths, weight [1].
    int ref_open(const char* skeleton_url, const char* clip_url, const char* map_url, const char* mask_url,
                 float clip_len)
    {
        auto s = std::make_unique<Session>();
        s->bs.clip_count               = 1;
        s->bs.blend_clip_file_path     = {clip_url};
        s->bs.blend_clip_file_length   = {clip_len};
        s->bs.blend_anim_skel_map_path = {map_url};
        s->bs.blend_weight             = {1.f};
        s->bs.blend_mask_file_path     = {mask_url, mask_url};
        s->bs.blend_ratio              = {0.f};
        // must use the very string the mask file stores in skeleton_file_path, else
        // animation_system.cpp:116 dereferences a null cache entry
        s->skel_data = AnimationManager::tryLoadSkeleton(skeleton_url);
        s->skeleton.buildSkeleton(*s->skel_data);
        g_sessions.push_back(std::move(s));
        return (int)g_sessions.size() - 1;
    }

    int ref_bone_count(int h) { return g_sessions[h]->skeleton.getBonesCount(); }

    // One applyAnimation + outputAnimationResult per phase.
    int ref_pose(int h, int n, const float* phases, float* out_mats, int* out_index, float* out_derived)
    {
        Session& s = *g_sessions[h];
        int      J = s.skeleton.getBonesCount();
        for (int i = 0; i < n; i++)
        {
            s.bs.blend_ratio[0] = phases[i];
            s.skeleton.applyAnimation(AnimationManager::getBlendStateWithClipData(s.bs));
            AnimationResult r = s.skeleton.outputAnimationResult();
            if ((int)r.node.size() != J)
                return -1;
            dumpResult(r, out_mats ? out_mats + (size_t)i * J * 16 : nullptr, i == 0 ? out_index : nullptr);
            if (out_derived)
                dumpDerived(s.skeleton, out_derived + (size_t)i * J * 10);
        }
        return 0;
    }

    // AnimationComponent::tick n times from *phase (animation_component.cpp:17-25); palette of every tick.
    int ref_tick(int h, int n_ticks, float dt, float* phase_io, float* out_mats, float* out_phases)
    {
        Session& s          = *g_sessions[h];
        int      J          = s.skeleton.getBonesCount();
        s.bs.blend_ratio[0] = *phase_io;
        for (int i = 0; i < n_ticks; i++)
        {
            s.bs.blend_ratio[0] += (dt / s.bs.blend_clip_file_length[0]);
            s.bs.blend_ratio[0] -= floor(s.bs.blend_ratio[0]);
            s.skeleton.applyAnimation(AnimationManager::getBlendStateWithClipData(s.bs));
            AnimationResult r = s.skeleton.outputAnimationResult();
            if (out_mats)
                dumpResult(r, out_mats + (size_t)i * J * 16, nullptr);
            if (out_phases)
                out_phases[i] = s.bs.blend_ratio[0];
        }
        *phase_io = s.bs.blend_ratio[0];
        return 0;
    }

    // CPU baseline: n_threads host threads, each with its own Skeleton over the warmed read-only
    // AnimationManager caches; every thread animates `per_thread` instances (distinct phases).
    // full_tick=1 adds getBlendStateWithClipData (what AnimationComponent::tick really does);
    // full_tick=0 reuses one pre-resolved BlendStateWithClipData (pose only).
    // Returns instances per second over all threads (wall clock of the slowest thread).
    double ref_bench_pose(int h, int n_threads, int per_thread, int full_tick, double* checksum)
    {
        Session& s0 = *g_sessions[h];
        (void)AnimationManager::getBlendStateWithClipData(s0.bs); // warm every cache
        std::vector<std::thread> th;
        std::vector<double>      secs(n_threads, 0.0), sums(n_threads, 0.0);
        std::atomic<int>         ready {0};
        std::atomic<bool>        go {false};
        for (int t = 0; t < n_threads; t++)
        {
            th.emplace_back([&, t] {
                Skeleton sk;
                sk.buildSkeleton(*s0.skel_data);
                BlendState             bs  = s0.bs;
                BlendStateWithClipData pre = AnimationManager::getBlendStateWithClipData(bs);
                ready++;
                while (!go.load())
                    std::this_thread::yield();
                auto   t0  = std::chrono::steady_clock::now();
                double acc = 0;
                for (int i = 0; i < per_thread; i++)
                {
                    float ph = (float)((t * per_thread + i) * 0.6180339887);
                    ph -= std::floor(ph);
                    if (full_tick)
                    {
                        bs.blend_ratio[0] = ph;
                        sk.applyAnimation(AnimationManager::getBlendStateWithClipData(bs));
                    }
                    else
                    {
                        pre.blend_ratio[0] = ph;
                        sk.applyAnimation(pre);
                    }
                    AnimationResult r = sk.outputAnimationResult();
                    acc += r.node.back().transform.v3;
                }
                secs[t] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
                sums[t] = acc;
            });
        }
        while (ready.load() < n_threads)
            std::this_thread::yield();
        go = true;
        for (auto& x : th)
            x.join();
        double worst = 0, sum = 0;
        for (int t = 0; t < n_threads; t++)
        {
            worst = std::max(worst, secs[t]);
            sum += sums[t];
        }
        if (checksum)
            *checksum = sum;
        return (double)n_threads * per_thread / worst;
    }

    // The per-bone normalised blend weights AnimationManager::getBlendStateWithClipData computes every tick
    // (animation_system.cpp:117-149) for a two-clip blend state with two different masks.  out: [2][J].
    int ref_blend_weights(const char* skeleton_url, const char* clip_url, const char* map_url, const char* mask0_url,
                          const char* mask1_url, float w0, float w1, float* out)
    {
        BlendState bs;
        bs.clip_count               = 2;
        bs.blend_clip_file_path     = {clip_url, clip_url};
        bs.blend_clip_file_length   = {1.f, 1.f};
        bs.blend_anim_skel_map_path = {map_url, map_url};
        bs.blend_weight             = {w0, w1};
        bs.blend_mask_file_path     = {mask0_url, mask1_url};
        bs.blend_ratio              = {0.f, 0.f};
        auto skel = AnimationManager::tryLoadSkeleton(skeleton_url); // the masks name it by this very url
        BlendStateWithClipData r = AnimationManager::getBlendStateWithClipData(bs);
        size_t J = skel->bones_map.size();
        if (r.blend_weight.size() != 2 || r.blend_weight[0].blend_weight.size() != J)
            return -1;
        for (int c = 0; c < 2; c++)
            std::memcpy(out + c * J, r.blend_weight[c].blend_weight.data(), J * sizeof(float));
        return (int)J;
    }

    // ---- skinning: mesh.vert:72-111 with the reference's CPU matrix operators -----------------
    // palette16: n_slots row-major 4x4 (slot 0 = identity, slot b+1 = bone b; mesh_component.cpp:72-79).
    // w4_raw is renormalised exactly like render_resource.cpp:714-725 before use.
    static void lbsRange(int v0, int v1, const float* pos, const float* nrm, const float* tan, const int* idx4,
                         const float* w4_raw, const float* palette16, float* out_pos, float* out_nrm, float* out_tan)
    {
        for (int v = v0; v < v1; v++)
        {
            const float* wr  = w4_raw + 4 * v;
            float        inv = wr[0] + wr[1] + wr[2] + wr[3];
            inv              = (inv != 0.0) ? 1 / inv : 1.0;
            float     w[4]   = {wr[0] * inv, wr[1] * inv, wr[2] * inv, wr[3] * inv};
            Matrix4x4 B(0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0);
            for (int k = 0; k < 4; k++)
            {
                int id = idx4[4 * v + k];
                if (w[k] > 0.0f && id > 0)
                {
                    float m[16];
                    std::memcpy(m, palette16 + 16 * (size_t)id, sizeof(m));
                    B = B + Matrix4x4(m) * w[k];
                }
            }
            Vector4 p = B * Vector4(pos[3 * v], pos[3 * v + 1], pos[3 * v + 2], 1.0f);
            out_pos[3 * v] = p.x, out_pos[3 * v + 1] = p.y, out_pos[3 * v + 2] = p.z;
            if (out_nrm)
            {
                Vector4 n = B * Vector4(nrm[3 * v], nrm[3 * v + 1], nrm[3 * v + 2], 0.0f);
                float   l = std::sqrt(n.x * n.x + n.y * n.y + n.z * n.z); // GLSL normalize(): v / length(v)
                out_nrm[3 * v] = n.x / l, out_nrm[3 * v + 1] = n.y / l, out_nrm[3 * v + 2] = n.z / l;
            }
            if (out_tan)
            {
                Vector4 t = B * Vector4(tan[3 * v], tan[3 * v + 1], tan[3 * v + 2], 0.0f);
                float   l = std::sqrt(t.x * t.x + t.y * t.y + t.z * t.z);
                out_tan[3 * v] = t.x / l, out_tan[3 * v + 1] = t.y / l, out_tan[3 * v + 2] = t.z / l;
            }
        }
    }

    int ref_lbs(int V, const float* pos, const float* nrm, const float* tan, const int* idx4, const float* w4_raw,
                const float* palette16, float* out_pos, float* out_nrm, float* out_tan)
    {
        lbsRange(0, V, pos, nrm, tan, idx4, w4_raw, palette16, out_pos, out_nrm, out_tan);
        return 0;
    }

    // n_inst instances (palettes [n_inst][n_slots][16]) skinned by n_threads threads; returns instances/s.
    double ref_bench_lbs(int V, const float* pos, const float* nrm, const float* tan, const int* idx4,
                         const float* w4_raw, int n_slots, const float* palettes, int n_inst, int n_threads,
                         double* checksum)
    {
        std::vector<std::thread> th;
        std::vector<double>      secs(n_threads, 0.0), sums(n_threads, 0.0);
        std::atomic<int>         next {0};
        auto                     t_begin = std::chrono::steady_clock::now();
        for (int t = 0; t < n_threads; t++)
        {
            th.emplace_back([&, t] {
                std::vector<float> op(3 * (size_t)V), on(3 * (size_t)V), ot(3 * (size_t)V);
                double             acc = 0;
                for (;;)
                {
                    int i = next.fetch_add(1);
                    if (i >= n_inst)
                        break;
                    lbsRange(0, V, pos, nrm, tan, idx4, w4_raw, palettes + (size_t)i * n_slots * 16, op.data(),
                             on.data(), ot.data());
                    acc += op[3 * (size_t)(V - 1)] + on[0] + ot[1];
                }
                sums[t] = acc;
            });
        }
        for (auto& x : th)
            x.join();
        double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
        double sum  = 0;
        for (double x : sums)
            sum += x;
        if (checksum)
            *checksum = sum;
        return n_inst / wall;
    }
}
