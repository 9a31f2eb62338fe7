// Driver for the host C++ mirror, used by tests/test_host_*.py.  It talks to the classes the way the
// reference's callers do (GObject::load/tick, MeshComponent::tick, LevelDebugger::drawBones) and dumps
// what they would see as raw little-endian arrays: each record = int32 tag, int32 count, payload.
#include "piccolo_anim.h"

#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <chrono>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

using namespace Piccolo;

static FILE* g_out = nullptr;
static void  putf(int tag, const std::vector<float>& v)
{
    int32_t h[2] = {tag, (int32_t)v.size()};
    fwrite(h, 4, 2, g_out);
    fwrite(v.data(), 4, v.size(), g_out);
}
static void puti(int tag, const std::vector<int32_t>& v)
{
    int32_t h[2] = {tag | 0x40000000, (int32_t)v.size()};
    fwrite(h, 4, 2, g_out);
    fwrite(v.data(), 4, v.size(), g_out);
}
static void putResult(int tag, const AnimationResult& r)
{
    std::vector<float>   m;
    std::vector<int32_t> idx;
    for (const auto& n : r.node)
    {
        idx.push_back(n.index);
        const float* p = &n.transform.v0;
        m.insert(m.end(), p, p + 16);
    }
    putf(tag, m);
    puti(tag, idx);
}

static int cmd_load(int argc, char** argv)
{
    if (argc < 8)
        return 2;
    AnimationLoader::setAssetRoot(argv[2]);
    if (argc > 8) // optional: binary asset cache directory
        AnimationLoader::setCacheDir(argv[8]);
    AnimationLoader L;
    auto            sk = L.loadSkeletonData(argv[3]);
    auto            cl = L.loadAnimationClipData(argv[4]);
    auto            mp = L.loadAnimSkelMap(argv[5]);
    auto            mk = L.loadSkeletonMask(argv[6]);
    g_out              = fopen(argv[7], "wb");
    std::vector<int32_t> meta = {(int32_t)sk->bones_map.size(), sk->is_flat, sk->in_topological_order, sk->root_index, cl->total_frame,
                                 cl->node_count, (int32_t)cl->node_channels.size()};
    puti(1, meta);
    std::vector<int32_t> idx, par;
    std::vector<float>   pos, rot, scl, tp;
    for (const RawBone& b : sk->bones_map)
    {
        idx.push_back(b.index), par.push_back(b.parent_index);
        pos.insert(pos.end(), {b.binding_pose.m_position.x, b.binding_pose.m_position.y, b.binding_pose.m_position.z});
        scl.insert(scl.end(), {b.binding_pose.m_scale.x, b.binding_pose.m_scale.y, b.binding_pose.m_scale.z});
        rot.insert(rot.end(), {b.binding_pose.m_rotation.w, b.binding_pose.m_rotation.x, b.binding_pose.m_rotation.y, b.binding_pose.m_rotation.z});
        const float* m = &b.tpose_matrix.v0;
        tp.insert(tp.end(), m, m + 16);
    }
    puti(2, idx), puti(3, par), putf(4, pos), putf(5, rot), putf(6, scl), putf(7, tp);
    std::vector<int32_t> counts;
    std::vector<float>   kp, kr, ks;
    for (const auto& ch : cl->node_channels)
    {
        counts.insert(counts.end(), {(int32_t)ch.position_keys.size(), (int32_t)ch.rotation_keys.size(), (int32_t)ch.scaling_keys.size()});
        for (auto& k : ch.position_keys) kp.insert(kp.end(), {k.x, k.y, k.z});
        for (auto& k : ch.rotation_keys) kr.insert(kr.end(), {k.w, k.x, k.y, k.z});
        for (auto& k : ch.scaling_keys) ks.insert(ks.end(), {k.x, k.y, k.z});
    }
    puti(8, counts), putf(9, kp), putf(10, kr), putf(11, ks);
    puti(12, std::vector<int32_t>(mp->convert.begin(), mp->convert.end()));
    puti(13, std::vector<int32_t>(mk->enabled.begin(), mk->enabled.end()));
    // a missing file leaves a default-constructed asset and raises nothing (asset_manager.h:25-29)
    auto none = L.loadSkeletonData("does/not/exist.json");
    puti(14, {(int32_t)none->bones_map.size(), none->is_flat});
    fclose(g_out);
    std::printf("skeleton_path_in_mask=%s\n", mk->skeleton_file_path.c_str());
    return 0;
}

// offline-converter conventions: name -> index map and the first-free-of-four binding rule
static int cmd_bind(int argc, char** argv)
{
    if (argc < 6)
        return 2;
    AnimationLoader::setAssetRoot(argv[2]);
    auto            sk = AnimationLoader().loadSkeletonData(argv[3]);
    AnimNodeMap     nm;
    // channel names = the clip's node_channels names, which is what AnimationAsset.node_map.convert lists
    auto cl = AnimationLoader().loadAnimationClipData(argv[4]);
    for (auto& ch : cl->node_channels)
        nm.convert.push_back(ch.name);
    nm.convert.push_back("no such bone");
    AnimSkelMap m = buildAnimSkelMap(nm, *sk);
    g_out         = fopen(argv[5], "wb");
    puti(1, std::vector<int32_t>(m.convert.begin(), m.convert.end()));
    MeshData md;
    md.bind.resize(3);
    bool ok[6] = {addBoneBind(md, 4, 0, 0.5f), addBoneBind(md, 9, 0, 0.25f), addBoneBind(md, 0, 0, 0.125f), addBoneBind(md, 57, 0, 0.0625f),
                  addBoneBind(md, 3, 0, 1.f), addBoneBind(md, 3, 7, 1.f)};
    puti(2, {md.bind[0].index0, md.bind[0].index1, md.bind[0].index2, md.bind[0].index3, ok[0], ok[1], ok[2], ok[3], ok[4], ok[5], md.bind[1].index0});
    putf(3, {md.bind[0].weight0, md.bind[0].weight1, md.bind[0].weight2, md.bind[0].weight3});
    fclose(g_out);
    return 0;
}

// AnimationManager::getBlendStateWithClipData of a two-clip, two-mask blend state: the per-bone weights
static int cmd_weights(int argc, char** argv)
{
    if (argc < 11)
        return 2;
    AnimationLoader::setAssetRoot(argv[2]);
    BlendState bs;
    bs.clip_count               = 2;
    bs.blend_clip_file_path     = {argv[4], argv[4]};
    bs.blend_clip_file_length   = {1.f, 1.f};
    bs.blend_anim_skel_map_path = {argv[5], argv[5]};
    bs.blend_weight             = {(float)std::atof(argv[8]), (float)std::atof(argv[9])};
    bs.blend_mask_file_path     = {argv[6], argv[7]};
    bs.blend_ratio              = {0.f, 0.f};
    AnimationManager::tryLoadSkeleton(argv[3]);
    BlendStateWithClipData r = AnimationManager::getBlendStateWithClipData(bs);
    g_out                    = fopen(argv[10], "wb");
    for (auto& w : r.blend_weight)
        putf(1, w.blend_weight);
    puti(2, {(int32_t)r.blend_clip.size(), (int32_t)r.blend_anim_skel_map.size(), r.clip_count, (int32_t)r.blend_clip[0].node_channels.size()});
    fclose(g_out);
    return 0;
}

// AnimationComponent as GObject::load + n x GObject::tick would drive it
static int cmd_tick(int argc, char** argv)
{
    if (argc < 7)
        return 2;
    AnimationLoader::setAssetRoot(argv[2]);
    auto               res = AnimationLoader().loadAnimationComponentRes(argv[3]);
    const int          n   = std::atoi(argv[4]);
    const float        dt  = (float)std::atof(argv[5]);
    g_out                  = fopen(argv[6], "wb");
    { // the component (and its device batch) must be gone before the runtime is shut down
    AnimationComponent comp;
    comp.setAnimationRes(*res);
    comp.postLoadResource(std::weak_ptr<GObject>());
    std::vector<float> phases;
    for (int t = 0; t < n; t++)
    {
        comp.tick(dt);
        phases.push_back(comp.getAnimationRes().blend_state.blend_ratio[0]);
        if (t == 0 || t == 1 || t == 45 || t == 46 || t == n - 1)
            putResult(100 + t, comp.getResult());
    }
    putf(1, phases);
    // what LevelDebugger::drawBones reads (level_debugger.cpp:110-146)
    const Skeleton&      sk    = comp.getSkeleton();
    const Bone*          bones = sk.getBones();
    std::vector<float>   der;
    std::vector<int32_t> parent, ids;
    for (int i = 0; i < sk.getBonesCount(); i++)
    {
        const Vector3&    p = bones[i]._getDerivedPosition();
        const Quaternion& q = bones[i]._getDerivedOrientation();
        const Vector3&    s = bones[i]._getDerivedScale();
        der.insert(der.end(), {p.x, p.y, p.z, q.w, q.x, q.y, q.z, s.x, s.y, s.z});
        const Node* par = bones[i].getParent();
        parent.push_back(par ? (int32_t)(static_cast<const Bone*>(par) - bones) : -1);
        ids.push_back((int32_t)bones[i].getID());
    }
    putf(2, der), puti(3, parent), puti(4, ids);
    std::printf("bones=%d first=%s last=%s ctx=%s\n", sk.getBonesCount(), sk.getBonesCount() ? bones[0].getName().c_str() : "-",
                sk.getBonesCount() ? bones[sk.getBonesCount() - 1].getName().c_str() : "-", Runtime::context() ? "gpu" : "none");
    }
    fclose(g_out);
    std::fflush(stdout);
    Runtime::shutdown();
    return 0;
}

// crowd: N components in one AnimationBatch, then skinning
static int cmd_batch(int argc, char** argv)
{
    if (argc < 9)
        return 2;
    AnimationLoader::setAssetRoot(argv[2]);
    auto          res = AnimationLoader().loadAnimationComponentRes(argv[3]);
    const int64_t N   = std::atoll(argv[4]);
    const int     n   = std::atoi(argv[5]);
    const float   dt  = (float)std::atof(argv[6]);
    auto          mesh = AnimationLoader().loadMeshData(argv[7]);
    g_out              = fopen(argv[8], "wb");
    {
    AnimationBatch batch;
    if (!batch.create(*res, N))
    {
        std::printf("create failed: %s\n", Runtime::lastError());
        return 1;
    }
    const int          K = res->blend_state.clip_count;
    std::vector<float> ph((size_t)N * K);
    for (int64_t i = 0; i < N; i++)
        for (int k = 0; k < K; k++)
            ph[i * K + k] = (float)i / (float)N;
    batch.setPhases(ph);
    for (int t = 0; t < n; t++)
        batch.tick(dt);
    putf(1, batch.getPhases());
    std::vector<float> mats;
    batch.getResults(0, N, mats);
    putf(2, mats);
    putResult(3, batch.getResult(N - 1));
    if (!mesh->vertex_buffer.empty())
    {
        const int64_t cnt = N < 8 ? N : 8;
        if (!batch.attachMesh(*mesh, cnt, true) || !batch.skin(N - cnt, cnt))
        {
            std::printf("skin failed: %s\n", Runtime::lastError());
            return 1;
        }
        std::vector<float> pos, nt;
        batch.downloadSkinned(0, cnt, pos, nt);
        putf(4, pos), putf(5, nt);
    }
    fclose(g_out);
    std::printf("batch ok N=%lld J=%d\n", (long long)N, batch.boneCount());
    }
    std::fflush(stdout);
    Runtime::shutdown();
    return 0;
}

// n AnimationComponents of one level, driven like GObject::tick drives them: every frame all components tick, then every
// consumer (MeshComponent::tick, mesh_component.cpp:75) reads its result.  Dumps every component's phases and results after
// `ticks` frames, the derived transforms of the last one, and a stand-alone Skeleton's answer for component 0's inputs.
static int cmd_multi(int argc, char** argv)
{
    if (argc < 8)
        return 2;
    AnimationLoader::setAssetRoot(argv[2]);
    auto        res   = AnimationLoader().loadAnimationComponentRes(argv[3]);
    const int   n     = std::atoi(argv[4]);
    const int   ticks = std::atoi(argv[5]);
    const float dt    = (float)std::atof(argv[6]);
    g_out             = fopen(argv[7], "wb");
    {
        std::vector<std::unique_ptr<AnimationComponent>> comps;
        for (int i = 0; i < n; i++)
        {
            comps.emplace_back(new AnimationComponent());
            AnimationComponentRes r = *res;
            r.blend_state.blend_ratio[0] = (float)i / (float)n;
            comps.back()->setAnimationRes(r);
            comps.back()->postLoadResource(std::weak_ptr<GObject>());
        }
        for (int t = 0; t < ticks; t++)
        {
            for (auto& c : comps)
                c->tick(dt);
            for (auto& c : comps)
                (void)c->getResult();
        }
        std::vector<float> phases;
        for (int i = 0; i < n; i++)
        {
            phases.push_back(comps[i]->getAnimationRes().blend_state.blend_ratio[0]);
            putResult(100 + i, comps[i]->getResult());
        }
        putf(1, phases);
        const Skeleton&    sk    = comps[n - 1]->getSkeleton();
        const Bone*        bones = sk.getBones();
        std::vector<float> der;
        for (int i = 0; i < sk.getBonesCount(); i++)
        {
            const Vector3&    p = bones[i]._getDerivedPosition();
            const Quaternion& q = bones[i]._getDerivedOrientation();
            const Vector3&    s = bones[i]._getDerivedScale();
            der.insert(der.end(), {p.x, p.y, p.z, q.w, q.x, q.y, q.z, s.x, s.y, s.z});
        }
        putf(2, der);
        // the reference's own pair of calls on a Skeleton used directly
        Skeleton solo;
        solo.buildSkeleton(*AnimationManager::tryLoadSkeleton(res->skeleton_file_path));
        solo.applyAnimation(AnimationManager::getBlendStateWithClipData(comps[0]->getAnimationRes().blend_state));
        putResult(3, solo.outputAnimationResult());
        // a component that joins late and one that leaves: slots are recycled, the others' results do not move
        if (n > 1)
            comps.erase(comps.begin() + n / 2);
        comps.emplace_back(new AnimationComponent());
        comps.back()->setAnimationRes(*res);
        comps.back()->postLoadResource(std::weak_ptr<GObject>());
        for (auto& c : comps)
            c->tick(0.f);
        putResult(4, comps[0]->getResult());
        putResult(5, comps.back()->getResult());
        std::printf("multi ok n=%d\n", n);
    }
    fclose(g_out);
    std::fflush(stdout);
    Runtime::shutdown();
    return 0;
}

// per-frame cost of n components through the drop-in classes: n x tick, then n x getResult (wall clock)
static int cmd_latency(int argc, char** argv)
{
    if (argc < 6)
        return 2;
    AnimationLoader::setAssetRoot(argv[2]);
    auto      res    = AnimationLoader().loadAnimationComponentRes(argv[3]);
    const int n      = std::atoi(argv[4]);
    const int frames = std::atoi(argv[5]);
    {
        std::vector<std::unique_ptr<AnimationComponent>> comps;
        for (int i = 0; i < n; i++)
        {
            comps.emplace_back(new AnimationComponent());
            AnimationComponentRes r = *res;
            r.blend_state.blend_ratio[0] = (float)i / (float)n;
            comps.back()->setAnimationRes(r);
            comps.back()->postLoadResource(std::weak_ptr<GObject>());
        }
        double sink = 0;
        auto   frame = [&]() {
            for (auto& c : comps)
                c->tick(1.f / 60.f);
            for (auto& c : comps)
                sink += c->getResult().node[7].transform.v3;
        };
        for (int t = 0; t < 50; t++)
            frame();
        std::vector<double> us;
        for (int t = 0; t < frames; t++)
        {
            auto t0 = std::chrono::steady_clock::now();
            frame();
            us.push_back(std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count());
        }
        std::sort(us.begin(), us.end());
        double mean = 0;
        for (double x : us)
            mean += x;
        std::printf("{\"components\": %d, \"frames\": %d, \"us_per_frame_median\": %.2f, \"us_per_frame_mean\": %.2f, \"us_per_frame_p90\": %.2f, \"sink\": %.3f}\n",
                    n, frames, us[us.size() / 2], mean / us.size(), us[us.size() * 9 / 10], sink);
    }
    std::fflush(stdout);
    Runtime::shutdown();
    return 0;
}

int main(int argc, char** argv)
{
    if (argc < 2)
        return 2;
    if (!std::strcmp(argv[1], "load"))
        return cmd_load(argc, argv);
    if (!std::strcmp(argv[1], "bind"))
        return cmd_bind(argc, argv);
    if (!std::strcmp(argv[1], "weights"))
        return cmd_weights(argc, argv);
    if (!std::strcmp(argv[1], "tick"))
        return cmd_tick(argc, argv);
    if (!std::strcmp(argv[1], "batch"))
        return cmd_batch(argc, argv);
    if (!std::strcmp(argv[1], "multi"))
        return cmd_multi(argc, argv);
    if (!std::strcmp(argv[1], "latency"))
        return cmd_latency(argc, argv);
    return 2;
}
