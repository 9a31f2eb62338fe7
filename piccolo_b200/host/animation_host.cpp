// Runtime, AnimationManager, Skeleton, AnimationComponent and AnimationBatch over the C ABI.
// Citations are relative to /root/reference/engine/source/runtime.
#include "piccolo_anim.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <tuple>
#include <algorithm>

namespace Piccolo
{
    // ---------------------------------------------------------------------------------------------
    // value types
    // ---------------------------------------------------------------------------------------------
    Matrix4x4::Matrix4x4()
    {
        for (int r = 0; r < 4; r++)
            for (int c = 0; c < 4; c++)
                m_mat[r][c] = (r == c) ? 1.f : 0.f;
    }
    Matrix4x4::Matrix4x4(const Matrix4x4_& m) { std::memcpy(m_mat, &m.v0, sizeof(m_mat)); } // core/math/matrix4.h:79-97 (row-major)
    Matrix4x4_ Matrix4x4::toMatrix4x4_() const
    {
        Matrix4x4_ r;
        std::memcpy(&r.v0, m_mat, sizeof(m_mat));
        return r;
    }

    // ---------------------------------------------------------------------------------------------
    // Runtime
    // ---------------------------------------------------------------------------------------------
    namespace
    {
        pcl_ctx*    g_ctx       = nullptr;
        bool        g_ctx_tried = false;
        std::string g_ctx_error;
        bool        g_blend_extension = false;
    } // namespace
    namespace detail
    {
        void destroyPoseGroups();
    }
    void Runtime::enableBlendExtension(bool on) { g_blend_extension = on; }
    bool Runtime::blendExtensionEnabled() { return g_blend_extension; }

    pcl_ctx* Runtime::context()
    {
        if (!g_ctx_tried)
        {
            g_ctx_tried     = true;
            const char* dev = std::getenv("PICCOLO_B200_DEVICE");
            if (pcl_create(dev ? std::atoi(dev) : 0, &g_ctx) != PCL_OK)
            {
                g_ctx_error = pcl_last_error(nullptr);
                g_ctx       = nullptr;
                std::fprintf(stderr, "[piccolo_b200] %s\n", g_ctx_error.c_str());
            }
        }
        return g_ctx;
    }
    const char* Runtime::lastError() { return g_ctx ? pcl_last_error(g_ctx) : g_ctx_error.c_str(); }
    void        Runtime::shutdown()
    {
        detail::destroyPoseGroups();
        AnimationManager::clearCaches();
        if (g_ctx)
            pcl_destroy(g_ctx);
        g_ctx = nullptr, g_ctx_tried = false;
    }

    pcl_skel uploadSkeleton(const SkeletonData& d)
    {
        pcl_ctx* c = Runtime::context();
        if (!c || d.bones_map.empty())
            return -1;
        const size_t         J = d.bones_map.size();
        std::vector<int32_t> index(J), parent(J);
        std::vector<float>   pos(3 * J), rot(4 * J), scl(3 * J), inv(16 * J);
        for (size_t i = 0; i < J; i++)
        {
            const RawBone& b = d.bones_map[i];
            index[i] = b.index, parent[i] = b.parent_index;
            pos[3 * i] = b.binding_pose.m_position.x, pos[3 * i + 1] = b.binding_pose.m_position.y, pos[3 * i + 2] = b.binding_pose.m_position.z;
            scl[3 * i] = b.binding_pose.m_scale.x, scl[3 * i + 1] = b.binding_pose.m_scale.y, scl[3 * i + 2] = b.binding_pose.m_scale.z;
            rot[4 * i] = b.binding_pose.m_rotation.w, rot[4 * i + 1] = b.binding_pose.m_rotation.x;
            rot[4 * i + 2] = b.binding_pose.m_rotation.y, rot[4 * i + 3] = b.binding_pose.m_rotation.z;
            std::memcpy(&inv[16 * i], &b.tpose_matrix.v0, 64);
        }
        pcl_skel s = -1;
        if (pcl_upload_skeleton(c, (int32_t)J, index.data(), parent.data(), pos.data(), rot.data(), scl.data(), inv.data(), d.is_flat,
                                d.in_topological_order, &s) != PCL_OK)
            return -1; // the reference builds nothing and says nothing (skeleton.cpp:24-28)
        return s;
    }

    pcl_mesh uploadMesh(const MeshData& mesh, int32_t max_slot)
    {
        pcl_ctx* c = Runtime::context();
        const size_t V = mesh.vertex_buffer.size();
        if (!c || V == 0 || mesh.bind.size() < V)
            return -1;
        std::vector<float>   pos(3 * V), nrm(3 * V), tan(3 * V), w(4 * V);
        std::vector<int32_t> idx(4 * V);
        for (size_t v = 0; v < V; v++)
        {
            const Vertex& x = mesh.vertex_buffer[v];
            pos[3 * v] = x.px, pos[3 * v + 1] = x.py, pos[3 * v + 2] = x.pz;
            nrm[3 * v] = x.nx, nrm[3 * v + 1] = x.ny, nrm[3 * v + 2] = x.nz;
            tan[3 * v] = x.tx, tan[3 * v + 1] = x.ty, tan[3 * v + 2] = x.tz;
            const SkeletonBinding& b = mesh.bind[v];
            idx[4 * v] = b.index0, idx[4 * v + 1] = b.index1, idx[4 * v + 2] = b.index2, idx[4 * v + 3] = b.index3;
            w[4 * v] = b.weight0, w[4 * v + 1] = b.weight1, w[4 * v + 2] = b.weight2, w[4 * v + 3] = b.weight3;
        }
        pcl_mesh m = -1;
        if (pcl_upload_mesh(c, (int32_t)V, pos.data(), nrm.data(), tan.data(), idx.data(), w.data(), max_slot, &m) != PCL_OK)
            return -1;
        return m;
    }

    // ---------------------------------------------------------------------------------------------
    // asset-side conventions
    // ---------------------------------------------------------------------------------------------
    int find_index_by_name(const SkeletonData& skeleton, const std::string& name)
    {
        for (const RawBone& b : skeleton.bones_map)
            if (b.name == name)
                return b.index;
        return std::numeric_limits<int>::max();
    }
    AnimSkelMap buildAnimSkelMap(const AnimNodeMap& anim, const SkeletonData& skeleton)
    {
        AnimSkelMap m;
        for (const std::string& n : anim.convert)
            m.convert.push_back(find_index_by_name(skeleton, n));
        return m;
    }
    bool addBoneBind(MeshData& mesh_bind, size_t bone_index, size_t vertex_index, float weight)
    {
        if (vertex_index >= mesh_bind.bind.size())
            return false;
        SkeletonBinding& b   = mesh_bind.bind[vertex_index];
        int*             idx[4] = {&b.index0, &b.index1, &b.index2, &b.index3};
        float*           w[4]   = {&b.weight0, &b.weight1, &b.weight2, &b.weight3};
        for (int k = 0; k < 4; k++)
            if (*idx[k] == 0) // 0 = empty slot, which is why stored indices are bone_index + 1
            {
                *idx[k] = (int)bone_index + 1;
                *w[k]   = weight;
                return true;
            }
        return false;
    }

    // ---------------------------------------------------------------------------------------------
    // AnimationManager (function/animation/animation_system.cpp)
    // ---------------------------------------------------------------------------------------------
    std::map<std::string, std::shared_ptr<SkeletonData>>  AnimationManager::m_skeleton_definition_cache;
    std::map<std::string, std::shared_ptr<AnimationClip>> AnimationManager::m_animation_data_cache;
    std::map<std::string, std::shared_ptr<AnimSkelMap>>   AnimationManager::m_animation_skeleton_map_cache;
    std::map<std::string, std::shared_ptr<BoneBlendMask>> AnimationManager::m_skeleton_mask_cache;
    namespace
    {
        std::map<std::tuple<pcl_skel, std::string, std::string>, pcl_clip> g_dev_clips;
        std::map<std::pair<pcl_skel, std::string>, pcl_mask>               g_dev_masks;
        std::map<std::pair<pcl_skel, uint64_t>, pcl_clip>                  g_dev_clips_by_content;
        std::map<uint64_t, pcl_skel>                                       g_dev_skels_by_content;
        uint64_t fnv1a(uint64_t h, const void* p, size_t n)
        {
            const unsigned char* b = (const unsigned char*)p;
            for (size_t i = 0; i < n; i++)
                h = (h ^ b[i]) * 1099511628211ull;
            return h;
        }
    } // namespace

    // Every AnimationComponent of a level names the same few skeleton files: one device copy per distinct CONTENT (what
    // pcl_upload_skeleton reads), so that components share step records, clips and masks - and the kernels' constant bank
    // is not rewritten when two characters take turns.
    pcl_skel AnimationManager::deviceSkeleton(const SkeletonData& d)
    {
        uint64_t h = 1469598103934665603ull;
        const int32_t flags[2] = {d.is_flat, d.in_topological_order};
        h = fnv1a(h, flags, sizeof(flags));
        for (const RawBone& b : d.bones_map)
        {
            const int32_t ip[2] = {b.index, b.parent_index};
            h = fnv1a(h, ip, sizeof(ip));
            h = fnv1a(h, &b.binding_pose.m_position, sizeof(Vector3));
            h = fnv1a(h, &b.binding_pose.m_rotation, sizeof(Quaternion));
            h = fnv1a(h, &b.binding_pose.m_scale, sizeof(Vector3));
            h = fnv1a(h, &b.tpose_matrix.v0, 64);
        }
        auto it = g_dev_skels_by_content.find(h);
        if (it != g_dev_skels_by_content.end())
            return it->second;
        pcl_skel s = uploadSkeleton(d);
        if (s >= 0)
            g_dev_skels_by_content.emplace(h, s);
        return s;
    }

    template<typename T, typename F>
    static std::shared_ptr<T> tryLoad(std::map<std::string, std::shared_ptr<T>>& cache, const std::string& path, F load)
    {
        auto found = cache.find(path);
        if (found != cache.end())
            return found->second;
        std::shared_ptr<T> res = load(path);
        cache.emplace(path, res);
        return res;
    }
    std::shared_ptr<SkeletonData> AnimationManager::tryLoadSkeleton(std::string p)
    {
        return tryLoad(m_skeleton_definition_cache, p, [](const std::string& u) { return AnimationLoader().loadSkeletonData(u); });
    }
    std::shared_ptr<AnimationClip> AnimationManager::tryLoadAnimation(std::string p)
    {
        return tryLoad(m_animation_data_cache, p, [](const std::string& u) { return AnimationLoader().loadAnimationClipData(u); });
    }
    std::shared_ptr<AnimSkelMap> AnimationManager::tryLoadAnimationSkeletonMap(std::string p)
    {
        return tryLoad(m_animation_skeleton_map_cache, p, [](const std::string& u) { return AnimationLoader().loadAnimSkelMap(u); });
    }
    std::shared_ptr<BoneBlendMask> AnimationManager::tryLoadSkeletonMask(std::string p)
    {
        return tryLoad(m_skeleton_mask_cache, p, [](const std::string& u) { return AnimationLoader().loadSkeletonMask(u); });
    }

    // animation_system.cpp:83-151.  Same outputs (clips and maps copied by value, per-bone normalised weights);
    // the reference's crash paths (blend_mask_file_path shorter than clip_count, mask naming a skeleton that was
    // never loaded) return what has been resolved so far instead.
    BlendStateWithClipData AnimationManager::getBlendStateWithClipData(const BlendState& bs)
    {
        BlendStateWithClipData out;
        out.clip_count  = bs.clip_count;
        out.blend_ratio = bs.blend_ratio;
        for (const auto& p : bs.blend_clip_file_path)
            out.blend_clip.push_back(*tryLoadAnimation(p));
        for (const auto& p : bs.blend_anim_skel_map_path)
            out.blend_anim_skel_map.push_back(*tryLoadAnimationSkeletonMap(p));
        std::vector<std::shared_ptr<BoneBlendMask>> masks;
        for (const auto& p : bs.blend_mask_file_path)
            masks.push_back(tryLoadSkeletonMask(p));
        out.source_clip_path = bs.blend_clip_file_path, out.source_map_path = bs.blend_anim_skel_map_path;
        out.source_mask_path = bs.blend_mask_file_path, out.source_weight = bs.blend_weight;
        if (masks.empty() || bs.clip_count < 0 || (size_t)bs.clip_count > masks.size() || (size_t)bs.clip_count > bs.blend_weight.size())
            return out;
        auto skel = m_skeleton_definition_cache.find(masks[0]->skeleton_file_path);
        if (skel == m_skeleton_definition_cache.end() || !skel->second)
            return out;
        const size_t bones = skel->second->bones_map.size();
        out.blend_weight.resize(bs.clip_count);
        for (auto& w : out.blend_weight)
            w.blend_weight.resize(bones);
        for (size_t b = 0; b < bones; b++)
        {
            float sum = 0;
            for (int c = 0; c < bs.clip_count; c++)
                if (b < masks[c]->enabled.size() && masks[c]->enabled[b])
                    sum += bs.blend_weight[c];
            for (int c = 0; c < bs.clip_count; c++)
                out.blend_weight[c].blend_weight[b] = (b < masks[c]->enabled.size() && masks[c]->enabled[b]) ? bs.blend_weight[c] / sum : 0.f;
        }
        return out;
    }

    static pcl_clip uploadClip(pcl_skel skel, const AnimationClip& clip, const AnimSkelMap& map)
    {
        pcl_ctx* c = Runtime::context();
        if (!c)
            return -1;
        const int            C = (int)clip.node_channels.size();
        std::vector<int32_t> counts(3 * (size_t)std::max(C, 1), 0);
        std::vector<float>   pos, rot, scl;
        for (int n = 0; n < C; n++)
        {
            const AnimationChannel& ch = clip.node_channels[n];
            counts[3 * n] = (int)ch.position_keys.size(), counts[3 * n + 1] = (int)ch.rotation_keys.size(), counts[3 * n + 2] = (int)ch.scaling_keys.size();
            for (const Vector3& k : ch.position_keys)
                pos.insert(pos.end(), {k.x, k.y, k.z});
            for (const Quaternion& k : ch.rotation_keys)
                rot.insert(rot.end(), {k.w, k.x, k.y, k.z});
            for (const Vector3& k : ch.scaling_keys)
                scl.insert(scl.end(), {k.x, k.y, k.z});
        }
        pos.push_back(0), rot.push_back(0), scl.push_back(0);
        std::vector<int32_t> conv(map.convert.begin(), map.convert.end());
        conv.push_back(0);
        pcl_clip out = -1;
        if (pcl_upload_clip(c, skel, clip.total_frame, clip.node_count, C, counts.data(), pos.data(), rot.data(), scl.data(),
                            (int32_t)map.convert.size(), conv.data(), &out) != PCL_OK)
            return -1;
        return out;
    }

    pcl_clip AnimationManager::deviceClip(pcl_skel skel, const std::string& clip_path, const std::string& map_path)
    {
        auto key = std::make_tuple(skel, clip_path, map_path);
        auto it  = g_dev_clips.find(key);
        if (it != g_dev_clips.end())
            return it->second;
        pcl_clip h = uploadClip(skel, *tryLoadAnimation(clip_path), *tryLoadAnimationSkeletonMap(map_path));
        g_dev_clips.emplace(key, h);
        return h;
    }

    pcl_clip AnimationManager::deviceClipFromData(pcl_skel skel, const AnimationClip& clip, const AnimSkelMap& map)
    {
        uint64_t h   = 1469598103934665603ull; // FNV-1a over everything the sampler reads
        auto     mix = [&h](const void* p, size_t n) {
            const unsigned char* b = (const unsigned char*)p;
            for (size_t i = 0; i < n; i++)
                h = (h ^ b[i]) * 1099511628211ull;
        };
        mix(&clip.total_frame, 4), mix(&clip.node_count, 4);
        for (const auto& ch : clip.node_channels)
        {
            size_t n[3] = {ch.position_keys.size(), ch.rotation_keys.size(), ch.scaling_keys.size()};
            mix(n, sizeof(n));
            if (n[0]) mix(ch.position_keys.data(), n[0] * sizeof(Vector3));
            if (n[1]) mix(ch.rotation_keys.data(), n[1] * sizeof(Quaternion));
            if (n[2]) mix(ch.scaling_keys.data(), n[2] * sizeof(Vector3));
        }
        if (!map.convert.empty())
            mix(map.convert.data(), map.convert.size() * sizeof(int));
        auto key = std::make_pair(skel, h);
        auto it  = g_dev_clips_by_content.find(key);
        if (it != g_dev_clips_by_content.end())
            return it->second;
        pcl_clip id = uploadClip(skel, clip, map);
        g_dev_clips_by_content.emplace(key, id);
        return id;
    }

    pcl_mask AnimationManager::deviceMask(pcl_skel skel, const std::string& mask_path)
    {
        auto key = std::make_pair(skel, mask_path);
        auto it  = g_dev_masks.find(key);
        if (it != g_dev_masks.end())
            return it->second;
        pcl_ctx* c = Runtime::context();
        pcl_mask h = -1;
        auto     m = tryLoadSkeletonMask(mask_path);
        int32_t  J = 0;
        if (c && pcl_skeleton_bone_count(c, skel, &J) == PCL_OK)
        {
            std::vector<int32_t> en(m->enabled.begin(), m->enabled.end());
            if ((int32_t)en.size() < J || pcl_upload_mask(c, skel, (int32_t)en.size(), en.data(), &h) != PCL_OK)
                h = -1;
        }
        g_dev_masks.emplace(key, h);
        return h;
    }

    // Drops the host caches AND the device copies behind them (a level unload): clips first, then every skeleton nothing
    // refers to any more (a skeleton some live Skeleton / AnimationBatch still poses with stays until that object is gone).
    void AnimationManager::clearCaches()
    {
        m_skeleton_definition_cache.clear(), m_animation_data_cache.clear(), m_animation_skeleton_map_cache.clear(), m_skeleton_mask_cache.clear();
        if (pcl_ctx* c = Runtime::context())
        {
            std::vector<pcl_clip> clips;
            for (auto& kv : g_dev_clips)
                clips.push_back(kv.second);
            for (auto& kv : g_dev_clips_by_content)
                clips.push_back(kv.second);
            std::sort(clips.begin(), clips.end());
            clips.erase(std::unique(clips.begin(), clips.end()), clips.end());
            for (pcl_clip h : clips)
                if (h >= 0)
                    pcl_clip_release(c, h);
            for (auto& kv : g_dev_skels_by_content)
                pcl_skeleton_release(c, kv.second); // refused while in use: then it stays
        }
        g_dev_clips.clear(), g_dev_masks.clear(), g_dev_clips_by_content.clear(), g_dev_skels_by_content.clear();
    }

    // ---------------------------------------------------------------------------------------------
    // PoseGroup: the hidden batch behind all AnimationComponents of one (skeleton, clip count)
    //
    // The reference ticks every GObject's AnimationComponent in turn, each tick a complete pose on the CPU
    // (function/framework/object/object.cpp:41-50).  One launch + one synchronisation per character would cost more than the
    // reference's 18 us per tick, so components only RECORD their frame inputs here; the first result anybody asks for
    // (MeshComponent::tick reads getResult(), mesh_component.cpp:75) poses every recorded member with ONE pcl_batch_frame
    // call.  Results are identical to the batch-of-one path: an instance's pose does not depend on batch composition.
    // ---------------------------------------------------------------------------------------------
    namespace detail
    {
        struct PoseGroup
        {
            pcl_skel   skel = -1;
            int        K = 1, J = 0;
            pcl_batch* batch = nullptr;
            int64_t    capacity = 0;
            bool       dirty = false, derived = false;
            std::vector<Skeleton*> members; // slot -> owner, nullptr = free
            std::vector<int32_t>   clip, mask;
            std::vector<float>     phase, weight;
            std::vector<float>     rows; // [capacity][J][12] of the last flush

            int acquire(Skeleton* who)
            {
                for (size_t i = 0; i < members.size(); i++)
                    if (!members[i])
                    {
                        members[i] = who;
                        return (int)i;
                    }
                members.push_back(who);
                const size_t n = members.size() * K;
                if (clip.size() < n)
                    clip.resize(n, -1), mask.resize(n, -1), phase.resize(n, 0.f), weight.resize(n, 1.f);
                return (int)members.size() - 1;
            }
            void release(int slot)
            {
                if (slot >= 0 && slot < (int)members.size())
                    members[slot] = nullptr;
            }
            void record(int slot, const int32_t* c, const float* ph, const float* w, const int32_t* m)
            {
                for (int k = 0; k < K; k++)
                    clip[(size_t)slot * K + k] = c[k], phase[(size_t)slot * K + k] = ph[k], weight[(size_t)slot * K + k] = w[k], mask[(size_t)slot * K + k] = m[k];
                dirty = true;
            }
            bool flush()
            {
                if (!dirty && batch)
                    return true;
                pcl_ctx* c = Runtime::context();
                if (!c || members.empty())
                    return false;
                if ((int64_t)members.size() > capacity) // grow: batches have a fixed size
                {
                    if (batch)
                        pcl_batch_destroy(batch), batch = nullptr;
                    // as many instances as there are components (a frame poses all of them): 1, 2, 4, 8, then multiples of 8
                    int64_t want = 1;
                    while (want < (int64_t)members.size() && want < 8)
                        want *= 2;
                    if (want < (int64_t)members.size())
                        want = ((int64_t)members.size() + 7) / 8 * 8;
                    capacity = want;
                    if (pcl_batch_create(c, skel, capacity, K, &batch) != PCL_OK)
                    {
                        batch = nullptr, capacity = 0;
                        return false;
                    }
                    if (derived)
                        pcl_batch_keep_derived(batch, 1);
                }
                // slots nobody has recorded into yet (free, or not ticked so far) pose some valid clip at phase 0
                int32_t any_clip = -1;
                for (int32_t v : clip)
                    if (v >= 0)
                    {
                        any_clip = v;
                        break;
                    }
                if (any_clip < 0)
                    return false;
                const size_t n = (size_t)capacity * K;
                if (clip.size() < n)
                    clip.resize(n, -1), mask.resize(n, -1), phase.resize(n, 0.f), weight.resize(n, 1.f);
                for (size_t i = 0; i < n; i++)
                    if (clip[i] < 0)
                        clip[i] = any_clip, mask[i] = -1, phase[i] = 0.f, weight[i] = 1.f;
                rows.resize((size_t)capacity * J * 12);
                if (pcl_batch_frame(batch, clip.data(), phase.data(), weight.data(), mask.data(), rows.data()) != PCL_OK)
                    return false;
                dirty = false;
                for (Skeleton* m : members)
                    if (m)
                        m->m_derived_stale = true;
                return true;
            }
        };
        namespace
        {
            std::map<std::pair<pcl_skel, int>, PoseGroup*> g_groups;
        }
        PoseGroup* poseGroup(pcl_skel skel, int K, int J)
        {
            auto key = std::make_pair(skel, K);
            auto it  = g_groups.find(key);
            if (it != g_groups.end())
                return it->second;
            PoseGroup* g = new PoseGroup();
            g->skel = skel, g->K = K, g->J = J;
            g_groups.emplace(key, g);
            return g;
        }
        void destroyPoseGroups()
        {
            for (auto& kv : g_groups)
            {
                for (Skeleton* m : kv.second->members)
                    if (m)
                        m->m_group = nullptr, m->m_slot = -1;
                if (kv.second->batch)
                    pcl_batch_destroy(kv.second->batch);
                delete kv.second;
            }
            g_groups.clear();
        }
    } // namespace detail

    // ---------------------------------------------------------------------------------------------
    // Skeleton (function/animation/skeleton.cpp)
    // ---------------------------------------------------------------------------------------------
    Skeleton::~Skeleton()
    {
        leaveGroup();
        if (m_batch)
            pcl_batch_destroy(m_batch);
        delete[] m_bones;
    }

    void Skeleton::leaveGroup()
    {
        if (m_group)
            m_group->release(m_slot);
        m_group = nullptr, m_slot = -1;
    }

    void Skeleton::buildSkeleton(const SkeletonData& def)
    {
        m_is_flat = def.is_flat;
        leaveGroup();
        if (m_batch)
            pcl_batch_destroy(m_batch), m_batch = nullptr, m_batch_k = 0;
        delete[] m_bones;
        m_bones = nullptr, m_bone_count = 0, m_skel = -1;
        if (!m_is_flat || !def.in_topological_order)
            return; // skeleton.cpp:24-28
        pcl_skel s = AnimationManager::deviceSkeleton(def);
        if (s < 0)
            return;
        m_skel       = s;
        m_bone_count = (int)def.bones_map.size();
        m_bones      = new Bone[m_bone_count];
        std::vector<int32_t> parent(m_bone_count);
        pcl_skeleton_parents(Runtime::context(), m_skel, parent.data());
        m_index.resize(m_bone_count);
        pcl_skeleton_indices(Runtime::context(), m_skel, m_index.data());
        for (int i = 0; i < m_bone_count; i++)
        {
            const RawBone& d = def.bones_map[i];
            Bone&          b = m_bones[i];
            b.m_name = d.name, b.m_has_definition = true, b.m_index = d.index;
            b.m_parent = parent[i] >= 0 ? &m_bones[parent[i]] : nullptr;
            // Bone::initialize, node.cpp:220-234 (setOrientation normalises; NaN -> identity)
            Quaternion q = d.binding_pose.m_rotation;
            if (std::isnan(q.w) || std::isnan(q.x) || std::isnan(q.y) || std::isnan(q.z))
                q = Quaternion();
            else
            {
                float f = 1.0f / std::sqrt(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
                q.w *= f, q.x *= f, q.y *= f, q.z *= f;
            }
            b.m_initial_orientation = q, b.m_initial_position = d.binding_pose.m_position, b.m_initial_scale = d.binding_pose.m_scale;
            b.m_inverse_Tpose = Matrix4x4(d.tpose_matrix);
        }
        m_derived_stale = true;
    }

    void Skeleton::ensureBatch(int k)
    {
        if (m_batch && m_batch_k == k)
            return;
        if (m_batch)
            pcl_batch_destroy(m_batch), m_batch = nullptr;
        if (pcl_batch_create(Runtime::context(), m_skel, 1, k, &m_batch) != PCL_OK)
        {
            m_batch = nullptr, m_batch_k = 0;
            return;
        }
        pcl_batch_keep_derived(m_batch, 1); // LevelDebugger reads the bones' derived TRS through getBones()
        m_batch_k = k;
    }

    // skeleton.cpp:39-117: clip 0 with weight forced to 1, whatever clip_count says (:46, :63-64).  With
    // Runtime::enableBlendExtension(true), clip_count in 2..4 blends with the masks and raw weights the blend state names.
    void Skeleton::applyAnimation(const BlendStateWithClipData& bs)
    {
        if (!m_bones || m_skel < 0)
            return;
        int K = Runtime::blendExtensionEnabled() ? bs.clip_count : 1;
        if (K < 1 || K > 4 || (size_t)K > bs.blend_clip.size() || (size_t)K > bs.blend_anim_skel_map.size() || (size_t)K > bs.blend_ratio.size())
            K = (bs.blend_clip.empty() || bs.blend_anim_skel_map.empty() || bs.blend_ratio.empty()) ? 0 : 1;
        if (K == 0)
            return;
        const bool has_src = bs.source_clip_path.size() >= (size_t)K && bs.source_map_path.size() >= (size_t)K;
        int32_t    clip[4], mask[4];
        float      phase[4], weight[4];
        for (int k = 0; k < K; k++)
        {
            clip[k] = has_src ? AnimationManager::deviceClip(m_skel, bs.source_clip_path[k], bs.source_map_path[k]) :
                                AnimationManager::deviceClipFromData(m_skel, bs.blend_clip[k], bs.blend_anim_skel_map[k]);
            if (clip[k] < 0)
                return;
            phase[k]  = bs.blend_ratio[k];
            weight[k] = (K > 1 && bs.source_weight.size() > (size_t)k) ? bs.source_weight[k] : 1.f;
            mask[k]   = (K > 1 && bs.source_mask_path.size() > (size_t)k) ? AnimationManager::deviceMask(m_skel, bs.source_mask_path[k]) : -1;
        }
        applyResolved(K, clip, phase, weight, mask);
    }

    void Skeleton::applyResolved(int K, const int32_t* clip, const float* phase, const float* weight, const int32_t* mask)
    {
        if (m_grouped)
        {
            if (m_group && m_group->K != K)
                leaveGroup();
            if (!m_group)
            {
                m_group = detail::poseGroup(m_skel, K, m_bone_count);
                m_slot  = m_group->acquire(this);
            }
            m_group->record(m_slot, clip, phase, weight, mask);
            m_derived_stale = true;
            return;
        }
        ensureBatch(K);
        if (!m_batch)
            return;
        pcl_batch_frame(m_batch, clip, phase, weight, mask, nullptr); // asynchronous: outputAnimationResult waits for the rows
        m_derived_stale = true;
    }

    // Matrix4x4::toMatrix4x4_ (core/math/matrix4.h:98-119) of an affine product: rows 0..2 come from the device, row 3 is 0 0 0 1
    void Skeleton::fillResult(const float* rows12, AnimationResult& r) const
    {
        r.node.resize(m_bone_count);
        for (int i = 0; i < m_bone_count; i++)
        {
            r.node[i].index = m_index[i];
            float* m = &r.node[i].transform.v0;
            std::memcpy(m, rows12 + (size_t)i * 12, 48);
            m[12] = 0.f, m[13] = 0.f, m[14] = 0.f, m[15] = 1.f;
        }
    }

    AnimationResult Skeleton::outputAnimationResult()
    {
        AnimationResult r;
        if (!m_bones)
            return r;
        if (m_grouped)
        {
            if (m_group && m_group->flush())
                fillResult(m_group->rows.data() + (size_t)m_slot * m_bone_count * 12, r);
            return r;
        }
        if (!m_batch)
            return r;
        std::vector<float> rows((size_t)m_bone_count * 12);
        if (pcl_batch_frame_result(m_batch, rows.data()) == PCL_OK)
            fillResult(rows.data(), r);
        return r;
    }

    void Skeleton::refreshDerived()
    {
        if (!m_derived_stale || !m_bones)
            return;
        std::vector<float> d((size_t)m_bone_count * 10);
        if (m_grouped)
        {
            if (!m_group)
                return;
            if (!m_group->derived) // first LevelDebugger-style reader of this group: keep derived transforms from now on
            {
                m_group->derived = true;
                if (m_group->batch)
                    pcl_batch_keep_derived(m_group->batch, 1);
                m_group->dirty = true;
            }
            if (!m_group->flush() || pcl_batch_download_derived(m_group->batch, m_slot, 1, d.data()) != PCL_OK)
                return;
        }
        else if (!m_batch || pcl_batch_download_derived(m_batch, 0, 1, d.data()) != PCL_OK)
            return;
        for (int i = 0; i < m_bone_count; i++)
        {
            const float* o = &d[(size_t)i * 10];
            m_bones[i].m_derived_position    = Vector3 {o[0], o[1], o[2]};
            m_bones[i].m_derived_orientation = Quaternion {o[3], o[4], o[5], o[6]};
            m_bones[i].m_derived_scale       = Vector3 {o[7], o[8], o[9]};
        }
        m_derived_stale = false;
    }

    // resetSkeleton only touches local TRS in the reference (node.cpp:197-206); derived values change at the next update()
    void Skeleton::resetSkeleton() {}

    const Bone* Skeleton::getBones() const
    {
        const_cast<Skeleton*>(this)->refreshDerived();
        return m_bones;
    }
    int32_t Skeleton::getBonesCount() const { return m_bone_count; }

    // ---------------------------------------------------------------------------------------------
    // AnimationComponent (function/framework/component/animation/animation_component.cpp)
    // ---------------------------------------------------------------------------------------------
    void AnimationComponent::postLoadResource(std::weak_ptr<GObject> parent_object)
    {
        m_parent_object   = parent_object;
        auto skeleton_res = AnimationManager::tryLoadSkeleton(m_animation_res.skeleton_file_path);
        m_skeleton.buildSkeleton(*skeleton_res);
        m_skeleton.m_grouped = true; // components pose together (detail::PoseGroup)
        m_resolved_paths.clear();
    }

    // animation_component.cpp:17-25.  The phase advances exactly as there; what the reference then does in
    // getBlendStateWithClipData (four map look-ups AND a deep copy of every clip and map, animation_system.cpp:104,108) is
    // reduced to the look-ups, and those to a comparison of the path strings once the handles are known; the pose itself is
    // recorded and computed when the result is first read.
    void AnimationComponent::tick(float delta_time)
    {
        BlendState& bs = m_animation_res.blend_state;
        if (bs.blend_ratio.empty() || bs.blend_clip_file_length.empty())
            return;
        bs.blend_ratio[0] += (delta_time / bs.blend_clip_file_length[0]);
        bs.blend_ratio[0] -= std::floor(bs.blend_ratio[0]);
        if (!m_skeleton.m_bones || m_skeleton.m_skel < 0)
            return;
        int K = Runtime::blendExtensionEnabled() ? bs.clip_count : 1;
        if (K < 1 || K > 4 || (size_t)K > bs.blend_clip_file_path.size() || (size_t)K > bs.blend_anim_skel_map_path.size() || (size_t)K > bs.blend_ratio.size())
            K = (bs.blend_clip_file_path.empty() || bs.blend_anim_skel_map_path.empty()) ? 0 : 1;
        if (K == 0)
            return;
        // resolve (and cache) the device handles of the paths this blend state names; a tick whose paths have not changed
        // only compares strings
        static const std::string no_mask;
        bool same = m_resolved_paths.size() == 3 * (size_t)K;
        for (int k = 0; k < K && same; k++)
        {
            const std::string& mp = (K > 1 && bs.blend_mask_file_path.size() > (size_t)k) ? bs.blend_mask_file_path[k] : no_mask;
            same = m_resolved_paths[3 * k] == bs.blend_clip_file_path[k] && m_resolved_paths[3 * k + 1] == bs.blend_anim_skel_map_path[k] &&
                   m_resolved_paths[3 * k + 2] == mp;
        }
        if (!same)
        {
            m_resolved_paths.clear();
            for (int k = 0; k < K; k++)
            {
                const std::string& mp = (K > 1 && bs.blend_mask_file_path.size() > (size_t)k) ? bs.blend_mask_file_path[k] : no_mask;
                m_clip[k] = AnimationManager::deviceClip(m_skeleton.m_skel, bs.blend_clip_file_path[k], bs.blend_anim_skel_map_path[k]);
                m_mask[k] = mp.empty() ? -1 : AnimationManager::deviceMask(m_skeleton.m_skel, mp);
                if (m_clip[k] < 0)
                {
                    m_resolved_paths.clear();
                    return;
                }
                m_resolved_paths.push_back(bs.blend_clip_file_path[k]);
                m_resolved_paths.push_back(bs.blend_anim_skel_map_path[k]);
                m_resolved_paths.push_back(mp);
            }
        }
        float phase[4], weight[4];
        for (int k = 0; k < K; k++)
        {
            phase[k]  = bs.blend_ratio[k];
            weight[k] = (K > 1 && bs.blend_weight.size() > (size_t)k) ? bs.blend_weight[k] : 1.f;
        }
        m_skeleton.applyResolved(K, m_clip, phase, weight, m_mask);
        m_result_stale = true;
    }

    const AnimationResult& AnimationComponent::getResult() const
    {
        if (m_result_stale)
        {
            // (the grouped path fills the component's own vector in place: no allocation per character and frame)
            if (m_skeleton.m_grouped && m_skeleton.m_group && m_skeleton.m_group->flush())
                m_skeleton.fillResult(m_skeleton.m_group->rows.data() + (size_t)m_skeleton.m_slot * m_skeleton.m_bone_count * 12,
                                      m_animation_res.animation_result);
            else
                m_animation_res.animation_result = m_skeleton.outputAnimationResult();
            m_result_stale = false;
        }
        return m_animation_res.animation_result;
    }
    const Skeleton& AnimationComponent::getSkeleton() const { return m_skeleton; }

    // ---------------------------------------------------------------------------------------------
    // AnimationBatch
    // ---------------------------------------------------------------------------------------------
    AnimationBatch::~AnimationBatch()
    {
        if (m_skinbuf)
            pcl_skinbuf_destroy(m_skinbuf);
        if (m_batch)
            pcl_batch_destroy(m_batch);
    }

    bool AnimationBatch::create(const AnimationComponentRes& res, int64_t n)
    {
        pcl_ctx* c = Runtime::context();
        if (!c || n < 1)
            return false;
        auto def = AnimationManager::tryLoadSkeleton(res.skeleton_file_path);
        if (!def->is_flat || !def->in_topological_order)
            return false;
        m_skel = AnimationManager::deviceSkeleton(*def);
        if (m_skel < 0)
            return false;
        const BlendState& bs = res.blend_state;
        const int         K  = bs.clip_count;
        if (K < 1 || K > 4 || bs.blend_clip_file_path.size() < (size_t)K || bs.blend_anim_skel_map_path.size() < (size_t)K ||
            bs.blend_ratio.size() < (size_t)K || bs.blend_clip_file_length.size() < (size_t)K)
            return false;
        if (pcl_batch_create(c, m_skel, n, K, &m_batch) != PCL_OK)
            return false;
        m_n = n, m_k = K, m_j = (int32_t)def->bones_map.size();
        m_index.resize(m_j);
        pcl_skeleton_indices(c, m_skel, m_index.data());
        std::vector<int32_t> clip((size_t)n * K), mask((size_t)n * K);
        std::vector<float>   phase((size_t)n * K), weight((size_t)n * K), len((size_t)n * K);
        for (int k = 0; k < K; k++)
        {
            const pcl_clip ck = AnimationManager::deviceClip(m_skel, bs.blend_clip_file_path[k], bs.blend_anim_skel_map_path[k]);
            const pcl_mask mk = (K > 1 && bs.blend_mask_file_path.size() > (size_t)k) ? AnimationManager::deviceMask(m_skel, bs.blend_mask_file_path[k]) : -1;
            if (ck < 0)
                return false;
            for (int64_t i = 0; i < n; i++)
            {
                clip[i * K + k] = ck, mask[i * K + k] = mk, phase[i * K + k] = bs.blend_ratio[k];
                weight[i * K + k] = bs.blend_weight.size() > (size_t)k ? bs.blend_weight[k] : 1.f;
                len[i * K + k]    = bs.blend_clip_file_length[k];
            }
        }
        return pcl_batch_set(m_batch, 0, n, clip.data(), phase.data(), weight.data(), len.data(), mask.data()) == PCL_OK;
    }

    void AnimationBatch::setPhases(const std::vector<float>& r)
    {
        if (m_batch && r.size() == (size_t)m_n * m_k)
            pcl_batch_set(m_batch, 0, m_n, nullptr, r.data(), nullptr, nullptr, nullptr);
    }
    void AnimationBatch::setWeights(const std::vector<float>& w)
    {
        if (m_batch && w.size() == (size_t)m_n * m_k)
            pcl_batch_set(m_batch, 0, m_n, nullptr, nullptr, w.data(), nullptr, nullptr);
    }
    void AnimationBatch::tick(float dt)
    {
        if (m_batch)
            pcl_batch_tick(m_batch, dt, PCL_ADVANCE_FIRST_ONLY); // only blend_ratio[0] moves, animation_component.cpp:19-21
    }
    std::vector<float> AnimationBatch::getPhases() const
    {
        std::vector<float> p((size_t)m_n * m_k);
        if (m_batch)
            pcl_batch_get_phases(m_batch, 0, m_n, p.data());
        return p;
    }
    void AnimationBatch::getResults(int64_t first, int64_t count, std::vector<float>& mats) const
    {
        mats.assign((size_t)count * m_j * 16, 0.f);
        if (!m_batch)
            return;
        // rows 0..2 cross the host link (48 B per bone), row 3 = 0 0 0 1 is filled in here (matrix4.h:98-119)
        std::vector<float> rows((size_t)count * m_j * 12);
        if (pcl_batch_download_palette3x4(m_batch, first, count, rows.data()) != PCL_OK)
            return;
        for (size_t e = 0; e < (size_t)count * m_j; e++)
        {
            std::memcpy(&mats[e * 16], &rows[e * 12], 48);
            mats[e * 16 + 15] = 1.f;
        }
    }
    AnimationResult AnimationBatch::getResult(int64_t i) const
    {
        AnimationResult    r;
        std::vector<float> mats;
        if (!m_batch || i < 0 || i >= m_n)
            return r;
        getResults(i, 1, mats);
        r.node.resize(m_j);
        for (int b = 0; b < m_j; b++)
        {
            r.node[b].index = m_index[b];
            std::memcpy(&r.node[b].transform.v0, &mats[(size_t)b * 16], 64);
        }
        return r;
    }
    bool AnimationBatch::attachMesh(const MeshData& mesh, int64_t cap, bool nt)
    {
        if (!m_batch)
            return false;
        m_mesh = uploadMesh(mesh, m_j);
        if (m_mesh < 0)
            return false;
        m_v = (int32_t)mesh.vertex_buffer.size(), m_nt = nt;
        if (m_skinbuf)
            pcl_skinbuf_destroy(m_skinbuf), m_skinbuf = nullptr;
        return pcl_skinbuf_create(Runtime::context(), m_mesh, cap, nt ? PCL_SKIN_POS_NRM_TAN : PCL_SKIN_POS, &m_skinbuf) == PCL_OK;
    }
    bool AnimationBatch::skin(int64_t first, int64_t count)
    {
        return m_batch && m_skinbuf && pcl_batch_skin(m_batch, m_mesh, m_skinbuf, first, count, 0) == PCL_OK;
    }
    void AnimationBatch::downloadSkinned(int64_t slot_first, int64_t count, std::vector<float>& pos, std::vector<float>& nt) const
    {
        pos.assign((size_t)count * m_v * 3, 0.f);
        nt.assign(m_nt ? (size_t)count * m_v * 6 : 0, 0.f);
        if (m_skinbuf)
            pcl_skinbuf_download(m_skinbuf, slot_first, count, pos.data(), m_nt ? nt.data() : nullptr);
    }
} // namespace Piccolo
