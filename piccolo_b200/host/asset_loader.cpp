// JSON -> asset structs with the semantics of the reference's generated Serializer::read
// (engine/template/allSerializer.ipp.mustache:28-33): a field is read only if its key is present
// (otherwise the C++ default stays), arrays are resize()d then read element by element, numbers are
// doubles cast to float / int (runtime/core/meta/serializer/serializer.cpp:21-52).  JSON key = member
// name minus a leading "m_" (engine/source/meta_parser/parser/meta/meta_utils.cpp:144-152).
// A missing or malformed file leaves a default-constructed asset, like AssetManager::loadAsset
// (runtime/resource/asset_manager/asset_manager.h:20-46), and the message goes to stderr.
#include "json.h"
#include "piccolo_anim.h"

#include <cstdio>

using pcljson::Value;

namespace Piccolo
{
    namespace
    {
        std::string g_root;

        std::string fullPath(const std::string& url) { return g_root.empty() ? url : g_root + "/" + url; }

        bool loadJson(const std::string& url, Value& out)
        {
            const std::string path = fullPath(url);
            std::string       err;
            if (!pcljson::parse_file(path, out, err))
            {
                std::fprintf(stderr, "[piccolo_b200] loadAsset %s: %s\n", url.c_str(), err.c_str());
                return false;
            }
            return true;
        }
        inline bool has(const Value& j, const char* k) { return !j[k].is_null(); }
        inline void rd(const Value& j, const char* k, float& v)
        {
            if (has(j, k))
                v = static_cast<float>(j[k].num);
        }
        inline void rd(const Value& j, const char* k, int& v)
        {
            if (has(j, k))
                v = static_cast<int>(j[k].num);
        }
        inline void rd(const Value& j, const char* k, bool& v)
        {
            if (has(j, k))
                v = j[k].b;
        }
        inline void rd(const Value& j, const char* k, std::string& v)
        {
            if (has(j, k))
                v = j[k].str;
        }
        void rd(const Value& j, Vector3& v) { rd(j, "x", v.x), rd(j, "y", v.y), rd(j, "z", v.z); }
        void rd(const Value& j, Quaternion& q) { rd(j, "w", q.w), rd(j, "x", q.x), rd(j, "y", q.y), rd(j, "z", q.z); }
        void rd(const Value& j, Matrix4x4_& m)
        {
            float* p = &m.v0;
            for (int i = 0; i < 16; i++)
                rd(j, ("v" + std::to_string(i)).c_str(), p[i]);
        }
        void rd(const Value& j, Transform& t)
        {
            if (has(j, "position"))
                rd(j["position"], t.m_position);
            if (has(j, "scale"))
                rd(j["scale"], t.m_scale);
            if (has(j, "rotation"))
                rd(j["rotation"], t.m_rotation);
        }
        void rd(const Value& j, RawBone& b);
        void rd(const Value& j, AnimationChannel& c);
        void rd(const Value& j, Vertex& v);
        void rd(const Value& j, SkeletonBinding& b);
        template<typename T>
        void rdArray(const Value& j, const char* k, std::vector<T>& out)
        {
            if (!has(j, k))
                return;
            const Value& a = j[k];
            out.resize(a.size());
            for (size_t i = 0; i < a.size(); i++)
                rd(a[i], out[i]);
        }
        void rdInts(const Value& j, const char* k, std::vector<int>& out)
        {
            if (!has(j, k))
                return;
            const Value& a = j[k];
            out.resize(a.size());
            for (size_t i = 0; i < a.size(); i++)
                out[i] = static_cast<int>(a[i].num);
        }
        void rdFloats(const Value& j, const char* k, std::vector<float>& out)
        {
            if (!has(j, k))
                return;
            const Value& a = j[k];
            out.resize(a.size());
            for (size_t i = 0; i < a.size(); i++)
                out[i] = static_cast<float>(a[i].num);
        }
        void rdStrings(const Value& j, const char* k, std::vector<std::string>& out)
        {
            if (!has(j, k))
                return;
            const Value& a = j[k];
            out.resize(a.size());
            for (size_t i = 0; i < a.size(); i++)
                out[i] = a[i].str;
        }
        void rd(const Value& j, RawBone& b)
        {
            rd(j, "name", b.name), rd(j, "index", b.index), rd(j, "parent_index", b.parent_index);
            if (has(j, "binding_pose"))
                rd(j["binding_pose"], b.binding_pose);
            if (has(j, "tpose_matrix"))
                rd(j["tpose_matrix"], b.tpose_matrix);
        }
        void rd(const Value& j, AnimationChannel& c)
        {
            rd(j, "name", c.name);
            rdArray(j, "position_keys", c.position_keys);
            rdArray(j, "rotation_keys", c.rotation_keys);
            rdArray(j, "scaling_keys", c.scaling_keys);
        }
        void rd(const Value& j, Vertex& v)
        {
            rd(j, "px", v.px), rd(j, "py", v.py), rd(j, "pz", v.pz), rd(j, "nx", v.nx), rd(j, "ny", v.ny), rd(j, "nz", v.nz);
            rd(j, "tx", v.tx), rd(j, "ty", v.ty), rd(j, "tz", v.tz), rd(j, "u", v.u), rd(j, "v", v.v);
        }
        void rd(const Value& j, SkeletonBinding& b)
        {
            rd(j, "index0", b.index0), rd(j, "index1", b.index1), rd(j, "index2", b.index2), rd(j, "index3", b.index3);
            rd(j, "weight0", b.weight0), rd(j, "weight1", b.weight1), rd(j, "weight2", b.weight2), rd(j, "weight3", b.weight3);
        }
        void rd(const Value& j, BlendState& s)
        {
            rd(j, "clip_count", s.clip_count);
            rdStrings(j, "blend_clip_file_path", s.blend_clip_file_path);
            rdFloats(j, "blend_clip_file_length", s.blend_clip_file_length);
            rdStrings(j, "blend_anim_skel_map_path", s.blend_anim_skel_map_path);
            rdFloats(j, "blend_weight", s.blend_weight);
            rdStrings(j, "blend_mask_file_path", s.blend_mask_file_path);
            rdFloats(j, "blend_ratio", s.blend_ratio);
        }
    } // namespace

    void               AnimationLoader::setAssetRoot(const std::string& root) { g_root = root; }
    const std::string& AnimationLoader::assetRoot() { return g_root; }

    std::shared_ptr<SkeletonData> AnimationLoader::loadSkeletonData(std::string url)
    {
        SkeletonData data;
        Value        j;
        if (detail::cacheLoad(fullPath(url), data))
            return std::make_shared<SkeletonData>(std::move(data));
        if (loadJson(url, j))
        {
            rdArray(j, "bones_map", data.bones_map);
            rd(j, "is_flat", data.is_flat), rd(j, "root_index", data.root_index), rd(j, "in_topological_order", data.in_topological_order);
            detail::cacheStore(fullPath(url), data);
        }
        return std::make_shared<SkeletonData>(data);
    }

    // reads an AnimationAsset and keeps only clip_data (function/animation/animation_loader.cpp:74-79)
    std::shared_ptr<AnimationClip> AnimationLoader::loadAnimationClipData(std::string url)
    {
        AnimationAsset asset;
        Value          j;
        if (detail::cacheLoad(fullPath(url), asset.clip_data))
            return std::make_shared<AnimationClip>(std::move(asset.clip_data));
        if (loadJson(url, j))
        {
            if (has(j, "node_map"))
                rdStrings(j["node_map"], "convert", asset.node_map.convert);
            if (has(j, "clip_data"))
            {
                const Value& c = j["clip_data"];
                rd(c, "total_frame", asset.clip_data.total_frame), rd(c, "node_count", asset.clip_data.node_count);
                rdArray(c, "node_channels", asset.clip_data.node_channels);
            }
            rd(j, "skeleton_file_path", asset.skeleton_file_path);
            detail::cacheStore(fullPath(url), asset.clip_data);
        }
        return std::make_shared<AnimationClip>(asset.clip_data);
    }

    std::shared_ptr<AnimSkelMap> AnimationLoader::loadAnimSkelMap(std::string url)
    {
        AnimSkelMap data;
        Value       j;
        if (detail::cacheLoad(fullPath(url), data))
            return std::make_shared<AnimSkelMap>(std::move(data));
        if (loadJson(url, j))
        {
            rdInts(j, "convert", data.convert);
            detail::cacheStore(fullPath(url), data);
        }
        return std::make_shared<AnimSkelMap>(data);
    }

    std::shared_ptr<BoneBlendMask> AnimationLoader::loadSkeletonMask(std::string url)
    {
        BoneBlendMask data;
        Value         j;
        if (detail::cacheLoad(fullPath(url), data))
            return std::make_shared<BoneBlendMask>(std::move(data));
        if (loadJson(url, j))
        {
            rd(j, "skeleton_file_path", data.skeleton_file_path);
            rdInts(j, "enabled", data.enabled);
            detail::cacheStore(fullPath(url), data);
        }
        return std::make_shared<BoneBlendMask>(data);
    }

    std::shared_ptr<MeshData> AnimationLoader::loadMeshData(std::string url)
    {
        MeshData data;
        Value    j;
        if (detail::cacheLoad(fullPath(url), data))
            return std::make_shared<MeshData>(std::move(data));
        if (loadJson(url, j))
        {
            rdArray(j, "vertex_buffer", data.vertex_buffer);
            rdInts(j, "index_buffer", data.index_buffer);
            rdArray(j, "bind", data.bind);
            detail::cacheStore(fullPath(url), data);
        }
        return std::make_shared<MeshData>(data);
    }

    // Finds the AnimationComponent entry of an object file ("components": [{"$typeName": "AnimationComponent",
    // "$context": {"animation_res": {...}}}], engine/asset/objects/character/player/player.object.json:24-56);
    // a bare {"animation_res": {...}} or the AnimationComponentRes object itself are accepted too.
    std::shared_ptr<AnimationComponentRes> AnimationLoader::loadAnimationComponentRes(std::string url)
    {
        AnimationComponentRes res;
        Value                 j;
        if (loadJson(url, j))
        {
            const Value* r = nullptr;
            if (has(j, "components"))
                for (size_t i = 0; i < j["components"].size(); i++)
                {
                    const Value& c = j["components"][i];
                    if (c["$typeName"].str == "AnimationComponent")
                        r = &c["$context"]["animation_res"];
                }
            if (!r)
                r = has(j, "animation_res") ? &j["animation_res"] : &j;
            rd(*r, "skeleton_file_path", res.skeleton_file_path);
            rd(*r, "frame_position", res.frame_position);
            if (has(*r, "blend_state"))
                rd((*r)["blend_state"], res.blend_state);
        }
        return std::make_shared<AnimationComponentRes>(res);
    }
} // namespace Piccolo
