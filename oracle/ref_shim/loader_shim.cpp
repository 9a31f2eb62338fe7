// TEST INFRASTRUCTURE ONLY -- part of the oracle, never linked into the product library.
//
// Stand-in for the one reference translation unit that cannot be compiled headless:
// engine/source/runtime/function/animation/animation_loader.cpp needs the libclang-generated
// "_generated/serializer/all_serializer.h", spdlog and g_runtime_global_context
// (animation_loader.cpp:4-11).  This file defines the four AnimationLoader::load* members
// (declared in the reference's own animation_loader.h:13-20) on top of the vendored json11,
// mimicking what the generated reader does (engine/template/allSerializer.ipp.mustache:28-33):
//   * a field is read only when its key is present (otherwise the C++ default stays),
//   * vectors are resize()d to the JSON array length, then each element is read in place,
//   * numbers are parsed as double and cast (runtime/core/meta/serializer/serializer.cpp:21-52).
// JSON key = member name minus a leading "m_" (meta_parser/parser/meta/meta_utils.cpp:144-152).
//
// Everything else in oracle/_ref is the UNMODIFIED reference code compiled in place.
#include "runtime/function/animation/animation_loader.h"

#include "json11.hpp"

#include <cstdio>
#include <fstream>
#include <sstream>
#include <string>

using json11::Json;

namespace
{
    std::string g_asset_root; // prefix joined in front of every url (AssetManager::getFullPath stand-in)

    bool loadJson(const std::string& url, Json& out)
    {
        std::string   path = g_asset_root.empty() ? url : (g_asset_root + "/" + url);
        std::ifstream f(path);
        if (!f)
        {
            std::fprintf(stderr, "[ref shim] open file: %s failed!\n", path.c_str());
            return false; // the reference logs and leaves a default-constructed asset (asset_manager.h:25-29)
        }
        std::stringstream ss;
        ss << f.rdbuf();
        std::string err;
        out = Json::parse(ss.str(), err);
        if (!err.empty())
        {
            std::fprintf(stderr, "[ref shim] parse json file %s failed: %s\n", path.c_str(), err.c_str());
            return false;
        }
        return true;
    }

    inline bool has(const Json& j, const char* k) { return !j[k].is_null(); }
    inline void rd(const Json& j, const char* k, float& v)
    {
        if (has(j, k))
            v = static_cast<float>(j[k].number_value());
    }
    inline void rd(const Json& j, const char* k, int& v)
    {
        if (has(j, k))
            v = static_cast<int>(j[k].number_value());
    }
    inline void rd(const Json& j, const char* k, bool& v)
    {
        if (has(j, k))
            v = j[k].bool_value();
    }
    inline void rd(const Json& j, const char* k, std::string& v)
    {
        if (has(j, k))
            v = j[k].string_value();
    }
    void rdVec3(const Json& j, Piccolo::Vector3& v)
    {
        rd(j, "x", v.x);
        rd(j, "y", v.y);
        rd(j, "z", v.z);
    }
    void rdQuat(const Json& j, Piccolo::Quaternion& q)
    {
        rd(j, "w", q.w);
        rd(j, "x", q.x);
        rd(j, "y", q.y);
        rd(j, "z", q.z);
    }
    void rdMat(const Json& j, Piccolo::Matrix4x4_& m)
    {
        float*      p[16] = {&m.v0, &m.v1, &m.v2, &m.v3, &m.v4, &m.v5, &m.v6, &m.v7,
                             &m.v8, &m.v9, &m.v10, &m.v11, &m.v12, &m.v13, &m.v14, &m.v15};
        for (int i = 0; i < 16; i++)
        {
            std::string k = "v" + std::to_string(i);
            rd(j, k.c_str(), *p[i]);
        }
    }
    void rdTransform(const Json& j, Piccolo::Transform& t)
    {
        if (has(j, "position"))
            rdVec3(j["position"], t.m_position);
        if (has(j, "scale"))
            rdVec3(j["scale"], t.m_scale);
        if (has(j, "rotation"))
            rdQuat(j["rotation"], t.m_rotation);
    }
} // namespace

extern "C" void ref_set_asset_root(const char* root) { g_asset_root = root ? root : ""; }

namespace Piccolo
{
    std::shared_ptr<SkeletonData> AnimationLoader::loadSkeletonData(std::string url)
    {
        SkeletonData data {};
        Json         j;
        if (loadJson(url, j))
        {
            rd(j, "is_flat", data.is_flat);
            rd(j, "root_index", data.root_index);
            rd(j, "in_topological_order", data.in_topological_order);
            if (has(j, "bones_map"))
            {
                const Json::array& arr = j["bones_map"].array_items();
                data.bones_map.resize(arr.size()); // value-init: index = parent_index = 0
                for (size_t i = 0; i < arr.size(); i++)
                {
                    RawBone& b = data.bones_map[i];
                    rd(arr[i], "name", b.name);
                    rd(arr[i], "index", b.index);
                    rd(arr[i], "parent_index", b.parent_index);
                    if (has(arr[i], "binding_pose"))
                        rdTransform(arr[i]["binding_pose"], b.binding_pose);
                    if (has(arr[i], "tpose_matrix"))
                        rdMat(arr[i]["tpose_matrix"], b.tpose_matrix);
                }
            }
        }
        return std::make_shared<SkeletonData>(data);
    }

    std::shared_ptr<AnimationClip> AnimationLoader::loadAnimationClipData(std::string url)
    {
        AnimationAsset asset {};
        Json           j;
        if (loadJson(url, j) && has(j, "clip_data"))
        {
            const Json&    c    = j["clip_data"];
            AnimationClip& clip = asset.clip_data;
            rd(c, "total_frame", clip.total_frame);
            rd(c, "node_count", clip.node_count);
            if (has(c, "node_channels"))
            {
                const Json::array& arr = c["node_channels"].array_items();
                clip.node_channels.resize(arr.size());
                for (size_t i = 0; i < arr.size(); i++)
                {
                    AnimationChannel& ch = clip.node_channels[i];
                    rd(arr[i], "name", ch.name);
                    if (has(arr[i], "position_keys"))
                    {
                        const Json::array& k = arr[i]["position_keys"].array_items();
                        ch.position_keys.resize(k.size());
                        for (size_t n = 0; n < k.size(); n++)
                            rdVec3(k[n], ch.position_keys[n]);
                    }
                    if (has(arr[i], "rotation_keys"))
                    {
                        const Json::array& k = arr[i]["rotation_keys"].array_items();
                        ch.rotation_keys.resize(k.size());
                        for (size_t n = 0; n < k.size(); n++)
                            rdQuat(k[n], ch.rotation_keys[n]);
                    }
                    if (has(arr[i], "scaling_keys"))
                    {
                        const Json::array& k = arr[i]["scaling_keys"].array_items();
                        ch.scaling_keys.resize(k.size());
                        for (size_t n = 0; n < k.size(); n++)
                            rdVec3(k[n], ch.scaling_keys[n]);
                    }
                }
            }
        }
        return std::make_shared<AnimationClip>(asset.clip_data);
    }

    std::shared_ptr<AnimSkelMap> AnimationLoader::loadAnimSkelMap(std::string url)
    {
        AnimSkelMap data {};
        Json        j;
        if (loadJson(url, j) && has(j, "convert"))
        {
            const Json::array& arr = j["convert"].array_items();
            data.convert.resize(arr.size());
            for (size_t i = 0; i < arr.size(); i++)
                data.convert[i] = static_cast<int>(arr[i].number_value());
        }
        return std::make_shared<AnimSkelMap>(data);
    }

    std::shared_ptr<BoneBlendMask> AnimationLoader::loadSkeletonMask(std::string url)
    {
        BoneBlendMask data {};
        Json          j;
        if (loadJson(url, j))
        {
            rd(j, "skeleton_file_path", data.skeleton_file_path);
            if (has(j, "enabled"))
            {
                const Json::array& arr = j["enabled"].array_items();
                data.enabled.resize(arr.size());
                for (size_t i = 0; i < arr.size(); i++)
                    data.enabled[i] = static_cast<int>(arr[i].number_value());
            }
        }
        return std::make_shared<BoneBlendMask>(data);
    }
} // namespace Piccolo
