// Host-side C++ mirror of Piccolo's runtime animation interface, implemented over the C ABI
// (include/piccolo_b200.h).  Same class names, member names, argument meaning and (silent) error
// behaviour as the reference, so that AnimationComponent / MeshComponent / LevelDebugger callers
// compile and behave unchanged; the work happens in CUDA kernels.
//
// Mirrors (file:line relative to /root/reference/engine/source/runtime):
//   resource/res_type/data/skeleton_data.h:9-35, animation_clip.h:9-50, animation_skeleton_node_map.h:9-16,
//   skeleton_mask.h:9-16, blend_state.h:10-45, mesh_data.h:9-53, resource/res_type/components/animation.h:12-43
//   function/animation/animation_loader.h:13-20, animation_system.h:15-31, skeleton.h:12-28, node.h:10-115
//   function/framework/component/animation/animation_component.h:10-30
// plus AnimationBatch, the crowd form (N components sharing a skeleton in one launch), which has no
// counterpart in the single-threaded reference.
#pragma once
#include "piccolo_b200.h"

#include <cstdint>
#include <limits>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace Piccolo
{
    // ---- core/math value types (layout and defaults of vector3.h:17-19, quaternion.h:19, matrix4.h:50-65) ----
    struct Vector3
    {
        float x {0.f}, y {0.f}, z {0.f};
    };
    struct Quaternion
    {
        float w {1.f}, x {0.f}, y {0.f}, z {0.f};
    };
    struct Matrix4x4_
    {
        float v0 {1.f}, v1 {0}, v2 {0}, v3 {0}, v4 {0}, v5 {1.f}, v6 {0}, v7 {0}, v8 {0}, v9 {0}, v10 {1.f}, v11 {0}, v12 {0}, v13 {0},
            v14 {0}, v15 {1.f};
    };
    struct Matrix4x4
    {
        float m_mat[4][4];
        Matrix4x4();
        Matrix4x4(const Matrix4x4_& m);
        Matrix4x4_ toMatrix4x4_() const;
    };
    struct Transform // core/math/transform.h:10-29
    {
        Vector3    m_position;
        Vector3    m_scale {1.f, 1.f, 1.f};
        Quaternion m_rotation;
    };

    // ---- asset schema -------------------------------------------------------------------------
    struct RawBone
    {
        std::string name;
        int         index {0};
        Transform   binding_pose;
        Matrix4x4_  tpose_matrix;
        int         parent_index {0};
    };
    struct SkeletonData
    {
        std::vector<RawBone> bones_map;
        bool                 is_flat {false};
        int                  root_index {0};
        bool                 in_topological_order {false};
    };
    struct AnimNodeMap
    {
        std::vector<std::string> convert;
    };
    struct AnimationChannel
    {
        std::string             name;
        std::vector<Vector3>    position_keys;
        std::vector<Quaternion> rotation_keys;
        std::vector<Vector3>    scaling_keys;
    };
    struct AnimationClip
    {
        int                           total_frame {0};
        int                           node_count {0};
        std::vector<AnimationChannel> node_channels;
    };
    struct AnimationAsset
    {
        AnimNodeMap   node_map;
        AnimationClip clip_data;
        std::string   skeleton_file_path;
    };
    struct AnimSkelMap
    {
        std::vector<int> convert;
    };
    struct BoneBlendMask
    {
        std::string      skeleton_file_path;
        std::vector<int> enabled;
    };
    struct BoneBlendWeight
    {
        std::vector<float> blend_weight;
    };
    struct BlendStateWithClipData
    {
        int                          clip_count {0};
        std::vector<AnimationClip>   blend_clip;
        std::vector<AnimSkelMap>     blend_anim_skel_map;
        std::vector<BoneBlendWeight> blend_weight;
        std::vector<float>           blend_ratio;
        // extension: where the data came from, so the device copies cached by AnimationManager are reused
        // instead of re-uploading the clip on every tick.  Empty when a caller fills the struct by hand.
        std::vector<std::string> source_clip_path, source_map_path, source_mask_path;
        std::vector<float>       source_weight;
    };
    struct BlendState
    {
        int                      clip_count {0};
        std::vector<std::string> blend_clip_file_path;
        std::vector<float>       blend_clip_file_length;
        std::vector<std::string> blend_anim_skel_map_path;
        std::vector<float>       blend_weight;
        std::vector<std::string> blend_mask_file_path;
        std::vector<float>       blend_ratio;
    };
    struct AnimationResultElement
    {
        int        index {0};
        Matrix4x4_ transform;
    };
    struct AnimationResult
    {
        std::vector<AnimationResultElement> node;
    };
    struct AnimationComponentRes
    {
        std::string     skeleton_file_path;
        BlendState      blend_state;
        float           frame_position {0.f};
        AnimationResult animation_result;
    };
    struct Vertex
    {
        float px {0}, py {0}, pz {0}, nx {0}, ny {0}, nz {0}, tx {0}, ty {0}, tz {0}, u {0}, v {0};
    };
    struct SkeletonBinding
    {
        int   index0 {0}, index1 {0}, index2 {0}, index3 {0};
        float weight0 {0}, weight1 {0}, weight2 {0}, weight3 {0};
    };
    struct MeshData
    {
        std::vector<Vertex>          vertex_buffer;
        std::vector<int>             index_buffer;
        std::vector<SkeletonBinding> bind;
    };

    // ---- function/animation/animation_loader.h ---------------------------------------------------
    class AnimationLoader
    {
    public:
        std::shared_ptr<AnimationClip> loadAnimationClipData(std::string animation_clip_url);
        std::shared_ptr<SkeletonData>  loadSkeletonData(std::string skeleton_data_url);
        std::shared_ptr<AnimSkelMap>   loadAnimSkelMap(std::string anim_skel_map_url);
        std::shared_ptr<BoneBlendMask> loadSkeletonMask(std::string skeleton_mask_file_url);
        // additions (the reference loads these through AssetManager::loadAsset<T>)
        std::shared_ptr<MeshData>              loadMeshData(std::string mesh_url);
        std::shared_ptr<AnimationComponentRes> loadAnimationComponentRes(std::string object_url);
        // AssetManager::getFullPath stand-in: prefix joined in front of every url
        static void               setAssetRoot(const std::string& root);
        static const std::string& assetRoot();
        // Binary cache of the parsed assets (asset_cache.cpp).  Empty (default) = off, like the reference.
        // With a directory set, a load first looks for a flat image of the asset there (validated against the JSON file's
        // size and modification time, and by a checksum) and writes one after parsing otherwise.
        static void               setCacheDir(const std::string& dir);
        static const std::string& cacheDir();
    };
    namespace detail // asset_cache.cpp
    {
        bool cacheLoad(const std::string& full_path, SkeletonData& out);
        bool cacheLoad(const std::string& full_path, AnimationClip& out);
        bool cacheLoad(const std::string& full_path, AnimSkelMap& out);
        bool cacheLoad(const std::string& full_path, BoneBlendMask& out);
        bool cacheLoad(const std::string& full_path, MeshData& out);
        void cacheStore(const std::string& full_path, const SkeletonData& v);
        void cacheStore(const std::string& full_path, const AnimationClip& v);
        void cacheStore(const std::string& full_path, const AnimSkelMap& v);
        void cacheStore(const std::string& full_path, const BoneBlendMask& v);
        void cacheStore(const std::string& full_path, const MeshData& v);
    } // namespace detail

    namespace detail
    {
        struct PoseGroup;
        void destroyPoseGroups();
    }

    // ---- function/animation/node.h (the read side LevelDebugger uses) -------------------------------
    class Node
    {
    public:
        enum class TransformSpace { LOCAL, AREN, OBJECT };
        virtual ~Node() = default;
        const std::string& getName() const { return m_name; }
        virtual Node*      getParent() const { return m_parent; }
        virtual const Quaternion& _getDerivedOrientation() const { return m_derived_orientation; }
        virtual const Vector3&    _getDerivedPosition() const { return m_derived_position; }
        virtual const Vector3&    _getDerivedScale() const { return m_derived_scale; }
        virtual const Matrix4x4&  _getInverseTpose() const { return m_inverse_Tpose; }
        virtual const Vector3&    getInitialPosition() const { return m_initial_position; }
        virtual const Quaternion& getInitialOrientation() const { return m_initial_orientation; }
        virtual const Vector3&    getInitialScale() const { return m_initial_scale; }

    protected:
        friend class Skeleton;
        Node*       m_parent {nullptr};
        std::string m_name;
        Quaternion  m_derived_orientation;
        Vector3     m_derived_position;
        Vector3     m_derived_scale {1.f, 1.f, 1.f};
        Vector3     m_initial_position;
        Quaternion  m_initial_orientation;
        Vector3     m_initial_scale {1.f, 1.f, 1.f};
        Matrix4x4   m_inverse_Tpose;
    };
    class Bone : public Node
    {
    public:
        size_t getID() const { return m_has_definition ? (size_t)m_index : std::numeric_limits<size_t>::max(); }

    protected:
        friend class Skeleton;
        bool m_has_definition {false};
        int  m_index {0};
    };

    // ---- function/animation/skeleton.h -------------------------------------------------------------
    class Skeleton
    {
    public:
        Skeleton() = default;
        ~Skeleton();
        Skeleton(const Skeleton&)            = delete; // the reference's copy would double free (raw owning pointer)
        Skeleton& operator=(const Skeleton&) = delete;

        void            buildSkeleton(const SkeletonData& skeleton_definition);
        void            applyAnimation(const BlendStateWithClipData& blend_state);
        AnimationResult outputAnimationResult();
        void            resetSkeleton();
        const Bone*     getBones() const;
        int32_t         getBonesCount() const;

    private:
        friend class AnimationComponent;
        friend struct detail::PoseGroup;
        friend void detail::destroyPoseGroups();
        bool       m_is_flat {false};
        int        m_bone_count {0};
        Bone*      m_bones {nullptr};
        pcl_skel   m_skel {-1};
        // Where the pose is computed.  A Skeleton used on its own (the reference's applyAnimation -> outputAnimationResult
        // pair) owns a batch of one; a Skeleton that belongs to an AnimationComponent shares ONE hidden batch with every
        // other component of the same skeleton and clip count (detail::PoseGroup): applyAnimation then only records the
        // frame's inputs, and the first result anybody asks for poses the whole group with one launch and one read-back.
        pcl_batch*         m_batch {nullptr};
        int                m_batch_k {0};
        bool               m_grouped {false};
        detail::PoseGroup* m_group {nullptr};
        int                m_slot {-1};
        bool               m_derived_stale {true};
        std::vector<int32_t> m_index; // AnimationResultElement::index per bone (getID() + 1)
        void ensureBatch(int k);
        void refreshDerived();
        void leaveGroup();
        void applyResolved(int K, const int32_t* clip, const float* phase, const float* weight, const int32_t* mask);
        void fillResult(const float* rows12, AnimationResult& out) const;
    };

    // ---- function/animation/animation_system.h ------------------------------------------------------
    class AnimationManager
    {
    private:
        static std::map<std::string, std::shared_ptr<SkeletonData>>  m_skeleton_definition_cache;
        static std::map<std::string, std::shared_ptr<AnimationClip>> m_animation_data_cache;
        static std::map<std::string, std::shared_ptr<AnimSkelMap>>   m_animation_skeleton_map_cache;
        static std::map<std::string, std::shared_ptr<BoneBlendMask>> m_skeleton_mask_cache;

    public:
        static std::shared_ptr<SkeletonData>  tryLoadSkeleton(std::string file_path);
        static std::shared_ptr<AnimationClip> tryLoadAnimation(std::string file_path);
        static std::shared_ptr<AnimSkelMap>   tryLoadAnimationSkeletonMap(std::string file_path);
        static std::shared_ptr<BoneBlendMask> tryLoadSkeletonMask(std::string file_path);
        static BlendStateWithClipData         getBlendStateWithClipData(const BlendState& blend_state);
        AnimationManager() = default;

        // device-side caches (extension): one upload per (skeleton, clip path, map path) / (skeleton, mask path)
        static pcl_clip deviceClip(pcl_skel skel, const std::string& clip_path, const std::string& map_path);
        static pcl_mask deviceMask(pcl_skel skel, const std::string& mask_path);
        static pcl_clip deviceClipFromData(pcl_skel skel, const AnimationClip& clip, const AnimSkelMap& map);
        static pcl_skel deviceSkeleton(const SkeletonData& def); // one upload per distinct skeleton content
        static void     clearCaches(); // tests
    };

    // ---- the process-wide device context the host classes share ------------------------------------
    class Runtime
    {
    public:
        // created on first use on device $PICCOLO_B200_DEVICE (default 0); nullptr when no GPU is present
        // (every host class then behaves like the reference does on an invalid skeleton: silently inert)
        static pcl_ctx*    context();
        static const char* lastError();
        static void        shutdown();
        // The reference plays clip 0 with weight 1 whatever clip_count says (skeleton.cpp:46,63-64), and so do Skeleton and
        // AnimationComponent here by default.  Switching the blend extension on makes them blend clip_count = 2..4 clips
        // with the blend state's weights and masks (DESIGN.md A4 - no reference behaviour to compare with).  AnimationBatch,
        // which has no counterpart in the reference, always blends.
        static void enableBlendExtension(bool on);
        static bool blendExtensionEnabled();
    };

    // ---- function/framework/component (only what this path needs) ------------------------------------
    class GObject;
    class Component
    {
    protected:
        std::weak_ptr<GObject> m_parent_object;

    public:
        Component()          = default;
        virtual ~Component() = default;
        virtual void postLoadResource(std::weak_ptr<GObject> parent_object) { m_parent_object = parent_object; }
        virtual void tick(float) {}
    };

    class AnimationComponent : public Component
    {
    public:
        AnimationComponent() = default;
        void                   postLoadResource(std::weak_ptr<GObject> parent_object) override;
        void                   tick(float delta_time) override;
        const AnimationResult& getResult() const;
        const Skeleton&        getSkeleton() const;
        // the reference fills m_animation_res through reflection when the object JSON is loaded
        void                         setAnimationRes(const AnimationComponentRes& res) { m_animation_res = res; }
        const AnimationComponentRes& getAnimationRes() const { return m_animation_res; }

    protected:
        mutable AnimationComponentRes m_animation_res; // animation_result is filled in when it is first asked for after a tick
        mutable Skeleton              m_skeleton;
        mutable bool                  m_result_stale {false};
        // handles resolved from the blend state's paths, re-resolved only when a path changes (no per-tick clip copies)
        std::vector<std::string> m_resolved_paths;
        int32_t                  m_clip[4] {-1, -1, -1, -1}, m_mask[4] {-1, -1, -1, -1};
    };

    // ---- crowd form: N AnimationComponents that share a skeleton, one launch per tick ----------------
    class AnimationBatch
    {
    public:
        AnimationBatch() = default;
        ~AnimationBatch();
        AnimationBatch(const AnimationBatch&)            = delete;
        AnimationBatch& operator=(const AnimationBatch&) = delete;

        // every instance starts from `res` (same skeleton / clips / masks); per-instance phases and weights
        // are then set with setPhases / setWeights.  Returns false (and stays inert) on invalid assets.
        bool create(const AnimationComponentRes& res, int64_t instance_count);
        void setPhases(const std::vector<float>& blend_ratio /*[N][clip_count]*/);
        void setWeights(const std::vector<float>& blend_weight /*[N][clip_count]*/);
        void tick(float delta_time);                       // AnimationComponent::tick for all N instances
        AnimationResult getResult(int64_t instance) const; // one instance's AnimationResult (synchronous)
        void            getResults(int64_t first, int64_t count, std::vector<float>& mats16) const; // [count][J][16]
        std::vector<float> getPhases() const;
        // skinning (mesh.vert) of instances [first, first+count) into ring slots [0, count)
        bool     attachMesh(const MeshData& mesh, int64_t ring_capacity, bool normals_and_tangents);
        bool     skin(int64_t first, int64_t count);
        void     downloadSkinned(int64_t slot_first, int64_t count, std::vector<float>& pos, std::vector<float>& normal_tangent) const;
        int64_t  size() const { return m_n; }
        int32_t  boneCount() const { return m_j; }
        pcl_batch* handle() const { return m_batch; }

    private:
        pcl_skel     m_skel {-1};
        pcl_batch*   m_batch {nullptr};
        pcl_mesh     m_mesh {-1};
        pcl_skinbuf* m_skinbuf {nullptr};
        int64_t      m_n {0};
        int32_t      m_j {0}, m_k {0}, m_v {0};
        bool         m_nt {false};
        std::vector<int32_t> m_index;
    };

    // ---- asset-side conventions (function/animation/utilities.cpp:46-53, animation_loader.cpp:30-71) -------------
    // The reference keeps these in an anonymous namespace of its offline converter and never calls them at run time;
    // they are what produced the bundled anim.skeleton_map.json / robot.mesh_bind.json and are provided so that new
    // clips and meshes can be bound the same way.
    int         find_index_by_name(const SkeletonData& skeleton, const std::string& name); // INT_MAX when absent
    AnimSkelMap buildAnimSkelMap(const AnimNodeMap& anim, const SkeletonData& skeleton);    // channel name -> bone index
    // first free of the vertex's four slots gets (bone_index + 1, weight); returns false when all four are taken
    bool addBoneBind(MeshData& mesh_bind, size_t bone_index, size_t vertex_index, float weight);

    // RawBone arrays for pcl_upload_skeleton (shared by Skeleton and AnimationBatch)
    pcl_skel uploadSkeleton(const SkeletonData& def);
    pcl_mesh uploadMesh(const MeshData& mesh, int32_t max_slot);
} // namespace Piccolo
