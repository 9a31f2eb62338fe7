/*
 * piccolo_b200 -- C ABI of the B200-native batch skeletal-animation + skinning engine.
 *
 * Drop-in scope: Piccolo's runtime animation path
 *   AnimationComponent::tick -> AnimationManager::getBlendStateWithClipData -> Skeleton::applyAnimation
 *   -> Node/Bone local->model update -> Skeleton::outputAnimationResult -> mesh.vert 4-weight LBS.
 * The host-side C++ mirror of the reference classes (piccolo_b200/host/) and any foreign binding sit on
 * top of exactly these entry points.  Plain pointers and sizes only; no torch / STL types.
 *
 * Conventions
 *   - every function returns PCL_OK (0) or a negative pcl_status; pcl_last_error() gives the text.
 *   - the caller owns every host array; the library owns all device memory it allocates.
 *   - one context per CUDA device.  Calls on one context are NOT thread-safe; distinct contexts are
 *     independent (one per GPU / per rank).  Work is enqueued on the context's streams and is
 *     asynchronous unless stated; pcl_sync() waits for it.
 *   - there is no CPU fallback: without a CUDA device pcl_create fails with PCL_ERR_NO_DEVICE.
 *   - quaternions are (w,x,y,z) like Piccolo::Quaternion (core/math/quaternion.h:19); matrices are
 *     row-major, column-vector convention, translation in [r][3] (core/math/matrix4.h:98-119).
 *
 * file:line citations are relative to /root/reference/engine/source/runtime unless they start with
 * "engine/".
 */
#ifndef PICCOLO_B200_H
#define PICCOLO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum pcl_status
{
    PCL_OK              = 0,
    PCL_ERR_INVALID     = -1, /* bad argument / malformed asset                                        */
    PCL_ERR_CUDA        = -2, /* a CUDA runtime call failed (text in pcl_last_error)                   */
    PCL_ERR_UNSUPPORTED = -3, /* well-formed input this engine does not handle (see each function)     */
    PCL_ERR_OOM         = -4,
    PCL_ERR_NO_DEVICE   = -5
} pcl_status;

typedef struct pcl_ctx     pcl_ctx;     /* one per device                                  */
typedef struct pcl_batch   pcl_batch;   /* N animated instances of one skeleton            */
typedef struct pcl_skinbuf pcl_skinbuf; /* device ring of skinned vertex streams           */
typedef int32_t            pcl_skel;    /* handles are small integers scoped to a context  */
typedef int32_t            pcl_clip;
typedef int32_t            pcl_mask;
typedef int32_t            pcl_mesh;

/* pcl_batch_advance / pcl_batch_tick */
#define PCL_ADVANCE_FIRST_ONLY 1 /* reference behaviour: only blend_ratio[0] moves (animation_component.cpp:19-21) */
#define PCL_ADVANCE_ALL 2        /* every clip slot advances with its own clip length                               */
/* pcl_skinbuf_create / pcl_batch_skin */
#define PCL_SKIN_POS 1           /* positions only (shadow / pick shader variants)                                  */
#define PCL_SKIN_POS_NRM_TAN 7   /* positions + normals + tangents (engine/shader/glsl/mesh.vert)                   */

/* ---- lifecycle ------------------------------------------------------------------------------ */
int         pcl_version(void);
int         pcl_device_count(void);
int         pcl_create(int device, pcl_ctx** out);
int         pcl_destroy(pcl_ctx* ctx);
const char* pcl_last_error(const pcl_ctx* ctx); /* ctx == NULL: error of the last failed pcl_create */
int         pcl_sync(pcl_ctx* ctx);
/* number of this library's kernels launched on the context so far */
int64_t     pcl_launch_count(const pcl_ctx* ctx);
/* pinned host memory for the download / set calls (cudaHostAlloc); optional, any host memory works.
 * Blocks not freed with pcl_host_free are released by pcl_destroy. */
int         pcl_host_alloc(pcl_ctx* ctx, size_t bytes, void** out);
int         pcl_host_free(pcl_ctx* ctx, void* p);
/* device timing on the context's compute stream: begin records an event, end waits for all the
 * context's streams, records, synchronises and returns the elapsed milliseconds */
int         pcl_timer_begin(pcl_ctx* ctx);
int         pcl_timer_end(pcl_ctx* ctx, float* ms);

/* ---- assets: repacked once into SoA device buffers --------------------------------------------
 * pcl_upload_skeleton replaces Skeleton::buildSkeleton (function/animation/skeleton.cpp:17-37) with
 * find_by_index (function/animation/utilities.cpp:7-27: parent_index == INT_MAX or >= own position
 * => root) and Bone::initialize (function/animation/node.cpp:220-234: the bind quaternion is
 * normalised, NaN -> identity).  Arrays follow SkeletonData::bones_map order
 * (resource/res_type/data/skeleton_data.h:9-35).
 *   PCL_ERR_UNSUPPORTED when !is_flat || !in_topological_order (the reference silently builds
 *   nothing, skeleton.cpp:24-28) or when an inverse-bind matrix is not affine (row 3 != 0 0 0 1).
 *   PCL_ERR_INVALID for J < 1 or a negative parent_index (undefined behaviour in the reference). */
int pcl_upload_skeleton(pcl_ctx* ctx, int32_t J, const int32_t* index /*[J] RawBone.index*/,
                        const int32_t* parent_index /*[J]*/, const float* bind_pos /*[3J]*/,
                        const float* bind_rot_wxyz /*[4J]*/, const float* bind_scl /*[3J]*/,
                        const float* inv_bind /*[16J] RawBone.tpose_matrix*/, int32_t is_flat,
                        int32_t in_topological_order, pcl_skel* out);
int pcl_skeleton_bone_count(pcl_ctx* ctx, pcl_skel skel, int32_t* J);
/* AnimationResultElement.index per bone = RawBone.index + 1 (skeleton.cpp:127, node.cpp:237-244) */
int pcl_skeleton_indices(pcl_ctx* ctx, pcl_skel skel, int32_t* out /*[J]*/);
/* parent position per bone after find_by_index (-1 = root) */
int pcl_skeleton_parents(pcl_ctx* ctx, pcl_skel skel, int32_t* out /*[J]*/);

/* One AnimationClip (resource/res_type/data/animation_clip.h:18-39) together with the AnimSkelMap it
 * is played through (resource/res_type/data/animation_skeleton_node_map.h:9-16), i.e. one entry of
 * BlendStateWithClipData::blend_clip / blend_anim_skel_map (resource/res_type/data/blend_state.h:19-30).
 * Tracks are ragged: key_counts[3c+0..2] = number of position / rotation / scaling keys of channel c,
 * stored back to back in pos / rot_wxyz / scl in channel order.  Only channels
 * n < min(node_count, convert_len) are played (skeleton.cpp:57-59); convert[n] == -1 skips a channel
 * (skeleton.cpp:69-73).  The key-count clamp that carries over from one channel to the next
 * (skeleton.cpp:75-87) is reproduced.
 *   PCL_ERR_INVALID: node_count > C, convert[n] outside {-1} U [0,J), or a played channel with an
 *   empty track (the reference would index [-1]). */
int pcl_upload_clip(pcl_ctx* ctx, pcl_skel skel, int32_t total_frame, int32_t node_count, int32_t C,
                    const int32_t* key_counts /*[3C]*/, const float* pos, const float* rot_wxyz, const float* scl,
                    int32_t convert_len, const int32_t* convert, pcl_clip* out);
/* BoneBlendMask::enabled (resource/res_type/data/skeleton_mask.h:9-16); used only when a batch blends
 * K > 1 clips (the reference forces weight = 1 for its single clip, skeleton.cpp:63-64). */
int pcl_upload_mask(pcl_ctx* ctx, pcl_skel skel, int32_t J, const int32_t* enabled, pcl_mask* out);
/* Vertex streams + joint binding of one skinned mesh: MeshData (resource/res_type/data/mesh_data.h:9-53)
 * as RenderResourceBase::loadMeshData lays it out (function/render/render_resource_base.cpp:96-148).
 * idx4 holds palette slots (bone position + 1, 0 = empty, function/animation/animation_loader.cpp:40-71);
 * w4_raw is renormalised here exactly like RenderResource::updateVertexBuffer
 * (function/render/render_resource.cpp:714-725).  The mesh must be fully expanded (one binding per
 * vertex, the case in which the reference's per-index binding upload is self-consistent,
 * render_resource.cpp:705-712).  PCL_ERR_INVALID when an idx4 entry is < 0 or > max_slot. */
int pcl_upload_mesh(pcl_ctx* ctx, int32_t V, const float* pos /*[3V]*/, const float* nrm /*[3V]*/,
                    const float* tan /*[3V]*/, const int32_t* idx4 /*[4V]*/, const float* w4_raw /*[4V]*/,
                    int32_t max_slot, pcl_mesh* out);

/* Releasing assets: what a level unload does to the caches of AnimationManager (animation_system.cpp:10-13 - the reference
 * never drops them; a long-running host must).  A released handle is invalid from then on and is not handed out again.
 * pcl_clip_release also gives the clip's (clip, frame) buckets back, so that reloading levels does not grow the bucket count
 * past what the frame-coherent order supports.  A skeleton can be released once no batch, pool or clip refers to it
 * (PCL_ERR_INVALID otherwise).  Masks live with their skeleton.  Each call waits for the context's work first. */
int pcl_clip_release(pcl_ctx* ctx, pcl_clip clip);
int pcl_mesh_release(pcl_ctx* ctx, pcl_mesh mesh);
int pcl_skeleton_release(pcl_ctx* ctx, pcl_skel skel);

/* ---- instances -------------------------------------------------------------------------------
 * A batch = N independent AnimationComponents sharing one skeleton, each holding K clip slots
 * (BlendState::clip_count, resource/res_type/data/blend_state.h:32-45): clip id, blend_ratio (phase),
 * blend_weight, blend_clip_file_length and mask id per slot.  K == 1 is the reference's executed path;
 * K in 2..4 is this engine's documented blend extension (DESIGN.md "A4").  K > 4: PCL_ERR_UNSUPPORTED. */
int pcl_batch_create(pcl_ctx* ctx, pcl_skel skel, int64_t N, int32_t K, pcl_batch** out);
int pcl_batch_destroy(pcl_batch* b);
/* Host arrays are [count][K]; a NULL array leaves that field unchanged.  Defaults after create:
 * clip 0, phase 0, weight 1, length 1, mask -1 (= every bone enabled).  Synchronous. */
int pcl_batch_set(pcl_batch* b, int64_t first, int64_t count, const int32_t* clip_ids, const float* phase,
                  const float* weight, const float* clip_len, const int32_t* mask_ids);
/* The per-frame part of pcl_batch_set - the fields gameplay code rewrites every frame, BlendState::blend_ratio and
 * blend_weight (resource/res_type/data/blend_state.h:32-45) - without stalling the caller: phase / weight ([count][K], either may
 * be NULL) are copied into a pinned staging buffer owned by the batch and enqueued on the compute stream, ordered before the
 * next pose; the call returns without waiting for the GPU (it waits only for the previous staging copy to have left the
 * buffer), and the caller's arrays may be reused as soon as it returns. */
int pcl_batch_set_frame_async(pcl_batch* b, int64_t first, int64_t count, const float* phase, const float* weight);
int pcl_batch_get_phases(pcl_batch* b, int64_t first, int64_t count, float* phase /*[count][K]*/);
/* ratio += dt / length; ratio -= floor(ratio)   (animation_component.cpp:19-21), on the device */
int pcl_batch_advance(pcl_batch* b, float dt, int32_t mode);

/* ---- hot path --------------------------------------------------------------------------------
 * pcl_batch_pose = Skeleton::applyAnimation + outputAnimationResult for every instance
 * (skeleton.cpp:39-156): keyframe sampling (lerp / lerp / nlerp shortest path), bind-relative local
 * pose, level-ordered local->model update, model x inverse-bind.  Product: the device palette
 * [N][pitch][3] float4 = rows 0..2 of each 4x4 (row 3 is 0 0 0 1), slot 0 = identity, slot i+1 = bone i
 * (the layout MeshComponent::tick hands to the renderer, function/framework/component/mesh/
 * mesh_component.cpp:72-79).  pcl_batch_tick = pcl_batch_advance + pcl_batch_pose in one launch
 * (AnimationComponent::tick, animation_component.cpp:17-25). */
int pcl_batch_pose(pcl_batch* b);
int pcl_batch_tick(pcl_batch* b, float dt, int32_t mode);
/* when enabled, pose/tick also keep the derived (model-space) TRS per bone that LevelDebugger reads
 * through Skeleton::getBones() (function/framework/level/level_debugger.cpp:110-146) */
int pcl_batch_keep_derived(pcl_batch* b, int32_t enable);

/* Skinned vertex streams live in a device ring of `capacity` instance slots:
 *   positions        [capacity][Vp][3] float   (MeshVertex::VulkanMeshVertexPostition,
 *   normal+tangent   [capacity][Vp][6] float    ...VaryingEnableBlending, function/render/render_mesh.h:14-23)
 * Vp = V rounded up to a multiple of 8 (keeps every instance and every 8-vertex run on 32-byte sectors). */
int pcl_skinbuf_create(pcl_ctx* ctx, pcl_mesh mesh, int64_t capacity, int32_t flags, pcl_skinbuf** out);
int pcl_skinbuf_destroy(pcl_skinbuf* s);
/* engine/shader/glsl/mesh.vert:72-111 for instances [first, first+count) of the batch, written to ring
 * slots [slot_first, slot_first+count).  Reads the palette of the last pcl_batch_pose / _tick. */
int pcl_batch_skin(pcl_batch* b, pcl_mesh mesh, pcl_skinbuf* out, int64_t first, int64_t count, int64_t slot_first);

/* One frame of a SMALL batch in ONE call - the reference's everyday case, a handful of characters each ticked once per
 * frame (function/framework/object/object.cpp:41-50 -> component/animation/animation_component.cpp:17-25 ->
 * component/mesh/mesh_component.cpp:68-79): sets the per-frame inputs of ALL instances ([N][K] arrays like pcl_batch_set;
 * NULL = keep the previous frame's; weight, mask default to 1 / none), samples + poses them, and returns their
 * AnimationResult rows (host12: [N][J][12], layout of pcl_batch_download_palette3x4) - one kernel launch and one
 * synchronisation: inputs and rows live in a pinned block the kernel reads and writes in place.  host12 == NULL returns
 * without waiting; pcl_batch_frame_result waits and copies the rows out.  The device palette is valid afterwards
 * (pcl_batch_skin, pcl_batch_export_render...).  Inputs given here are separate from pcl_batch_set's device arrays;
 * phases are advanced by the caller (as AnimationComponent::tick does on the host).  PCL_ERR_UNSUPPORTED when
 * N * (K * 20 + J * 48) exceeds PCL_FRAME_MAX_BYTES. */
#define PCL_FRAME_MAX_BYTES (64u << 20)
int pcl_batch_frame(pcl_batch* b, const int32_t* clip_ids, const float* phase, const float* weight, const int32_t* mask_ids, float* host12);
int pcl_batch_frame_result(pcl_batch* b, float* host12);

/* ---- readback / interop ----------------------------------------------------------------------- */
/* AnimationResult for instances [first, first+count): [count][J][16] floats, bones_map order, no
 * identity slot (resource/res_type/components/animation.h:12-29).  Synchronous. */
int pcl_batch_download_palette(pcl_batch* b, int64_t first, int64_t count, float* host16);
/* Same, enqueued on the context's copy stream after the last pose; returns at once.  host16 must stay
 * valid until pcl_sync().  Overlaps with a following pcl_batch_skin. */
int pcl_batch_download_palette_async(pcl_batch* b, int64_t first, int64_t count, float* host16);
/* Compact read-back: rows 0..2 of every joint matrix, [count][J][12] floats (48 B per bone instead of 64), bones_map order,
 * no identity slot.  Row 3 of every AnimationResult matrix is the constant 0 0 0 1 (Matrix4x4::toMatrix4x4_ of an affine
 * product, core/math/matrix4.h:98-119; the inverse binds are checked to be affine at upload), so the caller fills it in
 * on the host - the host mirror's AnimationBatch does - and a quarter of the bytes never crosses the host link.
 * Asynchronous like pcl_batch_download_palette_async: enqueued on the copy stream behind the last pose, host12 must
 * stay valid until pcl_sync(); the next pose waits until the palette has been read.  How the bytes travel is the
 * "export_mode" option (pcl_set_option): 1 = one strided DMA straight out of the device palette (no kernel, no staging
 * buffer), 2 = a kernel writes into host12 itself, which must then be pinned (pcl_host_alloc / cudaHostAlloc /
 * cudaHostRegister; pageable memory falls back to 3), 3 = kernel -> device staging buffer -> linear DMA; 0 = auto. */
int pcl_batch_download_palette3x4_async(pcl_batch* b, int64_t first, int64_t count, float* host12);
/* the same, synchronous */
int pcl_batch_download_palette3x4(pcl_batch* b, int64_t first, int64_t count, float* host12);
/* derived TRS [count][J][10] = P.xyz, Q.wxyz, S.xyz (needs pcl_batch_keep_derived) */
int pcl_batch_download_derived(pcl_batch* b, int64_t first, int64_t count, float* host10);
/* device pointer + geometry of the compact palette, for consumers living on the GPU */
int pcl_batch_palette_dev(pcl_batch* b, const void** dptr, int32_t* pitch_slots, int32_t* floats_per_slot);
int pcl_skinbuf_dev(pcl_skinbuf* s, const void** pos_dev, const void** nt_dev, int32_t* Vp);
/* tight host copies: pos [count][V][3], nt [count][V][6] (nt may be NULL) */
int pcl_skinbuf_download(pcl_skinbuf* s, int64_t slot_first, int64_t count, float* host_pos, float* host_nt);

/* No counterpart in the reference.  Same result as pcl_batch_download_palette, for callers that must not
 * allocate device memory: a render or simulation thread inside its frame loop, where cudaMalloc/cudaFree
 * would stall every stream of the device and take an unbounded time, or a host application that gave the
 * whole GPU memory budget to its own allocator at start-up, so that a later cudaMalloc can fail.
 * pcl_batch_download_palette grows a device staging buffer with cudaMalloc when a larger range is asked
 * for; this variant stages through a fixed array of PCL_STATIC_EXPORT_MAX_BYTES that is part of the
 * library image (every GPU that loads the library holds it until the process exits) and calls neither
 * cudaMalloc nor cudaFree nor any other allocating routine.
 *   - limit: count * J * 64 bytes <= PCL_STATIC_EXPORT_MAX_BYTES (2259 instances of the 58-bone player
 *     skeleton), otherwise PCL_ERR_UNSUPPORTED and nothing is written; split the range into several calls.
 *   - the first call on a device looks up the array's device address (a one-time set-up); make that first
 *     call before starting any stream capture on the device.
 *   - synchronous.  All calls on one GPU share the one array, from every context and thread: calls that
 *     overlap on the GPU race and corrupt each other's result, so serialise them.  The array keeps the last
 *     call's matrices; every call overwrites the part it uses, nothing has to start at zero. */
#define PCL_STATIC_EXPORT_MAX_BYTES (8u << 20)
int pcl_batch_download_palette_static(pcl_batch* b, int64_t first, int64_t count, float* host16);

/* ---- render hand-off ------------------------------------------------------------------------------
 * The per-draw-call joint SSBO of the reference's mesh passes: mat4 joint_matrices[1024 * 64] per draw,
 * joint_matrices[1024 * instance_in_draw + slot], row-major 4x4, slot 0 = identity, <= 64 instances per draw
 * (function/render/passes/main_camera_pass.cpp:2284-2298, function/render/render_common.h:19-20,70-73,
 * engine/shader/include/constants.h:3-7).  pcl_batch_export_render writes instances [first, first+count) into draws
 * draw_first + i/64 at instance slot i%64, touching only the J+1 live slots of each instance like the reference's copy
 * loop; the buffer (4 MiB per draw, zeroed at creation) is what a Vulkan/CUDA interop SSBO would alias - the interop
 * itself is out of scope (no graphics stack on this box). */
typedef struct pcl_renderbuf pcl_renderbuf;
int pcl_renderbuf_create(pcl_ctx* ctx, int64_t draws, pcl_renderbuf** out);
int pcl_renderbuf_destroy(pcl_renderbuf* r);
int pcl_batch_export_render(pcl_batch* b, int64_t first, int64_t count, pcl_renderbuf* r, int64_t draw_first);
int pcl_renderbuf_dev(pcl_renderbuf* r, const void** dptr);
int pcl_renderbuf_download(pcl_renderbuf* r, int64_t draw, float* host /*[64][1024][16]*/);

/* ---- palette pool: the crowd's joint palettes gathered on ONE GPU -----------------------
 * The reference hands every animated object's AnimationResult to one consumer, the render system of the one GPU that
 * draws the frame (MeshComponent::tick, function/framework/component/mesh/mesh_component.cpp:68-100 ->
 * RenderSystem::processSwapData, function/render/render_system.cpp:302-310 -> the per-draw joint buffer,
 * function/render/passes/main_camera_pass.cpp:2284-2298).  With the crowd sharded over GPUs that consumer needs the
 * palettes of all shards.  A pool is one device buffer [capacity][pitch][3] float4 (the compact palette layout) on the
 * consumer's GPU; a batch attached to rows [first, first + N) of it makes the pose kernel write its palettes THERE -
 * straight over NVLink when the pool lives on another GPU (peer-to-peer stores from the kernel, no staging copy and no
 * collective).  Same process, other context: pass the owner's pool (peer access is enabled on attach).  Other process:
 * pcl_pool_export on the owner, carry the PCL_POOL_HANDLE_BYTES to the producer (torch.distributed broadcast, a pipe ...),
 * pcl_pool_open there.  Ordering across GPUs is the caller's: producers pcl_sync() and signal (a barrier) before the
 * consumer calls pcl_batch_adopt_palettes on its own batch attached to the same rows and skins / exports from it.
 * A pool must outlive the batches attached to it (pcl_pool_destroy refuses while any is attached). */
typedef struct pcl_pool pcl_pool;
#define PCL_POOL_HANDLE_BYTES 64
int pcl_pool_create(pcl_ctx* ctx, pcl_skel skel, int64_t capacity, pcl_pool** out);
int pcl_pool_export(pcl_pool* pool, void* handle /*[PCL_POOL_HANDLE_BYTES]*/);
int pcl_pool_open(pcl_ctx* ctx, pcl_skel skel, int64_t capacity, const void* handle, pcl_pool** out);
int pcl_pool_destroy(pcl_pool* pool);
/* The batch's own palette buffer is released; its palettes live at pool rows [first, first + N) from now on. */
int pcl_batch_attach_pool(pcl_batch* b, pcl_pool* pool, int64_t first);
/* Consumer side: declare the attached rows valid (they were written by other batches / other GPUs). */
int pcl_batch_adopt_palettes(pcl_batch* b);
/* The other way to fill a pool: the batch keeps its own palette buffer and this call copies it to pool rows
 * [first, first + N) with the copy engine (cudaMemcpyAsync on the context's copy stream, ordered after the last pose;
 * completes by pcl_sync).  Measured on two B200s (tools/pool_gather_check.py, 1 M remote instances = 2.88 GB): pose +
 * push 6.6 ms (1.5 ms pose, then the copy engine at ~570 GB/s, which can overlap the next tick), pose with the pool
 * attached 8.1 ms (the kernel's 128- and 64-byte store runs reach 355 GB/s over NVLink); 100 k remote instances: 0.72 vs
 * 0.83 ms.  Attach when a second palette copy in HBM or the extra hop is unwelcome, push for the largest crowds. */
int pcl_batch_push_to_pool(pcl_batch* b, pcl_pool* pool, int64_t first);

/* ---- tuning knobs (benchmarks / tests; defaults are chosen by the library) --------------------- */
/* key: "pose_warps_per_block" (0 = auto, 1,2,4,8); "pose_generic" (1 = always use the channel-ordered pose kernel,
 * 0 = use the step-ordered one whenever the skeleton's clips allow it); "pose_sort" (-1 = auto, 0 = never, 1 = always: let
 * the step-ordered kernel visit the instances grouped by the (clip, frame) bucket of their phase - same results, coherent
 * key loads; three extra small launches per pose); "pose_min_blocks" (0 = auto, 2, 3, 4: resident blocks per SM the step-ordered
 * kernel's registers are allocated for - 255 / 168 / 128 registers per thread); "pose_level" (-1 = auto: one-clip crowds small
 * enough to leave SMs idle, 0 = never, 1 = always when applicable: the level-parallel pose kernel, 8 warps per 32 instances);
 * "pose_split" (-1 = auto: one-clip crowds beyond the level kernel's range that leave the SMs fewer than 8 warps each, 0 = never,
 * 1 = always when the skeleton has a balanced cut: two warps per 32 instances walk one half of the skeleton each; same bits);
 * "lbs_variant" (-1 = auto; 0 = lane-per-vertex kernel with
 * coalesced 128-bit stores, 1 = same with TMA bulk stores, 2 = lane-per-instance crowd kernel, 3 = crowd kernel with a
 * warp-specialised TMA store pipeline);
 * "lbs_warps" (0 = auto, 8, 12: resident warps per SM of the crowd kernel); "lbs_vsplit" (0 = auto);
 * "lbs_inst_tile" (0 = auto, <= 16; lane-per-vertex kernels only); "lbs_cta_tile" (0 = auto, 32, 16, 8: instances per
 * CTA of variant 3 - auto takes the largest whose palettes fit in shared memory: 32 up to 82 bones, 16 up to ~165, 8 up to ~340);
 * "export_mode" (0 = auto, 1, 2, 3: see pcl_batch_download_palette3x4_async); "lbs_compact" (-1 = auto: when whole palettes
 * would force variant 3 below 32 instances per CTA, 0 = never, 1 = whenever it fits: variant 3 keeps, per instance, only the
 * palette rows the current block of 1024 vertices references).
 * "pool_bulk" (-1 / 1 = a batch attached to a pool in ANOTHER GPU's memory sends its palette rows as asynchronous bulk copies, 0 = as
 * plain stores from the lanes).
 * PCL_ERR_INVALID for unknown keys. */
int pcl_set_option(pcl_ctx* ctx, const char* key, int32_t value);

#ifdef __cplusplus
}
#endif
#endif /* PICCOLO_B200_H */
