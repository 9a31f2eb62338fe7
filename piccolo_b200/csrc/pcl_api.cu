// C ABI of piccolo_b200 (include/piccolo_b200.h): context, asset repacking into SoA device buffers,
// kernel launches.  No CPU fallback anywhere: every compute entry point launches a CUDA kernel or fails.
// Citations are relative to /root/reference/engine/source/runtime.
#include "../../include/piccolo_b200.h"
#include "pcl_kernels.cuh"

#include <algorithm>
#include <atomic>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

using namespace pcl;

namespace
{
    // pcl_create failures of concurrent threads: the text is per thread (pcl_last_error(NULL) reads the caller's own)
    thread_local std::string g_create_error;

    struct SkelRec
    {
        int32_t              J = 0, pitch = 0, nslot = 0;
        int32_t              split_base[2] = {0, 0}, split_n[2] = {0, 0}; // the two halves' step records in h_steps (pose_simple_kernel<SPLIT>); 0 = none
        std::vector<int32_t> index, parent, order;
        std::vector<StepRec> h_steps;       // host copy: source of the __constant__ upload
        int32_t              n_clips = 0, n_simple = 0;
        int32_t              users = 0;     // batches, pools and skin launches referring to it: a skeleton in use cannot be released
        bool                 released = false;
        StepRec*             d_steps = nullptr;
        int32_t*             d_sched = nullptr; // [n_rounds][POSE_L_WARPS] level-parallel schedule of pose_level_kernel
        int32_t              n_rounds = 0;
        int32_t*             d_masks = nullptr; // [n_masks][J]
        int32_t              n_masks = 0, cap_masks = 0;
    };
    struct ClipHost
    {
        pcl_skel skel;
        void*    d_blob = nullptr;
        ClipRec  rec {};
        int32_t  frames = 0;     // (clip, frame) buckets it occupies
        bool     released = false;
    };
    struct MaskHost
    {
        pcl_skel skel;
        int32_t  local;
    };
    struct MeshRec
    {
        int32_t V = 0, Vp = 0, Vt = 0, max_slot = 0;
        bool    released = false;
        float * d_pos = nullptr, *d_nrm = nullptr, *d_tan = nullptr;
        int4*   d_idx = nullptr;
        float4* d_w   = nullptr;
        VRec*   d_vrec = nullptr;        // lane = instance form: slot-stable influences, see lbs_inst_kernel
        VRec*   d_vrec_cta[3] = {nullptr, nullptr, nullptr}; // the same packed along lbs_cta_kernel's visiting order for S = 1, 2, 4
                                         // vertex streams per warp (built on first use: which S runs depends on the skeleton)
        // compact form of lbs_cta_kernel (see LbsCParams::blk_sc): per block of LBS_C_BLK_VERTS vertices the palette slots it
        // references, and record streams carrying block-local row numbers; blk_mmax < 0 = not analysed yet
        VRec*     d_vrec_blk[3]  = {nullptr, nullptr, nullptr};
        uint16_t* d_blk_slots    = nullptr;
        int32_t*  d_blk_cnt      = nullptr;
        int32_t   blk_mmax = -1, blk_cap = 0, n_blocks = 0;
        std::vector<std::vector<uint16_t>> h_blk_slots;
        std::vector<float>  h_pos, h_nrm, h_tan; // gated, padded host copies the record streams are packed from
        std::vector<int4>   h_idx;
        std::vector<float4> h_w;
    };
} // namespace

// Fixed staging area of pcl_batch_download_palette_static: lives in every GPU that loads this library,
// from load until the process exits, and is shared by every call on that GPU.
static __device__ __align__(256) unsigned char g_static_export[PCL_STATIC_EXPORT_MAX_BYTES];
#define PCL_STATIC_MAX_DEVICES 64
#define PCL_SMALL_EXPORT_BYTES (32u << 10) /* measured: ahead up to ~8 player instances, behind the DMA path from 32 */
#define PCL_FRAME_DIRECT_BYTES (24u << 10) /* pcl_batch_frame: rows written by the kernel into pinned memory up to 8 player instances */
static std::atomic<void*> g_static_export_addr[PCL_STATIC_MAX_DEVICES]; // device address of the array, per device (0 = not looked up yet)

// Which skeleton's step records currently sit in the device's __constant__ bank (pose_simple_kernel).  The bank is
// per device, shared by every context on it.
struct ConstOwner
{
    pcl_ctx* ctx  = nullptr;
    int      skel = -1;
};
static ConstOwner g_const_owner[PCL_STATIC_MAX_DEVICES];
static std::mutex g_const_mu; // contexts on one device may be driven from different threads; the bank is shared

struct pcl_ctx
{
    int          device = 0;
    cudaStream_t compute = nullptr, copy = nullptr;
    cudaEvent_t  ev_pose = nullptr, ev_copy = nullptr, t0 = nullptr, t1 = nullptr;
    bool         copy_pending = false;
    std::string  err;
    int64_t      launches = 0;
    std::vector<SkelRec>  skels;
    std::vector<ClipHost> clips;
    std::vector<MaskHost> masks;
    std::vector<MeshRec>  meshes;
    ClipRec*     d_clips = nullptr;
    int32_t      cap_clips = 0;
    int32_t      opt_pose_wpb = 0, opt_pose_generic = 0, opt_lbs_variant = -1, opt_lbs_vsplit = 0, opt_lbs_inst_tile = 0, opt_lbs_warps = 0, opt_lbs_cta_tile = 0, opt_pose_sort = -1, opt_pose_min_blocks = 0, opt_pose_level = -1, opt_export_mode = 0, opt_lbs_compact = -1, opt_pool_bulk = -1, opt_pose_split = -1;
    void*        h_small = nullptr;  // pinned staging of small synchronous palette read-backs
    bool         debug_sync = false; // PICCOLO_B200_DEBUG_SYNC=1: synchronise after every launch
    int32_t      n_buckets = 0; // (clip, frame) buckets of all uploaded clips, see order_* kernels
    int          max_smem_optin = 0, sm_count = 0;
    std::vector<void*> host_allocs;           // pcl_host_alloc blocks still alive (released by pcl_destroy at the latest)
};

struct pcl_batch
{
    pcl_ctx*  ctx;
    pcl_skel  skel;
    long long N;
    int32_t   K;
    int32_t * d_clip_ids = nullptr, *d_mask_ids = nullptr;
    float *   d_phase = nullptr, *d_weight = nullptr, *d_len = nullptr;
    float*    d_phase_alt = nullptr; // second phase array of the two-half pose form (allocated at its first advancing launch)
    float4*   d_palette = nullptr;
    float*    d_derived = nullptr;
    float4*   d_export  = nullptr;
    long long export_cap = 0; // instances
    float4*   d_export12 = nullptr; // staging of the compact (3x4) read-back, export_mode 3
    long long export12_cap = 0;
    float*      h_stage = nullptr;  // pinned [2][N*K] staging of pcl_batch_set_frame_async
    cudaEvent_t ev_stage = nullptr; // the last staging copy has left h_stage
    // pcl_batch_frame: one pinned, device-visible block = the frame's inputs in device layout ([K][N] x 4 arrays) + its result
    // rows [N][J][12]; the pose kernel reads and writes it in place (no staging copies, no export kernel, no DMA)
    unsigned char* h_io = nullptr;
    cudaEvent_t    ev_frame = nullptr; // the last pcl_batch_frame's work is through
    bool           frame_pending = false;
    pcl_pool* pool = nullptr;     // d_palette points into this pool (not owned) when attached
    int32_t*  d_order = nullptr;  // [N] frame-coherent visiting order of the pose kernel (allocated on first use)
    int32_t*  d_ohist = nullptr;  // [2][ORDER_MAX_BUCKETS] histogram (kept zero between launches) and scatter cursors
    bool      posed = false;
};

struct pcl_pool
{
    pcl_ctx*  ctx;          // the context that created or opened it
    long long capacity;     // instance rows
    int32_t   pitch, J;     // slots per row / bones of the skeleton the rows are laid out for
    float4*   base = nullptr;
    bool      owner = true; // cudaMalloc'ed here; false: mapped from an IPC handle
    int       attached = 0;
};

struct pcl_renderbuf
{
    pcl_ctx*  ctx;
    long long draws;
    float4*   d_mats = nullptr; // [draws][64][1024][4] float4
};

struct pcl_skinbuf
{
    pcl_ctx*  ctx;
    pcl_mesh  mesh;
    long long capacity;
    int32_t   flags, V, Vp;
    float *   d_pos = nullptr, *d_nt = nullptr;
    bool        has_tmaps = false; // tensor maps of the two streams for lbs_cta_kernel's TMA stores, boxes of 32 / 16 / 8 instances
    CUtensorMap tm_pos[3], tm_nt[3];
};

namespace
{
    int fail(pcl_ctx* c, int code, const std::string& msg)
    {
        if (c)
            c->err = msg;
        else
            g_create_error = msg;
        return code;
    }
#define PCL_CK(ctx, call)                                                                                               \
    do                                                                                                                  \
    {                                                                                                                   \
        cudaError_t e__ = (call);                                                                                       \
        if (e__ != cudaSuccess)                                                                                         \
        {                                                                                                               \
            (void)cudaGetLastError(); /* a recoverable error must not resurface at the next launch check */             \
            return fail(ctx, e__ == cudaErrorMemoryAllocation ? PCL_ERR_OOM : PCL_ERR_CUDA,                             \
                        std::string(#call) + ": " + cudaGetErrorString(e__));                                           \
        }                                                                                                               \
    } while (0)

    // After every launch: the launch error, and - with PICCOLO_B200_DEBUG_SYNC=1 in the environment when the context is created -
    // a synchronisation, so that an asynchronous fault is reported by the call that caused it, with the kernel's name.
#define PCL_POST_LAUNCH(ctx, stream, name)                                                                                  \
    do                                                                                                                      \
    {                                                                                                                       \
        cudaError_t e__ = cudaGetLastError();                                                                               \
        if (e__ == cudaSuccess && (ctx)->debug_sync)                                                                        \
            e__ = cudaStreamSynchronize(stream);                                                                            \
        if (e__ != cudaSuccess)                                                                                             \
        {                                                                                                                   \
            (void)cudaGetLastError();                                                                                       \
            return fail(ctx, PCL_ERR_CUDA, std::string(name) + ": " + cudaGetErrorString(e__));                             \
        }                                                                                                                   \
    } while (0)

    // cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (device, kernel, size class) instead of on every launch
    int ensure_smem(pcl_ctx* c, const void* func, size_t bytes);

    // Host -> device copies and memsets go through the context's COMPUTE stream, never the legacy default stream: the
    // context's streams are non-blocking, so they are not ordered against legacy-stream work, and cudaMemcpy from pageable
    // memory returns once the data is STAGED, not once it has reached the device - a kernel launched right after an upload
    // (the record streams lbs_cta_kernel builds on first use) could read the buffer before its contents had arrived.
    cudaError_t copy_h2d(pcl_ctx* c, void* dst, const void* src, size_t bytes)
    {
        cudaError_t e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->compute);
        return e != cudaSuccess ? e : cudaStreamSynchronize(c->compute); // the host buffer may be freed when this returns
    }

    template<typename T>
    int upload(pcl_ctx* c, T** dptr, const std::vector<T>& h)
    {
        PCL_CK(c, cudaMalloc((void**)dptr, std::max<size_t>(h.size(), 1) * sizeof(T)));
        if (!h.empty())
        {
            const cudaError_t e = copy_h2d(c, *dptr, h.data(), h.size() * sizeof(T));
            if (e != cudaSuccess) // nothing of a failed upload stays behind
            {
                cudaFree(*dptr);
                *dptr = nullptr;
                (void)cudaGetLastError();
                return fail(c, PCL_ERR_CUDA, std::string("upload: ") + cudaGetErrorString(e));
            }
        }
        return PCL_OK;
    }

    // Slot plan for one processing order: which bones must park their derived TRS in shared memory
    // (those with a child that is not the very next step) and for how long.
    int plan_slots(const std::vector<int32_t>& parent, const std::vector<int32_t>& order, std::vector<int32_t>* psrc,
                   std::vector<int32_t>* store)
    {
        const int            J = (int)parent.size();
        std::vector<int32_t> step_of(J), last_child_step(J, -1), slot_of(J, -1);
        std::vector<char>    needs(J, 0);
        for (int s = 0; s < J; s++)
            step_of[order[s]] = s;
        for (int b = 0; b < J; b++)
            if (parent[b] >= 0)
            {
                last_child_step[parent[b]] = std::max(last_child_step[parent[b]], step_of[b]);
                if (step_of[b] != step_of[parent[b]] + 1)
                    needs[parent[b]] = 1;
            }
        std::vector<int32_t> free_list;
        int                  nslot = 0;
        psrc->assign(J, -1), store->assign(J, -1);
        std::vector<std::vector<int32_t>> release(J + 1);
        for (int s = 0; s < J; s++)
        {
            for (int32_t sl : release[s])
                free_list.push_back(sl);
            const int b = order[s], p = parent[b];
            if (p >= 0)
                (*psrc)[s] = (step_of[p] == s - 1) ? -2 : slot_of[p];
            if (needs[b])
            {
                int sl;
                if (!free_list.empty())
                    sl = free_list.back(), free_list.pop_back();
                else
                    sl = nslot++;
                slot_of[b]  = sl;
                (*store)[s] = sl;
                release[last_child_step[b] + 1].push_back(sl);
            }
        }
        return nslot;
    }

    void dfs_order(const std::vector<int32_t>& parent, std::vector<int32_t>* order)
    {
        const int                         J = (int)parent.size();
        std::vector<std::vector<int32_t>> kids(J);
        std::vector<int32_t>              roots, stack;
        for (int b = 0; b < J; b++)
            (parent[b] >= 0 ? kids[parent[b]] : roots).push_back(b);
        order->clear();
        for (auto r = roots.rbegin(); r != roots.rend(); ++r)
            stack.push_back(*r);
        while (!stack.empty())
        {
            int b = stack.back();
            stack.pop_back();
            order->push_back(b);
            for (auto k = kids[b].rbegin(); k != kids[b].rend(); ++k)
                stack.push_back(*k);
        }
    }

    // Slot-stable packing of vertex v (or padding when v >= V): influences whose bone is already resident keep their
    // slot, newcomers take free slots, unused slots keep the resident bone with weight 0; the mask says which slots
    // differ from the previously VISITED record.
    void pack_vrec(const MeshRec& m, VRec& r, size_t v, int cur[4], const VRec* prev)
    {
        auto  slot_of = [](const VRec& x, int sl) { return ((sl < 2 ? (x.i01 >> (16 * sl)) : (x.i23 >> (16 * (sl - 2)))) & 0xffffu); };
        float slot_w[4] = {0.f, 0.f, 0.f, 0.f};
        if (v < (size_t)m.V)
        {
            const int   src_i[4] = {m.h_idx[v].x, m.h_idx[v].y, m.h_idx[v].z, m.h_idx[v].w};
            const float src_w[4] = {m.h_w[v].x, m.h_w[v].y, m.h_w[v].z, m.h_w[v].w};
            bool        taken[4] = {false, false, false, false}, placed[4] = {false, false, false, false};
            for (int k = 0; k < 4; k++)
                if (src_w[k] != 0.f)
                    for (int sl = 0; sl < 4; sl++)
                        if (!taken[sl] && cur[sl] == src_i[k])
                        {
                            taken[sl] = true, placed[k] = true, slot_w[sl] = src_w[k];
                            break;
                        }
            for (int k = 0; k < 4; k++)
                if (src_w[k] != 0.f && !placed[k])
                    for (int sl = 0; sl < 4; sl++)
                        if (!taken[sl])
                        {
                            taken[sl] = true, cur[sl] = src_i[k], slot_w[sl] = src_w[k];
                            break;
                        }
            r.px = m.h_pos[3 * v], r.py = m.h_pos[3 * v + 1], r.pz = m.h_pos[3 * v + 2];
            r.nx = m.h_nrm[3 * v], r.ny = m.h_nrm[3 * v + 1], r.nz = m.h_nrm[3 * v + 2];
            r.tx = m.h_tan[3 * v], r.ty = m.h_tan[3 * v + 1], r.tz = m.h_tan[3 * v + 2];
        }
        r.w0 = slot_w[0], r.w1 = slot_w[1], r.w2 = slot_w[2], r.w3 = slot_w[3];
        r.i01  = (uint32_t)cur[0] * 3u | (((uint32_t)cur[1] * 3u) << 16);
        r.i23  = (uint32_t)cur[2] * 3u | (((uint32_t)cur[3] * 3u) << 16);
        r.mask = 0;
        for (int sl = 0; sl < 4; sl++)
            if (!prev || slot_of(r, sl) != slot_of(*prev, sl))
                r.mask |= 1u << sl;
    }

    // lbs_cta_kernel's record stream for T = 32 / 16 / 8 instances per tile (S = 1 / 2 / 4 vertex streams per warp):
    // 8 S interleaved visiting orders, sequence vw = vertices scv * s + 4 vw + {0..3}, s = 0, 1, ...
    int ensure_vrec_cta(pcl_ctx* c, MeshRec& m, int T)
    {
        const int si = T == 32 ? 0 : (T == 16 ? 1 : 2);
        if (m.d_vrec_cta[si])
            return PCL_OK;
        const size_t      scv  = (size_t)lbs_c_scv(T);
        const size_t      n_sc = ((size_t)m.Vp + scv - 1) / scv;
        const size_t      pad  = (size_t)lbs_c_rsc(T); // a record stage may overrun the last superchunk
        std::vector<VRec> hvc((n_sc + pad) * scv);
        std::memset(hvc.data(), 0, hvc.size() * sizeof(VRec));
        for (size_t vw = 0; vw < scv / LBS_C_VPW; vw++)
        {
            int         cw[4] = {0, 0, 0, 0};
            const VRec* prev  = nullptr;
            for (size_t sc = 0; sc < n_sc + pad; sc++)
                for (int j = 0; j < LBS_C_VPW; j++)
                {
                    const size_t v = sc * scv + vw * LBS_C_VPW + j;
                    pack_vrec(m, hvc[v], v, cw, prev);
                    prev = &hvc[v];
                }
        }
        return upload(c, &m.d_vrec_cta[si], hvc);
    }

    // Which palette slots does each block of LBS_C_BLK_VERTS consecutive vertices reference?  (Once per mesh; decides whether the
    // compact form of lbs_cta_kernel pays: a coherent 256-bone character touches a few dozen bones per 1024 vertices.)
    int analyse_blocks(pcl_ctx* c, MeshRec& m)
    {
        if (m.blk_mmax >= 0)
            return PCL_OK;
        const int nb = (m.Vp + LBS_C_BLK_VERTS - 1) / LBS_C_BLK_VERTS + 1; // + 1: a record stage may overrun into a padding block
        m.h_blk_slots.assign(nb, {});
        int mmax = 1;
        for (int b = 0; b < nb; b++)
        {
            std::vector<char> used((size_t)m.max_slot + 1, 0);
            for (size_t v = (size_t)b * LBS_C_BLK_VERTS; v < (size_t)(b + 1) * LBS_C_BLK_VERTS && v < (size_t)m.V; v++)
            {
                const int   id[4] = {m.h_idx[v].x, m.h_idx[v].y, m.h_idx[v].z, m.h_idx[v].w};
                const float w[4]  = {m.h_w[v].x, m.h_w[v].y, m.h_w[v].z, m.h_w[v].w};
                for (int k = 0; k < 4; k++)
                    if (w[k] != 0.f)
                        used[id[k]] = 1;
            }
            used[0] = 1; // slot 0 (the identity) is what gated and padding influences name
            for (int sl = 0; sl <= m.max_slot; sl++)
                if (used[sl])
                    m.h_blk_slots[b].push_back((uint16_t)sl);
            mmax = std::max(mmax, (int)m.h_blk_slots[b].size());
        }
        m.n_blocks = nb, m.blk_cap = mmax;
        std::vector<uint16_t> flat((size_t)nb * mmax, 0);
        std::vector<int32_t>  cnt(nb);
        for (int b = 0; b < nb; b++)
        {
            cnt[b] = (int32_t)m.h_blk_slots[b].size();
            std::copy(m.h_blk_slots[b].begin(), m.h_blk_slots[b].end(), flat.begin() + (size_t)b * mmax);
        }
        int rc;
        if ((rc = upload(c, &m.d_blk_slots, flat)) || (rc = upload(c, &m.d_blk_cnt, cnt)))
            return rc;
        m.blk_mmax = mmax;
        return PCL_OK;
    }

    // the record stream of ensure_vrec_cta with block-local row numbers: at every block start each (warp, stream) sequence starts
    // cold (all four slots reload, resident bones forgotten)
    int ensure_vrec_cta_compact(pcl_ctx* c, MeshRec& m, int T)
    {
        const int si = T == 32 ? 0 : (T == 16 ? 1 : 2);
        if (m.d_vrec_blk[si])
            return PCL_OK;
        const size_t      scv  = (size_t)lbs_c_scv(T);
        const size_t      bsc  = LBS_C_BLK_VERTS / scv; // superchunks per block
        const size_t      n_sc = ((size_t)m.Vp + scv - 1) / scv;
        const size_t      pad  = (size_t)lbs_c_rsc(T);
        std::vector<VRec> hvc((n_sc + pad) * scv);
        std::memset(hvc.data(), 0, hvc.size() * sizeof(VRec));
        for (size_t vw = 0; vw < scv / LBS_C_VPW; vw++)
        {
            int  cw[4] = {0, 0, 0, 0};
            VRec last {}; // the previously visited record, in GLOBAL slot numbers (what the reload masks are derived from)
            bool have_last = false;
            for (size_t sc = 0; sc < n_sc + pad; sc++)
            {
                const size_t blk = std::min<size_t>(sc / bsc, (size_t)m.n_blocks - 1);
                if (sc % bsc == 0)
                    cw[0] = cw[1] = cw[2] = cw[3] = 0, have_last = false; // slot 0 is in every block's list
                const std::vector<uint16_t>& list = m.h_blk_slots[blk];
                auto loc = [&list](uint32_t off3) -> uint32_t {
                    const uint16_t slot = (uint16_t)(off3 / 3u);
                    auto           it   = std::lower_bound(list.begin(), list.end(), slot);
                    return (it != list.end() && *it == slot) ? (uint32_t)(it - list.begin()) * 3u : 0u;
                };
                for (int j = 0; j < LBS_C_VPW; j++)
                {
                    const size_t v = sc * scv + vw * LBS_C_VPW + j;
                    VRec         g {};
                    pack_vrec(m, g, v, cw, have_last ? &last : nullptr);
                    last = g, have_last = true;
                    // global slot numbers -> rows of this block's list
                    g.i01  = loc(g.i01 & 0xffffu) | (loc(g.i01 >> 16) << 16);
                    g.i23  = loc(g.i23 & 0xffffu) | (loc(g.i23 >> 16) << 16);
                    hvc[v] = g;
                }
            }
        }
        return upload(c, &m.d_vrec_blk[si], hvc);
    }

    bool valid_skel(pcl_ctx* c, pcl_skel s) { return c && s >= 0 && s < (int)c->skels.size() && !c->skels[s].released; }

    // 2-D tensor map over a skinned stream [capacity][row_floats], box = 32 rows x 32 floats (128 B), 128-byte swizzle.
    // The encoder lives in the driver; it is fetched through the runtime so that the library has no link-time
    // dependency on libcuda (it must load on machines without a GPU).
    bool make_stream_tmap(CUtensorMap* tm, float* base, long long capacity, long long row_floats, unsigned box_rows)
    {
        typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                      const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
        static std::atomic<encode_fn> encode_slot {nullptr}; // contexts may be driven from different threads
        encode_fn                     encode = encode_slot.load(std::memory_order_acquire);
        if (!encode)
        {
            void*                           fn = nullptr;
            cudaDriverEntryPointQueryResult qr;
            if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) != cudaSuccess || !fn)
            {
                (void)cudaGetLastError();
                return false;
            }
            encode = (encode_fn)fn;
            encode_slot.store(encode, std::memory_order_release);
        }
        if (row_floats >= (1LL << 32) || capacity >= (1LL << 32))
            return false;
        const cuuint64_t dims[2]    = {(cuuint64_t)row_floats, (cuuint64_t)capacity};
        const cuuint64_t strides[1] = {(cuuint64_t)row_floats * 4};
        const cuuint32_t box[2]     = {32, box_rows};
        const cuuint32_t estr[2]    = {1, 1};
        return encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                      CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
    }

    // The attribute belongs to the (device, kernel) pair, not to a context: two contexts on one device share it, and setting it
    // is not monotone (a second context asking for less would shrink what the first one relies on).  One process-wide record.
    int ensure_smem(pcl_ctx* c, const void* func, size_t bytes)
    {
        static std::mutex                                        mu;
        static std::map<std::pair<int, const void*>, size_t>     have_all;
        std::lock_guard<std::mutex>                              lk(mu);
        size_t& have = have_all[std::make_pair(c->device, func)];
        if (bytes > have)
        {
            PCL_CK(c, cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
            have = bytes;
        }
        return PCL_OK;
    }
} // namespace

extern "C"
{
    int pcl_version(void) { return 100; }

    int pcl_device_count(void)
    {
        int n = 0;
        if (cudaGetDeviceCount(&n) != cudaSuccess)
        {
            (void)cudaGetLastError();
            return 0;
        }
        return n;
    }

    int pcl_create(int device, pcl_ctx** out)
    {
        if (!out)
            return fail(nullptr, PCL_ERR_INVALID, "pcl_create: out == NULL");
        *out  = nullptr;
        int n = pcl_device_count();
        if (n <= 0)
            return fail(nullptr, PCL_ERR_NO_DEVICE, "pcl_create: no CUDA device (this engine has no CPU path)");
        if (device < 0 || device >= n)
            return fail(nullptr, PCL_ERR_INVALID, "pcl_create: device index out of range");
        PCL_CK(nullptr, cudaSetDevice(device));
        cudaDeviceProp prop;
        PCL_CK(nullptr, cudaGetDeviceProperties(&prop, device));
        if (prop.major < 10)
            return fail(nullptr, PCL_ERR_UNSUPPORTED, "pcl_create: kernels are built for sm_100a (Blackwell) only");
        pcl_ctx* c        = new pcl_ctx();
        c->device         = device;
        c->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
        c->sm_count       = prop.multiProcessorCount;
        const char* dbg   = std::getenv("PICCOLO_B200_DEBUG_SYNC");
        c->debug_sync     = dbg && dbg[0] == '1';
        cudaError_t e     = cudaStreamCreateWithFlags(&c->compute, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->copy, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_pose, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_copy, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreate(&c->t0);
        if (e == cudaSuccess) e = cudaEventCreate(&c->t1);
        if (e != cudaSuccess)
        {
            if (c->t1) cudaEventDestroy(c->t1);
            if (c->t0) cudaEventDestroy(c->t0);
            if (c->ev_copy) cudaEventDestroy(c->ev_copy);
            if (c->ev_pose) cudaEventDestroy(c->ev_pose);
            if (c->copy) cudaStreamDestroy(c->copy);
            if (c->compute) cudaStreamDestroy(c->compute);
            delete c;
            return fail(nullptr, PCL_ERR_CUDA, std::string("pcl_create: ") + cudaGetErrorString(e));
        }
        *out = c;
        return PCL_OK;
    }

    int pcl_destroy(pcl_ctx* c)
    {
        if (!c)
            return PCL_ERR_INVALID;
        cudaSetDevice(c->device);
        cudaDeviceSynchronize();
        {
            std::lock_guard<std::mutex> lk(g_const_mu);
            if (c->device < PCL_STATIC_MAX_DEVICES && g_const_owner[c->device].ctx == c)
                g_const_owner[c->device] = ConstOwner();
        }
        for (auto& s : c->skels)
            cudaFree(s.d_steps), cudaFree(s.d_masks), cudaFree(s.d_sched);
        for (auto& cl : c->clips)
            cudaFree(cl.d_blob);
        for (auto& m : c->meshes)
            cudaFree(m.d_pos), cudaFree(m.d_nrm), cudaFree(m.d_tan), cudaFree(m.d_idx), cudaFree(m.d_w), cudaFree(m.d_vrec), cudaFree(m.d_vrec_cta[0]), cudaFree(m.d_vrec_cta[1]), cudaFree(m.d_vrec_cta[2]), cudaFree(m.d_vrec_blk[0]), cudaFree(m.d_vrec_blk[1]), cudaFree(m.d_vrec_blk[2]), cudaFree(m.d_blk_slots), cudaFree(m.d_blk_cnt);
        cudaFree(c->d_clips);
        for (void* p : c->host_allocs)
            cudaFreeHost(p);
        if (c->h_small)
            cudaFreeHost(c->h_small);
        cudaEventDestroy(c->ev_pose), cudaEventDestroy(c->ev_copy), cudaEventDestroy(c->t0), cudaEventDestroy(c->t1);
        cudaStreamDestroy(c->compute), cudaStreamDestroy(c->copy);
        delete c;
        return PCL_OK;
    }

    const char* pcl_last_error(const pcl_ctx* c) { return c ? c->err.c_str() : g_create_error.c_str(); }

    int pcl_sync(pcl_ctx* c)
    {
        if (!c)
            return PCL_ERR_INVALID;
        PCL_CK(c, cudaSetDevice(c->device));
        PCL_CK(c, cudaStreamSynchronize(c->copy));
        PCL_CK(c, cudaStreamSynchronize(c->compute));
        return PCL_OK;
    }

    int64_t pcl_launch_count(const pcl_ctx* c) { return c ? c->launches : 0; }

    int pcl_host_alloc(pcl_ctx* c, size_t bytes, void** out)
    {
        if (!c || !out)
            return PCL_ERR_INVALID;
        PCL_CK(c, cudaSetDevice(c->device));
        PCL_CK(c, cudaHostAlloc(out, bytes, cudaHostAllocDefault));
        c->host_allocs.push_back(*out);
        return PCL_OK;
    }
    int pcl_host_free(pcl_ctx* c, void* p)
    {
        if (!c)
            return PCL_ERR_INVALID;
        auto it = std::find(c->host_allocs.begin(), c->host_allocs.end(), p);
        if (it == c->host_allocs.end())
            return fail(c, PCL_ERR_INVALID, "pcl_host_free: pointer was not returned by pcl_host_alloc of this context");
        c->host_allocs.erase(it);
        PCL_CK(c, cudaFreeHost(p));
        return PCL_OK;
    }

    int pcl_timer_begin(pcl_ctx* c)
    {
        if (!c)
            return PCL_ERR_INVALID;
        PCL_CK(c, cudaSetDevice(c->device));
        PCL_CK(c, cudaEventRecord(c->t0, c->compute));
        return PCL_OK;
    }
    int pcl_timer_end(pcl_ctx* c, float* ms)
    {
        if (!c || !ms)
            return PCL_ERR_INVALID;
        PCL_CK(c, cudaSetDevice(c->device));
        PCL_CK(c, cudaEventRecord(c->ev_copy, c->copy)); // join the copy stream into the timed region
        PCL_CK(c, cudaStreamWaitEvent(c->compute, c->ev_copy, 0));
        PCL_CK(c, cudaEventRecord(c->t1, c->compute));
        PCL_CK(c, cudaEventSynchronize(c->t1));
        PCL_CK(c, cudaEventElapsedTime(ms, c->t0, c->t1));
        return PCL_OK;
    }

    int pcl_set_option(pcl_ctx* c, const char* key, int32_t v)
    {
        if (!c || !key)
            return PCL_ERR_INVALID;
        std::string k(key);
        if (k == "pose_warps_per_block" && (v == 0 || v == 1 || v == 2 || v == 4 || v == 8))
            c->opt_pose_wpb = v;
        else if (k == "pose_generic" && (v == 0 || v == 1))
            c->opt_pose_generic = v;
        else if (k == "pose_sort" && v >= -1 && v <= 1)
            c->opt_pose_sort = v;
        else if (k == "pose_min_blocks" && (v == 0 || (v >= 2 && v <= 4)))
            c->opt_pose_min_blocks = v;
        else if (k == "pose_level" && v >= -1 && v <= 1)
            c->opt_pose_level = v;
        else if (k == "pose_split" && v >= -1 && v <= 1)
            c->opt_pose_split = v;
        else if (k == "lbs_variant" && v >= -1 && v <= 3)
            c->opt_lbs_variant = v;
        else if (k == "lbs_warps" && (v == 0 || v == 8 || v == 12))
            c->opt_lbs_warps = v;
        else if (k == "lbs_vsplit" && v >= 0)
            c->opt_lbs_vsplit = v;
        else if (k == "lbs_inst_tile" && v >= 0 && v <= 16)
            c->opt_lbs_inst_tile = v;
        else if (k == "lbs_cta_tile" && (v == 0 || v == 8 || v == 16 || v == 32))
            c->opt_lbs_cta_tile = v;
        else if (k == "export_mode" && v >= 0 && v <= 3)
            c->opt_export_mode = v;
        else if (k == "lbs_compact" && v >= -1 && v <= 1)
            c->opt_lbs_compact = v;
        else if (k == "pool_bulk" && v >= -1 && v <= 1)
            c->opt_pool_bulk = v;
        else
            return fail(c, PCL_ERR_INVALID, "pcl_set_option: unknown key or bad value: " + k);
        return PCL_OK;
    }

    // ------------------------------------------------------------------------------------------
    // assets
    // ------------------------------------------------------------------------------------------
    int pcl_upload_skeleton(pcl_ctx* c, int32_t J, const int32_t* index, const int32_t* parent_index, const float* bind_pos,
                            const float* bind_rot, const float* bind_scl, const float* inv_bind, int32_t is_flat,
                            int32_t in_topo, pcl_skel* out)
    {
        if (!c || !out || J < 1 || !index || !parent_index || !bind_pos || !bind_rot || !bind_scl || !inv_bind)
            return fail(c, PCL_ERR_INVALID, "pcl_upload_skeleton: bad argument");
        if (!is_flat || !in_topo)
            return fail(c, PCL_ERR_UNSUPPORTED,
                        "pcl_upload_skeleton: skeleton must be flat and in topological order (the reference builds nothing "
                        "otherwise, skeleton.cpp:24-28)");
        PCL_CK(c, cudaSetDevice(c->device));
        SkelRec s;
        s.J = J;
        s.index.assign(index, index + J);
        s.parent.resize(J);
        for (int i = 0; i < J; i++)
        {
            int key = parent_index[i];
            if (key < 0)
                return fail(c, PCL_ERR_INVALID, "pcl_upload_skeleton: negative parent_index");
            s.parent[i] = (key == INT_MAX || key >= i) ? -1 : key; // find_by_index, utilities.cpp:7-27
            const float* m = inv_bind + 16 * (size_t)i;
            if (!(m[12] == 0.f && m[13] == 0.f && m[14] == 0.f && m[15] == 1.f))
                return fail(c, PCL_ERR_UNSUPPORTED, "pcl_upload_skeleton: inverse bind matrix is not affine (row 3 != 0 0 0 1)");
        }
        // processing order: bones_map order unless that needs many parked parents, then DFS pre-order
        std::vector<int32_t> order(J), psrc, store;
        for (int i = 0; i < J; i++)
            order[i] = i;
        int nslot = plan_slots(s.parent, order, &psrc, &store);
        if (nslot > 12)
        {
            std::vector<int32_t> o2, p2, s2;
            dfs_order(s.parent, &o2);
            int n2 = plan_slots(s.parent, o2, &p2, &s2);
            if (n2 < nslot)
                order = o2, psrc = p2, store = s2, nslot = n2;
        }
        if ((size_t)pose_smem_floats_per_warp(nslot) * 4 > (size_t)c->max_smem_optin)
            return fail(c, PCL_ERR_UNSUPPORTED, "pcl_upload_skeleton: hierarchy needs more parked parents than shared memory holds");
        s.nslot = nslot;
        s.order = order;
        s.pitch = (J + 1 + 3) & ~3; // slots per instance, a multiple of 4 => every instance starts on a 64-byte boundary (the pose kernel flushes 4 slots = 192 bytes at a time)
        std::vector<int32_t> step_of(J);
        for (int st = 0; st < J; st++)
            step_of[order[st]] = st;
        std::vector<StepRec> steps(J);
        for (int st = 0; st < J; st++)
        {
            const int b  = order[st];
            StepRec&  r  = steps[st];
            r.joint = b, r.psrc = psrc[st], r.store = store[st];
            r.pstep = s.parent[b] < 0 ? -1 : step_of[s.parent[b]];
            r.bA = make_float4(bind_pos[3 * b], bind_pos[3 * b + 1], bind_scl[3 * b], bind_scl[3 * b + 1]);
            r.bC = make_float4(1.f, bind_scl[3 * b + 2], bind_pos[3 * b + 2], 0.f);
            // Node::setOrientation, node.cpp:71-85: NaN -> identity, else normalise (factor = 1/length; q*factor)
            float w = bind_rot[4 * b], x = bind_rot[4 * b + 1], y = bind_rot[4 * b + 2], z = bind_rot[4 * b + 3];
            if (std::isnan(w) || std::isnan(x) || std::isnan(y) || std::isnan(z))
                w = 1.f, x = y = z = 0.f;
            else
            {
                float factor = 1.0f / std::sqrt(w * w + x * x + y * y + z * z);
                w = w * factor, x = x * factor, y = y * factor, z = z * factor;
            }
            r.q01 = make_float4(w, x, y, z);
            r.q23 = make_float4(-x, w, -z, y);
            const float* m = inv_bind + 16 * (size_t)b;
            r.ib0 = make_float4(m[0], m[1], m[2], m[3]);
            r.ib1 = make_float4(m[4], m[5], m[6], m[7]);
            r.ib2 = make_float4(m[8], m[9], m[10], m[11]);
        }
        // Two halves for small crowds (pose_simple_kernel<SPLIT>): a set T of whole subtrees is chosen so that {T + its ancestors}
        // and {everything outside T} are as equal as they get (greedy: the best single subtree first, then further subtrees
        // while the larger half shrinks); each half is an ancestor-closed bone list in processing order with a parking plan of
        // its own.  Kept only when it shortens the walk by a quarter and all records fit the constant bank.
        // (Player: T = right clavicle's + right thigh's subtrees: 31 + 33 steps instead of 58.)
        {
            std::vector<std::vector<int32_t>> kids(J);
            for (int bn = 0; bn < J; bn++)
                if (s.parent[bn] >= 0)
                    kids[s.parent[bn]].push_back(bn);
            std::vector<char> in_sub(J, 0), anc(J, 0);
            auto              halves = [&](const std::vector<char>& T, int* na, int* nb) {
                std::vector<char> up(J, 0);
                int               nt = 0, nu = 0;
                for (int bn = 0; bn < J; bn++)
                    if (T[bn])
                    {
                        nt++;
                        for (int q = s.parent[bn]; q >= 0 && !T[q] && !up[q]; q = s.parent[q])
                            up[q] = 1, nu++;
                    }
                *na = nt + nu, *nb = J - nt;
                return std::max(*na, *nb);
            };
            int cost = J;
            for (bool better = true; better && J >= 16;)
            {
                better = false;
                std::vector<char> bestT;
                for (int v = 0; v < J; v++)
                {
                    if (s.parent[v] < 0 || in_sub[v])
                        continue;
                    std::vector<char>    T = in_sub;
                    std::vector<int32_t> stack(1, v);
                    while (!stack.empty())
                    {
                        const int x = stack.back();
                        stack.pop_back();
                        T[x] = 1;
                        for (int k : kids[x])
                            stack.push_back(k);
                    }
                    int na, nb;
                    if (halves(T, &na, &nb) < cost)
                        cost = std::max(na, nb), bestT = T, better = true;
                }
                if (better)
                    in_sub = bestT;
            }
            int na = 0, nb = 0;
            if (cost < J)
                halves(in_sub, &na, &nb);
            if (cost < J && cost * 4 <= J * 3 && (size_t)J + (size_t)na + (size_t)nb + 2 * POSE_G <= (size_t)POSE_CMAX)
            {
                for (int bn = 0; bn < J; bn++)
                    if (in_sub[bn])
                        for (int q = s.parent[bn]; q >= 0; q = s.parent[q])
                            if (!in_sub[q])
                                anc[q] = 1;
                int  nslot_h = nslot;
                auto half    = [&](int which) {
                    std::vector<int32_t> bones, local(J, -1), par_l, ord_l, ps, stv;
                    for (int st = 0; st < J; st++)
                    {
                        const int b = order[st];
                        if (which == 0 ? (in_sub[b] || anc[b]) : !in_sub[b])
                            local[b] = (int32_t)bones.size(), bones.push_back(b);
                    }
                    for (size_t i = 0; i < bones.size(); i++)
                        par_l.push_back(s.parent[bones[i]] < 0 ? -1 : local[s.parent[bones[i]]]), ord_l.push_back((int32_t)i);
                    nslot_h             = std::max(nslot_h, plan_slots(par_l, ord_l, &ps, &stv));
                    s.split_base[which] = (int32_t)steps.size(), s.split_n[which] = (int32_t)bones.size();
                    for (size_t i = 0; i < bones.size(); i++)
                    {
                        StepRec r = steps[step_of[bones[i]]];
                        r.psrc = ps[i], r.store = stv[i], r.pstep = step_of[bones[i]];
                        steps.push_back(r);
                    }
                    for (int i = 0; i < POSE_G; i++) // the key prefetch runs up to two steps past the end of a whole group
                        steps.push_back(steps.back());
                };
                half(0), half(1);
                if ((size_t)pose_smem_floats_per_warp(nslot_h) * 4 > (size_t)c->max_smem_optin)
                    steps.resize(J), s.split_n[0] = s.split_n[1] = 0;
                else
                    s.nslot = nslot = nslot_h;
            }
        }
        int rc = upload(c, &s.d_steps, steps);
        if (rc != PCL_OK)
            return rc;
        s.h_steps = steps;
        {
            // pose_level_kernel's schedule: level by level (depth in the hierarchy), POSE_L_WARPS bones per round
            std::vector<int32_t> depth(J, 0);
            int                  levels = 0;
            for (int st = 0; st < J; st++) // parents precede children in every processing order
            {
                const int b = order[st];
                depth[b]    = s.parent[b] < 0 ? 0 : depth[s.parent[b]] + 1;
                levels      = std::max(levels, depth[b] + 1);
            }
            std::vector<int32_t> sched;
            for (int l = 0; l < levels; l++)
            {
                int in_round = 0;
                for (int st = 0; st < J; st++)
                    if (depth[order[st]] == l)
                    {
                        sched.push_back(st);
                        if (++in_round == POSE_L_WARPS)
                            in_round = 0;
                    }
                while (in_round != 0 && in_round++ < POSE_L_WARPS)
                    sched.push_back(-1);
            }
            s.n_rounds = (int32_t)(sched.size() / POSE_L_WARPS);
            if ((rc = upload(c, &s.d_sched, sched)) != PCL_OK)
            {
                cudaFree(s.d_steps), cudaFree(s.d_sched);
                return rc;
            }
        }
        c->skels.push_back(std::move(s));
        *out = (pcl_skel)c->skels.size() - 1;
        return PCL_OK;
    }

    int pcl_skeleton_bone_count(pcl_ctx* c, pcl_skel s, int32_t* J)
    {
        if (!valid_skel(c, s) || !J)
            return fail(c, PCL_ERR_INVALID, "pcl_skeleton_bone_count: bad handle");
        *J = c->skels[s].J;
        return PCL_OK;
    }
    int pcl_skeleton_indices(pcl_ctx* c, pcl_skel s, int32_t* out)
    {
        if (!valid_skel(c, s) || !out)
            return fail(c, PCL_ERR_INVALID, "pcl_skeleton_indices: bad handle");
        for (int i = 0; i < c->skels[s].J; i++)
            out[i] = c->skels[s].index[i] + 1; // getID() + 1, skeleton.cpp:127
        return PCL_OK;
    }
    int pcl_skeleton_parents(pcl_ctx* c, pcl_skel s, int32_t* out)
    {
        if (!valid_skel(c, s) || !out)
            return fail(c, PCL_ERR_INVALID, "pcl_skeleton_parents: bad handle");
        std::memcpy(out, c->skels[s].parent.data(), sizeof(int32_t) * c->skels[s].J);
        return PCL_OK;
    }

    int pcl_upload_clip(pcl_ctx* c, pcl_skel skel, int32_t total_frame, int32_t node_count, int32_t C, const int32_t* key_counts,
                        const float* pos, const float* rot, const float* scl, int32_t convert_len, const int32_t* convert,
                        pcl_clip* out)
    {
        if (!valid_skel(c, skel) || !out || C < 0 || node_count < 0 || convert_len < 0 || (C > 0 && !key_counts))
            return fail(c, PCL_ERR_INVALID, "pcl_upload_clip: bad argument");
        if (node_count > C)
            return fail(c, PCL_ERR_INVALID, "pcl_upload_clip: node_count exceeds the number of channels");
        if (total_frame < 1)
            return fail(c, PCL_ERR_INVALID, "pcl_upload_clip: total_frame < 1");
        if (convert_len > 0 && !convert)
            return fail(c, PCL_ERR_INVALID, "pcl_upload_clip: convert == NULL");
        {
            long long np = 0, nr = 0, ns = 0;
            for (int n = 0; n < C; n++)
                np += std::max(key_counts[3 * n], 0), nr += std::max(key_counts[3 * n + 1], 0), ns += std::max(key_counts[3 * n + 2], 0);
            if ((np > 0 && !pos) || (nr > 0 && !rot) || (ns > 0 && !scl))
                return fail(c, PCL_ERR_INVALID, "pcl_upload_clip: a key array is NULL although its tracks have keys");
        }
        PCL_CK(c, cudaSetDevice(c->device));
        const int J   = c->skels[skel].J;
        const int NCH = std::min(node_count, convert_len); // skeleton.cpp:57-59
        // track offsets
        std::vector<size_t> op(C + 1, 0), orr(C + 1, 0), os(C + 1, 0);
        int                 F = 1;
        for (int n = 0; n < C; n++)
        {
            const int np = key_counts[3 * n], nr = key_counts[3 * n + 1], ns = key_counts[3 * n + 2];
            if (np < 0 || nr < 0 || ns < 0)
                return fail(c, PCL_ERR_INVALID, "pcl_upload_clip: negative key count");
            op[n + 1] = op[n] + np, orr[n + 1] = orr[n] + nr, os[n + 1] = os[n] + ns;
            if (n < NCH)
                F = std::max(F, std::max(np, std::max(nr, ns)));
        }
        std::vector<int32_t> first(J, -1), next(std::max(C, 1), -1), hic(std::max(C, 1), 0), last(J, -1);
        int                  run = INT_MAX;
        for (int n = 0; n < NCH; n++)
        {
            const int bone = convert[n];
            if (bone == -1)
                continue; // size_t(-1) == numeric_limits<size_t>::max(): skipped before the clamp, skeleton.cpp:69-73
            if (bone < 0 || bone >= J)
                return fail(c, PCL_ERR_INVALID, "pcl_upload_clip: convert[] entry outside [0,J) (undefined behaviour in the reference)");
            const int shortest = std::min(key_counts[3 * n], std::min(key_counts[3 * n + 1], key_counts[3 * n + 2]));
            if (shortest < 1)
                return fail(c, PCL_ERR_INVALID, "pcl_upload_clip: a played channel has an empty track (the reference would index [-1])");
            run    = std::min(run, shortest - 1); // carry-over clamp, skeleton.cpp:75-87
            hic[n] = run;
            if (first[bone] < 0)
                first[bone] = n;
            else
                next[last[bone]] = n;
            last[bone] = n;
        }
        std::vector<float4> keys((size_t)std::max(C, 1) * F * 3, make_float4(0, 0, 0, 0));
        for (int n = 0; n < NCH; n++)
        {
            for (int k = 0; k < key_counts[3 * n] && k < F; k++)
            {
                const float* p = pos + 3 * (op[n] + k);
                keys[((size_t)n * F + k) * 3 + 0] = make_float4(p[0], p[1], p[2], 0.f);
            }
            for (int k = 0; k < key_counts[3 * n + 1] && k < F; k++)
            {
                const float* q = rot + 4 * (orr[n] + k);
                keys[((size_t)n * F + k) * 3 + 1] = make_float4(q[0], q[1], q[2], q[3]);
            }
            for (int k = 0; k < key_counts[3 * n + 2] && k < F; k++)
            {
                const float* s = scl + 3 * (os[n] + k);
                keys[((size_t)n * F + k) * 3 + 2] = make_float4(s[0], s[1], s[2], 0.f);
            }
        }
        // step-ordered form for pose_simple_kernel: possible when no bone is driven by two played channels
        SkelRec&             sk     = c->skels[skel];
        bool                 simple = true;
        for (int n = 0; n < NCH && simple; n++)
            if (convert[n] != -1 && next[n] >= 0)
                simple = false;
        const int            Jp = pose_jp(J);
        std::vector<float4>   skeys;
        std::vector<uint32_t> sanim;
        if (simple)
        {
            skeys.assign(((size_t)Jp * F * 5 + 1) / 2, make_float4(0, 0, 0, 0)); // a | q | c tables (16, 16, 8 bytes per key)
            float4* const ta = skeys.data();
            float4* const tq = ta + (size_t)Jp * F;
            float2* const tc = reinterpret_cast<float2*>(tq + (size_t)Jp * F);
            sanim.assign((Jp + 31) / 32, 0u);
            for (int st = 0; st < Jp; st++)
            {
                const int ch = st < J ? first[sk.order[st]] : -1;
                if (ch >= 0)
                    sanim[st >> 5] |= 1u << (st & 31);
                for (int k = 0; k < F; k++)
                {
                    // {p.x, p.y, s.x, s.y} | {q.wxyz} | {p.z, s.z}: one table per component (KeyPair, pcl_kernels.cuh); bones without a channel get the identity track; beyond its
                    // clamp (the carried-over prefix minimum, skeleton.cpp:75-87) a track continues with copies of the clamp key
                    float4 kp = make_float4(0, 0, 0, 0), kq = make_float4(1, 0, 0, 0), ks = make_float4(1, 1, 1, 0);
                    if (ch >= 0)
                    {
                        const int kk = std::min(k, hic[ch]);
                        kp = keys[((size_t)ch * F + kk) * 3], kq = keys[((size_t)ch * F + kk) * 3 + 1], ks = keys[((size_t)ch * F + kk) * 3 + 2];
                    }
                    ta[(size_t)st * F + k] = make_float4(kp.x, kp.y, ks.x, ks.y);
                    tq[(size_t)st * F + k] = kq;
                    tc[(size_t)st * F + k] = make_float2(kp.z, ks.z);
                }
            }
        }
        // one device blob: keys | first | next | hi_clamp | step-ordered keys | animated-step bits
        const size_t kb = keys.size() * sizeof(float4), fb = (size_t)J * 4, nb = next.size() * 4, hb = hic.size() * 4;
        const size_t skb = skeys.size() * sizeof(float4), shb = sanim.size() * 4;
        auto         up16 = [](size_t x) { return (x + 15) & ~(size_t)15; };
        const size_t total = up16(kb) + up16(fb) + up16(nb) + up16(hb) + up16(skb) + up16(shb) + 16;
        ClipHost     ch;
        ch.skel = skel;
        PCL_CK(c, cudaMalloc(&ch.d_blob, total));
        unsigned char* d  = (unsigned char*)ch.d_blob;
        unsigned char* d2 = d + up16(kb) + up16(fb) + up16(nb) + up16(hb);
        {
            cudaError_t e = copy_h2d(c, d, keys.data(), kb);
            if (e == cudaSuccess) e = copy_h2d(c, d + up16(kb), first.data(), fb);
            if (e == cudaSuccess) e = copy_h2d(c, d + up16(kb) + up16(fb), next.data(), nb);
            if (e == cudaSuccess) e = copy_h2d(c, d + up16(kb) + up16(fb) + up16(nb), hic.data(), hb);
            if (e == cudaSuccess && simple) e = copy_h2d(c, d2, skeys.data(), skb);
            if (e == cudaSuccess && simple) e = copy_h2d(c, d2 + up16(skb), sanim.data(), shb);
            if (e != cudaSuccess) // nothing of a failed upload stays behind
            {
                cudaFree(ch.d_blob);
                (void)cudaGetLastError();
                return fail(c, PCL_ERR_CUDA, std::string("pcl_upload_clip: ") + cudaGetErrorString(e));
            }
        }
        ch.rec.keys           = (const float4*)d;
        ch.rec.first_chan     = (const int32_t*)(d + up16(kb));
        ch.rec.next_chan      = (const int32_t*)(d + up16(kb) + up16(fb));
        ch.rec.hi_clamp       = (const int32_t*)(d + up16(kb) + up16(fb) + up16(nb));
        ch.rec.F              = F;
        ch.rec.total_frame_m1 = total_frame - 1;
        ch.rec.C              = C;
        ch.rec.simple         = simple ? 1 : 0;
        ch.rec.keys_step      = simple ? (const float4*)d2 : nullptr;
        ch.rec.anim_step      = simple ? (const uint32_t*)(d2 + up16(skb)) : nullptr;
        ch.rec.bucket_base    = c->n_buckets;
        ch.frames             = total_frame;
        c->n_buckets += total_frame; // lo = floor(phase * (total_frame - 1)) is one of total_frame values
        sk.n_clips++;
        sk.n_simple += simple ? 1 : 0;
        c->clips.push_back(ch);
        const int id = (int)c->clips.size() - 1;
        if (id >= c->cap_clips)
        {
            const int ncap = std::max(16, c->cap_clips * 2);
            ClipRec*  nd   = nullptr;
            PCL_CK(c, cudaStreamSynchronize(c->compute));
            PCL_CK(c, cudaMalloc((void**)&nd, ncap * sizeof(ClipRec)));
            std::vector<ClipRec> all;
            for (auto& x : c->clips)
                all.push_back(x.rec);
            PCL_CK(c, copy_h2d(c, nd, all.data(), all.size() * sizeof(ClipRec)));
            cudaFree(c->d_clips);
            c->d_clips = nd, c->cap_clips = ncap;
        }
        else
            PCL_CK(c, copy_h2d(c, c->d_clips + id, &ch.rec, sizeof(ClipRec)));
        *out = id;
        return PCL_OK;
    }

    int pcl_upload_mask(pcl_ctx* c, pcl_skel skel, int32_t J, const int32_t* enabled, pcl_mask* out)
    {
        if (!valid_skel(c, skel) || !out || !enabled)
            return fail(c, PCL_ERR_INVALID, "pcl_upload_mask: bad argument");
        SkelRec& s = c->skels[skel];
        if (J < s.J)
            return fail(c, PCL_ERR_INVALID, "pcl_upload_mask: mask shorter than the skeleton (animation_system.cpp:122 reads enabled[bone])");
        PCL_CK(c, cudaSetDevice(c->device));
        if (s.n_masks >= s.cap_masks)
        {
            const int ncap = std::max(8, s.cap_masks * 2);
            int32_t*  nd   = nullptr;
            PCL_CK(c, cudaStreamSynchronize(c->compute));
            PCL_CK(c, cudaMalloc((void**)&nd, (size_t)ncap * s.J * 4));
            if (s.n_masks)
            {
                cudaError_t e = cudaMemcpyAsync(nd, s.d_masks, (size_t)s.n_masks * s.J * 4, cudaMemcpyDeviceToDevice, c->compute);
                if (e == cudaSuccess)
                    e = cudaStreamSynchronize(c->compute);
                if (e != cudaSuccess)
                {
                    cudaFree(nd);
                    (void)cudaGetLastError();
                    return fail(c, PCL_ERR_CUDA, std::string("pcl_upload_mask: ") + cudaGetErrorString(e));
                }
            }
            cudaFree(s.d_masks);
            s.d_masks = nd, s.cap_masks = ncap;
        }
        PCL_CK(c, copy_h2d(c, s.d_masks + (size_t)s.n_masks * s.J, enabled, (size_t)s.J * 4));
        c->masks.push_back(MaskHost {skel, s.n_masks});
        s.n_masks++;
        *out = (pcl_mask)c->masks.size() - 1;
        return PCL_OK;
    }

    int pcl_upload_mesh(pcl_ctx* c, int32_t V, const float* pos, const float* nrm, const float* tan, const int32_t* idx4,
                        const float* w4, int32_t max_slot, pcl_mesh* out)
    {
        if (!c || !out || V < 1 || !pos || !idx4 || !w4 || max_slot < 0)
            return fail(c, PCL_ERR_INVALID, "pcl_upload_mesh: bad argument");
        PCL_CK(c, cudaSetDevice(c->device));
        MeshRec m;
        m.V = V, m.Vp = (V + 7) & ~7, m.Vt = (V + LBS_VT - 1) / LBS_VT * LBS_VT, m.max_slot = max_slot;
        // (Vp is also the row pitch of the skin buffers.  tools/tma_store_peak.cu shows the store stream alone gaining 3 % when the
        // pitch is not a multiple of 4 KiB - 16 416 instead of 16 384 vertices - but the full kernel did not move, so no padding.)
        std::vector<float>  hp((size_t)m.Vt * 3, 0.f), hn((size_t)m.Vt * 3, 0.f), ht((size_t)m.Vt * 3, 0.f);
        std::vector<int4>   hi(m.Vt, make_int4(0, 0, 0, 0));
        std::vector<float4> hw(m.Vt, make_float4(0, 0, 0, 0));
        std::memcpy(hp.data(), pos, (size_t)V * 12);
        if (nrm)
            std::memcpy(hn.data(), nrm, (size_t)V * 12);
        if (tan)
            std::memcpy(ht.data(), tan, (size_t)V * 12);
        for (int v = 0; v < V; v++)
        {
            const float* wr = w4 + 4 * (size_t)v;
            // RenderResource::updateVertexBuffer, function/render/render_resource.cpp:714-725
            float inv = wr[0] + wr[1] + wr[2] + wr[3];
            inv       = (inv != 0.0f) ? 1 / inv : 1.0f;
            int   id[4];
            float w[4];
            for (int k = 0; k < 4; k++)
            {
                id[k] = idx4[4 * (size_t)v + k];
                w[k]  = wr[k] * inv;
                if (id[k] < 0 || id[k] > max_slot)
                    return fail(c, PCL_ERR_INVALID, "pcl_upload_mesh: joint index outside [0, max_slot]");
                // mesh.vert:77-102 gate (weight > 0 && index > 0) folded into the data: a gated influence
                // becomes (slot 0, weight 0) and contributes an exact zero
                if (!(w[k] > 0.0f && id[k] > 0))
                    id[k] = 0, w[k] = 0.f;
            }
            hi[v] = make_int4(id[0], id[1], id[2], id[3]);
            hw[v] = make_float4(w[0], w[1], w[2], w[3]);
        }
        // lane = instance forms.  Each vertex's (gated) influences are packed into four register slots so that
        // a bone that was already resident for the previously VISITED vertex keeps its slot; unused slots keep the
        // resident bone with weight 0 (an exact zero contribution).  The blend is then summed in slot order,
        // which differs from the shader's x,y,z,w order only by a permutation of the same four products.
        // d_vrec: visiting order 0,1,2,... (lbs_inst_kernel: a warp walks a contiguous run); the interleaved streams of
        // lbs_cta_kernel are packed by ensure_vrec_cta on first use.
        int rc;
        if ((rc = upload(c, &m.d_pos, hp)) || (rc = upload(c, &m.d_nrm, hn)) || (rc = upload(c, &m.d_tan, ht)) ||
            (rc = upload(c, &m.d_idx, hi)) || (rc = upload(c, &m.d_w, hw)))
            return rc;
        m.h_pos = std::move(hp), m.h_nrm = std::move(hn), m.h_tan = std::move(ht), m.h_idx = std::move(hi), m.h_w = std::move(hw);
        if (max_slot * 3 <= 0xffff)
        {
            std::vector<VRec> hv((size_t)std::max(m.Vt, m.Vp) + LBS_I_SCH * LBS_I_CHUNK); // tail padding: a record stage may overrun the last chunk
            std::memset(hv.data(), 0, hv.size() * sizeof(VRec));
            int cur[4] = {0, 0, 0, 0};
            for (size_t v = 0; v < hv.size(); v++)
                pack_vrec(m, hv[v], v, cur, v ? &hv[v - 1] : nullptr);
            if ((rc = upload(c, &m.d_vrec, hv)))
                return rc;
        }
        c->meshes.push_back(m);
        *out = (pcl_mesh)c->meshes.size() - 1;
        return PCL_OK;
    }

    // ------------------------------------------------------------------------------------------
    // releasing assets (a level unload / reload must not leave its device copies behind)
    // ------------------------------------------------------------------------------------------
    int pcl_clip_release(pcl_ctx* c, pcl_clip clip)
    {
        if (!c || clip < 0 || clip >= (int)c->clips.size() || c->clips[clip].released)
            return fail(c, PCL_ERR_INVALID, "pcl_clip_release: bad handle");
        PCL_CK(c, cudaSetDevice(c->device));
        PCL_CK(c, cudaStreamSynchronize(c->compute)); // no launch may still read it
        ClipHost& ch = c->clips[clip];
        SkelRec&  sk = c->skels[ch.skel];
        cudaFree(ch.d_blob);
        ch.d_blob = nullptr, ch.released = true;
        sk.n_clips--, sk.n_simple -= ch.rec.simple ? 1 : 0;
        // give its (clip, frame) buckets back: the frame-coherent order only exists while all buckets fit its counters
        int32_t base = 0;
        for (auto& x : c->clips)
            if (!x.released)
                x.rec.bucket_base = base, base += x.frames;
        c->n_buckets = base;
        ch.rec = ClipRec {};
        std::vector<ClipRec> all;
        for (auto& x : c->clips)
            all.push_back(x.rec);
        if (c->d_clips)
            PCL_CK(c, copy_h2d(c, c->d_clips, all.data(), all.size() * sizeof(ClipRec)));
        return PCL_OK;
    }
    int pcl_mesh_release(pcl_ctx* c, pcl_mesh mesh)
    {
        if (!c || mesh < 0 || mesh >= (int)c->meshes.size() || c->meshes[mesh].released)
            return fail(c, PCL_ERR_INVALID, "pcl_mesh_release: bad handle");
        PCL_CK(c, cudaSetDevice(c->device));
        PCL_CK(c, cudaStreamSynchronize(c->compute));
        MeshRec& m = c->meshes[mesh];
        cudaFree(m.d_pos), cudaFree(m.d_nrm), cudaFree(m.d_tan), cudaFree(m.d_idx), cudaFree(m.d_w), cudaFree(m.d_vrec);
        for (int i = 0; i < 3; i++)
            cudaFree(m.d_vrec_cta[i]), cudaFree(m.d_vrec_blk[i]);
        cudaFree(m.d_blk_slots), cudaFree(m.d_blk_cnt);
        m = MeshRec {};
        m.released = true;
        return PCL_OK;
    }
    int pcl_skeleton_release(pcl_ctx* c, pcl_skel skel)
    {
        if (!valid_skel(c, skel))
            return fail(c, PCL_ERR_INVALID, "pcl_skeleton_release: bad handle");
        SkelRec& s = c->skels[skel];
        if (s.users > 0 || s.n_clips > 0)
            return fail(c, PCL_ERR_INVALID, "pcl_skeleton_release: batches, pools or clips still refer to the skeleton");
        PCL_CK(c, cudaSetDevice(c->device));
        PCL_CK(c, cudaStreamSynchronize(c->compute));
        {
            std::lock_guard<std::mutex> lk(g_const_mu);
            if (c->device < PCL_STATIC_MAX_DEVICES && g_const_owner[c->device].ctx == c && g_const_owner[c->device].skel == skel)
                g_const_owner[c->device] = ConstOwner();
        }
        cudaFree(s.d_steps), cudaFree(s.d_sched), cudaFree(s.d_masks);
        s.d_steps = nullptr, s.d_sched = nullptr, s.d_masks = nullptr, s.n_masks = s.cap_masks = 0;
        s.h_steps.clear(), s.h_steps.shrink_to_fit();
        s.released = true;
        return PCL_OK;
    }

    // ------------------------------------------------------------------------------------------
    // instances
    // ------------------------------------------------------------------------------------------
    int pcl_batch_create(pcl_ctx* c, pcl_skel skel, int64_t N, int32_t K, pcl_batch** out)
    {
        if (!valid_skel(c, skel) || !out || N < 1 || K < 1)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_create: bad argument");
        if (K > 4)
            return fail(c, PCL_ERR_UNSUPPORTED, "pcl_batch_create: at most 4 clip slots per instance");
        PCL_CK(c, cudaSetDevice(c->device));
        pcl_batch* b = new pcl_batch();
        b->ctx = c, b->skel = skel, b->N = N, b->K = K;
        const size_t   n = (size_t)N * K;
        const SkelRec& s = c->skels[skel];
        std::vector<float> ones(n, 1.f);
        cudaError_t        e    = cudaSuccess;
        auto               step = [&e](cudaError_t r) { if (e == cudaSuccess) e = r; };
        step(cudaMalloc((void**)&b->d_clip_ids, n * 4));
        step(cudaMalloc((void**)&b->d_mask_ids, n * 4));
        step(cudaMalloc((void**)&b->d_phase, n * 4));
        step(cudaMalloc((void**)&b->d_weight, n * 4));
        step(cudaMalloc((void**)&b->d_len, n * 4));
        step(cudaMalloc((void**)&b->d_palette, (size_t)N * s.pitch * 48));
        if (e == cudaSuccess)
        {
            step(cudaMemsetAsync(b->d_clip_ids, 0, n * 4, c->compute));
            step(cudaMemsetAsync(b->d_mask_ids, 0xff, n * 4, c->compute));
            step(cudaMemsetAsync(b->d_phase, 0, n * 4, c->compute));
            step(copy_h2d(c, b->d_weight, ones.data(), n * 4));
            step(copy_h2d(c, b->d_len, ones.data(), n * 4));
            step(cudaMemsetAsync(b->d_palette, 0, (size_t)N * s.pitch * 48, c->compute));
        }
        if (e != cudaSuccess) // a crowd too big for the GPU must leave nothing behind
        {
            (void)cudaGetLastError();
            cudaFree(b->d_clip_ids), cudaFree(b->d_mask_ids), cudaFree(b->d_phase), cudaFree(b->d_phase_alt), cudaFree(b->d_weight), cudaFree(b->d_len), cudaFree(b->d_palette);
            delete b;
            return fail(c, e == cudaErrorMemoryAllocation ? PCL_ERR_OOM : PCL_ERR_CUDA, std::string("pcl_batch_create: ") + cudaGetErrorString(e));
        }
        c->skels[skel].users++;
        *out = b;
        return PCL_OK;
    }

    int pcl_batch_destroy(pcl_batch* b)
    {
        if (!b)
            return PCL_ERR_INVALID;
        cudaSetDevice(b->ctx->device);
        b->ctx->skels[b->skel].users--;
        cudaStreamSynchronize(b->ctx->compute), cudaStreamSynchronize(b->ctx->copy);
        cudaFree(b->d_clip_ids), cudaFree(b->d_mask_ids), cudaFree(b->d_phase), cudaFree(b->d_phase_alt), cudaFree(b->d_weight), cudaFree(b->d_len);
        if (b->pool)
            b->pool->attached--;
        else
            cudaFree(b->d_palette);
        cudaFree(b->d_derived), cudaFree(b->d_export), cudaFree(b->d_export12), cudaFree(b->d_order), cudaFree(b->d_ohist);
        if (b->h_stage)
            cudaFreeHost(b->h_stage);
        if (b->ev_stage)
            cudaEventDestroy(b->ev_stage);
        if (b->h_io)
            cudaFreeHost(b->h_io);
        if (b->ev_frame)
            cudaEventDestroy(b->ev_frame);
        delete b;
        return PCL_OK;
    }

    // ------------------------------------------------------------------------------------------
    // palette pool
    // ------------------------------------------------------------------------------------------
    static_assert(sizeof(cudaIpcMemHandle_t) == PCL_POOL_HANDLE_BYTES, "PCL_POOL_HANDLE_BYTES must match cudaIpcMemHandle_t");

    int pcl_pool_create(pcl_ctx* c, pcl_skel skel, int64_t capacity, pcl_pool** out)
    {
        if (!valid_skel(c, skel) || !out || capacity < 1)
            return fail(c, PCL_ERR_INVALID, "pcl_pool_create: bad argument");
        PCL_CK(c, cudaSetDevice(c->device));
        const SkelRec& s = c->skels[skel];
        pcl_pool*      p = new pcl_pool();
        p->ctx = c, p->capacity = capacity, p->pitch = s.pitch, p->J = s.J, p->owner = true;
        const size_t bytes = (size_t)capacity * s.pitch * 48;
        cudaError_t  e     = cudaMalloc((void**)&p->base, bytes);
        if (e == cudaSuccess)
            e = cudaMemsetAsync(p->base, 0, bytes, c->compute);
        if (e == cudaSuccess)
            e = cudaStreamSynchronize(c->compute); // peers may write the pool as soon as they have its handle
        if (e != cudaSuccess)
        {
            cudaFree(p->base);
            delete p;
            return fail(c, PCL_ERR_CUDA, std::string("pcl_pool_create: ") + cudaGetErrorString(e));
        }
        *out = p;
        return PCL_OK;
    }
    int pcl_pool_export(pcl_pool* p, void* handle)
    {
        if (!p || !handle)
            return PCL_ERR_INVALID;
        pcl_ctx* c = p->ctx;
        if (!p->owner)
            return fail(c, PCL_ERR_INVALID, "pcl_pool_export: only the pool's owner can export it");
        PCL_CK(c, cudaSetDevice(c->device));
        cudaIpcMemHandle_t h;
        PCL_CK(c, cudaIpcGetMemHandle(&h, p->base));
        std::memcpy(handle, &h, sizeof(h));
        return PCL_OK;
    }
    int pcl_pool_open(pcl_ctx* c, pcl_skel skel, int64_t capacity, const void* handle, pcl_pool** out)
    {
        if (!valid_skel(c, skel) || !out || !handle || capacity < 1)
            return fail(c, PCL_ERR_INVALID, "pcl_pool_open: bad argument");
        PCL_CK(c, cudaSetDevice(c->device));
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handle, sizeof(h));
        void* base = nullptr;
        PCL_CK(c, cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess)); // maps the owner's buffer, enabling peer access if it is another GPU
        const SkelRec& s = c->skels[skel];
        pcl_pool*      p = new pcl_pool();
        p->ctx = c, p->capacity = capacity, p->pitch = s.pitch, p->J = s.J, p->owner = false, p->base = (float4*)base;
        *out = p;
        return PCL_OK;
    }
    int pcl_pool_destroy(pcl_pool* p)
    {
        if (!p)
            return PCL_ERR_INVALID;
        pcl_ctx* c = p->ctx;
        if (p->attached > 0)
            return fail(c, PCL_ERR_INVALID, "pcl_pool_destroy: batches are still attached to the pool");
        PCL_CK(c, cudaSetDevice(c->device));
        PCL_CK(c, cudaDeviceSynchronize());
        if (p->owner)
            cudaFree(p->base);
        else
            cudaIpcCloseMemHandle(p->base);
        delete p;
        return PCL_OK;
    }
    int pcl_batch_attach_pool(pcl_batch* b, pcl_pool* p, int64_t first)
    {
        if (!b || !p)
            return PCL_ERR_INVALID;
        pcl_ctx*       c = b->ctx;
        const SkelRec& s = c->skels[b->skel];
        if (b->pool)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_attach_pool: the batch is already attached to a pool");
        if (p->pitch != s.pitch || p->J != s.J)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_attach_pool: the pool was laid out for a skeleton with another bone count");
        if (first < 0 || first + b->N > p->capacity)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_attach_pool: rows outside the pool");
        PCL_CK(c, cudaSetDevice(c->device));
        if (p->ctx->device != c->device) // same process, pool on another GPU: the pose kernel will store to it over NVLink / PCIe
        {
            int can = 0;
            PCL_CK(c, cudaDeviceCanAccessPeer(&can, c->device, p->ctx->device));
            if (!can)
                return fail(c, PCL_ERR_UNSUPPORTED, "pcl_batch_attach_pool: no peer access between the batch's GPU and the pool's GPU");
            cudaError_t e = cudaDeviceEnablePeerAccess(p->ctx->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
                return fail(c, PCL_ERR_CUDA, std::string("pcl_batch_attach_pool: ") + cudaGetErrorString(e));
            (void)cudaGetLastError();
        }
        PCL_CK(c, cudaStreamSynchronize(c->compute));
        PCL_CK(c, cudaStreamSynchronize(c->copy));
        cudaFree(b->d_palette);
        b->d_palette = p->base + (size_t)first * s.pitch * 3;
        b->pool      = p;
        b->posed     = false;
        p->attached++;
        return PCL_OK;
    }
    int pcl_batch_push_to_pool(pcl_batch* b, pcl_pool* p, int64_t first)
    {
        if (!b || !p)
            return PCL_ERR_INVALID;
        pcl_ctx*       c = b->ctx;
        const SkelRec& s = c->skels[b->skel];
        if (b->pool)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_push_to_pool: the batch is attached to a pool (its palettes are already there)");
        if (!b->posed)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_push_to_pool: pcl_batch_pose / pcl_batch_tick has not run");
        if (p->pitch != s.pitch || p->J != s.J)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_push_to_pool: the pool was laid out for a skeleton with another bone count");
        if (first < 0 || first + b->N > p->capacity)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_push_to_pool: rows outside the pool");
        PCL_CK(c, cudaSetDevice(c->device));
        PCL_CK(c, cudaStreamWaitEvent(c->copy, c->ev_pose, 0));
        PCL_CK(c, cudaMemcpyAsync(p->base + (size_t)first * s.pitch * 3, b->d_palette, (size_t)b->N * s.pitch * 48, cudaMemcpyDefault, c->copy));
        PCL_CK(c, cudaEventRecord(c->ev_copy, c->copy));
        c->copy_pending = true; // the next pose must not overwrite the palette while the copy engine reads it
        return PCL_OK;
    }
    int pcl_batch_adopt_palettes(pcl_batch* b)
    {
        if (!b)
            return PCL_ERR_INVALID;
        if (!b->pool)
            return fail(b->ctx, PCL_ERR_INVALID, "pcl_batch_adopt_palettes: the batch is not attached to a pool");
        b->posed = true;
        return PCL_OK;
    }

    int pcl_batch_set(pcl_batch* b, int64_t first, int64_t count, const int32_t* clip_ids, const float* phase, const float* weight,
                      const float* clip_len, const int32_t* mask_ids)
    {
        if (!b || first < 0 || count < 0 || first + count > b->N)
            return fail(b ? b->ctx : nullptr, PCL_ERR_INVALID, "pcl_batch_set: range outside the batch");
        pcl_ctx* c = b->ctx;
        PCL_CK(c, cudaSetDevice(c->device));
        const int K = b->K;
        std::vector<int32_t> ti;
        std::vector<float>   tf;
        auto put_i = [&](int32_t* dst, const int32_t* src, bool is_mask) -> int {
            ti.resize((size_t)count * K);
            for (int64_t i = 0; i < count; i++)
                for (int k = 0; k < K; k++)
                {
                    int32_t v = src[i * K + k];
                    if (is_mask)
                    {
                        if (v >= 0)
                        {
                            if (v >= (int)c->masks.size() || c->masks[v].skel != b->skel)
                                return fail(c, PCL_ERR_INVALID, "pcl_batch_set: mask id does not belong to the batch's skeleton");
                            v = c->masks[v].local;
                        }
                        else
                            v = -1;
                    }
                    else if (v < 0 || v >= (int)c->clips.size() || c->clips[v].skel != b->skel || c->clips[v].released)
                        return fail(c, PCL_ERR_INVALID, "pcl_batch_set: clip id does not belong to the batch's skeleton");
                    ti[(size_t)k * count + i] = v;
                }
            for (int k = 0; k < K; k++)
                PCL_CK(c, cudaMemcpyAsync(dst + (size_t)k * b->N + first, ti.data() + (size_t)k * count, (size_t)count * 4,
                                          cudaMemcpyHostToDevice, c->compute));
            PCL_CK(c, cudaStreamSynchronize(c->compute));
            return PCL_OK;
        };
        auto put_f = [&](float* dst, const float* src) -> int {
            if (K == 1)
            {
                PCL_CK(c, cudaMemcpyAsync(dst + first, src, (size_t)count * 4, cudaMemcpyHostToDevice, c->compute));
            }
            else
            {
                tf.resize((size_t)count * K);
                for (int64_t i = 0; i < count; i++)
                    for (int k = 0; k < K; k++)
                        tf[(size_t)k * count + i] = src[i * K + k];
                for (int k = 0; k < K; k++)
                    PCL_CK(c, cudaMemcpyAsync(dst + (size_t)k * b->N + first, tf.data() + (size_t)k * count, (size_t)count * 4,
                                              cudaMemcpyHostToDevice, c->compute));
            }
            PCL_CK(c, cudaStreamSynchronize(c->compute));
            return PCL_OK;
        };
        int rc;
        if (clip_ids && (rc = put_i(b->d_clip_ids, clip_ids, false)))
            return rc;
        if (mask_ids && (rc = put_i(b->d_mask_ids, mask_ids, true)))
            return rc;
        if (phase && (rc = put_f(b->d_phase, phase)))
            return rc;
        if (weight && (rc = put_f(b->d_weight, weight)))
            return rc;
        if (clip_len && (rc = put_f(b->d_len, clip_len)))
            return rc;
        return PCL_OK;
    }

    int pcl_batch_set_frame_async(pcl_batch* b, int64_t first, int64_t count, const float* phase, const float* weight)
    {
        if (!b || first < 0 || count < 0 || first + count > b->N)
            return fail(b ? b->ctx : nullptr, PCL_ERR_INVALID, "pcl_batch_set_frame_async: range outside the batch");
        pcl_ctx* c = b->ctx;
        PCL_CK(c, cudaSetDevice(c->device));
        const int    K = b->K;
        const size_t n = (size_t)b->N * K;
        if (!b->h_stage)
        {
            PCL_CK(c, cudaHostAlloc((void**)&b->h_stage, 2 * n * sizeof(float), cudaHostAllocDefault));
            PCL_CK(c, cudaEventCreateWithFlags(&b->ev_stage, cudaEventDisableTiming));
        }
        else
            PCL_CK(c, cudaEventSynchronize(b->ev_stage)); // the previous frame's inputs have been read out of the staging buffer
        const float* src[2] = {phase, weight};
        float*       dst[2] = {b->d_phase, b->d_weight};
        for (int a = 0; a < 2; a++)
        {
            if (!src[a])
                continue;
            float* st = b->h_stage + a * n;
            if (K == 1)
                std::memcpy(st, src[a], (size_t)count * sizeof(float));
            else
                for (int64_t i = 0; i < count; i++)
                    for (int k = 0; k < K; k++)
                        st[(size_t)k * count + i] = src[a][i * K + k]; // [count][K] -> [K][count], the device layout
            for (int k = 0; k < K; k++)
                PCL_CK(c, cudaMemcpyAsync(dst[a] + (size_t)k * b->N + first, st + (size_t)k * count, (size_t)count * sizeof(float),
                                          cudaMemcpyHostToDevice, c->compute));
        }
        PCL_CK(c, cudaEventRecord(b->ev_stage, c->compute));
        return PCL_OK;
    }

    int pcl_batch_get_phases(pcl_batch* b, int64_t first, int64_t count, float* phase)
    {
        if (!b || !phase || first < 0 || count < 0 || first + count > b->N)
            return fail(b ? b->ctx : nullptr, PCL_ERR_INVALID, "pcl_batch_get_phases: range outside the batch");
        pcl_ctx* c = b->ctx;
        PCL_CK(c, cudaSetDevice(c->device));
        PCL_CK(c, cudaStreamSynchronize(c->compute));
        std::vector<float> t((size_t)count * b->K);
        for (int k = 0; k < b->K; k++)
            PCL_CK(c, cudaMemcpy(t.data() + (size_t)k * count, b->d_phase + (size_t)k * b->N + first, (size_t)count * 4, cudaMemcpyDeviceToHost));
        for (int64_t i = 0; i < count; i++)
            for (int k = 0; k < b->K; k++)
                phase[i * b->K + k] = t[(size_t)k * count + i];
        return PCL_OK;
    }

    int pcl_batch_advance(pcl_batch* b, float dt, int32_t mode)
    {
        if (!b || (mode != PCL_ADVANCE_FIRST_ONLY && mode != PCL_ADVANCE_ALL))
            return fail(b ? b->ctx : nullptr, PCL_ERR_INVALID, "pcl_batch_advance: bad mode");
        pcl_ctx* c = b->ctx;
        PCL_CK(c, cudaSetDevice(c->device));
        const long long total = (mode == PCL_ADVANCE_ALL) ? b->N * b->K : b->N; // slot 0 is row 0 of the [K][N] arrays
        advance_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->compute>>>(b->d_phase, b->d_len, 0, total, dt);
        c->launches++;
        PCL_POST_LAUNCH(c, c->compute, "advance_kernel");
        return PCL_OK;
    }

    // where a launch takes its per-instance inputs from / additionally writes its result rows to (pcl_batch_frame)
    struct FrameIO
    {
        const int32_t* clip_ids;
        float*         phase;
        const float*   weight;
        const float*   len;
        const int32_t* mask_ids;
        float4*        out12;
        bool           out12_written;
    };
    static int launch_pose(pcl_batch* b, float dt, int32_t advance, FrameIO* io = nullptr)
    {
        pcl_ctx* c = b->ctx;
        PCL_CK(c, cudaSetDevice(c->device));
        const SkelRec& s = c->skels[b->skel];
        if (c->clips.empty())
            return fail(c, PCL_ERR_INVALID, "pcl_batch_pose: no clip uploaded");
        if (c->copy_pending) // an async palette export may still be reading the palette
        {
            PCL_CK(c, cudaStreamWaitEvent(c->compute, c->ev_copy, 0));
            c->copy_pending = false;
        }
        PoseParams p {};
        p.steps = s.d_steps, p.clips = c->d_clips, p.masks = s.d_masks;
        p.clip_ids = b->d_clip_ids, p.phase = b->d_phase, p.weight = b->d_weight, p.clip_len = b->d_len;
        p.mask_ids = (b->K > 1 && s.n_masks > 0) ? b->d_mask_ids : nullptr;
        if (io)
        {
            p.clip_ids = io->clip_ids, p.phase = io->phase, p.weight = io->weight, p.clip_len = io->len;
            p.mask_ids = (b->K > 1 && s.n_masks > 0) ? io->mask_ids : nullptr;
        }
        p.palette = b->d_palette, p.derived = b->d_derived;
        p.N = b->N, p.J = s.J, p.pitch = s.pitch, p.nslot = s.nslot;
        p.n_steps_padded = (s.J + POSE_G - 1) / POSE_G * POSE_G;
        p.split_base[0] = s.split_base[0], p.split_base[1] = s.split_base[1], p.split_n[0] = s.split_n[0], p.split_n[1] = s.split_n[1];
        p.dt = dt, p.advance = advance;
        const long long warps = (b->N + 31) / 32;
        const size_t    per_warp = pose_smem_floats_per_warp(s.nslot) * 4;
        int             wpb = c->opt_pose_wpb;
        if (wpb == 0) // enough blocks for every SM first, then larger blocks
            wpb = warps >= 8LL * 4 * c->sm_count ? 4 : (warps >= 8LL * 2 * c->sm_count ? 2 : 1);
        const size_t steps_smem = PCL_POSE_STEPS_SMEM ? (size_t)s.J * sizeof(StepRec) : 0; // (pose_simple_kernel's copy of the step records)
        while (wpb > 1 && per_warp * wpb + steps_smem > (size_t)c->max_smem_optin)
            wpb >>= 1;
        // step-ordered kernel when every clip of this skeleton has that form and the step records fit the constant bank
        const bool simple = c->opt_pose_generic == 0 && s.J <= POSE_CMAX && s.n_clips > 0 && s.n_simple == s.n_clips &&
                            c->device < PCL_STATIC_MAX_DEVICES && per_warp + steps_smem <= (size_t)c->max_smem_optin;
        if (simple)
            wpb = std::min(wpb, POSE_S_THREADS / 32);
        size_t         smem = per_warp * wpb + (simple && PCL_POSE_STEPS_SMEM ? (size_t)s.J * sizeof(StepRec) : 0);
        const unsigned grid = (unsigned)((warps + wpb - 1) / wpb);
        // the bank's contents must stay put from the upload until this launch is enqueued
        std::unique_lock<std::mutex> bank_lock(g_const_mu, std::defer_lock);
        if (simple && !PCL_POSE_STEPS_SMEM)
        {
            bank_lock.lock();
            ConstOwner& own = g_const_owner[c->device];
            if (own.ctx != c || own.skel != b->skel)
            {
                if (own.ctx && own.ctx != c) // another context's kernels may still be reading the bank
                    PCL_CK(c, cudaStreamSynchronize(own.ctx->compute));
                PCL_CK(c, cudaMemcpyToSymbolAsync(c_steps, s.h_steps.data(), s.h_steps.size() * sizeof(StepRec), 0, cudaMemcpyHostToDevice, c->compute));
                own.ctx = c, own.skel = b->skel;
            }
        }
        // frame-coherent visiting order: worth three small launches (~45 us) once the crowd is large; measured with
        // tools/pose_bench.py: 1 M instances 1.63 -> 1.51 ms (K = 1), 2.43 -> 2.18 ms (K = 2); 100 k instances, K = 2: 0.27 -> 0.30 ms
        // (not when the palettes go to another GPU's pool: rows visited in random order over gigabytes of peer memory ran at
        // 69 GB/s against 350 GB/s in instance order, tools/pool_gather_check.py on two B200s)
        const bool remote = b->pool && (!b->pool->owner || b->pool->ctx->device != c->device);
        const bool bulk_remote = remote && c->opt_pool_bulk != 0; // palettes in another GPU's memory leave as bulk copies (pose_simple_kernel, BULK)
        // (thresholds re-measured on the final kernels, order_* launches with 2 048 items per block: one clip per instance 30 k / 40 k /
        // 60 k / 80 k / 100 k instances: +35 / 0 / -7 / -13 / -13 % with the order; two clips, 100 k / 150 k / 200 k: -6 / -4 / -5 %
        // with 8 mixed clips; when all instances blend the same two clips - bench.py's crowd - the unsorted gathers already fall
        // into three or four lines per load and the order buys -3.5 % alone, +1.5 % inside the bench's step: two clips from 128 k;
        // three and four clips: not re-measured)
        const long long sort_from = b->K == 1 ? 48 * 1024 : b->K == 2 ? 128 * 1024 : 262144;
        // One wave or two?  With 2 blocks per SM the lane kernel holds 8 warps per SM; a crowd of 8 to 12 warps per SM would run a
        // second, nearly empty wave (40 000 one-clip instances: 75 us against 48 us for 36 000).  The 168-register form holds 12:
        // one wave, at a third more per wave.  Measured (tools/pose_bench.py --min-blocks 2 / 3, 8 mixed clips): one clip, 40 k /
        // 48 k / 56 k instances 75 / 87 / 88 us -> 53 / 61 / 66 us (the frame-coherent order with either form: 65-93 us), 20 k /
        // 30 k: 33 / 40 -> 30 / 36 us; two clips, 40 k / 48 k / 56 k: 115 / 117 / 117 -> 79 / 92 / 97 us, 30 k: 57 -> 63 us.
        const bool one_wave3 = simple && b->K <= 2 && c->opt_pose_min_blocks == 0 && warps <= 12LL * c->sm_count &&
                               (b->K == 1 || warps > 8LL * c->sm_count);
        bool sorted = simple && !io && c->n_buckets <= ORDER_MAX_BUCKETS && b->N < (1LL << 31) &&
                      (c->opt_pose_sort == 1 || (c->opt_pose_sort < 0 && b->N >= sort_from && !remote && !one_wave3));
        if (sorted)
        {
            if (!b->d_order)
            {
                PCL_CK(c, cudaMalloc((void**)&b->d_order, (size_t)b->N * 4));
                PCL_CK(c, cudaMalloc((void**)&b->d_ohist, (size_t)2 * ORDER_MAX_BUCKETS * 4));
                PCL_CK(c, cudaMemsetAsync(b->d_ohist, 0, (size_t)2 * ORDER_MAX_BUCKETS * 4, c->compute));
            }
            OrderParams q {};
            q.clips = c->d_clips, q.clip_ids = b->d_clip_ids, q.phase = b->d_phase, q.clip_len = b->d_len;
            q.hist = b->d_ohist, q.cursor = b->d_ohist + ORDER_MAX_BUCKETS, q.order = b->d_order;
            q.N = b->N, q.nb = c->n_buckets, q.dt = dt, q.advance = advance;
            q.levels = (b->K >= 2 && (long long)c->n_buckets * c->n_buckets <= ORDER_MAX_BUCKETS) ? 2 : 1;
            q.nbk    = q.levels == 2 ? c->n_buckets * c->n_buckets : c->n_buckets;
            const unsigned ogrid = (unsigned)((b->N + ORDER_THREADS * ORDER_PER_THREAD - 1) / (ORDER_THREADS * ORDER_PER_THREAD));
            order_hist_kernel<<<ogrid, ORDER_THREADS, (size_t)q.nbk * 4, c->compute>>>(q);
            order_scan_kernel<<<1, 1024, 0, c->compute>>>(q);
            { int rc__ = ensure_smem(c, (const void*)order_scatter_kernel, (size_t)q.nbk * 8); if (rc__) return rc__; }
            order_scatter_kernel<<<ogrid, ORDER_THREADS, (size_t)q.nbk * 8, c->compute>>>(q);
            c->launches += 3;
            PCL_POST_LAUNCH(c, c->compute, "order_*_kernel");
            p.order = b->d_order;
            if (b->K == 1)
                smem += (size_t)wpb * pose_wkeys_floats(s.J) * 4; // the warps' bucket keys (pose_simple_kernel<1, true, *>)
        }
        // small one-clip crowds: the level-parallel kernel (8 warps share 32 instances), while it leaves no SM oversubscribed
        {
            const size_t smem_l = pose_l_smem_bytes(s.J, s.n_rounds);
            const bool   fits_l = simple && b->K == 1 && !sorted && s.d_sched && smem_l <= (size_t)c->max_smem_optin;
            // measured (one clip, player skeleton): 14.6 us against 27-29 us (lane kernel) / 20.5 us (its two-half form) up to 4 000
            // instances = one CTA per SM; 8 000: 24.8 against 27.0 / 21.8 - so beyond one wave only where there is no two-half form
            const bool   has_split = s.split_n[0] > 0 && c->opt_pose_split != 0 && !p.derived && !remote;
            const bool   small     = warps <= (has_split ? 1LL : 2LL) * c->sm_count;
            if (fits_l && (c->opt_pose_level == 1 || (c->opt_pose_level < 0 && (small || io))))
            {
                p.sched = s.d_sched, p.n_rounds = s.n_rounds;
                if (io && io->out12)
                    p.out12 = io->out12, io->out12_written = true;
                { int rc__ = ensure_smem(c, (const void*)pose_level_kernel, smem_l); if (rc__) return rc__; }
                pose_level_kernel<<<(unsigned)warps, POSE_L_WARPS * 64, smem_l, c->compute>>>(p);
                c->launches++;
                PCL_POST_LAUNCH(c, c->compute, "pose_level_kernel");
                PCL_CK(c, cudaEventRecord(c->ev_pose, c->compute));
                b->posed = true;
                return PCL_OK;
            }
        }
        // registers vs resident warps (see pose_simple_kernel): 2 blocks per SM, except where 3 save a second wave (one_wave3 above).
        // (Round 1 preferred 3 blocks for two-clip crowds beyond ~300 k instances; with the packed, straight-line groups the
        // 168-register form spills: 1.98 against 1.86 ms at 1 M.)
        // one-clip crowds too large for the level kernel, too small to give every scheduler two warps: two warps per 32 instances,
        // half a skeleton each (pose_simple_kernel<SPLIT>); measured with tools/pose_bench.py --clips 1: 8 000 instances 27.0 -> 21.8 us
        // (level kernel 24.8), 10 000: 28.8 -> 24.8, 16 000: 29.9 -> 25.0, 20 000 (more warps than the SMs hold at once): 33 -> 40
        if (simple && b->K == 1 && !sorted && !io && !p.derived && !remote && s.split_n[0] > 0 &&
            (c->opt_pose_split == 1 || (c->opt_pose_split < 0 && warps * 2 <= 8LL * c->sm_count)))
        {
            if (advance != 0 && !b->d_phase_alt)
                PCL_CK(c, cudaMalloc((void**)&b->d_phase_alt, (size_t)b->N * 4));
            p.phase_out = b->d_phase_alt;
            { int rc__ = ensure_smem(c, (const void*)pose_simple_kernel<1, false, 2, false, false, true>, smem); if (rc__) return rc__; }
            pose_simple_kernel<1, false, 2, false, false, true><<<dim3(grid, 2), wpb * 32, smem, c->compute>>>(p);
            if (advance != 0)
                std::swap(b->d_phase, b->d_phase_alt); // both halves read the old phases; the first one wrote the new ones here
            c->launches++;
            PCL_POST_LAUNCH(c, c->compute, "pose_simple_kernel<split>");
            PCL_CK(c, cudaEventRecord(c->ev_pose, c->compute));
            b->posed = true;
            return PCL_OK;
        }
        int mb = c->opt_pose_min_blocks;
        if (mb == 0)
            mb = (one_wave3 && !sorted) ? 3 : 2;
        if (b->K > 2)
            mb = 2;
#define PCL_LAUNCH_POSE_SD(KK, SS, MM, DD)                                                                                  \
    {                                                                                                                       \
        if (KK == 2 && MM == 2 && !DD && !p.mask_ids) /* two clips, no bone masks: blend weights hoisted out of the walk */  \
        {                                                                                                                   \
            { int rc__ = ensure_smem(c, (const void*)pose_simple_kernel<2, SS, 2, false, false, false, false>, smem); if (rc__) return rc__; } \
            pose_simple_kernel<2, SS, 2, false, false, false, false><<<grid, wpb * 32, smem, c->compute>>>(p);              \
        }                                                                                                                   \
        else                                                                                                                \
        {                                                                                                                   \
            { int rc__ = ensure_smem(c, (const void*)pose_simple_kernel<KK, SS, MM, DD>, smem); if (rc__) return rc__; }    \
            pose_simple_kernel<KK, SS, MM, DD><<<grid, wpb * 32, smem, c->compute>>>(p);                                    \
        }                                                                                                                   \
    }
#define PCL_LAUNCH_POSE_S(KK, SS, MM)                                                                                       \
    {                                                                                                                       \
        if (p.derived) PCL_LAUNCH_POSE_SD(KK, SS, 2, true) /* (the read-back form exists for 2 blocks per SM only) */       \
        else if (bulk_remote && KK <= 2 && !SS)                                                                             \
        {                                                                                                                   \
            { int rc__ = ensure_smem(c, (const void*)pose_simple_kernel<(KK <= 2 ? KK : 1), false, 2, false, true>, smem); if (rc__) return rc__; } \
            pose_simple_kernel<(KK <= 2 ? KK : 1), false, 2, false, true><<<grid, wpb * 32, smem, c->compute>>>(p);          \
        }                                                                                                                   \
        else PCL_LAUNCH_POSE_SD(KK, SS, MM, false)                                                                          \
    }
#define PCL_LAUNCH_POSE_MB(KK, SS)                                                                                          \
    {                                                                                                                       \
        if (KK > 2 || mb == 2) PCL_LAUNCH_POSE_S(KK, SS, 2)                                                                 \
        else if (mb == 3) PCL_LAUNCH_POSE_S((KK > 2 ? 1 : KK), SS, 3)                                                       \
        else PCL_LAUNCH_POSE_S((KK > 2 ? 1 : KK), SS, 4)                                                                    \
    }
#define PCL_LAUNCH_POSE(KK)                                                                                                 \
    {                                                                                                                       \
        if (simple && sorted) PCL_LAUNCH_POSE_MB(KK, true)                                                                  \
        else if (simple) PCL_LAUNCH_POSE_MB(KK, false)                                                                      \
        else                                                                                                                \
        {                                                                                                                   \
            { int rc__ = ensure_smem(c, (const void*)pose_kernel<KK>, smem); if (rc__) return rc__; }                       \
            pose_kernel<KK><<<grid, wpb * 32, smem, c->compute>>>(p);                                                       \
        }                                                                                                                   \
    }
        switch (b->K)
        {
            case 1: PCL_LAUNCH_POSE(1) break;
            case 2: PCL_LAUNCH_POSE(2) break;
            case 3: PCL_LAUNCH_POSE(3) break;
            default: PCL_LAUNCH_POSE(4) break;
        }
#undef PCL_LAUNCH_POSE_S
#undef PCL_LAUNCH_POSE_SD
#undef PCL_LAUNCH_POSE_MB
#undef PCL_LAUNCH_POSE
        c->launches++;
        PCL_POST_LAUNCH(c, c->compute, "pose kernel");
        PCL_CK(c, cudaEventRecord(c->ev_pose, c->compute));
        b->posed = true;
        return PCL_OK;
    }

    int pcl_batch_pose(pcl_batch* b)
    {
        if (!b)
            return PCL_ERR_INVALID;
        return launch_pose(b, 0.f, 0);
    }
    int pcl_batch_tick(pcl_batch* b, float dt, int32_t mode)
    {
        if (!b || (mode != PCL_ADVANCE_FIRST_ONLY && mode != PCL_ADVANCE_ALL))
            return fail(b ? b->ctx : nullptr, PCL_ERR_INVALID, "pcl_batch_tick: bad mode");
        return launch_pose(b, dt, mode);
    }

    // ------------------------------------------------------------------------------------------
    // one frame of a SMALL batch in one call (the reference's everyday case: a handful of AnimationComponents)
    // ------------------------------------------------------------------------------------------
    static size_t frame_io_bytes(const pcl_batch* b, int J) { return (size_t)b->N * b->K * 4 * 5 + (size_t)b->N * J * 48 + 64; }

    int pcl_batch_frame(pcl_batch* b, const int32_t* clip_ids, const float* phase, const float* weight, const int32_t* mask_ids, float* host12)
    {
        if (!b)
            return PCL_ERR_INVALID;
        pcl_ctx*       c = b->ctx;
        const SkelRec& s = c->skels[b->skel];
        if (frame_io_bytes(b, s.J) > (size_t)PCL_FRAME_MAX_BYTES)
            return fail(c, PCL_ERR_UNSUPPORTED, "pcl_batch_frame: the batch is too large for the one-call frame path (PCL_FRAME_MAX_BYTES); use pcl_batch_set_frame_async + pcl_batch_tick + pcl_batch_download_palette3x4_async");
        PCL_CK(c, cudaSetDevice(c->device));
        const size_t n = (size_t)b->N * b->K;
        if (!b->h_io)
        {
            if (!clip_ids || !phase)
                return fail(c, PCL_ERR_INVALID, "pcl_batch_frame: the first call on a batch must pass clip_ids and phase");
            PCL_CK(c, cudaHostAlloc((void**)&b->h_io, frame_io_bytes(b, s.J), cudaHostAllocDefault)); // unified addressing: the device sees it
            PCL_CK(c, cudaEventCreateWithFlags(&b->ev_frame, cudaEventDisableTiming));
            float*   wl = reinterpret_cast<float*>(b->h_io) + 2 * n;
            int32_t* mk = reinterpret_cast<int32_t*>(b->h_io) + 4 * n;
            for (size_t i = 0; i < n; i++)
                wl[i] = 1.f, wl[n + i] = 1.f, mk[i] = -1; // weight, clip length, mask: the defaults of pcl_batch_create
        }
        else if (b->frame_pending)
        {
            PCL_CK(c, cudaEventSynchronize(b->ev_frame)); // the previous frame's kernel has read its inputs and written its rows
            b->frame_pending = false;
        }
        int32_t* h_clip  = reinterpret_cast<int32_t*>(b->h_io);
        float*   h_phase = reinterpret_cast<float*>(b->h_io) + n;
        float*   h_wgt   = reinterpret_cast<float*>(b->h_io) + 2 * n;
        float*   h_len   = reinterpret_cast<float*>(b->h_io) + 3 * n;
        int32_t* h_mask  = reinterpret_cast<int32_t*>(b->h_io) + 4 * n;
        float4*  h_out   = reinterpret_cast<float4*>(b->h_io + ((n * 20 + 63) & ~(size_t)63));
        const int       K = b->K;
        const long long N = b->N;
        for (long long i = 0; i < N; i++) // [N][K] -> the kernels' [K][N]
            for (int k = 0; k < K; k++)
            {
                if (clip_ids)
                {
                    const int32_t v = clip_ids[i * K + k];
                    if (v < 0 || v >= (int)c->clips.size() || c->clips[v].skel != b->skel || c->clips[v].released)
                        return fail(c, PCL_ERR_INVALID, "pcl_batch_frame: clip id does not belong to the batch's skeleton");
                    h_clip[(size_t)k * N + i] = v;
                }
                if (phase)
                    h_phase[(size_t)k * N + i] = phase[i * K + k];
                if (weight)
                    h_wgt[(size_t)k * N + i] = weight[i * K + k];
                if (mask_ids)
                {
                    int32_t v = mask_ids[i * K + k];
                    if (v >= 0)
                    {
                        if (v >= (int)c->masks.size() || c->masks[v].skel != b->skel)
                            return fail(c, PCL_ERR_INVALID, "pcl_batch_frame: mask id does not belong to the batch's skeleton");
                        v = c->masks[v].local;
                    }
                    else
                        v = -1;
                    h_mask[(size_t)k * N + i] = v;
                }
            }
        // A few characters' rows are written by the pose kernel itself, straight into the pinned block (48-byte pieces over
        // the host link: fine for kilobytes - one launch, no DMA set-up; measured 6 GB/s, so not beyond).  Larger frames leave
        // through one strided DMA out of the device palette, like pcl_batch_download_palette3x4.
        const bool direct = (size_t)N * s.J * 48 <= PCL_FRAME_DIRECT_BYTES;
        FrameIO    io {h_clip, h_phase, h_wgt, h_len, h_mask, direct ? h_out : nullptr, false};
        int        rc = launch_pose(b, 0.f, 0, &io);
        if (rc != PCL_OK)
            return rc;
        if (!direct)
        {
            const size_t row = (size_t)s.J * 48;
            PCL_CK(c, cudaMemcpy2DAsync(h_out, row, reinterpret_cast<const char*>(b->d_palette) + 48, (size_t)s.pitch * 48, row, (size_t)N,
                                        cudaMemcpyDeviceToHost, c->compute));
        }
        else if (!io.out12_written) // the launcher could not use the kernel that writes the rows itself: one more small launch
        {
            const long long total = N * s.J * 3;
            export3x4_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->compute>>>(b->d_palette, h_out, 0, N, s.J, s.pitch);
            c->launches++;
            PCL_POST_LAUNCH(c, c->compute, "export3x4_kernel (frame block)");
        }
        PCL_CK(c, cudaEventRecord(b->ev_frame, c->compute));
        b->frame_pending = true;
        return host12 ? pcl_batch_frame_result(b, host12) : PCL_OK;
    }

    int pcl_batch_frame_result(pcl_batch* b, float* host12)
    {
        if (!b || !host12)
            return PCL_ERR_INVALID;
        pcl_ctx* c = b->ctx;
        if (!b->h_io)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_frame_result: pcl_batch_frame has not run");
        if (b->frame_pending)
        {
            PCL_CK(c, cudaEventSynchronize(b->ev_frame));
            b->frame_pending = false;
        }
        const SkelRec& s = c->skels[b->skel];
        const size_t   n = (size_t)b->N * b->K;
        std::memcpy(host12, b->h_io + ((n * 20 + 63) & ~(size_t)63), (size_t)b->N * s.J * 48);
        return PCL_OK;
    }

    int pcl_batch_keep_derived(pcl_batch* b, int32_t enable)
    {
        if (!b)
            return PCL_ERR_INVALID;
        pcl_ctx* c = b->ctx;
        PCL_CK(c, cudaSetDevice(c->device));
        PCL_CK(c, cudaStreamSynchronize(c->compute));
        if (enable && !b->d_derived)
            PCL_CK(c, cudaMalloc((void**)&b->d_derived, (size_t)b->N * c->skels[b->skel].J * 40));
        if (!enable && b->d_derived)
        {
            cudaFree(b->d_derived);
            b->d_derived = nullptr;
        }
        return PCL_OK;
    }

    int pcl_skinbuf_create(pcl_ctx* c, pcl_mesh mesh, int64_t capacity, int32_t flags, pcl_skinbuf** out)
    {
        if (!c || !out || mesh < 0 || mesh >= (int)c->meshes.size() || c->meshes[mesh].released || capacity < 1 ||
            (flags != PCL_SKIN_POS && flags != PCL_SKIN_POS_NRM_TAN))
            return fail(c, PCL_ERR_INVALID, "pcl_skinbuf_create: bad argument");
        PCL_CK(c, cudaSetDevice(c->device));
        pcl_skinbuf* s = new pcl_skinbuf();
        s->ctx = c, s->mesh = mesh, s->capacity = capacity, s->flags = flags;
        s->V = c->meshes[mesh].V, s->Vp = c->meshes[mesh].Vp;
        cudaError_t e = cudaMalloc((void**)&s->d_pos, (size_t)capacity * s->Vp * 12);
        if (e == cudaSuccess && flags == PCL_SKIN_POS_NRM_TAN)
            e = cudaMalloc((void**)&s->d_nt, (size_t)capacity * s->Vp * 24);
        if (e != cudaSuccess)
        {
            cudaFree(s->d_pos);
            delete s;
            (void)cudaGetLastError();
            return fail(c, PCL_ERR_OOM, std::string("pcl_skinbuf_create: ") + cudaGetErrorString(e));
        }
        s->has_tmaps = true;
        for (int i = 0; i < 3; i++)
        {
            const unsigned rows = 32u >> i;
            s->has_tmaps = s->has_tmaps && make_stream_tmap(&s->tm_pos[i], s->d_pos, capacity, (long long)s->Vp * 3, rows) &&
                           (!s->d_nt || make_stream_tmap(&s->tm_nt[i], s->d_nt, capacity, (long long)s->Vp * 6, rows));
            if (!s->d_nt)
                s->tm_nt[i] = s->tm_pos[i];
        }
        *out = s;
        return PCL_OK;
    }
    int pcl_skinbuf_destroy(pcl_skinbuf* s)
    {
        if (!s)
            return PCL_ERR_INVALID;
        cudaSetDevice(s->ctx->device);
        cudaStreamSynchronize(s->ctx->compute);
        cudaFree(s->d_pos), cudaFree(s->d_nt);
        delete s;
        return PCL_OK;
    }

    int pcl_batch_skin(pcl_batch* b, pcl_mesh mesh, pcl_skinbuf* out, int64_t first, int64_t count, int64_t slot_first)
    {
        if (!b || !out)
            return PCL_ERR_INVALID;
        pcl_ctx* c = b->ctx;
        if (out->ctx != c || out->mesh != mesh || mesh < 0 || mesh >= (int)c->meshes.size() || c->meshes[mesh].released)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_skin: skin buffer was created for another mesh / context");
        if (first < 0 || count < 1 || first + count > b->N || slot_first < 0 || slot_first + count > out->capacity)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_skin: range outside the batch or the ring");
        if (!b->posed)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_skin: pcl_batch_pose / pcl_batch_tick has not run");
        const SkelRec& s = c->skels[b->skel];
        const MeshRec& m = c->meshes[mesh];
        if (m.max_slot > s.J)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_skin: mesh references palette slots the skeleton does not have");
        PCL_CK(c, cudaSetDevice(c->device));
        const bool nt = out->flags == PCL_SKIN_POS_NRM_TAN;
        // crowd form (lane = instance) whenever a warp's 32 palettes fit in shared memory and there are
        // enough instances to fill the lanes; otherwise the lane = vertex kernel
        {
            const int    prow   = (s.pitch * 3) | 1; // odd number of 16-byte units per palette row
            int          warps  = c->opt_lbs_warps ? c->opt_lbs_warps : 8; // measured: 8 warps + whole-chunk flush beats 12 warps + half flush
            bool         half   = false;
            if (warps == 12 && lbs_i_smem_bytes(prow, 12, false) > (size_t)c->max_smem_optin)
                half = true; // 12 warps only fit with the smaller normal/tangent staging rows
            if (warps == 12 && half && lbs_i_smem_bytes(prow, 12, true) > (size_t)c->max_smem_optin)
                warps = 8, half = false;
            const size_t smem_i = lbs_i_smem_bytes(prow, warps, half);
            const bool   fits   = m.d_vrec && smem_i <= (size_t)c->max_smem_optin;
            int          var    = c->opt_lbs_variant;
            if (var == 2 && !fits)
                return fail(c, PCL_ERR_UNSUPPORTED, "pcl_batch_skin: lbs_variant 2 needs 32 palettes in shared memory");
            // lbs_cta_kernel: the largest instance tile (32, 16, 8) whose palettes leave room for two staging tiles
            int cT = 0;
            for (int cand = 32; cand >= 8 && !cT; cand >>= 1)
                if ((c->opt_lbs_cta_tile == 0 || c->opt_lbs_cta_tile == cand) && lbs_c_smem_bytes(prow, nt, 2, cand) <= (size_t)c->max_smem_optin)
                    cT = cand;
            // ... or, when whole palettes force a small tile, the compact form: per block of 1024 vertices only the rows that block
            // references (a coherent 256-bone character: a few dozen), which brings the tile back to 32 or 16 instances
            bool compact = false;
            int  prow_c  = prow;
            if (m.max_slot * 3 <= 0xffff && out->has_tmaps && c->opt_lbs_compact != 0 && (cT < 32 || c->opt_lbs_compact == 1))
            {
                MeshRec& mm = c->meshes[mesh];
                { int rc__ = analyse_blocks(c, mm); if (rc__) return rc__; }
                prow_c  = (mm.blk_mmax * 3) | 1;
                int cTc = 0;
                for (int cand = 32; cand >= 8 && !cTc; cand >>= 1)
                    if ((c->opt_lbs_cta_tile == 0 || c->opt_lbs_cta_tile == cand) && lbs_c_smem_bytes(prow_c, nt, 2, cand) <= (size_t)c->max_smem_optin)
                        cTc = cand;
                if (cTc > cT || (c->opt_lbs_compact == 1 && cTc >= cT && cTc > 0))
                    compact = true, cT = cTc;
                else
                    prow_c = prow;
            }
            const bool fits3 = cT && m.max_slot * 3 <= 0xffff && out->has_tmaps;
            if (var < 0)
            {
                // measured (tools/lbs_sweep.py; player, 120- and 256-bone skeletons): with the finer vertex splits small launches
                // get, the TMA-store kernel is level with the lane-per-vertex kernel at 16 instances and ahead from 32 (128
                // player instances: 19 us against 30 us); lbs_inst_kernel (no tensor maps available) keeps its ~35 us floor
                var = (fits3 && count >= 32) ? 3 : ((fits && !fits3 && count >= 160) ? 2 : 0);
            }
            if (var == 3)
            {
                if (!fits3)
                    return fail(c, PCL_ERR_UNSUPPORTED, "pcl_batch_skin: lbs_variant 3 needs 8 palettes in shared memory and tensor maps");
                const int    ctiles = lbs_c_smem_bytes(prow_c, nt, 3, cT) <= (size_t)c->max_smem_optin ? 3 : 2;
                const size_t smem_c = lbs_c_smem_bytes(prow_c, nt, ctiles, cT);
                const int    ti     = cT == 32 ? 0 : (cT == 16 ? 1 : 2);
                const long long full = (count + cT - 1) / cT; // tiles; the last one may be ragged
                // rows the stores may touch: the ring's capacity for whole tiles; for a ragged last tile, maps of this launch's
                // own height, so that the TMA drops the rows of the instances that do not exist
                CUtensorMap tmp_pos = out->tm_pos[ti], tmp_nt = out->tm_nt[ti];
                if (count % cT != 0)
                {
                    if (!make_stream_tmap(&tmp_pos, out->d_pos, slot_first + count, (long long)out->Vp * 3, (unsigned)cT) ||
                        (nt && !make_stream_tmap(&tmp_nt, out->d_nt, slot_first + count, (long long)out->Vp * 6, (unsigned)cT)))
                        return fail(c, PCL_ERR_CUDA, "pcl_batch_skin: cuTensorMapEncodeTiled failed");
                    if (!nt)
                        tmp_nt = tmp_pos;
                }
                {
                    MeshRec& mm = c->meshes[mesh];
                    { int rc__ = compact ? ensure_vrec_cta_compact(c, mm, cT) : ensure_vrec_cta(c, mm, cT); if (rc__) return rc__; }
                    LbsCParams q {};
                    q.vrec = compact ? mm.d_vrec_blk[ti] : mm.d_vrec_cta[ti], q.palette = b->d_palette;
                    if (compact)
                        q.blk_sc = LBS_C_BLK_VERTS / lbs_c_scv(cT), q.blk_cap = mm.blk_cap, q.blk_slots = mm.d_blk_slots, q.blk_cnt = mm.d_blk_cnt;
                    q.first = first, q.slot_first = slot_first, q.count = count, q.n_tiles = (int32_t)full;
                    q.Vp = m.Vp, q.n_sc = (m.Vp + lbs_c_scv(cT) - 1) / lbs_c_scv(cT), q.pitch = s.pitch, q.prow = prow_c;
                    int vs = c->opt_lbs_vsplit;
                    if (vs == 0)
                    {
                        // at least 32 superchunks per CTA: finer splits measured slower (palette reloads, pipeline fills) - unless the
                        // launch is so small that it would leave SMs idle; then down to 8 superchunks per CTA
                        int min_sc = 32;
                        while (min_sc > 8 && full * std::min<long long>(64, q.n_sc / min_sc) < 2LL * c->sm_count)
                            min_sc >>= 1;
                        const int max_vs = (int)std::max<long long>(1, std::min<long long>(min_sc == 32 ? 16 : 64, q.n_sc / min_sc));
                        double    best   = -1.0;
                        for (int cand = 1; cand <= max_vs; cand++)
                        {
                            const double waves = (double)(full * cand) / c->sm_count;
                            const double eff   = waves / std::ceil(waves);
                            if (eff > best + 0.02)
                                best = eff, vs = cand;
                        }
                    }
                    q.vsplit = std::max(1, std::min(vs, q.n_sc));
                    q.per_sc = (q.n_sc + q.vsplit - 1) / q.vsplit;
                    if (compact) // runs start on block boundaries: choose the run length in whole blocks
                    {
                        const int n_blk = (q.n_sc + q.blk_sc - 1) / q.blk_sc;
                        if (c->opt_lbs_vsplit == 0)
                        {
                            double best = -1.0;
                            int    best_k = n_blk;
                            for (int k = n_blk; k >= 1; k--) // longer runs first: ties go to fewer palette re-stagings per CTA
                            {
                                const int    splits = (n_blk + k - 1) / k;
                                const double waves  = (double)(full * splits) / c->sm_count;
                                // the last split of every tile may be shorter: count the work, not the CTAs
                                const double eff = ((double)full * n_blk) / (std::ceil(waves) * c->sm_count * (double)k);
                                if (eff > best + 0.02)
                                    best = eff, best_k = k;
                            }
                            q.per_sc = best_k * q.blk_sc;
                        }
                        else
                            q.per_sc = (q.per_sc + q.blk_sc - 1) / q.blk_sc * q.blk_sc;
                        q.vsplit = (q.n_sc + q.per_sc - 1) / q.per_sc;
                    }
                    const unsigned grid = (unsigned)(full * q.vsplit);
#define PCL_LAUNCH_LBSC_C(NT_, TL_, T_, C_)                                                                                 \
    {                                                                                                                       \
        { int rc__ = ensure_smem(c, (const void*)lbs_cta_kernel<NT_, TL_, T_, C_>, smem_c); if (rc__) return rc__; }         \
        lbs_cta_kernel<NT_, TL_, T_, C_><<<grid, LBS_C_THREADS, smem_c, c->compute>>>(q, tmp_pos, tmp_nt);                    \
    }
#define PCL_LAUNCH_LBSC(NT_, TL_, T_)                                                                                        \
    {                                                                                                                       \
        if (compact) PCL_LAUNCH_LBSC_C(NT_, TL_, T_, true)                                                                  \
        else PCL_LAUNCH_LBSC_C(NT_, TL_, T_, false)                                                                         \
    }
#define PCL_LAUNCH_LBSC_T(T_)                                                                                               \
    {                                                                                                                       \
        if (nt && ctiles == 3) PCL_LAUNCH_LBSC(true, 3, T_)                                                                 \
        else if (nt) PCL_LAUNCH_LBSC(true, 2, T_)                                                                           \
        else if (ctiles == 3) PCL_LAUNCH_LBSC(false, 3, T_)                                                                 \
        else PCL_LAUNCH_LBSC(false, 2, T_)                                                                                  \
    }
                    if (cT == 32) PCL_LAUNCH_LBSC_T(32)
                    else if (cT == 16) PCL_LAUNCH_LBSC_T(16)
                    else PCL_LAUNCH_LBSC_T(8)
#undef PCL_LAUNCH_LBSC_T
#undef PCL_LAUNCH_LBSC
#undef PCL_LAUNCH_LBSC_C
                    c->launches++;
                    PCL_POST_LAUNCH(c, c->compute, "lbs_cta_kernel");
                }
                return PCL_OK;
            }
            if (var == 2)
            {
                LbsIParams q {};
                q.vrec = m.d_vrec, q.palette = b->d_palette, q.out_pos = out->d_pos, q.out_nt = out->d_nt;
                q.first = first, q.count = count, q.slot_first = slot_first;
                q.Vp = m.Vp, q.n_chunks = m.Vp / LBS_I_CHUNK, q.pitch = s.pitch, q.prow = prow;
                const long long tiles = (count + 31) / 32;
                int             vs    = c->opt_lbs_vsplit;
                if (vs == 0)
                {
                    // One CTA per SM is resident, so the grid runs in waves of sm_count CTAs: pick the vertex split
                    // whose last wave is fullest (ties: fewer splits, i.e. longer runs - every warp restarts cold),
                    // among splits that leave each warp at least 16 chunks and, for small launches, give every SM work.
                    const int max_vs = (int)std::max<long long>(1, std::min<long long>(16, q.n_chunks / (warps * 16)));
                    double    best   = -1.0;
                    for (int cand = 1; cand <= max_vs; cand++)
                    {
                        const double waves = (double)(tiles * cand) / c->sm_count;
                        const double eff   = waves / std::ceil(waves);
                        if (eff > best + 0.02)
                            best = eff, vs = cand;
                    }
                }
                vs       = std::max(1, std::min(vs, q.n_chunks));
                q.vsplit = vs;
                const unsigned grid = (unsigned)(tiles * vs);
#define PCL_LAUNCH_LBSI(NT_, W_, H_)                                                                                               \
    {                                                                                                                              \
        { int rc__ = ensure_smem(c, (const void*)lbs_inst_kernel<NT_, W_, H_>, smem_i); if (rc__) return rc__; }                  \
        lbs_inst_kernel<NT_, W_, H_><<<grid, W_ * 32, smem_i, c->compute>>>(q);                                                    \
    }
                if (nt && warps == 12 && half) PCL_LAUNCH_LBSI(true, 12, true)
                else if (nt && warps == 12) PCL_LAUNCH_LBSI(true, 12, false)
                else if (nt) PCL_LAUNCH_LBSI(true, 8, false)
                else if (warps == 12 && half) PCL_LAUNCH_LBSI(false, 12, true)
                else if (warps == 12) PCL_LAUNCH_LBSI(false, 12, false)
                else PCL_LAUNCH_LBSI(false, 8, false)
#undef PCL_LAUNCH_LBSI
                c->launches++;
                PCL_POST_LAUNCH(c, c->compute, "lbs_inst_kernel");
                return PCL_OK;
            }
        }
        const int lv_variant = c->opt_lbs_variant == 1 ? 1 : 0;
        LbsParams p {};
        p.pos = m.d_pos, p.nrm = m.d_nrm, p.tan = m.d_tan, p.idx = m.d_idx, p.w = m.d_w;
        p.palette = b->d_palette, p.out_pos = out->d_pos, p.out_nt = out->d_nt;
        p.first = first, p.count = count, p.slot_first = slot_first;
        p.V = m.V, p.Vp = m.Vp, p.n_vtiles = m.Vt / LBS_VT, p.pitch = s.pitch;
        int it = c->opt_lbs_inst_tile;
        if (it == 0)
            it = (int)std::max<size_t>(1, std::min<size_t>(16, (48 * 1024) / ((size_t)s.pitch * 48)));
        while (it > 1 && lbs_smem_bytes(it, s.pitch) > (size_t)c->max_smem_optin)
            it--;
        const size_t smem = lbs_smem_bytes(it, s.pitch);
        if (smem > (size_t)c->max_smem_optin)
            return fail(c, PCL_ERR_UNSUPPORTED, "pcl_batch_skin: one palette does not fit in shared memory");
        p.inst_tile = it;
        const long long tiles = (count + it - 1) / it;
        int             vs    = c->opt_lbs_vsplit;
        if (vs == 0)
            vs = (int)std::max<long long>(1, std::min<long long>(p.n_vtiles, (16LL * c->sm_count + tiles - 1) / tiles));
        vs       = std::min(vs, p.n_vtiles);
        p.vsplit = vs;
        const unsigned grid = (unsigned)(tiles * vs);
#define PCL_LAUNCH_LBS(NT_, TMA_)                                                                                          \
    {                                                                                                                      \
        { int rc__ = ensure_smem(c, (const void*)lbs_kernel<NT_, TMA_>, smem); if (rc__) return rc__; }                      \
        lbs_kernel<NT_, TMA_><<<grid, LBS_THREADS, smem, c->compute>>>(p);                                                 \
    }
        if (nt && lv_variant == 0) PCL_LAUNCH_LBS(true, false)
        else if (nt) PCL_LAUNCH_LBS(true, true)
        else if (lv_variant == 0) PCL_LAUNCH_LBS(false, false)
        else PCL_LAUNCH_LBS(false, true)
#undef PCL_LAUNCH_LBS
        c->launches++;
        PCL_POST_LAUNCH(c, c->compute, "lbs_kernel");
        return PCL_OK;
    }

    // ------------------------------------------------------------------------------------------
    // readback
    // ------------------------------------------------------------------------------------------
    static int export_palette(pcl_batch* b, int64_t first, int64_t count, float* host16, bool async)
    {
        if (!b || !host16 || first < 0 || count < 1 || first + count > b->N)
            return fail(b ? b->ctx : nullptr, PCL_ERR_INVALID, "pcl_batch_download_palette: range outside the batch");
        pcl_ctx* c = b->ctx;
        if (!b->posed)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_download_palette: pcl_batch_pose / pcl_batch_tick has not run");
        PCL_CK(c, cudaSetDevice(c->device));
        const SkelRec& s = c->skels[b->skel];
        // small synchronous read-backs (a handful of characters: the reference's everyday case): the export kernel writes straight
        // into a pinned, device-visible staging block of the context and the CPU copies it out - no device staging buffer and
        // no separate DMA launch (one instance: 35 -> 28 us for tick + download, tools/tick_latency.py)
        const size_t bytes = (size_t)count * s.J * 64;
        if (!async && bytes <= PCL_SMALL_EXPORT_BYTES)
        {
            if (!c->h_small)
                PCL_CK(c, cudaHostAlloc(&c->h_small, PCL_SMALL_EXPORT_BYTES, cudaHostAllocDefault)); // unified addressing: the device sees it
            PCL_CK(c, cudaStreamWaitEvent(c->copy, c->ev_pose, 0));
            const long long total = count * s.J * 4;
            export_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->copy>>>(b->d_palette, (float4*)c->h_small, first, count, s.J, s.pitch);
            c->launches++;
            PCL_POST_LAUNCH(c, c->copy, "export_kernel");
            PCL_CK(c, cudaStreamSynchronize(c->copy));
            std::memcpy(host16, c->h_small, bytes);
            return PCL_OK;
        }
        if (count > b->export_cap)
        {
            PCL_CK(c, cudaStreamSynchronize(c->copy));
            cudaFree(b->d_export);
            b->d_export = nullptr, b->export_cap = 0;
            PCL_CK(c, cudaMalloc((void**)&b->d_export, (size_t)count * s.J * 64));
            b->export_cap = count;
        }
        PCL_CK(c, cudaStreamWaitEvent(c->copy, c->ev_pose, 0));
        const long long total = count * s.J * 4;
        export_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->copy>>>(b->d_palette, b->d_export, first, count, s.J, s.pitch);
        c->launches++;
        PCL_POST_LAUNCH(c, c->copy, "export_kernel");
        PCL_CK(c, cudaEventRecord(c->ev_copy, c->copy));
        c->copy_pending = true;
        PCL_CK(c, cudaMemcpyAsync(host16, b->d_export, (size_t)count * s.J * 64, cudaMemcpyDeviceToHost, c->copy));
        if (!async)
            PCL_CK(c, cudaStreamSynchronize(c->copy));
        return PCL_OK;
    }
    int pcl_batch_download_palette(pcl_batch* b, int64_t first, int64_t count, float* host16)
    {
        return export_palette(b, first, count, host16, false);
    }
    int pcl_batch_download_palette_async(pcl_batch* b, int64_t first, int64_t count, float* host16)
    {
        return export_palette(b, first, count, host16, true);
    }

    // Compact read-back (rows 0..2).  The three ways the bytes can travel are kept selectable ("export_mode") because which
    // one is best depends on the host link, not on the GPU; measurements in profiles/README.md.
    static int export_palette3x4(pcl_batch* b, int64_t first, int64_t count, float* host12, bool async)
    {
        if (!b || !host12 || first < 0 || count < 1 || first + count > b->N)
            return fail(b ? b->ctx : nullptr, PCL_ERR_INVALID, "pcl_batch_download_palette3x4: range outside the batch");
        pcl_ctx* c = b->ctx;
        if (!b->posed)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_download_palette3x4: pcl_batch_pose / pcl_batch_tick has not run");
        PCL_CK(c, cudaSetDevice(c->device));
        const SkelRec& s     = c->skels[b->skel];
        const size_t   row   = (size_t)s.J * 48;
        const size_t   bytes = (size_t)count * row;
        int            mode  = c->opt_export_mode ? c->opt_export_mode : 1;
        float4*        dev_view = nullptr; // device-side address of host12 when it is pinned
        if (mode == 2)
        {
            cudaPointerAttributes at {};
            if (cudaPointerGetAttributes(&at, host12) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer)
                dev_view = (float4*)at.devicePointer;
            else
            {
                (void)cudaGetLastError();
                mode = 3;
            }
        }
        PCL_CK(c, cudaStreamWaitEvent(c->copy, c->ev_pose, 0));
        const long long total = count * s.J * 3;
        if (mode == 1)
        {
            // slots 1..J of every palette row are the J matrices in bones_map order: one 2-D copy, rows of J * 48 bytes
            const char* src = reinterpret_cast<const char*>(b->d_palette) + (size_t)first * s.pitch * 48 + 48;
            PCL_CK(c, cudaMemcpy2DAsync(host12, row, src, (size_t)s.pitch * 48, row, (size_t)count, cudaMemcpyDeviceToHost, c->copy));
            PCL_CK(c, cudaEventRecord(c->ev_copy, c->copy)); // the palette is being read until the copy is through
            c->copy_pending = true;
        }
        else if (mode == 2)
        {
            export3x4_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->copy>>>(b->d_palette, dev_view, first, count, s.J, s.pitch);
            c->launches++;
            PCL_POST_LAUNCH(c, c->copy, "export3x4_kernel (pinned destination)");
            PCL_CK(c, cudaEventRecord(c->ev_copy, c->copy));
            c->copy_pending = true;
        }
        else
        {
            if (count > b->export12_cap)
            {
                PCL_CK(c, cudaStreamSynchronize(c->copy));
                cudaFree(b->d_export12);
                b->d_export12 = nullptr, b->export12_cap = 0;
                PCL_CK(c, cudaMalloc((void**)&b->d_export12, bytes));
                b->export12_cap = count;
            }
            export3x4_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->copy>>>(b->d_palette, b->d_export12, first, count, s.J, s.pitch);
            c->launches++;
            PCL_POST_LAUNCH(c, c->copy, "export3x4_kernel");
            PCL_CK(c, cudaEventRecord(c->ev_copy, c->copy)); // the next pose only waits for the kernel, not for the DMA
            c->copy_pending = true;
            PCL_CK(c, cudaMemcpyAsync(host12, b->d_export12, bytes, cudaMemcpyDeviceToHost, c->copy));
        }
        if (!async)
            PCL_CK(c, cudaStreamSynchronize(c->copy));
        return PCL_OK;
    }
    int pcl_batch_download_palette3x4_async(pcl_batch* b, int64_t first, int64_t count, float* host12)
    {
        return export_palette3x4(b, first, count, host12, true);
    }
    int pcl_batch_download_palette3x4(pcl_batch* b, int64_t first, int64_t count, float* host12)
    {
        return export_palette3x4(b, first, count, host12, false);
    }

    int pcl_batch_download_derived(pcl_batch* b, int64_t first, int64_t count, float* host10)
    {
        if (!b || !host10 || first < 0 || count < 1 || first + count > b->N)
            return fail(b ? b->ctx : nullptr, PCL_ERR_INVALID, "pcl_batch_download_derived: range outside the batch");
        pcl_ctx* c = b->ctx;
        if (!b->d_derived || !b->posed)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_download_derived: enable pcl_batch_keep_derived before posing");
        PCL_CK(c, cudaSetDevice(c->device));
        PCL_CK(c, cudaStreamSynchronize(c->compute));
        const int J = c->skels[b->skel].J;
        PCL_CK(c, cudaMemcpy(host10, b->d_derived + (size_t)first * J * 10, (size_t)count * J * 40, cudaMemcpyDeviceToHost));
        return PCL_OK;
    }

    int pcl_batch_palette_dev(pcl_batch* b, const void** dptr, int32_t* pitch_slots, int32_t* floats_per_slot)
    {
        if (!b || !dptr)
            return PCL_ERR_INVALID;
        *dptr = b->d_palette;
        if (pitch_slots)
            *pitch_slots = b->ctx->skels[b->skel].pitch;
        if (floats_per_slot)
            *floats_per_slot = 12;
        return PCL_OK;
    }

    int pcl_skinbuf_dev(pcl_skinbuf* s, const void** pos_dev, const void** nt_dev, int32_t* Vp)
    {
        if (!s)
            return PCL_ERR_INVALID;
        if (pos_dev)
            *pos_dev = s->d_pos;
        if (nt_dev)
            *nt_dev = s->d_nt;
        if (Vp)
            *Vp = s->Vp;
        return PCL_OK;
    }

    int pcl_skinbuf_download(pcl_skinbuf* s, int64_t slot_first, int64_t count, float* host_pos, float* host_nt)
    {
        if (!s || slot_first < 0 || count < 1 || slot_first + count > s->capacity)
            return fail(s ? s->ctx : nullptr, PCL_ERR_INVALID, "pcl_skinbuf_download: range outside the ring");
        pcl_ctx* c = s->ctx;
        PCL_CK(c, cudaSetDevice(c->device));
        PCL_CK(c, cudaStreamSynchronize(c->compute));
        if (host_pos)
            PCL_CK(c, cudaMemcpy2D(host_pos, (size_t)s->V * 12, s->d_pos + (size_t)slot_first * s->Vp * 3, (size_t)s->Vp * 12,
                                   (size_t)s->V * 12, (size_t)count, cudaMemcpyDeviceToHost));
        if (host_nt)
        {
            if (!s->d_nt)
                return fail(c, PCL_ERR_INVALID, "pcl_skinbuf_download: the ring was created without normals/tangents");
            PCL_CK(c, cudaMemcpy2D(host_nt, (size_t)s->V * 24, s->d_nt + (size_t)slot_first * s->Vp * 6, (size_t)s->Vp * 24,
                                   (size_t)s->V * 24, (size_t)count, cudaMemcpyDeviceToHost));
        }
        return PCL_OK;
    }

    // ------------------------------------------------------------------------------------------
    // render hand-off
    // ------------------------------------------------------------------------------------------
    int pcl_renderbuf_create(pcl_ctx* c, int64_t draws, pcl_renderbuf** out)
    {
        if (!c || !out || draws < 1)
            return fail(c, PCL_ERR_INVALID, "pcl_renderbuf_create: bad argument");
        PCL_CK(c, cudaSetDevice(c->device));
        pcl_renderbuf* r = new pcl_renderbuf();
        r->ctx = c, r->draws = draws;
        const size_t bytes = (size_t)draws * RENDER_MAX_INSTANCES * RENDER_MAX_JOINTS * 64;
        cudaError_t  e     = cudaMalloc((void**)&r->d_mats, bytes);
        if (e != cudaSuccess)
        {
            delete r;
            (void)cudaGetLastError();
            return fail(c, PCL_ERR_OOM, std::string("pcl_renderbuf_create: ") + cudaGetErrorString(e));
        }
        PCL_CK(c, cudaMemsetAsync(r->d_mats, 0, bytes, c->compute));
        *out = r;
        return PCL_OK;
    }
    int pcl_renderbuf_destroy(pcl_renderbuf* r)
    {
        if (!r)
            return PCL_ERR_INVALID;
        cudaSetDevice(r->ctx->device);
        cudaStreamSynchronize(r->ctx->compute);
        cudaFree(r->d_mats);
        delete r;
        return PCL_OK;
    }
    int pcl_batch_export_render(pcl_batch* b, int64_t first, int64_t count, pcl_renderbuf* rb, int64_t draw_first)
    {
        if (!b || !rb || rb->ctx != b->ctx)
            return fail(b ? b->ctx : nullptr, PCL_ERR_INVALID, "pcl_batch_export_render: bad handle");
        pcl_ctx* c = b->ctx;
        if (first < 0 || count < 1 || first + count > b->N || draw_first < 0 ||
            draw_first + (count + RENDER_MAX_INSTANCES - 1) / RENDER_MAX_INSTANCES > rb->draws)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_export_render: range outside the batch or the render buffer");
        if (!b->posed)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_export_render: pcl_batch_pose / pcl_batch_tick has not run");
        const SkelRec& s = c->skels[b->skel];
        if (s.J + 1 > RENDER_MAX_JOINTS)
            return fail(c, PCL_ERR_UNSUPPORTED, "pcl_batch_export_render: more than 1023 bones (m_mesh_vertex_blending_max_joint_count)");
        PCL_CK(c, cudaSetDevice(c->device));
        const long long total = count * (s.J + 1) * 4;
        render_export_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->compute>>>(b->d_palette, rb->d_mats, first, count, s.J + 1, s.pitch, draw_first);
        c->launches++;
        PCL_POST_LAUNCH(c, c->compute, "render_export_kernel");
        return PCL_OK;
    }
    int pcl_renderbuf_dev(pcl_renderbuf* r, const void** dptr)
    {
        if (!r || !dptr)
            return PCL_ERR_INVALID;
        *dptr = r->d_mats;
        return PCL_OK;
    }
    int pcl_renderbuf_download(pcl_renderbuf* r, int64_t draw, float* host)
    {
        if (!r || !host || draw < 0 || draw >= r->draws)
            return fail(r ? r->ctx : nullptr, PCL_ERR_INVALID, "pcl_renderbuf_download: draw outside the buffer");
        pcl_ctx* c = r->ctx;
        PCL_CK(c, cudaSetDevice(c->device));
        PCL_CK(c, cudaStreamSynchronize(c->compute));
        const size_t per = (size_t)RENDER_MAX_INSTANCES * RENDER_MAX_JOINTS * 64;
        PCL_CK(c, cudaMemcpy(host, (const char*)r->d_mats + (size_t)draw * per, per, cudaMemcpyDeviceToHost));
        return PCL_OK;
    }

    // ------------------------------------------------------------------------------------------
    // palette read-back without any device allocation (no counterpart in the reference)
    // ------------------------------------------------------------------------------------------
    int pcl_batch_download_palette_static(pcl_batch* b, int64_t first, int64_t count, float* host16)
    {
        if (!b || !host16 || first < 0 || count < 1 || first + count > b->N)
            return fail(b ? b->ctx : nullptr, PCL_ERR_INVALID, "pcl_batch_download_palette_static: range outside the batch");
        pcl_ctx* c = b->ctx;
        if (!b->posed)
            return fail(c, PCL_ERR_INVALID, "pcl_batch_download_palette_static: pcl_batch_pose / pcl_batch_tick has not run");
        const SkelRec& s     = c->skels[b->skel];
        const size_t   bytes = (size_t)count * s.J * 64;
        if (bytes > (size_t)PCL_STATIC_EXPORT_MAX_BYTES)
            return fail(c, PCL_ERR_UNSUPPORTED,
                        "pcl_batch_download_palette_static: count * J * 64 bytes exceeds PCL_STATIC_EXPORT_MAX_BYTES (the fixed "
                        "staging array); download in smaller ranges or use pcl_batch_download_palette");
        if (c->device < 0 || c->device >= PCL_STATIC_MAX_DEVICES)
            return fail(c, PCL_ERR_UNSUPPORTED, "pcl_batch_download_palette_static: device index beyond the per-device address table");
        PCL_CK(c, cudaSetDevice(c->device));
        // One-time set-up per device, with that device current: in host code `g_static_export` is only a
        // stand-in for the symbol, not an address the GPU understands, so ask the runtime where this
        // device's copy of the array lives and hand THAT pointer to the kernel.
        // (Contexts may be driven from different threads: the slot is atomic; two first callers both look the address up and
        // store the same value.)
        void* addr = g_static_export_addr[c->device].load(std::memory_order_acquire);
        if (!addr)
        {
            PCL_CK(c, cudaGetSymbolAddress(&addr, g_static_export));
            g_static_export_addr[c->device].store(addr, std::memory_order_release);
        }
        float4* staging = reinterpret_cast<float4*>(addr);
        PCL_CK(c, cudaStreamWaitEvent(c->copy, c->ev_pose, 0));
        const long long total = count * s.J * 4; // every float4 of the range is overwritten: nothing to clear
        export_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->copy>>>(b->d_palette, staging, first, count, s.J, s.pitch);
        c->launches++;
        PCL_POST_LAUNCH(c, c->copy, "export_kernel (static staging)");
        PCL_CK(c, cudaEventRecord(c->ev_copy, c->copy));
        c->copy_pending = true;
        PCL_CK(c, cudaMemcpyAsync(host16, staging, bytes, cudaMemcpyDeviceToHost, c->copy));
        PCL_CK(c, cudaStreamSynchronize(c->copy)); // the shared array is free again when this returns
        return PCL_OK;
    }
}
