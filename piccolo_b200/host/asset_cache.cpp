// Binary cache of the parsed animation assets.
//
// The reference parses its JSON assets on every load (AssetManager::loadAsset, resource/asset_manager/asset_manager.h:17-40;
// the 437 KB clip of the bundled player costs milliseconds in json11).  With AnimationLoader::setCacheDir(dir) the loaders
// keep a flat little-endian image of each parsed asset in dir, keyed by the asset's full path and validated against the
// JSON file's size and modification time; a missing, stale, truncated or foreign file is ignored and rewritten after the
// JSON has been parsed, so the cache can never change what a load returns.  The image holds exactly the fields of the
// schema structs (piccolo_anim.h), i.e. what pcl_upload_* needs once flattened - no pointers, no padding games.
#include "piccolo_anim.h"

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <sys/stat.h>
#include <vector>

namespace Piccolo
{
    namespace
    {
        std::string g_cache_dir;

        constexpr uint32_t MAGIC = 0x424c4350u; // "PCLB"
        constexpr uint32_t VERSION = 1;

        struct Header
        {
            uint32_t magic, version, kind, reserved;
            int64_t  src_size, src_mtime_ns;
            uint64_t payload_bytes, payload_hash;
        };

        uint64_t fnv1a(const void* p, size_t n, uint64_t h = 1469598103934665603ull)
        {
            const unsigned char* b = static_cast<const unsigned char*>(p);
            for (size_t i = 0; i < n; i++)
                h = (h ^ b[i]) * 1099511628211ull;
            return h;
        }

        struct Writer
        {
            std::vector<unsigned char> buf;
            void raw(const void* p, size_t n)
            {
                const unsigned char* b = static_cast<const unsigned char*>(p);
                buf.insert(buf.end(), b, b + n);
            }
            template<typename T>
            void pod(const T& v)
            {
                raw(&v, sizeof(T));
            }
            void str(const std::string& s)
            {
                pod<uint32_t>((uint32_t)s.size());
                raw(s.data(), s.size());
            }
            template<typename T>
            void pods(const std::vector<T>& v) // T is a struct of 4-byte scalars only
            {
                pod<uint64_t>(v.size());
                if (!v.empty())
                    raw(v.data(), v.size() * sizeof(T));
            }
        };
        struct Reader
        {
            const unsigned char* p;
            const unsigned char* end;
            bool                 ok = true;
            bool raw(void* dst, size_t n)
            {
                if (!ok || (size_t)(end - p) < n)
                    return ok = false;
                std::memcpy(dst, p, n);
                p += n;
                return true;
            }
            template<typename T>
            bool pod(T& v)
            {
                return raw(&v, sizeof(T));
            }
            bool str(std::string& s)
            {
                uint32_t n = 0;
                if (!pod(n) || (size_t)(end - p) < n)
                    return ok = false;
                s.assign(reinterpret_cast<const char*>(p), n);
                p += n;
                return true;
            }
            template<typename T>
            bool pods(std::vector<T>& v)
            {
                uint64_t n = 0;
                if (!pod(n) || (uint64_t)(end - p) / sizeof(T) < n)
                    return ok = false;
                v.resize((size_t)n);
                return n == 0 || raw(v.data(), (size_t)n * sizeof(T));
            }
        };
        static_assert(sizeof(Vector3) == 12 && sizeof(Quaternion) == 16 && sizeof(Matrix4x4_) == 64, "schema PODs must be packed floats");
        static_assert(sizeof(Vertex) == 44 && sizeof(SkeletonBinding) == 32, "mesh PODs must be packed 4-byte scalars");

        // ---- per-type images ---------------------------------------------------------------------------------
        void put(Writer& w, const SkeletonData& d)
        {
            w.pod<uint64_t>(d.bones_map.size());
            for (const RawBone& b : d.bones_map)
            {
                w.str(b.name);
                w.pod<int32_t>(b.index), w.pod<int32_t>(b.parent_index);
                w.pod(b.binding_pose.m_position), w.pod(b.binding_pose.m_scale), w.pod(b.binding_pose.m_rotation);
                w.pod(b.tpose_matrix);
            }
            w.pod<int32_t>(d.is_flat), w.pod<int32_t>(d.root_index), w.pod<int32_t>(d.in_topological_order);
        }
        bool get(Reader& r, SkeletonData& d)
        {
            uint64_t n = 0;
            if (!r.pod(n) || n > (uint64_t)(r.end - r.p))
                return r.ok = false;
            d.bones_map.resize((size_t)n);
            for (RawBone& b : d.bones_map)
            {
                int32_t idx = 0, par = 0;
                r.str(b.name), r.pod(idx), r.pod(par);
                r.pod(b.binding_pose.m_position), r.pod(b.binding_pose.m_scale), r.pod(b.binding_pose.m_rotation);
                r.pod(b.tpose_matrix);
                b.index = idx, b.parent_index = par;
                if (!r.ok)
                    return false;
            }
            int32_t flat = 0, root = 0, topo = 0;
            r.pod(flat), r.pod(root), r.pod(topo);
            d.is_flat = flat != 0, d.root_index = root, d.in_topological_order = topo != 0;
            return r.ok;
        }
        void put(Writer& w, const AnimationClip& d)
        {
            w.pod<int32_t>(d.total_frame), w.pod<int32_t>(d.node_count);
            w.pod<uint64_t>(d.node_channels.size());
            for (const AnimationChannel& c : d.node_channels)
                w.str(c.name), w.pods(c.position_keys), w.pods(c.rotation_keys), w.pods(c.scaling_keys);
        }
        bool get(Reader& r, AnimationClip& d)
        {
            int32_t  tf = 0, nc = 0;
            uint64_t n = 0;
            if (!r.pod(tf) || !r.pod(nc) || !r.pod(n) || n > (uint64_t)(r.end - r.p))
                return r.ok = false;
            d.total_frame = tf, d.node_count = nc;
            d.node_channels.resize((size_t)n);
            for (AnimationChannel& c : d.node_channels)
                if (!r.str(c.name) || !r.pods(c.position_keys) || !r.pods(c.rotation_keys) || !r.pods(c.scaling_keys))
                    return false;
            return r.ok;
        }
        void put(Writer& w, const AnimSkelMap& d) { w.pods(d.convert); }
        bool get(Reader& r, AnimSkelMap& d) { return r.pods(d.convert); }
        void put(Writer& w, const BoneBlendMask& d) { w.str(d.skeleton_file_path), w.pods(d.enabled); }
        bool get(Reader& r, BoneBlendMask& d) { return r.str(d.skeleton_file_path) && r.pods(d.enabled); }
        void put(Writer& w, const MeshData& d) { w.pods(d.vertex_buffer), w.pods(d.index_buffer), w.pods(d.bind); }
        bool get(Reader& r, MeshData& d) { return r.pods(d.vertex_buffer) && r.pods(d.index_buffer) && r.pods(d.bind); }

        bool stat_source(const std::string& path, int64_t& size, int64_t& mtime_ns)
        {
            struct stat st;
            if (::stat(path.c_str(), &st) != 0)
                return false;
            size     = (int64_t)st.st_size;
            mtime_ns = (int64_t)st.st_mtim.tv_sec * 1000000000ll + (int64_t)st.st_mtim.tv_nsec;
            return true;
        }
        std::string cache_path(uint32_t kind, const std::string& full_path)
        {
            char name[64];
            std::snprintf(name, sizeof(name), "/%016llx.k%u.pclbin", (unsigned long long)fnv1a(full_path.data(), full_path.size()), kind);
            return g_cache_dir + name;
        }

        template<typename T>
        bool try_load(uint32_t kind, const std::string& full_path, T& out)
        {
            int64_t size = 0, mtime = 0;
            if (g_cache_dir.empty() || !stat_source(full_path, size, mtime))
                return false;
            FILE* f = std::fopen(cache_path(kind, full_path).c_str(), "rb");
            if (!f)
                return false;
            Header h {};
            bool   ok = std::fread(&h, sizeof(h), 1, f) == 1 && h.magic == MAGIC && h.version == VERSION && h.kind == kind && h.src_size == size &&
                      h.src_mtime_ns == mtime && h.payload_bytes < (1ull << 32);
            std::vector<unsigned char> buf;
            if (ok)
            {
                buf.resize((size_t)h.payload_bytes);
                ok = (buf.empty() || std::fread(buf.data(), 1, buf.size(), f) == buf.size()) && fnv1a(buf.data(), buf.size()) == h.payload_hash;
            }
            std::fclose(f);
            if (!ok)
                return false;
            Reader r {buf.data(), buf.data() + buf.size()};
            T      tmp;
            if (!get(r, tmp) || r.p != r.end)
                return false;
            out = std::move(tmp);
            return true;
        }
        template<typename T>
        void store(uint32_t kind, const std::string& full_path, const T& v)
        {
            int64_t size = 0, mtime = 0;
            if (g_cache_dir.empty() || !stat_source(full_path, size, mtime))
                return; // nothing is cached for a source that does not exist: the reference's "default object" stays uncached
            Writer w;
            put(w, v);
            Header            h {MAGIC, VERSION, kind, 0, size, mtime, (uint64_t)w.buf.size(), fnv1a(w.buf.data(), w.buf.size())};
            const std::string dst = cache_path(kind, full_path), tmp = dst + ".tmp";
            FILE*             f   = std::fopen(tmp.c_str(), "wb");
            if (!f)
                return;
            const bool ok = std::fwrite(&h, sizeof(h), 1, f) == 1 && (w.buf.empty() || std::fwrite(w.buf.data(), 1, w.buf.size(), f) == w.buf.size());
            std::fclose(f);
            if (ok)
                std::rename(tmp.c_str(), dst.c_str()); // readers never see a half-written image
            else
                std::remove(tmp.c_str());
        }
    } // namespace

    void               AnimationLoader::setCacheDir(const std::string& dir) { g_cache_dir = dir; }
    const std::string& AnimationLoader::cacheDir() { return g_cache_dir; }

    namespace detail
    {
        bool cacheLoad(const std::string& p, SkeletonData& o) { return try_load(1, p, o); }
        bool cacheLoad(const std::string& p, AnimationClip& o) { return try_load(2, p, o); }
        bool cacheLoad(const std::string& p, AnimSkelMap& o) { return try_load(3, p, o); }
        bool cacheLoad(const std::string& p, BoneBlendMask& o) { return try_load(4, p, o); }
        bool cacheLoad(const std::string& p, MeshData& o) { return try_load(5, p, o); }
        void cacheStore(const std::string& p, const SkeletonData& o) { store(1, p, o); }
        void cacheStore(const std::string& p, const AnimationClip& o) { store(2, p, o); }
        void cacheStore(const std::string& p, const AnimSkelMap& o) { store(3, p, o); }
        void cacheStore(const std::string& p, const BoneBlendMask& o) { store(4, p, o); }
        void cacheStore(const std::string& p, const MeshData& o) { store(5, p, o); }
    } // namespace detail
} // namespace Piccolo
