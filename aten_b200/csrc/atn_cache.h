// atn_cache.h -- host-side files of the traversal data.
//
//  * the reference's `.sbvh` tree files: sbvh::exportTree / sbvh::importTree
//    (src/libaten/accelerator/sbvh.cpp:1220-1436): a 44-byte header, the voxel-LOD material name table, then nodeNum
//    48-byte ThreadedSbvhNodes.  read_sbvh() applies what importTree applies (triangle offset, material ids re-mapped
//    through the name table); the node array is usable as one bottom-layer list of an AtnSceneBlob as it is.
//  * the ENCODED bottom trees (QuadNode / TriLeaf arrays + BottomInfo of atn_bvh.h), so that a large scene skips the
//    host-side re-encoding of its node lists on the next upload (2 s for 10 M triangles): keyed by a hash of exactly what
//    the encoding depends on -- the bottom node lists, the triangle parameters, the vertex positions and the encoder's
//    version.  The top layer (instances, matrices) is never cached: it is re-encoded in microseconds (encode_top) and is
//    what changes from frame to frame.
#pragma once

#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <atomic>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include "atn_bvh.h"

namespace atn {

// ---- .sbvh -------------------------------------------------------------------------------------------------------
#pragma pack(push, 1)
struct SbvhFileHeader {             // sbvh.cpp:1220-1235
    char magic[4];                  // "SBVH"
    uint8_t version[4];
    uint32_t nodeNum, maxDepth, cntMtrlForVoxel;
    float boxmin[3], boxmax[3];
};
#pragma pack(pop)
static_assert(sizeof(SbvhFileHeader) == 44, "SbvhFileHeader");

struct SbvhFile {
    SbvhFileHeader header{};
    std::map<int32_t, std::string> voxel_materials;     // id in the file -> material name
    std::vector<AtnBvhNode> nodes;
    std::string error;
};

// sbvh::importTree (sbvh.cpp:1345-1436).  `material_names` (optional): the scene's materials by index, for the
// re-mapping of voxel material ids; a name that is not found is an error, as the reference asserts.
inline SbvhFile read_sbvh(const char* path, int32_t offset_tri_idx, const char* const* material_names, uint32_t material_count)
{
    SbvhFile f;
    FILE* fp = fopen(path, "rb");
    if (!fp) { f.error = std::string("cannot open ") + path; return f; }
    auto fail = [&](const std::string& what) { f.error = std::string(path) + ": " + what; fclose(fp); f.nodes.clear(); return f; };
    if (fread(&f.header, sizeof(f.header), 1, fp) != 1) return fail("too short for an SBVH header");
    if (memcmp(f.header.magic, "SBVH", 4) != 0) return fail("bad magic");
    for (uint32_t i = 0; i < f.header.cntMtrlForVoxel; i++) {
        int32_t id = -1, len = 0;
        char buf[128];
        if (fread(&id, 4, 1, fp) != 1 || fread(&len, 4, 1, fp) != 1) return fail("truncated material table");
        if (!(0 < len && len < (int32_t)sizeof(buf))) return fail("bad material name length");
        if (fread(buf, 1, (size_t)len, fp) != (size_t)len) return fail("truncated material table");
        buf[len] = 0;
        f.voxel_materials[id] = buf;
    }
    f.nodes.resize(f.header.nodeNum);
    if (f.header.nodeNum && fread(f.nodes.data(), sizeof(AtnBvhNode), f.header.nodeNum, fp) != f.header.nodeNum) return fail("truncated node array");
    fclose(fp);
    if (offset_tri_idx > 0 || !f.voxel_materials.empty()) {
        for (AtnBvhNode& n : f.nodes) {
            if (n.primid >= 0) n.primid += (float)offset_tri_idx;            // ThreadedSbvhNode::triid
            if (n.meshid >= 0 && !f.voxel_materials.empty()) {               // ThreadedSbvhNode::mtrlid sits where meshid is
                auto it = f.voxel_materials.find((int32_t)n.meshid);
                if (it == f.voxel_materials.end()) { f.error = std::string(path) + ": a node refers to a voxel material that is not in the name table"; f.nodes.clear(); return f; }
                if (material_names) {
                    int32_t found = -1;
                    for (uint32_t k = 0; k < material_count; k++)
                        if (material_names[k] && it->second == material_names[k]) { found = (int32_t)k; break; }
                    if (found < 0) { f.error = std::string(path) + ": material '" + it->second + "' not found in the scene"; f.nodes.clear(); return f; }
                    n.meshid = (float)found;
                }
            }
        }
    }
    return f;
}

// sbvh::exportTree's layout (sbvh.cpp:1236-1340)
inline std::string write_sbvh(const char* path, const AtnBvhNode* nodes, uint32_t count, uint32_t max_depth,
                              const float boxmin[3], const float boxmax[3],
                              const int32_t* voxel_ids, const char* const* voxel_names, uint32_t voxel_count)
{
    FILE* fp = fopen(path, "wb");
    if (!fp) return std::string("cannot create ") + path;
    SbvhFileHeader h{};
    memcpy(h.magic, "SBVH", 4);
    h.version[0] = 0; h.version[1] = 0; h.version[2] = 0; h.version[3] = 1;
    h.nodeNum = count; h.maxDepth = max_depth; h.cntMtrlForVoxel = voxel_count;
    for (int a = 0; a < 3; a++) { h.boxmin[a] = boxmin[a]; h.boxmax[a] = boxmax[a]; }
    bool ok = fwrite(&h, sizeof(h), 1, fp) == 1;
    for (uint32_t i = 0; ok && i < voxel_count; i++) {
        const int32_t raw = (int32_t)strlen(voxel_names[i]), alen = (raw + 3) / 4 * 4;
        char buf[132] = { 0 };
        if (alen >= 128) { ok = false; break; }
        memcpy(buf, voxel_names[i], (size_t)raw);
        ok = fwrite(&voxel_ids[i], 4, 1, fp) == 1 && fwrite(&alen, 4, 1, fp) == 1 && fwrite(buf, 1, (size_t)alen, fp) == (size_t)alen;
    }
    if (ok && count) ok = fwrite(nodes, sizeof(AtnBvhNode), count, fp) == count;
    ok = (fclose(fp) == 0) && ok;
    return ok ? std::string() : std::string("write error on ") + path;
}

// ---- encoded bottom trees ----------------------------------------------------------------------------------------
constexpr uint32_t kEncoderVersion = 3;        // bump when the QuadNode / TriLeaf encoding changes

inline uint64_t hash_words(const void* data, size_t bytes, uint64_t h)
{
    const uint8_t* p = static_cast<const uint8_t*>(data);
    size_t n = bytes / 8;
    for (size_t i = 0; i < n; i++) {
        uint64_t w; memcpy(&w, p + 8 * i, 8);
        h = (h ^ w) * 0x9E3779B97F4A7C15ull;
        h ^= h >> 29;
    }
    uint64_t tail = 0;
    memcpy(&tail, p + 8 * n, bytes - 8 * n);
    h = (h ^ tail ^ (uint64_t)bytes) * 0xC2B2AE3D27D4EB4Full;
    return h ^ (h >> 32);
}

// Hash of what the bottom-tree encoding depends on: every bottom node list (in order, with its size), the triangle
// parameters, the vertex positions.  Lists are hashed in parallel and combined in list order.
inline uint64_t source_hash(const AtnSceneBlob* scene)
{
    const uint32_t nl = scene->node_list_count;
    std::vector<uint64_t> part(nl + 2, 0);
    std::atomic<uint32_t> next{ 1 };
    auto worker = [&]() {
        for (uint32_t k = next.fetch_add(1); k < nl + 2; k = next.fetch_add(1)) {
            if (k < nl) part[k] = hash_words(scene->node_lists[k], (size_t)scene->node_list_sizes[k] * sizeof(AtnBvhNode), 0x1234567ull + k);
            else if (k == nl) part[k] = hash_words(scene->triangles, (size_t)scene->triangle_count * sizeof(AtnTriangleParameter), 0xabcdefull);
            else part[k] = hash_words(scene->positions, (size_t)scene->vertex_count * sizeof(AtnVec4), 0x13579bdfull);
        }
    };
    const unsigned nthreads = std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u);
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < nthreads; t++) pool.emplace_back(worker);
    worker();
    for (auto& th : pool) th.join();
    uint64_t h = 0x9ae16a3b2f90404full ^ ((uint64_t)kEncoderVersion << 40) ^ nl;
    for (uint32_t k = 1; k < nl + 2; k++) h = hash_words(&part[k], 8, h);
    return h;
}

#pragma pack(push, 1)
struct EncodedFileHeader {
    char magic[8];                  // "ATNENC01"
    uint32_t encoder_version, node_list_count;
    uint64_t source_hash, quad_nodes, tri_leaves;
    uint64_t slots_used, replaced, not_nested;
};
#pragma pack(pop)

// Returns "" and fills `out` when `path` holds the encoding of exactly this source (hash, list count, version);
// otherwise the reason (no file, other scene, other version, truncated) -- the caller then encodes and rewrites it.
inline std::string read_encoded(const char* path, uint64_t hash, uint32_t node_list_count, CachedBottom* out)
{
    FILE* fp = fopen(path, "rb");
    if (!fp) return "no cache file";
    EncodedFileHeader h{};
    std::string err;
    auto rd = [&](void* dst, size_t bytes) { return bytes == 0 || fread(dst, 1, bytes, fp) == bytes; };
    if (!rd(&h, sizeof(h)) || memcmp(h.magic, "ATNENC01", 8) != 0) err = "not an encoded-tree file";
    else if (h.encoder_version != kEncoderVersion) err = "written by another encoder version";
    else if (h.source_hash != hash || h.node_list_count != node_list_count) err = "encodes another scene";
    else {
        out->info.roots.resize(node_list_count); out->info.need.resize(node_list_count); out->info.box.resize((size_t)node_list_count * 6);
        out->nodes.resize(h.quad_nodes); out->stack_need.resize(h.quad_nodes); out->tris.resize(h.tri_leaves);
        out->info.node_count = (int32_t)h.quad_nodes;
        out->slots_used = h.slots_used; out->replaced = h.replaced; out->not_nested = h.not_nested;
        if (!rd(out->info.roots.data(), 4 * (size_t)node_list_count) || !rd(out->info.need.data(), 4 * (size_t)node_list_count) ||
            !rd(out->info.box.data(), 24 * (size_t)node_list_count) || !rd(out->nodes.data(), sizeof(QuadNode) * h.quad_nodes) ||
            !rd(out->stack_need.data(), 4 * h.quad_nodes) || !rd(out->tris.data(), sizeof(TriLeaf) * h.tri_leaves))
            err = "truncated";
        uint64_t check = 0;
        if (err.empty() && (!rd(&check, 8) || check != (hash ^ h.quad_nodes ^ (h.tri_leaves << 1)))) err = "truncated";
    }
    fclose(fp);
    return err;
}

// Writes to `path`.tmp and renames, so that a reader never sees half a file.
inline std::string write_encoded(const char* path, uint64_t hash, uint32_t node_list_count, const BottomInfo& info,
                                 const QuadNode* nodes, const int32_t* stack_need, uint64_t n_nodes, const TriLeaf* tris, uint64_t n_tris,
                                 uint64_t slots_used, uint64_t replaced, uint64_t not_nested)
{
    const std::string tmp = std::string(path) + ".tmp";
    FILE* fp = fopen(tmp.c_str(), "wb");
    if (!fp) return "cannot create " + tmp;
    EncodedFileHeader h{};
    memcpy(h.magic, "ATNENC01", 8);
    h.encoder_version = kEncoderVersion; h.node_list_count = node_list_count; h.source_hash = hash;
    h.quad_nodes = n_nodes; h.tri_leaves = n_tris; h.slots_used = slots_used; h.replaced = replaced; h.not_nested = not_nested;
    auto wr = [&](const void* src, size_t bytes) { return bytes == 0 || fwrite(src, 1, bytes, fp) == bytes; };
    const uint64_t check = hash ^ n_nodes ^ (n_tris << 1);
    bool ok = wr(&h, sizeof(h)) && wr(info.roots.data(), 4 * (size_t)node_list_count) && wr(info.need.data(), 4 * (size_t)node_list_count) &&
              wr(info.box.data(), 24 * (size_t)node_list_count) && wr(nodes, sizeof(QuadNode) * n_nodes) && wr(stack_need, 4 * n_nodes) &&
              wr(tris, sizeof(TriLeaf) * n_tris) && wr(&check, 8);
    ok = (fclose(fp) == 0) && ok;
    if (!ok) { remove(tmp.c_str()); return "write error on " + tmp; }
    if (rename(tmp.c_str(), path) != 0) { remove(tmp.c_str()); return std::string("cannot rename to ") + path; }
    return std::string();
}

}  // namespace atn
