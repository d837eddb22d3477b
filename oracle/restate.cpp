// oracle/restate.cpp -- TEST INFRASTRUCTURE ONLY (see restate.h).
//
// Serial CPU restatement of the reference hot path.  Paths cited are relative to
// /root/reference/src/InfiniTAM-Global-Consistency-h/InfiniTAM/ITMLib/ ;
//   DA/  = Engine/DeviceAgnostic/        CPU/ = Engine/DeviceSpecific/CPU/
//   CUDA/ = Engine/DeviceSpecific/CUDA/
// Build with -ffp-contract=off: every float expression below is written in the reference's
// operation order and must not be fused.
#include "restate.h"

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <vector>

namespace {

struct V3 { float x, y, z; };
struct V4 { float x, y, z, w; };

// Matrix4 * Vector4, ORUtils/Matrix.h:128-135 (column-major, left-to-right sums)
inline V4 mulM(const float *m, V4 v)
{
    V4 r;
    r.x = m[0] * v.x + m[4] * v.y + m[8] * v.z + m[12] * v.w;
    r.y = m[1] * v.x + m[5] * v.y + m[9] * v.z + m[13] * v.w;
    r.z = m[2] * v.x + m[6] * v.y + m[10] * v.z + m[14] * v.w;
    r.w = m[3] * v.x + m[7] * v.y + m[11] * v.z + m[15] * v.w;
    return r;
}

// hashIndex, DA/ITMRepresentationAccess.h:10-12 (mask generalised to bucket_num-1)
inline int hashIndex(int x, int y, int z, int bucketNum)
{
    return (int)((((unsigned)x * 73856093u) ^ ((unsigned)y * 19349669u) ^ ((unsigned)z * 83492791u)) &
                 (unsigned)(bucketNum - 1));
}

inline bool posEq(const dsb200_hash_entry &e, int x, int y, int z)
{
    return e.pos[0] == x && e.pos[1] == y && e.pos[2] == z;
}

inline void voxelReset(dsb200_voxel &v)
{ // ITMVoxel_s_rgb::reset, Utils/ITMLibDefines.h:150-156
    v.sdf = 32767; v.w_depth = 0; v.clr[0] = v.clr[1] = v.clr[2] = 0; v.w_color = 0;
}

struct Snapshot { std::vector<int> ids; size_t frameIdx; };
typedef std::shared_ptr<Snapshot> SnapPtr;

} // namespace

struct rst_recon {
    int numBlocks, bucketNum, excessSize, total;
    std::vector<unsigned char> allocType;   // entriesAllocType, CPU/ITMSceneReconstructionEngine_CPU.h:19
    std::vector<short> blockCoords;         // blockCoords (Vector4s), :20
    dsb200_weight_params weights;           // Engine/ITMSceneReconstructionEngine.h:21-25
    // history queues, CUDA/ITMSceneReconstructionEngine_CUDA.h:41-48
    std::deque<SnapPtr> decayQ, slideQ, decayDefusionQ, slideDefusionQ;
    std::map<double, SnapPtr> snapshots;    // mDefusionBlockDataBase
    size_t frameIdx;
    long long totalDecayed;
    dsb200_shard shard;                     // sharded scene (include/dsb200.h): count <= 1 = unsharded
};

extern "C" {

rst_recon *rst_recon_create(int num_blocks, int bucket_num, int excess_size)
{
    rst_recon *e = new rst_recon();
    e->numBlocks = num_blocks; e->bucketNum = bucket_num; e->excessSize = excess_size;
    e->total = bucket_num + excess_size;
    e->allocType.assign(e->total, 0);
    e->blockCoords.assign((size_t)e->total * 4, 0);
    e->weights.depth_weighting = 0; e->weights.max_new_w = 10; e->weights.max_distance = 100;
    e->frameIdx = 0; e->totalDecayed = 0;
    e->shard.rank = 0; e->shard.count = 1; e->shard.brick_log2 = 2; e->shard.halo = 0;
    return e;
}
void rst_recon_destroy(rst_recon *e) { delete e; }
void rst_set_fusion_weight_params(rst_recon *e, const dsb200_weight_params *p) { e->weights = *p; }
size_t rst_decayed_block_count(const rst_recon *e) { return (size_t)e->totalDecayed; }
void rst_set_shard(rst_recon *e, const dsb200_shard *s) { e->shard = *s; if (e->shard.count < 1) e->shard.count = 1; }
// block ownership of the sharded scene (no reference counterpart; definition in include/dsb200.h)
int rst_shard_owner(int bx, int by, int bz, int count, int brickLog2)
{
    if (count <= 1) return 0;
    unsigned h = ((unsigned)(bx >> brickLog2) * 73856093u) ^ ((unsigned)(by >> brickLog2) * 19349669u) ^
                 ((unsigned)(bz >> brickLog2) * 83492791u);
    h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
    return (int)(h % (unsigned)count);
}
// held = owned, or (halo) a +1 neighbour of an owned block: some b - d, d in {0,1}^3, is owned
int rst_shard_holds(int bx, int by, int bz, const dsb200_shard *s)
{
    if (s->count <= 1) return 1;
    for (int d = 0; d < (s->halo ? 8 : 1); d++)
        if (rst_shard_owner(bx - (d & 1), by - ((d >> 1) & 1), bz - (d >> 2), s->count, s->brick_log2) == s->rank) return 1;
    return 0;
}
void rst_history_sizes(const rst_recon *e, int *a, int *b, int *c, int *d, int *s)
{
    if (a) *a = (int)e->decayQ.size();
    if (b) *b = (int)e->slideQ.size();
    if (c) *c = (int)e->decayDefusionQ.size();
    if (d) *d = (int)e->slideDefusionQ.size();
    if (s) *s = (int)e->snapshots.size();
}

// ResetScene: CPU/ITMSceneReconstructionEngine_CPU.cpp:25-45, CUDA .cu:111-148 (queues cleared)
void rst_reset_scene(rst_recon *e, dsb200_scene *scene)
{
    size_t nv = (size_t)scene->num_blocks * DSB200_BLOCK_SIZE3;
    for (size_t i = 0; i < nv; i++) { voxelReset(scene->voxels[i]); scene->voxels[i].pad_ = 0; }
    for (int i = 0; i < scene->num_blocks; i++) scene->alloc_list[i] = i;
    scene->last_free_block = scene->num_blocks - 1;
    dsb200_hash_entry blank; memset(&blank, 0, sizeof(blank)); blank.ptr = -2;
    int total = scene->bucket_num + scene->excess_size;
    for (int i = 0; i < total; i++) scene->hash[i] = blank;
    for (int i = 0; i < scene->excess_size; i++) scene->excess_list[i] = i;
    scene->last_free_excess = scene->excess_size - 1;
    if (e) {
        e->totalDecayed = 0;
        e->decayQ.clear(); e->slideQ.clear(); e->snapshots.clear();
    }
}

// Matrix4::inv, ORUtils/Matrix.h:175-246 (cofactor expansion, same grouping)
int rst_mat4_inverse(const float *M, float *dst)
{
    float tmp[12], src[16], det;
    for (int i = 0; i < 4; i++) {
        src[i] = M[i * 4]; src[i + 4] = M[i * 4 + 1]; src[i + 8] = M[i * 4 + 2]; src[i + 12] = M[i * 4 + 3];
    }
    tmp[0] = src[10] * src[15]; tmp[1] = src[11] * src[14]; tmp[2] = src[9] * src[15]; tmp[3] = src[11] * src[13];
    tmp[4] = src[9] * src[14]; tmp[5] = src[10] * src[13]; tmp[6] = src[8] * src[15]; tmp[7] = src[11] * src[12];
    tmp[8] = src[8] * src[14]; tmp[9] = src[10] * src[12]; tmp[10] = src[8] * src[13]; tmp[11] = src[9] * src[12];
    dst[0] = (tmp[0] * src[5] + tmp[3] * src[6] + tmp[4] * src[7]) - (tmp[1] * src[5] + tmp[2] * src[6] + tmp[5] * src[7]);
    dst[1] = (tmp[1] * src[4] + tmp[6] * src[6] + tmp[9] * src[7]) - (tmp[0] * src[4] + tmp[7] * src[6] + tmp[8] * src[7]);
    dst[2] = (tmp[2] * src[4] + tmp[7] * src[5] + tmp[10] * src[7]) - (tmp[3] * src[4] + tmp[6] * src[5] + tmp[11] * src[7]);
    dst[3] = (tmp[5] * src[4] + tmp[8] * src[5] + tmp[11] * src[6]) - (tmp[4] * src[4] + tmp[9] * src[5] + tmp[10] * src[6]);
    det = src[0] * dst[0] + src[1] * dst[1] + src[2] * dst[2] + src[3] * dst[3];
    if (det == 0.0f) return 0;
    dst[4] = (tmp[1] * src[1] + tmp[2] * src[2] + tmp[5] * src[3]) - (tmp[0] * src[1] + tmp[3] * src[2] + tmp[4] * src[3]);
    dst[5] = (tmp[0] * src[0] + tmp[7] * src[2] + tmp[8] * src[3]) - (tmp[1] * src[0] + tmp[6] * src[2] + tmp[9] * src[3]);
    dst[6] = (tmp[3] * src[0] + tmp[6] * src[1] + tmp[11] * src[3]) - (tmp[2] * src[0] + tmp[7] * src[1] + tmp[10] * src[3]);
    dst[7] = (tmp[4] * src[0] + tmp[9] * src[1] + tmp[10] * src[2]) - (tmp[5] * src[0] + tmp[8] * src[1] + tmp[11] * src[2]);
    tmp[0] = src[2] * src[7]; tmp[1] = src[3] * src[6]; tmp[2] = src[1] * src[7]; tmp[3] = src[3] * src[5];
    tmp[4] = src[1] * src[6]; tmp[5] = src[2] * src[5]; tmp[6] = src[0] * src[7]; tmp[7] = src[3] * src[4];
    tmp[8] = src[0] * src[6]; tmp[9] = src[2] * src[4]; tmp[10] = src[0] * src[5]; tmp[11] = src[1] * src[4];
    dst[8] = (tmp[0] * src[13] + tmp[3] * src[14] + tmp[4] * src[15]) - (tmp[1] * src[13] + tmp[2] * src[14] + tmp[5] * src[15]);
    dst[9] = (tmp[1] * src[12] + tmp[6] * src[14] + tmp[9] * src[15]) - (tmp[0] * src[12] + tmp[7] * src[14] + tmp[8] * src[15]);
    dst[10] = (tmp[2] * src[12] + tmp[7] * src[13] + tmp[10] * src[15]) - (tmp[3] * src[12] + tmp[6] * src[13] + tmp[11] * src[15]);
    dst[11] = (tmp[5] * src[12] + tmp[8] * src[13] + tmp[11] * src[14]) - (tmp[4] * src[12] + tmp[9] * src[13] + tmp[10] * src[14]);
    dst[12] = (tmp[2] * src[10] + tmp[5] * src[11] + tmp[1] * src[9]) - (tmp[4] * src[11] + tmp[0] * src[9] + tmp[3] * src[10]);
    dst[13] = (tmp[8] * src[11] + tmp[0] * src[8] + tmp[7] * src[10]) - (tmp[6] * src[10] + tmp[9] * src[11] + tmp[1] * src[8]);
    dst[14] = (tmp[6] * src[9] + tmp[11] * src[11] + tmp[3] * src[8]) - (tmp[10] * src[11] + tmp[2] * src[8] + tmp[7] * src[9]);
    dst[15] = (tmp[10] * src[10] + tmp[4] * src[8] + tmp[9] * src[9]) - (tmp[8] * src[9] + tmp[11] * src[10] + tmp[5] * src[8]);
    float s = 1 / det;
    for (int i = 0; i < 16; i++) dst[i] *= s;
    return 1;
}

// Matrix4 * Matrix4, ORUtils/Matrix.h:114-120 : r(x,y) += lhs(k,y)*rhs(x,k), (col,row) -> m[col*4+row]
void rst_mat4_multiply(const float *A, const float *B, float *out)
{
    float r[16];
    for (int i = 0; i < 16; i++) r[i] = 0.0f;
    for (int x = 0; x < 4; x++) for (int y = 0; y < 4; y++) for (int k = 0; k < 4; k++)
        r[x * 4 + y] += A[k * 4 + y] * B[x * 4 + k];
    memcpy(out, r, sizeof(r));
}

} // extern "C"

// ---------------------------------------------------------------------------------------------
// Allocation + visible list
// ---------------------------------------------------------------------------------------------
namespace {

// checkPointVisibility<useSwapping>, DA/ITMSceneReconstructionEngine.h:491-517
inline void pointVisibility(bool &vis, bool &visEnl, V4 p, const float *M, const float *proj, int W, int H,
                            bool useSwapping)
{
    V4 q = mulM(M, p);
    if (q.z < 1e-10f) return;
    float u = proj[0] * q.x / q.z + proj[2];
    float v = proj[1] * q.y / q.z + proj[3];
    if (u >= 0 && u < W && v >= 0 && v < H) { vis = true; visEnl = true; }
    else if (useSwapping) {
        int lx = -W / 8, ly = W + W / 8, lz = -H / 8, lw = H + H / 8;
        if (u >= lx && u < ly && v >= lz && v < lw) visEnl = true;
    }
}

// checkBlockVisibility<useSwapping>, DA/ITMSceneReconstructionEngine.h:519-569.
// The corner walk is an in-place +=/-= sequence on the running point; keep it that way.
inline void blockVisibility(bool &vis, bool &visEnl, const int16_t *pos, const float *M, const float *proj,
                            float voxelSize, int W, int H, bool useSwapping)
{
    float f = (float)DSB200_BLOCK_SIZE * voxelSize;
    V4 p;
    vis = false; visEnl = false;
    p.x = (float)pos[0] * f; p.y = (float)pos[1] * f; p.z = (float)pos[2] * f; p.w = 1.0f;
    pointVisibility(vis, visEnl, p, M, proj, W, H, useSwapping); if (vis) return;   // 000
    p.z += f; pointVisibility(vis, visEnl, p, M, proj, W, H, useSwapping); if (vis) return;   // 001
    p.y += f; pointVisibility(vis, visEnl, p, M, proj, W, H, useSwapping); if (vis) return;   // 011
    p.x += f; pointVisibility(vis, visEnl, p, M, proj, W, H, useSwapping); if (vis) return;   // 111
    p.z -= f; pointVisibility(vis, visEnl, p, M, proj, W, H, useSwapping); if (vis) return;   // 110
    p.y -= f; pointVisibility(vis, visEnl, p, M, proj, W, H, useSwapping); if (vis) return;   // 100
    p.x -= f; p.y += f; pointVisibility(vis, visEnl, p, M, proj, W, H, useSwapping); if (vis) return; // 010
    p.x += f; p.y -= f; p.z += f; pointVisibility(vis, visEnl, p, M, proj, W, H, useSwapping);        // 101
}

// buildHashAllocAndVisibleTypePP, DA/ITMSceneReconstructionEngine.h:287-399
inline void allocRequestForPixel(unsigned char *allocType, unsigned char *visType, short *blockCoords, int x, int y,
                                 const float *depth, const float *invM, const float *invProj, float mu, int W,
                                 float oneOverBlock, const dsb200_hash_entry *hash, int bucketNum, float vfMin,
                                 float vfMax, const dsb200_shard &shard)
{
    float d = depth[x + y * W];
    if (d <= 0 || (d - mu) < 0 || (d - mu) < vfMin || (d + mu) > vfMax) return;
    V3 pc;
    pc.z = d;
    pc.x = pc.z * (((float)x - invProj[2]) * invProj[0]);
    pc.y = pc.z * (((float)y - invProj[3]) * invProj[1]);
    float norm = sqrtf(pc.x * pc.x + pc.y * pc.y + pc.z * pc.z);
    V4 t;
    t.x = pc.x * (1.0f - mu / norm); t.y = pc.y * (1.0f - mu / norm); t.z = pc.z * (1.0f - mu / norm); t.w = 1.0f;
    V4 a = mulM(invM, t);
    V3 point = {a.x * oneOverBlock, a.y * oneOverBlock, a.z * oneOverBlock};
    t.x = pc.x * (1.0f + mu / norm); t.y = pc.y * (1.0f + mu / norm); t.z = pc.z * (1.0f + mu / norm);
    V4 b = mulM(invM, t);
    V3 pointE = {b.x * oneOverBlock, b.y * oneOverBlock, b.z * oneOverBlock};
    V3 dir = {pointE.x - point.x, pointE.y - point.y, pointE.z - point.z};
    norm = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
    int noSteps = (int)ceilf(2.0f * norm);
    float div = (float)(noSteps - 1);
    dir.x /= div; dir.y /= div; dir.z /= div;

    for (int i = 0; i < noSteps; i++) {
        short bx = (short)floorf(point.x), by = (short)floorf(point.y), bz = (short)floorf(point.z);
        if (!rst_shard_holds(bx, by, bz, &shard)) {
            point.x += dir.x; point.y += dir.y; point.z += dir.z;      // another shard's block
            continue;
        }
        int h = hashIndex(bx, by, bz, bucketNum);
        bool found = false;
        dsb200_hash_entry e = hash[h];
        if (posEq(e, bx, by, bz) && e.ptr >= -1) { visType[h] = (e.ptr == -1) ? 2 : 1; found = true; }
        if (!found) {
            bool isExcess = false;
            if (e.ptr >= -1) {
                while (e.offset >= 1) {
                    h = bucketNum + e.offset - 1;
                    e = hash[h];
                    if (posEq(e, bx, by, bz) && e.ptr >= -1) { visType[h] = (e.ptr == -1) ? 2 : 1; found = true; break; }
                }
                isExcess = true;
            }
            if (!found) {
                allocType[h] = isExcess ? 2 : 1;
                if (!isExcess) visType[h] = 1;
                blockCoords[h * 4 + 0] = bx; blockCoords[h * 4 + 1] = by; blockCoords[h * 4 + 2] = bz; blockCoords[h * 4 + 3] = 1;
            }
        }
        point.x += dir.x; point.y += dir.y; point.z += dir.z;
    }
}

} // namespace

extern "C" {

// AllocateSceneFromDepth: CPU/ITMSceneReconstructionEngine_CPU.cpp:137-342 plus the history
// bookkeeping that only the CUDA engine has (CUDA/ITMSceneReconstructionEngine_CUDA.cu:257-358).
int rst_allocate_from_depth(rst_recon *e, dsb200_scene *scene, const dsb200_view *view,
                            const dsb200_tracking *tracking, dsb200_render_state *rs, int onlyUpdateVisibleList,
                            int isDefusion)
{
    const int W = view->depth_w, H = view->depth_h;
    const float voxelSize = scene->params.voxel_size, mu = scene->params.mu;
    const float *M = tracking->M_d, *invM = tracking->invM_d;
    float invProj[4] = {1.0f / view->intr_d[0], 1.0f / view->intr_d[1], view->intr_d[2], view->intr_d[3]};
    const int total = scene->bucket_num + scene->excess_size, bucketNum = scene->bucket_num;
    float oneOverBlock = 1.0f / (voxelSize * DSB200_BLOCK_SIZE);
    int lastFreeBlock = scene->last_free_block, lastFreeExcess = scene->last_free_excess;
    unsigned char *allocType = e->allocType.data(), *visType = rs->visible_type;
    short *blockCoords = e->blockCoords.data();
    dsb200_hash_entry *hash = scene->hash;
    bool useSwapping = scene->use_swapping != 0;

    memset(allocType, 0, total);
    for (int i = 0; i < rs->no_visible; i++) visType[rs->visible_ids[i]] = 3;

    for (int loc = 0; loc < W * H; loc++) {
        int y = loc / W, x = loc - y * W;
        allocRequestForPixel(allocType, visType, blockCoords, x, y, view->depth, invM, invProj, mu, W, oneOverBlock,
                             hash, bucketNum, scene->params.vf_min, scene->params.vf_max, e->shard);
    }

    if (onlyUpdateVisibleList) useSwapping = false;
    if (!onlyUpdateVisibleList) {
        for (int t = 0; t < total; t++) {
            int vbaIdx, exlIdx;
            switch (allocType[t]) {
            case 1:
                vbaIdx = lastFreeBlock; lastFreeBlock--;
                if (vbaIdx >= 0) {
                    dsb200_hash_entry ne; memset(&ne, 0, sizeof(ne));
                    ne.pos[0] = blockCoords[t * 4]; ne.pos[1] = blockCoords[t * 4 + 1]; ne.pos[2] = blockCoords[t * 4 + 2];
                    ne.ptr = scene->alloc_list[vbaIdx]; ne.offset = 0;
                    ne.allocated_time = 0;   // reference leaves this uninitialised; defined as 0
                    hash[t] = ne;
                }
                break;
            case 2:
                vbaIdx = lastFreeBlock; lastFreeBlock--;
                exlIdx = lastFreeExcess; lastFreeExcess--;
                if (vbaIdx >= 0 && exlIdx >= 0) {
                    dsb200_hash_entry ne; memset(&ne, 0, sizeof(ne));
                    ne.pos[0] = blockCoords[t * 4]; ne.pos[1] = blockCoords[t * 4 + 1]; ne.pos[2] = blockCoords[t * 4 + 2];
                    ne.ptr = scene->alloc_list[vbaIdx]; ne.offset = 0; ne.allocated_time = 0;
                    int exlOffset = scene->excess_list[exlIdx];
                    hash[t].offset = exlOffset + 1;
                    hash[bucketNum + exlOffset] = ne;
                    visType[bucketNum + exlOffset] = 1;
                }
                break;
            default: break;
            }
        }
    }

    int noVisible = 0;
    for (int t = 0; t < total; t++) {
        unsigned char vt = visType[t];
        if (vt == 3) {
            bool vis, visEnl;
            blockVisibility(vis, visEnl, hash[t].pos, M, view->intr_d, voxelSize, W, H, useSwapping);
            if (useSwapping) { if (!visEnl) vt = 0; } else { if (!vis) vt = 0; }
            visType[t] = vt;
        }
        // (swapStates bookkeeping of the swapping engine is outside the boundary: SURVEY 2.1 #22)
        if (vt > 0) { if (noVisible < rs->visible_capacity) rs->visible_ids[noVisible] = t; noVisible++; }
    }

    if (useSwapping) { // reallocate swapped-out-but-visible, CPU .cpp:319-336
        for (int t = 0; t < total; t++) {
            if (visType[t] > 0 && hash[t].ptr == -1) {
                int vbaIdx = lastFreeBlock; lastFreeBlock--;
                if (vbaIdx >= 0) hash[t].ptr = scene->alloc_list[vbaIdx];
            }
        }
    }

    rs->no_visible = noVisible;
    scene->last_free_block = lastFreeBlock;
    scene->last_free_excess = lastFreeExcess;

    // --- CUDA-only tail: snapshot the visible list for decay / slide-window / de-fusion (.cu:257-317)
    SnapPtr snap;
    if (noVisible > 0) {
        snap = std::make_shared<Snapshot>();
        snap->ids.assign(rs->visible_ids, rs->visible_ids + noVisible);
    }
    if (!isDefusion) {
        if (!snap) snap = std::make_shared<Snapshot>();
        snap->frameIdx = e->frameIdx;
        e->snapshots[view->timestamp] = snap;
        e->decayQ.push_back(snap);
        e->slideQ.push_back(snap);
        e->frameIdx++;
    } else {
        std::map<double, SnapPtr>::iterator it = e->snapshots.find(view->timestamp);
        if (it != e->snapshots.end()) {
            if (!snap) snap = std::make_shared<Snapshot>();
            snap->frameIdx = it->second->frameIdx;
            e->decayDefusionQ.push_back(snap);
            e->slideDefusionQ.push_back(snap);
            e->snapshots[view->timestamp] = snap;
        }
    }
    if (scene->last_free_block < 0) return DSB200_ERR_VBA_EXHAUSTED;       // .cu:350-353
    if (scene->last_free_excess < 0) return DSB200_ERR_EXCESS_EXHAUSTED;   // .cu:354-358
    return DSB200_OK;
}

} // extern "C"

// ---------------------------------------------------------------------------------------------
// Integration / de-integration
// ---------------------------------------------------------------------------------------------
namespace {

inline float sdfToFloat(float x) { return (float)(x) / 32767.0f; }        // SDF_valueToFloat, ITMLibDefines.h:131
// SDF_floatToValue, :132.  gcc lowers (short)(float) to cvttss2si + truncation, i.e. (short)(int); spelled out so
// the GPU (whose direct float->short conversion saturates) can be written identically.
inline short sdfToShort(float x) { return (short)(int)((x) * 32767.0f); }

// float -> int as the CUDA hardware converts (cvt.rzi.s32.f32: NaN -> 0, out of range saturates).  The only implementation of
// de-integration in the reference is its CUDA engine, and de-fusing the last colour observation divides by zero there
// (DeviceAgnostic/ITMSceneReconstructionEngine.h:204-206), so the device's conversion of inf / NaN IS the reference behaviour
// (pinned by tests/test_gpu_refcuda.py); for finite in-range values it is the plain C cast.
inline int deviceFloatToInt(float v)
{
    if (v != v) return 0;
    if (v >= 2147483648.0f) return 2147483647;
    if (v <= -2147483648.0f) return (-2147483647 - 1);
    return (int)v;
}
// float -> unsigned char as nvcc lowers the reference's `(uchar)newW` (:214): cvt.rzi.u32.f32 (negative -> 0), low byte kept
inline unsigned char deviceFloatToUChar(float v)
{
    if (v != v || v <= 0.0f) return 0;
    if (v >= 4294967296.0f) return 255;
    return (unsigned char)(unsigned)v;
}

// ROUND (ORUtils/MathUtils.h:24) then (int), then CLAMP 0..255: Vector3::toUChar, ORUtils/Vector.h:245-247
inline unsigned char toUChar(float v)
{
    int i = deviceFloatToInt((v < 0) ? (v - 0.5f) : (v + 0.5f));
    int lo = (255 < i) ? 255 : i;       // MIN(255, i)
    return (unsigned char)((0 < lo) ? lo : 0);
}

// interpolateBilinear<Vector4u>, DA/ITMPixelUtils.h:11-39 (rgb channels only are consumed)
inline void bilinearRGB(const uint8_t *src, float px, float py, int W, float out[3])
{
    int ix = (int)floorf(px), iy = (int)floorf(py);
    float dx = px - (float)ix, dy = py - (float)iy;
    const uint8_t zero[4] = {0, 0, 0, 0};
    const uint8_t *a = src + 4 * ((size_t)ix + (size_t)iy * W), *b = zero, *c = zero, *d = zero;
    if (dx != 0) b = src + 4 * ((size_t)(ix + 1) + (size_t)iy * W);
    if (dy != 0) c = src + 4 * ((size_t)ix + (size_t)(iy + 1) * W);
    if (dx != 0 && dy != 0) d = src + 4 * ((size_t)(ix + 1) + (size_t)(iy + 1) * W);
    for (int k = 0; k < 3; k++)
        out[k] = ((float)a[k] * (1.0f - dx) * (1.0f - dy) + (float)b[k] * dx * (1.0f - dy) +
                  (float)c[k] * (1.0f - dx) * dy + (float)d[k] * dx * dy);
}

// computeUpdatedVoxelDepthInfo / computeDeUpdatedVoxelDepthInfo, DA/ITMSceneReconstructionEngine.h:17-77,117-181
template <bool DE>
inline float updateDepth(dsb200_voxel &vox, V4 pt, const float *M, const float *proj, float mu, int maxW,
                         const float *depth, int W, int H, const dsb200_weight_params &wp)
{
    V4 pc = mulM(M, pt);
    if (pc.z <= 0) return -1;
    float u = proj[0] * pc.x / pc.z + proj[2];
    float v = proj[1] * pc.y / pc.z + proj[3];
    if ((u < 1) || (u > W - 2) || (v < 1) || (v > H - 2)) return -1;
    float dm = depth[(int)(u + 0.5f) + (int)(v + 0.5f) * W];
    if (dm <= 0.0) return -1;
    float eta = dm - pc.z;
    if (eta < -mu) return eta;
    float oldF = sdfToFloat(vox.sdf); int oldW = vox.w_depth;
    float newF = (1.0f < eta / mu) ? 1.0f : eta / mu;        // MIN(1.0f, eta/mu)
    int newW;
    if (wp.depth_weighting) {
        newW = (int)((float)wp.max_distance / dm + 0.5);     // double add, :54
        if (newW < 1) newW = 1;
        if (newW > wp.max_new_w) newW = wp.max_new_w;
    } else newW = 1;
    if (!DE) { newF = oldW * oldF + newW * newF; newW = oldW + newW; }
    else     { newF = oldW * oldF - newW * newF; newW = oldW - newW; }
    if (DE && newW == 0) {
        // the reference divides by zero here and converts NaN/inf to short (UB); the
        // voxel is reset a few lines later anyway because w_depth becomes 0 -> define as reset.
        voxelReset(vox);
        return eta;
    }
    newF /= newW;
    newW = (newW < maxW) ? newW : maxW;
    vox.sdf = sdfToShort(newF);
    vox.w_depth = (unsigned char)newW;
    if (DE && vox.w_depth == 0) voxelReset(vox);
    return eta;
}

// computeUpdatedVoxelColorInfo / computeDeUpdatedVoxelColorInfo, DA/...:80-113,184-217
template <bool DE>
inline void updateColour(dsb200_voxel &vox, V4 pt, const float *Mrgb, const float *proj, unsigned char maxW,
                         const uint8_t *rgb, int W, int H)
{
    float oldW = (float)vox.w_color;
    float oldC[3] = {(float)vox.clr[0] / 255.0f, (float)vox.clr[1] / 255.0f, (float)vox.clr[2] / 255.0f};
    V4 pc = mulM(Mrgb, pt);
    float u = proj[0] * pc.x / pc.z + proj[2];
    float v = proj[1] * pc.y / pc.z + proj[3];
    if ((u < 1) || (u > W - 2) || (v < 1) || (v > H - 2)) return;
    float m[3]; bilinearRGB(rgb, u, v, W, m);
    float newW = 1;
    float newC[3];
    for (int k = 0; k < 3; k++) {
        float meas = m[k] / 255.0f;
        newC[k] = DE ? (oldC[k] * oldW - meas * newW) : (oldC[k] * oldW + meas * newW);
    }
    newW = DE ? (oldW - newW) : (oldW + newW);
    // DE with oldW == 1 divides by zero (+-inf, or NaN for 0/0) and DE with oldW == 0 leaves newW = -1: both are converted as
    // the reference's CUDA build converts them (see deviceFloatToInt / deviceFloatToUChar): a channel becomes 255, 0 or 0, and
    // the negative weight becomes 0
    for (int k = 0; k < 3; k++) newC[k] /= newW;
    newW = (newW < maxW) ? newW : (float)maxW;
    for (int k = 0; k < 3; k++) vox.clr[k] = toUChar(newC[k] * 255.0f);
    vox.w_color = deviceFloatToUChar(newW);
}

template <bool DE>
void fuseList(const rst_recon *e, dsb200_scene *scene, const dsb200_view *view, const dsb200_tracking *tracking,
              const int *ids, int n)
{
    const float voxelSize = scene->params.voxel_size, mu = scene->params.mu;
    const int maxW = scene->params.max_w;
    const bool stopMaxW = scene->params.stop_integrating_at_max_w != 0;
    const bool approx = !tracking->requires_full_rendering;          // CUDA .cu:407-420, 1431-1432
    float Mrgb[16];
    rst_mat4_multiply(view->rgb_to_depth_inv, tracking->M_d, Mrgb);   // CPU .cpp:67
    for (int i = 0; i < n; i++) {
        const dsb200_hash_entry &he = scene->hash[ids[i]];
        if (he.ptr < 0) continue;
        int gx = he.pos[0] * DSB200_BLOCK_SIZE, gy = he.pos[1] * DSB200_BLOCK_SIZE, gz = he.pos[2] * DSB200_BLOCK_SIZE;
        dsb200_voxel *blk = scene->voxels + (size_t)he.ptr * DSB200_BLOCK_SIZE3;
        for (int z = 0; z < 8; z++) for (int y = 0; y < 8; y++) for (int x = 0; x < 8; x++) {
            dsb200_voxel &vox = blk[x + y * 8 + z * 64];
            if (stopMaxW && vox.w_depth == maxW) continue;
            if (approx && vox.w_depth != 0) continue;
            V4 pt = {(float)(gx + x) * voxelSize, (float)(gy + y) * voxelSize, (float)(gz + z) * voxelSize, 1.0f};
            float eta = updateDepth<DE>(vox, pt, tracking->M_d, view->intr_d, mu, maxW, view->depth, view->depth_w,
                                        view->depth_h, e->weights);
            if ((eta > mu) || (fabsf(eta / mu) > 0.25f)) continue;    // DA/...:247
            updateColour<DE>(vox, pt, Mrgb, view->intr_rgb, (unsigned char)maxW, view->rgb, view->rgb_w, view->rgb_h);
        }
    }
}

} // namespace

extern "C" {

// IntegrateIntoScene: CPU .cpp:50-133; WeightParams + approximateIntegration as in CUDA .cu:363-421,1409-1442
void rst_integrate(rst_recon *e, dsb200_scene *scene, const dsb200_view *view, const dsb200_tracking *tracking,
                   const dsb200_render_state *rs)
{
    fuseList<false>(e, scene, view, tracking, rs->visible_ids, rs->no_visible);
}

// DeIntegrateIntoScene: CUDA .cu:426-489, 1483-1517 (replays the list saved for view->timestamp)
void rst_deintegrate(rst_recon *e, dsb200_scene *scene, const dsb200_view *view, const dsb200_tracking *tracking)
{
    std::map<double, SnapPtr>::iterator it = e->snapshots.find(view->timestamp);
    if (it == e->snapshots.end() || it->second->ids.empty()) return;
    fuseList<true>(e, scene, view, tracking, it->second->ids.data(), (int)it->second->ids.size());
}

} // extern "C"

// ---------------------------------------------------------------------------------------------
// Visualisation: visible blocks, expected depths, raycast, ICP maps, shading
// ---------------------------------------------------------------------------------------------
namespace {

const float FAR_AWAY = 999999.9f, VERY_CLOSE = 0.05f;   // DA/ITMVisualisationEngine.h:13-19
const int MAX_RENDERING_BLOCKS = 65536 * 4;             // :24
const int SUBSAMPLE = 8;                                 // minmaximg_subsample :26

struct BlockCache { int bx, by, bz, ptr; BlockCache() : bx(0x7fffffff), by(0x7fffffff), bz(0x7fffffff), ptr(-1) {} };

// Sharded scene (rst_raycast_sharded): while set, every voxel read goes to the scene of the shard that owns the block.
struct ShardSet { int n, brickLog2; const dsb200_scene *const *scenes; };
static ShardSet g_shardSet = {0, 0, nullptr};

// pointToVoxelBlockPos + readVoxel (cached), DA/ITMRepresentationAccess.h:14-19,176-210
inline dsb200_voxel readVoxel(const dsb200_scene *s, int px, int py, int pz, bool &found, BlockCache &c)
{
    int bx = ((px < 0) ? px - 8 + 1 : px) / 8, by = ((py < 0) ? py - 8 + 1 : py) / 8, bz = ((pz < 0) ? pz - 8 + 1 : pz) / 8;
    int lin = (px - bx * 8) + (py - by * 8) * 8 + (pz - bz * 8) * 64;
    if (g_shardSet.n > 1) s = g_shardSet.scenes[rst_shard_owner(bx, by, bz, g_shardSet.n, g_shardSet.brickLog2)];
    if (bx == c.bx && by == c.by && bz == c.bz) { found = true; return s->voxels[c.ptr + lin]; }
    int h = hashIndex(bx, by, bz, s->bucket_num);
    while (true) {
        const dsb200_hash_entry &e = s->hash[h];
        if (posEq(e, bx, by, bz) && e.ptr >= 0) {
            found = true; c.bx = bx; c.by = by; c.bz = bz; c.ptr = e.ptr * DSB200_BLOCK_SIZE3;
            return s->voxels[c.ptr + lin];
        }
        if (e.offset < 1) break;
        h = s->bucket_num + e.offset - 1;
    }
    found = false;
    dsb200_voxel v; voxelReset(v); v.pad_ = 0; return v;
}
inline dsb200_voxel readVoxelNC(const dsb200_scene *s, int px, int py, int pz, bool &found)
{ BlockCache c; return readVoxel(s, px, py, pz, found, c); }

inline float roundHalfAway(float x) { return (x < 0) ? (x - 0.5f) : (x + 0.5f); }  // ROUND

// readFromSDF_float_uninterpolated, DA/ITMRepresentationAccess.h:243-249
inline float sdfNearest(const dsb200_scene *s, V3 p, bool &found, BlockCache &c)
{
    dsb200_voxel v = readVoxel(s, (int)roundHalfAway(p.x), (int)roundHalfAway(p.y), (int)roundHalfAway(p.z), found, c);
    return sdfToFloat(v.sdf);
}

// readFromSDF_float_interpolated, :251-276 (mixes raw shorts, divides at the end)
inline float sdfTrilinear(const dsb200_scene *s, V3 p, bool &found, BlockCache &c)
{
    float fx = floorf(p.x), fy = floorf(p.y), fz = floorf(p.z);
    float cx = p.x - fx, cy = p.y - fy, cz = p.z - fz;
    int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    float res1, res2, v1, v2;
    v1 = readVoxel(s, ix, iy, iz, found, c).sdf;         v2 = readVoxel(s, ix + 1, iy, iz, found, c).sdf;
    res1 = (1.0f - cx) * v1 + cx * v2;
    v1 = readVoxel(s, ix, iy + 1, iz, found, c).sdf;     v2 = readVoxel(s, ix + 1, iy + 1, iz, found, c).sdf;
    res1 = (1.0f - cy) * res1 + cy * ((1.0f - cx) * v1 + cx * v2);
    v1 = readVoxel(s, ix, iy, iz + 1, found, c).sdf;     v2 = readVoxel(s, ix + 1, iy, iz + 1, found, c).sdf;
    res2 = (1.0f - cx) * v1 + cx * v2;
    v1 = readVoxel(s, ix, iy + 1, iz + 1, found, c).sdf; v2 = readVoxel(s, ix + 1, iy + 1, iz + 1, found, c).sdf;
    res2 = (1.0f - cy) * res2 + cy * ((1.0f - cx) * v1 + cx * v2);
    found = true;
    return sdfToFloat((1.0f - cz) * res1 + cz * res2);
}

// analysis hook (profiles/): when set, castRay adds its loop-trip count per pixel here
static int *g_rayStepCounts = nullptr; static int g_rayStepWidth = 0;
extern "C" void rst_debug_ray_steps(int *counts, int width) { g_rayStepCounts = counts; g_rayStepWidth = width; }
// ... and, per pixel, how many of those trips (0) found no block (8-voxel step), (1) took the minimum one-voxel step, (2) sampled
// tri-linearly, (3) found no block right after a trip that found none either, (4) stepped exactly mu / voxelSize voxels (an empty
// allocated block: sdf = 1), (5) did so right after a trip of kind (4)
static int *g_rayStepKinds = nullptr;
constexpr int kRayKinds = 6;
extern "C" void rst_debug_ray_kinds(int *kinds6) { g_rayStepKinds = kinds6; }

// castRay, DA/ITMVisualisationEngine.h:124-196
inline bool castRay(float *out, int x, int y, const dsb200_scene *s, const float *invM, const float *invProj,
                    float oneOverVoxel, float mu, float zmin, float zmax)
{
    V4 pc; V3 ps, pe, dir, pr;
    bool ptFound, hashFound;
    float sdf = 1.0f, totalLength, stepLength, totalLengthMax, stepScale = mu * oneOverVoxel;
    pc.z = zmin;
    pc.x = pc.z * (((float)x - invProj[2]) * invProj[0]);
    pc.y = pc.z * (((float)y - invProj[3]) * invProj[1]);
    pc.w = 1.0f;
    totalLength = sqrtf(pc.x * pc.x + pc.y * pc.y + pc.z * pc.z) * oneOverVoxel;
    V4 t = mulM(invM, pc);
    ps.x = t.x * oneOverVoxel; ps.y = t.y * oneOverVoxel; ps.z = t.z * oneOverVoxel;
    pc.z = zmax;
    pc.x = pc.z * (((float)x - invProj[2]) * invProj[0]);
    pc.y = pc.z * (((float)y - invProj[3]) * invProj[1]);
    pc.w = 1.0f;
    totalLengthMax = sqrtf(pc.x * pc.x + pc.y * pc.y + pc.z * pc.z) * oneOverVoxel;
    t = mulM(invM, pc);
    pe.x = t.x * oneOverVoxel; pe.y = t.y * oneOverVoxel; pe.z = t.z * oneOverVoxel;
    dir.x = pe.x - ps.x; dir.y = pe.y - ps.y; dir.z = pe.z - ps.z;
    float dn = 1.0f / sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
    dir.x *= dn; dir.y *= dn; dir.z *= dn;
    pr = ps;
    BlockCache cache;
    int prevKind = -1;     // analysis hook only
    while (totalLength < totalLengthMax) {
        if (g_rayStepCounts) g_rayStepCounts[x + y * g_rayStepWidth]++;
        sdf = sdfNearest(s, pr, hashFound, cache);
        int *kinds = g_rayStepKinds ? g_rayStepKinds + kRayKinds * (x + y * g_rayStepWidth) : nullptr;
        if (!hashFound) { stepLength = DSB200_BLOCK_SIZE; if (kinds) { kinds[0]++; if (prevKind == 0) kinds[3]++; prevKind = 0; } }
        else {
            if ((sdf <= 0.1f) && (sdf >= -0.5f)) { sdf = sdfTrilinear(s, pr, hashFound, cache); if (kinds) kinds[2]++; }
            if (!(sdf > 0.0f)) break;
            stepLength = (sdf * stepScale < 1.0f) ? 1.0f : sdf * stepScale;   // MAX(sdf*stepScale, 1.0f)
            if (kinds) {
                if (stepLength == 1.0f) kinds[1]++;
                if (stepLength == stepScale) { kinds[4]++; if (prevKind == 4) kinds[5]++; prevKind = 4; } else prevKind = -1;
            }
        }
        pr.x += stepLength * dir.x; pr.y += stepLength * dir.y; pr.z += stepLength * dir.z;
        totalLength += stepLength;
    }
    if (sdf <= 0.0f) {
        stepLength = sdf * stepScale;
        pr.x += stepLength * dir.x; pr.y += stepLength * dir.y; pr.z += stepLength * dir.z;
        sdf = sdfTrilinear(s, pr, ptFound, cache);
        stepLength = sdf * stepScale;
        pr.x += stepLength * dir.x; pr.y += stepLength * dir.y; pr.z += stepLength * dir.z;
    } else ptFound = false;
    out[0] = pr.x; out[1] = pr.y; out[2] = pr.z; out[3] = ptFound ? 1.0f : 0.0f;
    return ptFound;
}

// GenericRaycast, CPU/ITMVisualisationEngine_CPU.cpp:296-333
void genericRaycast(const dsb200_scene *s, int W, int H, const float *invM, const float *intr,
                    const dsb200_render_state *rs)
{
    float invProj[4] = {1.0f / intr[0], 1.0f / intr[1], intr[2], intr[3]};
    float oneOverVoxel = 1.0f / s->params.voxel_size;
    for (int loc = 0; loc < W * H; loc++) {
        int y = loc / W, x = loc - y * W;
        int loc2 = (int)floorf((float)x / SUBSAMPLE) + (int)floorf((float)y / SUBSAMPLE) * W;
        castRay(rs->raycast_result + 4 * (size_t)loc, x, y, s, invM, invProj, oneOverVoxel, s->params.mu,
                rs->range_image[2 * loc2], rs->range_image[2 * loc2 + 1]);
    }
}

// ProjectSingleBlock, DA/ITMVisualisationEngine.h:36-85
inline bool projectBlock(const int16_t *pos, const float *M, const float *intr, int W, int H, float voxelSize,
                         int ul[2], int lr[2], float zr[2])
{
    ul[0] = W / SUBSAMPLE; ul[1] = H / SUBSAMPLE; lr[0] = -1; lr[1] = -1; zr[0] = FAR_AWAY; zr[1] = VERY_CLOSE;
    for (int corner = 0; corner < 8; ++corner) {
        short tx = pos[0], ty = pos[1], tz = pos[2];
        tx += (corner & 1) ? 1 : 0; ty += (corner & 2) ? 1 : 0; tz += (corner & 4) ? 1 : 0;
        V4 p = {(float)tx * (float)DSB200_BLOCK_SIZE * voxelSize, (float)ty * (float)DSB200_BLOCK_SIZE * voxelSize,
                (float)tz * (float)DSB200_BLOCK_SIZE * voxelSize, 1.0f};
        p = mulM(M, p);
        if (p.z < 1e-6) continue;                       // double literal, SURVEY C.4
        float px = (intr[0] * p.x / p.z + intr[2]) / SUBSAMPLE;
        float py = (intr[1] * p.y / p.z + intr[3]) / SUBSAMPLE;
        if (ul[0] > floorf(px)) ul[0] = (int)floorf(px);
        if (lr[0] < ceilf(px)) lr[0] = (int)ceilf(px);
        if (ul[1] > floorf(py)) ul[1] = (int)floorf(py);
        if (lr[1] < ceilf(py)) lr[1] = (int)ceilf(py);
        if (zr[0] > p.z) zr[0] = p.z;
        if (zr[1] < p.z) zr[1] = p.z;
    }
    if (ul[0] < 0) ul[0] = 0;
    if (ul[1] < 0) ul[1] = 0;
    if (lr[0] >= W) lr[0] = W - 1;
    if (lr[1] >= H) lr[1] = H - 1;
    if (ul[0] > lr[0]) return false;
    if (ul[1] > lr[1]) return false;
    if (zr[0] < VERY_CLOSE) zr[0] = VERY_CLOSE;
    if (zr[1] < VERY_CLOSE) return false;
    return true;
}

} // namespace

extern "C" {

// FindVisibleBlocks, CPU/ITMVisualisationEngine_CPU.cpp:40-78
void rst_find_visible_blocks(const dsb200_scene *scene, const float *M, const float *intr, dsb200_render_state *rs)
{
    int total = scene->bucket_num + scene->excess_size, n = 0;
    for (int t = 0; t < total; t++) {
        const dsb200_hash_entry &e = scene->hash[t];
        bool vis = false, visEnl = false;
        if (e.ptr >= 0) blockVisibility(vis, visEnl, e.pos, M, intr, scene->params.voxel_size, rs->width, rs->height, false);
        if (vis) { if (n < rs->visible_capacity) rs->visible_ids[n] = t; n++; }
    }
    rs->no_visible = n;
}

// CountVisibleBlocks, :81-97
int rst_count_visible_blocks(const dsb200_scene *scene, const dsb200_render_state *rs, int minId, int maxId)
{
    int ret = 0;
    for (int i = 0; i < rs->no_visible; i++) {
        int id = scene->hash[rs->visible_ids[i]].ptr;
        if (id >= minId && id <= maxId) ++ret;
    }
    return ret;
}

// CreateExpectedDepths, :116-180 (+ CreateRenderingBlocks DA/...:92-112 folded into the fill loop)
void rst_create_expected_depths(const dsb200_scene *scene, const float *M, const float *intr, dsb200_render_state *rs)
{
    const int W = rs->width, H = rs->height;
    for (int i = 0; i < W * H; i++) { rs->range_image[2 * i] = FAR_AWAY; rs->range_image[2 * i + 1] = VERY_CLOSE; }
    int numRenderingBlocks = 0;
    for (int b = 0; b < rs->no_visible; b++) {
        const dsb200_hash_entry &e = scene->hash[rs->visible_ids[b]];
        int ul[2], lr[2]; float zr[2];
        if (e.ptr < 0) continue;
        if (!projectBlock(e.pos, M, intr, W, H, scene->params.voxel_size, ul, lr, zr)) continue;
        int nx = (int)ceilf((float)(lr[0] - ul[0] + 1) / 16.0f), ny = (int)ceilf((float)(lr[1] - ul[1] + 1) / 16.0f);
        int required = nx * ny;
        if (numRenderingBlocks + required >= MAX_RENDERING_BLOCKS) continue;   // :158
        numRenderingBlocks += required;
        // the tiles of one block jointly cover exactly [ul, lr]; min/max fill is tile-order independent
        for (int y = ul[1]; y <= lr[1]; y++) for (int x = ul[0]; x <= lr[0]; x++) {
            float *px = rs->range_image + 2 * ((size_t)x + (size_t)y * W);
            if (px[0] > zr[0]) px[0] = zr[0];
            if (px[1] < zr[1]) px[1] = zr[1];
        }
    }
}

// FindSurface, :562-567
void rst_find_surface(const dsb200_scene *scene, const float *invM, const float *intr, dsb200_render_state *rs)
{
    genericRaycast(scene, rs->width, rs->height, invM, intr, rs);
}

} // extern "C"

// ---------------------------------------------------------------------------------------------
// ICP maps + free-view shading
// ---------------------------------------------------------------------------------------------
namespace {

// computeNormalAndAngle<useSmoothing>, DA/ITMVisualisationEngine.h:243-312
template <bool SMOOTH>
inline void normalAndAngleImage(bool &found, int x, int y, const float *pts, const float light[3], float voxelSize,
                                int W, int H, float n[3], float &angle)
{
    if (!found) return;
    const float *xp, *yp, *xm, *ym;
    if (SMOOTH) {
        if (y <= 2 || y >= H - 3 || x <= 2 || x >= W - 3) { found = false; return; }
        xp = pts + 4 * ((x + 2) + y * W); yp = pts + 4 * (x + (y + 2) * W);
        xm = pts + 4 * ((x - 2) + y * W); ym = pts + 4 * (x + (y - 2) * W);
    } else {
        if (y <= 1 || y >= H - 2 || x <= 1 || x >= W - 2) { found = false; return; }
        xp = pts + 4 * ((x + 1) + y * W); yp = pts + 4 * (x + (y + 1) * W);
        xm = pts + 4 * ((x - 1) + y * W); ym = pts + 4 * (x + (y - 1) * W);
    }
    float dx[3] = {0, 0, 0}, dy[3] = {0, 0, 0};
    bool doPlus1 = false;
    if (xp[3] <= 0 || yp[3] <= 0 || xm[3] <= 0 || ym[3] <= 0) doPlus1 = true;
    else {
        for (int k = 0; k < 3; k++) { dx[k] = xp[k] - xm[k]; dy[k] = yp[k] - ym[k]; }
        float a = dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2], b = dy[0] * dy[0] + dy[1] * dy[1] + dy[2] * dy[2];
        float lengthDiff = (a < b) ? b : a;
        if (lengthDiff * voxelSize * voxelSize > (0.15f * 0.15f)) doPlus1 = true;
    }
    if (doPlus1) {
        if (SMOOTH) {
            xp = pts + 4 * ((x + 1) + y * W); yp = pts + 4 * (x + (y + 1) * W);
            xm = pts + 4 * ((x - 1) + y * W); ym = pts + 4 * (x + (y - 1) * W);
            for (int k = 0; k < 3; k++) { dx[k] = xp[k] - xm[k]; dy[k] = yp[k] - ym[k]; }
        }
        if (xp[3] <= 0 || yp[3] <= 0 || xm[3] <= 0 || ym[3] <= 0) { found = false; return; }
    }
    n[0] = -(dx[1] * dy[2] - dx[2] * dy[1]);
    n[1] = -(dx[2] * dy[0] - dx[0] * dy[2]);
    n[2] = -(dx[0] * dy[1] - dx[1] * dy[0]);
    float sc = 1.0f / sqrtf(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
    n[0] *= sc; n[1] *= sc; n[2] *= sc;
    angle = n[0] * light[0] + n[1] * light[1] + n[2] * light[2];
    if (!(angle > 0.0)) found = false;
}

inline unsigned char greyOf(float angle) { return (unsigned char)((0.8f * angle + 0.2f) * 255.0f); }  // drawPixelGrey :314-318

// computeSingleNormalFromSDF, DA/ITMRepresentationAccess.h:314-427
inline void normalFromSDF(const dsb200_scene *s, const float p[3], float ret[3])
{
    bool f;
    float flx = floorf(p[0]), fly = floorf(p[1]), flz = floorf(p[2]);
    float cx = p[0] - flx, cy = p[1] - fly, cz = p[2] - flz;
    int X = (int)flx, Y = (int)fly, Z = (int)flz;
    float nx = 1.0f - cx, ny = 1.0f - cy, nz = 1.0f - cz;
#define RV(dx, dy, dz) ((float)readVoxelNC(s, X + (dx), Y + (dy), Z + (dz), f).sdf)
    float fr[4] = {RV(0, 0, 0), RV(1, 0, 0), RV(0, 1, 0), RV(1, 1, 0)};
    float bk[4] = {RV(0, 0, 1), RV(1, 0, 1), RV(0, 1, 1), RV(1, 1, 1)};
    float t[4], p1, p2, v1;
    // x
    p1 = fr[0] * ny * nz + fr[2] * cy * nz + bk[0] * ny * cz + bk[2] * cy * cz;
    t[0] = RV(-1, 0, 0); t[1] = RV(-1, 1, 0); t[2] = RV(-1, 0, 1); t[3] = RV(-1, 1, 1);
    p2 = t[0] * ny * nz + t[1] * cy * nz + t[2] * ny * cz + t[3] * cy * cz;
    v1 = p1 * cx + p2 * nx;
    p1 = fr[1] * ny * nz + fr[3] * cy * nz + bk[1] * ny * cz + bk[3] * cy * cz;
    t[0] = RV(2, 0, 0); t[1] = RV(2, 1, 0); t[2] = RV(2, 0, 1); t[3] = RV(2, 1, 1);
    p2 = t[0] * ny * nz + t[1] * cy * nz + t[2] * ny * cz + t[3] * cy * cz;
    ret[0] = sdfToFloat(p1 * nx + p2 * cx - v1);
    // y
    p1 = fr[0] * nx * nz + fr[1] * cx * nz + bk[0] * nx * cz + bk[1] * cx * cz;
    t[0] = RV(0, -1, 0); t[1] = RV(1, -1, 0); t[2] = RV(0, -1, 1); t[3] = RV(1, -1, 1);
    p2 = t[0] * nx * nz + t[1] * cx * nz + t[2] * nx * cz + t[3] * cx * cz;
    v1 = p1 * cy + p2 * ny;
    p1 = fr[2] * nx * nz + fr[3] * cx * nz + bk[2] * nx * cz + bk[3] * cx * cz;
    t[0] = RV(0, 2, 0); t[1] = RV(1, 2, 0); t[2] = RV(0, 2, 1); t[3] = RV(1, 2, 1);
    p2 = t[0] * nx * nz + t[1] * cx * nz + t[2] * nx * cz + t[3] * cx * cz;
    ret[1] = sdfToFloat(p1 * ny + p2 * cy - v1);
    // z
    p1 = fr[0] * nx * ny + fr[1] * cx * ny + fr[2] * nx * cy + fr[3] * cx * cy;
    t[0] = RV(0, 0, -1); t[1] = RV(1, 0, -1); t[2] = RV(0, 1, -1); t[3] = RV(1, 1, -1);
    p2 = t[0] * nx * ny + t[1] * cx * ny + t[2] * nx * cy + t[3] * cx * cy;
    v1 = p1 * cz + p2 * nz;
    p1 = bk[0] * nx * ny + bk[1] * cx * ny + bk[2] * nx * cy + bk[3] * cx * cy;
    t[0] = RV(0, 0, 2); t[1] = RV(1, 0, 2); t[2] = RV(0, 1, 2); t[3] = RV(1, 1, 2);
    p2 = t[0] * nx * ny + t[1] * cx * ny + t[2] * nx * cy + t[3] * cx * cy;
    ret[2] = sdfToFloat(p1 * nz + p2 * cz - v1);
#undef RV
}

// computeNormalAndAngle<TVoxel,TIndex>, DA/ITMVisualisationEngine.h:220-234
inline void normalAndAngleSDF(bool &found, const float p[3], const dsb200_scene *s, const float light[3], float n[3],
                              float &angle)
{
    if (!found) return;
    normalFromSDF(s, p, n);
    float sc = 1.0f / sqrtf(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
    n[0] *= sc; n[1] *= sc; n[2] *= sc;
    angle = n[0] * light[0] + n[1] * light[1] + n[2] * light[2];
    if (!(angle > 0.0)) found = false;
}

// readFromSDF_color4u_interpolated, DA/ITMRepresentationAccess.h:278-312
inline void colourTrilinear(const dsb200_scene *s, const float p[3], float out[4])
{
    BlockCache c; bool f;
    float flx = floorf(p[0]), fly = floorf(p[1]), flz = floorf(p[2]);
    float cx = p[0] - flx, cy = p[1] - fly, cz = p[2] - flz;
    int X = (int)flx, Y = (int)fly, Z = (int)flz;
    float ret[3] = {0.0f, 0.0f, 0.0f};
    const int off[8][3] = {{0,0,0},{1,0,0},{0,1,0},{1,1,0},{0,0,1},{1,0,1},{0,1,1},{1,1,1}};
    for (int i = 0; i < 8; i++) {
        dsb200_voxel v = readVoxel(s, X + off[i][0], Y + off[i][1], Z + off[i][2], f, c);
        float w = (off[i][0] ? cx : (1.0f - cx)) * (off[i][1] ? cy : (1.0f - cy)) * (off[i][2] ? cz : (1.0f - cz));
        for (int k = 0; k < 3; k++) ret[k] += w * (float)v.clr[k];
    }
    out[0] = ret[0] / 255.0f; out[1] = ret[1] / 255.0f; out[2] = ret[2] / 255.0f; out[3] = 255.0f / 255.0f;
}

} // namespace

extern "C" {

// CreateICPMaps, CPU/ITMVisualisationEngine_CPU.cpp:432-459 ; processPixelICP<true> DA/...:391-439
void rst_create_icp_maps(const dsb200_scene *scene, const dsb200_view *view, dsb200_tracking *tracking,
                         dsb200_render_state *rs)
{
    genericRaycast(scene, rs->width, rs->height, tracking->invM_d, view->intr_d, rs);
    rst_icp_maps_from_raycast(scene, view, tracking, rs->raycast_result, rs->raycast_image, rs->width, rs->height);
}

// second half of CreateICPMaps on a given point image (the sharded path composites the shards' raycasts first)
void rst_icp_maps_from_raycast(const dsb200_scene *scene, const dsb200_view *view, dsb200_tracking *tracking,
                               const float *raycastResult, uint8_t *outRendering, int W, int H)
{
    (void)view;
    const float *invM = tracking->invM_d;
    float light[3] = {-invM[8], -invM[9], -invM[10]};     // -invM.getColumn(2)
    float vs = scene->params.voxel_size;
    for (int y = 0; y < H; y++) for (int x = 0; x < W; x++) {
        int loc = x + y * W;
        const float *pt = raycastResult + 4 * (size_t)loc;
        bool found = pt[3] > 0.0f;
        float n[3], angle = 0.0f;
        normalAndAngleImage<true>(found, x, y, raycastResult, light, vs, W, H, n, angle);
        float *P = tracking->icp_points + 4 * (size_t)loc, *N = tracking->icp_normals + 4 * (size_t)loc;
        uint8_t *R = outRendering + 4 * (size_t)loc;
        if (found) {
            unsigned char g = greyOf(angle);
            R[0] = R[1] = R[2] = R[3] = g;
            P[0] = pt[0] * vs; P[1] = pt[1] * vs; P[2] = pt[2] * vs; P[3] = 1.0f;
            N[0] = n[0]; N[1] = n[1]; N[2] = n[2]; N[3] = 0.0f;
        } else {
            P[0] = P[1] = P[2] = 0.0f; P[3] = -1.0f;
            N[0] = N[1] = N[2] = 0.0f; N[3] = -1.0f;
            R[0] = R[1] = R[2] = R[3] = 0;
        }
    }
}

// Raycast of the whole sharded scene (definition in include/dsb200.h, dsb200_raycast_sharded): expected depths merged
// by min / max over the shards' range images, then the unchanged castRay with owner-routed voxel reads.
void rst_raycast_sharded(int n, const dsb200_scene *const *scenes, const float *const *rangeImages, int brickLog2,
                         const float *invM, const float *intr, int W, int H, float *mergedRange, float *out)
{
    const int count = W * (H / SUBSAMPLE + 1);
    for (int i = 0; i < count; i++) {
        float lo = FAR_AWAY, hi = VERY_CLOSE;
        for (int r = 0; r < n; r++) {
            float a = rangeImages[r][2 * i], b = rangeImages[r][2 * i + 1];
            if (a < lo) lo = a;
            if (b > hi) hi = b;
        }
        mergedRange[2 * i] = lo; mergedRange[2 * i + 1] = hi;
    }
    dsb200_render_state rs; memset(&rs, 0, sizeof(rs));
    rs.range_image = mergedRange; rs.raycast_result = out; rs.width = W; rs.height = H;
    g_shardSet.n = n; g_shardSet.brickLog2 = brickLog2; g_shardSet.scenes = scenes;
    genericRaycast(scenes[0], W, H, invM, intr, &rs);
    g_shardSet.n = 0; g_shardSet.scenes = nullptr;
}

// nearest-hit composite of the shards' raycast results (definition in include/dsb200.h, dsb200_composite_nearest)
void rst_composite_nearest(int n, const float *const *results, const float *invM, float voxelSize, int W, int H,
                           int rowBegin, int rowEnd, float *out)
{
    (void)H;
    const float k = 1.0f / voxelSize;
    const float cx = invM[12] * k, cy = invM[13] * k, cz = invM[14] * k;
    for (int i = rowBegin * W; i < rowEnd * W; i++) {
        float best[4] = {0.0f, 0.0f, 0.0f, 0.0f}, bestD = 0.0f;
        bool have = false;
        for (int r = 0; r < n; r++) {
            const float *q = results[r] + 4 * (size_t)i;
            if (q[3] > 0.0f) {
                float dx = q[0] - cx, dy = q[1] - cy, dz = q[2] - cz;
                float d = dx * dx + dy * dy + dz * dz;
                if (!have || d < bestD) { memcpy(best, q, 16); bestD = d; have = true; }
            }
        }
        memcpy(out + 4 * (size_t)i, best, 16);
    }
}

// RenderImage: CPU :339-405; RENDER_DEPTH_MAP as in CUDA/ITMVisualisationEngine_CUDA.cu:418-434
// (depth render falls through into the grey render).
void rst_render_image(const dsb200_scene *scene, const float *M, const float *invM, const float *intr,
                      dsb200_render_state *rs, uint8_t *outRGBA, float *outDepth, int type)
{
    const int W = rs->width, H = rs->height;
    genericRaycast(scene, W, H, invM, intr, rs);
    float light[3] = {-invM[8], -invM[9], -invM[10]};
    float vs = scene->params.voxel_size;
    if (type == DSB200_RENDER_COLOURCODED) type = DSB200_RENDER_SHADED_GREYSCALE;   // !TIndex::hasColorCoding
    for (int loc = 0; loc < W * H; loc++) {
        const float *pt = rs->raycast_result + 4 * (size_t)loc;
        bool found = pt[3] > 0;
        float n[3], angle = 0.0f;
        uint8_t *o = outRGBA + 4 * (size_t)loc;
        if (type == DSB200_RENDER_DEPTH_MAP && outDepth) {
            // processPixelColourDepth / drawPixelDepth, DA/...:340-357,527-540
            if (found) {
                V4 ph = {pt[0] * vs, pt[1] * vs, pt[2] * vs, 1.0f};
                V4 pc = mulM(M, ph);
                outDepth[loc] = pc.z / pc.w;
            } else outDepth[loc] = 0.0f;
        }
        bool f2 = found;
        normalAndAngleSDF(f2, pt, scene, light, n, angle);
        if (!f2) { o[0] = o[1] = o[2] = o[3] = 0; continue; }
        switch (type) {
        case DSB200_RENDER_COLOUR_FROM_VOLUME: {
            float c[4]; colourTrilinear(scene, pt, c);
            o[0] = (uint8_t)(c[0] * 255.0f); o[1] = (uint8_t)(c[1] * 255.0f); o[2] = (uint8_t)(c[2] * 255.0f); o[3] = 255;
            break; }
        case DSB200_RENDER_COLOUR_FROM_NORMAL:    // drawPixelNormal :320-325 (alpha untouched by the reference)
            o[0] = (uint8_t)((0.3f + (-n[0] + 1.0f) * 0.35f) * 255.0f);
            o[1] = (uint8_t)((0.3f + (-n[1] + 1.0f) * 0.35f) * 255.0f);
            o[2] = (uint8_t)((0.3f + (-n[2] + 1.0f) * 0.35f) * 255.0f);
            break;
        default: { unsigned char g = greyOf(angle); o[0] = o[1] = o[2] = o[3] = g; }
        }
    }
}

} // extern "C"

// ---------------------------------------------------------------------------------------------
// Voxel decay + sliding window (reference: CUDA only; parity UNPINNED by reference outputs)
//
// Deterministic schedule (DESIGN.md "pruning schedule"): the reference launches one CTA per
// snapshot entry; each CTA resets voxels, counts empties and, if the block emptied, thread 0
// unlinks it under a per-bucket try-lock (CUDA .cu:501-598, 601-754, 769-847).  The result
// depends on CTA interleaving.  We fix ONE valid interleaving: phase A = every CTA runs its voxel
// part against the table as it was at launch; phase B = the deletions run one at a time in
// snapshot order (so no lock is ever contended and no deletion is skipped).
// ---------------------------------------------------------------------------------------------
namespace {

// findVoxel(..., outHashIdx, outPrevHashIdx), DA/ITMRepresentationAccess.h:87-131
inline int findBlockWithPrev(const dsb200_scene *s, int bx, int by, int bz, int &prev)
{
    prev = -1;
    int h = hashIndex(bx, by, bz, s->bucket_num);
    while (true) {
        const dsb200_hash_entry &e = s->hash[h];
        if (posEq(e, bx, by, bz) && e.ptr >= 0) return h;
        if (e.offset < 1) break;
        prev = h;
        h = s->bucket_num + e.offset - 1;
    }
    return -1;
}

// deleteBlock, CUDA .cu:501-598 (lock elided: serial)
inline bool deleteBlock(dsb200_scene *s, unsigned char *visType, int bx, int by, int bz, int &lastFree)
{
    int prev, idx = findBlockWithPrev(s, bx, by, bz, prev);
    if (idx < 0) return false;   // reference would index hashTable[-1]; cannot happen under this schedule
    dsb200_hash_entry *hash = s->hash;
    lastFree++;
    s->alloc_list[lastFree] = hash[idx].ptr;
    // excess-list slots are NOT returned (reference TODO .cu:560-561)
    if (prev == -1) {
        if (hash[idx].offset >= 1) {
            int next = s->bucket_num + hash[idx].offset - 1;
            hash[idx] = hash[next];
            visType[idx] = visType[next]; visType[next] = 0;
            hash[next].offset = 0; hash[next].ptr = -2;
        } else {
            hash[idx].ptr = -2; visType[idx] = 0;
        }
    } else {
        hash[prev].offset = hash[idx].offset;
        hash[idx].offset = 0; hash[idx].ptr = -2;
        visType[prev] = visType[idx]; visType[idx] = 0;
    }
    return true;
}

// PartialDecay (.cu:880-909) and PartialSlideWindow (.cu:850-877) share this body.
void pruneSnapshot(rst_recon *e, dsb200_scene *s, dsb200_render_state *rs, const Snapshot &snap, bool slide,
                   int minAge, int maxWeight)
{
    struct Pending { short x, y, z; };
    std::vector<Pending> pending;
    int currentFrame = (int)e->frameIdx;
    // phase A
    for (size_t i = 0; i < snap.ids.size(); i++) {
        int entryId = snap.ids[i];
        if (entryId < 0) continue;
        const dsb200_hash_entry &cur = s->hash[entryId];
        if (cur.ptr < 0) continue;
        int prev, idx = findBlockWithPrev(s, cur.pos[0], cur.pos[1], cur.pos[2], prev);
        if (idx < 0) continue;
        bool safe = true;
        if (!slide) { int age = currentFrame - s->hash[idx].allocated_time; if (age < minAge) safe = false; }  // .cu:630-636
        int empties = 0;
        if (safe) {
            dsb200_voxel *blk = s->voxels + (size_t)s->hash[idx].ptr * DSB200_BLOCK_SIZE3;
            for (int v = 0; v < DSB200_BLOCK_SIZE3; v++) {
                if (slide) { voxelReset(blk[v]); empties++; }                      // deleteVoxel .cu:715-727
                else {                                                              // decayVoxel .cu:638-650
                    if (blk[v].w_depth <= maxWeight && blk[v].w_depth > 0) voxelReset(blk[v]);
                    if (blk[v].w_depth == 0) empties++;
                }
            }
        }
        if (safe && empties == DSB200_BLOCK_SIZE3) { Pending p = {cur.pos[0], cur.pos[1], cur.pos[2]}; pending.push_back(p); }
    }
    // phase B
    int lastFree = s->last_free_block;
    for (size_t i = 0; i < pending.size(); i++)
        deleteBlock(s, rs->visible_type, pending[i].x, pending[i].y, pending[i].z, lastFree);
    s->last_free_block = lastFree;
}

void pruneFromQueue(rst_recon *e, dsb200_scene *s, dsb200_render_state *rs, std::deque<SnapPtr> &q, long threshold,
                    bool slide, int minAge, int maxWeight)
{
    int oldLastFree = s->last_free_block;
    if ((long)q.size() > threshold) {
        SnapPtr snap = q.front(); q.pop_front();
        if (!snap->ids.empty()) pruneSnapshot(e, s, rs, *snap, slide, minAge, maxWeight);
    }
    e->totalDecayed += s->last_free_block - oldLastFree;
}

} // namespace

extern "C" {

// Decay, CUDA .cu:913-952 (forceAllVoxels accepted and ignored)
void rst_decay(rst_recon *e, dsb200_scene *s, dsb200_render_state *rs, int maxWeight, int minAge, int)
{ pruneFromQueue(e, s, rs, e->decayQ, minAge, false, minAge, maxWeight); }
// DecayDefusionPart, .cu:957-996
void rst_decay_defusion(rst_recon *e, dsb200_scene *s, dsb200_render_state *rs, int maxWeight, int minAge, int)
{ pruneFromQueue(e, s, rs, e->decayDefusionQ, 0, false, minAge, maxWeight); }
// SlideWindow, .cu:1001-1037
void rst_slide_window(rst_recon *e, dsb200_scene *s, dsb200_render_state *rs, int maxAge)
{ pruneFromQueue(e, s, rs, e->slideQ, maxAge, true, 0, 0); }
// SlideWindowDefusionPart, .cu:1041-1077
void rst_slide_window_defusion(rst_recon *e, dsb200_scene *s, dsb200_render_state *rs, int maxAge, int maxSize)
{ (void)maxAge; pruneFromQueue(e, s, rs, e->slideDefusionQ, maxSize, true, 0, 0); }

// ForwardRender, CPU/ITMVisualisationEngine_CPU.cpp:464-540 (only reached with useApproximateRaycast)
void rst_forward_render(const dsb200_scene *scene, const dsb200_view *view, dsb200_tracking *tracking,
                        dsb200_render_state *rs, float *fwd, int *missing, int *noMissing)
{
    const int W = rs->width, H = rs->height;
    const float *M = tracking->M_d, *invM = tracking->invM_d, *proj = view->intr_d;
    float invProj[4] = {1.0f / proj[0], 1.0f / proj[1], proj[2], proj[3]};
    float light[3] = {-invM[8], -invM[9], -invM[10]};
    float vs = scene->params.voxel_size;
    memset(fwd, 0, sizeof(float) * 4 * (size_t)W * H);
    for (int loc = 0; loc < W * H; loc++) {
        const float *px = rs->raycast_result + 4 * (size_t)loc;
        V4 p = {px[0] * vs, px[1] * vs, px[2] * vs, 1.0f};      // forwardProjectPixel DA/...:202-215 (pixel.w = 1)
        p = mulM(M, p);
        float u = proj[0] * p.x / p.z + proj[2], v = proj[1] * p.y / p.z + proj[3];
        if (!(u >= 0 && u <= W - 1 && v >= 0 && v <= H - 1)) continue;   // also skips NaN (reference: undefined index)
        int ln = (int)(u + 0.5f) + (int)(v + 0.5f) * W;
        memcpy(fwd + 4 * (size_t)ln, px, 16);
    }
    int n = 0;
    for (int y = 0; y < H; y++) for (int x = 0; x < W; x++) {
        int loc = x + y * W;
        int loc2 = (int)floorf((float)x / SUBSAMPLE) + (int)floorf((float)y / SUBSAMPLE) * W;
        const float *f = fwd + 4 * (size_t)loc;
        float zmin = rs->range_image[2 * loc2], zmax = rs->range_image[2 * loc2 + 1], d = view->depth[loc];
        if ((f[3] <= 0) && ((f[0] == 0 && f[1] == 0 && f[2] == 0) || (d >= 0)) && (zmin < zmax)) missing[n++] = loc;
    }
    *noMissing = n;
    for (int i = 0; i < n; i++) {
        int loc = missing[i], y = loc / W, x = loc - y * W;
        int loc2 = (int)floorf((float)x / SUBSAMPLE) + (int)floorf((float)y / SUBSAMPLE) * W;
        castRay(fwd + 4 * (size_t)loc, x, y, scene, invM, invProj, 1.0f / vs, scene->params.mu, rs->range_image[2 * loc2],
                rs->range_image[2 * loc2 + 1]);
    }
    for (int y = 0; y < H; y++) for (int x = 0; x < W; x++) {     // processPixelForwardRender<true> DA/...:441-456
        int loc = x + y * W;
        bool found = fwd[4 * (size_t)loc + 3] > 0.0f;
        float nrm[3], angle = 0.0f;
        normalAndAngleImage<true>(found, x, y, fwd, light, vs, W, H, nrm, angle);
        uint8_t *R = rs->raycast_image + 4 * (size_t)loc;
        unsigned char g = found ? greyOf(angle) : 0;
        R[0] = R[1] = R[2] = R[3] = g;
    }
}

} // extern "C"

// ---------------------------------------------------------------- ICP depth tracker (device half)
namespace {

// interpolateBilinear_withHoles, DA/ITMPixelUtils.h:41-71 (Vector4f source)
inline V4 bilinearHoles(const V4 *src, float px, float py, int W)
{
    short ix = (short)floorf(px), iy = (short)floorf(py);
    float dx = px - (float)ix, dy = py - (float)iy;
    V4 a = src[ix + iy * W], b = src[(ix + 1) + iy * W], c = src[ix + (iy + 1) * W], d = src[(ix + 1) + (iy + 1) * W];
    V4 r;
    if (a.w < 0 || b.w < 0 || c.w < 0 || d.w < 0) { r.x = 0; r.y = 0; r.z = 0; r.w = -1.0f; return r; }
    r.x = a.x * (1.0f - dx) * (1.0f - dy) + b.x * dx * (1.0f - dy) + c.x * (1.0f - dx) * dy + d.x * dx * dy;
    r.y = a.y * (1.0f - dx) * (1.0f - dy) + b.y * dx * (1.0f - dy) + c.y * (1.0f - dx) * dy + d.y * dx * dy;
    r.z = a.z * (1.0f - dx) * (1.0f - dy) + b.z * dx * (1.0f - dy) + c.z * (1.0f - dx) * dy + d.z * dx * dy;
    r.w = a.w * (1.0f - dx) * (1.0f - dy) + b.w * dx * (1.0f - dy) + c.w * (1.0f - dx) * dy + d.w * dx * dy;
    return r;
}

// computePerPointGH_Depth_Ab, DA/ITMDepthTracker.h:9-74.  mode: 0 both (6 parameters), 1 rotation only, 2 translation only
inline bool perPointAb(int mode, float *A, float &b, int x, int y, float depth, const float *vIntr, int sW, int sH,
                       const float *sIntr, const float *approxInvPose, const float *scenePose, const V4 *points,
                       const V4 *normals, float distThresh)
{
    if (depth <= 1e-8f) return false;
    V4 p;
    p.x = depth * (((float)x - vIntr[2]) / vIntr[0]);
    p.y = depth * (((float)y - vIntr[3]) / vIntr[1]);
    p.z = depth; p.w = 1.0f;
    p = mulM(approxInvPose, p);
    p.w = 1.0f;
    V4 q = mulM(scenePose, p);
    if (q.z <= 0.0f) return false;
    float u = sIntr[0] * q.x / q.z + sIntr[2];
    float v = sIntr[1] * q.y / q.z + sIntr[3];
    if (!((u >= 0.0f) && (u <= sW - 2) && (v >= 0.0f) && (v <= sH - 2))) return false;
    V4 c = bilinearHoles(points, u, v, sW);
    if (c.w < 0.0f) return false;
    float dx = c.x - p.x, dy = c.y - p.y, dz = c.z - p.z;
    float dist = dx * dx + dy * dy + dz * dz;
    if (dist > distThresh) return false;
    V4 n = bilinearHoles(normals, u, v, sW);
    b = n.x * dx + n.y * dy + n.z * dz;
    if (mode == 2) { A[0] = n.x; A[1] = n.y; A[2] = n.z; }
    else {
        A[0] = +p.z * n.y - p.y * n.z;
        A[1] = -p.z * n.x + p.x * n.z;
        A[2] = +p.y * n.x - p.x * n.y;
        if (mode == 0) { A[3] = n.x; A[4] = n.y; A[5] = n.z; }
    }
    return true;
}

template <typename Acc>
int trackerSums(int mode, const float *depth, int vW, int vH, const float *vIntr, const V4 *points, const V4 *normals, int sW,
                int sH, const float *sIntr, const float *scenePose, const float *approxInvPose, float distThresh, Acc *sumH,
                Acc *sumN, Acc &sumF)
{
    const int noPara = mode == 0 ? 6 : 3, noParaSQ = mode == 0 ? 21 : 6;
    int valid = 0; sumF = 0;
    for (int i = 0; i < noParaSQ; i++) sumH[i] = 0;
    for (int i = 0; i < noPara; i++) sumN[i] = 0;
    // serial, row-major pixel order: DeviceSpecific/CPU/ITMDepthTracker_CPU.cpp:35-69
    for (int y = 0; y < vH; y++) for (int x = 0; x < vW; x++) {
        float A[6], b;
        if (!perPointAb(mode, A, b, x, y, depth[x + y * vW], vIntr, sW, sH, sIntr, approxInvPose, scenePose, points, normals, distThresh))
            continue;
        // computePerPointGH_Depth, DA/ITMDepthTracker.h:76-104
        valid++; sumF += (Acc)(b * b);
        for (int r = 0, counter = 0; r < noPara; r++) {
            sumN[r] += (Acc)(b * A[r]);
            for (int c = 0; c <= r; c++, counter++) sumH[counter] += (Acc)(A[r] * A[c]);
        }
    }
    return valid;
}

} // namespace

extern "C" {

// ITMLowLevelEngine_CPU::FilterSubsampleWithHoles(ITMFloatImage), CPU/ITMLowLevelEngine_CPU.cpp:50-62 with
// filterSubsampleWithHoles, DA/ITMLowLevelEngine.h:26-47
void rst_filter_subsample_with_holes(const float *in, int w, int h, float *out)
{
    const int nw = w / 2, nh = h / 2;
    for (int y = 0; y < nh; y++) for (int x = 0; x < nw; x++) {
        float acc = 0.0f, good = 0.0f, v;
        v = in[(2 * x + 0) + (2 * y + 0) * w]; if (v > 0.0f) { acc += v; good++; }
        v = in[(2 * x + 1) + (2 * y + 0) * w]; if (v > 0.0f) { acc += v; good++; }
        v = in[(2 * x + 0) + (2 * y + 1) * w]; if (v > 0.0f) { acc += v; good++; }
        v = in[(2 * x + 1) + (2 * y + 1) * w]; if (v > 0.0f) { acc += v; good++; }
        if (good > 0) acc /= good;
        out[x + y * nw] = acc;
    }
}

// ITMDepthTracker_CPU::ComputeGandH, CPU/ITMDepthTracker_CPU.cpp:13-79.  double_sums = 0 follows the reference's float
// accumulators (bit-exact with it); double_sums = 1 accumulates the same float terms in double (the libdsb200 contract).
// sums (optional, 32 doubles): hessian lower triangle by rows, nabla, sum b^2, valid count.
int rst_tracker_compute_g_and_h(const float *depth, int vW, int vH, const float *vIntr, const float *points,
                                const float *normals, int sW, int sH, const float *sIntr, const float *scenePose,
                                const float *approxInvPose, float distThresh, int iterationType, int doubleSums, float *f,
                                float *nabla, float *hessian, double *sums)
{
    if (iterationType == DSB200_TRACKER_ITERATION_NONE) return 0;
    const int mode = iterationType == DSB200_TRACKER_ITERATION_ROTATION ? 1 : iterationType == DSB200_TRACKER_ITERATION_TRANSLATION ? 2 : 0;
    const int noPara = mode == 0 ? 6 : 3, noParaSQ = mode == 0 ? 21 : 6;
    const V4 *P = reinterpret_cast<const V4 *>(points), *N = reinterpret_cast<const V4 *>(normals);
    float sumH[21], sumN[6], sumF; int valid;
    double dH[21] = {0}, dN[6] = {0}, dF = 0;
    if (doubleSums) {
        valid = trackerSums<double>(mode, depth, vW, vH, vIntr, P, N, sW, sH, sIntr, scenePose, approxInvPose, distThresh, dH, dN, dF);
        for (int i = 0; i < noParaSQ; i++) sumH[i] = (float)dH[i];
        for (int i = 0; i < noPara; i++) sumN[i] = (float)dN[i];
        sumF = (float)dF;
    } else {
        valid = trackerSums<float>(mode, depth, vW, vH, vIntr, P, N, sW, sH, sIntr, scenePose, approxInvPose, distThresh, sumH, sumN, sumF);
        for (int i = 0; i < noParaSQ; i++) dH[i] = sumH[i];
        for (int i = 0; i < noPara; i++) dN[i] = sumN[i];
        dF = sumF;
    }
    for (int r = 0, counter = 0; r < noPara; r++) for (int c = 0; c <= r; c++, counter++) hessian[r + c * 6] = sumH[counter];
    for (int r = 0; r < noPara; ++r) for (int c = r + 1; c < noPara; c++) hessian[r + c * 6] = hessian[c + r * 6];
    memcpy(nabla, sumN, noPara * sizeof(float));
    *f = (valid > 100) ? sqrtf(sumF) / valid : 1e5f;
    if (sums) {
        for (int i = 0; i < 32; i++) sums[i] = 0;
        for (int i = 0; i < noParaSQ; i++) sums[i] = dH[i];
        for (int i = 0; i < noPara; i++) sums[noParaSQ + i] = dN[i];
        sums[noParaSQ + noPara] = dF; sums[noParaSQ + noPara + 1] = valid;
    }
    return valid;
}

} // extern "C"

// ---------------------------------------------------------------- ICP depth tracker (host loop)
namespace {

struct P3 { float x, y, z; };
inline float dotP(P3 a, P3 b) { float r = 0; r += a.x * b.x; r += a.y * b.y; r += a.z * b.z; return r; }   // ORUtils/Vector.h:811-816
inline P3 crossP(P3 a, P3 b) { P3 r = {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; return r; }

// ITMPose::SetModelViewFromParams, Objects/ITMPose.cpp:82-180
void poseToMatrix(P3 t, P3 w, float *M)
{
    float one_6th = 1.0f / 6.0f, one_20th = 1.0f / 20.0f;
    float theta_sq = dotP(w, w), theta = sqrtf(theta_sq), A, B;
    P3 T, cv = crossP(w, t);
    if (theta_sq < 1e-8f) {
        A = 1.0f - one_6th * theta_sq; B = 0.5f;
        T.x = t.x + 0.5f * cv.x; T.y = t.y + 0.5f * cv.y; T.z = t.z + 0.5f * cv.z;
    } else {
        float Cc;
        if (theta_sq < 1e-6f) {
            Cc = one_6th * (1.0f - one_20th * theta_sq); A = 1.0f - theta_sq * Cc; B = 0.5f - 0.25f * one_6th * theta_sq;
        } else {
            float inv_theta = 1.0f / theta;
            A = sinf(theta) * inv_theta; B = (1.0f - cosf(theta)) * (inv_theta * inv_theta); Cc = (1.0f - A) * (inv_theta * inv_theta);
        }
        P3 c2 = crossP(w, cv);
        T.x = t.x + B * cv.x + Cc * c2.x; T.y = t.y + B * cv.y + Cc * c2.y; T.z = t.z + B * cv.z + Cc * c2.z;
    }
    float wx2 = w.x * w.x, wy2 = w.y * w.y, wz2 = w.z * w.z, a, b;
    M[0] = 1.0f - B * (wy2 + wz2); M[5] = 1.0f - B * (wx2 + wz2); M[10] = 1.0f - B * (wx2 + wy2);
    a = A * w.z; b = B * (w.x * w.y); M[4] = b - a; M[1] = b + a;
    a = A * w.y; b = B * (w.x * w.z); M[8] = b + a; M[2] = b - a;
    a = A * w.x; b = B * (w.y * w.z); M[9] = b - a; M[6] = b + a;
    M[12] = T.x; M[13] = T.y; M[14] = T.z;
    M[3] = 0.0f; M[7] = 0.0f; M[11] = 0.0f; M[15] = 1.0f;
}

// ITMPose::SetParamsFromModelView, Objects/ITMPose.cpp:183-262
void matrixToPose(const float *M, P3 &tOut, P3 &wOut)
{
#define RM(r, c) M[(r) + 4 * (c)]
    P3 T = {M[12], M[13], M[14]}, rot;
    float cos_angle = (RM(0, 0) + RM(1, 1) + RM(2, 2) - 1.0f) * 0.5f;
    rot.x = (RM(2, 1) - RM(1, 2)) * 0.5f; rot.y = (RM(0, 2) - RM(2, 0)) * 0.5f; rot.z = (RM(1, 0) - RM(0, 1)) * 0.5f;
    float sin_angle_abs = sqrtf(dotP(rot, rot));
    if (cos_angle > M_SQRT1_2) {
        if (sin_angle_abs) { float p = asinf(sin_angle_abs) / sin_angle_abs; rot.x *= p; rot.y *= p; rot.z *= p; }
    } else if (cos_angle > -M_SQRT1_2) {
        float p = acosf(cos_angle) / sin_angle_abs; rot.x *= p; rot.y *= p; rot.z *= p;
    } else {
        float angle = (float)M_PI - asinf(sin_angle_abs);
        float d0 = RM(0, 0) - cos_angle, d1 = RM(1, 1) - cos_angle, d2 = RM(2, 2) - cos_angle;
        P3 r2;
        if (fabsf(d0) > fabsf(d1) && fabsf(d0) > fabsf(d2)) { r2.x = d0; r2.y = (RM(1, 0) + RM(0, 1)) * 0.5f; r2.z = (RM(0, 2) + RM(2, 0)) * 0.5f; }
        else if (fabsf(d1) > fabsf(d2)) { r2.x = (RM(1, 0) + RM(0, 1)) * 0.5f; r2.y = d1; r2.z = (RM(2, 1) + RM(1, 2)) * 0.5f; }
        else { r2.x = (RM(0, 2) + RM(2, 0)) * 0.5f; r2.y = (RM(2, 1) + RM(1, 2)) * 0.5f; r2.z = d2; }
        if (dotP(r2, rot) < 0.0f) { r2.x *= -1.0f; r2.y *= -1.0f; r2.z *= -1.0f; }
        float len = sqrtf(dotP(r2, r2));
        if (len == 0) { r2.x = r2.y = r2.z = 0; } else { r2.x /= len; r2.y /= len; r2.z /= len; }
        rot.x = angle * r2.x; rot.y = angle * r2.y; rot.z = angle * r2.z;
    }
#undef RM
    float shtot = 0.5f, theta = sqrtf(dotP(rot, rot));
    if (theta > 0.00001f) shtot = sinf(theta * 0.5f) / theta;
    float H[16];
    P3 zero = {0.0f, 0.0f, 0.0f}, hw = {rot.x * -0.5f, rot.y * -0.5f, rot.z * -0.5f};
    poseToMatrix(zero, hw, H);
    P3 rt;   // Matrix3 * Vector3, ORUtils/Matrix.h:310-316
    rt.x = H[0] * T.x + H[4] * T.y + H[8] * T.z;
    rt.y = H[1] * T.x + H[5] * T.y + H[9] * T.z;
    rt.z = H[2] * T.x + H[6] * T.y + H[10] * T.z;
    if (theta > 0.001f) {
        float denom = dotP(rot, rot);
        float param = dotP(T, rot) * (1 - 2 * shtot) / denom;
        rt.x -= rot.x * param; rt.y -= rot.y * param; rt.z -= rot.z * param;
    } else {
        float param = dotP(T, rot) / 24;
        rt.x -= rot.x * param; rt.y -= rot.y * param; rt.z -= rot.z * param;
    }
    rt.x /= 2 * shtot; rt.y /= 2 * shtot; rt.z /= 2 * shtot;
    tOut = rt; wOut = rot;
}

// ORUtils::Cholesky<float>, ORUtils/Cholesky.h:11-72
struct Chol {
    std::vector<float> a; int n;
    Chol(const float *mat, int size) : a(mat, mat + size * size), n(size)
    {
        for (int c = 0; c < n; c++) {
            float inv_diag = 1;
            for (int r = c; r < n; r++) {
                float val = a[c + r * n];
                for (int c2 = 0; c2 < c; c2++) val -= a[c + c2 * n] * a[c2 + r * n];
                if (r == c) { a[c + r * n] = val; inv_diag = 1.0f / val; }
                else { a[r + c * n] = val; a[c + r * n] = val * inv_diag; }
            }
        }
    }
    void backsub(float *result, const float *v) const
    {
        std::vector<float> y(n);
        for (int i = 0; i < n; i++) { float val = v[i]; for (int j = 0; j < i; j++) val -= a[j + i * n] * y[j]; y[i] = val; }
        for (int i = 0; i < n; i++) y[i] /= a[i + i * n];
        for (int i = n - 1; i >= 0; i--) { float val = y[i]; for (int j = i + 1; j < n; j++) val -= a[i + j * n] * result[j]; result[i] = val; }
    }
};

} // namespace

extern "C" {

// ITMPose::Coerce (Objects/ITMPose.cpp:345-349) on a raw matrix; params = (tx, ty, tz, rx, ry, rz)
void rst_pose_coerce(float *M, float *params)
{
    P3 t, w; matrixToPose(M, t, w); poseToMatrix(t, w, M);
    if (params) { params[0] = t.x; params[1] = t.y; params[2] = t.z; params[3] = w.x; params[4] = w.y; params[5] = w.z; }
}
void rst_pose_from_params(const float *p, float *M)
{
    P3 t = {p[0], p[1], p[2]}, w = {p[3], p[4], p[5]}; poseToMatrix(t, w, M);
}

// ITMDepthTracker::TrackCamera, Engine/ITMDepthTracker.cpp:153-211 with the ctor's schedules (:12-40).
// double_sums selects the accumulation of the G&H sums (0 = the reference CPU engine's float accumulators).
// hessian_good / nabla_good start zeroed here (uninitialised in the reference) and, like libdsb200, a level whose first
// evaluation fails before any good evaluation exists is left without a step.
void rst_track_camera(const dsb200_tracker_config *cfg, const dsb200_view *view, dsb200_tracking *trk, const float *scenePose,
                      int sceneW, int sceneH, int doubleSums, dsb200_tracker_report *report)
{
    const int L = cfg->levels;
    std::vector<std::vector<float> > pyr(L);
    std::vector<const float *> depth(L); std::vector<int> w(L), h(L); float intr[DSB200_TRACKER_MAX_LEVELS][4];
    depth[0] = view->depth; w[0] = view->depth_w; h[0] = view->depth_h;
    for (int k = 0; k < 4; k++) intr[0][k] = view->intr_d[k];
    for (int i = 1; i < L; i++) {   // PrepareForEvaluation :66-76
        w[i] = w[i - 1] / 2; h[i] = h[i - 1] / 2; pyr[i].resize((size_t)w[i] * h[i]);
        rst_filter_subsample_with_holes(depth[i - 1], w[i - 1], h[i - 1], pyr[i].data()); depth[i] = pyr[i].data();
        for (int k = 0; k < 4; k++) intr[i][k] = intr[i - 1][k] * 0.5f;
    }
    float distThresh[DSB200_TRACKER_MAX_LEVELS]; int noIter[DSB200_TRACKER_MAX_LEVELS];
    noIter[0] = 2; for (int i = 1; i < L; i++) noIter[i] = noIter[i - 1] + 2;
    float distThreshStep = cfg->dist_thresh / L;
    distThresh[L - 1] = cfg->dist_thresh;
    for (int i = L - 2; i >= 0; i--) distThresh[i] = distThresh[i + 1] - distThreshStep;

    float M[16]; memcpy(M, trk->M_d, sizeof(M));
    float f_old = 1e10, f_new = 0;
    float hessian_good[36] = {0}, hessian_new[36] = {0}, A[36], nabla_good[6] = {0}, nabla_new[6] = {0}, step[6];
    bool haveGood = false;
    if (report) memset(report, 0, sizeof(*report));
    (void)f_old;
    for (int levelId = L - 1; levelId >= cfg->no_icp_run_till_level; levelId--) {
        int iterationType = cfg->regime[levelId];
        if (iterationType == DSB200_TRACKER_ITERATION_NONE) continue;
        float approxInvPose[16], lastGood[16];
        rst_mat4_inverse(M, approxInvPose); memcpy(lastGood, M, sizeof(M));
        f_old = 1e20f; float lambda = 1.0;
        for (int iterNo = 0; iterNo < noIter[levelId]; iterNo++) {
            int n = rst_tracker_compute_g_and_h(depth[levelId], w[levelId], h[levelId], intr[levelId], trk->icp_points,
                                                trk->icp_normals, sceneW, sceneH, view->intr_d, scenePose, approxInvPose,
                                                distThresh[levelId], iterationType, doubleSums, &f_new, nabla_new, hessian_new, nullptr);
            if (report) { report->evaluations++; report->last_valid_points = n; report->last_f = f_new; }
            if (n <= 0 || f_new > f_old) {
                memcpy(M, lastGood, sizeof(M)); rst_mat4_inverse(M, approxInvPose); lambda *= 10.0f;
            } else {
                memcpy(lastGood, M, sizeof(M)); f_old = f_new;
                for (int i = 0; i < 36; ++i) hessian_good[i] = hessian_new[i] / n;
                for (int i = 0; i < 6; ++i) nabla_good[i] = nabla_new[i] / n;
                lambda /= 10.0f; haveGood = true;
            }
            if (!haveGood) break;
            for (int i = 0; i < 36; ++i) A[i] = hessian_good[i];
            for (int i = 0; i < 6; ++i) A[i + i * 6] *= 1.0f + lambda;
            // ComputeDelta :86-109
            for (int i = 0; i < 6; i++) step[i] = 0;
            if (iterationType != DSB200_TRACKER_ITERATION_BOTH) {
                float sm[9]; for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) sm[r + c * 3] = A[r + c * 6];
                Chol ch(sm, 3); ch.backsub(step, nabla_good);
            } else { Chol ch(A, 6); ch.backsub(step, nabla_good); }
            // ApplyDelta :122-151
            float s[6] = {0, 0, 0, 0, 0, 0};
            if (iterationType == DSB200_TRACKER_ITERATION_ROTATION) { s[0] = step[0]; s[1] = step[1]; s[2] = step[2]; }
            else if (iterationType == DSB200_TRACKER_ITERATION_TRANSLATION) { s[3] = step[0]; s[4] = step[1]; s[5] = step[2]; }
            else for (int i = 0; i < 6; i++) s[i] = step[i];
            float Tinc[16];
            Tinc[0] = 1.0f; Tinc[4] = s[2]; Tinc[8] = -s[1]; Tinc[12] = s[3];
            Tinc[1] = -s[2]; Tinc[5] = 1.0f; Tinc[9] = s[0]; Tinc[13] = s[4];
            Tinc[2] = s[1]; Tinc[6] = -s[0]; Tinc[10] = 1.0f; Tinc[14] = s[5];
            Tinc[3] = 0.0f; Tinc[7] = 0.0f; Tinc[11] = 0.0f; Tinc[15] = 1.0f;
            float next[16]; rst_mat4_multiply(Tinc, approxInvPose, next);
            rst_mat4_inverse(next, M);          // pose_d->SetInvM
            rst_pose_coerce(M, nullptr);        // pose_d->Coerce
            rst_mat4_inverse(M, approxInvPose); // pose_d->GetInvM
            if (report) report->steps++;
            float stepLength = 0.0f;            // HasConverged :111-120
            for (int i = 0; i < 6; i++) stepLength += step[i] * step[i];
            if (sqrtf(stepLength) / 6 < cfg->termination_threshold) break;
        }
    }
    memcpy(trk->M_d, M, sizeof(M)); rst_mat4_inverse(M, trk->invM_d);
}

} // extern "C"

// ---------------------------------------------------------------- view builder: bilateral depth filter
extern "C" {

// ITMViewBuilder_CPU::ConvertDisparityToDepth, CPU/ITMViewBuilder_CPU.cpp:94-106 with convertDisparityToDepth, DA/ITMViewBuilder.h:7-20
void rst_convert_disparity_to_depth(const short *d_in, float *d_out, int W, int H, float paramX, float paramY, float fx_depth)
{
    for (int i = 0; i < W * H; i++) {
        short disparity = d_in[i];
        float disparity_tmp = paramX - (float)(disparity);
        float depth;
        if (disparity_tmp == 0) depth = 0.0;
        else depth = 8.0f * paramY * fx_depth / disparity_tmp;
        d_out[i] = (depth > 0) ? depth : -1.0f;
    }
}

// ITMViewBuilder_CPU::UpdateView's useBilateralFilter branch, CPU/ITMViewBuilder_CPU.cpp:49-58 with DepthFiltering :119-130
// and filterDepth, DA/ITMViewBuilder.h:30-53.  exp on a float argument is the float overload (glibc expf), as in the reference.
void rst_view_filter_depth(float *depth, int W, int H)
{
    std::vector<float> tmp((size_t)W * H);
    float *bufs[2] = {depth, tmp.data()};
    for (int pass = 0; pass < 5; pass++) {
        const float *in = bufs[pass & 1]; float *out = bufs[(pass & 1) ^ 1];
        memset(out, 0, sizeof(float) * (size_t)W * H);
        for (int y = 2; y < H - 2; y++) for (int x = 2; x < W - 2; x++) {
            float z, tmpz, dz, final_depth = 0.0f, w, w_sum = 0.0f;
            z = in[x + y * W];
            if (z < 0.0f) { out[x + y * W] = -1.0f; continue; }
            float sigma_z = 1.0f / (0.0012f + 0.0019f * (z - 0.4f) * (z - 0.4f) + 0.0001f / sqrtf(z) * 0.25f);
            for (int i = -2; i <= 2; i++) for (int j = -2; j <= 2; j++) {
                tmpz = in[(x + j) + (y + i) * W];
                if (tmpz < 0.0f) continue;
                dz = (tmpz - z); dz *= dz;
                w = expf(-0.5f * ((abs(i) + abs(j)) * 1.2232f * 1.2232f + dz * sigma_z * sigma_z));
                w_sum += w;
                final_depth += w * tmpz;
            }
            final_depth /= w_sum;
            out[x + y * W] = final_depth;
        }
    }
    memcpy(depth, tmp.data(), sizeof(float) * (size_t)W * H);   // five passes end in the staging image; SetFrom copies it back
}

// ITMViewBuilder_CPU::ComputeNormalAndWeights, CPU/ITMViewBuilder_CPU.cpp:132-143 with computeNormalAndWeight,
// DA/ITMViewBuilder.h:55-112.  sqrt and acos on float arguments are the float overloads (glibc sqrtf / acosf), PI is
// float(3.14159...) (ORUtils/MathUtils.h:28).  Only the interior is written; a rejected pixel gets normal.w = -1 and sigmaZ = -1.
void rst_compute_normal_and_weights(const float *depth_in, float *normal_out, float *sigmaZ_out, int W, int H, const float *intrinparam)
{
    const float ix = intrinparam[0], iy = intrinparam[1], iz = intrinparam[2], iw = intrinparam[3];
    const float kPI = float(3.1415926535897932384626433832795);
    for (int y = 2; y < H - 2; y++) for (int x = 2; x < W - 2; x++) {
        const int idx = x + y * W;
        float *n4 = normal_out + 4 * (size_t)idx;
        float z = depth_in[idx];
        if (z < 0.0f) { n4[3] = -1.0f; sigmaZ_out[idx] = -1; continue; }
        float xp[3], xm[3], yp[3], ym[3];
        xp[2] = depth_in[(x + 1) + y * W]; yp[2] = depth_in[x + (y + 1) * W];
        xm[2] = depth_in[(x - 1) + y * W]; ym[2] = depth_in[x + (y - 1) * W];
        if (xp[2] <= 0 || yp[2] <= 0 || xm[2] <= 0 || ym[2] <= 0) { n4[3] = -1.0f; sigmaZ_out[idx] = -1; continue; }
        xp[0] = xp[2] * ((x + 1.0f) - iz) * ix; xp[1] = xp[2] * (y - iw) * iy;
        xm[0] = xm[2] * ((x - 1.0f) - iz) * ix; xm[1] = xm[2] * (y - iw) * iy;
        yp[0] = yp[2] * (x - iz) * ix; yp[1] = yp[2] * ((y + 1.0f) - iw) * iy;
        ym[0] = ym[2] * (x - iz) * ix; ym[1] = ym[2] * ((y - 1.0f) - iw) * iy;
        float dx[3], dy[3], n[3];
        for (int k = 0; k < 3; k++) { dx[k] = xp[k] - xm[k]; dy[k] = yp[k] - ym[k]; }
        n[0] = (dx[1] * dy[2] - dx[2] * dy[1]);
        n[1] = (dx[2] * dy[0] - dx[0] * dy[2]);
        n[2] = (dx[0] * dy[1] - dx[1] * dy[0]);
        if (n[0] == 0.0f && n[1] == 0 && n[2] == 0) { n4[3] = -1.0f; sigmaZ_out[idx] = -1; continue; }
        float norm = 1.0f / sqrtf(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
        n[0] *= norm; n[1] *= norm; n[2] *= norm;
        n4[0] = n[0]; n4[1] = n[1]; n4[2] = n[2]; n4[3] = 1.0f;
        float theta = acosf(n[2]);
        float theta_diff = theta / (kPI * 0.5f - theta);
        sigmaZ_out[idx] = (0.0012f + 0.0019f * (z - 0.4f) * (z - 0.4f) + 0.0001f / sqrtf(z) * theta_diff * theta_diff);
    }
}

// ---------------------------------------------------------------- swapping engine
// swap-state pass of AllocateSceneFromDepth with useSwapping, CPU/ITMSceneReconstructionEngine_CPU.cpp:287-299
void rst_swap_mark_visible(dsb200_swap_cache *cache, const dsb200_render_state *rs)
{
    for (int targetIdx = 0; targetIdx < cache->total_entries; targetIdx++)
        if (rs->visible_type[targetIdx] > 0 && cache->states[targetIdx] != 2) cache->states[targetIdx] = 1;
}

namespace {
// combineVoxelDepthInformation + combineVoxelColorInformation, DA/ITMSwappingEngine.h:7-45 (ITMVoxel_s_rgb has colour)
inline void combineVoxel(const dsb200_voxel &src, dsb200_voxel &dst, int maxW)
{
    {
        int newW = dst.w_depth, oldW = src.w_depth;
        float newF = sdfToFloat(dst.sdf), oldF = sdfToFloat(src.sdf);
        if (oldW != 0) {
            newF = oldW * oldF + newW * newF;
            newW = oldW + newW;
            newF /= newW;
            newW = (newW < maxW) ? newW : maxW;
            dst.w_depth = (uint8_t)newW;
            dst.sdf = sdfToShort(newF);
        }
    }
    {
        int newW = dst.w_color, oldW = src.w_color;
        float newC[3], oldC[3];
        for (int k = 0; k < 3; k++) { newC[k] = (float)dst.clr[k] / 255.0f; oldC[k] = (float)src.clr[k] / 255.0f; }
        if (oldW != 0) {
            for (int k = 0; k < 3; k++) newC[k] = oldC[k] * (float)oldW + newC[k] * (float)newW;
            newW = oldW + newW;
            for (int k = 0; k < 3; k++) newC[k] /= (float)newW;
            newW = (newW < maxW) ? newW : maxW;
            for (int k = 0; k < 3; k++) dst.clr[k] = toUChar(newC[k] * 255.0f);
            dst.w_color = (uint8_t)newW;
        }
    }
}
}

// LoadFromGlobalMemory + IntegrateGlobalIntoLocal, CPU/ITMSwappingEngine_CPU.cpp:22-131
int rst_swap_integrate_global_into_local(dsb200_swap_cache *cache, dsb200_scene *scene)
{
    std::vector<int> needed;
    for (int entryId = 0; entryId < cache->total_entries; entryId++) {
        if ((int)needed.size() >= cache->transfer_blocks) break;
        if (cache->states[entryId] == 1) needed.push_back(entryId);
    }
    const int maxW = scene->params.max_w;
    for (size_t i = 0; i < needed.size(); i++) {
        const int entryDestId = needed[i];
        const int ptr = scene->hash[entryDestId].ptr;
        if (cache->has_stored[entryDestId] && ptr >= 0) {      // ptr < 0: the reference would index the volume with it
            const dsb200_voxel *srcVB = cache->stored + (size_t)entryDestId * DSB200_BLOCK_SIZE3;
            dsb200_voxel *dstVB = scene->voxels + (size_t)ptr * DSB200_BLOCK_SIZE3;
            for (int vIdx = 0; vIdx < DSB200_BLOCK_SIZE3; vIdx++) combineVoxel(srcVB[vIdx], dstVB[vIdx], maxW);
        }
        cache->states[entryDestId] = 2;
    }
    return (int)needed.size();
}

// SaveToGlobalMemory(scene) / SaveToGlobalMemory(scene, renderState), CPU/ITMSwappingEngine_CPU.cpp:133-297
int rst_swap_save_to_global_memory(dsb200_swap_cache *cache, dsb200_scene *scene, const dsb200_render_state *rs)
{
    int noNeededEntries = 0;
    int noAllocatedVoxelEntries = scene->last_free_block;
    for (int entryDestId = 0; entryDestId < cache->total_entries; entryDestId++) {
        if (noNeededEntries >= cache->transfer_blocks) break;
        const int localPtr = scene->hash[entryDestId].ptr;
        if (cache->states[entryDestId] == 2 && localPtr >= 0 && (!rs || rs->visible_type[entryDestId] == 0)) {
            dsb200_voxel *loc = scene->voxels + (size_t)localPtr * DSB200_BLOCK_SIZE3;
            cache->has_stored[entryDestId] = 1;                 // SetStoredData, ITMGlobalCache.h:30-34 (via the staging copy)
            memcpy(cache->stored + (size_t)entryDestId * DSB200_BLOCK_SIZE3, loc, sizeof(dsb200_voxel) * DSB200_BLOCK_SIZE3);
            cache->states[entryDestId] = 0;
            const int vbaIdx = noAllocatedVoxelEntries;
            // the reference compares with SDF_BUCKET_NUM - 1; the second test only matters where it would write past the list
            if (vbaIdx < scene->bucket_num - 1 && vbaIdx + 1 < scene->num_blocks) {
                noAllocatedVoxelEntries++;
                scene->alloc_list[vbaIdx + 1] = localPtr;
                scene->hash[entryDestId].ptr = -1;
                for (int i = 0; i < DSB200_BLOCK_SIZE3; i++) {
                    dsb200_voxel v; memset(&v, 0, sizeof(v)); v.sdf = 32767;       // TVoxel()
                    loc[i] = v;
                }
            }
            noNeededEntries++;
        }
    }
    scene->last_free_block = noAllocatedVoxelEntries;
    return noNeededEntries;
}

} // extern "C"


// ================================================================ meshing
// ITMMeshingEngine_CPU<TVoxel,ITMVoxelBlockHash>::MeshScene, CPU/ITMMeshingEngine_CPU.cpp:20-83; buildVertList /
// findPointNeighbors / sdfInterp, DeviceAgnostic/ITMMeshingEngine.h:90-160; vertex colours
// readFromSDF_color4u_interpolated_noalpha, DeviceAgnostic/ITMRepresentationAccess.h:699-736.
#include "../denseslam-global-consistency-h_b200/csrc/mc_tables.h"
namespace {

struct F3 { float x, y, z; };

// sdfInterp :83-90
inline F3 sdfInterp(F3 p1, F3 p2, float valp1, float valp2)
{
    if (fabsf(0.0f - valp1) < 0.00001f) return p1;
    if (fabsf(0.0f - valp2) < 0.00001f) return p2;
    if (fabsf(valp1 - valp2) < 0.00001f) return p1;
    const float t = (0.0f - valp1) / (valp2 - valp1);
    F3 r = {p1.x + t * (p2.x - p1.x), p1.y + t * (p2.y - p1.y), p1.z + t * (p2.z - p1.z)};
    return r;
}

// cube corners in the order of findPointNeighbors :40-80
const int kCorner[8][3] = {{0, 0, 0}, {1, 0, 0}, {1, 1, 0}, {0, 1, 0}, {0, 0, 1}, {1, 0, 1}, {1, 1, 1}, {0, 1, 1}};
// cube edges (corner pairs) in the order of buildVertList :104-115
const int kEdge[12][2] = {{0, 1}, {1, 2}, {2, 3}, {3, 0}, {4, 5}, {5, 6}, {6, 7}, {7, 4}, {0, 4}, {1, 5}, {2, 6}, {3, 7}};

inline int buildVertList(F3 *vertList, const dsb200_scene *s, int gx, int gy, int gz)
{
    F3 pts[8]; float sdf[8];
    for (int i = 0; i < 8; i++) {
        const int x = gx + kCorner[i][0], y = gy + kCorner[i][1], z = gz + kCorner[i][2];
        bool found;
        pts[i].x = (float)x; pts[i].y = (float)y; pts[i].z = (float)z;
        sdf[i] = sdfToFloat(readVoxelNC(s, x, y, z, found).sdf);
        if (!found || sdf[i] == 1.0f) return -1;
    }
    int cube = 0;
    for (int i = 0; i < 8; i++) if (sdf[i] < 0) cube |= 1 << i;
    const unsigned edges = dsb::mc_edge_mask(dsb::kMcTrianglesHost[cube]);
    if (edges == 0) return -1;
    for (int e = 0; e < 12; e++)
        if (edges & (1u << e)) vertList[e] = sdfInterp(pts[kEdge[e][0]], pts[kEdge[e][1]], sdf[kEdge[e][0]], sdf[kEdge[e][1]]);
    return cube;
}

} // namespace

extern "C" {

void rst_mc_tables(int *edgeTable, int *triangleTable)
{
    for (int c = 0; c < 256; c++) {
        const uint64_t w = dsb::kMcTrianglesHost[c];
        edgeTable[c] = (int)dsb::mc_edge_mask(w);
        for (int i = 0; i < 16; i++) { const int e = (int)((w >> (4 * i)) & 0xF); triangleTable[c * 16 + i] = e == 0xF ? -1 : e; }
    }
}

unsigned rst_mesh_scene(const dsb200_scene *scene, dsb200_triangle *triangles, unsigned maxTriangles)
{
    const int total = scene->bucket_num + scene->excess_size;
    const float factor = scene->params.voxel_size;
    memset(triangles, 0, sizeof(dsb200_triangle) * (size_t)maxTriangles);      // mesh->triangles->Clear()
    unsigned n = 0;
    for (int entry = 0; entry < total; entry++) {
        const dsb200_hash_entry &e = scene->hash[entry];
        if (e.ptr < 0) continue;
        const int gx = e.pos[0] * DSB200_BLOCK_SIZE, gy = e.pos[1] * DSB200_BLOCK_SIZE, gz = e.pos[2] * DSB200_BLOCK_SIZE;
        for (int z = 0; z < DSB200_BLOCK_SIZE; z++) for (int y = 0; y < DSB200_BLOCK_SIZE; y++) for (int x = 0; x < DSB200_BLOCK_SIZE; x++) {
            F3 v[12];
            const int cube = buildVertList(v, scene, gx + x, gy + y, gz + z);
            if (cube < 0) continue;
            const uint64_t w = dsb::kMcTrianglesHost[cube];
            for (int i = 0; i < 15 && ((w >> (4 * i)) & 0xF) != 0xF; i += 3) {
                dsb200_triangle &t = triangles[n];
                for (int k = 0; k < 3; k++) {
                    const F3 p = v[(w >> (4 * (i + k))) & 0xF];
                    float *dst = k == 0 ? t.p0 : (k == 1 ? t.p1 : t.p2), *col = k == 0 ? t.c0 : (k == 1 ? t.c1 : t.c2);
                    dst[0] = p.x * factor; dst[1] = p.y * factor; dst[2] = p.z * factor;
                    const float pv[3] = {p.x, p.y, p.z};
                    float c4[4];
                    colourTrilinear(scene, pv, c4);
                    col[0] = c4[0]; col[1] = c4[1]; col[2] = c4[2];
                }
                if (n < maxTriangles - 1) n++;
            }
        }
    }
    return n;
}

} // extern "C"
