// capi.cu -- the C ABI of include/dsb200.h: engine objects, history queues, host logic.
//
// Host-side control flow follows the reference's CUDA engines
// (CUDA/ITMSceneReconstructionEngine_CUDA.cu:152-359, 363-489, 850-1082 and
// CUDA/ITMVisualisationEngine_CUDA.cu:150-251, 351-484) with these structural changes:
//  * counts (visible blocks, free-list tops) live on the device between stages; the host reads
//    them once per call (or once per fused frame), not after every kernel;
//  * the per-frame visible-list snapshots used by decay / slide-window / de-fusion live in a
//    recycled device arena (no cudaMalloc per frame, no 4x over-allocation, no dangling aliases:
//    SURVEY.md Appendix B Q2);
//  * errors are returned as codes (+ dsb200_last_error) instead of exit(-1)/throw.
#include "kernels.cuh"

#include <cmath>
#include <cstdio>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace dsb {
unsigned long long g_launches = 0;
}
using namespace dsb;

static thread_local std::string g_error;

static int fail(int code, const std::string &msg) { g_error = msg; return code; }
namespace dsb { int fail_with(int code, const char *msg) { return fail(code, msg); } }   // for tracker_host.cpp
static int cuda_fail(cudaError_t e, const char *what)
{
    return fail(e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? DSB200_ERR_NO_DEVICE : DSB200_ERR_CUDA,
                std::string(what) + ": " + cudaGetErrorString(e));
}
#define CU(call)                                                     \
    do {                                                             \
        cudaError_t e_ = (call);                                     \
        if (e_ != cudaSuccess) return cuda_fail(e_, #call);          \
    } while (0)

// ---------------------------------------------------------------- snapshot arena
namespace {

struct Chunk {
    int *base = nullptr;
    size_t capacity = 0, used = 0;   // in ints
};

struct Arena {
    std::vector<Chunk *> freeChunks;
    std::shared_ptr<Chunk> current;
    size_t chunkInts = (size_t)8 << 20;   // 32 MiB

    std::shared_ptr<Chunk> acquire(size_t minInts)
    {
        Chunk *c = nullptr;
        for (size_t i = 0; i < freeChunks.size(); i++)
            if (freeChunks[i]->capacity >= minInts) { c = freeChunks[i]; freeChunks.erase(freeChunks.begin() + i); break; }
        if (!c) {
            c = new Chunk();
            c->capacity = minInts > chunkInts ? minInts : chunkInts;
            if (cudaMalloc(&c->base, c->capacity * sizeof(int)) != cudaSuccess) { delete c; return nullptr; }
        }
        c->used = 0;
        Arena *self = this;
        return std::shared_ptr<Chunk>(c, [self](Chunk *p) { self->freeChunks.push_back(p); });
    }
    // reserve `ints` ints (64-int aligned); returns the owning chunk
    std::shared_ptr<Chunk> reserve(size_t ints, int **out)
    {
        ints = (ints + 63) & ~(size_t)63;
        if (!current || current->used + ints > current->capacity) current = acquire(ints);
        if (!current) return nullptr;
        *out = current->base + current->used;
        current->used += ints;
        return current;
    }
    // called where the caller blocks anyway (counter read-back): have the next chunk ready before the current one runs out, so
    // that no cudaMalloc (implicitly synchronising) lands between two asynchronous frames
    void ensure_spare()
    {
        if (!freeChunks.empty() || !current || current->used * 2 < current->capacity) return;
        Chunk *c = new Chunk();
        c->capacity = chunkInts;
        if (cudaMalloc(&c->base, c->capacity * sizeof(int)) != cudaSuccess) { cudaGetLastError(); delete c; return; }
        freeChunks.push_back(c);
    }
    void release_all()
    {
        current.reset();
        for (Chunk *c : freeChunks) { cudaFree(c->base); delete c; }
        freeChunks.clear();
    }
};

struct Snapshot {
    int *region = nullptr;     // arena reservation: [0] = device count, ids start at +64
    size_t regionInts = 0;
    int *ids = nullptr;        // device, in the arena
    int *countDev = nullptr;   // device count (valid as soon as k_visible_list has run)
    int countHost = -1;        // -1 until the host has synchronised
    size_t frameIdx = 0;
    std::shared_ptr<Chunk> chunk;
};
typedef std::shared_ptr<Snapshot> SnapPtr;

void fill_mat(Mat4 &dst, const float *src) { memcpy(dst.m, src, 64); }

} // namespace

struct dsb200_recon {
    cudaStream_t stream = nullptr;
    int numBlocks = 0, bucketNum = 0, excessSize = 0, total = 0;
    unsigned *reqKey = nullptr;       // order keys of allocation requests (entriesAllocType + blockCoords equivalent)
    unsigned *slotMark = nullptr;     // pruning: per-slot rank marks, then per-bucket leaders
    int *pruneScratch = nullptr; int pruneCap = 0;
    FrameState fs{};
    Counters *hostCounters = nullptr; // pinned
    dsb200_weight_params weights{0, 10, 100};
    bool async = false, deviceCountersValid = false;
    int lastKnownVisible = 0;         // visible count at the last host sync: sizes the snapshot reservations made while asynchronous
    const uint8_t *visBitsOwner = nullptr;   // the render state's visibility-flag array that fs.visBits currently describes
    cudaStream_t sideStream = nullptr; cudaEvent_t evFork = nullptr, evJoin = nullptr;   // fused frame: expected depths || integration
    // look-ahead (dsb200_recon_prefetch_allocation): first half of the next frame's allocation on its own stream
    cudaStream_t prefetchStream = nullptr; cudaEvent_t evSceneWritten = nullptr, evPrefetchDone = nullptr;
    unsigned long long modEpoch = 0, epochSceneWritten = ~0ull;   // scene-modifying calls so far / value when evSceneWritten was recorded
    bool prefetchPending = false;
    struct PrefetchKey { const float *depth; int w, h; float M[16], invM[16], intr[4]; const uint8_t *visType;
                         uint8_t *visTypeRW; const int *visibleIds; const dsb200_hash_entry *hash; int useSwapping;
                         bool deep; dsb200_hash_entry *hashRW; } prefetchKey{};
    // deep look-ahead (recon_alloc.cu): undo / commit list of the ordered allocation, second visible-list buffer
    int *deepBuf = nullptr; int *nextVisible = nullptr; int nextVisibleCap = 0;
    // SDF shadow: one short per voxel of the scene this engine maintains, kept in step by every kernel that writes voxels
    // (reset, integrate, de-integrate, pruning) and read by the fused frame's raycast instead of the 8-byte voxel records.
    short *sdfShadow = nullptr; const void *shadowFor = nullptr; bool shadowEnabled = true;
    unsigned char *weightImage = nullptr; size_t weightBytes = 0;   // per-pixel depth weights of the frame being fused (flat kernel)
    dsb200_shard shard{0, 1, 2, 0};      // sharded scene: this engine's block ownership (count <= 1: everything)
    bool forceSlowDiv = false;        // debug: run the plain-division kernel variants (dsb200_debug_force_plain_division)
    // history (CUDA/ITMSceneReconstructionEngine_CUDA.h:41-48)
    std::deque<SnapPtr> decayQ, slideQ, decayDefusionQ, slideDefusionQ;
    std::map<double, SnapPtr> snapshots;   // mDefusionBlockDataBase
    std::vector<SnapPtr> unsynced;         // snapshots whose host count is not known yet (async mode)
    size_t frameIdx = 0;
    long long totalDecayed = 0;
    Arena arena;
    cudaEvent_t evAlloc[DSB200_PROFILE_ALLOC_STAGES + 1] = {}; bool evAllocRecorded = false;
    bool profile = false; cudaEvent_t ev[DSB200_PROFILE_STAGES + 1] = {}; bool evRecorded[DSB200_PROFILE_STAGES + 1] = {};
};

struct dsb200_vis {
    cudaStream_t stream = nullptr;
    FrameState fs{};
    int scanTilesAlloc = 0;
    Counters *hostCounters = nullptr;
    int *countScratch = nullptr;      // [0] visible count for stand-alone calls, [1] CountVisibleBlocks result
    int *hostInts = nullptr;          // pinned [4]
    float *fwdProjection = nullptr; int *fwdMissing = nullptr; int fwdPixels = 0;
    RayQueue rayQueue{};              // suspended rays of the bounded first raycast pass (capacity = width * height of creation)
    int *tileOrder = nullptr; int tileOrderCap = 0;   // fused frame: raycast tiles in descending order of expected-depth span
};

extern "C" {

const char *dsb200_last_error(void) { return g_error.c_str(); }

int dsb200_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

unsigned long long dsb200_kernel_launch_count(void) { return g_launches; }

int dsb200_set_device(int device) { CU(cudaSetDevice(device)); return DSB200_OK; }

// ---------------------------------------------------------------- engine lifetime
int dsb200_recon_create(dsb200_recon **out, int32_t num_blocks, int32_t bucket_num, int32_t excess_size, void *stream)
{
    if (!out || num_blocks <= 0 || bucket_num <= 0 || (bucket_num & (bucket_num - 1)) || excess_size < 0)
        return fail(DSB200_ERR_INVALID, "recon_create: capacities must be positive and bucket_num a power of two");
    if (dsb200_device_count() == 0) return fail(DSB200_ERR_NO_DEVICE, "no CUDA device: libdsb200 has no CPU fallback");
    dsb200_recon *e = new dsb200_recon();
    e->stream = (cudaStream_t)stream;
    e->numBlocks = num_blocks; e->bucketNum = bucket_num; e->excessSize = excess_size; e->total = bucket_num + excess_size;
    e->fs.scanTiles = div_up(e->total, 1024);
    e->fs.bitTiles = div_up(e->total, 32768);
    CU(cudaMalloc(&e->fs.reqBits, (size_t)e->fs.bitTiles * 8192));
    CU(cudaMemset(e->fs.reqBits, 0, (size_t)e->fs.bitTiles * 8192));
    e->fs.reqExBits = e->fs.reqBits + (size_t)e->fs.bitTiles * 1024;
    CU(cudaMalloc(&e->fs.visBits, (size_t)e->fs.bitTiles * 4096));
    CU(cudaMemset(e->fs.visBits, 0, (size_t)e->fs.bitTiles * 4096));
    const int pruneTiles = div_up(num_blocks, 1024);
    size_t keyInts = (size_t)e->fs.scanTiles * 1024;
    CU(cudaMalloc(&e->reqKey, keyInts * 4));
    CU(cudaMemset(e->reqKey, 0, keyInts * 4));
    CU(cudaMalloc(&e->slotMark, (keyInts + bucket_num) * 4));
    CU(cudaMemset(e->slotMark, 0, (keyInts + bucket_num) * 4));
    CU(cudaMalloc(&e->fs.counters, sizeof(Counters)));
    CU(cudaMemset(e->fs.counters, 0, sizeof(Counters)));
    CU(cudaMalloc(&e->fs.scanState, (size_t)(2 * e->fs.scanTiles + pruneTiles + 8) * 8));
    CU(cudaMemset(e->fs.scanState, 0, (size_t)(2 * e->fs.scanTiles + pruneTiles + 8) * 8));
    CU(cudaMalloc(&e->fs.tickets, 4 * sizeof(int)));
    CU(cudaMemset(e->fs.tickets, 0, 4 * sizeof(int)));
    CU(cudaMallocHost(&e->hostCounters, sizeof(Counters)));
    memset(e->hostCounters, 0, sizeof(Counters));
    {   // the side stream's kernels are small and sit beside the integration, whose grid fills the GPU: highest priority, so that they
        // get the first SM slots that free up instead of queueing behind the whole grid
        int lo = 0, hi = 0;
        CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CU(cudaStreamCreateWithPriority(&e->sideStream, cudaStreamNonBlocking, hi));
    }
    CU(cudaEventCreateWithFlags(&e->evFork, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&e->evJoin, cudaEventDisableTiming));
    CU(cudaStreamCreateWithFlags(&e->prefetchStream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&e->evSceneWritten, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&e->evPrefetchDone, cudaEventDisableTiming));
    // 2 bytes per voxel = +25 % of the voxel array; if it does not fit the engine simply runs without it
    if (cudaMalloc(&e->sdfShadow, (size_t)num_blocks * DSB200_BLOCK_SIZE3 * sizeof(short)) != cudaSuccess) { cudaGetLastError(); e->sdfShadow = nullptr; }
    e->pruneCap = num_blocks;
    CU(cudaMalloc(&e->pruneScratch, (size_t)3 * e->pruneCap * 4));
    *out = e;
    return DSB200_OK;
}

int dsb200_recon_destroy(dsb200_recon *e)
{
    if (!e) return DSB200_OK;
    cudaStreamSynchronize(e->stream);
    if (e->sideStream) { cudaStreamSynchronize(e->sideStream); cudaStreamDestroy(e->sideStream); cudaEventDestroy(e->evFork); cudaEventDestroy(e->evJoin); }
    if (e->prefetchStream) { cudaStreamSynchronize(e->prefetchStream); cudaStreamDestroy(e->prefetchStream); cudaEventDestroy(e->evSceneWritten); cudaEventDestroy(e->evPrefetchDone); }
    e->decayQ.clear(); e->slideQ.clear(); e->decayDefusionQ.clear(); e->slideDefusionQ.clear();
    e->snapshots.clear(); e->unsynced.clear();
    e->arena.release_all();
    cudaFree(e->reqKey); cudaFree(e->slotMark); cudaFree(e->pruneScratch); cudaFree(e->fs.counters);
    cudaFree(e->fs.scanState); cudaFree(e->fs.tickets); cudaFreeHost(e->hostCounters);
    cudaFree(e->fs.reqBits); cudaFree(e->fs.visBits); cudaFree(e->sdfShadow); cudaFree(e->weightImage);
    cudaFree(e->deepBuf); cudaFree(e->nextVisible);
    delete e;
    return DSB200_OK;
}

int dsb200_recon_set_async(dsb200_recon *e, int async) { e->async = async != 0; return DSB200_OK; }

int dsb200_set_fusion_weight_params(dsb200_recon *e, const dsb200_weight_params *p) { e->weights = *p; return DSB200_OK; }
int dsb200_recon_set_shard(dsb200_recon *e, const dsb200_shard *s)
{
    if (!e || !s) return fail(DSB200_ERR_INVALID, "set_shard: null argument");
    if (s->count > 1 && (s->rank < 0 || s->rank >= s->count || s->brick_log2 < 0 || s->brick_log2 > 12 || s->count > kMaxShards))
        return fail(DSB200_ERR_INVALID, "set_shard: need 0 <= rank < count <= 16 and 0 <= brick_log2 <= 12");
    e->shard = *s;
    if (e->shard.count < 1) e->shard.count = 1;
    return DSB200_OK;
}
int dsb200_shard_owner(int bx, int by, int bz, int count, int brick_log2)
{
    return count > 1 ? shard_owner(bx, by, bz, count, brick_log2) : 0;
}
int dsb200_shard_holds(int bx, int by, int bz, const dsb200_shard *s)
{
    return (s->count <= 1 || shard_holds(bx, by, bz, s->rank, s->count, s->brick_log2, s->halo)) ? 1 : 0;
}
size_t dsb200_decayed_block_count(const dsb200_recon *e) { return (size_t)e->totalDecayed; }
int dsb200_history_sizes(const dsb200_recon *e, int *a, int *b, int *c, int *d, int *s)
{
    if (a) *a = (int)e->decayQ.size();
    if (b) *b = (int)e->slideQ.size();
    if (c) *c = (int)e->decayDefusionQ.size();
    if (d) *d = (int)e->slideDefusionQ.size();
    if (s) *s = (int)e->snapshots.size();
    return DSB200_OK;
}

} // extern "C"

// Pull the device counters to the host and refresh the caller's mirrors.
static int sync_counters(dsb200_recon *e, dsb200_scene *scene, dsb200_render_state *rs)
{
    CU(cudaMemcpyAsync(e->hostCounters, e->fs.counters, sizeof(Counters), cudaMemcpyDeviceToHost, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    if (scene) { scene->last_free_block = e->hostCounters->lastFreeBlock; scene->last_free_excess = e->hostCounters->lastFreeExcess; }
    if (rs) rs->no_visible = e->hostCounters->noVisible;
    e->lastKnownVisible = e->hostCounters->noVisible;
    if (e->hostCounters->snapshotOverflow) {
        e->hostCounters->snapshotOverflow = 0;
        CU(cudaMemcpyAsync(&e->fs.counters->snapshotOverflow, &e->hostCounters->snapshotOverflow, 4, cudaMemcpyHostToDevice, e->stream));
        return fail(DSB200_ERR_INVALID, "a visible-list snapshot outgrew its asynchronous reservation (the visible set more than doubled "
                                        "since the last dsb200_recon_sync): synchronise more often");
    }
    e->totalDecayed = e->hostCounters->totalFreed;     // accumulated on the device: exact whatever ran between two syncs
    if (!e->unsynced.empty()) {
        // host counts of snapshots taken since the last sync
        std::vector<int> counts(e->unsynced.size());
        for (size_t i = 0; i < e->unsynced.size(); i++)
            CU(cudaMemcpyAsync(&counts[i], e->unsynced[i]->countDev, 4, cudaMemcpyDeviceToHost, e->stream));
        CU(cudaStreamSynchronize(e->stream));
        for (size_t i = 0; i < e->unsynced.size(); i++) {
            Snapshot &s = *e->unsynced[i];
            s.countHost = counts[i];
            // give back the unused tail if this snapshot is still the newest reservation of its chunk
            size_t need = 64 + (((size_t)s.countHost + 63) & ~(size_t)63);
            if (s.chunk && s.chunk->base + s.chunk->used == s.region + s.regionInts && need < s.regionInts) {
                s.chunk->used -= (s.regionInts - need);
                s.regionInts = need;
            }
        }
        e->unsynced.clear();
    }
    e->arena.ensure_spare();
    e->deviceCountersValid = true;
    return DSB200_OK;
}

extern "C" int dsb200_recon_sync(dsb200_recon *e, dsb200_scene *scene, dsb200_render_state *rs)
{
    return sync_counters(e, scene, rs);
}

// the same, but only when the DEVICE counters are the authoritative ones (asynchronous engine with frames in flight); between
// synchronous calls, or after dsb200_recon_notify_scene_modified, the host mirrors already are and must not be overwritten
extern "C" int dsb200_recon_refresh_mirrors(dsb200_recon *e, dsb200_scene *scene, dsb200_render_state *rs)
{
    if (!e) return fail(DSB200_ERR_INVALID, "recon_refresh_mirrors: null engine");
    if (!(e->async && e->deviceCountersValid)) { CU(cudaStreamSynchronize(e->stream)); return DSB200_OK; }
    return sync_counters(e, scene, rs);
}

// ---------------------------------------------------------------- reconstruction engine
static void make_alloc_params(AllocParams &P, const dsb200_recon *e, const dsb200_scene *scene, const dsb200_view *view,
                              const dsb200_tracking *trk)
{
    P.shardRank = e->shard.rank; P.shardCount = e->shard.count; P.shardBrickLog2 = e->shard.brick_log2; P.shardHalo = e->shard.halo;
    fill_mat(P.invM, trk->invM_d);
    P.invProj[0] = 1.0f / view->intr_d[0]; P.invProj[1] = 1.0f / view->intr_d[1];
    P.invProj[2] = view->intr_d[2]; P.invProj[3] = view->intr_d[3];
    P.mu = scene->params.mu;
    P.oneOverBlock = 1.0f / (scene->params.voxel_size * DSB200_BLOCK_SIZE);
    P.vfMin = scene->params.vf_min; P.vfMax = scene->params.vf_max;
    P.W = view->depth_w; P.H = view->depth_h;
    P.bucketNum = scene->bucket_num; P.bucketMask = scene->bucket_num - 1;
}

static bool in_window(float v, int loExp, int hiExp) { float a = fabsf(v); return a >= ldexpf(1.0f, loExp) && a <= ldexpf(1.0f, hiExp); }
static bool significand_all_ones(float v) { unsigned u; memcpy(&u, &v, 4); return (u & 0x7fffffu) == 0x7fffffu; }

static void make_fuse_params(FuseParams &P, const dsb200_recon *e, const dsb200_scene *scene, const dsb200_view *view,
                             const dsb200_tracking *trk)
{
    fill_mat(P.M_d, trk->M_d);
    dsb200_mat4_multiply(view->rgb_to_depth_inv, trk->M_d, P.M_rgb.m);   // calib_inv * M_d (CPU .cpp:67)
    memcpy(P.proj_d, view->intr_d, 16); memcpy(P.proj_rgb, view->intr_rgb, 16);
    P.voxelSize = scene->params.voxel_size; P.mu = scene->params.mu; P.maxW = scene->params.max_w;
    P.W = view->depth_w; P.H = view->depth_h; P.Wrgb = view->rgb_w; P.Hrgb = view->rgb_h;
    P.stopMaxW = scene->params.stop_integrating_at_max_w;
    P.approximate = trk->requires_full_rendering ? 0 : 1;
    P.depthWeighting = e->weights.depth_weighting; P.maxNewW = e->weights.max_new_w; P.maxDistance = e->weights.max_distance;
    // exact-division support (exactdiv.cuh).  Host IEEE divisions give the correctly rounded reciprocals; fastDiv states
    // that no intermediate of the FMA correction sequences can overflow or go denormal for ANY voxel of ANY block:
    // |pt| <= 2^18 * voxelSize, so with these windows |proj * pc.xy| stays below 2^68.
    P.rcpMu = 1.0f / P.mu; P.rcp255 = 1.0f / 255.0f; P.rcp32767 = 1.0f / 32767.0f;
    // Upper side only.  A numerator BELOW the window (|proj * pc.xy| < 2^-84: a voxel within a hair of the optical axis times a
    // pose entry like cos(pi/2) = 6e-17, which synthetic and axis-aligned trajectories are full of) makes the correction
    // sequence inexact, but its quotient is then below 2^-64 in magnitude (|pc.z| >= 2^-20 is tested per voxel) and the only use
    // of the quotient is `q + principal point`, which rounds to the principal point for ANY |q| < 2^-40 when that is >= 1 and
    // fails the `u >= 1` image test either way when it is smaller: the voxel's fate and value do not depend on it.  (Round 2
    // until r02ac had a per-factor lower window here that sent every such frame -- most of C1 and C3 -- to the plain-division
    // kernels; tests/test_gpu_bench_config.py runs those sequences against the CPU checker bit for bit.)
    bool ok = in_window(P.voxelSize, -20, 10) && in_window(P.mu, -20, 20) && !significand_all_ones(P.mu);
    for (int i = 0; i < 16 && ok; i++) ok = fabsf(P.M_d.m[i]) <= ldexpf(1.0f, 20) && fabsf(P.M_rgb.m[i]) <= ldexpf(1.0f, 20);
    for (int i = 0; i < 4 && ok; i++) ok = fabsf(P.proj_d[i]) <= ldexpf(1.0f, 20) && fabsf(P.proj_rgb[i]) <= ldexpf(1.0f, 20);
    P.fastDiv = (ok && !e->forceSlowDiv) ? 1 : 0;
}

// the shadow to hand to a kernel that writes / reads this scene's voxels (nullptr when disabled); brought up to date first if
// the engine last maintained it for another scene (one engine serves all sub-maps in the reference, ITMMapGraphManager.h:60)
static short *shadow_for(dsb200_recon *e, const dsb200_scene *scene)
{
    if (!e->sdfShadow || !e->shadowEnabled) return nullptr;
    if (e->shadowFor != scene->voxels) {
        launch_rebuild_shadow(e->stream, scene->voxels, (size_t)scene->num_blocks * DSB200_BLOCK_SIZE3, e->sdfShadow);
        e->shadowFor = scene->voxels;
    }
    return e->sdfShadow;
}

// deep look-ahead buffers: 8 ints per voxel block of capacity (slot, parent, 6 saved words) -- a frame cannot allocate more
// blocks than the free list holds -- and a second visible list as large as the render state's
static DeepList deep_list(const dsb200_recon *e)
{
    DeepList d; d.slot = e->deepBuf; d.parent = e->deepBuf + e->numBlocks; d.old = e->deepBuf + 2 * (size_t)e->numBlocks; d.capacity = e->numBlocks;
    return d;
}
static bool deep_buffers(dsb200_recon *e, int visibleCapacity)
{
    if (!e->deepBuf && cudaMalloc(&e->deepBuf, (size_t)e->numBlocks * 8 * sizeof(int)) != cudaSuccess) { cudaGetLastError(); e->deepBuf = nullptr; return false; }
    if (visibleCapacity > e->nextVisibleCap) {
        cudaFree(e->nextVisible); e->nextVisible = nullptr; e->nextVisibleCap = 0;
        if (cudaMalloc(&e->nextVisible, (size_t)visibleCapacity * sizeof(int)) != cudaSuccess) { cudaGetLastError(); return false; }
        e->nextVisibleCap = visibleCapacity;
    }
    return true;
}

// Drop a pending look-ahead: wait for its kernels, then undo what they wrote (request keys / bitmaps, visibility flags) so
// that the next, un-prefetched allocation starts from the state the look-ahead found.
static int abandon_prefetch(dsb200_recon *e)
{
    if (!e->prefetchPending) return DSB200_OK;
    const dsb200_recon::PrefetchKey &k = e->prefetchKey;
    CU(cudaStreamWaitEvent(e->stream, e->evPrefetchDone, 0));
    if (k.deep) launch_abandon_allocation(e->stream, e->fs, k.hashRW, deep_list(e));   // the table is bit for bit what the look-ahead found
    launch_abandon_requests(e->stream, e->fs, k.hash, k.visibleIds, k.visTypeRW, e->reqKey, e->total);
    CU(cudaGetLastError());
    e->prefetchPending = false;
    return DSB200_OK;
}

// W * H bytes of engine scratch for the per-pixel depth weights (nullptr if the allocation fails: the kernels then compute them per voxel)
static unsigned char *weight_scratch(dsb200_recon *e, const dsb200_view *view)
{
    const size_t need = (size_t)view->depth_w * view->depth_h;
    if (need > e->weightBytes) {
        cudaFree(e->weightImage); e->weightImage = nullptr; e->weightBytes = 0;
        if (cudaMalloc(&e->weightImage, need) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        e->weightBytes = need;
    }
    return e->weightImage;
}

// Every scene-modifying entry point passes through here.  consumesPrefetch: the call is an allocation that may pick up a
// pending look-ahead; any other call while one is pending would invalidate what the look-ahead already wrote.
static int check_scene(dsb200_recon *e, const dsb200_scene *scene, bool consumesPrefetch = false)
{
    if (!e || !scene || !scene->voxels || !scene->hash || !scene->alloc_list || !scene->excess_list)
        return fail(DSB200_ERR_INVALID, "null engine/scene pointer");
    if (scene->num_blocks != e->numBlocks || scene->bucket_num != e->bucketNum || scene->excess_size != e->excessSize)
        return fail(DSB200_ERR_INVALID, "scene capacities differ from the capacities the engine was created with");
    if (e->prefetchPending && !consumesPrefetch)
        return fail(DSB200_ERR_INVALID, "a prefetched allocation is pending: the next scene call must be the allocation of that frame");
    e->modEpoch++;
    return DSB200_OK;
}

// enqueue: setToType3 + requests + ordered allocation + ordered visible list (+ snapshot)
static int enqueue_allocate(dsb200_recon *e, dsb200_scene *scene, const dsb200_view *view, const dsb200_tracking *trk,
                            dsb200_render_state *rs, int onlyVisible, int isDefusion)
{
    AllocParams AP; make_alloc_params(AP, e, scene, view, trk);
    if ((long long)AP.W * AP.H >= (1ll << 24)) return fail(DSB200_ERR_INVALID, "depth image too large for the 24-bit pixel order key");
    // noSteps = ceil(2 * |segment| / blockSize) with |segment| = 2*mu (+ rounding): must fit the 8-bit step field
    if (4.0f * AP.mu * AP.oneOverBlock + 2.0f > 250.0f) return fail(DSB200_ERR_INVALID, "mu / block size ratio too large for the 8-bit step order key");
    const bool useHost = !(e->async && e->deviceCountersValid);
    bool deepDone = false;
    if (rs->visible_type != e->visBitsOwner) {     // a render state this engine has not been tracking: derive its bitmap
        launch_rebuild_visbits(e->stream, e->fs, rs->visible_type, e->total);
        e->visBitsOwner = rs->visible_type;
    }
#define DSB_MARK_A(i) do { if (e->profile) cudaEventRecord(e->evAlloc[i], e->stream); } while (0)
    DSB_MARK_A(0);
    VisParams VP; fill_mat(VP.M, trk->M_d); memcpy(VP.proj, view->intr_d, 16);
    VP.voxelSize = scene->params.voxel_size; VP.W = AP.W; VP.H = AP.H;
    VP.useSwapping = (scene->use_swapping && !onlyVisible) ? 1 : 0;
    if (e->prefetchPending) {
        const dsb200_recon::PrefetchKey &k = e->prefetchKey;
        if (k.depth != view->depth || k.w != AP.W || k.h != AP.H || k.visType != rs->visible_type || memcmp(k.M, trk->M_d, 64) ||
            memcmp(k.invM, trk->invM_d, 64) || memcmp(k.intr, view->intr_d, 16) || useHost || k.useSwapping != VP.useSwapping ||
            (k.deep && (onlyVisible || k.hashRW != scene->hash))) {
            // the look-ahead is abandoned and undone (the scene is as it found it); the caller may allocate again
            int rc = abandon_prefetch(e); if (rc) return rc;
            return fail(DSB200_ERR_INVALID, "the pending prefetched allocation was made for a different view / pose / render state "
                                            "(it has been dropped and undone)");
        }
        e->prefetchPending = false;
        CU(cudaStreamWaitEvent(e->stream, e->evPrefetchDone, 0));     // setToType3 + requests + frustum re-test already ran on the look-ahead stream
        deepDone = k.deep;                                            // ... and so did the ordered allocation and the visible-list build
        DSB_MARK_A(1);
    } else {
        launch_begin_frame(e->stream, e->fs, scene->last_free_block, scene->last_free_excess, rs->no_visible, useHost ? 1 : 0,
                           rs->visible_ids, rs->visible_type, useHost ? rs->no_visible : rs->visible_capacity);
        DSB_MARK_A(1);
        launch_alloc_requests(e->stream, AP, e->fs, view->depth, scene->hash, rs->visible_type, e->reqKey);
        // Frustum re-test of last frame's blocks that no depth ray re-observed (flag 3).  It reads only flags and block positions,
        // so it runs BEFORE the ordered allocation (which then overrides the flag of a stale slot it re-allocates) -- and with
        // the look-ahead it moves under the previous frame's raycast together with the request pass.
        launch_recheck_visibility(e->stream, VP, e->fs, scene->hash, rs->visible_ids, rs->visible_type);
    }
    // The host mirrors seeded the device counters above (first asynchronous frame of an engine that has never synchronised);
    // from here on the device copies are the authoritative ones until the next dsb200_recon_sync.
    if (e->async) e->deviceCountersValid = true;
    DSB_MARK_A(2);
    if (!deepDone)
        launch_allocate(e->stream, AP, e->fs, view->depth, scene->hash, scene->alloc_list, scene->excess_list, rs->visible_type,
                        e->reqKey, e->total, onlyVisible ? 0 : 1);
    DSB_MARK_A(3);
    // snapshot reservation for decay / slide window / de-fusion (CUDA .cu:257-317)
    SnapPtr snap = std::make_shared<Snapshot>();
    // Synchronous calls trim the reservation to the real count on return.  While asynchronous the count is unknown until
    // the next sync, so reserve twice the last known visible count (+64 Ki) instead of the full capacity (8 MB per frame at
    // the bench capacities); k_visible_list bounds its writes and flags an overflow, reported by the next sync.
    size_t snapCap = (size_t)rs->visible_capacity;
    if (e->async && e->deviceCountersValid && 2 * (size_t)e->lastKnownVisible + 65536 < snapCap) snapCap = 2 * (size_t)e->lastKnownVisible + 65536;
    snap->regionInts = 64 + ((snapCap + 63) & ~(size_t)63);
    snap->chunk = e->arena.reserve(snap->regionInts, &snap->region);
    if (!snap->chunk) return fail(DSB200_ERR_CUDA, "out of device memory for the visible-list snapshot arena");
    snap->countDev = snap->region; snap->ids = snap->region + 64;
    DSB_MARK_A(4);
    if (deepDone)      // pending entries -> entries, the list built beside the last raycast -> the render state (+ snapshot), counters
        launch_commit_allocation(e->stream, e->fs, scene->hash, deep_list(e), e->nextVisible, rs->visible_ids, rs->visible_capacity,
                                 snap->ids, snap->countDev, (int)snapCap);
    else
        launch_visible_list(e->stream, VP, e->fs, scene->hash, scene->alloc_list, rs->visible_type, rs->visible_ids,
                            rs->visible_capacity, snap->ids, snap->countDev, (int)snapCap, e->total);
    DSB_MARK_A(5);
    if (e->profile) e->evAllocRecorded = true;
    e->unsynced.push_back(snap);
    if (!isDefusion) {
        snap->frameIdx = e->frameIdx;
        e->snapshots[view->timestamp] = snap;
        e->decayQ.push_back(snap); e->slideQ.push_back(snap);
        e->frameIdx++;
    } else {
        std::map<double, SnapPtr>::iterator it = e->snapshots.find(view->timestamp);
        if (it != e->snapshots.end()) {
            snap->frameIdx = it->second->frameIdx;
            e->decayDefusionQ.push_back(snap); e->slideDefusionQ.push_back(snap);
            e->snapshots[view->timestamp] = snap;
        }
    }
    return DSB200_OK;
}

static int finish_call(dsb200_recon *e, dsb200_scene *scene, dsb200_render_state *rs, bool checkExhaustion)
{
    CU(cudaGetLastError());
    if (e->async) return DSB200_OK;
    int rc = sync_counters(e, scene, rs);
    if (rc != DSB200_OK) return rc;
    if (checkExhaustion) {
        if (scene->last_free_block < 0)
            return fail(DSB200_ERR_VBA_EXHAUSTED, "Invalid free voxel block ID. InfiniTAM has run out of space in the Voxel Block Array.");
        if (scene->last_free_excess < 0)
            return fail(DSB200_ERR_EXCESS_EXHAUSTED, "Invalid free excess list slot ID. InfiniTAM has run out of slots in the hash table excess list. "
                                                     "Consider increasing the size of the excess list or the number of buckets.");
    }
    return DSB200_OK;
}

extern "C" {

int dsb200_reset_scene(dsb200_recon *e, dsb200_scene *scene)
{
    if (e && e->prefetchPending) {       // a reset abandons a pending look-ahead: let its kernels finish, then wipe what they wrote
        if (scene && e->prefetchKey.hashRW != scene->hash) {      // ... made for ANOTHER scene of this engine: undone there, exactly
            int rc = abandon_prefetch(e); if (rc) return rc;
        } else {
            CU(cudaStreamWaitEvent(e->stream, e->evPrefetchDone, 0));
            e->prefetchPending = false;
        }
    }
    int rc = check_scene(e, scene); if (rc) return rc;
    e->totalDecayed = 0;
    e->decayQ.clear(); e->slideQ.clear(); e->snapshots.clear();   // CUDA .cu:116-131
    launch_reset_scene(e->stream, scene->voxels, (size_t)scene->num_blocks * DSB200_BLOCK_SIZE3, scene->hash, e->total,
                       scene->alloc_list, scene->num_blocks, scene->excess_list, scene->excess_size, nullptr, e->reqKey,
                       e->fs.counters, (e->sdfShadow && e->shadowEnabled) ? e->sdfShadow : nullptr);
    if (e->sdfShadow && e->shadowEnabled) e->shadowFor = scene->voxels;
    CU(cudaMemsetAsync(e->fs.reqBits, 0, (size_t)e->fs.bitTiles * 8192, e->stream));   // the keys were cleared: so are their bits
    CU(cudaGetLastError());
    scene->last_free_block = scene->num_blocks - 1;
    scene->last_free_excess = scene->excess_size - 1;
    e->deviceCountersValid = true;       // k_reset_scene initialised every device counter
    e->lastKnownVisible = 0;
    if (!e->async) CU(cudaStreamSynchronize(e->stream));
    return DSB200_OK;
}

int dsb200_allocate_from_depth(dsb200_recon *e, dsb200_scene *scene, const dsb200_view *view, const dsb200_tracking *trk,
                               dsb200_render_state *rs, int onlyVisible, int isDefusion)
{
    int rc = check_scene(e, scene, true); if (rc) return rc;
    if (!view || !trk || !rs || !view->depth) return fail(DSB200_ERR_INVALID, "allocate: null view/tracking/render state");
    rc = enqueue_allocate(e, scene, view, trk, rs, onlyVisible, isDefusion); if (rc) return rc;
    return finish_call(e, scene, rs, true);
}

int dsb200_integrate(dsb200_recon *e, dsb200_scene *scene, const dsb200_view *view, const dsb200_tracking *trk,
                     const dsb200_render_state *rs)
{
    int rc = check_scene(e, scene); if (rc) return rc;
    FuseParams FP; make_fuse_params(FP, e, scene, view, trk);
    const int *countPtr = &e->fs.counters->noVisible;
    if (!(e->async && e->deviceCountersValid)) {
        // host mirror is authoritative between synchronous calls
        e->hostCounters->noVisible = rs->no_visible;
        CU(cudaMemcpyAsync(&e->fs.counters->noVisible, &e->hostCounters->noVisible, 4, cudaMemcpyHostToDevice, e->stream));
    }
    launch_fuse(e->stream, FP, false, scene->voxels, scene->hash, rs->visible_ids, countPtr, view->depth, view->rgb, shadow_for(e, scene),
                weight_scratch(e, view));
    CU(cudaGetLastError());
    if (!e->async) CU(cudaStreamSynchronize(e->stream));
    return DSB200_OK;
}

int dsb200_deintegrate(dsb200_recon *e, dsb200_scene *scene, const dsb200_view *view, const dsb200_tracking *trk)
{
    int rc = check_scene(e, scene); if (rc) return rc;
    std::map<double, SnapPtr>::iterator it = e->snapshots.find(view->timestamp);
    if (it == e->snapshots.end()) return DSB200_OK;            // silent no-op, CUDA .cu:437-440
    if (it->second->countHost == 0) return DSB200_OK;
    FuseParams FP; make_fuse_params(FP, e, scene, view, trk);
    launch_fuse(e->stream, FP, true, scene->voxels, scene->hash, it->second->ids, it->second->countDev, view->depth, view->rgb,
                shadow_for(e, scene));
    CU(cudaGetLastError());
    if (!e->async) CU(cudaStreamSynchronize(e->stream));
    return DSB200_OK;
}

} // extern "C"

// Decay / DecayDefusionPart / SlideWindow / SlideWindowDefusionPart share the queue policy of CUDA .cu:913-1077
static int prune_from_queue(dsb200_recon *e, dsb200_scene *scene, dsb200_render_state *rs, std::deque<SnapPtr> &q,
                            long threshold, bool slide, int minAge, int maxWeight)
{
    int rc = check_scene(e, scene); if (rc) return rc;
    if ((long)q.size() > threshold) {
        SnapPtr snap = q.front(); q.pop_front();
        if (snap->countHost != 0) {
            if (!(e->async && e->deviceCountersValid)) {
                e->hostCounters->lastFreeBlock = scene->last_free_block;
                CU(cudaMemcpyAsync(&e->fs.counters->lastFreeBlock, &e->hostCounters->lastFreeBlock, 4, cudaMemcpyHostToDevice, e->stream));
            }
            int bound = snap->countHost >= 0 ? snap->countHost : (int)(snap->regionInts - 64);
            if (bound > e->pruneCap) return fail(DSB200_ERR_INVALID, "snapshot larger than the pruning scratch");
            launch_prune(e->stream, slide, scene->voxels, scene->hash, scene->bucket_num, snap->ids, snap->countDev, e->pruneCap,
                         minAge, maxWeight, (int)e->frameIdx, scene->alloc_list, rs->visible_type,
                         rs->visible_type == e->visBitsOwner ? e->fs.visBits : nullptr, e->slotMark, e->pruneScratch,
                         e->fs.counters, e->fs, shadow_for(e, scene));
        }
    }
    return finish_call(e, scene, rs, false);
}

extern "C" {

int dsb200_decay(dsb200_recon *e, dsb200_scene *s, dsb200_render_state *rs, int maxWeight, int minAge, int)
{ return prune_from_queue(e, s, rs, e->decayQ, minAge, false, minAge, maxWeight); }
int dsb200_decay_defusion(dsb200_recon *e, dsb200_scene *s, dsb200_render_state *rs, int maxWeight, int minAge, int)
{ return prune_from_queue(e, s, rs, e->decayDefusionQ, 0, false, minAge, maxWeight); }
int dsb200_slide_window(dsb200_recon *e, dsb200_scene *s, dsb200_render_state *rs, int maxAge)
{ return prune_from_queue(e, s, rs, e->slideQ, maxAge, true, 0, 0); }
int dsb200_slide_window_defusion(dsb200_recon *e, dsb200_scene *s, dsb200_render_state *rs, int maxAge, int maxSize)
{ (void)maxAge; return prune_from_queue(e, s, rs, e->slideDefusionQ, maxSize, true, 0, 0); }

} // extern "C"

// ---------------------------------------------------------------- visualisation engine
static int vis_ensure_scan(dsb200_vis *v, int total)
{
    int tiles = div_up(total, 1024);
    if (tiles <= v->scanTilesAlloc) { v->fs.scanTiles = tiles; return DSB200_OK; }
    if (v->fs.scanState) CU(cudaFree(v->fs.scanState));
    CU(cudaMalloc(&v->fs.scanState, (size_t)(2 * tiles + 8) * 8));
    CU(cudaMemset(v->fs.scanState, 0, (size_t)(2 * tiles + 8) * 8));
    v->scanTilesAlloc = tiles; v->fs.scanTiles = tiles;
    return DSB200_OK;
}

static void make_ray_params(RayParams &P, const dsb200_scene *scene, const float *M, const float *invM, const float *intr,
                            int W, int H)
{
    fill_mat(P.invM, invM);
    if (M) fill_mat(P.M, M); else memset(P.M.m, 0, 64);
    P.invProj[0] = 1.0f / intr[0]; P.invProj[1] = 1.0f / intr[1]; P.invProj[2] = intr[2]; P.invProj[3] = intr[3];
    P.oneOverVoxel = 1.0f / scene->params.voxel_size; P.mu = scene->params.mu; P.voxelSize = scene->params.voxel_size;
    P.light[0] = -invM[8]; P.light[1] = -invM[9]; P.light[2] = -invM[10];      // -invM.getColumn(2)
    P.W = W; P.H = H; P.bucketNum = scene->bucket_num; P.bucketMask = scene->bucket_num - 1;
}

static void make_proj_params(ProjParams &P, const dsb200_scene *scene, const float *M, const float *intr, int W, int H)
{
    fill_mat(P.M, M); memcpy(P.proj, intr, 16);
    P.voxelSize = scene->params.voxel_size; P.W = W; P.H = H; P.maxRenderBlocks = 65536 * 4;   // MAX_RENDERING_BLOCKS
}

// stand-alone (synchronous) vis calls take the visible count from the caller's host mirror
static int vis_push_count(dsb200_vis *v, int n)
{
    v->hostInts[0] = n;
    CU(cudaMemcpyAsync(&v->fs.counters->noVisible, &v->hostInts[0], 4, cudaMemcpyHostToDevice, v->stream));
    return DSB200_OK;
}

extern "C" {

int dsb200_vis_create(dsb200_vis **out, int32_t max_visible, int32_t width, int32_t height, void *stream)
{
    (void)max_visible;
    if (!out || width <= 0 || height <= 0) return fail(DSB200_ERR_INVALID, "vis_create: bad arguments");
    if (dsb200_device_count() == 0) return fail(DSB200_ERR_NO_DEVICE, "no CUDA device: libdsb200 has no CPU fallback");
    dsb200_vis *v = new dsb200_vis();
    v->stream = (cudaStream_t)stream;
    CU(cudaMalloc(&v->fs.counters, sizeof(Counters)));
    CU(cudaMemset(v->fs.counters, 0, sizeof(Counters)));
    CU(cudaMalloc(&v->fs.tickets, 4 * sizeof(int)));
    CU(cudaMemset(v->fs.tickets, 0, 4 * sizeof(int)));
    CU(cudaMalloc(&v->countScratch, 4 * sizeof(int)));
    CU(cudaMallocHost(&v->hostCounters, sizeof(Counters)));
    CU(cudaMallocHost(&v->hostInts, 4 * sizeof(int)));
    // ray queue: one slot per pixel (36 bytes); a larger image than this simply runs the single unbounded raycast pass
    const size_t npx = (size_t)width * height;
    CU(cudaMalloc(&v->rayQueue.ctl, 4 * sizeof(int)));
    CU(cudaMemset(v->rayQueue.ctl, 0, 4 * sizeof(int)));
    CU(cudaMalloc(&v->rayQueue.pixel, npx * sizeof(int)));
    CU(cudaMalloc(&v->rayQueue.posLen, npx * sizeof(float4)));
    CU(cudaMalloc(&v->rayQueue.dirMax, npx * sizeof(float4)));
    v->tileOrderCap = div_up(width, 8) * div_up(height, 4);
    CU(cudaMalloc(&v->tileOrder, (size_t)(4 * v->tileOrderCap + 4) * sizeof(int)));      // 4 class counters + 4 lists (vis.cu: k_tile_order)
    CU(cudaMemset(v->tileOrder, 0, 4 * sizeof(int)));                    // k_tile_order's two counters (k_range_init zeroes them per frame)
    v->rayQueue.capacity = (int)npx;
    *out = v;
    return DSB200_OK;
}

int dsb200_vis_destroy(dsb200_vis *v)
{
    if (!v) return DSB200_OK;
    cudaStreamSynchronize(v->stream);
    cudaFree(v->fs.counters); cudaFree(v->fs.tickets); cudaFree(v->fs.scanState); cudaFree(v->countScratch);
    cudaFree(v->fwdProjection); cudaFree(v->fwdMissing);
    cudaFree(v->rayQueue.ctl); cudaFree(v->rayQueue.pixel); cudaFree(v->rayQueue.posLen); cudaFree(v->rayQueue.dirMax);
    cudaFreeHost(v->hostCounters); cudaFreeHost(v->hostInts); cudaFree(v->tileOrder);
    delete v;
    return DSB200_OK;
}

int dsb200_find_visible_blocks(dsb200_vis *v, const dsb200_scene *scene, const float *M, const float *intr, dsb200_render_state *rs)
{
    if (!v || !scene || !rs) return fail(DSB200_ERR_INVALID, "find_visible_blocks: null argument");
    const int total = scene->bucket_num + scene->excess_size;
    int rc = vis_ensure_scan(v, total); if (rc) return rc;
    VisParams P; fill_mat(P.M, M); memcpy(P.proj, intr, 16);
    P.voxelSize = scene->params.voxel_size; P.W = rs->width; P.H = rs->height; P.useSwapping = 0;
    launch_find_visible(v->stream, P, v->fs, scene->hash, rs->visible_ids, rs->visible_capacity, total);
    CU(cudaMemcpyAsync(v->hostCounters, v->fs.counters, sizeof(Counters), cudaMemcpyDeviceToHost, v->stream));
    CU(cudaStreamSynchronize(v->stream));
    rs->no_visible = v->hostCounters->noVisible;
    return DSB200_OK;
}

int dsb200_count_visible_blocks(dsb200_vis *v, const dsb200_scene *scene, const dsb200_render_state *rs, int minId, int maxId, int *count)
{
    int rc = vis_push_count(v, rs->no_visible); if (rc) return rc;
    launch_count_visible(v->stream, scene->hash, rs->visible_ids, &v->fs.counters->noVisible, minId, maxId, v->countScratch + 1);
    CU(cudaMemcpyAsync(&v->hostInts[1], v->countScratch + 1, 4, cudaMemcpyDeviceToHost, v->stream));
    CU(cudaStreamSynchronize(v->stream));
    *count = v->hostInts[1];
    return DSB200_OK;
}

int dsb200_create_expected_depths(dsb200_vis *v, const dsb200_scene *scene, const float *M, const float *intr, dsb200_render_state *rs)
{
    int rc = vis_push_count(v, rs->no_visible); if (rc) return rc;
    ProjParams P; make_proj_params(P, scene, M, intr, rs->width, rs->height);
    launch_expected_depths(v->stream, P, scene->hash, rs->visible_ids, &v->fs.counters->noVisible, rs->range_image, nullptr, v->fs.counters);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(v->stream));
    return DSB200_OK;
}

int dsb200_find_surface(dsb200_vis *v, const dsb200_scene *scene, const float *invM, const float *intr, dsb200_render_state *rs)
{
    RayParams P; make_ray_params(P, scene, nullptr, invM, intr, rs->width, rs->height);
    launch_raycast(v->stream, P, scene->voxels, scene->hash, rs->range_image, rs->raycast_result, v->fs.tickets + 2, nullptr, &v->rayQueue);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(v->stream));
    return DSB200_OK;
}

int dsb200_create_icp_maps(dsb200_vis *v, const dsb200_scene *scene, const dsb200_view *view, dsb200_tracking *trk, dsb200_render_state *rs)
{
    if (!trk->icp_points || !trk->icp_normals) return fail(DSB200_ERR_INVALID, "create_icp_maps: tracking state has no point cloud buffers");
    RayParams P; make_ray_params(P, scene, trk->M_d, trk->invM_d, view->intr_d, rs->width, rs->height);
    launch_raycast(v->stream, P, scene->voxels, scene->hash, rs->range_image, rs->raycast_result, v->fs.tickets + 2, nullptr, &v->rayQueue);
    launch_icp_maps(v->stream, P, rs->raycast_result, trk->icp_points, trk->icp_normals, rs->raycast_image);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(v->stream));
    return DSB200_OK;
}

int dsb200_render_image(dsb200_vis *v, const dsb200_scene *scene, const float *M, const float *invM, const float *intr,
                        dsb200_render_state *rs, uint8_t *outRGBA, float *outDepth, int type)
{
    if (!outRGBA) return fail(DSB200_ERR_INVALID, "render_image: null output image");
    if (type == DSB200_RENDER_COLOURCODED) type = DSB200_RENDER_SHADED_GREYSCALE;   // !TIndex::hasColorCoding
    RayParams P; make_ray_params(P, scene, M, invM, intr, rs->width, rs->height);
    launch_raycast(v->stream, P, scene->voxels, scene->hash, rs->range_image, rs->raycast_result, v->fs.tickets + 2, nullptr, &v->rayQueue);
    launch_shade(v->stream, P, type, scene->voxels, scene->hash, rs->raycast_result, outRGBA, outDepth);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(v->stream));
    return DSB200_OK;
}

int dsb200_composite_nearest(dsb200_vis *v, int n, const float *const *results, const float *invM, float voxel_size,
                             int width, int height, int row_begin, int row_end, float *out)
{
    if (!v || !results || !invM || !out || n < 1 || n > kMaxShards) return fail(DSB200_ERR_INVALID, "composite_nearest: bad arguments (1 <= n <= 16)");
    if (row_begin < 0 || row_end > height || row_begin > row_end) return fail(DSB200_ERR_INVALID, "composite_nearest: bad row range");
    for (int i = 0; i < n; i++) if (!results[i]) return fail(DSB200_ERR_INVALID, "composite_nearest: null raycast result");
    // camera centre in voxel units, as castRay forms its points: (invM * p) * (1 / voxelSize)
    const float oneOverVoxel = 1.0f / voxel_size;
    launch_composite_nearest(v->stream, n, results, invM[12] * oneOverVoxel, invM[13] * oneOverVoxel, invM[14] * oneOverVoxel,
                             width, row_begin, row_end, out);
    CU(cudaGetLastError());
    return DSB200_OK;     // stream-ordered: the caller synchronises (the sharded path keeps the frame in flight)
}

int dsb200_raycast_sharded(dsb200_vis *v, const dsb200_scene *scene, const dsb200_shard_tables *t, const float *invM,
                           const float *intr, int width, int height, float *merged_range, float *out_raycast)
{
    if (!v || !scene || !t || !invM || !intr || !merged_range || !out_raycast) return fail(DSB200_ERR_INVALID, "raycast_sharded: null argument");
    if (t->count < 1 || t->count > kMaxShards || t->rank < 0 || t->rank >= t->count || t->brick_log2 < 0 || t->brick_log2 > 12)
        return fail(DSB200_ERR_INVALID, "raycast_sharded: need 0 <= rank < count <= 16 and 0 <= brick_log2 <= 12");
    for (int i = 0; i < t->count; i++)
        if (!t->hash[i] || !t->voxels[i] || !t->range_image[i]) return fail(DSB200_ERR_INVALID, "raycast_sharded: null table pointer");
    RayParams P; make_ray_params(P, scene, nullptr, invM, intr, width, height);
    launch_raycast_sharded(v->stream, P, t->count, t->rank, t->brick_log2, t->hash, t->voxels, t->range_image, merged_range,
                           out_raycast, v->fs.tickets + 2, &v->rayQueue);
    CU(cudaGetLastError());
    return DSB200_OK;
}

int dsb200_icp_maps_from_raycast(dsb200_vis *v, const dsb200_scene *scene, const dsb200_view *view, dsb200_tracking *trk,
                                 const float *raycast_result, uint8_t *out_rendering, int width, int height)
{
    if (!v || !scene || !view || !trk || !raycast_result || !out_rendering) return fail(DSB200_ERR_INVALID, "icp_maps_from_raycast: null argument");
    if (!trk->icp_points || !trk->icp_normals) return fail(DSB200_ERR_INVALID, "icp_maps_from_raycast: tracking state has no point cloud buffers");
    RayParams P; make_ray_params(P, scene, trk->M_d, trk->invM_d, view->intr_d, width, height);
    launch_icp_maps(v->stream, P, raycast_result, trk->icp_points, trk->icp_normals, out_rendering);
    CU(cudaGetLastError());
    return DSB200_OK;
}

int dsb200_ipc_export(void *p, uint8_t handle[DSB200_IPC_HANDLE_BYTES])
{
    static_assert(sizeof(cudaIpcMemHandle_t) == DSB200_IPC_HANDLE_BYTES, "IPC handle size");
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, p));
    memcpy(handle, &h, sizeof(h));
    return DSB200_OK;
}
int dsb200_ipc_open(const uint8_t handle[DSB200_IPC_HANDLE_BYTES], void **p)
{
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    CU(cudaIpcOpenMemHandle(p, h, cudaIpcMemLazyEnablePeerAccess));
    return DSB200_OK;
}
int dsb200_ipc_close(void *p) { CU(cudaIpcCloseMemHandle(p)); return DSB200_OK; }

int dsb200_forward_render(dsb200_vis *v, const dsb200_scene *scene, const dsb200_view *view, dsb200_tracking *trk, dsb200_render_state *rs)
{
    // only reached with useApproximateRaycast, which DenseSLAM never sets; built for interface completeness
    if (!v || !scene || !view || !trk || !rs || !view->depth) return fail(DSB200_ERR_INVALID, "forward_render: null argument");
    const int px = rs->width * rs->height;
    if (px > v->fwdPixels) {
        if (v->fwdProjection) { CU(cudaFree(v->fwdProjection)); CU(cudaFree(v->fwdMissing)); }
        CU(cudaMalloc(&v->fwdProjection, (size_t)px * 16));
        CU(cudaMalloc(&v->fwdMissing, (size_t)px * 4));
        v->fwdPixels = px;
    }
    RayParams P; make_ray_params(P, scene, trk->M_d, trk->invM_d, view->intr_d, rs->width, rs->height);
    launch_forward_render(v->stream, P, view->intr_d, scene->voxels, scene->hash, rs->range_image, rs->raycast_result, view->depth,
                          v->fwdProjection, v->fwdMissing, v->countScratch + 2, rs->raycast_image);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(&v->hostInts[2], v->countScratch + 2, 4, cudaMemcpyDeviceToHost, v->stream));
    CU(cudaStreamSynchronize(v->stream));
    return DSB200_OK;
}

int dsb200_forward_projection(dsb200_vis *v, const float **fwd, int *no_missing)
{
    if (!v || !v->fwdProjection) return fail(DSB200_ERR_INVALID, "forward_projection: ForwardRender has not run");
    if (fwd) *fwd = v->fwdProjection;
    if (no_missing) *no_missing = v->hostInts[2];
    return DSB200_OK;
}

int dsb200_create_point_cloud(dsb200_vis *)
{
    return fail(DSB200_ERR_UNSUPPORTED, "CreatePointCloud serves the colour tracker, which DenseSLAM never selects");
}

int dsb200_view_from_raw(const int16_t *raw_depth, const uint8_t *bgr, float scale, float offset, dsb200_view *view, void *stream)
{
    if (!raw_depth || !view || !view->depth || (bgr && !view->rgb)) return fail(DSB200_ERR_INVALID, "view_from_raw: null argument");
    if (bgr && (view->depth_w != view->rgb_w || view->depth_h != view->rgb_h))
        return fail(DSB200_ERR_INVALID, "view_from_raw: depth and colour images must have the same size");
    if (((uintptr_t)raw_depth & 7) || ((uintptr_t)bgr & 3) || ((uintptr_t)view->depth & 15) || ((uintptr_t)view->rgb & 15))
        return fail(DSB200_ERR_INVALID, "view_from_raw: buffers must be 16-byte aligned (raw depth 8, bgr 4)");
    launch_view_from_raw((cudaStream_t)stream, raw_depth, bgr, scale, offset, const_cast<float *>(view->depth),
                         const_cast<uint8_t *>(view->rgb), view->depth_w * view->depth_h);
    CU(cudaGetLastError());
    return DSB200_OK;
}

// ---------------------------------------------------------------- ICP depth tracker (device half)
struct dsb200_tracker {
    cudaStream_t stream = nullptr;
    int levels = 0, width = 0, height = 0;
    float *owned[DSB200_TRACKER_MAX_LEVELS] = {};        // levels 1.. (level 0 is the caller's image)
    const float *depth[DSB200_TRACKER_MAX_LEVELS] = {};
    int w[DSB200_TRACKER_MAX_LEVELS] = {}, h[DSB200_TRACKER_MAX_LEVELS] = {};
    float intr[DSB200_TRACKER_MAX_LEVELS][4] = {};
    const float *points = nullptr, *normals = nullptr;
    int sceneW = 0, sceneH = 0;
    float sceneIntr[4] = {}, scenePose[16] = {};
    double *partials = nullptr, *result = nullptr, *hostResult = nullptr;
    unsigned *ticket = nullptr;
    int lastPara = 0;
    double seq = 0.0;                  // evaluation counter echoed by the kernel into hostResult[slots - 1]
};

int dsb200_tracker_create(dsb200_tracker **out, int32_t width, int32_t height, int32_t levels, void *stream)
{
    if (!out || width <= 0 || height <= 0 || levels < 1 || levels > DSB200_TRACKER_MAX_LEVELS)
        return fail(DSB200_ERR_INVALID, "tracker_create: bad size or level count");
    std::unique_ptr<dsb200_tracker> t(new dsb200_tracker);
    t->stream = (cudaStream_t)stream; t->levels = levels; t->width = width; t->height = height;
    t->w[0] = width; t->h[0] = height;
    for (int i = 1; i < levels; i++) {
        t->w[i] = t->w[i - 1] / 2; t->h[i] = t->h[i - 1] / 2;
        if (t->w[i] <= 0 || t->h[i] <= 0) return fail(DSB200_ERR_INVALID, "tracker_create: too many levels for this image size");
        CU(cudaMalloc(&t->owned[i], sizeof(float) * (size_t)t->w[i] * t->h[i]));
        t->depth[i] = t->owned[i];
    }
    CU(cudaMalloc(&t->partials, sizeof(double) * (size_t)tracker_gh_max_ctas() * tracker_gh_slots()));
    CU(cudaMalloc(&t->ticket, sizeof(unsigned)));
    CU(cudaMemset(t->ticket, 0, sizeof(unsigned)));
    // the last CTA writes the 30 sums straight into mapped pinned memory: no copy call between the kernel and the host's read
    CU(cudaHostAlloc(&t->hostResult, sizeof(double) * tracker_gh_slots(), cudaHostAllocMapped));
    CU(cudaHostGetDevicePointer(&t->result, t->hostResult, 0));
    memset(t->hostResult, 0, sizeof(double) * tracker_gh_slots());
    *out = t.release();
    return DSB200_OK;
}

int dsb200_tracker_destroy(dsb200_tracker *t)
{
    if (!t) return DSB200_OK;
    for (int i = 1; i < t->levels; i++) cudaFree(t->owned[i]);
    cudaFree(t->partials); cudaFree(t->ticket); cudaFreeHost(t->hostResult);
    delete t;
    return DSB200_OK;
}

int dsb200_tracker_set_evaluation_data(dsb200_tracker *t, const float *depth, const float *viewIntr, const float *points,
                                       const float *normals, int32_t sceneW, int32_t sceneH, const float *sceneIntr,
                                       const float *scenePose)
{
    if (!t || !depth || !viewIntr || !points || !normals || !sceneIntr || !scenePose || sceneW < 2 || sceneH < 2)
        return fail(DSB200_ERR_INVALID, "tracker_set_evaluation_data: null argument or empty scene maps");
    if (((uintptr_t)points & 15) || ((uintptr_t)normals & 15))
        return fail(DSB200_ERR_INVALID, "tracker_set_evaluation_data: the point and normal maps must be 16-byte aligned");
    if (sceneW > 32767 || sceneH > 32767)
        return fail(DSB200_ERR_INVALID, "tracker_set_evaluation_data: scene maps beyond 32767 pixels a side (the reference indexes them with shorts)");
    t->depth[0] = depth;
    for (int k = 0; k < 4; k++) t->intr[0][k] = viewIntr[k];
    for (int i = 1; i < t->levels; i++) {
        launch_filter_subsample_holes(t->stream, t->depth[i - 1], t->w[i - 1], t->h[i - 1], t->owned[i]);
        for (int k = 0; k < 4; k++) t->intr[i][k] = t->intr[i - 1][k] * 0.5f;
    }
    CU(cudaGetLastError());
    t->points = points; t->normals = normals; t->sceneW = sceneW; t->sceneH = sceneH;
    memcpy(t->sceneIntr, sceneIntr, sizeof(t->sceneIntr));
    memcpy(t->scenePose, scenePose, sizeof(t->scenePose));
    return DSB200_OK;
}

static int tracker_gh(dsb200_tracker *t, const float *depth, int vW, int vH, const float *vIntr, const float *points,
                      const float *normals, int sW, int sH, const float *sIntr, const float *scenePose, int iterationType,
                      float distThresh, const float *approxInvPose, float *f, float *nabla, float *hessian, int32_t *noValid)
{
    if (iterationType == DSB200_TRACKER_ITERATION_NONE) { *noValid = 0; return DSB200_OK; }   // ITMDepthTracker_CPU.cpp:25
    if (iterationType != DSB200_TRACKER_ITERATION_ROTATION && iterationType != DSB200_TRACKER_ITERATION_TRANSLATION &&
        iterationType != DSB200_TRACKER_ITERATION_BOTH)
        return fail(DSB200_ERR_INVALID, "tracker_compute_g_and_h: unknown iteration type");
    launch_tracker_gh(t->stream, iterationType, approxInvPose, scenePose, vIntr, sIntr, distThresh, depth, vW, vH, points, normals,
                      sW, sH, t->partials, t->ticket, t->result, t->seq += 1.0);
    CU(cudaGetLastError());
    // t->result aliases t->hostResult (mapped pinned memory).  Poll the echoed sequence number; fall back to the stream's
    // status every few thousand polls so a failed launch surfaces as an error instead of a spin.
    {
        volatile double *flag = t->hostResult + tracker_gh_slots() - 1;
        for (unsigned spins = 0; *flag != t->seq; ) {
            if ((++spins & 0xfff) == 0) {
                cudaError_t q = cudaStreamQuery(t->stream);
                if (q == cudaSuccess) { if (*flag != t->seq) CU(cudaStreamSynchronize(t->stream)); if (*flag != t->seq) return fail(DSB200_ERR_CUDA, "tracker: result never arrived"); break; }
                if (q != cudaErrorNotReady) return cuda_fail(q, "tracker_compute_g_and_h");
            }
#if defined(__x86_64__)
            __builtin_ia32_pause();
#endif
        }
        __atomic_thread_fence(__ATOMIC_ACQUIRE);      // the sums below are read after the flag
    }
    const bool shortIt = iterationType != DSB200_TRACKER_ITERATION_BOTH;
    const int noPara = shortIt ? 3 : 6, noParaSQ = shortIt ? 6 : 21;
    t->lastPara = noPara;
    const double *S = t->hostResult;
    // ITMDepthTracker_CPU.cpp:71-77
    for (int r = 0, counter = 0; r < noPara; r++) for (int c = 0; c <= r; c++, counter++) hessian[r + c * 6] = (float)S[counter];
    for (int r = 0; r < noPara; ++r) for (int c = r + 1; c < noPara; c++) hessian[r + c * 6] = hessian[c + r * 6];
    for (int r = 0; r < noPara; r++) nabla[r] = (float)S[noParaSQ + r];
    const int n = (int)S[noParaSQ + noPara + 1];
    const float sumF = (float)S[noParaSQ + noPara];
    *f = (n > 100) ? sqrtf(sumF) / (float)n : 1e5f;
    *noValid = n;
    return DSB200_OK;
}

int dsb200_tracker_compute_g_and_h(dsb200_tracker *t, int32_t level, int32_t iterationType, float distThresh,
                                   const float *approxInvPose, float *f, float *nabla, float *hessian, int32_t *noValid)
{
    if (!t || !approxInvPose || !f || !nabla || !hessian || !noValid) return fail(DSB200_ERR_INVALID, "tracker_compute_g_and_h: null argument");
    if (level < 0 || level >= t->levels) return fail(DSB200_ERR_INVALID, "tracker_compute_g_and_h: no such pyramid level");
    if (!t->depth[0] || !t->points) return fail(DSB200_ERR_INVALID, "tracker_compute_g_and_h: set_evaluation_data has not been called");
    return tracker_gh(t, t->depth[level], t->w[level], t->h[level], t->intr[level], t->points, t->normals, t->sceneW, t->sceneH,
                      t->sceneIntr, t->scenePose, iterationType, distThresh, approxInvPose, f, nabla, hessian, noValid);
}

int dsb200_tracker_compute_g_and_h_on(dsb200_tracker *t, const float *depth, int32_t vW, int32_t vH, const float *vIntr,
                                      const float *points, const float *normals, int32_t sW, int32_t sH, const float *sIntr,
                                      const float *scenePose, int32_t iterationType, float distThresh, const float *approxInvPose,
                                      float *f, float *nabla, float *hessian, int32_t *noValid)
{
    if (!t || !depth || !vIntr || !points || !normals || !sIntr || !scenePose || !approxInvPose || !f || !nabla || !hessian || !noValid)
        return fail(DSB200_ERR_INVALID, "tracker_compute_g_and_h_on: null argument");
    if (vW <= 0 || vH <= 0 || sW < 2 || sH < 2 || sW > 32767 || sH > 32767 || ((uintptr_t)points & 15) || ((uintptr_t)normals & 15))
        return fail(DSB200_ERR_INVALID, "tracker_compute_g_and_h_on: bad image size or unaligned maps");
    return tracker_gh(t, depth, vW, vH, vIntr, points, normals, sW, sH, sIntr, scenePose, iterationType, distThresh, approxInvPose, f,
                      nabla, hessian, noValid);
}

int dsb200_tracker_last_sums(const dsb200_tracker *t, double *sums)
{
    if (!t || !sums) return fail(DSB200_ERR_INVALID, "tracker_last_sums: null argument");
    memcpy(sums, t->hostResult, sizeof(double) * 32);
    sums[31] = 0.0;                    // the kernel's sequence echo is not part of the result
    return DSB200_OK;
}

int dsb200_tracker_level(const dsb200_tracker *t, int32_t level, const float **depth, int32_t *w, int32_t *h, float *intr)
{
    if (!t || level < 0 || level >= t->levels) return fail(DSB200_ERR_INVALID, "tracker_level: no such pyramid level");
    if (depth) *depth = t->depth[level];
    if (w) *w = t->w[level];
    if (h) *h = t->h[level];
    if (intr) memcpy(intr, t->intr[level], sizeof(float) * 4);
    return DSB200_OK;
}

int dsb200_filter_subsample_with_holes(const float *in, int32_t w, int32_t h, float *out, void *stream)
{
    if (!in || !out || w < 2 || h < 2) return fail(DSB200_ERR_INVALID, "filter_subsample_with_holes: null argument or image smaller than 2x2");
    launch_filter_subsample_holes((cudaStream_t)stream, in, w, h, out);
    CU(cudaGetLastError());
    return DSB200_OK;
}

int dsb200_depth_filter_pass(const float *in, float *out, int32_t w, int32_t h, void *stream)
{
    if (!in || !out || in == out) return fail(DSB200_ERR_INVALID, "depth_filter_pass: null or aliased images");
    if (w < 5 || h < 5) return fail(DSB200_ERR_INVALID, "depth_filter_pass: image smaller than the 5x5 window");
    launch_filter_depth((cudaStream_t)stream, in, out, w, h);
    CU(cudaGetLastError());
    return DSB200_OK;
}

int dsb200_convert_disparity_to_depth(const int16_t *disparity, float *depth, int32_t w, int32_t h, float param_x, float param_y,
                                      float fx_depth, void *stream)
{
    if (!disparity || !depth || w <= 0 || h <= 0) return fail(DSB200_ERR_INVALID, "convert_disparity_to_depth: null argument");
    launch_disparity_to_depth((cudaStream_t)stream, disparity, depth, param_x, param_y, fx_depth, w * h);
    CU(cudaGetLastError());
    return DSB200_OK;
}

int dsb200_compute_normal_and_weights(const float *depth, float *normal4, float *sigma_z, int32_t w, int32_t h, const float *intrinsics,
                                      void *stream)
{
    if (!depth || !normal4 || !sigma_z || !intrinsics) return fail(DSB200_ERR_INVALID, "compute_normal_and_weights: null argument");
    if (w <= 0 || h <= 0) return fail(DSB200_ERR_INVALID, "compute_normal_and_weights: empty image");
    if (w < 5 || h < 5) return DSB200_OK;                      // no interior pixel: the reference's loops do not run either
    launch_normal_and_weight((cudaStream_t)stream, depth, normal4, sigma_z, w, h, intrinsics);
    CU(cudaGetLastError());
    return DSB200_OK;
}

int dsb200_view_filter_depth(dsb200_view *view, float *scratch, void *stream)
{
    if (!view || !view->depth || !scratch) return fail(DSB200_ERR_INVALID, "view_filter_depth: null argument");
    const int W = view->depth_w, H = view->depth_h;
    if (W < 5 || H < 5) return fail(DSB200_ERR_INVALID, "view_filter_depth: image smaller than the 5x5 window");
    float *depth = const_cast<float *>(view->depth);
    cudaStream_t st = (cudaStream_t)stream;
    // 5 steps of bilateral filtering, CPU/ITMViewBuilder_CPU.cpp:49-58
    launch_filter_depth(st, depth, scratch, W, H);
    launch_filter_depth(st, scratch, depth, W, H);
    launch_filter_depth(st, depth, scratch, W, H);
    launch_filter_depth(st, scratch, depth, W, H);
    launch_filter_depth(st, depth, scratch, W, H);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(depth, scratch, sizeof(float) * (size_t)W * H, cudaMemcpyDeviceToDevice, st));
    return DSB200_OK;
}

int dsb200_selftest_expf(const float *x, float *y, int32_t n)
{
    if (!x || !y || n <= 0) return fail(DSB200_ERR_INVALID, "selftest_expf: null argument");
    if (dsb200_device_count() == 0) return fail(DSB200_ERR_NO_DEVICE, "no CUDA device: libdsb200 has no CPU fallback");
    float *dx = nullptr, *dy = nullptr;
    CU(cudaMalloc(&dx, sizeof(float) * (size_t)n)); CU(cudaMalloc(&dy, sizeof(float) * (size_t)n));
    CU(cudaMemcpy(dx, x, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice));
    launch_expf_selftest(nullptr, dx, dy, n);
    CU(cudaGetLastError());
    CU(cudaMemcpy(y, dy, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost));
    cudaFree(dx); cudaFree(dy);
    return DSB200_OK;
}

int dsb200_selftest_acosf(const float *x, float *y, int32_t n)
{
    if (!x || !y || n <= 0) return fail(DSB200_ERR_INVALID, "selftest_acosf: null argument");
    if (dsb200_device_count() == 0) return fail(DSB200_ERR_NO_DEVICE, "no CUDA device: libdsb200 has no CPU fallback");
    float *dx = nullptr, *dy = nullptr;
    CU(cudaMalloc(&dx, sizeof(float) * (size_t)n)); CU(cudaMalloc(&dy, sizeof(float) * (size_t)n));
    CU(cudaMemcpy(dx, x, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice));
    launch_acosf_selftest(nullptr, dx, dy, n);
    CU(cudaGetLastError());
    CU(cudaMemcpy(y, dy, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost));
    cudaFree(dx); cudaFree(dy);
    return DSB200_OK;
}

// ---------------------------------------------------------------- volume checkpoint
namespace {

// one file per array, MemoryBlockPersister's layout (ORUtils/MemoryBlockPersister.h:88-140): int32 element count, raw elements
int write_block(const std::string &path, const void *dev, size_t elemSize, size_t count, void *pinned, size_t pinnedBytes)
{
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) return fail(DSB200_ERR_INVALID, "checkpoint: could not open " + path + " for writing");
    const int32_t n = (int32_t)count;
    bool ok = fwrite(&n, 4, 1, f) == 1;
    const char *src = (const char *)dev;
    for (size_t off = 0, total = elemSize * count; ok && off < total; off += pinnedBytes) {
        const size_t chunk = total - off < pinnedBytes ? total - off : pinnedBytes;
        if (cudaMemcpy(pinned, src + off, chunk, cudaMemcpyDeviceToHost) != cudaSuccess) { fclose(f); return cuda_fail(cudaGetLastError(), "checkpoint copy"); }
        ok = fwrite(pinned, 1, chunk, f) == chunk;
    }
    fclose(f);
    return ok ? DSB200_OK : fail(DSB200_ERR_INVALID, "checkpoint: short write to " + path);
}

int read_block(const std::string &path, void *dev, size_t elemSize, size_t count, void *pinned, size_t pinnedBytes)
{
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) return fail(DSB200_ERR_INVALID, "checkpoint: could not open " + path + " for reading");
    int32_t n = 0;
    bool ok = fread(&n, 4, 1, f) == 1;
    if (ok && (size_t)n != count) { fclose(f); return fail(DSB200_ERR_INVALID, "checkpoint: " + path + " holds a block of a different size (capacities differ)"); }
    char *dst = (char *)dev;
    for (size_t off = 0, total = elemSize * count; ok && off < total; off += pinnedBytes) {
        const size_t chunk = total - off < pinnedBytes ? total - off : pinnedBytes;
        ok = fread(pinned, 1, chunk, f) == chunk;
        if (ok && cudaMemcpy(dst + off, pinned, chunk, cudaMemcpyHostToDevice) != cudaSuccess) { fclose(f); return cuda_fail(cudaGetLastError(), "checkpoint copy"); }
    }
    fclose(f);
    return ok ? DSB200_OK : fail(DSB200_ERR_INVALID, "checkpoint: short read from " + path);
}

} // namespace

int dsb200_scene_save(const dsb200_scene *scene, const dsb200_render_state *rs, const char *directory)
{
    if (!scene || !directory || !scene->voxels || !scene->hash) return fail(DSB200_ERR_INVALID, "scene_save: null argument");
    CU(cudaDeviceSynchronize());
    const size_t pinnedBytes = (size_t)32 << 20;
    void *pinned = nullptr;
    CU(cudaMallocHost(&pinned, pinnedBytes));
    const std::string d = std::string(directory) + "/";
    const size_t total = (size_t)scene->bucket_num + scene->excess_size;
    {   // a deep look-ahead's pending entries belong to a frame that has not happened: never into a checkpoint
        int *cnt = static_cast<int *>(pinned), *dcnt = nullptr;
        *cnt = 0;
        if (cudaMalloc(&dcnt, 4) != cudaSuccess) { cudaFreeHost(pinned); return cuda_fail(cudaGetLastError(), "scene_save"); }
        cudaMemset(dcnt, 0, 4);
        launch_count_pending(nullptr, scene->hash, (int)total, dcnt);
        cudaMemcpy(cnt, dcnt, 4, cudaMemcpyDeviceToHost);
        cudaFree(dcnt);
        if (*cnt != 0) {
            cudaFreeHost(pinned);
            return fail(DSB200_ERR_INVALID, "scene_save: the hash table holds pending entries of a look-ahead; call dsb200_recon_abandon_prefetch "
                                            "(or the consuming allocation) first");
        }
    }
    int rc = write_block(d + "voxel.dat", scene->voxels, sizeof(dsb200_voxel), (size_t)scene->num_blocks * DSB200_BLOCK_SIZE3, pinned, pinnedBytes);
    if (!rc) rc = write_block(d + "alloc.dat", scene->alloc_list, 4, scene->num_blocks, pinned, pinnedBytes);
    if (!rc) rc = write_block(d + "hash.dat", scene->hash, sizeof(dsb200_hash_entry), total, pinned, pinnedBytes);
    if (!rc) rc = write_block(d + "excess.dat", scene->excess_list, 4, scene->excess_size, pinned, pinnedBytes);
    if (!rc && rs) {
        rc = write_block(d + "visible.dat", rs->visible_ids, 4, (size_t)(rs->no_visible > 0 ? rs->no_visible : 0), pinned, pinnedBytes);
        if (!rc) rc = write_block(d + "vistype.dat", rs->visible_type, 1, total, pinned, pinnedBytes);
    }
    cudaFreeHost(pinned);
    if (rc) return rc;
    FILE *f = fopen((d + "counters.txt").c_str(), "w");
    if (!f) return fail(DSB200_ERR_INVALID, "checkpoint: could not write counters.txt");
    fprintf(f, "%d %d %d %d %d %d\n%.9g %.9g %d %.9g %.9g %d\n", scene->num_blocks, scene->bucket_num, scene->excess_size,
            scene->last_free_block, scene->last_free_excess, rs ? rs->no_visible : -1, scene->params.voxel_size, scene->params.mu,
            scene->params.max_w, scene->params.vf_min, scene->params.vf_max, scene->params.stop_integrating_at_max_w);
    fclose(f);
    return DSB200_OK;
}

int dsb200_scene_load(dsb200_recon *e, dsb200_scene *scene, dsb200_render_state *rs, const char *directory)
{
    if (!scene || !directory || !scene->voxels || !scene->hash) return fail(DSB200_ERR_INVALID, "scene_load: null argument");
    const std::string d = std::string(directory) + "/";
    FILE *f = fopen((d + "counters.txt").c_str(), "r");
    if (!f) return fail(DSB200_ERR_INVALID, "checkpoint: could not read counters.txt");
    int nb, bn, ex, lfb, lfe, nv, maxW, stop; float vs, mu, vmin, vmax;
    const int got = fscanf(f, "%d %d %d %d %d %d %g %g %d %g %g %d", &nb, &bn, &ex, &lfb, &lfe, &nv, &vs, &mu, &maxW, &vmin, &vmax, &stop);
    fclose(f);
    if (got != 12) return fail(DSB200_ERR_INVALID, "checkpoint: malformed counters.txt");
    if (nb != scene->num_blocks || bn != scene->bucket_num || ex != scene->excess_size)
        return fail(DSB200_ERR_INVALID, "checkpoint: capacities differ from the scene's");
    if (rs && nv < 0) return fail(DSB200_ERR_INVALID, "checkpoint: saved without a render state");
    if (rs && nv > rs->visible_capacity) return fail(DSB200_ERR_INVALID, "checkpoint: visible list larger than the render state's capacity");
    CU(cudaDeviceSynchronize());
    const size_t pinnedBytes = (size_t)32 << 20;
    void *pinned = nullptr;
    CU(cudaMallocHost(&pinned, pinnedBytes));
    const size_t total = (size_t)bn + ex;
    int rc = read_block(d + "voxel.dat", scene->voxels, sizeof(dsb200_voxel), (size_t)nb * DSB200_BLOCK_SIZE3, pinned, pinnedBytes);
    if (!rc) rc = read_block(d + "alloc.dat", scene->alloc_list, 4, nb, pinned, pinnedBytes);
    if (!rc) rc = read_block(d + "hash.dat", scene->hash, sizeof(dsb200_hash_entry), total, pinned, pinnedBytes);
    if (!rc) rc = read_block(d + "excess.dat", scene->excess_list, 4, ex, pinned, pinnedBytes);
    if (!rc && rs) {
        rc = read_block(d + "visible.dat", rs->visible_ids, 4, (size_t)nv, pinned, pinnedBytes);
        if (!rc) rc = read_block(d + "vistype.dat", rs->visible_type, 1, total, pinned, pinnedBytes);
    }
    cudaFreeHost(pinned);
    if (rc) return rc;
    scene->last_free_block = lfb; scene->last_free_excess = lfe;
    scene->params.voxel_size = vs; scene->params.mu = mu; scene->params.max_w = maxW; scene->params.vf_min = vmin; scene->params.vf_max = vmax;
    scene->params.stop_integrating_at_max_w = stop;
    if (rs) rs->no_visible = nv;
    if (e) {
        // the engine's derived state describes another scene now: counters come from the host mirrors again, the visibility
        // bitmap and the sdf shadow are rebuilt on first use, the pruning history starts empty, a pending look-ahead is dropped
        if (e->prefetchPending) { CU(cudaStreamSynchronize(e->prefetchStream)); e->prefetchPending = false; }
        e->deviceCountersValid = false; e->visBitsOwner = nullptr; e->shadowFor = nullptr; e->lastKnownVisible = nv > 0 ? nv : 0;
        e->decayQ.clear(); e->slideQ.clear(); e->decayDefusionQ.clear(); e->slideDefusionQ.clear(); e->snapshots.clear(); e->unsynced.clear();
        CU(cudaMemsetAsync(e->reqKey, 0, sizeof(unsigned) * (size_t)e->total, e->stream));
        CU(cudaMemsetAsync(e->fs.reqBits, 0, (size_t)e->fs.bitTiles * 8192, e->stream));
        CU(cudaStreamSynchronize(e->stream));
    }
    return DSB200_OK;
}

// ---------------------------------------------------------------- look-ahead
int dsb200_recon_prefetch_allocation(dsb200_recon *e, dsb200_scene *scene, const dsb200_view *view, const dsb200_tracking *trk,
                                     dsb200_render_state *rs, void *waitEvent)
{
    if (!e || !scene || !view || !trk || !rs || !view->depth) return fail(DSB200_ERR_INVALID, "prefetch_allocation: null argument");
    if (scene->num_blocks != e->numBlocks || scene->bucket_num != e->bucketNum || scene->excess_size != e->excessSize)
        return fail(DSB200_ERR_INVALID, "scene capacities differ from the capacities the engine was created with");
    if (e->prefetchPending) return fail(DSB200_ERR_INVALID, "prefetch_allocation: a prefetched allocation is already pending");
    if (!(e->async && e->deviceCountersValid))
        return fail(DSB200_ERR_INVALID, "prefetch_allocation: asynchronous mode with at least one frame submitted is required");
    if (e->profile) return DSB200_OK;                               // per-stage timing keeps every kernel on the main stream
    AllocParams AP; make_alloc_params(AP, e, scene, view, trk);
    if ((long long)AP.W * AP.H >= (1ll << 24)) return fail(DSB200_ERR_INVALID, "depth image too large for the 24-bit pixel order key");
    if (4.0f * AP.mu * AP.oneOverBlock + 2.0f > 250.0f) return fail(DSB200_ERR_INVALID, "mu / block size ratio too large for the 8-bit step order key");
    // Wait only for the last kernel that wrote the scene when that was the integration of the frame just submitted (its
    // raycast then overlaps the look-ahead); after any other scene call, wait for everything submitted so far.
    const bool behindIntegration = e->epochSceneWritten == e->modEpoch;
    if (!behindIntegration) CU(cudaEventRecord(e->evSceneWritten, e->stream));
    CU(cudaStreamWaitEvent(e->prefetchStream, e->evSceneWritten, 0));
    if (waitEvent) CU(cudaStreamWaitEvent(e->prefetchStream, (cudaEvent_t)waitEvent, 0));
    if (rs->visible_type != e->visBitsOwner) {     // a render state this engine has not been tracking: derive its bitmap
        launch_rebuild_visbits(e->prefetchStream, e->fs, rs->visible_type, e->total);
        e->visBitsOwner = rs->visible_type;
    }
    launch_begin_frame(e->prefetchStream, e->fs, 0, 0, 0, 0, rs->visible_ids, rs->visible_type, rs->visible_capacity);
    launch_alloc_requests(e->prefetchStream, AP, e->fs, view->depth, scene->hash, rs->visible_type, e->reqKey);
    VisParams VP; fill_mat(VP.M, trk->M_d); memcpy(VP.proj, view->intr_d, 16);
    VP.voxelSize = scene->params.voxel_size; VP.W = AP.W; VP.H = AP.H;
    VP.useSwapping = scene->use_swapping ? 1 : 0;       // the consuming call must be a full allocation (onlyUpdateVisibleList = false)
    launch_recheck_visibility(e->prefetchStream, VP, e->fs, scene->hash, rs->visible_ids, rs->visible_type);
    // deep look-ahead (recon_alloc.cu): ordered allocation with PENDING entries + the visible list into the second buffer.  Not with
    // swapping on (the visible-list pass then hands blocks back to swapped-out entries, which the running raycast would see).
    // Nor when the look-ahead starts behind everything submitted so far (a pruning pass followed the frame): nothing is left to overlap
    // with, and the pending / commit detour would only add work.
    const bool deep = tunables().deep_lookahead != 0 && !VP.useSwapping && behindIntegration && deep_buffers(e, rs->visible_capacity);
    if (deep) {
        const DeepList dl = deep_list(e);
        launch_allocate(e->prefetchStream, AP, e->fs, view->depth, scene->hash, scene->alloc_list, scene->excess_list, rs->visible_type,
                        e->reqKey, e->total, 1, &dl);
        launch_visible_list(e->prefetchStream, VP, e->fs, scene->hash, scene->alloc_list, rs->visible_type, e->nextVisible,
                            rs->visible_capacity, nullptr, nullptr, 0, e->total, 1);
    }
    CU(cudaGetLastError());
    CU(cudaEventRecord(e->evPrefetchDone, e->prefetchStream));
    dsb200_recon::PrefetchKey &k = e->prefetchKey;
    k.depth = view->depth; k.w = AP.W; k.h = AP.H; k.visType = rs->visible_type;
    k.visTypeRW = rs->visible_type; k.visibleIds = rs->visible_ids; k.hash = scene->hash; k.useSwapping = VP.useSwapping;
    memcpy(k.M, trk->M_d, 64); memcpy(k.invM, trk->invM_d, 64); memcpy(k.intr, view->intr_d, 16);
    k.deep = deep; k.hashRW = scene->hash;
    e->prefetchPending = true;
    return DSB200_OK;
}

int dsb200_recon_abandon_prefetch(dsb200_recon *e)
{
    if (!e) return fail(DSB200_ERR_INVALID, "abandon_prefetch: null engine");
    { int rc = abandon_prefetch(e); if (rc) return rc; }
    if (!e->async) CU(cudaStreamSynchronize(e->stream));
    return DSB200_OK;
}

int dsb200_recon_join_prefetch(dsb200_recon *e)
{
    if (!e) return fail(DSB200_ERR_INVALID, "join_prefetch: null engine");
    if (e->prefetchPending) CU(cudaStreamWaitEvent(e->stream, e->evPrefetchDone, 0));
    return DSB200_OK;
}

// ---------------------------------------------------------------- fused frame
int dsb200_process_frame(dsb200_recon *e, dsb200_vis *v, dsb200_scene *scene, const dsb200_view *view, dsb200_tracking *trk,
                         dsb200_render_state *rs, int stages)
{
    int rc = check_scene(e, scene, (stages & DSB200_FRAME_ALLOCATE) != 0); if (rc) return rc;
    if ((stages & (DSB200_FRAME_EXPECTED_DEPTHS | DSB200_FRAME_ICP_MAPS | DSB200_FRAME_RAYCAST)) && (!v || v->stream != e->stream))
        return fail(DSB200_ERR_INVALID, "process_frame: the two engines must share one stream");
    const int *countPtr = &e->fs.counters->noVisible;
#define DSB_MARK(i) do { if (e->profile) { cudaEventRecord(e->ev[i], e->stream); e->evRecorded[i] = true; } } while (0)
    if (e->profile) for (int i = 0; i <= DSB200_PROFILE_STAGES; i++) e->evRecorded[i] = false;
    DSB_MARK(0);
    if (stages & DSB200_FRAME_ALLOCATE) {
        rc = enqueue_allocate(e, scene, view, trk, rs, 0, 0); if (rc) return rc;
    } else if (!(e->async && e->deviceCountersValid)) {
        e->hostCounters->noVisible = rs->no_visible;
        CU(cudaMemcpyAsync(&e->fs.counters->noVisible, &e->hostCounters->noVisible, 4, cudaMemcpyHostToDevice, e->stream));
    }
    DSB_MARK(1);
    // Integration (writes voxels) and the expected-depth image (reads hash + visible list, writes the range image) are
    // independent: outside profiling mode the latter runs on a side stream, forked after the visible list and joined
    // before the raycast.
    // raycast tiles are handed out longest expected-depth span first; the order is derived from this call's range image
    const int nTiles = div_up(rs->width, 8) * div_up(rs->height, 4);
    int *tileOrder = (v && (stages & DSB200_FRAME_EXPECTED_DEPTHS) && nTiles <= v->tileOrderCap) ? v->tileOrder : nullptr;
    const bool fork = !e->profile && (stages & DSB200_FRAME_INTEGRATE) && (stages & DSB200_FRAME_EXPECTED_DEPTHS) && e->sideStream;
    if (fork) {
        CU(cudaEventRecord(e->evFork, e->stream));
        CU(cudaStreamWaitEvent(e->sideStream, e->evFork, 0));
        ProjParams P; make_proj_params(P, scene, trk->M_d, view->intr_d, rs->width, rs->height);
        launch_expected_depths(e->sideStream, P, scene->hash, rs->visible_ids, countPtr, rs->range_image, tileOrder, e->fs.counters);
        CU(cudaEventRecord(e->evJoin, e->sideStream));
    }
    if (stages & DSB200_FRAME_INTEGRATE) {
        FuseParams FP; make_fuse_params(FP, e, scene, view, trk);
        launch_fuse(e->stream, FP, false, scene->voxels, scene->hash, rs->visible_ids, countPtr, view->depth, view->rgb, shadow_for(e, scene),
                    weight_scratch(e, view));
    }
    if (e->async && !e->profile) {     // nothing after this point writes the scene: a look-ahead may start here
        CU(cudaEventRecord(e->evSceneWritten, e->stream));
        e->epochSceneWritten = e->modEpoch;
    }
    DSB_MARK(2);
    if (fork) CU(cudaStreamWaitEvent(e->stream, e->evJoin, 0));
    else if (stages & DSB200_FRAME_EXPECTED_DEPTHS) {
        ProjParams P; make_proj_params(P, scene, trk->M_d, view->intr_d, rs->width, rs->height);
        launch_expected_depths(e->stream, P, scene->hash, rs->visible_ids, countPtr, rs->range_image, tileOrder, e->fs.counters);
    }
    DSB_MARK(3);
    if (stages & DSB200_FRAME_ICP_MAPS) {
        if (!trk->icp_points || !trk->icp_normals) return fail(DSB200_ERR_INVALID, "process_frame: tracking state has no point cloud buffers");
        RayParams P; make_ray_params(P, scene, trk->M_d, trk->invM_d, view->intr_d, rs->width, rs->height);
        launch_raycast(e->stream, P, scene->voxels, scene->hash, rs->range_image, rs->raycast_result, v->fs.tickets + 2, shadow_for(e, scene), &v->rayQueue, tileOrder);
        DSB_MARK(4);
        launch_icp_maps(e->stream, P, rs->raycast_result, trk->icp_points, trk->icp_normals, rs->raycast_image);
    } else if (stages & DSB200_FRAME_RAYCAST) {
        RayParams P; make_ray_params(P, scene, trk->M_d, trk->invM_d, view->intr_d, rs->width, rs->height);
        launch_raycast(e->stream, P, scene->voxels, scene->hash, rs->range_image, rs->raycast_result, v->fs.tickets + 2, shadow_for(e, scene), &v->rayQueue, tileOrder);
        DSB_MARK(4);
    } else DSB_MARK(4);
    DSB_MARK(5);
    return finish_call(e, scene, rs, (stages & DSB200_FRAME_ALLOCATE) != 0);
}

int dsb200_recon_set_sdf_shadow(dsb200_recon *e, int on)
{
    if (!e) return fail(DSB200_ERR_INVALID, "set_sdf_shadow: null engine");
    e->shadowEnabled = on != 0;
    e->shadowFor = nullptr;       // re-derived from the voxels at the next use
    return DSB200_OK;
}
int dsb200_recon_notify_scene_modified(dsb200_recon *e)
{
    if (!e) return fail(DSB200_ERR_INVALID, "notify_scene_modified: null engine");
    { int rc = abandon_prefetch(e); if (rc) return rc; }   // a pending look-ahead is dropped AND undone (keys, bitmaps, flags)
    CU(cudaStreamSynchronize(e->stream));
    e->deviceCountersValid = false;        // the next call seeds the device counters from the scene's host mirrors
    e->shadowFor = nullptr;                // the sdf shadow is re-derived from the voxels at the next use
    e->modEpoch++;
    return DSB200_OK;
}

// the swapping engine's notification: it changed voxel blocks, entry pointers and scene->last_free_block and nothing else, so
// only that one device counter is re-seeded -- an asynchronous engine's visible count (whose host mirror the swapping engine
// never refreshed: it has no render state) and excess-list top stay the device's
int dsb200_recon_notify_volume_swapped(dsb200_recon *e, const dsb200_scene *scene)
{
    if (!e || !scene) return fail(DSB200_ERR_INVALID, "notify_volume_swapped: null argument");
    { int rc = abandon_prefetch(e); if (rc) return rc; }
    if (e->async && e->deviceCountersValid) {
        e->hostCounters->lastFreeBlock = scene->last_free_block;
        CU(cudaMemcpyAsync(&e->fs.counters->lastFreeBlock, &e->hostCounters->lastFreeBlock, 4, cudaMemcpyHostToDevice, e->stream));
    }
    CU(cudaStreamSynchronize(e->stream));
    e->shadowFor = nullptr;
    e->modEpoch++;
    return DSB200_OK;
}

int dsb200_recon_invalidate_sdf_shadow(dsb200_recon *e)
{
    if (!e) return fail(DSB200_ERR_INVALID, "invalidate_sdf_shadow: null engine");
    e->shadowFor = nullptr;
    return DSB200_OK;
}

int dsb200_debug_force_plain_division(dsb200_recon *e, int on) { e->forceSlowDiv = on != 0; return DSB200_OK; }
int dsb200_debug_set_tunable(const char *name, int value)
{
    if (!name || !set_tunable(name, value)) return fail(DSB200_ERR_INVALID, "debug_set_tunable: unknown name");
    return DSB200_OK;
}

int dsb200_selftest_division(unsigned long long n, unsigned long long seed, unsigned long long *mismatches)
{
    if (!mismatches) return fail(DSB200_ERR_INVALID, "selftest_division: null output");
    if (dsb200_device_count() == 0) return fail(DSB200_ERR_NO_DEVICE, "no CUDA device: libdsb200 has no CPU fallback");
    unsigned long long *d = nullptr;
    CU(cudaMalloc(&d, 8));
    CU(cudaMemset(d, 0, 8));
    launch_selftest_division(nullptr, n, seed, d);
    CU(cudaGetLastError());
    CU(cudaMemcpy(mismatches, d, 8, cudaMemcpyDeviceToHost));
    CU(cudaFree(d));
    return DSB200_OK;
}

int dsb200_profile_enable(dsb200_recon *e, int on)
{
    if (on && !e->ev[0]) {
        for (int i = 0; i <= DSB200_PROFILE_STAGES; i++) CU(cudaEventCreate(&e->ev[i]));
        for (int i = 0; i <= DSB200_PROFILE_ALLOC_STAGES; i++) CU(cudaEventCreate(&e->evAlloc[i]));
    }
    e->profile = on != 0;
    return DSB200_OK;
}
int dsb200_profile_read_allocate(dsb200_recon *e, float *ms)
{
    if (!e->profile || !e->evAllocRecorded) return fail(DSB200_ERR_INVALID, "profile_read_allocate: nothing recorded");
    CU(cudaEventSynchronize(e->evAlloc[DSB200_PROFILE_ALLOC_STAGES]));
    for (int i = 0; i < DSB200_PROFILE_ALLOC_STAGES; i++) CU(cudaEventElapsedTime(&ms[i], e->evAlloc[i], e->evAlloc[i + 1]));
    return DSB200_OK;
}
int dsb200_profile_read(dsb200_recon *e, float *ms)
{
    if (!e->profile || !e->evRecorded[DSB200_PROFILE_STAGES]) return fail(DSB200_ERR_INVALID, "profile_read: nothing recorded");
    CU(cudaEventSynchronize(e->ev[DSB200_PROFILE_STAGES]));
    for (int i = 0; i < DSB200_PROFILE_STAGES; i++) CU(cudaEventElapsedTime(&ms[i], e->ev[i], e->ev[i + 1]));
    return DSB200_OK;
}

// ---------------------------------------------------------------- host helpers
int dsb200_scene_alloc(dsb200_scene *s, int32_t num_blocks, int32_t bucket_num, int32_t excess_size)
{
    if (!s) return fail(DSB200_ERR_INVALID, "scene_alloc: null");
    if (dsb200_device_count() == 0) return fail(DSB200_ERR_NO_DEVICE, "no CUDA device: libdsb200 has no CPU fallback");
    s->num_blocks = num_blocks; s->bucket_num = bucket_num; s->excess_size = excess_size;
    CU(cudaMalloc(&s->voxels, (size_t)num_blocks * DSB200_BLOCK_SIZE3 * sizeof(dsb200_voxel)));
    CU(cudaMalloc(&s->hash, (size_t)(bucket_num + excess_size) * sizeof(dsb200_hash_entry)));
    CU(cudaMalloc(&s->alloc_list, (size_t)num_blocks * 4));
    CU(cudaMalloc(&s->excess_list, (size_t)(excess_size > 0 ? excess_size : 1) * 4));
    s->last_free_block = num_blocks - 1; s->last_free_excess = excess_size - 1;
    return DSB200_OK;
}
int dsb200_scene_free(dsb200_scene *s)
{
    if (!s) return DSB200_OK;
    cudaFree(s->voxels); cudaFree(s->hash); cudaFree(s->alloc_list); cudaFree(s->excess_list);
    s->voxels = nullptr; s->hash = nullptr; s->alloc_list = nullptr; s->excess_list = nullptr;
    return DSB200_OK;
}
int dsb200_render_state_alloc(dsb200_render_state *rs, int32_t total_entries, int32_t visible_capacity, int32_t width, int32_t height)
{
    if (!rs) return fail(DSB200_ERR_INVALID, "render_state_alloc: null");
    rs->visible_capacity = visible_capacity; rs->no_visible = 0; rs->width = width; rs->height = height;
    const size_t px = (size_t)width * height;
    CU(cudaMalloc(&rs->visible_ids, (size_t)visible_capacity * 4));
    CU(cudaMalloc(&rs->visible_type, (size_t)((total_entries + 3) & ~3)));
    CU(cudaMemset(rs->visible_type, 0, (size_t)((total_entries + 3) & ~3)));
    CU(cudaMalloc(&rs->range_image, px * 8));
    CU(cudaMalloc(&rs->raycast_result, px * 16));
    CU(cudaMalloc(&rs->raycast_image, px * 4));
    CU(cudaMemset(rs->raycast_result, 0, px * 16));
    CU(cudaMemset(rs->raycast_image, 0, px * 4));
    // ITMRenderState ctor pre-fills the range image with (vf_min, vf_max); callers overwrite it
    CU(cudaMemset(rs->range_image, 0, px * 8));
    return DSB200_OK;
}
int dsb200_render_state_free(dsb200_render_state *rs)
{
    if (!rs) return DSB200_OK;
    cudaFree(rs->visible_ids); cudaFree(rs->visible_type); cudaFree(rs->range_image); cudaFree(rs->raycast_result);
    cudaFree(rs->raycast_image);
    memset(rs, 0, sizeof(*rs));
    return DSB200_OK;
}
int dsb200_device_alloc(void **p, size_t bytes) { CU(cudaMalloc(p, bytes)); return DSB200_OK; }
int dsb200_device_free(void *p) { CU(cudaFree(p)); return DSB200_OK; }
int dsb200_host_alloc_pinned(void **p, size_t bytes) { CU(cudaMallocHost(p, bytes)); return DSB200_OK; }
int dsb200_host_free_pinned(void *p) { CU(cudaFreeHost(p)); return DSB200_OK; }
int dsb200_memcpy_h2d(void *dst, const void *src, size_t bytes, void *stream)
{
    CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return DSB200_OK;
}
int dsb200_memcpy_d2h(void *dst, const void *src, size_t bytes, void *stream)
{
    CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    return DSB200_OK;
}
int dsb200_stream_synchronize(void *stream) { CU(cudaStreamSynchronize((cudaStream_t)stream)); return DSB200_OK; }

} // extern "C"
