// kernels.cuh -- parameter blocks and host-side launchers of every kernel in libdsb200.
#pragma once
#include "common.cuh"

namespace dsb {

// Engine scratch that several kernels of one frame share (device pointers, passed by value).
struct FrameState {
    Counters *counters;
    unsigned long long *scanState;   // [2 * scanTiles]: look-back words of k_allocate, then k_visible_list
    int *tickets;                    // [4] dynamic tile tickets
    int scanTiles;                   // ceil(totalEntries / 1024)
    // slot bitmaps (reconstruction engine only), padded to bitTiles * 1024 words
    unsigned *reqBits;               // slots holding an allocation request this frame (all zero between frames)
    unsigned *reqExBits;             // ... of which excess-list requests
    unsigned *visBits;               // superset of the slots whose visibility flag may be non-zero
    int bitTiles;                    // ceil(totalEntries / 32768)
};

struct AllocParams {
    Mat4 invM;
    float invProj[4];                // (1/fx, 1/fy, cx, cy)
    float mu, oneOverBlock, vfMin, vfMax;
    int W, H, bucketNum, bucketMask;
    int shardRank, shardCount, shardBrickLog2, shardHalo;   // shardCount <= 1: unsharded
};

struct VisParams {                   // block visibility test
    Mat4 M;
    float proj[4];
    float voxelSize;
    int W, H, useSwapping;
};

struct FuseParams {                  // integrate / de-integrate
    Mat4 M_d, M_rgb;
    float proj_d[4], proj_rgb[4];
    float voxelSize, mu;
    int maxW, W, H, Wrgb, Hrgb;
    int stopMaxW, approximate;       // template flags of the reference kernel (CUDA .cu:407-420)
    int depthWeighting, maxNewW, maxDistance;
    // exact-division support (exactdiv.cuh): host-computed correctly rounded reciprocals and the operand-range verdict
    float rcpMu, rcp255, rcp32767;
    int fastDiv;
};

struct RayParams {                   // raycast (castRay) + ICP maps
    Mat4 invM, M;
    float invProj[4];
    float oneOverVoxel, mu, voxelSize;
    float light[3];
    int W, H, bucketNum, bucketMask;
};

struct ProjParams {                  // CreateExpectedDepths
    Mat4 M;
    float proj[4];
    float voxelSize;
    int W, H, maxRenderBlocks;
};

// Launch-shape knobs (defaults measured on B200, profiles/README.md).  Initialised from the environment variable of the same name
// in upper case with the DSB200_ prefix (DSB200_RAY_STEPS ...); dsb200_debug_set_tunable changes them at run time (tools/tune.py).
// None of them changes a result: every setting is covered by the same bit-exact parity tests.
struct Tunables {
    int ray_steps;        // loop bodies per ray in the first raycast pass before the ray is suspended (0: one unbounded pass)
    int ray_group;        // lanes per suspended ray in the finishing pass: 2, 4, 8, 16 or 32
    int ray_ctas;         // CTAs per SM of the first raycast pass
    int ray_finish_ctas;  // CTAs per SM of the finishing pass
    int ray_bottom_up;    // hand the image tiles out from the last row upwards
    int fuse_kernel;      // integration kernel: 1 warp per block (round 1), 2 + straight-line depth part, 3 + TMA bulk block loads (default)
    int fuse_ctas;        // CTAs per SM of the flat integration kernel
    int fuse_minb;        // ... and its __launch_bounds__ minimum CTAs per SM (variant 2: 4, 6 or 8; variant 3: 4 or 5)
    int deep_lookahead;   // dsb200_recon_prefetch_allocation also runs the ordered allocation and the visible-list build (1 = on)
    int ray_lpt;          // fused frame: raycast tiles dealt by class of expected-depth span: >= this many 8-voxel steps, half, a quarter, rest (0 = off)
};
Tunables &tunables();
bool set_tunable(const char *name, int value);

// Rays the bounded first raycast pass suspends, and the finishing pass completes (vis.cu).  Engine scratch of dsb200_vis.
struct RayQueue {
    int *ctl;          // [0] rays queued, [1] hand-out ticket of the finishing pass, [2] warps of the finishing pass that are done
    int *pixel;        // x + y * W
    float4 *posLen;    // rx, ry, rz, totalLength (voxel units, at the top of the march loop)
    float4 *dirMax;    // dx, dy, dz, totalLengthMax
    int capacity;      // rays (>= W * H of any image raycast through it)
};

// recon_alloc.cu
void launch_begin_frame(cudaStream_t st, const FrameState &fs, int hostLastFreeBlock, int hostLastFreeExcess,
                        int hostNoVisible, int useHost, const int *visibleIds, uint8_t *visType, int prevUpperBound);
void launch_alloc_requests(cudaStream_t st, const AllocParams &P, const FrameState &fs, const float *depth,
                           const dsb200_hash_entry *hash, uint8_t *visType, unsigned *reqKey);
void launch_abandon_requests(cudaStream_t st, const FrameState &fs, const dsb200_hash_entry *hash, const int *visibleIds,
                             uint8_t *visType, unsigned *reqKey, int total);
void launch_recheck_visibility(cudaStream_t st, const VisParams &P, const FrameState &fs, const dsb200_hash_entry *hash,
                               const int *prevVisibleIds, uint8_t *visType);
void launch_rebuild_visbits(cudaStream_t st, const FrameState &fs, const uint8_t *visType, int total);
// deep look-ahead (recon_alloc.cu): what k_allocate wrote, by allocation rank
constexpr int kPendingBase = -3;     // a pending entry's ptr = kPendingBase - block  (<= -3; -1 = swapped out, -2 = unallocated)
struct DeepList {
    int *slot;         // the entry written for this rank (ordered slot or excess child); -1: nothing written
    int *parent;       // excess allocations: the chain end whose offset was set; else -1
    int *old;          // 6 ints per rank: the entry's previous 5 words + the parent's previous offset
    int capacity;
};
void launch_allocate(cudaStream_t st, const AllocParams &P, const FrameState &fs, const float *depth,
                     dsb200_hash_entry *hash, const int *allocList, const int *excessList, uint8_t *visType,
                     unsigned *reqKey, int total, int doAllocate, const DeepList *deep = nullptr);
void launch_visible_list(cudaStream_t st, const VisParams &P, const FrameState &fs, dsb200_hash_entry *hash,
                         const int *allocList, uint8_t *visType, int *visibleIds, int visibleCapacity, int *snapshotIds,
                         int *snapshotCount, int snapshotCapacity, int total, int defer = 0);
void launch_commit_allocation(cudaStream_t st, const FrameState &fs, dsb200_hash_entry *hash, const DeepList &deep, const int *nextIds,
                              int *visibleIds, int visibleCapacity, int *snapshotIds, int *snapshotCount, int snapshotCapacity);
void launch_abandon_allocation(cudaStream_t st, const FrameState &fs, dsb200_hash_entry *hash, const DeepList &deep);
void launch_count_pending(cudaStream_t st, const dsb200_hash_entry *hash, int total, int *count);

// recon_integrate.cu
void launch_fuse(cudaStream_t st, const FuseParams &P, bool deintegrate, dsb200_voxel *voxels,
                 const dsb200_hash_entry *hash, const int *ids, const int *count, const float *depth,
                 const uint8_t *rgb, short *sdfShadowOrNull, unsigned char *weightScratchOrNull = nullptr /* W * H bytes */);
void launch_reset_scene(cudaStream_t st, dsb200_voxel *voxels, size_t numVoxels, dsb200_hash_entry *hash, int total,
                        int *allocList, int numBlocks, int *excessList, int excessSize, uint8_t *visTypeOrNull,
                        unsigned *reqKey, Counters *counters, short *sdfShadowOrNull);
void launch_rebuild_shadow(cudaStream_t st, const dsb200_voxel *voxels, size_t numVoxels, short *sdfShadow);

void launch_selftest_division(cudaStream_t st, unsigned long long n, unsigned long long seed, unsigned long long *mismatches);

// recon_prune.cu
void launch_prune(cudaStream_t st, bool slide, dsb200_voxel *voxels, dsb200_hash_entry *hash, int bucketNum,
                  const int *ids, const int *count, int countUpperBound, int minAge, int maxWeight, int currentFrame,
                  int *allocList, uint8_t *visType, unsigned *visBitsOrNull, unsigned *slotMark, int *pendingList,
                  Counters *counters, const FrameState &fs, short *sdfShadowOrNull);

// vis.cu
void launch_find_visible(cudaStream_t st, const VisParams &P, const FrameState &fs, const dsb200_hash_entry *hash,
                         int *visibleIds, int visibleCapacity, int total);
void launch_count_visible(cudaStream_t st, const dsb200_hash_entry *hash, const int *ids, const int *count, int minId,
                          int maxId, int *out);
void launch_expected_depths(cudaStream_t st, const ProjParams &P, const dsb200_hash_entry *hash, const int *ids,
                            const int *count, float *rangeImage, int *tileOrderOrNull /* 4 counters + 4 lists of ceil(W/8) * ceil(H/4) ints: raycast tiles by depth-span class */,
                            Counters *counters);
void launch_raycast(cudaStream_t st, const RayParams &P, const dsb200_voxel *voxels, const dsb200_hash_entry *hash,
                    const float *rangeImage, float *raycastResult, int *ticketPair /* two zeroed ints, re-armed by the kernel */,
                    const short *sdfShadowOrNull = nullptr, const RayQueue *queueOrNull = nullptr, const int *tileOrderOrNull = nullptr);
void launch_raycast_sharded(cudaStream_t st, const RayParams &P, int n, int rank, int brickLog2,
                            const dsb200_hash_entry *const *hash, const dsb200_voxel *const *voxels,
                            const float *const *rangeImages, float *mergedRange, float *raycastResult, int *ticketPair,
                            const RayQueue *queueOrNull = nullptr);
void launch_icp_maps(cudaStream_t st, const RayParams &P, const float *raycastResult, float *points, float *normals,
                     uint8_t *rendering);
void launch_composite_nearest(cudaStream_t st, int n, const float *const *resultsDev, float cx, float cy, float cz, int W,
                              int rowBegin, int rowEnd, float *out);
void launch_view_from_raw(cudaStream_t st, const int16_t *raw, const uint8_t *bgr, float scale, float offset, float *depth,
                          uint8_t *rgb, int n);
void launch_shade(cudaStream_t st, const RayParams &P, int renderType, const dsb200_voxel *voxels,
                  const dsb200_hash_entry *hash, const float *raycastResult, uint8_t *outRGBA, float *outDepth);
void launch_forward_render(cudaStream_t st, const RayParams &P, const float *proj, const dsb200_voxel *voxels,
                           const dsb200_hash_entry *hash, const float *rangeImage, const float *raycastResult,
                           const float *depth, float *fwd, int *owner, int *noMissing, uint8_t *rendering);

// view_filter.cu
void launch_filter_depth(cudaStream_t st, const float *in, float *out, int W, int H);
void launch_disparity_to_depth(cudaStream_t st, const short *in, float *out, float paramX, float paramY, float fx, int n);
void launch_normal_and_weight(cudaStream_t st, const float *depth, float *normal4, float *sigmaZ, int W, int H, const float *intr);
void launch_acosf_selftest(cudaStream_t st, const float *x, float *y, int n);
void launch_expf_selftest(cudaStream_t st, const float *x, float *y, int n);
// tracker.cu
void launch_filter_subsample_holes(cudaStream_t st, const float *in, int oldW, int oldH, float *out);
int tracker_gh_max_ctas();
int tracker_gh_slots();
void launch_tracker_gh(cudaStream_t st, int iterationType, const float *approxInvPose, const float *scenePose,
                       const float *viewIntr, const float *sceneIntr, float distThresh, const float *depth, int viewW, int viewH,
                       const float *points, const float *normals, int sceneW, int sceneH, double *partials, unsigned *ticket,
                       double *result, double seq);

} // namespace dsb
