/* dsb200.h -- C ABI of the B200-native volumetric-fusion hot path.
 *
 * One entry point per virtual of the two InfiniTAM engine interfaces that DenseSLAM's
 * InfiniTamDriver drives (paths relative to
 * /root/reference/src/InfiniTAM-Global-Consistency-h/InfiniTAM/):
 *   ITMLib/Engine/ITMSceneReconstructionEngine.h:45-96   (reconstruction engine)
 *   ITMLib/Engine/ITMVisualisationEngine.h:71-124        (visualisation engine)
 * and, for the two neighbours of that path, one per device step of
 *   ITMLib/Engine/ITMViewBuilder.h:20-55                 (depth conversions, bilateral filter)
 *   ITMLib/Engine/ITMDepthTracker.h:26-70                (ComputeGandH; plus a one-call TrackCamera)
 * Additions without a counterpart in the reference are marked as such: the sharded scene, the fused frame, the look-ahead
 * allocation, the volume checkpoint.
 * The adapter subclasses shown in INTEGRATION.md pull raw DEVICE pointers out of the
 * reference objects (ITMScene, ITMRenderState_VH, ITMView, ITMTrackingState) into the POD
 * structs below and forward.  No torch / C++ types cross this boundary.
 *
 * Conventions (same as the reference): 4x4 matrices are column-major float[16]
 * (ORUtils/Matrix.h:7-11); intrinsics are (fx, fy, cx, cy) (Objects/ITMIntrinsics.h:21-43);
 * every call is synchronous on return unless dsb200_recon_set_async() was used; host mirrors
 * (last_free_block, last_free_excess, no_visible) are valid on return.
 * All functions return DSB200_OK (0) or a negative dsb200_status; dsb200_last_error() gives text.
 */
#ifndef DSB200_H
#define DSB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- data contract: byte-identical to the reference structs at default capacities ---- */
#define DSB200_BLOCK_SIZE 8            /* SDF_BLOCK_SIZE   ITMLib/Utils/ITMLibDefines.h:36 */
#define DSB200_BLOCK_SIZE3 512         /* SDF_BLOCK_SIZE3  :37 */
#define DSB200_DEFAULT_LOCAL_BLOCKS 0x40000   /* SDF_LOCAL_BLOCK_NUM :38 */
#define DSB200_DEFAULT_BUCKETS 0x100000       /* SDF_BUCKET_NUM :44 */
#define DSB200_DEFAULT_EXCESS 0x60000         /* SDF_EXCESS_LIST_SIZE :46 */

/* ITMHashEntry, ITMLibDefines.h:57-74 (20 bytes) */
typedef struct dsb200_hash_entry {
    int16_t pos[3];
    int16_t pad_;
    int32_t offset;         /* 0 = chain end, else next = bucket_num + offset - 1 */
    int32_t ptr;            /* >=0 VBA block, -1 swapped out, -2 free */
    int32_t allocated_time; /* never written by the reference allocators: we write 0 */
} dsb200_hash_entry;

/* ITMVoxel_s_rgb, ITMLibDefines.h:128-158 (8 bytes) */
typedef struct dsb200_voxel {
    int16_t sdf;       /* (short)(f*32767), init 32767 */
    uint8_t w_depth;
    uint8_t clr[3];
    uint8_t w_color;
    uint8_t pad_;
} dsb200_voxel;

typedef enum dsb200_status {
    DSB200_OK = 0,
    DSB200_ERR_CUDA = -1,
    DSB200_ERR_INVALID = -2,
    DSB200_ERR_VBA_EXHAUSTED = -3,    /* "run out of space in the Voxel Block Array"  (CUDA .cu:350-353) */
    DSB200_ERR_EXCESS_EXHAUSTED = -4, /* "run out of slots in the hash table excess list" (.cu:354-358) */
    DSB200_ERR_UNSUPPORTED = -5,
    DSB200_ERR_NO_DEVICE = -6
} dsb200_status;

/* ITMSceneParams, Objects/ITMSceneParams.h:18-50 */
typedef struct dsb200_scene_params {
    float voxel_size, mu;
    int32_t max_w;
    float vf_min, vf_max;
    int32_t stop_integrating_at_max_w;
} dsb200_scene_params;

/* Device pointers pulled from ITMScene (Objects/ITMScene.h, ITMLocalVBA.h:28-36, ITMVoxelBlockHash.h:70-84) */
typedef struct dsb200_scene {
    dsb200_voxel *voxels;          /* [num_blocks * 512] */
    dsb200_hash_entry *hash;       /* [bucket_num + excess_size] */
    int32_t *alloc_list;           /* [num_blocks] VBA free list, stack, pops from the high end */
    int32_t *excess_list;          /* [excess_size] */
    int32_t num_blocks, bucket_num /* power of two */, excess_size;
    int32_t last_free_block;       /* host mirror (in/out) */
    int32_t last_free_excess;      /* host mirror (in/out) */
    dsb200_scene_params params;
    int32_t use_swapping;          /* only changes visibility semantics (enlarged frustum), SURVEY A4/Q11 */
} dsb200_scene;

/* Device pointers pulled from ITMRenderState_VH (Objects/ITMRenderState.h:34-57, ITMRenderState_VH.h:26-65) */
typedef struct dsb200_render_state {
    int32_t *visible_ids;          /* [visible_capacity] hash-slot indices, ascending */
    uint8_t *visible_type;         /* [bucket_num + excess_size] 0/1/2/3 */
    int32_t visible_capacity;
    int32_t no_visible;            /* host mirror (in/out) */
    float *range_image;            /* Vector2f[W*H]; meaningful at 1/8 res with stride W */
    float *raycast_result;         /* Vector4f[W*H], voxel units, w in {0,1} */
    uint8_t *raycast_image;        /* Vector4u[W*H] grey rendering written by CreateICPMaps */
    int32_t width, height;
} dsb200_render_state;

/* ITMView + calib (Objects/ITMView.h:20-36, ITMRGBDCalib.h) */
typedef struct dsb200_view {
    const float *depth;            /* float[W*H] metres, <=0 invalid */
    const uint8_t *rgb;            /* Vector4u[Wrgb*Hrgb] */
    int32_t depth_w, depth_h, rgb_w, rgb_h;
    float intr_d[4], intr_rgb[4];
    float rgb_to_depth_inv[16];    /* calib->trafo_rgb_to_depth.calib_inv */
    double timestamp;              /* view->mTimeStamp, key of the de-fusion snapshot map */
} dsb200_view;

/* ITMTrackingState (Objects/ITMTrackingState.h:29-46) */
typedef struct dsb200_tracking {
    float M_d[16];                 /* pose_d->GetM(): world -> camera */
    float invM_d[16];              /* pose_d->GetInvM() (host cofactor inverse, ORUtils/Matrix.h:175-246) */
    int32_t requires_full_rendering;
    float *icp_points;             /* pointCloud->locations Vector4f[W*H], metres, w=1 / -1 */
    float *icp_normals;            /* pointCloud->colours   Vector4f[W*H], w=0 / -1 */
} dsb200_tracking;

/* WeightParams, Engine/ITMSceneReconstructionEngine.h:21-25 */
typedef struct dsb200_weight_params { int32_t depth_weighting, max_new_w, max_distance; } dsb200_weight_params;

/* IITMVisualisationEngine::RenderImageType, Engine/ITMVisualisationEngine.h:21-28 (same values) */
typedef enum dsb200_render_type {
    DSB200_RENDER_SHADED_GREYSCALE = 0,
    DSB200_RENDER_COLOUR_FROM_VOLUME = 1,
    DSB200_RENDER_COLOUR_FROM_NORMAL = 2,
    DSB200_RENDER_DEPTH_MAP = 3,      /* the fork inserts the depth mode before COLOURCODED (:21-28) */
    DSB200_RENDER_COLOURCODED = 4
} dsb200_render_type;

typedef struct dsb200_recon dsb200_recon;   /* engine-owned scratch + history queues */
typedef struct dsb200_vis dsb200_vis;

const char *dsb200_last_error(void);
int dsb200_device_count(void);
/* selects the CUDA device for this thread's subsequent calls (one process per GPU under torchrun) */
int dsb200_set_device(int device);

/* ---- reconstruction engine (ITMSceneReconstructionEngine.h:45-96) ---- */
/* ctor: sizes the engine scratch (entriesAllocType/blockCoords equivalents) for these capacities. */
int dsb200_recon_create(dsb200_recon **out, int32_t num_blocks, int32_t bucket_num, int32_t excess_size,
                        void *cuda_stream /* cudaStream_t or NULL */);
int dsb200_recon_destroy(dsb200_recon *e);
/* async != 0: calls only enqueue work; host mirrors are refreshed by dsb200_recon_sync(). */
int dsb200_recon_set_async(dsb200_recon *e, int async);
int dsb200_recon_sync(dsb200_recon *e, dsb200_scene *scene, dsb200_render_state *rs);

int dsb200_reset_scene(dsb200_recon *e, dsb200_scene *scene);                                  /* ResetScene :50 */
int dsb200_allocate_from_depth(dsb200_recon *e, dsb200_scene *scene, const dsb200_view *view,  /* AllocateSceneFromDepth :57 */
                               const dsb200_tracking *tracking, dsb200_render_state *rs,
                               int only_update_visible_list, int is_defusion);
int dsb200_integrate(dsb200_recon *e, dsb200_scene *scene, const dsb200_view *view,            /* IntegrateIntoScene :63 */
                     const dsb200_tracking *tracking, const dsb200_render_state *rs);
int dsb200_deintegrate(dsb200_recon *e, dsb200_scene *scene, const dsb200_view *view,          /* DeIntegrateIntoScene :66 */
                       const dsb200_tracking *tracking);
int dsb200_decay(dsb200_recon *e, dsb200_scene *scene, dsb200_render_state *rs,                /* Decay :70 */
                 int max_weight, int min_age, int force_all_voxels);
int dsb200_decay_defusion(dsb200_recon *e, dsb200_scene *scene, dsb200_render_state *rs,       /* DecayDefusionPart :73 */
                          int max_weight, int min_age, int force_all_voxels);
int dsb200_slide_window(dsb200_recon *e, dsb200_scene *scene, dsb200_render_state *rs, int max_age);          /* SlideWindow :76 */
int dsb200_slide_window_defusion(dsb200_recon *e, dsb200_scene *scene, dsb200_render_state *rs,               /* SlideWindowDefusionPart :77 */
                                 int max_age, int max_size);
size_t dsb200_decayed_block_count(const dsb200_recon *e);                                      /* GetDecayedBlockCount :79 */
int dsb200_set_fusion_weight_params(dsb200_recon *e, const dsb200_weight_params *p);           /* SetFusionWeightParams :78-84 */
/* SDF shadow (no reference counterpart): the engine keeps one int16 per voxel (the sdf field alone, +25 % memory) in step with
 * every voxel write it performs, and the fused frame (dsb200_process_frame) raycasts from it instead of the 8-byte voxel
 * records -- a quarter of the bytes, so the visible set stays resident in the 126 MB L2 between integration and raycast.
 * Results are identical.  On by default; the stand-alone visualisation calls always read the voxel array.  If the caller
 * modifies the voxels behind the engine's back (loading a map, another engine instance) call invalidate: the shadow is
 * rebuilt from the voxels at its next use (it is also rebuilt when the engine is handed a different scene). */
int dsb200_recon_set_sdf_shadow(dsb200_recon *e, int on);
int dsb200_recon_invalidate_sdf_shadow(dsb200_recon *e);
/* introspection of the history queues (frameVisibleBlocks etc., CUDA/..._CUDA.h:41-48) */
int dsb200_history_sizes(const dsb200_recon *e, int *decay_q, int *slide_q, int *decay_defusion_q,
                         int *slide_defusion_q, int *snapshots);

/* ---- visualisation engine (ITMVisualisationEngine.h:71-124) ---- */
int dsb200_vis_create(dsb200_vis **out, int32_t max_visible, int32_t width, int32_t height, void *cuda_stream);
int dsb200_vis_destroy(dsb200_vis *v);
int dsb200_find_visible_blocks(dsb200_vis *v, const dsb200_scene *scene, const float *M, const float *intr,   /* FindVisibleBlocks :86 */
                               dsb200_render_state *rs);
int dsb200_count_visible_blocks(dsb200_vis *v, const dsb200_scene *scene, const dsb200_render_state *rs,      /* CountVisibleBlocks :91 */
                                int min_block_id, int max_block_id, int *count);
int dsb200_create_expected_depths(dsb200_vis *v, const dsb200_scene *scene, const float *M, const float *intr, /* CreateExpectedDepths :96 */
                                  dsb200_render_state *rs);
int dsb200_find_surface(dsb200_vis *v, const dsb200_scene *scene, const float *invM, const float *intr,       /* FindSurface :107 */
                        dsb200_render_state *rs);
int dsb200_create_icp_maps(dsb200_vis *v, const dsb200_scene *scene, const dsb200_view *view,                 /* CreateICPMaps :117 */
                           dsb200_tracking *tracking, dsb200_render_state *rs);
int dsb200_render_image(dsb200_vis *v, const dsb200_scene *scene, const float *M, const float *invM,          /* RenderImage :101 */
                        const float *intr, dsb200_render_state *rs, uint8_t *out_rgba, float *out_depth,
                        int render_type);
int dsb200_forward_render(dsb200_vis *v, const dsb200_scene *scene, const dsb200_view *view,                  /* ForwardRender :122 */
                          dsb200_tracking *tracking, dsb200_render_state *rs);
/* renderState->forwardProjection (Vector4f[W*H], device, engine-owned) and noFwdProjMissingPoints of the last ForwardRender */
int dsb200_forward_projection(dsb200_vis *v, const float **forward_projection, int *no_missing_points);
int dsb200_create_point_cloud(dsb200_vis *v);                                                                 /* CreatePointCloud :112 -> UNSUPPORTED */

/* ---- view building: ITMViewBuilder::UpdateView for the TRAFO_AFFINE calibration DenseSLAM sets up
 *      (DS/InfiniTamDriver.cpp:78-81, 280-287; convertDepthAffineToFloat, Engine/DeviceAgnostic/ITMViewBuilder.h:22-28),
 *      fused with the driver's per-pixel BGR -> RGBA host loop (DS/InfiniTamDriver.cpp:84-104).
 * raw_depth: int16[W*H] (millimetres for DenseSLAM), bgr: packed 8-bit B,G,R [W*H*3] or NULL (rgb left untouched); both are
 * DEVICE staging buffers (5 bytes per pixel cross PCIe instead of 8).  Writes view->depth (float, <=0 or >32000 raw -> -1,
 * else raw * scale + offset) and view->rgb (R,G,B,255).  Stream-ordered on cuda_stream, no host sync. ---- */
int dsb200_view_from_raw(const int16_t *raw_depth, const uint8_t *bgr, float scale, float offset, dsb200_view *view,
                         void *cuda_stream);

/* ---- fused frame: ITMDenseMapper::ProcessFrame + ITMTrackingController::Prepare in one submission
 *      (ITMLib/Engine/ITMDenseMapper.cpp:56-81, ITMTrackingController.cpp:21-58) ---- */
#define DSB200_FRAME_ALLOCATE 1
#define DSB200_FRAME_INTEGRATE 2
#define DSB200_FRAME_EXPECTED_DEPTHS 4
#define DSB200_FRAME_ICP_MAPS 8
#define DSB200_FRAME_ALL 15
int dsb200_process_frame(dsb200_recon *e, dsb200_vis *v, dsb200_scene *scene, const dsb200_view *view,
                         dsb200_tracking *tracking, dsb200_render_state *rs, int stages);

/* raycast only (FindSurface with the tracking pose), no ICP maps: the sharded path composites first */
#define DSB200_FRAME_RAYCAST 16

/* ---- sharded scene (no counterpart in the reference, which is single-GPU) ----
 * The voxel-block space is partitioned by block-coordinate ownership: bricks of 2^brick_log2 blocks
 * per axis are dealt to `count` shards by a hash of the brick coordinate.  A reconstruction engine
 * with a shard set only raises allocation requests for blocks it owns
 * (buildHashAllocAndVisibleTypePP, DeviceAgnostic/ITMSceneReconstructionEngine.h:328-396, with one
 * extra `continue`); everything downstream (ordered allocation, visible list, integration, expected
 * depths, raycast) then works on the shard's own hash table unchanged.  count <= 1 disables. */
typedef struct dsb200_shard {
    int32_t rank, count, brick_log2;
    int32_t halo;   /* 1: also hold (allocate + integrate) the blocks b with an owned block among b - {0,1}^3, i.e. the +1
                       neighbours the trilinear reads of castRay touch (ITMRepresentationAccess.h:258-272), so a ray inside
                       an owned block never reads a missing neighbour; replication factor ((B+1)/B)^3 for bricks of B blocks */
} dsb200_shard;
int dsb200_recon_set_shard(dsb200_recon *e, const dsb200_shard *s);
/* owner of block (bx,by,bz): the same function the kernels evaluate (host helper for tests / routing) */
int dsb200_shard_owner(int bx, int by, int bz, int count, int brick_log2);
/* does shard s hold block (bx,by,bz): owned, or (halo) a +1 neighbour of an owned block */
int dsb200_shard_holds(int bx, int by, int bz, const dsb200_shard *s);
/* Raycast of the WHOLE sharded scene, split by image tiles: this shard marches the 8x4-pixel tiles t with
 * t % count == rank; every block lookup goes to the hash table / voxel array of the shard that owns the block
 * (tables[] entries of the other shards are peer-GPU memory mapped with dsb200_ipc_open, so probes and voxel reads
 * travel over NVLink).  The march is castRay (DeviceAgnostic/ITMVisualisationEngine.h:124-196) probe for probe, so with
 * equal block sets the image equals the single-scene raycast bit for bit.  The expected-depth image used is the
 * element-wise min/max merge of the shards' range images (written to merged_range, Vector2f[W*H] scratch).
 * out_raycast (Vector4f[W*H]) may be the root shard's buffer: each shard stores only its own tiles.
 * Stream-ordered (no host sync): the caller orders it after every shard's integration and before the next frame. */
typedef struct dsb200_shard_tables {
    int32_t count, rank, brick_log2, pad_;
    const dsb200_hash_entry *hash[16];
    const dsb200_voxel *voxels[16];
    const float *range_image[16];
} dsb200_shard_tables;
int dsb200_raycast_sharded(dsb200_vis *v, const dsb200_scene *scene /* params + capacities, same on every shard */,
                           const dsb200_shard_tables *tables, const float *invM, const float *intr, int width, int height,
                           float *merged_range, float *out_raycast);
/* Per-pixel nearest-hit composite of n raycast results (Vector4f[W*H] each, voxel units, w = found):
 * out[p] = the hit closest to the camera centre (-> invM * (0,0,0,1) / voxel_size) among results[0..n),
 * ties to the lowest index, (0,0,0,0) when no shard hit.  results[i] may be peer-GPU memory mapped with
 * dsb200_ipc_open (the loads then travel over NVLink); row_begin/row_end restrict the rows written. */
int dsb200_composite_nearest(dsb200_vis *v, int n, const float *const *results, const float *invM,
                             float voxel_size, int width, int height, int row_begin, int row_end, float *out);
/* CreateICPMaps' second half (processPixelICP<true>) on an already raycast / composited point image */
int dsb200_icp_maps_from_raycast(dsb200_vis *v, const dsb200_scene *scene, const dsb200_view *view,
                                 dsb200_tracking *tracking, const float *raycast_result, uint8_t *out_rendering,
                                 int width, int height);
/* CUDA IPC plumbing for the peer-memory composite: export a device allocation, map a peer's */
#define DSB200_IPC_HANDLE_BYTES 64
int dsb200_ipc_export(void *device_ptr, uint8_t handle[DSB200_IPC_HANDLE_BYTES]);
int dsb200_ipc_open(const uint8_t handle[DSB200_IPC_HANDLE_BYTES], void **device_ptr);
int dsb200_ipc_close(void *device_ptr);

/* ---- host helpers (what ITMScene / ITMRenderState_VH / ITMView ctors do with cudaMalloc) ---- */
int dsb200_scene_alloc(dsb200_scene *scene, int32_t num_blocks, int32_t bucket_num, int32_t excess_size);
int dsb200_scene_free(dsb200_scene *scene);
int dsb200_render_state_alloc(dsb200_render_state *rs, int32_t total_entries, int32_t visible_capacity,
                              int32_t width, int32_t height);
int dsb200_render_state_free(dsb200_render_state *rs);
int dsb200_device_alloc(void **ptr, size_t bytes);
int dsb200_device_free(void *ptr);
int dsb200_host_alloc_pinned(void **ptr, size_t bytes);
int dsb200_host_free_pinned(void *ptr);
int dsb200_memcpy_h2d(void *dst, const void *src, size_t bytes, void *cuda_stream);
int dsb200_memcpy_d2h(void *dst, const void *src, size_t bytes, void *cuda_stream);
int dsb200_stream_synchronize(void *cuda_stream);
/* ORUtils Matrix4::inv restated (cofactor order preserved), ORUtils/Matrix.h:175-246 */
int dsb200_mat4_inverse(const float *M, float *out);
void dsb200_mat4_multiply(const float *A, const float *B, float *out);

/* ITMViewBuilder::UpdateView's useBilateralFilter branch (CPU/ITMViewBuilder_CPU.cpp:49-58, filterDepth
 * DeviceAgnostic/ITMViewBuilder.h:30-53): five passes of the 5x5 bilateral depth filter over view->depth, in place;
 * scratch = DEVICE buffer of depth_w * depth_h floats.  Off in DenseSLAM (ITMLibSettings.cpp:37).  Float arithmetic is the
 * reference's; its exp() is evaluated in double and rounded (see csrc/view_filter.cu for the parity statement).  Asynchronous
 * on cuda_stream like dsb200_view_from_raw. */
int dsb200_view_filter_depth(dsb200_view *view, float *scratch, void *cuda_stream);
/* one pass, ITMViewBuilder::DepthFiltering(image_out, image_in) (CPU/ITMViewBuilder_CPU.cpp:119-130); device pointers, in != out */
int dsb200_depth_filter_pass(const float *in, float *out, int32_t w, int32_t h, void *cuda_stream);
/* ITMViewBuilder::ConvertDisparityToDepth (CPU/ITMViewBuilder_CPU.cpp:94-106, convertDisparityToDepth
 * DeviceAgnostic/ITMViewBuilder.h:7-20): depth = 8 * param_y * fx_depth / (param_x - disparity), non-positive -> -1 */
int dsb200_convert_disparity_to_depth(const int16_t *disparity, float *depth, int32_t w, int32_t h, float param_x,
                                      float param_y, float fx_depth, void *cuda_stream);

/* Look-ahead for streaming callers (asynchronous mode only; no counterpart in the reference).  Enqueues the first half
 * of the NEXT frame's AllocateSceneFromDepth -- setToType3 over the current visible list and the per-pixel request pass
 * (ITMSceneReconstructionEngine_CPU.cpp:185-201), which only READ the hash table -- on an internal stream that waits for
 * the last scene write of the frame just submitted with dsb200_process_frame, not for its raycast: the two overlap.
 * The next scene call must be dsb200_process_frame (with DSB200_FRAME_ALLOCATE) or dsb200_allocate_from_depth for exactly
 * this view (same depth pointer, size, intrinsics), pose and render state; it picks the work up and continues with the
 * ordered allocation.  Any other scene call in between, or a mismatch, returns DSB200_ERR_INVALID.  Results are bit-identical
 * to the un-prefetched call sequence.  wait_cuda_event (optional): a cudaEvent_t after which next_view->depth is valid
 * (e.g. recorded after its upload).  A no-op while per-stage profiling is enabled. */
int dsb200_recon_prefetch_allocation(dsb200_recon *e, dsb200_scene *scene, const dsb200_view *next_view,
                                     const dsb200_tracking *next_tracking, dsb200_render_state *rs, void *wait_cuda_event);
/* Orders whatever is submitted next on the engine's stream after the pending look-ahead, without consuming it (for callers
 * that record timing events or flush caches between frames: bench.py).  dsb200_reset_scene abandons a pending look-ahead. */
int dsb200_recon_join_prefetch(dsb200_recon *e);
/* DEEP look-ahead (the default when the scene does not swap; tunable "deep_lookahead"): the look-ahead also runs the ordered
 * allocation and the visible-list build of the next frame beside this frame's raycast.  Its new hash entries are PENDING until the
 * consuming call: ptr = -3 - block, which every reader of the table (this library's and the reference's: they test ptr >= 0 or
 * ptr >= -1) takes for "no block here" -- the state the running frame must see; the next visible list is built in an engine-owned
 * buffer, and no counter of the current frame changes.  The consuming call commits all of it with one small kernel.  So between
 * dsb200_recon_prefetch_allocation and the consuming call the caller's buffers hold frame k's visible list and counters, and frame
 * k's table plus pending entries.  To read the table as frame k left it (checkpoints, digests), drop the look-ahead first: */
/* Drops a pending look-ahead and UNDOES it: hash entries, parent offsets, request keys and bitmaps are restored bit for bit,
 * visibility flags to "non-zero exactly for the members of the current visible list".  The next allocation then runs
 * un-prefetched.  A no-op without a pending look-ahead.  (dsb200_recon_notify_scene_modified does the same and also re-seeds the
 * device counters from the host mirrors.) */
int dsb200_recon_abandon_prefetch(dsb200_recon *e);

/* Volume checkpoint (the fork never calls ORUtils/MemoryBlockPersister.h, so the reference
 * has no save / resume): writes voxel.dat, alloc.dat, hash.dat, excess.dat -- and visible.dat, vistype.dat when rs is given --
 * into an existing directory, each file in MemoryBlockPersister's layout (int32 element count, raw elements; :88-140),
 * plus counters.txt (capacities, free-list tops, visible count, scene parameters).  Synchronises the device.
 * save refuses (DSB200_ERR_INVALID) a table that holds pending entries of a deep look-ahead: dsb200_recon_abandon_prefetch first.
 * load: the scene (and render state) must have the saved capacities; pass the engine that will keep working on the scene so
 * that its derived state (device counters, visibility bitmap, sdf shadow, pruning history, pending look-ahead) is reset. */
int dsb200_scene_save(const dsb200_scene *scene, const dsb200_render_state *rs, const char *directory);
int dsb200_scene_load(dsb200_recon *e /* may be NULL */, dsb200_scene *scene, dsb200_render_state *rs /* may be NULL */,
                      const char *directory);

/* ---- ICP depth tracker: the device-specific half ----
 * The reference's tracker is a host Levenberg-Marquardt loop (ITMLib/Engine/ITMDepthTracker.cpp:153-211, shared by
 * its CPU and CUDA builds) around ONE device virtual, ComputeGandH (ITMDepthTracker.h:59;
 * DeviceSpecific/CPU/ITMDepthTracker_CPU.cpp:13-79, DeviceSpecific/CUDA/ITMDepthTracker_CUDA.cu), plus the low-level
 * engine's FilterSubsampleWithHoles for the depth pyramid (ITMDepthTracker.cpp:66-76).  The adapter in INTEGRATION.md
 * subclasses ITMDepthTracker and forwards exactly these; the host loop stays the reference's. */
#define DSB200_TRACKER_ITERATION_ROTATION 1      /* TrackerIterationType, ITMLib/Utils/ITMLibDefines.h:289-295 */
#define DSB200_TRACKER_ITERATION_TRANSLATION 2
#define DSB200_TRACKER_ITERATION_BOTH 3
#define DSB200_TRACKER_ITERATION_NONE 4
#define DSB200_TRACKER_MAX_LEVELS 8
typedef struct dsb200_tracker dsb200_tracker;
/* ITMDepthTracker ctor (ITMDepthTracker.cpp:12-40): owns levels 1.. of the view-depth pyramid and the reduction scratch */
int dsb200_tracker_create(dsb200_tracker **out, int32_t width, int32_t height, int32_t levels, void *cuda_stream);
int dsb200_tracker_destroy(dsb200_tracker *t);
/* SetEvaluationData + PrepareForEvaluation (ITMDepthTracker.cpp:43-76).  depth: DEVICE float image of the tracker's
 * size, used in place as level 0; levels 1.. are rebuilt (FilterSubsampleWithHoles), intrinsics halve per level.
 * points / normals: DEVICE float4 maps of the last raycast (trackingState->pointCloud locations / colours), NOT
 * subsampled (SetEvaluationParams always selects scene level 0, :78-84); scene_pose = pose_pointCloud->GetM(). */
int dsb200_tracker_set_evaluation_data(dsb200_tracker *t, const float *depth, const float *view_intrinsics,
                                       const float *points, const float *normals, int32_t scene_w, int32_t scene_h,
                                       const float *scene_intrinsics, const float *scene_pose);
/* ComputeGandH at pyramid level `level`.  hessian: 36 floats, filled like the reference (hessian[r + c*6] for the
 * noPara x noPara block, symmetric; other entries untouched); nabla: 6 floats (first noPara written);
 * *f = sqrt(sum b^2) / valid if valid > 100 else 1e5; *no_valid_points = the return value of the reference.
 * Sums are accumulated in double along a fixed tree (deterministic).  Synchronous. */
int dsb200_tracker_compute_g_and_h(dsb200_tracker *t, int32_t level, int32_t iteration_type, float dist_thresh,
                                   const float *approx_inv_pose, float *f, float *nabla, float *hessian,
                                   int32_t *no_valid_points);
/* The same evaluation on caller-owned images (what a subclass of the reference's ITMDepthTracker forwards its protected
 * evaluation state to: viewHierarchyLevel->depth / intrinsics, sceneHierarchyLevel->pointsMap / normalsMap / intrinsics,
 * scenePose, distThresh[levelId], iterationType); the tracker handle only lends its reduction scratch and stream. */
int dsb200_tracker_compute_g_and_h_on(dsb200_tracker *t, const float *depth, int32_t view_w, int32_t view_h,
                                      const float *view_intrinsics, const float *points, const float *normals,
                                      int32_t scene_w, int32_t scene_h, const float *scene_intrinsics,
                                      const float *scene_pose, int32_t iteration_type, float dist_thresh,
                                      const float *approx_inv_pose, float *f, float *nabla, float *hessian,
                                      int32_t *no_valid_points);
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  The ComputeGandH evaluation of
 * dsb200_tracker_compute_g_and_h_on without a dsb200_tracker: it allocates NOTHING (no cudaMalloc / cudaFree / cudaMallocHost, no
 * allocating library routine), its reduction scratch (CTA partials + ticket) is a file-scope `static __device__` array of
 * DSB200_TRACKER_STATIC_SCRATCH_BYTES that every GPU loading libdsb200 holds until exit, and it is ASYNCHRONOUS on cuda_stream.
 * All image / map pointers are DEVICE pointers as above; the matrices and intrinsics are HOST floats read at the call.
 * sums: caller-owned DEVICE buffer of 32 doubles, 8-byte aligned, written by the kernel: the raw sums in the layout of
 * dsb200_tracker_last_sums (bit-identical to it: same kernel, same grid, same fixed summation tree); the slots after the valid
 * count (3-parameter iterations use 11 of them, BOTH uses 29) are 0.
 * Limits: view_w * view_h <= DSB200_TRACKER_STATIC_MAX_PIXELS (larger -> DSB200_ERR_INVALID); iteration type NONE is rejected
 * (nothing to launch).  ALL calls on one GPU share the one array: calls that may overlap on the GPU (different streams, or
 * concurrent graph branches) RACE -- serialise them on one stream.  The ticket word is cleared by a cudaMemsetAsync on the same
 * stream before every launch; the partials need no clearing (every slot read was written by the same launch). */
#define DSB200_TRACKER_STATIC_MAX_PIXELS (1 << 22)          /* 2048 x 2048 */
#define DSB200_TRACKER_STATIC_SCRATCH_BYTES (296 * 32 * 8 + 256)   /* 76,032 B of device memory per GPU */
int dsb200_tracker_compute_g_and_h_static(const float *depth, int32_t view_w, int32_t view_h, const float *view_intrinsics,
                                          const float *points, const float *normals, int32_t scene_w, int32_t scene_h,
                                          const float *scene_intrinsics, const float *scene_pose, int32_t iteration_type,
                                          float dist_thresh, const float *approx_inv_pose, double *sums /* DEVICE [32] */,
                                          void *cuda_stream);
/* raw double sums of the last compute_g_and_h: [0,NH) hessian lower triangle by rows, [NH,NH+NP) nabla, sum b^2, count */
int dsb200_tracker_last_sums(const dsb200_tracker *t, double *sums /* [32] */);
/* pyramid level access (device pointer owned by the tracker; level 0 is the caller's image) */
int dsb200_tracker_level(const dsb200_tracker *t, int32_t level, const float **depth, int32_t *w, int32_t *h,
                         float *intrinsics);
/* ITMLowLevelEngine::FilterSubsampleWithHoles(ITMFloatImage) (DeviceAgnostic/ITMLowLevelEngine.h:26-47):
 * out is (w/2) x (h/2), device pointers */
int dsb200_filter_subsample_with_holes(const float *in, int32_t w, int32_t h, float *out, void *cuda_stream);

/* ITMDepthTracker::TrackCamera (ITMDepthTracker.cpp:153-211) as one call, for callers without the reference's
 * host loop: pyramid, Levenberg-Marquardt iterations per level (2, 4, 6, ... from the finest; distance thresholds
 * dist_thresh * (level + 1) / levels, :20-30), Cholesky step (ORUtils/Cholesky.h), pose re-orthogonalised through its
 * six parameters after every step (ITMPose::Coerce, Objects/ITMPose.cpp:82-262).  Host arithmetic is float without
 * contraction, like the reference.  In: tracking->M_d = start pose, tracking->icp_points / icp_normals = the maps of
 * the last raycast (scene_w x scene_h), scene_pose = pose_pointCloud->GetM().  Out: tracking->M_d / invM_d.
 * Where the reference would step along an uninitialised Hessian (no valid evaluation yet), the pose is left alone. */
typedef struct dsb200_tracker_config {
    int32_t levels;                                  /* noHierarchyLevels (ITMLibSettings.cpp:62-69: 5) */
    int32_t regime[DSB200_TRACKER_MAX_LEVELS];       /* trackingRegime, level 0 = finest */
    int32_t no_icp_run_till_level;                   /* ITMLibSettings.cpp:72-73: 0 */
    float dist_thresh;                               /* depthTrackerICPThreshold, :13: 0.1^2 */
    float termination_threshold;                     /* depthTrackerTerminationThreshold, :16: 1e-3 */
} dsb200_tracker_config;
typedef struct dsb200_tracker_report {
    int32_t evaluations, steps, last_valid_points;
    float last_f;
} dsb200_tracker_report;
int dsb200_tracker_track_camera(dsb200_tracker *t, const dsb200_tracker_config *cfg, const dsb200_view *view,
                                dsb200_tracking *tracking, const float *scene_pose, int32_t scene_w, int32_t scene_h,
                                dsb200_tracker_report *report /* may be NULL */);
/* ITMPose::Coerce on a column-major 4x4 (in place); params (optional) receives (tx, ty, tz, rx, ry, rz) */
int dsb200_pose_coerce(float *M, float *params);
/* ITMPose::SetFrom(tx..rz) -> GetM() (SetModelViewFromParams, Objects/ITMPose.cpp:82-180) */
int dsb200_pose_from_params(const float *params, float *M);

/* per-stage device timing of dsb200_process_frame (CUDA events on the engine's stream).  Stage order:
 * 0 allocate(+visible list) 1 integrate 2 expected depths 3 raycast 4 ICP maps.  read() synchronises. */
#define DSB200_PROFILE_STAGES 5
int dsb200_profile_enable(dsb200_recon *e, int on);
int dsb200_profile_read(dsb200_recon *e, float *stage_ms /* [DSB200_PROFILE_STAGES] */);
/* kernels of the last AllocateSceneFromDepth: 0 begin-frame 1 requests 2 ordered allocation 3 visibility re-test 4 visible list */
#define DSB200_PROFILE_ALLOC_STAGES 5
int dsb200_profile_read_allocate(dsb200_recon *e, float *stage_ms /* [DSB200_PROFILE_ALLOC_STAGES] */);

/* ITMViewBuilder::ComputeNormalAndWeights (CPU/ITMViewBuilder_CPU.cpp:132-143, computeNormalAndWeight
 * DeviceAgnostic/ITMViewBuilder.h:55-112; CUDA/ITMViewBuilder_CUDA.cu:157-171), UpdateView's modelSensorNoise step: per-pixel
 * depth normal (float4, w = 1 valid / -1 rejected) from central differences of the un-projected neighbours and the depth
 * uncertainty sigma_z = 0.0012 + 0.0019 (z - 0.4)^2 + 0.0001 / sqrt(z) * (theta / (pi/2 - theta))^2, theta = acos(normal.z).
 * depth, normal4 (w*h*4 floats) and sigma_z (w*h floats) are DEVICE pointers; intrinsics = HOST {fx, fy, cx, cy}
 * (projectionParamsSimple.all).  Like the reference it writes only the interior [2, w-2) x [2, h-2), and for a rejected pixel only
 * normal.w and sigma_z.  Bit-exact with the reference CPU builder (acos restates glibc's acosf, csrc/view_filter.cu).  Consumed
 * only by the weighted-ICP tracker (ITMWeightedICPTracker.cpp:67), which DenseSLAM does not select. */
int dsb200_compute_normal_and_weights(const float *depth, float *normal4, float *sigma_z, int32_t w, int32_t h,
                                      const float *intrinsics, void *cuda_stream);

/* ---- host <-> device block swapping: ITMSwappingEngine on ITMGlobalCache ----
 * ITMGlobalCache (Objects/ITMGlobalCache.h:18-129): the per-entry swap states and the host-side block store.
 * states: ITMHashSwapState (ITMLibDefines.h:76-85), one byte per hash entry: 0 = most recent data on the host, 1 = on the host
 * and in the volume, not combined yet, 2 = most recent data in the volume.  In a cache obtained from dsb200_swap_cache_of the
 * states are a DEVICE array, has_stored is HOST memory and stored is NULL: the library keeps the stored blocks in a compact arena
 * of pinned chunks that its kernels read and write in place (dsb200_swap_stored_block returns a block's address). */
typedef struct dsb200_swap_cache {
    uint8_t *states;               /* [total_entries] */
    uint8_t *has_stored;           /* HOST bool[total_entries]            (hasStoredData) */
    dsb200_voxel *stored;          /* HOST [total_entries * 512]          (storedVoxelBlocks); NULL from dsb200_swap_cache_of */
    int32_t total_entries;         /* bucket_num + excess_size */
    int32_t transfer_blocks;       /* SDF_TRANSFER_BLOCK_NUM (0x1000, ITMLibDefines.h:42): most blocks moved by one call */
} dsb200_swap_cache;
#define DSB200_DEFAULT_TRANSFER_BLOCKS 0x1000
typedef struct dsb200_swap dsb200_swap;
/* ITMGlobalCache() + ITMSwappingEngine_CUDA(): states zeroed, store empty (it grows by pinned 64 MB chunks as blocks are
 * swapped out). */
int dsb200_swap_create(dsb200_swap **out, int32_t total_entries, int32_t transfer_blocks, void *cuda_stream);
int dsb200_swap_destroy(dsb200_swap *s);
int dsb200_swap_cache_of(dsb200_swap *s, dsb200_swap_cache *out);
/* ITMGlobalCache::GetStoredVoxelBlock (Objects/ITMGlobalCache.h:36): HOST address of the entry's 512 stored voxels, NULL when
 * HasStoredData(entry) is false */
const dsb200_voxel *dsb200_swap_stored_block(const dsb200_swap *s, int32_t entry);
/* The swap-state pass of AllocateSceneFromDepth when scene->useSwapping (CPU/ITMSceneReconstructionEngine_CPU.cpp:287-299,
 * CUDA .cu:1730-1733): every entry whose visibility flag is non-zero and whose state is not 2 goes to state 1.  Call it after
 * dsb200_recon_allocate_from_depth (the flags it reads are that call's final flags). */
int dsb200_swap_mark_visible(dsb200_swap *s, const dsb200_render_state *rs);
/* ITMSwappingEngine::IntegrateGlobalIntoLocal (CPU/ITMSwappingEngine_CPU.cpp:22-131, CUDA .cu:263-306): the first
 * transfer_blocks entries in state 1 (ascending entry index) get their stored block, if any, combined into the volume
 * (combineVoxelDepthInformation / combineVoxelColorInformation, DeviceAgnostic/ITMSwappingEngine.h:7-45) and go to state 2.
 * *moved = entries handled.  An entry whose ptr is negative is not combined (the reference indexes the volume with it). */
int dsb200_swap_integrate_global_into_local(dsb200_swap *s, dsb200_scene *scene, int32_t *moved);
/* ITMSwappingEngine::SaveToGlobalMemory (CPU .cpp:133-297, CUDA .cu:308-390): the first transfer_blocks entries (ascending)
 * in state 2 with a block in the volume -- and, with a render state, a visibility flag of 0 -- are copied to the host store,
 * go to state 0, and while the free list has room (the reference's test vbaIdx < SDF_BUCKET_NUM - 1) their block is reset,
 * pushed on the free list and the entry's ptr becomes -1.  rs = NULL is the overload DenseSLAM calls to swap a whole local map
 * out (DS/DenseSlam.h:248-251).  scene->last_free_block is read and updated: synchronise an asynchronous engine first
 * (dsb200_recon_sync), and tell the engine afterwards (dsb200_recon_notify_scene_modified). */
int dsb200_swap_save_to_global_memory(dsb200_swap *s, dsb200_scene *scene, const dsb200_render_state *rs, int32_t *moved);
/* Names the reconstruction engine that works on the same scene (NULL detaches).  While attached, IntegrateGlobalIntoLocal and
 * SaveToGlobalMemory call dsb200_recon_refresh_mirrors before they read the free-list top (an ASYNCHRONOUS engine's host mirrors are stale
 * until then) and dsb200_recon_notify_volume_swapped after they changed the volume (free-list top re-seeded on the device, pending
 * look-ahead undone, sdf shadow rebuilt on next use).  ITMSwappingEngine_B200 attaches the engine it is constructed with. */
/* dsb200_recon_sync when the device counters are the authoritative ones (asynchronous engine with frames in flight), otherwise
 * only a stream synchronisation: the host mirrors are then already current and are left alone */
int dsb200_recon_refresh_mirrors(dsb200_recon *e, dsb200_scene *scene, dsb200_render_state *rs);
int dsb200_swap_attach_recon(dsb200_swap *s, dsb200_recon *recon);
/* The host block store is a compact arena of PINNED 64 MB chunks that grows with the number of blocks ever swapped out (the
 * reference's is a lazily touched pageable malloc of the whole table, Objects/ITMGlobalCache.h:75-105).  max_blocks bounds it
 * (0 = no bound but the table size; DenseSLAM parks many local maps, one cache each).  A SaveToGlobalMemory call that meets the
 * bound -- or cannot pin another chunk -- moves the candidates that still fit and reports that number; when none fits it
 * returns DSB200_ERR_CUDA with nothing changed. */
int dsb200_swap_set_store_budget(dsb200_swap *s, int32_t max_blocks);
/* ITMGlobalCache::SaveToFile / ReadFromFile (Objects/ITMGlobalCache.h:75-105): bool[total] then total blocks */
int dsb200_swap_save_file(dsb200_swap *s, const char *path);
int dsb200_swap_load_file(dsb200_swap *s, const char *path);
/* Voxels, hash pointers or the free list were changed behind the reconstruction engine's back (swapping): device counters
 * are re-seeded from the scene's host mirrors, the sdf shadow is rebuilt on first use, a pending look-ahead is dropped. */
int dsb200_recon_notify_scene_modified(dsb200_recon *e);
/* The narrower notification the swapping engine sends (it changes voxel blocks, entry pointers and scene->last_free_block only):
 * an asynchronous engine keeps its device-side visible count and excess-list top, takes scene->last_free_block, drops a pending
 * look-ahead and rebuilds the sdf shadow on next use.  (The host mirror of the visible count lives in a render state the
 * swapping engine never sees, so re-seeding every counter from host mirrors would lose the visible list.) */
int dsb200_recon_notify_volume_swapped(dsb200_recon *e, const dsb200_scene *scene);

/* ---- diagnostics of the exact-division fast paths (csrc/exactdiv.cuh) ----
 * force_plain_division: run the integrate / de-integrate kernel variants that use the compiler's `/` everywhere.
 * selftest_division: evaluates every fast-path helper against `/` on n pseudo-random operand tuples drawn from
 * the operand windows the kernels guarantee (plus the out-of-window fall-backs); *mismatches = number of
 * bit-level disagreements (must be 0). */
int dsb200_debug_force_plain_division(dsb200_recon *e, int on);
/* launch-shape knobs (csrc/kernels.cuh: Tunables; "ray_steps", "ray_group", "ray_ctas", "ray_finish_ctas", "ray_bottom_up", "ray_lpt",
 * "fuse_kernel", "fuse_ctas", "fuse_minb", "deep_lookahead"): process-wide, results never depend on them (tools/tune.py sweeps them on the GPU) */
int dsb200_debug_set_tunable(const char *name, int value);
int dsb200_selftest_division(unsigned long long n, unsigned long long seed, unsigned long long *mismatches);
/* the bilateral filter's exp (csrc/view_filter.cu, a restatement of glibc's expf) on n HOST floats: y[i] = expf(x[i]);
 * tests compare it bit for bit with the host C library's expf */
int dsb200_selftest_expf(const float *x, float *y, int32_t n);
/* the sensor-noise model's acos (csrc/view_filter.cu, a restatement of glibc's fdlibm acosf) on n HOST floats */
int dsb200_selftest_acosf(const float *x, float *y, int32_t n);

/* ---- meshing: ITMMeshingEngine<ITMVoxel,ITMVoxelBlockHash>::MeshScene, ITMLib/Engine/ITMMeshingEngine.h:19-26;
 * CPU ITMLib/Engine/DeviceSpecific/CPU/ITMMeshingEngine_CPU.cpp:20-83, CUDA .../CUDA/ITMMeshingEngine_CUDA.cu:60-195; math in
 * ITMLib/Engine/DeviceAgnostic/ITMMeshingEngine.h (buildVertList, sdfInterp) and ITMRepresentationAccess.h:699-736 (vertex colours).
 * Marching cubes over every allocated block: triangles in the CPU engine's order -- hash slot ascending, voxel z / y / x,
 * table order -- (the reference CUDA engine appends with atomicAdd, i.e. in schedule order), vertices in metres, per-vertex
 * colours in [0, 1].  `triangles` is DEVICE memory for max_triangles records (ITMMesh::Triangle layout, 72 bytes; the reference
 * sizes it SDF_LOCAL_BLOCK_NUM * 64); the CPU engine's cap rule is kept: once max_triangles - 1 are stored every further
 * triangle overwrites the last slot.  *no_triangles = ITMMesh::noTotalTriangles.  Synchronous. */
typedef struct dsb200_triangle { float p0[3], p1[3], p2[3], c0[3], c1[3], c2[3]; } dsb200_triangle;
typedef struct dsb200_mesher dsb200_mesher;
int dsb200_mesher_create(dsb200_mesher **out, int32_t num_blocks, void *stream);
int dsb200_mesher_destroy(dsb200_mesher *m);
int dsb200_mesh_scene(dsb200_mesher *m, const dsb200_scene *scene, dsb200_triangle *triangles, uint32_t max_triangles,
                      uint32_t *no_triangles);

/* counters for bench.py: kernels launched by this library since process start */
unsigned long long dsb200_kernel_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* DSB200_H */
