/* oracle/restate.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement ("oracle") of the reference's volumetric-fusion hot path.  Same POD
 * structs as include/dsb200.h, but every pointer is a HOST pointer and every function is a
 * serial loop.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this; the product (libdsb200.so) never does.
 *
 * Pinning: oracle/_ref/libitmref.so (the unmodified reference CPU engines) is compared
 * bit-for-bit with this file by tests/test_oracle_vs_ref.py, and the committed fixtures under
 * tests/golden/ were generated from the reference.  Decay / slide-window / de-integration have
 * NO CPU implementation in the reference (CPU/ITMSceneReconstructionEngine_CPU.cpp:345-396
 * throw) and no reference test: for those functions parity is UNPINNED by reference outputs;
 * they restate CUDA/ITMSceneReconstructionEngine_CUDA.cu:426-1082 under the deterministic
 * schedule described in DESIGN.md.
 */
#ifndef RESTATE_H
#define RESTATE_H
#include "../include/dsb200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rst_recon rst_recon;

rst_recon *rst_recon_create(int num_blocks, int bucket_num, int excess_size);
void rst_recon_destroy(rst_recon *e);
void rst_set_fusion_weight_params(rst_recon *e, const dsb200_weight_params *p);
size_t rst_decayed_block_count(const rst_recon *e);
void rst_history_sizes(const rst_recon *e, int *decay_q, int *slide_q, int *decay_defusion_q,
                       int *slide_defusion_q, int *snapshots);

void rst_set_shard(rst_recon *e, const dsb200_shard *s);
int rst_shard_owner(int bx, int by, int bz, int count, int brick_log2);
int rst_shard_holds(int bx, int by, int bz, const dsb200_shard *s);
void rst_raycast_sharded(int n, const dsb200_scene *const *scenes, const float *const *range_images, int brick_log2,
                         const float *invM, const float *intr, int W, int H, float *merged_range, float *out);
void rst_composite_nearest(int n, const float *const *results, const float *invM, float voxel_size, int W, int H,
                           int row_begin, int row_end, float *out);
void rst_icp_maps_from_raycast(const dsb200_scene *scene, const dsb200_view *view, dsb200_tracking *tracking,
                               const float *raycast_result, uint8_t *out_rendering, int W, int H);

void rst_reset_scene(rst_recon *e, dsb200_scene *scene);
int rst_allocate_from_depth(rst_recon *e, dsb200_scene *scene, const dsb200_view *view,
                            const dsb200_tracking *tracking, dsb200_render_state *rs,
                            int only_update_visible_list, int is_defusion);
void rst_integrate(rst_recon *e, dsb200_scene *scene, const dsb200_view *view,
                   const dsb200_tracking *tracking, const dsb200_render_state *rs);
void rst_deintegrate(rst_recon *e, dsb200_scene *scene, const dsb200_view *view,
                     const dsb200_tracking *tracking);
void rst_decay(rst_recon *e, dsb200_scene *scene, dsb200_render_state *rs, int max_weight, int min_age,
               int force_all_voxels);
void rst_decay_defusion(rst_recon *e, dsb200_scene *scene, dsb200_render_state *rs, int max_weight,
                        int min_age, int force_all_voxels);
void rst_slide_window(rst_recon *e, dsb200_scene *scene, dsb200_render_state *rs, int max_age);
void rst_slide_window_defusion(rst_recon *e, dsb200_scene *scene, dsb200_render_state *rs, int max_age,
                               int max_size);

void rst_find_visible_blocks(const dsb200_scene *scene, const float *M, const float *intr,
                             dsb200_render_state *rs);
int rst_count_visible_blocks(const dsb200_scene *scene, const dsb200_render_state *rs, int min_id, int max_id);
void rst_create_expected_depths(const dsb200_scene *scene, const float *M, const float *intr,
                                dsb200_render_state *rs);
void rst_find_surface(const dsb200_scene *scene, const float *invM, const float *intr,
                      dsb200_render_state *rs);
void rst_create_icp_maps(const dsb200_scene *scene, const dsb200_view *view, dsb200_tracking *tracking,
                         dsb200_render_state *rs);
void rst_render_image(const dsb200_scene *scene, const float *M, const float *invM, const float *intr,
                      dsb200_render_state *rs, uint8_t *out_rgba, float *out_depth, int render_type);
void rst_forward_render(const dsb200_scene *scene, const dsb200_view *view, dsb200_tracking *tracking,
                        dsb200_render_state *rs, float *forward_projection, int *missing_points,
                        int *no_missing);

/* ICP depth tracker, device half (CPU/ITMDepthTracker_CPU.cpp:13-79, CPU/ITMLowLevelEngine_CPU.cpp:50-62) */
void rst_filter_subsample_with_holes(const float *in, int w, int h, float *out);
int rst_tracker_compute_g_and_h(const float *depth, int view_w, int view_h, const float *view_intrinsics,
                                const float *points, const float *normals, int scene_w, int scene_h,
                                const float *scene_intrinsics, const float *scene_pose, const float *approx_inv_pose,
                                float dist_thresh, int iteration_type, int double_sums, float *f, float *nabla,
                                float *hessian, double *sums);

void rst_convert_disparity_to_depth(const short *disparity, float *depth, int w, int h, float param_x, float param_y, float fx_depth);
void rst_view_filter_depth(float *depth, int w, int h);
void rst_compute_normal_and_weights(const float *depth, float *normal4, float *sigma_z, int w, int h, const float *intrinsics);
/* swapping engine on host arrays (cache->states is a HOST array here) */
void rst_swap_mark_visible(dsb200_swap_cache *cache, const dsb200_render_state *rs);
int rst_swap_integrate_global_into_local(dsb200_swap_cache *cache, dsb200_scene *scene);
int rst_swap_save_to_global_memory(dsb200_swap_cache *cache, dsb200_scene *scene, const dsb200_render_state *rs);
void rst_pose_coerce(float *M, float *params);
void rst_pose_from_params(const float *params, float *M);
void rst_track_camera(const dsb200_tracker_config *cfg, const dsb200_view *view, dsb200_tracking *tracking,
                      const float *scene_pose, int scene_w, int scene_h, int double_sums, dsb200_tracker_report *report);

int rst_mat4_inverse(const float *M, float *out);
void rst_mat4_multiply(const float *A, const float *B, float *out);

/* ITMMeshingEngine_CPU::MeshScene (CPU/ITMMeshingEngine_CPU.cpp:20-83) on a host scene; returns noTotalTriangles.
 * PINNED: tests/test_meshing.py compares it with the reference CPU mesher bit for bit. */
unsigned rst_mesh_scene(const dsb200_scene *scene, dsb200_triangle *triangles, unsigned max_triangles);
/* the marching-cubes tables the restatement uses, decoded (for the table tests) */
void rst_mc_tables(int *edge_table_256, int *triangle_table_256x16);

#ifdef __cplusplus
}
#endif
#endif
