/*
 * tygrit_b200.h — C ABI of the B200-native collision-checking engine that sits
 * behind TyGrit's MotionPlanner / CollisionCheckFn / Scene / robot-filter surface.
 *
 * There is no FFI in the reference: its planner reaches the collision arithmetic
 * through pybind11 into third-party VAMP (TyGrit/planning/fetch/vamp_preview.py).
 * Each entry point below names the reference call it replaces (file:line relative
 * to the TyGrit tree).  INTEGRATION.md shows the ctypes binding a maintainer adds.
 *
 * Conventions
 *  - every pointer marked DEVICE is a CUDA device pointer on the env's device
 *    (e.g. torch.Tensor.data_ptr()); pointers marked HOST are ordinary host memory;
 *  - the library never frees caller memory and never keeps a caller pointer after
 *    the call has been enqueued;
 *  - `stream` is a cudaStream_t passed as void* (0 = default stream); calls are
 *    asynchronous unless documented otherwise;
 *  - return value: 0 = ok, < 0 = error (TG_E_*); tg_last_error() gives the text;
 *  - base poses are [x, y, theta] (vamp_preview.py:103-104,162), arm is the 8
 *    planning joints [torso, shoulder_pan, ..., wrist_roll] (constants.py:54-63);
 *  - verdict bytes: 1 = collision-free, 0 = in collision
 *    (validate_config semantics, vamp_preview.py:159-163).
 */
#ifndef TYGRIT_B200_H
#define TYGRIT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TG_OK 0
#define TG_E_ARG (-1)      /* bad argument                                     */
#define TG_E_CUDA (-2)     /* a CUDA runtime call failed (see tg_last_error)   */
#define TG_E_CAPACITY (-3) /* cloud larger than the env was created for        */
#define TG_E_STATE (-4)    /* call not valid in the env's current state        */

#define TG_N_SPHERES 111 /* Fetch collision spheres (constants.py:84-225)       */
#define TG_N_LINKS 15    /* links returned by forward_kinematics (fk_numpy.py)  */
#define TG_MAX_OBSTACLES 64
#define TG_MAX_ATTACHED 64

typedef struct tg_env tg_env; /* opaque: device buffers of one collision world  */

/* ---- library ---------------------------------------------------------------- */
const char* tg_last_error(void);
int tg_version(void);

/* ---- environment (replaces vamp_preview.Environment, vamp_preview.py:46) ----- */
/* max_points: capacity of the cloud; voxels_per_metre: resolution of the
 * distance-bound grid (50 = 2 cm); point_radius: radius of every cloud point
 * (planning/config.py:13).  max_voxels bounds the grid memory (0 = default). */
int tg_env_create(tg_env** out, int device, int64_t max_points, float voxels_per_metre,
                  float point_radius, int64_t max_voxels);
int tg_env_destroy(tg_env* env);

/* Replace the cloud: clear_pointclouds + add_pointcloud (vamp_preview.py:183-240).
 * xyz: DEVICE, n x 3 f32 (n may be 0 = empty world).  Builds the sorted cell table
 * and the distance-bound grid on `stream`.  Synchronises the stream once (cloud
 * bounds are read back to size the grids) unless fixed bounds were set.          */
int tg_env_build(tg_env* env, const float* xyz, int64_t n, void* stream);
/* Same, xyz on the HOST (pageable or pinned).                                    */
int tg_env_build_host(tg_env* env, const float* xyz, int64_t n, void* stream);
/* Fix the workspace box [lo, hi] (HOST, 3 floats each): tg_env_build then never
 * synchronises (CUDA-graph capturable); points outside are ignored and counted
 * in tg_env_stats.  Pass NULL, NULL to return to automatic bounds.               */
int tg_env_set_bounds(tg_env* env, const float* lo, const float* hi);
/* Resolution of the narrow-phase cell grid used by the next build (cells per metre, the same on all
 * three axes; the default is anisotropic: 64 cells per metre along x, 14 along y and z).  Coarser cells mean fewer rows per ambiguous sphere but more candidate
 * points per row.                                                                */
int tg_env_set_cells_per_metre(tg_env* env, float cells_per_metre);
/* Anisotropic cells: a cell row runs along x and is one contiguous point range, so the x resolution
 * only tightens a sphere's chord (fewer candidate points) while the y/z resolution sets how many
 * rows a sphere touches (two cell_start loads each).                                             */
int tg_env_set_cell_shape(tg_env* env, float cells_per_metre_x, float cells_per_metre_yz);
/* point radius used by the next build (add_pointcloud(point_radius=...), :205-209) */
int tg_env_set_point_radius(tg_env* env, float point_radius);

/* add_sphere / clear_spheres (vamp_preview.py:242-255): xyzr HOST, n x 4 f32.    */
int tg_env_set_obstacles(tg_env* env, const float* xyzr, int n, void* stream);
/* attach_to_eef / detach_from_eef (vamp_preview.py:257-304): spheres xyzr HOST
 * (n x 4, positions in the attachment frame), offset position (3) and quaternion
 * [x,y,z,w] (4) relative to gripper_link, HOST.  n = 0 detaches.                 */
int tg_env_set_attachment(tg_env* env, const float* xyzr, int n, const float* offset_pos,
                          const float* offset_quat, void* stream);

/* Optional self-collision term of validate_whole_body_config (VAMP also tests the robot's
 * spheres against each other; its pair list is not available, so the link pairs are the reference's own
 * SELF_COLLISION_PAIRS, TyGrit/rl/train.py:78-107): when on, a configuration / edge sample also collides if two
 * spheres of a listed link pair touch (|c_i - c_j| <= r_i + r_j), and the clearance output becomes
 * min(cloud clearance, self clearance).  Off by default.                              */
int tg_env_set_self_collision(tg_env* env, int on);

/* Structure hand-off for multi-GPU (cloud structure replicated on every rank):
 * export gives a contiguous DEVICE blob + its size; import adopts a blob that was
 * broadcast into `blob` on this device.                                         */
int tg_env_export(tg_env* env, void** blob, size_t* bytes);
int tg_env_import(tg_env* env, const void* blob, size_t bytes, void* stream);
size_t tg_env_blob_capacity(const tg_env* env);
/* DEVICE address of the env's own blob storage (valid for the env's lifetime): a
 * receiving rank lets NCCL write straight into it, then calls tg_env_import on it. */
int tg_env_blob_ptr(tg_env* env, void** blob);

/* stats (HOST out, 8 doubles): n_points, n_dropped, grid nx, ny, nz, cells,
 * voxels_per_metre actually used, slack [m].                                    */
int tg_env_stats(const tg_env* env, double* out8);

/* Layout of the built structure (HOST out, 20 doubles): [0] empty flag, [1..3] distance-grid nx, ny, nz, [4..6] index bias
 * per axis (voxel index = round(p * vpm + bias) - 2^23), [7] voxels per metre, [8] metres per code, [9] slack, [10] window
 * in voxels, [11] byte offset of the distance grid in the blob, [12] byte offset of the coarse far-field grid,
 * [13..15] its dims, [16] voxels per coarse block edge, [17] metres per coarse code.  For tests and tools.          */
int tg_env_grid_info(const tg_env* env, double* out20);

/* ---- forward kinematics (fk_numpy.py:32-157, fk_torch.py:119-283) ------------- */
/* World centres of the 111 spheres in the reference's sphere order; head2 may be
 * NULL (= 0, 0: the planner's model).  arm B x 8, base B x 3, head2 B x 2,
 * centres B x 111 x 3, all DEVICE f32.                                          */
int tg_fk_spheres(const float* arm, const float* base, const float* head2, int64_t B,
                  float* centres, void* stream);
/* Link frames in base_link: q10 B x 10 -> frames B x 15 x 4 x 4 (row-major) f32,
 * link order = fk_numpy's (see tygrit_b200/fetch_model.py LINK_NAMES).           */
int tg_fk_frames(const float* q10, int64_t B, float* frames, void* stream);

/* ---- verdicts ------------------------------------------------------------------ */
/* validate_whole_body_config for a batch (vamp_preview.py:159-163,328-338).
 * valid: DEVICE B bytes.  clearance: DEVICE B f32 or NULL; when given it receives
 * min(exact clearance, clearance_cap) (cap = 0.25 m), within 3e-6 m of fp64.  The first
 * call with a clearance output after a build / import also builds a half-resolution
 * distance grid on `stream` (3 launches; its buffer is allocated on first use, so make
 * that call outside a stream capture).                                           */
int tg_check_configs(tg_env* env, const float* arm, const float* base, int64_t B,
                     uint8_t* valid, float* clearance, void* stream);
/* tg_check_configs for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  Same verdicts as
 * tg_check_configs (no clearance output); the queue between its two kernels lives in a static device array of the library
 * (4 MiB + 256 B, held by every GPU that loads the library until exit) instead of the handle's cudaMalloc'ed queue, and the
 * call makes no cudaMalloc / cudaFree or other allocating call.  B <= TG_NOALLOC_MAX_CONFIGS, else TG_E_CAPACITY with a
 * message in tg_last_error.  All calls on one device share the array: calls that overlap on the GPU (different streams or
 * handles) race; calls on one stream are fine.  The first call per device looks the array's address up
 * (cudaGetSymbolAddress): make it outside a stream capture.                                                             */
#define TG_NOALLOC_QUEUE_ITEMS 65536   /* 64-byte items of the static queue; overflow is resolved inline, never wrong */
#define TG_NOALLOC_MAX_CONFIGS 262144  /* configurations per call                                                     */
int tg_check_configs_noalloc(tg_env* env, const float* arm, const float* base, int64_t B,
                             uint8_t* valid, void* stream);
/* Motion-edge validation at fixed resolution (VAMP validate_motion, used inside
 * planner_func / simplify, planning/fetch/algorithms.py:37,51): edge e runs from
 * (arm0,base0)[e] to (arm1,base1)[e]; n = max(ceil(|delta| * resolution / rake),1)
 * blocks of `rake` samples, t_k = k/(rake n), k = 1..rake n.  valid: DEVICE E bytes.
 * n_checks (DEVICE int64[1] or NULL) receives sum_e rake*n_e.                    */
int tg_check_edges(tg_env* env, const float* arm0, const float* base0, const float* arm1,
                   const float* base1, int64_t E, int rake, float resolution, uint8_t* valid,
                   int64_t* n_checks, void* stream);

/* Verdict bytes <-> bits (bit i & 7 of byte i / 8 = verdict i): a sharded job gathers 1/8 of the bytes.  All DEVICE;
 * bits holds ceil(n / 8) bytes.                                                                                   */
int tg_pack_bits(const uint8_t* valid, int64_t n, uint8_t* bits, void* stream);
int tg_unpack_bits(const uint8_t* bits, int64_t n, uint8_t* valid, void* stream);

/* HOST-buffer variants (the reference-facing calls: numpy in, numpy out).  They
 * stage through pinned memory in chunks, overlap copies with the kernels, and
 * return after the results are in `valid` (synchronous).                        */
int tg_check_configs_host(tg_env* env, const float* arm, const float* base, int64_t B,
                          uint8_t* valid, float* clearance);
int tg_check_edges_host(tg_env* env, const float* arm0, const float* base0, const float* arm1,
                        const float* base1, int64_t E, int rake, float resolution,
                        uint8_t* valid, int64_t* n_checks);

/* counters of the last check call on this env (HOST out, 4 int64): phase-1
 * definite collisions, ambiguous items resolved by the exact narrow phase,
 * queue overflows handled inline, kernels launched.                              */
int tg_env_counters(tg_env* env, int64_t* out4, void* stream);

/* Narrow-phase work counters of a -DTG_STATS profiling build (HOST out, 16 int64; all zero in
 * the product build): [1] spheres listed, [2] spheres walked, [3] row batches, [4] 32-point scan
 * steps, [5..7] walked spheres by radius class, [8] hits, [9] candidate points.              */
int tg_debug_stats(int64_t* out16, int reset);

/* Device-time instrumentation for benchmarking: when on, tg_check_configs brackets its
 * phase-1 (FK + grid classification) and phase-2 (exact narrow phase) launches with CUDA
 * events on the launching stream.  tg_env_timing_read waits for them and returns
 * {phase-1 ms, phase-2 ms, launch pairs} accumulated since the last read (HOST out, 3 doubles). */
int tg_env_set_timing(tg_env* env, int on);
/* Test / tuning hook: the configuration kernel proves most spheres of a free configuration free from a coarse far-field
 * distance grid held in shared memory (used from 2^18 configurations per launch).  0 switches that variant off; the
 * verdicts do not depend on it.                                                                                    */
int tg_env_set_coarse(tg_env* env, int on);
/* Test hook: use only the first `items` entries of the ambiguous queue (0 = all) so that the
 * overflow path (items resolved inline by their own lane) can be exercised with small batches. */
int tg_env_set_queue_limit(tg_env* env, int64_t items);
int tg_env_timing_read(tg_env* env, double* out3);

/* ---- robot self-filter (kinematics/robot_filter.py:13-64) ---------------------- */
/* keep[i] = 1 iff no centre c has |p_i - c|^2 < radius^2 (f32, no FMA contraction:
 * bit-identical to the reference).  points N x 3, centres S x 3, keep N bytes,
 * all DEVICE.                                                                    */
int tg_filter_points(const float* points, int64_t N, const float* centres, int S, float radius,
                     uint8_t* keep, void* stream);
/* Batched depth self-filter: for every robot state s (states S x 13 DEVICE f32:
 * x, y, theta, q0..q7, head_pan, head_tilt) back-project depth[s] (S x H x W DEVICE
 * f32, stride-decimated like utils/depth.py:32-49 with K HOST 3x3 f64), transform by
 * the head-camera pose (kinematics/fetch/camera.py:25-58) and test against the 111
 * spheres at `radius`.  mask: DEVICE S x H' x W' bytes: 1 = keep (valid depth in
 * [zmin, zmax] and outside the robot), 0 = dropped.  world (DEVICE S x H' x W' x 3
 * f32 or NULL) receives the world points.                                        */
int tg_self_filter_depth(const float* depth, int H, int W, int stride, const double* K,
                         const float* states, int64_t S, float radius, float zmin, float zmax,
                         uint8_t* mask, float* world, void* stream);

/* ---- device-resident ingest and belief merge (belief_state/pointcloud_scene.py:92-268) ------ */
/* Order-preserving compaction: out receives the points whose flag byte is non-zero, in input
 * order (numpy boolean-mask order); *count (DEVICE int) the number kept (may exceed capacity:
 * only the first `capacity` are written).  All pointers DEVICE.                              */
int tg_compact_points(const float* points, const uint8_t* flags, int64_t N, float* out,
                      int64_t capacity, int* count, void* stream);
/* First point of every voxel over vstack(base, add), voxel key floor(p / voxel) evaluated in f32
 * (utils/pointcloud.py:13-64).  keep_add[j] = 1 iff add[j] is the first point of a voxel no base
 * point falls in.  voxel_downsample(points) is the nb = 0 case; merge_dedup(base, add) keeps all
 * of base plus the flagged add points.  Points with |coordinate / voxel| >= 2^20 or non-finite
 * are never merged (kept) and counted in *unpackable (DEVICE int or NULL).                     */
int tg_voxel_first(const float* base, int64_t nb, const float* add, int64_t na, float voxel,
                   uint8_t* keep_add, int* unpackable, void* stream);
/* One depth frame -> filtered world cloud, fused: back-projection (utils/depth.py:13-84, stride-
 * decimated, K HOST 3x3 f64, T_wc HOST 4x4 f64 camera-to-world), z range, XY crop around the camera
 * position (crop_radius < 0: off), ground filter z > ground_z (NaN: off), robot self-filter against
 * `centres` (DEVICE n x 3 f32, strict d2 < r2 removes; n = 0: off).  out_points DEVICE capacity x 3
 * in row-major pixel order, *count DEVICE int.                                                   */
int tg_ingest_depth(const float* depth, int H, int W, int stride, const double* K, const double* T_wc,
                    float zmin, float zmax, float crop_radius, float ground_z, const float* centres,
                    int n_centres, float robot_radius, float* out_points, int64_t capacity, int* count,
                    void* stream);
/* Carve flags for the accumulated map: mode 0 = points_in_frustum_mask (utils/pointcloud.py:93-133),
 * mode 1 = free-space ray test against `depth` (height x width DEVICE f32,
 * pointcloud_scene.py:218-262).  keep[i] = 0 for the points the reference removes.               */
int tg_carve_flags(const float* points, int64_t N, int mode, const double* K, const double* T_wc,
                   float zmin, float zmax, int width, int height, const float* depth,
                   float merge_radius, uint8_t* keep, void* stream);

/* ---- callers next to the path --------------------------- */
/* fp64 FK: q10 DEVICE S x 10 (torso, 7 arm joints, head pan, head tilt) -> frames DEVICE
 * S x 15 x 12 (3x4 row-major, base_link frame).  Replaces kinematics/fetch/fk_numpy.py:32-157
 * for the occlusion check below.                                                        */
int tg_fk_frames_f64(const double* q10, int64_t S, double* frames, void* stream);
/* Batched self-occlusion check: replaces checker/occlusion.py:48-116
 * (check_self_occlusion(link_poses, target_center, target_radius, intrinsics, image_shape)).
 * occluded[s] = 1 iff the cv2.circle discs of the robot spheres and of the target sphere
 * share a pixel of the H x W image.  target DEVICE n_targets x 4 f64 (x, y, z, r in
 * base_link; n_targets = 1 or S), K HOST 3x3 f64.  `links_present` bit i = link i of the
 * 15-link table is in the caller's dict (missing head_tilt_link -> 0 everywhere).       */
int tg_check_occlusion_frames(const double* frames, uint32_t links_present, const double* target,
                              int64_t n_targets, int64_t S, const double* K, int H, int W,
                              uint8_t* occluded, void* stream);
int tg_check_occlusion(const double* q10, const double* target, int64_t n_targets, int64_t S,
                       const double* K, int H, int W, uint8_t* occluded, void* stream);
/* first[g] = index within group g (consecutive `group` entries of valid[B]) of the first
 * non-zero verdict, or -1: the serial "first grasp x IK attempt that passes" of
 * subgoal_generator/samplers/grasp_sampler.py:109-137 after one tg_check_configs call.  */
int tg_first_valid(const uint8_t* valid, int64_t B, int group, int32_t* first, void* stream);

/* ---- device-resident planners (replace the VAMP planners TyGrit calls in
 *      planning/fetch/algorithms.py:23-232).  Configuration vectors are 11 floats [x, y, theta, q0..q7]; theta is an
 *      angle (shortest arc).  Trees, lattice, roadmap, candidate edges and verdicts stay in device memory; the host
 *      enqueues kernels and polls a status word in mapped pinned memory; a finished path is the only device->host copy. */
typedef struct tg_planner tg_planner;
int tg_planner_create(tg_planner** out, tg_env* env, int max_nodes, int batch);
int tg_planner_destroy(tg_planner* p);
/* plan_arm_rrtc / planner_func(start, goal, env, settings, sampler) (algorithms.py:23-37): batched RRT-Connect with a
 * Halton sampler (vamp_preview.py:58-59).  start, goal, lo, hi: HOST 11 floats; a dimension with lo == hi is not sampled
 * (arm-only planning fixes the base).  path_out: HOST max_path x 11.  *n_path = 0: no path within max_iters / timeout. */
int tg_planner_rrt_connect(tg_planner* p, const float* start, const float* goal, const float* lo, const float* hi, float range,
                           int max_iters, uint32_t halton_skip, int rake, float resolution, double timeout_s, float* path_out,
                           int max_path, int* n_path, int* iterations, void* stream);
/* module.simplify / whole_body_simplify (algorithms.py:51,124): all waypoint pairs validated in one batch, then the
 * shortest path through the valid skips; an edge on which the base would drive backwards costs reverse_weight x its
 * length (1 = plain length).  path: HOST n x 11 in/out (n <= 128).                                                 */
int tg_planner_shortcut(tg_planner* p, float* path, int* n_io, int rake, float resolution, float reverse_weight, void* stream);
/* Base layer of multilayer_rrtc (algorithms.py:77-147, HybridAStarConfig.reverse_penalty = 50): SE(2) lattice search
 * (grid `res` metres x 16 headings; forward / backward x reverse_penalty / turn in place), cost-to-go by parallel
 * relaxation, lazy whole-body edge validation with the arm held at arm8.  start, goal: HOST [x, y, theta]; bounds4 HOST
 * {x0, x1, y0, y1}; path_out HOST max_path x 3; stats3 HOST {sweeps, validation rounds, lattice nodes} or NULL.      */
int tg_planner_base_path(tg_planner* p, const float* start, const float* goal, const float* arm8, const float* bounds4, float res,
                         float reverse_penalty, float turn_cost, int rake, float resolution, int max_rounds, float* path_out,
                         int max_path, int* n_path, double* stats3, void* stream);
/* fcit_wb (algorithms.py:155-232: FCITSettings batch_size / max_samples / max_iterations / reverse_weight, XY bounds):
 * fully connected roadmap over Halton samples: every pair of valid samples closer than max_edge is validated in ONE
 * tg_check_edges call, then the cost-to-go is relaxed in parallel.  max_rounds bounds the relaxation sweeps.  stats4 HOST
 * {samples drawn, roadmap nodes, edges validated, relaxation sweeps} or NULL.                                        */
int tg_planner_fcit(tg_planner* p, const float* start, const float* goal, const float* lo, const float* hi, int batch_size,
                    int max_samples, int max_rounds, float reverse_weight, float max_edge, uint32_t halton_skip, int rake,
                    float resolution, double timeout_s, float* path_out, int max_path, int* n_path, double* stats4, void* stream);

/* ---- receding-horizon step as one CUDA graph (BASELINE config 5) ------------------- */
/* One scheduler step of the reference = update_environment (full rebuild, vamp_preview.py:183-240) + planning, i.e.
 * edge validation (planning/fetch/algorithms.py:37,51).  tg_c5_run captures [device-generated fresh edges ->
 * tg_env_build(clouds[i]) -> tg_check_edges] once per cloud buffer as a CUDA graph (fixed bounds required:
 * tg_env_set_bounds) and replays graph[i % n_clouds] for `steps` steps, timing with CUDA events from C - no host loop
 * in Python, no per-kernel launch overhead.  clouds: HOST array of n_clouds DEVICE pointers (n x 3 f32) with their point
 * counts; E arm-only edges per step at random bases inside base_box4 = {x0, x1, y0, y1} (HOST).  out8 (HOST doubles):
 * ms per step, ms per build-only replay, valid-edge fraction, configuration checks per step, graphs captured, nodes
 * per step graph, min / max step ms (sampled every 10th step).  Synchronous.  The blob header of the env is not
 * meaningful afterwards (rebuild before exporting).                                                              */
int tg_c5_run(tg_env* env, const float* const* clouds, const int64_t* n_points, int n_clouds, int64_t E, int steps,
              int warmup, uint64_t seed, const float* base_box4, int rake, float resolution, double* out8);

/* ---- synthetic input (benchmark support, not part of the reference surface) ----- */
/* Analytic head-camera depth frames: rays against the 8 x 6 x 2.6 m room shell, `nb`
 * furniture boxes (DEVICE nb x 6: lo xyz, hi xyz; nb <= 64) and the robot's own 111
 * spheres, ~2 % pixels zeroed.  states DEVICE S x 13, K HOST 3x3 f64, depth DEVICE
 * S x H x W f32.  Stands in for the ManiSkill/SAPIEN renderer (no graphics stack).   */
int tg_synth_depth(const float* states, int64_t S, const float* boxes, int nb, int H, int W,
                   const double* K, float reserved, uint64_t seed, float* depth, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TYGRIT_B200_H */
