/*
 * sspp_gpu.h — C ABI of the B200 geometric-query engine behind the sspp planner API.
 *
 * One data-parallel hot path of kucars/sspp lives behind this boundary: batched segment / ray
 * queries against the occupancy map (edge connection with swept-robot collision checks, node
 * validity, coverage visibility).  Everything is `extern "C"`, plain pointers and sizes; every call
 * returns an int status (0 = SSPP_OK) and never throws.  The text of the last failure is kept per
 * context (sspp_gpu_last_error).  A context is bound to one CUDA device and one stream and is NOT
 * thread-safe — the same contract as the reference, whose callbacks are serialised by ros::spin()
 * (reference src/reactive_planner_server_node.cpp:42).
 *
 * Citations are `file:line` in the reference tree (kucars/sspp) unless marked "upstream", which
 * means the un-vendored third-party code the reference calls (OctoMap, volumetric_mapping
 * octomap_world; reference package.xml:26-27).
 *
 * Pointer conventions: unless a name ends in `_device`, pointers are HOST pointers and the call
 * copies in/out itself (this is the path `bench.py` times as e2e).  `_device` variants take CUDA
 * device addresses valid on the context's device and leave results resident.
 */
#ifndef SSPP_GPU_H
#define SSPP_GPU_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sspp_gpu_ctx sspp_gpu_ctx;

enum {
  SSPP_OK = 0,
  SSPP_ERR_INVALID = 1, /* bad argument (null pointer, negative size, non-finite value) */
  SSPP_ERR_CUDA = 2,    /* CUDA runtime failure; text in sspp_gpu_last_error */
  SSPP_ERR_STATE = 3,   /* call made before its prerequisite (no map, no nodes, no result) */
  SSPP_ERR_RANGE = 4    /* size beyond what the engine can index (int32 columns, key range) */
};

/* volumetric_mapping::WorldBase::CellStatus (upstream), as compared at src/distance_heuristic.cpp:111 */
enum { SSPP_CELL_FREE = 0, SSPP_CELL_OCCUPIED = 1, SSPP_CELL_UNKNOWN = 2 };

/* ---- context ------------------------------------------------------------------------------ */

/* Replaces: construction of volumetric_mapping::OctomapManager (src/reactive_planner_server_node.cpp:40)
 * as the world model handed to DistanceHeuristic (include/sspp/distance_heuristic.h:37). */
int sspp_gpu_create(sspp_gpu_ctx** out, int device);
void sspp_gpu_destroy(sspp_gpu_ctx* ctx);
const char* sspp_gpu_last_error(const sspp_gpu_ctx* ctx);
/* CUDA stream the context launches on (a cudaStream_t), for callers that time with CUDA events. */
void* sspp_gpu_stream(sspp_gpu_ctx* ctx);
/* Blocks until everything queued on the context's stream has finished. */
int sspp_gpu_synchronize(sspp_gpu_ctx* ctx);

/* ---- map ingest ---------------------------------------------------------------------------
 * The octomap is flattened into a dense two-plane bit grid (known, occupied) over the key window
 * [key_min, key_min + dims) plus a derived "blocked" plane and a 4x4x4 any-blocked pyramid.
 * Keys are OctoMap keys: floor(coord / resolution) + 32768 (upstream OcTreeBaseImpl::coordToKey). */

/* Allocates an all-unknown map.  dims are rounded up internally (x to 128, y/z to 4).
 * Replaces: OctomapWorld construction with `resolution` (launch/reactive_planner_test.launch:14,25)
 * and `treat_unknown_as_occupied` (launch/reactive_planner_test.launch:29). */
int sspp_gpu_map_create(sspp_gpu_ctx* ctx, double resolution, const int32_t key_min[3], const int32_t dims[3],
                        int treat_unknown_as_occupied);
int sspp_gpu_map_set_treat_unknown(sspp_gpu_ctx* ctx, int treat_unknown_as_occupied);

/* Leaf boxes (centre xyz, edge length) as OctomapManager::getAllOccupiedBoxes returns them
 * (src/reactive_planner_server.cpp:60; include/sspp/reactive_planner_server.h:61) — and the free
 * leaves in the same format.  occupied[i] != 0 marks box i occupied, 0 marks it known-free. */
int sspp_gpu_map_upload_boxes(sspp_gpu_ctx* ctx, const double* centres_xyz, const double* sizes,
                              const uint8_t* occupied, int64_t n);

/* Replaces: OctomapManager::setFree(position, box) (src/reactive_planner_server.cpp:42) and its twin
 * setOccupied: every voxel met by a resolution-spaced lattice over the box grown by 0.001. */
int sspp_gpu_map_set_box(sspp_gpu_ctx* ctx, const double position[3], const double box[3], int occupied);

/* Sensor model of the incremental map update: OctomapParameters of volumetric_mapping (upstream);
 * the defaults are that library's (0.65, 0.4, 0.12, 0.97, 0.7). */
typedef struct sspp_sensor_model {
  double probability_hit, probability_miss, threshold_min, threshold_max, threshold_occupancy;
} sspp_sensor_model;
int sspp_gpu_map_set_sensor_model(sspp_gpu_ctx* ctx, const sspp_sensor_model* model);

/* Replaces: OctomapManager::insertPointcloudWithTf (src/reactive_planner_server.cpp:238) after its TF
 * step, i.e. OctomapWorld::insertPointcloudIntoMapImpl + castRay + updateOccupancy (upstream) for points
 * already in the world frame: one ray per point from the sensor origin (free cells along it, the end
 * cell occupied; beyond sensor_max_range only the free part up to the range), a point whose voxel an
 * earlier point of the same cloud already marked occupied casts no ray, every voxel updated at most
 * once per cloud (occupied wins), log-odds += hit / miss with clamping.  sensor_max_range < 0 = unlimited.
 * Keeps a float log-odds grid next to the bit planes (allocated on first use).  Cells outside the
 * map window are dropped. */
int sspp_gpu_map_insert_pointcloud(sspp_gpu_ctx* ctx, const float* points_world_xyz, int64_t n,
                                   const double sensor_origin[3], double sensor_max_range);
/* Log-odds of every voxel of the window (un-rounded dims, x fastest, then y, then z); NaN = never observed. */
int sspp_gpu_map_download_logodds(sspp_gpu_ctx* ctx, float* logodds);

/* Dense bit planes over the window given to map_create (its un-rounded dims), x fastest, rows of
 * ceil(dims[0]/32) 32-bit words, then y, then z.  Bit = (word >> (x & 31)) & 1. */
int sspp_gpu_map_upload_dense(sspp_gpu_ctx* ctx, const uint32_t* occupied_bits, const uint32_t* known_bits);
int sspp_gpu_map_download_dense(sspp_gpu_ctx* ctx, uint32_t* occupied_bits, uint32_t* known_bits);

/* ---- search-space nodes -------------------------------------------------------------------
 * Node positions in LIST ORDER: index 0 is the head of SearchSpace::searchspace
 * (include/sspp/searchspace.h:37; head insertion at src/searchspace.cpp:98-101).  xyz is n x 3
 * doubles (geometry_msgs::Pose::position of SearchSpaceNode::location,
 * include/sspp/searchspacenode.h:38). */
int sspp_gpu_set_nodes(sspp_gpu_ctx* ctx, const double* xyz, int64_t n);
int sspp_gpu_set_nodes_device(sspp_gpu_ctx* ctx, const double* xyz_device, int64_t n);

/* ---- edge connection ----------------------------------------------------------------------
 * Replaces: PathPlanner::connectNodes (src/pathplanner.cpp:527-567) [position_mode 0] and
 * PathPlanner::connectClustersInternalNodes (src/pathplanner.cpp:612-656) [position_mode 1],
 * with the per-pair test DistanceHeuristic::isConnectionConditionSatisfied
 * (src/distance_heuristic.cpp:99-127) -> OctomapWorld::getLineStatusBoundingBox (upstream).
 * Result: CSR adjacency over list indices, rows and columns both in list order. */
typedef struct sspp_connect_params {
  double radius;         /* regGridConRadius (include/sspp/pathplanner.h:60); pairs with Dist <= radius */
  double box[3];         /* swept robot box; the reference hard-codes (0.5,0.5,0.3), distance_heuristic.cpp:107 */
  int32_t use_map;       /* 0 = the manager==NULL case: every in-radius pair accepted (distance_heuristic.cpp:101,126) */
  int32_t position_mode; /* 0: exact != on x,y,z (pathplanner.cpp:545); 1: isPositionEqual 1e-12 (pathplanner.cpp:630) */
  int64_t row_begin;     /* source rows [row_begin,row_end) of the list; 0,0 = all rows (multi-GPU sharding) */
  int64_t row_end;
  int32_t use_pyramid;   /* 1: rows whose whole reach is free in the coarse pyramid skip the DDA (same result) */
  int32_t count_steps;   /* 1: also count the voxel lookups the reference semantics require (instrumented run:
                            disables the pyramid skip; for the byte model of the roofline, not for timing) */
  int32_t count_only;    /* 1: stop after the neighbour count (stats: rows, in_radius_pairs, tested_pairs, pyramid_rows);
                            no edge test, no CSR — what sspp_gpu_balanced_cuts runs */
  int32_t reserved;      /* 0 */
} sspp_connect_params;

typedef struct sspp_connect_stats {
  int64_t rows;            /* source rows processed */
  int64_t edges;           /* accepted directed edges E (MAXNODES, pathplanner.cpp:566) */
  int64_t in_radius_pairs; /* ordered pairs with Dist <= radius, self excluded */
  int64_t tested_pairs;    /* of those, pairs at different positions = pairs handed to the collision test (pathplanner.cpp:547) */
  int64_t rays_per_pair;   /* size of the ray lattice nx*ny*nz */
  int64_t pyramid_rows;    /* rows answered by the pyramid */
  int64_t tile_rows;       /* reserved, always 0 (kept so the struct layout does not change) */
  int64_t box_pairs;       /* tested pairs accepted because the key box they sweep is free */
  int64_t lattice_pairs;   /* tested pairs that ran the ray lattice */
  int64_t voxel_steps;     /* with count_steps: lookups up to and including the first non-free voxel of the first
                              failing ray (lattice order) of every tested pair — what getLineStatusBoundingBox performs */
} sspp_connect_stats;

/* Runs the whole connection pass on the device; the CSR stays resident until the next connect. */
int sspp_gpu_connect(sspp_gpu_ctx* ctx, const sspp_connect_params* params, sspp_connect_stats* stats);
/* Copies the last CSR to host: row_ptr has rows+1 entries (relative to row_begin), col_idx has `edges`. */
int sspp_gpu_connect_fetch(sspp_gpu_ctx* ctx, int64_t* row_ptr, int32_t* col_idx);
/* Diagnostics for the parity tests (no counterpart in the reference): how each row of the last connect
 * was answered, one byte per row — 0 the row has no in-radius pair; 1 every pair accepted because the
 * row's whole reach is free in the coarse pyramid (or no map is used); 2 reserved; 3 the row was decided
 * pair by pair (start voxels, pair key boxes, ray lattice).  The tests use it to draw oracle checks from
 * every class. */
int sspp_gpu_connect_row_classes(sspp_gpu_ctx* ctx, uint8_t* row_class);
/* Same, into DEVICE buffers of the caller (e.g. the send buffers of an NCCL all-gather). */
int sspp_gpu_connect_fetch_device(sspp_gpu_ctx* ctx, int64_t* row_ptr_device, int32_t* col_idx_device);
/* Multi-GPU assembly (SURVEY.md §8 e): after an all-gather of every rank's padded blocks —
 * row_ptr blocks of `row_stride` int64 (entry [rows_g] holds the block's edge count) and column blocks
 * of `col_stride` int32 — rebases and packs them into the whole-graph CSR on the context's stream, no
 * host round trip.  row_base_device: world+1 first-row indices.  *overflow_device receives the largest
 * edge count that did not fit `col_stride` (0 = all fitted); the caller then regathers with larger slots.
 * The concatenation equals what PathPlanner::connectNodes (src/pathplanner.cpp:527-567) builds in one piece. */
int sspp_gpu_assemble_csr_blocks(sspp_gpu_ctx* ctx, const int64_t* row_ptr_blocks_device, const int32_t* col_blocks_device,
                                 int world, int64_t row_stride, int64_t col_stride, const int64_t* row_base_device,
                                 int64_t* row_ptr_out_device, int32_t* col_out_device, int64_t* overflow_device);
/* Multi-GPU assembly through NVLink peer memory — the compaction result and its all-gather as one
 * kernel (no counterpart in the reference, which is one process; it realises SURVEY.md §8 e's exchange
 * step).  Every rank owns one whole-graph CSR (row_ptr: total_rows+1 int64, col_idx: col_capacity
 * int32) that its peers can address:
 *   sspp_gpu_peer_buffer_create  cudaMalloc + cudaIpcGetMemHandle: the 64-byte handle is what the ranks
 *                                exchange (any transport: an all-gather of 64 bytes per rank);
 *   sspp_gpu_peer_buffer_open    cudaIpcOpenMemHandle of a peer's handle (another process, same box);
 *   sspp_gpu_peer_buffer_release frees a created buffer / closes an opened one (destroy does it too).
 * sspp_gpu_connect_push_csr then writes this rank's last CSR block, rebased by the edge counts of the
 * lower ranks, into the buffers of all `world` ranks (entry [rank] = this rank's own buffer; host
 * arrays of device pointers) on the context's stream.  edge_counts_device: `world` int64, the
 * all-gathered edge counts of this connect; row_base_device: world+1 first-row indices.  If the
 * graph's edge total exceeds col_capacity nothing is written and *overflow_device receives the total
 * (every rank takes the same decision from the same counts).  The pushes of all ranks are complete —
 * and the CSR may be read — after a barrier that follows this call on every rank; a rank must not start
 * the next push before all ranks have finished reading (the all-gather of the next counts orders that).
 * The result is byte-identical to the CSR of a one-GPU connect over all rows. */
#define SSPP_MAX_PEERS 16
#define SSPP_PEER_HANDLE_BYTES 64
int sspp_gpu_peer_buffer_create(sspp_gpu_ctx* ctx, int64_t bytes, void** device_ptr, uint8_t handle[SSPP_PEER_HANDLE_BYTES]);
int sspp_gpu_peer_buffer_open(sspp_gpu_ctx* ctx, const uint8_t handle[SSPP_PEER_HANDLE_BYTES], void** device_ptr);
int sspp_gpu_peer_buffer_release(sspp_gpu_ctx* ctx, void* device_ptr);
int sspp_gpu_connect_push_csr(sspp_gpu_ctx* ctx, int world, int rank, const int64_t* edge_counts_device,
                              const int64_t* row_base_device, int64_t* const* row_ptr_peers, int32_t* const* col_idx_peers,
                              int64_t col_capacity, int64_t* overflow_device);
/* Multi-GPU, transfer overlapped with the edge test: sspp_gpu_connect, with the edge test and the
 * compaction run in row chunks (sizes 4:3:2:1 above 65 536 rows) and every finished chunk stored through
 * NVLink peer pointers (sspp_gpu_peer_buffer_open) into slot [rank] of EVERY rank's block buffers by a small
 * push kernel on a second, high-priority stream while the next chunk's edge kernel runs (which leaves half
 * the SMs one CTA slot free for it).  The push kernel takes the chunk's edge range from device memory, so the
 * host queues the whole pass without waiting for anything in between.  Block buffers as
 * sspp_gpu_assemble_csr_blocks reads them: row_ptr blocks of row_stride int64 per rank (relative to the
 * block, entry [rows] = the block's edge count), column blocks of col_stride int32 per rank (col_stride a
 * multiple of 4, buffers 16-byte aligned).  Returns when this rank's stores have been issued and retired;
 * after a barrier over the ranks, sspp_gpu_assemble_csr_blocks on the local buffers yields the whole-graph CSR.
 * SSPP_ERR_RANGE when the block's edges exceed col_stride (stats->edges = size needed): the chunks that
 * still fitted HAVE been sent, so the peers' slots then hold a partial block — the caller must not pack
 * them, but agree on larger slots with its peers (a MAX all-reduce of stats->edges) and run the pass again.
 * No counterpart in the reference (one process). */
typedef struct sspp_peer_blocks {
  int32_t world, rank;
  int64_t* const* row_ptr_blocks; /* host array of `world` device addresses: rank p's row_ptr block buffer */
  int32_t* const* col_blocks;     /* likewise, column block buffer */
  int64_t row_stride, col_stride;
} sspp_peer_blocks;
int sspp_gpu_connect_to_peers(sspp_gpu_ctx* ctx, const sspp_connect_params* params, const sspp_peer_blocks* peers,
                              sspp_connect_stats* stats);
/* Multi-GPU, exchange FUSED into the edge kernel (no counterpart in the reference; SURVEY.md §8 e's exchange step done
 * while the edge test is still running).  Every rank owns, per buffer set, a slot-format image of the WHOLE graph that
 * its peers can address: `slots` (total_rows x slot int32) and `deg` (total_rows int32).  sspp_gpu_connect_fused runs this
 * rank's rows of connectNodes with every row — rows whose reach is free in the pyramid included — going through the
 * edge kernel, whose warps store each finished row (accepted columns, already in list order, and the row's degree)
 * straight into row first_row + r of ALL ranks' images through NVLink peer pointers.  The transfer so rides along
 * with the computation row by row; nothing is left to send when the pass ends, and no local CSR is built
 * (sspp_gpu_connect_fetch is not available after it; stats->edges is the shard's edge count).  After a barrier over
 * the ranks, sspp_gpu_pack_slots turns a complete image into the whole-graph CSR (scan of the degrees + copy), which is
 * byte-identical to the CSR of a one-GPU connect.  Needs use_map = 1 and the pair-slot layout: SSPP_ERR_RANGE when a
 * row of this shard holds more than `slot` in-radius pairs (fall back to sspp_gpu_connect + sspp_gpu_connect_push_csr).
 * Alternate two buffer sets from call to call: a rank may already be writing the next graph into its peers' images
 * while they still pack the last one. */
typedef struct sspp_fused_peers {
  int32_t world, rank;
  int32_t slot;      /* 32 or 64: ints per row in the images (the same on every rank) */
  int32_t reserved;  /* 0 */
  int64_t first_row; /* global index of this rank's row 0 (= params->row_begin when the ranks share one node list) */
  int32_t* const* slot_peers; /* host array of `world` device addresses: rank p's `slots` image */
  int32_t* const* deg_peers;  /* likewise, rank p's `deg` image */
} sspp_fused_peers;
int sspp_gpu_connect_fused(sspp_gpu_ctx* ctx, const sspp_connect_params* params, const sspp_fused_peers* peers,
                           sspp_connect_stats* stats);
/* slots_device / deg_device: a complete image (all ranks have passed the barrier); row_ptr_out_device: n_rows + 1 int64;
 * col_out_device: col_capacity int32.  If the graph has more edges than col_capacity nothing is copied and
 * *overflow_device receives the edge count.  Runs on the context's stream. */
int sspp_gpu_pack_slots(sspp_gpu_ctx* ctx, const int32_t* slots_device, const int32_t* deg_device, int64_t n_rows, int32_t slot,
                        int64_t* row_ptr_out_device, int32_t* col_out_device, int64_t col_capacity, int64_t* overflow_device);
/* Candidate-balanced row partition (SURVEY.md §8 e "Partitioning"; the outer loop being split is
 * src/pathplanner.cpp:536): `world` + 1 cut points over the list indices, cuts[0] = 0, cuts[world] = n, such that
 * every block [cuts[g], cuts[g+1]) carries about the same WORK rather than the same number of rows.  Work of a
 * row = its in-radius pairs, times `work_row_weight` when the row has to go through the edge test (its reach is
 * not free in the coarse pyramid; 8 is the measured cost ratio on the S4 scenes).  Runs the neighbour count over
 * all rows (no edge test), so it is a set-up step per node set, not part of a connect pass; every rank that holds
 * the same nodes and map computes the same cuts, no message needed. */
int sspp_gpu_balanced_cuts(sspp_gpu_ctx* ctx, const sspp_connect_params* params, int world, int work_row_weight, int64_t* cuts);
/* Device addresses of the last CSR (valid until the next connect on this context). */
int sspp_gpu_connect_result_device(sspp_gpu_ctx* ctx, const int64_t** row_ptr_device, const int32_t** col_idx_device);

/* Convenience for the drop-in PathPlanner::connectNodes: set_nodes + connect + fetch in one call,
 * host buffers in, host CSR out (col_idx capacity `col_capacity`; SSPP_ERR_RANGE if too small,
 * with stats->edges telling the size needed). */
int sspp_gpu_connect_nodes_host(sspp_gpu_ctx* ctx, const double* xyz, int64_t n, const sspp_connect_params* params,
                                int64_t* row_ptr, int32_t* col_idx, int64_t col_capacity, sspp_connect_stats* stats);
/* The same on the nodes already set (sspp_gpu_set_nodes / _device): connect + host CSR out.  Above 65 536
 * rows the edge test and the compaction run in six row chunks of falling size and every finished chunk is
 * copied to the host buffers (pin them for the copies to be asynchronous) while the next one runs. */
int sspp_gpu_connect_to_host(sspp_gpu_ctx* ctx, const sspp_connect_params* params, int64_t* row_ptr, int32_t* col_idx,
                             int64_t col_capacity, sspp_connect_stats* stats);

/* ---- node store: batched SearchSpace::insertNode --------------------------------------------
 * Replaces the per-pose list scan of SearchSpace::insertNode / nodeExists (src/searchspace.cpp:69-105,
 * 164-175) with samePosition / isEqual (include/sspp/utils.h:386-390, 231-239): poses (n x 7 doubles:
 * x y z qx qy qz qw) are offered in insertion order; pose i is stored unless an already stored pose
 * equals it in all seven components, fabs(a-b) < tolerance each (the reference's constant is 1E-12).
 * keep[i] = 1 when pose i is stored.  dup_of (may be NULL): -1 for stored poses, else the insertion
 * index of the node insertNode returns (nodeExists walks from the head, i.e. the LATEST stored match).
 * *stored (may be NULL) = number of poses stored.  Same verdicts as the sequential scan, including
 * the non-transitive chains a ~ b ~ c.  Positions must be finite (SSPP_ERR_INVALID otherwise). */
int sspp_gpu_dedup_poses(sspp_gpu_ctx* ctx, const double* poses, int64_t n, double tolerance, uint8_t* keep, int64_t* dup_of,
                         int64_t* stored);

/* ---- batched segment / cell / ray queries --------------------------------------------------- */

/* Replaces: n calls of OctomapManager::getLineStatusBoundingBox(start,end,box) (upstream; called at
 * src/distance_heuristic.cpp:109).  status[i] is the CellStatus of segment i. */
int sspp_gpu_check_segments(sspp_gpu_ctx* ctx, const double* starts_xyz, const double* ends_xyz, int64_t n,
                            const double box[3], uint8_t* status);

/* Allocation-free twin of sspp_gpu_check_segments.  No counterpart in the reference: it exists for
 * callers that must not, or cannot, obtain device memory at query time — a replanning callback that
 * runs after some other component (an inference engine, a memory pool) has claimed the rest of the
 * GPU, or a control loop that cannot afford the device-wide stall a cudaMalloc/cudaFree causes.
 * Same arguments, same results; the difference is where the working memory comes from: the device
 * copies of `starts`/`ends`/`status` and the ray-offset table live in one fixed array compiled into
 * the library (about 0.8 MB, resident on every GPU that loads the library until the process exits),
 * so the call performs no cudaMalloc, cudaFree or other allocating call.
 *   - n may not exceed SSPP_NOALLOC_MAX_SEGMENTS; a larger batch is refused with SSPP_ERR_RANGE
 *     (split it, or use sspp_gpu_check_segments).
 *   - The first call on each device looks up the array's device address, and — the CUDA runtime
 *     loads kernels lazily — is also the call that loads the query kernel onto the device, which
 *     the driver may back with memory of its own.  Make one warm-up call per device at start-up,
 *     before starting any stream capture on that device and before the phase in which the device
 *     must not be asked for memory; from the second call on nothing is loaded or allocated.
 *   - The array is shared by every context on the device and keeps the previous call's contents.
 *     Each call overwrites everything it later reads, so nothing needs clearing between calls, but
 *     two calls whose GPU work overlaps (different contexts or threads) would race on it: serialise them. */
#define SSPP_NOALLOC_MAX_SEGMENTS 16384
int sspp_gpu_check_segments_noalloc(sspp_gpu_ctx* ctx, const double* starts_xyz, const double* ends_xyz, int64_t n,
                                    const double box[3], uint8_t* status);

/* Node validity.  Replaces: the hook Heuristic::isFilteringConditionSatisfied
 * (include/sspp/heuristic_interface.h:44; called at src/pathplanner.cpp:160), which the only
 * heuristic in the reference answers `true` (src/distance_heuristic.cpp:134-138).  Defined here as
 * OctomapWorld::getCellStatusBoundingBox(point, box) in key space (DESIGN.md "validity spec"). */
int sspp_gpu_validity(sspp_gpu_ctx* ctx, const double* points_xyz, int64_t n, const double box[3], uint8_t* status);

/* Replaces: n calls of octomap::OcTree::castRay(origin, direction, end, ignoreUnknown, maxRange)
 * (upstream).  hit[i]=1 when it returned true; hit_keys (n x 3, may be NULL) the key it stopped in;
 * max_range may be NULL (= -1, unlimited). */
int sspp_gpu_cast_rays(sspp_gpu_ctx* ctx, const float* origins_xyz, const float* directions_xyz, int64_t n,
                       int ignore_unknown, const double* max_range, uint8_t* hit, int32_t* hit_keys);

/* ---- coverage visibility ------------------------------------------------------------------
 * Replaces: the per-viewpoint visible-surface extraction a CoverageHeuristic performs inside
 * Heuristic::calculateHeuristic(Node*) (include/sspp/heuristic_interface.h:40; called at
 * src/astar.cpp:131,318) with the sensor model of include/sspp/sensors.h:57-68.  The reference
 * snapshot holds no CoverageHeuristic source; the semantics are the authored spec in DESIGN.md
 * ("visibility spec v1"): six-plane frustum test in float, then one castRay per in-frustum point. */
typedef struct sspp_frustum {
  double hfov_deg, vfov_deg, near_dist, far_dist; /* Sensors::hFOV, vFOV, near, far (include/sspp/sensors.h:61) */
  int32_t count_steps; /* 1: instrumented run, also counts the voxel lookups castRay performs (stats->voxel_steps) */
  int32_t reserved;    /* 0 */
} sspp_frustum;

typedef struct sspp_visibility_stats {
  int64_t visible_pairs; /* total (viewpoint, point) pairs visible */
  int64_t rays;          /* occlusion rays cast = in-frustum pairs */
  int64_t voxel_steps;   /* with count_steps: cells castRay looked up over all rays (the k of SURVEY.md §8 d's byte model) */
} sspp_visibility_stats;

/* Model points (float xyz), kept resident and binned for frustum queries. */
int sspp_gpu_set_model_points(sspp_gpu_ctx* ctx, const float* xyz, int64_t n);
/* sensor_poses: nv x 7 doubles (position xyz, quaternion xyzw) — the poses
 * Sensors::robot2sensorTransformation yields (src/sensors.cpp:47-92). */
int sspp_gpu_visibility(sspp_gpu_ctx* ctx, const double* sensor_poses, int64_t nv, const sspp_frustum* frustum,
                        sspp_visibility_stats* stats);
/* view_ptr: nv+1 entries; point_idx: visible_pairs entries, ascending within a viewpoint. */
int sspp_gpu_visibility_fetch(sspp_gpu_ctx* ctx, int64_t* view_ptr, int32_t* point_idx);
/* Multi-GPU visibility (the hook runs per way-point — include/sspp/heuristic_interface.h:40, src/astar.cpp:318 —
 * and is independent per viewpoint, so it shards like the rows of connectNodes; SURVEY.md §8 e): every rank
 * runs sspp_gpu_visibility on a contiguous block of the viewpoints, with the model points and the map
 * replicated, and this call writes the block — view_ptr rebased by the visible-pair counts of the lower
 * ranks — into the whole-result buffers of all `world` ranks through NVLink peer memory, exactly as
 * sspp_gpu_connect_push_csr does for the adjacency (same arguments, same overflow rule, same barrier
 * contract).  The assembled (view_ptr, point_idx) is byte-identical to one visibility call over all viewpoints. */
int sspp_gpu_visibility_push(sspp_gpu_ctx* ctx, int world, int rank, const int64_t* pair_counts_device,
                             const int64_t* view_base_device, int64_t* const* view_ptr_peers, int32_t* const* point_idx_peers,
                             int64_t idx_capacity, int64_t* overflow_device);
/* Milliseconds the kernels of the last sspp_gpu_visibility spent on the stream (CUDA events):
 * [0] total, [1] point binning (0 when the cached binning was reused), [2] viewpoint set-up + frustum
 * gather (count and fill), [3] the occlusion rays (k_view_rays), [4] compaction + per-viewpoint sort. */
int sspp_gpu_visibility_timings(sspp_gpu_ctx* ctx, float ms[8]);

/* ---- measurement helpers ------------------------------------------------------------------ */
/* Reads `bytes` of an L2-resident buffer `iters` times and returns the achieved GB/s — the
 * denominator north_star asks for ("achievable L2 bandwidth"), not in MEASURED_PEAKS.json. */
int sspp_gpu_measure_l2_bandwidth(sspp_gpu_ctx* ctx, int64_t bytes, int iters, double* gbps);
/* Milliseconds the kernels of the last sspp_gpu_connect spent on the stream (CUDA events):
 * [0] total, [1] binning+scan, [2] count+scan, [3] edge phase (pair fill, accept rows, edge kernel),
 * [4] CSR scan+compaction, [5] the edge kernel (k_edges) alone, [6..7] reserved. */
int sspp_gpu_connect_timings(sspp_gpu_ctx* ctx, float ms[8]);
/* Number of kernels this library has launched on the context since creation. */
int64_t sspp_gpu_launch_count(const sspp_gpu_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* SSPP_GPU_H */
