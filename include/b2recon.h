/* b2recon.h -- C-ABI of libb2recon.so: the B200-native RGB-D reconstruction core.
 *
 * This is the drop-in boundary for the one hot path of mishn-0/AzureKinect3DReconstruction: the
 * Open3D 0.19.0 legacy calls its scripts make (the scripts are pure call sites; the arithmetic is in
 * open3d==0.19.0, pinned at /root/reference/pyproject.toml:15).  Every entry point names the reference
 * call it replaces.  The reference binds this path through Open3D's pybind11 module; the ctypes binding
 * a maintainer would add instead is shown in INTEGRATION.md.
 *
 * Conventions
 *  - All image / point / mesh pointers are DEVICE pointers (torch CUDA tensors) unless the name says
 *    "_host".  4x4 matrices are row-major f64.  Intrinsics are (fx, fy, cx, cy) f64.
 *  - Every call is asynchronous on `stream` unless it returns a count (then it synchronises `stream`).
 *  - Return value: 0 = OK, <0 = error (codes below); b2r_last_error() gives the message (thread-local).
 *  - The caller owns every buffer it passes; the library owns only what lives behind a handle.
 *  - There is no CPU fallback: without a CUDA device every compute call fails with B2R_ERR_CUDA.
 */
#ifndef B2RECON_H
#define B2RECON_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __CUDA_RUNTIME_H__
typedef struct CUstream_st* cudaStream_t;
#endif

#define B2R_OK 0
#define B2R_ERR_BAD_ARG (-1)
#define B2R_ERR_CUDA (-2)
#define B2R_ERR_CAPACITY (-3) /* output / pool too small; the required size is reported */
#define B2R_ERR_FORMAT (-4)   /* upstream utility::LogError("Unsupported image format.") conditions */
#define B2R_ERR_SINGULAR (-5)
#define B2R_ERR_RANGE (-6)

#define B2R_COLOR_NONE 0   /* TSDFVolumeColorType.NoColor */
#define B2R_COLOR_RGB8 1   /* TSDFVolumeColorType.RGB8    */
#define B2R_COLOR_GRAY32 2 /* TSDFVolumeColorType.Gray32  */

typedef struct b2r_tsdf b2r_tsdf;

int b2r_version(void);
const char* b2r_last_error(void);

/* f64 pose algebra shared by every caller (host).  Upstream: Eigen Matrix4d::inverse(), SURVEY A.3. */
int b2r_m4_inverse(const double* in16, double* out16);
void b2r_m4_mul(const double* a16, const double* b16, double* out16);

/* multi-GPU sharding rule (host): rank that owns volume unit (kx,ky,kz) among `world` ranks; -1 if out of range */
int b2r_unit_owner(int kx, int ky, int kz, int world);

/* ---- a1: RGBDImage.create_from_color_and_depth pieces (Depth2PC.py:60-66,
 *      GoodColorReconstruction.py:131-137, RGBD_Odometry.py:15-18).  Normally fused into consumers. */
int b2r_depth_to_float(const uint16_t* depth, int64_t n, float depth_scale, float depth_trunc, float* out,
                       cudaStream_t stream);
/* the other 1-channel depth types upstream's ConvertDepthToFloatImage accepts: dtype 0 = u16, 1 = f32, 2 = u8 (/255 first) */
int b2r_depth_to_float_typed(const void* depth, int dtype, int64_t n, float depth_scale, float depth_trunc, float* out,
                             cudaStream_t stream);
int b2r_rgb_to_intensity(const uint8_t* rgb, int64_t n, int channels, int bgr, float* out, cudaStream_t stream);

/* ---- a3/a4/a5/a6: PointCloud.create_from_rgbd_image / create_from_depth_image (+ fused a1 conversion,
 *      BGRA swizzle, colour-to-depth gather and output transform), batched over B same-shaped frames.
 *      Reference: Depth2PC.py:69-93, BetterVisualizer2CamWorksCOLOR.py:515-541,634-640,
 *      InitialReconstruction-CUDA-2cam:639-646,743-755.
 *  depth        u16 (depth_is_u16=1; scaled by depth_scale, zeroed at >= depth_trunc) or f32 metres
 *  color        u8, 3 or 4 channels, same WxH as depth -- or, when color_K and d2c_host are given, a
 *               Wc x Hc image in its own camera (gather). color_bgr=1 swaps channel 0 and 2. May be NULL.
 *  gray         f32 intensity image (colour of a convert_rgb_to_intensity RGBDImage). May be NULL.
 *  depth_K      HOST (fx,fy,cx,cy); color_K HOST or NULL; d2c_host HOST 16 f64 depth->colour or NULL
 *  T_dev        DEVICE B*16 f64: camera -> output transform (= extrinsic^-1, optionally pre-multiplied
 *               by the flip / per-camera pose)
 *  out_f64      1: xyz/rgb are f64 and follow the upstream f64 chain exactly; 0: f32 throughput path
 *  valid_only   project_valid_depth_only
 *  xyz,rgb      [B][capacity][3]; src_index / gather_index (optional) [B][capacity] int32: raster index
 *               of the source depth pixel / gathered colour texel (-1 = outside the colour image)
 *  counts       DEVICE int32[B]; scratch DEVICE int32[b2r_unproject_scratch_words(...)], 8-byte aligned
 *  The f32 path is ONE kernel (tiles in raster order by ticket, offsets by decoupled look-back); the f64 path is
 *  count -> scan -> fill.  The gather chain (a6) is the f32 FMA chain defined in oracle/oracle.cpp "a6".  */
int64_t b2r_unproject_scratch_words(int W, int H, int B, int stride, int out_f64);
int b2r_unproject(const void* depth, int depth_is_u16, float depth_scale, float depth_trunc, int W, int H, int B,
                  const uint8_t* color, int color_channels, int color_bgr, int Wc, int Hc, const float* gray,
                  const double* depth_K, const double* color_K, const double* d2c_host, const double* T_dev,
                  int stride, int out_f64, int valid_only, void* xyz, void* rgb, int32_t* src_index,
                  int32_t* gather_index, int64_t capacity, int32_t* counts, int32_t* scratch, cudaStream_t stream);

/* pcd.transform(T) / mesh.transform(T)  (Depth2PC.py:87-93, bundlefusion_reconstruction.py:323-324) */
int b2r_transform_points(void* xyz, void* normals, int64_t n, int is_f64, const double* T_host, cudaStream_t stream);

/* ---- a7: ScalableTSDFVolume(voxel_length, sdf_trunc, color_type, volume_unit_resolution=16,
 *      depth_sampling_stride=4)  (GoodColorReconstruction.py:111-115, bundlefusion_reconstruction.py:108-112).
 *      capacity_units: size of the zeroed unit pool (80 KiB per RGB8 unit).  rank/world: this handle
 *      keeps only the units whose key hashes to `rank`; 0/1 for a single GPU. */
b2r_tsdf* b2r_tsdf_create(double voxel_length, double sdf_trunc, int color_type, int unit_resolution,
                          int depth_sampling_stride, int64_t capacity_units, int rank, int world);
void b2r_tsdf_destroy(b2r_tsdf* h);
int b2r_tsdf_reset(b2r_tsdf* h, cudaStream_t stream); /* volume.reset(), bundlefusion_reconstruction.py:309 */
double b2r_tsdf_voxel_length(b2r_tsdf* h);
double b2r_tsdf_sdf_trunc(b2r_tsdf* h);
int b2r_tsdf_color_type(b2r_tsdf* h);
int64_t b2r_tsdf_capacity(b2r_tsdf* h);

/* ---- a8+a9: volume.integrate(rgbd, intrinsic, extrinsic)  (GoodColorReconstruction.py:270,
 *      BetterVisualizer2CamWorksCOLOR.py:319, capture_frames.py:157, WORKSkinda.py:546-551).
 *      extrinsic: HOST 16 f64 world->camera.  Images as in b2r_unproject (aligned colour only). */
int b2r_tsdf_integrate(b2r_tsdf* h, const void* depth, int depth_is_u16, float depth_scale, float depth_trunc,
                       const uint8_t* color, int color_channels, int color_bgr, const float* gray, int W, int H,
                       double fx, double fy, double cx, double cy, const double* extrinsic_host, cudaStream_t stream);
/* n frames stored back to back (the reset + re-integrate loop, bundlefusion_reconstruction.py:309-315) */
int b2r_tsdf_integrate_batch(b2r_tsdf* h, int n, const void* depth, int depth_is_u16, float depth_scale,
                             float depth_trunc, const uint8_t* color, int color_channels, int color_bgr, int W,
                             int H, double fx, double fy, double cx, double cy, const double* extrinsics_host,
                             cudaStream_t stream);
/* n same-shaped frames NOT stored back to back: HOST arrays of n DEVICE pointers, fused in array order -- the slots of
 * a per-tick all-gather buffer taken tick-major (config 5: several cameras into one volume,
 * BetterVisualizerD2camerasWORKSkinda.py:546-551 integrates camera 0 then camera 1 of every tick) */
int b2r_tsdf_integrate_frames(b2r_tsdf* h, int n, const void* const* depth_ptrs, int depth_is_u16, float depth_scale,
                              float depth_trunc, const uint8_t* const* color_ptrs, int color_channels, int color_bgr,
                              int W, int H, double fx, double fy, double cx, double cy,
                              const double* extrinsics_host, cudaStream_t stream);
/* same with HOST images: H2D copies overlap fusion through an internal 6-deep device ring of 8-frame chunks
 * (env B2R_HOST_CHUNK).  The host buffers must be fully written when the call is made; they may be reused or freed
 * as soon as it returns (the call waits for its last copy, not for the kernels). */
int b2r_tsdf_integrate_host(b2r_tsdf* h, int n, const void* depth_host, int depth_is_u16, float depth_scale,
                            float depth_trunc, const uint8_t* color_host, int color_channels, int color_bgr, int W,
                            int H, double fx, double fy, double cx, double cy, const double* extrinsics_host,
                            cudaStream_t stream);

/* Frames are packed and their blocks allocated when integrate is called; the per-voxel update of up to
 * max_batch (default 32, env B2R_TSDF_BATCH, <= 32) pending frames runs as ONE pass over the touched
 * units when the batch is full or when anything reads the volume (stats, export, extract, ...).  Results
 * are bit-identical to frame-by-frame integration.  b2r_tsdf_flush forces the pass; b2r_tsdf_set_batch(1)
 * gives strictly per-call fusion. */
int b2r_tsdf_flush(b2r_tsdf* h, cudaStream_t stream);
int b2r_tsdf_set_batch(b2r_tsdf* h, int max_batch, cudaStream_t stream);

/* out[0]=units allocated, out[1]=error bits, out[2]=units touched by the last fused batch,
 * out[3]=cumulative updated voxels (N_upd), out[4]=cumulative (unit, frame) pairs fused.  Synchronises. */
int b2r_tsdf_stats(b2r_tsdf* h, int64_t* out5_host, cudaStream_t stream);
int64_t b2r_tsdf_num_units(b2r_tsdf* h, cudaStream_t stream);
/* per-kernel CUDA-event timing for the roofline line of bench.py: enable for up to max_launches kernel
 * launches (0 = off); read: out5 = {frames packed, total ms pack+alloc kernel, total ms fuse kernel,
 * pack+alloc launches, fuse launches} */
int b2r_tsdf_profile(b2r_tsdf* h, int max_launches);
int b2r_tsdf_profile_read(b2r_tsdf* h, double* out5_host, cudaStream_t stream);
/* Host-only: bytes c for which the f32 unprojection path's two-term constant misrounds c/255 (must return 0). */
int b2r_selftest_u8_div255(void);
/* GPU self-test of the shared-reciprocal IEEE division used by the fusion kernel: number of inputs whose
 * quotient differs bitwise from div.rn.f32 (must be 0) */
int b2r_selftest_div(int64_t n_per_mode, uint64_t seed, int64_t* mismatches_host, cudaStream_t stream);
/* Same for the f64 division by a per-call constant (fx, fy, unit length) in the allocation samples of a8: number of
 * random numerators whose quotient by `b` differs bitwise from div.rn.f64 (must be 0) */
int b2r_selftest_ddiv(double b, int64_t n, uint64_t seed, int64_t* mismatches_host, cudaStream_t stream);
/* parity / checkpoint hooks: keys int32[n][3]; planes tsdf[n][4096], weight[n][4096], rgb[n][3][4096],
 * voxel index x*256 + y*16 + z (upstream IndexOf) */
int b2r_tsdf_export_keys(b2r_tsdf* h, int32_t* keys3, int64_t n, cudaStream_t stream);
int b2r_tsdf_export_units(b2r_tsdf* h, int64_t first, int64_t count, float* tsdf, float* weight, float* rgb,
                          cudaStream_t stream);
int b2r_tsdf_import_units(b2r_tsdf* h, const int32_t* keys3, int64_t n, const float* tsdf, const float* weight,
                          const float* rgb, cudaStream_t stream);

/* ---- a10: volume.extract_point_cloud()  (GoodColorReconstruction.py:305,542).  f64 outputs.
 *      Two-call protocol: xyz==NULL or capacity too small -> *count set, returns B2R_ERR_CAPACITY. */
int b2r_tsdf_extract_points(b2r_tsdf* h, double* xyz, double* rgb, double* nrm, int32_t* ids4, int64_t capacity,
                            int64_t* count_host, cudaStream_t stream);
/* ---- a11: volume.extract_triangle_mesh()  (GoodColorReconstruction.py:285,536) */
int b2r_tsdf_extract_mesh(b2r_tsdf* h, double* vtx, double* vcol, int32_t* tri, int32_t* ids4, int64_t capV,
                          int64_t capF, int64_t* nV_host, int64_t* nF_host, cudaStream_t stream);
/* Multi-GPU extraction; integration shards by unit-key hash, extraction re-partitions by kx SLABS (cubes and
 * normals straddle unit borders, and a slab has only thin ghost layers).  The handle holds every unit with kx in
 * [kx_lo, kx_hi] plus the ghost layers kx_lo - 1, kx_hi + 1, kx_hi + 2; a10/a11 emit what the slab's units own.  Vertex ids are
 * global: vertex_base (the vertex count of the slabs below, known after every rank's sizing call) + the key-order numbering,
 * which runs on into the ghost layer above -- the slab-order concatenation of the ranks' outputs is the single-GPU mesh. */
int b2r_tsdf_extract_mesh_slab(b2r_tsdf* h, int kx_lo, int kx_hi, int64_t vertex_base, double* vtx, double* vcol, int32_t* tri,
                               int32_t* ids4, int64_t capV, int64_t capF, int64_t* nV_host, int64_t* nF_host, cudaStream_t stream);
int b2r_tsdf_extract_points_slab(b2r_tsdf* h, int kx_lo, int kx_hi, double* xyz, double* rgb, double* nrm, int32_t* ids4,
                                 int64_t capacity, int64_t* count_host, cudaStream_t stream);
/* units ids[0..n) (DEVICE pool indices) in export layout (keys3 may be NULL) */
int b2r_tsdf_export_selected(b2r_tsdf* h, const int32_t* ids, int64_t n, int32_t* keys3, float* tsdf, float* weight, float* rgb,
                             cudaStream_t stream);
/* the same units as pool-layout records [n][planes * 4096] floats: what the re-partition sends; the receiver lets the
 * all-to-all deposit them straight into its pool (b2r_tsdf_pool) and registers them with b2r_tsdf_adopt_units */
int b2r_tsdf_export_records(b2r_tsdf* h, const int32_t* ids, int64_t n, int32_t* keys3, float* units, cudaStream_t stream);
float* b2r_tsdf_pool(b2r_tsdf* h, int64_t* floats_per_unit);
int b2r_tsdf_adopt_units(b2r_tsdf* h, const int32_t* keys3, int64_t first, int64_t n, cudaStream_t stream);
/* ---- multi-GPU frame exchange over peer memory (one process per GPU; SURVEY 8(e) collective (1), north_star "each frame's
 *      depth/color/pose is broadcast"; reference precedent for many cameras into one volume:
 *      BetterVisualizerD2camerasWORKSkinda.py:546-551).  A rank keeps its uploaded frames in a b2r_peer_alloc buffer that the
 *      other ranks map with b2r_peer_open; b2r_tsdf_integrate_frames then takes pointers INTO THE PEERS' BUFFERS and K1's TMA
 *      loads pull the pixels over NVLink.  Ordering between ranks: 32-bit counters written / awaited in stream order
 *      (cuStreamWriteValue32 / cuStreamWaitValue32 -- no kernel, no SM).  handle64: 64 bytes, exchanged by any transport. */
int b2r_peer_alloc(int64_t bytes, void** dptr, unsigned char* handle64);
int b2r_peer_free(void* dptr);
int b2r_peer_open(const unsigned char* handle64, void** dptr);
int b2r_peer_close(void* dptr);
int b2r_peer_copy(void* dst, const void* src, int64_t bytes, cudaStream_t stream);   /* copy engine; either side may be a mapped peer buffer */
int b2r_stream_write32(cudaStream_t stream, void* addr, uint32_t value);
int b2r_stream_wait32_geq(cudaStream_t stream, void* addr, uint32_t value);
int b2r_stream_write32_multi(cudaStream_t stream, void* const* addrs, int n, uint32_t value);
int b2r_stream_wait32_geq_multi(cudaStream_t stream, void* const* addrs, int n, uint32_t value);
/* wait *ready >= value, copy-engine copy, *done = value: one peer's share of a super-chunk */
int b2r_peer_pull(cudaStream_t stream, void* ready, void* dst, const void* src, int64_t bytes, void* done, uint32_t value);
/* mesh.compute_vertex_normals()  (GoodColorReconstruction.py:286) */
int b2r_mesh_vertex_normals(const double* vtx, int64_t nV, const int32_t* tri, int64_t nF, double* out,
                            cudaStream_t stream);
/* Allocation-free variant of b2r_mesh_vertex_normals.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same inputs, outputs and result (per-vertex sums in ascending triangle order); the scratch -- per-vertex
 * incidence lists -- lives in a file-scope device array of 8.4 MB that every GPU loading the library holds until exit.
 * Limits: nV <= B2R_NOALLOC_MAX_VERTICES, nF <= B2R_NOALLOC_MAX_TRIANGLES, else B2R_ERR_CAPACITY (-3) and a message.  Calls no
 * cudaMalloc / cudaFree / allocating library routine.  All calls on a device share the array: calls that overlap on the GPU
 * (different streams) race; calls on one stream are safe (the stale counts are cleared at the start of every call). */
#define B2R_NOALLOC_MAX_VERTICES 262144
#define B2R_NOALLOC_MAX_TRIANGLES 524288
int b2r_mesh_vertex_normals_noalloc(const double* vtx, int64_t nV, const int32_t* tri, int64_t nF, double* out,
                                    cudaStream_t stream);

/* ---- a12: o3d.pipelines.odometry.compute_rgbd_odometry  (RGBD_Odometry.py:26-35,
 *      capture_frames.py:148-151, bundlefusion_reconstruction.py:224-227).
 *      Images: f32 intensity + f32 depth (metres), W x H, DEVICE.  jac_kind 0 = colour term, 1 = hybrid.
 *      iters_host[n_levels]: iterations per pyramid level, coarsest first ([20,10,5]).
 *      out_T_host 16 f64 (source->target), out_info_host 36 f64, *success_host 0/1.  Synchronises. */
int b2r_odometry(const float* src_gray, const float* src_depth, const float* tgt_gray, const float* tgt_depth, int W,
                 int H, double fx, double fy, double cx, double cy, const double* init_host, int jac_kind,
                 const int* iters_host, int n_levels, double depth_diff_max, double depth_min, double depth_max,
                 double* out_T_host, double* out_info_host, int* success_host, cudaStream_t stream);
/* n independent pairs through ONE launch sequence (config 4: the consecutive frames of a sequence, odo_init = I each,
 * RGBD_Odometry.py:26-35 / capture_frames.py:148-151 in a loop): images = HOST arrays of n DEVICE pointers; init_host = n x 16
 * or NULL (identity); outputs n x 16, n x 36, n.  Pair p's result is bit-identical to b2r_odometry on that pair alone. */
int b2r_odometry_batch(int n, const float* const* src_gray, const float* const* src_depth, const float* const* tgt_gray,
                       const float* const* tgt_depth, int W, int H, double fx, double fy, double cx, double cy,
                       const double* init_host, int jac_kind, const int* iters_host, int n_levels, double depth_diff_max,
                       double depth_min, double depth_max, double* out_T_host, double* out_info_host, int* success_host,
                       cudaStream_t stream);

/* image building blocks of a12, exposed for parity tests (Image::Filter Gaussian3/Sobel3Dx/Sobel3Dy =
 * kind 0/1/2, Image::Downsample) */
int b2r_image_filter(const float* in, int W, int H, int kind, float* out, cudaStream_t stream);
int b2r_image_downsample(const float* in, int W, int H, float* out, cudaStream_t stream);

/* ---- a13: registration_icp(..., TransformationEstimationPointToPlane(), ICPConvergenceCriteria(...))
 *      (GoodColorReconstruction.py:194-198, BetterVisualizer2CamWorksCOLOR.py:255-259).
 *      src_xyz N*3, tgt_xyz/tgt_nrm M*3 f64 DEVICE.  corr (optional) DEVICE int32[N]: target index or -1.
 *      Returns iterations executed (>= 0) or an error code; synchronises. */
int b2r_icp_point_to_plane(const double* src_xyz, int64_t N, const double* tgt_xyz, const double* tgt_nrm, int64_t M,
                           double max_dist, const double* init_host, int max_iter, double rel_fitness,
                           double rel_rmse, double* out_T_host, double* fitness_host, double* rmse_host,
                           int32_t* corr, cudaStream_t stream);
/* result.correspondence_set (GoodColorReconstruction.py:200-204 reads the result's fields): the matched (source, target) index pairs,
 * ascending by source, from the per-source target indices the ICP entry points write; pairs_out: DEVICE int32[n][2] */
int b2r_correspondence_set(const int32_t* corr, int64_t n, int32_t* pairs_out, int64_t* count_host, cudaStream_t stream);

/* registration_icp with TransformationEstimationPointToPoint(False) (Works2CamerasCOLORVisualizer.py:252-257): the same
 * on-device loop, update = closed-form rigid fit (Horn) of the correspondences.  No normals needed. */
int b2r_icp_point_to_point(const double* src_xyz, int64_t N, const double* tgt_xyz, int64_t M, double max_dist,
                           const double* init_host, int max_iter, double rel_fitness, double rel_rmse, double* out_T_host,
                           double* fitness_host, double* rmse_host, int32_t* corr, cudaStream_t stream);

/* ---- f4 (part): o3d.pipelines.registration.registration_colored_icp(src, tgt, max_dist, init,
 *      TransformationEstimationForColoredICP(), ICPConvergenceCriteria(...))  (GoodColorReconstruction.py:246-254).
 *      DEVICE f64 arrays, colours in [0, 1]; the target's colour gradients (SearchHybrid(2 * max_dist, 30) neighbours) are
 *      computed inside; grad_out optional device [M][3].  Missing normals / colours: B2R_ERR_FORMAT with upstream's text.
 *      Returns the iteration count. */
int b2r_icp_colored(const double* src_xyz, const double* src_rgb, int64_t N, const double* tgt_xyz, const double* tgt_nrm,
                    const double* tgt_rgb, int64_t M, double max_dist, double lambda_geometric, const double* init_host,
                    int max_iter, double rel_fitness, double rel_rmse, double* out_T_host, double* fitness_host,
                    double* rmse_host, int32_t* corr, double* grad_out, cudaStream_t stream);

/* ---- f4: o3d.pipelines.registration.compute_fpfh_feature(pcd, KDTreeSearchParamHybrid(radius, max_nn))
 *      (BetterVisualizer2CamWorksCOLOR.py:211-214).  DEVICE f64; out [n][33] point-major (Feature.data transposed).
 *      radius <= 0: KDTreeSearchParamKNN(max_nn).  No normals: B2R_ERR_FORMAT with upstream's text. */
int b2r_compute_fpfh(const double* xyz, const double* normals, int64_t n, int max_nn, double radius, double* out,
                     cudaStream_t stream);

/* corres_ij of registration_ransac_based_on_feature_matching: nearest target feature (squared L2, ties -> lower index) of
 * every query feature.  DEVICE f64 point-major [n][dim], dim <= 40. */
int b2r_match_features(const double* query_feat, int64_t nq, const double* target_feat, int64_t nt, int dim, int32_t* out_idx,
                       cudaStream_t stream);
/* registration_ransac_based_on_correspondence(source, target, corres, max_dist, TransformationEstimationPointToPoint(False),
 * ransac_n, [CorrespondenceCheckerBasedOnEdgeLength(edge_thr), CorrespondenceCheckerBasedOnDistance(dist_thr)],
 * RANSACConvergenceCriteria(max_iteration, confidence))  (BetterVisualizer2CamWorksCOLOR.py:230-243).
 * DEVICE: src [N][3], tgt [M][3], corres int32 [C][2], corr_out (optional) int32 [N] = nearest target of every source point
 * under the winner, -1 beyond max_dist.  edge_thr / dist_thr <= 0: checker absent.  Deterministic for a given seed
 * (csrc/features.cu).  HOST outputs; report4 (optional) = {iterations consumed, hypotheses validated, winning iteration,
 * its inlier-correspondence count}.  Nothing acceptable: identity, fitness 0 (upstream's empty RegistrationResult). */
int b2r_ransac_correspondence(const double* src_xyz, int64_t N, const double* tgt_xyz, int64_t M, const int32_t* corres, int64_t C,
                              double max_dist, int ransac_n, double edge_thr, double dist_thr, int64_t max_iteration,
                              double confidence, uint64_t seed, double* out_T_host, double* fitness_host, double* rmse_host,
                              int64_t* report4_host, int32_t* corr_out, cudaStream_t stream);
typedef struct b2r_knn b2r_knn;   /* the exact-neighbour index of row f1, declared below */
/* The validation step on its own: count[h] = source points with a target point (indexed by `h`) closer than max_dist after
 * T[h] (DEVICE [H][16]), sumd2[h] = sum of those squared distances.  DEVICE outputs. */
int b2r_knn_evaluate_transforms(const b2r_knn* h, const double* src_xyz, int64_t N, const double* T, int64_t H, double max_dist,
                                int64_t* count, double* sumd2, cudaStream_t stream);

/* ---- f1: point-cloud conditioning on either side of ICP.  All arrays DEVICE f64 [n][3] unless noted.
 *      Reference: pcd.remove_statistical_outlier(20, 2.0) / voxel_down_sample(v) / estimate_normals(Hybrid(r, max_nn)) /
 *      orient_normals_towards_camera_location -- GoodColorReconstruction.py:141-153,
 *      BetterVisualizer2CamWorksCOLOR.py:531,537,643.  Exact neighbour search through a uniform grid (pointcloud.cu). */
typedef struct b2r_knn b2r_knn;
/* index of n points.  cell > 0: grid cell edge (the search radius for hybrid queries); cell <= 0: chosen for k_hint-NN
 * queries.  Synchronises.  NULL on error. */
b2r_knn* b2r_knn_create(const double* xyz, int64_t n, double cell, int k_hint, cudaStream_t stream);
void b2r_knn_destroy(b2r_knn* h);
int b2r_knn_info(const b2r_knn* h, double* cell, int64_t* n_cells);
/* KDTreeFlann::SearchKNN (radius <= 0) / SearchHybrid (radius > 0, at most k with d2 < radius^2).  query NULL = every
 * indexed point (itself first).  idx [nq][k] ascending (distance, index), -1 padded; d2 optional [nq][k]; counts [nq]. */
int b2r_knn_search(const b2r_knn* h, const double* query, int64_t nq, int k, double radius, int32_t* idx, double* d2,
                   int32_t* counts, cudaStream_t stream);
/* PointCloud::EstimateNormals(param, fast_normal_computation=true); radius <= 0: KNN(k), else Hybrid(radius, k).
 * normals in/out; has_normals=1 orients along the existing normals; cov_out optional [n][6] (xx,xy,xz,yy,yz,zz). */
int b2r_pcd_estimate_normals(const b2r_knn* h, const double* xyz, int k, double radius, double* normals, int has_normals,
                             double* cov_out, cudaStream_t stream);
/* PointCloud::OrientNormalsTowardsCameraLocation(cam) */
int b2r_pcd_orient_normals(const double* xyz, int64_t n, const double* cam3_host, double* normals, cudaStream_t stream);
/* PointCloud::RemoveStatisticalOutliers: keep u8[n], inlier_idx int32[n] (first *n_inliers_host valid, ascending),
 * avg_out optional f64[n].  Synchronises. */
int b2r_pcd_statistical_outlier(const b2r_knn* h, int nb_neighbors, double std_ratio, uint8_t* keep, int32_t* inlier_idx,
                                int64_t* n_inliers_host, double* avg_out, cudaStream_t stream);
/* select_by_index: dst[r][:] = src[idx[r]][:], rows of `width` f64 (or int32 when is_i32) */
int b2r_pcd_gather_rows(const void* src, const int32_t* idx, int64_t m, int width, int is_i32, void* dst, cudaStream_t stream);
/* PointCloud::VoxelDownSample(voxel): outputs in ascending voxel-index order, `capacity` rows (n always suffices);
 * col / nrm / out_key optional.  *m_host = voxels.  Synchronises. */
int b2r_pcd_voxel_down_sample(const double* xyz, const double* col, const double* nrm, int64_t n, double voxel, double* out_xyz,
                              double* out_col, double* out_nrm, int32_t* out_key, int64_t capacity, int64_t* m_host,
                              cudaStream_t stream);
/* get_min_bound / get_max_bound: out6_host = {min xyz, max xyz}.  Synchronises. */
int b2r_pcd_bounds(const double* xyz, int64_t n, double* out6_host, cudaStream_t stream);

/* ---- f3: o3d.pipelines.registration.global_optimization(pose_graph, GlobalOptimizationLevenbergMarquardt(),
 *      GlobalOptimizationConvergenceCriteria(), GlobalOptimizationOption(...))
 *      (CodeThatKindaWorks/capture_frames.py:74-89; bundlefusion_reconstruction.py:260-273,292-304).
 *      HOST arrays, row-major f64 (a small sparse LM); no CUDA device needed.
 *      poses[n_nodes][16] in/out; transformation[n_edges][16]; information[n_edges][36]; confidence in/out;
 *      keep[n_edges] out (upstream replaces the graph by the pruned one: drop the edges with keep == 0);
 *      option4 = {max_correspondence_distance, edge_prune_threshold, preference_loop_closure, reference_node};
 *      criteria8 = {max_iteration, min_relative_increment, min_relative_residual_increment, min_right_term,
 *                   min_residual, max_iteration_lm, upper_scale_factor, lower_scale_factor};
 *      report8 (may be NULL) = {valid, iterations / LM steps of pass 1, of pass 2, residual before, after, edges kept}.
 *      An invalid graph (unconnected, bad node id, certain edge with confidence != 1) is left untouched as
 *      upstream does (warning text in b2r_last_error, report8[0] = 0, returns B2R_OK). */
int b2r_posegraph_optimize(int n_nodes, double* poses, int n_edges, const int32_t* src, const int32_t* dst,
                           const double* transformation, const double* information, const uint8_t* uncertain,
                           double* confidence, uint8_t* keep, const double* option4, const double* criteria8,
                           double* report8);

#ifdef __cplusplus
}
#endif
#endif /* B2RECON_H */
