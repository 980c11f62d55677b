/*
 * lodgs.h — C ABI of the B200-native (sm_100a) hierarchical-LOD Gaussian hot path.
 *
 * Drop-in boundary for FelixWindisch/hierarchical-LOD-gaussians. Every entry point takes plain
 * device/host pointers and sizes, an explicit CUDA stream (passed as void*), never allocates device
 * memory (the caller owns all tensors and workspaces) and returns 0 on success or a negative
 * LODGS_E_* code; lodgs_last_error() gives the text. No torch types appear here.
 *
 * Each function cites the reference interface it replaces (paths under /root/reference):
 *   HR = submodules/hierarchy-rasterizer, GH = submodules/gaussianhierarchy, AR = submodules/alt-rasterizer
 */
#ifndef LODGS_H_INCLUDED
#define LODGS_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

#define LODGS_OK 0
#define LODGS_E_ARG (-1)       /* bad argument (null pointer, negative size, ...) */
#define LODGS_E_WORKSPACE (-2) /* caller workspace too small */
#define LODGS_E_CUDA (-3)      /* CUDA runtime error (text in lodgs_last_error) */
#define LODGS_E_IO (-4)        /* file not found / short read */

/* Text of the last error on this thread ("" if none). */
const char* lodgs_last_error(void);
/* Library version, and the SM architecture the kernels were compiled for (100). */
int lodgs_version(void);
int lodgs_compiled_sm(void);
/* Number of kernels launched by this library since load (bench.py's gpu_launches). */
uint64_t lodgs_launch_count(void);
/* Per-kernel device timing for bench.py's roofline: lodgs_profile(1) starts recording a CUDA event after every
 * kernel launch on the launching stream (and at every API entry), lodgs_profile(0) stops and clears.
 * lodgs_profile_dump synchronises the device and writes "kernel_name total_ms launches\n" lines into buf. */
int lodgs_profile(int enable);
int lodgs_profile_dump(char* buf, size_t nbytes);

/* ------------------------------------------------------------------------------------------------
 * LOD cut.  Replaces Switching::expandToSize (GH/runtime_switching.cu:997-1030:
 * markNodesForSize :495-529 + cub InclusiveSum + putRenderIndices :68-93) as bound by
 * gaussian_hierarchy._C.expand_to_size (GH/torch/torch_interface.cpp:148-167).
 *   nodes  [N,7] i32  {depth,parent,start,count_leafs,count_merged,start_children,count_children} (GH/types.h:83-92)
 *   boxes  [N,2,4] f32 {min.xyz,w ; max.xyz,w}                                              (GH/types.h:51-55)
 *   viewpoint_dev: 3 floats on the DEVICE (the reference dereferences it in the kernel, :1019)
 * Outputs are written in ascending node-id order into caller buffers (capacity `capacity` entries;
 * entries beyond capacity are dropped and LODGS_E_WORKSPACE is returned when the count is read back).
 * count_host: if non-NULL the count is copied back and the stream synchronised (the reference's
 * blocking cudaMemcpy, :1027-1028); if NULL the call is fully asynchronous and the count stays in
 * the first int32 of `ws`.
 * mode: 0 = single-pass flat rule (reference-equivalent work, 60 B/node), 1 = level-synchronous
 * warp-cooperative traversal (requires lodgs_cut_prepare to have validated monotonicity; falls back to 0).
 */
size_t lodgs_cut_workspace_bytes(int N);
int lodgs_expand_to_size(const int32_t* nodes, const float* boxes, int N, float target_size,
                         const float* viewpoint_dev, int32_t* render_indices, int32_t* parent_indices,
                         int32_t* nodes_for_render_indices, int capacity, void* ws, size_t ws_bytes,
                         int mode, int32_t* count_host, void* stream);

/* One-time hierarchy check for the traversal cut: verifies on the device that along every
 * parent->child edge box(parent) contains box(child) and w(parent) >= w(child), which makes the
 * top-down traversal select exactly the flat rule's set. Writes 1/0 to *monotone_host (synchronises). */
int lodgs_cut_prepare(const int32_t* nodes, const float* boxes, int N, void* ws, size_t ws_bytes,
                      int32_t* monotone_host, void* stream);

/* Interpolation weights. Replaces Switching::getTsIndexed (GH/runtime_switching.cu:687-714,
 * computeTsIndexed :588-634) as bound by gaussian_hierarchy._C.get_interpolation_weights
 * (torch_interface.cpp:200-220). viewpoint is passed BY VALUE (host floats) as in the reference. */
int lodgs_get_interpolation_weights(const int32_t* node_indices, int K, float target_size,
                                    const int32_t* nodes, const float* boxes, float vx, float vy, float vz,
                                    float* ts, int32_t* num_siblings, void* stream);

/* Cut and weights on the DYNAMIC hierarchy format (the one this fork's creator writes and create_from_hier loads):
 * replaces Switching::expandToSizeDynamic (GH/runtime_switching.cu:748-782, kernels :533-582, :95-108; bound as
 * expand_to_size_dynamic, GH/torch/torch_interface.cpp:171-194) and Switching::getTsIndexedDynamic (:637-684; bound as
 * get_interpolation_weights_dynamic, torch_interface.cpp:222-244). nodes = HierarchyNode[N] (6 x i32: depth, parent,
 * child_count, first_child, next_sibling, max_side_length; GH/types.h:60-79), positions f32[N,3], scales f32[N,3].
 * A node is on the cut when it lies inside the view cone (cos between viewpoint-position and zdir < -0.5), has depth >= 0 and is
 * either a leaf with max(scale)/distance >= target_size or the first node below target_size on its path. Entries are node ids
 * in ascending order; parent_indices[k] is written only for entries with a parent (the reference leaves the caller's value
 * for roots). Workspace and count conventions as lodgs_expand_to_size. */
int lodgs_expand_to_size_dynamic(const int32_t* nodes, const float* positions, const float* scales, int N,
                                 float target_size, const float* viewpoint_dev, float zdir_x, float zdir_y, float zdir_z,
                                 int32_t* render_indices, int32_t* parent_indices, int32_t* nodes_for_render_indices,
                                 int capacity, void* ws, size_t ws_bytes, int32_t* count_host, void* stream);
int lodgs_get_interpolation_weights_dynamic(const int32_t* node_indices, int K, float target_size, const int32_t* nodes,
                                            const float* positions, const float* scales, float vx, float vy, float vz,
                                            float* ts, int32_t* num_siblings, void* stream);

/* Cut through the selected SPTs. Replaces Switching::getSPTCut (GH/runtime_switching.cu:881-994, kernels :784-852) as bound by
 * gaussian_hierarchy._C.get_spt_cut_cuda (GH/torch/torch_interface.cpp:287-316). Table t is the run
 * [SPT_starts[t], SPT_starts[t+1]) of (gaussian_indices, SPT_min, SPT_max) entries, sorted by SPT_max descending; slot i selects
 * table SPT_indices[i] at distance SPT_distances[i]. The result is the reference's entry for entry: ids of the entries with
 * SPT_min < distance inside the bisected prefix of each run (with the reference's one-entry shift of that window, see
 * csrc/cut_spt.cu), id 0 dropped, in slot order; counts_out[S] = exclusive prefix sum of the per-slot kept entries.
 * Two calls because the caller allocates the result: lodgs_spt_cut_count fills counts_out and the workspace and returns the
 * number of ids through *count_host (one 12-byte read-back; pass NULL to stay asynchronous), lodgs_spt_cut_write then writes at
 * most `capacity` ids. n_entries = length of the three entry arrays; an entry index >= n_entries is treated as not kept. */
size_t lodgs_spt_cut_workspace_bytes(int S);
int lodgs_spt_cut_count(int S, const int32_t* gaussian_indices, int n_entries, const int32_t* SPT_starts, const float* SPT_max,
                        const float* SPT_min, const int32_t* SPT_indices, const float* SPT_distances, int32_t* counts_out,
                        void* ws, size_t ws_bytes, int32_t* count_host, void* stream);
int lodgs_spt_cut_write(int S, const int32_t* gaussian_indices, int n_entries, const int32_t* SPT_starts, const float* SPT_min,
                        const int32_t* SPT_indices, const float* SPT_distances, const void* ws, int32_t* cut_out, int capacity,
                        void* stream);

/* Cut of the "upper tree" that precedes the SPT cut — one level-synchronous kernel. Replaces the per-frame torch code
 * GaussianModel.cut_hierarchy_on_condition (scene/gaussian_model.py:354-394) with its cull GaussianModel.frustum_cull_spheres (:80-103)
 * and LOD condition `min_distance_squared[i] > |campos - xyz[i]|^2 * distance_multiplier` (train_post.py:336-339), plus the split of
 * the result that feeds get_spt_cut_cuda (train_post.py:349-356). nodes [N,6] i32 = upper_tree_nodes {depth, parent, child_count,
 * first_child (leaves: SPT index, or <= 0 for none), next_sibling, gaussian index}; xyz [N,3]; bounds [N] bounding-sphere radii or NULL
 * (no culling); planes [n_planes,4] normalised (nx,ny,nz,d) as extract_frustum_planes builds them (:55-78), 16-byte aligned;
 * min_dist_sq [N]; campos [3]; all on the device. Outputs in the reference's order (per level: leaves of the culled stack, then the
 * nodes failing the condition; next stack = first children then their next siblings): cut [capacity] node ids; for the cut's leaves,
 * in cut order: spt_indices / spt_nodes / spt_distances (leaves with first_child >= 0: SPT id, node id, |xyz - campos| *
 * distance_multiplier) and direct (gaussian index of leaves with first_child <= 0) — each [capacity] or NULL. counts_host (3 ints:
 * cut, SPT slots, direct leaves) is filled through a mapped pinned word when non-NULL (NULL: nothing is read back, the counts
 * stay in the first 12 bytes of ws). */
size_t lodgs_upper_cut_workspace_bytes(int N);
int lodgs_upper_tree_cut(const int32_t* nodes, int N, int root, const float* xyz, const float* bounds, const float* planes, int n_planes,
                         const float* min_dist_sq, const float* campos, float distance_multiplier, int32_t* cut, int capacity,
                         int32_t* spt_indices, int32_t* spt_nodes, float* spt_distances, int32_t* direct, void* ws, size_t ws_bytes,
                         int32_t* counts_host, void* stream);

/* CPU <-> GPU paging of parameter and Adam-moment rows between training iterations. Replaces train_post.py:440-479 (write-back of the
 * rows leaving the render set into host storage, `tensor[keep_mask]`, `cat(kept, storage[load_indices].cuda())`, for 6 parameter and 12
 * moment tensors). lodgs_page_plan scans the keep mask ONCE: dst[i] = slot of resident row i among the kept rows (>= 0), or -1 - slot
 * among the evicted rows, both in the original order (what boolean-mask indexing yields); counts_host[0..1] = kept, evicted (through a
 * mapped pinned word; NULL = not read back, they stay in the first 8 bytes of ws). lodgs_page_apply, per tensor of `width` floats per
 * row: out[dst] = kept rows, out[n_keep + j] = loaded[j] (the rows just copied up from host storage, device staging [n_load,width]),
 * evict[-1 - dst] = evicted rows (device staging [evicted,width], NULL to drop them). The transfers between device staging and pinned
 * host staging are one cudaMemcpyAsync each (the caller's); lodgs_host_scatter_rows / lodgs_host_gather_rows move whole rows between
 * pinned staging and the big host storage tensors with memcpy over the host cores (storage[indices[i]] = staging[i] and the reverse). */
size_t lodgs_page_workspace_bytes(int64_t n);
int lodgs_page_plan(const uint8_t* keep_mask, int64_t n, int32_t* dst, void* ws, size_t ws_bytes, int32_t* counts_host, void* stream);
/* lodgs_page_plan for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). No
 * counterpart in the reference. Same dst[] as lodgs_page_plan; the scan's scratch is a static
 * __device__ array inside the library (64.4 KiB, held by every GPU that loads the library until exit), so n is limited to
 * lodgs_page_plan_static_max_rows() = 16 777 216 rows — more returns LODGS_E_WORKSPACE with a message naming the limit. The totals
 * {kept, evicted} go to counts_dev, a caller-owned DEVICE buffer of 2 ints (no host read-back). The call allocates nothing: no cudaMalloc /
 * cudaFree / pinned memory / events, only kernel launches on `stream`. All calls on a device share the one scratch array: every word a call
 * reads it has written first (nothing needs clearing between calls), but two calls that overlap on the GPU (different streams, no ordering)
 * race on it — serialise them. */
int64_t lodgs_page_plan_static_max_rows(void);
int lodgs_page_plan_static(const uint8_t* keep_mask, int64_t n, int32_t* dst, int32_t* counts_dev, void* stream);
int lodgs_page_apply(const float* src, int64_t n, int width, const int32_t* dst, float* out, float* evict, const float* loaded,
                     int64_t n_load, int64_t n_keep, void* stream);
int lodgs_host_scatter_rows(void* storage, int64_t storage_rows, int64_t row_bytes, const int64_t* indices, int64_t n, const void* staging);
int lodgs_host_gather_rows(const void* storage, int64_t storage_rows, int64_t row_bytes, const int64_t* indices, int64_t n, void* staging);

/* Fused L1 + SSIM image loss, forward and backward. Replaces the loss of the training loops, train_post.py:558-559 / train_single.py:
 * 106-110: `(1 - lambda) * l1_loss(image, gt) + lambda * (1 - fused_ssim(image[None], gt[None]))` with l1_loss / ssim of utils/loss_utils.py:
 * 17-18,33-63 (11x11 Gaussian window, sigma 1.5, zero padding, C1 = 0.01^2, C2 = 0.03^2; `fused_ssim` is an external package computing
 * the same SSIM). img, gt [C,H,W] on the device. out3 (device, 3 floats) = { a_l1 * mean|img - gt| + a_ssim * mean(ssim) + bias, mean|img - gt|,
 * mean(ssim) }; dL_dimg [C,H,W] (optional) = gradient of out3[0] with respect to img. The training loss is a_l1 = 1 - lambda,
 * a_ssim = -lambda, bias = lambda; fused_ssim alone is a_l1 = 0, a_ssim = 1, bias = 0. ws: lodgs_l1_ssim_workspace_bytes (3 derivative
 * maps + block partial sums). Sums are reduced in a fixed order in double: the result is deterministic. */
size_t lodgs_l1_ssim_workspace_bytes(int C, int H, int W);
int lodgs_l1_ssim_loss(const float* img, const float* gt, int C, int H, int W, float a_l1, float a_ssim, float bias, float* out3,
                       float* dL_dimg, void* ws, size_t ws_bytes, void* stream);

/* Morton codes. Replaces MORTON::getMortonCode (GH/morton.cu:9-60) as bound by gaussian_hierarchy._C.get_morton_indices
 * (torch_interface.cpp GetMortonCode; scene/gaussian_model.py:574). xyz f32[P,3]; min3/max3 = 3 floats each ON THE DEVICE;
 * codes i64[P]: bits 0..20 of trunc((c - min) / (max - min) * 2^21) per axis, interleaved x, y, z from bit 0. */
int lodgs_get_morton_codes(const float* xyz, const float* min3_dev, const float* max3_dev, int64_t* codes, int P, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Hierarchy file I/O (host only). Replaces HierarchyLoader::load (GH/hierarchy_loader.cpp:26-128)
 * behind gaussian_hierarchy._C.load_hierarchy (torch_interface.cpp:9-40), and HierarchyWriter::write
 * (GH/hierarchy_writer.cpp:27-119). Two-call protocol: lodgs_hier_probe gives P and N, then the caller
 * allocates pos[P,3] shs[P,48] alpha[P] scales[P,3] rot[P,4] nodes[N,7] boxes[N,8] and calls lodgs_hier_load.
 * Half-compressed files (negative P on disk) are expanded to fp32.
 */
int lodgs_hier_probe(const char* filename, int32_t* P, int32_t* N);
int lodgs_hier_load(const char* filename, float* pos, float* shs, float* alpha, float* scales, float* rot,
                    int32_t* nodes, float* boxes);
int lodgs_hier_write(const char* filename, int P, int N, const float* pos, const float* shs, const float* alpha,
                     const float* scales, const float* rot, const int32_t* nodes, const float* boxes,
                     int compressed);

/* GPU construction of the hierarchy. Replaces the reference's CPU creator pipeline (GH/mainHierarchyCreator.cpp:149-183):
 * PointbasedKdTreeGenerator::generate (GH/PointbasedKdTreeGenerator.cpp:16-75), ClusterMerger::merge (GH/ClusterMerger.cpp:23-173),
 * RotationAligner::align (GH/rotation_aligner.cpp:24-114) and the flattening of Writer::makeDynamicHierarchy (GH/writer.cpp:98-183) /
 * Writer::makeHierarchy (GH/writer.cpp:21-96, upstream child emission). Inputs are the L leaf Gaussians as the creator holds them after
 * loading and filtering (GH/loader.cpp:105-124): activated scales, normalised rotations (w,x,y,z), activated opacities, sh_floats SH
 * coefficients per Gaussian (48 in the creator). Outputs, all on the device, for the N = 2L-1 nodes: one Gaussian per node in PRE-ORDER
 * (out_pos[N,3], out_rot[N,4], out_scale[N,3] activated, out_opacity[N], out_shs[N,sh_floats]); dyn_nodes[N,6] (HierarchyNode,
 * GH/types.h:60-67: node i owns Gaussian i; max_side_length = index of the input Gaussian for a leaf, -1 otherwise) and/or the static
 * format nodes[N,7] + boxes[N,2,4] (GH/types.h:83-92,51-55: node.start = index of its Gaussian, box w = longest side for inner nodes,
 * 0 for leaves). Any of dyn_nodes / nodes / boxes may be NULL. Equal coordinates at a split fall as a stable sort leaves them
 * (std::nth_element leaves them unspecified); the eigen-decomposition is a float Jacobi iteration (the reference's is Eigen's). */
size_t lodgs_hier_build_workspace_bytes(int L);
/* Host side of the creator around lodgs_hier_build, with the libm calls of the reference so that a .dhier written from the GPU build is
 * byte-identical to the creator's: lodgs_creator_activate = Loader::loadPly<LessRichPoint> (GH/loader.cpp:105-124: sigmoid, exp,
 * normalise, DC only) + the validity filter of GH/mainHierarchyCreator.cpp:87-149 on `count` 26-float PLY records (keep[i] = 1 for the
 * Gaussians the creator keeps); lodgs_logf_host = the `scale.array().log()` of GH/writer.cpp:119,130. Host pointers, no stream. */
int lodgs_creator_activate(const float* records, int64_t count, float* pos, float* scale, float* rot, float* opacity, float* shs48,
                           uint8_t* keep);
int lodgs_logf_host(const float* in, int64_t n, float* out);
int lodgs_hier_build(int L, int sh_floats, const float* pos, const float* scale, const float* rot, const float* opacity,
                     const float* shs, float* out_pos, float* out_rot, float* out_scale, float* out_opacity, float* out_shs,
                     int32_t* dyn_nodes, int32_t* nodes, float* boxes, void* ws, size_t ws_bytes, void* stream);

/* Dynamic hierarchy files (host only). Replaces HierarchyLoader::loadDynamic (GH/hierarchy_loader.cpp:129-189) behind
 * gaussian_hierarchy._C.load_dynamic_hierarchy (torch_interface.cpp:43-75) and HierarchyWriter::writeDynamic
 * (GH/hierarchy_writer.cpp:122-168) behind write_dynamic_hierarchy (torch_interface.cpp:107-134). Layout: int P, int sh_degree,
 * pos[P,3], rot[P,4], log-scales[P,3], opacities[P], SH[P, 3*(deg+1)^2], int N, HierarchyNode[N] (6 x i32). As in the reference
 * the loader returns exactly P nodes whatever N says; lodgs_dhier_probe reports P, the SH degree and the N stored in the file,
 * the caller allocates pos[P,3] shs[P,(deg+1)^2,3] alpha[P] scales[P,3] rot[P,4] nodes[P,6] and calls lodgs_dhier_load. */
int lodgs_dhier_probe(const char* filename, int32_t* P, int32_t* sh_degree, int32_t* N_in_file);
int lodgs_dhier_load(const char* filename, float* pos, float* shs, float* alpha, float* scales, float* rot, int32_t* nodes);
int lodgs_dhier_write(const char* filename, int P, int N, int sh_degree, const float* pos, const float* shs, const float* alpha,
                      const float* scales, const float* rot, const int32_t* nodes);

/* Depth-target cut (host only). Replaces Traversal::expandToTarget (GH/traversal.cpp:15-39) behind
 * gaussian_hierarchy._C.expand_to_target (torch_interface.cpp:136-142): pre-order walk of nodes[N,7] from node 0, Gaussian ids
 * of the leaves of every visited node plus the merged Gaussians of nodes at depth <= target (which are not descended).
 * out = NULL: only *count is produced. */
int lodgs_expand_to_target(const int32_t* nodes, int N, int target, int32_t* out, int64_t capacity, int64_t* count);

/* ------------------------------------------------------------------------------------------------
 * Rasterizer. Replaces CudaRasterizer::Rasterizer::forward / backward / markVisible
 * (HR/cuda_rasterizer/rasterizer.h:20-109, rasterizer_impl.cu:203-517) as bound by
 * diff_gaussian_rasterization._C.rasterize_gaussians[_backward] (HR/rasterize_points.cu:36-245).
 *
 * Scratch is three opaque caller-owned byte buffers (the reference's geomBuffer / binningBuffer /
 * imgBuffer, HR/rasterize_points.cu:85-92); sizes come from the *_bytes functions. The layout is
 * private to this library but fixed between forward and backward; lodgs_*_layout expose array offsets
 * for parity tests.
 *
 * P     = length of the render list (K if `indices` is non-NULL else fullP)
 * D, M  = active SH degree, SH coefficients stored per Gaussian
 * All pointer arguments are DEVICE pointers unless named *_host. Optional inputs are NULL when absent
 * (the reference's empty-tensor convention).
 */
size_t lodgs_geom_bytes(int P);
size_t lodgs_image_bytes(int W, int H);
size_t lodgs_binning_bytes(int R);

/* Stage 1: preprocessCUDA (HR forward.cu:218-445). Writes radii[P], the geometry state (one 48-byte blend record,
 * cov3D, clamp bits, tile box and depth bits per Gaussian) and the number of (Gaussian,tile) instances R; then orders the
 * Gaussians by depth bits (onesweep radix sort over P u32 keys) and scans their tile counts in that order, which
 * replaces the index-order InclusiveSum of rasterizer_impl.cu:322. If num_rendered_host is non-NULL, R is copied back
 * through a pinned staging word and the call waits for it (rasterizer_impl.cu:330-331). colors_precomp, when given, is
 * folded into the blend records here. */
int lodgs_forward_preprocess(int P, int D, int M, int W, int H,
                             const int32_t* indices, const int32_t* parent_indices, const float* ts,
                             const float* means3D, const float* shs, const float* colors_precomp,
                             const float* opacities, const float* scales, float scale_modifier,
                             const float* rotations, const float* cov3D_precomp,
                             const float* viewmatrix, const float* projmatrix, const float* campos,
                             float tan_fovx, float tan_fovy, int32_t* radii,
                             void* geom_buffer, size_t geom_bytes, int32_t* num_rendered_host, void* stream);

/* Same as lodgs_forward_preprocess with the conventions of the reference's second rasterizer
 * (submodules/alt-rasterizer, AR) selected by `flags`:
 *   LODGS_RASTER_SQUARE_RECT   tile rectangle from the square of the largest 3-sigma radius (AR/cuda_rasterizer/forward.cu:249,
 *                              auxiliary.h:45-55) instead of the per-axis rectangle (HR forward.cu:400-404)
 *   LODGS_RASTER_NO_ANTIALIAS  no opacity compensation (AR forward.cu:226-229 with antialiasing=false); pass the same flag to
 *                              lodgs_backward_ex.
 * and, for lodgs_backward_ex only, two quirks of AR's per-Gaussian backward that a drop-in has to reproduce:
 *   LODGS_RASTER_BACKGROUND_TWICE    the background term of dL/dalpha is counted twice: AR backward.cu:576 starts the running
 *                                    colour from minus the RENDERED pixel colour, which already contains T_final*bg
 *                                    (forward.cu:422, rasterizer_impl.cu:495), and :617 subtracts T_final/(1-alpha)*bg.dL again
 *   LODGS_RASTER_CLAMPED_ALPHA_GRAD  pairs whose opacity*G exceeds 0.99 keep their alpha gradient (AR backward.cu:596 has no
 *                                    counterpart of HR backward.cu:620,687 `nullalpha`). */
#define LODGS_RASTER_SQUARE_RECT 1u
#define LODGS_RASTER_NO_ANTIALIAS 2u
#define LODGS_RASTER_BACKGROUND_TWICE 4u
#define LODGS_RASTER_CLAMPED_ALPHA_GRAD 8u
int lodgs_forward_preprocess_ex(int P, int D, int M, int W, int H,
                                const int32_t* indices, const int32_t* parent_indices, const float* ts,
                                const float* means3D, const float* shs, const float* colors_precomp,
                                const float* opacities, const float* scales, float scale_modifier,
                                const float* rotations, const float* cov3D_precomp,
                                const float* viewmatrix, const float* projmatrix, const float* campos,
                                float tan_fovx, float tan_fovy, int32_t* radii,
                                void* geom_buffer, size_t geom_bytes, int32_t* num_rendered_host, unsigned flags, void* stream);

/* lodgs_forward_preprocess_ex for hierarchical-LOD rendering with the sibling counts known at stage 1 (`kids` [P] i32, the
 * num_node_kids of GaussianRasterizationSettings; requires ts): the interpolation weight and 1/kids are stored into the blend
 * records by the preprocess kernel itself, so stage 2 has no separate pass over the records. Call lodgs_forward_render with the
 * same ts and kids = NULL afterwards ("records complete"). kids = NULL here is lodgs_forward_preprocess_ex. */
int lodgs_forward_preprocess_lod(int P, int D, int M, int W, int H,
                                 const int32_t* indices, const int32_t* parent_indices, const float* ts, const int32_t* kids,
                                 const float* means3D, const float* shs, const float* colors_precomp,
                                 const float* opacities, const float* scales, float scale_modifier,
                                 const float* rotations, const float* cov3D_precomp,
                                 const float* viewmatrix, const float* projmatrix, const float* campos,
                                 float tan_fovx, float tan_fovy, int32_t* radii,
                                 void* geom_buffer, size_t geom_bytes, int32_t* num_rendered_host, unsigned flags, void* stream);

/* Stage 2: replaces duplicateWithKeys (rasterizer_impl.cu:70-115) + cub::DeviceRadixSort::SortPairs on the 64-bit
 * tile|depth keys (:358-363) + identifyTileRanges (:120-142) + renderCUDA (forward.cu:450-596). Stage 1 has already
 * ordered the Gaussians by depth bits (onesweep radix sort over P u32 keys); stage 2 emits the (tile, Gaussian) instances
 * in that order, sorts them stably by tile id alone (2 onesweep passes at 1080p instead of 6 over 64-bit keys), which
 * yields exactly the reference's sorted list, then finds the tile ranges and blends; the blend kernels gather the
 * per-Gaussian records through the sorted id list, so no per-instance record or key array is written.
 * R is the value stage 1 reported; ts non-NULL selects the hierarchical-LOD alpha: with kids the records get t and 1/kids here
 * (one extra pass), with kids = NULL they must already hold them (stage 1 = lodgs_forward_preprocess_lod).
 * out_color [3,H,W]; out_invdepth [H,W] or NULL; seen [P] i32 or NULL (must be zero-filled by the caller,
 * as the reference's torch::full does). */
int lodgs_forward_render(int P, int W, int H, int R, const float* background,
                         const float* ts, const int32_t* kids, const float* colors_precomp,
                         const int32_t* radii, void* geom_buffer, void* binning_buffer, size_t binning_bytes,
                         void* image_buffer, size_t image_bytes,
                         float* out_color, float* out_invdepth, int32_t* seen, void* stream);

/* Backward: renderCUDA bwd (HR backward.cu:498-721) with tile-level gradient reduction, then the
 * fused computeCov2DCUDA + preprocessCUDA bwd (backward.cu:147-326, :398-495).
 * All dL_d* outputs are [fullP,...] and must be zero-filled by the caller (rasterize_points.cu:178-187).
 * dL_dmean2D [fullP,3], dL_dconic [fullP,4], dL_dopacity [fullP], dL_dcolor [fullP,3],
 * dL_dinvdepth [fullP] or NULL, dL_dmean3D [fullP,3], dL_dcov3D [fullP,6], dL_dsh [fullP,M,3],
 * dL_dscale [fullP,3], dL_drot [fullP,4]. */
int lodgs_backward(int P, int fullP, int D, int M, int R, int W, int H, const float* background,
                   const int32_t* indices, const int32_t* parent_indices, const float* ts, const int32_t* kids,
                   const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
                   const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
                   const float* viewmatrix, const float* projmatrix, const float* campos,
                   float tan_fovx, float tan_fovy, const int32_t* radii,
                   void* geom_buffer, void* binning_buffer, void* image_buffer,
                   const float* dL_dpix, const float* dL_dinvdepth_pix,
                   float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                   float* dL_dinvdepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                   float* dL_dscale, float* dL_drot, void* stream);

/* lodgs_backward with the LODGS_RASTER_* flags the forward was run with (AR/cuda_rasterizer/rasterizer_impl.cu:537-621). */
int lodgs_backward_ex(int P, int fullP, int D, int M, int R, int W, int H, const float* background,
                      const int32_t* indices, const int32_t* parent_indices, const float* ts, const int32_t* kids,
                      const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
                      const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
                      const float* viewmatrix, const float* projmatrix, const float* campos,
                      float tan_fovx, float tan_fovy, const int32_t* radii,
                      void* geom_buffer, void* binning_buffer, void* image_buffer,
                      const float* dL_dpix, const float* dL_dinvdepth_pix,
                      float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                      float* dL_dinvdepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                      float* dL_dscale, float* dL_drot, unsigned flags, void* stream);

/* lodgs_backward_ex with the colour gradient exposed in its factored form: dL_drgb [fullP,3] receives, per Gaussian, the gradient of
 * the loss with respect to its view-dependent colour AFTER the clamp mask (HR/cuda_rasterizer/backward.cu:45-50), i.e. the factor
 * from which computeColorFromSH's backward forms dL_dsh[k] = basis_k(direction) * dL_drgb (backward.cu:60-131). With dL_drgb given,
 * dL_dsh may be NULL: the 12*M-byte SH gradient row is then neither written nor re-read, and lodgs_sh_adam_from_rgb applies the
 * optimizer step to the SH coefficients directly from dL_drgb (of one view, or of the views of several GPUs: multi-view training
 * exchanges 12 bytes per Gaussian and view instead of an all-reduce over 12*M). Rows of culled Gaussians are set to zero when the
 * call has no index list and no parents (as lodgs_backward_overwrites_geometry reports); otherwise the caller zero-fills dL_drgb.
 * Children with parents (LOD redirect) leave their row untouched, like their dL_dsh row (backward.cu:489-493). */
int lodgs_backward_rgb(int P, int fullP, int D, int M, int R, int W, int H, const float* background,
                       const int32_t* indices, const int32_t* parent_indices, const float* ts, const int32_t* kids,
                       const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
                       const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
                       const float* viewmatrix, const float* projmatrix, const float* campos,
                       float tan_fovx, float tan_fovy, const int32_t* radii,
                       void* geom_buffer, void* binning_buffer, void* image_buffer,
                       const float* dL_dpix, const float* dL_dinvdepth_pix,
                       float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                       float* dL_dinvdepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                       float* dL_dscale, float* dL_drot, float* dL_drgb, unsigned flags, void* stream);

/* The alt-rasterizer's bucket count, exactly (second value of AR's rasterize_gaussians 9-tuple, AR/rasterize_points.cu:116): after
 * lodgs_forward_render of a LODGS_RASTER_SQUARE_RECT forward, out2_host[0] = sum over tiles of ceil(n_t / 32) and out2_host[1] = sum of n_t,
 * n_t = instances of tile t that survive the reference's per-tile opacity test (AR/cuda_rasterizer/rasterizer_impl.cu:56-100,143-187,
 * 223-232). ws: device scratch of at least 4 * tiles + 256 bytes. */
int lodgs_ar_bucket_count(int P, int W, int H, int R, void* geom_buffer, void* binning_buffer, void* ws, size_t ws_bytes,
                          int32_t* out2_host, void* stream);

/* 1 when lodgs_backward will write every row of dL_dsh[0..P) itself (no index list, SH rows a multiple of 16 bytes, 16-byte
 * aligned pointers): the caller may then pass an uninitialised dL_dsh instead of the zero-filled one the reference
 * allocates (rasterize_points.cu:183). */
int lodgs_backward_overwrites_dsh(int M, int has_indices, int has_parents, const float* shs, const float* dL_dsh);

/* 1 when lodgs_backward stores every row [0,P) of dL_dmean3D and dL_dcov3D — and of dL_dscale / dL_drot when scales are
 * given — itself, zeros for culled Gaussians (no index list, no parents): the caller may then pass these four tensors
 * uninitialised instead of zero-filled (rasterize_points.cu:178-187). The tensors the blend accumulates into with
 * atomics (dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dinvdepth) must always be zero-filled. */
int lodgs_backward_overwrites_geometry(int has_indices, int has_parents);

/* MCMC relocation, replaces UTILS::ComputeRelocation (HR/cuda_rasterizer/utils.cu:6-51, bound as _C.compute_relocation):
 * opacity_new[P], scale_new[P,3] of Gaussians split into N[i] copies; binoms is the [n_max,n_max] binomial table. */
int lodgs_compute_relocation(int P, const float* opacity_old, const float* scale_old, const int32_t* N, const float* binoms,
                             int n_max, float* opacity_new, float* scale_new, void* stream);

/* checkFrustum (HR rasterizer_impl.cu:54-66,145-157): present[i] = (view-space z > 0.2). present is u8[P]. */
int lodgs_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                       uint8_t* present, void* stream);

/* Parity-test accessors: byte offsets of the arrays inside the opaque buffers.
 * geom:   [0]=depth bits u32[P] (view-space depth as f32 bits; 0xffffffff for culled Gaussians)
 *         [1]=[2]=[3]=blend records f32[P,12], rows of visible Gaussians only:
 *             mean2D.x mean2D.y conic.xx conic.xy | conic.yy opacity r g | b 1/depth t-or-cut 1/kids
 *         [4]=cov3D f32[P,6] [5]=clamped u8[P] (bit c = channel c clamped) [6]=tiles_touched u32[P]
 *         [7]=[8]=rects i32[P,2]
 * binning:[0]=tiles_unsorted u32[R] [1]=values_unsorted u32[R]  (instances emitted in (depth bits, index) order of their
 *         Gaussians, carrying only the tile id) [2]=tiles_sorted u32[R] [3]=[4]=point_list u32[R]
 * image:  [0]=final_T f32[W*H] [1]=n_contrib u32[W*H] [2]=ranges u32[tiles,2] */
int lodgs_geom_layout(int P, size_t* offsets9);
int lodgs_binning_layout(int R, size_t* offsets5);
int lodgs_image_layout(int W, int H, size_t* offsets3);

/* Work counters of the blend for roofline reports (bench.py): after a forward, counts over all pixels the pairs
 * whose quadratic form is evaluated (N_test), that reach the exponential (N_alpha) and that are blended (N_blend)
 * into counters3_dev[0..2] (device u64). Not used by any product path. */
int lodgs_blend_stats(int P, int W, int H, int R, int lod, void* geom_buffer, void* binning_buffer, void* image_buffer,
                      uint64_t* counters3_dev, void* stream);

/* Parity-test accessor: the reference's sorted 64-bit key list (tile << 32 | depth bits, rasterizer_impl.cu:358-363),
 * rebuilt after a forward from the sorted tile ids, the sorted id list and the depth bits into keys_out u64[R].
 * The product path never materialises these keys. */
int lodgs_sorted_keys(int P, int R, void* geom_buffer, void* binning_buffer, uint64_t* keys_out, void* stream);

/* Stand-alone stable LSD radix sort of (u64 key, u32 value) pairs on bits [0,end_bit) — the onesweep
 * kernel used inside lodgs_forward_render, exported for the CUB cross-check tests.
 * ws must be lodgs_sort_workspace_bytes(n). Result is in keys_out/values_out. */
size_t lodgs_sort_workspace_bytes(int n);
int lodgs_sort_pairs(const uint64_t* keys_in, const uint32_t* values_in, uint64_t* keys_out,
                     uint32_t* values_out, uint64_t* keys_tmp, uint32_t* values_tmp, int n, int end_bit,
                     void* ws, size_t ws_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Sparse Adam. Replaces scene/OurAdam.py Adam.step(relevant) / _single_tensor_adam (:106-176, :251-354):
 * Adam with bias correction restricted to the rows listed in `relevant` (i64 row indices, or NULL = all
 * rows, which is _single_tensor_adam2 :357-452). One launch per parameter tensor of `width` floats per row.
 * step_size = lr / (1 - beta1^step); denom = sqrt(v)/sqrt(1 - beta2^step) + eps (bias terms evaluated
 * on the host in double as Python does; lr/betas/eps are doubles because Python floats are).
 * bias_correction = 0 gives AR's adamUpdateCUDA (AR/cuda_rasterizer/adam.cu:9-38: no bias correction,
 * rows masked by `visible` u8[rows] instead of an index list).
 */
int lodgs_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq,
                    const int64_t* relevant, int64_t n_relevant, const uint8_t* visible, int64_t rows,
                    int width, double lr, double beta1, double beta2, double eps, int64_t step,
                    int bias_correction, void* stream);

/* The same update for all parameter tensors of one optimizer step in ONE call (launches issued back to back from C): tensor t has
 * rows[t] rows of widths[t] floats and its own lr / betas / eps / step; the row list or visibility mask is shared. */
int lodgs_adam_step_multi(int n_tensors, float* const* params, const float* const* grads, float* const* exp_avgs,
                          float* const* exp_avg_sqs, const int64_t* relevant, int64_t n_relevant, const uint8_t* visible,
                          const int64_t* rows, const int32_t* widths, const double* lrs, const double* betas1, const double* betas2,
                          const double* epss, const int64_t* steps, int bias_correction, void* stream);

/* Sparse Adam on the SH coefficients straight from the colour gradient(s) of lodgs_backward_rgb: the fusion of the preprocess
 * backward's SH part (HR/cuda_rasterizer/backward.cu:23-142) with the optimizer step (scene/OurAdam.py:251-354; AR adamUpdateCUDA,
 * AR/cuda_rasterizer/adam.cu:9-38) — the gradient row dL_dsh[row] = sum over views of basis(normalize(mean - campos_v)) (x)
 * dL_drgb[v][row] is rebuilt in registers and never touches HBM.
 *   dL_drgb [n_views] blocks of [P,3], view_stride floats apart (3*P when packed; an all-gather of every rank's block in multi-view
 *   training), campos [n_views,3] (device), means3D [P,3] = the means the forward used.
 *   dc / rest: the two SH parameter tensors of a Gaussian model (scene/gaussian_model.py _features_dc [P,1,3], _features_rest
 *   [P,15,3]) with their first and second moments in the same layout; *_stride = floats between consecutive rows, so one fused
 *   [P,16,3] tensor is dc = shs, rest = shs + 3, both strides 48. rest_floats = 3 * (coefficients in rest), 0 allowed.
 *   Rows: `relevant` index list (OurAdam) or all `rows`, optionally masked by `visible` u8[P] (AR). Coefficients above degree D get
 *   a zero gradient (their moments still decay), as the zero-filled dL_dsh of the reference does.
 * Arithmetic, bias_correction and step as lodgs_adam_step. Views are added in index order on every rank: replicas stay bit-identical. */
int lodgs_sh_adam_from_rgb(int64_t rows, int D, int n_views, const float* dL_drgb, int64_t view_stride, const float* campos,
                           const float* means3D, const int64_t* relevant, int64_t n_relevant, const uint8_t* visible,
                           float* dc, float* dc_m, float* dc_v, int dc_stride, float* rest, float* rest_m, float* rest_v,
                           int rest_stride, int rest_floats, double lr_dc, double lr_rest, double beta1, double beta2, double eps,
                           int64_t step, int bias_correction, void* stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* LODGS_H_INCLUDED */
