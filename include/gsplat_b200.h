/* gsplat_b200.h — C ABI of the B200 (sm_100a) backend for gsplat's rasterization() hot path.
 *
 * This is the drop-in boundary: the Python wrappers in gsplat/b200_impl/_wrapper.py bind these
 * symbols with ctypes exactly where the reference binds its own kernels
 *   - gsplat/triton_impl/_wrapper.py (Triton launchers) and
 *   - gsplat/cuda/csrc/ext.cpp:31-56 (pybind `_C` op list of the `cuda` backend).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in `_host`;
 *   - the caller owns every buffer (outputs, scratch, sort temp): nothing here allocates;
 *   - all floating tensors are contiguous float32, ids int32/int64 as noted;
 *   - `stream` is the CUDA stream to launch on (torch.cuda.current_stream().cuda_stream);
 *   - return value: GSB_OK (0) or a negative error code; gsb_last_error() gives the text.
 *   - no host synchronisation happens inside any entry point.
 */
#ifndef GSPLAT_B200_H
#define GSPLAT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __CUDA_RUNTIME_H__
typedef struct CUstream_st *cudaStream_t;
#endif

#define GSB_OK 0
#define GSB_ERR_INVALID (-1)
#define GSB_ERR_CUDA (-2)
#define GSB_ERR_UNSUPPORTED (-3)

#define GSB_CAMERA_PINHOLE 0
#define GSB_CAMERA_ORTHO 1
#define GSB_CAMERA_FISHEYE 2

/* library identity / diagnostics */
const char *gsb_last_error(void);
int gsb_abi_version(void);
/* number of kernels launched by this library since load (or since gsb_reset_launch_count) */
int64_t gsb_launch_count(void);
void gsb_reset_launch_count(void);
/* kernel-variant switches, for A/B measurements only (the defaults are the tuned path):
 *   GSB_OPT_FWD_VARIANT  1 = rasterize_fwd.cu (warp per 8x4 block, scalar FP32)
 *                        2 = rasterize_fwd2.cu (half-warp per block, 2 pixels per lane, packed FFMA2) [default, D <= 4]
 *                        3 = the same compositor fed by a producer warp (double-buffered staging, named barriers)
 *   GSB_OPT_BWD_VARIANT  1 = rasterize_bwd.cu (warp per block, two-phase moment reduction)
 *                        2 = rasterize_bwd2.cu (half-warp per block, 2 pixels per lane, packed FFMA2) [default, D <= 4]
 *   GSB_OPT_ISECT_BIN_RADIX  1 = gsb_isect_bin ranks every chunk with its radix fallback (tests; default 0)
 *   GSB_OPT_SH_STAGING   0 = gsb_sh_colors_fwd / _bwd stage the coefficient rows with ordinary loads and a CTA barrier
 *                        1 = cp.async.bulk + mbarrier staging, persistent CTAs (degree 3, K = 16)
 *                        2 = cp.async (16-byte pieces, warp-cooperative) staging, persistent CTAs [default; degree 3,
 *                            K = 16, 16-byte aligned rows; anything else takes 0]
 *   GSB_OPT_FINALIZE_VEC 0 = gsb_isect_finalize handles one intersection per thread [default]
 *                        1 = four per thread with 16-byte accesses (measured slower: 61 vs 47 us at cfg2)
 *   GSB_OPT_SH_CTAS      n > 0 caps the persistent grid of the staged backward colour kernel at n CTAs per SM (0 = all
 *                        that fit); GSB_OPT_SH_FWD_STAGING / GSB_OPT_SH_FWD_CTAS: the same two switches for the forward
 *                        kernel (a persistent CTA never gives its SM slot back: beside another stream's kernels the
 *                        caps decide who waits)
 *   GSB_OPT_CHAINED_LAUNCH  mask of kernel groups launched with programmatic stream serialization (a grid becomes
 *                        resident while its predecessor in the stream drains and waits at `griddepcontrol.wait`):
 *                        1 = intersection pipeline incl. the radix sort passes, 2 = projection, 4 = staged colour
 *                        kernels, 8 = splat records, 16 = forward compositor v2, 32 = backward compositor v2.
 *                        Default 37 (1 + 4 + 32), measured at cfg2: 0 -> 1.624 ms per step, 1 -> 1.619, 5 -> 1.606,
 *                        37 -> 1.603, 13 -> 1.637, 63 -> 1.628 */
#define GSB_OPT_FWD_VARIANT 0
#define GSB_OPT_BWD_VARIANT 1
#define GSB_OPT_ISECT_BIN_RADIX 2
#define GSB_OPT_SH_STAGING 3
#define GSB_OPT_FINALIZE_VEC 4
#define GSB_OPT_SH_CTAS 5
#define GSB_OPT_SH_FWD_STAGING 6
#define GSB_OPT_SH_FWD_CTAS 7
#define GSB_OPT_CHAINED_LAUNCH 8
#define GSB_OPT_COUNT 9
int gsb_set_option(int which, int value);
int gsb_get_option(int which);

/* ---- fully_fused_projection ------------------------------------------------------------------
 * replaces triton_impl/fused_projection_fwd.py:232-300 (launcher) / :18-228 (kernel) and
 * cuda/csrc/ProjectionEWA3DGSFused.cu:16-260. means[N,3]; covars[N,6] (triu) or NULL;
 * quats[N,4] wxyz, scales[N,3] (used when covars==NULL); viewmats[C,16]; Ks[C,9].
 * outputs: radii int32[C,N], means2d[C,N,2], depths[C,N], conics[C,N,3], compensations[C,N] or NULL.
 * slots with radii==0 get zeros in means2d/conics/compensations (depths always written). */
int gsb_projection_fwd(int C, int N, const float *means, const float *covars, const float *quats,
                       const float *scales, const float *viewmats, const float *Ks, int width, int height,
                       float eps2d, float near_plane, float far_plane, float radius_clip, int camera_model,
                       int32_t *radii, float *means2d, float *depths, float *conics, float *compensations,
                       cudaStream_t stream);

/* replaces triton_impl/fused_projection_bwd.py:367-450 / :26-363 and ProjectionEWA3DGSFused.cu:266-500.
 * Any of v_means/v_covars/v_quats/v_scales/v_viewmats may be NULL (not needed). Outputs are
 * fully written (no pre-zeroing needed; v_viewmats is zeroed internally). */
int gsb_projection_bwd(int C, int N, const float *means, const float *covars, const float *quats,
                       const float *scales, const float *viewmats, const float *Ks, int width, int height,
                       float eps2d, int camera_model, const int32_t *radii, const float *conics,
                       const float *compensations, const float *v_means2d, const float *v_depths,
                       const float *v_conics, const float *v_compensations, float *v_means, float *v_covars,
                       float *v_quats, float *v_scales, float *v_viewmats, cudaStream_t stream);

/* packed (COO) projection: replaces cuda/csrc/ProjectionEWA3DGSPacked.cu:17-233 and
 * Projection.cpp:254-395. Two passes around a caller-side scan of block_counts
 * (gsb_scan_i32_to_i64): block_counts int32[C * ceil(N/256)], camera-major. Row order of
 * the outputs is row-major (camera, then Gaussian). camera_ids/gaussian_ids are int64. */
int gsb_projection_packed_count(int C, int N, const float *means, const float *covars, const float *quats,
                                const float *scales, const float *viewmats, const float *Ks, int width,
                                int height, float eps2d, float near_plane, float far_plane, float radius_clip,
                                int camera_model, int32_t *block_counts, cudaStream_t stream);
int gsb_projection_packed_write(int C, int N, const float *means, const float *covars, const float *quats,
                                const float *scales, const float *viewmats, const float *Ks, int width,
                                int height, float eps2d, float near_plane, float far_plane, float radius_clip,
                                int camera_model, const int64_t *block_offsets, int64_t *camera_ids,
                                int64_t *gaussian_ids, int32_t *radii, float *means2d, float *depths,
                                float *conics, float *compensations, cudaStream_t stream);
/* replaces ProjectionEWA3DGSPacked.cu:349-560. dense: acc is a float[N,9] scratch and the
 * outputs are [N,*]; sparse_grad: outputs are per-row [nnz,*] (COO values), acc unused. */
int gsb_projection_packed_bwd(int C, int N, int64_t nnz, const float *means, const float *covars,
                              const float *quats, const float *scales, const float *viewmats, const float *Ks,
                              int width, int height, float eps2d, int camera_model, const int64_t *camera_ids,
                              const int64_t *gaussian_ids, const float *conics, const float *compensations,
                              const float *v_means2d, const float *v_depths, const float *v_conics,
                              const float *v_compensations, int sparse_grad, float *acc, float *v_means,
                              float *v_covars, float *v_quats, float *v_scales, float *v_viewmats,
                              cudaStream_t stream);

/* ---- spherical_harmonics ----------------------------------------------------------------------
 * replaces triton_impl/sh_fwd.py:194-230 / sh_bwd.py:383-430 and SphericalHarmonicsCUDA.cu:374/448.
 * dirs[n_batch, n_rows, 3], masks uint8[n_batch, n_rows] or NULL, colors[n_batch, n_rows, 3];
 * coeffs[n_rows, K, 3] is shared by all n_batch entries (the stride-0 camera broadcast of
 * rendering.py:402); per-view coefficients are passed as n_batch=1, n_rows=C*N.
 * Masked-out entries produce 0. */
int gsb_sh_fwd(int degree, int64_t n_rows, int n_batch, int K, const float *dirs, const float *coeffs,
               const uint8_t *masks, float *colors, cudaStream_t stream);
/* v_coeffs[n_rows,K,3] fully written (sum over the batch); v_dirs[n_batch,n_rows,3] or NULL. */
int gsb_sh_bwd(int degree, int64_t n_rows, int n_batch, int K, const float *dirs, const float *coeffs,
               const uint8_t *masks, const float *v_colors, float *v_coeffs, float *v_dirs,
               cudaStream_t stream);

/* Fused view-dependent colour = the torch glue of rendering.py:384-408 folded into the SH kernels (unpacked rows,
 * per-Gaussian coefficients): dirs = means[n] - campos[c]; colors[c,n] = max(SH(dirs, coeffs[n]) + 0.5, 0), with
 * SH := 0 where radii[c,n] <= 0 (masks = radii > 0). campos[C,3] = torch.inverse(viewmats)[:, :3, 3]
 * (gsb_camera_centers for affine matrices). Backward: v_coeffs[N,K,3] and v_means[N,3] (or NULL), summed over
 * cameras, fully written; the clamp passes gradient where SH + 0.5 >= 0. */
int gsb_camera_centers(int C, const float *viewmats, float *campos, cudaStream_t stream);
int gsb_sh_colors_fwd(int degree, int64_t N, int C, int K, const float *means, const float *campos,
                      const float *coeffs, const int32_t *radii, float *colors, cudaStream_t stream);
int gsb_sh_colors_bwd(int degree, int64_t N, int C, int K, const float *means, const float *campos,
                      const float *coeffs, const int32_t *radii, const float *v_colors, float *v_coeffs,
                      float *v_means, cudaStream_t stream);

/* ---- isect_tiles / isect_offset_encode --------------------------------------------------------
 * replaces triton_impl/isect_tiles.py:14-133, radix_sort/radix_sort.cu:10-62, isect_offset.py:9-63
 * and cuda/csrc/IntersectTile.cu:24-344. P = C*N (unpacked) or nnz (packed). */
int64_t gsb_isect_num_blocks(int64_t P); /* rows of block_sums / block_offsets */
int gsb_isect_count(int64_t P, const float *means2d, const int32_t *radii, int tile_size, int tile_width,
                    int tile_height, int32_t *tiles_per_gauss, int32_t *block_sums, int64_t *block_offsets,
                    int64_t *n_isects_out, cudaStream_t stream);
int gsb_scan_i32_to_i64(int64_t n, const int32_t *in, int64_t *out, int64_t *total_out, cudaStream_t stream);
int gsb_isect_emit(int64_t P, int N, int packed, const float *means2d, const int32_t *radii,
                   const float *depths, const int64_t *camera_ids, int tile_size, int tile_width,
                   int tile_height, int tile_n_bits, const int32_t *tiles_per_gauss,
                   const int64_t *block_offsets, int64_t *isect_ids, int32_t *flatten_ids, cudaStream_t stream);
int64_t gsb_sort_pairs_temp_bytes(int64_t n);
int gsb_sort_pairs(int64_t n, int end_bit, int64_t *keys, int64_t *keys_alt, int32_t *vals, int32_t *vals_alt,
                   void *temp, int64_t temp_bytes, int *result_in_alt_host, cudaStream_t stream);
int gsb_isect_offsets(int64_t n_isects, const int64_t *isect_ids, int C, int tile_width, int tile_height,
                      int tile_n_bits, int32_t *offsets, cudaStream_t stream);

/* Depth-presorted isect_tiles (sort=True), bit-identical to a stable radix sort of the emitted 64-bit keys:
 *   1. gsb_isect_row_keys: key = camera << 32 | depth bits per ROW, rows = 0..P-1; the caller sorts them with
 *      gsb_sort_pairs over 32 + camera_bits bits (P rows, not n_isects intersections). keys may be NULL (only
 *      rows is written): with one camera the depth bit patterns themselves are the keys (gsb_sort_pairs_u32);
 *   2. gsb_isect_count_perm / gsb_isect_emit_perm: tile counts and (camera | tile) 32-bit words + row ids emitted
 *      in that row order (perm = the sorted rows);
 *   3. gsb_sort_pairs_u32 over tile_bits + camera_bits bits only (2 radix passes at 1080p instead of 6);
 *   4. gsb_isect_finalize: rebuilds the 64-bit isect_ids (word << 32 | depth bits of the row) and writes
 *      isect_offsets (the output of isect_offset_encode) in the same pass.
 * Depths must be positive (as projection produces; a negative depth sign-extends into the tile bits in the
 * reference's own key). tile_bits + camera_bits must be <= 32. */
int gsb_isect_row_keys(int64_t P, int N, int packed, const float *depths, const int64_t *camera_ids, int64_t *keys,
                       int32_t *rows, cudaStream_t stream);
/* natural-order (coalesced) part of steps 1-2: rows[i] = i, tiles_per_gauss[i] = tile count of row i and (optional)
 * row_records int32[P,4] = (x0 | y0 << 16, width, camera << tile_n_bits, count) for gsb_isect_emit_perm_cap. With it,
 * gsb_isect_count_perm may be called with means2d = radii = NULL and only gathers the 4-byte counts. */
int gsb_isect_row_tiles(int64_t P, int N, int packed, const float *means2d, const int32_t *radii,
                        const int64_t *camera_ids, int tile_size, int tile_width, int tile_height, int tile_n_bits,
                        int32_t *rows, int32_t *tiles_per_gauss, int32_t *row_records, cudaStream_t stream);
int gsb_isect_count_perm(int64_t P, const int32_t *perm, const float *means2d, const int32_t *radii, int tile_size,
                         int tile_width, int tile_height, int32_t *tiles_per_gauss, int32_t *block_sums,
                         int64_t *block_offsets, int64_t *n_isects_out, cudaStream_t stream);
int gsb_isect_emit_perm(int64_t P, int N, int packed, const int32_t *perm, const float *means2d,
                        const int32_t *radii, const int64_t *camera_ids, int tile_size, int tile_width,
                        int tile_height, int tile_n_bits, const int32_t *tiles_per_gauss,
                        const int64_t *block_offsets, int32_t *tile_keys, int32_t *flatten_ids, cudaStream_t stream);
int64_t gsb_sort_pairs_u32_temp_bytes(int64_t n);
int gsb_sort_pairs_u32(int64_t n, int end_bit, int32_t *keys, int32_t *keys_alt, int32_t *vals, int32_t *vals_alt,
                       void *temp, int64_t temp_bytes, int *result_in_alt_host, cudaStream_t stream);
int gsb_isect_finalize(int64_t n_isects, const int32_t *tile_keys, const int32_t *flatten_ids, const float *depths,
                       int C, int tile_width, int tile_height, int tile_n_bits, int64_t *isect_ids, int32_t *offsets,
                       cudaStream_t stream);
/* Speculatively sized variants (no host read-back of n_isects before the compositor is launched): buffers hold
 * `capacity` slots, the real count stays on the device (`n_isects_dev`, the total written by gsb_isect_count_perm).
 * emit writes only slots < capacity and pads [n_isects, capacity) with 0xffffffff keys (they sort last); finalize
 * treats min(*n_isects_dev, capacity) as the count (isect_ids or offsets may be NULL: only the other output is
 * written, so the ids, which the compositor does not read, can be produced on a second stream). The caller reads the count back later and, if it exceeded
 * the capacity, redoes the step with exact sizes. row_records (or NULL): the per-row records of
 * gsb_isect_row_tiles; with them emit gathers one 16-byte record per row instead of means2d/radii/camera_ids. */
int gsb_isect_emit_perm_cap(int64_t P, int N, int packed, const int32_t *perm, const float *means2d,
                            const int32_t *radii, const int64_t *camera_ids, int tile_size, int tile_width,
                            int tile_height, int tile_n_bits, const int32_t *tiles_per_gauss,
                            const int64_t *block_offsets, int32_t *tile_keys, int32_t *flatten_ids,
                            int64_t capacity, const int64_t *n_isects_dev, const int32_t *row_records,
                            cudaStream_t stream);
int gsb_isect_finalize_cap(int64_t n_isects, const int64_t *n_isects_dev, const int32_t *tile_keys,
                           const int32_t *flatten_ids, const float *depths, int C, int tile_width, int tile_height,
                           int tile_n_bits, int64_t *isect_ids, int32_t *offsets, cudaStream_t stream);

/* Tile binning (csrc/isect_bin.cu): steps 2-4 above without a sort over the intersections, for C * tiles <=
 * gsb_isect_bin_max_bins() (8192) bins; same isect_ids / flatten_ids / offsets, bit for bit
 * (isect_tiles.py:14-133 + isect_offset.py:9-63; IntersectTile.cu:24-344).
 *   gsb_isect_row_tiles_keys      = gsb_isect_row_tiles + row_keys[i] = depth bits of row i, or 0xffffffff for a row
 *                                   without tiles: sorting (row_keys, rows) over 32 bits puts the emitting rows first;
 *   gsb_isect_count_perm_records  gathers the row records into that order (sorted_records int32[P,4]) and scans the
 *                                   tile counts per 256 rows (block_offsets int64[blocks+1], total -> n_isects_out);
 *   gsb_isect_bin                 counting sort with one bin per (camera, tile): per-chunk counts, scan, then every
 *                                   intersection is written once at its final place. `capacity` = slots of isect_ids /
 *                                   flatten_ids; the real count may stay on the device (n_isects_dev, or NULL when
 *                                   capacity is the count). temp: gsb_isect_bin_temp_bytes(capacity, C * tiles). */
int gsb_isect_row_tiles_keys(int64_t P, int N, int packed, const float *means2d, const int32_t *radii,
                             const float *depths, const int64_t *camera_ids, int tile_size, int tile_width,
                             int tile_height, int tile_n_bits, int32_t *rows, int32_t *tiles_per_gauss,
                             int32_t *row_records, int32_t *row_keys, void *key_hist, cudaStream_t stream);
/* Producers that count the digits of the keys they write, so that the sort that follows skips its histogram pass
 * (gsb_radix_prepare + gsb_radix_sort_pairs_u32_prepared): key_hist = the sort's temp buffer (or NULL). row_tiles_keys
 * counts the 4 x 8-bit digits of row_keys; emit_perm_hist those of the key_bits-bit (camera | tile) words it emits
 * (otherwise = gsb_isect_emit_perm_cap). */
int gsb_isect_emit_perm_hist(int64_t P, int N, int packed, const int32_t *perm, const float *means2d,
                             const int32_t *radii, const int64_t *camera_ids, int tile_size, int tile_width,
                             int tile_height, int tile_n_bits, const int32_t *tiles_per_gauss,
                             const int64_t *block_offsets, int32_t *tile_keys, int32_t *flatten_ids,
                             int64_t capacity, const int64_t *n_isects_dev, const int32_t *row_records,
                             void *key_hist, int key_bits, cudaStream_t stream);
int gsb_isect_count_perm_records(int64_t P, const int32_t *perm, const int32_t *row_records, int32_t *sorted_records,
                                 int32_t *block_sums, int64_t *block_offsets, int64_t *n_isects_out,
                                 cudaStream_t stream);
/* The emission step of the default pipeline on the sorted records of gsb_isect_count_perm_records: chunks of 8192
 * slots of the emission order, owner rows from a marker max-scan instead of a search per slot; writes the same
 * tile_keys / flatten_ids as gsb_isect_emit_perm_cap and (key_hist = the tile sort's prepared temp buffer, or NULL)
 * counts the digits of the key_bits-bit words. */
int gsb_isect_emit_chunks(int64_t P, int64_t capacity, const int64_t *n_isects_dev, const int32_t *sorted_records,
                          const int32_t *perm, const int64_t *block_offsets, int tile_width, int32_t *tile_keys,
                          int32_t *flatten_ids, void *key_hist, int key_bits, cudaStream_t stream);
int gsb_isect_bin_max_bins(void);
int64_t gsb_isect_bin_temp_bytes(int64_t capacity, int bins);
int gsb_isect_bin(int64_t P, int64_t capacity, const int64_t *n_isects_dev, const int32_t *sorted_records,
                  const int32_t *perm, const int32_t *sorted_keys, const int64_t *block_offsets, int C, int tile_width,
                  int tile_height, int tile_n_bits, void *temp, int64_t temp_bytes, int64_t *isect_ids,
                  int32_t *flatten_ids, int32_t *offsets, cudaStream_t stream);

/* ---- rasterize_to_pixels ----------------------------------------------------------------------
 * replaces triton_impl/rasterize_to_pixels_fwd.py:200-275 / :21-196, rasterize_to_pixels_bwd.py:341-430
 * / :21-337 and cuda/csrc/RasterizeToPixels3DGSFwd.cu:18-188, RasterizeToPixels3DGSBwd.cu:16-278.
 * P rows of projected Gaussians (C*N or nnz); D colour channels, D in {1,2,3,4,8,16,32}
 * (the Python wrapper zero-pads other widths like the reference does); tile_size must be 16.
 * gsb_rasterize_pack builds the 2 x float4 splat records the compositor gathers. */
int gsb_rasterize_pack(int64_t P, const float *means2d, const float *conics, const float *opacities,
                       float *recA, float *recB, cudaStream_t stream);
/* backgrounds[C,D] or NULL; masks uint8[C,th,tw] or NULL (true = render, false = background only);
 * outputs render_colors[C,H,W,D], render_alphas[C,H,W], last_ids int32[C,H,W]. */
int gsb_rasterize_fwd(int C, int64_t n_isects, int D, const float *recA, const float *recB, const float *colors,
                      const float *backgrounds, const uint8_t *masks, int width, int height, int tile_size,
                      int tile_width, int tile_height, const int32_t *isect_offsets, const int32_t *flatten_ids,
                      float *render_colors, float *render_alphas, int32_t *last_ids, cudaStream_t stream);
/* same, for speculatively sized intersection buffers: n_isects = capacity of flatten_ids, the real count is
 * min(*n_isects_dev, n_isects) (NULL: n_isects is exact). See gsb_isect_emit_perm_cap.
 * block_masks uint8[n_isects] or NULL: receives, for every intersection the pass stages, which of the tile's eight
 * 8x4 pixel blocks the Gaussian can reach (0 = culled); gsb_rasterize_bwd reuses it instead of recomputing. */
int gsb_rasterize_fwd_cap(int C, int64_t n_isects, const int64_t *n_isects_dev, int D, const float *recA,
                          const float *recB, const float *colors, const float *backgrounds, const uint8_t *masks,
                          int width, int height, int tile_size, int tile_width, int tile_height,
                          const int32_t *isect_offsets, const int32_t *flatten_ids, float *render_colors,
                          float *render_alphas, int32_t *last_ids, uint8_t *block_masks, cudaStream_t stream);
/* diagnostics: counts[C,H,W] = number of Gaussians composited into each pixel (ones: float[P] of 1.0f);
 * same traversal as gsb_rasterize_fwd with D = 1. warp_pairs (zero-filled int64[1] or NULL) receives the number of
 * (warp, splat) pairs the kernel walked: x 32 lanes = the pixel-Gaussian evaluations it EXECUTED (bench.py's
 * executed-work roofline). */
int gsb_rasterize_count(int C, int64_t n_isects, const float *recA, const float *recB, const float *ones, int width,
                        int height, int tile_size, int tile_width, int tile_height, const int32_t *isect_offsets,
                        const int32_t *flatten_ids, float *counts, float *render_alphas, int32_t *last_ids,
                        int64_t *warp_pairs, cudaStream_t stream);
/* Any other tile size (1 <= tile_size^2 <= 1024; the reference's tests use 4): the plain one-thread-per-pixel
 * compositor of csrc/rasterize_generic.cu on the unpacked means2d / conics / opacities (no splat records). Same
 * contracts as gsb_rasterize_fwd / gsb_rasterize_bwd (gradient buffers zero-filled by the caller). */
int gsb_rasterize_fwd_generic(int C, int64_t n_isects, int D, const float *means2d, const float *conics,
                              const float *opacities, const float *colors, const float *backgrounds,
                              const uint8_t *masks, int width, int height, int tile_size, int tile_width,
                              int tile_height, const int32_t *isect_offsets, const int32_t *flatten_ids,
                              float *render_colors, float *render_alphas, int32_t *last_ids, cudaStream_t stream);
int gsb_rasterize_bwd_generic(int C, int64_t n_isects, int D, const float *means2d, const float *conics,
                              const float *opacities, const float *colors, const float *backgrounds,
                              const uint8_t *masks, int width, int height, int tile_size, int tile_width,
                              int tile_height, const int32_t *isect_offsets, const int32_t *flatten_ids,
                              const float *render_alphas, const int32_t *last_ids, const float *v_render_colors,
                              const float *v_render_alphas, float *v_means2d_abs, float *v_means2d, float *v_conics,
                              float *v_colors, float *v_opacities, cudaStream_t stream);
/* diagnostics: iterations of the v2 forward kernel's inner loop (x 64 lanes = executed evaluations), D = 1 */
int gsb_rasterize_fwd2_count(int C, int64_t n_isects, const float *recA, const float *recB, const float *ones,
                             int width, int height, int tile_width, int tile_height, const int32_t *isect_offsets,
                             const int32_t *flatten_ids, float *scratch_colors, float *scratch_alphas,
                             int32_t *scratch_last, int64_t *warp_iters, cudaStream_t stream);
/* v_means2d[P,2], v_conics[P,3], v_colors[P,D], v_opacities[P] must be ZERO-filled by the caller
 * (the kernel accumulates with atomics); v_means2d_abs[P,2] (zero-filled) or NULL.
 * block_masks: what gsb_rasterize_fwd_cap wrote for the same inputs, or NULL (recomputed). */
int gsb_rasterize_bwd(int C, int64_t n_isects, int D, const float *recA, const float *recB, const float *colors,
                      const float *backgrounds, const uint8_t *masks, int width, int height, int tile_size,
                      int tile_width, int tile_height, const int32_t *isect_offsets, const int32_t *flatten_ids,
                      const float *render_alphas, const int32_t *last_ids, const uint8_t *block_masks,
                      const float *v_render_colors, const float *v_render_alphas, float *v_means2d_abs,
                      float *v_means2d, float *v_conics, float *v_colors, float *v_opacities, cudaStream_t stream);

/* ---- rasterize_to_indices_in_range (cuda/_wrapper.py:577-648, RasterizeToIndices3DGS.cu:14-170) ---------------
 * the (Gaussian, pixel) pairs the compositor blends for sorted-intersection batches [range_start, range_end) of
 * every tile (batch = tile_size^2 intersections), starting from transmittances[C,H,W]. Pass 1 writes counts[C*H*W];
 * the caller scans them (exclusive, int64) into pixel_starts and sizes the outputs; pass 2 writes
 * gaussian_ids (within-camera id) and indices (camera*H*W + pixel), pixel-major, sorted order inside a pixel.
 * Any tile_size (the grid of isect_offsets must cover the image). */
int gsb_rasterize_indices_count(int C, int N, int64_t n_isects, int range_start, int range_end,
                                const float *transmittances, const float *means2d, const float *conics,
                                const float *opacities, int width, int height, int tile_size, int tile_width,
                                int tile_height, const int32_t *isect_offsets, const int32_t *flatten_ids,
                                int32_t *counts, cudaStream_t stream);
int gsb_rasterize_indices_write(int C, int N, int64_t n_isects, int range_start, int range_end,
                                const float *transmittances, const float *means2d, const float *conics,
                                const float *opacities, int width, int height, int tile_size, int tile_width,
                                int tile_height, const int32_t *isect_offsets, const int32_t *flatten_ids,
                                const int64_t *pixel_starts, int64_t *gaussian_ids, int64_t *indices,
                                cudaStream_t stream);

/* ---- selective fused Adam (optimizers/selective_adam.py, AdamCUDA.cu:13-52) ------------------------------------
 * in-place update of param / exp_avg / exp_avg_sq [N, M] for the Gaussians with visibility[n] != 0 (uint8[N], or
 * NULL = all): m = b1 m + (1-b1) g, v = b2 v + (1-b2) g^2, p -= lr m / (sqrt(v) + eps); no bias correction. */
int gsb_selective_adam(int64_t N, int64_t M, float *param, const float *grad, float *exp_avg, float *exp_avg_sq,
                       const uint8_t *visibility, float lr, float b1, float b2, float eps, cudaStream_t stream);

/* ---- MCMC relocation (relocation.py:10-50, RelocationCUDA.cu:13-45) ---------------------------------------------
 * per Gaussian split into ratios[n] (1..n_max) copies: new_opacities [N], new_scales [N,3]; binoms [n_max,n_max]. */
int gsb_relocation(int64_t N, const float *opacities, const float *scales, const int32_t *ratios,
                   const float *binoms, int n_max, float *new_opacities, float *new_scales, cudaStream_t stream);

/* ---- stable LSD radix sort of (key, int32 value) pairs, hand-written onesweep (csrc/radix_sort.cu) ------------
 * replaces triton_impl/radix_sort/radix_sort.cu:10-62 / cub::DeviceRadixSort in IntersectTile.cu. Sorts over key
 * bits [0, end_bit), 8 bits per pass, one kernel per pass + one histogram kernel + one memset of `temp`
 * (gsb_radix_temp_bytes(n, end_bit) bytes, caller-allocated). n < 2^30. `n_dev` (NULL or a device int64): the
 * number of valid leading items is min(n, *n_dev) - the rest of the buffers is neither read nor written, so a
 * speculatively sized caller pays for the real count only. keys_alt / vals_alt: ping-pong buffers of n items;
 * *result_in_alt_host (host int, written at call time) tells which pair of buffers holds the result. */
int64_t gsb_radix_temp_bytes(int64_t n, int end_bit);
int gsb_radix_sort_pairs_u32(int64_t n, const int64_t *n_dev, int end_bit, int32_t *keys, int32_t *keys_alt,
                             int32_t *vals, int32_t *vals_alt, void *temp, int64_t temp_bytes,
                             int *result_in_alt_host, cudaStream_t stream);
int gsb_radix_sort_pairs_u64(int64_t n, const int64_t *n_dev, int end_bit, int64_t *keys, int64_t *keys_alt,
                             int32_t *vals, int32_t *vals_alt, void *temp, int64_t temp_bytes,
                             int *result_in_alt_host, cudaStream_t stream);
/* The producer of the keys may count the digits itself and spare the sort its histogram pass: gsb_radix_prepare clears
 * `temp`; the producer adds, for every key and every pass p < ceil(end_bit / 8), one count to
 * ((uint32_t *)temp)[p * 256 + ((key >> 8 p) & mask_p)] (mask_p = 255, or the remaining bits in the last pass);
 * gsb_radix_sort_pairs_u32_prepared then runs the passes only (no memset, no histogram kernel). */
int gsb_radix_prepare(int64_t n, int end_bit, void *temp, int64_t temp_bytes, cudaStream_t stream);
int gsb_radix_sort_pairs_u32_prepared(int64_t n, const int64_t *n_dev, int end_bit, int32_t *keys, int32_t *keys_alt,
                                      int32_t *vals, int32_t *vals_alt, void *temp, int64_t temp_bytes,
                                      int *result_in_alt_host, cudaStream_t stream);

/* The 32-bit sort for callers that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc). NO COUNTERPART IN THE REFERENCE.
 * Same result as gsb_radix_sort_pairs_u32, but the scratch is a file-scope device array inside the library (2.1 MB per
 * GPU that loads it, held until exit) instead of a caller-provided `temp`: the call allocates nothing (no cudaMalloc /
 * cudaFree, no allocating library routine). Limit: n <= GSB_RADIX_NOALLOC_MAX_ITEMS pairs, larger n is rejected with
 * GSB_ERR_INVALID. The caller still passes the keys / values and their ping-pong buffers. Every call on a device shares
 * the one array (cleared by the call itself on `stream`): calls that may OVERLAP on the GPU (different streams without
 * ordering) race on it. The current device must be the one the buffers live on. */
#define GSB_RADIX_NOALLOC_MAX_ITEMS (1 << 21)
int gsb_radix_sort_pairs_u32_noalloc(int64_t n, const int64_t *n_dev, int end_bit, int32_t *keys, int32_t *keys_alt,
                                     int32_t *vals, int32_t *vals_alt, int *result_in_alt_host, cudaStream_t stream);

/* ---- distributed exchange rows (rendering.py:413-494, distributed.py:170-257) -----------------
 * One fused row per projected Gaussian: means2d(2) depths(1) conics(3) opacities(1) colors(D)
 * radii(1, int32 bit pattern) = (8+D) 4-byte words, so a single all-to-all per direction
 * replaces the reference's cat/split + separate int32 exchange. */
int gsb_a2a_pack(int64_t rows, int D, const float *means2d, const float *depths, const float *conics,
                 const float *opacities, const float *colors, const int32_t *radii, float *packed,
                 cudaStream_t stream);
int gsb_a2a_unpack(int64_t rows, int D, const float *packed, float *means2d, float *depths, float *conics,
                   float *opacities, float *colors, int32_t *radii, cudaStream_t stream);

/* ---- fused pack + peer-store exchange over NVLink (one process per GPU) -----------------------
 * replaces the NCCL all_to_all + cat/split/reshape_view sequence of rendering.py:453-494 /
 * distributed.py:170-257 for ranks of one NVLink domain. Each rank owns a "mailbox" allocation
 * (gsb_peer_alloc: the one place this library allocates, because CUDA IPC needs its own cudaMalloc
 * block) whose IPC handle (gsb_peer_handle_bytes bytes) the caller distributes to the peers
 * (torch.distributed all_gather); peers map it with gsb_peer_open_handle. Layout of a mailbox:
 * gsb_peer_header_bytes() of flags, then 4 regions of equal capacity, region_offset =
 * header + (2 * dir + (epoch & 1)) * capacity, dir 0 = forward, 1 = backward.
 * A region holding `rows` rows is a structure of arrays: means2d(2) depths(1) conics(3) opacities(1)
 * colours(D) [radii(1), forward only], field f starting at float index gsb_a2a_field_offset(rows, D, f)
 * (each field 16-byte aligned; f = 6 or 5 gives the region's total floats with / without radii).
 *   gsb_a2a_put_fwd: every (camera, field) run of the local [world*C_local, N_local] projection is stored
 *     directly into the mailbox of the rank owning the camera, at rows [c_local, n_base + n] of its
 *     [C_local, N_total] table; the last CTA publishes `epoch` (> 0, increasing) to every peer.
 *   gsb_a2a_put_bwd: gradient runs of the local [C_local, N_total] table go to the rank owning each Gaussian
 *     (n_bases_host[world+1]: first global Gaussian of every rank), at rows [rank*C_local + c_local, n - base]
 *     of its [world*C_local, N_r] table. NULL gradient pointers send zeros.
 *   gsb_a2a_wait_copy: waits (device side, bounded: *timeout_flag, a host-mapped int from
 *     gsb_peer_flag_alloc, becomes 1 + the rank that never arrived) for all `world` flags of direction
 *     `dir` to reach `epoch`, then copies the local region (n_floats, multiple of 4) into `out`, whose
 *     field views are the SoA tensors.
 * `peer_bases_host`: host array of `world` device pointers (own mailbox at [rank]). */
#define GSB_MAX_PEERS 16
int64_t gsb_peer_header_bytes(void);
int gsb_peer_handle_bytes(void);
int gsb_peer_alloc(int64_t bytes, void **dptr_out_host);
int gsb_peer_free(void *dptr);
int gsb_peer_get_handle(void *dptr, void *handle_out_host);
int gsb_peer_open_handle(const void *handle_host, void **dptr_out_host);
int gsb_peer_close_handle(void *dptr);
int gsb_peer_flag_alloc(void **flag_out_host);
int gsb_peer_flag_free(void *flag_host);
int gsb_a2a_put_fwd(int world, int rank, int C_local, int64_t N_local, int64_t N_total, int64_t n_base, int D,
                    const float *means2d, const float *depths, const float *conics, const float *opacities,
                    const float *colors, const int32_t *radii, const void *const *peer_bases_host,
                    int64_t region_offset, int64_t epoch, cudaStream_t stream);
int gsb_a2a_put_bwd(int world, int rank, int C_local, int64_t N_total, const int64_t *n_bases_host, int D,
                    const float *v_means2d, const float *v_depths, const float *v_conics, const float *v_opacities,
                    const float *v_colors, const void *const *peer_bases_host, int64_t region_offset,
                    int64_t epoch, cudaStream_t stream);
int gsb_a2a_wait_copy(int world, int dir, int64_t epoch, const void *local_base, int64_t region_offset,
                      int64_t n_floats, float *out, int *timeout_flag, cudaStream_t stream);
/* Split exchange: the same puts restricted to a subset of the fields (fields_mask: bit f = field f in the order
 * means2d, depths, conics, opacities, colours, radii) and published on their own flag channel (0 forward geometry /
 * all, 1 backward geometry / all, 2 forward colours, 3 backward colours), so that the colours can travel on a side
 * stream under the receiver's intersection pipeline, which needs the geometry only. gsb_a2a_wait_copy_part waits for
 * `channel` and copies the float ranges [begin0, end0) and [begin1, end1) of the region (multiples of 4 floats:
 * field offsets are) into the same positions of `out`. */
int gsb_a2a_put_fwd_part(int world, int rank, int C_local, int64_t N_local, int64_t N_total, int64_t n_base, int D,
                         const float *means2d, const float *depths, const float *conics, const float *opacities,
                         const float *colors, const int32_t *radii, int fields_mask, int channel,
                         const void *const *peer_bases_host, int64_t region_offset, int64_t epoch,
                         cudaStream_t stream);
int gsb_a2a_put_bwd_part(int world, int rank, int C_local, int64_t N_total, const int64_t *n_bases_host, int D,
                         const float *v_means2d, const float *v_depths, const float *v_conics,
                         const float *v_opacities, const float *v_colors, int fields_mask, int channel,
                         const void *const *peer_bases_host, int64_t region_offset, int64_t epoch,
                         cudaStream_t stream);
int gsb_a2a_wait_copy_part(int world, int channel, int64_t epoch, const void *local_base, int64_t region_offset,
                           int64_t begin0, int64_t end0, int64_t begin1, int64_t end1, float *out,
                           int *timeout_flag, cudaStream_t stream);
int64_t gsb_a2a_field_offset(int64_t rows, int D, int field);
/* Push exchange (unpacked rows): instead of a separate put kernel, gsb_projection_fwd_push / gsb_sh_colors_fwd_push
 * store every row ALSO into the mailbox region of the rank that owns its camera (same layout and positions as
 * gsb_a2a_put_fwd would; opacities[N] x compensation goes into the opacity field), and gsb_a2a_publish, queued
 * after them on the same stream, raises this rank's epoch flag at every peer. The receiver waits with
 * gsb_a2a_wait_copy as before. */
int gsb_projection_fwd_push(int C, int N, const float *means, const float *covars, const float *quats,
                            const float *scales, const float *viewmats, const float *Ks, int width, int height,
                            float eps2d, float near_plane, float far_plane, float radius_clip, int camera_model,
                            int32_t *radii, float *means2d, float *depths, float *conics, float *compensations,
                            const float *opacities, int world, int C_local, int64_t N_total, int64_t n_base, int D,
                            const void *const *peer_bases_host, int64_t region_offset, cudaStream_t stream);
int gsb_sh_colors_fwd_push(int degree, int64_t N, int C, int K, const float *means, const float *campos,
                           const float *coeffs, const int32_t *radii, float *colors, int world, int C_local,
                           int64_t N_total, int64_t n_base, const void *const *peer_bases_host, int64_t region_offset,
                           cudaStream_t stream);
int gsb_a2a_publish(int world, int rank, int channel, int64_t epoch, const void *const *peer_bases_host,
                    cudaStream_t stream);
/* Packed (COO) rows over the same mailboxes (reference rendering.py:413-451: bincount + all_to_all of counts,
 * all_to_all of the rows, a second one for the rebased ids). The sender's packed rows are camera-major, so the rows
 * for destination d are the run after those for d-1: send_counts_host[d] rows go to rows
 * [dst_row_base_host[d], + count) of d's table of dst_rows_host[d] rows (a region with two extra int32 fields after
 * radii: camera id local to d = id - cam_base_host[d], Gaussian id global = id + gauss_base; field offsets from
 * gsb_a2a_field_offset with field 6, 7, 8). All *_host arrays have `world` entries and live on the host.
 * gsb_a2a_put_packed_bwd returns the gradient rows held for sender r (recv_counts_host[r], in sender order) to rows
 * [dst_row_base_host[r], + count) of r's table of dst_rows_host[r] = nnz_r rows. */
int gsb_a2a_put_packed_fwd(int world, int rank, int D, const int64_t *send_counts_host,
                           const int64_t *dst_row_base_host, const int64_t *dst_rows_host,
                           const int64_t *cam_base_host, int64_t gauss_base, const float *means2d,
                           const float *depths, const float *conics, const float *opacities, const float *colors,
                           const int32_t *radii, const int64_t *camera_ids, const int64_t *gaussian_ids,
                           const void *const *peer_bases_host, int64_t region_offset, int64_t epoch,
                           cudaStream_t stream);
int gsb_a2a_put_packed_bwd(int world, int rank, int D, const int64_t *recv_counts_host,
                           const int64_t *dst_row_base_host, const int64_t *dst_rows_host, const float *v_means2d,
                           const float *v_depths, const float *v_conics, const float *v_opacities,
                           const float *v_colors, const void *const *peer_bases_host, int64_t region_offset,
                           int64_t epoch, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* GSPLAT_B200_H */
