/*
 * origami_b200 -- C ABI of the B200 (sm_100a) page-geometry kernels.
 *
 * Drop-in boundary for the hot path of poke1024/origami.  The
 * reference has no FFI for this path: it calls OpenCV / SciPy / scikit-image from Python.
 * Each entry point below names the reference call site (paths relative to the reference
 * root) whose arithmetic it reproduces bit for bit.  INTEGRATION.md shows the ctypes
 * binding a maintainer adds on the reference side.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless the parameter is documented as host;
 *  - the caller owns and allocates every buffer; the library never allocates device
 *    memory, never frees and never keeps a pointer after the call returns;
 *  - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream);
 *    calls only enqueue work, they do not synchronise;
 *  - return value: 0 = ok, < 0 = argument error (OG_E_*), > 0 = a cudaError_t value;
 *    og_last_error() returns a thread-local message for the last non-zero return;
 *  - images are row-major 8-bit gray with a row pitch in BYTES; maps and grids are
 *    contiguous fp32 with the last axis (x, y), exactly the layout of dewarp.zip's
 *    data.npy (reference docs/formats.md:298-306).
 */
#ifndef ORIGAMI_B200_H
#define ORIGAMI_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OG_OK 0
#define OG_E_ARG (-1)      /* null pointer, non-positive size, bad enum            */
#define OG_E_RANGE (-2)    /* dimension beyond what OpenCV itself accepts (>=32767) */
#define OG_E_ALIGN (-3)    /* pointer / pitch alignment the kernel cannot serve     */
#define OG_E_WINDOW (-4)   /* Sauvola window even, 1, or larger than supported      */

#define OG_INTER_NEAREST 0 /* == cv2.INTER_NEAREST */
#define OG_INTER_LINEAR 1  /* == cv2.INTER_LINEAR  */
/* OR-ed into og_remap_page's `interp`: copy each tile's source window to shared memory first
 * (same bits; measured slower than the direct gather on B200, kept as a tuning switch). */
#define OG_FLAG_STAGE_SMEM 0x100

#define OG_BIN_NONE 0
#define OG_BIN_OTSU 1
#define OG_BIN_SAUVOLA 2
/* OR-ed into og_extract_lines' `binarize`: use the map-gather engines even without bounding boxes
 * (what a non-NULL `map` selects anyway; kept as an explicit switch for tests and comparisons). */
#define OG_FLAG_MAP_GATHER 0x100

/* One text line in DEWARPED page space.  Mirrors the fields of the reference's Line
 * (origami/core/block.py:35-55; lines.zip: docs/formats.md:223-246) plus the result of
 * Line._position (block.py:195-221), which the host evaluates because it is scalar work.
 * All vectors are (x, y).  112 bytes, 8-byte aligned. */
typedef struct og_line {
	double p[2];      /* bottom-left corner of the line box (un-clipped)                 */
	double right[2];  /* baseline vector (un-clipped)                                    */
	double up[2];     /* ascender vector, points to smaller y                            */
	double cp[2];     /* origin after table-column clipping (== p without a column)      */
	double cright[2]; /* baseline vector after clipping      (== right without a column) */
	double xres;      /* samples per warped pixel along the line, after column rescale   */
	double yres;      /* samples per warped pixel across the line                        */
	int32_t xsteps;   /* > 0: fixed sample count; 0: derive from the rough 2x2 extent    */
	int32_t ysteps;   /* > 0: fixed (target_height); 0: derive                           */
	int32_t flags;    /* reserved, 0                                                     */
	int32_t pad_;
} og_line;

/* Message for the last non-zero return on this thread (never NULL). */
const char *og_last_error(void);

/* Library / build identification, e.g. "origami_b200 0.1 sm_100a". */
const char *og_version(void);

/* A2  Grid.points("full")  -- origami/core/dewarp.py:514-526 (2x cv2.resize INTER_LINEAR of
 * the x and y planes, np.dstack).  grid (gh, gw, 2) fp32 -> map_out (gh*cell, gw*cell, 2)
 * fp32, bit-identical to OpenCV's generic (cv2.setUseOptimized(False)) code path. */
int og_expand_map(const float *grid, int gh, int gw, int cell, float *map_out, void *stream);

/* A2+A3+A4  Dewarper.dewarped / Dewarper.dewarp_image -- origami/core/dewarp.py:608-627
 * (cv2.remap of the page through Grid.points("full")), fused with the map expansion.
 * src (H, W) u8 -> dst (gh*cell, gw*cell) u8.  The sampling coordinates are the dense map
 * times (scale_x, scale_y) in fp32 (1.0f, 1.0f for Dewarper.dewarped; w1/w0, h1/h0 rounded
 * to fp32 for dewarp_image).  map_out, if not NULL, receives the UNSCALED dense map
 * (same bits as og_expand_map); dst may be NULL when only the map is wanted.
 * interp: OG_INTER_LINEAR (OpenCV's 1/32-px fixed-point bilinear, BORDER_CONSTANT 0) or
 * OG_INTER_NEAREST. */
int og_remap_page(const uint8_t *src, int H, int W, int64_t src_pitch,
	const float *grid, int gh, int gw, int cell,
	uint8_t *dst, int64_t dst_pitch, float *map_out,
	float scale_x, float scale_y, int interp, void *stream);

/* A5  Grid.inverse_yx -- origami/core/dewarp.py:561-567 (SciPy RegularGridInterpolator,
 * linear, extrapolating) evaluated at n points.  pts_yx (n, 2) f64 dewarped (y, x) ->
 * out_xy (n, 2) f64 warped (x, y).  The dense map is not read: its nodes are recomputed
 * from the sparse grid with the A2 arithmetic, then blended in float64 in SciPy's order. */
int og_inverse_yx(const float *grid, int gh, int gw, int cell,
	const double *pts_yx, int64_t n, double *out_xy, void *stream);

/* A8 (step counts) + A21  Line.warped_grid step derivation -- origami/core/block.py:226-236.
 * For every line: steps_out[i] = (xsteps, ysteps); pix_offsets_out[i] = sum over j<i of
 * xsteps*ysteps (n+1 entries, element offsets of each strip in a packed buffer);
 * chunk_offsets_out (n+2 entries) = n+1 prefix sums of the work items the extraction kernels
 * split each line into (64 strip columns each), followed by one work-ticket word that this call
 * zeroes and og_extract_lines / og_sample_labels use while they run (and leave at zero).  A plan
 * therefore serves one such call at a time.  Two small kernels, no host synchronisation.
 * A line whose extent is not finite, or whose strip would exceed 2^31 pixels, gets steps (0, 0)
 * and no work items. */
int og_line_steps(const float *grid, int gh, int gw, int cell,
	const og_line *lines, int n,
	int32_t *steps_out, int64_t *pix_offsets_out, int32_t *chunk_offsets_out, void *stream);

/* The three line entry points below take `map`: the dense map (gh*cell, gw*cell, 2) fp32 that
 * og_expand_map / og_remap_page wrote for the SAME grid, or NULL.  With a map the engines that
 * evaluate every sample's float64 blend (warped grids, bounding boxes, label counts) gather the
 * four nodes from it (the reference, too, interpolates in its cached Grid.points("full")); with
 * NULL they are recomputed from the sparse grid on chip.  Both paths produce identical bits. */

/* A8  Line.warped_grid -- origami/core/block.py:223-248.  Writes the fp32 (ysteps, xsteps, 2)
 * warped sampling grid of every line, packed at 2*pix_offsets[i] floats.  capacity is in
 * grid points (pixels); lines that would not fit are skipped (check pix_offsets[n]). */
int og_warped_grids(const float *grid, int gh, int gw, int cell, const float *map,
	const og_line *lines, int n,
	const int32_t *steps, const int64_t *pix_offsets, const int32_t *chunk_offsets,
	float *grids_out, int64_t capacity, void *stream);

/* A9+A10+A12+A21  Line.image = Line.dewarped_image + binarizer(im) -- origami/core/block.py:129-143,
 * 250-259, origami/core/binarize.py:11-43, batched over all lines of a page in one launch (replaces
 * LineExtractor's ThreadPool, origami/batch/core/lines.py:148-149).  src is the WARPED page.  Strip i
 * is written as (ysteps, xsteps) u8, row 0 = top of the text (the reference's pixels[::-1]), at byte
 * offset pix_offsets[i] of strips_out.  capacity = bytes of strips_out (and of gray_out); strips that
 * do not fit are skipped whole (compare pix_offsets[n] with the capacity).
 *   binarize   OG_BIN_NONE: strips_out receives the gray strips.
 *              OG_BIN_OTSU: strips_out receives (gray > t) * 255 with t the strip's own Otsu threshold
 *              (thresholds_out, n int32, may be NULL).  gray_out, if not NULL, additionally keeps the
 *              gray strips.
 *              OG_BIN_SAUVOLA: as og_binarize_strips (window odd, 3..127; k = 0.2, r = 127.5); the
 *              window needs the whole gray strip, so gray_out is required and must not alias strips_out.
 *   bbox_out   if not NULL, receives per line (floor min x, floor min y, ceil max x, ceil max y) of its
 *              fp32 warped grid as int32 (origami/core/mask.py:10-12 convention).
 * With `map` (the dense map of the SAME grid) the samples' nodes are gathered from it and the
 * binarizer runs as a second pass over the L2-resident strips; with map == NULL the dense-map nodes
 * under every line are recomputed on chip from the sparse grid by CTAs that own a line each, and Otsu
 * is applied by the CTA that sampled the line.  Identical bits either way.  Like og_sample_labels the
 * call uses the plan's work-ticket word while it runs (one call per plan at a time). */
int og_extract_lines(const uint8_t *src, int H, int W, int64_t src_pitch,
	const float *grid, int gh, int gw, int cell, const float *map,
	const og_line *lines, int n,
	const int32_t *steps, const int64_t *pix_offsets, int32_t *chunk_offsets,
	uint8_t *strips_out, int64_t capacity, int32_t *bbox_out,
	int binarize, int window, uint8_t *gray_out, int32_t *thresholds_out, void *stream);

/* A12  binarizers on packed strips -- origami/core/binarize.py:11-43 as applied per line by
 * Line.image (block.py:140-141).  gray and out hold n strips packed at pix_offsets; they
 * must not alias for OG_BIN_SAUVOLA (window odd, 3..127; k = 0.2, r = 127.5 as in the
 * reference).  capacity = bytes of gray and of out; strips that do not fit are skipped whole, as
 * in og_extract_lines.  thresholds_out (n int32, OTSU only, may be NULL) receives each threshold.
 * chunk_offsets is the n + 2 entry array og_line_steps wrote: like og_extract_lines, the Sauvola pass
 * hands its work out through the plan's work-ticket word (entry n + 1) while it runs and leaves it at
 * zero, so a plan serves one such call at a time. */
int og_binarize_strips(const uint8_t *gray, const int32_t *steps, const int64_t *pix_offsets,
	const int32_t *chunk_offsets, int n, int mode, int window, uint8_t *out, int64_t capacity,
	int32_t *thresholds_out, void *stream);

/* A12  page-scale Sauvola -- skimage.filters.threshold_sauvola (0.16.2) + `> t` as used by
 * Page.binarized (origami/core/page.py:101-104) and Regions.binarized
 * (origami/batch/detect/layout.py:157-169).  out = 255 where src > t else 0. */
int og_sauvola(const uint8_t *src, int h, int w, int64_t src_pitch, int window,
	double k, double r, uint8_t *dst, int64_t dst_pitch, void *stream);

/* A12  page-scale Otsu -- skimage.filters.threshold_otsu (0.16.2) via origami/core/
 * binarize.py:18-26.  hist_scratch: 256 uint32 (zeroed by the call). */
int og_otsu(const uint8_t *src, int h, int w, int64_t src_pitch,
	uint8_t *dst, int64_t dst_pitch, uint32_t *hist_scratch, int32_t *threshold_out,
	void *stream);

/* og_otsu without caller scratch.  This entry point has no counterpart in the reference: it
 * exists for callers that cannot allocate device memory -- an Origami stage plugin living in
 * the same process as the TensorFlow segmentation nets (TensorFlow reserves the device for its
 * own arena, so cudaMalloc fails afterwards), a worker that was handed a fixed set of buffers,
 * or code that must not trigger the device-wide synchronisation of cudaMalloc / cudaFree.
 * The 256-counter histogram lives in a static device array inside the library (1 KiB per GPU
 * that loads it, held until exit); the call allocates nothing.  Limits and rules:
 *  - h * w <= 2^30 pixels (OG_E_RANGE beyond that; use og_otsu with your own scratch);
 *  - all calls on a device share the array: it is cleared at the start of each call, and calls
 *    whose GPU work overlaps (other streams / unordered host threads) race -- serialise them;
 *  - the first call on each device (with that device current) looks the array's address up
 *    once; make that first call before any stream capture.
 * Output is identical to og_otsu. */
int og_otsu_noalloc(const uint8_t *src, int h, int w, int64_t src_pitch,
	uint8_t *dst, int64_t dst_pitch, int32_t *threshold_out, void *stream);

/* (f)4  strip hand-off to the recogniser -- origami/batch/detect/ocr.py:238-268: the OCR stage drops
 * strips smaller than min_w x min_h (6 x 6) and feeds the rest to its model in batches padded to the
 * widest line.  strips / steps / pix_offsets as written by og_line_steps + og_extract_lines (capacity
 * = bytes of strips; a strip the extraction skipped becomes an all-padding slot).  Kept strips
 * (xsteps >= min_w, ysteps >= min_h, and fitting the batch: ysteps <= H, xsteps <= w_max), in line
 * order, are copied into batch_out (n, H, w_max) u8 -- slot j at byte j * H * w_max, right and bottom
 * padded with pad_value; widths_out[j] = its xsteps, kept_index_out[j] = its line index (both n
 * int32), *n_kept_out = the number of slots used.  slot_scratch: n int32 the call may overwrite. */
int og_pack_strips(const uint8_t *strips, const int32_t *steps, const int64_t *pix_offsets, int n, int64_t capacity,
	int min_w, int min_h, int pad_value, int H, int w_max,
	uint8_t *batch_out, int32_t *slot_scratch, int32_t *widths_out, int32_t *kept_index_out, int32_t *n_kept_out, void *stream);

/* (f)4  binarized strips as one bit per pixel, for callers that take the strips to the host: eight
 * times fewer bytes over PCIe than the 0 / 255 bytes Line.image returns (origami/core/block.py:140-141,
 * consumed by origami/batch/detect/ocr.py:238-249 and origami/batch/export/lines.py:57-67, which writes
 * them as 1-bit images).  No counterpart in the reference: a transport format.  dst[i] bit (7 - j) =
 * (src[8 i + j] != 0), numpy.packbits order; n pixels in, (n + 7) / 8 bytes out, the last byte padded
 * with zero bits.  Strip i of a packed strip buffer starts at bit pix_offsets[i]. */
int og_pack_bits(const uint8_t *src, int64_t n, uint8_t *dst, void *stream);

/* A triangulation with vertex values on the device (all pointers are DEVICE pointers): what
 * og_tri_interp takes as separate arguments, as one struct for the entry points that need two of them. */
typedef struct og_tri_tables {
	const int32_t *cell_start;  /* (gy * gx + 1) offsets into cell_tris                              */
	const int32_t *cell_tris;   /* triangle indices per bucket                                       */
	const double *transform;    /* (m, 3, 2) scipy.spatial.Delaunay.transform                        */
	const int32_t *simplices;   /* (m, 3) vertex indices                                             */
	const double *values;       /* (n_vertices, 2)                                                   */
	double ox, oy, cell, eps;   /* bucket grid origin, bucket size; barycentric tolerance            */
	int32_t gx, gy;
} og_tri_tables;

/* A scalar angle field as origami/core/dewarp.py:108-122 builds it with lingrid.lininterp: values
 * component 0 = phi; `inter` over the samples (has_inter = 0 when they are collinear), `extra` over the
 * hull / box / corner points, queries clipped to bounds = (minx, miny, maxx, maxy); with fewer than three
 * samples the field is the constant phi0 (constant = 1, the tables are ignored).  HOST struct. */
typedef struct og_field {
	og_tri_tables inter, extra;
	double bounds[4];
	double phi0;
	int32_t has_inter, constant;
} og_field;

/* (f)1  GridFactory's Euler integration -- origami/core/dewarp.py:123-141 (Field.estimate_extent),
 * :341-419 (extend_border_h / _v), :421-439 (grid_h): n_points float32 points advance n_steps times by
 * `step` pixels along (cos phi, sin phi) of the field (evaluated in float32 as Field.get does).
 * start_xy (n_points, 2) f32 -> out_xy (n_steps, n_points, 2) f32, out[k] = positions before step k. */
int og_field_march(const og_field *field, const float *start_xy, int n_points, int n_steps, float step,
	float *out_xy, void *stream);

/* (f)1  GridFactory.grid_hv's vertical lines -- origami/core/dewarp.py:441-470 with BatchIntersections
 * (:176-263): for columns col0 .. col0 + ncols - 1 of the lattice, starting on h-line 0, a ray of
 * max_step_len along the vertical field is intersected with the polyline of the next h-line (rows
 * further down are tried when it is missed) and the nearest crossing becomes the next node.  grid_h
 * (gh, gw, 2) f32 with the end points of every h-line already pushed outwards as the reference's
 * make_row_slice does; grid_hv_out (gh, gw, 2) f32 receives the columns; *status_out is set to
 * 1 + column where no crossing was found (the reference raises there), left untouched otherwise. */
int og_grid_vlines(const og_field *field_v, const float *grid_h, int gh, int gw, int col0, int ncols,
	float max_step_len, float *grid_hv_out, int32_t *status_out, void *stream);

/* (f)2  Grid.transformer / lininterp evaluation -- origami/core/dewarp.py:143-161, 555-559 and
 * origami/core/lingrid.py:185-206, 266-275 (scipy.interpolate.LinearNDInterpolator over the warped
 * lattice nodes), applied to all contour vertices of a page by origami/batch/detect/dewarp.py:16-39.
 * pts (n, 2) f64 -> out (n, 2) f64: the query is clipped to clip_box (HOST array minx, miny, maxx,
 * maxy), located in the triangulation and blended from the three vertex values; NaN where no
 * triangle holds it.  The triangulation comes from the host (scipy.spatial.Delaunay, the oracle's
 * own): simplices (m, 3) int32, transform (m, 3, 2) f64 = Delaunay.transform, values (n_vertices, 2)
 * f64, plus a bucket grid over the triangles' bounding boxes: bucket (cy, cx) of size `cell` at
 * (ox, oy) lists cell_tris[cell_start[cy*gx+cx] .. cell_start[cy*gx+cx+1]).  eps: barycentric
 * tolerance of the inside test.  only_nan != 0: only entries of out that hold NaN are computed (the
 * hull-edge extrapolation pass of lingrid's Interpolator). */
int og_tri_interp(const double *pts, int64_t n, const double *clip_box,
	const int32_t *cell_start, const int32_t *cell_tris, int gx, int gy, double ox, double oy, double cell,
	const double *transform, const int32_t *simplices, const double *values, double eps,
	double *out, int only_nan, void *stream);

/* A15  ConfidenceSampler.__call__ -- origami/batch/detect/lines.py:34-54.  For every line:
 * warped grid at (xres, yres) from the descriptors, scaled to the label map by
 * fp32(lw/W), fp32(lh/H), NEAREST lookup (OOB -> label 0), histogram.  counts_out
 * (n, n_classes) int64, zeroed by the call; labels >= n_classes are counted in
 * overflow_out[i] (n int64, may be NULL; np.bincount would have grown the array).  Like
 * og_extract_lines it uses the plan's work-ticket word while it runs. */
int og_sample_labels(const uint8_t *labels, int lh, int lw, int64_t labels_pitch,
	int H, int W,
	const float *grid, int gh, int gw, int cell, const float *map,
	const og_line *lines, int n,
	const int32_t *steps, int32_t *chunk_offsets,
	int n_classes, int64_t *counts_out, int64_t *overflow_out, void *stream);

/* A18+A19  label argmax and class masks -- origami/core/predict.py:281-285,
 * origami/core/contours.py:488.  probs (M, h, w, C) fp32 -> labels (h, w) u8 =
 * argmax_c of the in-order fp32 sum over models (first maximum wins); masks_out, if not
 * NULL, (n_mask, h, w) u8 with masks_out[k] = (labels == mask_values[k]) (mask_values is a
 * HOST array of n_mask bytes). */
int og_argmax_labels(const float *probs, int M, int C, int h, int w,
	uint8_t *labels_out, uint8_t *masks_out, const uint8_t *mask_values, int n_mask,
	void *stream);

/* A18  per-tile label assembly -- origami/core/predict.py:43-63 (Tile.write_inner) and :275-285: the
 * network runs on overlapping tiles; each tile's labels = argmax_c sum_m probs[m, y, x, c], and only
 * the tile's inner rectangle is copied into the page label map.  probs (M, th, tw, C) fp32 of ONE
 * tile; the rectangle (inner_dx, inner_dy, inner_w, inner_h) in tile coordinates is written to
 * labels (lh, lw) u8 at (x0, y0) -- one launch, nothing else of the map is touched. */
int og_argmax_tile(const float *probs, int M, int C, int th, int tw,
	int inner_dx, int inner_dy, int inner_w, int inner_h,
	uint8_t *labels, int lh, int lw, int64_t labels_pitch, int x0, int y0, void *stream);

/* A19 alone: masks from an existing label map. */
int og_class_masks(const uint8_t *labels, int h, int w, int64_t pitch,
	uint8_t *masks_out, const uint8_t *mask_values, int n_mask, void *stream);

/* A20  Regions.binarized helpers -- origami/batch/detect/layout.py:157-186.
 * og_label_mask_ne: out = 255 where labels != background else 0 (layout.py:175-176, the separator mask
 * that is then dewarped with og_remap_page(OG_INTER_NEAREST)).
 * og_or_dilate3: out (h, w) fp32 = 1.0 where bin != 0 or the 3x3 neighbourhood of mask holds a non-zero
 * (skimage binary_dilation with square(3), zero outside the image), else 0.0 (layout.py:179-186);
 * mask may be NULL. */
int og_label_mask_ne(const uint8_t *labels, int h, int w, int64_t pitch, int background,
	uint8_t *out, int64_t out_pitch, void *stream);
int og_or_dilate3(const uint8_t *bin, int64_t bin_pitch, const uint8_t *mask, int64_t mask_pitch,
	int h, int w, float *out, void *stream);

/* (f)3  SplitDetector's column statistic -- origami/batch/detect/layout.py:915-931:
 * scipy.signal.periodogram(pixels, axis=0) (mean-free columns, boxcar window, one-sided density)
 * followed by np.quantile(dens, quantile, axis=0).  pixels (h, w): uint8 (scaled by 1/255 in float32
 * first, as the reference does) or float32; out (w) float32.  The spectrum is evaluated in float64
 * (the reference's is complex64 pocketfft): agreement to float32 round-off, h <= 8192. */
int og_column_psd_quantile(const void *pixels, int is_u8, int h, int w, int64_t pitch_bytes,
	double quantile, float *out, void *stream);

/* (f)3  the same statistic for ALL regions of a page in one launch -- FixSpillOverH / FixSpillOverV,
 * origami/batch/detect/layout.py:947-1075, call the detector once per contour on crop(regions.binarized,
 * contour) (:886-894) resp. its transpose.  pixels: the (H, W) page image (uint8 or float32) on the device;
 * regions: DEVICE array of m records (x0, y0, w, h, transposed) int32, each rectangle inside the image;
 * a record yields w series of h samples (its columns) or, transposed, h series of w samples (its rows);
 * offsets: DEVICE array of m + 1 int32 prefix sums of the series counts, total = offsets[m], max_len = the
 * longest series (<= 8192).  out: total float32, region r's values at offsets[r]. */
int og_region_psd_quantile(const void *pixels, int is_u8, int H, int W, int64_t pitch_bytes,
	const int32_t *regions, const int32_t *offsets, int m, int total, int max_len, double quantile, float *out, void *stream);

/* A11  polygon mask / cut-out -- origami/core/mask.py:8-74 (Mask.__init__: cairo fill of the polygon,
 * `R > 0`; Mask._extract: pixels outside the mask become `background`), used by
 * Line._extract_deskewed (origami/core/block.py:174-186), Line.masked_image (:143-153) and
 * Block.image (:396-401).  verts_xy: DEVICE array of k (x, y) f64 vertices of a closed ring, relative
 * to the mask's origin (the caller rounds them to integers as Mask.__init__ does).  A pixel belongs to
 * the mask when the polygon covers at least 1/510 of it (cairo's 8-bit alpha rounds to non-zero); the
 * coverage is exact polygon-area arithmetic, cairo's own rasteriser is not reproduced (unpinned).
 * mask_out (h, w) u8 255/0 and/or img_inout (h, w) u8 whose pixels outside the polygon are set to
 * background; either may be NULL. */
int og_polygon_mask(const double *verts_xy, int k, int w, int h,
	uint8_t *mask_out, int64_t mask_pitch, uint8_t *img_inout, int64_t img_pitch, int background, void *stream);

/* A11  warpAffine stage of Line._extract_deskewed -- origami/core/block.py:155-172
 * (--do-not-dewarp path; the cairo polygon cut-out that follows is not reproduced).
 * M is a HOST 2x3 float64 forward matrix as returned by cv2.getAffineTransform. */
int og_warp_affine(const uint8_t *src, int H, int W, int64_t src_pitch,
	const double *M, uint8_t *dst, int width, int height, int64_t dst_pitch,
	int border_value, void *stream);

#ifdef __cplusplus
}
#endif
#endif
