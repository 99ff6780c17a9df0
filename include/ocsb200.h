/*
 * ocsb200.h -- C ABI of libocsb200.so: the B200-native raster hot path of OCSMesh.
 *
 * The reference (noaa-ocs-modeling/OCSMesh) is 100 % Python and has NO
 * plugin / FFI layer on this path: the arithmetic is inline numpy/scipy inside
 * Python methods.  Each entry point below therefore replaces the body of one
 * reference loop ("per window: ...") and cites the reference file:line it
 * stands in for.  INTEGRATION.md shows the ctypes stub a maintainer would add
 * at those call sites.
 *
 * Conventions
 *   - every function returns an int32 status: 0 = OK, <0 = error
 *     (OCSB_EINVAL, OCSB_ENOMEM, OCSB_ECUDA, OCSB_EOVERFLOW); the text of the
 *     last error on the calling thread is returned by ocsb_last_error().
 *   - no exception crosses the boundary, no ownership is transferred: the
 *     caller owns every buffer.  Pointers may be HOST or DEVICE memory (the
 *     library asks cudaPointerGetAttributes); host buffers are staged through
 *     device scratch inside the call (that is the "e2e" path of bench.py).
 *   - a call returns after its work has been *enqueued* on the context's
 *     stream when all buffers are device buffers, and after completion when
 *     any buffer was a host buffer.  ocsb_ctx_sync() waits for the stream.
 *   - "None" for an optional double argument is passed as NaN (OCSB_NONE).
 *   - one ocsb_ctx per (process, device); calls on one ctx are not re-entrant.
 *   - pixel buffers are row-major windows: element (row, col) of a window lives
 *     at base[row * row_stride + col]; `base` points at the window's top-left
 *     pixel, so a window inside a larger raster is base = raster + row_off *
 *     W + col_off, row_stride = W.
 */
#ifndef OCSB200_H
#define OCSB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OCSB_OK          0
#define OCSB_EINVAL     (-1)
#define OCSB_ENOMEM     (-2)
#define OCSB_ECUDA      (-3)
#define OCSB_ENCCL      (-4)
#define OCSB_EOVERFLOW  (-5)

#define OCSB_ABI_VERSION 5

/* DEM element types (the hfun raster itself is always float32,
 * ocsmesh/hfun/raster.py:76) */
#define OCSB_F32 0
#define OCSB_F64 1

/* Mesh.interpolate methods (ocsmesh/mesh/mesh.py:2306-2336) */
#define OCSB_INTERP_LINEAR  0
#define OCSB_INTERP_NEAREST 1
#define OCSB_INTERP_SPLINE  2

typedef struct ocsb_ctx ocsb_ctx;

/*
 * One semantic raster window, as the reference sees it:
 *   (x0, y0, x1, y1) = Raster.get_window_bounds(window)          ocsmesh/raster.py:1642-1662
 *                      = (west, south, east, north) pixel-EDGE bounds
 *   pixel column k has x = np.linspace(x0, x1, width)[k]          ocsmesh/raster.py:452-453
 *   pixel row    j has y = np.linspace(y1, y0, height)[j]         ocsmesh/raster.py:479-480
 *
 * row_begin / row_end select the DEVICE TILE of the semantic window that this
 * call processes: only pixel rows [row_begin, row_end) are read-modified-
 * written (row_end == 0 means `height`).  Coordinates, the radius law and the
 * one-sided gradient differences still refer to the WHOLE window, so a window
 * split into row bands over several GPUs (or several calls) gives bit-identical
 * results to one call (semantic windows != device tiles).
 * Buffers are always addressed relative to the window's top-left pixel; a rank
 * that holds only its band (+1 halo row of the DEM for the gradient stencil)
 * passes `band_buffer - first_held_row * row_stride`.
 */
typedef struct ocsb_win {
    double  x0, y0, x1, y1;
    int32_t width, height;
    int32_t row_begin, row_end;
} ocsb_win;

/* ---- library / context -------------------------------------------------- */

int32_t     ocsb_abi_version(void);
const char *ocsb_last_error(void);

/* device: CUDA ordinal.  stream: a cudaStream_t; NULL = a non-blocking stream
 * the context creates and owns.  ocsb_ctx_set_stream() takes the handle as
 * is, so NULL there selects the legacy default stream (what torch uses unless
 * told otherwise).  */
int32_t ocsb_ctx_create(int32_t device, void *stream, ocsb_ctx **out);
int32_t ocsb_ctx_destroy(ocsb_ctx *ctx);
int32_t ocsb_ctx_set_stream(ocsb_ctx *ctx, void *stream);
int32_t ocsb_ctx_sync(ocsb_ctx *ctx);
/* number of kernels this context has launched so far (bench "gpu_launches") */
int64_t ocsb_ctx_launch_count(const ocsb_ctx *ctx);
/* release cached device scratch */
int32_t ocsb_ctx_trim(ocsb_ctx *ctx);

/* Per-kernel-class device timing (CUDA events recorded on the context's
 * stream around every launch of the class while enabled).  get_timing
 * synchronises the stream, returns the summed duration and the launch count
 * of `kclass` since the last reset; reset_timing drops all records.          */
#define OCSB_KC_FILL              0
#define OCSB_KC_CLAMP             1
#define OCSB_KC_MIN_COMBINE       2
#define OCSB_KC_COORDS            3
#define OCSB_KC_LIMITER           4
#define OCSB_KC_FEATURE_REFINE    5
#define OCSB_KC_FEATURE_EVAL      6
#define OCSB_KC_INTERP            7
#define OCSB_KC_SPLINE_PREFILTER  8
#define OCSB_KC_SPLINE_EVAL       9
#define OCSB_KC_NEAREST          10
#define OCSB_KC_BUCKET           11
#define OCSB_KC_CONSTRAINT       12
#define OCSB_KC_CONTOUR          13
#define OCSB_KC_SAMPLE           14
#define OCSB_KC_PASS_TILE        15
#define OCSB_KC_PASS_LIMITER     16   /* split pass: limiter sweep of all tiles     */
#define OCSB_KC_PASS_FEATURE     17   /* split pass: feature op on candidate tiles  */
#define OCSB_KC_COUNT            18
int32_t     ocsb_ctx_set_timing(ocsb_ctx *ctx, int32_t enabled);
int32_t     ocsb_ctx_get_timing(ocsb_ctx *ctx, int32_t kclass, double *total_ms,
                                int64_t *launches);
int32_t     ocsb_ctx_reset_timing(ocsb_ctx *ctx);
const char *ocsb_kernel_class_name(int32_t kclass);

/* ---- A1: window list ------------------------------------------------------
 * get_iter_windows(width, height, chunk_size, overlap)   ocsmesh/raster.py:1891-1939
 * Raster.iter_windows (chunk_size in [0, None, False] -> one full window)
 *                                                        ocsmesh/raster.py:1577-1608
 * out4: cap x 4 int32 (col_off, row_off, width, height), row-major order.
 * *n receives the number of windows (also when out4 == NULL or cap is too
 * small, in which case OCSB_EOVERFLOW is returned).  Host-only, no ctx.       */
int32_t ocsb_windows(int32_t width, int32_t height, int32_t chunk_size,
                     int32_t overlap, int32_t *out4, int64_t cap, int64_t *n);

/* ---- A2: window coordinates -----------------------------------------------
 * Raster.get_x / get_y                                   ocsmesh/raster.py:428-480
 * Writes np.linspace(x0, x1, width) to x_out[width] and np.linspace(y1, y0,
 * height) to y_out[height], computed ON THE DEVICE with the same
 * fl(fl(k*step)+start) / pinned-endpoint rule the kernels use in-register
 * (the reference's (h*w,2) get_xy array is never materialised).              */
int32_t ocsb_window_coords(ocsb_ctx *ctx, const ocsb_win *win,
                           double *x_out, double *y_out);

/* ---- A4: hfun initial fill -------------------------------------------------
 * HfunInputRaster.__set__: every pixel = finfo(float32).max   ocsmesh/hfun/raster.py:57-86 */
int32_t ocsb_fill_f32(ocsb_ctx *ctx, float *dst, int64_t n, float value);

/* ---- A10: global clamp re-pass ---------------------------------------------
 * HfunRaster.apply_constraints with no custom constraints     ocsmesh/hfun/raster.py:509-568
 * (values < hmin -> hmin, values > hmax -> hmax, float32 compares)           */
int32_t ocsb_hfun_clamp(ocsb_ctx *ctx, float *hfun, int64_t row_stride,
                        int32_t width, int32_t height,
                        double hmin, double hmax);

/* ---- A6 + A7: feature / contour distance sizing ----------------------------
 * Body of the window loop of HfunRaster.add_feature            ocsmesh/hfun/raster.py:1230-1257
 *   tree = cKDTree(points); d = tree.query(xy[, distance_upper_bound=r])
 *   v = float32(rate*target*d + target); clamp [hmin,hmax]; new = min(old, v)
 * verts_xy : nv x 2 float64, the resampled feature vertices ("points", :1230)
 * hfun_old : window of the PRE-op hfun (the reference reads self.src), or
 *            NULL = the raster is still in its A4 initial state (every pixel
 *            finfo(float32).max, ocsmesh/hfun/raster.py:75-82): the kernel
 *            then neither needs the fill pass nor reads 4 B/px
 * hfun_new : window of the POST-op hfun (may alias hfun_old)
 * hmin/hmax: the HfunRaster's global bounds, NaN = None.  hmax "truthy"
 *            (not None, != 0) selects the bounded query, exactly as :1240.
 * fuse_clamp != 0 additionally applies the A10 clamp to the result (what
 *            apply_constraints_wrap does right after the op when no custom
 *            constraint is registered, ocsmesh/features/constraint.py:420-423).
 * dist_out : optional window (row_stride = width) of float64 distances as
 *            used by the size law (r where scipy returned inf), or NULL
 * idx_out  : optional window (row_stride = width) of int64 nearest-vertex
 *            indices (nv where scipy returned nv), or NULL                    */
int32_t ocsb_hfun_feature(ocsb_ctx *ctx, const ocsb_win *win,
                          const double *verts_xy, int64_t nv,
                          const float *hfun_old, int64_t old_stride,
                          float *hfun_new, int64_t new_stride,
                          double expansion_rate, double target_size,
                          double hmin, double hmax, int32_t fuse_clamp,
                          double *dist_out, int64_t *idx_out);

/* Nearest-vertex query alone (second cKDTree site, Constraint._apply_rate
 * ocsmesh/features/constraint.py:104-105): for `nq` arbitrary query points.
 * r = NaN -> unbounded.  dist_out[nq] float64 (inf on miss), idx_out[nq]
 * int64 (nv on miss).                                                         */
int32_t ocsb_nearest_vertex(ocsb_ctx *ctx, const double *verts_xy, int64_t nv,
                            const double *query_xy, int64_t nq, double r,
                            double *dist_out, int64_t *idx_out);

/* The same query without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  The vertex grid lives in one fixed file-scope device array (~4 MiB, held
 * by every GPU that loads the library until the process exits), so:
 *   - nv <= OCSB_STATIC_MAX_VERTS, more -> OCSB_EINVAL with a message;
 *   - every buffer must be DEVICE memory (nothing is staged), verts_xy and
 *     query_xy 16-byte aligned;
 *   - no cudaMalloc / cudaFree / pool or library allocation on any path;
 *   - all calls share the array (it keeps the last call's contents; what must
 *     start at zero is cleared by each call): calls that overlap on the GPU
 *     (two streams or two contexts of one device) race -- serialise them.
 * Results are bit-identical to ocsb_nearest_vertex; the call is stream-ordered
 * after one internal wait for the vertices' bounding box.                      */
#define OCSB_STATIC_MAX_VERTS 262144
int32_t ocsb_nearest_vertex_static(ocsb_ctx *ctx, const double *verts_xy, int64_t nv,
                                   const double *query_xy, int64_t nq, double r,
                                   double *dist_out, int64_t *idx_out);

/* ---- A8 (+A9): subtidal flow limiter / gradient stencil --------------------
 * Body of the window loop of HfunRaster.add_subtidal_flow_limiter
 *                                                         ocsmesh/hfun/raster.py:1396-1435
 * dem      : window of the DEM band (dem_dtype OCSB_F32 / OCSB_F64)
 * hfun_old : as for ocsb_hfun_feature (NULL = A4 initial state)
 * dx, dy   : Raster.dx, Raster.dy (np.gradient(z, dx, dy): dx spaces axis 0!)
 * hmin,hmax,lower_bound,upper_bound : the method's arguments, NaN = None
 * ghmin,ghmax : the HfunRaster's global bounds (self._hmin/_hmax), NaN = None
 * Zero-gradient pixels follow the masked-array arithmetic of the reference
 * (raw value 1/3).  The window needs width,height >= 2
 * (np.gradient raises otherwise -> OCSB_EINVAL).                              */
int32_t ocsb_hfun_flow_limiter(ocsb_ctx *ctx, const ocsb_win *win,
                               const void *dem, int32_t dem_dtype,
                               int64_t dem_stride,
                               const float *hfun_old, int64_t old_stride,
                               float *hfun_new, int64_t new_stride,
                               double dx, double dy,
                               double hmin, double hmax,
                               double lower_bound, double upper_bound,
                               double ghmin, double ghmax,
                               int32_t fuse_clamp);

/* Host half of the contour path: join the segments of ocsb_contour_segments
 * that share an edge id into polylines (what matplotlib's path walk +
 * shapely.ops.linemerge do in the reference, ocsmesh/raster.py:1426-1470,
 * 1340-1372).  Pure host code, no ctx.  Open lines come first (walked from
 * their lower edge id end), then closed rings (first vertex repeated at the
 * end); lines are ordered by their start node.
 *   xy_out      : cap_xy x 2 float64, the polylines' vertices back to back
 *   offsets_out : cap_lines + 1 int64, line k = xy_out[offsets[k] : offsets[k+1]]
 * cap_xy = 2 n and cap_lines = n always suffice.                               */
int32_t ocsb_stitch_segments(const int64_t *edges, const double *pts, int64_t n,
                             double *xy_out, int64_t cap_xy,
                             int64_t *offsets_out, int64_t cap_lines,
                             int64_t *n_xy, int64_t *n_lines);

/* Feature prologue of add_feature (host code): repartition every polyline into
 * chunks of at most max_verts vertices that share their end vertex
 * (utils.repartition_features, ocsmesh/utils.py:247-277) and place a point every
 * target_size metres along each chunk, plus both chunk ends
 * (utils.transform_linestring, ocsmesh/utils.py:280-312): the `points` array of
 * ocsmesh/hfun/raster.py:1219-1230.
 *   xy, offsets : n_lines polylines, line k = xy[offsets[k] : offsets[k+1]] (x, y)
 *   out         : cap x 2 float64 or NULL (count only); *n_out = number of points;
 *                 cap < *n_out: OCSB_EOVERFLOW, nothing usable written          */
int32_t ocsb_feature_points(const double *xy, const int64_t *offsets,
                            int64_t n_lines, double target_size,
                            int32_t max_verts, double *out, int64_t cap,
                            int64_t *n_out);

/* ---- 8(f) next-2: add_constant_value ----------------------------------------
 * Body of the window loop of HfunRaster.add_constant_value  ocsmesh/hfun/raster.py:1489-1498
 *   v = old; v[(dem > lower_bound) & (dem < upper_bound)] = value; new = min(old, v)
 * The compares run in the DEM dtype, `value` is rounded to float32 (numpy's
 * casting of Python scalars).  lower/upper_bound: +-inf = None.  hfun_old,
 * ghmin/ghmax/fuse_clamp as for ocsb_hfun_feature.                             */
int32_t ocsb_hfun_constant_value(ocsb_ctx *ctx, const ocsb_win *win,
                                 const void *dem, int32_t dem_dtype,
                                 int64_t dem_stride,
                                 const float *hfun_old, int64_t old_stride,
                                 float *hfun_new, int64_t new_stride,
                                 double value, double lower_bound,
                                 double upper_bound, double ghmin, double ghmax,
                                 int32_t fuse_clamp);

/* ---- 8(f) next-2: topography constraints ------------------------------------
 * TopoConstConstraint.apply / TopoFuncConstraint.apply     ocsmesh/features/constraint.py:164-177, 207-224
 * followed by Constraint._apply_rate                       ocsmesh/features/constraint.py:83-121
 *   mask = (dem > lower_bound) & (dem < upper_bound) & ~satisfies(hfun, val)
 *   hfun[mask] = val
 *   rate != NaN: every non-mask pixel p with nearest mask pixel m (cKDTree of
 *   the window's mask pixels, Euclidean, window coordinates):
 *       t = hfun[m] * (1 + |p - m| * rate * sign);  hfun[p] = t unless
 *       satisfies(hfun[p], t)          (sign = -1 for MIN, +1 for MAX)
 * satisfies = np.greater for OCSB_VT_MIN, np.less for OCSB_VT_MAX.
 * values   : NULL -> the constant `value` (TopoConst; compared in float32),
 *            else a window of per-pixel values (TopoFunc: func(dem) evaluated
 *            by the caller; values_dtype OCSB_F32 / OCSB_F64)
 * hfun     : read-modify-written in place (the global [hmin, hmax] clamp of
 *            apply_constraints is a separate ocsb_hfun_clamp call)
 * A rate needs the whole window on one device (row_begin/row_end unset); for a
 * row band use ocsb_hfun_topo_mask_band + ocsb_hfun_apply_rate below.           */
#define OCSB_VT_MIN 0
#define OCSB_VT_MAX 1
int32_t ocsb_hfun_topo_constraint(ocsb_ctx *ctx, const ocsb_win *win,
                                  const void *dem, int32_t dem_dtype,
                                  int64_t dem_stride,
                                  float *hfun, int64_t hfun_stride,
                                  const void *values, int32_t values_dtype,
                                  int64_t values_stride, double value,
                                  double lower_bound, double upper_bound,
                                  int32_t value_type, double rate);

/* The same constraint WITH a rate when the window is split into row bands over
 * several devices: the nearest mask pixel of a band pixel may lie in another
 * band, so the call is split around an exchange of mask-boundary pixels.
 *   1. ocsb_hfun_topo_mask_band on every band: applies the mask step to the band
 *      (hfun[mask] = val), writes the band's mask (device, (row_end-row_begin) x
 *      width, compact) and returns the band's mask-boundary pixels as vertices
 *      (window coordinates, row-major) with their values.  Mask pixels on a band
 *      edge that is not a window edge count as boundary (their outer neighbour
 *      is unknown) -- a superset that cannot change a nearest-pixel answer.
 *      verts_out / vals_out NULL: count only (*n_out); cap < *n_out:
 *      OCSB_EOVERFLOW.  The call is idempotent.
 *   2. the caller concatenates all bands' vertices and values in band order
 *      (an all-gather over ranks, ocsmesh_b200.distributed.all_gather_ragged)
 *   3. ocsb_hfun_apply_rate on every band with the full vertex set: the
 *      _apply_rate step (ocsmesh/features/constraint.py:99-117) for the band's
 *      non-mask pixels.  nv == 0 (no mask pixel anywhere): no-op
 *      (`if not np.any(mask): return values`, :87-88).
 * hfun: (virtual) address of window row 0, device memory in step 3.             */
int32_t ocsb_hfun_topo_mask_band(ocsb_ctx *ctx, const ocsb_win *win,
                                 const void *dem, int32_t dem_dtype,
                                 int64_t dem_stride,
                                 float *hfun, int64_t hfun_stride,
                                 const void *values, int32_t values_dtype,
                                 int64_t values_stride, double value,
                                 double lower_bound, double upper_bound,
                                 int32_t value_type, uint8_t *mask_out,
                                 double *verts_out, float *vals_out, int64_t cap,
                                 int64_t *n_out);
int32_t ocsb_hfun_apply_rate(ocsb_ctx *ctx, const ocsb_win *win,
                             float *hfun, int64_t hfun_stride,
                             const uint8_t *mask, const double *verts_xy,
                             const float *vert_values, int64_t nv, double rate,
                             int32_t value_type);

/* CourantNumConstraint.apply                               ocsmesh/features/constraint.py:254-313
 * (utils.py:1677-1835: linear-wave particle velocity, characteristic
 * velocity, element size for a Courant number).  Arithmetic in the DEM dtype
 * like numpy; pixels whose approximate Courant number does not satisfy
 * `courant` (np.greater for MIN, np.less for MAX) get the size that does.     */
int32_t ocsb_hfun_courant_constraint(ocsb_ctx *ctx, const ocsb_win *win,
                                     const void *dem, int32_t dem_dtype,
                                     int64_t dem_stride,
                                     float *hfun, int64_t hfun_stride,
                                     double courant, double timestep,
                                     double wave_amplitude, int32_t value_type);

/* ---- 8(f) next-1: contour extraction, the per-pixel half --------------------
 * Raster.get_contour -> _get_raster_contour_windowed        ocsmesh/raster.py:1340-1372, 1426-1470
 * (matplotlib/contourpy tracing restated as marching squares: linear
 * interpolation along cell edges, cell mean decides saddles).  One sweep of
 * the DEM window classifies every 2x2 cell and compacts its 0/1/2 segments in
 * cell (row-major) order:
 *   edges : n x 2 int64, the two GLOBAL EDGE IDS of each segment
 *           (horizontal (j,i)-(j,i+1): j*(width-1)+i; vertical (j,i)-(j+1,i):
 *           -(j*width+i)-1) -- segments sharing an id are joined by the caller
 *   pts   : n x 4 float64 (xa, ya, xb, yb), the crossing points in window
 *           coordinates (the window's linspace axes, A2)
 * *n receives the segment count.  edges == NULL (or pts == NULL): count only.
 * cap < n: OCSB_EOVERFLOW, nothing written.  win->row_begin/row_end select
 * the cell rows [row_begin, min(row_end, height-1)); the DEM pointer is the
 * (virtual) address of window row 0 and row min(row_end, height-1) is read.   */
int32_t ocsb_contour_segments(ocsb_ctx *ctx, const ocsb_win *win,
                              const void *dem, int32_t dem_dtype,
                              int64_t dem_stride, double level, int64_t cap,
                              int64_t *edges, double *pts, int64_t *n);

/* ---- multi-GPU: blocks of the per-step exchange of a row-band run -----------
 * (contour-vertex all-gather + 1-row DEM halo.)  Every rank
 * contributes ONE fixed-size block to ONE all-gather per step:
 *   [first owned DEM row | last owned DEM row] (2 x width float32, padded to
 *   16 bytes) followed by cap vertices (its n contour vertices, then repeats of
 *   the last one -- duplicates cannot change a nearest-vertex distance; a rank
 *   without vertices fills with (far_coord, far_coord), |far_coord| < 1e14).
 * ocsb_band_pack builds a rank's block, ocsb_band_unpack turns the gathered
 * world x block array into the contiguous (world * cap, 2) vertex array and the
 * two halo rows (NULL where the rank has no neighbour).  The collective itself
 * is the caller's (torch.distributed / NCCL).  Device memory only.             */
int32_t ocsb_band_pack(ocsb_ctx *ctx, const double *verts_xy, int64_t n, int64_t cap,
                       const float *row_first, const float *row_last, int32_t width,
                       void *send_block, double far_coord);
int32_t ocsb_band_unpack(ocsb_ctx *ctx, const void *recv_blocks, int32_t world,
                         int32_t rank, int64_t cap, int32_t width, double *verts_out,
                         float *halo_above, float *halo_below);

/* ---- 8(f) next-2: polygon regions on the raster grid -----------------------
 * The 0 / 1 raster of a polygon set on a window, as float32:
 *   all_touched = 0 : pixel (r, c) is 1 when its window coordinate
 *       (np.linspace, Raster.get_xy) lies inside the rings -- even-odd crossing
 *       number over ALL edges (holes, several polygons) -- what
 *       RegionConstraint.apply tests with geopandas sjoin(predicate='within')
 *       (ocsmesh/features/constraint.py:357-380).  A RegionConstraint is then a
 *       TopoConstConstraint on this raster (0.5 < raster).
 *   all_touched = 1 : pixel is 1 when its AREA (edge-to-edge rectangle of the
 *       window) is touched: centre inside or a ring edge intersects the closed
 *       rectangle -- the mask of HfunRaster.add_patch
 *       (rasterio.mask.raster_geometry_mask(all_touched=True),
 *       ocsmesh/hfun/raster.py:890-897).
 * shapely / GDAL are not importable here: both rules are restatements, parity
 * with the reference UNPINNED.
 *   edges    : n_edges x 4 float64 (x0, y0, x1, y1), every ring closed
 *   mask_out : (virtual) address of window row 0, rows row_begin..row_end written */
int32_t ocsb_region_mask(ocsb_ctx *ctx, const ocsb_win *win,
                         const double *edges, int64_t n_edges, int32_t all_touched,
                         float *mask_out, int64_t mask_stride);

/* ---- 8(f) next-1 on the device: polylines and the feature prologue ---------
 * Raster.get_contour (ocsmesh/raster.py:1340-1372, 1426-1470: contourpy +
 * shapely linemerge) and the resampling prologue of add_feature
 * (utils.repartition_features / transform_linestring, ocsmesh/utils.py:247-312,
 * called at ocsmesh/hfun/raster.py:1219-1225) without leaving the device: the
 * segments of ocsb_contour_segments are joined by list ranking (pointer
 * jumping over half-edges) into the SAME polylines, vertex for vertex, as
 * ocsb_stitch_segments gives on the host, and resampled chunk by chunk into the
 * SAME vertex array as ocsb_feature_points (running sums in np.cumsum order).
 * The tracing rules themselves (cell-mean saddles, "on the level is not
 * above") restate contourpy's documented behaviour; contourpy / shapely are not
 * importable here, so polyline parity with the reference is UNPINNED.
 *   dem : device memory.  win->row_begin/row_end: the cell rows of a device band.
 * Results stay on the device until taken:
 *   ocsb_contour_lines  -> slot 0: n_xy x 2 float64 vertices, slot 1:
 *                          n_lines + 1 int64 offsets (open lines by start edge
 *                          id first, then rings; a ring repeats its first vertex)
 *   ocsb_contour_points -> slot 0: n_points x 2 float64 = the array handed to
 *                          the nearest-vertex search (hfun/raster.py:1230), all
 *                          levels in order
 *   ocsb_take_result(slot, dst, dst_bytes): copy the kept array to dst (host
 *                          or device memory) and release it.                   */
int32_t ocsb_contour_lines(ocsb_ctx *ctx, const ocsb_win *win,
                           const void *dem, int32_t dem_dtype, int64_t dem_stride,
                           double level, int64_t *n_xy, int64_t *n_lines);
int32_t ocsb_contour_points(ocsb_ctx *ctx, const ocsb_win *win,
                            const void *dem, int32_t dem_dtype, int64_t dem_stride,
                            const double *levels, int32_t n_levels,
                            double target_size, int32_t max_verts,
                            int64_t *n_points);
int32_t ocsb_take_result(ocsb_ctx *ctx, int32_t slot, void *dst, int64_t dst_bytes);

/* ---- 8(f) next-3: sampling side of HfunRaster.msh_t ---------------------------
 * source points of the background size mesh                 ocsmesh/hfun/raster.py:345-377
 *   coords = get_xy(window)[::stride, ::stride]  (row-major, (ny*nx, 2) float64)
 *   values = float64(hfun[::stride, ::stride]), clamped to [hmin, hmax] in f64
 * with nx = ceil(width/stride), ny = ceil(height/stride).  coords or values
 * may be NULL.  hmin/hmax: NaN = None.                                         */
int32_t ocsb_hfun_sample_source(ocsb_ctx *ctx, const ocsb_win *win,
                                const float *hfun, int64_t hfun_stride,
                                int32_t stride, double hmin, double hmax,
                                double *coords_out, double *values_out);

/* utils.interpolate(src, dst, method='nearest')             ocsmesh/utils.py:1122-1140
 * = scipy griddata nearest = cKDTree(src.coords).query(dst.coords): every
 * node gets the (clamped, float64) value of its Euclidean-nearest SOURCE point
 * of ocsb_hfun_sample_source, without materialising the source arrays.
 * xy: K x 2 float64 (16-byte aligned if on the device), values_out: K float64. */
int32_t ocsb_hfun_sample_nearest(ocsb_ctx *ctx, const ocsb_win *win,
                                 const float *hfun, int64_t hfun_stride,
                                 int32_t stride, double hmin, double hmax,
                                 const double *xy, int64_t K,
                                 double *values_out);

/* ---- 8(e): the per-step band exchange as stores into peer memory -------------
 * A row-band run exchanges, every step, each rank's contour vertices (the
 * vertex all-gather of the north star; the reference computes them in one
 * process, ocsmesh/hfun/raster.py:991-1002) and its first / last owned DEM row
 * (the 1-row halos of np.gradient, ocsmesh/hfun/raster.py:1399).
 * ocsb_band_pack + NCCL all-gather + ocsb_band_unpack do that with a
 * collective; these entry points do it with NVLink stores from a kernel:
 *   ocsb_band_p2p_bytes : size of a rank's exchange buffer (two generations of
 *                         `world` slots [2 rows | cap vertices] + 256 B of flags)
 *   ocsb_p2p_alloc      : cudaMalloc + zero + CUDA IPC handle (64 bytes) to hand
 *                         to the other ranks; ocsb_p2p_open maps a peer's buffer,
 *                         ocsb_p2p_close / ocsb_p2p_free undo the two
 *   ocsb_band_push      : pack this rank's block into slot `rank` of EVERY
 *                         rank's buffer (peer_bufs_dev: device array of `world`
 *                         buffer addresses, own buffer at index `rank`), then
 *                         publish `epoch` (1, 2, 3, ... one per step) in every
 *                         rank's flag word; done_counter_dev: one zeroed device
 *                         word kept by the caller across steps
 *   ocsb_band_pull      : wait (bounded, ~2 s -> *status_dev = 1) until all ranks
 *                         published `epoch` in own_buf, then unpack as
 *                         ocsb_band_unpack does
 * Same results as the collective form, bit for bit (bench.py parity check at
 * N > 1).  A rank may run one step ahead of a peer, never two.                  */
int64_t ocsb_band_p2p_bytes(int32_t world, int64_t cap, int32_t width);
int32_t ocsb_p2p_alloc(ocsb_ctx *ctx, int64_t bytes, void **ptr, uint8_t *handle64);
int32_t ocsb_p2p_open(ocsb_ctx *ctx, const uint8_t *handle64, void **ptr);
int32_t ocsb_p2p_close(ocsb_ctx *ctx, void *ptr);
int32_t ocsb_p2p_free(ocsb_ctx *ctx, void *ptr);
int32_t ocsb_band_push(ocsb_ctx *ctx, const double *verts, int64_t n, int64_t cap,
                       const float *row_first, const float *row_last, int32_t width,
                       const void *const *peer_bufs_dev, int32_t world, int32_t rank,
                       int64_t epoch, double far, unsigned long long *done_counter_dev);
int32_t ocsb_band_pull(ocsb_ctx *ctx, const void *own_buf, int32_t world, int32_t rank,
                       int64_t cap, int32_t width, int64_t epoch, double *verts_out,
                       float *halo_above, float *halo_below, int32_t *status_dev);

/* ---- 8(f) next-4: byte decoders of the windowed GeoTIFF read -----------------
 * Raster.get_values -> rasterio windowed read                ocsmesh/raster.py:502-531
 * (GDAL / libtiff, outside the reference tree: restated from the TIFF 6.0
 * specification).  compression: 5 = LZW (MSB-first codes, early change),
 * 32773 = PackBits.  One strip or tile per call; *n_out = decoded bytes
 * (<= cap).  Pure host code, no ctx.                                           */
int32_t ocsb_tiff_decode(int32_t compression, const uint8_t *src, int64_t n,
                         uint8_t *dst, int64_t cap, int64_t *n_out);

/* ---- 8(f) next-4: nearest warp onto the "fast" collector's big raster ---------
 * HfunCollector._create_big_raster -> rasterio.warp.reproject(source,
 * destination, resampling=Resampling.nearest, init_dest_nodata=False)
 *                                           ocsmesh/hfun/collector.py:2419-2438
 * for two north-up rasters of the SAME projected CRS (the geographic / UTM
 * reprojection branch is out of scope with the rest of the pyproj paths).
 * Rule (restated from GDAL's nearest warp kernel; GDAL is absent here, parity
 * UNPINNED): the centre of destination pixel (r, c) is mapped through the two
 * affine transforms to source pixel coordinates (u, v); the pixel takes
 * float32(src[floor(v + 1e-10), floor(u + 1e-10)]) when that lies inside the
 * source and is not masked, and is left untouched otherwise.
 *   src            : source band (OCSB_F32 / OCSB_F64), src_stride elements per row
 *   src_transform  : affine (a, b, c, d, e, f): x = a*col + c, y = e*row + f; b = d = 0
 *                    (host array of 6 doubles, like dst_transform)
 *   src_mask       : optional uint8 nodata mask of the source (non-zero = nodata)
 *   dst            : float32 destination, updated in place
 * Host or device pointers.                                                      */
int32_t ocsb_resample_nearest(ocsb_ctx *ctx,
                              const void *src, int32_t src_dtype, int64_t src_stride,
                              int32_t src_width, int32_t src_height,
                              const double *src_transform,
                              const uint8_t *src_mask, int64_t mask_stride,
                              float *dst, int64_t dst_stride,
                              int32_t dst_width, int32_t dst_height,
                              const double *dst_transform);

/* ---- A4 + A6..A10 as ONE streamed pass --------------------------------------
 * The reference applies every add_* to the whole raster before the next one
 * starts, rewriting a GeoTIFF each time (ocsmesh/hfun/raster.py:1257-1260,
 * :1432-1435).  Every op on this path is pixel-local given the DEM (+1 halo
 * row) and the vertex set, so the same result is obtained band by band:
 * ocsb_hfun_pass runs the whole op list on one row band after the other and,
 * for HOST buffers, overlaps the H2D copy of band k+1, the kernels of band k
 * and the D2H copy of band k-1 on three streams (full-duplex PCIe), with two
 * device band buffers instead of whole-raster staging.  It is also the way to
 * process rasters that do not fit in HBM.
 *   ops[i]   : OCSB_OP_FEATURE (verts_xy, nv, expansion_rate, target_size) =
 *              one ocsb_hfun_feature call; OCSB_OP_LIMITER (hmin, hmax,
 *              lower_bound, upper_bound) = one ocsb_hfun_flow_limiter call;
 *              fuse_clamp as in those calls.  Applied in list order.
 *   hfun_in  : pre-pass raster window, NULL = A4 initial state
 *   hfun_out : post-pass raster window (may alias hfun_in)
 *   dem      : DEM window (only needed when a limiter op is present)
 *   ghmin/ghmax : the HfunRaster's global bounds, NaN = None
 *   band_rows: rows per streamed band, 0 = automatic (~64 MB of hfun)
 * win->row_begin/row_end restrict the pass to a device tile as everywhere.
 * Bit-identical to the op-by-op calls (tests/test_pass_gpu.py).              */
#define OCSB_OP_FEATURE 0
#define OCSB_OP_LIMITER 1
typedef struct ocsb_hfun_op {
    int32_t       kind;
    int32_t       fuse_clamp;
    const double *verts_xy;       /* FEATURE */
    int64_t       nv;
    double        expansion_rate, target_size;
    double        hmin, hmax;     /* LIMITER: the method's arguments */
    double        lower_bound, upper_bound;
} ocsb_hfun_op;

int32_t ocsb_hfun_pass(ocsb_ctx *ctx, const ocsb_win *win,
                       const void *dem, int32_t dem_dtype, int64_t dem_stride,
                       const float *hfun_in, int64_t in_stride,
                       float *hfun_out, int64_t out_stride,
                       double dx, double dy, double ghmin, double ghmax,
                       const ocsb_hfun_op *ops, int32_t n_ops,
                       int32_t band_rows);

/* The whole op list over a LIST of semantic windows, every buffer resident on
 * the device: ONE tiled sweep (12 B/px for the list: DEM read, hfun read, hfun
 * write) in which the window table is a table of units of work of the same
 * launches -- the window loops of ocsmesh/hfun/raster.py:1212-1260, 1393-1436
 * over Raster.iter_windows (ocsmesh/raster.py:1577-1608) without a launch per
 * window.  Every window keeps its own np.linspace coordinates, search radius
 * and one-sided border differences.
 *   wins[k]     : window k (row_begin/row_end = the rows this device owns)
 *   dem         : element (0, 0) of the stored DEM block (dem_rows x dem_cols,
 *                 row stride dem_stride); window k's pixel (0, 0) lies
 *                 dem_off[k] elements into it (may be negative: a window that
 *                 starts above a band's stored rows)
 *   hfun_off[k] : element offset of window k's pixel (0, 0) in hfun_in /
 *                 hfun_out (one layout: in_stride == out_stride)
 *   keep_wh     : optional [n_wins][2]: only pixels (r < keep_h, c < keep_w) of
 *                 window k are written -- for overlapping windows (overlap > 0,
 *                 where the later window wins) with ONE op and distinct in /
 *                 out buffers; NULL = whole windows
 *   status_dev  : NULL = synchronous (candidate-pool overflow is retried inside
 *                 the call).  Device int32: ASYNCHRONOUS -- nothing waits on the
 *                 host; the pass is a no-op on the device if *status_dev != 0
 *                 on entry and sets it to 1 if a pool overflowed (the size
 *                 raster is then untouched), so a sequence of passes stops at
 *                 its first failure and can be redone from there.
 * Vertex sets must be device memory.  Bit-identical to the per-window,
 * per-op calls (tests/test_pass_gpu.py).                                     */
int32_t ocsb_hfun_pass_windows(ocsb_ctx *ctx, const ocsb_win *wins,
                               const int64_t *dem_off, const int64_t *hfun_off,
                               const int32_t *keep_wh, int32_t n_wins,
                               const void *dem, int32_t dem_dtype, int64_t dem_stride,
                               int32_t dem_rows, int32_t dem_cols,
                               const float *hfun_in, int64_t in_stride,
                               float *hfun_out, int64_t out_stride,
                               double dx, double dy, double ghmin, double ghmax,
                               const ocsb_hfun_op *ops, int32_t n_ops,
                               int32_t *status_dev);

/* One op of the list above on a LIST of semantic windows -- the window loops of
 * HfunRaster.add_feature (ocsmesh/hfun/raster.py:1212-1260) and
 * add_subtidal_flow_limiter (:1393-1436) over Raster.iter_windows
 * (ocsmesh/raster.py:1577-1608) -- with every buffer resident on the device.
 * Window k is wins[k]; its (virtual) row-0 / column-0 pixel lies dem_off[k]
 * elements into `dem` and hfun_off[k] elements into hfun_old / hfun_new (which
 * share one layout).  Windows are independent, so their launches are spread
 * over internal worker streams and overlap; results are those of n_wins
 * single-window calls.  hfun_old NULL = initial state.  Overlapping windows
 * (overlap > 0) must be given distinct old / new buffers by the caller, as for
 * the single-window calls, and in_order = 1: pixels shared by two windows are
 * then written in list order (the later window wins, as in the reference's
 * sequential loop): a window waits for the earlier windows it overlaps, the
 * others still run side by side.                                             */
int32_t ocsb_hfun_op_windows(ocsb_ctx *ctx, const ocsb_win *wins,
                             const int64_t *dem_off, const int64_t *hfun_off,
                             int32_t n_wins,
                             const void *dem, int32_t dem_dtype, int64_t dem_stride,
                             const float *hfun_old, int64_t old_stride,
                             float *hfun_new, int64_t new_stride,
                             double dx, double dy, double ghmin, double ghmax,
                             const ocsb_hfun_op *op, int32_t in_order);

/* ---- A12: pixel-wise min of co-registered hfun windows (build-defined) ---- */
int32_t ocsb_min_combine(ocsb_ctx *ctx, float *dst, int64_t dst_stride,
                         const float *src, int64_t src_stride,
                         int32_t width, int32_t height);

/* ---- A14: Mesh.interpolate worker ------------------------------------------
 * _mesh_interpolate_worker, one (raster, window)             ocsmesh/mesh/mesh.py:2285-2347
 * dem         : window of the raster band (dem_dtype), row-major
 * nodata_mask : optional uint8 window (1 = nodata), NULL = no nodata
 * xy          : K x 2 float64 node coordinates
 * method      : OCSB_INTERP_LINEAR / NEAREST / SPLINE
 * touched     : K uint8; set to 1 for every node this raster updates
 *               (= the boolean mask the worker returns), left alone otherwise
 * values      : K float64; values[k] is overwritten for touched nodes only,
 *               so calling this once per raster in list order reproduces the
 *               reference's last-wins scatter               ocsmesh/mesh/mesh.py:372-376 */
int32_t ocsb_interp(ocsb_ctx *ctx, const ocsb_win *win,
                    const void *dem, int32_t dem_dtype, int64_t dem_stride,
                    const uint8_t *nodata_mask, int64_t mask_stride,
                    const double *xy, int64_t K, int32_t method,
                    uint8_t *touched, double *values);

/* ---- A13: EuclideanMesh2D.interpolate over a raster LIST -------------------
 * values[mask] = vals for every raster in list order, last raster wins
 *                                                        ocsmesh/mesh/mesh.py:365-376
 * One (raster, window) pair of the list; same fields as ocsb_interp.          */
typedef struct ocsb_interp_raster {
    ocsb_win       win;
    const void    *dem;
    int32_t        dem_dtype;
    int64_t        dem_stride;
    const uint8_t *nodata_mask;   /* NULL = no nodata */
    int64_t        mask_stride;
    /* spline only, optional: the window's coefficients from an earlier
     * ocsb_spline_prefilter call (device memory, height x width doubles): the
     * prefilter (16 B/px, the bulk of a spline call) is then skipped.  The
     * coefficients depend on the window's DEM values and shape only.          */
    const double  *spline_coef;
} ocsb_interp_raster;

/* Equivalent to calling ocsb_interp once per entry in list order, but the node
 * coordinates are read (and the values written) once: every node walks the
 * list from the last entry to the first and stops at the first entry that
 * yields a value.  source (optional, K int32) receives the index of that
 * entry and is left alone for untouched nodes.  Linear / nearest run as ONE
 * kernel launch; spline entries are prefiltered and evaluated one by one.
 * perm (optional, K int32, linear / nearest only): see ocsb_bucket_nodes.     */
int32_t ocsb_interp_multi(ocsb_ctx *ctx, const ocsb_interp_raster *rasters,
                          int32_t n_rasters, const double *xy, int64_t K,
                          int32_t method, uint8_t *touched, double *values,
                          int32_t *source, const int32_t *perm);

/* Spatial bucketing of nodes (BASELINE.json north_star: "mesh nodes are sharded
 * by spatial bucket"): counting sort of the K nodes by the cell of a uniform
 * grid of `cell_size` over bbox = (xmin, ymin, xmax, ymax).  perm_out[i] (K
 * int32) is the original index of the i-th node in bucket order and
 * xy_sorted_out (K x 2 float64) the coordinates in that order.  Passing both to
 * ocsb_interp_multi (xy = xy_sorted, perm = perm) makes a warp's gathers hit
 * neighbouring texels while touched / values / source are still written at
 * the ORIGINAL node index, so results are independent of the bucketing.      */
int32_t ocsb_bucket_nodes(ocsb_ctx *ctx, const double *xy, int64_t K,
                          const double *bbox4, double cell_size,
                          int32_t *perm_out, double *xy_sorted_out);

/* Spline prefilter for OCSB_INTERP_SPLINE: builds the tensor-product
 * not-a-knot cubic B-spline coefficients of RectBivariateSpline(kx=ky=3, s=0)
 * (ocsmesh/mesh/mesh.py:2306-2314) for one window into coef[width*height]
 * float64 (device or host).  ocsb_interp(..., SPLINE) calls it internally;
 * exported for tests.                                                          */
int32_t ocsb_spline_prefilter(ocsb_ctx *ctx, const ocsb_win *win,
                              const void *dem, int32_t dem_dtype,
                              int64_t dem_stride, double *coef);

#ifdef __cplusplus
}
#endif
#endif /* OCSB200_H */
