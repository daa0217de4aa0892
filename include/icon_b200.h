/* include/icon_b200.h -- C ABI of the B200-native icon_rt hot path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++
 * or torch types.  It replaces what icon_rt's host code does through
 * Pipeline/OWL in the reference:
 *
 *   reference (file:line under /root/reference)             | this ABI
 *   --------------------------------------------------------+---------------------------
 *   Buffer<ICONCell> upload            hostCode.cu:717-810  | icb_set_grid
 *   volbounds/sphericalBounds/dataRange hostCode.cu:792-808 | icb_get_bounds
 *   build{Triangle,UserGeom,Shell}Accel hostCode.cu:440-682 | (inside icb_set_grid)
 *   transfunc.* launch params          hostCode.cu:947-951  | icb_set_transfunc
 *   camera.* launch params             hostCode.cu:936-945  | icb_set_camera
 *   unitDistance, ambient.., accumID  hostCode.cu:924-958  | icb_set_params
 *   pl.setRayGen + pl.launch()  hostCode.cu:860,963; pipeline.cu:1038-1075 | icb_render
 *   pl.present() D2H copy              pipeline.cu:1077-1091| icb_render_host
 *   sampleVolume()                     deviceCode.cu:58-125 | icb_locate (parity hook)
 *
 * All functions return ICB_OK (0) or a negative icb_error; nothing in the library
 * calls exit()/abort() (the reference prints-and-continues, hostCode.cu:19-34).
 * A handle is not thread-safe; distinct handles are independent.  Device work is
 * stream-ordered on the stream passed in (a cudaStream_t cast to void*, 0 = default).
 * "dev" pointers are CUDA device pointers on the handle's device; "host" pointers
 * are ordinary host memory.  Inputs may be freed as soon as a set_* call returns.
 */
#ifndef ICON_B200_H
#define ICON_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct icb_context *icb_handle;

typedef enum {
  ICB_OK = 0,
  ICB_ERR_INVALID = -1,      /* bad argument */
  ICB_ERR_CUDA = -2,         /* a CUDA call failed; see icb_last_error */
  ICB_ERR_NO_GRID = -3,      /* render/locate before icb_set_grid */
  ICB_ERR_NO_TRANSFUNC = -4, /* render before icb_set_transfunc */
  ICB_ERR_BAD_RECORD = -5,   /* numLayers outside [0,31] (ICONGrid.h:57,68-75) */
  ICB_ERR_NO_DEVICE = -6,    /* no usable CUDA device: the library has no CPU fallback */
  ICB_ERR_NCCL = -7          /* NCCL missing (libnccl.so.2 is loaded at run time, by icb_group_* only) or an NCCL call failed */
} icb_error;

#define ICB_RECORD_BYTES 284 /* sizeof(ICONCell), ICONGrid.h:59-76 */

/* icb_set_params flags */
#define ICB_FLAG_JITTER      1u /* generateRay's rnd() jitter on (deviceCode.cu:42-43); off: rnd():=0 */
/* icb_locate flags */
#define ICB_LOCATE_WALK      1u /* consecutive positions share walk state (64 per thread) */
#define ICB_LOCATE_EXACT     2u /* candidate-bin brute force only (no walk, no margins) */

typedef struct {
  float volLo[3], volHi[3];   /* Cartesian bounds, ICONCell::getBounds union (hostCode.cu:805) */
  float sphLo[3], sphHi[3];   /* (radius, lat, lon) bounds (hostCode.cu:799-804) */
  float dataRange[2];         /* hostCode.cu:806 */
  float unitDistance;         /* 10^(floor(log10(r_inner))-3), hostCode.cu:838-840 */
} icb_bounds;

typedef struct {
  int64_t numRecords, numColumns, numVertices;
  int64_t numBins, numBinEntries;
  int64_t numGenericColumns;      /* columns that always take the exact candidate-bin path */
  int64_t numGlobalRecords;       /* degenerate records (unbounded acceptance region) checked for every sample */
  int64_t deviceBytes;
  double  setupSeconds;           /* host preprocessing + upload of the last icb_set_grid */
  /* counters of the last icb_render / icb_locate (device atomics, read back on demand) */
  uint64_t steps;                 /* sample positions evaluated */
  uint64_t hits;                  /* positions located in a record */
  uint64_t slowSteps;             /* positions that left the in-prism fast path */
  uint64_t exactQueries;          /* positions resolved by the exact candidate-bin query */
  uint64_t raysMarched;           /* rays with at least one shell segment */
  uint64_t totalSteps, totalHits; /* summed over all frames since icb_reset_totals (no per-frame readback needed) */
  int32_t  kernelLaunches;        /* kernels launched by the last render/locate call */
  float    lastKernelMs;          /* device time of the last render/locate kernel (events) */
} icb_stats_t;

/* ---- lifetime ------------------------------------------------------------------- */
int icb_create(int device, icb_handle *out);
int icb_destroy(icb_handle h);
const char *icb_last_error(icb_handle h);     /* h may be NULL: last creation error */
const char *icb_version(void);

/* ---- scene ------------------------------------------------------------------------ */
/* records: host pointer to n packed 284-byte ICONCell records, bytes identical to a
 * `.ic` file (convert_icon.cpp:383-387).  Record id == index in this array.
 * Synchronous; the records are only read and may be freed on return.  Host memory beyond the caller's records:
 * about 0.3 KB per record at the peak (the 256-byte lines per record are streamed to the device in chunks). */
int icb_set_grid(icb_handle h, const void *records, size_t n);
/* Replace the scalars only (time series): values[n*32] host floats laid out like
 * ICONCell::value of each record, record order as in icb_set_grid. */
int icb_set_scalars(icb_handle h, const float *values, size_t n);
/* same, from a device buffer (n*32 floats), stream-ordered */
int icb_set_scalars_dev(icb_handle h, const float *values_dev, size_t n, void *stream);
int icb_get_bounds(icb_handle h, icb_bounds *out);
/* rgba: host pointer to size*4 floats (Transfunc::rgbaLUT, transfunc.h:29-49) */
int icb_set_transfunc(icb_handle h, const float *rgba, int size, float lo, float hi, float opacityScale);
int icb_set_camera(icb_handle h, const float org[3], const float dir_00[3],
                   const float dir_du[3], const float dir_dv[3]);
/* dt<=0 selects unitDistance/2; termAlpha<=0 selects 0.99 */
int icb_set_params(icb_handle h, float unitDistance, float dt, const float ambientColor[3],
                   float ambientRadiance, int accumID, unsigned flags, float termAlpha);
/* locator tuning: marginMeters<0 keeps the default; see DESIGN.md "margins" */
int icb_set_locator(icb_handle h, float marginMeters, int maxWalkSteps);
/* skip-ahead over samples that provably stay in the current prism (default on; results are identical either way) */
int icb_set_skip_ahead(icb_handle h, int enabled);
/* Marcher: split rays.  A ray's samples are composited in chunks of 2^chunkSamplesLog2 consecutive sample indices per shell
 * segment, the chunks merged in order with `over` (CPU statement: oracle/ref_driver.cpp march_pixel).  For the frame's
 * heaviest tiles -- up to maxHeavyTiles tiles whose cost in an earlier frame was >= 2^minTileCostLog2 lane iterations --
 * the chunks of a ray are marched by different lanes (a pass in front of the pixel pass), so that the frame's critical
 * path is one chunk long instead of one limb ray long; WHICH rays are split never changes a bit of the frame.
 * mode: ICB_SPLIT_NEVER (default), ICB_SPLIT_ALWAYS, or ICB_SPLIT_AUTO (split when the frame is a sort-first share: a tile
 * list of at most half the frame's tiles).  Measured on C3 (profiles/r02_scale_*.json): the heaviest tile alone drops from
 * 3.1 ms to 1.1 ms, but the speculative chunks behind the sample at which a ray terminates cost more than the tail they
 * remove once frames are kept in flight -- slower at every N from 1 to 8 -- so the pass is off unless asked for.
 * Defaults 10, 384, 16, NEVER; -1 keeps a value.
 * The compositing ARITHMETIC follows the mode, not the per-frame decision: with ICB_SPLIT_NEVER a ray is composited by the
 * plain A.3 recurrence (chunkSamplesLog2 has no effect); with ALWAYS or AUTO every ray of every frame is composited in chunks,
 * whether or not any ray of that frame is actually split.  Frames of the two arithmetics differ in fp32 rounding (inside the
 * 1e-3 image tolerance, not bit for bit).  Handles whose frames are compared or assembled bit for bit -- every rank of an
 * icb_group and a single-GPU handle rendering the same view -- must therefore share mode and chunkSamplesLog2;
 * icb_group_begin refuses (ICB_ERR_INVALID) local handles that disagree; ranks in other processes are the caller's to keep equal. */
enum { ICB_SPLIT_NEVER = 0, ICB_SPLIT_ALWAYS = 1, ICB_SPLIT_AUTO = 2 };
int icb_set_ray_split(icb_handle h, int chunkSamplesLog2, int maxHeavyTiles, int minTileCostLog2, int mode);
/* Which ray-gen program icb_render / icb_render_host run -- the analogue of Pipeline::setRayGen(ptx, name)
 * (hostCode.cu:860, pipeline.h:88-104):
 *   ICB_INTEGRATOR_MARCH     front-to-back ray marcher with early termination (default)
 *   ICB_INTEGRATOR_WOODCOCK  the reference's woodcockTrackingAE (deviceCode.cu:239-275): one stochastic collision
 *                            per pixel per frame (majorant 1, boxTest against the volume bounds), running mean
 *                            over accumID.  Bit-identical to the reference's CPU path on the same host libm
 *                            (glibc's logf is evaluated on the device once the first call has verified, over all 2^24
 *                            possible arguments, that the device sequence reproduces the calling host's libm; otherwise,
 *                            or with ICB_LOG_TABLE=1, it is read from a table built by that libm).
 * Tile lists (icb_set_tiles, icb_render_tiles) and icb_render_trace apply to the marcher only.
 *   ICB_INTEGRATOR_WOODCOCK_GRID  the reference's DEFAULT ray-gen woodcockTrackingWithAccel (deviceCode.cu:281-341,
 *                            selected at hostCode.cu:860) in GRID_ACCEL_MODE: dda3 (DDA.h:35-136) over a macrocell grid on
 *                            the volume bounds, woodcockTracking per macrocell with majorant maxOpacities[leafID].  The
 *                            grid is built on the device: per-macrocell value ranges by float atomicMin/Max
 *                            (hostCode.cu:36-56, 205-255), max opacities per transfer function (hostCode.cu:400-434);
 *                            rebuilt after icb_set_grid / icb_set_scalars*, re-classified after icb_set_transfunc.
 *                            Bit-identical to the reference's CPU path.
 *   ICB_INTEGRATOR_WOODCOCK_SPHERE  the same ray-gen in SPHERE_ACCEL_MODE, the accel mode the reference STARTS in
 *                            (hostCode.cu:75, :915): sdda (ShellAccel.h:83-199) over the (rad, lat, lon) shell accel on
 *                            the spherical bounds, value ranges by buildShell_ICON (hostCode.cu:298-334), max opacities
 *                            by computeMaxOpacities(ShellAccel) (:362-397).  sdda's cell ids come from toSpherical
 *                            (ICONGrid.h:36-42: asinf, atan2f); the device evaluates glibc's operation sequences for
 *                            both (csrc/fdlibm_trig.cuh: glibc 2.39's sequences, an assumption about the host libm that
 *                            tests/test_fdlibm_trig.py checks -- every 251st bit pattern + 8M atan2f pairs by default), so this too is
 *                            bit-identical to the reference's CPU path. */
enum { ICB_INTEGRATOR_MARCH = 0, ICB_INTEGRATOR_WOODCOCK = 1, ICB_INTEGRATOR_WOODCOCK_GRID = 2, ICB_INTEGRATOR_WOODCOCK_SPHERE = 3 };
int icb_set_integrator(icb_handle h, int integrator);
/* shell accel of ICB_INTEGRATOR_WOODCOCK_SPHERE: (rad, lat, lon) cells; the reference uses 1 x 1024 x 1024
 * (hostCode.cu:654), the default here */
int icb_set_accel_shell(icb_handle h, int nRad, int nLat, int nLon);
/* parity hook: the shell accel as the ray-gen reads it (layout as icb_get_majorants; cell = (lon*nLat + lat)*nRad + rad) */
int icb_get_shell_majorants(icb_handle h, float *maxOpacities_host, float *valueRanges_host, size_t numCells);
/* macrocell grid of ICB_INTEGRATOR_WOODCOCK_GRID; the reference uses 256^3 (hostCode.cu:670), the default here */
int icb_set_accel_grid(icb_handle h, int nx, int ny, int nz);
/* parity hook: the grid as the ray-gen reads it -- maxOpacities (nx*ny*nz floats) and, if not NULL, the value
 * ranges (2 floats per macrocell: lower, upper; empty = {FLT_MAX,-FLT_MAX}) -- copied to host memory */
int icb_get_majorants(icb_handle h, float *maxOpacities_host, float *valueRanges_host, size_t numMacrocells);
/* parity hook: logf(1 - k/2^24), k = 0..2^24-1 -- every argument a free flight of the Woodcock ray-gens can take
 * (deviceCode.cu:163) -- as the kernels compute it (csrc/glibc_logf.cuh: glibc's operation sequence, both of its x86-64
 * builds) against the calling host's libm.  mismatches[0]: the FMA form (the one the kernels use), [1]: the SSE2 form. */
int icb_selftest_logf(icb_handle h, int64_t mismatches[2]);
/* Which sampler gives a located sample its value -- the analogue of Volume::mode (Params.h:52-67;
 * CPU / USER_GEOM / TRIANGLE modes all end in ICONCell::getValue, CUBQL_MODE in intersectWedgeEXT):
 *   ICB_SAMPLER_CELL   value[layer] of the located record, piecewise constant (ICONGrid.h:146-153; default)
 *   ICB_SAMPLER_PRISM  barycentric-in-triangle x linear-in-height interpolation inside the located prism: the wedge
 *                      shape functions of UElems.h:176-187 at the closed-form parametric coordinates that
 *                      intersectWedgeEXT (UElems.h:214-311) finds by Newton iteration, over vertex scalars =
 *                      mid-level values (cf. hostCode.cu:574-575) averaged over the columns sharing the vertex.
 *                      Record ids are those of ICB_SAMPLER_CELL (location is unchanged).  Allocates
 *                      numVertices x (layers of the tallest column + 1) floats; recomputed by icb_set_scalars*. */
enum { ICB_SAMPLER_CELL = 0, ICB_SAMPLER_PRISM = 1 };
int icb_set_sampler(icb_handle h, int sampler);

/* ---- hot path --------------------------------------------------------------------- */
/* Frame buffers are caller-owned device memory shaped like dvr_course::Frame (fb.h:27-48):
 * fbPointer w*h RGBA8 (rows bottom-up), accumBuffer w*h float4, fbDepth w*h float or NULL. */
int icb_render(icb_handle h, uint32_t *fbPointer_dev, float *accumBuffer_dev, float *fbDepth_dev,
               int width, int height, void *stream);
/* Same through host buffers: params up, render, frame down (either pointer may be NULL; with
 * accumBuffer_host == NULL the running mean stays on the device between frames, as in the reference
 * where present() copies fbPointer only, pipeline.cu:1077-1091).  Pinned buffers (icb_host_alloc) make
 * the copies asynchronous DMA. */
int icb_render_host(icb_handle h, uint32_t *fbPointer_host, float *accumBuffer_host,
                    int width, int height);
/* Per-pixel trace of the last-configured frame: perPixel_dev[w*h*4] int32 =
 * {steps, hits, idHashLo, idHashHi} (FNV-1a over the located record ids, -1 = miss). */
int icb_render_trace(icb_handle h, uint32_t *fbPointer_dev, float *accumBuffer_dev,
                     int32_t *perPixel_dev, int width, int height, void *stream);
/* Located-id sequence of one pixel (host output); returns the number of steps (>=0) */
int icb_trace_pixel(icb_handle h, int x, int y, int width, int height,
                    int32_t *ids_host, int maxIds, float rgba_host[4]);
/* icb_trace_pixel without a single allocation -- no counterpart in the reference (which has no per-pixel trace at all);
 * it exists for hosts that cannot let the library call cudaMalloc / cudaFree once they are running:
 *   - a present loop with frames in flight (icb_render_host_async, icb_group_render): icb_trace_pixel frees its three
 *     staging buffers on return, and cudaFree waits for everything outstanding on the device, so a probe stalls the loop;
 *   - a process whose device memory is fully committed (an R2B10 replica plus the host application's own pool), where
 *     an allocation may fail at any moment, or whose embedding forbids driver allocations after start-up.
 * Same arguments, same results, bit for bit, as icb_trace_pixel; the ids, the step count and the RGBA are staged in an
 * array that is part of the library's device image (64.25 KB on every GPU that loads the library, held until exit).
 * Limit: maxIds <= ICB_TRACE_STATIC_MAX_IDS, otherwise ICB_ERR_INVALID with the reason in icb_last_error (the return
 * value still counts every step of the ray, as icb_trace_pixel's does; only the first maxIds ids are stored).
 * The array is shared by ALL handles on a GPU and keeps the previous call's contents (the header is zeroed before each
 * launch): two calls that can overlap on one GPU -- two handles, two host threads -- race and must be serialised by the
 * caller.  The first call on each GPU looks the array's device address up (and may load the module there): make that
 * call during start-up, before any stream capture. */
#define ICB_TRACE_STATIC_MAX_IDS 16384
int icb_trace_pixel_static(icb_handle h, int x, int y, int width, int height,
                           int32_t *ids_host, int maxIds, float rgba_host[4]);
/* positions -> record id (-1 = none) and value; pos n*3 floats */
int icb_locate(icb_handle h, const float *pos_dev, size_t n, int32_t *ids_dev, float *values_dev,
               unsigned flags, void *stream);
int icb_locate_host(icb_handle h, const float *pos_host, size_t n, int32_t *ids_host,
                    float *values_host, unsigned flags);

/* ---- sort-first tiles (multi-GPU) --------------------------------------------------- */
#define ICB_TILE 16 /* screen tiles are ICB_TILE x ICB_TILE pixels */
/* Restrict rendering to the listed tiles (ids in row-major tile order); n==0 -> all tiles.
 * DEPRECATED in favour of icb_set_tiles_async (below), kept for callers that render one frame at a time: this form copies
 * the list synchronously and is NOT ordered against frames still running -- call it only between frames, after the previous
 * frame's stream work has completed.  A loop with frames in flight must use icb_set_tiles_async (icb_group_* does). */
int icb_set_tiles(icb_handle h, const int32_t *tileIds_host, int n);
/* Render the assigned tiles into a compact buffer: tile i of the list occupies
 * rgba8_dev[i*256 .. i*256+255] (and accum_dev[i*1024 ..] if non-NULL), row-major inside the tile. */
int icb_render_tiles(icb_handle h, uint32_t *rgba8_dev, float *accum_dev, int width, int height, void *stream);
/* Scatter compact tiles (as produced by icb_render_tiles) into a full frame. */
int icb_scatter_tiles(icb_handle h, const uint32_t *rgba8_dev, const float *accum_dev,
                      const int32_t *tileIds_dev, int n, uint32_t *fbPointer_dev, float *accumBuffer_dev,
                      int width, int height, void *stream);
/* per-tile step counts of the last render (cost estimate for the next frame), device int32[numTiles] */
int icb_tile_costs(icb_handle h, int32_t *costs_dev, int numTiles, void *stream);
/* Several frames in flight.  The handle keeps its per-frame device state (tile list, work cursor, counters, per-tile costs,
 * timing events) in ICB_FRAME_SLOTS (16) slots; every call that touches that state (icb_set_tiles*, icb_render*, icb_tile_costs,
 * icb_set_tile_costs, icb_stats) uses the slot selected here (default 0).  A caller that alternates slot AND stream AND
 * frame buffers from frame to frame lets the GPU fill the tail of one frame (a few long limb rays: a serial chain of
 * dependent fetches) with the head of the next.  icb_stats' totalSteps/totalHits sum over all slots. */
#define ICB_FRAME_SLOTS 16
int icb_set_frame_slot(icb_handle h, int slot);
/* how many frames the caller keeps in flight (1..ICB_FRAME_SLOTS, default 1): each frame's persistent kernel is then
 * launched with its share of the device's resident CTA slots, so that consecutive frames' kernels are co-resident */
int icb_set_frames_in_flight(icb_handle h, int n);
/* icb_render_host without the wait (a present loop with two frames in flight): renders the current slot's frame on the
 * slot's own stream and enqueues the download of the RGBA8 frame into page-locked host memory (icb_host_alloc) behind it.
 * icb_frame_wait(h, slot) returns once that slot's frame is in host memory.  The running mean stays on the device, in
 * the SLOT's own accumulation buffer: a progressive sequence (accumID > 0) must keep to one slot, and the call returns
 * ICB_ERR_INVALID for accumID > 0 while icb_set_frames_in_flight(n > 1) is in force. */
int icb_render_host_async(icb_handle h, uint32_t *fbPointer_pinned, int width, int height);
int icb_frame_wait(icb_handle h, int slot);
/* multi-GPU: hand the library the all-reduced per-tile costs of the last frame, so that a rank orders its NEW tile set
 * heaviest-first with the costs measured by whichever rank rendered each tile (stream-ordered) */
int icb_set_tile_costs(icb_handle h, const int32_t *costs_dev, int numTiles, void *stream);
/* icb_set_tiles without the host synchronisation: the copy is ordered on `stream` (after the frame that still reads the
 * previous list); tileIds_pinned must be page-locked and stay untouched until the copy has run */
int icb_set_tiles_async(icb_handle h, const int32_t *tileIds_pinned, int n, void *stream);

/* ---- multi-GPU: sort-first frames behind the C ABI (SURVEY.md §8b icb_group_create / icb_group_render, §8e) ----------
 * The reference has no multi-GPU path (one GPU, one thread per pixel, owlLaunch2D pipeline.cu:1064), so
 * nothing is replaced here; the contract is SURVEY.md §8e: the image is cut into ICB_TILE x ICB_TILE tiles, every rank
 * renders a disjoint tile set against its own full replica of the grid, finished tiles are gathered on rank 0 with
 * ncclSend/ncclRecv over NVLink, the per-tile costs are all-reduced, and the next partition is cut from them -- identically
 * on every rank.  The assembled frame is bit-identical to the single-GPU frame (no cross-tile blending).
 * A group is this process's share of `world` ranks: ONE rank per process (icb_group_create; ranks exchange the id made
 * by icb_group_unique_id on rank 0 through whatever launcher they have: torchrun's store, MPI, a file) or ALL ranks in one
 * process (icb_group_create_local).  Handles stay owned by the caller and must outlive the group. */
typedef struct icb_group_s *icb_group;
#define ICB_GROUP_ID_BYTES 128                 /* sizeof(ncclUniqueId) */
int icb_group_unique_id(void *id /* ICB_GROUP_ID_BYTES */);
int icb_group_create(icb_handle h, int rank, int world, const void *id, icb_group *out);
int icb_group_create_local(const icb_handle *handles, int nDev, icb_group *out);
int icb_group_destroy(icb_group g);
int icb_group_size(icb_group g);
const char *icb_group_last_error(icb_group g);   /* g may be NULL: last creation error */
/* The grid is built ONCE, by rank 0 (icb_set_grid on its handle), and the finished device arrays reach the other ranks by
 * ncclBroadcast over NVLink.  Ranks other than 0 pass records == NULL. */
int icb_group_set_grid(icb_group g, const void *records, size_t n);
/* icb_set_transfunc / icb_set_camera / icb_set_params on every local handle */
int icb_group_set_transfunc(icb_group g, const float *rgba, int size, float lo, float hi, float opacityScale);
int icb_group_set_camera(icb_group g, const float org[3], const float dir_00[3], const float dir_du[3], const float dir_dv[3]);
int icb_group_set_params(icb_group g, float unitDistance, float dt, const float ambientColor[3], float ambientRadiance,
                         int accumID, unsigned flags, float termAlpha);
/* frame size and how many frames the group keeps in flight (1..ICB_FRAME_SLOTS); allocates the per-slot buffers */
int icb_group_begin(icb_group g, int width, int height, int framesInFlight);
/* One sort-first frame (marcher) with the handles' current camera / parameters, enqueued WITHOUT waiting for the device
 * (the one wait is for the tile costs of the frame that occupied the slot `framesInFlight` frames ago).
 * fbPointer_pinned: rank 0 only, may be NULL -- page-locked host memory that receives the assembled RGBA8 frame.
 * Returns the slot (>= 0) the frame occupies, or a negative icb_error. */
int icb_group_render(icb_group g, uint32_t *fbPointer_pinned);
/* returns once the frame in `slot` (slot < 0: every slot) has left the device */
int icb_group_wait(icb_group g, int slot);
/* rank 0's assembled frame of `slot` (device memory owned by the group; NULL where rank 0 is not local) */
int icb_group_frame(icb_group g, int slot, uint32_t **fbPointer_dev);
typedef struct {
  int32_t world, localRanks, framesInFlight, pad;
  int64_t frames;              /* icb_group_render calls since icb_group_begin / icb_group_reset_stats */
  double hostMsPerFrame;       /* host time inside icb_group_render, per frame ... */
  double waitMsPerFrame;       /* ... of which waiting for the device (tile costs of an earlier frame): 0 = the host is the limit */
} icb_group_stats_t;
int icb_group_stats(icb_group g, icb_group_stats_t *out);
int icb_group_reset_stats(icb_group g);
/* Time series (BASELINE configs[3]): a timestep's scalars (n*32 floats, icb_set_scalars layout, page-locked) are uploaded
 * 1/world per rank over each rank's own PCIe link and exchanged by one ncclAllGather over equal, padded slabs on a
 * communicator of its own; double-buffered: values_pinned must stay untouched until the next call but one.
 * icb_group_swap_scalars then swaps the oldest streamed timestep in: after every frame in flight, before every frame
 * enqueued afterwards.  Neither call waits for the device.  At most TWO streamed timesteps may be pending (staged, not yet
 * swapped in): a third icb_group_stream_scalars before a swap returns ICB_ERR_INVALID instead of overwriting the oldest. */
int icb_group_stream_scalars(icb_group g, const float *values_pinned, size_t n);
/* Which floats of values_pinned icb_group_stream_scalars reads in THIS process: [*firstFloat, *firstFloat + *numFloats) of the
 * n*32 (one rank per process: that rank's slab; icb_group_create_local: all of them).  Nothing outside that range is ever
 * dereferenced, so a rank may keep only its slab of every timestep in pinned memory and pass slab_pointer - *firstFloat. */
int icb_group_scalar_slab(icb_group g, size_t n, size_t *firstFloat, size_t *numFloats);
int icb_group_swap_scalars(icb_group g);
/* The partition itself (host only, no GPU needed): tile ids (row-major tile index) rank-major in tileIds[tilesX*tilesY],
 * rank r owns tileIds[offsets[r] .. offsets[r+1]).  costs: per-tile costs of an earlier frame indexed by tile id, or NULL. */
int icb_partition_tiles(int tilesX, int tilesY, const int32_t *costs, int nRanks, int32_t *tileIds, int32_t *offsets);

int icb_stats(icb_handle h, icb_stats_t *out);
/* Developer counters of the last icb_render_trace frame: {warp iterations of the march loop, lane iterations,
   lane iterations that stayed in their prism, changed layer, fetched a node, came back undecided, skipped
   nothing, 0}. */
int icb_debug_counters(icb_handle h, uint64_t out[8]);
int icb_reset_totals(icb_handle h);

/* ---- device memory helpers (so that a host without CUDA headers can own a Frame, fb.h:27-48) ---- */
int icb_dev_alloc(icb_handle h, size_t bytes, void **out_dev);
int icb_dev_free(icb_handle h, void *ptr_dev);
int icb_dev_upload(icb_handle h, void *dst_dev, const void *src_host, size_t bytes);
int icb_dev_download(icb_handle h, void *dst_host, const void *src_dev, size_t bytes);
int icb_dev_sync(icb_handle h);
/* page-locked host memory for frames / scalar time steps that are streamed every frame */
int icb_host_alloc(icb_handle h, size_t bytes, void **out_host);
int icb_host_free(icb_handle h, void *ptr_host);

/* ---- host mirror helpers (no CUDA needed), C++ originals in icon-ray-tracing_b200/host/icon_host.hpp ---- */
/* Camera::viewAll + getScreen -> launch-param camera (camera.h:86-104, hostCode.cu:936-945) */
void icb_camera_view_all(const float volLo[3], const float volHi[3], float fovy, float aspect, int width, int height,
                         float org[3], float dir_00[3], float dir_du[3], float dir_dv[3]);
/* Camera::setOrientation + getScreen (camera.h:34-96) */
void icb_camera_orient(const float eye[3], const float poi[3], const float up[3], float fovy, float aspect,
                       int width, int height, float org[3], float dir_00[3], float dir_du[3], float dir_dv[3]);
/* resampleLUT (dvr_course-common.h:44-70): src nsrc*4 floats -> dst ndst*4 floats */
void icb_resample_lut(const float *src, int nsrc, float *dst, int ndst);
/* `.ic` loader with the reference's filtering (hostCode.cu:717-758). Returns the number of surviving
 * records (<= capacity written to records), or a negative icb_error.  records may be NULL to count. */
int64_t icb_load_ic(const char *path, int64_t maxNumCells, float latLo, float latHi, float lonLo, float lonHi,
                    void *records, int64_t capacity);

/* ---- convert_icon's array -> `.ic` half (no CUDA, no NetCDF needed; tools/convert_icon/convert_icon.cpp:317-391) ------
 * The caller hands over what the converter holds after reading its files: clon_vertices / clat_vertices (double, 3 per
 * cell), HSURF (double per cell), numHhl HHL level fields and numData data level fields (float per cell, each with its
 * `height` level index).  Levels are sorted by `height` descending, numLayers = min(numData, maxLayers) (maxLayers < 0:
 * no cap; the converter's default is 5), columns are cut into records of <= 31 layers.
 *   ICB_INGEST_REFERENCE  statement for statement, defects included: the last record of a column gets
 *                         (numLayers % 32) - 1 layers (90 levels -> 31 + 31 + 25; 32 levels -> 31 and a record with -1)
 *   ICB_INGEST_FIXED      the last record gets the layers that are left (31 + 31 + 28)
 *   ICB_INGEST_SURFACE_LEVEL (flag, or-ed to either) the column starts at the lowest half level, H[0] = R + HHL_bottom - hsurf,
 *                         instead of the converter's R + hsurf (which lies ABOVE its own H[1] = R over land and shifts every
 *                         data level one layer down); consumes one more HHL field
 * Unused height/value entries are zero (the converter writes uninitialised stack there).  Returns the number of records
 * (cells * ceil(numLayers / 31); records == NULL only counts) or a negative icb_error -- ICB_ERR_INVALID also when the
 * level fields handed over are fewer than the records would consume (the converter reads past the end there). */
enum { ICB_INGEST_REFERENCE = 0, ICB_INGEST_FIXED = 1, ICB_INGEST_SURFACE_LEVEL = 2 };
int64_t icb_ingest_num_records(int64_t cells, int numDataLevels, int maxLayers);
int64_t icb_ingest_columns(int64_t cells, const double *clon_vertices, const double *clat_vertices, const double *hsurf,
                           int numHhl, const int32_t *hhlHeights, const float *const *hhlValues,
                           int numData, const int32_t *dataHeights, const float *const *dataValues,
                           int maxLayers, int mode, void *records, int64_t capacity);
/* a data level normalised to [0,1] by its own extrema, in double, narrowed to float (convert_icon.cpp:317-333) */
int icb_ingest_normalize(const double *var, int64_t n, float *out);

/* ---- synthetic grid generator (no CUDA needed) ------------------------------------ */
int64_t icb_gen_num_columns(int n, int k);
int64_t icb_gen_num_records(int n, int k, int levels);
/* writes icb_gen_num_records() packed records; timestep shifts the scalar field in longitude */
int icb_gen_grid(int n, int k, int levels, float exaggeration, int timestep, void *records, int64_t capacity);
/* the scalars alone (32 floats per record, icb_set_scalars layout) of records [firstRecord, firstRecord + numRecords) for
 * ntimesteps consecutive timesteps: values[t][numRecords * 32], bit-identical to icb_gen_grid's value arrays.  A rank of
 * the time-series workload generates its own slab of every timestep with it (icb_group_scalar_slab). */
int icb_gen_values(int n, int k, int levels, int timestep0, int ntimesteps, int64_t firstRecord, int64_t numRecords, float *values);

#ifdef __cplusplus
}
#endif
#endif
