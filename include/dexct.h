/*
 * dexct.h -- C ABI of libdexct.so: the B200 (sm_100a) implementation of the data-parallel hot path of
 * gjadick/dex-ct-sim.  Plain pointers and sizes only; no C++/torch types cross this boundary.
 *
 * The reference has no FFI of its own -- its "operator API" for this path is three Python call sites:
 *     xtomosim.forward_project.get_sino(ct, phantom, spec)            /root/reference/main.py:120
 *     matdecomp.get_basismat_sinos(ct, s1, s2, spec1, spec2, n_iters)  /root/reference/main.py:153, matdecomp.py:167
 *     xtomosim.back_project.get_recon(sino, ct, spec, N, FOV, ramp)    /root/reference/main.py:134,168
 * Each entry point below states which of those (and which lines of matdecomp.py) it replaces.  The ctypes
 * binding a maintainer would add on the reference side is shown in INTEGRATION.md and lives, working, in
 * dex-ct-sim_b200/dexct/_native.py.
 *
 * Conventions
 *   - Every data pointer is a DEVICE pointer unless its name ends in _host.  The caller owns all buffers.
 *   - Work is enqueued on the given CUDA stream (a cudaStream_t passed as void*; NULL = default stream) and the
 *     call returns without synchronising, except the *_host helpers which synchronise that stream.
 *   - Return value: 0 on success, negative DEXCT_E* on failure; dexct_last_error() gives the message of the last
 *     failure on the calling thread.  No global mutable state besides that thread-local string.
 *   - dtype: DEXCT_F32 or DEXCT_F64 selects the arithmetic of the kernels and the element type of every
 *     "void*" floating-point buffer of that call.
 */
#ifndef DEXCT_H
#define DEXCT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DEXCT_VERSION 100

#define DEXCT_F32 0
#define DEXCT_F64 1

#define DEXCT_OK            0
#define DEXCT_EINVAL      (-1)   /* bad argument (null pointer, non-positive size, unsupported combination) */
#define DEXCT_ECUDA       (-2)   /* CUDA runtime error, text in dexct_last_error() */
#define DEXCT_ENODEVICE   (-3)   /* no sm_100 device: there is NO CPU fallback */
#define DEXCT_ECAPACITY   (-4)   /* tables too large for on-chip staging */

#define DEXCT_MAX_PEERS         16   /* ranks of one node whose memory a rank can map (peer section below) */
#define DEXCT_PEER_HANDLE_BYTES 64   /* size of an exported memory handle */

/* Scan geometry: third-generation equiangular fan (cone when n_rows > 1), source on a circle in the z = 0 plane.
 * Mirrors the fields of the reference's FanBeamGeometry(N_channels, N_proj, gamma_fan, SID, SDD, h_iso, ...)
 * (/root/reference/plots.py:109-111, input/params.txt:18-27).
 *   beta_v  = beta0  + v * dbeta      source angle of view v;   S = SID (cos beta, sin beta, 0)
 *   gamma_c = gamma0 + c * dgamma     fan angle of channel c;   ray direction -(cos(beta+gamma), sin(beta+gamma))
 *   zdet_r  = z0     + r * dzdet      z of detector row r at the detector (distance SDD from the source)      */
typedef struct DexctScan {
    int32_t n_views;      /* views in the whole scan (size of the angle table) */
    int32_t n_rows;       /* detector rows (1 = single-slice fan beam) */
    int32_t n_ch;         /* detector channels */
    int32_t reserved;
    double sid, sdd;      /* source-isocentre and source-detector distance [cm] */
    double beta0, dbeta;
    double gamma0, dgamma;
    double z0, dzdet;
} DexctScan;

/* Voxel phantom: uint8 material labels, labels[(k*ny + j)*nx + i], centred on the origin, voxel size in cm
 * (input/params.txt:10-15; VoxelPhantom at plots.py:124).                                                   */
typedef struct DexctVolume {
    int32_t nx, ny, nz;
    int32_t reserved;
    double dx, dy, dz;
} DexctVolume;

int         dexct_version(void);
const char *dexct_last_error(void);
/* Number of CUDA devices with compute capability 10.x visible to the process (0 => every call fails). */
int         dexct_device_count(void);

/* ---- forward projection: replaces get_sino (main.py:120) ------------------------------------------------ */

/* Bytes of device scratch needed by dexct_siddon_prepare for this scan + volume:
 *   - the angle tables (cos/sin of the view and channel angles, detector row heights; float64);
 *   - two minor-axis-contiguous "pair" layouts of the label volume, one per possible major axis (2 B per voxel each,
 *     one-voxel zero aprons): what the trace kernels and the K > 128 fallback walk;
 *   - four "skip" layouts (major axis x sign of the minor slope, 2 B per voxel each): ordinary pairs or homogeneous
 *     stretches of up to 127 columns -- what the production kernels walk;
 *   - for volumes with more than one slice, two z-homogeneity radius volumes (1 B per voxel each) used to build them.
 * About 14 B per voxel for a 3-D volume: 1 MiB for a 512 x 512 slice, 1.8 GiB for 512^3, 14 GiB for 1024^3.           */
size_t dexct_siddon_workspace_bytes(const DexctScan *scan, const DexctVolume *vol);

/* Fill the workspace: angle tables (device double precision), pair layouts and skip layouts.  Must be called again
 * whenever labels, scan or vol change.  Cost: a few passes over the volume (0.1 ms for a 512 x 512 slice, 10 ms for
 * 512^3, 60 ms for 1024^3).  Refuses scans whose rays climb more than one slice per column (thin slices under a tall
 * detector): the column routine assumes at most one z-plane crossing per column.                              */
int dexct_siddon_prepare(const DexctScan *scan, const DexctVolume *vol, const uint8_t *labels,
                         void *workspace, void *stream);

/* Copy the angle tables the device uses to host arrays (each may be NULL): cos/sin beta [n_views],
 * cos/sin gamma [n_ch], zdet [n_rows].  Synchronises.  Lets a checker trace exactly the rays the kernel traces. */
int dexct_siddon_tables_host(const DexctScan *scan, const void *workspace, double *cos_beta_host,
                             double *sin_beta_host, double *cos_gamma_host, double *sin_gamma_host,
                             double *zdet_host, void *stream);

/* Siddon traversal + fused polyenergetic Beer-Lambert detection for views [view_begin, view_begin+view_count).
 *   mu_KxE   [K][nE]        linear attenuation of material label m at energy bin e [1/cm]
 *   W_SxE    [nSpec][nE]    detected weight of spectrum s in bin e (I0 * eta * (E if energy-integrating) * dE,
 *                           the reference's i0, matdecomp.py:146-150)
 *   counts   [nSpec][view_count][n_rows][n_ch]   sum_e W exp(-sum_m L_m mu_m)            (sino_raw)
 *   logs     same shape or NULL                   -ln(counts / sum_e W)                   (sino_log)
 *   L_out    [view_count][n_rows][n_ch][K] or NULL  per-material path length [cm]
 * K <= 256 (uint8 labels); labels >= K are a caller error: the kernels index their accumulators with the label and do
 * not check it (the Python binding does, once per prepared phantom: dexct.ops.Projector.project).               */
int dexct_siddon_poly(const DexctScan *scan, const DexctVolume *vol, const void *workspace,
                      int32_t view_begin, int32_t view_count,
                      const void *mu_KxE, int32_t K, int32_t nE, const void *W_SxE, int32_t nSpec,
                      void *counts, void *logs, void *L_out, int32_t dtype, void *stream);

/* Detection only, from stored path lengths L [n_rays][K] (second spectrum of the same scan: main.py:109-120
 * calls get_sino twice on the same ct/phantom).  counts/logs [nSpec][n_rays].                               */
int dexct_poly_detect(const void *L, int64_t n_rays, const void *mu_KxE, int32_t K, int32_t nE,
                      const void *W_SxE, int32_t nSpec, void *counts, void *logs, int32_t dtype, void *stream);

/* Debug: ordered voxel crossing sequence of selected rays.  ray_ids [n_rays] = (view*n_rows + row)*n_ch + ch.
 * For ray t the first min(n_seg[t], max_seg) entries of vox/lab/len rows t are valid: linear voxel index
 * (k*ny + j)*nx + i, its label, and the intersection length [cm] (float or double per dtype).
 * vox, lab and len may all be NULL to obtain only the segment counts n_seg (the unit of the bytes-per-ray
 * roofline).                                                                                                */
int dexct_siddon_trace(const DexctScan *scan, const DexctVolume *vol, const void *workspace,
                       const int64_t *ray_ids, int32_t n_rays, int32_t max_seg,
                       int64_t *vox, uint8_t *lab, void *len, int32_t *n_seg, int32_t dtype, void *stream);

/* Debug: the steps of the PRODUCTION walk over the skip layouts (the routine dexct_siddon_poly inlines for phantoms
 * with labels < 128), recorded by a sink inside that routine.  For ray t the first min(n_steps[t], max_steps) entries
 * of rows t are valid, in order of increasing column index (the walk's order, whatever the ray's direction):
 *   col  [n_rays][max_steps]      first column of the step (index along the ray's major axis)
 *   run  [n_rays][max_steps]      0: one ordinary column, split into its pieces;  S > 0: a homogeneous stretch of S columns
 *   vox / lab / len [n_rays][max_steps][3]   the pieces in order along the major axis: linear voxel index
 *        (k*ny + j)*nx + i (-1: apron, outside the volume; for a run: the voxel it starts in), label, length [cm]
 *   meta [n_rays]   bit 0: x is the major axis, bit 1: the ray travels towards +major, bit 2: non-negative minor slope,
 *                   bit 3: the ray meets the volume
 *   L_out [n_rays][K]   per-material lengths [cm] summed by the walk (what the production kernel feeds its epilogue)
 * col .. len may all be NULL to obtain only n_steps, meta and L_out.  Needs 1 <= K <= 128.
 * The oracle's voxel sequence of the same ray (north_star: "voxel-crossing index sequences and per-ray material IDs must
 * match bit-exactly") is compared with these steps in tests/test_gpu_siddon.py.                                       */
int dexct_siddon_walk_trace(const DexctScan *scan, const DexctVolume *vol, const void *workspace,
                            const int64_t *ray_ids, int32_t n_rays, int32_t max_steps, int32_t K, int32_t *col, int32_t *run,
                            int64_t *vox, uint8_t *lab, void *len, int32_t *n_steps, int32_t *meta, void *L_out,
                            int32_t dtype, void *stream);

/* ---- Gauss-Newton two-basis decomposition: replaces optimize_sino_cpu (matdecomp.py:87-127) --------------- */

/* Per pixel p (P = views*channels, any order): n_iters full-Newton steps on the Poisson likelihood of the two
 * measurements g[0][p], g[1][p] under nu_m = sum_e i0[m][e] exp(-(a1 mu[0][e] + a2 mu[1][e])), from a = 1e-6.
 *   g_2xP   [2][P] measured counts;  i0_2xE [2][nE] (the reference tiles it over channels, matdecomp.py:151);
 *   mus_2xE [2][nE] basis attenuation;  a_out [P][2].
 * mask_thresh > 0 additionally zeroes both outputs of pixels with g[0][p] >= mask_thresh * max_p g[0][p]
 * (get_basismat_sinos, matdecomp.py:195-205).
 * scratch = dexct_gn_scratch_bytes(P, dtype) bytes of device memory: the air maximum and, in single precision, the list
 * of pixels whose iterates left the float range (photon-starved or metal-crossing rays: the reference's exponent clip,
 * matdecomp.py:116, is what keeps them finite in float64) -- those are redone in float64 inside the same call, so the
 * result is finite wherever the reference's is -- and the list of the pixels that are not air (with a mask the air pixels are
 * zeroed up front and only the others are decomposed).  32 + 8 P bytes in single precision, 16 in double.  May be NULL only
 * for DEXCT_F64 with mask_thresh <= 0.
 * scratch_has_max != 0: scratch already holds the maximum as written by dexct_gn_air_max (pixel-sharded runs reduce it
 * over ranks first); 0: it is computed here from g[0][0..P).                                                     */
size_t dexct_gn_scratch_bytes(int64_t P, int32_t dtype);
int dexct_gn_decompose(const void *g_2xP, const void *i0_2xE, const void *mus_2xE, int32_t nE, int64_t P,
                       int32_t n_iters, double mask_thresh, void *a_out_Px2, void *scratch,
                       int32_t scratch_has_max, int32_t dtype, void *stream);

/* optimize_sino_cpu as the reference wrote it (matdecomp.py:87-127), beyond what get_basismat_sinos feeds it: nMeas
 * measurements g [nMeas][P], nMats basis materials mus [nMats][nE] (1..4 each; :95,125 invert an nMats x nMats system
 * per pixel) and, with i0_per_bin != 0, a spectrum per detector bin i0 [nMeas][n_bins][nE] (:101-105,118) -- pixel p
 * then belongs to bin p % n_bins (P a multiple of n_bins); with i0_per_bin == 0, i0 is [nMeas][nE].  a_out [P][nMats].
 * Arithmetic in float64 with the reference's exponent clip whatever `dtype` (which names the type of g, i0, mus, a_out);
 * no air mask (get_basismat_sinos applies it for two materials only).                                              */
int dexct_gn_decompose_general(const void *g, const void *i0, const void *mus, int32_t nMeas, int32_t nMats, int32_t nE,
                               int64_t n_bins, int32_t i0_per_bin, int64_t P, int32_t n_iters, void *a_out, int32_t dtype,
                               void *stream);

/* Single-precision dexct_gn_decompose for callers that may not allocate device memory (code under CUDA graph capture,
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Arguments as for dexct_gn_decompose with dtype = DEXCT_F32 and no scratch: the scratch block (air-maximum
 * key, flagged-pixel count and list) lives in a file-scope device array of the library, 4 MiB + 256 B per GPU that loads
 * it, held until the process exits.  P <= DEXCT_GN_STATIC_MAX_PIXELS; larger calls return DEXCT_ECAPACITY with a message.
 * The call makes no cudaMalloc / cudaFree and calls no allocating library routine.  All calls share the array, which keeps
 * the last call's contents: what must start from zero is reset at the start of every call, and calls that overlap on one
 * GPU (two streams, two threads) race on it -- serialise them.  Runs the FP32-pipe kernel with its table row in shared
 * memory (not the tensor-core kernel, whose table staging buffer is allocated on first use), then the float64 pass over
 * flagged pixels: results agree with dexct_gn_decompose to single-precision rounding (1e-5 scale-relative).              */
#define DEXCT_GN_STATIC_MAX_PIXELS 1048576
int64_t dexct_gn_static_max_pixels(void);
int dexct_gn_decompose_static(const void *g_2xP, const void *i0_2xE, const void *mus_2xE, int32_t nE, int64_t P,
                              int32_t n_iters, double mask_thresh, void *a_out_Px2, void *stream);

/* max_p g1[p] (the first measurement, matdecomp.py:196) into scratch[0..8) as an order-preserving integer key:
 * int32 (DEXCT_F32) or int64 (DEXCT_F64), so that ranks holding different pixels can combine their values with an
 * integer MAX all-reduce before calling dexct_gn_decompose(..., scratch_has_max = 1).                          */
int dexct_gn_air_max(const void *g1, int64_t P, void *scratch, int32_t dtype, void *stream);

/* ---- fan-beam filtered back-projection: replaces get_recon (main.py:134,168) ------------------------------ */

/* Equiangular fan-beam FBP (Kak & Slaney) of the views [view_begin, view_begin + view_count) of a full-rotation scan,
 * as a gather over views: cosine pre-weight, ramp filter with a sinc window cut at
 * ramp * Nyquist (input/params.txt:35), 1/L^2-weighted back-projection with linear interpolation between channels.
 *   sino    [view_count][n_rows][n_ch] (dtype)   line integrals of THOSE views (a view shard starts at its own row 0)
 *   image   [n_rows][n_matrix][n_matrix] (dtype) one slice per detector row, pixel (j, i) centred at
 *           x = (i - (n-1)/2) fov/n, y = (j - (n-1)/2) fov/n -- the phantom's axes; units of the sinogram per cm
 *   accumulate != 0: add to the image instead of overwriting it (view chunks, or partial images that ranks holding
 *           different view ranges sum with an all-reduce)
 *   scratch dexct_fbp_workspace_bytes(scan, view_count, dtype) bytes of device memory (filtered lines + tables)
 * The whole square must lie inside the source circle (fov / sqrt(2) < SID).                                      */
size_t dexct_fbp_workspace_bytes(const DexctScan *scan, int32_t view_count, int32_t dtype);
int dexct_fbp_fan(const DexctScan *scan, const void *sino, int32_t view_begin, int32_t view_count,
                  int32_t n_matrix, double fov, double ramp, void *image, int32_t accumulate,
                  void *scratch, int32_t dtype, void *stream);

/* hu[i] = 1000 (image[i] - mu_water) / mu_water: the Hounsfield image get_recon returns beside the raw one
 * (main.py:134-140).  In place (hu == image) is allowed.                                                        */
int dexct_fbp_hu(const void *image, int64_t n, double mu_water, void *hu, int32_t dtype, void *stream);

/* ---- quantum noise and virtual monoenergetic images -------------------------- */

/* counts[i] <- one Poisson variate of mean counts[i], i in [0, n): the photon statistics of a scan whose spectra were
 * scaled to photons per pixel per view (main.py:68).  Generator: Philox4x32-10 keyed by `seed`, counter
 * (offset + i, block, stream_id) -- `offset` is the element's index in the WHOLE sinogram, so view-sharded ranks draw
 * exactly the numbers a single GPU would; `stream_id` separates the acquisitions of one run (one per spectrum).
 * Sampler: CDF inversion below mean 10, Hoermann's PTRS rejection above (exact at every mean).
 * logs (same length) may be NULL; otherwise logs[i] <- -ln(max(counts[i], floor_counts) / air_counts).          */
int dexct_poisson_noise(void *counts, void *logs, int64_t n, uint64_t seed, uint32_t stream_id, uint64_t offset,
                        double air_counts, double floor_counts, int32_t dtype, void *stream);

/* Virtual monoenergetic image from two basis-material images (make_vmi, plots.py:136-144):
 * out[i] = u1 m1[i] + u2 m2[i], and with hu != 0 then 1000 (out[i] - u_water) / u_water; float32 out like the
 * reference's .astype(np.float32).  u1, u2, u_water: mass attenuation of the bases and of water at the energy.   */
int dexct_vmi(const void *m1, const void *m2, int64_t n, double u1, double u2, double u_water, int32_t hu,
              float *out, int32_t dtype, void *stream);

/* ---- peer memory: view-sharded scans on one NVLink/NVSwitch node -------------------- *
 * One process per GPU.  Each rank allocates its copy of a sinogram buffer with dexct_peer_alloc, exports it, and opens
 * the handles of all other ranks (exchanged by the host program, e.g. torch.distributed.all_gather_object); results
 * are then STORED into every rank's copy by the producing kernels -- dexct_gn_decompose_mirrored's epilogue, or
 * dexct_peer_push for a finished buffer -- in place of an all-gather.  Synchronisation uses a small control block per
 * rank (dexct_peer_ctrl_bytes, allocated and exchanged the same way, zero-initialised by dexct_peer_alloc):
 * "ctrl_ptrs_host" below is the host array of the world_size control-block addresses AS MAPPED IN THE CALLING PROCESS,
 * own block at index `rank`.  `epoch` must be 1, 2, 3, ... over successive calls, the same on every rank (separate
 * sequences for dexct_peer_barrier and dexct_peer_max_key).  A wait longer than timeout_s gives up and is reported
 * by dexct_peer_status (0 = healthy, r + 1 = rank r never arrived).                                               */
size_t dexct_peer_ctrl_bytes(void);
int dexct_peer_alloc(size_t bytes, void **ptr_host);          /* cudaMalloc + zero fill (exportable allocation) */
int dexct_peer_free(void *ptr);
int dexct_peer_export(const void *ptr, void *handle64_host);   /* DEXCT_PEER_HANDLE_BYTES bytes out */
int dexct_peer_open(const void *handle64_host, void **ptr_host);
int dexct_peer_close(void *ptr);
/* All ranks' earlier work on `stream` (including stores into peer memory) is complete and visible when the barrier
 * kernel of every rank has finished.                                                                             */
int dexct_peer_barrier(void *const *ctrl_ptrs_host, int32_t rank, int32_t world_size, uint64_t epoch, double timeout_s,
                       void *stream);
/* Integer MAX all-reduce of the air-maximum key written by dexct_gn_air_max (key = its scratch, int32 for DEXCT_F32,
 * int64 for DEXCT_F64), in place; then dexct_gn_decompose(..., scratch_has_max = 1).                               */
int dexct_peer_max_key(void *const *ctrl_ptrs_host, int32_t rank, int32_t world_size, uint64_t epoch, void *key,
                       int32_t dtype, double timeout_s, void *stream);
int dexct_peer_status(const void *ctrl_local, int32_t *status_host, void *stream);   /* synchronises */
/* Copy src[0, bytes) to dst_ptrs_host[d] for every d != skip (skip = -1: all); pointers 16-byte aligned, bytes a
 * multiple of 16.  max_ctas bounds the grid (0 = one CTA per SM) so the copy can share the GPU with a compute kernel. */
int dexct_peer_push(const void *src, size_t bytes, void *const *dst_ptrs_host, int32_t n_dst, int32_t skip,
                    int32_t max_ctas, void *stream);
/* The same transfer on the copy engines (one DMA per destination; no alignment requirement, no SM time). */
int dexct_peer_push_dma(const void *src, size_t bytes, void *const *dst_ptrs_host, int32_t n_dst, int32_t skip, void *stream);
/* dexct_gn_decompose whose epilogue also stores each pixel's result to the same index of n_mirrors further arrays
 * (the other ranks' copies of the basis sinogram, offset to this rank's pixel range).                             */
int dexct_gn_decompose_mirrored(const void *g_2xP, const void *i0_2xE, const void *mus_2xE, int32_t nE, int64_t P,
                                int32_t n_iters, double mask_thresh, void *a_out_Px2, void *const *mirrors_host,
                                int32_t n_mirrors, void *scratch, int32_t scratch_has_max, int32_t dtype, void *stream);
/* ... and that also carries a finished buffer to the peers: the kernel's threads copy push_src[0, push_bytes) into each
 * push_dst_host[d] (the other ranks' copies of the count/log shard, 16-byte aligned) one slice per Newton iteration, so
 * the transfer is spread over all SMs and over the kernel's run time -- one kernel computes and communicates.       */
int dexct_gn_decompose_fused_push(const void *g_2xP, const void *i0_2xE, const void *mus_2xE, int32_t nE, int64_t P,
                                  int32_t n_iters, double mask_thresh, void *a_out_Px2, void *const *mirrors_host,
                                  int32_t n_mirrors, const void *push_src, size_t push_bytes, void *const *push_dst_host,
                                  int32_t n_push_dst, void *scratch, int32_t scratch_has_max, int32_t dtype, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* DEXCT_H */
