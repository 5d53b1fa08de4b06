/* libvistaf — C ABI of the B200-native VISTAF multimodal sensing hot path.
 *
 * The reference (rimelq/VISTAF-RoboSkin) has no FFI of its own: its boundary for this path is the
 * Python module API consumed by Code/multimodal_sensor.py (shape_ftp.main :1428,
 * force_sensor.depth_map_to_volume_cm3 :93 / predict_force_from_volume :149,
 * temperature_sensor.main :749).  The Python modules of the same names in this repo bind the entry
 * points below through ctypes (INTEGRATION.md shows the stub).  Conventions:
 *   - plain pointers and sizes only; the caller owns every input / output / workspace buffer,
 *     the library owns only the opaque plan handles;
 *   - every call takes DEVICE pointers and a cudaStream_t (as void*) and is asynchronous; host staging is the
 *     caller's (PyTorch pinned buffers in the Python modules, multimodal_sensor.MultimodalPlan);
 *   - a plan is read-only during `*_batch` calls, so several threads / streams may run batches on one plan with
 *     their own workspaces; vistaf_ftp_set_reference writes the plan and must not overlap `*_finish` / `_batch`
 *     calls that are still being ENQUEUED (stream order takes care of the device side);
 *   - return 0 on success; vistaf_last_error() describes the last failure of the calling THREAD;
 *   - repeated batch calls with identical arguments are replayed from a CUDA graph recorded on the second call
 *     (vistaf_set_graphs(0) or VISTAF_GRAPHS=0 turns that off);
 *   - per-frame soft failures (empty ROI, no carrier peak, phase residues ...) are reported in the
 *     `status` word of the per-frame record, never through the return code.
 */
#ifndef VISTAF_H
#define VISTAF_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define VISTAF_MAX_TERMS 35
#define VISTAF_MAX_ISO 64

/* per-frame status bits */
#define VISTAF_ST_EMPTY_ROI        0x001  /* temperature_sensor.py:443  / shape_ftp.py:1677 */
#define VISTAF_ST_NO_PEAK          0x002  /* temperature_sensor.py:457 */
#define VISTAF_ST_SATURATED        0x004  /* saturated pixels present: wide map quantised + Telea-filled (T10, temperature_sensor.py:546) */
#define VISTAF_ST_ROTATED_CARRIER  0x008  /* informational: carrier angle != 0, the oriented blur took the warpAffine path (T12, temperature_sensor.py:705) */
#define VISTAF_ST_RESIDUES         0x010  /* informational: phase residues inside the reliable mask, the unwrap ran as the reference's sequential quality-guided flood (F10) */
#define VISTAF_ST_FEW_RELIABLE     0x020  /* < 200 reliable pixels: polynomial detrend skipped (shape_ftp.py:1103) */
#define VISTAF_ST_SIGN_FLIPPED     0x040  /* auto sign flip taken (shape_ftp.py:1759) */
#define VISTAF_ST_CONTACT_P95      0x080  /* contact threshold fell back to p95 / p98 (shape_ftp.py:1733-1738) */
#define VISTAF_ST_BG_FALLBACK      0x100  /* background < 15 % of reliable: detrend on reliable (shape_ftp.py:1744) */
#define VISTAF_ST_SPLIT_RELIABLE   0x400  /* informational: reliable mask not connected after the erosion, the flood does not reach every component (as in the reference) */
#define VISTAF_ST_HOLES_FILLED     0x800  /* internal holes of the reliable mask were Telea-filled (F14, shape_ftp.py:1776-1801) */
#define VISTAF_ST_TELEA_OVERFLOW   0x8000 /* a Telea cluster did not fit the march workspace (hole area > half the frame): those pixels are not filled */
#define VISTAF_ST_RERUN_ALONE      0x200  /* more than an eighth of one chunk's frames are saturated (T10) or have a tilted carrier (T12): this frame
                                             found no slot and skipped that step -- run it again in a batch of its own (the Python
                                             modules do: TempPlan.redo_flagged, MultimodalPlan.records_host) */
#define VISTAF_ST_SAT_UNFILLED     VISTAF_ST_RERUN_ALONE

/* StandardScaler -> PolynomialFeatures -> linear model [-> isotonic], temperature_sensor.py:236 */
typedef struct {
  int32_t n_features;                 /* 3 (L,a,b) or 4 (L,a,b,gray) */
  int32_t n_terms;
  int32_t n_iso;                      /* 0 = no isotonic calibrator */
  int32_t pad_;
  double mean[4], scale[4];
  int32_t powers[VISTAF_MAX_TERMS][4];
  double coef[VISTAF_MAX_TERMS];
  double intercept;
  double iso_x[VISTAF_MAX_ISO], iso_y[VISTAF_MAX_ISO];
} VistafPolyModel;

/* temperature_sensor.py:34-98 */
typedef struct {
  int32_t height, width;
  double roi_cx, roi_cy, roi_r;       /* outer circle, float64 (temperature_sensor.py:157) */
  float color_t_min, color_t_max, guard_band, switch_margin, final_t_min, final_t_max;
  int32_t band_radius, dc_exclusion;
  float peak_max_dy_frac, illum_sigma;
  int32_t sat_thresh, sat_dilate;
  int32_t close_kx, close_ky, open_kx, open_ky;
  float chroma_min;
  int32_t support_dilate;
  float sigma_across, sigma_along;
  VistafPolyModel color_model, wide_model;
  int32_t lab_gamma[256];             /* OpenCV 8-bit BGR2Lab tables (tools/make_lab_lut.py) */
  int32_t lab_cbrt[3072];
} VistafTempParams;

typedef struct {
  float mean_C, median_C, std_C, min_C, max_C;   /* multimodal_sensor.py:555-567 */
  int32_t valid_pixels;
  int32_t peak_x, peak_y;                        /* carrier peak, fftshift-ed pixel coordinates */
  float phi0, mean_gray_a, mean_gray_b, median_gray;
  int32_t sat_pixels, roi_eff_pixels, dark_pixels, color_ok_pixels;
  int32_t status;
  int32_t pad_[3];
} VistafTempRecord;                               /* 80 bytes */

typedef struct {
  float* fused;        /* [B][H][W]  temperature_map_fused  (NaN outside ROI) */
  float* final_map;    /* [B][H][W]  temperature_map_final */
  uint8_t* sat;        /* [B][H][W]  0/1 masks; any may be NULL */
  uint8_t* dark;
  uint8_t* light;
  uint8_t* color_support;
  uint8_t* color_ok;
  uint8_t* source;     /* 0 wide / 128 blend / 255 colour (debug_fused_source.png) */
  VistafTempRecord* records; /* [B] */
  int32_t background_valid;  /* != 0: every map / mask plane above already holds its background (NaN / 0) outside the ROI
                                bounding box -- e.g. the same buffers were filled by an earlier call of this plan -- so the
                                ~76 % of each plane that lies outside the box is not written again */
  int32_t pad_;
} VistafTempOutputs;

typedef struct VistafTempPlan VistafTempPlan;

const char* vistaf_last_error(void);
int vistaf_version(void);

/* Stage timers (measurement only; off by default).  When enabled every pipeline stage boundary records a CUDA
 * event on the launch stream; vistaf_profile_read synchronises and returns the accumulated milliseconds, call
 * count and kernel-launch count per stage since the previous read.  vistaf_launch_count() = kernels this
 * process has launched through the library so far (always counted). */
typedef struct {
  char name[48];
  double ms;
  int32_t calls, launches;
} VistafStageTime;
int vistaf_profile_enable(int on);
/* CUDA-graph replay of repeated batch calls (default on; off while the stage timers run) and how many calls were replays */
int vistaf_set_graphs(int on);
long long vistaf_graph_replays(void);
int vistaf_profile_read(VistafStageTime* out, int max_stages, int* n_out);
long long vistaf_launch_count(void);

/* Temperature path (replaces temperature_sensor.main :749 numeric body, batched). */
int vistaf_temp_plan_create(const VistafTempParams* params, VistafTempPlan** plan);
void vistaf_temp_plan_destroy(VistafTempPlan* plan);
/* bytes of device workspace needed to process `chunk` frames at a time */
size_t vistaf_temp_workspace_bytes(const VistafTempPlan* plan, int chunk);
/* frames_bgr: DEVICE uint8 [batch][H][W][3]; outputs: DEVICE pointers. */
int vistaf_temp_batch(VistafTempPlan* plan, const uint8_t* frames_bgr, int batch, const VistafTempOutputs* out,
                      void* workspace, size_t workspace_bytes, void* cuda_stream);
/* copies the plan's ROI mask (uint8 [H][W]) to a DEVICE buffer */
int vistaf_temp_roi_mask(const VistafTempPlan* plan, uint8_t* roi_dev, void* cuda_stream);
/* debugging / parity: name -> byte offset and size of an intermediate plane of frame 0 in the workspace */
int vistaf_temp_workspace_plane(const VistafTempPlan* plan, int chunk, const char* name, size_t* offset, size_t* bytes);

/* ------------------------------------------------------------------ FTP shape + force path */
/* calibrated curve families shared by shape_ftp.model_predict (:682) and force_sensor (:129-140) */
#define VISTAF_MODEL_LINEAR0 0
#define VISTAF_MODEL_LINEAR 1
#define VISTAF_MODEL_POLY2 2
#define VISTAF_MODEL_SAT_EXP 3
#define VISTAF_MODEL_GROWTH 4
#define VISTAF_MODEL_HINGE_SAT 5

typedef struct {
  int32_t type;          /* VISTAF_MODEL_* */
  int32_t pad_;
  double p[3];           /* (a) | (a,b) | (c2,c1,c0) | (a,b) | (a,b) | (a,b,c) */
} VistafCurveModel;

/* shape_ftp.py:16-218 (constants read by the numeric path) on an (crop_h x crop_w) gray crop */
typedef struct {
  int32_t crop_h, crop_w;
  int32_t cx, cy, r;                 /* ROI circle in crop coordinates (shape_ftp.py:1513-1516) */
  int32_t roi_erode_px, apod_taper;
  int32_t patch_half, dc_exclusion, n_peaks, fft_pad;
  float peak_max_dy_frac;
  float illum_sigma, pre_blur_sigma, quality_sigma, smooth_sigma, unrel_sigma;
  float amp_percentile;
  int32_t close_kernel, edge_margin;
  float bad_int_pct, bad_grad_pct;
  int32_t bad_dilate, bad_inpaint_radius;
  float core_pct, contact_pct;
  int32_t dilate_kernel, dilate_iters;
  float min_contact_frac, max_contact_frac;
  float frontier_band;
  int32_t use_negated_height;        /* calibration JSON use_negated_height_for_fit */
  double blob_min_peak_mm, blob_rel_frac;   /* shape_ftp.py:1215 (python floats: 0.1, 1/3) */
  VistafCurveModel height_model;     /* phase -> mm   (shape_ftp.py:672-705) */
  VistafCurveModel force_model;      /* cm^3 -> N     (force_sensor.py:149-167) */
  double grating_pitch_mm, depth_eps_mm;   /* force_sensor.py:32-33 */
  /* F14, shape_ftp.py:139-145: FILL_INTERNAL_HOLES_IN_RELIABLE, HOLE_NEIGHBORHOOD_PX, HOLE_KNOWN_FRACTION,
   * HOLE_MIN_DIST_FROM_RELIABLE_EDGE_PX, INPAINT_RADIUS */
  int32_t fill_internal_holes, hole_neighborhood;
  float hole_known_frac, hole_min_dist;
  int32_t hole_inpaint_radius, pad2_;
} VistafFtpParams;

typedef struct {
  double force_N, volume_cm3, contact_area_mm2, max_depth_mm;   /* multimodal_sensor.py:412-419 */
  double mm_per_px, period_px;
  float amp_thr, contact_thr, bg_median, core_median;
  int32_t reliable_px, output_reliable_px, contact_dilated_px, contact_kept_px, bad_px, residues;
  int32_t seed_x, seed_y;
  int32_t status;
  int32_t pad_;
} VistafFtpRecord;                                   /* 104 bytes */

typedef struct {
  float* height_mm;          /* [B][h][w] depth map in mm, NaN outside the ROI (height_map_mm_crop) */
  uint8_t* reliable;         /* [B][h][w] 0/1; any mask may be NULL */
  uint8_t* output_reliable;
  uint8_t* contact_dilated;
  uint8_t* contact_kept;
  VistafFtpRecord* records;  /* [B] */
  uint8_t* hole_candidates;  /* [B][h][w] 0/1, may be NULL (shape_ftp.py:1776-1784) */
} VistafFtpOutputs;

typedef struct {
  float peak_x, peak_y;      /* refined carrier peak (fftshift-ed bins of the padded FFT) */
  int32_t fft_h, fft_w;
  float kx, ky, period_px;
  int32_t bad_px;
} VistafFtpReferenceInfo;

typedef struct VistafFtpPlan VistafFtpPlan;

int vistaf_ftp_plan_create(const VistafFtpParams* params, VistafFtpPlan** plan);
void vistaf_ftp_plan_destroy(VistafFtpPlan* plan);
size_t vistaf_ftp_workspace_bytes(const VistafFtpPlan* plan, int chunk);
/* crop + cv2 BGR2GRAY of full frames (shape_ftp.py:1508-1511): frames DEVICE uint8 [B][H][W][3] -> gray DEVICE uint8 [B][h][w] */
int vistaf_crop_gray(const uint8_t* frames_bgr, int batch, int H, int W, int x1, int y1, int h, int w, uint8_t* gray,
                     void* cuda_stream);
/* reference-frame half of shape_ftp.py:1632-1639 (dense FFT, carrier search, demodulation); cached in the plan */
int vistaf_ftp_set_reference(VistafFtpPlan* plan, const uint8_t* ref_gray, void* workspace, size_t workspace_bytes,
                             VistafFtpReferenceInfo* info, void* cuda_stream);
/* deformed frames: gray DEVICE uint8 [B][h][w] (already aligned); shape_ftp.py:1641-1865 + force_sensor.py:93-167 */
int vistaf_ftp_batch(VistafFtpPlan* plan, const uint8_t* def_gray, int batch, const VistafFtpOutputs* out, void* workspace,
                     size_t workspace_bytes, void* cuda_stream);
/* vistaf_ftp_batch for ONE chunk in two calls on the same workspace (batch must fit it whole): _begin runs the stages
 * that do not read the reference frame (shape_ftp.py:625-666, :821-848: bad pixels, Telea fill, illumination
 * normalisation, apodisation, median) and may be issued before vistaf_ftp_set_reference; _finish runs the rest. */
int vistaf_ftp_batch_begin(VistafFtpPlan* plan, const uint8_t* def_gray, int batch, void* workspace, size_t workspace_bytes,
                           void* cuda_stream);
int vistaf_ftp_batch_finish(VistafFtpPlan* plan, const uint8_t* def_gray, int batch, const VistafFtpOutputs* out,
                            void* workspace, size_t workspace_bytes, void* cuda_stream);
/* F14 on its own (shape_ftp.py:1153-1203, 1776-1801): height [B][h][w] in/out with NaN where no height is known,
 * reliable in; output_reliable and hole_candidates out (DEVICE uint8 [B][h][w]).  workspace as for `batch` frames. */
int vistaf_ftp_fill_holes(VistafFtpPlan* plan, float* height, const uint8_t* reliable, int batch, uint8_t* output_reliable,
                          uint8_t* hole_candidates, void* workspace, size_t workspace_bytes, void* cuda_stream);
/* F10 on its own (shape_ftp.py:1043-1080): wrapped / quality float [B][h][w], reliable uint8 -> unwrapped float (NaN where
 * the flood does not reach); residues[b] = residue count.  mode 0: as the chain does (parallel wrap-count field, the
 * sequential flood only for frames with residues), 1: sequential flood on every frame. */
int vistaf_ftp_unwrap(VistafFtpPlan* plan, const float* wrapped, const float* quality, const uint8_t* reliable, int batch,
                      float* unwrapped, int32_t* residues, int mode, void* workspace, size_t workspace_bytes, void* cuda_stream);
/* bit 0: run the sequential flood unwrap on every frame of later batches (parity tests) */
int vistaf_ftp_set_debug(VistafFtpPlan* plan, int flags);
int vistaf_ftp_masks(const VistafFtpPlan* plan, uint8_t* roi_dev, uint8_t* circ_dev, float* apo_dev, void* cuda_stream);
int vistaf_ftp_workspace_plane(const VistafFtpPlan* plan, int chunk, const char* name, size_t* offset, size_t* bytes);
/* force_sensor.depth_map_to_volume_cm3 (:93-123) on DEVICE maps: out3 = DEVICE float64 [B][3] (volume_cm3, area_mm2, max_depth_mm) */
int vistaf_depth_to_volume(const float* height_mm, const uint8_t* roi_or_null, int batch, long long npix, double mm_per_px,
                           double depth_eps_mm, double* out3, void* cuda_stream);
/* vistaf_depth_to_volume WITHOUT any allocation.  No counterpart in the reference: added
   for callers that cannot call cudaMalloc (stream capture, tight serving loops, harnesses that forbid it).
   Same arguments, same kernels and same results as vistaf_depth_to_volume; the per-frame accumulators live in a
   file-scope `static __device__` array of VISTAF_VOL_STATIC_MAX_BATCH entries (160 KiB, held by every GPU that loads
   the library until exit) instead of a cudaMallocAsync block.  batch > VISTAF_VOL_STATIC_MAX_BATCH is rejected with an
   error (split the call).  No cudaMalloc / cudaFree / allocating library routine is called.  Every call on a device
   shares the array (cleared on `cuda_stream` at the start of each call): calls that overlap on the GPU -- different
   streams with no ordering between them -- race on it; calls on one stream, or ordered by events, are safe. */
#define VISTAF_VOL_STATIC_MAX_BATCH 4096
int vistaf_depth_to_volume_static(const float* height_mm, const uint8_t* roi_or_null, int batch, long long npix, double mm_per_px,
                                  double depth_eps_mm, double* out3, void* cuda_stream);
int vistaf_depth_to_volume_static_max_batch(void);
/* host-side scalar: force_sensor.predict_force_from_volume (:149-167); returns NaN for an unknown type */
double vistaf_curve_predict(const VistafCurveModel* model, double x);

#ifdef __cplusplus
}
#endif
#endif
