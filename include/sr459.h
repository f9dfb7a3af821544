/* sr459.h -- C ABI of the B200-native SAA / SAA+IBP super-resolution engine.
 *
 * This is the drop-in boundary for the hot path of the reference's four
 * run_sr.py scripts (benedikthoward/ENPH459-Super-Resolution).  The reference
 * has no FFI layer of its own -- its "operator API" is a set of module-level
 * Python functions -- so every entry point below names the reference function
 * (file:line in mono_cal_target/run_sr.py unless noted) it replaces.  The
 * Python mirror of that surface (enph459-super-resolution_b200/ops.py) binds
 * these symbols with ctypes; INTEGRATION.md shows the stub a maintainer of the
 * reference would add.
 *
 * Conventions
 *   - return 0 on success, a negative SR459_E* code otherwise; never throws.
 *     sr459_last_error() returns a thread-local human-readable message.
 *   - every image pointer named d_* is CALLER-OWNED DEVICE memory: float32,
 *     row-major, pitch == width (tensor.data_ptr() of a contiguous CUDA
 *     tensor).  Pointers named h_* are host memory.
 *   - stream is a cudaStream_t passed as void* (0 = legacy default stream).
 *     Functions enqueue work and return without synchronising unless the
 *     description says otherwise.
 *   - a plan owns its scratch buffers and tables; one plan must not be used
 *     from two host threads at once; distinct plans are independent.
 *   - there is no CPU fallback anywhere behind this ABI.
 */
#ifndef SR459_H
#define SR459_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)   /* the library is built with -fvisibility=hidden */
#endif

#define SR459_OK            0
#define SR459_EINVAL       -1   /* bad argument (shape mismatch, null pointer, ...) */
#define SR459_ECUDA        -2   /* a CUDA runtime call or kernel launch failed      */
#define SR459_ENOMEM       -3
#define SR459_EUNSUPPORTED -4

typedef struct sr459_plan sr459_plan;

/* ---- library ---------------------------------------------------------- */
int         sr459_version(void);                 /* 100*major + minor */
const char* sr459_last_error(void);
/* hash of the sources this binary was built from (build.py: source_stamp());
 * the Python binding refuses a library whose stamp differs from the sources on disk. */
const char* sr459_source_stamp(void);
/* number of CUDA kernels this library has launched since load (per process);
 * bench.py reports the delta over its timed region as gpu_launches.          */
uint64_t    sr459_launch_count(void);

/* Debugging aid (alloc.cu): pad every later device allocation of the library with `bytes` of a fixed pattern on
 * both sides (0 = off; also SR459_GUARD=<bytes> in the environment at load time).  sr459_debug_check_guards()
 * synchronises, reads the bands back and returns how many allocations had theirs overwritten (0 = clean). */
int sr459_debug_guard_bands(int bytes);
int sr459_debug_check_guards(void);
int sr459_debug_live_allocations(void);
int sr459_debug_poke_guard(int restore);   /* checker self-test: damage / repair one guard byte */

/* ---- PSF construction ------------------------------------------------- */
/* make_gaussian_psf(size, sigma)                     run_sr.py:104-111.
 * out must hold (2*(size/2)+1)^2 doubles (the reference builds mgrid[-hw:hw+1]). */
int sr459_psf_gaussian(int size, double sigma, double* h_out);

/* load_measured_psf(psf_dir)                         run_sr.py:114-152,
 * with the PNG decode done by the caller.  d_imgs: n_img images, uint8
 * (is_f32==0) or float32 (is_f32==1), each H x W, contiguous one after the
 * other.  For every image the first-maximum argmax is found on the device
 * (run_sr.py:127), images whose peak is within halfwidth+6 of an edge are
 * skipped (:128-131), the (2*(halfwidth+6)+1)^2 patches are averaged (:137), the
 * centre (2*halfwidth+1)^2 is cut (:141-142), the mean of the four 3x3 corner
 * blocks subtracted (:144-148), clipped at 0 and normalised (:149-150).
 * h_peaks (2*n_img ints, may be NULL) receives (row, col) per image, or (-1,-1)
 * for skipped ones; *h_n_used the number of patches.  Synchronises the stream. */
int sr459_psf_measured(const void* d_imgs, int is_f32, int n_img, int H, int W,
                       int halfwidth, double* h_out, int* h_peaks, int* h_n_used,
                       void* stream);

/* ---- frame ingest / egress (load_gray, extract_red, rep-mean, uint8 cast) -- */
/* load_gray for a mono PNG: uint8 -> float32           run_sr.py:73-75        */
int sr459_u8_to_f32(const uint8_t* d_in, int64_t n, float* d_out, void* stream);
/* extract_red: out[i,j] = in[2i+row_off, 2j+col_off]  rgb_cal_target/run_sr.py:73-75
 * in is H x W uint8 (is_f32==0) or float32; out is ceil((H-row_off)/2) x ceil((W-col_off)/2) f32 */
int sr459_extract_red(const void* d_in, int is_f32, int H, int W, int row_off, int col_off,
                      float* d_out, void* stream);
/* np.mean(frames, axis=0) over n_frames device images  run_sr.py:274, rgb_cal_target/run_sr.py:107-108 */
int sr459_mean_frames(const float* const* h_d_frames, int n_frames, int64_t n_px,
                      float* d_out, void* stream);
/* One-pass ingest of a frame set from its uint8 files: load_gray (run_sr.py:73-75), optional Bayer
 * lattice gather in[step*i+row_off, step*j+col_off] (extract_red, rgb_cal_target/run_sr.py:73-75;
 * step 1 = every pixel), mean over the `reps` repetitions of each of the n_groups frames
 * (rgb_cal_target/run_sr.py:107-108) and np.mean over the frames (run_sr.py:274), in float64 and in
 * numpy's order of operations.  h_d_imgs: n_groups*reps device images (H x W uint8), group-major.
 * h_d_frames: n_groups outputs (float32, ceil((H-row_off)/step) x ceil((W-col_off)/step)).
 * d_mean_f32 / d_mean_u8 (either may be NULL): the frame mean and its clip+truncate uint8 cast (:308). */
int sr459_ingest_groups(const uint8_t* const* h_d_imgs, int n_groups, int reps, int H, int W,
                        int step, int row_off, int col_off, float* const* h_d_frames,
                        float* d_mean_f32, uint8_t* d_mean_u8, void* stream);
/* The same with a repetition count per frame (h_reps[n_groups], any counts >= 1: the reference's
 * np.stack(reps).mean(axis=0) per corner accepts unequal counts); images are frame-major. */
int sr459_ingest_groups_v(const uint8_t* const* h_d_imgs, int n_groups, const int* h_reps, int H, int W,
                          int step, int row_off, int col_off, float* const* h_d_frames,
                          float* d_mean_f32, uint8_t* d_mean_u8, void* stream);
/* np.clip(img,0,255).astype(np.uint8)  (truncation)    run_sr.py:303          */
int sr459_f32_to_u8(const float* d_in, int64_t n, uint8_t* d_out, void* stream);

/* ---- stand-alone operators (the reference's module-level functions) ---- */
/* blur(img, kernel) = fftconvolve(img, kernel, 'same')  run_sr.py:157-158.
 * h_psf is kh x kw doubles; zero-padded true convolution.                    */
int sr459_blur(const float* d_img, int H, int W, const double* h_psf, int kh, int kw,
               float* d_out, void* stream);
/* scipy.ndimage.zoom(a, factor, order=3), defaults      call sites run_sr.py:185, :279.
 * out is round(h*factor) x round(w*factor).                                   */
int sr459_zoom(const float* d_in, int h, int w, int factor, float* d_out, void* stream);
/* scipy.ndimage.zoom(img, zoom, order=order) for any output size H x W = round(h * zoom_y) x round(w * zoom_x) (the caller
 * rounds as scipy does) and spline order 0, 1 or 3 -- what shift_and_add(lr_list, shifts, factor, order) hands to zoom,
 * run_sr.py:185; other arguments scipy's defaults (mode='constant', grid_mode=False).  Synchronises the stream.        */
int sr459_zoom_to(const float* d_in, int h, int w, int H, int W, int order, float* d_out, void* stream);
/* One axis of back_project's  shift(zero_insert(err, factor), s, order=3, mode='nearest')  (run_sr.py:170-176) as a
 * banded matrix: out[o] = sum_{a<20} h_w[o*20 + a] * err[h_i0[o] + a], o = 0 .. factor*n_lr-1; `shift` is the HR
 * shift scipy receives (-factor*d).  Host only (no device work): the tables the FIR back-projection kernel uses.
 * Returns SR459_EUNSUPPORTED when n_lr < 20.                                                                  */
int sr459_fir_bp_axis(int n_lr, int factor, double shift, int* h_i0, float* h_w);
/* scipy.ndimage.shift(a, (sy,sx), order=3, mode='nearest')  call sites :163,:176,:186 */
int sr459_shift(const float* d_in, int H, int W, double sy, double sx, float* d_out, void* stream);
/* forward_model(hr, kernel, shift_yx, factor)           run_sr.py:161-165.
 * out is ceil(H/factor) x ceil(W/factor).                                     */
int sr459_forward_model(const float* d_hr, int H, int W, const double* h_psf, int kh, int kw,
                        double dy, double dx, int factor, float* d_out_lr, void* stream);
/* back_project(err, kernel, shift_yx, factor, hr_shape)  run_sr.py:168-178    */
int sr459_back_project(const float* d_err, int h, int w, const double* h_psf, int kh, int kw,
                       double dy, double dx, int factor, int H, int W, float* d_out_hr,
                       void* stream);

/* ---- plans: Native-2x, SAA, SAA+IBP on a fixed geometry ---------------- */
/* Geometry: n_frames LR frames of h x w, HR = (factor*h) x (factor*w).
 * h_shifts_yx: 2*n_frames doubles, (dy,dx) per frame in LR pixels (the
 * reference's shifts_lr).  h_psf: kh x kw doubles.  step / clip_lo / clip_hi are
 * ibp()'s step and the literals 0 and 255 of run_sr.py:205.                   */
int sr459_plan_create(sr459_plan** out, int device, int n_frames, int h, int w, int factor,
                      const double* h_shifts_yx, const double* h_psf, int kh, int kw,
                      double step, double clip_lo, double clip_hi);
void sr459_plan_destroy(sr459_plan* plan);

/* 1 if every factor*shift is an integer (closed-form fused IBP kernels are
 * used), 0 if the cubic-spline warp path is used.                            */
int sr459_plan_is_integer(const sr459_plan* plan);
/* Force a code path, for tests: 0 = automatic, 1 = generic stage-by-stage path,
 * 2 = dense fused integer kernel, 3 = polyphase fused kernel (any 7x7 PSF),
 * 4 = TMA-staged separable-PSF fused kernel, warp-synchronous form (row bands
 * of several lattices), 7 = the register-streaming kernels (one warp per strip,
 * TMA row ring; rank-1 PSF; one lattice: k_ibp_stream, two to four lattices:
 * k_ibp_stream_ml; the default when they apply), 8 = the lattice kernels for PSFs
 * up to 31 x 31.  (5 and 6, two superseded separable kernels, were removed.)
 * SR459_EUNSUPPORTED if the plan's geometry cannot run the requested path.    */
int sr459_plan_set_path(sr459_plan* plan, int path);
/* Bytes of device scratch the plan holds.                                     */
int64_t sr459_plan_scratch_bytes(const sr459_plan* plan);

/* Native-2x: ndi_zoom(mean_lr, factor, order=3)         run_sr.py:279         */
int sr459_native2x(sr459_plan* plan, const float* d_mean_lr, float* d_hr, void* stream);
/* shift_and_add(lr_list, shifts, factor, order=3)       run_sr.py:181-187.
 * h_d_lr: host array of n_frames device pointers.                             */
int sr459_saa(sr459_plan* plan, const float* const* h_d_lr, float* d_hr, void* stream);
/* ibp(lr_list, shifts, kernel, hr_init, factor, n_iter, step)  run_sr.py:190-209.
 * d_hr_inout holds hr_init on entry and the result on exit (the Python mirror
 * copies hr_init first, as the reference does at :191).  If h_errors != NULL it
 * receives the n_iter per-iteration MSE values (errors list, :206) and the call
 * synchronises the stream; with NULL nothing is synchronised and the values
 * can be fetched later with sr459_ibp_errors().                               */
int sr459_ibp(sr459_plan* plan, const float* const* h_d_lr, float* d_hr_inout, int n_iter,
              double* h_errors, void* stream);
int sr459_ibp_errors(sr459_plan* plan, int n_iter, double* h_errors, void* stream);
/* SAA followed by IBP in one call (process_session, run_sr.py:284-292):
 * d_hr receives the SAA+IBP result; d_saa (may be NULL) a copy of the SAA image. */
int sr459_saa_ibp(sr459_plan* plan, const float* const* h_d_lr, float* d_hr, float* d_saa,
                  int n_iter, double* h_errors, void* stream);

/* ---- row bands: ONE large HR image split across GPUs (BASELINE config 4) ---- */
/* The reference has no multi-device path; a band plan runs shift_and_add (run_sr.py:181-187) and ibp()'s
 * iteration (run_sr.py:197-205) on the HR rows [2a, 2b) that rank `rank` of `nranks` owns (LR rows [a, b) = an
 * even split of h), keeping `halo` further hr rows on each side: 2*reach+2 for integer HR shifts (fused kernels),
 * 2*(reach+19) rounded up to 16 for fractional shifts (rgb_cal_target/run_sr.py:88-96; stage-by-stage iteration on
 * the local image).  Factor 2 only.  The band's two hr buffers and its flag words live in one device block ("arena")
 * that the neighbouring ranks map, so the halo rows travel as peer-memory stores from a kernel (csrc/bands.cu) --
 * there is no collective on the data path; the host layer (bands.py) only moves the 64-byte handles once and
 * all-reduces the K squared-error sums at the end.
 * Local buffers: hr buffers hold global HR rows [hr_lo, hr_hi); the caller's LR frames hold the global LR rows
 * [lr_ext_lo, lr_ext_hi) = hr_lo/2 - 32 .. hr_hi/2 + 32 clipped to the image (the margin feeds zoom's prefilter). */
int sr459_band_plan_create(sr459_plan** out, int device, int n_frames, int h, int w, int factor,
                           const double* h_shifts_yx, const double* h_psf, int kh, int kw,
                           double step, double clip_lo, double clip_hi, int rank, int nranks);
/* h_out8 = { own_y0, own_y1, hr_lo, hr_hi, lr_lo, lr_hi, halo_rows, nranks } (global row numbers; lr_lo..lr_hi are
 * the LR rows the two-call form below takes)                                                                      */
int sr459_band_layout(const sr459_plan* plan, int* h_out8);
/* h_out4 = { lr_ext_lo, lr_ext_hi, fractional (0/1), lr_margin }                                                  */
int sr459_band_layout_ext(const sr459_plan* plan, int* h_out4);
/* device pointers of the band's current image buffer and of the other one ((hr_hi-hr_lo) x W float32 each); the
 * current one is where sr459_band_saa writes, where a caller-made initial image goes (halo rows included) and where
 * the result of sr459_band_run is found afterwards (ask again: the two swap with every iteration).                 */
int sr459_band_buffers(sr459_plan* plan, void** d_cur, void** d_other, int64_t* arena_bytes);
/* index (0/1) of the buffer holding the current image; set_to >= 0 sets it first (replaying a captured run)         */
int sr459_band_current(sr459_plan* plan, int set_to);
/* the arena's cudaIpcMemHandle_t (64 bytes; for a neighbour in another process) and/or its raw device pointer (for
 * a neighbour in this process); either output may be NULL                                                          */
int sr459_band_export(sr459_plan* plan, unsigned char* h_handle64, void** h_arena_ptr);
/* map the arenas of the band above (rank-1) and below (rank+1): per direction EITHER a handle from another process
 * OR the raw pointer + device of a plan in this process; NULL, NULL for a direction without a neighbour            */
int sr459_band_attach(sr459_plan* plan, const unsigned char* up_handle64, void* up_arena, int up_device,
                      const unsigned char* down_handle64, void* down_arena, int down_device);
/* shift_and_add restricted to this band (halo rows included) from LR rows only -> the current buffer             */
int sr459_band_saa(sr459_plan* plan, const float* const* h_d_lr_ext, void* stream);
/* optional: allocate / grow what sr459_band_run(n_iter) needs now (growing frees device memory = a device-wide sync) */
int sr459_band_reserve(sr459_plan* plan, const float* const* h_d_lr_ext, int n_iter, void* stream);
/* n_iter iterations on the current buffer; after each one a kernel stores the band's boundary rows into the
 * neighbours' halo rows and waits for theirs.  Every rank must make the same calls.  Enqueues only (capturable).   */
int sr459_band_run(sr459_plan* plan, const float* const* h_d_lr_ext, int n_iter, void* stream);
/* the iterations [it0, it1) of a run of n_iter (it0 == 0 prepares; sr459_band_run_close ends it): one process driving
 * several bands of ONE GPU interleaves the bands iteration by iteration, so that no kernel ever waits for one that is
 * enqueued behind it (streams of one device may share a hardware queue)                                            */
int sr459_band_run_range(sr459_plan* plan, const float* const* h_d_lr_ext, int it0, int it1, int n_iter, void* stream);
int sr459_band_run_close(sr459_plan* plan, void* stream);
/* two-call form with caller-owned buffers (LR rows [lr_lo, lr_hi)) and the exchange left to the host layer:
 * once per ibp() call ...                                                                                          */
int sr459_band_begin(sr459_plan* plan, const float* const* h_d_lr_local, int n_iter, void* stream);
/* ... then iteration `it`: d_nxt[owned rows] = update(d_cur); halo rows of d_nxt are NOT written                   */
int sr459_band_iterate(sr459_plan* plan, const float* const* h_d_lr_local, const float* d_cur,
                       float* d_nxt, int it, void* stream);
/* h_sums[it] = sum over the frames and the sites this rank owns of err^2 (un-normalised; the host layer all-reduces
 * and divides by n*h*w as run_sr.py:202,206 do).  Synchronises the stream; SR459_ECUDA if an exchange timed out.   */
int sr459_band_errors(sr459_plan* plan, int n_iter, double* h_sums, void* stream);

/* ---- shift calibration / PSF characterisation ------------------------------------------------- */
/* find_peak(img, smooth_sigma=2.0)                data_collection/psf_mtf_utils.py:48-51, batched: for each of the
 * n_img images (uint8 or float32, H x W, contiguous one after the other) the flat index row*W+col of the FIRST maximum
 * of scipy's gaussian_filter(img, sigma) (float64, boundary 'reflect', truncate 4).  sigma <= 3.  d_smooth_out (may be
 * NULL): n_img*H*W doubles receiving the smoothed images.  Synchronises the stream.                                 */
int sr459_find_peaks(const void* d_imgs, int is_f32, int n_img, int H, int W, double sigma, int* h_peaks,
                     double* d_smooth_out, void* stream);
/* find_pinhole_center(img)        calibration_beam_shift/calibrate_shift_grid.py:66-96, batched, from the peaks on:
 * extract_psf (psf_mtf_utils.py:54-71: ROI of +-radius around the peak, clipped at the image, minus the bg_percentile
 * of the ring >= 0.6*radius from the ROI centre, clamped at 0), fit_gaussian_psf (:115-131: 7-parameter rotated Gaussian,
 * same start point and bounds; projected Levenberg-Marquardt, at most max_iter accepted steps), and the centre-of-mass
 * fall-back (:88-96).  h_out: 12 doubles per image = cx, cy (image coordinates), amp, x0, y0, sigma_x, sigma_y, theta,
 * offset (ROI coordinates; NaN without a fit), status (0 fit, 1 centre of mass, 2 peak), peak row, peak col.
 * h_roi (may be NULL): 31*31 doubles per image, the background-subtracted ROI with row pitch = its width;
 * h_roi_dims (may be NULL): rows, cols, first row, first col per image.  radius <= 15.  is_f32 == 2: float64 images;
 * bg_percentile < 0: no background removal (fit_gaussian_psf on a ready-made ROI).  Synchronises the stream.           */
int sr459_pinhole_fit(const void* d_imgs, int is_f32, int n_img, int H, int W, const int* h_peaks, int radius,
                      double bg_percentile, int max_iter, double* h_out, double* h_roi, int* h_roi_dims, void* stream);
/* compute_mtf(psf)                                 psf_mtf_utils.py:134-170: the PSF (host, float64, ph x pw) centred in
 * a pad x pad zero array (pad even), normalised to unit sum, |DFT2| with the origin at the centre, normalised to its
 * maximum; h_radial: pad/2 ring means about (pad/2, pad/2); h_mtf2d (may be NULL): pad*pad.  Synchronises.          */
int sr459_psf_mtf(const double* h_psf, int ph, int pw, int pad, double* h_mtf2d, double* h_radial, void* stream);

/* ---- quality analytics on HR outputs ------------------------------------------------------------ */
/* slanted_edge_esf(roi, side) + esf_to_mtf + mtf_at_fraction   mono_cal_target/analysis.ipynb cells 7 and 10, for the
 * ROI [ry0, ry0+rh) x [rx0, rx0+rw) of a device image (uint8 or float32, row pitch in elements).  right_side: 0 = 'left'.
 * Outputs (host, each may be NULL, capacity sr459_edge_mtf_capacity() doubles): esf_x, esf_y, lsf (n_esf values), freq
 * and mtf (n_esf / 2 values, cycles per pixel).  h_scal (8 doubles): status, n_esf, edge angle in degrees, MTF50 and
 * MTF10 in cycles/mm at pixel pitch pitch_mm (NaN where the notebook's function returns nan), number of edge pixels,
 * number on the chosen side, gradient threshold.  The notebook's RuntimeErrors come back as SR459_EINVAL with the same
 * text in sr459_last_error().  Synchronises the stream.                                                               */
int sr459_edge_mtf(const void* d_img, int is_f32, int pitch, int ry0, int rx0, int rh, int rw, int right_side,
                   double pitch_mm, double* h_esf_x, double* h_esf_y, double* h_lsf, double* h_freq, double* h_mtf,
                   double* h_scal, void* stream);
int sr459_edge_mtf_capacity(void);
/* local_contrast(profile, window=20)               analysis.ipynb cell 4 (host arrays in and out)                    */
int sr459_local_contrast(const double* h_profile, int n, int window, double* h_out, void* stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* SR459_H */
