/* radler_b200.h — C ABI of the B200-native Radler minor-loop library.
 *
 * Drop-in boundary: the reference's plugin interface
 *   radler::algorithms::DeconvolutionAlgorithm
 *   (cpp/algorithms/deconvolution_algorithm.h:60-210; virtuals
 *    ExecuteMajorIteration :69-71 and Clone :73)
 * The reference has no C ABI of its own; the closest precedent is the flat-array
 * plugin of PythonDeconvolution (cpp/algorithms/python_deconvolution.cc:171-275).
 * Every entry point below names the reference interface it replaces. All
 * pointers are plain; no C++/torch types cross this boundary. Status codes
 * replace exceptions; rdl_last_error() returns the message a C++ shim re-throws
 * as std::runtime_error.
 *
 * There is NO CPU fallback: every compute entry point fails with
 * RDL_ERR_NO_DEVICE when no CUDA device (sm_100) is usable.
 */
#ifndef RADLER_B200_H_
#define RADLER_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rdl_handle rdl_handle;

typedef enum {
  RDL_OK = 0,
  RDL_ERR_INVALID_ARGUMENT = 1, /* std::runtime_error in the reference (dimension mismatch ...) */
  RDL_ERR_NO_DEVICE = 2,
  RDL_ERR_CUDA = 3,
  RDL_ERR_UNSUPPORTED = 4, /* settings the reference rejects or this build does not cover */
  RDL_ERR_OUT_OF_MEMORY = 5
} rdl_status;

/* radler::AlgorithmType (cpp/settings.h) subset on the hot path */
typedef enum { RDL_GENERIC_CLEAN = 0, RDL_MULTISCALE = 1, RDL_IUWT = 2 } rdl_algorithm;
/* radler::MultiscaleShape (cpp/settings.h) */
typedef enum { RDL_SHAPE_TAPERED_QUADRATIC = 0, RDL_SHAPE_GAUSSIAN = 1 } rdl_multiscale_shape;
/* schaapcommon::fitters::SpectralFittingMode subset */
typedef enum { RDL_FIT_NONE = 0, RDL_FIT_POLYNOMIAL = 1 } rdl_spectral_mode;
typedef enum { RDL_MEM_HOST = 0, RDL_MEM_DEVICE = 1 } rdl_memory_space;

/* Construction-time configuration: the constructor arguments of GenericClean
 * (generic_clean.h: use_sub_minor_optimization), MultiScaleAlgorithm
 * (multiscale_algorithm.cc:153-163: Settings::Multiscale, beam size, pixel
 * scales, track_components) and IuwtDeconvolution. */
typedef struct {
  int32_t algorithm;                  /* rdl_algorithm */
  int32_t device;                     /* CUDA device ordinal (new knob, not in the reference) */
  int32_t use_sub_minor_optimization; /* Settings::Generic, settings.h:527-532 */
  /* Settings::Multiscale, settings.h:465-524 */
  int32_t ms_fast_sub_minor_loop;
  double ms_sub_minor_loop_gain;
  double ms_scale_bias;
  uint64_t ms_max_scales;
  double ms_convolution_padding;
  const double* ms_scale_list;
  uint64_t ms_n_scale_list;
  int32_t ms_shape; /* rdl_multiscale_shape */
  double beam_size;
  double pixel_scale_x, pixel_scale_y;
  int32_t track_components;
} rdl_config;

/* Fills the reference's defaults (settings.h; deconvolution_algorithm.h:189-202). */
void rdl_config_default(rdl_config* cfg);

/* Describes an ImageSet (cpp/image_set.h:19-381): images ordered
 * [deconvolution channel][polarization], all width x height fp32 row-major. */
typedef struct {
  uint64_t width, height;
  uint64_t n_channels;           /* NDeconvolutionChannels() == PsfCount() */
  uint64_t n_pol;                /* images per channel */
  const float* channel_weights;  /* ImageSet::weights_ [n_channels]; NULL = all 1 */
  const uint8_t* pol_linked;     /* [n_pol] 1 if in linked_polarizations (or the set is empty); NULL = all */
  float pol_norm;                /* polarization_normalization_factor_ (image_set.h:286-313) */
  int32_t squared_joins;         /* Settings::squared_joins */
  int32_t memory_space;          /* rdl_memory_space of residual/model/psf pointers */
} rdl_cube_desc;

/* DeconvolutionResult (deconvolution_algorithm.h:31-58) */
typedef struct {
  int32_t has_starting_peak; /* OptionalNumber<float>::has_value */
  float starting_peak_value;
  float final_peak_value;
  int32_t another_iteration_required;
  int32_t is_diverging;
  uint64_t iteration_number; /* IterationNumber() after the call */
} rdl_result;

/* One recorded component, for parity checks of peak sequences. */
typedef struct {
  uint32_t iteration;
  uint32_t scale_index;
  uint32_t x, y;
  float value;
} rdl_trace_entry;

/* MultiScaleAlgorithm::ScaleInfo (multiscale_algorithm.h:58-77) */
typedef struct {
  float scale, psf_peak, kernel_peak, bias_factor, gain;
  float max_normalized_image_value, max_unnormalized_image_value, rms;
  uint64_t max_x, max_y;
  int32_t is_active;
  uint64_t n_components_cleaned;
  float total_flux_cleaned;
} rdl_scale_info;

/* ---- lifetime: constructors / Clone() / destructor ------------------------- */
rdl_status rdl_create(const rdl_config* cfg, rdl_handle** out);
/* DeconvolutionAlgorithm::Clone (deconvolution_algorithm.h:73): deep copy incl. state */
rdl_status rdl_clone(const rdl_handle* h, rdl_handle** out);
void rdl_destroy(rdl_handle* h);
/* Device buffers of 1 MB and more that the library releases are kept in a per-process cache and
 * reused by later calls (a multiscale major iteration would otherwise spend 0.1-0.4 s in
 * cudaMalloc / cudaFree). The cache is emptied when the last handle is destroyed, when an
 * allocation fails, or by this call. RDL_POOL_MAX_GB (environment) caps it; default 96. */
void rdl_release_cached_memory(void);
/* message of the last failing call on this handle (never NULL) */
const char* rdl_last_error(const rdl_handle* h);

/* ---- setters mirroring DeconvolutionAlgorithm (deconvolution_algorithm.h:75-166) */
rdl_status rdl_set_max_iterations(rdl_handle* h, uint64_t n);             /* SetMaxIterations */
rdl_status rdl_set_threshold(rdl_handle* h, float v);                     /* SetThreshold */
rdl_status rdl_set_major_iteration_threshold(rdl_handle* h, float v);     /* SetMajorIterationThreshold */
rdl_status rdl_set_minor_loop_gain(rdl_handle* h, float v);               /* SetMinorLoopGain */
rdl_status rdl_set_major_loop_gain(rdl_handle* h, float v);               /* SetMajorLoopGain */
rdl_status rdl_set_allow_negative_components(rdl_handle* h, int32_t v);   /* SetAllowNegativeComponents */
rdl_status rdl_set_stop_on_negative_components(rdl_handle* h, int32_t v); /* SetStopOnNegativeComponents */
rdl_status rdl_set_clean_border_ratio(rdl_handle* h, float v);            /* SetCleanBorderRatio */
rdl_status rdl_set_divergence_limit(rdl_handle* h, float v);              /* SetDivergenceLimit */
rdl_status rdl_set_iteration_number(rdl_handle* h, uint64_t n);           /* SetIterationNumber */
/* SetCleanMask: non-owning const bool* in the reference; here the mask (width*height
 * bytes, host OR device memory -- e.g. what rdl_auto_mask_from_model wrote) is copied at the
 * call, with a blocking copy that orders after work on the handle's own stream. NULL clears it. */
rdl_status rdl_set_clean_mask(rdl_handle* h, const uint8_t* mask, uint64_t width, uint64_t height);
/* SetRmsFactorImage (moved-in aocommon::Image); NULL clears it. Host or device memory (e.g. the
 * plane rdl_rms_make_factor_image converted), copied at the call. */
rdl_status rdl_set_rms_factor_image(rdl_handle* h, const float* image, uint64_t width, uint64_t height);
/* SetSpectralFitter(unique_ptr<SpectralFitter>, n_polarizations)
 * (deconvolution_algorithm.h:137-142; fitter built in radler.cc:318-331). */
rdl_status rdl_set_spectral_fitter(rdl_handle* h, int32_t mode, uint64_t n_terms,
                                   const double* frequencies, const float* weights,
                                   uint64_t n_channels, uint64_t n_polarizations);
/* NOT SUPPORTED (no entry point): SetSpectrallyForcedImages (deconvolution_algorithm.h:155-158,
 * the forced-terms fit: RDL_FIT modes other than polynomial are rejected with
 * RDL_ERR_UNSUPPORTED) and SetLogReceiver (:113-115; the library does not log -- status codes
 * and rdl_last_error carry what the reference would report). */

/* Radler passes the same psf_images to every major iteration of a run (radler.cc:175-268): with
 * rdl_set_psfs_constant(h, 1) the caller promises that PSF planes passed again under the same
 * pointers are unchanged. RDL_MEM_HOST calls then copy them to the device once instead of on
 * every call (any other pointer, plane size or plane count uploads again), and a multiscale
 * handle (either memory space) keeps what it derives from the PSFs alone -- scale-convolved and
 * twice-convolved PSFs, kernel spectra, PSF peaks -- from one call to the next, which holds HBM
 * between calls (K4: up to ~60 GB) until the option is cleared or the handle destroyed. Results
 * are the same bits with and without the option. */
rdl_status rdl_set_psfs_constant(rdl_handle* h, int32_t constant);
/* RDL_MEM_HOST calls only. With rdl_set_pin_host_buffers(h, 1) the caller's residual / model /
 * PSF planes are registered (cudaHostRegister) the first time they are seen, so that pageable
 * aocommon::Image buffers are copied at PCIe rate; they are unregistered by rdl_destroy or by
 * switching the option off. Planes that cannot be registered are copied as they are. The caller
 * must keep a registered plane allocated until then. */
rdl_status rdl_set_pin_host_buffers(rdl_handle* h, int32_t enable);

/* MultiScaleAlgorithm::SetAutoMaskMode (multiscale_algorithm.h:41-45) */
rdl_status rdl_set_auto_mask_mode(rdl_handle* h, int32_t track_per_scale_masks,
                                  int32_t use_per_scale_masks);
/* record up to `capacity` components of the next calls (0 disables) */
rdl_status rdl_set_trace_capacity(rdl_handle* h, uint64_t capacity);

/* ---- getters ----------------------------------------------------------------- */
uint64_t rdl_iteration_number(const rdl_handle* h);             /* IterationNumber */
uint64_t rdl_scale_count(const rdl_handle* h);                  /* ScaleCount */
rdl_status rdl_scale_info_get(const rdl_handle* h, uint64_t i, rdl_scale_info* out); /* ScaleSize & co */
/* ---- component list (cpp/component_list.h; rdl_config.track_components = settings.save_source_list)
 * MultiScaleAlgorithm::GetComponentList() (multiscale_algorithm.h:46-48): the components of every
 * major iteration since creation / the last clear, per scale, after ComponentList::MergeDuplicates
 * (component_list.h:113-131): one record per position, values summed in the order the components
 * were found, one value per image of the cube. rdl_component_get copies `count` records starting
 * at `first`: x[count], y[count], values[count * n_images]. */
uint64_t rdl_component_scale_count(const rdl_handle* h);
rdl_status rdl_component_count(rdl_handle* h, uint64_t scale_index, uint64_t* count);
rdl_status rdl_component_get(rdl_handle* h, uint64_t scale_index, uint64_t first, uint64_t count,
                             uint64_t* x, uint64_t* y, float* values);
/* MultiScaleAlgorithm::ClearComponentList() */
void rdl_component_list_clear(rdl_handle* h);

/* GetScaleMask(i): copies width*height bytes to host memory */
rdl_status rdl_scale_mask_get(const rdl_handle* h, uint64_t i, uint8_t* out);
uint64_t rdl_trace_size(const rdl_handle* h);
uint64_t rdl_trace_get(const rdl_handle* h, rdl_trace_entry* out, uint64_t capacity);
/* number of pixels the last sub-minor run selected (diagnostic) */
uint64_t rdl_last_selected_count(const rdl_handle* h);
/* outer minor-loop passes of the last multiscale call (diagnostic) */
uint64_t rdl_minor_iteration_count(const rdl_handle* h);
/* number of kernels this handle has launched so far (bench.py "gpu_launches") */
uint64_t rdl_kernel_launch_count(const rdl_handle* h);
/* How many schaapcommon::math::Convolve calls the reference makes for the work of the last
 * rdl_execute_major_iteration, by frame size: up to `capacity` (width, height, count) triples;
 * returns the number of distinct sizes. (The device path caches kernel spectra and twice-convolved
 * PSFs and runs fewer transforms; this is the count a CPU baseline has to be priced with,
 * multiscale_algorithm.cc:29-88,331-354,432-461,578-634, subminor_loop.cc:195-218.) */
uint64_t rdl_reference_convolutions(const rdl_handle* h, uint64_t* triples, uint64_t capacity);

/* ---- the hot path: DeconvolutionAlgorithm::ExecuteMajorIteration ------------
 * (deconvolution_algorithm.h:69-71). residual[i], model[i] (i < n_channels*n_pol)
 * are updated in place; psfs[c] (c < n_channels) are read-only, same size as the
 * images. With RDL_MEM_HOST the images are staged to HBM and copied back inside
 * the call; with RDL_MEM_DEVICE they are used where they are. `stream` is a
 * cudaStream_t (NULL = the handle's own stream).
 * RDL_MEM_HOST + GenericClean without the sub-minor loop (one GPU, no facets): the model planes
 * are neither uploaded nor downloaded -- the call records its components (pixel, one value per
 * image) on the device and adds them to the caller's planes before it returns, in order and with
 * the same fp32 additions, so the result is the same bits (used while the component log stays
 * below 256 MB, i.e. max_iterations - iteration_number below ~4 M for 16 images). */
rdl_status rdl_execute_major_iteration(rdl_handle* h, const rdl_cube_desc* cube,
                                       float* const* residual, float* const* model,
                                       const float* const* psfs, void* stream, rdl_result* out);

/* ---- facets: ParallelDeconvolution::ExecuteMajorIteration ------------------------------------
 * (cpp/algorithms/parallel_deconvolution.cc:494-654) over a grid_width x grid_height grid of
 * sub-images (Settings::Parallel::grid_width/height, settings.h). The handle is the prototype
 * algorithm (SetAlgorithm clones it once per facet; the clones and their iteration counters live
 * in the handle across calls); its setters reach every facet like ParallelDeconvolution's
 * (:244-292). Dividing paths are computed on the host from the integrated residual
 * (math/dijkstra_splitter.cc), everything else runs on the device: per-facet trims, the
 * peak-only pass, the common major-iteration threshold start_peak * (1 - mgain), deconvolution,
 * divergence rollback, masked copy-back and model accumulation. A 1 x 1 grid is
 * ExecuteSingleThreadedRun. Facets run in index order (the reference with one thread). */
rdl_status rdl_parallel_execute_major_iteration(rdl_handle* h, uint64_t grid_width,
                                                uint64_t grid_height, const rdl_cube_desc* cube,
                                                float* const* residual, float* const* model,
                                                const float* const* psfs, void* stream,
                                                rdl_result* out);
/* The same with direction-dependent PSFs (ParallelDeconvolution::ExecuteMajorIteration's
 * psf_images[dd][channel] + psf_offsets, parallel_deconvolution.cc:494-503): psfs holds
 * n_dd * n_channels planes ordered [dd][channel], each psf_width x psf_height (not larger than
 * the image), psf_offset_x/y[n_dd] the pixel each PSF set was made for. Every facet uses the set
 * NearestPsfIndex picks for its centre (:34-55, :163), resized to the facet like
 * aocommon::Image::Resize (:329); a 1 x 1 grid uses the set nearest to the image centre, padded to
 * the image size (:518-546). */
rdl_status rdl_parallel_execute_major_iteration_dd(
    rdl_handle* h, uint64_t grid_width, uint64_t grid_height, const rdl_cube_desc* cube,
    float* const* residual, float* const* model, const float* const* psfs, uint64_t n_dd,
    uint64_t psf_width, uint64_t psf_height, const int64_t* psf_offset_x,
    const int64_t* psf_offset_y, void* stream, rdl_result* out);
/* SubImage of the last facet run (parallel_deconvolution.h:22-35) */
typedef struct {
  uint64_t x, y, width, height;
  double peak;
  int32_t reached_major_threshold;
  uint64_t iteration_number; /* of this facet's algorithm clone */
} rdl_facet_info;
uint64_t rdl_parallel_facet_count(const rdl_handle* h);
/* Wall-clock split of the last rdl_parallel_execute_major_iteration call on this handle: the
 * host search for the facet dividers (DijkstraSplitter, cpp/math/dijkstra_splitter.cc) and the
 * whole call, in milliseconds. Diagnostics only; the reference has no counterpart. */
rdl_status rdl_parallel_timing(const rdl_handle* h, double* host_geometry_ms, double* total_ms);
/* boundary_mask (optional): width*height bytes, the pixels this facet owns */
rdl_status rdl_parallel_facet_get(const rdl_handle* h, uint64_t i, rdl_facet_info* out,
                                  uint8_t* boundary_mask);

/* Facet geometry on the host (no device needed): radler::math::DijkstraSplitter
 * (cpp/math/dijkstra_splitter.h:40-110) and MakeSubImages (parallel_deconvolution.cc:69-168).
 * dividing_path = DivideVertically (vertical != 0; [lo,hi) are columns) / DivideHorizontally;
 * flood_area = FloodVerticalArea / FloodHorizontalArea; bounding_mask = GetBoundingMask with
 * box = {x, y, width, height}; facet_grid fills grid_width*grid_height boxes, row-major. */
rdl_status rdl_facet_dividing_path(const float* image, float* output, uint64_t width,
                                   uint64_t height, uint64_t lo, uint64_t hi, int32_t vertical);
rdl_status rdl_facet_flood_area(const float* dividers, uint64_t width, uint64_t height,
                                uint64_t inside, int32_t vertical, uint8_t* mask, uint64_t* start,
                                uint64_t* extent);
rdl_status rdl_facet_bounding_mask(const uint8_t* vertical_mask, uint64_t vertical_mask_x,
                                   uint64_t vertical_mask_width, const uint8_t* horizontal_mask,
                                   uint8_t* mask, uint64_t width, uint64_t height, uint64_t* box);
rdl_status rdl_facet_grid(const float* image, uint64_t width, uint64_t height,
                          const uint8_t* user_mask, uint64_t grid_width, uint64_t grid_height,
                          rdl_facet_info* out);

/* ---- kernel-level entry points (device pointers; used by tests and the roofline
 * harness, and what a host-side caller can compose itself) ------------------------ */

/* peak_finder::Find / FindWithMask (cpp/math/peak_finder.h:80-127) fused with
 * ImageSet::GetLinearIntegrated (mode 0) / GetSquareIntegrated (mode 1)
 * (cpp/image_set.h:131-155) and the optional RMS-factor multiply
 * (generic_clean.cc:255-277). images: device pointers. found/x/y/value: host outputs. */
rdl_status rdl_find_peak(rdl_handle* h, const rdl_cube_desc* cube, const float* const* images,
                         int32_t integration_mode, int32_t allow_negative, uint64_t start_y,
                         uint64_t end_y, uint64_t horizontal_border, uint64_t vertical_border,
                         const uint8_t* device_mask, const float* device_rms_factor,
                         void* stream, int32_t* found, uint64_t* x, uint64_t* y, float* value);

/* ThreadedDeconvolutionTools::SubtractImage for all images of a cube in one launch
 * (threaded_deconvolution_tools.cc:18-28 -> simple_clean.cc:61-96):
 * images[i] -= psfs[i / n_pol] shifted to (x, y), times factors[i]. Device pointers;
 * factors is a host array of n_channels*n_pol floats. */
rdl_status rdl_subtract_image(rdl_handle* h, const rdl_cube_desc* cube, float* const* images,
                              const float* const* psfs, uint64_t x, uint64_t y,
                              const float* factors, void* stream);

/* ImageSet::GetLinearIntegrated / GetSquareIntegrated materialised (device pointers). */
rdl_status rdl_integrate(rdl_handle* h, const rdl_cube_desc* cube, const float* const* images,
                         int32_t integration_mode, float* device_dest, void* stream);

/* schaapcommon::math::Convolve (called at multiscale_transforms.cc:21,
 * subminor_loop.cc:210): in-place circular convolution of n_images device images
 * (width x height) with one device kernel image prepared by PrepareConvolutionKernel /
 * PrepareSmallConvolutionKernel. */
rdl_status rdl_convolve(rdl_handle* h, float* const* images, uint64_t n_images,
                        const float* device_kernel, uint64_t width, uint64_t height, void* stream);

/* schaapcommon::math::PaddedConvolution as used by CorrectResidualDirty
 * (subminor_loop.cc:195-218): out = Trim(Convolve(Untrim(image, padded), PrepareConvolutionKernel(
 * Untrim(psf, padded)))). image / psf / out are width x height device images; out may alias
 * image. The padded size is free (GenericClean: even(ceil(1.1f * N)), generic_clean.cc:63-66);
 * sizes with large prime factors are evaluated through a larger smooth frame (same result). */
rdl_status rdl_padded_convolution(rdl_handle* h, const float* device_image, const float* device_psf,
                                  float* device_out, uint64_t width, uint64_t height,
                                  uint64_t padded_width, uint64_t padded_height, void* stream);

/* MultiScaleTransforms::Transform (multiscale_transforms.cc:11-23). */
rdl_status rdl_multiscale_transform(rdl_handle* h, float* const* images, uint64_t n_images,
                                    uint64_t width, uint64_t height, float scale, int32_t shape,
                                    void* stream);

/* ---- multi-GPU: one process per GPU, rows of the cube sharded over the ranks ----------------
 * The reference has no multi-device path; its closest analogue is the row split of
 * SubtractImage / FindPeak over threads (threaded_deconvolution_tools.cc:18-28,
 * peak_finder.h start_y/end_y). Here every rank holds the rows [y_offset, y_offset + height)
 * of every image (cube->height is the band's row count) plus the full-frame PSFs; FindPeak and
 * SubtractImage are local to the band, and the only exchange of a minor iteration is the joined
 * arg-max: an NCCL all-gather of (64-bit peak key with the frame's row-major index, per-image
 * values at that pixel) per rank, after which every rank runs the identical loop control.
 * Results are bit-identical to the single-GPU loop. GenericClean without the sub-minor
 * optimisation only. */
/* ncclGetUniqueId: call on one rank, ship the 128 bytes to all ranks (e.g. torch.distributed). */
rdl_status rdl_dist_unique_id(uint8_t* out128);
/* ncclCommInitRank on the handle's device (collective over all ranks). */
rdl_status rdl_dist_init(rdl_handle* h, const uint8_t* id128, int32_t rank, int32_t world);
/* The next rdl_execute_major_iteration calls work on rows [y_offset, y_offset + cube->height) of
 * a width x global_height frame; PSFs, clean mask and RMS image stay full-frame.
 * global_height = 0 returns to whole frames. */
rdl_status rdl_dist_set_row_band(rdl_handle* h, uint64_t y_offset, uint64_t global_height);
/* Cyclic variant (after rdl_dist_init): rank r owns the row blocks b with b % world == r, block b
 * = rows [b * block_rows, (b + 1) * block_rows), stored one after the other; cube->height =
 * global_height / world. A PSF window covers half the frame, so contiguous bands leave some
 * ranks with three times the work of others; cyclic blocks balance it. global_height must be a
 * multiple of block_rows * world. */
rdl_status rdl_dist_set_row_blocks(rdl_handle* h, uint64_t block_rows, uint64_t global_height);
rdl_status rdl_dist_finalize(rdl_handle* h);

/* ---- IUWT decomposition (cpp/algorithms/iuwt/iuwt_decomposition.{h,cc}); device pointers ----
 * One B3-spline a-trous smoothing, taps {1,4,6,4,1}/16 at (2^scale - 1) * {-2..2}, scale >= 1.
 * flavour 0: IuwtDecomposition::convolveMT (iuwt_decomposition.cc:55-67, used by DecomposeMt);
 * flavour 1: IuwtDecomposition::convolve (iuwt_decomposition.h:249-267, used by Recompose and
 * Convolve). Out of place. */
rdl_status rdl_iuwt_smooth(rdl_handle* h, const float* device_in, float* device_out,
                           uint64_t width, uint64_t height, int32_t scale, int32_t flavour,
                           void* stream);
/* IuwtDecomposition::Decompose / DecomposeMt (iuwt_decomposition.cc:9-53): writes n_scales
 * wavelet planes (+ the smooth remainder as plane n_scales when include_largest) of
 * width*height floats each into device_coefficients. */
rdl_status rdl_iuwt_decompose(rdl_handle* h, const float* device_input, uint64_t width,
                              uint64_t height, int32_t n_scales, int32_t include_largest,
                              float* device_coefficients, void* stream);
/* The same decomposition for callers that may not allocate device memory (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc). No counterpart in the
 * reference. The two ping-pong planes that
 * rdl_iuwt_decompose keeps in cudaMalloc'ed buffers of the handle are a file-scope device array
 * of the library instead (8 MiB, held by every GPU that loads the library until exit); the call
 * runs no cudaMalloc, cudaFree or allocating library routine. Limit: width * height <=
 * RDL_STATIC_SCRATCH_MAX_PIXELS, larger images fail with RDL_ERR_INVALID_ARGUMENT and a message.
 * All calls on a device share the array, which keeps the last call's contents (nothing here
 * needs it cleared): calls that overlap on the GPU race, so serialise them. Results are the bits
 * of rdl_iuwt_decompose. Asynchronous on `stream` like rdl_iuwt_decompose. */
#define RDL_STATIC_SCRATCH_MAX_PIXELS (1u << 20)
rdl_status rdl_iuwt_decompose_static(rdl_handle* h, const float* device_input, uint64_t width,
                                     uint64_t height, int32_t n_scales, int32_t include_largest,
                                     float* device_coefficients, void* stream);
/* IuwtDecomposition::Decompose(x, x, false): the input aliased to the scratch image, which is
 * how IuwtDeconvolutionAlgorithm calls it (iuwt_deconvolution_algorithm.cc:344,385,685,780) and
 * which changes scale 0 (iuwt_decomposition.cc:19-26,58-61: a recursive row pass, and the
 * difference taken against the scratch content). n_scales planes; device_x is clobbered. */
rdl_status rdl_iuwt_decompose_aliased(rdl_handle* h, float* device_x, uint64_t width,
                                      uint64_t height, int32_t n_scales,
                                      float* device_coefficients, void* stream);

/* IuwtDecomposition::ApplyMask (iuwt_decomposition.h:162-169; device_masks = n_scales planes of
 * width*height bytes, NULL = no mask) followed by Recompose (iuwt_decomposition.h:121-147). */
rdl_status rdl_iuwt_recompose(rdl_handle* h, const float* device_coefficients, uint64_t width,
                              uint64_t height, int32_t n_scales, int32_t include_largest,
                              const uint8_t* device_masks, float* device_out, void* stream);

/* ---- IUWT structure selection (cpp/algorithms/iuwt/image_analysis.{h,cc}); device pointers ----
 * device_coefficients: n_scales planes of width*height floats (IuwtDecomposition);
 * device_masks: n_scales planes of width*height bytes (IuwtMask), updated in place;
 * thresholds: n_scales HOST floats. A negative threshold t selects v < t or v > -t
 * (ExceedsThreshold, image_analysis.cc:9-15). clean_border in [0, 0.5). These calls return
 * when the result is on the host (they synchronise the stream).
 *
 * image_analysis::SelectStructures (image_analysis.cc:208-239): marks every voxel of scales
 * [min_scale, end_scale) that exceeds its threshold, is not marked yet, lies inside
 * device_prior_mask (width*height bytes, NULL = everywhere) and inside the clean border;
 * *area_size = number of voxels newly marked. */
rdl_status rdl_iuwt_select_structures(rdl_handle* h, const float* device_coefficients,
                                      uint8_t* device_masks, uint64_t width, uint64_t height,
                                      int32_t n_scales, const float* thresholds, int32_t min_scale,
                                      int32_t end_scale, float clean_border,
                                      const uint8_t* device_prior_mask, uint64_t* area_size,
                                      void* stream);
/* image_analysis::Floodfill (image_analysis.cc:86-149) and, with a prior mask, MaskedFloodfill
 * (:151-206): marks the 6-connected (x, y, scale) structure around the seed voxel;
 * *area_size = voxels filled, the seed included. */
rdl_status rdl_iuwt_floodfill(rdl_handle* h, const float* device_coefficients,
                              uint8_t* device_masks, uint64_t width, uint64_t height,
                              int32_t n_scales, const float* thresholds, int32_t min_scale,
                              int32_t end_scale, uint64_t seed_x, uint64_t seed_y,
                              int32_t seed_scale, float clean_border,
                              const uint8_t* device_prior_mask, uint64_t* area_size, void* stream);
/* image_analysis::FloodFill2D (image_analysis.cc:241-333): use_abs = 0 is the ExceedsThreshold
 * overload that returns the area size (:241-283), use_abs = 1 the fabs(v) > threshold overload
 * that collects the area (:285-333; its pixels are the bytes newly set in device_mask). */
rdl_status rdl_iuwt_floodfill_2d(rdl_handle* h, const float* device_image, uint8_t* device_mask,
                                 uint64_t width, uint64_t height, float threshold, int32_t use_abs,
                                 uint64_t seed_x, uint64_t seed_y, uint64_t* area_size,
                                 void* stream);
/* IuwtDeconvolutionAlgorithm::Mad (cpp/algorithms/iuwt_deconvolution_algorithm.cc:104-110):
 * element n/2 of the sorted |v| of a plane of n floats, divided by 0.674559. */
rdl_status rdl_iuwt_mad(rdl_handle* h, const float* device_image, uint64_t n, float* mad,
                        void* stream);
/* IuwtDeconvolutionAlgorithm::GetMaxAbsWithoutMask / GetMaxAbsWithMask
 * (iuwt_deconvolution_algorithm.cc:112-167): first row-major maximum of |v| (allow_negative)
 * or v inside the clean border and, when given, inside device_mask (width*height bytes). When
 * nothing qualifies *x = width, *y = height and *value = -FLT_MAX, as in the reference. */
rdl_status rdl_iuwt_max_abs(rdl_handle* h, const float* device_image, uint64_t width,
                            uint64_t height, float clean_border, int32_t allow_negative,
                            const uint8_t* device_mask, float* value, uint64_t* x, uint64_t* y,
                            void* stream);

/* utils::CalculateGoodFFTSize / GetConvolutionSize (cpp/utils/fft_size_calculations.h:15-50) */
/* ---- local-RMS maps (cpp/math/rms_image.cc), the producer of the image that
 * rdl_set_rms_factor_image consumes; device pointers ------------------------------------------
 * SlidingMinimum / SlidingMaximum (rms_image.cc:39-84): separable minimum over the half-open
 * windows [x - w/2, x + w/2) clipped to the image, exactly the reference's bounds. Out of place. */
rdl_status rdl_rms_sliding_minimum(rdl_handle* h, const float* device_in, float* device_out,
                                   uint64_t width, uint64_t height, uint64_t window_size,
                                   int32_t maximum, void* stream);
/* MakeRmsFactorImage (rms_image.cc:97-123), in place: value = (min / value)^strength for non-zero
 * pixels (strength 0: all ones); *stddev = the minimum. Negative pixels: RDL_ERR_INVALID_ARGUMENT. */
rdl_status rdl_rms_make_factor_image(rdl_handle* h, float* device_image, uint64_t n_pixels,
                                     double local_rms_strength, double* stddev, void* stream);
/* Make / MakeWithNegativityLimit (rms_image.cc:16-37, 86-95): sqrt(norm * (image^2 convolved with
 * the Gaussian window)), optionally max'ed with |sliding minimum| * 0.3. The window is
 * schaapcommon::math::RestoreImage in the reference (not in its tree): here a circular FFT
 * convolution with a unit-peak elliptical Gaussian of FWHM beam * window_size, position angle
 * (radians) from +y towards +x; that part is restated, not pinned. */
rdl_status rdl_rms_make(rdl_handle* h, const float* device_in, float* device_out, uint64_t width,
                        uint64_t height, double window_size, double beam_major, double beam_minor,
                        double beam_pa, double pixel_scale_l, double pixel_scale_m,
                        int32_t with_negativity_limit, void* stream);

/* ---- the steps of Radler::Perform either side of ExecuteMajorIteration,
 * for a caller that keeps the cube in device memory between major iterations. Device pointers
 * only; `device_inputs` / `device_outputs` / `device_models` are host arrays of device pointers. */
/* integrated.MedianAndStdDevFromMAD() (radler.cc:162-169; aocommon, restated): median of the
 * finite pixels and 1.4826 x the median of |pixel - median|; n_finite = pixels that took part
 * (0: both results 0). Any output pointer may be null. Synchronises the stream. */
rdl_status rdl_median_and_stddev_from_mad(rdl_handle* h, const float* device_image,
                                         uint64_t n_pixels, float* median, float* stddev,
                                         uint64_t* n_finite, void* stream);
/* ImageSet::LoadAndAverage (image_set.cc:105-140) after the loads: output[output_index[i]] +=
 * input[i] * weights[i] in the order given (weight 0: input not read), then every output is
 * scaled by 1 / (sum of its weights). The reference's order is: per deconvolution channel, its
 * original channels in turn, polarisations innermost; output_index = channel * n_pol + pol. */
rdl_status rdl_load_and_average(rdl_handle* h, const float* const* device_inputs,
                                const double* weights, const uint32_t* output_index,
                                uint64_t n_inputs, float* const* device_outputs,
                                uint64_t n_outputs, uint64_t n_pixels, void* stream);
/* ImageSet::LoadAndAveragePsfs (image_set.cc:142-205) after the loads, one PSF per original
 * channel: the same grouping (output_index = deconvolution channel), but the reference keeps the
 * weights double there -- psf = float(double(psf) + double(input) * weight), then
 * psf = float(double(psf) / sum of weights); inputs of weight 0 are read (x 0), an output without
 * weight is scaled by 0. */
rdl_status rdl_load_and_average_psfs(rdl_handle* h, const float* const* device_inputs,
                                     const double* weights, const uint32_t* output_index,
                                     uint64_t n_inputs, float* const* device_outputs,
                                     uint64_t n_outputs, uint64_t n_pixels, void* stream);
/* The auto-mask derived from the model planes when the second auto-masking phase starts
 * (radler.cc:254-266): mask = 1 where any model plane is finite and non-zero, else 0. */
rdl_status rdl_auto_mask_from_model(rdl_handle* h, const float* const* device_models,
                                    uint64_t n_models, uint8_t* device_mask, uint64_t n_pixels,
                                    void* stream);
/* ImageSet::InterpolateAndStoreModel (image_set.cc:208-283) for one polarisation, the case of
 * fewer deconvolution channels than original ones: per pixel, the polynomial fitter set with
 * rdl_set_spectral_fitter (its channels = the n_models planes) is fitted to the model planes and
 * evaluated at each of output_frequencies (all-zero spectra give zeros). Other fitting modes and
 * more than 8 terms: RDL_ERR_UNSUPPORTED. With as many deconvolution channels as original ones
 * the reference stores the planes as they are: no call needed. Synchronises the stream. */
rdl_status rdl_interpolate_model(rdl_handle* h, const float* const* device_models,
                                 uint64_t n_models, const double* output_frequencies,
                                 uint64_t n_outputs, float* const* device_outputs,
                                 uint64_t n_pixels, void* stream);

uint64_t rdl_good_fft_size(uint64_t minimum_size);
uint64_t rdl_convolution_size(double convolution_scale, uint64_t original_size, double padding);

/* library / device introspection */
const char* rdl_version(void);
int32_t rdl_device_count(void);

#ifdef __cplusplus
}
#endif
#endif /* RADLER_B200_H_ */
