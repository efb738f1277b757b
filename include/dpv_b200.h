/* dpv_b200.h — C ABI of libdpvb200.so: the B200-native (sm_100a) depth-probability-volume (DPV)
 * + light-curtain (LC) hot path of CMU-Light-Curtains/DepthEstimation.
 *
 * Conventions:
 *   - every `const float*` / `float*` / `int32_t*` argument is a DEVICE pointer owned by the caller
 *     unless its comment says "host"; tensors are dense, row-major, fp32;
 *   - `stream` is a cudaStream_t passed as void*; every function is asynchronous on it and never
 *     synchronises (CUDA-graph capturable) except the *_create functions;
 *   - return 0 on success, a DPV_E* code otherwise; dpv_last_error() gives the thread-local text;
 *   - the library allocates only inside the opaque handles (LcDevice, LcPlanner).
 * File:line citations name the reference function each entry point replaces
 * (paths relative to the reference repository root).
 */
#ifndef DPV_B200_H_
#define DPV_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DPV_OK 0
#define DPV_EINVAL 1   /* bad shape / argument (reference: Python Exception / std::invalid_argument) */
#define DPV_ECUDA 2    /* CUDA runtime error */
#define DPV_EUNSUPPORTED 3

int dpv_version(void);
const char* dpv_last_error(void);
/* Number of kernels this library has launched in the calling process (bench.py: gpu_launches). */
long long dpv_launch_count(void);

/* ---------------------------------------------------------------------------------------------
 * (a) Plane sweep.  warping/homography.py:98-135 est_swp_volume_v4 (+ _back_warp_homo_parallel
 * :170-198, img_dis_L2_pard :80-86), batched over B frames (the reference loops, models/models.py:617).
 *   ref [B,C,h,w]  src [B,V,C,h,w]  K [B,3,3]  R [B,V,3,3]  t [B,V,3]  rays [B,3,h*w] (or [1,3,h*w]
 *   with rays_batch_stride = 0)  center [B,2] = (cx,cy) of the numpy intrinsics (homography.py:194)
 *   d_candi [D]  ->  cost [B,D,h,w] = sum_v sum_c dist(bilinear(src_v, H_d p), ref) / sigma
 *   metric: 0 = 'L2', 1 = 'L1'.
 */
int dpv_plane_sweep(const float* ref, const float* src, int B, int V, int C, int h, int w,
                    const float* K, const float* R, const float* t, const float* rays,
                    long long rays_batch_stride, const float* center, const float* d_candi, int D,
                    float sigma, int metric, float* cost, void* stream);

/* The same with explicit element strides between batch items, for inputs that are views into larger tensors: the
 * reference takes both views out of one [B,T,C,h,w] feature tensor (models/models.py:607-617: ref = [:, -1],
 * src = [:, :-1]) and the poses out of [B,T,3,3] / [B,T,3].  Within an item the layout is dense. */
int dpv_plane_sweep_strided(const float* ref, long long ref_batch_stride, const float* src, long long src_batch_stride,
                            int B, int V, int C, int h, int w, const float* K, const float* R, long long R_batch_stride,
                            const float* t, long long t_batch_stride, const float* rays, long long rays_batch_stride,
                            const float* center, const float* d_candi, int D, float sigma, int metric, float* cost,
                            void* stream);

/* warping/homography.py:137-168 warp_feature: out[b,v,i] = warp of channel i of view v onto plane i.
 *   src [B,T,C,h,w] with C >= D  ->  out [B,T,D,h,w]   (reference requires C == D) */
int dpv_warp_feature(const float* src, int B, int T, int C, int h, int w, const float* K,
                     const float* R, const float* t, const float* rays, long long rays_batch_stride,
                     const float* center, const float* d_candi, int D, float* out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * (b) DPV normalisation + statistics in one pass.  F.log_softmax(x [+ add], dim=1) at
 * models/models.py:351,435,590,671,743,777,816; expectation utils/img_utils.py:197-215
 * (dpv_to_depthmap); variance trainer/default_trainer.py:446-450; entropy (north_star);
 * temporal carry F.interpolate(.., 0.25, 'nearest') trainer/default_trainer.py:284,289.
 *   x, add(nullable) [B,D,H,W] -> logp [B,D,H,W]; mean,var,entropy [B,H,W] (each nullable);
 *   carry (nullable) [B,D,H/4,W/4].
 */
int dpv_logsoftmax_stats(const float* x, const float* add, int B, int D, int H, int W,
                         const float* d_candi, float* logp, float* mean, float* var, float* entropy,
                         float* carry, void* stream);

/* utils/img_utils.py:197-215 dpv_to_depthmap, batched: mean[b,y,x] = sum_d d_candi[d] * p. */
int dpv_expectation(const float* dpv, int B, int D, int H, int W, const float* d_candi, int bv_log,
                    float* mean, void* stream);

/* utils/img_utils.py:408-436 spread_dpv_hack: N-tap box SUM along D (zero pad) then / sum_d. */
int dpv_spread(const float* p, int B, int D, int H, int W, int N, float* out, void* stream);

/* utils/img_utils.py:438-449 upsample_dpv: bilinear along D (align_corners=False) + renormalise. */
int dpv_resample_depth(const float* dpv, int B, int D, int H, int W, int Dout, int bv_log, float* out,
                       void* stream);

/* models/models.py:1096-1122 (and ros/ros_all.py:169-203): clamp(exp(logp)) -> for each likelihood
 * volume: NaN->0, clamp, multiply, renormalise -> spread x spread_iter -> clamp -> log.
 *   logp_in [B,D,H,W]; liks [B,nlik,D,H,W]; -> logp_out [B,D,H,W], mean_out (nullable) [B,H,W]. */
int dpv_bayes_fuse(const float* logp_in, const float* liks, int nlik, int B, int D, int H, int W,
                   const float* d_candi, int spread_iter, int spread_n, float* logp_out, float* mean_out,
                   void* stream);

/* ---------------------------------------------------------------------------------------------
 * (d1) DPV -> top-down uncertainty field.  utils/img_utils.py:475-565 gen_ufield
 * (+ convert_flowfield :343-349, depth_to_pts :269-293).
 *   dpv [B,D,H,W], mean [B,H,W] (= dpv_expectation of the same dpv), K_up [B,3,3]
 *   mask (nullable) [B,H,W]  ->  uf [B,D,W] (NaN where a column has no band pixel),
 *   depth_zero (nullable) [B,H,W].
 */
typedef struct DpvUfieldParams {
  int pshift;          /* img_utils.py:485/491 */
  float ystart, yend;  /* camera-Y band ("zstart","zend" in the reference) */
  float maxd, mind;
  int quash;           /* img_utils.py:532-539 */
  int bv_log;
} DpvUfieldParams;
/* scratch: device memory of at least dpv_ufield_scratch_bytes(B,H,W) bytes, 4-byte aligned, owned by the caller
 * (band-mask words, per-column counts and minima; contents are not needed after the call). */
size_t dpv_ufield_scratch_bytes(int B, int H, int W);
int dpv_ufield(const float* dpv, const float* mean, const float* mask, int B, int D, int H, int W,
               const float* d_candi, const float* K_up, const DpvUfieldParams* prm /*host*/, float* uf,
               float* depth_zero, void* scratch, size_t scratch_bytes, void* stream);
/* dpv_ufield for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.
 * Same inputs, outputs and results as dpv_ufield, but the scratch is a 1 MiB (DPV_UFIELD_STATIC_MAX_BYTES) file-scope
 * device array inside the library -- every GPU that loads the library holds it until the process exits -- and the call
 * allocates nothing (no cudaMalloc / cudaFree, no allocating library routine).  Shapes with
 * dpv_ufield_scratch_bytes(B,H,W) > DPV_UFIELD_STATIC_MAX_BYTES are rejected with DPV_EINVAL (C2, 8 x 256 x 768, needs
 * 240 KiB; use dpv_ufield with caller scratch beyond the limit).  The device to run on must be current; the array's
 * device address is resolved once per device on the first call there (make that call outside stream capture).
 * Every call on a device shares the one array: it keeps the last call's contents (each call resets what it needs
 * zeroed), and calls that overlap on the GPU -- different streams, or graphs replayed concurrently -- race. */
#define DPV_UFIELD_STATIC_MAX_BYTES (1u << 20)
int dpv_ufield_static(const float* dpv, const float* mean, const float* mask, int B, int D, int H, int W,
                      const float* d_candi, const float* K_up, const DpvUfieldParams* prm /*host*/, float* uf,
                      float* depth_zero, void* stream);

/* ---------------------------------------------------------------------------------------------
 * (d2) Planner input.  lc/light_curtain.py:646-708 plan_default up to the planner calls
 * (FieldWarp.preprocess :79-85, create_mean_kernel :213-218, normalize :207-210, FieldWarp.warp
 * :47-58, invert :221-225).
 *   uf [B,D,W] (not modified; the reference edits it in place), flow [R,W,2] (polar z->range flow),
 *   steps host [nsteps]  ->  umaps [B,1+2*nsteps,R,W], argmax [B,W], field_pre (nullable) [B,R,W].
 */
int lc_plan_prepare(const float* uf, int B, int D, int W, int R, const float* flow, const float* steps,
                    int nsteps, float* umaps, int32_t* argmax, float* field_pre, void* stream);
/* The planning field alone (plan_sweep / plan_empty lc/light_curtain.py:396-503, plan_m1 :505-552, which do not run
 * the DP on plan_default's three maps): uf [B,D,W] -> field_range [B,R,W] (polar-warped), field_pre [B,R,W] (before the
 * warp).  LC_FIELD_DEFAULT reproduces plan_default's field.  Like the reference's 5x5 smoothing kernel (zeros
 * outside its middle column, 0 * NaN = NaN), a column within two of one that still holds a NaN or an infinity turns
 * NaN; without LC_FIELD_NORMALIZE that is applied to the whole neighbouring column, not only to the 5x5 window. */
#define LC_FIELD_NAN_COLUMNS 1 /* a column holding a NaN becomes (eps,...,eps,1): :658-662, :517-521 */
#define LC_FIELD_NORMALIZE 2   /* per-column min-max: :207-210 */
#define LC_FIELD_NAN_TO_EPS 4  /* :681 */
#define LC_FIELD_DEFAULT 7
int lc_plan_field(const float* uf, int B, int D, int W, int R, const float* flow, int flags, float* field_range,
                  float* field_pre, void* stream);
/* The same preparation for an LC camera posed or sized differently from the RGB camera (lc/light_curtain.py:665-685
 * with the :675-676 branch): uf [B,D,Win] is resized on BOTH axes to [R,Wout] (FieldWarp.preprocess :79-85), smoothed,
 * warped RGB -> LC by flow_T [R,Wout,2] (FieldWarp.transformZTheta :118-155 / :178-187; NULL skips it), then
 * normalised / NaN-filled per `flags` (LC_FIELD_*), polar-warped by flow [R,Wout,2] and split like lc_plan_prepare.
 * nsteps = 0: umaps [B,1,R,Wout] is the polar field alone (argmax unused).  scratch: 2*B*R*Wout floats (device). */
int lc_plan_prepare_transformed(const float* uf, int B, int D, int Win, int R, int Wout, const float* flow_T,
                                const float* flow, const float* steps, int nsteps, int flags, float* umaps,
                                int32_t* argmax, float* field_pre, float* scratch, void* stream);

/* ---------------------------------------------------------------------------------------------
 * lcsim device model.  external/lcsim/python/sim.py:29-211 LCDevice,
 * external/lcsim/lcsim/src/dprocessor.cpp:41-235 (setSensors, createNormalMap, computeLaserParams).
 * All arguments host; K row-major 3x3, lTc row-major 4x4 (camera -> laser).
 */
typedef struct LcDevice LcDevice;
int lc_device_create(int W, int H, const float* K, const float* lTc, float fov, float laser_timestep,
                     float laser_limit, float thickness, float divergence, int hit_mode, float limit,
                     LcDevice** out);
void lc_device_destroy(LcDevice* dev);
/* host copies of per-column camera angles [W] (degrees) and the 1000-entry hit LUT */
int lc_device_angles(const LcDevice* dev, float* out_host);
int lc_device_hit_lut(const LcDevice* dev, float* out_host);

/* external/lcsim/lcsim/src/algo.cpp:696-741 Algo::generateCameraAngles (zero distortion). host in/out */
int lc_camera_angles(const float* K, int W, int H, float* out_host);

/* pylc.fitBSpline = Algo::fitBSpline (external/lcsim/lcsim/src/algo.cpp:449-589; lc/light_curtain.py:580-582, plan_m1):
 * the cubic curve through n_pts control points pts host[n_pts,3] = (x, y, weight), sampled every 0.01 units
 * -> out host[*n_out,2].  Host only.  out = NULL asks for the number of points (*n_out). */
int lc_fit_bspline(const float* pts, int n_pts, float* out, long long cap, long long* n_out);

/* (d3) DP planner.  external/lcsim/lcsim/src/planning.cpp:14-126 (+ Trajectory :162-217),
 * python/planner.py:162-189 PlannerRT.  ranges host [R]. */
typedef struct LcPlanner LcPlanner;
int lc_planner_create(const LcDevice* dev, const float* ranges, int R, LcPlanner** out);
void lc_planner_destroy(LcPlanner* pl);
long long lc_planner_edge_count(const LcPlanner* pl);
/* umaps [n,R,W] -> ks [n,W] (range index per ray, -1 past the end), xz [n,W,2], counts [n] (design points of the
 * path; -1 = the solve of that map was aborted on the device because its producer CTA never delivered, i.e. a fault or
 * a failed launch: the wait is bounded instead of hanging the GPU).  Solves on one stream are ordered and share a log
 * inside the handle; solves on different streams use separate logs and may overlap.  A log grows on the first solve of
 * a given size on its stream: not while that stream is being captured (DPV_EINVAL; run one solve of that size first). */
int lc_planner_solve(LcPlanner* pl, const float* umaps, int n, int32_t* ks, float* xz, int32_t* counts,
                     void* stream);
/* How many maps one producer / consumer CTA pair solves (no reference counterpart: a scheduling choice of this
 * library; results are identical).  1 (default): one map per pair, lowest latency of a solve (n maps hold 2n SMs).
 * 2: two maps per pair, each CTA running two independent teams of 512 threads -- n maps hold n SMs instead of 2n for
 * a longer solve (C2 on a B200: 1.15 ms instead of 0.71 ms).  Measured in the bench's throughput mode it does not pay
 * (a 1024-thread CTA takes its SM's whole register file, so nothing else runs beside it: -1 to -5 %); it is there for
 * callers that must leave SMs free.  Falls back to 1 where two teams' shared memory does not fit an SM. */
int lc_planner_set_pack(LcPlanner* pl, int pack);

/* (d4) Curtain image formation.  dprocessor.cpp:995-1067 processPointsT = checkPoints :238-280 ->
 * findCameraIntersectionsOpt :514-634 -> calculateSurface :735-925 -> computeDepthHits :942-993.
 *   xz [n,npts,2] design points, counts [n] (valid points per curtain; nullable = all npts)
 *   -> planes [n,W,12] camera-frame plane coefficients (main, +divergence, -divergence),
 *      design (nullable) [n,W,4].
 */
int lc_curtain_planes(const LcDevice* dev, const float* xz, const int32_t* counts, int n, int npts,
                      float* planes, float* design, void* stream);
/* LCDevice.get_return (sim.py:168-211): curtains n = B*nc, depth [B,H,W], noise (nullable) [n,H,W]
 *   -> image [n,H,W,4] (x,y,z,intensity 0..255; NaN = no return), thick [n,H,W]. */
int lc_sense_image(const LcDevice* dev, const float* planes, const float* depth, const float* noise,
                   int B, int nc, float* image, float* thick, void* stream);
/* LightCurtain.sense_high tail (lc/light_curtain.py:847-888: mapping :228-244, mixed_model :247-252,
 * utils/img_utils.py:176-195)  -> lik [n,D,H,W]. */
int lc_sense_likelihood(const LcDevice* dev, const float* planes, const float* depth, const float* noise,
                        int B, int nc, const float* d_candi, int D, float* lik, void* stream);
/* One LC iteration's sensing + Bayesian update fused (models/models.py:1064-1122): no likelihood
 * volume is materialised.  logp_in [B,D,H,W], depth [B,H,W], planes [B,nc,W,12]
 *   -> logp_out [B,D,H,W] (may alias logp_in), mean_out (nullable) [B,H,W].
 * flags: 0 = log-probability in and out (the reference's format); LC_FUSE_IN_PROB / LC_FUSE_OUT_PROB keep the
 * volume as clamped probability between the iterations of one lc_process call (no exp/log per bin);
 * LC_FUSE_IN_CLAMPED (with LC_FUSE_IN_PROB) promises that the input is a previous call's LC_FUSE_OUT_PROB output,
 * i.e. already inside [eps, 1], so the entry clamp of models.py:1097 is the identity and is not evaluated. */
#define LC_FUSE_IN_PROB 1
#define LC_FUSE_OUT_PROB 2
#define LC_FUSE_IN_CLAMPED 4
/* with LC_FUSE_OUT_PROB: store the normalised probabilities without the final clamp of models.py:1119, for a call
 * that hands the volume to another call with more curtains of the same iteration (that call takes it with
 * LC_FUSE_IN_PROB | LC_FUSE_IN_CLAMPED, i.e. without an entry clamp either; spread_iter goes to the last call) */
#define LC_FUSE_OUT_RAW 8
int lc_sense_fuse(const LcDevice* dev, const float* logp_in, const float* depth, const float* planes,
                  const float* noise, int B, int nc, const float* d_candi, int D, int spread_iter,
                  int spread_n, int flags, float* logp_out, float* mean_out, void* stream);

/* ---- real-sensor steps either side of the fusion -------------------------------------- */

/* pylc.transformPoints = Depth::transformPoints (external/lcsim/lcsim/src/depth.cpp:143-222; bound at
 * pylc_lib.cpp; called by lc/light_curtain.py:751,:831 and transform_measurement :1076-1089): n point sets
 * pts [n,N,4] (x,y,z,intensity in the light-curtain camera), thick [n,N], intr host[3x3] (target camera), M host[4x4]
 * (source -> target camera)  ->  maps [n,3,height,width] = (depth, intensity, thickness) of the nearest point per
 * pixel (0 where none), pixels failing the `filtering`-neighbourhood outlier test zeroed (0 = no filter).
 * scratch: device memory of lc_transform_points_scratch_bytes(...) bytes, 8-byte aligned, owned by the caller. */
size_t lc_transform_points_scratch_bytes(int n, int width, int height, int filtering);
int lc_transform_points(const float* pts, const float* thick, int n, int N, const float* intr, const float* M,
                        int width, int height, int filtering, float* maps, void* scratch, size_t scratch_bytes,
                        void* stream);
/* F.interpolate(x, scale_factor=scale, mode='nearest') for an integer factor (the ground-truth depth of mode "high",
 * models/models.py:986): in [n,h,w] -> out [n,h*scale,w*scale], out(y,x) = in(y/scale, x/scale). */
int dpv_upsample_nearest(const float* in, int n, int h, int w, int scale, float* out, void* stream);
/* img_utils.depth_to_pts (utils/img_utils.py:269-293) in the point layout lc_transform_points takes
 * (lc/light_curtain.py:826-830 concatenates a row of ones and transposes): depth [n,H,W], intr host[3x3]
 *   -> pts [n,H*W,4] = ((u - cx) / fx * d, (v - cy) / fy * d, d, w4). */
int lc_depth_to_points(const float* depth, int n, int H, int W, const float* intr, float w4, float* pts, void* stream);
/* The tail of sense_high / sense_low (lc/light_curtain.py:869-888) for a sensor whose returns had to be re-projected
 * (rTc != I, :850-858): maps [B*nc,3,H,W] = lc_transform_points output in the RGB camera, depth [B,H,W] (the
 * ground-truth depth whose > 0 mask gates the pixel)  ->  lik [B*nc,D,H,W], the same mixture as lc_sense_likelihood. */
int lc_likelihood_maps(const float* maps, const float* depth, int B, int nc, int H, int W, const float* d_candi, int D,
                       float* lik, void* stream);
/* BaseModel.lc_process_real's sampling of a recorded sweep (models/models.py:924-927):
 * sampling_depth [B,nc,H,W], feat_z / feat_int [B,K,H,W] -> out [B,nc,H,W] = feat_int at the slice whose depth is
 * nearest to sampling_depth (torch.argmin semantics: first minimum, NaN first), inds (nullable) [B,nc,H,W]. */
int lc_sweep_sample(const float* sampling_depth, const float* feat_z, const float* feat_int, int B, int nc, int K,
                    int H, int W, float* out, int32_t* inds, void* stream);

/* ---- callers either side of the fused path ----------------------------- */

/* img_utils.gen_dpv_withmask (utils/img_utils.py:567-582): soft label of variance `var` around dmaps where
 * masks = 1, uniform 1/D where masks = 0, clamped to [eps, 1].  dmaps, masks [B,H,W] -> out [B,D,H,W]. */
int dpv_truth_dpv(const float* dmaps, const float* masks, int B, int H, int W, const float* d_candi, int D,
                  float var, float* out, void* stream);
/* img_utils.lc_intensities_to_dist (utils/img_utils.py:144-157), the likelihood of LightCurtain.gen_lc_dpv_true
 * (lc/light_curtain.py:1106-1125): mean_i[d] = exp(-(|d - placement| / inten_sigma)^2) * peak,
 * lik = softmax_d log N(intensity; mean_i[d], noise_sigma).  Per-pixel inputs [npix]; peak may be NULL
 * (peak_const is used).  Outputs are depth-major [D,npix]; mean_out may be NULL. */
int lc_likelihood_true(const float* placement, const float* intensity, const float* inten_sigma, const float* peak,
                       float peak_const, float noise_sigma, long long npix, const float* d_candi, int D,
                       float* mean_out, float* lik_out, void* stream);
/* utils/inverse_warp.py:174-210 (pixel2cam :26-40, cam2pixel :43-69, F.grid_sample align_corners=False).
 * img [B,C,H,W], depth [B,H,W], Kinv [B,3,3] = intrinsics^-1, P [B,3,4] = intrinsics @ pose
 *   -> out [B,C,H,W], valid (nullable) [B,H,W] bytes.  mode 0 bilinear / 1 nearest; padding 0 zeros / 1 border /
 * 2 reflection. */
int dpv_inverse_warp(const float* img, const float* depth, const float* Kinv, const float* P, int B, int C, int H,
                     int W, int mode, int padding, float* out, uint8_t* valid, void* stream);
/* Backward of dpv_inverse_warp: what torch autograd derives in the reference, where inverse_warp is only ever called
 * under autograd (losses/loss_blocks.py:116,135,168-170): grid_sample backward -> cam2pixel :43-69 -> pixel2cam :26-40.
 * grad_out [B,C,H,W] -> grad_img [B,C,H,W], grad_depth [B,H,W], grad_P [B,3,4] (each nullable; grad_img and grad_P are
 * cleared first, then accumulated with atomics).  The gradient w.r.t. the pose vector / 4x4 follows from grad_P through
 * the small pose algebra the caller keeps in torch. */
int dpv_inverse_warp_backward(const float* grad_out, const float* img, const float* depth, const float* Kinv,
                              const float* P, int B, int C, int H, int W, int mode, int padding, float* grad_img,
                              float* grad_depth, float* grad_P, void* stream);
/* models/correlation_package (correlation_cuda.cc:21-38 shapes, correlation_cuda_kernel.cu:44-112 forward):
 * out[b, (tj,ti), y, x] = mean over the kernel window and channels of pad(in1)[.., y1+j, x1+i] * pad(in2)[.., y2+j, x2+i],
 * x1 = x*stride1 + max_displacement, x2 = x1 + ti*stride2.  in1, in2 [B,C,H,W] -> out [B,outC,outH,outW]. */
int dpv_correlation_shape(int H, int W, int pad_size, int kernel_size, int max_displacement, int stride1,
                          int stride2, int* outC, int* outH, int* outW);
int dpv_correlation(const float* in1, const float* in2, int B, int C, int H, int W, int pad_size, int kernel_size,
                    int max_displacement, int stride1, int stride2, float* out, void* stream);
/* correlation_cuda_kernel.cu:116-303 backward (stride1 = 1, the only stride the reference's launch :486 supports):
 * grad_out [B,outC,outH,outW], in1, in2 [B,C,H,W] -> grad_in1, grad_in2 [B,C,H,W]. */
int dpv_correlation_backward(const float* grad_out, const float* in1, const float* in2, int B, int C, int H, int W,
                             int pad_size, int kernel_size, int max_displacement, int stride1, int stride2,
                             float* grad_in1, float* grad_in2, void* stream);
/* depthError (external/deval_lib/src/evaluate_depth.h:20-118): pred, gt [n,H,W] (gt < 0 = invalid)
 *   -> out [n,10]: mae, rmse, inverse mae, inverse rmse, log mae, log rmse, scale invariant log, abs relative,
 *      squared relative, number of valid pixels. */
int dpv_depth_errors(const float* pred, const float* gt, int n_images, int H, int W, float* out, void* stream);
/* img_utils.compute_unc_rmse (utils/img_utils.py:371-382): fields [D,W] (probabilities) -> out[1]. */
int dpv_unc_rmse(const float* field_truth, const float* field_pred, int D, int W, const float* d_candi, float* out,
                 void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DPV_B200_H_ */
