/* raygun_b200.h -- C ABI of libraygun_b200.so (sm_100a only).
 *
 * This is the drop-in boundary of the B200-native hot path of htem/raygun: the 3D
 * convolution forward/backward, normalisation, activation, pooling, skip and loss
 * arithmetic that the reference obtains from torch.nn (-> cuDNN / ATen).  The
 * reference has no FFI of its own (it is pure Python); the seam is the set of
 * torch.nn call sites inside raygun.torch.networks / losses / predict.worker.
 * Every entry point below cites the reference call site it replaces
 * (paths relative to /root/reference/src/raygun/torch unless noted).
 *
 * Conventions
 *  - plain pointers + sizes only; all pointers are DEVICE pointers unless named h_*.
 *  - every function returns int: 0 = ok, negative = RG_E_*; text via rg_last_error().
 *  - every launch is asynchronous on the caller's cudaStream_t (passed as void*).
 *  - the library never allocates user-visible memory; the caller owns all tensors.
 *  - tensors are described by rg_tensor5 (logical N,C,D,H,W + element strides), so the
 *    same entry points serve NCDHW fp32 and NDHWC (channels-last-3d) bf16 storage.
 *  - DETERMINISM: no entry point finishes a floating-point reduction with atomics.  Reducing kernels write one
 *    partial row per block into a caller-provided fp32 workspace (`ws` / `workspace`, uninitialised device memory of
 *    the size the matching *_workspace_bytes / rg_reduce_ws_floats query returns) and a second kernel adds the rows in
 *    a fixed order: same inputs -> bit-identical outputs, run after run (the reference's CPU path has that property).
 */
#ifndef RAYGUN_B200_H
#define RAYGUN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RG_VERSION 201

enum { RG_OK = 0, RG_E_BADARG = -1, RG_E_UNSUPPORTED = -2, RG_E_CUDA = -3, RG_E_ARCH = -4 };
enum { RG_F32 = 0, RG_BF16 = 1, RG_U8 = 2 };
enum { RG_ACT_NONE = 0, RG_ACT_RELU = 1, RG_ACT_LEAKY = 2, RG_ACT_TANH = 3, RG_ACT_SIGMOID = 4 };
enum { RG_PAD_ZEROS = 0, RG_PAD_REFLECT = 1, RG_PAD_REPLICATE = 2 };
enum { RG_NORM_INSTANCE = 0, RG_NORM_BATCH = 1 };
enum { RG_LOSS_MSE_CONST = 0, RG_LOSS_SMOOTHL1 = 1, RG_LOSS_L1 = 2, RG_LOSS_MEAN = 3 };

/* Logical (N,C,D,H,W) tensor view: sizes + strides in ELEMENTS + dtype. */
typedef struct rg_tensor5 {
  int64_t n, c, d, h, w;
  int64_t sn, sc, sd, sh, sw;
  int32_t dtype; /* RG_F32 | RG_BF16 | RG_U8 */
  int32_t _pad;
} rg_tensor5;

/* Geometry of one Conv3d / ConvTranspose3d (zero padding only: reflect / replicate
 * shells are materialised by rg_pad3d, as the reference does with F.pad / *Pad3d). */
typedef struct rg_conv_geom {
  int32_t kernel[3];  /* kD,kH,kW */
  int32_t stride[3];
  int32_t pad[3];     /* zero padding, per side */
  int32_t groups;     /* UNet.py:75-83 x_init_map uses groups=min(Cin,Cout) */
  int32_t act;        /* fused epilogue activation RG_ACT_* (fprop only) */
  float slope;        /* LeakyReLU negative slope */
} rg_conv_geom;

/* ---- library ------------------------------------------------------------------- */
int rg_version(void);
const char* rg_last_error(void);          /* thread-local message of the last failure */
int rg_device_check(int device);          /* RG_E_ARCH unless the device is sm_100   */

/* ---- direct (CUDA-core, fp32-accumulate) convolutions ---------------------------
 * Replace nn.Conv3d / nn.ConvTranspose3d forward+backward for ANY shape
 * (UNet.py:61-69,75-83,150-165,265-269; ResNet.py:271-338,407-416;
 *  NLayerDiscriminator.py:137-174; systems/BaseSystem.py:209-224).  Weights are the
 * checkpoint-layout fp32 tensors: Conv3d [Cout,Cin/g,kD,kH,kW]; ConvTranspose3d is
 * expressed through the same three kernels (its forward is dgrad of the conv whose
 * weight it shares: [Cin_T,Cout_T,k...] == [Cout_conv,Cin_conv,k...]).
 * x / y / dy / dx may be f32 or bf16 in any strided layout; accumulation is fp32. */
int rg_conv3d_direct_fprop(const rg_tensor5* xd, const void* x, const float* w, const float* bias,
                           const rg_tensor5* yd, void* y, const rg_conv_geom* g, void* stream);
int rg_conv3d_direct_dgrad(const rg_tensor5* dyd, const void* dy, const float* w, const float* bias,
                           const rg_tensor5* dxd, void* dx, const rg_conv_geom* g, void* stream);
/* dw (fp32, [Cout,Cin/g,kD,kH,kW]) is overwritten, or added to when accumulate != 0; dbias (may be NULL) is overwritten;
 * ws = rg_conv3d_direct_wgrad_workspace_bytes(...) bytes (split-K partial rows). */
size_t rg_conv3d_direct_wgrad_workspace_bytes(const rg_tensor5* xd, const rg_tensor5* dyd, const rg_conv_geom* g);
int rg_conv3d_direct_wgrad(const rg_tensor5* xd, const void* x, const rg_tensor5* dyd, const void* dy,
                           float* dw, float* dbias, const rg_conv_geom* g, int accumulate, float* ws, void* stream);

/* Weight gradient of small-reduction convolutions, Cin * kD * kH * kW <= 64 (smallk_wgrad.cu): the 1 -> ngf stem of a
 * U-Net (UNet.py:61-69) and the first, stride-2 convolution of NLayerDiscriminator (NLayerDiscriminator.py:137-145).
 * dy bf16, channels innermost, Cout % 64 == 0, strides multiples of 16 bytes; x bf16 or f32, any strides; zero padding,
 * any stride.  dw fp32 [Cout,Cin,kD,kH,kW] overwritten, or added to when accumulate != 0; dbias (may be NULL) overwritten;
 * ws = rg_smallk_wgrad_workspace_bytes(Cout, Cin*kD*kH*kW) bytes of partial rows (ordered two-stage sum).          */
int rg_smallk_wgrad_supported(const rg_tensor5* xd, const rg_tensor5* dyd, const rg_conv_geom* g);
size_t rg_smallk_wgrad_workspace_bytes(int cout, int k);
int rg_smallk_wgrad(const rg_tensor5* xd, const void* x, const rg_tensor5* dyd, const void* dy, const rg_conv_geom* g,
                    float* dw, float* dbias, float* ws, int accumulate, void* stream);

/* ---- tcgen05 / TMEM / TMA implicit-GEMM convolutions (bf16 NDHWC, fp32 accumulate) --
 * The performance path for the same call sites when Cin%64==0 and Cout%16==0.
 * A plan owns the CUtensorMap descriptors of one (layer, shape) pair. */
typedef struct rg_plan rg_plan;

typedef struct rg_igemm_desc {
  int32_t mode;        /* 0 = fprop, 1 = dgrad (also ConvTranspose3d forward), 2 = wgrad */
  int32_t n, cin, cout;          /* of the CONVOLUTION (fprop direction)             */
  int32_t in_dhw[3], out_dhw[3]; /* fprop input / output spatial extents             */
  int32_t kernel[3], stride[3], pad[3];
  /* element strides of the bf16 NDHWC buffers (channel stride must be 1)            */
  int64_t x_strides[4];  /* n,d,h,w of the fprop-input-side tensor  (x or dx)        */
  int64_t y_strides[4];  /* n,d,h,w of the fprop-output-side tensor (y or dy)        */
  int32_t act; float slope;      /* fused epilogue activation (fprop / convT fwd)    */
  int32_t has_bias, want_stats;  /* epilogue: +bias[c]; accumulate sum / sumsq [n,c] */
  int32_t out_f32;               /* write fp32 instead of bf16 (unused: reserved)    */
} rg_igemm_desc;

int rg_plan_create(const rg_igemm_desc* desc, rg_plan** out);
int rg_plan_destroy(rg_plan* plan);
/* bytes of `workspace` rg_igemm_run (only when statistics are requested) / rg_igemm_wgrad need for this plan */
size_t rg_plan_workspace_bytes(const rg_plan* plan);
/* Re-pack fp32 checkpoint weights [Cout,Cin,kD,kH,kW] into the bf16 K-major layouts the
 * plans read: fprop pack [Cout][tap][Cin]; dgrad pack [Cin][tap][Cout].             */
int rg_pack_weights(const float* w, int cout, int cin, int taps, void* w_fprop_bf16, void* w_dgrad_bf16,
                    void* stream);
/* in -> out of the plan's mode.  fprop: in = x, out = y.  dgrad: in = dy, out = dx.
 * stats_sum/sumsq: fp32 [n*cout] outputs (overwritten) or NULL; they need `workspace`. */
int rg_igemm_run(const rg_plan* plan, const void* in, const void* w_packed, const float* bias, void* out,
                 float* stats_sum, float* stats_sqsum, void* workspace, void* stream);
/* wgrad: dw fp32 [Cout,Cin,kD,kH,kW] overwritten, or added to when accumulate != 0 (gradient buckets);
 * dbias must be NULL (rg_norm_stats(kind = batch) on dy produces it).                 */
int rg_igemm_wgrad(const rg_plan* plan, const void* x, const void* dy, float* dw, float* dbias,
                   void* workspace, int accumulate, void* stream);

/* ---- slice-ring convolution (stride 1; 64-channel produced blocks; ring_conv_sm100.cu) -------------------
 * Same role as rg_igemm_run for fprop (ConvPass / ResnetGenerator3D head, UNet.py:52-87, ResNet.py:324-327) and
 * stride-1 dgrad, for layers where the plain implicit GEMM is L2-feed bound.  The weights are packed per
 * (n-tile, k-chunk, kh, kw) as [kd reversed][64 produced][64 reduction] bf16 by rg_ring_pack_weights
 * (mode 0 = fprop, 1 = dgrad: taps flipped, channel roles swapped); same element count as the fp32 weights.
 * rg_ring_supported: 1 when rg_ring_run accepts `desc`.  rg_ring_config: {S, lag, slice buffers, weight stages,
 * segment length, segments, smem bytes, slice pitch} the host picks (introspection for tests / docs).        */
int rg_ring_supported(const rg_igemm_desc* desc);
int rg_ring_config(const rg_igemm_desc* desc, int* out8);
int rg_ring_pack_weights(const float* w, int cout, int cin, const int* kernel3, int mode, void* out_bf16, void* stream);
size_t rg_ring_workspace_bytes(const rg_igemm_desc* desc);     /* needed only when statistics are requested */
int rg_ring_run(const rg_igemm_desc* desc, const void* in, const void* w_ring, const float* bias, void* out,
                float* stats_sum, float* stats_sqsum, void* workspace, void* stream);
/* The generator's last convolution with the network head folded into its epilogue (BaseSystem.py:209-224:
 * Sequential(net, Conv3d(ngf -> output_nc, 1), final_activation); UNet.py:52-87 for the producing ConvPass layer):
 *   head_out[n, o, d, h, w] = head_act(head_b[o] + sum_c head_w[o][c] * bf16(act(conv(in) + bias)[n, c, d, h, w]))
 * `desc` describes the producing layer (mode 0, cout == 64, its activation in desc->act); its 64-channel output is
 * never written.  head_w fp32 [head_n][64] (16-byte aligned), head_b fp32 [head_n] or NULL, 1 <= head_n <= 4;
 * head_out bf16, channels innermost, element strides head_strides = {n, d, h, w}.  Inference only.            */
int rg_ring_run_head(const rg_igemm_desc* desc, const void* in, const void* w_ring, const float* bias,
                     const float* head_w, const float* head_b, int head_n, int head_act, float head_slope,
                     void* head_out, const int64_t* head_strides, void* stream);

/* ---- small-K stem convolution (stem_conv_sm100.cu) ---------------------------------------------------------
 * Forward of a stride-1 Conv3d whose whole reduction Cin*kD*kH*kW is <= 64 (the 1 -> ngf 3^3 convolution that
 * opens a U-Net, UNet.py:52-87 with input_nc = 1): im2col rows are gathered straight into the tensor-core operand
 * tile in shared memory, so the 1-channel input and the output cross HBM once.  `x` is bf16 with arbitrary element
 * strides (desc->x_strides = n,d,h,w; x_channel_stride for c); `w64` = bf16 [Cout][64], row co = the flattened
 * [Cin][kD][kH][kW] weights of that output channel, zero padded; `y` bf16, channels innermost.  `pad_mode`
 * (RG_PAD_*): how the desc->pad taps outside the input are read -- zeros, or mirrored / clamped indices, i.e. the
 * reflect / replicate 'same' padding of UNet.py:61-69 without a padded copy of the input (forward only).        */
int rg_stem_conv_supported(const rg_igemm_desc* desc);
int rg_stem_conv_run(const rg_igemm_desc* desc, const void* x, long long x_channel_stride, const void* w64,
                     const float* bias, void* y, int pad_mode, void* stream);

/* ---- ConvTranspose3d, kernel = stride = 2 (upconv2_sm100.cu) -------------------------------------------------
 * Forward of the U-Net up-sampling (`Upsample.up`, UNet.py:265-269) as one GEMM launch: every input voxel produces
 * its 2x2x2 output voxels, D[voxel][(parity class, co)] = X[voxel][ci] . W[ci][(class, co)].  x: bf16 [n][Cin][d,h,w]
 * channels innermost; y: bf16 [n][Cout][2d,2h,2w], channels innermost, any 16-byte aligned strides (e.g. a channel
 * slice of the skip-concat buffer).  w_packed: bf16 [8][Cout][Cin], class = 4*pd + 2*ph + pw, i.e. the
 * nn.ConvTranspose3d weight [Cin][Cout][2][2][2] permuted (2,3,4,1,0).                                          */
typedef struct rg_upconv2_desc {
  int32_t n, cin, cout;
  int32_t in_dhw[3];
  int64_t x_strides[4];  /* n,d,h,w element strides of x */
  int64_t y_strides[4];  /* n,d,h,w element strides of y */
  int32_t act; float slope;
  int32_t has_bias, _pad;
} rg_upconv2_desc;
int rg_upconv2_supported(const rg_upconv2_desc* desc);
int rg_upconv2_run(const rg_upconv2_desc* desc, const void* x, const void* w_packed, const float* bias, void* y, void* stream);

/* ---- halo-resident weight gradient (stride 1; halo_wgrad_sm100.cu) ------------------------------------------
 * Same role as rg_igemm_wgrad (autograd of Conv3d, reference: torch.nn.Conv3d call sites UNet.py:69-87,
 * ResNet.py:262-327) for stride-1 layers: one X plane with its (kh,kw) halo per 128-voxel chunk serves every tap
 * of a pass through row-shifted UMMA descriptors.  desc->mode must be 2.  dw fp32 [Cout,Cin,kD,kH,kW] is overwritten,
 * or added to when accumulate != 0; workspace = rg_halo_wgrad_workspace_bytes(desc) bytes of device scratch (one
 * [tap][ci][co] slab per voxel split, summed in split order).                                                   */
int rg_halo_wgrad_supported(const rg_igemm_desc* desc);
size_t rg_halo_wgrad_workspace_bytes(const rg_igemm_desc* desc);
int rg_halo_wgrad(const rg_igemm_desc* desc, const void* x, const void* dy, float* dw, void* workspace, int accumulate,
                  void* stream);

/* ---- 1x1x1 convolutions onto <= 4 channels (BaseSystem.py:209-224: Conv3d(ngf, output_nc, 1) + Tanh) -------
 * HBM-bound streaming kernels over a bf16 NDHWC x (Cin/8 a power of two <= 32); w fp32 [Cout][Cin].
 * fprop fuses bias + activation; wgrad overwrites dw (adds when accumulate != 0) and dbias, ws =
 * rg_pw_conv_wgrad_workspace_bytes(cin, cout) bytes.                                                       */
int rg_pw_conv_fprop(const rg_tensor5* xd, const void* x, const float* w, const float* bias, int act, float slope,
                     const rg_tensor5* yd, void* y, void* stream);
int rg_pw_conv_dgrad(const rg_tensor5* dyd, const void* dy, const float* w, const rg_tensor5* dxd, void* dx, void* stream);
/* Data gradient of a convolution with <= 4 input channels and any kernel / stride / padding (the discriminator's first
 * layer when the generator loss back-propagates through it, NLayerDiscriminator.py:137-145): dy bf16 channels
 * innermost (Cout % 8 == 0), w fp32 [Cout][Cin][kD][kH][kW], dx any layout / dtype.                               */
int rg_smallcin_dgrad(const rg_tensor5* dyd, const void* dy, const float* w, const rg_tensor5* dxd, void* dx,
                      const rg_conv_geom* g, void* stream);
size_t rg_pw_conv_wgrad_workspace_bytes(int cin, int cout);
int rg_pw_conv_wgrad(const rg_tensor5* xd, const void* x, const rg_tensor5* dyd, const void* dy, float* dw, float* dbias,
                     float* ws, int accumulate, void* stream);

/* ---- data movement: strided copy / cast / crop / concat-slice ------------------------
 * Replaces the reference's center `crop` + torch.cat (UNet.py:327-351), ResNet crop
 * (ResNet.py:436-443), dtype/layout casts and .float() (predict/worker.py:101-107).
 * dst[i] (+)= src[i] over the logical shape of dst (src must have the same shape).  */
int rg_copy5(const rg_tensor5* dd, void* dst, const rg_tensor5* sd, const void* src, int accumulate,
             void* stream);

/* Small-Cin stems (ResNet.py:283 Conv3d(input_nc=1, ngf, 7); UNet.py:61-69 first ConvPass conv):
 * y[n][(ci*KH+kh)*KW+kw][d][h][w] = x[n][ci][d][h+kh-ph][w+kw-pw] (0 outside), channels zero-filled up to
 * yd->c.  Folding the H/W taps into >= 64 channels turns the layer into a (KD,1,1) convolution that the
 * tcgen05 implicit-GEMM kernels run at tensor-core speed.  y must be bf16.                                */
int rg_im2col_hw(const rg_tensor5* xd, const void* x, int kh, int kw, int ph, int pw, const rg_tensor5* yd, void* y,
                 void* stream);

/* ---- padding (nn.ReflectionPad3d / ReplicationPad3d / F.pad; ResNet.py:270-279,392-401,
 *      UNet.py:61-69 padding_mode="reflect") and its adjoint ----------------------------*/
int rg_pad3d_fwd(const rg_tensor5* xd, const void* x, const rg_tensor5* yd, void* y,
                 const int32_t lo[3], int mode, void* stream);
/* dx = fold(dy): the adjoint of rg_pad3d_fwd (every dx element is written; no zero-initialisation needed) */
int rg_pad3d_bwd(const rg_tensor5* dyd, const void* dy, const rg_tensor5* dxd, void* dx,
                 const int32_t lo[3], int mode, void* stream);
/* In-place adjoint of the padding for a padded bf16 NDHWC gradient: folds the shell (lo/hi voxels per side) into the
 * interior voxels it mirrors (reflect) or clamps to (replicate); afterwards the interior view of `g` is the gradient of
 * the pad's input (autograd adjoint of F.pad / padding_mode="reflect", UNet.py:61-69) at O(surface) bytes.          */
int rg_pad3d_fold(const rg_tensor5* gd, void* g, const int32_t lo[3], const int32_t hi[3], int mode, void* stream);
/* In-place fill of the `margin`-voxel shell of a padded bf16 NDHWC buffer whose interior is already written (the
 * producing convolution's epilogue wrote it through strides): the reference's explicit padded copy
 * (F.pad / padding_mode="reflect", UNet.py:61-69, ResNet.py:262-270) reduced to O(surface) bytes.               */
int rg_pad_border_fill(const rg_tensor5* bufd, void* buf, const int32_t margin[3], int mode, void* stream);

/* ---- normalisation (nn.InstanceNorm3d / nn.BatchNorm3d; ResNet.py:236,284,299,334,408,417;
 *      NLayerDiscriminator.py:112,146-168; UNet.py:84-88) --------------------------------
 * stats: sum[g], sqsum[g] (fp32, overwritten) with g = n*C+c (instance) or c (batch);
 * ws = rg_reduce_ws_floats(N, C) fp32 elements of scratch (shared by rg_norm_bwd / rg_act_bwd_bias). */
long long rg_reduce_ws_floats(long long n, long long c);
/* out[g] (+)= sum over r < rows of part[r * groups + g], r ascending (the second stage, for partial rows the caller owns) */
int rg_rows_sum(const float* part, int rows, long long groups, float* out, int accumulate, void* stream);
int rg_norm_stats(const rg_tensor5* xd, const void* x, int kind, float* sum, float* sqsum, float* ws, void* stream);
/* rg_norm_stats without a workspace argument and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  The partial rows live in a library-owned `static __device__` array of
 * RG_NORM_STATS_STATIC_MAX_FLOATS fp32 elements (4 MiB per GPU that loads the library, held until exit); shapes whose
 * rg_reduce_ws_floats(N, C) exceeds that are rejected with RG_E_BADARG and a message (e.g. N = 2: C <= 439).  The device
 * owning `x` must be current.  Every call on a device shares the array (it keeps the last call's contents; nothing in it
 * needs clearing): calls that overlap on the GPU -- different streams without an ordering event -- RACE.  Same results as
 * rg_norm_stats, bit for bit.                                                                                     */
#define RG_NORM_STATS_STATIC_MAX_FLOATS (1 << 20)
int rg_norm_stats_static(const rg_tensor5* xd, const void* x, int kind, float* sum, float* sqsum, void* stream);
/* mean/rstd from sum/sqsum: count elements per group, biased variance, eps inside sqrt. */
int rg_norm_finalize(const float* sum, const float* sqsum, int groups, double count, float eps,
                     float* mean, float* rstd, float* var_out, void* stream);
/* y = act((x-mean)*rstd*gamma+beta (+ residual)); gamma/beta/residual may be NULL.
 * kind selects how mean/rstd are indexed (per (n,c) or per c).                       */
int rg_norm_apply(const rg_tensor5* xd, const void* x, int kind, const float* mean, const float* rstd,
                  const float* gamma, const float* beta, const rg_tensor5* rd, const void* residual,
                  int act, float slope, const rg_tensor5* yd, void* y, void* stream);
/* Backward of y = act(norm(x)): needs x, dy and (for act!=NONE) y.  Two launches inside:
 * reductions (sum g, sum g*xhat per group; dgamma/dbeta) then the dx pass.
 * red: fp32 [2*groups] (overwritten); ws as for rg_norm_stats.  training=0 -> statistics are constants. */
int rg_norm_bwd(const rg_tensor5* xd, const void* x, const rg_tensor5* dyd, const void* dy,
                const rg_tensor5* yd, const void* y, int kind, int training, const float* mean,
                const float* rstd, const float* gamma, int act, float slope, float* red,
                float* dgamma, float* dbeta, const rg_tensor5* dxd, void* dx, float* ws, void* stream);

/* ---- activations (nn.ReLU / LeakyReLU(0.2) / Tanh / Sigmoid; UNet.py:41-47,
 *      NLayerDiscriminator.py:138, BaseSystem.py:175-176) -------------------------------*/
int rg_act_fwd(const rg_tensor5* xd, const void* x, int act, float slope, const rg_tensor5* yd, void* y,
               void* stream);
/* dx = dy * act'(.) evaluated from the OUTPUT y (valid for all four supported kinds). */
int rg_act_bwd(const rg_tensor5* yd, const void* y, const rg_tensor5* dyd, const void* dy, int act,
               float slope, const rg_tensor5* dxd, void* dx, void* stream);

/* dx = dy * act'(y) and bias_sum[c] = sum_{n,voxels} dx in one pass (the activation derivative and the bias gradient of
 * a Conv3d + activation layer read the same dy): dense NDHWC bf16 only (RG_E_UNSUPPORTED otherwise); bias_sum fp32 [C]
 * (overwritten); ws = rg_reduce_ws_floats(1, C) fp32 elements.                                                     */
int rg_act_bwd_bias(const rg_tensor5* yd, const void* y, const rg_tensor5* dyd, const void* dy, int act, float slope,
                    const rg_tensor5* dxd, void* dx, float* bias_sum, float* ws, void* stream);

/* ---- pooling / nearest up-sampling (UNet.py:209-218 MaxPool3d; UNet.py:273 nn.Upsample) --*/
int rg_maxpool3d_fwd(const rg_tensor5* xd, const void* x, const int32_t f[3], const rg_tensor5* yd, void* y,
                     void* stream);
/* dx must be zeroed by the caller when the input is not an exact multiple of f. First max wins. */
int rg_maxpool3d_bwd(const rg_tensor5* xd, const void* x, const rg_tensor5* yd, const void* y,
                     const rg_tensor5* dyd, const void* dy, const int32_t f[3], const rg_tensor5* dxd,
                     void* dx, void* stream);
int rg_upsample_nearest_fwd(const rg_tensor5* xd, const void* x, const int32_t f[3], const rg_tensor5* yd,
                            void* y, void* stream);
int rg_upsample_nearest_bwd(const rg_tensor5* dyd, const void* dy, const int32_t f[3], const rg_tensor5* dxd,
                            void* dx, void* stream);

/* ---- sampling bottleneck (models/CycleModel.py:24-30: F.interpolate(..., mode="trilinear", align_corners=True), down by
 *      scale_factor then back up to the original size) ---------------------------------------------------------------
 * y = trilinear resize of x to y's spatial extents (align_corners = True); bwd = its adjoint as a deterministic gather:
 * dx[i] = sum over the dy voxels that sampled x[i].  Any strides, f32 / bf16.                                      */
int rg_resize_trilinear_fwd(const rg_tensor5* xd, const void* x, const rg_tensor5* yd, void* y, void* stream);
int rg_resize_trilinear_bwd(const rg_tensor5* dyd, const void* dy, const rg_tensor5* dxd, void* dx, void* stream);

/* ---- losses (losses/GANLoss.py:24-25,48-64 MSELoss vs constant label;
 *      losses/LinkCycleLoss.py:22,99-105 SmoothL1Loss / L1Loss), warp-shuffle reductions ----
 * loss[0] = scale_loss * mean(f(a, b|target)); grad_a = scale_grad * d mean/da (may be NULL).
 * b is ignored for MSE_CONST / MEAN (target used instead).  a,b share a logical shape.
 * ws: RG_LOSS_WS_FLOATS fp32 elements (one partial per block).                              */
#define RG_LOSS_WS_FLOATS 1184
int rg_loss_fwd_grad(int kind, const rg_tensor5* ad, const void* a, const rg_tensor5* bd, const void* b,
                     float target, float beta, float scale_loss, float* loss, float scale_grad,
                     const rg_tensor5* gd, void* grad_a, float* ws, void* stream);

/* ---- blockwise prediction arithmetic (predict/worker.py:100-114 and :133-147) ---------- */
/* y = (float(u8) - 0) / 255 * scale + shift */
int rg_u8_normalise(const uint8_t* x, int64_t count, float scale, float shift, int out_dtype, void* y,
                    void* stream);
/* out = uint8(trunc(clamp(x*255, 0, 255))) over a cropped strided view */
int rg_denorm_clamp_u8(const rg_tensor5* xd, const void* x, uint8_t* out, void* stream);

/* Block transfers of the worker loop (predict/worker.py:100 `source.to_ndarray(block.read_roi)`, :160
 * `destination[block.write_roi] = out`): the (d,h,w) = ext sub-block at `origin` of a uint8 (D,H,W) = vol_dhw volume
 * <-> a dense device block, as ONE strided 3-D DMA on `stream`.  h_vol: page-locked host memory (asynchronous) or a
 * device pointer (resident volume).  rg_d2h_block copies the sub-block at blk_origin of a dense device block of
 * extents blk_dhw (the halo is cropped by the copy itself).                                              */
int rg_h2d_block(const uint8_t* h_vol, const int64_t vol_dhw[3], const int64_t origin[3], const int64_t ext[3],
                 uint8_t* d_block, void* stream);
int rg_d2h_block(const uint8_t* d_block, const int64_t blk_dhw[3], const int64_t blk_origin[3], uint8_t* h_vol,
                 const int64_t vol_dhw[3], const int64_t origin[3], const int64_t ext[3], void* stream);

/* ---- optimiser (torch.optim.Adam call sites losses/LinkCycleLoss.py:78,128) -------------
 * One fused pass over a flat fp32 parameter bucket: m,v,p update (no amsgrad, no decay). */
int rg_adam_step(float* p, const float* g, float* m, float* v, int64_t count, float lr, float beta1,
                 float beta2, float eps, int step, float grad_scale, void* stream);
/* The same update with the step counter in device memory (int64; incremented first, then read): the form a captured
 * CUDA graph can replay.  16-byte aligned buckets; grad_scale folds the 1/world of a summed all-reduce.  */
int rg_adam_flat(float* p, const float* g, float* m, float* v, int64_t count, float lr, float beta1,
                 float beta2, float eps, long long* step_dev, float grad_scale, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RAYGUN_B200_H */
