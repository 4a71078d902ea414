/* ppm_b200.h — C ABI of libppm_sm100.so: the B200 (sm_100a) drop-in for the PPMStereo
 * correlation + Pick-and-Play-Memory hot path.
 *
 * The reference (cocowy1/PPMStereo) is pure Python/PyTorch and has no FFI for this path;
 * each entry point below replaces a Python call surface of the reference, cited as
 * file:line under /root/reference.  All pointers are DEVICE pointers unless stated, all
 * tensors are caller-allocated and densely packed as described, outputs are fully
 * overwritten, nothing is retained across calls, no call synchronises the host and every
 * kernel is launched on `stream` (a cudaStream_t passed as void*), so sequences of calls are
 * CUDA-graph capturable.  Every function returns 0 on success or a PPM_ERR_* code;
 * ppm_last_error() then returns a thread-local description.  There is no CPU fallback.
 */
#ifndef PPM_B200_H_
#define PPM_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PPM_ABI_VERSION 1
/* The library is built with -fvisibility=hidden: exactly the functions declared here are exported. */
#define PPM_API __attribute__((visibility("default")))

enum ppm_dtype { PPM_F32 = 0, PPM_BF16 = 1 };

enum ppm_status {
  PPM_OK = 0,
  PPM_ERR_ARG = 1,         /* bad shape / dtype / null pointer / misalignment */
  PPM_ERR_CUDA = 2,        /* a CUDA runtime or driver call failed */
  PPM_ERR_UNSUPPORTED = 3  /* valid request the kernels do not cover (e.g. T > 64) */
};

PPM_API int ppm_abi_version(void);
PPM_API const char* ppm_last_error(void);

/* ---------------------------------------------------------------------------------------
 * K1  correlation pyramid build.
 * Replaces CorrBlock1D.__init__ + CorrBlock1D.corr (models/core/corr.py:56-72, :96-104):
 *   L0[(b,h,i), j] = sum_d fmap1[b,d,h,i] * fmap2[b,d,h,j] / sqrt(D)
 *   L(l+1)[r, j]   = (L(l)[r, 2j] + L(l)[r, 2j+1]) / 2, j < floor(W(l)/2)
 * fmap1 (B,D,H,ld1), fmap2 (B,D,H,ld2): NCHW with row pitch ld >= W (elements), the first W
 * entries of each row valid, dtype `in_dtype`.
 * levels[l], l = 0..n_levels: (B*H*W1, W2 >> l) row-major, dtype `out_dtype`.
 * PPM_BF16 inputs run on tcgen05 kind::f16 (bf16 x bf16 -> fp32); PPM_F32 inputs on
 * kind::tf32.  With out_dtype == PPM_BF16 the result is rounded the way the reference's
 * autocast path rounds it (accumulator -> bf16, / sqrt(D) -> bf16, every level -> bf16).
 * Requires ld % 8 == 0 for bf16 (4 for f32) and 16-byte aligned fmaps (TMA row pitch),
 * W2 >> n_levels >= 1, n_levels <= 4.  `levels` is a HOST array of device pointers.
 * (The library reads no environment variable.  The PPM_CB_* timing ablations behind profiles/r01_k1_ablation.md
 * exist only in an experiments build compiled with -DPPM_EXPERIMENTS; see scripts/.) */
PPM_API int ppm_corr_build(const void* fmap1, const void* fmap2, int in_dtype, int B, int D, int H,
                   int W1, int W2, int ld1, int ld2, void* const* levels, int n_levels,
                   int out_dtype, void* stream);

/* ---------------------------------------------------------------------------------------
 * K2  radius-r correlation lookup.
 * Replaces CorrBlock1D.__call__ + bilinear_sampler (models/core/corr.py:74-94, :10-27):
 *   x_l = (x + flow[b,0,y,x]) / 2^l
 *   out[b, l*(2r+1) + k + r, y, x] = lerp(L_l row (b,y,x), x_l + k), zero outside [0, W_l-1]
 * levels[l], l = 0..n_levels-1 as written by ppm_corr_build (dtype vol_dtype); `levels` is a
 * HOST array of device pointers.  flow (B,2,H,W1) fp32 (only channel 0 is read).
 * out (B, n_levels*(2r+1), H, W1) fp32.  radius <= 4, n_levels <= 4. */
PPM_API int ppm_corr_lookup(const void* const* levels, int vol_dtype, const float* flow, int B, int H,
                    int W1, int W2, int radius, int n_levels, float* out, void* stream);

/* K2b  the same lookup (radius 4, 4 levels, bf16 volume) sampled from LEVEL 0 ONLY: the 80-element level-0 span
 * under the level-3 window of a pixel contains the windows of all four levels; the pooled taps are rebuilt in
 * registers with ppm_corr_build's arithmetic, so the result equals ppm_corr_lookup on a pyramid written by
 * ppm_corr_build bit for bit while touching ~2.2 instead of ~4.6 128-byte lines per pixel (B200 fetches a
 * whole line for every random sector).  level0 (B*H*W1, W2) bf16, 32-byte aligned, W2 % 8 == 0. */
PPM_API int ppm_corr_lookup_l0(const void* level0, const float* flow, int B, int H, int W1, int W2, float* out,
                       void* stream);

/* ---------------------------------------------------------------------------------------
 * K3-pre  QAM query/key frame similarity, once per cascade level (SURVEY.md 8a row a6;
 * PPMStereo.compute_qk_similarity, models/core/ppmstereo.py:397-423).
 *   x_[b,t,:]  = mean over the 128 channels of AdaptiveMaxPool2d((h/4, w/4))(x[b,:,t])   x = query, key
 *   sim[b,i,j] = cos(k_[b,i], q_[b,j]) with F.cosine_similarity's eps = 1e-8 clamp on each norm
 * query, key indexed src[b*sb + c*sc + t*st + y*w + x], dtype `dtype`; pooled_ws: 2*b*T*(h/4)*(w/4) fp32
 * scratch; sim (b,T,T) fp32.  fp32 arithmetic, deterministic reduction order.
 * dtype rule: dtype == PPM_BF16 is the reference's autocast regime (ppmstereo.py:617), where `.mean(dim=1)` of the
 * bf16 pooled tensor returns bf16 (mean is not an autocast-fp32 op) and only F.cosine_similarity up-casts: the
 * pooled channel mean x_ is rounded to bf16 (fp32 accumulation, one rounding).  PPM_F32: nothing is rounded. */
PPM_API int ppm_qk_similarity(const void* query, const void* key, int dtype, int64_t sb, int64_t sc, int64_t st,
                      int b, int T, int h, int w, float* pooled_ws, float* sim, void* stream);

/* ---------------------------------------------------------------------------------------
 * K3a  QAM frame scoring + top-k pick (models/core/ppmstereo.py:501-513, :532-533).
 *   conf[b,j]      = mean_{hw} uncertainty[b*T + j]
 *   score[b,i,j]   = exp(-S[b,i,j] / (sum_j S[b,i,j] + T)) * sim[b,i,j] + conf[b,j]
 *   picked[b,i,:]  = indices of the top_k largest score[b,i,:] in ASCENDING index order
 *                    (ties -> lower index); -1 padded when T < top_k
 *   S[b,i,j]      += 1 for picked j                       (strive_time, in/out)
 *   score_norm[b,i,n] = score[b,i,picked n] / mean_{b,n} score[b,i,picked n]
 * sim, strive, frame_score: (b,T,T) fp32; uncertainty (b*T, hw) dtype unc_dtype;
 * picked (b,T,top_k) int32; score_norm (b,T,top_k) fp32; mask (b,T,T) uint8 (may be NULL).
 * conf_ws: (b*T) fp32 scratch.  T <= 64, top_k <= 16.
 * dtype rule: unc_dtype == PPM_BF16 is the reference's autocast regime, where `uncertainty_score...mean(-1)`
 * (ppmstereo.py:506) returns a bf16 tensor: conf = bf16(sum_fp32 * (1/hw)), promoted back to fp32 by the add at
 * :507.  PPM_F32: conf = sum * (1/hw), unrounded.  Means are sum * (1/n), as ATen's CUDA mean computes them. */
PPM_API int ppm_qam_pick(const float* sim, float* strive, const void* uncertainty, int unc_dtype, int b,
                 int T, int hw, int top_k, float* frame_score, int32_t* picked,
                 float* score_norm, uint8_t* mask, float* conf_ws, void* stream);

/* ---------------------------------------------------------------------------------------
 * K3a, allocation-free variant.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc or provide scratch (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid allocation).  Same arithmetic, arguments and results as ppm_qam_pick, minus `conf_ws`: the
 * frame-confidence scratch is a static device array inside the library (PPM_QAM_STATIC_MAX_FRAMES floats = 16 KB,
 * held by every GPU that loads the library until the process exits).  Calls no cudaMalloc / cudaFree / allocating
 * routine.  b*T > PPM_QAM_STATIC_MAX_FRAMES is rejected with PPM_ERR_UNSUPPORTED and a message naming the limit.
 * All calls on one device share the array (it keeps the last call's contents; nothing in it needs to start at zero:
 * each call overwrites the b*T entries it then reads).  Calls that OVERLAP on the GPU (different streams without
 * ordering) race on it -- serialise them, or use ppm_qam_pick with per-caller conf_ws. */
#define PPM_QAM_STATIC_MAX_FRAMES 4096
PPM_API int ppm_qam_pick_static(const float* sim, float* strive, const void* uncertainty, int unc_dtype, int b,
                 int T, int hw, int top_k, float* frame_score, int32_t* picked,
                 float* score_norm, uint8_t* mask, void* stream);

/* ---------------------------------------------------------------------------------------
 * K3b  operand packing for the memory read (models/core/ppmstereo.py:518-548).
 * Source tensors are indexed src[b*sb + c*sc + t*st + p] (element strides, pixel stride 1),
 * c = 128 channels, dtype src_dtype.
 *   ppm_pack_query : Q[(b*T+t), p, c]            = bf16(query + pe[t,c])            (:518-522)
 *   ppm_pack_key   : K[((b*T+i)*Kp+n), p, c]     = bf16(key[b,c,j,p]*score_norm[b,i,n] + pe[j,c]),
 *                    j = picked[b,i,n]; rows with j < 0 are zero-filled             (:524-548)
 *   ppm_pack_value : V[(b*T+t), c, p_pad]        = bf16(value), p_pad = round_up(hw, 8) (:529)
 * pe: (T,128) fp32. */
PPM_API int ppm_pack_query(const void* query, int src_dtype, int64_t sb, int64_t sc, int64_t st,
                   const float* pe, int b, int T, int hw, void* q_out, void* stream);
PPM_API int ppm_pack_key(const void* key, int src_dtype, int64_t sb, int64_t sc, int64_t st,
                 const float* pe, const int32_t* picked, const float* score_norm, int b, int T,
                 int Kp, int hw, void* k_out, void* stream);
PPM_API int ppm_pack_value(const void* value, int src_dtype, int64_t sb, int64_t sc, int64_t st, int b,
                   int T, int hw, void* v_out, void* stream);

/* ---------------------------------------------------------------------------------------
 * K3c  query-adaptive softmax aggregation over the picked frames + residual
 * (flash_attn_func call at models/core/ppmstereo.py:550, residual :551-555).
 *   hidden[b,i,p,:] = bf16( softmax_{n,p'}(scale * Q[b,i,p] . K[b,i,n,p']) V[picked[b,i,n], :, p'] )
 *   out[(b*T+i), c, p] = mf[b*mb + c*mc + i*mt + p] + beta[0] * hidden[b,i,p,c]
 * Q, K, V as written by the pack calls; mf dtype mf_dtype; beta: device fp32 scalar;
 * out (b*T, 128, hw) fp32.  bf16 operands, fp32 scores/softmax, P rounded to bf16 before
 * P.V (FlashAttention-2 rounding), tcgen05 MMAs with TMEM accumulators. */
PPM_API int ppm_memory_attend(const void* Q, const void* K, const void* V, const int32_t* picked,
                      const void* mf, int mf_dtype, int64_t mb, int64_t mc, int64_t mt,
                      const float* beta, float scale, int b, int T, int Kp, int hw, float* out,
                      void* stream);

/* Row-band form of ppm_memory_attend (optional split of ONE clip over GPUs): only the
 * query pixels [q_begin, q_begin + q_count) of every frame are computed and stored; Q, K, V, mf and
 * out keep the full-frame layouts above, keys and values of ALL hw pixels take part.  A rank that
 * owns a band of image rows calls this after gathering the value bands of the other ranks.
 * ppm_memory_attend == ppm_memory_attend_band(..., 0, hw, ...). */
PPM_API int ppm_memory_attend_band(const void* Q, const void* K, const void* V, const int32_t* picked,
                           const void* mf, int mf_dtype, int64_t mb, int64_t mc, int64_t mt,
                           const float* beta, float scale, int b, int T, int Kp, int hw, int q_begin,
                           int q_count, float* out, void* stream);

/* ---------------------------------------------------------------------------------------
 * K4  convex upsampling of the flow (models/core/ppmstereo.py:185-197, called
 * per iteration at :577).
 *   p[n,k,i,j,y,x]          = softmax_k mask[n, (k*rate + i)*rate + j, y, x]          k = 0..8
 *   out[n,c,rate*y+i,rate*x+j] = sum_k p * rate * flow[n, c, y + k/3 - 1, x + k%3 - 1]   (zero outside)
 * flow (N,2,H,W) fp32; mask (N, 9*rate*rate, H, W) dtype mask_dtype (for PPM_BF16 the softmax result is
 * rounded to bf16 before the product, as torch.softmax does for a bf16 input); out (N,2,rate*H,rate*W)
 * fp32, 16-byte aligned.  rate must be 4. */
PPM_API int ppm_convex_upsample(const float* flow, const void* mask, int mask_dtype, int N, int H, int W,
                        int rate, float* out, void* stream);

/* ---------------------------------------------------------------------------------------
 * K5c  SequenceUpdateBlock3D.forward (models/core/ppmtereo_update.py:971-1003) for a block without update-time /
 * update-space attention and with the 2-D convex mask (PPMStereo's update_block08 / update_block04), one call:
 *   net = gru(net, cat[inp, motion_features, motion_features_global]);  delta_flow = flow_head(net);
 *   mask = 0.25 * mask_2d(net)      (mask_2d = Conv3x3(128,256) -> ReLU -> Conv1x1(256, n_mask), :~905)
 * The 384-channel cat is never materialised (x_parts[3] are packed straight into the GRU's operand) and all tensors
 * are frame-major (b*T, C, H, W) contiguous as the reference holds them outside the unit (no rearrange copies).
 * net (b*T,128,H,W) fp32 or bf16 (type promotion as ppm_sepconv_gru3d); x_parts[i] (b*T,128,H,W), x_dtypes[i].
 * gru_weights / gru_biases, flow_*: as ppm_sepconv_gru3d / ppm_flow_head3d.  mask_w1: bf16 [9][256][128] (tap order
 * ih*3 + iw), mask_b1 fp32 [256]; mask_w2: bf16 [1][192][256] with rows n_mask..191 zero, mask_b2 fp32 [192].
 * Outputs: net_out (b*T,128,H,W) net_dtype; mask_out (b*T,n_mask,H,W) bf16 (the 0.25 included: exact);
 * dflow_out (b*T,2,H,W) bf16.  workspace: ppm_update_block3d_workspace bytes, 256-byte aligned. */
PPM_API int ppm_update_block3d_workspace(int b, int T, int H, int W, size_t* bytes);
PPM_API int ppm_update_block3d(const void* net, int net_dtype, const void* const* x_parts, const int* x_dtypes,
                       const void* const* gru_weights, const float* const* gru_biases, const void* flow_w1,
                       const float* flow_b1, const void* flow_w2, const float* flow_b2, const void* mask_w1,
                       const float* mask_b1, const void* mask_w2, const float* mask_b2, int n_mask, int b, int T,
                       int H, int W, void* workspace, size_t workspace_bytes, void* net_out, void* mask_out,
                       void* dflow_out, void* stream);

/* ---------------------------------------------------------------------------------------
 * K6  lookup fused into the motion encoder's front end.
 * Replaces `cor = F.gelu(self.convc1(corr))` of BasicMotionEncoder_v2.forward (models/core/ppmtereo_update.py:466),
 * convc1 = PCBlock4_Deep_nopool_res(36, 256, k_conv=[1, 7]) (:91-116), together with the CorrBlock1D.__call__ that
 * produced `corr` (models/core/corr.py:74-94), in the bf16 autocast regime: the (B,36,H,W1) fp32 lookup result is
 * never written.  Rounding: 1x1 conv = bf16(bf16(acc) + bias), depthwise conv = bf16(acc + bias), GELU of a bf16
 * tensor = bf16, residual stream fp32 (torch / cuDNN's flow, identified on the B200).
 * level0 / flow / W2: as ppm_corr_lookup_l0 -- or corr != NULL: start from a materialised (B,36,H,W1) fp32 lookup.
 * weights_a: PPM_MF_A_FLOATS fp32 (bf16-representable values): ffn1.0 W[54][36], b[54] (+2 pad), ffn1.2 W[36][56]
 *   (rows padded from 54), b[36], conv_list.0 w[36], b[36].
 * weights_b: PPM_MF_B_FLOATS fp32: conv_list.1 w[36][49], b[36], pw W[36][36], b[36], ffn2.0 W[54][36], b[54] (+2 pad),
 *   ffn2.2 W TRANSPOSED [54][256], b[256].
 * x2_ws: B*36*H*W1 floats of scratch.  cor: (B,256,H,W1) bf16 = gelu(convc1(corr)). */
#define PPM_MF_A_FLOATS 4124
#define PPM_MF_B_FLOATS 19212
PPM_API int ppm_motion_corr_front(const void* level0, const float* flow, const float* corr, const float* weights_a,
                          const float* weights_b, int B, int H, int W1, int W2, float* x2_ws, void* cor,
                          void* stream);

/* ---------------------------------------------------------------------------------------
 * K7  BasicMotionEncoder_v2.forward (models/core/ppmtereo_update.py:458-482) as one call, bf16 autocast regime:
 *   cor = gelu(convc1(lookup))  (K6 + one K = 64 GEMM)          flo = relu(convf2(relu(convf1(flow))))
 *   cor = relu(convc2(cor))                                      out = relu(final_conv(cat[cor, flo, mhs]))
 * convf1 (7x7 on the two flow channels) is an im2col pack + one K = 98 (padded to 128) GEMM; the three 3x3
 * convolutions run on the K5 shifted-GEMM kernel (9 offsets, tcgen05); their producers write one
 * channels-last [B*H*W][320] buffer, so the reference's torch.cat is never executed, and the final convolution's
 * epilogue performs the torch.split: channels 0..125 -> motion_features[:, 0:126] (fp32: the reference's cat with
 * the fp32 flow promotes), channels 126..189 -> mhs_out (bf16).  motion_features[:, 126:128] (= flow) is the
 * caller's to fill.  init_conv (first iteration only) stays with the caller: mhs (B,64,H,W) fp32 or bf16.
 * flow (B,2,H,W) fp32, always required; level0 / corr / W2 / front_a / front_b: as ppm_motion_corr_front
 * (weights_a / weights_b).  wd: bf16 [1][256][64] = convc1.ffn2.2.weight.reshape(256, 54), columns 54..63 zero: the
 * block's last layer and the outer GELU run as a GEMM on the hidden vector K6b leaves channels-last (its bias is read
 * from weights_b).  wf1: bf16 [1][128][128] = convf1.weight.reshape(128, 98), columns 98..127 zero; bf1 fp32 [128].
 * wc2: bf16 [9][192][256], wf2: bf16 [9][64][128], wfin: bf16 [9][192][320] (rows 190, 191 zero), tap = ih*3 + iw;
 * bc2 fp32 [192], bf2 fp32 [64], bfin fp32 [192].  Rounding per convolution: bf16(bf16(acc) + bias), then ReLU.
 * workspace: ppm_motion_encoder_workspace bytes, 256-byte aligned. */
PPM_API int ppm_motion_encoder_workspace(int B, int H, int W, size_t* bytes);
PPM_API int ppm_motion_encoder(const void* level0, const float* flow, const float* corr, int W2, const float* front_a,
                       const float* front_b, const void* wd, const void* wf1, const float* bf1, const void* mhs,
                       int mhs_dtype, const void* wc2, const float* bc2, const void* wf2, const float* bf2, const void* wfin,
                       const float* bfin, int B, int H, int W, void* workspace, size_t workspace_bytes,
                       float* motion_features, void* mhs_out, void* stream);

/* ---------------------------------------------------------------------------------------
 * K8  value projection + uncertainty head: the two lines of the update iteration between the motion encoder and the
 * memory read (models/core/ppmstereo.py:492-495), bf16 autocast regime:
 *   value       = aggregator.to_v(motion_features)           Conv2d(128,128,1,bias=False)   ppmtereo_update.py:646, :948
 *   uncertainty = update_block.uncertainty(cat[net, value])   Conv2d(256,128,3) ReLU Conv2d(128,1,1) Sigmoid   :889-893
 * Three GEMMs on the K5 kernel; the value GEMM's epilogue stores NCHW bf16 (the memory read's V operand, used in place)
 * and channels-last (second operand segment of the 3x3 convolution: torch.cat is never executed).
 * motion_features, net: (B,128,H,W) fp32 or bf16.  wv: bf16 [1][128][128], bv fp32 [128] (zeros: to_v has no bias);
 * wu1: bf16 [9][128][256] (tap = ih*3 + iw; input channels net 0..127, value 128..255), bu1 fp32 [128];
 * wu2: bf16 [1][64][128] (row 0 = the weight, rows 1..63 zero), bu2 fp32 [64].
 * Outputs: value (B,128,H,W) bf16, uncertainty (B,1,H,W) bf16 = bf16(sigmoid(bf16 conv)).
 * workspace: ppm_value_uncertainty_workspace bytes, 256-byte aligned. */
PPM_API int ppm_value_uncertainty_workspace(int B, int H, int W, size_t* bytes);
PPM_API int ppm_value_uncertainty(const void* motion_features, int mf_dtype, const void* net, int net_dtype, const void* wv,
                          const float* bv, const void* wu1, const float* bu1, const void* wu2, const float* bu2, int B,
                          int H, int W, void* workspace, size_t workspace_bytes, void* value, void* uncertainty,
                          void* stream);

/* ---------------------------------------------------------------------------------------
 * K5  SKSepConvGRU3D (models/core/ppmtereo_update.py:254-312, called from
 * SequenceUpdateBlock3D.forward :971-1003), in the reference's bf16 autocast regime:
 *   per stage s = 1 (along W), 2 (along H), 3 (along T):
 *     z = sigmoid(convz_s(cat[h, x])); r = sigmoid(convr_s(cat[h, x])); q = tanh(convq_s(cat[r*h, x]));
 *     h = (1 - z) * h + z * q          (stage 1: convz / convr = Conv(15 taps) -> GELU -> Conv(5 taps))
 *   every elementwise result is rounded to bf16 once and every conv result is bf16(bf16(acc) + bias), as cuDNN's bf16
 *   conv3d + torch's bias add and the elementwise ops of bf16 tensors do in the reference; fp32 accumulation on
 *   tcgen05 tensor cores.
 * h (b,128,T,H,W) and x (b,384,T,H,W) are indexed src[b*sb + c*sc + t*st + y*W + x] (element strides), fp32 or bf16.
 * weights[8]: bf16, packed [tap][Cout][Cin] (Cin order = cat[h, x]):
 *   0: convz1.0 | convr1.0 stacked on Cout (15, 256, 512)   1: convz1.2 (5,128,128)   2: convr1.2 (5,128,128)
 *   3: convq1 (5,128,512)   4: convz2 | convr2 (5,256,512)   5: convq2   6: convz3 | convr3   7: convq3
 * biases[8]: fp32 arrays of Cout values (the bf16-rounded biases), same order.
 *   h_dtype selects the reference's type promotion: a bf16 h makes r*h, (1-z)*h and the final sum bf16 ops; an fp32 h
 *   (what PPMStereo hands over: its SST blocks return fp32) makes (1-z)*h and the sum fp32 ops, so the state keeps fp32
 *   precision from iteration to iteration (the convs still see bf16(h)).  Pass out_dtype = h_dtype for the reference's result.
 * workspace: ppm_sepconv_gru3d_workspace bytes, 256-byte aligned (channels-last bf16 copies of x, h, r*h, z and the
 * stage-1 GELU tensor, plus the fp32 state).  out: (b,128,T,H,W) contiguous, out_dtype. */
PPM_API int ppm_sepconv_gru3d_workspace(int b, int T, int H, int W, size_t* bytes);
PPM_API int ppm_sepconv_gru3d(const void* h, int h_dtype, int64_t hsb, int64_t hsc, int64_t hst, const void* x,
                      int x_dtype, int64_t xsb, int64_t xsc, int64_t xst, const void* const* weights,
                      const float* const* biases, int b, int T, int H, int W, void* workspace,
                      size_t workspace_bytes, void* out, int out_dtype, void* stream);
/* Byte offset, inside a ppm_sepconv_gru3d workspace, of the bf16 channels-last copy [b*T*H*W][128] of the unit's RESULT
 * (valid after the call): ppm_flow_head3d can take it as `x_cl` instead of re-packing the (b,128,T,H,W) tensor. */
PPM_API int ppm_sepconv_gru3d_state_offset(int b, int T, int H, int W, size_t* offset);

/* ---------------------------------------------------------------------------------------
 * K5b  FlowHead3D (models/core/ppmtereo_update.py:670-678): conv2(relu(conv1(x))), two 3x3x3
 * Conv3d (128 -> 256 -> 2) in the bf16 autocast regime (each conv = bf16(bf16(acc) + bias), relu in bf16), on the K5
 * shifted-GEMM kernel.  x (b,128,T,H,W) indexed x[b*sb + c*sc + t*st + y*W + x], fp32 or bf16 -- or x_cl != NULL: the same
 * tensor already as bf16 channels-last [b*T*H*W][128] (then x may be NULL).  w1: bf16 [27][256][128]; b1: fp32 [256];
 * w2: bf16 [1][64][256], row tap*2 + c = conv2.weight[c, :, tap] (54 rows, rows 54..63 zero): the second layer contracts
 * the 256 channels once per position for all 27 x 2 (offset, output) pairs and then adds the 27 shifted fp32 planes;
 * b2: fp32 [64] (entries 2..63 ignored).  Tap order (it*3 + ih)*3 + iw.
 * workspace: ppm_flow_head3d_workspace bytes, 256-byte aligned.  out: (b,2,T,H,W) contiguous, out_dtype. */
PPM_API int ppm_flow_head3d_workspace(int b, int T, int H, int W, size_t* bytes);
PPM_API int ppm_flow_head3d(const void* x, int x_dtype, int64_t xsb, int64_t xsc, int64_t xst, const void* x_cl,
                    const void* w1, const float* b1, const void* w2, const float* b2, int b, int T, int H, int W,
                    void* workspace, size_t workspace_bytes, void* out, int out_dtype, void* stream);


/* Bytes of the packed operands for (b, T, Kp, hw): q, k, v in that order. */
PPM_API int ppm_pack_bytes(int b, int T, int Kp, int hw, size_t* q_bytes, size_t* k_bytes,
                   size_t* v_bytes);

#ifdef __cplusplus
}
#endif
#endif /* PPM_B200_H_ */
