/* alexcap.h -- C ABI of libalexcap_b200.so: the B200-native (sm_100a) caption-decoder hot path of
 * AlexIMOVI/ImageCaptioning (AlexCap LSTM and attention-LSTM decoders + CustomCrossEntropyLoss).
 *
 * The reference is pure Python/PyTorch and has no FFI; the boundary it exposes is the nn.Module
 * surface of AlexCap/LanguageModule.py, AlexCap/AttentionLanguageModule.py and AlexCap/CustomLoss.py.
 * Each entry point below names the reference interface it replaces (paths relative to the reference
 * repository root).  The Python mirror of that surface lives in imagecaptioning_b200/modules.py and
 * binds this header with ctypes (imagecaptioning_b200/_lib.py); INTEGRATION.md shows the stub.
 *
 * Conventions
 *   - plain pointers + sizes only; every buffer a call reads or writes is a DEVICE buffer owned by the caller (torch),
 *     and the library keeps no caller pointer after return.  The training / decode entry points never allocate or free
 *     device memory per call; the only device memory the library owns is one process-lifetime scratch per GPU for the
 *     attention kernels' cross-CTA combine (grown on the first call that needs more, never freed or moved, so captured
 *     graphs stay valid -- do the first call of a shape outside stream capture), and the operand-copy arena of the
 *     test / micro-benchmark hook alexcap_gemm_f32 with engine >= 300 (that hook may synchronise the device when it
 *     grows and is not capturable then).
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*), asynchronously, with no
 *     device-wide synchronisation: a call is capturable into a CUDA graph.
 *   - return 0 on success; <0 on failure (ALEXCAP_E_*), message via alexcap_last_error().
 *   - row-major contiguous tensors; weights in nn.Linear layout [out,in]; LSTM gate blocks stacked
 *     i,f,g,o along dim 0 (torch.nn.LSTMCell / nn.LSTM); fp32 master parameters and fp32 gradients.
 *   - token ids are int32; 0 = padding/ignore, START = V+1, END = V+2, classes Vp = V+3
 *     (AlexCap/AttentionLanguageModule.py:43-44, AlexCap/my_model_preprocess.py:115-134).
 *   - H, D, E must be multiples of 4 (128-bit loads); pointers 16-byte aligned.
 *   - no CPU fallback: a device that is not sm_100 fails with ALEXCAP_E_UNSUPPORTED.
 */
#ifndef ALEXCAP_H_
#define ALEXCAP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ALEXCAP_ABI_VERSION 2

#if defined(__GNUC__)
#define ALEXCAP_API __attribute__((visibility("default")))
#else
#define ALEXCAP_API
#endif

#define ALEXCAP_E_BADARG      (-1)
#define ALEXCAP_E_UNSUPPORTED (-2)
#define ALEXCAP_E_CUDA        (-3)
#define ALEXCAP_E_WORKSPACE   (-4)

#define ALEXCAP_F32  0
#define ALEXCAP_BF16 1   /* bf16 activations / tensor-core operands, fp32 accumulate + fp32 master weights */

typedef struct alexcap_dims {
  int32_t B, L, D, H, E, Vp, S;   /* batch, regions, feature dim, hidden, embed, classes V+3, steps T+1 */
  int32_t num_layers;              /* plain LSTM only (nn.LSTM num_layers); ignored by the attention decoder */
  int32_t dtype;                   /* ALEXCAP_F32 | ALEXCAP_BF16 : arithmetic type of the GEMM operands */
  int32_t need_dfeat;              /* 1: also produce d loss / d features (CNN fine-tuning) */
  int32_t ignore_index;            /* CustomLoss.py:7 -> 0 */
  float   label_smoothing;         /* CustomLoss.py:7 -> 0.1 */
  float   reg_weight;              /* LSTMwAttentionModel.py:70-71 -> 1.0 (0 disables the regulariser) */
  float   dropout_p;               /* AttentionLanguageModule.py:54 -> 0.5 in train(), 0 in eval() */
  uint64_t seed, offset;           /* Philox counter for the dropout mask when drop_mask == NULL and philox_state == NULL */
  uint64_t* philox_state;          /* DEVICE {seed, offset} or NULL.  When set (and drop_mask == NULL) the mask kernel reads
                                      the counter from the device and the call advances `offset` ON the device afterwards, so
                                      a captured CUDA graph draws a fresh mask at every replay (the reference's train() mode:
                                      nn.Dropout() p = 0.5, AttentionLanguageModule.py:54,79). */
} alexcap_dims;

/* Parameters of AlexCap.AttentionLanguageModule.Decoder (AttentionLanguageModule.py:46-58), state_dict order.
 * The same struct type, with every member pointing at an fp32 buffer of the same shape, receives gradients. */
typedef struct alexcap_attn_params {
  float *init_h_w, *init_h_b;     /* [H,D],[H]   */
  float *init_c_w, *init_c_b;     /* [H,D],[H]   */
  float *f_beta_w, *f_beta_b;     /* [D,H],[D]   */
  float *deep_w, *deep_b;         /* [Vp,H],[Vp] */
  float *att_U_w, *att_U_b;       /* [H,H],[H]   */
  float *att_W_w, *att_W_b;       /* [H,D],[H]   */
  float *att_v_w, *att_v_b;       /* [1,H],[1]   */
  float *emb_w;                   /* [Vp,E]      */
  float *lstm_w_ih, *lstm_w_hh;   /* [4H,E+D],[4H,H] */
  float *lstm_b_ih, *lstm_b_hh;   /* [4H],[4H]   */
} alexcap_attn_params;

/* Parameters of AlexCap.LanguageModule.LanguageModel (LanguageModule.py:35-49).  Per-layer LSTM tensors are
 * arrays of num_layers pointers (layer l input size is E for l=0, H above). */
#define ALEXCAP_MAX_LAYERS 8
typedef struct alexcap_lstm_params {
  float *enc_w, *enc_b;           /* image_encoder.encode [E,D],[E] */
  float *emb_w;                   /* lookup_table [Vp,E] */
  float *out_w, *out_b;           /* rnn.linear [Vp,H],[Vp] */
  float *w_ih[ALEXCAP_MAX_LAYERS], *w_hh[ALEXCAP_MAX_LAYERS];   /* [4H,in],[4H,H] */
  float *b_ih[ALEXCAP_MAX_LAYERS], *b_hh[ALEXCAP_MAX_LAYERS];   /* [4H],[4H] */
} alexcap_lstm_params;

ALEXCAP_API int alexcap_abi_version(void);
ALEXCAP_API const char* alexcap_last_error(void);   /* thread-local, static storage */
/* 0 if the current CUDA device is usable (compute capability 10.x), else ALEXCAP_E_UNSUPPORTED / _CUDA. */
ALEXCAP_API int alexcap_device_check(void);
/* Kernel launches enqueued by this library since load (host-side counter; memcpy/memset nodes not counted). */
ALEXCAP_API unsigned long long alexcap_launch_count(void);
/* GEMM engine used by the training entry points below (default 3).  0 = fp32 CUDA-core tiles; 1 = tcgen05/TMEM/TMA tiles
 * in 3xTF32 (operands split into tf32 hi/lo in shared memory, ~2^-21; the decode entry points always use it: exact token
 * ids); 2 = tcgen05 bf16 operands / fp32 accumulate (what dims.dtype = ALEXCAP_BF16 selects for the duration of a call);
 * 3 = tcgen05 bf16x3 (operands split into bf16 hi/lo pairs, three kind::f16 MMAs, ~2^-17: 1e-5-level gradients, 1.3x
 * faster than engine 1).  Inside the training entry points the LARGE GEMMs of engines 2 and 3 (M >= 256, >= 1 GFLOP) run on
 * the pre-split path: each fp32 operand is split once into bf16 hi (+ lo) copies in the workspace and a persistent
 * TMA -> tcgen05 kernel consumes them (gemm_ps.cu; ALEXCAP_GEMM_PRESPLIT=0 disables).
 * Calls whose operands TMA cannot map (ld % 4 != 0) use engine 0.  Process-wide. */
ALEXCAP_API int alexcap_set_gemm_engine(int engine);
ALEXCAP_API int alexcap_get_gemm_engine(void);

/* ------------------------------------------------------------------ attention decoder (train) ----------
 * Workspace (bytes) the three calls below need for `dims`; the caller allocates it once and passes it in.
 * forward() leaves what backward() needs inside the workspace, so the SAME workspace must be handed to the
 * matching backward(). */
ALEXCAP_API size_t alexcap_attn_workspace_bytes(const alexcap_dims* dims);

/* Decoder half of AlexCapModel.forward_train (LSTMwAttentionModel.py:67-72): teacher-forced forward
 * (AttentionLanguageModule.py:62-84), CustomCrossEntropyLoss (CustomLoss.py:9-15), the attention regulariser
 * and the full backward, in one call.  Logits are never handed to the caller.
 *   feat [B,L,D] fp32, gt [B,S-1] int32 (0-padded captions, the reference's data.gt_labels)
 *   grads: every member OVERWRITTEN with d loss/d param;  dfeat [B,L,D] or NULL (needs dims.need_dfeat)
 *   loss_out[4] (device): {loss_ce + reg, loss_ce, reg, n_valid}
 *   alphas_out [B,S,L] or NULL;  drop_mask [S,B,H] (0 or 1/(1-p), time-major) or NULL -> Philox(seed,offset)
 *   ce_denominator (device, 1 float) or NULL: overrides N_valid as the CE divisor (data-parallel global mean). */
ALEXCAP_API int alexcap_attn_train_fused(const alexcap_dims* dims, const alexcap_attn_params* params,
                             const float* feat, const int32_t* gt,
                             const alexcap_attn_params* grads, float* dfeat,
                             float* loss_out, float* alphas_out, const float* drop_mask,
                             const float* ce_denominator,
                             void* workspace, size_t workspace_bytes, void* stream);

/* The same step in two calls, so that a data-parallel caller can start the all-reduce of the vocabulary-projection
 * gradients (the largest tensors, final after phase 1) while phase 2 runs (the reference is
 * single-GPU and has no counterpart).  phase 1: forward + loss + vocab backward; phase 2: reverse-time chain and all
 * other weight gradients (same arguments and workspace as phase 1); phase 3: both (== alexcap_attn_train_fused).
 * Phase 2 may itself be run as phase 4 (reverse-time chain, then the lstm.* and embedding gradients) followed by phase 8
 * (all remaining gradients), so that the second-largest gradient group can be exchanged while phase 8 runs.
 * Phase 1 may itself be run as phase 16 (targets + the teacher-forced forward recurrence) followed by phase 32 (vocabulary
 * projection, loss, vocabulary backward, regulariser): `ce_denominator` is only read by phase 32, so a data-parallel caller
 * can all-reduce N_valid (alexcap_count_valid_targets) on a side stream while phase 16 runs.
 * Phase 32 may itself be run as phase 128 (everything of 32 except d deep_output.weight / .bias, which nothing later in the step
 * reads) and phase 64 (those two gradients, restricted to the SMs the reverse-time loop leaves idle): 64 may be enqueued on a
 * SECOND stream, ordered after 128, concurrently with phase 4 on the first -- a data-parallel caller runs 64 and the all-reduce of
 * deep_output.* on its side stream underneath the reverse loop.  Where the split does not apply (exact-fp32 engines, a batch that
 * fills the GPU) 128 does all of 32 and 64 does nothing.  The one-call forms (phase 3, alexcap_attn_train_fused) fork the
 * same work onto a library-owned stream and join it before returning. */
ALEXCAP_API int alexcap_attn_train_phase(const alexcap_dims* dims, const alexcap_attn_params* params,
                             const float* feat, const int32_t* gt,
                             const alexcap_attn_params* grads, float* dfeat,
                             float* loss_out, float* alphas_out, const float* drop_mask,
                             const float* ce_denominator,
                             void* workspace, size_t workspace_bytes, int32_t phase, void* stream);

/* Decoder.forward, train branch (AttentionLanguageModule.py:62-84): materialises preds [B,S,Vp] and
 * alphas [B,S,L] exactly like the reference, for callers that build their own loss on top. */
ALEXCAP_API int alexcap_attn_forward(const alexcap_dims* dims, const alexcap_attn_params* params,
                         const float* feat, const int32_t* gt,
                         float* preds, float* alphas, const float* drop_mask,
                         void* workspace, size_t workspace_bytes, void* stream);

/* Autograd backward of the call above: consumes d preds [B,S,Vp] and d alphas [B,S,L] (either may be NULL = 0). */
ALEXCAP_API int alexcap_attn_backward(const alexcap_dims* dims, const alexcap_attn_params* params,
                          const float* feat, const float* dpreds, const float* dalphas,
                          const alexcap_attn_params* grads, float* dfeat,
                          void* workspace, size_t workspace_bytes, void* stream);

/* Test / introspection hook: byte offset and size of a named region of the attention training workspace for `dims`
 * ("MASK": the [S,B,H] dropout mask the last step drew; "ALPHA": [B,S,L]; "H": [S+1,B,D+H] slots [gctx_t | h_{t-1}]; "LOGITS": the fp32 logits chunk of the NON-fused vocabulary path
 * -- absent (ALEXCAP_E_BADARG) whenever the fused projection + cross-entropy epilogue is used, i.e. logits are never materialised).
 * Returns 0, or ALEXCAP_E_BADARG for an unknown name / a region this `dims` does not have. */
ALEXCAP_API int alexcap_attn_workspace_region(const alexcap_dims* dims, const char* name, size_t* offset, size_t* bytes);

/* m[b,:] = mean_l feat[b,l,:]  (Decoder.get_init_lstm_state, AttentionLanguageModule.py:117).  feat [B,L,D], m [B,D]. */
ALEXCAP_API int alexcap_mean_over_L(const float* feat, int32_t B, int32_t L, int32_t D, float* m, void* stream);

/* Selects how the forward recurrence of the attention decoder's training step runs (AttentionLanguageModule.py:72-82):
 *   0 (default) = four launches per time step (query GEMM, fused attention step, gates GEMM, LSTM pointwise), chained with
 *       programmatic dependent launch inside the caller's CUDA graph;
 *   1 = ONE persistent cooperative kernel for all S steps (time_loop.cu): recurrent weights resident in shared memory as bf16
 *       hi/lo pairs, the LSTM pointwise math done in the tcgen05 epilogue of the gates GEMM straight out of TMEM, the next
 *       step's query GEMM behind a grid barrier instead of a kernel boundary.  Shapes whose weight slices do not fit
 *       (B > 128, H > 512, ...) use 0 regardless.  Must be set BEFORE alexcap_attn_workspace_bytes is queried for a step
 *       (the workspace carries the loop's bf16 operand buffers).  Returns the previous setting.  Process-wide.
 * Measured at cfg2 on B200: 0 is ~4 % faster per training step; ALEXCAP_TIME_LOOP=1 sets the default. */
ALEXCAP_API int alexcap_set_time_loop(int32_t on);

/* Debug hook of the persistent time-loop kernels (time_loop.cu).  With ALEXCAP_LOOP_PROF=1 in the environment the kernels record
 * %globaltimer stamps (ns) of CTA 0 at every phase boundary; this call synchronises the device and copies them to the HOST array
 * host_out as [direction (0 forward, 1 backward)][step < 64][8 stamps].  Returns the number of uint64 written (0 = profiling off). */
ALEXCAP_API int alexcap_time_loop_profile(unsigned long long* host_out, int32_t max_entries);

/* ------------------------------------------------------------------ plain LSTM decoder (train) ---------- */
ALEXCAP_API size_t alexcap_lstm_workspace_bytes(const alexcap_dims* dims);

/* Decoder half of AlexCapModel.forward_train for the plain LSTM (LSTMModel.py:70-73; LanguageModule.py:99-107). */
ALEXCAP_API int alexcap_lstm_train_fused(const alexcap_dims* dims, const alexcap_lstm_params* params,
                             const float* feat, const int32_t* gt,
                             const alexcap_lstm_params* grads, float* dfeat,
                             float* loss_out, const float* drop_mask, const float* ce_denominator,
                             void* workspace, size_t workspace_bytes, void* stream);
/* LanguageModel.forward train branch (LanguageModule.py:99-107): preds [B,S,Vp]. */
ALEXCAP_API int alexcap_lstm_forward(const alexcap_dims* dims, const alexcap_lstm_params* params,
                         const float* feat, const int32_t* gt, float* preds, const float* drop_mask,
                         void* workspace, size_t workspace_bytes, void* stream);
ALEXCAP_API int alexcap_lstm_backward(const alexcap_dims* dims, const alexcap_lstm_params* params,
                          const float* feat, const float* dpreds,
                          const alexcap_lstm_params* grads, float* dfeat,
                          void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------ targets and loss -------------------
 * get_target (AttentionLanguageModule.py:197-212 == LanguageModule.py:113-128), both variants at once:
 *   in_tok[B,T+1] = [START, gt...];   tgt[B,T+1] = [gt..., 0] with the first 0 at index>=1 replaced by END.
 * Either output may be NULL. */
ALEXCAP_API int alexcap_make_target(const int32_t* gt, int32_t* in_tok, int32_t* tgt, int32_t B, int32_t T,
                        int32_t start_token, int32_t end_token, void* stream);

/* Number of non-ignored targets get_target(gt, make_target=True) yields for gt [B,T] (ignore_index 0): nnz(gt) + B (every
 * row gets exactly one END).  n_valid_out: 1 device float.  The divisor of CustomLoss.py:7's reduction='mean'; data-parallel
 * callers all-reduce it and hand N_valid_global / world to the training entry points as `ce_denominator`. */
ALEXCAP_API int alexcap_count_valid_targets(const int32_t* gt, int32_t B, int32_t T, float* n_valid_out, void* stream);

/* CustomCrossEntropyLoss.forward (CustomLoss.py:9-15) on logits [N,C] (row stride ld), targets [N] int32.
 * loss_out[2] (device) = {mean loss over non-ignored rows, n_valid}; dlogits [N,C] (row stride ld) or NULL
 * receives d loss/d logits (may alias logits).  scratch: >= 2*N floats. */
ALEXCAP_API int alexcap_cross_entropy(const float* logits, const int32_t* tgt, int32_t N, int32_t C, int64_t ld,
                          float label_smoothing, int32_t ignore_index,
                          float* loss_out, float* dlogits, float* scratch, void* stream);

/* alexcap_cross_entropy without the caller-provided scratch, for callers that cannot allocate device memory (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid cudaMalloc).  NO counterpart in the reference.
 * Same arithmetic, same kernels, same outputs as alexcap_cross_entropy; the scratch is a
 * file-scope __device__ array inside the library (csrc/no_alloc.cu; 2 * ALEXCAP_CE_STATIC_MAX_ROWS floats = 512 KB on every
 * GPU that loads the library, held until the process exits).  The call allocates and frees nothing.
 *   N > ALEXCAP_CE_STATIC_MAX_ROWS is rejected with ALEXCAP_E_UNSUPPORTED (message via alexcap_last_error()).
 *   Every call on a device shares the one array, which keeps the previous call's contents (the part a call uses is cleared
 *   on `stream` first): calls that may OVERLAP on the GPU (different streams without an ordering between them) race --
 *   serialise them, or use alexcap_cross_entropy with separate scratch buffers. */
#define ALEXCAP_CE_STATIC_MAX_ROWS 65536
ALEXCAP_API int32_t alexcap_cross_entropy_static_max_rows(void);
ALEXCAP_API int alexcap_cross_entropy_static(const float* logits, const int32_t* tgt, int32_t N, int32_t C, int64_t ld,
                          float label_smoothing, int32_t ignore_index,
                          float* loss_out, float* dlogits, void* stream);

/* ------------------------------------------------------------------ decoding ---------------------------
 * Decoder.forward greedy branch (AttentionLanguageModule.py:88-115), batched (the reference's own code only
 * works for B=1, :108-109).  ids [B,S] int64, alphas [B,S,L]; rows keep decoding until ALL rows emitted END
 * (:111), exactly like the reference loop; steps never reached stay 0.
 * The early exit is evaluated on the device: no host synchronisation. */
ALEXCAP_API size_t alexcap_attn_decode_workspace_bytes(const alexcap_dims* dims, int32_t beam);
ALEXCAP_API int alexcap_attn_decode_greedy(const alexcap_dims* dims, const alexcap_attn_params* params, const float* feat,
                               int64_t* ids, float* alphas, void* workspace, size_t workspace_bytes, void* stream);
/* Decoder.caption (AttentionLanguageModule.py:128-195) for B images at once, beam `beam` per image:
 * cumulative RAW-logit scores, step-1 top-k from beam row 0, finished beams shrink k, result = completed
 * beam with the highest score.  ids [B,S] int64 (without START, END included, 0 after), lens [B] int32
 * (0 = no beam completed: the reference raises ValueError at :191), scores [B], alphas [B,S,L] of the winner. */
ALEXCAP_API int alexcap_attn_decode_beam(const alexcap_dims* dims, const alexcap_attn_params* params, const float* feat,
                             int32_t beam, int64_t* ids, int32_t* lens, float* scores, float* alphas,
                             void* workspace, size_t workspace_bytes, void* stream);
/* LanguageModel.predict_caption (LanguageModule.py:130-150): ids [B,S] int64. */
ALEXCAP_API size_t alexcap_lstm_decode_workspace_bytes(const alexcap_dims* dims);
ALEXCAP_API int alexcap_lstm_decode_greedy(const alexcap_dims* dims, const alexcap_lstm_params* params, const float* feat,
                               int64_t* ids, void* workspace, size_t workspace_bytes, void* stream);
/* LanguageModel.caption (LanguageModule.py:152-198) for B images at once, beam `beam` per image.  The RAW features are
 * run through the LSTM (no image_encoder, :160): dims->D must equal dims->E (else ALEXCAP_E_BADARG).  Cumulative
 * RAW-logit scores, step-1 top-k from beam row 0, finished beams shrink k, at most seq_length+2 = S+1 steps (:190-192),
 * result = completed beam with the highest score.  ids [B,S+1] int64 (without the leading START the reference keeps,
 * END included, 0 after), lens [B] int32 (0 = no beam completed: the reference's max([]) raises ValueError at :194),
 * scores [B]. */
ALEXCAP_API size_t alexcap_lstm_decode_beam_workspace_bytes(const alexcap_dims* dims, int32_t beam);
ALEXCAP_API int alexcap_lstm_decode_beam(const alexcap_dims* dims, const alexcap_lstm_params* params, const float* feat,
                             int32_t beam, int64_t* ids, int32_t* lens, float* scores, void* workspace,
                             size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------ optimizer step ----------
 * Global-norm gradient clipping + Adam over ALL parameter tensors in three launches.  Replaces, with identical arithmetic,
 * AlexCapModel.clip_gradient (AlexCap/LSTMModel.py:50-51 = torch.nn.utils.clip_grad_norm_(params, max_norm), called from
 * train_LSTM.py:86-87 / train_LSTMwAttention.py:87-88) followed by optim.step() of torch.optim.Adam(params, lr, betas, eps,
 * weight_decay) (train_LSTMwAttention.py:60, :109).  The learning rate is passed per call, so the reference's LambdaLR
 * warm-up/cosine schedule (train_LSTMwAttention.py:58-76) keeps driving it from the host.
 *   params / grads / exp_avg / exp_avg_sq: HOST arrays of n_tensors device pointers (fp32, numel[i] elements each)
 *   max_norm <= 0: no clipping.  step: 1-based Adam step count (bias correction).  grads are scaled in place like
 *   clip_grad_norm_ does.  norm_out (device, 1 float) or NULL: the total gradient norm BEFORE clipping.
 *   scratch: device buffer of alexcap_clip_adam_scratch_floats() floats. */
#define ALEXCAP_OPT_MAX_TENSORS 48
ALEXCAP_API size_t alexcap_clip_adam_scratch_floats(void);
ALEXCAP_API int alexcap_clip_adam_step(int32_t n_tensors, float* const* params, float* const* grads, float* const* exp_avg,
                           float* const* exp_avg_sq, const int64_t* numel, float max_norm, float lr, float beta1,
                           float beta2, float eps, float weight_decay, int64_t step, float* norm_out, float* scratch,
                           void* stream);

/* The CNN hand-off either side of the path.  The backbone's output is NCHW [B, D, h, w];
 * the reference makes the decoder's grid with permute(0, 2, 3, 1).view(B, -1, C) (AlexCap/LSTMwAttentionModel.py:64-66,
 * :79-80).  alexcap_features_from_nchw writes feat[b, l, d] = nchw[b, d, l] (L = h*w) in one pass -- e.g. straight into
 * the buffer passed as `feat` to the training entry points; alexcap_dfeat_to_nchw is the way back for the gradient of the
 * fine-tune phase (train_LSTMwAttention.py:51-55): dnchw[b, d, l] (+)= dfeat[b, l, d]. */
ALEXCAP_API int alexcap_features_from_nchw(int32_t B, int32_t D, int32_t L, const float* nchw, float* feat, void* stream);
ALEXCAP_API int alexcap_dfeat_to_nchw(int32_t B, int32_t D, int32_t L, const float* dfeat, float* dnchw, int32_t accumulate,
                          void* stream);

/* ------------------------------------------------------------------ kernel-level entry points ----------
 * The building blocks, exported so the parity tests can pin each kernel on its own.
 * C[M,N] = act(alpha * op(A) op(B) + bias[N] + bias2[N]) (+ C if accumulate), fp32 SIMT or tensor-core path.
 *   transa=0: A is [M,K] (lda);  transa=1: A is stored [K,M] (lda).   transb=0: B is [K,N];  transb=1: B is [N,K].
 *   act: 0 none, 1 tanh, 2 relu.   engine: 0 = fp32 CUDA-core tiles, 1 = tcgen05 3xTF32, 100 = tcgen05 bf16, 200 = tcgen05 bf16x3,
 *   n in 2..99 = tcgen05 3xTF32 with split-K forced to n (100 + n, 200 + n likewise);
 *   300 (+n) = bf16x3 over operands pre-split in HBM (persistent kernel), 400 (+n) = bf16 pre-split; 500 / 600 = the same two
 *   but reusing the previous call's operand copies when the pointers match (times the GEMM kernel alone; test hook). */
ALEXCAP_API int alexcap_gemm_f32(int32_t transa, int32_t transb, int32_t M, int32_t N, int32_t K, float alpha,
                     const float* A, int64_t lda, const float* B, int64_t ldb,
                     float* C, int64_t ldc, const float* bias, const float* bias2,
                     int32_t accumulate, int32_t act, int32_t engine, void* stream);

/* One attention step, Attention.forward with W(features) precomputed (AttentionLanguageModule.py:15-22) and the
 * beta gate of Decoder.forward (:74-75) fused:  alpha[b,:] (row stride alpha_ld), ctx[B,D], beta[B,D],
 * gctx[b,:] (row stride gctx_ld) = sigmoid(bpre) * ctx.   q [B,H] (q_ld), bpre [B,D] (bpre_ld).
 * rows_bf16 (both step entry points) is a bit field.  Bit 0: Ws and feat point at BF16 rows of the same shapes (the
 * ALEXCAP_BF16 storage of the streamed grids, half the bytes per launch; D, H multiples of 8); everything else stays fp32.
 * Bit 1: Ws holds 2^(clamp(2 log2(e) W(features), +-60)) = e^(2 Ws) instead of the pre-activation -- the form the training
 * entry points keep it in (one MUFU op per score element: tanh(w + q) = 1 - 2 / (1 + e^(2w) e^(2q))); streaming kernels only. */
ALEXCAP_API int alexcap_attn_step_fwd(int32_t B, int32_t L, int32_t D, int32_t H,
                          const float* Ws, const float* feat, const float* q, int64_t q_ld,
                          const float* bpre, int64_t bpre_ld, const float* w_v,
                          float* alpha, int64_t alpha_ld, float* ctx, float* beta,
                          float* gctx, int64_t gctx_ld, int32_t rows_bf16, void* stream);

/* Backward of one attention step.  In: dgctx [B,D] (ld), saved beta/ctx/alpha/q, optional extra d alpha terms
 * (dalpha_ext row stride ext_ld; reg_r [B,L]).  Out: dq [B,H] (ld), dbpre [B,D] (ld), de[b,:] (de_ld),
 * dwv_part [B,H] += per-image partial of d w_v, dfeat [B,L,D] += alpha*dctx (or NULL). */
ALEXCAP_API int alexcap_attn_step_bwd(int32_t B, int32_t L, int32_t D, int32_t H,
                          const float* Ws, const float* feat, const float* q, int64_t q_ld, const float* w_v,
                          const float* alpha, int64_t alpha_ld, const float* ctx, const float* beta,
                          const float* dgctx, int64_t dgctx_ld,
                          const float* dalpha_ext, int64_t ext_ld, const float* reg_r,
                          float* dq, int64_t dq_ld, float* dbpre, int64_t dbpre_ld,
                          float* de, int64_t de_ld, float* dwv_part, float* dfeat, int32_t rows_bf16, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ALEXCAP_H_ */
