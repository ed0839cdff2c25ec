/*
 * mingpt_b200.h -- C ABI of libmingpt_b200.so: the sm_100a (B200) kernels behind the training step of
 * aponte411/minGPT-distributed (GPT forward/backward in mingpt/model.py, optimizer + gradient clipping in
 * mingpt/trainer.py).
 *
 * The reference has no FFI of its own (it is pure Python on top of torch), so every entry point below
 * cites the reference line -- or the torch call that line makes -- whose arithmetic it replaces.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless stated otherwise; the library never allocates, frees or
 *     retains memory past the call (the caller -- PyTorch's caching allocator in this repo -- owns it);
 *   - `stream` is a cudaStream_t passed as void*; all work is asynchronous on that stream;
 *   - return value 0 = ok, negative = error (mgb_last_error() gives a thread-local message);
 *     nothing throws across the boundary and nothing falls back to another implementation;
 *   - "bf16" buffers are __nv_bfloat16, row-major with an explicit leading dimension in ELEMENTS;
 *   - entry points are re-entrant: forward ops are called from the Python thread, backward ops from
 *     autograd's device thread.
 */
#ifndef MINGPT_B200_H_
#define MINGPT_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* mgb_stream_t;

enum {
    MGB_OK = 0,
    MGB_ERR_BAD_SHAPE = -1,
    MGB_ERR_BAD_ALIGN = -2,
    MGB_ERR_BAD_DTYPE = -3,
    MGB_ERR_UNSUPPORTED = -4,
    MGB_ERR_LAUNCH = -5,
    MGB_ERR_DRIVER = -6,
    MGB_ERR_TOO_LARGE = -7
};

enum { MGB_ACT_NONE = 0, MGB_ACT_GELU_TANH = 1, MGB_ACT_GELU_ERF = 2,
       MGB_GATE_MUL = 3,            /* gate_kind only: the gate tensor holds the factor itself */
       MGB_ACT_STORE_GRAD = 256 };  /* OR-ed into act: preact receives act'(v) instead of v */
enum { MGB_GEMM_ENGINE_SINGLE = 1024, MGB_GEMM_ENGINE_PAIR = 2048 };

/* attention mask modes.
 *   CAUSAL         scores[i,j] = -inf for j > i  (the README / north_star "CausalSelfAttention")
 *   ADD_TRIL_ONES  scores[i,j] += 1.0 for j <= i, nothing masked: what mingpt/model.py:142-145,164 really
 *                  computes, because a FLOAT attn_mask is ADDED by nn.MultiheadAttention
 *                  (torch/nn/functional.py:6638-6643)
 *   NONE           plain bidirectional softmax */
enum { MGB_MASK_NONE = 0, MGB_MASK_CAUSAL = 1, MGB_MASK_ADD_TRIL_ONES = 2 };

const char* mgb_last_error(void);
int mgb_version(void);
/* number of kernels this library has launched in the calling process (bench.py's gpu_launches) */
int64_t mgb_launch_count(void);
/* Tuning / test knobs (process-wide). Keys: "gemm_single" (1 = always use the single-CTA tile engine),
 * "gemm_direct_store" (1 = CTA-pair engine writes D from registers instead of through TMA), "attn_mma"
 * (1 = head size 64 also uses the mma.sync attention kernels instead of the tcgen05 ones), "ln_unstaged"
 * (1 = LayerNorm backward loads its rows straight into registers instead of staging them through shared
 * memory with bulk copies), "ce_unstaged" (1 = cross-entropy streams its rows from global memory, two
 * exponentials per logit, instead of staging them in shared memory), "pdl" (default 1 = launch the GEMM,
 * LayerNorm and attention kernels with programmatic stream serialization; 0 = ordinary launches, so that a kernel's prologue --
 * barrier set-up, TMEM allocation -- overlaps the tail of the kernel before it; every such kernel waits on
 * griddepcontrol before it reads global memory), "gemm_generic" (1 = the CTA-pair GEMM always runs its generic
 * epilogue instantiation, which decides every epilogue term at run time, instead of the ones specialised at compile
 * time for the combinations the training step uses), "attn_poly" (default 1: share of the forward softmax's
 * exponentials computed by a cubic polynomial on the FMA pipe instead of the special-function unit, in quarters of
 * the 128 keys of a block -- 0, 1 or 2; relative error of the polynomial 7.5e-5), "reserved_sms" (default 0; n = size the persistent grids -- GEMM,
 * attention backward, cross-entropy, LayerNorm backward -- for n fewer SMs than the device has, leaving them to a
 * concurrent communication kernel: the data-parallel trainer sets it while gradient all-reduces are in flight, so no
 * CTA pair of a persistent GEMM waits a whole tile wave for SMs that NCCL occupies). Returns the
 * previous value, or -1 for an unknown key. Every alternative computes the same result. */
int mgb_set_option(const char* key, int value);

/* ---------------------------------------------------------------------------------------------------
 * GEMM: D[m,n] = epilogue(alpha * sum_k A(m,k) B(n,k)), bf16 inputs, fp32 accumulate (tcgen05 + TMEM).
 * Replaces nn.Linear forward/backward: mingpt/model.py:138 (c_proj), :180-181 (mlp.0, mlp.1), :249 (head),
 * torch/nn/functional.py:6478 (MHA in-proj), :6653 (MHA out_proj), and autograd's dgrad/wgrad of each
 * (mingpt/trainer.py:127).
 *   a_mn_major = 0: A stored [m][k] (k contiguous, lda between rows of m)
 *   a_mn_major = 1: A stored [k][m] (m contiguous, lda between rows of k)       -- same for B with n
 * epilogue, in this order:  v = alpha*acc;  v += bias[n];  preact[m,n] = bf16(v) (if preact; bf16(act'(v))
 *   when act carries MGB_ACT_STORE_GRAD: the forward GEMM then leaves exactly what the backward needs and
 *   its tanh / erf is evaluated once);  v = act(v);  v *= act'(gate[m,n]) (if gate; gate_kind selects the
 *   derivative, MGB_GATE_MUL multiplies by gate[m,n] itself);  v += residual[m,n];
 *   D = v  (bf16 or fp32)   or   D += v (fp32, atomic; required for split_k > 1)
 * The epilogue is applied exactly once per output element: with `accumulate` the reduction is split over K only
 * when there is NO epilogue term (bias, act, gate, residual, preact, colsum all absent -- the weight-gradient
 * case); with any of them present an automatic split (split_k = 0) is forced to 1 and an explicit split_k > 1 is
 * rejected (MGB_ERR_UNSUPPORTED), because every split unit would otherwise run the epilogue on its partial sum.
 * D rows are written in 16-byte units: when n is not a multiple of 8 (bf16) / 4 (fp32) AND ldd leaves room
 * after column n, the rest of the last unit (at most 7 / 3 elements) is overwritten too (with the epilogue of
 * a zero accumulator). With ldd == n (a dense D) nothing outside the matrix is ever touched.
 * ------------------------------------------------------------------------------------------------- */
typedef struct mgb_gemm_args {
    const void* a; int64_t lda; int32_t a_mn_major;
    const void* b; int64_t ldb; int32_t b_mn_major;
    void* d; int64_t ldd; int32_t d_f32; int32_t accumulate;
    int64_t m, n, k;
    const float* bias;
    const void* residual; int64_t ldr;
    int32_t act;
    void* preact; int64_t ldp;
    const void* gate; int64_t ldg; int32_t gate_kind;
    int32_t split_k;   /* 0 = auto */
    int32_t block_n;   /* 0 = auto; else tile width 64/128/192/256, optionally OR-ed with an engine request:
                          MGB_GEMM_ENGINE_SINGLE (128-row tiles, one CTA) or MGB_GEMM_ENGINE_PAIR (256-row tiles,
                          two CTAs, tcgen05 cta_group::2). Both engines compute the same result. */
    float alpha;       /* 0 is read as 1 */
    float* colsum;     /* optional, bf16 D only: colsum[n] += sum_m D[m,n] (fp32, atomic) -- the bias gradient of the
                          layer below when D is an output gradient; taken from the accumulators before rounding */
    /* optional "row dot" side output (CTA-pair engine, bf16 D, n a multiple of 64, no gate / residual / preact):
     * rowdot_out[(b * (n/64) + h) * rowdot_t + t] = sum over the 64 columns of head h of D[m, .] * rowdot_x[m, .] for
     * row m = b * rowdot_t + t, taken from the fp32 values before D is rounded. With D = dO (the gradient of the
     * attention output, produced by the out-projection's dgrad) and rowdot_x = O this is flash attention's
     * delta = rowsum(dO * O) [B, H, T], which the attention backward otherwise computes in a pass of its own. */
    const void* rowdot_x; int64_t ld_rowdot;
    float* rowdot_out; int64_t rowdot_t;
    /* optional (CTA-pair engine, bf16 D, no activation): dropout between the bias and the residual,
     *   D = residual + (keep ? (alpha * A B^T + bias) / (1 - drop_p) : 0),
     * with the mask mgb_dropout_bf16 draws for (seed, site) over a contiguous [m, n] tensor (element = row * n +
     * col; n % 16 == 0): c_proj -> resid_drop -> residual add of the reference's blocks (mingpt/model.py:140,167,183,
     * 187) in the projection's epilogue. drop_p = 0 or drop_seed = NULL: off. */
    const uint64_t* drop_seed; float drop_p; uint32_t drop_site;
} mgb_gemm_args;
int mgb_gemm_bf16(const mgb_gemm_args* args, mgb_stream_t stream);

/* ---------------------------------------------------------------------------------------------------
 * LayerNorm (nn.LayerNorm, eps 1e-5, biased variance, affine): mingpt/model.py:176,178,248 applied at
 * :187-188,312.  x,y bf16 [rows, cols]; gamma/beta fp32; mean/rstd fp32 [rows] saved for backward.
 * ------------------------------------------------------------------------------------------------- */
int mgb_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean, float* rstd,
                      int64_t rows, int32_t cols, float eps, mgb_stream_t stream);
/* dx = (dres ? dres : 0) + LN'(dy); dgamma += sum_rows(dy * xhat); dbeta += sum_rows(dy)  (fp32, atomic).
 * `dres` fuses the residual-stream gradient of Block.forward (mingpt/model.py:187-188). dx may alias dres. */
int mgb_layernorm_bwd(const void* dy, const void* x, const float* gamma, const float* mean, const float* rstd,
                      const void* dres, void* dx, float* dgamma, float* dbeta, int64_t rows, int32_t cols,
                      mgb_stream_t stream);
/* Same, and additionally dres_colsum[c] += sum over rows of dres[row, c] (fp32, atomic): dres is the gradient of
 * the residual stream, i.e. also the output gradient of the linear layer that wrote into it (c_proj of the
 * attention or of the MLP, mingpt/model.py:138,181), so this is that layer's bias gradient, taken from the read
 * this kernel makes anyway instead of a separate pass over dres. dres must not be NULL. */
int mgb_layernorm_bwd_colsum(const void* dy, const void* x, const float* gamma, const float* mean, const float* rstd,
                             const void* dres, void* dx, float* dgamma, float* dbeta, float* dres_colsum,
                             int64_t rows, int32_t cols, mgb_stream_t stream);

/* ---------------------------------------------------------------------------------------------------
 * Embedding (GPTEmbedding.forward, mingpt/model.py:222-231 as repaired: wte[idx] + wpe[:, :T]).
 * idx int64 [B,T]; wte bf16 [vocab, C]; wpe bf16 [block_size, C]; out bf16 [B,T,C].
 * An index outside [0, vocab) sets *err_flag (int32, device, may be NULL) and produces zeros.
 * ------------------------------------------------------------------------------------------------- */
int mgb_embedding_fwd(const int64_t* idx, const void* wte, const void* wpe, void* out, int64_t B, int64_t T,
                      int32_t C, int64_t vocab, int32_t* err_flag, mgb_stream_t stream);
/* Input pipeline: cuts a batch of training windows out of a byte-coded corpus that already lives on the device --
 * CharDataset.__getitem__ (mingpt/char_dataset.py:35-47) for B window starts at once, instead of the reference's
 * per-item Python list comprehension + DataLoader collation + two blocking H2D copies (mingpt/trainer.py:141-142).
 * codes: uint8 [n_codes] (vocabulary index of every corpus byte); starts: int64 [B];
 * x[b,t] = codes[starts[b] + t], y[b,t] = codes[starts[b] + t + 1], both int64 [B,T] (the dtype GPT.forward takes).
 * A window that does not fit (start < 0 or start + T + 1 > n_codes) sets *err_flag = 3 (may be NULL) and is skipped. */
int mgb_token_windows(const void* codes, int64_t n_codes, const int64_t* starts, int64_t B, int32_t T,
                      int64_t* x, int64_t* y, int32_t* err_flag, mgb_stream_t stream);
/* dwte[idx[b,t], :] += dx[b,t,:] (fp32 atomics);  dwpe[t,:] += sum_b dx[b,t,:] */
int mgb_embedding_bwd(const int64_t* idx, const void* dx, float* dwte, float* dwpe, int64_t B, int64_t T,
                      int32_t C, int64_t vocab, mgb_stream_t stream);

/* ---------------------------------------------------------------------------------------------------
 * Fused attention (flash-style: online softmax, no T x T tensor in HBM). Replaces
 * torch/nn/functional.py:6630-6647 (scale, baddbmm with mask, softmax, bmm) called from
 * mingpt/model.py:160-165.
 * qkv bf16 [B, T, 3, H, d] (the in-proj output, untouched); out bf16 [B, T, H, d]; lse fp32 [B, H, T]
 * (natural-log sum-exp of the masked/biased scaled scores).  d in {16, 32, 64}.
 * ------------------------------------------------------------------------------------------------- */
int mgb_attention_fwd(const void* qkv, void* out, float* lse, int32_t B, int32_t T, int32_t H, int32_t d,
                      int32_t mask_mode, float scale, mgb_stream_t stream);
/* OR-ed into mask_mode of the backward entry points: `delta` already holds rowsum(dO * O) -- the out-projection's dgrad
 * GEMM produced it in its epilogue (mgb_gemm_args.rowdot_*) -- so the backward skips its own pass over O and dO. */
#define MGB_ATTN_DELTA_GIVEN 0x100
/* Backward with recompute. delta fp32 [B,H,T] and dq_accum fp32 [B,T,H,d] are caller-provided scratch
 * (dq_accum is zeroed by the call); dqkv bf16 [B,T,3,H,d] is fully overwritten. */
int mgb_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                      float* dq_accum, void* dqkv, int32_t B, int32_t T, int32_t H, int32_t d, int32_t mask_mode,
                      float scale, mgb_stream_t stream);
/* Same, and additionally dqkv_colsum[0 .. 3*H*d) += column sums of dqkv (fp32, atomic) = the bias gradient of the
 * qkv projection (torch/nn/functional.py:6478 in_proj_bias), produced where the values are already in registers
 * (head size 64: the main kernel's per-key-block epilogue and the dQ conversion) instead of by a pass over dqkv. */
int mgb_attention_bwd_colsum(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                             float* dq_accum, void* dqkv, float* dqkv_colsum, int32_t B, int32_t T, int32_t H,
                             int32_t d, int32_t mask_mode, float scale, mgb_stream_t stream);

/* Incremental decode for GPT.generate (mingpt/model.py:322-356; the reference re-runs the whole window for every
 * new token, :334-339): attention of ONE new position per sequence against cached keys / values.
 * qkv_cache bf16 [B, Tcap, 3, H, d] -- the packed in-proj layout; row `pos` of every sequence holds the new token's
 * q / k / v (the caller's qkv GEMM writes it there), rows 0..pos-1 the earlier tokens' k / v. Causal attention of the
 * last position sees every cached key, so there is no mask. out bf16 [B, H*d]. fp32 softmax. */
int mgb_attention_decode(const void* qkv_cache, void* out, int32_t B, int32_t Tcap, int32_t pos, int32_t H,
                         int32_t d, float scale, mgb_stream_t stream);

/* Dropout (the reference trains with p = 0.1 at four sites and never calls eval(): mingpt/model.py:49-51,140,
 * 150,183,215). Masks are counter based (Philox4x32, 7 rounds, over (seed, site, element)): the backward
 * regenerates them, or reads the packed keep words of mgb_attention_dropout_bits; no mask tensor is kept. *seed_dev is a DEVICE uint64 the caller changes every step; `site` tells the
 * dropouts of one step apart. An element is kept with probability 1 - round(p * 65536) / 65536.
 *   mgb_dropout_bf16:   y = (res ? res : 0) + (keep ? x / (1-p) : 0)     (embedding / residual dropouts, and
 *                       their backward: call it on the incoming gradient with res = NULL)
 *   mgb_attention_*_dropout: mgb_attention_fwd / _bwd with dropout on the softmax probabilities
 *                       (torch/nn/functional.py:6644-6645); p_drop = 0 forwards to the plain entry points
 *   mgb_attention_dropout_mask: the uint8 keep mask [B,H,T,T] those kernels use (for tests) */
int mgb_dropout_bf16(const void* x, const void* res, void* y, int64_t n, float p, const uint64_t* seed_dev,
                     uint32_t site, mgb_stream_t stream);
/* mgb_dropout_bf16 (res = NULL) over a [rows, cols] gradient, plus colsum[c] += sum over rows of y[:, c]: the
 * backward of a residual-branch dropout and, in the same pass, the bias gradient of the linear layer in front of
 * it. Same mask as mgb_dropout_bf16 over the contiguous tensor (element = row * cols + col); cols % 8 == 0. */
int mgb_dropout_bf16_colsum(const void* x, void* y, int64_t rows, int32_t cols, float p, const uint64_t* seed_dev,
                            uint32_t site, float* colsum, mgb_stream_t stream);
int mgb_attention_fwd_dropout(const void* qkv, void* out, float* lse, int32_t B, int32_t T, int32_t H, int32_t d,
                              int32_t mask_mode, float scale, float p_drop, const uint64_t* seed_dev, uint32_t site,
                              mgb_stream_t stream);
int mgb_attention_bwd_dropout(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                              float* dq_accum, void* dqkv, int32_t B, int32_t T, int32_t H, int32_t d,
                              int32_t mask_mode, float scale, float p_drop, const uint64_t* seed_dev, uint32_t site,
                              mgb_stream_t stream);
int mgb_attention_dropout_mask(void* keep, int32_t B, int32_t T, int32_t H, float p_drop, const uint64_t* seed_dev,
                               uint32_t site, mgb_stream_t stream);
/* The same dropout with the keep bits generated ahead of the attention kernels (head size 64 only). The softmax
 * warps of those kernels are bound by their instruction stream, and the inline Philox calls doubled it; a separate,
 * well-occupied generator produces the identical bits once per layer and step, and forward AND backward read them.
 *   mgb_attention_dropout_bits_words: *words = size of the buffer in 32-bit words,
 *                                     B*H*ceil(T/32) * (roundup(T,16) + roundup(T,32))
 *   mgb_attention_dropout_bits:       fills it: first [B*H][T/32][roundup(T,16)] words of 32 keys of one query (key
 *                                     k at bit 16*(k&1) + 15 - (k>>1)) for the forward, then the same bits
 *                                     transposed, [B*H][T/32][roundup(T,32)] words of 32 queries of one key (query
 *                                     q at bit q%32) for the backward; for MGB_MASK_CAUSAL the words right of the
 *                                     diagonal are left untouched
 *   mgb_attention_fwd_dropout_bits / _bwd_dropout_bits: mgb_attention_fwd_dropout / _bwd_dropout reading that buffer;
 *                                     the backward also takes dqkv_colsum (or null) like mgb_attention_bwd_colsum.
 * Same results as the seed-taking entry points above (same mask, same arithmetic). */
int mgb_attention_dropout_bits_words(int32_t B, int32_t T, int32_t H, int64_t* words);
int mgb_attention_dropout_bits(uint32_t* bits, int32_t B, int32_t T, int32_t H, int32_t mask_mode, float p_drop,
                               const uint64_t* seed_dev, uint32_t site, mgb_stream_t stream);
int mgb_attention_fwd_dropout_bits(const void* qkv, void* out, float* lse, int32_t B, int32_t T, int32_t H, int32_t d,
                                   int32_t mask_mode, float scale, float p_drop, const uint32_t* bits,
                                   mgb_stream_t stream);
int mgb_attention_bwd_dropout_bits(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                                   float* dq_accum, void* dqkv, int32_t B, int32_t T, int32_t H, int32_t d,
                                   int32_t mask_mode, float scale, float p_drop, const uint32_t* bits,
                                   float* dqkv_colsum, mgb_stream_t stream);

/* No counterpart in the reference. mgb_attention_bwd for callers that have no scratch to give: a plugin
 * inside a process whose device pool is fixed at start-up, or an FFI stub that only forwards pointers and
 * owns no allocator. delta and the dQ accumulator live in a fixed 17 MiB __device__ array inside the library
 * (held by every GPU that loads it, until exit); the call allocates nothing. Limits and rules:
 *   - B*T*H*d <= 4194304, otherwise MGB_ERR_TOO_LARGE (use mgb_attention_bwd);
 *   - the first call on each device looks up the array's address; make that call before any stream capture;
 *   - all calls share the one array (it is cleared as needed at the start of each call), so two calls that
 *     overlap on the GPU -- different streams, or different threads on one device -- race: serialise them. */
int mgb_attention_bwd_noalloc(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                              int32_t B, int32_t T, int32_t H, int32_t d, int32_t mask_mode, float scale,
                              mgb_stream_t stream);

/* ---------------------------------------------------------------------------------------------------
 * Cross entropy (F.cross_entropy(..., ignore_index=-1), mean over valid targets: mingpt/model.py:316-318).
 * In place on a chunk of bf16 logits [rows, ld] (only the first `vocab` columns are meaningful):
 *   row_loss[r] = lse - logit[target]   (0 for ignored rows)
 *   logits[r,:] <- (softmax - onehot) * grad_scale / max(*n_valid, 1)    (0 for ignored rows), bf16
 * n_valid is a DEVICE int32 holding the number of non-ignored targets of the WHOLE batch.
 * With write_grad = 0 the logits are left untouched (evaluation).
 * ------------------------------------------------------------------------------------------------- */
int mgb_cross_entropy_fwd_bwd(void* logits, int64_t ld, const int64_t* targets, float* row_loss,
                              const int32_t* n_valid, int64_t rows, int64_t vocab, float grad_scale,
                              int32_t write_grad, int32_t* err_flag, mgb_stream_t stream);
int mgb_count_valid_targets(const int64_t* targets, int64_t n, int32_t* n_valid, mgb_stream_t stream);
/* loss = sum(row_loss) / max(*n_valid, 1); deterministic single-block tree reduction */
int mgb_loss_finalize(const float* row_loss, int64_t rows, const int32_t* n_valid, float* loss,
                      mgb_stream_t stream);

/* ---------------------------------------------------------------------------------------------------
 * Small HBM-bound helpers of the backward pass.
 * ------------------------------------------------------------------------------------------------- */
/* out[c] += sum_r x[r, c]   (bias gradients; fp32 atomics) */
int mgb_colsum_bf16(const void* x, int64_t ld, int64_t rows, int64_t cols, float* out, mgb_stream_t stream);
/* dz = dy * act'(z)   (nn.GELU backward when the activation is not fused into a GEMM epilogue) */
int mgb_act_bwd_bf16(const void* dy, const void* z, void* dz, int64_t n, int32_t act, mgb_stream_t stream);
/* y = x + act(z) or act(z) (elementwise forward used by tests and the unfused path) */
int mgb_act_fwd_bf16(const void* z, const void* res, void* y, int64_t n, int32_t act, mgb_stream_t stream);
int mgb_cast_f32_to_bf16(const float* src, void* dst, int64_t n, mgb_stream_t stream);
int mgb_cast_bf16_to_f32(const void* src, float* dst, int64_t n, mgb_stream_t stream);
/* dst += alpha * src (fp32) */
int mgb_axpy_f32(float* dst, const float* src, float alpha, int64_t n, mgb_stream_t stream);
/* Same with alpha read from DEVICE memory, and x *= *alpha for bf16: they apply the upstream gradient of the
 * scalar loss (loss.backward() at mingpt/trainer.py:127 hands it over as a device tensor) without a host sync. */
int mgb_axpy_f32_dev(float* dst, const float* src, const float* alpha, int64_t n, mgb_stream_t stream);
int mgb_scale_bf16_dev(void* x, int64_t n, const float* alpha, mgb_stream_t stream);

/* ---------------------------------------------------------------------------------------------------
 * Gradient clipping + AdamW.
 *   clip_grad_norm: mingpt/trainer.py:129-130 -> torch/nn/utils/clip_grad.py (global L2 norm,
 *                   coef = max_norm / (norm + 1e-6) clamped to 1)
 *   AdamW step:     mingpt/trainer.py:132 -> torch/optim/adam.py (decoupled weight decay, eps 1e-8,
 *                   bias correction), groups from create_optimizer (mingpt/model.py:62-122)
 * mgb_sumsq_f32 ACCUMULATES sum(x^2) into *out (fp32, device; zero it first with cudaMemsetAsync).
 * mgb_adamw_multi updates every tensor described by `table` in ONE launch.  The clip coefficient is
 * computed on the device from *gradnorm_sq (may be NULL = no clipping) so no host sync is needed:
 *   g = grad * grad_scale;  norm = sqrt(*gradnorm_sq) * grad_scale;  g *= min(1, max_norm/(norm+1e-6))
 * step_dev (DEVICE int32, may be NULL): when given, the step count for the bias corrections is read from it
 * instead of `step`, so a captured CUDA graph of the training step can be replayed (host arguments are frozen
 * at capture time; the caller increments *step_dev inside the graph).
 * ------------------------------------------------------------------------------------------------- */
typedef struct mgb_adamw_chunk {
    int64_t offset;      /* element offset into the flat fp32 param / grad / m / v buffers */
    int32_t numel;       /* elements in this chunk (<= chunk size chosen by the caller) */
    float weight_decay;  /* this tensor's decay (0 for the no-decay group) */
} mgb_adamw_chunk;
int mgb_sumsq_f32(const float* x, int64_t n, float* out, mgb_stream_t stream);
int mgb_adamw_multi(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, void* params_bf16,
                    const mgb_adamw_chunk* table /*device*/, int32_t num_chunks, float lr, float beta1, float beta2,
                    float eps, int32_t step, float grad_scale, const float* gradnorm_sq, float max_norm,
                    const int32_t* step_dev, mgb_stream_t stream);
/* grads *= min(1, max_norm / (sqrt(*gradnorm_sq) + 1e-6))  -- for callers that keep a foreign optimizer */
int mgb_clip_scale_f32(float* grads, int64_t n, const float* gradnorm_sq, float max_norm, mgb_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* MINGPT_B200_H_ */
