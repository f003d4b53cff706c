/* b200ml — C ABI of libb200ml.so (sm_100a kernels + host runtime).
 *
 * Every entry point takes raw device pointers and a CUDA stream, launches asynchronously on that stream and
 * returns 0 on success or -1 on failure (text from b200ml_last_error()).  No entry point allocates device
 * memory unless its comment says so; scratch is passed in by the caller (size from the *_scratch_bytes helper).
 * bf16 tensors are row-major with an explicit leading dimension in elements where one is taken.
 *
 * The "ref:" notes cite the reference behaviour each kernel reproduces (paths relative to the reference repo).
 */
#ifndef B200ML_H_
#define B200ML_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __CUDA_RUNTIME_H__
typedef struct CUstream_st* cudaStream_t;
#endif

/* ------------------------------------------------------------------------------------------------ runtime */
const char* b200ml_last_error(void);
int b200ml_version(void);
int b200ml_device_info(int* sm_count, int* cc_major, int* cc_minor, size_t* smem_optin);

/* ------------------------------------------------------------------------------------------------ K4/K5/K13 GEMM
 * C[M,N] = epilogue(alpha * A·Bᵀ).  tcgen05/TMEM/TMA kernel; SIMT fallback when a leading dimension is not a
 * multiple of 8 elements or a base pointer is not 16-byte aligned (or force_simt != 0).
 *   a_mn_major = 0: A is [M,K] (K contiguous)      1: A is stored [K,M] (M contiguous)
 *   b_mn_major = 0: B is [N,K] (K contiguous)      1: B is stored [K,N] (N contiguous)
 * epilogue order: v = alpha*acc (+bias[col]) -> act (0 none, 1 ReLU, 2 sigmoid) -> dropout(dropout_p) ->
 *   aux_mode (0 none; 1: v = aux>0 ? v*aux_scale : 0; 2: v += aux; 3: v *= aux*(1-aux)) -> store (bf16 or fp32;
 *   fp32 may accumulate into C).
 * ref: every nn.Linear + ReLU/Sigmoid/Dropout neighbour (transformer.py:71-72,107-117,177-179,272;
 *      distributed_multilayer_perceptron.py:47-53). */
typedef struct {
    int M, N, K;
    const void* A; long long lda; int a_mn_major;
    const void* B; long long ldb; int b_mn_major;
    void* C; long long ldc; int c_fp32;
    const float* bias;
    int act;
    const void* aux; long long ldaux; int aux_mode; float aux_scale;
    float alpha;
    const float* alpha_dev; /* optional device scalar multiplied into alpha */
    int accumulate;
    float dropout_p; const void* rng_state; unsigned rng_site;
    int tile_n;      /* 0 = auto, or 64 / 128 / 256 */
    int force_simt;  /* testing: run the SIMT fallback */
    int split_k;     /* 0 = auto, 1 = never, n > 1 = request n K-splits (plain epilogue only; needs workspace) */
    void* workspace; size_t workspace_bytes; /* fp32 scratch for split-K partial tiles (optional) */
    int epilogue_direct; /* testing: 1 = register->global stores even where the TMA-store epilogue applies */
    int max_ctas;    /* > 0: use at most this many CTAs of the persistent tcgen05 kernel (a leaf GEMM that should leave
                        SMs to kernels on the critical path of other streams); 0 = one per SM */
    void* row_stats; /* optional float2 [M][row_stats_ld]: softmax statistics of every output row per column block, taken
                        from the (bf16-rounded) values the epilogue holds in registers: entry [row][2 * (n0 / BN) + h] =
                        (max, sum exp(v - max)) over the h-th half of the tile's columns.  Only with the bf16 TMA-store
                        epilogue (tile_n = 256 or 512, bf16 C with 16-byte rows); row_stats_ld >= 2 * ceil(N / 256).
                        Lets the cross-entropy that follows skip its max / sum passes (b200ml_xent_lse). */
    int row_stats_ld;
} b200ml_gemm_args;
int b200ml_gemm_bf16(const b200ml_gemm_args* args, cudaStream_t stream);

/* ------------------------------------------------------------------------------------------------ K11/K12 LayerNorm
 * y = LN(x + dropout(a)); either of a / x may be NULL.  z_out (bf16, optional) keeps the pre-norm sum for backward.
 * d must be a multiple of 8 and <= 1024.  ref: transformer.py:86-101,130-139,209-224. */
int b200ml_layernorm_fwd(const void* a, const void* x, const float* gamma, const float* beta, void* y, void* z_out,
                         float* mean, float* rstd, int rows, int d, float eps, float dropout_p, const void* rng_state,
                         unsigned rng_site, cudaStream_t stream);
size_t b200ml_layernorm_bwd_scratch_bytes(int rows, int d);
/* dz = gradient w.r.t. the pre-norm sum (i.e. the residual branch); da (optional) = dz through the dropout mask.
 * dgamma/dbeta are overwritten (accumulate = 0) or added to.  scratch holds per-CTA partial sums. */
int b200ml_layernorm_bwd(const void* dy, const void* z, const float* mean, const float* rstd, const float* gamma,
                         void* dz, void* da, float* dgamma, float* dbeta, void* scratch, size_t scratch_bytes, int rows,
                         int d, float dropout_p, const void* rng_state, unsigned rng_site, int accumulate,
                         cudaStream_t stream);

/* The same, plus dbias_a[d] = column sums of d_a (of dz when there is no dropout): the bias gradient of the Linear whose
 * output `a` was, produced from the registers that hold d_a anyway instead of by a separate pass over it.  scratch:
 * b200ml_layernorm_bwd_scratch_bytes.  ref: the bias of every Linear that feeds a residual LayerNorm
 * (transformer.py:108,130-139; 177-179,209-224). */
int b200ml_layernorm_bwd_bias(const void* dy, const void* z, const float* mean, const float* rstd, const float* gamma,
                              void* dz, void* da, float* dgamma, float* dbeta, float* dbias_a, void* scratch,
                              size_t scratch_bytes, int rows, int d, float dropout_p, const void* rng_state,
                              unsigned rng_site, int accumulate, int accumulate_bias, cudaStream_t stream);

/* LayerNorm backward WITHOUT caller scratch and without any allocation (no counterpart in the reference; added for
 * hosts that may not allocate).  Same arguments and results as b200ml_layernorm_bwd minus (scratch, scratch_bytes): the
 * per-CTA partial sums of dgamma/dbeta go to a ~2.3 MiB array that is compiled into the library and lives on every GPU
 * that loads it until the process exits.  Meant for callers whose device memory is owned by their own pool and who
 * must not call cudaMalloc/cudaFree after start-up (the call can fail when the pool holds the rest of HBM, and it
 * synchronises the device), and for C/FFI callers with no allocator to borrow a workspace from.
 *   - limit: d % 8 == 0 and d <= b200ml_layernorm_bwd_noalloc_max_d() (1024); larger d is refused with an error —
 *     use the scratch-taking entry point for it.  rows is unlimited, but the grid is capped at 296 CTAs (the
 *     scratch-taking entry point uses up to 4 per SM, 592 on a B200), so large inputs run somewhat slower here.
 *   - the FIRST call on each device looks up the array's device address (cudaGetSymbolAddress, with that device
 *     current); make that first call before any stream capture begins.
 *   - all calls on a device share the one array.  Nothing in it needs clearing between calls (each call overwrites
 *     what it later reads), but two calls whose kernels overlap in time on the same GPU — different streams, or
 *     several host threads — race on it: serialise them on one stream or with events. */
size_t b200ml_layernorm_bwd_noalloc_max_d(void);
int b200ml_layernorm_bwd_noalloc(const void* dy, const void* z, const float* mean, const float* rstd,
                                 const float* gamma, void* dz, void* da, float* dgamma, float* dbeta, int rows, int d,
                                 float dropout_p, const void* rng_state, unsigned rng_site, int accumulate,
                                 cudaStream_t stream);

/* ------------------------------------------------------------------------------------------------ K7/K8 embedding
 * ref: transformer.py:27-62 (sinusoidal PE + embedding + dropout), distributed_lstm.py:115,128 (padding_idx). */
int b200ml_pe_table(float* pe, int S, int d, cudaStream_t stream);
int b200ml_embedding_fwd(const long long* tokens, const void* table_bf16, const float* pe, void* out, int out_fp32,
                         int n_tokens, int S, int d, int V, float dropout_p, const void* rng_state, unsigned rng_site,
                         cudaStream_t stream);
/* dtable (fp32, [V,d]) += scatter of dout (fp32 atomicAdd: the order of the additions, and so the last bits, vary from
 * run to run); the caller zeroes dtable once per step.  padding_idx < 0 disables it.  Kept for token counts above
 * b200ml_embedding_index_max_tokens(); the framework's own path is the sorted pair below. */
int b200ml_embedding_bwd(const long long* tokens, const void* dout, int dout_fp32, float* dtable, int n_tokens, int d,
                         int V, long long padding_idx, float dropout_p, const void* rng_state, unsigned rng_site,
                         cudaStream_t stream);
/* Deterministic backward in two launches.  b200ml_embedding_index (one CTA; n_tokens <=
 * b200ml_embedding_index_max_tokens() = 16384; depends on the token ids only, so it can run at forward time) sorts the
 * keys (token id << 32 | position) of one batch and writes, into a caller buffer of b200ml_embedding_index_bytes()
 * bytes, the sorted keys, the start of every segment of equal tokens, the list of segments with more than 64 tokens
 * and one byte per table row telling whether the batch touches it.  b200ml_embedding_bwd_sorted then WRITES every row
 * of dtable exactly once: the sum over the row's tokens in position order, zeros for untouched rows and the padding
 * row (accumulate = 0; the table needs no clearing), or adds to what is there (accumulate = 1).  Long segments are
 * summed in 32 interleaved ways through `scratch` (b200ml_embedding_bwd_sorted_scratch_bytes; it starts with counters
 * that must be ZERO before the first call and are left zero by every call; calls sharing a scratch must not overlap).
 * ref: nn.Embedding backward (transformer.py:52-61; distributed_lstm.py:115,128). */
int b200ml_embedding_index_max_tokens(void);
size_t b200ml_embedding_index_bytes(int n_tokens, int V);
int b200ml_embedding_index(const long long* tokens, int n_tokens, int V, void* index_out, unsigned char* touched_out,
                           cudaStream_t stream);   /* touched_out (optional, V bytes): a second copy of the touched-row flags */
size_t b200ml_embedding_bwd_sorted_scratch_bytes(int n_tokens, int d);
int b200ml_embedding_bwd_sorted(const void* index, const void* dout, int dout_fp32, float* dtable, int n_tokens, int d,
                                int V, long long padding_idx, float dropout_p, const void* rng_state,
                                unsigned rng_site, int accumulate, void* scratch, size_t scratch_bytes,
                                cudaStream_t stream);

/* ------------------------------------------------------------------------------------------------ K6 cross-entropy
 * ref: nn.CrossEntropyLoss (distributed_cnn.py:141) and ignore_index=0 + masked mean
 *      (pytorch_machine_translator.py:125,182-188). */
int b200ml_xent_count(const long long* targets, int rows, long long ignore_index, float* inv_count2 /*[2]*/,
                      cudaStream_t stream);
/* logits are overwritten with d(loss)/d(logits) when write_grad != 0.  grad scale = *grad_scale_dev if non-NULL. */
int b200ml_xent_fwd_bwd(void* logits, int logits_fp32, long long ld_logits, const long long* targets, float* row_loss,
                        const float* grad_scale_dev, float grad_scale_host, int rows, int V, long long ignore_index,
                        int write_grad, cudaStream_t stream);
/* Large-vocabulary variant (bf16 logits, rows 16-byte aligned and padded to a multiple of 8 columns, V <= 20480): one
 * HBM read of each row, gradient written in place, and optionally colsum_out[V rounded up to 8] = column sums of that gradient (the
 * bias gradient of the projection; the padding entries are zero; colsum_accumulate != 0 adds to colsum_out, for callers that walk the rows in chunks) through `scratch` (b200ml_xent_rows_scratch_bytes(V) bytes). */
size_t b200ml_xent_rows_scratch_bytes(int V);
int b200ml_xent_rows(void* logits_bf16, long long ld_logits, const long long* targets, float* row_loss,
                     const float* grad_scale_dev, float grad_scale_host, int rows, int V, long long ignore_index,
                     int write_grad, float* colsum_out, int colsum_accumulate, void* scratch, size_t scratch_bytes,
                     cudaStream_t stream);
/* The same with the log-sum-exp of every row supplied (lse[rows]): one pass over each row, no max / sum reductions.
 * b200ml_xent_lse merges the per-tile (max, sum-exp) pairs that b200ml_gemm_bf16 leaves in `row_stats` (float2
 * [rows][ld], n_parts used per row) into lse - the softmax statistics of the translator's 19 214-wide head come out
 * of the projection GEMM's epilogue instead of two more passes over its 246 MB of logits. */
int b200ml_xent_rows_lse(void* logits_bf16, long long ld_logits, const long long* targets, float* row_loss,
                         const float* grad_scale_dev, float grad_scale_host, int rows, int V, long long ignore_index,
                         int write_grad, float* colsum_out, int colsum_accumulate, void* scratch, size_t scratch_bytes,
                         const float* lse, cudaStream_t stream);
int b200ml_xent_lse(const void* stats_float2, long long ld, int n_parts, int rows, float* lse, cudaStream_t stream);

int b200ml_loss_reduce(const float* row_loss, int rows, const float* inv_count_dev, float inv_host, float* out,
                       cudaStream_t stream);

/* ------------------------------------------------------------------------------------------------ K15/K16/K17
 * kind 1 = SGD, 2 = Adam (torch defaults semantics).  p16 (optional) receives the bf16 working copy.
 * ref: distributed_cnn.py:138,183; distributed_lstm.py:141,194; pytorch_machine_translator.py:129,196. */
int b200ml_optim_step(float* p, const void* g, int g_bf16, float* m, float* v, void* p16, long long n, int kind,
                      float lr, float beta1, float beta2, float eps, float weight_decay, const int* step_dev,
                      float grad_scale, cudaStream_t stream);
int b200ml_cast_f32_bf16(const float* src, void* dst, long long n, cudaStream_t stream);
size_t b200ml_colsum_scratch_bytes(int rows, int cols);
int b200ml_colsum_bf16(const void* x, long long ldx, int rows, int cols, float* out, int accumulate, void* scratch,
                       size_t scratch_bytes, cudaStream_t stream);
int b200ml_counters_advance(void* rng_state /*{u64 seed, u64 step}*/, int* step_dev, cudaStream_t stream);
/* ref: accuracy_fn / argmax (distributed_cnn.py:144-147,211-212) */
int b200ml_argmax_accuracy(const void* logits, int logits_fp32, long long ld, const long long* targets, int rows, int C,
                           int* pred, int* correct_count, cudaStream_t stream);

/* ------------------------------------------------------------------------------------------------ K10 attention
 * One head of q/k/v/o lives in 64 consecutive columns of a row-major bf16 matrix [B*S, ld] starting at column
 * col0 + h*head_stride, so the interleaved qkv / kv Linear outputs are consumed (and their gradients produced) in place.
 * head_dim = 64; forward needs S_kv <= 256, backward S_q, S_kv <= 256.  lse: fp32 [B, H, S_q].
 * mask_mode: 0 none, 1 reference (+1.0 where key < query), 2 causal (-inf where key > query).
 * ref: transformer.py:12-25,74-83,179-191. */
typedef struct {
    const void* ptr;
    long long ld;
    int col0;
    int head_stride;
} b200ml_attn_operand;
int b200ml_attn_fwd(const b200ml_attn_operand* q, const b200ml_attn_operand* k, const b200ml_attn_operand* v,
                    const b200ml_attn_operand* out, float* lse, int B, int H, int Sq, int Skv, float scale,
                    int mask_mode, cudaStream_t stream);
int b200ml_attn_bwd(const b200ml_attn_operand* q, const b200ml_attn_operand* k, const b200ml_attn_operand* v,
                    const b200ml_attn_operand* out, const b200ml_attn_operand* dout, const float* lse,
                    const b200ml_attn_operand* dq, const b200ml_attn_operand* dk, const b200ml_attn_operand* dv, int B,
                    int H, int Sq, int Skv, float scale, int mask_mode, cudaStream_t stream);

/* ------------------------------------------------------------------------------------------------ K1/K2/K3 conv
 * relu(conv2d(x, w, bias, stride 1, pad 1)) on fp32 NCHW, optionally followed by max_pool2d(2) (pool != 0: y and
 * pool_idx are [B,Cout,H/2,W/2], pool_idx holds the 2-bit arg-max).  Cin, Cout <= 64.
 * Backward takes the layer OUTPUT's gradient / saved value (pooled when pool != 0); dx may be NULL.
 * ref: FashionMNISTModel (distributed_cnn.py:51-73). */
int b200ml_conv3x3_relu_fwd(const float* x, const float* w, const float* bias, float* y, unsigned char* pool_idx, int B,
                            int Cin, int Cout, int H, int W, int pool, cudaStream_t stream);
/* 0 = automatic (tcgen05 implicit GEMM once 9 * C >= 64), 1 = direct CUDA-core kernels, 2 = implicit GEMM; also the
 * B200ML_CONV_ALGO environment variable.  The weight gradient always uses the direct partial-sum kernel. */
void b200ml_conv_set_algo(int algo);
size_t b200ml_conv3x3_bwd_scratch_bytes(int B, int Cin, int Cout, int H, int W);
int b200ml_conv3x3_relu_bwd(const float* dy, const float* ysaved, const unsigned char* pool_idx, const float* x,
                            const float* w, float* dx, float* dw, float* db, void* scratch, size_t scratch_bytes, int B,
                            int Cin, int Cout, int H, int W, int pool, int accumulate, cudaStream_t stream);

/* ------------------------------------------------------------------------------------------------ K9 LSTM
 * The sequential part of one LSTM layer as a persistent kernel per direction of time (hidden must be 32, fp32):
 *   forward : pre [B,T,128] = X·W_ihᵀ + b_ih + b_hh (hoisted GEMM) -> h_out [B,T,32], h_n, c_n [B,32];
 *             gates_out [B,T,128] (i,f,g,o after their nonlinearity) and c_out [B,T,32] are saved for backward
 *             (both NULL at inference).
 *   backward: dh_seq [B,T,32] (may be NULL), dh_n / dc_n -> dg_out [B,T,128] bf16 (pre-activation gate gradients, the
 *             operand of the hoisted dW_ih / dW_hh / dX / db GEMMs), dh0, dc0.
 * b200ml_dropout_cast / _bwd: (inter-layer) dropout + fp32->bf16 cast and its backward with the mask regenerated.
 * ref: nn.LSTM inside the classifier (distributed_lstm.py:118-119,131). */
int b200ml_lstm_rec_fwd(const float* pre, const float* w_hh, const float* h0, const float* c0, float* h_out,
                        float* gates_out, float* c_out, float* h_n, float* c_n, int B, int T, int hidden,
                        cudaStream_t stream);
int b200ml_lstm_rec_bwd(const float* w_hh, const float* c0, const float* gates, const float* c_seq,
                        const float* dh_seq, const float* dh_n, const float* dc_n, void* dg_out_bf16, float* dh0,
                        float* dc0, int B, int T, int hidden, cudaStream_t stream);
/* Two-layer wavefront (hidden 32): both layers of nn.LSTM(32, 32, num_layers=2, dropout=p) advance in one persistent
 * kernel per direction, layer 1 one step behind layer 0; the inter-layer dropout + bf16 cast, the layer-1 input
 * projection and their backward run inside it.  pre0 [B,T,128] is layer 0's hoisted input projection; per layer l the
 * kernel writes h_out_l [B,T,32], gates_l [B,T,128], c_seq_l [B,T,32] (gates / c_seq of both layers NULL together at
 * inference); x1_bf16 [B*T,32] = bf16(dropout(h_out_0)); h_n / c_n, h0 / c0, dh_n / dc_n, dh0 / dc0 are [2,B,32]
 * (optional).  Backward writes dg0 / dg1 [B*T,128] bf16, the operands of the hoisted weight-gradient GEMMs.
 * pre0, gates, c_seq and d_out1 must be 16-byte aligned (they are staged by bulk-async copies). */
int b200ml_lstm2_fwd(const float* pre0, const float* w_hh0, const float* w_ih1, const float* bias1, const float* w_hh1,
                     const float* h0, const float* c0, float* h_out0, float* gates0, float* c_seq0, float* h_out1,
                     float* gates1, float* c_seq1, void* x1_bf16, float* h_n, float* c_n, int B, int T, int hidden,
                     float dropout_p, const void* rng_state, unsigned rng_site1, cudaStream_t stream);
int b200ml_lstm2_bwd(const float* w_hh0, const float* w_ih1, const float* w_hh1, const float* c0, const float* gates0,
                     const float* c_seq0, const float* gates1, const float* c_seq1, const float* d_out1,
                     const float* dh_n, const float* dc_n, void* dg0_bf16, void* dg1_bf16, float* dh0, float* dc0, int B,
                     int T, int hidden, float dropout_p, const void* rng_state, unsigned rng_site1, cudaStream_t stream);
int b200ml_dropout_cast(const float* x, void* out_bf16, long long n, float p, const void* rng_state, unsigned rng_site,
                        cudaStream_t stream);
int b200ml_dropout_bwd(const void* g, int g_fp32, float* dx, long long n, float p, const void* rng_state,
                       unsigned rng_site, cudaStream_t stream);

/* ------------------------------------------------------------------------------------------------ K18 all-reduce
 * Symmetric buffers: b200ml_comm_alloc ALLOCATES (cudaMalloc + zero fill) — it is the set-up-time allocator for
 * buffers that peers map through CUDA IPC (export on the owner, import on every peer).
 * b200ml_allreduce_update: grads[r] / params[r] / w16[r] / signals[r] are rank r's flat gradient (fp32), master
 * weight (fp32), working weight (bf16) and signal-pad (2*320*8 uint32, zero-initialised) buffers as mapped in THIS
 * process; counter is a local zero-initialised uint32[2].  Elements [start, start+n) are summed over ranks in rank
 * order, divided by world, and (kind 1 SGD / 2 Adam) applied to the weights; kind 0 writes the mean to `out`.
 * algo: 0 auto, 1 one-shot, 2 two-shot with SM-issued peer loads, 3 two-shot over NVLS (multimem.ld_reduce /
 * multimem.st; auto picks it for large buckets when the multicast addresses are set), 4 two-shot with the bulk-async
 * (TMA) ring - what auto uses for buckets above 512 KB whose start / n are multiples of 8 -, 5 = 4 with per-thread
 * weight stores instead of TMA stores.  In the two-shot variants rank r owns slice r of the bucket (8-element
 * granularity when start and n are multiples of 8, else 4) and is the only rank that updates m / v for it.
 * Every rank must issue the same sequence of calls, and CTA i of a launch waits for CTA i on every peer: a
 * communicator may have only ONE of these kernels in flight at a time (use one stream); the launcher clamps the grid
 * to the number of CTAs that are resident at once, so a launch can never wait for a CTA that is not scheduled.
 * ref: DDP bucket all-reduce + optimizer.step (distributed_cnn.py:156,181-183). */
typedef struct {
    int world, rank;
    float* grads[8];
    float* params[8];
    void* w16[8];
    unsigned* signals[8];
    unsigned* counter;
    float* mc_grads;   /* NVLS multicast addresses of the gradient / master / bf16 buffers, or NULL */
    float* mc_params;
    void* mc_w16;
} b200ml_comm;
int b200ml_comm_alloc(size_t bytes, void** ptr);
int b200ml_comm_free(void* ptr);
int b200ml_ipc_export(const void* ptr, unsigned char* handle64);
int b200ml_ipc_import(const unsigned char* handle64, void** ptr);
int b200ml_ipc_close(void* ptr);
/* Row-sparse part of a bucket: [start, start + rows * row_elems) of the flat buffer is a table whose gradient rows are
 * exactly zero on rank r wherever touched[r][row] == 0 (an embedding table: b200ml_embedding_index writes the flags,
 * here into peer-mapped byte arrays).  The bulk-async transport then fetches a peer's 4 KB chunk only when that peer
 * touched one of its rows; the result is bit-identical to the dense reduction.  Other transports ignore it. */
typedef struct {
    long long start;
    int rows, row_elems;
    const unsigned char* touched[8];   /* rank r's flags as mapped in this process */
} b200ml_sparse_rows;
int b200ml_allreduce_update_sparse(const b200ml_comm* cm, long long start, long long n, int algo, int kind, float lr,
                                   float beta1, float beta2, float eps, float weight_decay, const int* step_dev,
                                   float* m, float* v, float* out, int push_master, int max_blocks,
                                   const b200ml_sparse_rows* sparse, cudaStream_t stream);
/* Same, with a second range [start2, start2 + n2) riding in the launch (bulk-async two-shot path only, n2 = 0: none):
 * sliced and updated like the bucket, its fp32 masters always pushed to every rank.  The engine sends its small fp32
 * parameters (biases, LayerNorm vectors) along with the last big bucket of a step this way: one launch, one pair of
 * cross-GPU barriers, instead of two. */
int b200ml_allreduce_update_tail(const b200ml_comm* cm, long long start, long long n, long long start2, long long n2,
                                 int algo, int kind, float lr, float beta1, float beta2, float eps, float weight_decay,
                                 const int* step_dev, float* m, float* v, float* out, int push_master, int max_blocks,
                                 const b200ml_sparse_rows* sparse, cudaStream_t stream);
int b200ml_allreduce_update(const b200ml_comm* cm, long long start, long long n, int algo, int kind, float lr,
                            float beta1, float beta2, float eps, float weight_decay, const int* step_dev, float* m,
                            float* v, float* out, int push_master, int max_blocks, cudaStream_t stream);

/* out[i] = sum over ranks of bufs[r][i] (float64, rank order: identical bits on every rank); bufs[r] is rank r's
 * buffer as mapped in this process, n even.  Same barrier protocol and constraints as b200ml_allreduce_update.
 * ref: the treeAggregate of MLlib's distributed L-BFGS cost function (mllib_multilayer_perceptron_classifier.py:35-39). */
int b200ml_allreduce_sum_f64(const b200ml_comm* cm, const double* const* bufs, long long n, double* out,
                             cudaStream_t stream);

/* ------------------------------------------------------------------------------------------------ K19 MLPC loss + gradient
 * float64 forward / softmax cross-entropy / backward of the MLlib feed-forward topology (Affine -> Sigmoid -> ... ->
 * Affine -> Softmax) over one rank's rows.  `weights` is MLlib's flat vector (per layer [W (in x out, row-major) ‖ b]);
 * layers[n_layers] are the widths; X is [rows, layers[0]] with row pitch ldx; row_w[i] = 1 / size of row i's block
 * (MLlib averages per block of blockSize rows).  out_grad_loss receives the gradient in the same flat layout followed by
 * the weighted loss: num_weights + 1 doubles.  Deterministic (no atomics).  scratch: b200ml_mlpc_scratch_bytes.
 * ref: MultilayerPerceptronClassifier(...).fit (mllib_multilayer_perceptron_classifier.py:35-39). */
size_t b200ml_mlpc_scratch_bytes(const int* layers, int n_layers, int rows);
int b200ml_mlpc_loss_grad(const double* weights, const int* layers, int n_layers, const double* X, long long ldx,
                          const long long* y, const double* row_w, int rows, double* out_grad_loss, void* scratch,
                          size_t scratch_bytes, cudaStream_t stream);

/* ------------------------------------------------------------------------------------------------ K20 L-BFGS direction
 * ref: MultilayerPerceptronClassifier(solver='l-bfgs') (mllib_multilayer_perceptron_classifier.py:35), Breeze LBFGS.
 * Two-loop recursion over k <= 32 history pairs as one cooperative kernel; float64, rank-reproducible.
 * S, Y: [k][ld] oldest first.  n <= 148 * 256 * 16 (606 208) on a B200; d must not alias g. */
size_t b200ml_lbfgs_scratch_bytes(void);
int b200ml_lbfgs_direction(const double* g, const double* S, const double* Y, int k, long long n, long long ld,
                           double* d, void* scratch, size_t scratch_bytes, cudaStream_t stream);

/* ------------------------------------------------------------------------------------------------ gradient buckets
 * Host-side bucket manager of the data-parallel engine (csrc/runtime/reducer.cpp): the role of torch's c10d Reducer
 * behind DistributedDataParallel (distributed_cnn.py:156).  offsets[n_params + 1] are element offsets into the flat
 * gradient buffer; group_ends[n_groups] are parameter-index boundaries buckets must not cross. */
typedef struct b200ml_reducer b200ml_reducer;
b200ml_reducer* b200ml_reducer_create(int n_params, const long long* offsets, int n_groups, const int* group_ends,
                                      long long cap_elems);
void b200ml_reducer_destroy(b200ml_reducer* r);
int b200ml_reducer_num_buckets(const b200ml_reducer* r);
int b200ml_reducer_bucket(const b200ml_reducer* r, int b, long long* start, long long* numel, int* first_param,
                          int* n_params, int* group);
int b200ml_reducer_bucket_of(const b200ml_reducer* r, int param);
void b200ml_reducer_begin_step(b200ml_reducer* r);
int b200ml_reducer_mark_ready(b200ml_reducer* r, int param); /* completed bucket index, or -1 */
int b200ml_reducer_flush(b200ml_reducer* r, int* out_buckets, int cap);

#ifdef __cplusplus
}
#endif
#endif /* B200ML_H_ */
