/*
 * b200attn.h -- C ABI of the B200-native attention hot path for nano-vllm-v1.
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch types, no C++
 * exceptions.  The reference is pure Python; what it binds for this path today are
 * (a) a Triton kernel and (b) two flash-attn entry points plus three inductor-compiled
 * pointwise modules.  Each function below names the reference interface it replaces.
 * A maintainer binds them with ctypes (INTEGRATION.md shows the stub); the shipped
 * host-side mirror is nano-vllm-v1_b200/b200attn/.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - element type is bf16 (2 bytes) for q/k/v/caches/out, int32 for index arrays,
 *     int64 for positions, fp32 for the RoPE cache (the reference's own dtypes,
 *     SURVEY.md 2.4); head_dim must be 128 and the KV page 256 tokens;
 *   - *_stride arguments are in ELEMENTS; inner (head_dim) stride is 1 and the head
 *     stride is head_dim, exactly what attention.py:121-122 asserts;
 *   - `stream` is a cudaStream_t passed as void*; all work is stream-ordered, nothing is
 *     allocated, freed or retained, no host synchronisation happens inside;
 *   - return value: 0 = ok; negative = argument error (B200_ERR_*); positive = cudaError_t
 *     of a failed launch.  b200_status_string() describes either.
 *   - graph-capture safe: every length is read from device memory; launch geometry
 *     depends only on host integers.
 */
#ifndef B200ATTN_H_
#define B200ATTN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200ATTN_VERSION 100 /* 0.1.0 */

#define B200_OK 0
#define B200_ERR_NULL (-1)       /* required pointer is NULL */
#define B200_ERR_SHAPE (-2)      /* unsupported head_dim / page size / head counts */
#define B200_ERR_ALIGN (-3)      /* pointer or stride not 16-byte aligned */
#define B200_ERR_WORKSPACE (-4)  /* workspace too small for (batch, heads, splits) */
#define B200_ERR_RANGE (-5)      /* negative or overflowing count */
#define B200_ERR_UNAVAILABLE (-6)/* requested kernel variant not built / no driver entry point */
#define B200_ERR_SCRATCH_LIMIT (-7) /* problem needs more than the library's built-in scratch array */

int b200_version(void);
const char* b200_status_string(int status);

/* Number of SMs of the current device (cached); helper for split planning. */
int b200_sm_count(void);

/* ---- (f3) small-message tensor-parallel all-reduce over NVLink peer memory ----------------------
 * Replaces dist.all_reduce(y) in RowParallelLinear.forward (nanovllm/layers/linear.py:407-410) for
 * bf16 messages up to the `max_msg_bytes` the communicator was created with; larger messages stay on
 * the reference's NCCL call.  One process per GPU: every rank b200_ar_alloc()s one buffer, the 64-byte
 * IPC handles are exchanged by the host (the drop-in uses torch.distributed.all_gather_object on the
 * reference's own process group), peers are mapped with b200_ar_open(), and b200_ar_create() builds the
 * communicator (peer_ptrs[rank] = the local buffer).  b200_allreduce_bf16 sums `data` (n_elems bf16,
 * multiple of 8, 16-byte aligned) over all ranks IN PLACE, one kernel launch, stream-ordered, CUDA-graph
 * capturable; all ranks must call it in the same order with the same n_elems.  two_shot: 0 = one-shot, 1 = reduce-scatter +
 * all-gather with peer loads, 2 = the same with peer stores (push), 3 = one-shot with
 * peer stores (needs world x message <= 2 x max_msg_bytes); 1 and 2 need world > 2.  Results are bit-identical on all ranks. */
size_t b200_ar_buffer_bytes(size_t max_msg_bytes);
int b200_ar_alloc(size_t bytes, void** dev_ptr, void* ipc_handle_out64);
int b200_ar_open(const void* ipc_handle64, void** peer_ptr);
int b200_ar_create(int rank, int world, void* const* peer_ptrs, size_t max_msg_bytes, void** comm_out);
int b200_allreduce_bf16(void* comm, void* data, int64_t n_elems, int two_shot, void* stream);
/* Tear-down, the counterpart of the reference's dist.destroy_process_group (engine/model_runner.py:187-193):
 * closes the peers' IPC mappings, frees the local buffer (from b200_ar_alloc) and the communicator.  Synchronise
 * the device and the group first.  b200_ar_status: 0 = healthy, 1 = a wait inside some launch exceeded
 * B200ATTN_AR_TIMEOUT_S (default 30 s): the kernels give up and flag it instead of trapping the context;
 * synchronises the device (tests, tools, shutdown).  All launches are cooperative (co-resident blocks). */
int b200_ar_destroy(void* comm);
int b200_ar_status(void* comm, int32_t* status_out);

/* ---- (f3) all-reduce FUSED with the residual add + RMSNorm that follows it, on NVLS multicast memory ----
 * Replaces, in ONE launch per rank, dist.all_reduce(y) of RowParallelLinear.forward (layers/linear.py:407-410)
 * AND the RMSNorm.add_rms_forward that consumes y (layers/layernorm.py:95-141, called at models/qwen3.py:370,374,440):
 *     s        = sum over ranks of x            (in-switch: multimem.ld_reduce, fp32 accumulate, one bf16 rounding,
 *                                                 i.e. the value dist.all_reduce would have produced)
 *     v        = float(s) + float(residual);   residual <- bf16(v)           (in place)
 *     out      = bf16( bf16(v * rsqrt(mean(v^2) + eps)) * weight )
 * `local_buf` / `multicast_buf`: the unicast and the multicast mapping of ONE symmetric allocation of
 * b200_nvls_buffer_bytes(max_msg_bytes) bytes per rank (cuMemCreate + cuMulticastBindMem; the drop-in gets both
 * from torch.distributed._symmetric_memory on the reference's own process group).  world == 1 needs no multicast
 * mapping (pass NULL).  After b200_nvls_create on every rank the host must run a group barrier before the first
 * call.  x [T, hidden] bf16: this rank's partial sums (any device memory); residual [T, hidden] bf16 updated in
 * place, or NULL (then v = float(s)); weight [hidden] bf16, or NULL together with residual for a plain all-reduce
 * (out <- s).  peer_bufs[world]: the unicast mappings of every rank's buffer (peer_bufs[rank] == local_buf), or NULL
 * (then form 3 is unavailable).  b200_nvls_create zeroes the local buffer.
 * form: 0 = reduce-scatter (multimem.ld_reduce) + multicast all-gather (multimem.st), two barriers;
 *       1 = every rank multimem.ld_reduces every row itself, one barrier;
 *       2 = "LL" one-shot: no barrier -- every rank multicasts {data, epoch} packets, readers spin on the epoch that
 *           rides in the same 16-byte store as the data; sums in rank order (needs 2 x world x message <= one LL slot);
 *       3 = "LL" two-shot: no barrier -- reduce-scatter by peer stores of packets, owner sums in rank order and
 *           multicasts result packets (messages <= 4 MB).
 *     Forms 2/3 move twice the bytes and save every barrier (~6 us each); their results are bit-identical on all
 *     ranks and equal the fp32 sum in rank order rounded once; forms 0/1 round the switch's own fp32 sum once.
 *     All ranks must pass the same shapes and form.
 * Stream-ordered, CUDA-graph capturable (slots alternate with a device-resident epoch), cooperative launch. */
size_t b200_nvls_buffer_bytes(size_t max_msg_bytes);
int b200_nvls_create(int rank, int world, void* local_buf, void* multicast_buf, void* const* peer_bufs,
                     size_t max_msg_bytes, void** comm_out);
int b200_nvls_allreduce_add_rmsnorm(void* comm, const void* x, void* residual, const void* weight, float eps, void* out,
                                    int32_t num_tokens, int32_t hidden, int32_t form, void* stream);
int b200_nvls_status(void* comm, int32_t* status_out);
int b200_nvls_destroy(void* comm);

/* Debug/profiling aid (no reference counterpart): while `dev_buf` (zeroed int64[6][max_tiles][8] in
 * device memory) is non-NULL, b200_attn_prefill_planned runs an instrumented build whose CTA 0 stamps
 * clock64 at the phase boundaries of its first `max_tiles` tiles (rows 0/1: softmax warpgroups: wait,
 * S ready, S loaded, max done, exp done, P published; rows 2/3: MMA issuer: wait P, P seen, issued,
 * unit boundaries; rows 4/5: epilogue phases of a unit's last tile).
 * Pass NULL to return to the product kernel.  tools/trace_prefill.py reads it. */
int b200_debug_prefill_trace(void* dev_buf, int max_tiles);

/* ---------------------------------------------------------------------------------
 * (1) KV-cache scatter.  Replaces store_kvcache_kernel + store_kvcache,
 *     nanovllm/layers/attention.py:34-90 and :93-127.
 *     k_cache.view(-1, Hkv*D)[slot_mapping[i]] = key[i]; same for value; slot -1 skipped.
 *     key/value: [T, Hkv, D] with row strides (value is a strided view of qkv).
 *     Bit-exact copy; 64-bit addressing (the Triton kernel's 32-bit `slot * D` is not kept).
 */
int b200_kv_scatter(const void* key, const void* value,
                    int64_t key_row_stride, int64_t value_row_stride,
                    void* k_cache, void* v_cache, const int32_t* slot_mapping,
                    int32_t num_tokens, int32_t num_kv_heads, int32_t head_dim,
                    void* stream);

/* ---------------------------------------------------------------------------------
 * (3) Paged decode attention (split-K + reduce).  Replaces flash_attn_with_kvcache as
 *     called at attention.py:229-232, and flash_attn_varlen_func (attention.py:242-248)
 *     for batches whose every row has one query token (eager decode).
 *     q/out: [B, Hq, D] (token stride given); caches [num_blocks, 256, Hkv, D];
 *     context_lens int32[B] (0 => row is graph padding, output 0);
 *     block_tables int32[B, >=ceil(len/256)] with row stride bt_row_stride; entries past
 *     the context (-1 eager, 0 graph) are never dereferenced.
 *     num_splits: from b200_decode_plan (host-only heuristic => static under CUDA graphs).
 *     workspace: b200_decode_workspace_bytes(B, Hq, num_splits) bytes, may be NULL iff
 *     num_splits == 1.  Its first B200_DECODE_TICKET_BYTES bytes are arrival tickets (one int32 per
 *     (row, kv head)): they must be ZERO before the first launch that uses the buffer and every launch
 *     leaves them zero, so one zero-initialised buffer serves all calls and CUDA-graph replays.  The
 *     split that arrives last at a (row, kv head) merges all of them with LSE weights inside the same
 *     launch; there is no separate reduce launch (B200ATTN_DECODE_FUSED_REDUCE=0 restores it; it is
 *     also used when B * Hkv tickets do not fit the region).
 */
#define B200_DECODE_TICKET_BYTES (64 * 1024)
int b200_decode_plan(int32_t batch, int32_t num_kv_heads, int32_t max_context_len);
size_t b200_decode_workspace_bytes(int32_t batch, int32_t num_heads, int32_t num_splits);
int b200_attn_decode(const void* q, int64_t q_token_stride,
                     const void* k_cache, const void* v_cache,
                     void* out, int64_t out_token_stride,
                     const int32_t* context_lens,
                     const int32_t* block_tables, int32_t bt_row_stride,
                     int32_t batch, int32_t num_heads, int32_t num_kv_heads,
                     int32_t head_dim, int32_t page_size, float softmax_scale,
                     int32_t num_splits, void* workspace, size_t workspace_bytes,
                     void* stream);

/* (3, no-allocation form) The same paged decode for a caller that has no device memory to give as workspace.
 *     No counterpart in the reference: there, flash-attn asks torch's caching allocator for its split
 *     accumulators on every call, which a torch-free host cannot do.  This entry point is for the callers of
 *     THIS library that cannot allocate: an engine that has already turned all free HBM into KV cache and can no
 *     longer rely on cudaMalloc succeeding; a cgo / JNI / plain-ctypes binding that owns no device allocator;
 *     a request loop that cannot afford the device-wide stall of cudaMalloc / cudaFree.
 *     Arguments and results are those of b200_attn_decode minus (workspace, workspace_bytes).  The split-K
 *     scratch is an array of B200_DECODE_STATIC_SCRATCH_BYTES (8 MiB) built into the library: every GPU the
 *     library is loaded on holds those bytes until the process exits, whether or not this function is called.
 *     Limits and rules:
 *       - b200_decode_workspace_bytes(batch, num_heads, num_splits) must be <= B200_DECODE_STATIC_SCRATCH_BYTES,
 *         else B200_ERR_SCRATCH_LIMIT is returned and nothing is launched (lower num_splits, or use
 *         b200_attn_decode with your own workspace).  num_splits == 1 uses no scratch and always fits.
 *       - The function itself allocates and frees nothing (no cudaMalloc/cudaFree, no allocating library call).
 *       - The first call on each device looks up that device's address of the array (cudaGetSymbolAddress, with
 *         the device current) and opts the kernel into its shared-memory size.  Make that first call outside any
 *         stream capture.
 *       - All calls on a device share the one array, and it keeps the previous call's contents.  Each call
 *         clears the arrival tickets it is about to use (a small stream-ordered memset); everything else is
 *         overwritten before it is read.  Two calls whose GPU work can overlap (different streams without an
 *         ordering between them, or two host threads) race on the array and give wrong results: order them. */
#define B200_DECODE_STATIC_SCRATCH_BYTES (8 * 1024 * 1024)
int b200_attn_decode_static_scratch(const void* q, int64_t q_token_stride,
                                    const void* k_cache, const void* v_cache,
                                    void* out, int64_t out_token_stride,
                                    const int32_t* context_lens,
                                    const int32_t* block_tables, int32_t bt_row_stride,
                                    int32_t batch, int32_t num_heads, int32_t num_kv_heads,
                                    int32_t head_dim, int32_t page_size, float softmax_scale,
                                    int32_t num_splits, void* stream);

/* ---------------------------------------------------------------------------------
 * (2) Varlen causal prefill / chunked-prefill attention.  Replaces
 *     flash_attn_varlen_func as called at attention.py:242-248:
 *       block_tables == NULL : k, v are packed [Tk, Hkv, D] with token strides
 *                              (sub-paths A/B of SURVEY 3.4; v is a strided view);
 *       block_tables != NULL : k, v are the caches [num_blocks, 256, Hkv, D]
 *                              (token strides ignored), rows gathered through the table.
 *     Bottom-right aligned causal mask (query i sees keys j <= i + Lk - Lq).
 *     cu_seqlens_q/k int32[B+1] on device; max_seqlen_q/k host ints as in the reference.
 *     variant: 0 = auto, 1 = mma.sync kernel, 2 = tcgen05/TMEM/TMA kernel.
 */
int b200_attn_prefill(const void* q, int64_t q_token_stride,
                      const void* k, const void* v,
                      int64_t k_token_stride, int64_t v_token_stride,
                      void* out, int64_t out_token_stride,
                      const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                      const int32_t* block_tables, int32_t bt_row_stride,
                      int32_t batch, int32_t total_q, int32_t max_seqlen_q, int32_t max_seqlen_k,
                      int32_t num_heads, int32_t num_kv_heads, int32_t head_dim,
                      int32_t page_size, int64_t num_cache_blocks, float softmax_scale,
                      int32_t variant, void* stream);

/* ---------------------------------------------------------------------------------
 * (2'/3') Planned varlen attention for mixed batches.  The reference hands
 *     flash_attn_varlen_func one call per layer whatever the mix of rows (attention.py:242-248);
 *     here a step plan built ON THE DEVICE from cu_seqlens routes
 *       rows with q_len > 1  -> persistent tcgen05 prefill kernel (per-unit work table sorted heaviest-first),
 *       rows with q_len == 1 -> split-K decode kernel (only when a block table exists),
 *     and is reused by every layer of the step.  No host reads, graph-capturable.
 *     plan: b200_attn_plan_bytes(batch, total_q) bytes of device memory, 16-byte aligned; it is built for
 *     one (num_heads, num_kv_heads) pair (the query-tile height depends on the GQA group being even).
 *     b200_attn_prefill_planned / b200_attn_decode_varlen write disjoint rows of `out`.
 */
size_t b200_attn_plan_bytes(int32_t batch, int32_t total_q);
int b200_attn_plan(const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k, int32_t batch, int32_t total_q,
                   int32_t num_heads, int32_t num_kv_heads, int32_t route_decode,
                   void* plan, size_t plan_bytes, void* stream);
int b200_attn_prefill_planned(const void* q, int64_t q_token_stride,
                              const void* k, const void* v,
                              int64_t k_token_stride, int64_t v_token_stride,
                              void* out, int64_t out_token_stride,
                              const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                              const int32_t* block_tables, int32_t bt_row_stride,
                              int32_t batch, int32_t total_q, int32_t max_seqlen_q, int32_t max_seqlen_k,
                              int32_t num_heads, int32_t num_kv_heads, int32_t head_dim,
                              int32_t page_size, int64_t num_cache_blocks, float softmax_scale,
                              const void* plan, size_t plan_bytes, void* stream);
int b200_attn_decode_varlen(const void* q, int64_t q_token_stride,
                            const void* k_cache, const void* v_cache,
                            void* out, int64_t out_token_stride,
                            const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                            const int32_t* block_tables, int32_t bt_row_stride,
                            int32_t batch, int32_t num_heads, int32_t num_kv_heads,
                            int32_t head_dim, int32_t page_size, float softmax_scale,
                            int32_t num_splits, void* workspace, size_t workspace_bytes,
                            const void* plan, size_t plan_bytes, void* stream);

/* ---------------------------------------------------------------------------------
 * (4a) Fused per-head RMSNorm + RoPE on q and k.  Replaces q_norm/k_norm
 *      (layers/layernorm.py:57-93 via models/qwen3.py:191-193) followed by
 *      RotaryEmbedding.forward (layers/rotary_embedding.py:114-153), keeping the two
 *      bf16 roundings between them.  q [T,Hq,D], k [T,Hkv,D] are views with token strides;
 *      q_out/k_out are contiguous [T,H,D]; positions int64[T]; cos_sin fp32 [max_pos, D].
 *      q_weight / k_weight NULL => that tensor is only rotated (stand-alone RotaryEmbedding.forward).
 */
int b200_qk_norm_rope(const void* q, int64_t q_token_stride,
                      const void* k, int64_t k_token_stride,
                      const int64_t* positions, const float* cos_sin_cache,
                      const void* q_weight, const void* k_weight, float eps,
                      void* q_out, void* k_out,
                      int32_t num_tokens, int32_t num_heads, int32_t num_kv_heads,
                      int32_t head_dim, void* stream);

/* (4a + 1) The same q/k RMSNorm + RoPE, and in the SAME launch the KV-cache scatter of (1)
 *      (SURVEY.md 8 f2): the rotated k row goes to k_out (may be NULL when nothing reads fresh k:
 *      paged and decode calls read the cache) AND to k_cache[slot_mapping[t]]; the untouched v row
 *      is copied to v_cache[slot_mapping[t]]; slot -1 skips both cache writes
 *      (layers/attention.py:79-80).  Replaces, in one pass over qkv, models/qwen3.py:191-195 plus
 *      store_kvcache (layers/attention.py:93-127, called at :219-220).  Caches are
 *      [num_slots, Hkv, D] contiguous.  The launch completes (stream order) before any attention
 *      kernel reads the cache, so same-step prefix sharing is unaffected.
 */
int b200_qk_norm_rope_store(const void* q, int64_t q_token_stride,
                            const void* k, int64_t k_token_stride,
                            const void* v, int64_t v_token_stride,
                            const int64_t* positions, const float* cos_sin_cache,
                            const void* q_weight, const void* k_weight, float eps,
                            void* q_out, void* k_out,
                            void* k_cache, void* v_cache, const int32_t* slot_mapping,
                            int32_t num_tokens, int32_t num_heads, int32_t num_kv_heads,
                            int32_t head_dim, void* stream);

/* (4a') RMSNorm over the hidden size with optional fused residual add.  Replaces
 *      RMSNorm.rms_forward / add_rms_forward (layernorm.py:57-141).  residual may be NULL;
 *      when given it is updated IN PLACE to round(x + residual), as the reference returns. */
int b200_rmsnorm(const void* x, void* residual, const void* weight, float eps,
                 void* out, int32_t num_tokens, int32_t hidden, void* stream);

/* (4b) SiluAndMul.  Replaces layers/activation.py:36-58: out[T,I] = silu(x[:, :I]) * x[:, I:],
 *      fp32 math, one rounding (the compiled form the reference runs on the GPU). */
int b200_silu_and_mul(const void* x, void* out, int32_t num_tokens, int32_t inter, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200ATTN_H_ */
