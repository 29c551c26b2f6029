/* C ABI of lib/libb200.so — the sm_100a kernel library of accelerate_hpc_test_b200.
 *
 * Every entry point returns 0 on success, a positive cudaError_t, or a negative B200_ERR_* code
 * (b200_error_string() turns either into text).  Pointers are device pointers unless stated; `stream`
 * is a cudaStream_t.  bf16 tensors are row-major with 16-byte aligned rows.  The Python layer
 * (accelerate_hpc_test_b200/ops/_lib.py) binds exactly these symbols with ctypes.
 *
 * Reference parity: the reference (hang-zou/accelerate-hpc-test) ships no native code; each group below
 * names the SURVEY.md §2.6 row (K* kernels, C* collectives) it stands in for.
 */
#ifndef B200_OPS_H
#define B200_OPS_H

#ifdef __cplusplus
extern "C" {
#endif

#define B200_ERR_BAD_ARG (-1)     /* shape / alignment not supported by the kernel */
#define B200_ERR_UNSUPPORTED (-2) /* configuration not supported (e.g. head_dim != 128) */
#define B200_ERR_TOO_LARGE (-3)   /* size exceeds the built-in limit of the entry point */
#define B200_ERR_DRIVER (-4)      /* driver entry point / tensor-map encode failed */

/* ---- runtime (csrc/runtime.cu) */
int b200_version(void);
const char* b200_error_string(int code);
unsigned long long b200_launch_count(void); /* kernels launched by this library in this process */
void b200_reset_launch_count(void);
int b200_device_info(int* sm_count, int* cc_major, int* cc_minor, long long* total_mem);
int b200_symm_alloc(void** ptr, long long bytes); /* cudaMalloc + zero: one symmetric buffer per rank */
int b200_symm_free(void* ptr);
int b200_ipc_get_handle(void* ptr, void* handle64);       /* 64-byte CUDA IPC handle (host memory) */
int b200_ipc_open_handle(const void* handle64, void** ptr);
int b200_ipc_close_handle(void* ptr);
int b200_memset_async(void* ptr, int value, long long bytes, void* stream);
int b200_memcpy_async(void* dst, const void* src, long long bytes, void* stream);

/* ---- K1: bf16 GEMM on tcgen05 + TMEM + TMA (csrc/gemm_tcgen05.cu)
 * D[M,N] (+)= sum_k A(m,k) B(n,k) (+ bias[n]).  a_mn_major==0: A is [M,K] (row stride lda); 1: A is [K,M].
 * b_mn_major==0: B is [N,K]; 1: B is [K,N].  D is bf16 or fp32 (out_fp32), row stride ldc.
 * max_ctas > 0 caps the persistent grid. */
int b200_gemm_bf16(const void* A, const void* B, void* C, const void* bias, int M, int N, int K, long long lda, long long ldb,
                   long long ldc, int a_mn_major, int b_mn_major, int out_fp32, int accumulate, int max_ctas, void* stream);

/* Same contract on an SM pair: cluster of 2 CTAs, tcgen05.mma.cta_group::2, 256x256 tiles, 6 TMA stages
 * (csrc/gemm2_tcgen05.cu).  The Python layer uses this one for plain GEMMs. */
int b200_gemm2_bf16(const void* A, const void* B, void* C, const void* bias, int M, int N, int K, long long lda, long long ldb,
                    long long ldc, int a_mn_major, int b_mn_major, int out_fp32, int accumulate, void* stream);

/* ---- tensor-parallel fused compute+collective GEMMs over peer memory (csrc/gemm_tcgen05.cu)
 * all-gather -> GEMM: every rank owns `rows` rows of A at byte offset a_off inside its symmetric workspace
 * (peer_ptrs[r] = base of rank r's workspace mapped here).  A gather warp inside the GEMM pulls the remote shards
 * into a_full with TMA bulk copies while the tensor cores already work on the local shard; ready[s] is bumped to
 * `epoch` when shard s has landed.  C[world*rows, N] = gathered(A) @ B^T. */
int b200_gemm_bf16_allgather_a(const void* const* peer_ptrs, int rank, int world, long long a_off, void* a_full, int rows,
                               const void* B, void* C, int N, int K, long long ldb, long long ldc, int b_mn_major,
                               unsigned* ready, unsigned long long epoch, void* stream);
/* GEMM -> reduce-scatter: the epilogue stores the fp32 tile rows owned by rank d straight into d's slot `rank`
 * (at byte offset out_off of d's workspace); after a device barrier each rank sums its `world` slots. */
int b200_gemm_bf16_scatter(const void* A, const void* B, void* const* peer_ptrs, int rank, int world, long long out_off,
                           int rows, int N, int K, long long lda, long long ldb, int a_mn_major, int b_mn_major, void* stream);
int b200_reduce_slots(const void* slots, void* out_bf16, int world, long long elems_per_slot, void* stream);

/* SM-pair versions of the two fused tensor-parallel GEMMs (csrc/gemm2_tcgen05.cu); same contracts as above */
int b200_gemm2_bf16_allgather_a(const void* const* peer_ptrs, int rank, int world, long long a_off, void* a_full, int rows,
                                const void* B, void* C, int N, int K, long long ldb, long long ldc, int b_mn_major,
                                unsigned* ready, unsigned long long epoch, void* stream);
int b200_gemm2_bf16_scatter(const void* A, const void* B, void* const* peer_ptrs, int rank, int world, long long out_off,
                            int rows, int N, int K, long long lda, long long ldb, int a_mn_major, int b_mn_major, void* stream);

/* ---- K2: fp8 (e4m3) with one scale per tensor (csrc/fp8.cu, csrc/gemm2_tcgen05.cu)
 * amax[0] = max(amax[0], max|x|); q = sat(x * 448/amax), qt = q^T when qt != NULL, deq[0] = amax/448;
 * D = deq_a * deq_b * A_q B_q^T on tcgen05.mma.kind::f8f6f4 (both operands K-major byte matrices). */
int b200_amax_bf16(const void* x, long long n, void* amax, void* stream);
int b200_quantize_e4m3(const void* x, long long ldx, int R, int C, const void* amax, void* q, long long ldq, void* qt,
                       long long ldqt, void* deq, void* stream);
int b200_gemm2_e4m3(const void* A, const void* B, void* C, const void* bias, const void* deq_a, const void* deq_b, int M, int N,
                    int K, long long lda, long long ldb, long long ldc, int out_fp32, int accumulate, void* stream);

/* ---- K3: MXFP8 — e4m3 elements, one UE8M0 scale per 32 elements along K, applied inside the tensor core
 * (tcgen05.mma.kind::mxf8f6f4.block_scale; csrc/gemm_mxfp8.cu).  Scale buffers: 512-byte blocks, one per (128 rows, 128 k):
 * byte (r % 32) * 16 + ((r / 32) % 4) * 4 + g holds the scale of row r, k-group g.  K % 128 == 0.
 * b200_gemm2_mxfp8 is the SM-pair kernel (256x256 tiles); its scale buffers must cover whole 256-row tiles. */
long long b200_mx_sf_bytes(long long rows, long long k);
int b200_quantize_mx_rowwise(const void* x, long long ldx, int R, int C, void* q, long long ldq, void* sf, void* stream);
int b200_quantize_mx_colwise(const void* x, long long ldx, int R, int C, void* qt, long long ldqt, void* sft, void* stream);
int b200_gemm_mxfp8(const void* A, const void* B, const void* sfa, const void* sfb, void* C, const void* bias, int M, int N, int K,
                    long long lda, long long ldb, long long ldc, int out_fp32, int accumulate, void* stream);
int b200_gemm2_mxfp8(const void* A, const void* B, const void* sfa, const void* sfb, void* C, const void* bias, int M, int N, int K,
                     long long lda, long long ldb, long long ldc, int out_fp32, int accumulate, void* stream);

/* ---- C6: Ulysses sequence <-> head exchange (csrc/comm.cu): the all-to-all with both layout changes folded into the copy.
 * mode 0: peers hold [B*S_loc, src_pitch]; column block `rank` of every segment lands at row (b*S + p*S_loc + r) of out.
 * mode 1: peers hold [B*S, src_pitch]; rows (b*S + rank*S_loc + r) land in column block p of every segment of out.
 * segs: host array of nseg x {src_base, dst_base, bytes} (all multiples of 16). */
int b200_seq_head_exchange(const void* const* peer_ptrs, int rank, int world, long long src_off, void* out, int mode, int B,
                           int S_loc, long long src_pitch, long long dst_pitch, const long long* segs, int nseg, int ch_arrive,
                           int ch_done, unsigned epoch, void* counter, int ctas, void* stream);

/* ---- K4: flash attention, head_dim 128, GQA, causal or full (csrc/flash_attn.cu)
 * q/k/v/o are [B*S, heads*128] views (row strides in elements); lse/delta are fp32 [B, Hq, S]. */
int b200_flash_attn_fwd(const void* q, const void* k, const void* v, void* o, void* lse, int B, int S, int Hq, int Hkv, int D,
                        long long q_stride, long long k_stride, long long v_stride, long long o_stride, float scale, int causal,
                        void* stream);
int b200_flash_attn_bwd_prep(const void* dout, const void* out, void* delta, int B, int S, int Hq, long long do_stride,
                             long long o_stride, void* stream);
/* dq_acc: fp32 [B, Hq, S_pad, 128] scratch (S_pad = S rounded up to 64), zero-initialised by the caller; every 64-row
 * dQ tile is added with one cp.reduce.async.bulk; b200_dq_finalize scales it and writes token-major bf16. */
int b200_flash_attn_bwd(const void* q, const void* k, const void* v, const void* dout, const void* lse, const void* delta,
                        void* dq_acc, void* dk, void* dv, int B, int S, int Hq, int Hkv, int D, long long q_stride,
                        long long k_stride, long long v_stride, long long do_stride, long long dkv_stride, float scale,
                        int causal, void* stream);
int b200_dq_finalize(const void* acc, void* dst, int B, int S, int Hq, long long dst_stride, float scale, void* stream);

/* ---- K5 / K6 / K7 / K11 / K12: bandwidth-bound fused ops (csrc/elementwise.cu) */
int b200_rmsnorm_fwd(const void* x, const void* res, const void* w, void* y, void* sum_out, void* rstd, long long T, int H,
                     float eps, void* stream);
int b200_rmsnorm_bwd_parts(long long T); /* rows of fp32 scratch b200_rmsnorm_bwd needs: [parts, H] */
int b200_rmsnorm_bwd(const void* dy, const void* x, const void* w, const void* rstd, const void* dres, void* dx, void* dw,
                     void* dw_partial_scratch, long long T, int H, int accumulate_dw, void* stream);
/* No counterpart in the reference.  Same result as b200_rmsnorm_bwd for hosts that are not allowed to allocate
 * where they call it (device heap owned by another allocator with a fixed budget, or a call made from inside that
 * allocator's callback): the [parts, H] scratch is a fixed 8 MiB __device__ array inside the library instead of a
 * caller buffer.  H <= 8192, else B200_ERR_TOO_LARGE; at most 256 partial rows.  The first call on a device looks up
 * the array's device address (cudaGetSymbolAddress) and must come before any stream capture on that device.  All calls
 * share the array: nothing needs clearing between calls, but calls that overlap on the GPU race on it. */
int b200_rmsnorm_bwd_noalloc(const void* dy, const void* x, const void* w, const void* rstd, const void* dres, void* dx, void* dw,
                             long long T, int H, int accumulate_dw, void* stream);
int b200_rope_inplace(void* buf, const void* cos_tab, const void* sin_tab, const void* pos, long long T, long long row_stride,
                      int n_heads, int D, int backward, void* stream);
int b200_swiglu_fwd(const void* gate_up, void* act, long long T, int I, void* stream);
int b200_swiglu_bwd(const void* gate_up, const void* dact, void* dgate_up, long long T, int I, void* stream);
int b200_embedding_fwd(const void* ids, const void* table, void* out, long long T, int H, long long vocab, void* stream);
int b200_embedding_bwd(const void* ids, const void* dout, void* dtable, long long T, int H, long long vocab, void* stream);
int b200_scale_cast(const void* in, void* out, long long n, float scale, int kind, void* stream); /* 0 f32->bf16 1 bf16->f32 2 f32 3 bf16 */
int b200_add_inplace_bf16(void* y, const void* a, long long n, void* stream);
/* out (+)= in * (*scale), bf16, n % 8 == 0, 16-byte aligned pointers; `scale` is a device fp32 scalar (no host sync) */
int b200_scale_by_scalar_bf16(const void* in, void* out, long long n, const void* scale, int accumulate, void* stream);

/* ---- K8: cross-entropy over a logits chunk, forward + in-place gradient (csrc/cross_entropy.cu) */
int b200_ce_fwd_bwd(void* logits, const void* labels, void* loss_rows, void* lse_rows, long long T, long long ld, int V,
                    const void* inv_count, float scale, long long ignore_index, int write_grad, void* stream);
int b200_count_valid(const void* labels, long long n, long long ignore_index, void* inv_count, void* count_out, void* stream);
int b200_loss_reduce(const void* rows, long long n, const void* inv_count, float scale, void* out, int accumulate, void* stream);

/* ---- K9 / K10: fused AdamW, gradient norm, clip coefficient (csrc/optim.cu) */
int b200_adamw_flat(void* master, void* m, void* v, const void* grad, void* shadow_bf16, long long n, int grad_is_bf16, float lr,
                    float beta1, float beta2, float eps, float weight_decay, int step, float grad_scale, const void* clip_coef,
                    const void* found_inf, void* stream);
int b200_adamw_multi(const void* table, int ntensors, long long total_chunks, int grad_is_bf16, float lr, float beta1, float beta2,
                     float eps, float weight_decay, int step, float grad_scale, const void* clip_coef, const void* found_inf,
                     void* stream);
int b200_sumsq(const void* g, long long n, int is_bf16, void* out2, void* stream);
int b200_clip_coef(const void* sumsq, float max_norm, void* norm_out, void* coef_out, void* stream);

/* ---- C1 / C2 / C3 / C6: collectives over peer-mapped memory (csrc/comm.cu)
 * peer_ptrs: HOST array of `world` device pointers, the base of each rank's symmetric buffer as mapped
 * into this process.  Channels index the signal pad at the start of every buffer; epochs grow by one
 * per use of a channel.  `counter` is a zero-initialised uint32 in local device memory per channel. */
int b200_comm_limits(int* max_ranks, int* num_channels, long long* signal_pad_bytes);
int b200_comm_set_timeout_ms(long long ms);
int b200_barrier_all(const void* const* peer_ptrs, int rank, int world, int channel, unsigned epoch, void* stream);
int b200_allgather_pull(const void* const* peer_ptrs, int rank, int world, long long shard_off, long long shard_bytes, void* out,
                        int wait_channel, unsigned wait_epoch, int ctas, void* stream);
int b200_reduce_scatter_adamw(const void* const* peer_ptrs, int rank, int world, long long grad_off, long long n, void* master,
                              void* m, void* v, long long param_off, void* grad_out, float lr, float beta1, float beta2, float eps,
                              float weight_decay, int step, float grad_scale, int do_adam, const void* clip_coef, int ch_arrive,
                              int ch_done, unsigned epoch, void* counter, int ctas, void* stream);
int b200_allreduce(const void* const* peer_ptrs, int rank, int world, long long off, long long n, int is_bf16, float scale,
                   int ch_arrive, int ch_done, unsigned epoch, void* counter, int ctas, void* stream);
int b200_alltoall_pull(const void* const* peer_ptrs, int rank, int world, long long send_off, long long block_bytes, void* out,
                       int ch_arrive, int ch_done, unsigned epoch, void* counter, int ctas, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_OPS_H */
