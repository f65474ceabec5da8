/*
 * b200zero — C ABI of the sm_100a kernel library (libb200zero.so).
 *
 * Every entry point takes raw device pointers plus a CUDA stream handle (as
 * void*; NULL = legacy default stream), launches asynchronously on that stream
 * and returns 0 on success. On failure it returns non-zero and
 * bz_last_error() describes why (thread-local string). Unless stated otherwise
 * an entry point allocates nothing and synchronises nothing.
 *
 * Python binds these with ctypes (ops/lib.py); C/C++ callers can link the .so
 * directly. Citations are to the reference repo the framework replaces
 * (SURVEY.md); the reference itself contains no native code.
 */
#ifndef B200ZERO_H_
#define B200ZERO_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime ---------------------------------------------------------- */
const char* bz_last_error(void);
int bz_version(void);
int bz_sm_count(void);

/* ---- K4: bf16 GEMM on tcgen05/TMEM/TMA --------------------------------
 * C[M,N] (+)= A*B with fp32 accumulation.
 *   a_mn_major=0: A stored [M,K] row-major (lda = row stride in elements)
 *   a_mn_major=1: A stored [K,M] row-major
 *   b_mn_major=0: B stored [N,K] row-major (nn.Linear weight layout)
 *   b_mn_major=1: B stored [K,N] row-major
 *   out_mode: 0 bf16 store, 1 bf16 accumulate (C += AB), 2 fp32 store, 3 fp32 accumulate
 *   cta_group: 1 = one CTA per tile (128x256), 2 = CTA pair (256x256, cta_group::2), 0 = auto
 *   max_ctas:  0 = all SMs; otherwise cap the persistent grid (leave SMs for a
 *              concurrent communication kernel)
 * Requirements: pointers 16-byte aligned; lda, ldb multiples of 8.
 * Replaces the cuBLAS calls a DeepSpeed user script would make.
 */
int bz_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                 int a_mn_major, int b_mn_major, int out_mode, int cta_group, int max_ctas, void* stream);

/* ---- K5: RMSNorm (optionally fused with the preceding residual add) ------
 * fwd: xo = x + delta (only when delta != NULL; written to x_out), y = xo*rstd*w.
 *      rstd (fp32 [T]) may be NULL. bf16 tensors, D % 8 == 0, D <= 8192.
 * bwd: dx = dres + d(rmsnorm)/dx ; dw (+)= sum over rows. `workspace` must hold
 *      bz_rmsnorm_bwd_workspace_floats(T, D) floats when dw != NULL. rstd may be
 *      NULL (recomputed). Replaces torch-eager RMSNorm. */
int bz_rmsnorm_fwd(const void* x, const void* delta, const void* w, void* x_out, void* y, float* rstd, int T,
                   int D, float eps, void* stream);
int bz_rmsnorm_bwd_workspace_floats(int T, int D);
int bz_rmsnorm_bwd(const void* x, const void* w, const float* rstd, const void* dy, const void* dres, void* dx,
                   void* dw, int dw_is_f32, int dw_accumulate, float* workspace, int T, int D, float eps,
                   void* stream);

/* RMSNorm backward for callers that may not allocate (no counterpart in the reference, which has
 * no kernels). Same maths, launch geometry and results as bz_rmsnorm_bwd, but the dw partial-sum
 * scratch is a device array built into the library (BZ_RMSNORM_BWD_STATIC_SCRATCH_FLOATS floats,
 * 5 MiB per GPU for the life of the process) instead of a caller workspace. Meant for step loops
 * that must not hit cudaMalloc's device-wide synchronisation, jobs with no HBM slack left, and
 * C/C++ hosts without an allocator to lend a workspace. Refuses problems needing more than the
 * built-in scratch (min(T, 2*SMs)*D floats). The first call on each device resolves the array's
 * address — make it during set-up, before any stream capture. All calls on a device share the
 * array: calls that can overlap on the GPU race; serialise them on one stream or with events.
 * Never calls cudaMalloc/cudaFree. */
#define BZ_RMSNORM_BWD_STATIC_SCRATCH_FLOATS 1310720
int bz_rmsnorm_bwd_static_scratch(const void* x, const void* w, const float* rstd, const void* dy, const void* dres,
                                  void* dx, void* dw, int dw_is_f32, int dw_accumulate, int T, int D, float eps,
                                  void* stream);

/* ---- K7: RoPE, in place on the first n_rot_heads heads of each row --------
 * cos/sin: fp32 [S, head_dim/2]; position = row % S; inverse=1 applies the
 * transpose rotation (backward). */
int bz_rope(void* buf, const float* cos_t, const float* sin_t, int T, int S, int n_rot_heads, int head_dim, int ld,
            int inverse, void* stream);

/* ---- K8: SwiGLU on a fused [T, 2F] gate|up buffer ---- */
int bz_swiglu_fwd(const void* gate_up, void* act, long long T, int F, void* stream);
int bz_swiglu_bwd(const void* gate_up, const void* dact, void* dgate_up, long long T, int F, void* stream);

/* ---- K10: embedding gather / scatter-add ------------ */
int bz_embedding_fwd(const void* ids, int ids_are_i64, const void* table, void* out, long long T, int D, int V,
                     void* stream);
int bz_embedding_bwd(const void* ids, int ids_are_i64, const void* dout, void* dtable, int dtable_is_f32,
                     long long T, int D, int V, void* stream);

/* ---- K9 (part): cross-entropy fwd+bwd on a logits chunk -------------------
 * loss[t] = lse(z_t) - z_t[y_t]; when write_grad, logits are overwritten by
 * (softmax - onehot) * grad_scale. targets < 0 are ignored. V % 8 == 0.
 *  */
int bz_cross_entropy(void* logits, const void* targets, int targets_are_i64, float* loss, float* lse, int T, int V,
                     long long ld, float grad_scale, int write_grad, void* stream);

int bz_add_bf16(const void* a, const void* b, void* out, long long n, void* stream);

/* ---- K11 / F9 / F10: optimizer-side kernels --------------------------------
 * bz_adamw: torch.optim.AdamW semantics on flat fp32 master/m/v; grad bf16 or
 *   fp32, multiplied by grad_scale * (*grad_scale_dev if non-NULL); writes the
 *   bf16 copy of the updated weights when param_bf16 != NULL; does nothing when
 *   *skip_flag != 0 (non-finite step). `step` is 1-based.
 * bz_sumsq: out (+)= sum(g^2); sets *nonfinite_flag when the sum is not finite.
 *   `workspace` must hold bz_sumsq_workspace_floats() floats.
 * bz_clip_coef: coef = min(1, max_norm/(sqrt(sumsq)+1e-6)) * extra_scale on device.
 * (DeepSpeed FusedAdam / grad-norm path: SURVEY.md §2.3 F7, F9, F10) */
int bz_adamw(float* master, float* m, float* v, const void* grad, int grad_is_f32, void* param_bf16, long long n,
             float lr, float beta1, float beta2, float eps, float weight_decay, int step, float grad_scale,
             const float* grad_scale_dev, const int* skip_flag, void* stream);
int bz_sumsq_workspace_floats(void);
int bz_sumsq(const void* g, int g_is_f32, long long n, float* workspace, float* out, int accumulate,
             int* nonfinite_flag, void* stream);
int bz_clip_coef(const float* sumsq, float max_norm, float extra_scale, float* coef_out, float* norm_out,
                 int* nonfinite_flag, void* stream);
int bz_cast_f32_to_bf16(const float* in, void* out, long long n, void* stream);
int bz_cast_bf16_to_f32(const void* in, float* out, long long n, void* stream);

/* ---- H2: peer-memory runtime (CUDA IPC; one process per GPU) --------------------
 * bz_symm_alloc: zero-filled device allocation suitable for IPC export (cudaMalloc —
 * set-up time only). Handles are 64 opaque bytes exchanged out of band
 * (torch.distributed all_gather in parallel/symm.py). New scope: the reference has no
 * communication backend at all. */
int bz_comm_max_world(void);
int bz_comm_signal_pad_bytes(void);
int bz_comm_segment_bytes(void);
int bz_fused_comm_workspace_bytes(void);
int bz_symm_alloc(unsigned long long bytes, void** out);
int bz_symm_free(void* p);
int bz_ipc_export(void* p, unsigned char* handle64);
int bz_ipc_import(const unsigned char* handle64, void** out);
int bz_ipc_close(void* p);
int bz_enable_peer_access(int peer_device);

/* ---- H2, second backend: CUDA virtual-memory management + NVSwitch multicast objects (csrc/vmm.cu) ----
 * One exportable physical allocation per rank (cuMemCreate, POSIX file descriptor), mapped by every rank of the node;
 * optionally bound into one multicast object whose mapping is what multimem.ld_reduce / multimem.st address.
 * Set-up order, all ranks: bz_vmm_round_size -> bz_vmm_alloc -> exchange fds (unix sockets, parallel/symm.py) ->
 * bz_vmm_import per peer. Multicast: rank 0 bz_mc_create, others bz_mc_import; everyone bz_mc_add_device; barrier;
 * everyone bz_mc_bind_and_map. Driver entry points are resolved at run time (no libcuda link dependency).
 * New scope: the reference has no communication layer. */
int bz_vmm_supported(int* vmm, int* multicast);
int bz_vmm_round_size(unsigned long long bytes, int world, unsigned long long* out);
int bz_vmm_alloc(unsigned long long size, void** ptr, unsigned long long* handle, int* fd);
int bz_vmm_import(int fd, unsigned long long size, void** ptr, unsigned long long* handle);
int bz_vmm_unmap(void* ptr, unsigned long long size, unsigned long long handle);
int bz_mc_create(int world, unsigned long long size, unsigned long long* mc_handle, int* fd);
int bz_mc_import(int fd, unsigned long long* mc_handle);
int bz_mc_add_device(unsigned long long mc_handle);
int bz_mc_bind_and_map(unsigned long long mc_handle, unsigned long long mem_handle, unsigned long long size,
                       void** mc_ptr);
int bz_mc_unmap(void* mc_ptr, unsigned long long size, unsigned long long mc_handle);

/* ---- K1: fused reduce-scatter + scale + accumulate + ‖g‖² + isfinite ---------------
 * Segment record (bz_comm_segment_bytes() = 32 bytes): {int64 full_off, int64 numel,
 * int64 shard_off, int32 partitioned, int32 pad}; element offsets, multiples of 16.
 * use_barrier: 0 = none (single process / emulated ranks), 1 = every CTA waits for all ranks' buffers (phase 0) and the
 * last CTA for all ranks' reads (phase 1), 3 = phase 0 runs as a one-warp launch in front of the reduce kernel.
 * grad_ptrs[ws] / signal_ptrs[ws]: peer-mapped pointers, index = rank. workspace:
 * bz_fused_comm_workspace_bytes() bytes, zeroed once. epoch must increase by one per
 * call per channel on every rank. use_barrier=0 skips the cross-rank flags (single-GPU
 * emulation of several ranks: never run mutually-waiting kernels on one device).
 * Replaces ncclReduceScatter + cast + norm + isfinite (SURVEY.md §2.3 X1,X4,X5,X6,F9). */
int bz_fused_reduce_scatter(void* const* grad_ptrs, void* const* signal_ptrs, const void* segs, int nseg,
                            void* shard_grads, int shard_is_f32, float scale, float* workspace, float* sumsq_part,
                            float* sumsq_repl, int* nonfinite_flag, unsigned int epoch, int channel, int rank, int ws,
                            int use_barrier, int ctas, void* stream);

/* ---- K3: AdamW on the shard + push all-gather of the updated bf16 rows -------------
 * push=1: rows are stored into every rank's gathered parameter buffer (param_ptrs);
 * shard_params (optional) receives the local bf16 copy. Predicated on *skip_flag.
 * Replaces FusedAdam + cast + ncclAllGather (SURVEY.md §2.3 F7,F10,X3). */
int bz_fused_adam_push(void* const* param_ptrs, void* const* signal_ptrs, const void* segs, int nseg, float* master,
                       float* m, float* v, void* shard_grads, int grads_are_f32, void* shard_params, float lr,
                       float beta1, float beta2, float eps, float weight_decay, int step,
                       const float* grad_scale_dev, const int* skip_flag, float* workspace, unsigned int epoch,
                       int channel, int rank, int ws, int push, int use_barrier, int zero_grads, int ctas,
                       void* stream);

/* ---- X5/X6/X8 without NCCL (opt-in): sum four device floats over all ranks through the signal pads
 * (16-byte peer stores + epoch flag, rank-ordered sum: bit-identical on every rank). One call per step on every
 * rank with the same growing epoch and a channel of its own. */
int bz_peer_allreduce4(void* const* signal_ptrs, float* vals4, unsigned int epoch, int channel, int rank, int ws,
                       void* stream);

/* ---- K1, NVLS flavour (opt-in): the gradient buffer is additionally mapped through an NVSwitch
 * multicast object (grad_mc = multicast VA of the buffer whose per-rank VAs are in grad_ptrs).
 * Partitioned tensors are summed inside the switch (multimem.ld_reduce); replicated tensors,
 * barriers and outputs are exactly as in bz_fused_reduce_scatter. */
int bz_fused_reduce_scatter_mc(void* const* grad_ptrs, void* const* signal_ptrs, void* grad_mc, const void* segs,
                               int nseg, void* shard_grads, int shard_is_f32, float scale, float* workspace,
                               float* sumsq_part, float* sumsq_repl, int* nonfinite_flag, unsigned int epoch,
                               int channel, int rank, int ws, int use_barrier, int ctas, void* stream);

/* ---- pull all-gather of a unit's rows from every peer's shard (ZeRO-3, X2) ---------- */
int bz_pull_allgather(void* const* shard_param_ptrs, const void* segs, int nseg, void* dst_unit, int rank, int ws,
                      int ctas, void* stream);
/* v2 (opt-in): walks the vector index and unrolls over the peers — no integer division per vector. */
int bz_pull_allgather_v2(void* const* shard_param_ptrs, const void* segs, int nseg, void* dst_unit, int rank, int ws,
                         int ctas, void* stream);

/* ---- K6: causal flash attention forward (head_dim 128, GQA) on tcgen05 -----------
 * qkv: fused [B*S, (Hq+2*Hkv)*128] bf16 with RoPE applied; out: [B*S, Hq*128] bf16;
 * lse (optional): [Hq, B*S] fp32 natural-log-sum-exp of the scaled scores.
 * S % 128 == 0. */
int bz_flash_attn_fwd(const void* qkv, void* out, float* lse, int B, int S, int Hq, int Hk, int head_dim,
                      float softmax_scale, void* stream);

/* ---- K2: all-gather -> GEMM in one kernel ------------------------------------------
 * Y[M,N] = X[M,K] · W[N,K]^T with W row-partitioned over `ws` ranks. shard_ptrs[p] is
 * rank p's [N/ws, K] slice (peer-mapped; entry `rank` local). Communication warps in
 * every CTA pull W chunk by chunk over NVLink and write it through to Wfull [N,K];
 * GEMM tiles consume chunks as they arrive. chunk_ready: >= ceil(N/256) uint32 scratch.
 * Replaces ncclAllGather + cuBLAS GEMM (SURVEY.md §2.3 X2 + F1). */
int bz_ag_gemm_max_tiles(void);
int bz_ag_gemm_bf16(const void* X, void* const* shard_ptrs, void* Wfull, void* Y, void* chunk_ready, int M, int N,
                    int K, int ldx, int ldc, int rank, int ws, int cta_group, void* stream);

/* ---- K2, dgrad flavour: dX[M,K] = dY[M,N] · W[N,K] with W row-partitioned over `ws` ranks (the gathered
 * dimension is the reduction dimension). Same peer/chunk protocol as bz_ag_gemm_bf16; the k-loop consumes the
 * 256-row chunks in arrival order, W is read MN-major (no transposed copy). chunk_ready: >= ceil(N/256) uint32. */
int bz_ag_gemm_dgrad_bf16(const void* dY, void* const* shard_ptrs, void* Wfull, void* dX, void* chunk_ready, int M,
                          int N, int K, int ldy, int ldc, int rank, int ws, int cta_group, void* stream);

/* ---- K6: flash attention backward ---------------------------------------------------
 * dqkv (same fused layout as qkv) from qkv, out, lse (from the forward) and dout.
 * workspace: bz_flash_attn_bwd_workspace_floats(B,S,Hq) floats (fp32 dQ accumulator +
 * Delta); zeroed by the call. Five tcgen05 GEMMs per (kv block, q tile); dK/dV accumulate
 * in TMEM; dQ via cp.reduce.async.bulk add.f32. */
long long bz_flash_attn_bwd_workspace_floats(int B, int S, int Hq);
int bz_flash_attn_bwd(const void* qkv, const void* out, const float* lse, const void* dout, void* dqkv,
                      float* workspace, int B, int S, int Hq, int Hk, int head_dim, float softmax_scale,
                      void* stream);

/* ---- K6 forward, current kernel (bz_flash_attn_fwd forwards to it): P double-buffered so exp/pack of block j+1 overlaps O += P_j V_j; same
 * contract as bz_flash_attn_fwd. error_flag (device int, may be NULL) becomes 1 if the kernel declined to run
 * because its dynamic shared-memory segment was not 1024-aligned (it needs 226.25 KB of the 227 KB).
 * poly_stride in {0, 2, 4, 8}: every poly_stride-th exponential is evaluated on the FMA pipe (degree-3
 * polynomial, rel. error 7.5e-5) instead of the SFU; 0 = all on the SFU. */
int bz_flash_attn_fwd_v3(const void* qkv, void* out, float* lse, int* error_flag, int B, int S, int Hq, int Hk,
                         int head_dim, float softmax_scale, int poly_stride, void* stream);

/* ---- K5, v2 (opt-in): RMSNorm forward / backward with the next row's loads issued before the current row's
 * reductions (twice the bytes in flight per SM); same contracts, workspace and bit-identical results. */
int bz_rmsnorm_fwd_v2(const void* x, const void* delta, const void* w, void* x_out, void* y, float* rstd, int T, int D,
                      float eps, void* stream);
int bz_rmsnorm_bwd_v2(const void* x, const void* w, const float* rstd, const void* dy, const void* dres, void* dx,
                      void* dw, int dw_is_f32, int dw_accumulate, float* workspace, int T, int D, float eps,
                      void* stream);

/* ---- K8, v2 (opt-in): SwiGLU forward / backward indexed by (row, vector) — no 64-bit integer division per vector
 * — with the correctly rounded reciprocal (bit-identical to v1) or, fast != 0, rcp.approx (2 ulp before bf16 rounding). */
int bz_swiglu_fwd_v2(const void* gate_up, void* act, long long T, int F, int fast, void* stream);
int bz_swiglu_bwd_v2(const void* gate_up, const void* dact, void* dgate_up, long long T, int F, int fast, void* stream);

/* ---- GPT-2 family element-wise ops (csrc/gpt2_ops.cu) ----------
 * LayerNorm: y = (x - mean) * rstd * w + b over rows of width D (D % 8 == 0, D <= 16384); mean / rstd are
 * fp32 [T] saved for backward. Backward also produces dw = sum_rows dy * xhat and db = sum_rows dy
 * (bf16 or fp32, optionally accumulated) through a per-CTA partial workspace of
 * bz_gpt2_bwd_workspace_floats(T, D) floats. bias + GELU(tanh): y = gelu(x + bias); backward returns dx and
 * dbias = column sums of the bf16-rounded dx. */
int bz_layernorm_fwd(const void* x, const void* w, const void* b, void* y, float* mean, float* rstd, int T, int D,
                     float eps, void* stream);
long long bz_gpt2_bwd_workspace_floats(int T, int D);
int bz_layernorm_bwd(const void* x, const void* w, const float* mean, const float* rstd, const void* dy, void* dx,
                     void* dw, void* db, int dwb_is_f32, int accumulate, float* workspace, int T, int D, void* stream);
int bz_bias_gelu_fwd(const void* x, const void* bias, void* y, long long T, int F, void* stream);
int bz_bias_gelu_bwd(const void* x, const void* bias, const void* dy, void* dx, void* dbias, int db_is_f32,
                     int accumulate, float* workspace, int T, int F, void* stream);

/* ---- K6 backward, current kernel (bz_flash_attn_bwd forwards to it): S^T of tile i+1 issued behind dV of tile i, dP^T of tile i+1 behind dS(i),
 * dQ^T of tile i-1 staged after dS(i); same contract and workspace as bz_flash_attn_bwd. */
int bz_flash_attn_bwd_v3(const void* qkv, const void* out, const float* lse, const void* dout, void* dqkv,
                         float* workspace, int* error_flag, int B, int S, int Hq, int Hk, int head_dim,
                         float softmax_scale, void* stream);

/* ==== libb200zero_host.so (plain C++, no CUDA) ========================================= */

/* ---- H1: AdamW on the host for ZeRO-Offload (AVX-512/AVX2 multiversioned, thread pool).
 * In place over [0, n); step is 1-based; out (bf16 bits) may be NULL; nthreads <= 0 -> cores/2.
 *  */
int bz_host_adamw(float* master, float* m, float* v, const void* grad, int grad_is_bf16, unsigned short* out,
                  long long n, float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                  float grad_scale, int nthreads);
int bz_host_threads(void);

/* ---- H1b: asynchronous file I/O for the NVMe tier of the optimizer offload ------------
 * Requests are split into block_bytes slices serviced by a pool of pread/pwrite workers; each
 * request returns a ticket; bz_aio_wait returns 0, an errno, or -1 for a short read. */
void* bz_aio_create(int nthreads, long long block_bytes);
void bz_aio_destroy(void* h);
int bz_aio_open(const char* path, int create, long long size_bytes, int direct, int* direct_used);
int bz_aio_close(int fd);
int bz_aio_fsync(int fd);
long long bz_aio_read(void* h, int fd, void* buf, long long nbytes, long long offset);
long long bz_aio_write(void* h, int fd, const void* buf, long long nbytes, long long offset);
int bz_aio_wait(void* h, long long ticket);
int bz_aio_wait_all(void* h);
long long bz_aio_bytes_read(void* h);
long long bz_aio_bytes_written(void* h);

#ifdef __cplusplus
}
#endif
#endif /* B200ZERO_H_ */
