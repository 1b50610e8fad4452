/* mpit_b200 -- C ABI of the sm_100a kernel library (libmpit_kernels.so).
 *
 * Conventions
 *   - every entry point returns 0 on success or a cudaError_t value (negative values are
 *     library-level errors, see MPIT_E_*);
 *   - pointers are device pointers unless the name says host; `stream` is a cudaStream_t
 *     passed as void*; nothing here synchronises the stream;
 *   - "flat" tensors are in the reference's getParameters() order
 *     (/root/reference/asyncsgd/goot.lua:41), "packed" tensors are in the tensor-core
 *     operand layouts described in DESIGN.md; `*map` arrays translate flat index -> packed
 *     offset.
 *
 * The reference's native layer is a Lua binding of MPI (/root/reference/lua-mpi.h,
 * /root/reference/mpifuncs.c); it has no device code.  The host message layer that takes
 * its place is declared in mpit_msg.h; this header is the device side.
 */
#ifndef MPIT_B200_H_
#define MPIT_B200_H_

#ifdef __cplusplus
extern "C" {
#endif

#define MPIT_E_UNSUPPORTED (-2)
#define MPIT_E_TIMEOUT (-3)
#define MPIT_E_ARG (-4)
#define MPIT_E_SCRATCH (-5)   /* request exceeds the built-in scratch of a *_noalloc entry point */

/* Programmatic dependent launch for the training-step kernels (boundary, prep, igemm, wgrad, loss,
 * reduces): when on (default) each of them may begin its set-up while the previous kernel on the
 * stream is still running and blocks in `griddepcontrol.wait` before touching data.
 * enabled: 1 on, 0 off, <0 query only.  Returns the previous setting.  Takes effect for launches
 * (and graph captures) made after the call.  No reference counterpart. */
int mpit_set_pdl(int enabled);
/* Precision of the tensor-core layer kernels launched from now on (mpit_umma_igemm, mpit_umma_wgrad): 0 = TF32
 * products (default), 1 / 2 = 3xTF32: every operand tile is split in shared memory into the part the tensor core
 * keeps and the exact remainder and three MMAs are issued per k-step (hi*lo, lo*hi, hi*hi), which gives products
 * of fp32 accuracy on the TF32 pipe -- the reference computes in fp32 (Torch FloatTensor / CudaTensor SGEMM).
 * 1 assumes the tensor core truncates its fp32 input, 2 rounds the hi part to nearest explicitly.  Returns the
 * previous mode; a negative argument only queries. */
int mpit_set_tf32x3(int mode);

/* ---- optimizer rules (csrc/kernels/optim.cu) ------------------------------------------ */

/* Rule ids for mpit_rule_step (server rules of /root/reference/BiCNN/pserver.lua:123-198 and
 * the single-worker rules of /root/reference/BiCNN/optim-*-single.lua). */
enum { MPIT_RULE_ADD = 0, MPIT_RULE_RMSPROP = 1, MPIT_RULE_ADAM = 2, MPIT_RULE_ADAMAX = 3,
       MPIT_RULE_ADAGRAD = 4, MPIT_RULE_ADADELTA = 5 };

/* Nesterov SGD halves fused into one pass (/root/reference/asyncsgd/optim-msgd.lua:20-39,
 * /root/reference/asyncsgd/optim-eamsgd.lua:26-47,70):
 *   do_apply:     g' = g[gmap] + l2wd*x ; x -= clr*g' ; v -= clr*g' ; x -= d
 *   do_lookahead: v *= mom_next ; x += v
 *   wpk != NULL:  wpk[wmap[m*n+i]] = x[i] for m < nmaps, skipping -1 (refresh the tensor-core
 *                 operand copies: forward layout, data-gradient layout) */
int mpit_local_step(float* x, float* v, const float* g, const int* gmap, const float* d, float* wpk,
                    const int* wmap, int nmaps, long long n, float clr, float l2wd, float mom_next, int do_apply,
                    int use_mom, int do_lookahead, void* stream);

/* One fused pass of a server/single rule. hyper6 is a HOST array {a,b,c,d,eps,step}
 * (see csrc/kernels/optim_math.cuh). Optional gradient prologue: gscale, L1, L2, clamp
 * (/root/reference/BiCNN/bicnn.lua:398-409). */
int mpit_rule_step(int rule, float* p, const float* g, const int* gmap, float* s0, float* s1, float* s2,
                   float* wpk, const int* wmap, int nmaps, long long n, const float* hyper6, float l1,
                   float l2, float clip, float gscale, void* stream);

/* Downpour client accumulate (/root/reference/asyncsgd/optim-downpour.lua:26-45). */
int mpit_downpour_accum(float* x, float* acc, const float* g, const int* gmap, float* wpk, const int* wmap,
                        int nmaps, long long n, float neg_lr, float l2wd, int move_local, void* stream);

int mpit_grad_reg_clip(float* g, const float* p, long long n, float l1, float l2, float clip, void* stream);

/* dst[i] = src[map[i]] (scatter=0) or dst[map[i]] = src[i] (scatter=1). */
int mpit_map_copy(float* dst, const float* src, const int* map, long long n, int scatter, void* stream);

/* ---- fp32 CUDA-core reference layers (csrc/kernels/nn_ref.cu) -------------------------- */
/* geom12 is a HOST array {B,SH,SW,H,W,C,KH,KW,KWP,OH,OW,CO}.  `rng` is a DEVICE array
 * {seed, step} keying the Philox dropout stream; mpit_nll_fwd_bwd increments rng_inc[1]. */
int mpit_conv_fwd_ref(const int* geom12, const float* x, const float* w, const float* bias, const int* crop,
                      float* out, unsigned char* pool_idx, int relu, int pool, int dropout,
                      const unsigned long long* rng, void* stream);
int mpit_conv_dgrad_ref(const int* geom12, const float* dz, const float* w, float* dx, void* stream);
int mpit_conv_wgrad_ref(const int* geom12, const float* dz, const float* x, const int* crop, float* dw,
                        void* stream);
int mpit_act_bwd(const float* dout, const float* out, const unsigned char* pool_idx, float* dz, int B, int OH,
                 int OW, int C, int pool, float scale, void* stream);
int mpit_colsum(const float* dz, float* db, long long rows, int C, void* stream);
/* uint8 NCHW images -> pair-expanded NHWC8 floats in [0,1] (goot.lua:80-81 normalisation).  Optionally copies the
 * batch's labels [B] and crop offsets [B][2] (device pointers) into the engine's buffers in the same launch. */
int mpit_prep_u8(const unsigned char* src, float* dst, int B, int SH, int SW, int cin, const int* labels_src,
                 int* labels_dst, const int* crop_src, int* crop_dst, void* stream);
/* The same launch also forwards fwd_n (<= 256) floats fwd_src -> fwd_dst after the preceding work on the stream has
 * completed; fwd_dst may be host-mapped pinned memory (the end-to-end step returns the previous step's loss this
 * way instead of through a copy-engine operation on the compute stream). */
int mpit_prep_u8_fwd(const unsigned char* src, float* dst, int B, int SH, int SW, int cin, const int* labels_src,
                     int* labels_dst, const int* crop_src, int* crop_dst, const float* fwd_src, float* fwd_dst,
                     int fwd_n, void* stream);
/* loss_sum = (accumulate ? loss_sum : 0) + sum of per-sample NLL; dlogits (may be NULL) = gscale * dNLL.
 * dbias (optional, needs dlogits and B <= 4096): dbias[c] = sum over samples of dlogits[.][c], c < ld --
 * the bias gradient of the classifier layer, in a fixed summation order. */
int mpit_nll_fwd_bwd(const float* logits, const int* labels, float* dlogits, float* loss_sum, int* confusion,
                     int B, int ncls, int ld, float gscale, unsigned long long* rng_inc, int accumulate,
                     float* dbias, void* stream);

/* Fused classifier head (csrc/kernels/head.cu): Linear(K -> ncls) + LogSoftMax + ClassNLLCriterion forward and
 * their backward down to the pre-activation gradient of the layer below, in ONE launch
 * (/root/reference/asyncsgd/models/7-layers-cnn.lua:22-25, goot.lua:40,200-206):
 *   logits[B][ld] = x[B][K] w[ncls][K]^T + bias;  loss_sum (+)= sum NLL;  confusion[y][argmax]++;
 *   dlogits = gscale * dNLL;  db_cls[ld] = colsum(dlogits);
 *   dx[B][K] = (dlogits w) * (x > 0 ? in_scale : 0);  db_in[K] = colsum(dx).
 * dlogits == NULL: forward only.  Limits: ncls <= ld <= 32, K in {64,128,256,512}, else MPIT_E_UNSUPPORTED
 * (callers fall back to mpit_umma_igemm + mpit_nll_fwd_bwd).  rng_inc as for mpit_nll_fwd_bwd.
 * scratch: mpit_classifier_head_scratch_floats(K) floats, zero-filled ONCE by the caller; rows are spread over
 * several CTAs, each leaves partial sums there and the last one to finish adds them in a fixed order (and
 * re-arms the ticket), so results are bit-reproducible.  Calls sharing a scratch must not overlap. */
int mpit_classifier_head(const float* x, const float* w, const float* bias, const int* labels, float* logits,
                         float* dlogits, float* dx, float* loss_sum, int* confusion, float* db_cls, float* db_in,
                         unsigned long long* rng_inc, float* scratch, int B, int K, int ncls, int ld, float gscale,
                         float in_scale, int accumulate, void* stream);
int mpit_classifier_head_scratch_floats(int K);

/* ---- NVLink exchange kernels (csrc/kernels/comm.cu) -------------------------------------- */
/* `shard_ptrs` is a HOST array of nranks device addresses: entry s is rank s's shard of the
 * centre variable mapped into this process (symmetric heap).  The flat vector of n floats is
 * partitioned as in /root/reference/asyncsgd/pclient.lua:118-128 (floor(n/S) each, last takes
 * the remainder); shard buffers must be 16-byte aligned and padded to a multiple of 4 floats. */
int mpit_signal_words(void);
int mpit_max_ranks(void);

/* Get: dst[i] = centre[i]  (pclient.lua:74-84); optionally also refreshes the packed operand copy.  The index
 * tables cover the elements [map_base, n) (nmaps tables of n - map_base entries): parameters below map_base,
 * e.g. an embedding table that is read in flat order, have no operand copy. */
int mpit_param_pull(const unsigned long long* shard_ptrs, int nranks, int rank, float* dst, long long n,
                    float* wpk, const int* wmap, int nmaps, long long map_base, void* stream);
/* Put: centre[i] = src[i]  (seed, pclient.lua:62-72,131-133; single-mode push). */
int mpit_param_put(const unsigned long long* shard_ptrs, int nranks, int rank, const float* src, long long n,
                   void* stream);
/* Accumulate: centre[i] += scale * clip(reg(src[i]))  (pserver.lua:91), optionally zeroing src
 * (optim-downpour.lua:42). `params` is only read when l1/l2 != 0. */
int mpit_push_add(const unsigned long long* shard_ptrs, int nranks, int rank, float* src, const float* params,
                  long long n, float scale, float l1, float l2, float clip, int zero_src, void* stream);
/* d = alpha*(x - centre); centre += d  (optim-eamsgd.lua:58-65). */
int mpit_elastic_exchange(const unsigned long long* shard_ptrs, int nranks, int rank, const float* x, float* d,
                          long long n, float alpha, void* stream);
/* The whole EAMSGD step boundary in one pass: apply(t-1) [x -= clr g; v -= clr g; x -= d_in],
 * elastic(t) when do_sync [d = alpha (x - centre); centre += d], lookahead(t) [v *= mom; x += v],
 * and refresh of the packed operand copy (optim-eamsgd.lua:26-47,58-70).  la_in: the previous step was finished by
 * mpit_multi_wgrad_reduce_opt, i.e. x holds the parameters before the pending look-ahead and v is already scaled;
 * the kernel starts from x + v. */
int mpit_eamsgd_fused_step(const unsigned long long* shard_ptrs, int nranks, int rank, float* x, float* v,
                           const float* g, const int* gmap, float* d, float* wpk, const int* wmap,
                           int nmaps, long long n, float alpha, float clr, float l2wd, float mom_next, int do_apply,
                           int use_mom, int do_lookahead, int do_sync, int has_d_in, int la_in, void* stream);
/* SUM/AVG all-reduce over symmetric buffers (test/testreduceall.lua:28, asyncsgd/testreduceall.lua:72-77).
 * algo 0 = one-shot (result in the private `local_out`), 1 = two-shot (result in every rank's `out` buffer).
 * `in_mc`/`out_mc` are NVLS multicast addresses (0 = use peer loads/stores).  x_elastic != NULL fuses
 * x -= alpha*(x - mean) into the same kernel; the two-shot form then needs `flag_ptrs`, a symmetric uint32
 * buffer of mpit_allreduce_flag_words(capacity) words (per-chunk arrival flags).  The signal pad must not be
 * shared with a collective running concurrently on another stream.  All ranks call with identical n/algo. */
long long mpit_allreduce_flag_words(long long capacity_floats);
int mpit_allreduce(const unsigned long long* in_ptrs, unsigned long long in_mc,
                   const unsigned long long* out_ptrs, unsigned long long out_mc,
                   const unsigned long long* sig_ptrs, const unsigned long long* flag_ptrs, int nranks, int rank,
                   float* local_out, float* x_elastic, long long n, float scale, float alpha, int algo,
                   int max_blocks, void* stream);
/* The whole Downpour step boundary in one pass (optim-downpour.lua:24-54): g' = neg_lr (g + l2wd x); acc += g';
 * do_sync: centre += acc on the owners (accumulate that returns the previous value), x = that value + acc (push
 * followed by pull), acc = 0; otherwise x += g'.  Refreshes the packed operand copy. */
int mpit_downpour_fused_step(const unsigned long long* shard_ptrs, int nranks, int rank, float* x, float* acc,
                             const float* g, const int* gmap, float* wpk, const int* wmap, int nmaps, long long n,
                             float neg_lr, float l2wd, int do_sync, void* stream);
/* Gradient push of a row-sharded Linear weight W [N][K] (ops/pulled_linear.py): owner(n)'s rows += scale * g[n][:]
 * and the same values into its transposed copy [K][N/nranks] (read by the pulled data-gradient GEMM). */
int mpit_pulled_linear_push(const unsigned long long* wf_ptrs, const unsigned long long* wd_ptrs, int nranks, int rank,
                            const float* g, int N, int K, float scale, void* stream);
/* Boundary of a synchronous-EASGD step in ONE kernel: mean over the ranks of the symmetric input (each rank's
 * parameters after its local update), x -= alpha (x - mean), then the Nesterov look-ahead of the next step
 * (v *= mom; x += v when v != NULL) and the refresh of the packed operand copy (wpk / wmap as in mpit_local_step).
 * The all-reduce/average path of asyncsgd/testreduceall.lua:72-77 fused with the elastic update of
 * optim-eamsgd.lua:62-70. */
int mpit_sync_easgd_step(const unsigned long long* in_ptrs, unsigned long long in_mc,
                         const unsigned long long* out_ptrs, unsigned long long out_mc,
                         const unsigned long long* sig_ptrs, const unsigned long long* flag_ptrs, int nranks, int rank,
                         float* x, float* v, float* wpk, const int* wmap, int nmaps, long long n, float alpha, float mom,
                         int algo, int max_blocks, void* stream);
int mpit_bcast(const unsigned long long* buf_ptrs, unsigned long long buf_mc, const unsigned long long* sig_ptrs,
               int nranks, int rank, int root, long long n, int max_blocks, void* stream);
int mpit_device_barrier(const unsigned long long* sig_ptrs, int nranks, int rank, void* stream);
/* One-sided window operation on a peer's symmetric memory (MPI_Put / MPI_Get / MPI_Accumulate(SUM),
 * /root/reference/mpifuncs.c:1656,1131,9): op 0 remote[i] = scale*local[i], 1 local[i] = remote[i],
 * 2 remote[i] += scale*local[i] (red.add, fire and forget).  `remote` = peer base + displacement. */
int mpit_peer_rma(int op, float* remote, float* local, long long n, float scale, void* stream);
/* ---- asynchronous parameter server with server-side rules (csrc/kernels/ps_async.cu) ----------------
 * Device path of BiCNN/pserver.lua:115-205 + BiCNN/optim-{adam,adamax,adagrad,adadelta,rmsprop}.lua: every rank is
 * a client and the owner of one shard.  Symmetric buffers: `ctl` (mpit_ps_ctl_words() uint32, zeroed), `slots`
 * ([nranks][cap] floats on every owner), the centre shard `p`.  hyper10 (HOST doubles) =
 * {a, b, c, d, eps, lr, beta1, beta2, lrd, step_div} (a..eps as in mpit_rule_step's hyper6).
 * mpit_ps_push: slot_ptrs[s] = address of MY slot inside owner s.  Gradient element i comes from gflat[i], or from
 * gpk[gmap[i - map_base]] when gmap != NULL and i >= map_base; then L1/L2/clamp (skipped when *active == 0), gscale,
 * su-step accumulation into acc; do_push = 0 only accumulates.  reg_loss[2] += {l1*|x|_1, l2*|x|^2/2}.
 * mpit_ps_drain: poll + apply on `stream` (applies whatever has arrived, in arrival order; max_take > 0 limits
 * the number of gradients per drain).  Neither call waits for a peer except push's bounded flow-control wait. */
int mpit_ps_ctl_words(void);
int mpit_ps_counter(const void* ctl_local, int which, unsigned* out);
int mpit_ps_push(const unsigned long long* slot_ptrs, const unsigned long long* ctl_ptrs, int nranks, int rank,
                 const float* gflat, const float* gpk, const int* gmap, long long map_base, const float* x, float* acc,
                 const int* active, float* reg_loss, long long n, float l1, float l2, float clip, float gscale,
                 int do_push, void* stream);
int mpit_ps_drain(const unsigned long long* ctl_ptrs, int nranks, int rank, int rule, const double* hyper10, float* p,
                  float* s0, float* s1, float* s2, const float* slots, long long cap, long long size, int max_take,
                  void* stream);
/* Native server thread: keeps draining on a high-priority stream of `dev` every interval_us until stopped. */
int mpit_ps_server_start(int dev, const unsigned long long* ctl_ptrs, int nranks, int rank, int rule,
                         const double* hyper10, float* p, float* s0, float* s1, float* s2, const float* slots,
                         long long cap, long long size, int interval_us, void** out_handle);
int mpit_ps_server_stop(void* handle, long long* rounds_out);

/* ---- device transport of the message layer (csrc/kernels/mailbox.cu) --------------------------------------
 * Two-sided tagged messages between GPUs behind mpiT.Send/Recv/Isend/Irecv/Iprobe/Test/Wait/Cancel for CUDA tensors
 * (/root/reference/lua-mpi.h:70-78 hands the device pointer to MPI; /root/reference/init.lua:41-108 polls
 * Iprobe/Test).  Payloads move GPU to GPU (eager: into a slot of the receiver's symmetric memory; rendezvous: the
 * receiver pulls from the sender's tensor through CUDA IPC); 128-byte descriptors live in a POSIX shared-memory
 * segment registered with CUDA, ring layout [dst][src][K]: the sender's GPU publishes `seq` in stream order, all
 * probing / matching / cancelling / acknowledging is CPU atomics on the same words.  See mpit_b200/msg/device.py. */
int mpit_mbx_desc_bytes(void);
int mpit_shm_map(const char* name, long long bytes, int create, void** host_ptr, void** dev_ptr);
int mpit_shm_unmap(const char* name, void* host_ptr, long long bytes, int unlink_it);
int mpit_mbx_copy(void* dst, const void* src, long long nbytes, void* stream);
int mpit_mbx_publish(void* desc_dev, unsigned seq, void* stream);
int mpit_mbx_fill(void* desc_host, int tag, long long nbytes, int kind, unsigned long long rv_offset,
                  const void* ipc64);
int mpit_mbx_publish_host(void* desc_host, unsigned seq);
int mpit_mbx_word(const void* desc_host, int which);
int mpit_mbx_set_done(void* desc_host, unsigned seq);
int mpit_mbx_cas_state(void* desc_host, unsigned expect, unsigned val);
int mpit_mbx_scan(const void* ring, int nsrc, int K, const unsigned* base, const unsigned long long* mask,
                  int want_src, int want_tag, int start_src, int* out5, long long* out_bytes);
int mpit_mbx_desc_info(const void* desc_host, unsigned long long* rv_offset, void* ipc64);
int mpit_ipc_export_range(unsigned long long ptr, void* handle64, unsigned long long* offset);

/* ---- symmetric heap (csrc/symm/symm_heap.cu) --------------------------------------------- */
/* Takes the place of the reference's CUDA-aware-MPI pointer bridge (/root/reference/lua-mpi.h:70-78).
 * All sizes in bytes; pointers are returned as integers. VMM calls return 1000+CUresult on a
 * driver error. */
int mpit_symm_malloc(unsigned long long bytes, unsigned long long* out_ptr);
int mpit_symm_free(unsigned long long ptr);
int mpit_symm_ipc_export(unsigned long long ptr, void* handle64);
int mpit_symm_ipc_import(const void* handle64, unsigned long long* out_ptr);
int mpit_symm_ipc_close(unsigned long long ptr);
int mpit_enable_peer_access(int dev, int peer);
long long mpit_vmm_granularity(int dev, int nranks, int with_multicast);
int mpit_vmm_multicast_supported(int dev);
int mpit_vmm_alloc(int dev, unsigned long long bytes, unsigned long long* out_ptr, int* out_fd);
int mpit_fd_from_pid(int pid, int remote_fd);
int mpit_vmm_import(int dev, int fd, unsigned long long bytes, unsigned long long* out_ptr);
int mpit_mc_create(int nranks, unsigned long long bytes, unsigned long long* out_handle, int* out_fd);
int mpit_mc_import(int fd, unsigned long long* out_handle);
int mpit_mc_add_device(unsigned long long mc_handle, int dev);
int mpit_mc_bind(unsigned long long mc_handle, unsigned long long local_ptr, unsigned long long bytes);
int mpit_mc_map(unsigned long long mc_handle, int dev, unsigned long long bytes, unsigned long long* out_ptr);
int mpit_vmm_free(unsigned long long ptr);
int mpit_close_fd(int fd);

/* ---- tcgen05 / TMEM / TMA layers (csrc/kernels/umma_igemm.cu) ---------------------------------- */
/* Encode a rank-`rank` fp32 tiled tensor map (dims/box innermost-first, strides_bytes for dims
 * 1..rank-1). out128 is a HOST buffer of 128 bytes. */
int mpit_tmap_encode_f32(void* out128, const void* gptr, int rank, const unsigned long long* dims,
                         const unsigned long long* strides_bytes, const unsigned* box, int swizzle128);
/* Implicit-GEMM conv / linear, forward (sgn=+1) or data gradient (sgn=-1) with fused epilogue:
 * mode 0 bias(+ReLU)(+dropout) store, 1 bias+2x2 maxpool(+ReLU)+argmax, 2 ReLU-mask of a gradient,
 * 3 un-pool + ReLU-mask of a gradient (7-layers-cnn.lua:8-24 forward/backward).
 * geom17 (HOST) = {nsub,tw,th,tn,tiles_w,tiles_h,kw_it,cb,num_kb,sgn,B,OH,OW,N,mode,relu,dropout}.
 * colpart (optional, modes 2 and 3): [m_tiles][N] floats, row t = column sums of the values tile t wrote
 * (m_tiles = ceil(ceil(B/tn)*tiles_w*tiles_h / nsub)); summing the rows gives the bias gradient of the
 * layer whose pre-activation gradient this call produces. */
int mpit_umma_igemm(const void* tmap_a128, const void* tmap_b128, const int* geom17, float scale,
                    const float* bias, float* out, unsigned char* idx, const float* aux,
                    const unsigned char* aux_idx, const int* crop, const unsigned long long* rng, int* err,
                    float* colpart, int block_n, void* stream);
/* Resident-input convolution (forward, sgn=+1) or data gradient (sgn=-1): one CTA per image loads the image's
 * (zero-padded) input ONCE into shared memory; each kernel tap is the same plane at a shifted descriptor start row,
 * only the weights are streamed.  For layers whose input per image is a few tens of KB (conv2/conv3 of
 * /root/reference/asyncsgd/models/7-layers-cnn.lua); mpit_umma_igemm re-fetches the window for every tap.
 * geom15 (HOST) = {RW, RH, x_org, y_org, cb, KH, KW, sgn, ntiles, B, OH, OW, N, mode, relu}: plane of RH x RW pixels
 * whose element (0,0) is input pixel (y_org, x_org) (negative = zero padding); output pixels q = y*RW + x, y < OH,
 * in ntiles tiles of 128 rows; x >= OW is discarded.  tmap_in: 4-D NHWC map, box (32, RW, RH, 1), zero OOB fill;
 * tmap_w: [N][K] map, box (32, block_n), K ordered (kh, kw, channel chunk).  Modes 0, 2, 3 of mpit_umma_igemm;
 * colpart has B*ntiles rows.  MPIT_E_UNSUPPORTED when the planes do not fit in shared memory / TMEM. */
int mpit_umma_conv_resident(const void* tmap_in128, const void* tmap_w128, const int* geom15, float scale,
                            const float* bias, float* out, const float* aux, const unsigned char* aux_idx, int* err,
                            float* colpart, int block_n, void* stream);
/* Parameter pull fused with the GEMM that consumes the weights (/root/reference/asyncsgd/pclient.lua:74-84
 * followed by the model's first Linear): the weight rows stay in the NVLink-sharded centre and the
 * kernel's TMA loads read each tile from its owner's shard. tmaps_b: HOST array of nranks x 128 bytes. */
int mpit_umma_linear_pull(const void* tmap_a128, const void* tmaps_b, int nranks, int rows_per_owner,
                          const int* geom17, const float* bias, float* out, int* err, int block_n, void* stream);
/* The same fusion with reuse inside the CTA and a backward: y = act(x W^T + b) (by_k = 0) or dx = dy W (by_k = 1),
 * W row-sharded over the owners' symmetric memory.  One CTA owns mt (1, 2, 4) 128-row M-tiles, so each weight tile
 * crosses NVLink once per CTA and feeds mt accumulators.  by_k = 1 reads the owners' TRANSPOSED shards
 * ([out][rows_per_owner]); the reduction then runs across owners, chosen per k-block.  M rows of x / dy, reduction
 * length K, N output columns; tmap_a128: 4-D map [K,1,1,Mpad] box [32,1,1,128]; tmaps_b: HOST array nranks x 128 B.
 * Weight tiles travel in their own ring (up to 16 tiles in flight per CTA), ahead of the activation ring.
 * ksplit > 1 (by_k = 1 only, no bias / activation, K / 32 divisible by ksplit): split-K; `out` is zeroed on the
 * stream and the ksplit slices of an output tile add into it (red.add.v4.f32; summation order not fixed). */
int mpit_umma_pulled_gemm(const void* tmap_a128, const void* tmaps_b, int nranks, int rows_per_owner, int by_k, int M,
                          int K, int N, int relu, const float* bias, float* out, int* err, int block_n, int mt,
                          int ksplit, void* stream);

/* ---- tcgen05 weight gradients (csrc/kernels/umma_wgrad.cu) ------------------------------------- */
/* dW = Xcol^T dZ over all output pixels, pixel dimension split over `nsplit` CTAs that each write a
 * partial tile into the scratch workspace `part` (nsplit * N * m_pad floats, laid out
 * part[split][n][m] with m_pad = ceil(total_chunks/4)*128 the pitch of one n row); mpit_wgrad_reduce sums
 * the splits into the packed gradient gw[n*ldw + m]
 * (accGradParameters of SpatialConvolution/Linear, asyncsgd/goot.lua:206).
 * geom11 (HOST) = {tw,th,tn,tiles_w,tiles_h,kw_it,cb,total_chunks,num_kb,B,N}. */
int mpit_umma_wgrad(const void* tmap_x128, const void* tmap_dz128, const int* geom11, float* part,
                    const int* crop, int* err, int block_n, int nsplit, void* stream);
int mpit_wgrad_reduce(const float* part, float* gw, int nsplit, int m_pad, int K, int N, int ldw, void* stream);
/* Same result as mpit_umma_wgrad followed by mpit_wgrad_reduce (gw[n*ldw + m], n < N,
 * m < 32*total_chunks), for callers that cannot allocate device memory: a plain-C host or another
 * language binding with no device allocator of its own, code sharing the GPU with a framework
 * allocator that already owns the whole heap, or a latency-critical loop that must avoid the
 * device-wide synchronisation of cudaMalloc/cudaFree.  The split-K workspace is a fixed array inside
 * the library (MPIT_NOALLOC_SCRATCH_FLOATS floats = 8 MiB, resident on every GPU that loads the
 * library until the process exits); the call allocates and frees nothing.
 *   - Limit: nsplit * ceil(total_chunks/4)*128 * N <= MPIT_NOALLOC_SCRATCH_FLOATS, otherwise the call
 *     prints the required size to stderr, launches nothing and returns MPIT_E_SCRATCH.
 *   - The first call on each device looks up the array's device address and sets kernel attributes
 *     for that device; make that first call with the device current and before any stream capture
 *     is started on it.
 *   - All calls on a device share the one array, which keeps the previous call's contents. The
 *     kernels overwrite every word they later read, so no clearing is needed, but two calls whose
 *     GPU work overlaps (different streams, or different host threads) race on it: serialise them.
 * This entry point has no counterpart in the reference. */
#define MPIT_NOALLOC_SCRATCH_FLOATS (2u * 1024u * 1024u)
int mpit_umma_wgrad_noalloc(const void* tmap_x128, const void* tmap_dz128, const int* geom11, float* gw, int ldw,
                            const int* crop, int* err, int block_n, int nsplit, void* stream);
/* mpit_multi_wgrad_reduce: the split-K reduction of up to 16 segments in ONE launch (HOST arrays of
 * nseg entries): gw_i[n*K_i + m] = sum_s part_i[(s*N_i + n)*m_pad_i + m].  The bias-gradient partial rows
 * written by mpit_umma_igemm (colpart, T tiles x G*C columns) are the segment {nsplit=T*G, m_pad=1, K=1,
 * N=C}.  mpit_multi_colsum: bias gradients (column sums) of up to 8 [rows][C] matrices in
 * one launch; srcs/dsts/rows/cols are HOST arrays; C % 4 == 0 and 256 % (C/4) == 0. */
int mpit_multi_wgrad_reduce(int nseg, const unsigned long long* parts, const unsigned long long* gws,
                            const int* nsplit, const int* m_pad, const int* K, const int* N, void* stream);
/* The reduction as the LAST kernel of a training step that has no exchange: besides summing the partials it applies
 * every gradient element to its parameter (Nesterov / EAMSGD apply, asyncsgd/optim-msgd.lua:31-37,
 * optim-eamsgd.lua:37-44,70; d subtracted when has_d), scales v by the next momentum and writes the next step's
 * look-ahead point x + v into the operand copies -- the boundary kernel of the next step is then not launched.  x is
 * left at the parameters proper; la_in = the step started from that same state.  Output segments must lie inside the
 * packed gradient buffer gpk; pk2flat: packed index -> flat parameter index (-1 none); wmap2: flat index -> slot of
 * the data-gradient operand layout (-1 none). */
int mpit_multi_wgrad_reduce_opt(int nseg, const unsigned long long* parts, const unsigned long long* gws,
                                const int* nsplit, const int* m_pad, const int* K, const int* N, float* x, float* v,
                                const float* d, const int* pk2flat, float* wpk, const int* wmap2, const float* gpk,
                                float clr, float l2wd, float mom_next, int use_mom, int has_d, int la_in, void* stream);
int mpit_multi_colsum(int nseg, const unsigned long long* srcs, const unsigned long long* dsts,
                      const long long* rows, const int* cols, void* stream);

/* ---- BiCNN sentence-CNN kernels (csrc/kernels/bicnn.cu) ---------------------------------------- */
/* LookupTable forward / tied backward (/root/reference/BiCNN/bicnn.lua:30-40). dims multiples of 4. */
int mpit_embedding_gather(const float* E, const int* tokens, float* out, long long rows, int dim, int ldE, int ld,
                          void* stream);
int mpit_embedding_scatter_add(float* dE, const int* tokens, const float* g, const int* row_valid, long long rows,
                               int dim, int ldE, int ld, void* stream);
/* Finish of TemporalConvolution -> Max -> ReLU -> Normalize(2) (bicnn.lua:59,76-84; Normalize.lua:11-38):
 * unpack the (value, argmax) keys written by the max-over-time epilogue of mpit_umma_igemm (mode 4). */
int mpit_encoder_finish(const unsigned long long* mx, float* r, unsigned char* arg, float* y, float* normp, int S,
                        int F, int ld, float eps, void* stream);
/* GESD similarity of P index pairs (bicnn.lua:98-105); stats = {dot, l2} per pair. */
int mpit_gesd_pairs(const float* y, const int* qi, const int* ai, float* sim, float* stats, int P, int F, int ld,
                    void* stream);
/* MarginRankingCriterion (bicnn.lua:121) over triples (pairs 2t, 2t+1) fused with the GESD gradient. */
int mpit_gesd_margin_bwd(const float* y, const int* qi, const int* ai, const float* sim, const float* stats, float* dy,
                         float* loss_sum, int triples, int F, int ld, float margin, void* stream);
/* Backward of Normalize (closed form, cf. Normalize.lua:53-69) -> ReLU -> Max into the dense frame gradient
 * dC[S][T][ld]: every element is written (zeros off the arg-max frame), no clearing needed.  gmax (optional,
 * [S][ld]): the per-sentence gradient itself; its column sums are the temporal convolution's bias gradient. */
int mpit_encoder_bwd(const float* dy, const float* r, const unsigned char* arg, const float* normp, float* dC,
                     float* gmax, int S, int F, int ld, int T, void* stream);

/* Device-side mini-batch assembly for a BiCNN training step (bicnn.lua:295-359): sentence rows of the question,
 * the positive answer and K uniformly drawn wrong answers per sample (rejection against the sample's targets,
 * bicnn.lua:327-334; Philox keyed by seed, *step and the row).  ptrs18 is a HOST array of device pointers
 * {q_tok, q_len, a_tok, a_len, targets, ans_tok, ans_len, batch, step, tokens, lens, frames, row_valid, neg_idx,
 * 0, 0, 0, 0}; dims8 (HOST) = {tmax, A, Lk, kw, B, K, S, 0}.  batch[b] < 0 marks an empty slot. */
int mpit_bicnn_build_batch(const unsigned long long* ptrs18, const int* dims8, unsigned long long seed, void* stream);
/* First margin violator among the K candidates of every sample (bicnn.lua:351-358) from the similarities of the
 * B positive pairs followed by the B*K candidate pairs; writes the (q,a+),(q,a-) pair list of the backward pass
 * (inactive samples get the similarity pair (1,0)), *active = any, counters += {triples, steps}, and advances
 * *step. */
int mpit_bicnn_select(const float* sim, const float* stats, const int* batch, int B, int K, float margin, int* ai2,
                      float* sim2, float* stats2, int* active, float* counters, unsigned long long* step, void* stream);

/* ---- bandwidth-path exchange through the bulk-copy engine (csrc/kernels/comm_bulk.cu) ----------- */
/* Same semantics as mpit_param_pull / mpit_push_add / mpit_elastic_exchange, moved in 16 KiB
 * cp.async.bulk / cp.reduce.async.bulk transactions. Require n >= 4096*nranks, n % (4*nranks) == 0 and
 * 16-byte aligned buffers; return MPIT_E_UNSUPPORTED otherwise (callers then use the per-thread kernels). */
int mpit_bulk_supported(long long n, int nranks);
int mpit_param_pull_bulk(const unsigned long long* shard_ptrs, int nranks, int rank, float* dst, long long n,
                         void* stream);
int mpit_push_add_bulk(const unsigned long long* shard_ptrs, int nranks, int rank, float* src, long long n,
                       float scale, int zero_src, void* stream);
int mpit_elastic_exchange_bulk(const unsigned long long* shard_ptrs, int nranks, int rank, const float* x, float* d,
                               long long n, float alpha, void* stream);

/* ---- host side of an end-to-end training step (csrc/kernels/stepper.cu) ------------------------------
 * stage(): H2D of one mini-batch (pinned host pointers) on an internal copy stream into the caller's
 * double-buffered device staging, then mpit_prep_u8 on `main_stream`; hands back the summed loss of the step
 * issued `lag` steps earlier.  The caller then enqueues its optimizer step on `main_stream` and calls
 * finish(), which queues the 4-byte loss read-back into the caller's pinned `loss_host_pinned` ([lag+1][2]
 * floats) and returns once the staged inputs have left the host buffers.  Allocates no device memory.
 * Counterpart in the reference: the Lua step loop of /root/reference/asyncsgd/goot.lua:172-225. */
int mpit_stepper_create(void** handle, int lag, void* d_img0, void* d_img1, int* d_lab0, int* d_lab1, int* d_crop0,
                        int* d_crop1, float* loss_host_pinned, long long img_bytes, int B, int SH, int SW, int cin);
int mpit_stepper_stage(void* handle, const void* img, const int* labels, const int* crop, float* x0_dst,
                       int* labels_dst, int* crop_dst, void* main_stream, float* prev_loss_sum, int* have_prev);
int mpit_stepper_finish(void* handle, const float* loss_dev, void* main_stream);
int mpit_stepper_loss(void* handle, long long index, float* loss_sum);
int mpit_stepper_counters(void* handle, long long* issued, long long* returned, int mark_returned);
int mpit_stepper_destroy(void* handle);

/* ---- derived-datatype pack / unpack on the device (csrc/kernels/dtpack.cu) ------------------------
 * Device-side twin of mpit_dt_pack / mpit_dt_unpack (mpit_msg.h): moves `nbytes` between the contiguous
 * `stream_buf` and replicas, `extent` bytes apart, of a type made of `nblk` blocks.  d_disp[i] = byte
 * displacement of block i, d_pre[i] = stream offset of block i inside one replica (d_pre[nblk] = size);
 * both are DEVICE arrays.  `unit` (1,2,4,8,16) must divide every displacement, length, the extent, nbytes
 * and both base addresses.  to_stream: 1 gather (pack), 0 scatter (unpack).  The reference leaves this
 * to its CUDA-aware MPI library (/root/reference/lua-mpi.h:70-78). */
int mpit_dt_move_device(void* strided_base, void* stream_buf, long long nbytes, long long size, long long extent,
                        int nblk, const long long* d_disp, const long long* d_pre, int unit, int to_stream,
                        void* stream);
/* The same for a REGULAR block list (size/blen blocks of `blen` bytes, block i at d0 + i*bstride: vectors, 2-d
 * sub-arrays): no tables, four independent words in flight per thread. */
int mpit_dt_move_device_regular(void* strided_base, void* stream_buf, long long nbytes, long long size,
                                long long extent, long long d0, long long blen, long long bstride, int unit,
                                int to_stream, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MPIT_B200_H_ */
