/*
 * dctr.h -- C-ABI of the B200-native deep-ctr hot path (libdctr.so, sm_100a).
 *
 * The reference (Atomu2014/deep-ctr) is pure Python on top of Theano / TensorFlow 0.x and has NO
 * FFI of its own; every entry point below therefore names the reference
 * ENGINE CALL (file:line under /root/reference) it replaces.  INTEGRATION.md shows the ctypes stub
 * a maintainer would add on the reference side.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer borrowed for the call unless the name ends in `_host`;
 *     nothing is allocated, freed or synchronised inside (stream-ordered, CUDA-graph capturable);
 *   - `stream` is a cudaStream_t passed as void*;
 *   - return 0 on success, a negative DCTR_E* code otherwise; the message is kept per thread and
 *     read back with dctr_last_error(); no C++ exception crosses the boundary;
 *   - embedding table: float32 [X, ld], ld % 4 == 0, row = [w | v_1..v_r | zero pad], d = r + 1;
 *   - ids: int32 GLOBAL row ids [N, F] (row-major, sample x field); wts: float32 [N, F] or NULL (=1);
 *   - all reductions run in a fixed order: two runs on the same inputs are bit-identical.
 */
#ifndef DCTR_H_
#define DCTR_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DCTR_VERSION 2

enum { DCTR_OK = 0, DCTR_EINVAL = -1, DCTR_EWORKSPACE = -2, DCTR_ECUDA = -3 };
enum { DCTR_ACT_NONE = 0, DCTR_ACT_TANH = 1, DCTR_ACT_RELU = 2, DCTR_ACT_SIGMOID = 3 };
enum { DCTR_CROSS_NONE = 0, DCTR_CROSS_IP = 1, DCTR_CROSS_OP = 2 };
/* duplicate-row semantics of the sparse update */
enum { DCTR_SCATTER_SUM = 0,      /* TF: row -= lr * sum of its gradients (FNN.py:46-49 densified grad) */
       DCTR_SCATTER_SEQ_DECAY = 1 /* Theano: per occurrence, in sample order, row = row*c - lr*g (FNN_wnzh.py:300-306) */ };
/* GEMM operand layouts / epilogues */
enum { DCTR_GEMM_NN = 0, DCTR_GEMM_NT = 1, DCTR_GEMM_TN = 2 };
enum { DCTR_EPI_NONE = 0, DCTR_EPI_FWD = 1, DCTR_EPI_DGRAD = 2 };

int dctr_version(void);
const char* dctr_last_error(void);
/* number of kernels this library has launched from the calling process (bench.py's gpu_launches) */
long long dctr_launch_count(void);

/* How the calling thread's next launches are issued (captured CUDA graphs keep the mode they were captured with):
 *   0  plain stream-ordered launches (default)
 *   1  programmatic dependent launch for every kernel: the next kernel's CTAs are scheduled while the current one still
 *      runs and block in griddepcontrol.wait until it has completed -- hides the launch gap of short kernels
 *   2  as 1, except the tensor-core GEMM (a parked CTA of it holds a whole SM)
 * Results are identical in every mode.  The DCTR_PDL environment variable, when set to 0 / 1 / 2, overrides the call. */
int dctr_set_launch_mode(int mode);

/* Row ids come from data files.  Every kernel that takes ids compares them with X: an id outside [0, X) is READ as row 0 by
 * the forward kernels, sorted behind the table by dctr_sort_*, never written by the folds, and counted in the two device
 * words registered here for the calling thread: status[0] += 1 per offending id, status[1] = one offending value (captured
 * CUDA graphs keep the pointer they were captured with; NULL = do not report).  TensorFlow raises InvalidArgument and the
 * Theano script KeyError for such ids (FM.py:57 embedding_lookup_sparse, FNN_wnzh.py:95); the host side of the build
 * raises after the step when status[0] != 0. */
int dctr_set_id_status(int32_t* status);

/* ---- K1: per-field embedding gather + concat (+ fused inner/outer product layer) -------------
 * replaces feats_to_layer_one_array (FNN_wnzh.py:87-96, data_fm.py:46-54), the 13 rank-1 matmuls +
 * 26 embedding_lookup_sparse + concat of FNN.forward (FNN.py:76-85), the N*P scalar matmuls of
 * FNN_IP_L3.forward (FNN_IP_L3.py:91-97) and the outer-product sum of FNN_OP_L3.forward
 * (FNN_OP_L3.py:93-100).
 *   e[n,f,:] = wts[n,f] * table[ids[n,f], 0:d]
 *   bias_first=1 (Theano): z[n,:] = [ bias | e[n,0,:] .. e[n,F-1,:] ]
 *   bias_first=0 (TF):     z[n,:] = [ e[n,0,:] .. e[n,F-1,:] | cross | bias ]
 *   cross IP: p[pi(i,j)] = <e_i,e_j>, i<j, pi = i*F - i(i+1)/2 + j-i-1;  OP: q[a*d+b] = sum_{i<j} e_i[a] e_j[b]
 *   out[n,c] = act0(z[n,c]) for c < D0; columns D0..ldz-1 are written as 0.
 * bias: device pointer to ONE float (w_0 / b). */
int dctr_gather_concat_fwd(const int32_t* ids, const float* wts, const float* table, int64_t X, int ld,
                           int N, int F, int d, int cross, int bias_first, const float* bias, int act0,
                           float* out, int ldz, void* stream);

/* ---- K0: feature list -> one slot per field (the slot logic of feats_to_layer_one_array, FNN_wnzh.py:91-95) ----
 * feats: int32 [N, J] feature ids, negative = absent; feat_field: int32 [X] field of every feature (fm.model.txt,
 * FNN_wnzh.py:82-84).  ids[n, field(feat)] = feat, wts = 1; fields a sample does not have get (pad_row, 0); a later
 * feature of the same field overwrites an earlier one, as in the reference loop. */
int dctr_arrange_fields(const int32_t* feats, const int32_t* feat_field, int N, int J, int F, int32_t pad_row,
                        int32_t* ids, float* wts, void* stream);
/* The same, plus the sample's UPDATE list: the reference's scatter loops over EVERY feature of the sample (`for feat in ft`,
 * FNN_wnzh.py:300-306), so a feature whose slot was overwritten by a later feature of the same field, or that occurs twice,
 * is still updated -- from the gradient of its field's slot.  upd_rows / upd_pos: int32 [N, J]; entry j = (feature id,
 * n * F + field(feature)); absent (negative) and unknown features become (pad_row, n * F) -- sort them with dctr_sort_pairs
 * (payload = upd_pos) and fold with dctr_scatter_sgd_masked(skip_row = pad_row).  X = number of entries of feat_field: a feature
 * id >= X is reported through dctr_set_id_status (the reference raises KeyError there) and treated as absent. */
int dctr_arrange_update_list(const int32_t* feats, const int32_t* feat_field, int N, int J, int F, int64_t X, int32_t pad_row,
                             int32_t* ids, float* wts, int32_t* upd_rows, int32_t* upd_pos, void* stream);

/* ---- eval metrics on the device (baseline.py:246-259 watch_train: roc_auc_score, mean_squared_error; log_loss) --------
 * out (device, 4 doubles) = { ROC AUC with ties by average rank, -1 when one class is absent (the reference's ValueError
 * branch); mean log-loss (predictions clipped to [1e-15, 1 - 1e-15]); RMSE; number of positives }.
 * pred / y: float32 [N] on the device, labels are 0 / 1, N < 2^30.  Deterministic (integer rank sums, fixed-order doubles). */
size_t dctr_eval_metrics_workspace(int64_t N);
int dctr_eval_metrics(const float* pred, const float* y, int64_t N, double* out, void* workspace, size_t workspace_bytes,
                      void* stream);

/* HOST helper (no device work, no stream): pack one batch of the TF-form feed into the [N, F] staging blocks,
 * F = n_num + n_cat.  ids[n, j] = j, wts[n, j] = v_wts[n, j] for the numeric fields (FNN.py:41-43);
 * ids[n, n_num + j] = c_ids[n, j] + offsets[j] (the offset add of FNN.py:65-69), wts = c_wts (NULL -> 1).
 * c_ids: dense [N, n_cat] int32 (id_bytes 4) or int64 (id_bytes 8); rows are split over <= threads host threads. */
int dctr_host_pack_fields(const void* c_ids, int id_bytes, const int64_t* offsets, const float* v_wts,
                          const float* c_wts, int64_t N, int n_num, int n_cat, int32_t* ids, float* wts, int threads);

/* ---- row-sharded embedding table (BASELINE config 5).  NO COUNTERPART IN THE REFERENCE, which is single-process (SURVEY section 2:
 * "Parallelism strategies / communication backend in the reference: none"); the entry points from here to dctr_cross_bwd_push --
 * except dctr_cross_bwd, the autodiff of FNN_IP_L3.py:91-97 / FNN_OP_L3.py:93-100 -- implement SURVEY 8(e), third row. ----
 * row gather: out[i, 0:ld) = table[ids[i], 0:ld), i < M.  The owner-side lookup of the row-sharded table (SURVEY 8(e):
 * rows live on rank row % world) and the permutation of per-slot gradient rows into all-to-all order. */
int dctr_gather_rows(const int32_t* ids, const float* table, int64_t X, int ld, int64_t M, float* out, void* stream);
/* fixed-capacity exchange plan of the row-sharded table: the M slots of a batch, ordered by owner rank (perm /
 * owner_sorted: the stable sort of the slots on the key ids % world), are laid out as `world` buckets of C entries:
 *   send_ids[o*C + k] = ids[slot] / world (the owner-local row), src_of_dest[o*C + k] = slot, inv[slot] = o*C + k;
 * unused entries are -1 / 0.  Equal-sized buckets make the three all-to-alls of a step static (no split sizes read on the host,
 * CUDA-graph capturable).  status[0] += slots that did not fit (the step is then invalid), status[1] = max(status[1], largest fill). */
int dctr_shard_pack(const int32_t* ids, const int32_t* perm, const int32_t* owner_sorted, int64_t M, int world, int64_t C,
                    int32_t* send_ids, int32_t* src_of_dest, int32_t* inv, int32_t* status, void* stream);
/* ... and the delivery of the id buckets in the same launch: bucket o, its unused tail included, is also stored as block `rank` of
 * peers[o] ([world*C] int32, memory of rank o mapped with dctr_ipc_open; peers[rank] = the local list).  Replaces the id all-to-all;
 * the caller orders it against the readers with a barrier. */
int dctr_shard_pack_push(const int32_t* ids, const int32_t* perm, const int32_t* owner_sorted, int64_t M, int world, int64_t C,
                         int32_t* send_ids, int32_t* src_of_dest, int32_t* inv, int32_t* status, int rank, int32_t* const* peers,
                         void* stream);

/* backward of the cross layer (the autodiff of FNN_IP_L3.py:91-97 / FNN_OP_L3.py:93-100 and of the gather FNN.py:76-85) into a
 * per-(sample,field) embedding-gradient buffer
 *   ge[n,f,c] = wts[n,f] * ( dz[n, f*d+c] + d(cross)/d(e_f[c]) ),  ge: float32 [N, F, ldg], ldg%4==0
 * dz is dL/dz (the activation derivative already applied), TF layout (bias last).
 * cross = NONE just copies/scales the embedding block.  Reads the PRE-update table. */
int dctr_cross_bwd(const int32_t* ids, const float* wts, const float* table, int64_t X, int ld,
                   int N, int F, int d, int cross, const float* dz, int ldz,
                   float* ge, int ldg, void* stream);
/* the same, written in a caller-given row order: slot (n, f) goes to row dest[n*F + f] of ge -- the row-sharded step writes the
 * per-slot gradients straight into the owner buckets of its all-to-all (dest = the inv of dctr_shard_pack) */
int dctr_cross_bwd_to(const int32_t* ids, const float* wts, const float* table, int64_t X, int ld,
                      int N, int F, int d, int cross, const float* dz, int ldz,
                      float* ge, int ldg, const int32_t* dest, void* stream);

/* ---- the row-sharded exchanges as kernels that store into PEER memory (NVLink / NVSwitch), SURVEY 8(e) -------------------------
 * Exchange buffers other ranks of the node can map: dctr_ipc_alloc (cudaMalloc + cudaIpcGetMemHandle, zero-filled; handle64 = the
 * 64-byte CUDA IPC handle to pass to the other processes), dctr_ipc_open (cudaIpcOpenMemHandle with lazy peer access) on the
 * receiving side, dctr_ipc_close / dctr_ipc_free at teardown.  peers = DEVICE array of `world` base pointers (own buffer included).
 * dctr_gather_rows_push: the owner-side lookup + delivery in one kernel: entry k of bucket s of ids [world * C] (the owner-local
 *   rows rank s asked for; -1 = unused, not transmitted) is stored to row rank * C + k of rank s's row buffer peers[s] ([world*C, ld]).
 * dctr_cross_bwd_push: dctr_cross_bwd_to whose destination bucket o = dest / C lives on rank o: the per-slot gradient row goes to
 *   row rank * C + dest % C of peers[o] ([world*C, ldg]).
 * Ordering is the caller's: a collective (any NCCL call all ranks take part in) between the last reader of a buffer and the next
 * remote writer, and one after the pushing kernel before the owner reads. */
int dctr_ipc_alloc(size_t bytes, void** ptr, unsigned char* handle64);
int dctr_ipc_open(const unsigned char* handle64, void** ptr);
int dctr_ipc_close(void* ptr);
int dctr_ipc_free(void* ptr);
int dctr_gather_rows_push(const int32_t* ids, const float* table, int64_t X, int ld, int64_t C, int world, int rank,
                          float* const* peers, void* stream);
/* models without a cross layer (TF layout, embedding block first): the same delivery walked in BUCKET order, so that every remote
 * store streams contiguous rows: entry i (slot src_of_dest[i], valid where send_ids[i] >= 0; both from dctr_shard_pack) =
 * wts[slot] * dz[n, f*d .. f*d+d) -> row rank * C + i % C of peers[i / C] ([world*C, ldg]) */
int dctr_slot_grads_push(const int32_t* send_ids, const int32_t* src_of_dest, const float* wts, const float* dz, int64_t ldz,
                         int F, int d, int ldg, int64_t C, int world, int rank, float* const* peers, void* stream);
int dctr_cross_bwd_push(const int32_t* ids, const float* wts, const float* table, int64_t X, int ld,
                        int N, int F, int d, int cross, const float* dz, int ldz, int ldg, const int32_t* dest,
                        float* const* peers, int64_t C, int rank, void* stream);

/* ---- K1fm: fused gather + FM forward + sigmoid cross-entropy (FM.py:55-64, 33-38) -------------
 *   s[n,c] = sum_f x_f*row_f[c];  logit = s[n,0] + b + 0.5*sum_{c>=1}( s[n,c]^2 - sum_f x_f^2 row_f[c]^2 )
 *   pred = sigmoid(logit); loss_n = max(z,0) - z*y + log1p(exp(-|z|)); g[n] = (pred - y)*gscale
 * y == NULL: inference only (loss_n, g untouched).  s: float32 [N, lds] (lds%4==0) or NULL. */
int dctr_fm_fwd(const int32_t* ids, const float* wts, const float* table, int64_t X, int ld,
                int N, int F, int d, const float* bias, const float* y, float gscale,
                float* logit, float* pred, float* loss_n, float* g, float* s, int lds, void* stream);

/* ---- K6a: stable LSD radix sort of the (row, position) list ----------------------------------
 * keys = ids[0..M), payload = position p = n*F+f.  Stable, so equal rows keep sample order: this is
 * what makes the duplicate-row fold deterministic and lets SEQ_DECAY follow FNN_wnzh.py:300-306. */
size_t dctr_sort_rows_workspace(int64_t M);
int dctr_sort_rows(const int32_t* ids, int64_t M, int64_t X, int32_t* sorted_rows, int32_t* sorted_pos,
                   void* workspace, size_t workspace_bytes, void* stream);
/* The same with an explicit payload per key (NULL = the element index): sorted_pos receives payload[i] of the i-th smallest
 * key, ties in input order.  Keys outside [0, X) are reported through dctr_set_id_status, sort behind every valid row and are
 * never folded. */
int dctr_sort_pairs(const int32_t* ids, const int32_t* payload, int64_t M, int64_t X, int32_t* sorted_rows,
                    int32_t* sorted_pos, void* workspace, size_t workspace_bytes, void* stream);

/* ---- K6b: segmented fold + SGD into the touched rows only -------------------------------------
 * dense source (FNN*, Theano FNN):  grad(p=n*F+f, c) = (wts? wts[p] : 1) * g[n*g_ld_n + f*g_ld_f + c]
 *   SUM:        row -= lr * sum_p grad(p,:)
 *   SEQ_DECAY:  for p in segment order: row = row*decay - lr*grad(p,:)
 * The fold is flat over the sorted list: 4-lane groups fold DCTR_FOLD_CHUNK consecutive elements, pieces that
 * cross chunk / tile borders are composed in list order (fixed association => bit-reproducible).
 * rows / pos / table / workspace must be 16-byte aligned.
 * decay is a DOUBLE: the Theano factor 1 - 2*lambda_fm*lr/N (FNN_wnzh.py:305) is 1 - 4.9e-8 at the headline configuration, below
 * float32's resolution at 1.0; the kernels apply it as x - (1-decay)*x and decay^m as exp(m*log(decay)).
 * workspace: dctr_scatter_workspace(M, ld) bytes (one carry-in and one carry-out partial per tile). */
#ifndef DCTR_FOLD_CHUNK
#define DCTR_FOLD_CHUNK 4    /* sorted elements folded serially by one 4-lane group (4 or 8; 4: <= 64 registers, 4 CTAs per SM) */
#endif
#define DCTR_FOLD_TILE (64 * DCTR_FOLD_CHUNK)   /* sorted elements per CTA (64 groups) */
size_t dctr_scatter_workspace(int64_t M, int ld);
int dctr_scatter_sgd(int mode, const int32_t* sorted_rows, const int32_t* sorted_pos, int64_t M, int F, int d,
                     const float* g, int64_t g_ld_n, int g_ld_f, const float* wts,
                     float lr, double decay, float* table, int64_t X, int ld,
                     void* workspace, size_t workspace_bytes, void* stream);
/* The same fold for 0 / 1 weights where 0 only occurs on ONE row -- the all-zero pad row that stands in for a field a sample
 * does not have (feats_to_layer_one_array leaves x = 0 there, FNN_wnzh.py:91-95): every slot has weight 1 and the segment of
 * `skip_row` is never written, so the per-slot weight gather (a random 4-byte read per slot: 20 % of the fold's time on
 * lists larger than L2) disappears.  skip_row = -1 skips nothing. */
int dctr_scatter_sgd_masked(int mode, const int32_t* sorted_rows, const int32_t* sorted_pos, int64_t M, int F, int d,
                            const float* g, int64_t g_ld_n, int g_ld_f, int32_t skip_row,
                            float lr, double decay, float* table, int64_t X, int ld,
                            void* workspace, size_t workspace_bytes, void* stream);
/* FM source (FM.py:55-64 differentiated):
 *   grad(p,0) = g[n]*x_p ; grad(p,c>=1) = g[n]*( x_p*s[n,c] - x_p^2*row_old[c] ) ; row = row*decay - lr*sum grad
 *   (decay = 1 - lr*lambda realises the L2 term of FM.py:39 on the touched rows). */
int dctr_scatter_sgd_fm(const int32_t* sorted_rows, const int32_t* sorted_pos, int64_t M, int F, int d,
                        const float* g, const float* s, int lds, const float* wts,
                        float lr, double decay, float* table, int64_t X, int ld,
                        void* workspace, size_t workspace_bytes, void* stream);

/* Gradient-only form of the FM fold (for optimizers that need the densified gradient, i.e. Adam): rows are read from
 * `table`, and for every touched row   grad_out[row, c] = sum_p g x s_c - table[row, c] * sum_p g x^2   (c >= 1),
 * grad_out[row, 0] = sum_p g x   is WRITTEN (not accumulated); untouched rows of grad_out are left alone. */
int dctr_scatter_grad_fm(const int32_t* sorted_rows, const int32_t* sorted_pos, int64_t M, int F, int d,
                         const float* g, const float* s, int lds, const float* wts, const float* table, int64_t X, int ld,
                         float* grad_out, void* workspace, size_t workspace_bytes, void* stream);
/* dense L2 sweep for FM.py:39 on the rows the batch did NOT touch: table[r,:] *= decay unless r is in
 * sorted_rows (binary search).  Pure streaming read-modify-write over X*ld floats. */
int dctr_decay_sweep(float* table, int64_t X, int ld, float decay,
                     const int32_t* sorted_rows, int64_t M, void* stream);

/* ---- K4: fp32 GEMM with fused MLP epilogues ---------------------------------------------------
 * replaces tf.matmul + bias + activate + dropout (FNN.py:86-94) and T.dot/T.tanh (FNN_wnzh.py:147-170)
 * and their autodiff.  C[M,N] = op(A)*op(B):
 *   NN: A[M,K] (lda), B[K,N] (ldb)      forward
 *   NT: A[M,K] (lda), B[N,K] (ldb)      dgrad  (dY * W^T)
 *   TN: A[K,M] (lda), B[K,N] (ldb)      wgrad  (act^T * dY), split-K over `splits` slabs
 * epilogues (aux: [M,N] with ldaux, mask: [M,N] with ldmask, ldmask==0 broadcasts one row, NULL = 1):
 *   FWD:   C = act(acc + bias[n]) * mask
 *   DGRAD: C = acc * mask * act'(aux/mask)      (aux = the stored forward output of that layer)
 * TN writes `splits` partial slabs into C (slab stride M*ldc) when splits > 1; fold them with
 * dctr_splitk_reduce (fixed order). */
int dctr_gemm(int layout, int M, int N, int K, const float* A, int lda, const float* B, int ldb,
              float* C, int ldc, int epilogue, int act, const float* bias,
              const float* aux, int ldaux, const float* mask, int ldmask, int splits, void* stream);
int dctr_splitk_reduce(const float* partials, int splits, int M, int N, int ldc, float* out, int ldo, void* stream);
/* the same fold with the SGD step applied: W = W * decay - lr * (sum of the slabs); the weight gradient never goes to memory
 * (plain SGD of FNN_wnzh.py:176-183 / GradientDescentOptimizer on one GPU) */
int dctr_splitk_reduce_sgd(const float* partials, int splits, int M, int N, int ldc, float* W, int ldw, float lr, float decay,
                           void* stream);
/* 3xTF32 error-compensated tensor-core variant (tcgen05.mma kind::tf32, TMEM accumulators, operands split into
 * tf32 hi + lo parts on their way into shared memory: acc += A_lo*B_hi + A_hi*B_lo + A_hi*B_hi).  Same contract,
 * layouts, epilogues and split-K slab convention as dctr_gemm; results agree with it to ~1e-6 relative. */
int dctr_gemm_tc(int layout, int M, int N, int K, const float* A, int lda, const float* B, int ldb,
                 float* C, int ldc, int epilogue, int act, const float* bias,
                 const float* aux, int ldaux, const float* mask, int ldmask, int splits, void* stream);

/* ---- K5: output unit + loss (FNN.py:54-63, FNN_wnzh.py:169-173) --------------------------------
 *   logit[n] = <a[n,0:H], w> + b ; pred = sigmoid ; loss_n = stable xent ; dlogit[n] = (pred-y)*gscale */
/* loss_sum != NULL additionally writes loss_scale * sum_n loss_n in the same launch (CTA partials folded in CTA order by
 * the CTA that takes the last ticket).  workspace: dctr_reduce_workspace(N, 1) bytes, ZERO-FILLED ONCE before first use;
 * the kernels leave it ready for the next launch (CUDA-graph replayable). */
size_t dctr_reduce_workspace(int N, int H);
int dctr_out_xent_fwd(const float* a, int lda, int N, int H, const float* w, const float* b, const float* y,
                      float gscale, float* logit, float* pred, float* loss_n, float* dlogit,
                      float loss_scale, float* loss_sum, void* workspace, size_t workspace_bytes, void* stream);
/*   da[n,h] = dlogit[n]*w[h] * mask * act'(a/mask)   (backward through the last hidden activation) */
int dctr_out_bwd(const float* dlogit, const float* w, const float* a, int lda, int N, int H, int act,
                 const float* mask, int ldmask, float* da, int ldda, void* stream);
/*   the same plus the output unit's own gradients in ONE launch: gw[h] = sum_n dlogit[n]*a[n,h], gb = sum_n dlogit[n]
 *   (fixed-order two-level sum).  workspace: dctr_reduce_workspace(N, H) bytes, zero-filled once. */
int dctr_out_bwd_fused(const float* dlogit, const float* w, const float* a, int lda, int N, int H, int act,
                       const float* mask, int ldmask, float* da, int ldda, float* gw, float* gb,
                       void* workspace, size_t workspace_bytes, void* stream);

/* The whole output unit of a TRAIN step in one pass over the last activation a [N, H] (H <= 1024): what dctr_out_xent_fwd
 * (with y and loss_sum) followed by dctr_out_bwd_fused computes -- logit / pred / loss_n / dlogit (each may be NULL),
 * loss_sum = loss_scale * sum loss_n (may be NULL), da, gw [H], gb [1].  workspace: dctr_reduce_workspace(N, H + 1) bytes,
 * zero-filled once.  Fixed summation order (rows per warp, warps per CTA, CTAs by ticket). */
int dctr_out_unit_train(const float* a, int lda, int N, int H, const float* w, const float* b, const float* y,
                        float gscale, int act, const float* mask, int ldmask, float* logit, float* pred,
                        float* loss_n, float* dlogit, float loss_scale, float* loss_sum, float* da, int ldda,
                        float* gw, float* gb, void* workspace, size_t workspace_bytes, void* stream);
/*   out[h] = sum_n (wv? wv[n]:1) * A[n,h]   single-launch fixed-order column reduction (bias / last-layer grads);
 *   workspace: dctr_col_wsum_workspace(N,H) bytes, zero-filled once */
size_t dctr_col_wsum_workspace(int N, int H);
int dctr_col_wsum(const float* A, int lda, int N, int H, const float* wv, float* out,
                  void* workspace, size_t workspace_bytes, void* stream);
/*   the same with the SGD step of a bias applied: param[h] = param[h] * decay - lr * sum_n (wv? wv[n]:1) * A[n,h] */
int dctr_col_wsum_sgd(const float* A, int lda, int N, int H, const float* wv, float* param, float lr, float decay,
                      void* workspace, size_t workspace_bytes, void* stream);
/*   dctr_col_wsum for callers that cannot allocate (code under CUDA graph capture, serving loops, harnesses that forbid
 *   cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same result as dctr_col_wsum,
 *   bit for bit; the scratch is a device array built into the library (4.63 MiB per GPU that loads it, held until exit), so
 *   H <= dctr_col_wsum_static_max_h() = 1024 -- wider matrices are rejected with DCTR_EINVAL and a message.  The call itself
 *   allocates nothing (no cudaMalloc / cudaFree / allocating library routine).  Every call shares that array: calls are
 *   ordered by their stream; calls that overlap on the GPU (streams without an ordering between them, several host threads on
 *   one device) race on it.  The ticket words are cleared on `stream` before each launch. */
int dctr_col_wsum_static_max_h(void);
int dctr_col_wsum_static(const float* A, int lda, int N, int H, const float* wv, float* out, void* stream);
/*   out[0] = sum_n v[n]  (fixed order, one CTA) -- loss = reduce_sum(log_loss) (FM.py:36) */
int dctr_vec_sum(const float* v, int N, float scale, float* out, void* stream);

/*   out[0] += scale * ( sum(v[0:n)^2) + sum(extra[0:n_extra)^2) ) / 2 -- the regulariser lam * (l2_loss(W) + l2_loss(V) + l2_loss(b)) that
 *   FM.py:39 / LR.py:37 add to the loss the caller fetches (tf.nn.l2_loss = sum(t^2) / 2).  n % 4 == 0, v 16-byte aligned; partial
 *   sums in double, fixed order.  workspace: dctr_reduce_workspace(1, 1) bytes, zero-filled once. */
int dctr_l2_loss_add(const float* v, int64_t n, const float* extra, int n_extra, float scale, float* out,
                     void* workspace, size_t workspace_bytes, void* stream);

/* tf.nn.dropout masks (FNN.py:87-89) for all layer inputs of a step in one launch: masks[i] = Bernoulli(keep) / keep from
 * Philox-4x32-10 keyed by (seed, state[0]); state = {step counter, ticket} (2 words, zero before the first step), advanced on
 * the device so that a captured CUDA graph draws fresh masks on every replay.  apply != NULL: apply[0:apply_n) *= masks[0:apply_n)
 * (the first layer's input).  n, apply_n multiples of 4; 16-byte aligned buffers. */
int dctr_dropout_masks(float* masks, int64_t n, float keep, unsigned seed, unsigned* state, float* apply, int64_t apply_n,
                       void* stream);

/* ---- K7: dense SGD over a flat parameter buffer (tf_util.py:27 / FNN_wnzh.py:180-182) ---------- */
int dctr_dense_sgd(float* param, const float* grad, int64_t n, float lr, float decay, void* stream);

/* Adam in TensorFlow's form (tf_util.py:18-21 AdamOptimizer(lr, epsilon); baseline.py:93 uses it for every Criteo model).
 * dctr_adam_tick advances state = [t, lr_t, beta1^t, beta2^t] (4 floats, zero before the first step) on the device:
 *   lr_t = lr * sqrt(1 - beta2^t) / (1 - beta1^t)          -- once per step, graph-replay safe.
 * dctr_dense_adam, elementwise over n floats (n % 4 == 0, 16-byte aligned buffers):
 *   g += l2 * p;  m = beta1 m + (1-beta1) g;  v = beta2 v + (1-beta2) g^2;  p -= lr_t * m / (sqrt(v) + eps);
 *   zero_grad != 0 clears grad afterwards (the densified table gradient is rebuilt by dctr_scatter_sgd every step). */
int dctr_adam_tick(float* state, float lr, float beta1, float beta2, void* stream);
int dctr_dense_adam(float* param, float* grad, float* m, float* v, int64_t n, const float* state, float beta1, float beta2,
                    float eps, float l2, int zero_grad, void* stream);

/* FTRL-proximal in TensorFlow's ApplyFtrl form (tf_util.py:22-24 FtrlOptimizer(lr): lr_power = -0.5, accumulators start
 * at 0.1, l1 = l2 = 0; TensorFlow itself is not under /root/reference, the update below is its published kernel).
 * Elementwise over n floats (n % 4 == 0, 16-byte aligned buffers), accum > 0 on entry:
 *   g += grad_l2 * p                                   (the loss's own L2 term, FM.py:39 / LR.py:37)
 *   n' = accum + g^2;  linear += g - (n'^(-lr_power) - accum^(-lr_power)) / lr * p
 *   p = |linear| > l1 ? (l1 * sign(linear) - linear) / (n'^(-lr_power) / lr + 2 * l2) : 0;  accum = n'
 * Dense, as TF applies it to these models: an element with g = 0 is re-derived from `linear` alone, so a weight that has
 * not seen a gradient yet becomes 0 on the first step.  zero_grad != 0 clears grad afterwards. */
int dctr_dense_ftrl(float* param, float* grad, float* accum, float* linear, int64_t n, float lr, float lr_power,
                    float l1, float l2, float grad_l2, int zero_grad, void* stream);

/* ---- fused train step of a two-hidden-layer MLP at small batches (csrc/mlp2.cu) -------------------------------------------
 * replaces the whole Theano graph of FNN_wnzh.py:147-182 (and FNN.py:86-94 at small widths) -- forward, loss, T.grad, the SGD
 * updates -- with two launches instead of ~14 per-layer ones.  fp32 FFMA (the layers are far too small for tensor-core tiles).
 *   x [N, ld0] = act0(z) of the embedding layer (dctr_gather_concat_fwd); m0 / m1 / m2: dropout masks on x / a1 / a2
 *   ([N, ld] with ldm = row stride, or one broadcast row with ldm = 0, or NULL); W1 [D0, ldw1], W2 [H1, ldw2], w3 [H2].
 * dctr_mlp2_train: a1 = m1*act1(x W1 + b1), a2 = m2*act2(a1 W2 + b2), logit = a2 w3 + b3, pred, loss_n, dlogit = (pred-y)*gscale,
 *   loss_sum = loss_scale * sum loss_n, then dz2, dz1 [N, ldd] (gradients w.r.t. the pre-activations) and dz0 [N, ld0] (the
 *   gradient w.r.t. z: Theano's gx, what the sparse update folds).  x is overwritten by x*m0 when m0 != NULL; a1 is stored.
 *   W1t [H1, ld0] / W2t [H2, ld1] are the TRANSPOSED weights (dctr_mlp2_transpose) the backward GEMMs stream.
 *   sgd != 0: the kernel's last CTA applies w3 = w3*decay_out - lr*gw3 (b3 likewise); else gw3 [H2], gb3 [1] are written.
 * dctr_mlp2_wgrad: gW1 = x^T dz1, gb1 = colsum dz1, gW2 = a1^T dz2, gb2 = colsum dz2 in one split-K launch (fixed order);
 *   sgd != 0: W = W*decay - lr*gW in place (biases likewise), else the gradients are stored (same leading dimensions as W).
 * workspace: dctr_mlp2_workspace bytes, zero-filled once, shared by the three entry points.  Supported: dctr_mlp2_supported. */
int dctr_mlp2_supported(int D0, int H1, int H2);
size_t dctr_mlp2_workspace(int N, int D0, int H1, int H2);
int dctr_mlp2_transpose(int D0, int H1, int H2, const float* W1, int ldw1, const float* W2, int ldw2, float* W1t, int ld0,
                        float* W2t, int ld1, void* stream);
int dctr_mlp2_train(int N, int D0, int H1, int H2, float* x, int ld0, const float* m0, int ldm0, int act0,
                    const float* W1, int ldw1, const float* b1, int act1, const float* m1, int ldm1,
                    const float* W2, int ldw2, const float* b2, int act2, const float* m2, int ldm2,
                    const float* W1t, const float* W2t, float* w3, float* b3, const float* y, float gscale, float loss_scale,
                    float* a1, int ld1, float* dz1, float* dz2, int ldd, float* dz0,
                    float* logit, float* pred, float* loss_n, float* dlogit, float* loss_sum,
                    float* gw3, float* gb3, int sgd, float lr, float decay_out,
                    void* workspace, size_t workspace_bytes, void* stream);
int dctr_mlp2_wgrad(int N, int D0, int H1, int H2, const float* x, int ld0, const float* a1, int ld1,
                    const float* dz1, const float* dz2, int ldd, float* W1, int ldw1, float* b1, float* W2, int ldw2,
                    float* b2, float* gW1, float* gb1, float* gW2, float* gb2, int sgd, float lr, float decay,
                    void* workspace, size_t workspace_bytes, void* stream);

/* ---- the same fused step on the tensor cores (csrc/mlp2tc.cu): one CTA per 128 samples chains x -> a1 -> a2 -> output unit ->
 * dz2 -> dz1 -> dz0 through tcgen05 (3xTF32, accumulators in tensor memory, activations handed from an epilogue to the next
 * phase's A tiles in shared memory).  Same contract and outputs as dctr_mlp2_train, except that the hidden-layer weights are
 * read from `packed`: dctr_mlp2tc_pack of W1 [D0, ldw1] and W2 [H1, ldw2] (split into tf32 hi / lo and laid out as the swizzled
 * shared-memory tiles of the four GEMM phases in consumption order; dctr_mlp2tc_packed_bytes bytes, 128-byte aligned; to be
 * redone whenever W1 / W2 change) -- every weight tile is one contiguous bulk copy, no transposed copies, no tensor maps.
 * tanh / sigmoid are evaluated through ex2.approx / rcp.approx (absolute error < 3e-7).  The weight gradients are the caller's
 * (dctr_mlp2_wgrad or the TN GEMMs).  Limits: dctr_mlp2tc_supported (D0 <= 192, H1 <= 320, H2 <= 128).
 * workspace: dctr_mlp2tc_workspace bytes, zero-filled once.  DCTR_M2TC_TIMING=1 prints a %globaltimer timeline of CTA 0. */
int dctr_mlp2tc_supported(int D0, int H1, int H2);
size_t dctr_mlp2tc_workspace(int N, int H2);
size_t dctr_mlp2tc_packed_bytes(int D0, int H1, int H2);
int dctr_mlp2tc_pack(int D0, int H1, int H2, const float* W1, int ldw1, const float* W2, int ldw2, void* packed,
                     size_t packed_bytes, void* stream);
int dctr_mlp2tc_train(int N, int D0, int H1, int H2, float* x, int ld0, const float* m0, int ldm0, int act0,
                      const void* packed, const float* b1, int act1, const float* m1, int ldm1,
                      const float* b2, int act2, const float* m2, int ldm2,
                      float* w3, float* b3, const float* y, float gscale, float loss_scale,
                      float* a1, int ld1, float* dz1, float* dz2, int ldd, float* dz0,
                      float* logit, float* pred, float* loss_n, float* dlogit, float* loss_sum,
                      float* gw3, float* gb3, int sgd, float lr, float decay_out,
                      void* workspace, size_t workspace_bytes, void* stream);

/* ---- embedding bag: the first layer of the SNN models (get_fi_h1_y, SNN_DAE.py:234-260, SNN_RBM.py:238-262)
 *   out[n, c] = act( b0[c] + sum_f wts[n,f] * table[ids[n,f], c] ),  c < H;  columns H..ldo-1 are written as 0
 * table: float32 [X, ld] (ld % 4 == 0, ld >= H; the reference's ww0 is 133465 x 200), b0: [H] or NULL, act: DCTR_ACT_*
 * (the reference applies the sigmoid).  Ragged samples pad their slots with any valid row and weight 0.
 * Its backward is the existing path: d/d(pre-activation) comes out of the first dgrad GEMM's epilogue, b0's gradient is
 * dctr_col_wsum of it, and the touched-row update `ww0[feat] -= lr * gx*x*(1-x)` (SNN_DAE.py:282-288) is dctr_scatter_sgd in
 * SUM mode with g_ld_f = 0: every feature of a sample receives that sample's gradient row. */
int dctr_bag_fwd(const int32_t* ids, const float* wts, const float* table, int64_t X, int ld, int N, int F, int H,
                 const float* b0, int act, float* out, int ldo, void* stream);

/* ---- device-side ingestion of the reference's text formats -----------------------------------------
 * The raw file bytes are uploaded as they are; these entry points replace the Python parsers on the step's critical path:
 * linecache.getline + replace(':',' ').split() (get_fxy, FNN_wnzh.py:224-253), load_ipinyou_data (ipinyou.py:42-65) and
 * get_fxy (baseline.py:190-198).  text: uint8 [nbytes] on the device, 16-byte aligned.
 *
 * dctr_text_index_lines: line_start[i] = byte offset of the i-th NON-EMPTY line (a byte that is not '\n' and follows a '\n'
 * or the start of the buffer), in file order; n_lines[0] (device) = how many there are in total.  line_start has room for
 * max_lines + 1 entries: entry min(n_lines, max_lines) bounds the last line that is kept (nbytes, or the start of the first
 * line that did not fit).  workspace: dctr_text_index_workspace(nbytes) bytes, 8-byte aligned. */
size_t dctr_text_index_workspace(int64_t nbytes);
int dctr_text_index_lines(const uint8_t* text, int64_t nbytes, int64_t* line_start, int64_t max_lines, int64_t* n_lines,
                          void* workspace, size_t workspace_bytes, void* stream);
/* `y [skip tokens] id:val id:val ...` -- train.fm.txt (skip = 0: y = int(s[0]), ids at the odd positions, values ignored;
 * FNN_wnzh.py:248-252) and *.yzx.txt (skip = 1: fields[2:], x.split(':')[0]; ipinyou.py:50-59).  Per line i < min(n_lines,
 * max_lines):  y[i];  feats[i, j] = id of the j-th feature, wts[i, j] = 1 (values are forced to 1, ipinyou.py:55) for
 * j < min(nnz, J);  the remaining slots get (pad_id, 0) -- pad_id = -1 feeds dctr_arrange_fields, pad_id = X_dim - 1 is the
 * padding of ipinyou.py:58-59;  status[i] = nnz (may exceed J: the line was truncated), -1 for a whitespace-only line,
 * -2 for a malformed token.  n_lines is read on the device (no host sync between indexing and parsing); wts may be NULL. */
int dctr_text_parse_sparse(const uint8_t* text, int64_t nbytes, const int64_t* line_start, const int64_t* n_lines,
                           int64_t max_lines, int skip, int J, int32_t pad_id, int32_t* feats, float* wts, float* y,
                           int32_t* status, void* stream);
/* `y \t n_1 .. n_K \t c_1 .. c_C` (Criteo indexed TSV written by clean.py:148-184, read by get_fxy, baseline.py:190-198),
 * F = n_num + n_cat.  Output is the engines' [N, F] block (what dctr_host_pack_fields builds from the TF feed):
 *   ids[i, k] = k, wts[i, k] = (float)(n_k / max_vals[k]) (double division, as Python's);  ids[i, n_num + j] =
 *   c_j + offsets[j], wts = 1;  y[i];  status[i] = number of tab-separated fields (1 + F when well formed), -2 if a field is
 *   empty or not a number.  offsets: int64 [n_cat], max_vals: double [n_num], both on the device. */
int dctr_text_parse_criteo(const uint8_t* text, int64_t nbytes, const int64_t* line_start, const int64_t* n_lines,
                           int64_t max_lines, int n_num, int n_cat, const int64_t* offsets, const double* max_vals,
                           int32_t* ids, float* wts, float* y, int32_t* status, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DCTR_H_ */
