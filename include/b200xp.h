/*
 * b200xp.h — C-ABI of libb200xp.so, the B200-native (sm_100a) experience-path engine.
 *
 * This is the drop-in boundary for the hot path of AgileRL's agilerl/components/
 * (replay_buffer.py, segment_tree.py, rollout_buffer.py).  Every entry point names the
 * reference interface it replaces (file:line relative to the AgileRL source tree).
 *
 * Conventions
 *   - All `dev` pointers are DEVICE pointers owned by the caller (the PyTorch caching
 *     allocator in the Python host).  The library owns no device memory and never allocates.
 *   - Every device op is ASYNCHRONOUS on the `stream` argument (a cudaStream_t passed as
 *     void*; NULL = legacy default stream).  No entry point synchronises.
 *   - Return value: XP_OK (0) or a negative xp_status_t; the message is in xp_last_error()
 *     (thread-local).  The Python host maps XP_ERR_INDEX -> AssertionError (index range,
 *     replay_buffer.py:319; power-of-two capacity, segment_tree.py:20-23) and
 *     XP_ERR_INVALID -> ValueError.
 *   - One host thread per buffer (the reference is single-threaded per buffer); different
 *     buffers may use different streams.
 *   - Integer / index / tree results are bit-exact against the reference; floating point
 *     follows the reference's rounding sequence (no FMA contraction) — see DESIGN.md.
 */
#ifndef B200XP_H_
#define B200XP_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int xp_status_t;
#define XP_OK 0
#define XP_ERR_INVALID (-1) /* bad argument (shape, null pointer, size)            */
#define XP_ERR_INDEX (-2)   /* index out of range / capacity not a power of two     */
#define XP_ERR_CUDA (-3)    /* CUDA runtime error; text in xp_last_error()          */
#define XP_ERR_STATE (-4)   /* call made in the wrong state                         */

#define XP_MAX_KEYS 24   /* keys (columns) per ring/gather launch                 */
#define XP_MAX_NSTEP 32  /* n of the n-step fold                                   */

/* ---- library ---------------------------------------------------------------------- */
int xp_version(void);               /* ABI version, bumped on any signature change  */
const char *xp_last_error(void);    /* thread-local, never NULL                     */
/* Number of CUDA kernels this library has launched since it was loaded (all threads).  */
uint64_t xp_launch_count(void);
/* Device facts used for grid sizing (SM count, compute capability). */
xp_status_t xp_device_info(int device, int *sm_count, int *cc_major, int *cc_minor);

/* Host staging: asynchronous copy of a small host array to device memory (or back) on `stream`.
 * Replaces the `.to(device)` of the per-call host inputs of the reference: the torch.rand draws
 * (replay_buffer.py:377), the priorities / indices that arrive as numpy arrays (:411-428,
 * algorithms/dqn_rainbow.py:487-488).  Truly asynchronous only for pinned host memory; the
 * caller keeps the host array alive and unmodified until the stream has passed this point.  */
xp_status_t xp_copy_to_device(void *dst_dev, const void *src_host, int64_t nbytes, void *stream);
xp_status_t xp_copy_to_host(void *dst_host, const void *src_dev, int64_t nbytes, void *stream);

/* ---- keys: one column of a transition table ---------------------------------------- */
/* A replay storage is struct-of-arrays: one [max_size, row_bytes] device array per key
 * (replay_buffer.py:60-70).  `src`/`dst` meaning depends on the op.                     */
typedef struct xp_key {
  const void *src;   /* device pointer                                                 */
  void *dst;         /* device pointer                                                 */
  int64_t row_bytes; /* bytes per row of this key                                      */
} xp_key;

/* ---- ring insert / indexed gather (ReplayBuffer.add / storage[idx]) ------------------ */
/* replaces ReplayBuffer.add's circular slice-assign (replay_buffer.py:99-107):
 * for r in [0,n_rows): dst[(cursor + r) % max_size] = src[r], for every key, one launch. */
xp_status_t xp_ring_insert(const xp_key *keys, int n_keys, int64_t n_rows, int64_t cursor,
                           int64_t max_size, void *stream);
/* replaces TensorDict index-gather `storage[indices]` + `.clone()`
 * (replay_buffer.py:126,204,345-346): dst[b] = src[idx[b]] for every key, one launch.
 * idx_dev: int64[B]. Rows with idx outside [0,n_src_rows) set *err_flag_dev (if not NULL)
 * to 1 and are written as zeros.                                                        */
xp_status_t xp_gather(const xp_key *keys, int n_keys, const int64_t *idx_dev, int64_t B,
                      int64_t n_src_rows, int *err_flag_dev, void *stream);

/* ---- uniform sampling without replacement (ReplayBuffer.sample) ---------------------- */
/* replaces `torch.randperm(size)[:batch]` (replay_buffer.py:125, rollout_buffer.py:570):
 * given the raw 32-bit mt19937 draws x_0..x_{B-1} torch would consume, produce the first
 * B entries of the forward Fisher-Yates permutation of range(n), exactly, in O(B log B).
 * P independent agents per launch: raw_dev is u32[P*B], idx_out_dev int64[P*B].
 * Requires 1 <= B <= n, B <= 4096.                                                      */
xp_status_t xp_randperm_prefix(const uint32_t *raw_dev, int64_t B, int64_t P, int64_t n,
                               int64_t *idx_out_dev, void *stream);
/* Counter-based raw draws on the device (throughput mode; Philox4x32-10). out u32[n].   */
xp_status_t xp_philox_u32(uint64_t seed, uint64_t offset, uint32_t *out_dev, int64_t n,
                          void *stream);

/* ---- segment trees (SumSegmentTree / MinSegmentTree, segment_tree.py:5-182) ---------- */
/* A tree of capacity `cap` (power of two) is 2*cap fp64 nodes in heap order: node 1 is the
 * root, leaves are cap..2cap-1 (segment_tree.py:25,89-95).  Node i of a tree lives at
 * base[i*stride] (stride in doubles) so that a (sum,min) pair can be interleaved
 * (stride 2, min = sum+1) or stored separately (stride 1).  Either tree pointer may be
 * NULL to operate on one tree only.                                                     */
typedef struct xp_tree {
  double *sum;     /* device; node i at sum[i*stride]; init 0.0   (segment_tree.py:122) */
  double *min;     /* device; node i at min[i*stride]; init +inf  (segment_tree.py:170) */
  int64_t stride;  /* in doubles                                                        */
  int64_t cap;     /* power of two                                                      */
  int32_t *aux;    /* device int32[2*cap] scratch owned by the library after xp_tree_init
                      (which zeroes it): [cap,2cap) last-writer tickets (zero between calls),
                      [0,cap) bucket counters / bucketed items of large batches          */
} xp_tree;

/* fill all nodes with the initial values (SegmentTree.__init__, segment_tree.py:25).    */
xp_status_t xp_tree_init(const xp_tree *t, void *stream);
/* replaces a batch of `tree[idx] = val` (SegmentTree.__setitem__, segment_tree.py:81-95)
 * applied SEQUENTIALLY: duplicates resolve last-writer-wins, every ancestor ends up as
 * op(left,right) of its final children — bit-identical to the sequential loop.
 * idx_dev int64[B] in [0,cap); val_dev fp64[B].  Out-of-range sets *err_flag_dev.        */
xp_status_t xp_tree_set_batch(const xp_tree *t, const int64_t *idx_dev, const double *val_dev,
                              int64_t B, int *err_flag_dev, void *stream);
/* replaces SegmentTree.operate(start,end) (segment_tree.py:28-79) with the reference's
 * recursive association order; result written to out_dev[0] (sum) and out_dev[1] (min).   */
xp_status_t xp_tree_query(const xp_tree *t, int64_t start, int64_t end, double *out_dev,
                          void *stream);
/* replaces SumSegmentTree.retrieve (segment_tree.py:136-156) for a batch of upper bounds. */
xp_status_t xp_tree_retrieve(const xp_tree *t, const double *ub_dev, int64_t B,
                             int64_t *idx_out_dev, void *stream);

/* ---- prioritized replay (PrioritizedReplayBuffer, replay_buffer.py:261-428) ---------- */
/* device-resident scalar state of one PER buffer */
typedef struct xp_per_state {
  double max_priority; /* replay_buffer.py:284,329 */
} xp_per_state;

/* replaces the per-row loop of PrioritizedReplayBuffer.add (replay_buffer.py:306-309):
 * leaves (tree_ptr + r) % max_size, r in [0,n) := max_priority ** alpha in both trees.   */
xp_status_t xp_per_add(const xp_tree *t, xp_per_state *state_dev, int64_t tree_ptr, int64_t n,
                       int64_t max_size, double alpha, void *stream);
/* replaces _sample_proportional (replay_buffer.py:357-381) and _calculate_weights
 * (:383-409) for P agents x B samples in one launch.  u_dev fp32[P*B] are the torch.rand
 * draws; sample j uses stratum i = j % B.  idx_out_dev int64[P*B]; w_out_dev fp32[P*B] or
 * NULL (indices only).  `size` is the replay size (len(buffer)).                         */
xp_status_t xp_per_sample(const xp_tree *t, const float *u_dev, int64_t B, int64_t P,
                          int64_t size, double beta, int64_t *idx_out_dev, float *w_out_dev,
                          void *stream);
/* _calculate_weights alone (replay_buffer.py:383-409) for given indices int64[B].        */
xp_status_t xp_per_weights(const xp_tree *t, const int64_t *idx_dev, int64_t B, int64_t size,
                           double beta, float *w_out_dev, void *stream);
/* replaces update_priorities (replay_buffer.py:411-428): sequential semantics
 * p = max(float(prio), 1e-5); leaf = p ** alpha; max_priority = max(max_priority, p).
 * idx_dev int64[B] must lie in [0,max_size) (else *err_flag_dev = 1, row skipped);
 * prio_dev fp32[B].                                                                     */
xp_status_t xp_per_update(const xp_tree *t, xp_per_state *state_dev, const int64_t *idx_dev,
                          const float *prio_dev, int64_t B, int64_t max_size, double alpha,
                          int *err_flag_dev, void *stream);
/* The two calls above for ONE agent's batch of B <= XP_HOST_VEC_MAX rows whose fp32 vector is still in HOST memory
 * — the torch.rand draws of _sample_proportional (replay_buffer.py:377) and the ndarray update_priorities
 * receives (replay_buffer.py:411-428, from dqn_rainbow.py:487-488): the vector is read during the call and travels
 * inside the launch (a kernel parameter), so no host-to-device copy precedes the kernel and the caller may reuse
 * the host array on return.  Same arithmetic, same results as xp_per_sample (P = 1) / xp_per_update.          */
#define XP_HOST_VEC_MAX 1024
xp_status_t xp_per_sample_hostu(const xp_tree *t, const float *u_host, int64_t B, int64_t size, double beta,
                                int64_t *idx_out_dev, float *w_out_dev, void *stream);
xp_status_t xp_per_update_hostp(const xp_tree *t, xp_per_state *state_dev, const int64_t *idx_dev,
                                const float *prio_host, int64_t B, int64_t max_size, double alpha,
                                int *err_flag_dev, void *stream);
/* priority ** alpha as CPython computes it (glibc pow), batch, fp64 in/out
 * (replay_buffer.py:322,401,406).                                                       */
xp_status_t xp_pow_batch(const double *x_dev, double y, double *out_dev, int64_t n,
                         void *stream);

/* ---- n-step fold (MultiStepReplayBuffer._get_n_step_info, replay_buffer.py:206-258) --- */
/* The last n env-steps live in a device stage ring of n slots x N envs per key; `slot0`
 * is the slot of the OLDEST step.  For every env e writes one row at ring position
 * (cursor + e) % max_size of the replay storage:
 *   keys[k].src = stage base [n_step, N, row_bytes], keys[k].dst = storage base;
 *   role[k]: XP_ROLE_FIRST  -> copied from the oldest step
 *            XP_ROLE_LAST   -> copied from step m (next_obs, done)
 *            XP_ROLE_REWARD -> fp32 fold R = r_0; R += r_i * gpow[i], i = 1..m (no FMA)
 *   m = first i >= 1 whose done flag is set for ANY env, else n_step-1 (:252).
 * done_key indexes keys[]; done_elem_bytes is 1 (bool/uint8) or 4 (fp32).               */
#define XP_ROLE_FIRST 0
#define XP_ROLE_LAST 1
#define XP_ROLE_REWARD 2
xp_status_t xp_nstep_fold_insert(const xp_key *keys, const int *roles, int n_keys, int done_key,
                                 int done_elem_bytes, int n_step, int slot0, const float *gpow,
                                 int64_t N, int64_t cursor, int64_t max_size, void *stream);

/* ---- rollout returns (RolloutBuffer.compute_returns_and_advantages) ------------------- */
/* replaces the GAE branch (rollout_buffer.py:448-467) for P agents in one launch.
 * rewards/values fp32 [P][T][N] (agent_stride elements between agents, row stride N),
 * dones u8 same layout; last_value/last_done fp64 [P][N]; adv/ret fp32 out, same layout
 * as rewards.  Carry in fp64; gamma*v[t+1] rounded in fp32 for t < T-1 (:455-456), fp64 at
 * t = T-1 (:452-453); ret = fp32(adv) + v in fp32 (:467).                                */
xp_status_t xp_gae(const float *rewards, const float *values, const uint8_t *dones,
                   const double *last_value, const double *last_done, float *adv, float *ret,
                   int64_t T, int64_t N, int64_t P, int64_t agent_stride, double gamma,
                   double gae_lambda, void *stream);
/* replaces the all-fp32 GAE loop of the legacy "experiences" learn path (algorithms/ppo.py:686-713,
 * algorithms/ippo.py:703-720): torch fp32 tensors with Python-float scalars, i.e. every op rounded
 * in fp32, gamma and gamma*lambda entering as fp32; next_value / next_done fp32 [P][N] bootstrap
 * t = T-1; dones u8 (the reference's dones.long()).  ret = adv + values.                       */
xp_status_t xp_gae_f32(const float *rewards, const float *values, const uint8_t *dones,
                       const float *next_value, const float *next_done, float *adv, float *ret,
                       int64_t T, int64_t N, int64_t P, int64_t agent_stride, double gamma,
                       double gae_lambda, void *stream);
/* replaces the Monte-Carlo branch (rollout_buffer.py:468-477). */
xp_status_t xp_mc_returns(const float *rewards, const float *values, const uint8_t *dones,
                          const double *last_value, const double *last_done, float *adv,
                          float *ret, int64_t T, int64_t N, int64_t P, int64_t agent_stride,
                          double gamma, void *stream);

/* ---- token- and turn-level GAE of the LLM PPO path ------------------- */
/* All arrays are [B][S] row-major with TIME contiguous (batch, seq_len), fp32; mask is the
 * reference's action_mask.float().
 * xp_gae_tokens replaces the loop of PPO._compute_gae_returns_token (algorithms/ppo_llm.py:851-867):
 *   adv[b][t] = last_gae * mask[b][t],  ret = (adv + values) * mask,  every op rounded in fp32;
 *   final_mask != 0 also applies the closing `token_advantages * mask` (:870, the no-whitening
 *   return); with whitening the caller runs xp_masked_whiten(times_mask = 1) on adv instead.   */
xp_status_t xp_gae_tokens(const float *rewards, const float *values, const float *mask, float *adv,
                          float *ret, int64_t B, int64_t S, double gamma, double gae_lambda,
                          int final_mask, void *stream);
/* xp_gae_turns replaces pool_by_turns (utils/llm_utils.py:256-312: rewards mean-pooled, values
 * mean-pooled or, final_value != 0, taken at the last token of the turn) and the turn chain of
 * PPO._compute_gae_returns (ppo_llm.py:786-813).  turn_ids int64 [B][S], -1 = not an action token;
 * num_turns = max id + 1 (<= XP_MAX_TURNS).  Outputs [B][num_turns] fp32: pooled turn values,
 * turn advantages, turn mask (1.0 where the sequence has the turn).
 * xp_gae_turn_broadcast replaces ppo_llm.py:822-829: ret = (turn_adv + turn_values)[turn_id] * valid,
 * adv = turn_adv_pg[turn_id] * valid (turn_adv_pg = the whitened advantages, or turn_adv).      */
#define XP_MAX_TURNS 2048
xp_status_t xp_gae_turns(const float *rewards, const float *values, const int64_t *turn_ids,
                         float *turn_values, float *turn_adv, float *turn_mask, int64_t B, int64_t S,
                         int64_t num_turns, int final_value, double gamma, double gae_lambda,
                         void *stream);
xp_status_t xp_gae_turn_broadcast(const int64_t *turn_ids, const float *turn_values,
                                  const float *turn_adv, const float *turn_adv_pg, float *ret,
                                  float *adv, int64_t B, int64_t S, int64_t num_turns, void *stream);
/* masked_whiten (utils/llm_utils.py:215-254, shift_mean = True, unbiased variance):
 *   out = (x - mean) * rsqrt(var + 1e-8) [* mask when times_mask], mean / var over mask-weighted x.
 * The two sums are accumulated in fp64 in a fixed order and rounded once (torch sums in fp32 in
 * an unspecified order); everything else is the reference's fp32 op sequence.  scratch_dev:
 * XP_WHITEN_SCRATCH doubles, zero before the first call (reusable afterwards).  *err_flag_dev is
 * set when fewer than 2 values are unmasked (the reference raises ValueError).                 */
#define XP_WHITEN_SCRATCH 520
xp_status_t xp_masked_whiten(const float *x, const float *mask, float *out, int64_t n,
                             int times_mask, double *scratch_dev, int *err_flag_dev, void *stream);

/* Global advantage normalisation of PPO's flat learn step (algorithms/ppo.py:829-833, SURVEY.md 8f
 * row 2): out = (x - x.mean()) / (x.std() + 1e-8), std unbiased with fp64 accumulation as
 * torch.std; out may alias x.  n == 1 gives NaN like torch.  scratch_dev as for xp_masked_whiten. */
xp_status_t xp_standardize(const float *x, float *out, int64_t n, double *scratch_dev, void *stream);
/* xp_standardize without caller-owned scratch.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  Same kernels, same fixed summation order, hence the same
 * bits as xp_standardize.  The XP_WHITEN_SCRATCH doubles of scratch (4 160 bytes) live in a file-scope
 * `static __device__` array inside the library — every GPU that loads the library holds it until exit —
 * whose device address is looked up once per device (cudaGetSymbolAddress, with that device current) on
 * the first call made on it.  The call allocates nothing: no cudaMalloc / cudaFree, no allocating
 * library routine; it clears the one scratch word that must start at zero with a stream-ordered
 * cudaMemsetAsync.  n > XP_STANDARDIZE_STATIC_MAX_N is refused with XP_ERR_INVALID.  ALL calls on a
 * device share the one array: calls that overlap on the GPU (different streams, or different host
 * threads / processes sharing a context) race — serialise them on one stream, or use xp_standardize.  */
#define XP_STANDARDIZE_STATIC_MAX_N (1ll << 26) /* elements (256 MiB of fp32) */
xp_status_t xp_standardize_static(const float *x, float *out, int64_t n, void *stream);

/* ---- truncated-BPTT sequence preparation (RolloutBuffer.prepare_sequence_tensors) ------ */
/* replaces the per-env Python walk of rollout_buffer.py:736-762 + _get_complete_sequences
 * (:677-722) + extract_sequences_from_episode (utils/algo_utils.py:214-265).
 * dones u8 [T][N] (the first T = buffer_size stored steps).  An episode ends where a done flag
 * is set and at step T-1.  An episode of len >= max_seq_len (max_seq_len > 0) is cut into
 * sequences of max_seq_len steps: CHUNKED max(1, len / max_seq_len) non-overlapping chunks (the
 * remainder is dropped, as in the reference), MAXIMUM every window, FIFTY_PERCENT_OVERLAP windows
 * every max_seq_len / 2 steps; shorter episodes (or max_seq_len == 0, "None") are one sequence.
 * Sequences are numbered env-major, then by start step.  Outputs (device):
 *   seq_t0/seq_env/seq_len int32[>= T*N], seq_off int64[>= T*N] (offset of the sequence in the
 *   concatenated unpadded step list), header int64[4] = {num_sequences, longest sequence,
 *   longest episode, total unpadded steps}; env_scratch int32[2N + 2].                          */
#define XP_BPTT_CHUNKED 0
#define XP_BPTT_MAXIMUM 1
#define XP_BPTT_FIFTY_PERCENT_OVERLAP 2
xp_status_t xp_bptt_plan(const uint8_t *dones_dev, int64_t T, int64_t N, int64_t max_seq_len,
                         int seq_type, int32_t *env_scratch_dev, int32_t *seq_t0_dev,
                         int32_t *seq_env_dev, int32_t *seq_len_dev, int64_t *seq_off_dev,
                         int64_t *header_dev, void *stream);
/* The padded layout as row indices for xp_gather (replaces _pad_sequences, :627-675):
 *   pad_idx int64[num_seq][pad_len] = flat step t*N + env, or -1 where the sequence is shorter
 *   (xp_gather writes zero rows for negative indices == the reference's zero padding);
 *   pad_mask u8 same shape (:767-770); unp_idx int64[total unpadded steps] = the concatenated
 *   `unpadded_slices` (:752-762); first_idx int64[num_seq] (may be NULL) = first step of every
 *   sequence (initial hidden states, :818-826).                                              */
xp_status_t xp_bptt_indices(const int32_t *seq_t0_dev, const int32_t *seq_env_dev,
                            const int32_t *seq_len_dev, const int64_t *seq_off_dev, int64_t num_seq,
                            int64_t pad_len, int64_t N, int64_t *pad_idx_dev, uint8_t *pad_mask_dev,
                            int64_t *unp_idx_dev, int64_t *first_idx_dev, void *stream);

/* ---- replicated-replay exchange over NVLink peer memory ---------------------------------- */
/* The reference's buffers take part in no collective (one `memory` object serves the whole population
 * in one process: training/train_off_policy.py:46,343-345,363-376).  This exchange is the build's own
 * (BASELINE.json north_star: replicated shared replay, one process per GPU): every replica applies every
 * rank's update_priorities (components/replay_buffer.py:411-428) in rank-major order, which needs every
 * rank's (indices, priorities).  Rank r keeps its block in peer-mapped memory at peer_base[r];
 * peer_flags[r] is a
 * uint32[world] array in r's peer-mapped memory (zero before the first call).  The kernel signals
 * `epoch` to every peer, waits for every peer's `epoch`, and copies segment k of every rank's
 * block (peer_base[r] + src_off, `bytes` long) to segs[k].dst + r*bytes (local memory): a
 * rank-major all-gather.  Call with epoch = 1, 2, 3, ... on every rank, and alternate between two
 * source blocks so that a block is not rewritten while a slower peer may still read it.
 * peer_base / peer_flags / segs are HOST arrays.  A peer that does not arrive within the time-out
 * (xp_peer_set_timeout_cycles, default ~30 s) sets *err_flag_dev instead of hanging; the data of that
 * call is then NOT valid and the caller must treat the replica as lost.                          */
#define XP_MAX_PEERS 16
#define XP_MAX_PEER_SEGS 4
typedef struct xp_peer_seg {
  int64_t src_off; /* byte offset of the segment inside every rank's block */
  void *dst;       /* local device memory, world * bytes long              */
  int64_t bytes;   /* per rank                                             */
} xp_peer_seg;
xp_status_t xp_peer_allgather(const void *const *peer_base, void *const *peer_flags, int rank,
                              int world, uint32_t epoch, const xp_peer_seg *segs, int n_segs,
                              int *err_flag_dev, void *stream);
/* The same exchange with the epoch kept on the device, so that the call's parameters never change and
 * it can be captured in a CUDA graph together with the sample / gather / update launches of a step:
 * epoch_state_dev is uint32[2], zero before the first call ([0] = exchanges completed, [1] = scratch);
 * exchange number e = 1, 2, 3, ... publishes block (e - 1) & 1 (peer_base0 / peer_base1 of every rank)
 * and the last CTA of the launch stores e for the next launch to read.  Every rank must issue the same
 * sequence of calls on one stream.                                                               */
xp_status_t xp_peer_allgather_graph(const void *const *peer_base0, const void *const *peer_base1,
                                    void *const *peer_flags, int rank, int world,
                                    uint32_t *epoch_state_dev, const xp_peer_seg *segs, int n_segs,
                                    int *err_flag_dev, void *stream);
/* xp_peer_allgather_graph for a SMALL block (all segments together <= 32 KB) that the exchange kernel fills itself:
 * segment q of this rank's block is copied from pub_src_dev[q] (device memory; NULL: the caller has written it) and
 * segment pub_host_seg (or -1) from the fp32 HOST vector pub_host (<= XP_HOST_VEC_MAX values, read during the call,
 * carried inside the launch) — what ReplicatedReplay.update_priorities publishes per agent: the sampled indices on
 * the device, the learner's priorities as a host ndarray (dqn_rainbow.py:487-488).  No copy call precedes the kernel. */
xp_status_t xp_peer_publish_allgather_graph(const void *const *peer_base0, const void *const *peer_base1,
                                            void *const *peer_flags, int rank, int world,
                                            uint32_t *epoch_state_dev, const xp_peer_seg *segs, int n_segs,
                                            const void *const *pub_src_dev, int pub_host_seg,
                                            const float *pub_host, int *err_flag_dev, void *stream);
/* Bounded wait of the handshake, in SM clock cycles (default 6e10, ~30 s).                       */
void xp_peer_set_timeout_cycles(int64_t cycles);

/* Exchange + insert in one kernel: the per-step broadcast of newly collected transitions into the
 * replicated shared memory (one `memory` object for the whole population,
 * training/train_off_policy.py:46,343-345; ReplayBuffer.add, components/replay_buffer.py:72-112).
 * Every rank has written its n_rows new rows of key k at peer_base[r] + segs[k].src_off
 * (row-major, row_bytes each).  After the same signal/wait handshake as xp_peer_allgather, the
 * rank-major row q = r * n_rows + i is written to ring slot (cursor + q) % ring_rows of
 * segs[k].dst — what ReplayBuffer.add does with the concatenated batch (:100-107, wrap
 * included), without materialising the concatenation.  The caller advances its cursor by
 * world * n_rows.  Requires world * n_rows <= ring_rows.                                      */
#define XP_MAX_PEER_RING_SEGS 16
typedef struct xp_peer_ring_seg {
  int64_t src_off;   /* byte offset of this key's rows inside every rank's block */
  void *dst;         /* local ring storage of the key, ring_rows * row_bytes     */
  int64_t row_bytes;
} xp_peer_ring_seg;
xp_status_t xp_peer_ring_insert(const void *const *peer_base, void *const *peer_flags, int rank,
                                int world, uint32_t epoch, const xp_peer_ring_seg *segs,
                                int n_segs, int64_t n_rows, int64_t cursor, int64_t ring_rows,
                                int *err_flag_dev, void *stream);

/* ---- host-side RNG mirror (no GPU needed) -------------------------------------------- */
/* torch's CPU generator is mt19937; these reproduce the raw 32-bit stream so that the
 * engine consumes "identical uniform draws" (replay_buffer.py:125,377).
 * state: 624 words + index, as torch.get_rng_state() stores them.                        */
typedef struct xp_mt19937 {
  uint32_t mt[624];
  int32_t idx; /* number of words already consumed from mt[] (0..624) */
} xp_mt19937;
void xp_mt19937_seed(xp_mt19937 *g, uint32_t seed);
void xp_mt19937_draw(xp_mt19937 *g, uint32_t *out, int64_t n);
void xp_mt19937_skip(xp_mt19937 *g, int64_t n);
/* host restatement of the O(B) randperm prefix for small cases / index validation.      */
xp_status_t xp_host_randperm_prefix(const uint32_t *raw, int64_t B, int64_t n, int64_t *out);

#ifdef __cplusplus
}
#endif
#endif /* B200XP_H_ */
