/*
 * b200dp.h -- C ABI of libb200dp.so: the B200-native (sm_100a) data-parallel hot path.
 *
 * The path (BASELINE.json north_star): every training step the
 * reference scripts hand ~214 fp32 gradient tensors (Keras ResNet-50, 25,583,592
 * elements) to a cross-replica SUM and then to the optimizer.  The reference reaches
 * that through three glue layers, none of which it implements itself:
 *   - tf.distribute.MirroredStrategy()            imagenet-resnet50-mirror.py:21   (fit :81)
 *   - MultiWorkerMirroredStrategy(NCCL)           imagenet-resnet50-multiworkers.py:20-25 (fit :95)
 *   - hvd.DistributedOptimizer(Adam)              imagenet-resnet50-hvd.py:99-101  (fit :120)
 *   - ParameterServerStrategy push/pull           imagenet-resnet50-ps.py:75-82    (fit :142)
 *   - compile(optimizer='adam')                   imagenet-resnet50-mirror.py:76, imagenet-resnet50.py:62
 * This library is what a binding for those layers calls instead of
 * {concat, ncclAllReduce, split, 214 x ResourceApplyAdam}.  Each entry point below
 * cites the reference interface it stands in for.
 *
 * Conventions
 *   - plain C, no exceptions; every function returns b2dp_status (0 ok, <0 error) and
 *     b2dp_last_error() gives a thread-local message for the last failure;
 *   - tensor memory that crosses the boundary is caller-owned device memory; the library
 *     owns only its flat symmetric buffers, flag words, optimizer-state shards and tables;
 *   - work is enqueued on the caller's cudaStream_t (passed as void*) and the call returns
 *     without a host sync;
 *   - there is no CPU fallback: contexts can only be created on an sm_100 device.
 *   - "rank" r of a context is GPU r of the job.  A *local* context drives all ranks from
 *     one process (MirroredStrategy's shape); a *rank* context drives one (Horovod /
 *     MultiWorkerMirroredStrategy's shape, one process per GPU).  `dev` below is the index
 *     of the rank inside the context: 0..n-1 for a local context, always 0 for a rank one.
 *   - collective entry points (step, allreduce, all_gather, reduce_scatter_apply, broadcast,
 *     gather_state, allreduce_small) must be called for every rank, in the same order, like
 *     NCCL calls.  Unlike NCCL communicators, ALL collectives of one plan share one block of
 *     flag words and one per-rank epoch counter: on every rank they must be ordered on ONE
 *     stream (or serialised with events).  Two collectives of the same plan in flight on
 *     different streams can satisfy each other's waits and read buckets that are not packed
 *     yet.  Use separate plans for traffic that must overlap (e.g. metrics on a side stream).
 *   - a local context driven by ONE host thread: between the first and the last rank's call of a
 *     collective the host must not block on any of the devices (cudaFree, cudaDeviceSynchronize, a
 *     synchronous memcpy, b2dp_plan_destroy of another plan): the kernels already launched wait on
 *     the GPU for the ranks still to be launched.  This is what ncclGroupStart/End is for in NCCL;
 *     here the calls themselves never block, so it is enough not to put anything blocking between
 *     them (Python: `with b200dp.issue_group():`, which also keeps the garbage collector from
 *     running a destructor there).  One host thread per device has no such constraint.
 *   - if a cross-GPU flag wait gives up (a peer never reached the collective within the
 *     context's timeout), the waiting blocks leave their slice untouched, the rank's status
 *     word is set, and every later collective call of that rank returns B2DP_ERR_TIMEOUT
 *     without launching until b2dp_plan_check has reported and cleared the condition.
 */
#ifndef B200DP_H_
#define B200DP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  B2DP_OK = 0,
  B2DP_ERR_INVALID = -1,   /* bad argument / mismatched shapes (the glue raises ValueError) */
  B2DP_ERR_CUDA = -2,      /* a CUDA runtime or driver call failed */
  B2DP_ERR_NO_DEVICE = -3, /* no usable CUDA device */
  B2DP_ERR_ARCH = -4,      /* device is not sm_100 */
  B2DP_ERR_STATE = -5,     /* call made in the wrong state (e.g. peers not attached) */
  B2DP_ERR_TIMEOUT = -6,   /* a cross-GPU flag wait gave up */
  B2DP_ERR_SYS = -7,       /* socket / OS failure in the fd exchange helper */
  B2DP_ERR_TOO_LARGE = -8  /* size beyond a documented fixed maximum */
} b2dp_status;

typedef enum { B2DP_F32 = 0, B2DP_BF16 = 1 } b2dp_dtype;

typedef enum { B2DP_OPT_NONE = 0, B2DP_OPT_ADAM = 1, B2DP_OPT_SGDM = 2 } b2dp_opt;

/* Which flat buffer an accessor / unpack refers to. */
typedef enum {
  B2DP_BUF_GRADS = 0,   /* wire-format gradient bucket buffer (f32 or bf16), symmetric */
  B2DP_BUF_WEIGHTS = 1, /* fp32 mirrored weights, symmetric */
  B2DP_BUF_REDUCED = 2, /* fp32 all-reduced gradients (== GRADS when the wire is f32) */
  B2DP_BUF_STATE0 = 3,  /* this rank's shard of Adam m / SGD momentum accumulator */
  B2DP_BUF_STATE1 = 4,  /* this rank's shard of Adam v */
  B2DP_BUF_STATS = 5    /* 4 doubles, see b2dp_step */
} b2dp_buf;

/* Hyper-parameters are per-call: the scripts' callbacks mutate the learning rate while
 * training (ReduceLROnPlateau imagenet-resnet50-mirror.py:78; LearningRateWarmupCallback
 * imagenet-resnet50-hvd.py:115) and Adam's bias correction depends on the step. */
typedef struct {
  float lr;           /* Adam / SGD learning rate */
  float beta1;        /* Adam */
  float beta2;        /* Adam */
  float eps;          /* Adam (Keras default 1e-7) */
  float momentum;     /* SGD mu */
  float weight_decay; /* SGD decoupled weight decay lambda (0 = off) */
  int32_t nesterov;   /* SGD */
  int32_t want_stats; /* !=0: also produce the B2DP_BUF_STATS words for this rank's shard */
  int64_t step;       /* 1-based optimizer step (Adam bias correction) */
  float postscale;    /* multiplies the reduced sum before the optimizer (1/n for hvd Average) */
  float reserved;
} b2dp_hyper;

typedef struct b2dp_layout b2dp_layout;
typedef struct b2dp_ctx b2dp_ctx;
typedef struct b2dp_plan b2dp_plan;

const char* b2dp_last_error(void);
/* ABI version of this header; bumps when a signature changes. */
int b2dp_abi_version(void);

/* ---------------------------------------------------------------- layout (host only) --
 * The gradient bucketing layer.  Stands in for tf.distribute's _pack_tensors /
 * group_by_size(bytes_per_pack) [CommunicationOptions, imagenet-resnet50-multiworkers.py:20-21]
 * and Horovod's fusion buffer (HOROVOD_FUSION_THRESHOLD) [imagenet-resnet50-hvd.py:101].
 * K tensors in the caller's order are laid end to end, each start aligned to 32 elements;
 * a bucket closes when the next tensor would push it past bucket_bytes (0 = one bucket,
 * the reference default); each bucket is padded so it cuts into nranks equal slices of a
 * multiple of 256 elements.  Rank r owns slice r of every bucket.  Pure host arithmetic:
 * usable (and tested) without a GPU. */
int b2dp_layout_create(int32_t num_tensors, const int64_t* numel, int32_t nranks,
                       int32_t wire_dtype, int64_t bucket_bytes, b2dp_layout** out);
void b2dp_layout_destroy(b2dp_layout* l);
int32_t b2dp_layout_num_tensors(const b2dp_layout* l);
int32_t b2dp_layout_num_buckets(const b2dp_layout* l);
int32_t b2dp_layout_nranks(const b2dp_layout* l);
int64_t b2dp_layout_total_elems(const b2dp_layout* l);  /* padded flat length */
int64_t b2dp_layout_shard_elems(const b2dp_layout* l);  /* optimizer-state elements per rank */
int64_t b2dp_layout_tensor_offset(const b2dp_layout* l, int32_t k); /* flat element offset */
int32_t b2dp_layout_tensor_bucket(const b2dp_layout* l, int32_t k);
/* info[0]=first tensor, [1]=tensor count, [2]=flat offset, [3]=padded elems,
 * [4]=slice elems (per rank), [5]=offset of this bucket inside a rank's state shard */
int b2dp_layout_bucket_info(const b2dp_layout* l, int32_t b, int64_t info[6]);

/* ------------------------------------------------------------------------- context --
 * Local context: one process, ndev GPUs, peer access through one address space --
 * MirroredStrategy() (imagenet-resnet50-mirror.py:21).  Listing the same device more
 * than once gives *emulated ranks* on one GPU: the collective kernels then run without
 * cross-GPU flag waits and the caller must issue each collective call for ranks 0..n-1
 * back to back on one stream (used to check the n-rank arithmetic on a 1-GPU box). */
int b2dp_ctx_create_local(int32_t ndev, const int32_t* devices, b2dp_ctx** out);
/* Rank context: this process drives GPU `device` as rank `rank` of `nranks` --
 * hvd.init() + local_rank pinning (imagenet-resnet50-hvd.py:16,36-41) and one MWMS
 * worker (imagenet-resnet50-multiworkers.py:24-25). */
int b2dp_ctx_create_rank(int32_t rank, int32_t nranks, int32_t device, b2dp_ctx** out);
void b2dp_ctx_destroy(b2dp_ctx* c);
int32_t b2dp_ctx_nranks(const b2dp_ctx* c);
int32_t b2dp_ctx_num_local(const b2dp_ctx* c);
/* Seconds a cross-GPU flag wait may spin before the kernel gives up (default 10). */
int b2dp_ctx_set_timeout(b2dp_ctx* c, double seconds);

/* ---------------------------------------------------------------------------- plan --
 * Binds a layout to a context: allocates, per local rank, one peer-mappable block
 * holding flags | gradient buckets | mirrored weights | reduced gradients | state shards.
 * For a local context the peer mapping is done here.  For a rank context the caller
 * moves one POSIX fd per rank between processes (b2dp_fd_exchange below, or any
 * SCM_RIGHTS / pidfd_getfd transport) and attaches each peer. */
int b2dp_plan_create(b2dp_ctx* c, const b2dp_layout* l, b2dp_plan** out);
void b2dp_plan_destroy(b2dp_plan* p);
int b2dp_plan_export_fd(b2dp_plan* p, int* fd_out);            /* rank context only */
int b2dp_plan_attach_peer(b2dp_plan* p, int32_t peer_rank, int fd); /* rank context only */
int32_t b2dp_plan_ready(const b2dp_plan* p);                   /* 1 once every peer is mapped */
/* Device pointer of one of the plan's flat buffers on local rank `dev` (aliasable: a
 * harness may make its variables views of B2DP_BUF_WEIGHTS and skip unpack). */
int b2dp_plan_buffer(b2dp_plan* p, int32_t dev, int32_t which, void** ptr, int64_t* elems);
/* Reads and clears the device-side status word of rank `dev` (host sync on `stream`):
 * B2DP_OK or B2DP_ERR_TIMEOUT. */
int b2dp_plan_check(b2dp_plan* p, int32_t dev, void* stream);

/* Path B -- NVSwitch multicast (NVLS).  The default collective path (B2DP_PATH_PEER) pulls
 * slices with peer loads and sums them in rank order: bit-identical to TF's AddN.  When the
 * box has an NVSwitch with multicast, a plan can instead bind every rank's block to one
 * multicast object; b2dp_step / b2dp_allreduce then use multimem.ld_reduce (the switch
 * adds the n copies) and multimem.st (the switch replicates the store): about S bytes per
 * GPU per direction instead of 2(n-1)/n*S, at the price of the switch's own summation
 * order (parity within (n-1)*2^-24*sum|g_i|, not bitwise) -- the same trade NCCL's NVLS
 * all-reduce makes under MultiWorkerMirroredStrategy(NCCL) (-multiworkers.py:20-21).
 * With a bf16 wire the switch accumulates the n bf16 copies in fp32 and returns the sum
 * converted ONCE more to bf16 (multimem.ld_reduce .acc::f32 .bf16x2; the rounding of that
 * conversion is not documented, so one bf16 ulp is allowed): within
 * 2^-7*|sum| + (n-1)*2^-24*sum|b_i| of the bf16-wire oracle (b_i = RNE bf16 of g_i) and
 * 3*2^-8*sum|g_i| of the fp32 oracle; the weights still travel as fp32.  Local context: multicast_create does everything (fd_out = -1).
 * Rank contexts: rank 0 calls multicast_create and ships the fd to every rank; every rank
 * calls multicast_join(fd); after ALL ranks joined, every rank calls multicast_bind; after
 * ALL ranks bound, set_path(B2DP_PATH_MULTICAST). */
typedef enum { B2DP_PATH_PEER = 0, B2DP_PATH_MULTICAST = 1 } b2dp_path;
int b2dp_plan_multicast_create(b2dp_plan* p, int* fd_out);
int b2dp_plan_multicast_join(b2dp_plan* p, int fd);
int b2dp_plan_multicast_bind(b2dp_plan* p);
int b2dp_plan_set_path(b2dp_plan* p, int32_t path);
int32_t b2dp_plan_path(const b2dp_plan* p);
/* Launch-geometry knobs of the collective kernels (every rank must set the same values:
 * all ranks launch the same grid).  Defaults come from sweeps on 8xB200 (profiles/). */
typedef enum {
  B2DP_TUNE_GRID_CAP = 0,    /* peer path: max blocks (<= 1024, also bounded by co-residency) */
  B2DP_TUNE_MC_GRID_CAP = 1, /* multicast path: max blocks */
  B2DP_TUNE_MC_UNROLL = 2,   /* multicast path: vectors in flight per thread (1, 2, 4) */
  B2DP_TUNE_PEER_UNROLL = 3, /* peer path: 0 = automatic, 1 or 2 */
  B2DP_TUNE_PEER_VEC = 4,    /* peer path, f32 wire, n in {2,4,8}: 4 = 128-bit or 8 = 256-bit accesses */
  B2DP_TUNE_MC_BF16_GRID_CAP = 5, /* multicast path, bf16 wire: max blocks */
  B2DP_TUNE_SKIP_OWN = 6     /* pack_step, peer path, f32 wire: 1 = the rank's own slice is not packed but read in
                                place by the step kernel (bit-identical; measured no faster); 0 (default) = pack everything */
} b2dp_tune;
int b2dp_plan_set_tuning(b2dp_plan* p, int32_t key, int32_t value);

/* --------------------------------------------------------------------- the hot path --
 * (1) pack: gather + cast + prescale the bucket's gradient tensors into the wire buffer.
 * One launch over a device-resident (pointer, offset, count) tile table.  Replaces
 * cross_device_ops._pack_tensors/concat (MirroredStrategy, -mirror.py:21) and Horovod's
 * batched_memcpy_k + prescale into the fusion buffer (-hvd.py:101).  `grads` is a HOST
 * array of K device pointers (all tensors of the plan, index = tensor id; only the
 * bucket's are read); bucket < 0 packs every bucket. */
int b2dp_pack(b2dp_plan* p, int32_t dev, int32_t bucket, const void* const* grads,
              int32_t grad_dtype, float prescale, void* stream);

/* (2)+(3) step: reduce-scatter(SUM) the bucket across ranks -- rank r pulls slice r from
 * every rank's wire buffer and adds in rank order 0..n-1 in fp32, which is TF's
 * ReductionToOneDevice/AddN order -- then applies the optimizer to rank r's shard of
 * (w, state) in registers, and all-gathers by storing the updated weights into every
 * rank's weight buffer.  One kernel; flag barriers at entry and exit.  Replaces
 * batch_reduce(SUM) + 214 ResourceApplyAdam launches (-mirror.py:76,81), the
 * CollectiveReduceV2 path (-multiworkers.py:20-25,95) and, with slices as PS shards,
 * the PS push/apply/pull (-ps.py:75-82,142).  With n == 1 it is the optimizer apply
 * alone (imagenet-resnet50.py:62,67).  opt == B2DP_OPT_NONE is rejected (use allreduce).
 * hyper->want_stats: B2DP_BUF_STATS[0] = sum of squares of the reduced gradient over
 * this rank's shard of the buckets THIS CALL covers (bucket < 0: all of them, accumulated
 * across the per-bucket launches), [1] = count of non-finite reduced-gradient elements. */
int b2dp_step(b2dp_plan* p, int32_t dev, int32_t bucket, int32_t opt, const b2dp_hyper* hyper,
              void* stream);

/* pack + step as one call -- what a binding normally issues once per training step.  With
 * one rank and an f32 wire (imagenet-resnet50.py:62,67: no reduction, 214 x ResourceApplyAdam)
 * it is a single kernel that reads each gradient in place through the tile table and never
 * materialises the bucket (28 instead of 36 bytes of HBM traffic per element); with n > 1 it
 * is b2dp_pack followed by b2dp_step -- except that with several buckets bucket b+1 is packed on an
 * internal stream under bucket b's step, and that with B2DP_TUNE_SKIP_OWN (peer path, f32 wire,
 * 2/4/8 ranks) the rank's own slice of each bucket is left unpacked and read in place by the step
 * kernel.  The gradient tensors must stay untouched until the call's work has run.
 * Results are bit-identical to the two-call form. */
int b2dp_pack_step(b2dp_plan* p, int32_t dev, int32_t bucket, const void* const* grads,
                   int32_t grad_dtype, float prescale, int32_t opt, const b2dp_hyper* hyper,
                   void* stream);

/* hvd-compatible allreduce: same reduce-scatter + all-gather with no optimizer; the
 * result (sum * postscale, fp32) lands in B2DP_BUF_REDUCED of every rank.  Replaces
 * hvd.allreduce(op=Average|Sum) under hvd.DistributedOptimizer (-hvd.py:101) and
 * CrossDeviceOps.batch_reduce_implementation for non-variable destinations. */
int b2dp_allreduce(b2dp_plan* p, int32_t dev, int32_t bucket, float postscale, void* stream);

/* The two halves of step as separate launches (PS push = reduce_scatter_apply, pull =
 * all_gather; -ps.py:75-82).  all_gather pushes this rank's weight slice to all peers. */
int b2dp_reduce_scatter_apply(b2dp_plan* p, int32_t dev, int32_t bucket, int32_t opt,
                              const b2dp_hyper* hyper, void* stream);
int b2dp_all_gather(b2dp_plan* p, int32_t dev, int32_t bucket, void* stream);

/* unpack: scatter a flat buffer (B2DP_BUF_WEIGHTS or B2DP_BUF_REDUCED) back into the
 * caller's K tensors (fp32).  Replaces array_ops.split after the all-reduce
 * (-mirror.py:81) and Horovod's memcpy-out of the fusion buffer (-hvd.py:101).
 * pack's inverse for initial weights: b2dp_load_weights gathers K fp32 tensors into
 * B2DP_BUF_WEIGHTS (hvd BroadcastGlobalVariables-style initial sync is a step-0 concern
 * of the glue, -hvd.py:111). */
int b2dp_unpack(b2dp_plan* p, int32_t dev, int32_t bucket, int32_t which, void* const* outs,
                void* stream);
int b2dp_load_weights(b2dp_plan* p, int32_t dev, const void* const* weights, void* stream);

/* Step-0 and per-epoch helpers of the Horovod surface (collective calls, like step):
 * broadcast -- the root rank's flat buffer (B2DP_BUF_WEIGHTS or B2DP_BUF_REDUCED) is stored
 * into every rank's copy: hvd.callbacks.BroadcastGlobalVariablesCallback(0)
 * (imagenet-resnet50-hvd.py:111).
 * allreduce_small -- latency-bound one-shot all-reduce of count <= 1024 fp32 values in
 * caller-owned device memory (out = rank-order sum * postscale): metric averaging,
 * hvd.callbacks.MetricAverageCallback (imagenet-resnet50-hvd.py:113).  More than 1024 values
 * returns B2DP_ERR_TOO_LARGE (use b2dp_allreduce). */
int b2dp_broadcast(b2dp_plan* p, int32_t dev, int32_t which, int32_t root, void* stream);
/* gather_state -- collective: every rank's shard of B2DP_BUF_STATE0 / STATE1 (Adam m / v, SGD
 * accumulator) is assembled into B2DP_BUF_REDUCED of every rank in flat tensor order; follow
 * with b2dp_unpack(B2DP_BUF_REDUCED) to get per-variable tensors.  What the checkpoint at the
 * end of every script, model.save(...) (imagenet-resnet50-mirror.py:83-86), needs from an
 * optimizer whose state is sharded across ranks.  Overwrites B2DP_BUF_REDUCED (for an f32
 * wire that is the gradient bucket: call it between steps). */
int b2dp_gather_state(b2dp_plan* p, int32_t dev, int32_t which, void* stream);
int b2dp_allreduce_small(b2dp_plan* p, int32_t dev, const float* in, float* out, int32_t count,
                         float postscale, void* stream);

/* Scalar the Adam kernel multiplies by: lr*sqrt(1-beta2^t)/(1-beta1^t) in fp32, exposed
 * so a checker can use the library's exact value. */
float b2dp_adam_alpha(float lr, float beta1, float beta2, int64_t step);

/* ------------------------------------------------- entry point that allocates nothing --
 * No counterpart in the reference: TensorFlow's ResourceApplyAdam takes its temporaries
 * from the framework allocator.  This one is for callers that may not allocate device
 * memory at all -- a framework op whose memory must all come from the framework's
 * allocator (a TF / Horovod custom-op Compute()), a latency-critical loop, a process with
 * a fixed memory budget.  It is the n == 1 fused Adam apply of b2dp_step on CALLER-OWNED
 * arrays: w, m, v are updated in place from g (all fp32 device pointers on the current
 * device, 16-byte aligned, numel elements; same one-rounding-per-operation arithmetic, so
 * w, m, v come out bit-identical to b2dp_step), and stats_out (2 doubles, device memory)
 * receives [sum of squares of the scaled gradient, count of non-finite gradient elements].
 * The scratch the statistics reduction needs is a static __device__ array inside the
 * library (128.25 KB on every GPU that loads it, held until exit) instead of an
 * allocation, which fixes the largest call: numel > B2DP_NOALLOC_MAX_ELEMS returns
 * B2DP_ERR_TOO_LARGE.  No cudaMalloc/cudaFree, no plan, no context.
 *   - The first call on each device resolves the scratch array's device address
 *     (cudaGetSymbolAddress) and the launch geometry; make that first call outside any
 *     stream capture.
 *   - Every call on a device shares the one scratch array.  Calls ordered on one stream
 *     (or otherwise serialised) are fine; calls that overlap on the GPU race on it.
 *   - hyper: lr, beta1, beta2, eps, step (>= 1), postscale are used; the rest ignored. */
#define B2DP_NOALLOC_MAX_ELEMS 67108864LL /* 2^26 elements per call */
int b2dp_adam_apply_noalloc(float* w, float* m, float* v, const float* g, int64_t numel,
                            const b2dp_hyper* hyper, double* stats_out, void* stream);

/* ----------------------------------------------------------- rank bootstrap helper --
 * Moves one fd per rank between the processes of a job over abstract unix datagram
 * sockets (SCM_RIGHTS).  `session` must be the same string on every rank and unique to
 * the job (e.g. MASTER_PORT + a counter).  peer_fds[r] receives rank r's fd
 * (peer_fds[rank] = dup of my_fd).  Host-only; works for any fd. */
int b2dp_fd_exchange(int32_t rank, int32_t nranks, const char* session, int my_fd,
                     int* peer_fds, int32_t timeout_ms);

#ifdef __cplusplus
}
#endif
#endif /* B200DP_H_ */
