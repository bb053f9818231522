/* C ABI of libaar_cuda.so — the sm_100a data plane of akka_allreduce_b200.
 *
 * Everything here is callable from C, ctypes or a torch extension.  Functions return 0 on
 * success or a negative AAR_E* code; aar_last_error() gives the text of the last failure on
 * the calling thread.  Nothing in this header allocates device memory unless it says so.
 */
#ifndef AAR_B200_H_
#define AAR_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AAR_MAX_RANKS 16

enum { AAR_OK = 0, AAR_EINVAL = -1, AAR_ECUDA = -2, AAR_EUNSUPPORTED = -3, AAR_ETOOBIG = -4 };
enum { AAR_F32 = 0, AAR_BF16 = 1 };
/* AAR_ALGO_DIRECT: the one-shot idea for LARGE messages on two ranks (stream kernel).  With P = 2 a two-shot round and
 * a one-shot round put the same number of bytes on the link (N each way), but the one-shot form has one hop instead
 * of two: every rank streams its whole message into the peer's one-shot row chunk by chunk while the other warps sum
 * "mine + the peer's" for every chunk that has landed, straight into the output.  Needs thresholds of 1, world = 2,
 * a region laid out with os_cap >= n. */
enum { AAR_ALGO_TWOSHOT = 0, AAR_ALGO_ONESHOT = 1, AAR_ALGO_NVLS = 2, AAR_ALGO_DIRECT = 3 };
/* The round is a reduce-scatter (AllreduceWorker.scala:212-238 scatter + ScatteredDataBuffer.scala:20-32 reduce)
 * followed by an all-gather (AllreduceWorker.scala:252-268 broadcast + ReducedDataBuffer.scala:26-53 flush); each half
 * can be run on its own (stream kernel only):
 *   AAR_OP_REDUCE_SCATTER  src = the whole message, out / counts = THIS rank's block only (block-sized buffers)
 *   AAR_OP_ALL_GATHER      src = THIS rank's block, out / counts = the whole message (count 1 where a block arrived) */
enum { AAR_OP_ALLREDUCE = 0, AAR_OP_REDUCE_SCATTER = 1, AAR_OP_ALL_GATHER = 2 };

/* One fused scatter -> reduce -> broadcast round (reference hot path: AllreduceWorker.scala:
 * 212-238 scatter, ScatteredDataBuffer.scala:20-32 reduce, AllreduceWorker.scala:252-268
 * broadcast, ReducedDataBuffer.scala:26-53 flush).  In emulation mode (`emu` != 0) one
 * cooperative launch plays all `world` ranks on one GPU: rank = blockIdx.y, and src/out/
 * counts/sym/scratch are indexed by rank.  In real mode only index 0 of src/out/counts/
 * scratch is used and sym[] holds the peer-mapped bases of every rank's region. */
typedef struct AarLaunch {
  int32_t rank, world, emu;
  int32_t algo;             /* AAR_ALGO_* */
  int32_t op;               /* AAR_OP_*: the whole round, or one of its halves */
  int32_t in_dtype;         /* dtype of src and out: AAR_F32 | AAR_BF16 */
  int32_t wire_dtype;       /* dtype of the symmetric slots */
  int32_t ctas;             /* work CTAs per rank (one control CTA is added); clamped to aar_max_ctas() */
  int32_t use_tma;          /* stage tiles with cp.async.bulk when alignment allows */
  int32_t fast_fences;      /* experiments only: gpu-scope instead of system-scope fences before flags */
  int32_t role_s, role_c;   /* generic kernel: work CTAs dedicated to the scatter / complete phases; stream kernel:
                               warps (of the 16 per CTA) that scatter / complete; 0 = pick automatically */
  int32_t nvls_unroll;      /* multimem.ld_reduce requests in flight per thread: 0/8 = 8, or 16 */
  int32_t part_kb;          /* largest work item in KB of wire data; 0 = 256 (generic), 32 (stream: reduce tiles) */
  int32_t part_sc_kb;       /* stream kernel: largest scatter / completion tile in KB; 0 = 8 */
  int32_t overlap_flush;    /* 1: flush already-delivered chunks between reduce items (overlaps phases R and C) */
  int32_t oneshot_pull;     /* one-shot only. 0: push copies to every peer; 1: stage locally, peers pull (in-switch
                               multimem.ld_reduce when mc_base is set); 2: stage locally, unicast pulls only */
  int32_t kernel;           /* two-shot / NVLS: 0 = pick (the stream kernel whenever pointers, block step and chunk
                               size are 16-byte aligned and src dtype == wire dtype), 1 = generic kernel, 2 = stream kernel */
  int32_t self_publish;     /* stream kernel: who raises the per-chunk flags.  1 = the publisher warp of the control CTA
                               (scans part counters, one fence per batch), 2 = the data warp that completes a chunk's
                               last part (no scan latency, one fence per chunk), 0 = pick */
  int32_t push_chunks;      /* NVLS + stream kernel: the last `push_chunks` chunks of every block take the unicast push
                               path instead of the in-switch reduction (a hybrid round); 0 = all through the switch */
  int32_t mc_bcast;         /* two-shot: 1 = broadcast reduced tiles with one multimem.st through mc_base instead of
                               P unicast stores (whenever no destination is skipped); NVLS always does */
  int64_t n, max_chunk;     /* dataSize, maxChunkSize (elements) */
  int64_t os_cap;           /* one-shot capacity the region was laid out with */
  int64_t block_step;       /* 0 = reference partition ceil(n/world); else an aligned step >= that */
  int64_t n_cap;            /* 0, or the LARGER message length the region and the scratch were laid out for: this
                               round then moves n <= n_cap elements through the same region (strides and offsets
                               from the capacity, partition from n) — how one heap serves buckets of several sizes */
  int64_t block_step_cap;   /* with n_cap: the block step of the capacity layout (block_step must be set too) */
  int32_t L, slot;          /* ring slots (maxLag+1), round % L */
  uint32_t tag;             /* round + 1 */
  int32_t req_reduce;       /* int(thReduce * P) */
  int32_t req_complete;     /* int(thComplete * totalChunks) */
  int32_t gate_count;       /* peers that must have finished round-1 before scattering; 0 = no gate */
  int32_t normalize;        /* divide every chunk by its contributor count */
  float scale;              /* source scale fused into the scatter */
  uint32_t skip_dest_mask[AAR_MAX_RANKS]; /* straggler injection: rank r does not scatter to these dests */
  uint32_t mute_mask;       /* straggler injection: these ranks do not scatter at all this round */
  uint64_t delay_ns[AAR_MAX_RANKS];       /* straggler injection: rank r idles this long first */
  uint64_t timeout_ns;      /* bound on every flag wait; on expiry the round is force-completed */
  const void* src[AAR_MAX_RANKS];
  void* out[AAR_MAX_RANKS];          /* dataSize elements of in_dtype */
  int32_t* counts[AAR_MAX_RANKS];    /* dataSize int32 contributor counts, or NULL */
  uint8_t* sym[AAR_MAX_RANKS];       /* base of every rank's symmetric region */
  uint8_t* mc_base;                  /* NVLS multicast alias of the region, or NULL */
  uint8_t* scratch[AAR_MAX_RANKS];   /* local scratch, aar_scratch_bytes(), zeroed once */
  volatile uint32_t* host_flags;     /* mapped pinned, written by the kernel only: [1] done tag, [2] tag of
                                        the last round closed by a timeout; 4 words per rank in emulation
                                        mode; may be NULL.  (The force-complete tag is a DEVICE word:
                                        aar_scratch_force_offset() bytes into each rank's scratch.) */
  uint64_t* trace;                   /* optional device buffer [ranks][ctas+1][16] of globaltimer stamps
                                        (start, gate passed, S done, R done, C done, ns waiting in R,
                                        ns waiting in C, ns publishing in S); NULL = no tracing */
} AarLaunch;

/* Sizes of the symmetric region / local scratch for a configuration. */
int64_t aar_sym_bytes(int64_t n, int64_t max_chunk, int world, int L, int wire_dtype, int64_t os_cap,
                      int64_t block_step);
int64_t aar_sym_flag_bytes(int64_t n, int64_t max_chunk, int world, int L, int wire_dtype, int64_t os_cap,
                           int64_t block_step);
/* Byte offsets of the staging (scat) and sink slot arrays inside the region; slot s, row r starts
 * (s*world + r)*block_pad elements further (block_pad = block step rounded up to 128 elements). */
int64_t aar_sym_scat_offset(int64_t n, int64_t max_chunk, int world, int L, int wire_dtype, int64_t os_cap,
                            int64_t block_step);
int64_t aar_sym_sink_offset(int64_t n, int64_t max_chunk, int world, int L, int wire_dtype, int64_t os_cap,
                            int64_t block_step);
int64_t aar_scratch_bytes(int64_t n, int64_t max_chunk, int world, int64_t block_step);
/* Byte offset, inside a rank's scratch, of the u32 force tag: rounds with tag <= that value close
 * every wait immediately (device form of the catch-up path, AllreduceWorker.scala:100-106). */
int64_t aar_scratch_force_offset(int64_t n, int64_t max_chunk, int world, int64_t block_step);

/* Byte offset, inside a rank's scratch, of the u32 bit mask of peers known dead (bit q = rank q).  The host
 * sets it when the coordinator reports a worker gone (AllreduceMaster.scala:46-52 drops it from the map,
 * AllreduceWorker.scala:213-216 then skips it): the kernels neither push to nor wait for those ranks. */
int64_t aar_scratch_dead_offset(int64_t n, int64_t max_chunk, int world, int64_t block_step);

/* sizeof(AarLaunch) as this library was compiled: bindings compare it with their own mirror of the struct. */
int aar_abi_sizeof_launch(void);

/* Launches one round on `stream` (a cudaStream_t).  Returns the number of kernels launched
 * (>0) or a negative error. */
int aar_allreduce_round(const AarLaunch* launch, void* stream);

/* Same round as aar_allreduce_round(), for callers that cannot allocate device memory.  (No counterpart
 * in the reference: it has no device memory to manage.)
 *
 * Who that is: a process that handed the whole GPU to its own memory pool when it started (an
 * inference server that reserves HBM up front leaves nothing for a later cudaMalloc); a control loop
 * that must not stall (cudaMalloc synchronises the device); a host that embeds this library through
 * the C ABI alone and has no allocator of its own to give it buffers.  For them the kernel's scratch —
 * the tile counters and per-chunk decisions that aar_allreduce_round() takes from launch->scratch[] —
 * lives in the library image: a fixed array of 16 KB per rank (16 ranks: 256 KB) in the global memory
 * of every GPU that loads the library, held until the process exits.  launch->scratch[] is ignored;
 * inputs, outputs and the symmetric region are still the caller's.
 *
 * - Limit: aar_scratch_bytes() of the configuration must be <= 16384 (roughly world x chunks-per-block
 *   <= 3000); beyond that the call fails with AAR_ETOOBIG and a message naming both numbers.
 * - It calls no cudaMalloc/cudaFree and no allocating library routine; the array is cleared with a
 *   cudaMemsetAsync on `stream` before each launch, so nothing carries over from the previous call
 *   (this also means the host-written words inside the scratch — the force tag and the dead-peer mask — are
 *   always 0 here: rounds close by threshold, skip notice or timeout).
 * - The first call on each device looks the array's device address up (that device must be current);
 *   make that first call before you start capturing any stream.
 * - All calls in the process share the one array per device.  Two rounds that overlap on the GPU
 *   (different streams, or two engines) would corrupt each other's counters: serialise them. */
int aar_allreduce_round_static_scratch(const AarLaunch* launch, void* stream);

/* Co-residency limit: the kernels wait on flags written by their own control CTA and by their peers, so
 * every CTA of a launch must be resident.  Largest `ctas` for which (ctas+1) x world_on_device CTAs fit on
 * `device`; kernel: 1 = generic (one CTA per SM), 2 = stream (two per SM).  aar_allreduce_round() refuses
 * launches beyond it.  aar_max_emu_ctas(world, device) = aar_max_ctas(1, world, device). */
int aar_max_ctas(int kernel, int world_on_device, int device);
int aar_max_emu_ctas(int world, int device);
/* Which kernel aar_allreduce_round() would run for this launch: 1 generic, 2 stream, 3 one-shot; < 0 = error. */
int aar_round_kernel(const AarLaunch* launch);
/* Work CTAs that launch would really use: launch->ctas clamped to aar_max_ctas() of its kernel. */
int aar_round_ctas(const AarLaunch* launch);

/* Symmetric heap (csrc/cuda/sym_heap.cu): CUDA VMM allocation shared with every peer process
 * through POSIX file descriptors passed over unix sockets, plus an NVLS multicast alias when
 * the driver supports it.  `uds_dir` must be a directory all ranks can reach. */
typedef struct AarHeap AarHeap;
int aar_heap_create(AarHeap** out, int rank, int world, int device, int64_t bytes, const char* uds_dir,
                    int want_multicast);
int aar_heap_ptrs(AarHeap* h, uint8_t** peer_bases /* [world] */, uint8_t** mc_base);
int64_t aar_heap_bytes(AarHeap* h);
int aar_heap_destroy(AarHeap* h);

const char* aar_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* AAR_B200_H_ */
