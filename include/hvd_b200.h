/* hvd_b200.h - C ABI of the native libraries (bound from Python with ctypes; usable from C/C++).
 *
 *   libhvd_core.so  host runtime   (g++; CUDA driver resolved with dlopen, loads on CPU-only hosts)
 *   libhvd_cuda.so  sm_100a kernels (nvcc -gencode arch=compute_100a,code=sm_100a)
 *
 * Every function returns 0 on success unless noted.  Device pointers are raw addresses; `stream`
 * is a cudaStream_t.  The reference (weikaolun/horovod-distributed-example) has no native code of
 * its own: each block cites the script line / Horovod component whose capability it provides.
 */
#ifndef HVD_B200_H_
#define HVD_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ libhvd_core.so */

/* Static bucket plan - replaces Horovod's per-cycle tensor fusion (controller.cc, fusion_buffer_manager.cc)
 * behind hvd.DistributedOptimizer (reference mnist_keras.py:87).  out: 4 int64 per bucket
 * (start, end, first_tensor, last_tensor).  Returns the bucket count. */
int64_t hvd_plan_buckets(const int64_t* offsets, const int64_t* numels, int64_t n, int64_t slab_elems,
                         int64_t threshold_bytes, const int64_t* cuts, int64_t n_cuts, int64_t* out);
uint64_t hvd_fnv1a(uint64_t h, const uint8_t* data, int64_t n);
uint32_t hvd_crc32c(const uint8_t* p, int64_t n);

/* TF event files - what keras.callbacks.TensorBoard writes (reference mnist_keras.py:105). */
void* hvd_ev_open(const char* path);
int hvd_ev_scalar(void* h, const char* tag, double value, int64_t step, double wall_time);
void hvd_ev_flush(void* h);
void hvd_ev_close(void* h);

/* Chrome-trace timeline - HOROVOD_TIMELINE (Horovod timeline.cc). ph: 'B','E','i','X'. */
void* hvd_tl_open(const char* path);
int hvd_tl_event(void* h, const char* tensor, const char* activity, char ph, int64_t ts_us, int64_t dur_us);
int64_t hvd_tl_now_us(void* h);
void hvd_tl_close(void* h);

/* Levelled logging (Horovod logging.cc; HVD_LOG_LEVEL / HOROVOD_LOG_LEVEL read on first use) and NVTX ranges
 * (header-only NVTX3: free unless a profiler is attached). */
enum { HVD_LOG_TRACE = 0, HVD_LOG_DEBUG = 1, HVD_LOG_INFO = 2, HVD_LOG_WARNING = 3, HVD_LOG_ERROR = 4, HVD_LOG_FATAL = 5 };
void hvd_log_set_level(int lvl);
int hvd_log_level(void);
void hvd_log_set_rank(int rank);
int hvd_log_parse_level(const char* s);
void hvd_log(int lvl, const char* msg);
int hvd_nvtx_push(const char* name);
int hvd_nvtx_pop(void);
void hvd_nvtx_mark(const char* name);

/* Minimal HDF5 writer/reader - the checkpoint-{epoch}.h5 files of ModelCheckpoint / model.save /
 * load_model (reference mnist_keras.py:101-104,118-119,124).  dtype: 0 f32, 1 i64, 2 f64, 3 i32, 4 string. */
const char* hvd_h5_last_error(void);
void* hvd_h5w_create(void);
void hvd_h5w_free(void* w);
int hvd_h5w_group(void* w, const char* path);
int hvd_h5w_dataset(void* w, const char* path, int dtype, int rank, const int64_t* dims, const void* data);
int hvd_h5w_attr_str(void* w, const char* obj, const char* name, const char* data, int64_t len);
int hvd_h5w_attr_strs(void* w, const char* obj, const char* name, const char* packed, int64_t n, int64_t width);
int hvd_h5w_attr_num(void* w, const char* obj, const char* name, int dtype, int rank, const int64_t* dims, const void* data);
int hvd_h5w_write(void* w, const char* filename);
void* hvd_h5r_open(const char* filename);
void hvd_h5r_close(void* r);
int hvd_h5r_kind(void* r, const char* path); /* 0 missing, 1 group, 2 dataset */
int64_t hvd_h5r_list(void* r, const char* path, char* out, int64_t cap);
int hvd_h5r_dataset_info(void* r, const char* path, int* dtype, int* rank, int64_t* dims, int64_t* elsize);
int hvd_h5r_dataset_read(void* r, const char* path, void* out, int64_t nbytes);
int64_t hvd_h5r_attr_list(void* r, const char* obj, char* out, int64_t cap);
int hvd_h5r_attr_info(void* r, const char* obj, const char* name, int* dtype, int* rank, int64_t* dims, int64_t* elsize);
int hvd_h5r_attr_read(void* r, const char* obj, const char* name, void* out, int64_t nbytes);

/* Symmetric (peer-mapped) memory on the CUDA VMM driver API - the exportable replacement of Horovod's
 * fusion buffer; communicator setup behind hvd.init() (reference mnist_keras.py:30). */
const char* hvd_symm_last_error(void);
int hvd_symm_available(void);
int hvd_symm_multicast_supported(int device);
int hvd_symm_handle_type_supported(int device);
long long hvd_symm_round_size(int device, long long bytes, int nranks_for_multicast);
int hvd_symm_create(int device, long long bytes, unsigned long long* handle, int* fd);
int hvd_symm_import(int fd, unsigned long long* handle);
int hvd_symm_map(int device, unsigned long long handle, long long bytes, unsigned long long* ptr);
int hvd_symm_unmap(unsigned long long ptr, long long bytes);
int hvd_symm_release(unsigned long long handle);
int hvd_symm_mc_create(int ndev, long long bytes, unsigned long long* mc, int* fd);
int hvd_symm_mc_add_device(unsigned long long mc, int device);
int hvd_symm_mc_bind(unsigned long long mc, unsigned long long mem, long long bytes);

/* Process-level runtime state (csrc/core/runtime.cpp) - Horovod's C API horovod_init / rank / size / local_rank /
 * shutdown behind hvd.init(), hvd.rank(), hvd.size(), hvd.local_rank() (reference mnist_keras.py:30,35,42).
 * hvd_core_init reads the launcher's environment (HVD_*, OMPI_COMM_WORLD_*, PMI_*, torchrun; none = a world of one);
 * the getters return -1 before it.  The registry records where every rank's copy of each symmetric buffer is mapped
 * in this process (what the collective kernels address); hvd_registry_find maps a local address back to its buffer. */
int hvd_core_discover_env(int* out4, const char** contract);
int hvd_core_init(void);
int hvd_core_is_initialized(void);
int hvd_core_rank(void);
int hvd_core_size(void);
int hvd_core_local_rank(void);
int hvd_core_local_size(void);
const char* hvd_core_launcher(void);
void hvd_core_shutdown(void);
int hvd_registry_add(const uint64_t* ptrs, int n, uint64_t bytes, uint64_t mc_ptr, int local_index);
int hvd_registry_find(uint64_t p, uint64_t* byte_offset);   /* newest matching entry */
int hvd_registry_remove(int id);
uint64_t hvd_registry_peer_ptr(int id, int r, uint64_t* mc_ptr, uint64_t* bytes);
int hvd_registry_count(void);
void hvd_registry_clear(void);

/* Bootstrap channel (csrc/core/bootstrap.cpp): fixed-layout messages + POSIX fd passing (SCM_RIGHTS) between the
 * ranks of one host - the start-up handle exchange behind hvd.init() (reference mnist_keras.py:30; Horovod:
 * MPI_Bcast of the ncclUniqueId, message.cc / wire/message.fbs).  kind: 1 = memory handle, 2 = multicast handle. */
typedef struct HvdBootMsg {
  uint32_t magic;    /* 'HVDB' */
  uint16_t version;
  uint16_t kind;
  int32_t src_rank;
  int32_t tag;       /* ordinal of the symmetric allocation the handle belongs to */
  uint64_t payload;  /* size in bytes of that allocation (consistency check) */
} HvdBootMsg;
void* hvd_boot_open(const char* dir, int rank, int size, double timeout_s);
int hvd_boot_send(void* h, int dst, int kind, int tag, uint64_t payload, int fd /* -1: none */);
int hvd_boot_recv(void* h, HvdBootMsg* out, int* fd);
void hvd_boot_close(void* h);
const char* hvd_boot_last_error(void);

/* ------------------------------------------------------------------ libhvd_cuda.so */

struct HvdCommCtx;   /* csrc/comm/comm.cuh  */
struct HvdFusedArgs; /* csrc/comm/fused_allreduce.cu */

/* Fused allreduce + 1/N + optimizer update over peer memory - the data path of
 * hvd.DistributedOptimizer / MetricAverageCallback (reference mnist_keras.py:87,
 * tensorflow2_keras_mnist.py:77).  emulate_ranks > 0: one cooperative launch, blockIdx.y = rank. */
int hvd_fused_allreduce_launch(const struct HvdCommCtx* c, const struct HvdFusedArgs* a, int ctas, int threads,
                               void* stream, int emulate_ranks);
/* Broadcast a range from root's arena - BroadcastGlobalVariablesCallback (reference mnist_keras.py:97). */
int hvd_bcast_launch(const struct HvdCommCtx* c, int slot, int root, long long off, long long n, int ctas, int threads,
                     void* stream, int emulate_ranks);
int hvd_pad_bytes(void);
int hvd_epoch_bytes(void);
int hvd_cuda_malloc(void** p, long long bytes);
int hvd_cuda_free(void* p);
int hvd_ipc_get_handle(void* p, void* out64);
int hvd_ipc_open(const void* h64, void** p);
int hvd_ipc_close(void* p);
int hvd_can_access_peer(int dev, int peer);
int hvd_memcpy_d2d(void* dst, const void* src, long long bytes, void* stream);
int hvd_memcpy_async(void* dst, const void* src, long long bytes, void* stream);   /* any direction (cudaMemcpyDefault) */
int hvd_host_is_pinned(const void* p);
const char* hvd_cuda_error_string(int e);

/* tcgen05 TF32 GEMM, D[M,N] = A*B.  A: [M][K] (a_mn=0) or [K][M] (a_mn=1); B: [N][K] or [K][N].
 * epilogue: 0 store, 1 atomic add (split-K), 2 bias+relu, 3 maxpool/dropout/relu-backward scatter.
 * The Dense(128) layer's forward / dgrad / wgrad (reference mnist_keras.py:79). */
int hvd_tc_gemm(const float* A, const float* B, float* D, int M, int N, int K, int a_mn, int b_mn, int bn, int k_split,
                int epilogue, const float* bias, long long ldd, void* stream, unsigned lbo_a, unsigned sbo_a,
                unsigned lbo_b, unsigned sbo_b, const uint8_t* mask, float scale);

/* Same GEMM with the two hooks of the pipelined data-parallel step (csrc/comm/pipeline.cuh; either may be NULL):
 * rs   - epilogue 4 (EPI_RS_PUSH): every 128x128 output tile is written into the receive area of the rank that
 *        owns those rows, local memory or a peer GPU over NVLink (D may be NULL) - the Dense(128) weight gradient's
 *        reduce-scatter fused into the GEMM that produces it (hvd.DistributedOptimizer, reference mnist_keras.py:87);
 * wait - the producer warp first waits for the version flags of the weights it is about to read. */
struct HvdRsPush;
struct HvdParamWait;
struct HvdTailArgs;
struct HvdAllgatherArgs;
struct HvdLLArgs;
int hvd_tc_gemm_ex(const float* A, const float* B, float* D, int M, int N, int K, int a_mn, int b_mn, int bn, int k_split,
                   int epilogue, const float* bias, long long ldd, void* stream, unsigned lbo_a, unsigned sbo_a,
                   unsigned lbo_b, unsigned sbo_b, const uint8_t* mask, float scale, const struct HvdRsPush* rs,
                   const struct HvdParamWait* wait);

/* Pipelined data-parallel step (csrc/comm/pipeline_tail.cu) - what hvd.DistributedOptimizer's allreduce + 1/N +
 * optimizer update (reference mnist_keras.py:84-87, tensorflow2_keras_mnist.py:55-58) becomes on the fused engine.
 * hvd_rs_flags_launch: raise the "tile landed" flags of the EPI_RS_PUSH GEMM just enqueued on `stream`.
 * hvd_tail_launch: LL one-shot allreduce + update of the small tensors and the owner's reduce + update of its rows.
 * hvd_allgather_launch: deliver the owned rows to every replica and stamp version flags (use_tma: TMA bulk copies).
 * hvd_param_wait_launch: make `stream` wait until every peer's rows of the current version have landed.
 * emulate_ranks > 0: all ranks on one GPU (tests); the tail then runs as one cooperative launch. */
int hvd_rs_flags_launch(const struct HvdRsPush* rs, void* stream);
int hvd_tail_launch(const struct HvdCommCtx* c, const struct HvdTailArgs* a, int ctas, void* stream, int emulate_ranks);
int hvd_allgather_launch(const struct HvdCommCtx* c, const struct HvdAllgatherArgs* a, int ctas, void* stream,
                         int emulate_ranks, int use_tma);
int hvd_param_wait_launch(const struct HvdParamWait* w, void* stream);
int hvd_globaltimer_probe(unsigned long long* out, void* stream);   /* one %globaltimer reading (HVD_TIMELINE calibration) */
/* Barrier-free LL one-shot allreduce of a small message (<= 128 KB by default): flagged 16-byte packets pushed to
 * every peer, no barrier.  The packed metric vector of MetricAverageCallback (reference tensorflow2_keras_mnist.py:77)
 * and any small tensor-level hvd.allreduce. */
int hvd_ll_allreduce_launch(const struct HvdCommCtx* c, const struct HvdLLArgs* a, void* stream, int emulate_ranks);

/* Dense(128) forward + classifier head in ONE launch (reference layers mnist_keras.py:79-81, loss / metric :89-91):
 * h_pre[B][128] += x[B][K] * W[K][128] over k_split slices of K; the slice CTAs then wait for each other on `counter`
 * (device uint64, starts at 0, never reset) and the first ceil(B/8) of them do hvd_head's work on 8 rows each.
 * Needs B <= 128, k_split * 8 >= B, k_split <= 148; h_pre zero on entry (and zero again on exit). */
int hvd_dense1_fwd_head(const float* x, const float* W, int B, int K, int k_split, float* h_pre, const float* b1,
                        const float* w2, const float* b2, const long long* labels, float* dh_pre, float* g_w2,
                        float* g_b2, float* g_b1, float* probs, void* step_state, int training, int backward, float keep,
                        unsigned long long* counter, const struct HvdParamWait* wait, void* stream);

/* Fused MNIST-CNN engine kernels (reference model mnist_keras.py:71-81). */
/* conv1 forward; also the step prologue: clears [zero_a, +n_a) and [zero_b, +n_b) (the gradient ranges
 * later kernels red.add into; multiples of 4 floats, may be NULL), clears the per-step loss/accuracy words
 * of step_state and, if advance_step, bumps its dropout step counter. */
int hvd_conv1_fwd(const float* x, const float* w, const float* bias, float* a1, int B, float* zero_a, int n_a,
                  float* zero_b, int n_b, void* step_state, int advance_step, void* stream);
int hvd_conv2_fprop(const float* a1, const float* w2, const float* bias2, float* p2, uint8_t* mask,
                    const void* step_state, int B, int training, float keep, int ctas, void* stream);
/* _ex variants for the pipelined step: conv1's grid is sized for blocks_per_sm resident blocks and its first
 * ag_blocks blocks deliver last step's updated Dense(128) rows to the peers (comm / ag may be NULL, ag_blocks 0);
 * conv2's spare warp waits for the peers' rows (wait may be NULL). */
int hvd_conv1_fwd_ex(const float* x, const float* w, const float* bias, float* a1, int B, float* zero_a, int n_a,
                     float* zero_b, int n_b, void* step_state, int advance_step, void* stream, int blocks_per_sm,
                     const struct HvdCommCtx* comm, const struct HvdAllgatherArgs* ag, int ag_blocks);
int hvd_conv2_fprop_ex(const float* a1, const float* w2, const float* bias2, float* p2, uint8_t* mask,
                       const void* step_state, int B, int training, float keep, int ctas, void* stream,
                       const struct HvdParamWait* wait);
/* g_w2 / g_b2 / g_b1 and the loss/accuracy words are ACCUMULATED into (zeroed by hvd_conv1_fwd's prologue). */
int hvd_head(float* h_pre, const float* b1, const float* w2, const float* b2, const long long* labels, float* dh_pre,
             float* g_w2, float* g_b2, float* g_b1, float* probs, void* step_state, int B, int training, int backward,
             float keep, void* stream);
/* da1 may be NULL.  With x != NULL the conv1 weight/bias gradients are accumulated (red.add) into
 * g_w1 [9][32] / g_b1 [32] by the same kernel (tensor-core GEMM over staged tiles; no da1 round trip). */
int hvd_conv2_dgrad(const float* dy2, const float* w2, const float* a1, float* da1, int B, int ctas, void* stream,
                    const float* x, float* g_w1, float* g_b1);
/* g_w [3][3][32][64] and g_b [64] are ACCUMULATED (red.add): zero them first.  ctas <= 0 picks a default. */
int hvd_conv2_wgrad(const float* a1, const float* dy2, float* g_w, float* g_b, int B, int ctas, void* stream);
/* Same, and CTA 0 also publishes the training step's result: it copies (last_loss, last_correct, step) from the
 * step-state block (written by hvd_head earlier on the stream) into slot step % 16 of `host_ring`, a page-locked
 * host buffer of 16 x 4 floats, as one 16-byte store.  The host polls the step word. */
int hvd_conv2_wgrad_ex(const float* a1, const float* dy2, float* g_w, float* g_b, int B, int ctas, void* stream,
                       const void* step_state, float* host_ring);
/* hvd_conv2_wgrad_ex followed by hvd_conv2_dgrad (with the fused conv1 gradients, no da1 output) as ONE kernel:
 * the first `wgrad_ctas` CTAs reduce the wgrad strips and then take `wgrad_cta_tiles` dgrad tiles each, the other
 * CTAs split the remaining dgrad tiles.  All four gradient outputs are accumulated (zero them first). */
int hvd_conv2_bwd(const float* dy2, const float* w2, const float* a1, const float* x, float* g_w2, float* g_b2,
                  float* g_w1, float* g_b1, int B, int ctas, void* stream, const void* step_state, float* host_ring,
                  int wgrad_ctas /* <= 0: default */, int wgrad_cta_tiles /* < 0: default */);
int hvd_conv1_wgrad(const float* x, const float* da1, float* g_w, float* g_b, int B, void* stream);
int hvd_colsum64(const float* m, float* out, long long rows, void* stream);

/* Bring-up / profiling aids (bench/mma_probe.py, bench/desc_probe.py, bench/conv_bwd_timing.py).
 * hvd_mma_probe: cycles for `iters` tcgen05.mma of shape M x N x 8 (MN = (M << 16) | N) rotating over n_acc
 * accumulators; out[cta] = elapsed clock64.  hvd_conv2_bwd_debug_read: the 64 phase timers CTA 0 of the last
 * conv2 dgrad launch recorded. */
int hvd_mma_probe(int MN, int iters, int n_acc, int a_bytes_step, int ctas, long long* out, void* stream);
int hvd_conv2_bwd_debug_read(long long* out64);
int hvd_desc_probe(const float* Ag, const float* Bg, float* D, int a_mn, int rows_a, int shift_rows, unsigned lbo,
                   unsigned sbo, int base_offset, void* stream);

/* Dense layer forward y = act(x*W + bias) as a split-K tcgen05 GEMM (keras Dense, reference
 * mnist_keras.py:79,81).  x: [M][K], W: [K][N] (Keras layout), y: [M][N]; relu: 0/1; K and N must be
 * multiples of 4 (TMA row strides are 16-byte granular; returns -2 otherwise).  The k_split CTAs
 * per tile accumulate with red.add into `scratch` (M*N floats, caller-allocated device memory that must
 * be zero on entry); the finalize pass adds bias/activation into y and zeroes scratch again. */
int hvd_dense_forward(const float* x, const float* W, const float* bias, float* y, int M, int N, int K, int relu,
                      int k_split, float* scratch, void* stream);

/* Dense forward for callers that may not allocate device memory (plugins inside a foreign memory
 * pool, latency-critical serving loops that must never hit cudaMalloc's device-wide lock, hosts with a
 * fixed memory budget).  No counterpart in the reference: it exists only because hvd_dense_forward
 * needs an M*N scratch accumulator that such callers cannot provide.  Same maths and arguments as
 * hvd_dense_forward minus `scratch`: the accumulator is a 1 MiB array built into libhvd_cuda.so (held
 * by every GPU that loads the library until exit).  Limit: M*N <= hvd_dense_forward_noalloc_max_elems()
 * (262144); larger sizes return -100 with a message in hvd_dense_forward_noalloc_error().  The first
 * call on a device does a one-time address look-up - make it before any stream capture.  The array
 * used is that of the device that is current at the call, and `stream` must belong to that device (a
 * stream of another device is refused with -101).  All calls on a device share the array: calls whose
 * GPU work may overlap race.  Every refusal or failure leaves a message in ..._noalloc_error(). */
int hvd_dense_forward_noalloc(const float* x, const float* W, const float* bias, float* y, int M, int N, int K,
                              int relu, int k_split, void* stream);
long long hvd_dense_forward_noalloc_max_elems(void);
const char* hvd_dense_forward_noalloc_error(void);

#ifdef __cplusplus
}
#endif
#endif /* HVD_B200_H_ */
