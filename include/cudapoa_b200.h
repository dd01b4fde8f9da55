/*
 * cudapoa_b200.h -- C ABI of the B200-native batched partial-order-alignment engine.
 *
 * This is the drop-in boundary for the cudapoa path of GenomeWorks 0.5.0.  Every entry point
 * below stands in for one member of the reference's C++ plugin interface, which is what the
 * reference's own FFI (the pygenomeworks Cython module) binds:
 *
 *   reference interface (file:line, relative to the GenomeWorks tree)        C entry point
 *   ------------------------------------------------------------------------ -----------------------------
 *   cudapoa::Init()                          cudapoa/include/.../cudapoa.hpp:41    cudapoa_b200_init
 *   BatchSize(3-arg ctor)                    cudapoa/include/.../batch.hpp:75      cudapoa_b200_batch_size_default
 *   BatchSize(6-arg ctor)                    cudapoa/include/.../batch.hpp:99      cudapoa_b200_batch_size_explicit
 *   create_batch(...)                        cudapoa/include/.../batch.hpp:220     cudapoa_b200_create_batch
 *   Batch::~Batch                            cudapoa/include/.../batch.hpp:137     cudapoa_b200_destroy_batch
 *   Batch::add_poa_group                     cudapoa/include/.../batch.hpp:153     cudapoa_b200_add_poa_group(_flat)
 *   Batch::get_total_poas                    cudapoa/include/.../batch.hpp:159     cudapoa_b200_get_total_poas
 *   Batch::generate_poa                      cudapoa/include/.../batch.hpp:162     cudapoa_b200_generate_poa
 *   Batch::get_consensus                     cudapoa/include/.../batch.hpp:174     cudapoa_b200_get_consensus + result accessors
 *   Batch::get_msa                           cudapoa/include/.../batch.hpp:186     cudapoa_b200_get_msa       + result accessors
 *   Batch::get_graphs                        cudapoa/include/.../batch.hpp:196     cudapoa_b200_get_graphs    + result accessors
 *   Batch::batch_id                          cudapoa/include/.../batch.hpp:202     cudapoa_b200_batch_id
 *   Batch::reset                             cudapoa/include/.../batch.hpp:205     cudapoa_b200_reset
 *   get_multi_batch_sizes                    cudapoa/include/.../utils.hpp:48      cudapoa_b200_get_multi_batch_sizes
 *   (pygenomeworks CudaPoaBatch binds exactly these: pygenomeworks/genomeworks/cudapoa/cudapoa.pxd:54-90)
 *
 * Plain pointers and sizes only; no C++ or torch types cross this boundary.  Status codes are the
 * numeric values of cudapoa::StatusType (cudapoa.hpp:26-38); a negative return means a C++ exception
 * was caught at the boundary (-1 std::invalid_argument, -2 std::runtime_error, -3 anything else) and
 * cudapoa_b200_last_error() holds its what() for the calling thread.
 *
 * There is no CPU fallback behind this interface: without a CUDA device create_batch fails.
 */
#ifndef CUDAPOA_B200_H
#define CUDAPOA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Symbol prefix.  The same C-ABI translation unit is also compiled against the reference's own
 * headers and library on the GPU box (oracle/Makefile) to obtain a bit-exactness oracle with an
 * identical calling convention; that build sets CUDAPOA_CABI_PREFIX to cudapoa_ref_. */
#ifndef CUDAPOA_CABI_PREFIX
#define CUDAPOA_CABI_PREFIX cudapoa_b200_
#endif
#define CUDAPOA_CABI_CAT2(a, b) a##b
#define CUDAPOA_CABI_CAT(a, b) CUDAPOA_CABI_CAT2(a, b)
#define CUDAPOA_CABI(name) CUDAPOA_CABI_CAT(CUDAPOA_CABI_PREFIX, name)

/* cudapoa::StatusType, same order and values (cudapoa.hpp:26-38). */
enum
{
    CUDAPOA_B200_SUCCESS                                   = 0,
    CUDAPOA_B200_EXCEEDED_MAXIMUM_POAS                     = 1,
    CUDAPOA_B200_EXCEEDED_MAXIMUM_SEQUENCE_SIZE            = 2,
    CUDAPOA_B200_EXCEEDED_MAXIMUM_SEQUENCES_PER_POA        = 3,
    CUDAPOA_B200_NODE_COUNT_EXCEEDED_MAXIMUM_GRAPH_SIZE    = 4,
    CUDAPOA_B200_EDGE_COUNT_EXCEEDED_MAXIMUM_GRAPH_SIZE    = 5,
    CUDAPOA_B200_SEQ_LEN_EXCEEDED_MAXIMUM_NODES_PER_WINDOW = 6,
    CUDAPOA_B200_LOOP_COUNT_EXCEEDED_UPPER_BOUND           = 7,
    CUDAPOA_B200_OUTPUT_TYPE_UNAVAILABLE                   = 8,
    CUDAPOA_B200_GENERIC_ERROR                             = 9
};

/* cudapoa::OutputType bit mask (cudapoa.hpp:44-48). */
enum
{
    CUDAPOA_B200_OUTPUT_CONSENSUS = 1,
    CUDAPOA_B200_OUTPUT_MSA       = 2
};

/* Mirror of cudapoa::BatchSize, fields in declaration order (batch.hpp:53-73). */
typedef struct cudapoa_b200_batch_size
{
    int32_t max_sequence_size;
    int32_t max_consensus_size;
    int32_t max_nodes_per_window;
    int32_t max_nodes_per_window_banded;
    int32_t max_matrix_graph_dimension;
    int32_t max_matrix_graph_dimension_banded;
    int32_t max_matrix_sequence_dimension;
    int32_t alignment_band_width;
    int32_t max_sequences_per_poa;
} cudapoa_b200_batch_size;

/* Opaque batch handle (one cudapoa::Batch). */
typedef struct cudapoa_b200_batch cudapoa_b200_batch;

/* Per-batch counters of the last generate_poa(); filled only by the native engine. */
typedef struct cudapoa_b200_stats
{
    int64_t windows;        /* windows processed                                             */
    int64_t reads_aligned;  /* sum over windows of (reads - 1)                               */
    int64_t dp_cells;       /* sum of N_s * Wc (SURVEY.md 8(d) work unit)                    */
    int64_t graph_rows;     /* sum of N_s                                                    */
    int64_t exact_path;     /* alignments that took the exact-arithmetic slow path           */
    int64_t h2d_bytes;      /* bytes copied host->device by generate_poa                     */
    int64_t d2h_bytes;      /* bytes copied device->host by the last get_*                   */
    double kernel_ms;       /* CUDA-event time of the device work of generate_poa            */
    double graph_kernel_ms; /* the POA graph-build kernel alone                              */
    int64_t widened;        /* adaptive banding: alignments repeated with a wider band       */
    int64_t kernel_launches; /* kernels enqueued by generate_poa (and by bench_relaunch since)   */
} cudapoa_b200_stats;

const char* CUDAPOA_CABI(last_error)(void);
const char* CUDAPOA_CABI(implementation)(void); /* "b200-native" or "genomeworks-reference" */

int32_t CUDAPOA_CABI(init)(void);

int32_t CUDAPOA_CABI(batch_size_default)(cudapoa_b200_batch_size* out, int32_t max_seq_sz, int32_t max_seq_per_poa,
                                         int32_t band_width);
int32_t CUDAPOA_CABI(batch_size_explicit)(cudapoa_b200_batch_size* out, int32_t max_seq_sz, int32_t max_consensus_sz,
                                          int32_t max_nodes_per_w, int32_t max_nodes_per_w_banded, int32_t band_width,
                                          int32_t max_seq_per_poa);

/* stream is a cudaStream_t passed as void* (NULL = default stream); the caller keeps ownership.
 * cuda_banded_alignment: 0 / 1 as the reference's bool; 2 (native engine only, an extension) = adaptive banding:
 * static band first, repeated with the band doubled (up to 4x) while the alignment runs along a band edge. */
cudapoa_b200_batch* CUDAPOA_CABI(create_batch)(int32_t device_id, void* stream, size_t max_mem, int8_t output_mask,
                                               const cudapoa_b200_batch_size* batch_size, int16_t gap_score,
                                               int16_t mismatch_score, int16_t match_score,
                                               int32_t cuda_banded_alignment);
void CUDAPOA_CABI(destroy_batch)(cudapoa_b200_batch* batch);

/* One group.  weights may be NULL, and any weights[i] may be NULL (= all ones).  per_seq_status has room for
 * n_entries values; *n_status receives how many were written (0 when the group itself was refused, matching
 * the reference, which clears the vector only after the score buffer was reserved). */
int32_t CUDAPOA_CABI(add_poa_group)(cudapoa_b200_batch* batch, const char* const* seqs, const int8_t* const* weights,
                                    const int32_t* lengths, int32_t n_entries, int32_t* per_seq_status,
                                    int32_t* n_status);

/* Many groups from flat host buffers: bases/weights are the reads of all windows back to back, read_lengths one
 * entry per read, window_n_reads one entry per window.  Adds windows in order by calling add_poa_group and stops
 * at the first group the batch refuses; *n_added = windows accepted.  per_seq_status (optional) gets one value
 * per read of the accepted windows.  Returns the status of the last add_poa_group call. */
int32_t CUDAPOA_CABI(add_poa_groups_flat)(cudapoa_b200_batch* batch, const uint8_t* bases, const int8_t* weights,
                                          const int32_t* read_lengths, const int32_t* window_n_reads,
                                          int32_t n_windows, int32_t* n_added, int32_t* per_seq_status);

int32_t CUDAPOA_CABI(get_total_poas)(const cudapoa_b200_batch* batch);
int32_t CUDAPOA_CABI(generate_poa)(cudapoa_b200_batch* batch);
int32_t CUDAPOA_CABI(batch_id)(const cudapoa_b200_batch* batch);
int32_t CUDAPOA_CABI(reset)(cudapoa_b200_batch* batch);

/* get_*: run the reference-named call and keep its outputs inside the handle until the next get_* / destroy.
 * The return value is what the C++ call returned (get_graphs returns void there: always 0 here). */
int32_t CUDAPOA_CABI(get_consensus)(cudapoa_b200_batch* batch);
int32_t CUDAPOA_CABI(get_msa)(cudapoa_b200_batch* batch);
int32_t CUDAPOA_CABI(get_graphs)(cudapoa_b200_batch* batch);

/* Accessors over the outputs of the last get_*.  Pointers stay valid until the next get_* or destroy. */
int32_t CUDAPOA_CABI(result_count)(const cudapoa_b200_batch* batch);       /* size of the consensus/msa/graphs vector */
int32_t CUDAPOA_CABI(result_status_count)(const cudapoa_b200_batch* batch); /* size of output_status                  */
int32_t CUDAPOA_CABI(result_status)(const cudapoa_b200_batch* batch, int32_t i);
int32_t CUDAPOA_CABI(result_consensus)(const cudapoa_b200_batch* batch, int32_t i, const char** seq,
                                       const uint16_t** coverage); /* returns length */
int32_t CUDAPOA_CABI(result_msa_rows)(const cudapoa_b200_batch* batch, int32_t i);
int32_t CUDAPOA_CABI(result_msa_row)(const cudapoa_b200_batch* batch, int32_t i, int32_t row,
                                     const char** seq); /* returns length */
/* Graph i as arrays: node labels (one char per node id 0..n-1; 0 for ids without a label) and an edge list
 * sorted by (sink, source).  Each call fills up to cap entries and returns the total count. */
int32_t CUDAPOA_CABI(result_graph_nodes)(const cudapoa_b200_batch* batch, int32_t i, char* labels, int32_t cap);
int32_t CUDAPOA_CABI(result_graph_edges)(const cudapoa_b200_batch* batch, int32_t i, int32_t* src, int32_t* dst,
                                         int32_t* weight, int32_t cap);
/* Whole-batch digest helpers used by the benchmark so that result decoding stays out of Python:
 * total characters over all consensus strings (or MSA rows), and an FNV-1a 64 hash of
 * status|length|bytes(|coverage) over all windows in order. */
int64_t CUDAPOA_CABI(result_total_chars)(const cudapoa_b200_batch* batch);
uint64_t CUDAPOA_CABI(result_digest)(const cudapoa_b200_batch* batch);

/* get_multi_batch_sizes over flat host buffers (lengths only are needed).  Writes up to cap batch sizes and,
 * per window, the index of the batch size it was assigned to (-1 if none).  Returns the number of batch sizes. */
int32_t CUDAPOA_CABI(get_multi_batch_sizes)(const int32_t* read_lengths, const int32_t* window_n_reads,
                                            int32_t n_windows, int32_t banded_alignment, int32_t msa_flag,
                                            int32_t band_width, float gpu_memory_usage_quota, int32_t mismatch_score,
                                            int32_t gap_score, int32_t match_score, cudapoa_b200_batch_size* out,
                                            int32_t cap, int32_t* window_to_batch);

/* Native-engine extras (absent from the reference build of this ABI). */
int32_t CUDAPOA_CABI(get_stats)(const cudapoa_b200_batch* batch, cudapoa_b200_stats* out);
int32_t CUDAPOA_CABI(max_poas)(const cudapoa_b200_batch* batch);

/* Native library only (no reference counterpart): out[i] = MUFU.RCP(float(i)) for i < n, evaluated on `device_id`.
 * The reference is built with -use_fast_math, so its band geometry divides through this instruction
 * (cudapoa_nw_banded.cuh:206); tests/golden/make_rcp_table.py stores the table so that the CPU oracle can
 * reproduce the device's band starts bit for bit.  Returns 0, or a cudaError_t value. */
int32_t cudapoa_b200_debug_rcp_table(int32_t device_id, float* out, int32_t n);

/* Native library only: DPX issue-rate micro-benchmark used as the roofline denominator.
 * Runs independent __viaddmax_s16x2 (kind 0) or __viaddmax_s32 (kind 1) chains, 8 accumulators per thread, on every
 * SM of `device_id` and returns thread-level instructions per second in *instr_per_s and the kernel time in *ms. */
int32_t cudapoa_b200_debug_dpx_peak(int32_t device_id, int32_t kind, double* instr_per_s, double* ms);

/* Native library only.  NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * Banded consensus (what create_batch(..., OutputType::consensus, BatchSize(max_sequence_size, max_sequences_per_poa,
 * alignment_band_width), gap, mismatch, match, true) + add_poa_group + generate_poa + get_consensus compute) for
 * n_windows windows whose inputs and outputs the CALLER holds in device memory; the function allocates nothing -- no
 * cudaMalloc, cudaFree or allocating library call -- and only enqueues work on `stream` (no synchronisation).
 * All per-window state lives in one file-scope device array of CUDAPOA_B200_NOALLOC_SCRATCH_BYTES that every GPU
 * loading the library holds until the process exits.  Consequences:
 *   - a request whose state exceeds that size is refused (-3; the message names the bytes needed, the maximum and the
 *     bytes per window): 12 windows of BatchSize(640, 30, 256), 32 of BatchSize(256, 8, 256);
 *   - the array is shared by all calls and keeps the last call's contents (what must start at zero is cleared by every
 *     call): calls that overlap on the GPU -- two streams, two host threads -- RACE; serialise them;
 *   - int16 configurations only (max_sequence_size below ~1490 with scores 8/-6/-8): others are refused (-2);
 *   - the FIRST call per device looks up the array's device address and the launch limits (cudaGetSymbolAddress,
 *     cudaFuncSetAttribute, an occupancy query; with lazy module loading it may also load the module, which is what
 *     reserves the array): make one call per device OUTSIDE stream capture before using the entry point under capture;
 *   - graph nodes are always ordered by the reference's default (Kahn) topological sort: the SPOA_ACCURATE build flag
 *     and the CUDAPOA_B200_SPOA_ACCURATE environment switch of create_batch do not apply here;
 *   - a window whose read count is negative or above max_sequences_per_poa makes the offsets of all LATER windows
 *     unknowable: it and every window after it get status generic_error (9).
 * Inputs (device): bases_d = the reads of all windows back to back; read_lengths_d = one int32 per read;
 * window_n_reads_d = one int32 per window.  No weights (every base weighs 1).  A window with no reads, more reads than
 * max_sequences_per_poa or a read longer than max_sequence_size gets status generic_error (9).
 * Outputs (device), C = 2 * max_sequence_size: consensus_d[n_windows * C] (start -> end, NUL terminated when shorter
 * than C), coverage_d[n_windows * C], length_d[n_windows], status_d[n_windows] (StatusType values).
 * Returns 0, or a negative code with a message in error_msg (may be NULL): -1 bad argument, -2 needs 32-bit scores,
 * -3 scratch too small, -4 CUDA error. */
#define CUDAPOA_B200_NOALLOC_SCRATCH_BYTES (32 * 1024 * 1024)
int32_t cudapoa_b200_consensus_noalloc(int32_t device_id, void* stream, const uint8_t* bases_d,
                                       const int32_t* read_lengths_d, const int32_t* window_n_reads_d,
                                       int32_t n_windows, int32_t max_sequences_per_poa, int32_t max_sequence_size,
                                       int32_t alignment_band_width, int16_t gap_score, int16_t mismatch_score,
                                       int16_t match_score, uint8_t* consensus_d, uint16_t* coverage_d,
                                       int32_t* length_d, int32_t* status_d, char* error_msg, int32_t error_msg_cap);

/* Native library only: benchmark hook.  Re-runs the device work of the last generate_poa() `iterations` times on
 * the inputs already resident in device memory (no host<->device copies) and returns the CUDA-event time of all
 * iterations in milliseconds, measured on the batch's own stream; negative if nothing has been staged. */
double cudapoa_b200_bench_relaunch(cudapoa_b200_batch* batch, int32_t iterations);

/* Native library only (tests, A/B measurements): how banded batches are scheduled on the device from now on, process
 * wide.  0 = default (one launch per phase and read -- poa_fill_kernel / poa_graph_kernel -- unless the environment
 * says CUDAPOA_B200_SCHEDULE=fused), 1 = always the single launch (poa_window_kernel), 2 = always phase-split.
 * Adding 4 or 8 also fixes the traceback kernel of the phase-split schedule: 4 = one warp per window
 * (poa_trace_kernel), 8 = one thread per window (poa_trace_lanes_kernel; by default batches of 12 288 windows and more).
 * Results are identical either way. */
void cudapoa_b200_debug_set_schedule(int32_t mode);

/* Native library only: the same device work once more with a CUDA event between the kernel launches.  *fill_ms and
 * *graph_ms receive the time spent in the score-fill launches and in the graph launches (traceback, fusion,
 * topological sort, consensus) of the phase-split schedule; returns the number of launches, negative on error. */
int32_t cudapoa_b200_bench_relaunch_phases(cudapoa_b200_batch* batch, double* fill_ms, double* graph_ms);

/* Native library only: bench_relaunch and bench_relaunch_phases in one -- `iterations` passes with a CUDA event between
 * the launches.  Returns the event time of all passes in milliseconds (negative on error), the time spent in the fill
 * and in the graph launches in *fill_ms / *graph_ms (sums over the passes) and the number of launches in *launches. */
double cudapoa_b200_bench_relaunch_timed(cudapoa_b200_batch* batch, int32_t iterations, double* fill_ms,
                                         double* graph_ms, int32_t* launches);

/* Native library only: per-phase totals of the last bench_relaunch_timed call, in milliseconds: out[0] score fill,
 * out[1] traceback, out[2] fusion into the graph, out[3] topological sort, out[4] consensus / MSA.  Returns how many
 * were written (at most `cap`), negative on error.  No reference counterpart (measurement only). */
int32_t cudapoa_b200_bench_phase_ms(cudapoa_b200_batch* batch, double* out, int32_t cap);

/* Native library only (development aid): copy a window's device state to the host after generate_poa.  region 0 = the
 * window's arena slice (graph arrays, row records, scratch), 1 = its score area (move codes, spilled score rows).
 * Returns the region's size in bytes (copies min(size, cap)), negative on error. */
int64_t cudapoa_b200_debug_read_window(cudapoa_b200_batch* batch, int32_t window, int32_t region, void* out, int64_t cap);

/* ---- Window pipeline (native library only; no reference counterpart) ------------------------------------------
 * The stream scheduler / sharding layer above the Batch interface (the reference drives several Batch objects from
 * std::async tasks in cudapoa/benchmarks/multi_batch.hpp:157-170; it has no multi-GPU code).  Each worker thread owns
 * one stream and one Batch on one device and pulls chunks of windows from a shared queue; results land in slots
 * indexed by window, so output order is deterministic.  devices may name the same device several times. */
typedef struct cudapoa_b200_pipeline cudapoa_b200_pipeline;
const char* cudapoa_b200_pipeline_last_error(void);
cudapoa_b200_pipeline* cudapoa_b200_pipeline_create(const int32_t* devices, int32_t n_devices,
                                                    int32_t workers_per_device, size_t mem_per_worker,
                                                    int8_t output_mask, const cudapoa_b200_batch_size* batch_size,
                                                    int16_t gap_score, int16_t mismatch_score, int16_t match_score,
                                                    int32_t cuda_banded_alignment);
void cudapoa_b200_pipeline_destroy(cudapoa_b200_pipeline* p);
/* Blocking: all windows (flat host buffers as in add_poa_groups_flat) through the workers.  chunk_windows = windows per
 * batch round (0: automatic).  Returns 0 or a negative exception code (see last_error). */
int32_t cudapoa_b200_pipeline_run(cudapoa_b200_pipeline* p, const uint8_t* bases, const int8_t* weights,
                                  const int32_t* read_lengths, const int32_t* window_n_reads, int32_t n_windows,
                                  int32_t chunk_windows);
int32_t cudapoa_b200_pipeline_workers(const cudapoa_b200_pipeline* p);
int32_t cudapoa_b200_pipeline_capacity(const cudapoa_b200_pipeline* p); /* smallest max_poas over the workers */
double cudapoa_b200_pipeline_wall_ms(const cudapoa_b200_pipeline* p);   /* wall time of the last run            */
int32_t cudapoa_b200_pipeline_stats(const cudapoa_b200_pipeline* p, cudapoa_b200_stats* out); /* summed over rounds */
int32_t cudapoa_b200_pipeline_status(const cudapoa_b200_pipeline* p, int32_t window);
int32_t cudapoa_b200_pipeline_consensus(const cudapoa_b200_pipeline* p, int32_t window, const char** seq,
                                        const uint16_t** coverage); /* returns length */
int32_t cudapoa_b200_pipeline_msa_rows(const cudapoa_b200_pipeline* p, int32_t window);
int32_t cudapoa_b200_pipeline_msa_row(const cudapoa_b200_pipeline* p, int32_t window, int32_t row, const char** seq);
int64_t cudapoa_b200_pipeline_total_chars(const cudapoa_b200_pipeline* p);
/* Same hash as cudapoa_b200_result_digest of one batch holding windows [first, first + count). */
uint64_t cudapoa_b200_pipeline_digest(const cudapoa_b200_pipeline* p, int32_t first, int32_t count);

#ifdef __cplusplus
}
#endif

#endif /* CUDAPOA_B200_H */
