/* bsmm.h — C-ABI of the B200-native block-sparse matmul library (libbsmm.so).
 *
 * This is the drop-in boundary for ONE path of tgale96/sputnik: the block-sparse
 * SDD / DSD / DDS matmuls at 128x128 blocks over BCSR topology.  Every entry point names
 * the reference interface it replaces (paths relative to the reference repository).
 *
 * Conventions (same as the reference, sputnik/block/arguments.h:48-162):
 *   - all data / metadata pointers are DEVICE pointers owned by the caller;
 *   - rows, cols, nonzeros are in ELEMENTS (nonzeros = nnz_blocks * 128 * 128);
 *   - dense matrices are row-major, leading dimension == their column count; a transpose
 *     flag means "stored transposed";
 *   - offsets / offsets_t / block_offsets are int32, indices / indices_t / row_indices int16;
 *   - alpha = 1, beta = 0; fp32 accumulation; output rounded to nearest 16-bit value;
 *   - work is enqueued on `stream` (a cudaStream_t passed as void*) and is fully
 *     asynchronous — including the transposed-metadata build, which the reference does on
 *     the host with blocking copies (sputnik/block/transpose/transpose.cu:69-125).
 *
 * Unlike the reference (which aborts the process on programmer errors, sputnik/logging.cc:14-18)
 * the C-ABI never aborts: it returns a bsmm_status.  The C++ shim include/sputnik/sputnik.h maps
 * these back onto the reference's cudaError_t / abort behaviour.
 */
#ifndef BSMM_H_
#define BSMM_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  BSMM_OK = 0,
  BSMM_ERR_BLOCK_SIZE = 1,   /* block_size != 128 (reference: cudaErrorNotSupported, sdd.cu:11-14) */
  BSMM_ERR_MISALIGNED = 2,   /* m/n/k not multiples of 8, sparse dims not multiples of 128, or a
                                pointer not 16-byte aligned (reference: can_implement failure) */
  BSMM_ERR_SHAPE = 3,        /* shapes disagree (reference: ValidMatmul, arguments.h:217-231) */
  BSMM_ERR_MISSING_META = 4, /* a required metadata pointer is null (reference: SPUTNIK_CHECK abort) */
  BSMM_ERR_CUDA = 5,         /* a CUDA runtime / driver call failed; see bsmm_last_cuda_error */
  BSMM_ERR_DTYPE = 6,        /* dtype is not BSMM_F16 / BSMM_BF16 (/ BSMM_F8E4M3 where supported) */
  BSMM_ERR_RANGE = 7,        /* a dimension exceeds what int16 indices can address */
  BSMM_ERR_HINT = 8,         /* BSMM_VALIDATE_HINTS=1 only: a topology hint flag of the call does
                                not hold for the topology it was passed with */
  BSMM_ERR_STATIC_LIMIT = 9  /* bsmm_dsd_static only: the sparse operand exceeds
                                BSMM_STATIC_MAX_BLOCKS / BSMM_STATIC_MAX_BLOCK_COLS */
} bsmm_status;

typedef enum {
  BSMM_F16 = 0, /* the reference's only type (cutlass::half_t in every instantiation) */
  BSMM_BF16 = 1, /* extension */
  /* extension, SDD / DSD / DDS only: every INPUT operand — dense matrices and
   * the values of a sparse operand — holds FP8 e4m3 (1 byte per element, torch.float8_e4m3fn;
   * dense operand dims multiples of 16), products are exact, accumulation is fp32
   * (tcgen05.mma.kind::f8f6f4) and the OUTPUT (dense C, or the values of a sparse C) is bf16.
   * Scaling factors are the caller's business, as with any fp8 GEMM. */
  BSMM_F8E4M3 = 2
} bsmm_dtype;

/* Field-for-field mirror of sputnik::block::BlockMatrix (sputnik/block/arguments.h:48-68),
 * minus `bitmask` (the reference's DSS needs it; ours does not — see bsmm_bitmask) and
 * `create_metadata` (a per-call flag here). */
typedef struct {
  int rows, cols, nonzeros;
  int block_size;
  void* data;
  void* offsets;       /* int32 [rows/128 + 1] */
  void* indices;       /* int16 [nnz_blocks] block-column ids, any order within a row */
  void* offsets_t;     /* int32 [cols/128 + 1] */
  void* indices_t;     /* int16 [nnz_blocks] block-row id of the i-th block in column order */
  void* block_offsets; /* int32 [nnz_blocks] position of that block in `data` */
  void* row_indices;   /* int16 [nnz_blocks] (sparse outputs) */
} bsmm_block_matrix;

/* Mirror of sputnik::block::Matrix (sputnik/block/arguments.h:155-162). */
typedef struct {
  int rows, cols;
  void* data;
} bsmm_matrix;

/* flags */
#define BSMM_FLAG_NONE 0u
/* Build offsets_t / indices_t / block_offsets on the device before a transposed-sparse matmul
 * (what Matmul() does by default, BlockMatrix::create_metadata, arguments.h:67-68).  Without it
 * the call behaves like MatmulEx (dsd.cu:20-27, dds.cu:17-24): metadata is taken as valid. */
#define BSMM_FLAG_CREATE_METADATA 1u

/* Topology hints (extensions; a wrong hint gives wrong results, like stale MatmulEx metadata).
 * TWIN_ROWS: block rows 2i and 2i+1 of the sparse operand have identical column-index lists (in
 * the same order) for every i; TWIN_COLS: the same for block columns 2i, 2i+1 of the transposed
 * view (indices_t).  dMoE block-diagonal topologies with an even number of block rows per expert
 * (resp. an even ffn/128) satisfy them.  With the hint that matches the walk direction of the
 * call, DSD / DDS run the 2-SM kernel (tcgen05.mma.cta_group::2 over CTA pairs), which halves the
 * dense-operand traffic per SM. */
/* Debug validation: with BSMM_VALIDATE_HINTS=1 in the environment (read on every call) each
 * product entry point first checks, on the device, every hint flag it would act on and returns
 * BSMM_ERR_HINT instead of launching when one does not hold.  This makes the call synchronous
 * (4 bytes come back to the host); it is skipped while `stream` is being captured. */
#define BSMM_FLAG_TWIN_ROWS 2u
#define BSMM_FLAG_TWIN_COLS 4u
/* EVEN_ROWS: every block row holds an even number of blocks.  SDD runs its 2-SM kernel (2 x 2
 * cliques of output blocks) when given TWIN_ROWS | EVEN_ROWS. */
#define BSMM_FLAG_EVEN_ROWS 8u
/* DEVICE_NNZ (extension): the sparse matrix is sized for a CAPACITY — `rows`, `nonzeros` and the
 * buffers are upper bounds chosen on the host — while the live block count is what the device
 * reads at offsets[rows / 128]; block rows past the live ones are empty.  SDD computes only the
 * live blocks (DSD / DDS need no flag: empty rows already produce zeros / contribute nothing).
 * This is what lets a dMoE step whose token counts exist only on the device run without a host
 * synchronisation: see bsmm_moe_topology. */
#define BSMM_FLAG_DEVICE_NNZ 16u

/* Fused epilogues (extensions; the reference's kernels are instantiated with
 * LinearCombination alpha = 1, beta = 0 and nothing else, e.g.
 * sdd_mixed_b128_128x128x32x5_nn_align8.cu:83).  Accepted by every product entry point.
 * ACCUMULATE: C += op(A) op(B) instead of C = ...: the product is rounded to the 16-bit type and
 *   added to the value already in C, rounded again — bit for bit what "product into a temporary,
 *   then an elementwise add" stores, without the temporary and its extra passes over C (weight-
 *   gradient accumulation over micro-batches: dw += dsd(h^T, dy), dw += dds(x^T, dh)).
 * RELU / GELU: C = act(op(A) op(B)), applied to the fp32 accumulator before the one rounding
 *   (GELU in its tanh form, the dMoE MLP's activation on the SDD output).  With ACCUMULATE:
 *   C += act(...).  bsmm_dsd_scatter ignores ACCUMULATE. */
#define BSMM_FLAG_ACCUMULATE 32u
#define BSMM_FLAG_RELU 64u
#define BSMM_FLAG_GELU 128u
/* bsmm_sdd only (extension): gated-linear-unit epilogue.  op(B) has TWICE the
 * columns of c: for output block column j its block column 2j is the gate panel and 2j + 1 the up
 * panel (weights interleaved at 128-column granularity).  Each nonzero block of c receives
 * silu(op(A) gate) * (op(A) up) — both products accumulated in fp32 in one 128 x 256 tile, the
 * gate / up intermediates never reach memory.  Not combinable with ACCUMULATE / RELU / GELU
 * (BSMM_ERR_SHAPE), nor accepted by any other product (BSMM_ERR_SHAPE).  With BSMM_FLAG_TWIN_ROWS the 2-SM kernel is used (a CTA pair per pair of block
 * rows, one CTA loading the gate panel and the other the up panel), otherwise the 1-SM kernel. */
#define BSMM_FLAG_SWIGLU 256u

/* SDD: C(sparse, sampled at c's topology) = op(A) * op(B).
 * Replaces sputnik::block::Matmul(Matrix, bool, Matrix, bool, BlockMatrix, cudaStream_t)
 * (sputnik/block/sdd/sdd.h:10-15, sdd.cu:7-15).  Requires c->row_indices. */
int bsmm_sdd(const bsmm_matrix* a, int transpose_a, const bsmm_matrix* b, int transpose_b,
             const bsmm_block_matrix* c, int dtype, unsigned flags, void* stream);

/* DSD: C(dense) = op(A sparse) * op(B).
 * Replaces Matmul / MatmulEx(BlockMatrix, bool, Matrix, bool, Matrix, cudaStream_t)
 * (sputnik/block/dsd/dsd.h:10-22, dsd.cu:9-27).  transpose_a needs the transposed metadata. */
int bsmm_dsd(const bsmm_block_matrix* a, int transpose_a, const bsmm_matrix* b, int transpose_b,
             const bsmm_matrix* c, int dtype, unsigned flags, void* stream);

/* DSD that allocates nothing, ever (no counterpart in the reference: added
 * for callers that cannot call cudaMalloc — code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same product and arguments as bsmm_dsd through the
 * 1-SM kernel (topology hint flags and BSMM_FLAG_CREATE_METADATA are ignored), except that every
 * byte of scratch the call needs lives in ONE file-scope `static __device__` array of the library
 * (about 105 KiB, held by every GPU that loads the library until the process exits):
 *   - the tile counter of the dynamic scheduler (bsmm_dsd takes it from cudaMalloc'ed scratch);
 *   - with transpose_a, the transposed metadata (offsets_t / indices_t / block_offsets), which
 *     the call builds itself from a->offsets / a->indices: the caller's a->offsets_t, a->indices_t
 *     and a->block_offsets are neither needed nor read, so no transpose buffers have to exist.
 * Limits (BSMM_ERR_STATIC_LIMIT beyond them, checked before any CUDA call):
 *   a->nonzeros / 128^2 <= BSMM_STATIC_MAX_BLOCKS and a->cols / 128 <= BSMM_STATIC_MAX_BLOCK_COLS.
 * The call makes no cudaMalloc / cudaFree / stream-ordered allocation; its first use on a device
 * resolves the array's device address and opts the kernels in to their shared-memory size (do it
 * once outside a capture).  The array is shared by every call and keeps the last call's
 * contents; the call clears the counter on `stream` first.  Two calls whose GPU work overlaps
 * (different streams, or a graph replayed beside an eager call) therefore RACE: serialise them. */
#define BSMM_STATIC_MAX_BLOCKS 16384
#define BSMM_STATIC_MAX_BLOCK_COLS 2048
int bsmm_dsd_static(const bsmm_block_matrix* a, int transpose_a, const bsmm_matrix* b,
                    int transpose_b, const bsmm_matrix* c, int dtype, unsigned flags, void* stream);

/* DSD with scattered output rows (extension; SURVEY.md 8e/8f: the un-permute / un-pad / return
 * all-to-all that follows y = dsd(h, w2) in a dMoE layer, fused into the epilogue).  Row i of
 * op(A) * op(B) (c_cols elements) is written at c_row_ptrs[i]; a null entry drops the row.
 * c_row_ptrs is a DEVICE array of c_rows pointers, each 16-byte aligned.  The pointers may be
 * peer-GPU addresses (P2P-mapped / symmetric memory): the epilogue's 128 B line stores then go
 * straight over NVLink and replace the collective — the caller orders them against the peers'
 * reads with its own barrier.  Everything else as bsmm_dsd. */
int bsmm_dsd_scatter(const bsmm_block_matrix* a, int transpose_a, const bsmm_matrix* b,
                     int transpose_b, int c_rows, int c_cols, void* const* c_row_ptrs, int dtype,
                     unsigned flags, void* stream);

/* Row gather, the mirror of bsmm_dsd_scatter on the way in (the dMoE token dispatch: permute,
 * pad and — when the sources are peer-GPU addresses — the all-to-all, in one pass): row i of the
 * row-major dst [rows][cols] (16-bit elements, cols a multiple of 8, dst 16-byte aligned) is copied
 * from src_row_ptrs[i] (DEVICE array; each 16-byte aligned); a null entry gives a zero row.
 * HBM / NVLink-bound copy: 2 * cols bytes read + written per row. */
int bsmm_gather_rows(void* const* src_row_ptrs, int rows, int cols, void* dst, void* stream);

/* DDS: C(dense) = op(A) * op(B sparse).
 * Replaces Matmul / MatmulEx(Matrix, bool, BlockMatrix, bool, Matrix, cudaStream_t)
 * (sputnik/block/dds/dds.h:10-22, dds.cu:7-24).  NON-transposed B needs the transposed metadata
 * (dds_mixed_b128_256x128x32x3_nn_align8.cu:74-84). */
int bsmm_dds(const bsmm_matrix* a, int transpose_a, const bsmm_block_matrix* b, int transpose_b,
             const bsmm_matrix* c, int dtype, unsigned flags, void* stream);

/* ---- sparse-output / sparse x sparse siblings ----
 * flags: BSMM_FLAG_CREATE_METADATA builds the transposed metadata of EVERY sparse input that is
 * walked by block column (a when transpose_a; b when !transpose_b) before the product; without
 * it the metadata is taken as valid (the reference's MatmulEx).  The reference additionally needs
 * caller-allocated bitmasks for DSS (dss_mixed_b128_128x128x32x5_nn_align8.cu:76-77) which it
 * fills through a host round trip on every call (:90-101); this library intersects the index
 * lists on the device inside the product and needs none. */

/* SSD: C(sparse, sampled at c's topology) = op(A sparse) * op(B).
 * Replaces Matmul / MatmulEx(BlockMatrix, bool, Matrix, bool, BlockMatrix, cudaStream_t)
 * (sputnik/block/ssd/ssd.h, ssd.cu:7-24).  Requires c->row_indices; transpose_a needs a's
 * transposed metadata (ssd_mixed_b128_128x128x32x5_tn_align8.cu:75-77). */
int bsmm_ssd(const bsmm_block_matrix* a, int transpose_a, const bsmm_matrix* b, int transpose_b,
             const bsmm_block_matrix* c, int dtype, unsigned flags, void* stream);

/* SDS: C(sparse, sampled at c's topology) = op(A) * op(B sparse).
 * Replaces Matmul / MatmulEx(Matrix, bool, BlockMatrix, bool, BlockMatrix, cudaStream_t)
 * (sputnik/block/sds/sds.h, sds.cu:7-24).  Requires c->row_indices; NON-transposed b needs b's
 * transposed metadata. */
int bsmm_sds(const bsmm_matrix* a, int transpose_a, const bsmm_block_matrix* b, int transpose_b,
             const bsmm_block_matrix* c, int dtype, unsigned flags, void* stream);

/* DSS: C(dense) = op(A sparse) * op(B sparse): per output tile the contraction runs over the
 * intersection of op(A)'s block row and op(B)'s block column.
 * Replaces Matmul / MatmulEx(BlockMatrix, bool, BlockMatrix, bool, Matrix, cudaStream_t)
 * (sputnik/block/dss/dss.h, dss.cu:9-27).  K <= 262144 (reference: 32768, index_merge.h:103). */
int bsmm_dss(const bsmm_block_matrix* a, int transpose_a, const bsmm_block_matrix* b,
             int transpose_b, const bsmm_matrix* c, int dtype, unsigned flags, void* stream);

/* K-contiguous bitmask of a topology: uint64 words [walk rows][ceil(walk cols / 64)], bit (c % 64)
 * of word c / 64 set for every block (r, c).  The transposed view (offsets_t / indices_t) is
 * walked when m->offsets_t is non-null, exactly like the reference.  Replaces
 * sputnik::block::Bitmask (sputnik/block/bitmask/bitmask.cu:7-45, host code with blocking copies;
 * layout bit_matrix.h:12-62) with a device kernel; bit-identical.  Not needed by bsmm_dss. */
int bsmm_bitmask(const bsmm_block_matrix* m, void* bitmask, void* stream);
/* Bytes of the caller-owned bitmask buffer (AllocateBitmaskBuffers, bitmask.h:16-23). */
int bsmm_bitmask_bytes(const bsmm_block_matrix* m, size_t* bytes);

/* row_indices[offsets[r] .. offsets[r+1]) = r.
 * Replaces sputnik::block::RowIndices (sputnik/block/row_indices/row_indices.h:10,
 * row_indices.cu:7-36). */
int bsmm_row_indices(const bsmm_block_matrix* a, short* row_indices, void* stream);

/* Fill a->offsets_t, a->indices_t, a->block_offsets from a->offsets, a->indices: a stable sort
 * of the blocks by block column, bit-identical to the host code in
 * sputnik/block/transpose/transpose.cu:69-125, but as device kernels (no stream stall). */
int bsmm_transpose_metadata(const bsmm_block_matrix* a, void* stream);

/* Bytes of caller-owned buffers the transposed metadata needs
 * (AllocateTransposeBuffers, sputnik/block/arguments.h:233-245). */
int bsmm_transpose_buffer_bytes(const bsmm_block_matrix* a, size_t* offsets_t_bytes,
                                size_t* indices_t_bytes, size_t* block_offsets_bytes);

/* dMoE topology built on the device from the router histogram (no reference
 * counterpart: MegaBlocks builds it in Python around the reference and synchronises with the host
 * for the sizes).  One kernel launch, no host round trip:
 *   tokens_per_expert  device int32[num_experts]
 *   expert e gets n_e = ceil(tokens_e / (128 * row_align_blocks)) * row_align_blocks block rows
 *   (row_align_blocks = 2 makes every topology TWIN_ROWS) and the block columns
 *   [e * ffn_blocks, (e + 1) * ffn_blocks), all n_e x ffn_blocks blocks present.
 * `out` describes the CAPACITY: rows = cap_block_rows * 128, cols = num_experts * ffn_blocks * 128,
 * nonzeros = cap_block_rows * ffn_blocks * 128 * 128; it must hold offsets, indices and
 * row_indices, and either all or none of offsets_t / indices_t / block_offsets.  Block rows past
 * sum(n_e) are empty, the tails of the index arrays are zero-filled: use the result with
 * BSMM_FLAG_DEVICE_NNZ.  The metadata equals, bit for bit, what bsmm_row_indices and
 * bsmm_transpose_metadata produce for the same topology.  Experts that do not fit in the capacity
 * are truncated (no out-of-bounds write).  padded_bins (optional, device int32[num_experts + 1]):
 * first padded token row of each expert and, last, the unclamped total — compare it with
 * out->rows to detect truncation after the fact. */
int bsmm_moe_topology(const int* tokens_per_expert, int num_experts, int ffn_blocks,
                      int row_align_blocks, const bsmm_block_matrix* out, int* padded_bins,
                      void* stream);

/* Exchange tables of an expert-sharded dMoE layer, built on the device (the other half of
 * bsmm_moe_topology; no reference counterpart).  counts: device int32 [world][num_experts], the
 * token rows every rank holds per global expert (each rank's rows sorted by expert) — e.g. the
 * all-gathered router histograms.  This rank owns experts [rank * E / world, (rank + 1) * E / world);
 * its local rows are expert-major, source-rank-minor, padded per expert to 128 * row_align_blocks —
 * the row layout of bsmm_moe_topology for tokens_per_expert, which is written here (device int32
 * [E / world], may be null).  For each of the 128 * cap_block_rows local rows the kernel writes, into
 * each of the num_tables (<= 4) device tables (uint64 [128 * cap_block_rows]), the address
 * peer_bases[t * world + s] + r * row_pitch_bytes of the row it corresponds to (source rank s, row r
 * of that rank's input), or 0 for padding rows: tables for bsmm_gather_rows (token dispatch) and
 * bsmm_dsd_scatter (return), which differ only in their base addresses.  peer_bases is a HOST
 * array of num_tables * world pointers (peer mappings of the same symmetric buffer).  One launch,
 * no host synchronisation: a training step whose routing exists only on the device. */
int bsmm_moe_exchange(const int* counts, int world, int num_experts, int rank, int row_align_blocks,
                      int cap_block_rows, size_t row_pitch_bytes, int num_tables,
                      const void* const* peer_bases, void* const* tables, int* tokens_per_expert,
                      void* stream);

/* Computes the topology hint flags a matrix qualifies for: BSMM_FLAG_TWIN_ROWS and
 * BSMM_FLAG_EVEN_ROWS from offsets / indices, and BSMM_FLAG_TWIN_COLS from offsets_t / indices_t
 * when those are non-null (they must be valid: build them first).  An analysis step in the
 * cuSPARSE sense — call it when a topology is created, pass the flags to every product on it.
 * SYNCHRONOUS: runs two small kernels on `stream` and waits for the 4-byte result.  Never called
 * by the matmul entry points. */
int bsmm_analyze_topology(const bsmm_block_matrix* a, unsigned* hint_flags, void* stream);

/* Shape / alignment / metadata validation only (no launch): what the reference's
 * can_launch_* + ValidMatmul decide (e.g. sdd_mixed_b128_128x128x32x5_nn_align8.cu:48-68). */
int bsmm_sdd_check(const bsmm_matrix* a, int transpose_a, const bsmm_matrix* b, int transpose_b,
                   const bsmm_block_matrix* c, int dtype);
int bsmm_dsd_check(const bsmm_block_matrix* a, int transpose_a, const bsmm_matrix* b,
                   int transpose_b, const bsmm_matrix* c, int dtype, unsigned flags);
int bsmm_dds_check(const bsmm_matrix* a, int transpose_a, const bsmm_block_matrix* b,
                   int transpose_b, const bsmm_matrix* c, int dtype, unsigned flags);

int bsmm_ssd_check(const bsmm_block_matrix* a, int transpose_a, const bsmm_matrix* b,
                   int transpose_b, const bsmm_block_matrix* c, int dtype, unsigned flags);
int bsmm_sds_check(const bsmm_matrix* a, int transpose_a, const bsmm_block_matrix* b,
                   int transpose_b, const bsmm_block_matrix* c, int dtype, unsigned flags);
int bsmm_dss_check(const bsmm_block_matrix* a, int transpose_a, const bsmm_block_matrix* b,
                   int transpose_b, const bsmm_matrix* c, int dtype, unsigned flags);

/* Optional one-time set-up for the CURRENT device: allocates and zeroes the library's per-device
 * scratch (scheduler tile counters and gate words, 48 KiB) and resolves the driver entry point
 * used to encode TMA descriptors.  The product entry points do the same at first use; call this
 * once per device before capturing CUDA graphs (a first use INSIDE a capture cannot allocate and
 * falls back to the static tile schedule for that launch) or before timing. */
int bsmm_init(void);

const char* bsmm_status_string(int status);
/* cudaError_t (as int) of the last failing CUDA call on this host thread, 0 if none. */
int bsmm_last_cuda_error(void);
/* Source line (csrc/bsmm_api.cu) of that failing call: a debugging aid. */
int bsmm_last_cuda_error_line(void);
/* Number of kernels this library has launched from this process (bench.py's gpu_launches). */
unsigned long long bsmm_launch_count(void);
const char* bsmm_version(void);

#ifdef __cplusplus
}
#endif
#endif /* BSMM_H_ */
