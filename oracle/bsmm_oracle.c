/* bsmm_oracle.c — CPU restatement of the reference's host-side algorithms for the block-sparse
 * SDD / DSD / DDS path.  TEST INFRASTRUCTURE ONLY: nothing under sputnik_b200/ may call this;
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
 *
 * Parity pin: the reference holds NO golden vectors (unseeded absl::BitGen, sdd_test.cu:66).
 * This restatement is pinned instead against the reference's own code compiled from
 * /root/reference (oracle/_ref, see oracle/Makefile): tests/test_oracle_vs_ref.py and the
 * committed fixtures under tests/golden/ (made by oracle/make_golden.py through oracle/_ref).
 *
 * All matrices are fp32 row-major; callers pass values already rounded to the 16-bit type the
 * device sees.  Citations are paths inside the reference repository.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define BLK 128

/* sputnik/matrix_utils.h:376-391 — Matrix::operator*: naive i-j-l loop, product in float,
 * accumulation in double, result cast to float. */
void oracle_matmul(const float* a, const float* b, float* out, int m, int k, int n) {
  for (int i = 0; i < m; ++i) {
    for (int j = 0; j < n; ++j) {
      double acc = 0.0;
      for (int l = 0; l < k; ++l) {
        float prod = a[(size_t)i * k + l] * b[(size_t)l * n + j];
        acc += prod;
      }
      out[(size_t)i * n + j] = (float)acc;
    }
  }
}

/* Same arithmetic, restricted to output rows [row0, row1): the bounded sample the CPU baseline
 * times at large shapes. */
void oracle_matmul_rows(const float* a, const float* b, float* out, int m, int k, int n, int row0,
                        int row1) {
  (void)m;
  for (int i = row0; i < row1; ++i) {
    for (int j = 0; j < n; ++j) {
      double acc = 0.0;
      for (int l = 0; l < k; ++l) {
        float prod = a[(size_t)i * k + l] * b[(size_t)l * n + j];
        acc += prod;
      }
      out[(size_t)i * n + j] = (float)acc;
    }
  }
}

/* sputnik/matrix_utils.cu:635-643 — Matrix::T(): out(j, i) = in(i, j). */
void oracle_transpose(const float* in, float* out, int rows, int cols) {
  for (int i = 0; i < rows; ++i)
    for (int j = 0; j < cols; ++j) out[(size_t)j * rows + i] = in[(size_t)i * cols + j];
}

/* sputnik/matrix_utils.h:364-374 — Matrix::Mul(): elementwise product (the SDD mask). */
void oracle_mul(float* inout, const float* other, size_t n) {
  for (size_t i = 0; i < n; ++i) inout[i] *= other[i];
}

/* sputnik/block/matrix_utils.h:81-112 — ToMatrix(): densify BCSR.  values is
 * [nnz_blocks][128][128]; block j of block row r lands at (r*128, indices[j]*128). */
void oracle_densify(const float* values, const int32_t* offsets, const int16_t* indices,
                    int block_rows, int block_cols, float* dense) {
  const size_t cols = (size_t)block_cols * BLK;
  memset(dense, 0, sizeof(float) * (size_t)block_rows * BLK * cols);
  for (int r = 0; r < block_rows; ++r) {
    for (int j = offsets[r]; j < offsets[r + 1]; ++j) {
      const int c = indices[j];
      const float* blk = values + (size_t)j * BLK * BLK;
      for (int y = 0; y < BLK; ++y)
        memcpy(dense + ((size_t)r * BLK + y) * cols + (size_t)c * BLK, blk + (size_t)y * BLK,
               sizeof(float) * BLK);
    }
  }
}

/* Inverse of densify at a topology: gather the blocks of `dense` (used for the SDD expectation,
 * sdd_test.cu:170-171,193: expected = (A*B) masked by the densified all-ones output). */
void oracle_sample(const float* dense, const int32_t* offsets, const int16_t* indices,
                   int block_rows, int block_cols, float* values) {
  const size_t cols = (size_t)block_cols * BLK;
  for (int r = 0; r < block_rows; ++r) {
    for (int j = offsets[r]; j < offsets[r + 1]; ++j) {
      const int c = indices[j];
      float* blk = values + (size_t)j * BLK * BLK;
      for (int y = 0; y < BLK; ++y)
        memcpy(blk + (size_t)y * BLK, dense + ((size_t)r * BLK + y) * cols + (size_t)c * BLK,
               sizeof(float) * BLK);
    }
  }
}

/* sputnik/block/row_indices/row_indices.cu:7-19. */
void oracle_row_indices(const int32_t* offsets, int block_rows, int16_t* row_indices) {
  for (int r = 0; r < block_rows; ++r)
    for (int j = offsets[r]; j < offsets[r + 1]; ++j) row_indices[j] = (int16_t)r;
}

/* sputnik/block/transpose/transpose.cu:69-125 — Transpose(): order = stable argsort of the
 * column indices (:11-19); indices_t = row_of_block[order] (:21-43); block_offsets = order;
 * offsets_t = exclusive cumsum of the column histogram (:51-66). */
typedef struct { int16_t key; int32_t pos; } keypos;
static void merge_sort(keypos* a, keypos* tmp, int n) { /* stable */
  if (n < 2) return;
  int h = n / 2;
  merge_sort(a, tmp, h);
  merge_sort(a + h, tmp, n - h);
  int i = 0, j = h, o = 0;
  while (i < h && j < n) tmp[o++] = (a[j].key < a[i].key) ? a[j++] : a[i++];
  while (i < h) tmp[o++] = a[i++];
  while (j < n) tmp[o++] = a[j++];
  memcpy(a, tmp, sizeof(keypos) * n);
}

void oracle_transpose_metadata(const int32_t* offsets, const int16_t* indices, int block_rows,
                               int block_cols, int32_t* offsets_t, int16_t* indices_t,
                               int32_t* block_offsets) {
  const int nb = offsets[block_rows];
  keypos* kp = (keypos*)malloc(sizeof(keypos) * (nb + 1) * 2);
  int16_t* rows = (int16_t*)malloc(sizeof(int16_t) * (nb + 1));
  for (int j = 0; j < nb; ++j) { kp[j].key = indices[j]; kp[j].pos = j; }
  merge_sort(kp, kp + nb + 1, nb);
  oracle_row_indices(offsets, block_rows, rows);
  for (int i = 0; i < nb; ++i) {
    indices_t[i] = rows[kp[i].pos];
    block_offsets[i] = kp[i].pos;
  }
  for (int c = 0; c <= block_cols; ++c) offsets_t[c] = 0;
  for (int j = 0; j < nb; ++j) offsets_t[indices[j] + 1]++;
  for (int c = 0; c < block_cols; ++c) offsets_t[c + 1] += offsets_t[c];
  free(kp);
  free(rows);
}

/* ------------------------------------------------------------------------------------------
 * Block-aware variants: the SAME arithmetic as oracle_matmul on the densified operands
 * (float product, double accumulation in ascending contraction order, one cast at the end) but
 * visiting nonzero blocks only — the skipped terms are exact zeros, so results are identical.
 * Needed because the dense host matmul is O(M*N*K).  OpenMP over independent
 * output rows / blocks.  `a` / `b` are the operands AS STORED; transpose flags follow
 * sputnik/block/arguments.h:164-215 (MatmulShape).
 * ------------------------------------------------------------------------------------------ */

typedef struct { int kb; int blk; } walk_item;
static int cmp_walk(const void* x, const void* y) {
  const walk_item* a = (const walk_item*)x; const walk_item* b = (const walk_item*)y;
  if (a->kb != b->kb) return a->kb - b->kb;
  return a->blk - b->blk;
}

/* For each block row (by_col = 0) or block column (by_col = 1) of the stored sparse matrix,
 * the list of (other coordinate, block id) sorted by that coordinate. */
static void build_walk(const int32_t* offsets, const int16_t* indices, int block_rows,
                       int block_cols, int by_col, int32_t* woff, walk_item* items) {
  const int nb = offsets[block_rows];
  const int nmaj = by_col ? block_cols : block_rows;
  for (int i = 0; i <= nmaj; ++i) woff[i] = 0;
  for (int r = 0; r < block_rows; ++r)
    for (int j = offsets[r]; j < offsets[r + 1]; ++j) woff[(by_col ? indices[j] : r) + 1]++;
  for (int i = 0; i < nmaj; ++i) woff[i + 1] += woff[i];
  int32_t* cur = (int32_t*)malloc(sizeof(int32_t) * (nmaj + 1));
  memcpy(cur, woff, sizeof(int32_t) * (nmaj + 1));
  for (int r = 0; r < block_rows; ++r)
    for (int j = offsets[r]; j < offsets[r + 1]; ++j) {
      const int maj = by_col ? indices[j] : r;
      items[cur[maj]].kb = by_col ? r : indices[j];
      items[cur[maj]].blk = j;
      cur[maj]++;
    }
  for (int i = 0; i < nmaj; ++i) qsort(items + woff[i], woff[i + 1] - woff[i], sizeof(walk_item), cmp_walk);
  (void)nb;
  free(cur);
}

static float* materialize_op(const float* x, int rows, int cols, int trans) {
  /* returns op(x) as a fresh row-major array; stored shape is rows x cols */
  float* out = (float*)malloc(sizeof(float) * (size_t)rows * cols + 16);
  if (trans) oracle_transpose(x, out, rows, cols);
  else memcpy(out, x, sizeof(float) * (size_t)rows * cols);
  return out;
}

/* DSD: C[m x n] = op(A sparse) * op(B).  A stored block_rows x block_cols blocks. */
void oracle_dsd_blocked(const float* values, const int32_t* offsets, const int16_t* indices,
                        int block_rows, int block_cols, int trans_a, const float* b, int trans_b,
                        float* out, int n) {
  const int nb = offsets[block_rows];
  const int mblocks = trans_a ? block_cols : block_rows;
  const int k = (trans_a ? block_rows : block_cols) * BLK;
  int32_t* woff = (int32_t*)malloc(sizeof(int32_t) * (mblocks + 1));
  walk_item* items = (walk_item*)malloc(sizeof(walk_item) * (nb + 1));
  build_walk(offsets, indices, block_rows, block_cols, trans_a, woff, items);
  float* bop = materialize_op(b, trans_b ? n : k, trans_b ? k : n, trans_b); /* k x n */
#pragma omp parallel
  {
    double* acc = (double*)malloc(sizeof(double) * (n + 1));
#pragma omp for schedule(dynamic, 8)
    for (int i = 0; i < mblocks * BLK; ++i) {
      const int R = i / BLK, y = i % BLK;
      for (int j = 0; j < n; ++j) acc[j] = 0.0;
      for (int w = woff[R]; w < woff[R + 1]; ++w) {
        const float* v = values + (size_t)items[w].blk * BLK * BLK;
        for (int x = 0; x < BLK; ++x) {
          const float av = trans_a ? v[x * BLK + y] : v[y * BLK + x];
          const float* brow = bop + ((size_t)items[w].kb * BLK + x) * n;
          for (int j = 0; j < n; ++j) { float p = av * brow[j]; acc[j] += p; }
        }
      }
      for (int j = 0; j < n; ++j) out[(size_t)i * n + j] = (float)acc[j];
    }
    free(acc);
  }
  free(bop); free(items); free(woff);
}

/* DDS: C[m x n] = op(A) * op(B sparse).  B stored block_rows x block_cols blocks. */
void oracle_dds_blocked(const float* a, int trans_a, int m, const float* values,
                        const int32_t* offsets, const int16_t* indices, int block_rows,
                        int block_cols, int trans_b, float* out) {
  const int nb = offsets[block_rows];
  const int nblocks = trans_b ? block_rows : block_cols;
  const int k = (trans_b ? block_cols : block_rows) * BLK;
  const int n = nblocks * BLK;
  int32_t* woff = (int32_t*)malloc(sizeof(int32_t) * (nblocks + 1));
  walk_item* items = (walk_item*)malloc(sizeof(walk_item) * (nb + 1));
  build_walk(offsets, indices, block_rows, block_cols, !trans_b, woff, items);
  float* aop = materialize_op(a, trans_a ? k : m, trans_a ? m : k, trans_a); /* m x k */
#pragma omp parallel for schedule(dynamic, 8)
  for (int i = 0; i < m; ++i) {
    double acc[BLK];
    for (int C = 0; C < nblocks; ++C) {
      for (int j = 0; j < BLK; ++j) acc[j] = 0.0;
      for (int w = woff[C]; w < woff[C + 1]; ++w) {
        const float* v = values + (size_t)items[w].blk * BLK * BLK;
        const float* arow = aop + (size_t)i * k + (size_t)items[w].kb * BLK;
        for (int x = 0; x < BLK; ++x) {
          const float av = arow[x];
          if (trans_b) { for (int j = 0; j < BLK; ++j) { float p = av * v[j * BLK + x]; acc[j] += p; } }
          else         { for (int j = 0; j < BLK; ++j) { float p = av * v[x * BLK + j]; acc[j] += p; } }
        }
      }
      for (int j = 0; j < BLK; ++j) out[(size_t)i * n + (size_t)C * BLK + j] = (float)acc[j];
    }
  }
  free(aop); free(items); free(woff);
}

/* SDD: C(sparse values [nb][128][128]) = op(A)[m x k] * op(B)[k x n] sampled at the topology. */
void oracle_sdd_blocked(const float* a, int trans_a, const float* b, int trans_b, int k,
                        const int32_t* offsets, const int16_t* indices, int block_rows,
                        int block_cols, float* values) {
  const int nb = offsets[block_rows];
  const int m = block_rows * BLK, n = block_cols * BLK;
  int16_t* rows = (int16_t*)malloc(sizeof(int16_t) * (nb + 1));
  oracle_row_indices(offsets, block_rows, rows);
  float* aop = materialize_op(a, trans_a ? k : m, trans_a ? m : k, trans_a); /* m x k */
  float* bop = materialize_op(b, trans_b ? n : k, trans_b ? k : n, trans_b); /* k x n */
#pragma omp parallel for schedule(dynamic, 16) collapse(2)
  for (int t = 0; t < nb; ++t) {
    for (int y = 0; y < BLK; ++y) {
      double acc[BLK];
      const float* arow = aop + ((size_t)rows[t] * BLK + y) * k;
      const float* bcol = bop + (size_t)indices[t] * BLK;
      for (int j = 0; j < BLK; ++j) acc[j] = 0.0;
      for (int l = 0; l < k; ++l) {
        const float av = arow[l];
        const float* brow = bcol + (size_t)l * n;
        for (int j = 0; j < BLK; ++j) { float p = av * brow[j]; acc[j] += p; }
      }
      float* o = values + ((size_t)t * BLK + y) * BLK;
      for (int j = 0; j < BLK; ++j) o[j] = (float)acc[j];
    }
  }
  free(aop); free(bop); free(rows);
}

/* Bitmask of a topology: sputnik/block/bitmask/bitmask.cu:7-45 filling a BitMatrix
 * (sputnik/block/bitmask/bit_matrix.h:12-62): uint64 words, [walk_rows][ceil(walk_cols / 64)],
 * Set(i, j): word[i * entries + j / 64] |= 1 << (j % 64).  The caller passes the view the
 * reference walks (offsets_t / indices_t when the transposed buffers exist, bitmask.cu:8-10). */
void oracle_bitmask(const int32_t* offsets, const int16_t* indices, int walk_rows, int walk_cols,
                    uint64_t* out) {
  const int entries = (walk_cols + 63) / 64;
  for (size_t i = 0; i < (size_t)entries * (size_t)walk_rows; ++i) out[i] = 0;
  for (int i = 0; i < walk_rows; ++i)
    for (int o = offsets[i]; o < offsets[i + 1]; ++o) {
      const int j = indices[o];
      out[(size_t)i * entries + j / 64] |= (uint64_t)1 << (j % 64);
    }
}

#ifdef _OPENMP
#include <omp.h>
int oracle_num_threads(void) { return omp_get_max_threads(); }
void oracle_set_threads(int n) { omp_set_num_threads(n); }
#else
int oracle_num_threads(void) { return 1; }
void oracle_set_threads(int n) { (void)n; }
#endif
