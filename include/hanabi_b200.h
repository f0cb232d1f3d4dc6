/*
 * hanabi_b200.h -- C ABI of the B200-native hot path of vsois/hanabi-agents.
 *
 * Plain pointers and sizes only (no torch / pybind types).  Every entry point
 * names the reference interface it replaces (path:line under /root/reference).
 *
 * Conventions
 *   - all functions return HB_OK (0) or a negative HB_ERR_* code;
 *     hb_last_error() gives a human readable message for the calling thread.
 *   - pointers named d_* are DEVICE pointers, h_* are HOST pointers.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *     Device-pointer entry points are stream ordered and do not synchronise.
 *   - *_host entry points take host buffers, copy in, run, copy out and
 *     synchronise: they are what a pybind11 / ctypes drop-in binds.
 *   - there is NO CPU fallback: without a usable CUDA device every compute
 *     entry point fails with HB_ERR_CUDA.
 */
#ifndef HANABI_B200_H_
#define HANABI_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HB_OK 0
#define HB_ERR_INVALID (-1) /* bad argument (NULL, negative size, bad config) */
#define HB_ERR_CUDA (-2)    /* CUDA runtime / driver failure, no device        */
#define HB_ERR_INDEX (-3)   /* index out of range (maps to Python IndexError)  */
#define HB_ERR_STATE (-4)   /* call not valid in the object's current state    */

const char* hb_last_error(void);
/* Library version and the SM architecture the kernels were compiled for. */
int hb_version(void);
const char* hb_build_arch(void);

/* ------------------------------------------------------------------------ */
/* Sum tree  (replaces sum_tree/sum_tree/include/sum_tree.h:22-130 and the   */
/* pybind module sum_tree/sum_tree/src/sum_tree_py.cc:7-27)                  */
/* ------------------------------------------------------------------------ */
typedef struct hb_sumtree hb_sumtree;

/* SumTree<float>::SumTree(capacity)  sum_tree.h:25-30 (+populateTree :107-123).
 * depth = ceil(log2(capacity)); 2^depth zero leaves. */
int hb_st_create(int64_t capacity, int device, hb_sumtree** out);
int hb_st_destroy(hb_sumtree* t);
/* getCapacity()  sum_tree.h:76  -- the rounded-up power of two. */
int64_t hb_st_capacity(const hb_sumtree* t);
int hb_st_depth(const hb_sumtree* t);
/* Device pointer to the flat 1-indexed heap (2*capacity floats; [0] unused,
 * [1] root, leaves at [capacity, 2*capacity)).  For tests and fused callers. */
float* hb_st_heap(hb_sumtree* t);

/* updateValues(indices, values)  sum_tree.h:38-44 -> updateValue :32-35 ->
 * updateValue_ :80-90, with the sequential (OMP_NUM_THREADS=1) semantics:
 * elements applied in batch order, diff = value - leaf, every node on the path
 * `+= diff` in fp32.  Out-of-range indices set the sticky error flag and are
 * skipped (the reference segfaults there). */
int hb_st_update(hb_sumtree* t, const int64_t* d_idx, const float* d_val, int64_t n, void* stream);
/* Same, for the contiguous ascending leaf range [start, start+n) -- the shape
 * PriorityBuffer.add_transitions produces (priority_buffer.py:29-31).
 * d_val == NULL means every element takes `fill`. */
int hb_st_update_range(hb_sumtree* t, int64_t start, int64_t n, const float* d_val, float fill,
                       void* stream);
/* getIndices(quantiles)  sum_tree.h:52-59 -> getIndex :46-50 -> getIndex_ :92-105 */
int hb_st_get_indices(hb_sumtree* t, const float* d_q, int64_t* d_out, int64_t n, void* stream);
/* getValues(indices)  sum_tree.h:65-72 (.at() bounds check -> sticky error flag) */
int hb_st_get_values(hb_sumtree* t, const int64_t* d_idx, float* d_out, int64_t n, void* stream);
/* getTotalVal()  sum_tree.h:74 */
int hb_st_total(hb_sumtree* t, float* d_out, void* stream);
/* get_indices + get_values + get_total_val in one launch: the three calls of
 * PriorityBuffer.sample_batch  priority_buffer.py:41-42.  d_total may be NULL. */
int hb_st_sample(hb_sumtree* t, const float* d_q, int64_t* d_idx_out, float* d_val_out,
                 float* d_total, int64_t n, void* stream);
/* Reads and clears the sticky out-of-range flag (synchronises `stream`).
 * Returns HB_OK or HB_ERR_INDEX. */
int hb_st_check(hb_sumtree* t, void* stream);

/* Host-buffer forms: exactly the pybind surface of sum_tree_py.cc:11-18. */
int hb_st_update_host(hb_sumtree* t, const int64_t* h_idx, const float* h_val, int64_t n);
int hb_st_get_indices_host(hb_sumtree* t, const float* h_q, int64_t* h_out, int64_t n);
int hb_st_get_values_host(hb_sumtree* t, const int64_t* h_idx, float* h_out, int64_t n);
int hb_st_total_host(hb_sumtree* t, float* h_out);
/* Copies the whole heap (2*capacity floats) to the host: tests compare every
 * node value with the oracle. */
int hb_st_heap_host(hb_sumtree* t, float* h_out);

/* @@ENV@@ */

#ifdef __cplusplus
}
#endif
#endif /* HANABI_B200_H_ */
