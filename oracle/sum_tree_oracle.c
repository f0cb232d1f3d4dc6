/*
 * TEST INFRASTRUCTURE -- not part of the product.
 *
 * CPU restatement of the reference sum tree, used only by tests/, by
 * __graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference
 * legs as the checker / reported CPU baseline.  Nothing under
 * hanabi_agents_b200/ may import, link or execute it.
 *
 * Follows /root/reference/sum_tree/sum_tree/include/sum_tree.h with the
 * sequential (one OpenMP thread) order of evaluation.  The reference's pointer
 * tree is stored here as a flat 1-indexed heap (root at [1], leaves at
 * [cap, 2cap)); the arithmetic on every node is the reference's.
 *
 * Parity is PINNED: tests/test_sum_tree_oracle.py checks this file against the
 * known answers of sum_tree/src/test_sum_tree.cc:45-117, against the heap
 * vectors of tests/rlax_dqn/test_sum_tree.py:26,38,80-89, and node-by-node
 * against the reference header itself compiled into oracle/_ref.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
  int depth;
  int64_t cap;
  float* heap;
} so_tree;

/* SumTree(capacity): depth_ = ceil(log2(capacity)); populateTree  sum_tree.h:25-30,107-123 */
so_tree* so_create(int64_t capacity) {
  if (capacity < 1) return NULL;
  so_tree* t = (so_tree*)malloc(sizeof(so_tree));
  if (!t) return NULL;
  t->depth = (int)ceil(log2((double)capacity));
  t->cap = (int64_t)1 << t->depth;
  t->heap = (float*)calloc((size_t)(2 * t->cap), sizeof(float));
  if (!t->heap) { free(t); return NULL; }
  return t;
}

void so_destroy(so_tree* t) {
  if (!t) return;
  free(t->heap);
  free(t);
}

int64_t so_capacity(const so_tree* t) { return t->cap; }       /* sum_tree.h:76 */
float so_total(const so_tree* t) { return t->heap[1]; }        /* sum_tree.h:74 */
const float* so_heap(const so_tree* t) { return t->heap; }

/* updateValue + updateValue_  sum_tree.h:32-35, 80-90.  Returns -1 if out of range. */
int so_update_value(so_tree* t, int64_t index, float value) {
  if (index < 0 || index >= t->cap) return -1;
  int64_t node = t->cap + index;
  volatile float diff = value - t->heap[node]; /* :33 */
  for (;;) {
    volatile float s = t->heap[node] + diff;   /* :84 node->value += diff */
    t->heap[node] = s;
    if (node == 1) break;                      /* :86-88 root reached */
    node >>= 1;                                /* :89 parent */
  }
  return 0;
}

/* updateValues, one thread: batch order  sum_tree.h:38-44 */
int so_update_values(so_tree* t, const int64_t* idx, const float* val, int64_t n) {
  int rc = 0;
  for (int64_t k = 0; k < n; ++k)
    if (so_update_value(t, idx[k], val[k]) != 0) rc = -1;
  return rc;
}

/* getIndex + getIndex_  sum_tree.h:46-50, 92-105 */
int64_t so_get_index(const so_tree* t, float quantile) {
  volatile float q = quantile * t->heap[1]; /* :47 */
  int64_t node = 1;
  for (int l = 0; l < t->depth; ++l) {
    const int64_t left = node << 1;
    const float lv = t->heap[left];
    if (q < lv) {         /* :99 */
      node = left;
    } else {              /* :104 */
      q = q - lv;
      node = left + 1;
    }
  }
  return node - t->cap;
}

void so_get_indices(const so_tree* t, const float* q, int64_t* out, int64_t n) {
  for (int64_t k = 0; k < n; ++k) out[k] = so_get_index(t, q[k]);
}

/* getValue(s)  sum_tree.h:61-72 ; returns -1 when .at() would throw */
int so_get_values(const so_tree* t, const int64_t* idx, float* out, int64_t n) {
  for (int64_t k = 0; k < n; ++k) {
    if (idx[k] < 0 || idx[k] >= t->cap) return -1;
    out[k] = t->heap[t->cap + idx[k]];
  }
  return 0;
}
