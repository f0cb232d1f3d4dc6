// TEST INFRASTRUCTURE -- not part of the product.
//
// C-ABI over the UNMODIFIED reference header, compiled from where it lies:
//   g++ -I/root/reference/sum_tree/sum_tree/include ref_sum_tree_shim.cc ...
// (recipe: oracle/build.py; output only into oracle/_ref/).  The header names
// SumTreeTestProxy<Precision> as a friend (sum_tree.h:8-9,129), which is how the
// reference's own gtest reads the private nodes (test_sum_tree.cc:7-20); this
// shim uses the same door to dump every node value into a flat 1-indexed heap
// so the CUDA tree and the C restatement can be compared node by node.
#include <cstdint>
#include <stdexcept>
#include <vector>
#include "sum_tree.h"

template <typename Precision>
class SumTreeTestProxy {
 public:
  explicit SumTreeTestProxy(size_t capacity) : st(capacity) {}
  std::shared_ptr<SumTreeNode<Precision>> getRoot() { return st.root_; }
  size_t depth() const { return st.depth_; }
  SumTree<Precision> st;
};

typedef SumTreeTestProxy<float> Ref;

static void dump(const std::shared_ptr<SumTreeNode<float>>& node, int64_t pos, float* heap) {
  heap[pos] = node->value;
  if (!node->left) return;
  dump(node->left, 2 * pos, heap);
  dump(node->right, 2 * pos + 1, heap);
}

extern "C" {

void* ref_st_create(int64_t capacity) { return new Ref((size_t)capacity); }
void ref_st_destroy(void* h) { delete (Ref*)h; }
int64_t ref_st_capacity(void* h) { return (int64_t)((Ref*)h)->st.getCapacity(); }
int ref_st_depth(void* h) { return (int)((Ref*)h)->depth(); }
float ref_st_total(void* h) { return ((Ref*)h)->st.getTotalVal(); }

void ref_st_update_value(void* h, int index, float value) { ((Ref*)h)->st.updateValue(index, value); }

void ref_st_update_values(void* h, const int64_t* idx, const float* val, int64_t n) {
  std::vector<size_t> i(idx, idx + n);
  std::vector<float> v(val, val + n);
  ((Ref*)h)->st.updateValues(i, v);
}

void ref_st_get_indices(void* h, const float* q, int64_t* out, int64_t n) {
  std::vector<float> qq(q, q + n);
  std::vector<size_t> r = ((Ref*)h)->st.getIndices(qq);
  for (int64_t k = 0; k < n; ++k) out[k] = (int64_t)r[k];
}

int ref_st_get_values(void* h, const int64_t* idx, float* out, int64_t n) {
  try {
    std::vector<size_t> i(idx, idx + n);
    std::vector<float> r = ((Ref*)h)->st.getValues(i);
    for (int64_t k = 0; k < n; ++k) out[k] = r[k];
  } catch (const std::out_of_range&) {
    return -1;
  }
  return 0;
}

// heap must hold 2*capacity floats; [0] is left untouched.
void ref_st_dump_heap(void* h, float* heap) { dump(((Ref*)h)->getRoot(), 1, heap); }

}  // extern "C"
