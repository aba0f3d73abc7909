// Parameter-file layout helpers shared between translation units.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../include/b200gru.h"

namespace b200gru {
struct TensorSpec {
  std::string name;
  int64_t offset, size;
  int32_t rows, cols;
};
std::vector<TensorSpec> build_layout(const b200gru_config &c);
int validate_config(const b200gru_config *c);

// Index helpers into build_layout()'s vector (ref:375-407 order).
struct LayoutIndex {
  int L;
  explicit LayoutIndex(int layers) : L(layers) {}
  int emb() const { return 0; }
  int w_i(int layer, int gate) const { return 1 + 3 * layer + gate; }
  int w_h(int layer, int gate) const { return 1 + 3 * L + 3 * layer + gate; }
  int b_i(int layer, int gate) const { return 1 + 6 * L + 3 * layer + gate; }
  int b_h(int layer, int gate) const { return 1 + 9 * L + 3 * layer + gate; }
  int w_fc() const { return 1 + 12 * L; }
  int b_fc() const { return 2 + 12 * L; }
};
}  // namespace b200gru
