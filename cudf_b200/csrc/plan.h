// Host-side plan of one groupby call: validation, key packing layout, accumulator set, output recipes.
// Mirrors the host logic of the reference path:
//   verify_valid_requests / size check          /root/reference/cpp/src/groupby/groupby.cu:186-229
//   can_use_hash_groupby                        /root/reference/cpp/src/groupby/hash/groupby.cu:63-131
//   extract_single_pass_aggs (decomposition)    /root/reference/cpp/src/groupby/hash/extract_single_pass_aggs.cpp:38-177
//   result_column_creator (nullability)         /root/reference/cpp/src/groupby/hash/output_utils.cu:49-113
#pragma once
#include <string>
#include <vector>

#include "bgb_internal.cuh"

namespace bgb {

struct KeyPlan {
  bgb_column_view v;
  uint8_t width, fclass, shift;
  int8_t null_bit;
  bool nullable_out;
  // range-compressed image (key repacking, groupby.cu): the field holds (value - base) in `bits` bits; bits == 0: the element itself
  uint64_t base = 0;
  uint8_t bits  = 0;
};

struct ValuePlan {
  bgb_column_view v;      // DICTIONARY32 values: the view of the INDICES child under the parent's offset / mask, typed like the KEYS child
  const uint8_t* dict = nullptr;  // DICTIONARY32: first element of the keys child (the dictionary), else nullptr
  uint8_t idx_width   = 0;        // DICTIONARY32: bytes of one index
  uint8_t src;
  bool has_nulls;   // column_view::has_nulls() == null_count > 0
  bool has_mask;    // nullable()
};

struct AccPlan {
  int value_col;    // index into values, -1 for COUNT_ALL
  uint8_t op;
  void* cells = nullptr;
  int aux_acc = -1;  // ACC_ARGROW_*: index of the MIN / MAX accumulator it reads in the second pass
};

struct OutPlan {
  int request, agg;
  int32_t type_id, scale;
  bool nullable;
  uint8_t op;       // OutOp
  uint8_t src;
  uint8_t out_width;
  uint8_t acc_op;
  int acc_a = -1, acc_b = -1, acc_cnt = -1;
  int32_t ddof = 1;
};

struct Plan {
  int64_t num_rows = 0;
  bool include_null_keys = false;
  bool skip_null_keys = false;
  int key_words = 1;
  bool has_dict = false;                // some value column is a dictionary: generic kernels only (the element is keys[indices[row]])
  bool wide = false;                    // key image wider than 128 bits: row-number slots, generic kernels only (groupby.cu)
  std::vector<KeyPlan> keys;
  std::vector<ValuePlan> values;
  std::vector<AccPlan> accs;
  std::vector<OutPlan> outs;            // request-major
  std::vector<int> request_first_out;   // outs index of (request r, agg 0)
  std::vector<int> request_num_aggs;
};

int type_width(int32_t type_id);                 // 0 when not fixed-width <= 8 bytes
uint8_t type_src(int32_t type_id);               // Src or S_BAD
int32_t target_type(int32_t type_id, int32_t kind);  // -1 invalid

// status: BGB_OK / BGB_INVALID_ARGUMENT / BGB_UNSUPPORTED (message via set_error)
// rank_independent: the key image layout and the accumulator set must not depend on which columns happen to carry a null
// mask on THIS rank (multi-GPU exchange: every rank packs keys and lays out records identically); every value column is
// then treated as nullable and, under INCLUDE, every key column gets a null flag bit.
int make_plan(const bgb_column_view* keys, int32_t num_keys, int32_t include_null_keys, const bgb_request* requests,
              int32_t num_requests, Plan& plan, bool rank_independent = false, bool allow_wide = false);
// allow_wide: key images wider than 128 bits are planned anyway (key_words > 2): only valid as the input of key repacking

}  // namespace bgb
