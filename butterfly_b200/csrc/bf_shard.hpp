// bf_shard.hpp — block-row sharding of a flattened operator (host only): ownership of leaves (bf_shard.cpp).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "bf_cuda.h"

namespace bfc {

constexpr int kMaxShardRanks = 16; // one bit per rank in the 16-bit reader masks of the plan compiler

struct ShardedOps {
  int world = 1;
  std::vector<BfCudaRawOp> ops;     // the operator with tall / diagonal leaves cut at ownership boundaries
  std::vector<int32_t> owner;       // [ops] rank that applies the leaf
  std::vector<int64_t> yCuts;       // [world + 1] rank r owns rows [yCuts[r], yCuts[r + 1]) of the result
  std::vector<int64_t> originOp;    // [ops] index of the leaf in the ORIGINAL op list ...
  std::vector<int64_t> originRow;   // ... and the first acting row of it this (slab of a) leaf holds
};

// Decide who applies what.  fractions: NULL (equal shares) or `world` relative shares of Y.
BfCudaStatus shardOps(const BfCudaPlanDesc &desc, int world, const double *fractions, ShardedOps &out, std::string &err);
// Rank `rank`'s op list: the leaves of other ranks carry BF_CUDA_LEAF_REMOTE (and no values), every leaf its owner
// in flag bits BF_CUDA_LEAF_OWNER_SHIFT ...
void maskForRank(const ShardedOps &S, int rank, std::vector<BfCudaRawOp> &out);

} // namespace bfc
