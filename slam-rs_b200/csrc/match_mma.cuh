// Tensor-core (tcgen05 kind::i8) Hamming matcher: declarations shared by match.cu and match_mma.cu.
#pragma once
#include "common.cuh"

namespace orbx {

constexpr int kMmaRowBytes = 320;  // expanded int8 row: 256 bits as +-16, 32 aux bytes per role
constexpr int kMmaTileN = 256;     // train rows per accumulator tile; expanded sets are padded to a multiple of it

struct MmaMatchArgs {
  const uint8_t* ea;  // expanded query sets: problem p at ea + p * ea_stride
  const uint8_t* eb;  // expanded train sets
  size_t ea_stride, eb_stride;
  const int* nq_dev;  // per-problem row counts (device) or null -> *_fixed
  const int* nt_dev;
  int nq_fixed, nt_fixed;
  int max_rows, max_rows_t;  // clamps for the device counts; max_rows is also the row pitch of `part`
  int npairs, mtiles;        // 128-row query tiles per problem
  int nsplits, tiles_per_split;  // the train tiles of a problem are cut into nsplits ranges (large train sets)
  int K;                         // 1 or 2 keys per (problem, split, query row)
  unsigned long long* part;      // [problem][split][max_rows][K] packed (dist << 32 | train_idx), ~0 = none
  unsigned long long* rev;       // optional [problem][rev_rows] reverse best-1 (dist << 32 | query_idx), preset ~0
  int rev_rows;
};

size_t mma_expanded_bytes(int rows_padded);  // bytes of one expanded set of rows_padded (multiple of 256) rows
// desc sets (32-byte rows, set s at desc + s * desc_stride, n_dev[s] or n_fixed valid rows) -> expanded int8 sets
int launch_expand_desc(const uint8_t* desc, size_t desc_stride, const int* n_dev, int n_fixed, int max_rows, int nsets,
                       int rows_padded, uint8_t* out, size_t out_stride, cudaStream_t s);
int launch_hamming_mma(const MmaMatchArgs& a, int num_sms, cudaStream_t s);
// true: the launch runs as CTA pairs that share their train tiles; a work unit then covers TWO query tiles and there are
// num_sms / 2 workers (the caller's choice of nsplits goes by that)
bool mma_use_pairs(int mtiles, bool reverse);

}  // namespace orbx
