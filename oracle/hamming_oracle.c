/* TEST INFRASTRUCTURE ONLY — plain-C restatement of cv::batchDistance(NORM_HAMMING) + K-best insertion,
 * i.e. what BFMatcher::knnMatchImpl runs when called from /root/reference/src/feature/matcher.rs:34.
 * Semantics: scan train rows in ascending index, keep the K best sorted
 * ascending with STRICT comparisons, so equal distances keep ascending train index.
 * Built by oracle/Makefile into oracle/_build/libhamming_oracle.so; used only by tests/ as the checker
 * for cases too large for the numpy restatement. */
#include <stdint.h>
#include <string.h>

static inline int hamming256(const uint64_t* a, const uint64_t* b) {
  return __builtin_popcountll(a[0] ^ b[0]) + __builtin_popcountll(a[1] ^ b[1]) +
         __builtin_popcountll(a[2] ^ b[2]) + __builtin_popcountll(a[3] ^ b[3]);
}

/* idx, dist: nq x k (filled with -1 where fewer than k train rows exist) */
void hamming_knn(const uint8_t* q, int nq, const uint8_t* t, int nt, int k, int32_t* idx, int32_t* dist) {
  for (int i = 0; i < nq; ++i) {
    uint64_t qa[4];
    memcpy(qa, q + (size_t)i * 32, 32);
    int32_t* bi = idx + (size_t)i * k;
    int32_t* bd = dist + (size_t)i * k;
    int have = 0;
    for (int j = 0; j < k; ++j) bi[j] = bd[j] = -1;
    for (int j = 0; j < nt; ++j) {
      uint64_t ta[4];
      memcpy(ta, t + (size_t)j * 32, 32);
      const int d = hamming256(qa, ta);
      if (have == k && !(d < bd[k - 1])) continue; /* strict: a tie never displaces an earlier row */
      int pos = have < k ? have : k - 1;
      while (pos > 0 && bd[pos - 1] > d) {
        bd[pos] = bd[pos - 1];
        bi[pos] = bi[pos - 1];
        --pos;
      }
      bd[pos] = d;
      bi[pos] = j;
      if (have < k) ++have;
    }
  }
}
