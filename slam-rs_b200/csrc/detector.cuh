// Detector-side shared declarations: device geometry block passed by value to every kernel, the
// handle, and the per-stage launchers (one .cu per stage group).
#pragma once
#include "common.cuh"

namespace orbx {

constexpr int kEdge = 31;        // edgeThreshold: keypoints live in [31, w-31) x [31, h-31)
constexpr int kFastTileW = 62;   // interior pixels per FAST tile (score region = 64 x 32)
constexpr int kFastTileH = 30;
constexpr int kBlurTileW = 128;  // one warp blurs a 128 x 35 tile (35 = 5 x the 7-row register window)
constexpr int kBlurTileH = 35;
constexpr int kResizeTileW = 128;
constexpr int kResizeTileH = 32;

// candidate word: score << 24 | y << 12 | x   (level coordinates, < 4096)
__host__ __device__ inline uint32_t pack_cand(int x, int y, int score) {
  return ((uint32_t)score << 24) | ((uint32_t)y << 12) | (uint32_t)x;
}

struct DevLevel {
  const uint8_t* img;   // un-blurred level, frame f at img + f*frame_stride
  uint8_t* blur;        // blurred level (same pitch / stride as the internal pyramid)
  size_t frame_stride;  // bytes
  size_t blur_stride;
  int w, h, pitch, blur_pitch;
  int aligned4;         // img base, pitch and stride are 4-byte aligned
  float scale;          // s_l
  int quota;            // N_l
  uint32_t* cand;       // [B][cand_cap]
  uint32_t* key;        // [B][cand_cap] order-preserving Harris key, 0 = not selected
  int cand_cap;
  unsigned long long* sel;  // [B][sel_cap] sorted (key << 32 | 0xFFFFFF - pos)
  int sel_cap;
  int fast_tiles_x, fast_tiles_y, fast_tile_begin;  // FAST tile table
  int blur_tiles_x, blur_tiles_y, blur_tile_begin;
};

struct DevGeom {
  DevLevel lv[kMaxLevels];
  int nlevels, batch;
  int fast_threshold;
  int fast_tiles_total, blur_tiles_total;
  const uint32_t* fast_tab;  // [fast_tiles_total] level | x0 << 4 | y0 << 16 (first interior pixel of the tile)
  uint32_t* cand_count;  // [B][8]
  uint32_t* hist;        // [B][8][256]
  uint32_t* sel_count;   // [B][8]
  uint32_t* sel_off;     // [B][8] offset of the level's sorted selection inside `spill`, or kSelInline
  int* flags;            // [0] != 0: the spill arena is exhausted (entries needed); [1] = spill cursor
  unsigned long long* spill;  // tie floods: selections larger than sel_cap / the shared-memory sort live here
  uint32_t spill_cap;    // entries
  int* sticky;           // survives the per-batch reset: [0] arena exhausted, [1] largest per-frame count that did not
                         // fit the caller's cap on the asynchronous path (read and cleared by orbx_detector_sync)
};
constexpr uint32_t kSelInline = 0xFFFFFFFFu;

struct OutArgs {
  orbx_keypoint* kps;  // [B][cap]
  uint8_t* desc;       // [B][cap][32] or null
  uint32_t* kp_lvl;    // [B][lvl_cap]  level << 24 | y << 12 | x
  int* n;              // [B]
  int cap, lvl_cap;
  int clamp_n;         // asynchronous path: n[frame] = min(total, cap), the overflow goes to DevGeom::sticky[1]
};

struct ResizeArgs {
  const uint8_t* src;
  uint8_t* dst;
  size_t src_stride, dst_stride;
  int sw, sh, spitch, dw, dh, dpitch;
  const uint32_t* xtab;  // [dw]  ofs << 16 | c1
  const uint32_t* ytab;  // [dh]
  int src_align;       // 16, 4 or 1: common alignment of the source base, pitch and frame stride
  int smem_w, smem_h;  // staged source region (bytes, rows)
};

// stage launchers -------------------------------------------------------------------------------
int launch_bgr_to_gray(const uint8_t* src, size_t spitch, size_t sstride, uint8_t* dst, int dpitch, size_t dstride,
                       int w, int h, int batch, cudaStream_t s);
int launch_repack_gray(const uint8_t* src, size_t spitch, size_t sstride, uint8_t* dst, int dpitch, size_t dstride, int w,
                       int h, int batch, cudaStream_t s);
int launch_resize(const ResizeArgs& a, int batch, cudaStream_t s);
int launch_fast(const DevGeom& g, cudaStream_t s);
int launch_harris(const DevGeom& g, cudaStream_t s);
int launch_select(const DevGeom& g, int sort_cap, cudaStream_t s);
int launch_orient(const DevGeom& g, const OutArgs& o, cudaStream_t s);
int launch_blur(const DevGeom& g, cudaStream_t s);
int launch_brief(const DevGeom& g, const OutArgs& o, cudaStream_t s);

}  // namespace orbx
