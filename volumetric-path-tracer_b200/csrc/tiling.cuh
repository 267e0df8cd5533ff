// tiling.cuh — thread -> pixel mapping. A warp covers an 8x4 pixel tile so that
// its 32 camera rays (and most first bounces) walk the same BVH nodes.
#pragma once

#include "vmath.cuh"

namespace vpt {

VPT_FN int tile_slot_count(int width, int height) {
  return ((width + 7) / 8) * ((height + 3) / 4) * 32;
}
// pixel index of thread `tid`, or -1 if the tile hangs over the image edge
VPT_FN int tile_pixel_index(int tid, int width, int height) {
  int tiles_x = (width + 7) / 8;
  int tile = tid >> 5, lane = tid & 31;
  int px = (tile % tiles_x) * 8 + (lane & 7);
  int py = (tile / tiles_x) * 4 + (lane >> 3);
  if (px >= width || py >= height) return -1;
  return py * width + px;
}

}  // namespace vpt
