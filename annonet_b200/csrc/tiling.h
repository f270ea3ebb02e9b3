// Host-side helpers of the hot path: the overlapping-tile generator (tiling/tiling.cpp:7-87 with the
// dlib rectangle adaptor tiling/dlib-wrapper.h:18-39) and set_weights (annonet_train.h:20-83).
#pragma once
#include <cstdint>
#include <vector>

namespace ab {

struct Tile {
    int full[4];             // inclusive left, top, right, bottom
    int non_overlapping[4];
    int unique[4];           // area not covered by the full rect of any other tile
};

int find_starting_center(int full_dimension, int tile_dimension, int overlap);
std::vector<Tile> get_tiles(int width, int height, int max_tile_w, int max_tile_h, int overlap_x, int overlap_y);

// One rank's share of a sharded annonet_infer(): a contiguous band of tile ROWS.
struct Band {
    int t0 = 0, t1 = 0;        // tiles [t0, t1)
    int own0 = 0, own1 = -1;   // image rows this rank owns (union of its tiles' non-overlapping rectangles): it runs their argmax
    int f0 = 0, f1 = -1;       // image rows its tiles touch (union of the full rectangles): the extent of its private canvas
    bool empty() const { return t1 <= t0; }
};
std::vector<Band> plan_bands(const std::vector<Tile>& tiles, int world, int levels);

// weights[i] = 0 for ignored pixels (label 65535), else class weight (avg_count/count)^cw normalised so that
// the weights sum to total * (area/total)^iw
void set_weights(const uint16_t* labels, int h, int w, double class_weight, double image_weight, float* weights_out);

}  // namespace ab
