#include "tiling.h"

#include <algorithm>
#include <cmath>
#include <stdexcept>

namespace ab {

int find_starting_center(int full_dimension, int tile_dimension, int overlap)
{
    if (tile_dimension <= overlap) throw std::invalid_argument("tile dimension must exceed the overlap");
    const int stride = tile_dimension - overlap;
    int c = full_dimension / 2;
    while (c - stride >= 0) c -= stride;
    if (c > tile_dimension / 2) c -= stride / 2;
    return c;
}

std::vector<Tile> get_tiles(int width, int height, int max_w, int max_h, int ov_x, int ov_y)
{
    const int sx = find_starting_center(width, max_w, ov_x), sy = find_starting_center(height, max_h, ov_y);
    const int stride_x = max_w - ov_x, stride_y = max_h - ov_y;
    std::vector<Tile> tiles;
    for (int cy = sy; cy < height; cy += stride_y) {
        const bool top_row = cy == sy, bottom_row = cy + stride_y >= height;
        for (int cx = sx; cx < width; cx += stride_x) {
            const bool left_col = cx == sx, right_col = cx + stride_x >= width;
            const int dl = cx - max_w / 2, dt = cy - max_h / 2;
            const int left = std::max(0, dl), top = std::max(0, dt);
            const int right = std::min(dl + max_w, width), bottom = std::min(dt + max_h, height);  // exclusive
            Tile t;
            t.full[0] = left; t.full[1] = top; t.full[2] = right - 1; t.full[3] = bottom - 1;
            t.non_overlapping[0] = left_col ? left : left + (ov_x + 1) / 2;
            t.non_overlapping[1] = top_row ? top : top + (ov_y + 1) / 2;
            t.non_overlapping[2] = (right_col ? right : right - ov_x / 2) - 1;
            t.non_overlapping[3] = (bottom_row ? bottom : bottom - ov_y / 2) - 1;
            t.unique[0] = left_col ? left : left + ov_x;
            t.unique[1] = top_row ? top : top + ov_y;
            t.unique[2] = (right_col ? right : right - ov_x) - 1;
            t.unique[3] = (bottom_row ? bottom : bottom - ov_y) - 1;
            tiles.push_back(t);
        }
    }
    return tiles;
}

void set_weights(const uint16_t* labels, int h, int w, double class_weight, double image_weight, float* out)
{
    const size_t n = (size_t)h * w;
    std::vector<size_t> counts;
    // the histogram grows as index*2+16 (annonet_train.h:30-35) and its allocated size enters average_count
    for (size_t i = 0; i < n; ++i) {
        const uint16_t l = labels[i];
        if (l == 65535) continue;
        if (l >= counts.size()) counts.resize((size_t)l * 2 + 16);
        ++counts[l];
    }
    size_t total = 0;
    for (size_t c : counts) total += c;
    std::vector<double> lw(counts.size(), 0.0);
    if (total > 0) {
        const double average = total / (double)counts.size();
        double total_unnorm = 0.0;
        for (size_t l = 0; l < counts.size(); ++l)
            if (counts[l] > 0) {
                lw[l] = std::pow(average / counts[l], class_weight);
                total_unnorm += counts[l] * lw[l];
            }
        const double target = total * std::pow((double)h * w / (double)total, image_weight);
        for (double& v : lw) v *= target / total_unnorm;
    }
    for (size_t i = 0; i < n; ++i) out[i] = labels[i] == 65535 ? 0.f : (float)lw[labels[i]];
}

}  // namespace ab
