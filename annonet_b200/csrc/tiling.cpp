#include "tiling.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
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


namespace {
// smallest size >= minimum that survives `levels` pool/up-conv round trips: 2^levels * k + (2^levels - 1) (NetPimpl.cpp:246-285)
int band_input_dimension(int minimum, int levels)
{
    const int m = 1 << levels;
    int k = 1;
    while (m * k + m - 1 < minimum) ++k;
    return m * k + m - 1;
}
}  // namespace

// Tiles arrive row by row (tiling.cpp:57-83).  Each rank gets a contiguous run of them -- a BAND -- of near-equal COST (the pixels
// the net actually runs: every tile is enlarged to 64k + 63); neighbouring bands share a seam, so a band talks to at most two
// others.  Every image row is owned (argmax) by exactly one rank.
//
// Tile granularity (the default where it applies): a band may begin and end in the middle of a tile row.  A split row is OWNED
// by the rank that holds its first tile; the other rank sends what its tiles of that row added -- the whole row's strip, not just
// the 127-row overlap -- to the owner through the same strip exchange.  That is a few hundred MB over NVLink against whole
// tiles of compute, and it takes the balance from "19 rows over 8 ranks = 3 rows on the busiest against 2.375 on average" to
// within one tile.  The exchange stays with the two neighbours as long as every band contains at least one whole tile row
// (then the rows above and below a split row are owned by the two ranks that split it); where that does not hold -- few rows
// per rank -- or with AB_BAND_ROWS=1 the bands are whole rows as before.
static bool plan_bands_by_tile(const std::vector<Tile>& tiles, const std::vector<int>& row_start, const std::vector<double>& cost, int world,
                               std::vector<Band>& bands)
{
    const int rows = (int)row_start.size() - 1, ntiles = (int)tiles.size();
    if (world < 2 || ntiles < world) return false;
    double total = 0;
    for (double c : cost) total += c;
    std::vector<int> cut(world + 1, 0);
    cut[world] = ntiles;
    int j = 0;
    double done = 0;
    for (int r = 1; r < world; ++r) {   // the cut whose cumulative cost is closest to r / world of the total
        const double target = total * r / world;
        while (j < ntiles && std::abs(done + cost[j] - target) <= std::abs(done - target)) done += cost[j++];
        cut[r] = j;
    }
    auto row_of = [&](int tile) { return (int)(std::upper_bound(row_start.begin(), row_start.end(), tile) - row_start.begin()) - 1; };
    for (int r = 0; r < world; ++r) {
        if (cut[r + 1] <= cut[r]) return false;
        // first row that STARTS inside the band, and the row of its last tile
        const int first_row = row_of(cut[r]) + (row_start[row_of(cut[r])] == cut[r] ? 0 : 1), last_row = row_of(cut[r + 1] - 1);
        // a whole row inside the band: starts at or after the band's first tile and ends at or before its last
        bool whole = false;
        for (int q = first_row; q <= last_row && q < rows; ++q) whole = whole || (row_start[q] >= cut[r] && row_start[q + 1] <= cut[r + 1]);
        if (!whole || first_row > last_row) return false;
        Band& b = bands[r];
        b.t0 = cut[r]; b.t1 = cut[r + 1];
        b.own0 = tiles[row_start[first_row]].non_overlapping[1];      // rows it owns: from its first whole row ...
        b.own1 = tiles[b.t1 - 1].non_overlapping[3];                  // ... to the row of its last tile (a split row goes to its first holder)
        b.f0 = tiles[b.t0].full[1]; b.f1 = tiles[b.t1 - 1].full[3];
    }
    return true;
}

std::vector<Band> plan_bands(const std::vector<Tile>& tiles, int world, int levels)
{
    std::vector<int> row_start;   // first tile of each tile row
    std::vector<double> row_cost, tile_cost(tiles.size());
    for (size_t i = 0; i < tiles.size(); ++i) {
        if (i == 0 || tiles[i].full[1] != tiles[i - 1].full[1]) { row_start.push_back((int)i); row_cost.push_back(0); }
        const int fw = tiles[i].full[2] - tiles[i].full[0] + 1, fh = tiles[i].full[3] - tiles[i].full[1] + 1;
        tile_cost[i] = (double)band_input_dimension(fw, levels) * band_input_dimension(fh, levels);
        row_cost.back() += tile_cost[i];
    }
    const int rows = (int)row_start.size();
    row_start.push_back((int)tiles.size());
    double total = 0;
    for (double c : row_cost) total += c;
    std::vector<Band> bands(world);
    static const bool rows_only = getenv("AB_BAND_ROWS") && atoi(getenv("AB_BAND_ROWS")) != 0;
    if (!rows_only && plan_bands_by_tile(tiles, row_start, tile_cost, world, bands)) return bands;
    bands.assign(world, Band{});
    int j = 0;
    double done = 0;
    for (int r = 0; r < world; ++r) {
        const int ja = j;
        const double target = total * (r + 1) / world;
        // take rows while that brings the running cost closer to this rank's share of the total; leave enough rows for the rest
        while (j < rows && (rows - j > world - 1 - r) && (j == ja ? true : std::abs(done + row_cost[j] - target) <= std::abs(done - target))) done += row_cost[j++];
        if (r == world - 1) while (j < rows) done += row_cost[j++];
        if (j == ja && j < rows && rows - j >= world - r) done += row_cost[j++];
        Band& b = bands[r];
        b.t0 = row_start[ja]; b.t1 = row_start[j];
        if (b.empty()) continue;
        b.own0 = tiles[b.t0].non_overlapping[1]; b.own1 = tiles[b.t1 - 1].non_overlapping[3];
        b.f0 = tiles[b.t0].full[1]; b.f1 = tiles[b.t1 - 1].full[3];
    }
    return bands;
}

}  // namespace ab
