// plan.cuh -- the transform plan shared by tree.cu (builder) and transform.cu (executor).
//
// The RAHT tree is the max-Cartesian tree of level[]: row r >= 1 is merged, as the
// right sibling, at bit level level[r], with the nearest earlier row s whose level is higher (or
// row 0); the weights are w0 = r - s and w1 = t - r, t the nearest later row with a higher level
// (or n).  A butterfly can run inside a tile of consecutive entries iff s and t lie in the tile.
// Stage 0 tiles the rows themselves; the entries whose butterfly does not fit ("deferred", plus the
// root) are compacted -- order preserved, so they form the same kind of tree -- into the next stage.
#pragma once
#include "common.cuh"

constexpr int PLAN_MAX_STAGES = 40;
constexpr int PLAN_MAX_ROUNDS = 64;

struct __align__(8) TileHdr {
    uint16_t n_rows;                        // entries in this tile
    uint16_t n_bf;                          // butterflies scheduled in this tile
    uint16_t n_rounds;                      // dependency rounds (distinct bit levels present)
    uint16_t pad;
    uint16_t round_end[PLAN_MAX_ROUNDS];    // exclusive end of each round in the schedule
};
static_assert(sizeof(TileHdr) == 136, "TileHdr layout");

// schedule entry: bits 0-9 local index of the right row r, 10-19 local index of the left row s,
// 20-25 bit level.  ab = (a, b) = (sqrt(w0/(w0+w1)), sqrt(w1/(w0+w1)))  (RAHT.py:321-322).
struct PlanStage {
    int64_t n_entries;
    int64_t n_tiles;
    uint32_t *rows;      // original row of each entry; nullptr in stage 0 (identity)
    uint8_t *to_next;    // 1 = entry is carried into the next stage
    uint32_t *sched;     // n_tiles * tile_rows
    double2 *ab;         // n_tiles * tile_rows
    TileHdr *hdr;        // n_tiles
};

struct raht_plan {
    int64_t n;
    int tile_rows;
    int n_stages;
    size_t bytes;
    PlanStage st[PLAN_MAX_STAGES];
};
