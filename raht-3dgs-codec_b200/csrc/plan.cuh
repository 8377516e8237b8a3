// plan.cuh -- the transform plan shared by tree.cu (builder) and transform.cu (executor).
//
// The RAHT tree is the max-Cartesian tree of level[]: row r >= 1 is merged, as the
// right sibling, at bit level level[r], with the nearest earlier row s whose level is higher (or
// row 0); the weights are w0 = r - s and w1 = t - r, t the nearest later row with a higher level
// (or n).  A butterfly can run inside a tile of consecutive pending entries iff s and t lie in it.
//
// STATIC hierarchy (no data-dependent sizes, hence no host synchronisation and no device allocation):
//   stage 0   tiles of tile_rows consecutive rows;
//   stage k   tile j takes, in order, the entries the PLAN_FANIN tiles 8j..8j+7 of stage k-1 could
//             not finish ("deferred": s or t outside the tile), until one tile is left.
// A tile defers at most PLAN_DEFER_MAX = 129 entries whatever its size: between two pending entries of
// equal level there is always a pending entry of a higher level, so the entries without an earlier
// higher one are strictly increasing in level (<= 64), those without a later higher one strictly
// decreasing (<= 64), plus the root.  Hence an upper tile holds at most 8 * 129 = 1032 entries and the
// number of stages, every grid and every buffer size follow from (n, tile_rows) alone.
#pragma once
#include "common.cuh"

constexpr int PLAN_FANIN = 8;
constexpr int PLAN_DEFER_MAX = 129;
constexpr int PLAN_UP_CAP = PLAN_FANIN * PLAN_DEFER_MAX;      // 1032 entries in an upper tile at most
constexpr int PLAN_MAX_STAGES = 16;
constexpr int PLAN_MAX_ROUNDS = 64;

struct __align__(8) TileHdr {
    uint16_t n_rows;                        // entries in this tile
    uint16_t n_bf;                          // butterflies scheduled in this tile
    uint16_t n_rounds;                      // dependency rounds (distinct bit levels present)
    uint16_t n_def;                         // entries deferred to the next stage
    uint32_t bf_off;                        // first slot of the tile's schedule in sched[] / ab[]
    uint32_t pad;
    uint16_t round_end[PLAN_MAX_ROUNDS];    // exclusive end of each round in the schedule
};
static_assert(sizeof(TileHdr) == 144, "TileHdr layout");

// schedule entry: bits 0-11 local index of the right row r, 12-23 local index of the left row s,
// 24-29 bit level.  ab = (a, b) = (sqrt(w0/(w0+w1)), sqrt(w1/(w0+w1)))  (RAHT.py:321-322).
__host__ __device__ inline uint32_t sched_pack(int r, int s, int l) { return (uint32_t)r | ((uint32_t)s << 12) | ((uint32_t)l << 24); }

// device view of a plan (all pointers into the caller's plan workspace)
struct PlanDev {
    int64_t n;
    int tile_rows;                          // stage-0 tile size
    int n_stages;
    int64_t tiles[PLAN_MAX_STAGES];         // tiles per stage
    int64_t tile_base[PLAN_MAX_STAGES];     // index of the stage's first tile among all tiles
    int64_t tiles_total;
    int64_t slots0;                         // tiles[0] * tile_rows: schedule slots of stage 0
    TileHdr *hdr;                           // [tiles_total]
    uint32_t *def_rows;                     // [tiles_total][129] original row of each deferred entry
    uint16_t *def_local;                    // [tiles_total][129] its index among the tile's entries
    uint8_t *def_lev;                       // [tiles_total][129] its level
    uint32_t *sched;                        // [slots0 + n]
    double2 *ab;                            // [slots0 + n]
    unsigned int *bump;                     // next free upper-stage schedule slot (starts at slots0)
    unsigned int *arrive;                   // [tiles_total] child tiles that have signed in (plan build, forward transform); zero between launches
    unsigned int *done;                     // [tiles_total] epoch of the inverse transform that finished the tile
    unsigned int *ticket;                   // next work item of the inverse transform; zero between launches
};

struct raht_plan {
    PlanDev d;
    size_t bytes;
};
