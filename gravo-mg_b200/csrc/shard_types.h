// Definition of the opaque row-sharded build handle, shared by shard.cu and shard_solver.cu.
#pragma once

#include "hierarchy_types.h"

// Levels >= 1 with at most this many fine vertices keep the gathered P whole (36 B per row): the row-sharded
// V-cycle can then run them on every rank without any halo exchange.
constexpr int64_t kKeepWholeP = 4000000;

struct gmg_shard {
    gmg_ctx* ctx = nullptr;
    int rank = 0, world = 1;
    gmg_hierarchy_options opt{};
    std::vector<double> radius_override;
    // current level (replicated)
    int level = 0;
    int32_t n = 0;
    int64_t nnz = 0;
    int32_t lo = 0, hi = 0;          // owned fine vertices
    int32_t chunk = 0;               // slice length of the per-vertex exchange buffers (n_pad = chunk * world)
    const double* pts = nullptr;
    const int32_t* colptr = nullptr;
    const int32_t* rowidx = nullptr;
    const double* val = nullptr;     // stored values of the level's edge matrix
    DevBuf<double> len;              // Euclidean lengths (level 0: also the stored values)
    DevBuf<uint8_t> need;            // world > 1: columns whose lengths this rank computes
    bool val_is_len = false;
    int32_t gchunk = 0;              // column groups (32 columns) per rank
    DevBuf<double> len_partial;      // exchange: sum of the lengths of every column group
    double radius = 0.0;
    // exchange buffers (per-vertex, chunk * world entries; a rank writes [lo, hi) only)
    DevBuf<uint8_t> state, changed[2];
    DevBuf<unsigned long long> dist;
    DevBuf<int32_t> srcid, pred, parents;
    int parity = 0;
    ShardSampler sampler;
    // coarse side
    int32_t m = 0, clo = 0, chi = 0, cchunk = 0;
    DevBuf<int32_t> samples;
    Clusters clusters;
    DevBuf<int32_t> ccnt;            // exchange: distinct neighbours per coarse row, cchunk * world (+1)
    CoarseRows rows;
    DevBuf<int32_t> rowoff;          // exclusive scan of ccnt over ALL rows (m + 1) = global row-major offsets
    // Variable-size exchanges (packed coarse rows, rows of P, outer vectors of A_c): the exchange buffer IS the
    // final row-major array; rank r owns elements [v_off[r], v_off[r + 1]) of it (gmg_shard_buffer_v).
    DevBuf<int32_t> xQ;              // exchange: packed coarse rows (neighbour ids), all ranks
    DevBuf<double> xt;               // exchange: packed coarse rows (values)
    std::vector<int64_t> rows_off;   // world + 1 element offsets of xQ / xt, or of A_c's arrays while those are exchanged
    bool rows_are_ac = false;        // ROWS_ID / ROWS_VAL currently name the arrays of A_c
    StageStatus status;
    DevBuf<int64_t> xstatus;         // exchange: 8 status words per rank (LEVEL_STATUS / LEVEL_END)
    int64_t status_host[8] = {0};
    // Galerkin side (build_galerkin): the level's system matrix (replicated, pattern = any) and the exchange
    int32_t a_n = 0;
    const int32_t* a_colptr = nullptr;
    const int32_t* a_rowidx = nullptr;
    const double* a_val = nullptr;
    DevBuf<int32_t> pcnt;            // exchange: entries per row of P, chunk * world (+1)
    DeviceCsr Pfull;                 // all rows of P (transient); colidx / val are exchange buffers (P_COL / P_VAL)
    std::vector<int64_t> p_off;      // world + 1 element offsets of the ranks' rows in Pfull.colidx / val
    DeviceCsc ac_own;                // the owned outer vectors of A_c, indexed from clo
    // finished levels: replicated coarse data + the owned rows of P
    struct Done {
        gmg_level_info info{};
        int32_t lo = 0, hi = 0;
        DevBuf<int32_t> samples, parents;
        DeviceCsc coarse;
        DevBuf<double> coarse_pts;
        Triangles tri;
        DeviceCsr P;                 // rows [lo, hi)
        int32_t clo = 0, chi = 0;    // owned coarse vertices
        int32_t chunk = 0, cchunk = 0;   // slice lengths of the fine / coarse splits (owner of vertex j = j / chunk)
        DeviceCsr R;                 // build_galerkin: rows [clo, chi) of P^T (restriction of the owned coarse rows)
        DeviceCsr Pfull;             // build_galerkin, small levels (see kKeepWholeP): ALL rows of P, for a solver
                                     // that finishes the small levels redundantly on every rank
        DeviceCsc Ac;                // build_galerkin: A_c, replicated
    };
    std::vector<std::unique_ptr<Done>> done;
    std::unique_ptr<Done> cur;
};

