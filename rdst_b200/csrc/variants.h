// variants.h — the tile / occupancy variants of the K2 pass kernel.  In the
// reference the Tuner picks one of nine CPU algorithms per chunk
// (reference src/tuner.rs:10-35, src/tuners/standard_tuner.rs:10-64); here the
// only remaining choice is which tile shape of the one GPU pass kernel to run.
#pragma once

namespace b200 {

struct VariantInfo {
    const char* name;
    int threads;
    int items[4];  // records per thread for WORDS = 1, 2, 3, 4
    bool tma;      // stage tiles with bulk asynchronous copies (persistent, prefetching)
    bool match_hw; // MATCH.ANY instead of eight ballots
    int min_blocks;
    int algo;      // 0 = ballot-matching kernel (onesweep.cuh), 1 = OR-matching kernel (pass_or.cuh),
                   // 2 = half-warp OR-matching kernel (pass_or16.cuh)
    int sub;       // half-warp kernel: independent slot tables per half-warp (0 means 1)
};

constexpr int NUM_VARIANTS = 23;

constexpr VariantInfo kVariants[NUM_VARIANTS] = {
    {"t512_ldg", 512, {16, 16, 12, 8}, false, false, 1, 0},
    {"t512_tma", 512, {24, 16, 12, 8}, true, false, 1, 0},
    {"t256_tma_occ", 256, {16, 16, 12, 8}, true, false, 2, 0},
    {"t256_ldg_occ", 256, {16, 16, 12, 8}, false, false, 2, 0},
    {"t512_tma_matchhw", 512, {24, 16, 12, 8}, true, true, 1, 0},
    {"t1024_tma", 1024, {12, 8, 6, 4}, true, false, 1, 0},
    {"or256_i16", 256, {16, 16, 10, 8}, false, false, 3, 1},
    {"or256_i8", 256, {8, 8, 6, 4}, false, false, 4, 1},
    {"or512_i8", 512, {8, 8, 6, 4}, false, false, 2, 1},
    {"or384_i18", 384, {18, 12, 8, 6}, false, false, 2, 1},
    {"or256_i12", 256, {12, 12, 8, 6}, false, false, 4, 1},
    {"or256_i24", 256, {24, 20, 12, 10}, false, false, 2, 1},
    {"h256_i16", 256, {16, 16, 10, 8}, false, false, 3, 2},
    {"h256_i12", 256, {12, 12, 8, 6}, false, false, 4, 2},
    {"h384_i18", 384, {18, 12, 8, 6}, false, false, 2, 2},
    {"h256_i24", 256, {24, 20, 12, 10}, false, false, 2, 2},
    {"h256_i20_occ3", 256, {20, 20, 12, 10}, false, false, 3, 2},
    {"h256_i24_occ3", 256, {24, 24, 14, 12}, false, false, 3, 2},
    {"h256_i32_occ2", 256, {32, 32, 20, 16}, false, false, 2, 2},
    {"h512_i16_occ2", 512, {16, 16, 10, 8}, false, false, 2, 2},
    {"h512_i12_occ2", 512, {20, 12, 8, 6}, false, false, 2, 2},
    {"h256_i48_occ2", 256, {48, 40, 24, 20}, false, false, 2, 2},
    {"h256_i40_occ2", 256, {40, 36, 22, 18}, false, false, 2, 2},
};

// Defaults measured on B200 (profiles/r1h_sweep_k2_polish.txt, r1l_sweep_large_tiles.txt): best variant per record
// width in words for inputs that fill the GPU with large tiles; below kSmallInputTiles tiles of
// the large variant per SM slot the 4096-record variant keeps every SM busy.
constexpr int kDefaultVariant[4] = {18, 18, 21, 22};
constexpr int kSmallVariant = 12;
constexpr int kRemapVariant = 18;

}  // namespace b200
