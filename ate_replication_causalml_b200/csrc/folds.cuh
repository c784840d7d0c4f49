// folds.cuh -- device construction of the cross-fit's fold lists (folds.cu).
#pragma once
#include "common.cuh"

// work arrays for one ate_i_build_fold_lists call (all device; n = all rows, nloc = this rank's rows)
struct FoldScratch {
    int32_t *key_n, *pos_n;             // [n]
    int32_t *key_a, *key_b, *key_h;     // [n]
    int32_t *pos_loc;                   // [nloc]
    int *blockhist;                     // ate_i_rank_scratch_ints(n, max(F, 1)) ints
    long long *totals;                  // [65]
    int *bad;                           // error flags: 1 outer id out of range, 2 inner id out of range
};

size_t ate_i_rank_scratch_ints(long long n, int C);
// stable counting sort by a small key: d_pos[i] = number of earlier elements with the same key
// (keys 0..C; d_pos may be NULL: class sizes only); d_totals[c] = size of class c.
int ate_i_rank_by_key(ate_ctx *ctx, const int32_t *d_keys, long long n, int C, int32_t *d_pos, int *d_blockhist,
                      long long *d_totals, int *d_bad);
// stable group-by on top of it: d_list = indices 0..n-1 ordered by key (keys 1..C; key 0 is skipped)
int ate_i_group_by_key(ate_ctx *ctx, const int32_t *d_keys, long long n, int C, int32_t *d_pos, int *d_blockhist,
                       long long *d_totals, int *d_bad, int32_t *d_list);
int ate_i_build_fold_lists(ate_ctx *ctx, const int32_t *d_fold, long long n, long long lo, long long hi, int k, int K,
                           const int32_t *d_inner_a, const int32_t *d_inner_b, int F, FoldScratch &sc,
                           int32_t *list_a, int32_t *list_b, int32_t *list_h, long long *d_tot);
