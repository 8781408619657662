/*
 * voda_oracle.h — C API of the CPU oracle.
 *
 * TEST INFRASTRUCTURE, NOT PRODUCT.  The oracle is a CPU restatement of the reference's
 * algorithms for the hot path (heyfey/vodascheduler pkg/algorithm, pkg/placement and the
 * external heyfey/munkres module).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it, and only as the checker or the timed
 * CPU baseline — never on the product path (vodascheduler_b200/ never imports it).
 *
 * Parity pin: the reference has no tests, golden vectors or fixtures for this path and no
 * Go toolchain exists on this image, so the oracle is pinned by the hand-traced known-answer
 * tests of SURVEY.md Appendix B/C (tests/golden/kats.json) only; the munkres part is
 * "parity unpinned" (module source not in the reference tree; see oracle_munkres.cpp).
 *
 * The oracle reuses the batch structs of include/voda_accel.h with HOST pointers so that
 * the same arrays feed it and the CUDA library.
 */
#ifndef VODA_ORACLE_H
#define VODA_ORACLE_H

#include "../include/voda_accel.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Structure-faithful restatement: string-keyed hash maps, a stable sort per greedy pick,
 * the reference's own loop nests.  Single-threaded when threads <= 1, else rounds are split
 * statically over `threads` host threads. */
int voda_oracle_allocate(const voda_alloc_batch* b, int32_t* out_gpus, int32_t* out_status,
                         int32_t threads);
/* Dense restatement: arrays instead of maps, same results bit for bit. */
int voda_oracle_allocate_dense(const voda_alloc_batch* b, int32_t* out_gpus,
                               int32_t* out_status, int32_t threads);

int voda_oracle_munkres_max(int32_t n, int32_t batch, const int64_t* A,
                            int32_t* out_col_of_row, int32_t* out_status, int32_t threads);
/* step-4 / step-6 iteration counts of the last single-threaded munkres call (diagnostics) */
void voda_oracle_munkres_counters(int64_t* step4, int64_t* step6, int64_t* step5);

int voda_oracle_place(const voda_place_batch* b, const voda_place_out* out, int32_t threads);

/* AFS-L diagnostics of the last single-threaded faithful call: fold steps, champion changes */
void voda_oracle_afsl_counters(int64_t* fold_steps, int64_t* champion_changes, int64_t* picks);

#ifdef __cplusplus
}
#endif
#endif
