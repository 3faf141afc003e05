// Kernel-variant selectors for tools/ (sweeps) and tests/ (bit-identity of variants).  NOT part of the public
// C ABI (include/vslam_b200.h): process-global, not thread-safe, meant for single-threaded measurement scripts.
#pragma once
#ifdef __cplusplus
extern "C" {
#endif
/* kNN-2 tile / thread variant: 0 default <128 threads, 4 queries per thread, 512-row tiles, carry-save popcount>. */
int vslam_hamming_set_variant(int variant);
/* Feature response kernel: 0 automatic (by batch size), 1 warp-specialised (latency), 2 one thread per column. */
int vslam_features_set_variant(int variant);
/* Schur complement kernel of a BA problem: 1 sorted units over a precomputed G (default), 0 resident chunks
 * (Y = Hpl Hinv of a camera chunk in shared memory, partner blocks gathered; measured slower at C4: 455 vs 323 us). */
struct vslam_ba;
int vslam_ba_set_schur_variant(struct vslam_ba* ba, int variant);
#ifdef __cplusplus
}
#endif
