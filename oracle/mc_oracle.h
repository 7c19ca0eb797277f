/*
 * mc_oracle.h -- CPU oracle for the hand-ranking / showdown / equity hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may build, load or
 * call it, and only as the checker or as the timed CPU baseline.  The product (libmcpoker.so) never
 * links or calls it and has no CPU fallback.
 *
 * PARITY STATUS: the reference is Clojure and cannot be executed on this box (no JVM), so
 * this is a restatement in C of
 *     /root/reference/src/montecarlo/hand_evaluator.clj:6-133,156-172   (evaluator, compare-max, 7-card)
 *     /root/reference/src/montecarlo/card.clj:7-13                      (Card, COMPLETE-DECK order)
 *     /root/reference/src/montecarlo/gameplay.clj:118-133               (showdown ranking / tie groups)
 *     /root/reference/src/montecarlo/gameplay.clj:30-75, board.clj:148  (what one dealt hand is)
 * and of the two third-party functions the path calls that are not under /root/reference:
 *     org.clojure/math.combinatorics 0.0.4  combinations / permutations  (index-lexicographic order)
 *     org.clojure/clojure 1.4.0             compare on persistent vectors (APersistentVector.compareTo:
 *                                           length first, then element-wise Util.compare)
 * It is PINNED for the 5-card evaluator and its predicates by every assert of the reference's
 * test/montecarlo/hand_evaluator_test.clj (tests/golden/reference_vectors.json, tests/test_oracle.py).  The reference has
 * no test for 7-card ranking, showdown, or equity: those are "parity unpinned" by the reference and are
 * pinned here by composition (5-card value -> max under compare -> tie by =), with three independent
 * formulations (literal, rank-count best-of-21, direct bitmask) cross-checked exhaustively.
 *
 * Card encoding shared with the product (include/mcpoker.h):
 *   card id  = 13*suit + (rank-2) = index in COMPLETE-DECK (card.clj:10-13; suits hearts, diamonds,
 *              spades, clubs = 0..3; ranks 2..14)
 *   card set = uint64 mask, bit (16*suit + rank); rank 1..14 (rank 1 only exists in the reference's
 *              tests, hand_evaluator_test.clj:62-67,80)
 *   key      = cat<<20 | r1<<16 | r2<<12 | r3<<8 | r4<<4 | r5  with r1..r5 = (concat hit kick) of the
 *              Clojure value [cat hit kick]; unsigned order of keys == clojure.core/compare order.
 */
#ifndef MC_ORACLE_H
#define MC_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct { int suit; int rank; } orc_card;

/* [cat hit kick] exactly as hand_evaluator.clj:100-133 builds it. */
typedef struct {
    int cat;
    int nhit;  int hit[7];
    int nkick; int kick[7];
    int hit_is_lazyseq;     /* cat 6: slot 1 is (map :rank ...) -- a LazySeq, not Comparable */
} orc_value;

/* ---- literal restatement (ref_literal.c) ------------------------------------------------------ */
/* predicates: return 1 (truthy) / 0 (nil); *_hit variants also return the cards found */
int  orc_constant_coll(const int *xs, int n);                          /* hand_evaluator.clj:6-9   */
int  orc_straight_p(const orc_card *cards, int n);                     /* :32-40 */
int  orc_flush_p(const orc_card *cards, int n);                        /* :42-44 */
int  orc_n_of_a_kind_p(const orc_card *cards, int n, int k, orc_card *hit);      /* :46-55 */
int  orc_n_pairs_p(const orc_card *cards, int n, int k1, int k2,
                   orc_card *g1, int *n1, orc_card *g2, int *n2);      /* :71-79 */
int  orc_full_house_p(const orc_card *cards, int n);                   /* :88-90 */
int  orc_two_pair_p(const orc_card *cards, int n);                     /* :92-94 */
void orc_evaluator(const orc_card *cards5, orc_value *out);            /* :112-133 */
/* clojure.core/compare on two values; *threw set to 1 when the JVM would raise ClassCastException
 * (both cat 6); the returned sign is then the lexicographic ("intended") one. */
int  orc_compare(const orc_value *a, const orc_value *b, int *threw);
/* max-hand-value :156-160 (tie -> b) */
const orc_value *orc_max_hand_value(const orc_value *a, const orc_value *b, int *threw);
/* player->hand-value :162-172 on (concat community pocket) given in that order; n in 5..7.
 * returns 0, or -1 when n < 5 (the reference raises ArityException). */
int  orc_player_hand_value(const orc_card *cards, int n, orc_value *out, int *would_throw);
/* end-game ranking :127-132 + group-winners :118-120.  keys[] per seat (nibble keys); writes
 * group_of[seat] = index of its tie group (0 = best).  returns number of groups. */
int  orc_group_winners(const uint32_t *keys, int n_seats, int *group_of);

/* update-stacks :104-116: winner_sets[g] = seat bitmask of tie group g (best first); stack[seat] += share */
int  orc_update_stacks(const unsigned *winner_sets, int n_groups, const long long *pot_bet, const int *pot_n,
                       const unsigned *pot_players, int n_pots, long long *stack);

uint32_t orc_value_to_key(const orc_value *v);
void     orc_key_to_value(uint32_t key, orc_value *v);

/* ---- independent fast formulations (ref_fast.c) ----------------------------------------------- */
uint32_t orc_rc5_key(const int *ranks5, const int *suits5);            /* rank-count 5-card evaluator */
uint32_t orc_best_of_subsets_key(uint64_t mask);                       /* max of rc5 over all 5-subsets */
uint32_t orc_eval_mask(uint64_t mask);                                 /* direct bitmask, 5..7 cards; 0xFFFFFFFF on bad input */
uint32_t orc_eval_mask_literal(uint64_t mask, int *would_throw);       /* literal path on cards in ascending bit order */
void     orc_eval_batch(const uint64_t *masks, size_t n, uint32_t *keys, int which /*0 direct,1 best-of-subsets,2 literal*/, int threads);

/* exhaustive enumeration of all C(52,k) hands, k = 5,6,7, in colex order of card ids (index = sum_i C(c_i, i+1)).
 * which: 0 direct bitmask, 1 rank-count best-of-subsets.  hist[cat], key_sum, n_distinct may be NULL.
 * first/count select a sub-range of combination indices (count = 0 -> to the end). */
int  orc_enumerate(int k, int which, int threads, uint64_t first, uint64_t count,
                   uint64_t hist[9], uint64_t *key_sum, uint64_t *n_distinct);
int  orc_enumerate_masks(int k, uint64_t first, uint64_t count, int threads, uint64_t *masks);
/* same enumeration but compares the two fast formulations hand by hand; returns #mismatches */
uint64_t orc_enumerate_crosscheck(int k, int threads);

/* exhaustive equity: known holes, board_mask holds 0..5 known community cards, dead_mask removed. */
int  orc_equity_exhaustive(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                           int threads, int64_t *win, int64_t *tie, int64_t *loss, uint64_t *n_boards);

/* EV: the same enumeration, each board settled by group-winners (gameplay.clj:118-120) + the literal
 * update-stacks (:104-116, orc_update_stacks); chips[seat] = sum over the boards of what the seat receives */
int  orc_equity_exhaustive_ev(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask, int threads,
                              const long long *pot_bet, const int *pot_n, const unsigned *pot_players, int n_pots,
                              int64_t *win, int64_t *tie, int64_t *loss, int64_t *chips, uint64_t *n_boards);

/* Philox4x32-10 and the deal spec shared with the CUDA kernels (DESIGN.md "Deal spec") */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
uint64_t orc_random_hand_mask(uint64_t seed, uint64_t index, int n_cards);   /* cfg-1 hands */
void orc_random_hands(uint64_t seed, uint64_t first, size_t n, int n_cards, uint64_t *masks);
int  orc_deal_trial(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                    uint64_t seed, uint64_t trial, uint64_t *holes_out, uint64_t *board_out);
int  orc_equity_mc(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                   uint64_t seed, uint64_t first_trial, uint64_t n_trials, int threads,
                   int64_t *win, int64_t *tie, int64_t *loss);

int  orc_equity_mc_ev(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                      uint64_t seed, uint64_t first_trial, uint64_t n_trials, int threads,
                      const long long *pot_bet, const int *pot_n, const unsigned *pot_players, int n_pots,
                      int64_t *win, int64_t *tie, int64_t *loss, int64_t *chips);

/* Monte Carlo with a weighted hole range per seat (deal spec R1): weights = n_seats x 1326 by hole index
 * c0 + c1*(c1-1)/2 over card ids (rows of fixed seats are ignored); hole_masks[s] != 0 fixes seat s.  *n_void = trials in
 * which no attempt dealt disjoint holes (counted in none of win / tie / loss). */
int  orc_equity_mc_ranges(const int32_t *weights, const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                          uint64_t seed, uint64_t first_trial, uint64_t n_trials, int threads,
                          int64_t *win, int64_t *tie, int64_t *loss, uint64_t *n_void);

/* one cell of the 169x169 preflop matrix (class ids: see include/mcpoker.h) by brute force */
int  orc_preflop_cell(int class_a, int class_b, int threads, int64_t wtl[3]);

int  orc_max_threads(void);
void orc_set_threads(int threads);   /* overrides OMP_NUM_THREADS for the calls that follow */

#ifdef __cplusplus
}
#endif
#endif
