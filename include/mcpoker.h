/*
 * mcpoker.h -- C ABI of libmcpoker.so, the B200 (sm_100a) engine behind the reference's hand-ranking /
 * showdown path.  Only ints, pointers to primitive arrays and sizes cross this boundary (JNA / Panama /
 * ctypes friendly: no structs by value, no callbacks, no torch types).
 *
 * What each entry point replaces in the reference (/root/reference/src/montecarlo/...):
 *   mc_eval_cards        hand_evaluator.clj:112-133 `evaluator` (5 cards) and :162-172 `player->hand-value`
 *                        (5/6/7 cards = flop/turn/river), one hand
 *   mc_eval7_batch[_dev] the same, lifted over a batch (`(map #(player->hand-value board %) ids)`,
 *                        gameplay.clj:127-128)
 *   mc_eval_batch_packed48  the same batch with hands as 6-byte fields of card positions (compact wire format)
 *   mc_eval_batch_colex32   the same batch with 7-card hands as 4-byte colex ranks in, 2-byte class ids out (the most
 *                        compact wire format: the host-buffer calls are bound by the host link)
 *   mc_end_game          end-game in one call: ranking, tie groups, settlement of the pots (gameplay.clj:122-133)
 *   mc_showdown          gameplay.clj:127-132 `end-game` ranking + :118-120 `group-winners`
 *   mc_enumerate7*       player->hand-value over every C(52,7) hand (BASELINE config 2)
 *   mc_equity_exhaustive* end-game ranking over every completion of the board (config 3)
 *   mc_equity_mc*        deal (board.clj:148 shuffle, gameplay.clj:63-75 deal-hand, :41-54 flop/turn/river)
 *                        + end-game ranking per trial (config 4)
 *   mc_preflop_matrix    config 5
 *
 * Encodings
 *   card id   13*suit + (rank-2): index in COMPLETE-DECK (card.clj:10-13); suit 0..3 = hearts, diamonds,
 *             spades, clubs; rank 2..14
 *   card set  uint64 mask, bit (16*suit + rank); ranks 1..14 are accepted by the evaluation entry points
 *             (rank 1 exists only in the reference's tests); enumeration / equity use the 52-card deck
 *   key       uint32, MC_KEY_NIBBLE (default):  cat<<20 | r1<<16 | r2<<12 | r3<<8 | r4<<4 | r5 where
 *             r1..r5 = (concat hit kick) of the Clojure value [cat hit kick] (hand_evaluator.clj:100-133).
 *             Unsigned order of keys == clojure.core/compare order of the values; equal keys <=> `=` values.
 *             hit lengths by cat 0..8: 5,2,4,3,5,5,5,4,5.   0xFFFFFFFF marks an invalid input hand.
 *             MC_KEY_PACKED: cat<<28 | P<<14 | S with P, S 14-bit rank sets (bit r-1): same order, decodes
 *             to the same value (mc_key_packed_to_nibble).
 *   class id  uint16, MC_KEY_CLASS16: the index of the value in the ascending list of all 7,462 distinct five-card
 *             values (the best five of 6 or 7 cards is one of them; 4,892 occur for 7 cards).  Same order as the keys,
 *             so class ids compare directly; mc_class_keys() hands out the list (class id -> nibble key, 30 KB) once.
 *             0xFFFF marks an invalid input hand (and the reference tests' rank-1 hands, which have no class).
 *
 * Every function returns 0 (MC_OK) or a negative MC_ERR_*; mc_last_error() holds a thread-local message.
 * The caller owns every buffer it passes; the library owns device memory, streams and tables.  All entry
 * points are thread-safe (calls that touch the same device are serialised inside the library), and that includes
 * the lifecycle: mc_init* / mc_shutdown wait for the calls in flight and no call starts on a context that is being
 * torn down (a call that arrives after mc_shutdown fails with MC_ERR_NOT_INIT).  There is no CPU fallback: without a
 * usable sm_100 device every compute call fails with MC_ERR_NO_DEVICE.
 */
#ifndef MCPOKER_H
#define MCPOKER_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MC_OK               0
#define MC_ERR_INVALID     -1   /* bad argument (the reference would throw or mis-evaluate) */
#define MC_ERR_NOT_INIT    -2
#define MC_ERR_CUDA        -3
#define MC_ERR_NO_DEVICE   -4
#define MC_ERR_TOO_FEW     -5   /* fewer than 5 cards: reference raises ArityException */

#define MC_KEY_NIBBLE       0
#define MC_KEY_PACKED       1
/* 2, 3, 4: A/B variants of the nibble-key kernels (bench.py) */
#define MC_KEY_CLASS16      5   /* uint16 class ids instead of uint32 keys */
#define MC_KEY_INVALID      0xFFFFFFFFu
#define MC_CLASS_INVALID    0xFFFFu
#define MC_N_HAND_CLASSES   7462

/* input formats of mc_eval_batch */
#define MC_IN_MASK64        0   /* uint64 card masks, 8 B per hand (5..7 cards) */
#define MC_IN_PACKED48      1   /* seven 6-bit card positions, 6 B per hand (5..7 cards) */
#define MC_IN_COLEX32       2   /* uint32 colex rank of a 7-card hand over card ids, 4 B per hand */
#define MC_IN_COLEX27       3   /* the same ranks as a bit stream, 27 bits per hand: 32 hands in 27 words, 3.375 B per hand */

#define MC_MAX_SEATS        9
#define MC_MAX_POTS         8   /* side pots of the EV calls */
#define MC_N_CLASSES        169

/* ---- lifecycle ---------------------------------------------------------------------------------- */
/* flags: MC_INIT_NCCL = sum the int64 counts of the devices with one ncclAllReduce per device (single process,
 * ncclCommInitAll over the devices, NVLink) instead of on the host from pinned copies.  libnccl.so.2 is loaded at run
 * time only when the flag is set and more than one device is driven; it fails loudly if it cannot be.  Both ways give
 * identical counts; for the 27..36 words of this path the host sum is the faster one. */
#define MC_INIT_NCCL        1u
int  mc_init(int n_gpus /* 0 = all visible */, unsigned flags);
int  mc_collective(void);                                 /* 1 = NCCL all-reduce, 0 = host sum */
int  mc_init_devices(const int *device_ids, int n);      /* one process per GPU: pass {LOCAL_RANK} */
void mc_shutdown(void);
int  mc_device_count(void);                               /* devices this library instance drives */
const char *mc_last_error(void);
int  mc_abi_version(void);

/* ---- pure host helpers (bit shuffling only, no evaluation) ------------------------------------- */
uint64_t mc_card_bit(int card_id);                                        /* 0 on bad id */
uint32_t mc_key_packed_to_nibble(uint32_t packed_key);
int  mc_key_decode(uint32_t nibble_key, int32_t *cat, int32_t hit[5], int32_t *n_hit,
                   int32_t kick[5], int32_t *n_kick);                     /* -> [cat hit kick] */

/* ---- evaluation (K1) ----------------------------------------------------------------------------- */
int  mc_eval_cards(const int32_t *suit_id, const int32_t *rank, int n_cards /* 5..7 */, uint32_t *key);
int  mc_eval7_batch(const uint64_t *masks, size_t n, uint32_t *keys);                 /* host buffers */
int  mc_eval7_batch_fmt(const uint64_t *masks, size_t n, uint32_t *keys, int key_format);
/* The same batch in the compact wire format: 6 bytes per hand = a little-endian 48-bit field of seven 6-bit card
 * positions (16*suit + rank, i.e. the bit mc_card_bit() sets; any order; 0 = no card, so the 5- and 6-card hands of
 * flop and turn fit as well; bits 42..47 zero).  The call is bound by the host link, so 6 B instead of 8 B per hand
 * is 4/3 of the throughput.  A repeated card or a malformed field gives MC_KEY_INVALID for that hand.  Replaces the
 * same reference lines as mc_eval7_batch (hand_evaluator.clj:162-172 over a batch). */
int  mc_eval_batch_packed48(const uint8_t *hands /* 6*n bytes */, size_t n, uint32_t *keys);
/* The most compact wire format: a 7-card hand as the colex rank of its card ids, sum_i C(c_i, i+1) with c_0 < ... < c_6
 * (the combination index of mc_enumerate7_range; < C(52,7) = 133,784,560 < 2^27) in, its class id out: 4 B + 2 B per hand
 * over the host link instead of 8 + 4.  A rank >= C(52,7) gives MC_CLASS_INVALID.  Unranked on the device. */
int  mc_eval_batch_colex32(const uint32_t *ranks, size_t n, uint16_t *classes);
/* The same with the ranks packed: a little-endian bit stream of 27 bits per hand (C(52,7) < 2^27), hand j = bits
 * [27 j, 27 j + 27) of the stream, i.e. 32 hands in 27 uint32 words; `stream` holds ceil(27 n / 32) words. */
int  mc_eval_batch_colex27(const uint32_t *stream, size_t n, uint16_t *classes);
/* Every combination of the above: in_format MC_IN_*, key_format MC_KEY_* (out = uint32 keys, or uint16 class ids for
 * MC_KEY_CLASS16).  stage_cards: 5, 6 or 7 = the batch is flop / turn / river hands, 0 = decide from the first hand; it
 * only selects the kernel's table image -- hands of another size are still evaluated exactly, just slower. */
int  mc_eval_batch(const void *hands, int in_format, size_t n, void *out, int key_format, int stage_cards);
/* Allocation-free variant of the colex32 evaluation for callers that cannot allocate (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * d_ranks (uint32 colex ranks) and d_classes (uint16 class ids, 4-byte aligned) are the caller's device
 * buffers on the current device; the intermediate card masks live in a static device array inside the library
 * (MC_NOALLOC_MAX_HANDS x 8 B = 8 MiB, held by every GPU that loads the library until the process exits).  The call
 * allocates nothing and only enqueues two kernels on `stream`.  n > MC_NOALLOC_MAX_HANDS fails with MC_ERR_INVALID.
 * All calls on a device share the array: calls that overlap on the GPU (different streams, or graph replays in
 * flight together) race -- serialise them on one stream.  Nothing has to be cleared between calls: each call overwrites
 * the n entries it reads.  Same results as mc_eval_batch_colex32.  Needs mc_init for the current device. */
#define MC_NOALLOC_MAX_HANDS 1048576
int  mc_eval_batch_colex32_dev_noalloc(const uint32_t *d_ranks, size_t n, uint16_t *d_classes, void *stream);
/* class id -> nibble key, MC_N_HAND_CLASSES entries, ascending (so class ids order like keys).  Needs mc_init. */
int  mc_class_keys(uint32_t *keys, int n /* = MC_N_HAND_CLASSES */);
/* Page-locked host memory for the buffers of the host-buffer calls: they copy straight from / into the caller's
 * memory, which runs at the full host-link rate only when it is pinned (a JVM caller wraps the pointer in a direct
 * ByteBuffer).  Pageable buffers are accepted everywhere, just slower.  Needs mc_init. */
void *mc_alloc_host(size_t bytes);                                                     /* NULL on failure (mc_last_error) */
int  mc_free_host(void *p);
int  mc_eval7_batch_dev(const uint64_t *d_masks, size_t n, uint32_t *d_keys, void *stream);
int  mc_eval7_batch_dev_fmt(const uint64_t *d_masks, size_t n, uint32_t *d_keys, int key_format, void *stream);
/* device buffers, any key format (d_out: uint32 keys or uint16 class ids) and stage (5, 6: flop / turn image; else river) */
int  mc_eval_batch_dev(const uint64_t *d_masks, size_t n, void *d_out, int key_format, int stage_cards, void *stream);
int  mc_unrank_colex32_dev(const uint32_t *d_ranks, size_t n, uint64_t *d_masks, void *stream);   /* colex ranks -> masks */
int  mc_showdown(uint64_t board_mask, const uint64_t *hole_masks, int n_seats,
                 uint32_t *keys, uint32_t *winner_bitmask);
/* full end-game grouping: group_of[seat] = index of its tie group, 0 = best (gameplay.clj:118-120) */
int  mc_showdown_groups(uint64_t board_mask, const uint64_t *hole_masks, int n_seats,
                        uint32_t *keys, int32_t *group_of, int32_t *n_groups, uint32_t *winner_bitmask /* nullable */);

/* pot settlement of end-game: update-stacks gameplay.clj:104-116 over the Bet records of bet.clj:7-8.  Pure integer
 * host arithmetic (no evaluation): for every pot, the FIRST tie group (group_of ascending = best first) that holds one
 * of the pot's players wins it, and each of its members IN the pot receives int(bet * n / members) chips (:110-112,
 * truncating division; the remainder is dropped exactly as the reference drops it).  pot_players = bitmask of seats.
 * delta[seat] is accumulated into.  A pot none of whose players holds cards fails with MC_ERR_INVALID (the reference
 * divides by zero there). */
int  mc_settle_pots(const int32_t *group_of, int n_seats, const int64_t *pot_bet, const int32_t *pot_n,
                    const uint32_t *pot_players, int n_pots, int64_t *delta);
/* end-game (gameplay.clj:122-133) in one call: mc_showdown_groups followed by mc_settle_pots on its groups.  keys,
 * group_of, n_groups, winner_bitmask are nullable; delta[seat] is accumulated into. */
int  mc_end_game(uint64_t board_mask, const uint64_t *hole_masks, int n_seats, const int64_t *pot_bet, const int32_t *pot_n,
                 const uint32_t *pot_players, int n_pots, uint32_t *keys, int32_t *group_of, int32_t *n_groups,
                 uint32_t *winner_bitmask, int64_t *delta);

/* ---- synthetic inputs (BASELINE config 1: Philox-seeded random hands, identical on CPU and GPU) --- */
int  mc_random_hands(uint64_t seed, uint64_t first, size_t n, int n_cards, uint64_t *masks);
int  mc_random_hands_dev(uint64_t seed, uint64_t first, size_t n, int n_cards, uint64_t *d_masks, void *stream);

/* ---- exhaustive enumeration (K2) ---------------------------------------------------------------- */
/* combination index = colex rank over card ids: sum_i C(c_i, i+1), c_0 < ... < c_6 */
int  mc_enumerate7(uint64_t hist[9], uint64_t *key_sum /* nullable; sum of nibble keys */);
int  mc_enumerate7_range(uint64_t first, uint64_t count, uint64_t hist[9], uint64_t *key_sum);
/* d_out: 10 x uint64 on the current device = hist[0..8], key_sum; accumulated into (zero it first) */
int  mc_enumerate7_dev(uint64_t first, uint64_t count, uint64_t *d_out, void *stream);

/* ---- exhaustive equity (K3) --------------------------------------------------------------------- */
int  mc_equity_exhaustive(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                          int64_t *win, int64_t *tie, int64_t *loss, uint64_t *n_boards);
int  mc_equity_exhaustive_range(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                                uint64_t first_board, uint64_t n_boards_in,
                                int64_t *win, int64_t *tie, int64_t *loss);
/* d_counts: 3*MC_MAX_SEATS int64 on the current device, [win[9] tie[9] loss[9]], accumulated into */
int  mc_equity_exhaustive_dev(const uint64_t *hole_masks /* host */, int n_seats, uint64_t board_mask,
                              uint64_t dead_mask, uint64_t first_board, uint64_t n_boards_in,
                              int64_t *d_counts, void *stream);

/* EV: the same enumeration with the reference's pot settlement applied to every board (update-stacks
 * gameplay.clj:104-116 over the Bet records of bet.clj:7-8; pots exactly as in mc_settle_pots).  Every seat of a
 * simulation is at the showdown, so for each pot the seats of pot_players holding its best hand receive
 * int(bet * n / their number) each (:110-112, the remainder is dropped as the reference drops it).  chips[seat] = the sum
 * over all boards of what the seat receives: EV = chips / n_boards minus what the seat put in.  1 <= n_pots <=
 * MC_MAX_POTS, every pot's players a non-empty subset of the seats, all pots together fewer than 2^31 chips. */
int  mc_equity_exhaustive_ev(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                             const int64_t *pot_bet, const int32_t *pot_n, const uint32_t *pot_players, int n_pots,
                             int64_t *win, int64_t *tie, int64_t *loss, int64_t *chips, uint64_t *n_boards);

/* ---- Monte Carlo equity (K4) -------------------------------------------------------------------- */
/* hole mask 0 = random seat.  Trial t depends only on (seed, t): any split of [first, first+n) over
 * devices / processes / calls sums to the same counts. */
int  mc_equity_mc(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                  uint64_t seed, uint64_t first_trial, uint64_t n_trials,
                  int64_t *win, int64_t *tie, int64_t *loss);
int  mc_equity_mc_dev(const uint64_t *hole_masks /* host */, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                      uint64_t seed, uint64_t first_trial, uint64_t n_trials,
                      int64_t *d_counts /* 27 int64, accumulated into */, void *stream);

/* EV by Monte Carlo: as mc_equity_exhaustive_ev, per trial.  d_counts of the _dev form: 36 int64 = win[9] tie[9] loss[9]
 * chips[9], accumulated into. */
int  mc_equity_mc_ev(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                     uint64_t seed, uint64_t first_trial, uint64_t n_trials,
                     const int64_t *pot_bet, const int32_t *pot_n, const uint32_t *pot_players, int n_pots,
                     int64_t *win, int64_t *tie, int64_t *loss, int64_t *chips);
int  mc_equity_mc_ev_dev(const uint64_t *hole_masks /* host */, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                         uint64_t seed, uint64_t first_trial, uint64_t n_trials,
                         const int64_t *pot_bet, const int32_t *pot_n, const uint32_t *pot_players, int n_pots,
                         int64_t *d_counts /* 36 int64, accumulated into */, void *stream);

/* ---- heads-up range vs range on a (partial) board ------------------------------------- */
/* A hole is indexed by the colex rank of its two card ids: idx = c0 + c1*(c1-1)/2, c0 < c1 (0..1325).  Every pair of
 * disjoint holes (a from range A, b from range B, neither touching board | dead) with weight_a[a] * weight_b[b] != 0 is
 * played out over every completion of the board; wtl = {win, tie, loss} of side A and *total, all weighted by the
 * product of the two weights (weights 0..32767).  All devices of the library instance share the matchup list. */
#define MC_N_HOLES 1326
int  mc_equity_ranges(const int32_t *weight_a /* 1326 */, const int32_t *weight_b /* 1326 */, uint64_t board_mask,
                      uint64_t dead_mask, int64_t wtl[3], int64_t *total /* nullable */);

/* ---- Monte Carlo with a weighted hole range per seat ------------------------------------------------- */
/* Seat s is fixed (hole_masks[s] != 0) or ranged: weights[s * 1326 + hole index], 0..32767 (the rows of fixed seats are
 * ignored; holes that hold a board, dead or fixed card are dropped).  Per trial every ranged seat draws a hole exactly in
 * proportion to its weight (integer alias table, counter-based Philox: trial t depends only on (seed, t), so any split of
 * the trial range sums to the same counts); trials whose holes collide are dealt again, so the joint law is the product of
 * the weights restricted to disjoint holes; the missing community cards are uniform.  *n_void (nullable) = trials in
 * which 4096 attempts all collided (ranges that cannot be dealt together); they are in none of win / tie / loss.
 * What an equity query of a client with beliefs about its opponents asks (README.md:9); stages as deal-stage
 * gameplay.clj:56-61 (board_mask = the community cards so far). */
int  mc_equity_mc_ranges(const int32_t *weights /* n_seats * 1326 */, const uint64_t *hole_masks, int n_seats,
                         uint64_t board_mask, uint64_t dead_mask, uint64_t seed, uint64_t first_trial, uint64_t n_trials,
                         int64_t *win, int64_t *tie, int64_t *loss, uint64_t *n_void);

/* ---- 169 x 169 preflop matrix (K5) -------------------------------------------------------------- */
/* class id: x = rank-2; pair 13x+x; suited (hi>lo) 13*hi+lo; offsuit 13*lo+hi.  wtl[a][b] = {win,tie,loss} of a */
int  mc_preflop_matrix(int64_t *wtl /* 169*169*3 */);
/* share `part` of `n_parts` of the suit-canonical matchup list (one process per GPU: part = rank); the full matrix
 * is the element-wise sum of the parts (one all-reduce of 169*169*3 int64) */
int  mc_preflop_matrix_part(int part, int n_parts, int64_t *wtl /* 169*169*3 */, int32_t *n_matchups /* nullable */);
/* The same matrix by BOARDS: every suit-canonical 5-card board (134,459, weight = orbit size) evaluates
 * its 1,081 hole pairs once, sorts them and counts wins by class with a running scan, then removes the pairs of holes
 * that share a card.  mc_preflop_wins_by_boards returns share `part` of `n_parts` of win[a][b] (169*169 int64; the
 * full win matrix is the element-wise sum of the parts); mc_preflop_matrix_from_wins derives {win,tie,loss} from the
 * full win matrix (host arithmetic: loss = win transposed, tie = pairs * C(48,5) - win - loss);
 * mc_preflop_matrix_by_boards does both on the devices of this library instance.  Equal to mc_preflop_matrix cell
 * by cell. */
int  mc_preflop_boards_count(int32_t *n_boards /* nullable */, int64_t *weight_sum /* nullable */);
int  mc_preflop_wins_by_boards(int part, int n_parts, int64_t *win /* 169*169 */, int32_t *n_boards /* nullable */);
int  mc_preflop_matrix_from_wins(const int64_t *win /* 169*169 */, int64_t *wtl /* 169*169*3 */);
int  mc_preflop_matrix_by_boards(int64_t *wtl /* 169*169*3 */);
int  mc_preflop_cells(const int32_t *class_a, const int32_t *class_b, int n_cells, int64_t *wtl /* n*3 */);

#ifdef __cplusplus
}
#endif
#endif
