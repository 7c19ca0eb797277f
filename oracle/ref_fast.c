/*
 * ref_fast.c -- two independent fast formulations of the reference's hand value, plus the lifted
 * workloads (enumeration, exhaustive equity, Philox Monte Carlo, preflop cell) on the CPU.
 * TEST INFRASTRUCTURE ONLY (see mc_oracle.h).
 *
 *   orc_rc5_key              5-card value from rank counts; same category order and hit/kick shapes as
 *                            hand_evaluator.clj:112-133 (table in SURVEY 8.0)
 *   orc_best_of_subsets_key  max over the C(n,5) subsets = player->hand-value, hand_evaluator.clj:171-172,
 *                            with compare == unsigned key order
 *   orc_eval_mask            direct 5/6/7-card evaluation from per-rank counts (no subset loop)
 * The CUDA kernels use a third, bit-plane formulation (monte-carlo_b200/csrc/eval.cuh); none of the three
 * shares code.
 */
#include "mc_oracle.h"
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define VALID_MASK 0x7FFE7FFE7FFE7FFEULL      /* bits 1..14 of each 16-bit suit lane (ranks 1..14)  */
#define DECK_MASK  0x7FFC7FFC7FFC7FFCULL      /* ranks 2..14: the 52 cards of COMPLETE-DECK, card.clj:10-13 */

int orc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
/* the launcher (torchrun) exports OMP_NUM_THREADS=1; the timed CPU baseline wants every host core it may use */
void orc_set_threads(int t) {
#ifdef _OPENMP
    if (t > 0) omp_set_num_threads(t);
#else
    (void)t;
#endif
}
static int use_threads(int t) { int m = orc_max_threads(); return (t <= 0 || t > m) ? m : t; }

static uint32_t pack_key(int cat, const int *r) {
    return ((uint32_t)cat << 20) | ((uint32_t)r[0] << 16) | ((uint32_t)r[1] << 12) |
           ((uint32_t)r[2] << 8) | ((uint32_t)r[3] << 4) | (uint32_t)r[4];
}

/* ---------------------------------------------------------------------------------------------- */
/* rank-count 5-card evaluator                                                                    */
uint32_t orc_rc5_key(const int *ranks, const int *suits) {
    int r[5];
    for (int i = 0; i < 5; i++) r[i] = ranks[i];
    for (int i = 1; i < 5; i++) { int x = r[i], j = i - 1; while (j >= 0 && r[j] < x) { r[j + 1] = r[j]; j--; } r[j + 1] = x; }
    int flush = suits[0] == suits[1] && suits[1] == suits[2] && suits[2] == suits[3] && suits[3] == suits[4];
    int straight = r[0] == r[1] + 1 && r[1] == r[2] + 1 && r[2] == r[3] + 1 && r[3] == r[4] + 1;   /* no wheel */
    /* group sorted ranks into (count, rank) runs */
    int gr[5], gc[5], ng = 0;
    for (int i = 0; i < 5;) { int j = i; while (j < 5 && r[j] == r[i]) j++; gr[ng] = r[i]; gc[ng] = j - i; ng++; i = j; }
    int out[5] = {0, 0, 0, 0, 0};
    if (straight && flush) return pack_key(8, r);
    if (ng == 2) {
        int big = gc[0] >= gc[1] ? 0 : 1, small = 1 - big;
        if (gc[big] == 4) { out[0] = out[1] = out[2] = out[3] = gr[big]; out[4] = gr[small]; return pack_key(7, out); }
        out[0] = out[1] = out[2] = gr[big]; out[3] = out[4] = gr[small]; return pack_key(6, out);    /* t t t p p */
    }
    if (flush) return pack_key(5, r);
    if (straight) return pack_key(4, r);
    if (ng == 3) {
        int t = -1;
        for (int i = 0; i < 3; i++) if (gc[i] == 3) t = i;
        if (t >= 0) {                                           /* [3 [t t t] [k1 k2]] */
            int k = 3; out[0] = out[1] = out[2] = gr[t];
            for (int i = 0; i < 3; i++) if (i != t) out[k++] = gr[i];
            return pack_key(3, out);
        }
        int k = 0, kick = 0;                                    /* [2 [h h l l] [k]] */
        for (int i = 0; i < 3; i++) { if (gc[i] == 2) { out[k] = out[k + 1] = gr[i]; k += 2; } else kick = gr[i]; }
        out[4] = kick; return pack_key(2, out);
    }
    if (ng == 4) {                                              /* [1 [p p] [k1 k2 k3]] */
        int k = 2;
        for (int i = 0; i < 4; i++) { if (gc[i] == 2) out[0] = out[1] = gr[i]; else out[k++] = gr[i]; }
        return pack_key(1, out);
    }
    return pack_key(0, r);
}

static int mask_to_cards(uint64_t mask, int *ranks, int *suits) {
    int n = 0;
    if (mask & ~VALID_MASK) return -1;
    while (mask) {
        int b = __builtin_ctzll(mask); mask &= mask - 1;
        if (n == 7) return -1;
        suits[n] = b >> 4; ranks[n] = b & 15; n++;
    }
    return n;
}

uint32_t orc_best_of_subsets_key(uint64_t mask) {
    int ranks[7], suits[7];
    int n = mask_to_cards(mask, ranks, suits);
    if (n < 5) return 0xFFFFFFFFu;
    uint32_t best = 0;
    int idx[5];
    for (idx[0] = 0; idx[0] < n; idx[0]++)
    for (idx[1] = idx[0] + 1; idx[1] < n; idx[1]++)
    for (idx[2] = idx[1] + 1; idx[2] < n; idx[2]++)
    for (idx[3] = idx[2] + 1; idx[3] < n; idx[3]++)
    for (idx[4] = idx[3] + 1; idx[4] < n; idx[4]++) {
        int r5[5], s5[5];
        for (int i = 0; i < 5; i++) { r5[i] = ranks[idx[i]]; s5[i] = suits[idx[i]]; }
        uint32_t k = orc_rc5_key(r5, s5);
        if (k > best) best = k;
    }
    return best;
}

/* ---------------------------------------------------------------------------------------------- */
/* direct evaluator from per-rank counts, 5..7 cards                                              */
uint32_t orc_eval_mask(uint64_t mask) {
    if (mask & ~VALID_MASK) return 0xFFFFFFFFu;
    int n = __builtin_popcountll(mask);
    if (n < 5 || n > 7) return 0xFFFFFFFFu;
    int cnt[16] = {0}, scnt[4] = {0};
    for (int s = 0; s < 4; s++) {
        unsigned lane = (unsigned)(mask >> (16 * s)) & 0xFFFFu;
        scnt[s] = __builtin_popcount(lane);
        for (int r = 1; r <= 14; r++) cnt[r] += (lane >> r) & 1;
    }
    int out[5] = {0, 0, 0, 0, 0}, k;
    /* flush / straight flush: with <= 7 cards a flush excludes quads and full houses */
    for (int s = 0; s < 4; s++) if (scnt[s] >= 5) {
        unsigned lane = (unsigned)(mask >> (16 * s)) & 0xFFFFu;
        for (int hi = 14; hi >= 5; hi--) {
            unsigned run = 0x1Fu << (hi - 4);
            if ((lane & run) == run) { for (int i = 0; i < 5; i++) out[i] = hi - i; return pack_key(8, out); }
        }
        k = 0;
        for (int r = 14; r >= 1 && k < 5; r--) if ((lane >> r) & 1) out[k++] = r;
        return pack_key(5, out);
    }
    int quad = 0, trip1 = 0, trip2 = 0, pair[3] = {0, 0, 0}, np = 0;
    for (int r = 14; r >= 1; r--) {
        if (cnt[r] == 4) quad = r;
        else if (cnt[r] == 3) { if (!trip1) trip1 = r; else if (!trip2) trip2 = r; }
        else if (cnt[r] == 2) { if (np < 3) pair[np++] = r; }
    }
    if (quad) {
        int kick = 0;
        for (int r = 14; r >= 1; r--) if (cnt[r] && r != quad) { kick = r; break; }
        out[0] = out[1] = out[2] = out[3] = quad; out[4] = kick; return pack_key(7, out);
    }
    if (trip1 && (trip2 || np)) {
        int p = trip2 > pair[0] ? trip2 : pair[0];
        out[0] = out[1] = out[2] = trip1; out[3] = out[4] = p; return pack_key(6, out);
    }
    for (int hi = 14; hi >= 5; hi--) {
        int ok = 1;
        for (int i = 0; i < 5; i++) if (!cnt[hi - i]) { ok = 0; break; }
        if (ok) { for (int i = 0; i < 5; i++) out[i] = hi - i; return pack_key(4, out); }
    }
    if (trip1) {
        out[0] = out[1] = out[2] = trip1; k = 3;
        for (int r = 14; r >= 1 && k < 5; r--) if (cnt[r] && r != trip1) out[k++] = r;
        return pack_key(3, out);
    }
    if (np >= 2) {
        out[0] = out[1] = pair[0]; out[2] = out[3] = pair[1];
        for (int r = 14; r >= 1; r--) if (cnt[r] && r != pair[0] && r != pair[1]) { out[4] = r; break; }
        return pack_key(2, out);
    }
    if (np == 1) {
        out[0] = out[1] = pair[0]; k = 2;
        for (int r = 14; r >= 1 && k < 5; r--) if (cnt[r] && r != pair[0]) out[k++] = r;
        return pack_key(1, out);
    }
    k = 0;
    for (int r = 14; r >= 1 && k < 5; r--) if (cnt[r]) out[k++] = r;
    return pack_key(0, out);
}

void orc_eval_batch(const uint64_t *masks, size_t n, uint32_t *keys, int which, int threads) {
    int nt = use_threads(threads);
    (void)nt;
#pragma omp parallel for schedule(static) num_threads(nt)
    for (long long i = 0; i < (long long)n; i++) {
        if (which == 0) keys[i] = orc_eval_mask(masks[i]);
        else if (which == 1) keys[i] = orc_best_of_subsets_key(masks[i]);
        else keys[i] = orc_eval_mask_literal(masks[i], 0);
    }
}

/* ---------------------------------------------------------------------------------------------- */
/* combinations in colex order over positions 0..n-1:  index = sum_i C(c_i, i+1), c_0 < c_1 < ...   */
static uint64_t BINOM[65][9];
static int binom_ready = 0;
static void binom_init(void) {
    if (binom_ready) return;
    for (int n = 0; n <= 64; n++) for (int k = 0; k <= 8; k++)
        BINOM[n][k] = (k == 0) ? 1 : (n == 0 ? 0 : BINOM[n - 1][k - 1] + BINOM[n - 1][k]);
    binom_ready = 1;
}
static void colex_unrank(uint64_t idx, int k, int *c) {
    for (int i = k; i >= 1; i--) {
        int x = i - 1;
        while (BINOM[x + 1][i] <= idx) x++;
        c[i - 1] = x; idx -= BINOM[x][i];
    }
}
static int colex_next(int *c, int k, int n) {          /* lowest position moves fastest */
    for (int i = 0; i < k; i++) {
        int lim = (i + 1 < k) ? c[i + 1] : n;
        if (c[i] + 1 < lim) { c[i]++; for (int j = 0; j < i; j++) c[j] = j; return 1; }
    }
    return 0;
}

static uint64_t CARD_BIT[52];
static void cards_init(void) {
    for (int id = 0; id < 52; id++) CARD_BIT[id] = 1ULL << (16 * (id / 13) + (id % 13) + 2);
}

int orc_enumerate(int k, int which, int threads, uint64_t first, uint64_t count,
                  uint64_t hist[9], uint64_t *key_sum, uint64_t *n_distinct) {
    if (k < 5 || k > 7) return -1;
    binom_init(); cards_init();
    uint64_t total = BINOM[52][k];
    if (first > total) return -1;
    if (count == 0 || first + count > total) count = total - first;
    int nt = use_threads(threads);
    uint64_t H[9] = {0}, S = 0;
    unsigned char *seen = n_distinct ? calloc(1u << 24, 1) : 0;
    uint64_t chunk = (count + (uint64_t)nt - 1) / (uint64_t)nt;
#pragma omp parallel num_threads(nt)
    {
        uint64_t h[9] = {0}, s = 0;
#pragma omp for schedule(static, 1)
        for (int t = 0; t < nt; t++) {
            uint64_t lo = first + chunk * (uint64_t)t, hi = lo + chunk;
            if (hi > first + count) hi = first + count;
            if (lo >= hi) continue;
            int c[7]; colex_unrank(lo, k, c);
            for (uint64_t i = lo; i < hi; i++) {
                uint64_t m = 0;
                for (int j = 0; j < k; j++) m |= CARD_BIT[c[j]];
                uint32_t key = which == 0 ? orc_eval_mask(m) : orc_best_of_subsets_key(m);
                h[key >> 20]++; s += key;
                if (seen) seen[key] = 1;            /* benign race: only ever stores 1 */
                colex_next(c, k, 52);
            }
        }
#pragma omp critical
        { for (int i = 0; i < 9; i++) H[i] += h[i]; S += s; }
    }
    if (hist) memcpy(hist, H, sizeof H);
    if (key_sum) *key_sum = S;
    if (n_distinct) { uint64_t d = 0; for (size_t i = 0; i < (1u << 24); i++) d += seen[i]; *n_distinct = d; free(seen); }
    return 0;
}

/* the masks of combinations first .. first+count-1 (same colex order as orc_enumerate), for key-by-key checks */
int orc_enumerate_masks(int k, uint64_t first, uint64_t count, int threads, uint64_t *masks) {
    if (k < 1 || k > 7) return -1;
    binom_init(); cards_init();
    if (first > BINOM[52][k] || count > BINOM[52][k] - first) return -1;
    int nt = use_threads(threads);
    uint64_t chunk = (count + (uint64_t)nt - 1) / (uint64_t)nt;
#pragma omp parallel for num_threads(nt) schedule(static, 1)
    for (int t = 0; t < nt; t++) {
        uint64_t lo = chunk * (uint64_t)t, hi = lo + chunk > count ? count : lo + chunk;
        if (lo >= hi) continue;
        int c[7]; colex_unrank(first + lo, k, c);
        for (uint64_t i = lo; i < hi; i++) {
            uint64_t m = 0;
            for (int j = 0; j < k; j++) m |= CARD_BIT[c[j]];
            masks[i] = m;
            colex_next(c, k, 52);
        }
    }
    return 0;
}

uint64_t orc_enumerate_crosscheck(int k, int threads) {
    if (k < 5 || k > 7) return ~0ULL;
    binom_init(); cards_init();
    uint64_t total = BINOM[52][k], bad = 0;
    int nt = use_threads(threads);
    uint64_t chunk = (total + (uint64_t)nt - 1) / (uint64_t)nt;
#pragma omp parallel for schedule(static, 1) num_threads(nt) reduction(+ : bad)
    for (int t = 0; t < nt; t++) {
        uint64_t lo = chunk * (uint64_t)t, hi = lo + chunk;
        if (hi > total) hi = total;
        if (lo >= hi) continue;
        int c[7]; colex_unrank(lo, k, c);
        for (uint64_t i = lo; i < hi; i++) {
            uint64_t m = 0;
            for (int j = 0; j < k; j++) m |= CARD_BIT[c[j]];
            if (orc_eval_mask(m) != orc_best_of_subsets_key(m)) bad++;
            colex_next(c, k, 52);
        }
    }
    return bad;
}

/* ---------------------------------------------------------------------------------------------- */
/* showdown counting shared by the equity functions: win if sole max, tie if shared max, else loss  */
static void count_showdown(const uint32_t *keys, int n, int64_t *win, int64_t *tie, int64_t *loss) {
    uint32_t best = 0; int nbest = 0;
    for (int s = 0; s < n; s++) { if (keys[s] > best) { best = keys[s]; nbest = 1; } else if (keys[s] == best) nbest++; }
    for (int s = 0; s < n; s++) {
        if (keys[s] != best) loss[s]++;
        else if (nbest == 1) win[s]++;
        else tie[s]++;
    }
}

/* EV: the reference's settlement of one showdown -- group-winners gameplay.clj:118-120 on the keys, then the
 * literal update-stacks :104-116 (ref_literal.c) over the pots; chips[seat] += what the seat receives. */
typedef struct { const long long *bet; const int *n; const unsigned *players; int n_pots; } pots_t;
static int settle_showdown(const uint32_t *keys, int n_seats, const pots_t *P, long long *chips /* 32 seats */) {
    int group_of[9];
    const int ng = orc_group_winners(keys, n_seats, group_of);
    unsigned sets[9] = {0};
    for (int s = 0; s < n_seats; s++) sets[group_of[s]] |= 1u << s;
    return orc_update_stacks(sets, ng, P->bet, P->n, P->players, P->n_pots, chips);
}
static int check_pots(const pots_t *P, int n_seats) {
    if (P->n_pots < 1) return -1;
    for (int p = 0; p < P->n_pots; p++)
        if (P->players[p] == 0 || (P->players[p] >> n_seats) || P->bet[p] < 0 || P->n[p] < 0) return -1;
    return 0;
}

static int equity_exhaustive_impl(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                                  int threads, const pots_t *P, int64_t *win, int64_t *tie, int64_t *loss, int64_t *chips,
                                  uint64_t *n_boards) {
    if (n_seats < 1 || n_seats > 9) return -1;
    if (P && check_pots(P, n_seats)) return -1;
    binom_init();
    uint64_t used = board_mask | dead_mask;
    if (board_mask & ~DECK_MASK) return -1;
    int nb = __builtin_popcountll(board_mask);
    if (nb > 5) return -1;
    for (int s = 0; s < n_seats; s++) {
        if ((hole_masks[s] & ~DECK_MASK) || __builtin_popcountll(hole_masks[s]) != 2) return -1;
        if (hole_masks[s] & used) return -1;
        used |= hole_masks[s];
    }
    uint64_t deck[52]; int nd = 0;
    for (int b = 0; b < 64; b++) if (((DECK_MASK & ~used) >> b) & 1) deck[nd++] = 1ULL << b;
    int k = 5 - nb;
    if (nd < k) return -1;
    uint64_t total = BINOM[nd][k];
    int nt = use_threads(threads);
    int64_t W[9] = {0}, T[9] = {0}, L[9] = {0};
    long long CH[32] = {0};
    uint64_t chunk = (total + (uint64_t)nt - 1) / (uint64_t)nt;
#pragma omp parallel num_threads(nt)
    {
        int64_t w[9] = {0}, t_[9] = {0}, l[9] = {0};
        long long ch[32] = {0};
#pragma omp for schedule(static, 1)
        for (int t = 0; t < nt; t++) {
            uint64_t lo = chunk * (uint64_t)t, hi = lo + chunk;
            if (hi > total) hi = total;
            if (lo >= hi) continue;
            int c[5] = {0, 0, 0, 0, 0};
            if (k > 0) colex_unrank(lo, k, c);
            for (uint64_t i = lo; i < hi; i++) {
                uint64_t b = board_mask;
                for (int j = 0; j < k; j++) b |= deck[c[j]];
                uint32_t keys[9];
                for (int s = 0; s < n_seats; s++) keys[s] = orc_eval_mask(b | hole_masks[s]);
                count_showdown(keys, n_seats, w, t_, l);
                if (P) settle_showdown(keys, n_seats, P, ch);
                if (k > 0) colex_next(c, k, nd);
            }
        }
#pragma omp critical
        { for (int s = 0; s < 9; s++) { W[s] += w[s]; T[s] += t_[s]; L[s] += l[s]; CH[s] += ch[s]; } }
    }
    for (int s = 0; s < n_seats; s++) { win[s] = W[s]; tie[s] = T[s]; loss[s] = L[s]; if (chips) chips[s] = CH[s]; }
    if (n_boards) *n_boards = total;
    return 0;
}
int orc_equity_exhaustive(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                          int threads, int64_t *win, int64_t *tie, int64_t *loss, uint64_t *n_boards) {
    return equity_exhaustive_impl(hole_masks, n_seats, board_mask, dead_mask, threads, 0, win, tie, loss, 0, n_boards);
}
int orc_equity_exhaustive_ev(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask, int threads,
                             const long long *pot_bet, const int *pot_n, const unsigned *pot_players, int n_pots,
                             int64_t *win, int64_t *tie, int64_t *loss, int64_t *chips, uint64_t *n_boards) {
    const pots_t P = {pot_bet, pot_n, pot_players, n_pots};
    return equity_exhaustive_impl(hole_masks, n_seats, board_mask, dead_mask, threads, &P, win, tie, loss, chips, n_boards);
}

/* ---------------------------------------------------------------------------------------------- */
/* Philox4x32-10 (Salmon et al., SC'11), the counter-based generator named in BASELINE.json          */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

#define TAG_DEAL 0x4D434445u     /* counter word 3 for Monte Carlo trials */
#define TAG_HAND 0x4D434831u     /* counter word 3 for the cfg-1 random-hand generator */

/* Deal spec: the 6-bit values of block b are (w_j >> 6i) & 63 for j = 0..3, i = 0..4 of
 * Philox(key = seed, ctr = (index_lo, index_hi, b, tag)); a value is a bit position in the card mask and
 * is accepted iff it is a real card (bits 2..14 of a lane) that is not yet used. */
typedef struct { uint32_t key[2], ctr[4], w[4]; int pos; } deal_stream;
static void stream_init(deal_stream *st, uint64_t seed, uint64_t index, uint32_t tag) {
    st->key[0] = (uint32_t)seed; st->key[1] = (uint32_t)(seed >> 32);
    st->ctr[0] = (uint32_t)index; st->ctr[1] = (uint32_t)(index >> 32); st->ctr[2] = 0; st->ctr[3] = tag;
    st->pos = 20;
}
static int stream_next6(deal_stream *st) {
    if (st->pos == 20) { orc_philox4x32_10(st->ctr, st->key, st->w); st->ctr[2]++; st->pos = 0; }
    int v = (int)((st->w[st->pos / 5] >> (6 * (st->pos % 5))) & 63u);
    st->pos++;
    return v;
}
static uint64_t stream_card(deal_stream *st, uint64_t *used) {
    for (;;) {
        uint64_t bit = 1ULL << stream_next6(st);
        if (!(*used & bit)) { *used |= bit; return bit; }
    }
}

uint64_t orc_random_hand_mask(uint64_t seed, uint64_t index, int n_cards) {
    deal_stream st; stream_init(&st, seed, index, TAG_HAND);
    uint64_t used = ~DECK_MASK, m = 0;
    for (int i = 0; i < n_cards; i++) m |= stream_card(&st, &used);
    return m;
}
void orc_random_hands(uint64_t seed, uint64_t first, size_t n, int n_cards, uint64_t *masks) {
#pragma omp parallel for schedule(static)
    for (long long i = 0; i < (long long)n; i++) masks[i] = orc_random_hand_mask(seed, first + (uint64_t)i, n_cards);
}

/* one trial (deal spec v4): random seats (hole mask 0) receive cards in the reference's deal order --
 * one card per random seat, twice round the table (gameplay.clj:63-75) -- then the missing community cards
 * (gameplay.clj:41-54; burns are distributionally irrelevant).  The cards are the first entries of a Fisher-Yates
 * shuffle of the ascending list of live cards; seat cards are steps 0..2nr-1, community card k is step 2nr + k
 * (steps of community cards that are already known are skipped), step i swaps deck[i] with deck[j], j in [i, n):
 *   word w (the w-th 32-bit output of Philox blocks 0..B-1, block b = Philox4x32-10(key = seed,
 *   ctr = (trial_lo, trial_hi, b | attempt << 8, TAG_DEAL4))) serves steps 3w, 3w+1, 3w+2 by a multiply chain:
 *       r = word;  for each of its steps:  x = r * (n - i);  j = i + (x >> 32);  r = x & 0xFFFFFFFF
 *   which is floor(word * M / 2^32) in mixed radix, M = product of the (n - i) of its executed steps, so by Lemire's
 *   bounded sampler it is exact iff the final r >= 2^32 mod M; otherwise the whole trial is dealt again with
 *   attempt + 1. */
#define TAG_DEAL4 0x4D434434u
int orc_deal_trial(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                   uint64_t seed, uint64_t trial, uint64_t *holes_out, uint64_t *board_out) {
    uint64_t used = dead_mask | board_mask;
    int rseat[9], nr = 0;
    for (int s = 0; s < n_seats; s++) { used |= hole_masks[s]; holes_out[s] = hole_masks[s]; if (!hole_masks[s]) rseat[nr++] = s; }
    int live[52], n = 0;
    for (int b = 0; b < 64; b++) if (((DECK_MASK & ~used) >> b) & 1ULL) live[n++] = b;
    const int nb = 5 - __builtin_popcountll(board_mask);
    const int n_steps = 2 * nr + 5, n_words = (n_steps + 2) / 3, n_blocks = (n_words + 3) / 4;
    const uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    for (uint32_t attempt = 0;; attempt++) {
        uint32_t word[16];
        for (int b = 0; b < n_blocks; b++) {
            const uint32_t ctr[4] = {(uint32_t)trial, (uint32_t)(trial >> 32), (uint32_t)b | (attempt << 8), TAG_DEAL4};
            orc_philox4x32_10(ctr, key, word + 4 * b);
        }
        int deck[52], rejected = 0;
        for (int i = 0; i < n; i++) deck[i] = live[i];
        for (int w = 0; w < n_words && !rejected; w++) {
            uint64_t r = word[w], M = 1;
            for (int i = 3 * w; i < 3 * w + 3 && i < n_steps; i++) {
                if (i >= 2 * nr + nb) continue;                 /* community card already known */
                const uint64_t m = (uint64_t)(n - i), x = r * m;
                const int j = i + (int)(x >> 32);
                r = x & 0xFFFFFFFFULL; M *= m;
                const int c = deck[j]; deck[j] = deck[i]; deck[i] = c;
            }
            if (r < (1ULL << 32) % M) rejected = 1;
        }
        if (rejected) continue;
        for (int i = 0; i < 2 * nr; i++) holes_out[rseat[i % nr]] |= 1ULL << deck[i];
        uint64_t bm = board_mask;
        for (int k = 0; k < nb; k++) bm |= 1ULL << deck[2 * nr + k];
        *board_out = bm;
        return 0;
    }
}

static int check_mc_args(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask) {
    if (n_seats < 1 || n_seats > 9) return -1;
    if ((board_mask & ~DECK_MASK) || (dead_mask & ~DECK_MASK) || (board_mask & dead_mask)) return -1;
    if (__builtin_popcountll(board_mask) > 5) return -1;
    uint64_t used = board_mask | dead_mask; int need = 5 - __builtin_popcountll(board_mask);
    for (int s = 0; s < n_seats; s++) {
        if (!hole_masks[s]) { need += 2; continue; }
        if ((hole_masks[s] & ~DECK_MASK) || __builtin_popcountll(hole_masks[s]) != 2 || (hole_masks[s] & used)) return -1;
        used |= hole_masks[s];
    }
    if (52 - __builtin_popcountll(used) < need) return -1;
    return 0;
}

static int equity_mc_impl(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                          uint64_t seed, uint64_t first_trial, uint64_t n_trials, int threads, const pots_t *P,
                          int64_t *win, int64_t *tie, int64_t *loss, int64_t *chips) {
    if (check_mc_args(hole_masks, n_seats, board_mask, dead_mask)) return -1;
    if (P && check_pots(P, n_seats)) return -1;
    int nt = use_threads(threads);
    (void)nt;
    int64_t W[9] = {0}, T[9] = {0}, L[9] = {0};
    long long CH[32] = {0};
#pragma omp parallel num_threads(nt)
    {
        int64_t w[9] = {0}, t_[9] = {0}, l[9] = {0};
        long long ch[32] = {0};
#pragma omp for schedule(static)
        for (long long i = 0; i < (long long)n_trials; i++) {
            uint64_t holes[9], board; uint32_t keys[9];
            orc_deal_trial(hole_masks, n_seats, board_mask, dead_mask, seed, first_trial + (uint64_t)i, holes, &board);
            for (int s = 0; s < n_seats; s++) keys[s] = orc_eval_mask(board | holes[s]);
            count_showdown(keys, n_seats, w, t_, l);
            if (P) settle_showdown(keys, n_seats, P, ch);
        }
#pragma omp critical
        { for (int s = 0; s < 9; s++) { W[s] += w[s]; T[s] += t_[s]; L[s] += l[s]; CH[s] += ch[s]; } }
    }
    for (int s = 0; s < n_seats; s++) { win[s] = W[s]; tie[s] = T[s]; loss[s] = L[s]; if (chips) chips[s] = CH[s]; }
    return 0;
}
int orc_equity_mc(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                  uint64_t seed, uint64_t first_trial, uint64_t n_trials, int threads,
                  int64_t *win, int64_t *tie, int64_t *loss) {
    return equity_mc_impl(hole_masks, n_seats, board_mask, dead_mask, seed, first_trial, n_trials, threads, 0, win, tie, loss, 0);
}
int orc_equity_mc_ev(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                     uint64_t seed, uint64_t first_trial, uint64_t n_trials, int threads,
                     const long long *pot_bet, const int *pot_n, const unsigned *pot_players, int n_pots,
                     int64_t *win, int64_t *tie, int64_t *loss, int64_t *chips) {
    const pots_t P = {pot_bet, pot_n, pot_players, n_pots};
    return equity_mc_impl(hole_masks, n_seats, board_mask, dead_mask, seed, first_trial, n_trials, threads, &P, win, tie, loss, chips);
}

/* ---------------------------------------------------------------------------------------------- */
/* Monte Carlo with a weighted hole RANGE per seat (SURVEY 8f.3 tail; deal spec R1).  A seat is fixed (hole mask)
 * or ranged (1326 weights by hole index c0 + c1*(c1-1)/2 over card ids).  used0 = board | dead | fixed holes; a ranged
 * seat's list = its holes of non-zero weight that avoid used0, in ascending hole index; drawn through an integer alias
 * table (exactly proportional to the weights); a hole that hits a card already dealt in the trial fails the ATTEMPT, which
 * is then restarted with the next Philox attempt number (so the joint law is prod_s w_s(h_s) restricted to disjoint holes);
 * the missing community cards are drawn uniformly from the live cards, re-drawing a card that is already out. */
#define TAG_RANGE 0x4D435231u
#define RANGE_MAX_ATTEMPTS 4096
typedef struct { unsigned char c0, c1; uint16_t alias; uint32_t thr; } range_entry;
typedef struct { uint32_t key[2], ctr[4], w[4]; int pos; } word_stream;
static void ws_init(word_stream *st, uint64_t seed, uint64_t trial, uint32_t attempt) {
    st->key[0] = (uint32_t)seed; st->key[1] = (uint32_t)(seed >> 32);
    st->ctr[0] = (uint32_t)trial; st->ctr[1] = (uint32_t)(trial >> 32); st->ctr[2] = attempt << 16; st->ctr[3] = TAG_RANGE;
    st->pos = 4;
}
static uint32_t ws_next(word_stream *st) {
    if (st->pos == 4) { orc_philox4x32_10(st->ctr, st->key, st->w); st->ctr[2]++; st->pos = 0; }
    return st->w[st->pos++];
}
static uint32_t ws_below(word_stream *st, uint32_t n) {          /* exactly uniform in [0, n): Lemire, re-drawing the word */
    const uint32_t zone = (uint32_t)((1ULL << 32) % n);
    for (;;) {
        const uint64_t x = (uint64_t)ws_next(st) * n;
        if ((uint32_t)x >= zone) return (uint32_t)(x >> 32);
    }
}
/* integer alias table over n items of weights w (sum W): slot i keeps thr[i] of its capacity W for itself, the rest goes to
 * alias[i].  Vose's construction with two stacks filled in ascending index order (part of the spec: the GPU library
 * builds the same table). */
static void build_alias(const uint32_t *w, int n, uint32_t W, range_entry *e) {
    uint64_t *mass = malloc(sizeof(uint64_t) * (size_t)n);
    int *small = malloc(sizeof(int) * (size_t)n), *large = malloc(sizeof(int) * (size_t)n), ns = 0, nl = 0;
    for (int i = 0; i < n; i++) { mass[i] = (uint64_t)w[i] * (uint64_t)n; if (mass[i] < W) small[ns++] = i; else large[nl++] = i; }
    while (ns && nl) {
        const int s_ = small[--ns], l = large[--nl];
        e[s_].thr = (uint32_t)mass[s_]; e[s_].alias = (uint16_t)l;
        mass[l] -= (uint64_t)W - mass[s_];
        if (mass[l] < W) small[ns++] = l; else large[nl++] = l;
    }
    while (nl) { const int l = large[--nl]; e[l].thr = W; e[l].alias = (uint16_t)l; }
    while (ns) { const int s_ = small[--ns]; e[s_].thr = W; e[s_].alias = (uint16_t)s_; }     /* (only by rounding: never with integers) */
    free(mass); free(small); free(large);
}
int orc_equity_mc_ranges(const int32_t *weights, const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                         uint64_t seed, uint64_t first_trial, uint64_t n_trials, int threads,
                         int64_t *win, int64_t *tie, int64_t *loss, uint64_t *n_void) {
    if (n_seats < 1 || n_seats > 9) return -1;
    if ((board_mask & ~DECK_MASK) || (dead_mask & ~DECK_MASK) || (board_mask & dead_mask) || __builtin_popcountll(board_mask) > 5) return -1;
    cards_init();
    uint64_t used0 = board_mask | dead_mask;
    for (int s = 0; s < n_seats; s++) if (hole_masks[s]) {
        if ((hole_masks[s] & ~DECK_MASK) || __builtin_popcountll(hole_masks[s]) != 2 || (hole_masks[s] & used0)) return -1;
        used0 |= hole_masks[s];
    }
    range_entry *tab[9] = {0}; uint32_t n_of[9] = {0}, W_of[9] = {0};
    int rc = 0, n_ranged = 0;
    for (int s = 0; s < n_seats && !rc; s++) if (!hole_masks[s]) {
        uint32_t w[1326]; range_entry *e = malloc(sizeof(range_entry) * 1326); int n = 0; uint64_t W = 0;
        tab[s] = e;
        for (int c1 = 1, idx = 0; c1 < 52; c1++) for (int c0 = 0; c0 < c1; c0++, idx++) {
            const int32_t wt = weights[s * 1326 + idx];
            if (wt < 0 || wt > 32767) rc = -1;
            if (wt <= 0 || ((CARD_BIT[c0] | CARD_BIT[c1]) & used0)) continue;
            e[n].c0 = (unsigned char)__builtin_ctzll(CARD_BIT[c0]); e[n].c1 = (unsigned char)__builtin_ctzll(CARD_BIT[c1]);
            w[n++] = (uint32_t)wt; W += (uint32_t)wt;
        }
        if (n == 0) rc = -1;
        if (!rc) { build_alias(w, n, (uint32_t)W, e); n_of[s] = (uint32_t)n; W_of[s] = (uint32_t)W; n_ranged++; }
    }
    int live[52], n_live = 0;
    for (int b = 0; b < 64; b++) if (((DECK_MASK & ~used0) >> b) & 1ULL) live[n_live++] = b;
    const int need = 5 - __builtin_popcountll(board_mask);
    if (!rc && n_live < need + 2 * n_ranged) rc = -1;
    int64_t W_[9] = {0}, T_[9] = {0}, L_[9] = {0}; uint64_t V_ = 0;
    if (!rc) {
        int nt = use_threads(threads);
        (void)nt;
#pragma omp parallel num_threads(nt)
        {
            int64_t w[9] = {0}, t_[9] = {0}, l[9] = {0}; uint64_t v = 0;
#pragma omp for schedule(static)
            for (long long i = 0; i < (long long)n_trials; i++) {
                const uint64_t trial = first_trial + (uint64_t)i;
                uint64_t holes[9], bm = 0; int ok = 0;
                for (uint32_t attempt = 0; attempt < RANGE_MAX_ATTEMPTS && !ok; attempt++) {
                    word_stream st; ws_init(&st, seed, trial, attempt);
                    uint64_t used = used0; int fail = 0;
                    for (int s = 0; s < n_seats && !fail; s++) {
                        holes[s] = hole_masks[s];
                        if (hole_masks[s]) continue;
                        const uint32_t slot = ws_below(&st, n_of[s]), vv = ws_below(&st, W_of[s]);
                        const range_entry *e = &tab[s][slot];
                        if (vv >= e->thr) e = &tab[s][e->alias];
                        const uint64_t h = (1ULL << e->c0) | (1ULL << e->c1);
                        if (h & used) { fail = 1; break; }
                        used |= h; holes[s] = h;
                    }
                    if (fail) continue;
                    bm = board_mask;
                    for (int k = 0; k < need; k++) {
                        for (;;) {
                            const uint64_t bit = 1ULL << live[ws_below(&st, (uint32_t)n_live)];
                            if (!(used & bit)) { used |= bit; bm |= bit; break; }
                        }
                    }
                    ok = 1;
                }
                if (!ok) { v++; continue; }
                uint32_t keys[9];
                for (int s = 0; s < n_seats; s++) keys[s] = orc_eval_mask(bm | holes[s]);
                count_showdown(keys, n_seats, w, t_, l);
            }
#pragma omp critical
            { for (int s = 0; s < 9; s++) { W_[s] += w[s]; T_[s] += t_[s]; L_[s] += l[s]; } V_ += v; }
        }
    }
    for (int s = 0; s < 9; s++) free(tab[s]);
    if (rc) return rc;
    for (int s = 0; s < n_seats; s++) { win[s] = W_[s]; tie[s] = T_[s]; loss[s] = L_[s]; }
    if (n_void) *n_void = V_;
    return 0;
}

/* ---------------------------------------------------------------------------------------------- */
/* 169 starting-hand classes: x = rank-2 (0..12); pair 13x+x; suited hi>lo 13*hi+lo; offsuit 13*lo+hi */
static int class_holes(int cls, uint64_t *out) {
    int a = cls / 13, b = cls % 13, n = 0;
    if (a == b) {
        for (int s = 0; s < 4; s++) for (int t = s + 1; t < 4; t++)
            out[n++] = (1ULL << (16 * s + a + 2)) | (1ULL << (16 * t + a + 2));
    } else if (a > b) {                       /* suited */
        for (int s = 0; s < 4; s++) out[n++] = (1ULL << (16 * s + a + 2)) | (1ULL << (16 * s + b + 2));
    } else {                                  /* offsuit: hi = b, lo = a */
        for (int s = 0; s < 4; s++) for (int t = 0; t < 4; t++) if (s != t)
            out[n++] = (1ULL << (16 * s + b + 2)) | (1ULL << (16 * t + a + 2));
    }
    return n;
}
int orc_preflop_cell(int class_a, int class_b, int threads, int64_t wtl[3]) {
    if (class_a < 0 || class_a >= 169 || class_b < 0 || class_b >= 169) return -1;
    uint64_t ha[12], hb[12];
    int na = class_holes(class_a, ha), nb = class_holes(class_b, hb);
    wtl[0] = wtl[1] = wtl[2] = 0;
    for (int i = 0; i < na; i++) for (int j = 0; j < nb; j++) {
        if (ha[i] & hb[j]) continue;
        uint64_t holes[2] = {ha[i], hb[j]}; int64_t w[2], t[2], l[2]; uint64_t nbd;
        if (orc_equity_exhaustive(holes, 2, 0, 0, threads, w, t, l, &nbd)) return -1;
        wtl[0] += w[0]; wtl[1] += t[0]; wtl[2] += l[0];
    }
    return 0;
}
