/*
 * ref_literal.c -- function-by-function restatement of the reference's naive evaluator.
 * TEST INFRASTRUCTURE ONLY (see mc_oracle.h).  Deliberately keeps the reference's algorithm: C(5,n)
 * subset scans, the 120-permutation n-pairs?, compare-based max over the 21 five-card subsets -- it is
 * both the semantic anchor and the timed "reference CPU implementation" of bench.py.
 *
 * Every function cites the lines of /root/reference/src/montecarlo/hand_evaluator.clj it follows.
 */
#include "mc_oracle.h"
#include <string.h>

#define MAXC 7

/* ---- math.combinatorics 0.0.4 (not under /root/reference; published behaviour restated) ---------- */
/* combinations: index-lexicographic k-subsets.  idx[] must start as 0..k-1; returns 0 when exhausted. */
static int next_combination(int *idx, int n, int k) {
    int i = k - 1;
    while (i >= 0 && idx[i] == n - k + i) i--;
    if (i < 0) return 0;
    idx[i]++;
    for (int j = i + 1; j < k; j++) idx[j] = idx[j - 1] + 1;
    return 1;
}
/* permutations: index-lexicographic; perm[] starts as 0..n-1 */
static int next_permutation(int *p, int n) {
    int i = n - 2;
    while (i >= 0 && p[i] >= p[i + 1]) i--;
    if (i < 0) return 0;
    int j = n - 1;
    while (p[j] <= p[i]) j--;
    int t = p[i]; p[i] = p[j]; p[j] = t;
    for (int a = i + 1, b = n - 1; a < b; a++, b--) { t = p[a]; p[a] = p[b]; p[b] = t; }
    return 1;
}

static int card_eq(orc_card a, orc_card b) { return a.suit == b.suit && a.rank == b.rank; }

/* stable insertion sort ascending by rank == (sort-by :rank x) for these sizes */
static void sort_by_rank(orc_card *c, int n) {
    for (int i = 1; i < n; i++) {
        orc_card x = c[i]; int j = i - 1;
        while (j >= 0 && c[j].rank > x.rank) { c[j + 1] = c[j]; j--; }
        c[j + 1] = x;
    }
}
static void sort_ints_desc(int *a, int n) {
    for (int i = 1; i < n; i++) {
        int x = a[i], j = i - 1;
        while (j >= 0 && a[j] < x) { a[j + 1] = a[j]; j--; }
        a[j + 1] = x;
    }
}

/* hand_evaluator.clj:6-9  (every? true? (map = coll (rest coll))) */
int orc_constant_coll(const int *xs, int n) {
    for (int i = 0; i + 1 < n; i++) if (xs[i] != xs[i + 1]) return 0;
    return 1;
}
/* :11-14 constant-cards? on :rank / :suit */
static int constant_rank(const orc_card *c, int n) {
    for (int i = 0; i + 1 < n; i++) if (c[i].rank != c[i + 1].rank) return 0;
    return 1;
}
static int constant_suit(const orc_card *c, int n) {
    for (int i = 0; i + 1 < n; i++) if (c[i].suit != c[i + 1].suit) return 0;
    return 1;
}

/* :18-22 high-card: a iff rank a > rank b, else b */
static orc_card high_card(orc_card a, orc_card b) { return (a.rank > b.rank) ? a : b; }

/* :24-30 high-hand: hand-a iff (= (high-card high-a high-b) high-a), else hand-b */
static int high_hand_is_a(const orc_card *a, int na, const orc_card *b, int nb) {
    orc_card ha = a[0], hb = b[0];
    for (int i = 1; i < na; i++) ha = high_card(ha, a[i]);
    for (int i = 1; i < nb; i++) hb = high_card(hb, b[i]);
    return card_eq(high_card(ha, hb), ha);
}

/* :32-40 straight?: sorted ranks, each the inc of the previous.  No wheel. */
int orc_straight_p(const orc_card *cards, int n) {
    int r[MAXC];
    for (int i = 0; i < n; i++) r[i] = cards[i].rank;
    for (int i = 1; i < n; i++) { int x = r[i], j = i - 1; while (j >= 0 && r[j] > x) { r[j + 1] = r[j]; j--; } r[j + 1] = x; }
    for (int i = 0; i + 1 < n; i++) if (r[i] + 1 != r[i + 1]) return 0;
    return 1;
}

/* :42-44 flush? */
int orc_flush_p(const orc_card *cards, int n) { return constant_suit(cards, n); }

/* :46-55 n-of-a-kinds + n-of-a-kind?: all k-subsets (index-lex) with constant rank, reduced by high-hand */
int orc_n_of_a_kind_p(const orc_card *cards, int n, int k, orc_card *hit) {
    if (k > n) return 0;                       /* combinations -> nil */
    int idx[MAXC]; for (int i = 0; i < k; i++) idx[i] = i;
    orc_card best[MAXC], cur[MAXC]; int have = 0;
    do {
        for (int i = 0; i < k; i++) cur[i] = cards[idx[i]];
        if (!constant_rank(cur, k)) continue;
        if (!have) { memcpy(best, cur, sizeof(orc_card) * k); have = 1; }
        else if (!high_hand_is_a(best, k, cur, k)) memcpy(best, cur, sizeof(orc_card) * k);
    } while (next_combination(idx, n, k));
    if (have && hit) memcpy(hit, best, sizeof(orc_card) * k);
    return have;
}

/* :71-79 n-pairs?: first permutation (index-lex) whose first k1 and next k2 cards are rank-constant;
 * each group rank-sorted.  `distinct` cannot change which hit is first, so it is not materialised. */
int orc_n_pairs_p(const orc_card *cards, int n, int k1, int k2,
                  orc_card *g1, int *n1, orc_card *g2, int *n2) {
    int p[MAXC]; for (int i = 0; i < n; i++) p[i] = i;
    do {
        orc_card a[MAXC], b[MAXC]; int na = 0, nb = 0;
        for (int i = 0; i < k1 && i < n; i++) a[na++] = cards[p[i]];          /* (take k1 %)           */
        for (int i = k1; i < k1 + k2 && i < n; i++) b[nb++] = cards[p[i]];    /* (take k2 (drop k1 %)) */
        sort_by_rank(a, na); sort_by_rank(b, nb);
        if (constant_rank(a, na) && constant_rank(b, nb)) {
            if (g1) { memcpy(g1, a, sizeof(orc_card) * na); *n1 = na; }
            if (g2) { memcpy(g2, b, sizeof(orc_card) * nb); *n2 = nb; }
            return 1;
        }
    } while (next_permutation(p, n));
    return 0;
}
int orc_full_house_p(const orc_card *cards, int n) { return orc_n_pairs_p(cards, n, 3, 2, 0, 0, 0, 0); } /* :88-90 */
int orc_two_pair_p(const orc_card *cards, int n)   { return orc_n_pairs_p(cards, n, 2, 2, 0, 0, 0, 0); } /* :92-94 */

/* :96-98 cards->ranks */
static int cards_to_ranks(const orc_card *c, int n, int *out) {
    for (int i = 0; i < n; i++) out[i] = c[i].rank;
    sort_ints_desc(out, n);
    return n;
}
/* :100-102 ret: [value (cards->ranks hit) (cards->ranks (remove (set hit) cards))] */
static void ret(int value, const orc_card *cards, int n, const orc_card *hit, int nh, orc_value *out) {
    memset(out, 0, sizeof *out);
    out->cat = value;
    out->nhit = cards_to_ranks(hit, nh, out->hit);
    orc_card rest[MAXC]; int nr = 0;
    for (int i = 0; i < n; i++) {
        int in = 0;
        for (int j = 0; j < nh; j++) if (card_eq(cards[i], hit[j])) { in = 1; break; }
        if (!in) rest[nr++] = cards[i];
    }
    out->nkick = cards_to_ranks(rest, nr, out->kick);
}

/* :112-133 evaluator (5 cards) */
void orc_evaluator(const orc_card *cards, orc_value *out) {
    const int n = 5;
    int is_straight = orc_straight_p(cards, n);
    int is_flush = orc_flush_p(cards, n);
    orc_card hit[MAXC], g1[MAXC], g2[MAXC]; int n1, n2;
    memset(out, 0, sizeof *out);
    if (is_straight && is_flush) {                                   /* :116-118 */
        out->cat = 8; out->nhit = cards_to_ranks(cards, n, out->hit); return;
    }
    if (orc_n_of_a_kind_p(cards, n, 4, hit)) { ret(7, cards, n, hit, 4, out); return; }   /* :119-120 */
    if (orc_n_pairs_p(cards, n, 3, 2, g1, &n1, g2, &n2)) {           /* :121-122, ret-full-house :104-106 */
        out->cat = 6; out->hit_is_lazyseq = 1;                       /* (map :rank (flatten hit)) */
        for (int i = 0; i < n1; i++) out->hit[out->nhit++] = g1[i].rank;
        for (int i = 0; i < n2; i++) out->hit[out->nhit++] = g2[i].rank;
        return;
    }
    if (is_flush)    { out->cat = 5; out->nhit = cards_to_ranks(cards, n, out->hit); return; }   /* :123-124 */
    if (is_straight) { out->cat = 4; out->nhit = cards_to_ranks(cards, n, out->hit); return; }   /* :125-126 */
    if (orc_n_of_a_kind_p(cards, n, 3, hit)) { ret(3, cards, n, hit, 3, out); return; }           /* :127-128 */
    if (orc_n_pairs_p(cards, n, 2, 2, g1, &n1, g2, &n2)) {           /* :129-130, ret-two-pair :108-110 */
        orc_card flat[MAXC]; int nf = 0;
        for (int i = 0; i < n1; i++) flat[nf++] = g1[i];
        for (int i = 0; i < n2; i++) flat[nf++] = g2[i];
        ret(2, cards, n, flat, nf, out); return;
    }
    if (orc_n_of_a_kind_p(cards, n, 2, hit)) { ret(1, cards, n, hit, 2, out); return; }           /* :131-132 */
    ret(0, 0, 0, cards, n, out);                                      /* :133 (ret 0 [] cards): hit = cards */
}

/* clojure.core/compare on [cat hit kick] (APersistentVector.compareTo, Clojure 1.4.0) */
static int cmp_int(int a, int b) { return (a > b) - (a < b); }
static int cmp_vec(const int *a, int na, const int *b, int nb) {
    if (na != nb) return na < nb ? -1 : 1;
    for (int i = 0; i < na; i++) { int c = cmp_int(a[i], b[i]); if (c) return c; }
    return 0;
}
int orc_compare(const orc_value *a, const orc_value *b, int *threw) {
    int c = cmp_int(a->cat, b->cat);
    if (c) return c;
    /* slot 1: two distinct LazySeq objects -> ClassCastException; intended order kept */
    if ((a->hit_is_lazyseq || b->hit_is_lazyseq) && threw) *threw = 1;
    c = cmp_vec(a->hit, a->nhit, b->hit, b->nhit);
    if (c) return c;
    return cmp_vec(a->kick, a->nkick, b->kick, b->nkick);
}
/* :156-160 */
const orc_value *orc_max_hand_value(const orc_value *a, const orc_value *b, int *threw) {
    return orc_compare(a, b, threw) > 0 ? a : b;
}

/* :162-172 player->hand-value minus the println and the ref derefs */
int orc_player_hand_value(const orc_card *cards, int n, orc_value *out, int *would_throw) {
    if (would_throw) *would_throw = 0;
    if (n < 5 || n > MAXC) return -1;          /* (reduce f ()) -> ArityException, SURVEY Q3 */
    int idx[5] = {0, 1, 2, 3, 4};
    orc_value acc, cur; int first = 1, threw = 0;
    do {
        orc_card h[5];
        for (int i = 0; i < 5; i++) h[i] = cards[idx[i]];
        orc_evaluator(h, &cur);
        if (first) { acc = cur; first = 0; }
        else if (orc_max_hand_value(&acc, &cur, &threw) == &cur) acc = cur;
    } while (next_combination(idx, n, 5));
    *out = acc;
    if (would_throw) *would_throw = threw;
    return 0;
}

/* gameplay.clj:127-132 + :118-120: sort-by value, reverse, partition-by value -> tie groups best first */
int orc_group_winners(const uint32_t *keys, int n_seats, int *group_of) {
    int groups = 0;
    for (int s = 0; s < n_seats; s++) {
        int g = 0;                                 /* number of distinct values strictly greater */
        for (int t = 0; t < n_seats; t++) {
            if (keys[t] <= keys[s]) continue;
            int seen = 0;
            for (int u = 0; u < t; u++) if (keys[u] == keys[t]) { seen = 1; break; }
            if (!seen) g++;
        }
        group_of[s] = g;
        if (g + 1 > groups) groups = g + 1;
    }
    return groups;
}

/* update-stacks, gameplay.clj:104-116, statement by statement.  winners = the tie groups as sets of seats, best first
 * (group-winners); a pot = Bet{bet players n} (bet.clj:7-8), players as a seat set.
 *   pot-intersections = (map intersection (repeat (:players pot)) winners)        :107-109
 *   pot-winners       = (first (filter not-empty pot-intersections))               :110
 *   pot-share         = (int (/ (* bet n) (count pot-winners)))                    :111-112
 *   every p in pot-winners gets + pot-share                                        :113-116
 * returns 0, or -1 where the reference would throw (no intersection -> (count nil) = 0 -> divide by zero). */
int orc_update_stacks(const unsigned *winner_sets, int n_groups, const long long *pot_bet, const int *pot_n,
                      const unsigned *pot_players, int n_pots, long long *stack /* by seat */) {
    for (int p = 0; p < n_pots; p++) {
        unsigned pot_winners = 0;
        for (int g = 0; g < n_groups; g++) {
            const unsigned inter = pot_players[p] & winner_sets[g];
            if (inter) { pot_winners = inter; break; }
        }
        if (!pot_winners) return -1;
        const long long total_pot = pot_bet[p] * (long long)pot_n[p];
        const long long pot_share = total_pot / __builtin_popcount(pot_winners);     /* (int (/ a b)) truncates */
        for (int seat = 0; seat < 32; seat++) if ((pot_winners >> seat) & 1u) stack[seat] += pot_share;
    }
    return 0;
}

/* key <-> value (hit lengths per category) */
static const int HIT_LEN[9] = {5, 2, 4, 3, 5, 5, 5, 4, 5};
uint32_t orc_value_to_key(const orc_value *v) {
    uint32_t k = (uint32_t)v->cat;
    int n = 0;
    for (int i = 0; i < v->nhit; i++, n++)  k = (k << 4) | (uint32_t)(v->hit[i] & 15);
    for (int i = 0; i < v->nkick; i++, n++) k = (k << 4) | (uint32_t)(v->kick[i] & 15);
    for (; n < 5; n++) k <<= 4;
    return k;
}
void orc_key_to_value(uint32_t key, orc_value *v) {
    memset(v, 0, sizeof *v);
    v->cat = (int)(key >> 20);
    int r[5];
    for (int i = 0; i < 5; i++) r[i] = (int)((key >> (16 - 4 * i)) & 15);
    int hl = (v->cat >= 0 && v->cat <= 8) ? HIT_LEN[v->cat] : 5;
    for (int i = 0; i < 5; i++) { if (i < hl) v->hit[v->nhit++] = r[i]; else v->kick[v->nkick++] = r[i]; }
    v->hit_is_lazyseq = (v->cat == 6);
}

/* literal 7-card path on a mask: cards listed in ascending bit order (suit-major, rank ascending) */
uint32_t orc_eval_mask_literal(uint64_t mask, int *would_throw) {
    orc_card c[MAXC]; int n = 0;
    for (int b = 0; b < 64; b++) if ((mask >> b) & 1) {
        if (n == MAXC) return 0xFFFFFFFFu;
        c[n].suit = b >> 4; c[n].rank = b & 15; n++;
    }
    orc_value v;
    if (orc_player_hand_value(c, n, &v, would_throw)) return 0xFFFFFFFFu;
    return orc_value_to_key(&v);
}
