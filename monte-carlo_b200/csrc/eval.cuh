// eval.cuh -- device-side hand evaluation for sm_100a: bit-plane formulation, no subset loop.
//
// Replaces /root/reference/src/montecarlo/hand_evaluator.clj:112-133 (`evaluator`) and :162-172
// (`player->hand-value`, best of the 21 / 6 / 1 five-card subsets) for 5, 6 or 7 cards at once.
// Semantics are the reference's: category order 8..0, hit/kick shapes of `ret`,
// `ret-full-house`, `ret-two-pair`, NO ace-low straight (straight? :32-40 has none), full houses
// ordered lexicographically on (t t t p p).
//
// Input: the card set as two 32-bit words; lo = lanes of suits 0,1, hi = lanes of suits 2,3; each 16-bit
// lane has bit r set for rank r (1..14).
// Output ("packed key"): cat<<28 | P<<14 | S, P and S being 14-bit rank sets (bit r-1):
//   cat 8/4  P = 0            S = {top rank of the straight}
//   cat 7    P = {quad rank}  S = {kicker}
//   cat 6    P = {trips}      S = {pair}
//   cat 5/0  P = 0            S = the five ranks
//   cat 3    P = {trips}      S = two kickers
//   cat 2    P = two pairs    S = {kicker}
//   cat 1    P = {pair}       S = three kickers
// Within a category |P| and |S| are fixed, so unsigned order of packed keys == clojure.core/compare order
// of [cat hit kick] (sets of equal size compare like their descending rank lists).
#pragma once
#include <cstdint>

// The functions are plain integer code; tests/host_check.cpp compiles this header with g++ to check the
// formulation against the oracle exhaustively on a box without a GPU.  The product only ever calls them
// from kernels.
#if defined(__CUDACC__)
#define MCP_HD __host__ __device__ __forceinline__
#else
#define MCP_HD static inline
#endif

namespace mcp {

MCP_HD int popc32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}
MCP_HD int popc64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    return __popcll(x);
#else
    return __builtin_popcountll(x);
#endif
}
MCP_HD uint32_t top_bit_index(uint32_t x) {   // x != 0
#if defined(__CUDA_ARCH__)
    return 31 - __clz(x);
#else
    return 31 - __builtin_clz(x);
#endif
}

MCP_HD uint32_t drop_lowest(uint32_t x, int d) {   // clear the d (<= 2) lowest set bits
    if (d > 0) x &= x - 1;
    if (d > 1) x &= x - 1;
    return x;
}

// five-in-a-row detector: bit i of the result <=> ranks i..i+4 all present
MCP_HD uint32_t runs5(uint32_t m) {
    uint32_t t = m & (m >> 1);
    uint32_t u = t & (t >> 2);
    return u & (m >> 4);
}

// Generic evaluator: valid for 5, 6 and 7 cards (no check here; callers validate).
MCP_HD uint32_t eval_packed(uint32_t lo, uint32_t hi) {
    // rank-count planes from the four suit lanes (hand_evaluator.clj:46-55 n-of-a-kinds, :71-79 n-pairs?)
    const uint32_t A = lo & hi, O = lo | hi;
    const uint32_t Ah = A >> 16, Oh = O >> 16;
    const uint32_t o  = (O | Oh) & 0xFFFFu;                 // rank present
    const uint32_t c2 = (A | Ah | (O & Oh)) & 0xFFFFu;      // rank held >= 2 times
    const uint32_t c3 = (A & Oh) | (Ah & O);                // >= 3 times
    const uint32_t c4 = A & Ah;                             // 4 times
    // flush lane (flush? :42-44): at most one suit can hold 5 of <= 7 cards
    const uint32_t l0 = lo & 0xFFFFu, l1 = lo >> 16, l2 = hi & 0xFFFFu, l3 = hi >> 16;
    uint32_t fm = 0;
    if (popc32(l0) >= 5) fm = l0;
    if (popc32(l1) >= 5) fm = l1;
    if (popc32(l2) >= 5) fm = l2;
    if (popc32(l3) >= 5) fm = l3;
    const bool is_flush = fm != 0;

    // rank-pattern category (evaluator :119-133 order; quads/full house cannot coexist with flush/straight)
    const int npair = popc32(c2);
    int cat = npair < 2 ? npair : 2;
    uint32_t Pbase = c2;
    if (c3) { cat = npair >= 2 ? 6 : 3; if (npair >= 2) Pbase = c3; }
    if (c4) { cat = 7; Pbase = c4; }
    // hit: keep the top kp ranks of Pbase (two pairs: 2, otherwise 1; single-bit planes are unchanged)
    uint32_t P = drop_lowest(Pbase, popc32(Pbase) - (cat == 2 ? 2 : 1));
    if (cat == 0) P = 0;
    uint32_t Sbase = (cat == 6) ? c2 : o;
    // kick: top ks ranks of what is left.  ks by cat 0..7 = 5,3,1,2,-,-,1,1
    int ks = (0x11002135u >> (4 * cat)) & 15;
    uint32_t x = Sbase ^ P;

    // straight? :32-40 on the flush lane if there is one (straight flush), else on all ranks
    const uint32_t run = runs5(is_flush ? fm : o);
    if (is_flush) { cat = 5; P = 0; x = fm; ks = 5; }
    uint32_t S = drop_lowest(x, popc32(x) - ks);
    if (run) { cat = is_flush ? 8 : 4; P = 0; S = (run & ~(run >> 1)) << 4; }
    return ((uint32_t)cat << 28) | (P << 13) | (S >> 1);
}

MCP_HD uint32_t eval_packed(uint64_t m) {
    return eval_packed((uint32_t)m, (uint32_t)(m >> 32));
}

// true iff m is a legal 5..7-card set with ranks 1..14
MCP_HD bool valid_hand(uint64_t m) {
    const int n = popc64(m);
    return (m & 0x8001800180018001ull) == 0 && n >= 5 && n <= 7;
}

// packed key -> nibble key cat<<20 | r1..r5 (the public, directly decodable format)
MCP_HD uint32_t packed_to_nibble(uint32_t k) {
    const uint32_t cat = k >> 28;
    uint32_t P = (k >> 14) & 0x3FFFu, S = k & 0x3FFFu;
    // straights: S holds the top rank h -> h, h-1, .., h-4
    auto top = [](uint32_t &m) -> uint32_t { const uint32_t b = top_bit_index(m); m ^= 1u << b; return b + 1; };
    uint32_t out = cat << 20;
    if (cat == 8 || cat == 4) { const uint32_t h = top(S); return out | (h * 0x11111u - 0x01234u); }
    // multiplicity of each P rank by category: 7 -> 4, 6/3 -> 3, 2/1 -> 2 ; S ranks: 2 for full house else 1
    const uint32_t pm = (cat == 7) ? 4 : ((cat == 6 || cat == 3) ? 3 : 2);
    const uint32_t sm = (cat == 6) ? 2 : 1;
    int shift = 20;
    while (P) { const uint32_t r = top(P); for (uint32_t i = 0; i < pm; i++) { shift -= 4; out |= r << shift; } }
    while (S && shift > 0) { const uint32_t r = top(S); for (uint32_t i = 0; i < sm && shift > 0; i++) { shift -= 4; out |= r << shift; } }
    return out;
}

}  // namespace mcp
