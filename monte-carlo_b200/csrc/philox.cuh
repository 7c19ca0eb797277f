// philox.cuh -- Philox4x32-10 (Salmon et al., SC'11) and the deal spec shared by every kernel that draws
// cards.  The reference deals with an unseeded java.util.Collections.shuffle (board.clj:148,
// gameplay.clj:145) and has no reproducible stream, so the spec is this engine's own (DESIGN.md "Deal
// spec"); the CPU oracle implements the same spec independently (oracle/ref_fast.c).
//
//   block b of stream (seed, index, tag):  Philox(key = (seed_lo, seed_hi), ctr = (index_lo, index_hi, b, tag))
//   6-bit values of a block:               (w_j >> 6i) & 63 for j = 0..3, i = 0..4   (20 per block)
//   a value is a bit position in the 64-bit card mask; it is accepted iff that bit is not in `used`
//   (`used` starts as every non-card bit | dead | known cards), and is then added to `used`.
#pragma once
#include <cstdint>

namespace mcp {

constexpr uint32_t PHILOX_M0 = 0xD2511F53u, PHILOX_M1 = 0xCD9E8D57u;
constexpr uint32_t PHILOX_W0 = 0x9E3779B9u, PHILOX_W1 = 0xBB67AE85u;
constexpr uint32_t TAG_DEAL = 0x4D434445u;     // Monte Carlo trials
constexpr uint32_t TAG_HAND = 0x4D434831u;     // synthetic random hands (BASELINE config 1)
constexpr uint64_t DECK_MASK = 0x7FFC7FFC7FFC7FFCull;   // the 52 cards of COMPLETE-DECK (card.clj:10-13)
constexpr uint64_t VALID_MASK = 0x7FFE7FFE7FFE7FFEull;  // ranks 1..14 (rank 1 only in the reference's tests)

__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                              uint32_t k0, uint32_t k1, uint32_t (&out)[4]) {
#pragma unroll
    for (int r = 0; r < 10; r++) {
        const uint32_t hi0 = __umulhi(PHILOX_M0, c0), lo0 = PHILOX_M0 * c0;
        const uint32_t hi1 = __umulhi(PHILOX_M1, c2), lo1 = PHILOX_M1 * c2;
        c0 = hi1 ^ c1 ^ k0; c1 = lo1;
        c2 = hi0 ^ c3 ^ k1; c3 = lo0;
        k0 += PHILOX_W0; k1 += PHILOX_W1;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

}  // namespace mcp
