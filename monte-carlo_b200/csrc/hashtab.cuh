// hashtab.cuh -- additive-key / perfect-hash evaluator for 7-card hands of the 52-card deck.
//
// Same values as eval.cuh (hand_evaluator.clj:112-133,162-172 semantics) but built for kernels that
// assemble hands incrementally (enumeration, board completion, Monte Carlo deals): a hand is carried as
//     t   = sum over its cards of RKM[rank]        (32-bit wrap-around; identifies the rank multiset)
//     sn  = 0x3333 + sum of 1 << 4*suit            (four biased suit counters: bit 3 of a nibble <=> >= 5 cards)
//     m   = the card mask (only read when a flush is present)
// so adding a card is two integer adds, and a 7-card value is
//     no flush:  V[(t >> 16) + D[(t >> 1) & 8191]]            two shared-memory loads
//     flush:     F[lane of the flush suit >> 2]               (a flush outranks every rank pattern that can
//                                                              coexist with it in 7 cards: at most trips)
// RKM[r] = RK[r] * HT_MUL mod 2^32 where RK are thirteen integers whose 7-multiset sums (<= 4 per rank) are all
// distinct (tools/find_rank_keys.cpp; max sum 7,825,759 < 2^23), so t = (sum RK) * HT_MUL is a bijective
// scramble of the multiset id whose bits 1..13 pick a bucket and bits 16..31 a slot; D holds the per-bucket
// displacement found by first-fit (hash-and-displace, built at mc_init by hashtab_build.h).
//
// A table value ("class code") is cat << 12 | index of the value inside its category in ascending order, so
// unsigned order of codes == clojure.core/compare order of [cat hit kick], and cat = code >> 12.
#pragma once
#include <cstdint>
#include "eval.cuh"

namespace mcp {

constexpr uint32_t RK_RAW[13] = {0, 1, 5, 22, 98, 453, 2031, 8698, 22854, 83661, 262349, 636345, 1479181};
constexpr int HT_BUCKETS = 8192;                 // D entries (u16)
constexpr int HT_SLOTS = 65536 + 2048;           // V entries (u16): slot = (t >> 16) + displacement, no wrap
constexpr int HT_MAX_DISP = 2047;
constexpr int HT_FLUSH = 8192;                   // F entries (u16), index = suit lane >> 2 (ranks 2..14)
constexpr int HT_N_MULTISETS = 49205;            // 7-card rank multisets with <= 4 of a rank
constexpr uint32_t SN_BIAS = 0x3333u;

// one contiguous blob in global memory, copied to shared memory by the kernels that use it
struct HashBlob {
    uint16_t D[HT_BUCKETS];
    uint16_t F[HT_FLUSH];
    uint16_t V[HT_SLOTS];
};
static_assert(sizeof(HashBlob) % 16 == 0, "blob is copied with 16-byte loads");

MCP_HD uint32_t ht_bucket(uint32_t t) { return (t >> 1) & (HT_BUCKETS - 1); }
MCP_HD uint32_t ht_slot(uint32_t t, uint32_t disp) { return (t >> 16) + disp; }


// per-card constants by bit position p = 16*suit + rank: .x = premultiplied rank key, .y = suit nibble increment
struct CardKey { uint32_t rk, sn; };

// class code of a 7-card hand from its incremental state.  D, F, V: shared-memory (or host) tables.
MCP_HD uint32_t ht_class(uint32_t t, uint32_t sn, uint32_t mlo, uint32_t mhi,
                         const uint16_t *D, const uint16_t *F, const uint16_t *V) {
    const uint32_t flag = sn & 0x8888u;
    if (flag) {
        // flag is exactly one of 0x8, 0x80, 0x800, 0x8000 for <= 9 cards
        const uint32_t w = (flag & 0x8800u) ? mhi : mlo;           // suits 2,3 live in the high word
        const uint32_t lane = (flag & 0x8080u) ? (w >> 16) : (w & 0xFFFFu);
        return F[lane >> 2];
    }
    return V[ht_slot(t, D[ht_bucket(t)])];
}

// ---- K1 layout ("stream blob"): everything the streaming kernel needs, in one shared-memory image ------------
// [0, 64 KB)   256 rows of 256 B, row r = byte value r of a 16-bit rank set (a suit lane of the card mask, or a
//              rank-count bit plane taken over the four lanes, which is how eval7_stream_hash_kernel reads it):
//                  first half   TL[r] replicated for the 32 lanes (bank == lane: conflict-free for any mix of r);
//                               TL[r] = sum of RKM[rank] over the bits of r read as ranks 0..7 (only r % 4 == 0 occur)
//                  second half  TH[r] likewise for ranks 8..15 (only r < 128 occur)
//              the halves that can never be addressed hold the nibble keys (first halves of rows r % 4 != 0, 32 keys
//              each) and the flush table (second halves of rows r >= 128, 64 u16 each)
// [64 KB, +16 KB)  D as in HashBlob
// [80 KB, ...)     V', same slots as HashBlob::V but holding the BYTE ADDRESS (< 65536) of the hand's nibble key
// t is linear in the cards, so with count = c0 + 2*c1 + 4*c2 per rank, t = T(c0) + 2*T(c1) + 4*T(c2): a 7-card hand
// without quads costs 4 conflict-free loads (t), D, V' and the key itself.
constexpr uint32_t SB_ROWS_BYTES = 65536;
constexpr uint32_t SB_D_OFF = SB_ROWS_BYTES;
constexpr uint32_t SB_V_OFF = SB_D_OFF + HT_BUCKETS * 2;
constexpr uint32_t SB_BYTES = SB_V_OFF + HT_SLOTS * 2;
static_assert(SB_BYTES % 16 == 0, "stream blob is copied with 16-byte loads");
MCP_HD uint32_t sb_key_addr(uint32_t dense) { const uint32_t f = dense >> 5; return ((f / 3u) * 4u + f % 3u + 1u) * 256u + (dense & 31u) * 4u; }
MCP_HD uint32_t sb_flush_addr(uint32_t fi) { return (128u + (fi >> 6)) * 256u + 128u + (fi & 63u) * 2u; }

}  // namespace mcp
