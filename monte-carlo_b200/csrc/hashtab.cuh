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
// A table value ("class code") is the dense index of the value in the ascending list of the 4,892 values seven cards
// can make, so unsigned order of codes == clojure.core/compare order of [cat hit kick]; nib_keys[code] is the public
// key and cat_base[] (first code of each category) gives the category.
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
// The streaming kernel does not sum per-card keys.  It adds the four suit lanes of the mask into per-rank count bit
// planes (count = c0 + 2*c1 + 4*c2, sh_planes) and hashes the planes themselves: (c0, c1, c2) IS the rank multiset,
// so u = c0 | c1 << 16 (plus the rare quads plane c2) is an injective 32-bit name of it and two multiplications
// give the two hash words -- no table is read before the displacement:
//     t = u * m1 + c2 * m3        slot   = (t >> 16) + D[bucket]
//     b = u * m2 + c2 * m4        bucket = b >> 19
// [SB_KEY_OFF, +20 KB)  the 4,892 nibble keys, dense (ascending)
// [SB_F_OFF,   +16 KB)  flush table: u16 byte address of the key, by (flush lane >> 2)
// [SB_D_OFF,   +16 KB)  D' = 2 * displacement (in bytes of V'), by bucket
// [SB_V_OFF,   ...)     V', HT_SLOTS u16: byte address of the hand's nibble key
// so a 7-card hand costs three data-dependent loads (D', V', key), plus the flush entry when one lane holds >= 5.
struct StreamHash { uint32_t m1, m2, m3, m4; };
// Two images with the same structure:
//   Lay7   7-card hands (river): 49,205 rank multisets, 4,892 values; slot = (t >> 16) + D          184 KB
//   Lay56  5- and 6-card hands (flop / turn, hand_evaluator.clj:171 on 5 or 6 cards): 6,175 + 18,395 multisets, up to
//          7,462 values; slot = (t >> 17) + D                                                       ~130 KB
// Each image exists twice on the device: "key" (the first region holds nibble keys) and "class" (the same region holds
// the 16-bit global class id of the value: its index in the ascending list of all 7,462 five-card values,
// mc_class_keys), so the last load of the kernel delivers whichever output format the call asked for.
struct Lay7 {
    static constexpr int N_MIN = 7, N_MAX = 7;                   // card counts the image serves
    static constexpr uint32_t N_KEYS = 5120;                     // capacity of the first region (4,892 used)
    static constexpr uint32_t SLOT_MUL = 65536u;                 // slot = umulhi(t, SLOT_MUL) + displacement
    static constexpr uint32_t N_SLOTS = 65536u + 2048u;
    static constexpr uint32_t KEY_OFF = 0;
    static constexpr uint32_t F_OFF = KEY_OFF + 4u * N_KEYS;
    static constexpr uint32_t D_OFF = F_OFF + HT_FLUSH * 2;
    static constexpr uint32_t V_OFF = D_OFF + HT_BUCKETS * 2;
    static constexpr uint32_t BYTES = V_OFF + N_SLOTS * 2;
};
struct Lay56 {
    static constexpr int N_MIN = 5, N_MAX = 6;
    static constexpr uint32_t N_KEYS = 7552;                     // 7,462 used
    static constexpr uint32_t SLOT_MUL = 32768u;
    static constexpr uint32_t N_SLOTS = 32768u + 2048u;
    static constexpr uint32_t KEY_OFF = 0;
    static constexpr uint32_t F_OFF = KEY_OFF + 4u * N_KEYS;
    static constexpr uint32_t D_OFF = F_OFF + HT_FLUSH * 2;
    static constexpr uint32_t V_OFF = D_OFF + HT_BUCKETS * 2;
    static constexpr uint32_t BYTES = V_OFF + N_SLOTS * 2;
};
static_assert(Lay7::BYTES % 16 == 0 && Lay56::BYTES % 16 == 0, "stream images are copied in 16-byte units");
static_assert(Lay7::V_OFF < 65536u && Lay56::V_OFF < 65536u, "key addresses are 16-bit");
constexpr int N_HAND_CLASSES = 7462;                             // distinct five-card values
constexpr uint32_t SB_KEY_OFF = Lay7::KEY_OFF, SB_F_OFF = Lay7::F_OFF, SB_D_OFF = Lay7::D_OFF, SB_V_OFF = Lay7::V_OFF;
constexpr uint32_t SB_BYTES = Lay7::BYTES;
MCP_HD uint32_t sb_key_addr(uint32_t dense) { return SB_KEY_OFF + 4u * dense; }
MCP_HD uint32_t sb_flush_addr(uint32_t fi) { return SB_F_OFF + 2u * fi; }
// rank-count bit planes of a card mask (carry-save addition of the four 16-bit suit lanes); low 16 bits valid
MCP_HD void sh_planes(uint32_t lo, uint32_t hi, uint32_t *c0, uint32_t *c1, uint32_t *c2) {
    const uint32_t x = lo ^ hi, y = lo & hi, xs = x >> 16, ys = y >> 16;
    *c0 = (x ^ xs) & 0xFFFFu;
    *c1 = ((x & xs) ^ y ^ ys) & 0xFFFFu;
    *c2 = y & ys;
}
MCP_HD uint32_t sh_t(uint32_t c0, uint32_t c1, uint32_t c2, const StreamHash &S) { return (c0 | c1 << 16) * S.m1 + c2 * S.m3; }
MCP_HD uint32_t sh_b(uint32_t c0, uint32_t c1, uint32_t c2, const StreamHash &S) { return (c0 | c1 << 16) * S.m2 + c2 * S.m4; }
MCP_HD uint32_t sh_bucket(uint32_t b) { return b >> 19; }
template <class LAY> MCP_HD uint32_t sh_slot(uint32_t t) { return (uint32_t)(((uint64_t)t * LAY::SLOT_MUL) >> 32); }
static_assert(HT_BUCKETS == 1 << 13, "sh_bucket takes the top 13 bits");

}  // namespace mcp
