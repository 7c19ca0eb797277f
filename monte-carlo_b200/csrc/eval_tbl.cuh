// eval_tbl.cuh -- table-assisted evaluator: card set -> public nibble key in one pass.
//
// Same semantics as eval.cuh (hand_evaluator.clj:112-133,162-172) but built for the measured
// pipe mix of sm_100a (profiles/r01_pipebench.txt: ALU pipe 64 lanes/clk/SM is the binding resource, the
// FMA pipe and the LSU idle).  The "top-k ranks of a set" selections, the straight test and the rank ->
// nibble conversion all come from ONE shared-memory table indexed by a rank set, so the ALU work left is
// the rank-count planes and the category selects:
//
//   T[x]  (x = rank set)   bits 31..12  the ranks of x in descending order as nibbles r1..r5 (at most 5)
//                          bits  3..0   top rank of the highest 5-run in x (straight? :32-40, no wheel), else 0
//
//   key = base[cat] + top1(T[P]) * MA[cat] + top2nd(T[P]) * MB[cat] + (T[S] >> SH[cat]) * MS[cat]
//
// RB = 14: rank sets over ranks 1..14 (bit r of a lane -> index bit r-1), 16384 entries (64 KB)
// RB = 13: ranks 2..14 only (the 52-card deck), 8192 entries (32 KB)
#pragma once
#include <cstdint>
#include "eval.cuh"

namespace mcp {

template <int RB> struct NibTable {
    static constexpr int ENTRIES = 1 << RB;
    static constexpr int LOW = 15 - RB;                 // 1 (ranks 1..14) or 2 (ranks 2..14): lowest rank bit in a lane
    static constexpr size_t BYTES = (size_t)ENTRIES * 4 + 9 * 16;
};

// entry for rank set `idx` whose bit i stands for rank i + low
MCP_HD uint32_t nib_entry(uint32_t idx, int low) {
    uint32_t m = idx, v = 0;
    int shift = 28;
    while (m && shift >= 12) {
        const uint32_t b = top_bit_index(m);
        m ^= 1u << b;
        v |= (b + low) << shift;
        shift -= 4;
    }
    uint32_t run = idx & (idx >> 1) & (idx >> 2) & (idx >> 3) & (idx >> 4);   // bit i: ranks i..i+4 (+low)
    if (run) v |= top_bit_index(run) + low + 4;
    return v;
}

// per-category constants {MA, MB, MS | SH<<24, base}
MCP_HD void cat_consts(int cat, uint32_t (&c)[4]) {
    const uint32_t MA[9] = {0, 0x11000u, 0x11000u, 0x11100u, 0, 0, 0x11100u, 0x11110u, 0};
    const uint32_t MB[9] = {0, 0, 0x110u, 0, 0, 0, 0, 0, 0};
    const uint32_t MS[9] = {1, 1, 1, 1, 0x11111u, 1, 0x11u, 1, 0x11111u};
    const uint32_t SH[9] = {12, 20, 28, 24, 0, 12, 28, 28, 0};
    const uint32_t SUB[9] = {0, 0, 0, 0, 0x1234u, 0, 0, 0, 0x1234u};
    c[0] = MA[cat]; c[1] = MB[cat]; c[2] = MS[cat] | (SH[cat] << 24); c[3] = ((uint32_t)cat << 20) - SUB[cat];
}

#if defined(__CUDACC__)
template <int RB>
__device__ __forceinline__ void fill_nib_table_v2(uint32_t *T, uint4 *cc) {
    for (uint32_t x = threadIdx.x; x < (uint32_t)NibTable<RB>::ENTRIES; x += blockDim.x) T[x] = nib_entry(x, NibTable<RB>::LOW);
    if (threadIdx.x < 9) {
        uint32_t c[4];
        cat_consts(threadIdx.x, c);
        cc[threadIdx.x] = make_uint4(c[0], c[1], c[2], c[3]);
    }
}
#endif

// T lookup by a lane-aligned rank plane (bit r = rank r): byte offset = plane << (2 - LOW), no shift needed
template <int RB>
MCP_HD uint32_t tbl(const uint32_t *T, uint32_t plane) {
    constexpr int LOW = NibTable<RB>::LOW;
    return *reinterpret_cast<const uint32_t *>(reinterpret_cast<const char *>(T) + (LOW == 2 ? plane : plane * 2u));
}

// 5..7 cards -> nibble key.  cc = the nine uint4 category constants (as 4 words each).
template <int RB>
MCP_HD uint32_t eval_nibble_tbl(uint32_t lo, uint32_t hi, const uint32_t *T, const uint32_t *cc) {
    const uint32_t A = lo & hi, O = lo | hi;
    const uint32_t Ah = A >> 16, Oh = O >> 16;
    const uint32_t o  = (O | Oh) & 0xFFFFu;
    const uint32_t c2 = (A | Ah | (O & Oh)) & 0xFFFFu;
    const uint32_t c3 = (A & Oh) | (Ah & O);
    const uint32_t c4 = A & Ah;
    // flush lane: only the word holding >= 5 cards can contain it
    const int p01 = popc32(lo), p23 = popc32(hi);
    const bool up = p23 >= 5;
    const uint32_t w = up ? hi : lo;
    const int pw = up ? p23 : p01;
    const uint32_t wl = w & 0xFFFFu, wh = w >> 16;
    const int pl = popc32(wl), ph = pw - pl;
    const uint32_t fm = pl >= 5 ? wl : (ph >= 5 ? wh : 0u);

    // rank pattern (evaluator :119-133): planes with their lowest bit(s) cleared tell how many ranks they hold
    const uint32_t y2 = c2 & (c2 - 1), z2 = y2 & (y2 - 1), y3 = c3 & (c3 - 1);
    const bool full = (c3 != 0) && (y2 != 0);
    uint32_t P = z2 ? y2 : c2;                       // pair / two pair (top two of three) / lone trips (c2 == c3)
    int cat = (c2 != 0) + (y2 != 0);
    if (c3) cat = 3;
    if (full) { cat = 6; P = y3 ? y3 : c3; }
    if (c4) { cat = 7; P = c4; }
    const uint32_t x = (full && !c4 ? c2 : o) ^ P;

    const uint32_t b1 = tbl<RB>(T, fm ? fm : o);
    const uint32_t a = tbl<RB>(T, P), bx = tbl<RB>(T, x);
    const uint32_t h = b1 & 15u;                     // top of a straight in the flush lane / in all ranks
    if (fm) cat = 5;
    if (h) cat = fm ? 8 : 4;
    const uint32_t *c = cc + 4 * cat;
    const uint32_t ms = c[2];
    const uint32_t b = fm ? b1 : bx;
    const uint32_t Bv = h ? h : (b >> (ms >> 24));
    return c[3] + (a >> 28) * c[0] + ((a >> 24) & 15u) * c[1] + Bv * (ms & 0xFFFFFFu);
}

}  // namespace mcp
