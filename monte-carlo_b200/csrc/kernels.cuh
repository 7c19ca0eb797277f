// kernels.cuh -- launch wrappers of the sm_100a kernels (kernels.cu), called by the C ABI (capi.cu).
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>
#include "hashtab.cuh"

namespace mcp {

constexpr int MAX_SEATS = 9;

// K1 variants (DESIGN.md "K1"): how the public nibble key is produced
enum K1Mode : int {
    K1_PACKED = 0,        // packed key, no conversion
    K1_NIBBLE_TABLE = 1,  // nibble key through the shared-memory rank-set table (v2; unaligned buffers and A/B)
    K1_NIBBLE_FLO = 2,    // nibble key through FLO/clz loops, no table (A/B variant)
    K1_NIBBLE_HASH = 3,   // nibble key through the plane-hash tables (eval7_stream_hash_kernel; default for aligned buffers)
};

struct DeviceInfo { int sm_count; int device; };

cudaError_t launch_eval7_stream(const uint64_t *d_masks, size_t n, uint32_t *d_keys, int mode,
                                const DeviceInfo &di, cudaStream_t st);
struct DeviceTables;
// K1, plane-hash kernel.  Requires 16-byte aligned masks and 8-byte (keys) / 4-byte (classes) aligned output.
// class16: uint16 global class ids instead of uint32 nibble keys; stage_cards: 5 or 6 = the batch is flop / turn hands
// (Lay56 image), anything else = river hands (Lay7 image); hands of another size than the image serves are still
// evaluated correctly, on the slow path.
cudaError_t launch_eval_stream_hash(const uint64_t *d_masks, size_t n, void *d_out, bool class16, int stage_cards,
                                    const DeviceTables &tb, const DeviceInfo &di, cudaStream_t st);
// colex ranks of 7-card hands over card ids (sum_i C(c_i, i+1), < C(52,7)) -> card masks; a rank out of range gives
// the mask 1 (not a card), which every K1 variant maps to the invalid key / class
cudaError_t launch_unrank_colex32(const uint32_t *d_ranks, size_t n, uint64_t *d_masks, const DeviceInfo &di, cudaStream_t st);
// scratch of the allocation-free colex32 call: a file-scope __device__ array of NOALLOC_MAX_HANDS masks in kernels.cu;
// its device address on the CURRENT device (cudaGetSymbolAddress; the array's name is not an address in host code)
constexpr size_t NOALLOC_MAX_HANDS = (size_t)1 << 20;
cudaError_t noalloc_scratch_address(uint64_t **d_ptr);
// the same ranks as a 27-bit little-endian bit stream (hand j = bits [27 j, 27 j + 27)); d_stream 4-byte aligned
cudaError_t launch_unrank_colex27(const uint32_t *d_stream, size_t n, uint64_t *d_masks, const DeviceInfo &di, cudaStream_t st);
cudaError_t launch_random_hands(uint64_t seed, uint64_t first, size_t n, int n_cards, uint64_t *d_masks,
                                const DeviceInfo &di, cudaStream_t st);
cudaError_t launch_unpack48(const unsigned char *d_packed, size_t n, uint64_t *d_masks, const DeviceInfo &di, cudaStream_t st);
struct EquityArgs {
    uint64_t hole[MAX_SEATS];   // 0 = random seat (Monte Carlo only)
    uint64_t board;             // known community cards
    uint64_t dead;
    int n_seats;
};
// Side-pot structure for the EV variants of K3 / K4 (update-stacks gameplay.clj:104-116 over the Bet records of
// bet.clj:7-8): pot p holds total[p] = bet * n chips and is played for by the seats of players[p] (bitmask); per trial
// the seats of the pot that hold its best hand receive floor(total / their number) each (:110-112).
constexpr int MAX_POTS = 8;
struct PotArgs {
    int n_pots = 0;
    uint32_t players[MAX_POTS] = {};
    uint32_t total[MAX_POTS] = {};             // chips in the pot; the totals of all pots sum below 2^31
};
// one warp: keys per seat, winner bitmask, tie-group index per seat, number of groups
// d_out: uint32 keys[9], uint32 winners, int32 group_of[9], int32 n_groups  (20 words)
// d_out of these two may be device-mapped pinned host memory: after the results the kernel writes `ticket` into
// 32-bit word SMALL_TICKET_WORD of d_out (system-scope fence in between), so the host can poll instead of
// synchronising the stream
constexpr int SMALL_TICKET_WORD = 126;           // last 64-bit word of the 64-word scratch: no copy of counts ever lands there
cudaError_t launch_showdown(const EquityArgs &a, uint32_t *d_out, uint32_t ticket, cudaStream_t st);
cudaError_t launch_eval_one(uint64_t mask, uint32_t *d_out, uint32_t ticket, cudaStream_t st);

// Monte Carlo with a weighted hole range per seat (deal spec R1, DESIGN.md "Ranges per seat").  Alias-table entry of
// one hole of a seat's list: its two cards (bit positions), the slot it shares its capacity with, and how much of the
// capacity W is its own.
struct RangeEntry { unsigned char c0, c1; uint16_t alias; uint32_t thr; };
struct RangeSeat { uint32_t off, n, W, zone_n, zone_W; };      // entries [off, off + n); zone_x = 2^32 mod x (Lemire)
struct McRangeArgs {
    uint64_t hole[MAX_SEATS];        // fixed seats; 0 = ranged
    RangeSeat rs[MAX_SEATS];         // n = 0 for a fixed seat
    uint64_t board, used0;           // known community cards; board | dead | fixed holes
    unsigned char live[52];          // bit positions of the cards outside used0, ascending
    uint32_t n_live, zone_live;
    int n_seats;
};
constexpr int RANGE_MAX_ATTEMPTS = 4096;
// d_counts: win[9] tie[9] loss[9] void (28 int64, accumulated into)
cudaError_t launch_equity_mc_ranges(const McRangeArgs &a, const RangeEntry *d_entries, uint64_t seed, uint64_t first_trial,
                                    uint64_t n_trials, long long *d_counts, const DeviceTables &tb, const DeviceInfo &di, cudaStream_t st);

// preflop matrix: matchup list = (hole_a, hole_b, cell index, weight); each block takes matchups
struct Matchup { uint64_t hole_a, hole_b; int32_t cell_ab, cell_ba; int64_t weight; };

// ---- kernels on the additive-key / perfect-hash evaluator (kernels_hash.cu, hashtab.cuh) ----------------
struct CatBase { uint32_t v[10]; };          // dense class index of the first class of each category; v[9] = #classes
struct StreamImage { const void *d_key = nullptr, *d_cls = nullptr; StreamHash hash{}; };   // the two flavours of one layout
struct DeviceTables {
    const HashBlob *d_blob = nullptr;        // device: D / F / V
    const CardKey *d_card = nullptr;         // device: 64 entries by bit position
    const uint32_t *d_nib = nullptr;         // device: dense class index -> nibble key
    StreamImage img7, img56;                 // device: shared-memory images of the streaming kernel (K1): river; flop / turn
    const uint32_t *d_class_nib = nullptr;   // device: global class id -> nibble key (N_HAND_CLASSES, ascending)
    const CardKey *h_card = nullptr;         // host copy of d_card
    CatBase cat_base{};
    uint32_t n_classes = 0;
};
cudaError_t launch_enum7_hash(uint64_t first, uint64_t count, unsigned long long *d_out /* hist[9], nibble key sum */,
                              const DeviceTables &tb, const DeviceInfo &di, cudaStream_t st);
// d_counts: win[9] tie[9] loss[9] (27 int64, accumulated into); with pots (n_pots > 0) also chips[9] at [27..36)
cudaError_t launch_equity_exhaustive_hash(const EquityArgs &a, uint64_t first_board, uint64_t n_boards, long long *d_counts,
                                          const DeviceTables &tb, const DeviceInfo &di, cudaStream_t st,
                                          const PotArgs *pots = nullptr);
cudaError_t launch_equity_mc_fy(const EquityArgs &a, uint64_t seed, uint64_t first_trial, uint64_t n_trials,
                                long long *d_counts, const DeviceTables &tb, const DeviceInfo &di, cudaStream_t st,
                                const PotArgs *pots = nullptr);
cudaError_t launch_preflop_hash(const Matchup *d_matchups, int n_matchups, long long *d_wtl,
                                const DeviceTables &tb, const DeviceInfo &di, cudaStream_t st);

// d_out: win, tie, loss of side A and the total, weighted by Matchup::weight (4 x int64, accumulated into)
cudaError_t launch_range_equity(const Matchup *d_matchups, int n_matchups, uint64_t board, uint64_t dead, long long *d_out,
                                const DeviceTables &tb, const DeviceInfo &di, cudaStream_t st);

// config 5 by boards (kernels_boards.cu): d_win = 169 x 169 u64, win[a][b] += weight * #{hole of class a
// beats a disjoint hole of class b on the board}; d_sorted / d_info: PB_STRIDE words of scratch per board
struct PfBoard { unsigned char card[5]; unsigned char weight; unsigned char pad[2]; };   // bit positions, orbit size
constexpr int PB_STRIDE = 1088;
cudaError_t launch_preflop_boards(const PfBoard *d_boards, int n_boards, uint32_t *d_sorted, uint32_t *d_info,
                                  unsigned long long *d_win, const DeviceTables &tb, const DeviceInfo &di, cudaStream_t st);

uint64_t host_binom(int n, int k);

}  // namespace mcp
