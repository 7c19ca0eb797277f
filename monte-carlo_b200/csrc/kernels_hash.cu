// kernels_hash.cu -- the enumeration / equity kernels on the additive-key + perfect-hash evaluator (hashtab.cuh).
//   K2 enum7_hash          all C(52,7) hands -> category histogram + nibble-key checksum   (BASELINE config 2)
//   K3 equity_exh_hash     every completion of the board -> win/tie/loss per seat          (config 3; gameplay.clj:127-132)
//   K5 preflop_hash        suit-canonical matchups x all C(48,5) boards                    (config 5)
//   K4 equity_mc_fy        Philox Fisher-Yates deal per trial + showdown                   (config 4; board.clj:148,
//                                                                                           gameplay.clj:63-75,41-54,127-132)
// One block per SM: the 164 KB hash blob lives in shared memory; a hand is (t, sn, mask) and grows by one card
// with two integer adds; a 7-card value is two shared-memory loads (hashtab.cuh).  No tensor cores, no HBM
// traffic besides the table staging and <= 27 64-bit atomics per block.
#include "kernels.cuh"
#include "common.cuh"
#include "hashtab.cuh"
#include "philox.cuh"
#include <atomic>
#include <cstring>

namespace mcp {

struct HandState { uint32_t t, sn, mlo, mhi; };
__device__ __forceinline__ HandState with_card(const HandState &h, const uint4 &c) {
    return HandState{h.t + c.x, h.sn + c.y, h.mlo | c.z, h.mhi | c.w};
}
__host__ __device__ __forceinline__ uint4 live_entry(const CardKey &k, int pos) {
    return make_uint4(k.rk, k.sn, pos < 32 ? 1u << pos : 0u, pos >= 32 ? 1u << (pos - 32) : 0u);
}

// Walks K-subsets of the live list in colex order, keeping the partial hands P[i] = base + cards c[i..K-1] in
// registers: the common step (lowest card moves) costs one 16-byte shared load and four integer ops.
// LS = stride of the live list in entries: 1 for a plain list, 32 for a list replicated per lane ([entry][lane],
// `live` then points at the lane's column), which makes the 16-byte loads of a warp conflict-free.
template <int K, int LS = 1>
struct Walker {
    int c[K > 0 ? K : 1];
    HandState P[K + 1];

    __device__ __forceinline__ void init(uint32_t idx, int nl, const uint32_t *sbin, const uint4 *live, const HandState &base) {
        P[K] = base;
        if constexpr (K > 0) {
            colex_unrank<K>(idx, nl, sbin, c);
#pragma unroll
            for (int i = K - 1; i >= 0; i--) P[i] = with_card(P[i + 1], live[c[i] * LS]);
        }
    }
    template <int I>
    __device__ __forceinline__ void bump(const uint4 *live, int nl) {
        if constexpr (I < K) {
            c[I]++;
            if constexpr (I + 1 < K) {
                if (c[I] == c[I + 1]) { c[I] = I; bump<I + 1>(live, nl); }
            } else {
                if (c[I] >= nl) c[I] = nl;           // past the last subset: live[nl] is a zero entry
            }
            P[I] = with_card(P[I + 1], live[c[I] * LS]);
        }
    }
    __device__ __forceinline__ void next(const uint4 *live, int nl) { bump<0>(live, nl); }
    __device__ __forceinline__ const HandState &hand() const { return P[0]; }
};

// per-seat win / tie counters of one thread and their reduction into d_counts = win[9] tie[9] loss[9]
template <int NS>
struct Tally {
    uint32_t win[NS], tie[NS], units;
    __device__ __forceinline__ void clear() {
#pragma unroll
        for (int s = 0; s < NS; s++) { win[s] = 0; tie[s] = 0; }
        units = 0;
    }
    __device__ __forceinline__ void add(const uint32_t (&key)[NS], int n_seats) {
        uint32_t best = 0;
#pragma unroll
        for (int s = 0; s < NS; s++) if (NS == 2 || s < n_seats) best = max(best, key[s]);
        int nbest = 0;
#pragma unroll
        for (int s = 0; s < NS; s++) if (NS == 2 || s < n_seats) nbest += (key[s] == best);
        const bool solo = nbest == 1;
#pragma unroll
        for (int s = 0; s < NS; s++) if (NS == 2 || s < n_seats) {
            const bool b = key[s] == best;
            win[s] += (b && solo);
            tie[s] += (b && !solo);
        }
        units++;
    }
    // sred: 19 zeroed u64 in shared memory; seat_of[s] = output seat of counter s
    __device__ __forceinline__ void flush(int n_seats, unsigned long long *sred, long long *d_counts) {
        const int lane = threadIdx.x & 31;
#pragma unroll
        for (int s = 0; s < NS; s++) if (NS == 2 || s < n_seats) {
            const uint32_t w = warp_sum(win[s]), t = warp_sum(tie[s]);
            if (lane == 0) { if (w) atomicAdd(&sred[s], (unsigned long long)w); if (t) atomicAdd(&sred[9 + s], (unsigned long long)t); }
        }
        const uint32_t u = warp_sum(units);
        if (lane == 0 && u) atomicAdd(&sred[18], (unsigned long long)u);
        __syncthreads();
        if ((int)threadIdx.x < n_seats) {
            const int s = threadIdx.x;
            const unsigned long long w = sred[s], t = sred[9 + s], n = sred[18];
            unsigned long long *o = reinterpret_cast<unsigned long long *>(d_counts);
            if (w) atomicAdd(&o[s], w);
            if (t) atomicAdd(&o[9 + s], t);
            if (n - w - t) atomicAdd(&o[18 + s], n - w - t);
        }
    }
};

// ---------------------------------------------------------------------------------------------------
// EV: the reference's pot settlement per trial (update-stacks gameplay.clj:104-116; every seat of a simulation is at the
// showdown, so "the first tie group that intersects the pot" is the set of its seats holding the pot's best hand).
// Class codes are 16-bit: k1 = code | 0x10000 marks a seat, `in` (0x1FFFF or 0) keeps it if it plays for the pot, and
// masked + (0x20000 - best) carries into bit 17 exactly for the pot's seats at its maximum -- no compares.  The share
// floor(total / k) comes from a table built on the host (no division here); chips go to 32-bit per-thread accumulators
// that are folded into 64-bit shared counters before they can wrap (flush_every trials, computed from the totals).
// ---------------------------------------------------------------------------------------------------
struct PotParams {
    uint32_t in[MAX_POTS][MAX_SEATS];          // by kernel seat
    uint32_t share[MAX_POTS][MAX_SEATS + 1];   // share[p][k] = floor(total / k), k = 1..9
    uint32_t main_share[MAX_SEATS + 1];        // K4: the pots every seat plays for are settled from the showdown itself
    int n_pots, has_main;
    uint32_t flush_every;
};
template <int NSLOT>
__device__ __forceinline__ void pots_settle(const PotParams &pp, const uint32_t (&key)[NSLOT], uint32_t (&chips)[NSLOT]) {
    uint32_t k1[NSLOT];
#pragma unroll
    for (int s = 0; s < NSLOT; s++) k1[s] = key[s] | 0x10000u;
#pragma unroll
    for (int p = 0; p < MAX_POTS; p++) {
        if (p < pp.n_pots) {
            uint32_t mk[NSLOT], best = 0;
#pragma unroll
            for (int s = 0; s < NSLOT; s++) { mk[s] = k1[s] & pp.in[p][s]; best = max(best, mk[s]); }
            const uint32_t cb = 0x20000u - best;
            uint32_t at[NSLOT], k = 0;
#pragma unroll
            for (int s = 0; s < NSLOT; s++) { at[s] = (mk[s] + cb) >> 17; k += at[s]; }
            const uint32_t share = pp.share[p][k];
#pragma unroll
            for (int s = 0; s < NSLOT; s++) chips[s] += at[s] * share;
        }
    }
}
// per-thread fold inside a loop whose trip count differs between the lanes of a warp: plain shared atomics
template <int NSLOT>
__device__ __forceinline__ void chips_fold_thread(uint32_t (&chips)[NSLOT], unsigned long long *schips) {
#pragma unroll
    for (int s = 0; s < NSLOT; s++) { if (chips[s]) atomicAdd(&schips[s], (unsigned long long)chips[s]); chips[s] = 0; }
}
// final fold, every lane of the warp present: two 16-bit halves through REDUX, one shared atomic per seat and warp
template <int NSLOT>
__device__ __forceinline__ void chips_fold_warp(const uint32_t (&chips)[NSLOT], unsigned long long *schips) {
#pragma unroll
    for (int s = 0; s < NSLOT; s++) {
        const uint32_t lo = warp_sum(chips[s] & 0xFFFFu), hi = warp_sum(chips[s] >> 16);
        if ((threadIdx.x & 31) == 0 && (lo | hi)) atomicAdd(&schips[s], (unsigned long long)lo + ((unsigned long long)hi << 16));
    }
}
// host: kernel-seat view of the caller's pots.  seat_of[q] = caller seat of kernel seat q (n_kernel_seats of them)
// use_main: pots played for by EVERY seat go to main_share instead of a slot (their winners are the showdown's own)
static void make_pot_params(const PotArgs *pots, const int *seat_of, int n_kernel_seats, PotParams *pp, bool use_main = false) {
    memset(pp, 0, sizeof *pp);
    if (!pots) return;
    const uint32_t everybody = (1u << n_kernel_seats) - 1u;
    // pots played for by the same seats have the same winners: they share one slot, whose share table is the sum of
    // theirs (floor(a / k) + floor(b / k), each pot truncated on its own as update-stacks does)
    uint64_t sum = 0;
    uint32_t mask_of[MAX_POTS];
    for (int p = 0; p < pots->n_pots; p++) {
        if (use_main && pots->players[p] == everybody) {
            for (int k = 1; k <= MAX_SEATS; k++) pp->main_share[k] += pots->total[p] / (uint32_t)k;
            pp->has_main = 1;
            sum += pots->total[p];
            continue;
        }
        int slot = 0;
        while (slot < pp->n_pots && mask_of[slot] != pots->players[p]) slot++;
        if (slot == pp->n_pots) {
            mask_of[slot] = pots->players[p];
            for (int q = 0; q < n_kernel_seats; q++) pp->in[slot][q] = ((pots->players[p] >> seat_of[q]) & 1u) ? 0x1FFFFu : 0u;
            pp->n_pots++;
        }
        for (int k = 1; k <= MAX_SEATS; k++) pp->share[slot][k] += pots->total[p] / (uint32_t)k;
        sum += pots->total[p];
    }
    const uint64_t f = sum ? 0xFFFFFFFFull / sum : 0xFFFFFFFFull;
    pp->flush_every = (uint32_t)(f > 0x7FFFFFFFull ? 0x7FFFFFFFull : (f < 1 ? 1 : f));
}

constexpr int HASH_THREADS = 512;
constexpr int LIVE_MAX = 56;

// ht_class of hashtab.cuh over a HashBlob staged in shared memory at shared address hb (staged_table_base)
constexpr uint32_t HB_F = HT_BUCKETS * 2, HB_V = HB_F + HT_FLUSH * 2;
static_assert(sizeof(HashBlob) == HB_V + HT_SLOTS * 2, "HashBlob is D, F, V back to back");
// Branch-free: the displacement is always read; a hand with a flush then reads F instead of V through the SAME load
// (its address is selected), so the evaluations of several hands / seats overlap instead of being fenced by a
// divergent branch each.
__device__ __forceinline__ uint32_t hb_slot_addr(uint32_t t, uint32_t hb) {
    const uint32_t d = lds_u16(hb + (t & 0x3FFEu));                       // D[(t >> 1) & 8191]
    return hb + HB_V + 2u * ((t >> 16) + d);
}
__device__ __forceinline__ uint32_t hb_flush_addr(uint32_t lane_ranks, uint32_t hb) { return hb + HB_F + 2u * lane_ranks; }
__device__ __forceinline__ uint32_t hb_class(uint32_t t, uint32_t sn, uint32_t mlo, uint32_t mhi, uint32_t hb) {
    const uint32_t flag = sn & 0x8888u;                                   // one of 0x8, 0x80, 0x800, 0x8000 (or 0) for <= 9 cards
    const uint32_t w = (flag & 0x8800u) ? mhi : mlo;                      // suits 2,3 live in the high word
    const uint32_t lane = (flag & 0x8080u) ? (w >> 16) : (w & 0xFFFFu);
    const uint32_t va = hb_slot_addr(t, hb), fa = hb_flush_addr(lane >> 2, hb);
    return lds_u16(flag ? fa : va);
}

// ---------------------------------------------------------------------------------------------------
// K3: exhaustive equity over all K-card completions of the board
// ---------------------------------------------------------------------------------------------------
struct EqxHashParams {
    uint32_t h_t[MAX_SEATS], h_sn[MAX_SEATS], h_mlo[MAX_SEATS], h_mhi[MAX_SEATS];   // fixed hole cards per seat
    HandState base;                     // known community cards, sn biased
    uint64_t first, count;
    int n_seats, n_live;
    uint4 live[LIVE_MAX];               // ascending bit position; entries >= n_live are zero
    PotParams pots;                     // EV variant only
};

template <int NS> constexpr int eqx_threads() { return NS == 2 ? 1024 : HASH_THREADS; }   // heads-up needs <= 51 registers

template <int K, int NS, bool EV>
__global__ void __launch_bounds__(eqx_threads<NS>(), 1)
equity_exh_hash_kernel(const __grid_constant__ EqxHashParams prm, const HashBlob *__restrict__ g_blob,
                       long long *__restrict__ d_counts) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    HashBlob *H = reinterpret_cast<HashBlob *>(smem_raw);
    uint4 *slive = reinterpret_cast<uint4 *>(smem_raw + sizeof(HashBlob));
    uint32_t *sbin = reinterpret_cast<uint32_t *>(slive + LIVE_MAX);
    __shared__ unsigned long long sred[19 + MAX_SEATS];
    __shared__ unsigned long long mbar;
    BulkStage stage;
    stage.start(&mbar, H, g_blob, (uint32_t)sizeof(HashBlob));
    fill_binom(sbin);
    if (threadIdx.x < LIVE_MAX) slive[threadIdx.x] = prm.live[threadIdx.x];
    if (threadIdx.x < 19 + MAX_SEATS) sred[threadIdx.x] = 0ull;
    __syncthreads();                     // binomials and live list are in place; the hash image is still on its way
    const int n_seats = prm.n_seats, nl = prm.n_live;
    Tally<NS> tl; tl.clear();
    uint32_t chips[NS], since = 0;
#pragma unroll
    for (int s = 0; s < NS; s++) chips[s] = 0;
    const uint64_t nthreads = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t per = (prm.count + nthreads - 1) / nthreads;
    const uint64_t me = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t lo = me * per < prm.count ? me * per : prm.count;
    const uint64_t hi = lo + per < prm.count ? lo + per : prm.count;
    // the colex unranking of the thread's first board needs no hash table: it runs while the TMA engine stages the image
    Walker<K> w;
    if (lo < hi) w.init((uint32_t)(prm.first + lo), nl, sbin, slive, prm.base);
    stage.wait();
    const uint32_t tbl = staged_table_base(H);
    if (lo < hi) {
        for (uint64_t i = lo; i < hi; i++) {
            const HandState b = w.hand();
            uint32_t key[NS];
#pragma unroll
            for (int s = 0; s < NS; s++) {
                key[s] = 0;
                if (NS == 2 || s < n_seats)
                    key[s] = hb_class(b.t + prm.h_t[s], b.sn + prm.h_sn[s], b.mlo | prm.h_mlo[s], b.mhi | prm.h_mhi[s], tbl);
            }
            tl.add(key, n_seats);
            if constexpr (EV) {
                pots_settle<NS>(prm.pots, key, chips);
                if (++since == prm.pots.flush_every) { chips_fold_thread<NS>(chips, sred + 19); since = 0; }
            }
            w.next(slive, nl);
        }
    }
    if constexpr (EV) chips_fold_warp<NS>(chips, sred + 19);
    tl.flush(n_seats, sred, d_counts);
    if constexpr (EV) {           // (tl.flush has synchronised the block)
        if ((int)threadIdx.x < n_seats && sred[19 + threadIdx.x])
            atomicAdd(reinterpret_cast<unsigned long long *>(d_counts) + 27 + threadIdx.x, sred[19 + threadIdx.x]);
    }
}

static size_t eqx_smem() { return sizeof(HashBlob) + LIVE_MAX * sizeof(uint4) + BIN_N * BIN_K * sizeof(uint32_t); }

// (host helpers of the launchers: they walk the set bits only -- a heads-up launch is a few microseconds of device time, so
// the host side of the call counts)
static void fill_live(uint64_t live_mask, const CardKey *card, uint4 *live, int *n_live) {
    int nl = 0;
    for (uint64_t x = live_mask; x; x &= x - 1) { const int b = __builtin_ctzll(x); live[nl++] = live_entry(card[b], b); }
    *n_live = nl;
    for (int i = nl; i < LIVE_MAX; i++) live[i] = make_uint4(0, 0, 0, 0);
}
static HandState state_of(uint64_t mask, const CardKey *card, uint32_t sn0) {
    HandState h{0, sn0, (uint32_t)mask, (uint32_t)(mask >> 32)};
    for (uint64_t x = mask; x; x &= x - 1) { const int b = __builtin_ctzll(x); h.t += card[b].rk; h.sn += card[b].sn; }
    return h;
}

template <int K, int NS, bool EV>
static cudaError_t launch_eqx(const EqxHashParams &prm, const HashBlob *blob, long long *d_counts, int grid, cudaStream_t st) {
    auto kern = equity_exh_hash_kernel<K, NS, EV>;
    // the attribute is per function and device: set once per (instantiation, device), not on every launch
    static std::atomic<uint64_t> attr_done{0};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev >= 64 || !((attr_done.load(std::memory_order_relaxed) >> dev) & 1ull)) {
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)eqx_smem());
        if (e != cudaSuccess) return e;
        if (dev < 64) attr_done.fetch_or(1ull << dev, std::memory_order_relaxed);
    }
    kern<<<grid, eqx_threads<NS>(), eqx_smem(), st>>>(prm, blob, d_counts);
    return cudaGetLastError();
}

cudaError_t launch_equity_exhaustive_hash(const EquityArgs &a, uint64_t first_board, uint64_t n_boards, long long *d_counts,
                                          const DeviceTables &tb, const DeviceInfo &di, cudaStream_t st, const PotArgs *pots) {
    if (n_boards == 0) return cudaSuccess;
    EqxHashParams prm;
    const bool ev = pots && pots->n_pots > 0;
    int ident[MAX_SEATS];
    for (int s = 0; s < MAX_SEATS; s++) ident[s] = s;
    make_pot_params(ev ? pots : nullptr, ident, a.n_seats, &prm.pots);
    uint64_t used = a.board | a.dead;
    for (int s = 0; s < MAX_SEATS; s++) {
        const uint64_t h = s < a.n_seats ? a.hole[s] : 0ull;
        used |= h;
        const HandState hs = state_of(h, tb.h_card, 0);
        prm.h_t[s] = hs.t; prm.h_sn[s] = hs.sn; prm.h_mlo[s] = hs.mlo; prm.h_mhi[s] = hs.mhi;
    }
    prm.base = state_of(a.board, tb.h_card, SN_BIAS);
    prm.first = first_board; prm.count = n_boards; prm.n_seats = a.n_seats;
    fill_live(DECK_MASK & ~used, tb.h_card, prm.live, &prm.n_live);
    const int K = 5 - __builtin_popcountll(a.board);
    const bool two = a.n_seats == 2 && !ev;               // (the EV variant exists in the general shape only)
    const uint64_t tpb = two ? eqx_threads<2>() : eqx_threads<MAX_SEATS>();
    uint64_t want = (n_boards + tpb - 1) / tpb;
    const int grid = (int)(want > (uint64_t)di.sm_count ? (uint64_t)di.sm_count : want);
#define EQX_CASE(k) case k: return ev ? launch_eqx<k, MAX_SEATS, true>(prm, tb.d_blob, d_counts, grid, st) : two ? launch_eqx<k, 2, false>(prm, tb.d_blob, d_counts, grid, st) : launch_eqx<k, MAX_SEATS, false>(prm, tb.d_blob, d_counts, grid, st)
    switch (K) {
        EQX_CASE(0); EQX_CASE(1); EQX_CASE(2); EQX_CASE(3); EQX_CASE(4); EQX_CASE(5);
        default: return cudaErrorInvalidValue;
    }
#undef EQX_CASE
}

// ---------------------------------------------------------------------------------------------------
// K5: preflop matrix by brute force over suit-canonical matchups.  A block takes one matchup at a time; its
// threads split the C(48,5) boards into equal contiguous index ranges (one unranking per thread per matchup).
// ---------------------------------------------------------------------------------------------------
constexpr int PF_THREADS = 1024;          // one block per matchup; 32 warps hide the hash-lookup latency (512: 0.43 s per matrix)
__global__ void __launch_bounds__(PF_THREADS, 1)
preflop_hash_kernel(const Matchup *__restrict__ mus, int n_mu, const HashBlob *__restrict__ g_blob,
                    const CardKey *__restrict__ g_card, long long *__restrict__ wtl) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    HashBlob *H = reinterpret_cast<HashBlob *>(smem_raw);
    uint4 *slive = reinterpret_cast<uint4 *>(smem_raw + sizeof(HashBlob));                  // [LIVE_MAX][32 lanes]
    uint32_t *sbin = reinterpret_cast<uint32_t *>(slive + LIVE_MAX * 32);
    __shared__ uint4 slist[LIVE_MAX];
    __shared__ unsigned long long sred[3];
    __shared__ CardKey scard[64];
    __shared__ unsigned long long mbar;
    BulkStage stage;
    stage.start(&mbar, H, g_blob, (uint32_t)sizeof(HashBlob));
    fill_binom(sbin);
    if (threadIdx.x < 64) scard[threadIdx.x] = g_card[threadIdx.x];
    const uint4 *my_live = slive + (threadIdx.x & 31);
    stage.wait();
    const uint32_t tbl = staged_table_base(H);
    const uint32_t total = cbinom(48, 5);
    const uint32_t per = (total + PF_THREADS - 1) / PF_THREADS;
    for (int mi = blockIdx.x; mi < n_mu; mi += gridDim.x) {
        const Matchup mu = mus[mi];
        if (threadIdx.x < 3) sred[threadIdx.x] = 0ull;
        const uint64_t livem = DECK_MASK & ~(mu.hole_a | mu.hole_b);
        if (threadIdx.x < 64) {
            const int p = threadIdx.x;
            if ((livem >> p) & 1ull) slist[__popcll(livem & ((1ull << p) - 1ull))] = live_entry(scard[p], p);
            if (p >= 48 && p < LIVE_MAX) slist[p] = make_uint4(0, 0, 0, 0);
        }
        __syncthreads();
        for (int i = threadIdx.x; i < LIVE_MAX * 32; i += PF_THREADS) slive[i] = slist[i >> 5];   // one copy per lane
        // the two hands' fixed parts
        HandState ha{0, 0, (uint32_t)mu.hole_a, (uint32_t)(mu.hole_a >> 32)}, hb{0, 0, (uint32_t)mu.hole_b, (uint32_t)(mu.hole_b >> 32)};
        {
            uint64_t x = mu.hole_a;
            while (x) { const int p = __ffsll((long long)x) - 1; x &= x - 1; ha.t += scard[p].rk; ha.sn += scard[p].sn; }
            x = mu.hole_b;
            while (x) { const int p = __ffsll((long long)x) - 1; x &= x - 1; hb.t += scard[p].rk; hb.sn += scard[p].sn; }
        }
        __syncthreads();
        const uint32_t lo = per * threadIdx.x, hi = min(total, lo + per);
        uint32_t wa = 0, wb = 0, tie = 0;
        if (lo < hi) {
            Walker<5, 32> w;
            w.init(lo, 48, sbin, my_live, HandState{0, SN_BIAS, 0, 0});
            for (uint32_t i = lo; i < hi; i++) {
                const HandState b = w.hand();
                const uint32_t ka = hb_class(b.t + ha.t, b.sn + ha.sn, b.mlo | ha.mlo, b.mhi | ha.mhi, tbl);
                const uint32_t kb = hb_class(b.t + hb.t, b.sn + hb.sn, b.mlo | hb.mlo, b.mhi | hb.mhi, tbl);
                wa += (kb - ka) >> 31; wb += (ka - kb) >> 31;      // class codes are 16-bit: the sign bit of the difference
                w.next(my_live, 48);
            }
            tie = (hi - lo) - wa - wb;
        }
        const uint32_t sa = warp_sum(wa), sb = warp_sum(wb), stie = warp_sum(tie);
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(&sred[0], (unsigned long long)sa); atomicAdd(&sred[1], (unsigned long long)sb);
            atomicAdd(&sred[2], (unsigned long long)stie);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned long long *o = reinterpret_cast<unsigned long long *>(wtl);
            const unsigned long long wgt = (unsigned long long)mu.weight;
            atomicAdd(&o[3 * mu.cell_ab + 0], wgt * sred[0]);
            atomicAdd(&o[3 * mu.cell_ab + 1], wgt * sred[2]);
            atomicAdd(&o[3 * mu.cell_ab + 2], wgt * sred[1]);
            atomicAdd(&o[3 * mu.cell_ba + 0], wgt * sred[1]);
            atomicAdd(&o[3 * mu.cell_ba + 1], wgt * sred[2]);
            atomicAdd(&o[3 * mu.cell_ba + 2], wgt * sred[0]);
        }
        __syncthreads();
    }
}

cudaError_t launch_preflop_hash(const Matchup *d_matchups, int n_matchups, long long *d_wtl,
                                const DeviceTables &tb, const DeviceInfo &di, cudaStream_t st) {
    if (n_matchups <= 0) return cudaSuccess;
    const size_t smem = sizeof(HashBlob) + LIVE_MAX * 32 * sizeof(uint4) + BIN_N * BIN_K * sizeof(uint32_t);
    cudaError_t e = cudaFuncSetAttribute(preflop_hash_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const int grid = n_matchups < di.sm_count ? n_matchups : di.sm_count;
    preflop_hash_kernel<<<grid, PF_THREADS, smem, st>>>(d_matchups, n_matchups, tb.d_blob, tb.d_card, d_wtl);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------
// K6: heads-up range vs range on a (partial) board.  One WARP per weighted matchup: its lanes split the
// C(n_live, K) board completions into equal contiguous index ranges; each warp keeps its own live-card list.
// out = win, tie, loss of side A and the total, each weighted by the matchup weight (4 x int64, accumulated into).
// ---------------------------------------------------------------------------------------------------
constexpr int RNG_THREADS = 1024;         // 32 warps per SM, one matchup each (K5 gained 1.45x from 512 -> 1024)
constexpr int RNG_WARPS = RNG_THREADS / 32;

template <int K>
__global__ void __launch_bounds__(RNG_THREADS, 1)
range_equity_kernel(const Matchup *__restrict__ mus, int n_mu, uint64_t board, uint64_t dead,
                    const HashBlob *__restrict__ g_blob, const CardKey *__restrict__ g_card, long long *__restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    HashBlob *H = reinterpret_cast<HashBlob *>(smem_raw);
    uint4 *slive_all = reinterpret_cast<uint4 *>(smem_raw + sizeof(HashBlob));             // [RNG_WARPS][LIVE_MAX]
    uint32_t *sbin = reinterpret_cast<uint32_t *>(slive_all + RNG_WARPS * LIVE_MAX);
    __shared__ CardKey scard[64];
    __shared__ unsigned long long mbar;
    BulkStage stage;
    stage.start(&mbar, H, g_blob, (uint32_t)sizeof(HashBlob));
    fill_binom(sbin);
    if (threadIdx.x < 64) scard[threadIdx.x] = g_card[threadIdx.x];
    stage.wait();
    const uint32_t tbl = staged_table_base(H);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint4 *slive = slive_all + warp * LIVE_MAX;
    HandState base{0, SN_BIAS, (uint32_t)board, (uint32_t)(board >> 32)};
    {
        uint64_t x = board;
        while (x) { const int p = __ffsll((long long)x) - 1; x &= x - 1; base.t += scard[p].rk; base.sn += scard[p].sn; }
    }
    unsigned long long wa = 0, wb = 0, tie = 0, tot = 0;
    const int gw = blockIdx.x * RNG_WARPS + warp, nw = gridDim.x * RNG_WARPS;
    for (int mi = gw; mi < n_mu; mi += nw) {
        const Matchup mu = mus[mi];
        const uint64_t livem = DECK_MASK & ~(mu.hole_a | mu.hole_b | board | dead);
        const int nl = __popcll(livem);
        __syncwarp();
        for (int p = lane; p < 64; p += 32) {
            if ((livem >> p) & 1ull) slive[__popcll(livem & ((1ull << p) - 1ull))] = live_entry(scard[p], p);
        }
        if (lane < LIVE_MAX - nl && nl + lane < LIVE_MAX) slive[nl + lane] = make_uint4(0, 0, 0, 0);
        HandState ha{0, 0, (uint32_t)mu.hole_a, (uint32_t)(mu.hole_a >> 32)}, hb{0, 0, (uint32_t)mu.hole_b, (uint32_t)(mu.hole_b >> 32)};
        {
            uint64_t x = mu.hole_a;
            while (x) { const int p = __ffsll((long long)x) - 1; x &= x - 1; ha.t += scard[p].rk; ha.sn += scard[p].sn; }
            x = mu.hole_b;
            while (x) { const int p = __ffsll((long long)x) - 1; x &= x - 1; hb.t += scard[p].rk; hb.sn += scard[p].sn; }
        }
        __syncwarp();
        const uint32_t total = K == 0 ? 1u : sbin[nl * BIN_K + K];
        const uint32_t per = (total + 31u) / 32u;
        const uint32_t lo = min(total, per * (uint32_t)lane), hi = min(total, lo + per);
        uint32_t ca = 0, cb = 0, ct = 0;
        if (lo < hi) {
            Walker<K> w;
            w.init(lo, nl, sbin, slive, base);
            for (uint32_t i = lo; i < hi; i++) {
                const HandState b = w.hand();
                const uint32_t ka = hb_class(b.t + ha.t, b.sn + ha.sn, b.mlo | ha.mlo, b.mhi | ha.mhi, tbl);
                const uint32_t kb = hb_class(b.t + hb.t, b.sn + hb.sn, b.mlo | hb.mlo, b.mhi | hb.mhi, tbl);
                ca += (kb - ka) >> 31; cb += (ka - kb) >> 31;      // class codes are 16-bit: the sign bit of the difference
                w.next(slive, nl);
            }
            ct = (hi - lo) - ca - cb;
        }
        const unsigned long long wgt = (unsigned long long)mu.weight;
        wa += wgt * ca; wb += wgt * cb; tie += wgt * ct; tot += wgt * (hi - lo);
    }
    wa = warp_sum64(wa); wb = warp_sum64(wb); tie = warp_sum64(tie); tot = warp_sum64(tot);
    if (lane == 0) {
        unsigned long long *o = reinterpret_cast<unsigned long long *>(out);
        if (wa) atomicAdd(&o[0], wa);
        if (tie) atomicAdd(&o[1], tie);
        if (wb) atomicAdd(&o[2], wb);
        if (tot) atomicAdd(&o[3], tot);
    }
}

static size_t range_smem() { return sizeof(HashBlob) + RNG_WARPS * LIVE_MAX * sizeof(uint4) + BIN_N * BIN_K * sizeof(uint32_t); }

template <int K>
static cudaError_t launch_range_k(const Matchup *d_mu, int n_mu, uint64_t board, uint64_t dead, long long *d_out,
                                  const DeviceTables &tb, int grid, cudaStream_t st) {
    auto kern = range_equity_kernel<K>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)range_smem());
    if (e != cudaSuccess) return e;
    kern<<<grid, RNG_THREADS, range_smem(), st>>>(d_mu, n_mu, board, dead, tb.d_blob, tb.d_card, d_out);
    return cudaGetLastError();
}

cudaError_t launch_range_equity(const Matchup *d_matchups, int n_matchups, uint64_t board, uint64_t dead, long long *d_out,
                                const DeviceTables &tb, const DeviceInfo &di, cudaStream_t st) {
    if (n_matchups <= 0) return cudaSuccess;
    const int want = (n_matchups + RNG_WARPS - 1) / RNG_WARPS;
    const int grid = want < di.sm_count ? want : di.sm_count;
    switch (5 - __builtin_popcountll(board)) {
        case 0: return launch_range_k<0>(d_matchups, n_matchups, board, dead, d_out, tb, grid, st);
        case 1: return launch_range_k<1>(d_matchups, n_matchups, board, dead, d_out, tb, grid, st);
        case 2: return launch_range_k<2>(d_matchups, n_matchups, board, dead, d_out, tb, grid, st);
        case 3: return launch_range_k<3>(d_matchups, n_matchups, board, dead, d_out, tb, grid, st);
        case 4: return launch_range_k<4>(d_matchups, n_matchups, board, dead, d_out, tb, grid, st);
        case 5: return launch_range_k<5>(d_matchups, n_matchups, board, dead, d_out, tb, grid, st);
        default: return cudaErrorInvalidValue;
    }
}

// ---------------------------------------------------------------------------------------------------
// K2: all C(52,7) hands in colex order of card ids.  out = hist[9], sum of nibble keys (10 x u64).
// ---------------------------------------------------------------------------------------------------
constexpr int ENUM_FOLD = 32;         // hands between folds of the packed category counters (6-bit fields hold <= 32... 63)
constexpr int ENUM_THREADS = 1024;

__global__ void __launch_bounds__(ENUM_THREADS, 1)
enum7_hash_kernel(uint64_t first, uint64_t count, const HashBlob *__restrict__ g_blob, const CardKey *__restrict__ g_card,
                  const uint32_t *__restrict__ g_nib, int n_classes, const __grid_constant__ CatBase cb,
                  unsigned long long *__restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    HashBlob *H = reinterpret_cast<HashBlob *>(smem_raw);
    uint4 *sdeck = reinterpret_cast<uint4 *>(smem_raw + sizeof(HashBlob));                 // [LIVE_MAX][32 lanes]: one copy per lane
    uint4 *scat = sdeck + LIVE_MAX * 32;                              // 16 entries: {dense base, counter increments lo / hi, 0}
    unsigned long long *sred = reinterpret_cast<unsigned long long *>(scat + 16);   // 16 words
    uint32_t *sbin = reinterpret_cast<uint32_t *>(sred + 16);
    uint32_t *snib = sbin + BIN_N * BIN_K;
    __shared__ unsigned long long mbar;
    BulkStage stage;
    stage.start(&mbar, H, g_blob, (uint32_t)sizeof(HashBlob));
    fill_binom(sbin);
    for (int i = threadIdx.x; i < n_classes; i += blockDim.x) snib[i] = g_nib[i];
    for (int i = threadIdx.x; i < LIVE_MAX * 32; i += blockDim.x) {
        const int id = i >> 5;                                       // card id 13*suit + rank-2 -> bit position
        const int p = id + 2 + 3 * (id / 13);
        sdeck[i] = id < 52 ? live_entry(g_card[p], p) : make_uint4(0, 0, 0, 0);
    }
    const uint4 *my_deck = sdeck + (threadIdx.x & 31);
    if (threadIdx.x < 16) {
        const int c = threadIdx.x;                                   // categories 0..3: 8-bit fields of lo; 4..8: 6-bit fields of hi
        sred[c] = 0ull;
        scat[c] = make_uint4(c < 10 ? cb.v[c] : 0u, c < 4 ? 1u << (8 * c) : 0u, (c >= 4 && c < 9) ? 1u << (6 * (c - 4)) : 0u, 0u);
    }
    stage.wait();
    const uint32_t tbl = staged_table_base(H);

    uint32_t h[9];
#pragma unroll
    for (int i = 0; i < 9; i++) h[i] = 0;
    unsigned long long sumN = 0;
    const uint64_t nthreads = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t per = (count + nthreads - 1) / nthreads;
    const uint64_t me = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t lo = me * per < count ? me * per : count;
    const uint64_t hi = lo + per < count ? lo + per : count;
    if (lo < hi) {
        Walker<7, 32> w;
        w.init((uint32_t)(first + lo), 52, sbin, my_deck, HandState{0, SN_BIAS, 0, 0});
        for (uint64_t base = lo; base < hi; base += ENUM_FOLD) {
            const uint32_t len = (uint32_t)(hi - base < (uint64_t)ENUM_FOLD ? hi - base : (uint64_t)ENUM_FOLD);
            uint32_t acc_lo = 0, acc_hi = 0, nsum = 0;                // nsum <= 32 * 2^24
            for (uint32_t j = 0; j < len; j++) {
                const HandState s = w.hand();
                const uint32_t code = hb_class(s.t, s.sn, s.mlo, s.mhi, tbl);
                const uint32_t nib = snib[code];                      // class codes are dense indices; cat = key >> 20
                const uint4 ci = scat[nib >> 20];
                acc_lo += ci.y; acc_hi += ci.z;
                nsum += nib;
                w.next(my_deck, 52);
            }
#pragma unroll
            for (int i = 0; i < 4; i++) h[i] += (acc_lo >> (8 * i)) & 255u;
#pragma unroll
            for (int i = 4; i < 9; i++) h[i] += (acc_hi >> (6 * (i - 4))) & 63u;
            sumN += nsum;
        }
    }
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int i = 0; i < 9; i++) {
        const uint32_t s = warp_sum(h[i]);
        if (lane == 0 && s) atomicAdd(&sred[i], (unsigned long long)s);
    }
    sumN = warp_sum64(sumN);
    if (lane == 0) atomicAdd(&sred[9], sumN);
    __syncthreads();
    if (threadIdx.x < 10 && sred[threadIdx.x]) atomicAdd(&out[threadIdx.x], sred[threadIdx.x]);
}

// ---------------------------------------------------------------------------------------------------
// K2, warp-cooperative walk (A/B variant, MCP_K2_VARIANT=warp; measured slower than the per-thread walk, see the launcher).  In colex order the hands that share their four highest cards c3 < c4 < c5 < c6
// are one contiguous run of C(c3, 3) indices, and inside the run the index IS the colex rank of the three low cards
// (c0, c1, c2 < c3): C(c2, 3) + pair index.  A warp takes chunks of 1,024 consecutive indices and walks the runs inside; its lanes
// stride the three-card index, so every lane does the same thing every step -- no colex carry diverges a warp (the
// per-thread walk pays the carry path of ANY lane in 98 % of its steps: 61.5 executed instructions per hand against 41
// on the common step).  Per hand: one 8-byte load of the pair's {rank-key sum, suit counters} from a table of all 1,326
// pairs (consecutive lanes read consecutive entries), two adds onto the run's partial hand + card c2 (kept per lane,
// advanced when the three-card index crosses C(c2 + 1, 3)), the two hash loads; the pair's mask words are only loaded by
// the lanes whose hand holds a flush.  Counting is one shared-memory reduction per hand into a histogram of the 4,892 class
// codes; categories and the key checksum come out of it once per block.  Lane utilisation over the whole enumeration:
// 97.2 % (runs end mid-warp).
// ---------------------------------------------------------------------------------------------------
constexpr int N_PAIRS = 1326;
constexpr uint32_t EW_CHUNK = 4096;           // hands per unit of work: chunks are dealt round-robin to the warps, so the slow
                                              // stretches of the enumeration (low indices: runs of a few hands) spread over all of them
constexpr uint32_t EW_HEAD = 1u << 21;        // ... and the slowest stretch, the first 2 Mi indices, goes out in chunks of 256
__global__ void __launch_bounds__(ENUM_THREADS, 1)
enum7_warp_kernel(uint64_t first, uint64_t count, const HashBlob *__restrict__ g_blob, const CardKey *__restrict__ g_card,
                  const uint32_t *__restrict__ g_nib, int n_classes, unsigned long long *__restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    HashBlob *H = reinterpret_cast<HashBlob *>(smem_raw);
    uint4 *sdeck = reinterpret_cast<uint4 *>(smem_raw + sizeof(HashBlob));            // [52] by card id: {rk, sn, mlo, mhi}
    uint2 *pairA = reinterpret_cast<uint2 *>(sdeck + 52);                             // [1326] {rk0 + rk1, sn0 + sn1} by colex pair index
    uint2 *pairB = pairA + N_PAIRS;                                                   // [1326] {mlo, mhi}
    unsigned long long *sred = reinterpret_cast<unsigned long long *>(pairB + N_PAIRS);   // 16 words
    uint32_t *sbin = reinterpret_cast<uint32_t *>(sred + 16);
    uint32_t *shist = sbin + BIN_N * BIN_K;                                           // [n_classes] hands per class code
    __shared__ unsigned long long mbar;
    BulkStage stage;
    stage.start(&mbar, H, g_blob, (uint32_t)sizeof(HashBlob));
    fill_binom(sbin);
    for (int i = threadIdx.x; i < n_classes; i += blockDim.x) shist[i] = 0u;
    for (int p = threadIdx.x; p < N_PAIRS; p += blockDim.x) {
        int c1 = 1;
        while ((c1 + 1) * c1 / 2 <= p) c1++;                                          // largest c1 with C(c1, 2) <= p
        const int c0 = p - c1 * (c1 - 1) / 2;
        const int p0 = c0 + 2 + 3 * (c0 / 13), p1 = c1 + 2 + 3 * (c1 / 13);           // card id -> bit position
        const uint4 a = live_entry(g_card[p0], p0), b = live_entry(g_card[p1], p1);
        pairA[p] = make_uint2(a.x + b.x, a.y + b.y);
        pairB[p] = make_uint2(a.z | b.z, a.w | b.w);
    }
    if (threadIdx.x < 52) { const int id = threadIdx.x, p = id + 2 + 3 * (id / 13); sdeck[id] = live_entry(g_card[p], p); }
    if (threadIdx.x < 16) sred[threadIdx.x] = 0ull;
    stage.wait();
    const uint32_t tbl = staged_table_base(H);
    const uint32_t hist32 = staged_table_base(shist);

    const int lane = threadIdx.x & 31;
    const uint64_t nw = (uint64_t)gridDim.x * (ENUM_THREADS / 32), gw = (uint64_t)blockIdx.x * (ENUM_THREADS / 32) + (threadIdx.x >> 5);
    // work list: [0, head) in chunks of 256, the rest in chunks of EW_CHUNK, both dealt round-robin
    const uint64_t head = first < EW_HEAD ? (EW_HEAD - first < count ? EW_HEAD - first : count) : 0;
    const uint64_t n_head = (head + 255) >> 8, n_all = n_head + (count - head + EW_CHUNK - 1) / EW_CHUNK;
    // (lane-constant) where index `lane` of a run sits: the run's three-card index starts at 0 in all but a chunk's first run
    uint32_t c2_lane = 2;
    while (sbin[(c2_lane + 1) * BIN_K + 3] <= (uint32_t)lane) c2_lane++;
    for (uint64_t ch = gw; ch < n_all; ch += nw) {
        const uint64_t lo = ch < n_head ? ch << 8 : head + (ch - n_head) * EW_CHUNK;
        const uint64_t top = ch < n_head ? head : count;
        const uint64_t len = ch < n_head ? 256 : EW_CHUNK;
        const uint64_t hi = lo + len < top ? lo + len : top;
        // the run the chunk starts in: its four high cards and the three-card index inside it (warp-uniform; four
        // branch-free binary searches: the largest x with C(x, k) <= rest is below the previous card by itself)
        uint32_t idx = (uint32_t)(first + lo);
        int c[7];
#pragma unroll
        for (int k = 7; k >= 4; k--) {
            uint32_t x = (uint32_t)k - 1u;
#pragma unroll
            for (uint32_t step = 32; step >= 1; step >>= 1) { const uint32_t y = min(x + step, 52u); if (sbin[y * BIN_K + k] <= idx) x = y; }
            c[k - 1] = (int)x;
            idx -= sbin[x * BIN_K + k];
        }
        uint32_t i0 = idx, remaining = (uint32_t)(hi - lo);
        for (;;) {
            const uint32_t size3 = sbin[c[3] * BIN_K + 3];
            const uint32_t iend = min(size3, i0 + remaining);
            const uint4 d3 = sdeck[c[3]], d4 = sdeck[c[4]], d5 = sdeck[c[5]], d6 = sdeck[c[6]];
            const HandState P{d3.x + d4.x + d5.x + d6.x, SN_BIAS + d3.y + d4.y + d5.y + d6.y, d3.z | d4.z | d5.z | d6.z, d3.w | d4.w | d5.w | d6.w};
            uint32_t i = i0 + (uint32_t)lane;
            const uint32_t rounds = (iend - i0 + 31u) >> 5;           // warp-uniform: every lane runs the same trip count
            uint32_t c2 = c2_lane;                                   // largest c2 with C(c2, 3) <= i; c2 < c3 <= 48
            if (i0 != 0) {                                           // (uniform) a chunk's first run starts anywhere
                c2 = 2;
#pragma unroll
                for (uint32_t step = 32; step >= 1; step >>= 1) { const uint32_t y = min(c2 + step, 51u); if (sbin[y * BIN_K + 3] <= i) c2 = y; }
            }
            c2 = min(c2, (uint32_t)c[3] - 1u);                       // (a lane past the end of a short run idles on a valid card)
            uint32_t t3c = sbin[c2 * BIN_K + 3], t3n = sbin[(c2 + 1) * BIN_K + 3];
            uint4 e2 = sdeck[c2];
            HandState Q{P.t + e2.x, P.sn + e2.y, P.mlo | e2.z, P.mhi | e2.w};
            for (uint32_t r = 0; r < rounds; r++, i += 32) {
                const bool live = i < iend;
                while (live && i >= t3n) {                           // the three-card index entered the next c2
                    c2++; t3c = t3n; t3n = sbin[(c2 + 1) * BIN_K + 3];
                    e2 = sdeck[c2];
                    Q = HandState{P.t + e2.x, P.sn + e2.y, P.mlo | e2.z, P.mhi | e2.w};
                }
                const uint32_t p = live ? i - t3c : 0u;
                const uint2 A = pairA[p];
                const uint32_t t = Q.t + A.x, sn = Q.sn + A.y, flag = sn & 0x8888u;
                uint32_t addr = hb_slot_addr(t, tbl);
                if (flag) {                                          // one of 0x8, 0x80, 0x800, 0x8000: the flush suit's lane
                    const uint2 B = pairB[p];
                    const uint32_t w = (flag & 0x8800u) ? (Q.mhi | B.y) : (Q.mlo | B.x);
                    addr = hb_flush_addr(((flag & 0x8080u) ? (w >> 16) : (w & 0xFFFFu)) >> 2, tbl);
                }
                const uint32_t code = lds_u16(addr);
                // one shared-memory reduction per hand into the histogram of the 4,892 class codes; categories and the
                // key checksum are derived from it once per block
                if (live) asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(hist32 + 4u * code) : "memory");
            }
            remaining -= iend - i0;
            if (remaining == 0) break;
            i0 = 0;                                                  // next run: colex successor of (c3, c4, c5, c6)
            c[3]++;
            if (c[3] == c[4]) { c[3] = 3; c[4]++; if (c[4] == c[5]) { c[4] = 4; c[5]++; if (c[5] == c[6]) { c[5] = 5; c[6]++; } } }
        }
    }
    __syncthreads();
    // class histogram -> categories (key >> 20) and key checksum
    for (int cl = threadIdx.x; cl < n_classes; cl += blockDim.x) {
        const uint32_t n = shist[cl];
        if (n) {
            const uint32_t nib = g_nib[cl];
            atomicAdd(&sred[nib >> 20], (unsigned long long)n);
            atomicAdd(&sred[9], (unsigned long long)n * nib);
        }
    }
    __syncthreads();
    if (threadIdx.x < 10 && sred[threadIdx.x]) atomicAdd(&out[threadIdx.x], sred[threadIdx.x]);
}
static size_t enum_warp_smem(int n_classes) {
    return sizeof(HashBlob) + 52 * sizeof(uint4) + 2 * N_PAIRS * sizeof(uint2) + 16 * sizeof(unsigned long long) +
           (BIN_N * BIN_K + (size_t)n_classes) * sizeof(uint32_t);
}

// ---------------------------------------------------------------------------------------------------
// K2, pair walk (default).  The per-thread colex walk with its two lowest cards stepped as ONE index: inside
// a fixed (c2 .. c6) the hands are the C(c2, 2) pairs c0 < c1 < c2 in colex order, i.e. pair index 0, 1, 2, ...  Per hand
// the thread adds the pair's {rank-key sum, suit counters} (one 8-byte load from the table of all 1,326 pairs) onto the
// partial hand of its five high cards; only when the pair index reaches C(c2, 2) does the colex walk of the high cards
// (Walker::bump<2>) run.  The frequent carry of the plain walk (c0 reaching c1, every ~c1 hands per lane, so in 98 % of a
// warp's steps) is gone: a lane carries every C(c2, 2) hands (~190 on average).  ncu (profiles/r02_k2_pair_ncu.txt): 44
// executed instructions per hand against 60.5; the first build was then bound by the shared-memory pipe (19.3 wavefronts
// per 32 hands: pair entry, D, V, key, and a 16-byte category entry), so categories are now counted arithmetically in one
// 64-bit word of 7-bit fields (15 wavefronts), and thread ranges are dealt round-robin over the blocks (block 0 used to own
// the slow low indices): 0.404 -> 0.382 -> 0.319 ms for all C(52,7) hands.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(ENUM_THREADS, 1)
enum7_pair_kernel(uint64_t first, uint64_t count, const HashBlob *__restrict__ g_blob, const CardKey *__restrict__ g_card,
                  const uint32_t *__restrict__ g_nib, int n_classes, unsigned long long *__restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    HashBlob *H = reinterpret_cast<HashBlob *>(smem_raw);
    uint4 *sdeck = reinterpret_cast<uint4 *>(smem_raw + sizeof(HashBlob));            // [LIVE_MAX] by card id; ids >= 52 are zero entries
    uint2 *pairA = reinterpret_cast<uint2 *>(sdeck + LIVE_MAX);                              // [1326] {rk0 + rk1, sn0 + sn1} by colex pair index
    uint2 *pairB = pairA + N_PAIRS;                                                   // [1326] {mlo, mhi}
    unsigned long long *sred = reinterpret_cast<unsigned long long *>(pairB + N_PAIRS);   // 16 words
    uint32_t *sbin = reinterpret_cast<uint32_t *>(sred + 16);
    uint32_t *snib = sbin + BIN_N * BIN_K;
    __shared__ unsigned long long mbar;
    BulkStage stage;
    stage.start(&mbar, H, g_blob, (uint32_t)sizeof(HashBlob));
    fill_binom(sbin);
    for (int i = threadIdx.x; i < n_classes; i += blockDim.x) snib[i] = g_nib[i];
    for (int p = threadIdx.x; p < N_PAIRS; p += blockDim.x) {
        int c1 = 1;
        while ((c1 + 1) * c1 / 2 <= p) c1++;                                          // largest c1 with C(c1, 2) <= p
        const int c0 = p - c1 * (c1 - 1) / 2;
        const int p0 = c0 + 2 + 3 * (c0 / 13), p1 = c1 + 2 + 3 * (c1 / 13);           // card id -> bit position
        const uint4 a = live_entry(g_card[p0], p0), b = live_entry(g_card[p1], p1);
        pairA[p] = make_uint2(a.x + b.x, a.y + b.y);
        pairB[p] = make_uint2(a.z | b.z, a.w | b.w);
    }
    if (threadIdx.x < LIVE_MAX) {
        const int id = threadIdx.x, p = id + 2 + 3 * (id / 13);
        sdeck[id] = id < 52 ? live_entry(g_card[p], p) : make_uint4(0, 0, 0, 0);
    }
    if (threadIdx.x < 16) sred[threadIdx.x] = 0ull;
    stage.wait();
    const uint32_t tbl = staged_table_base(H);

    uint32_t h[9];
#pragma unroll
    for (int i = 0; i < 9; i++) h[i] = 0;
    unsigned long long sumN = 0;
    const uint64_t nthreads = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t per = (count + nthreads - 1) / nthreads;
    // ranges are dealt round-robin over the blocks: the slow stretch of the enumeration (low indices: small c2, frequent
    // carries) is spread over all SMs instead of landing on block 0
    const uint64_t me = (uint64_t)threadIdx.x * gridDim.x + blockIdx.x;
    const uint64_t lo = me * per < count ? me * per : count;
    const uint64_t hi = lo + per < count ? lo + per : count;
    if (lo < hi) {
        Walker<7, 1> w;                                              // P[2] = the partial hand of c2 .. c6
        w.init((uint32_t)(first + lo), 52, sbin, sdeck, HandState{0, SN_BIAS, 0, 0});
        uint32_t p = (uint32_t)(w.c[0] + w.c[1] * (w.c[1] - 1) / 2);
        uint32_t pend = (uint32_t)(w.c[2] * (w.c[2] - 1) / 2);
        HandState Q = w.P[2];
        for (uint64_t base = lo; base < hi; base += ENUM_FOLD) {
            const uint32_t len = (uint32_t)(hi - base < (uint64_t)ENUM_FOLD ? hi - base : (uint64_t)ENUM_FOLD);
            unsigned long long acc = 0;                               // nine 7-bit category counters (<= 32 per fold)
            uint32_t nsum = 0;                                        // <= 32 * 2^24
            for (uint32_t j = 0; j < len; j++) {
                const uint2 A = pairA[p];
                const uint32_t t = Q.t + A.x, sn = Q.sn + A.y, flag = sn & 0x8888u;
                uint32_t addr = hb_slot_addr(t, tbl);
                if (flag) {                                          // one of 0x8, 0x80, 0x800, 0x8000: the flush suit's lane
                    const uint2 B = pairB[p];
                    const uint32_t wd = (flag & 0x8800u) ? (Q.mhi | B.y) : (Q.mlo | B.x);
                    addr = hb_flush_addr(((flag & 0x8080u) ? (wd >> 16) : (wd & 0xFFFFu)) >> 2, tbl);
                }
                const uint32_t nib = snib[lds_u16(addr)];             // class codes are dense indices; cat = key >> 20
                acc += 1ull << (7u * (nib >> 20));                    // (arithmetic: the shared-memory pipe is the busy one)
                nsum += nib;
                if (++p == pend) {                                   // every pair below c2 done: colex successor of (c2 .. c6)
                    w.template bump<2>(sdeck, 52);
                    Q = w.P[2];
                    p = 0;
                    pend = (uint32_t)(w.c[2] * (w.c[2] - 1) / 2);
                }
            }
#pragma unroll
            for (int i = 0; i < 9; i++) h[i] += (uint32_t)(acc >> (7 * i)) & 127u;
            sumN += nsum;
        }
    }
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int i = 0; i < 9; i++) {
        const uint32_t s = warp_sum(h[i]);
        if (lane == 0 && s) atomicAdd(&sred[i], (unsigned long long)s);
    }
    sumN = warp_sum64(sumN);
    if (lane == 0) atomicAdd(&sred[9], sumN);
    __syncthreads();
    if (threadIdx.x < 10 && sred[threadIdx.x]) atomicAdd(&out[threadIdx.x], sred[threadIdx.x]);
}
static size_t enum_pair_smem(int n_classes) {
    return sizeof(HashBlob) + LIVE_MAX * sizeof(uint4) + 2 * N_PAIRS * sizeof(uint2) + 16 * sizeof(unsigned long long) +
           (BIN_N * BIN_K + (size_t)n_classes) * sizeof(uint32_t);
}

// ---------------------------------------------------------------------------------------------------
// K2, lockstep walk (default for large ranges).  The pair walk's biggest shared-memory cost is the pair entry itself: 32
// lanes at 32 unrelated pair indices cost ~8 wavefronts per LDS.64.  Here all lanes of a warp share c2 and step the pair
// index p = 0 .. C(c2, 2) - 1 TOGETHER, each lane on its own four high cards (c3 < c4 < c5 < c6 above c2): the pair entry
// is one broadcast load.  A unit of work is (c2, 32 consecutive four-card tops in colex order over the cards above c2, a
// chunk of at most 64 pair indices): a warp's whole share is ~880 hand-steps, so whole pair runs (up to 1,081 steps) would
// leave the grid badly balanced; units are dealt round-robin over all warps of the grid.  The enumeration ORDER is therefore not the colex order -- the outputs are sums, any order gives the
// same histogram and checksum -- but index ranges are still honoured: a hand's colex index is base(c2 .. c6) + p, the
// kernel only visits tops whose c6 the range can contain and counts a hand only if its index lies in [first, first + count).
// ---------------------------------------------------------------------------------------------------
constexpr uint32_t EL_PCHUNK = 64;
template <bool FULL>
__global__ void __launch_bounds__(ENUM_THREADS, 1)
enum7_lock_kernel(uint32_t first, uint32_t count, int c6lo, int c6hi, const HashBlob *__restrict__ g_blob,
                  const CardKey *__restrict__ g_card, const uint32_t *__restrict__ g_nib, int n_classes,
                  unsigned long long *__restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    HashBlob *H = reinterpret_cast<HashBlob *>(smem_raw);
    uint4 *sdeck = reinterpret_cast<uint4 *>(smem_raw + sizeof(HashBlob));            // [LIVE_MAX] by card id
    uint2 *pairA = reinterpret_cast<uint2 *>(sdeck + LIVE_MAX);                       // [1326] {rk0 + rk1, sn0 + sn1} by colex pair index
    uint2 *pairB = pairA + N_PAIRS;                                                   // [1326] {mlo, mhi}
    unsigned long long *sred = reinterpret_cast<unsigned long long *>(pairB + N_PAIRS);   // 16 words
    uint32_t *sbin = reinterpret_cast<uint32_t *>(sred + 16);
    uint32_t *snib = sbin + BIN_N * BIN_K;
    __shared__ uint32_t sfirst[64], sqlo[64], sqhi[64];                               // by c2: first unit, first / last + 1 top index
    __shared__ unsigned long long mbar;
    BulkStage stage;
    stage.start(&mbar, H, g_blob, (uint32_t)sizeof(HashBlob));
    fill_binom(sbin);
    for (int i = threadIdx.x; i < n_classes; i += blockDim.x) snib[i] = g_nib[i];
    for (int p = threadIdx.x; p < N_PAIRS; p += blockDim.x) {
        int c1 = 1;
        while ((c1 + 1) * c1 / 2 <= p) c1++;                                          // largest c1 with C(c1, 2) <= p
        const int c0 = p - c1 * (c1 - 1) / 2;
        const int p0 = c0 + 2 + 3 * (c0 / 13), p1 = c1 + 2 + 3 * (c1 / 13);           // card id -> bit position
        const uint4 a = live_entry(g_card[p0], p0), b = live_entry(g_card[p1], p1);
        pairA[p] = make_uint2(a.x + b.x, a.y + b.y);
        pairB[p] = make_uint2(a.z | b.z, a.w | b.w);
    }
    if (threadIdx.x < LIVE_MAX) {
        const int id = threadIdx.x, p = id + 2 + 3 * (id / 13);
        sdeck[id] = id < 52 ? live_entry(g_card[p], p) : make_uint4(0, 0, 0, 0);
    }
    if (threadIdx.x < 16) sred[threadIdx.x] = 0ull;
    stage.wait();
    const uint32_t tbl = staged_table_base(H);
    // tops per c2: d3 < d4 < d5 < d6 over the 51 - c2 cards above c2 (c_i = c2 + 1 + d_i), colex index q; a range can only
    // hold hands whose c6 lies in [c6lo, c6hi]: q in [C(c6lo - c2 - 1, 4), C(c6hi - c2, 4))
    if (threadIdx.x < 64) {
        const int c2 = threadIdx.x;
        uint32_t lo = 0, hi = 0;
        if (c2 >= 2 && c2 <= 47) {
            const int dl = c6lo - c2 - 1, dh = c6hi - c2;
            lo = dl > 0 ? sbin[dl * BIN_K + 4] : 0u;
            hi = dh > 0 ? sbin[(dh > 51 - c2 ? 51 - c2 : dh) * BIN_K + 4] : 0u;
            if (hi < lo) hi = lo;
        }
        sqlo[c2] = lo; sqhi[c2] = hi;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t run = 0;
        for (int c2 = 0; c2 < 64; c2++) {
            sfirst[c2] = (c2 >= 2 && c2 <= 47) ? run : (c2 < 2 ? 0u : 0xFFFFFFFFu);
            if (c2 >= 2 && c2 <= 47) run += ((sqhi[c2] - sqlo[c2] + 31u) >> 5) * (((uint32_t)(c2 * (c2 - 1) / 2) + EL_PCHUNK - 1u) / EL_PCHUNK);
        }
        sfirst[48] = run;                                                             // total number of units; 49.. stay 2^32 - 1
    }
    __syncthreads();
    const uint32_t n_units = sfirst[48];

    uint32_t h[9];
#pragma unroll
    for (int i = 0; i < 9; i++) h[i] = 0;
    unsigned long long sumN = 0;
    const int lane = threadIdx.x & 31;
    const uint32_t nw = gridDim.x * (ENUM_THREADS / 32), gw = blockIdx.x * (ENUM_THREADS / 32) + (threadIdx.x >> 5);
    for (uint32_t unit = gw; unit < n_units; unit += nw) {
        // c2 of the unit: the largest c2 <= 47 with sfirst[c2] <= unit (units of empty c2 have zero width and are skipped by
        // taking the largest)
        uint32_t c2 = 2;
#pragma unroll
        for (uint32_t step = 32; step >= 1; step >>= 1) { const uint32_t y = c2 + step; if (y <= 47u && sfirst[y] <= unit) c2 = y; }
        const uint32_t qend = sqhi[c2];
        const uint32_t npairs = c2 * (c2 - 1u) / 2u, nchunks = (npairs + EL_PCHUNK - 1u) / EL_PCHUNK;
        const uint32_t u = unit - sfirst[c2], qb = u / nchunks, p0 = (u - qb * nchunks) * EL_PCHUNK, p1 = min(npairs, p0 + EL_PCHUNK);
        uint32_t q = sqlo[c2] + (qb << 5) + (uint32_t)lane;
        const bool valid = q < qend;
        q = valid ? q : qend - 1u;                                                    // (an idle lane walks a valid top, uncounted)
        int d[4];
        colex_unrank<4>(q, 0, sbin, d);
        const int c3 = d[0] + (int)c2 + 1, c4 = d[1] + (int)c2 + 1, c5 = d[2] + (int)c2 + 1, c6 = d[3] + (int)c2 + 1;
        const uint4 e2 = sdeck[c2], e3 = sdeck[c3], e4 = sdeck[c4], e5 = sdeck[c5], e6 = sdeck[c6];
        const HandState Q{e2.x + e3.x + e4.x + e5.x + e6.x, SN_BIAS + e2.y + e3.y + e4.y + e5.y + e6.y,
                          e2.z | e3.z | e4.z | e5.z | e6.z, e2.w | e3.w | e4.w | e5.w | e6.w};
        // colex index of the hand with pair index p: base + p
        const uint32_t base = sbin[c6 * BIN_K + 7] + sbin[c5 * BIN_K + 6] + sbin[c4 * BIN_K + 5] + sbin[c3 * BIN_K + 4] + sbin[c2 * BIN_K + 3];
        const uint32_t rel = base - first;                                            // (wraps below the range: then rel + p >= count)
        if (!FULL) {
            // does [base, base + npairs) meet [first, first + count) for any lane?  (all below 2^28: no wrap)
            const bool overlap = valid && base + p0 < first + count && base + p1 > first;
            if (!__any_sync(0xFFFFFFFFu, overlap)) continue;
        }
        unsigned long long acc = 0;                                   // nine 7-bit category counters (<= 32 per fold)
        uint32_t nsum = 0;                                            // <= 32 * 2^24
        for (uint32_t p = p0; p < p1; p++) {
            const uint2 A = pairA[p];                                 // the same entry for every lane: one broadcast
            const uint32_t t = Q.t + A.x, sn = Q.sn + A.y, flag = sn & 0x8888u;
            uint32_t addr = hb_slot_addr(t, tbl);
            if (flag) {                                              // one of 0x8, 0x80, 0x800, 0x8000: the flush suit's lane
                const uint2 B = pairB[p];
                const uint32_t wd = (flag & 0x8800u) ? (Q.mhi | B.y) : (Q.mlo | B.x);
                addr = hb_flush_addr(((flag & 0x8080u) ? (wd >> 16) : (wd & 0xFFFFu)) >> 2, tbl);
            }
            const uint32_t nib = snib[lds_u16(addr)];                 // class codes are dense indices; cat = key >> 20
            const bool ok = valid && (FULL || rel + p < count);
            acc += (unsigned long long)ok << (7u * (nib >> 20));
            nsum += ok ? nib : 0u;
            if ((p & 31u) == 31u) {
#pragma unroll
                for (int i = 0; i < 9; i++) h[i] += (uint32_t)(acc >> (7 * i)) & 127u;
                sumN += nsum;
                acc = 0; nsum = 0;
            }
        }
#pragma unroll
        for (int i = 0; i < 9; i++) h[i] += (uint32_t)(acc >> (7 * i)) & 127u;
        sumN += nsum;
    }
#pragma unroll
    for (int i = 0; i < 9; i++) {
        const uint32_t s = warp_sum(h[i]);
        if (lane == 0 && s) atomicAdd(&sred[i], (unsigned long long)s);
    }
    sumN = warp_sum64(sumN);
    if (lane == 0) atomicAdd(&sred[9], sumN);
    __syncthreads();
    if (threadIdx.x < 10 && sred[threadIdx.x]) atomicAdd(&out[threadIdx.x], sred[threadIdx.x]);
}

static size_t enum_smem(int n_classes) {
    return sizeof(HashBlob) + (LIVE_MAX * 32 + 16) * sizeof(uint4) + 16 * sizeof(unsigned long long) +
           (BIN_N * BIN_K + (size_t)n_classes) * sizeof(uint32_t);
}

cudaError_t launch_enum7_hash(uint64_t first, uint64_t count, unsigned long long *d_out,
                              const DeviceTables &tb, const DeviceInfo &di, cudaStream_t st) {
    if (count == 0) return cudaSuccess;
    // Four walks with identical results (all C(52,7) hands on a B200; tools/k2_time.py):
    //   lockstep (>= 2^22 hands) 0.235 ms  35 executed instructions and 12.9 shared-memory wavefronts per hand (profiles/r02_k2_lock_ncu.txt)
    //   pair (smaller ranges)   0.317 ms  44 executed instructions per hand, bound by the shared-memory pipe (15 wavefronts per 32 hands)
    //   MCP_K2_VARIANT=thread   0.404 ms  the plain per-thread colex walk: 60.5 instructions per hand (the c0 / c1 carry diverges 98 % of the steps)
    //   MCP_K2_VARIANT=warp     0.479 ms  warp-cooperative: 62-68 instructions per hand at IPC 2.2 (profiles/r02_k2_warp_ncu.txt)
    //   lockstep                see enum7_lock_kernel: the default for ranges of at least 2^22 hands (MCP_K2_VARIANT=lock forces it)
    static const char variant = [] { const char *e = getenv("MCP_K2_VARIANT"); return e ? e[0] : 'a'; }();      // 'a' = automatic
    const bool per_thread = variant != 'w';
    if (variant == 'l' || (variant == 'a' && count >= ((uint64_t)1 << 22))) {
        const size_t lsmem = enum_pair_smem((int)tb.n_classes);
        const uint64_t total = cbinom(52, 7);
        const bool full = first == 0 && count == total;
        auto c6_of = [](uint64_t idx) { int x = 6; while (x < 51 && cbinom(x + 1, 7) <= idx) x++; return x; };
        const int c6lo = full ? 0 : c6_of(first), c6hi = full ? 51 : c6_of(first + count - 1);
        auto kern = full ? enum7_lock_kernel<true> : enum7_lock_kernel<false>;
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lsmem);
        if (e != cudaSuccess) return e;
        kern<<<di.sm_count, ENUM_THREADS, lsmem, st>>>((uint32_t)first, (uint32_t)count, c6lo, c6hi, tb.d_blob, tb.d_card, tb.d_nib, (int)tb.n_classes, d_out);
        return cudaGetLastError();
    }
    if (variant == 'p' || variant == 'a') {
        const size_t psmem = enum_pair_smem((int)tb.n_classes);
        cudaError_t e = cudaFuncSetAttribute(enum7_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)psmem);
        if (e != cudaSuccess) return e;
        uint64_t want = (count + ENUM_THREADS - 1) / ENUM_THREADS;
        const int grid = (int)(want > (uint64_t)di.sm_count ? (uint64_t)di.sm_count : want);
        enum7_pair_kernel<<<grid, ENUM_THREADS, psmem, st>>>(first, count, tb.d_blob, tb.d_card, tb.d_nib, (int)tb.n_classes, d_out);
        return cudaGetLastError();
    }
    if (!per_thread) {
        const size_t wsmem = enum_warp_smem((int)tb.n_classes);
        cudaError_t e = cudaFuncSetAttribute(enum7_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsmem);
        if (e != cudaSuccess) return e;
        uint64_t want = (count + EW_CHUNK - 1) / EW_CHUNK;            // one chunk per warp before another block is worth it
        want = (want + ENUM_THREADS / 32 - 1) / (ENUM_THREADS / 32);
        const int grid = (int)(want > (uint64_t)di.sm_count ? (uint64_t)di.sm_count : (want < 1 ? 1 : want));
        enum7_warp_kernel<<<grid, ENUM_THREADS, wsmem, st>>>(first, count, tb.d_blob, tb.d_card, tb.d_nib, (int)tb.n_classes, d_out);
        return cudaGetLastError();
    }
    const size_t smem = enum_smem((int)tb.n_classes);
    cudaError_t e = cudaFuncSetAttribute(enum7_hash_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    uint64_t want = (count + ENUM_THREADS - 1) / ENUM_THREADS;
    const int grid = (int)(want > (uint64_t)di.sm_count ? (uint64_t)di.sm_count : want);
    enum7_hash_kernel<<<grid, ENUM_THREADS, smem, st>>>(first, count, tb.d_blob, tb.d_card, tb.d_nib, (int)tb.n_classes, tb.cat_base, d_out);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------
// K4: Monte Carlo equity.  One thread per trial (grid-stride).  Deal spec v4: the first cards of a
// Fisher-Yates shuffle of the ascending list of live cards.  One 32-bit Philox word serves three consecutive
// steps by a multiply chain (x = r * m: the high word is the draw, the low word is the next r), which equals
// Lemire's exact bounded sampler on the product of the three ranges; a trial in which a word ends in the
// sampler's rejection zone (p ~ 2e-4 per trial) is dealt again from the next Philox attempt.  The thread's deck is a byte array in a
// bank-private shared-memory column (13 words, stride S = 1024 threads for the 768-thread block), so every deck
// access of a warp is conflict-free whatever the indices, and the byte offset of position j is
// (j * (S + 1) & (15 * 4S | 3)) | 4 * tid.  768 threads x 80 registers per SM: the 52 KB of decks, the 8 KB card
// table and the 164 KB hash blob fill the 227 KB of shared memory (measured 512 / 640 / 768 threads: 40.7 / 42.8 /
// 43.9 G trials/s, tools/k4_ab.py); tables of up to three random seats need <= 64 registers and run 1024 threads.
// Seats are processed in kernel order: random seats first (cards q and q + NR of the deal, the order of
// deal-hand gameplay.clj:63-75), then the fixed seats; seat_of[] maps back to the caller's seat numbers.
// A card's table entry is {premultiplied rank key, suit-counter increment | rank bit << 16}: the low half of the
// second word counts suits, the high half rebuilds the flush lane without touching the card mask.
// ---------------------------------------------------------------------------------------------------
constexpr int MC_REP = 16;                    // lanes the card table is replicated over: an LDS.64 is served per half-warp
constexpr uint32_t TAG_DEAL4 = 0x4D434434u;
constexpr int MC_MAX_STEPS = 2 * MAX_SEATS + 5;

struct McFyParams {
    uint32_t h_t[MAX_SEATS], h_sn[MAX_SEATS], h_mlo[MAX_SEATS], h_mhi[MAX_SEATS];   // fixed seats, kernel order
    HandState base;                           // known community cards, sn biased
    uint64_t first, count;
    int n_fixed, n_board, n_live;
    int seat_of[MAX_SEATS];                   // kernel seat -> caller seat
    uint32_t deck_init[13];                   // live cards (bit positions, ascending), four per word
    uint32_t m[MC_MAX_STEPS];                 // n_live - i of deal step i
    uint32_t thr[(MC_MAX_STEPS + 2) / 3];     // 2^32 mod (product of the m[i] of the steps word w serves)
    uint32_t rk0[10], rk1[10];                // Philox round keys: seed_lo + r * W0, seed_hi + r * W1
    PotParams pots;                           // EV variant only (kernel seat order)
};

template <int NR> struct McShape {
    static constexpr int STEPS = 2 * NR + 5;                       // seat cards, then up to 5 community cards
    static constexpr int WORDS = (STEPS + 2) / 3;                  // one 32-bit Philox word per three steps
    static constexpr int BLOCKS = (WORDS + 3) / 4;
};

__device__ __forceinline__ void philox_block(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, const McFyParams &prm,
                                             uint32_t *o) {
#pragma unroll
    for (int r = 0; r < 10; r++) {
        const uint64_t p0 = (uint64_t)PHILOX_M0 * c0, p1 = (uint64_t)PHILOX_M1 * c2;
        c0 = (uint32_t)(p1 >> 32) ^ c1 ^ prm.rk0[r]; c1 = (uint32_t)p1;
        c2 = (uint32_t)(p0 >> 32) ^ c3 ^ prm.rk1[r]; c3 = (uint32_t)p0;
    }
    o[0] = c0; o[1] = c1; o[2] = c2; o[3] = c3;
}

// NFX = fixed seats compiled in: MAX_SEATS - NR in general, 1 for the tables with at most one fixed seat (their
// showdown then has NR + 1 slots instead of nine)
template <int NR, int MC_THREADS, int NFX, bool EV>
__global__ void __launch_bounds__(MC_THREADS, 1)
equity_mc_fy_kernel(const __grid_constant__ McFyParams prm, const HashBlob *__restrict__ g_blob,
                    const CardKey *__restrict__ g_card, long long *__restrict__ d_counts) {
    using SH = McShape<NR>;
    constexpr int NSLOT = NR + NFX;                                                          // seats compiled in
    constexpr int MC_STRIDE = MC_THREADS <= 512 ? 512 : 1024;                                // deck column stride in words (power of two)
    constexpr int MC_T4 = 4 * MC_STRIDE;
    constexpr uint32_t J_MUL = MC_STRIDE + 1, J_MASK = (15u * MC_T4) | 3u;                   // j * J_MUL & J_MASK = (j >> 2) * MC_T4 + (j & 3)
    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned char *sdeck = smem_raw;                                                         // [13 words][stride]
    unsigned char *srep_b = smem_raw + 13 * MC_T4;                                           // [64 cards][16 lanes] x 8 B
    uint2 *srep = reinterpret_cast<uint2 *>(srep_b);
    HashBlob *H = reinterpret_cast<HashBlob *>(srep_b + 64 * MC_REP * sizeof(uint2));
    unsigned long long *sred = reinterpret_cast<unsigned long long *>(H + 1);                // 19 words, the mbarrier, 9 chip sums
    BulkStage stage;
    stage.start(sred + 19, H, g_blob, (uint32_t)sizeof(HashBlob));
    for (int i = threadIdx.x; i < 64 * MC_REP; i += blockDim.x) {
        const int p = i / MC_REP, r = p & 15;
        const CardKey k = g_card[p];
        srep[i] = make_uint2(k.rk, k.sn | ((r >= 2 && r <= 14) ? 1u << (16 + r - 2) : 0u));
    }
    if (threadIdx.x < 19) sred[threadIdx.x] = 0ull;
    if (EV && threadIdx.x < MAX_SEATS) sred[20 + threadIdx.x] = 0ull;
    stage.wait();
    const uint32_t tbl = staged_table_base(H);

    const uint32_t tid4 = 4u * threadIdx.x;
    unsigned char *dk = sdeck + tid4;
    // table entry of card a for this lane: shared address a * 128 + (srep + (lane & 15) * 8), one IMAD (the FMA pipe is the
    // idle one); the addend goes through staged_table_base so that the compiler keeps it whole in a register
    const uint32_t rep_lane = staged_table_base(srep_b + 8u * (threadIdx.x & (MC_REP - 1)));
    auto rep_of = [&](uint32_t a) -> uint2 { return lds_u32x2(a * (8u * MC_REP) + rep_lane); };
    const int nb = prm.n_board, nf = prm.n_fixed;
    uint32_t dinit[13];
#pragma unroll
    for (int k = 0; k < 13; k++) dinit[k] = prm.deck_init[k];
    uint32_t acc[NSLOT > 0 ? NSLOT : 1];                  // per kernel seat: wins << 16 | times at the maximum
#pragma unroll
    for (int s = 0; s < NSLOT; s++) acc[s] = 0;
    uint32_t units = 0;
    uint32_t chips[EV && NSLOT > 0 ? NSLOT : 1], since = 0;
#pragma unroll
    for (int s = 0; s < (EV ? NSLOT : 1); s++) chips[s] = 0;

    const uint64_t nthreads = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t tr = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; tr < prm.count; tr += nthreads) {
        const uint64_t trial = prm.first + tr;
        uint32_t card[SH::STEPS];
        for (uint32_t attempt = 0;; attempt++) {
            uint32_t w[4 * SH::BLOCKS];
#pragma unroll
            for (int b = 0; b < SH::BLOCKS; b++)
                philox_block((uint32_t)trial, (uint32_t)(trial >> 32), (uint32_t)b | (attempt << 8), TAG_DEAL4, prm, w + 4 * b);
#pragma unroll
            for (int k = 0; k < 13; k++) *reinterpret_cast<uint32_t *>(dk + k * MC_T4) = dinit[k];
            bool rej = false;
            uint32_t r = 0;
#pragma unroll
            for (int i = 0; i < SH::STEPS; i++) {
                if (i % 3 == 0) r = w[i / 3];
                if (i < 2 * NR || i - 2 * NR < nb) {                     // (else: community card already on the table)
                    uint32_t hi;                                         // r * m: hi = offset in [0, m), lo = the next r
                    asm("mul.hi.u32 %0, %1, %2;" : "=r"(hi) : "r"(r), "r"(prm.m[i]));
                    r *= prm.m[i];
                    const uint32_t jm = hi * J_MUL + (uint32_t)i * J_MUL;
                    unsigned char *pj = sdeck + ((jm & J_MASK) | tid4);       // word (j >> 2) of the column, byte j & 3
                    const uint32_t a = *pj;
                    // deck[j] = deck[i]: nothing has been swapped before step 0, nobody reads after the last step
                    if (i == 0) *pj = (unsigned char)dinit[0];
                    else if (i < SH::STEPS - 1) *pj = dk[(i >> 2) * MC_T4 + (i & 3)];
                    card[i] = a;
                }
                if (i % 3 == 2 || i == SH::STEPS - 1) rej |= r < prm.thr[i / 3];   // Lemire's zone of the word's product
            }
            if (!rej || attempt >= 63u) break;                          // (the cap only bounds the loop)
        }

        // community cards: hand state of the board, and the flush lane of its candidate suit
        uint2 eb[5];
        HandState b = prm.base;
#pragma unroll
        for (int k = 0; k < 5; k++) {
            eb[k] = make_uint2(0u, 0u);
            if (k < nb) { eb[k] = rep_of(card[2 * NR + k]); b.t += eb[k].x; b.sn += eb[k].y; }      // (high half of the sum is never read)
        }
        // a seat can only flush in a suit the board holds >= 3 of: cbit = that suit's counter bit (0 if none)
        const uint32_t cbit = ((b.sn + 0x2222u) & 0x8888u) >> 3;
        const uint32_t blane_w = (cbit & 0x1100u) ? b.mhi : b.mlo;      // known community cards
        uint32_t lb = cbit ? (((cbit & 0x1010u) ? (blane_w >> 16) : (blane_w & 0xFFFFu)) >> 2) : 0u;
#pragma unroll
        for (int k = 0; k < 5; k++) if (k < nb) lb |= (eb[k].y & cbit) ? (eb[k].y >> 16) : 0u;

        uint32_t key[NSLOT > 0 ? NSLOT : 1];
#pragma unroll
        for (int q = 0; q < NR; q++) {
            const uint2 e1 = rep_of(card[q]), e2 = rep_of(card[q + NR]);
            const uint32_t t = b.t + e1.x + e2.x, sn = b.sn + e1.y + e2.y;
            // (a divergent branch per seat measures faster here than the branch-free select of hb_class: 38.3 vs 37.5 G trials/s)
            uint32_t k = lds_u16(hb_slot_addr(t, tbl));
            if (sn & 0x8888u) {
                const uint32_t fm = lb | ((e1.y & cbit) ? (e1.y >> 16) : 0u) | ((e2.y & cbit) ? (e2.y >> 16) : 0u);
                k = lds_u16(hb_flush_addr(fm, tbl));
            }
            key[q] = k;
        }
#pragma unroll
        for (int s = 0; s < NFX; s++) {
            key[NR + s] = 0;
            if (s < nf) {
                const uint32_t t = b.t + prm.h_t[s], sn = b.sn + prm.h_sn[s];
                uint32_t k = lds_u16(hb_slot_addr(t, tbl));
                if (sn & 0x8888u) {
                    const uint32_t hw = (cbit & 0x1100u) ? prm.h_mhi[s] : prm.h_mlo[s];
                    k = lds_u16(hb_flush_addr(lb | (((cbit & 0x1010u) ? (hw >> 16) : (hw & 0xFFFFu)) >> 2), tbl));
                }
                key[NR + s] = k;
            }
        }
        // showdown (gameplay.clj:127-132): sole maximum wins, shared maximum ties.  Class codes are 16-bit, so
        // key + (0x10000 - best) carries into bit 16 exactly for the seats at the maximum: no compares, no predicates
        uint32_t best = 0;
#pragma unroll
        for (int s = 0; s < NSLOT; s++) if (s < NR || s - NR < nf) best = max(best, key[s]);
        const uint32_t cb = 0x10000u - best;
        uint32_t at[NSLOT > 0 ? NSLOT : 1], nbest = 0;
#pragma unroll
        for (int s = 0; s < NSLOT; s++) {
            at[s] = 0;
            if (s < NR || s - NR < nf) { at[s] = (key[s] + cb) >> 16; nbest += at[s]; }
        }
        const uint32_t inc = nbest == 1 ? 0x10001u : 1u;
#pragma unroll
        for (int s = 0; s < NSLOT; s++) if (s < NR || s - NR < nf) acc[s] += at[s] * inc;
        units++;
        if constexpr (EV && NSLOT > 0) {
            if (prm.pots.has_main) {                     // (uniform) the main pot: its winners are the seats at the maximum
                const uint32_t share = prm.pots.main_share[nbest];
#pragma unroll
                for (int s = 0; s < NSLOT; s++) chips[s] += at[s] * share;
            }
            pots_settle<NSLOT>(prm.pots, key, chips);
            if (++since == prm.pots.flush_every) { chips_fold_thread<NSLOT>(chips, sred + 20); since = 0; }
        }
    }
    if constexpr (EV && NSLOT > 0) chips_fold_warp<NSLOT>(chips, sred + 20);

    // per-thread 16-bit halves (the launcher keeps a thread below 65536 trials) -> warp -> block -> global
    const int lane = threadIdx.x & 31, n_seats = NR + nf;
#pragma unroll
    for (int s = 0; s < NSLOT; s++) if (s < n_seats) {
        const uint32_t w = warp_sum(acc[s] >> 16), e = warp_sum(acc[s] & 0xFFFFu);
        if (lane == 0) { if (w) atomicAdd(&sred[s], (unsigned long long)w); if (e) atomicAdd(&sred[9 + s], (unsigned long long)e); }
    }
    const uint32_t u = warp_sum(units);
    if (lane == 0 && u) atomicAdd(&sred[18], (unsigned long long)u);
    __syncthreads();
    if ((int)threadIdx.x < n_seats) {
        const int s = threadIdx.x, os = prm.seat_of[s];
        const unsigned long long w = sred[s], e = sred[9 + s], n = sred[18];
        unsigned long long *o = reinterpret_cast<unsigned long long *>(d_counts);
        if (w) atomicAdd(&o[os], w);
        if (e - w) atomicAdd(&o[9 + os], e - w);
        if (n - e) atomicAdd(&o[18 + os], n - e);
        if (EV && sred[20 + s]) atomicAdd(&o[27 + os], sred[20 + s]);
    }
}

static size_t mc_smem(int threads) { return 13 * 4 * (size_t)(threads <= 512 ? 512 : 1024) + 64 * MC_REP * sizeof(uint2) + sizeof(HashBlob) + 32 * sizeof(unsigned long long); }
// threads per block by number of random seats: tables of up to three random seats fit 64 registers and run 1024
// threads, the others 768 x 80 registers.  MCP_K4_THREADS=512|768|1024 overrides (A/B knob of tools/k4_ab.py; 1024 only
// where the variant exists).
static int mc_threads(int nr) {
    static const int forced = [] { const char *e = getenv("MCP_K4_THREADS"); const int v = e ? atoi(e) : 0; return v == 512 || v == 768 || v == 1024 ? v : 0; }();
    if (forced == 1024) return nr <= 3 ? 1024 : 768;
    if (forced) return forced;
    return nr <= 3 ? 1024 : 768;
}

template <int NR, int MC_THREADS, int NFX, bool EV = false>
static cudaError_t launch_mc_nfx(const McFyParams &prm, const DeviceTables &tb, long long *d_counts, int grid, cudaStream_t st) {
    auto kern = equity_mc_fy_kernel<NR, MC_THREADS, NFX, EV>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mc_smem(MC_THREADS));
    if (e != cudaSuccess) return e;
    kern<<<grid, MC_THREADS, mc_smem(MC_THREADS), st>>>(prm, tb.d_blob, tb.d_card, d_counts);
    return cudaGetLastError();
}
template <int NR, int MC_THREADS>
static cudaError_t launch_mc(const McFyParams &prm, const DeviceTables &tb, long long *d_counts, int grid, cudaStream_t st) {
    constexpr int NF_ALL = MAX_SEATS - NR;
    // the lean instantiation exists for the default block size of the shape only
    constexpr bool lean = NF_ALL > 1 && MC_THREADS == (NR <= 3 ? 1024 : 768);
    if constexpr (lean) { if (prm.n_fixed <= 1) return launch_mc_nfx<NR, MC_THREADS, 1>(prm, tb, d_counts, grid, st); }
    return launch_mc_nfx<NR, MC_THREADS, NF_ALL>(prm, tb, d_counts, grid, st);
}

// EV variant: the general seat shape only
// 768 threads (the chip accumulators then squeeze the kernel into 80 registers, at most 16 bytes of spill) or, with
// MCP_K4_EV_THREADS=640, 640 threads x 96 registers (A/B arm)
static int mc_ev_threads(int nr) {
    static const bool t640 = [] { const char *e = getenv("MCP_K4_EV_THREADS"); return e && atoi(e) == 640; }();
    return (t640 && nr >= 4) ? 640 : 768;
}
template <int NR>
static cudaError_t launch_mc_ev(const McFyParams &prm, const DeviceTables &tb, long long *d_counts, int grid, cudaStream_t st) {
    if constexpr (NR >= 4) { if (mc_ev_threads(NR) == 640) return launch_mc_nfx<NR, 640, MAX_SEATS - NR, true>(prm, tb, d_counts, grid, st); }
    return launch_mc_nfx<NR, 768, MAX_SEATS - NR, true>(prm, tb, d_counts, grid, st);
}

cudaError_t launch_equity_mc_fy(const EquityArgs &a, uint64_t seed, uint64_t first_trial, uint64_t n_trials,
                                long long *d_counts, const DeviceTables &tb, const DeviceInfo &di, cudaStream_t st,
                                const PotArgs *pots) {
    if (n_trials == 0) return cudaSuccess;
    McFyParams prm;
    memset(&prm, 0, sizeof prm);
    uint64_t used = a.board | a.dead;
    int nr = 0, nf = 0;
    for (int s = 0; s < a.n_seats; s++) { used |= a.hole[s]; if (!a.hole[s]) nr++; }
    for (int s = 0, q = 0; s < a.n_seats; s++) {
        if (!a.hole[s]) { prm.seat_of[q++] = s; continue; }
        const HandState hs = state_of(a.hole[s], tb.h_card, 0);
        prm.h_t[nf] = hs.t; prm.h_sn[nf] = hs.sn; prm.h_mlo[nf] = hs.mlo; prm.h_mhi[nf] = hs.mhi;
        prm.seat_of[nr + nf] = s;
        nf++;
    }
    prm.base = state_of(a.board, tb.h_card, SN_BIAS);
    prm.n_fixed = nf; prm.n_board = 5 - __builtin_popcountll(a.board);
    for (int r = 0; r < 10; r++) { prm.rk0[r] = (uint32_t)seed + (uint32_t)r * PHILOX_W0; prm.rk1[r] = (uint32_t)(seed >> 32) + (uint32_t)r * PHILOX_W1; }
    int nl = 0;
    for (int b = 0; b < 64; b++) if (((DECK_MASK & ~used) >> b) & 1ull) { prm.deck_init[nl >> 2] |= (uint32_t)b << (8 * (nl & 3)); nl++; }
    prm.n_live = nl;
    for (int i = 0; i < MC_MAX_STEPS; i++) { const int m = nl - i; prm.m[i] = m > 0 ? (uint32_t)m : 1u; }
    for (int w = 0; w < (MC_MAX_STEPS + 2) / 3; w++) {
        uint64_t prod = 1;
        for (int i = 3 * w; i < 3 * w + 3 && i < 2 * nr + 5; i++)
            if (i < 2 * nr || i - 2 * nr < prm.n_board) prod *= prm.m[i];
        prm.thr[w] = (uint32_t)((1ull << 32) % prod);
    }
    const bool ev = pots && pots->n_pots > 0;
    make_pot_params(ev ? pots : nullptr, prm.seat_of, a.n_seats, &prm.pots, true);
    const int grid = di.sm_count;
    const int MC_THREADS = ev ? mc_ev_threads(nr) : mc_threads(nr);
    const uint64_t per_launch = 65535ull * (uint64_t)grid * MC_THREADS;       // 16-bit per-thread counter halves
    for (uint64_t done = 0; done < n_trials; done += per_launch) {
        prm.first = first_trial + done;
        prm.count = (n_trials - done) < per_launch ? (n_trials - done) : per_launch;
        cudaError_t e;
        if (ev) {
            switch (nr) {
                case 0: e = launch_mc_ev<0>(prm, tb, d_counts, grid, st); break;
                case 1: e = launch_mc_ev<1>(prm, tb, d_counts, grid, st); break;
                case 2: e = launch_mc_ev<2>(prm, tb, d_counts, grid, st); break;
                case 3: e = launch_mc_ev<3>(prm, tb, d_counts, grid, st); break;
                case 4: e = launch_mc_ev<4>(prm, tb, d_counts, grid, st); break;
                case 5: e = launch_mc_ev<5>(prm, tb, d_counts, grid, st); break;
                case 6: e = launch_mc_ev<6>(prm, tb, d_counts, grid, st); break;
                case 7: e = launch_mc_ev<7>(prm, tb, d_counts, grid, st); break;
                case 8: e = launch_mc_ev<8>(prm, tb, d_counts, grid, st); break;
                case 9: e = launch_mc_ev<9>(prm, tb, d_counts, grid, st); break;
                default: return cudaErrorInvalidValue;
            }
            if (e != cudaSuccess) return e;
            continue;
        }
        switch (nr) {
            case 0: e = MC_THREADS == 1024 ? launch_mc<0, 1024>(prm, tb, d_counts, grid, st) : MC_THREADS == 768 ? launch_mc<0, 768>(prm, tb, d_counts, grid, st) : launch_mc<0, 512>(prm, tb, d_counts, grid, st); break;
            case 1: e = MC_THREADS == 1024 ? launch_mc<1, 1024>(prm, tb, d_counts, grid, st) : MC_THREADS == 768 ? launch_mc<1, 768>(prm, tb, d_counts, grid, st) : launch_mc<1, 512>(prm, tb, d_counts, grid, st); break;
            case 2: e = MC_THREADS == 1024 ? launch_mc<2, 1024>(prm, tb, d_counts, grid, st) : MC_THREADS == 768 ? launch_mc<2, 768>(prm, tb, d_counts, grid, st) : launch_mc<2, 512>(prm, tb, d_counts, grid, st); break;
            case 3: e = MC_THREADS == 1024 ? launch_mc<3, 1024>(prm, tb, d_counts, grid, st) : MC_THREADS == 768 ? launch_mc<3, 768>(prm, tb, d_counts, grid, st) : launch_mc<3, 512>(prm, tb, d_counts, grid, st); break;
            case 4: e = MC_THREADS == 768 ? launch_mc<4, 768>(prm, tb, d_counts, grid, st) : launch_mc<4, 512>(prm, tb, d_counts, grid, st); break;
            case 5: e = MC_THREADS == 768 ? launch_mc<5, 768>(prm, tb, d_counts, grid, st) : launch_mc<5, 512>(prm, tb, d_counts, grid, st); break;
            case 6: e = MC_THREADS == 768 ? launch_mc<6, 768>(prm, tb, d_counts, grid, st) : launch_mc<6, 512>(prm, tb, d_counts, grid, st); break;
            case 7: e = MC_THREADS == 768 ? launch_mc<7, 768>(prm, tb, d_counts, grid, st) : launch_mc<7, 512>(prm, tb, d_counts, grid, st); break;
            case 8: e = MC_THREADS == 768 ? launch_mc<8, 768>(prm, tb, d_counts, grid, st) : launch_mc<8, 512>(prm, tb, d_counts, grid, st); break;
            case 9: e = MC_THREADS == 768 ? launch_mc<9, 768>(prm, tb, d_counts, grid, st) : launch_mc<9, 512>(prm, tb, d_counts, grid, st); break;
            default: return cudaErrorInvalidValue;
        }
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// ---------------------------------------------------------------------------------------------------
// Monte Carlo with a weighted hole range per seat ("hand ranges per seat (weights), which is what an AI
// client of this server would query", README.md:9; stage states of deal-stage gameplay.clj:56-61).  Deal spec R1:
// per trial and attempt a word stream Philox(key = seed, ctr = (trial_lo, trial_hi, block | attempt << 16, "MCR1")); a
// ranged seat draws a slot and a share of the slot's capacity (both exactly uniform: Lemire's bounded sampler, a word in
// its rejection zone is drawn again) from its integer alias table, i.e. a hole exactly in proportion to its weight; a hole
// that hits a card already out fails the attempt, which restarts with attempt + 1 (so the joint law is the product of the
// weights restricted to disjoint holes); the missing community cards are uniform draws from the live list, a card that
// is already out being drawn again.  Trial t depends only on (seed, t).  One thread per trial, everything in registers;
// the alias tables (8 B per hole, up to 95 KB) are read from global memory (L2), the hash image from shared memory.
// ---------------------------------------------------------------------------------------------------
constexpr uint32_t TAG_RANGE = 0x4D435231u;
constexpr int MCR_THREADS = 512;
struct WordStream {
    uint32_t k0, k1, c0, c1, c2, w[4];
    int pos;
    __device__ __forceinline__ void init(uint32_t seed_lo, uint32_t seed_hi, uint64_t trial, uint32_t attempt) {
        k0 = seed_lo; k1 = seed_hi; c0 = (uint32_t)trial; c1 = (uint32_t)(trial >> 32); c2 = attempt << 16; pos = 4;
    }
    __device__ __forceinline__ uint32_t next() {
        if (pos == 4) { philox4x32_10(c0, c1, c2, TAG_RANGE, k0, k1, w); c2++; pos = 0; }
        const uint32_t v = pos == 0 ? w[0] : (pos == 1 ? w[1] : (pos == 2 ? w[2] : w[3]));
        pos++;
        return v;
    }
    __device__ __forceinline__ uint32_t below(uint32_t n, uint32_t zone) {       // exactly uniform in [0, n)
        for (;;) {
            const uint32_t u = next();
            const uint32_t lo = u * n;
            if (lo >= zone) return __umulhi(u, n);
        }
    }
};

__global__ void __launch_bounds__(MCR_THREADS, 1)
equity_mc_ranges_kernel(const __grid_constant__ McRangeArgs a, const RangeEntry *__restrict__ entries, uint32_t seed_lo, uint32_t seed_hi,
                        uint64_t first, uint64_t count, const HashBlob *__restrict__ g_blob, const CardKey *__restrict__ g_card,
                        long long *__restrict__ d_counts) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    HashBlob *H = reinterpret_cast<HashBlob *>(smem_raw);
    __shared__ CardKey scard[64];
    __shared__ unsigned char slive[52];
    __shared__ unsigned long long sred[20];
    __shared__ unsigned long long mbar;
    BulkStage stage;
    stage.start(&mbar, H, g_blob, (uint32_t)sizeof(HashBlob));
    if (threadIdx.x < 64) scard[threadIdx.x] = g_card[threadIdx.x];
    if (threadIdx.x < 52) slive[threadIdx.x] = a.live[threadIdx.x];
    if (threadIdx.x < 20) sred[threadIdx.x] = 0ull;
    stage.wait();
    const uint32_t tbl = staged_table_base(H);

    // hand state of the known community cards
    HandState base{0, SN_BIAS, (uint32_t)a.board, (uint32_t)(a.board >> 32)};
    {
        uint64_t x = a.board;
        while (x) { const int p = __ffsll((long long)x) - 1; x &= x - 1; base.t += scard[p].rk; base.sn += scard[p].sn; }
    }
    const int need = 5 - __popcll(a.board), n_seats = a.n_seats;
    Tally<MAX_SEATS> tl; tl.clear();
    uint32_t voids = 0;
    const uint64_t nthreads = (uint64_t)gridDim.x * blockDim.x;
    // One ATTEMPT per loop iteration and lane: a lane whose holes collide retries the same trial with the next attempt number
    // in its next iteration while its neighbours go on to their next trials, so a warp never waits for its unluckiest lane
    // (with a restart loop per trial the warp ran max-over-lanes attempts, 3-4 instead of 1.3).  (trial, attempt) streams
    // are what they were: same counts.
    uint64_t tr = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t attempt = 0;
    while (tr < count) {
        const uint64_t trial = first + tr;
        uint64_t hole[MAX_SEATS], used = a.used0, bm = a.board;
        WordStream ws; ws.init(seed_lo, seed_hi, trial, attempt);
        bool fail = false;
#pragma unroll
        for (int s = 0; s < MAX_SEATS; s++) {
            if (s < n_seats && !fail) {
                hole[s] = a.hole[s];
                if (a.rs[s].n) {
                    const uint32_t slot = ws.below(a.rs[s].n, a.rs[s].zone_n), vv = ws.below(a.rs[s].W, a.rs[s].zone_W);
                    RangeEntry e = entries[a.rs[s].off + slot];
                    if (vv >= e.thr) e = entries[a.rs[s].off + e.alias];
                    const uint64_t h = (1ull << e.c0) | (1ull << e.c1);
                    fail = (h & used) != 0ull;
                    used |= h; hole[s] = h;
                }
            }
        }
        if (fail) {
            if (++attempt == (uint32_t)RANGE_MAX_ATTEMPTS) { voids++; tr += nthreads; attempt = 0; }
            continue;
        }
        for (int k = 0; k < need; k++) {
            for (;;) {
                const uint64_t bit = 1ull << slive[ws.below(a.n_live, a.zone_live)];
                if (!(used & bit)) { used |= bit; bm |= bit; break; }
            }
        }
        tr += nthreads; attempt = 0;
        // the board of this trial, then the seats
        HandState b = base;
        {
            uint64_t x = bm & ~a.board;
            while (x) { const int p = __ffsll((long long)x) - 1; x &= x - 1; b.t += scard[p].rk; b.sn += scard[p].sn; }
            b.mlo = (uint32_t)bm; b.mhi = (uint32_t)(bm >> 32);
        }
        uint32_t key[MAX_SEATS];
#pragma unroll
        for (int s = 0; s < MAX_SEATS; s++) {
            key[s] = 0;
            if (s < n_seats) {
                const uint64_t h = hole[s];
                const int p1 = __ffsll((long long)h) - 1, p2 = 63 - __clzll((long long)h);
                key[s] = hb_class(b.t + scard[p1].rk + scard[p2].rk, b.sn + scard[p1].sn + scard[p2].sn,
                                  b.mlo | (uint32_t)h, b.mhi | (uint32_t)(h >> 32), tbl);
            }
        }
        tl.add(key, n_seats);
    }
    const uint32_t v = warp_sum(voids);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(&sred[19], (unsigned long long)v);
    tl.flush(n_seats, sred, d_counts);                     // (synchronises the block)
    if (threadIdx.x == 0 && sred[19]) atomicAdd(reinterpret_cast<unsigned long long *>(d_counts) + 27, sred[19]);
}

cudaError_t launch_equity_mc_ranges(const McRangeArgs &a, const RangeEntry *d_entries, uint64_t seed, uint64_t first_trial,
                                    uint64_t n_trials, long long *d_counts, const DeviceTables &tb, const DeviceInfo &di, cudaStream_t st) {
    if (n_trials == 0) return cudaSuccess;
    const size_t smem = sizeof(HashBlob);
    cudaError_t e = cudaFuncSetAttribute(equity_mc_ranges_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    uint64_t want = (n_trials + MCR_THREADS - 1) / MCR_THREADS;
    const int grid = (int)(want > (uint64_t)di.sm_count ? (uint64_t)di.sm_count : want);
    // (the per-thread counters are 32-bit: one launch covers at most 2^31 trials per thread-slot; the caller's trial
    // counts are far below grid * threads * 2^31)
    equity_mc_ranges_kernel<<<grid, MCR_THREADS, smem, st>>>(a, d_entries, (uint32_t)seed, (uint32_t)(seed >> 32), first_trial, n_trials,
                                                            tb.d_blob, tb.d_card, d_counts);
    return cudaGetLastError();
}

}  // namespace mcp
