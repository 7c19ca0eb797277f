// kernels_boards.cu -- the 169 x 169 preflop matrix by BOARDS: sort-and-count instead of playing
// every matchup over every board.
//   For one 5-card board the 1,081 hole pairs of the other 47 cards are evaluated ONCE (K7a) and sorted by strength.
//   wins(A, B) over that board = #{(ha in A, hb in B) : ha stronger than hb, ha and hb disjoint}
//                              = #{(ha, hb) : ha stronger}  -  #{(ha, hb) sharing a card : ha stronger}.
//   The first term is a scan of the sorted holes with a running per-class count of the strictly weaker ones
//   (row class(ha) += running); the second is a walk over the 47 x C(46,2) pairs of holes with a common card (K7b).
//   Boards are folded by the 24 suit permutations (134,459 canonical boards, weight = orbit size): the cell of a
//   pair of hole classes does not depend on suit names.  Ties and losses follow on the host from the win matrix:
//   loss(A,B) = win(B,A), tie(A,B) = pairs(A,B) * C(48,5) - win(A,B) - win(B,A).
// Counts are exact integers; the result must equal the brute-force matrix of K5 cell by cell (tests).
#include "kernels.cuh"
#include "common.cuh"
#include "hashtab.cuh"
#include <algorithm>

namespace mcp {

constexpr int PB_THREADS = 1024;
constexpr int PB_LIVE = 47;
constexpr int PB_HOLES = PB_LIVE * (PB_LIVE - 1) / 2;      // 1,081
constexpr int PB_CLASSES = 169;
constexpr int PB_WORDS = (PB_CLASSES + 1) / 2;             // a row of 16-bit counters, two per word
constexpr int PB_BINS = 4896;                              // dense strengths (4,892 classes), padded to 1024 x 5 - ...
constexpr int PB_FLUSH = 16;                               // boards between folds of the 16-bit counters: 16 * 24 * 144 < 65536
constexpr int PB_SCAN_THREADS = 96;                        // three warps scan (85 lanes own two classes each); the rest correct
constexpr uint64_t PB_DECK = 0x7FFC7FFC7FFC7FFCull;

__device__ __forceinline__ uint32_t pb_class_of(int p1, int p2) {                       // 169-class id of two bit positions
    const int r1 = (p1 & 15) - 2, r2 = (p2 & 15) - 2;
    const int hi = max(r1, r2), lo = min(r1, r2);
    return (uint32_t)(((p1 >> 4) == (p2 >> 4) || hi == lo) ? 13 * hi + lo : 13 * lo + hi);
}

// colex pair table: entry h = i | j << 8 with i < j, h = j (j - 1) / 2 + i
__device__ __forceinline__ void pb_fill_pairs(uint16_t *spair) {
    for (int h = threadIdx.x; h < PB_HOLES; h += blockDim.x) {
        int j = (int)((1.0f + sqrtf(1.0f + 8.0f * (float)h)) * 0.5f);
        while (j * (j - 1) / 2 > h) j--;
        while ((j + 1) * j / 2 <= h) j++;
        spair[h] = (uint16_t)((h - j * (j - 1) / 2) | (j << 8));
    }
}

// ---------------------------------------------------------------------------------------------------
// K7a: per board, info[h] = dense strength << 8 | hole class for every hole h (colex index over the live cards in
// ascending bit position), and the same words sorted by strength (counting sort over the 4,892 dense classes).
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(PB_THREADS, 1)
pf_board_eval_kernel(const PfBoard *__restrict__ boards, int nb, const HashBlob *__restrict__ g_blob,
                     const CardKey *__restrict__ g_card, const __grid_constant__ CatBase cb,
                     uint32_t *__restrict__ out_sorted, uint32_t *__restrict__ out_info) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    HashBlob *H = reinterpret_cast<HashBlob *>(smem_raw);
    uint32_t *shist = reinterpret_cast<uint32_t *>(smem_raw + sizeof(HashBlob));        // PB_BINS
    __shared__ CardKey scard[64];
    __shared__ uint16_t spair[PB_HOLES + 1];
    __shared__ unsigned char slive[64];
    __shared__ uint32_t swarp[32];
    __shared__ unsigned long long mbar;
    BulkStage stage;
    stage.start(&mbar, H, g_blob, (uint32_t)sizeof(HashBlob));
    pb_fill_pairs(spair);
    if (threadIdx.x < 64) scard[threadIdx.x] = g_card[threadIdx.x];
    stage.wait();
    const uint32_t tbl = staged_table_base(H);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    for (int bi = blockIdx.x; bi < nb; bi += gridDim.x) {
        const PfBoard B = boards[bi];
        uint64_t bm = 0;
        uint32_t bt = 0, bsn = SN_BIAS;
#pragma unroll
        for (int k = 0; k < 5; k++) { bm |= 1ull << B.card[k]; bt += scard[B.card[k]].rk; bsn += scard[B.card[k]].sn; }
        const uint64_t livem = PB_DECK & ~bm;
        if (threadIdx.x < 64 && ((livem >> threadIdx.x) & 1ull))
            slive[__popcll(livem & ((1ull << threadIdx.x) - 1ull))] = (unsigned char)threadIdx.x;
        for (int i = threadIdx.x; i < PB_BINS; i += PB_THREADS) shist[i] = 0u;
        __syncthreads();
        // 1,081 holes over 1,024 threads: two rounds
        uint32_t info[2] = {0u, 0u}, dense[2] = {0u, 0u};
#pragma unroll
        for (int r = 0; r < 2; r++) {
            const int h = threadIdx.x + r * PB_THREADS;
            if (h < PB_HOLES) {
                const uint32_t pr = spair[h];
                const int p1 = slive[pr & 255u], p2 = slive[pr >> 8];
                const uint64_t m = bm | (1ull << p1) | (1ull << p2);
                const uint32_t t = bt + scard[p1].rk + scard[p2].rk, sn = bsn + scard[p1].sn + scard[p2].sn;
                const uint32_t flag = sn & 0x8888u;
                const uint32_t w = (flag & 0x8800u) ? (uint32_t)(m >> 32) : (uint32_t)m;
                const uint32_t ln = (flag & 0x8080u) ? (w >> 16) : (w & 0xFFFFu);
                const uint32_t d = lds_u16(tbl + (t & 0x3FFEu));
                const uint32_t va = tbl + HT_BUCKETS * 2 + HT_FLUSH * 2 + 2u * ((t >> 16) + d), fa = tbl + HT_BUCKETS * 2 + 2u * (ln >> 2);
                const uint32_t code = lds_u16(flag ? fa : va);
                dense[r] = code;                                   // class codes are dense indices
                info[r] = dense[r] << 8 | pb_class_of(p1, p2);
                atomicAdd(&shist[dense[r]], 1u);
            }
        }
        __syncthreads();
        // exclusive prefix sum of the histogram: five consecutive bins per thread, warp scan, scan of the warp totals
        {
            const int base = threadIdx.x * 5;
            uint32_t v[5], s = 0;
#pragma unroll
            for (int k = 0; k < 5; k++) { v[k] = (base + k < PB_BINS) ? shist[base + k] : 0u; s += v[k]; }
            uint32_t inc = s;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t n = __shfl_up_sync(0xFFFFFFFFu, inc, o); if (lane >= o) inc += n; }
            if (lane == 31) swarp[warp] = inc;
            __syncthreads();
            if (warp == 0) {
                uint32_t t = swarp[lane], ti = t;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const uint32_t n = __shfl_up_sync(0xFFFFFFFFu, ti, o); if (lane >= o) ti += n; }
                swarp[lane] = ti - t;
            }
            __syncthreads();
            uint32_t run = swarp[warp] + inc - s;
#pragma unroll
            for (int k = 0; k < 5; k++) { if (base + k < PB_BINS) shist[base + k] = run; run += v[k]; }
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < 2; r++) {
            const int h = threadIdx.x + r * PB_THREADS;
            if (h < PB_HOLES) {
                const uint32_t pos = atomicAdd(&shist[dense[r]], 1u);
                // sorted word: strength << 16 | byte offset of the hole class's row of 16-bit counters (K7b adds there)
                out_sorted[(size_t)bi * PB_STRIDE + pos] = dense[r] << 16 | ((info[r] & 255u) * (PB_WORDS * 4u));
                out_info[(size_t)bi * PB_STRIDE + h] = info[r];
            }
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------
// K7b: the win matrix.  A block keeps 169 x 169 32-bit counters and a 16-bit copy (two classes per word) that the
// boards add into and that is folded into the 32-bit counters every PB_FLUSH boards.  Four boards are in flight per
// block, eight warps each: three warps scan the sorted holes (a scan step is a short dependent chain, so one board
// alone leaves the SM idle), five warps walk the pairs of holes with a common card.  All updates are word-wide shared
// atomics, so scans and corrections of different boards run concurrently in any order: a 16-bit field may go
// "negative" in between, the word arithmetic stays exact and every field is back in range at the fold.
// ---------------------------------------------------------------------------------------------------
constexpr int PB_GROUPS = 4, PB_GROUP_THREADS = PB_THREADS / PB_GROUPS;   // boards in flight per block, threads per board
constexpr int PB_CORR_WARPS = PB_GROUP_THREADS / 32 - PB_SCAN_THREADS / 32;
static_assert(PB_FLUSH % PB_GROUPS == 0, "a fold follows a whole number of rounds");
static_assert(PB_WORDS * 4 * (PB_CLASSES - 1) < 65536, "row offsets fit the low half of a sorted word");

__device__ __forceinline__ void pb_group_sync(int group) {
    asm volatile("bar.sync %0, %1;" ::"r"(1 + group), "r"(PB_GROUP_THREADS) : "memory");
}

__global__ void __launch_bounds__(PB_THREADS, 1)
pf_board_count_kernel(const PfBoard *__restrict__ boards, int nb, const uint32_t *__restrict__ g_sorted,
                      const uint32_t *__restrict__ g_info, unsigned long long *__restrict__ g_win) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t *W32 = reinterpret_cast<uint32_t *>(smem_raw);                 // [169][169]
    uint32_t *W16 = W32 + PB_CLASSES * PB_CLASSES;                          // [169][PB_WORDS]
    uint32_t *sbuf = W16 + PB_CLASSES * PB_WORDS;                           // per group: sorted words, info words
    uint32_t *srow = sbuf + PB_GROUPS * 2 * PB_STRIDE;                      // per correction warp: the 46 holes of one card
    __shared__ uint16_t spair[PB_HOLES + 1];
    pb_fill_pairs(spair);
    for (int i = threadIdx.x; i < PB_CLASSES * PB_CLASSES; i += PB_THREADS) W32[i] = 0u;
    for (int i = threadIdx.x; i < PB_CLASSES * PB_WORDS; i += PB_THREADS) W16[i] = 0u;
    __syncthreads();

    const int group = threadIdx.x / PB_GROUP_THREADS, gt = threadIdx.x % PB_GROUP_THREADS, lane = threadIdx.x & 31;
    uint32_t *sp = sbuf + group * 2 * PB_STRIDE, *ip = sp + PB_STRIDE;
    const uint32_t w16 = (uint32_t)__cvta_generic_to_shared(W16);
    // a block takes contiguous batches of PB_FLUSH boards; group g plays boards g, g + 4, ... of the batch
    for (int batch = blockIdx.x * PB_FLUSH; batch < nb; batch += gridDim.x * PB_FLUSH) {
        for (int bi = batch + group; bi < min(nb, batch + PB_FLUSH); bi += PB_GROUPS) {
            const uint32_t wgt = boards[bi].weight;
            for (int i = gt; i < PB_HOLES; i += PB_GROUP_THREADS) {
                sp[i] = g_sorted[(size_t)bi * PB_STRIDE + i];
                ip[i] = g_info[(size_t)bi * PB_STRIDE + i];
            }
            pb_group_sync(group);
            if (gt < PB_SCAN_THREADS) {
                // ascending strength: `running` = holes strictly weaker than the current one, by class (this lane's two classes)
                if (gt < PB_WORDS) {
                    const uint32_t my = w16 + 4u * gt, own0 = 2u * gt * (PB_WORDS * 4u), own1 = own0 + PB_WORDS * 4u;
                    uint32_t running = 0, pending = 0, prev = 0xFFFFFFFFu;
#pragma unroll 4
                    for (int i = 0; i < PB_HOLES; i++) {
                        const uint32_t e = sp[i], d = e >> 16, ro = e & 0xFFFFu;
                        if (d != prev) { running += pending; pending = 0u; prev = d; }
                        asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(my + ro), "r"(running) : "memory");
                        pending += ro == own0 ? wgt : (ro == own1 ? wgt << 16 : 0u);
                    }
                }
            } else {
                // pairs of holes with a common card were counted above as if they could meet: take them out again.
                // A warp takes one common card at a time: its 46 holes go to a row, then the C(46,2) pairs are compared.
                const int cw = (gt - PB_SCAN_THREADS) >> 5;
                uint32_t *row = srow + (group * PB_CORR_WARPS + cw) * 48;
                for (int c = cw; c < PB_LIVE; c += PB_CORR_WARPS) {
                    for (int u = lane; u < PB_LIVE - 1; u += 32) {
                        const int li = u + (u >= c);
                        row[u] = ip[li < c ? c * (c - 1) / 2 + li : li * (li - 1) / 2 + c];
                    }
                    __syncwarp();
                    for (int p = lane; p < (PB_LIVE - 1) * (PB_LIVE - 2) / 2; p += 32) {
                        const uint32_t pr = spair[p];
                        const uint32_t x = row[pr & 255u], y = row[pr >> 8];
                        if ((x >> 8) != (y >> 8)) {
                            const bool xw = (x >> 8) > (y >> 8);
                            const uint32_t a = (xw ? x : y) & 255u, b = (xw ? y : x) & 255u;
                            atomicSub(&W16[a * PB_WORDS + (b >> 1)], wgt << (16u * (b & 1u)));
                        }
                    }
                    __syncwarp();
                }
            }
            pb_group_sync(group);
        }
        __syncthreads();
        for (int i = threadIdx.x; i < PB_CLASSES * PB_WORDS; i += PB_THREADS) {
            const uint32_t w = W16[i];
            if (w) {
                W16[i] = 0u;
                const int a = i / PB_WORDS, l = i - a * PB_WORDS;
                W32[a * PB_CLASSES + 2 * l] += w & 0xFFFFu;
                if (2 * l + 1 < PB_CLASSES) W32[a * PB_CLASSES + 2 * l + 1] += w >> 16;
            }
        }
        __syncthreads();
    }
    for (int i = threadIdx.x; i < PB_CLASSES * PB_CLASSES; i += PB_THREADS)
        if (W32[i]) atomicAdd(&g_win[i], (unsigned long long)W32[i]);
}

cudaError_t launch_preflop_boards(const PfBoard *d_boards, int n_boards, uint32_t *d_sorted, uint32_t *d_info,
                                  unsigned long long *d_win, const DeviceTables &tb, const DeviceInfo &di, cudaStream_t st) {
    if (n_boards <= 0) return cudaSuccess;
    const size_t smem_a = sizeof(HashBlob) + PB_BINS * sizeof(uint32_t);
    const size_t smem_b = ((size_t)PB_CLASSES * PB_CLASSES + (size_t)PB_CLASSES * PB_WORDS + (size_t)PB_GROUPS * 2 * PB_STRIDE +
                           (size_t)PB_GROUPS * PB_CORR_WARPS * 48) * sizeof(uint32_t);
    cudaError_t e = cudaFuncSetAttribute(pf_board_eval_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_a);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(pf_board_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_b);
    if (e != cudaSuccess) return e;
    const int grid = n_boards < di.sm_count ? n_boards : di.sm_count;
    pf_board_eval_kernel<<<grid, PB_THREADS, smem_a, st>>>(d_boards, n_boards, tb.d_blob, tb.d_card, tb.cat_base, d_sorted, d_info);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    const int grid_b = std::min(di.sm_count, (n_boards + PB_FLUSH - 1) / PB_FLUSH);
    pf_board_count_kernel<<<grid_b, PB_THREADS, smem_b, st>>>(d_boards, n_boards, d_sorted, d_info, d_win);
    return cudaGetLastError();
}

}  // namespace mcp
