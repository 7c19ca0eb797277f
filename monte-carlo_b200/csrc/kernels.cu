// kernels.cu -- hand-written sm_100a kernels for the hand-ranking / showdown / equity path.
//   K1 eval7_stream        batch of card masks -> keys            (player->hand-value, hand_evaluator.clj:162-172)
//   K2 enum7               all C(52,7) hands -> category histogram + key checksums (BASELINE config 2)
//   K3 equity_exhaustive   every completion of the board -> win/tie/loss per seat (end-game, gameplay.clj:127-132)
//   K4 equity_mc           Philox deal per trial + showdown      (board.clj:148, gameplay.clj:63-75,41-54,127-132)
//   K5 preflop             canonical matchups x all boards        (BASELINE config 5)
// Integer / lookup work: no tensor cores, no TMA (nothing is GEMM-shaped; inputs are either a linear stream
// or generated in registers).  Grids are multiples of the SM count; per-thread register counters are reduced
// with warp REDUX / shuffles -> shared memory -> a handful of 64-bit global atomics per block.
#include "kernels.cuh"
#include "common.cuh"
#include "eval.cuh"
#include "eval_tbl.cuh"
#include "hashtab.cuh"
#include "philox.cuh"

namespace mcp {

uint64_t host_binom(int n, int k) { return cbinom(n, k); }

// card id (13*suit + rank-2) <-> bit position (16*suit + rank)
__device__ __forceinline__ int id_to_pos(int id) { return id + 2 + 3 * (id / 13); }
__device__ __forceinline__ int pos_succ(int p) { return ((p & 15) == 14) ? p + 4 : p + 1; }
__device__ __forceinline__ uint64_t pos_bit(int p) { return 1ull << p; }

// Rank-set -> nibble table T[x], x a 14-bit rank set (bit r-1): the ranks of x in descending order as
// nibbles, left-aligned at bits 16..19, at most five.  64 KB of shared memory; turns the public nibble
// key into two conflict-tolerant LDS + three IMAD instead of FLO/clz loops.
constexpr int NIB_ENTRIES = 1 << 14;
__device__ __forceinline__ void fill_nib_table(uint32_t *T) {
    for (uint32_t x = threadIdx.x; x < NIB_ENTRIES; x += blockDim.x) {
        uint32_t m = x, v = 0;
        int shift = 16;
        while (m && shift >= 0) {
            const uint32_t b = 31 - __clz(m);
            m ^= 1u << b;
            v |= (b + 1) << shift;
            shift -= 4;
        }
        T[x] = v;
    }
}
// per-category constants of the conversion: key = cat<<20 + a1*MA + a2*MB + (b >> SH)*MS - SUB
//   a = T[P] (a1, a2 its first two nibbles), b = T[S]
__constant__ uint4 CAT_CONST[9] = {
    /*0 high card */ {0u, 0u, 1u, 0u | (0u << 16)},
    /*1 pair      */ {0x11000u, 0u, 1u, 0u | (8u << 16)},
    /*2 two pair  */ {0x11000u, 0x110u, 1u, 0u | (16u << 16)},
    /*3 trips     */ {0x11100u, 0u, 1u, 0u | (12u << 16)},
    /*4 straight  */ {0u, 0u, 0x11111u, 0x1234u | (16u << 16)},
    /*5 flush     */ {0u, 0u, 1u, 0u | (0u << 16)},
    /*6 full house*/ {0x11100u, 0u, 0x11u, 0u | (16u << 16)},
    /*7 quads     */ {0x11110u, 0u, 1u, 0u | (16u << 16)},
    /*8 str flush */ {0u, 0u, 0x11111u, 0x1234u | (16u << 16)},
};
__device__ __forceinline__ uint32_t nibble_from_packed_tbl(uint32_t k, const uint32_t *T, const uint4 *cc) {
    const uint32_t cat = k >> 28;
    const uint32_t a = T[(k >> 14) & 0x3FFFu], b = T[k & 0x3FFFu];
    const uint4 c = cc[cat];
    return (cat << 20) + (a >> 16) * c.x + ((a >> 12) & 15u) * c.y + (b >> (c.w >> 16)) * c.z - (c.w & 0xFFFFu);
}

// ---------------------------------------------------------------------------------------------------
// K1: streaming batch evaluation.  8 B in, 4 B out per hand; every LDG.128 / STG.64 of a warp is one
// contiguous 512 B / 256 B segment.  Invalid hands (not 5..7 cards, or a non-card bit) -> 0xFFFFFFFF.
// ---------------------------------------------------------------------------------------------------
constexpr int K1_THREADS = 512;

template <int MODE>
__device__ __forceinline__ uint32_t k1_key(uint64_t m, const uint32_t *T, const uint4 *cc) {
    const uint32_t lo = (uint32_t)m, hi = (uint32_t)(m >> 32);
    uint32_t k;
    // lanes are clipped to the rank bits so that an invalid mask can never index outside the table
    if (MODE == K1_NIBBLE_TABLE) k = eval_nibble_tbl<14>(lo & 0x7FFE7FFEu, hi & 0x7FFE7FFEu, T, reinterpret_cast<const uint32_t *>(cc));
    else k = eval_packed(lo, hi);
    if (MODE == K1_NIBBLE_FLO) k = packed_to_nibble(k);
    const uint32_t n = (uint32_t)(__popc(lo) + __popc(hi));
    const bool ok = (((lo | hi) & 0x80018001u) == 0u) && (n - 5u <= 2u);
    return ok ? k : 0xFFFFFFFFu;
}

template <int MODE>
__global__ void __launch_bounds__(K1_THREADS)
eval7_stream_kernel(const uint64_t *__restrict__ masks, uint32_t *__restrict__ keys, size_t n) {
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t *T = smem;
    uint4 *cc = reinterpret_cast<uint4 *>(smem + NIB_ENTRIES);
    if (MODE == K1_NIBBLE_TABLE) {
        fill_nib_table_v2<14>(T, cc);
        __syncthreads();
    }
    const size_t npair = n >> 1;
    const ulonglong2 *in = reinterpret_cast<const ulonglong2 *>(masks);
    uint2 *out = reinterpret_cast<uint2 *>(keys);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npair; i += 2 * stride) {
        const size_t j = i + stride;
        const bool two = j < npair;
        const ulonglong2 a = __ldcs(in + i);
        ulonglong2 b = make_ulonglong2(0ull, 0ull);
        if (two) b = __ldcs(in + j);
        uint2 ka, kb;
        ka.x = k1_key<MODE>(a.x, T, cc);
        ka.y = k1_key<MODE>(a.y, T, cc);
        kb.x = k1_key<MODE>(b.x, T, cc);
        kb.y = k1_key<MODE>(b.y, T, cc);
        __stcs(out + i, ka);
        if (two) __stcs(out + j, kb);
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) keys[n - 1] = k1_key<MODE>(masks[n - 1], T, cc);
}

// scalar variant for buffers that are not 16-byte aligned
template <int MODE>
__global__ void __launch_bounds__(K1_THREADS)
eval7_scalar_kernel(const uint64_t *__restrict__ masks, uint32_t *__restrict__ keys, size_t n) {
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t *T = smem;
    uint4 *cc = reinterpret_cast<uint4 *>(smem + NIB_ENTRIES);
    if (MODE == K1_NIBBLE_TABLE) {
        fill_nib_table_v2<14>(T, cc);
        __syncthreads();
    }
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        keys[i] = k1_key<MODE>(masks[i], T, cc);
}

template <int MODE>
static cudaError_t launch_k1(const uint64_t *d_masks, size_t n, uint32_t *d_keys, const DeviceInfo &di, cudaStream_t st) {
    const size_t smem = (MODE == K1_NIBBLE_TABLE) ? (NIB_ENTRIES * 4 + 9 * 16) : 0;
    const bool aligned = ((reinterpret_cast<uintptr_t>(d_masks) & 15) == 0) && ((reinterpret_cast<uintptr_t>(d_keys) & 7) == 0);
    auto kern = aligned ? eval7_stream_kernel<MODE> : eval7_scalar_kernel<MODE>;
    if (smem) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    const int per_sm = (MODE == K1_NIBBLE_TABLE) ? 3 : 4;      // 3 x (64 KB table) / SM, or 4 x 512 threads
    const size_t per_thread = aligned ? 4 : 1;
    size_t want = (n + (size_t)K1_THREADS * per_thread - 1) / ((size_t)K1_THREADS * per_thread);
    size_t cap = (size_t)di.sm_count * per_sm;
    int grid = (int)(want < 1 ? 1 : (want > cap ? cap : want));
    kern<<<grid, K1_THREADS, smem, st>>>(d_masks, d_keys, n);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------
// K1 v3: streaming evaluation through the additive-key / perfect-hash tables (hashtab.cuh "K1 layout").
// A valid 7-card hand of the 52-card deck costs eight conflict-free byte-table loads (one PRMT each builds the
// address: row = a byte of the mask, column = the lane), the displacement, the slot and the key; the ALU only
// sums, finds a flush lane and checks validity.  Anything else (5/6 cards, the tests' rank 1, invalid masks)
// takes the bit-plane evaluator of eval.cuh on a divergent slow path.
// ---------------------------------------------------------------------------------------------------
constexpr int K1H_THREADS = 1024;

__device__ __forceinline__ uint32_t k1h_key(uint64_t m, const unsigned char *sb, uint32_t lane4) {
    const uint32_t lo = (uint32_t)m, hi = (uint32_t)(m >> 32);
    auto row = [&](uint32_t w, uint32_t sel, uint32_t half) -> uint32_t {
        return *reinterpret_cast<const uint32_t *>(sb + __byte_perm(w, lane4, sel) + half);
    };
    const uint32_t t = (row(lo, 0x5504, 0) + row(lo, 0x5514, 128) + row(lo, 0x5524, 0) + row(lo, 0x5534, 128)) +
                       (row(hi, 0x5504, 0) + row(hi, 0x5514, 128) + row(hi, 0x5524, 0) + row(hi, 0x5534, 128));
    // flush lane: only the word holding >= 5 cards can contain it
    const int p01 = __popc(lo), p23 = __popc(hi);
    const bool up = p23 >= 5;
    const uint32_t w = up ? hi : lo;
    const int pw = up ? p23 : p01;
    const uint32_t wl = w & 0xFFFFu, wh = w >> 16;
    const int pl = __popc(wl), ph = pw - pl;
    const uint32_t fm = pl >= 5 ? wl : (ph >= 5 ? wh : 0u);
    const uint32_t d = *reinterpret_cast<const uint16_t *>(sb + SB_D_OFF + (t & 0x3FFEu));
    uint32_t ka = *reinterpret_cast<const uint16_t *>(sb + SB_V_OFF + 2u * ((t >> 16) + d));
    if (fm) ka = *reinterpret_cast<const uint16_t *>(sb + sb_flush_addr((fm >> 2) & 0x1FFFu));   // masked: an invalid mask must stay inside the table
    uint32_t key = *reinterpret_cast<const uint32_t *>(sb + ka);
    const bool deck7 = (p01 + p23 == 7) && (((lo | hi) & 0x80038003u) == 0u);
    if (!deck7) {
        const uint32_t n = (uint32_t)(p01 + p23);
        const bool ok = (((lo | hi) & 0x80018001u) == 0u) && (n - 5u <= 2u);
        key = ok ? packed_to_nibble(eval_packed(lo, hi)) : 0xFFFFFFFFu;
    }
    return key;
}

__global__ void __launch_bounds__(K1H_THREADS, 1)
eval7_stream_hash_kernel(const uint64_t *__restrict__ masks, uint32_t *__restrict__ keys, size_t n,
                         const uint4 *__restrict__ g_blob) {
    extern __shared__ __align__(16) unsigned char sb[];
    {
        uint4 *s = reinterpret_cast<uint4 *>(sb);
        for (int i = threadIdx.x; i < (int)(SB_BYTES / 16); i += blockDim.x) s[i] = __ldg(g_blob + i);
    }
    __syncthreads();
    const uint32_t lane4 = (threadIdx.x & 31u) * 4u;
    const size_t npair = n >> 1;
    const ulonglong2 *in = reinterpret_cast<const ulonglong2 *>(masks);
    uint2 *out = reinterpret_cast<uint2 *>(keys);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < npair; i += 2 * stride) {
        const size_t j = i + stride;
        const bool two = j < npair;
        const ulonglong2 a = __ldcs(in + i);
        ulonglong2 b = make_ulonglong2(0ull, 0ull);
        if (two) b = __ldcs(in + j);
        uint2 ka, kb;
        ka.x = k1h_key(a.x, sb, lane4);
        ka.y = k1h_key(a.y, sb, lane4);
        kb.x = k1h_key(b.x, sb, lane4);
        kb.y = k1h_key(b.y, sb, lane4);
        __stcs(out + i, ka);
        if (two) __stcs(out + j, kb);
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) keys[n - 1] = k1h_key(masks[n - 1], sb, lane4);
}

cudaError_t launch_eval7_stream_hash(const uint64_t *d_masks, size_t n, uint32_t *d_keys, const void *d_stream_blob,
                                     const DeviceInfo &di, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(eval7_stream_hash_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SB_BYTES);
    if (e != cudaSuccess) return e;
    size_t want = (n + (size_t)K1H_THREADS * 4 - 1) / ((size_t)K1H_THREADS * 4);
    const int grid = (int)(want > (size_t)di.sm_count ? (size_t)di.sm_count : want);
    eval7_stream_hash_kernel<<<grid, K1H_THREADS, SB_BYTES, st>>>(d_masks, d_keys, n, reinterpret_cast<const uint4 *>(d_stream_blob));
    return cudaGetLastError();
}

cudaError_t launch_eval7_stream(const uint64_t *d_masks, size_t n, uint32_t *d_keys, int mode,
                                const DeviceInfo &di, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    switch (mode) {
        case K1_PACKED: return launch_k1<K1_PACKED>(d_masks, n, d_keys, di, st);
        case K1_NIBBLE_FLO: return launch_k1<K1_NIBBLE_FLO>(d_masks, n, d_keys, di, st);
        default: return launch_k1<K1_NIBBLE_TABLE>(d_masks, n, d_keys, di, st);
    }
}

// ---------------------------------------------------------------------------------------------------
// Philox card stream (philox.cuh spec).  take() returns the next accepted card's bit position.
// ---------------------------------------------------------------------------------------------------
struct CardStream {
    uint32_t k0, k1, c0, c1, blk, tag;
    uint32_t w[4];
    int pos;
    __device__ __forceinline__ void init(uint64_t seed, uint64_t index, uint32_t tag_) {
        k0 = (uint32_t)seed; k1 = (uint32_t)(seed >> 32);
        c0 = (uint32_t)index; c1 = (uint32_t)(index >> 32);
        blk = 0; tag = tag_; pos = 20;
    }
    __device__ __forceinline__ int next6() {
        if (pos == 20) { philox4x32_10(c0, c1, blk, tag, k0, k1, w); blk++; pos = 0; }
        const int j = pos / 5, i = pos % 5;
        const uint32_t word = j == 0 ? w[0] : (j == 1 ? w[1] : (j == 2 ? w[2] : w[3]));
        pos++;
        return (int)((word >> (6 * i)) & 63u);
    }
    __device__ __forceinline__ int take(uint64_t &used) {
        for (;;) {
            const int v = next6();
            const uint64_t bit = 1ull << v;
            if (!(used & bit)) { used |= bit; return v; }
        }
    }
};

__global__ void __launch_bounds__(256)
random_hands_kernel(uint64_t seed, uint64_t first, size_t n, int n_cards, uint64_t *__restrict__ masks) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        CardStream cs; cs.init(seed, first + i, TAG_HAND);
        uint64_t used = ~DECK_MASK, m = 0;
        for (int c = 0; c < n_cards; c++) m |= 1ull << cs.take(used);
        masks[i] = m;
    }
}
cudaError_t launch_random_hands(uint64_t seed, uint64_t first, size_t n, int n_cards, uint64_t *d_masks,
                                const DeviceInfo &di, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    size_t want = (n + 255) / 256, cap = (size_t)di.sm_count * 8;
    int grid = (int)(want > cap ? cap : want);
    random_hands_kernel<<<grid, 256, 0, st>>>(seed, first, n, n_cards, d_masks);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------
// K2: exhaustive enumeration of C(52,7) in colex order of card ids.  A thread takes chunks of ENUM_CHUNK
// consecutive combination indices: one unranking per chunk, then the successor rule with the OR of the six
// outer cards kept in registers.  No global memory traffic except 11 atomics per block.
// ---------------------------------------------------------------------------------------------------
constexpr int ENUM_THREADS = 256;
constexpr int ENUM_CHUNK = 64;

template <bool NIBSUM>
__global__ void __launch_bounds__(ENUM_THREADS)
enum7_kernel(uint64_t first, uint64_t count, unsigned long long *__restrict__ out) {
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t *sbin = smem;                                   // BIN_N*BIN_K words
    uint32_t *T = smem + BIN_N * BIN_K;                      // NIBSUM: 16384 words
    uint4 *cc = reinterpret_cast<uint4 *>(T + (NIBSUM ? NIB_ENTRIES : 0));
    __shared__ unsigned long long sred[11];
    fill_binom(sbin);
    if (NIBSUM) { fill_nib_table(T); if (threadIdx.x < 9) cc[threadIdx.x] = CAT_CONST[threadIdx.x]; }
    if (threadIdx.x < 11) sred[threadIdx.x] = 0ull;
    __syncthreads();

    uint32_t h[9];
#pragma unroll
    for (int i = 0; i < 9; i++) h[i] = 0;
    unsigned long long sumP = 0, sumN = 0;

    const uint64_t nchunks = (count + ENUM_CHUNK - 1) / ENUM_CHUNK;
    const uint64_t nthreads = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t ch = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; ch < nchunks; ch += nthreads) {
        const uint64_t off = ch * ENUM_CHUNK;
        const uint32_t len = (uint32_t)((count - off) < (uint64_t)ENUM_CHUNK ? (count - off) : (uint64_t)ENUM_CHUNK);
        int c[7];
        colex_unrank<7>((uint32_t)(first + off), 52, sbin, c);
        int p0 = id_to_pos(c[0]), p1 = id_to_pos(c[1]), p2 = id_to_pos(c[2]), p3 = id_to_pos(c[3]),
            p4 = id_to_pos(c[4]), p5 = id_to_pos(c[5]), p6 = id_to_pos(c[6]);
        uint64_t U6 = pos_bit(p6), U5 = U6 | pos_bit(p5), U4 = U5 | pos_bit(p4), U3 = U4 | pos_bit(p3),
                 U2 = U3 | pos_bit(p2), U1 = U2 | pos_bit(p1);
        unsigned long long acc = 0;                           // nine 7-bit category counters (len <= 64)
        for (uint32_t j = 0; j < len; j++) {
            const uint64_t m = U1 | pos_bit(p0);
            const uint32_t k = eval_packed(m);
            acc += 1ull << (7 * (k >> 28));
            sumP += k;
            if (NIBSUM) sumN += nibble_from_packed_tbl(k, T, cc);
            // colex successor: lowest card moves fastest
            p0 = pos_succ(p0);
            if (p0 == p1) {
                p0 = 2; p1 = pos_succ(p1);
                if (p1 == p2) {
                    p1 = 3; p2 = pos_succ(p2);
                    if (p2 == p3) {
                        p2 = 4; p3 = pos_succ(p3);
                        if (p3 == p4) {
                            p3 = 5; p4 = pos_succ(p4);
                            if (p4 == p5) {
                                p4 = 6; p5 = pos_succ(p5);
                                if (p5 == p6) { p5 = 7; p6 = pos_succ(p6); U6 = pos_bit(p6 & 63); }
                                U5 = U6 | pos_bit(p5);
                            }
                            U4 = U5 | pos_bit(p4);
                        }
                        U3 = U4 | pos_bit(p3);
                    }
                    U2 = U3 | pos_bit(p2);
                }
                U1 = U2 | pos_bit(p1);
            }
        }
#pragma unroll
        for (int i = 0; i < 9; i++) h[i] += (uint32_t)(acc >> (7 * i)) & 127u;
    }
    // warp REDUX -> shared -> 11 global atomics per block
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int i = 0; i < 9; i++) {
        const uint32_t s = warp_sum(h[i]);
        if (lane == 0 && s) atomicAdd(&sred[i], (unsigned long long)s);
    }
    sumN = warp_sum64(sumN);
    sumP = warp_sum64(sumP);
    if (lane == 0) { atomicAdd(&sred[9], sumN); atomicAdd(&sred[10], sumP); }
    __syncthreads();
    if (threadIdx.x < 11 && sred[threadIdx.x]) atomicAdd(&out[threadIdx.x], sred[threadIdx.x]);
}

cudaError_t launch_enum7(uint64_t first, uint64_t count, unsigned long long *d_out, bool nib_sum,
                         const DeviceInfo &di, cudaStream_t st) {
    if (count == 0) return cudaSuccess;
    const uint64_t nchunks = (count + ENUM_CHUNK - 1) / ENUM_CHUNK;
    uint64_t want = (nchunks + ENUM_THREADS - 1) / ENUM_THREADS;
    const int per_sm = nib_sum ? 3 : 8;
    uint64_t cap = (uint64_t)di.sm_count * per_sm;
    const int grid = (int)(want > cap ? cap : want);
    if (nib_sum) {
        const size_t smem = (BIN_N * BIN_K + NIB_ENTRIES) * 4 + 9 * 16;
        cudaError_t e = cudaFuncSetAttribute(enum7_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        enum7_kernel<true><<<grid, ENUM_THREADS, smem, st>>>(first, count, d_out);
    } else {
        const size_t smem = (BIN_N * BIN_K) * 4 + 16;
        enum7_kernel<false><<<grid, ENUM_THREADS, smem, st>>>(first, count, d_out);
    }
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------
// K3: exhaustive equity.  The K missing community cards run over all K-subsets (colex order of positions
// in the ascending list of live cards).
// ---------------------------------------------------------------------------------------------------
constexpr int EQX_THREADS = 256;
constexpr int EQX_CHUNK = 32;

struct EqxParams {
    EquityArgs a;
    uint64_t first, count;
    int n_live;
    unsigned char live[52];        // bit positions of the live cards, ascending
};

template <int K>
__global__ void __launch_bounds__(EQX_THREADS)
equity_exhaustive_kernel(const __grid_constant__ EqxParams prm, long long *__restrict__ d_counts) {
    __shared__ uint32_t sbin[BIN_N * BIN_K];
    __shared__ unsigned long long sred[19];
    __shared__ unsigned char slive[64];
    fill_binom(sbin);
    if (threadIdx.x < 52) slive[threadIdx.x] = prm.live[threadIdx.x];
    if (threadIdx.x < 19) sred[threadIdx.x] = 0ull;
    __syncthreads();
    const int n_seats = prm.a.n_seats, nl = prm.n_live;
    SeatCounters sc; sc.clear();
    uint32_t units = 0;
    const uint64_t nchunks = (prm.count + EQX_CHUNK - 1) / EQX_CHUNK;
    const uint64_t nthreads = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t ch = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; ch < nchunks; ch += nthreads) {
        const uint64_t off = ch * EQX_CHUNK;
        const uint32_t len = (uint32_t)((prm.count - off) < (uint64_t)EQX_CHUNK ? (prm.count - off) : (uint64_t)EQX_CHUNK);
        int c[K > 0 ? K : 1];
        if (K > 0) colex_unrank<(K > 0 ? K : 1)>((uint32_t)(prm.first + off), nl, sbin, c);
        for (uint32_t j = 0; j < len; j++) {
            uint64_t board = prm.a.board;
#pragma unroll
            for (int i = 0; i < K; i++) board |= pos_bit(slive[c[i]]);
            uint32_t key[MAX_SEATS];
#pragma unroll
            for (int s = 0; s < MAX_SEATS; s++) key[s] = (s < n_seats) ? eval_packed(board | prm.a.hole[s]) : 0u;
            sc.add(key, n_seats);
            units++;
            // colex successor over indices into the live list
            if (K > 0) {
                bool carry = true;
#pragma unroll
                for (int i = 0; i < K; i++) {
                    if (carry) {
                        const int lim = (i + 1 < K) ? c[(i + 1 < K) ? i + 1 : 0] : nl;
                        if (c[i] + 1 < lim) { c[i]++; carry = false; }
                        else c[i] = i;
                    }
                }
            }
        }
    }
    flush_counters(sc, units, n_seats, sred, d_counts);
}

cudaError_t launch_equity_exhaustive(const EquityArgs &a, uint64_t first_board, uint64_t n_boards,
                                     long long *d_counts, const DeviceInfo &di, cudaStream_t st) {
    if (n_boards == 0) return cudaSuccess;
    EqxParams prm;
    prm.a = a; prm.first = first_board; prm.count = n_boards;
    uint64_t used = a.board | a.dead;
    for (int s = 0; s < a.n_seats; s++) used |= a.hole[s];
    int nl = 0;
    for (int b = 0; b < 64; b++) if (((DECK_MASK & ~used) >> b) & 1ull) prm.live[nl++] = (unsigned char)b;
    for (int i = nl; i < 52; i++) prm.live[i] = 0;
    prm.n_live = nl;
    const int K = 5 - __builtin_popcountll(a.board);
    const uint64_t nchunks = (n_boards + EQX_CHUNK - 1) / EQX_CHUNK;
    uint64_t want = (nchunks + EQX_THREADS - 1) / EQX_THREADS, cap = (uint64_t)di.sm_count * 4;
    const int grid = (int)(want > cap ? cap : want);
    switch (K) {
        case 0: equity_exhaustive_kernel<0><<<grid, EQX_THREADS, 0, st>>>(prm, d_counts); break;
        case 1: equity_exhaustive_kernel<1><<<grid, EQX_THREADS, 0, st>>>(prm, d_counts); break;
        case 2: equity_exhaustive_kernel<2><<<grid, EQX_THREADS, 0, st>>>(prm, d_counts); break;
        case 3: equity_exhaustive_kernel<3><<<grid, EQX_THREADS, 0, st>>>(prm, d_counts); break;
        case 4: equity_exhaustive_kernel<4><<<grid, EQX_THREADS, 0, st>>>(prm, d_counts); break;
        case 5: equity_exhaustive_kernel<5><<<grid, EQX_THREADS, 0, st>>>(prm, d_counts); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------
// K4: Monte Carlo equity.  One thread per trial (grid-stride).  The accepted cards of a trial are parked in
// a conflict-free shared-memory column (slot-major, thread-minor) so that every lane can run its own
// rejection loop and the seats are then assembled with warp-uniform slot indices.
// ---------------------------------------------------------------------------------------------------
constexpr int MC_THREADS = 256;
constexpr int MC_MAX_DRAWS = 2 * MAX_SEATS + 5;     // 23

struct McParams {
    EquityArgs a;
    uint64_t seed, first, count;
    uint64_t used0;                // non-card bits | dead | board | fixed holes
    int n_random;                  // seats with hole == 0
    int n_board;                   // community cards to draw
    signed char ridx[MAX_SEATS];   // index of seat s among the random seats, -1 if fixed
};

__global__ void __launch_bounds__(MC_THREADS)
equity_mc_kernel(const __grid_constant__ McParams prm, long long *__restrict__ d_counts) {
    __shared__ uint32_t slots[MC_MAX_DRAWS * MC_THREADS];
    __shared__ unsigned long long sred[19];
    if (threadIdx.x < 19) sred[threadIdx.x] = 0ull;
    __syncthreads();
    const int n_seats = prm.a.n_seats, nr = prm.n_random, need = 2 * nr + prm.n_board;
    const uint32_t k0 = (uint32_t)prm.seed, k1 = (uint32_t)(prm.seed >> 32);
    const uint32_t used0_lo = (uint32_t)prm.used0, used0_hi = (uint32_t)(prm.used0 >> 32);
    uint32_t *my = slots + threadIdx.x;
    SeatCounters sc; sc.clear();
    uint32_t units = 0;
    const uint64_t nthreads = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < prm.count; t += nthreads) {
        const uint64_t trial = prm.first + t;
        uint32_t used_lo = used0_lo, used_hi = used0_hi;
        int cnt = 0;
        uint32_t blk = 0;
        while (cnt < need) {
            uint32_t w[4];
            philox4x32_10((uint32_t)trial, (uint32_t)(trial >> 32), blk, TAG_DEAL, k0, k1, w);
            blk++;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                uint32_t word = w[j];
#pragma unroll
                for (int i = 0; i < 5; i++) {
                    const uint32_t v = word & 63u;
                    word >>= 6;
                    const uint32_t bit = 1u << (v & 31u);
                    const bool up = v >= 32u;
                    const uint32_t cur = up ? used_hi : used_lo;
                    if (!(cur & bit) && cnt < need) {
                        if (up) used_hi |= bit; else used_lo |= bit;
                        my[cnt * MC_THREADS] = v;
                        cnt++;
                    }
                }
            }
        }
        // community cards: slots 2*nr .. need-1
        uint32_t b_lo = (uint32_t)prm.a.board, b_hi = (uint32_t)(prm.a.board >> 32);
#pragma unroll
        for (int i = 0; i < 5; i++) if (i < prm.n_board) {
            const uint32_t v = my[(2 * nr + i) * MC_THREADS];
            const uint32_t bit = 1u << (v & 31u);
            if (v >= 32u) b_hi |= bit; else b_lo |= bit;
        }
        uint32_t key[MAX_SEATS];
#pragma unroll
        for (int s = 0; s < MAX_SEATS; s++) {
            key[s] = 0;
            if (s < n_seats) {
                uint32_t lo = b_lo | (uint32_t)prm.a.hole[s], hi = b_hi | (uint32_t)(prm.a.hole[s] >> 32);
                const int r = prm.ridx[s];
                if (r >= 0) {       // random seat: its cards are accepted draws r and r + nr (deal order, gameplay.clj:63-75)
                    const uint32_t v1 = my[r * MC_THREADS], v2 = my[(r + nr) * MC_THREADS];
                    const uint32_t bit1 = 1u << (v1 & 31u), bit2 = 1u << (v2 & 31u);
                    if (v1 >= 32u) hi |= bit1; else lo |= bit1;
                    if (v2 >= 32u) hi |= bit2; else lo |= bit2;
                }
                key[s] = eval_packed(lo, hi);
            }
        }
        sc.add(key, n_seats);
        units++;
    }
    flush_counters(sc, units, n_seats, sred, d_counts);
}

cudaError_t launch_equity_mc(const EquityArgs &a, uint64_t seed, uint64_t first_trial, uint64_t n_trials,
                             long long *d_counts, const DeviceInfo &di, cudaStream_t st) {
    McParams prm;
    prm.a = a; prm.seed = seed;
    uint64_t used = ~DECK_MASK | a.dead | a.board;
    int nr = 0;
    for (int s = 0; s < MAX_SEATS; s++) {
        prm.ridx[s] = -1;
        if (s < a.n_seats) { used |= a.hole[s]; if (!a.hole[s]) prm.ridx[s] = (signed char)nr++; }
    }
    prm.used0 = used; prm.n_random = nr; prm.n_board = 5 - __builtin_popcountll(a.board);
    const int grid = di.sm_count * 3;
    const uint64_t per_launch = 1ull << 38;                   // keeps the 32-bit per-thread counters safe
    for (uint64_t done = 0; done < n_trials; done += per_launch) {
        prm.first = first_trial + done;
        prm.count = (n_trials - done) < per_launch ? (n_trials - done) : per_launch;
        equity_mc_kernel<<<grid, MC_THREADS, 0, st>>>(prm, d_counts);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// ---------------------------------------------------------------------------------------------------
// single showdown: one warp, lane s = seat s (end-game ranking + group-winners, gameplay.clj:118-132)
// ---------------------------------------------------------------------------------------------------
__global__ void showdown_kernel(const __grid_constant__ EquityArgs a, uint32_t *__restrict__ out) {
    const int lane = threadIdx.x;
    const bool live = lane < a.n_seats;
    const uint64_t m = live ? (a.board | a.hole[lane]) : 0ull;
    const bool ok = live && valid_hand(m);
    const uint32_t pk = ok ? eval_packed(m) : 0u;
    const uint32_t best = __reduce_max_sync(0xFFFFFFFFu, pk);
    const uint32_t winners = __ballot_sync(0xFFFFFFFFu, ok && pk == best);
    // tie group of a seat = number of distinct keys strictly above its own
    const uint32_t same = __match_any_sync(0xFFFFFFFFu, pk);
    const bool leader = live && ((same & ((1u << lane) - 1u)) == 0u);
    int above = 0;
    for (int s = 0; s < a.n_seats; s++) {
        const uint32_t ks = __shfl_sync(0xFFFFFFFFu, pk, s);
        const bool ls = __shfl_sync(0xFFFFFFFFu, (int)leader, s) != 0;
        above += (ls && ks > pk);
    }
    const uint32_t n_groups = __popc(__ballot_sync(0xFFFFFFFFu, leader));
    if (live) {
        out[lane] = ok ? packed_to_nibble(pk) : 0xFFFFFFFFu;
        out[10 + lane] = (uint32_t)above;
    }
    if (lane == 0) { out[9] = winners; out[19] = n_groups; }
}
cudaError_t launch_showdown(const EquityArgs &a, uint32_t *d_out, cudaStream_t st) {
    showdown_kernel<<<1, 32, 0, st>>>(a, d_out);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------
// K5: preflop matrix by brute force over suit-canonical matchups.  A block takes one matchup at a time;
// its threads split the C(48,5) boards into equal contiguous index ranges (one unranking per thread per
// matchup), and the three counts are scaled by the orbit weight into both mirrored cells.
// ---------------------------------------------------------------------------------------------------
constexpr int PF_THREADS = 256;

__global__ void __launch_bounds__(PF_THREADS)
preflop_kernel(const Matchup *__restrict__ mus, int n_mu, long long *__restrict__ wtl) {
    __shared__ uint32_t sbin[BIN_N * BIN_K];
    __shared__ unsigned char slive[64];
    __shared__ unsigned long long sred[3];
    fill_binom(sbin);
    __syncthreads();
    const uint32_t total = cbinom(48, 5);
    const uint32_t per = (total + PF_THREADS - 1) / PF_THREADS;
    for (int mi = blockIdx.x; mi < n_mu; mi += gridDim.x) {
        const Matchup mu = mus[mi];
        if (threadIdx.x < 3) sred[threadIdx.x] = 0ull;
        if (threadIdx.x == 0) {
            int nl = 0;
            const uint64_t livem = DECK_MASK & ~(mu.hole_a | mu.hole_b);
            for (int b = 0; b < 64; b++) if ((livem >> b) & 1ull) slive[nl++] = (unsigned char)b;
        }
        __syncthreads();
        const uint32_t lo = per * threadIdx.x, hi = min(total, lo + per);
        uint32_t wa = 0, wb = 0, tie = 0;
        if (lo < hi) {
            int c[5];
            colex_unrank<5>(lo, 48, sbin, c);
            uint64_t U4 = pos_bit(slive[c[4]]), U3 = U4 | pos_bit(slive[c[3]]), U2 = U3 | pos_bit(slive[c[2]]),
                     U1 = U2 | pos_bit(slive[c[1]]);
            for (uint32_t i = lo; i < hi; i++) {
                const uint64_t board = U1 | pos_bit(slive[c[0]]);
                const uint32_t ka = eval_packed(board | mu.hole_a), kb = eval_packed(board | mu.hole_b);
                wa += ka > kb; wb += kb > ka; tie += ka == kb;
                c[0]++;
                if (c[0] == c[1]) {
                    c[0] = 0; c[1]++;
                    if (c[1] == c[2]) {
                        c[1] = 1; c[2]++;
                        if (c[2] == c[3]) {
                            c[2] = 2; c[3]++;
                            if (c[3] == c[4]) { c[3] = 3; c[4]++; U4 = pos_bit(slive[c[4] < 48 ? c[4] : 0]); }
                            U3 = U4 | pos_bit(slive[c[3]]);
                        }
                        U2 = U3 | pos_bit(slive[c[2]]);
                    }
                    U1 = U2 | pos_bit(slive[c[1]]);
                }
            }
        }
        const uint32_t sa = warp_sum(wa), sb = warp_sum(wb), stie = warp_sum(tie);
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(&sred[0], (unsigned long long)sa); atomicAdd(&sred[1], (unsigned long long)sb);
            atomicAdd(&sred[2], (unsigned long long)stie);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned long long *o = reinterpret_cast<unsigned long long *>(wtl);
            const unsigned long long w = (unsigned long long)mu.weight;
            atomicAdd(&o[3 * mu.cell_ab + 0], w * sred[0]);
            atomicAdd(&o[3 * mu.cell_ab + 1], w * sred[2]);
            atomicAdd(&o[3 * mu.cell_ab + 2], w * sred[1]);
            atomicAdd(&o[3 * mu.cell_ba + 0], w * sred[1]);
            atomicAdd(&o[3 * mu.cell_ba + 1], w * sred[2]);
            atomicAdd(&o[3 * mu.cell_ba + 2], w * sred[0]);
        }
        __syncthreads();
    }
}
cudaError_t launch_preflop(const Matchup *d_matchups, int n_matchups, long long *d_wtl,
                           const DeviceInfo &di, cudaStream_t st) {
    if (n_matchups <= 0) return cudaSuccess;
    int grid = di.sm_count * 8;
    if (grid > n_matchups) grid = n_matchups;
    preflop_kernel<<<grid, PF_THREADS, 0, st>>>(d_matchups, n_matchups, d_wtl);
    return cudaGetLastError();
}

}  // namespace mcp
