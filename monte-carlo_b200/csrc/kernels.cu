// kernels.cu -- hand-written sm_100a kernels for the hand-ranking / showdown / equity path.
//   K1 eval7_stream        batch of card masks -> keys            (player->hand-value, hand_evaluator.clj:162-172)
//   showdown               one table: keys, winners, tie groups   (end-game + group-winners, gameplay.clj:118-132)
//   random_hands           BASELINE config 1 input generator
// (K2-K5, the enumeration / equity kernels, live in kernels_hash.cu)
// Integer / lookup work: no tensor cores (nothing is GEMM-shaped).  Grids are multiples of the SM count.
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

// ---------------------------------------------------------------------------------------------------
// K1: streaming batch evaluation.  8 B in, 4 B out per hand; every LDG.128 / STG.64 of a warp is one
// contiguous 512 B / 256 B segment.  Invalid hands (not 5..7 cards, or a non-card bit) -> 0xFFFFFFFF.
// ---------------------------------------------------------------------------------------------------
constexpr int K1_THREADS = 512;
constexpr int NIB_ENTRIES = 1 << 14;        // rank-set table of eval_tbl.cuh (K1 v2): 64 KB

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
    __shared__ unsigned long long mbar;
    BulkStage stage;
    stage.start(&mbar, sb, g_blob, SB_BYTES);
    stage.wait();
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

}  // namespace mcp
