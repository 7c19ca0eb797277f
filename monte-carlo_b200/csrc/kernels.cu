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
#include <type_traits>

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
// K1 v5: streaming evaluation by hashing the rank-count bit planes (hashtab.cuh "K1 layout").
// The four suit lanes of the mask are added carry-save into per-rank count planes (7 logic ops); the planes ARE
// the rank multiset, so two multiplications of u = c0 | c1 << 16 (plus the rare quads plane) give bucket and slot
// with no table read before the displacement.  A 7-card hand of the 52-card deck costs three data-dependent
// shared loads (D', V', key) plus the flush entry; shifts and address arithmetic run on the FMA pipe (IMAD,
// IMAD.HI) and the suit counts on the XU pipe (POPC), because ALU and the l1tex data pipe are the busy ones.
// Anything else (5/6 cards, the tests' rank 1, invalid masks) is redone by the bit-plane evaluator of eval.cuh
// behind one divergent branch per loop iteration; inputs are software-pipelined one iteration ahead.
// ---------------------------------------------------------------------------------------------------
constexpr int K1H_THREADS = 1024;
constexpr int K1H_RING_DEFAULT = 6;           // measured on a B200: 3: 482, 4: 508, 5: 515, 6: 532, 8: 525 G evals/s


// The value of a hand the image serves (LAY::N_MIN..N_MAX cards of the 52-card deck): its nibble key or, from the
// class flavour of the image, its global class id.  `odd` is non-zero for every other mask (bits outside the deck,
// another card count), whose result is then meaningless but whose loads all stay inside the tables.
// sb32 = shared address of the stream image.
template <class LAY>
__device__ __forceinline__ uint32_t k1h_fast(uint64_t m, uint32_t sb32, const StreamHash &S, uint32_t &odd) {
    const uint32_t lo = (uint32_t)m, hi = (uint32_t)(m >> 32);
    // per-rank card counts over the four suit lanes as bit planes: count = c0 + 2*c1 + 4*c2 (sh_planes)
    const uint32_t x = lo ^ hi, y = lo & hi;                   // per lane pair: sum bit, carry bit
    const uint32_t xs = __umulhi(x, 65536u), ys = __umulhi(y, 65536u);   // >> 16 on the FMA pipe
    const uint32_t c0 = x ^ xs;
    uint32_t c1;
    asm("lop3.b32 %0, %1, %2, %3, 0x6A;" : "=r"(c1) : "r"(x), "r"(xs), "r"(y));   // (x & xs) ^ y in one LOP3
    c1 ^= ys;
    const uint32_t c2 = y & ys;                                // quads plane (ys has no high half)
    const uint32_t u = __byte_perm(c0, c1, 0x5410);            // low halves: c0 | c1 << 16
    const uint32_t t = u * S.m1 + c2 * S.m3;
    const uint32_t b = u * S.m2 + c2 * S.m4;
    // flush lane: only the word holding >= 5 cards can contain it (POPC runs on the XU pipe, cheaper here than selects)
    const int p01 = __popc(lo), p23 = __popc(hi);
    const uint32_t w = p23 >= 5 ? hi : lo;
    const uint32_t wm = w & 0x7FFC7FFCu;                       // rank bits only: the flush address stays inside F for any mask
    const uint32_t wl = wm & 0xFFFFu;
    const int pw = __popc(w), pl = __popc(wl);
    const bool low5 = pl >= 5;
    const uint32_t fm = low5 ? wl : __umulhi(wm, 65536u);
    const bool flush = low5 || (pw - pl >= 5);
    if (LAY::N_MIN == LAY::N_MAX) odd = ((lo | hi) & 0x80038003u) | (uint32_t)(p01 + p23 - LAY::N_MIN);
    else odd = ((lo | hi) & 0x80038003u) | (uint32_t)((uint32_t)(p01 + p23 - LAY::N_MIN) > (uint32_t)(LAY::N_MAX - LAY::N_MIN));
    // D' holds 2 * displacement, so the byte address of the slot is one IMAD: 2 * slot + D'[bucket]
    const uint32_t d2 = lds_u16(sb32 + LAY::D_OFF + __umulhi(b, 8192u) * 2u);
    uint32_t ka = lds_u16(sb32 + LAY::V_OFF + (__umulhi(t, LAY::SLOT_MUL) * 2u + d2));
    if (flush) ka = lds_u16(sb32 + LAY::F_OFF + (fm >> 1));    // F index = flush lane >> 2, two bytes each
    return lds_u32(sb32 + ka);
}

// everything the image does not serve: other card counts, the tests' rank 1, invalid masks.  CLS: the global class id
// (binary search of the nibble key in the ascending class list, global memory) instead of the key; 0xFFFF = invalid
template <bool CLS>
__device__ __noinline__ uint32_t k1h_slow(uint64_t m, const uint32_t *__restrict__ class_nib) {
    const uint32_t lo = (uint32_t)m, hi = (uint32_t)(m >> 32);
    const uint32_t n = (uint32_t)(__popc(lo) + __popc(hi));
    const bool ok = (((lo | hi) & 0x80018001u) == 0u) && (n - 5u <= 2u);
    if (!ok) return CLS ? 0xFFFFu : 0xFFFFFFFFu;
    const uint32_t key = packed_to_nibble(eval_packed(lo, hi));
    if (!CLS) return key;
    int a = 0, b = N_HAND_CLASSES - 1;                         // a rank-1 hand (test input only) has no class: 0xFFFF
    while (a < b) { const int mid = (a + b) >> 1; if (class_nib[mid] < key) a = mid + 1; else b = mid; }
    return class_nib[a] == key ? (uint32_t)a : 0xFFFFu;
}

// one launch covers at most 2^31 hands, so every index fits 32 bits (the launcher cuts longer batches)
// OUT = uint32_t: nibble keys (key flavour of the image);  OUT = uint16_t: global class ids (class flavour)
template <int K1H_RING, class LAY, typename OUT>
__global__ void __launch_bounds__(K1H_THREADS, 1)
eval_stream_hash_kernel(const uint64_t *__restrict__ masks, OUT *__restrict__ keys, uint32_t n,
                        const uint4 *__restrict__ g_blob, const __grid_constant__ StreamHash S,
                        const uint32_t *__restrict__ class_nib) {
    constexpr bool CLS = sizeof(OUT) == 2;
    using OUT2 = typename std::conditional<CLS, uint32_t, uint2>::type;       // two results per store
    extern __shared__ __align__(16) unsigned char sb[];
    __shared__ unsigned long long mbar;
    BulkStage stage;
    stage.start(&mbar, sb, g_blob, LAY::BYTES);
    stage.wait();
    const uint32_t sb32 = staged_table_base(sb);
    const uint32_t npair = n >> 1;
    const ulonglong2 *in = reinterpret_cast<const ulonglong2 *>(masks);
    OUT2 *out = reinterpret_cast<OUT2 *>(keys);
    const uint32_t stride = gridDim.x * blockDim.x;
    // software pipeline: a thread evaluates two hands (16 bytes) per step and keeps the loads of its next RING - 1
    // steps in flight (80 B per thread, 80 KB per SM at RING = 6: HBM bandwidth x latency needs well over 32 KB
    // per SM).  The ring is unrolled so that it stays in registers; prefetch indices are clamped, not predicated.
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < npair) {
        const uint32_t last = npair - 1u;
        constexpr int RING = K1H_RING;                        // RING - 1 steps in flight
        ulonglong2 q[RING];
#pragma unroll
        for (int k = 0; k < RING - 1; k++) q[k] = __ldcs(in + min(i + k * stride, last));
        while (true) {
#pragma unroll
            for (int k = 0; k < RING; k++) {
                q[(k + RING - 1) % RING] = __ldcs(in + min(i + (RING - 1) * stride, last));
                uint32_t o0, o1;
                uint32_t kx = k1h_fast<LAY>(q[k].x, sb32, S, o0);
                uint32_t ky = k1h_fast<LAY>(q[k].y, sb32, S, o1);
                if (o0 | o1) {                                // never for the hands the image serves: one divergent branch
                    if (o0) kx = k1h_slow<CLS>(q[k].x, class_nib);
                    if (o1) ky = k1h_slow<CLS>(q[k].y, class_nib);
                }
                if constexpr (CLS) __stcs(out + i, kx | ky << 16);
                else __stcs(out + i, make_uint2(kx, ky));
                i += stride;
                if (i >= npair) break;
            }
            if (i >= npair) break;
        }
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
        const uint64_t m = masks[n - 1];
        uint32_t o;
        const uint32_t k = k1h_fast<LAY>(m, sb32, S, o);
        keys[n - 1] = (OUT)(o ? k1h_slow<CLS>(m, class_nib) : k);
    }
}

template <class LAY, typename OUT>
static cudaError_t launch_stream_hash_t(const uint64_t *d_masks, size_t n, OUT *d_keys, const StreamImage &im,
                                        const uint32_t *d_class_nib, const DeviceInfo &di, cudaStream_t st) {
    auto kern = eval_stream_hash_kernel<K1H_RING_DEFAULT, LAY, OUT>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LAY::BYTES);
    if (e != cudaSuccess) return e;
    constexpr size_t CHUNK = (size_t)1 << 30;                 // hands per launch: even, so every chunk keeps the alignment
    for (size_t off = 0; off < n; off += CHUNK) {
        const size_t m = n - off < CHUNK ? n - off : CHUNK;
        size_t want = (m + (size_t)K1H_THREADS * 4 - 1) / ((size_t)K1H_THREADS * 4);
        const int grid = (int)(want > (size_t)di.sm_count ? (size_t)di.sm_count : want);
        kern<<<grid, K1H_THREADS, LAY::BYTES, st>>>(d_masks + off, d_keys + off, (uint32_t)m,
                                                    reinterpret_cast<const uint4 *>(sizeof(OUT) == 2 ? im.d_cls : im.d_key), im.hash, d_class_nib);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

cudaError_t launch_eval_stream_hash(const uint64_t *d_masks, size_t n, void *d_out, bool class16, int stage_cards,
                                    const DeviceTables &tb, const DeviceInfo &di, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const bool river = stage_cards != 5 && stage_cards != 6;
    if (river) {
        return class16 ? launch_stream_hash_t<Lay7>(d_masks, n, static_cast<uint16_t *>(d_out), tb.img7, tb.d_class_nib, di, st)
                       : launch_stream_hash_t<Lay7>(d_masks, n, static_cast<uint32_t *>(d_out), tb.img7, tb.d_class_nib, di, st);
    }
    return class16 ? launch_stream_hash_t<Lay56>(d_masks, n, static_cast<uint16_t *>(d_out), tb.img56, tb.d_class_nib, di, st)
                   : launch_stream_hash_t<Lay56>(d_masks, n, static_cast<uint32_t *>(d_out), tb.img56, tb.d_class_nib, di, st);
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
// packed48 -> masks (mc_eval_batch_packed48): a hand is a little-endian 48-bit field of seven 6-bit card positions
// (16 * suit + rank, the bit mc_card_bit sets; 0 = no card, so 5- and 6-card hands fit too).  6 B per hand cross
// the host link instead of 8.  A thread expands four hands: three 8-byte loads, two 16-byte stores.  A repeated
// card or a set bit above bit 41 yields the mask 1 (not a card), which every K1 variant maps to MC_KEY_INVALID.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t expand48(uint64_t h) {
    uint64_t m = 0, dup = h >> 42;
#pragma unroll
    for (int k = 0; k < 7; k++) {
        const uint32_t f = (uint32_t)(h >> (6 * k)) & 63u;
        const uint64_t bit = f ? 1ull << f : 0ull;
        dup |= m & bit;
        m |= bit;
    }
    return dup ? 1ull : m;
}
__global__ void __launch_bounds__(256)
unpack48_kernel(const unsigned char *__restrict__ pk, size_t n, uint64_t *__restrict__ masks) {
    const size_t stride = (size_t)gridDim.x * blockDim.x, groups = n / 4;
    constexpr uint64_t M48 = 0xFFFFFFFFFFFFull;
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += stride) {
        const uint64_t *p = reinterpret_cast<const uint64_t *>(pk + 24 * g);
        const uint64_t a = __ldcs(p), b = __ldcs(p + 1), c = __ldcs(p + 2);
        ulonglong2 lo, hi;
        lo.x = expand48(a & M48);
        lo.y = expand48((a >> 48 | b << 16) & M48);
        hi.x = expand48((b >> 32 | c << 32) & M48);
        hi.y = expand48(c >> 16);
        ulonglong2 *o = reinterpret_cast<ulonglong2 *>(masks + 4 * g);
        o[0] = lo; o[1] = hi;
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {                  // the last n % 4 hands, byte by byte
        const size_t i = 4 * groups + threadIdx.x;
        uint64_t h = 0;
        for (int k = 0; k < 6; k++) h |= (uint64_t)pk[6 * i + k] << (8 * k);
        masks[i] = expand48(h);
    }
}
// d_packed 8-byte aligned, d_masks 16-byte aligned
cudaError_t launch_unpack48(const unsigned char *d_packed, size_t n, uint64_t *d_masks, const DeviceInfo &di, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    size_t want = (n / 4 + 255) / 256 + 1, cap = (size_t)di.sm_count * 8;
    unpack48_kernel<<<(int)(want > cap ? cap : want), 256, 0, st>>>(d_packed, n, d_masks);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------
// colex32 -> masks (mc_eval_batch_colex32): a 7-card hand travels as its colex rank over card ids, sum_i C(c_i, i+1)
// < C(52,7) < 2^27 -- 4 B per hand over the host link instead of 8.  Unranking = seven branch-free binary searches
// (six steps each) in a shared table of binomials padded with 0xFFFFFFFF: the largest x with C(x, k) <= rest is
// automatically below the previous card, so no bound is carried.  A rank >= C(52,7) yields the mask 1 (not a card).
// ---------------------------------------------------------------------------------------------------
// colex27: the same ranks as a little-endian bit stream, 27 bits per hand (C(52,7) < 2^27): hand j is bits [27 j, 27 j + 27),
// so 32 hands travel in 27 words -- 3.375 B per hand.  The stream is read with two 4-byte loads and a funnel shift.
constexpr int UNR_COLS = 72;
template <bool PACKED27>
__global__ void __launch_bounds__(256)
unrank_colex_kernel(const uint32_t *__restrict__ ranks, size_t n, uint64_t *__restrict__ masks) {
    __shared__ uint32_t sb[8 * UNR_COLS];                         // sb[k * 72 + x] = C(x, k) for x <= 52, else 2^32 - 1
    for (int i = threadIdx.x; i < 8 * UNR_COLS; i += blockDim.x) {
        const int k = i / UNR_COLS, x = i % UNR_COLS;
        unsigned long long r = (k <= x) ? 1ull : 0ull;
        for (int j = 1; j <= k && k <= x; j++) r = r * (unsigned long long)(x - k + j) / (unsigned long long)j;
        sb[i] = x <= 52 ? (uint32_t)r : 0xFFFFFFFFu;
    }
    __syncthreads();
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint32_t rest;
        if (PACKED27) {
            const size_t bit = 27 * i, w = bit >> 5;
            const uint32_t lo = __ldg(ranks + w), hi = ((bit & 31) > 5) ? __ldg(ranks + w + 1) : 0u;   // (the next word only when the field crosses into it)
            rest = __funnelshift_r(lo, hi, (uint32_t)(bit & 31)) & 0x7FFFFFFu;
        } else {
            rest = __ldcs(ranks + i);
        }
        uint64_t m = 1ull;
        if (rest < cbinom(52, 7)) {
            m = 0ull;
#pragma unroll
            for (int k = 7; k >= 1; k--) {
                const uint32_t *row = sb + k * UNR_COLS;
                uint32_t x = (uint32_t)k - 1u;                    // C(k-1, k) = 0 <= rest
#pragma unroll
                for (uint32_t step = 32; step >= 1; step >>= 1) { const uint32_t y = x + step; if (row[y] <= rest) x = y; }
                rest -= row[x];
                m |= 1ull << id_to_pos((int)x);
            }
        }
        masks[i] = m;
    }
}
cudaError_t launch_unrank_colex32(const uint32_t *d_ranks, size_t n, uint64_t *d_masks, const DeviceInfo &di, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    size_t want = (n + 255) / 256, cap = (size_t)di.sm_count * 8;
    unrank_colex_kernel<false><<<(int)(want > cap ? cap : want), 256, 0, st>>>(d_ranks, n, d_masks);
    return cudaGetLastError();
}
cudaError_t launch_unrank_colex27(const uint32_t *d_stream, size_t n, uint64_t *d_masks, const DeviceInfo &di, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    size_t want = (n + 255) / 256, cap = (size_t)di.sm_count * 8;
    unrank_colex_kernel<true><<<(int)(want > cap ? cap : want), 256, 0, st>>>(d_stream, n, d_masks);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------
// Scratch of the allocation-free entry point mc_eval_batch_colex32_dev_noalloc (capi.cu): the card masks the colex
// ranks are unranked into before K1 reads them.  A file-scope array instead of a cudaMalloc'ed stage, so the call can be
// made where allocation is forbidden (CUDA graph capture, serving loops).  8 MiB; every GPU that loads the library
// holds it until the process exits.  Every call on a device shares it: calls that overlap on the GPU race.  Nothing in
// it has to start at zero -- the unranking kernel overwrites the first n entries, which are the only ones K1 reads.
// In host code `g_noalloc_masks` is only a stand-in for the symbol, not a device address: the address comes from
// cudaGetSymbolAddress (below, called once per device at mc_init with that device current).
// ---------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned long long g_noalloc_masks[NOALLOC_MAX_HANDS];

cudaError_t noalloc_scratch_address(uint64_t **d_ptr) {
    void *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_masks);
    if (e != cudaSuccess) return e;
    *d_ptr = static_cast<uint64_t *>(p);
    return cudaSuccess;
}

// ---------------------------------------------------------------------------------------------------
// single showdown: one warp, lane s = seat s (end-game ranking + group-winners, gameplay.clj:118-132)
// ---------------------------------------------------------------------------------------------------
__global__ void showdown_kernel(const __grid_constant__ EquityArgs a, uint32_t *__restrict__ out, uint32_t ticket) {
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
    // `out` may be mapped host memory that the host polls: results first, then the ticket
    __threadfence_system();
    __syncwarp();
    if (lane == 0) *reinterpret_cast<volatile uint32_t *>(out + SMALL_TICKET_WORD) = ticket;
}
cudaError_t launch_showdown(const EquityArgs &a, uint32_t *d_out, uint32_t ticket, cudaStream_t st) {
    showdown_kernel<<<1, 32, 0, st>>>(a, d_out, ticket);
    return cudaGetLastError();
}

// one hand, mask by kernel parameter (evaluator / player->hand-value for a single call of the host)
__global__ void eval_one_kernel(uint64_t m, uint32_t *__restrict__ out, uint32_t ticket) {
    out[0] = k1_key<K1_NIBBLE_FLO>(m, nullptr, nullptr);
    __threadfence_system();
    *reinterpret_cast<volatile uint32_t *>(out + SMALL_TICKET_WORD) = ticket;
}
cudaError_t launch_eval_one(uint64_t mask, uint32_t *d_out, uint32_t ticket, cudaStream_t st) {
    eval_one_kernel<<<1, 1, 0, st>>>(mask, d_out, ticket);
    return cudaGetLastError();
}

}  // namespace mcp
