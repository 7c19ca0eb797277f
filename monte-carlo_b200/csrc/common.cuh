// common.cuh -- device helpers shared by kernels.cu and kernels_hash.cu: binomial table + colex unranking,
// warp reductions, the per-seat showdown counters and their block-level flush into 64-bit global counts.
#pragma once
#include <cstdint>
#include <cstdio>
#include "kernels.cuh"

namespace mcp {

// ---------------------------------------------------------------------------------------------------
// small helpers
// ---------------------------------------------------------------------------------------------------
__host__ __device__ constexpr uint32_t cbinom(int n, int k) {
    if (k < 0 || k > n) return 0;
    unsigned long long r = 1;
    for (int i = 1; i <= k; i++) r = r * (unsigned long long)(n - k + i) / (unsigned long long)i;
    return (uint32_t)r;
}

constexpr int BIN_N = 53, BIN_K = 8;

// C(n, k) for n < 53, k < 8, computed at compile time: the kernels copy it into shared memory (1.7 KB, one coalesced load
// per thread) instead of computing it with 64-bit divisions in every block's prologue (~2 us of a 5 us launch floor)
struct BinomTable {
    uint32_t v[BIN_N * BIN_K];
    constexpr BinomTable() : v() {
        for (int n = 0; n < BIN_N; n++) for (int k = 0; k < BIN_K; k++) v[n * BIN_K + k] = cbinom(n, k);
    }
};
static __device__ const BinomTable g_binom_table = BinomTable();
// shared-memory copy of the binomial table used by colex unranking
__device__ __forceinline__ void fill_binom(uint32_t *sbin) {
    for (int i = threadIdx.x; i < BIN_N * BIN_K; i += blockDim.x) sbin[i] = g_binom_table.v[i];
}
// colex unrank: idx = sum_i C(c_i, i+1), c_0 < c_1 < ... < c_{K-1} < n.  Per level a branch-free binary search (six
// steps) for the largest x with C(x, i) <= idx: idx < C(n, K) makes it < n, and what is left after a level is < C(x, i - 1),
// so the next card is below the previous one by itself -- no bound is carried.
template <int K>
__device__ __forceinline__ void colex_unrank(uint32_t idx, int n, const uint32_t *sbin, int (&c)[K]) {
    (void)n;
#pragma unroll
    for (int i = K; i >= 1; i--) {
        uint32_t x = (uint32_t)i - 1u;                        // C(i - 1, i) = 0 <= idx
#pragma unroll
        for (uint32_t step = 32; step >= 1; step >>= 1) {
            const uint32_t y = min(x + step, (uint32_t)(BIN_N - 1));
            if (sbin[y * BIN_K + i] <= idx) x = y;
        }
        c[i - 1] = (int)x;
        idx -= sbin[x * BIN_K + i];
    }
}

// ---------------------------------------------------------------------------------------------------
// Table staging by the TMA engine: one elected thread arms an mbarrier with the byte count and issues 1-D bulk
// copies global -> shared (cp.async.bulk, SASS UBLKCP); the other threads keep filling the small per-block tables
// and everybody waits on the barrier's phase 0.  bytes % 16 == 0, dst and src 16-byte aligned.
// ---------------------------------------------------------------------------------------------------
struct BulkStage {
    uint32_t bar;
    __device__ __forceinline__ void start(unsigned long long *mbar, void *smem_dst, const void *gsrc, uint32_t bytes) {
        bar = (uint32_t)__cvta_generic_to_shared(mbar);
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");          // the async proxy sees the initialised barrier
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(smem_dst);
            const char *src = static_cast<const char *>(gsrc);
            constexpr uint32_t CHUNK = 32768;
            for (uint32_t off = 0; off < bytes; off += CHUNK) {
                const uint32_t n = bytes - off < CHUNK ? bytes - off : CHUNK;
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(dst + off), "l"(src + off), "r"(n), "r"(bar) : "memory");
            }
        }
    }
    // every thread of the block; includes the block-wide barrier that publishes the other threads' table fills
    __device__ __forceinline__ void wait() {
        __syncthreads();                      // the mbarrier is initialised and armed before anybody polls it
        uint32_t done = 0;
        while (!done)
            asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}"
                         : "=r"(done) : "r"(bar) : "memory");
    }
};

// Loads from a staged (read-only) shared-memory table by 32-bit shared address.  Plain asm, so the compiler may
// schedule them freely; u16 entries arrive zero-extended in a full register (a C++ uint16_t load is kept as a
// 16-bit value and merged / masked with extra PRMT and LOP3 around every select or compare).
// MCP_CHECK build (make check -> libmcpoker_check.so; compute-sanitizer is closed on the GPU pool): every table load
// verifies that its bytes lie inside the block's shared-memory allocation and are naturally aligned, and traps otherwise.
#ifdef MCP_CHECK
__device__ __forceinline__ void lds_check(uint32_t addr, uint32_t bytes) {
    uint32_t total;
    asm("mov.u32 %0, %%total_smem_size;" : "=r"(total));
    if (addr + bytes > total || (addr & (bytes - 1u))) {
        printf("MCP_CHECK: shared load of %u bytes at %u outside the %u-byte allocation or misaligned (block %d thread %d)\n",
               bytes, addr, total, (int)blockIdx.x, (int)threadIdx.x);
        __trap();
    }
}
#else
__device__ __forceinline__ void lds_check(uint32_t, uint32_t) {}
#endif
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) { lds_check(addr, 4); uint32_t v; asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr)); return v; }
__device__ __forceinline__ uint2 lds_u32x2(uint32_t addr) { lds_check(addr, 8); uint2 v; asm("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr)); return v; }
__device__ __forceinline__ uint32_t lds_u16(uint32_t addr) { lds_check(addr, 2); uint32_t v; asm("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(addr)); return v; }
// shared address of a table, to be taken AFTER BulkStage::wait(): every lds_* through it depends on this volatile
// asm, so none can move above the wait, and the compiler keeps the base in a register instead of rebuilding it
__device__ __forceinline__ uint32_t staged_table_base(const void *smem_table) {
    uint32_t a = (uint32_t)__cvta_generic_to_shared(smem_table);
    asm volatile("" : "+r"(a) : : "memory");
    return a;
}

__device__ __forceinline__ uint32_t warp_sum(uint32_t v) { return __reduce_add_sync(0xFFFFFFFFu, v); }
__device__ __forceinline__ unsigned long long warp_sum64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    return v;
}

// ---------------------------------------------------------------------------------------------------
// showdown of up to nine evaluated seats held in registers (gameplay.clj:127-132): win if sole maximum,
// tie if the maximum is shared, else loss.  Packed keys order exactly like clojure.core/compare.
// ---------------------------------------------------------------------------------------------------
struct SeatCounters {
    uint32_t win[MAX_SEATS], tie[MAX_SEATS];
    __device__ __forceinline__ void clear() {
#pragma unroll
        for (int s = 0; s < MAX_SEATS; s++) { win[s] = 0; tie[s] = 0; }
    }
    __device__ __forceinline__ void add(const uint32_t (&key)[MAX_SEATS], int n_seats) {
        uint32_t best = 0;
#pragma unroll
        for (int s = 0; s < MAX_SEATS; s++) if (s < n_seats) best = max(best, key[s]);
        int nbest = 0;
#pragma unroll
        for (int s = 0; s < MAX_SEATS; s++) if (s < n_seats) nbest += (key[s] == best);
        const bool solo = nbest == 1;
#pragma unroll
        for (int s = 0; s < MAX_SEATS; s++) if (s < n_seats) {
            const bool b = key[s] == best;
            win[s] += (b && solo);
            tie[s] += (b && !solo);
        }
    }
};

// block-level reduction of the per-thread counters into d_counts = win[9] tie[9] loss[9] (loss = n - win - tie)
__device__ __forceinline__ void flush_counters(const SeatCounters &sc, uint32_t my_units, int n_seats,
                                               unsigned long long *sred /*19*/, long long *d_counts) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int s = 0; s < MAX_SEATS; s++) if (s < n_seats) {
        const uint32_t w = warp_sum(sc.win[s]), t = warp_sum(sc.tie[s]);
        if (lane == 0) { if (w) atomicAdd(&sred[s], (unsigned long long)w); if (t) atomicAdd(&sred[9 + s], (unsigned long long)t); }
    }
    const uint32_t u = warp_sum(my_units);
    if (lane == 0) atomicAdd(&sred[18], (unsigned long long)u);
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

}  // namespace mcp
