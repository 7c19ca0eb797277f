// capi.cu -- the extern "C" boundary of libmcpoker.so (include/mcpoker.h).  Host-side plumbing only:
// argument checks, device contexts, staging pipelines, range sharding over the devices of one process and
// the final sum of the per-device int64 counts.  All hand evaluation happens in kernels.cu; there is no CPU
// evaluation path in this library.
#include "../../include/mcpoker.h"
#include "eval.cuh"
#include "kernels.cuh"
#include "philox.cuh"
#include "hashtab_build.h"

#include <nccl.h>          // types only: the library is bound at run time (dlopen), see NcclApi
#include <dlfcn.h>

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <thread>
#include <vector>

using namespace mcp;

namespace {

thread_local std::string g_err;
int fail(int code, const std::string &msg) { g_err = msg; return code; }
int cuda_fail(cudaError_t e, const char *what) {
    return fail(MC_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
#define CU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return cuda_fail(e_, #x); } while (0)

constexpr int N_SLOTS = 3;
constexpr size_t SLOT_HANDS = 1u << 22;           // 4 Mi hands per pipeline stage: 32 MiB in, 16 MiB out
constexpr int SMALL_WORDS = 64;
static_assert(SMALL_TICKET_WORD < 2 * SMALL_WORDS && SMALL_TICKET_WORD >= 2 * 32, "ticket word inside the scratch, above every counts copy");

struct DevCtx {
    int device = -1;
    DeviceInfo info{};
    std::mutex mu;
    cudaStream_t stream = nullptr;                // small jobs
    cudaStream_t slot_stream[N_SLOTS] = {};
    uint64_t *d_in[N_SLOTS] = {};
    uint32_t *d_out[N_SLOTS] = {};
    unsigned char *d_pk[N_SLOTS] = {};            // packed48 input stages (6 B per hand), allocated on first use
    unsigned long long *d_small = nullptr;        // SMALL_WORDS x u64 scratch
    unsigned long long *h_small = nullptr;        // pinned mirror (mapped)
    unsigned long long *hd_small = nullptr;       // the device's address of h_small
    Matchup *d_matchups = nullptr; size_t matchup_cap = 0;
    RangeEntry *d_range = nullptr;                // alias tables of mc_equity_mc_ranges: MC_MAX_SEATS x 1326 entries
    uint64_t *d_noalloc_masks = nullptr;          // this device's address of the static scratch of the allocation-free call (not owned)
    long long *d_wtl = nullptr;
    PfBoard *d_pb_boards = nullptr; size_t pb_board_cap = 0;            // config 5 by boards: board list, scratch, win matrix
    uint32_t *d_pb_scratch = nullptr;
    unsigned long long *d_pb_win = nullptr;
    DeviceTables tb;                              // perfect-hash evaluator tables (hashtab.cuh), built at init
    void *d_tables = nullptr;                     // one allocation: blob | card keys | nibble keys

    ~DevCtx() {
        if (device < 0) return;
        cudaSetDevice(device);
        for (int i = 0; i < N_SLOTS; i++) {
            if (d_in[i]) cudaFree(d_in[i]);
            if (d_out[i]) cudaFree(d_out[i]);
            if (d_pk[i]) cudaFree(d_pk[i]);
            if (slot_stream[i]) cudaStreamDestroy(slot_stream[i]);
        }
        if (d_small) cudaFree(d_small);
        if (h_small) cudaFreeHost(h_small);
        if (d_matchups) cudaFree(d_matchups);
        if (d_range) cudaFree(d_range);
        if (d_wtl) cudaFree(d_wtl);
        if (d_pb_boards) cudaFree(d_pb_boards);
        if (d_pb_scratch) cudaFree(d_pb_scratch);
        if (d_pb_win) cudaFree(d_pb_win);
        if (d_tables) cudaFree(d_tables);
        if (stream) cudaStreamDestroy(stream);
    }
};

// Lifetime of the device contexts.  Every entry point that touches g_ctx holds g_life SHARED for its whole duration
// (struct Use below); mc_init* / mc_shutdown hold it EXCLUSIVE while they swap or destroy the contexts, so they wait for
// the calls in flight and no call can start on a context that is being torn down.  Work on one device is serialised by
// DevCtx::mu as before.  (An entry point takes the shared hold exactly once: nested public calls go through the
// innermost function that takes it.)
std::shared_mutex g_life;
std::vector<std::unique_ptr<DevCtx>> g_ctx;

// The one collective of the path: the int64 counts of the devices of this library instance are summed with a
// single ncclAllReduce per device inside one group, over NVLink.  Single process, ncclCommInitAll.  NCCL is bound with
// dlopen when mc_init is asked for it (MC_INIT_NCCL), so the library has no link-time dependency on it and a host that
// drives one device per process (bench.py under torchrun) never loads it.  Without the flag the counts are summed on the
// host from pinned copies: for 27 words that is the faster way (tools/collective_ab.py).
struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    std::vector<ncclComm_t> comm;                   // one per context, same order as g_ctx
    bool active() const { return !comm.empty(); }
    std::string load() {
        if (lib) return "";
        lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!lib) return std::string("libnccl.so.2 could not be loaded: ") + dlerror();
        auto sym = [&](const char *n) { return dlsym(lib, n); };
        CommInitAll = reinterpret_cast<decltype(CommInitAll)>(sym("ncclCommInitAll"));
        CommDestroy = reinterpret_cast<decltype(CommDestroy)>(sym("ncclCommDestroy"));
        AllReduce = reinterpret_cast<decltype(AllReduce)>(sym("ncclAllReduce"));
        GroupStart = reinterpret_cast<decltype(GroupStart)>(sym("ncclGroupStart"));
        GroupEnd = reinterpret_cast<decltype(GroupEnd)>(sym("ncclGroupEnd"));
        GetErrorString = reinterpret_cast<decltype(GetErrorString)>(sym("ncclGetErrorString"));
        if (!CommInitAll || !CommDestroy || !AllReduce || !GroupStart || !GroupEnd || !GetErrorString) return "libnccl.so.2 lacks a required symbol";
        return "";
    }
    void destroy() {
        for (ncclComm_t c : comm) if (c) CommDestroy(c);
        comm.clear();
    }
} g_nccl;
std::atomic<uint32_t> g_ticket_seq{1};              // tickets of the single-call kernels: process-wide, never 0, so a
                                                    // re-initialised context can never match a word left in pinned memory
uint32_t next_ticket() { uint32_t t; do t = g_ticket_seq.fetch_add(1, std::memory_order_relaxed); while (t == 0); return t; }

int make_info(int dev, DeviceInfo *out) {
    int major = 0, sms = 0;
    CU(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (major != 10) return fail(MC_ERR_NO_DEVICE, "device " + std::to_string(dev) + " is not sm_100 (B200); this library has no other code path");
    out->device = dev; out->sm_count = sms;
    return MC_OK;
}

// Hash tables of hashtab.cuh.  The representative hands are evaluated on the device with the bit-plane
// evaluator (K1, packed keys); the host only lays the results out (hashtab_build.h).
std::unique_ptr<HashTables> g_tables;
std::unique_ptr<StreamImages> g_images;

// packed keys of `masks` by the bit-plane kernel
int eval_reps_on_device(DevCtx &c, const std::vector<uint64_t> &masks, std::vector<uint32_t> *packed) {
    const size_t n = masks.size();
    packed->resize(n);
    uint64_t *d_m = nullptr; uint32_t *d_k = nullptr;
    CU(cudaMalloc(&d_m, n * sizeof(uint64_t)));
    cudaError_t e = cudaMalloc(&d_k, n * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_m, masks.data(), n * sizeof(uint64_t), cudaMemcpyHostToDevice, c.stream);
    if (e == cudaSuccess) e = launch_eval7_stream(d_m, n, d_k, K1_PACKED, c.info, c.stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(packed->data(), d_k, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c.stream);
    cudaFree(d_m); cudaFree(d_k);
    if (e != cudaSuccess) return cuda_fail(e, "evaluating the table representatives");
    return MC_OK;
}

int build_tables(DevCtx &c) {
    std::vector<HashRep> reps;
    ht_representatives(HT_MUL_CANDIDATES[0], &reps);
    std::vector<uint64_t> masks(reps.size());
    std::vector<uint32_t> packed;
    for (size_t i = 0; i < reps.size(); i++) masks[i] = reps[i].mask;
    int rc = eval_reps_on_device(c, masks, &packed);
    if (rc) return rc;
    std::unique_ptr<HashTables> T(new HashTables());
    std::string err = "no multiplier tried";
    for (uint32_t mul : HT_MUL_CANDIDATES) {
        ht_representatives(mul, &reps);
        err = ht_build(reps, packed, mul, T.get());
        if (err.empty()) break;
    }
    if (!err.empty()) return fail(MC_ERR_CUDA, "hash table construction failed: " + err);
    // the streaming kernel's images (river, flop / turn; key and class flavours) and the global class list
    std::vector<StreamRep> sreps;
    ht_stream_representatives(&sreps);
    masks.resize(sreps.size());
    for (size_t i = 0; i < sreps.size(); i++) masks[i] = sreps[i].mask;
    rc = eval_reps_on_device(c, masks, &packed);
    if (rc) return rc;
    std::unique_ptr<StreamImages> I(new StreamImages());
    err = ht_build_stream_images(sreps, packed, I.get());
    if (!err.empty()) return fail(MC_ERR_CUDA, "stream image construction failed: " + err);
    g_tables = std::move(T);
    g_images = std::move(I);
    return MC_OK;
}

int upload_tables(DevCtx &c) {
    const HashTables &T = *g_tables;
    const StreamImages &I = *g_images;
    auto up = [](size_t x) { return (x + 255) & ~(size_t)255; };
    const size_t off_card = sizeof(HashBlob), off_nib = off_card + sizeof(T.card);
    const size_t off_cls = up(off_nib + T.nib_keys.size() * sizeof(uint32_t));
    const size_t off_k7 = up(off_cls + I.class_nib.size() * sizeof(uint32_t)), off_c7 = off_k7 + up(Lay7::BYTES);
    const size_t off_k56 = off_c7 + up(Lay7::BYTES), off_c56 = off_k56 + up(Lay56::BYTES);
    const size_t total = off_c56 + up(Lay56::BYTES);
    CU(cudaMalloc(&c.d_tables, total));
    unsigned char *base = static_cast<unsigned char *>(c.d_tables);
    CU(cudaMemcpyAsync(base, &T.blob, sizeof(HashBlob), cudaMemcpyHostToDevice, c.stream));
    CU(cudaMemcpyAsync(base + off_card, T.card, sizeof(T.card), cudaMemcpyHostToDevice, c.stream));
    CU(cudaMemcpyAsync(base + off_nib, T.nib_keys.data(), T.nib_keys.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, c.stream));
    CU(cudaMemcpyAsync(base + off_cls, I.class_nib.data(), I.class_nib.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, c.stream));
    CU(cudaMemcpyAsync(base + off_k7, I.key7.data(), Lay7::BYTES, cudaMemcpyHostToDevice, c.stream));
    CU(cudaMemcpyAsync(base + off_c7, I.cls7.data(), Lay7::BYTES, cudaMemcpyHostToDevice, c.stream));
    CU(cudaMemcpyAsync(base + off_k56, I.key56.data(), Lay56::BYTES, cudaMemcpyHostToDevice, c.stream));
    CU(cudaMemcpyAsync(base + off_c56, I.cls56.data(), Lay56::BYTES, cudaMemcpyHostToDevice, c.stream));
    CU(cudaStreamSynchronize(c.stream));
    c.tb.d_blob = reinterpret_cast<const HashBlob *>(base);
    c.tb.d_card = reinterpret_cast<const CardKey *>(base + off_card);
    c.tb.d_nib = reinterpret_cast<const uint32_t *>(base + off_nib);
    c.tb.d_class_nib = reinterpret_cast<const uint32_t *>(base + off_cls);
    c.tb.img7 = StreamImage{base + off_k7, base + off_c7, I.h7};
    c.tb.img56 = StreamImage{base + off_k56, base + off_c56, I.h56};
    c.tb.h_card = T.card;
    for (int i = 0; i < 10; i++) c.tb.cat_base.v[i] = T.cat_base[i];
    c.tb.n_classes = (uint32_t)T.nib_keys.size();
    return MC_OK;
}

int init_devices(const int *ids, int n, unsigned flags) {
    std::unique_lock<std::shared_mutex> lk(g_life);          // waits for every call in flight
    int visible = 0;
    cudaError_t e = cudaGetDeviceCount(&visible);
    if (e != cudaSuccess || visible == 0) { cudaGetLastError(); return fail(MC_ERR_NO_DEVICE, "no CUDA device visible; libmcpoker has no CPU fallback"); }
    std::vector<int> want;
    if (ids) want.assign(ids, ids + n);
    else { if (n <= 0 || n > visible) n = visible; for (int i = 0; i < n; i++) want.push_back(i); }
    for (int d : want) if (d < 0 || d >= visible) return fail(MC_ERR_INVALID, "device id out of range");
    for (size_t i = 0; i < want.size(); i++) for (size_t j = i + 1; j < want.size(); j++)
        if (want[i] == want[j]) return fail(MC_ERR_INVALID, "a device id is listed twice");
    // the new contexts are built aside and swapped in only when every device succeeded: a failure half way leaves the
    // library exactly as it was
    std::vector<std::unique_ptr<DevCtx>> fresh;
    for (int d : want) {
        std::unique_ptr<DevCtx> c(new DevCtx());
        int rc = make_info(d, &c->info);
        if (rc) return rc;
        CU(cudaSetDevice(d));
        c->device = d;
        CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
        CU(cudaMalloc(&c->d_small, SMALL_WORDS * sizeof(unsigned long long)));
        CU(cudaHostAlloc(&c->h_small, SMALL_WORDS * sizeof(unsigned long long), cudaHostAllocMapped | cudaHostAllocPortable));
        std::memset(c->h_small, 0, SMALL_WORDS * sizeof(unsigned long long));      // the ticket word must not hold a stale ticket
        CU(cudaHostGetDevicePointer(reinterpret_cast<void **>(&c->hd_small), c->h_small, 0));
        if (!g_tables) { rc = build_tables(*c); if (rc) return rc; }
        rc = upload_tables(*c);
        if (rc) return rc;
        CU(noalloc_scratch_address(&c->d_noalloc_masks));     // device d is current: the symbol's address on THIS device
        fresh.push_back(std::move(c));
    }
    // the collective, if asked for and if there is anything to reduce over
    std::vector<ncclComm_t> comms;
    if ((flags & MC_INIT_NCCL) && fresh.size() > 1) {
        const std::string err = g_nccl.load();
        if (!err.empty()) return fail(MC_ERR_CUDA, "MC_INIT_NCCL: " + err);
        comms.assign(fresh.size(), nullptr);
        const ncclResult_t r = g_nccl.CommInitAll(comms.data(), (int)want.size(), want.data());
        if (r != ncclSuccess) return fail(MC_ERR_CUDA, std::string("ncclCommInitAll: ") + g_nccl.GetErrorString(r));
    }
    g_nccl.destroy();                                         // communicators of the contexts that go away
    g_nccl.comm = comms;
    g_ctx.swap(fresh);                                        // the old contexts are destroyed with `fresh`
    return MC_OK;
}

// Shared hold on the device contexts for the duration of one entry point.
struct Use {
    std::shared_lock<std::shared_mutex> lk{g_life};
    int rc;
    Use() { rc = g_ctx.empty() ? fail(MC_ERR_NOT_INIT, "mc_init has not been called") : MC_OK; }
};

// If a call returns early (a CUDA error half way), no asynchronous copy into the caller's or the call's own host
// buffers may stay in flight: drain the streams of every context on the way out.  Free when nothing is pending.
struct DrainOnExit {
    bool armed = true;
    ~DrainOnExit() {
        if (!armed) return;
        for (auto &c : g_ctx)
            if (cudaSetDevice(c->device) == cudaSuccess) {
                (void)cudaStreamSynchronize(c->stream);
                if (c->d_in[0]) for (int s = 0; s < N_SLOTS; s++) (void)cudaStreamSynchronize(c->slot_stream[s]);
            }
        (void)cudaGetLastError();
    }
};

// Wait for a single-call kernel (launch_showdown / launch_eval_one writing into the mapped h_small): poll the ticket
// word it writes last instead of paying the stream synchronisation; a kernel that never delivers (launch failure,
// sticky error) is caught by the bounded spin falling back to cudaStreamSynchronize, which reports the error.
int await_ticket(DevCtx &c, uint32_t ticket) {
    volatile uint32_t *flag = reinterpret_cast<volatile uint32_t *>(c.h_small) + SMALL_TICKET_WORD;
    for (int spin = 0; spin < 200000; spin++) {
        if (*flag == ticket) { std::atomic_thread_fence(std::memory_order_acquire); return MC_OK; }
        __builtin_ia32_pause();
    }
    CU(cudaStreamSynchronize(c.stream));
    if (*flag != ticket) return fail(MC_ERR_CUDA, "single-call kernel finished without delivering its result");
    return MC_OK;
}

// context of the calling thread's current device (for the *_dev entry points)
int current_info(DeviceInfo *out, const DeviceTables **tb = nullptr) {
    int dev = 0;
    CU(cudaGetDevice(&dev));
    for (auto &c : g_ctx) if (c->device == dev) { *out = c->info; if (tb) *tb = &c->tb; return MC_OK; }
    return make_info(dev, out);
}

// tables of the calling thread's current device (the hash kernels need mc_init to have run for it)
int current_tables(DeviceInfo *di, const DeviceTables **tb) {
    int dev = 0;
    CU(cudaGetDevice(&dev));
    for (auto &c : g_ctx) if (c->device == dev) { *di = c->info; *tb = &c->tb; return MC_OK; }
    return fail(MC_ERR_NOT_INIT, "mc_init / mc_init_devices has not been called for the current device");
}

// K1 dispatch.  key_format: 0 nibble (the plane-hash kernel for aligned buffers, else the rank-set-table kernel),
// 1 packed, 2 nibble via FLO loops, 3 nibble via the rank-set table, 4 nibble via the plane hash (A/B variants),
// 5 = MC_KEY_CLASS16 (uint16 global class ids; plane-hash kernel only).  stage_cards 5 / 6 selects the flop / turn image.
// The plane-hash kernel serves every aligned batch: measured (tools/k1_crossover.py) it wins at every size (6 us floor:
// 184 KB staged by TMA; the rank-set-table kernel computes its 64 KB table in every block, 12 us)
cudaError_t launch_k1(const uint64_t *d_masks, size_t n, void *d_out, int key_format, int stage_cards, const DeviceInfo &di,
                      const DeviceTables *tb, cudaStream_t st) {
    const bool cls = key_format == MC_KEY_CLASS16;
    const bool aligned = ((reinterpret_cast<uintptr_t>(d_masks) & 15) == 0) && ((reinterpret_cast<uintptr_t>(d_out) & (cls ? 3 : 7)) == 0);
    const bool can_hash = tb && tb->img7.d_key && aligned;
    if (cls || key_format == 4 || (key_format == 0 && can_hash)) {
        if (!can_hash) return cudaErrorInvalidValue;
        return launch_eval_stream_hash(d_masks, n, d_out, cls, stage_cards, *tb, di, st);
    }
    const int mode = key_format == MC_KEY_PACKED ? K1_PACKED : (key_format == 2 ? K1_NIBBLE_FLO : K1_NIBBLE_TABLE);
    return launch_eval7_stream(d_masks, n, static_cast<uint32_t *>(d_out), mode, di, st);
}

int ensure_slots(DevCtx &c) {
    if (c.d_in[0]) return MC_OK;
    for (int i = 0; i < N_SLOTS; i++) {
        CU(cudaStreamCreateWithFlags(&c.slot_stream[i], cudaStreamNonBlocking));
        CU(cudaMalloc(&c.d_in[i], SLOT_HANDS * sizeof(uint64_t)));
        CU(cudaMalloc(&c.d_out[i], SLOT_HANDS * sizeof(uint32_t)));
    }
    return MC_OK;
}

int ensure_packed_slots(DevCtx &c) {
    if (c.d_pk[0]) return MC_OK;
    for (int i = 0; i < N_SLOTS; i++) CU(cudaMalloc(&c.d_pk[i], SLOT_HANDS * 6));
    return MC_OK;
}

int check_equity_args(const uint64_t *hole_masks, int n_seats, uint64_t board, uint64_t dead, bool allow_random,
                      EquityArgs *out, int *need_cards) {
    if (!hole_masks) return fail(MC_ERR_INVALID, "hole_masks is null");
    if (n_seats < 1 || n_seats > MC_MAX_SEATS) return fail(MC_ERR_INVALID, "n_seats must be 1..9");
    if ((board & ~DECK_MASK) || (dead & ~DECK_MASK)) return fail(MC_ERR_INVALID, "board/dead mask has a bit that is not one of the 52 cards");
    if (board & dead) return fail(MC_ERR_INVALID, "board and dead overlap");
    const int nb = __builtin_popcountll(board);
    if (nb > 5) return fail(MC_ERR_INVALID, "more than 5 community cards");
    uint64_t used = board | dead;
    int need = 5 - nb;
    std::memset(out, 0, sizeof *out);
    for (int s = 0; s < n_seats; s++) {
        const uint64_t h = hole_masks[s];
        if (h == 0) {
            if (!allow_random) return fail(MC_ERR_INVALID, "a seat has no hole cards");
            need += 2;
        } else {
            if ((h & ~DECK_MASK) || __builtin_popcountll(h) != 2) return fail(MC_ERR_INVALID, "a hole mask must hold exactly two of the 52 cards");
            if (h & used) return fail(MC_ERR_INVALID, "a card is used twice");
            used |= h;
        }
        out->hole[s] = h;
    }
    if (52 - __builtin_popcountll(used) < need) return fail(MC_ERR_INVALID, "not enough live cards");
    out->board = board; out->dead = dead; out->n_seats = n_seats;
    if (need_cards) *need_cards = need;
    return MC_OK;
}

// pots of the EV calls -> PotArgs (kernels.cuh)
int check_pots(const int64_t *pot_bet, const int32_t *pot_n, const uint32_t *pot_players, int n_pots, int n_seats, PotArgs *out) {
    if (n_pots < 1 || n_pots > MC_MAX_POTS) return fail(MC_ERR_INVALID, "n_pots must be 1..MC_MAX_POTS");
    if (!pot_bet || !pot_n || !pot_players) return fail(MC_ERR_INVALID, "null pot array");
    *out = PotArgs{};
    out->n_pots = n_pots;
    int64_t sum = 0;
    for (int p = 0; p < n_pots; p++) {
        if (pot_bet[p] < 0 || pot_n[p] < 0) return fail(MC_ERR_INVALID, "negative pot");
        const unsigned __int128 tot = (unsigned __int128)(uint64_t)pot_bet[p] * (uint64_t)pot_n[p];
        if (tot > 0x7FFFFFFFull) return fail(MC_ERR_INVALID, "the pots of a simulation must hold fewer than 2^31 chips in total");
        sum += (int64_t)tot;
        if (sum > 0x7FFFFFFFll) return fail(MC_ERR_INVALID, "the pots of a simulation must hold fewer than 2^31 chips in total");
        if (pot_players[p] == 0 || (pot_players[p] >> n_seats)) return fail(MC_ERR_INVALID, "a pot's players must be a non-empty set of the seats (the reference divides by zero on an empty one, gameplay.clj:112)");
        out->players[p] = pot_players[p];
        out->total[p] = (uint32_t)tot;
    }
    return MC_OK;
}

// run fn(ctx, lo, hi) on every device over an even split of [0, total), then sum `words` u64 of d_small
template <typename F>
int sharded_small_job(uint64_t total, int words, unsigned long long *sum, F launch) {
    Use use;
    if (use.rc) return use.rc;
    int rc;
    const int nd = (int)g_ctx.size();
    std::vector<std::unique_lock<std::mutex>> locks;
    for (auto &c : g_ctx) locks.emplace_back(c->mu);
    DrainOnExit drain;
    const uint64_t per = (total + nd - 1) / nd;
    for (int d = 0; d < nd; d++) {
        DevCtx &c = *g_ctx[d];
        const uint64_t lo = std::min<uint64_t>(total, per * d), hi = std::min<uint64_t>(total, lo + per);
        CU(cudaSetDevice(c.device));
        CU(cudaMemsetAsync(c.d_small, 0, SMALL_WORDS * sizeof(unsigned long long), c.stream));
        if (hi > lo) { rc = launch(c, lo, hi - lo); if (rc) return rc; }
        if (!g_nccl.active()) CU(cudaMemcpyAsync(c.h_small, c.d_small, words * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c.stream));
    }
    if (g_nccl.active()) {
        // one all-reduce per device inside one group, each on the stream its kernel ran on; device 0 delivers the sums
        ncclResult_t r = g_nccl.GroupStart();
        for (int d = 0; d < nd && r == ncclSuccess; d++)
            r = g_nccl.AllReduce(g_ctx[d]->d_small, g_ctx[d]->d_small, (size_t)words, ncclUint64, ncclSum, g_nccl.comm[d], g_ctx[d]->stream);
        const ncclResult_t e = g_nccl.GroupEnd();
        if (r == ncclSuccess) r = e;
        if (r != ncclSuccess) return fail(MC_ERR_CUDA, std::string("ncclAllReduce: ") + g_nccl.GetErrorString(r));
        CU(cudaSetDevice(g_ctx[0]->device));
        CU(cudaMemcpyAsync(g_ctx[0]->h_small, g_ctx[0]->d_small, words * sizeof(unsigned long long), cudaMemcpyDeviceToHost, g_ctx[0]->stream));
    }
    for (int i = 0; i < words; i++) sum[i] = 0;
    for (int d = 0; d < nd; d++) {
        DevCtx &c = *g_ctx[d];
        CU(cudaSetDevice(c.device));
        CU(cudaStreamSynchronize(c.stream));
        if (d == 0 || !g_nccl.active()) for (int i = 0; i < words; i++) sum[i] += c.h_small[i];
    }
    return MC_OK;
}

}  // namespace

extern "C" {

int mc_abi_version(void) { return 2; }
const char *mc_last_error(void) { return g_err.c_str(); }

int mc_init(int n_gpus, unsigned flags) { return init_devices(nullptr, n_gpus, flags); }
int mc_init_devices(const int *device_ids, int n) {
    if (!device_ids || n <= 0) return fail(MC_ERR_INVALID, "device_ids is empty");
    return init_devices(device_ids, n, 0);
}
int mc_collective(void) {
    std::shared_lock<std::shared_mutex> lk(g_life);
    return g_nccl.active() ? 1 : 0;
}
void mc_shutdown(void) {
    std::unique_lock<std::shared_mutex> lk(g_life);          // waits for every call in flight
    g_nccl.destroy();
    g_ctx.clear();
    g_tables.reset();
    g_images.reset();
}
int mc_device_count(void) {
    std::shared_lock<std::shared_mutex> lk(g_life);
    return (int)g_ctx.size();
}

uint64_t mc_card_bit(int card_id) {
    if (card_id < 0 || card_id >= 52) return 0;
    return 1ull << (16 * (card_id / 13) + card_id % 13 + 2);
}
uint32_t mc_key_packed_to_nibble(uint32_t k) { return k == MC_KEY_INVALID ? k : packed_to_nibble(k); }
int mc_key_decode(uint32_t key, int32_t *cat, int32_t hit[5], int32_t *n_hit, int32_t kick[5], int32_t *n_kick) {
    static const int HIT_LEN[9] = {5, 2, 4, 3, 5, 5, 5, 4, 5};     // SURVEY 7 / hand_evaluator.clj:100-133
    const int c = (int)(key >> 20);
    if (key == MC_KEY_INVALID || c > 8) return fail(MC_ERR_INVALID, "not a hand key");
    int nh = 0, nk = 0;
    for (int i = 0; i < 5; i++) {
        const int r = (int)((key >> (16 - 4 * i)) & 15u);
        if (i < HIT_LEN[c]) hit[nh++] = r; else kick[nk++] = r;
    }
    *cat = c; *n_hit = nh; *n_kick = nk;
    return MC_OK;
}

// ---- K1 --------------------------------------------------------------------------------------------
namespace {
bool key_format_ok(int f) { return f >= 0 && f <= MC_KEY_CLASS16; }
}
int mc_eval_batch_dev(const uint64_t *d_masks, size_t n, void *d_out, int key_format, int stage_cards, void *stream) {
    if (n && (!d_masks || !d_out)) return fail(MC_ERR_INVALID, "null device buffer");
    if (!key_format_ok(key_format)) return fail(MC_ERR_INVALID, "unknown key format");
    std::shared_lock<std::shared_mutex> life(g_life);
    DeviceInfo di; const DeviceTables *tb = nullptr;
    int rc = current_info(&di, &tb);
    if (rc) return rc;
    if (key_format == MC_KEY_CLASS16 && !tb) return fail(MC_ERR_NOT_INIT, "mc_init / mc_init_devices has not been called for the current device");
    CU(launch_k1(d_masks, n, d_out, key_format, stage_cards, di, tb, (cudaStream_t)stream));
    return MC_OK;
}
int mc_eval7_batch_dev_fmt(const uint64_t *d_masks, size_t n, uint32_t *d_keys, int key_format, void *stream) {
    if (key_format == MC_KEY_CLASS16) return fail(MC_ERR_INVALID, "class ids are 16-bit: use mc_eval_batch_dev");
    return mc_eval_batch_dev(d_masks, n, d_keys, key_format, 7, stream);
}
int mc_eval7_batch_dev(const uint64_t *d_masks, size_t n, uint32_t *d_keys, void *stream) {
    return mc_eval_batch_dev(d_masks, n, d_keys, MC_KEY_NIBBLE, 7, stream);
}
int mc_unrank_colex32_dev(const uint32_t *d_ranks, size_t n, uint64_t *d_masks, void *stream) {
    if (n && (!d_ranks || !d_masks)) return fail(MC_ERR_INVALID, "null device buffer");
    std::shared_lock<std::shared_mutex> life(g_life);
    DeviceInfo di;
    int rc = current_info(&di);
    if (rc) return rc;
    CU(launch_unrank_colex32(d_ranks, n, d_masks, di, (cudaStream_t)stream));
    return MC_OK;
}
// The allocation-free variant of the colex32 batch evaluation: ranks and class ids in the caller's device memory, the
// intermediate masks in the library's static device array.  No cudaMalloc / cudaFree / allocating routine on this path.
int mc_eval_batch_colex32_dev_noalloc(const uint32_t *d_ranks, size_t n, uint16_t *d_classes, void *stream) {
    if (n > NOALLOC_MAX_HANDS)
        return fail(MC_ERR_INVALID, "mc_eval_batch_colex32_dev_noalloc: n = " + std::to_string(n) + " exceeds MC_NOALLOC_MAX_HANDS = " +
                                    std::to_string(NOALLOC_MAX_HANDS) + " (the size of the static scratch); split the batch");
    if (n && (!d_ranks || !d_classes)) return fail(MC_ERR_INVALID, "null device buffer");
    if (reinterpret_cast<uintptr_t>(d_classes) & 3) return fail(MC_ERR_INVALID, "d_classes must be 4-byte aligned");
    std::shared_lock<std::shared_mutex> life(g_life);
    int dev = 0;
    CU(cudaGetDevice(&dev));
    DevCtx *c = nullptr;
    for (auto &x : g_ctx) if (x->device == dev) c = x.get();
    if (!c) return fail(MC_ERR_NOT_INIT, "mc_init / mc_init_devices has not been called for the current device");
    if (n == 0) return MC_OK;
    CU(launch_unrank_colex32(d_ranks, n, c->d_noalloc_masks, c->info, (cudaStream_t)stream));
    CU(launch_eval_stream_hash(c->d_noalloc_masks, n, d_classes, true, 7, c->tb, c->info, (cudaStream_t)stream));
    return MC_OK;
}
int mc_class_keys(uint32_t *keys, int n) {
    if (!keys || n != N_HAND_CLASSES) return fail(MC_ERR_INVALID, "keys must hold MC_N_HAND_CLASSES entries");
    Use use;
    if (use.rc) return use.rc;
    std::memcpy(keys, g_images->class_nib.data(), (size_t)n * sizeof(uint32_t));
    return MC_OK;
}

namespace {
size_t stage_hands() {        // hands per pipeline stage of the host-buffer calls (<= SLOT_HANDS, the staging buffers' size)
    static const size_t v = [] {
        const char *e = getenv("MCP_STAGE_LOG2");
        const int l = e ? atoi(e) : 0;
        return (l >= 12 && l <= 22) ? (size_t)1 << l : SLOT_HANDS;
    }();
    return v;
}
// bytes of the first `hands` hands of a batch in the wire format (colex27: whole 32-bit words of the bit stream; the
// pipeline cuts batches at multiples of 32 hands, where the stream is word-aligned)
size_t wire_bytes(int in_format, size_t hands) {
    if (in_format == MC_IN_COLEX27) return 4 * ((27 * hands + 31) / 32);
    return hands * (in_format == MC_IN_MASK64 ? 8 : in_format == MC_IN_PACKED48 ? 6 : 4);
}
// how many cards the FIRST hand of a host batch holds (the host-buffer calls pick the K1 image by it; hands of another
// size are still evaluated exactly, on the kernel's slow path)
int first_hand_cards(const unsigned char *src, int in_format) {
    if (in_format == MC_IN_COLEX32 || in_format == MC_IN_COLEX27) return 7;
    uint64_t h = 0;
    std::memcpy(&h, src, in_format == MC_IN_MASK64 ? 8 : 6);
    if (in_format == MC_IN_MASK64) return __builtin_popcountll(h);
    int n = 0;
    for (int k = 0; k < 7; k++) n += ((h >> (6 * k)) & 63u) != 0;
    return n;
}
// host buffers -> keys / classes.  The input is masks (8 B per hand), packed48 hands (6 B) or colex32 ranks (4 B), the
// compact formats being expanded to masks on the device; the output nibble keys (4 B) or class ids (2 B).  Contiguous
// slice per device, each streamed through N_SLOTS staging buffers on their own streams so that the H2D copy of stage
// i+1, the kernel(s) of stage i and the D2H copy of stage i-1 overlap.
int eval_batch_pipeline(const unsigned char *src, int in_format, size_t n, unsigned char *dst, int key_format, int stage_cards) {
    int rc;
    const int nd = (int)g_ctx.size();
    const size_t out_bytes = key_format == MC_KEY_CLASS16 ? 2 : 4;
    // slices start at multiples of 32 hands: there the 27-bit stream is word-aligned (and every format keeps the alignment
    // of the two-hand loads / stores)
    const size_t per_dev = ((n + nd - 1) / nd + 31) & ~(size_t)31;
    std::vector<size_t> pos(nd), end(nd);
    for (int d = 0; d < nd; d++) {
        pos[d] = std::min(n, per_dev * d); end[d] = std::min(n, pos[d] + per_dev);
        if (end[d] > pos[d]) {
            CU(cudaSetDevice(g_ctx[d]->device));
            rc = ensure_slots(*g_ctx[d]);
            if (!rc && in_format != MC_IN_MASK64) rc = ensure_packed_slots(*g_ctx[d]);
            if (rc) return rc;
        }
    }
    bool more = true;
    for (size_t round = 0; more; round++) {
        more = false;
        for (int d = 0; d < nd; d++) {
            if (pos[d] >= end[d]) continue;
            DevCtx &c = *g_ctx[d];
            const int s = (int)(round % N_SLOTS);
            const size_t cnt = std::min(stage_hands(), end[d] - pos[d]);
            CU(cudaSetDevice(c.device));
            const size_t in_off = wire_bytes(in_format, pos[d]), in_len = wire_bytes(in_format, pos[d] + cnt) - in_off;
            if (in_format == MC_IN_MASK64) {
                CU(cudaMemcpyAsync(c.d_in[s], src + in_off, in_len, cudaMemcpyHostToDevice, c.slot_stream[s]));
            } else {
                CU(cudaMemcpyAsync(c.d_pk[s], src + in_off, in_len, cudaMemcpyHostToDevice, c.slot_stream[s]));
                if (in_format == MC_IN_PACKED48) CU(launch_unpack48(c.d_pk[s], cnt, c.d_in[s], c.info, c.slot_stream[s]));
                else if (in_format == MC_IN_COLEX32) CU(launch_unrank_colex32(reinterpret_cast<const uint32_t *>(c.d_pk[s]), cnt, c.d_in[s], c.info, c.slot_stream[s]));
                else CU(launch_unrank_colex27(reinterpret_cast<const uint32_t *>(c.d_pk[s]), cnt, c.d_in[s], c.info, c.slot_stream[s]));
            }
            CU(launch_k1(c.d_in[s], cnt, c.d_out[s], key_format, stage_cards, c.info, &c.tb, c.slot_stream[s]));
            CU(cudaMemcpyAsync(dst + pos[d] * out_bytes, c.d_out[s], cnt * out_bytes, cudaMemcpyDeviceToHost, c.slot_stream[s]));
            pos[d] += cnt;
            if (pos[d] < end[d]) more = true;
        }
    }
    for (int d = 0; d < nd; d++) {
        DevCtx &c = *g_ctx[d];
        if (!c.d_in[0]) continue;
        CU(cudaSetDevice(c.device));
        for (int s = 0; s < N_SLOTS; s++) CU(cudaStreamSynchronize(c.slot_stream[s]));
    }
    return MC_OK;
}
int eval_batch_host(const void *in, int in_format, size_t n, void *out, int key_format, int stage_cards) {
    if (in_format < MC_IN_MASK64 || in_format > MC_IN_COLEX27) return fail(MC_ERR_INVALID, "unknown input format");
    if (!key_format_ok(key_format)) return fail(MC_ERR_INVALID, "unknown key format");
    Use use;
    if (use.rc) return use.rc;
    if (n == 0) return MC_OK;
    if (!in || !out) return fail(MC_ERR_INVALID, "null buffer");
    const unsigned char *src = static_cast<const unsigned char *>(in);
    if (stage_cards == 0) stage_cards = first_hand_cards(src, in_format);
    std::vector<std::unique_lock<std::mutex>> locks;
    for (auto &c : g_ctx) locks.emplace_back(c->mu);
    DrainOnExit drain;          // a failed call must not leave copies into the caller's buffers in flight
    return eval_batch_pipeline(src, in_format, n, static_cast<unsigned char *>(out), key_format, stage_cards);
}
}  // namespace

int mc_eval_batch(const void *hands, int in_format, size_t n, void *out, int key_format, int stage_cards) {
    return eval_batch_host(hands, in_format, n, out, key_format, stage_cards);
}
int mc_eval7_batch_fmt(const uint64_t *masks, size_t n, uint32_t *keys, int key_format) {
    if (key_format == MC_KEY_CLASS16) return fail(MC_ERR_INVALID, "class ids are 16-bit: use mc_eval_batch");
    return eval_batch_host(masks, MC_IN_MASK64, n, keys, key_format, 0);
}
int mc_eval_batch_packed48(const uint8_t *hands, size_t n, uint32_t *keys) {
    return eval_batch_host(hands, MC_IN_PACKED48, n, keys, MC_KEY_NIBBLE, 0);
}
int mc_eval_batch_colex32(const uint32_t *ranks, size_t n, uint16_t *classes) {
    return eval_batch_host(ranks, MC_IN_COLEX32, n, classes, MC_KEY_CLASS16, 7);
}
int mc_eval_batch_colex27(const uint32_t *stream, size_t n, uint16_t *classes) {
    return eval_batch_host(stream, MC_IN_COLEX27, n, classes, MC_KEY_CLASS16, 7);
}
void *mc_alloc_host(size_t bytes) {
    Use use;
    if (use.rc) return nullptr;
    void *p = nullptr;
    if (cudaSetDevice(g_ctx[0]->device) != cudaSuccess || cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) {
        fail(MC_ERR_CUDA, "cudaHostAlloc failed");
        (void)cudaGetLastError();
        return nullptr;
    }
    return p;
}
int mc_free_host(void *p) {
    if (!p) return MC_OK;
    CU(cudaFreeHost(p));
    return MC_OK;
}
int mc_eval7_batch(const uint64_t *masks, size_t n, uint32_t *keys) {
    return mc_eval7_batch_fmt(masks, n, keys, MC_KEY_NIBBLE);
}

int mc_eval_cards(const int32_t *suit_id, const int32_t *rank, int n_cards, uint32_t *key) {
    Use use;
    if (use.rc) return use.rc;
    int rc;
    if (!suit_id || !rank || !key) return fail(MC_ERR_INVALID, "null argument");
    if (n_cards < 5) return fail(MC_ERR_TOO_FEW, "fewer than 5 cards: the reference raises ArityException here (hand_evaluator.clj:171-172)");
    if (n_cards > 7) return fail(MC_ERR_INVALID, "more than 7 cards");
    uint64_t m = 0;
    for (int i = 0; i < n_cards; i++) {
        if (suit_id[i] < 0 || suit_id[i] > 3) return fail(MC_ERR_INVALID, "suit id must be 0..3");
        if (rank[i] < 1 || rank[i] > 14) return fail(MC_ERR_INVALID, "rank must be 1..14");
        const uint64_t b = 1ull << (16 * suit_id[i] + rank[i]);
        if (m & b) return fail(MC_ERR_INVALID, "duplicate card");
        m |= b;
    }
    DevCtx &c = *g_ctx[0];
    std::lock_guard<std::mutex> lk(c.mu);
    CU(cudaSetDevice(c.device));
    // the mask travels as a kernel parameter and the key is written straight into mapped pinned host memory:
    // one launch and one stream sync, no copies
    const uint32_t ticket = next_ticket();
    CU(launch_eval_one(m, reinterpret_cast<uint32_t *>(c.hd_small), ticket, c.stream));
    rc = await_ticket(c, ticket);
    if (rc) return rc;
    *key = reinterpret_cast<const uint32_t *>(c.h_small)[0];
    return MC_OK;
}

namespace {
// one table through the showdown kernel: keys, tie groups, winner bitmask (the caller holds the shared hold)
int showdown_locked(uint64_t board_mask, const uint64_t *hole_masks, int n_seats, uint32_t *keys, int32_t *group_of,
                    int32_t *n_groups, uint32_t *winner_bitmask) {
    EquityArgs a;
    int rc = check_equity_args(hole_masks, n_seats, board_mask, 0, false, &a, nullptr);
    if (rc) return rc;
    if (__builtin_popcountll(board_mask) < 3) return fail(MC_ERR_TOO_FEW, "fewer than 5 cards per player: the reference raises ArityException here");
    DevCtx &c = *g_ctx[0];
    std::lock_guard<std::mutex> lk(c.mu);
    CU(cudaSetDevice(c.device));
    const uint32_t ticket = next_ticket();
    CU(launch_showdown(a, reinterpret_cast<uint32_t *>(c.hd_small), ticket, c.stream));   // seats by parameter, result into mapped host memory
    rc = await_ticket(c, ticket);
    if (rc) return rc;
    const uint32_t *h = reinterpret_cast<const uint32_t *>(c.h_small);
    for (int s = 0; s < n_seats; s++) { if (keys) keys[s] = h[s]; if (group_of) group_of[s] = (int32_t)h[10 + s]; }
    if (n_groups) *n_groups = (int32_t)h[19];
    if (winner_bitmask) *winner_bitmask = h[9];
    return MC_OK;
}
int settle_pots(const int32_t *group_of, int n_seats, const int64_t *pot_bet, const int32_t *pot_n,
                const uint32_t *pot_players, int n_pots, int64_t *delta) {
    if (!group_of || !delta || n_seats < 1 || n_seats > MC_MAX_SEATS || n_pots < 0) return fail(MC_ERR_INVALID, "bad argument");
    if (n_pots && (!pot_bet || !pot_n || !pot_players)) return fail(MC_ERR_INVALID, "null pot array");
    for (int p = 0; p < n_pots; p++) {
        int best_group = -1;                                     // lowest group index among the pot's players
        for (int s = 0; s < n_seats; s++)
            if ((pot_players[p] >> s) & 1u) { if (group_of[s] < 0) continue; if (best_group < 0 || group_of[s] < best_group) best_group = group_of[s]; }
        if (best_group < 0) return fail(MC_ERR_INVALID, "a pot has no player at the showdown (the reference divides by zero here, gameplay.clj:112)");
        int members = 0;
        for (int s = 0; s < n_seats; s++) if (((pot_players[p] >> s) & 1u) && group_of[s] == best_group) members++;
        const int64_t share = (pot_bet[p] * (int64_t)pot_n[p]) / members;          // int(/ total-pot (count pot-winners))
        for (int s = 0; s < n_seats; s++) if (((pot_players[p] >> s) & 1u) && group_of[s] == best_group) delta[s] += share;
    }
    return MC_OK;
}
}  // namespace

int mc_showdown_groups(uint64_t board_mask, const uint64_t *hole_masks, int n_seats,
                       uint32_t *keys, int32_t *group_of, int32_t *n_groups, uint32_t *winner_bitmask) {
    Use use;
    if (use.rc) return use.rc;
    return showdown_locked(board_mask, hole_masks, n_seats, keys, group_of, n_groups, winner_bitmask);
}
int mc_showdown(uint64_t board_mask, const uint64_t *hole_masks, int n_seats, uint32_t *keys, uint32_t *winner_bitmask) {
    Use use;
    if (use.rc) return use.rc;
    return showdown_locked(board_mask, hole_masks, n_seats, keys, nullptr, nullptr, winner_bitmask);
}
// end-game (gameplay.clj:122-133) in one call: ranking, tie groups and the settlement of the pots
int mc_end_game(uint64_t board_mask, const uint64_t *hole_masks, int n_seats, const int64_t *pot_bet, const int32_t *pot_n,
                const uint32_t *pot_players, int n_pots, uint32_t *keys, int32_t *group_of, int32_t *n_groups,
                uint32_t *winner_bitmask, int64_t *delta) {
    if (!delta) return fail(MC_ERR_INVALID, "delta is null");
    Use use;
    if (use.rc) return use.rc;
    int32_t groups[MC_MAX_SEATS];
    int rc = showdown_locked(board_mask, hole_masks, n_seats, keys, groups, n_groups, winner_bitmask);
    if (rc) return rc;
    if (group_of) for (int s = 0; s < n_seats; s++) group_of[s] = groups[s];
    return settle_pots(groups, n_seats, pot_bet, pot_n, pot_players, n_pots, delta);
}

int mc_settle_pots(const int32_t *group_of, int n_seats, const int64_t *pot_bet, const int32_t *pot_n,
                   const uint32_t *pot_players, int n_pots, int64_t *delta) {
    return settle_pots(group_of, n_seats, pot_bet, pot_n, pot_players, n_pots, delta);
}

// ---- synthetic hands ------------------------------------------------------------------------------
int mc_random_hands_dev(uint64_t seed, uint64_t first, size_t n, int n_cards, uint64_t *d_masks, void *stream) {
    if (n_cards < 1 || n_cards > 7) return fail(MC_ERR_INVALID, "n_cards must be 1..7");
    if (n && !d_masks) return fail(MC_ERR_INVALID, "null device buffer");
    std::shared_lock<std::shared_mutex> life(g_life);
    DeviceInfo di;
    int rc = current_info(&di);
    if (rc) return rc;
    CU(launch_random_hands(seed, first, n, n_cards, d_masks, di, (cudaStream_t)stream));
    return MC_OK;
}
int mc_random_hands(uint64_t seed, uint64_t first, size_t n, int n_cards, uint64_t *masks) {
    Use use;
    if (use.rc) return use.rc;
    int rc;
    if (n_cards < 1 || n_cards > 7) return fail(MC_ERR_INVALID, "n_cards must be 1..7");
    if (n && !masks) return fail(MC_ERR_INVALID, "null buffer");
    DevCtx &c = *g_ctx[0];
    std::lock_guard<std::mutex> lk(c.mu);
    CU(cudaSetDevice(c.device));
    rc = ensure_slots(c);
    if (rc) return rc;
    for (size_t pos = 0; pos < n; pos += SLOT_HANDS) {
        const size_t cnt = std::min(SLOT_HANDS, n - pos);
        CU(launch_random_hands(seed, first + pos, cnt, n_cards, c.d_in[0], c.info, c.slot_stream[0]));
        CU(cudaMemcpyAsync(masks + pos, c.d_in[0], cnt * sizeof(uint64_t), cudaMemcpyDeviceToHost, c.slot_stream[0]));
        CU(cudaStreamSynchronize(c.slot_stream[0]));
    }
    return MC_OK;
}

// ---- K2 ---------------------------------------------------------------------------------------------
int mc_enumerate7_dev(uint64_t first, uint64_t count, uint64_t *d_out, void *stream) {
    const uint64_t total = host_binom(52, 7);
    if (first > total || count > total - first) return fail(MC_ERR_INVALID, "combination range out of bounds");
    if (!d_out) return fail(MC_ERR_INVALID, "null device buffer");
    std::shared_lock<std::shared_mutex> life(g_life);
    DeviceInfo di; const DeviceTables *tb;
    int rc = current_tables(&di, &tb);
    if (rc) return rc;
    CU(launch_enum7_hash(first, count, reinterpret_cast<unsigned long long *>(d_out), *tb, di, (cudaStream_t)stream));
    return MC_OK;
}
int mc_enumerate7_range(uint64_t first, uint64_t count, uint64_t hist[9], uint64_t *key_sum) {
    const uint64_t total = host_binom(52, 7);
    if (first > total || count > total - first) return fail(MC_ERR_INVALID, "combination range out of bounds");
    if (!hist) return fail(MC_ERR_INVALID, "hist is null");
    unsigned long long sum[10];
    int rc = sharded_small_job(count, 10, sum, [&](DevCtx &c, uint64_t lo, uint64_t cnt) -> int {
        CU(launch_enum7_hash(first + lo, cnt, c.d_small, c.tb, c.info, c.stream));
        return MC_OK;
    });
    if (rc) return rc;
    for (int i = 0; i < 9; i++) hist[i] = sum[i];
    if (key_sum) *key_sum = sum[9];
    return MC_OK;
}
int mc_enumerate7(uint64_t hist[9], uint64_t *key_sum) { return mc_enumerate7_range(0, host_binom(52, 7), hist, key_sum); }

// ---- K3 ---------------------------------------------------------------------------------------------
static uint64_t board_count(const EquityArgs &a) {
    uint64_t used = a.board | a.dead;
    for (int s = 0; s < a.n_seats; s++) used |= a.hole[s];
    return host_binom(52 - __builtin_popcountll(used), 5 - __builtin_popcountll(a.board));
}
int mc_equity_exhaustive_dev(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                             uint64_t first_board, uint64_t n_boards_in, int64_t *d_counts, void *stream) {
    EquityArgs a;
    int rc = check_equity_args(hole_masks, n_seats, board_mask, dead_mask, false, &a, nullptr);
    if (rc) return rc;
    const uint64_t total = board_count(a);
    if (first_board > total || n_boards_in > total - first_board) return fail(MC_ERR_INVALID, "board range out of bounds");
    if (!d_counts) return fail(MC_ERR_INVALID, "null device buffer");
    std::shared_lock<std::shared_mutex> life(g_life);
    DeviceInfo di; const DeviceTables *tb;
    rc = current_tables(&di, &tb);
    if (rc) return rc;
    CU(launch_equity_exhaustive_hash(a, first_board, n_boards_in, reinterpret_cast<long long *>(d_counts), *tb, di, (cudaStream_t)stream));
    return MC_OK;
}
int mc_equity_exhaustive_range(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                               uint64_t first_board, uint64_t n_boards_in, int64_t *win, int64_t *tie, int64_t *loss) {
    EquityArgs a;
    int rc = check_equity_args(hole_masks, n_seats, board_mask, dead_mask, false, &a, nullptr);
    if (rc) return rc;
    const uint64_t total = board_count(a);
    if (first_board > total || n_boards_in > total - first_board) return fail(MC_ERR_INVALID, "board range out of bounds");
    unsigned long long sum[27];
    rc = sharded_small_job(n_boards_in, 27, sum, [&](DevCtx &c, uint64_t lo, uint64_t cnt) -> int {
        CU(launch_equity_exhaustive_hash(a, first_board + lo, cnt, reinterpret_cast<long long *>(c.d_small), c.tb, c.info, c.stream));
        return MC_OK;
    });
    if (rc) return rc;
    for (int s = 0; s < n_seats; s++) {
        if (win) win[s] = (int64_t)sum[s];
        if (tie) tie[s] = (int64_t)sum[9 + s];
        if (loss) loss[s] = (int64_t)sum[18 + s];
    }
    return MC_OK;
}
int mc_equity_exhaustive(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                         int64_t *win, int64_t *tie, int64_t *loss, uint64_t *n_boards) {
    EquityArgs a;
    int rc = check_equity_args(hole_masks, n_seats, board_mask, dead_mask, false, &a, nullptr);
    if (rc) return rc;
    const uint64_t total = board_count(a);
    if (n_boards) *n_boards = total;
    return mc_equity_exhaustive_range(hole_masks, n_seats, board_mask, dead_mask, 0, total, win, tie, loss);
}

int mc_equity_exhaustive_ev(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                            const int64_t *pot_bet, const int32_t *pot_n, const uint32_t *pot_players, int n_pots,
                            int64_t *win, int64_t *tie, int64_t *loss, int64_t *chips, uint64_t *n_boards) {
    EquityArgs a;
    int rc = check_equity_args(hole_masks, n_seats, board_mask, dead_mask, false, &a, nullptr);
    if (rc) return rc;
    PotArgs pots;
    rc = check_pots(pot_bet, pot_n, pot_players, n_pots, n_seats, &pots);
    if (rc) return rc;
    const uint64_t total = board_count(a);
    if (n_boards) *n_boards = total;
    unsigned long long sum[36];
    rc = sharded_small_job(total, 36, sum, [&](DevCtx &c, uint64_t lo, uint64_t cnt) -> int {
        CU(launch_equity_exhaustive_hash(a, lo, cnt, reinterpret_cast<long long *>(c.d_small), c.tb, c.info, c.stream, &pots));
        return MC_OK;
    });
    if (rc) return rc;
    for (int s = 0; s < n_seats; s++) {
        if (win) win[s] = (int64_t)sum[s];
        if (tie) tie[s] = (int64_t)sum[9 + s];
        if (loss) loss[s] = (int64_t)sum[18 + s];
        if (chips) chips[s] = (int64_t)sum[27 + s];
    }
    return MC_OK;
}

// ---- K4 ---------------------------------------------------------------------------------------------
int mc_equity_mc_ev_dev(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                        uint64_t seed, uint64_t first_trial, uint64_t n_trials,
                        const int64_t *pot_bet, const int32_t *pot_n, const uint32_t *pot_players, int n_pots,
                        int64_t *d_counts, void *stream) {
    EquityArgs a;
    int rc = check_equity_args(hole_masks, n_seats, board_mask, dead_mask, true, &a, nullptr);
    if (rc) return rc;
    PotArgs pots;
    rc = check_pots(pot_bet, pot_n, pot_players, n_pots, n_seats, &pots);
    if (rc) return rc;
    if (!d_counts) return fail(MC_ERR_INVALID, "null device buffer");
    std::shared_lock<std::shared_mutex> life(g_life);
    DeviceInfo di; const DeviceTables *tb;
    rc = current_tables(&di, &tb);
    if (rc) return rc;
    if (n_trials) CU(launch_equity_mc_fy(a, seed, first_trial, n_trials, reinterpret_cast<long long *>(d_counts), *tb, di, (cudaStream_t)stream, &pots));
    return MC_OK;
}
int mc_equity_mc_ev(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                    uint64_t seed, uint64_t first_trial, uint64_t n_trials,
                    const int64_t *pot_bet, const int32_t *pot_n, const uint32_t *pot_players, int n_pots,
                    int64_t *win, int64_t *tie, int64_t *loss, int64_t *chips) {
    EquityArgs a;
    int rc = check_equity_args(hole_masks, n_seats, board_mask, dead_mask, true, &a, nullptr);
    if (rc) return rc;
    PotArgs pots;
    rc = check_pots(pot_bet, pot_n, pot_players, n_pots, n_seats, &pots);
    if (rc) return rc;
    unsigned long long sum[36];
    rc = sharded_small_job(n_trials, 36, sum, [&](DevCtx &c, uint64_t lo, uint64_t cnt) -> int {
        CU(launch_equity_mc_fy(a, seed, first_trial + lo, cnt, reinterpret_cast<long long *>(c.d_small), c.tb, c.info, c.stream, &pots));
        return MC_OK;
    });
    if (rc) return rc;
    for (int s = 0; s < n_seats; s++) {
        if (win) win[s] = (int64_t)sum[s];
        if (tie) tie[s] = (int64_t)sum[9 + s];
        if (loss) loss[s] = (int64_t)sum[18 + s];
        if (chips) chips[s] = (int64_t)sum[27 + s];
    }
    return MC_OK;
}

int mc_equity_mc_dev(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                     uint64_t seed, uint64_t first_trial, uint64_t n_trials, int64_t *d_counts, void *stream) {
    EquityArgs a;
    int rc = check_equity_args(hole_masks, n_seats, board_mask, dead_mask, true, &a, nullptr);
    if (rc) return rc;
    if (!d_counts) return fail(MC_ERR_INVALID, "null device buffer");
    std::shared_lock<std::shared_mutex> life(g_life);
    DeviceInfo di; const DeviceTables *tb;
    rc = current_tables(&di, &tb);
    if (rc) return rc;
    if (n_trials) CU(launch_equity_mc_fy(a, seed, first_trial, n_trials, reinterpret_cast<long long *>(d_counts), *tb, di, (cudaStream_t)stream));
    return MC_OK;
}
int mc_equity_mc(const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                 uint64_t seed, uint64_t first_trial, uint64_t n_trials, int64_t *win, int64_t *tie, int64_t *loss) {
    EquityArgs a;
    int rc = check_equity_args(hole_masks, n_seats, board_mask, dead_mask, true, &a, nullptr);
    if (rc) return rc;
    unsigned long long sum[27];
    rc = sharded_small_job(n_trials, 27, sum, [&](DevCtx &c, uint64_t lo, uint64_t cnt) -> int {
        CU(launch_equity_mc_fy(a, seed, first_trial + lo, cnt, reinterpret_cast<long long *>(c.d_small), c.tb, c.info, c.stream));
        return MC_OK;
    });
    if (rc) return rc;
    for (int s = 0; s < n_seats; s++) {
        if (win) win[s] = (int64_t)sum[s];
        if (tie) tie[s] = (int64_t)sum[9 + s];
        if (loss) loss[s] = (int64_t)sum[18 + s];
    }
    return MC_OK;
}

// ---- Monte Carlo with a range per seat (deal spec R1) ----------------------------------------------------
namespace {
// integer alias table (Vose, two stacks filled in ascending index order -- the order is part of the deal spec, the CPU
// oracle builds the same table): slot i keeps thr[i] of its capacity W for its own hole, the rest belongs to alias[i]
void build_alias(const std::vector<uint32_t> &w, uint32_t W, RangeEntry *e) {
    const size_t n = w.size();
    std::vector<uint64_t> mass(n);
    std::vector<int> small, large;
    for (size_t i = 0; i < n; i++) { mass[i] = (uint64_t)w[i] * n; (mass[i] < W ? small : large).push_back((int)i); }
    while (!small.empty() && !large.empty()) {
        const int s = small.back(), l = large.back();
        small.pop_back(); large.pop_back();
        e[s].thr = (uint32_t)mass[s]; e[s].alias = (uint16_t)l;
        mass[l] -= (uint64_t)W - mass[s];
        (mass[l] < W ? small : large).push_back(l);
    }
    for (int l : large) { e[l].thr = W; e[l].alias = (uint16_t)l; }
    for (int s : small) { e[s].thr = W; e[s].alias = (uint16_t)s; }
}
}  // namespace

int mc_equity_mc_ranges(const int32_t *weights, const uint64_t *hole_masks, int n_seats, uint64_t board_mask, uint64_t dead_mask,
                        uint64_t seed, uint64_t first_trial, uint64_t n_trials,
                        int64_t *win, int64_t *tie, int64_t *loss, uint64_t *n_void) {
    if (!weights || !hole_masks) return fail(MC_ERR_INVALID, "null argument");
    if (n_seats < 1 || n_seats > MC_MAX_SEATS) return fail(MC_ERR_INVALID, "n_seats must be 1..9");
    if ((board_mask & ~DECK_MASK) || (dead_mask & ~DECK_MASK) || (board_mask & dead_mask)) return fail(MC_ERR_INVALID, "bad board / dead mask");
    if (__builtin_popcountll(board_mask) > 5) return fail(MC_ERR_INVALID, "more than 5 community cards");
    McRangeArgs a;
    std::memset(&a, 0, sizeof a);
    a.board = board_mask; a.n_seats = n_seats;
    uint64_t used0 = board_mask | dead_mask;
    for (int s = 0; s < n_seats; s++) if (hole_masks[s]) {
        const uint64_t h = hole_masks[s];
        if ((h & ~DECK_MASK) || __builtin_popcountll(h) != 2) return fail(MC_ERR_INVALID, "a hole mask must hold exactly two of the 52 cards");
        if (h & used0) return fail(MC_ERR_INVALID, "a card is used twice");
        used0 |= h; a.hole[s] = h;
    }
    a.used0 = used0;
    std::vector<RangeEntry> entries;
    int n_ranged = 0;
    for (int s = 0; s < n_seats; s++) {
        if (hole_masks[s]) continue;
        std::vector<uint32_t> w;
        const size_t off = entries.size();
        uint64_t W = 0;
        for (int c1 = 1, idx = 0; c1 < 52; c1++)
            for (int c0 = 0; c0 < c1; c0++, idx++) {
                const int32_t wt = weights[(size_t)s * MC_N_HOLES + idx];
                if (wt < 0 || wt > 32767) return fail(MC_ERR_INVALID, "weights must be 0..32767");
                const uint64_t h = mc_card_bit(c0) | mc_card_bit(c1);
                if (wt == 0 || (h & used0)) continue;
                RangeEntry e{};
                e.c0 = (unsigned char)__builtin_ctzll(mc_card_bit(c0)); e.c1 = (unsigned char)__builtin_ctzll(mc_card_bit(c1));
                entries.push_back(e);
                w.push_back((uint32_t)wt); W += (uint32_t)wt;
            }
        if (w.empty()) return fail(MC_ERR_INVALID, "a ranged seat has no hole left (every weighted hole holds a board, dead or fixed card)");
        build_alias(w, (uint32_t)W, entries.data() + off);
        a.rs[s] = RangeSeat{(uint32_t)off, (uint32_t)w.size(), (uint32_t)W, (uint32_t)((1ull << 32) % w.size()), (uint32_t)((1ull << 32) % W)};
        n_ranged++;
    }
    for (int b = 0; b < 64; b++) if (((DECK_MASK & ~used0) >> b) & 1ull) a.live[a.n_live++] = (unsigned char)b;
    if ((int)a.n_live < 5 - __builtin_popcountll(board_mask) + 2 * n_ranged) return fail(MC_ERR_INVALID, "not enough live cards");
    a.zone_live = (uint32_t)((1ull << 32) % a.n_live);
    unsigned long long sum[28];
    int rc = sharded_small_job(n_trials, 28, sum, [&](DevCtx &c, uint64_t lo, uint64_t cnt) -> int {
        if (!c.d_range) CU(cudaMalloc(&c.d_range, (size_t)MC_MAX_SEATS * MC_N_HOLES * sizeof(RangeEntry)));
        if (!entries.empty()) CU(cudaMemcpyAsync(c.d_range, entries.data(), entries.size() * sizeof(RangeEntry), cudaMemcpyHostToDevice, c.stream));
        CU(launch_equity_mc_ranges(a, c.d_range, seed, first_trial + lo, cnt, reinterpret_cast<long long *>(c.d_small), c.tb, c.info, c.stream));
        return MC_OK;
    });
    if (rc) return rc;
    for (int s = 0; s < n_seats; s++) {
        if (win) win[s] = (int64_t)sum[s];
        if (tie) tie[s] = (int64_t)sum[9 + s];
        if (loss) loss[s] = (int64_t)sum[18 + s];
    }
    if (n_void) *n_void = sum[27];
    return MC_OK;
}

// ---- K5 ---------------------------------------------------------------------------------------------
namespace {
int hole_class(uint64_t h) {                       // 169-class id of a two-card hole mask
    const int b1 = __builtin_ctzll(h), b2 = 63 - __builtin_clzll(h);
    const int r1 = (b1 & 15) - 2, r2 = (b2 & 15) - 2, s1 = b1 >> 4, s2 = b2 >> 4;
    const int hi = std::max(r1, r2), lo = std::min(r1, r2);
    if (hi == lo) return 13 * hi + hi;
    return s1 == s2 ? 13 * hi + lo : 13 * lo + hi;
}
uint64_t permute_suits(uint64_t m, const int *perm) {
    uint64_t r = 0;
    for (int s = 0; s < 4; s++) r |= ((m >> (16 * s)) & 0xFFFFull) << (16 * perm[s]);
    return r;
}
// every unordered pair of disjoint hole-card pairs, folded by the 24 suit permutations
void build_matchups_uncached(const std::vector<std::pair<int, int>> *only_cells, std::vector<Matchup> *out);
void build_matchups(const std::vector<std::pair<int, int>> *only_cells, std::vector<Matchup> *out) {
    if (only_cells) { build_matchups_uncached(only_cells, out); return; }
    static std::mutex mu;
    static std::vector<Matchup> all;                 // the full list never changes: build it once per process
    std::lock_guard<std::mutex> lk(mu);
    if (all.empty()) build_matchups_uncached(nullptr, &all);
    *out = all;
}
void build_matchups_uncached(const std::vector<std::pair<int, int>> *only_cells, std::vector<Matchup> *out) {
    std::vector<uint64_t> holes;
    for (int a = 0; a < 52; a++) for (int b = a + 1; b < 52; b++) holes.push_back(mc_card_bit(a) | mc_card_bit(b));
    int perms[24][4], np = 0;
    int p[4] = {0, 1, 2, 3};
    do { std::memcpy(perms[np++], p, sizeof p); } while (std::next_permutation(p, p + 4));
    std::map<std::pair<uint64_t, uint64_t>, int64_t> orbit;
    for (size_t i = 0; i < holes.size(); i++) for (size_t j = i + 1; j < holes.size(); j++) {
        if (holes[i] & holes[j]) continue;
        if (only_cells) {
            const int ca = hole_class(holes[i]), cb = hole_class(holes[j]);
            bool hit = false;
            for (auto &c : *only_cells) if ((c.first == ca && c.second == cb) || (c.first == cb && c.second == ca)) { hit = true; break; }
            if (!hit) continue;
        }
        std::pair<uint64_t, uint64_t> best(~0ull, ~0ull);
        for (int q = 0; q < 24; q++) {
            uint64_t x = permute_suits(holes[i], perms[q]), y = permute_suits(holes[j], perms[q]);
            if (x > y) std::swap(x, y);
            best = std::min(best, std::make_pair(x, y));
        }
        orbit[best]++;
    }
    out->clear();
    for (auto &kv : orbit) {
        Matchup m;
        m.hole_a = kv.first.first; m.hole_b = kv.first.second;
        m.cell_ab = hole_class(m.hole_a) * MC_N_CLASSES + hole_class(m.hole_b);
        m.cell_ba = hole_class(m.hole_b) * MC_N_CLASSES + hole_class(m.hole_a);
        m.weight = kv.second;
        out->push_back(m);
    }
}
int run_matchups(const std::vector<Matchup> &mus, std::vector<long long> *wtl_full) {
    Use use;
    if (use.rc) return use.rc;
    const int nd = (int)g_ctx.size();
    const size_t cells = (size_t)MC_N_CLASSES * MC_N_CLASSES * 3;
    std::vector<std::unique_lock<std::mutex>> locks;
    for (auto &c : g_ctx) locks.emplace_back(c->mu);
    // every matchup costs the same (C(48,5) boards), so a round-robin deal is already balanced
    std::vector<std::vector<Matchup>> part(nd);
    for (size_t i = 0; i < mus.size(); i++) part[i % nd].push_back(mus[i]);
    std::vector<std::vector<long long>> host(nd, std::vector<long long>(cells));
    DrainOnExit drain;                       // `host` must outlive every copy into it, also on an early error return
    for (int d = 0; d < nd; d++) {
        DevCtx &c = *g_ctx[d];
        CU(cudaSetDevice(c.device));
        if (!c.d_wtl) CU(cudaMalloc(&c.d_wtl, cells * sizeof(long long)));
        if (c.matchup_cap < part[d].size()) {
            if (c.d_matchups) cudaFree(c.d_matchups);
            CU(cudaMalloc(&c.d_matchups, part[d].size() * sizeof(Matchup)));
            c.matchup_cap = part[d].size();
        }
        CU(cudaMemsetAsync(c.d_wtl, 0, cells * sizeof(long long), c.stream));
        if (!part[d].empty()) {
            CU(cudaMemcpyAsync(c.d_matchups, part[d].data(), part[d].size() * sizeof(Matchup), cudaMemcpyHostToDevice, c.stream));
            CU(launch_preflop_hash(c.d_matchups, (int)part[d].size(), c.d_wtl, c.tb, c.info, c.stream));
        }
        CU(cudaMemcpyAsync(host[d].data(), c.d_wtl, cells * sizeof(long long), cudaMemcpyDeviceToHost, c.stream));
    }
    wtl_full->assign(cells, 0);
    for (int d = 0; d < nd; d++) {
        CU(cudaSetDevice(g_ctx[d]->device));
        CU(cudaStreamSynchronize(g_ctx[d]->stream));
        for (size_t i = 0; i < cells; i++) (*wtl_full)[i] += host[d][i];
    }
    return MC_OK;
}
}  // namespace

int mc_equity_ranges(const int32_t *weight_a, const int32_t *weight_b, uint64_t board_mask, uint64_t dead_mask,
                     int64_t wtl[3], int64_t *total) {
    Use use;
    if (use.rc) return use.rc;
    if (!weight_a || !weight_b || !wtl) return fail(MC_ERR_INVALID, "null argument");
    if ((board_mask & ~DECK_MASK) || (dead_mask & ~DECK_MASK) || (board_mask & dead_mask)) return fail(MC_ERR_INVALID, "bad board / dead mask");
    if (__builtin_popcountll(board_mask) > 5) return fail(MC_ERR_INVALID, "more than 5 community cards");
    std::vector<std::pair<uint64_t, int32_t>> ra, rb;
    for (int c1 = 1, idx = 0; c1 < 52; c1++)
        for (int c0 = 0; c0 < c1; c0++, idx++) {
            const uint64_t h = mc_card_bit(c0) | mc_card_bit(c1);
            if (weight_a[idx] < 0 || weight_a[idx] > 32767 || weight_b[idx] < 0 || weight_b[idx] > 32767) return fail(MC_ERR_INVALID, "weights must be 0..32767");
            if (h & (board_mask | dead_mask)) continue;
            if (weight_a[idx]) ra.emplace_back(h, weight_a[idx]);
            if (weight_b[idx]) rb.emplace_back(h, weight_b[idx]);
        }
    std::vector<Matchup> mus;
    for (auto &a : ra) for (auto &b : rb) {
        if (a.first & b.first) continue;
        Matchup m; m.hole_a = a.first; m.hole_b = b.first; m.cell_ab = 0; m.cell_ba = 0; m.weight = (int64_t)a.second * b.second;
        mus.push_back(m);
    }
    if (52 - __builtin_popcountll(board_mask | dead_mask) - 4 < 5 - __builtin_popcountll(board_mask)) return fail(MC_ERR_INVALID, "not enough live cards");
    {   // the weighted totals are int64: refuse ranges whose sum of weight products x boards cannot fit
        const uint64_t boards = host_binom(52 - __builtin_popcountll(board_mask | dead_mask) - 4, 5 - __builtin_popcountll(board_mask));
        unsigned __int128 tot = 0;
        for (auto &m : mus) tot += (unsigned __int128)(uint64_t)m.weight * boards;
        if (tot > (unsigned __int128)INT64_MAX) return fail(MC_ERR_INVALID, "sum of weight products x boards exceeds int64: scale the weights down");
    }
    const int nd = (int)g_ctx.size();
    DrainOnExit drain;
    std::vector<std::unique_lock<std::mutex>> locks;
    for (auto &c : g_ctx) locks.emplace_back(c->mu);
    const size_t per = (mus.size() + nd - 1) / nd;
    for (int d = 0; d < nd; d++) {
        DevCtx &c = *g_ctx[d];
        const size_t lo = std::min(mus.size(), per * d), hi = std::min(mus.size(), lo + per);
        CU(cudaSetDevice(c.device));
        CU(cudaMemsetAsync(c.d_small, 0, SMALL_WORDS * sizeof(unsigned long long), c.stream));
        if (hi > lo) {
            if (c.matchup_cap < hi - lo) {
                if (c.d_matchups) cudaFree(c.d_matchups);
                c.d_matchups = nullptr; c.matchup_cap = 0;
                CU(cudaMalloc(&c.d_matchups, (hi - lo) * sizeof(Matchup)));
                c.matchup_cap = hi - lo;
            }
            CU(cudaMemcpyAsync(c.d_matchups, mus.data() + lo, (hi - lo) * sizeof(Matchup), cudaMemcpyHostToDevice, c.stream));
            CU(launch_range_equity(c.d_matchups, (int)(hi - lo), board_mask, dead_mask, reinterpret_cast<long long *>(c.d_small), c.tb, c.info, c.stream));
        }
        CU(cudaMemcpyAsync(c.h_small, c.d_small, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c.stream));
    }
    int64_t sum[4] = {0, 0, 0, 0};
    for (int d = 0; d < nd; d++) {
        DevCtx &c = *g_ctx[d];
        CU(cudaSetDevice(c.device));
        CU(cudaStreamSynchronize(c.stream));
        for (int i = 0; i < 4; i++) sum[i] += (int64_t)c.h_small[i];
    }
    wtl[0] = sum[0]; wtl[1] = sum[1]; wtl[2] = sum[2];
    if (total) *total = sum[3];
    return MC_OK;
}

int mc_preflop_matrix(int64_t *wtl) {
    if (!wtl) return fail(MC_ERR_INVALID, "wtl is null");
    std::vector<Matchup> mus;
    build_matchups(nullptr, &mus);
    std::vector<long long> full;
    int rc = run_matchups(mus, &full);
    if (rc) return rc;
    for (size_t i = 0; i < full.size(); i++) wtl[i] = full[i];
    return MC_OK;
}
int mc_preflop_matrix_part(int part, int n_parts, int64_t *wtl, int32_t *n_matchups) {
    if (!wtl || n_parts < 1 || part < 0 || part >= n_parts) return fail(MC_ERR_INVALID, "bad part / n_parts");
    std::vector<Matchup> all, mine;
    build_matchups(nullptr, &all);
    for (size_t i = (size_t)part; i < all.size(); i += (size_t)n_parts) mine.push_back(all[i]);
    std::vector<long long> full;
    int rc = run_matchups(mine, &full);
    if (rc) return rc;
    for (size_t i = 0; i < full.size(); i++) wtl[i] = full[i];
    if (n_matchups) *n_matchups = (int32_t)mine.size();
    return MC_OK;
}
// ---- config 5 by boards (K7) -----------------------------------------------------------
namespace {
constexpr int PB_CHUNK = 16384;                       // boards per launch pair: 2 x 16384 x 1088 x 4 B = 143 MB of scratch
// 5-card boards folded by the 24 suit permutations: one representative (the lexicographically smallest image as a
// sorted list of card ids) per orbit, weight = orbit size.  134,459 boards, weights sum to C(52,5).
const std::vector<PfBoard> &canonical_boards() {
    static std::mutex mu;
    static std::vector<PfBoard> all;
    std::lock_guard<std::mutex> lk(mu);
    if (!all.empty()) return all;
    int perms[24][4], np = 0;
    int p[4] = {0, 1, 2, 3};
    do { std::memcpy(perms[np++], p, sizeof p); } while (std::next_permutation(p, p + 4));
    const int nt = (int)std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    std::vector<std::vector<PfBoard>> found(48);     // by first card id, so that the concatenation is ordered
    auto work = [&](int t) {
        for (int c0 = t; c0 < 48; c0 += nt)
            for (int c1 = c0 + 1; c1 < 49; c1++) for (int c2 = c1 + 1; c2 < 50; c2++)
                for (int c3 = c2 + 1; c3 < 51; c3++) for (int c4 = c3 + 1; c4 < 52; c4++) {
                    const int c[5] = {c0, c1, c2, c3, c4};
                    int stab = 0;
                    bool canonical = true;
                    for (int q = 0; q < 24 && canonical; q++) {
                        int img[5];
                        for (int k = 0; k < 5; k++) img[k] = 13 * perms[q][c[k] / 13] + c[k] % 13;
                        std::sort(img, img + 5);
                        int cmp = 0;
                        for (int k = 0; k < 5 && !cmp; k++) cmp = img[k] - c[k];
                        if (cmp < 0) canonical = false;
                        stab += cmp == 0;
                    }
                    if (!canonical) continue;
                    PfBoard b{};
                    for (int k = 0; k < 5; k++) b.card[k] = (unsigned char)(16 * (c[k] / 13) + c[k] % 13 + 2);
                    b.weight = (unsigned char)(24 / stab);
                    found[c0].push_back(b);
                }
    };
    std::vector<std::thread> th;
    for (int t = 0; t < nt; t++) th.emplace_back(work, t);
    for (auto &t : th) t.join();
    for (auto &v : found) all.insert(all.end(), v.begin(), v.end());
    return all;
}
}  // namespace

int mc_preflop_boards_count(int32_t *n_boards, int64_t *weight_sum) {
    const std::vector<PfBoard> &all = canonical_boards();
    if (n_boards) *n_boards = (int32_t)all.size();
    if (weight_sum) { int64_t s = 0; for (auto &b : all) s += b.weight; *weight_sum = s; }
    return MC_OK;
}

int mc_preflop_wins_by_boards(int part, int n_parts, int64_t *win, int32_t *n_boards) {
    if (!win || n_parts < 1 || part < 0 || part >= n_parts) return fail(MC_ERR_INVALID, "bad part / n_parts");
    Use use;
    if (use.rc) return use.rc;
    const std::vector<PfBoard> &all = canonical_boards();
    std::vector<PfBoard> mine;
    for (size_t i = (size_t)part; i < all.size(); i += (size_t)n_parts) mine.push_back(all[i]);
    if (n_boards) *n_boards = (int32_t)mine.size();
    const int nd = (int)g_ctx.size();
    const size_t cells = (size_t)MC_N_CLASSES * MC_N_CLASSES;
    std::vector<std::unique_lock<std::mutex>> locks;
    for (auto &c : g_ctx) locks.emplace_back(c->mu);
    std::vector<std::vector<PfBoard>> share(nd);
    for (size_t i = 0; i < mine.size(); i++) share[i % nd].push_back(mine[i]);
    std::vector<std::vector<unsigned long long>> host(nd, std::vector<unsigned long long>(cells));
    DrainOnExit drain;                       // `host` must outlive every copy into it, also on an early error return
    for (int d = 0; d < nd; d++) {
        DevCtx &c = *g_ctx[d];
        CU(cudaSetDevice(c.device));
        if (!c.d_pb_win) CU(cudaMalloc(&c.d_pb_win, cells * sizeof(unsigned long long)));
        if (!c.d_pb_scratch) CU(cudaMalloc(&c.d_pb_scratch, (size_t)2 * PB_CHUNK * PB_STRIDE * sizeof(uint32_t)));
        if (c.pb_board_cap < share[d].size()) {
            if (c.d_pb_boards) cudaFree(c.d_pb_boards);
            CU(cudaMalloc(&c.d_pb_boards, share[d].size() * sizeof(PfBoard)));
            c.pb_board_cap = share[d].size();
        }
        CU(cudaMemsetAsync(c.d_pb_win, 0, cells * sizeof(unsigned long long), c.stream));
        if (!share[d].empty())
            CU(cudaMemcpyAsync(c.d_pb_boards, share[d].data(), share[d].size() * sizeof(PfBoard), cudaMemcpyHostToDevice, c.stream));
        for (size_t off = 0; off < share[d].size(); off += PB_CHUNK) {
            const int n = (int)std::min<size_t>(PB_CHUNK, share[d].size() - off);
            CU(launch_preflop_boards(c.d_pb_boards + off, n, c.d_pb_scratch, c.d_pb_scratch + (size_t)PB_CHUNK * PB_STRIDE,
                                     c.d_pb_win, c.tb, c.info, c.stream));
        }
        CU(cudaMemcpyAsync(host[d].data(), c.d_pb_win, cells * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c.stream));
    }
    for (size_t i = 0; i < cells; i++) win[i] = 0;
    for (int d = 0; d < nd; d++) {
        CU(cudaSetDevice(g_ctx[d]->device));
        CU(cudaStreamSynchronize(g_ctx[d]->stream));
        for (size_t i = 0; i < cells; i++) win[i] += (int64_t)host[d][i];
    }
    return MC_OK;
}

int mc_preflop_matrix_from_wins(const int64_t *win, int64_t *wtl) {
    if (!win || !wtl) return fail(MC_ERR_INVALID, "null argument");
    // ordered pairs of disjoint holes per cell; each meets on C(48,5) boards
    static std::vector<int64_t> pairs;
    static std::mutex mu;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (pairs.empty()) {
            pairs.assign((size_t)MC_N_CLASSES * MC_N_CLASSES, 0);
            std::vector<uint64_t> holes;
            for (int a = 0; a < 52; a++) for (int b = a + 1; b < 52; b++) holes.push_back(mc_card_bit(a) | mc_card_bit(b));
            std::vector<int> cls(holes.size());
            for (size_t i = 0; i < holes.size(); i++) cls[i] = hole_class(holes[i]);
            for (size_t i = 0; i < holes.size(); i++) for (size_t j = 0; j < holes.size(); j++)
                if (!(holes[i] & holes[j])) pairs[(size_t)cls[i] * MC_N_CLASSES + cls[j]]++;
        }
    }
    const int64_t boards = (int64_t)host_binom(48, 5);
    for (int a = 0; a < MC_N_CLASSES; a++) for (int b = 0; b < MC_N_CLASSES; b++) {
        const size_t ab = (size_t)a * MC_N_CLASSES + b, ba = (size_t)b * MC_N_CLASSES + a;
        wtl[3 * ab + 0] = win[ab];
        wtl[3 * ab + 1] = pairs[ab] * boards - win[ab] - win[ba];
        wtl[3 * ab + 2] = win[ba];
    }
    return MC_OK;
}

int mc_preflop_matrix_by_boards(int64_t *wtl) {
    if (!wtl) return fail(MC_ERR_INVALID, "wtl is null");
    std::vector<int64_t> win((size_t)MC_N_CLASSES * MC_N_CLASSES);
    int rc = mc_preflop_wins_by_boards(0, 1, win.data(), nullptr);
    if (rc) return rc;
    return mc_preflop_matrix_from_wins(win.data(), wtl);
}

int mc_preflop_cells(const int32_t *class_a, const int32_t *class_b, int n_cells, int64_t *wtl) {
    if (!class_a || !class_b || !wtl || n_cells <= 0) return fail(MC_ERR_INVALID, "bad cell list");
    std::vector<std::pair<int, int>> cells;
    for (int i = 0; i < n_cells; i++) {
        if (class_a[i] < 0 || class_a[i] >= MC_N_CLASSES || class_b[i] < 0 || class_b[i] >= MC_N_CLASSES) return fail(MC_ERR_INVALID, "class id out of range");
        cells.emplace_back(class_a[i], class_b[i]);
    }
    std::vector<Matchup> mus;
    build_matchups(&cells, &mus);
    std::vector<long long> full;
    int rc = run_matchups(mus, &full);
    if (rc) return rc;
    for (int i = 0; i < n_cells; i++)
        for (int k = 0; k < 3; k++) wtl[3 * i + k] = full[((size_t)class_a[i] * MC_N_CLASSES + class_b[i]) * 3 + k];
    return MC_OK;
}

}  // extern "C"
