// host_check.cpp -- TEST HARNESS: compiles the product's evaluator header (plain integer code) with g++ and
// checks it against the CPU oracle, so the formulation can be verified exhaustively on a box without a GPU.
// Not part of the product; the product calls these functions from CUDA kernels only.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "../monte-carlo_b200/csrc/eval.cuh"
#include "../monte-carlo_b200/csrc/eval_tbl.cuh"
#include "../monte-carlo_b200/csrc/hashtab_build.h"
#include <vector>
#include "../oracle/mc_oracle.h"

static uint64_t BINOM[65][9];
static void unrank(uint64_t idx, int k, int *c) {
    for (int i = k; i >= 1; i--) { int x = i - 1; while (BINOM[x + 1][i] <= idx) x++; c[i - 1] = x; idx -= BINOM[x][i]; }
}
static int next(int *c, int k, int n) {
    for (int i = 0; i < k; i++) { int lim = (i + 1 < k) ? c[i + 1] : n; if (c[i] + 1 < lim) { c[i]++; for (int j = 0; j < i; j++) c[j] = j; return 1; } }
    return 0;
}
// one hand through a stream image, step by step as k1h_fast (kernels.cu) does it
template <class LAY>
static uint32_t stream_lookup(const std::vector<unsigned char> &img, const mcp::StreamHash &SH, uint64_t m) {
    auto rd32 = [&](uint32_t a) { uint32_t v; memcpy(&v, img.data() + a, 4); return v; };
    auto rd16 = [&](uint32_t a) { uint16_t v; memcpy(&v, img.data() + a, 2); return (uint32_t)v; };
    uint32_t c0, c1, c2;
    mcp::sh_planes((uint32_t)m, (uint32_t)(m >> 32), &c0, &c1, &c2);
    const uint32_t t = mcp::sh_t(c0, c1, c2, SH), bk = mcp::sh_bucket(mcp::sh_b(c0, c1, c2, SH));
    uint32_t fm = 0;
    for (int s4 = 0; s4 < 4; s4++) { uint32_t lane = (uint32_t)(m >> (16 * s4)) & 0xFFFFu; if (__builtin_popcount(lane) >= 5) fm = lane; }
    uint32_t ka = rd16(LAY::V_OFF + 2u * mcp::sh_slot<LAY>(t) + rd16(LAY::D_OFF + 2u * bk));   // D' is in bytes
    if (fm) ka = rd16(LAY::F_OFF + 2u * (fm >> 2));
    return rd32(ka);
}
// mode 2: the additive-key / perfect-hash evaluator (hashtab.cuh + hashtab_build.h) on every 7-card hand.
// The representatives are evaluated here with eval.cuh compiled for the host; in the product mc_init does it on the GPU.
static int check_hash() {
    static mcp::HashTables T;
    std::vector<mcp::HashRep> reps;
    std::string err = "no multiplier tried";
    for (uint32_t mul : mcp::HT_MUL_CANDIDATES) {
        mcp::ht_representatives(mul, &reps);
        std::vector<uint32_t> packed(reps.size());
        for (size_t i = 0; i < reps.size(); i++) packed[i] = mcp::eval_packed(reps[i].mask);
        err = mcp::ht_build(reps, packed, mul, &T);
        if (err.empty()) break;
    }
    if (!err.empty()) { fprintf(stderr, "ht_build: %s\n", err.c_str()); return 1; }
    printf("hash tables: mul=%08x classes=%zu reps=%zu\n", T.mul, T.nib_keys.size(), reps.size());
    // the shared-memory images of the streaming kernel (K1), read here exactly as k1h_fast reads them
    std::vector<mcp::StreamRep> sreps;
    mcp::ht_stream_representatives(&sreps);
    std::vector<uint32_t> spacked(sreps.size());
    for (size_t i = 0; i < sreps.size(); i++) spacked[i] = mcp::eval_packed(sreps[i].mask);
    static mcp::StreamImages IM;
    err = mcp::ht_build_stream_images(sreps, spacked, &IM);
    if (!err.empty()) { fprintf(stderr, "ht_build_stream_images: %s\n", err.c_str()); return 1; }
    printf("stream hash multipliers: %08x %08x %08x %08x / %08x %08x %08x %08x, %zu classes\n", IM.h7.m1, IM.h7.m2, IM.h7.m3, IM.h7.m4,
           IM.h56.m1, IM.h56.m2, IM.h56.m3, IM.h56.m4, IM.class_nib.size());
    const mcp::StreamHash SH = IM.h7;
    const std::vector<unsigned char> &sb = IM.key7;
    auto stream_key = [&](uint64_t m) -> uint32_t { return stream_lookup<mcp::Lay7>(sb, SH, m); };
    // flop / turn images: every 5- and 6-card hand through Lay56, key flavour and class flavour
    for (int k = 5; k <= 6; k++) {
        uint64_t tot = BINOM[52][k], bad56 = 0;
#pragma omp parallel for reduction(+ : bad56) schedule(static, 1)
        for (int th = 0; th < 64; th++) {
            uint64_t chunk = (tot + 63) / 64, lo = chunk * th, hi = lo + chunk > tot ? tot : lo + chunk;
            if (lo >= hi) continue;
            int c[7]; unrank(lo, k, c);
            for (uint64_t i = lo; i < hi; i++) {
                uint64_t m = 0;
                for (int j = 0; j < k; j++) m |= 1ull << (16 * (c[j] / 13) + c[j] % 13 + 2);
                const uint32_t want = orc_eval_mask(m);
                const uint32_t gid = stream_lookup<mcp::Lay56>(IM.cls56, IM.h56, m);
                if (stream_lookup<mcp::Lay56>(IM.key56, IM.h56, m) != want || gid >= IM.class_nib.size() || IM.class_nib[gid] != want) {
                    if (bad56 < 5) fprintf(stderr, "LAY56 MISMATCH mask=%016llx\n", (unsigned long long)m);
                    bad56++;
                }
                next(c, k, 52);
            }
        }
        printf("k=%d deck=52 hands=%llu mismatches=%llu (Lay56 stream images)\n", k, (unsigned long long)tot, (unsigned long long)bad56);
        if (bad56) return 1;
    }
    for (size_t i = 1; i < IM.class_nib.size(); i++) if (IM.class_nib[i - 1] >= IM.class_nib[i]) { fprintf(stderr, "class keys not ascending\n"); return 1; }
    uint64_t total = BINOM[52][7], kbad = 0;
#pragma omp parallel for reduction(+ : kbad) schedule(static, 1)
    for (int th = 0; th < 64; th++) {
        uint64_t chunk = (total + 63) / 64, lo = chunk * th, hi = lo + chunk > total ? total : lo + chunk;
        if (lo >= hi) continue;
        int c[7]; unrank(lo, 7, c);
        for (uint64_t i = lo; i < hi; i++) {
            uint64_t m = 0; uint32_t t = 0, sn = mcp::SN_BIAS;
            for (int j = 0; j < 7; j++) { const int p = 16 * (c[j] / 13) + c[j] % 13 + 2; m |= 1ull << p; t += T.card[p].rk; sn += T.card[p].sn; }
            const uint32_t code = mcp::ht_class(t, sn, (uint32_t)m, (uint32_t)(m >> 32), T.blob.D, T.blob.F, T.blob.V);
            const uint32_t got = T.nib_keys[code];
            const uint32_t want = orc_eval_mask(m);
            if (stream_key(m) != want) { if (kbad < 5) fprintf(stderr, "STREAM BLOB MISMATCH mask=%016llx\n", (unsigned long long)m); kbad++; }
            const uint32_t gid = stream_lookup<mcp::Lay7>(IM.cls7, SH, m);
            if (gid >= IM.class_nib.size() || IM.class_nib[gid] != want) { if (kbad < 5) fprintf(stderr, "CLASS IMAGE MISMATCH mask=%016llx\n", (unsigned long long)m); kbad++; }
            if (got != want) { if (kbad < 5) fprintf(stderr, "HASH MISMATCH mask=%016llx want=%06x got=%06x code=%04x\n", (unsigned long long)m, want, got, code); kbad++; }
            next(c, 7, 52);
        }
    }
    printf("k=7 deck=52 hands=%llu mismatches=%llu (hash evaluator)\n", (unsigned long long)total, (unsigned long long)kbad);
    return kbad ? 1 : 0;
}
int main(int argc, char **argv) {
    int use56 = argc > 1 ? atoi(argv[1]) : 0;     // 1: 56-card deck incl. rank 1 (k = 5 only)
    for (int n = 0; n <= 64; n++) for (int k = 0; k <= 8; k++) BINOM[n][k] = k == 0 ? 1 : (n == 0 ? 0 : BINOM[n - 1][k - 1] + BINOM[n - 1][k]);
    if (use56 == 2) return check_hash();
    int ncards = use56 ? 56 : 52, per = use56 ? 14 : 13, base = use56 ? 1 : 2;
    uint64_t bad = 0, total_all = 0;
    // the shared-memory table of eval_tbl.cuh, built on the host with the same entry function
    std::vector<uint32_t> T14(1 << 14), T13(1 << 13), CC(36);
    for (uint32_t x = 0; x < (1u << 14); x++) T14[x] = mcp::nib_entry(x, 1);
    for (uint32_t x = 0; x < (1u << 13); x++) T13[x] = mcp::nib_entry(x, 2);
    for (int c = 0; c < 9; c++) { uint32_t v[4]; mcp::cat_consts(c, v); for (int i = 0; i < 4; i++) CC[4 * c + i] = v[i]; }
    for (int k = 5; k <= (use56 ? 6 : 7); k++) {
        uint64_t total = BINOM[ncards][k];
        uint64_t kbad = 0;
#pragma omp parallel for reduction(+ : kbad) schedule(static, 1)
        for (int t = 0; t < 64; t++) {
            uint64_t chunk = (total + 63) / 64, lo = chunk * t, hi = lo + chunk > total ? total : lo + chunk;
            if (lo >= hi) continue;
            int c[7]; unrank(lo, k, c);
            for (uint64_t i = lo; i < hi; i++) {
                uint64_t m = 0;
                for (int j = 0; j < k; j++) m |= 1ull << (16 * (c[j] / per) + c[j] % per + base);
                uint32_t want = orc_eval_mask(m);
                uint32_t pk = mcp::eval_packed(m);
                uint32_t got = mcp::packed_to_nibble(pk);
                uint32_t t14 = mcp::eval_nibble_tbl<14>((uint32_t)m, (uint32_t)(m >> 32), T14.data(), CC.data());
                uint32_t t13 = use56 ? want : mcp::eval_nibble_tbl<13>((uint32_t)m, (uint32_t)(m >> 32), T13.data(), CC.data());
                if (t14 != want || t13 != want) got = t14 != want ? t14 : t13;
                if (got != want || !mcp::valid_hand(m)) { if (kbad < 5) fprintf(stderr, "MISMATCH k=%d mask=%016llx want=%06x got=%06x packed=%08x\n", k, (unsigned long long)m, want, got, pk); kbad++; }
                next(c, k, ncards);
            }
        }
        printf("k=%d deck=%d hands=%llu mismatches=%llu\n", k, ncards, (unsigned long long)total, (unsigned long long)kbad);
        bad += kbad; total_all += total;
    }
    return bad ? 1 : 0;
}
