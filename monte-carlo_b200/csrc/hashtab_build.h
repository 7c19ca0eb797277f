// hashtab_build.h -- host-side construction of the perfect-hash tables of hashtab.cuh.  Pure table layout work:
// it never evaluates a hand.  The caller supplies, for every 7-card rank multiset and every 5..7-card flush lane,
// the packed key of a representative hand (mc_init computes them ON THE GPU with the bit-plane evaluator of
// eval.cuh; tests/host_check.cpp computes them with the same header compiled for the host), and gets back
//   blob      D / F / V of hashtab.cuh, values = class codes (dense index of the value in ascending order)
//   nib_keys  dense class index -> public nibble key (ascending)
//   cat_base  dense index of the first class of each category (cat_base[9] = number of classes)
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>
#include "hashtab.cuh"

namespace mcp {

// premultiplied key of rank r (2..14) for multiplier mul
inline uint32_t rank_key(int r, uint32_t mul) { return RK_RAW[r - 2] * mul; }

struct HashRep { uint64_t mask; uint32_t t; uint32_t flush_index; };   // flush_index = 0xFFFFFFFF for rank-pattern reps

// representative hands: one per 7-card rank multiset (suits dealt round-robin so that no suit holds more than two
// cards), then one per rank set of 5, 6 or 7 cards in a single suit
inline void ht_representatives(uint32_t mul, std::vector<HashRep> *out) {
    out->clear();
    int cnt[13];
    struct Rec {
        static void go(int r, int left, int *cnt, uint32_t mul, std::vector<HashRep> *out) {
            if (r == 13) {
                if (left) return;
                HashRep h{0, 0, 0xFFFFFFFFu};
                int k = 0;
                for (int i = 0; i < 13; i++)
                    for (int c = 0; c < cnt[i]; c++, k++) {
                        h.mask |= 1ull << (16 * (k & 3) + i + 2);
                        h.t += rank_key(i + 2, mul);
                    }
                out->push_back(h);
                return;
            }
            for (int c = 0; c <= (left < 4 ? left : 4); c++) { cnt[r] = c; go(r + 1, left - c, cnt, mul, out); }
            cnt[r] = 0;
        }
    };
    std::memset(cnt, 0, sizeof cnt);
    Rec::go(0, 7, cnt, mul, out);
    for (uint32_t x = 0; x < (uint32_t)HT_FLUSH; x++) {
        const int n = __builtin_popcount(x);
        if (n < 5 || n > 7) continue;
        out->push_back(HashRep{(uint64_t)x << 2, 0, x});
    }
}

struct HashTables {
    HashBlob blob;
    std::vector<uint32_t> nib_keys;
    uint32_t cat_base[10];
    uint32_t mul;
    CardKey card[64];                 // by bit position
};

// packed[i] = packed key (eval.cuh) of reps[i].  Returns "" or an error description.
inline std::string ht_build(const std::vector<HashRep> &reps, const std::vector<uint32_t> &packed, uint32_t mul,
                            HashTables *T) {
    if (reps.size() != packed.size()) return "size mismatch";
    std::vector<uint32_t> keys(packed);
    std::sort(keys.begin(), keys.end());
    keys.erase(std::unique(keys.begin(), keys.end()), keys.end());
    std::vector<uint16_t> code(keys.size());
    T->nib_keys.resize(keys.size());
    for (int c = 0; c <= 9; c++) T->cat_base[c] = (uint32_t)keys.size();
    if (keys.size() >= 65535) return "more than 65534 classes";
    uint32_t prev_cat = 99;
    for (size_t i = 0; i < keys.size(); i++) {
        const uint32_t cat = keys[i] >> 28;
        if (cat > 8) return "a representative evaluated to an invalid key";
        if (cat != prev_cat) { prev_cat = cat; T->cat_base[cat] = (uint32_t)i; }
        code[i] = (uint16_t)i;                       // class code = dense index of the value in ascending order
        T->nib_keys[i] = packed_to_nibble(keys[i]);
    }
    for (int c = 8; c >= 0; c--) if (T->cat_base[c] == keys.size() && c < 9) T->cat_base[c] = T->cat_base[c + 1];
    auto code_of = [&](uint32_t pk) { return code[std::lower_bound(keys.begin(), keys.end(), pk) - keys.begin()]; };

    std::memset(&T->blob, 0, sizeof T->blob);
    std::vector<std::vector<std::pair<uint32_t, uint16_t>>> bucket(HT_BUCKETS);     // (t >> 16, code)
    size_t n_multi = 0;
    for (size_t i = 0; i < reps.size(); i++) {
        if (reps[i].flush_index != 0xFFFFFFFFu) { T->blob.F[reps[i].flush_index] = code_of(packed[i]); continue; }
        bucket[ht_bucket(reps[i].t)].push_back({reps[i].t >> 16, code_of(packed[i])});
        n_multi++;
    }
    if (n_multi != (size_t)HT_N_MULTISETS) return "unexpected number of rank multisets";
    std::vector<int> order(HT_BUCKETS);
    for (int i = 0; i < HT_BUCKETS; i++) order[i] = i;
    std::sort(order.begin(), order.end(), [&](int a, int b) { return bucket[a].size() != bucket[b].size() ? bucket[a].size() > bucket[b].size() : a < b; });
    std::vector<char> used(HT_SLOTS, 0);
    for (int b : order) {
        auto &items = bucket[b];
        if (items.empty()) break;
        for (size_t i = 0; i < items.size(); i++)
            for (size_t j = i + 1; j < items.size(); j++)
                if (items[i].first == items[j].first) return "two multisets share bucket and slot: pick another multiplier";
        uint32_t d = 0;
        for (; d <= (uint32_t)HT_MAX_DISP; d++) {
            bool ok = true;
            for (auto &it : items) if (used[it.first + d]) { ok = false; break; }
            if (ok) break;
        }
        if (d > (uint32_t)HT_MAX_DISP) return "displacement search failed: pick another multiplier";
        T->blob.D[b] = (uint16_t)d;
        for (auto &it : items) { used[it.first + d] = 1; T->blob.V[it.first + d] = it.second; }
    }
    T->mul = mul;
    for (int p = 0; p < 64; p++) {
        const int r = p & 15, s = p >> 4;
        T->card[p].rk = (r >= 2 && r <= 14) ? rank_key(r, mul) : 0u;
        T->card[p].sn = 1u << (4 * s);
    }
    return "";
}

// ---- the shared-memory images of the streaming kernel (layouts and hash: hashtab.cuh "K1 layout") -------------
// Representative hands, in a fixed order: every 5-, 6- and 7-card rank multiset (<= 4 of a rank; suits dealt
// round-robin, so no suit holds more than two cards), then one hand per rank set of 5, 6 or 7 cards in suit 0.
// The caller evaluates them (mc_init: on the GPU with the bit-plane evaluator; tests/host_check.cpp: the same header
// on the host) and hands the packed keys to ht_build_stream_images.
struct StreamRep { uint64_t mask; int n_cards; uint32_t flush_index; };       // flush_index = 0xFFFFFFFF for multisets
inline void ht_stream_representatives(std::vector<StreamRep> *out) {
    out->clear();
    struct Rec {
        static void go(int r, int left, int K, int *cnt, std::vector<StreamRep> *out) {
            if (r == 13) {
                if (left) return;
                StreamRep h{0, K, 0xFFFFFFFFu};
                int k = 0;
                for (int i = 0; i < 13; i++)
                    for (int c = 0; c < cnt[i]; c++, k++) h.mask |= 1ull << (16 * (k & 3) + i + 2);
                out->push_back(h);
                return;
            }
            for (int c = 0; c <= (left < 4 ? left : 4); c++) { cnt[r] = c; go(r + 1, left - c, K, cnt, out); }
            cnt[r] = 0;
        }
    };
    int cnt[13];
    for (int K = 5; K <= 7; K++) { std::memset(cnt, 0, sizeof cnt); Rec::go(0, K, K, cnt, out); }
    for (uint32_t x = 0; x < (uint32_t)HT_FLUSH; x++) {
        const int n = __builtin_popcount(x);
        if (n >= 5 && n <= 7) out->push_back(StreamRep{(uint64_t)x << 2, n, x});
    }
}

struct StreamImages {
    std::vector<unsigned char> key7, cls7, key56, cls56;      // Lay7 / Lay56 images, nibble-key and class-id flavour
    StreamHash h7{}, h56{};
    std::vector<uint32_t> class_nib;                           // global class id -> nibble key (7,462, ascending)
};

// one image: hash-and-displace over the plane names of `items`; the first region holds, per value of the image
// (dense, ascending), the nibble key (key image) or the global class id (class image)
struct StreamItem { uint32_t c0, c1, c2, packed; };
template <class LAY>
inline std::string ht_build_stream_image(const std::vector<StreamItem> &items,
                                         const std::vector<std::pair<uint32_t, uint32_t>> &flush /* (F index, packed key) */,
                                         const std::vector<uint32_t> &global_packed /* ascending */,
                                         std::vector<unsigned char> *key_img, std::vector<unsigned char> *cls_img, StreamHash *S) {
    std::vector<uint32_t> vals;
    for (auto &it : items) vals.push_back(it.packed);
    for (auto &f : flush) vals.push_back(f.second);
    std::sort(vals.begin(), vals.end());
    vals.erase(std::unique(vals.begin(), vals.end()), vals.end());
    if (vals.size() > LAY::N_KEYS) return "too many values for the first region of the image";
    key_img->assign(LAY::BYTES, 0);
    cls_img->assign(LAY::BYTES, 0);
    auto put32 = [&](std::vector<unsigned char> *B, uint32_t addr, uint32_t v) { std::memcpy(B->data() + addr, &v, 4); };
    auto put16 = [&](std::vector<unsigned char> *B, uint32_t addr, uint16_t v) { std::memcpy(B->data() + addr, &v, 2); };
    for (size_t d = 0; d < vals.size(); d++) {
        const auto g = std::lower_bound(global_packed.begin(), global_packed.end(), vals[d]);
        if (g == global_packed.end() || *g != vals[d]) return "a value of the image is not a five-card value";
        put32(key_img, LAY::KEY_OFF + 4u * (uint32_t)d, packed_to_nibble(vals[d]));
        put32(cls_img, LAY::KEY_OFF + 4u * (uint32_t)d, (uint32_t)(g - global_packed.begin()));
    }
    auto addr_of = [&](uint32_t pk) { return (uint16_t)(LAY::KEY_OFF + 4u * (uint32_t)(std::lower_bound(vals.begin(), vals.end(), pk) - vals.begin())); };
    for (auto &f : flush) { put16(key_img, LAY::F_OFF + 2u * f.first, addr_of(f.second)); put16(cls_img, LAY::F_OFF + 2u * f.first, addr_of(f.second)); }
    uint64_t rng = 0x9E3779B97F4A7C15ull;                     // splitmix64: a fixed sequence of odd multipliers
    auto next_odd = [&]() { rng += 0x9E3779B97F4A7C15ull; uint64_t z = rng; z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
                            z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31; return (uint32_t)(z >> 32) | 1u; };
    constexpr uint32_t MAX_DISP = LAY::N_SLOTS - LAY::SLOT_MUL - 1;
    for (int attempt = 0; attempt < 256; attempt++) {
        StreamHash H{next_odd(), next_odd(), next_odd(), next_odd()};
        std::vector<std::vector<std::pair<uint32_t, uint16_t>>> bucket(HT_BUCKETS);     // (slot before displacement, key address)
        for (auto &it : items) bucket[sh_bucket(sh_b(it.c0, it.c1, it.c2, H))].push_back({sh_slot<LAY>(sh_t(it.c0, it.c1, it.c2, H)), addr_of(it.packed)});
        std::vector<int> order(HT_BUCKETS);
        for (int i = 0; i < HT_BUCKETS; i++) order[i] = i;
        std::sort(order.begin(), order.end(), [&](int a, int b) { return bucket[a].size() != bucket[b].size() ? bucket[a].size() > bucket[b].size() : a < b; });
        std::vector<char> used(LAY::N_SLOTS, 0);
        std::vector<uint16_t> D(HT_BUCKETS, 0), V(LAY::N_SLOTS, 0);
        bool ok = true;
        for (int b : order) {
            auto &its = bucket[b];
            if (its.empty()) break;
            for (size_t i = 0; i < its.size() && ok; i++)
                for (size_t j = i + 1; j < its.size(); j++) if (its[i].first == its[j].first) { ok = false; break; }
            if (!ok) break;
            uint32_t d = 0;
            for (; d <= MAX_DISP; d++) {
                bool fits = true;
                for (auto &it : its) if (used[it.first + d]) { fits = false; break; }
                if (fits) break;
            }
            if (d > MAX_DISP) { ok = false; break; }
            D[b] = (uint16_t)(2u * d);                                            // D' = byte displacement
            for (auto &it : its) { used[it.first + d] = 1; V[it.first + d] = it.second; }
        }
        if (!ok) continue;
        // an empty slot keeps address 0 (a valid key address; never read for a valid hand)
        for (auto *B : {key_img, cls_img}) {
            std::memcpy(B->data() + LAY::D_OFF, D.data(), HT_BUCKETS * 2);
            std::memcpy(B->data() + LAY::V_OFF, V.data(), LAY::N_SLOTS * 2);
        }
        *S = H;
        return "";
    }
    return "no multiplier set gave a displacement table for the streaming hash";
}

// packed[i] = packed key (eval.cuh) of reps[i] (ht_stream_representatives order)
inline std::string ht_build_stream_images(const std::vector<StreamRep> &reps, const std::vector<uint32_t> &packed, StreamImages *out) {
    if (reps.size() != packed.size()) return "size mismatch";
    // the global class list = every five-card value: the 5-card multisets and the 5-rank flushes
    std::vector<uint32_t> global;
    for (size_t i = 0; i < reps.size(); i++) if (reps[i].n_cards == 5) global.push_back(packed[i]);
    std::sort(global.begin(), global.end());
    global.erase(std::unique(global.begin(), global.end()), global.end());
    if (global.size() != (size_t)N_HAND_CLASSES) return "unexpected number of five-card values";
    out->class_nib.resize(global.size());
    for (size_t i = 0; i < global.size(); i++) out->class_nib[i] = packed_to_nibble(global[i]);
    std::vector<StreamItem> it7, it56;
    std::vector<std::pair<uint32_t, uint32_t>> fl7, fl56;
    for (size_t i = 0; i < reps.size(); i++) {
        if (reps[i].flush_index != 0xFFFFFFFFu) {
            fl7.push_back({reps[i].flush_index, packed[i]});
            if (reps[i].n_cards <= 6) fl56.push_back({reps[i].flush_index, packed[i]});
            continue;
        }
        StreamItem it;
        sh_planes((uint32_t)reps[i].mask, (uint32_t)(reps[i].mask >> 32), &it.c0, &it.c1, &it.c2);
        it.packed = packed[i];
        (reps[i].n_cards == 7 ? it7 : it56).push_back(it);
    }
    if (it7.size() != (size_t)HT_N_MULTISETS) return "unexpected number of 7-card rank multisets";
    std::string err = ht_build_stream_image<Lay7>(it7, fl7, global, &out->key7, &out->cls7, &out->h7);
    if (!err.empty()) return "Lay7: " + err;
    err = ht_build_stream_image<Lay56>(it56, fl56, global, &out->key56, &out->cls56, &out->h56);
    if (!err.empty()) return "Lay56: " + err;
    return "";
}

constexpr uint32_t HT_MUL_CANDIDATES[] = {0x165667B1u, 0x27D4EB2Fu, 0x85EBCA6Bu, 0xC2B2AE35u, 0x9E3779B1u};

}  // namespace mcp
