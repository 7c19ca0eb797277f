// find_rank_keys.cpp -- offline search for the 13 per-rank keys K[r] used by the incremental evaluator
// (csrc/tables): the sum of the keys of a hand's ranks must identify its rank multiset, i.e. all multisets of
// the same size (<= 7 cards, <= 4 per rank) must have distinct sums.  Greedy: each new rank takes the smallest
// key that keeps every size class collision-free.   g++ -O2 -o find_rank_keys find_rank_keys.cpp && ./find_rank_keys
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <vector>
int main() {
    const int MAXS = 7, NR = 13;
    const uint32_t LIMIT = 1u << 23;
    std::vector<std::vector<uint32_t>> S(MAXS + 1);
    S[0].push_back(0);
    std::vector<uint32_t> K;
    for (int r = 0; r < NR; r++) {
        std::vector<bool> bad(LIMIT, false);
        for (int s = 1; s <= MAXS; s++)
            for (int t1 = 1; t1 <= std::min(4, s); t1++)
                for (int t2 = 0; t2 < t1; t2++) {
                    const auto &A = S[s - t1], &B = S[s - t2];
                    const uint32_t d = t1 - t2;
                    for (uint32_t a : A) for (uint32_t b : B) if (b > a) { uint32_t q = b - a; if (q % d == 0 && q / d < LIMIT) bad[q / d] = true; }
                }
        uint32_t x = K.empty() ? 0 : K.back() + 1;
        while (x < LIMIT && bad[x]) x++;
        if (r == 0) x = 0;
        K.push_back(x);
        std::vector<std::vector<uint32_t>> N(MAXS + 1);
        for (int s = 0; s <= MAXS; s++) {
            for (int t = 0; t <= std::min(4, s); t++) for (uint32_t a : S[s - t]) N[s].push_back(a + t * x);
            std::sort(N[s].begin(), N[s].end());
            if (std::adjacent_find(N[s].begin(), N[s].end()) != N[s].end()) { printf("collision at rank %d size %d\n", r, s); return 1; }
        }
        S.swap(N);
        printf("K[%d] = %u   (|S7| = %zu, max sum7 = %u)\n", r, x, S[7].size(), S[7].empty() ? 0 : S[7].back());
        fflush(stdout);
    }
    printf("keys:");
    for (uint32_t k : K) printf(" %u,", k);
    printf("\n");
    return 0;
}
