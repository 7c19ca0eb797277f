/* A plain C99 consumer of include/mcpoker.h, the way a JNA / Panama / cgo binding sees the library: compiled by
 * tests/test_host.py with gcc -std=c99 -pedantic and linked against libmcpoker.so.  It only uses entry points that
 * need no device (helpers, error convention); the compute calls are exercised by the GPU tests. */
#include <stdio.h>
#include <string.h>
#include "mcpoker.h"

int main(void) {
    int32_t cat = -1, hit[5], n_hit = 0, kick[5], n_kick = 0;
    uint64_t hist[9];
    if (mc_abi_version() != 2) return 1;
    /* card id 12 = ace of hearts (COMPLETE-DECK index): suit lane 0, rank 14 */
    if (mc_card_bit(12) != ((uint64_t)1 << 14)) return 2;
    if (mc_card_bit(52) != 0) return 3;
    /* [1 [9 9] [14 8 2]]: a pair of nines */
    if (mc_key_decode((1u << 20) | (9u << 16) | (9u << 12) | (14u << 8) | (8u << 4) | 2u, &cat, hit, &n_hit, kick, &n_kick) != MC_OK) return 4;
    if (cat != 1 || n_hit != 2 || n_kick != 3 || hit[0] != 9 || hit[1] != 9 || kick[0] != 14 || kick[1] != 8 || kick[2] != 2) return 5;
    /* no device was initialised: compute calls fail loudly with a message, they never fall back */
    if (mc_enumerate7(hist, NULL) != MC_ERR_NOT_INIT) return 6;
    if (strlen(mc_last_error()) == 0) return 7;
    printf("abi consumer ok: version %d, %d exported seats max\n", mc_abi_version(), MC_MAX_SEATS);
    return 0;
}
