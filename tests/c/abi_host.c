/* Plain-C client of libchacra_b200.so: proves the boundary is a C ABI (no C++ or torch types in the signatures) and
 * exercises the host-only entry points.  Built and run by tests/test_abi.py with gcc.
 *   gcc -std=c99 -Wall -Werror -I include tests/c/abi_host.c -o abi_host -L chacra_b200/csrc -lchacra_b200 */
#define _POSIX_C_SOURCE 200809L
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "chacra_b200.h"

static int fail(const char* what) { fprintf(stderr, "FAIL: %s\n", what); return 1; }

int main(int argc, char** argv) {
    if (argc < 2) return fail("usage: abi_host <scratch.tsv>");
    if (chc_abi_version() != 1) return fail("abi version");
    if (sizeof(chc_contact) != 32 || sizeof(chc_near) != 32 || sizeof(chc_xtc_frame) != 64) return fail("record sizes");

    char text[32];
    if (chc_format_fixed3(3.0625f, text) != 5 || strcmp(text, "3.062") != 0) return fail("fixed3 tie");
    if (chc_format_fixed3(2.8714f, text) != 5 || strcmp(text, "2.871") != 0) return fail("fixed3");

    /* three atoms, two frames, one duplicate record, one pair given in reverse label order */
    const char* labels = "A:ASP:1:OD1" "A:LYS:2:NZ" "A:ALA:3:CB";
    const int64_t off[4] = {0, 11, 21, 31};
    const int32_t rank[3] = {1, 2, 0};                       /* string order: ALA < ASP < LYS */
    const char* itypes = "sb" "vdw";
    const int32_t ioff[3] = {0, 2, 5};
    const int32_t block[2] = {0, 4};
    chc_contact rec[4] = {
        {1, 1, 1, 2, -1, -1, 3.5f, 0}, {0, 0, 1, 0, -1, -1, 2.8714f, 0}, {0, 0, 1, 0, -1, -1, 2.8714f, 0}, {0, 1, 0, 2, -1, -1, 4.0f, 0}};
    FILE* fh = fopen(argv[1], "w");
    if (!fh) return fail("open scratch");
    fputs("# total_frames:2 beg:0 end:1 stride:1 interaction_types:sb,vdw\n", fh);
    fflush(fh);
    const int64_t lines = chc_write_contacts_tsv(fileno(fh), rec, 4, labels, off, rank, 3, 1, itypes, ioff, block, 2, 0, 1, 1, 2);
    fclose(fh);
    if (lines != 3) return fail("line count (duplicate dropped)");
    fh = fopen(argv[1], "r");
    char got[512]; size_t n = fread(got, 1, sizeof got - 1, fh); got[n] = 0; fclose(fh);
    const char* want = "# total_frames:2 beg:0 end:1 stride:1 interaction_types:sb,vdw\n"
                       "0\tsb\tA:ASP:1:OD1\tA:LYS:2:NZ\t2.871\n"
                       "0\tvdw\tA:ALA:3:CB\tA:ASP:1:OD1\t4.000\n"
                       "1\tvdw\tA:ALA:3:CB\tA:LYS:2:NZ\t3.500\n";
    if (strcmp(got, want) != 0) { fputs(got, stderr); return fail("tsv text"); }

    chc_tsv_counts* res = NULL;
    if (chc_count_tsv(argv[1], "sb,vdw", 0, 2, &res) != CHC_OK) return fail("count_tsv");
    int64_t n_res, n_pairs, bytes, total;
    chc_tsv_counts_sizes(res, &n_res, &n_pairs, &bytes, &total);
    if (n_res != 3 || n_pairs != 3 || total != 2) return fail("count sizes");
    char* names = malloc((size_t)bytes + 1);
    int64_t* noff = malloc(sizeof(int64_t) * (size_t)(n_res + 1));
    int32_t* r1 = malloc(sizeof(int32_t) * (size_t)n_pairs); int32_t* r2 = malloc(sizeof(int32_t) * (size_t)n_pairs);
    int64_t* cnt = malloc(sizeof(int64_t) * (size_t)n_pairs);
    chc_tsv_counts_get(res, names, noff, r1, r2, cnt);
    for (int64_t k = 0; k < n_pairs; ++k) if (cnt[k] != 1 || r1[k] == r2[k]) return fail("counts");
    chc_tsv_counts_free(res);
    free(names); free(noff); free(r1); free(r2); free(cnt);

    /* device entry points reject bad arguments before touching CUDA */
    if (chc_submit_frames(NULL, 0, NULL, 1, 0, 0, NULL) != CHC_E_INVALID) return fail("null context");
    if (chc_dcd_unpack(NULL, 0, 0, 1, 10, 0, NULL, 0, 0, NULL) != CHC_E_INVALID) return fail("dcd null");
    if (chc_permuted_pca(NULL, 20, 10, 1, NULL, NULL, 0, NULL, NULL, 0, NULL) != CHC_E_INVALID) return fail("pca null");
    puts("abi_host ok");
    return 0;
}
