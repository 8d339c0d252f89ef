/* A C host of the library, as INTEGRATION.md §4 describes it: nothing but the public header and the .so.
 * Exercises only host-side entry points (no GPU needed): tests/test_abi.py builds and runs it. */
#include <stdio.h>
#include <string.h>

#include "fasttalk_b200.h"

int main(void) {
    int splits = ft_gemm_pick_splits(2048, 64, 2048, 1);
    if (splits < 1) { printf("bad splits %d\n", splits); return 1; }
    ft_engine_config cfg;
    ft_weights w;
    memset(&cfg, 0, sizeof cfg);
    memset(&w, 0, sizeof w);
    ft_engine* e = ft_engine_create(&cfg, &w);          /* an all-zero configuration must be refused, not crash */
    if (e != NULL) { printf("engine created from an empty config\n"); return 2; }
    const char* err = ft_last_error();
    if (err == NULL || strlen(err) == 0) { printf("no error text\n"); return 3; }
    printf("C_HOST_OK splits=%d error=\"%s\" epilogue=%zu config=%zu\n", splits, err, sizeof(ft_gemm_epilogue),
           sizeof(ft_engine_config));
    return 0;
}
