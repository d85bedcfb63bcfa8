/* Plain-C user of include/cassmantle_b200.h: proves the boundary needs nothing but a C compiler
 * and the shared library.  Built and run by tests/test_gpu_c_abi.py on the GPU box.
 *   gcc -std=c11 -I include tests/c/abi_smoke.c -o abi_smoke cassmantle_b200/libcassmantle_b200.so -lm */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "cassmantle_b200.h"

#define CHECK(call)                                                              \
    do {                                                                         \
        int rc_ = (call);                                                        \
        if (rc_ != CM_OK) {                                                      \
            fprintf(stderr, "%s -> %d: %s\n", #call, rc_, cm_last_error());     \
            return 1;                                                            \
        }                                                                        \
    } while (0)

int main(void) {
    enum { V = 1000, D = 300, B = 5 };
    cm_engine *e = NULL;
    CHECK(cm_engine_create(&e, 0, V, D));
    /* rows: row r = e_(r mod D) scaled by (r+1), so cosines are 0 or 1 by construction */
    float *rows = calloc((size_t)V * D, sizeof(float));
    for (int r = 0; r < V; ++r) rows[(size_t)r * D + r % D] = (float)(r + 1);
    CHECK(cm_engine_upload_rows(e, rows, 0, V));
    int32_t answers[2] = {7, -1};                       /* row 7 = 8*e_7 ; an unknown answer */
    CHECK(cm_set_answers(e, answers, 2, NULL));
    CHECK(cm_engine_synchronize(e));
    int32_t idx[B] = {307, 8, -1, 7, 307};              /* same direction, orthogonal, OOV, itself, vs unknown */
    int32_t len[B] = {1, 1, 1, 1, 1};
    int32_t slot[B] = {0, 0, 0, 0, 1};
    uint8_t exact[B] = {0, 0, 0, 1, 0};
    float scores[B];
    int32_t centi[B];
    CHECK(cm_score_batch_host(e, idx, len, B, 1, slot, exact, 0.01f, scores, centi));
    const float want[B] = {1.0f, 0.01f, 0.01f, 1.0f, 0.01f};
    for (int i = 0; i < B; ++i) {
        printf("score[%d] = %.8g (centi %d)\n", i, scores[i], centi[i]);
        if (fabsf(scores[i] - want[i]) > 2e-6f) { fprintf(stderr, "mismatch at %d\n", i); return 2; }
    }
    if (cm_score_batch_host(e, idx, len, B, 0, slot, exact, 0.01f, scores, NULL) != CM_ERR_INVALID) return 3;
    printf("error text: %s\n", cm_last_error());
    CHECK(cm_engine_destroy(e));
    free(rows);
    printf("abi_smoke ok (library %s, %lld kernel launches)\n", cm_version(), (long long)cm_kernel_launch_count());
    return 0;
}
