// stubs.cu -- entry points not implemented yet (fail loudly, never fall back).
#include "sr_internal.cuh"
#define NOTYET(ctx, name) return sr_fail(ctx, SR_E_BADARG, name ": not implemented yet")
extern "C" {
int sr_band_rows(int h, int rank, int nranks, int* row0, int* row1) {
    if (h < 0 || nranks < 1 || rank < 0 || rank >= nranks || !row0 || !row1) return SR_E_BADARG;
    *row0 = (int)((long long)h * rank / nranks);
    *row1 = (int)((long long)h * (rank + 1) / nranks);
    return SR_OK;
}
}
