// knn_l2_tc.cu -- placeholder until the tcgen05 ranking kernel lands (next commit).
#include "mpl_ctx.h"
bool   mpl_l2tc_supported(int, int, int) { return false; }
size_t mpl_l2tc_scratch_bytes(int, int, int, int) { return 16; }
int    mpl_l2tc_top2(const float*, int, size_t, const float*, int, size_t, int, int, long long, void*, size_t, mpl_key64*, int, cudaStream_t)
{ mpl_set_error("tensor-core L2 path not built"); return MPL_ERR_BAD_ARG; }
